#!/usr/bin/env python
"""Builds the scene archives (scenes/_built/*.vptscene) the tests and bench render.

Runs only where /root/reference exists (this container): each reference test
scene is loaded by the reference's own loaders through oracle/_ref/vpt_export
(JSON, PLY/OBJ, PNG/HDR, .sdf grids, Catmull-Clark tessellation) and written as
a named-array archive that both the CUDA path and the reference CPU renderer
(oracle/_ref/ref_render) read. Six of the seven reference scenes lack assets
(.MISSING_LARGE_BLOBS); deterministic stand-ins are generated into
scenes/_work/<scene>/ next to symlinks of the shipped files (SURVEY.md F2,
section 8(d)). Nothing from the reference is copied into git: _work and _built
are ignored, but travel to the GPU box with the snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
REF_TESTS = "/root/reference/tests"
WORK = os.path.join(HERE, "_work")
BUILT = os.path.join(HERE, "_built")
EXPORT = os.path.join(ROOT, "oracle", "_ref", "vpt_export")


def link_tree(src: str, dst: str) -> None:
    """Mirror `src` under `dst` with symlinks to the files (directories are real)."""
    for base, _dirs, files in os.walk(src):
        rel = os.path.relpath(base, src)
        os.makedirs(os.path.join(dst, rel), exist_ok=True)
        for f in files:
            target = os.path.join(dst, rel, f)
            if not os.path.lexists(target):
                os.symlink(os.path.join(base, f), target)


def export(scene_dir: str, json_name: str, out_name: str, force: bool = False) -> str:
    out = os.path.join(BUILT, out_name + ".vptscene")
    if os.path.exists(out) and not force:
        return out
    os.makedirs(BUILT, exist_ok=True)
    subprocess.run([EXPORT, os.path.join(scene_dir, json_name), out + ".tmp"], check=True)
    os.replace(out + ".tmp", out)
    return out


def build_03_volume(force=False):
    # the one scene whose assets are all shipped
    return export(os.path.join(REF_TESTS, "03_volume"), "volume.json", "03_volume", force)


def main(argv):
    force = "--force" in argv
    only = [a for a in argv if not a.startswith("--")]
    import standins  # generators for the missing assets

    jobs = {
        "03_volume": build_03_volume,
        "01_surface": standins.build_01_surface,
        "02_rollingteapot": standins.build_02_rollingteapot,
        "04_head1": standins.build_04_head1,
        "05_head1ss": standins.build_05_head1ss,
        "06_gridsdf": standins.build_06_gridsdf,
        "07_sdfunction": standins.build_07_sdfunction,
    }
    for name, fn in jobs.items():
        if only and name not in only:
            continue
        print(f"[make_scenes] {name}", flush=True)
        fn(force)
    return 0


if __name__ == "__main__":
    if not os.path.isdir(REF_TESTS) or not os.path.exists(EXPORT):
        print("[make_scenes] reference or oracle/_ref/vpt_export absent: nothing to do")
        sys.exit(0)
    sys.exit(main(sys.argv[1:]))
