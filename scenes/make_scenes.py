#!/usr/bin/env python
"""Builds the scene archives (scenes/_built/*.vptscene) the tests and bench render.

Runs only where /root/reference exists (this container): each reference test
scene is loaded by the reference's own loaders through tools/vpt_export
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
EXPORT = os.path.join(ROOT, "tools", "vpt_export")


def link_tree(src: str, dst: str) -> None:
    """Mirror `src` under `dst` with symlinks to the files (directories are real)."""
    for base, _dirs, files in os.walk(src):
        rel = os.path.relpath(base, src)
        os.makedirs(os.path.join(dst, rel), exist_ok=True)
        for f in files:
            target = os.path.join(dst, rel, f)
            if not os.path.lexists(target):
                os.symlink(os.path.join(base, f), target)


def export(scene_dir: str, json_name: str, out_name: str, force: bool = False,
           keep_subdivs: bool = False) -> str:
    """Reference loaders (+ reference tessellation unless keep_subdivs) -> archive. The control
    cages of the scene's subdivision surfaces also go to <out_name>.subdivs.vptarch, a small file
    from which tests re-tessellate on the device and compare with the reference's result."""
    out = os.path.join(BUILT, out_name + ".vptscene")
    if os.path.exists(out) and not force:
        return out
    os.makedirs(BUILT, exist_ok=True)
    cmd = [EXPORT, os.path.join(scene_dir, json_name), out + ".tmp"]
    if keep_subdivs:
        cmd.append("--keep-subdivs")
    else:
        cmd += ["--dump-subdivs", os.path.join(BUILT, out_name + ".subdivs.vptarch")]
    subprocess.run(cmd, check=True)
    os.replace(out + ".tmp", out)
    return out


def build_03_volume(force=False):
    # the one scene whose assets are all shipped
    return export(os.path.join(REF_TESTS, "03_volume"), "volume.json", "03_volume", force)


JSON = os.path.join(HERE, "_json")


def build_json_scenes(force=False):
    """Scenes in the reference's own JSON format for the integration test (the adapter of
    integration/ is driven through the reference's load_scene): scenes/_json/ is git-ignored
    and travels to the GPU box, where /root/reference does not exist.
      03_volume/  volume.json + exactly the assets it names (all shipped), copied
      sdf_mini/   the analytic half of tests/07_sdfunction (no voxel grids) plus a sphere and a
                  bbox: every `sdfunctions` type the adapter has to recover from the JSON"""
    import json
    import shutil

    src = os.path.join(REF_TESTS, "03_volume")
    dst = os.path.join(JSON, "03_volume")
    doc = json.load(open(os.path.join(src, "volume.json")))
    uris = ["volume.json"] + [e["uri"] for group in doc.values() if isinstance(group, list)
                              for e in group if isinstance(e, dict) and "uri" in e]
    for uri in uris:
        target = os.path.join(dst, uri)
        if force or not os.path.exists(target):
            os.makedirs(os.path.dirname(target), exist_ok=True)
            shutil.copyfile(os.path.join(src, uri), target)
    src = os.path.join(REF_TESTS, "07_sdfunction")
    dst = os.path.join(JSON, "sdf_mini")
    doc = json.load(open(os.path.join(src, "sdfunction.json")))
    doc.pop("volumes", None), doc.pop("vol_instances", None)
    ident = [1, 0, 0, 0, 1, 0, 0, 0, 1]
    doc["sdfunctions"] += [
        {"name": "sphere1", "type": "sphere", "radius": 0.07, "material": 3,
         "frame": ident + [-0.35, -0.07, -0.1]},
        {"name": "bbox1", "type": "bbox", "thickness": 0.008, "whd": [0.07, 0.07, 0.07],
         "material": 5, "frame": ident + [0.35, -0.075, 0.1]}]
    os.makedirs(os.path.join(dst, "textures"), exist_ok=True)
    for e in doc.get("textures", []):
        target = os.path.join(dst, e["uri"])
        if force or not os.path.exists(target):
            shutil.copyfile(os.path.join(src, e["uri"]), target)
    with open(os.path.join(dst, "sdf_mini.json"), "w") as f:
        json.dump(doc, f, indent=1)
    return dst


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
        "json": build_json_scenes,
    }
    for name, fn in jobs.items():
        if only and name not in only:
            continue
        print(f"[make_scenes] {name}", flush=True)
        fn(force)
    return 0


if __name__ == "__main__":
    if not os.path.isdir(REF_TESTS) or not os.path.exists(EXPORT):
        print("[make_scenes] reference or tools/vpt_export absent: nothing to do")
        sys.exit(0)
    sys.exit(main(sys.argv[1:]))
