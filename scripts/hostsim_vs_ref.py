#!/usr/bin/env python
"""Device headers on the host (tests/hostsim) vs the reference CPU path (oracle/_ref/ref_render):
bit-exact image sums and generator states on the built scenes. A CPU-only pre-flight for kernel
changes that touch traversal order or arithmetic.

    python scripts/hostsim_vs_ref.py [scene shader res spp bounces] ...
"""
import os, subprocess, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as entry
import util
vpt = entry.load_package()
CASES = [("03_volume", "volpathtrace", 240, 2, 16), ("01_surface", "pathtrace", 240, 2, 4),
         ("05_head1ss", "volpathtrace", 160, 1, 8), ("mixed", "volpathtrace", 96, 2, 8),
         ("mixed", "pathtrace", 96, 2, 4), ("02_rollingteapot", "pathtrace", 200, 2, 4),
         ("04_head1", "volpathtrace", 160, 1, 8)]
def main():
    cases = CASES
    if len(sys.argv) > 1:
        a = sys.argv[1:]
        cases = [(a[i], a[i + 1], int(a[i + 2]), int(a[i + 3]), int(a[i + 4])) for i in range(0, len(a), 5)]
    bad_total = 0
    for scene, shader, res, spp, bounces in cases:
        sp = util.scene_path(scene)
        if sp is None:
            print(f"{scene}: not built, skipped"); continue
        with tempfile.TemporaryDirectory() as tmp:
            a, b = os.path.join(tmp, "a"), os.path.join(tmp, "b")
            args = ["--shader", shader, "--samples", str(spp), "--resolution", str(res), "--bounces", str(bounces)]
            subprocess.run([util.REF_RENDER, sp, a, *args, "--dump-state"], check=True, stdout=subprocess.DEVNULL)
            r = subprocess.run([os.path.join(ROOT, "tests/hostsim/hostsim"), sp, b, *args], check=True, capture_output=True, text=True)
            ra, rb = vpt.archive.load(a), vpt.archive.load(b)
            bad = int(util.bit_mismatch(ra["image_sum"], rb["image_sum"]).sum())
            rng_ok = np.array_equal(ra["rngs_after"], rb["rngs_after"])
            bad_total += bad + (0 if rng_ok else 1)
            print(f"{scene} {shader} {res}px {spp}spp b{bounces}: {bad} pixels differ, generators {'equal' if rng_ok else 'DIFFER'}")
            for line in r.stdout.splitlines():
                if "traversal" in line or "node_visits" in line: print("   ", line[:400])
    sys.exit(1 if bad_total else 0)
main()
