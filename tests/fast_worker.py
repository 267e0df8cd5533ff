"""Helper of tests/test_gpu_fast.py: runs in a subprocess with VPT_LIB_DIR pointing at the
statistical (-DVPT_FAST) build, because a process can hold only one build of the libraries.
usage: fast_worker.py render scene shader spp res bounces out.npy
       fast_worker.py hits scene rays.npy out.npz"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as entry  # noqa: E402
import util  # noqa: E402

vpt = entry.load_package()
b = vpt.binding
assert os.path.basename(os.path.dirname(b.LIB_CUDA_PATH)) == "fast", b.LIB_CUDA_PATH
b.set_device(0)
mode, scene = sys.argv[1], sys.argv[2]
sc = b.Scene(util.scene_path(scene))
if mode == "render":
    shader, spp, res, bounces, out = sys.argv[3], int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6]), sys.argv[7]
    r = b.Render(sc, b.make_params(shader=shader, samples=spp, resolution=res, bounces=bounces))
    r.trace(spp)
    np.save(out, r.image(normalized=False))
    r.close()
else:
    rays = np.load(sys.argv[3])
    ids, uvd = sc.intersect(rays)
    np.savez(sys.argv[4], ids=ids, uvd=uvd)
sc.close()
