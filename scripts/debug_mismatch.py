"""Where does a deterministic shader differ from the oracle? usage: debug_mismatch.py scene shader spp res"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as e
import util
vpt = e.load_package(); b = vpt.binding
scene, shader, spp, res = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
ref = util.oracle_render(vpt, scene, shader, spp, res)
sc = b.Scene(util.scene_path(scene))
r = b.Render(sc, b.make_params(shader=shader, samples=spp, resolution=res)); r.trace(spp)
img = r.image(False)
bad = util.bit_mismatch(img, ref["image_sum"])
print("mismatching pixels", int(bad.sum()), "of", bad.size, "rng equal", np.array_equal(r.rngs(), ref["rngs_after"]))
# first-hit instance per pixel via the 'color'-independent route: camera rays are not exposed, so
# report neighbours' values instead
ys, xs = np.nonzero(bad)
for y, x in list(zip(ys, xs))[:20]:
    g, o = img[y, x], ref["image_sum"][y, x]
    ulps = (g.view(np.int32).astype(np.int64) - o.view(np.int32).astype(np.int64))
    print((x, y), "gpu", g, "ref", o, "ulp diff", ulps)
