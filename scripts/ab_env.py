"""A/B of kernel variants selected by an environment variable on one GPU (AB_VAR, default VPT_WF_POOL: 0 = ray
per lane, 1 = pooled extend; AB_MODES = the values to run). Device time of a 64 spp call per workload, images
compared bit for bit."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import __graft_entry__ as e
b = e.load_package().binding
b.set_device(0)
cases = [("03_volume", "volpathtrace", 1280, 64, 64), ("01_surface", "pathtrace", 720, 4, 64),
         ("05_head1ss", "volpathtrace", 1280, 64, 64), ("02_rollingteapot", "pathtrace", 720, 4, 64),
         ("07_sdfunction", "implicit", 1920, 4, 16), ("06_gridsdf", "implicit", 1280, 4, 16),
         ("04_head1", "volpathtrace", 1280, 64, 16)]
VAR = os.environ.get("AB_VAR", "VPT_WF_POOL")
only = sys.argv[1:]
for scene, shader, res, bounces, spp in cases:
    if only and scene not in only: continue
    sc = b.Scene(f"scenes/_built/{scene}.vptscene")
    imgs = {}
    for pool in os.environ.get("AB_MODES", "0 1").split():
        os.environ[VAR] = pool
        r = b.Render(sc, b.make_params(shader=shader, samples=4096, resolution=res, bounces=bounces))
        r.trace(4)
        ms = []
        for _ in range(3):
            r.trace(spp); st = r.stats(); ms.append(st.device_ms)
        imgs[pool] = r.image(normalized=False).copy()
        print(f"{scene} {shader} {VAR}={pool}: {spp} spp {min(ms):.1f} ms -> {st.paths/min(ms)/1e3:.1f} Mpaths/s; "
              f"timed extend {st.top_kernel_ms:.3f} ms / {st.top_kernel_units} rays", flush=True)
        r.close()
    if len(imgs) == 2:
        a, c = imgs.values()
        print(f"  images bit-identical: {np.array_equal(a.view(np.uint32), c.view(np.uint32))}", flush=True)
    sc.close()
