import sys, time, numpy as np
import __graft_entry__ as e
p = e.load_package(); b = p.binding
what = sys.argv[1]
t0=time.time()
b.set_device(0); print("set_device", time.time()-t0, flush=True)
sc = b.Scene("scenes/_built/03_volume.vptscene"); print("scene", time.time()-t0, flush=True)
if what == "isect":
    rays = np.array([[0,0.3,1.2, 0,-0.3,-1, 1e-4, 3e38]]*64, dtype=np.float32)
    print(sc.intersect(rays)[0][:2], time.time()-t0, flush=True)
else:
    shader, res, spp, bounces = what, int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    r = b.Render(sc, b.make_params(shader=shader, samples=4096, resolution=res, bounces=bounces, pipeline="mega"))
    print("render created", time.time()-t0, flush=True)
    r.trace(1)
    r.trace(spp); print("traced", r.stats().device_ms, "ms", time.time()-t0, flush=True)
    print(r.image()[..., :3].mean())
import ctypes
flags = (ctypes.c_int32*16)()
b.libs()[0].vpt_cuda_debug_flags(flags); print("guard flags", list(flags), flush=True)
