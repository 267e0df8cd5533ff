import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as e
p = e.load_package(); b = p.binding
b.set_device(0)
cases = [
    ("03_volume", "volpathtrace", 1280, 64, 64),
    ("03_volume", "pathtrace", 1280, 4, 64),
    ("01_surface", "pathtrace", 720, 4, 64),
    ("07_sdfunction", "implicit", 1280, 4, 16),
    ("06_gridsdf", "implicit", 1280, 4, 16),
]
for scene, shader, res, bounces, spp in cases:
  sc = b.Scene(f"scenes/_built/{scene}.vptscene")
  for pipe in ("mega", "wavefront"):
    try:
        r = b.Render(sc, b.make_params(shader=shader, samples=4096, resolution=res, bounces=bounces, pipeline=pipe))
        r.trace(2)
        t0 = time.time(); r.trace(spp); dt = time.time() - t0
        st = r.stats()
        print(f"{scene} {shader} {r.width}x{r.height} b{bounces} {pipe}: {spp} spp device {st.device_ms:.1f} ms wall {dt*1e3:.1f} ms -> {st.paths/st.device_ms/1e3:.1f} Mpaths/s; launches {st.kernel_launches} extend rays {st.rays_extend} pdf {st.rays_lightpdf} first-extend {st.top_kernel_ms:.3f} ms", flush=True)
        r.close()
    except Exception as ex:
        print(scene, shader, pipe, "FAILED", ex, flush=True)
  sc.close()
