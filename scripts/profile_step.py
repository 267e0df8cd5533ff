"""One wavefront render call for profilers: `ncu ... python scripts/profile_step.py [scene shader res bounces spp]`.
Numbers printed under a profiler are not bench values."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as e

b = e.load_package().binding
b.set_device(0)
a = sys.argv[1:]
scene, shader = (a + ["03_volume", "volpathtrace"])[:2] if len(a) < 2 else a[:2]
res, bounces, spp = (int(x) for x in (a[2:5] if len(a) >= 5 else ("1280", "64", "16")))
pipe = a[5] if len(a) > 5 else "wavefront"
sc = b.Scene(f"scenes/_built/{scene}.vptscene")
r = b.Render(sc, b.make_params(shader=shader, samples=4096, resolution=res, bounces=bounces, pipeline=pipe))
r.trace(spp)
st = r.stats()
print(f"{scene} {shader} {r.width}x{r.height} b{bounces} {pipe}: {spp} spp {st.device_ms:.1f} ms, "
      f"{st.kernel_launches} launches, extend rays {st.rays_extend}, pdf rays {st.rays_lightpdf}")
