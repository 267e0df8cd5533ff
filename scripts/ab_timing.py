"""A/B timing of render calls: one warm-up call, then N timed calls; prints the median device time.
usage: ab_timing.py scene shader res bounces spp [calls]   (env: VPT_LIB_DIR, VPT_WF_* select the variant)"""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as e
b = e.load_package().binding
b.set_device(0)
scene, shader, res, bounces, spp = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
calls = int(sys.argv[6]) if len(sys.argv) > 6 else 5
sc = b.Scene(f"scenes/_built/{scene}.vptscene")
r = b.Render(sc, b.make_params(shader=shader, samples=4096, resolution=res, bounces=bounces, pipeline="wavefront"))
r.trace(spp)
ms, prev = [], r.stats().device_ms
for _ in range(calls):
    r.trace(spp)
    st = r.stats()
    ms.append(st.device_ms)
med = statistics.median(ms)
print(f"{scene} {shader} {r.width}x{r.height} b{bounces} {spp} spp: median {med:.1f} ms (min {min(ms):.1f}, max {max(ms):.1f}) "
      f"-> {r.width * r.height * spp / med / 1e3:.1f} Mpaths/s; {st.kernel_launches} kernels, {st.graph_launches} graph launches; "
      f"tag={os.environ.get('VPT_LIB_DIR', '')} {' '.join(k + '=' + v for k, v in os.environ.items() if k.startswith("VPT_"))}", flush=True)
