"""Throughput vs samples per call (the tail is amortised over the batch)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as e
b = e.load_package().binding
b.set_device(0)
sc = b.Scene("scenes/_built/03_volume.vptscene")
r = b.Render(sc, b.make_params(shader="volpathtrace", samples=1 << 20, resolution=1280, bounces=64))
r.trace(4)
for spp in (16, 64, 128, 256, 512, 1024):
    r.trace(spp)
    st = r.stats()
    print(f"{spp:5d} spp/call: {st.device_ms:8.1f} ms  {st.paths / st.device_ms / 1e3:7.1f} Mpaths/s  launches {st.kernel_launches}", flush=True)
