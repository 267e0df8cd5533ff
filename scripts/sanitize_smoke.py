"""Small renders that touch every scheduling path in a few seconds (written for compute-sanitizer, which is closed
on this GPU pool; run plainly it is a quick crash / NaN check). Covers the two-pipe wavefront, the drain hand-over, parked sphere-tracing rays, the megakernel, device-built
lights and the sRGB output stage on the committed fixtures."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as e

b = e.load_package().binding
b.set_device(0)
G = os.path.join(e.ROOT, "tests", "golden")
for scene, shader, pipe, env in [("mixed", "volpathtrace", "wavefront", {"VPT_WF_PIPES": "2"}),
                                 ("mixed", "pathtrace", "wavefront", {"VPT_WF_PIPES": "3", "VPT_WF_DRAIN": "0"}),
                                 ("mixed", "naive", "mega", {}),
                                 ("sdf_small", "implicit", "wavefront", {"VPT_WF_SUSPEND": "24"}),
                                 ("sdf_small", "implicit_normal", "mega", {})]:
    os.environ.update(env)
    sc = b.Scene(os.path.join(G, scene + ".vptscene"))
    sc.use_device_lights()
    r = b.Render(sc, b.make_params(shader=shader, samples=3, resolution=96, bounces=8, pipeline=pipe))
    r.trace(3)
    img = r.image()
    r.image_srgb8()
    print(scene, shader, pipe, env, float(img[..., :3].mean()), flush=True)
    r.close(); sc.close()
    for k in env:
        del os.environ[k]
print("sanitize_smoke: done")
