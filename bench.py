#!/usr/bin/env python
"""bench.py — Mpaths/s of the path-tracing hot path on N B200s (one process per GPU).

    python bench.py --gpus N --steps K --warmup W [--impl reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

Workload (config.workload): BASELINE.json configs[1] — tests/03_volume rendered with
shade_volpathtrace at 1280x533, --bounces 64 (scripts/run-highres.sh:3 of the reference).
One *step* = 64 samples per pixel over the whole image (43.7 M camera paths); the default
K = 16 steps is exactly the config's 1024 spp. Mpaths/s = W*H*spp / seconds.

  value    device-resident: scene, per-pixel generators and accumulation image already in HBM;
           each step is timed with CUDA events on the launching stream; L2 is flushed between
           steps by an untimed 256 MiB memset.
  e2e      the same steps through the reference-facing API with HOST buffers: every step
           uploads the pixel generators from pinned host memory, traces, and downloads the
           accumulation image (wall clock around the three C-ABI calls).
  roofline fp32 CUDA-core issue (north_star: "FP32 peak divided by the measured box- and
           triangle-test FLOPs per ray"): algorithmic FLOPs per path measured by
           tests/hostsim work counters x paths per step / step time. HBM and tensor cores
           do not bound this path (working set is L2-resident, no dense contraction).
  cpu_baseline / --impl reference: the UNMODIFIED reference CPU path (oracle/_ref/ref_render,
           parallel_for over all host cores) on a bounded sample of the same workload.
Multi-GPU (torchrun): every rank renders the same image with a disjoint stretch of each
pixel's random stream (weak scaling: per-GPU work fixed); the accumulation images are summed
with one NCCL reduce. Time = max over ranks, measured on the device.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: scene, shader, resolution, bounces, spp per step, (w, h)
    "03_volume": dict(scene="03_volume", shader="volpathtrace", resolution=1280, bounces=64,
                      spp_per_step=64, size=(1280, 533)),
    "01_surface": dict(scene="01_surface", shader="pathtrace", resolution=720, bounces=4,
                       spp_per_step=64, size=(720, 300)),
    "05_head1ss": dict(scene="05_head1ss", shader="volpathtrace", resolution=1280, bounces=64,
                       spp_per_step=16, size=(1280, 1280)),
    "06_gridsdf": dict(scene="06_gridsdf", shader="implicit", resolution=1280, bounces=4,
                       spp_per_step=16, size=(1280, 533)),
    "07_sdfunction": dict(scene="07_sdfunction", shader="implicit", resolution=1920, bounces=4,
                          spp_per_step=16, size=(1920, 800)),
}

# Algorithmic work per camera path, measured with tests/hostsim (VPT_WORK_COUNTERS; identical
# arithmetic to the kernels) at 360 px / 4 spp, and the FLOP cost of each unit.
WORK = {
    "03_volume": dict(box_tests=102.83, tri_tests=29.72, inst_transforms=20.63, bounces=2.712,
                      sdf_scene_evals=0, sdf_fn_evals=0, sdf_grid_evals=0, trilinear=0),
    "01_surface": dict(box_tests=49.44, tri_tests=9.19, inst_transforms=13.58, bounces=2.044,
                       sdf_scene_evals=0, sdf_fn_evals=0, sdf_grid_evals=0, trilinear=0),
    "05_head1ss": dict(box_tests=61.12, tri_tests=7.17, inst_transforms=0.93, bounces=1.502,
                       sdf_scene_evals=0, sdf_fn_evals=0, sdf_grid_evals=0, trilinear=0),
    "06_gridsdf": dict(box_tests=0, tri_tests=0, inst_transforms=0, bounces=1.866,
                       sdf_scene_evals=143.89, sdf_fn_evals=366.74, sdf_grid_evals=576.62,
                       trilinear=33.70),
    "07_sdfunction": dict(box_tests=0, tri_tests=0, inst_transforms=0, bounces=1.850,
                          sdf_scene_evals=129.25, sdf_fn_evals=595.43, sdf_grid_evals=259.04,
                          trilinear=14.94),
}
FLOPS = dict(box_tests=26, tri_tests=52, inst_transforms=30, bounces=300, sdf_scene_evals=0,
             sdf_fn_evals=45, sdf_grid_evals=43, trilinear=30)
SM_COUNT, FP32_LANES = 148, 128

# The dominant kernel is k_wf_extend_persistent (one closest-hit query per queued ray). Its algorithmic
# work per RAY: the counters above minus the share spent in light-pdf instance queries
# (hostsim iq_* counters), divided by the rays per path; bytes = nodes 32 B, primitives 48/64 B,
# instance records 64 B, plus the ray read (32 B) and hit write (24 B) that must touch HBM.
EXTEND = {
    "03_volume": dict(box_tests=36.37, tri_tests=10.52, inst_transforms=6.06, prim_bytes=64),
    "01_surface": dict(box_tests=22.45, tri_tests=4.02, inst_transforms=4.90, prim_bytes=48),
    "05_head1ss": dict(box_tests=40.69, tri_tests=4.78, inst_transforms=0.62, prim_bytes=48),
}
# dram__bytes_read.sum + dram__bytes_write.sum of one full-occupancy k_wf_extend_persistent
# launch (one of the two pipes: 341 k rays), from profiles/r01d_ncu_full.brief.txt (ncu --set full),
# per launch like `achieved`
NCU_TRAFFIC = {"03_volume": 46.1e6 + 6.7e6}


def extend_flops_per_ray(name):
    w = EXTEND.get(name)
    return None if w is None else sum(w[k] * FLOPS[k] for k in ("box_tests", "tri_tests", "inst_transforms"))


def extend_bytes_per_ray(name):
    w = EXTEND[name]
    return w["box_tests"] * 32 + w["tri_tests"] * w["prim_bytes"] + w["inst_transforms"] * 64 + 56


def flops_per_path(name):
    return sum(WORK[name][k] * FLOPS[k] for k in FLOPS)


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""

    QUERY = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu, self.stop_flag, self.rows = gpu_index, threading.Event(), []

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.QUERY}",
                                      "--format=csv,noheader,nounits"], capture_output=True,
                                     text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=3)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        mhz = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None,
                "sm_max_mhz": float(self.rows[0][1]) if self.rows[0][1].replace(".", "").isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


def run_reference_cpu(wl, samples, resolution=None):
    """Times the unmodified reference CPU path (oracle/_ref/ref_render) on `samples` spp."""
    ref = os.path.join(ROOT, "oracle", "_ref", "ref_render")
    scene = os.path.join(ROOT, "scenes", "_built", wl["scene"] + ".vptscene")
    if not os.path.exists(ref) or not os.path.exists(scene):
        return None
    import __graft_entry__ as entry

    pkg = entry.load_package()
    with tempfile.TemporaryDirectory() as tmp:
        out = os.path.join(tmp, "ref.vptarch")
        subprocess.run([ref, scene, out, "--shader", wl["shader"], "--samples", str(samples),
                        "--resolution", str(resolution or wl["resolution"]), "--bounces",
                        str(wl["bounces"])], check=True, stdout=subprocess.DEVNULL)
        arch = pkg.archive.load(out)
    w, h, spp, threads = (int(v) for v in arch["meta"])
    seconds = float(arch["seconds"][0])
    return {"mpaths": w * h * spp / seconds / 1e6, "seconds": seconds, "threads": threads,
            "paths": w * h * spp, "size": (w, h), "spp": spp}


def reference_arm(args, wl, rank):
    """--impl reference: the reference's own CPU implementation on the host cores, one bounded
    sample (2 spp of the full-resolution workload) per step."""
    if rank != 0:
        return 0
    spp = 2
    times, info = [], None
    for i in range(args.warmup + args.steps):
        info = run_reference_cpu(wl, spp)
        if info is None:
            print(json.dumps({"impl": "reference", "unavailable":
                              "oracle/_ref/ref_render or the scene archive is missing"}))
            return 0
        if i >= args.warmup:
            times.append(info["seconds"])
    total = sum(times)
    value = info["paths"] * len(times) / total / 1e6
    sample = f"{spp} spp of {info['size'][0]}x{info['size'][1]} per step ({info['paths']} paths)"
    line = {"impl": "reference", "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, wl),
            "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": info["threads"],
                             "kind": "reference", "sample": sample},
            "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def workload_config(args, wl):
    return {"workload": f"tests/{wl['scene']} {wl['shader']} {wl['size'][0]}x{wl['size'][1]} "
                        f"bounces {wl['bounces']}, {wl['spp_per_step']} spp per step",
            "scene": wl["scene"], "shader": wl["shader"], "resolution": wl["resolution"],
            "bounces": wl["bounces"], "spp_per_step": wl["spp_per_step"],
            "paths_per_step_per_gpu": wl["size"][0] * wl["size"][1] * wl["spp_per_step"],
            "pipeline": args.pipeline,
            "l2": "flushed between steps by an untimed 256 MiB memset",
            "parallelism": f"sample-range x{args.gpus}, one NCCL reduce"}


def gpu_arm(args, wl, rank, world, local_rank):
    import numpy as np
    import torch

    import __graft_entry__ as entry

    pkg = entry.load_package()
    b = pkg.binding
    if not torch.cuda.is_available() or b.device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: the B200 path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    b.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    scene_path = os.path.join(ROOT, "scenes", "_built", wl["scene"] + ".vptscene")
    if not os.path.exists(scene_path):
        raise SystemExit(f"{scene_path} is missing: run __graft_entry__.build() where the reference is")
    spp = wl["spp_per_step"]
    total_spp = (args.warmup + args.steps) * spp + 2
    t0 = time.time()
    scene = b.Scene(scene_path)
    host_build_s = time.time() - t0
    params = b.make_params(shader=wl["shader"], samples=total_spp, resolution=wl["resolution"],
                           bounces=wl["bounces"], pipeline=args.pipeline)
    r = b.Render(scene, params)
    # sample-range sharding: rank r owns a disjoint stretch of every pixel's PCG32 stream
    rng_offset = pkg.sharding.rank_rng_offset(rank)
    r.reset(rng_offset)
    paths_per_step = r.width * r.height * spp
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    t0 = time.time()
    r.trace(1)  # first call uploads the scene
    upload_s = time.time() - t0
    r.reset(rng_offset)
    for _ in range(args.warmup):
        r.trace(spp)
    if dist is not None:
        # warm-up of the one collective too (NCCL sets up its channels on first use)
        scratch = torch.zeros(r.height * r.width * 4, dtype=torch.float32, device="cuda")
        pkg.sharding.reduce_images(scratch, dst=0)
        del scratch

    # ---- device-resident timed region ------------------------------------------------------
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    step_ms, launches = [], 0
    top_ms, top_units, top_launches = 0.0, 0, 0
    for _ in range(args.steps):
        flush.zero_()
        torch.cuda.synchronize()
        r.trace(spp)
        st = r.stats()
        step_ms.append(st.device_ms)
        top_ms += st.top_kernel_ms
        top_units += st.top_kernel_units
        top_launches += st.top_kernel_launches
        launches += st.kernel_launches
    barrier()
    clocks = sampler.summary()
    total_ms = float(sum(step_ms))

    # ---- one reduce of the accumulation images (NCCL over NVLink) ---------------------------
    reduce_ms = 0.0
    image_mean = None
    done_spp = r.samples
    if dist is not None:
        acc = torch.empty(r.height * r.width * 4, dtype=torch.float32, device="cuda")
        r.export_image(acc.data_ptr())
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        pkg.sharding.reduce_images(acc, dst=0)
        e1.record()
        torch.cuda.synchronize()
        reduce_ms = e0.elapsed_time(e1)
        if rank == 0:
            r.import_image(acc.data_ptr(), done_spp * world)
    if rank == 0:
        img = r.image(normalized=True)
        image_mean = float(np.clip(img[..., :3], 0, 4).mean())

    # ---- end to end through the host API with host buffers ------------------------------------
    e2e_steps = max(1, min(args.steps, 8))
    barrier()
    t0 = time.time()
    for _ in range(e2e_steps):
        r.reset(rng_offset)          # H2D: per-pixel generators from (pinned) host memory
        r.trace(spp)
        host_img = r.image(normalized=False)  # D2H: the accumulation image
    torch.cuda.synchronize()
    e2e_s = time.time() - t0
    del host_img
    pix = r.width * r.height

    # ---- the dominant kernel on its own (untimed extra call, one pipe) ---------------------------
    # In the timed region two wavefront pipes share the device, so the event pair around an extend
    # launch also spans the other pipe's kernels. One more call with a single pipe gives the
    # kernel's own rate; it is reported beside the timed-region figure, never instead of it.
    alone = None
    if rank == 0 and "VPT_WF_PIPES" not in os.environ:
        os.environ["VPT_WF_PIPES"] = "1"
        r.reset(rng_offset)
        r.trace(spp)
        st = r.stats()
        del os.environ["VPT_WF_PIPES"]
        if st.top_kernel_launches > 0 and st.top_kernel_ms > 0:
            alone = (st.top_kernel_ms, st.top_kernel_units, st.top_kernel_launches)

    # ---- max over ranks -----------------------------------------------------------------------
    times = torch.tensor([total_ms + reduce_ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    worst_ms, worst_e2e_ms = (float(v) for v in times.tolist())

    if rank == 0:
        value = world * paths_per_step * args.steps / (worst_ms / 1e3) / 1e6
        e2e_value = world * paths_per_step * e2e_steps / (worst_e2e_ms / 1e3) / 1e6
        peaks = measured_peaks()
        sm_max = float(peaks.get("sm_max_mhz", 1965.0))
        fp32_peak = SM_COUNT * FP32_LANES * 2 * sm_max * 1e6 / 1e12
        fpp = flops_per_path(args.workload)
        ms_per_step = worst_ms / args.steps
        step_tflops = fpp * paths_per_step / (ms_per_step / 1e3) / 1e12
        wavefront = args.pipeline != "mega" and extend_flops_per_ray(args.workload) is not None \
            and top_launches > 0 and wl["shader"] in ("volpathtrace", "pathtrace", "naive")
        if wavefront:
            # dominant kernel, timed live with CUDA events on its own stream (one launch per
            # counter poll, ray count known exactly): flops per ray x rays / duration
            kernel = "k_wf_extend_persistent"
            achieved = extend_flops_per_ray(args.workload) * top_units / (top_ms / 1e3) / 1e12
            per_launch = {"kernel": kernel, "sampled_launches": top_launches,
                          "avg_launch_us": 1e3 * top_ms / top_launches,
                          "avg_rays_per_launch": top_units / top_launches,
                          "grays_per_s": top_units / (top_ms / 1e3) / 1e9,
                          "flops_per_ray": extend_flops_per_ray(args.workload),
                          "algorithmic_bytes_per_ray": extend_bytes_per_ray(args.workload),
                          "note": "timed region: the other wavefront pipe's kernels share the device "
                                  "during these launches"}
            if alone is not None:
                a_ms, a_units, a_n = alone
                a_tflops = extend_flops_per_ray(args.workload) * a_units / (a_ms / 1e3) / 1e12
                per_launch["alone"] = {"sampled_launches": a_n, "avg_launch_us": 1e3 * a_ms / a_n,
                                       "avg_rays_per_launch": a_units / a_n,
                                       "grays_per_s": a_units / (a_ms / 1e3) / 1e9,
                                       "achieved": a_tflops, "frac": a_tflops / fp32_peak,
                                       "note": "one untimed extra call with a single pipe "
                                               "(VPT_WF_PIPES=1): the kernel with the device to itself"}
        else:
            kernel, achieved, per_launch = "whole step", step_tflops, None
        line = {
            "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload_config(args, wl),
            "e2e": {"value": e2e_value, "unit": "Mpaths/s", "h2d_bytes_per_step": pix * 16,
                    "d2h_bytes_per_step": pix * 16, "steps": e2e_steps},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": fp32_peak,
                         "unit": "TFLOP/s", "frac": achieved / fp32_peak,
                         "traffic": NCU_TRAFFIC.get(args.workload) if wavefront else None,
                         "kernel": kernel, "per_launch": per_launch,
                         "whole_step": {"achieved": step_tflops, "frac": step_tflops / fp32_peak,
                                        "flops_per_path": fpp},
                         "note": "compute roofline of north_star (FP32 peak / algorithmic flops; "
                                 "HBM and tensor cores do not bound this path): algorithmic work from "
                                 "tests/hostsim counters; peak = 148 SM x 128 lanes x 2 x sm_max_mhz "
                                 "of MEASURED_PEAKS.json, which holds no fp32 figure; traffic = ncu "
                                 "dram bytes of one full-occupancy launch (profiles/)"},
            "reduce_ms": reduce_ms, "image_mean": image_mean,
            "host_build_s": host_build_s, "scene_upload_s": upload_s,
        }
        if world == 1 and not args.no_cpu_baseline:
            info = run_reference_cpu(wl, 8)
            if info is not None:
                line["cpu_baseline"] = {
                    "value": info["mpaths"], "unit": "Mpaths/s", "cores": info["threads"],
                    "kind": "reference",
                    "sample": f"8 spp of {info['size'][0]}x{info['size'][1]} ({info['paths']} paths, "
                              f"{info['seconds']:.2f} s) by oracle/_ref/ref_render"}
        print(json.dumps(line), flush=True)
    r.close()
    scene.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="03_volume", choices=sorted(WORKLOADS))
    ap.add_argument("--pipeline", default="auto", choices=["auto", "mega", "wavefront"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    wl = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return reference_arm(args, wl, rank)
    if world != args.gpus and world == 1 and args.gpus > 1:
        # not under torchrun: relaunch ourselves with one process per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
               f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1", "--master-port",
               str(29500 + os.getpid() % 1000), os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    return gpu_arm(args, wl, rank, world, local_rank)


if __name__ == "__main__":
    sys.exit(main())
