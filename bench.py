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
  workloads short runs (--side-steps, default 3) of the other six tests/ scenes, so that one
           driver-run line covers every configuration: value, e2e, whole-step roofline fraction and
           (N = 1) the reference CPU baseline of each.
  strong   BASELINE.json's split configurations (07_sdfunction 1920x800 4096 spp, 03_volume
           1280x533 1024 spp): the TOTAL spp divided over the ranks, clock from scene upload to the
           one download of the summed image. The driver derives efficiency from the per-N lines.
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
    # the two remaining tests/ scenes at the size of the reference's scripts/run.sh:2,4
    "02_rollingteapot": dict(scene="02_rollingteapot", shader="pathtrace", resolution=720, bounces=4,
                             spp_per_step=64, size=(720, 300)),
    "04_head1": dict(scene="04_head1", shader="volpathtrace", resolution=720, bounces=4,
                     spp_per_step=32, size=(720, 720)),
}
# what `data` says: tests/03_volume ships all its assets; the other six scenes lack meshes /
# textures / grids in the reference checkout (.MISSING_LARGE_BLOBS) and run on deterministic
# stand-ins (scenes/standins.py) inside the reference's own scene JSON
DATA = {name: ("real: the reference's tests/03_volume scene as shipped" if name == "03_volume" else
               "synthetic stand-in assets in the reference's scene JSON (scenes/standins.py)")
        for name in WORKLOADS}
# BASELINE.json configs the strong-scaling leg renders: total spp split over the GPUs
STRONG = {"07_sdfunction": 4096, "03_volume": 1024}

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
    # hostsim with the binary tree (VPT_NO_WIDE=1): box tests after hoisting / instance culling,
    # i.e. slightly BELOW the reference's own count, so these fractions are conservative
    "02_rollingteapot": dict(box_tests=22.82, tri_tests=6.27, inst_transforms=5.96, bounces=1.968,
                             sdf_scene_evals=0, sdf_fn_evals=0, sdf_grid_evals=0, trilinear=0),
    "04_head1": dict(box_tests=30.67, tri_tests=4.26, inst_transforms=0.63, bounces=1.199,
                     sdf_scene_evals=0, sdf_fn_evals=0, sdf_grid_evals=0, trilinear=0),
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
def ncu_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of one full-occupancy k_wf_extend_persistent
    launch (one of the two pipes), read from the NEWEST committed ncu --set full brief under
    profiles/ (scripts/ncu_brief.py output of the 03_volume capture). -> (bytes, file) or (None, None)"""
    if workload != "03_volume":
        return None, None
    import glob

    best = None
    for path in glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_full*.brief.txt")):
        if "sdf" in os.path.basename(path):
            continue
        try:
            lines = open(path).read().splitlines()
            cols = lines[0].split()[1:]
            col = next(i for i, c in enumerate(cols) if c.startswith("k_wf_extend_persistent"))
            val = {}
            for ln in lines[1:]:
                for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    if ln.startswith(key):
                        unit = ln[ln.index("[") + 1:ln.index("]")]
                        scale = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}[unit]
                        val[key] = float(ln[60:].split()[col]) * scale
            if len(val) == 2 and (best is None or os.path.basename(path) > os.path.basename(best[1])):
                best = (val["dram__bytes_read.sum"] + val["dram__bytes_write.sum"], path)
        except Exception:
            continue
    return (best[0], os.path.relpath(best[1], ROOT)) if best else (None, None)


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
        # NVML in-process when the binding is here (a query costs microseconds); else the nvidia-smi
        # binary. Spawning nvidia-smi five times a second from rank 0 delayed that rank by some
        # 10 ms per step on an 8-GPU box, which the other ranks then waited for inside the reduce.
        nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            handle = pynvml.nvmlDeviceGetHandleByIndex(self.gpu)
            nvml = pynvml
        except Exception:
            nvml = None
        while not self.stop_flag.is_set():
            try:
                if nvml is not None:
                    sm = nvml.nvmlDeviceGetClockInfo(handle, nvml.NVML_CLOCK_SM)
                    top = nvml.nvmlDeviceGetMaxClockInfo(handle, nvml.NVML_CLOCK_SM)
                    try:
                        bits = nvml.nvmlDeviceGetCurrentClocksEventReasons(handle)
                    except Exception:
                        bits = nvml.nvmlDeviceGetCurrentClocksThrottleReasons(handle)
                    flag = lambda mask: "Active" if bits & mask else "Not Active"
                    # hw_slowdown 0x8, hw_thermal_slowdown 0x40, sw_thermal_slowdown 0x20, sw_power_cap 0x4
                    self.rows.append([str(sm), str(top), flag(0x8), flag(0x40), flag(0x20), flag(0x4)])
                else:
                    out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.QUERY}",
                                          "--format=csv,noheader,nounits"], capture_output=True,
                                         text=True, timeout=5).stdout.strip()
                    if out:
                        self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.1 if nvml is not None else 0.5)

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
    config = workload_config(args, wl)
    # the bounded sample this arm really runs per step (throughput per path does not depend on
    # spp: the reference does one parallel_for barrier per sample)
    config["spp_per_step"] = spp
    config["paths_per_step_per_gpu"] = info["paths"]
    config["workload"] += f" [reference arm: {spp} spp per step]"
    line = {"impl": "reference", "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": DATA[args.workload],
            "config": config,
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


class Bench:
    """Shared plumbing of the GPU arm: package, device, process group, L2 flush buffer."""

    def __init__(self, args, rank, world, local_rank):
        import torch

        import __graft_entry__ as entry

        self.args, self.rank, self.world, self.local_rank = args, rank, world, local_rank
        self.torch = torch
        self.pkg = entry.load_package()
        self.b = self.pkg.binding
        if not torch.cuda.is_available() or self.b.device_count() < 1:
            raise SystemExit("bench.py needs a CUDA device: the B200 path has no CPU fallback")
        torch.cuda.set_device(local_rank)
        self.b.set_device(local_rank)
        self.dist = None
        if world > 1:
            import torch.distributed as dist

            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            self.dist = dist
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        self.rng_offset = self.pkg.sharding.rank_rng_offset(rank)

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, values):
        t = self.torch.tensor(values, dtype=self.torch.float64, device="cuda")
        if self.dist is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

    def gather_over_ranks(self, value):
        t = self.torch.zeros(self.world, dtype=self.torch.float64, device="cuda")
        t[self.rank] = value
        if self.dist is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return [round(float(v), 3) for v in t.tolist()]

    def reduce_to_rank0(self, r, total_samples):
        """The one collective of the path: sum of the accumulation images on rank 0 (NCCL)."""
        if self.dist is None:
            return 0.0
        torch = self.torch
        acc = torch.empty(r.height * r.width * 4, dtype=torch.float32, device="cuda")
        r.export_image(acc.data_ptr())
        torch.cuda.synchronize()
        # ranks enter the collective together: its device time is then the reduce itself, not the
        # wait for a rank whose host thread fell behind between steps
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        self.pkg.sharding.reduce_images(acc, dst=0)
        e1.record()
        torch.cuda.synchronize()
        if self.rank == 0:
            r.import_image(acc.data_ptr(), total_samples)
        return e0.elapsed_time(e1)

    def scene(self, name):
        path = os.path.join(ROOT, "scenes", "_built", WORKLOADS[name]["scene"] + ".vptscene")
        if not os.path.exists(path):
            raise SystemExit(f"{path} is missing: run __graft_entry__.build() where the reference is")
        t0 = time.time()
        sc = self.b.Scene(path)
        return sc, time.time() - t0


def fp32_peak():
    sm_max = float(measured_peaks().get("sm_max_mhz", 1965.0))
    return SM_COUNT * FP32_LANES * 2 * sm_max * 1e6 / 1e12


def measure(bn, name, steps, warmup, full):
    """Device-resident value, end-to-end value and whole-step roofline fraction of one workload.
    `full`: the headline run (clock sampling, per-launch roofline of the dominant kernel)."""
    args, b, torch, np = bn.args, bn.b, bn.torch, __import__("numpy")
    wl = WORKLOADS[name]
    spp = wl["spp_per_step"]
    scene, host_build_s = bn.scene(name)
    params = b.make_params(shader=wl["shader"], samples=(warmup + steps) * spp + 2,
                           resolution=wl["resolution"], bounces=wl["bounces"], pipeline=args.pipeline)
    r = b.Render(scene, params)
    # sample-range sharding: rank r owns a disjoint stretch of every pixel's PCG32 stream
    r.reset(bn.rng_offset)
    paths_per_step = r.width * r.height * spp
    t0 = time.time()
    r.trace(1)  # first call uploads the scene
    upload_s = time.time() - t0
    r.reset(bn.rng_offset)
    for _ in range(warmup):
        r.trace(spp)
    if full and bn.dist is not None:
        # warm-up of the one collective too (NCCL sets up its channels on first use)
        scratch = torch.zeros(r.height * r.width * 4, dtype=torch.float32, device="cuda")
        bn.pkg.sharding.reduce_images(scratch, dst=0)
        del scratch

    # ---- device-resident timed region ------------------------------------------------------
    bn.barrier()
    sampler = ClockSampler(bn.local_rank) if full else None
    if sampler:
        sampler.start()
    step_ms, launches, graphs = [], 0, 0
    top_ms, top_units, top_launches = 0.0, 0, 0
    for _ in range(steps):
        bn.flush.zero_()
        torch.cuda.synchronize()
        r.trace(spp)
        st = r.stats()
        step_ms.append(st.device_ms)
        top_ms += st.top_kernel_ms
        top_units += st.top_kernel_units
        top_launches += st.top_kernel_launches
        launches += st.kernel_launches
        graphs += st.graph_launches
    bn.barrier()
    clocks = sampler.summary() if sampler else None
    total_ms = float(sum(step_ms))

    # ---- one reduce of the accumulation images (NCCL over NVLink) ---------------------------
    done_spp = r.samples
    reduce_ms = bn.reduce_to_rank0(r, done_spp * bn.world)
    image_mean = None
    if bn.rank == 0:
        img = r.image(normalized=True)
        image_mean = float(np.clip(img[..., :3], 0, 4).mean())

    r.pin()  # page-locked host mirrors for the transfer loop below (outside the clock, like cudaMallocHost)
    # ---- end to end through the host API with host buffers -----------------------------------
    # per step: generators up from pinned host memory, trace, (N > 1: the NCCL reduce,) the
    # accumulation image down to the host
    e2e_steps = max(1, min(steps, 8))
    bn.barrier()
    t0 = time.time()
    for _ in range(e2e_steps):
        r.reset(bn.rng_offset)          # H2D
        r.trace(spp)
        bn.reduce_to_rank0(r, spp * bn.world)
        host_img = r.image_view()  # D2H into the render's page-locked host image (pathtrace_state::image)
    torch.cuda.synchronize()
    e2e_s = time.time() - t0
    del host_img
    pix = r.width * r.height

    # ---- the dominant kernel on its own (untimed extra call, one pipe) ------------------------
    alone = None
    if full and bn.rank == 0 and "VPT_WF_PIPES" not in os.environ:
        os.environ["VPT_WF_PIPES"] = "1"
        r.reset(bn.rng_offset)
        r.trace(spp)
        st = r.stats()
        del os.environ["VPT_WF_PIPES"]
        if st.top_kernel_launches > 0 and st.top_kernel_ms > 0:
            alone = (st.top_kernel_ms, st.top_kernel_units, st.top_kernel_launches)
    r.close()
    scene.close()

    worst_ms, worst_e2e_ms = bn.max_over_ranks([total_ms + reduce_ms, e2e_s * 1e3])
    per_rank_ms = bn.gather_over_ranks(total_ms / steps)  # every rank's own device time per step
    value = bn.world * paths_per_step * steps / (worst_ms / 1e3) / 1e6
    e2e_value = bn.world * paths_per_step * e2e_steps / (worst_e2e_ms / 1e3) / 1e6
    ms_per_step = worst_ms / steps
    fpp = flops_per_path(name)
    step_tflops = fpp * paths_per_step / (ms_per_step / 1e3) / 1e12
    return dict(value=value, e2e_value=e2e_value, e2e_steps=e2e_steps, ms_per_step=ms_per_step,
                step_tflops=step_tflops, fpp=fpp, launches=launches, graphs=graphs, clocks=clocks,
                top=(top_ms, top_units, top_launches), alone=alone, reduce_ms=reduce_ms,
                image_mean=image_mean, host_build_s=host_build_s, upload_s=upload_s, pix=pix,
                per_rank_ms=per_rank_ms,
                size=(wl["size"][0], wl["size"][1]), paths_per_step=paths_per_step)


def strong_leg(bn, name, total_spp):
    """BASELINE.json's split configuration: `total_spp` samples of the whole image divided over
    the ranks (rank r renders sample_ranges[r] from its own stretch of the streams), clock =
    scene upload to the device + first (cold) call + render + the NCCL reduce + ONE download of
    the summed image on rank 0; scene file load and host BVH / light builds are outside."""
    args, b, torch = bn.args, bn.b, bn.torch
    wl = WORKLOADS[name]
    scene, host_build_s = bn.scene(name)
    first, mine = bn.pkg.sharding.sample_ranges(total_spp, bn.world)[bn.rank]
    params = b.make_params(shader=wl["shader"], samples=max(mine, 2), resolution=wl["resolution"],
                           bounces=wl["bounces"], pipeline=args.pipeline)
    r = b.Render(scene, params)
    r.reset(bn.rng_offset)
    bn.barrier()
    t0 = time.time()
    done = 0
    calls = []
    while done < mine:
        n = min(64, mine - done)
        tc = time.time()
        r.trace(n)  # the first call uploads the scene
        calls.append(time.time() - tc)
        done += n
    t_render = time.time() - t0
    bn.reduce_to_rank0(r, total_spp)
    t_reduce = time.time() - t0 - t_render
    if bn.rank == 0:
        img = r.image(normalized=False)
        del img
    torch.cuda.synchronize()
    seconds = time.time() - t0
    # where the clock goes on the slowest rank: first (cold) call, the other calls, reduce (includes
    # waiting for the slowest rank), download
    worst, first_call, other_calls, reduce_s = bn.max_over_ranks(
        [seconds, calls[0], sum(calls[1:]), t_reduce])
    paths = wl["size"][0] * wl["size"][1] * total_spp
    r.close()
    scene.close()
    return {"workload": f"tests/{wl['scene']} {wl['shader']} {wl['size'][0]}x{wl['size'][1]} "
                        f"{total_spp} spp in total",
            "n_gpus": bn.world, "spp_total": total_spp, "spp_this_rank": mine,
            "seconds": worst, "value": paths / worst / 1e6, "unit": "Mpaths/s", "scaling": "strong",
            "breakdown_s": {"first_call": first_call, "other_calls": other_calls, "calls": len(calls),
                            "reduce_and_wait": reduce_s, "download": max(0.0, seconds - t_render - t_reduce)},
            "clock": "scene upload + cold first call + render + NCCL reduce + one D2H of the sum; "
                     "max over ranks (wall clock around blocking calls)"}


def roofline_block(name, m, wl):
    """north_star's compute roofline: FP32 peak / algorithmic flops (HBM and tensor cores do not
    bound this path). Dominant kernel timed live with CUDA events on its own stream."""
    peak = fp32_peak()
    top_ms, top_units, top_launches = m["top"]
    wavefront = extend_flops_per_ray(name) is not None and top_launches > 0 and top_ms > 0 \
        and wl["shader"] in ("volpathtrace", "pathtrace", "naive")
    traffic, traffic_src = ncu_traffic(name) if wavefront else (None, None)
    if wavefront:
        kernel = "k_wf_extend_persistent"
        achieved = extend_flops_per_ray(name) * top_units / (top_ms / 1e3) / 1e12
        per_launch = {"kernel": kernel, "sampled_launches": top_launches,
                      "avg_launch_us": 1e3 * top_ms / top_launches,
                      "avg_rays_per_launch": top_units / top_launches,
                      "grays_per_s": top_units / (top_ms / 1e3) / 1e9,
                      "flops_per_ray": extend_flops_per_ray(name),
                      "algorithmic_bytes_per_ray": extend_bytes_per_ray(name),
                      "note": "timed region: the other wavefront pipe's kernels share the device "
                              "during these launches"}
        if m["alone"] is not None:
            a_ms, a_units, a_n = m["alone"]
            a_tflops = extend_flops_per_ray(name) * a_units / (a_ms / 1e3) / 1e12
            per_launch["alone"] = {"sampled_launches": a_n, "avg_launch_us": 1e3 * a_ms / a_n,
                                   "avg_rays_per_launch": a_units / a_n,
                                   "grays_per_s": a_units / (a_ms / 1e3) / 1e9,
                                   "achieved": a_tflops, "frac": a_tflops / peak,
                                   "note": "one untimed extra call with a single pipe "
                                           "(VPT_WF_PIPES=1): the kernel with the device to itself"}
    else:
        kernel, achieved, per_launch = "whole step", m["step_tflops"], None
    return {"bound": "fp32", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
            "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
            "kernel": kernel, "per_launch": per_launch,
            "whole_step": {"achieved": m["step_tflops"], "frac": m["step_tflops"] / peak,
                           "flops_per_path": m["fpp"]},
            "note": "compute roofline of north_star (FP32 peak / algorithmic flops; HBM and tensor "
                    "cores do not bound this path): algorithmic work = the reference algorithm's, "
                    "from tests/hostsim counters; peak = 148 SM x 128 lanes x 2 x sm_max_mhz of "
                    "MEASURED_PEAKS.json, which holds no fp32 figure; traffic = ncu dram bytes of "
                    "one full-occupancy launch, read from the newest profiles/*brief.txt"}


def gpu_arm(args, wl, rank, world, local_rank):
    bn = Bench(args, rank, world, local_rank)
    m = measure(bn, args.workload, args.steps, args.warmup, full=True)
    line = None
    if rank == 0:
        line = {
            "metric": "Mpaths/s", "value": m["value"], "unit": "Mpaths/s", "n_gpus": world,
            "build": "fast (statistical parity)" if os.path.basename(os.environ.get("VPT_LIB_DIR", "")) == "fast"
                     else "exact (bit-identical to the reference)",
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": m["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": DATA[args.workload], "config": workload_config(args, wl),
            "e2e": {"value": m["e2e_value"], "unit": "Mpaths/s", "h2d_bytes_per_step": m["pix"] * 16,
                    "d2h_bytes_per_step": m["pix"] * 16, "steps": m["e2e_steps"],
                    "includes": "generator upload, trace, " + ("NCCL reduce, " if world > 1 else "") +
                                "accumulation image download"},
            "gpu_launches": m["launches"],
            "host_submissions": {"graph_launches": m["graphs"],
                                 "kernels_inside_graphs": m["graphs"] * 40,
                                 "note": "gpu_launches counts every kernel executed; a batch of 8 "
                                         "iterations (40 kernels) reaches the device as one CUDA "
                                         "graph launch"},
            "clocks": m["clocks"], "roofline": roofline_block(args.workload, m, wl),
            "reduce_ms": m["reduce_ms"], "image_mean": m["image_mean"],
            "per_rank_ms_per_step": m["per_rank_ms"],
            "host_build_s": m["host_build_s"], "scene_upload_s": m["upload_s"],
        }
        if world == 1 and not args.no_cpu_baseline:
            info = run_reference_cpu(wl, 8)
            if info is not None:
                line["cpu_baseline"] = {
                    "value": info["mpaths"], "unit": "Mpaths/s", "cores": info["threads"],
                    "kind": "reference",
                    "sample": f"8 spp of {info['size'][0]}x{info['size'][1]} ({info['paths']} paths, "
                              f"{info['seconds']:.2f} s) by oracle/_ref/ref_render"}
    # ---- the other configurations, short runs (so that the driver's record covers them) ----------
    if not args.no_workloads:
        others = {}
        for name in WORKLOADS:
            if name == args.workload:
                continue
            scene_file = os.path.join(ROOT, "scenes", "_built", WORKLOADS[name]["scene"] + ".vptscene")
            if not os.path.exists(scene_file):
                continue
            w = measure(bn, name, args.side_steps, 1, full=False)
            if rank == 0:
                entry = {"value": w["value"], "unit": "Mpaths/s", "ms_per_step": w["ms_per_step"],
                         "steps": args.side_steps, "e2e": w["e2e_value"],
                         "frac": w["step_tflops"] / fp32_peak(), "achieved_tflops": w["step_tflops"],
                         "gpu_launches": w["launches"], "data": DATA[name],
                         "workload": workload_config(args, WORKLOADS[name])["workload"]}
                if world == 1 and not args.no_cpu_baseline:
                    info = run_reference_cpu(WORKLOADS[name], 2)
                    if info is not None:
                        entry["cpu_baseline"] = {"value": info["mpaths"], "unit": "Mpaths/s",
                                                 "cores": info["threads"], "kind": "reference",
                                                 "sample": f"2 spp, {info['seconds']:.2f} s"}
                others[name] = entry
        if rank == 0:
            line["workloads"] = others
    # ---- the statistical build next to the exact one: what bit-exactness costs ------------------
    fast_dir = os.path.join(ROOT, "volumetric-path-tracer_b200", "fast")
    if rank == 0 and world == 1 and not args.no_fast and "VPT_LIB_DIR" not in os.environ \
            and os.path.exists(os.path.join(fast_dir, "libvpt_cuda.so")):
        # a process holds one build of the libraries: the fast one runs in a child
        cmd = [sys.executable, os.path.abspath(__file__), "--workload", args.workload, "--steps",
               str(min(args.steps, 6)), "--warmup", "3", "--pipeline", args.pipeline, "--no-workloads",
               "--no-strong", "--no-cpu-baseline", "--no-fast"]
        out = subprocess.run(cmd, capture_output=True, text=True, env=dict(os.environ, VPT_LIB_DIR=fast_dir))
        try:
            f = json.loads(out.stdout.strip().splitlines()[-1])
            fr = f["roofline"]
            line["fast"] = {"build": "-DVPT_FAST: FMA contraction, approximate division / square root, "
                                     "CUDA libm, 1-FFMA slab tests; statistical parity only "
                                     "(tests/test_gpu_fast.py)",
                            "value": f["value"], "e2e": f["e2e"]["value"], "ms_per_step": f["ms_per_step"],
                            "steps": f["steps"], "exact_over_fast": line["value"] / f["value"]}
            line["roofline"]["exact"] = {"frac": line["roofline"]["frac"],
                                         "whole_step_frac": line["roofline"]["whole_step"]["frac"]}
            line["roofline"]["fast"] = {"frac": fr["frac"], "achieved": fr["achieved"],
                                        "whole_step_frac": fr["whole_step"]["frac"],
                                        "per_launch": fr["per_launch"]}
        except Exception as ex:  # the headline must not depend on the side run
            line["fast"] = {"unavailable": f"{type(ex).__name__}: {out.stderr[-300:]}"}
    # ---- strong scaling: BASELINE.json's split configurations --------------------------------
    if not args.no_strong:
        strong = {}
        for name, total_spp in STRONG.items():
            scene_file = os.path.join(ROOT, "scenes", "_built", WORKLOADS[name]["scene"] + ".vptscene")
            if os.path.exists(scene_file):
                strong[name] = strong_leg(bn, name, total_spp)
        if rank == 0:
            line["strong"] = strong
    if rank == 0:
        print(json.dumps(line), flush=True)
    if bn.dist is not None:
        bn.dist.barrier()
        bn.dist.destroy_process_group()
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
    ap.add_argument("--no-workloads", action="store_true", help="skip the short runs of the other scenes")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling leg")
    ap.add_argument("--no-fast", action="store_true", help="skip the run of the statistical build")
    ap.add_argument("--side-steps", type=int, default=3)
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
