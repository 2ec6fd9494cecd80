#!/usr/bin/env python
"""bench.py — Mrays/s of the render hot path on N B200s (one process per GPU).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload benchmark|simple|mesh_hdri|volumes]
  python bench.py --impl reference ...     # the CPU restatement on the box's host cores

A step = one render() of the workload (BASELINE.json configs[1] by default: scenes/benchmark.json
at 1920x1080, 1024 spp).  `value` = rays/s with the scene resident in HBM (render kernels + the
cross-rank reduce + resolve, CUDA events on the launching stream, max over ranks).  `e2e` = the
same metric through the reference-facing call: scene upload (H2D) + render + RGBA8 D2H with host
buffers, wall clock around synchronised calls.  N>1: every rank renders its own spp-sized sample
slice (weak scaling), one NCCL reduce of the fp32 accumulation buffers, resolve on rank 0.
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (scene recipe, W, H, spp, description)
    "benchmark": ("benchmark", 1920, 1080, 1024, "scenes/benchmark.json 1920x1080x1024spp (5 spheres, BVH off as shipped)"),
    "benchmark_bvh": ("benchmark_bvh", 1920, 1080, 1024, "scenes/benchmark.json with enable_bvh_tree forced on, 1920x1080x1024spp (SURVEY C2: both settings reported)"),
    "simple": ("simple", 1920, 1080, 1024, "scenes/simple.json 1920x1080x1024spp (102 spheres, BVH on)"),
    "mesh_hdri": ("mesh_hdri", 1920, 1080, 256, "house+plane+fancy_monkey meshes, synthetic 4096x2048 HDRI, 1920x1080x256spp"),
    "volumes": ("volumes", 1920, 1080, 64, "mesh/sphere-bounded media over plane, synthetic HDRI, 1920x1080x64spp"),
    "bvh_stress": ("bvh_stress", 2048, 2048, 16, "C3(ii): 1M random spheres (seed 42), pinhole 2048x2048 from (0,0,120), Lambertian, depth 6, 16spp"),
}


def make_scene(name):
    import _scenes
    if name == "mesh_hdri":
        return _scenes.mesh_hdri(4096, 2048)
    if name == "volumes":
        return _scenes.volumes(2048, 1024)
    if name == "bvh_stress":
        return _scenes.random_spheres(1_000_000)
    if name == "benchmark_bvh":
        s = _scenes.benchmark()
        s.settings.enable_bvh_tree = True
        return s
    return getattr(_scenes, name)()


def algorithmic_ops(st):
    """SURVEY §8(d) issue-slot roofline: FP32 work per unit with 1/d hoisted — slab test 21 ops, Moeller-Trumbore 51,
    sphere 33, env lookup 40 (normalize + acos + atan2), and 150 per ray for the path bookkeeping + scatter
    (~2 Philox4x32-10 draws)."""
    return 21 * st["nodes_visited"] + 51 * st["tri_tests"] + 33 * st["sphere_tests"] + 40 * st["env_lookups"] + 150 * st["rays"]


def algorithmic_bytes(st):
    """32/box tested + 48/triangle test + 16/sphere test + 16/env texel + 32/ray."""
    return 32 * st["nodes_visited"] + 48 * st["tri_tests"] + 16 * st["sphere_tests"] + 16 * st["env_lookups"] + 32 * st["rays"]


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._stop_evt = index, [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.reasons |= {n for bit, n in names.items() if r & bit}
            except Exception:
                pass
            time.sleep(0.05)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# bounded CPU sample per workload (W, H, spp factor): the reference-shaped CPU walk tests every pierced subtree, which
# on the 1 M-sphere scene is ~150x slower per ray than on the others — keep each sample to seconds of CPU work
# (16 host cores: benchmark ~6 s, simple ~4 s, mesh_hdri ~5 s, volumes ~9 s, bvh_stress ~1 s per sample)
CPU_SAMPLE = {"bvh_stress": (256, 256, 1 / 16.0), "benchmark": (1920, 1080, 16.0), "benchmark_bvh": (1920, 1080, 16.0),
              "simple": (1920, 1080, 4.0)}


def cpu_sample_for(workload, W, H, spp_sample):
    w, h, f = CPU_SAMPLE.get(workload, (W, H, 1.0))
    return w, h, max(1, int(round(spp_sample * f)))


def cpu_render(scene, W, H, spp_sample, seed=0, count_work=False):
    """Oracle (CPU restatement) render of a bounded sample of the workload, all host threads."""
    import raytracingrust_b200 as rt
    from oracle import orc
    o = orc.OracleScene(scene)
    p = rt.make_params(scene, W, H, seed=seed, sample_count=spp_sample)
    # all host cores this process may run on (torchrun exports OMP_NUM_THREADS=1, so do not rely on the OpenMP default)
    threads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    _, _, st = o.render(p, threads=threads, count_work=count_work, want_rgba=True, want_accum=False)
    return st


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name, W, H, spp, desc = WORKLOADS[args.workload]
    scene = make_scene(name)
    W, H, spp_sample = cpu_sample_for(args.workload, W, H, args.cpu_spp)
    for _ in range(args.warmup):
        cpu_render(scene, W, H, 1)
    t_ms, rays, samples, threads = 0.0, 0, 0, 1
    for k in range(args.steps):
        st = cpu_render(scene, W, H, spp_sample, seed=k)
        t_ms += st["device_ms"]; rays += st["rays"]; samples += st["samples"]; threads = st["kernel_launches"]
    v = rays / (t_ms * 1e-3) / 1e6
    sample = f"{W}x{H} at {spp_sample} spp of {spp} per step (rates are spp-independent)"
    print(json.dumps({
        "impl": "reference", "metric": "Mrays/s", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "msamples_per_s": samples / (t_ms * 1e-3) / 1e6,
        "config": {"workload": desc, "note": "CPU restatement of the reference (oracle port, OpenMP over pixels); "
                   "the Rust crate cannot be built here (no cargo/rustc)"},
        "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="benchmark", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel (debug only)")
    ap.add_argument("--cpu-spp", type=int, default=16, help="spp of the bounded CPU sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--kernel", type=int, default=0)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: every rank renders the workload's spp (image has N x spp samples); strong: the workload's "
                         "spp is split across ranks (BASELINE config 5 sweep)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import raytracingrust_b200 as rt
    from raytracingrust_b200 import _abi as A

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    name, W, H, spp, desc = WORKLOADS[args.workload]
    if args.spp:
        spp = args.spp
        desc += f" [spp overridden to {spp}]"
    scene = make_scene(name)
    scene.settings.samples_per_pixel = spp

    # a real (non-NULL) stream shared by torch (events, NCCL) and the library's launches
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    ctx = rt.Context(local, stream=ctypes.c_void_p(stream.cuda_stream))
    dev_scene = ctx.upload(scene)
    info = dev_scene.info()
    accum = torch.empty((H, W, 4), dtype=torch.float32, device="cuda")
    rgba = torch.empty((H, W, 4), dtype=torch.uint8, device="cuda")
    # weak scaling: rank r renders global samples [r*spp, (r+1)*spp); the image has world*spp samples.
    # strong scaling: the image keeps spp samples, rank r renders its balanced slice of them.
    from raytracingrust_b200.distributed import sample_slice
    total_spp = spp * world if args.scaling == "weak" else spp
    s_begin, s_count = (rank * spp, spp) if args.scaling == "weak" else sample_slice(spp, rank, world)
    params = rt.make_params(scene, W, H, seed=0, sample_begin=s_begin, sample_count=s_count, kernel=args.kernel)

    def step():
        dev_scene.render_accumulate_device(params, accum.data_ptr(), want_stats=False)
        if world > 1:
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)  # one NCCL reduce over NVLink
        if rank == 0:
            ctx.resolve_device(accum.data_ptr(), W, H, total_spp, rgba.data_ptr())

    # rays per step and work counters (untimed, counting variant of the kernel)
    pc = rt.make_params(scene, W, H, seed=0, sample_begin=s_begin, sample_count=s_count, kernel=args.kernel,
                        flags=A.FLAG_COUNT_WORK | A.FLAG_REFERENCE_TREE)  # algorithmic work is DEFINED on the reference topology
    st_count = dev_scene.render_accumulate_device(pc, accum.data_ptr(), want_stats=True)
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    ev0.record(stream)
    for _ in range(args.steps):
        step()
    ev1.record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()
    ms = ev0.elapsed_time(ev1)
    # dominant kernel alone (megakernel launches only), CUDA events inside the library, same stream
    kern_ms = []
    for _ in range(args.steps):
        kern_ms.append(dev_scene.render_accumulate_device(params, accum.data_ptr(), want_stats=True)["device_ms"])
    t = torch.tensor([ms, float(st_count["rays"])], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms, rays_step = float(tmax[0]), float(tsum[1])
    else:
        rays_step = float(t[1])
    samples_step = float(W) * H * total_spp

    # end to end through the reference-facing call: upload (H2D) + render + RGBA8 D2H, host buffers
    host_rgba = torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()
    lowered = scene.lower()
    pe = rt.make_params(scene, W, H, seed=0, sample_begin=s_begin, sample_count=s_count, kernel=args.kernel)
    e2e_steps = max(1, min(args.steps, 3))
    ds = ctx.upload(lowered)  # one untimed warm-up pass of the same call sequence (grows the library's memory pool once)
    ds.render_into(pe, host_rgba.data_ptr())
    ds.close()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ds = ctx.upload(lowered)
        ds.render_into(pe, host_rgba.data_ptr())
        ds.close()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te[0])

    # which of our kernels the megakernel selector resolved to (capi.cpp render_into): the all-sphere BVH-off scan
    # kernel (csrc/scan.cu) when every object is a sphere and the BVH is off, as in scenes/benchmark.json
    used = {1: "megakernel", 2: "wavefront"}[st_count["kernel_used"]]
    kernel_fn = "render_megakernel" if used == "megakernel" else "wavefront wave sequence (wf_extend dominant)"
    if (used == "megakernel" and not scene.settings.enable_bvh_tree and scene.settings.max_ray_depth >= 1
            and os.environ.get("RTR_SCAN_KERNEL", "1") != "0" and all(isinstance(o, rt.Sphere) for o in scene.objects)):
        used, kernel_fn = "megakernel (all-sphere scan kernel)", "render_scan_kernel"
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak, peak_src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
        k_ms = float(np.mean(kern_ms))
        sm_count = torch.cuda.get_device_properties(local).multi_processor_count
        alg = algorithmic_bytes(st_count)
        achieved = alg / (k_ms * 1e-3) / 1e9
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload)
        except Exception:
            pass
        out = {
            "metric": "Mrays/s", "value": rays_step * args.steps / (ms * 1e-3) / 1e6, "unit": "Mrays/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "msamples_per_s": samples_step * args.steps / (ms * 1e-3) / 1e6,
            "rays_per_step": rays_step, "samples_per_step": samples_step,
            "config": {"workload": desc, "kernel": used, "per_gpu_spp": s_count, "image_spp": total_spp, "parallelism": f"sample-slices x{world} + ncclReduce",
                       "l2_policy": "each step rewrites the 33 MB accumulation buffer and traces fresh Philox paths; scene working set is L2-resident by design (SURVEY 8d)",
                       "scene_device_bytes": info["device_bytes"], "bvh_nodes": info["n_nodes"]},
            "clocks": clocks,
            "e2e": {"value": rays_step * e2e_steps / e2e_s / 1e6, "unit": "Mrays/s",
                    "h2d_bytes_per_step": info["h2d_bytes"], "d2h_bytes_per_step": W * H * 4 + 64, "steps": e2e_steps},
            # our kernels per step: the render kernels (megakernel, or the wavefront's per-wave sequence) + resolve_kernel
            # (NCCL's reduce kernel is not ours)
            "gpu_launches": args.steps * (st_count["kernel_launches"] + 1),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src, "kernel": kernel_fn,
                         "kernel_ms": k_ms, "algorithmic_bytes_per_launch": alg,
                         "bytes_per_ray": alg / max(st_count["rays"], 1),
                         "traffic_note": "dram__bytes_read+write per launch of the dominant kernel from profiles/ (ncu --set full); "
                                         "far BELOW the algorithmic bytes because the scene is L2/L1-resident",
                         # secondary, physically binding roofline: thread-instructions
                         "issue": {"achieved": algorithmic_ops(st_count) / (k_ms * 1e-3) / 1e12, "unit": "T thread-op/s",
                                   "peak": sm_count * 4 * 32 * (clocks["sm_mhz"] or 1965.0) * 1e6 / 1e12,
                                   "frac": algorithmic_ops(st_count) / (k_ms * 1e-3) / (sm_count * 4 * 32 * (clocks["sm_mhz"] or 1965.0) * 1e6)},
                         "note": "scene is L2-resident: the binding units are the L1 data pipe (divergent node fetches) and SM "
                                 "issue slots under SIMT divergence, not HBM (ncu: DESIGN.md 6, profiles/)"},
        }
        if not args.no_cpu:
            cw, ch, cs = cpu_sample_for(args.workload, W, H, args.cpu_spp)
            st = cpu_render(scene, cw, ch, cs)
            out["cpu_baseline"] = {"value": st["rays"] / (st["device_ms"] * 1e-3) / 1e6, "unit": "Mrays/s",
                                   "cores": st["kernel_launches"], "kind": "port",
                                   "sample": f"{cw}x{ch} at {cs} spp (oracle port, OpenMP, all host threads)"}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
