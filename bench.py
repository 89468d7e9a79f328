#!/usr/bin/env python
"""bench.py — headline benchmark of the galaxy volume integrator (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A "step" is one frame: K1 (ray march) + K2 (post-process) of a 2048x2048 render of Spiral.gax
(arms + dust + stars; 10/9/10 noise octaves hard-coded per class) at the reference step count
(rayStep 0.025f) from the oblique exterior orbit of SURVEY.md §8(d) C2b.  Frame f of the orbit has
azimuth 45deg + f * 1deg (frame 0 is exactly C2b); with N GPUs step s renders frames s*N .. s*N+N-1,
one per rank (weak scaling, no data-path collective).  `value` = ray samples/s (loop iterations of
Rasterizer::getIntensity, counted exactly by the instrumented kernel; same integer as the CPU
oracle) with the scene resident on the device; `e2e` = the same through the host-buffer C-ABI
calls (scene upload, D2H of the float framebuffer and the 8-bit image inside the timed region).
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

METRIC = "ray samples/s (and ms/frame) at 2048x2048"
UNIT = "ray samples/s"
SIZE = 2048
FP32_LANES_PER_SM = 128


def orbit_view(frame_index: int, radius=0.9, elevation=math.asin(1 / math.sqrt(3)), fov=60.0):
    """Camera of orbit frame f: |cam| = 0.9, elevation 35.26deg, azimuth 45deg + f deg. Frame 0 = C2b."""
    if frame_index == 0:
        return ((0.5196, 0.5196, 0.5196), (0, 0, 0), (0, 1, 0), fov)
    az = math.radians(45.0 + frame_index)
    c = radius * math.cos(elevation)
    return ((c * math.cos(az), radius * math.sin(elevation), c * math.sin(az)), (0, 0, 0), (0, 1, 0), fov)


WORKLOADS = {
    # name: (galaxy, size, view or None for the orbit)
    "spiral_2048_oblique_orbit": ("Spiral", 2048, None),
    "spiral_2048_cli_view": ("Spiral", 2048, "cli"),
    "spiral_512_cli_view": ("Spiral", 512, "cli"),
    "sombrero_4096_edge": ("Sombrero", 4096, "edge"),
}


def make_scene(workload: str, frame_index: int, size=None):
    import scenes
    gal, sz, view = WORKLOADS[workload]
    return scenes.single_galaxy_scene(gal, size or sz, orbit_view(frame_index) if view is None else view)


def algorithmic_flop(cnt: dict) -> float:
    """100 per simplex octave + 64 per component evaluation + 32 per sample + 18 per gate."""
    return 100.0 * cnt["octaves"] + 64.0 * cnt["comp_evals"] + 32.0 * cnt["samples"] + 18.0 * cnt["gates"]


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) < 9:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = sorted(s for s in sm if s > 0.5 * max(sm)) if sm else []
        return {"sm_mhz": busy[len(busy) // 2] if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


def cpu_sample_scene(workload: str, row_stride: int):
    """Bounded sample of frame 0 of the workload for the CPU legs: every row_stride-th image row."""
    sc = make_scene(workload, 0)
    n = sc.frame.full_size
    return sc.tile(0, 0, n, n // row_stride, band_rows=1, band_stride=row_stride)


class CpuArm:
    """The CPU side of the comparison on a bounded sample (every row_stride-th row of frame 0 of the workload).
    kind "reference": the reference's OWN render path (rasterizer.cpp, galaxycomponent*.cpp, noise/*.cpp ...)
    compiled in place with its own flags by oracle/Makefile into oracle/_ref/libgamer_ref_path_fast.so
    (-Ofast -fopenmp, OpenMP over pixels like RenderPixelsOMP) — used whenever that library travelled here and
    the host has AVX2. kind "port": the oracle restatement built on this machine with -Ofast -march=native
    -fopenmp. Ray samples are counted by the oracle on the same rows (its march is bit-identical to the
    reference's, tests/test_ref_path.py); the count is outside the timed region."""

    def __init__(self, workload: str, row_stride: int, threads: int = 0, force_port: bool = False):
        import oracle_lib
        import ref_path_lib
        import scenes
        from gamer_b200.model import GalaxyInstance
        self.workload, self.row_stride = workload, row_stride
        self.scene = cpu_sample_scene(workload, row_stride)
        self.oracle = oracle_lib.load(fast=True)
        self.rows, self.size = self.scene.frame.height, self.scene.frame.full_size
        self.ref = None
        self.threads = threads
        if ref_path_lib.available("fast") and not force_port:
            name, size, view = WORKLOADS[workload]
            view = orbit_view(0) if view is None else view
            rp = scenes.rendering_params(size, view)
            gi = GalaxyInstance(scenes.load_galaxy(name), name, (0, 0, 0), (0, 1, 0), 1.0, 0.0)
            self.ref = ref_path_lib.RefRasterizer(rp, [gi], "fast")
            self.ref.render_rows(0, row_stride, 0, 1, prepare=True)      # Rasterizer::Prepare, once, untimed
        self.kind = "reference" if self.ref else "port"
        _, cnt = self.oracle.render(self.scene)
        self.samples = cnt["samples"]
        self.cores = threads or self.oracle.max_threads()

    def step(self) -> float:
        t0 = time.perf_counter()
        if self.ref:
            _, self.cores = self.ref.render_rows(0, self.row_stride, self.rows, self.threads, prepare=False)
        else:
            self.oracle.render(self.scene, self.threads)
        return time.perf_counter() - t0

    def describe(self) -> str:
        what = ("the reference's own code (oracle/_ref/libgamer_ref_path_fast.so: -Ofast -march=x86-64-v3 -fopenmp)"
                if self.ref else "oracle port (-Ofast -march=native -fopenmp)")
        return (f"frame 0 of {self.workload}, every {self.row_stride}th row ({self.rows} of {self.size} rows, "
                f"{self.samples} ray samples), pixel loop only, {what}")

    def close(self):
        if self.ref:
            self.ref.close()


def time_cpu(workload: str, row_stride: int, repeats: int = 1):
    """cpu_baseline leg. Returns (samples/s, cores, sample description, seconds, kind)."""
    arm = CpuArm(workload, row_stride)
    best = min(arm.step() for _ in range(repeats))
    out = (arm.samples / best, arm.cores, arm.describe() + f", best of {repeats}", best, arm.kind, port_value(arm))
    arm.close()
    return out


def port_value(arm: "CpuArm"):
    """When the timed arm is the reference's own code, also time the oracle port on the same sample: the port
    scales linearly with threads, the reference does not (every calculateIntensity call stores m_currentGI into
    component objects all threads share, galaxycomponent.cpp:48 — profiles/r01_cpu_scaling.jsonl), so the port is
    the STRONGER CPU baseline and is reported beside the reference."""
    if arm.kind != "reference":
        return None
    port = CpuArm(arm.workload, arm.row_stride, force_port=True)
    dt = min(port.step() for _ in range(2))
    return {"value": port.samples / dt, "unit": UNIT, "cores": port.cores, "kind": "port"}


# dram__bytes_read.sum + dram__bytes_write.sum of one k1_march launch, from the `ncu --set full` capture
# summarised in profiles/r01_k1_ncu_summary.csv (the 50 MB framebuffer is still in the 126 MB L2 when the
# kernel ends, so DRAM sees ~8 MB; the algorithmic bytes are 12 B/pixel = 50 MB)
K1_DRAM_TRAFFIC = {"spiral_2048_oblique_orbit": 74752 + 7794432, "spiral_2048_cli_view": 87552 + 9464832}
# the same for one k2_tonemap_v4 launch at 2048^2 (profiles/r01_k2_ncu_summary.csv): the 50 MB frame is read from DRAM,
# the two 16.8 MB images it writes are still in L2 when it ends (algorithmic bytes: 84 MB)
K2_DRAM_TRAFFIC = {2048: 50363648 + 3948032}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from gamer_b200.rasterizer import Context

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ctx = Context([local])
    K, Wm = args.steps, args.warmup
    wl = args.workload
    size = WORKLOADS[wl][1]
    tiles = args.mode == "tiles"
    if tiles:       # strong scaling: every rank renders its interleaved bands of the SAME frame, then one gather
        from gamer_b200 import parallel
        full = [make_scene(wl, s) for s in range(Wm + K)]
        scenes_ = [parallel.band_tile(sc, rank, world) for sc in full]
    else:           # weak scaling: step s renders frames s*world .. s*world+world-1, one per rank
        frames = [s * world + rank for s in range(Wm + K)]
        scenes_ = [make_scene(wl, f) for f in frames]

    # exact work of every timed frame (instrumented kernel, outside the timed region)
    counts = []
    for sc in scenes_[Wm:]:
        ctx.upload(sc)
        counts.append(ctx.render_counted())
    my_samples = sum(c["samples"] for c in counts)
    my_flop = sum(algorithmic_flop(c) for c in counts)

    stream = torch.cuda.ExternalStream(ctx.stream(0), device=torch.device("cuda", local))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")       # > 126 MB L2
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]

    gathered = [None, None]

    def step_device(i, timed_idx=None):
        ctx.upload(scenes_[i])                      # host-side: scene -> kernel parameter block
        with torch.cuda.stream(stream):
            flush.zero_()
            if timed_idx is not None:
                ev[timed_idx][0].record(stream)
            ctx.render_device()                     # K1
            ctx.tonemap_device(1.0, 1.0, 1.0)       # K2
            if tiles and world > 1:                 # K3: the one exchange step — gather the disjoint bands
                rgb_t, _, img_t = ctx.device_tensors()
                gathered[0] = parallel.gather_bands(rgb_t, size, rank, world)
                gathered[1] = parallel.gather_bands(img_t, size, rank, world)
            if timed_idx is not None:
                ev[timed_idx][1].record(stream)

    for i in range(Wm):
        step_device(i)
    clocks = ClockSampler(local)
    barrier()
    if rank == 0:
        clocks.start()
    l0 = ctx.launches()
    k1_ms, k2_ms = [], []
    t_wall0 = time.perf_counter()
    for k in range(K):
        step_device(Wm + k, k)
        a, b = ctx.kernel_ms()
        k1_ms.append(a)
        k2_ms.append(b)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = ctx.launches() - l0
    clk = clocks.stop() if rank == 0 else None
    dev_ms = sum(e0.elapsed_time(e1) for e0, e1 in ev)

    # ---- e2e: host buffers through the C ABI (upload + K1 + D2H float frame + K2 + D2H 8-bit image) ----
    host_rgb = torch.empty((size, size, 3), dtype=torch.float32, pin_memory=True)
    host_img = torch.empty((size, size), dtype=torch.int32, pin_memory=True)

    def step_e2e(i):
        ctx.upload(scenes_[i])
        if tiles and world > 1:                     # bands stay on the device, gather over NVLink, rank 0 reads the frame
            step_device(i)
            if rank == 0:
                with torch.cuda.stream(stream):
                    host_rgb.copy_(gathered[0], non_blocking=True)
                    host_img.copy_(gathered[1], non_blocking=True)
            stream.synchronize()
        else:
            ctx.render_ptr(host_rgb.data_ptr())
            ctx.tonemap_ptr(1.0, 1.0, 1.0, None, host_img.data_ptr())

    # Two independent contexts, one host thread each, take alternate frames: the blocking calls of one
    # (D2H of frame k) overlap the kernels of the other (frame k+1). Every frame still goes through
    # upload -> render -> tonemap with its own host<->device copies inside the timed region.
    pipelined = not (tiles and world > 1) and args.e2e_contexts > 1
    if pipelined:
        ctxs = [ctx] + [Context([local]) for _ in range(args.e2e_contexts - 1)]
        bufs = [(host_rgb, host_img)] + [(torch.empty_like(host_rgb).pin_memory(), torch.empty_like(host_img).pin_memory())
                                         for _ in range(args.e2e_contexts - 1)]

        def worker(w, first, count):
            c, (hr, hi) = ctxs[w], bufs[w]
            for i in range(first + w, first + count, len(ctxs)):
                c.upload(scenes_[i])
                c.render_ptr(hr.data_ptr())
                c.tonemap_ptr(1.0, 1.0, 1.0, None, hi.data_ptr())

        def run_e2e(first, count):
            th = [threading.Thread(target=worker, args=(w, first, count)) for w in range(len(ctxs))]
            for t_ in th:
                t_.start()
            for t_ in th:
                t_.join()
    else:
        def run_e2e(first, count):
            for i in range(first, first + count):
                step_e2e(i)

    run_e2e(0, min(Wm, 3))
    barrier()
    t0 = time.perf_counter()
    run_e2e(Wm, K)
    barrier()
    e2e_s = time.perf_counter() - t0
    checksum = float(host_rgb[::64, ::64].double().sum())
    if pipelined:
        for c in ctxs[1:]:
            c.close()

    # ---- K2 alone, cold L2, for its HBM roofline ----
    k2_cold = []
    for _ in range(5):
        with torch.cuda.stream(stream):
            flush.zero_()
        ctx.tonemap_device(1.0, 1.0, 1.0)
        k2_cold.append(ctx.kernel_ms()[1])

    # ---- reduce over ranks: time = max, work = sum ----
    t = torch.tensor([dev_ms, e2e_s * 1e3, t_wall * 1e3], dtype=torch.float64, device="cuda")
    w = torch.tensor([float(my_samples), my_flop, float(launches)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(w, op=dist.ReduceOp.SUM)
    dev_ms_max, e2e_ms_max, wall_ms_max = t.tolist()
    tot_samples, tot_flop, tot_launches = w.tolist()

    if rank == 0:
        peaks, peak_src = measured_peaks()
        sm_count = torch.cuda.get_device_properties(local).multi_processor_count
        sm_max = float(peaks.get("sm_max_mhz", 1965.0))
        fp32_peak = sm_count * FP32_LANES_PER_SM * 2 * sm_max * 1e6 / 1e12
        k1_avg_ms = sum(k1_ms) / len(k1_ms)
        k1_tflops = (my_flop / K) / (k1_avg_ms * 1e-3) / 1e12
        px = size * size
        k2_best = min(k2_cold)
        k2_bytes = px * (12 + 4 + 4)
        out = {
            "metric": METRIC, "value": tot_samples / (dev_ms_max * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": K, "warmup": Wm, "ms_per_step": dev_ms_max / K, "higher_is_better": True, "scaling": "strong" if tiles else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl, "mode": args.mode, "galaxy": WORKLOADS[wl][0] + ".gax (tests/golden/gax)", "size": size,
                       "ray_step": "0.025f", "frames_per_step": 1 if tiles else world,
                       "parallelism": (f"interleaved {parallel.BAND}-row bands x{world} + NCCL gather to rank 0" if tiles
                                       else f"frames x{world}"),
                       "l2": "256 MB device buffer zeroed between timed iterations (outside the event bracket); "
                             "inputs are a few KB of scene parameters",
                       "ray_samples_per_frame": my_samples / K, "octaves_per_frame": sum(c["octaves"] for c in counts) / K},
            "gpu_launches": int(tot_launches),
            "e2e": {"value": tot_samples / (e2e_ms_max * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": world * ctx_scene_bytes(scenes_[0]), "d2h_bytes_per_step": (1 if tiles else world) * px * 16,
                    "ms_per_step": e2e_ms_max / K, "checksum": checksum,
                    "contexts": args.e2e_contexts if pipelined else 1},
            "roofline": {"kernel": "k1_march", "bound": "fp32", "achieved": k1_tflops, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": k1_tflops / fp32_peak, "traffic": K1_DRAM_TRAFFIC.get(wl),
                         "peak_source": f"{sm_count} SMs x 128 FP32 lanes x 2 flop x {sm_max:.0f} MHz (sm_max_mhz of "
                                        f"MEASURED_PEAKS.json, {peak_src}); the path is FP32/SFU-pipe bound, not HBM or tensor",
                         "algorithmic_flop_per_launch": my_flop / K, "launch_ms": k1_avg_ms},
            "roofline_k2": {"kernel": "k2_tonemap_v4", "bound": "hbm", "achieved": k2_bytes / (k2_best * 1e-3) / 1e9,
                            "peak": float(peaks.get("hbm_gbs", 6650.0)), "unit": "GB/s",
                            "frac": k2_bytes / (k2_best * 1e-3) / 1e9 / float(peaks.get("hbm_gbs", 6650.0)),
                            "traffic": K2_DRAM_TRAFFIC.get(size) if args.mode == "frames" else None, "algorithmic_bytes_per_launch": k2_bytes, "launch_ms": k2_best,
                            "peak_source": f"hbm_gbs of MEASURED_PEAKS.json ({peak_src})"},
            "kernel_ms": {"k1_avg": k1_avg_ms, "k2_avg_in_step": sum(k2_ms) / len(k2_ms), "k2_cold_l2": k2_best,
                          "wall_ms_per_step_incl_flush": wall_ms_max / K},
            "clocks": clk,
        }
        if world == 1 and not args.no_cpu:
            v, cores, desc, secs, kind, port = time_cpu(wl, args.cpu_row_stride)
            out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": desc, "seconds": secs}
            if port:
                out["cpu_baseline"]["oracle_port"] = port
        print(json.dumps(out), flush=True)
    # torch tensors that were used on the context's stream must be released while that stream still exists
    gathered[:] = [None, None]
    del flush, host_rgb, host_img
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()


def ctx_scene_bytes(scene) -> int:
    import ctypes as C
    return C.sizeof(scene.frame) + C.sizeof(scene.galaxies) + C.sizeof(scene.components)


def run_reference(args):
    """Reference arm: the reference's CPU implementation of the path on the host cores (class CpuArm: the
    reference's own sources compiled in place when oracle/_ref travelled here, else the oracle port), all host
    threads, each step a bounded sample of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    wl = args.workload
    arm = CpuArm(wl, args.ref_row_stride)
    sc, samples = arm.scene, arm.samples
    times = []
    for i in range(args.warmup + args.steps):
        dt = arm.step()
        if i >= args.warmup:
            times.append(dt)
    total = sum(times)
    value = samples * len(times) / total
    size = WORKLOADS[wl][1]
    frac = sc.frame.height / sc.frame.full_size
    desc = f"each step = {arm.describe()}; a whole frame is ~{1 / frac:.0f}x that"
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f64/f32 (reference mix)", "data": "synthetic",
           "config": {"workload": wl, "galaxy": WORKLOADS[wl][0] + ".gax (tests/golden/gax)", "size": size, "ray_step": "0.025f",
                      "sample": desc, "est_ms_per_full_frame": 1e3 * total / len(times) / frac},
           "cpu_baseline": {"value": value, "unit": UNIT, "cores": arm.cores, "kind": arm.kind, "sample": desc,
                            **({"oracle_port": port_value(arm)} if arm.kind == "reference" else {})},
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="spiral_2048_oblique_orbit", choices=list(WORKLOADS))
    ap.add_argument("--mode", default="frames", choices=["frames", "tiles"],
                    help="N>1: frames = one frame per rank per step (weak, no collective); tiles = one frame split "
                         "into interleaved row bands + NCCL gather (strong)")
    ap.add_argument("--e2e-contexts", type=int, default=3,
                    help="independent gamer_ctx (one host thread each) the e2e leg alternates frames over; 1 = strictly serial")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-row-stride", type=int, default=16, help="cpu_baseline sample: every n-th row")
    ap.add_argument("--ref-row-stride", type=int, default=64, help="--impl reference step: every n-th row")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:           # convenience: relaunch under torchrun
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
                   "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
            sys.exit(subprocess.call(cmd))
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if args.warmup < 3:
        args.warmup = 3
    run_ours(args)


if __name__ == "__main__":
    main()
