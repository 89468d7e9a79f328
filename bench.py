#!/usr/bin/env python
"""bench.py — headline benchmark of the galaxy volume integrator (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME] [--mode frames|tiles]

A "step" is one frame: K1 (ray march) + K2 (post-process) of a 2048x2048 render of Spiral.gax
(arms + dust + stars; 10/9/10 noise octaves hard-coded per class) at the reference step count
(rayStep 0.025f) from the oblique exterior orbit of SURVEY.md §8(d) C2b.  Frame f of the orbit has
azimuth 45deg + f * 1deg (frame 0 is exactly C2b); with N GPUs step s renders frames s*N .. s*N+N-1,
one per rank (weak scaling, no data-path collective).  `value` = ray samples/s (loop iterations of
Rasterizer::getIntensity, counted exactly by the instrumented kernel; same integer as the CPU
oracle) with the scene resident on the device; `e2e` = the same through the host-buffer C-ABI
calls (scene upload, D2H of the float framebuffer and the 8-bit image inside the timed region).

Beside the headline the line carries, under "also", the other BASELINE configs as measured workloads:
  tiles_sombrero_4096   C3: one 4096^2 edge-on frame split into interleaved 32-row bands over the ranks + the final gather
  sequence_1920x1080    C4: the 360-frame rotating-camera sequence, frame k on rank k mod N
  batch_256_1024        C5: 256 seeded random galaxies at 1024^2, longest-job-first over the ranks
and, at N = 1, "parity": the rows the cpu_baseline leg renders with the reference's own code compared with the same
rows of the GPU frame, and the C2a in-disk view ("cli_view").
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
FP32_LANES_PER_SM = 128
RTOL, ATOL = 1e-3, 1e-5          # the parity bar on linear intensity (tests/test_gpu_parity.py)


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
L2_NOTE = ("GPU arm: a 256 MB device buffer is zeroed between timed iterations (outside the event bracket), inputs are a few KB "
           "of scene parameters; CPU arm: not applicable")


def make_scene(workload: str, frame_index: int, size=None):
    import scenes
    gal, sz, view = WORKLOADS[workload]
    return scenes.single_galaxy_scene(gal, size or sz, orbit_view(frame_index) if view is None else view)


def workload_config(wl: str, mode: str, world: int) -> dict:
    """The `config` object: what is rendered and how it is partitioned. Built the same way by both arms."""
    from gamer_b200 import parallel
    tiles = mode == "tiles"
    return {"workload": wl, "mode": mode, "galaxy": WORKLOADS[wl][0] + ".gax (tests/golden/gax)", "size": WORKLOADS[wl][1],
            "ray_step": "0.025f", "frames_per_step": 1 if tiles else world,
            "parallelism": (f"interleaved {parallel.BAND}-row bands x{world} + gather to rank 0" if tiles else f"frames x{world}"),
            "l2": L2_NOTE}


def algorithmic_flop(cnt: dict) -> float:
    """100 per simplex octave + 64 per component evaluation + 32 per sample + 18 per gate."""
    return 100.0 * cnt["octaves"] + 64.0 * cnt["comp_evals"] + 32.0 * cnt["samples"] + 18.0 * cnt["gates"]


def host_threads() -> int:
    """The host cores this process may use. Asked for explicitly everywhere: under torch.distributed.run
    OMP_NUM_THREADS is 1, which silently reduced the reference arm of round 1 to one thread at N > 1."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


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
    reference's, tests/test_ref_path.py); the count is outside the timed region.
    Nothing here loads libgamer_b200.so: the scene's camera matrix and galaxy rotation come from the oracle's host math."""

    def __init__(self, workload: str, row_stride: int, threads: int = 0, force_port: bool = False):
        import oracle_lib
        import ref_path_lib
        from gamer_b200 import model
        self.oracle = oracle_lib.load(fast=True)
        model.set_host_math(self.oracle.camera_inv_vp, self.oracle.rotation_to)
        try:
            import scenes
            self.workload, self.row_stride = workload, row_stride
            self.scene = cpu_sample_scene(workload, row_stride)
            self.rows, self.size = self.scene.frame.height, self.scene.frame.full_size
            self.ref = None
            self.threads = threads or host_threads()
            if ref_path_lib.available("fast") and not force_port:
                name, size, view = WORKLOADS[workload]
                view = orbit_view(0) if view is None else view
                rp = scenes.rendering_params(size, view)
                gi = model.GalaxyInstance(scenes.load_galaxy(name), name, (0, 0, 0), (0, 1, 0), 1.0, 0.0)
                self.ref = ref_path_lib.RefRasterizer(rp, [gi], "fast")
                self.ref.render_rows(0, row_stride, 0, 1, prepare=True)      # Rasterizer::Prepare, once, untimed
        finally:
            model.set_host_math()
        self.kind = "reference" if self.ref else "port"
        # the exact work and the parity pixels come from the STRICT oracle build (IEEE arithmetic, bit-equal to the
        # reference's own sources on the golden frames); the timed builds use the reference's -Ofast flags, whose
        # fast-math changes roundings (and drops the `intensity > 0.001` test, see gamer_b200.h) — a few samples in 7e7
        self.strict_pixels, cnt = oracle_lib.load().render(self.scene, self.threads)
        self.counts = cnt
        self.samples = cnt["samples"]
        self.cores = self.threads

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
                f"{self.samples} ray samples), pixel loop only, {self.cores} threads, {what}")

    def close(self):
        if self.ref:
            self.ref.close()


def port_value(arm: "CpuArm"):
    """When the timed arm is the reference's own code, also time the oracle port on the same sample: the port
    scales linearly with threads, the reference does not (every calculateIntensity call stores m_currentGI into
    component objects all threads share, galaxycomponent.cpp:48 — profiles/r01_cpu_scaling.jsonl), so the port is
    the STRONGER CPU baseline and is reported beside the reference."""
    if arm.kind != "reference":
        return None
    port = CpuArm(arm.workload, arm.row_stride, arm.threads, force_port=True)
    dt = min(port.step() for _ in range(2))
    return {"value": port.samples / dt, "unit": UNIT, "cores": port.cores, "kind": "port"}


def parity_against_cpu_rows(ctx, arm: "CpuArm") -> dict:
    """Parity AT THE BENCHMARKED SIZE: the rows of the cpu_baseline sample rendered by the strict CPU oracle (bit-equal
    to the reference's own sources on the golden frames, tests/test_ref_path.py) against the same rows of the GPU frame,
    through the C ABI; the GPU's exact counters for those rows against the oracle's; and the reference harness's own
    criterion — identical 8-bit values (tools/galaxy_repro.py:988-995) — as the fraction of equal channel values."""
    ctx.upload(arm.scene)
    cnt = ctx.render_counted()
    gpu = ctx.render()
    _, gpu_img = ctx.tonemap(1.0, 1.0, 1.0)
    cpu = arm.strict_pixels.reshape(gpu.shape)
    _, cpu_img = arm.oracle.tonemap(cpu)
    err = np.abs(gpu.astype(np.float64) - cpu) / (RTOL * np.abs(cpu.astype(np.float64)) + ATOL)
    g8 = gpu_img.view(np.uint8).reshape(-1, 4)[:, :3]
    c8 = cpu_img.view(np.uint8).reshape(-1, 4)[:, :3]
    return {"against": "CPU oracle, strict IEEE build (oracle/gamer_oracle.cpp)", "rows": int(arm.rows), "row_stride": int(arm.row_stride), "size": int(arm.size),
            "max_err_over_bound": float(err.max()), "bound": f"|gpu-cpu| <= {RTOL}*|cpu| + {ATOL} per channel",
            "samples_equal": bool(cnt["samples"] == arm.counts["samples"] and cnt["rays"] == arm.counts["rays"]
                                  and cnt["gates"] == arm.counts["gates"]),
            "ray_samples": int(cnt["samples"]), "eight_bit_equal_frac": float((g8 == c8).mean()),
            "eight_bit_max_diff": int(np.abs(g8.astype(np.int16) - c8.astype(np.int16)).max())}


class SharedHostFrame:
    """One page-locked host buffer that every rank of the node maps (a file in /dev/shm, cudaHostRegister'ed by each
    process): the destination of gamer_b200_copy_bands, so the tile partition's frame is assembled on the host by all
    ranks in parallel, each over its own PCIe link, instead of one rank copying the gathered 268 MB alone."""

    def __init__(self, tag: str, nbytes: int, rank: int, barrier):
        import torch
        self.rank, self.barrier, self.nbytes = rank, barrier, nbytes
        self.path = f"/dev/shm/gamer_b200_{tag}_{os.environ.get('MASTER_PORT', str(os.getppid()))}"
        if rank == 0:
            with open(self.path, "wb") as fh:
                fh.truncate(nbytes)
        barrier()
        self.mm = np.memmap(self.path, dtype=np.uint8, mode="r+", shape=(nbytes,))
        self.ptr = self.mm.ctypes.data
        rc = torch.cuda.cudart().cudaHostRegister(self.ptr, nbytes, 0)
        self.registered = int(rc) == 0

    def close(self):
        import torch
        if self.registered:
            torch.cuda.cudart().cudaHostUnregister(self.ptr)
        del self.mm
        self.barrier()
        if self.rank == 0 and os.path.exists(self.path):
            os.unlink(self.path)


def run_jobs(ctx, jobs, my_ids, rank, world, barrier, reduce_max, reduce_sum, unit_name, e2e_contexts=3, queue=None):
    """A list of independent render jobs (frames of a sequence, galaxies of a batch). Static: this rank renders jobs[i]
    for i in my_ids, in that order. With `queue` (parallel.JobQueue, one shared counter for all ranks) the device leg takes
    its jobs from the queue instead - a rank asks for the next one when its GPU has finished the last (render_device
    returns after the kernel) - and my_ids only says whose exact work this rank COUNTS beforehand (the totals are summed
    over the ranks either way); the e2e leg then replays the jobs this rank ended up with. No collective in the data path.
    Device time over the whole list with CUDA events on the context's stream (max over ranks); e2e through host buffers,
    frames alternating over e2e_contexts contexts."""
    import torch
    from gamer_b200.rasterizer import Context
    mine = [jobs[i] for i in my_ids]
    counts = {"samples": 0, "octaves": 0}
    for sc in mine:                                   # exact work, instrumented kernel, untimed
        ctx.upload(sc)
        c = ctx.render_counted()
        counts["samples"] += c["samples"]
        counts["octaves"] += c["octaves"]
    stream = torch.cuda.ExternalStream(ctx.stream(0), device=torch.device("cuda", torch.cuda.current_device()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for sc in mine[:3]:                               # warm-up
        ctx.upload(sc); ctx.render_device(); ctx.tonemap_device(1.0, 1.0, 1.0)
    barrier()
    l0 = ctx.launches()
    e0.record(stream)
    if queue is None:
        for sc in mine:
            ctx.upload(sc)
            ctx.render_device()
            ctx.tonemap_device(1.0, 1.0, 1.0)
    else:
        taken = []
        for j in queue:
            taken.append(j)
            ctx.upload(jobs[j])
            ctx.render_device()
            ctx.tonemap_device(1.0, 1.0, 1.0)
        mine = [jobs[j] for j in taken]
    e1.record(stream)
    barrier()
    dev_ms = e0.elapsed_time(e1) if mine else 0.0
    launches = ctx.launches() - l0
    # e2e: upload + K1 + D2H float frame + K2 + D2H image per job, host buffers sized for the largest job
    px = max((sc.frame.width * sc.frame.height for sc in mine), default=1)
    nctx = max(1, min(e2e_contexts, len(mine)))
    ctxs = [ctx] + [Context([torch.cuda.current_device()]) for _ in range(nctx - 1)]
    bufs = [(torch.empty(px * 3, dtype=torch.float32).pin_memory(), torch.empty(px, dtype=torch.int32).pin_memory()) for _ in ctxs]

    def worker(w):
        c, (hr, hi) = ctxs[w], bufs[w]
        for sc in mine[w::nctx]:
            c.upload(sc)
            c.render_ptr(hr.data_ptr())
            c.tonemap_ptr(1.0, 1.0, 1.0, None, hi.data_ptr())
    barrier()
    t0 = time.perf_counter()
    th = [threading.Thread(target=worker, args=(w,)) for w in range(nctx)]
    for t_ in th:
        t_.start()
    for t_ in th:
        t_.join()
    e2e_mine = time.perf_counter() - t0
    barrier()
    for c in ctxs[1:]:
        c.close()
    del bufs
    dmax, emax = reduce_max([dev_ms, e2e_mine * 1e3])
    dsum, nsum, ssum, lsum, bsum = reduce_sum([dev_ms, float(len(mine)), float(counts["samples"]), float(launches),
                                               float(sum(sc.frame.width * sc.frame.height * 16 for sc in mine))])
    return {"unit": unit_name + "/s", "value": nsum / (dmax * 1e-3) if dmax else None, "n_units": int(nsum), "ms_total": dmax,
            "rank_ms_max": dmax, "rank_ms_mean": dsum / world, "ray_samples_per_s": ssum / (dmax * 1e-3) if dmax else None,
            "e2e": {"value": nsum / (emax * 1e-3) if emax else None, "unit": unit_name + "/s", "ms_total": emax, "contexts": nctx,
                    "d2h_bytes_total": int(bsum)},
            "gpu_launches": int(lsum)}


# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the dominant kernel, from an `ncu --set full` capture;
# it belongs to the library build the capture was taken with and says so (traffic_source). The 50 MB framebuffer is
# still in the 126 MB L2 when K1 ends, so DRAM sees ~8 MB; the algorithmic bytes are 12 B/pixel = 50 MB.
K1_DRAM_TRAFFIC = {"spiral_2048_oblique_orbit": (425728 + 3480576, "profiles/r02_k1_ncu_summary.csv"),
                   "spiral_2048_cli_view": (116992 + 4035840, "profiles/r02_k1_cli_ncu_summary.csv")}
K2_DRAM_TRAFFIC = {2048 * 2048: (50342400 + 2795776, "profiles/r02b_k2_ncu_summary.csv")}


def time_frames(ctx, scenes_, Wm, K, stream, flush, barrier, after_kernels=None):
    """W untimed + K timed frames on the device: per frame upload (host) -> [flush L2] -> event -> K1 -> K2 -> [after_kernels]
    -> event. Returns (sum of the K event intervals in ms, per-frame K1 ms list, per-frame K2 ms list, launches)."""
    import torch
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]

    def step(i, k=None):
        if scenes_[i] is None:                      # a rank without rows in tiles mode
            if after_kernels:
                after_kernels(None)
            return
        ctx.upload(scenes_[i])
        with torch.cuda.stream(stream):
            flush.zero_()
            if k is not None:
                ev[k][0].record(stream)
            ctx.render_device()
            ctx.tonemap_device(1.0, 1.0, 1.0)
            if after_kernels:
                after_kernels(ctx)
            if k is not None:
                ev[k][1].record(stream)
    for i in range(Wm):
        step(i)
    barrier()
    l0 = ctx.launches()
    k1, k2 = [], []
    for k in range(K):
        step(Wm + k, k)
        if scenes_[Wm + k] is not None:
            a, b = ctx.kernel_ms()
            k1.append(a); k2.append(b)
    barrier()
    dev_ms = sum(e0.elapsed_time(e1) for (e0, e1), sc in zip(ev, scenes_[Wm:]) if sc is not None)
    return dev_ms, k1, k2, ctx.launches() - l0


def time_e2e(ctx, scenes_, first, count, nctx, size, barrier):
    """count frames end to end through host buffers (upload, K1, D2H 12 B/px, K2, D2H 4 B/px), alternating over nctx
    independent contexts with one host thread each (nctx = 1: strictly serial). Returns (seconds, checksum)."""
    import torch
    from gamer_b200.rasterizer import Context
    dev = torch.cuda.current_device()
    ctxs = [ctx] + [Context([dev]) for _ in range(nctx - 1)]
    bufs = [(torch.empty((size, size, 3), dtype=torch.float32).pin_memory(), torch.empty((size, size), dtype=torch.int32).pin_memory())
            for _ in ctxs]

    def worker(w, lo, n):
        c, (hr, hi) = ctxs[w], bufs[w]
        for i in range(lo + w, lo + n, nctx):
            c.upload(scenes_[i])
            c.render_ptr(hr.data_ptr())
            c.tonemap_ptr(1.0, 1.0, 1.0, None, hi.data_ptr())

    def run(lo, n):
        th = [threading.Thread(target=worker, args=(w, lo, n)) for w in range(nctx)]
        for t_ in th:
            t_.start()
        for t_ in th:
            t_.join()
    run(0, min(first, 3))
    barrier()
    t0 = time.perf_counter()
    run(first, count)
    barrier()
    secs = time.perf_counter() - t0
    checksum = float(bufs[0][0][::64, ::64].double().sum())
    for c in ctxs[1:]:
        c.close()
    return secs, checksum


def run_ours(args):
    import torch
    import torch.distributed as dist
    from gamer_b200 import parallel
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

    def reduce_(vals, op):
        t = torch.tensor(vals, dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=op)
        return t.tolist()
    reduce_max = lambda v: reduce_(v, dist.ReduceOp.MAX)
    reduce_sum = lambda v: reduce_(v, dist.ReduceOp.SUM)

    ctx = Context([local])
    stream = torch.cuda.ExternalStream(ctx.stream(0), device=torch.device("cuda", local))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")       # > 126 MB L2
    flush_ro = torch.zeros(64 << 20, dtype=torch.int32, device="cuda")     # 256 MB that are only ever READ (K2's cold-L2 leg)
    peaks, peak_src = measured_peaks()
    sm_count = torch.cuda.get_device_properties(local).multi_processor_count
    sm_max = float(peaks.get("sm_max_mhz", 1965.0))
    fp32_peak = sm_count * FP32_LANES_PER_SM * 2 * sm_max * 1e6 / 1e12
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))

    def leg(wl, mode, K, Wm, want_clocks=False, e2e_contexts=(1,)):
        """One workload in one partition mode -> the numbers of a bench line (rank 0 gets the dict, others None)."""
        size = WORKLOADS[wl][1]
        tiles = mode == "tiles"
        if tiles:       # strong scaling: every rank renders its interleaved bands of the SAME frame, then one gather
            scenes_ = [parallel.band_tile(make_scene(wl, s), rank, world) for s in range(Wm + K)]
        else:           # weak scaling: step s renders frames s*world .. s*world+world-1, one per rank
            scenes_ = [make_scene(wl, s * world + rank) for s in range(Wm + K)]
        counts = []
        for sc in scenes_[Wm:]:     # exact work of every timed frame (instrumented kernel, outside the timed region)
            if sc is not None:
                ctx.upload(sc)
                counts.append(ctx.render_counted())
        my_samples = sum(c["samples"] for c in counts)
        my_flop = sum(algorithmic_flop(c) for c in counts)
        my_px = scenes_[Wm].frame.width * scenes_[Wm].frame.height if scenes_[Wm] is not None else 0

        gathered = [None, None]

        def gather(c):              # K3: the one exchange step — the disjoint bands to rank 0 over NVLink
            if c is None:
                rgb_t = torch.empty((0, size, 3), dtype=torch.float32, device="cuda")
                img_t = torch.empty((0, size), dtype=torch.int32, device="cuda")
            else:
                rgb_t, _, img_t = c.device_tensors()
            gathered[0] = parallel.gather_bands(rgb_t, size, rank, world)
            gathered[1] = parallel.gather_bands(img_t, size, rank, world)
        clocks = ClockSampler(local)
        if want_clocks and rank == 0:
            clocks.start()
        dev_ms, k1_ms, k2_ms, launches = time_frames(ctx, scenes_, Wm, K, stream, flush, barrier,
                                                     gather if tiles and world > 1 else None)
        clk = clocks.stop() if want_clocks and rank == 0 else None
        gathered[:] = [None, None]

        # ---- e2e: host buffers through the C ABI ----
        e2e = {}
        if tiles and world > 1:
            # every rank copies its own bands straight into ONE page-locked host frame all ranks map
            # (gamer_b200_copy_bands): no device collective and no serial copy on the host path
            shm_rgb = SharedHostFrame("rgb", size * size * 12, rank, barrier)
            shm_img = SharedHostFrame("img", size * size * 4, rank, barrier)

            def e2e_step(i):
                if scenes_[i] is not None:
                    ctx.upload(scenes_[i])
                    ctx.render_device()
                    ctx.tonemap_device(1.0, 1.0, 1.0)
                    ctx.copy_bands(0, shm_rgb.ptr)
                    ctx.copy_bands(2, shm_img.ptr)
            for i in range(min(Wm, 3)):
                e2e_step(i)
            barrier()
            t0 = time.perf_counter()
            for i in range(Wm, Wm + K):
                e2e_step(i)
                barrier()                           # the frame is complete on the host when every rank has copied
            secs = time.perf_counter() - t0
            chk = float(np.frombuffer(shm_rgb.mm, dtype=np.float32).reshape(size, size, 3)[::64, ::64].astype(np.float64).sum()) if rank == 0 else 0.0
            e2e[1] = (secs, chk, "each rank copies its bands into one shared page-locked host frame (gamer_b200_copy_bands)",
                      shm_rgb.registered and shm_img.registered)
            shm_rgb.close(); shm_img.close()
        else:
            for nctx in e2e_contexts:
                secs, chk = time_e2e(ctx, scenes_, Wm, K, nctx, size, barrier)
                e2e[nctx] = (secs, chk, "serial: one context, one host thread" if nctx == 1 else
                             f"{nctx} contexts, one host thread each: one frame's D2H overlaps the next frame's kernels", True)

        # ---- K2 alone, cold L2, for its HBM roofline (this rank's pixels) ----
        k2_cold = []
        if scenes_[Wm] is not None:
            ctx.upload(scenes_[Wm]); ctx.render_device()
            for _ in range(5):
                # cold AND clean: after the write sweep the L2 holds 126 MB of DIRTY lines of the flush buffer, and a kernel
                # as short as K2 (15-20 us) would be timed writing those back as it allocates (it measured 20.4 us that way,
                # 15.5 us under ncu, whose flush also writes back). The read sweep of a second buffer evicts them first.
                with torch.cuda.stream(stream):
                    flush.zero_()
                    flush_ro.sum()
                ctx.tonemap_device(1.0, 1.0, 1.0)
                k2_cold.append(ctx.kernel_ms()[1])
        e2e_keys = sorted(e2e)
        dev_ms_max, *e2e_ms_max = reduce_max([dev_ms] + [e2e[k][0] * 1e3 for k in e2e_keys])
        tot_samples, tot_flop, tot_launches = reduce_sum([float(my_samples), my_flop, float(launches)])
        if rank != 0:
            return None
        k1_avg_ms = sum(k1_ms) / max(len(k1_ms), 1)
        k1_tflops = (my_flop / max(len(counts), 1)) / (k1_avg_ms * 1e-3) / 1e12 if k1_ms else 0.0
        k2_best = min(k2_cold) if k2_cold else None
        k2_bytes = my_px * (12 + 4 + 4)
        best = min(range(len(e2e_keys)), key=lambda j: e2e_ms_max[j])
        e2e_objs = {}
        for j, kx in enumerate(e2e_keys):
            e2e_objs[kx] = {"value": tot_samples / (e2e_ms_max[j] * 1e-3), "unit": UNIT,
                            "h2d_bytes_per_step": (1 if tiles else world) * ctx_scene_bytes(scenes_[Wm]) if scenes_[Wm] is not None else 0,
                            "d2h_bytes_per_step": (1 if tiles else world) * size * size * 16,
                            "ms_per_step": e2e_ms_max[j] / K, "checksum": e2e[kx][1], "contexts": kx, "how": e2e[kx][2],
                            "host_buffers_page_locked": bool(e2e[kx][3])}
        traffic = K1_DRAM_TRAFFIC.get(wl) if mode == "frames" else None
        k2_traffic = K2_DRAM_TRAFFIC.get(my_px)
        out = {
            "metric": METRIC, "value": tot_samples / (dev_ms_max * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": K, "warmup": Wm, "ms_per_step": dev_ms_max / K, "higher_is_better": True, "scaling": "strong" if tiles else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(wl, mode, world),
            "work": {"ray_samples_per_frame": tot_samples / K / (1 if tiles else world),
                     "octaves_per_frame_rank0": sum(c["octaves"] for c in counts) / max(len(counts), 1)},
            "gpu_launches": int(tot_launches),
            "e2e": e2e_objs[e2e_keys[best]],
            "roofline": {"kernel": "k1_march", "bound": "fp32", "achieved": k1_tflops, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": k1_tflops / fp32_peak, "traffic": traffic[0] if traffic else None,
                         "traffic_source": (traffic[1] + " (one ncu --set full capture; not re-measured by this run)") if traffic else None,
                         "peak_source": f"{sm_count} SMs x 128 FP32 lanes x 2 flop x {sm_max:.0f} MHz (sm_max_mhz of "
                                        f"MEASURED_PEAKS.json, {peak_src}); the path is FP32/SFU-pipe bound, not HBM or tensor",
                         "algorithmic_flop_per_launch": my_flop / max(len(counts), 1), "launch_ms": k1_avg_ms},
            "kernel_ms": {"k1_avg": k1_avg_ms, "k2_avg_in_step": sum(k2_ms) / max(len(k2_ms), 1), "k2_cold_l2": k2_best},
        }
        if len(e2e_keys) > 1:
            out["e2e_serial"] = e2e_objs[1]
        if k2_best:
            out["roofline_k2"] = {"kernel": "k2_tonemap_v4", "bound": "hbm", "achieved": k2_bytes / (k2_best * 1e-3) / 1e9,
                                  "peak": hbm_peak, "unit": "GB/s", "frac": k2_bytes / (k2_best * 1e-3) / 1e9 / hbm_peak,
                                  "traffic": k2_traffic[0] if k2_traffic else None, "traffic_source": k2_traffic[1] if k2_traffic else None,
                                  "algorithmic_bytes_per_launch": k2_bytes, "pixels_this_rank": my_px, "launch_ms": k2_best,
                                  "l2": "cold and clean: a 256 MB write sweep, then a 256 MB read sweep of another buffer, before each timed launch",
                                  "peak_source": f"hbm_gbs of MEASURED_PEAKS.json ({peak_src})"}
        if clk:
            out["clocks"] = clk
        return out

    K, Wm, wl = args.steps, args.warmup, args.workload
    out = leg(wl, args.mode, K, Wm, want_clocks=True, e2e_contexts=(1, args.e2e_contexts) if args.e2e_contexts > 1 else (1,))

    # ---- the other BASELINE configs, each a short measured workload of its own ----
    also = {}
    if not args.no_extras:
        import scenes
        t_ex = leg("sombrero_4096_edge", "tiles", 3, 3)
        if rank == 0:
            also["tiles_sombrero_4096"] = {k: t_ex[k] for k in ("value", "unit", "ms_per_step", "scaling", "config", "e2e", "gpu_launches", "kernel_ms")}
        n_seq = args.sequence_frames
        seq = [scenes.hd_frame_scene(k, n_frames=360) for k in range(n_seq)]
        r = run_jobs(ctx, seq, parallel.units_of(rank, world, n_seq), rank, world, barrier, reduce_max, reduce_sum, "frames", args.e2e_contexts)
        if rank == 0:
            r["config"] = {"workload": "sequence_1920x1080", "frames": n_seq, "of": 360, "galaxy": "Spiral.gax", "size": "1920x1080 (rows 420..1499 of the 1920^2 image)",
                           "parallelism": f"frame k -> rank k mod {world}", "l2": "frames differ (camera moves 1 deg per frame); no flush between them"}
            also["sequence_1920x1080"] = r
        n_gal = args.batch_galaxies
        batch = [scenes.random_galaxy_scene(i, 1024) for i in range(n_gal)]
        # cost of a job without an oracle: its exact algorithmic work at 1/256 of the pixels (64^2, counting kernel,
        # ~0.1 ms each). Every rank measures all of them itself, so the table needs no exchange. The static proxy
        # (pixels x octaves per in-slab sample, parallel.job_cost) ignores the view and left 9 % imbalance at N = 2.
        costs = []
        for i in range(n_gal):
            ctx.upload(scenes.random_galaxy_scene(i, 64))
            costs.append(algorithmic_flop(ctx.render_counted()) + 1.0)
        # one queue for all ranks in longest-job-first order (parallel.JobQueue): the static table of
        # parallel.longest_job_first balances the estimated costs and left 6.7 % between the slowest rank and the mean
        # at N = 8 (232 ms against 217.5); a rank that finishes early now takes the next job instead
        order = parallel.ljf_order(costs)
        r = run_jobs(ctx, batch, parallel.units_of(rank, world, n_gal), rank, world, barrier, reduce_max, reduce_sum, "galaxies",
                     args.e2e_contexts, queue=parallel.JobQueue(order, "batch"))
        if rank == 0:
            r["config"] = {"workload": "batch_256_1024", "galaxies": n_gal, "of": 256, "size": 1024, "seed": 20261104,
                           "parallelism": f"one job queue over {world} ranks, longest job first by each galaxy's exact work at 64^2 "
                                          "(parallel.JobQueue: a shared counter in the c10d store, control plane only)",
                           "l2": "every job is a different galaxy; no flush between them"}
            also["batch_256_1024"] = r
        if world == 1:
            c = leg("spiral_2048_cli_view", "frames", 3, 3)
            also["cli_view"] = {k: c[k] for k in ("value", "unit", "ms_per_step", "config", "roofline", "gpu_launches")}

    if rank == 0:
        if world == 1 and not args.no_cpu:
            arm = CpuArm(wl, args.cpu_row_stride)
            best = min(arm.step() for _ in range(1))
            out["cpu_baseline"] = {"value": arm.samples / best, "unit": UNIT, "cores": arm.cores, "kind": arm.kind,
                                   "sample": arm.describe(), "seconds": best}
            out["parity"] = parity_against_cpu_rows(ctx, arm)
            port = port_value(arm)
            if port:
                out["cpu_baseline"]["oracle_port"] = port
            arm.close()
        if also:
            out["also"] = also
        print(json.dumps(out), flush=True)
    del flush, flush_ro
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
    reference's own sources compiled in place when oracle/_ref travelled here, else the oracle port), ALL host
    threads (asked for explicitly), each step a bounded sample of the same workload. Runs on rank 0 only; loads neither
    torch nor libgamer_b200.so."""
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
    frac = sc.frame.height / sc.frame.full_size
    desc = (f"each step = {arm.describe()}; a whole frame is ~{1 / frac:.0f}x that "
            f"(~{1e3 * total / len(times) / frac:.0f} ms per full frame at this rate)")
    loaded = [l.split()[-1] for l in open("/proc/self/maps") if "libgamer_b200" in l]
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True,
           "scaling": "strong" if args.mode == "tiles" else "weak",
           "vs_baseline": None, "dtype": "f64/f32 (reference mix)", "data": "synthetic",
           "config": workload_config(wl, args.mode, world),
           "cpu_baseline": {"value": value, "unit": UNIT, "cores": arm.cores, "kind": arm.kind, "sample": desc,
                            **({"oracle_port": port_value(arm)} if arm.kind == "reference" else {})},
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0, "product_library_loaded": bool(loaded)}
    arm.close()
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
                         "into interleaved row bands + the final gather (strong)")
    ap.add_argument("--e2e-contexts", type=int, default=3,
                    help="independent gamer_ctx (one host thread each) the pipelined e2e leg alternates frames over; the "
                         "strictly serial figure (1 context) is always measured as well")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline and parity legs")
    ap.add_argument("--no-extras", action="store_true", help="skip the C3 / C4 / C5 / C2a workloads carried under 'also'")
    ap.add_argument("--sequence-frames", type=int, default=360, help="frames of the C4 sequence to render (of 360)")
    ap.add_argument("--batch-galaxies", type=int, default=256, help="galaxies of the C5 batch to render (of 256)")
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
