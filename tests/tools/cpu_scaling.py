"""Thread scaling of the two CPU arms of bench.py on this host: the reference's own code
(oracle/_ref/libgamer_ref_path_fast.so) and the oracle port. Prints one JSON line per (arm, threads)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402

if __name__ == "__main__":
    wl = sys.argv[1] if len(sys.argv) > 1 else "spiral_2048_oblique_orbit"
    stride = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    ncpu = os.cpu_count() or 1
    for force_port in (False, True):
        for t in sorted({1, 2, 4, 8, 16, 32, ncpu}):
            if t > ncpu:
                continue
            arm = bench.CpuArm(wl, stride, threads=t, force_port=force_port)
            dt = min(arm.step() for _ in range(2))
            print(json.dumps({"arm": arm.kind, "threads": t, "samples_per_s": arm.samples / dt, "seconds": dt,
                              "host_cpus": ncpu}), flush=True)
            arm.close()
