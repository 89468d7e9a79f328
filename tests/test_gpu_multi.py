"""Multi-GPU parity (-m gpu, needs >= 2 devices; skipped on a 1-GPU box): sharding the frame over ranks
(interleaved bands + NCCL gather) and over the devices of one multi-device context must give the
single-GPU frame bit for bit — every pixel is computed by the same kernel with the same arithmetic."""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest

import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch
    return torch.cuda.device_count()


WORKER = textwrap.dedent("""
    import os, sys
    sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "tests"))
    import numpy as np, torch, torch.distributed as dist
    import scenes
    from gamer_b200 import parallel
    from gamer_b200.rasterizer import Context
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    ctx = Context([local])
    stream = torch.cuda.ExternalStream(ctx.stream(0), device=torch.device("cuda", local))
    for name, size, view in (("Spiral", 200, "oblique"), ("Sombrero", 136, "edge")):
        sc = scenes.single_galaxy_scene(name, size, view)
        ctx.upload(parallel.band_tile(sc, rank, world))
        with torch.cuda.stream(stream):
            ctx.render_device()
            ctx.tonemap_device(1.0, 1.0, 1.0)
            rgb, _, img = ctx.device_tensors()
            frame = parallel.gather_bands(rgb, size, rank, world)
            image = parallel.gather_bands(img, size, rank, world)
        stream.synchronize()
        # the host-side gather: every rank copies its bands into ONE frame in shared memory (gamer_b200_copy_bands)
        path = "/dev/shm/gamer_b200_test_%s_%s" % (os.environ["MASTER_PORT"], name)
        if rank == 0:
            with open(path, "wb") as fh:
                fh.truncate(size * size * 16)
        dist.barrier()
        mm = np.memmap(path, dtype=np.uint8, mode="r+", shape=(size * size * 16,))
        ctx.copy_bands(0, mm[: size * size * 12].ctypes.data)
        ctx.copy_bands(2, mm[size * size * 12:].ctypes.data)
        dist.barrier()
        if rank == 0:
            ctx.upload(sc)
            whole = ctx.render()
            argb, shadow = ctx.tonemap(1.0, 1.0, 1.0)
            assert np.array_equal(frame.cpu().numpy(), whole), name + ": sharded frame differs from single-GPU frame"
            assert np.array_equal(image.cpu().numpy().view(np.uint32), shadow), name + ": sharded image differs"
            assert np.array_equal(np.frombuffer(mm[: size * size * 12], dtype=np.float32).reshape(size, size, 3), whole), name + ": copy_bands frame differs"
            assert np.array_equal(np.frombuffer(mm[size * size * 12:], dtype=np.uint32).reshape(size, size), shadow), name + ": copy_bands image differs"
        del mm
        dist.barrier()
        if rank == 0:
            os.unlink(path)
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()
    print("rank %d ok" % rank, flush=True)   # one write for the text: the ranks share the pipe
""")


def test_nccl_band_gather_equals_single_gpu(tmp_path):
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={min(n, 4)}",
           "--master-addr", "127.0.0.1", "--master-port", "29541", str(script)]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "rank 0 ok" in p.stdout


def test_multi_device_context_equals_single_gpu():
    """One process driving several GPUs through one gamer_ctx (the C ABI's own band split)."""
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    from gamer_b200.rasterizer import Context
    sc = scenes.single_galaxy_scene("Spiral", 200, "oblique")
    one = Context([0])
    one.upload(sc)
    ref = one.render()
    ref_img = one.tonemap(1.0, 1.0, 1.0)
    one.close()
    many = Context(list(range(min(n, 4))))
    many.upload(sc)
    got = many.render()
    got_img = many.tonemap(1.0, 1.0, 1.0)
    many.close()
    assert np.array_equal(got, ref)
    assert np.array_equal(got_img[0], ref_img[0]) and np.array_equal(got_img[1], ref_img[1])
