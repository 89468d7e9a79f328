"""N>1 host logic on CPU: world_size-2 gloo processes exercise the band partition + gather and the
frame sharding used by bench.py / the multi-GPU path (no CUDA involved)."""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest

import scenes
from gamer_b200 import parallel

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("height", [1, 31, 32, 100, 2048])
def test_band_rows_partition(world, height):
    rows = [parallel.band_rows(r, world, height) for r in range(world)]
    flat = sorted(y for r in rows for y in r)
    assert flat == list(range(height))                       # disjoint cover, bit-exact tile indexing
    sizes = [len(r) for r in rows]
    assert max(sizes) - min(sizes) <= parallel.BAND


def test_units_of():
    assert parallel.units_of(1, 4, 10) == [1, 5, 9]
    assert sorted(sum((parallel.units_of(r, 8, 360) for r in range(8)), [])) == list(range(360))


def test_band_tile_matches_oracle_rows(oracle):
    """A rank's band tile renders exactly the rows band_rows() says (checked with the CPU oracle)."""
    sc = scenes.single_galaxy_scene("Spiral", 80, "oblique")
    whole, _ = oracle.render(sc)
    for world in (2, 3):
        for r in range(world):
            t = parallel.band_tile(sc, r, world)
            part, _ = oracle.render(t)
            assert np.array_equal(part, whole[parallel.band_rows(r, world, 80)])


WORKER = textwrap.dedent("""
    import os, sys
    sys.path.insert(0, {root!r})
    import torch, torch.distributed as dist
    from gamer_b200 import parallel
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    H, W = 100, 7
    rows = parallel.band_rows(rank, world, H)
    local = torch.tensor(rows, dtype=torch.float32)[:, None, None].expand(len(rows), W, 3).contiguous() * 10 + rank
    out = parallel.gather_bands(local, H, rank, world)
    if rank == 0:
        owner = torch.tensor([(y // parallel.BAND) % world for y in range(H)], dtype=torch.float32)
        expect = (torch.arange(H, dtype=torch.float32) * 10 + owner)[:, None, None].expand(H, W, 3)
        assert torch.equal(out, expect), "gathered frame differs"
    else:
        assert out is None
    mx = parallel.max_over_ranks([float(rank)], "cpu")[0]
    sm = parallel.sum_over_ranks([1.0], "cpu")[0]
    assert mx == world - 1 and sm == world
    assert parallel.units_of(rank, world, 5) == list(range(rank, 5, world))
    dist.destroy_process_group()
    print("rank %d ok" % rank, flush=True)   # one write for the text: the ranks share the pipe
""")


def test_gloo_world2_gather(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", str(script)]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "rank 0 ok" in p.stdout and "rank 1 ok" in p.stdout
