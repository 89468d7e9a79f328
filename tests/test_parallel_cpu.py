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
    # the shared job queue: every index of the order handed out exactly once over the two ranks
    import time, random
    order = parallel.ljf_order([float((7 * i) % 13) for i in range(40)])
    random.seed(rank)
    mine = []
    for j in parallel.JobQueue(order, "t1"):
        mine.append(j); time.sleep(random.random() * 2e-3)
    got = [None, None]
    dist.all_gather_object(got, mine)
    assert sorted(got[0] + got[1]) == list(range(40)), got
    assert all(order.index(a) < order.index(b) for m in got for a, b in zip(m, m[1:])), "a rank's jobs are not in queue order"
    assert len(got[0]) > 5 and len(got[1]) > 5, got
    assert list(parallel.JobQueue(order, "t1")) == []          # the same queue is spent; a new name is a new queue
    dist.barrier()
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


def test_band_tile_of_a_frame_smaller_than_the_band_fan():
    """height < BAND * world: some ranks own no band. band_tile returns None for them (they neither upload nor render)
    and the others still cover the frame exactly (it used to produce a height-0 tile that upload rejects)."""
    sc = scenes.single_galaxy_scene("Spiral", 64, "oblique")
    tiles = [parallel.band_tile(sc, r, 4) for r in range(4)]
    assert tiles[2] is None and tiles[3] is None
    assert [t.frame.height for t in tiles[:2]] == [32, 32]
    rows = sorted(y for r in range(4) for y in parallel.band_rows(r, 4, 64))
    assert rows == list(range(64))


def test_longest_job_first_balances_and_is_deterministic():
    """The galaxy batch / frame sequence queue: greedy longest-processing-time on a cost proxy."""
    rng = np.random.default_rng(0)
    costs = list(rng.uniform(1, 100, 256))
    for world in (1, 2, 4, 8):
        jobs = parallel.longest_job_first(costs, world)
        assert sorted(j for r in jobs for j in r) == list(range(256))                 # every job exactly once
        loads = [sum(costs[j] for j in r) for r in jobs]
        assert max(loads) - min(loads) <= max(costs)                                   # LPT bound on the spread
        assert max(loads) <= sum(costs) / world * 4 / 3 + 1e-9                         # Graham's 4/3 bound (vs the mean)
        assert all(costs[a] >= costs[b] for r in jobs for a, b in zip(r, r[1:]))       # each rank runs its longest first
        assert jobs == parallel.longest_job_first(costs, world)                        # same table on every rank
    # modulo assignment of a cost-sorted batch can be far worse: that is what the queue is for
    skew = [100.0] * 8 + [1.0] * 56
    lpt = max(sum(skew[j] for j in r) for r in parallel.longest_job_first(skew, 8))
    mod = max(sum(skew[j] for j in parallel.units_of(r, 8, 64)) for r in range(8))
    assert lpt <= mod


def test_job_queue_single_process_is_the_order_itself():
    order = parallel.ljf_order([3.0, 9.0, 1.0, 9.0])
    assert order == [1, 3, 0, 2]
    assert list(parallel.JobQueue(order, "solo")) == order


def test_job_cost_counts_octaves_of_active_components():
    sc = scenes.single_galaxy_scene("Spiral", 64, "oblique")
    # Spiral.gax: bulge, disk(10), dust2(9), [inactive bulge], stars(10), disk(10)
    assert parallel.job_cost(sc) == 64 * 64 * (1 + 10 + 9 + 10 + 10)
