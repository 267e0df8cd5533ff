"""CPU suite: the multi-GPU host logic with world_size 2 over gloo (no GPU needed).

Each rank builds the same host state, jumps its per-pixel generators to its own stretch of
the streams, and the per-rank accumulation images meet in ONE reduce — the same calls
bench.py makes over NCCL.
"""
import os
import socket
import sys

import numpy as np
import pytest

import util


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world))
    sys.path.insert(0, util.ROOT)
    import torch
    import torch.distributed as dist

    import __graft_entry__ as entry

    vpt = entry.load_package()
    dist.init_process_group("gloo", rank=rank, world_size=world)
    b, sh = vpt.binding, vpt.sharding
    sc = b.Scene(util.scene_path("mixed"))
    r = b.Render(sc, b.make_params(shader="pathtrace", samples=8, resolution=32))
    base = r.rngs().copy()
    r.reset(sh.rank_rng_offset(rank))
    mine = r.rngs()
    # the jump is PCG's advance by rank << 40 draws, same increment
    for px in (0, 1, base.shape[0] - 1):
        want = sh.pcg_advance(int(base[px, 0]), int(base[px, 1]), sh.rank_rng_offset(rank))
        assert int(mine[px, 0]) == want and mine[px, 1] == base[px, 1]
    # stand-in for the rank's accumulation image: depends on rank and pixel, 2 samples each
    h, w = r.height, r.width
    local = torch.full((h * w * 4,), float(rank + 1)) * torch.arange(h * w * 4, dtype=torch.float32)
    total = sh.reduce_images(local.clone(), dst=0)
    gathered = [torch.zeros(2, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(gathered, torch.tensor([int(mine[0, 0] >> 1), rank], dtype=torch.int64))
    if rank == 0:
        np.save(os.path.join(out_dir, "sum.npy"), total.numpy())
        np.save(os.path.join(out_dir, "states.npy"), torch.stack(gathered).numpy())
    dist.barrier()
    dist.destroy_process_group()
    r.close()
    sc.close()


def test_two_rank_sample_sharding_over_gloo(vpt, tmp_path):
    import torch.multiprocessing as mp

    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    total = np.load(tmp_path / "sum.npy")
    n = total.size
    assert np.array_equal(total, (1.0 + 2.0) * np.arange(n, dtype=np.float32))
    states = np.load(tmp_path / "states.npy")
    assert states[0, 1] == 0 and states[1, 1] == 1
    assert states[0, 0] != states[1, 0], "ranks must not share a stream position"


def test_sample_ranges_are_disjoint_and_cover(vpt):
    sh = vpt.sharding
    for total, world in ((4096, 8), (1024, 3), (5, 8), (0, 2), (7, 1)):
        ranges = sh.sample_ranges(total, world)
        assert len(ranges) == world
        assert sum(n for _, n in ranges) == total
        pos = 0
        for start, n in ranges:
            assert start == pos and n >= 0
            pos += n
        assert max(n for _, n in ranges) - min(n for _, n in ranges) <= 1
    with pytest.raises(ValueError):
        sh.sample_ranges(8, 0)


def test_rank_offsets_do_not_overlap(vpt):
    sh = vpt.sharding
    offs = [sh.rank_rng_offset(r) for r in range(8)]
    assert offs[0] == 0, "a single rank replays the reference's streams"
    # one pixel draws well under 2^40 numbers even at 4096 spp x 64 bounces x ~10 draws
    assert all(b - a == 1 << 40 for a, b in zip(offs, offs[1:]))
    assert 4096 * 64 * 16 < 1 << 40
    # advance composes: jumping a then b equals jumping a + b
    s, inc = 10473825460370060855, 1450249601
    assert sh.pcg_advance(sh.pcg_advance(s, inc, 12345), inc, 1 << 40) == sh.pcg_advance(s, inc, 12345 + (1 << 40))
    # and matches stepping the generator
    step = s
    for _ in range(5):
        step = (step * sh.PCG_MULT + inc) & sh.MASK64
    assert sh.pcg_advance(s, inc, 5) == step
