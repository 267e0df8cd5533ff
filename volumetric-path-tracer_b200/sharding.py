"""Sample-range sharding of one render over the GPUs of a node.

Every rank renders the whole image but owns a disjoint stretch of each pixel's
PCG32 stream: rank r starts `r << RNG_STRIDE_LOG2` draws into the stream the
reference seeds for that pixel (yocto_pathtrace.cpp:1116-1121 seeds, the jump is
PCG's O(log n) advance). The per-rank accumulation images are summed with ONE
reduce (NCCL on the GPUs, gloo in the CPU tests); there is no other data-path
collective. Rank 0 of a 1-rank job reproduces the reference's streams exactly.
"""
from __future__ import annotations

RNG_STRIDE_LOG2 = 40  # 2^40 draws per rank: far beyond what one rank consumes per pixel
PCG_MULT = 6364136223846793005
MASK64 = (1 << 64) - 1


def rank_rng_offset(rank: int) -> int:
    return rank << RNG_STRIDE_LOG2


def sample_ranges(total_samples: int, world: int):
    """[(first_sample, count)] per rank: contiguous, disjoint, covering [0, total_samples)."""
    if world <= 0 or total_samples < 0:
        raise ValueError("world must be positive and total_samples non-negative")
    base, extra = divmod(total_samples, world)
    out, start = [], 0
    for r in range(world):
        n = base + (1 if r < extra else 0)
        out.append((start, n))
        start += n
    return out


def pcg_advance(state: int, inc: int, delta: int):
    """PCG32 jump-ahead (the host / device k_advance_rngs do the same): state after `delta` draws."""
    acc_mult, acc_plus, cur_mult, cur_plus = 1, 0, PCG_MULT, inc
    while delta > 0:
        if delta & 1:
            acc_mult = (acc_mult * cur_mult) & MASK64
            acc_plus = (acc_plus * cur_mult + cur_plus) & MASK64
        cur_plus = ((cur_mult + 1) * cur_plus) & MASK64
        cur_mult = (cur_mult * cur_mult) & MASK64
        delta >>= 1
    return (acc_mult * state + acc_plus) & MASK64


def reduce_images(local_sum, dst: int = 0):
    """Sum the per-rank accumulation images on `dst` (in place there). One collective."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(local_sum, dst=dst, op=dist.ReduceOp.SUM)
    return local_sum
