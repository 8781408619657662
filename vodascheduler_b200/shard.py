"""Multi-GPU plumbing: rounds are independent units, so they shard across ranks with NO
data-path collective.  One process per GPU (torchrun); each rank takes a
contiguous range of rounds, computes it on its own device through the C ABI, and the only
exchange is an optional gather of the outputs plus a 4-number summary per rank
(rounds, GPUs handed out, rounds that panicked, a digest of all the outputs) for a cross-rank check.

`compute` is whatever maps a host AllocBatch to (gpus, status): Context.allocate on a GPU
rank.  The host logic here is device-agnostic so it can be tested with the gloo backend.
"""
from __future__ import annotations

import numpy as np

from .batch import AllocBatch


def shard_bounds(rounds: int, world: int, rank: int):
    """Contiguous split of [0, rounds) into `world` nearly equal ranges."""
    return rounds * rank // world, rounds * (rank + 1) // world


def digest64(arr: np.ndarray) -> int:
    """63-bit digest of the WHOLE array (BLAKE2b, 8-byte digest): equal outputs <=> equal digests
    for any practical purpose, at memory speed."""
    import hashlib
    h = hashlib.blake2b(np.ascontiguousarray(arr).view(np.uint8).data, digest_size=8).digest()
    return int.from_bytes(h, "little") & 0x7FFFFFFFFFFFFFFF


def summary(gpus: np.ndarray, status: np.ndarray) -> np.ndarray:
    """Per-rank summary: rounds, GPUs handed out, rounds with a non-zero status, digest of all the
    per-job outputs."""
    ok = status == 0
    return np.array([status.size, int(gpus.astype(np.int64).sum()), int((~ok).sum()), digest64(gpus)], np.int64)


def run_sharded(batch: AllocBatch, compute, rank: int, world: int, dist=None, gather: bool = True):
    """Compute this rank's rounds; optionally all-gather every rank's outputs.

    Returns (gpus, status, summaries): full-batch arrays when gather=True (identical on all
    ranks), else this rank's slice; summaries is a [world, 4] int64 array (rank-local row only
    when dist is None)."""
    r0, r1 = shard_bounds(batch.rounds, world, rank)
    sub = batch.slice_rounds(r0, r1)
    g, s = compute(sub)
    g = np.asarray(g, np.int32)
    s = np.asarray(s, np.int32)
    mine = summary(g, s)
    if dist is None or world == 1:
        return g, s, mine[None, :]
    import torch
    sums = [torch.zeros(4, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(sums, torch.from_numpy(mine))
    sums = np.stack([t.numpy() for t in sums])
    if not gather:
        return g, s, sums
    # outputs: pad to the largest shard so all_gather sees equal shapes
    jobs_per = [int(batch.job_off[shard_bounds(batch.rounds, world, k)[1]]) -
                int(batch.job_off[shard_bounds(batch.rounds, world, k)[0]]) for k in range(world)]
    rounds_per = [shard_bounds(batch.rounds, world, k)[1] - shard_bounds(batch.rounds, world, k)[0]
                  for k in range(world)]
    gp = torch.zeros(max(jobs_per + [1]), dtype=torch.int32)
    sp = torch.zeros(max(rounds_per + [1]), dtype=torch.int32)
    gp[:g.size] = torch.from_numpy(g)
    sp[:s.size] = torch.from_numpy(s)
    gl = [torch.zeros_like(gp) for _ in range(world)]
    sl = [torch.zeros_like(sp) for _ in range(world)]
    dist.all_gather(gl, gp)
    dist.all_gather(sl, sp)
    full_g = np.concatenate([gl[k].numpy()[:jobs_per[k]] for k in range(world)])
    full_s = np.concatenate([sl[k].numpy()[:rounds_per[k]] for k in range(world)])
    return full_g, full_s, sums
