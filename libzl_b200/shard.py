"""Multi-GPU partitioning of the render path: one process per GPU, independent units sharded by rank
with no data-path collective; a collective only where ONE stereo mix spans GPUs (the reference's JACK graph sums the
SamplerChannel ports of all tracks into one master, lib/JackPassthrough.cpp:94-101 + the JACK port mix).

Host logic only; the kernels never see the world size.  On the GPUs the collective is the library's own
(zl_comm_init / zl_mix_reduce over the NCCL C API, include/zl_b200.h); reduce_mix below is the torch.distributed form of
the same reduce for hosts that already run a process group (and for the gloo CPU test)."""
from __future__ import annotations


def units_for_rank(n_units: int, rank: int, world: int) -> list[int]:
    """Unit (session / track / clip) ids owned by `rank`: unit_id mod world == rank (round-robin)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return list(range(rank, n_units, world))


def owner_of(unit: int, world: int) -> tuple[int, int]:
    """(rank, local index) of a global unit id under the round-robin partition."""
    return unit % world, unit // world


def mix_groups(world: int, span: int) -> list[list[int]]:
    """Groups of `span` consecutive ranks that share one stereo mix (span must divide world)."""
    if span < 1 or world % span:
        raise ValueError("span must divide the world size")
    return [list(range(g, g + span)) for g in range(0, world, span)]


def reduce_mix(master, group=None, dst=None):
    """Sum the per-rank partial stereo buses of one mix (float32 [sessions, 2, frames]) over the ranks of `group`.
    dst=None: all-reduce (every rank ends with the mix); else reduce to global rank `dst` (ncclReduce).
    The summation order differs from the reference's serial += over tracks, so parity is 1e-5 of full scale, not bit-exact;
    peak meters of a spanned mix must be taken AFTER this reduce (max is not linear in the partial sums)."""
    import torch.distributed as dist
    if dst is None:
        dist.all_reduce(master, op=dist.ReduceOp.SUM, group=group)
    else:
        dist.reduce(master, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return master
