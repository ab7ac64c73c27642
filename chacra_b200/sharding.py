"""Multi-GPU work split for the contact path.

Frames are independent and so are states; the only coupling is that the hydrophobic-carbon set is defined per
100-frame fragment from the fragment's first frame (``hydrophobics.py:73-77`` through the per-fragment cutoff dict,
``compute_contacts.py:260-261``).  Work items are therefore (state, fragment); every rank takes a contiguous,
fragment-aligned slice of every state, accumulates its own ``int32 [states][pairs]`` count table and the tables are
merged with ONE all-reduce (integer sum: order-free, bit-exact for any rank count).  The reference's analogue is the
fragment queue feeding ``cores`` worker processes (``compute_contacts.py:250-281``).
"""
from __future__ import annotations

from typing import List, Tuple

TRAJ_FRAG_SIZE = 100


def fragment_slice(n_frames: int, world: int, rank: int, fragment: int = TRAJ_FRAG_SIZE) -> Tuple[int, int]:
    """(first_index, count) of the frames rank ``rank`` processes out of ``n_frames`` selected frames.
    Slices are contiguous, start on a fragment boundary, are disjoint and cover everything; fragment counts per
    rank differ by at most one."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world %d/%d" % (rank, world))
    n_frag = (n_frames + fragment - 1) // fragment
    base, extra = divmod(n_frag, world)
    f0 = rank * base + min(rank, extra)
    f1 = f0 + base + (1 if rank < extra else 0)
    first = min(f0 * fragment, n_frames)
    last = min(f1 * fragment, n_frames)
    return first, last - first


def state_slices(frames_per_state, world: int, rank: int, fragment: int = TRAJ_FRAG_SIZE) -> List[Tuple[int, int, int]]:
    """[(state, first_index, count)] for this rank: a slice of EVERY state (balanced even when the number of states
    does not divide by the number of GPUs, e.g. 20 states on 8 GPUs)."""
    out = []
    for s, n in enumerate(frames_per_state):
        first, cnt = fragment_slice(int(n), world, rank, fragment)
        if cnt:
            out.append((s, first, cnt))
    return out


def merge_count_tables(counts, group=None):
    """Sum the per-rank count tables in place (``torch.distributed.all_reduce``; NCCL over NVLink for CUDA tensors,
    gloo for the CPU tests).  A no-op when torch.distributed is not initialised (single GPU)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
    return counts
