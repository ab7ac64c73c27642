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


def state_slices(frames_per_state, world: int, rank: int, fragment: int = TRAJ_FRAG_SIZE, balance: str = "global") -> List[Tuple[int, int, int]]:
    """[(state, first_index, count)] for this rank.

    ``balance="global"`` (default): the fragments of all states form one sequence (state 0's fragments, then state 1's, ...)
    and every rank takes a contiguous 1/world of it -- fragment counts per rank differ by at most one whatever the number
    of states (20 states on 8 GPUs: 250 fragments each), and a rank touches few states, so its launches stay large.
    ``balance="per_state"``: a slice of EVERY state on every rank (``fragment_slice`` per state)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world %d/%d" % (rank, world))
    if balance == "per_state":
        out = []
        for s, n in enumerate(frames_per_state):
            first, cnt = fragment_slice(int(n), world, rank, fragment)
            if cnt:
                out.append((s, first, cnt))
        return out
    frags = [(int(n) + fragment - 1) // fragment for n in frames_per_state]
    total = sum(frags)
    g0, g1 = total * rank // world, total * (rank + 1) // world
    out, base = [], 0
    for s, (n, k) in enumerate(zip(frames_per_state, frags)):
        lo, hi = max(g0, base), min(g1, base + k)
        if lo < hi:
            first = (lo - base) * fragment
            out.append((s, first, min((hi - base) * fragment, int(n)) - first))
        base += k
    return out


COMPACT_ABOVE_BYTES = 64 << 20       # dense tables larger than this go through the compacted exchange


def merge_count_tables(counts, group=None, compact=None, info=None):
    """Sum the per-rank count tables ``int32 [states][pairs]`` in place (``torch.distributed.all_reduce``; NCCL over
    NVLink for CUDA tensors, gloo for the CPU tests).  A no-op when torch.distributed is not initialised (single GPU).

    Large tables (``compact=None``: above 64 MiB; BASELINE config 4 has 22.5 M residue pairs x 32 states = 2.9 GB) are
    almost empty -- a residue touches a few dozen others -- so they are exchanged in compacted form (the candidate-pair
    index SURVEY 8(e) proposes, made exact): (1) one byte per pair "any rank, any state non-zero", all-reduced with MAX;
    (2) the ``[states][K]`` columns of the K flagged pairs, all-reduced with SUM; (3) scattered back.  Integer sums in
    both forms: bit-exact for any rank count.  ``info`` (a dict) receives the bytes each form would move."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1):
        return counts
    dense_bytes = counts.numel() * counts.element_size()
    if compact is None:
        compact = dense_bytes > COMPACT_ABOVE_BYTES
    if info is not None:
        info.update(dense_bytes=int(dense_bytes), compact=bool(compact))
    if not compact or counts.dim() != 2:
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
        return counts
    flags = (counts != 0).any(dim=0).to(torch.uint8)
    dist.all_reduce(flags, op=dist.ReduceOp.MAX, group=group)
    idx = flags.nonzero(as_tuple=True)[0]
    cols = counts.index_select(1, idx).contiguous()
    dist.all_reduce(cols, op=dist.ReduceOp.SUM, group=group)
    counts.index_copy_(1, idx, cols)
    if info is not None:
        info.update(compact_bytes=int(flags.numel() + cols.numel() * cols.element_size()), pairs_nonzero=int(idx.numel()))
    return counts


def bind_to_gpu_numa_node(device: int) -> dict:
    """Pin this process to the CPUs next to ``cuda:device`` (sysfs ``local_cpulist`` of the GPU's PCI function) so that
    pinned host buffers allocated afterwards land on that NUMA node: with one rank per GPU every rank then feeds its
    GPU from local memory instead of all ranks pulling from the node the launcher happened to start on.
    Returns what was done; a no-op (reason given) when the topology is not exposed or has one node."""
    import os
    info = {"bound": False}
    try:
        import torch
        pr = torch.cuda.get_device_properties(device)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        base = "/sys/bus/pci/devices/" + bdf
        node = int(open(base + "/numa_node").read().strip())
        info.update(pci=bdf, node=node)
        if node < 0:
            info["reason"] = "numa_node not reported"
            return info
        cpus = set()
        for part in open(base + "/local_cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = os.sched_getaffinity(0)
        use = cpus & allowed
        if not use or use == allowed:
            info["reason"] = "no narrower CPU set available (%d local, %d allowed)" % (len(cpus), len(allowed))
            return info
        os.sched_setaffinity(0, use)
        info.update(bound=True, cpus=len(use))
    except (OSError, ValueError, AttributeError, RuntimeError, AssertionError) as exc:
        info["reason"] = "%s: %s" % (type(exc).__name__, exc)
    return info

