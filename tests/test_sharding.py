"""Multi-rank plumbing on the CPU (gloo, world_size 2 and 3): fragment-aligned slices and the count all-reduce."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from chacra_b200.sharding import fragment_slice, merge_count_tables, state_slices


@pytest.mark.parametrize("n_frames,world", [(10000, 8), (20000, 8), (5000, 4), (250, 2), (99, 3), (100, 3), (1, 2), (0, 2), (730, 8)])
def test_slices_cover_disjoint_and_aligned(n_frames, world):
    seen = np.zeros(n_frames, np.int32)
    frags = []
    for r in range(world):
        first, cnt = fragment_slice(n_frames, world, r)
        assert first % 100 == 0 or cnt == 0
        seen[first:first + cnt] += 1
        frags.append((cnt + 99) // 100)
    assert (seen == 1).all()
    assert max(frags) - min(frags) <= 1


def test_state_slices_every_state_on_every_rank():
    work = [state_slices([10000] * 20, 8, r) for r in range(8)]
    assert all(len(w) == 20 for w in work)                     # 20 states on 8 GPUs: nobody idles
    per_rank = [sum(c for _, _, c in w) for w in work]
    assert sum(per_rank) == 200000 and max(per_rank) - min(per_rank) <= 20 * 100


def _fragment_counts(state, fragment, n_pairs):
    """Stand-in for the pair kernel: a deterministic integer contribution of one (state, fragment) work item."""
    rng = np.random.Generator(np.random.PCG64(1000 * state + fragment))
    return rng.integers(0, 100, n_pairs, dtype=np.int32)


def _worker(rank, world, port, frames_per_state, n_pairs, out_dir):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    counts = torch.zeros((len(frames_per_state), n_pairs), dtype=torch.int32)
    for state, first, cnt in state_slices(frames_per_state, world, rank):
        for frag in range(first // 100, (first + cnt + 99) // 100):
            counts[state] += torch.from_numpy(_fragment_counts(state, frag, n_pairs))
    merge_count_tables(counts)
    np.save(os.path.join(out_dir, "rank%d.npy" % rank), counts.numpy())
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_count_tables_merge_bit_exact(tmp_path, world):
    frames_per_state, n_pairs = [250, 1000, 730], 501
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(world, port, frames_per_state, n_pairs, str(tmp_path)), nprocs=world, join=True)
    want = np.zeros((len(frames_per_state), n_pairs), np.int64)
    for s_, n in enumerate(frames_per_state):
        for frag in range((n + 99) // 100):
            want[s_] += _fragment_counts(s_, frag, n_pairs)
    for r in range(world):
        assert np.array_equal(np.load(tmp_path / ("rank%d.npy" % r)).astype(np.int64), want)


def test_merge_is_noop_without_process_group():
    t = torch.arange(6, dtype=torch.int32)
    assert merge_count_tables(t) is t and t.tolist() == [0, 1, 2, 3, 4, 5]
