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
    work = [state_slices([10000] * 20, 8, r, balance="per_state") for r in range(8)]
    assert all(len(w) == 20 for w in work)                     # 20 states on 8 GPUs: nobody idles
    per_rank = [sum(c for _, _, c in w) for w in work]
    assert sum(per_rank) == 200000 and max(per_rank) - min(per_rank) <= 20 * 100


@pytest.mark.parametrize("lens,world", [([10000] * 20, 8), ([250, 1000, 730], 3), ([99, 1, 0, 100, 101], 4), ([20000] * 20, 8), ([5000] * 32, 2), ([130], 5)])
def test_global_state_slices_cover_once_aligned_and_balanced(lens, world):
    seen = [np.zeros(n, np.int32) for n in lens]
    frags = []
    for r in range(world):
        work = state_slices(lens, world, r)
        for s, first, cnt in work:
            assert first % 100 == 0 and cnt > 0
            seen[s][first:first + cnt] += 1
        frags.append(sum((c + 99) // 100 for _, _, c in work))
    assert all((v == 1).all() for v in seen)
    assert max(frags) - min(frags) <= 1
    if lens == [10000] * 20:
        assert frags == [250] * 8 and max(len(state_slices(lens, world, r)) for r in range(world)) <= 3


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


def _compact_worker(rank, world, port, n_states, n_pairs, out_dir):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.Generator(np.random.PCG64(77 + rank))
    counts = torch.zeros((n_states, n_pairs), dtype=torch.int32)
    cols = torch.from_numpy(rng.choice(n_pairs, 40, replace=False))         # a residue touches few others: sparse columns
    counts[:, cols] = torch.from_numpy(rng.integers(0, 100, (n_states, 40)).astype(np.int32))
    np.save(os.path.join(out_dir, "in%d.npy" % rank), counts.numpy())
    dense, info = counts.clone(), {}
    merge_count_tables(dense, compact=False)
    merge_count_tables(counts, compact=True, info=info)
    assert torch.equal(dense, counts) and info["compact"] and info["compact_bytes"] < info["dense_bytes"] / 4
    assert info["pairs_nonzero"] <= 40 * world
    np.save(os.path.join(out_dir, "out%d.npy" % rank), counts.numpy())
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_compacted_exchange_equals_dense_all_reduce(tmp_path, world):
    """The candidate-pair exchange for large tables (config 4: 22.5 M pairs x 32 states) must be the dense integer sum."""
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_compact_worker, args=(world, port, 5, 4001, str(tmp_path)), nprocs=world, join=True)
    want = sum(np.load(tmp_path / ("in%d.npy" % r)).astype(np.int64) for r in range(world))
    for r in range(world):
        assert np.array_equal(np.load(tmp_path / ("out%d.npy" % r)).astype(np.int64), want)


@pytest.mark.parametrize("world", [1, 2, 8])
def test_plan_states_reference_frame_selection(world):
    """Every rank's share of every state covers the reference's frame selection exactly once (compute_contacts.py:247-254:
    clamped beg / end, last fragment NOT clamped to --end)."""
    from chacra_b200.states import plan_states
    lens = [10000, 250, 99, 1, 730]
    for beg, end, stride in ((0, None, 1), (3, 15, 4), (120, 500, 3), (0, 10 ** 9, 7)):
        seen = [np.zeros(n, np.int32) for n in lens]
        for rank in range(world):
            sel, work = plan_states(lens, beg, end, stride, world, rank)
            for s, first, cnt in work:
                assert first % 100 == 0
                b = sel[s][0]
                seen[s][b + first * stride: b + (first + cnt - 1) * stride + 1: stride] += 1
        for s, n in enumerate(lens):
            b = max(min(beg, n - 1), 0)
            e = min(max(10 ** 18 if end is None else end, b), n - 1)
            nfrag = -(-(e - b + 1) // (100 * stride))
            last = min(b + nfrag * 100 * stride - 1, n - 1)
            want = np.zeros(n, np.int32)
            want[b:last + 1:stride] = 1
            assert np.array_equal(seen[s], want), (s, beg, end, stride)
            assert sel[s][2] == -(-(e - b + 1) // stride)
