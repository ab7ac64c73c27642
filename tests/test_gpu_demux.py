"""Device-side replica -> state demultiplexing: reading frames through a source index gives exactly the
counts of gathering the same frames first (what get_state_coordinates_from_replica does, process_hremd.py:184-230) --
for the bulk-copy binning kernel, the register and gather variants, and the in-place pair path."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ALL = ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"]


@pytest.fixture(scope="module")
def replicas():
    import torch
    import bench
    from chacra_b200 import compile_topology
    from chacra_b200.synthetic import FrameGenerator
    path, atoms = bench.make_topology(1)
    topo = compile_topology(atoms)
    n_rep, n_frames = 4, 250
    xs = [FrameGenerator(atoms.xyz, 3000 + r, amplitude=1.0 + 0.2 * r, device="cuda").frames(n_frames) for r in range(n_rep)]
    rng = np.random.default_rng(5)
    table = np.stack([rng.permutation(n_rep) for _ in range(n_frames)])
    return topo, torch.stack(xs).contiguous(), table


def _engine(topo, n_states):
    from chacra_b200.engine import ContactEngine
    return ContactEngine(topo, ALL, n_states=n_states)


@pytest.mark.parametrize("variant,force_large", [(None, False), ("0", False), ("1", False), ("4", False), (None, True)])
def test_indexed_submission_equals_gather_then_submit(replicas, monkeypatch, variant, force_large):
    import torch
    from chacra_b200.demux import demux_contacts, state_sources
    topo, reps, table = replicas
    if variant is not None:
        monkeypatch.setenv("CHACRA_BIN_VARIANT", variant)
    if force_large:
        monkeypatch.setenv("CHACRA_FORCE_LARGE", "1")
    n_rep, n_frames = reps.shape[:2]
    eng = _engine(topo, n_rep)
    src = demux_contacts(eng, reps, list(table))
    eng.synchronize()
    got = [eng.counts(s) for s in range(n_rep)]
    eng.close()
    assert np.array_equal(src, state_sources(list(table)))
    ref = _engine(topo, n_rep)
    f = torch.arange(n_frames, device=reps.device)
    for s in range(n_rep):
        state_traj = reps[torch.from_numpy(src[s]).to(reps.device), f].contiguous()       # the state trajectory, materialised
        ref.submit(state_traj, s, 0)
    ref.synchronize()
    for s in range(n_rep):
        want = ref.counts(s)
        assert want.sum() > 0 and np.array_equal(got[s], want)
    ref.close()
    assert not np.array_equal(got[0], got[n_rep - 1])                 # the states really differ (amplitude per replica)


def test_indexed_submission_validates(replicas):
    import torch
    from chacra_b200.engine import ChacraError
    topo, reps, table = replicas
    eng = _engine(topo, 1)
    pool = reps.view(-1, reps.shape[2], 3)
    with pytest.raises(ChacraError, match="outside the pool"):
        eng.submit_indexed(pool, np.array([0, pool.shape[0]]))
    with pytest.raises(ChacraError, match="pool"):
        eng.submit_indexed(pool.cpu(), np.array([0]))
    eng.submit_indexed(pool, np.zeros(0, np.int64))                  # empty: accepted, nothing happens
    eng.submit_indexed(pool, torch.tensor([3, 3, 3], device=pool.device))     # repeats are allowed
    eng.synchronize()
    three = eng.counts(0)
    eng.reset_counts()
    eng.submit(pool[3:4].contiguous(), 0, 0)
    eng.synchronize()
    assert np.array_equal(three, 3 * eng.counts(0))
    eng.close()


def test_replica_files_to_frequency_tables(tmp_path, replicas):
    """r{j}.dcd + state table -> freqs_state_{i}.tsv (the extension script), against state trajectories gathered on the
    host and pushed through the ordinary path; windows that cut the run at fragment boundaries change nothing."""
    import torch
    from test_gpu_dcd import write_dcd
    from chacra_b200.demux import state_sources
    from chacra_b200.frequencies import counts_to_table
    from chacra_b200.getcontacts import get_replica_frequencies
    import bench
    topo, reps, table = replicas
    n_rep, n_frames = reps.shape[:2]
    host = reps.cpu().numpy()
    paths = []
    for r in range(n_rep):
        paths.append(str(tmp_path / ("r%d.dcd" % r)))
        write_dcd(paths[-1], host[r], "<", r % 2 == 0)
    np.save(str(tmp_path / "states.npy"), np.repeat(table, 2, axis=0))          # two cycles per saved frame
    top, _ = bench.make_topology(1)
    out = str(tmp_path / "freqs")
    get_replica_frequencies.main(argv=["--topology", top, "--replicas"] + paths + ["--state_data", str(tmp_path / "states.npy"),
                                       "--save_interval", "2", "--output_dir", out, "--window", "100"])
    src = state_sources(list(table))
    ref = _engine(topo, n_rep)
    for s in range(n_rep):
        ref.submit(np.ascontiguousarray(host[src[s], np.arange(n_frames)]), s, 0)
    ref.synchronize()
    for s in range(n_rep):
        want = counts_to_table(topo.residue_label, ref.counts(s))
        rows = [l.rstrip("\n").split("\t") for l in open(os.path.join(out, "freqs_state_%d.tsv" % s)).read().splitlines()[2:]]
        got = {(a, b): f for a, b, f in rows}
        assert len(got) == len(want) > 50
        for k, c in want.items():
            assert got[k] == "%.3f" % (c / n_frames)
    head = open(os.path.join(out, "freqs_state_0.tsv")).readline()
    assert head.startswith("#\ttotal_frames:%d\t" % n_frames)
    ref.close()


def test_demultiplexed_counts_against_the_oracle(replicas):
    """The device path against the ORACLE, not against itself: 3 replicas x 60 frames of one 5xnd tile; the state
    trajectories are gathered on the host with the reference's index logic (sort_replica_trajectories, pinned in
    tests/test_demux.py) and pushed through the numpy oracle; the indexed submission must give the same count tables."""
    import bench
    import getcontacts_oracle as o
    from parallel import run_parallel
    from chacra_b200.demux import demux_contacts, state_sources
    topo, reps, table = replicas
    reps, table = reps[:3, :60].contiguous(), np.argsort(np.argsort(table[:60, :3], axis=1), axis=1)   # permutations of 0..2
    eng = _engine(topo, 3)
    src = demux_contacts(eng, reps, list(table))
    eng.synchronize()
    assert np.array_equal(src, state_sources(list(table)))
    host = reps.cpu().numpy()
    top, _ = bench.make_topology(1)
    rank = {lab: k for k, lab in enumerate(topo.residue_label)}
    for s in range(3):
        res = run_parallel(top, np.ascontiguousarray(host[src[s], np.arange(60)]), ALL)
        want = np.zeros(eng.n_pairs, np.int64)
        for (r1, r2), c in o.intended_counts(res).items():
            want[int(eng.pair_index(np.array(rank[r1]), np.array(rank[r2])))] = c
        assert want.sum() > 0 and np.array_equal(eng.counts(s).astype(np.int64), want), s
    eng.close()
