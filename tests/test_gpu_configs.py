"""Parity on the shapes the published numbers are quoted on (BASELINE.json configs 2, 3 and 5 at full size), through the
C ABI.  Mirrors the reference's own integration test, which asserts exact content on what it runs
(external/getcontacts/tests/test_get_dynamic_contacts.py:11-45).

  config 2  the bench shape itself (bench.make_topology(4): 6 516 atoms, 3 272 pair atoms, bin_frames_tma_kernel<256,13,1> +
            the shared-memory pair path), 200 frames = two fragments at the amplitude of the hottest state; oracle run live
  config 3  16 290 atoms, 101 frames (fragment boundary on the in-place path); committed oracle digest
  config 5  200 k atoms with explicit TIP3 water (grid B, water bridges); committed oracle records of one frame
"""
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

pytestmark = pytest.mark.gpu
ALL = ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"]
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))


def _drain(eng):
    from chacra_b200.engine import NEAR_KINDS
    eng.synchronize()
    rec = eng.drain_contacts()
    near, total = eng.drain_near() if eng.track_near else (np.zeros(0), 0)
    assert total == len(near)
    nears = {(int(r["frame"]), int(r["kind"]), min(int(r["atom1"]), int(r["atom2"])), max(int(r["atom1"]), int(r["atom2"]))) for r in near}
    return rec, nears


def _dense(idx, val, n_pairs):
    t = np.zeros(n_pairs, np.int64)
    t[idx] = val
    return t


def test_config2_bench_shape_against_oracle():
    import torch
    import bench
    from parallel import run_parallel
    import getcontacts_oracle as o
    from chacra_b200 import compile_topology
    from chacra_b200.engine import ContactEngine, ITYPE_NAMES, NEAR_KINDS
    from chacra_b200.synthetic import FrameGenerator
    path, atoms = bench.make_topology(4)
    topo = compile_topology(atoms)                                   # exactly what bench.py builds
    frames = FrameGenerator(atoms.xyz, 2019, amplitude=bench.state_amplitude(19, 20), device="cpu").frames(200).numpy()
    res = run_parallel(path, frames, ALL, collect_near=True)
    eng = ContactEngine(topo, ALL, emit_contacts=True, track_near=True, contact_capacity=1 << 21)
    st = eng.stats()
    assert (st["path"], st["bin_variant"], st["bin_apt"], len(topo.pair_atom)) == (1, 3, 13, 3272)     # the benchmarked instantiation
    eng.submit(torch.from_numpy(frames).cuda(), 0, 0)
    rec, nears = _drain(eng)
    got = {(int(r["frame"]), ITYPE_NAMES[r["itype"]], min(int(r["atom1"]), int(r["atom2"])), max(int(r["atom1"]), int(r["atom2"]))) for r in rec}
    assert len(got) == len(rec)
    want = {(c[0], c[1], min(c[2], c[3]), max(c[2], c[3])) for c in res["contacts"]}
    for it in sorted({c[1] for c in want | got}):
        w, g = {c for c in want if c[1] == it}, {c for c in got if c[1] == it}
        assert g == w, "%s: missing %s extra %s" % (it, sorted(w - g)[:5], sorted(g - w)[:5])
    assert len(want) > 200 * 4000 and {c[0] for c in want} == set(range(200))
    assert nears == {(f, NEAR_KINDS.index(k), min(a, b), max(a, b)) for f, k, a, b, _ in res["near"]}
    ref = o.intended_counts(res)
    rank = {lab: k for k, lab in enumerate(topo.residue_label)}
    table = np.zeros(eng.n_pairs, np.int64)
    for (r1, r2), c in ref.items():
        table[int(eng.pair_index(np.array(rank[r1]), np.array(rank[r2])))] = c
    assert np.array_equal(eng.counts(0).astype(np.int64), table)
    eng.close()
    # the counts-only instantiation the benchmark times (no records, no near channel), device and host entry points
    for host in (False, True):
        e2 = ContactEngine(topo, ALL)
        e2.submit(frames if host else torch.from_numpy(frames).cuda(), 0, 0)
        e2.synchronize()
        assert np.array_equal(e2.counts(0).astype(np.int64), table)
        e2.close()


def test_config3_101_frames_against_committed_oracle_digest(tmp_path):
    import torch
    from make_config_goldens import frames_digest, record_digest
    from chacra_b200 import read_pdb, compile_topology
    from chacra_b200.engine import ContactEngine, ITYPE_NAMES
    from chacra_b200.synthetic import FrameGenerator, tile_pdb
    g = np.load(os.path.join(GOLDEN, "c3_tiled10_101frames.npz"))
    pdb = str(tmp_path / "c3.pdb")
    tile_pdb(os.path.join(GOLDEN, "5xnd_topology.pdb"), pdb, int(g["copies"]))
    atoms = read_pdb(pdb)
    n = int(g["n_frames"])
    frames = FrameGenerator(atoms.xyz, int(g["seed"]), amplitude=float(g["amplitude"]), device="cpu").frames(n).numpy()
    assert frames_digest(frames) == int(g["frames_digest"]), "the synthetic frames are not the ones the golden was made from"
    topo = compile_topology(atoms, frames[0])
    eng = ContactEngine(topo, ALL, emit_contacts=True, track_near=True, contact_capacity=1 << 21)
    assert eng.stats()["path"] == 2
    eng.submit(torch.from_numpy(frames).cuda(), 0, 0)
    rec, nears = _drain(eng)
    per = np.zeros((n, len(ITYPE_NAMES)), np.int32)
    np.add.at(per, (rec["frame"], rec["itype"]), 1)
    assert np.array_equal(per, g["per_frame_itype"])
    dig = record_digest(rec["frame"].astype(np.int64), rec["itype"].astype(np.int64), rec["atom1"].astype(np.int64), rec["atom2"].astype(np.int64),
                        rec["water1"].astype(np.int64), rec["water2"].astype(np.int64), n)
    assert np.array_equal(dig, g["digest"])
    assert np.array_equal(eng.counts(0).astype(np.int64), _dense(g["count_idx"], g["count_val"], eng.n_pairs))
    assert nears == {tuple(int(v) for v in row) for row in g["near"]}
    eng.close()


def test_config5_200k_atoms_against_committed_oracle_records(tmp_path):
    import torch
    from make_config_goldens import frames_digest
    from chacra_b200 import read_pdb, compile_topology
    from chacra_b200.engine import ContactEngine
    from chacra_b200.synthetic import FrameGenerator, solvate_pdb, tile_pdb
    g = np.load(os.path.join(GOLDEN, "c5_solv200k_frame.npz"))
    prot, pdb = str(tmp_path / "prot.pdb"), str(tmp_path / "c5.pdb")
    tile_pdb(os.path.join(GOLDEN, "5xnd_topology.pdb"), prot, 4)
    solvate_pdb(prot, pdb, int(g["target"]))
    atoms = read_pdb(pdb)
    assert atoms.n == int(g["n_atoms"])
    frames = FrameGenerator(atoms.xyz, int(g["seed"]), amplitude=float(g["amplitude"]), device="cpu").frames(1).numpy()
    assert frames_digest(frames) == int(g["frames_digest"]), "the synthetic frame is not the one the golden was made from"
    topo = compile_topology(atoms, frames[0])
    assert topo.n_pair_a < len(topo.pair_atom) and topo.n_solvent_pair > 60000          # grid B with the bulk solvent
    eng = ContactEngine(topo, ALL, emit_contacts=True, track_near=True)
    assert eng.stats()["path"] == 2
    eng.submit(torch.from_numpy(frames).cuda(), 0, 0)
    rec, nears = _drain(eng)

    def key(it, a, b, w1, w2):
        if b < a:
            a, b, w1, w2 = b, a, (w2 if w2 >= 0 else w1), (w1 if w2 >= 0 else w2)
        return (it, a, b, w1, w2)
    got = sorted(key(int(r["itype"]), int(r["atom1"]), int(r["atom2"]), int(r["water1"]), int(r["water2"])) for r in rec)
    want = sorted(key(*(int(v) for v in row[1:])) for row in g["records"])
    assert len(want) > 4000 and any(k[0] == 14 for k in want) and any(k[0] == 15 for k in want)     # wb and wb2 present
    assert got == want
    assert np.array_equal(eng.counts(0).astype(np.int64), _dense(g["count_idx"], g["count_val"], eng.n_pairs))
    assert nears == {tuple(int(v) for v in row) for row in g["near"]}
    eng.close()
