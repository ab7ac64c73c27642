"""The C++/OpenMP CPU twin (oracle/getcontacts_cpu.cpp: bench.py's cpu_baseline and --impl reference arm) against the numpy
oracle, which is itself pinned to the reference's golden contact lists (tests/test_oracle_golden.py): identical contact
multisets per (frame, itype) and identical residue-pair frame counts."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

ALL = ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"]


def _sets(rec):
    from cpu_twin import ITYPE_NAMES
    return {(int(f), ITYPE_NAMES[it], int(a), int(b)) for f, it, a, b in rec}


def _oracle_sets(res):
    return {(c[0], c[1], min(c[2], c[3]), max(c[2], c[3])) for c in res["contacts"]}


def test_twin_matches_oracle_on_5xnd(oracle_5xnd):
    import getcontacts_oracle as o
    from cpu_twin import Twin
    tw = Twin(os.path.join(GOLDEN, "5xnd_topology.pdb"), xyz0=oracle_5xnd["frames"][0])
    rec, counts = tw.run(oracle_5xnd["frames"])
    got, want = _sets(rec), _oracle_sets(oracle_5xnd)
    assert len(got) == len(rec)
    for it in sorted({c[1] for c in want | got}):
        w, g = {c for c in want if c[1] == it}, {c for c in got if c[1] == it}
        assert g == w, "%s: missing %s extra %s" % (it, sorted(w - g)[:5], sorted(g - w)[:5])
    assert tw.count_dict(counts) == o.intended_counts(oracle_5xnd)
    _, again = tw.run(oracle_5xnd["frames"], want_contacts=False, threads=1)     # counts-only leg (what the bench times), one thread
    assert np.array_equal(again, counts)


def test_twin_fragments_selections_and_geometry(oracle_5xnd):
    import getcontacts_oracle as o
    from cpu_twin import Twin
    top = os.path.join(GOLDEN, "5xnd_topology.pdb")
    base = oracle_5xnd["frames"]
    frames = np.ascontiguousarray(base[(np.arange(130) * 3) % 20])
    frames[100:] = base[(np.arange(30) + 11) % 20]
    s1, s2 = "protein and resid 1 to 60", "protein and resid 45 to 112"
    geom = dict(o.DEFAULT_GEOM)
    geom.update({"HBOND_CUTOFF_ANGLE": 120, "VDW_RES_DIFF": 3, "PI_STACK_CUTOFF_DISTANCE": 9.0, "T_STACK_CUTOFF_ANGLE": 60,
                 "PI_CATION_CUTOFF_DISTANCE": 8.0, "SALT_BRIDGE_CUTOFF_DISTANCE": 5.0})
    res = o.run(top, top, ALL, geom=geom, frames=frames, sele1=s1, sele2=s2)
    tw = Twin(top, s1, s2, xyz0=frames[0], geom=geom)
    rec, counts = tw.run(frames)
    got, want = _sets(rec), _oracle_sets(res)
    for it in sorted({c[1] for c in want | got}):
        w, g = {c for c in want if c[1] == it}, {c for c in got if c[1] == it}
        assert g == w, "%s: missing %s extra %s" % (it, sorted(w - g)[:5], sorted(g - w)[:5])
    assert tw.count_dict(counts) == o.intended_counts(res)


def test_twin_on_the_bench_shape():
    """Config-2 shape (4 tiled copies, 6 516 atoms): 8 frames against the numpy oracle."""
    import bench
    from parallel import run_parallel
    import getcontacts_oracle as o
    from cpu_twin import Twin
    from chacra_b200.synthetic import FrameGenerator
    path, atoms = bench.make_topology(4)
    frames = FrameGenerator(atoms.xyz, 2019, amplitude=bench.state_amplitude(19, 20), device="cpu").frames(8).numpy()
    res = run_parallel(path, frames, ALL)
    tw = Twin(path, xyz0=atoms.xyz)
    rec, counts = tw.run(frames)
    assert _sets(rec) == _oracle_sets(res)
    assert tw.count_dict(counts) == o.intended_counts(res)
