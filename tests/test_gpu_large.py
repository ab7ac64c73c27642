"""The in-place path (frames too large for shared memory: BASELINE.json configs 3-5) against the oracle."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
ALL = ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"]


def _tiled(tmp_path, copies):
    from chacra_b200.synthetic import tile_pdb
    from chacra_b200 import read_pdb
    pdb = str(tmp_path / ("tiled_%d.pdb" % copies))
    tile_pdb(os.path.join(GOLDEN, "5xnd_topology.pdb"), pdb, copies)
    return pdb, read_pdb(pdb)


def _run(atoms, frames, **kw):
    import torch
    from chacra_b200 import compile_topology
    from chacra_b200.engine import ContactEngine
    topo = compile_topology(atoms, frames[0])
    eng = ContactEngine(topo, ALL, emit_contacts=True, contact_capacity=1 << 22, **kw)
    eng.submit(torch.from_numpy(frames).cuda(), 0, 0)
    eng.synchronize()
    rec = eng.drain_contacts()
    got = {(int(r["frame"]), int(r["itype"]), min(int(r["atom1"]), int(r["atom2"])), max(int(r["atom1"]), int(r["atom2"]))) for r in rec}
    assert len(got) == len(rec)
    return topo, eng, got


def test_config3_shape_16k_atoms(tmp_path):
    """10 tiled copies (16 290 atoms, 8 180 pair atoms, 1 100 residues): the record no longer fits in shared memory."""
    import getcontacts_oracle as o
    from chacra_b200.engine import ITYPE_NAMES
    from chacra_b200.synthetic import FrameGenerator
    pdb, atoms = _tiled(tmp_path, 10)
    frames = FrameGenerator(atoms.xyz, 3000, device="cpu").frames(2).numpy()
    topo, eng, got = _run(atoms, frames)
    assert eng.stats()["path"] == 2
    res = o.run(pdb, pdb, ALL, frames=frames)
    want = {(c[0], ITYPE_NAMES.index(c[1]), min(c[2], c[3]), max(c[2], c[3])) for c in res["contacts"]}
    assert got == want
    ref = o.intended_counts(res)
    counts = eng.counts(0)
    assert int(counts.sum()) == sum(ref.values()) and int((counts > 0).sum()) == len(ref)


def test_config4_shape_100k_atoms_golden(tmp_path):
    """61 tiled copies (99 369 atoms, 6 710 residues -> 22.5 M residue pairs): committed oracle output."""
    g = np.load(os.path.join(GOLDEN, "c4_tiled61_frame.npz"))
    pdb, atoms = _tiled(tmp_path, int(g["copies"]))
    topo, eng, got = _run(atoms, g["frames"])
    assert eng.stats()["path"] == 2
    want = {tuple(int(v) for v in row) for row in g["contacts"]}
    assert got == want
    counts = eng.counts(0)
    assert int(counts.sum()) == int(g["pair_frames"]) and int((counts > 0).sum()) == int(g["n_pairs"])


def test_stamped_pair_set_survives_generation_wrap(monkeypatch, oracle_5xnd):
    """600 frames on 2 CTAs of the in-place path: each CTA passes generation 255 and wipes its table once.
    The 20 fixture frames are cycled, so every count must be exactly 30x the single-pass count."""
    import torch
    from chacra_b200 import read_pdb, compile_topology
    from chacra_b200.engine import ContactEngine
    monkeypatch.setenv("CHACRA_FORCE_LARGE", "1")
    monkeypatch.setenv("CHACRA_STAMPED_SET", "1")          # mid-size systems would keep a plain set in shared memory
    monkeypatch.setenv("CHACRA_PAIR_GRID", "2")
    base = oracle_5xnd["frames"]
    topo = compile_topology(read_pdb(os.path.join(GOLDEN, "5xnd_topology.pdb")), base[0])
    eng = ContactEngine(topo, ALL)
    assert eng.stats()["path"] == 2 and eng.stats()["grid"] == 2
    eng.submit(torch.from_numpy(base).cuda(), 0, 0)
    once = eng.counts(0).astype(np.int64)
    eng.reset_counts()
    eng.submit(torch.from_numpy(np.ascontiguousarray(base[np.arange(600) % 20])).cuda(), 0, 0)
    assert np.array_equal(eng.counts(0).astype(np.int64), 30 * once)
