"""Water-mediated hydrogen bonds, ligands and dual selections on the GPU: the reference's static 5Y0T interface
golden (external/getcontacts/tests/test_get_static_contacts.py:11-45 and test_5y0t_interface_expected.tsv)."""
import os
import sys
from collections import Counter

import pytest

from conftest import GOLDEN, ROOT, read_golden_tsv

pytestmark = pytest.mark.gpu
GC = os.path.join(ROOT, "chacra_b200", "getcontacts")


def _norm(t):          # the golden file predates the water swap of compute_contacts.py:346-348
    t = tuple(t)
    return t[:4] + tuple(sorted(t[4:]))


def test_static_5y0t_interface_matches_golden(pdb_5y0t, tmp_path):
    if GC not in sys.path:
        sys.path.insert(0, GC)
    import get_static_contacts as cli
    out = tmp_path / "iface.tsv"
    cli.main(["--structure", pdb_5y0t, "--output", str(out), "--itypes", "all", "--sele", "chain A", "--sele2", "chain B"])
    lines = [ln.rstrip("\n") for ln in open(out)]
    assert len(lines) == 271                                    # test_get_static_contacts.py:20
    body = [ln.split("\t") for ln in lines[2:]]
    assert {t[1] for t in body} == {"hbbb", "hbsb", "hbss", "hp", "lwb", "lwb2", "ps", "sb", "vdw", "wb", "wb2"}    # :29
    for t in body:                                              # :33 every contact spans the two chains
        assert {t[2][0], t[3][0]} == {"A", "B"}
    assert len({tuple(t) for t in body}) == len(body)           # :36-43 no duplicates
    gold = read_golden_tsv(os.path.join(GOLDEN, "test_5y0t_interface_expected.tsv"))
    assert Counter(_norm(t) for t in body) == Counter(_norm(t) for t in gold)


def test_static_5y0t_whole_structure_vs_oracle(pdb_5y0t):
    """Single selection (everything bridges with everything): many more waters, wb and wb2 lines than the interface."""
    import torch
    import getcontacts_oracle as o
    from chacra_b200 import read_pdb, compile_topology
    from chacra_b200.engine import ContactEngine, ITYPE_NAMES
    geom = dict(o.DEFAULT_GEOM)
    geom["HBOND_CUTOFF_ANGLE"] = 180
    res = o.run(pdb_5y0t, pdb_5y0t, ["hb"], geom=geom, sele1="protein", sele2="protein")
    atoms = read_pdb(pdb_5y0t)
    topo = compile_topology(atoms, atoms.xyz, "protein", "protein")
    eng = ContactEngine(topo, ["hb"], {"HBOND_CUTOFF_ANGLE": 180}, emit_contacts=True)
    eng.submit(torch.from_numpy(atoms.xyz[None].copy()).cuda(), 0, 0)
    eng.synchronize()
    rec = eng.drain_contacts()
    lab = atoms.atom_labels()

    def canon(it, a, b, ws):
        la, lb = lab[a], lab[b]
        if lb < la:
            la, lb, ws = lb, la, ws[::-1] if len(ws) == 2 else ws
        return (it, la, lb) + tuple(lab[w] for w in ws)
    got = Counter(canon(ITYPE_NAMES[r["itype"]], int(r["atom1"]), int(r["atom2"]), [int(w) for w in (r["water1"], r["water2"]) if w >= 0])
                  for r in rec)
    want = Counter(canon(c[1], c[2], c[3], list(c[4:])) for c in res["contacts"])
    assert sum(1 for k in want if k[0] == "wb2") > 50 and sum(1 for k in want if k[0] == "wb") > 50
    assert set(got) == set(want)
    ref = o.intended_counts(res)
    counts = eng.counts(0)
    assert int(counts.sum()) == sum(ref.values()) and int((counts > 0).sum()) == len(ref)


@pytest.mark.parametrize("n_atoms,force_large", [(6000, False), (6000, True), (30000, False)])
def test_solvated_synthetic_vs_oracle(tmp_path, monkeypatch, n_atoms, force_large):
    """5xnd in a TIP3 droplet (BASELINE.json config-5 recipe, small): every interaction type incl. water bridges,
    through the shared-memory path (second cell list in shared memory) and the in-place path."""
    import torch
    import getcontacts_oracle as o
    from chacra_b200 import read_pdb, compile_topology
    from chacra_b200.engine import ContactEngine, ITYPE_NAMES
    from chacra_b200.synthetic import FrameGenerator, solvate_pdb
    if force_large:
        monkeypatch.setenv("CHACRA_FORCE_LARGE", "1")
    pdb = str(tmp_path / "solv.pdb")
    solvate_pdb(os.path.join(GOLDEN, "5xnd_topology.pdb"), pdb, n_atoms)
    atoms = read_pdb(pdb)
    frames = FrameGenerator(atoms.xyz, 5000, amplitude=0.3, device="cpu").frames(3).numpy()
    its = ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"]
    res = o.run(pdb, pdb, its, frames=frames)
    topo = compile_topology(atoms, frames[0])
    eng = ContactEngine(topo, its, emit_contacts=True)
    assert eng.stats()["path"] == (2 if (force_large or n_atoms > 20000) else 1)
    eng.submit(torch.from_numpy(frames).cuda(), 0, 0)
    eng.synchronize()
    rec = eng.drain_contacts()

    def key(f, it, a, b, ws):
        if b < a:
            a, b, ws = b, a, ws[::-1] if len(ws) == 2 else ws
        return (f, it, a, b) + tuple(ws)
    got = Counter(key(int(r["frame"]), ITYPE_NAMES[r["itype"]], int(r["atom1"]), int(r["atom2"]),
                      [int(w) for w in (r["water1"], r["water2"]) if w >= 0]) for r in rec)
    want = Counter(key(c[0], c[1], c[2], c[3], list(c[4:])) for c in res["contacts"])
    assert any(k[1] == "wb" for k in want) and any(k[1] == "wb2" for k in want)
    assert got == want
    ref = o.intended_counts(res)
    counts = eng.counts(0)
    assert int(counts.sum()) == sum(ref.values()) and int((counts > 0).sum()) == len(ref)
