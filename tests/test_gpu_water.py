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
