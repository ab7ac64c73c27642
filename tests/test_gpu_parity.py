"""CUDA path vs the oracle through the C ABI (run with -m gpu on a B200).

Bit-exact bar: identical contact sets per (frame, itype) and identical integer residue-pair frame
counts (SURVEY A.12 levels L1 and L2); the near-cutoff side channel must list the same observations.
"""
import os
from collections import Counter

import numpy as np
import pytest

from conftest import GOLDEN, read_golden_tsv

pytestmark = pytest.mark.gpu
ALL = ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"]


def _engine_run(top_path, frames, itypes, sele1="protein", sele2=None, geometry=None, host=False, **kw):
    import torch
    from chacra_b200 import read_pdb, compile_topology
    from chacra_b200.engine import ContactEngine, ITYPE_NAMES, NEAR_KINDS
    atoms = read_pdb(top_path)
    g = geometry or {}
    topo = compile_topology(atoms, frames[0], sele1, sele2, vdw_res_diff=g.get("VDW_RES_DIFF", 2),
                            hbond_res_diff=g.get("HBOND_RES_DIFF", 1))
    eng = ContactEngine(topo, itypes, geometry, emit_contacts=True, track_near=True, **kw)
    for s in range(0, len(frames), 300):
        chunk = frames[s:s + 300]
        eng.submit(chunk if host else torch.from_numpy(chunk).cuda(), 0, s)
    eng.synchronize()
    rec = eng.drain_contacts()
    near, near_total = eng.drain_near()
    counts = eng.counts(0)
    contacts = set()
    for r in rec:
        a, b = int(r["atom1"]), int(r["atom2"])
        contacts.add((int(r["frame"]), ITYPE_NAMES[r["itype"]], min(a, b), max(a, b)))
    assert len(contacts) == len(rec), "the kernel emitted a duplicate record"
    nears = {(int(r["frame"]), NEAR_KINDS[r["kind"]], min(int(r["atom1"]), int(r["atom2"])),
              max(int(r["atom1"]), int(r["atom2"]))) for r in near}
    return dict(engine=eng, topo=topo, records=rec, contacts=contacts, near=nears, counts=counts, near_total=near_total)


def _oracle_sets(res):
    contacts = set()
    for c in res["contacts"]:
        if len(c) == 4:
            contacts.add((c[0], c[1], min(c[2], c[3]), max(c[2], c[3])))
    nears = {(f, kind, min(a, b), max(a, b)) for f, kind, a, b, _ in res["near"]}   # the oracle lists (i,j) and (j,i)
    return contacts, nears


def _check_counts(run, oracle_res):
    import getcontacts_oracle as o
    topo, eng = run["topo"], run["engine"]
    want = o.intended_counts(oracle_res)
    rank = {lab: k for k, lab in enumerate(topo.residue_label)}
    table = np.zeros(eng.n_pairs, np.int64)
    for (r1, r2), c in want.items():
        table[int(eng.pair_index(np.array(rank[r1]), np.array(rank[r2])))] = c
    assert np.array_equal(run["counts"].astype(np.int64), table)


def test_5xnd_all_itypes(oracle_5xnd):
    frames = oracle_5xnd["frames"]
    run = _engine_run(os.path.join(GOLDEN, "5xnd_topology.pdb"), frames, ALL)
    want, want_near = _oracle_sets(oracle_5xnd)
    got = run["contacts"]
    for it in sorted({c[1] for c in want | got}):
        w = {c for c in want if c[1] == it}
        g = {c for c in got if c[1] == it}
        assert g == w, "%s: missing %s extra %s" % (it, sorted(w - g)[:5], sorted(g - w)[:5])
    assert run["near"] == want_near
    _check_counts(run, oracle_5xnd)
    # --distout: float32 sqrtf(d2) of the two labelled atoms (compute_contacts.py:161-167)
    for r in run["records"][:2000]:
        d = frames[r["frame"]][r["atom1"]] - frames[r["frame"]][r["atom2"]]
        d2 = np.float32(np.float32(d[0] * d[0] + d[1] * d[1]) + d[2] * d[2])
        assert np.float32(np.sqrt(d2)) == r["distance"]


def test_5xnd_matches_reference_golden_file(oracle_5xnd):
    """Straight against the reference's own expected output (atom labels, not indices)."""
    from chacra_b200.engine import ITYPE_NAMES
    run = _engine_run(os.path.join(GOLDEN, "5xnd_topology.pdb"), oracle_5xnd["frames"], ALL, host=True)
    lab = run["topo"].atoms.atom_labels()
    got = Counter()
    for r in run["records"]:
        a, b = sorted((lab[r["atom1"]], lab[r["atom2"]]))
        got[(str(r["frame"]), ITYPE_NAMES[r["itype"]], a, b)] += 1
    gold = Counter(tuple(x[:4]) for x in read_golden_tsv(os.path.join(GOLDEN, "test_5xnd_expected.tsv.gz")))
    assert got == gold


def test_5xnd_salt_bridges_subset_selection():
    import getcontacts_oracle as o
    top = os.path.join(GOLDEN, "5xnd_topology.pdb")
    res = o.run(top, os.path.join(GOLDEN, "5xnd_trajectory.dcd"), ["sb"], sele1="resname ASP LYS", sele2="resname ASP LYS")
    run = _engine_run(top, res["frames"], ["sb"], "resname ASP LYS", "resname ASP LYS")
    assert run["contacts"] == _oracle_sets(res)[0]
    assert {c[0] for c in run["contacts"]} == {1, 2, 3, 4, 5, 8, 12, 13, 15, 16, 18, 19}


def test_fragment_boundaries_and_chunking(oracle_5xnd):
    """250 frames (the 20 fixture frames cycled) = 3 fragments; the hydrophobic set is rebuilt per fragment."""
    import getcontacts_oracle as o
    base = oracle_5xnd["frames"]
    frames = np.ascontiguousarray(base[np.arange(250) % 20][:, :, :])
    frames[100:200] = base[(np.arange(100) + 7) % 20]
    top = os.path.join(GOLDEN, "5xnd_topology.pdb")
    res = o.run(top, top, ["hp", "vdw", "hb"], frames=frames)
    run = _engine_run(top, frames, ["hp", "vdw", "hb"], host=True)
    assert run["contacts"] == _oracle_sets(res)[0]
    _check_counts(run, res)


def test_binning_kernel_variants_agree(oracle_5xnd, monkeypatch):
    """Every binning kernel (bulk-copy 3: two CTAs per SM, 4: double-buffered with a bulk store; register 1/2; gather 0) must hand the pair kernel the same
    frame: identical contacts and counts over 3 fragments (the fused and the stand-alone hydrophobic query included),
    with device chunks that start on unaligned frame addresses (1 629 atoms: 12 N is not a multiple of 16)."""
    base = oracle_5xnd["frames"]
    frames = np.ascontiguousarray(base[np.arange(250) % 20])
    frames[100:200] = base[(np.arange(100) + 7) % 20]
    top = os.path.join(GOLDEN, "5xnd_topology.pdb")
    runs = {}
    for variant in ("3", "0", "1", "2", "4"):
        monkeypatch.setenv("CHACRA_BIN_VARIANT", variant)
        runs[variant] = _engine_run(top, frames, ALL)
        assert runs[variant]["contacts"] == runs["3"]["contacts"], variant
        assert np.array_equal(runs[variant]["counts"], runs["3"]["counts"]), variant
        assert runs[variant]["near"] == runs["3"]["near"], variant
    # the first 20 frames are the fixture itself
    want, _ = _oracle_sets(oracle_5xnd)
    assert {c for c in runs["3"]["contacts"] if c[0] < 20} == want


@pytest.mark.parametrize("first,last", [(0, 1), (0, 2), (3, 6), (0, 12), (40, 75), (100, 112)])
def test_tiny_and_odd_sized_systems(tmp_path, oracle_5xnd, first, last):
    """Cut-outs of the 5xnd fixture (residues first..last-1: from one residue to a few dozen; fewer pair atoms than a
    warp, a handful of cells, possibly no ring / ion / hydrogen-bond partner at all) against the oracle, on both pair
    paths and two binning variants."""
    import getcontacts_oracle as o
    src = os.path.join(GOLDEN, "5xnd_topology.pdb")
    keep, lines, seen = [], [], []
    k = 0
    for ln in open(src):
        if ln.startswith(("ATOM  ", "HETATM")):
            key = (ln[21], ln[22:27])
            if key not in seen:
                seen.append(key)
            if first <= seen.index(key) < last:
                keep.append(k)
                lines.append(ln)
            k += 1
    top = str(tmp_path / "cut.pdb")
    open(top, "w").write("".join(lines) + "END\n")
    frames = np.ascontiguousarray(oracle_5xnd["frames"][:, keep, :])
    res = o.run(top, top, ALL, frames=frames, collect_near=True)
    want, want_near = _oracle_sets(res)
    for env in ({}, {"CHACRA_FORCE_LARGE": "1"}, {"CHACRA_FORCE_LARGE": "1", "CHACRA_STAMPED_SET": "1", "CHACRA_RING_GRID": "1"}, {"CHACRA_BIN_VARIANT": "0"},
                {"CHACRA_BIN_VARIANT": "0", "CHACRA_COUNT16": "1"}, {"CHACRA_BIN_VARIANT": "4"}):
        os.environ.update(env)
        try:
            run = _engine_run(top, frames, ALL)
        finally:
            for name in env:
                os.environ.pop(name, None)
        assert run["contacts"] == want, (env, sorted(want - run["contacts"])[:3], sorted(run["contacts"] - want)[:3])
        assert run["near"] == want_near
        _check_counts(run, res)


def test_far_from_the_origin_coarse_float_grid(oracle_5xnd):
    """The same frames shifted by thousands of Angstrom: float32 coordinates then sit on a 5e-4 A grid, exact distance
    ties and near-cutoff events become common -- the float32 recipes must still agree bit for bit."""
    import getcontacts_oracle as o
    top = os.path.join(GOLDEN, "5xnd_topology.pdb")
    frames = (oracle_5xnd["frames"][:8].astype(np.float64) + np.array([4000.0, -7000.0, 9000.0])).astype(np.float32)
    res = o.run(top, top, ALL, frames=frames, collect_near=True)
    want, want_near = _oracle_sets(res)
    run = _engine_run(top, frames, ALL)
    assert run["contacts"] == want, (sorted(want - run["contacts"])[:3], sorted(run["contacts"] - want)[:3])
    assert run["near"] == want_near and len(want_near) > 0
    _check_counts(run, res)


@pytest.mark.parametrize("change", [
    {"SALT_BRIDGE_CUTOFF_DISTANCE": 8.0, "HBOND_CUTOFF_DISTANCE": 5.0, "VDW_EPSILON": 2.0},          # wide: larger cells
    {"SALT_BRIDGE_CUTOFF_DISTANCE": 2.5, "HBOND_CUTOFF_DISTANCE": 2.8, "VDW_EPSILON": 0.0, "HBOND_CUTOFF_ANGLE": 20},
    {"HBOND_CUTOFF_ANGLE": 180, "PI_CATION_CUTOFF_ANGLE": 90, "PI_STACK_CUTOFF_ANGLE": 90, "T_STACK_CUTOFF_ANGLE": 90,
     "PI_STACK_PSI_ANGLE": 90, "T_STACK_PSI_ANGLE": 90, "PI_CATION_CUTOFF_DISTANCE": 9.5, "PI_STACK_CUTOFF_DISTANCE": 9.5,
     "T_STACK_CUTOFF_DISTANCE": 6.0},
    {"HBOND_RES_DIFF": 0, "VDW_RES_DIFF": 0},
    {"HBOND_RES_DIFF": 4, "VDW_RES_DIFF": 6, "HBOND_CUTOFF_ANGLE": 0},
])
def test_non_default_geometric_criteria(oracle_5xnd, change):
    """Every --*_cutoff_* / --*_res_diff flag of the reference's parser (argparsers.py:80-125) away from its default."""
    import getcontacts_oracle as o
    top = os.path.join(GOLDEN, "5xnd_topology.pdb")
    frames = np.ascontiguousarray(oracle_5xnd["frames"][:5])
    geom = dict(o.DEFAULT_GEOM)
    geom.update(change)
    res = o.run(top, top, ALL, geom=geom, frames=frames, collect_near=True)
    want, want_near = _oracle_sets(res)
    run = _engine_run(top, frames, ALL, geometry=geom)
    for it in sorted({c[1] for c in want | run["contacts"]}):
        w = {c for c in want if c[1] == it}
        g = {c for c in run["contacts"] if c[1] == it}
        assert g == w, "%s: missing %s extra %s" % (it, sorted(w - g)[:5], sorted(g - w)[:5])
    assert run["near"] == want_near
    _check_counts(run, res)


def test_unequal_overlapping_selections_over_two_fragments(oracle_5xnd):
    """--sele and --sele2 different and overlapping, 130 frames = two fragments: the stand-alone hydrophobic-query kernel
    (subset rule for sele1, equality rule for sele2, hydrophobics.py:24,39) and both directions of every pair rule."""
    import getcontacts_oracle as o
    top = os.path.join(GOLDEN, "5xnd_topology.pdb")
    base = oracle_5xnd["frames"]
    frames = np.ascontiguousarray(base[(np.arange(130) * 3) % 20])
    frames[100:] = base[(np.arange(30) + 11) % 20]
    s1, s2 = "protein and resid 1 to 60", "protein and resid 45 to 112"
    res = o.run(top, top, ALL, frames=frames, sele1=s1, sele2=s2, collect_near=True)
    want, want_near = _oracle_sets(res)
    for env in ({}, {"CHACRA_FORCE_LARGE": "1"}):
        os.environ.update(env)
        try:
            run = _engine_run(top, frames, ALL, s1, s2)
        finally:
            for name in env:
                os.environ.pop(name, None)
        for it in sorted({c[1] for c in want | run["contacts"]}):
            w = {c for c in want if c[1] == it}
            g = {c for c in run["contacts"] if c[1] == it}
            assert g == w, "%s %s: missing %s extra %s" % (env, it, sorted(w - g)[:5], sorted(g - w)[:5])
        assert run["near"] == want_near
        _check_counts(run, res)


def test_crafted_frames_on_every_plain_distance_cutoff(oracle_5xnd):
    """Near-cutoff channel for the distance cutoffs whose <= / < direction SURVEY 8(c) calls weakly pinned: one atom of the
    5xnd fixture is slid until a chosen pair sits 3e-5 A below / above the salt-bridge cutoff (salt_bridges.py:76-80, 4.0 A),
    the hydrogen-bond distance (hbonds.py:75-82, 3.5 A) and the soft `measure contacts` radii of pc / ps / ts
    (pi_cation.py:22,70; aromatics.py:26-27,166).  The kernel must list the same observations as the oracle and decide the
    contacts the same way on both sides of every boundary."""
    import itertools
    import getcontacts_oracle as o
    from conftest import tune_pair_distance
    top = os.path.join(GOLDEN, "5xnd_topology.pdb")
    x = oracle_5xnd["frames"][0]
    T, S = oracle_5xnd["top"], oracle_5xnd["setup"]
    an, ca, ring = np.nonzero(S.anion)[0], np.nonzero(S.cation)[0], np.nonzero(S.ring)[0]
    d = np.linalg.norm(x[an][:, None] - x[ca][None], axis=-1)
    k = np.argwhere((d > 3.2) & (d < 4.8))[0]
    sb = (int(an[k[0]]), int(ca[k[1]]))
    hb = next((c[2], c[3]) for c in oracle_5xnd["contacts"] if c[0] == 0 and c[1] == "hbbb")
    pc = None
    for c2 in np.nonzero(S.prot_cation)[0]:
        for r in ring:
            tri = o._aromatic_triplet(S, int(r))
            if len(tri) != 3 or T.reslabel(int(c2)) == T.reslabel(int(r)):
                continue
            dd = [float(np.linalg.norm(x[c2] - x[t])) for t in tri]
            cen = np.mean([x[t] for t in tri], axis=0)
            if 9.0 < max(dd) < 9.9 and np.linalg.norm(x[c2] - cen) < 9.3:
                pc = (int(c2), tri[int(np.argmax(dd))])
                break
        if pc:
            break
    closest = {}
    for i, j in itertools.combinations([int(r) for r in ring], 2):
        ri, rj = T.reslabel(i), T.reslabel(j)
        if ri != rj:
            dd = float(np.linalg.norm(x[i] - x[j]))
            if (ri, rj) not in closest or dd < closest[(ri, rj)][0]:
                closest[(ri, rj)] = (dd, i, j)
    ps = next(v for v in closest.values() if 9.0 < v[0] < 9.8)
    ts = next(v for v in closest.values() if 5.2 < v[0] < 5.9)
    frames = []
    for (i, j), target in ((sb, 4.0), (hb, 3.5), (pc, 10.0), (ps[1:], 10.0), (ts[1:], 6.0)):
        frames += [tune_pair_distance(x, i, j, target, False), tune_pair_distance(x, i, j, target, True)]
    frames = np.stack(frames)
    geom = dict(o.DEFAULT_GEOM)
    geom.update({"PI_CATION_CUTOFF_DISTANCE": 9.5, "PI_STACK_CUTOFF_DISTANCE": 12.0, "T_STACK_CUTOFF_DISTANCE": 9.0})
    res = o.run(top, top, ALL, geom=geom, frames=frames, collect_near=True)
    want, want_near = _oracle_sets(res)
    kinds = {(n[0], n[1]) for n in want_near}
    assert {(0, "sb_dist"), (1, "sb_dist"), (2, "hb_dist"), (3, "hb_dist"), (4, "pc_soft"), (5, "pc_soft"), (6, "ps_soft"),
            (7, "ps_soft"), (8, "ts_soft"), (9, "ts_soft")} <= kinds
    assert (0, "sb", min(sb), max(sb)) in want and (1, "sb", min(sb), max(sb)) not in want      # the boundary really decides
    for env in ({}, {"CHACRA_FORCE_LARGE": "1"}):
        os.environ.update(env)
        try:
            run = _engine_run(top, frames, ALL, geometry=geom)
        finally:
            for name in env:
                os.environ.pop(name, None)
        assert run["contacts"] == want, (env, sorted(want - run["contacts"])[:3], sorted(run["contacts"] - want)[:3])
        assert run["near"] == want_near, (env, sorted(want_near - run["near"])[:5], sorted(run["near"] - want_near)[:5])
        _check_counts(run, res)
    # only the sb / hb types enabled: the prefilter radius is then the cutoff itself and must still cover the band above it
    for its in (["sb"], ["hb"]):
        r2 = o.run(top, top, its, geom=geom, frames=frames[:4], collect_near=True)
        w2, n2 = _oracle_sets(r2)
        run = _engine_run(top, frames[:4], its, geometry=geom)
        assert run["contacts"] == w2 and run["near"] == n2 and len(n2) >= 2


def _pdb_with_zinc(tmp_path, oracle_5xnd):
    """The 5xnd fixture plus one HETATM zinc ion 2.1 A from an aspartate carboxylate oxygen: a charged ligand inside the
    selection (contact_utils.py:605-665 makes metal-labelled ligand atoms cations)."""
    x = oracle_5xnd["frames"]
    T, S = oracle_5xnd["top"], oracle_5xnd["setup"]
    od = next(int(i) for i in np.nonzero(S.anion)[0] if T.resname[i] == "ASP")
    cg = next(int(i) for i in range(T.n) if T.residue[i] == T.residue[od] and T.name[i] == "CG")
    u = x[0][od] - x[0][cg]
    u /= np.linalg.norm(u)
    zn0 = x[0][od] + 2.1 * u
    lines = [ln for ln in open(os.path.join(GOLDEN, "5xnd_topology.pdb")) if ln.startswith(("ATOM  ", "HETATM"))]
    lines.append("HETATM%5d ZN    ZN A 900    %8.3f%8.3f%8.3f  1.00  0.00          ZN\n" % (len(lines) + 1, zn0[0], zn0[1], zn0[2]))
    top = str(tmp_path / "zn.pdb")
    open(top, "w").write("".join(lines) + "END\n")
    # the ion rides on the carboxylate oxygen in every frame
    zn = x[:, od, :] + (zn0 - x[0][od])[None].astype(np.float32)
    return top, np.ascontiguousarray(np.concatenate([x, zn[:, None, :]], axis=1), dtype=np.float32)


def test_charged_ligand_salt_bridges_emit_the_plain_sb_line_too(tmp_path, oracle_5xnd):
    """salt_bridges.py:90-97 (if / if / if-else): a protein-ligand salt bridge is written as `sbpl` (or `sblp`) AND as a plain
    `sb` line -- the one get_contact_frequencies counts.  Contact list, count table and the text file against the oracle."""
    import getcontacts_oracle as o
    top, frames = _pdb_with_zinc(tmp_path, oracle_5xnd)
    sel = "protein or resname ZN"
    res = o.run(top, top, ["sb", "vdw", "hb"], frames=frames, sele1=sel, sele2=sel, collect_near=True)
    want, want_near = _oracle_sets(res)
    assert any(c[1] == "sbpl" for c in want) and sum(c[1] == "sb" for c in want) > sum(c[1] == "sbpl" for c in want)
    for env in ({}, {"CHACRA_FORCE_LARGE": "1"}):
        os.environ.update(env)
        try:
            run = _engine_run(top, frames, ["sb", "vdw", "hb"], sel, sel)
        finally:
            for name in env:
                os.environ.pop(name, None)
        for it in sorted({c[1] for c in want | run["contacts"]}):
            w, g = {c for c in want if c[1] == it}, {c for c in run["contacts"] if c[1] == it}
            assert g == w, "%s %s: missing %s extra %s" % (env, it, sorted(w - g)[:5], sorted(g - w)[:5])
        assert run["near"] == want_near
        _check_counts(run, res)


def test_unaligned_growing_submissions_keep_the_hydrophobic_set(oracle_5xnd):
    """chc_submit_frames may continue a fragment; the mask buffer then grows while slot 0 must keep the continued fragment's
    hydrophobic set (250 frames at index 0, then 350 at index 250: the second call needs more fragment slots than the
    first).  Device tensors, every binning variant with its own mask path; against one aligned submission."""
    import torch
    from chacra_b200 import read_pdb, compile_topology
    from chacra_b200.engine import ContactEngine
    base = oracle_5xnd["frames"]
    frames = np.ascontiguousarray(base[(np.arange(600) * 7) % 20])
    for k in range(6):                                   # every fragment opens on a different frame -> different hp sets
        frames[100 * k] = base[(3 * k + 1) % 20]
    x = torch.from_numpy(frames).cuda()
    topo = compile_topology(read_pdb(os.path.join(GOLDEN, "5xnd_topology.pdb")), frames[0])
    for variant in ("3", "0"):
        os.environ["CHACRA_BIN_VARIANT"] = variant
        try:
            outs = []
            for cuts in ([0, 600], [0, 250, 600], [0, 30, 130, 131, 460, 600]):
                eng = ContactEngine(topo, ["hp", "vdw"], emit_contacts=True, contact_capacity=1 << 21)
                for a, b in zip(cuts[:-1], cuts[1:]):
                    eng.submit(x[a:b], 0, a)
                eng.synchronize()
                rec = eng.drain_contacts()
                # atom_1 / atom_2 of a record follow the cell-sorted order, which is not fixed inside a cell: compare unordered
                key = np.stack([rec["frame"], rec["itype"], np.minimum(rec["atom1"], rec["atom2"]), np.maximum(rec["atom1"], rec["atom2"])], 1)
                outs.append((eng.counts(0), key[np.lexsort(key.T[::-1])]))
                eng.close()
            assert (outs[0][1][:, 1] == 18).sum() > 600 * 50             # hp records present in every fragment
            for got in outs[1:]:
                assert np.array_equal(got[0], outs[0][0]) and np.array_equal(got[1], outs[0][1]), variant
        finally:
            os.environ.pop("CHACRA_BIN_VARIANT", None)


def test_fused_kernel_option_matches_oracle(oracle_5xnd, monkeypatch):
    """CHACRA_FUSED=1: the pair kernel sorts every frame itself from the caller's coordinates (no binned record; only the
    fragment-opening frames pass through the binning kernel for the hydrophobic masks).  Same contacts, counts and near-cutoff
    list over three fragments, device and host entry points."""
    import getcontacts_oracle as o
    monkeypatch.setenv("CHACRA_FUSED", "1")
    base = oracle_5xnd["frames"]
    frames = np.ascontiguousarray(base[np.arange(250) % 20])
    frames[100:200] = base[(np.arange(100) + 7) % 20]
    top = os.path.join(GOLDEN, "5xnd_topology.pdb")
    res = o.run(top, top, ALL, frames=frames, collect_near=True)
    want, want_near = _oracle_sets(res)
    for host in (False, True):
        run = _engine_run(top, frames, ALL, host=host)
        assert run["engine"].stats()["fused"] == 1
        assert run["contacts"] == want and run["near"] == want_near
        _check_counts(run, res)


def test_ring_task_list_overflow_falls_back_to_every_task(oracle_5xnd, monkeypatch):
    """The ring prefilter lists the (cation, ring) / (ring, ring) tasks worth a float64 evaluation; a list that overflows
    (CHACRA_RLIST_CAP=4 here) must make the kernel evaluate every task instead -- same pc / ps / ts contacts, on the
    shared-memory path, on the in-place path and on the in-place path with the coarse ring grid."""
    import getcontacts_oracle as o
    top = os.path.join(GOLDEN, "5xnd_topology.pdb")
    frames = oracle_5xnd["frames"]
    res = o.run(top, top, ALL, frames=frames, collect_near=True)
    want, want_near = _oracle_sets(res)
    monkeypatch.setenv("CHACRA_RLIST_CAP", "4")
    for env in ({}, {"CHACRA_FORCE_LARGE": "1"}, {"CHACRA_FORCE_LARGE": "1", "CHACRA_RING_GRID": "1"}):
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        run = _engine_run(top, frames, ALL)
        for k in env:
            monkeypatch.delenv(k)
        assert run["contacts"] == want and run["near"] == want_near, env
        _check_counts(run, res)

