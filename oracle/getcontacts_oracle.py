"""CPU oracle: a numpy restatement of getcontacts' contact detection (TEST INFRASTRUCTURE ONLY).

This file is the *checker* for the CUDA path in ``chacra_b200``.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may
import it; the product package never does.

What it restates (citations relative to ``/root/reference/external/getcontacts``):

* ``contact_calc/compute_contacts.py:50-103``   per-frame dispatch order sb, pc, ps, ts, vdw, hb, hp
* ``contact_calc/compute_contacts.py:187-293``  setup, frame clamping, 100-frame fragments
* ``contact_calc/compute_contacts.py:306-359``  canonicalisation + per-fragment de-duplication + TSV
* ``contact_calc/salt_bridges.py:33-99``, ``pi_cation.py:34-129``, ``aromatics.py:116-257``,
  ``vanderwaals.py:24-96``, ``hbonds.py:29-206``, ``hydrophobics.py:8-123``
* ``contact_calc/atom.py:5-118`` labels, vdW radii, backbone names, element inference
* ``contact_calc/contact_utils.py:396-409`` disulfides, ``:411-579`` solv/lipid/ligand macros,
  ``:605-665`` ligand ions, ``:958-1061`` float64 ring geometry

The geometry primitives live in VMD (vmd-python 3.1.4, third party, not under /root/reference).
Their published behaviour is restated from SURVEY.md Appendix A (``measure contacts``,
``measure hbonds``, ``within``, ``measure bond``, the distance-based bond search and the
selection keywords).  PARITY PIN: this oracle is checked against the reference's own golden
contact lists ``tests/test_5xnd_expected.tsv`` (22 587 contacts, 20 frames, 7 itypes) and
``tests/test_5y0t_interface_expected.tsv`` (269 contacts incl. water bridges) and the line-count /
frame-set assertions of ``tests/test_get_dynamic_contacts.py`` / ``test_get_static_contacts.py``
(see ``tests/test_oracle_golden.py``).

Arithmetic recipe (the part that decides ties): float32 coordinates; ``d2 = (dx*dx + dy*dy) +
dz*dz`` in float32 with separate multiply/add; ``sqrtf`` in float32; hbond angle
``57.2958f * (float)atan2((double)|u x v|, (double)(u.v))`` with u x v, u.v in float32; ring
geometry in float64 on the float32 values.
"""
from __future__ import annotations

import math
import re
import struct
from collections import defaultdict
from itertools import product

import numpy as np

try:  # candidate generation only; every candidate is re-tested with the exact float32 recipe
    from scipy.spatial import cKDTree
except Exception:  # pragma: no cover
    cKDTree = None

F32 = np.float32
TRAJ_FRAG_SIZE = 100  # compute_contacts.py:43

# ----------------------------------------------------------------------------------------------
# atom.py restatement
# ----------------------------------------------------------------------------------------------
BB_NAMES = {"O", "N", "OP1", "OP2", "O1P", "O2P", "O3'", "O2'", "O4'", "O5'"}  # atom.py:107
VDW_RADII = {  # atom.py:111-118
    'H': 1.20, 'HE': 1.40, 'LI': 1.82, 'BE': 1.53, 'B': 1.92, 'C': 1.70, 'N': 1.55, 'O': 1.52, 'F': 1.47,
    'NE': 1.54, 'NA': 2.27, 'MG': 1.73, 'AL': 1.84, 'SI': 2.10, 'P': 1.80, 'S': 1.80, 'CL': 1.75, 'AR': 1.88,
    'K': 2.75, 'CA': 2.31, 'SC': 2.11, 'NI': 1.63, 'CU': 1.40, 'ZN': 1.39, 'GA': 1.87, 'GE': 2.11, 'AS': 1.85,
    'SE': 1.90, 'BR': 1.85, 'KR': 2.02, 'RB': 3.03, 'SR': 2.49, 'PD': 1.63, 'AG': 1.72, 'CD': 1.58, 'IN': 1.93,
    'SN': 2.17, 'SB': 2.06, 'TE': 2.06, 'I': 1.98, 'XE': 2.16, 'CS': 3.43, 'BA': 2.68, 'PT': 1.75, 'AU': 1.66,
    'HG': 1.55, 'TL': 1.96, 'PB': 2.02, 'BI': 2.07, 'PO': 1.97, 'AT': 2.02, 'RN': 2.20, 'FR': 3.48, 'RA': 2.83,
    'U': 1.86, 'GD': 2.79}
ELEMENT_NAMES = set(
    "H HE LI BE B C N O F NE NA MG AL SI P S CL AR K CA SC TI V CR MN FE CO NI CU ZN GA GE AS SE BR KR RB SR Y "
    "ZR NB MO TC RU RH PD AG CD IN SN SB TE I XE CS BA LA CE PR ND PM SM EU GD TB DY HO ER TM YB LU HF TA W RE "
    "OS IR PT AU HG TL PB BI PO AT RN FR RA AC TH PA U NP PU AM CM BK CF ES FM MD NO LR RF DB SG BH HS MT DS RG "
    "CN NH FL MC LV TS OG".split())  # atom.py:122-145 (keys)


def infer_element(resname: str, name: str) -> str:
    """atom.py:54-103."""
    m = re.search(r'[a-zA-Z]+', name)
    if m is None:
        return "?"
    letters = m.group(0).upper()
    if len(letters) == 1:
        return letters if letters in ELEMENT_NAMES else "?"
    e1, e2 = letters[0], letters[0:2]
    if e1 in ELEMENT_NAMES and e2 in ELEMENT_NAMES:
        return e2 if resname == name else e1
    if e1 in ELEMENT_NAMES:
        return e1
    if e2 in ELEMENT_NAMES:
        return e2
    return "?"


# ----------------------------------------------------------------------------------------------
# Topology (PDB) and VMD-side static state (Appendix A.2, A.3)
# ----------------------------------------------------------------------------------------------
class Topology:
    def __init__(self, path: str):
        names, resnames, chains, resids, icodes, elems, segs, alts, xyz, serials = [], [], [], [], [], [], [], [], [], []
        conect = []
        with open(path) as fh:
            for line in fh:
                rec = line[0:6]
                if rec in ("ATOM  ", "HETATM"):
                    serials.append(line[6:11].strip())
                    names.append(line[12:16].strip())
                    alts.append(line[16:17].strip())
                    resnames.append(line[17:21].strip())
                    chains.append(line[21:22].strip())
                    resids.append(int(line[22:26]))
                    icodes.append(line[26:27].strip())
                    xyz.append((float(line[30:38]), float(line[38:46]), float(line[46:54])))
                    segs.append(line[72:76].strip() if len(line) > 72 else "")
                    elems.append(line[76:78].strip() if len(line) > 76 else "")
                elif rec == "CONECT":
                    f = [line[6 + 5 * k: 11 + 5 * k].strip() for k in range(5)]
                    f = [x for x in f if x]
                    if len(f) >= 2:
                        conect.append(f)
                elif rec.startswith("END") and rec.strip() in ("END", "ENDMDL"):
                    if rec.strip() == "ENDMDL":
                        break
        self.n = n = len(names)
        self.name = np.array(names, dtype=object)
        self.resname = np.array(resnames, dtype=object)
        self.chain = np.array(chains, dtype=object)
        self.resid = np.array(resids, dtype=np.int64)
        self.icode = np.array(icodes, dtype=object)
        self.segname = np.array(segs, dtype=object)
        self.altloc = np.array(alts, dtype=object)
        self.xyz0 = np.array(xyz, dtype=np.float64).astype(F32)  # VMD stores float32
        # atom.py:25-35: upper-case; "X" (VMD's value for a missing element column) -> inferred
        el = []
        for i in range(n):
            e = elems[i].upper() if elems[i] else "X"
            if e == "X":
                e = infer_element(resnames[i], names[i])
            el.append(e)
        self.element = np.array(el, dtype=object)
        self.vdw = np.array([VDW_RADII.get(e, 1.7) for e in el], dtype=np.float64)
        self.is_h = np.array([_vmd_is_hydrogen(nm) for nm in names], dtype=bool)
        self.index = np.arange(n)
        # residues: consecutive atoms sharing (chain, segname, resid, icode)
        key = list(zip(chains, segs, resids, icodes))
        res_of = np.zeros(n, dtype=np.int64)
        r = 0
        for i in range(1, n):
            if key[i] != key[i - 1]:
                r += 1
            res_of[i] = r
        self.residue = res_of
        self.n_res = r + 1 if n else 0
        # bonds (Appendix A.3)
        ser2idx = {}
        for i, s in enumerate(serials):
            ser2idx.setdefault(s, i)
        self.bonds = _guess_bonds(self)
        for f in conect:
            if f[0] not in ser2idx:
                continue
            a = ser2idx[f[0]]
            for s in f[1:]:
                if s in ser2idx and ser2idx[s] != a:
                    b = ser2idx[s]
                    self.bonds.add((min(a, b), max(a, b)))
        nb = defaultdict(list)
        for a, b in self.bonds:
            nb[a].append(b)
            nb[b].append(a)
        self.neigh = nb
        self._protein = _vmd_residue_class(self, {"N", "CA", "C", "O"}, {"OT1", "OT2", "OXT", "O1", "O2"})
        self._nucleic = _vmd_residue_class(
            self, {"P", "O1P", "O2P", "OP1", "OP2", "C3*", "C3'", "O3*", "O3'", "C4*", "C4'", "C5*", "C5'", "O5*", "O5'"}, set())
        self._nucleic &= ~self._protein

    def label(self, i: int, with_index: bool = False) -> str:
        """atom.py:40-48 (the consumer strips the trailing index, compute_contacts.py:338-341)."""
        s = "%s:%s:%d%s:%s" % (self.chain[i], self.resname[i], self.resid[i], self.icode[i], self.name[i])
        return s + ":%d" % i if with_index else s

    def reslabel(self, i: int) -> str:
        return "%s:%s:%d%s" % (self.chain[i], self.resname[i], self.resid[i], self.icode[i])

    def bonded(self, a: int, b: int) -> bool:
        return (min(a, b), max(a, b)) in self.bonds


def _vmd_is_hydrogen(name: str) -> bool:
    return bool(name) and (name[0] == 'H' or (len(name) > 1 and name[0].isdigit() and name[1] == 'H'))


def _vmd_default_radius(name: str) -> float:
    s = name.lstrip("0123456789")
    c = s[0].upper() if s else ""
    return {"H": 1.0, "C": 1.5, "N": 1.4, "O": 1.3, "F": 1.2, "S": 1.9}.get(c, 1.5)


def _guess_bonds(top: Topology):
    """Appendix A.3: i-j bonded iff 0.001 <= d2 and d < 0.6*(ri+rj), never H-H; float32."""
    n = top.n
    rad = np.array([_vmd_default_radius(nm) for nm in top.name], dtype=F32)
    x = top.xyz0
    bonds = set()
    if n < 2:
        return bonds
    rmax = float(rad.max())
    pairs = _candidate_pairs(x, x, 1.2 * rmax + 1e-3, same=True)
    if len(pairs) == 0:
        return bonds
    i, j = pairs[:, 0], pairs[:, 1]
    d2 = _d2(x[i], x[j])
    cut = (F32(0.6) * (rad[i] + rad[j])).astype(F32)
    ok = (d2 >= F32(0.001)) & (d2 < cut * cut) & ~(top.is_h[i] & top.is_h[j])
    alt_i, alt_j = top.altloc[i], top.altloc[j]
    ok &= ~((alt_i != alt_j) & (alt_i != "") & (alt_j != ""))
    for a, b in zip(i[ok], j[ok]):
        bonds.add((int(min(a, b)), int(max(a, b))))
    return bonds


def _vmd_residue_class(top: Topology, bbnames, terminal):
    """Appendix A.2: a residue is protein (nucleic) when it holds >= 4 backbone-named atoms."""
    cnt = np.zeros(top.n_res, dtype=np.int64)
    isbb = np.array([nm in bbnames for nm in top.name])
    np.add.at(cnt, top.residue[isbb], 1)
    if terminal:
        ist = np.array([nm in terminal for nm in top.name])
        np.add.at(cnt, top.residue[ist], 1)
    # all four distinct names must be present for protein
    if bbnames == {"N", "CA", "C", "O"}:
        has = np.zeros((top.n_res, 4), dtype=bool)
        for k, nm in enumerate(("N", "CA", "C")):
            has[top.residue[top.name == nm], k] = True
        for nm in ("O",) + tuple(terminal):
            has[top.residue[top.name == nm], 3] = True
        return has.all(axis=1)[top.residue]
    return (cnt >= 4)[top.residue]


# ----------------------------------------------------------------------------------------------
# float32 primitives (Appendix A.1, A.4)
# ----------------------------------------------------------------------------------------------
def _d2(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    d = (a.astype(F32) - b.astype(F32)).astype(F32)
    return ((d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]).astype(F32) + d[..., 2] * d[..., 2]).astype(F32)


def _candidate_pairs(xa: np.ndarray, xb: np.ndarray, r: float, same: bool) -> np.ndarray:
    """Index pairs (into xa, xb) with distance <= r (+ slack); exact test is done by the caller."""
    if len(xa) == 0 or len(xb) == 0:
        return np.zeros((0, 2), dtype=np.int64)
    rr = float(r) * (1 + 1e-5) + 1e-4
    if cKDTree is not None and (len(xa) * len(xb) > 4_000_000):
        ta = cKDTree(xa.astype(np.float64))
        if same:
            p = ta.query_pairs(rr, output_type='ndarray')
            return p.astype(np.int64)
        tb = cKDTree(xb.astype(np.float64))
        sp = ta.sparse_distance_matrix(tb, rr, output_type='coo_matrix')
        return np.stack([sp.row, sp.col], axis=1).astype(np.int64)
    out = []
    step = max(1, 2_000_000 // max(1, len(xb)))
    xb64 = xb.astype(np.float64)
    for s in range(0, len(xa), step):
        d = xa[s:s + step, None, :].astype(np.float64) - xb64[None, :, :]
        m = (d * d).sum(-1) <= rr * rr
        ii, jj = np.nonzero(m)
        ii = ii + s
        if same:
            k = ii < jj
            ii, jj = ii[k], jj[k]
        out.append(np.stack([ii, jj], axis=1))
    return np.concatenate(out, axis=0).astype(np.int64) if out else np.zeros((0, 2), dtype=np.int64)


# ----------------------------------------------------------------------------------------------
# VMD atom-selection language (the subset getcontacts and its tests use; Appendix A.2)
# ----------------------------------------------------------------------------------------------
_ION_RESN = set("AL BA CA Ca CAL CD CES CLA CL Cl CO CS CU Cu CU1 CUA HG IN IOD K MG MN3 MO3 MO4 MO5 MO6 NA Na NAW "
                "OC7 PB POT PT RB SOD TB TL WO4 YB ZN ZN1 ZN2".split())
_WATER_RESN = set("H2O HH0 OHH HOH OH2 SOL WAT TIP TIP2 TIP3 TIP4".split())
_TOKEN = re.compile(r"\s*(\(|\)|'[^']*'|\"[^\"]*\"|[^\s()]+)")


class Selector:
    """Recursive-descent evaluator: or / and / not / ( ) / keyword value-lists / within R of."""
    KEYS = {"chain", "resname", "name", "resid", "index", "element", "segname", "residue", "type", "altloc",
            "insertion", "serial"}

    def __init__(self, top: Topology, macros=None, xyz=None):
        self.top = top
        self.macros = dict(macros or {})
        self.xyz = top.xyz0 if xyz is None else xyz

    def __call__(self, text: str) -> np.ndarray:
        self.tok = _TOKEN.findall(text)
        self.pos = 0
        m = self._or()
        if self.pos != len(self.tok):
            raise ValueError("selection parse error near %r in %r" % (self.tok[self.pos], text))
        return m

    def _peek(self):
        return self.tok[self.pos] if self.pos < len(self.tok) else None

    def _or(self):
        m = self._and()
        while self._peek() == "or":
            self.pos += 1
            m = m | self._and()
        return m

    def _and(self):
        m = self._not()
        while self._peek() == "and":
            self.pos += 1
            m = m & self._not()
        return m

    def _not(self):
        if self._peek() == "not":
            self.pos += 1
            return ~self._not()
        return self._atom()

    def _values(self):
        vals = []
        while self._peek() is not None and self._peek() not in ("and", "or", "not", ")", "("):
            v = self.tok[self.pos]
            if v[0] in "'\"":
                v = v[1:-1]
            vals.append(v)
            self.pos += 1
        return vals

    def _atom(self):
        t = self._peek()
        top = self.top
        if t is None:
            raise ValueError("unexpected end of selection")
        if t == "(":
            self.pos += 1
            m = self._or()
            if self._peek() != ")":
                raise ValueError("missing )")
            self.pos += 1
            return m
        self.pos += 1
        if t in self.macros:
            mac = self.macros[t]
            if isinstance(mac, np.ndarray):
                return mac.copy()
            return Selector(top, self.macros, self.xyz)(mac)
        if t == "all":
            return np.ones(top.n, dtype=bool)
        if t == "none":
            return np.zeros(top.n, dtype=bool)
        if t == "protein":
            return top._protein.copy()
        if t == "nucleic":
            return top._nucleic.copy()
        if t in ("water", "waters"):
            return np.array([r in _WATER_RESN for r in top.resname])
        if t == "hydrogen":
            return top.is_h.copy()
        if t == "noh":
            return ~top.is_h
        if t == "ion":
            return np.array([r in _ION_RESN for r in top.resname])
        if t == "carbon":
            return np.array([nm.startswith("C") for nm in top.name]) & ~np.array([r in _ION_RESN for r in top.resname])
        if t == "sulfur":
            return np.array([nm.startswith("S") for nm in top.name]) & ~np.array([r in _ION_RESN for r in top.resname])
        if t == "backbone":
            return top._protein & np.array([nm in ("N", "CA", "C", "O") for nm in top.name])
        if t == "within":
            r = float(self.tok[self.pos])
            self.pos += 1
            if self._peek() != "of":
                raise ValueError("expected 'of'")
            self.pos += 1
            inner = self._not()
            return within(self.xyz, np.ones(top.n, dtype=bool), inner, r)
        if t in self.KEYS:
            vals = self._values()
            return self._keyword(t, vals)
        raise ValueError("unsupported selection keyword %r" % t)

    def _keyword(self, key, vals):
        top = self.top
        if key in ("resid", "index", "residue", "serial"):
            arr = {"resid": top.resid, "index": top.index, "residue": top.residue, "serial": top.index + 1}[key]
            m = np.zeros(top.n, dtype=bool)
            k = 0
            while k < len(vals):
                if k + 2 < len(vals) and vals[k + 1] == "to":
                    m |= (arr >= int(vals[k])) & (arr <= int(vals[k + 2]))
                    k += 3
                else:
                    m |= arr == int(vals[k])
                    k += 1
            return m
        arr = {"chain": top.chain, "resname": top.resname, "name": top.name, "element": top.element,
               "segname": top.segname, "type": top.name, "altloc": top.altloc, "insertion": top.icode}[key]
        vs = set(vals)
        return np.array([a in vs for a in arr], dtype=bool)


def within(xyz: np.ndarray, cand: np.ndarray, ref: np.ndarray, r: float) -> np.ndarray:
    """Appendix A.6: candidates with float32 d2 <= r*r to any atom of ref (ref atoms included)."""
    ci, ri = np.nonzero(cand)[0], np.nonzero(ref)[0]
    out = np.zeros(len(xyz), dtype=bool)
    if len(ci) == 0 or len(ri) == 0:
        return out
    p = _candidate_pairs(xyz[ci], xyz[ri], r, same=False)
    if len(p) == 0:
        return out
    d2 = _d2(xyz[ci[p[:, 0]]], xyz[ri[p[:, 1]]])
    ok = d2 <= F32(r) * F32(r)
    out[ci[p[ok, 0]]] = True
    return out


# ----------------------------------------------------------------------------------------------
# One-off setup (compute_contacts.py:226-238)
# ----------------------------------------------------------------------------------------------
DEFAULT_GEOM = {  # argparsers.py:80-125 (dynamic defaults)
    "SALT_BRIDGE_CUTOFF_DISTANCE": 4.0, "PI_CATION_CUTOFF_DISTANCE": 6.0, "PI_CATION_CUTOFF_ANGLE": 60,
    "PI_STACK_CUTOFF_DISTANCE": 7.0, "PI_STACK_CUTOFF_ANGLE": 30, "PI_STACK_PSI_ANGLE": 45,
    "T_STACK_CUTOFF_DISTANCE": 5.0, "T_STACK_CUTOFF_ANGLE": 30, "T_STACK_PSI_ANGLE": 45,
    "HBOND_CUTOFF_DISTANCE": 3.5, "HBOND_CUTOFF_ANGLE": 70, "HBOND_RES_DIFF": 1,
    "VDW_EPSILON": 0.5, "VDW_RES_DIFF": 2}
ALL_ITYPES = ["hp", "sb", "pc", "ps", "ts", "vdw", "hb"]  # get_dynamic_contacts.py:68

_HIS = {"HIS", "HSD", "HSE", "HSP", "HIE", "HIP", "HID"}
_SOLV_AUTO = set("H2O HH0 OHH HOH OH2 SOL WAT TIP TIP2 TIP3 TIP4 T3P IP3".split())  # contact_utils.py:448
_LIPID_AUTO = set("DLPE DMPC DPPC GPC LPPC PALM PC PGCL POPC POPE".split())          # contact_utils.py:503
_HP_RESN = set("ALA PHE GLY ILE LEU PRO VAL TRP".split())                             # hydrophobics.py:9
_RING_NAMES = {"PHE": ("CG", "CE1", "CE2"), "TYR": ("CG", "CE1", "CE2"), "TRP": ("CD2", "CZ2", "CZ3")}
for _h in _HIS:
    _RING_NAMES[_h] = ("CG", "CE1", "CD2")


class Setup:
    """Everything compute_contacts derives once per call from topology + frame 0."""

    def __init__(self, top: Topology, sele1="protein", sele2=None, solv_sele="", lipid_sele="", ligand_sele="",
                 xyz0=None):
        self.top = top
        sele2 = sele1 if sele2 is None else sele2
        self.sele1_str, self.sele2_str = sele1, sele2
        self.same = sele1 == sele2  # string equality, e.g. hbonds.py:61
        xyz0 = top.xyz0 if xyz0 is None else xyz0
        base = Selector(top, {}, xyz0)
        # configure_solv / configure_lipid / configure_ligand (contact_utils.py:411-579)
        if solv_sele:
            solv = base(solv_sele)
        else:
            solv = np.array([r in _SOLV_AUTO for r in top.resname], dtype=bool)
        if lipid_sele:
            lipid = Selector(top, {"solv": solv}, xyz0)(lipid_sele)
        else:
            lipid = np.array([r in _LIPID_AUTO for r in top.resname], dtype=bool)
        macros = {"solv": solv, "lipid": lipid}
        if ligand_sele:
            ligand = Selector(top, macros, xyz0)(ligand_sele)
        else:
            ligand = ~(lipid | solv | top._protein | top._nucleic)
        macros["ligand"] = ligand
        self.solv, self.lipid, self.ligand = solv, lipid, ligand
        self.macros = macros
        sel = Selector(top, macros, xyz0)
        self.sele1 = sel(sele1)
        self.sele2 = sel(sele2)
        self.disulfide = self._find_disulfide(xyz0)
        self.ligand_anions, self.ligand_cations = self._ligand_features(xyz0)
        n = top.n
        rn, nm = top.resname, top.name
        # salt_bridges.py:25-30
        anion = np.zeros(n, dtype=bool)
        cation = np.zeros(n, dtype=bool)
        ring = np.zeros(n, dtype=bool)
        nucl_resn = set("C U G A DC DT DG DA".split())
        for i in range(n):
            r, a = rn[i], nm[i]
            if (r == "ASP" and a in ("OD1", "OD2")) or (r == "GLU" and a in ("OE1", "OE2")) or \
                    (r in nucl_resn and a in ("OP1", "OP2", "O1P", "O2P")):
                anion[i] = True
            if (r in _HIS and a in ("ND1", "NE2")) or (r == "LYS" and a == "NZ") or (r == "ARG" and a in ("NH1", "NH2")):
                cation[i] = True
            if r in _RING_NAMES and a in _RING_NAMES[r]:
                ring[i] = True
            elif top._nucleic[i] and a in ("C4", "C2", "C6"):
                ring[i] = True
        self.prot_cation = cation.copy()           # pi_cation.py:24-26 has no ligand cations
        self.anion = anion.copy()
        self.cation = cation.copy()
        for i in self.ligand_anions:
            self.anion[i] = True
        for i in self.ligand_cations:
            self.cation[i] = True
        self.ring = ring
        # hbond static classes (hbonds.py:62/67): not (carbon or sulfur or solv)
        self.carbon = sel("carbon")
        self.sulfur = sel("sulfur")

    # contact_utils.py:396-409
    def _find_disulfide(self, xyz0):
        top = self.top
        pairs = []
        cys = top.resname == "CYS"
        sg = np.array([a == "SG" for a in top.name], dtype=bool)
        for rid in sorted(set(top.resid[cys].tolist())):
            ref = sg & (top.resid == rid)
            got = within(xyz0, sg, ref, 2.1)
            s = set(top.resid[got].tolist())
            if len(s) == 2 and s not in pairs:
                pairs.append(s)
        return pairs

    # contact_utils.py:605-665
    def _ligand_features(self, xyz0):
        top = self.top
        lig = np.nonzero(self.ligand)[0]
        anions, cations = [], []
        if len(lig) == 0:
            return anions, cations
        metal = ["MG", "MN", "RH", "ZN", "FE", "BI", "AS", "AG"]
        allm = np.ones(top.n, dtype=bool)
        for a in lig:
            one = np.zeros(top.n, dtype=bool)
            one[a] = True
            nb = [int(k) for k in np.nonzero(within(xyz0, allm, one, 1.95))[0] if k != a]
            lab = top.label(int(a), with_index=True)
            if any(m in lab for m in metal):
                cations.append(int(a))
                continue
            if top.element[a] == 'C' and len(nb) == 3:
                ang = _vmd_angle(xyz0[nb[0]], xyz0[a], xyz0[nb[1]])
                if not ((120. - 5) < ang < (120. + 5)):
                    continue
                els = [top.element[k] for k in nb]
                if not (els.count('C') == 1 and els.count('O') == 2):
                    continue
                anions += [k for k in nb if top.element[k] == 'O']
        return anions, cations


def _vmd_angle(p1, p2, p3) -> float:
    """``measure angle`` (float32 acos form); only used for ligand carboxylate detection."""
    a = (p1 - p2).astype(F32)
    b = (p3 - p2).astype(F32)
    cosang = float(np.dot(a, b)) / (float(np.sqrt(np.dot(a, a))) * float(np.sqrt(np.dot(b, b))))
    return math.degrees(math.acos(max(-1.0, min(1.0, cosang))))


# ----------------------------------------------------------------------------------------------
# VMD measure primitives on one frame
# ----------------------------------------------------------------------------------------------
def measure_contacts(top: Topology, xyz, m1, m2, cutoff: float, one_selection=False):
    """Appendix A.4.  Returns an (P,2) int array of (i in s1, j in s2).  Bonded pairs and i==j are
    excluded.  ``one_selection`` -> each unordered pair once (i<j)."""
    i1 = np.nonzero(m1)[0]
    i2 = i1 if one_selection else np.nonzero(m2)[0]
    p = _candidate_pairs(xyz[i1], xyz[i2], cutoff, same=one_selection)
    if len(p) == 0:
        return np.zeros((0, 2), dtype=np.int64)
    a, b = i1[p[:, 0]], i2[p[:, 1]]
    c = F32(cutoff)
    ok = (_d2(xyz[a], xyz[b]) <= c * c) & (a != b)
    a, b = a[ok], b[ok]
    keep = np.array([not top.bonded(int(u), int(v)) for u, v in zip(a, b)], dtype=bool) if len(a) else np.zeros(0, bool)
    return np.stack([a[keep], b[keep]], axis=1)


NEAR_DIST_TOL = 1e-4     # Angstrom (north star: pairs within 1e-4 A / 1e-3 degree of a cutoff are listed separately)
NEAR_ANGLE_TOL = 1e-3


def contacts_band(top: Topology, xyz, m1, m2, cutoff: float, one_selection=False):
    """Oracle-only companion of ``measure_contacts`` for the near-cutoff side channel (not reference behaviour):
    every non-bonded pair (i in s1, j in s2, i != j) whose float32 ``sqrtf(d2)`` is below ``cutoff + 1e-4``, as
    (i, j, delta) with delta = (double)sqrtf(d2) - cutoff.  |delta| < 1e-4 marks the pairs whose membership in
    ``measure contacts cutoff`` rests on the weakly pinned <= / < at the float32 boundary."""
    i1 = np.nonzero(m1)[0]
    i2 = i1 if one_selection else np.nonzero(m2)[0]
    p = _candidate_pairs(xyz[i1], xyz[i2], cutoff + 4 * NEAR_DIST_TOL, same=one_selection)
    out = []
    if len(p) == 0:
        return out
    a, b = i1[p[:, 0]], i2[p[:, 1]]
    d = np.sqrt(_d2(xyz[a], xyz[b])).astype(F32).astype(np.float64) - float(cutoff)
    for u, v, dl in zip(a.tolist(), b.tolist(), d.tolist()):
        if u != v and dl < NEAR_DIST_TOL and not top.bonded(u, v):
            out.append((u, v, dl))
    return out


def hbond_angle_f32(xd, xh, xa) -> np.float32:
    """Appendix A.5: 57.2958f * (float)atan2(|u x v|, u.v), u = h-d, v = a-h, float32 products."""
    u = (xh - xd).astype(F32)
    v = (xa - xh).astype(F32)
    cx = F32(F32(u[1] * v[2]) - F32(u[2] * v[1]))
    cy = F32(F32(u[2] * v[0]) - F32(u[0] * v[2]))
    cz = F32(F32(u[0] * v[1]) - F32(u[1] * v[0]))
    psin = F32(np.sqrt(F32(F32(F32(cx * cx) + F32(cy * cy)) + F32(cz * cz))))
    pcos = F32(F32(F32(u[0] * v[0]) + F32(u[1] * v[1])) + F32(u[2] * v[2]))
    return F32(F32(57.2958) * F32(math.atan2(float(psin), float(pcos))))


def measure_hbonds(top: Topology, xyz, mD, mA, cutoff: float, ang: float, one_selection=False, near=None,
                   near_filter=None):
    """Appendix A.5.  Returns a set of (donor, acceptor) heavy-atom index pairs.

    ``near`` (oracle-only side channel, not reference behaviour): every (donor, H, acceptor) angle within
    1e-3 degree of the cutoff is recorded, for every hydrogen of the donor, when ``near_filter(d, a)``
    holds (the caller passes the non-geometric filters so that only decisive pairs are listed)."""
    iD = np.nonzero(mD & ~top.is_h)[0]
    iA = iD if one_selection else np.nonzero(mA & ~top.is_h)[0]
    out = set()
    p = _candidate_pairs(xyz[iD], xyz[iA], cutoff, same=one_selection)
    if len(p) == 0:
        return out
    a, b = iD[p[:, 0]], iA[p[:, 1]]
    c = F32(cutoff)
    ok = (_d2(xyz[a], xyz[b]) <= c * c) & (a != b)
    angf = F32(ang)
    for u, v in zip(a[ok].tolist(), b[ok].tolist()):
        if top.bonded(u, v):
            continue
        for d, acc in (((u, v), (v, u)) if one_selection else ((u, v),)):
            for h in top.neigh.get(d, ()):
                if not top.is_h[h]:
                    continue
                t = hbond_angle_f32(xyz[d], xyz[h], xyz[acc])
                if near is not None and abs(float(t) - float(angf)) < 1e-3 and (near_filter is None or near_filter(d, acc)):
                    near.append(("hb_angle", d, acc, float(t) - float(angf)))
                if t < angf:
                    out.add((d, acc))
    return out


# ----------------------------------------------------------------------------------------------
# float64 ring helpers (contact_utils.py:958-1061)
# ----------------------------------------------------------------------------------------------
def _angle_between(v1, v2) -> float:
    c = float(np.dot(v1, v2)) / (math.sqrt(float(np.dot(v1, v1))) * math.sqrt(float(np.dot(v2, v2))))
    return math.degrees(math.acos(c))


def _centroid(p1, p2, p3):
    return (p1 + p2 + p3) / 3


def _normal(p1, p2, p3):
    return np.cross(p3 - p1, p2 - p1)


def _fold(t: float) -> float:
    return min(math.fabs(t - 0), math.fabs(t - 180))


def _psi(c1, c2, n) -> float:
    return _fold(_angle_between(n, c1 - c2))


# ----------------------------------------------------------------------------------------------
# Per-frame detectors.  Each returns a list of (itype, a, b[, w1[, w2]]) with atom indices.
# ----------------------------------------------------------------------------------------------
class FragmentState(dict):
    """The per-fragment cache the reference keeps inside geom_criteria (Appendix A.8)."""


def compute_salt_bridges(S: Setup, xyz, geom, near=None):
    top = S.top
    cut = geom['SALT_BRIDGE_CUTOFF_DISTANCE']
    if near is not None:     # salt_bridges.py:76-80: the 4.0 A `measure contacts` boundary
        band = contacts_band(top, xyz, S.anion & S.sele1, S.cation & S.sele2, cut)
        if not S.same:
            band += contacts_band(top, xyz, S.anion & S.sele2, S.cation & S.sele1, cut)
        near += [("sb_dist", a, c, dl) for a, c, dl in band if abs(dl) < NEAR_DIST_TOL]
    p12 = measure_contacts(top, xyz, S.anion & S.sele1, S.cation & S.sele2, cut)
    pairs = set(map(tuple, p12.tolist()))
    if not S.same:
        p21 = measure_contacts(top, xyz, S.anion & S.sele2, S.cation & S.sele1, cut)
        pairs |= set(map(tuple, p21.tolist()))
    out = []
    for a, c in sorted(pairs):
        al, cl = bool(S.ligand[a]), bool(S.ligand[c])
        if al and not cl:
            out.append(("sblp", a, c))
        if (not al) and cl:
            out.append(("sbpl", a, c))
        if al and cl:
            out.append(("sbll", a, c))
        else:
            out.append(("sb", a, c))  # salt_bridges.py:94-97 (if/if/if-else)
    return out


def _ring_rep(top, trip):
    """contact_utils.py:335-347: smallest atom NAME of the triplet."""
    return min(trip, key=lambda i: top.name[i])


def compute_pi_cation(S: Setup, xyz, geom, near=None):
    top = S.top
    cd, ca = geom['PI_CATION_CUTOFF_DISTANCE'], geom['PI_CATION_CUTOFF_ANGLE']
    pairs = set(map(tuple, measure_contacts(top, xyz, S.prot_cation & S.sele1, S.ring & S.sele2, 10.0).tolist()))
    if not S.same:
        pairs |= set(map(tuple, measure_contacts(top, xyz, S.prot_cation & S.sele2, S.ring & S.sele1, 10.0).tolist()))
    groups = defaultdict(set)
    for c, r in pairs:
        groups[(c, top.reslabel(r))].add(r)
    out = []
    x64 = xyz.astype(np.float64)
    if near is not None:
        # pi_cation.py:22,70: the soft 10 A `measure contacts` decides whether a ring's three atoms all reach the group.
        # Listed only where it can change the outcome: all three atoms inside 10 A + 1e-4 and the centroid within the
        # hard distance cutoff (+ 1e-4).
        band = contacts_band(top, xyz, S.prot_cation & S.sele1, S.ring & S.sele2, 10.0)
        if not S.same:
            band += contacts_band(top, xyz, S.prot_cation & S.sele2, S.ring & S.sele1, 10.0)
        wide = defaultdict(dict)
        for c, r, dl in band:
            wide[(c, top.reslabel(r))][r] = dl
        for (c, _), ring in wide.items():
            if len(ring) != 3 or not any(abs(dl) < NEAR_DIST_TOL for dl in ring.values()):
                continue
            p = sorted(ring)
            cen = _centroid(x64[p[0]], x64[p[1]], x64[p[2]])
            if float(np.linalg.norm(x64[c] - cen)) - cd < NEAR_DIST_TOL:
                near += [("pc_soft", c, r, dl) for r, dl in ring.items() if abs(dl) < NEAR_DIST_TOL]
    for (c, _), ring in sorted(groups.items(), key=lambda kv: (kv[0][0], min(kv[1]))):
        if len(ring) != 3:
            continue
        p = sorted(ring, key=lambda i: top.label(i, True))  # pi_cation.py:101
        p1, p2, p3 = x64[p[0]], x64[p[1]], x64[p[2]]
        cen = _centroid(p1, p2, p3)
        d = float(np.linalg.norm(x64[c] - cen))
        if near is not None and abs(d - cd) < 1e-4:
            near.append(("pc_dist", c, _ring_rep(top, p), d - cd))
        if d > cd:
            continue
        t = _fold(_angle_between(_normal(p1, p2, p3), x64[c] - cen))
        if near is not None and abs(t - ca) < 1e-3:
            near.append(("pc_angle", c, _ring_rep(top, p), t - ca))
        if t > ca:
            continue
        out.append(("pc", c, _ring_rep(top, p)))
    return out


def _aromatic_triplet(S: Setup, i: int):
    """aromatics.py:71-113: same chain, resname, resid (icode NOT part of the query), ring names."""
    top = S.top
    rn = top.resname[i]
    names = set(_RING_NAMES[rn]) if rn in _RING_NAMES else {"C2", "C4", "C6"}
    m = (top.chain == top.chain[i]) & (top.resname == rn) & (top.resid == top.resid[i])
    return [int(k) for k in np.nonzero(m)[0] if top.name[k] in names]


def compute_aromatics(S: Setup, xyz, itype, soft, dist_cut, ang_cut, psi_cut, near=None):
    top = S.top
    sel = S.ring & (S.sele1 | S.sele2)
    pairs = measure_contacts(top, xyz, sel, sel, soft, one_selection=True)
    res_pairs = []
    seen = set()
    trip = {}
    for a, b in pairs.tolist():
        if not ((S.sele1[a] and S.sele2[b]) or (S.sele2[a] and S.sele1[b])):
            continue
        ra, rb = top.reslabel(a), top.reslabel(b)
        if ra == rb:
            continue
        if ra not in trip:
            trip[ra] = _aromatic_triplet(S, a)
        if rb not in trip:
            trip[rb] = _aromatic_triplet(S, b)
        if (ra, rb) not in seen and (rb, ra) not in seen:
            seen.add((ra, rb))
            res_pairs.append((ra, rb))
    out = []
    x64 = xyz.astype(np.float64)
    if near is not None:
        # aromatics.py:26-27,166: the soft `measure contacts` (ps 10 A, ts 6 A) finds a residue pair through its CLOSEST
        # selected ring-atom pair; listed when that closest distance is within 1e-4 A of the soft cutoff and the centroids
        # are within the hard distance cutoff (+ 1e-4).
        best = {}
        for a, b, dl in contacts_band(top, xyz, sel, sel, soft, one_selection=True):
            if not ((S.sele1[a] and S.sele2[b]) or (S.sele2[a] and S.sele1[b])):
                continue
            ra, rb = top.reslabel(a), top.reslabel(b)
            if ra == rb:
                continue
            key = (min(ra, rb), max(ra, rb))
            if key not in best or dl < best[key][0]:
                best[key] = (dl, a, b)
        for (ra, rb), (dl, a, b) in best.items():
            if abs(dl) >= NEAR_DIST_TOL:
                continue
            t1, t2 = _aromatic_triplet(S, a), _aromatic_triplet(S, b)
            if len(t1) != 3 or len(t2) != 3:
                continue
            c1, c2 = _centroid(*(x64[k] for k in t1)), _centroid(*(x64[k] for k in t2))
            if float(np.linalg.norm(c1 - c2)) - dist_cut < NEAR_DIST_TOL:
                near.append((itype + "_soft", _ring_rep(top, t1), _ring_rep(top, t2), dl))
    for ra, rb in res_pairs:
        t1, t2 = trip[ra], trip[rb]
        if len(t1) != 3 or len(t2) != 3:
            continue
        a1, a2, a3 = (x64[k] for k in t1)
        b1, b2, b3 = (x64[k] for k in t2)
        c1, c2 = _centroid(a1, a2, a3), _centroid(b1, b2, b3)
        d = float(np.linalg.norm(c1 - c2))
        if near is not None and abs(d - dist_cut) < 1e-4:
            near.append((itype + "_dist", _ring_rep(top, t1), _ring_rep(top, t2), d - dist_cut))
        if d > dist_cut:
            continue
        n1, n2 = _normal(a1, a2, a3), _normal(b1, b2, b3)
        ang = _angle_between(n1, n2)
        v = _fold(ang) if itype == "ps" else math.fabs(ang - 90)
        if near is not None and abs(v - ang_cut) < 1e-3:
            near.append((itype + "_angle", _ring_rep(top, t1), _ring_rep(top, t2), v - ang_cut))
        if v > ang_cut:
            continue
        psi = min(_psi(c1, c2, n1), _psi(c2, c1, n2))
        if near is not None and abs(psi - psi_cut) < 1e-3:
            near.append((itype + "_psi", _ring_rep(top, t1), _ring_rep(top, t2), psi - psi_cut))
        if psi > psi_cut:
            continue
        out.append((itype, _ring_rep(top, t1), _ring_rep(top, t2)))
    return out


def _local(top, a, b, k) -> bool:
    return top.chain[a] == top.chain[b] and abs(int(top.resid[a]) - int(top.resid[b])) < k


def _is_disulfide(S: Setup, a, b) -> bool:
    top = S.top
    if top.resname[a] == "CYS" and top.resname[b] == "CYS":
        return {int(top.resid[a]), int(top.resid[b])} in S.disulfide
    return False


def _vdw_accept(top, xyz, a, b, eps) -> tuple:
    d = float(F32(np.sqrt(_d2(xyz[a], xyz[b]))))  # measure bond: sqrtf in float32, parsed as double
    cut = top.vdw[a] + top.vdw[b] + eps            # double sum, vanderwaals.py:92
    return d < cut, d - cut


def compute_vanderwaals(S: Setup, xyz, geom, frag: FragmentState, near=None):
    top = S.top
    eps, rd = geom['VDW_EPSILON'], geom['VDW_RES_DIFF']
    if 'soft_vdw_cutoff' not in frag:  # vanderwaals.py:24-31 (max over ALL atoms of each selection)
        frag['soft_vdw_cutoff'] = float(top.vdw[S.sele1].max()) + float(top.vdw[S.sele2].max()) + eps
    soft = frag['soft_vdw_cutoff']
    pairs = measure_contacts(top, xyz, S.sele1 & ~top.is_h, S.sele2 & ~top.is_h, soft)
    out = []
    for a, b in pairs.tolist():
        if _local(top, a, b, rd):
            continue
        if _is_disulfide(S, a, b):
            continue
        ok, delta = _vdw_accept(top, xyz, a, b, eps)
        if near is not None and abs(delta) < 1e-4:
            near.append(("vdw_dist", a, b, delta))
        if ok:
            out.append(("vdw", a, b))
    return out


def _prepare_hp(S: Setup, xyz0):
    """hydrophobics.py:8-41 on the fragment's first frame."""
    top = S.top
    base = (np.array([r in _HP_RESN for r in top.resname]) | S.ligand) & (top.element == 'C')
    allm = np.ones(top.n, dtype=bool)
    h1, h2 = [], []
    ch = {"C", "H"}
    cache = {}
    for which, sele, dst in ((1, S.sele1, h1), (2, S.sele2, h2)):
        for c in np.nonzero(base & sele)[0].tolist():
            if c not in cache:
                one = np.zeros(top.n, dtype=bool)
                one[c] = True
                cache[c] = set(top.element[within(xyz0, allm, one, 1.7)].tolist())
            els = cache[c]
            if which == 1:
                if els | ch == ch:
                    dst.append(c)
            elif els == ch:
                dst.append(c)
    return h1, h2


def compute_hydrophobics(S: Setup, xyz, geom, frag: FragmentState, near=None):
    top = S.top
    eps, rd = geom['VDW_EPSILON'], geom['VDW_RES_DIFF']
    if 'hp1' not in frag:
        frag['hp1'], frag['hp2'] = _prepare_hp(S, frag['xyz0'])
    h1, h2 = frag['hp1'], frag['hp2']
    if not h1 or not h2:
        return []
    m1 = np.zeros(top.n, dtype=bool)
    m1[h1] = True
    if S.same:
        pairs = measure_contacts(top, xyz, m1, m1, eps + 2 * 1.7, one_selection=True)
    else:
        m2 = np.zeros(top.n, dtype=bool)
        m2[h2] = True
        pairs = measure_contacts(top, xyz, m1, m2, eps + 2 * 1.7)
    out = []
    for a, b in pairs.tolist():
        if _local(top, a, b, rd):
            continue
        if _is_disulfide(S, a, b):
            continue
        ok, delta = _vdw_accept(top, xyz, a, b, eps)
        if near is not None and abs(delta) < 1e-4:
            near.append(("hp_dist", a, b, delta))
        if not ok:
            continue
        la, lb = bool(S.ligand[a]), bool(S.ligand[b])
        out.append(("hp" if not (la or lb) else ("hplp" if la != lb else "hpll"), a, b))
    return out


def compute_hydrogen_bonds(S: Setup, xyz, geom, near=None):
    top = S.top
    cd, ca, rd = geom['HBOND_CUTOFF_DISTANCE'], geom['HBOND_CUTOFF_ANGLE'], geom['HBOND_RES_DIFF']
    selany = S.sele1 | S.sele2
    U = selany & ~(S.carbon | S.sulfur | S.solv)
    W = within(xyz, S.solv, selany, 3.5) if S.solv.any() else np.zeros(top.n, dtype=bool)
    dual = lambda a, b: (S.sele1[a] and S.sele2[b]) or (S.sele2[a] and S.sele1[b])
    SS = measure_hbonds(top, xyz, U, U, cd, ca, one_selection=True, near=near,
                        near_filter=lambda d, a: dual(d, a) and not _local(top, d, a, rd))
    has_h = getattr(top, "_has_h", None)
    if has_h is None:
        has_h = np.zeros(top.n, dtype=bool)
        for i, nb in top.neigh.items():
            if any(top.is_h[h] for h in nb):
                has_h[i] = True
        top._has_h = has_h
    heavy = ~top.is_h
    if near is not None:     # hbonds.py:75-82: the donor-acceptor distance of `measure hbonds` (heavy pairs with a hydrogen)
        near += [("hb_dist", u, v, dl) for u, v, dl in contacts_band(top, xyz, U & heavy, U & heavy, cd, one_selection=True)
                 if abs(dl) < NEAR_DIST_TOL and (has_h[u] or has_h[v]) and dual(u, v) and not _local(top, u, v, rd)]
    out = []
    for d, a in sorted(SS):
        if not dual(d, a):
            continue
        if _local(top, d, a, rd):
            continue
        dbb, abb = top.name[d] in BB_NAMES, top.name[a] in BB_NAMES
        dl, al = bool(S.ligand[d]), bool(S.ligand[a])
        if dl and al:
            t = "hbll"
        elif dl:
            t = "hblb" if abb else "hbls"
        elif al:
            t = "hblb" if dbb else "hbls"
        elif dbb and abb:
            t = "hbbb"
        elif not dbb and not abb:
            t = "hbss"
        else:
            t = "hbsb"
        out.append((t, d, a))
    if near is not None and S.solv.any():
        # water edges (hbonds.py:64-80): a solvent heavy atom matters through its hydrogen bond to a selection atom; that
        # bond's 3.5 A distance test is also what puts the water into the `within 3.5` shell, one float32 comparison.
        same_res = lambda u, v: top.chain[u] == top.chain[v] and top.resid[u] == top.resid[v]
        near += [("hb_dist", u, v, dl) for u, v, dl in contacts_band(top, xyz, U & heavy, S.solv & heavy, cd)
                 if abs(dl) < NEAR_DIST_TOL and (has_h[u] or has_h[v]) and not same_res(u, v)]
    if not W.any():
        return out
    SW = measure_hbonds(top, xyz, U, W, cd, ca) | measure_hbonds(top, xyz, W, U, cd, ca)
    WW = measure_hbonds(top, xyz, W, W, cd, ca, one_selection=True)
    wd = defaultdict(set)
    for d, a in sorted(SW | WW):
        if top.chain[d] == top.chain[a] and top.resid[d] == top.resid[a]:
            continue
        if S.solv[d]:
            wd[d].add(a)
        if S.solv[a]:
            wd[a].add(d)
    if near is not None:
        # water-water bonds decide a two-water bridge only between waters that each hold a non-solvent partner
        marked = np.zeros(top.n, dtype=bool)
        for w, nb in wd.items():
            if any(not S.solv[k] for k in nb):
                marked[w] = True
        near += [("hb_dist", u, v, dl) for u, v, dl in contacts_band(top, xyz, marked & heavy, marked & heavy, cd, one_selection=True)
                 if abs(dl) < NEAR_DIST_TOL and (has_h[u] or has_h[v]) and not same_res(u, v)]
    for w in sorted(wd):
        for n1, n2 in product(sorted(wd[w]), repeat=2):
            s1, s2 = bool(S.solv[n1]), bool(S.solv[n2])
            if s1 and s2:
                continue
            if s1:
                n1, n2 = n2, n1
                s1, s2 = s2, s1
            l1 = bool(S.ligand[n1])
            if s2:
                if n2 <= w:
                    continue
                for m in sorted(wd.get(n2, ())):
                    if S.solv[m]:
                        continue
                    if not dual(n1, m):
                        continue
                    if _local(top, n1, m, rd):
                        continue
                    out.append(("lwb2" if (l1 or S.ligand[m]) else "wb2", n1, m, w, n2))
            else:
                if not dual(n1, n2):
                    continue
                if _local(top, n1, n2, rd):
                    continue
                out.append(("lwb" if (l1 or S.ligand[n2]) else "wb", n1, n2, w))
    return out


def compute_frame_contacts(S: Setup, xyz, itypes, geom, frag: FragmentState, near=None):
    """compute_contacts.py:50-103, fixed order sb, pc, ps, ts, vdw, hb, hp."""
    xyz = np.ascontiguousarray(xyz, dtype=F32)
    if 'xyz0' not in frag:
        frag['xyz0'] = xyz
    out = []
    if "sb" in itypes:
        out += compute_salt_bridges(S, xyz, geom, near)
    if "pc" in itypes:
        out += compute_pi_cation(S, xyz, geom, near)
    if "ps" in itypes:
        out += compute_aromatics(S, xyz, "ps", 10.0, geom['PI_STACK_CUTOFF_DISTANCE'],
                                 geom['PI_STACK_CUTOFF_ANGLE'], geom['PI_STACK_PSI_ANGLE'], near)
    if "ts" in itypes:
        out += compute_aromatics(S, xyz, "ts", 6.0, geom['T_STACK_CUTOFF_DISTANCE'],
                                 geom['T_STACK_CUTOFF_ANGLE'], geom['T_STACK_PSI_ANGLE'], near)
    if "vdw" in itypes:
        out += compute_vanderwaals(S, xyz, geom, frag, near)
    if "hb" in itypes:
        out += compute_hydrogen_bonds(S, xyz, geom, near)
    if "hp" in itypes:
        out += compute_hydrophobics(S, xyz, geom, frag, near)
    return out


# ----------------------------------------------------------------------------------------------
# Trajectory input (Appendix B.2) -- DCD only; a PDB "trajectory" is its own coordinates
# ----------------------------------------------------------------------------------------------
def read_dcd(path: str) -> np.ndarray:
    """Returns float32 [F][N][3] (Angstrom)."""
    with open(path, "rb") as fh:
        raw = fh.read()
    if struct.unpack("<i", raw[:4])[0] == 84:
        e = "<"
    elif struct.unpack(">i", raw[:4])[0] == 84:
        e = ">"
    else:
        raise ValueError("not a DCD file")
    icntrl = struct.unpack(e + "20i", raw[8:88])
    has_cell = icntrl[10] == 1
    pos = 92
    ln = struct.unpack(e + "i", raw[pos:pos + 4])[0]
    pos += 8 + ln
    ln = struct.unpack(e + "i", raw[pos:pos + 4])[0]
    natoms = struct.unpack(e + "i", raw[pos + 4:pos + 8])[0]
    pos += 8 + ln
    per = (56 if has_cell else 0) + 3 * (8 + 4 * natoms)
    nfr = (len(raw) - pos) // per
    out = np.empty((nfr, natoms, 3), dtype=F32)
    for f in range(nfr):
        p = pos + f * per + (56 if has_cell else 0)
        for k in range(3):
            out[f, :, k] = np.frombuffer(raw, dtype=e + "f4", count=natoms, offset=p + 4)
            p += 8 + 4 * natoms
    return out


def load_frames(top: Topology, traj: str) -> np.ndarray:
    if traj.lower().endswith(".dcd"):
        return read_dcd(traj)
    if traj.lower().endswith(".pdb"):
        return Topology(traj).xyz0[None] if traj != getattr(top, "path", None) else top.xyz0[None]
    raise ValueError("oracle reads .dcd and .pdb only")


# ----------------------------------------------------------------------------------------------
# compute_contacts (compute_contacts.py:187-359) on indices, then text
# ----------------------------------------------------------------------------------------------
def run(top_path, traj_path, itypes, geom=None, beg=0, end=None, stride=1, sele1="protein", sele2=None,
        solv="", lipid="", ligand="", frames=None, collect_near=False):
    """Returns dict(total_frames, beg, end, stride, itypes, contacts=[(frame, itype, a, b, w1, w2)], top, setup).
    De-duplicated per fragment like the consumer (on canonical label lines)."""
    geom = dict(DEFAULT_GEOM if geom is None else geom)
    top = Topology(top_path)
    if frames is None:
        frames = top.xyz0[None] if traj_path == top_path else load_frames(top, traj_path)
    S = Setup(top, sele1, sele2, solv, lipid, ligand, xyz0=frames[0])
    if not S.sele1.any() or not S.sele2.any():
        raise SystemExit(1)
    sim_length = len(frames)
    end = (2 ** 63 - 1) if end is None else end
    beg = max(min(beg, sim_length - 1), 0)
    end = min(max(end, beg), sim_length - 1)
    stride = max(1, stride)
    contacts, near_all = [], []
    for beg_frame in range(beg, end + 1, TRAJ_FRAG_SIZE * stride):
        end_frame = min(beg_frame + TRAJ_FRAG_SIZE * stride - 1, sim_length - 1)
        frag = FragmentState()
        seen = set()
        for f in range(beg_frame, end_frame + 1, stride):
            near = [] if collect_near else None
            for c in compute_frame_contacts(S, frames[f], itypes, geom, frag, near):
                rec = canonical(top, f, c)
                if rec[0] not in seen:
                    seen.add(rec[0])
                    contacts.append(rec[1])
            if collect_near:
                near_all += [(f,) + t for t in near]
    return dict(total_frames=math.ceil((end - beg + 1) / stride), beg=beg, end=end, stride=stride, itypes=itypes,
                contacts=contacts, top=top, setup=S, frames=frames, near=near_all)


def canonical(top: Topology, frame: int, c):
    """compute_contacts.py:338-348: strip index, order atom_1/atom_2 (and the waters) by label."""
    itype, a, b = c[0], c[1], c[2]
    w = list(c[3:])
    la, lb = top.label(a), top.label(b)
    if lb < la:
        a, b, la, lb = b, a, lb, la
        if len(w) == 2:
            w = [w[1], w[0]]
    text = "\t".join([str(frame), itype, la, lb] + [top.label(k) for k in w])
    return text, (frame, itype, a, b) + tuple(w)


def write_tsv(res, output, distout=False):
    top, frames = res["top"], res["frames"]
    with open(output, "w") as fd:
        fd.write("# total_frames:%d beg:%d end:%d stride:%d interaction_types:%s\n" %
                 (res["total_frames"], res["beg"], res["end"], res["stride"], ",".join(res["itypes"])))
        fd.write("# Columns: frame, interaction_type, atom_1, atom_2[, atom_3[, atom_4]]%s\n" %
                 (", distance_1-2" if distout else ""))
        for rec in res["contacts"]:
            f, itype, a, b = rec[:4]
            cols = [str(f), itype, top.label(a), top.label(b)] + [top.label(k) for k in rec[4:]]
            if distout:
                cols.append("%.3f" % float(F32(np.sqrt(_d2(frames[f][a], frames[f][b])))))
            fd.write("\t".join(cols) + "\n")


# ----------------------------------------------------------------------------------------------
# Frequency semantics (Appendix A.12)
# ----------------------------------------------------------------------------------------------
FREQ_ITYPES = ["sb", "pc", "ps", "ts", "vdw", "hb", "lhb", "hbbb", "hbsb", "hbss", "wb", "wb2", "hbls", "hblb",
               "lwb", "lwb2", "br"]  # get_contact_frequencies.py:107-109


def intended_counts(res, itypes=None):
    """L2: count[(res1,res2)] = #frames with any counted-itype line joining the two residues."""
    its = set(FREQ_ITYPES if itypes is None else itypes)
    top = res["top"]
    per = defaultdict(set)
    for rec in res["contacts"]:
        f, it, a, b = rec[:4]
        if it not in its:
            continue
        ra, rb = top.reslabel(a), top.reslabel(b)
        per[(min(ra, rb), max(ra, rb))].add(f)
    return {k: len(v) for k, v in per.items()}
