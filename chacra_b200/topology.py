"""Topology compiler: PDB -> the static integer/float tables the CUDA contact kernels consume.

In the reference every one of these facts is re-derived per frame through Tcl strings
(``atomselect``, ``$sel get ...``); here they are computed once on the host and uploaded once.
Citations are relative to ``/root/reference/external/getcontacts``:

* atom records and labels ............ ``contact_calc/atom.py:5-51`` and ``contact_utils.py:201-258``
* element inference, vdW radii ........ ``contact_calc/atom.py:54-118``
* solvent / lipid / ligand macros ..... ``contact_calc/contact_utils.py:411-579``
* disulfide bridges ................... ``contact_calc/contact_utils.py:396-409``
* ligand anions / cations ............. ``contact_calc/contact_utils.py:605-665``
* ion, ring and hydrophobic classes ... ``salt_bridges.py:25-30``, ``pi_cation.py:24-31``,
  ``aromatics.py:71-113``, ``hydrophobics.py:8-41``
* VMD's own static state (hydrogen flag, ``protein``/``nucleic`` residue classes, distance-based
  bond search on the topology coordinates + CONECT) follows SURVEY.md Appendix A.2/A.3.
"""
from __future__ import annotations

import re
from dataclasses import dataclass, field

import numpy as np

from .selection import Selection

F32 = np.float32

# contact_calc/atom.py:107
BACKBONE_NAMES = ("O", "N", "OP1", "OP2", "O1P", "O2P", "O3'", "O2'", "O4'", "O5'")
# contact_calc/atom.py:111-118
VDW_RADII = dict(
    H=1.20, HE=1.40, LI=1.82, BE=1.53, B=1.92, C=1.70, N=1.55, O=1.52, F=1.47, NE=1.54, NA=2.27, MG=1.73, AL=1.84,
    SI=2.10, P=1.80, S=1.80, CL=1.75, AR=1.88, K=2.75, CA=2.31, SC=2.11, NI=1.63, CU=1.40, ZN=1.39, GA=1.87, GE=2.11,
    AS=1.85, SE=1.90, BR=1.85, KR=2.02, RB=3.03, SR=2.49, PD=1.63, AG=1.72, CD=1.58, IN=1.93, SN=2.17, SB=2.06,
    TE=2.06, I=1.98, XE=2.16, CS=3.43, BA=2.68, PT=1.75, AU=1.66, HG=1.55, TL=1.96, PB=2.02, BI=2.07, PO=1.97,
    AT=2.02, RN=2.20, FR=3.48, RA=2.83, U=1.86, GD=2.79)
DEFAULT_VDW_RADIUS = 1.7  # atom.py:31-35
_PERIODIC = frozenset(
    "H HE LI BE B C N O F NE NA MG AL SI P S CL AR K CA SC TI V CR MN FE CO NI CU ZN GA GE AS SE BR KR RB SR Y ZR NB "
    "MO TC RU RH PD AG CD IN SN SB TE I XE CS BA LA CE PR ND PM SM EU GD TB DY HO ER TM YB LU HF TA W RE OS IR PT AU "
    "HG TL PB BI PO AT RN FR RA AC TH PA U NP PU AM CM BK CF ES FM MD NO LR RF DB SG BH HS MT DS RG CN NH FL MC LV TS "
    "OG".split())

HIS_FAMILY = ("HIS", "HSD", "HSE", "HSP", "HIE", "HIP", "HID")
SOLVENT_RESNAMES = "H2O HH0 OHH HOH OH2 SOL WAT TIP TIP2 TIP3 TIP4 T3P IP3".split()      # contact_utils.py:448
LIPID_RESNAMES = "DLPE DMPC DPPC GPC LPPC PALM PC PGCL POPC POPE".split()                # contact_utils.py:503
HYDROPHOBIC_RESNAMES = "ALA PHE GLY ILE LEU PRO VAL TRP".split()                         # hydrophobics.py:9
NUCLEIC_RESNAMES = "C U G A DC DT DG DA".split()
RING_ATOM_NAMES = {"PHE": ("CG", "CE1", "CE2"), "TYR": ("CG", "CE1", "CE2"), "TRP": ("CD2", "CZ2", "CZ3")}
RING_ATOM_NAMES.update({h: ("CG", "CE1", "CD2") for h in HIS_FAMILY})
NUCLEIC_RING_NAMES = ("C4", "C2", "C6")
_VMD_BOND_RADIUS = {"H": 1.0, "C": 1.5, "N": 1.4, "O": 1.3, "F": 1.2, "S": 1.9}


def element_from_name(resname: str, name: str) -> str:
    """Element symbol from the atom name when the PDB element column is empty (atom.py:54-103)."""
    m = re.search(r"[a-zA-Z]+", name)
    if not m:
        return "?"
    s = m.group(0).upper()
    one, two = s[:1], s[:2]
    if len(s) == 1:
        return one if one in _PERIODIC else "?"
    if one in _PERIODIC and two in _PERIODIC:
        return two if resname == name else one
    if one in _PERIODIC:
        return one
    return two if two in _PERIODIC else "?"


# ------------------------------------------------------------------------------------------------
# geometry helpers on float32 coordinates (host side, setup only)
# ------------------------------------------------------------------------------------------------
def dist2_f32(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """(dx*dx + dy*dy) + dz*dz with every operation rounded to binary32."""
    d = (a.astype(F32) - b.astype(F32)).astype(F32)
    return ((d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]).astype(F32) + d[..., 2] * d[..., 2]).astype(F32)


def close_pairs(xa: np.ndarray, xb: np.ndarray, r: float, same: bool = False) -> np.ndarray:
    """Candidate index pairs with distance <= r (slightly padded); callers re-test exactly."""
    if len(xa) == 0 or len(xb) == 0:
        return np.zeros((0, 2), np.int64)
    pad = float(r) * (1 + 1e-5) + 1e-4
    try:
        from scipy.spatial import cKDTree
    except Exception:  # pragma: no cover - scipy ships with the image
        cKDTree = None
    if cKDTree is not None:
        ta = cKDTree(np.asarray(xa, np.float64))
        if same:
            return ta.query_pairs(pad, output_type="ndarray").astype(np.int64)
        tb = cKDTree(np.asarray(xb, np.float64))
        coo = ta.sparse_distance_matrix(tb, pad, output_type="coo_matrix")
        return np.stack([coo.row, coo.col], 1).astype(np.int64)
    chunks = []
    xb64 = np.asarray(xb, np.float64)
    step = max(1, 2_000_000 // len(xb))
    for s in range(0, len(xa), step):
        d = np.asarray(xa[s:s + step], np.float64)[:, None, :] - xb64[None]
        ii, jj = np.nonzero((d * d).sum(-1) <= pad * pad)
        ii += s
        if same:
            keep = ii < jj
            ii, jj = ii[keep], jj[keep]
        chunks.append(np.stack([ii, jj], 1))
    return np.concatenate(chunks).astype(np.int64)


def within_mask(xyz, candidates, reference, r: float) -> np.ndarray:
    """VMD ``<candidates> and within r of <reference>``: float32 d2 <= r*r (Appendix A.6)."""
    out = np.zeros(len(xyz), bool)
    ci, ri = np.nonzero(candidates)[0], np.nonzero(reference)[0]
    if len(ci) == 0 or len(ri) == 0:
        return out
    p = close_pairs(xyz[ci], xyz[ri], r)
    if len(p):
        ok = dist2_f32(xyz[ci[p[:, 0]]], xyz[ri[p[:, 1]]]) <= F32(r) * F32(r)
        out[ci[p[ok, 0]]] = True
    return out


# ------------------------------------------------------------------------------------------------
# PDB -> AtomTable
# ------------------------------------------------------------------------------------------------
@dataclass
class AtomTable:
    """Per-atom columns in VMD index order (= order of ATOM/HETATM records)."""
    name: np.ndarray
    resname: np.ndarray
    chain: np.ndarray
    resid: np.ndarray
    icode: np.ndarray
    segname: np.ndarray
    altloc: np.ndarray
    element: np.ndarray
    xyz: np.ndarray                      # float32 [N][3], topology coordinates
    conect: list = field(default_factory=list)
    serial: np.ndarray = None

    def __post_init__(self):
        n = self.n = len(self.name)
        first = np.array([s[:1] for s in self.name])
        second = np.array([s[1:2] for s in self.name])
        self.is_hydrogen = (first == "H") | (np.char.isdigit(first) & (second == "H"))       # Appendix A.2
        self.vdw_radius = np.array([VDW_RADII.get(e, DEFAULT_VDW_RADIUS) for e in self.element], np.float64)
        # residues = runs of consecutive atoms with the same (chain, segname, resid, icode)
        if n:
            brk = np.ones(n, bool)
            brk[1:] = (self.chain[1:] != self.chain[:-1]) | (self.segname[1:] != self.segname[:-1]) | \
                      (self.resid[1:] != self.resid[:-1]) | (self.icode[1:] != self.icode[:-1])
            self.residue = np.cumsum(brk) - 1
        else:
            self.residue = np.zeros(0, np.int64)
        self.n_residues = int(self.residue[-1]) + 1 if n else 0
        self.is_protein = self._residues_holding(("N", "CA", "C"), ("O", "OT1", "OT2", "OXT", "O1", "O2"))
        nuc = np.isin(self.name, np.array(["P", "O1P", "O2P", "OP1", "OP2", "C3*", "C3'", "O3*", "O3'", "C4*", "C4'",
                                           "C5*", "C5'", "O5*", "O5'"]))
        cnt = np.bincount(self.residue[nuc], minlength=self.n_residues)
        self.is_nucleic = (cnt >= 4)[self.residue] & ~self.is_protein if n else np.zeros(0, bool)
        self.bonds = self._bond_search()
        self._neighbours = None

    def _residues_holding(self, required, any_of):
        ok = np.ones(self.n_residues, bool)
        for nm in required:
            has = np.zeros(self.n_residues, bool)
            has[self.residue[self.name == nm]] = True
            ok &= has
        has = np.zeros(self.n_residues, bool)
        has[self.residue[np.isin(self.name, np.array(any_of))]] = True
        return (ok & has)[self.residue] if self.n else np.zeros(0, bool)

    def _bond_search(self) -> np.ndarray:
        """VMD distance bond search on the topology coordinates + CONECT (Appendix A.3).
        Returns a sorted unique int64 [B][2] array with i < j."""
        n = self.n
        rad = np.empty(n, F32)
        for i, nm in enumerate(self.name):
            s = nm.lstrip("0123456789")
            rad[i] = _VMD_BOND_RADIUS.get(s[:1].upper(), 1.5)
        pairs = np.zeros((0, 2), np.int64)
        if n > 1:
            cand = close_pairs(self.xyz, self.xyz, 1.2 * float(rad.max()) + 1e-3, same=True)
            if len(cand):
                i, j = cand[:, 0], cand[:, 1]
                d2 = dist2_f32(self.xyz[i], self.xyz[j])
                cut = (F32(0.6) * (rad[i] + rad[j])).astype(F32)
                ok = (d2 >= F32(0.001)) & (d2 < cut * cut) & ~(self.is_hydrogen[i] & self.is_hydrogen[j])
                ai, aj = self.altloc[i], self.altloc[j]
                ok &= ~((ai != aj) & (ai != "") & (aj != ""))
                pairs = np.stack([np.minimum(i, j)[ok], np.maximum(i, j)[ok]], 1)
        extra = []
        if self.conect:
            first_of = {}
            for k, s in enumerate(self.serial):
                first_of.setdefault(s, k)
            for rec in self.conect:
                a = first_of.get(rec[0])
                if a is None:
                    continue
                for s in rec[1:]:
                    b = first_of.get(s)
                    if b is not None and b != a:
                        extra.append((min(a, b), max(a, b)))
        if extra:
            pairs = np.concatenate([pairs, np.array(extra, np.int64)])
        return np.unique(pairs, axis=0) if len(pairs) else pairs

    @property
    def neighbours(self):
        """CSR adjacency (start[N+1], adj[2B]) of the static bond list."""
        if self._neighbours is None:
            b = self.bonds
            src = np.concatenate([b[:, 0], b[:, 1]]) if len(b) else np.zeros(0, np.int64)
            dst = np.concatenate([b[:, 1], b[:, 0]]) if len(b) else np.zeros(0, np.int64)
            order = np.lexsort((dst, src))
            start = np.zeros(self.n + 1, np.int64)
            np.cumsum(np.bincount(src, minlength=self.n), out=start[1:])
            self._neighbours = (start, dst[order])
        return self._neighbours

    def atom_label(self, i: int) -> str:
        """``chain:resname:resid+icode:name`` -- atom.py:40-48 without the trailing VMD index."""
        return "%s:%s:%d%s:%s" % (self.chain[i], self.resname[i], self.resid[i], self.icode[i], self.name[i])

    def atom_labels(self) -> np.ndarray:
        return np.array([self.atom_label(i) for i in range(self.n)], dtype=object)

    def residue_labels(self) -> np.ndarray:
        return np.array(["%s:%s:%d%s" % (self.chain[i], self.resname[i], self.resid[i], self.icode[i])
                         for i in range(self.n)], dtype=object)


def read_pdb(path: str) -> AtomTable:
    """Fixed-column PDB reader; stops at the first ENDMDL."""
    cols = dict(name=[], resname=[], chain=[], resid=[], icode=[], segname=[], altloc=[], element=[], serial=[])
    xyz, conect = [], []
    with open(path) as fh:
        for line in fh:
            rec = line[:6]
            if rec in ("ATOM  ", "HETATM"):
                cols["serial"].append(line[6:11].strip())
                cols["name"].append(line[12:16].strip())
                cols["altloc"].append(line[16:17].strip())
                cols["resname"].append(line[17:21].strip())
                cols["chain"].append(line[21:22].strip())
                cols["resid"].append(int(line[22:26]))
                cols["icode"].append(line[26:27].strip())
                xyz.append((float(line[30:38]), float(line[38:46]), float(line[46:54])))
                cols["segname"].append(line[72:76].strip())
                el = line[76:78].strip().upper() or "X"
                if el == "X":                                      # atom.py:25-29
                    el = element_from_name(cols["resname"][-1], cols["name"][-1])
                cols["element"].append(el)
            elif rec == "CONECT":
                f = [line[6 + 5 * k: 11 + 5 * k].strip() for k in range(5)]
                f = [x for x in f if x]
                if len(f) >= 2:
                    conect.append(f)
            elif rec.strip() == "ENDMDL":
                break
    serial = cols.pop("serial")
    arr = {k: (np.array(v, np.int64) if k == "resid" else np.array(v, dtype=str)) for k, v in cols.items()}
    for k in arr:                                                   # keep zero-length strings comparable
        if arr[k].dtype.kind == "U" and arr[k].dtype.itemsize == 0:
            arr[k] = arr[k].astype("U1")
    return AtomTable(xyz=np.array(xyz, np.float64).astype(F32).reshape(-1, 3), conect=conect, serial=serial, **arr)


# ------------------------------------------------------------------------------------------------
# Compiled tables
# ------------------------------------------------------------------------------------------------
# bits of PairAtoms.info (must match csrc/chacra_b200.cu)
INFO_ELEM_MASK = 0xF
INFO_S1 = 1 << 4
INFO_S2 = 1 << 5
INFO_HBU = 1 << 6        # heavy atom inside the hydrogen-bond selection U (hbonds.py:62/67)
INFO_HAS_H = 1 << 7      # ... that has at least one bonded hydrogen
INFO_ANION = 1 << 8
INFO_CATION = 1 << 9
INFO_HPB1 = 1 << 10      # hydrophobic base carbon in sele1 (hydrophobics.py:9-11)
INFO_HPB2 = 1 << 11      # ... in sele2
INFO_VDW = 1 << 12       # takes part in the vdW search (heavy, in sele1|sele2)
INFO_SOLV = 1 << 13      # solvent heavy atom (water-bridge partner, hbonds.py:64-82)
INFO_LIGAND = 1 << 14
INFO_BB = 1 << 15        # atom.py:37-38 is_bb
INFO_CYS = 1 << 16
INFO_XBOND = 1 << 17     # has a bond that the residue-distance filters do not already exclude


@dataclass
class CompiledTopology:
    atoms: AtomTable
    sele1: np.ndarray
    sele2: np.ndarray
    same_selection: bool
    solvent: np.ndarray
    lipid: np.ndarray
    ligand: np.ndarray
    # pair atoms ("P atoms"): every heavy atom the cell-list pair search must see
    pair_atom: np.ndarray = None        # int32 [NP] atom index
    pair_info: np.ndarray = None        # uint32 [NP]
    pair_lkey: np.ndarray = None        # int32 [NP]; same chain & |dresid| < k  <=>  |dlkey| < k
    pair_res: np.ndarray = None         # int32 [NP] residue-label rank (string order)
    pair_resid: np.ndarray = None       # int32 [NP]
    pair_h_start: np.ndarray = None     # int32 [NP+1] CSR into aux slots holding the bonded hydrogens
    pair_bond_start: np.ndarray = None  # int32 [NP+1]
    pair_bond_adj: np.ndarray = None    # int32 [..] atom indices bonded to the pair atom
    aux_atom: np.ndarray = None         # int32 [NA] atoms copied verbatim to shared memory (polar H, rings, cations)
    ring_aux: np.ndarray = None         # int32 [NR][3] aux slots of the geometry triplet, index order (aromatics.py:71-113)
    ring_pc_aux: np.ndarray = None      # int32 [NR][3] aux slots in label order (pi_cation.py:101) or -1
    ring_sel_aux: np.ndarray = None     # int32 [NR][3] aux slots of ring atoms inside sele1|sele2 (-1 padded)
    ring_sel_flags: np.ndarray = None   # uint8 [NR][3] bit0 sele1, bit1 sele2
    ring_res: np.ndarray = None         # int32 [NR] residue-label rank
    ring_rep: np.ndarray = None         # int32 [NR] atom with the smallest name (contact_utils.py:335-347)
    ring_rep_aux: np.ndarray = None     # int32 [NR] aux slot of ring_rep
    ring_atoms: np.ndarray = None       # int32 [NR][3] atom indices of ring_sel_aux
    cation_aux: np.ndarray = None       # int32 [NC] aux slot
    cation_atom: np.ndarray = None      # int32 [NC]
    cation_flags: np.ndarray = None     # uint8 [NC] bit0 sele1, bit1 sele2
    cation_res: np.ndarray = None       # int32 [NC]
    cation_excl_ring: np.ndarray = None # int32 [NC] ring owning an atom bonded to the cation (never a pc partner) or -1
    disulfide: np.ndarray = None        # int32 [ND][2] resid pairs
    elem_symbols: list = None
    elem_radius: np.ndarray = None      # float64 [K]
    residue_label: np.ndarray = None    # object [R] sorted label strings (rank -> label)
    atom_res: np.ndarray = None         # int32 [N] residue-label rank of every atom
    hp_base_atoms: np.ndarray = None    # int32 [NHB] pair-atom indices (into pair_atom) of hydrophobic base carbons
    elem_is_c: np.ndarray = None        # uint8 [N] 1 = element C, 2 = element H, 0 = other (for the 1.7 A rule)
    elem_class: np.ndarray = None       # uint8 [K] the same code per element class
    pair_solv_ord: np.ndarray = None    # int32 [NP] ordinal among solvent pair atoms or -1
    n_solvent_pair: int = 0
    n_pair_a: int = 0                   # pair atoms [0, n_pair_a) live in grid A, the rest (unselected solvent) in grid B

    @property
    def n_atoms(self):
        return self.atoms.n

    @property
    def n_residues(self):
        return len(self.residue_label)


def compile_topology(atoms: AtomTable, xyz0: np.ndarray | None = None, sele1: str = "protein", sele2: str | None = None,
                     solv_sele: str = "", lipid_sele: str = "", ligand_sele: str = "",
                     vdw_res_diff: int = 2, hbond_res_diff: int = 1) -> CompiledTopology:
    """Everything ``compute_contacts`` derives once per call (compute_contacts.py:226-238).

    ``xyz0`` is frame 0 of the *trajectory* (the reference evaluates its setup selections on frames
    0..1 of the trajectory, contact_utils.py:89-116); bonds come from the topology coordinates.
    """
    n = atoms.n
    xyz0 = atoms.xyz if xyz0 is None else np.ascontiguousarray(xyz0, F32)
    sele2 = sele1 if sele2 is None else sele2
    base = Selection(atoms, {}, xyz0)
    solvent = base(solv_sele) if solv_sele else np.isin(atoms.resname, np.array(SOLVENT_RESNAMES))
    lipid = Selection(atoms, {"solv": solvent}, xyz0)(lipid_sele) if lipid_sele else \
        np.isin(atoms.resname, np.array(LIPID_RESNAMES))
    macros = {"solv": solvent, "lipid": lipid}
    ligand = Selection(atoms, macros, xyz0)(ligand_sele) if ligand_sele else \
        ~(lipid | solvent | atoms.is_protein | atoms.is_nucleic)
    macros["ligand"] = ligand
    sel = Selection(atoms, macros, xyz0)
    m1, m2 = sel(sele1), sel(sele2)
    ct = CompiledTopology(atoms=atoms, sele1=m1, sele2=m2, same_selection=(sele1 == sele2), solvent=solvent,
                          lipid=lipid, ligand=ligand)
    either = m1 | m2
    heavy = ~atoms.is_hydrogen
    nm, rn = atoms.name, atoms.resname

    # --- residue-label ranks: frequency keys are the label strings, ordered as Python strings ----------
    # (only residues that own a selected atom can appear in a contact line; solvent outside the selection would
    # otherwise blow the pair table up)
    rlab = atoms.residue_labels()
    uniq = sorted(set(rlab[either].tolist()))
    rank = {s: k for k, s in enumerate(uniq)}
    ct.residue_label = np.array(uniq, dtype=object)
    ct.atom_res = np.array([rank.get(s, -1) for s in rlab], np.int32)

    # --- ion classes (salt_bridges.py:25-30) + ligand ions (contact_utils.py:605-665) -------------------
    anion = ((rn == "ASP") & np.isin(nm, ["OD1", "OD2"])) | ((rn == "GLU") & np.isin(nm, ["OE1", "OE2"])) | \
            (np.isin(rn, NUCLEIC_RESNAMES) & np.isin(nm, ["OP1", "OP2", "O1P", "O2P"]))
    prot_cation = (np.isin(rn, HIS_FAMILY) & np.isin(nm, ["ND1", "NE2"])) | ((rn == "LYS") & (nm == "NZ")) | \
                  ((rn == "ARG") & np.isin(nm, ["NH1", "NH2"]))
    cation = prot_cation.copy()
    lig_an, lig_cat = _ligand_ions(atoms, ligand, xyz0)
    anion[lig_an] = True
    cation[lig_cat] = True

    # --- ring atoms (pi_cation.py:27-31) -----------------------------------------------------------
    ring = np.zeros(n, bool)
    for r_, names in RING_ATOM_NAMES.items():
        ring |= (rn == r_) & np.isin(nm, names)
    ring |= atoms.is_nucleic & np.isin(nm, NUCLEIC_RING_NAMES) & ~np.isin(rn, list(RING_ATOM_NAMES))

    # --- hydrogen-bond selection U (hbonds.py:62/67) -------------------------------------------------
    hbu = either & ~(sel("carbon") | sel("sulfur") | solvent) & heavy
    start, adj = atoms.neighbours
    nh = np.zeros(n, np.int64)
    for i in np.nonzero(hbu | (solvent & heavy))[0]:
        nh[i] = int(atoms.is_hydrogen[adj[start[i]:start[i + 1]]].sum())

    # --- hydrophobic base carbons (hydrophobics.py:8-41) ---------------------------------------------
    hp_base = (np.isin(rn, HYDROPHOBIC_RESNAMES) | ligand) & (atoms.element == "C")

    # --- disulfides (contact_utils.py:396-409) --------------------------------------------------------
    ct.disulfide = _disulfides(atoms, xyz0)

    # --- element classes & thresholds ------------------------------------------------------------
    vdw_atoms = either & heavy
    # atoms that are neither carbon nor hydrogen by ELEMENT decide the 1.7 A hydrophobic-carbon rule whether or
    # not they are selected, so they join the grid too (they carry no selection bit and never form a contact)
    not_ch = ~np.isin(atoms.element, ["C", "H"])
    # solvent heavy atoms can mediate water bridges (hbonds.py:64-82) whether or not they are selected
    solv_heavy = solvent & heavy
    pmask = vdw_atoms | hbu | ((anion | cation) & either & heavy) | (hp_base & either) | not_ch | solv_heavy
    # grid B = solvent atoms outside both selections: they only ever pair with non-solvent atoms (as bridge waters),
    # so the kernels keep them in a second cell list that is searched from the other side only.  They go last.
    gridb = solv_heavy & ~either
    pidx = np.concatenate([np.nonzero(pmask & ~gridb)[0], np.nonzero(gridb)[0]])
    ct.n_pair_a = int((pmask & ~gridb).sum())
    symbols = sorted(set(atoms.element[pidx].tolist())) or ["C"]
    if len(symbols) > 15:
        raise ValueError("more than 15 distinct elements among selected heavy atoms (%d)" % len(symbols))
    sym_id = {s: k for k, s in enumerate(symbols)}
    ct.elem_symbols = symbols
    ct.elem_radius = np.array([VDW_RADII.get(s, DEFAULT_VDW_RADIUS) for s in symbols], np.float64)

    # --- locality key: same chain and |resid_i - resid_j| < k  <=>  |lkey_i - lkey_j| < k ------------------
    kmax = max(int(vdw_res_diff), int(hbond_res_diff), 1) + 1
    lkey = np.zeros(n, np.int64)
    off = 0
    for ch in sorted(set(atoms.chain.tolist())):
        msk = atoms.chain == ch
        lo, hi = int(atoms.resid[msk].min()), int(atoms.resid[msk].max())
        lkey[msk] = atoms.resid[msk] - lo + off
        off += (hi - lo) + kmax + 1
    if off >= 2 ** 31:
        raise ValueError("residue id range too large for the 32-bit locality key")

    # --- non-local bonds: bonds not already excluded by same-chain |dresid| < 1 (the weakest filter) ----
    xbond = np.zeros(n, bool)
    b = atoms.bonds
    if len(b):
        nonlocal_b = ~((atoms.chain[b[:, 0]] == atoms.chain[b[:, 1]]) &
                       (np.abs(atoms.resid[b[:, 0]] - atoms.resid[b[:, 1]]) < min(int(vdw_res_diff), int(hbond_res_diff))))
        xbond[b[nonlocal_b, 0]] = True
        xbond[b[nonlocal_b, 1]] = True

    info = np.zeros(n, np.uint32)
    info[pidx] = np.array([sym_id[e] for e in atoms.element[pidx]], np.uint32)
    for bit, mask in ((INFO_S1, m1), (INFO_S2, m2), (INFO_HBU, hbu), (INFO_HAS_H, (hbu | solv_heavy) & (nh > 0)),
                      (INFO_ANION, anion & either), (INFO_CATION, cation & either), (INFO_HPB1, hp_base & m1),
                      (INFO_HPB2, hp_base & m2), (INFO_LIGAND, ligand), (INFO_BB, np.isin(nm, BACKBONE_NAMES)),
                      (INFO_CYS, rn == "CYS"), (INFO_XBOND, xbond), (INFO_VDW, vdw_atoms), (INFO_SOLV, solv_heavy)):
        info[mask] |= np.uint32(bit)
    ct.pair_atom = pidx.astype(np.int32)
    ct.pair_info = info[pidx]
    ct.pair_lkey = lkey[pidx].astype(np.int32)
    ct.pair_res = ct.atom_res[pidx]
    ct.pair_resid = atoms.resid[pidx].astype(np.int32)
    ct.hp_base_atoms = np.nonzero((ct.pair_info & (INFO_HPB1 | INFO_HPB2)) != 0)[0].astype(np.int32)
    is_sv = (ct.pair_info & INFO_SOLV) != 0
    ct.pair_solv_ord = np.where(is_sv, np.cumsum(is_sv) - 1, -1).astype(np.int32)
    ct.n_solvent_pair = int(is_sv.sum())
    el = atoms.element
    ct.elem_is_c = np.where(el == "C", 1, np.where(el == "H", 2, 0)).astype(np.uint8)
    ct.elem_class = np.array([1 if s == "C" else (2 if s == "H" else 0) for s in symbols], np.uint8)

    # --- aux atoms: polar hydrogens, ring atoms, pi-cation cations -------------------------------------
    # slots [0, NH) hold the donor hydrogens in CSR order (pair_h_start), the rest are handed out on demand
    aux, slot = [], {}
    h_start = np.zeros(len(pidx) + 1, np.int32)
    for k, a in enumerate(pidx):
        if hbu[a] or solv_heavy[a]:
            aux.extend(int(h) for h in adj[start[a]:start[a + 1]] if atoms.is_hydrogen[h])
        h_start[k + 1] = len(aux)
    ct.pair_h_start = h_start
    n_h_slots = len(aux)

    def other_slot(a):
        a = int(a)
        if a not in slot:
            slot[a] = len(aux)
            aux.append(a)
        return slot[a]

    bs = np.zeros(len(pidx) + 1, np.int32)
    badj = []
    for k, a in enumerate(pidx):
        nb = adj[start[a]:start[a + 1]]
        badj.extend(int(x) for x in nb)
        bs[k + 1] = len(badj)
    ct.pair_bond_start, ct.pair_bond_adj = bs, np.array(badj, np.int32)

    # rings: one entry per residue label that owns ring-named atoms
    rings = {}
    for a in np.nonzero(ring)[0]:
        rings.setdefault(rlab[a], []).append(int(a))
    r_aux, r_pc, r_sel, r_self, r_res, r_rep, r_atoms = [], [], [], [], [], [], []
    for lab in sorted(rings, key=lambda s: rings[s][0]):
        members = rings[lab]
        a0 = members[0]
        # geometry triplet of ps/ts: chain + resname + resid (insertion code ignored), ring names, index order
        names = RING_ATOM_NAMES.get(rn[a0], NUCLEIC_RING_NAMES)
        trip = np.nonzero((atoms.chain == atoms.chain[a0]) & (rn == rn[a0]) & (atoms.resid == atoms.resid[a0]) &
                          np.isin(nm, names))[0].tolist()
        sel_members = [a for a in members if either[a]]
        if not sel_members:
            continue
        if len(sel_members) > 3 or len(members) > 3:
            raise ValueError("residue %s has more than three ring atoms (alternate locations?); unsupported" % lab)
        pc_order = sorted(sel_members, key=lambda a: "%s:%d" % (atoms.atom_label(a), a))       # pi_cation.py:101
        r_aux.append([other_slot(a) for a in trip] if len(trip) == 3 else [-1, -1, -1])
        r_pc.append([other_slot(a) for a in pc_order] if len(sel_members) == 3 else [-1, -1, -1])
        pad = sel_members + [-1] * (3 - len(sel_members))
        r_sel.append([other_slot(a) if a >= 0 else -1 for a in pad])
        r_self.append([(int(m1[a]) | (int(m2[a]) << 1)) if a >= 0 else 0 for a in pad])
        r_atoms.append(pad)
        r_res.append(int(ct.atom_res[a0]))
        rep_pool = trip if len(trip) == 3 else sel_members
        r_rep.append(min(rep_pool, key=lambda a: nm[a]))
    ct.ring_aux = np.array(r_aux, np.int32).reshape(-1, 3)
    ct.ring_pc_aux = np.array(r_pc, np.int32).reshape(-1, 3)
    ct.ring_sel_aux = np.array(r_sel, np.int32).reshape(-1, 3)
    ct.ring_sel_flags = np.array(r_self, np.uint8).reshape(-1, 3)
    ct.ring_atoms = np.array(r_atoms, np.int32).reshape(-1, 3)
    ct.ring_res = np.array(r_res, np.int32)
    ct.ring_rep = np.array(r_rep, np.int32)
    ct.ring_rep_aux = np.array([other_slot(a) for a in r_rep], np.int32)

    cat = np.nonzero(prot_cation & either)[0]
    ct.cation_atom = cat.astype(np.int32)
    ct.cation_aux = np.array([other_slot(a) for a in cat], np.int32)
    ct.cation_flags = (m1[cat].astype(np.uint8) | (m2[cat].astype(np.uint8) << 1)).astype(np.uint8)
    ct.cation_res = ct.atom_res[cat]
    # measure contacts drops bonded pairs, so a ring with an atom bonded to the cation (HIS ND1/NE2 and
    # its own imidazole) can never collect its three atoms (pi_cation.py:70-98)
    ring_of_atom = {int(a): k for k in range(len(r_atoms)) for a in r_atoms[k] if a >= 0}
    excl = np.full(len(cat), -1, np.int32)
    for k, a in enumerate(cat):
        owners = {ring_of_atom[int(x)] for x in adj[start[a]:start[a + 1]] if int(x) in ring_of_atom}
        if len(owners) > 1:
            raise ValueError("cation %s is bonded to more than one aromatic ring; unsupported" % atoms.atom_label(a))
        if owners:
            excl[k] = owners.pop()
    ct.cation_excl_ring = excl
    for a_, b_ in atoms.bonds:
        if int(a_) in ring_of_atom and int(b_) in ring_of_atom and ring_of_atom[int(a_)] != ring_of_atom[int(b_)]:
            raise ValueError("covalently linked aromatic rings (%s - %s) are unsupported" %
                             (atoms.atom_label(a_), atoms.atom_label(b_)))
    ct.aux_atom = np.array(aux, np.int32)
    ct.n_h_slots = n_h_slots
    return ct


def _disulfides(atoms: AtomTable, xyz0) -> np.ndarray:
    sg = atoms.name == "SG"
    cys = atoms.resname == "CYS"
    found = []
    for rid in sorted(set(atoms.resid[cys].tolist())):
        got = within_mask(xyz0, sg, sg & (atoms.resid == rid), 2.1)
        s = sorted(set(atoms.resid[got].tolist()))
        if len(s) == 2 and s not in found:
            found.append(s)
    return np.array(found, np.int32).reshape(-1, 2)


def _ligand_ions(atoms: AtomTable, ligand, xyz0):
    """contact_utils.py:605-665: metal-labelled ligand atoms are cations; carboxylate oxygens anions."""
    anions, cations = [], []
    metals = ("MG", "MN", "RH", "ZN", "FE", "BI", "AS", "AG")
    everything = np.ones(atoms.n, bool)
    for a in np.nonzero(ligand)[0]:
        label = "%s:%d" % (atoms.atom_label(a), a)
        if any(m in label for m in metals):
            cations.append(int(a))
            continue
        if atoms.element[a] != "C":
            continue
        one = np.zeros(atoms.n, bool)
        one[a] = True
        nb = [int(k) for k in np.nonzero(within_mask(xyz0, everything, one, 1.95))[0] if k != a]
        if len(nb) != 3:
            continue
        u = (xyz0[nb[0]] - xyz0[a]).astype(F32)
        v = (xyz0[nb[1]] - xyz0[a]).astype(F32)
        c = float(np.dot(u, v)) / (float(np.sqrt(np.dot(u, u))) * float(np.sqrt(np.dot(v, v))))
        ang = np.degrees(np.arccos(max(-1.0, min(1.0, c))))
        if not (115.0 < ang < 125.0):
            continue
        els = [atoms.element[k] for k in nb]
        if els.count("C") == 1 and els.count("O") == 2:
            anions += [k for k in nb if atoms.element[k] == "O"]
    return np.array(anions, np.int64), np.array(cations, np.int64)


def vdw_threshold_tables(radius: np.ndarray, epsilon: float, soft_cutoff: float):
    """Per element-pair acceptance thresholds for ``distance < r_i + r_j + epsilon``.

    The reference parses VMD's float32 ``sqrtf(d2)`` into a Python float and compares in double
    (vanderwaals.py:90-93).  ``sqrtf`` is monotone, so for every element pair there is a largest
    binary32 d2 that is still accepted; the kernels compare d2 against it.
    The soft-cutoff test ``d2 <= float32(soft)^2`` of ``measure contacts`` is folded in.
    Returns (d2max float32 [K][K], cut float64 [K][K])."""
    k = len(radius)
    cut = radius[:, None] + radius[None, :] + float(epsilon)
    soft2 = F32(soft_cutoff) * F32(soft_cutoff)
    d2max = np.zeros((k, k), F32)
    for i in range(k):
        for j in range(k):
            c = float(cut[i, j])
            t = F32(c * c)
            # walk to the largest float32 t with (double)sqrtf(t) < c
            while float(np.sqrt(t, dtype=F32)) < c:
                t = np.nextafter(t, F32(np.inf), dtype=F32)
            while not (float(np.sqrt(t, dtype=F32)) < c):
                t = np.nextafter(t, F32(-np.inf), dtype=F32)
            d2max[i, j] = min(t, soft2)
    return d2max, cut
