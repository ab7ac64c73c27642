"""Test infrastructure / CPU baseline: ctypes driver of the C++/OpenMP twin (oracle/getcontacts_cpu.cpp).

Static tables come from the numpy oracle's own ``Topology`` / ``Setup`` (nothing from the product package), the per-frame
work runs in ``libgc_cpu.so`` on all host cores.  ``run`` returns the layout of ``getcontacts_oracle.run`` restricted to
what the twin computes (no solvent / water bridges: a topology with solvent raises).  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import getcontacts_oracle as o  # noqa: E402

LIB = os.path.join(HERE, "_build", "libgc_cpu.so")
ITYPE_NAMES = ["sb", "sblp", "sbpl", "sbll", "pc", "ps", "ts", "vdw", "hbbb", "hbsb", "hbss", "hbll", "hblb", "hbls",
               "wb", "wb2", "lwb", "lwb2", "hp", "hplp", "hpll"]          # enum chc_itype of include/chacra_b200.h
ITYPE_BITS = {"sb": 1, "pc": 2, "ps": 4, "ts": 8, "vdw": 16, "hb": 32, "hp": 64}
i32p, u8p, f64p = C.POINTER(C.c_int32), C.POINTER(C.c_uint8), C.POINTER(C.c_double)


class GcTopo(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n_atoms", "n_reskeys", "same", "n_dis", "n_ring_groups", "n_ring_res", "reserved0", "reserved1")] + [
        ("flags", u8p), ("flags2", u8p), ("vdw", f64p), ("chain", i32p), ("resid", i32p), ("reskey", i32p), ("bond_start", i32p),
        ("bond_adj", i32p), ("dis", i32p), ("rg_start", i32p), ("rg_atoms", i32p), ("rg_reskey", i32p), ("rr_reskey", i32p),
        ("rr_trip", i32p), ("rr_rep", i32p), ("rr_sel_start", i32p), ("rr_sel", i32p)]


class GcParams(C.Structure):
    _fields_ = [("itypes", C.c_uint32)] + [(n, C.c_int32) for n in ("hb_rd", "vdw_rd", "fragment", "threads", "reserved")] + [
        (n, C.c_double) for n in ("sb_dist", "pc_dist", "pc_angle", "ps_dist", "ps_angle", "ps_psi", "ts_dist", "ts_angle", "ts_psi",
                                  "hb_dist", "hb_angle", "eps", "soft")]


def build(force=False):
    src = os.path.join(HERE, "getcontacts_cpu.cpp")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.run(["make", "-s", "-C", HERE], check=True)
    return LIB


_lib = None


def load():
    global _lib
    if _lib is None:
        lib = C.CDLL(build())
        lib.gc_run.restype = C.c_int
        lib.gc_run.argtypes = [C.POINTER(GcTopo), C.POINTER(GcParams), C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64,
                               C.POINTER(C.c_int64)]
        lib.gc_threads.restype = C.c_int
        _lib = lib
    return _lib


class Twin:
    """Static tables of one (topology, selections) pair, ready for ``contacts`` / ``counts`` calls."""

    def __init__(self, top_path, sele1="protein", sele2=None, xyz0=None, geom=None, itypes=("sb", "pc", "ps", "ts", "vdw", "hb", "hp")):
        top = o.Topology(top_path)
        xyz0 = top.xyz0 if xyz0 is None else np.ascontiguousarray(xyz0, dtype=np.float32)
        S = o.Setup(top, sele1, sele2, "", "", "", xyz0=xyz0)
        if (S.solv & (S.sele1 | S.sele2)).any() or S.solv.any():
            raise NotImplementedError("the C++ twin has no water-bridge path; use the numpy oracle for solvated systems")
        self.top, self.S = top, S
        self.geom = dict(o.DEFAULT_GEOM if geom is None else geom)
        self.itypes = list(itypes)
        n = top.n
        keep = {}

        def arr(name, v, dt):
            keep[name] = np.ascontiguousarray(v, dtype=dt)
            return keep[name]
        selany = S.sele1 | S.sele2
        U = selany & ~(S.carbon | S.sulfur | S.solv)
        flags = (top.is_h * 1 + S.sele1 * 2 + S.sele2 * 4 + S.anion * 8 + S.cation * 16 + S.prot_cation * 32 + U * 64 + S.ligand * 128)
        hp_base = (np.array([r in o._HP_RESN for r in top.resname]) | S.ligand) & (top.element == "C")
        flags2 = (np.array([nm in o.BB_NAMES for nm in top.name]) * 1 + (top.resname == "CYS") * 2 + hp_base * 4 +
                  (top.element == "C") * 8 + (top.element == "H") * 16)
        chains = {c: k for k, c in enumerate(sorted(set(top.chain.tolist())))}
        labels = [top.reslabel(i) for i in range(n)]
        self.res_labels = sorted(set(labels))
        rank = {lab: k for k, lab in enumerate(self.res_labels)}
        reskey = np.array([rank[lab] for lab in labels], np.int32)
        start = np.zeros(n + 1, np.int32)
        adj = []
        for i in range(n):
            nb = sorted(top.neigh.get(i, ()))
            adj += nb
            start[i + 1] = len(adj)
        dis = [sorted(s) for s in S.disulfide]
        # pi-cation groups: ring atoms by residue label, ordered by label:index (pi_cation.py:83-101)
        ring_idx = np.nonzero(S.ring)[0].tolist()
        groups = {}
        for r in ring_idx:
            groups.setdefault(labels[r], []).append(r)
        rg_start, rg_atoms, rg_reskey = [0], [], []
        for lab in sorted(groups):
            rg_atoms += sorted(groups[lab], key=lambda i: top.label(i, True))
            rg_start.append(len(rg_atoms))
            rg_reskey.append(rank[lab])
        # aromatic residues for ps / ts: triplet by (chain, resname, resid) (aromatics.py:71-113), selected ring atoms
        rr_reskey, rr_trip, rr_rep, rr_sel_start, rr_sel = [], [], [], [0], []
        for lab in sorted(groups):
            first = groups[lab][0]
            trip = o._aromatic_triplet(S, first)
            rr_reskey.append(rank[lab])
            rr_trip += trip if len(trip) == 3 else [-1, -1, -1]
            rr_rep.append(o._ring_rep(top, trip) if trip else -1)
            rr_sel += [i for i in groups[lab] if selany[i]]
            rr_sel_start.append(len(rr_sel))
        t = GcTopo()
        t.n_atoms, t.n_reskeys, t.same, t.n_dis = n, len(self.res_labels), int(S.same), len(dis)
        t.n_ring_groups, t.n_ring_res = len(rg_reskey), len(rr_reskey)
        t.flags = arr("flags", flags, np.uint8).ctypes.data_as(u8p)
        t.flags2 = arr("flags2", flags2, np.uint8).ctypes.data_as(u8p)
        t.vdw = arr("vdw", top.vdw, np.float64).ctypes.data_as(f64p)
        for name, v in (("chain", [chains[c] for c in top.chain]), ("resid", top.resid), ("reskey", reskey), ("bond_start", start),
                        ("bond_adj", adj or [0]), ("dis", np.array(dis, np.int32).ravel() if dis else [0, 0]), ("rg_start", rg_start),
                        ("rg_atoms", rg_atoms or [0]), ("rg_reskey", rg_reskey or [0]), ("rr_reskey", rr_reskey or [0]),
                        ("rr_trip", rr_trip or [0]), ("rr_rep", rr_rep or [0]), ("rr_sel_start", rr_sel_start), ("rr_sel", rr_sel or [0])):
            setattr(t, name, arr(name, v, np.int32).ctypes.data_as(i32p))
        self._t, self._keep = t, keep
        self.n_pairs = len(self.res_labels) * (len(self.res_labels) + 1) // 2

    def _params(self, threads):
        g = self.geom
        p = GcParams()
        p.itypes = sum(ITYPE_BITS[x] for x in self.itypes)
        p.hb_rd, p.vdw_rd, p.fragment, p.threads = int(g["HBOND_RES_DIFF"]), int(g["VDW_RES_DIFF"]), o.TRAJ_FRAG_SIZE, int(threads or 0)
        p.sb_dist = g["SALT_BRIDGE_CUTOFF_DISTANCE"]
        p.pc_dist, p.pc_angle = g["PI_CATION_CUTOFF_DISTANCE"], g["PI_CATION_CUTOFF_ANGLE"]
        p.ps_dist, p.ps_angle, p.ps_psi = g["PI_STACK_CUTOFF_DISTANCE"], g["PI_STACK_CUTOFF_ANGLE"], g["PI_STACK_PSI_ANGLE"]
        p.ts_dist, p.ts_angle, p.ts_psi = g["T_STACK_CUTOFF_DISTANCE"], g["T_STACK_CUTOFF_ANGLE"], g["T_STACK_PSI_ANGLE"]
        p.hb_dist, p.hb_angle, p.eps = g["HBOND_CUTOFF_DISTANCE"], g["HBOND_CUTOFF_ANGLE"], g["VDW_EPSILON"]
        top, S = self.top, self.S
        p.soft = float(top.vdw[S.sele1].max()) + float(top.vdw[S.sele2].max()) + g["VDW_EPSILON"]      # vanderwaals.py:24-31
        return p

    def run(self, frames, want_contacts=True, want_counts=True, threads=None):
        """frames: float32 [F][N][3] (a whole number of fragments from frame 0).  Returns (contacts [(frame, itype, a, b)], counts)."""
        frames = np.ascontiguousarray(frames, dtype=np.float32)
        assert frames.ndim == 3 and frames.shape[1] == self.top.n
        lib = load()
        p = self._params(threads)
        counts = np.zeros(self.n_pairs, np.int32) if want_counts else None
        n_out = C.c_int64(0)
        cap = 0
        rec = None
        if want_contacts:
            cap = max(1024, int(len(frames) * self.top.n * 1.2))
            rec = np.empty((cap, 4), np.int32)
        rc = lib.gc_run(C.byref(self._t), C.byref(p), frames.ctypes.data, len(frames), counts.ctypes.data if want_counts else None,
                        rec.ctypes.data if want_contacts else None, cap, C.byref(n_out))
        if rc == -1:
            rec = np.empty((n_out.value, 4), np.int32)
            counts = np.zeros(self.n_pairs, np.int32) if want_counts else None
            rc = lib.gc_run(C.byref(self._t), C.byref(p), frames.ctypes.data, len(frames), counts.ctypes.data if want_counts else None,
                            rec.ctypes.data, len(rec), C.byref(n_out))
        assert rc == 0
        return (rec[:n_out.value] if want_contacts else None), counts

    def count_dict(self, counts):
        """{(res1, res2): frames} like getcontacts_oracle.intended_counts."""
        out = {}
        r = len(self.res_labels)
        idx = np.nonzero(counts)[0]
        lo = np.zeros(len(idx), np.int64)
        # invert idx = lo*r - lo*(lo-1)/2 + (hi-lo)
        starts = np.array([k * r - (k * (k - 1)) // 2 for k in range(r)], np.int64)
        lo = np.searchsorted(starts, idx, side="right") - 1
        hi = lo + (idx - starts[lo])
        for a, b, c in zip(lo.tolist(), hi.tolist(), counts[idx].tolist()):
            out[(self.res_labels[a], self.res_labels[b])] = c
        return out
