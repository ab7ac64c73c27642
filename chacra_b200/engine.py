"""ContactEngine: the host-side handle on one ``chc_ctx`` (one CUDA device, one topology).

Python keeps only buffers and bookkeeping here (torch tensors for device memory, numpy for the
record arrays); every frame is processed by the CUDA kernels behind the C ABI.  Mirrors the work
split of the reference: setup once (``compute_contacts.py:226-238``), then frames in 100-frame
fragments (``compute_contacts.py:250-261``).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .topology import CompiledTopology, vdw_threshold_tables

# argparsers.py:80-125 defaults of get_dynamic_contacts (keys of get_dynamic_contacts.py:24-39)
DEFAULT_GEOMETRY = {
    "SALT_BRIDGE_CUTOFF_DISTANCE": 4.0, "PI_CATION_CUTOFF_DISTANCE": 6.0, "PI_CATION_CUTOFF_ANGLE": 60,
    "PI_STACK_CUTOFF_DISTANCE": 7.0, "PI_STACK_CUTOFF_ANGLE": 30, "PI_STACK_PSI_ANGLE": 45,
    "T_STACK_CUTOFF_DISTANCE": 5.0, "T_STACK_CUTOFF_ANGLE": 30, "T_STACK_PSI_ANGLE": 45,
    "HBOND_CUTOFF_DISTANCE": 3.5, "HBOND_CUTOFF_ANGLE": 70, "HBOND_RES_DIFF": 1,
    "VDW_EPSILON": 0.5, "VDW_RES_DIFF": 2}
ITYPE_BITS = {"sb": 1, "pc": 2, "ps": 4, "ts": 8, "vdw": 16, "hb": 32, "hp": 64}
COMPUTE_ORDER = ["sb", "pc", "ps", "ts", "vdw", "hb", "hp"]          # compute_contacts.py:85-101
ITYPE_NAMES = ["sb", "sblp", "sbpl", "sbll", "pc", "ps", "ts", "vdw", "hbbb", "hbsb", "hbss", "hbll", "hblb", "hbls",
               "wb", "wb2", "lwb", "lwb2", "hp", "hplp", "hpll"]
NEAR_KINDS = ["vdw_dist", "hp_dist", "hb_angle", "pc_dist", "pc_angle", "ps_dist", "ps_angle", "ps_psi", "ts_dist",
              "ts_angle", "ts_psi", "sb_dist", "hb_dist", "pc_soft", "ps_soft", "ts_soft"]
TRAJ_FRAG_SIZE = 100                                                  # compute_contacts.py:43

CONTACT_DTYPE = np.dtype([("frame", "<i4"), ("itype", "<i4"), ("atom1", "<i4"), ("atom2", "<i4"), ("water1", "<i4"),
                          ("water2", "<i4"), ("distance", "<f4"), ("reserved", "<i4")])
NEAR_DTYPE = np.dtype([("frame", "<i4"), ("kind", "<i4"), ("atom1", "<i4"), ("atom2", "<i4"), ("delta", "<f8"),
                       ("reserved", "<i8")])


class ChacraError(RuntimeError):
    pass


def _ptr(arr, ctype):
    return arr.ctypes.data_as(C.POINTER(ctype))


class ContactEngine:
    def __init__(self, topo: CompiledTopology, itypes=("sb", "pc", "ps", "ts", "vdw", "hb", "hp"), geometry=None,
                 n_states: int = 1, emit_contacts: bool = False, track_near: bool = False,
                 contact_capacity: int = 1 << 22, near_capacity: int = 1 << 16, device: int = 0,
                 fragment_frames: int = TRAJ_FRAG_SIZE):
        import torch  # device buffers + streams only
        if not torch.cuda.is_available():
            raise ChacraError("chacra_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self._torch = torch
        self.lib = _lib.load()
        self.topo = topo
        self.device = int(device)
        self.n_states = int(n_states)
        g = dict(DEFAULT_GEOMETRY)
        g.update(geometry or {})
        self.geometry = g
        self.itypes = [t for t in COMPUTE_ORDER if t in set(itypes)]
        unknown = set(itypes) - set(COMPUTE_ORDER)
        if unknown:
            raise ChacraError("unsupported interaction types: %s" % sorted(unknown))
        a = topo.atoms
        eps = float(g["VDW_EPSILON"])
        soft = float(a.vdw_radius[topo.sele1].max()) + float(a.vdw_radius[topo.sele2].max()) + eps   # vanderwaals.py:24-31
        d2max, cut = vdw_threshold_tables(topo.elem_radius, eps, 1e9)
        # keep every array alive for the duration of chc_create
        keep = {}

        def arr(name, value, dtype):
            keep[name] = np.ascontiguousarray(value, dtype=dtype)
            return keep[name]

        t = _lib.ChcTopology()
        t.n_atoms, t.n_pair, t.n_aux, t.n_h_slots = a.n, len(topo.pair_atom), len(topo.aux_atom), topo.n_h_slots
        t.n_rings, t.n_cations, t.n_disulfide = len(topo.ring_res), len(topo.cation_atom), len(topo.disulfide)
        t.n_elem, t.n_residues, t.n_hp_base = len(topo.elem_symbols), topo.n_residues, len(topo.hp_base_atoms)
        t.n_bond_adj, t.same_selection = len(topo.pair_bond_adj), int(topo.same_selection)
        t.n_pair_a, t.n_solvent = int(topo.n_pair_a), int(topo.n_solvent_pair)
        for name, ct in (("pair_atom", C.c_int32), ("pair_lkey", C.c_int32), ("pair_res", C.c_int32),
                         ("pair_resid", C.c_int32), ("pair_h_start", C.c_int32), ("pair_bond_start", C.c_int32),
                         ("pair_bond_adj", C.c_int32), ("aux_atom", C.c_int32), ("ring_aux", C.c_int32),
                         ("ring_pc_aux", C.c_int32), ("ring_sel_aux", C.c_int32), ("ring_atoms", C.c_int32),
                         ("ring_res", C.c_int32), ("ring_rep", C.c_int32), ("ring_rep_aux", C.c_int32),
                         ("cation_aux", C.c_int32), ("cation_atom", C.c_int32), ("cation_res", C.c_int32),
                         ("cation_excl_ring", C.c_int32), ("disulfide", C.c_int32), ("pair_solv_ord", C.c_int32)):
            setattr(t, name, _ptr(arr(name, getattr(topo, name), np.int32), ct))
        t.pair_info = _ptr(arr("pair_info", topo.pair_info, np.uint32), C.c_uint32)
        t.ring_sel_flags = _ptr(arr("ring_sel_flags", topo.ring_sel_flags, np.uint8), C.c_uint8)
        t.cation_flags = _ptr(arr("cation_flags", topo.cation_flags, np.uint8), C.c_uint8)
        t.vdw_d2max = _ptr(arr("d2max", d2max, np.float32), C.c_float)
        t.vdw_cut = _ptr(arr("cut", cut, np.float64), C.c_double)
        t.hp_base = _ptr(arr("hp_base", topo.hp_base_atoms, np.int32), C.c_int32)
        t.atom_class = _ptr(arr("atom_class", topo.elem_is_c, np.uint8), C.c_uint8)
        t.elem_class = _ptr(arr("elem_class", topo.elem_class, np.uint8), C.c_uint8)

        p = _lib.ChcParams()
        p.itype_mask = sum(ITYPE_BITS[x] for x in self.itypes)
        p.n_states, p.fragment_frames = self.n_states, int(fragment_frames)
        p.emit_contacts, p.track_near = int(emit_contacts), int(track_near)
        p.hbond_res_diff, p.vdw_res_diff = int(g["HBOND_RES_DIFF"]), int(g["VDW_RES_DIFF"])
        p.contact_capacity = int(contact_capacity) if emit_contacts else 1
        p.near_capacity = int(near_capacity) if track_near else 1
        p.sb_dist = g["SALT_BRIDGE_CUTOFF_DISTANCE"]
        p.pc_dist, p.pc_angle = g["PI_CATION_CUTOFF_DISTANCE"], g["PI_CATION_CUTOFF_ANGLE"]
        p.ps_dist, p.ps_angle, p.ps_psi = g["PI_STACK_CUTOFF_DISTANCE"], g["PI_STACK_CUTOFF_ANGLE"], g["PI_STACK_PSI_ANGLE"]
        p.ts_dist, p.ts_angle, p.ts_psi = g["T_STACK_CUTOFF_DISTANCE"], g["T_STACK_CUTOFF_ANGLE"], g["T_STACK_PSI_ANGLE"]
        p.hb_dist, p.hb_angle = g["HBOND_CUTOFF_DISTANCE"], g["HBOND_CUTOFF_ANGLE"]
        p.vdw_epsilon, p.vdw_soft_cutoff, p.hp_dist = eps, soft, eps + 2 * 1.7                 # hydrophobics.py:83
        p.near_dist_tol, p.near_angle_tol = 1e-4, 1e-3
        self.emit_contacts, self.track_near = bool(emit_contacts), bool(track_near)
        self.contact_capacity, self.near_capacity = int(p.contact_capacity), int(p.near_capacity)
        self.fragment_frames = int(fragment_frames)

        self._ctx = C.c_void_p()
        rc = self.lib.chc_create(C.byref(self._ctx), C.byref(t), C.byref(p), self.device)
        if rc != 0:
            msg = self.lib.chc_last_error(self._ctx).decode() if self._ctx else "allocation failed"
            if self._ctx:
                self.lib.chc_destroy(self._ctx)
                self._ctx = C.c_void_p()
            raise ChacraError("chc_create failed (%d): %s" % (rc, msg))
        self.n_pairs = int(self.lib.chc_n_pairs(self._ctx))
        with torch.cuda.device(self.device):
            self.counts_tensor = torch.zeros((self.n_states, self.n_pairs), dtype=torch.int32, device="cuda:%d" % self.device)
        self._check(self.lib.chc_bind_counts(self._ctx, self.counts_tensor.data_ptr(), self.counts_tensor.numel()))

    # ------------------------------------------------------------------------------------------
    def _check(self, rc):
        if rc != 0:
            raise ChacraError("libchacra_b200 error %d: %s" % (rc, self.lib.chc_last_error(self._ctx).decode()))

    def _stream(self):
        return C.c_void_p(self._torch.cuda.current_stream(self.device).cuda_stream)

    def submit(self, xyz, state: int = 0, first_index: int = 0):
        """Process frames ``xyz`` ([F][N][3] float32; torch CUDA tensor or host numpy / pinned torch tensor)."""
        torch = self._torch
        n = self.topo.n_atoms
        if isinstance(xyz, torch.Tensor) and xyz.is_cuda:
            if xyz.dtype != torch.float32 or xyz.dim() != 3 or xyz.shape[1] != n or xyz.shape[2] != 3 or not xyz.is_contiguous():
                raise ChacraError("device frames must be a contiguous float32 [F][%d][3] tensor" % n)
            if xyz.device.index != self.device:
                raise ChacraError("frames live on cuda:%d, engine on cuda:%d" % (xyz.device.index, self.device))
            self._check(self.lib.chc_submit_frames(self._ctx, state, xyz.data_ptr(), xyz.shape[0], first_index, 0, self._stream()))
            return
        if isinstance(xyz, torch.Tensor):
            if xyz.dtype != torch.float32 or not xyz.is_contiguous() or tuple(xyz.shape[1:]) != (n, 3):
                raise ChacraError("host frames must be a contiguous float32 [F][%d][3] tensor" % n)
            ptr, nf = xyz.data_ptr(), xyz.shape[0]
        else:
            xyz = np.ascontiguousarray(xyz, dtype=np.float32)
            if xyz.ndim != 3 or xyz.shape[1:] != (n, 3):
                raise ChacraError("host frames must have shape [F][%d][3]" % n)
            ptr, nf = xyz.ctypes.data, xyz.shape[0]
        self._check(self.lib.chc_submit_frames_host(self._ctx, state, ptr, nf, first_index, self._stream()))
        # asynchronous H2D copies read the caller's (pinned) buffer after this call returns: every submitted host
        # array stays referenced until synchronize() / the next blocking read
        self._keepalive = getattr(self, "_keepalive", [])
        if not any(k is xyz for k in self._keepalive):
            self._keepalive.append(xyz)

    def submit_indexed(self, pool, source_index, state: int = 0, first_index: int = 0):
        """Frame k of the submission is ``pool[source_index[k]]`` (replica -> state demultiplexing).
        ``pool``: contiguous CUDA float32 [M][N][3]; ``source_index``: int64, CUDA tensor or host array."""
        torch = self._torch
        n = self.topo.n_atoms
        if not (isinstance(pool, torch.Tensor) and pool.is_cuda and pool.dtype == torch.float32 and pool.dim() == 3
                and tuple(pool.shape[1:]) == (n, 3) and pool.is_contiguous() and pool.device.index == self.device):
            raise ChacraError("frame pool must be a contiguous float32 [M][%d][3] tensor on cuda:%d" % (n, self.device))
        idx = source_index if isinstance(source_index, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(source_index, dtype=np.int64))
        idx = idx.to(device=pool.device, dtype=torch.int64).contiguous()
        if idx.dim() != 1:
            raise ChacraError("source_index must be one-dimensional")
        if idx.numel() and (int(idx.min()) < 0 or int(idx.max()) >= pool.shape[0]):
            raise ChacraError("source_index outside the pool of %d frames" % pool.shape[0])
        self._check(self.lib.chc_submit_frames_indexed(self._ctx, state, pool.data_ptr(), idx.data_ptr(), idx.numel(), first_index, 0, self._stream()))
        self._keepalive_idx = getattr(self, "_keepalive_idx", [])[-63:] + [idx]      # read by kernels still in the stream

    def synchronize(self):
        self._check(self.lib.chc_synchronize(self._ctx, self._stream()))
        self._keepalive = []

    def counts(self, state: int = 0) -> np.ndarray:
        """int32 [P] residue-pair frame counts of one state (device -> host)."""
        out = np.empty(self.n_pairs, np.int32)
        self._check(self.lib.chc_read_counts(self._ctx, state, out.ctypes.data, self.n_pairs, self._stream()))
        return out

    def reset_counts(self):
        self._check(self.lib.chc_reset_counts(self._ctx, self._stream()))

    def drain_contacts(self, copy: bool = True) -> np.ndarray:
        """Records emitted since the last drain.  They land in a pinned host buffer the engine keeps; ``copy=False``
        returns a view of it that is valid until the next drain (the streaming TSV path uses that)."""
        if getattr(self, "_drain_buf", None) is None:
            pinned = self._torch.empty(self.contact_capacity * CONTACT_DTYPE.itemsize, dtype=self._torch.uint8, pin_memory=True)
            self._drain_pin = pinned
            self._drain_buf = pinned.numpy().view(CONTACT_DTYPE)
        buf = self._drain_buf
        n = C.c_int64(0)
        self._check(self.lib.chc_drain_contacts(self._ctx, buf.ctypes.data, self.contact_capacity, C.byref(n), self._stream()))
        st = self.stats()
        if st["contacts_dropped"]:
            raise ChacraError("contact buffer overflow: %d records dropped; raise contact_capacity or submit fewer frames "
                              "between drains" % st["contacts_dropped"])
        return buf[:n.value].copy() if copy else buf[:n.value]

    def drain_near(self):
        buf = np.empty(self.near_capacity, NEAR_DTYPE)
        n, total = C.c_int64(0), C.c_int64(0)
        self._check(self.lib.chc_drain_near_cutoff(self._ctx, buf.ctypes.data, self.near_capacity, C.byref(n), C.byref(total), self._stream()))
        return buf[:n.value].copy(), int(total.value)

    def set_timing(self, on: bool):
        self._check(self.lib.chc_set_timing(self._ctx, int(on)))

    def stats(self) -> dict:
        s = _lib.ChcStats()
        self._check(self.lib.chc_get_stats(self._ctx, C.byref(s)))
        return {k: getattr(s, k) for k, _ in s._fields_}

    def pair_index(self, ra, rb):
        """Index of residue pair (ra, rb) (ranks into ``topo.residue_label``) in a count table."""
        lo, hi = np.minimum(ra, rb).astype(np.int64), np.maximum(ra, rb).astype(np.int64)
        r = self.topo.n_residues
        return lo * r - (lo * (lo - 1)) // 2 + (hi - lo)

    def close(self):
        if getattr(self, "_ctx", None):
            self.lib.chc_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
