"""ctypes binding of ``libchacra_b200.so`` (C ABI declared in ``include/chacra_b200.h``).

There is no CPU fallback: if the shared library has not been built the import of the engine fails
loudly (``ChacraLibraryError``), and ``chc_create`` refuses any device that is not sm_100.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.environ.get("CHACRA_LIB") or os.path.join(CSRC, "libchacra_b200.so")    # CHACRA_LIB: A/B runs of build variants
SOURCES = ["chacra_b200.cu"]
HEADERS = ["device_types.cuh", "bin_tma_kernel.cuh", "xtc_decode.cuh", "pca_perm.cuh", "dcd_unpack.cuh", "tsv_writer.h", "tsv_counts.h", "frame_kernel.cuh", "frame_phases.cuh", "bin_kernel.cuh", "pair_kernel.cuh", "water_phase.cuh",
           os.path.join("..", "..", "include", "chacra_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC"]


class ChacraLibraryError(RuntimeError):
    pass


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile the CUDA sources in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, s) for s in SOURCES]
    deps = srcs + [os.path.join(CSRC, h) for h in HEADERS]
    if not force and os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(d) for d in deps):
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH] + srcs
    res = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if res.returncode != 0:
        raise ChacraLibraryError("nvcc failed:\n%s\n%s" % (" ".join(cmd), res.stderr))
    if verbose:
        sys.stderr.write(res.stderr)
    return LIB_PATH


i32p, u32p, u8p = C.POINTER(C.c_int32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint8)
f32p, f64p = C.POINTER(C.c_float), C.POINTER(C.c_double)


class ChcTopology(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "n_atoms", "n_pair", "n_aux", "n_h_slots", "n_rings", "n_cations", "n_disulfide", "n_elem", "n_residues",
        "n_hp_base", "n_bond_adj", "same_selection", "n_pair_a", "n_solvent")] + [
        ("pair_atom", i32p), ("pair_info", u32p), ("pair_lkey", i32p), ("pair_res", i32p), ("pair_resid", i32p),
        ("pair_h_start", i32p), ("pair_bond_start", i32p), ("pair_bond_adj", i32p), ("aux_atom", i32p),
        ("ring_aux", i32p), ("ring_pc_aux", i32p), ("ring_sel_aux", i32p), ("ring_sel_flags", u8p),
        ("ring_atoms", i32p), ("ring_res", i32p), ("ring_rep", i32p), ("ring_rep_aux", i32p),
        ("cation_aux", i32p), ("cation_atom", i32p), ("cation_flags", u8p), ("cation_res", i32p),
        ("cation_excl_ring", i32p), ("disulfide", i32p), ("vdw_d2max", f32p), ("vdw_cut", f64p), ("hp_base", i32p),
        ("atom_class", u8p), ("elem_class", u8p), ("pair_solv_ord", i32p)]


class ChcParams(C.Structure):
    _fields_ = [("itype_mask", C.c_uint32), ("n_states", C.c_int32), ("fragment_frames", C.c_int32),
                ("emit_contacts", C.c_int32), ("hbond_res_diff", C.c_int32), ("vdw_res_diff", C.c_int32),
                ("track_near", C.c_int32), ("reserved0", C.c_int32), ("contact_capacity", C.c_int64),
                ("near_capacity", C.c_int64)] + [(n, C.c_double) for n in (
                    "sb_dist", "pc_dist", "pc_angle", "ps_dist", "ps_angle", "ps_psi", "ts_dist", "ts_angle", "ts_psi",
                    "hb_dist", "hb_angle", "vdw_epsilon", "vdw_soft_cutoff", "hp_dist", "near_dist_tol",
                    "near_angle_tol")]


class ChcStats(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("frames_submitted", "kernel_launches", "contacts_emitted",
                                         "contacts_dropped", "near_total")] + \
               [(n, C.c_int32) for n in ("path", "smem_bytes", "threads", "grid", "bin_smem_bytes", "bin_grid")] + \
               [("chunk_frames", C.c_int64), ("record_bytes", C.c_int64), ("pair_ms_total", C.c_double),
                ("bin_ms_total", C.c_double), ("timed_chunks", C.c_int64), ("bin_variant", C.c_int32), ("bin_apt", C.c_int32), ("fused", C.c_int32), ("fused_smem_bytes", C.c_int32)]


# every symbol include/chacra_b200.h declares: name -> (restype, argtypes)
vp = C.c_void_p
SYMBOLS = {
    "chc_abi_version": (C.c_int, []),
    "chc_build_tag": (C.c_char_p, []),
    "chc_create": (C.c_int, [C.POINTER(vp), C.POINTER(ChcTopology), C.POINTER(ChcParams), C.c_int]),
    "chc_destroy": (None, [vp]),
    "chc_last_error": (C.c_char_p, [vp]),
    "chc_bind_counts": (C.c_int, [vp, vp, C.c_int64]),
    "chc_n_pairs": (C.c_int64, [vp]),
    "chc_reset_counts": (C.c_int, [vp, vp]),
    "chc_counts_device": (C.c_int, [vp, C.POINTER(vp), C.POINTER(C.c_int64)]),
    "chc_read_counts": (C.c_int, [vp, C.c_int, vp, C.c_int64, vp]),
    "chc_submit_frames": (C.c_int, [vp, C.c_int, vp, C.c_int64, C.c_int64, C.c_int64, vp]),
    "chc_submit_frames_host": (C.c_int, [vp, C.c_int, vp, C.c_int64, C.c_int64, vp]),
    "chc_submit_frames_indexed": (C.c_int, [vp, C.c_int, vp, vp, C.c_int64, C.c_int64, C.c_int64, vp]),
    "chc_synchronize": (C.c_int, [vp, vp]),
    "chc_drain_contacts": (C.c_int, [vp, vp, C.c_int64, C.POINTER(C.c_int64), vp]),
    "chc_drain_near_cutoff": (C.c_int, [vp, vp, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64), vp]),
    "chc_get_stats": (C.c_int, [vp, C.POINTER(ChcStats)]),
    "chc_set_timing": (C.c_int, [vp, C.c_int]),
    "chc_xtc_decode": (C.c_int, [vp, vp, C.c_int64, C.c_int32, vp, C.c_int64, C.c_int, vp]),
    "chc_xtc_decode_ws": (C.c_int, [vp, vp, C.c_int64, C.c_int32, vp, C.c_int64, vp, C.c_int64, C.c_int, vp]),
    "chc_xtc_work_bytes": (C.c_int64, [C.c_int64, C.c_int32]),
    "chc_write_contacts_tsv": (C.c_int64, [C.c_int, vp, C.c_int64, vp, vp, vp, C.c_int32, C.c_int32, vp, vp, vp, C.c_int32, C.c_int64,
                                         C.c_int64, C.c_int32, C.c_int32]),
    "chc_format_fixed3": (C.c_int, [C.c_float, vp]),
    "chc_count_tsv": (C.c_int, [C.c_char_p, C.c_char_p, C.c_int32, C.c_int32, C.POINTER(vp)]),
    "chc_tsv_counts_sizes": (None, [vp, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "chc_tsv_counts_get": (None, [vp, vp, vp, vp, vp, vp]),
    "chc_tsv_counts_error": (C.c_char_p, [vp]),
    "chc_tsv_counts_free": (None, [vp]),
    "chc_dcd_unpack": (C.c_int, [vp, C.c_int64, C.c_int64, C.c_int64, C.c_int32, C.c_int, vp, C.c_int64, C.c_int, vp]),
    "chc_permuted_pca": (C.c_int, [vp, C.c_int32, C.c_int32, C.c_int32, vp, vp, C.c_uint64, vp, vp, C.c_int, vp]),
    "chc_permuted_pca_static": (C.c_int, [vp, C.c_int32, C.c_int32, C.c_int32, vp, vp, C.c_uint64, vp, C.c_int, vp]),
}

_lib = None


def load() -> C.CDLL:
    """dlopen the in-tree library and bind every declared symbol; raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ChacraLibraryError(
            "%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). chacra_b200 has no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError here = header/library mismatch
        fn.restype, fn.argtypes = res, args
    if lib.chc_abi_version() != 1:
        raise ChacraLibraryError("ABI version mismatch")
    _lib = lib
    return lib
