/* chacra_b200.h -- C ABI of the B200 contact-detection engine (libchacra_b200.so).
 *
 * The reference (Dan-Burns/ChACRA -> external/getcontacts) has no FFI: its hot path is Python that
 * drives VMD through Tcl strings.  The entry points below are therefore the surface a maintainer
 * binds with ctypes from the two reference functions that own the path:
 *
 *   chc_create            <- the one-off setup of compute_contacts()
 *                            (external/getcontacts/contact_calc/compute_contacts.py:226-238)
 *   chc_submit_frames*    <- compute_fragment_contacts() + compute_frame_contacts()
 *                            (compute_contacts.py:106-180 and :50-103; the per-itype detectors
 *                            salt_bridges.py:33-99, pi_cation.py:34-129, aromatics.py:116-257,
 *                            vanderwaals.py:34-96, hbonds.py:29-206, hydrophobics.py:44-123)
 *   chc_drain_contacts    <- the result queue feeding contact_consumer() (compute_contacts.py:306-359)
 *   chc_read_counts       <- res_contacts_xl()/gen_counts() of the frequency step
 *                            (external/getcontacts/contact_calc/transformations.py:70-150, :446-470;
 *                            driver external/getcontacts/get_contact_frequencies.py:58-134)
 *   chc_drain_near_cutoff <- new: the "within 1e-4 A / 1e-3 deg of a cutoff" side channel
 *   chc_write_contacts_tsv <- contact_consumer()'s text (compute_contacts.py:306-359), host threads
 *   chc_count_tsv         <- the contact-file reader of the frequency step (transformations.py:70-150), host threads
 *   chc_submit_frames_indexed, chc_dcd_unpack, chc_xtc_decode, chc_permuted_pca
 *                         <- the rows SURVEY 8(f) ranks next: replica -> state demultiplexing
 *                            (chacra/trajectories/process_hremd.py:184-307), trajectory input (VMD molfile plugins behind
 *                            contact_utils.py:89-116), ContactPCA's permutation test (chacra/ContactFrequencies.py:631-656)
 *
 * Conventions: plain C types only; 0 = success, negative = error (chc_last_error gives the text);
 * nothing throws or exits; the caller owns every buffer it passes in; the library owns what it
 * allocates until chc_destroy.  A context is bound to one CUDA device and is not thread-safe.
 * All device work is enqueued on the stream the caller passes (NULL = the legacy default stream).
 */
#ifndef CHACRA_B200_H
#define CHACRA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CHC_ABI_VERSION 1

/* interaction-type bits of chc_params.itype_mask (compute order of compute_contacts.py:85-101) */
#define CHC_IT_SB  (1u << 0)
#define CHC_IT_PC  (1u << 1)
#define CHC_IT_PS  (1u << 2)
#define CHC_IT_TS  (1u << 3)
#define CHC_IT_VDW (1u << 4)
#define CHC_IT_HB  (1u << 5)
#define CHC_IT_HP  (1u << 6)
#define CHC_IT_ALL 0x7Fu

/* chc_contact.itype codes; the text tokens are the ones of SURVEY.md Appendix B.4 */
enum chc_itype {
    CHC_SB = 0, CHC_SBLP, CHC_SBPL, CHC_SBLL, CHC_PC, CHC_PS, CHC_TS, CHC_VDW,
    CHC_HBBB, CHC_HBSB, CHC_HBSS, CHC_HBLL, CHC_HBLB, CHC_HBLS,
    CHC_WB, CHC_WB2, CHC_LWB, CHC_LWB2, CHC_HP, CHC_HPLP, CHC_HPLL, CHC_N_ITYPES
};

/* kinds of chc_near records */
enum chc_near_kind {
    CHC_NEAR_VDW_DIST = 0, CHC_NEAR_HP_DIST, CHC_NEAR_HB_ANGLE, CHC_NEAR_PC_DIST, CHC_NEAR_PC_ANGLE,
    CHC_NEAR_PS_DIST, CHC_NEAR_PS_ANGLE, CHC_NEAR_PS_PSI, CHC_NEAR_TS_DIST, CHC_NEAR_TS_ANGLE,
    CHC_NEAR_TS_PSI, CHC_NEAR_SB_DIST, CHC_NEAR_HB_DIST,
    /* soft `measure contacts` cutoffs of the ring detectors (pi_cation.py:22,70; aromatics.py:26-27,166) */
    CHC_NEAR_PC_SOFT, CHC_NEAR_PS_SOFT, CHC_NEAR_TS_SOFT, CHC_N_NEAR_KINDS
};

/* error codes */
#define CHC_OK              0
#define CHC_E_INVALID      -1   /* bad argument */
#define CHC_E_CUDA         -2   /* CUDA runtime error */
#define CHC_E_TOO_LARGE    -3   /* topology does not fit any kernel path */
#define CHC_E_ORDER        -4   /* frames submitted off a fragment boundary */
#define CHC_E_NOMEM        -5

/* One emitted contact (32 bytes).  atom1/atom2 (+ water1/water2 for bridges) are 0-based atom
 * indices of the topology; `frame` is the index the caller passed (first_index + k). */
typedef struct chc_contact {
    int32_t frame;
    int32_t itype;      /* enum chc_itype */
    int32_t atom1;
    int32_t atom2;
    int32_t water1;     /* -1 when unused */
    int32_t water2;     /* -1 when unused */
    float   distance;   /* float32 sqrtf(d2) between atom1 and atom2 (--distout) */
    int32_t reserved;
} chc_contact;

/* One near-cutoff observation (32 bytes): value - cutoff, Angstrom or degree */
typedef struct chc_near {
    int32_t frame;
    int32_t kind;       /* enum chc_near_kind */
    int32_t atom1;
    int32_t atom2;
    double  delta;
    int64_t reserved;
} chc_near;

/* Static tables produced by the host-side topology compiler (chacra_b200/topology.py).
 * All pointers are HOST pointers; chc_create copies them to the device. */
typedef struct chc_topology {
    int32_t n_atoms;          /* atoms per frame (every stored atom, hydrogens and solvent included) */
    int32_t n_pair;           /* NP: heavy atoms the cell-list pair search sees */
    int32_t n_aux;            /* NA: atoms copied verbatim next to the cell list (donor H, rings, cations) */
    int32_t n_h_slots;        /* the first n_h_slots aux entries are donor hydrogens (CSR pair_h_start) */
    int32_t n_rings;
    int32_t n_cations;
    int32_t n_disulfide;
    int32_t n_elem;           /* K <= 15 element classes among pair atoms */
    int32_t n_residues;       /* R distinct residue labels, ranked in string order */
    int32_t n_hp_base;
    int32_t n_bond_adj;
    int32_t same_selection;   /* 1 when the --sele and --sele2 strings are equal */
    int32_t n_pair_a;         /* pair atoms [0, n_pair_a) are searched against each other (grid A); the rest are
                                 unselected solvent heavy atoms, searched from the non-solvent side only (grid B) */
    int32_t n_solvent;        /* solvent heavy atoms among the pair atoms (water-bridge candidates) */
    const int32_t*  pair_atom;        /* [NP] atom index */
    const uint32_t* pair_info;        /* [NP] element class | flag bits (topology.py INFO_*) */
    const int32_t*  pair_lkey;        /* [NP] locality key: same chain & |dresid|<k <=> |dlkey|<k */
    const int32_t*  pair_res;         /* [NP] residue rank */
    const int32_t*  pair_resid;       /* [NP] resid (disulfide rule) */
    const int32_t*  pair_h_start;     /* [NP+1] CSR into aux slots [0, n_h_slots) */
    const int32_t*  pair_bond_start;  /* [NP+1] */
    const int32_t*  pair_bond_adj;    /* [n_bond_adj] bonded atom indices */
    const int32_t*  aux_atom;         /* [NA] */
    const int32_t*  ring_aux;         /* [NR][3] geometry triplet (aux slots) or -1 */
    const int32_t*  ring_pc_aux;      /* [NR][3] pi-cation triplet in label order or -1 */
    const int32_t*  ring_sel_aux;     /* [NR][3] ring atoms inside sele1|sele2, -1 padded */
    const uint8_t*  ring_sel_flags;   /* [NR][3] bit0 sele1, bit1 sele2 */
    const int32_t*  ring_atoms;       /* [NR][3] atom indices matching ring_sel_aux */
    const int32_t*  ring_res;         /* [NR] residue rank */
    const int32_t*  ring_rep;         /* [NR] representative atom (smallest name) */
    const int32_t*  ring_rep_aux;     /* [NR] its aux slot */
    const int32_t*  cation_aux;       /* [NC] */
    const int32_t*  cation_atom;      /* [NC] */
    const uint8_t*  cation_flags;     /* [NC] bit0 sele1, bit1 sele2 */
    const int32_t*  cation_res;       /* [NC] */
    const int32_t*  cation_excl_ring; /* [NC] ring holding an atom bonded to the cation, or -1 */
    const int32_t*  disulfide;        /* [ND][2] resid pairs */
    const float*    vdw_d2max;        /* [K][K] largest accepted float32 d2 (pure r1+r2+eps rule) */
    const double*   vdw_cut;          /* [K][K] r1+r2+eps */
    const int32_t*  hp_base;          /* [n_hp_base] pair indices of hydrophobic base carbons */
    const uint8_t*  atom_class;       /* [n_atoms] 1 = element C, 2 = element H, 0 = other */
    const uint8_t*  elem_class;       /* [K] the same code per element class of the pair atoms */
    const int32_t*  pair_solv_ord;    /* [NP] ordinal among the solvent pair atoms, -1 for the others */
} chc_topology;

typedef struct chc_params {
    uint32_t itype_mask;
    int32_t  n_states;
    int32_t  fragment_frames;     /* 100 = TRAJ_FRAG_SIZE (compute_contacts.py:43) */
    int32_t  emit_contacts;       /* 1: fill the contact-record buffer */
    int32_t  hbond_res_diff;
    int32_t  vdw_res_diff;
    int32_t  track_near;          /* 1: fill the near-cutoff buffer */
    int32_t  reserved0;
    int64_t  contact_capacity;    /* records */
    int64_t  near_capacity;       /* records */
    double sb_dist, pc_dist, pc_angle, ps_dist, ps_angle, ps_psi, ts_dist, ts_angle, ts_psi;
    double hb_dist, hb_angle, vdw_epsilon;
    double vdw_soft_cutoff;       /* max radius(sele1) + max radius(sele2) + epsilon (vanderwaals.py:24-31) */
    double hp_dist;               /* epsilon + 2*1.7 (hydrophobics.py:83) */
    double near_dist_tol;         /* 1e-4 */
    double near_angle_tol;        /* 1e-3 */
} chc_params;

typedef struct chc_stats {
    int64_t frames_submitted;
    int64_t kernel_launches;      /* binning + pair kernels launched so far */
    int64_t contacts_emitted;     /* including dropped ones */
    int64_t contacts_dropped;     /* beyond contact_capacity */
    int64_t near_total;
    int32_t path;                 /* 1 = frame staged in shared memory by bulk async copy, 2 = read in place (large systems) */
    int32_t smem_bytes;           /* pair kernel, dynamic shared memory per CTA */
    int32_t threads;              /* pair kernel */
    int32_t grid;                 /* pair kernel */
    int32_t bin_smem_bytes;       /* binning kernel */
    int32_t bin_grid;
    int64_t chunk_frames;         /* frames per (binning, pair) launch pair */
    int64_t record_bytes;         /* binned record per frame */
    double  pair_ms_total;        /* CUDA-event time of the pair kernels since chc_set_timing(1) */
    double  bin_ms_total;         /* ... of the binning kernels */
    int64_t timed_chunks;
    int32_t bin_variant;          /* binning kernel: 3 / 4 = bulk-copy (two CTAs per SM / double buffered), 1 / 2 = registers, 0 = gather */
    int32_t bin_apt;              /* pair atoms per thread of the bulk-copy binning kernel (template instantiation) */
    int32_t fused;                /* 1: chc_submit_frames runs the fused kernel (the pair kernel sorts every frame itself; the binning kernel
                                     only sees fragment-opening frames, for the hydrophobic masks); indexed submissions use the two-kernel path */
    int32_t fused_smem_bytes;     /* fused kernel, dynamic shared memory per CTA */
} chc_stats;

typedef struct chc_ctx chc_ctx;

int  chc_abi_version(void);
/* Name of this build of the CONTACT kernels (CHC_BUILD_TAG in chacra_b200.cu: bumped whenever the binning / pair kernels or their
 * launch plan change); profiles/ncu summaries quote it so that a measured DRAM-traffic figure is only ever attached to the build
 * of those kernels it was captured from. */
const char* chc_build_tag(void);
int  chc_create(chc_ctx** out, const chc_topology* topo, const chc_params* params, int device);
void chc_destroy(chc_ctx* ctx);
const char* chc_last_error(const chc_ctx* ctx);

/* Use a caller-owned device buffer int32 [n_states][n_pairs] for the count tables (so that the host
 * side can all-reduce it with NCCL).  n_elems must equal n_states * chc_n_pairs(). */
int     chc_bind_counts(chc_ctx* ctx, int32_t* d_counts, int64_t n_elems);
int64_t chc_n_pairs(const chc_ctx* ctx);     /* R*(R+1)/2, index = lo*R - lo*(lo-1)/2 + (hi-lo) */
int     chc_reset_counts(chc_ctx* ctx, void* stream);
int     chc_counts_device(chc_ctx* ctx, int32_t** d_counts, int64_t* n_pairs);
int     chc_read_counts(chc_ctx* ctx, int state, int32_t* h_counts, int64_t n_pairs, void* stream);

/* d_xyz: DEVICE float32 [n_frames][n_atoms][3] (Angstrom), frame k at d_xyz + k*frame_stride floats
 * (frame_stride 0 = 3*n_atoms).  first_index: position of the first frame in the caller's selected
 * frame sequence; a chunk must start on a multiple of fragment_frames or continue the previous chunk. */
int chc_submit_frames(chc_ctx* ctx, int state, const float* d_xyz, int64_t n_frames, int64_t first_index,
                      int64_t frame_stride, void* stream);
/* Same with HOST coordinates: the library stages them through pinned memory and copies them to
 * the device itself (double buffered). */
int chc_submit_frames_host(chc_ctx* ctx, int state, const float* h_xyz, int64_t n_frames, int64_t first_index,
                           void* stream);
/* Replica -> state demultiplexing without a state trajectory (replaces
 * get_state_coordinates_from_replica / write_state_trajectory, chacra/trajectories/process_hremd.py:184-230):
 * frame k of the submission is read from d_xyz + d_source_index[k]*frame_stride floats.  d_source_index is a
 * DEVICE array of n_frames entries and must stay valid until the work has run; everything else as above. */
int chc_submit_frames_indexed(chc_ctx* ctx, int state, const float* d_xyz, const int64_t* d_source_index,
                              int64_t n_frames, int64_t first_index, int64_t frame_stride, void* stream);
int chc_synchronize(chc_ctx* ctx, void* stream);

/* Copy out (and clear) the records accumulated so far.  *n = records copied. */
int chc_drain_contacts(chc_ctx* ctx, chc_contact* host_buf, int64_t cap, int64_t* n, void* stream);
int chc_drain_near_cutoff(chc_ctx* ctx, chc_near* host_buf, int64_t cap, int64_t* n, int64_t* n_total, void* stream);
/* ---- trajectory input: XTC frames decoded on the device -----------------------------------
 * Replaces VMD's gromacs molfile plugin on the path (contact_utils.py:89-116 load_traj -> molecule.read(.., 'xtc', ..))
 * for the files chacra/scripts/process_hremd_output.py:147 passes (state_{i}.xtc).  The host walks the frame headers
 * and fills one chc_xtc_frame per frame; the compressed bit streams travel to the device as they are in the file. */
typedef struct chc_xtc_frame {
    int64_t offset;          /* of the frame's bit stream in the blob, bytes, multiple of 4 */
    int32_t nbytes;          /* bytes of the bit stream (raw frames: 12 * n_atoms) */
    float   precision;       /* as stored (usually 1000: 0.001 nm grid) */
    int32_t minint[3], maxint[3];
    int32_t smallidx;
    int32_t raw;             /* 1: n_atoms <= 9, the frame holds big-endian float32 nm */
    int32_t bitsize;         /* bits of one packed "large" triple, 0 = three fields of bitsizeint bits (xdrfile sizeofints) */
    int32_t bitsizeint[3];
} chc_xtc_frame;

/* d_blob: DEVICE bytes (4-byte aligned, 128 readable bytes after the last stream); d_meta: DEVICE array of n_frames
 * entries; d_xyz: DEVICE float32 [n_frames][frame_stride] (frame_stride 0 = 3*n_atoms), Angstrom.  Synchronises the
 * stream; CHC_E_INVALID when a bit stream is inconsistent with n_atoms. */
int chc_xtc_decode(const void* d_blob, const chc_xtc_frame* d_meta, int64_t n_frames, int32_t n_atoms, float* d_xyz,
                   int64_t frame_stride, int device, void* stream);
/* The same with a caller-owned workspace (DEVICE, 16-byte aligned; no allocation inside): the decoder runs in two passes --
 * group boundaries of every frame by one thread per frame, then one thread per group -- and keeps 8 bytes per group (at most one
 * group per atom) between them.  chc_xtc_work_bytes(n_frames, n_atoms) is the size that takes all frames in one batch; a smaller
 * workspace (at least chc_xtc_work_bytes(1, n_atoms)) makes the call go through the frames in several batches, each of which
 * costs the serial boundary walk of one frame (about 1 ms for 6 500 atoms) again.  chc_xtc_decode allocates and frees such a
 * workspace itself (at most 2 GB). */
int     chc_xtc_decode_ws(const void* d_blob, const chc_xtc_frame* d_meta, int64_t n_frames, int32_t n_atoms, float* d_xyz,
                          int64_t frame_stride, void* d_work, int64_t work_bytes, int device, void* stream);
int64_t chc_xtc_work_bytes(int64_t n_frames, int32_t n_atoms);

/* Contact-list TSV writer (host code, no GPU work; replaces contact_consumer, gc/contact_calc/compute_contacts.py:306-359).
 * rec[n]: records from chc_drain_contacts covering WHOLE frames (reordered in place).  Labels: atom a's text is
 * label_text[label_off[a] .. label_off[a+1]) ("chain:resname:resid:name"), label_rank[a] its position in string order
 * (equal strings share a rank; ranks_unique = no two atoms share one).  itype_text / itype_off / block_rank: token and
 * block position of every chc_itype.  Output line: frame*stride+beg, itype, atom_1, atom_2[, water_1[, water_2]][, "%.3f"
 * distance], TAB separated; atoms ordered by label, identical lines once, frames ascending.  Writes to file descriptor
 * fd with up to n_threads formatting threads.  Returns the number of lines, or CHC_E_INVALID / CHC_E_IO. */
#define CHC_E_IO -6
int64_t chc_write_contacts_tsv(int fd, chc_contact* rec, int64_t n, const char* label_text, const int64_t* label_off,
                               const int32_t* label_rank, int32_t n_atoms, int32_t ranks_unique, const char* itype_text,
                               const int32_t* itype_off, const int32_t* block_rank, int32_t n_itypes, int64_t beg,
                               int64_t stride, int32_t distout, int32_t n_threads);
/* printf("%.3f") of a float32 distance as the writer formats it (exposed for tests); returns the length */
int chc_format_fixed3(float value, char* out32);

/* Contact-list TSV -> residue-pair frame counts (host code; replaces the line loop of get_contact_frequencies.py:58-134 /
 * contact_calc/transformations.py:70-150, 262-304, 446-470).  itypes_csv: comma-separated interaction types to count
 * (NULL or "" = all lines).  shipped != 0 reproduces the reference reader's behaviour (first line of every new frame
 * dropped, last frame never flushed); 0 counts every line.  Returns CHC_OK and a result handle, or
 * CHC_E_IO (cannot read), CHC_E_INVALID (malformed line; chc_tsv_counts_error gives the reference's message), or
 * CHC_E_ORDER (frame numbers not strictly increasing: not a file get_dynamic_contacts wrote; use a line-by-line reader). */
typedef struct chc_tsv_counts chc_tsv_counts;
int  chc_count_tsv(const char* path, const char* itypes_csv, int32_t shipped, int32_t n_threads, chc_tsv_counts** out);
/* sizes: residues, pairs, bytes of all residue names together, total_frames of the header */
void chc_tsv_counts_sizes(const chc_tsv_counts* r, int64_t* n_residues, int64_t* n_pairs, int64_t* name_bytes, int64_t* total_frames);
/* name_text[name_off[k] .. name_off[k+1]) = residue k; pair p joins res1[p] <= res2[p] (string order), count[p] frames */
void chc_tsv_counts_get(const chc_tsv_counts* r, char* name_text, int64_t* name_off, int32_t* res1, int32_t* res2, int64_t* count);
const char* chc_tsv_counts_error(const chc_tsv_counts* r);
void chc_tsv_counts_free(chc_tsv_counts* r);

/* Raw DCD frames -> interleaved coordinates (replaces VMD's dcd molfile plugin behind load_traj,
 * contact_calc/contact_utils.py:89-116).  d_raw: DEVICE copy of file bytes; the x record payload of frame k starts at
 * d_raw + first_x_offset + k*frame_pitch (4-byte aligned), y and z follow at +(4*n_atoms+8) bytes each, as in the file.
 * big_endian: the file's byte order differs from the device's.  d_xyz: DEVICE float32, frame k at k*frame_stride floats
 * (0 = 3*n_atoms).  Asynchronous on `stream`. */
int chc_dcd_unpack(const void* d_raw, int64_t first_x_offset, int64_t frame_pitch, int64_t n_frames, int32_t n_atoms,
                   int big_endian, float* d_xyz, int64_t frame_stride, int device, void* stream);

/* ContactPCA permutation test (replaces run_permuted_pca + de_correlate,
 * chacra/ContactFrequencies.py:631-656, called from ContactPCA.permuted_pca :910-964).
 * d_vals: DEVICE float64 [n_samples][n_features] row major (the contact-frequency table, states x contacts),
 * n_samples <= 64.  For each of n_perm permutations every column gets its rows reordered independently and the PCA
 * explained-variance ratios (min(n_samples, n_features) values, descending) go to d_evr [n_perm][n_comp] (DEVICE).
 * d_perms_in (DEVICE uint8 [n_perm][n_features][n_samples], entry i = the source row of row i; may be NULL): use
 * these row orders instead of generating them from `seed`.  d_perms_out (same shape, may be NULL): receives the
 * row orders used.  The call is asynchronous on `stream`; d_work (DEVICE, 8*n_samples*n_features bytes) is scratch. */
int chc_permuted_pca(const double* d_vals, int32_t n_samples, int32_t n_features, int32_t n_perm,
                     const uint8_t* d_perms_in, uint8_t* d_perms_out, uint64_t seed, double* d_work, double* d_evr,
                     int device, void* stream);

/* The same permutation test for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Identical to chc_permuted_pca except that the work table lives in a static device array inside the library
 * (CHC_PCA_STATIC_MAX_ELEMS float64 = 4 MiB, held by every GPU that loads the library until exit) instead of d_work; the
 * call performs no cudaMalloc / cudaFree.  n_samples * n_features > CHC_PCA_STATIC_MAX_ELEMS returns CHC_E_TOO_LARGE (and
 * says so on stderr).  All calls on a device share the one array: calls that overlap on the GPU (different streams or
 * threads) race; calls ordered on one stream are fine.  Results are bit-identical to chc_permuted_pca. */
#define CHC_PCA_STATIC_MAX_ELEMS (1 << 19)
int chc_permuted_pca_static(const double* d_vals, int32_t n_samples, int32_t n_features, int32_t n_perm,
                            const uint8_t* d_perms_in, uint8_t* d_perms_out, uint64_t seed, double* d_evr,
                            int device, void* stream);

int chc_get_stats(chc_ctx* ctx, chc_stats* out);
int chc_set_timing(chc_ctx* ctx, int on);

#ifdef __cplusplus
}
#endif
#endif
