// device_types.cuh -- device-side views of the topology tables, the cutoffs and the output buffers.
// Built with -fmad=false: every float multiply/add below rounds separately, which is what the
// float32 recipes of SURVEY.md Appendix A (VMD's measure contacts / hbonds / bond) require.
// Fused multiply-adds are used only where written explicitly (__fmaf_rn) for conservative prefilters.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/chacra_b200.h"

// bits of the per-pair-atom info word (mirror of chacra_b200/topology.py INFO_*)
#define I_ELEM   0xFu
#define I_S1     (1u << 4)
#define I_S2     (1u << 5)
#define I_HBU    (1u << 6)
#define I_HASH   (1u << 7)
#define I_ANION  (1u << 8)
#define I_CATION (1u << 9)
#define I_HP1    (1u << 10)   // static: base carbon in sele1; in the binned copy: member of H1 for this fragment
#define I_HP2    (1u << 11)   // static: base carbon in sele2; in the binned copy: member of H2 for this fragment
#define I_VDW    (1u << 12)
#define I_SOLV   (1u << 13)
// the bits below live only in the static table (looked up by pair index in the emit stage)
#define I_LIG    (1u << 14)
#define I_BB     (1u << 15)
#define I_CYS    (1u << 16)
#define I_XBOND  (1u << 17)
#define I_PBOND  (1u << 18)   // set by chc_create: bonded to a polar pair atom (hydrogen-bond selection, ion, solvent) and polar itself
#define I_PIDX_SHIFT 14        // the binned copy keeps the pair index in the bits above the dynamic flags
#define I_WMASK  ((1u << I_PIDX_SHIFT) - 1u)
#define MAX_PAIR_ATOMS (1 << 18)

// pair classes handed from the exact-distance stage to the emit stage
#define C_VDW   1u
#define C_HP    2u
#define C_SB    4u
#define C_HB    8u
#define C_NEARV 16u   // within near_dist_tol of the vdW cutoff
#define C_NEARH 32u   // ... of the hydrophobic (same) cutoff
#define C_HBW   64u   // hydrogen-bond candidate between a non-solvent atom and a solvent atom (water-bridge edge)
#define C_NEARS 128u  // ion pair within near_dist_tol of the salt-bridge cutoff
#define C_NEARB 256u  // donor / acceptor heavy pair (or water edge) within near_dist_tol of the hydrogen-bond distance cutoff

struct DevTopo {
    int n_atoms, n_pair, n_aux, n_h, n_rings, n_cat, n_dis, K, R, n_hp_base, same;
    int n_pair_a, n_solv;  // grid A pair atoms; solvent heavy atoms among all pair atoms
    const int* pair_solv_ord;
    const int* pair_atom; const unsigned* pair_info; const int* pair_lkey; const int* pair_res; const int* pair_resid;
    const int* pair_h_start; const int* pair_bond_start; const int* pair_bond_adj;
    const int* aux_atom; const unsigned short* aux_off16;   // float offsets 3 a + d of the aux coordinates (small frames)
    const int* ring_aux; const int* ring_pc_aux; const int* ring_sel_aux; const unsigned char* ring_sel_flags;
    const int* ring_atoms; const int* ring_res; const int* ring_rep; const int* ring_rep_aux;
    const int* cat_aux; const int* cat_atom; const unsigned char* cat_flags; const int* cat_res; const int* cat_excl;
    const int* disulfide;
    const float* tv;       // [K][K] min(vdw_d2max, soft^2)
    const float* th;       // [K][K] min(vdw_d2max, hp^2)
    const double* vcut;    // [K][K]
    const int* hp_base; const unsigned char* atom_class;
};

struct DevParams {
    unsigned itypes;
    unsigned elem_ch_mask; // bit e set when element class e is carbon or hydrogen (hydrophobics.py:24,39)
    int vdw_rd, hb_rd;
    int emit, track_near;
    float cell;            // cell edge (>= every pair cutoff, padded)
    float pre2;            // padded cmax^2 for the fused prefilter
    float sb2, hb2, hp2, soft2;   // float32(c)*float32(c)
    float hb_angle;        // float32 cutoff of measure hbonds
    float hb_tan_lo, hb_tan_hi; int hb_fast;   // tangent bracket of the cutoff (frame_phases.cuh hb_direction)
    float ring_pre2;       // conservative centroid prefilter for pc/ps/ts
    float ps_soft2, ts_soft2, pc_soft2;
    double ps_soft, ts_soft, pc_soft;
    double pc_dist, pc_angle, ps_dist, ps_angle, ps_psi, ts_dist, ts_angle, ts_psi;
    double sb_dist, hb_dist, hp_dist, soft;
    double near_d, near_a;
};

struct DevOut {
    int* counts;                 // [P] of the state being processed
    long long n_pairs;
    chc_contact* contacts; unsigned long long* contact_n; long long contact_cap;
    chc_near* near; unsigned long long* near_n; long long near_cap;
    unsigned* status;            // sticky error bits (ST_*), checked by the host
};
#define ST_WATER_DEGREE 1u       // a solvent atom had more than WADJ hydrogen-bonded non-solvent neighbours

// ---- binned frames: what the streaming/binning kernel hands to the pair kernel -----------------------
// One record per frame in a device buffer, every section 16-byte aligned:
//   BinHeader | spos float4[NP] (cell-sorted x, y, z, flags|pair_index<<14) | aux float[3*NA] | cells u16|int[ncell+1]
// `cells` are exclusive starts over a padded (nc+2)^3 grid (one guard layer), x fastest.
struct BinHeader {
    float org[3]; float inv;
    int nc[3]; int ncell;
};
struct BinLayout {
    int off_spos, off_aux, off_cells;   // bytes from the start of the record
    int fixed_bytes;                    // header + spos + aux (multiple of 16): one bulk copy
    int max_cells;                      // capacity of the cells section (entries incl. the end sentinel)
    int cell16;                         // cell starts are unsigned 16-bit (records the pair kernel keeps in shared memory)
    int count16;                        // gather binning variant: packed 16-bit cell counters in shared memory (n_pair < 65 536)
    long long stride;                   // bytes per record
};

__device__ __forceinline__ float dist2_exact(float ax, float ay, float az, float bx, float by, float bz) {
    // (dx*dx + dy*dy) + dz*dz, each operation rounded to binary32 (no contraction: -fmad=false)
    float dx = ax - bx, dy = ay - by, dz = az - bz;
    return (dx * dx + dy * dy) + dz * dz;
}

__device__ __forceinline__ long long pair_index(int ra, int rb, int R) {
    int lo = min(ra, rb), hi = max(ra, rb);
    return (long long)lo * R - ((long long)lo * (lo - 1)) / 2 + (hi - lo);
}
