// pair_kernel.cuh -- every interaction type of one frame from its binned record (bin_kernel.cuh).
//
// Persistent CTAs, one frame at a time per CTA.  SMALL = the record fits in shared memory: one bulk async copy
// (cp.async.bulk + mbarrier) brings header | spos | aux, a second one the cell starts; the per-frame residue-pair
// set is a plain shared-memory bitset.  !SMALL = everything is read in place from global memory / L2 and the
// residue-pair set is a generation-stamped bitset private to the CTA (frame_kernel.cuh count_pair).
//
// Pair phase, per warp (tiles of 32 cell-sorted atoms are handed out through a shared counter):
//   E1  lane = atom i of the tile; every lane walks its own half-shell ranges (five runs of three x-adjacent cells,
//       contiguous in the sorted array) 32 slots at a time, exact float32 d2 against the largest cutoff, hits kept as a
//       per-lane bit mask (no warp votes in the inner loop); the masks are expanded into the warp's queue afterwards
//   E2  32 survivors at a time: class tests (selection, locality, vdW/hp thresholds, ion and hbond classes)
//   E3  32 classified pairs at a time: bonded / disulfide / hydrogen-angle tests, residue-pair set, counts, records
// The three stages form one resumable loop so that E2 and E3 are emitted exactly once in the SASS (the first
// version inlined them at ~45 sites and stalled on instruction fetch, profiles/r01_v1_*).
#pragma once
#include <type_traits>
#include "frame_phases.cuh"
#include "water_phase.cuh"

#ifndef QE_CAP
#define QE_CAP 64           // E1 -> E2 entries per warp: E1 runs while fewer than 32 are queued and adds at most 32
#endif
#ifndef Q2_CAP
#define Q2_CAP 64           // classified pairs per warp and queue (plain, polar)
#endif
#ifndef E1_TARGET
#define E1_TARGET 32        // E1 runs until this many survivors are queued (one F batch)
#endif
static_assert(QE_CAP - E1_TARGET >= 32 && E1_TARGET >= 32, "one entry per lane must fit behind E1_TARGET queued entries");
#ifndef E1_UNROLL
#define E1_UNROLL 4
#endif
#define CHC_PRAGMA(x) _Pragma(#x)
#define CHC_UNROLL(n) CHC_PRAGMA(unroll n)

struct PairLayout {         // shared-memory plan of the pair kernel (bytes)
    int off_rec;            // SMALL: header | spos | aux  (B.fixed_bytes)
    int off_cells;          // SMALL: cell starts
    int off_lkey;           // SMALL: locality keys by pair index
    int off_bits;           // SMALL: residue-pair bitset
    int off_q1, off_qf, off_q2, off_tv, off_th, off_misc;
    // SMALL: the sorted atoms as slot pairs (x0 x1 y0 y1 | z0 z1), and per slot the flag word, the filter key, the residue rank.
    // The float4 image the bulk copy lands (inside off_rec) is dead once these are built: the queues live there.
    int off_exy, off_ez, off_ek, off_sw, off_sres;      // off_ez: float4 z0 z1 lkey0 lkey1 per slot pair; off_ek: partner bit arrays
    int off_plkf, off_pres, off_pst;                    // SMALL, by pair index (loaded once per launch): locality key (float), residue rank, static bits
    int off_rc, off_rlist, rlist_cap;                   // ring centroids (float[3 n_rings]); ring tasks that passed the prefilter
    int off_rgrid, off_rsort;                           // !SMALL: coarse grid over the ring centroids (RG_CAP + 1 counters), ring ids sorted by its cells; 0 = none
    int bits_plain;                                     // !SMALL: the residue-pair set is a plain bitset in shared memory (off_bits) as on the small path
    int off_aux, off_cnt;                               // FUSED: aux atoms copied out of the frame; transient cell counters
    long long soa_stride;                               // !SMALL: bytes of one CTA's slot-pair arrays in the global scratch (off_exy .. count from its start)                   // ring centroids (float[3 n_rings]) and the ring tasks that passed the prefilter
    int bits_words;         // words of the residue-pair set (plain: P/32, stamped: P/24)
    int total;
};

// ---- packed float32 pairs (Blackwell FADD2 / FMUL2 / FFMA2): two neighbour slots per instruction in the E1 prefilter ----
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) { f32x2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

// The cheap SUFFICIENT test for "no class" (most close pairs sit in the same or adjacent residues and only matter when both atoms
// are polar) runs inside the E1 loop: the locality key rides along as a float next to z, sign(dl^2 - (vdw_res_diff - 1/2)^2) marks a LOCAL neighbour, and
// "polar partner" comes from a bit per slot: partner set c of a lane = (c & 1 ? hydrogen-bond selection : 0) | (c & 2 ? cations : 0)
// | (c & 4 ? anions : 0) | solvent, set 8 = everything but solvent (for a solvent atom): 9 bit arrays over the sorted slots.
#define E1_PAD_PAIRS 20        // slot pairs past the last atom that the unmasked E1 loads may touch
#define E1_PARTNER_SETS 9
#ifndef E1_TAIL_TILES
#define E1_TAIL_TILES 16       // the last tiles of a frame are handed out range by range
#endif
__device__ __forceinline__ int e1_partner_set(unsigned w) {
    return (w & I_SOLV) ? 8 : (int)(((w >> 6) & 1u) | (((w >> 8) & 1u) << 1) | (((w >> 9) & 1u) << 2));   // hbu | anion -> wants cations | cation -> wants anions
}

#define RG_CAP 1024            // cells of the coarse ring grid (large path; shared memory there is taken from the L1 the slot arrays live in)
struct PairMisc {
    unsigned long long bar;     // mbarrier of the bulk copies
    int cursor;                 // next tile of 32 sorted atoms
    int ring_cursor;            // next unit of 32 ring tasks
    int nmark;                  // waters with at least one hydrogen bond to a non-solvent atom (this frame)
    int qcur[32];               // SMALL: entries in every warp's E2 queue (lanes reserve room with one shared atomic each)
    int rn_pc, rn_st;           // ring tasks that passed the centroid prefilter: pi-cation from the front of the list, stacking from its end
    // FUSED (the CTA sorts its own frame): bounding box as ordered ints, warp totals of the scan, the grid of this frame
    int bb_min[2][3], bb_max[2][3];
    int carry[32];
    BinHeader hdr;
};

// TN: records mode -- contact records and / or the near-cutoff side channel are compiled in (chc_params.emit_contacts,
//     track_near); the counts-only production path carries none of it.
// FUSED (with SMALL): no binned record at all.  The CTA reads its frame straight from the caller's coordinates (the frame it
//     needs next is prefetched into L2 one frame ahead with a bulk prefetch), sorts the pair atoms into the cell list itself
//     (shared-memory counting sort: cell + rank by one shared atomic, packed 16-bit scan, scatter into the slot-pair arrays)
//     and goes on to the pair phase: every coordinate crosses HBM once and nothing is written back but counts.
#define FUSED_APT 4             // pair atoms per thread the in-kernel sort can take (n_pair <= FUSED_APT * THREADS)
#define FUSED_AUX 3             // aux coordinates per thread (3 n_aux <= FUSED_AUX * THREADS)
template <int THREADS, bool SMALL, bool TN, bool FUSED>
__global__ void __launch_bounds__(THREADS, 1)
pair_frames_kernel(DevTopo T, DevParams P, DevOut O, PairLayout L, BinLayout B, const unsigned char* __restrict__ binned,
                   int n_frames, long long first_index, int frag_frames, long long frag0,
                   const unsigned char* __restrict__ hpmask, unsigned* __restrict__ gstamp, unsigned* __restrict__ gstate,
                   int* __restrict__ wscratch, const float* __restrict__ xyz, long long frame_stride, unsigned char* __restrict__ gsoa) {
    static_assert(!FUSED || SMALL, "the in-kernel sort feeds the shared-memory path");
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float* tv = reinterpret_cast<float*>(smem + L.off_tv);
    float* th = reinterpret_cast<float*>(smem + L.off_th);
    PairMisc* M = reinterpret_cast<PairMisc*>(smem + L.off_misc);
    // queue entries: the pair (i, j) of sorted slots is one word on the shared-memory path (slots < 65 536), two otherwise
    constexpr int QW = SMALL ? 1 : 2;
    using cell_t = typename std::conditional<SMALL, unsigned short, int>::type;     // BinLayout::cell16 <=> SMALL
    // E1 -> E2 entries (atom slot, first neighbour slot, survivor mask): 8 bytes on the shared-memory path, 16 otherwise
    uint2* qe = reinterpret_cast<uint2*>(smem + L.off_qf) + warp * QE_CAP;
    uint4* qe4 = reinterpret_cast<uint4*>(smem + L.off_qf) + warp * QE_CAP;
    auto ent_put = [&](int at, unsigned i, unsigned first, unsigned m) {
        if (SMALL) qe[at] = make_uint2(i | (first << 16), m); else qe4[at] = make_uint4(i, first, m, 0u);
    };
    auto ent_get = [&](int at, unsigned& i, unsigned& first, unsigned& m) {
        if (SMALL) { const uint2 e = qe[at]; i = e.x & 0xffffu; first = e.x >> 16; m = e.y; }
        else { const uint4 e = qe4[at]; i = e.x; first = e.y; m = e.z; }
    };
    auto ent_mask = [&](int at, unsigned m) { if (SMALL) qe[at].y = m; else qe4[at].z = m; };
    unsigned* q2 = reinterpret_cast<unsigned*>(smem + L.off_q2) + warp * (2 * (QW + 1) * Q2_CAP);  // plain | polar: pair, class
    auto qput = [](unsigned* q, int cap, int at, unsigned i, unsigned j) {
        if (SMALL) q[at] = i | (j << 16); else { q[at] = i; q[cap + at] = j; }
    };
    auto qget = [](const unsigned* q, int cap, int at, unsigned& i, unsigned& j) {
        if (SMALL) { const unsigned w = q[at]; i = w & 0xffffu; j = w >> 16; } else { i = q[at]; j = q[cap + at]; }
    };
    for (int k = tid; k < T.K * T.K; k += THREADS) { tv[k] = T.tv[k]; th[k] = T.th[k]; }
    unsigned* bits;
    unsigned gen = 0;
    float* p_lkf = reinterpret_cast<float*>(smem + L.off_plkf);
    unsigned short* p_res = reinterpret_cast<unsigned short*>(smem + L.off_pres);
    unsigned char* p_st = reinterpret_cast<unsigned char*>(smem + L.off_pst);
    if (SMALL) {
        for (int k = tid; k < T.n_pair; k += THREADS) {
            if (!FUSED) p_lkf[k] = (float)T.pair_lkey[k];
            p_res[k] = (unsigned short)T.pair_res[k]; p_st[k] = (unsigned char)((T.pair_info[k] >> 14) & 0x1fu);
        }
        bits = reinterpret_cast<unsigned*>(smem + L.off_bits);
        if (!FUSED && tid == 0) { mbar_init(&M->bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
        if (FUSED) {
            // the padding slots behind the last atom never change; all cell counters start at zero; the box starts empty
            float4* e0 = reinterpret_cast<float4*>(smem + L.off_exy); float4* z0 = reinterpret_cast<float4*>(smem + L.off_ez);
            unsigned* w0 = reinterpret_cast<unsigned*>(smem + L.off_sw);
            const int np2 = (T.n_pair + 1) >> 1;
            for (int q = tid; q < np2 + E1_PAD_PAIRS; q += THREADS) {
                e0[q] = make_float4(3.0e18f, 3.0e18f, 3.0e18f, 3.0e18f); z0[q] = make_float4(3.0e18f, 3.0e18f, -3.0e7f, -3.0e7f);
                w0[2 * q] = 0u; w0[2 * q + 1] = 0u;
            }
            unsigned* cw = reinterpret_cast<unsigned*>(smem + L.off_cnt);
            for (int q = tid; q < 4 * ((B.max_cells + 8) / 8); q += THREADS) cw[q] = 0u;
            if (tid < 6) { M->bb_min[tid / 3][tid % 3] = 0x7fffffff; M->bb_max[tid / 3][tid % 3] = (int)0x80000000; }
        }
    } else {
        if (L.bits_plain) bits = reinterpret_cast<unsigned*>(smem + L.off_bits);      // mid-size systems: the plain set fits beside the queues
        else {
            bits = gstamp + (size_t)blockIdx.x * L.bits_words;
            gen = gstate[blockIdx.x];          // generation carried across launches (0 on the first one)
        }
    }
    __syncthreads();
    unsigned phase = 0;
    // FUSED: this thread's pair atoms p = tid + k THREADS of the frame to be sorted next (coordinates, locality key, flag word)
    // and its share of the aux coordinates; loaded by fused_load, which also merges the atoms into the bounding box M->bb_*[box].
    float rx[FUSED ? FUSED_APT : 1], ry[FUSED ? FUSED_APT : 1], rz[FUSED ? FUSED_APT : 1], rlk[FUSED ? FUSED_APT : 1], raux[FUSED ? FUSED_AUX : 1];
    unsigned rinfo[FUSED ? FUSED_APT : 1];
    int par = 0;
    auto fused_load = [&](int fn, int box) {
        if (!FUSED) return;
        // every register is (re)defined on every call, also past the last frame: nothing stays live through the pair phase
        const bool live = fn < n_frames;
        const float* X = xyz + (size_t)(live ? fn : 0) * (size_t)frame_stride;
        float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
#pragma unroll
        for (int k = 0; k < (FUSED ? FUSED_APT : 1); ++k) {
            const int p = tid + k * THREADS;
            rx[k] = 0.f; ry[k] = 0.f; rz[k] = 0.f; rlk[k] = 0.f; rinfo[k] = 0u;
            if (live && p < T.n_pair) {
                const float* q = X + 3 * (size_t)__ldg(&T.pair_atom[p]);
                rx[k] = __ldcs(q); ry[k] = __ldcs(q + 1); rz[k] = __ldcs(q + 2);
                rlk[k] = (float)__ldg(&T.pair_lkey[p]);
                rinfo[k] = (__ldg(&T.pair_info[p]) & I_WMASK) | ((unsigned)p << I_PIDX_SHIFT);
                lo[0] = fminf(lo[0], rx[k]); hi[0] = fmaxf(hi[0], rx[k]);
                lo[1] = fminf(lo[1], ry[k]); hi[1] = fmaxf(hi[1], ry[k]);
                lo[2] = fminf(lo[2], rz[k]); hi[2] = fmaxf(hi[2], rz[k]);
            }
        }
#pragma unroll
        for (int k = 0; k < (FUSED ? FUSED_AUX : 1); ++k) {
            const int e = tid + k * THREADS;
            raux[k] = 0.f;
            if (live && e < 3 * T.n_aux) { const int a = e / 3; raux[k] = __ldcs(X + 3 * (size_t)__ldg(&T.aux_atom[a]) + (e - 3 * a)); }
        }
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            const int a = __reduce_min_sync(FULL, f2ord(lo[d])), c = __reduce_max_sync(FULL, f2ord(hi[d]));
            if (live && lane == 0) { atomicMin(&M->bb_min[box][d], a); atomicMax(&M->bb_max[box][d], c); }
        }
    };
    if (FUSED) {
        if (tid == 0 && (int)(blockIdx.x + gridDim.x) < n_frames) {
            const unsigned long long a = reinterpret_cast<unsigned long long>(xyz + (size_t)(blockIdx.x + gridDim.x) * (size_t)frame_stride);
            const unsigned long long a0 = a & ~15ull;
            const unsigned bytes = ((unsigned)(a - a0) + 12u * (unsigned)T.n_atoms) & ~15u;
            if (bytes) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a0), "r"(bytes) : "memory");
        }
        fused_load((int)blockIdx.x, 0);
        __syncthreads();
    }

    for (int f = blockIdx.x; f < n_frames; f += gridDim.x) {
        const unsigned char* rec = binned + (size_t)f * B.stride;
        const long long gidx = first_index + f;
        const float4* spos; const float* aux; const cell_t* cells; const BinHeader* hdr;
        if (FUSED) {
            // ---- S: this CTA's frame into the cell-sorted slot-pair arrays.  The coordinates (and the static words) of this
            //      thread's pair atoms are already in registers: every warp loaded them at the end of its pair phase of the
            //      previous frame (fused_load below: the global-memory latency hides in the frame tail), and the bounding box
            //      of the frame is complete in M->bb_*[par]. ----
            const int NPs = T.n_pair;
            const bool gridb = T.n_pair_a < NPs;
            float4* e_xy = reinterpret_cast<float4*>(smem + L.off_exy);
            float4* e_zk = reinterpret_cast<float4*>(smem + L.off_ez);
            unsigned* e_w = reinterpret_cast<unsigned*>(smem + L.off_sw);
            float* auxs = reinterpret_cast<float*>(smem + L.off_aux);
            unsigned short* cnt16 = reinterpret_cast<unsigned short*>(smem + L.off_cells);
            unsigned* cntw = reinterpret_cast<unsigned*>(smem + L.off_cnt);          // transient counters: all zero between frames
            if (tid == 0) {
#ifndef NO_L2_PREFETCH
                if (f + 2 * (int)gridDim.x < n_frames) {   // two frames ahead: the next one is read in this frame's tail
                    const unsigned long long a = reinterpret_cast<unsigned long long>(xyz + (size_t)(f + 2 * (int)gridDim.x) * (size_t)frame_stride);
                    const unsigned long long a0 = a & ~15ull;
                    const unsigned bytes = ((unsigned)(a - a0) + 12u * (unsigned)T.n_atoms) & ~15u;     // rounded down: stays inside the frame
                    if (bytes) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a0), "r"(bytes) : "memory");
                }
#endif
                M->cursor = 0; M->ring_cursor = 0; M->nmark = 0; M->rn_pc = 0; M->rn_st = 0;
            }
            float org[3], inv; int nc[3];
            {
                float ext[3];
#pragma unroll
                for (int d = 0; d < 3; ++d) { org[d] = ord2f(M->bb_min[par][d]); ext[d] = ord2f(M->bb_max[par][d]) - org[d]; }
                grid_dims(ext, P.cell, B.max_cells, gridb, inv, nc);
            }
            const int fpx = nc[0] + 2, fpxy = fpx * (nc[1] + 2), fncell = fpxy * (nc[2] + 2);
            const int nkey = gridb ? 2 * fncell : fncell;
            // S2: cell of every atom; the shared atomic on the packed 16-bit counters returns the atom's rank inside its cell.
            //     The aux atoms (donor hydrogens, ring atoms, cations) and the empty residue-pair set ride in the same interval.
            unsigned r_cr[FUSED_APT];
#pragma unroll
            for (int k = 0; k < FUSED_APT; ++k) {
                const int p = tid + k * THREADS;
                if (p < NPs) {
                    int c = (cell_of(rz[k], org[2], inv, nc[2]) + 1) * fpxy + (cell_of(ry[k], org[1], inv, nc[1]) + 1) * fpx + cell_of(rx[k], org[0], inv, nc[0]) + 1;
                    if (p >= T.n_pair_a) c += fncell;
                    const unsigned sh = ((unsigned)c & 1u) << 4;
                    const unsigned old = atomicAdd(&cntw[c >> 1], 1u << sh);
                    r_cr[k] = (unsigned)c | (((old >> sh) & 0xffffu) << 16);
                } else r_cr[k] = 0u;
            }
#pragma unroll
            for (int k = 0; k < FUSED_AUX; ++k) { const int e = tid + k * THREADS; if (e < 3 * T.n_aux) auxs[e] = raux[k]; }
            for (int w = tid; w < L.bits_words; w += THREADS) bits[w] = 0u;
            __syncthreads();
            if (tid < 3) { M->bb_min[par ^ 1][tid] = 0x7fffffff; M->bb_max[par ^ 1][tid] = (int)0x80000000; }   // the box this frame's tail fills
            // S3: exclusive scan of counters 0..nkey into the cell starts; the counters go back to zero.  Thread t owns counters
            //     [8 groups t, 8 groups (t + 1)).
            {
                const int groups = (nkey + 8 * THREADS) / (8 * THREADS);
                uint4 v[4];
                unsigned s2 = 0u;         // even counters in the low half, odd ones in the high half (sums <= n_pair < 65 536)
#pragma unroll
                for (int g = 0; g < 4; ++g)
                    if (g < groups) {
                        const bool have = 8 * (tid * groups + g) <= nkey;      // only the counters 0..nkey exist
                        v[g] = have ? reinterpret_cast<const uint4*>(cntw)[tid * groups + g] : make_uint4(0u, 0u, 0u, 0u);
                        if (have) reinterpret_cast<uint4*>(cntw)[tid * groups + g] = make_uint4(0u, 0u, 0u, 0u);
                        s2 += (v[g].x + v[g].y) + (v[g].z + v[g].w);
                    }
                const int sum = (int)((s2 & 0xffffu) + (s2 >> 16));
                int incl = sum;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
                if (lane == 31) M->carry[warp] = incl;
                __syncthreads();
                const int w0 = lane < (THREADS >> 5) ? M->carry[lane] : 0;
                int w = w0;
#pragma unroll
                for (int o = 1; o < (THREADS >> 5); o <<= 1) { const int t = __shfl_up_sync(FULL, w, o); if (lane >= o) w += t; }
                unsigned run = (unsigned)(__shfl_sync(FULL, w - w0, warp) + incl - sum);
                auto ex2 = [&](unsigned word) {      // two counters -> their two exclusive starts
                    const unsigned lo2 = word & 0xffffu, e = run | ((run + lo2) << 16);
                    run += lo2 + (word >> 16);
                    return e;
                };
#pragma unroll
                for (int g = 0; g < 4; ++g)
                    if (g < groups && 8 * (tid * groups + g) <= nkey) {
                        uint4 e;
                        e.x = ex2(v[g].x); e.y = ex2(v[g].y); e.z = ex2(v[g].z); e.w = ex2(v[g].w);
                        reinterpret_cast<uint4*>(smem + L.off_cells)[tid * groups + g] = e;
                    }
            }
            __syncthreads();
            // S4: scatter into the slot-pair arrays (x0 x1 y0 y1 | z0 z1 lkey0 lkey1) and the flag words
            {
                float* fxy = reinterpret_cast<float*>(e_xy); float* fzk = reinterpret_cast<float*>(e_zk);
#pragma unroll
                for (int k = 0; k < FUSED_APT; ++k) {
                    const int p = tid + k * THREADS;
                    if (p < NPs) {
                        const int dst = (int)cnt16[r_cr[k] & 0xffffu] + (int)(r_cr[k] >> 16);
                        const int o4 = 4 * (dst >> 1) + (dst & 1);
                        fxy[o4] = rx[k]; fxy[o4 + 2] = ry[k];
                        fzk[o4] = rz[k]; fzk[o4 + 2] = rlk[k];
                        e_w[dst] = rinfo[k];
                    }
                }
                if (tid == 0) {
                    BinHeader h;
                    h.org[0] = org[0]; h.org[1] = org[1]; h.org[2] = org[2]; h.inv = inv;
                    h.nc[0] = nc[0]; h.nc[1] = nc[1]; h.nc[2] = nc[2]; h.ncell = fncell;
                    M->hdr = h;
                }
            }
            hdr = &M->hdr; spos = nullptr; aux = auxs; cells = reinterpret_cast<const cell_t*>(smem + L.off_cells);
        } else if (SMALL) {
            if (tid == 0) {
                const int ncell = __ldcg(&reinterpret_cast<const BinHeader*>(rec)->ncell);
                const unsigned cbytes = (unsigned)((sizeof(cell_t) * ((T.n_pair_a < T.n_pair ? 2 : 1) * ncell + 1) + 15) & ~15);
                // the previous frame's generic-proxy reads of these buffers are ordered before the async-proxy writes
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_expect_tx(&M->bar, (unsigned)B.fixed_bytes + cbytes);
                bulk_g2s(smem + L.off_rec, rec, (unsigned)B.fixed_bytes, &M->bar);
                bulk_g2s(smem + L.off_cells, rec + B.off_cells, cbytes, &M->bar);
                // the record this CTA needs next: on its way to L2 while this frame is enumerated (large chunks spill to HBM)
#ifndef NO_L2_PREFETCH
                if (f + (int)gridDim.x < n_frames) {
                    const unsigned char* nxt = rec + (size_t)gridDim.x * B.stride;
                    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(nxt), "r"((unsigned)B.fixed_bytes) : "memory");
                    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(nxt + B.off_cells), "r"(cbytes) : "memory");
                }
#endif
                M->cursor = 0; M->ring_cursor = 0; M->nmark = 0; M->rn_pc = 0; M->rn_st = 0;
            }
            for (int w = tid; w < L.bits_words; w += THREADS) bits[w] = 0u;
            mbar_wait(&M->bar, phase);
            phase ^= 1u;
            hdr = reinterpret_cast<const BinHeader*>(smem + L.off_rec);
            spos = reinterpret_cast<const float4*>(smem + L.off_rec + B.off_spos);
            aux = reinterpret_cast<const float*>(smem + L.off_rec + B.off_aux);
            cells = reinterpret_cast<const cell_t*>(smem + L.off_cells);
        } else {
            if (tid == 0) { M->cursor = 0; M->ring_cursor = 0; M->nmark = 0; M->rn_pc = 0; M->rn_st = 0; }
            if (L.bits_plain) {
                for (int w = tid; w < L.bits_words; w += THREADS) bits[w] = 0u;
            } else {
                gen = gen + 1u;
                if (gen > 255u) {               // generations wrapped: wipe the table once per 255 frames
                    for (int w = tid; w < L.bits_words; w += THREADS) bits[w] = 0u;
                    gen = 1u;
                }
            }
            hdr = reinterpret_cast<const BinHeader*>(rec);
            spos = reinterpret_cast<const float4*>(rec + B.off_spos);
            aux = reinterpret_cast<const float*>(rec + B.off_aux);
            cells = reinterpret_cast<const cell_t*>(rec + B.off_cells);
        }
        // ring centroids (float) for the prefilter below, in the barrier interval that ends the frame set-up
        float* rc = reinterpret_cast<float*>(smem + L.off_rc);
        const int n_pc = (P.itypes & CHC_IT_PC) ? T.n_cat * T.n_rings : 0;
        const int n_st = (P.itypes & (CHC_IT_PS | CHC_IT_TS)) ? T.n_rings * (T.n_rings - 1) / 2 : 0;
        if (n_pc + n_st > 0)
            for (int rr = tid; rr < T.n_rings; rr += THREADS) {
                const int* t1 = T.ring_aux + 3 * rr;
                float cx = 3.0e18f, cy = 0.f, cz = 0.f;                       // a ring without its three atoms takes part in nothing
                if (t1[0] >= 0) {
                    const float* p1 = aux + 3 * t1[0]; const float* p2 = aux + 3 * t1[1]; const float* p3 = aux + 3 * t1[2];
                    cx = (p1[0] + p2[0] + p3[0]) * (1.0f / 3.0f); cy = (p1[1] + p2[1] + p3[1]) * (1.0f / 3.0f); cz = (p1[2] + p2[2] + p3[2]) * (1.0f / 3.0f);
                }
                rc[3 * rr] = cx; rc[3 * rr + 1] = cy; rc[3 * rr + 2] = cz;
            }
        __syncthreads();
        // ---- R1: ring tasks (cation x ring, ring x ring) through a float prefilter on the ring centroids; the few that can reach
        //      a hard distance cutoff are listed and evaluated densely (float64) after the pair phase.  The prefilter radius is the
        //      largest hard cutoff + 0.05 A: a superset of what ring_pc_task / ring_stack_task accept or report as near-cutoff.
        //      Runs in the same barrier interval as the conversion below (both only read the record). ----
        unsigned* rlist = reinterpret_cast<unsigned*>(smem + L.off_rlist);
        if (!SMALL && L.off_rgrid > 0 && n_pc + n_st > 0) {
            // Large systems: the all-pairs prefilter below is quadratic (config 4: 2.8 M (cation, ring) and (ring, ring) tasks per
            // frame).  The ring centroids are binned into a coarse grid laid over the pair grid's box -- cells of k x k x k pair
            // cells, k edges >= the prefilter radius, at most RG_CAP cells -- and every cation / ring looks at its 27 cells only.
            unsigned* gcnt = reinterpret_cast<unsigned*>(smem + L.off_rgrid);
            unsigned short* gring = reinterpret_cast<unsigned short*>(smem + L.off_rsort);
            int k = max(1, (int)ceilf(sqrtf(P.ring_pre2) * hdr->inv * 1.001f + 0.01f));
            int gx, gy, gz;
            for (;; ++k) {
                gx = (hdr->nc[0] + k - 1) / k; gy = (hdr->nc[1] + k - 1) / k; gz = (hdr->nc[2] + k - 1) / k;
                if ((long long)gx * gy * gz <= RG_CAP) break;
            }
            const int gcells = gx * gy * gz;
            const unsigned kmagic = 0xffffffffu / (unsigned)k + 1u;           // pair-cell index / k (indices < 2^16)
            auto coarse = [&](float x, float y, float z, int& cx, int& cy, int& cz) {
                cx = (int)__umulhi((unsigned)cell_of(x, hdr->org[0], hdr->inv, hdr->nc[0]), kmagic);
                cy = (int)__umulhi((unsigned)cell_of(y, hdr->org[1], hdr->inv, hdr->nc[1]), kmagic);
                cz = (int)__umulhi((unsigned)cell_of(z, hdr->org[2], hdr->inv, hdr->nc[2]), kmagic);
            };
            for (int c = tid; c <= gcells; c += THREADS) gcnt[c] = 0u;
            __syncthreads();
            for (int rr = tid; rr < T.n_rings; rr += THREADS) {
                if (rc[3 * rr] > 1.0e18f) continue;
                int cx, cy, cz; coarse(rc[3 * rr], rc[3 * rr + 1], rc[3 * rr + 2], cx, cy, cz);
                atomicAdd(&gcnt[(cz * gy + cy) * gx + cx], 1u);
            }
            __syncthreads();
            {   // exclusive scan of the counts (thread t owns RG_CAP / THREADS consecutive cells)
                constexpr int PER = (RG_CAP + THREADS - 1) / THREADS;
                unsigned v[PER], sum = 0u;
#pragma unroll
                for (int q = 0; q < PER; ++q) { const int c = tid * PER + q; v[q] = c < gcells ? gcnt[c] : 0u; sum += v[q]; }
                unsigned incl = sum;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
                if (lane == 31) M->carry[warp] = (int)incl;
                __syncthreads();
                unsigned base = 0u;
                for (int w = 0; w < warp; ++w) base += (unsigned)M->carry[w];
                unsigned run = base + incl - sum;
#pragma unroll
                for (int q = 0; q < PER; ++q) { const int c = tid * PER + q; if (c < gcells) gcnt[c] = run; run += v[q]; }
            }
            __syncthreads();
            // fill: the running start of a cell ends up as its END; cell c then spans [c ? gcnt[c - 1] : 0, gcnt[c])
            for (int rr = tid; rr < T.n_rings; rr += THREADS) {
                if (rc[3 * rr] > 1.0e18f) continue;
                int cx, cy, cz; coarse(rc[3 * rr], rc[3 * rr + 1], rc[3 * rr + 2], cx, cy, cz);
                gring[atomicAdd(&gcnt[(cz * gy + cy) * gx + cx], 1u)] = (unsigned short)rr;
            }
            __syncthreads();
            const int n_cat_on = n_pc > 0 ? T.n_cat : 0, n_ring_on = n_st > 0 ? T.n_rings : 0;
            for (int sidx = tid; sidx < n_cat_on + n_ring_on; sidx += THREADS) {
                const bool pc = sidx < n_cat_on;
                const int me = pc ? sidx : sidx - n_cat_on;
                float ax, ay, az;
                if (pc) { const float* xc = aux + 3 * T.cat_aux[me]; ax = xc[0]; ay = xc[1]; az = xc[2]; }
                else { ax = rc[3 * me]; ay = rc[3 * me + 1]; az = rc[3 * me + 2]; if (ax > 1.0e18f) continue; }
                const int excl = pc ? T.cat_excl[me] : -1;
                int cx, cy, cz; coarse(ax, ay, az, cx, cy, cz);
                const int x0 = max(cx - 1, 0), x1 = min(cx + 1, gx - 1);
                for (int zz = max(cz - 1, 0); zz <= min(cz + 1, gz - 1); ++zz)
                    for (int yy = max(cy - 1, 0); yy <= min(cy + 1, gy - 1); ++yy) {
                        const int row = (zz * gy + yy) * gx;
                        const int s0 = row + x0 > 0 ? (int)gcnt[row + x0 - 1] : 0, s1 = (int)gcnt[row + x1];
                        for (int sl = s0; sl < s1; ++sl) {
                            const int rr = (int)gring[sl];
                            if (pc ? rr == excl : rr <= me) continue;
                            const float dx = ax - rc[3 * rr], dy = ay - rc[3 * rr + 1], dz = az - rc[3 * rr + 2];
                            if (!(dx * dx + dy * dy + dz * dz <= P.ring_pre2)) continue;
                            const int at = atomicAdd(pc ? &M->rn_pc : &M->rn_st, 1);
                            if (at < L.rlist_cap) {
                                if (pc) rlist[at] = (unsigned)(me * T.n_rings + rr);
                                else rlist[L.rlist_cap - 1 - at] = (unsigned)(rr * (rr - 1) / 2 + me);
                            }
                        }
                    }
            }
        } else if (n_pc + n_st > 0) {
            const unsigned nr_m = 0xffffffffu / (unsigned)max(T.n_rings, 1) + 1u;
            for (int t = tid; t < n_pc + n_st; t += THREADS) {
                float ax, ay, az, bx, by, bz;
                if (t < n_pc) {
                    int c = (int)__umulhi((unsigned)t, nr_m);
                    if (c * T.n_rings > t) --c;                               // rounded-up reciprocal: exact below 2^16, corrected above
                    const int rr = t - c * T.n_rings;
                    if (T.cat_excl[c] == rr) continue;
                    const float* xc = aux + 3 * T.cat_aux[c];
                    ax = xc[0]; ay = xc[1]; az = xc[2];
                    bx = rc[3 * rr]; by = rc[3 * rr + 1]; bz = rc[3 * rr + 2];
                } else {
                    const int q = t - n_pc;
                    int r2 = (int)((1.0f + sqrtf(1.0f + 8.0f * (float)q)) * 0.5f);
                    while (r2 * (r2 - 1) / 2 > q) --r2;
                    while ((r2 + 1) * r2 / 2 <= q) ++r2;
                    const int r1 = q - r2 * (r2 - 1) / 2;
                    ax = rc[3 * r1]; ay = rc[3 * r1 + 1]; az = rc[3 * r1 + 2];
                    bx = rc[3 * r2]; by = rc[3 * r2 + 1]; bz = rc[3 * r2 + 2];
                    if (ax > 1.0e18f || bx > 1.0e18f) continue;
                }
                const float dx = ax - bx, dy = ay - by, dz = az - bz;
                if (!(dx * dx + dy * dy + dz * dz <= P.ring_pre2)) continue;
                const bool pc = t < n_pc;
                const int at = atomicAdd(pc ? &M->rn_pc : &M->rn_st, 1);
                if (pc) { if (at < L.rlist_cap) rlist[at] = (unsigned)t; }
                else if (at < L.rlist_cap) rlist[L.rlist_cap - 1 - at] = (unsigned)(t - n_pc);
            }
        }
        // the slot-pair arrays: shared memory (small path) or this CTA's slice of the global scratch (large path: L2 / L1)
        unsigned char* soa = SMALL ? smem : gsoa + (size_t)blockIdx.x * (size_t)L.soa_stride;
        float4* exy = reinterpret_cast<float4*>(soa + L.off_exy);
        float4* ezk = reinterpret_cast<float4*>(soa + L.off_ez);
        unsigned* partner = reinterpret_cast<unsigned*>(soa + L.off_ek);      // [E1_PARTNER_SETS][pwords] bit per sorted slot
        const int pwords = ((T.n_pair + 2 * E1_PAD_PAIRS + 31) >> 5) + 2;
        unsigned* sw = reinterpret_cast<unsigned*>(soa + L.off_sw);
        {
            // The float4 image becomes slot PAIRS (E1 tests two neighbours per packed instruction): x0 x1 y0 y1 | z0 z1 lkey0 lkey1,
            // plus the flag word per slot.
            const int NPs = T.n_pair, np2 = (NPs + 1) >> 1;
            if (!FUSED) for (int q = tid; q < np2 + E1_PAD_PAIRS; q += THREADS) {
                const int j0 = 2 * q, j1 = j0 + 1;
                const float4 far = make_float4(3.0e18f, 3.0e18f, 3.0e18f, 0.f);
                const float4 a = j0 < NPs ? spos[j0] : far, b = j1 < NPs ? spos[j1] : far;
                exy[q] = make_float4(a.x, b.x, a.y, b.y);
                const unsigned wa = __float_as_uint(a.w), wb = __float_as_uint(b.w);
                const int pa = wa >> I_PIDX_SHIFT, pb = wb >> I_PIDX_SHIFT;
                const float la = j0 < NPs ? (SMALL ? p_lkf[pa] : (float)__ldg(&T.pair_lkey[pa])) : -3.0e7f;
                const float lb = j1 < NPs ? (SMALL ? p_lkf[pb] : (float)__ldg(&T.pair_lkey[pb])) : -3.0e7f;
                ezk[q] = make_float4(a.z, b.z, la, lb);
                sw[j0] = wa; sw[j1] = wb;
            }
            for (int wd = warp; wd < pwords; wd += THREADS / 32) {        // partner bits, one word of 32 slots per warp round
                const int j = wd * 32 + lane;
                const unsigned w = j < NPs ? (FUSED ? sw[j] : __float_as_uint(spos[j].w)) : 0u;
                const unsigned hbu = __ballot_sync(FULL, (w & I_HBU) != 0u), cat = __ballot_sync(FULL, (w & I_CATION) != 0u),
                               an = __ballot_sync(FULL, (w & I_ANION) != 0u), solv = __ballot_sync(FULL, (w & I_SOLV) != 0u);
                if (lane < E1_PARTNER_SETS)
                    partner[lane * pwords + wd] = lane == 8 ? ~solv : (((lane & 1) ? hbu : 0u) | ((lane & 2) ? cat : 0u) | ((lane & 4) ? an : 0u) | solv);
            }
            __syncthreads();               // small path: from here on the float4 image is dead, its shared memory holds the (empty) queues
        }
        const int px = hdr->nc[0] + 2, pxy = px * (hdr->nc[1] + 2), ncell = hdr->ncell;
        FrameCtxT<SMALL> F{exy, ezk, sw, p_res, p_st, aux, bits, gen, tv, th,
                           hpmask + (size_t)(gidx / frag_frames - frag0) * T.n_pair, (int)gidx,
                           water_scratch(wscratch, blockIdx.x, T.n_solv), &M->nmark};
        // ---- E: pair interactions ------------------------------------------------------------------------------
        const float cut2p = P.pre2 * 1.000002f;       // largest cutoff^2 of the enabled pair types, padded for the fused filter
        int qfn = 0, q2n = 0, q3n = 0;            // entries queued for E2; classified pairs queued for E3 (plain, polar)
        // E1 state.  A tile = 32 consecutive cell-sorted atoms of grid A, one per lane (atom i).  Lane ranges:
        // 0..4 = half shell of grid A (own row from the slot after i to the end of cell x+1, then rows (y+1,z) (y-1,z+1)
        // (y,z+1) (y+1,z+1) from x-1 to x+1); 5..13 (only with unselected solvent) = the full 3x3 rows of grid B.
        const int nA = T.n_pair_a;
        const int nr = T.n_pair_a < T.n_pair ? 14 : 5;
        const unsigned nr_magic = 0xffffffffu / (unsigned)nr + 1u;
        const int ntiles = (nA + 31) >> 5;
        const int n_whole = max(ntiles - E1_TAIL_TILES, 0), nunits_s = n_whole + (ntiles - n_whole) * nr;   // SMALL: unit numbering
        bool whole = false;
        int r = 0, tile = -1, k0 = 0, maxlen = 0;      // warp-uniform: current range and tile, chunk offset, longest lane range
        int my_i = -1, my_c = 0, beg = 0, len = 0;     // this lane's atom, its cell, its current range
        float ix = 0.f, iy = 0.f, iz = 0.f;
        int mbase = 0;                                 // first slot of the chunk this lane tests: neighbour j = mbase + bit
        bool exhausted = false;
        // SMALL: this lane's locality key (float) and partner bit array; lead = its range starts on an odd slot
        float my_lk = 0.f; const unsigned* my_partner = partner; int lead = 0;
        const float loc_thr = P.vdw_rd > 0 ? -((float)P.vdw_rd - 0.5f) * ((float)P.vdw_rd - 0.5f) : 1.0f;   // dl^2 + loc_thr < 0 <=> |dl| < vdw_res_diff
        for (;;) {
            // ---- E1 ----
            {
                // Slot pairs, packed float32 arithmetic; distance AND locality / polar-partner filter in one pass; the survivor masks
                // go to the warp's entry queue (qfn = entries queued; 32 entries hold at least one E2 batch).
                while (qfn < E1_TARGET && !exhausted) {
                    k0 += 32;
                    if (k0 >= maxlen) {
                        // next range of a whole-tile unit, or the next unit of work: whole tiles while many remain (one shared
                        // atomic per five ranges), single (tile, range) units for the last E1_TAIL_TILES tiles (short frame tail)
                        int t = tile;
                        if (whole && r + 1 < nr) ++r;
                        else {
                            int u = 0;
                            if (lane == 0) u = atomicAdd(&M->cursor, 1);
                            u = __shfl_sync(FULL, u, 0);
                            if (u >= nunits_s) { exhausted = true; break; }
                            whole = u < n_whole;
                            if (whole) { t = u; r = 0; }
                            else { const int v = u - n_whole; const int q = (int)__umulhi((unsigned)v, nr_magic); t = n_whole + q; r = v - q * nr; }
                        }
                        if (t != tile) {
                            tile = t;
                            my_i = t * 32 + lane;
                            if (my_i < nA) {
                                const float4 p = F.at(my_i);
                                ix = p.x; iy = p.y; iz = p.z;
                                my_lk = (float)F.lk(my_i, 0u);
                                my_partner = partner + e1_partner_set(__float_as_uint(p.w)) * pwords;
                                my_c = (cell_of(iz, hdr->org[2], hdr->inv, hdr->nc[2]) + 1) * pxy +
                                       (cell_of(iy, hdr->org[1], hdr->inv, hdr->nc[1]) + 1) * px +
                                       cell_of(ix, hdr->org[0], hdr->inv, hdr->nc[0]) + 1;
                            } else my_i = -1;
                        }
                        len = 0; lead = 0;
                        if (my_i >= 0) {
                            if (r == 0) { beg = my_i + 1; len = (int)cells[my_c + 2] - beg; }
                            else {
                                int base;
                                if (r < 5) base = my_c + ((r == 1 || r == 4) - (r == 2)) * px + (r >= 2 ? pxy : 0);
                                else { const int q = r - 5; base = my_c + ncell + (q % 3 - 1) * px + (q / 3 - 1) * pxy; }
                                beg = (int)cells[base - 1];
                                len = (int)cells[base + 2] - beg;
                            }
                            lead = beg & 1;            // ranges are walked from an even slot; the slot before an odd start is masked off
                            len = len > 0 ? len + lead : 0;
                            beg -= lead;
                        }
                        maxlen = __reduce_max_sync(FULL, len);
                        k0 = 0;
                        if (maxlen == 0) continue;
                    }
                    // up to 16 slot pairs of every lane's range: fused-multiply-add d2 of two neighbours at a time against the padded
                    // largest cutoff (a superset filter: every class test in E2 compares the EXACT float32 d2 with its own cutoff).
                    // The sign bit of cut - d2 is shifted into the miss mask (first slot ends up highest).  Slots past the lane's own
                    // range are computed too (inside the padded arrays) and masked off afterwards.
                    const int npair = (min(32, maxlen - k0) + 1) >> 1;
                    mbase = beg + k0;
                    const float4* pxy2 = exy + (mbase >> 1);
                    const float4* pzk2 = ezk + (mbase >> 1);
                    const f32x2 X = pk2(ix, ix), Y = pk2(iy, iy), Z = pk2(iz, iz), C2 = pk2(cut2p, cut2p), LK = pk2(my_lk, my_lk),
                                TH = pk2(loc_thr, loc_thr);
                    unsigned miss = 0u, loc = 0u;
CHC_UNROLL(E1_UNROLL)
                    for (int q = 0; q < npair; ++q) {
                        const float4 a = pxy2[q];
                        const float4 b = pzk2[q];
                        const f32x2 dx = sub2(X, pk2(a.x, a.y)), dy = sub2(Y, pk2(a.z, a.w)), dz = sub2(Z, pk2(b.x, b.y));
                        const f32x2 sg = sub2(C2, fma2(dz, dz, fma2(dy, dy, mul2(dx, dx))));
                        const f32x2 dl = sub2(LK, pk2(b.z, b.w));
                        const f32x2 lc = fma2(dl, dl, TH);
                        miss = __funnelshift_l((unsigned)sg, miss, 1);
                        loc = __funnelshift_l((unsigned)lc, loc, 1);
                        miss = __funnelshift_l((unsigned)(sg >> 32), miss, 1);
                        loc = __funnelshift_l((unsigned)(lc >> 32), loc, 1);
                    }
                    // slot k of the chunk sits at bit 2 npair - 1 - k: bring it to bit k; a hit survives when it is not local or when
                    // the neighbour is a polar partner of this lane's atom; keep [lead, len) of the lane's range
                    const int sh = 32 - 2 * npair;
                    const unsigned hits = ~__brev(miss << sh), local = __brev(loc << sh);
                    const int wq = mbase >> 5, wr = mbase & 31;
                    const unsigned part = __funnelshift_r(my_partner[wq], my_partner[wq + 1], wr);
                    const int lo = max(lead - k0, 0), hi = min(len - k0, 32);
                    const unsigned valid = hi > lo ? ((hi >= 32 ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u)) : 0u;
                    // One queue entry per lane with survivors: (atom slot | first slot << 16, survivor mask), placed by ballot rank
                    // (no per-survivor loop, no atomics); E2 deals the survivors of the entries out to its 32 lanes.  qfn < 32 here
                    // and at most 32 entries are added: QE_CAP = 64 always holds them.
                    const unsigned surv = hits & (~local | part) & valid;
                    const unsigned bal = __ballot_sync(FULL, surv != 0u);
                    if (surv) ent_put(qfn + __popc(bal & ((1u << lane) - 1u)), (unsigned)my_i, (unsigned)mbase, surv);
                    qfn += __popc(bal);
                }
                if (qfn == 0 && q2n == 0 && q3n == 0) break;           // only reachable when exhausted
                __syncwarp();
            }
            // ---- E2: 32 filtered survivors; classified pairs go to the plain queue (vdW / hydrophobic only) or to the
            //      polar queue (salt bridge, hydrogen bond, water edge), so that an E3 batch runs one kind of code ----
            const bool drain1 = exhausted;
            if (qfn >= 32 || (drain1 && qfn > 0)) {
                unsigned cls = 0u; float d2 = 0.f; unsigned ei = 0u, ej = 0u;
                {
                    // Deal out up to 32 survivors from the top entries of the queue (lane L looks at entry qfn - 1 - L): prefix sum
                    // of the entries' survivor counts, the entry of survivor s from the entries' end marks (below), its bit by
                    // clearing the lower ones.  Entries used up leave the queue; the one cut by the batch keeps its unused bits.
                    const int nw = min(qfn, 32);
                    unsigned ent_i = 0u, ent_first = 0u, ent_m = 0u;
                    if (lane < nw) ent_get(qfn - 1 - lane, ent_i, ent_first, ent_m);
                    const int c = __popc(ent_m);
                    int incl = c;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
                    const int S = min(32, __shfl_sync(FULL, incl, 31));          // survivors in this batch
                    // entry e ends with survivor incl_e - 1: the OR of those end marks over the warp tells every survivor how many
                    // entries end before it (= its entry) and where its entry starts (one past the last end mark below it)
                    const unsigned endm = __reduce_or_sync(FULL, (lane < nw && incl <= 32) ? 1u << (incl - 1) : 0u);
                    const unsigned endb = endm & ((1u << lane) - 1u);
                    const int e = __popc(endb) & 31;
                    const int before = 32 - __clz(endb);
                    const unsigned src_i = __shfl_sync(FULL, ent_i, e), src_first = __shfl_sync(FULL, ent_first, e), em = __shfl_sync(FULL, ent_m, e);
                    unsigned bitpos = 0u;
                    if (lane < S) {
                        unsigned m = em;
                        for (int n = lane - before; n > 0; --n) m &= m - 1u;
                        bitpos = (unsigned)(__ffs(m) - 1);
                        ei = src_i; ej = src_first + bitpos;
                    }
                    // the entry holding survivor S - 1 is the last one touched: keep what the batch did not use of it
                    const int lb = __popc(endm & ((1u << (S - 1)) - 1u));
                    const unsigned last = __shfl_sync(FULL, bitpos, S - 1);               // bit of survivor S - 1 inside that entry
                    int keep = 0;
                    if (lane == lb) {
                        const unsigned m = ent_m & ~((2u << last) - 1u);
                        if (m) { ent_mask(qfn - 1 - lane, m); keep = 1; }
                    }
                    qfn -= lb + 1 - __shfl_sync(FULL, keep, lb);
                    if (lane < S) cls = e2_classify<TN>(T, P, F, (int)ei, (int)ej, d2);
                }
                const bool polar = (cls & (C_SB | C_HB | C_HBW | C_NEARS | C_NEARB)) != 0u;
                const unsigned balp = __ballot_sync(FULL, polar), balv = __ballot_sync(FULL, cls != 0u && !polar);
                if (cls) {
                    const unsigned below = (1u << lane) - 1u;
                    unsigned* q = polar ? q2 + (QW + 1) * Q2_CAP : q2;
                    const int at = polar ? q3n + __popc(balp & below) : q2n + __popc(balv & below);
                    qput(q, Q2_CAP, at, ei, ej); q[QW * Q2_CAP + at] = cls;
                }
                q2n += __popc(balv); q3n += __popc(balp);
                __syncwarp();
            }
            // ---- E3: 32 classified pairs of one queue (or the rest when nothing else is left) ----
            const bool drain = drain1 && qfn == 0;
            while (q2n >= 32 || q3n >= 32 || (drain && (q2n | q3n) != 0)) {
                const bool plain = q2n >= 32 || (q3n < 32 && q2n > 0);
                int& qn = plain ? q2n : q3n;
                const unsigned* q = plain ? q2 : q2 + (QW + 1) * Q2_CAP;
                const int nb = min(qn, 32);
                qn -= nb;
                if (lane < nb) {
                    unsigned ei, ej;
                    qget(q, Q2_CAP, qn + lane, ei, ej);
                    e3_finish<TN>(T, P, O, F, (int)ei, (int)ej, q[QW * Q2_CAP + qn + lane]);
                }
                __syncwarp();
            }
        }
        // ---- R2: the listed ring tasks (float64 geometry), 32 per unit, handed out through a shared counter AFTER the pair phase:
        //      warps that run out of atom tiles early fill the end-of-frame tail.  If the list overflowed (rlist_cap) every task is
        //      evaluated the slow way instead. ----
        if (n_pc + n_st > 0) {
            const int rn_pc = M->rn_pc, rn_st = M->rn_st;
            const bool listed = rn_pc + rn_st <= L.rlist_cap;
            const int m_pc = listed ? rn_pc : n_pc, m_st = listed ? rn_st : n_st;
            const int u_pc = (m_pc + 31) >> 5, u_all = u_pc + ((m_st + 31) >> 5);
            for (;;) {
                int u = 0;
                if (lane == 0) u = atomicAdd(&M->ring_cursor, 1);
                u = __shfl_sync(FULL, u, 0);
                if (u >= u_all) break;
                if (u < u_pc) {
                    const int k = u * 32 + lane;
                    if (k < m_pc) {
                        const int t = listed ? (int)rlist[k] : k;
                        ring_pc_task<TN>(T, P, O, F, t / T.n_rings, t % T.n_rings);
                    }
                } else {
                    const int k = (u - u_pc) * 32 + lane;
                    if (k < m_st) {
                        const int t = listed ? (int)rlist[L.rlist_cap - 1 - k] : k;
                        int r2 = (int)((1.0f + sqrtf(1.0f + 8.0f * (float)t)) * 0.5f);
                        while (r2 * (r2 - 1) / 2 > t) --r2;
                        while ((r2 + 1) * r2 / 2 <= t) ++r2;
                        ring_stack_task<TN>(T, P, O, F, t - r2 * (r2 - 1) / 2, r2);
                    }
                }
                __syncwarp();
            }
        }
        // FUSED: a warp that is done with this frame fetches its atoms of the frame the CTA sorts next (registers + bounding box)
        fused_load(f + (int)gridDim.x, par ^ 1);
        par ^= 1;
        __syncthreads();
        // ---- W: water bridges from the (non-solvent, solvent) hydrogen bonds the pair phase recorded ----------------
        if (T.n_solv > 0 && (P.itypes & CHC_IT_HB)) {
            WaterScratch W = water_scratch(wscratch, blockIdx.x, T.n_solv);
            water_bridges<TN>(T, P, O, F, W, cells, px, pxy, ncell, hdr, &M->nmark, tid, THREADS);
            __syncthreads();
        }
    }
    if (!SMALL && !L.bits_plain && tid == 0) gstate[blockIdx.x] = gen;
}
