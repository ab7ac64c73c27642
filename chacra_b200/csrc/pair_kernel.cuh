// pair_kernel.cuh -- every interaction type of one frame from its binned record (bin_kernel.cuh).
//
// Persistent CTAs, one frame at a time per CTA.  SMALL = the record fits in shared memory: one bulk async copy
// (cp.async.bulk + mbarrier) brings header | spos | aux, a second one the cell starts; the per-frame residue-pair
// set is a plain shared-memory bitset.  !SMALL = everything is read in place from global memory / L2 and the
// residue-pair set is a generation-stamped bitset private to the CTA (frame_kernel.cuh count_pair).
//
// Pair phase, per warp (cells are handed out 32 at a time through a shared counter):
//   E1  one cell's atoms (i) against the concatenated half-shell ranges (j): lane-per-j, loop over i, exact float32
//       d2 against the largest cutoff; survivors are compacted into the warp's queue with ballot/popc
//   E2  32 survivors at a time: class tests (selection, locality, vdW/hp thresholds, ion and hbond classes)
//   E3  32 classified pairs at a time: bonded / disulfide / hydrogen-angle tests, residue-pair set, counts, records
// The three stages form one resumable loop so that E2 and E3 are emitted exactly once in the SASS (the first
// version inlined them at ~45 sites and stalled on instruction fetch, profiles/r01_v1_*).
#pragma once
#include "frame_phases.cuh"
#include "water_phase.cuh"

#define Q1_CAP 64           // (i, j) survivors per warp
#define Q2_CAP 64           // classified pairs per warp
#define CELL_BLK 8          // cells a warp claims at a time (small blocks keep the end-of-frame tail short)

struct PairLayout {         // shared-memory plan of the pair kernel (bytes)
    int off_rec;            // SMALL: header | spos | aux  (B.fixed_bytes)
    int off_cells;          // SMALL: cell starts
    int off_lkey;           // SMALL: locality keys by pair index
    int off_bits;           // SMALL: residue-pair bitset
    int off_q1, off_q2, off_tv, off_th, off_misc;
    int bits_words;         // words of the residue-pair set (plain: P/32, stamped: P/24)
    int total;
};

struct PairMisc {
    unsigned long long bar;     // mbarrier of the bulk copies
    int cursor;                 // next block of CELL_BLK cells
    int nmark;                  // waters with at least one hydrogen bond to a non-solvent atom (this frame)
};

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    asm volatile(
        "{\n .reg .pred p;\n"
        "WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n bra WAIT_%=;\n"
        "DONE_%=:\n}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// 1-D bulk async copy global -> shared, completion counted in bytes on the mbarrier (TMA engine, no tensor map)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

template <int THREADS, bool SMALL>
__global__ void __launch_bounds__(THREADS, 1)
pair_frames_kernel(DevTopo T, DevParams P, DevOut O, PairLayout L, BinLayout B, const unsigned char* __restrict__ binned,
                   int n_frames, long long first_index, int frag_frames, long long frag0,
                   const unsigned char* __restrict__ hpmask, unsigned* __restrict__ gstamp, unsigned* __restrict__ gstate,
                   int* __restrict__ wscratch) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float* tv = reinterpret_cast<float*>(smem + L.off_tv);
    float* th = reinterpret_cast<float*>(smem + L.off_th);
    PairMisc* M = reinterpret_cast<PairMisc*>(smem + L.off_misc);
    unsigned* q1 = reinterpret_cast<unsigned*>(smem + L.off_q1) + warp * (2 * Q1_CAP);       // i[Q1_CAP], j[Q1_CAP]
    unsigned* q2 = reinterpret_cast<unsigned*>(smem + L.off_q2) + warp * (4 * Q2_CAP);       // i, j, cls, d2
    for (int k = tid; k < T.K * T.K; k += THREADS) { tv[k] = T.tv[k]; th[k] = T.th[k]; }
    int* s_lkey = reinterpret_cast<int*>(smem + L.off_lkey);
    unsigned* bits;
    unsigned gen = 0;
    if (SMALL) {
        for (int k = tid; k < T.n_pair; k += THREADS) s_lkey[k] = T.pair_lkey[k];
        bits = reinterpret_cast<unsigned*>(smem + L.off_bits);
        if (tid == 0) { mbar_init(&M->bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    } else {
        bits = gstamp + (size_t)blockIdx.x * L.bits_words;
        gen = gstate[blockIdx.x];          // generation carried across launches (0 on the first one)
    }
    __syncthreads();
    unsigned phase = 0;

    for (int f = blockIdx.x; f < n_frames; f += gridDim.x) {
        const unsigned char* rec = binned + (size_t)f * B.stride;
        const long long gidx = first_index + f;
        const float4* spos; const float* aux; const int* cells; const BinHeader* hdr;
        if (SMALL) {
            if (tid == 0) {
                const int ncell = __ldcg(&reinterpret_cast<const BinHeader*>(rec)->ncell);
                const unsigned cbytes = (unsigned)((4 * ((T.n_pair_a < T.n_pair ? 2 : 1) * ncell + 1) + 15) & ~15);
                // the previous frame's generic-proxy reads of these buffers are ordered before the async-proxy writes
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_expect_tx(&M->bar, (unsigned)B.fixed_bytes + cbytes);
                bulk_g2s(smem + L.off_rec, rec, (unsigned)B.fixed_bytes, &M->bar);
                bulk_g2s(smem + L.off_cells, rec + B.off_cells, cbytes, &M->bar);
                M->cursor = 0; M->nmark = 0;
            }
            for (int w = tid; w < L.bits_words; w += THREADS) bits[w] = 0u;
            mbar_wait(&M->bar, phase);
            phase ^= 1u;
            hdr = reinterpret_cast<const BinHeader*>(smem + L.off_rec);
            spos = reinterpret_cast<const float4*>(smem + L.off_rec + B.off_spos);
            aux = reinterpret_cast<const float*>(smem + L.off_rec + B.off_aux);
            cells = reinterpret_cast<const int*>(smem + L.off_cells);
        } else {
            if (tid == 0) { M->cursor = 0; M->nmark = 0; }
            gen = gen + 1u;
            if (gen > 255u) {               // generations wrapped: wipe the table once per 255 frames
                for (int w = tid; w < L.bits_words; w += THREADS) bits[w] = 0u;
                gen = 1u;
            }
            hdr = reinterpret_cast<const BinHeader*>(rec);
            spos = reinterpret_cast<const float4*>(rec + B.off_spos);
            aux = reinterpret_cast<const float*>(rec + B.off_aux);
            cells = reinterpret_cast<const int*>(rec + B.off_cells);
        }
        __syncthreads();
        const int px = hdr->nc[0] + 2, pxy = px * (hdr->nc[1] + 2), ncell = hdr->ncell;
        FrameCtx F{spos, SMALL ? s_lkey : T.pair_lkey, aux, bits, gen, tv, th,
                   hpmask + (size_t)(gidx / frag_frames - frag0) * T.n_pair, (int)gidx,
                   water_scratch(wscratch, blockIdx.x, T.n_solv), &M->nmark};
        // ---- R: ring interactions (small lists; float prefilter, float64 geometry) ------------------------------
        if (P.itypes & CHC_IT_PC) {
            const int n = T.n_cat * T.n_rings;
            for (int t = tid; t < n; t += THREADS) ring_pc_task(T, P, O, F, t / T.n_rings, t % T.n_rings);
        }
        if (P.itypes & (CHC_IT_PS | CHC_IT_TS)) {
            const int n = T.n_rings * (T.n_rings - 1) / 2;
            for (int t = tid; t < n; t += THREADS) {
                int r2 = (int)((1.0f + sqrtf(1.0f + 8.0f * (float)t)) * 0.5f);
                while (r2 * (r2 - 1) / 2 > t) --r2;
                while ((r2 + 1) * r2 / 2 <= t) ++r2;
                ring_stack_task(T, P, O, F, t - r2 * (r2 - 1) / 2, r2);
            }
        }
        // ---- E: pair interactions ------------------------------------------------------------------------------
        const float cut2 = P.pre2;            // largest exact float32 cutoff^2 of the enabled pair types
        int q1n = 0, q2n = 0;
        // E1 iteration state (warp-uniform): current cell [i, ce), its j ranges, the j block
        int blk_base = 0; unsigned blk_mask = 0u; int my_cs = 0, my_ce = 0;
        int i = 0, ce = 0, cs = 0, J = 0, jj = 0;
        int rb = 0, rlen = 0;                 // lanes 0..nr-1: begin / length of neighbour range `lane`
        // ranges 0..4: half shell of grid A; ranges 5..13 (only with unselected solvent): the full 3x3 rows of grid B
        const int nr = T.n_pair_a < T.n_pair ? 14 : 5;
        int j = -1; float4 pj = make_float4(0.f, 0.f, 0.f, 0.f);
        bool exhausted = false;
        for (;;) {
            // ---- E1 ----
            while (q1n < 32 && !exhausted) {
                if (i >= ce) {
                    jj += 32;
                    if (jj >= J) {                     // next non-empty cell
                        while (blk_mask == 0u) {
                            int base = 0;
                            if (lane == 0) base = atomicAdd(&M->cursor, CELL_BLK);
                            base = __shfl_sync(FULL, base, 0);
                            if (base >= ncell) { exhausted = true; break; }
                            blk_base = base;
                            const int c = base + lane;
                            const bool mine = lane < CELL_BLK && c < ncell;
                            my_cs = mine ? cells[c] : 0;
                            my_ce = mine ? cells[c + 1] : 0;
                            blk_mask = __ballot_sync(FULL, my_ce > my_cs);
                        }
                        if (exhausted) break;
                        const int b = __ffs(blk_mask) - 1;
                        blk_mask &= blk_mask - 1u;
                        cs = __shfl_sync(FULL, my_cs, b);
                        ce = __shfl_sync(FULL, my_ce, b);
                        const int c = blk_base + b;
                        // half shell: own row from the cell itself to the end of cell x+1, then rows
                        // (y+1,z) (y-1,z+1) (y,z+1) (y+1,z+1) from x-1 to x+1
                        if (lane < 5) {
                            const int dy = (lane == 1 || lane == 4) - (lane == 2), dz = lane >= 2;
                            const int base = c + dy * px + dz * pxy;
                            rb = lane == 0 ? cs : cells[base - 1];
                            rlen = cells[base + 2] - rb;
                        } else if (lane < nr) {
                            const int r = lane - 5;
                            const int base = c + ncell + (r % 3 - 1) * px + (r / 3 - 1) * pxy;
                            rb = cells[base - 1];
                            rlen = cells[base + 2] - rb;
                        }
                        J = 0;
#pragma unroll
                        for (int r = 0; r < 5; ++r) J += __shfl_sync(FULL, rlen, r);
                        for (int r = 5; r < nr; ++r) J += __shfl_sync(FULL, rlen, r);
                        jj = 0;
                    }
                    // this lane's j of the block
                    int k = jj + lane;
                    j = -1;
#pragma unroll
                    for (int r = 0; r < 5; ++r) {
                        const int len = __shfl_sync(FULL, rlen, r), beg = __shfl_sync(FULL, rb, r);
                        if (j < 0 && k >= 0 && k < len) j = beg + k;
                        k -= len;
                    }
                    for (int r = 5; r < nr; ++r) {          // grid B rows (only with unselected solvent)
                        const int len = __shfl_sync(FULL, rlen, r), beg = __shfl_sync(FULL, rb, r);
                        if (j < 0 && k >= 0 && k < len) j = beg + k;
                        k -= len;
                    }
                    if (j >= 0) pj = spos[j];
                    i = cs;
                }
                const float4 pi = spos[i];
                const float d2 = dist2_exact(pi.x, pi.y, pi.z, pj.x, pj.y, pj.z);
                const bool hit = j > i && d2 <= cut2;
                const unsigned bal = __ballot_sync(FULL, hit);
                if (hit) {
                    const int at = q1n + __popc(bal & ((1u << lane) - 1u));
                    q1[at] = (unsigned)i; q1[Q1_CAP + at] = (unsigned)j;
                }
                q1n += __popc(bal);
                ++i;
            }
            if (q1n == 0 && q2n == 0) break;           // only reachable when exhausted
            __syncwarp();
            // ---- E2: up to 32 survivors ----
            {
                const int nb = min(q1n, 32);
                q1n -= nb;
                unsigned cls = 0u; float d2 = 0.f; unsigned ei = 0u, ej = 0u;
                if (lane < nb) {
                    ei = q1[q1n + lane]; ej = q1[Q1_CAP + q1n + lane];
                    cls = e2_classify(T, P, F, spos[ei], spos[ej], d2);
                }
                const unsigned bal = __ballot_sync(FULL, cls != 0u);
                if (cls) {
                    const int at = q2n + __popc(bal & ((1u << lane) - 1u));
                    q2[at] = ei; q2[Q2_CAP + at] = ej; q2[2 * Q2_CAP + at] = cls; q2[3 * Q2_CAP + at] = __float_as_uint(d2);
                }
                q2n += __popc(bal);
                __syncwarp();
            }
            // ---- E3: 32 classified pairs (or the rest when nothing else is left) ----
            if (q2n >= 32 || (exhausted && q1n == 0 && q2n > 0)) {
                const int nb = min(q2n, 32);
                q2n -= nb;
                if (lane < nb)
                    e3_finish(T, P, O, F, (int)q2[q2n + lane], (int)q2[Q2_CAP + q2n + lane], q2[2 * Q2_CAP + q2n + lane],
                              __uint_as_float(q2[3 * Q2_CAP + q2n + lane]));
                __syncwarp();
            }
        }
        __syncthreads();
        // ---- W: water bridges from the (non-solvent, solvent) hydrogen bonds the pair phase recorded ----------------
        if (T.n_solv > 0 && (P.itypes & CHC_IT_HB)) {
            WaterScratch W = water_scratch(wscratch, blockIdx.x, T.n_solv);
            water_bridges(T, P, O, F, W, cells, px, pxy, ncell, hdr, &M->nmark, tid, THREADS);
            __syncthreads();
        }
    }
    if (!SMALL && tid == 0) gstate[blockIdx.x] = gen;
}
