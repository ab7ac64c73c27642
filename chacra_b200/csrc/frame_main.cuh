// frame_main.cuh -- the persistent one-CTA-per-frame contact kernel (see frame_kernel.cuh for the phases).
//
// Code-size note: the first version inlined E2/E3 at ~45 sites (36k SASS instructions) and spent 82 % of
// its cycles in instruction-fetch stalls (profiles/r01_ncu_v0_summary.txt).  The pair phase is therefore
// written as one loop nest in which the prefilter, the exact stage and the emit stage each appear at
// exactly ONE place; the "final" pass of the outer loop only flushes the queues.
#pragma once
#include "frame_phases.cuh"

#define LQ_CAP 48          // per-lane survivor slots (lane-strided, conflict-free)

__device__ __forceinline__ int cell_of(float x, float origin, float inv, int n) {
    int c = (int)((x - origin) * inv);
    return min(max(c, 0), n - 1);
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS, 1)
frame_contacts_kernel(DevTopo T, DevParams P, DevOut O, SmemLayout L, const float* __restrict__ xyz,
                      long long frame_stride, int n_frames, long long first_index, int frag_frames, long long frag0,
                      const unsigned char* __restrict__ hpmask, unsigned* gbits) {
    extern __shared__ __align__(16) unsigned char smem[];
    float4* spos = reinterpret_cast<float4*>(smem + L.off_spos);
    int* slk = reinterpret_cast<int*>(smem + L.off_slk);
    unsigned* tmp = reinterpret_cast<unsigned*>(smem + L.off_tmp);
    float* aux = reinterpret_cast<float*>(smem + L.off_aux);
    int* cells = reinterpret_cast<int*>(smem + L.off_cells);
    float* tv = reinterpret_cast<float*>(smem + L.off_tv);
    float* th = reinterpret_cast<float*>(smem + L.off_th);
    FrameMisc* M = reinterpret_cast<FrameMisc*>(smem + L.off_misc);
    unsigned* bits = L.bits_in_smem ? reinterpret_cast<unsigned*>(smem + L.off_bits)
                                    : gbits + (size_t)blockIdx.x * L.bits_words;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int NP = T.n_pair;
    unsigned short* lq = reinterpret_cast<unsigned short*>(smem + L.off_q1) + warp * (LQ_CAP * 32) + lane;
    unsigned* q2ij = reinterpret_cast<unsigned*>(smem + L.off_q2) + warp * (3 * Q2_CAP);
    unsigned* q2cls = q2ij + Q2_CAP;
    float* q2d = reinterpret_cast<float*>(q2cls + Q2_CAP);
    for (int k = tid; k < T.K * T.K; k += THREADS) { tv[k] = T.tv[k]; th[k] = T.th[k]; }

    for (int f = blockIdx.x; f < n_frames; f += gridDim.x) {
        const float* __restrict__ X = xyz + (size_t)f * frame_stride;
        const long long gidx = first_index + f;
        const unsigned char* __restrict__ hm = hpmask + (size_t)(gidx / frag_frames - frag0) * NP;
        if (tid < 3) { M->bb_min[tid] = 0x7fffffff; M->bb_max[tid] = (int)0x80000000; }
        if (tid == 0) { M->chunk = 0; }
        __syncthreads();
        // ---- A: bounding box ----------------------------------------------------------------------
        {
            float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
            for (int p = tid; p < NP; p += THREADS) {
                const float* q = X + 3 * (size_t)T.pair_atom[p];
#pragma unroll
                for (int d = 0; d < 3; ++d) { float v = q[d]; lo[d] = fminf(lo[d], v); hi[d] = fmaxf(hi[d], v); }
            }
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                int a = __reduce_min_sync(FULL, f2ord(lo[d])), b = __reduce_max_sync(FULL, f2ord(hi[d]));
                if (lane == 0) { atomicMin(&M->bb_min[d], a); atomicMax(&M->bb_max[d], b); }
            }
        }
        __syncthreads();
        float org[3], inv; int nc[3];
        {
            float ext[3];
#pragma unroll
            for (int d = 0; d < 3; ++d) { org[d] = ord2f(M->bb_min[d]); ext[d] = ord2f(M->bb_max[d]) - org[d]; }
            float cell = P.cell;
            for (;;) {
                inv = 1.0f / cell;
#pragma unroll
                for (int d = 0; d < 3; ++d) nc[d] = (int)(ext[d] * inv) + 1;
                if ((long long)(nc[0] + 2) * (nc[1] + 2) * (nc[2] + 2) + 2 <= L.max_cells) break;
                cell *= 1.25f;
            }
        }
        const int px = nc[0] + 2, pxy = px * (nc[1] + 2), ncell = pxy * (nc[2] + 2);
        for (int c = tid; c <= ncell; c += THREADS) cells[c] = 0;
        __syncthreads();
        // ---- B: count -------------------------------------------------------------------------------
        for (int p = tid; p < NP; p += THREADS) {
            const float* q = X + 3 * (size_t)T.pair_atom[p];
            int c = (cell_of(q[2], org[2], inv, nc[2]) + 1) * pxy + (cell_of(q[1], org[1], inv, nc[1]) + 1) * px +
                    cell_of(q[0], org[0], inv, nc[0]) + 1;
            unsigned rank = atomicAdd(&cells[c], 1);
            tmp[p] = (rank << 14) | (unsigned)c;
        }
        __syncthreads();
        // ---- C: exclusive scan over the cells -------------------------------------------------------------
        {
            const int per = (ncell + 1 + THREADS - 1) / THREADS;
            const int b = min(tid * per, ncell + 1), e = min(b + per, ncell + 1);
            int s = 0;
            for (int c = b; c < e; ++c) s += cells[c];
            int incl = s;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
            if (lane == 31) M->scan_carry[warp] = incl;
            __syncthreads();
            if (warp == 0) {
                int v = lane < (THREADS >> 5) ? M->scan_carry[lane] : 0, w = v;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(FULL, w, o); if (lane >= o) w += t; }
                M->scan_carry[lane] = w - v;
            }
            __syncthreads();
            int run = M->scan_carry[warp] + incl - s;
            for (int c = b; c < e; ++c) { int v = cells[c]; cells[c] = run; run += v; }
        }
        __syncthreads();
        // ---- D: scatter, aux copy, clear the residue-pair set ---------------------------------------------
        for (int p = tid; p < NP; p += THREADS) {
            const float* q = X + 3 * (size_t)T.pair_atom[p];
            unsigned t = tmp[p];
            int dst = cells[t & 0x3fff] + (int)(t >> 14);
            unsigned info = T.pair_info[p];
            unsigned h = hm[p];      // bit0: in H1, bit1: in H2 for this fragment
            info = (info & ~(I_HP1 | I_HP2)) | ((info & I_HP1) && (h & 1u) ? I_HP1 : 0u) | ((info & I_HP2) && (h & 2u) ? I_HP2 : 0u);
            spos[dst] = make_float4(q[0], q[1], q[2], __uint_as_float((info & ((1u << I_PIDX_SHIFT) - 1)) | ((unsigned)p << I_PIDX_SHIFT)));
            slk[dst] = T.pair_lkey[p];
        }
        for (int s = tid; s < T.n_aux; s += THREADS) {
            const float* q = X + 3 * (size_t)T.aux_atom[s];
            aux[3 * s] = q[0]; aux[3 * s + 1] = q[1]; aux[3 * s + 2] = q[2];
        }
        for (int w = tid; w < L.bits_words; w += THREADS) bits[w] = 0u;
        __syncthreads();
        // ---- R: ring interactions ----------------------------------------------------------------------
        FrameCtx F{spos, slk, aux, bits, tv, th, (int)gidx};
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
        // ---- E: pair interactions ------------------------------------------------------------------------
        const float pre2 = P.pre2;
        int q2n = 0;
        bool final_pass = false;
        do {
            int chunk = 0;
            if (lane == 0) chunk = atomicAdd(&M->chunk, 1);
            chunk = __shfl_sync(FULL, chunk, 0);
            final_pass = chunk * 32 >= NP;
            const int i = chunk * 32 + lane;
            const bool valid = !final_pass && i < NP;
            const float4 me = valid ? spos[i] : make_float4(0.f, 0.f, 0.f, 0.f);
            const int lki = valid ? slk[i] : 0;
            int c0 = 0;
            if (valid)
                c0 = (cell_of(me.z, org[2], inv, nc[2]) + 1) * pxy + (cell_of(me.y, org[1], inv, nc[1]) + 1) * px +
                     cell_of(me.x, org[0], inv, nc[0]) + 1;
            int cnt = 0;
            int r = 0;
#pragma unroll 1
            do {
                // half shell: own row from i+1 to the end of cell x+1, then rows (y+1,z) (y-1,z+1) (y,z+1) (y+1,z+1)
                int jb = 0, len = 0;
                if (valid) {
                    const int dy = (r == 1 || r == 4) - (r == 2), dz = r >= 2;
                    const int base = c0 + dy * px + dz * pxy;
                    jb = r == 0 ? i + 1 : cells[base - 1];
                    len = cells[base + 2] - jb;
                }
                const int maxlen = __reduce_max_sync(FULL, len);
                int off = 0;
#pragma unroll 1
                do {
                    // ---- E1: fused-multiply prefilter, survivors into the lane's own slots
                    const int n = min(32, len - off);
                    const float4* pj = spos + jb + off;
#pragma unroll 4
                    for (int k = 0; k < n; ++k) {
                        const float4 o = pj[k];
                        const float dx = o.x - me.x, dy = o.y - me.y, dz = o.z - me.z;
                        const float d2 = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, dx * dx));
                        if (d2 <= pre2) { lq[cnt * 32] = (unsigned short)(jb + off + k); ++cnt; }
                    }
                    off += 32;
                    const bool last = off >= maxlen && r == 4;
                    if (last || __any_sync(FULL, cnt > LQ_CAP - 32)) {
                        // ---- E2: exact float32 d2 + class tests, one row of survivors per iteration
                        const int rows = __reduce_max_sync(FULL, cnt);
                        const bool flush = final_pass && last;
#pragma unroll 1
                        for (int s = 0; s < rows || (flush && s == 0); ++s) {
                            unsigned cls = 0; float d2 = 0.f; int j = 0;
                            if (s < cnt) { j = lq[s * 32]; cls = e2_classify(T, P, F, me, lki, j, d2); }
                            const unsigned bal = __ballot_sync(FULL, cls != 0);
                            if (cls) {
                                const int idx = q2n + __popc(bal & ((1u << lane) - 1));
                                q2ij[idx] = ((unsigned)i << 16) | (unsigned)j; q2cls[idx] = cls; q2d[idx] = d2;
                            }
                            q2n += __popc(bal);
                            __syncwarp();
                            // ---- E3: bonded / disulfide / hydrogen-angle tests, dedup, counts, records
                            if (q2n >= 32 || (flush && s + 1 >= rows && q2n > 0)) {
                                const int nb = min(q2n, 32);
                                q2n -= nb;
                                if (lane < nb) {
                                    const unsigned e = q2ij[q2n + lane];
                                    e3_finish(T, P, O, F, e >> 16, e & 0xffff, q2cls[q2n + lane], q2d[q2n + lane]);
                                }
                                __syncwarp();
                            }
                        }
                        cnt = 0;
                    }
                } while (off < maxlen);
                ++r;
            } while (r < 5);
        } while (!final_pass);
        __syncthreads();
    }
}
