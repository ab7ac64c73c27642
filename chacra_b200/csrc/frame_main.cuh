// frame_main.cuh -- the persistent one-CTA-per-frame contact kernel (see frame_kernel.cuh for the phases).
#pragma once
#include "frame_phases.cuh"

struct WarpQueues {
    unsigned* q1;          // prefilter survivors: (i << 16) | j, sorted slots
    unsigned* q2ij; unsigned* q2cls; float* q2d;
    int qn, q2n, lane;

    __device__ __forceinline__ void e3_batch(const DevTopo& T, const DevParams& P, const DevOut& O, const FrameCtx& F,
                                             int first, int n) {
        if (lane < n) {
            unsigned e = q2ij[first + lane];
            e3_finish(T, P, O, F, e >> 16, e & 0xffff, q2cls[first + lane], q2d[first + lane]);
        }
        __syncwarp();
    }
    __device__ __forceinline__ void e2_batch(const DevTopo& T, const DevParams& P, const DevOut& O, const FrameCtx& F,
                                             int first, int n) {
        unsigned e = 0, cls = 0; float d2 = 0.f;
        if (lane < n) {
            e = q1[first + lane];
            cls = e2_classify(T, P, F, e >> 16, e & 0xffff, d2);
        }
        unsigned bal = __ballot_sync(FULL, cls != 0);
        if (cls) {
            int idx = q2n + __popc(bal & ((1u << lane) - 1));
            q2ij[idx] = e; q2cls[idx] = cls; q2d[idx] = d2;
        }
        q2n += __popc(bal);
        __syncwarp();
        if (q2n >= 32) { q2n -= 32; e3_batch(T, P, O, F, q2n, 32); }
    }
    __device__ __forceinline__ void drain_q1(const DevTopo& T, const DevParams& P, const DevOut& O, const FrameCtx& F) {
        while (qn >= 32) { qn -= 32; e2_batch(T, P, O, F, qn, 32); }
        if (qn > 0) { int n = qn; qn = 0; e2_batch(T, P, O, F, 0, n); }
    }
    __device__ __forceinline__ void drain_all(const DevTopo& T, const DevParams& P, const DevOut& O, const FrameCtx& F) {
        drain_q1(T, P, O, F);
        if (q2n > 0) { int n = q2n; q2n = 0; e3_batch(T, P, O, F, 0, n); }
    }
    // warp-uniform call: every lane brings the survivor bitmask of its atom i against slots jbase..jbase+31
    __device__ __forceinline__ void push(const DevTopo& T, const DevParams& P, const DevOut& O, const FrameCtx& F,
                                         unsigned mask, int i, int jbase) {
        int cnt = __popc(mask), incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
        int total = __shfl_sync(FULL, incl, 31);
        if (total == 0) return;
        if (qn + total > Q1_CAP) drain_q1(T, P, O, F);
        if (total > Q1_CAP) {           // very dense neighbourhood: every lane finishes its own survivors
            while (mask) {
                int k = __ffs(mask) - 1; mask &= mask - 1;
                float d2; unsigned cls = e2_classify(T, P, F, i, jbase + k, d2);
                if (cls) e3_finish(T, P, O, F, i, jbase + k, cls, d2);
            }
            __syncwarp();
            return;
        }
        int pos = qn + incl - cnt;
        while (mask) { int k = __ffs(mask) - 1; mask &= mask - 1; q1[pos++] = ((unsigned)i << 16) | (unsigned)(jbase + k); }
        qn += total;
        __syncwarp();
        while (qn >= 32) { qn -= 32; e2_batch(T, P, O, F, qn, 32); }
    }
};

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
    WarpQueues Q;
    Q.q1 = reinterpret_cast<unsigned*>(smem + L.off_q1) + warp * Q1_CAP;
    Q.q2ij = reinterpret_cast<unsigned*>(smem + L.off_q2) + warp * (3 * Q2_CAP);
    Q.q2cls = Q.q2ij + Q2_CAP;
    Q.q2d = reinterpret_cast<float*>(Q.q2cls + Q2_CAP);
    Q.lane = lane;
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
        Q.qn = 0; Q.q2n = 0;
        const float pre2 = P.pre2;
        for (;;) {
            int chunk = 0;
            if (lane == 0) chunk = atomicAdd(&M->chunk, 1);
            chunk = __shfl_sync(FULL, chunk, 0);
            const int i = chunk * 32 + lane;
            if (chunk * 32 >= NP) break;
            const bool valid = i < NP;
            float4 me = valid ? spos[i] : make_float4(0.f, 0.f, 0.f, 0.f);
            int rb[5], re[5];
            if (valid) {
                int c = (cell_of(me.z, org[2], inv, nc[2]) + 1) * pxy + (cell_of(me.y, org[1], inv, nc[1]) + 1) * px +
                        cell_of(me.x, org[0], inv, nc[0]) + 1;
                rb[0] = i + 1;                 re[0] = cells[c + 2];
                rb[1] = cells[c + px - 1];     re[1] = cells[c + px + 2];
                int u = c + pxy;
                rb[2] = cells[u - px - 1];     re[2] = cells[u - px + 2];
                rb[3] = cells[u - 1];          re[3] = cells[u + 2];
                rb[4] = cells[u + px - 1];     re[4] = cells[u + px + 2];
            } else {
#pragma unroll
                for (int r = 0; r < 5; ++r) { rb[r] = 0; re[r] = 0; }
            }
#pragma unroll
            for (int r = 0; r < 5; ++r) {
                const int len = re[r] - rb[r];
                const int maxlen = __reduce_max_sync(FULL, len);
                for (int off = 0; off < maxlen; off += 32) {
                    const int jb = rb[r] + off;
                    const int n = min(32, len - off);
                    unsigned mask = 0;
                    for (int k = 0; k < n; ++k) {
                        float4 o = spos[jb + k];
                        float dx = o.x - me.x, dy = o.y - me.y, dz = o.z - me.z;
                        float d2 = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, dx * dx));
                        mask |= (d2 <= pre2 ? 1u : 0u) << k;
                    }
                    Q.push(T, P, O, F, mask, i, jb);
                }
            }
        }
        Q.drain_all(T, P, O, F);
        __syncthreads();
    }
}
