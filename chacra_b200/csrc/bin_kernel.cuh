// bin_kernel.cuh -- the coordinate-streaming kernel: one CTA per frame reads the frame's float32 coordinates
// from HBM and writes its "binned" record (BinHeader | cell-sorted float4 pair atoms | aux atoms | cell starts).
//
// This replaces, per frame, what the reference asks VMD to do through `atomselect` + `measure contacts`
// (uniform-grid pair search): the grid is built ONCE per frame here and shared by every
// interaction type in pair_kernel.cuh.  Phases (cells live in shared memory, counting sort):
//   1  bounding box of the pair atoms             (gather through T.pair_atom, block reduce)
//   2  per-cell counts                            (shared-memory atomics)
//   3  exclusive scan                             (block scan)
//   4  scatter to the sorted float4 array in HBM/L2, aux atoms, cell starts, header
//   5  only on the first frame of a 100-frame fragment: hydrophobic carbon membership (hydrophobics.py:8-41),
//      a 1.7 A radius query on the freshly built grid.
// Algorithmic HBM traffic: 12 B x n_atoms read per frame (every 32-byte sector of the frame is touched once:
// heavy atoms and hydrogens interleave), plus the binned record written once.
#pragma once
#include "frame_kernel.cuh"

__device__ __forceinline__ int cell_of(float x, float origin, float inv, int n) {
    int c = (int)((x - origin) * inv);
    return min(max(c, 0), n - 1);
}

struct BinShared {
    int bb_min[3], bb_max[3];
    int scan_carry[33];
};

template <int THREADS>
__global__ void __launch_bounds__(THREADS)
bin_frames_kernel(DevTopo T, DevParams P, BinLayout B, const float* __restrict__ xyz, long long frame_stride, int n_frames,
                  long long first_index, int frag_frames, long long frag0, unsigned char* __restrict__ binned,
                  unsigned char* __restrict__ hpmask) {
    extern __shared__ __align__(16) int cells[];          // [B.max_cells]
    __shared__ BinShared M;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int NP = T.n_pair;
    for (int f = blockIdx.x; f < n_frames; f += gridDim.x) {
        const float* __restrict__ X = xyz + (size_t)f * frame_stride;
        unsigned char* rec = binned + (size_t)f * B.stride;
        float4* spos = reinterpret_cast<float4*>(rec + B.off_spos);
        float* aux = reinterpret_cast<float*>(rec + B.off_aux);
        int* cells_out = reinterpret_cast<int*>(rec + B.off_cells);
        if (tid < 3) { M.bb_min[tid] = 0x7fffffff; M.bb_max[tid] = (int)0x80000000; }
        __syncthreads();
        // ---- 1: bounding box --------------------------------------------------------------------------
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
                if (lane == 0) { atomicMin(&M.bb_min[d], a); atomicMax(&M.bb_max[d], b); }
            }
        }
        __syncthreads();
        float org[3], inv; int nc[3];
        {
            float ext[3];
#pragma unroll
            for (int d = 0; d < 3; ++d) { org[d] = ord2f(M.bb_min[d]); ext[d] = ord2f(M.bb_max[d]) - org[d]; }
            float cell = P.cell;
            for (;;) {                    // the cell edge only ever grows: every pair within the cutoff stays in adjacent cells
                inv = 1.0f / cell;
#pragma unroll
                for (int d = 0; d < 3; ++d) nc[d] = (int)(ext[d] * inv) + 1;
                if ((long long)(nc[0] + 2) * (nc[1] + 2) * (nc[2] + 2) + 1 <= B.max_cells) break;
                cell *= 1.25f;
            }
        }
        const int px = nc[0] + 2, pxy = px * (nc[1] + 2), ncell = pxy * (nc[2] + 2);
        for (int c = tid; c < ncell; c += THREADS) cells[c] = 0;
        __syncthreads();
        // ---- 2: count ---------------------------------------------------------------------------------
        for (int p = tid; p < NP; p += THREADS) {
            const float* q = X + 3 * (size_t)T.pair_atom[p];
            int c = (cell_of(q[2], org[2], inv, nc[2]) + 1) * pxy + (cell_of(q[1], org[1], inv, nc[1]) + 1) * px +
                    cell_of(q[0], org[0], inv, nc[0]) + 1;
            atomicAdd(&cells[c], 1);
        }
        __syncthreads();
        // ---- 3: exclusive scan over the cells --------------------------------------------------------------
        {
            const int per = (ncell + THREADS - 1) / THREADS;
            const int b = min(tid * per, ncell), e = min(b + per, ncell);
            int s = 0;
            for (int c = b; c < e; ++c) s += cells[c];
            int incl = s;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
            if (lane == 31) M.scan_carry[warp] = incl;
            __syncthreads();
            if (warp == 0) {
                int v = lane < (THREADS >> 5) ? M.scan_carry[lane] : 0, w = v;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(FULL, w, o); if (lane >= o) w += t; }
                M.scan_carry[lane] = w - v;
            }
            __syncthreads();
            int run = M.scan_carry[warp] + incl - s;
            for (int c = b; c < e; ++c) { int v = cells[c]; cells[c] = run; run += v; }
        }
        __syncthreads();
        // ---- 4: scatter (cells[c] becomes the END of cell c), aux atoms, header ----------------------------------
        for (int p = tid; p < NP; p += THREADS) {
            const float* q = X + 3 * (size_t)T.pair_atom[p];
            const float x = q[0], y = q[1], z = q[2];
            int c = (cell_of(z, org[2], inv, nc[2]) + 1) * pxy + (cell_of(y, org[1], inv, nc[1]) + 1) * px +
                    cell_of(x, org[0], inv, nc[0]) + 1;
            int dst = atomicAdd(&cells[c], 1);
            spos[dst] = make_float4(x, y, z, __uint_as_float((T.pair_info[p] & I_WMASK) | ((unsigned)p << I_PIDX_SHIFT)));
        }
        for (int s = tid; s < T.n_aux; s += THREADS) {
            const float* q = X + 3 * (size_t)T.aux_atom[s];
            aux[3 * s] = q[0]; aux[3 * s + 1] = q[1]; aux[3 * s + 2] = q[2];
        }
        if (tid == 0) {
            BinHeader h;
            h.org[0] = org[0]; h.org[1] = org[1]; h.org[2] = org[2]; h.inv = inv;
            h.nc[0] = nc[0]; h.nc[1] = nc[1]; h.nc[2] = nc[2]; h.ncell = ncell;
            *reinterpret_cast<BinHeader*>(rec) = h;
            cells_out[0] = 0;
        }
        __syncthreads();
        for (int c = tid; c < ncell; c += THREADS) cells_out[c + 1] = cells[c];
        // ---- 5: hydrophobic carbon sets of the fragment this frame opens ------------------------------------------
        const long long gidx = first_index + f;
        if ((P.itypes & CHC_IT_HP) && gidx % frag_frames == 0) {
            unsigned char* hm = hpmask + (size_t)(gidx / frag_frames - frag0) * NP;
            const float r2 = 1.7f * 1.7f;
            for (int b = tid; b < T.n_hp_base; b += THREADS) {
                const int p = T.hp_base[b];
                const float* q = X + 3 * (size_t)T.pair_atom[p];
                const float x = q[0], y = q[1], z = q[2];
                const int c0 = (cell_of(z, org[2], inv, nc[2]) + 1) * pxy + (cell_of(y, org[1], inv, nc[1]) + 1) * px +
                               cell_of(x, org[0], inv, nc[0]) + 1;
                bool other = false;       // an atom that is neither carbon nor hydrogen within 1.7 A
                for (int dz = -1; dz <= 1; ++dz)
                    for (int dy = -1; dy <= 1; ++dy) {
                        const int base = c0 + dy * px + dz * pxy;
                        const int jb = base >= 2 ? cells[base - 2] : 0, je = cells[base + 1];   // cells[] holds ENDS: start of base-1 .. end of base+1
                        for (int j = jb; j < je; ++j) {
                            const float4 o = __ldcg(spos + j);
                            if (dist2_exact(o.x, o.y, o.z, x, y, z) <= r2 && !((P.elem_ch_mask >> (__float_as_uint(o.w) & I_ELEM)) & 1u))
                                other = true;
                        }
                    }
                unsigned sub = other ? 0u : 1u;
                unsigned eq = 0u;
                if (!T.same && sub) {
                    // sele2 wants the element set to EQUAL {C, H} (hydrophobics.py:39): a hydrogen must be in range too.
                    // Hydrogens are not in the grid; rare path (ChACRA always passes equal selections): linear scan.
                    bool hyd = false;
                    for (int a = 0; a < T.n_atoms && !hyd; ++a)
                        if (T.atom_class[a] == 2) {
                            const float* h = X + 3 * (size_t)a;
                            hyd = dist2_exact(h[0], h[1], h[2], x, y, z) <= r2;
                        }
                    eq = hyd ? 2u : 0u;
                }
                hm[p] = (unsigned char)(sub | eq);
            }
        }
        __syncthreads();
    }
}
