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
// hp_mask_kernel (below) then runs a 1.7 A radius query on the grid of every frame that opens a 100-frame
// fragment: the hydrophobic carbon membership of that fragment (hydrophobics.py:8-41).
// Algorithmic HBM traffic: 12 B x n_atoms read per frame (every 32-byte sector of the frame is touched once:
// heavy atoms and hydrogens interleave), plus the binned record written once.
#pragma once
#include "frame_kernel.cuh"

struct BinShared {
    int bb_min[3], bb_max[3];
    int scan_carry[33];
};

// APT > 0: every thread keeps its (at most APT) pair atoms' coordinates in registers, so a frame is read from HBM
// exactly once with 3*APT independent loads in flight per thread; APT == 0: three gather passes (large systems; the
// second and third pass hit L2).
template <int THREADS, int APT, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
bin_frames_kernel(DevTopo T, DevParams P, BinLayout B, const float* __restrict__ xyz, long long frame_stride, int n_frames,
                  long long first_index, int frag_frames, long long frag0, unsigned char* __restrict__ binned,
                  unsigned char* __restrict__ hpmask, const long long* __restrict__ src_index) {
    extern __shared__ __align__(16) int cells[];          // [B.max_cells]
    __shared__ BinShared M;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int NP = T.n_pair;
    // grid B: pair atoms [n_pair_a, NP) (unselected solvent) are sorted into a second set of cells, keys ncell..2*ncell-1
    const bool gridb = T.n_pair_a < NP;
    constexpr int NR = APT > 0 ? APT : 1;
    int r_atom[NR];
    float rx[NR], ry[NR], rz[NR];
    if (APT > 0) {
#pragma unroll
        for (int k = 0; k < NR; ++k) {
            const int p = tid + k * THREADS;
            r_atom[k] = p < NP ? T.pair_atom[p] : -1;
        }
    }
    for (int f = blockIdx.x; f < n_frames; f += gridDim.x) {
        const float* __restrict__ X = frame_ptr(xyz, src_index, f, frame_stride);
        unsigned char* rec = binned + (size_t)f * B.stride;
        float4* spos = reinterpret_cast<float4*>(rec + B.off_spos);
        float* aux = reinterpret_cast<float*>(rec + B.off_aux);
        int* cells_out = reinterpret_cast<int*>(rec + B.off_cells);
        if (tid < 3) { M.bb_min[tid] = 0x7fffffff; M.bb_max[tid] = (int)0x80000000; }
        __syncthreads();
        // ---- 1: bounding box --------------------------------------------------------------------------
        {
            float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
            if (APT > 0) {
#pragma unroll
                for (int k = 0; k < NR; ++k)
                    if (r_atom[k] >= 0) {
                        const float* q = X + 3 * (size_t)r_atom[k];
                        rx[k] = q[0]; ry[k] = q[1]; rz[k] = q[2];
                    }
#pragma unroll
                for (int k = 0; k < NR; ++k)
                    if (r_atom[k] >= 0) {
                        lo[0] = fminf(lo[0], rx[k]); hi[0] = fmaxf(hi[0], rx[k]);
                        lo[1] = fminf(lo[1], ry[k]); hi[1] = fmaxf(hi[1], ry[k]);
                        lo[2] = fminf(lo[2], rz[k]); hi[2] = fmaxf(hi[2], rz[k]);
                    }
            } else {
#pragma unroll 4
                for (int p = tid; p < NP; p += THREADS) {
                    const float* q = X + 3 * (size_t)T.pair_atom[p];
#pragma unroll
                    for (int d = 0; d < 3; ++d) { float v = q[d]; lo[d] = fminf(lo[d], v); hi[d] = fmaxf(hi[d], v); }
                }
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
            grid_dims(ext, P.cell, B.max_cells, gridb, inv, nc);
        }
        const int px = nc[0] + 2, pxy = px * (nc[1] + 2), ncell = pxy * (nc[2] + 2);
        const int nkey = gridb ? 2 * ncell : ncell;
        // gather variant on systems below 65 536 pair atoms: two 16-bit counters per word (twice the cells in the same shared
        // memory: the cells stay at the cutoff size for up to ~65 k pair atoms); no half can overflow, every value is <= n_pair
        const bool pk = APT == 0 && B.count16 != 0;
        for (int c = tid; c < (pk ? (nkey + 1) >> 1 : nkey); c += THREADS) cells[c] = 0;
        __syncthreads();
        // ---- 2: count ---------------------------------------------------------------------------------
        int r_cell[NR];
        if (APT > 0) {
#pragma unroll
            for (int k = 0; k < NR; ++k)
                if (r_atom[k] >= 0) {
                    r_cell[k] = (cell_of(rz[k], org[2], inv, nc[2]) + 1) * pxy + (cell_of(ry[k], org[1], inv, nc[1]) + 1) * px +
                                cell_of(rx[k], org[0], inv, nc[0]) + 1 + (tid + k * THREADS >= T.n_pair_a ? ncell : 0);
                    atomicAdd(&cells[r_cell[k]], 1);
                }
        } else {
#pragma unroll 4
            for (int p = tid; p < NP; p += THREADS) {
                const float* q = X + 3 * (size_t)T.pair_atom[p];
                int c = (cell_of(q[2], org[2], inv, nc[2]) + 1) * pxy + (cell_of(q[1], org[1], inv, nc[1]) + 1) * px +
                        cell_of(q[0], org[0], inv, nc[0]) + 1 + (p >= T.n_pair_a ? ncell : 0);
                if (pk) atomicAdd(&cells[c >> 1], 1 << ((c & 1) << 4)); else atomicAdd(&cells[c], 1);
            }
        }
        __syncthreads();
        // ---- 3: exclusive scan over the cells --------------------------------------------------------------
        {
            const int nscan = pk ? (nkey + 1) >> 1 : nkey;            // words to scan
            const int per = (nscan + THREADS - 1) / THREADS;
            const int b = min(tid * per, nscan), e = min(b + per, nscan);
            int s = 0;
            if (pk) for (int c = b; c < e; ++c) { const unsigned w = (unsigned)cells[c]; s += (int)((w & 0xffffu) + (w >> 16)); }
            else for (int c = b; c < e; ++c) s += cells[c];
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
            if (pk) for (int c = b; c < e; ++c) {
                const unsigned w = (unsigned)cells[c], lo2 = w & 0xffffu;
                cells[c] = (int)((unsigned)run | (((unsigned)run + lo2) << 16));
                run += (int)(lo2 + (w >> 16));
            }
            else for (int c = b; c < e; ++c) { int v = cells[c]; cells[c] = run; run += v; }
        }
        __syncthreads();
        // ---- 4: scatter (cells[c] becomes the END of cell c), aux atoms, header ----------------------------------
        if (APT > 0) {
#pragma unroll
            for (int k = 0; k < NR; ++k)
                if (r_atom[k] >= 0) {
                    const int p = tid + k * THREADS;
                    const int dst = atomicAdd(&cells[r_cell[k]], 1);
                    spos[dst] = make_float4(rx[k], ry[k], rz[k], __uint_as_float((T.pair_info[p] & I_WMASK) | ((unsigned)p << I_PIDX_SHIFT)));
                }
        } else {
#pragma unroll 4
            for (int p = tid; p < NP; p += THREADS) {
                const float* q = X + 3 * (size_t)T.pair_atom[p];
                const float x = q[0], y = q[1], z = q[2];
                int c = (cell_of(z, org[2], inv, nc[2]) + 1) * pxy + (cell_of(y, org[1], inv, nc[1]) + 1) * px +
                        cell_of(x, org[0], inv, nc[0]) + 1 + (p >= T.n_pair_a ? ncell : 0);
                int dst;
                if (pk) { const int sh = (c & 1) << 4; dst = (atomicAdd(&cells[c >> 1], 1 << sh) >> sh) & 0xffff; }
                else dst = atomicAdd(&cells[c], 1);
                spos[dst] = make_float4(x, y, z, __uint_as_float((T.pair_info[p] & I_WMASK) | ((unsigned)p << I_PIDX_SHIFT)));
            }
        }
#pragma unroll 4
        for (int s = tid; s < T.n_aux; s += THREADS) {
            const float* q = X + 3 * (size_t)T.aux_atom[s];
            aux[3 * s] = q[0]; aux[3 * s + 1] = q[1]; aux[3 * s + 2] = q[2];
        }
        if (tid == 0) {
            BinHeader h;
            h.org[0] = org[0]; h.org[1] = org[1]; h.org[2] = org[2]; h.inv = inv;
            h.nc[0] = nc[0]; h.nc[1] = nc[1]; h.nc[2] = nc[2]; h.ncell = ncell;
            *reinterpret_cast<BinHeader*>(rec) = h;
            if (B.cell16) reinterpret_cast<unsigned short*>(cells_out)[0] = 0; else cells_out[0] = 0;
        }
        __syncthreads();
        if (pk) {
            for (int c = tid; c < nkey; c += THREADS) {
                const int v = (int)(((unsigned)cells[c >> 1] >> ((c & 1) << 4)) & 0xffffu);
                if (B.cell16) reinterpret_cast<unsigned short*>(cells_out)[c + 1] = (unsigned short)v; else cells_out[c + 1] = v;
            }
        }
        else if (B.cell16) for (int c = tid; c < nkey; c += THREADS) reinterpret_cast<unsigned short*>(cells_out)[c + 1] = (unsigned short)cells[c];
        else for (int c = tid; c < nkey; c += THREADS) cells_out[c + 1] = cells[c];
        __syncthreads();
    }
}

// ---- hydrophobic carbon sets -------------------------------------------------------------------------------
// The reference recomputes, for every 100-frame fragment, which carbons of ALA PHE GLY ILE LEU PRO VAL TRP (or ligand)
// have only C/H atoms within 1.7 A in that fragment's FIRST frame (hydrophobics.py:8-41); sele1 uses a
// subset test, sele2 an equality test against {C, H}.  Every atom that is neither C nor H by element is a pair atom
// (topology.py), so the subset test is a radius query on the frame's grid: one warp per base carbon, lanes over the
// nine 3-cell rows around it.  blockIdx.x = k-th fragment-opening frame of the chunk.
__global__ void __launch_bounds__(256)
hp_mask_kernel(DevTopo T, DevParams P, BinLayout B, const float* __restrict__ xyz, long long frame_stride, int n_frames,
               long long first_index, int frag_frames, long long frag0, const unsigned char* __restrict__ binned,
               unsigned char* __restrict__ hpmask, const long long* __restrict__ src_index) {
    const int f0 = (int)((frag_frames - first_index % frag_frames) % frag_frames);
    const long long f = f0 + (long long)blockIdx.x * frag_frames;
    if (f >= n_frames) return;
    const long long gidx = first_index + f;
    const float* __restrict__ X = frame_ptr(xyz, src_index, f, frame_stride);
    const unsigned char* rec = binned + (size_t)f * B.stride;
    const BinHeader h = *reinterpret_cast<const BinHeader*>(rec);
    const float4* spos = reinterpret_cast<const float4*>(rec + B.off_spos);
    const int* cells32 = reinterpret_cast<const int*>(rec + B.off_cells);
    const unsigned short* cells16 = reinterpret_cast<const unsigned short*>(rec + B.off_cells);
    const bool c16 = B.cell16 != 0;
    unsigned char* hm = hpmask + (size_t)(gidx / frag_frames - frag0) * T.n_pair;
    const int px = h.nc[0] + 2, pxy = px * (h.nc[1] + 2);
    const int lane = threadIdx.x & 31, warps = blockDim.x >> 5;
    const float r2 = 1.7f * 1.7f;
    for (int b = blockIdx.y * warps + (threadIdx.x >> 5); b < T.n_hp_base; b += gridDim.y * warps) {
        const int p = T.hp_base[b];
        const float* q = X + 3 * (size_t)T.pair_atom[p];
        const float x = q[0], y = q[1], z = q[2];
        const int c0 = (cell_of(z, h.org[2], h.inv, h.nc[2]) + 1) * pxy + (cell_of(y, h.org[1], h.inv, h.nc[1]) + 1) * px +
                       cell_of(x, h.org[0], h.inv, h.nc[0]) + 1;
        unsigned other = 0u;          // an atom that is neither carbon nor hydrogen within 1.7 A
        const int ngrid = T.n_pair_a < T.n_pair ? 2 : 1;          // unselected solvent lives in the second set of cells
        for (int g = 0; g < ngrid; ++g) {
            int jb[9], je[9];
#pragma unroll
            for (int r = 0; r < 9; ++r) {               // all eighteen bounds first: independent loads
                const int base = c0 + g * h.ncell + (r % 3 - 1) * px + (r / 3 - 1) * pxy;
                jb[r] = c16 ? (int)cells16[base - 1] : cells32[base - 1];
                je[r] = c16 ? (int)cells16[base + 2] : cells32[base + 2];
            }
#pragma unroll
            for (int r = 0; r < 9; ++r)
                for (int j = jb[r] + lane; j < je[r]; j += 32) {
                    const float4 o = spos[j];
                    if (dist2_exact(o.x, o.y, o.z, x, y, z) <= r2 && !((P.elem_ch_mask >> (__float_as_uint(o.w) & I_ELEM)) & 1u))
                        other = 1u;
                }
        }
        other = __reduce_or_sync(FULL, other);
        const unsigned sub = other ? 0u : 1u;
        unsigned eq = 0u;
        if (!T.same && sub) {
            // sele2 wants the element set to EQUAL {C, H} (hydrophobics.py:39): a hydrogen must be in range too.
            // Hydrogens are not in the grid; ChACRA always passes equal selections, so this scan is the rare path.
            unsigned hyd = 0u;
            for (int a = lane; a < T.n_atoms; a += 32)
                if (T.atom_class[a] == 2) {
                    const float* hq = X + 3 * (size_t)a;
                    if (dist2_exact(hq[0], hq[1], hq[2], x, y, z) <= r2) hyd = 1u;
                }
            eq = __reduce_or_sync(FULL, hyd) ? 2u : 0u;
        }
        if (lane == 0) hm[p] = (unsigned char)(sub | eq);
    }
}
