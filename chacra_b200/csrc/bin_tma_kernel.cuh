// bin_tma_kernel.cuh -- the coordinate-streaming kernel for frames that fit in shared memory.
//
// Same job and same output record as bin_frames_kernel (bin_kernel.cuh), organised around the HBM pipe instead of
// around the sort.  Whole frames arrive through 1-D bulk async copies (cp.async.bulk -> SASS UBLKCP, completion counted
// on an mbarrier): the coordinate stream is perfectly coalesced and costs no registers and no issue slots.  Two
// shapes of the same code (template NBUF):
//   NBUF = 1  two 256-thread CTAs per SM, one frame buffer each.  A CTA copies its frame to registers (P1), releases
//             the buffer and has the NEXT frame's bulk copy in flight while it sorts the current one; the two CTAs of
//             an SM are independent instruction streams, so the barrier-separated phases of one hide behind the other.
//   NBUF = 2  one 512-thread CTA per SM, two frame buffers, the sorted record leaves through a bulk async store.
// Per frame (<= APT pair atoms per thread):
//   P1  gather this thread's pair atoms from the shared-memory frame into registers, bounding box (warp redux +
//       shared atomics), aux atoms straight to the record (one float per thread, coalesced)          -> barrier
//   P2  cell of every atom; shared atomicAdd returns the atom's rank inside its cell                -> barrier
//   P3  exclusive scan of the counters in use (int4 groups per thread)                              -> barrier
//   P4  scatter x, y, z, flags|index to start[cell] + rank; cell starts to the record (coalesced)
//   P5  (frames that open a 100-frame fragment, equal selections) hydrophobic carbon membership: 1.7 A query on the
//       grid just built (hydrophobics.py:8-41), thread = base carbon, sorted atoms staged in the frame buffer
// Everything that is the same for all threads of a frame (grid origin, cell edge, grid dimensions) costs issue slots in
// every warp, so the CTAs are few fat warps (up to 16 atoms per thread) rather than many thin ones.
// Algorithmic HBM traffic: 12 B x n_atoms read per frame; the record (written once) is consumed from L2.
#pragma once
#include "frame_kernel.cuh"

#define BT_MAXG 8               // uint4 groups (8 16-bit counters each) per thread in the scan

struct BinTmaLayout {
    int off_auxtab;             // uint16 [3 n_aux]: float offset inside the frame of every aux coordinate
    int off_buf0, off_buf1;     // frame buffers, bytes from the start of dynamic shared memory (cell counters come first)
    int buf_bytes;              // one frame buffer: >= 12 n_atoms + 28 and >= 32 + 16 n_pair + 4 n_hp_base (P5 / record image)
    int cap_groups;             // capacity of the counter array in uint4 groups (8 counters) per thread
    int max_cells;              // cells the grid may use (<= BinLayout::max_cells and < 8 cap_groups THREADS)
    int open_phase;             // first_index % fragment_frames of the launch (set per launch)
    int total;
};

struct BinTmaShared {
    int bb_min[2][3], bb_max[2][3];
    int carry[32];
    int nlist;                  // P5: base carbons queued
    unsigned long long full[2];
};


template <int THREADS, int APT, int NBUF>
__global__ void __launch_bounds__(THREADS, NBUF == 1 ? 2 : 1)
bin_frames_tma_kernel(DevTopo T, DevParams P, BinLayout B, BinTmaLayout Y, const float* __restrict__ xyz, long long frame_stride,
                      int n_frames, int frag_frames, long long frag_slot0, unsigned char* __restrict__ binned,
                      unsigned char* __restrict__ hpmask, int fuse_hp, const long long* __restrict__ src_index) {
    constexpr bool STAGE = NBUF == 2;     // record image in shared memory + bulk store, or direct global stores
    extern __shared__ __align__(128) unsigned char bsm[];      // counters u16[8 cap_groups THREADS] | aux table | frame buffer(s)
    __shared__ BinTmaShared M;
    // cell counters / starts are 16 bit (counts and starts are at most n_pair <= 8 192): half the shared memory and
    // half the scan traffic of 32-bit ones; atomics work on the containing word
    unsigned short* cells = reinterpret_cast<unsigned short*>(bsm);
    unsigned* cellw = reinterpret_cast<unsigned*>(bsm);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int NP = T.n_pair;
    const bool gridb = T.n_pair_a < NP;
    const unsigned frame_bytes = 12u * (unsigned)T.n_atoms;
    const int dummy = 8 * Y.cap_groups * THREADS - 1;          // counter nobody reads: padding slots count there
    // bulk copy of frame f into buffer b: source rounded down / size rounded up to 16 bytes (the consumer skips the
    // leading floats); the rounding never leaves the 16-byte granule the caller's allocation ends in
    auto issue = [&](int f, int b) {
        const unsigned long long a = reinterpret_cast<unsigned long long>(frame_ptr(xyz, src_index, f, frame_stride));
        const unsigned long long a0 = a & ~15ull;
        const unsigned bytes = ((unsigned)(a - a0) + frame_bytes + 15u) & ~15u;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(&M.full[b], bytes);
        bulk_g2s(bsm + (b ? Y.off_buf1 : Y.off_buf0), reinterpret_cast<const void*>(a0), bytes, &M.full[b]);
    };
    if (tid == 0) {                       // the first frame(s) are on their way before anything else happens
        mbar_init(&M.full[0], 1); mbar_init(&M.full[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#pragma unroll
        for (int k = 0; k < NBUF; ++k)
            if (blockIdx.x + k * gridDim.x < n_frames) issue(blockIdx.x + k * gridDim.x, k);
        M.nlist = 0;
    }
    if (tid < 3) {
        M.bb_min[0][tid] = M.bb_min[1][tid] = 0x7fffffff;
        M.bb_max[0][tid] = M.bb_max[1][tid] = (int)0x80000000;
    }
    // this thread's pair atoms p = tid + k THREADS (float offset inside the frame); padding slots repeat the thread's
    // first atom (or atom 0), so that P1 and P2 run without branches: they change neither the box nor any real counter
    int r_atom[APT];
#pragma unroll
    for (int k = 0; k < APT; ++k) {
        const int p = tid + k * THREADS;
        r_atom[k] = 3 * T.pair_atom[p < NP ? p : min(tid, NP - 1)];
    }
    {                                     // float offsets of the aux coordinates: a shared-memory copy (read every frame)
        const uint4* src = reinterpret_cast<const uint4*>(T.aux_off16);
        uint4* dst = reinterpret_cast<uint4*>(bsm + Y.off_auxtab);
        for (int q = tid; q < (3 * T.n_aux + 7) / 8; q += THREADS) dst[q] = __ldg(&src[q]);
    }
    const unsigned short* auxtab = reinterpret_cast<const unsigned short*>(bsm + Y.off_auxtab);
    const int n_aux3 = 3 * T.n_aux;
    for (int g = 0; g < Y.cap_groups; ++g) reinterpret_cast<uint4*>(bsm)[g * THREADS + tid] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    int it = 0, used_groups = 0;          // int4 groups per thread the previous frame dirtied
    int fphase = (Y.open_phase + (int)blockIdx.x) % frag_frames;      // (open_phase + f) % frag_frames, kept incrementally
    int refill_f = -1, refill_b = 0;      // buffer whose record image is being stored: refill it once the store has read it
    for (int f = blockIdx.x; f < n_frames; f += gridDim.x, ++it) {
        const int b = NBUF == 2 ? (it & 1) : 0, bi = it & 1;
        unsigned char* rec = binned + (size_t)f * B.stride;
        unsigned char* fbuf = bsm + (b ? Y.off_buf1 : Y.off_buf0);
        const float* Xs = reinterpret_cast<const float*>(fbuf) +
                          (int)((reinterpret_cast<unsigned long long>(frame_ptr(xyz, src_index, f, frame_stride)) & 15ull) >> 2);
        const int fidx = Y.open_phase + f;                      // position inside the fragment sequence of this launch
        if (it > 0) { fphase += (int)gridDim.x; while (fphase >= frag_frames) fphase -= frag_frames; }
        const bool opens = fuse_hp && fphase == 0;
        // counters the previous frame used go back to zero (nobody reads them any more: S6)
        for (int g = 0; g < used_groups; ++g) reinterpret_cast<uint4*>(bsm)[tid * used_groups + g] = make_uint4(0u, 0u, 0u, 0u);
        if (tid == 0 && refill_f >= 0) {
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            if (refill_f < n_frames) issue(refill_f, refill_b);
        }
        refill_f = -1;
        mbar_wait(&M.full[b], (unsigned)(NBUF == 2 ? (it >> 1) : it) & 1u);
        // ---- P1: registers <- shared frame, bounding box, aux atoms -----------------------------------------------------
        float rx[APT], ry[APT], rz[APT];
        {
            float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
#pragma unroll
            for (int k = 0; k < APT; ++k) {
                const float* q = Xs + r_atom[k];
                rx[k] = q[0]; ry[k] = q[1]; rz[k] = q[2];
            }
#pragma unroll
            for (int k = 0; k < APT; ++k) {
                lo[0] = fminf(lo[0], rx[k]); hi[0] = fmaxf(hi[0], rx[k]);
                lo[1] = fminf(lo[1], ry[k]); hi[1] = fmaxf(hi[1], ry[k]);
                lo[2] = fminf(lo[2], rz[k]); hi[2] = fmaxf(hi[2], rz[k]);
            }
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                const int a = __reduce_min_sync(FULL, f2ord(lo[d])), c = __reduce_max_sync(FULL, f2ord(hi[d]));
                if (lane == 0) { atomicMin(&M.bb_min[bi][d], a); atomicMax(&M.bb_max[bi][d], c); }
            }
            // aux atoms (donor hydrogens, ring atoms, cations) straight to the record, one float per thread: coalesced
            float* aux = reinterpret_cast<float*>(rec + B.off_aux);
#pragma unroll 4
            for (int e = tid; e < n_aux3; e += THREADS) aux[e] = Xs[auxtab[e]];
        }
        __syncthreads();                                                                    // S1: the frame lives in registers
        if (!STAGE && tid == 0 && !opens && f + NBUF * (int)gridDim.x < n_frames) issue(f + NBUF * gridDim.x, b);
        if (tid >= 32 && tid < 35) {      // bounding box of the next frame starts empty (read again only after later barriers)
            M.bb_min[bi ^ 1][tid - 32] = 0x7fffffff; M.bb_max[bi ^ 1][tid - 32] = (int)0x80000000;
        }
        float org[3], inv; int nc[3];
        {
            float ext[3];
#pragma unroll
            for (int d = 0; d < 3; ++d) { org[d] = ord2f(M.bb_min[bi][d]); ext[d] = ord2f(M.bb_max[bi][d]) - org[d]; }
            grid_dims(ext, P.cell, Y.max_cells, gridb, inv, nc);
        }
        const int px = nc[0] + 2, pxy = px * (nc[1] + 2), ncell = pxy * (nc[2] + 2);
        const int nkey = gridb ? 2 * ncell : ncell;
        const int groups = (nkey + 8 * THREADS) / (8 * THREADS);        // covers counters 0..nkey
        used_groups = groups;
        // ---- P2: cell of every atom; the shared atomic returns the atom's rank inside its cell -----------------------
        unsigned r_cr[APT];               // cell | rank << 16   (max_cells <= 16 384 < 65 536, n_pair <= 8 192 on this path)
#pragma unroll
        for (int k = 0; k < APT; ++k) {
            const int p = tid + k * THREADS;
            int c = (cell_of(rz[k], org[2], inv, nc[2]) + 1) * pxy + (cell_of(ry[k], org[1], inv, nc[1]) + 1) * px + cell_of(rx[k], org[0], inv, nc[0]) + 1;
            r_cr[k] = (unsigned)(p < NP ? c : dummy);
        }
        if (gridb) {                      // unselected solvent sorts into the second set of cells
#pragma unroll
            for (int k = 0; k < APT; ++k) {
                const int p = tid + k * THREADS;
                if (p >= T.n_pair_a && p < NP) r_cr[k] += (unsigned)ncell;
            }
        }
#pragma unroll
        for (int k = 0; k < APT; ++k) {
            const unsigned sh = (r_cr[k] & 1u) << 4;
            const unsigned old = atomicAdd(&cellw[r_cr[k] >> 1], 1u << sh);
            r_cr[k] |= ((old >> sh) & 0xffffu) << 16;
        }
        __syncthreads();                                                                    // S2
        // ---- P3: exclusive scan; thread t owns counters [8 groups t, 8 groups (t + 1)) -----------------------------------
        {
            uint4 v[BT_MAXG];
            unsigned s2 = 0u;             // even counters in the low half, odd ones in the high half (no carry: sums <= 8 192)
#pragma unroll
            for (int g = 0; g < BT_MAXG; ++g)
                if (g < groups) {
                    v[g] = reinterpret_cast<const uint4*>(bsm)[tid * groups + g];
                    s2 += (v[g].x + v[g].y) + (v[g].z + v[g].w);
                }
            const int s = (int)((s2 & 0xffffu) + (s2 >> 16));
            int incl = s;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
            if (lane == 31) M.carry[warp] = incl;
            __syncthreads();                                                                // S3
            // every warp scans the (at most 32) warp totals itself: no second barrier
            const int w0 = lane < (THREADS >> 5) ? M.carry[lane] : 0;
            int w = w0;
#pragma unroll
            for (int o = 1; o < (THREADS >> 5); o <<= 1) { const int t = __shfl_up_sync(FULL, w, o); if (lane >= o) w += t; }
            unsigned run = (unsigned)(__shfl_sync(FULL, w - w0, warp) + incl - s);
            auto ex2 = [&](unsigned word) {      // two counters -> their two exclusive starts
                const unsigned lo = word & 0xffffu, e = run | ((run + lo) << 16);
                run += lo + (word >> 16);
                return e;
            };
#pragma unroll
            for (int g = 0; g < BT_MAXG; ++g)
                if (g < groups) {
                    uint4 e;
                    e.x = ex2(v[g].x); e.y = ex2(v[g].y); e.z = ex2(v[g].z); e.w = ex2(v[g].w);
                    reinterpret_cast<uint4*>(bsm)[tid * groups + g] = e;
                }
        }
        __syncthreads();                                                                    // S5
        // ---- P4: scatter (into the record image in the frame buffer, or straight to the record) -----------------------
        unsigned char* img = (STAGE || opens) ? fbuf : rec;
        float4* stage = reinterpret_cast<float4*>(fbuf + B.off_spos);
        int* list = reinterpret_cast<int*>(fbuf + B.off_aux);           // P5 work list: sorted slots of the base carbons
        {
            int dst[APT]; unsigned info[APT];
#pragma unroll
            for (int k = 0; k < APT; ++k) {
                const int p = tid + k * THREADS;
                dst[k] = (int)cells[r_cr[k] & 0xffffu] + (int)(r_cr[k] >> 16);
                info[k] = p < NP ? ((__ldg(&T.pair_info[p]) & I_WMASK) | ((unsigned)p << I_PIDX_SHIFT)) : 0u;
            }
            float4* out = reinterpret_cast<float4*>(img + B.off_spos);
#pragma unroll
            for (int k = 0; k < APT; ++k)
                if (tid + k * THREADS < NP) out[dst[k]] = make_float4(rx[k], ry[k], rz[k], __uint_as_float(info[k]));
            if (opens) {
#pragma unroll
                for (int k = 0; k < APT; ++k)
                    if (tid + k * THREADS < NP && (info[k] & (I_HP1 | I_HP2))) list[atomicAdd(&M.nlist, 1)] = dst[k];
            }
            // cell starts 0..nkey (the last one = total) to the record, coalesced
            if (B.cell16) {                 // the packed scan result as it is
                uint4* cells_out = reinterpret_cast<uint4*>(rec + B.off_cells);
                for (int q = tid; 8 * q <= nkey; q += THREADS) cells_out[q] = reinterpret_cast<const uint4*>(bsm)[q];
            } else {
                int4* cells_out = reinterpret_cast<int4*>(rec + B.off_cells);
                for (int q = tid; 4 * q <= nkey; q += THREADS) {
                    const uint2 w = reinterpret_cast<const uint2*>(bsm)[q];
                    cells_out[q] = make_int4((int)(w.x & 0xffffu), (int)(w.x >> 16), (int)(w.y & 0xffffu), (int)(w.y >> 16));
                }
            }
        }
        if (tid == 0) {
            BinHeader h;
            h.org[0] = org[0]; h.org[1] = org[1]; h.org[2] = org[2]; h.inv = inv;
            h.nc[0] = nc[0]; h.nc[1] = nc[1]; h.nc[2] = nc[2]; h.ncell = ncell;
            *reinterpret_cast<BinHeader*>(img) = h;
        }
        if (STAGE || opens) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // image writes -> async proxy
        __syncthreads();                                                                    // S6
        if (STAGE || opens) {
            if (tid == 0) {
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                             ::"l"(rec), "r"(smem_u32(fbuf)), "r"((unsigned)B.off_aux) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            refill_f = f + NBUF * (int)gridDim.x; refill_b = b;
        }
        // ---- P5: hydrophobic carbon membership of the fragment this frame opens (equal selections) ---------------------
        // thread = base carbon; the 1.7 A ball touches at most 2 x 2 x 2 cells (cell edge >= 3.5 A)
        if (opens) {
            unsigned char* hm = hpmask + (size_t)(frag_slot0 + fidx / frag_frames) * NP;
            const float r2 = 1.7f * 1.7f, reach = 1.75f;
            const int nlist = M.nlist;
            for (int e = tid; e < nlist; e += THREADS) {
                const float4 me = stage[list[e]];
                const int x0 = cell_of(me.x - reach, org[0], inv, nc[0]), x1 = cell_of(me.x + reach, org[0], inv, nc[0]);
                const int y0 = cell_of(me.y - reach, org[1], inv, nc[1]), y1 = cell_of(me.y + reach, org[1], inv, nc[1]);
                const int z0 = cell_of(me.z - reach, org[2], inv, nc[2]), z1 = cell_of(me.z + reach, org[2], inv, nc[2]);
                unsigned other = 0u;          // an atom that is neither carbon nor hydrogen within 1.7 A
                for (int g = 0; g < (gridb ? 2 : 1); ++g)
                    for (int z = z0; z <= z1; ++z)
                        for (int y = y0; y <= y1; ++y) {
                            const int row = g * ncell + (z + 1) * pxy + (y + 1) * px + 1;
                            const int je = (int)cells[row + x1 + 1];
                            for (int j = (int)cells[row + x0]; j < je; ++j) {
                                const float4 o = stage[j];
                                if (dist2_exact(o.x, o.y, o.z, me.x, me.y, me.z) <= r2 &&
                                    !((P.elem_ch_mask >> (__float_as_uint(o.w) & I_ELEM)) & 1u)) other = 1u;
                            }
                        }
                hm[__float_as_uint(me.w) >> I_PIDX_SHIFT] = other ? 0u : 1u;
            }
            __syncthreads();              // the next frame re-zeroes the counters and refills this buffer
            if (tid == 0) M.nlist = 0;
        }
    }
    if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");       // the last record image has left shared memory
}
