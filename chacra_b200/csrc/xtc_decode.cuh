// xtc_decode.cuh -- GROMACS XTC frames -> float32 [F][N][3] Angstrom on the device.
//
// Replaces, for `state_{i}.xtc` (chacra/scripts/process_hremd_output.py:147), VMD's gromacs molfile plugin that the
// reference reaches through contact_utils.py:89-116.  The host only walks the frame headers (chacra_b200/xtc.py) and
// ships the COMPRESSED bit streams (about a third of the float32 bytes) over PCIe; this kernel undoes xdrfile's
// xdr3dfcoord coding: per atom three "large" integers (mixed radix, offset by minint) or small differences to the
// previous atom in runs (flag bit + 5 bits; the first two atoms of a run swapped), the small radix index stepping
// through `magicints`.  The bit stream of a frame is inherently sequential, frames are independent: one thread per
// frame.  Parity: unpinned against real files (oracle/xtc_oracle.py header); bit-exact against the oracle codec.
#pragma once
#include "device_types.cuh"

struct XtcFrameMeta {           // mirror of chc_xtc_frame (include/chacra_b200.h)
    long long offset;           // bytes from the start of the blob, multiple of 4
    int nbytes;                 // compressed bytes (raw frames: 12 n_atoms)
    float precision;
    int minint[3], maxint[3];
    int smallidx;
    int raw;                    // 1: n_atoms <= 9, big-endian float32 nm instead of a bit stream
    int bitsize;                // bits of one packed large triple (0: three separate fields of bitsizeint bits), host-computed
    int bitsizeint[3];
};

__device__ const int xtc_magicints[73] = {
    0, 0, 0, 0, 0, 0, 0, 0, 0, 8, 10, 12, 16, 20, 25, 32, 40, 50, 64, 80, 101, 128, 161, 203, 256, 322, 406, 512, 645, 812, 1024, 1290,
    1625, 2048, 2580, 3250, 4096, 5060, 6501, 8192, 10321, 13003, 16384, 20642, 26007, 32768, 41285, 52015, 65536, 82570, 104031,
    131072, 165140, 208063, 262144, 330280, 416127, 524287, 660561, 832255, 1048576, 1321122, 1664510, 2097152, 2642245, 3329021,
    4194304, 5284491, 6658042, 8388607, 10568983, 13316085, 16777216};
#define XTC_FIRSTIDX 9
#define XTC_LASTIDX 72

struct XtcBits {                // MSB-first bit stream over big-endian 32-bit words
    const unsigned* w; unsigned long long acc; int avail;
    __device__ __forceinline__ unsigned get(int k) {           // 0 <= k <= 32
        if (k == 0) return 0u;
        if (avail < k) { acc = (acc << 32) | __byte_perm(*w++, 0u, 0x0123); avail += 32; }
        avail -= k;
        return (unsigned)(acc >> avail) & (k == 32 ? 0xffffffffu : ((1u << k) - 1u));
    }
};

// v / d for v < 2^52 (exact in binary64) and d < 2^24 with a precomputed reciprocal: the estimate is off by at most one
__device__ __forceinline__ unsigned long long xtc_div(unsigned long long v, unsigned d, double inv) {
    unsigned long long q = __double2ull_rz(__ull2double_rn(v) * inv);
    const long long rem = (long long)(v - q * d);
    if (rem < 0) --q; else if (rem >= (long long)d) ++q;
    return q;
}

// the byte order of xdrfile's receiveints: whole bytes least significant first, the remaining bits on top.  W = the
// nbits bits as one MSB-first number (nbits <= 64)
__device__ __forceinline__ unsigned long long xtc_reorder(unsigned long long W, int nbits) {
    const int r = nbits & 7, k = nbits >> 3;
    if (k == 0) return W;
    const unsigned long long rest = r ? (W & ((1ull << r) - 1ull)) : 0ull;
    const unsigned long long top = W >> r;                                   // k bytes, first byte most significant
    const unsigned lo = (unsigned)top, hi = (unsigned)(top >> 32);
    const unsigned long long swapped = ((unsigned long long)__byte_perm(lo, 0u, 0x0123) << 32) | __byte_perm(hi, 0u, 0x0123);
    const unsigned long long rev = swapped >> (64 - 8 * k);
    return k == 8 ? rev : (rev | (rest << (8 * k)));
}

// three mixed-radix digits (radices sz[1], sz[2] < 2^24 matter) from `nbits` bits (xdrfile receiveints); inv1, inv2 =
// 1.0 / sz[1], 1.0 / sz[2]
__device__ __forceinline__ void xtc_ints(XtcBits& B, int nbits, unsigned sz1, unsigned sz2, double inv1, double inv2, int out[3]) {
    const unsigned sz[3] = {0u, sz1, sz2};
    if (nbits <= 64) {
        unsigned long long W;
        if (nbits <= 32) W = B.get(nbits);
        else { const unsigned hi = B.get(nbits - 32); W = ((unsigned long long)hi << 32) | B.get(32); }
        const unsigned long long v = xtc_reorder(W, nbits);
        unsigned long long q, q1;
        if (nbits <= 52) { q = xtc_div(v, sz[2], inv2); q1 = xtc_div(q, sz[1], inv1); }
        else { q = v / sz[2]; q1 = q / sz[1]; }
        out[2] = (int)(v - q * sz[2]);
        out[1] = (int)(q - q1 * sz[1]);
        out[0] = (int)q1;
    } else {                    // up to 72 bits: three 32-bit limbs, schoolbook division by a radix < 2^24
        unsigned limb[3] = {0u, 0u, 0u};
        int sh = 0, n = nbits;
        while (n >= 8) { limb[sh >> 5] |= B.get(8) << (sh & 31); sh += 8; n -= 8; }
        if (n) limb[sh >> 5] |= B.get(n) << (sh & 31);
#pragma unroll
        for (int d = 2; d >= 1; --d) {
            unsigned long long r = 0ull;
#pragma unroll
            for (int k = 2; k >= 0; --k) {
                const unsigned long long cur = (r << 32) | limb[k];
                const unsigned long long q = cur / sz[d];
                limb[k] = (unsigned)q;
                r = cur - q * sz[d];
            }
            out[d] = (int)r;
        }
        out[0] = (int)limb[0];
    }
}

__global__ void __launch_bounds__(128)
xtc_decode_kernel(const unsigned* __restrict__ blob, const XtcFrameMeta* __restrict__ meta, int n_frames, int n_atoms,
                  float* __restrict__ xyz, long long frame_stride, unsigned* __restrict__ status) {
    __shared__ double inv_magic[XTC_LASTIDX + 1];
    for (int k = threadIdx.x; k <= XTC_LASTIDX; k += blockDim.x) inv_magic[k] = xtc_magicints[k] > 0 ? 1.0 / (double)xtc_magicints[k] : 0.0;
    __syncthreads();
    for (int f = blockIdx.x * blockDim.x + threadIdx.x; f < n_frames; f += gridDim.x * blockDim.x) {
        const XtcFrameMeta m = meta[f];
        float* out = xyz + (size_t)f * frame_stride;
        const unsigned* src = blob + (m.offset >> 2);
        if (m.raw) {
            for (int k = 0; k < 3 * n_atoms; ++k) out[k] = __uint_as_float(__byte_perm(src[k], 0u, 0x0123)) * 10.0f;
            continue;
        }
        const float inv = 1.0f / m.precision;
        unsigned sizeint[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) sizeint[k] = (unsigned)(m.maxint[k] - m.minint[k] + 1);
        int smallidx = m.smallidx;
        if (smallidx < XTC_FIRSTIDX || smallidx > XTC_LASTIDX) { atomicOr(status, 1u); continue; }
        int smaller = xtc_magicints[max(XTC_FIRSTIDX, smallidx - 1)] / 2;
        int smallnum = xtc_magicints[smallidx] / 2;
        const double inv1 = 1.0 / (double)sizeint[1], inv2 = 1.0 / (double)sizeint[2];
        double invs = inv_magic[smallidx];
        XtcBits B{src, 0ull, 0};
        const unsigned* end = src + ((m.nbytes + 3) >> 2) + 2;
        // One atom per loop iteration, whichever kind it is (a "large" atom or one small difference of a run), so that the
        // 32 frames of a warp, each somewhere else in its own stream, stay on the same instructions.
        int i = 0, run = 0;            // atoms written; length of the last announced run (in integers, a multiple of 3)
        int remaining = 0, pend = 0;   // small atoms still to come in this run; radix step to apply when it ends
        bool first_small = false;      // the next small atom is the first of its run (stored swapped with the large one)
        int prev[3] = {0, 0, 0};
        unsigned ssmall = (unsigned)xtc_magicints[smallidx];
        bool bad = false;
        while (i < n_atoms) {
            const bool small = remaining > 0;
            int v[3];
            if (!small && m.bitsize == 0) {
#pragma unroll
                for (int k = 0; k < 3; ++k) v[k] = (int)B.get(m.bitsizeint[k]);
            } else {
                xtc_ints(B, small ? smallidx : m.bitsize, small ? ssmall : sizeint[1], small ? ssmall : sizeint[2],
                         small ? invs : inv1, small ? invs : inv2, v);
            }
            bool step = false;         // a run (or a lone large atom) has just ended: the small radix may move
            if (small) {
#pragma unroll
                for (int c = 0; c < 3; ++c) v[c] += prev[c] - smallnum;
                if (first_small) {     // the first two atoms of a run are stored in swapped order (water: H O H)
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        out[3 * i + c] = ((float)v[c] * inv) * 10.0f;
                        out[3 * (i + 1) + c] = ((float)prev[c] * inv) * 10.0f;
                    }
                    i += 2;
                    first_small = false;
                } else {
#pragma unroll
                    for (int c = 0; c < 3; ++c) out[3 * i + c] = ((float)v[c] * inv) * 10.0f;
                    ++i;
                }
#pragma unroll
                for (int c = 0; c < 3; ++c) prev[c] = v[c];
                step = --remaining == 0;
            } else {
#pragma unroll
                for (int k = 0; k < 3; ++k) v[k] += m.minint[k];
                int is_smaller = 0;
                if (B.get(1)) {
                    run = (int)B.get(5);
                    is_smaller = run % 3;
                    run -= is_smaller;
                    --is_smaller;
                }
                if (i + 1 + run / 3 > n_atoms || B.w > end) { bad = true; break; }
                pend = is_smaller;
                if (run > 0) {
#pragma unroll
                    for (int c = 0; c < 3; ++c) prev[c] = v[c];
                    remaining = run / 3;
                    first_small = true;
                } else {
#pragma unroll
                    for (int c = 0; c < 3; ++c) out[3 * i + c] = ((float)v[c] * inv) * 10.0f;
                    ++i;
                    step = true;
                }
            }
            if (step && pend != 0) {
                smallidx += pend;
                if (smallidx < XTC_FIRSTIDX || smallidx > XTC_LASTIDX) { bad = true; break; }
                if (pend < 0) {
                    smallnum = smaller;
                    smaller = smallidx > XTC_FIRSTIDX ? xtc_magicints[smallidx - 1] / 2 : 0;
                } else {
                    smaller = smallnum;
                    smallnum = xtc_magicints[smallidx] / 2;
                }
                ssmall = (unsigned)xtc_magicints[smallidx];
                invs = inv_magic[smallidx];
                pend = 0;
            }
        }
        if (bad) atomicOr(status, 1u);
    }
}

// ---- two-pass decoder ------------------------------------------------------------------------------------------------
// The stream of a frame is a sequence of GROUPS: one "large" atom (fixed number of bits), a flag bit (+ 5 bits that change
// the run length and step the small radix), then run / 3 small atoms of `smallidx` bits each.  Only the group boundaries
// are serial, and finding them needs no arithmetic on the coordinates:
//   pass 1 (xtc_index_kernel, one thread per frame): walk the boundaries -- skip the large bits, read the flag and the 5
//          bits, skip the small bits -- and write one 8-byte record per group (bit position, first atom, small radix
//          index, number of small atoms); raw (<= 9 atom) frames are written out here.  ~25 instructions per group
//          instead of ~100 per atom on the serial chain.
//   pass 2 (xtc_groups_kernel, one thread per group): decode the group's large atom and its run, write the coordinates.
// Group records: bits 0..27 bit position, 28..31 small atoms (0..8), 32..57 first atom, 58..63 smallidx - 9.
__device__ __forceinline__ unsigned xtc_peek(const unsigned* src, long long bit, int k) {        // k <= 25 bits at any bit position
    const unsigned* w = src + (bit >> 5);
    const unsigned long long two = ((unsigned long long)__byte_perm(w[0], 0u, 0x0123) << 32) | __byte_perm(w[1], 0u, 0x0123);
    return (unsigned)(two >> (64 - (int)(bit & 31) - k)) & ((1u << k) - 1u);
}

__global__ void __launch_bounds__(32)
xtc_index_kernel(const unsigned* __restrict__ blob, const XtcFrameMeta* __restrict__ meta, int frame0, int n_frames, int n_atoms,
                 float* __restrict__ xyz, long long frame_stride, unsigned long long* __restrict__ groups, int* __restrict__ gcount,
                 unsigned* __restrict__ status) {
    for (int fl = blockIdx.x * blockDim.x + threadIdx.x; fl < n_frames; fl += gridDim.x * blockDim.x) {
        const int f = frame0 + fl;
        const XtcFrameMeta m = meta[f];
        const unsigned* src = blob + (m.offset >> 2);
        gcount[fl] = 0;
        if (m.raw) {
            float* out = xyz + (size_t)f * frame_stride;
            for (int k = 0; k < 3 * n_atoms; ++k) out[k] = __uint_as_float(__byte_perm(src[k], 0u, 0x0123)) * 10.0f;
            continue;
        }
        int smallidx = m.smallidx;
        const long long limit = 8ll * m.nbytes + 64;                  // the stream may not run past its bytes (+ the reader's slack)
        if (smallidx < XTC_FIRSTIDX || smallidx > XTC_LASTIDX || limit >= (1ll << 28)) { atomicOr(status, 1u); continue; }
        const int large_bits = m.bitsize ? m.bitsize : m.bitsizeint[0] + m.bitsizeint[1] + m.bitsizeint[2];
        unsigned long long* rec = groups + (size_t)fl * n_atoms;
        long long bit = 0;
        int i = 0, g = 0, run = 0;
        bool bad = false;
        while (i < n_atoms) {
            const long long start = bit;
            // the walk is one dependent load per group: ask for the sectors it will need ~40 groups from now
            asm volatile("prefetch.global.L1 [%0];" ::"l"(src + (bit >> 5) + 128));
            bit += large_bits;
            if (bit >= limit) { bad = true; break; }
            int step = 0;
            const unsigned six = xtc_peek(src, bit, 6);                // flag + the 5 bits that follow it when it is set
            if (six & 32u) {
                run = (int)(six & 31u);
                step = run % 3;
                run -= step;
                --step;
                bit += 6;
            } else bit += 1;
            const int ns = run / 3;
            if (i + 1 + ns > n_atoms) { bad = true; break; }
            rec[g++] = (unsigned long long)start | ((unsigned long long)ns << 28) | ((unsigned long long)i << 32) |
                       ((unsigned long long)(smallidx - XTC_FIRSTIDX) << 58);
            bit += (long long)ns * smallidx;
            i += 1 + ns;
            smallidx += step;
            if (smallidx < XTC_FIRSTIDX || smallidx > XTC_LASTIDX || bit > limit) { bad = true; break; }
        }
        if (bad) atomicOr(status, 1u); else gcount[fl] = g;
    }
}

__global__ void __launch_bounds__(128)
xtc_groups_kernel(const unsigned* __restrict__ blob, const XtcFrameMeta* __restrict__ meta, int frame0, int n_atoms,
                  float* __restrict__ xyz, long long frame_stride, const unsigned long long* __restrict__ groups,
                  const int* __restrict__ gcount) {
    const int fl = blockIdx.y, f = frame0 + fl;
    const int ng = gcount[fl];
    if ((int)(blockIdx.x * blockDim.x) >= ng) return;
    const XtcFrameMeta m = meta[f];
    const unsigned* src = blob + (m.offset >> 2);
    float* out = xyz + (size_t)f * frame_stride;
    const float inv = 1.0f / m.precision;
    const unsigned sz1 = (unsigned)(m.maxint[1] - m.minint[1] + 1), sz2 = (unsigned)(m.maxint[2] - m.minint[2] + 1);
    const double inv1 = 1.0 / (double)sz1, inv2 = 1.0 / (double)sz2;
    const unsigned long long* rec = groups + (size_t)fl * n_atoms;
    for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < ng; g += gridDim.x * blockDim.x) {
        const unsigned long long r = rec[g];
        const unsigned start = (unsigned)r & 0x0fffffffu;
        const int ns = (int)((r >> 28) & 15ull);
        int i = (int)((r >> 32) & 0x3ffffffull);
        const int smallidx = (int)(r >> 58) + XTC_FIRSTIDX;
        XtcBits B{src + (start >> 5), 0ull, 0};
        B.get((int)(start & 31u));
        int v[3];
        if (m.bitsize == 0) {
#pragma unroll
            for (int k = 0; k < 3; ++k) v[k] = (int)B.get(m.bitsizeint[k]);
        } else xtc_ints(B, m.bitsize, sz1, sz2, inv1, inv2, v);
#pragma unroll
        for (int k = 0; k < 3; ++k) v[k] += m.minint[k];
        if (ns == 0) {
#pragma unroll
            for (int c = 0; c < 3; ++c) out[3 * i + c] = ((float)v[c] * inv) * 10.0f;
            continue;
        }
        if (B.get(1)) B.get(5);
        const unsigned ssmall = (unsigned)xtc_magicints[smallidx];
        const int smallnum = (int)(ssmall / 2u);
        const double invs = 1.0 / (double)ssmall;
        int prev[3] = {v[0], v[1], v[2]};
        for (int s = 0; s < ns; ++s) {
            xtc_ints(B, smallidx, ssmall, ssmall, invs, invs, v);
#pragma unroll
            for (int c = 0; c < 3; ++c) v[c] += prev[c] - smallnum;
            if (s == 0) {                   // the first two atoms of a run are stored in swapped order (water: H O H)
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    out[3 * i + c] = ((float)v[c] * inv) * 10.0f;
                    out[3 * (i + 1) + c] = ((float)prev[c] * inv) * 10.0f;
                }
                i += 2;
            } else {
#pragma unroll
                for (int c = 0; c < 3; ++c) out[3 * i + c] = ((float)v[c] * inv) * 10.0f;
                ++i;
            }
#pragma unroll
            for (int c = 0; c < 3; ++c) prev[c] = v[c];
        }
    }
}
