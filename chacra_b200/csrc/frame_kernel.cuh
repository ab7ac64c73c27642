// frame_kernel.cuh -- helpers shared by the binning kernel (bin_kernel.cuh) and the pair kernel (pair_kernel.cuh):
// ordered-float keys, record emission, per-frame residue-pair de-duplication, bond / disulfide lookups,
// the float32 hydrogen-bond angle and the float64 ring geometry.
// Reference semantics: SURVEY.md Appendix A / D; file:line citations are in include/chacra_b200.h.
#pragma once
#include "device_types.cuh"

#define FULL 0xffffffffu

__device__ __forceinline__ int f2ord(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

// ---- mbarrier + bulk async copy (TMA engine, no tensor map) ------------------------------------------------
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

// grid coordinate of x along one axis; the binning kernel and the pair kernel must agree bit for bit (-fmad=false)
// frame f of a submission: consecutive, or picked through a source index (replica -> state demultiplexing)
__device__ __forceinline__ const float* frame_ptr(const float* xyz, const long long* __restrict__ src, long long f, long long frame_stride) {
    return xyz + (size_t)(src ? __ldg(src + f) : f) * (size_t)frame_stride;
}

// Grid over a frame's bounding box: the cell edge starts at the cutoff and only ever grows (every pair within the cutoff
// stays in adjacent cells) until the padded grid fits `max_cells` counters.  Safe for any input: a NaN or infinite extent
// counts as zero (such atoms are clamped into the border cells by cell_of), per-axis counts are capped before they are
// multiplied, so nothing overflows and the loop ends (at most ~400 steps for an extent of 3e38).
__device__ __forceinline__ void grid_dims(const float ext_in[3], float cell, long long max_cells, bool gridb, float& inv, int nc[3]) {
    float ext[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) ext[d] = (ext_in[d] >= 0.f && ext_in[d] <= 3.0e38f) ? ext_in[d] : 0.f;
    for (;;) {
        inv = 1.0f / cell;
#pragma unroll
        for (int d = 0; d < 3; ++d) nc[d] = (int)fminf(ext[d] * inv, 1.0e6f) + 1;
        if ((long long)(nc[0] + 2) * (nc[1] + 2) * (nc[2] + 2) * (gridb ? 2 : 1) + 1 <= max_cells) break;
        cell *= 1.25f;
    }
}

__device__ __forceinline__ int cell_of(float x, float origin, float inv, int n) {
    int c = (int)((x - origin) * inv);
    return min(max(c, 0), n - 1);
}

// ---- output helpers ---------------------------------------------------------------------------
__device__ __forceinline__ void emit_contact(const DevOut& O, int frame, int itype, int a1, int a2, float dist) {
    unsigned m = __activemask();
    int lane = threadIdx.x & 31;
    int leader = __ffs(m) - 1;
    unsigned long long base = 0;
    if (lane == leader) base = atomicAdd(O.contact_n, (unsigned long long)__popc(m));
    base = __shfl_sync(m, base, leader);
    unsigned long long pos = base + __popc(m & ((1u << lane) - 1));
    if (pos < (unsigned long long)O.contact_cap) {
        int4* dst = reinterpret_cast<int4*>(O.contacts + pos);
        dst[0] = make_int4(frame, itype, a1, a2);
        dst[1] = make_int4(-1, -1, __float_as_int(dist), 0);
    }
}

// water-bridge record: atom1, atom2 and the one or two mediating solvent atoms (hbonds.py:168-204)
__device__ __forceinline__ void emit_bridge(const DevOut& O, int frame, int itype, int a1, int a2, int w1, int w2, float dist) {
    unsigned long long pos = atomicAdd(O.contact_n, 1ull);
    if (pos < (unsigned long long)O.contact_cap) {
        int4* dst = reinterpret_cast<int4*>(O.contacts + pos);
        dst[0] = make_int4(frame, itype, a1, a2);
        dst[1] = make_int4(w1, w2, __float_as_int(dist), 0);
    }
}

// Per-CTA scratch of the water-bridge phase (global memory): for every solvent pair atom (by solvent ordinal) the
// number of hydrogen-bonded non-solvent neighbours found in this frame, their sorted slots, the solvent atom's own
// sorted slot, and the list of solvent ordinals that have at least one neighbour.  `cnt` is all zero between frames.
#define WADJ 8
struct WaterScratch { int* cnt; int* adj; int* slot; int* list; };
__device__ __forceinline__ WaterScratch water_scratch(int* base, int cta, int n_solv) {
    WaterScratch W{nullptr, nullptr, nullptr, nullptr};
    if (base && n_solv > 0) {
        int* p = base + (size_t)cta * n_solv * (3 + WADJ);
        W.cnt = p; W.adj = p + n_solv; W.slot = p + (size_t)n_solv * (1 + WADJ); W.list = p + (size_t)n_solv * (2 + WADJ);
    }
    return W;
}

__device__ __forceinline__ void emit_near(const DevOut& O, int frame, int kind, int a1, int a2, double delta) {
    unsigned long long pos = atomicAdd(O.near_n, 1ull);
    if (pos < (unsigned long long)O.near_cap) {
        chc_near r; r.frame = frame; r.kind = kind; r.atom1 = a1; r.atom2 = a2; r.delta = delta; r.reserved = 0;
        O.near[pos] = r;
    }
}

// Per-frame residue-pair set.  Two encodings of the same idea (first setter of a pair bumps its count):
//  * plain bitset (shared memory, cleared per frame):            word = idx / 32, bit = idx % 32
//  * generation-stamped bitset (global memory, one per CTA, never cleared between frames): every word holds
//    an 8-bit generation in its top byte and 24 pair bits; a word whose generation is stale counts as empty.
//    The generation runs 1..255 and the table is wiped when it wraps (pair_kernel.cuh).
__device__ __forceinline__ void count_pair(const DevOut& O, unsigned* bits, unsigned gen, int ra, int rb, int R) {
    if (gen == 0u) {
        // plain shared-memory bitset: few residues (the set fits in shared memory), 32-bit index arithmetic; most contacts of a frame
        // repeat a residue pair that is already set: look before the atomic
        const unsigned lo = (unsigned)min(ra, rb), hi = (unsigned)max(ra, rb);
        const unsigned idx = lo * (unsigned)R - (lo * (lo - 1u)) / 2u + (hi - lo);
        const unsigned bit = 1u << (idx & 31u);
        if (*reinterpret_cast<volatile unsigned*>(&bits[idx >> 5]) & bit) return;
        const unsigned old = atomicOr(&bits[idx >> 5], bit);
        if (!(old & bit)) atomicAdd(&O.counts[idx], 1);
    } else {
        long long idx = pair_index(ra, rb, R);
        unsigned* w = bits + idx / 24;
        unsigned bit = 1u << (unsigned)(idx % 24);
        unsigned old = *w;
        for (;;) {
            unsigned cur = (old >> 24) == gen ? old : (gen << 24);
            if (cur & bit) return;
            unsigned seen = atomicCAS(w, old, cur | bit);
            if (seen == old) break;
            old = seen;
        }
        atomicAdd(&O.counts[idx], 1);
    }
}

__device__ __forceinline__ bool is_bonded(const DevTopo& T, int pi, int atom_j) {
    int b = T.pair_bond_start[pi], e = T.pair_bond_start[pi + 1];
    for (int k = b; k < e; ++k) if (T.pair_bond_adj[k] == atom_j) return true;
    return false;
}

__device__ __forceinline__ bool is_disulfide(const DevTopo& T, int resid_i, int resid_j) {
    for (int k = 0; k < T.n_dis; ++k) {
        int a = T.disulfide[2 * k], b = T.disulfide[2 * k + 1];
        if ((a == resid_i && b == resid_j) || (a == resid_j && b == resid_i)) return true;
    }
    return false;
}

// measure hbonds angle: 57.2958f * (float)atan2(|u x v|, u.v), u = h-d, v = a-h, float32 products
__device__ __forceinline__ float hbond_angle(float dx, float dy, float dz, float hx, float hy, float hz,
                                             float ax, float ay, float az) {
    float ux = hx - dx, uy = hy - dy, uz = hz - dz;
    float vx = ax - hx, vy = ay - hy, vz = az - hz;
    float cx = uy * vz - uz * vy, cy = uz * vx - ux * vz, cz = ux * vy - uy * vx;
    float psin = sqrtf((cx * cx + cy * cy) + cz * cz);
    float pcos = (ux * vx + uy * vy) + uz * vz;
    return 57.2958f * (float)atan2((double)psin, (double)pcos);
}

// ---- float64 ring geometry (contact_utils.py:958-1061) ---------------------------------------------
struct D3 { double x, y, z; };
__device__ __forceinline__ D3 d3(const float* p) { return D3{(double)p[0], (double)p[1], (double)p[2]}; }
__device__ __forceinline__ D3 sub3(D3 a, D3 b) { return D3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ double dot3(D3 a, D3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ D3 cross3(D3 a, D3 b) { return D3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ D3 centroid3(D3 a, D3 b, D3 c) { return D3{((a.x + b.x) + c.x) / 3.0, ((a.y + b.y) + c.y) / 3.0, ((a.z + b.z) + c.z) / 3.0}; }
__device__ __forceinline__ D3 normal3(D3 p1, D3 p2, D3 p3) { return cross3(sub3(p3, p1), sub3(p2, p1)); }
__device__ __forceinline__ double angle_deg(D3 a, D3 b) {
    double c = dot3(a, b) / (sqrt(dot3(a, a)) * sqrt(dot3(b, b)));
    return acos(c) * (180.0 / 3.14159265358979323846);
}
__device__ __forceinline__ double fold180(double t) { return fmin(fabs(t), fabs(t - 180.0)); }
