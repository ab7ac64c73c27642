// chacra_b200.cu -- C ABI of libchacra_b200.so (see include/chacra_b200.h).  sm_100a only.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdarg>
#include <cmath>
#include <vector>
#include <algorithm>
#include <mutex>
#include "bin_kernel.cuh"
#include "bin_tma_kernel.cuh"
#include "pair_kernel.cuh"
#include "xtc_decode.cuh"
#include "pca_perm.cuh"
#include "dcd_unpack.cuh"
#include "tsv_writer.h"
#include "tsv_counts.h"

#ifndef PAIR_THREADS
#define PAIR_THREADS 1024
#endif
#define SMEM_LIMIT (227 * 1024)

struct chc_ctx {
    int device = 0;
    int sm_count = 148;
    chc_params prm{};
    DevTopo T{};
    DevParams P{};
    PairLayout L{};
    PairLayout LF{};             // shared-memory plan of the fused kernel (the CTA sorts its own frame); valid when fused_ok
    bool fused_ok = false;
    long long* open_idx = nullptr; int64_t open_cap = 0;     // fragment-opening frames of a chunk (hydrophobic masks ahead of the fused kernel)
    BinLayout B{};
    BinTmaLayout Y{};
    bool small_path = true, bin_tma = false; int tma_apt = 8;
    int lkey_max = 0;
    int bin_smem = 0, bin_grid = 0, pair_grid = 0, bin_threads = 256, bin_variant = 0;
    int64_t chunk_frames = 0;
    unsigned char* binned = nullptr;
    unsigned* gstamp = nullptr; unsigned* gstate = nullptr;
    unsigned char* gsoa = nullptr;       // large path: per-CTA slot-pair arrays
    int* wscratch = nullptr; unsigned* status = nullptr;
    std::vector<void*> dev_allocs;
    int32_t* counts = nullptr; bool counts_owned = false; int64_t n_pairs = 0;
    chc_contact* contacts = nullptr; unsigned long long* contact_n = nullptr;
    chc_near* near = nullptr; unsigned long long* near_n = nullptr;
    unsigned char* hpmask = nullptr; int64_t hpmask_frags = 0;
    float* dev_stage[2] = {nullptr, nullptr}; float* pin_stage[2] = {nullptr, nullptr};
    int64_t stage_frames = 0; cudaEvent_t stage_done[2] = {nullptr, nullptr}; cudaEvent_t copied[2] = {nullptr, nullptr}; cudaStream_t copy_stream = nullptr;
    bool timing = false;
    struct Ev3 { cudaEvent_t a, b, c; };
    std::vector<Ev3> ev_pending, ev_free;   // CUDA events around (bin kernel, pair kernel) of a chunk
    int64_t last_state = -1, next_index = -1;          // for fragment continuation checks
    int hp_last_slot = 0;
    chc_stats st{};
    char err[512] = {0};
};

static int fail(chc_ctx* c, int code, const char* fmt, ...) {
    if (c) { va_list ap; va_start(ap, fmt); vsnprintf(c->err, sizeof(c->err), fmt, ap); va_end(ap); }
    return code;
}
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(ctx, CHC_E_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

template <typename Tp>
static int upload(chc_ctx* ctx, const Tp* host, size_t n, const Tp** dev) {
    *dev = nullptr;
    size_t bytes = std::max<size_t>(n, 1) * sizeof(Tp);
    void* d = nullptr;
    CK(cudaMalloc(&d, bytes));
    ctx->dev_allocs.push_back(d);
    if (n) { if (!host) return fail(ctx, CHC_E_INVALID, "null table with %zu entries", n); CK(cudaMemcpy(d, host, n * sizeof(Tp), cudaMemcpyHostToDevice)); }
    else CK(cudaMemset(d, 0, bytes));
    *dev = static_cast<const Tp*>(d);
    return CHC_OK;
}

// apt: atoms per thread of the bulk-copy binning kernels (chosen by plan_bin_tma in chc_create)
static const void* bin_kernel_fn(int variant, int apt, int* threads) {
    switch (variant) {
        case 3:               // two 256-thread CTAs per SM, one frame buffer each
            if (threads) *threads = 256;
            switch (apt) {
                case 2: return (const void*)bin_frames_tma_kernel<256, 2, 1>;
                case 4: return (const void*)bin_frames_tma_kernel<256, 4, 1>;
                case 6: return (const void*)bin_frames_tma_kernel<256, 6, 1>;
                case 8: return (const void*)bin_frames_tma_kernel<256, 8, 1>;
                case 10: return (const void*)bin_frames_tma_kernel<256, 10, 1>;
                case 12: return (const void*)bin_frames_tma_kernel<256, 12, 1>;
                case 13: return (const void*)bin_frames_tma_kernel<256, 13, 1>;
                case 14: return (const void*)bin_frames_tma_kernel<256, 14, 1>;
                default: return (const void*)bin_frames_tma_kernel<256, 16, 1>;
            }
        case 4:               // one 512-thread CTA per SM, two frame buffers, bulk store of the record image
            if (threads) *threads = 512;
            return apt <= 8 ? (const void*)bin_frames_tma_kernel<512, 8, 2> : (const void*)bin_frames_tma_kernel<512, 16, 2>;
        case 1: if (threads) *threads = 512; return (const void*)bin_frames_kernel<512, 8, 2>;
        case 2: if (threads) *threads = 1024; return (const void*)bin_frames_kernel<1024, 8, 1>;
        default: if (threads) *threads = 1024; return (const void*)bin_frames_kernel<1024, 0, 1>;
    }
}

// small: the frame lives in shared memory; rec: records mode (contact records and / or near-cutoff channel); fused: the CTA
// sorts its frame itself from the caller's coordinates (no binned record)
static const void* pair_kernel_fn(bool small, bool rec, bool fused = false) {
    if (fused) return rec ? (const void*)pair_frames_kernel<PAIR_THREADS, true, true, true> : (const void*)pair_frames_kernel<PAIR_THREADS, true, false, true>;
    if (small) return rec ? (const void*)pair_frames_kernel<PAIR_THREADS, true, true, false> : (const void*)pair_frames_kernel<PAIR_THREADS, true, false, false>;
    return rec ? (const void*)pair_frames_kernel<PAIR_THREADS, false, true, false> : (const void*)pair_frames_kernel<PAIR_THREADS, false, false, false>;
}

__global__ void iota_stride_kernel(long long* out, long long first, long long step, int n) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) out[k] = first + (long long)k * step;
}

static size_t al16(size_t v) { return (v + 15) & ~size_t(15); }

// Decide where a frame lives while its pairs are enumerated (shared memory or in place in global memory) and lay
// out the binned record and the pair kernel's shared memory.
static int plan(chc_ctx* ctx, size_t cells_cap = 49152) {
    const DevTopo& T = ctx->T;
    PairLayout& L = ctx->L;
    BinLayout& B = ctx->B;
    const int warps = PAIR_THREADS / 32;
    if (T.n_pair > MAX_PAIR_ATOMS) return CHC_E_TOO_LARGE;
    B.off_spos = (int)sizeof(BinHeader);
    B.off_aux = B.off_spos + 16 * T.n_pair;
    B.fixed_bytes = (int)al16((size_t)B.off_aux + 12 * (size_t)std::max(T.n_aux, 1));
    B.off_cells = B.fixed_bytes;
    auto queue_bytes = [&](bool small, size_t* q1, size_t* qf, size_t* q2) {
        const size_t qw = small ? 1 : 2;
        // E1 -> E2 entries (atom slot, first neighbour slot, survivor mask): 8 bytes each on the shared-memory path, 16 otherwise
        *q1 = 0; *qf = al16((small ? 8 : 16) * (size_t)warps * QE_CAP);
        *q2 = al16(8 * (qw + 1) * (size_t)warps * Q2_CAP);
        return *q1 + *qf + *q2;
    };
    size_t q1b, qfb, q2b;
    const size_t queues = queue_bytes(true, &q1b, &qfb, &q2b);
    // ring phase: float centroids + the list of (cation, ring) / (ring, ring) tasks that pass the centroid prefilter
    const size_t ring_bytes = al16(12 * (size_t)std::max(T.n_rings, 1));
    int rlist_cap = T.n_rings <= 128 ? 1024 : 8192;
    if (const char* e = getenv("CHACRA_RLIST_CAP")) rlist_cap = std::max(4, std::min(rlist_cap, atoi(e)));   // test hook: overflow -> every ring task the slow way
    // ring task numbers (cation * n_rings + ring, triangular ring pairs) are 32-bit
    if ((long long)T.n_cat * T.n_rings + (long long)T.n_rings * (T.n_rings - 1) / 2 > 0x7fffffffll) return CHC_E_TOO_LARGE;
    const size_t tables = 2 * al16(4 * (size_t)T.K * T.K) + al16(sizeof(PairMisc)) + ring_bytes + 4 * (size_t)rlist_cap;
    // ~3 cells per pair atom keeps the cells at the cutoff size for protein-like densities
    if (const char* e = getenv("CHACRA_MAX_CELLS")) cells_cap = std::min<size_t>(cells_cap, (size_t)std::max(2048, atoi(e)));   // test hook
    // systems the gather binning variant takes (more than 8 192 pair atoms) below 65 536 pair atoms: packed 16-bit counters,
    // twice the cells in the binning kernel's shared memory
    const bool gather = T.n_pair > 1024 * 8;
    B.count16 = T.n_pair <= 65535 && (gather || getenv("CHACRA_COUNT16")) && !getenv("CHACRA_COUNT32") ? 1 : 0;
    if (gather && B.count16 && cells_cap == 49152) cells_cap = 98304;
    size_t want_cells = std::min<size_t>(cells_cap, std::max<size_t>(2048, 3 * (size_t)T.n_pair));
    const size_t plain_bits = al16(4 * (size_t)((ctx->n_pairs + 31) / 32));
    // shared-memory path: the sorted atoms as slot pairs (x0 x1 y0 y1 | z0 z1) + flag word, filter key and residue rank per
    // slot (padded); the queues live in the float4 image of the record once that has been converted (when they fit there)
    const size_t e_pairs = ((size_t)T.n_pair + 1) / 2 + E1_PAD_PAIRS;
    const size_t partner_bytes = al16(4 * (size_t)E1_PARTNER_SETS * ((((size_t)T.n_pair + 2 * E1_PAD_PAIRS + 31) >> 5) + 2));
    const size_t bypidx_bytes = al16(4 * (size_t)T.n_pair) + al16(2 * (size_t)T.n_pair) + al16((size_t)T.n_pair);
    const size_t e_bytes = 2 * al16(16 * e_pairs) + partner_bytes + al16(4 * 2 * e_pairs) + bypidx_bytes;
    const bool alias_queues = queues <= 16 * (size_t)T.n_pair;
    ctx->small_path = false;
    for (size_t cells = std::min<size_t>(want_cells, 16384); cells >= 2048 && cells >= (size_t)T.n_pair; cells -= 1024) {
        size_t need = 256 + (size_t)B.fixed_bytes + al16(2 * cells) + plain_bits + (alias_queues ? 0 : queues) + tables + e_bytes;
        if (need <= SMEM_LIMIT) { ctx->small_path = true; want_cells = cells; break; }
    }
    if (getenv("CHACRA_FORCE_LARGE")) ctx->small_path = false;     // test hook: exercise the in-place path on small systems
    if (T.R > 65535) ctx->small_path = false;   // residue ranks are 16 bit there
    B.max_cells = (int)want_cells;
    B.cell16 = ctx->small_path ? 1 : 0;     // n_pair < 65 536 there by a wide margin; the large path keeps 32-bit starts
    const size_t cell_bytes = B.cell16 ? 2 : 4;
    B.stride = (long long)((B.off_cells + al16(cell_bytes * (size_t)B.max_cells) + 127) & ~size_t(127));
    size_t off = 0;
    queue_bytes(ctx->small_path, &q1b, &qfb, &q2b);
    bool queues_placed = false;
    if (ctx->small_path) {
        L.off_rec = (int)off;   off += al16((size_t)B.fixed_bytes); off = (off + 127) & ~size_t(127);
        L.off_cells = (int)off; off += al16(cell_bytes * (size_t)B.max_cells);
        L.off_lkey = 0;
        L.off_bits = (int)off;  off += plain_bits;
        L.bits_words = (int)((ctx->n_pairs + 31) / 32);
        L.off_exy = (int)off;   off += al16(16 * e_pairs);
        L.off_ez = (int)off;    off += al16(16 * e_pairs);
        L.off_ek = (int)off;    off += partner_bytes;
        L.off_sw = (int)off;    off += al16(4 * 2 * e_pairs);
        L.off_sres = 0;
        L.off_plkf = (int)off;  off += al16(4 * (size_t)T.n_pair);
        L.off_pres = (int)off;  off += al16(2 * (size_t)T.n_pair);
        L.off_pst = (int)off;   off += al16((size_t)T.n_pair);
        if (alias_queues) {
            L.off_q1 = L.off_qf = L.off_rec + B.off_spos;
            L.off_q2 = L.off_qf + (int)qfb;
            queues_placed = true;
        }
    } else {
        L.off_rec = L.off_cells = L.off_lkey = L.off_bits = L.off_sres = 0;
        L.off_plkf = L.off_pres = L.off_pst = 0;
        // the slot-pair arrays of the large path live in a per-CTA slice of a global scratch (L2 / L1), offsets from its start
        size_t g = 0;
        L.off_exy = (int)g; g += al16(16 * e_pairs);
        L.off_ez = (int)g;  g += al16(16 * e_pairs);
        L.off_sw = (int)g;  g += al16(4 * 2 * e_pairs);
        L.off_ek = (int)g;  g += partner_bytes;
        L.soa_stride = (long long)((g + 127) & ~size_t(127));
        if ((ctx->n_pairs + 23) / 24 > 0x7fffffff) return CHC_E_TOO_LARGE;
        L.bits_words = (int)((ctx->n_pairs + 23) / 24);
    }
    if (!queues_placed) {
        L.off_q1 = (int)off; off += q1b;
        L.off_qf = (int)off; off += qfb;
        L.off_q2 = (int)off; off += q2b;
    }
    L.off_tv = (int)off; off += al16(4 * (size_t)T.K * T.K);
    L.off_th = (int)off; off += al16(4 * (size_t)T.K * T.K);
    L.off_misc = (int)off; off += al16(sizeof(PairMisc));
    L.off_rc = (int)off; off += ring_bytes;
    L.off_rlist = (int)off; off += 4 * (size_t)rlist_cap;
    L.rlist_cap = rlist_cap;
    L.off_rgrid = L.off_rsort = 0;
    // coarse grid over the ring centroids (pair_kernel.cuh R1) once the all-pairs prefilter would dominate a frame
    const long long ring_tasks = (long long)T.n_cat * T.n_rings + (long long)T.n_rings * (T.n_rings - 1) / 2;
    if (!ctx->small_path && T.n_rings <= 65535 && (ring_tasks > 200000 || getenv("CHACRA_RING_GRID")) && !getenv("CHACRA_RING_ALLPAIRS")) {
        L.off_rgrid = (int)off; off += al16(4 * (size_t)(RG_CAP + 1));
        L.off_rsort = (int)off; off += al16(2 * (size_t)std::max(T.n_rings, 1));
    }
    L.bits_plain = 0;
    if (!ctx->small_path && ctx->n_pairs < (1ll << 31) && off + plain_bits + 1024 <= SMEM_LIMIT && !getenv("CHACRA_STAMPED_SET")) {
        // mid-size systems: the plain per-frame residue-pair set fits in shared memory beside the queues (no stamped global set)
        L.bits_plain = 1;
        L.off_bits = (int)off; off += plain_bits;
        L.bits_words = (int)((ctx->n_pairs + 31) / 32);
    }
    L.total = (int)off;
    // ---- fused kernel: no landing zone for a record; queues in their own space; aux atoms copied out of the frame ----
    ctx->fused_ok = false;
    if (ctx->small_path && T.n_pair <= FUSED_APT * PAIR_THREADS && 3 * T.n_aux <= FUSED_AUX * PAIR_THREADS && B.max_cells + 1 < 4 * 8 * PAIR_THREADS && getenv("CHACRA_FUSED") && !getenv("CHACRA_BIN_VARIANT")) {   // opt-in (CHACRA_FUSED=1): measured 7 % slower than the two-kernel path, DESIGN.md 3
        PairLayout& G = ctx->LF;
        G = L;
        size_t o = 0;
        G.off_rec = 0;
        G.off_cells = (int)o; o += al16(cell_bytes * (size_t)B.max_cells + 32);
        G.off_cnt = (int)o;   o += al16(cell_bytes * (size_t)B.max_cells + 32);
        G.off_bits = (int)o;  o += plain_bits;
        G.off_exy = (int)o;   o += al16(16 * e_pairs);
        G.off_ez = (int)o;    o += al16(16 * e_pairs);
        G.off_ek = (int)o;    o += partner_bytes;
        G.off_sw = (int)o;    o += al16(4 * 2 * e_pairs);
        G.off_plkf = 0;
        G.off_pres = (int)o;  o += al16(2 * (size_t)T.n_pair);
        G.off_pst = (int)o;   o += al16((size_t)T.n_pair);
        G.off_aux = (int)o;   o += al16(12 * (size_t)std::max(T.n_aux, 1));
        G.off_q1 = G.off_qf = (int)o; o += qfb;
        G.off_q2 = (int)o;    o += q2b;
        G.off_tv = (int)o;    o += al16(4 * (size_t)T.K * T.K);
        G.off_th = (int)o;    o += al16(4 * (size_t)T.K * T.K);
        G.off_misc = (int)o;  o += al16(sizeof(PairMisc));
        G.off_rc = (int)o;    o += ring_bytes;
        G.off_rlist = (int)o; o += 4 * (size_t)rlist_cap;
        G.total = (int)o;
        ctx->fused_ok = (size_t)G.total <= SMEM_LIMIT;
    }
    ctx->bin_smem = (int)al16((gather && B.count16 ? 2 : 4) * (size_t)B.max_cells + 16);
    if ((size_t)L.total > SMEM_LIMIT || (size_t)ctx->bin_smem + 1024 > SMEM_LIMIT) return CHC_E_TOO_LARGE;
    return CHC_OK;
}

// names the build of the contact kernels (bin_kernel.cuh, bin_tma_kernel.cuh, pair_kernel.cuh, frame_*.cuh, water_phase.cuh, plan()):
// bump it with every change to those, and re-capture profiles/ncu_traffic.json (tools/ncu_traffic.py)
#ifndef CHC_BUILD_TAG
#define CHC_BUILD_TAG "r02_v6"
#endif
extern "C" int chc_abi_version(void) { return CHC_ABI_VERSION; }
extern "C" const char* chc_build_tag(void) { return CHC_BUILD_TAG; }
extern "C" const char* chc_last_error(const chc_ctx* ctx) { return ctx ? ctx->err : "null context"; }
extern "C" int64_t chc_n_pairs(const chc_ctx* ctx) { return ctx ? ctx->n_pairs : 0; }

extern "C" void chc_destroy(chc_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    for (void* p : ctx->dev_allocs) cudaFree(p);
    if (ctx->counts_owned) cudaFree(ctx->counts);
    cudaFree(ctx->contacts); cudaFree(ctx->contact_n); cudaFree(ctx->near); cudaFree(ctx->near_n);
    cudaFree(ctx->hpmask); cudaFree(ctx->binned); cudaFree(ctx->gstamp); cudaFree(ctx->gstate);
    cudaFree(ctx->wscratch); cudaFree(ctx->status); cudaFree(ctx->open_idx); cudaFree(ctx->gsoa);
    for (int k = 0; k < 2; ++k) {
        cudaFree(ctx->dev_stage[k]); if (ctx->pin_stage[k]) cudaFreeHost(ctx->pin_stage[k]);
        if (ctx->stage_done[k]) cudaEventDestroy(ctx->stage_done[k]);
        if (ctx->copied[k]) cudaEventDestroy(ctx->copied[k]);
    }
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    for (auto& e : ctx->ev_pending) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); cudaEventDestroy(e.c); }
    for (auto& e : ctx->ev_free) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); cudaEventDestroy(e.c); }
    delete ctx;
}

extern "C" int chc_create(chc_ctx** out, const chc_topology* t, const chc_params* p, int device) {
    if (!out) return CHC_E_INVALID;
    *out = nullptr;
    chc_ctx* ctx = new (std::nothrow) chc_ctx();
    if (!ctx) return CHC_E_NOMEM;
    *out = ctx;                     // handed back even on failure so that chc_last_error works; caller destroys
    if (!t || !p) return fail(ctx, CHC_E_INVALID, "null topology or params");
    if (t->n_atoms <= 0 || t->n_pair < 0 || t->n_elem < 1 || t->n_elem > 15 || t->n_residues < 1)
        return fail(ctx, CHC_E_INVALID, "bad topology sizes (atoms %d, pair %d, elem %d, residues %d)", t->n_atoms, t->n_pair, t->n_elem, t->n_residues);
    if (p->n_states < 1 || p->fragment_frames < 1) return fail(ctx, CHC_E_INVALID, "n_states and fragment_frames must be >= 1");
    ctx->device = device;
    ctx->prm = *p;
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return fail(ctx, CHC_E_CUDA, "device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major, prop.minor);
    ctx->sm_count = prop.multiProcessorCount;

    DevTopo& T = ctx->T;
    T.n_atoms = t->n_atoms; T.n_pair = t->n_pair; T.n_aux = t->n_aux; T.n_h = t->n_h_slots; T.n_rings = t->n_rings;
    T.n_cat = t->n_cations; T.n_dis = t->n_disulfide; T.K = t->n_elem; T.R = t->n_residues; T.n_hp_base = t->n_hp_base;
    T.same = t->same_selection;
    T.n_pair_a = t->n_pair_a; T.n_solv = t->n_solvent;
    if (T.n_pair_a < 0 || T.n_pair_a > T.n_pair) return fail(ctx, CHC_E_INVALID, "bad n_pair_a %d", T.n_pair_a);
    int rc;
    const size_t NP = t->n_pair, NR = t->n_rings, NC = t->n_cations, K2 = (size_t)t->n_elem * t->n_elem;
#define UP(field, src, n) if ((rc = upload(ctx, src, (size_t)(n), &T.field)) != CHC_OK) return rc
    // I_PBOND: a polar pair atom (hydrogen-bond selection, ion, solvent) bonded to another polar pair atom -- only such pairs ever
    // need the bonded-pair test of measure contacts / measure hbonds in the polar classes (frame_phases.cuh e3_finish)
    std::vector<uint32_t> info(t->pair_info, t->pair_info + NP);
    {
        const uint32_t polar = I_HBU | I_ANION | I_CATION | I_SOLV;
        std::vector<int> a2p((size_t)t->n_atoms, -1);
        for (size_t k = 0; k < NP; ++k) if (t->pair_atom[k] >= 0 && t->pair_atom[k] < t->n_atoms) a2p[t->pair_atom[k]] = (int)k;
        for (size_t k = 0; k < NP; ++k) {
            // I_XBOND, recomputed for the van der Waals / hydrophobic classes: a bond to a pair atom that the residue-distance
            // filter of those classes (|dlkey| < vdw_res_diff) does not already exclude (disulfides, links between chains)
            info[k] &= ~(I_PBOND | I_XBOND);
            for (int b = t->pair_bond_start[k]; b < t->pair_bond_start[k + 1]; ++b) {
                const int a = t->pair_bond_adj[b];
                const int q = (a >= 0 && a < t->n_atoms) ? a2p[a] : -1;
                if (q < 0) continue;
                if ((info[k] & polar) && (t->pair_info[q] & polar)) info[k] |= I_PBOND;
                if (std::abs(t->pair_lkey[k] - t->pair_lkey[q]) >= p->vdw_res_diff) info[k] |= I_XBOND;
            }
        }
    }
    UP(pair_atom, t->pair_atom, NP); UP(pair_info, info.data(), NP); UP(pair_lkey, t->pair_lkey, NP);
    UP(pair_res, t->pair_res, NP); UP(pair_resid, t->pair_resid, NP); UP(pair_h_start, t->pair_h_start, NP + 1);
    UP(pair_bond_start, t->pair_bond_start, NP + 1); UP(pair_bond_adj, t->pair_bond_adj, t->n_bond_adj);
    UP(aux_atom, t->aux_atom, t->n_aux); UP(pair_solv_ord, t->pair_solv_ord, NP);
    UP(ring_aux, t->ring_aux, 3 * NR); UP(ring_pc_aux, t->ring_pc_aux, 3 * NR); UP(ring_sel_aux, t->ring_sel_aux, 3 * NR);
    UP(ring_sel_flags, t->ring_sel_flags, 3 * NR); UP(ring_atoms, t->ring_atoms, 3 * NR); UP(ring_res, t->ring_res, NR);
    UP(ring_rep, t->ring_rep, NR); UP(ring_rep_aux, t->ring_rep_aux, NR);
    UP(cat_aux, t->cation_aux, NC); UP(cat_atom, t->cation_atom, NC); UP(cat_flags, t->cation_flags, NC); UP(cat_res, t->cation_res, NC); UP(cat_excl, t->cation_excl_ring, NC);
    UP(disulfide, t->disulfide, 2 * (size_t)t->n_disulfide);
    {
        std::vector<unsigned short> off((3 * (size_t)std::max(t->n_aux, 0) + 7) / 8 * 8 + 8, 0);
        for (int k = 0; k < t->n_aux; ++k)
            for (int d = 0; d < 3; ++d) off[3 * (size_t)k + d] = (unsigned short)(3 * (long long)t->aux_atom[k] + d);   // used only when 3 n_atoms < 65536
        UP(aux_off16, off.data(), off.size());
    }
    UP(vcut, t->vdw_cut, K2); UP(hp_base, t->hp_base, t->n_hp_base); UP(atom_class, t->atom_class, t->n_atoms);
    // fold the measure-contacts cutoffs into the per-element-pair acceptance thresholds
    DevParams& P = ctx->P;
    auto sq = [](double c) { float f = (float)c; return f * f; };
    P.soft2 = sq(p->vdw_soft_cutoff); P.hp2 = sq(p->hp_dist); P.sb2 = sq(p->sb_dist); P.hb2 = sq(p->hb_dist);
    {
        std::vector<float> tv(K2), th(K2);
        for (size_t k = 0; k < K2; ++k) { tv[k] = std::min(t->vdw_d2max[k], P.soft2); th[k] = std::min(t->vdw_d2max[k], P.hp2); }
        UP(tv, tv.data(), K2); UP(th, th.data(), K2);
    }
#undef UP
    P.itypes = p->itype_mask; P.vdw_rd = p->vdw_res_diff; P.hb_rd = p->hbond_res_diff; P.emit = p->emit_contacts; P.track_near = p->track_near;
    double cmax = 0.5;
    if (p->itype_mask & CHC_IT_VDW) cmax = std::max(cmax, p->vdw_soft_cutoff);
    if (p->itype_mask & CHC_IT_HP) cmax = std::max(cmax, p->hp_dist);
    if (p->itype_mask & CHC_IT_SB) cmax = std::max(cmax, p->sb_dist);
    if (p->itype_mask & CHC_IT_HB) cmax = std::max(cmax, p->hb_dist);
    P.cell = (float)(cmax * 1.0005 + 1e-3);
    P.pre2 = 0.f;                         // the largest exact float32 cutoff^2 among the enabled pair types
    if (p->itype_mask & CHC_IT_VDW) P.pre2 = std::max(P.pre2, P.soft2);
    if (p->itype_mask & CHC_IT_HP) P.pre2 = std::max(P.pre2, P.hp2);
    if (p->itype_mask & CHC_IT_SB) P.pre2 = std::max(P.pre2, P.sb2);
    if (p->itype_mask & CHC_IT_HB) P.pre2 = std::max(P.pre2, P.hb2);
    if (p->track_near) {                  // the near-cutoff band just ABOVE a distance cutoff must survive the prefilter too
        const double band = 4.0 * std::max(p->near_dist_tol, 0.0);
        P.pre2 = std::max(P.pre2, sq(cmax + band));
    }
    P.hb_angle = (float)p->hb_angle;
    P.hb_fast = P.hb_angle < 89.0f && P.hb_angle > 1.0f;
    if (P.hb_fast) {
        const double t = std::tan((double)P.hb_angle / (double)57.2958f);
        P.hb_tan_lo = (float)(t * (1.0 - 4e-4)); P.hb_tan_hi = (float)(t * (1.0 + 4e-4));
    }
    P.elem_ch_mask = 0;
    for (int k = 0; k < t->n_elem; ++k) if (t->elem_class && t->elem_class[k]) P.elem_ch_mask |= 1u << k;
    double rmax = std::max(p->pc_dist, std::max(p->ps_dist, p->ts_dist)) + 0.05;
    P.ring_pre2 = (float)(rmax * rmax);
    P.pc_soft2 = sq(10.0); P.ps_soft2 = sq(10.0); P.ts_soft2 = sq(6.0);   // pi_cation.py:22, aromatics.py:26-27
    P.pc_soft = 10.0; P.ps_soft = 10.0; P.ts_soft = 6.0;
    P.pc_dist = p->pc_dist; P.pc_angle = p->pc_angle; P.ps_dist = p->ps_dist; P.ps_angle = p->ps_angle; P.ps_psi = p->ps_psi;
    P.ts_dist = p->ts_dist; P.ts_angle = p->ts_angle; P.ts_psi = p->ts_psi;
    P.sb_dist = p->sb_dist; P.hb_dist = p->hb_dist; P.hp_dist = p->hp_dist; P.soft = p->vdw_soft_cutoff;
    P.near_d = p->near_dist_tol; P.near_a = p->near_angle_tol;

    ctx->n_pairs = (int64_t)T.R * (T.R + 1) / 2;
    for (size_t k = 0; k < NP; ++k) ctx->lkey_max = std::max(ctx->lkey_max, t->pair_lkey[k]);
    rc = plan(ctx);
    if (rc != CHC_OK) return fail(ctx, rc, "topology with %d pair atoms / %d residues does not fit any kernel path", T.n_pair, T.R);
    CK(cudaFuncSetAttribute(pair_kernel_fn(ctx->small_path, p->track_near != 0 || p->emit_contacts != 0), cudaFuncAttributeMaxDynamicSharedMemorySize, ctx->L.total));
    if (ctx->fused_ok)
        CK(cudaFuncSetAttribute(pair_kernel_fn(true, p->track_near != 0 || p->emit_contacts != 0, true), cudaFuncAttributeMaxDynamicSharedMemorySize, ctx->LF.total));
    // binning kernel variant: registers hold the frame's pair atoms when they fit (one HBM pass), else three gather passes
    ctx->bin_variant = T.n_pair <= 512 * 8 ? 1 : (T.n_pair <= 1024 * 8 ? 2 : 0);
    // frames that fit in shared memory stream through bulk async copies (bin_tma_kernel.cuh)
    auto plan_bin_tma = [&](int variant) -> bool {
        BinTmaLayout& Y = ctx->Y;
        const int threads = variant == 3 ? 256 : 512, nbuf = variant == 3 ? 1 : 2, ctas = variant == 3 ? 2 : 1;
        const int apt_need = (T.n_pair + threads - 1) / threads;
        if (T.n_pair < 1 || apt_need > 16 || 3 * (size_t)T.n_atoms >= 65536) return false;
        static const int apts[] = {2, 4, 6, 8, 10, 12, 13, 14, 16};
        int apt = 16;
        for (int a : apts) if (a >= apt_need) { apt = a; break; }
        if (variant == 4) apt = apt_need <= 8 ? 8 : 16;
        Y.buf_bytes = (int)al16(std::max<size_t>(12 * (size_t)T.n_atoms + 28, (size_t)ctx->B.off_aux + 4 * (size_t)std::max(T.n_hp_base, 0)));
        // as many counters as fit (a full grid needs B.max_cells of them; fewer only means coarser cells for wide frames)
        const size_t auxtab = al16(6 * (size_t)std::max(T.n_aux, 1));
        for (int cap = std::min(BT_MAXG, (ctx->B.max_cells + 8 * threads) / (8 * threads)); cap >= 1; --cap) {
            Y.cap_groups = cap;
            Y.max_cells = std::min(ctx->B.max_cells, 8 * cap * threads - 1);
            Y.off_auxtab = 16 * cap * threads;
            Y.off_buf0 = (int)(((size_t)Y.off_auxtab + auxtab + 127) & ~size_t(127));
            Y.off_buf1 = nbuf == 2 ? (int)(((size_t)Y.off_buf0 + Y.buf_bytes + 127) & ~size_t(127)) : Y.off_buf0;
            Y.total = Y.off_buf1 + Y.buf_bytes;
            // ~1.5 cells per pair atom keep the cell edge at the cutoff for protein-like densities
            if (Y.max_cells < 2048 || 2 * (long long)Y.max_cells < 3 * (long long)T.n_pair) break;
            if ((size_t)ctas * ((size_t)Y.total + 1024 + 512) <= SMEM_LIMIT) { ctx->tma_apt = apt; return true; }
        }
        return false;
    };
    if (plan_bin_tma(3)) ctx->bin_variant = 3;
    else if (plan_bin_tma(4)) ctx->bin_variant = 4;
    if (const char* e = getenv("CHACRA_BIN_VARIANT")) {           // test hook
        const int v = atoi(e);
        if (v < 3 || plan_bin_tma(v)) ctx->bin_variant = v;
    }
    if (ctx->bin_variant >= 3 && ctx->Y.max_cells + 1 < ctx->B.max_cells) {
        // the record need not hold more cell starts than this binning kernel can count
        const int v = ctx->bin_variant;
        const bool was_small = ctx->small_path;
        rc = plan(ctx, (size_t)ctx->Y.max_cells + 1);
        if (rc != CHC_OK || ctx->small_path != was_small || !plan_bin_tma(v)) return fail(ctx, CHC_E_INVALID, "internal: cell capacity replanning failed");
    }
    ctx->bin_tma = ctx->bin_variant >= 3;
    if (ctx->bin_tma) ctx->bin_smem = ctx->Y.total;
    const void* bin_fn = bin_kernel_fn(ctx->bin_variant, ctx->tma_apt, &ctx->bin_threads);
    CK(cudaFuncSetAttribute(bin_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, ctx->bin_smem));
    ctx->pair_grid = ctx->sm_count;
    if (const char* e = getenv("CHACRA_PAIR_GRID")) ctx->pair_grid = std::max(1, std::min(ctx->sm_count, atoi(e)));   // test hook
    {
        int per_sm = 1;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bin_fn, ctx->bin_threads, ctx->bin_smem));
        ctx->bin_grid = ctx->sm_count * std::max(1, per_sm);
    }
    // frames per (bin, pair) launch pair: large enough that launch latency, the kernels' prologues and their tails are
    // amortised (the pair kernel is issue bound and does not care whether the records come from L2 or HBM), bounded by
    // ~1.5 GB of binned records and 80 frames per SM
    {
        int64_t by_mem = (int64_t)((1500ull << 20) / (size_t)ctx->B.stride);
        int64_t g = ctx->pair_grid;
        ctx->chunk_frames = std::max<int64_t>(g, std::min<int64_t>(by_mem, 80 * g) / g * g);
        if (const char* e = getenv("CHACRA_CHUNK_FRAMES")) ctx->chunk_frames = std::max<int64_t>(1, atoll(e));
    }
    CK(cudaMalloc(&ctx->binned, (size_t)ctx->chunk_frames * (size_t)ctx->B.stride));
    CK(cudaMalloc(&ctx->status, 4)); CK(cudaMemset(ctx->status, 0, 4));
    if (T.n_solv > 0) {
        // the solvent shell of the reference (3.5 A, hbonds.py:24) is implied by the hydrogen-bond cutoff only up to 3.5 A
        if ((p->itype_mask & CHC_IT_HB) && p->hb_dist > 3.5)
            return fail(ctx, CHC_E_INVALID, "hbond cutoff %.3f A above the 3.5 A water-shell radius is not supported with solvent", p->hb_dist);
        size_t ints = (size_t)ctx->pair_grid * (size_t)T.n_solv * (3 + WADJ);
        CK(cudaMalloc(&ctx->wscratch, ints * 4)); CK(cudaMemset(ctx->wscratch, 0, ints * 4));
    }
    if (ctx->lkey_max >= (1 << 24))
        return fail(ctx, CHC_E_TOO_LARGE, "residue ids span more than 2^24 (locality keys are carried as exact floats)");
    if (!ctx->small_path) {
        CK(cudaMalloc(&ctx->gsoa, (size_t)ctx->pair_grid * (size_t)ctx->L.soa_stride));
        size_t words = (size_t)ctx->pair_grid * (size_t)ctx->L.bits_words;
        CK(cudaMalloc(&ctx->gstamp, words * 4)); CK(cudaMemset(ctx->gstamp, 0, words * 4));
        CK(cudaMalloc(&ctx->gstate, (size_t)ctx->pair_grid * 4)); CK(cudaMemset(ctx->gstate, 0, (size_t)ctx->pair_grid * 4));
    }
    CK(cudaMalloc(&ctx->counts, (size_t)p->n_states * ctx->n_pairs * sizeof(int32_t)));
    ctx->counts_owned = true;
    CK(cudaMemset(ctx->counts, 0, (size_t)p->n_states * ctx->n_pairs * sizeof(int32_t)));
    CK(cudaMalloc(&ctx->contact_n, 8)); CK(cudaMemset(ctx->contact_n, 0, 8));
    CK(cudaMalloc(&ctx->near_n, 8)); CK(cudaMemset(ctx->near_n, 0, 8));
    CK(cudaMalloc(&ctx->contacts, std::max<int64_t>(p->contact_capacity, 1) * sizeof(chc_contact)));
    CK(cudaMalloc(&ctx->near, std::max<int64_t>(p->near_capacity, 1) * sizeof(chc_near)));
    ctx->st.path = ctx->small_path ? 1 : 2; ctx->st.smem_bytes = ctx->L.total; ctx->st.threads = PAIR_THREADS;
    ctx->st.grid = ctx->pair_grid; ctx->st.bin_smem_bytes = ctx->bin_smem; ctx->st.bin_grid = ctx->bin_grid;
    ctx->st.chunk_frames = ctx->chunk_frames; ctx->st.record_bytes = ctx->B.stride;
    ctx->st.bin_variant = ctx->bin_variant; ctx->st.bin_apt = ctx->bin_tma ? ctx->tma_apt : 0;
    ctx->st.fused = ctx->fused_ok && (!((p->itype_mask & CHC_IT_HP) && t->n_hp_base > 0) || (ctx->bin_tma && t->same_selection)) ? 1 : 0;
    ctx->st.fused_smem_bytes = ctx->fused_ok ? ctx->LF.total : 0;
    return CHC_OK;
}

// ---- counts -------------------------------------------------------------------------------------------
extern "C" int chc_bind_counts(chc_ctx* ctx, int32_t* d_counts, int64_t n_elems) {
    if (!ctx || !d_counts) return CHC_E_INVALID;
    if (n_elems != (int64_t)ctx->prm.n_states * ctx->n_pairs)
        return fail(ctx, CHC_E_INVALID, "bind_counts: expected %lld elements, got %lld", (long long)ctx->prm.n_states * ctx->n_pairs, (long long)n_elems);
    CK(cudaSetDevice(ctx->device));
    if (ctx->counts_owned) { cudaFree(ctx->counts); ctx->counts_owned = false; }
    ctx->counts = d_counts;
    return CHC_OK;
}

extern "C" int chc_reset_counts(chc_ctx* ctx, void* stream) {
    if (!ctx) return CHC_E_INVALID;
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemsetAsync(ctx->counts, 0, (size_t)ctx->prm.n_states * ctx->n_pairs * sizeof(int32_t), (cudaStream_t)stream));
    return CHC_OK;
}

extern "C" int chc_counts_device(chc_ctx* ctx, int32_t** d_counts, int64_t* n_pairs) {
    if (!ctx) return CHC_E_INVALID;
    if (d_counts) *d_counts = ctx->counts;
    if (n_pairs) *n_pairs = ctx->n_pairs;
    return CHC_OK;
}

extern "C" int chc_read_counts(chc_ctx* ctx, int state, int32_t* h_counts, int64_t n_pairs, void* stream) {
    if (!ctx || !h_counts || state < 0 || state >= ctx->prm.n_states || n_pairs != ctx->n_pairs)
        return fail(ctx, CHC_E_INVALID, "read_counts: bad state or size");
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpyAsync(h_counts, ctx->counts + (size_t)state * ctx->n_pairs, (size_t)n_pairs * 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CK(cudaStreamSynchronize((cudaStream_t)stream));
    return CHC_OK;
}

// ---- frames ------------------------------------------------------------------------------------------
static int ensure_hpmask(chc_ctx* ctx, int64_t frags) {
    if (frags <= ctx->hpmask_frags) return CHC_OK;
    unsigned char* fresh = nullptr;
    size_t np = std::max(ctx->T.n_pair, 1);
    CK(cudaMalloc(&fresh, (size_t)frags * np));
    CK(cudaMemset(fresh, 0, (size_t)frags * np));
    if (ctx->hpmask) {
        CK(cudaDeviceSynchronize());       // earlier kernels may still read the old masks
        CK(cudaMemcpy(fresh, ctx->hpmask, np, cudaMemcpyDeviceToDevice));   // slot 0 carries a continued fragment
        cudaFree(ctx->hpmask);
    }
    ctx->hpmask = fresh; ctx->hpmask_frags = frags;
    return CHC_OK;
}

static int submit_impl(chc_ctx* ctx, int state, const float* d_xyz, const long long* d_src, int64_t n_frames, int64_t first_index,
                       int64_t frame_stride, void* stream_) {
    if (!ctx) return CHC_E_INVALID;
    if (n_frames == 0) return CHC_OK;
    if (!d_xyz || n_frames < 0 || first_index < 0 || state < 0 || state >= ctx->prm.n_states)
        return fail(ctx, CHC_E_INVALID, "submit_frames: bad arguments (state %d, frames %lld, first %lld)", state, (long long)n_frames, (long long)first_index);
    if (n_frames > 0x7fffffff) return fail(ctx, CHC_E_INVALID, "submit_frames: at most 2^31-1 frames per call");
    cudaStream_t stream = (cudaStream_t)stream_;
    CK(cudaSetDevice(ctx->device));
    const int F = ctx->prm.fragment_frames;
    if (frame_stride == 0) frame_stride = 3 * (int64_t)ctx->T.n_atoms;
    const int64_t frag0 = first_index / F, frag1 = (first_index + n_frames - 1) / F;
    const bool continues = (first_index % F) != 0;
    if (continues && !(ctx->last_state == state && ctx->next_index == first_index))
        return fail(ctx, CHC_E_ORDER, "submit_frames: chunk starts at %lld, inside a %d-frame fragment it does not continue", (long long)first_index, F);
    if (continues && ctx->hp_last_slot > 0) {
        // the fragment being continued was computed into slot hp_last_slot by the previous call; it moves to slot 0 BEFORE the
        // mask buffer may grow (ensure_hpmask carries slot 0 over into a fresh buffer)
        CK(cudaMemcpyAsync(ctx->hpmask, ctx->hpmask + (size_t)ctx->hp_last_slot * ctx->T.n_pair, ctx->T.n_pair, cudaMemcpyDeviceToDevice, stream));
        ctx->hp_last_slot = 0;
    }
    int rc = ensure_hpmask(ctx, frag1 - frag0 + 1);
    if (rc != CHC_OK) return rc;
    ctx->hp_last_slot = (int)(frag1 - frag0);
    DevOut O;
    O.counts = ctx->counts + (size_t)state * ctx->n_pairs; O.n_pairs = ctx->n_pairs;
    O.contacts = ctx->contacts; O.contact_n = ctx->contact_n; O.contact_cap = ctx->prm.contact_capacity;
    O.near = ctx->near; O.near_n = ctx->near_n; O.near_cap = ctx->prm.near_capacity;
    O.status = ctx->status;
    // chunks of frames: the binning kernel streams the coordinates and writes the chunk's binned records, the pair
    // kernel consumes them (from L2 when the chunk fits)
    for (int64_t c0 = 0; c0 < n_frames; c0 += ctx->chunk_frames) {
        const int n = (int)std::min<int64_t>(ctx->chunk_frames, n_frames - c0);
        const float* x = d_src ? d_xyz : d_xyz + (size_t)c0 * frame_stride;     // indexed: offsets count from the caller's base
        const long long* src = d_src ? d_src + c0 : nullptr;
        chc_ctx::Ev3 ev{nullptr, nullptr, nullptr};
        if (ctx->timing) {
            if (!ctx->ev_free.empty()) { ev = ctx->ev_free.back(); ctx->ev_free.pop_back(); }
            else { CK(cudaEventCreate(&ev.a)); CK(cudaEventCreate(&ev.b)); CK(cudaEventCreate(&ev.c)); }
            CK(cudaEventRecord(ev.a, stream));
        }
        const bool hp_on = (ctx->P.itypes & CHC_IT_HP) && ctx->T.n_hp_base > 0;
        const bool fuse_hp = hp_on && ctx->bin_tma && ctx->T.same;      // the binning CTA answers the 1.7 A query itself
        // Fused path: the pair kernel sorts every frame itself; only the frames that OPEN a fragment go through the binning
        // kernel first, for the hydrophobic masks of their fragments (the binning kernel with fragment length 1 and a source index).
        const bool fused = ctx->fused_ok && !d_src && (!hp_on || fuse_hp);
        auto launch_bin = [&](const float* bx, const long long* bsrc, int nn, long long fi, int ff, long long slot0, int open_phase, int fuse) -> int {
            long long fs = frame_stride, fr0 = frag0;
            BinTmaLayout Y = ctx->Y;
            Y.open_phase = open_phase;
            void* args_reg[] = {&ctx->T, &ctx->P, &ctx->B, &bx, &fs, &nn, &fi, &ff, &fr0, &ctx->binned, &ctx->hpmask, &bsrc};
            void* args_tma[] = {&ctx->T, &ctx->P, &ctx->B, &Y, &bx, &fs, &nn, &ff, &slot0, &ctx->binned, &ctx->hpmask, &fuse, &bsrc};
            CK(cudaLaunchKernel(bin_kernel_fn(ctx->bin_variant, ctx->tma_apt, nullptr), dim3(std::min(nn, ctx->bin_grid)), dim3(ctx->bin_threads),
                                ctx->bin_tma ? args_tma : args_reg, ctx->bin_smem, stream));
            CK(cudaGetLastError());
            ctx->st.kernel_launches++;
            return CHC_OK;
        };
        const long long fi0 = first_index + c0;
        if (fused) {
            const int64_t f0 = (F - fi0 % F) % F;                 // first fragment-opening frame of the chunk
            if (hp_on && f0 < n) {
                const int n_open = (int)((n - f0 + F - 1) / F);
                if (n_open > ctx->open_cap) {
                    CK(cudaStreamSynchronize(stream));
                    cudaFree(ctx->open_idx);
                    ctx->open_cap = std::max<int64_t>(256, 2 * (int64_t)n_open);
                    CK(cudaMalloc(&ctx->open_idx, (size_t)ctx->open_cap * sizeof(long long)));
                }
                iota_stride_kernel<<<(n_open + 255) / 256, 256, 0, stream>>>(ctx->open_idx, (long long)f0, (long long)F, n_open);
                CK(cudaGetLastError());
                int rcb = launch_bin(x, ctx->open_idx, n_open, fi0 + f0, 1, (fi0 + f0) / F - frag0, 0, 1);
                if (rcb != CHC_OK) return rcb;
            }
            if (ctx->timing) CK(cudaEventRecord(ev.b, stream));
            int nn = n; long long fi = fi0, fr0 = frag0, fs = frame_stride; int ff = F;
            unsigned* gs = nullptr; unsigned* gt = nullptr;
            void* args[] = {&ctx->T, &ctx->P, &O, &ctx->LF, &ctx->B, &ctx->binned, &nn, &fi, &ff, &fr0, &ctx->hpmask, &gs, &gt, &ctx->wscratch, &x, &fs, &ctx->gsoa};
            CK(cudaLaunchKernel(pair_kernel_fn(true, ctx->P.track_near != 0 || ctx->P.emit != 0, true), dim3(std::min(n, ctx->pair_grid)), dim3(PAIR_THREADS),
                                args, ctx->LF.total, stream));
            ctx->st.kernel_launches++;
        } else {
            int rcb = launch_bin(x, src, n, fi0, F, fi0 / F - frag0, (int)(fi0 % F), fuse_hp ? 1 : 0);
            if (rcb != CHC_OK) return rcb;
            if (ctx->timing) CK(cudaEventRecord(ev.b, stream));        // the hydrophobic query kernel (below) is timed with the pair kernel
            if (hp_on && !fuse_hp) {
                const int64_t f0 = (F - fi0 % F) % F;                 // first fragment-opening frame of the chunk
                if (f0 < n) {
                    const int n_open = (int)((n - f0 + F - 1) / F);
                    const int by = std::max(1, std::min(1024, (ctx->T.n_hp_base + 7) / 8));     // one base carbon per warp
                    hp_mask_kernel<<<dim3(n_open, by), 256, 0, stream>>>(ctx->T, ctx->P, ctx->B, x, frame_stride, n, fi0, F, frag0,
                                                                         ctx->binned, ctx->hpmask, src);
                    CK(cudaGetLastError());
                    ctx->st.kernel_launches++;
                }
            }
            const int grid = std::min(n, ctx->pair_grid);
            int nn = n; long long fi = fi0, fr0 = frag0, fs = frame_stride; int ff = F;
            unsigned* gs = ctx->small_path ? nullptr : ctx->gstamp; unsigned* gt = ctx->small_path ? nullptr : ctx->gstate;
            void* args[] = {&ctx->T, &ctx->P, &O, &ctx->L, &ctx->B, &ctx->binned, &nn, &fi, &ff, &fr0, &ctx->hpmask, &gs, &gt, &ctx->wscratch, &x, &fs, &ctx->gsoa};
            CK(cudaLaunchKernel(pair_kernel_fn(ctx->small_path, ctx->P.track_near != 0 || ctx->P.emit != 0), dim3(grid), dim3(PAIR_THREADS), args, ctx->L.total, stream));
            ctx->st.kernel_launches++;
        }
        CK(cudaGetLastError());
        if (ctx->timing) { CK(cudaEventRecord(ev.c, stream)); ctx->ev_pending.push_back(ev); }
    }
    ctx->st.frames_submitted += n_frames;
    ctx->last_state = state; ctx->next_index = first_index + n_frames;
    return CHC_OK;
}

extern "C" int chc_submit_frames(chc_ctx* ctx, int state, const float* d_xyz, int64_t n_frames, int64_t first_index,
                                 int64_t frame_stride, void* stream_) {
    return submit_impl(ctx, state, d_xyz, nullptr, n_frames, first_index, frame_stride, stream_);
}

// frame f of the state trajectory is the stored frame d_source_index[f] of d_xyz (all replicas' frames in
// one device buffer).  Replaces get_state_coordinates_from_replica + write_state_trajectory
// (chacra/trajectories/process_hremd.py:184-230): no state trajectory is ever materialised.
extern "C" int chc_submit_frames_indexed(chc_ctx* ctx, int state, const float* d_xyz, const int64_t* d_source_index, int64_t n_frames,
                                         int64_t first_index, int64_t frame_stride, void* stream_) {
    if (ctx && n_frames > 0 && !d_source_index) return fail(ctx, CHC_E_INVALID, "submit_frames_indexed: null source index");
    static_assert(sizeof(long long) == sizeof(int64_t), "index width");
    return submit_impl(ctx, state, d_xyz, reinterpret_cast<const long long*>(d_source_index), n_frames, first_index, frame_stride, stream_);
}

extern "C" int chc_submit_frames_host(chc_ctx* ctx, int state, const float* h_xyz, int64_t n_frames, int64_t first_index, void* stream_) {
    if (!ctx) return CHC_E_INVALID;
    if (n_frames == 0) return CHC_OK;
    if (!h_xyz || n_frames < 0) return fail(ctx, CHC_E_INVALID, "submit_frames_host: bad arguments");
    cudaStream_t stream = (cudaStream_t)stream_;
    CK(cudaSetDevice(ctx->device));
    const int F = ctx->prm.fragment_frames;
    const size_t fbytes = 12 * (size_t)ctx->T.n_atoms;
    // chunk = whole fragments, about 64 MiB: small enough that the copy of chunk k+1 hides behind the kernels of chunk k
    // from the first chunk on (the host path is PCIe bound, not kernel bound)
    int64_t chunk = std::max<int64_t>(1, (int64_t)((64u << 20) / fbytes) / F) * F;
    if (ctx->stage_frames < chunk) {
        CK(cudaDeviceSynchronize());
        for (int k = 0; k < 2; ++k) {
            cudaFree(ctx->dev_stage[k]); if (ctx->pin_stage[k]) cudaFreeHost(ctx->pin_stage[k]);
            CK(cudaMalloc(&ctx->dev_stage[k], chunk * fbytes));
            CK(cudaMallocHost(&ctx->pin_stage[k], chunk * fbytes));
            if (!ctx->stage_done[k]) CK(cudaEventCreateWithFlags(&ctx->stage_done[k], cudaEventDisableTiming));
        }
        ctx->stage_frames = chunk;
    }
    if (!ctx->copy_stream) {
        CK(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        for (int k = 0; k < 2; ++k) CK(cudaEventCreateWithFlags(&ctx->copied[k], cudaEventDisableTiming));
    }
    cudaPointerAttributes attr;
    bool pinned = cudaPointerGetAttributes(&attr, h_xyz) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError();
    int64_t done = 0; int buf = 0;
    // copies run on their own stream so that the H2D of chunk k+1 overlaps the kernels of chunk k;
    // a chunk that continues a fragment first runs up to the next fragment boundary
    while (done < n_frames) {
        int64_t idx = first_index + done;
        int64_t n = std::min<int64_t>(chunk, n_frames - done);
        if (idx % F) n = std::min<int64_t>(n, F - idx % F);
        const float* src = h_xyz + (size_t)done * 3 * ctx->T.n_atoms;
        if (!pinned) {
            CK(cudaEventSynchronize(ctx->stage_done[buf]));    // pinned bounce buffer free again
            memcpy(ctx->pin_stage[buf], src, n * fbytes);
            src = ctx->pin_stage[buf];
        }
        CK(cudaStreamWaitEvent(ctx->copy_stream, ctx->stage_done[buf], 0));   // device buffer no longer read
        CK(cudaMemcpyAsync(ctx->dev_stage[buf], src, n * fbytes, cudaMemcpyHostToDevice, ctx->copy_stream));
        CK(cudaEventRecord(ctx->copied[buf], ctx->copy_stream));
        CK(cudaStreamWaitEvent(stream, ctx->copied[buf], 0));
        int rc = chc_submit_frames(ctx, state, ctx->dev_stage[buf], n, idx, 0, stream);
        if (rc != CHC_OK) return rc;
        CK(cudaEventRecord(ctx->stage_done[buf], stream));
        done += n; buf ^= 1;
    }
    return CHC_OK;
}

extern "C" int chc_synchronize(chc_ctx* ctx, void* stream) {
    if (!ctx) return CHC_E_INVALID;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize((cudaStream_t)stream));
    unsigned st = 0;
    CK(cudaMemcpy(&st, ctx->status, 4, cudaMemcpyDeviceToHost));
    if (st & ST_WATER_DEGREE)
        return fail(ctx, CHC_E_TOO_LARGE, "a solvent atom is hydrogen bonded to more than %d non-solvent atoms in one frame; water bridges would be incomplete", WADJ);
    return CHC_OK;
}

// ---- record buffers -----------------------------------------------------------------------------------
template <typename Rec>
static int drain(chc_ctx* ctx, Rec* dev, unsigned long long* dev_n, int64_t dev_cap, Rec* host, int64_t cap, int64_t* n, int64_t* total, cudaStream_t s) {
    unsigned long long have = 0;
    CK(cudaMemcpyAsync(&have, dev_n, 8, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    int64_t stored = std::min<int64_t>((int64_t)have, dev_cap);
    int64_t take = std::min<int64_t>(stored, cap);
    if (take > 0 && host) CK(cudaMemcpyAsync(host, dev, (size_t)take * sizeof(Rec), cudaMemcpyDeviceToHost, s));
    CK(cudaMemsetAsync(dev_n, 0, 8, s));
    CK(cudaStreamSynchronize(s));
    if (n) *n = take;
    if (total) *total = (int64_t)have;
    return CHC_OK;
}

extern "C" int chc_drain_contacts(chc_ctx* ctx, chc_contact* host_buf, int64_t cap, int64_t* n, void* stream) {
    if (!ctx || cap < 0) return CHC_E_INVALID;
    CK(cudaSetDevice(ctx->device));
    int64_t total = 0;
    int rc = drain(ctx, ctx->contacts, ctx->contact_n, ctx->prm.contact_capacity, host_buf, cap, n, &total, (cudaStream_t)stream);
    if (rc != CHC_OK) return rc;
    ctx->st.contacts_emitted += total;
    ctx->st.contacts_dropped += std::max<int64_t>(0, total - std::min<int64_t>(cap, ctx->prm.contact_capacity));
    return CHC_OK;
}

extern "C" int chc_drain_near_cutoff(chc_ctx* ctx, chc_near* host_buf, int64_t cap, int64_t* n, int64_t* n_total, void* stream) {
    if (!ctx || cap < 0) return CHC_E_INVALID;
    CK(cudaSetDevice(ctx->device));
    int64_t total = 0;
    int rc = drain(ctx, ctx->near, ctx->near_n, ctx->prm.near_capacity, host_buf, cap, n, &total, (cudaStream_t)stream);
    if (rc != CHC_OK) return rc;
    ctx->st.near_total += total;
    if (n_total) *n_total = total;
    return CHC_OK;
}

extern "C" int chc_set_timing(chc_ctx* ctx, int on) {
    if (!ctx) return CHC_E_INVALID;
    ctx->timing = on != 0;
    if (on) { ctx->st.pair_ms_total = 0.0; ctx->st.bin_ms_total = 0.0; ctx->st.timed_chunks = 0; }
    return CHC_OK;
}

// ---- contact-list TSV (A13), host threads ------------------------------------------------------------------------
extern "C" int64_t chc_write_contacts_tsv(int fd, chc_contact* rec, int64_t n, const char* label_text, const int64_t* label_off,
                                          const int32_t* label_rank, int32_t n_atoms, int32_t ranks_unique, const char* itype_text,
                                          const int32_t* itype_off, const int32_t* block_rank, int32_t n_itypes, int64_t beg,
                                          int64_t stride, int32_t distout, int32_t n_threads) {
    static_assert(sizeof(tsvw::Rec) == sizeof(chc_contact), "record layout");
    tsvw::Tables T{label_text, label_off, label_rank, n_atoms, ranks_unique != 0, itype_text, itype_off, block_rank, n_itypes,
                   (long long)beg, (long long)stride, distout != 0};
    const int64_t r = tsvw::write_contacts(fd, reinterpret_cast<tsvw::Rec*>(rec), n, T, n_threads);
    return r == -1 ? CHC_E_INVALID : (r == -2 ? CHC_E_IO : (r == -3 ? CHC_E_NOMEM : r));
}

extern "C" int chc_format_fixed3(float value, char* out32) {
    char* e = tsvw::put_fixed3(out32, value);
    *e = 0;
    return (int)(e - out32);
}

// ---- contact-list TSV -> residue-pair counts (F1-F3), host threads ---------------------------------------------------
struct chc_tsv_counts { tsvc::Result R; };

extern "C" int chc_count_tsv(const char* path, const char* itypes_csv, int32_t shipped, int32_t n_threads, chc_tsv_counts** out) {
    if (!path || !out) return CHC_E_INVALID;
    std::vector<std::string> keep;
    if (itypes_csv && *itypes_csv) {
        const char* p = itypes_csv;
        while (true) {
            const char* c = strchr(p, ',');
            keep.emplace_back(p, c ? (size_t)(c - p) : strlen(p));
            if (!c) break;
            p = c + 1;
        }
    }
    chc_tsv_counts* r = new (std::nothrow) chc_tsv_counts();
    if (!r) return CHC_E_NOMEM;
    int rc;
    try { rc = tsvc::count_file(path, keep, shipped != 0, n_threads, r->R); }
    catch (const std::bad_alloc&) { delete r; return CHC_E_NOMEM; }
    *out = r;
    return rc == tsvc::OK ? CHC_OK : rc == tsvc::E_OPEN ? CHC_E_IO : rc == tsvc::E_PARSE ? CHC_E_INVALID : CHC_E_ORDER;
}

extern "C" void chc_tsv_counts_sizes(const chc_tsv_counts* r, int64_t* n_residues, int64_t* n_pairs, int64_t* name_bytes, int64_t* total_frames) {
    int64_t bytes = 0;
    for (const std::string& s : r->R.residues) bytes += (int64_t)s.size();
    if (n_residues) *n_residues = (int64_t)r->R.residues.size();
    if (n_pairs) *n_pairs = (int64_t)r->R.count.size();
    if (name_bytes) *name_bytes = bytes;
    if (total_frames) *total_frames = r->R.total_frames;
}

extern "C" void chc_tsv_counts_get(const chc_tsv_counts* r, char* name_text, int64_t* name_off, int32_t* res1, int32_t* res2, int64_t* count) {
    int64_t off = 0;
    for (size_t k = 0; k < r->R.residues.size(); ++k) {
        name_off[k] = off;
        memcpy(name_text + off, r->R.residues[k].data(), r->R.residues[k].size());
        off += (int64_t)r->R.residues[k].size();
    }
    name_off[r->R.residues.size()] = off;
    const size_t n = r->R.count.size();
    if (n) {
        memcpy(res1, r->R.r1.data(), n * sizeof(int32_t));
        memcpy(res2, r->R.r2.data(), n * sizeof(int32_t));
        memcpy(count, r->R.count.data(), n * sizeof(int64_t));
    }
}

extern "C" const char* chc_tsv_counts_error(const chc_tsv_counts* r) { return r ? r->R.error.c_str() : ""; }
extern "C" void chc_tsv_counts_free(chc_tsv_counts* r) { delete r; }

// ---- DCD frames (X1) ----------------------------------------------------------------------------------------------
extern "C" int chc_dcd_unpack(const void* d_raw, int64_t first, int64_t pitch, int64_t n_frames, int32_t n_atoms, int big_endian,
                              float* d_xyz, int64_t frame_stride, int device, void* stream_) {
    if (n_frames == 0) return CHC_OK;
    if (!d_raw || !d_xyz || n_frames < 0 || n_frames > 0x7fffffff || n_atoms < 1 || first < 0 || pitch < 0) return CHC_E_INVALID;
    if ((((uintptr_t)d_raw + (uintptr_t)first) & 3) != 0 || (pitch & 3) != 0) return CHC_E_INVALID;
    if (cudaSetDevice(device) != cudaSuccess) return CHC_E_CUDA;
    if (frame_stride == 0) frame_stride = 3 * (int64_t)n_atoms;
    const int ntiles = (n_atoms + DCD_TILE - 1) / DCD_TILE;
    const dim3 grid((unsigned)std::min(ntiles, 64), (unsigned)std::min<int64_t>(n_frames, 65535));
    dcd_unpack_kernel<<<grid, DCD_TILE, 0, (cudaStream_t)stream_>>>(static_cast<const unsigned char*>(d_raw), first, pitch, (int)n_frames,
                                                                     n_atoms, big_endian ? 1 : 0, d_xyz, frame_stride);
    return cudaGetLastError() == cudaSuccess ? CHC_OK : CHC_E_CUDA;
}

// ---- ContactPCA permutation test (X4) -------------------------------------------------------------------------
extern "C" int chc_permuted_pca(const double* d_vals, int32_t n, int32_t p, int32_t n_perm, const uint8_t* d_perms_in,
                                uint8_t* d_perms_out, uint64_t seed, double* d_work, double* d_evr, int device, void* stream_) {
    if (n_perm == 0) return CHC_OK;
    if (!d_vals || !d_work || !d_evr || n < 1 || n > PCA_MAXN || p < 1 || n_perm < 0) return CHC_E_INVALID;
    cudaStream_t stream = (cudaStream_t)stream_;
    if (cudaSetDevice(device) != cudaSuccess) return CHC_E_CUDA;
    pca_center_kernel<<<(p + 255) / 256, 256, 0, stream>>>(d_vals, n, p, d_work);
    const int npairs = n * (n + 1) / 2;
    const int k = (npairs + PCA_THREADS - 1) / PCA_THREADS;
    PcaSmem S;
    S.tile_cols = n <= 32 ? 256 : 128;
    size_t off = 0;
    S.off_v = (int)off;     off += al16(8 * (size_t)S.tile_cols * n);
    S.off_g = (int)off;     off += al16(8 * (size_t)n * n);
    S.off_perm = (int)off;  off += al16((size_t)S.tile_cols * n);
    S.off_pairs = (int)off; off += al16(2 * (size_t)npairs);
    S.total = (int)off;
    const void* fn = k <= 1 ? (const void*)pca_perm_kernel<1> : k <= 3 ? (const void*)pca_perm_kernel<3> : (const void*)pca_perm_kernel<9>;
    if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, S.total) != cudaSuccess) return CHC_E_CUDA;
    int sms = 148, per_sm = 1;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, PCA_THREADS, S.total);
    const int grid = std::min(n_perm, sms * std::max(per_sm, 1));
    const unsigned char* pin = d_perms_in; unsigned char* pout = d_perms_out; unsigned long long sd = seed;
    const double* xct = d_work;
    void* args[] = {&xct, &n, &p, &n_perm, &pin, &pout, &sd, &S, &d_evr};
    if (cudaLaunchKernel(fn, dim3(grid), dim3(PCA_THREADS), args, S.total, stream) != cudaSuccess) return CHC_E_CUDA;
    return cudaGetLastError() == cudaSuccess ? CHC_OK : CHC_E_CUDA;
}

// ---- the same test for callers that must not allocate (no counterpart in the reference).
// The centred table `xct` (the only scratch of the path) lives in a file-scope device array instead of a caller
// buffer: CHC_PCA_STATIC_MAX_ELEMS float64 = 4 MiB held by every device that loads the library.  In host code the array's name
// is only a stand-in; its device address is looked up once per device (with that device current) and handed to the kernels.
// pca_center_kernel rewrites every element the permutation kernel reads, so nothing has to start at zero; all calls share the
// array, so calls that overlap on the GPU (two streams, two threads) race.
static __device__ __align__(256) double g_pca_static_xct[CHC_PCA_STATIC_MAX_ELEMS];
static double* g_pca_static_ptr[64];
static std::mutex g_pca_static_mu;

extern "C" int chc_permuted_pca_static(const double* d_vals, int32_t n, int32_t p, int32_t n_perm, const uint8_t* d_perms_in,
                                       uint8_t* d_perms_out, uint64_t seed, double* d_evr, int device, void* stream_) {
    if (n_perm == 0) return CHC_OK;
    if (!d_vals || !d_evr || n < 1 || n > PCA_MAXN || p < 1 || n_perm < 0 || device < 0 || device >= 64) return CHC_E_INVALID;
    if ((int64_t)n * (int64_t)p > (int64_t)CHC_PCA_STATIC_MAX_ELEMS) {
        fprintf(stderr, "chc_permuted_pca_static: n_samples*n_features = %lld exceeds CHC_PCA_STATIC_MAX_ELEMS = %lld; "
                        "use chc_permuted_pca with a caller-owned work buffer\n", (long long)n * p, (long long)CHC_PCA_STATIC_MAX_ELEMS);
        return CHC_E_TOO_LARGE;
    }
    if (cudaSetDevice(device) != cudaSuccess) return CHC_E_CUDA;
    double* work = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_pca_static_mu);
        if (!g_pca_static_ptr[device]) {
            void* addr = nullptr;
            const cudaError_t e = cudaGetSymbolAddress(&addr, g_pca_static_xct);
            if (e != cudaSuccess || !addr) {
                fprintf(stderr, "chc_permuted_pca_static: cudaGetSymbolAddress failed on device %d: %s\n", device, cudaGetErrorString(e));
                return CHC_E_CUDA;
            }
            g_pca_static_ptr[device] = static_cast<double*>(addr);
        }
        work = g_pca_static_ptr[device];
    }
    return chc_permuted_pca(d_vals, n, p, n_perm, d_perms_in, d_perms_out, seed, work, d_evr, device, stream_);
}

extern "C" int chc_get_stats(chc_ctx* ctx, chc_stats* out) {
    if (!ctx || !out) return CHC_E_INVALID;
    if (!ctx->ev_pending.empty()) {
        CK(cudaSetDevice(ctx->device));
        for (auto& e : ctx->ev_pending) {
            CK(cudaEventSynchronize(e.c));
            float ms = 0.f;
            CK(cudaEventElapsedTime(&ms, e.a, e.b)); ctx->st.bin_ms_total += ms;
            CK(cudaEventElapsedTime(&ms, e.b, e.c)); ctx->st.pair_ms_total += ms;
            ctx->st.timed_chunks++;
            ctx->ev_free.push_back(e);
        }
        ctx->ev_pending.clear();
    }
    *out = ctx->st;
    return CHC_OK;
}

// ---- XTC decode ---------------------------------------------------------------------------------------
static_assert(sizeof(XtcFrameMeta) == sizeof(chc_xtc_frame) && sizeof(chc_xtc_frame) == 64, "chc_xtc_frame layout");

// bytes of workspace chc_xtc_decode_ws wants to take n_frames in one batch (any smaller multiple of a frame's share works)
static inline int64_t xtc_work_per_frame(int32_t n_atoms) { return 8 * (int64_t)n_atoms + 8; }

extern "C" int chc_xtc_decode_ws(const void* d_blob, const chc_xtc_frame* d_meta, int64_t n_frames, int32_t n_atoms, float* d_xyz,
                                 int64_t frame_stride, void* d_work, int64_t work_bytes, int device, void* stream_) {
    if (n_frames == 0) return CHC_OK;
    if (!d_blob || !d_meta || !d_xyz || !d_work || n_frames < 0 || n_frames > 0x7fffffff || n_atoms < 1) return CHC_E_INVALID;
    if (((uintptr_t)d_blob & 3) != 0 || ((uintptr_t)d_work & 15) != 0) return CHC_E_INVALID;
    cudaStream_t stream = (cudaStream_t)stream_;
    if (cudaSetDevice(device) != cudaSuccess) return CHC_E_CUDA;
    if (frame_stride == 0) frame_stride = 3 * (int64_t)n_atoms;
    // workspace: [status word, padding to 16 bytes][group counts, batch ints][group records, batch x n_atoms x 8 bytes]
    const int64_t batch = std::min<int64_t>(std::min<int64_t>(n_frames, 65535), (work_bytes - 64) / xtc_work_per_frame(n_atoms));
    if (batch < 1) return CHC_E_INVALID;
    unsigned* status = static_cast<unsigned*>(d_work);
    int* gcount = reinterpret_cast<int*>(static_cast<unsigned char*>(d_work) + 16);
    unsigned long long* groups = reinterpret_cast<unsigned long long*>(static_cast<unsigned char*>(d_work) + 16 + ((4 * batch + 15) & ~15ll));
    cudaMemsetAsync(status, 0, 4, stream);
    if (getenv("CHACRA_XTC_ONEPASS")) {          // test hook: the first decoder (one thread walks and decodes a whole frame)
        const char* tenv = getenv("CHACRA_XTC_THREADS");
        const int threads = tenv ? atoi(tenv) : 32;    // one frame per thread and the chain is serial: spread the warps over the SMs
        const int blocks = (int)std::min<int64_t>((n_frames + threads - 1) / threads, 148 * 64);
        xtc_decode_kernel<<<blocks, threads, 0, stream>>>(static_cast<const unsigned*>(d_blob), reinterpret_cast<const XtcFrameMeta*>(d_meta),
                                                        (int)n_frames, n_atoms, d_xyz, frame_stride, status);
    } else {
        // two passes (xtc_decode.cuh): group boundaries by one thread per frame, then one thread per group.  The boundary walk is
        // a serial chain (~1.1 ms for a 6 500-atom frame) however many frames a launch holds, so the frames go in as few batches
        // as the workspace for the group records allows (8 bytes, at most one per atom).
        const int gx = std::max(1, std::min((n_atoms + 127) / 128, 64));
        for (int64_t f0 = 0; f0 < n_frames; f0 += batch) {
            const int nb = (int)std::min<int64_t>(batch, n_frames - f0);
            xtc_index_kernel<<<(nb + 31) / 32, 32, 0, stream>>>(static_cast<const unsigned*>(d_blob), reinterpret_cast<const XtcFrameMeta*>(d_meta),
                                                               (int)f0, nb, n_atoms, d_xyz, frame_stride, groups, gcount, status);
            xtc_groups_kernel<<<dim3(gx, nb), 128, 0, stream>>>(static_cast<const unsigned*>(d_blob), reinterpret_cast<const XtcFrameMeta*>(d_meta),
                                                                (int)f0, n_atoms, d_xyz, frame_stride, groups, gcount);
        }
    }
    unsigned st = 0;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(&st, status, 4, cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) return CHC_E_CUDA;
    return st ? CHC_E_INVALID : CHC_OK;
}

extern "C" int64_t chc_xtc_work_bytes(int64_t n_frames, int32_t n_atoms) {
    if (n_frames < 1 || n_atoms < 1) return 64;
    return 64 + std::min<int64_t>(n_frames, 65535) * xtc_work_per_frame(n_atoms) + 16;
}

extern "C" int chc_xtc_decode(const void* d_blob, const chc_xtc_frame* d_meta, int64_t n_frames, int32_t n_atoms, float* d_xyz,
                              int64_t frame_stride, int device, void* stream_) {
    if (n_frames == 0) return CHC_OK;
    if (n_frames < 0 || n_atoms < 1) return CHC_E_INVALID;
    if (cudaSetDevice(device) != cudaSuccess) return CHC_E_CUDA;
    // the workspace for all frames at once, at most 2 GB, halved until it can be had
    int64_t bytes = std::min<int64_t>(chc_xtc_work_bytes(n_frames, n_atoms), 2048ll << 20);
    const int64_t least = 64 + xtc_work_per_frame(n_atoms) + 16;
    void* work = nullptr;
    for (;;) {
        bytes = std::max(bytes, least);
        if (cudaMalloc(&work, (size_t)bytes) == cudaSuccess) break;
        cudaGetLastError();
        if (bytes == least) return CHC_E_NOMEM;
        bytes /= 2;
    }
    const int rc = chc_xtc_decode_ws(d_blob, d_meta, n_frames, n_atoms, d_xyz, frame_stride, work, bytes, device, stream_);
    cudaFree(work);
    return rc;
}
