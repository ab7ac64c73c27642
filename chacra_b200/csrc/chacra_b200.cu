// chacra_b200.cu -- C ABI of libchacra_b200.so (see include/chacra_b200.h).  sm_100a only.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdarg>
#include <cmath>
#include <vector>
#include <algorithm>
#include "frame_main.cuh"
#include "prepass.cuh"

#define THREADS_SMALL 512
#define SMEM_LIMIT (227 * 1024)

struct chc_ctx {
    int device = 0;
    int sm_count = 148;
    chc_params prm{};
    DevTopo T{};
    DevParams P{};
    SmemLayout L{};
    std::vector<void*> dev_allocs;
    int32_t* counts = nullptr; bool counts_owned = false; int64_t n_pairs = 0;
    chc_contact* contacts = nullptr; unsigned long long* contact_n = nullptr;
    chc_near* near = nullptr; unsigned long long* near_n = nullptr;
    unsigned char* hpmask = nullptr; int64_t hpmask_frags = 0;
    unsigned* gbits = nullptr;
    float* dev_stage[2] = {nullptr, nullptr}; float* pin_stage[2] = {nullptr, nullptr};
    int64_t stage_frames = 0; cudaEvent_t stage_done[2] = {nullptr, nullptr}; cudaEvent_t copied[2] = {nullptr, nullptr}; cudaStream_t copy_stream = nullptr;
    bool timing = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev_pending, ev_free;   // CUDA-event pairs around main kernels
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

static int plan_small(chc_ctx* ctx) {
    const DevTopo& T = ctx->T;
    SmemLayout& L = ctx->L;
    const int warps = THREADS_SMALL / 32;
    auto al = [](size_t v) { return (v + 15) & ~size_t(15); };
    size_t off = 0;
    L.off_spos = (int)off; off = al(off + 16 * (size_t)T.n_pair);
    L.off_slk = (int)off;  off = al(off + 4 * (size_t)T.n_pair);
    L.off_tmp = (int)off;  off = al(off + 4 * (size_t)T.n_pair);
    L.off_aux = (int)off;  off = al(off + 12 * (size_t)std::max(T.n_aux, 1));
    L.off_q1 = (int)off;   off = al(off + 2 * (size_t)warps * LQ_CAP * 32);
    L.off_q2 = (int)off;   off = al(off + 12 * (size_t)warps * Q2_CAP);
    L.off_tv = (int)off;   off = al(off + 4 * (size_t)T.K * T.K);
    L.off_th = (int)off;   off = al(off + 4 * (size_t)T.K * T.K);
    L.off_misc = (int)off; off = al(off + sizeof(FrameMisc));
    if (T.n_pair > MAX_PAIR_SMALL || off + 4 * 1024 > SMEM_LIMIT) return CHC_E_TOO_LARGE;
    L.bits_words = (int)((ctx->n_pairs + 31) / 32);
    size_t bits_bytes = al(4 * (size_t)L.bits_words);
    size_t left = SMEM_LIMIT - off;
    // the cell array wants ~2 cells per pair atom (plus guard cells) to keep cells at the cutoff size
    size_t want_cells = std::min<size_t>(16384, std::max<size_t>(2048, 3 * (size_t)T.n_pair));
    L.bits_in_smem = (bits_bytes + 4 * std::min<size_t>(want_cells, 4096) <= left) ? 1 : 0;
    L.off_bits = (int)off;
    if (L.bits_in_smem) { off += bits_bytes; left -= bits_bytes; }
    L.off_cells = (int)off;
    size_t cells = std::min<size_t>(want_cells, left / 4);
    if (cells < 1024) return CHC_E_TOO_LARGE;
    L.max_cells = (int)cells;
    off = al(off + 4 * cells);
    L.total = (int)off;
    return CHC_OK;
}

extern "C" int chc_abi_version(void) { return CHC_ABI_VERSION; }
extern "C" const char* chc_last_error(const chc_ctx* ctx) { return ctx ? ctx->err : "null context"; }
extern "C" int64_t chc_n_pairs(const chc_ctx* ctx) { return ctx ? ctx->n_pairs : 0; }

extern "C" void chc_destroy(chc_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    for (void* p : ctx->dev_allocs) cudaFree(p);
    if (ctx->counts_owned) cudaFree(ctx->counts);
    cudaFree(ctx->contacts); cudaFree(ctx->contact_n); cudaFree(ctx->near); cudaFree(ctx->near_n);
    cudaFree(ctx->hpmask); cudaFree(ctx->gbits);
    for (int k = 0; k < 2; ++k) {
        cudaFree(ctx->dev_stage[k]); if (ctx->pin_stage[k]) cudaFreeHost(ctx->pin_stage[k]);
        if (ctx->stage_done[k]) cudaEventDestroy(ctx->stage_done[k]);
        if (ctx->copied[k]) cudaEventDestroy(ctx->copied[k]);
    }
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    for (auto& e : ctx->ev_pending) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
    for (auto& e : ctx->ev_free) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
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
    int rc;
    const size_t NP = t->n_pair, NR = t->n_rings, NC = t->n_cations, K2 = (size_t)t->n_elem * t->n_elem;
#define UP(field, src, n) if ((rc = upload(ctx, src, (size_t)(n), &T.field)) != CHC_OK) return rc
    UP(pair_atom, t->pair_atom, NP); UP(pair_info, t->pair_info, NP); UP(pair_lkey, t->pair_lkey, NP);
    UP(pair_res, t->pair_res, NP); UP(pair_resid, t->pair_resid, NP); UP(pair_h_start, t->pair_h_start, NP + 1);
    UP(pair_bond_start, t->pair_bond_start, NP + 1); UP(pair_bond_adj, t->pair_bond_adj, t->n_bond_adj);
    UP(aux_atom, t->aux_atom, t->n_aux);
    UP(ring_aux, t->ring_aux, 3 * NR); UP(ring_pc_aux, t->ring_pc_aux, 3 * NR); UP(ring_sel_aux, t->ring_sel_aux, 3 * NR);
    UP(ring_sel_flags, t->ring_sel_flags, 3 * NR); UP(ring_atoms, t->ring_atoms, 3 * NR); UP(ring_res, t->ring_res, NR);
    UP(ring_rep, t->ring_rep, NR); UP(ring_rep_aux, t->ring_rep_aux, NR);
    UP(cat_aux, t->cation_aux, NC); UP(cat_atom, t->cation_atom, NC); UP(cat_flags, t->cation_flags, NC); UP(cat_res, t->cation_res, NC); UP(cat_excl, t->cation_excl_ring, NC);
    UP(disulfide, t->disulfide, 2 * (size_t)t->n_disulfide);
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
    P.pre2 = (float)(cmax * cmax * (1.0 + 1e-5) + 1e-5);
    P.hb_angle = (float)p->hb_angle;
    double rmax = std::max(p->pc_dist, std::max(p->ps_dist, p->ts_dist)) + 0.05;
    P.ring_pre2 = (float)(rmax * rmax);
    P.pc_soft2 = sq(10.0); P.ps_soft2 = sq(10.0); P.ts_soft2 = sq(6.0);   // pi_cation.py:22, aromatics.py:26-27
    P.pc_dist = p->pc_dist; P.pc_angle = p->pc_angle; P.ps_dist = p->ps_dist; P.ps_angle = p->ps_angle; P.ps_psi = p->ps_psi;
    P.ts_dist = p->ts_dist; P.ts_angle = p->ts_angle; P.ts_psi = p->ts_psi;
    P.sb_dist = p->sb_dist; P.hb_dist = p->hb_dist; P.hp_dist = p->hp_dist; P.soft = p->vdw_soft_cutoff;
    P.near_d = p->near_dist_tol; P.near_a = p->near_angle_tol;

    ctx->n_pairs = (int64_t)T.R * (T.R + 1) / 2;
    rc = plan_small(ctx);
    if (rc != CHC_OK) return fail(ctx, rc, "topology with %d pair atoms / %d residues does not fit the one-CTA-per-frame path", T.n_pair, T.R);
    CK(cudaFuncSetAttribute(frame_contacts_kernel<THREADS_SMALL>, cudaFuncAttributeMaxDynamicSharedMemorySize, ctx->L.total));
    if (!ctx->L.bits_in_smem) {
        CK(cudaMalloc(&ctx->gbits, (size_t)ctx->sm_count * ctx->L.bits_words * 4));
    }
    CK(cudaMalloc(&ctx->counts, (size_t)p->n_states * ctx->n_pairs * sizeof(int32_t)));
    ctx->counts_owned = true;
    CK(cudaMemset(ctx->counts, 0, (size_t)p->n_states * ctx->n_pairs * sizeof(int32_t)));
    CK(cudaMalloc(&ctx->contact_n, 8)); CK(cudaMemset(ctx->contact_n, 0, 8));
    CK(cudaMalloc(&ctx->near_n, 8)); CK(cudaMemset(ctx->near_n, 0, 8));
    CK(cudaMalloc(&ctx->contacts, std::max<int64_t>(p->contact_capacity, 1) * sizeof(chc_contact)));
    CK(cudaMalloc(&ctx->near, std::max<int64_t>(p->near_capacity, 1) * sizeof(chc_near)));
    ctx->st.path = 1; ctx->st.smem_bytes = ctx->L.total; ctx->st.threads = THREADS_SMALL; ctx->st.staged = 0;
    ctx->st.bitset_in_smem = ctx->L.bits_in_smem;
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

extern "C" int chc_submit_frames(chc_ctx* ctx, int state, const float* d_xyz, int64_t n_frames, int64_t first_index,
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
    int rc = ensure_hpmask(ctx, frag1 - frag0 + 1);
    if (rc != CHC_OK) return rc;
    if (continues && ctx->hp_last_slot > 0) {
        // the fragment being continued was computed into slot hp_last_slot by the previous call
        CK(cudaMemcpyAsync(ctx->hpmask, ctx->hpmask + (size_t)ctx->hp_last_slot * ctx->T.n_pair, ctx->T.n_pair, cudaMemcpyDeviceToDevice, stream));
    }
    const int first_g = continues ? 1 : 0;
    const int n_new = (int)(frag1 - frag0 + 1) - first_g;
    if ((ctx->P.itypes & CHC_IT_HP) && ctx->T.n_hp_base > 0 && n_new > 0) {
        const float r = 1.7f;
        int by = std::max(1, std::min(64, (ctx->T.n_hp_base + 7) / 8));
        hp_prepass_kernel<<<dim3(n_new, by), 256, 0, stream>>>(ctx->T, r * r, d_xyz, frame_stride, first_index, F, frag0, first_g, ctx->hpmask);
        CK(cudaGetLastError());
        ctx->st.kernel_launches++;
    }
    ctx->hp_last_slot = (int)(frag1 - frag0);
    DevOut O;
    O.counts = ctx->counts + (size_t)state * ctx->n_pairs; O.n_pairs = ctx->n_pairs;
    O.contacts = ctx->contacts; O.contact_n = ctx->contact_n; O.contact_cap = ctx->prm.contact_capacity;
    O.near = ctx->near; O.near_n = ctx->near_n; O.near_cap = ctx->prm.near_capacity;
    int grid = (int)std::min<int64_t>(n_frames, ctx->sm_count);
    std::pair<cudaEvent_t, cudaEvent_t> ev{nullptr, nullptr};
    if (ctx->timing) {
        if (!ctx->ev_free.empty()) { ev = ctx->ev_free.back(); ctx->ev_free.pop_back(); }
        else { CK(cudaEventCreate(&ev.first)); CK(cudaEventCreate(&ev.second)); }
        CK(cudaEventRecord(ev.first, stream));
    }
    frame_contacts_kernel<THREADS_SMALL><<<grid, THREADS_SMALL, ctx->L.total, stream>>>(
        ctx->T, ctx->P, O, ctx->L, d_xyz, frame_stride, (int)n_frames, first_index, F, frag0, ctx->hpmask, ctx->gbits);
    CK(cudaGetLastError());
    if (ctx->timing) { CK(cudaEventRecord(ev.second, stream)); ctx->ev_pending.push_back(ev); }
    ctx->st.kernel_launches++;
    ctx->st.frames_submitted += n_frames;
    ctx->st.grid = grid;
    ctx->last_state = state; ctx->next_index = first_index + n_frames;
    return CHC_OK;
}

extern "C" int chc_submit_frames_host(chc_ctx* ctx, int state, const float* h_xyz, int64_t n_frames, int64_t first_index, void* stream_) {
    if (!ctx) return CHC_E_INVALID;
    if (n_frames == 0) return CHC_OK;
    if (!h_xyz || n_frames < 0) return fail(ctx, CHC_E_INVALID, "submit_frames_host: bad arguments");
    cudaStream_t stream = (cudaStream_t)stream_;
    CK(cudaSetDevice(ctx->device));
    const int F = ctx->prm.fragment_frames;
    const size_t fbytes = 12 * (size_t)ctx->T.n_atoms;
    // chunk = whole fragments, about 64 MiB
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
    if (on) { ctx->st.kernel_ms_total = 0.0; ctx->st.timed_launches = 0; }
    return CHC_OK;
}

extern "C" int chc_get_stats(chc_ctx* ctx, chc_stats* out) {
    if (!ctx || !out) return CHC_E_INVALID;
    if (!ctx->ev_pending.empty()) {
        CK(cudaSetDevice(ctx->device));
        for (auto& e : ctx->ev_pending) {
            CK(cudaEventSynchronize(e.second));
            float ms = 0.f;
            CK(cudaEventElapsedTime(&ms, e.first, e.second));
            ctx->st.last_kernel_ms = ms; ctx->st.kernel_ms_total += ms; ctx->st.timed_launches++;
            ctx->ev_free.push_back(e);
        }
        ctx->ev_pending.clear();
    }
    *out = ctx->st;
    return CHC_OK;
}
