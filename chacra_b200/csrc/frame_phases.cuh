// frame_phases.cuh -- ring phase (pc / ps / ts) and the two dense survivor stages E2 / E3.
#pragma once
#include "frame_kernel.cuh"

// Per-CTA view of the current frame.  The sorted atoms live as slot pairs (x0 x1 y0 y1 | z0 z1 lkey0 lkey1, locality key as an
// exact float) plus one flag word per slot: in shared memory on the small path, in a per-CTA global scratch (L2) on the large
// one.  TAB = the by-pair-index tables (residue rank, static bits) are in shared memory too (small path); otherwise they are
// read from the global topology tables.
template <bool TAB>
struct FrameCtxT {
    const float4* exy; const float4* ezk; const unsigned* sw;   // x0 x1 y0 y1 | z0 z1 lkey0 lkey1 (floats) | flags, pair index
    const unsigned short* pres; const unsigned char* pst;       // TAB, by pair index: residue rank; static bits (I_LIG .. I_PBOND) >> 14
    __device__ __forceinline__ float4 at(int j) const {
        const float4 a = exy[j >> 1]; const float4 b = ezk[j >> 1];
        const float w = __uint_as_float(sw[j]);
        return (j & 1) ? make_float4(a.y, a.w, b.y, w) : make_float4(a.x, a.z, b.x, w);
    }
    __device__ __forceinline__ int lk(int j, unsigned) const {
        const float4 b = ezk[j >> 1];
        return (int)((j & 1) ? b.w : b.z);              // exact: locality keys are below 2^24 (checked by chc_create)
    }
    __device__ __forceinline__ int res(const DevTopo& T, int pidx) const { return TAB ? (int)pres[pidx] : T.pair_res[pidx]; }
    __device__ __forceinline__ unsigned sinfo(const DevTopo& T, int pidx) const { return TAB ? ((unsigned)pst[pidx] << 14) : T.pair_info[pidx]; }
    const float* aux;        // [NA][3] donor hydrogens, ring atoms, cations
    unsigned* bits;          // residue-pair set of this frame
    unsigned gen;            // 0: plain bitset; 1..255: generation of the stamped bitset
    const float* tv; const float* th;
    const unsigned char* hm; // hydrophobic membership of this frame's fragment, by pair index (bit0 H1, bit1 H2)
    int frame;               // frame id reported in records
    WaterScratch W;          // water-bridge scratch of this CTA (null pointers without solvent)
    int* nmark;              // shared counter: entries of W.list
};

// ---- R: pi-cation ----------------------------------------------------------------------------------
template <bool TN, class FrameCtx>
__device__ __forceinline__ void ring_pc_task(const DevTopo& T, const DevParams& P, const DevOut& O, const FrameCtx& F,
                                             int c, int r) {
    const int* pa = T.ring_pc_aux + 3 * r;
    if (pa[0] < 0 || T.cat_excl[c] == r) return;     // bonded pairs never pass measure contacts
    const float* xc = F.aux + 3 * T.cat_aux[c];
    const float* p1 = F.aux + 3 * pa[0]; const float* p2 = F.aux + 3 * pa[1]; const float* p3 = F.aux + 3 * pa[2];
    // conservative float prefilter on the centroid distance
    float cx = (p1[0] + p2[0] + p3[0]) * (1.0f / 3.0f), cy = (p1[1] + p2[1] + p3[1]) * (1.0f / 3.0f),
          cz = (p1[2] + p2[2] + p3[2]) * (1.0f / 3.0f);
    float dx = xc[0] - cx, dy = xc[1] - cy, dz = xc[2] - cz;
    if (dx * dx + dy * dy + dz * dz > P.ring_pre2) return;
    // measure contacts 10.0: every ring atom must be selected against the cation, within the soft cutoff, unbonded
    unsigned cf = T.cat_flags[c];
    int catom = T.cat_atom[c];
    bool soft_fail = false;              // some ring atom lies just above the soft cutoff (inside the near band)
    int near_atom[3]; double near_dl[3]; int n_near = 0;
    for (int k = 0; k < 3; ++k) {
        // ring_pc_aux is a permutation of ring_sel_aux; find the matching selection flags / atom index
        int slot = pa[k], q = 0;
        while (q < 2 && T.ring_sel_aux[3 * r + q] != slot) ++q;
        unsigned rf = T.ring_sel_flags[3 * r + q];
        int ratom = T.ring_atoms[3 * r + q];
        bool dir = ((cf & 1u) && (rf & 2u)) || (!T.same && (cf & 2u) && (rf & 1u));
        if (!dir || ratom == catom) return;
        const float* pr = F.aux + 3 * slot;
        const float d2k = dist2_exact(xc[0], xc[1], xc[2], pr[0], pr[1], pr[2]);
        if (TN && P.track_near && P.track_near) {
            const double dl = (double)sqrtf(d2k) - P.pc_soft;
            if (fabs(dl) < P.near_d) { near_atom[n_near] = ratom; near_dl[n_near] = dl; ++n_near; }
            if (d2k > P.pc_soft2) { if (!(dl < P.near_d)) return; soft_fail = true; }
        } else if (d2k > P.pc_soft2) return;
    }
    D3 a = d3(p1), b = d3(p2), cc = d3(p3), x = d3(xc);
    D3 cen = centroid3(a, b, cc);
    D3 v = sub3(x, cen);
    double d = sqrt(dot3(v, v));
    int rep = T.ring_rep[r];
    if (n_near > 0 && d - P.pc_dist < P.near_d)       // the soft 10 A search decides this group (pi_cation.py:22,70,98)
        for (int k = 0; k < n_near; ++k) emit_near(O, F.frame, CHC_NEAR_PC_SOFT, catom, near_atom[k], near_dl[k]);
    if (soft_fail) return;
    if (TN && P.track_near && fabs(d - P.pc_dist) < P.near_d) emit_near(O, F.frame, CHC_NEAR_PC_DIST, catom, rep, d - P.pc_dist);
    if (d > P.pc_dist) return;
    double t = fold180(angle_deg(normal3(a, b, cc), v));
    if (TN && P.track_near && fabs(t - P.pc_angle) < P.near_a) emit_near(O, F.frame, CHC_NEAR_PC_ANGLE, catom, rep, t - P.pc_angle);
    if (t > P.pc_angle) return;
    count_pair(O, F.bits, F.gen, T.cat_res[c], T.ring_res[r], T.R);
    if (TN && P.emit) {
        const float* pr = F.aux + 3 * T.ring_rep_aux[r];
        emit_contact(O, F.frame, CHC_PC, catom, rep, sqrtf(dist2_exact(xc[0], xc[1], xc[2], pr[0], pr[1], pr[2])));
    }
}

// ---- R: pi-stacking and T-stacking ---------------------------------------------------------------------
template <bool TN, class FrameCtx>
__device__ __forceinline__ void ring_stack_task(const DevTopo& T, const DevParams& P, const DevOut& O, const FrameCtx& F,
                                                int r1, int r2) {
    if (T.ring_res[r1] == T.ring_res[r2]) return;
    const int* t1 = T.ring_aux + 3 * r1; const int* t2 = T.ring_aux + 3 * r2;
    // soft-cutoff stage on the selected ring atoms: which of the two searches finds this residue pair?
    bool hit_ps = false, hit_ts = false;
    float best = 3.0e38f;
    for (int i = 0; i < 3; ++i) {
        int si = T.ring_sel_aux[3 * r1 + i];
        if (si < 0) continue;
        unsigned fi = T.ring_sel_flags[3 * r1 + i];
        const float* pi = F.aux + 3 * si;
        for (int j = 0; j < 3; ++j) {
            int sj = T.ring_sel_aux[3 * r2 + j];
            if (sj < 0) continue;
            unsigned fj = T.ring_sel_flags[3 * r2 + j];
            if (!(((fi & 1u) && (fj & 2u)) || ((fi & 2u) && (fj & 1u)))) continue;
            const float* pj = F.aux + 3 * sj;
            best = fminf(best, dist2_exact(pi[0], pi[1], pi[2], pj[0], pj[1], pj[2]));
        }
    }
    hit_ps = (P.itypes & CHC_IT_PS) && best <= P.ps_soft2;
    hit_ts = (P.itypes & CHC_IT_TS) && best <= P.ts_soft2;
    // the closest selected ring-atom pair within the near band of a soft search cutoff (aromatics.py:26-27,166)
    double soft_ps = 1.0, soft_ts = 1.0;
    if (TN && P.track_near && best < 3.0e38f) {
        const double db = (double)sqrtf(best);
        if (P.itypes & CHC_IT_PS) soft_ps = db - P.ps_soft;
        if (P.itypes & CHC_IT_TS) soft_ts = db - P.ts_soft;
    }
    const bool near_ps = fabs(soft_ps) < P.near_d, near_ts = fabs(soft_ts) < P.near_d;
    if (!(hit_ps || hit_ts || near_ps || near_ts) || t1[0] < 0 || t2[0] < 0) return;
    D3 a1 = d3(F.aux + 3 * t1[0]), a2 = d3(F.aux + 3 * t1[1]), a3 = d3(F.aux + 3 * t1[2]);
    D3 b1 = d3(F.aux + 3 * t2[0]), b2 = d3(F.aux + 3 * t2[1]), b3 = d3(F.aux + 3 * t2[2]);
    D3 c1 = centroid3(a1, a2, a3), c2 = centroid3(b1, b2, b3);
    D3 cc = sub3(c1, c2);
    double d = sqrt(dot3(cc, cc));
    int rep1 = T.ring_rep[r1], rep2 = T.ring_rep[r2];
    D3 n1 = normal3(a1, a2, a3), n2 = normal3(b1, b2, b3);
    double ang = 0.0, psi = 0.0;
    bool have_ang = false;
    for (int which = 0; which < 2; ++which) {
        bool on = which == 0 ? hit_ps : hit_ts;
        double dcut = which == 0 ? P.ps_dist : P.ts_dist, acut = which == 0 ? P.ps_angle : P.ts_angle,
               pcut = which == 0 ? P.ps_psi : P.ts_psi;
        if ((which == 0 ? near_ps : near_ts) && d - dcut < P.near_d)
            emit_near(O, F.frame, which == 0 ? CHC_NEAR_PS_SOFT : CHC_NEAR_TS_SOFT, rep1, rep2, which == 0 ? soft_ps : soft_ts);
        if (!on) continue;
        if (TN && P.track_near && fabs(d - dcut) < P.near_d)
            emit_near(O, F.frame, which == 0 ? CHC_NEAR_PS_DIST : CHC_NEAR_TS_DIST, rep1, rep2, d - dcut);
        if (d > dcut) continue;
        if (!have_ang) {
            ang = angle_deg(n1, n2);
            psi = fmin(fold180(angle_deg(n1, cc)), fold180(angle_deg(n2, D3{-cc.x, -cc.y, -cc.z})));
            have_ang = true;
        }
        double v = which == 0 ? fold180(ang) : fabs(ang - 90.0);
        if (TN && P.track_near && fabs(v - acut) < P.near_a)
            emit_near(O, F.frame, which == 0 ? CHC_NEAR_PS_ANGLE : CHC_NEAR_TS_ANGLE, rep1, rep2, v - acut);
        if (v > acut) continue;
        if (TN && P.track_near && fabs(psi - pcut) < P.near_a)
            emit_near(O, F.frame, which == 0 ? CHC_NEAR_PS_PSI : CHC_NEAR_TS_PSI, rep1, rep2, psi - pcut);
        if (psi > pcut) continue;
        count_pair(O, F.bits, F.gen, T.ring_res[r1], T.ring_res[r2], T.R);
        if (TN && P.emit) {
            const float* q1 = F.aux + 3 * T.ring_rep_aux[r1]; const float* q2 = F.aux + 3 * T.ring_rep_aux[r2];
            emit_contact(O, F.frame, which == 0 ? CHC_PS : CHC_TS, rep1, rep2,
                         sqrtf(dist2_exact(q1[0], q1[1], q1[2], q2[0], q2[1], q2[2])));
        }
    }
}

// ---- E2: exact float32 d2 and class tests for one prefilter survivor ---------------------------------------
template <bool TN, class FrameCtx>
__device__ __forceinline__ unsigned e2_classify(const DevTopo& T, const DevParams& P, const FrameCtx& F, int si, int sj,
                                                float& d2_out) {
    const float4 a = F.at(si), b = F.at(sj);
    float d2 = dist2_exact(a.x, a.y, a.z, b.x, b.y, b.z);
    d2_out = d2;
    unsigned wi = __float_as_uint(a.w), wj = __float_as_uint(b.w);
    bool dual = ((wi & I_S1) && (wj & I_S2)) || ((wi & I_S2) && (wj & I_S1));
    // water-bridge edge: exactly one solvent atom, the other in the hydrogen-bond selection U; needs no dual selection
    // (hbonds.py:78-80) and is dropped only inside one residue (:131)
    const bool edge_pre = (P.itypes & CHC_IT_HB) && ((wi ^ wj) & I_SOLV) && (((wi & I_SOLV) ? wj : wi) & I_HBU) &&
                          ((wi | wj) & I_HASH);
    const bool edge = edge_pre && d2 <= P.hb2;
    // near-cutoff bands of the two plain distance cutoffs (salt_bridges.py:76-80, hbonds.py:64-82)
    bool band_sb = false, band_hb = false;
    if (TN && P.track_near && P.track_near) {
        const double dd = (double)sqrtf(d2);
        band_sb = fabs(dd - P.sb_dist) < P.near_d; band_hb = fabs(dd - P.hb_dist) < P.near_d;
    }
    const bool edge_near = edge_pre && band_hb;
    if (!dual && !edge && !edge_near) return 0u;
    int dl = abs(F.lk(si, wi) - F.lk(sj, wj));
    if (!dual) return dl != 0 ? ((edge ? C_HBW : 0u) | (edge_near ? C_NEARB : 0u)) : 0u;
    int e = (wi & I_ELEM) * T.K + (wj & I_ELEM);
    unsigned cls = 0;
    if (dl >= P.vdw_rd) {
        if ((P.itypes & CHC_IT_VDW) && (wi & wj & I_VDW)) {
            if (d2 <= F.tv[e]) cls |= C_VDW;
            if (TN && P.track_near && d2 <= P.soft2 && fabs((double)sqrtf(d2) - T.vcut[e]) < P.near_d) cls |= C_NEARV;
        }
        if ((P.itypes & CHC_IT_HP) && d2 <= P.hp2) {
            // static base-carbon bits first, then this fragment's membership (hydrophobics.py:8-41)
            bool base = T.same ? ((wi & wj & I_HP1) != 0)
                               : (((wi & I_HP1) && (wj & I_HP2)) || ((wj & I_HP1) && (wi & I_HP2)));
            if (base) {
                unsigned hi = F.hm[wi >> I_PIDX_SHIFT], hj = F.hm[wj >> I_PIDX_SHIFT];
                bool hp = T.same ? ((hi & hj & 1u) != 0)
                                 : (((wi & I_HP1) && (hi & 1u) && (wj & I_HP2) && (hj & 2u)) ||
                                    ((wj & I_HP1) && (hj & 1u) && (wi & I_HP2) && (hi & 2u)));
                if (hp) {
                    if (d2 <= F.th[e]) cls |= C_HP;
                    if (TN && P.track_near && fabs((double)sqrtf(d2) - T.vcut[e]) < P.near_d) cls |= C_NEARH;
                }
            }
        }
    }
    if ((P.itypes & CHC_IT_SB) && (((wi & I_ANION) && (wj & I_CATION)) || ((wj & I_ANION) && (wi & I_CATION)))) {
        if (d2 <= P.sb2) cls |= C_SB;
        if (band_sb) cls |= C_NEARS;
    }
    if ((P.itypes & CHC_IT_HB) && (wi & wj & I_HBU) && ((wi | wj) & I_HASH) && dl >= P.hb_rd) {
        if (d2 <= P.hb2) cls |= C_HB;
        if (band_hb) cls |= C_NEARB;
    }
    if (dl != 0) cls |= (edge ? C_HBW : 0u) | (edge_near ? C_NEARB : 0u);
    return cls;
}

// does donor d see acceptor a under the angle rule, for any of its hydrogens?
// The reference value is 57.2958f * (float)atan2((double)|u x v|, (double)(u.v)) < cutoff.  atan2 is
// monotone in |u x v| / (u.v), so away from the cutoff the decision follows from one multiply: P.hb_tan_lo / hb_tan_hi
// bracket tan(cutoff) by a relative 4e-4 (~7e-3 degree at 70 degrees, far above every rounding step of the exact recipe
// and above the 1e-3 degree near-cutoff band); only inside the bracket is the exact double-precision atan2 evaluated.
template <bool TN, class FrameCtx>
__device__ __forceinline__ bool hb_direction(const DevTopo& T, const DevParams& P, const DevOut& O, const FrameCtx& F,
                                             float4 d, float4 a, int pd, int atom_d, int atom_a, bool report = true) {
    bool ok = false;
    int hb = T.pair_h_start[pd], he = T.pair_h_start[pd + 1];
    for (int h = hb; h < he; ++h) {
        const float* ph = F.aux + 3 * h;
        const float hx = ph[0], hy = ph[1], hz = ph[2];
        float ux = hx - d.x, uy = hy - d.y, uz = hz - d.z;
        float vx = a.x - hx, vy = a.y - hy, vz = a.z - hz;
        float cx = uy * vz - uz * vy, cy = uz * vx - ux * vz, cz = ux * vy - uy * vx;
        float psin = sqrtf((cx * cx + cy * cy) + cz * cz);
        float pcos = (ux * vx + uy * vy) + uz * vz;
        if (P.hb_fast) {
            if (pcos > 0.f && psin < pcos * P.hb_tan_lo) { ok = true; continue; }      // surely below the cutoff
            if (pcos <= 0.f || psin > pcos * P.hb_tan_hi) continue;                    // surely above it
        }
        float ang = 57.2958f * (float)atan2((double)psin, (double)pcos);
        if (report && TN && P.track_near && fabs((double)ang - (double)P.hb_angle) < P.near_a)
            emit_near(O, F.frame, CHC_NEAR_HB_ANGLE, atom_d, atom_a, (double)ang - (double)P.hb_angle);
        ok |= ang < P.hb_angle;
    }
    return ok;
}

// ---- E3: everything that needs the global tables, for one classified pair ---------------------------------------
template <bool TN, class FrameCtx>
__device__ __forceinline__ void e3_finish(const DevTopo& T, const DevParams& P, const DevOut& O, const FrameCtx& F,
                                          int i, int j, unsigned cls) {
    float4 a = F.at(i), b = F.at(j);
    const float d2 = dist2_exact(a.x, a.y, a.z, b.x, b.y, b.z);      // as in e2_classify (same operands, same order)
    unsigned wi = __float_as_uint(a.w), wj = __float_as_uint(b.w);
    int pi = wi >> I_PIDX_SHIFT, pj = wj >> I_PIDX_SHIFT;
    const unsigned si = F.sinfo(T, pi), sj = F.sinfo(T, pj);         // static-only bits: ligand, backbone, CYS, xbond, pbond
    // directly bonded pairs are excluded by measure contacts / measure hbonds (Appendix A.4, A.5); the static flags say whether
    // the atom has ANY bond that could matter (I_XBOND: one the residue-distance filters do not exclude; I_PBOND: one to a polar atom)
    const bool need_bond = ((cls & (C_VDW | C_HP | C_NEARV | C_NEARH)) && (si & sj & I_XBOND)) ||
                           ((cls & (C_SB | C_HB | C_HBW | C_NEARS | C_NEARB)) && (si & sj & I_PBOND));
    if (need_bond && is_bonded(T, pi, T.pair_atom[pj])) return;
    int ai = 0, aj = 0;                                              // atom indices are only needed for records
    if (TN) { ai = T.pair_atom[pi]; aj = T.pair_atom[pj]; }
    int ri = F.res(T, pi), rj = F.res(T, pj);
    float dist = sqrtf(d2);
    if (cls & C_NEARS) emit_near(O, F.frame, CHC_NEAR_SB_DIST, ai, aj, (double)dist - P.sb_dist);
    if (cls & C_NEARB) emit_near(O, F.frame, CHC_NEAR_HB_DIST, ai, aj, (double)dist - P.hb_dist);
    if (cls & (C_VDW | C_HP | C_NEARV | C_NEARH)) {
        bool ss = (si & I_CYS) && (sj & I_CYS) && is_disulfide(T, T.pair_resid[pi], T.pair_resid[pj]);
        if (!ss) {
            int e = (wi & I_ELEM) * T.K + (wj & I_ELEM);
            if (cls & C_NEARV) emit_near(O, F.frame, CHC_NEAR_VDW_DIST, ai, aj, (double)dist - T.vcut[e]);
            if (cls & C_NEARH) emit_near(O, F.frame, CHC_NEAR_HP_DIST, ai, aj, (double)dist - T.vcut[e]);
            if (cls & C_VDW) {
                count_pair(O, F.bits, F.gen, ri, rj, T.R);
                if (TN && P.emit) emit_contact(O, F.frame, CHC_VDW, ai, aj, dist);
            }
            if (TN && (cls & C_HP) && P.emit) {     // hp is not a counted type (get_contact_frequencies.py:107-109)
                bool li = si & I_LIG, lj = sj & I_LIG;
                emit_contact(O, F.frame, (li || lj) ? ((li != lj) ? CHC_HPLP : CHC_HPLL) : CHC_HP, ai, aj, dist);
            }
        }
    }
    if (cls & C_SB) {
        // orient as (anion, cation); salt_bridges.py:90-97 appends a plain sb line unless both are ligand atoms
        bool ij = (wi & I_ANION) && (wj & I_CATION);
        unsigned wa = ij ? si : sj, wc = ij ? sj : si;
        int aa = ij ? ai : aj, ac = ij ? aj : ai;
        bool al = wa & I_LIG, cl = wc & I_LIG;
        if (!(al && cl)) count_pair(O, F.bits, F.gen, ri, rj, T.R);
        if (TN && P.emit) {
            emit_contact(O, F.frame, al ? (cl ? CHC_SBLL : CHC_SBLP) : (cl ? CHC_SBPL : CHC_SB), aa, ac, dist);
            // the reference's if / if / if-else also appends a plain `sb` line behind an `sblp` or `sbpl` one (salt_bridges.py:90-97):
            // it is the line get_contact_frequencies counts
            if (al != cl) emit_contact(O, F.frame, CHC_SB, aa, ac, dist);
        }
    }
    if (cls & (C_HB | C_HBW)) {
        bool ok = false;
#pragma unroll 1
        for (int dir = 0; dir < 2; ++dir) {            // both atoms are tried as donor (measure hbonds, both argument orders)
            const unsigned wd = dir ? wj : wi;
            if (!(wd & I_HASH)) continue;
            const float4 xd = dir ? b : a, xa = dir ? a : b;
            ok |= hb_direction<TN>(T, P, O, F, xd, xa, dir ? pj : pi, dir ? aj : ai, dir ? ai : aj, (cls & C_HB) != 0);
        }
        if (ok && (cls & C_HBW)) {
            // record the edge with the solvent atom: neighbour = sorted slot of the non-solvent atom (hbonds.py:126-142)
            const bool wi_solv = (wi & I_SOLV) != 0;
            const int ord = T.pair_solv_ord[wi_solv ? pi : pj];
            const int n = atomicAdd(&F.W.cnt[ord], 1);
            if (n < WADJ) F.W.adj[ord * WADJ + n] = wi_solv ? j : i;
            else atomicOr(O.status, ST_WATER_DEGREE);
            if (n == 0) { F.W.slot[ord] = wi_solv ? i : j; F.W.list[atomicAdd(F.nmark, 1)] = ord; }
        }
        if (ok && (cls & C_HB)) {
            bool li = si & I_LIG, lj = sj & I_LIG, bi = si & I_BB, bj = sj & I_BB;
            int t;
            if (li && lj) t = CHC_HBLL;
            else if (li) t = bj ? CHC_HBLB : CHC_HBLS;
            else if (lj) t = bi ? CHC_HBLB : CHC_HBLS;
            else if (bi && bj) t = CHC_HBBB;
            else if (!bi && !bj) t = CHC_HBSS;
            else t = CHC_HBSB;
            if (t != CHC_HBLL) count_pair(O, F.bits, F.gen, ri, rj, T.R);
            if (TN && P.emit) emit_contact(O, F.frame, t, ai, aj, dist);
        }
    }
}
