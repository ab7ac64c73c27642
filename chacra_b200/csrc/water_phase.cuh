// water_phase.cuh -- water-mediated hydrogen bonds (hbonds.py:126-204; SURVEY Appendix D "hb").
//
// The pair phase has recorded, for every solvent heavy atom w, the non-solvent atoms of the hydrogen-bond selection
// that are hydrogen bonded to it in this frame (frame_phases.cuh, class C_HBW): N(w), at most WADJ entries, plus the
// list of solvent atoms with a non-empty N ("marked").  A bridge needs a non-solvent partner on every water, and such
// a partner lies within the 3.5 A hydrogen-bond cutoff of the water, so the reference's solvent shell ("solv within
// 3.5 of the selection", hbonds.py:64) never decides anything beyond that: only marked waters can appear in a line.
//   wb / lwb    every unordered pair {a, b} of N(w) passing the dual-selection and residue-distance filters
//   wb2 / lwb2  every hydrogen bond between marked solvent atoms w, w2 (found here with a radius query on the grid,
//               taken once, from the atom with the lower index: hbonds.py:170) times N(w) x N(w2)
// One warp per marked solvent atom.
#pragma once
#include "frame_phases.cuh"
#include "bin_kernel.cuh"

template <class FrameCtx>
__device__ __forceinline__ void bridge_emit(const DevTopo& T, const DevParams& P, const DevOut& O, const FrameCtx& F,
                                            int sa, int sb, int atom_w1, int atom_w2) {
    const float4 xa = F.at(sa), xb = F.at(sb);
    const unsigned wa = __float_as_uint(xa.w), wb = __float_as_uint(xb.w);
    if (!(((wa & I_S1) && (wb & I_S2)) || ((wa & I_S2) && (wb & I_S1)))) return;          // filter_dual_selection
    const int pa = wa >> I_PIDX_SHIFT, pb = wb >> I_PIDX_SHIFT;
    if (abs(F.lk(sa, wa) - F.lk(sb, wb)) < P.hb_rd) return;                                   // same chain, |dresid| < res_diff
    const bool lig = ((T.pair_info[pa] | T.pair_info[pb]) & I_LIG) != 0;
    const int itype = atom_w2 < 0 ? (lig ? CHC_LWB : CHC_WB) : (lig ? CHC_LWB2 : CHC_WB2);
    count_pair(O, F.bits, F.gen, T.pair_res[pa], T.pair_res[pb], T.R);
    if (P.emit)
        emit_bridge(O, F.frame, itype, T.pair_atom[pa], T.pair_atom[pb], atom_w1, atom_w2,
                    sqrtf(dist2_exact(xa.x, xa.y, xa.z, xb.x, xb.y, xb.z)));
}

template <bool TN, class cell_t, class FrameCtx>
__device__ __forceinline__ void water_bridges(const DevTopo& T, const DevParams& P, const DevOut& O, const FrameCtx& F,
                                              const WaterScratch& W, const cell_t* cells, int px, int pxy, int ncell,
                                              const BinHeader* hdr, const int* nmark, int tid, int nthreads) {
    const int lane = tid & 31, nwarps = nthreads >> 5;
    const int nm = *nmark;
    const int ngrid = T.n_pair_a < T.n_pair ? 2 : 1;
    const float ox = hdr->org[0], oy = hdr->org[1], oz = hdr->org[2], inv = hdr->inv;
    const int n0 = hdr->nc[0], n1 = hdr->nc[1], n2 = hdr->nc[2];
    for (int m = tid >> 5; m < nm; m += nwarps) {
        const int ord = W.list[m];
        const int sw = W.slot[ord];
        const int nw = min(W.cnt[ord], WADJ);
        const int* Nw = W.adj + (size_t)ord * WADJ;
        const float4 xw = F.at(sw);
        const unsigned ww = __float_as_uint(xw.w);
        const int pw = ww >> I_PIDX_SHIFT;
        const int aw = T.pair_atom[pw];
        // ---- one water -------------------------------------------------------------------------------
        const int tri = nw * (nw - 1) / 2;
        const int npair = tri + (P.hb_rd == 0 ? nw : 0);        // with res_diff 0 the reference also pairs an atom with itself
        for (int t = lane; t < npair; t += 32) {
            int a = 0, b = 1;
            if (t < tri) { int r = t; while (r >= nw - 1 - a) { r -= nw - 1 - a; ++a; } b = a + 1 + r; }
            else a = b = t - tri;
            bridge_emit(T, P, O, F, Nw[a], Nw[b], aw, -1);
        }
        // ---- two waters ------------------------------------------------------------------------------
        const int c0 = (cell_of(xw.z, oz, inv, n2) + 1) * pxy + (cell_of(xw.y, oy, inv, n1) + 1) * px + cell_of(xw.x, ox, inv, n0) + 1;
        for (int g = 0; g < ngrid; ++g)
            for (int r = 0; r < 9; ++r) {
                const int base = c0 + g * ncell + (r % 3 - 1) * px + (r / 3 - 1) * pxy;
                const int jb = (int)cells[base - 1], je = (int)cells[base + 2];
                for (int j = jb + lane; j < je; j += 32) {
                    const float4 xj = F.at(j);
                    const unsigned wj = __float_as_uint(xj.w);
                    if (!(wj & I_SOLV)) continue;
                    const int pj = wj >> I_PIDX_SHIFT;
                    const int aj = T.pair_atom[pj];
                    if (aj <= aw) continue;                                     // each solvent-solvent bond once (hbonds.py:170)
                    const int ordj = T.pair_solv_ord[pj];
                    const int nj = min(W.cnt[ordj], WADJ);
                    if (nj == 0) continue;                                      // no non-solvent partner: cannot close a bridge
                    if (F.lk(sw, ww) == F.lk(j, wj)) continue;                     // same residue (hbonds.py:131)
                    const float d2w = dist2_exact(xw.x, xw.y, xw.z, xj.x, xj.y, xj.z);
                    const double dlw = (TN && P.track_near) ? (double)sqrtf(d2w) - P.hb_dist : 1.0;
                    if (d2w > P.hb2 && !(fabs(dlw) < P.near_d)) continue;
                    if (!((ww | wj) & I_HASH) || is_bonded(T, pw, aj)) continue;
                    if (fabs(dlw) < P.near_d) emit_near(O, F.frame, CHC_NEAR_HB_DIST, aw, aj, dlw);
                    if (d2w > P.hb2) continue;
                    bool ok = false;
                    if (ww & I_HASH) ok |= hb_direction<TN>(T, P, O, F, xw, xj, pw, aw, aj, false);
                    if (!ok && (wj & I_HASH)) ok |= hb_direction<TN>(T, P, O, F, xj, xw, pj, aj, aw, false);
                    if (!ok) continue;
                    const int* Nj = W.adj + (size_t)ordj * WADJ;
                    for (int a = 0; a < nw; ++a)
                        for (int b = 0; b < nj; ++b) bridge_emit(T, P, O, F, Nw[a], Nj[b], aw, aj);
                }
            }
    }
    __syncthreads();
    for (int m = tid; m < nm; m += nthreads) W.cnt[W.list[m]] = 0;      // all zero again for the next frame
}
