// getcontacts_cpu.cpp -- TEST INFRASTRUCTURE / CPU BASELINE, not part of the product.
//
// C++/OpenMP twin of the numpy oracle (oracle/getcontacts_oracle.py), i.e. of the reference's per-frame detectors
//   external/getcontacts/contact_calc/compute_contacts.py:50-103   (order sb, pc, ps, ts, vdw, hb, hp)
//   salt_bridges.py:33-99, pi_cation.py:34-129, aromatics.py:116-257, vanderwaals.py:34-96, hbonds.py:29-124 (no solvent),
//   hydrophobics.py:8-123, and the residue-pair frame counts of transformations.py:262-304,446-470 (intended semantics, L2)
// with the VMD primitives of SURVEY.md Appendix A restated in the same float32 / float64 recipes.  A competent all-core
// CPU implementation (cell list, one frame per OpenMP task) for bench.py's `cpu_baseline` / `--impl reference` arm:
// what a CPU user would run if getcontacts were not bound to VMD's Tcl interpreter.
// Parity: pinned to the numpy oracle (which is pinned to the reference's golden contact lists) by tests/test_cpu_twin.py.
// Not covered: solvent / water bridges (the wrapper refuses topologies with solvent; the numpy oracle covers those).
//
// Static per-atom tables come from the numpy oracle's Topology / Setup objects (oracle/cpu_twin.py).
// Build: make -C oracle   (g++ -O3 -fopenmp -ffp-contract=off: float32 products and sums must round separately)
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

extern "C" {

struct gc_topo {
    int32_t n_atoms, n_reskeys, same, n_dis, n_ring_groups, n_ring_res, reserved0, reserved1;
    const uint8_t* flags;        // [N] bit0 hydrogen (by name), 1 sele1, 2 sele2, 3 anion, 4 cation, 5 protein cation, 6 hbond set U,
                                 //     7 ligand  (second byte array below for the rest)
    const uint8_t* flags2;       // [N] bit0 backbone name, 1 CYS, 2 hydrophobic base carbon, 3 element C, 4 element H
    const double* vdw;           // [N] van der Waals radius by element
    const int32_t* chain;        // [N] chain id
    const int32_t* resid;        // [N]
    const int32_t* reskey;       // [N] rank of the residue label string chain:resname:resid+icode
    const int32_t* bond_start;   // [N+1] CSR of the static bond list
    const int32_t* bond_adj;
    const int32_t* dis;          // [2 n_dis] disulfide resid pairs
    // pi-cation groups: ring atoms grouped by residue label (pi_cation.py:83-91)
    const int32_t* rg_start;     // [n_ring_groups+1]
    const int32_t* rg_atoms;     // ring atoms of the group, in the order of their label:index strings (pi_cation.py:101)
    const int32_t* rg_reskey;    // [n_ring_groups]
    // aromatic residues for ps / ts (aromatics.py:71-113,170-205)
    const int32_t* rr_reskey;    // [n_ring_res]
    const int32_t* rr_trip;      // [3 n_ring_res] the triplet (-1 when the residue does not have exactly three ring atoms)
    const int32_t* rr_rep;       // [n_ring_res] ring atom with the smallest name (contact_utils.py:335-347)
    const int32_t* rr_sel_start; // [n_ring_res+1] selected ring atoms of the residue
    const int32_t* rr_sel;
};

struct gc_params {
    uint32_t itypes;             // bit0 sb, 1 pc, 2 ps, 3 ts, 4 vdw, 5 hb, 6 hp
    int32_t hb_rd, vdw_rd, fragment, threads, reserved;
    double sb_dist, pc_dist, pc_angle, ps_dist, ps_angle, ps_psi, ts_dist, ts_angle, ts_psi, hb_dist, hb_angle, eps, soft;
};

struct gc_contact { int32_t frame, itype, a, b; };

}  // extern "C"

namespace {

enum { F_H = 1, F_S1 = 2, F_S2 = 4, F_AN = 8, F_CAT = 16, F_PCAT = 32, F_U = 64, F_LIG = 128 };
enum { G_BB = 1, G_CYS = 2, G_HPB = 4, G_EC = 8, G_EH = 16 };
// chc_itype numbering of include/chacra_b200.h (shared with chacra_b200.engine.ITYPE_NAMES)
enum { T_SB = 0, T_SBLP, T_SBPL, T_SBLL, T_PC, T_PS, T_TS, T_VDW, T_HBBB, T_HBSB, T_HBSS, T_HBLL, T_HBLB, T_HBLS, T_WB, T_WB2, T_LWB, T_LWB2,
       T_HP, T_HPLP, T_HPLL };

inline float d2f(const float* a, const float* b) {
    const float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    return (dx * dx + dy * dy) + dz * dz;           // -ffp-contract=off: every product and sum rounds to binary32
}

struct D3 { double x, y, z; };
inline D3 d3(const float* p) { return D3{(double)p[0], (double)p[1], (double)p[2]}; }
inline D3 operator-(D3 a, D3 b) { return D3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline double dot(D3 a, D3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline D3 cross(D3 a, D3 b) { return D3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline D3 centroid(D3 a, D3 b, D3 c) { return D3{((a.x + b.x) + c.x) / 3.0, ((a.y + b.y) + c.y) / 3.0, ((a.z + b.z) + c.z) / 3.0}; }
inline double angle_deg(D3 a, D3 b) { return std::acos(dot(a, b) / (std::sqrt(dot(a, a)) * std::sqrt(dot(b, b)))) * (180.0 / 3.14159265358979323846); }
inline double fold(double t) { return std::min(std::fabs(t), std::fabs(t - 180.0)); }

struct Grid {
    float org[3], inv; int nc[3];
    std::vector<int> start, order;      // cell -> first slot; slot -> member index
    std::vector<int> cell_of;
    void build(const float* xyz, const int* members, int m, float edge) {
        float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f};
        for (int k = 0; k < m; ++k) for (int d = 0; d < 3; ++d) { const float v = xyz[3 * members[k] + d]; lo[d] = std::min(lo[d], v); hi[d] = std::max(hi[d], v); }
        if (m == 0) for (int d = 0; d < 3; ++d) lo[d] = hi[d] = 0.f;
        for (;;) {
            inv = 1.0f / edge;
            long long tot = 1;
            for (int d = 0; d < 3; ++d) { org[d] = lo[d]; nc[d] = (int)std::min((hi[d] - lo[d]) * inv, 1.0e6f) + 1; tot *= nc[d]; }
            if (tot <= 4000000) break;
            edge *= 1.5f;
        }
        const int ncell = nc[0] * nc[1] * nc[2];
        start.assign(ncell + 1, 0); cell_of.resize(m); order.resize(m);
        for (int k = 0; k < m; ++k) { cell_of[k] = cell(xyz + 3 * members[k]); start[cell_of[k] + 1]++; }
        for (int c = 0; c < ncell; ++c) start[c + 1] += start[c];
        std::vector<int> cur(start.begin(), start.end() - 1);
        for (int k = 0; k < m; ++k) order[cur[cell_of[k]]++] = k;
    }
    int axis(float v, int d) const { int c = (int)((v - org[d]) * inv); return std::min(std::max(c, 0), nc[d] - 1); }
    int cell(const float* p) const { return (axis(p[2], 2) * nc[1] + axis(p[1], 1)) * nc[0] + axis(p[0], 0); }
};

struct Ctx {
    const gc_topo& T; const gc_params& P;
    std::vector<int> pair_atoms;                 // heavy atoms of sele1 | sele2
    std::vector<int> all_atoms;
    std::vector<int> cations, pcations;
    float sb2, hb2, soft2, hp2, cmax;
    Ctx(const gc_topo& t, const gc_params& p) : T(t), P(p) {
        for (int i = 0; i < T.n_atoms; ++i) {
            all_atoms.push_back(i);
            const uint8_t f = T.flags[i];
            if (!(f & F_H) && (f & (F_S1 | F_S2))) pair_atoms.push_back(i);
            if (f & F_PCAT) pcations.push_back(i);
        }
        auto sq = [](double c) { const float f = (float)c; return f * f; };
        sb2 = sq(P.sb_dist); hb2 = sq(P.hb_dist); soft2 = sq(P.soft); hp2 = sq(P.eps + 2 * 1.7);
        double c = 0.5;
        if (P.itypes & 1) c = std::max(c, P.sb_dist);
        if (P.itypes & 16) c = std::max(c, P.soft);
        if (P.itypes & 32) c = std::max(c, P.hb_dist);
        if (P.itypes & 64) c = std::max(c, P.eps + 3.4);
        cmax = (float)(c * 1.0005 + 1e-3);
    }
    bool bonded(int a, int b) const {
        for (int k = T.bond_start[a]; k < T.bond_start[a + 1]; ++k) if (T.bond_adj[k] == b) return true;
        return false;
    }
    bool dual(int a, int b) const {
        const uint8_t fa = T.flags[a], fb = T.flags[b];
        return ((fa & F_S1) && (fb & F_S2)) || ((fa & F_S2) && (fb & F_S1));
    }
    bool local(int a, int b, int k) const { return T.chain[a] == T.chain[b] && std::abs(T.resid[a] - T.resid[b]) < k; }
    bool disulfide(int a, int b) const {
        if (!((T.flags2[a] & G_CYS) && (T.flags2[b] & G_CYS))) return false;
        const int ra = T.resid[a], rb = T.resid[b];
        for (int k = 0; k < T.n_dis; ++k)
            if ((T.dis[2 * k] == ra && T.dis[2 * k + 1] == rb) || (T.dis[2 * k] == rb && T.dis[2 * k + 1] == ra)) return true;
        return false;
    }
};

struct FrameOut {
    std::vector<gc_contact> rec;           // this frame's contacts (itype, a < b)
    void add(int frame, int itype, int a, int b) { rec.push_back(gc_contact{frame, itype, std::min(a, b), std::max(a, b)}); }
};

// hydrophobics.py:8-41 on the fragment's first frame: 1 = member of H1 (subset rule), 2 = member of H2 (equality rule)
void prepare_hp(const Ctx& C, const float* x0, std::vector<uint8_t>& hp, bool& any1, bool& any2) {
    const gc_topo& T = C.T;
    hp.assign(T.n_atoms, 0);
    Grid G;
    G.build(x0, C.all_atoms.data(), T.n_atoms, 1.75f);
    const float r2 = 1.7f * 1.7f;
    any1 = any2 = false;
    for (int c = 0; c < T.n_atoms; ++c) {
        if (!(T.flags2[c] & G_HPB) || !(T.flags[c] & (F_S1 | F_S2))) continue;
        bool other = false, has_h = false;
        const float* p = x0 + 3 * c;
        const int cx = G.axis(p[0], 0), cy = G.axis(p[1], 1), cz = G.axis(p[2], 2);
        for (int z = std::max(cz - 1, 0); z <= std::min(cz + 1, G.nc[2] - 1); ++z)
            for (int y = std::max(cy - 1, 0); y <= std::min(cy + 1, G.nc[1] - 1); ++y)
                for (int xx = std::max(cx - 1, 0); xx <= std::min(cx + 1, G.nc[0] - 1); ++xx) {
                    const int cell = (z * G.nc[1] + y) * G.nc[0] + xx;
                    for (int s = G.start[cell]; s < G.start[cell + 1]; ++s) {
                        const int j = G.order[s];
                        if (d2f(x0 + 3 * j, p) <= r2) {
                            if (T.flags2[j] & G_EH) has_h = true;
                            else if (!(T.flags2[j] & G_EC)) other = true;
                        }
                    }
                }
        if (other) continue;
        if (T.flags[c] & F_S1) { hp[c] |= 1; any1 = true; }
        if ((T.flags[c] & F_S2) && has_h) { hp[c] |= 2; any2 = true; }
    }
}

void frame_contacts(const Ctx& C, const float* x, int frame, const std::vector<uint8_t>& hp, bool hp_on, Grid& G, FrameOut& out) {
    const gc_topo& T = C.T; const gc_params& P = C.P;
    out.rec.clear();
    const bool same = T.same != 0;
    const float hb_ang = (float)P.hb_angle;
    // ---- rings: pi-cation (pi_cation.py:34-129) ----------------------------------------------------------------------------
    if (P.itypes & 2) {
        const float soft = 10.0f * 10.0f;
        for (int c : C.pcations) {
            const uint8_t fc = T.flags[c];
            for (int g = 0; g < T.n_ring_groups; ++g) {
                int got[4], n = 0;
                for (int k = T.rg_start[g]; k < T.rg_start[g + 1] && n < 4; ++k) {
                    const int r = T.rg_atoms[k];
                    const uint8_t fr = T.flags[r];
                    const bool dir = ((fc & F_S1) && (fr & F_S2)) || (!same && (fc & F_S2) && (fr & F_S1));
                    if (!dir || r == c || d2f(x + 3 * c, x + 3 * r) > soft || C.bonded(c, r)) continue;
                    got[n++] = r;
                }
                if (n != 3) continue;
                const D3 p1 = d3(x + 3 * got[0]), p2 = d3(x + 3 * got[1]), p3 = d3(x + 3 * got[2]), xc = d3(x + 3 * c);
                const D3 cen = centroid(p1, p2, p3), v = xc - cen;
                if (std::sqrt(dot(v, v)) > P.pc_dist) continue;
                if (fold(angle_deg(cross(p3 - p1, p2 - p1), v)) > P.pc_angle) continue;
                int rep = -1;                                  // smallest atom NAME of the triplet: precomputed per aromatic residue
                for (int q = 0; q < T.n_ring_res; ++q) if (T.rr_reskey[q] == T.rg_reskey[g]) { rep = T.rr_rep[q]; break; }
                out.add(frame, T_PC, c, rep);
            }
        }
    }
    // ---- rings: pi-stacking / T-stacking (aromatics.py:116-257) ------------------------------------------------------------------
    if (P.itypes & (4 | 8)) {
        for (int r1 = 0; r1 < T.n_ring_res; ++r1)
            for (int r2 = r1 + 1; r2 < T.n_ring_res; ++r2) {
                if (T.rr_reskey[r1] == T.rr_reskey[r2]) continue;
                float best = 3.0e38f;
                for (int a = T.rr_sel_start[r1]; a < T.rr_sel_start[r1 + 1]; ++a)
                    for (int b = T.rr_sel_start[r2]; b < T.rr_sel_start[r2 + 1]; ++b) {
                        const int i = T.rr_sel[a], j = T.rr_sel[b];
                        if (!C.dual(i, j) || C.bonded(i, j)) continue;
                        best = std::min(best, d2f(x + 3 * i, x + 3 * j));
                    }
                const bool hit_ps = (P.itypes & 4) && best <= 100.0f, hit_ts = (P.itypes & 8) && best <= 36.0f;
                if (!(hit_ps || hit_ts) || T.rr_trip[3 * r1] < 0 || T.rr_trip[3 * r2] < 0) continue;
                const D3 a1 = d3(x + 3 * T.rr_trip[3 * r1]), a2 = d3(x + 3 * T.rr_trip[3 * r1 + 1]), a3 = d3(x + 3 * T.rr_trip[3 * r1 + 2]);
                const D3 b1 = d3(x + 3 * T.rr_trip[3 * r2]), b2 = d3(x + 3 * T.rr_trip[3 * r2 + 1]), b3 = d3(x + 3 * T.rr_trip[3 * r2 + 2]);
                const D3 c1 = centroid(a1, a2, a3), c2 = centroid(b1, b2, b3), cc = c1 - c2;
                const double d = std::sqrt(dot(cc, cc));
                const D3 n1 = cross(a3 - a1, a2 - a1), n2 = cross(b3 - b1, b2 - b1);
                for (int which = 0; which < 2; ++which) {
                    if (!(which ? hit_ts : hit_ps)) continue;
                    if (d > (which ? P.ts_dist : P.ps_dist)) continue;
                    const double ang = angle_deg(n1, n2);
                    if ((which ? std::fabs(ang - 90.0) : fold(ang)) > (which ? P.ts_angle : P.ps_angle)) continue;
                    const double psi = std::min(fold(angle_deg(n1, cc)), fold(angle_deg(n2, D3{-cc.x, -cc.y, -cc.z})));
                    if (psi > (which ? P.ts_psi : P.ps_psi)) continue;
                    out.add(frame, which ? T_TS : T_PS, T.rr_rep[r1], T.rr_rep[r2]);
                }
            }
    }
    // ---- pair interactions over a cell list of the selected heavy atoms --------------------------------------------------------
    if (!(P.itypes & (1 | 16 | 32 | 64))) return;
    const int m = (int)C.pair_atoms.size();
    G.build(x, C.pair_atoms.data(), m, C.cmax);
    const float pre2 = std::max(std::max((P.itypes & 1) ? C.sb2 : 0.f, (P.itypes & 16) ? C.soft2 : 0.f),
                                std::max((P.itypes & 32) ? C.hb2 : 0.f, (P.itypes & 64) ? C.hp2 : 0.f));
    auto pair_test = [&](int i, int j) {
        const float d2 = d2f(x + 3 * i, x + 3 * j);
        if (d2 > pre2) return;
        const uint8_t fi = T.flags[i], fj = T.flags[j];
        if (!C.dual(i, j)) return;                 // every pair type needs one atom in each selection
        bool know_bond = false, is_b = false;
        auto bonded = [&]() { if (!know_bond) { is_b = C.bonded(i, j); know_bond = true; } return is_b; };
        // salt bridges (salt_bridges.py:76-97)
        if ((P.itypes & 1) && d2 <= C.sb2) {
            int a = -1, c = -1;
            if ((fi & F_AN) && (fj & F_CAT)) { a = i; c = j; } else if ((fj & F_AN) && (fi & F_CAT)) { a = j; c = i; }
            if (a >= 0) {
                const uint8_t fa = T.flags[a], fc = T.flags[c];
                const bool dir = ((fa & F_S1) && (fc & F_S2)) || (!same && (fa & F_S2) && (fc & F_S1));
                if (dir && !bonded()) {
                    const bool al = fa & F_LIG, cl = fc & F_LIG;
                    if (al && !cl) out.add(frame, T_SBLP, a, c);
                    if (!al && cl) out.add(frame, T_SBPL, a, c);
                    if (al && cl) out.add(frame, T_SBLL, a, c); else out.add(frame, T_SB, a, c);
                }
            }
        }
        // van der Waals and hydrophobics (vanderwaals.py:69-94, hydrophobics.py:83-121)
        const bool want_vdw = (P.itypes & 16) && d2 <= C.soft2;
        bool want_hp = false;
        if (hp_on && d2 <= C.hp2) {
            const uint8_t hi = hp[i], hj = hp[j];
            want_hp = same ? ((hi & hj & 1) != 0) : (((hi & 1) && (hj & 2)) || ((hj & 1) && (hi & 2)));
        }
        if ((want_vdw || want_hp) && !C.local(i, j, P.vdw_rd) && !bonded() && !C.disulfide(i, j)) {
            const double dist = (double)std::sqrt(d2);              // float32 sqrt, widened (measure bond)
            if (dist < (T.vdw[i] + T.vdw[j]) + P.eps) {
                if (want_vdw) out.add(frame, T_VDW, i, j);
                if (want_hp) {
                    const bool li = fi & F_LIG, lj = fj & F_LIG;
                    out.add(frame, (li || lj) ? ((li != lj) ? T_HPLP : T_HPLL) : T_HP, i, j);
                }
            }
        }
        // hydrogen bonds (hbonds.py:75-121; measure hbonds)
        if ((P.itypes & 32) && d2 <= C.hb2 && (fi & fj & F_U) && !C.local(i, j, P.hb_rd) && !bonded()) {
            bool ok = false;
            for (int dir = 0; dir < 2 && !ok; ++dir) {
                const int d = dir ? j : i, a = dir ? i : j;
                for (int k = T.bond_start[d]; k < T.bond_start[d + 1] && !ok; ++k) {
                    const int h = T.bond_adj[k];
                    if (!(T.flags[h] & F_H)) continue;
                    const float* xd = x + 3 * d; const float* xh = x + 3 * h; const float* xa = x + 3 * a;
                    const float ux = xh[0] - xd[0], uy = xh[1] - xd[1], uz = xh[2] - xd[2];
                    const float vx = xa[0] - xh[0], vy = xa[1] - xh[1], vz = xa[2] - xh[2];
                    const float cx = uy * vz - uz * vy, cy = uz * vx - ux * vz, cz = ux * vy - uy * vx;
                    const float psin = std::sqrt((cx * cx + cy * cy) + cz * cz);
                    const float pcos = (ux * vx + uy * vy) + uz * vz;
                    const float ang = 57.2958f * (float)std::atan2((double)psin, (double)pcos);
                    ok = ang < hb_ang;
                }
            }
            if (ok) {
                const bool li = fi & F_LIG, lj = fj & F_LIG, bi = T.flags2[i] & G_BB, bj = T.flags2[j] & G_BB;
                int t;
                if (li && lj) t = T_HBLL; else if (li) t = bj ? T_HBLB : T_HBLS; else if (lj) t = bi ? T_HBLB : T_HBLS;
                else if (bi && bj) t = T_HBBB; else if (!bi && !bj) t = T_HBSS; else t = T_HBSB;
                out.add(frame, t, i, j);
            }
        }
    };
    const int nx = G.nc[0], ny = G.nc[1], nz = G.nc[2];
    for (int z = 0; z < nz; ++z)
        for (int y = 0; y < ny; ++y)
            for (int xx = 0; xx < nx; ++xx) {
                const int c = (z * ny + y) * nx + xx;
                const int cb = G.start[c], ce = G.start[c + 1];
                if (cb == ce) continue;
                for (int s = cb; s < ce; ++s)                                  // inside the cell
                    for (int t = s + 1; t < ce; ++t) pair_test(C.pair_atoms[G.order[s]], C.pair_atoms[G.order[t]]);
                for (int dz = 0; dz <= 1; ++dz)                                // half shell of neighbour cells
                    for (int dy = (dz ? -1 : 0); dy <= 1; ++dy)
                        for (int dx = ((dz || dy) ? -1 : 1); dx <= 1; ++dx) {
                            const int X = xx + dx, Y = y + dy, Z = z + dz;
                            if (X < 0 || X >= nx || Y < 0 || Y >= ny || Z >= nz) continue;
                            const int o = (Z * ny + Y) * nx + X;
                            for (int s = cb; s < ce; ++s)
                                for (int t = G.start[o]; t < G.start[o + 1]; ++t) pair_test(C.pair_atoms[G.order[s]], C.pair_atoms[G.order[t]]);
                        }
            }
}

inline bool counted(int itype) {           // get_contact_frequencies.py:107-109: no hp*, no sblp / sbpl / sbll, no hbll
    return itype == T_SB || itype == T_PC || itype == T_PS || itype == T_TS || itype == T_VDW || itype == T_HBBB || itype == T_HBSB ||
           itype == T_HBSS || itype == T_HBLB || itype == T_HBLS || itype == T_WB || itype == T_WB2 || itype == T_LWB || itype == T_LWB2;
}

}  // namespace

extern "C" int gc_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// xyz: float32 [n_frames][N][3].  counts: int32 [R (R + 1) / 2] (index lo R - lo (lo - 1) / 2 + (hi - lo) over residue-label ranks), zeroed
// by the caller, or NULL.  out / cap: contact records (unique per frame), or cap = 0 for counts only.  Returns 0, or -1 when `cap` was too small
// (n_out then holds the number needed).
extern "C" int gc_run(const gc_topo* Tp, const gc_params* Pp, const float* xyz, int64_t n_frames, int32_t* counts, gc_contact* out,
                      int64_t cap, int64_t* n_out) {
    const gc_topo& T = *Tp; const gc_params& P = *Pp;
    Ctx C(T, P);
    const int frag = std::max(1, (int)P.fragment);
    const int64_t n_frag = (n_frames + frag - 1) / frag;
    const size_t fstride = 3 * (size_t)T.n_atoms;
    const bool hp_wanted = (P.itypes & 64) != 0;
    std::vector<std::vector<uint8_t>> hp(n_frag);
    std::vector<char> hp_on(n_frag, 0);
    const int threads = P.threads > 0 ? P.threads : gc_threads();
#ifdef _OPENMP
    // an explicit thread count wins over the environment (torchrun exports OMP_NUM_THREADS=1 to its workers)
    if (P.threads > 0) { omp_set_dynamic(0); omp_set_num_threads(threads); }
#endif
    if (hp_wanted) {
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
        for (int64_t k = 0; k < n_frag; ++k) {
            bool a1, a2;
            prepare_hp(C, xyz + (size_t)k * frag * fstride, hp[k], a1, a2);
            hp_on[k] = a1 && a2;                     // hydrophobics.py: `if not h1 or not h2: return []`
        }
    }
    const int64_t R = T.n_reskeys;
    std::vector<std::vector<gc_contact>> per_thread(threads);
    int64_t total = 0;
#pragma omp parallel num_threads(threads)
    {
#ifdef _OPENMP
        const int me = omp_get_thread_num();
#else
        const int me = 0;
#endif
        Grid G; FrameOut fo;
        std::vector<int64_t> pairs;
        std::vector<uint8_t> none;
#pragma omp for schedule(dynamic, 4)
        for (int64_t f = 0; f < n_frames; ++f) {
            const int64_t k = f / frag;
            frame_contacts(C, xyz + (size_t)f * fstride, (int)f, hp_wanted ? hp[k] : none, hp_wanted && hp_on[k], G, fo);
            // canonical records of one frame are unique per (itype, a, b) (compute_contacts.py:350-359)
            std::sort(fo.rec.begin(), fo.rec.end(), [](const gc_contact& u, const gc_contact& v) {
                return u.itype != v.itype ? u.itype < v.itype : (u.a != v.a ? u.a < v.a : u.b < v.b); });
            fo.rec.erase(std::unique(fo.rec.begin(), fo.rec.end(), [](const gc_contact& u, const gc_contact& v) {
                return u.itype == v.itype && u.a == v.a && u.b == v.b; }), fo.rec.end());
            if (counts) {
                pairs.clear();
                for (const gc_contact& c : fo.rec) {
                    if (!counted(c.itype)) continue;
                    const int64_t ra = T.reskey[c.a], rb = T.reskey[c.b], lo = std::min(ra, rb), hi = std::max(ra, rb);
                    pairs.push_back(lo * R - (lo * (lo - 1)) / 2 + (hi - lo));
                }
                std::sort(pairs.begin(), pairs.end());
                pairs.erase(std::unique(pairs.begin(), pairs.end()), pairs.end());
                for (int64_t p : pairs) {
#pragma omp atomic
                    counts[p] += 1;
                }
            }
            if (cap > 0) per_thread[me].insert(per_thread[me].end(), fo.rec.begin(), fo.rec.end());
        }
    }
    for (auto& v : per_thread) total += (int64_t)v.size();
    if (n_out) *n_out = total;
    if (cap > 0) {
        if (total > cap) return -1;
        int64_t at = 0;
        for (auto& v : per_thread) { if (!v.empty()) memcpy(out + at, v.data(), v.size() * sizeof(gc_contact)); at += (int64_t)v.size(); }
    }
    return 0;
}
