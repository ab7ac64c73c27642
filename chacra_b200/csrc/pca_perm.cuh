// pca_perm.cuh -- the ContactPCA permutation test on the device.
//
// Reference: chacra/ContactFrequencies.py:631-656 (`de_correlate`: every contact column gets its rows reordered
// independently; `run_permuted_pca`: sklearn PCA().fit(...).explained_variance_ratio_ per permutation) driven by
// `ContactPCA.permuted_pca` (:910-964), N_permutations x a full SVD of an [n_states x n_contacts] matrix (20 x ~4 300).
//
// Only the explained-variance RATIOS are used, and those are the eigenvalues of the n x n Gram matrix of the
// column-centred data divided by their sum.  Reordering the rows of a column does not change its mean, so the centred
// data are computed once (`pca_center_kernel`, transposed to [p][n]) and each permutation is
//   G[a][b] = sum_c x[pi_c(a)][c] * x[pi_c(b)][c]      (float64, one CTA per permutation, columns tiled through
//                                                      shared memory, one (a <= b) accumulator per thread)
// followed by a cyclic Jacobi eigenvalue iteration on G by one warp.  Permutations come either from the caller
// (parity tests feed the reference's own index arrays) or from a counter-based generator (Fisher-Yates per column,
// splitmix64 keyed by seed, permutation and column); they can be written out for checking.
#pragma once
#include <cstdint>

#define PCA_MAXN 64
#define PCA_THREADS 256

__device__ __forceinline__ unsigned long long pca_mix(unsigned long long z) {      // splitmix64 finaliser
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

// column means in row order (plain sequential sums in float64), centred values transposed to xct[p][n]
__global__ void pca_center_kernel(const double* __restrict__ vals, int n, int p, double* __restrict__ xct) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= p) return;
    double s = 0.0;
    for (int a = 0; a < n; ++a) s += vals[(size_t)a * p + c];
    const double mean = s / (double)n;
    for (int a = 0; a < n; ++a) xct[(size_t)c * n + a] = vals[(size_t)a * p + c] - mean;
}

struct PcaSmem {                 // byte offsets into dynamic shared memory
    int off_v, off_perm, off_g, off_pairs, total, tile_cols;
};

template <int K>                 // K = accumulators per thread = ceil(n (n + 1) / 2 / PCA_THREADS)
__global__ void __launch_bounds__(PCA_THREADS)
pca_perm_kernel(const double* __restrict__ xct, int n, int p, int n_perm, const unsigned char* __restrict__ perms_in,
                unsigned char* __restrict__ perms_out, unsigned long long seed, PcaSmem S, double* __restrict__ evr) {
    extern __shared__ __align__(16) unsigned char psm[];
    double* V = reinterpret_cast<double*>(psm + S.off_v);                    // [tile_cols][n] permuted centred columns
    unsigned char* perm = psm + S.off_perm;                                  // [tile_cols][n]
    double* G = reinterpret_cast<double*>(psm + S.off_g);                    // [n][n]
    unsigned char* pairs = psm + S.off_pairs;                                // (a, b) of accumulator q
    const int tid = threadIdx.x, lane = tid & 31;
    const int npairs = n * (n + 1) / 2, TC = S.tile_cols;
    for (int q = tid; q < npairs; q += PCA_THREADS) {                        // q -> (a <= b), row-major over the upper triangle
        int a = 0, rem = q;
        while (rem >= n - a) { rem -= n - a; ++a; }
        pairs[2 * q] = (unsigned char)a; pairs[2 * q + 1] = (unsigned char)(a + rem);
    }
    __syncthreads();
    int pa[K], pb[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const int q = tid + k * PCA_THREADS;
        pa[k] = q < npairs ? pairs[2 * q] : 0; pb[k] = q < npairs ? pairs[2 * q + 1] : 0;
    }
    for (int pm = blockIdx.x; pm < n_perm; pm += gridDim.x) {
        double acc[K];
#pragma unroll
        for (int k = 0; k < K; ++k) acc[k] = 0.0;
        for (int c0 = 0; c0 < p; c0 += TC) {
            const int tc = min(TC, p - c0);
            // ---- A: one column per thread: its row order, then the permuted column into the tile ----------------
            for (int c = tid; c < tc; c += PCA_THREADS) {
                unsigned char* pr = perm + c * n;
                const size_t col = (size_t)c0 + c;
                if (perms_in) {
                    const unsigned char* src = perms_in + ((size_t)pm * p + col) * n;
                    for (int i = 0; i < n; ++i) pr[i] = src[i];
                } else {
                    for (int i = 0; i < n; ++i) pr[i] = (unsigned char)i;
                    unsigned long long x = pca_mix(seed ^ pca_mix((unsigned long long)pm * (unsigned long long)p + col + 0x9e3779b97f4a7c15ull));
                    for (int i = n - 1; i > 0; --i) {                         // Fisher-Yates, j uniform in [0, i]
                        x = pca_mix(x + 0x9e3779b97f4a7c15ull);
                        const int j = (int)__umul64hi(x, (unsigned long long)(i + 1));
                        const unsigned char t = pr[i]; pr[i] = pr[j]; pr[j] = t;
                    }
                }
                if (perms_out) {
                    unsigned char* dst = perms_out + ((size_t)pm * p + col) * n;
                    for (int i = 0; i < n; ++i) dst[i] = pr[i];
                }
                const double* xc = xct + col * n;
                double* v = V + c * n;
                for (int i = 0; i < n; ++i) v[i] = __ldg(xc + pr[i]);        // row i takes the value of row perm[i] (de_correlate)
            }
            __syncthreads();
            // ---- B: Gram accumulation ------------------------------------------------------------------------------
#pragma unroll 4
            for (int c = 0; c < tc; ++c) {
                const double* v = V + c * n;
#pragma unroll
                for (int k = 0; k < K; ++k) acc[k] = fma(v[pa[k]], v[pb[k]], acc[k]);
            }
            __syncthreads();
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int q = tid + k * PCA_THREADS;
            if (q < npairs) { G[pa[k] * n + pb[k]] = acc[k]; G[pb[k] * n + pa[k]] = acc[k]; }
        }
        __syncthreads();
        // ---- C: eigenvalues of G, cyclic Jacobi by warp 0 ----------------------------------------------------------
        if (tid < 32) {
            double tr = 0.0;
            for (int k = 0; k < n; ++k) tr += G[k * n + k];
            const double tiny = tr * tr * 1e-34;
            for (int sweep = 0; sweep < 40; ++sweep) {
                double off = 0.0;
                for (int a = lane; a < n; a += 32)
                    for (int b = 0; b < n; ++b) if (b != a) off = fma(G[a * n + b], G[a * n + b], off);
#pragma unroll
                for (int o = 16; o; o >>= 1) off += __shfl_xor_sync(0xffffffffu, off, o);
                if (!(off > tiny)) break;
                for (int i = 0; i < n - 1; ++i)
                    for (int j = i + 1; j < n; ++j) {
                        const double apq = G[i * n + j];
                        if (apq == 0.0) continue;                            // uniform: every lane reads the same word
                        const double app = G[i * n + i], aqq = G[j * n + j];
                        const double theta = (aqq - app) / (2.0 * apq);
                        const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(fma(theta, theta, 1.0)));
                        const double cs = 1.0 / sqrt(fma(t, t, 1.0)), sn = t * cs;
                        __syncwarp();
                        for (int k = lane; k < n; k += 32) {
                            if (k == i || k == j) continue;
                            const double gki = G[k * n + i], gkj = G[k * n + j];
                            const double ni = cs * gki - sn * gkj, nj = sn * gki + cs * gkj;
                            G[k * n + i] = ni; G[i * n + k] = ni;
                            G[k * n + j] = nj; G[j * n + k] = nj;
                        }
                        if (lane == 0) {
                            G[i * n + i] = app - t * apq; G[j * n + j] = aqq + t * apq;
                            G[i * n + j] = 0.0; G[j * n + i] = 0.0;
                        }
                        __syncwarp();
                    }
            }
            // descending order, negatives (rounding) to zero, ratios
            const int ncomp = min(n, p);
            double sum = 0.0;
            for (int k = 0; k < n; ++k) sum += fmax(G[k * n + k], 0.0);
            for (int a = lane; a < n; a += 32) {
                const double la = fmax(G[a * n + a], 0.0);
                int rank = 0;
                for (int b = 0; b < n; ++b) {
                    const double lb = fmax(G[b * n + b], 0.0);
                    rank += (lb > la || (lb == la && b < a)) ? 1 : 0;
                }
                if (rank < ncomp) evr[(size_t)pm * ncomp + rank] = sum > 0.0 ? la / sum : 0.0;
            }
        }
        __syncthreads();
    }
}
