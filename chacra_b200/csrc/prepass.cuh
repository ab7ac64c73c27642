// prepass.cuh -- per-fragment hydrophobic carbon sets (hydrophobics.py:8-41).
//
// The reference recomputes, for every 100-frame fragment, which carbons of ALA PHE GLY ILE LEU PRO VAL TRP
// (or ligand) have only C/H atoms within 1.7 A *in that fragment's first frame*; sele1 uses a subset test,
// sele2 an equality test against {C, H}.  One warp per base carbon scans the whole frame (every stored
// atom counts, hydrogens and unselected atoms included) with the exact float32 d2 <= 1.7f*1.7f rule.
#pragma once
#include "device_types.cuh"

__global__ void __launch_bounds__(256)
hp_prepass_kernel(DevTopo T, float r2, const float* __restrict__ xyz, long long frame_stride, long long first_index,
                  int frag_frames, long long frag0, int first_g, unsigned char* __restrict__ hpmask) {
    const int g = first_g + blockIdx.x;
    const long long k = (frag0 + g) * (long long)frag_frames - first_index;   // frame of this call
    const float* __restrict__ X = xyz + (size_t)k * frame_stride;
    const int lane = threadIdx.x & 31, warps = blockDim.x >> 5;
    for (int b = blockIdx.y * warps + (threadIdx.x >> 5); b < T.n_hp_base; b += gridDim.y * warps) {
        const int p = T.hp_base[b];
        const float* c = X + 3 * (size_t)T.pair_atom[p];
        const float cx = c[0], cy = c[1], cz = c[2];
        unsigned seen = 0;
        for (int a = lane; a < T.n_atoms; a += 32) {
            const float* q = X + 3 * (size_t)a;
            if (dist2_exact(q[0], q[1], q[2], cx, cy, cz) <= r2) {
                unsigned cl = T.atom_class[a];
                seen |= cl == 1 ? 1u : (cl == 2 ? 2u : 4u);
            }
        }
        seen = __reduce_or_sync(0xffffffffu, seen);
        if (lane == 0) {
            unsigned sub = (seen & 4u) ? 0u : 1u;            // elements within 1.7 A are a subset of {C, H}
            unsigned eq = (sub && (seen & 2u) && (seen & 1u)) ? 2u : 0u;   // ... exactly {C, H}
            hpmask[(size_t)g * T.n_pair + p] = (unsigned char)(sub | eq);
        }
    }
}
