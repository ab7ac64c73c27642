// dcd_unpack.cuh -- raw DCD frame records -> float32 [F][N][3] on the device.
//
// A DCD frame stores x[N], y[N], z[N] as three Fortran records (4-byte length before and after each), optionally after
// a 56-byte unit-cell record; VMD's dcd plugin, which the reference reaches through contact_utils.py:89-116, turns that
// into interleaved coordinates on the host.  Here the file bytes cross PCIe as they are (chacra_b200/trajectory.py) and
// this kernel interleaves them: coalesced 4-byte loads of the three arrays, a shared-memory tile, coalesced stores.
// HBM bound: 12 N bytes read + 12 N written per frame.
#pragma once

#define DCD_TILE 256

__global__ void __launch_bounds__(DCD_TILE)
dcd_unpack_kernel(const unsigned char* __restrict__ raw, long long first, long long pitch, int n_frames, int n_atoms, int swap,
                  float* __restrict__ xyz, long long frame_stride) {
    __shared__ float s[3][DCD_TILE + 1];
    const int tid = threadIdx.x;
    const int rec = n_atoms + 2;                      // words from one coordinate array to the next (two record markers)
    const int ntiles = (n_atoms + DCD_TILE - 1) / DCD_TILE;
    for (int f = blockIdx.y; f < n_frames; f += gridDim.y) {
        const unsigned* X = reinterpret_cast<const unsigned*>(raw + first + (long long)f * pitch);
        float* out = xyz + (size_t)f * frame_stride;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const int i = tile * DCD_TILE + tid;
            if (i < n_atoms) {
#pragma unroll
                for (int d = 0; d < 3; ++d) {
                    unsigned w = __ldg(X + (size_t)d * rec + i);
                    if (swap) w = __byte_perm(w, 0u, 0x0123);
                    s[d][tid] = __uint_as_float(w);
                }
            }
            __syncthreads();
            const int cnt = 3 * min(DCD_TILE, n_atoms - tile * DCD_TILE);
            float* o = out + 3 * (size_t)tile * DCD_TILE;
            for (int k = tid; k < cnt; k += DCD_TILE) o[k] = s[k % 3][k / 3];
            __syncthreads();
        }
    }
}
