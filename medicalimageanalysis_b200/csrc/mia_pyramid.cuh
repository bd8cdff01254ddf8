// SURVEY N5 (first part): one level of a multi-resolution pyramid.  The reference delegates multi-resolution
// registration to ITK / Elastix (utils/deformable/simpleitk.py:150-163: "NumberOfResolutions"); the pyramid level
// built here is the block mean -- every output voxel is the average of the fx * fy * fz input voxels it covers
// (fewer on the high edges), an anti-aliasing box filter followed by the shrink in one pass.  HBM-bound streaming:
// every input voxel is read once, N_in * b + N_out * b bytes.
#pragma once
#include "mia_device.cuh"

namespace mia {

template <typename T>
__device__ __forceinline__ T mean_to_out(float s, float inv);
template <>
__device__ __forceinline__ float mean_to_out<float>(float s, float inv) { return s * inv; }
template <>
__device__ __forceinline__ int16_t mean_to_out<int16_t>(float s, float inv)
{
    return (int16_t)floorf(fminf(fmaxf(s * inv, -32768.0f), 32767.0f) + 0.5f);       // half-up, like the resampler
}

// One thread per output voxel, x fastest: a warp reads 32 * fx consecutive input elements per row (coalesced) and
// writes 32 consecutive outputs.  Sums are float32 in x, y, z order (exact for int16 blocks of up to 2^8 voxels).
template <typename T>
__global__ void __launch_bounds__(256) shrink_mean_kernel(const T *__restrict__ in, int nx, int ny, int nz, T *__restrict__ out,
                                                          int ox, int oy, int oz, int fx, int fy, int fz)
{
    const unsigned long long n = (unsigned long long)ox * oy * oz;
    for (unsigned long long o = (unsigned long long)blockIdx.x * 256u + threadIdx.x; o < n; o += (unsigned long long)gridDim.x * 256u) {
        const int i = (int)(o % (unsigned)ox);
        const unsigned long long r = o / (unsigned)ox;
        const int j = (int)(r % (unsigned)oy), k = (int)(r / (unsigned)oy);
        const int x0 = i * fx, y0 = j * fy, z0 = k * fz;
        const int x1 = min(nx, x0 + fx), y1 = min(ny, y0 + fy), z1 = min(nz, z0 + fz);
        float s = 0.0f;
        for (int z = z0; z < z1; ++z)
            for (int y = y0; y < y1; ++y) {
                const T *row = in + ((size_t)z * ny + y) * nx;
                for (int x = x0; x < x1; ++x) s += ldg_f32(row + x);
            }
        const float cnt = (float)((x1 - x0) * (y1 - y0) * (z1 - z0));
        out[o] = mean_to_out<T>(s, 1.0f / cnt);
    }
}

// Factor (2, 2, fz) on even nx with a 4-byte aligned int16 base / 8-byte aligned float base: one 2-voxel vector
// load per row instead of two scalar ones.
template <typename T, typename V2>
__global__ void __launch_bounds__(256) shrink_mean_x2_kernel(const T *__restrict__ in, int nx, int ny, int nz, T *__restrict__ out,
                                                             int ox, int oy, int oz, int fy, int fz)
{
    const unsigned long long n = (unsigned long long)ox * oy * oz;
    for (unsigned long long o = (unsigned long long)blockIdx.x * 256u + threadIdx.x; o < n; o += (unsigned long long)gridDim.x * 256u) {
        const int i = (int)(o % (unsigned)ox);
        const unsigned long long r = o / (unsigned)ox;
        const int j = (int)(r % (unsigned)oy), k = (int)(r / (unsigned)oy);
        const int y0 = j * fy, z0 = k * fz;
        const int y1 = min(ny, y0 + fy), z1 = min(nz, z0 + fz);
        float s = 0.0f;
        for (int z = z0; z < z1; ++z)
            for (int y = y0; y < y1; ++y) {
                const V2 v = __ldg(reinterpret_cast<const V2 *>(in + ((size_t)z * ny + y) * nx) + i);
                s += (float)v.x;
                s += (float)v.y;
            }
        const float cnt = (float)(2 * (y1 - y0) * (z1 - z0));
        out[o] = mean_to_out<T>(s, 1.0f / cnt);
    }
}

}  // namespace mia
