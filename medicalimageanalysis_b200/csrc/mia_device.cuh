// Device-side building blocks of the rigid resample + metric kernels (sm_100a).
//
// Data layout in HBM: volumes are dense [z][y][x]; int16 (CT/MR/PT as the reference reader produces,
// read/dicom.py:530) or float32.  One thread owns V consecutive output voxels along x so the output
// store (and the fixed-image load of the fused metrics) is one 8/16-byte vector access, warps cover
// 32*V contiguous output voxels => fully coalesced 256/512-byte row segments.
//
// Coordinates: the first voxel of a thread is mapped through A in float64 (12 DFMA per V voxels),
// split into an integer base and a float32 fraction; the remaining V-1 voxels advance by the float32
// x-column of A from that fraction.  All float32 coordinates are therefore LOCAL (|c| < V + 1), which
// keeps the weight error near 1e-7 voxel regardless of volume size (a global float32 coordinate at
// x = 1000 would carry 6e-5) and makes the +-0.5 border test agree with the float64 oracle.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mia {

constexpr int kV = 4;          // output voxels per thread along x
constexpr int kBlockX = 32;    // threads along x  -> 128 voxels
constexpr int kBlockY = 8;     // rows per block
constexpr int kThreads = kBlockX * kBlockY;
constexpr int kTileX = kBlockX * kV;

struct VolView {
    const void *p;
    int nx, ny, nz;
};

struct Mat34 {
    double a[12];
};

// Global tap -> float32 through the read-only path.  int16 is loaded sign-extended into a 32-bit register so that
// the conversion is the half-rate ALU I2FP.F32.S32 and not the quarter-rate XU I2F.S16 a C++ short load compiles to.
__device__ __forceinline__ float ldg_f32(const float *p) { return __ldg(p); }
__device__ __forceinline__ float ldg_f32(const int16_t *p)
{
    int v;
    asm("ld.global.nc.s16 %0, [%1];" : "=r"(v) : "l"(p));
    return (float)v;
}

template <typename T>
__device__ __forceinline__ float ld_tap(const T *__restrict__ p, size_t i)
{
    return ldg_f32(p + i);
}

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }

// Per-thread sampling context: integer base of the first voxel and the float32 increments.
struct RowCtx {
    int bx, by, bz;        // floor of the first voxel's continuous index
    float fx, fy, fz;      // its fractional part (exact to 2^-25)
    float ax, ay, az;      // d(c)/d(i): x column of A in float32
    float lox, loy, loz;   // -0.5 - base   (exact)
    float hix, hiy, hiz;   // N - 0.5 - base (exact)
};

__device__ __forceinline__ RowCtx make_row_ctx(const double *__restrict__ A, int i, int j, int k,
                                               int nx, int ny, int nz)
{
    RowCtx r;
    // association order of vtkImageResliceExecute: origin + k*z + j*y + i*x
    double cx = ((A[3] + (double)k * A[2]) + (double)j * A[1]) + (double)i * A[0];
    double cy = ((A[7] + (double)k * A[6]) + (double)j * A[5]) + (double)i * A[4];
    double cz = ((A[11] + (double)k * A[10]) + (double)j * A[9]) + (double)i * A[8];
    // clamp far-away coordinates so the int conversion cannot overflow; anything beyond is outside anyway
    const double big = 1.0e9;
    cx = fmin(fmax(cx, -big), big);
    cy = fmin(fmax(cy, -big), big);
    cz = fmin(fmax(cz, -big), big);
    double flx = floor(cx), fly = floor(cy), flz = floor(cz);
    r.bx = (int)flx; r.by = (int)fly; r.bz = (int)flz;
    r.fx = (float)(cx - flx); r.fy = (float)(cy - fly); r.fz = (float)(cz - flz);
    r.ax = (float)A[0]; r.ay = (float)A[4]; r.az = (float)A[8];
    r.lox = -0.5f - (float)r.bx; r.loy = -0.5f - (float)r.by; r.loz = -0.5f - (float)r.bz;
    r.hix = (float)nx - 0.5f - (float)r.bx;
    r.hiy = (float)ny - 0.5f - (float)r.by;
    r.hiz = (float)nz - 0.5f - (float)r.bz;
    return r;
}

// Trilinear sample at LOCAL coordinates (lx, ly, lz) relative to the integer base of `r`.  Returns the inside
// flag; value in `val` (unrounded float32).  When all 8 taps lie inside the volume (the common case) they are
// addressed from one base pointer; otherwise neighbour indices are clamped to the buffer.
template <typename TM, bool CLOSED>
__device__ __forceinline__ bool sample_local(const TM *__restrict__ mov, int nx, int ny, int nz, const RowCtx &r,
                                             float lx, float ly, float lz, float &val)
{
    bool in = (lx >= r.lox) & (ly >= r.loy) & (lz >= r.loz);
    if (CLOSED)
        in &= (lx <= r.hix) & (ly <= r.hiy) & (lz <= r.hiz);
    else
        in &= (lx < r.hix) & (ly < r.hiy) & (lz < r.hiz);
    if (!in) return false;
    const float flx = floorf(lx), fly = floorf(ly), flz = floorf(lz);
    const float wx = lx - flx, wy = ly - fly, wz = lz - flz;
    const int ix = r.bx + (int)flx, iy = r.by + (int)fly, iz = r.bz + (int)flz;
    float v000, v100, v010, v110, v001, v101, v011, v111;
    if (((unsigned)ix < (unsigned)(nx - 1)) & ((unsigned)iy < (unsigned)(ny - 1)) & ((unsigned)iz < (unsigned)(nz - 1))) {
        const TM *__restrict__ q = mov + (((size_t)iz * ny + iy) * nx + ix);
        const size_t py = (size_t)nx, pz = (size_t)nx * ny;
        v000 = ldg_f32(q); v100 = ldg_f32(q + 1);
        v010 = ldg_f32(q + py); v110 = ldg_f32(q + py + 1);
        v001 = ldg_f32(q + pz); v101 = ldg_f32(q + pz + 1);
        v011 = ldg_f32(q + pz + py); v111 = ldg_f32(q + pz + py + 1);
    } else {
        const int x0 = clampi(ix, 0, nx - 1), x1 = clampi(ix + 1, 0, nx - 1);
        const int y0 = clampi(iy, 0, ny - 1), y1 = clampi(iy + 1, 0, ny - 1);
        const int z0 = clampi(iz, 0, nz - 1), z1 = clampi(iz + 1, 0, nz - 1);
        const size_t r00 = ((size_t)z0 * ny + y0) * nx, r10 = ((size_t)z0 * ny + y1) * nx;
        const size_t r01 = ((size_t)z1 * ny + y0) * nx, r11 = ((size_t)z1 * ny + y1) * nx;
        v000 = ld_tap(mov, r00 + x0); v100 = ld_tap(mov, r00 + x1);
        v010 = ld_tap(mov, r10 + x0); v110 = ld_tap(mov, r10 + x1);
        v001 = ld_tap(mov, r01 + x0); v101 = ld_tap(mov, r01 + x1);
        v011 = ld_tap(mov, r11 + x0); v111 = ld_tap(mov, r11 + x1);
    }
    const float vx00 = fmaf(wx, v100 - v000, v000);
    const float vx10 = fmaf(wx, v110 - v010, v010);
    const float vx01 = fmaf(wx, v101 - v001, v001);
    const float vx11 = fmaf(wx, v111 - v011, v011);
    const float vxx0 = fmaf(wy, vx10 - vx00, vx00);
    const float vxx1 = fmaf(wy, vx11 - vx01, vx01);
    val = fmaf(wz, vxx1 - vxx0, vxx0);
    return true;
}

// Sample voxel number v (0..V-1) of the thread.  Returns inside flag; value in `val` (unrounded float32).
template <typename TM, bool CLOSED>
__device__ __forceinline__ bool sample_voxel(const TM *__restrict__ mov, int nx, int ny, int nz,
                                             const RowCtx &r, int v, float &val)
{
    return sample_local<TM, CLOSED>(mov, nx, ny, nz, r, fmaf((float)v, r.ax, r.fx), fmaf((float)v, r.ay, r.fy),
                                    fmaf((float)v, r.az, r.fz), val);
}

// Moving-image mask at the mapped point of voxel v: the mask voxel nearest to the continuous index (floor(c + 0.5),
// ITK's RoundHalfIntegerUp), clamped to the volume like the taps.
__device__ __forceinline__ bool nearest_mask_voxel(const unsigned char *__restrict__ mask, int nx, int ny, int nz,
                                                   const RowCtx &r, int v)
{
    const int ix = clampi(r.bx + (int)floorf(fmaf((float)v, r.ax, r.fx) + 0.5f), 0, nx - 1);
    const int iy = clampi(r.by + (int)floorf(fmaf((float)v, r.ay, r.fy) + 0.5f), 0, ny - 1);
    const int iz = clampi(r.bz + (int)floorf(fmaf((float)v, r.az, r.fz) + 0.5f), 0, nz - 1);
    return __ldg(mask + ((size_t)iz * ny + iy) * nx + ix) != 0;
}

// float -> output element.  ROUND: 0 = floor(v + 0.5) (VTK), 1 = trunc (ITK); clamp to the type range first.
template <typename TO>
__device__ __forceinline__ TO to_out(float v, int round_mode);

template <>
__device__ __forceinline__ float to_out<float>(float v, int) { return v; }

template <>
__device__ __forceinline__ int16_t to_out<int16_t>(float v, int round_mode)
{
    v = fminf(fmaxf(v, -32768.0f), 32767.0f);
    return (int16_t)(round_mode == 0 ? floorf(v + 0.5f) : truncf(v));
}

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ unsigned warp_sum_u(unsigned v)
{
    return __reduce_add_sync(0xffffffffu, v);
}

}  // namespace mia
