// displacement-vector-field kernels next to the rigid path.
//
//  * warp_dvf_kernel -- the second stage of Deformable.create_image (structure/deformable.py:797-806):
//    sitk.Resample(image, DisplacementFieldTransform(dvf), sitkLinear, default).  Output voxel idx sits at physical
//    p; the field (one 3-vector per output voxel, reference layout [z][y][x][3], physical units) moves it to
//    p + d, and the image is sampled there: c = A . (idx, 1) + B . d, A = W2P_mov . P2W_out (3x4), B = the 3x3
//    linear part of W2P_mov.  Inside rule, edge clamp and rounding are those of the rigid kernels (ITK / VTK mode).
//  * sample_field_points_kernel -- scipy.ndimage.map_coordinates(field[..., c], [z, y, x], order=1,
//    mode='nearest') for a point set (structure/deformable.py:868-878, :1095-1104: ROI mesh points through the
//    DVF): coordinates are clamped to the volume, then trilinear.
//
// The displacement is arbitrary, so no TMA box can be promised: taps are cached global loads (smooth fields keep
// neighbouring threads on the same lines).  Coordinates are assembled in float64 (the displaced coordinate has no
// row structure to exploit) and split into an integer base + float32 fraction, as make_row_ctx does.
#pragma once
#include "mia_device.cuh"

namespace mia {

struct DvfParams {
    VolView mov;                 // image that is sampled
    const float *dvf;            // [oz][oy][ox][3] displacement, physical units (x, y, z)
    void *out;
    int ox, oy, oz;
    Mat34 A;                     // output index -> moving continuous index (without displacement)
    double B[9];                 // physical displacement -> moving index displacement (row-major 3x3)
    float background;
    int round_mode;
};

__device__ __forceinline__ RowCtx make_point_ctx(double cx, double cy, double cz, int nx, int ny, int nz)
{
    RowCtx r;
    const double big = 1.0e9;                 // NaN / far-away coordinates: outside, and the int conversion is safe
    cx = fmin(fmax(cx, -big), big);
    cy = fmin(fmax(cy, -big), big);
    cz = fmin(fmax(cz, -big), big);
    const double flx = floor(cx), fly = floor(cy), flz = floor(cz);
    r.bx = (int)flx; r.by = (int)fly; r.bz = (int)flz;
    r.fx = (float)(cx - flx); r.fy = (float)(cy - fly); r.fz = (float)(cz - flz);
    r.ax = r.ay = r.az = 0.0f;
    r.lox = -0.5f - (float)r.bx; r.loy = -0.5f - (float)r.by; r.loz = -0.5f - (float)r.bz;
    r.hix = (float)nx - 0.5f - (float)r.bx;
    r.hiy = (float)ny - 0.5f - (float)r.by;
    r.hiz = (float)nz - 0.5f - (float)r.bz;
    return r;
}

constexpr int kDvfThreads = 256;
#ifndef MIA_DVF_TZ
#define MIA_DVF_TZ 4
#endif
#ifndef MIA_DVF_MINB
#define MIA_DVF_MINB 4
#endif
constexpr int kDvfTZ = MIA_DVF_TZ;   // planes per thread: the float64 part of the coordinate is paid once per kDvfTZ voxels

// Work item = 32 (x) x 8 (y) x 8 (z) output voxels per CTA pass: lane = x (coalesced field reads and stores),
// warp = y row, each thread walks 8 z planes.  The rigid part of the coordinate is mapped in float64 once per
// thread and split into an integer base and a float32 fraction; the plane step and the displacement B . d are
// added in float32 LOCAL coordinates (|l| of a few voxels), so the weights stay exact to ~1e-7 voxel.
template <typename TM, typename TO, bool CLOSED>
__global__ void __launch_bounds__(kDvfThreads, MIA_DVF_MINB) warp_dvf_kernel(const DvfParams p)
{
    const TM *__restrict__ mov = static_cast<const TM *>(p.mov.p);
    TO *__restrict__ out = static_cast<TO *>(p.out);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ntx = (p.ox + 31) / 32, nty = (p.oy + 7) / 8, ntz = (p.oz + kDvfTZ - 1) / kDvfTZ;
    const long long ntiles = (long long)ntx * nty * ntz;
    const TO bg = to_out<TO>(p.background, p.round_mode);
    const double *A = p.A.a;
    const float a2x = (float)A[2], a2y = (float)A[6], a2z = (float)A[10];
    const float b00 = (float)p.B[0], b01 = (float)p.B[1], b02 = (float)p.B[2];
    const float b10 = (float)p.B[3], b11 = (float)p.B[4], b12 = (float)p.B[5];
    const float b20 = (float)p.B[6], b21 = (float)p.B[7], b22 = (float)p.B[8];
    const size_t plane = (size_t)p.oy * p.ox;
    for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const int tz = (int)(t / ((long long)ntx * nty)), r2 = (int)(t % ((long long)ntx * nty));
        const int i = (r2 % ntx) * 32 + lane, j = (r2 / ntx) * 8 + warp, k0 = tz * kDvfTZ;
        if (i >= p.ox || j >= p.oy) continue;
        const int kcnt = min(kDvfTZ, p.oz - k0);
        const double cx = ((A[3] + (double)k0 * A[2]) + (double)j * A[1]) + (double)i * A[0];
        const double cy = ((A[7] + (double)k0 * A[6]) + (double)j * A[5]) + (double)i * A[4];
        const double cz = ((A[11] + (double)k0 * A[10]) + (double)j * A[9]) + (double)i * A[8];
        const RowCtx r = make_point_ctx(cx, cy, cz, p.mov.nx, p.mov.ny, p.mov.nz);
        size_t v = ((size_t)k0 * p.oy + j) * p.ox + i;
        float d[kDvfTZ][3];
#pragma unroll
        for (int kz = 0; kz < kDvfTZ; ++kz) {          // all field loads of the thread in flight before the gathers
            const bool live = kz < kcnt;
            const float *q = p.dvf + 3 * (v + (size_t)kz * plane);
            d[kz][0] = live ? __ldg(q) : 0.0f;
            d[kz][1] = live ? __ldg(q + 1) : 0.0f;
            d[kz][2] = live ? __ldg(q + 2) : 0.0f;
        }
#pragma unroll
        for (int kz = 0; kz < kDvfTZ; ++kz) {
            if (kz < kcnt) {
                const float lx = fmaf((float)kz, a2x, r.fx) + fmaf(b02, d[kz][2], fmaf(b01, d[kz][1], b00 * d[kz][0]));
                const float ly = fmaf((float)kz, a2y, r.fy) + fmaf(b12, d[kz][2], fmaf(b11, d[kz][1], b10 * d[kz][0]));
                const float lz = fmaf((float)kz, a2z, r.fz) + fmaf(b22, d[kz][2], fmaf(b21, d[kz][1], b20 * d[kz][0]));
                float val = 0.0f;
                const bool in = sample_local<TM, CLOSED>(mov, p.mov.nx, p.mov.ny, p.mov.nz, r, lx, ly, lz, val);
                out[v + (size_t)kz * plane] = in ? to_out<TO>(val, p.round_mode) : bg;
            }
        }
    }
}

// field: [nz][ny][nx][C] float32; pts: [n][3] (x, y, z) continuous voxel coordinates, float64; out: [n][C] float64
// (map_coordinates returns the input dtype; float64 keeps a float64 DVF exact to float32 taps).
__global__ void __launch_bounds__(kDvfThreads)
sample_field_points_kernel(const float *__restrict__ field, int nx, int ny, int nz, int C, const double *__restrict__ pts,
                           long long n, double scale, double *__restrict__ out)
{
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x) {
        // mode='nearest': the coordinate is clamped to [0, N-1] (NaN -> 0 through fmax)
        const double x = fmin(fmax(pts[3 * q], 0.0), (double)(nx - 1));
        const double y = fmin(fmax(pts[3 * q + 1], 0.0), (double)(ny - 1));
        const double z = fmin(fmax(pts[3 * q + 2], 0.0), (double)(nz - 1));
        const double fx = floor(x), fy = floor(y), fz = floor(z);
        const int x0 = (int)fx, y0 = (int)fy, z0 = (int)fz;
        const int x1 = min(x0 + 1, nx - 1), y1 = min(y0 + 1, ny - 1), z1 = min(z0 + 1, nz - 1);
        const double wx = x - fx, wy = y - fy, wz = z - fz;
        const size_t r00 = ((size_t)z0 * ny + y0) * nx, r10 = ((size_t)z0 * ny + y1) * nx;
        const size_t r01 = ((size_t)z1 * ny + y0) * nx, r11 = ((size_t)z1 * ny + y1) * nx;
        for (int c = 0; c < C; ++c) {
            const double v000 = __ldg(field + (r00 + x0) * C + c), v100 = __ldg(field + (r00 + x1) * C + c);
            const double v010 = __ldg(field + (r10 + x0) * C + c), v110 = __ldg(field + (r10 + x1) * C + c);
            const double v001 = __ldg(field + (r01 + x0) * C + c), v101 = __ldg(field + (r01 + x1) * C + c);
            const double v011 = __ldg(field + (r11 + x0) * C + c), v111 = __ldg(field + (r11 + x1) * C + c);
            const double a = v000 + wx * (v100 - v000), b = v010 + wx * (v110 - v010);
            const double d = v001 + wx * (v101 - v001), e = v011 + wx * (v111 - v011);
            const double lo = a + wy * (b - a), hi = d + wy * (e - d);
            out[q * C + c] = scale * (lo + wz * (hi - lo));
        }
    }
}

}  // namespace mia
