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
#include "mia_tile.cuh"

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
    unsigned magic_fold;         // staged kernel: see tile_compute (mia_tile.cuh)
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
// min / max that PROPAGATE NaN (fminf / fmaxf drop it): a NaN displacement must poison the column's coordinate range so
// that the column fails the "strictly interior" test and takes the per-voxel branch, where NaN compares as outside.
__device__ __forceinline__ float min_nan(float a, float b) { float r; asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float max_nan(float a, float b) { float r; asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float min3_nan(float a, float b, float c) { float r; asm("min.NaN.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }   // FMNMX3
__device__ __forceinline__ float max3_nan(float a, float b, float c) { float r; asm("max.NaN.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }

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


// ---------------------------------------------------------------------------------------------------------------
// Box-staged DVF warp.  A deformation field from a registration is smooth: inside one 32 x 8 x 8 output tile the
// displaced coordinates stay within a few voxels of the rigid ones, so the tile's taps fit a shared-memory box like
// the rigid tile kernel's -- only the box is not known before the field has been read.  Per tile a CTA therefore
// (1) loads its 2048 displacement vectors (lane = x: 384 contiguous bytes per warp and plane) and forms the local
// coordinates, (2) reduces the integer bounding box of the INSIDE voxels' taps over the CTA (REDUX + 8 shared
// slots), (3) has one thread fetch that box with one TMA copy (zero fill outside the volume = the edge clamp),
// (4) computes from shared memory with the packed arithmetic of tile_compute.  A tile whose taps do not fit the box
// (a tear in the field) takes the per-voxel gather of warp_dvf_kernel, so any field stays correct.  The phases of
// one CTA are serial; 3-4 resident CTAs per SM overlap them.
#ifndef MIA_DVF_STAGED_MINB
#define MIA_DVF_STAGED_MINB 3
#endif
#ifndef MIA_DVF_PREFETCH
#define MIA_DVF_PREFETCH 0
#endif
// DIAG: B is diagonal (image axes along the physical axes, the usual case): one FFMA per axis instead of three.
template <typename TM, typename TO, bool CLOSED, class BOX, bool DIAG>
__global__ void __launch_bounds__(kDvfThreads, MIA_DVF_STAGED_MINB)
warp_dvf_tile_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ DvfParams p)
{
    constexpr unsigned ES = sizeof(TM);
    constexpr int PY = BOX::kX * ES, PZ = BOX::kX * BOX::kY * ES;
    constexpr int xalign = 16 / (int)sizeof(TM);
    constexpr unsigned kBoxBytes = BOX::kX * BOX::kY * BOX::kZ * ES;
    extern __shared__ __align__(1024) unsigned char dsm[];
    __shared__ int s_box[8][6];
    __shared__ __align__(8) unsigned long long s_bar;

    const TM *__restrict__ mov = static_cast<const TM *>(p.mov.p);
    TO *__restrict__ out = static_cast<TO *>(p.out);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ntx = (p.ox + 31) / 32, nty = (p.oy + 7) / 8, ntz = (p.oz + kTZ - 1) / kTZ;
    const unsigned ntiles = (unsigned)ntx * nty * ntz;
    const TO bg = tile_to_out<TO, CLOSED>(p.background);
    const double *A = p.A.a;
    const float a2x = (float)A[2], a2y = (float)A[6], a2z = (float)A[10];
    const float b00 = (float)p.B[0], b01 = (float)p.B[1], b02 = (float)p.B[2];
    const float b10 = (float)p.B[3], b11 = (float)p.B[4], b12 = (float)p.B[5];
    const float b20 = (float)p.B[6], b21 = (float)p.B[7], b22 = (float)p.B[8];
    const unsigned plane = (unsigned)p.oy * (unsigned)p.ox;
    const int nx = p.mov.nx, ny = p.mov.ny, nz = p.mov.nz;
    const uint32_t bar = smem_u32(&s_bar);
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    unsigned phase = 0;
    const f32x2 magic2 = pk(kMagic, kMagic);

    // The field of tile t + gridDim.x can be requested while tile t is computed (MIA_DVF_PREFETCH): of the two serial
    // round trips per tile -- field, then box -- the first then overlaps the arithmetic of the previous tile.
    // Measured and left OFF: the 24 extra live registers cost more than the overlap gains -- 3 CTAs of 80 registers
    // spill (int16 410 -> 469 us), 2 CTAs of 128 registers lose the third CTA (442 us).
    float d[kTZ][3];
    auto fetch = [&](unsigned tt) {
        if (tt >= ntiles) return;
        const int tz = (int)(tt / (unsigned)(ntx * nty)), r2 = (int)(tt % (unsigned)(ntx * nty));
        const int i = (r2 % ntx) * 32 + lane, j = (r2 / ntx) * 8 + warp, k0 = tz * kTZ;
        const int kcnt = ((i < p.ox) & (j < p.oy)) ? min(kTZ, p.oz - k0) : 0;
        const float *q = p.dvf + 3 * (((size_t)k0 * p.oy + j) * p.ox + i);
        if (kcnt == kTZ) {                            // full column: no predicates, one stepped 64-bit pointer
#pragma unroll
            for (int kz = 0; kz < kTZ; ++kz) {        // all field loads in flight first
                asm("" : "+l"(q));
                d[kz][0] = __ldg(q);
                d[kz][1] = __ldg(q + 1);
                d[kz][2] = __ldg(q + 2);
                q += 3u * plane;
            }
        } else {
#pragma unroll
            for (int kz = 0; kz < kTZ; ++kz) {
                const bool live = kz < kcnt;
                d[kz][0] = live ? __ldg(q) : 0.0f;
                d[kz][1] = live ? __ldg(q + 1) : 0.0f;
                d[kz][2] = live ? __ldg(q + 2) : 0.0f;
                q += 3u * plane;
            }
        }
    };
    if (MIA_DVF_PREFETCH) fetch(blockIdx.x);

    for (unsigned t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const int tz = (int)(t / (unsigned)(ntx * nty)), r2 = (int)(t % (unsigned)(ntx * nty));
        const int i = (r2 % ntx) * 32 + lane, j = (r2 / ntx) * 8 + warp, k0 = tz * kTZ;
        const bool row_ok = (i < p.ox) & (j < p.oy);
        const int kcnt = row_ok ? min(kTZ, p.oz - k0) : 0;
        // rigid part in float64 once per thread, split into integer base + float32 fraction (as make_row_ctx)
        const double cx = ((A[3] + (double)k0 * A[2]) + (double)j * A[1]) + (double)i * A[0];
        const double cy = ((A[7] + (double)k0 * A[6]) + (double)j * A[5]) + (double)i * A[4];
        const double cz = ((A[11] + (double)k0 * A[10]) + (double)j * A[9]) + (double)i * A[8];
        const RowCtx r = make_point_ctx(cx, cy, cz, nx, ny, nz);
        const size_t v = ((size_t)k0 * p.oy + j) * p.ox + i;
        // local coordinates of the 8 voxels (relative to the thread base r.b*), inside flags, clamped like BORDER tiles
        float lx[kTZ], ly[kTZ], lz[kTZ];
        unsigned inbits = 0u;
        const float hix = CLOSED ? r.hix : prev_float(r.hix), hiy = CLOSED ? r.hiy : prev_float(r.hiy), hiz = CLOSED ? r.hiz : prev_float(r.hiz);
        const float clox = -(float)r.bx, cloy = -(float)r.by, cloz = -(float)r.bz;
        const float chix = (float)(nx - 1) - (float)r.bx, chiy = (float)(ny - 1) - (float)r.by, chiz = (float)(nz - 1) - (float)r.bz;
        // Fast path (ncu: the first cut spent 80 lane-instructions per voxel in this phase, half of the kernel, on
        // per-voxel inside tests, clamps and FRND / F2I for the bounding box): form the coordinates, take the thread's
        // float min / max, and if all 8 voxels are strictly interior (no clamp, no test needed) floor only those six
        // numbers.  Border columns, partial z tiles and fields that leave the volume take the per-voxel path.
        float fmnx = INFINITY, fmny = INFINITY, fmnz = INFINITY, fmxx = -INFINITY, fmxy = -INFINITY, fmxz = -INFINITY;
        {
            if (!MIA_DVF_PREFETCH) fetch(t);
#pragma unroll
            for (int kz = 0; kz < kTZ; ++kz) {
                const float rx = fmaf((float)kz, a2x, r.fx), ry = fmaf((float)kz, a2y, r.fy), rz = fmaf((float)kz, a2z, r.fz);
                if (DIAG) {
                    lx[kz] = fmaf(b00, d[kz][0], rx);
                    ly[kz] = fmaf(b11, d[kz][1], ry);
                    lz[kz] = fmaf(b22, d[kz][2], rz);
                } else {
                    lx[kz] = rx + fmaf(b02, d[kz][2], fmaf(b01, d[kz][1], b00 * d[kz][0]));
                    ly[kz] = ry + fmaf(b12, d[kz][2], fmaf(b11, d[kz][1], b10 * d[kz][0]));
                    lz[kz] = rz + fmaf(b22, d[kz][2], fmaf(b21, d[kz][1], b20 * d[kz][0]));
                }
                fmnx = min_nan(fmnx, lx[kz]); fmxx = max_nan(fmxx, lx[kz]);
                fmny = min_nan(fmny, ly[kz]); fmxy = max_nan(fmxy, ly[kz]);
                fmnz = min_nan(fmnz, lz[kz]); fmxz = max_nan(fmxz, lz[kz]);
            }
        }
        // NaN coordinates (kept by min_nan / max_nan) fail the comparisons below and take the per-voxel path, where they are outside
        const bool interior = (kcnt == kTZ) & (fmnx >= clox) & (fmxx <= chix) & (fmny >= cloy) & (fmxy <= chiy) &
                              (fmnz >= cloz) & (fmxz <= chiz);
        if (interior) {
            inbits = 0xffu;
        } else {
            fmnx = fmny = fmnz = INFINITY;
            fmxx = fmxy = fmxz = -INFINITY;
#pragma unroll
            for (int kz = 0; kz < kTZ; ++kz) {
                float x = lx[kz], y = ly[kz], z = lz[kz];
                const bool in = (kz < kcnt) & (x >= r.lox) & (x <= hix) & (y >= r.loy) & (y <= hiy) & (z >= r.loz) & (z <= hiz);
                x = fminf(fmaxf(x, clox), chix); y = fminf(fmaxf(y, cloy), chiy); z = fminf(fmaxf(z, cloz), chiz);
                lx[kz] = x; ly[kz] = y; lz[kz] = z;
                if (in) {
                    inbits |= 1u << kz;
                    fmnx = fminf(fmnx, x); fmxx = fmaxf(fmxx, x);
                    fmny = fminf(fmny, y); fmxy = fmaxf(fmxy, y);
                    fmnz = fminf(fmnz, z); fmxz = fmaxf(fmxz, z);
                }
            }
        }
        const int mnx = (int)floorf(fmnx), mxx = (int)floorf(fmxx), mny = (int)floorf(fmny), mxy = (int)floorf(fmxy),
                  mnz = (int)floorf(fmnz), mxz = (int)floorf(fmxz);           // garbage when inbits == 0: not used then
        // bounding box of the taps in moving index space: thread -> warp (REDUX) -> CTA
        const bool any = inbits != 0u;
        const int big = 0x7fffffff;
        const int wmnx = __reduce_min_sync(0xffffffffu, any ? r.bx + mnx : big), wmxx = __reduce_max_sync(0xffffffffu, any ? r.bx + mxx : -big);
        const int wmny = __reduce_min_sync(0xffffffffu, any ? r.by + mny : big), wmxy = __reduce_max_sync(0xffffffffu, any ? r.by + mxy : -big);
        const int wmnz = __reduce_min_sync(0xffffffffu, any ? r.bz + mnz : big), wmxz = __reduce_max_sync(0xffffffffu, any ? r.bz + mxz : -big);
        if (lane == 0) {
            s_box[warp][0] = wmnx; s_box[warp][1] = wmxx; s_box[warp][2] = wmny;
            s_box[warp][3] = wmxy; s_box[warp][4] = wmnz; s_box[warp][5] = wmxz;
        }
        __syncthreads();
        int bx0 = big, bx1 = -big, by0 = big, by1 = -big, bz0 = big, bz1 = -big;
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            bx0 = min(bx0, s_box[w][0]); bx1 = max(bx1, s_box[w][1]);
            by0 = min(by0, s_box[w][2]); by1 = max(by1, s_box[w][3]);
            bz0 = min(bz0, s_box[w][4]); bz1 = max(bz1, s_box[w][5]);
        }
        const bool none = bx0 == big;                          // CTA-uniform from here on
        bx0 &= ~(xalign - 1);                                   // TMA: inner coordinate * element size = 0 mod 16 bytes
        const bool fits = !none && (bx1 + 2 - bx0 <= BOX::kX) && (by1 + 2 - by0 <= BOX::kY) && (bz1 + 2 - bz0 <= BOX::kZ);
        TO *o = out + v;
        if (fits && threadIdx.x == 0) {
            mbar_arrive_expect_tx(bar, kBoxBytes);
            tma_load_3d(smem_u32(dsm), &tmap, bar, bx0, by0, bz0);
        }
        if (MIA_DVF_PREFETCH) fetch(t + gridDim.x);           // d[] is dead here: the coordinates are in lx / ly / lz
        if (none) {
#pragma unroll
            for (int kz = 0; kz < kTZ; ++kz)
                if (kz < kcnt) st_out(o + (size_t)kz * plane, bg);
        } else if (!fits) {
            // a tear in the field: per-voxel gather (coordinates were clamped above; weights beyond the edge are 0)
#pragma unroll
            for (int kz = 0; kz < kTZ; ++kz) {
                if (kz < kcnt) {
                    float val = 0.0f;
                    const bool in = (inbits >> kz) & 1u;
                    if (in) sample_local<TM, true>(mov, nx, ny, nz, r, lx[kz], ly[kz], lz[kz], val);
                    st_out(o + (size_t)kz * plane, in ? tile_to_out<TO, CLOSED>(val) : bg);
                }
            }
        } else {
            const int tbx = r.bx - bx0, tby = r.by - by0, tbz = r.bz - bz0;
            const uint32_t ubase = smem_u32(dsm) + (uint32_t)((tbz * BOX::kY + tby) * BOX::kX + tbx) * ES + p.magic_fold;
            // a voxel that is outside contributed nothing to the box, so it must not address beyond it either: its
            // (unused) taps are taken at box element (0, 0, 0)
            const float sx = (float)(-tbx), sy = (float)(-tby), sz = (float)(-tbz);
            mbar_wait<64>(bar, phase);
            phase ^= 1u;
#pragma unroll
            for (int kp = 0; kp < kTZ; kp += 2) {
                const bool i0 = (inbits >> kp) & 1u, i1 = (inbits >> (kp + 1)) & 1u;
                const f32x2 lx2 = pk(i0 ? lx[kp] : sx, i1 ? lx[kp + 1] : sx), ly2 = pk(i0 ? ly[kp] : sy, i1 ? ly[kp + 1] : sy),
                            lz2 = pk(i0 ? lz[kp] : sz, i1 ? lz[kp + 1] : sz);
                const f32x2 tx2 = add2_rm(lx2, magic2), ty2 = add2_rm(ly2, magic2), tz2 = add2_rm(lz2, magic2);
                const f32x2 wx2 = sub2(lx2, sub2(tx2, magic2)), wy2 = sub2(ly2, sub2(ty2, magic2)), wz2 = sub2(lz2, sub2(tz2, magic2));
                const uint32_t a0 = mad_u32(mad_u32(mad_u32((uint32_t)tz2, BOX::kY, (uint32_t)ty2), BOX::kX, (uint32_t)tx2), ES, ubase);
                const uint32_t a1 = mad_u32(mad_u32(mad_u32((uint32_t)(tz2 >> 32), BOX::kY, (uint32_t)(ty2 >> 32)), BOX::kX, (uint32_t)(tx2 >> 32)), ES, ubase);
                const f32x2 v000 = pk(Tap<TM, 0>::ld(a0), Tap<TM, 0>::ld(a1));
                const f32x2 v100 = pk(Tap<TM, ES>::ld(a0), Tap<TM, ES>::ld(a1));
                const f32x2 v010 = pk(Tap<TM, PY>::ld(a0), Tap<TM, PY>::ld(a1));
                const f32x2 v110 = pk(Tap<TM, PY + ES>::ld(a0), Tap<TM, PY + ES>::ld(a1));
                const f32x2 v001 = pk(Tap<TM, PZ>::ld(a0), Tap<TM, PZ>::ld(a1));
                const f32x2 v101 = pk(Tap<TM, PZ + ES>::ld(a0), Tap<TM, PZ + ES>::ld(a1));
                const f32x2 v011 = pk(Tap<TM, PZ + PY>::ld(a0), Tap<TM, PZ + PY>::ld(a1));
                const f32x2 v111 = pk(Tap<TM, PZ + PY + ES>::ld(a0), Tap<TM, PZ + PY + ES>::ld(a1));
                const f32x2 vx00 = lerp2(wx2, v000, v100), vx10 = lerp2(wx2, v010, v110);
                const f32x2 vx01 = lerp2(wx2, v001, v101), vx11 = lerp2(wx2, v011, v111);
                const f32x2 vxx0 = lerp2(wy2, vx00, vx10), vxx1 = lerp2(wy2, vx01, vx11);
                float val0, val1;
                upk(lerp2(wz2, vxx0, vxx1), val0, val1);
                const TO o0 = tile_to_out<TO, CLOSED>(val0), o1 = tile_to_out<TO, CLOSED>(val1);
                if (kp < kcnt) st_out(o, ((inbits >> kp) & 1u) ? o0 : bg);
                o += plane;
                if (kp + 1 < kcnt) st_out(o, ((inbits >> (kp + 1)) & 1u) ? o1 : bg);
                o += plane;
            }
        }
        __syncthreads();          // the box and s_box are free again
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Software-pipelined form of the box-staged warp: BOTH round trips of a tile (field, box) overlap arithmetic.
//   iteration for tile t:   [coordinates + bounding box of tile t+1, from its field buffer]     <- box(t) in flight
//                           [wait box(t); interpolate tile t; store]
//                           __syncthreads
//                           [one thread: request box(t+1) and the field of tile t+2 + ...]
// The coordinates of tile t+1 stay in registers while tile t is interpolated (24 floats more than the serial kernel),
// so this kernel runs TWO CTAs of up to 128 registers per SM instead of three of 80, with the field double-buffered
// (box + 2 x 24 KB).  One CTA barrier per tile instead of two.
constexpr unsigned kDvfFieldBytes = 96u * 8u * 8u * 4u;     // one tile of the field: 32 x 8 x 8 vectors of 3 floats

struct DvfTile {                 // what the coordinate phase leaves for the interpolation phase, per thread
    float lx[kTZ], ly[kTZ], lz[kTZ];      // local coordinates relative to (bx, by, bz), clamped on border columns
    unsigned inbits;                      // voxel kz is inside the moving volume
    int bx, by, bz;                       // thread base: floor of the rigid coordinate of the first voxel
    int kcnt;                             // planes of this column that exist (0: column outside the lattice)
    unsigned v;                           // output offset of the first voxel (< 2^31: checked by the host)
};

struct DvfConsts { float a2x, a2y, a2z, b[9]; };       // float32 images of A's z column and of B, formed once per thread

struct DvfWalk {                 // tile coordinates of the tiles t0, t0 + g, t0 + 2 g, ... without a division per tile
    int tix, tiy, tiz, step_x, step_y, step_z, ntx, nty;
    __device__ __forceinline__ void init(unsigned t0, unsigned g, int ntx_, int nty_)
    {
        ntx = ntx_; nty = nty_;
        const unsigned txy = (unsigned)(ntx * nty);
        tiz = (int)(t0 / txy); tiy = (int)(t0 % txy) / ntx; tix = (int)(t0 % txy) % ntx;
        step_z = (int)(g / txy); step_y = (int)(g % txy) / ntx; step_x = (int)(g % txy) % ntx;
    }
    __device__ __forceinline__ void advance()
    {
        tix += step_x;
        if (tix >= ntx) { tix -= ntx; ++tiy; }
        tiy += step_y;
        if (tiy >= nty) { tiy -= nty; ++tiz; }
        tiz += step_z;
    }
};

template <bool CLOSED, bool DIAG>
__device__ __forceinline__ void dvf_coords(const DvfParams &p, const DvfConsts &c, const float *fbuf, const DvfWalk &w, int lane,
                                           int warp, DvfTile &st, int (*sbox)[6])
{
    const double *A = p.A.a;
    const int nx = p.mov.nx, ny = p.mov.ny, nz = p.mov.nz;
    const int i = w.tix * 32 + lane, j = w.tiy * 8 + warp, k0 = w.tiz * kTZ;
    const bool row_ok = (i < p.ox) & (j < p.oy);
    st.kcnt = row_ok ? min(kTZ, p.oz - k0) : 0;
    const double cx = ((A[3] + (double)k0 * A[2]) + (double)j * A[1]) + (double)i * A[0];
    const double cy = ((A[7] + (double)k0 * A[6]) + (double)j * A[5]) + (double)i * A[4];
    const double cz = ((A[11] + (double)k0 * A[10]) + (double)j * A[9]) + (double)i * A[8];
    const RowCtx r = make_point_ctx(cx, cy, cz, nx, ny, nz);
    st.bx = r.bx; st.by = r.by; st.bz = r.bz;
    st.v = ((unsigned)k0 * (unsigned)p.oy + (unsigned)j) * (unsigned)p.ox + (unsigned)i;
    const float hix = CLOSED ? r.hix : prev_float(r.hix), hiy = CLOSED ? r.hiy : prev_float(r.hiy), hiz = CLOSED ? r.hiz : prev_float(r.hiz);
    const float clox = -(float)r.bx, cloy = -(float)r.by, cloz = -(float)r.bz;
    const float chix = (float)(nx - 1) - (float)r.bx, chiy = (float)(ny - 1) - (float)r.by, chiz = (float)(nz - 1) - (float)r.bz;
    float fmnx = INFINITY, fmny = INFINITY, fmnz = INFINITY, fmxx = -INFINITY, fmxy = -INFINITY, fmxz = -INFINITY;
    const float *q = fbuf + warp * 96 + 3 * lane;
    // two planes per step: packed FFMA2 for the coordinates, one 3-input FMNMX per range end (NaN kept, see min_nan)
    const f32x2 ax2 = pk(c.a2x, c.a2x), ay2 = pk(c.a2y, c.a2y), az2 = pk(c.a2z, c.a2z);
    const f32x2 fx2 = pk(r.fx, r.fx), fy2 = pk(r.fy, r.fy), fz2 = pk(r.fz, r.fz);
#pragma unroll
    for (int kz = 0; kz < kTZ; kz += 2) {
        const f32x2 d0 = pk(q[kz * 768], q[kz * 768 + 768]), d1 = pk(q[kz * 768 + 1], q[kz * 768 + 769]),
                    d2 = pk(q[kz * 768 + 2], q[kz * 768 + 770]);                           // zero past the field's edge
        const f32x2 kk = pk((float)kz, (float)(kz + 1));
        const f32x2 rx = fma2(kk, ax2, fx2), ry = fma2(kk, ay2, fy2), rz = fma2(kk, az2, fz2);
        f32x2 x2, y2, z2;
        if (DIAG) {
            x2 = fma2(pk(c.b[0], c.b[0]), d0, rx);
            y2 = fma2(pk(c.b[4], c.b[4]), d1, ry);
            z2 = fma2(pk(c.b[8], c.b[8]), d2, rz);
        } else {
            x2 = add2(rx, fma2(pk(c.b[2], c.b[2]), d2, fma2(pk(c.b[1], c.b[1]), d1, mul2(pk(c.b[0], c.b[0]), d0))));
            y2 = add2(ry, fma2(pk(c.b[5], c.b[5]), d2, fma2(pk(c.b[4], c.b[4]), d1, mul2(pk(c.b[3], c.b[3]), d0))));
            z2 = add2(rz, fma2(pk(c.b[8], c.b[8]), d2, fma2(pk(c.b[7], c.b[7]), d1, mul2(pk(c.b[6], c.b[6]), d0))));
        }
        upk(x2, st.lx[kz], st.lx[kz + 1]);
        upk(y2, st.ly[kz], st.ly[kz + 1]);
        upk(z2, st.lz[kz], st.lz[kz + 1]);
        fmnx = min3_nan(fmnx, st.lx[kz], st.lx[kz + 1]); fmxx = max3_nan(fmxx, st.lx[kz], st.lx[kz + 1]);
        fmny = min3_nan(fmny, st.ly[kz], st.ly[kz + 1]); fmxy = max3_nan(fmxy, st.ly[kz], st.ly[kz + 1]);
        fmnz = min3_nan(fmnz, st.lz[kz], st.lz[kz + 1]); fmxz = max3_nan(fmxz, st.lz[kz], st.lz[kz + 1]);
    }
    // strictly interior columns need no test and no clamp (a NaN range fails the comparisons)
    const bool interior = (st.kcnt == kTZ) & (fmnx >= clox) & (fmxx <= chix) & (fmny >= cloy) & (fmxy <= chiy) &
                          (fmnz >= cloz) & (fmxz <= chiz);
    if (interior) {
        st.inbits = 0xffu;
    } else {
        st.inbits = 0u;
        fmnx = fmny = fmnz = INFINITY;
        fmxx = fmxy = fmxz = -INFINITY;
#pragma unroll
        for (int kz = 0; kz < kTZ; ++kz) {
            float x = st.lx[kz], y = st.ly[kz], z = st.lz[kz];
            const bool in = (kz < st.kcnt) & (x >= r.lox) & (x <= hix) & (y >= r.loy) & (y <= hiy) & (z >= r.loz) & (z <= hiz);
            x = fminf(fmaxf(x, clox), chix); y = fminf(fmaxf(y, cloy), chiy); z = fminf(fmaxf(z, cloz), chiz);
            st.lx[kz] = x; st.ly[kz] = y; st.lz[kz] = z;
            if (in) {
                st.inbits |= 1u << kz;
                fmnx = fminf(fmnx, x); fmxx = fmaxf(fmxx, x);
                fmny = fminf(fmny, y); fmxy = fmaxf(fmxy, y);
                fmnz = fminf(fmnz, z); fmxz = fmaxf(fmxz, z);
            }
        }
    }
    const int mnx = (int)floorf(fmnx), mxx = (int)floorf(fmxx), mny = (int)floorf(fmny), mxy = (int)floorf(fmxy),
              mnz = (int)floorf(fmnz), mxz = (int)floorf(fmxz);           // garbage when inbits == 0: not used then
    const bool any = st.inbits != 0u;
    const int big = 0x7fffffff;
    const int wmnx = __reduce_min_sync(0xffffffffu, any ? r.bx + mnx : big), wmxx = __reduce_max_sync(0xffffffffu, any ? r.bx + mxx : -big);
    const int wmny = __reduce_min_sync(0xffffffffu, any ? r.by + mny : big), wmxy = __reduce_max_sync(0xffffffffu, any ? r.by + mxy : -big);
    const int wmnz = __reduce_min_sync(0xffffffffu, any ? r.bz + mnz : big), wmxz = __reduce_max_sync(0xffffffffu, any ? r.bz + mxz : -big);
    if (lane == 0) {
        sbox[warp][0] = wmnx; sbox[warp][1] = wmxx; sbox[warp][2] = wmny;
        sbox[warp][3] = wmxy; sbox[warp][4] = wmnz; sbox[warp][5] = wmxz;
    }
}

struct DvfBoxPlace { int bx0, by0, bz0; bool none, fits; };

// CTA bounding box from the 8 per-warp boxes: lane w < 8 takes warp w's six numbers, six REDUX finish the job (every
// thread reading all 48 numbers cost 52 instructions per thread and tile, 6 % of the kernel).
template <typename TM, class BOX>
__device__ __forceinline__ DvfBoxPlace dvf_place(const int (*sbox)[6], int lane)
{
    constexpr int xalign = 16 / (int)sizeof(TM);
    const int big = 0x7fffffff;
    const bool have = lane < 8;
    const int *mine = sbox[lane & 7];
    int bx0 = __reduce_min_sync(0xffffffffu, have ? mine[0] : big), bx1 = __reduce_max_sync(0xffffffffu, have ? mine[1] : -big);
    const int by0 = __reduce_min_sync(0xffffffffu, have ? mine[2] : big), by1 = __reduce_max_sync(0xffffffffu, have ? mine[3] : -big);
    const int bz0 = __reduce_min_sync(0xffffffffu, have ? mine[4] : big), bz1 = __reduce_max_sync(0xffffffffu, have ? mine[5] : -big);
    DvfBoxPlace b;
    b.none = bx0 == big;
    bx0 &= ~(xalign - 1);                                   // TMA: inner coordinate * element size = 0 mod 16 bytes
    b.fits = !b.none && (bx1 + 2 - bx0 <= BOX::kX) && (by1 + 2 - by0 <= BOX::kY) && (bz1 + 2 - bz0 <= BOX::kZ);
    b.bx0 = bx0; b.by0 = by0; b.bz0 = bz0;
    return b;
}

template <typename TM, typename TO, bool CLOSED, class BOX>
__device__ __forceinline__ void dvf_interp(const DvfParams &p, const DvfTile &st, const DvfBoxPlace &b, uint32_t box_smem,
                                           uint32_t bar, unsigned &phase)
{
    constexpr unsigned ES = sizeof(TM);
    constexpr int PY = BOX::kX * ES, PZ = BOX::kX * BOX::kY * ES;
    const unsigned plane = (unsigned)p.oy * (unsigned)p.ox;
    const TO bg = tile_to_out<TO, CLOSED>(p.background);
    TO *o = static_cast<TO *>(p.out) + st.v;
    const f32x2 magic2 = pk(kMagic, kMagic);
    if (b.none) {
#pragma unroll
        for (int kz = 0; kz < kTZ; ++kz)
            if (kz < st.kcnt) st_out(o + (size_t)kz * plane, bg);
    } else if (!b.fits) {
        // a tear in the field: per-voxel gather (coordinates were clamped; weights beyond the edge are 0)
        RowCtx r;
        r.bx = st.bx; r.by = st.by; r.bz = st.bz;
        r.fx = r.fy = r.fz = 0.0f; r.ax = r.ay = r.az = 0.0f;
        r.lox = -0.5f - (float)r.bx; r.loy = -0.5f - (float)r.by; r.loz = -0.5f - (float)r.bz;
        r.hix = (float)p.mov.nx - 0.5f - (float)r.bx; r.hiy = (float)p.mov.ny - 0.5f - (float)r.by; r.hiz = (float)p.mov.nz - 0.5f - (float)r.bz;
        const TM *__restrict__ mov = static_cast<const TM *>(p.mov.p);
#pragma unroll
        for (int kz = 0; kz < kTZ; ++kz) {
            if (kz < st.kcnt) {
                float val = 0.0f;
                const bool in = (st.inbits >> kz) & 1u;
                if (in) sample_local<TM, true>(mov, p.mov.nx, p.mov.ny, p.mov.nz, r, st.lx[kz], st.ly[kz], st.lz[kz], val);
                st_out(o + (size_t)kz * plane, in ? tile_to_out<TO, CLOSED>(val) : bg);
            }
        }
    } else {
        const int tbx = st.bx - b.bx0, tby = st.by - b.by0, tbz = st.bz - b.bz0;
        const uint32_t ubase = box_smem + (uint32_t)((tbz * BOX::kY + tby) * BOX::kX + tbx) * ES + p.magic_fold;
        // a voxel that is outside contributed nothing to the box: its (unused) taps are taken at box element (0, 0, 0)
        const float sx = (float)(-tbx), sy = (float)(-tby), sz = (float)(-tbz);
        mbar_wait<64>(bar, phase);
        phase ^= 1u;
#pragma unroll
        for (int kp = 0; kp < kTZ; kp += 2) {
            const bool i0 = (st.inbits >> kp) & 1u, i1 = (st.inbits >> (kp + 1)) & 1u;
            const f32x2 lx2 = pk(i0 ? st.lx[kp] : sx, i1 ? st.lx[kp + 1] : sx), ly2 = pk(i0 ? st.ly[kp] : sy, i1 ? st.ly[kp + 1] : sy),
                        lz2 = pk(i0 ? st.lz[kp] : sz, i1 ? st.lz[kp + 1] : sz);
            const f32x2 tx2 = add2_rm(lx2, magic2), ty2 = add2_rm(ly2, magic2), tz2 = add2_rm(lz2, magic2);
            const f32x2 wx2 = sub2(lx2, sub2(tx2, magic2)), wy2 = sub2(ly2, sub2(ty2, magic2)), wz2 = sub2(lz2, sub2(tz2, magic2));
            const uint32_t a0 = mad_u32(mad_u32(mad_u32((uint32_t)tz2, BOX::kY, (uint32_t)ty2), BOX::kX, (uint32_t)tx2), ES, ubase);
            const uint32_t a1 = mad_u32(mad_u32(mad_u32((uint32_t)(tz2 >> 32), BOX::kY, (uint32_t)(ty2 >> 32)), BOX::kX, (uint32_t)(tx2 >> 32)), ES, ubase);
#ifdef MIA_BOUNDS_CHECK
            // Debug build (compute-sanitizer is closed on this pool): all 8 taps of both voxels inside the staged box
            if (a0 < box_smem || a0 + PZ + PY + 2u * ES > box_smem + BOX::kX * BOX::kY * BOX::kZ * ES ||
                a1 < box_smem || a1 + PZ + PY + 2u * ES > box_smem + BOX::kX * BOX::kY * BOX::kZ * ES)
                __trap();
#endif
            const f32x2 v000 = pk(Tap<TM, 0>::ld(a0), Tap<TM, 0>::ld(a1));
            const f32x2 v100 = pk(Tap<TM, ES>::ld(a0), Tap<TM, ES>::ld(a1));
            const f32x2 v010 = pk(Tap<TM, PY>::ld(a0), Tap<TM, PY>::ld(a1));
            const f32x2 v110 = pk(Tap<TM, PY + ES>::ld(a0), Tap<TM, PY + ES>::ld(a1));
            const f32x2 v001 = pk(Tap<TM, PZ>::ld(a0), Tap<TM, PZ>::ld(a1));
            const f32x2 v101 = pk(Tap<TM, PZ + ES>::ld(a0), Tap<TM, PZ + ES>::ld(a1));
            const f32x2 v011 = pk(Tap<TM, PZ + PY>::ld(a0), Tap<TM, PZ + PY>::ld(a1));
            const f32x2 v111 = pk(Tap<TM, PZ + PY + ES>::ld(a0), Tap<TM, PZ + PY + ES>::ld(a1));
            const f32x2 vx00 = lerp2(wx2, v000, v100), vx10 = lerp2(wx2, v010, v110);
            const f32x2 vx01 = lerp2(wx2, v001, v101), vx11 = lerp2(wx2, v011, v111);
            const f32x2 vxx0 = lerp2(wy2, vx00, vx10), vxx1 = lerp2(wy2, vx01, vx11);
            float val0, val1;
            upk(lerp2(wz2, vxx0, vxx1), val0, val1);
            const TO o0 = tile_to_out<TO, CLOSED>(val0), o1 = tile_to_out<TO, CLOSED>(val1);
            if (kp < st.kcnt) st_out(o, i0 ? o0 : bg);
            o += plane;
            if (kp + 1 < st.kcnt) st_out(o, i1 ? o1 : bg);
            o += plane;
        }
    }
}

#ifndef MIA_DVF_PIPE_MINB
#define MIA_DVF_PIPE_MINB 2
#endif
template <class BOX, typename TM, int NBOX>
struct DvfPipeSmem {
    static constexpr unsigned kBoxBytes = BOX::kX * BOX::kY * BOX::kZ * (unsigned)sizeof(TM);
    static constexpr unsigned kBoxStage = (kBoxBytes + 127u) & ~127u;
    static constexpr unsigned kFieldOff = kBoxStage * NBOX;
    static constexpr unsigned kBytes = kFieldOff + 2u * kDvfFieldBytes;
};

// NBOX = 2 (int16: two 80 x 15 x 13 boxes fit beside the field buffers of two CTAs): the box of tile t+1 is requested as
// soon as its coordinates are reduced, BEFORE tile t is interpolated, into the other box buffer -- it has the whole
// interpolation of tile t and the coordinate phase of tile t+2 to arrive, and the barrier after the interpolation goes.
// NBOX = 1 (float32: 46 KB boxes): box(t+1) is requested after tile t has been interpolated and arrives during the
// coordinate phase of tile t+2.
template <typename TM, typename TO, bool CLOSED, class BOX, bool DIAG, int NBOX>
__global__ void __launch_bounds__(kDvfThreads, MIA_DVF_PIPE_MINB)
warp_dvf_pipe_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap fmap,
                     const __grid_constant__ DvfParams p)
{
    using L = DvfPipeSmem<BOX, TM, NBOX>;
    extern __shared__ __align__(1024) unsigned char dsm[];
    __shared__ int s_box[2][8][6];
    __shared__ __align__(8) unsigned long long s_bar[2], s_fbar[2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ntx = (p.ox + 31) / 32, nty = (p.oy + 7) / 8, ntz = (p.oz + kTZ - 1) / kTZ;
    const unsigned ntiles = (unsigned)ntx * nty * ntz;
    const uint32_t box_smem = smem_u32(dsm);
    const float *fbuf0 = reinterpret_cast<const float *>(dsm + L::kFieldOff);
    auto request_field = [&](unsigned tt, unsigned b) {       // one thread
        const int tz = (int)(tt / (unsigned)(ntx * nty)), r2 = (int)(tt % (unsigned)(ntx * nty));
        const uint32_t fb = smem_u32(&s_fbar[b]);
        mbar_arrive_expect_tx(fb, kDvfFieldBytes);
        tma_load_3d(smem_u32(fbuf0) + b * kDvfFieldBytes, &fmap, fb, (r2 % ntx) * 96, (r2 / ntx) * 8, tz * kTZ);
    };
    auto request_box = [&](const DvfBoxPlace &b, unsigned buf) {     // one thread
        const uint32_t bb = smem_u32(&s_bar[buf]);
        mbar_arrive_expect_tx(bb, L::kBoxBytes);
        tma_load_3d(box_smem + buf * L::kBoxStage, &tmap, bb, b.bx0, b.by0, b.bz0);
    };
    const unsigned t0 = blockIdx.x, g = gridDim.x;
    if (threadIdx.x == 0) {
        mbar_init(smem_u32(&s_bar[0]), 1);
        mbar_init(smem_u32(&s_bar[1]), 1);
        mbar_init(smem_u32(&s_fbar[0]), 1);
        mbar_init(smem_u32(&s_fbar[1]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (t0 < ntiles) request_field(t0, 0u);
        if (t0 + g < ntiles) request_field(t0 + g, 1u);
    }
    __syncthreads();
    if (t0 >= ntiles) return;
    unsigned ph0 = 0, ph1 = 0, fph0 = 0, fph1 = 0;
    DvfConsts c;
    c.a2x = (float)p.A.a[2]; c.a2y = (float)p.A.a[6]; c.a2z = (float)p.A.a[10];
#pragma unroll
    for (int q = 0; q < 9; ++q) c.b[q] = (float)p.B[q];

    // prologue: coordinates and box of the first tile
    DvfWalk walk;
    walk.init(t0, g, ntx, nty);
    DvfTile tileA, tileB;
    DvfBoxPlace placeA, placeB;
    mbar_wait<64>(smem_u32(&s_fbar[0]), fph0);
    fph0 ^= 1u;
    dvf_coords<CLOSED, DIAG>(p, c, fbuf0, walk, lane, warp, tileA, s_box[0]);
    walk.advance();
    __syncthreads();
    placeA = dvf_place<TM, BOX>(s_box[0], lane);
    if (threadIdx.x == 0) {
        if (placeA.fits) request_box(placeA, 0u);
        if (t0 + 2 * g < ntiles) request_field(t0 + 2 * g, 0u);
    }
    unsigned t = t0;
    // One tile: `cur` / `pc` were prepared by the previous step (field buffer and s_box slot CB, box buffer CB or 0); the
    // step prepares `nxt` / `pn` in the other slot.  Two steps with the roles swapped make one loop iteration, so the
    // 24 coordinates per thread are never copied between registers.
    auto step = [&](auto cb_tag, DvfTile &cur, DvfBoxPlace &pc, DvfTile &nxt, DvfBoxPlace &pn) -> bool {
        constexpr unsigned CB = decltype(cb_tag)::value, NB = 1u - CB;
        constexpr unsigned BB = NBOX == 2 ? CB : 0u;          // box buffer of THIS tile
        const bool more = t + g < ntiles;
        if (more) {                                           // coordinates of tile t + g while box(t) is in flight
            unsigned &fph = NB ? fph1 : fph0;
            mbar_wait<64>(smem_u32(&s_fbar[NB]), fph);
            fph ^= 1u;
            dvf_coords<CLOSED, DIAG>(p, c, fbuf0 + NB * (kDvfFieldBytes / 4u), walk, lane, warp, nxt, s_box[NB]);
            walk.advance();
        }
        if (NBOX == 2) {
            // s_box[NB] complete, field buffer NB read by everyone, and box buffer NB (tile t - g) long done with
            __syncthreads();
            if (more) {
                pn = dvf_place<TM, BOX>(s_box[NB], lane);
                if (threadIdx.x == 0) {
                    if (pn.fits) request_box(pn, NB);
                    if (t + 3 * g < ntiles) request_field(t + 3 * g, NB);
                }
            }
            dvf_interp<TM, TO, CLOSED, BOX>(p, cur, pc, box_smem + BB * L::kBoxStage, smem_u32(&s_bar[BB]), BB ? ph1 : ph0);
        } else {
            dvf_interp<TM, TO, CLOSED, BOX>(p, cur, pc, box_smem, smem_u32(&s_bar[0]), ph0);
            __syncthreads();      // the box is free, s_box[NB] is complete, field buffer NB has been read by everyone
            if (more) {
                pn = dvf_place<TM, BOX>(s_box[NB], lane);
                if (threadIdx.x == 0) {
                    if (pn.fits) request_box(pn, 0u);
                    if (t + 3 * g < ntiles) request_field(t + 3 * g, NB);
                }
            }
        }
        t += g;
        return more;
    };
    for (;;) {
        if (!step(std::integral_constant<unsigned, 0u>{}, tileA, placeA, tileB, placeB)) break;
        if (!step(std::integral_constant<unsigned, 1u>{}, tileB, placeB, tileA, placeA)) break;
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
