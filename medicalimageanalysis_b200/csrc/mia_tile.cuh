// TMA-staged tile kernel of the rigid resample + metric hot path (sm_100a).
//
// Why: the direct kernel in mia_kernels.cuh spends ~200 lane-instructions per output voxel (64-bit
// index math, six clamps, eight LDG, float64 setup per 4 voxels) and is issue-bound at 7 % of the HBM
// roofline.  Here the moving-volume box an output tile touches is staged in shared memory by TMA
// (cp.async.bulk.tensor.3d, zero instructions in the compute warps, zero fill outside the volume) and
// the gather becomes 8 LDS with compile-time offsets + 8 I2FP + 7 lerps on thread-local float32
// coordinates: ~50 lane-instructions per voxel.
//
// Shape: persistent CTAs of 9 warps.  Warp 8 (one lane) is the TMA producer: per tile it maps the tile
// origin through A in float64, classifies the tile (outside / interior / border / does-not-fit) and
// issues one 3-D box copy into a kStages-deep ring guarded by full/empty mbarriers.  Warps 0..7 are
// consumers: an output tile is 32 (x) x 8 (y) x 8 (z); lane = x (so LDS and the fixed/out accesses of a
// warp are contiguous), warp = y row, each thread walks the 8 z planes of the tile.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "mia_kernels.cuh"

namespace mia {

constexpr int kTX = 32, kTY = 8, kTZ = 8;       // output tile
constexpr int kConsumerWarps = kTY;
constexpr int kTileThreads = 32 * (kConsumerWarps + 1);

enum TileClass { kOutside = 0, kInterior = 1, kBorder = 2, kNoFit = 3 };

template <int X, int Y, int Z>
struct Box {
    static constexpr int kX = X, kY = Y, kZ = Z;
};

struct TileInfo {          // producer -> consumers, one per ring stage
    long long rx, ry, rz;  // (continuous moving index of the tile's first voxel - box origin) in Q31.32 fixed point
    int bx, by, bz;        // moving index of box element [0][0][0]
    int cls;
};

// ---- mbarrier / TMA primitives (inline PTX; no CUTLASS dependency) ------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}

// Wait for a phase of an mbarrier.  try_wait blocks in hardware until the phase completes or a time limit passes; the
// explicit suspend-time hint (nanoseconds) keeps a waiting warp off the issue port for that long.  The first cut looped
// over nanosleep + try_wait: ncu counted 12 % of all issued instructions in those loops (nanosleep returns after a few
// nanoseconds on this part), stolen from warps that had work.  NS = hint; the loop only re-arms after a time-out.
template <int NS>
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
        "@!p bra WAIT_%=;\n"
        "}\n" ::"r"(bar), "r"(parity), "r"((uint32_t)(NS < 1000 ? 20000 : 1000000))
        : "memory");
}

__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int x, int y, int z)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(z), "r"(bar)
        : "memory");
}

// floor through the mantissa: (l + 1.5*2^23) rounded toward -inf keeps floor(l) in the low mantissa bits
constexpr float kMagic = 12582912.0f;
constexpr uint32_t kMagicBits = 0x4B400000u;

__device__ __forceinline__ float add_rm(float a, float b)
{
    float r;
    asm("add.rm.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}

template <typename TO, bool VTK>
__device__ __forceinline__ TO tile_to_out(float v);

template <>
__device__ __forceinline__ float tile_to_out<float, true>(float v) { return v; }
template <>
__device__ __forceinline__ float tile_to_out<float, false>(float v) { return v; }
template <>
__device__ __forceinline__ int16_t tile_to_out<int16_t, true>(float v)   // floor(v + 0.5), saturated
{
    short r;
    asm("cvt.rmi.sat.s16.f32 %0, %1;" : "=h"(r) : "f"(v + 0.5f));
    return r;
}
template <>
__device__ __forceinline__ int16_t tile_to_out<int16_t, false>(float v)  // trunc(v), saturated
{
    short r;
    asm("cvt.rzi.sat.s16.f32 %0, %1;" : "=h"(r) : "f"(v));
    return r;
}

// Device-side kernel selection for matrices that live in device memory (graph-captured registration loops: the host
// never sees them).  Two kernels that serve the same call are BOTH enqueued with complementary gates; every CTA
// evaluates the same cheap predicate -- does a tile of (ex+1, ey+1, ez+1) voxels fit a box (bx, by, bz) under every
// matrix of the launch? -- and the kernel whose gate fails returns at once (a few microseconds), so no round trip
// through the host is needed and the sequence stays capturable.
struct FitGate {
    int mode;              // 0 = off (always run), 1 = run only if every matrix fits, 2 = run only if some matrix does not
    int ex, ey, ez;        // tile extents minus one
    int bx, by, bz;        // box
    int slack;             // rows / planes the fit must leave free (room for the spread of the probes of one iteration)
};

__host__ __device__ inline bool gate_fits(const double *A, const FitGate &g)
{
    const double ext[3] = {(double)g.ex, (double)g.ey, (double)g.ez};
    const int B[3] = {g.bx, g.by, g.bz};
    for (int a = 0; a < 3; ++a) {
        double n = 0.0, q = 0.0;
        for (int c = 0; c < 3; ++c) {
            const double d = A[a * 4 + c] * ext[c];
            if (d < 0.0) n += d; else q += d;
        }
        // + 2 taps + 1 for the floor of both ends, + 7 along x for the 16-byte TMA origin (int16)
        if (!(ceil(q - n) + 3.0 + (a == 0 ? 7.0 : (double)g.slack) <= (double)B[a])) return false;
    }
    return true;
}

__device__ __forceinline__ bool gate_pass(const FitGate &g, const double *dA, int n_mats)
{
    if (g.mode == 0) return true;
    bool all = true;
    for (int m = 0; m < n_mats; ++m) all &= gate_fits(dA + 12 * m, g);
    return g.mode == 1 ? all : !all;
}

// Extent of one output tile in moving index space: negative / positive reach from the tile origin.
struct Reach {
    double neg[3], pos[3];
};

__host__ __device__ inline Reach tile_reach(const double *A)
{
    Reach r;
    const double ext[3] = {(double)(kTX - 1), (double)(kTY - 1), (double)(kTZ - 1)};
    for (int a = 0; a < 3; ++a) {
        double n = 0.0, q = 0.0;
        for (int c = 0; c < 3; ++c) {
            const double d = A[a * 4 + c] * ext[c];
            if (d < 0.0) n += d; else q += d;
        }
        r.neg[a] = n;
        r.pos[a] = q;
    }
    return r;
}

struct TileAabb {           // bounding box of one tile in moving index space, padded and clipped to the volume
    double c0[3];           // continuous moving index of the tile's first voxel
    int lo[3], hi[3];       // first box element needed, one past the last (upper tap included)
    bool outside, interior;
};

__host__ __device__ inline TileAabb tile_aabb(const double *A, const Reach &r, int i0, int j0, int k0, int nx, int ny, int nz)
{
    TileAabb t;
    t.c0[0] = ((A[3] + (double)k0 * A[2]) + (double)j0 * A[1]) + (double)i0 * A[0];
    t.c0[1] = ((A[7] + (double)k0 * A[6]) + (double)j0 * A[5]) + (double)i0 * A[4];
    t.c0[2] = ((A[11] + (double)k0 * A[10]) + (double)j0 * A[9]) + (double)i0 * A[8];
    const int n[3] = {nx, ny, nz};
    t.outside = false;
    t.interior = true;
    for (int a = 0; a < 3; ++a) {
        t.lo[a] = t.hi[a] = 0;
        const double lo = t.c0[a] + r.neg[a], hi = t.c0[a] + r.pos[a];
        if (!(hi >= -0.5) || !(lo <= (double)n[a] - 0.5)) { t.outside = true; continue; }   // NaN -> outside
        if (!(lo >= 0.0 && hi <= (double)(n[a] - 1))) t.interior = false;
        const double top = (double)(n[a] - 1);
        double lc = lo < 0.0 ? 0.0 : (lo > top ? top : lo);
        double hc = hi < 0.0 ? 0.0 : (hi > top ? top : hi);
        lc -= 1.0e-3;                      // float32 thread coordinates may stray ~1e-6 beyond the AABB
        hc += 1.0e-3;
        t.lo[a] = (int)floor(lc);
        t.hi[a] = (int)floor(hc) + 2;
    }
    return t;
}

// Does the tile's AABB fit a box whose element [0][0][0] is moving index b?
template <class BOX>
__host__ __device__ inline bool aabb_fits(const TileAabb &t, int bx, int by, int bz)
{
    return t.lo[0] >= bx && t.lo[1] >= by && t.lo[2] >= bz && t.hi[0] - bx <= BOX::kX && t.hi[1] - by <= BOX::kY &&
           t.hi[2] - bz <= BOX::kZ;
}

// The record the consumers need for a tile whose box origin has been chosen.
template <class BOX>
__host__ __device__ inline TileInfo tile_place(const TileAabb &bb, int bx, int by, int bz)
{
    TileInfo t;
    t.bx = bx; t.by = by; t.bz = bz;
    t.cls = bb.outside ? kOutside : (!aabb_fits<BOX>(bb, bx, by, bz) ? kNoFit : (bb.interior ? kInterior : kBorder));
    t.rx = t.ry = t.rz = 0;
    if (t.cls == kInterior || t.cls == kBorder) {       // |c0 - b| is a few box widths at most here
        t.rx = (long long)floor((bb.c0[0] - (double)bx) * 4294967296.0);
        t.ry = (long long)floor((bb.c0[1] - (double)by) * 4294967296.0);
        t.rz = (long long)floor((bb.c0[2] - (double)bz) * 4294967296.0);
    }
    return t;
}

// Classify one tile and choose its own box.  Shared by the device producer and the host (launch heuristics,
// unit tests).
// TMA constraint measured on B200 (tests/cuda/tma_probe3.cu): with INTERLEAVE_NONE the innermost box
// coordinate times the element size must be a multiple of 16 bytes, otherwise the copy raises "illegal
// instruction".  The box origin along x is therefore rounded down to `xalign` (= 16 / sizeof(element)) voxels.
template <class BOX>
__host__ __device__ inline TileInfo classify_tile(const double *A, const Reach &r, int i0, int j0, int k0,
                                                  int nx, int ny, int nz, int xalign)
{
    const TileAabb bb = tile_aabb(A, r, i0, j0, k0, nx, ny, nz);
    return tile_place<BOX>(bb, bb.lo[0] & ~(xalign - 1), bb.lo[1], bb.lo[2]);   // & rounds toward -inf (two's complement)
}


__device__ __forceinline__ float prev_float(float h)
{
    const int b = __float_as_int(h);
    return __int_as_float(h > 0.0f ? b - 1 : (h < 0.0f ? b + 1 : (int)0x80000001));
}

// Shared-memory tap -> float32.  int16 is loaded sign-extended (LDS.S16) so that the conversion is the
// half-rate ALU I2FP.F32.S32 and not the quarter-rate XU I2F.S16 the compiler picks for a C++ short load.
template <int BYTE_OFF>
__device__ __forceinline__ float tap_i16(uint32_t addr)
{
    int v;
    asm volatile("ld.shared.s16 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(BYTE_OFF));
    return (float)v;
}
template <int BYTE_OFF>
__device__ __forceinline__ float tap_f32(uint32_t addr)
{
    float v;
    asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v) : "r"(addr), "n"(BYTE_OFF));
    return v;
}
template <typename TM, int BYTE_OFF>
struct Tap;
template <int BYTE_OFF>
struct Tap<int16_t, BYTE_OFF> {
    static __device__ __forceinline__ float ld(uint32_t a) { return tap_i16<BYTE_OFF>(a); }
};
template <int BYTE_OFF>
struct Tap<float, BYTE_OFF> {
    static __device__ __forceinline__ float ld(uint32_t a) { return tap_f32<BYTE_OFF>(a); }
};

__device__ __forceinline__ void st_out(int16_t *p, int16_t v) { asm volatile("st.global.cs.u16 [%0], %1;" ::"l"(p), "h"(v) : "memory"); }
__device__ __forceinline__ void st_out(float *p, float v) { asm volatile("st.global.cs.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory"); }

// ---- packed float32x2 arithmetic (Blackwell FADD2 / FFMA2): one issue slot for two voxels ------------
// The kernel is issue-bound, not FP32-pipe bound (ncu: 71 % issue slots, 31 % FMA pipe), so the z-adjacent
// voxel pair (kz, kz+1) of a thread shares every coordinate and lerp instruction.
typedef unsigned long long f32x2;

// a * b + c as ONE integer multiply-add (the compiler otherwise strength-reduces the box addressing into
// a longer shift / mask / add sequence)
__device__ __forceinline__ uint32_t mad_u32(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}

__device__ __forceinline__ f32x2 pk(float lo, float hi)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void upk(f32x2 v, float &lo, float &hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 add2_rm(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("add.rm.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 lerp2(f32x2 w, f32x2 a, f32x2 b) { return fma2(w, sub2(b, a), a); }

// Histogram bins of a voxel pair as raw bits kMagicBits + bin: the arithmetic of bin_of() -- (v - vmin) * scale
// with separate roundings, clamp, floor -- with the floor taken by a round-down add of 1.5 * 2^23 instead of
// FRND + F2I (both quarter-rate XU) and the pair sharing the packed instructions.  Clamping to the integers
// [0, bins-1] before the floor equals clamping after it; NaN goes to bin 0 through fmaxf as in bin_of().
__device__ __forceinline__ f32x2 bin_bits2(f32x2 v2, f32x2 vmin2, f32x2 scale2, float binmax, f32x2 magic2)
{
    float b0, b1;
    upk(mul2(sub2(v2, vmin2), scale2), b0, b1);
    b0 = fminf(fmaxf(b0, 0.0f), binmax);
    b1 = fminf(fmaxf(b1, 0.0f), binmax);
    return add2_rm(pk(b0, b1), magic2);
}
__device__ __forceinline__ void red_shared_inc(uint32_t addr)
{
    asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(addr) : "memory");
}

// (Warp-aggregated counter updates -- one MATCH.ALL per voxel, a single lane adding 32 when the whole warp lands in one
// bin pair, as it does in air -- were built and measured on the config-3 phantoms: 51.4 us per evaluation against
// 45.7 with plain per-lane atomics, same histogram.  The match costs more than the serialised wavefronts it removes.)

// Per-thread constants of a launch: (lane * A[:,0] + warp * A[:,1]) in Q31.32 and the z column in float32.
struct ThreadGeom {
    long long ox, oy, oz;
    float a2x, a2y, a2z;
};

__device__ __forceinline__ ThreadGeom make_thread_geom(const double *A, int lane, int warp)
{
    ThreadGeom g;
    g.ox = __double2ll_rn(((double)lane * A[0] + (double)warp * A[1]) * 4294967296.0);
    g.oy = __double2ll_rn(((double)lane * A[4] + (double)warp * A[5]) * 4294967296.0);
    g.oz = __double2ll_rn(((double)lane * A[8] + (double)warp * A[9]) * 4294967296.0);
    g.a2x = (float)A[2]; g.a2y = (float)A[6]; g.a2z = (float)A[10];
    return g;
}

// Fixed-image values of one tile for this thread (all 8 planes at once): the prologue of the consumer loop and
// the tiles that bypass tile_compute.
template <typename TF>
__device__ __forceinline__ void load_fixed(float (&fv)[kTZ], const TF *fp, int nk, unsigned plane)
{
#pragma unroll
    for (int kz = 0; kz < kTZ; ++kz) {
        asm("" : "+l"(fp));
        fv[kz] = kz < nk ? (float)__ldg(fp) : 0.0f;
        fp += plane;
    }
}

// Fixed-image mask of this thread's column in one tile: bit kz set = voxel (i, j, k0 + kz) takes part in the metric.
__device__ __forceinline__ unsigned load_mask_bits(const unsigned char *mp, int nk, unsigned plane)
{
    unsigned bits = 0u;
#pragma unroll
    for (int kz = 0; kz < kTZ; ++kz) {
        asm("" : "+l"(mp));
        if (kz < nk && __ldg(mp) != 0) bits |= 1u << kz;
        mp += plane;
    }
    return bits;
}

// One consumer thread's share of one tile whose box is resident in shared memory: the 8 z planes of
// output voxel column (i0 + lane, j0 + warp), as 4 packed pairs.  BORDER adds the inside test and the
// coordinate clamp (== VTK/ITK index clamp: the clamped coordinate has weight 0 on the tap beyond the edge,
// which TMA zero-filled); interior tiles need neither.
// FX says who owns the fixed-image registers fv[]: kFxLoadHere (loaded at the top from nfp / nk), kFxPrefetchNext
// (valid on entry, each pair refills its two registers with the NEXT tile's values from nfp / nk) or kFxKeep
// (valid on entry, left alone: the multi-matrix kernel reuses them for every matrix of the tile).
enum { kFxLoadHere = 0, kFxPrefetchNext = 1, kFxKeep = 2 };

// MASK: only the voxels whose bit is set in `mbits` (load_mask_bits) count for the metric.
template <typename TM, typename TF, typename TO, int METRIC, bool STORE, bool VTK, bool BORDER, class BOX, int FX, bool MASK = false>
__device__ __forceinline__ void tile_compute(const KParams &p, const unsigned char *stage, const TileInfo &ti,
                                             const ThreadGeom &g, int i0, int j0, int k0, int lane, int warp,
                                             double &acc, unsigned &cnt, unsigned *s_hist, float (&fv)[kTZ],
                                             const TF *nfp, int nk, unsigned mbits = 0xffu)
{
    static_assert(kTZ % 2 == 0, "voxels are processed in z pairs");
    constexpr unsigned ES = sizeof(TM);
    constexpr int PY = BOX::kX * ES, PZ = BOX::kX * BOX::kY * ES;
    const int i = i0 + lane, j = j0 + warp;
    const bool row_ok = (i < p.ox) & (j < p.oy);
    const int kcnt = min(kTZ, p.oz - k0);

    // thread base: 64-bit fixed point add, integer part = box element, top 24 fraction bits -> float32
    const long long tx64 = ti.rx + g.ox, ty64 = ti.ry + g.oy, tz64 = ti.rz + g.oz;
    const int tbx = (int)(tx64 >> 32), tby = (int)(ty64 >> 32), tbz = (int)(tz64 >> 32);
    const float fx = (float)((unsigned)tx64 >> 8) * 5.9604644775390625e-8f;
    const float fy = (float)((unsigned)ty64 >> 8) * 5.9604644775390625e-8f;
    const float fz = (float)((unsigned)tz64 >> 8) * 5.9604644775390625e-8f;
    // element index = ((iz*BY + iy)*BX + ix) with i* = thread base + floor(local); floor(local) arrives as the
    // raw bits of (local + kMagic), i.e. kMagicBits + floor: the constant part is folded into ubase (mod 2^32)
    // (p.magic_fold is a launch parameter on purpose: as a literal, ptxas pulls it back out of ubase and
    // re-adds it in front of every LDS -- 8 extra VIADD per voxel in the float32 kernels)
    const uint32_t ubase = smem_u32(stage) + (uint32_t)((tbz * BOX::kY + tby) * BOX::kX + tbx) * ES + p.magic_fold;

    float ilox = 0.f, ihix = 0.f, iloy = 0.f, ihiy = 0.f, iloz = 0.f, ihiz = 0.f;
    float clox = 0.f, chix = 0.f, cloy = 0.f, chiy = 0.f, cloz = 0.f, chiz = 0.f;
    if (BORDER) {
        const float gx = (float)(ti.bx + tbx), gy = (float)(ti.by + tby), gz = (float)(ti.bz + tbz);
        ilox = -0.5f - gx; iloy = -0.5f - gy; iloz = -0.5f - gz;
        ihix = (float)p.mov.nx - 0.5f - gx; ihiy = (float)p.mov.ny - 0.5f - gy; ihiz = (float)p.mov.nz - 0.5f - gz;
        if (!VTK) { ihix = prev_float(ihix); ihiy = prev_float(ihiy); ihiz = prev_float(ihiz); }
        clox = -gx; cloy = -gy; cloz = -gz;
        chix = (float)(p.mov.nx - 1) - gx; chiy = (float)(p.mov.ny - 1) - gy; chiz = (float)(p.mov.nz - 1) - gz;
    }

    // the host guarantees 8 * oy * ox < 2^31, so plane offsets inside a tile are 32-bit
    const unsigned plane = (unsigned)p.oy * (unsigned)p.ox;
    const size_t gofs = ((size_t)k0 * p.oy + j) * p.ox + i;
    TO *out = static_cast<TO *>(p.out) + gofs;
    const TO bgv = tile_to_out<TO, VTK>(p.background);

    // fv[] holds this thread's 8 fixed-image values of THIS tile.  kFxPrefetchNext: they were fetched while the
    // previous tile was computed, and each pair reloads its two registers with the next tile's values as soon as
    // it has used them, so the global-load latency hides behind a whole tile of arithmetic.  kFxLoadHere: loaded
    // now -- the 72-register int16 kernels spill with the longer live ranges and lose more than the latency costs.
    if (METRIC != kNone && FX == kFxLoadHere) load_fixed<TF>(fv, nfp, nk, plane);
    const f32x2 ax2 = pk(g.a2x, g.a2x), ay2 = pk(g.a2y, g.a2y), az2 = pk(g.a2z, g.a2z);
    const f32x2 fx2 = pk(fx, fx), fy2 = pk(fy, fy), fz2 = pk(fz, fz);
    const f32x2 magic2 = pk(kMagic, kMagic);
    f32x2 part2 = pk(0.0f, 0.0f);
    const uint32_t hrow = 4u * (uint32_t)p.bins;
    const uint32_t hconst = smem_u32(s_hist) - 4u * kMagicBits - hrow * kMagicBits;
    const float binmax = (float)(p.bins - 1);
    const f32x2 fmin2 = pk(p.fmin, p.fmin), fsc2 = pk(p.fscale, p.fscale);
    const f32x2 mmin2 = pk(p.mmin, p.mmin), msc2 = pk(p.mscale, p.mscale);
#pragma unroll
    for (int kp = 0; kp < kTZ; kp += 2) {
        const f32x2 kk = pk((float)kp, (float)(kp + 1));
        f32x2 lx2 = fma2(kk, ax2, fx2), ly2 = fma2(kk, ay2, fy2), lz2 = fma2(kk, az2, fz2);
        bool in0 = true, in1 = true;
        if (BORDER) {
            float x0, x1, y0, y1, z0, z1;
            upk(lx2, x0, x1); upk(ly2, y0, y1); upk(lz2, z0, z1);
            in0 = (x0 >= ilox) & (x0 <= ihix) & (y0 >= iloy) & (y0 <= ihiy) & (z0 >= iloz) & (z0 <= ihiz);
            in1 = (x1 >= ilox) & (x1 <= ihix) & (y1 >= iloy) & (y1 <= ihiy) & (z1 >= iloz) & (z1 <= ihiz);
            lx2 = pk(fminf(fmaxf(x0, clox), chix), fminf(fmaxf(x1, clox), chix));
            ly2 = pk(fminf(fmaxf(y0, cloy), chiy), fminf(fmaxf(y1, cloy), chiy));
            lz2 = pk(fminf(fmaxf(z0, cloz), chiz), fminf(fmaxf(z1, cloz), chiz));
        }
        const f32x2 tx2 = add2_rm(lx2, magic2), ty2 = add2_rm(ly2, magic2), tz2 = add2_rm(lz2, magic2);
        const f32x2 wx2 = sub2(lx2, sub2(tx2, magic2)), wy2 = sub2(ly2, sub2(ty2, magic2)), wz2 = sub2(lz2, sub2(tz2, magic2));
        const uint32_t txa = (uint32_t)tx2, txb = (uint32_t)(tx2 >> 32);
        const uint32_t tya = (uint32_t)ty2, tyb = (uint32_t)(ty2 >> 32);
        const uint32_t tza = (uint32_t)tz2, tzb = (uint32_t)(tz2 >> 32);
        const uint32_t a0 = mad_u32(mad_u32(mad_u32(tza, BOX::kY, tya), BOX::kX, txa), ES, ubase);
        const uint32_t a1 = mad_u32(mad_u32(mad_u32(tzb, BOX::kY, tyb), BOX::kX, txb), ES, ubase);
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
        const f32x2 val2 = lerp2(wz2, vxx0, vxx1);
        float val0, val1;
        upk(val2, val0, val1);
        const bool live0 = row_ok && kp < kcnt && (!MASK || ((mbits >> kp) & 1u));
        const bool live1 = row_ok && kp + 1 < kcnt && (!MASK || ((mbits >> (kp + 1)) & 1u));
        static_assert(!(MASK && STORE), "a masked metric does not store (the mask would punch holes in the output)");
        if (STORE) {
            // one 64-bit output pointer stepped by the plane pitch: 2 address instructions per store, not 5
            // (one mad.wide.u32 per store -- plane bytes x constant plane number + base -- was measured slower:
            // 230.0 against 225.6 us on C2)
            TO o0, o1;
            if (sizeof(TO) == 2 && sizeof(TM) == 2 && VTK) {
                // int16 in, int16 out, floor(v + 0.5): the interpolant of int16 taps stays inside the int16 range
                // (convex weights; the float32 lerp overshoots by an ulp at most), so the saturating F2I (quarter-rate
                // XU, one per voxel) is not needed: floor through the mantissa for the pair, low 16 bits are the result
                const f32x2 r2 = add2_rm(add2(val2, pk(0.5f, 0.5f)), magic2);
                o0 = (TO)(short)(unsigned short)(uint32_t)r2;
                o1 = (TO)(short)(unsigned short)(uint32_t)(r2 >> 32);
            } else {
                o0 = tile_to_out<TO, VTK>(val0);
                o1 = tile_to_out<TO, VTK>(val1);
            }
            asm("" : "+l"(out));
            if (live0) st_out(out, in0 ? o0 : bgv);
            out += plane;
            asm("" : "+l"(out));
            if (live1) st_out(out, in1 ? o1 : bgv);
            out += plane;
        }
        if (METRIC == kSsd) {
            f32x2 d2 = sub2(pk(fv[kp], fv[kp + 1]), val2);
            float d0, d1;
            upk(d2, d0, d1);
            const bool u0 = live0 && in0, u1 = live1 && in1;
            d2 = pk(u0 ? d0 : 0.0f, u1 ? d1 : 0.0f);
            part2 = fma2(d2, d2, part2);
            cnt += (u0 ? 1u : 0u) + (u1 ? 1u : 0u);
        }
        if (METRIC == kHist) {
            // counter address = hist + 4 * (fbin * bins + mbin); the magic bits of both bin words are folded into
            // hconst.  The voxel count is not kept here: it is the sum of the histogram (kernel epilogue).
            const f32x2 fb2 = bin_bits2(pk(fv[kp], fv[kp + 1]), fmin2, fsc2, binmax, magic2);
            const f32x2 mb2 = bin_bits2(val2, mmin2, msc2, binmax, magic2);
            const uint32_t h0 = mad_u32((uint32_t)mb2, 4u, mad_u32((uint32_t)fb2, hrow, hconst));
            const uint32_t h1 = mad_u32((uint32_t)(mb2 >> 32), 4u, mad_u32((uint32_t)(fb2 >> 32), hrow, hconst));
            if (live0 && in0) red_shared_inc(h0);
            if (live1 && in1) red_shared_inc(h1);
        }
        if (METRIC == kMattes) {
            // Parzen-window joint histogram (mattes_add, mia_kernels.cuh); the voxel count is the histogram sum / 1024
            const unsigned short *lut = reinterpret_cast<const unsigned short *>(s_hist + p.bins * p.bins);
            if (live0 && in0) mattes_add(s_hist, lut, fv[kp], val0, p.fmin, p.fscale, p.mmin, p.mscale, p.bins);
            if (live1 && in1) mattes_add(s_hist, lut, fv[kp + 1], val1, p.fmin, p.fscale, p.mmin, p.mscale, p.bins);
        }
        if (METRIC != kNone && FX == kFxPrefetchNext) {
            asm("" : "+l"(nfp));                     // keep ONE stepped 64-bit pointer (2 instructions per load)
            fv[kp] = kp < nk ? (float)__ldg(nfp) : 0.0f;
            nfp += plane;
            asm("" : "+l"(nfp));
            fv[kp + 1] = kp + 1 < nk ? (float)__ldg(nfp) : 0.0f;
            nfp += plane;
        }
    }
    if (METRIC == kSsd) {
        float s0, s1;
        upk(part2, s0, s1);
        acc += (double)(s0 + s1);
    }
}

// Tiles that need no box: entirely outside (background only) or too large for the box (per-voxel global
// gather, the arithmetic of the direct kernel).  Rare; kept simple.
template <typename TM, typename TF, typename TO, int METRIC, bool STORE, bool VTK>
__device__ __noinline__ void tile_compute_nobox(const KParams &p, const double *A, int cls, int i0, int j0, int k0,
                                                int lane, int warp, double &acc, unsigned &cnt, unsigned *s_hist)
{
    const int i = i0 + lane, j = j0 + warp;
    if (i >= p.ox || j >= p.oy) return;
    const int kcnt = min(kTZ, p.oz - k0);
    const TM *mov = static_cast<const TM *>(p.mov.p);
    const size_t plane = (size_t)p.oy * p.ox;
    const size_t gofs = ((size_t)k0 * p.oy + j) * p.ox + i;
    const TF *fixed = static_cast<const TF *>(p.fixed) + gofs;
    TO *out = static_cast<TO *>(p.out) + gofs;
    const TO bgv = tile_to_out<TO, VTK>(p.background);
    float part = 0.0f;
    for (int kz = 0; kz < kcnt; ++kz) {
        float val = 0.0f;
        bool in = false;
        if (cls != kOutside) {
            const RowCtx r = make_row_ctx(A, i, j, k0 + kz, p.mov.nx, p.mov.ny, p.mov.nz);
            in = sample_voxel<TM, VTK>(mov, p.mov.nx, p.mov.ny, p.mov.nz, r, 0, val);
        }
        if (STORE) out[(size_t)kz * plane] = in ? tile_to_out<TO, VTK>(val) : bgv;
        if (METRIC != kNone && in && p.fmask && __ldg(p.fmask + gofs + (size_t)kz * plane) == 0) in = false;
        if (METRIC != kNone && in) {
            const float f = (float)__ldg(fixed + (size_t)kz * plane);
            if (METRIC == kSsd) {
                const float d = f - val;
                part = fmaf(d, d, part);
            } else if (METRIC == kMattes) {
                mattes_add(s_hist, reinterpret_cast<const unsigned short *>(s_hist + p.bins * p.bins), f, val, p.fmin, p.fscale,
                           p.mmin, p.mscale, p.bins);
            } else {
                atomicAdd(&s_hist[bin_of(f, p.fmin, p.fscale, p.bins) * p.bins + bin_of(val, p.mmin, p.mscale, p.bins)], 1u);
            }
            ++cnt;
        }
    }
    if (METRIC == kSsd) acc += (double)part;
}


// Shared-memory carve-up of the tile kernel (dynamic): kStages boxes, 2*kStages mbarriers, kStages
// TileInfo records, then the privatised joint histogram (METRIC == kHist).
template <typename TM, class BOX, int NST>
struct TileSmem {
    static constexpr unsigned kBoxBytes = BOX::kX * BOX::kY * BOX::kZ * sizeof(TM);
    static constexpr unsigned kStageBytes = (kBoxBytes + 127u) & ~127u;
    static constexpr unsigned kBarOff = kStageBytes * NST;
    static constexpr unsigned kInfoOff = kBarOff + 16u * NST;
    static constexpr unsigned kHistOff = (kInfoOff + (unsigned)sizeof(TileInfo) * 64u + 15u) & ~15u;
    // the Parzen table of the Mattes kernels ([257][4] unsigned short) sits right behind the counters
    static size_t bytes(int bins, bool parzen = false)
    {
        return (size_t)kHistOff + sizeof(unsigned) * (size_t)bins * bins + (parzen ? 8u * (kParzenSteps + 1) + 8u : 0u);
    }
};

template <typename TM, typename TF, typename TO, int METRIC, bool STORE, bool VTK, class BOX, int NST, int MINB, bool PREFETCH, bool MASK = false>
__global__ void __launch_bounds__(kTileThreads, MINB)
tile_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ KParams p)
{
    using L = TileSmem<TM, BOX, NST>;
    extern __shared__ __align__(1024) unsigned char tsm[];
    __shared__ double s_red[kTileThreads / 32];
    __shared__ unsigned long long s_redc[kTileThreads / 32];
    __shared__ bool s_last;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int m = blockIdx.y;
    unsigned *s_hist = reinterpret_cast<unsigned *>(tsm + L::kHistOff);
    TileInfo *infos = reinterpret_cast<TileInfo *>(tsm + L::kInfoOff);
    const uint32_t bar0 = smem_u32(tsm + L::kBarOff);       // full[s] = bar0 + 8 s, empty[s] = bar0 + 8 (NST + s)

    double A[12];
    {
        const double *Ap = p.dA ? p.dA + 12 * m : p.A0.a;
#pragma unroll
        for (int q = 0; q < 12; ++q) A[q] = Ap[q];
    }

    if (tid == 0) {
        for (int s = 0; s < NST; ++s) {
            mbar_init(bar0 + 8 * s, 1);
            mbar_init(bar0 + 8 * (NST + s), kConsumerWarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (METRIC == kHist || METRIC == kMattes)
        for (int b = tid; b < p.bins * p.bins; b += kTileThreads) s_hist[b] = 0u;
    if (METRIC == kMattes) parzen_lut_fill(reinterpret_cast<unsigned short *>(s_hist + p.bins * p.bins), tid, kTileThreads);
    __syncthreads();

    double acc = 0.0;
    unsigned cnt = 0;

    // tiles are visited as tile = blockIdx.x + n * gridDim.x; (tix, tiy, tiz) advance incrementally
    const int txy = p.ntx * p.nty;
    const int step_z = (int)(gridDim.x / (unsigned)txy), step_r = (int)(gridDim.x % (unsigned)txy);
    const int step_y = step_r / p.ntx, step_x = step_r % p.ntx;
    int tiz = (int)(blockIdx.x / (unsigned)txy);
    int tiy = (int)(blockIdx.x % (unsigned)txy) / p.ntx, tix = (int)(blockIdx.x % (unsigned)txy) % p.ntx;
    const int ntile_mine = (int)((p.ntiles - 1 - blockIdx.x) / gridDim.x) + 1;
    auto advance = [&]() {
        tix += step_x;
        if (tix >= p.ntx) { tix -= p.ntx; ++tiy; }
        tiy += step_y;
        if (tiy >= p.nty) { tiy -= p.nty; ++tiz; }
        tiz += step_z;
    };

    if (warp == kConsumerWarps) {
        // ---------------- producer warp ----------------
        // Classification is SIMD over the warp: in one pass lane q classifies this CTA's tile number base + q,
        // so the float64 chain is paid once per 32 tiles and never sits on the critical path of a TMA issue.
        // The records go to a 64-entry queue (two passes deep); lane 0 then feeds the stage ring.
        const Reach reach = tile_reach(A);
        unsigned s = 0, ph = 0;
        for (int base = 0; base < ntile_mine; base += 32) {
            const int n = base + lane;
            TileInfo mine;
            mine.rx = mine.ry = mine.rz = 0;
            mine.bx = mine.by = mine.bz = 0;
            mine.cls = kOutside;
            if (n < ntile_mine) {
                const unsigned tile = blockIdx.x + (unsigned)n * gridDim.x;
                const int z = (int)(tile / (unsigned)txy), r2 = (int)(tile % (unsigned)txy);
                mine = classify_tile<BOX>(A, reach, (r2 % p.ntx) * kTX, (r2 / p.ntx) * kTY, z * kTZ, p.mov.nx, p.mov.ny, p.mov.nz,
                                          16 / (int)sizeof(TM));
            }
            infos[(base & 32) + lane] = mine;
            __syncwarp();
            const int cnt_pass = min(32, ntile_mine - base);
            // (An L2 prefetch of the boxes 1 / 2 / 4 tiles ahead of the ring -- cp.async.bulk.prefetch.tensor -- was
            // measured and made every kernel 2-17 % slower: the ring is short of L2->SM throughput, not of DRAM
            // latency, and the prefetch adds L2 lookups.)
            for (int q = 0; q < cnt_pass; ++q) {
                const int cls = __shfl_sync(0xffffffffu, mine.cls, q);
                const int bx = __shfl_sync(0xffffffffu, mine.bx, q);
                const int by = __shfl_sync(0xffffffffu, mine.by, q);
                const int bz = __shfl_sync(0xffffffffu, mine.bz, q);
                if (lane == 0) {
                    mbar_wait<200>(bar0 + 8 * (NST + s), ph ^ 1u);
                    if (cls == kInterior || cls == kBorder) {
                        mbar_arrive_expect_tx(bar0 + 8 * s, L::kBoxBytes);
                        tma_load_3d(smem_u32(tsm + L::kStageBytes * s), &tmap, bar0 + 8 * s, bx, by, bz);
                    } else {
                        mbar_arrive(bar0 + 8 * s);
                    }
                }
                if (++s == NST) { s = 0; ph ^= 1u; }
            }
            __syncwarp();
        }
    } else {
        // ---------------- consumers: 8 warps ----------------
        const ThreadGeom geom = make_thread_geom(A, lane, warp);
        const unsigned plane = (unsigned)p.oy * (unsigned)p.ox;
        // this thread's column of the fixed image in tile (a, b, c): pointer to its first plane, planes to read
        auto fixed_col = [&](int a, int b, int c, int &nk) -> const TF * {
            const int i = a * kTX + lane, j = b * kTY + warp, k0 = c * kTZ;
            nk = (i < p.ox && j < p.oy) ? min(kTZ, p.oz - k0) : 0;
            return static_cast<const TF *>(p.fixed) + (((size_t)k0 * p.oy + j) * p.ox + i);
        };
        float fv[kTZ];
        if (METRIC != kNone && PREFETCH) {
            int nk0 = 0;
            const TF *fp0 = ntile_mine > 0 ? fixed_col(tix, tiy, tiz, nk0) : nullptr;
            load_fixed<TF>(fv, fp0, nk0, plane);
        }
        unsigned s = 0, ph = 0;
        for (int n = 0; n < ntile_mine; ++n) {
            const int i0 = tix * kTX, j0 = tiy * kTY, k0 = tiz * kTZ;
            int nk = 0;
            const TF *nfp = nullptr;
            if (METRIC != kNone && !PREFETCH) nfp = fixed_col(tix, tiy, tiz, nk);
            advance();                                   // (tix, tiy, tiz) is now the NEXT tile
            if (METRIC != kNone && PREFETCH && n + 1 < ntile_mine) nfp = fixed_col(tix, tiy, tiz, nk);
            unsigned mbits = 0xffu;
            if (MASK) {
                const int i = i0 + lane, j = j0 + warp;
                const int mk = (i < p.ox && j < p.oy) ? min(kTZ, p.oz - k0) : 0;
                mbits = load_mask_bits(p.fmask + (((size_t)k0 * p.oy + j) * p.ox + i), mk, plane);
            }
            mbar_wait<32>(bar0 + 8 * s, ph);
            const TileInfo ti = infos[n & 63];
            const unsigned char *stage = tsm + L::kStageBytes * s;
            if (MASK && __all_sync(0xffffffffu, mbits == 0u)) {
                // nothing of this warp's row is inside the mask: no arithmetic, only keep the prefetch registers moving
                if (METRIC != kNone && PREFETCH) load_fixed<TF>(fv, nfp, nk, plane);
            } else if (ti.cls == kInterior) {
                tile_compute<TM, TF, TO, METRIC, STORE, VTK, false, BOX, PREFETCH ? kFxPrefetchNext : kFxLoadHere, MASK>(p, stage, ti, geom, i0, j0, k0, lane, warp, acc, cnt, s_hist,
                                                                         fv, nfp, nk, mbits);
            } else if (ti.cls == kBorder) {
                tile_compute<TM, TF, TO, METRIC, STORE, VTK, true, BOX, PREFETCH ? kFxPrefetchNext : kFxLoadHere, MASK>(p, stage, ti, geom, i0, j0, k0, lane, warp, acc, cnt, s_hist,
                                                                        fv, nfp, nk, mbits);
            } else {
                if (STORE || ti.cls == kNoFit)
                    tile_compute_nobox<TM, TF, TO, METRIC, STORE, VTK>(p, p.dA ? p.dA + 12 * m : p.A0.a, ti.cls, i0, j0, k0, lane, warp,
                                                                       acc, cnt, s_hist);
                if (METRIC != kNone && PREFETCH) load_fixed<TF>(fv, nfp, nk, plane);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar0 + 8 * (NST + s));
            if (++s == NST) { s = 0; ph ^= 1u; }
        }
    }

    if (METRIC == kNone) return;

    // ---- block reduction, fixed order => deterministic for a given grid ----
    if (METRIC == kSsd) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    }
    if (METRIC == kHist) {
        // flush the CTA's histogram; the number of voxels it counted is the sum of its bins
        __syncthreads();
        unsigned *gh = p.d_hist + (size_t)m * p.bins * p.bins;
        cnt = 0;
        for (int b = tid; b < p.bins * p.bins; b += kTileThreads) {
            const unsigned c = s_hist[b];
            if (c) atomicAdd(gh + b, c);
            cnt += c;
        }
    }
    unsigned long long wsum = 0;
    if (METRIC == kMattes) {
        // 64-bit global counters; every sample added kParzenOne in total, so the voxel count is the sum / kParzenOne
        // (taken per block: the host sizes the grid so that a block's sum stays below 2^32 * 2^10)
        __syncthreads();
        unsigned long long *gh = p.d_hist64 + (size_t)m * p.bins * p.bins;
        for (int b = tid; b < p.bins * p.bins; b += kTileThreads) {
            const unsigned c = s_hist[b];
            if (c) atomicAdd(gh + b, (unsigned long long)c);
            wsum += c;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, o);
        cnt = 0;
    }
    cnt = warp_sum_u(cnt);
    if (lane == 0) {
        s_red[warp] = acc;
        s_redc[warp] = METRIC == kMattes ? wsum : (unsigned long long)cnt;
    }
    __syncthreads();
    if (tid == 0) {
        double sacc = 0.0;
        unsigned long long c = 0;
        for (int w = 0; w < kTileThreads / 32; ++w) {
            sacc += s_red[w];
            c += s_redc[w];
        }
        if (METRIC == kMattes) c /= (unsigned long long)kParzenOne;
        p.ws_ssd[(size_t)m * gridDim.x + blockIdx.x] = sacc;
        p.ws_cnt[(size_t)m * gridDim.x + blockIdx.x] = (unsigned)c;
        __threadfence();
        const unsigned t = atomicAdd(p.ws_ticket + m, 1u);
        s_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;

    // last block of this matrix: fixed-order second stage over all block partials
    __threadfence();
    double sacc = 0.0;
    unsigned long long c = 0;
    for (unsigned b = tid; b < gridDim.x; b += kTileThreads) {
        sacc += __ldcg(p.ws_ssd + (size_t)m * gridDim.x + b);
        c += __ldcg(p.ws_cnt + (size_t)m * gridDim.x + b);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sacc += __shfl_xor_sync(0xffffffffu, sacc, o);
        c += __shfl_xor_sync(0xffffffffu, c, o);
    }
    __syncthreads();      // s_red / s_redc are reused
    if (lane == 0) {
        s_red[warp] = sacc;
        s_redc[warp] = c;
    }
    __syncthreads();
    if (tid == 0) {
        double tot = 0.0;
        unsigned long long tc = 0;
        for (int w = 0; w < kTileThreads / 32; ++w) {
            tot += s_red[w];
            tc += s_redc[w];
        }
        if (METRIC == kSsd) p.d_ssd[m] = tot;
        p.d_count[m] = tc;
        p.ws_ticket[m] = 0u;     // self-reset: the workspace is reusable without a memset
    }
}

}  // namespace mia
