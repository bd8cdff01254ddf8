// Second-generation TMA-staged tile kernel for int16 volumes (sm_100a): 64-wide tiles, one thread = one x-PAIR.
//
// Why (profiles/r01_tile_ncu_summary.txt): the 32 x 8 x 8 kernel of mia_tile.cuh is bound by issue
// slots (43.6 per voxel) and by shared-memory wavefronts (1.29 per LDS: a warp of 32 x-neighbours reads 64 bytes per
// tap, and every row / plane change inside the warp puts two lanes on one bank), and it stages a 96-wide box of
// which a 32-wide tile uses 43 columns (18 bytes of L2->SM traffic per output voxel).  Here
//   * a tile is 64 (x) x 8 (y) x 8 (z); lane l of warp w owns the voxel pair (i0 + 2l, i0 + 2l + 1) of row j0 + w and
//     walks the 8 planes.  The 32 lanes of a tap instruction are then 2 elements = one 32-bit word apart: with a row
//     pitch of 128 shorts (0 mod 32 words) every lane has its own bank wherever the rows and planes change --
//     1.01 wavefronts per LDS on the config-2 matrix (bank simulator, scripts/sim_banks.py);
//   * the x-pair shares every packed f32x2 instruction (coordinates, floors, lerps, rounding) and its two output voxels
//     are adjacent in memory (second store = first address + 2);
//   * the 128 x 16 x 12 box serves 4096 voxels: 12 bytes of L2->SM traffic per voxel, half the TMA issues and
//     barrier hand-overs per voxel;
//   * interior tiles with all 8 planes present run without a single per-voxel predicate: edge columns / rows are
//     masked by two per-tile predicates that guard the stores.
// Scope: the pure resample (Rigid.create_image, structure/rigid.py:685-693; sitk.Resample, structure/deformable.py:579).
// A fused-SSD instantiation of this kernel was built and measured slower than the 32-wide kernels on the fixed-lattice
// workload (296 against 232 us with store, 244 against 207 us without): at 96 registers it has no room to prefetch
// the fixed image, so the metric launches stay on mia_tile.cuh / mia_tile_multi.cuh.
// Arithmetic per voxel is that of mia_tile.cuh (Q31.32 thread base, float32 local coordinates, floor through the
// mantissa, 7 lerps), so the two kernels agree to the last bit on every voxel both can serve.
#pragma once
#include "mia_tile.cuh"

namespace mia {

// (An L2 prefetch of the boxes a few tiles ahead of the ring, cp.async.bulk.prefetch.tensor, was measured: no gain.)
// Registers: two CTAs of 9 warps put 5 warps on one scheduler, whose 16 K registers then allow 102 per thread --
// __launch_bounds__(288, 2) makes ptxas stop at 96 (a cap of 112 through __maxnreg__ left ONE resident CTA: 248 us
// instead of 209 on config 2).
constexpr int kXW = 64, kXH = 8, kXD = 8;            // tile: x, y, z
constexpr int kX2Consumers = kXH;
constexpr int kX2Threads = 32 * (kX2Consumers + 1);
#ifndef MIA_X2_PASS
#define MIA_X2_PASS 4
#endif
constexpr int kX2Pass = MIA_X2_PASS;                    // tiles per claim = records per pass (two passes resident), <= 32.  Config 2 / its
                                                        // z-slab of 8, same box: 16: 198 / 45 us, 8: 193 / 41, 4: 191 / 37, 2: 211 / 37, 1: 226 / 39
constexpr int kX2Skip = 4, kX2Done = 5;                 // record classes beside TileClass: slot past the last tile / no more tiles

__host__ __device__ inline Reach tile_reach_ext(const double *A, int ex, int ey, int ez)
{
    Reach r;
    const double ext[3] = {(double)ex, (double)ey, (double)ez};
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

template <class BOX>
__host__ __device__ inline TileInfo x2_classify(const double *A, const Reach &r, int i0, int j0, int k0, int nx, int ny, int nz)
{
    const TileAabb bb = tile_aabb(A, r, i0, j0, k0, nx, ny, nz);
    return tile_place<BOX>(bb, bb.lo[0] & ~7, bb.lo[1], bb.lo[2]);     // TMA: x origin on 16 bytes
}

// Per-thread constants of a launch: (2 lane * A[:,0] + warp * A[:,1]) in Q31.32, the x and z columns in float32.
struct X2Geom {
    long long ox, oy, oz;
    float a0x, a0y, a0z;      // one voxel along x: the second voxel of the pair
    float a2x, a2y, a2z;      // one plane along z
};

__device__ __forceinline__ X2Geom x2_make_geom(const double *A, int lane, int warp)
{
    X2Geom g;
    g.ox = __double2ll_rn(((double)(2 * lane) * A[0] + (double)warp * A[1]) * 4294967296.0);
    g.oy = __double2ll_rn(((double)(2 * lane) * A[4] + (double)warp * A[5]) * 4294967296.0);
    g.oz = __double2ll_rn(((double)(2 * lane) * A[8] + (double)warp * A[9]) * 4294967296.0);
    g.a0x = (float)A[0]; g.a0y = (float)A[4]; g.a0z = (float)A[8];
    g.a2x = (float)A[2]; g.a2y = (float)A[6]; g.a2z = (float)A[10];
    return g;
}

__device__ __forceinline__ void upk_u(f32x2 v, uint32_t &lo, uint32_t &hi) { asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v)); }

__device__ __forceinline__ float lds_s16f(uint32_t addr, int off)
{
    int v;
    asm volatile("ld.shared.s16 %0, [%1];" : "=r"(v) : "r"(addr + off));
    return (float)v;
}

__device__ __forceinline__ void stg_cs_u16(int16_t *p, uint32_t v)
{
    asm volatile("st.global.cs.u16 [%0], %1;" ::"l"(p), "h"((unsigned short)v) : "memory");
}

// Everything a consumer thread keeps for one tile.
struct X2Thread {
    uint32_t ubase;              // shared address of the thread's base element, magic bits folded (see tile_compute)
    uint32_t safe;               // BORDER: a conflict-free address inside the stage for lanes outside the volume
    f32x2 fx2, fy2, fz2;         // local coordinates of the pair on plane 0
    float ilox, ihix, iloy, ihiy, iloz, ihiz;      // BORDER: inside interval, relative to the thread base
    float clox, chix, cloy, chiy, cloz, chiz;      // BORDER: clamp interval
};

// Trilinear samples of the pair on plane `kf` (plane number as float32).  BORDER adds the inside test and the
// coordinate clamp of tile_compute (mia_tile.cuh); interior tiles need neither.
template <bool BORDER, class BOX>
__device__ __forceinline__ f32x2 x2_sample(const X2Thread &t, const X2Geom &g, float kf, bool &inA, bool &inB)
{
    constexpr int PY = BOX::kX * 2, PZ = BOX::kX * BOX::kY * 2;
    const f32x2 kk = pk(kf, kf), magic2 = pk(kMagic, kMagic);
    f32x2 lx2 = fma2(kk, pk(g.a2x, g.a2x), t.fx2), ly2 = fma2(kk, pk(g.a2y, g.a2y), t.fy2), lz2 = fma2(kk, pk(g.a2z, g.a2z), t.fz2);
    inA = inB = true;
    if (BORDER) {
        float x0, x1, y0, y1, z0, z1;
        upk(lx2, x0, x1); upk(ly2, y0, y1); upk(lz2, z0, z1);
        inA = (x0 >= t.ilox) & (x0 <= t.ihix) & (y0 >= t.iloy) & (y0 <= t.ihiy) & (z0 >= t.iloz) & (z0 <= t.ihiz);
        inB = (x1 >= t.ilox) & (x1 <= t.ihix) & (y1 >= t.iloy) & (y1 <= t.ihiy) & (z1 >= t.iloz) & (z1 <= t.ihiz);
        lx2 = pk(fminf(fmaxf(x0, t.clox), t.chix), fminf(fmaxf(x1, t.clox), t.chix));
        ly2 = pk(fminf(fmaxf(y0, t.cloy), t.chiy), fminf(fmaxf(y1, t.cloy), t.chiy));
        lz2 = pk(fminf(fmaxf(z0, t.cloz), t.chiz), fminf(fmaxf(z1, t.cloz), t.chiz));
    }
    const f32x2 tx2 = add2_rm(lx2, magic2), ty2 = add2_rm(ly2, magic2), tz2 = add2_rm(lz2, magic2);
    const f32x2 wx2 = sub2(lx2, sub2(tx2, magic2)), wy2 = sub2(ly2, sub2(ty2, magic2)), wz2 = sub2(lz2, sub2(tz2, magic2));
    // element = (z * BY + y) * BX + x on the raw bits kMagicBits + floor: the constant part sits in t.ubase
    uint32_t xa, xb, ya, yb, za, zb;
    upk_u(tx2, xa, xb); upk_u(ty2, ya, yb); upk_u(tz2, za, zb);
    uint32_t a0 = mad_u32(mad_u32(mad_u32(za, BOX::kY, ya), BOX::kX, xa), 2u, t.ubase);
    uint32_t a1 = mad_u32(mad_u32(mad_u32(zb, BOX::kY, yb), BOX::kX, xb), 2u, t.ubase);
#ifdef MIA_BOUNDS_CHECK
    // Debug build (compute-sanitizer is closed on this pool): every tap of both voxels must lie inside the staged box.
    // t.safe - 4 lane is the stage base.  A violation traps, which surfaces as a launch failure in the caller.
    {
        const uint32_t lo = t.safe - 4u * (threadIdx.x & 31u), hi = lo + BOX::kX * BOX::kY * BOX::kZ * 2u;
        const bool okA = !BORDER || inA, okB = !BORDER || inB;
        if ((okA && (a0 < lo || a0 + PZ + PY + 4u > hi)) || (okB && (a1 < lo || a1 + PZ + PY + 4u > hi))) __trap();
    }
#endif
    if (BORDER) {
        // A voxel outside the volume is stored as background whatever it loads, but its clamped coordinates pile the
        // outside lanes of a warp onto one x with different rows -- one bank, up to 7 wavefronts per LDS with a row
        // pitch of 0 mod 32 words (ncu: 2.9 on average over the border tiles of config 2).  Such a lane reads the
        // word of its own bank at the start of the stage instead.
        a0 = inA ? a0 : t.safe;
        a1 = inB ? a1 : t.safe;
    }
    const f32x2 v000 = pk(lds_s16f(a0, 0), lds_s16f(a1, 0));
    const f32x2 v100 = pk(lds_s16f(a0, 2), lds_s16f(a1, 2));
    const f32x2 v010 = pk(lds_s16f(a0, PY), lds_s16f(a1, PY));
    const f32x2 v110 = pk(lds_s16f(a0, PY + 2), lds_s16f(a1, PY + 2));
    const f32x2 v001 = pk(lds_s16f(a0, PZ), lds_s16f(a1, PZ));
    const f32x2 v101 = pk(lds_s16f(a0, PZ + 2), lds_s16f(a1, PZ + 2));
    const f32x2 v011 = pk(lds_s16f(a0, PZ + PY), lds_s16f(a1, PZ + PY));
    const f32x2 v111 = pk(lds_s16f(a0, PZ + PY + 2), lds_s16f(a1, PZ + PY + 2));
    const f32x2 vx00 = lerp2(wx2, v000, v100), vx10 = lerp2(wx2, v010, v110);
    const f32x2 vx01 = lerp2(wx2, v001, v101), vx11 = lerp2(wx2, v011, v111);
    const f32x2 vxx0 = lerp2(wy2, vx00, vx10), vxx1 = lerp2(wy2, vx01, vx11);
    return lerp2(wz2, vxx0, vxx1);
}

// float32 pair -> two int16 bit patterns in the low halves of two 32-bit words
template <bool VTK>
__device__ __forceinline__ void x2_to_i16(f32x2 val2, uint32_t &oA, uint32_t &oB)
{
    if (VTK) {
        // floor(v + 0.5) through the mantissa; an interpolant of int16 taps cannot leave the int16 range by more than
        // an ulp (0.004), so no saturation is needed and the low 16 bits are the result
        const f32x2 r2 = add2_rm(add2(val2, pk(0.5f, 0.5f)), pk(kMagic, kMagic));
        oA = (uint32_t)r2;
        oB = (uint32_t)(r2 >> 32);
    } else {
        float a, b;
        upk(val2, a, b);
        oA = (uint32_t)(unsigned short)tile_to_out<int16_t, false>(a);
        oB = (uint32_t)(unsigned short)tile_to_out<int16_t, false>(b);
    }
}

// Refinement of a BORDER tile per consumer warp (= one 64 x 1 x 8 row of the tile).  With 64-wide tiles 30 % of the
// tiles that carry a box touch a face of the moving volume, but most of their rows lie entirely on one side of it:
// the warp votes on the bounding box of each lane's 16 voxels (the map is affine: the extremes are at the corner
// voxels) and takes the interior path, skips the row (background only), or keeps the tested / clamped path.
// Coordinates are relative to the box origin (magnitude < 200: float32 error ~1e-5); the margins make both votes
// conservative.
template <bool VTK>
__device__ __forceinline__ int x2_refine(const KParams &p, const TileInfo &ti, const X2Geom &g, int i0, int j0, int lane, int warp)
{
    const long long tx64 = ti.rx + g.ox, ty64 = ti.ry + g.oy, tz64 = ti.rz + g.oz;
    const float rel[3] = {(float)((double)tx64 * 2.3283064365386963e-10), (float)((double)ty64 * 2.3283064365386963e-10),
                          (float)((double)tz64 * 2.3283064365386963e-10)};
    const float a0[3] = {g.a0x, g.a0y, g.a0z}, a2[3] = {g.a2x, g.a2y, g.a2z};
    const int b[3] = {ti.bx, ti.by, ti.bz}, n[3] = {p.mov.nx, p.mov.ny, p.mov.nz};
    const float eps = 2.0e-3f;
    bool interior = true, outside = false;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float d2 = (float)(kXD - 1) * a2[a];
        const float lo = rel[a] + fminf(a0[a], 0.0f) + fminf(d2, 0.0f), hi = rel[a] + fmaxf(a0[a], 0.0f) + fmaxf(d2, 0.0f);
        const float c0 = (float)(-b[a]), c1 = (float)(n[a] - 1 - b[a]);
        interior &= (lo >= c0 + eps) & (hi <= c1 - eps);
        outside |= (hi < c0 - 0.5f - eps) | (lo > c1 + 0.5f + eps);
    }
    // Lanes beyond the lattice edge store nothing, so they have no say in the outside vote; but they do execute the
    // loads of the interior path, whose unclamped coordinates must stay inside the staged box: they count there.
    const bool dead = (i0 + 2 * lane >= p.ox) | (j0 + warp >= p.oy);
    if (__all_sync(0xffffffffu, outside | dead)) return kOutside;
    if (__all_sync(0xffffffffu, interior)) return kInterior;
    return kBorder;
}

// One consumer thread's share of one tile whose box is resident: the pair (i0 + 2 lane, +1) of row j0 + warp on the
// tile's planes.  Tiles with all 8 planes run fully unrolled without a per-voxel predicate (the two edge predicates
// guard the stores directly); the last layer of a lattice whose depth is not a multiple of 8 takes the rolled loop.
template <bool VTK, bool BORDER, class BOX>
__device__ __forceinline__ void x2_tile(const KParams &p, uint32_t stage_addr, const TileInfo &ti, const X2Geom &g,
                                        int i0, int j0, int k0, int lane, int warp)
{
    const int i = i0 + 2 * lane, j = j0 + warp;
    const bool okA = (i < p.ox) & (j < p.oy), okB = (i + 1 < p.ox) & (j < p.oy);
    const int kcnt = min(kXD, p.oz - k0);

    const long long tx64 = ti.rx + g.ox, ty64 = ti.ry + g.oy, tz64 = ti.rz + g.oz;
    const int tbx = (int)(tx64 >> 32), tby = (int)(ty64 >> 32), tbz = (int)(tz64 >> 32);
    const float fx = (float)((unsigned)tx64 >> 8) * 5.9604644775390625e-8f;
    const float fy = (float)((unsigned)ty64 >> 8) * 5.9604644775390625e-8f;
    const float fz = (float)((unsigned)tz64 >> 8) * 5.9604644775390625e-8f;
    X2Thread t;
    t.ubase = stage_addr + (uint32_t)((tbz * BOX::kY + tby) * BOX::kX + tbx) * 2u + p.magic_fold;
    t.fx2 = pk(fx, fx + g.a0x); t.fy2 = pk(fy, fy + g.a0y); t.fz2 = pk(fz, fz + g.a0z);
    t.safe = stage_addr + 4u * (uint32_t)lane;
    if (BORDER) {
        const float gx = (float)(ti.bx + tbx), gy = (float)(ti.by + tby), gz = (float)(ti.bz + tbz);
        t.ilox = -0.5f - gx; t.iloy = -0.5f - gy; t.iloz = -0.5f - gz;
        t.ihix = (float)p.mov.nx - 0.5f - gx; t.ihiy = (float)p.mov.ny - 0.5f - gy; t.ihiz = (float)p.mov.nz - 0.5f - gz;
        if (!VTK) { t.ihix = prev_float(t.ihix); t.ihiy = prev_float(t.ihiy); t.ihiz = prev_float(t.ihiz); }
        t.clox = -gx; t.cloy = -gy; t.cloz = -gz;
        t.chix = (float)(p.mov.nx - 1) - gx; t.chiy = (float)(p.mov.ny - 1) - gy; t.chiz = (float)(p.mov.nz - 1) - gz;
    }

    const unsigned plane = (unsigned)p.oy * (unsigned)p.ox;
    int16_t *out = static_cast<int16_t *>(p.out) + (((size_t)k0 * p.oy + j) * p.ox + i);
    const uint32_t bg = (uint32_t)(unsigned short)tile_to_out<int16_t, VTK>(p.background);

    auto step = [&](float kf) {
        bool inA, inB;
        const f32x2 val2 = x2_sample<BORDER, BOX>(t, g, kf, inA, inB);
        uint32_t oA, oB;
        x2_to_i16<VTK>(val2, oA, oB);
        if (BORDER) { oA = inA ? oA : bg; oB = inB ? oB : bg; }
        asm("" : "+l"(out));                      // keep ONE stepped 64-bit pointer
        if (okA) stg_cs_u16(out, oA);
        if (okB) stg_cs_u16(out + 1, oB);
        out += plane;
    };
    if (kcnt == kXD) {
#pragma unroll
        for (int kz = 0; kz < kXD; ++kz) step((float)kz);
    } else {
#pragma unroll 1
        for (int kz = 0; kz < kcnt; ++kz) step((float)kz);
    }
}

// Tiles too large for the box (a rotation beyond what the box takes): per-voxel global gather with the arithmetic of
// the direct kernel.  Rare for the matrices this kernel is launched with (host matrices are fit-tested); kept simple.
template <bool VTK>
__device__ __noinline__ void x2_tile_nobox(const KParams &p, const double *A, int i0, int j0, int k0, int lane, int warp)
{
    const int j = j0 + warp;
    if (j >= p.oy) return;
    const int kcnt = min(kXD, p.oz - k0);
    const int16_t *mov = static_cast<const int16_t *>(p.mov.p);
    const size_t plane = (size_t)p.oy * p.ox;
    const int16_t bgv = tile_to_out<int16_t, VTK>(p.background);
    for (int v = 0; v < 2; ++v) {
        const int i = i0 + 2 * lane + v;
        if (i >= p.ox) break;
        int16_t *out = static_cast<int16_t *>(p.out) + (((size_t)k0 * p.oy + j) * p.ox + i);
        for (int kz = 0; kz < kcnt; ++kz) {
            float val = 0.0f;
            const RowCtx r = make_row_ctx(A, i, j, k0 + kz, p.mov.nx, p.mov.ny, p.mov.nz);
            const bool in = sample_voxel<int16_t, VTK>(mov, p.mov.nx, p.mov.ny, p.mov.nz, r, 0, val);
            out[(size_t)kz * plane] = in ? tile_to_out<int16_t, VTK>(val) : bgv;
        }
    }
}

// A row of a tile that lies outside the moving volume: background only.
template <bool VTK>
__device__ __forceinline__ void x2_fill_background(const KParams &p, int i0, int j0, int k0, int lane, int warp)
{
    const int i = i0 + 2 * lane, j = j0 + warp;
    if (j >= p.oy) return;
    const int kcnt = min(kXD, p.oz - k0);
    const unsigned plane = (unsigned)p.oy * (unsigned)p.ox;
    int16_t *out = static_cast<int16_t *>(p.out) + (((size_t)k0 * p.oy + j) * p.ox + i);
    const uint32_t bg = (uint32_t)(unsigned short)tile_to_out<int16_t, VTK>(p.background);
    const bool okA = i < p.ox, okB = i + 1 < p.ox;
    for (int kz = 0; kz < kcnt; ++kz) {
        if (okA) stg_cs_u16(out, bg);
        if (okB) stg_cs_u16(out + 1, bg);
        out += plane;
    }
}

template <class BOX, int NST>
struct X2Smem {
    static constexpr unsigned kBoxBytes = BOX::kX * BOX::kY * BOX::kZ * 2u;
    static constexpr unsigned kStageBytes = (kBoxBytes + 127u) & ~127u;
    static constexpr unsigned kBarOff = kStageBytes * NST;                 // full[NST], empty[NST], info_full[2], info_empty[2]
    static constexpr unsigned kInfoOff = kBarOff + 16u * NST + 32u;
    static constexpr unsigned kCoordOff = (kInfoOff + (unsigned)sizeof(TileInfo) * 2u * kX2Pass + 15u) & ~15u;   // int4 (i0, j0, k0, -) per record
    static constexpr unsigned kBytes = kCoordOff + 16u * 2u * kX2Pass;
};

// Resample-only kernel (Rigid.create_image / sitk.Resample): grid.x persistent CTAs stride over the 64 x 8 x 8 tiles.  Warps 0..7: consumers (warp =
// row of the tile), warp 8: TMA producer (classification SIMD over the warp, one cp.async.bulk.tensor.3d per tile).
template <bool VTK, class BOX, int NST, int MINB>
__global__ void __launch_bounds__(kX2Threads, MINB)
tile_x2_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ KParams p)
{
    using L = X2Smem<BOX, NST>;
    extern __shared__ __align__(1024) unsigned char xsm[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    TileInfo *infos = reinterpret_cast<TileInfo *>(xsm + L::kInfoOff);
    const uint32_t bar0 = smem_u32(xsm + L::kBarOff);       // full[s] = bar0 + 8 s, empty[s] = bar0 + 8 (NST + s)

    double A[12];
    {
        const double *Ap = p.dA ? p.dA : p.A0.a;
#pragma unroll
        for (int q = 0; q < 12; ++q) A[q] = Ap[q];
    }
    // Tile records travel in passes of kX2Pass (two passes resident) behind their own barrier pair, because only tiles that
    // need a box take a ring slot: a tile outside the volume (18.6 % of config 2's tiles, the auto-crop corners) or one
    // too large for the box costs the ring nothing, so the box of the tile AFTER it is requested as soon as a slot is
    // free instead of one (very short) tile ahead (206 -> 203 us on config 2).
    //
    // Tried and dropped: ONE CTA per SM with two groups of 8 consumer warps around one shared 4-stage ring, tiles
    // claimed dynamically (one shared-memory atomic + a named barrier per group and tile): 213 us.  The consumers'
    // quarter of their time on the full barrier is mostly skew between the 8 warps of a tile (interior, border and dead
    // rows in one tile), not late boxes -- L2 -> SM runs at 45 % of its cap -- and a barrier per tile removes the
    // little slack a ring leaves them.
    const uint32_t ibar = bar0 + 16 * NST;                  // info_full[b] = ibar + 8 b, info_empty[b] = ibar + 16 + 8 b
    if (tid == 0) {
        for (int s = 0; s < NST; ++s) {
            mbar_init(bar0 + 8 * s, 1);
            mbar_init(bar0 + 8 * (NST + s), kX2Consumers);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(ibar + 8 * b, 1);
            mbar_init(ibar + 16 + 8 * b, kX2Consumers);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // Tiles are handed out in chunks of kX2Pass consecutive tiles (one pass of records).  With a claim counter
    // (p.ws_ticket: {next tile, CTAs that are done}) a CTA takes the next chunk whenever its producer has room for one,
    // so a CTA that drew cheap tiles (background corners) simply takes more of them; without one the chunks go round-robin.
    // Static round-robin left ~11 % of the SM time idle at the end of config 2 (90 tiles per CTA of cost 0.1 .. 1.4:
    // the slowest of 296 sums sets the time): 201 -> 189 us.
    // (A guided tail -- the last chunk per CTA worth of tiles claimed ONE at a time from a second counter -- was measured
    // on config 2 and on its z-slabs of 2 / 4 / 8 replayed as graphs: no gain, 1-2 us lost with a longer tail zone.  A
    // slab kernel of an 8-GPU job runs 31 us where an eighth of the whole launch would be 24: the difference is the
    // launch itself and the first box, not the order of the tiles.)
    const int txy = p.ntx * p.nty;
    const unsigned ntiles = (unsigned)p.ntiles;
    unsigned *const claim = p.ws_ticket;
    int4 *coords = reinterpret_cast<int4 *>(xsm + L::kCoordOff);

    if (warp == kX2Consumers) {
        // ---------------- producer warp ----------------
        const Reach reach = tile_reach_ext(A, kXW - 1, kXH - 1, kXD - 1);
        unsigned s = 0, ph = 0;
        for (int pass = 0;; ++pass) {
            unsigned base = 0;
            // the first chunk of every CTA is its own (no atomic round trip before the first box can be requested)
            if (lane == 0)
                base = (claim && pass > 0) ? gridDim.x * (unsigned)kX2Pass + atomicAdd(claim, (unsigned)kX2Pass)
                                           : (blockIdx.x + (unsigned)pass * gridDim.x) * (unsigned)kX2Pass;
            base = __shfl_sync(0xffffffffu, base, 0);
            const bool done = base >= ntiles;
            TileInfo mine;
            mine.rx = mine.ry = mine.rz = 0;
            mine.bx = mine.by = mine.bz = 0;
            mine.cls = done ? kX2Done : kX2Skip;
            int4 c = make_int4(0, 0, 0, 0);
            if (lane < kX2Pass && base + (unsigned)lane < ntiles) {
                const unsigned tile = base + (unsigned)lane;
                const int z = (int)(tile / (unsigned)txy), r2 = (int)(tile % (unsigned)txy);
                c = make_int4((r2 % p.ntx) * kXW, (r2 / p.ntx) * kXH, z * kXD, 0);
                mine = x2_classify<BOX>(A, reach, c.x, c.y, c.z, p.mov.nx, p.mov.ny, p.mov.nz);
            }
            if (pass >= 2 && lane == 0) mbar_wait<200>(ibar + 16 + 8 * (pass & 1), (unsigned)(((pass - 2) >> 1) & 1));
            __syncwarp();                                   // the consumers are done with the records of pass - 2
            if (lane < kX2Pass) {
                infos[(pass & 1) * kX2Pass + lane] = mine;
                coords[(pass & 1) * kX2Pass + lane] = c;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(ibar + 8 * (pass & 1));
            if (done) break;
#pragma unroll 1
            for (int q = 0; q < kX2Pass; ++q) {
                const int cls = __shfl_sync(0xffffffffu, mine.cls, q);
                const int bx = __shfl_sync(0xffffffffu, mine.bx, q);
                const int by = __shfl_sync(0xffffffffu, mine.by, q);
                const int bz = __shfl_sync(0xffffffffu, mine.bz, q);
                if (cls != kInterior && cls != kBorder) continue;          // no box, no ring slot
                if (lane == 0) {
                    mbar_wait<200>(bar0 + 8 * (NST + s), ph ^ 1u);
                    mbar_arrive_expect_tx(bar0 + 8 * s, L::kBoxBytes);
                    tma_load_3d(smem_u32(xsm + L::kStageBytes * s), &tmap, bar0 + 8 * s, bx, by, bz);
                }
                if (++s == NST) { s = 0; ph ^= 1u; }
            }
            __syncwarp();
        }
        // This CTA claims nothing more.  The last CTA to say so puts the counter pair back to zero for the next launch
        // that is handed this slot (no CTA touches `next` after its own increment of `done`).
        if (claim && lane == 0) {
            __threadfence();
            if (atomicAdd(claim + 1, 1u) == gridDim.x - 1) {
                claim[0] = 0u;
                claim[1] = 0u;
            }
        }
    } else {
        // ---------------- consumers: 8 warps ----------------
        const X2Geom geom = x2_make_geom(A, lane, warp);
        unsigned s = 0, ph = 0;
        for (int n = 0;; ++n) {
            const int pass = n / kX2Pass, q = n % kX2Pass;
            if (q == 0) mbar_wait<32>(ibar + 8 * (pass & 1), (unsigned)((pass >> 1) & 1));      // records of this pass
            const TileInfo ti = infos[(pass & 1) * kX2Pass + q];
            if (ti.cls == kX2Done) break;
            const int4 c = coords[(pass & 1) * kX2Pass + q];
            const int i0 = c.x, j0 = c.y, k0 = c.z;
            if (ti.cls == kInterior || ti.cls == kBorder) {
                mbar_wait<32>(bar0 + 8 * s, ph);
                const uint32_t stage = smem_u32(xsm + L::kStageBytes * s);
                int cls = ti.cls;
                if (cls == kBorder) cls = x2_refine<VTK>(p, ti, geom, i0, j0, lane, warp);
                if (cls == kInterior) x2_tile<VTK, false, BOX>(p, stage, ti, geom, i0, j0, k0, lane, warp);
                else if (cls == kBorder) x2_tile<VTK, true, BOX>(p, stage, ti, geom, i0, j0, k0, lane, warp);
                else x2_fill_background<VTK>(p, i0, j0, k0, lane, warp);
                __syncwarp();
                if (lane == 0) mbar_arrive(bar0 + 8 * (NST + s));
                if (++s == NST) { s = 0; ph ^= 1u; }
            } else if (ti.cls == kOutside) {
                x2_fill_background<VTK>(p, i0, j0, k0, lane, warp);
            } else if (ti.cls == kNoFit) {
                x2_tile_nobox<VTK>(p, p.dA ? p.dA : p.A0.a, i0, j0, k0, lane, warp);
            }                                               // kX2Skip: past the last tile, nothing to do
            if (q == kX2Pass - 1) {
                __syncwarp();
                if (lane == 0) mbar_arrive(ibar + 16 + 8 * (pass & 1));
            }
        }
    }

}

}  // namespace mia
