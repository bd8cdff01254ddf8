// Multi-matrix SSD kernel for int16 pairs on the x-pair tile (sm_100a): the registration hot loop of configs 1 and 5.
//
// Structure of mia_tile_multi.cuh -- one staged box per tile position serves every probe of an optimiser iteration, the
// producer warp (multi_producer, shared with that file) partitions the probes into boxes, G consumer groups walk the
// matrices against the resident box with the fixed-image values held in registers and per-thread accumulators in shared
// memory -- on the tile and thread mapping of mia_tile_x2.cuh: 64 x 8 x 8 tiles, lane = x-pair, box rows of 128 shorts.
// What that buys over the 32-wide kernel: 33 instead of 43 instructions per voxel and probe, 1.02 instead of 1.27
// shared-memory wavefronts per tap load, half the boxes and barrier hand-overs per voxel.  The ring wait that costs the
// one-matrix x-pair kernel a quarter of its time does not arise here: a box is consumed 13 times before the next one
// is needed.
#pragma once
#include "mia_tile_multi.cuh"
#include "mia_tile_x2.cuh"

namespace mia {

template <class BOX, int NST, int G>
struct MultiX2Smem {
    static constexpr unsigned kBoxBytes = BOX::kX * BOX::kY * BOX::kZ * 2u;
    static constexpr unsigned kStageBytes = (kBoxBytes + 127u) & ~127u;
    static constexpr unsigned kBarOff = kStageBytes * NST;
    static constexpr unsigned kInfoOff = kBarOff + 16u * NST;                                       // TileInfo[NST][kMaxMulti]
    static constexpr unsigned kMatOff = kInfoOff + (unsigned)sizeof(TileInfo) * NST * kMaxMulti;   // double[kMaxMulti][12]
    static constexpr unsigned kHdrOff = kMatOff + 8u * 12u * kMaxMulti;                            // unsigned[NST]
    static constexpr unsigned kAccOff = kHdrOff + 64u;                                             // double[G][kMaxMulti][256]
    static constexpr unsigned kCntOff = kAccOff + 8u * G * kMaxMulti * kMultiLanes;                // unsigned[G][kMaxMulti][256]
    static constexpr unsigned kBytes = kCntOff + 4u * G * kMaxMulti * kMultiLanes;
};

__device__ __forceinline__ float x2_ldg_s16f(const int16_t *p)
{
    int v;
    asm volatile("ld.global.nc.s16 %0, [%1];" : "=r"(v) : "l"(p));
    return (float)v;
}

// SSD of one thread's pair over the planes of one tile against a resident box.  fv2[kz] = the pair's fixed-image values
// (zero where the voxel does not exist).  Interior tiles with all 8 planes accumulate without a predicate: the pair's
// two partial sums are masked once at the end by the edge predicates, the count follows from them.
template <bool VTK, bool BORDER, class BOX>
__device__ __forceinline__ void x2_tile_ssd(const KParams &p, uint32_t stage_addr, const TileInfo &ti, const X2Geom &g, int i0,
                                            int j0, int k0, int lane, int warp, double &acc, unsigned &cnt,
                                            const f32x2 (&fv2)[kXD])
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
    f32x2 part2 = pk(0.0f, 0.0f);
    auto step = [&](f32x2 fvk, float kf, bool live) {
        bool inA, inB;
        const f32x2 val2 = x2_sample<BORDER, BOX>(t, g, kf, inA, inB);
        f32x2 d2 = sub2(fvk, val2);
        if (BORDER) {
            float dA, dB;
            upk(d2, dA, dB);
            const bool uA = okA & inA & live, uB = okB & inB & live;
            d2 = pk(uA ? dA : 0.0f, uB ? dB : 0.0f);
            cnt += (uA ? 1u : 0u) + (uB ? 1u : 0u);
        }
        part2 = fma2(d2, d2, part2);
    };
    if (BORDER || kcnt == kXD) {
        // BORDER masks per voxel anyway, so a clipped last layer costs it nothing extra: planes beyond kcnt are dead
#pragma unroll
        for (int kz = 0; kz < kXD; ++kz) step(fv2[kz], (float)kz, kz < kcnt);
    } else {
        // interior tile of a clipped last layer: planes beyond kcnt carry fixed value 0 but a real sample -- leave them out
        const int16_t *fp = static_cast<const int16_t *>(p.fixed) + (((size_t)k0 * p.oy + j) * p.ox + i);
        const unsigned plane = (unsigned)p.oy * (unsigned)p.ox;
#pragma unroll 1
        for (int kz = 0; kz < kcnt; ++kz) {
            step(pk(okA ? x2_ldg_s16f(fp) : 0.0f, okB ? x2_ldg_s16f(fp + 1) : 0.0f), (float)kz, true);
            fp += plane;
        }
    }
    float sA, sB;
    upk(part2, sA, sB);
    if (BORDER) {
        acc += (double)(sA + sB);
    } else {
        acc += (double)((okA ? sA : 0.0f) + (okB ? sB : 0.0f));
        cnt += (unsigned)kcnt * ((okA ? 1u : 0u) + (okB ? 1u : 0u));
    }
}

// Misfit of the shared box (kNoFit): per-voxel gather for this thread's pair, the arithmetic of the direct kernel.
template <bool VTK>
__device__ __noinline__ void x2_tile_ssd_nobox(const KParams &p, const double *A, int i0, int j0, int k0, int lane, int warp,
                                               double &acc, unsigned &cnt)
{
    const int j = j0 + warp;
    if (j >= p.oy) return;
    const int kcnt = min(kXD, p.oz - k0);
    const int16_t *mov = static_cast<const int16_t *>(p.mov.p);
    const size_t plane = (size_t)p.oy * p.ox;
    float part = 0.0f;
    for (int v = 0; v < 2; ++v) {
        const int i = i0 + 2 * lane + v;
        if (i >= p.ox) break;
        const int16_t *fixed = static_cast<const int16_t *>(p.fixed) + (((size_t)k0 * p.oy + j) * p.ox + i);
        for (int kz = 0; kz < kcnt; ++kz) {
            float val = 0.0f;
            const RowCtx r = make_row_ctx(A, i, j, k0 + kz, p.mov.nx, p.mov.ny, p.mov.nz);
            if (sample_voxel<int16_t, VTK>(mov, p.mov.nx, p.mov.ny, p.mov.nz, r, 0, val)) {
                const float d = x2_ldg_s16f(fixed + (size_t)kz * plane) - val;
                part = fmaf(d, d, part);
                ++cnt;
            }
        }
    }
    acc += (double)part;
}

// One CTA per SM: G groups of 8 consumer warps (warp = row of the 64 x 8 x 8 tile, lane = x-pair) share every staged
// box -- group (g + n) mod G starts at matrix 0 of tile n -- and warp 8 G is the producer.
template <bool VTK, class BOX, int NST, int G>
__global__ void __launch_bounds__(32 * (kX2Consumers * G + 1), 1)
tile_multi_x2_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ KParams p, const int n_mats,
                     const FitGate gate)
{
    if (!gate_pass(gate, p.dA, n_mats)) return;      // the 32-wide kernel of the pair serves this launch (see FitGate)
    using L = MultiX2Smem<BOX, NST, G>;
    extern __shared__ __align__(1024) unsigned char msm[];
    __shared__ bool s_last;

    constexpr int kThreads = 32 * (kX2Consumers * G + 1), kProducer = kX2Consumers * G;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int grp = wid / kX2Consumers, warp = wid % kX2Consumers;
    TileInfo *infos = reinterpret_cast<TileInfo *>(msm + L::kInfoOff);
    double *s_A = reinterpret_cast<double *>(msm + L::kMatOff);
    volatile unsigned *s_hdr = reinterpret_cast<volatile unsigned *>(msm + L::kHdrOff);
    double *s_acc = reinterpret_cast<double *>(msm + L::kAccOff);
    unsigned *s_cnt = reinterpret_cast<unsigned *>(msm + L::kCntOff);
    const uint32_t bar0 = smem_u32(msm + L::kBarOff);       // full[s] = bar0 + 8 s, empty[s] = bar0 + 8 (NST + s)

    for (int q = tid; q < 12 * n_mats; q += kThreads) s_A[q] = p.dA[q];
    for (int q = tid; q < G * kMaxMulti * kMultiLanes; q += kThreads) {
        s_acc[q] = 0.0;
        s_cnt[q] = 0u;
    }
    if (tid == 0) {
        for (int s = 0; s < NST; ++s) {
            mbar_init(bar0 + 8 * s, 1);
            mbar_init(bar0 + 8 * (NST + s), kX2Consumers * G);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int txy = p.ntx * p.nty;
    TileWalk walk;
    walk.step_z = (int)(gridDim.x / (unsigned)txy);
    {
        const int step_r = (int)(gridDim.x % (unsigned)txy);
        walk.step_y = step_r / p.ntx;
        walk.step_x = step_r % p.ntx;
    }
    walk.tiz = (int)(blockIdx.x / (unsigned)txy);
    walk.tiy = (int)(blockIdx.x % (unsigned)txy) / p.ntx;
    walk.tix = (int)(blockIdx.x % (unsigned)txy) % p.ntx;
    walk.ntx = p.ntx;
    walk.nty = p.nty;
    const int ntile_mine = (int)((p.ntiles - 1 - blockIdx.x) / gridDim.x) + 1;

    if (wid == kProducer) {
        multi_producer<int16_t, BOX, NST, kXW, kXH, kXD>(p, &tmap, msm, L::kStageBytes, L::kBoxBytes, bar0, infos, s_hdr, s_A,
                                                         n_mats, lane, walk, ntile_mine);
    } else {
        double *g_acc = s_acc + (size_t)grp * kMaxMulti * kMultiLanes + warp * 32 + lane;
        unsigned *g_cnt = s_cnt + (size_t)grp * kMaxMulti * kMultiLanes + warp * 32 + lane;
        const unsigned plane = (unsigned)p.oy * (unsigned)p.ox;
        unsigned s = 0, ph = 0;
        for (int n = 0; n < ntile_mine; ++n) {
            const int i0 = walk.tix * kXW, j0 = walk.tiy * kXH, k0 = walk.tiz * kXD;
            walk.advance();
            // this thread's pair of the fixed image on the tile's planes: read once per tile position, used by every
            // matrix (requested before the wait on the box)
            f32x2 fv2[kXD];
            {
                const int i = i0 + 2 * lane, j = j0 + warp;
                const int nk = j < p.oy ? min(kXD, p.oz - k0) : 0;
                const bool hA = i < p.ox, hB = i + 1 < p.ox;
                const int16_t *fp = static_cast<const int16_t *>(p.fixed) + (((size_t)k0 * p.oy + j) * p.ox + i);
#pragma unroll
                for (int kz = 0; kz < kXD; ++kz) {
                    asm("" : "+l"(fp));
                    fv2[kz] = pk((hA && kz < nk) ? x2_ldg_s16f(fp) : 0.0f, (hB && kz < nk) ? x2_ldg_s16f(fp + 1) : 0.0f);
                    fp += plane;
                }
            }
            unsigned last;
            do {
                mbar_wait<32>(bar0 + 8 * s, ph);
                last = s_hdr[s];
                const uint32_t stage = smem_u32(msm + L::kStageBytes * s);
                for (int m = (grp + n) % G; m < n_mats; m += G) {
                    const TileInfo ti = infos[s * kMaxMulti + m];
                    if (ti.cls == kOutside) continue;                    // warp-uniform: outside, or served by another stage
                    double acc = 0.0;
                    unsigned cnt = 0;
                    if (ti.cls == kNoFit) {
                        x2_tile_ssd_nobox<VTK>(p, s_A + 12 * m, i0, j0, k0, lane, warp, acc, cnt);
                    } else {
                        const X2Geom geom = x2_make_geom(s_A + 12 * m, lane, warp);
                        int cls = ti.cls;
                        if (cls == kBorder) cls = x2_refine<VTK>(p, ti, geom, i0, j0, lane, warp);
                        if (cls == kInterior) x2_tile_ssd<VTK, false, BOX>(p, stage, ti, geom, i0, j0, k0, lane, warp, acc, cnt, fv2);
                        else if (cls == kBorder) x2_tile_ssd<VTK, true, BOX>(p, stage, ti, geom, i0, j0, k0, lane, warp, acc, cnt, fv2);
                    }
                    g_acc[m * kMultiLanes] += acc;
                    g_cnt[m * kMultiLanes] += cnt;
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(bar0 + 8 * (NST + s));
                if (++s == NST) { s = 0; ph ^= 1u; }
            } while (!last);
        }
    }

    // ---- block partial per matrix (fixed order), then the last block sums the block partials in block order ----
    __syncthreads();
    for (int m = wid; m < n_mats; m += kThreads / 32) {
        double sa = 0.0;
        unsigned c = 0;
        for (int g = 0; g < G; ++g)
            for (int q = lane; q < kMultiLanes; q += 32) {
                sa += s_acc[(g * kMaxMulti + m) * kMultiLanes + q];
                c += s_cnt[(g * kMaxMulti + m) * kMultiLanes + q];
            }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            sa += __shfl_xor_sync(0xffffffffu, sa, o);
            c += __shfl_xor_sync(0xffffffffu, c, o);
        }
        if (lane == 0) {
            p.ws_ssd[(size_t)m * gridDim.x + blockIdx.x] = sa;
            p.ws_cnt[(size_t)m * gridDim.x + blockIdx.x] = c;
            __threadfence();
        }
    }
    __syncthreads();
    if (tid == 0) {
        const unsigned t = atomicAdd(p.ws_ticket, 1u);
        s_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (int m = wid; m < n_mats; m += kThreads / 32) {
        double sa = 0.0;
        unsigned long long c = 0;
        for (unsigned b = lane; b < gridDim.x; b += 32) {
            sa += __ldcg(p.ws_ssd + (size_t)m * gridDim.x + b);
            c += __ldcg(p.ws_cnt + (size_t)m * gridDim.x + b);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            sa += __shfl_xor_sync(0xffffffffu, sa, o);
            c += __shfl_xor_sync(0xffffffffu, c, o);
        }
        if (lane == 0) {
            p.d_ssd[m] = sa;
            p.d_count[m] = c;
        }
    }
    if (tid == 0) p.ws_ticket[0] = 0u;     // self-reset: the workspace is reusable without a memset
}

}  // namespace mia
