// Multi-matrix SSD tile kernel: ONE staged box per output tile serves every candidate matrix of the launch.
//
// The 13 probes of an optimiser iteration (registration.py) differ by a fraction of a voxel to a few voxels,
// so the boxes their tiles need nearly coincide.  tile_kernel (mia_tile.cuh) runs one matrix per CTA
// (blockIdx.y) and therefore pulls every box 13 times through L2 -> SM (the float32 SSD kernel sat on the
// full-chip L2 cap, 12.3 TB/s) and reads the fixed image 13 times.  Here a persistent CTA walks tile
// positions; for each one the producer warp classifies the tile for all matrices at once (lane = matrix),
// chooses one box origin for the lot, issues one TMA load, and the consumers loop over the matrices with the
// fixed-image values held in registers.  Matrices whose tile does not fit the shared box fall back to the
// per-voxel gather for that tile, so unrelated matrices stay correct (just slow).
//
// Boxes per tile position.  The probes of one iteration drift apart with the distance from the rotation centre: on a
// 1024 x 1024 x 1200 lattice of 0.5 mm voxels a +-0.5 degree probe moves the far tiles by +-4.5 voxels, and one box
// with 2 voxels of slack in z then serves only 8.75 of the 13 matrices on average -- the other 4.25 took the per-voxel
// gather at a third of the speed (measured: 283 Gvoxel/s on config 5 against 425 on config 1).  The producer
// therefore partitions the matrices of a tile position into up to kMultiBoxes groups that each share a box, one ring
// stage per group (a header word per stage tells the consumers whether the tile position is finished); with the
// deeper int16 box (MultiBox) 0.2 matrices per tile are left for the gather on that lattice.
//
// Per-voxel arithmetic is tile_compute, unchanged: thread coordinates are taken relative to the box origin in
// Q31.32 fixed point, and an integer change of that origin does not change any fraction, so every voxel value
// is bit-identical to what tile_kernel computes for the same matrix.
#pragma once
#include "mia_tile.cuh"

namespace mia {

constexpr int kMaxMulti = 16;      // matrices per launch on this path (lane = matrix in the producer warp)
#ifndef MIA_MULTI_BOXES
#define MIA_MULTI_BOXES 3
#endif
constexpr int kMultiBoxes = MIA_MULTI_BOXES;     // boxes (ring stages) per tile position at most

// Position of a persistent CTA in the tile lattice: tiles are visited as tile = blockIdx.x + n * gridDim.x and
// (tix, tiy, tiz) advance incrementally.
struct TileWalk {
    int tix, tiy, tiz, step_x, step_y, step_z, ntx, nty;
    __device__ __forceinline__ void advance()
    {
        tix += step_x;
        if (tix >= ntx) { tix -= ntx; ++tiy; }
        tiy += step_y;
        if (tiy >= nty) { tiy -= nty; ++tiz; }
        tiz += step_z;
    }
};

// The producer warp of the multi-matrix kernels (this file and mia_tile_multi_x2.cuh): lane m owns matrix m.  Per tile
// position of TX x TY x TZ voxels it classifies the tile for every matrix, partitions the matrices into up to
// kMultiBoxes groups that share a box, and feeds one ring stage per group: TileInfo records for all matrices, a header
// word (1 = last stage of this tile position), one TMA load.
template <typename TM, class BOX, int NST, int TX, int TY, int TZ>
__device__ __forceinline__ void multi_producer(const KParams &p, const CUtensorMap *tmap, unsigned char *tsm, unsigned stage_bytes,
                                               unsigned box_bytes, uint32_t bar0, TileInfo *infos, volatile unsigned *s_hdr,
                                               const double *s_A, int n_mats, int lane, TileWalk walk, int ntile_mine)
{
    constexpr int xalign = 16 / (int)sizeof(TM);
    const bool have = lane < n_mats;
    double A[12];
#pragma unroll
    for (int q = 0; q < 12; ++q) A[q] = have ? s_A[12 * lane + q] : 0.0;
    Reach reach;
    {
        const double ext[3] = {(double)(TX - 1), (double)(TY - 1), (double)(TZ - 1)};
        for (int a = 0; a < 3; ++a) {
            double n = 0.0, q = 0.0;
            for (int c = 0; c < 3; ++c) {
                const double d = A[a * 4 + c] * ext[c];
                if (d < 0.0) n += d; else q += d;
            }
            reach.neg[a] = n;
            reach.pos[a] = q;
        }
    }
    unsigned s = 0, ph = 0;
    for (int n = 0; n < ntile_mine; ++n) {
        const TileAabb bb = tile_aabb(A, reach, walk.tix * TX, walk.tiy * TY, walk.tiz * TZ, p.mov.nx, p.mov.ny, p.mov.nz);
        const bool active = have && !bb.outside;
        unsigned rem = __ballot_sync(0xffffffffu, active);       // matrices of this tile position still without a box
        int nbox = 0;
        do {
            TileInfo mine;                                       // default: not in this stage (or outside) -> skipped
            mine.rx = mine.ry = mine.rz = 0;
            mine.bx = mine.by = mine.bz = 0;
            mine.cls = kOutside;
            unsigned boxmask = 0;
            int ox = 0, oy = 0, oz = 0;
            if (rem) {
                const bool inrem = (rem >> lane) & 1u;
                // candidate origin 1: the corner of the union of the remaining boxes
                const int big = 0x7fffffff;
                const int ux = __reduce_min_sync(0xffffffffu, inrem ? bb.lo[0] : big) & ~(xalign - 1);
                const int uy = __reduce_min_sync(0xffffffffu, inrem ? bb.lo[1] : big);
                const int uz = __reduce_min_sync(0xffffffffu, inrem ? bb.lo[2] : big);
                const unsigned fu = __ballot_sync(0xffffffffu, inrem && aabb_fits<BOX>(bb, ux, uy, uz));
                // candidate origin 2: the first remaining matrix's own box (the centre probe in a registration loop)
                const int src = __ffs(rem) - 1;
                const int ax = __shfl_sync(0xffffffffu, bb.lo[0], src) & ~(xalign - 1);
                const int ay = __shfl_sync(0xffffffffu, bb.lo[1], src);
                const int az = __shfl_sync(0xffffffffu, bb.lo[2], src);
                const unsigned fa = __ballot_sync(0xffffffffu, inrem && aabb_fits<BOX>(bb, ax, ay, az));
                const bool use_union = __popc(fu) >= __popc(fa);
                ox = use_union ? ux : ax; oy = use_union ? uy : ay; oz = use_union ? uz : az;
                boxmask = use_union ? fu : fa;
                // the last box of the position also carries whatever still does not fit (tile_place marks those
                // kNoFit: per-voxel gather); a matrix whose own tile is larger than the box ends the search
                const bool last_box = boxmask == 0u || nbox + 1 == kMultiBoxes || (rem & ~boxmask) == 0u;
                if (inrem && (last_box || ((boxmask >> lane) & 1u))) mine = tile_place<BOX>(bb, ox, oy, oz);
                rem = last_box ? 0u : (rem & ~boxmask);
            }
            if (lane == 0) mbar_wait<1000>(bar0 + 8 * (NST + s), ph ^ 1u);
            __syncwarp();
            if (lane < kMaxMulti) infos[s * kMaxMulti + lane] = mine;
            if (lane == 0) s_hdr[s] = rem == 0u ? 1u : 0u;
            __syncwarp();
            if (lane == 0) {
                if (boxmask) {
                    mbar_arrive_expect_tx(bar0 + 8 * s, box_bytes);
                    tma_load_3d(smem_u32(tsm + stage_bytes * s), tmap, bar0 + 8 * s, ox, oy, oz);
                } else {
                    mbar_arrive(bar0 + 8 * s);
                }
            }
            if (++s == NST) { s = 0; ph ^= 1u; }
            ++nbox;
        } while (rem);
        walk.advance();
    }
}

// One CTA per SM: G groups of 8 consumer warps share every staged box (group g takes matrices g, g+G, ...), one
// producer warp feeds the ring.  More resident warps than two independent CTAs could hold (their rings would not
// fit), and the shared-memory budget of a whole SM: a deeper ring plus per-thread accumulators, so that no
// cross-lane reduction sits between two tiles.
constexpr int kMultiLanes = kConsumerWarps * 32;          // threads of one consumer group

template <typename TM, class BOX, int NST, int G>
struct MultiSmem {
    static constexpr unsigned kBoxBytes = BOX::kX * BOX::kY * BOX::kZ * sizeof(TM);
    static constexpr unsigned kStageBytes = (kBoxBytes + 127u) & ~127u;
    static constexpr unsigned kBarOff = kStageBytes * NST;
    static constexpr unsigned kInfoOff = kBarOff + 16u * NST;                                  // TileInfo[NST][kMaxMulti]
    static constexpr unsigned kMatOff = kInfoOff + (unsigned)sizeof(TileInfo) * NST * kMaxMulti;   // double[kMaxMulti][12]
    static constexpr unsigned kHdrOff = kMatOff + 8u * 12u * kMaxMulti;                        // unsigned[NST]: 1 = last stage of its tile
    static constexpr unsigned kAccOff = kHdrOff + 64u;                                         // double[G][kMaxMulti][256]
    static constexpr unsigned kCntOff = kAccOff + 8u * G * kMaxMulti * kMultiLanes;            // unsigned[G][kMaxMulti][256]
    static constexpr unsigned kBytes = kCntOff + 4u * G * kMaxMulti * kMultiLanes;
};

template <typename TM, typename TF, bool VTK, class BOX, int NST, int G, bool MASK = false>
__global__ void __launch_bounds__(32 * (kConsumerWarps * G + 1), 1)
tile_multi_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ KParams p, const int n_mats,
                  const FitGate gate)
{
    if (!gate_pass(gate, p.dA, n_mats)) return;      // the other kernel of the pair serves this launch (see FitGate)
    using L = MultiSmem<TM, BOX, NST, G>;
    extern __shared__ __align__(1024) unsigned char tsm[];
    __shared__ bool s_last;

    constexpr int kThreads = 32 * (kConsumerWarps * G + 1), kProducer = kConsumerWarps * G;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int grp = wid / kConsumerWarps, warp = wid % kConsumerWarps;      // consumer group, y row inside the tile
    TileInfo *infos = reinterpret_cast<TileInfo *>(tsm + L::kInfoOff);
    double *s_A = reinterpret_cast<double *>(tsm + L::kMatOff);
    volatile unsigned *s_hdr = reinterpret_cast<volatile unsigned *>(tsm + L::kHdrOff);
    static_assert(NST <= 16, "header words");
    double *s_acc = reinterpret_cast<double *>(tsm + L::kAccOff);
    unsigned *s_cnt = reinterpret_cast<unsigned *>(tsm + L::kCntOff);
    const uint32_t bar0 = smem_u32(tsm + L::kBarOff);       // full[s] = bar0 + 8 s, empty[s] = bar0 + 8 (NST + s)

    for (int q = tid; q < 12 * n_mats; q += kThreads) s_A[q] = p.dA[q];
    for (int q = tid; q < G * kMaxMulti * kMultiLanes; q += kThreads) {
        s_acc[q] = 0.0;
        s_cnt[q] = 0u;
    }
    if (tid == 0) {
        for (int s = 0; s < NST; ++s) {
            mbar_init(bar0 + 8 * s, 1);
            mbar_init(bar0 + 8 * (NST + s), kConsumerWarps * G);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // tile positions are visited as tile = blockIdx.x + n * gridDim.x; (tix, tiy, tiz) advance incrementally
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

    if (wid == kProducer) {
        TileWalk walk = {tix, tiy, tiz, step_x, step_y, step_z, p.ntx, p.nty};
        multi_producer<TM, BOX, NST, kTX, kTY, kTZ>(p, &tmap, tsm, L::kStageBytes, L::kBoxBytes, bar0, infos, s_hdr, s_A, n_mats,
                                                    lane, walk, ntile_mine);
    } else {
        // ---------------- consumers: G groups of 8 warps, every group against the same box ----------------
        // accumulators are private to (group, matrix, thread): the groups drift apart by up to a ring stage and
        // take turns on every matrix (see the loop below), so they must not share a slot
        double *g_acc = s_acc + (size_t)grp * kMaxMulti * kMultiLanes + warp * 32 + lane;
        unsigned *g_cnt = s_cnt + (size_t)grp * kMaxMulti * kMultiLanes + warp * 32 + lane;
        const unsigned plane = (unsigned)p.oy * (unsigned)p.ox;
        unsigned s = 0, ph = 0;
        for (int n = 0; n < ntile_mine; ++n) {
            const int i0 = tix * kTX, j0 = tiy * kTY, k0 = tiz * kTZ;
            advance();
            // this thread's 8 fixed-image values, read once per tile (issued before the wait on the box)
            float fv[kTZ];
            {
                const int i = i0 + lane, j = j0 + warp;
                const int nk = (i < p.ox && j < p.oy) ? min(kTZ, p.oz - k0) : 0;
                const TF *fp = static_cast<const TF *>(p.fixed) + (((size_t)k0 * p.oy + j) * p.ox + i);
                load_fixed<TF>(fv, fp, nk, plane);
            }
            unsigned mbits = 0xffu;
            bool skip = false;
            if (MASK) {                                   // the mask of the tile position, shared by every matrix
                const int i = i0 + lane, j = j0 + warp;
                const int mk = (i < p.ox && j < p.oy) ? min(kTZ, p.oz - k0) : 0;
                mbits = load_mask_bits(p.fmask + (((size_t)k0 * p.oy + j) * p.ox + i), mk, plane);
                skip = __all_sync(0xffffffffu, mbits == 0u);     // nothing of this warp's row is in the mask
            }
            // one ring stage per box of this tile position (usually one); the header word says which is the last
            unsigned last;
            do {
                mbar_wait<32>(bar0 + 8 * s, ph);
                last = s_hdr[s];
                const unsigned char *stage = tsm + L::kStageBytes * s;
                // group (grp + n) mod G starts at matrix 0: with 13 matrices and 2 groups the 7 / 6 split alternates
                // from tile to tile and both groups do the same work over two tiles
                for (int m = (grp + n) % G; m < n_mats && !skip; m += G) {
                    const TileInfo ti = infos[s * kMaxMulti + m];
                    if (ti.cls == kOutside) continue;                    // warp-uniform: outside, or served by another stage
                    double acc = 0.0;
                    unsigned cnt = 0;
                    if (ti.cls == kNoFit) {
                        tile_compute_nobox<TM, TF, TM, kSsd, false, VTK>(p, s_A + 12 * m, ti.cls, i0, j0, k0, lane, warp, acc, cnt, nullptr);
                    } else {
                        const ThreadGeom geom = make_thread_geom(s_A + 12 * m, lane, warp);
                        if (ti.cls == kInterior)
                            tile_compute<TM, TF, TM, kSsd, false, VTK, false, BOX, kFxKeep, MASK>(p, stage, ti, geom, i0, j0, k0, lane, warp, acc, cnt,
                                                                                            nullptr, fv, nullptr, 0, mbits);
                        else
                            tile_compute<TM, TF, TM, kSsd, false, VTK, true, BOX, kFxKeep, MASK>(p, stage, ti, geom, i0, j0, k0, lane, warp, acc, cnt,
                                                                                           nullptr, fv, nullptr, 0, mbits);
                    }
                    // per-thread accumulators in shared memory: nothing between two tiles waits on another lane
                    g_acc[m * kMultiLanes] += acc;
                    g_cnt[m * kMultiLanes] += cnt;
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(bar0 + 8 * (NST + s));
                if (++s == NST) { s = 0; ph ^= 1u; }
            } while (!last);
        }
    }

    // ---- block partial per matrix (fixed order: 8 slots per lane, then a butterfly), then the last block sums
    // the block partials in block order ----
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
