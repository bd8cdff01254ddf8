// C-ABI implementation of include/mia_b200.h.  Host dispatch only; kernels live in mia_kernels.cuh.
#include "../../include/mia_b200.h"

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <mutex>
#include <string>
#include <vector>

#include <cstdlib>
#include <map>

#include "mia_kernels.cuh"
#include "mia_optimizer.cuh"
#include "mia_tile.cuh"
#include "mia_tile_multi.cuh"
#include "mia_tile_x2.cuh"
#include "mia_tile_multi_x2.cuh"
#include "mia_dvf.cuh"
#include "mia_dose.cuh"
#include "mia_pyramid.cuh"

using namespace mia;

namespace {

thread_local std::string g_err;

int fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CUDA_TRY(expr)                                                                          \
    do {                                                                                        \
        cudaError_t e_ = (expr);                                                                \
        if (e_ != cudaSuccess) return fail(MIA_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e_)); \
    } while (0)

int sm_count()
{
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (!cached[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

bool dims_ok(const int d[3]) { return d && d[0] > 0 && d[1] > 0 && d[2] > 0; }
bool dtype_ok(int t) { return t == MIA_I16 || t == MIA_F32; }
size_t esize(int t) { return t == MIA_I16 ? 2 : 4; }

constexpr int kMaxGridX = 148 * 16;   // upper bound of persistent grid.x used for workspace sizing

struct WsLayout {
    size_t off_A, off_ssd, off_cnt, off_ticket, total;
};

WsLayout ws_layout(int n_mats)
{
    WsLayout w;
    size_t o = 0;
    w.off_A = o;       o += sizeof(double) * 12 * MIA_MAX_MATS;
    w.off_ssd = o;     o += sizeof(double) * (size_t)n_mats * kMaxGridX;
    w.off_cnt = o;     o += sizeof(unsigned) * (size_t)n_mats * kMaxGridX;
    w.off_ticket = o;  o += sizeof(unsigned) * MIA_MAX_MATS;
    w.total = (o + 255) & ~(size_t)255;
    return w;
}

void fill_common(KParams &p, const void *d_mov, const int mdims[3], const int odims[3])
{
    memset(&p, 0, sizeof p);
    p.mov.p = d_mov;
    p.mov.nx = mdims[0]; p.mov.ny = mdims[1]; p.mov.nz = mdims[2];
    p.ox = odims[0]; p.oy = odims[1]; p.oz = odims[2];
    p.ntx = (p.ox + (kV - 1) + kTileX - 1) / kTileX;   // rows are shifted by up to kV-1 for alignment
    p.nty = (p.oy + kBlockY - 1) / kBlockY;
    p.ntiles = (long long)p.ntx * p.nty * p.oz;
}

template <typename TM, typename TO>
void launch_resample(const KParams &p, int mode, dim3 grid, cudaStream_t s)
{
    dim3 block(kBlockX, kBlockY);
    if (mode == MIA_MODE_VTK)
        resample_kernel<TM, int16_t, TO, kNone, true, true><<<grid, block, 0, s>>>(p);
    else
        resample_kernel<TM, int16_t, TO, kNone, true, false><<<grid, block, 0, s>>>(p);
}

template <typename TM, typename TF, int METRIC, bool STORE>
void launch_metric(const KParams &p, int mode, dim3 grid, size_t smem, cudaStream_t s)
{
    dim3 block(kBlockX, kBlockY);
    if (mode == MIA_MODE_VTK) {
        auto k = resample_kernel<TM, TF, TM, METRIC, STORE, true>;
        if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k<<<grid, block, smem, s>>>(p);
    } else {
        auto k = resample_kernel<TM, TF, TM, METRIC, STORE, false>;
        if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k<<<grid, block, smem, s>>>(p);
    }
}

template <int METRIC, bool STORE>
void dispatch_metric(int mov_dtype, int fixed_dtype, const KParams &p, int mode, dim3 grid, size_t smem,
                     cudaStream_t s)
{
    if (mov_dtype == MIA_I16 && fixed_dtype == MIA_I16)
        launch_metric<int16_t, int16_t, METRIC, STORE>(p, mode, grid, smem, s);
    else if (mov_dtype == MIA_I16 && fixed_dtype == MIA_F32)
        launch_metric<int16_t, float, METRIC, STORE>(p, mode, grid, smem, s);
    else if (mov_dtype == MIA_F32 && fixed_dtype == MIA_I16)
        launch_metric<float, int16_t, METRIC, STORE>(p, mode, grid, smem, s);
    else
        launch_metric<float, float, METRIC, STORE>(p, mode, grid, smem, s);
}


// ------------------------------------------------------------------------------------------------
// Tile kernel dispatch (mia_tile.cuh).  The tile path needs a TMA descriptor of the moving volume:
// 16-byte aligned base and row pitch.  Anything else goes to the direct kernel.
thread_local int g_last_path = 0;    // 0 = direct kernel, 1 = tile kernel (diagnostic, mia_debug_last_path)

// Per kernel family (element type x fused reduction): box shape, ring depth, resident CTAs per SM and whether
// the fixed image is prefetched one tile ahead.  Every entry is a measured choice (profiles/);
// tuning builds override them with -D.
//
// Bank rules behind the box shapes.  float32: 64 words make the row and plane pitch = 0 mod 32 banks, so a warp
// whose 32 x-neighbours run over several rows / planes of the box still hits 32 different banks (a 40-word pitch
// cost 2.4 wavefronts per LDS at config-3 angles).  int16: 96 shorts = 48 words with BY odd make both pitches
// = 16 mod 32; a warp's run covers at most 17 words, so its continuation on the next row / plane lands in the
// other half of the banks whatever the sign of the rotation (1.26 wavefronts per LDS against 1.75 for a 64-short
// pitch, where the two lanes that straddle the row change share a bank half of the time).  The joint-histogram
// kernels carry bins^2 counters beside the ring and use smaller boxes to keep their resident CTA count.
#ifndef MIA_CFG_I16_STORE
#define MIA_CFG_I16_STORE 96, 15, 13, 3, 2, false
#endif
#ifndef MIA_CFG_I16_SSD
#define MIA_CFG_I16_SSD 96, 15, 12, 2, 3, false
#endif
#ifndef MIA_CFG_I16_HIST
#define MIA_CFG_I16_HIST 64, 16, 12, 2, 3, false
#endif
#ifndef MIA_CFG_F32_STORE
#define MIA_CFG_F32_STORE 64, 15, 14, 2, 2, false
#endif
#ifndef MIA_CFG_F32_SSD
#define MIA_CFG_F32_SSD 64, 15, 14, 2, 2, false
#endif
#ifndef MIA_CFG_F32_HIST
#define MIA_CFG_F32_HIST 64, 14, 13, 2, 2, true
#endif
template <int BX, int BY, int BZ, int NST_, int MINB_, bool PREFETCH_>
struct TileCfgT {
    using BOX = Box<BX, BY, BZ>;
    static constexpr int NST = NST_, MINB = MINB_;
    static constexpr bool PREFETCH = PREFETCH_;
};
template <typename TM, int METRIC> struct TileCfg;
template <> struct TileCfg<int16_t, kNone> : TileCfgT<MIA_CFG_I16_STORE> {};   // 3-deep ring, 2 CTAs of 96 registers
template <> struct TileCfg<int16_t, kSsd> : TileCfgT<MIA_CFG_I16_SSD> {};
template <> struct TileCfg<int16_t, kHist> : TileCfgT<MIA_CFG_I16_HIST> {};
template <> struct TileCfg<float, kNone> : TileCfgT<MIA_CFG_F32_STORE> {};
template <> struct TileCfg<float, kSsd> : TileCfgT<MIA_CFG_F32_SSD> {};
template <> struct TileCfg<float, kHist> : TileCfgT<MIA_CFG_F32_HIST> {};
template <> struct TileCfg<int16_t, kMattes> : TileCfgT<MIA_CFG_I16_HIST> {};   // same shared-memory budget as the hard-binned histogram
template <> struct TileCfg<float, kMattes> : TileCfgT<MIA_CFG_F32_HIST> {};

// Mattes counters are 32-bit per CTA in units of 1 / 1024 sample: a CTA may see at most 2^22 voxels (2^21 kept as the
// limit), so large lattices get more blocks than are resident.
long long mattes_min_grid(const KParams &p)
{
    const long long nvox = (long long)p.ox * p.oy * p.oz;
    return (nvox + (1LL << 21) - 1) >> 21;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn()
{
    static EncodeTiledFn fn = [] {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            f = nullptr;
        return reinterpret_cast<EncodeTiledFn>(f);
    }();
    return fn;
}

bool tile_path_disabled()
{
    static const bool off = [] {
        const char *e = getenv("MIA_FORCE_DIRECT");
        return e && e[0] && e[0] != '0';
    }();
    return off;
}

// Builds the 3-D tensor map of the moving volume with the box of BoxFor<TM>.  False => not eligible.
template <typename TM, class BOX>
bool make_tensor_map(CUtensorMap *map, const void *d_mov, const int mdims[3])
{
    if (tile_path_disabled()) return false;
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return false;
    const size_t es = sizeof(TM);
    if (((uintptr_t)d_mov & 15u) != 0 || ((size_t)mdims[0] * es) % 16 != 0) return false;
    const cuuint64_t gdim[3] = {(cuuint64_t)mdims[0], (cuuint64_t)mdims[1], (cuuint64_t)mdims[2]};
    const cuuint64_t gstr[2] = {(cuuint64_t)mdims[0] * es, (cuuint64_t)mdims[0] * mdims[1] * es};
    const cuuint32_t box[3] = {BOX::kX, BOX::kY, BOX::kZ};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUtensorMapDataType dt = es == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    const CUresult r = enc(map, dt, 3, const_cast<void *>(d_mov), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

// Does a typical tile of this host matrix fit the box?  (Device matrices are classified per tile.)
template <typename TM, class BOX>
bool box_fits(const double *A)
{
    const Reach r = tile_reach(A);
    const int B[3] = {BOX::kX, BOX::kY, BOX::kZ};
    for (int a = 0; a < 3; ++a)
        if (std::ceil(r.pos[a] - r.neg[a]) + 3.0 + (a == 0 ? 16.0 / sizeof(TM) - 1.0 : 0.0) > (double)B[a]) return false;
    return true;
}

void fill_tile(KParams &p)
{
    p.ntx = (p.ox + kTX - 1) / kTX;
    p.nty = (p.oy + kTY - 1) / kTY;
    p.ntiles = (long long)p.ntx * p.nty * ((p.oz + kTZ - 1) / kTZ);
}

template <typename TM, typename TF, typename TO, int METRIC, bool STORE, bool VTK, bool MASK = false, class CFG = TileCfg<TM, METRIC>>
int launch_tile(const CUtensorMap &map, const KParams &p, int n_mats, cudaStream_t s)
{
    using BOX = typename CFG::BOX;
    constexpr int NST = CFG::NST, MINB = CFG::MINB;
    constexpr bool PREFETCH = CFG::PREFETCH;
    auto k = tile_kernel<TM, TF, TO, METRIC, STORE, VTK, BOX, NST, MINB, PREFETCH, MASK>;
    const size_t smem = TileSmem<TM, BOX, NST>::bytes((METRIC == kHist || METRIC == kMattes) ? p.bins : 0, METRIC == kMattes);
    static std::mutex mu;
    static std::map<std::pair<int, size_t>, int> occ;     // (device, smem) -> resident CTAs per SM
    static std::map<int, size_t> attr;                    // device -> dynamic shared memory the kernel is allowed
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    int nb = 0;
    {
        std::lock_guard<std::mutex> lock(mu);
        auto it = occ.find({dev, smem});
        if (it == occ.end()) {
            // only ever raise the limit: the histogram kernels come with several sizes (bins), and lowering it for a
            // small one would make the next cached large launch fail
            if (smem > attr[dev]) {
                CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                attr[dev] = smem;
            }
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k, kTileThreads, smem));
            if (nb < 1) return fail(MIA_ERR_UNSUPPORTED, "tile kernel does not fit on this device (%zu B smem)", smem);
            occ[{dev, smem}] = nb;
        } else {
            nb = it->second;
        }
    }
    long long gx = (long long)sm_count() * nb;
    if (METRIC == kMattes && gx < mattes_min_grid(p)) gx = mattes_min_grid(p);
    if (gx > p.ntiles) gx = p.ntiles;
    if (gx > kMaxGridX) gx = kMaxGridX;
    dim3 grid((unsigned)gx, (unsigned)n_mats, 1);
    k<<<grid, kTileThreads, smem, s>>>(map, p);
    CUDA_TRY(cudaGetLastError());
    g_last_path = 1;
    return MIA_OK;
}


// Second box family of the store kernel: rotations in the x-y plane beyond what the standard box takes (about 11 degrees
// for int16).  A 32 x 8 x 8 tile turned by any in-plane angle spans at most 32 rows and 42 columns of the moving
// volume, so a 36-row box serves every angle about z (with a few degrees of tilt); the per-voxel kernel remains for
// large tilts.  This is the interactive view rotation (Image.update_rotation) and Rigid.update_rotation.
template <typename TM> struct TallCfg;
template <> struct TallCfg<int16_t> : TileCfgT<72, 35, 11, 2, 2, false> {};     // 54 KB per stage; 36-word pitch: a warp that runs down a box column (90 degrees) hits 8 banks, not 1
template <> struct TallCfg<float> : TileCfgT<48, 36, 11, 2, 1, false> {};       // 76 KB per stage, one CTA per SM

// Third and fourth family: rotations about y (x and z mix: a 32 x 8 x 8 tile turned by 90 degrees spans 8 columns and 32
// planes) and everything else that is not larger than about 22 voxels along y and z (rotations about x on anisotropic
// lattices, compound rotations).  The z-heavy box keeps an odd row count so that the plane pitch is 4 (int16) or 28
// (float32) words mod 32: a warp that runs DOWN the planes (90 degrees about y) then hits 8 banks, not 1 -- the rule
// the tall family follows for rows.  Any matrix beyond these still takes the per-voxel kernel.
template <typename TM> struct DeepCfg;
template <> struct DeepCfg<int16_t> : TileCfgT<56, 13, 36, 2, 2, false> {};     // 51 KB per stage, plane pitch 364 words = 12 mod 32
template <> struct DeepCfg<float> : TileCfgT<44, 13, 36, 2, 1, false> {};       // 80 KB per stage, plane pitch 572 words = 28 mod 32
template <typename TM> struct CubeCfg;
template <> struct CubeCfg<int16_t> : TileCfgT<48, 25, 25, 2, 1, false> {};     // 59 KB per stage
template <> struct CubeCfg<float> : TileCfgT<40, 25, 25, 2, 1, false> {};       // 98 KB per stage

// One (element type, box) attempt: fit test for host matrices, tensor map, launch.
template <int METRIC, bool STORE, typename TM, class CFG = TileCfg<TM, METRIC>>
int try_tile_box(int mode, const void *d_mov, const int mdims[3], const double *A_host, int n_mats, KParams p,
                 cudaStream_t s, bool &launched)
{
    using BOX = typename CFG::BOX;
    launched = false;
    alignas(64) CUtensorMap map;
    for (int m = 0; A_host && m < n_mats; ++m)
        if (!box_fits<TM, BOX>(A_host + 12 * m)) return MIA_OK;
    if (!make_tensor_map<TM, BOX>(&map, d_mov, mdims)) return MIA_OK;
    fill_tile(p);
    p.magic_fold = 0u - kMagicBits * (unsigned)((BOX::kY + 1) * BOX::kX + 1) * (unsigned)sizeof(TM);
    if (p.ntiles > 0x7fffffffLL) return MIA_OK;
    if ((long long)p.ox * p.oy * kTZ >= 0x7fffffffLL) return MIA_OK;   // in-tile plane offsets are 32-bit
    if (p.fmask) {
        // masked metrics exist without a stored output only (SSD / joint histogram inside a registration loop)
        if constexpr (!STORE && METRIC != kNone) {
            launched = true;
            return mode == MIA_MODE_VTK ? launch_tile<TM, TM, TM, METRIC, false, true, true, CFG>(map, p, n_mats, s)
                                        : launch_tile<TM, TM, TM, METRIC, false, false, true, CFG>(map, p, n_mats, s);
        }
        return MIA_OK;
    }
    launched = true;
    return mode == MIA_MODE_VTK ? launch_tile<TM, TM, TM, METRIC, STORE, true, false, CFG>(map, p, n_mats, s)
                                : launch_tile<TM, TM, TM, METRIC, STORE, false, false, CFG>(map, p, n_mats, s);
}

// Multi-matrix SSD (mia_tile_multi.cuh): one box per tile position for all matrices of the launch, one CTA per SM
// with G consumer groups.  (ring depth, groups) per element type; tuning builds override with -D.  Measured per
// evaluation, 13 probes (int16 512x512x100 / float32 256^3): one matrix per CTA 66.8 / 43.7 us; G = 1: 76.7 / 47.7;
// G = 2: 59.8 / 39.1; G = 3: 64.1 / 39.8 (25 warps put 7 on one scheduler: 72 registers, spills).
// Tried and dropped: the same kernel with 8 privatised joint histograms per CTA (13 probes = launches of 8 + 5):
// 55 us per evaluation against 48 for one matrix per CTA with the fixed image prefetched; and prefetching the next
// tile position's fixed values here (39.0 against 38.1 us).
#ifndef MIA_MULTI_I16
#define MIA_MULTI_I16 3, 2
#endif
#ifndef MIA_MULTI_F32
#define MIA_MULTI_F32 2, 2
#endif
// The multi-matrix kernel has its own int16 box, two planes deeper than the one-matrix SSD box: it has the shared memory
// of a whole SM, and the z slack decides how many probes of a large isotropic lattice share a box (mia_tile_multi.cuh).
#ifndef MIA_MULTI_BOX_I16_Z
#define MIA_MULTI_BOX_I16_Z 14
#endif
#ifndef MIA_MULTI_BOX_I16
#define MIA_MULTI_BOX_I16 96, 15, MIA_MULTI_BOX_I16_Z
#endif
#ifndef MIA_MULTI_BOX_F32
#define MIA_MULTI_BOX_F32 64, 15, 14
#endif
template <typename TM> struct MultiBox;
template <> struct MultiBox<int16_t> { using BOX = Box<MIA_MULTI_BOX_I16>; };
template <> struct MultiBox<float> { using BOX = Box<MIA_MULTI_BOX_F32>; };
template <int NST_, int G_> struct MultiCfgT { static constexpr int NST = NST_, G = G_; };
template <typename TM> struct MultiCfg;
template <> struct MultiCfg<int16_t> : MultiCfgT<MIA_MULTI_I16> {};
template <> struct MultiCfg<float> : MultiCfgT<MIA_MULTI_F32> {};

template <typename TM, bool VTK, bool MASK = false>
int launch_multi(const CUtensorMap &map, const KParams &p, int n_mats, cudaStream_t s, FitGate gate = FitGate{0, 0, 0, 0, 0, 0, 0, 0})
{
    using BOX = typename MultiBox<TM>::BOX;
    constexpr int NST = MultiCfg<TM>::NST, G = MultiCfg<TM>::G;
    constexpr int threads = 32 * (kConsumerWarps * G + 1);
    auto k = tile_multi_kernel<TM, TM, VTK, BOX, NST, G, MASK>;
    const size_t smem = MultiSmem<TM, BOX, NST, G>::kBytes;
    static std::mutex mu;
    static std::map<int, int> occ;                         // device -> resident CTAs per SM
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    int nb = 0;
    {
        std::lock_guard<std::mutex> lock(mu);
        auto it = occ.find(dev);
        if (it == occ.end()) {
            CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k, threads, smem));
            if (nb < 1) return fail(MIA_ERR_UNSUPPORTED, "multi-matrix tile kernel does not fit on this device (%zu B smem)", smem);
            occ[dev] = nb;
        } else {
            nb = it->second;
        }
    }
    long long gx = (long long)sm_count() * nb;
    if (gx > p.ntiles) gx = p.ntiles;
    if (gx > kMaxGridX) gx = kMaxGridX;
    k<<<dim3((unsigned)gx, 1, 1), threads, smem, s>>>(map, p, n_mats, gate);
    CUDA_TRY(cudaGetLastError());
    g_last_path = 2;
    return MIA_OK;
}

bool multi_path_disabled()
{
    static const bool off = [] { const char *e = getenv("MIA_NO_MULTI"); return e && *e && *e != '0'; }();
    return off;
}

template <typename TM>
int try_multi_t(int mode, const void *d_mov, const int mdims[3], const double *A_host, int n_mats, KParams p,
                cudaStream_t s, bool &launched, FitGate gate = FitGate{0, 0, 0, 0, 0, 0, 0, 0})
{
    using BOX = typename MultiBox<TM>::BOX;
    launched = false;
    alignas(64) CUtensorMap map;
    for (int m = 0; A_host && m < n_mats; ++m)
        if (!box_fits<TM, BOX>(A_host + 12 * m)) return MIA_OK;
    if (!make_tensor_map<TM, BOX>(&map, d_mov, mdims)) return MIA_OK;
    fill_tile(p);
    p.magic_fold = 0u - kMagicBits * (unsigned)((BOX::kY + 1) * BOX::kX + 1) * (unsigned)sizeof(TM);
    if (p.ntiles > 0x7fffffffLL) return MIA_OK;
    if ((long long)p.ox * p.oy * kTZ >= 0x7fffffffLL) return MIA_OK;
    launched = true;
    if (p.fmask)
        return mode == MIA_MODE_VTK ? launch_multi<TM, true, true>(map, p, n_mats, s, gate) : launch_multi<TM, false, true>(map, p, n_mats, s, gate);
    return mode == MIA_MODE_VTK ? launch_multi<TM, true>(map, p, n_mats, s, gate) : launch_multi<TM, false>(map, p, n_mats, s, gate);
}

bool x2_path_disabled();      // MIA_NO_X2 (defined with the one-matrix x-pair kernel below)

// x-pair multi-matrix SSD kernel (mia_tile_multi_x2.cuh): box 128 x 16 x 15 (60 KB), 2 ring stages, 2 consumer groups.
// The box budget is what the per-thread accumulators (96 KB) leave of one SM's shared memory: BY * BZ <= 246.  16 x 15
// takes about 5 degrees about z and 4 degrees of tilt on an isotropic lattice, probe spread included; beyond that the
// 32-wide kernel serves the launch.
#ifndef MIA_MX2_BY
#define MIA_MX2_BY 16
#endif
#ifndef MIA_MX2_BZ
#define MIA_MX2_BZ 15
#endif
struct MultiX2Cfg {
    using BOX = Box<128, MIA_MX2_BY, MIA_MX2_BZ>;
    static constexpr int NST = 2, G = 2;
};

template <bool VTK>
int launch_multi_x2(const CUtensorMap &map, const KParams &p, int n_mats, cudaStream_t s, FitGate gate)
{
    using BOX = MultiX2Cfg::BOX;
    constexpr int NST = MultiX2Cfg::NST, G = MultiX2Cfg::G;
    constexpr int threads = 32 * (kX2Consumers * G + 1);
    auto k = tile_multi_x2_kernel<VTK, BOX, NST, G>;
    const size_t smem = MultiX2Smem<BOX, NST, G>::kBytes;
    static std::mutex mu;
    static std::map<int, int> occ;                         // device -> resident CTAs per SM
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    int nb = 0;
    {
        std::lock_guard<std::mutex> lock(mu);
        auto it = occ.find(dev);
        if (it == occ.end()) {
            CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k, threads, smem));
            if (nb < 1) return fail(MIA_ERR_UNSUPPORTED, "x-pair multi-matrix kernel does not fit on this device (%zu B smem)", smem);
            occ[dev] = nb;
        } else {
            nb = it->second;
        }
    }
    long long gx = (long long)sm_count() * nb;
    if (gx > p.ntiles) gx = p.ntiles;
    if (gx > kMaxGridX) gx = kMaxGridX;
    k<<<dim3((unsigned)gx, 1, 1), threads, smem, s>>>(map, p, n_mats, gate);
    CUDA_TRY(cudaGetLastError());
    g_last_path = 6;
    return MIA_OK;
}

int try_multi_x2(int mode, const void *d_mov, const int mdims[3], int n_mats, KParams p, cudaStream_t s, bool &launched,
                 FitGate gate)
{
    using BOX = MultiX2Cfg::BOX;
    launched = false;
    alignas(64) CUtensorMap map;
    if (!make_tensor_map<int16_t, BOX>(&map, d_mov, mdims)) return MIA_OK;
    p.ntx = (p.ox + kXW - 1) / kXW;
    p.nty = (p.oy + kXH - 1) / kXH;
    p.ntiles = (long long)p.ntx * p.nty * ((p.oz + kXD - 1) / kXD);
    p.magic_fold = 0u - kMagicBits * (unsigned)((BOX::kY + 1) * BOX::kX + 1) * 2u;
    if (p.ntiles > 0x7fffffffLL) return MIA_OK;
    if ((long long)p.ox * p.oy * kXD >= 0x7fffffffLL) return MIA_OK;
    launched = true;
    return mode == MIA_MODE_VTK ? launch_multi_x2<true>(map, p, n_mats, s, gate) : launch_multi_x2<false>(map, p, n_mats, s, gate);
}

// SSD of 2..kMaxMulti matrices without a stored output; p.dA is set (metric_common stages host matrices).
int try_multi(int dtype, int mode, const void *d_mov, const int mdims[3], const double *A_host, int n_mats, KParams p,
              cudaStream_t s, bool &launched)
{
    launched = false;
    if (n_mats < 2 || n_mats > kMaxMulti || !p.dA || tile_path_disabled() || multi_path_disabled()) return MIA_OK;
    if (dtype == MIA_F32) return try_multi_t<float>(mode, d_mov, mdims, A_host, n_mats, p, s, launched);
    // int16: the x-pair kernel when every probe fits its box, the 32-wide kernel otherwise.  Host matrices are tested
    // here; device matrices by the kernels themselves (FitGate: both are enqueued, one returns at once).
    using XB = MultiX2Cfg::BOX;
    // Slack 3: measured on a 1024 x 1024 x 304 lattice of 0.5 mm voxels, 13 probes +-0.5 degree / mm: around identity the
    // x-pair kernel runs 560 Gvoxel/s against 489, around (3, -2, 4) degrees 380 against 405 -- across a 64-wide tile the
    // probes drift apart twice as far and need several boxes per tile position -- so it takes the launch only while
    // the centre probe leaves three rows / planes free (about 2 degrees on an isotropic lattice).
    const FitGate gate_x2{1, kXW - 1, kXH - 1, kXD - 1, XB::kX, XB::kY, XB::kZ, 3};
    if (x2_path_disabled() || p.fmask) return try_multi_t<int16_t>(mode, d_mov, mdims, A_host, n_mats, p, s, launched);
    if (A_host) {
        bool fits = true;
        for (int m = 0; m < n_mats; ++m) fits &= gate_fits(A_host + 12 * m, gate_x2);
        if (!fits) return try_multi_t<int16_t>(mode, d_mov, mdims, A_host, n_mats, p, s, launched);
        return try_multi_x2(mode, d_mov, mdims, n_mats, p, s, launched, FitGate{0, 0, 0, 0, 0, 0, 0, 0});
    }
    bool l1 = false, l2 = false;
    int rc = try_multi_x2(mode, d_mov, mdims, n_mats, p, s, l1, gate_x2);
    if (rc != MIA_OK) return rc;
    FitGate other = gate_x2;
    other.mode = l1 ? 2 : 0;
    rc = try_multi_t<int16_t>(mode, d_mov, mdims, A_host, n_mats, p, s, l2, other);
    launched = l1 || l2;
    if (l1) g_last_path = 6;
    return rc;
}

// ------------------------------------------------------------------------------------------------
// x-pair tile kernel for int16 volumes (mia_tile_x2.cuh): 64 x 8 x 8 tiles, box 128 x 16 x 12 (48 KB per stage).
// (box, ring depth, resident CTAs per SM); tuning builds override with -D.
#ifndef MIA_X2_BY
#define MIA_X2_BY 16
#endif
#ifndef MIA_X2_BZ
#define MIA_X2_BZ 12
#endif
#ifndef MIA_X2_NST
#define MIA_X2_NST 2
#endif
#ifndef MIA_X2_MINB
#define MIA_X2_MINB 2
#endif
#define MIA_X2_CFG 128, MIA_X2_BY, MIA_X2_BZ, MIA_X2_NST, MIA_X2_MINB
template <int BX, int BY, int BZ, int NST_, int MINB_>
struct X2CfgT {
    using BOX = Box<BX, BY, BZ>;
    static constexpr int NST = NST_, MINB = MINB_;
};
using X2Cfg = X2CfgT<MIA_X2_CFG>;

bool x2_path_disabled()
{
    static const bool off = [] { const char *e = getenv("MIA_NO_X2"); return e && *e && *e != '0'; }();
    return off;
}

template <class BOX>
bool x2_box_fits(const double *A)
{
    const Reach r = tile_reach_ext(A, kXW - 1, kXH - 1, kXD - 1);
    const int B[3] = {BOX::kX, BOX::kY, BOX::kZ};
    for (int a = 0; a < 3; ++a)
        if (std::ceil(r.pos[a] - r.neg[a]) + 3.0 + (a == 0 ? 7.0 : 0.0) > (double)B[a]) return false;
    return true;
}

// Claim counters of the x-pair kernel's dynamic tile order: {next tile, CTAs done} per launch, zero when idle (the last
// CTA of a launch zeroes its pair).  A launch in flight must own its pair, so
//   * eager launches walk a pool of kX2EagerSlots pairs in launch order: two launches share a pair only when they are
//     kX2EagerSlots launches apart, far beyond what the launch queues of a process hold in flight;
//   * a launch under stream capture runs again with every replay of its graph, at any later time: it takes a pair
//     of its own out of a second pool for good (a replay cannot overlap itself: launches of one executable graph are
//     ordered).  When that pool is used up, captured launches keep the static round-robin order (claim == nullptr).
// MIA_X2_STATIC=1 keeps the static order everywhere (e.g. for two executable instances of ONE captured graph that are
// to run concurrently on two streams).
constexpr unsigned kX2EagerSlots = 1u << 16, kX2GraphSlots = 1u << 16;
std::atomic<int> g_x2_static{[] { const char *e = getenv("MIA_X2_STATIC"); return e && *e && *e != '0' ? 1 : 0; }()};
__device__ unsigned g_x2_claims[kX2EagerSlots + kX2GraphSlots][2];

int x2_claim_slot(cudaStream_t s, unsigned **claim)
{
    *claim = nullptr;
    if (g_x2_static.load()) return MIA_OK;
    static std::mutex mu;
    static unsigned *base[64] = {nullptr};
    static std::atomic<unsigned> seq_eager{0}, seq_graph{0};
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return MIA_OK;
    unsigned *b = nullptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!base[dev]) {
            void *ptr = nullptr;
            CUDA_TRY(cudaGetSymbolAddress(&ptr, g_x2_claims));
            base[dev] = static_cast<unsigned *>(ptr);
        }
        b = base[dev];
    }
    if (!b) return MIA_OK;
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &st) != cudaSuccess) {     // e.g. the legacy stream while another stream captures:
        cudaGetLastError();                                 // not this function's business -- deal the tiles statically
        return MIA_OK;
    }
    if (st == cudaStreamCaptureStatusNone) {
        *claim = b + 2u * (seq_eager.fetch_add(1u) % kX2EagerSlots);
    } else {
        const unsigned slot = seq_graph.fetch_add(1u);
        if (slot < kX2GraphSlots) *claim = b + 2u * (kX2EagerSlots + slot);
        else seq_graph.store(kX2GraphSlots);               // stay saturated instead of wrapping
    }
    return MIA_OK;
}

template <bool VTK>
int launch_x2(const CUtensorMap &map, const KParams &p, cudaStream_t s)
{
    using BOX = X2Cfg::BOX;
    constexpr int NST = X2Cfg::NST, MINB = X2Cfg::MINB;
    auto k = tile_x2_kernel<VTK, BOX, NST, MINB>;
    const size_t smem = X2Smem<BOX, NST>::kBytes;
    static std::mutex mu;
    static std::map<int, int> occ;                         // device -> resident CTAs per SM
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    int nb = 0;
    {
        std::lock_guard<std::mutex> lock(mu);
        auto it = occ.find(dev);
        if (it == occ.end()) {
            CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k, kX2Threads, smem));
            if (nb < 1) return fail(MIA_ERR_UNSUPPORTED, "x-pair tile kernel does not fit on this device (%zu B smem)", smem);
            occ[dev] = nb;
        } else {
            nb = it->second;
        }
    }
    static const int maxb = [] { const char *e = getenv("MIA_X2_MAXB"); return e && *e ? atoi(e) : 0; }();   // tuning knob
    if (maxb > 0 && nb > maxb) nb = maxb;
    const long long nchunks = (p.ntiles + kX2Pass - 1) / kX2Pass;
    long long gx = (long long)sm_count() * nb;
    if (gx > nchunks) gx = nchunks;
    if (gx > kMaxGridX) gx = kMaxGridX;
    KParams q = p;
    const int rc = x2_claim_slot(s, &q.ws_ticket);          // nullptr: chunks go round-robin
    if (rc) return rc;
    k<<<dim3((unsigned)gx, 1, 1), kX2Threads, smem, s>>>(map, q);
    CUDA_TRY(cudaGetLastError());
    g_last_path = 5;
    return MIA_OK;
}

// Pure resample, int16 in and out.  The host matrix is fit-tested against the box; anything else takes the older paths.
int try_x2(int mode, const void *d_mov, const int mdims[3], const double *A_host, KParams p, cudaStream_t s, bool &launched)
{
    using BOX = X2Cfg::BOX;
    launched = false;
    if (tile_path_disabled() || x2_path_disabled() || !A_host || !x2_box_fits<BOX>(A_host)) return MIA_OK;
    alignas(64) CUtensorMap map;
    if (!make_tensor_map<int16_t, BOX>(&map, d_mov, mdims)) return MIA_OK;
    p.ntx = (p.ox + kXW - 1) / kXW;
    p.nty = (p.oy + kXH - 1) / kXH;
    p.ntiles = (long long)p.ntx * p.nty * ((p.oz + kXD - 1) / kXD);
    p.magic_fold = 0u - kMagicBits * (unsigned)((BOX::kY + 1) * BOX::kX + 1) * 2u;
    if (p.ntiles > 0x7fffffffLL) return MIA_OK;
    if ((long long)p.ox * p.oy * kXD >= 0x7fffffffLL) return MIA_OK;   // in-tile plane offsets are 32-bit
    launched = true;
    return mode == MIA_MODE_VTK ? launch_x2<true>(map, p, s) : launch_x2<false>(map, p, s);
}

// Try the tile path for (dtype, mode); `launched` tells whether it took the call.  A_host may be null
// (device matrices: the kernel classifies every tile itself and falls back per tile).
template <int METRIC, bool STORE>
int try_tile(int dtype, int mode, const void *d_mov, const int mdims[3], const double *A_host, int n_mats,
             KParams p, cudaStream_t s, bool &launched)
{
    launched = false;
    if (dtype == MIA_F32) return try_tile_box<METRIC, STORE, float>(mode, d_mov, mdims, A_host, n_mats, p, s, launched);
    return try_tile_box<METRIC, STORE, int16_t>(mode, d_mov, mdims, A_host, n_mats, p, s, launched);
}

}  // namespace

extern "C" {

int mia_version(void) { return MIA_B200_VERSION; }

const char *mia_last_error(void) { return g_err.c_str(); }

int mia_debug_last_path(void) { return g_last_path; }

int mia_debug_x2_static(int on)
{
    return g_x2_static.exchange(on ? 1 : 0);
}

int mia_resample_rigid(const void *d_mov, int mov_dtype, const int mdims[3], void *d_out, int out_dtype,
                       const int odims[3], const double A[12], double background, int mode, void *stream)
{
    if (!d_mov || !d_out || !A) return fail(MIA_ERR_ARG, "mia_resample_rigid: null pointer");
    if (!dims_ok(mdims) || !dims_ok(odims)) return fail(MIA_ERR_ARG, "mia_resample_rigid: dims must be positive");
    if (!dtype_ok(mov_dtype) || !dtype_ok(out_dtype))
        return fail(MIA_ERR_DTYPE, "mia_resample_rigid: dtype must be MIA_I16 or MIA_F32");
    if (mode != MIA_MODE_VTK && mode != MIA_MODE_ITK) return fail(MIA_ERR_ARG, "mia_resample_rigid: bad mode");
    for (int i = 0; i < 12; ++i)
        if (!std::isfinite(A[i])) return fail(MIA_ERR_ARG, "mia_resample_rigid: non-finite matrix");
    KParams p;
    fill_common(p, d_mov, mdims, odims);
    p.out = d_out;
    memcpy(p.A0.a, A, sizeof(double) * 12);
    p.background = (float)background;
    p.round_mode = mode == MIA_MODE_VTK ? 0 : 1;
    if (p.ntiles > 0x7fffffffLL) return fail(MIA_ERR_UNSUPPORTED, "mia_resample_rigid: output too large");
    dim3 grid((unsigned)p.ntiles, 1, 1);
    cudaStream_t s = (cudaStream_t)stream;
    g_last_path = 0;
    if (mov_dtype == out_dtype) {
        bool launched = false;
        int rc = MIA_OK;
        if (mov_dtype == MIA_I16) {
            rc = try_x2(mode, d_mov, mdims, A, p, s, launched);
            if (launched) return rc;
        }
        rc = try_tile<kNone, true>(mov_dtype, mode, d_mov, mdims, A, 1, p, s, launched);
        if (launched) return rc;
        rc = mov_dtype == MIA_F32 ? try_tile_box<kNone, true, float, TallCfg<float>>(mode, d_mov, mdims, A, 1, p, s, launched)
                                  : try_tile_box<kNone, true, int16_t, TallCfg<int16_t>>(mode, d_mov, mdims, A, 1, p, s, launched);
        if (launched) {
            g_last_path = 4;
            return rc;
        }
        rc = mov_dtype == MIA_F32 ? try_tile_box<kNone, true, float, DeepCfg<float>>(mode, d_mov, mdims, A, 1, p, s, launched)
                                  : try_tile_box<kNone, true, int16_t, DeepCfg<int16_t>>(mode, d_mov, mdims, A, 1, p, s, launched);
        if (launched) {
            g_last_path = 7;
            return rc;
        }
        rc = mov_dtype == MIA_F32 ? try_tile_box<kNone, true, float, CubeCfg<float>>(mode, d_mov, mdims, A, 1, p, s, launched)
                                  : try_tile_box<kNone, true, int16_t, CubeCfg<int16_t>>(mode, d_mov, mdims, A, 1, p, s, launched);
        if (launched) {
            g_last_path = 8;
            return rc;
        }
    }
    if (mov_dtype == MIA_I16 && out_dtype == MIA_I16) launch_resample<int16_t, int16_t>(p, mode, grid, s);
    else if (mov_dtype == MIA_I16 && out_dtype == MIA_F32) launch_resample<int16_t, float>(p, mode, grid, s);
    else if (mov_dtype == MIA_F32 && out_dtype == MIA_I16) launch_resample<float, int16_t>(p, mode, grid, s);
    else launch_resample<float, float>(p, mode, grid, s);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

size_t mia_metric_workspace_bytes(const int odims[3], int n_mats, int bins)
{
    (void)odims;
    (void)bins;
    if (n_mats < 1) n_mats = 1;
    if (n_mats > MIA_MAX_MATS) n_mats = MIA_MAX_MATS;
    return ws_layout(n_mats).total;
}

static int metric_common(KParams &p, const char *who, const void *d_mov, int mov_dtype, const int mdims[3],
                         const void *d_fixed, int fixed_dtype, const int odims[3], int n_mats,
                         const double *A, const double *d_A, int mode, void *d_ws, size_t ws_bytes,
                         cudaStream_t s, WsLayout &w)
{
    if (!d_mov || !d_fixed || !d_ws) return fail(MIA_ERR_ARG, "%s: null pointer", who);
    if (!A && !d_A) return fail(MIA_ERR_ARG, "%s: need A (host) or d_A (device)", who);
    if (!dims_ok(mdims) || !dims_ok(odims)) return fail(MIA_ERR_ARG, "%s: dims must be positive", who);
    if (!dtype_ok(mov_dtype) || !dtype_ok(fixed_dtype)) return fail(MIA_ERR_DTYPE, "%s: bad dtype", who);
    if (n_mats < 1 || n_mats > MIA_MAX_MATS) return fail(MIA_ERR_ARG, "%s: n_mats must be 1..%d", who, MIA_MAX_MATS);
    if (mode != MIA_MODE_VTK && mode != MIA_MODE_ITK) return fail(MIA_ERR_ARG, "%s: bad mode", who);
    w = ws_layout(n_mats);
    if (ws_bytes < w.total) return fail(MIA_ERR_WORKSPACE, "%s: workspace %zu < %zu bytes", who, ws_bytes, w.total);
    if (((uintptr_t)d_ws & 15) != 0) return fail(MIA_ERR_WORKSPACE, "%s: workspace must be 16-byte aligned", who);
    fill_common(p, d_mov, mdims, odims);
    p.fixed = d_fixed;
    char *ws = static_cast<char *>(d_ws);
    p.ws_ssd = reinterpret_cast<double *>(ws + w.off_ssd);
    p.ws_cnt = reinterpret_cast<unsigned *>(ws + w.off_cnt);
    p.ws_ticket = reinterpret_cast<unsigned *>(ws + w.off_ticket);
    p.round_mode = mode == MIA_MODE_VTK ? 0 : 1;
    if (d_A) {
        p.dA = d_A;
    } else if (n_mats == 1) {
        for (int i = 0; i < 12; ++i)
            if (!std::isfinite(A[i])) return fail(MIA_ERR_ARG, "%s: non-finite matrix", who);
        memcpy(p.A0.a, A, sizeof(double) * 12);
        p.dA = nullptr;
    } else {
        for (int i = 0; i < 12 * n_mats; ++i)
            if (!std::isfinite(A[i])) return fail(MIA_ERR_ARG, "%s: non-finite matrix", who);
        CUDA_TRY(cudaMemcpyAsync(ws + w.off_A, A, sizeof(double) * 12 * n_mats, cudaMemcpyHostToDevice, s));
        p.dA = reinterpret_cast<const double *>(ws + w.off_A);
    }
    CUDA_TRY(cudaMemsetAsync(p.ws_ticket, 0, sizeof(unsigned) * MIA_MAX_MATS, s));
    return MIA_OK;
}

int mia_resample_ssd(const void *d_mov, int mov_dtype, const int mdims[3], const void *d_fixed,
                     int fixed_dtype, const int odims[3], void *d_out, int out_dtype, int n_mats,
                     const double *A, const double *d_A, double background, int mode, double *d_ssd,
                     unsigned long long *d_count, void *d_ws, size_t ws_bytes, void *stream)
{
    return mia_resample_ssd_masked(d_mov, mov_dtype, mdims, d_fixed, fixed_dtype, nullptr, nullptr, odims, d_out, out_dtype, n_mats, A,
                                   d_A, background, mode, d_ssd, d_count, d_ws, ws_bytes, stream);
}

int mia_resample_ssd_masked(const void *d_mov, int mov_dtype, const int mdims[3], const void *d_fixed,
                            int fixed_dtype, const unsigned char *d_fixed_mask, const unsigned char *d_moving_mask,
                            const int odims[3], void *d_out, int out_dtype, int n_mats, const double *A, const double *d_A,
                            double background, int mode, double *d_ssd, unsigned long long *d_count, void *d_ws,
                            size_t ws_bytes, void *stream)
{
    cudaStream_t s = (cudaStream_t)stream;
    if (d_out && (d_fixed_mask || d_moving_mask))
        return fail(MIA_ERR_UNSUPPORTED, "mia_resample_ssd_masked: a masked metric does not store");
    if (!d_ssd || !d_count) return fail(MIA_ERR_ARG, "mia_resample_ssd: null result pointer");
    if (d_out && n_mats != 1) return fail(MIA_ERR_ARG, "mia_resample_ssd: storing needs n_mats == 1");
    if (d_out && out_dtype != mov_dtype)
        return fail(MIA_ERR_UNSUPPORTED, "mia_resample_ssd: stored output must have the moving dtype");
    KParams p;
    WsLayout w;
    int rc = metric_common(p, "mia_resample_ssd", d_mov, mov_dtype, mdims, d_fixed, fixed_dtype, odims, n_mats,
                           A, d_A, mode, d_ws, ws_bytes, s, w);
    if (rc) return rc;
    p.fmask = d_fixed_mask;
    p.mmask = d_moving_mask;
    p.out = d_out;
    p.background = (float)background;
    p.d_ssd = d_ssd;
    p.d_count = d_count;
    long long gx = (long long)sm_count() * 8;
    if (gx > p.ntiles) gx = p.ntiles;
    if (gx > kMaxGridX) gx = kMaxGridX;
    dim3 grid((unsigned)gx, (unsigned)n_mats, 1);
    g_last_path = 0;
    if (mov_dtype == fixed_dtype && !p.mmask) {          // a moving mask runs on the per-voxel kernel
        bool launched = false;
        if (!d_out) {
            const int mrc = try_multi(mov_dtype, mode, d_mov, mdims, d_A ? nullptr : A, n_mats, p, s, launched);
            if (launched) return mrc;
        }
        const int trc = d_out ? try_tile<kSsd, true>(mov_dtype, mode, d_mov, mdims, d_A ? nullptr : A, n_mats, p, s, launched)
                              : try_tile<kSsd, false>(mov_dtype, mode, d_mov, mdims, d_A ? nullptr : A, n_mats, p, s, launched);
        if (launched) return trc;
    }
    if (d_out) dispatch_metric<kSsd, true>(mov_dtype, fixed_dtype, p, mode, grid, 0, s);
    else dispatch_metric<kSsd, false>(mov_dtype, fixed_dtype, p, mode, grid, 0, s);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

int mia_resample_jointhist(const void *d_mov, int mov_dtype, const int mdims[3], const void *d_fixed,
                           int fixed_dtype, const int odims[3], int n_mats, const double *A,
                           const double *d_A, int mode, float fmin, float fmax, float mmin, float mmax,
                           int bins, unsigned int *d_hist, unsigned long long *d_count, void *d_ws,
                           size_t ws_bytes, void *stream)
{
    return mia_resample_jointhist_masked(d_mov, mov_dtype, mdims, d_fixed, fixed_dtype, nullptr, nullptr, odims, n_mats, A, d_A, mode,
                                         fmin, fmax, mmin, mmax, bins, d_hist, d_count, d_ws, ws_bytes, stream);
}

int mia_resample_jointhist_masked(const void *d_mov, int mov_dtype, const int mdims[3], const void *d_fixed,
                                  int fixed_dtype, const unsigned char *d_fixed_mask, const unsigned char *d_moving_mask,
                                  const int odims[3], int n_mats, const double *A, const double *d_A, int mode, float fmin,
                                  float fmax, float mmin, float mmax, int bins, unsigned int *d_hist,
                                  unsigned long long *d_count, void *d_ws, size_t ws_bytes, void *stream)
{
    cudaStream_t s = (cudaStream_t)stream;
    if (!d_hist || !d_count) return fail(MIA_ERR_ARG, "mia_resample_jointhist: null result pointer");
    if (bins < 2 || bins > MIA_MAX_BINS) return fail(MIA_ERR_ARG, "mia_resample_jointhist: bins must be 2..%d", MIA_MAX_BINS);
    if (!(fmax > fmin) || !(mmax > mmin)) return fail(MIA_ERR_ARG, "mia_resample_jointhist: empty intensity range");
    KParams p;
    WsLayout w;
    int rc = metric_common(p, "mia_resample_jointhist", d_mov, mov_dtype, mdims, d_fixed, fixed_dtype, odims,
                           n_mats, A, d_A, mode, d_ws, ws_bytes, s, w);
    if (rc) return rc;
    p.fmask = d_fixed_mask;
    p.mmask = d_moving_mask;
    p.d_hist = d_hist;
    p.d_count = d_count;
    p.bins = bins;
    p.fmin = fmin; p.fscale = (float)bins / (fmax - fmin);
    p.mmin = mmin; p.mscale = (float)bins / (mmax - mmin);
    CUDA_TRY(cudaMemsetAsync(d_hist, 0, sizeof(unsigned) * (size_t)n_mats * bins * bins, s));
    const size_t smem = sizeof(unsigned) * (size_t)bins * bins;
    long long gx = (long long)sm_count() * (smem > 32 * 1024 ? 2 : 4);
    if (gx > p.ntiles) gx = p.ntiles;
    if (gx > kMaxGridX) gx = kMaxGridX;
    dim3 grid((unsigned)gx, (unsigned)n_mats, 1);
    g_last_path = 0;
    if (mov_dtype == fixed_dtype && !p.mmask) {
        bool launched = false;
        const int trc = try_tile<kHist, false>(mov_dtype, mode, d_mov, mdims, d_A ? nullptr : A, n_mats, p, s, launched);
        if (launched) return trc;
    }
    dispatch_metric<kHist, false>(mov_dtype, fixed_dtype, p, mode, grid, smem, s);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

int mia_mi_from_hist(const unsigned int *d_hist, int bins, int n_mats, double *d_mi, void *stream)
{
    if (!d_hist || !d_mi) return fail(MIA_ERR_ARG, "mia_mi_from_hist: null pointer");
    if (bins < 2 || bins > MIA_MAX_BINS || n_mats < 1) return fail(MIA_ERR_ARG, "mia_mi_from_hist: bad bins / n_mats");
    const size_t smem = sizeof(double) * (2 * (size_t)bins + 256);
    mi_kernel<unsigned><<<n_mats, 256, smem, (cudaStream_t)stream>>>(d_hist, bins, d_mi);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

int mia_mi_from_hist64(const unsigned long long *d_hist, int bins, int n_mats, double *d_mi, void *stream)
{
    if (!d_hist || !d_mi) return fail(MIA_ERR_ARG, "mia_mi_from_hist64: null pointer");
    if (bins < 2 || bins > MIA_MAX_BINS || n_mats < 1) return fail(MIA_ERR_ARG, "mia_mi_from_hist64: bad bins / n_mats");
    const size_t smem = sizeof(double) * (2 * (size_t)bins + 256);
    mi_kernel<unsigned long long><<<n_mats, 256, smem, (cudaStream_t)stream>>>(d_hist, bins, d_mi);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

int mia_resample_mattes(const void *d_mov, int mov_dtype, const int mdims[3], const void *d_fixed, int fixed_dtype,
                        const unsigned char *d_fixed_mask, const unsigned char *d_moving_mask, const int odims[3],
                        int n_mats, const double *A, const double *d_A, int mode, float fmin, float fmax, float mmin,
                        float mmax, int bins, unsigned long long *d_hist, unsigned long long *d_count, void *d_ws,
                        size_t ws_bytes, void *stream)
{
    cudaStream_t s = (cudaStream_t)stream;
    if (!d_hist || !d_count) return fail(MIA_ERR_ARG, "mia_resample_mattes: null result pointer");
    if (bins < 8 || bins > MIA_MAX_BINS) return fail(MIA_ERR_ARG, "mia_resample_mattes: bins must be 8..%d (4 of them are padding)", MIA_MAX_BINS);
    if (!(fmax > fmin) || !(mmax > mmin)) return fail(MIA_ERR_ARG, "mia_resample_mattes: empty intensity range");
    KParams p;
    WsLayout w;
    int rc = metric_common(p, "mia_resample_mattes", d_mov, mov_dtype, mdims, d_fixed, fixed_dtype, odims, n_mats, A, d_A,
                           mode, d_ws, ws_bytes, s, w);
    if (rc) return rc;
    p.fmask = d_fixed_mask;
    p.mmask = d_moving_mask;
    p.d_hist64 = d_hist;
    p.d_count = d_count;
    p.bins = bins;
    p.fmin = fmin; p.fscale = (float)(bins - 4) / (fmax - fmin);
    p.mmin = mmin; p.mscale = (float)(bins - 4) / (mmax - mmin);
    if (mattes_min_grid(p) > kMaxGridX) return fail(MIA_ERR_UNSUPPORTED, "mia_resample_mattes: lattice too large for 32-bit block counters");
    CUDA_TRY(cudaMemsetAsync(d_hist, 0, sizeof(unsigned long long) * (size_t)n_mats * bins * bins, s));
    const size_t smem = sizeof(unsigned) * (size_t)bins * bins + 8u * (kParzenSteps + 1) + 8u;
    long long gx = (long long)sm_count() * (smem > 32 * 1024 ? 2 : 4);
    if (gx < mattes_min_grid(p)) gx = mattes_min_grid(p);
    if (gx > p.ntiles) gx = p.ntiles;
    if (gx > kMaxGridX) gx = kMaxGridX;
    dim3 grid((unsigned)gx, (unsigned)n_mats, 1);
    g_last_path = 0;
    if (mov_dtype == fixed_dtype && !p.mmask) {
        bool launched = false;
        const int trc = try_tile<kMattes, false>(mov_dtype, mode, d_mov, mdims, d_A ? nullptr : A, n_mats, p, s, launched);
        if (launched) return trc;
    }
    dispatch_metric<kMattes, false>(mov_dtype, fixed_dtype, p, mode, grid, smem, s);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// mia_resample_ssd_static: the fused resample + SSD launch for callers that cannot allocate device memory
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The metric workspace (staged matrices, per-block partials, tickets)
// lives in a file-scope device array; every GPU that loads the library holds MIA_STATIC_WS_BYTES of it until exit.
// All calls on one device share it: it keeps the previous call's contents (the tickets are cleared per call on the
// caller's stream, nothing else needs to start at zero), and two calls that overlap on the GPU race.
namespace {

__device__ __align__(256) unsigned char g_static_metric_ws[MIA_STATIC_WS_BYTES];

// In host code the array's name is a host-side stand-in, not a device address: ask the runtime for the address once
// per device, with that device current.
int static_metric_ws(void **ptr)
{
    static std::mutex mu;
    static void *cached[64] = {nullptr};
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(MIA_ERR_UNSUPPORTED, "mia_resample_ssd_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!cached[dev]) {
        void *p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&p, g_static_metric_ws));
        if (!p) return fail(MIA_ERR_CUDA, "mia_resample_ssd_static: cudaGetSymbolAddress returned a null address");
        cached[dev] = p;
    }
    *ptr = cached[dev];
    return MIA_OK;
}

}  // namespace

extern "C" int mia_resample_ssd_static(const void *d_mov, int mov_dtype, const int mdims[3], const void *d_fixed,
                                       int fixed_dtype, const int odims[3], void *d_out, int out_dtype, int n_mats,
                                       const double *A, const double *d_A, double background, int mode, double *d_ssd,
                                       unsigned long long *d_count, void *stream)
{
    if (n_mats < 1 || n_mats > MIA_STATIC_MAX_MATS)
        return fail(MIA_ERR_ARG, "mia_resample_ssd_static: n_mats must be 1..%d (the static scratch holds %d bytes)",
                    MIA_STATIC_MAX_MATS, MIA_STATIC_WS_BYTES);
    const size_t need = ws_layout(n_mats).total;
    if (need > (size_t)MIA_STATIC_WS_BYTES)
        return fail(MIA_ERR_WORKSPACE, "mia_resample_ssd_static: %d matrices need %zu bytes of scratch, the static array holds %d",
                    n_mats, need, MIA_STATIC_WS_BYTES);
    void *ws = nullptr;
    const int rc = static_metric_ws(&ws);
    if (rc) return rc;
    return mia_resample_ssd_masked(d_mov, mov_dtype, mdims, d_fixed, fixed_dtype, nullptr, nullptr, odims, d_out, out_dtype,
                                   n_mats, A, d_A, background, mode, d_ssd, d_count, ws, (size_t)MIA_STATIC_WS_BYTES, stream);
}

#include "mia_host_api.inl"
#include "mia_dvf.inl"
#include "mia_dose.inl"
#include "mia_io.inl"
#include "mia_pyramid.inl"
