// C-ABI of the displacement-field kernels; see mia_dvf.cuh and include/mia_b200.h.

namespace {

template <typename TM, typename TO>
void launch_warp_dvf(const mia::DvfParams &p, int mode, unsigned grid, cudaStream_t s)
{
    if (mode == MIA_MODE_VTK) mia::warp_dvf_kernel<TM, TO, true><<<grid, mia::kDvfThreads, 0, s>>>(p);
    else mia::warp_dvf_kernel<TM, TO, false><<<grid, mia::kDvfThreads, 0, s>>>(p);
}

// Box-staged kernel (mia_dvf.cuh::warp_dvf_tile_kernel): needs a TMA descriptor of the image, like the rigid tile path.
using DvfBoxI16 = mia::Box<96, 15, 14>;       // 40 KB: bank-friendly 96-short pitch (see TileCfg), 3 CTAs per SM
using DvfBoxF32 = mia::Box<64, 14, 13>;       // 46.6 KB

bool dvf_staged_disabled()
{
    static const bool off = [] { const char *e = getenv("MIA_DVF_DIRECT"); return e && *e && *e != '0'; }();
    return off;
}

template <typename TM, class BOX>
bool try_warp_dvf_staged(mia::DvfParams p, const void *d_mov, const int mdims[3], int mode, cudaStream_t s)
{
    if (dvf_staged_disabled() || tile_path_disabled()) return false;
    alignas(64) CUtensorMap map;
    if (!make_tensor_map<TM, BOX>(&map, d_mov, mdims)) return false;
    const long long ntiles = (long long)((p.ox + 31) / 32) * ((p.oy + 7) / 8) * ((p.oz + mia::kTZ - 1) / mia::kTZ);
    if (ntiles > 0x7fffffffLL || (long long)p.ox * p.oy * mia::kTZ >= 0x7fffffffLL) return false;
    p.magic_fold = 0u - mia::kMagicBits * (unsigned)((BOX::kY + 1) * BOX::kX + 1) * (unsigned)sizeof(TM);
    constexpr size_t smem = (size_t)BOX::kX * BOX::kY * BOX::kZ * sizeof(TM);
    long long gx = (long long)sm_count() * MIA_DVF_STAGED_MINB;
    if (gx > ntiles) gx = ntiles;
    // the float32 image of B decides; the diagonal kernel differs from the general one by one float32 rounding of a
    // local coordinate (~1e-7 voxel)
    const bool diag = (float)p.B[1] == 0.0f && (float)p.B[2] == 0.0f && (float)p.B[3] == 0.0f && (float)p.B[5] == 0.0f &&
                      (float)p.B[6] == 0.0f && (float)p.B[7] == 0.0f;
    const unsigned g = (unsigned)gx;
    if (mode == MIA_MODE_VTK) {
        if (diag) mia::warp_dvf_tile_kernel<TM, TM, true, BOX, true><<<g, mia::kDvfThreads, smem, s>>>(map, p);
        else mia::warp_dvf_tile_kernel<TM, TM, true, BOX, false><<<g, mia::kDvfThreads, smem, s>>>(map, p);
    } else {
        if (diag) mia::warp_dvf_tile_kernel<TM, TM, false, BOX, true><<<g, mia::kDvfThreads, smem, s>>>(map, p);
        else mia::warp_dvf_tile_kernel<TM, TM, false, BOX, false><<<g, mia::kDvfThreads, smem, s>>>(map, p);
    }
    g_last_path = 3;
    return true;
}


// Software-pipelined kernel (mia_dvf.cuh::warp_dvf_pipe_kernel): the field arrives by TMA too, through a tensor map of
// the field seen as [oz][oy][3 ox] floats with one 96 x 8 x 8 box per tile -- needs ox % 4 == 0 (global strides in
// multiples of 16 bytes) and a 16-byte aligned field; anything else takes the serial staged kernel below.
using DvfPipeBoxI16 = mia::Box<80, 15, 13>;   // 31.2 KB, two of them: 40-word pitch (8 mod 32)
using DvfPipeBoxF32 = mia::Box<64, 14, 13>;   // 46.6 KB, one

bool dvf_pipe_disabled()
{
    static const bool off = [] { const char *e = getenv("MIA_DVF_NO_PIPE"); return e && *e && *e != '0'; }();
    return off;
}

template <typename TM, class BOX, int NBOX, bool VTK, bool DIAG>
bool dvf_launch_pipe(const CUtensorMap &map, const CUtensorMap &fmap, const mia::DvfParams &p, unsigned g, cudaStream_t s)
{
    auto k = mia::warp_dvf_pipe_kernel<TM, TM, VTK, BOX, DIAG, NBOX>;
    constexpr size_t smem = mia::DvfPipeSmem<BOX, TM, NBOX>::kBytes;
    static std::mutex mu;
    static bool done[64] = {false};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return false;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!done[dev]) {             // more than 48 KB of dynamic shared memory: opt in once per instantiation and device
            if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return false;
            done[dev] = true;
        }
    }
    k<<<g, mia::kDvfThreads, smem, s>>>(map, fmap, p);
    return true;
}

template <typename TM, class BOX, int NBOX>
bool try_warp_dvf_pipe(mia::DvfParams p, const void *d_mov, const int mdims[3], int mode, cudaStream_t s)
{
    if (dvf_staged_disabled() || dvf_pipe_disabled() || tile_path_disabled()) return false;
    if (p.ox % 4 != 0 || ((uintptr_t)p.dvf & 15u) != 0) return false;
    alignas(64) CUtensorMap map, fmap;
    if (!make_tensor_map<TM, BOX>(&map, d_mov, mdims)) return false;
    const long long ntiles = (long long)((p.ox + 31) / 32) * ((p.oy + 7) / 8) * ((p.oz + mia::kTZ - 1) / mia::kTZ);
    if (ntiles > 0x7fffffffLL || (long long)p.ox * p.oy * p.oz >= 0x7fffffffLL) return false;      // 32-bit output offsets
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return false;
    const cuuint64_t gdim[3] = {(cuuint64_t)p.ox * 3u, (cuuint64_t)p.oy, (cuuint64_t)p.oz};
    const cuuint64_t gstr[2] = {(cuuint64_t)p.ox * 12u, (cuuint64_t)p.ox * 12u * (cuuint64_t)p.oy};
    const cuuint32_t box[3] = {96, 8, (cuuint32_t)mia::kTZ};
    const cuuint32_t estr[3] = {1, 1, 1};
    if (enc(&fmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float *>(p.dvf), gdim, gstr, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return false;
    p.magic_fold = 0u - mia::kMagicBits * (unsigned)((BOX::kY + 1) * BOX::kX + 1) * (unsigned)sizeof(TM);
    long long gx = (long long)sm_count() * MIA_DVF_PIPE_MINB;
    if (gx > ntiles) gx = ntiles;
    const bool diag = (float)p.B[1] == 0.0f && (float)p.B[2] == 0.0f && (float)p.B[3] == 0.0f && (float)p.B[5] == 0.0f &&
                      (float)p.B[6] == 0.0f && (float)p.B[7] == 0.0f;
    const unsigned g = (unsigned)gx;
    const bool ok = mode == MIA_MODE_VTK
        ? (diag ? dvf_launch_pipe<TM, BOX, NBOX, true, true>(map, fmap, p, g, s) : dvf_launch_pipe<TM, BOX, NBOX, true, false>(map, fmap, p, g, s))
        : (diag ? dvf_launch_pipe<TM, BOX, NBOX, false, true>(map, fmap, p, g, s) : dvf_launch_pipe<TM, BOX, NBOX, false, false>(map, fmap, p, g, s));
    if (!ok) return false;
    g_last_path = 9;
    return true;
}

}  // namespace

extern "C" {

int mia_warp_dvf(const void *d_mov, int mov_dtype, const int mdims[3], const float *d_dvf, void *d_out, int out_dtype,
                 const int odims[3], const double A[12], const double B[9], double background, int mode, void *stream)
{
    if (!d_mov || !d_dvf || !d_out || !A || !B) return fail(MIA_ERR_ARG, "mia_warp_dvf: null pointer");
    if (!dims_ok(mdims) || !dims_ok(odims)) return fail(MIA_ERR_ARG, "mia_warp_dvf: dims must be positive");
    if (!dtype_ok(mov_dtype) || !dtype_ok(out_dtype)) return fail(MIA_ERR_DTYPE, "mia_warp_dvf: bad dtype");
    if (mov_dtype != out_dtype)
        return fail(MIA_ERR_UNSUPPORTED, "mia_warp_dvf: output must have the image dtype (sitk.Resample keeps the pixel type)");
    if (mode != MIA_MODE_VTK && mode != MIA_MODE_ITK) return fail(MIA_ERR_ARG, "mia_warp_dvf: bad mode");
    for (int i = 0; i < 12; ++i)
        if (!std::isfinite(A[i])) return fail(MIA_ERR_ARG, "mia_warp_dvf: non-finite matrix");
    for (int i = 0; i < 9; ++i)
        if (!std::isfinite(B[i])) return fail(MIA_ERR_ARG, "mia_warp_dvf: non-finite matrix");
    mia::DvfParams p;
    p.mov.p = d_mov;
    p.mov.nx = mdims[0]; p.mov.ny = mdims[1]; p.mov.nz = mdims[2];
    p.dvf = d_dvf;
    p.out = d_out;
    p.ox = odims[0]; p.oy = odims[1]; p.oz = odims[2];
    memcpy(p.A.a, A, sizeof(double) * 12);
    memcpy(p.B, B, sizeof(double) * 9);
    p.background = (float)background;
    p.round_mode = mode == MIA_MODE_VTK ? 0 : 1;
    const long long nwork = (long long)((odims[0] + 31) / 32) * ((odims[1] + 7) / 8) * ((odims[2] + mia::kDvfTZ - 1) / mia::kDvfTZ);
    long long gx = (long long)sm_count() * 8;
    if (gx > nwork) gx = nwork;
    cudaStream_t s = (cudaStream_t)stream;
    g_last_path = 0;
    bool staged = mov_dtype == MIA_I16 ? try_warp_dvf_pipe<int16_t, DvfPipeBoxI16, 2>(p, d_mov, mdims, mode, s)
                                       : try_warp_dvf_pipe<float, DvfPipeBoxF32, 1>(p, d_mov, mdims, mode, s);
    if (!staged)
        staged = mov_dtype == MIA_I16 ? try_warp_dvf_staged<int16_t, DvfBoxI16>(p, d_mov, mdims, mode, s)
                                      : try_warp_dvf_staged<float, DvfBoxF32>(p, d_mov, mdims, mode, s);
    if (staged) {
        CUDA_TRY(cudaGetLastError());
        return MIA_OK;
    }
    if (mov_dtype == MIA_I16) launch_warp_dvf<int16_t, int16_t>(p, mode, (unsigned)gx, s);
    else launch_warp_dvf<float, float>(p, mode, (unsigned)gx, s);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

int mia_sample_field_points(const float *d_field, const int dims[3], int channels, const double *d_pts, long long n,
                            double scale, double *d_out, void *stream)
{
    if (n < 0) return fail(MIA_ERR_ARG, "mia_sample_field_points: negative point count");
    if (!dims_ok(dims)) return fail(MIA_ERR_ARG, "mia_sample_field_points: dims must be positive");
    if (channels < 1 || channels > 16) return fail(MIA_ERR_ARG, "mia_sample_field_points: channels must be 1..16");
    if (n == 0) return MIA_OK;                 // an empty point set has no buffers to check
    if (!d_field || !d_pts || !d_out) return fail(MIA_ERR_ARG, "mia_sample_field_points: null pointer");
    long long gx = (n + mia::kDvfThreads - 1) / mia::kDvfThreads;
    const long long cap = (long long)sm_count() * 16;
    if (gx > cap) gx = cap;
    mia::sample_field_points_kernel<<<(unsigned)gx, mia::kDvfThreads, 0, (cudaStream_t)stream>>>(
        d_field, dims[0], dims[1], dims[2], channels, d_pts, n, scale, d_out);
    CUDA_TRY(cudaGetLastError());
    return MIA_OK;
}

}  // extern "C"
