/*
 * mia_b200 -- C ABI of the B200-native rigid resample + similarity-metric hot path.
 *
 * The reference (caleb-oconnor/MedicalImageAnalysis v2.1.92) has NO FFI layer: it is pure Python over
 * the VTK / SimpleITK wheels.  Each entry point below names the reference call site whose native
 * arithmetic it replaces (paths relative to the reference checkout); INTEGRATION.md shows the ctypes
 * stub a maintainer would drop into those sites.
 *
 * Conventions
 *   - return 0 on success, negative on error; message via mia_last_error() (thread-local).
 *   - volumes are dense, C-order [z][y][x] (Image.array layout, read/dicom.py:530-534); dims arrays are
 *     (x, y, z) counts.  d_* pointers are DEVICE pointers owned by the caller; the library never allocates
 *     or frees device memory on the d_* entry points (workspaces are caller supplied).
 *   - A is the row-major 3x4 matrix "output voxel index (i,j,k,1) -> continuous moving index (x,y,z)",
 *     float64, built on the host from spacing/origin/direction and the 4x4 rigid matrix.
 *   - every d_* call is asynchronous on `stream` (a cudaStream_t passed as void*); no implicit sync.
 *   - plain pointers and sizes only: no torch / CUDA types in the signatures.
 */
#ifndef MIA_B200_H
#define MIA_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MIA_B200_VERSION 100

/* element types of Image.array on this path (read/dicom.py:530: int16; utils/creation.py:257: any) */
#define MIA_I16 0
#define MIA_F32 1

/* border / rounding semantics */
#define MIA_MODE_VTK 0 /* vtkImageReslice: inside [-0.5, N-0.5], int out = floor(v + 0.5)   */
#define MIA_MODE_ITK 1 /* sitk.Resample:   inside [-0.5, N-0.5), int out = trunc(v)          */

#define MIA_OK 0
#define MIA_ERR_ARG (-1)
#define MIA_ERR_DTYPE (-2)
#define MIA_ERR_WORKSPACE (-3)
#define MIA_ERR_CUDA (-4)
#define MIA_ERR_UNSUPPORTED (-5)

#define MIA_MAX_MATS 32  /* candidate matrices per batched metric launch */
#define MIA_MAX_BINS 128 /* joint-histogram bins per axis (bins^2 uint32 counters live in shared memory) */

int mia_version(void);
const char *mia_last_error(void);
/* Diagnostic: which kernel the calling thread's last resample / metric call launched
 * (0 = direct per-voxel kernel, 1 = TMA-staged 32-wide tile kernel, 2 = multi-matrix tile kernel, 3 = DVF tile kernel,
 * 4 / 7 / 8 = the tile kernel with the tall / deep / cube box family, 5 = x-pair tile kernel, 6 = x-pair multi-matrix
 * kernel, 9 = software-pipelined DVF tile kernel).  Used by the parity tests. */
int mia_debug_last_path(void);
/* Diagnostic / escape hatch: the x-pair kernel hands its tiles out through a per-launch claim counter (CTAs that drew
 * cheap tiles take more of them).  on != 0 keeps the static round-robin order instead (the same voxels either way,
 * bit for bit); the environment variable MIA_X2_STATIC=1 sets the initial state.  Needed only by a caller that runs
 * two executable instances of ONE captured graph concurrently.  Returns the previous setting.  No counterpart in the
 * reference. */
int mia_debug_x2_static(int on);

/*
 * Rigid trilinear resample of a device-resident moving volume onto an output lattice.
 * Replaces the vtkImageReslice::Update() of Rigid.create_image (structure/rigid.py:685-693; twins
 * structure/image.py:189-196, structure/dose.py:205-212) with mode = MIA_MODE_VTK, background = -3001,
 * and sitk.Resample(mov, ref, T, sitkLinear, default, pixelID) of structure/deformable.py:579,788-795 /
 * utils/deformable/simpleitk.py:81-94 with mode = MIA_MODE_ITK.
 */
int mia_resample_rigid(const void *d_mov, int mov_dtype, const int mdims[3], void *d_out, int out_dtype,
                       const int odims[3], const double A[12], double background, int mode, void *stream);

/*
 * Same, HOST buffers in and out (the reference boundary: Image.array is a NumPy array and
 * Display.compute_reslice adopts the output as one, structure/rigid.py:672,194-199).  The library
 * stages through its own device buffers, pipelining H2D / kernel / D2H by z-slab on internal streams;
 * synchronous on return.  device = CUDA ordinal.
 */
int mia_resample_rigid_host(const void *h_mov, int mov_dtype, const int mdims[3], void *h_out,
                            int out_dtype, const int odims[3], const double A[12], double background,
                            int mode, int device);

/*
 * Output lattice of vtkImageReslice with AutoCropOutputOn + identity output direction
 * (structure/rigid.py:689-691): host-only helper, no GPU work.  M = 4x4 row-major fixed->moving physical,
 * P2W_mov = 4x4 row-major moving index->physical.  extent_rule 0 = Round (VTK), 1 = ceil.
 */
int mia_autocrop_extent(const double M[16], const double P2W_mov[16], const int mdims[3],
                        const double out_spacing[3], int extent_rule, int odims[3], double oorigin[3]);

/*
 * Exact axis permutation / flip copy: out = flip(transpose(in, perm), axes where flip[a] != 0), axes in
 * [z, y, x] order (NumPy semantics).  Replaces the np.rot90 / np.flip / transpose views of the
 * Feet-First-Supine normalisation (read/dicom.py:695-717, RTDOSE twin :2066-2088) that the reference
 * materialises later with array.flatten() (structure/rigid.py:672).  Bit exact.  idims = (x, y, z) of `in`.
 */
int mia_permute_axes(const void *d_in, int dtype, const int idims[3], void *d_out, const int perm[3],
                     const int flip[3], void *stream);

/* Bytes of caller-supplied device workspace the two metric launches need (16-byte aligned base). */
size_t mia_metric_workspace_bytes(const int odims[3], int n_mats, int bins);

/*
 * Fused resample + sum of squared differences over n_mats candidate matrices in ONE launch.
 * For matrix m: d_ssd[m] = sum over fixed voxels whose sample is inside of (F - mov(A_m v))^2 with the
 * UNROUNDED interpolated value, d_count[m] = number of inside voxels.  If d_out != NULL (n_mats must be 1)
 * the resampled volume is also stored (background outside).  Matrices come from the host (A) or, when
 * d_A != NULL, from device memory (for graph-captured optimiser loops).  No reference implementation
 * exists (SetMetricAsMeanSquares is only named at utils/deformable/simpleitk.py:115): spec = SURVEY M1.
 * The reduction is deterministic (fixed-order float64 second stage).
 */
int mia_resample_ssd(const void *d_mov, int mov_dtype, const int mdims[3], const void *d_fixed,
                     int fixed_dtype, const int odims[3], void *d_out, int out_dtype, int n_mats,
                     const double *A, const double *d_A, double background, int mode, double *d_ssd,
                     unsigned long long *d_count, void *d_ws, size_t ws_bytes, void *stream);

/*
 * Fused resample + joint histogram: d_hist[m][bf][bm] (uint32, zeroed by the call) over inside voxels,
 * bf = clamp(floor((F - fmin) * bins / (fmax - fmin)), 0, bins-1), bm likewise from the interpolated value.
 * Spec = SURVEY M2 ("AdvancedMattesMutualInformation" + NumberOfHistogramBins are only named at
 * utils/deformable/simpleitk.py:131,155-156).  Integer counts => bit-reproducible across shardings.
 */
int mia_resample_jointhist(const void *d_mov, int mov_dtype, const int mdims[3], const void *d_fixed,
                           int fixed_dtype, const int odims[3], int n_mats, const double *A,
                           const double *d_A, int mode, float fmin, float fmax, float mmin, float mmax,
                           int bins, unsigned int *d_hist, unsigned long long *d_count, void *d_ws,
                           size_t ws_bytes, void *stream);

/* d_mi[m] = sum p_ij ln(p_ij / (p_i. p_.j)) in float64 from integer counts (natural log). */
int mia_mi_from_hist(const unsigned int *d_hist, int bins, int n_mats, double *d_mi, void *stream);

/*
 * Device-side optimiser step (no host round trip): reads the 13 metric values of the current probe set,
 * applies one regular-step gradient-descent update with central differences and step halving, and writes
 * the next 13 candidate matrices to d_A.  See DESIGN.md "registration loop".  d_state is an opaque
 * device block of mia_optimizer_state_bytes() bytes initialised by mia_optimizer_init.
 */
size_t mia_optimizer_state_bytes(void);
int mia_optimizer_init(void *d_state, const double params0[6], const double center[3],
                       const double P2W_fixed[16], const double W2P_mov[16], double rot_step_deg,
                       double trans_step_mm, double min_step_scale, int maximize, double *d_A, void *stream);
int mia_optimizer_step(void *d_state, const double *d_metric, const unsigned long long *d_count,
                       int metric_is_ssd, double *d_A, void *stream);
/*
 * The same step for the sharded loops (z-slabs of one large pair, or the 13 probes dealt out over the ranks).
 * mia_metric_pack gathers this rank's partial results into ONE float64 buffer d_packed[2 * n_total]:
 * [0, n_total) the metric and [n_total, 2 n_total) the voxel counts as float64 (exact below 2^53), both zero outside
 * the rank's share [q0, q0 + n_local) -- so a single all-reduce(SUM) of 2 * n_total doubles per iteration combines the
 * ranks.  mia_optimizer_step_packed then reads that buffer (n_total = 13) and, when d_A_slab != NULL, also writes the
 * probe matrices shifted to the z-slab that starts at fixed plane k0 (A[:, 3] += A[:, 2] * k0).
 */
int mia_metric_pack(const double *d_metric, const unsigned long long *d_count, int n_local, int q0, int n_total,
                    double *d_packed, void *stream);
int mia_optimizer_step_packed(void *d_state, const double *d_packed, int metric_is_ssd, double *d_A, double *d_A_slab,
                              int k0, void *stream);
/*
 * The exchange of those 26 doubles WITHOUT a collective library call ("mia_allreduce_fused").  Every
 * rank owns a buffer of mia_fused_exchange_bytes() in symmetric, peer-mapped memory (zeroed, then a host barrier, before
 * the first use; e.g. torch.distributed._symmetric_memory).  mia_fused_push stores this rank's packed partials into slot
 * [parity][rank] of EVERY rank's buffer through NVLink (d_peer_bufs: device array of `world` buffer pointers, own one
 * included), fences and raises its flag in every buffer; mia_optimizer_step_fused waits for the `world` flags of the
 * iteration, adds the slots in rank order and steps as mia_optimizer_step_packed does.  Two small launches per
 * iteration, no host involvement, capturable.  A wait gives up after about two seconds (a dead peer must not hang the
 * GPU) and counts the event; mia_fused_timeouts reads the counter back (synchronises): non-zero = results invalid.
 */
size_t mia_fused_exchange_bytes(void);
int mia_fused_push(const void *d_state, const double *d_packed, void *const *d_peer_bufs, int rank, int world, void *stream);
int mia_optimizer_step_fused(void *d_state, void *d_my_buf, int world, int metric_is_ssd, double *d_A, double *d_A_slab,
                             int k0, void *stream);
int mia_fused_timeouts(const void *d_my_buf, unsigned long long *timeouts, void *stream);
/* Copy the optimiser state back (synchronises `stream`).  Any output pointer may be NULL; hist_* hold
 * min(iters, max_hist) entries: the centre cost, parameters [n][6] and step scale seen at each step. */
int mia_optimizer_read(const void *d_state, double params[6], double *scale, int *iters, int *converged,
                       int *n_reject, double *hist_cost, double *hist_params, double *hist_scale,
                       int max_hist, void *stream);

/*
 * SURVEY N1 -- displacement fields next to the rigid path.
 *
 * mia_warp_dvf replaces the second stage of Deformable.create_image (structure/deformable.py:797-806:
 * sitk.Resample(image, sitk.DisplacementFieldTransform(dvf), sitk.sitkLinear, -3001, pixel id)).  d_dvf holds
 * one physical displacement (x, y, z) per OUTPUT voxel, float32, reference layout [z][y][x][3] (self.dvf,
 * deformable.py:797).  Output voxel idx is sampled at moving index  A . (idx, 1) + B . d :
 * A = W2P_image . P2W_out (3x4, row-major), B = linear part of W2P_image (3x3, row-major).  Inside rule,
 * rounding and background as mia_resample_rigid; the output keeps the image's pixel type.  A displacement that is
 * NaN gives the background value.  Any sizes and pitches work; the fast kernel (image box and field tile staged by
 * TMA) needs an image row and a field row that are multiples of 16 bytes (odims[0] % 4 == 0 for the field) and
 * 16-byte aligned base pointers.
 */
int mia_warp_dvf(const void *d_mov, int mov_dtype, const int mdims[3], const float *d_dvf, void *d_out, int out_dtype,
                 const int odims[3], const double A[12], const double B[9], double background, int mode, void *stream);

/*
 * scipy.ndimage.map_coordinates(field[..., c], [z, y, x], order=1, mode='nearest') for a point set, all channels
 * in one pass (structure/deformable.py:868-878 and :1095-1104: ROI mesh vertices through the DVF).
 * d_field: float32 [z][y][x][channels]; d_pts: float64 [n][3] continuous voxel coordinates (x, y, z);
 * d_out: float64 [n][channels] = scale * interpolated value (the reference's `percent / 100`).
 */
int mia_sample_field_points(const float *d_field, const int dims[3], int channels, const double *d_pts, long long n,
                            double scale, double *d_out, void *stream);

/*
 * SURVEY N3 -- dose on an image lattice under an ROI mask, and its dose-volume statistics.
 *
 * Replaces Dose.compute_roi_dose_array (structure/dose.py:738-772: sitk.ResampleImageFilter, reference image =
 * the CT, sitkLinear, default pixel 0.0, then dose_arr[np.where(mask_arr > 0)]) and the NumPy reductions of
 * Dose.compute_roi_dose_statistics (structure/dose.py:774-816: min / max / mean / median / std, np.percentile,
 * counts below dose levels).
 *
 * mia_mask_scan counts the non-zero bytes of d_mask (uint8, [z][y][x] on the output lattice `odims`) per
 * 4096-voxel chunk, scans the counts into the workspace and writes the total to *d_total (device).
 * mia_resample_masked then samples ONLY the masked voxels (same arithmetic, inside rule and background as
 * mia_resample_rigid, value kept as unrounded float32) and writes them compacted in C order -- the order of
 * np.where -- to d_out[0 .. total).  The workspace must be the one mia_mask_scan filled for the same mask.
 */
size_t mia_mask_workspace_bytes(const int odims[3]);
int mia_mask_scan(const unsigned char *d_mask, const int odims[3], unsigned long long *d_total, void *d_ws,
                  size_t ws_bytes, void *stream);
int mia_resample_masked(const void *d_mov, int mov_dtype, const int mdims[3], const unsigned char *d_mask,
                        const int odims[3], const double A[12], double background, int mode, const void *d_ws,
                        size_t ws_bytes, float *d_out, void *stream);

/*
 * Statistics of a compact float32 device sample of n > 0 values.  levels (HOST, ascending, <= 64) and ranks
 * (HOST, 0-based order-statistic indices < n, <= 64) are copied by the call.  d_result (device, float64):
 *   [0] min  [1] max  [2] sum  [3] mean  [4] sum (x - mean)^2  [5] n
 *   [6 .. 6 + n_levels)                      number of values < levels[l]
 *   [6 + n_levels .. 6 + n_levels + n_ranks) the ranks[r]-th smallest value (exact: three-pass radix select)
 * np.percentile(x, q) = lerp of the order statistics floor / ceil of q / 100 * (n - 1); np.std = sqrt([4] / n).
 */
size_t mia_sample_stats_workspace_bytes(void);
int mia_sample_stats(const float *d_values, unsigned long long n, const float *levels, int n_levels,
                     const unsigned long long *ranks, int n_ranks, double *d_result, void *d_ws, size_t ws_bytes,
                     void *stream);

/*
 * SURVEY N4 -- raw volume payloads between files and device memory (MetaImage .raw, the data block of an .npy).
 * Replaces the host arrays of sitk.ReadImage + GetArrayFromImage (read/mhd.py:157-196), np.load / np.save of
 * array.npy (structure/image.py:709-735, 804-836) and vtkMetaImageWriter (structure/rigid.py:697-716) for
 * device-resident volumes.  The payload moves through page-locked 8 MiB staging buffers (four worker threads with two
 * each, per device), file I/O of one chunk overlapping the PCIe copy of another; both calls synchronise before they return.
 * `offset` is the byte offset of the payload inside the file (header size of an .npy, 0 for a .raw).
 */
int mia_read_file_to_device(const char *path, unsigned long long offset, unsigned long long bytes, void *d_dst, int device);
int mia_write_file_from_device(const char *path, unsigned long long offset, const void *d_src, unsigned long long bytes,
                               int truncate, int device);

/*
 * SURVEY N5 -- one level of a multi-resolution pyramid (what "NumberOfResolutions" asks of Elastix,
 * utils/deformable/simpleitk.py:150-163): d_out[k][j][i] = mean of the factors[0] x factors[1] x factors[2] block of
 * d_in starting at (i, j, k) * factors (fewer voxels on the high edges), int16 rounded half-up.  factors and idims
 * are (x, y, z); output dims = ceil(idims / factors).  The level's lattice: spacing * factors, origin moved by
 * half a (factor - 1) voxel along each axis (geometry.pyramid_grid).
 */
int mia_shrink_mean(const void *d_in, int dtype, const int idims[3], void *d_out, const int factors[3], void *stream);

/*
 * SURVEY N5 -- fixed-image masks for the fused metrics (SetMetricFixedMask / elastix SetFixedMask,
 * utils/deformable/simpleitk.py:116-117, 146-147).  Same arguments as mia_resample_ssd / mia_resample_jointhist
 * plus d_fixed_mask: uint8 [z][y][x] on the fixed lattice, 0 = the voxel takes no part in the metric (NULL = no
 * mask), and d_moving_mask (SetMetricMovingMask, :118-119, 148-149): uint8 on the MOVING lattice, looked up at the
 * voxel nearest to the mapped point (floor(c + 0.5)); a launch with a moving mask runs on the per-voxel kernel.
 * A masked SSD does not store the resampled volume (d_out must be NULL).
 */
int mia_resample_ssd_masked(const void *d_mov, int mov_dtype, const int mdims[3], const void *d_fixed,
                            int fixed_dtype, const unsigned char *d_fixed_mask, const unsigned char *d_moving_mask,
                            const int odims[3], void *d_out, int out_dtype, int n_mats, const double *A, const double *d_A,
                            double background, int mode, double *d_ssd, unsigned long long *d_count, void *d_ws,
                            size_t ws_bytes, void *stream);
int mia_resample_jointhist_masked(const void *d_mov, int mov_dtype, const int mdims[3], const void *d_fixed,
                                  int fixed_dtype, const unsigned char *d_fixed_mask, const unsigned char *d_moving_mask,
                                  const int odims[3], int n_mats, const double *A, const double *d_A, int mode, float fmin,
                                  float fmax, float mmin, float mmax, int bins, unsigned int *d_hist,
                                  unsigned long long *d_count, void *d_ws, size_t ws_bytes, void *stream);

/*
 * SURVEY N5 -- Mattes mutual information ("AdvancedMattesMutualInformation" + NumberOfHistogramBins,
 * utils/deformable/simpleitk.py:155-156; itk::MattesMutualInformationImageToImageMetricv4): `bins` bins per axis of
 * which 2 on either side are padding, box window on the fixed image, cubic B-spline Parzen window over 4 bins on
 * the moving image.  The B-spline is evaluated on a 1/256 sub-bin lattice with integer weights that sum to 1024 per
 * sample, so d_hist ([n][bins][bins] uint64, units of 1/1024 sample) is an INTEGER histogram: bit-identical for
 * every grid, sharding and GPU count; d_count[n] = samples.  mia_mi_from_hist64 turns it into MI (nats).
 * d_fixed_mask and d_moving_mask may be NULL (semantics as in mia_resample_ssd_masked).
 */
int mia_resample_mattes(const void *d_mov, int mov_dtype, const int mdims[3], const void *d_fixed, int fixed_dtype,
                        const unsigned char *d_fixed_mask, const unsigned char *d_moving_mask, const int odims[3],
                        int n_mats, const double *A, const double *d_A, int mode, float fmin, float fmax, float mmin,
                        float mmax, int bins, unsigned long long *d_hist, unsigned long long *d_count, void *d_ws,
                        size_t ws_bytes, void *stream);
int mia_mi_from_hist64(const unsigned long long *d_hist, int bins, int n_mats, double *d_mi, void *stream);

/*
 * mia_resample_ssd without any caller workspace and without any allocation.  This entry point has NO counterpart in
 * the reference: it was added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  Arguments and results are those of
 * mia_resample_ssd minus d_ws / ws_bytes; the scratch (staged matrices, per-block partials, tickets) is a file-scope
 * device array of MIA_STATIC_WS_BYTES that every GPU which loads the library holds until exit.  Limit: n_mats <=
 * MIA_STATIC_MAX_MATS, anything larger is rejected with MIA_ERR_ARG.  The call makes no cudaMalloc / cudaFree.
 * All calls on a device share the array and it keeps the last call's contents (what must start at zero is cleared
 * on `stream` by the call): calls of one process that overlap on the GPU (different streams or threads)
 * race; serialise them on one stream.
 */
#define MIA_STATIC_MAX_MATS 16
#define MIA_STATIC_WS_BYTES (512 * 1024)
int mia_resample_ssd_static(const void *d_mov, int mov_dtype, const int mdims[3], const void *d_fixed,
                            int fixed_dtype, const int odims[3], void *d_out, int out_dtype, int n_mats,
                            const double *A, const double *d_A, double background, int mode, double *d_ssd,
                            unsigned long long *d_count, void *stream);

/* Page-locked host buffers for the *_host entry point (a NumPy array can wrap them).  mia_host_register pins memory the
 * caller already owns -- e.g. its z-slab of ONE volume in a shared mapping that several ranks resample together
 * (Rigid.create_image on a volume cut over the GPUs of a box, structure/rigid.py:646-695). */
void *mia_host_alloc(size_t bytes);
int mia_host_free(void *p);
int mia_host_register(void *p, size_t bytes);
int mia_host_unregister(void *p);

#ifdef __cplusplus
}
#endif
#endif /* MIA_B200_H */
