/*
 * C / OpenMP restatement of the rigid trilinear resampler and the two metrics.
 *
 * TEST INFRASTRUCTURE + the timed CPU baseline of bench.py.  Never linked into the product library.
 * PARITY UNPINNED: the reference (caleb-oconnor/MedicalImageAnalysis v2.1.92) delegates this arithmetic
 * to VTK 9.3.1 vtkImageReslice (structure/rigid.py:685-693) and SimpleITK 2.4.1 ResampleImageFilter
 * (structure/deformable.py:579,788-795); neither is vendored nor installable here, so this file restates
 * their published algorithms (see oracle/reslice.py for the clause-by-clause description) and is itself
 * checked against oracle/reslice.py (NumPy float64) by tests/test_oracle.py.
 *
 * Arithmetic is float64 throughout, like vtkImageResliceExecute's double row buffer and ITK's
 * InternalComputationType.  Compile WITHOUT -ffast-math and with -ffp-contract=off so it matches NumPy.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define DT_I16 0
#define DT_F32 1
#define MODE_VTK 0 /* closed inside test, weights-product formula, floor(v+0.5)  */
#define MODE_ITK 1 /* half-open inside test, nested lerps, truncation            */

static inline double tap(const void *mov, int dt, size_t idx)
{
    return dt == DT_I16 ? (double)((const int16_t *)mov)[idx] : (double)((const float *)mov)[idx];
}

static inline int clampi(long v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : (int)v); }

/* One sample.  Returns 1 when inside (value in *val), 0 when outside. */
static inline int sample(const void *mov, int dt, const int md[3], const double c[3], int mode, double *val)
{
    long i0[3];
    int a0[3], a1[3];
    double f[3];
    for (int a = 0; a < 3; ++a) {
        double hi = md[a] - 0.5;
        if (c[a] < -0.5) return 0;
        if (mode == MODE_VTK ? (c[a] > hi) : (c[a] >= hi)) return 0;
        double fl = floor(c[a]);
        f[a] = c[a] - fl;
        i0[a] = (long)fl;
        a0[a] = clampi(i0[a], 0, md[a] - 1);
        a1[a] = clampi(i0[a] + (f[a] != 0.0), 0, md[a] - 1);
    }
    size_t sx = 1, sy = (size_t)md[0], sz = (size_t)md[0] * md[1];
    double v000 = tap(mov, dt, a0[2] * sz + a0[1] * sy + a0[0] * sx);
    double v100 = tap(mov, dt, a0[2] * sz + a0[1] * sy + a1[0] * sx);
    double v010 = tap(mov, dt, a0[2] * sz + a1[1] * sy + a0[0] * sx);
    double v110 = tap(mov, dt, a0[2] * sz + a1[1] * sy + a1[0] * sx);
    double v001 = tap(mov, dt, a1[2] * sz + a0[1] * sy + a0[0] * sx);
    double v101 = tap(mov, dt, a1[2] * sz + a0[1] * sy + a1[0] * sx);
    double v011 = tap(mov, dt, a1[2] * sz + a1[1] * sy + a0[0] * sx);
    double v111 = tap(mov, dt, a1[2] * sz + a1[1] * sy + a1[0] * sx);
    double fx = f[0], fy = f[1], fz = f[2];
    if (mode == MODE_VTK) {
        double rx = 1.0 - fx, ry = 1.0 - fy, rz = 1.0 - fz;
        double ryrz = ry * rz, fyrz = fy * rz, ryfz = ry * fz, fyfz = fy * fz;
        *val = (rx * (ryrz * v000 + ryfz * v001 + fyrz * v010 + fyfz * v011) +
                fx * (ryrz * v100 + ryfz * v101 + fyrz * v110 + fyfz * v111));
    } else {
        double vx00 = v000 + (v100 - v000) * fx;
        double vx10 = v010 + (v110 - v010) * fx;
        double vx01 = v001 + (v101 - v001) * fx;
        double vx11 = v011 + (v111 - v011) * fx;
        double vxx0 = vx00 + (vx10 - vx00) * fy;
        double vxx1 = vx01 + (vx11 - vx01) * fy;
        *val = vxx0 + (vxx1 - vxx0) * fz;
    }
    return 1;
}

static inline int16_t to_i16(double v, int mode)
{
    if (v < -32768.0) v = -32768.0;
    if (v > 32767.0) v = 32767.0;
    return (int16_t)(mode == MODE_VTK ? floor(v + 0.5) : trunc(v));
}

static inline void coord(const double *A, int i, int j, int k, double c[3])
{
    for (int a = 0; a < 3; ++a)
        c[a] = ((A[a * 4 + 3] + (double)k * A[a * 4 + 2]) + (double)j * A[a * 4 + 1]) + (double)i * A[a * 4 + 0];
}

int oracle_set_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
    return omp_get_max_threads();
#else
    (void)n;
    return 1;
#endif
}

/* out lattice odims = (x,y,z); both volumes [z][y][x] contiguous. */
int oracle_reslice(const void *mov, int mov_dt, const int md[3], void *out, int out_dt, const int od[3],
                   const double *A, double background, int mode)
{
    if (!mov || !out || !A) return -1;
    const int nx = od[0], ny = od[1], nz = od[2];
    const int16_t bg16 = to_i16(background, mode);
    const float bgf = (float)background;
#pragma omp parallel for collapse(2) schedule(static)
    for (int k = 0; k < nz; ++k)
        for (int j = 0; j < ny; ++j) {
            size_t row = ((size_t)k * ny + j) * nx;
            for (int i = 0; i < nx; ++i) {
                double c[3], v;
                coord(A, i, j, k, c);
                int ok = sample(mov, mov_dt, md, c, mode, &v);
                if (out_dt == DT_I16)
                    ((int16_t *)out)[row + i] = ok ? to_i16(v, mode) : bg16;
                else
                    ((float *)out)[row + i] = ok ? (float)v : bgf;
            }
        }
    return 0;
}

/* ssd over inside voxels, optional store of the resampled volume (out may be NULL). */
int oracle_reslice_ssd(const void *mov, int mov_dt, const int md[3], const void *fixed, int fixed_dt,
                       void *out, int out_dt, const int od[3], const double *A, double background,
                       int mode, double *ssd, unsigned long long *count)
{
    if (!mov || !fixed || !A || !ssd || !count) return -1;
    const int nx = od[0], ny = od[1], nz = od[2];
    const int16_t bg16 = to_i16(background, mode);
    const float bgf = (float)background;
    double total = 0.0;
    unsigned long long n = 0;
    /* per-row partials summed in row order => deterministic irrespective of thread count */
    double *rows = (double *)malloc(sizeof(double) * (size_t)nz * ny);
    unsigned *rown = (unsigned *)malloc(sizeof(unsigned) * (size_t)nz * ny);
    if (!rows || !rown) { free(rows); free(rown); return -2; }
#pragma omp parallel for collapse(2) schedule(static)
    for (int k = 0; k < nz; ++k)
        for (int j = 0; j < ny; ++j) {
            size_t row = ((size_t)k * ny + j) * nx;
            double s = 0.0;
            unsigned cnt = 0;
            for (int i = 0; i < nx; ++i) {
                double c[3], v;
                coord(A, i, j, k, c);
                int ok = sample(mov, mov_dt, md, c, mode, &v);
                if (ok) {
                    double d = tap(fixed, fixed_dt, row + i) - v;
                    s += d * d;
                    ++cnt;
                }
                if (out) {
                    if (out_dt == DT_I16)
                        ((int16_t *)out)[row + i] = ok ? to_i16(v, mode) : bg16;
                    else
                        ((float *)out)[row + i] = ok ? (float)v : bgf;
                }
            }
            rows[(size_t)k * ny + j] = s;
            rown[(size_t)k * ny + j] = cnt;
        }
    for (size_t r = 0; r < (size_t)nz * ny; ++r) { total += rows[r]; n += rown[r]; }
    free(rows);
    free(rown);
    *ssd = total;
    *count = n;
    return 0;
}

static inline int bin_of(float v, float vmin, float scale, int bins)
{
    float b = floorf((v - vmin) * scale);
    if (!(b >= 0.0f)) return 0;
    if (b > (float)(bins - 1)) return bins - 1;
    return (int)b;
}

/* hist[bf * bins + bm], uint64, zeroed here. */
int oracle_joint_hist(const void *mov, int mov_dt, const int md[3], const void *fixed, int fixed_dt,
                      const int od[3], const double *A, int mode, float fmin, float fmax, float mmin,
                      float mmax, int bins, unsigned long long *hist, unsigned long long *count)
{
    if (!mov || !fixed || !A || !hist || !count || bins < 1) return -1;
    const int nx = od[0], ny = od[1], nz = od[2];
    const float fscale = (float)bins / (fmax - fmin), mscale = (float)bins / (mmax - mmin);
    memset(hist, 0, sizeof(unsigned long long) * (size_t)bins * bins);
    unsigned long long n = 0;
#pragma omp parallel
    {
        unsigned long long *h = (unsigned long long *)calloc((size_t)bins * bins, sizeof(unsigned long long));
        unsigned long long ln = 0;
#pragma omp for collapse(2) schedule(static) nowait
        for (int k = 0; k < nz; ++k)
            for (int j = 0; j < ny; ++j) {
                size_t row = ((size_t)k * ny + j) * nx;
                for (int i = 0; i < nx; ++i) {
                    double c[3], v;
                    coord(A, i, j, k, c);
                    if (!sample(mov, mov_dt, md, c, mode, &v)) continue;
                    int bf = bin_of((float)tap(fixed, fixed_dt, row + i), fmin, fscale, bins);
                    int bm = bin_of((float)v, mmin, mscale, bins);
                    ++h[(size_t)bf * bins + bm];
                    ++ln;
                }
            }
#pragma omp critical
        {
            for (size_t b = 0; b < (size_t)bins * bins; ++b) hist[b] += h[b];
            n += ln;
        }
        free(h);
    }
    *count = n;
    return 0;
}
