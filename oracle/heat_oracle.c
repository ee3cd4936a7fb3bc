/*
 * heat_oracle.c -- CPU restatement of the reference solver loop (plain C).
 *
 * TEST INFRASTRUCTURE ONLY (see heat_oracle.h).  Build with
 *     gcc -O2 -fwrapv -ffp-contract=off [-fopenmp] -fPIC -shared
 * -ffp-contract=off is REQUIRED: the reference is built without FMA
 * contraction (mpi/Makefile:2,13 has no -O and no -march), and the only fused
 * operations allowed here are the two explicit fmaf() of ORACLE_F32_FMA.
 * Never -ffast-math.
 *
 * "M:" = /root/reference/mpi/mpi_heat_improved_persistent_stat.c
 * "C:" = /root/reference/cuda/cuda_heat.cu
 */
#include "heat_oracle.h"

#include <math.h>
#include <stdio.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

int oracle_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* M:315-321 / C:274-280.  The product is evaluated in `int` by the reference and
 * wraps for nx,ny >= 432; two's-complement wrap is reproduced with uint32 math. */
void oracle_inidat(int nx, int ny, float *u)
{
    for (int ix = 0; ix < nx; ix++) {
        uint32_t a = (uint32_t)ix * (uint32_t)(nx - ix - 1);
        for (int iy = 0; iy < ny; iy++) {
            uint32_t p = a * (uint32_t)iy * (uint32_t)(ny - iy - 1);
            u[(int64_t)ix * ny + iy] = (float)(int32_t)p;
        }
    }
}

/* M:326-341 / C:285-300: transposed, iy descending, "%6.1f" separated by one blank. */
int oracle_prtdat(int nx, int ny, const float *u, const char *fnam)
{
    FILE *fp = fopen(fnam, "w");
    if (!fp) return -1;
    for (int iy = ny - 1; iy >= 0; iy--)
        for (int ix = 0; ix < nx; ix++)
            fprintf(fp, ix != nx - 1 ? "%6.1f " : "%6.1f\n", u[(int64_t)ix * ny + iy]);
    return fclose(fp) ? -1 : 0;
}

/* C:155-161 built with -fmad=false: every operation rounded to float. */
static inline float upd_f32_nofma(float c, float n, float s, float e, float w, float cx, float cy)
{
    float t1 = (s + n) - 2.0f * c;
    float t2 = (e + w) - 2.0f * c;
    float a = cx * t1;
    float r = c + a;
    float b = cy * t2;
    return r + b;
}

/* C:155-161 as nvcc contracts it by default: c + cx*t1 -> FFMA, (..) + cy*t2 -> FFMA.
 * (2.0f*c is exact, so FFMA(-2,c,n+s) and FADD(n+s,-(c+c)) give the same bits.) */
static inline float upd_f32_fma(float c, float n, float s, float e, float w, float cx, float cy)
{
    float t1 = (s + n) - 2.0f * c;
    float t2 = (e + w) - 2.0f * c;
    return fmaf(cy, t2, fmaf(cx, t1, c));
}

/* M:168-174: the two neighbour sums are float adds; the literal 2.0 promotes the
 * rest to double; parms.cx/cy are floats promoted to double; one final rounding. */
static inline float upd_f64mix(float c, float n, float s, float e, float w, float cx, float cy)
{
    float ns = s + n;
    float ew = e + w;
    double r = (double)c + (double)cx * ((double)ns - 2.0 * (double)c)
                         + (double)cy * ((double)ew - 2.0 * (double)c);
    return (float)r;
}

/* Interior update; guards M:187 / C:57 (ring never written). */
void oracle_step(int arith, int nx, int ny, float cx, float cy, const float *in, float *out)
{
#define ROWLOOP(UPD)                                                              \
    _Pragma("omp parallel for schedule(static)")                                  \
    for (int i = 1; i < nx - 1; i++) {                                            \
        const float *c = in + (int64_t)i * ny;                                    \
        const float *n = c - ny, *s = c + ny;                                     \
        float *o = out + (int64_t)i * ny;                                         \
        for (int j = 1; j < ny - 1; j++)                                          \
            o[j] = UPD(c[j], n[j], s[j], c[j + 1], c[j - 1], cx, cy);             \
    }
    switch (arith) {
    case ORACLE_F32_FMA:    ROWLOOP(upd_f32_fma)   break;
    case ORACLE_F32_NOFMA:  ROWLOOP(upd_f32_nofma) break;
    default:                ROWLOOP(upd_f64mix)    break;
    }
#undef ROWLOOP
}

float oracle_max_abs_diff(int64_t n, const float *a, const float *b)
{
    float m = 0.0f;
#pragma omp parallel for reduction(max : m) schedule(static)
    for (int64_t k = 0; k < n; k++) {
        float d = fabsf(a[k] - b[k]);
        if (d > m) m = d;
    }
    return m;
}

int oracle_solve(int arith, int nx, int ny, int steps, float cx, float cy,
                 int conv_every, float conv_eps, int conv_phase, int step_convention,
                 float *grid, float *work, int *updates_done, int *converged)
{
    if (nx < 1 || ny < 1 || steps < 0 || !grid || !work) return -1;
    const int64_t n = (int64_t)nx * ny;
    /* both buffers start with the same ring: C:197-198 copies the field to both;
     * in M the second buffer's ring is static zero == inidat's ring. */
    memcpy(work, grid, (size_t)n * sizeof(float));
    float *buf[2] = { grid, work };
    int old = 0, done = 0, conv = 0;
    const int total = steps + (step_convention ? 1 : 0); /* M:159 vs C:204 */
    for (int it = 0; it < total; it++) {
        oracle_step(arith, nx, ny, cx, cy, buf[old], buf[1 - old]);
        old = 1 - old;                                   /* M:264 / C:211,217 */
        done = it + 1;
        if (conv_every > 0) {
            int check = conv_phase ? (it % conv_every == 0)          /* C:219 */
                                   : ((it + 1) % conv_every == 0);   /* M:157,236-238 */
            if (check && oracle_max_abs_diff(n, buf[0], buf[1]) < conv_eps) { /* M:245 / C:67 */
                conv = 1;
                break;
            }
        }
    }
    if (old == 1) memcpy(grid, work, (size_t)n * sizeof(float));
    if (updates_done) *updates_done = done;
    if (converged) *converged = conv;
    return 0;
}

/* ---------------------------------------------------------------------------------------------------------
 * Double-precision grids.  PARITY UNPINNED: the reference hard-codes `float u` (M:46,
 * C:170), so there is no reference build or golden vector for this mode; it is the reference's expression
 * (M:168-174 / C:155-161) restated with `double u` -- every value and operation in double, parms.cx/cy still
 * floats (M:29-32) promoted.  fused = 0: every operation rounded (what gcc makes of the MPI source);
 * fused = 1: the two multiply-adds contracted, as nvcc does for the CUDA source. */
void oracle_inidat_f64(int nx, int ny, double *u)
{
    for (int ix = 0; ix < nx; ix++) {
        uint32_t a = (uint32_t)ix * (uint32_t)(nx - ix - 1);
        for (int iy = 0; iy < ny; iy++)
            u[(int64_t)ix * ny + iy] = (double)(int32_t)(a * (uint32_t)iy * (uint32_t)(ny - iy - 1)); /* int product, M:320 */
    }
}

void oracle_step_f64(int fused, int nx, int ny, float cxf, float cyf, const double *in, double *out)
{
    const double cx = (double)cxf, cy = (double)cyf;
#pragma omp parallel for schedule(static)
    for (int i = 1; i < nx - 1; i++) {
        const double *c = in + (int64_t)i * ny, *n = c - ny, *s = c + ny;
        double *o = out + (int64_t)i * ny;
        for (int j = 1; j < ny - 1; j++) {
            const double t1 = (s[j] + n[j]) - 2.0 * c[j];
            const double t2 = (c[j + 1] + c[j - 1]) - 2.0 * c[j];
            if (fused) {
                o[j] = fma(cy, t2, fma(cx, t1, c[j]));
            } else {
                const double a = cx * t1;
                const double r = c[j] + a;
                const double b = cy * t2;
                o[j] = r + b;
            }
        }
    }
}

int oracle_solve_f64(int fused, int nx, int ny, int steps, float cx, float cy, int conv_every, float conv_eps, int conv_phase,
                     int step_convention, double *grid, double *work, int *updates_done, int *converged)
{
    if (nx < 1 || ny < 1 || steps < 0 || !grid || !work) return -1;
    const int64_t n = (int64_t)nx * ny;
    memcpy(work, grid, (size_t)n * sizeof(double));
    double *buf[2] = {grid, work};
    int old = 0, done = 0, conv = 0;
    const int total = steps + (step_convention ? 1 : 0);
    for (int it = 0; it < total; it++) {
        oracle_step_f64(fused, nx, ny, cx, cy, buf[old], buf[1 - old]);
        old = 1 - old;
        done = it + 1;
        if (conv_every > 0) {
            int check = conv_phase ? (it % conv_every == 0) : ((it + 1) % conv_every == 0);
            if (check) {
                double m = 0.0;
                for (int64_t k = 0; k < n; k++) {
                    double d = fabs(buf[0][k] - buf[1][k]);
                    if (d > m) m = d;
                }
                if (m < (double)conv_eps) {
                    conv = 1;
                    break;
                }
            }
        }
    }
    if (old == 1) memcpy(grid, work, (size_t)n * sizeof(double));
    if (updates_done) *updates_done = done;
    if (converged) *converged = conv;
    return 0;
}

/* ---------------------------------------------------------------------------------------------------------
 * 3-D, 7-point.  PARITY UNPINNED: the reference is 2-D only.  Its expression (M:168-174 /
 * C:155-161) with a third neighbour pair, evaluated left to right: u + cx*((X+ + X-) - 2u) + cy*((Y+ + Y-) - 2u)
 * + cz*((Z+ + Z-) - 2u), in the same three arithmetics.  u[ix][iy][iz], iz fastest; all six faces are the ring. */
void oracle_inidat3d(int nx, int ny, int nz, float *u)
{
    for (int ix = 0; ix < nx; ix++)
        for (int iy = 0; iy < ny; iy++) {
            uint32_t a = (uint32_t)ix * (uint32_t)(nx - ix - 1) * (uint32_t)iy * (uint32_t)(ny - iy - 1);
            for (int iz = 0; iz < nz; iz++)
                u[((int64_t)ix * ny + iy) * nz + iz] = (float)(int32_t)(a * (uint32_t)iz * (uint32_t)(nz - iz - 1));
        }
}

static inline float upd3_oracle(int arith, float c, float xs, float ys, float zs, float cx, float cy, float cz)
{
    if (arith == ORACLE_F32_FMA)
        return fmaf(cz, zs - 2.0f * c, fmaf(cy, ys - 2.0f * c, fmaf(cx, xs - 2.0f * c, c)));
    if (arith == ORACLE_F32_NOFMA) {
        float a = cx * (xs - 2.0f * c);
        float r = c + a;
        float b = cy * (ys - 2.0f * c);
        r = r + b;
        float d = cz * (zs - 2.0f * c);
        return r + d;
    }
    double cd = (double)c;
    double r = cd + (double)cx * ((double)xs - 2.0 * cd);
    r = r + (double)cy * ((double)ys - 2.0 * cd);
    r = r + (double)cz * ((double)zs - 2.0 * cd);
    return (float)r;
}

void oracle_step3d(int arith, int nx, int ny, int nz, float cx, float cy, float cz, const float *in, float *out)
{
    const int64_t sx = (int64_t)ny * nz, sy = nz;
#pragma omp parallel for schedule(static)
    for (int i = 1; i < nx - 1; i++)
        for (int j = 1; j < ny - 1; j++) {
            const float *c = in + i * sx + j * sy;
            float *o = out + i * sx + j * sy;
            for (int k = 1; k < nz - 1; k++) {
                float xs = c[k + sx] + c[k - sx];
                float ys = c[k + sy] + c[k - sy];
                float zs = c[k + 1] + c[k - 1];
                o[k] = upd3_oracle(arith, c[k], xs, ys, zs, cx, cy, cz);
            }
        }
}

int oracle_solve3d(int arith, int nx, int ny, int nz, int steps, float cx, float cy, float cz, int conv_every, float conv_eps,
                   int conv_phase, int step_convention, float *grid, float *work, int *updates_done, int *converged)
{
    if (nx < 1 || ny < 1 || nz < 1 || steps < 0 || !grid || !work) return -1;
    const int64_t n = (int64_t)nx * ny * nz;
    memcpy(work, grid, (size_t)n * sizeof(float));
    float *buf[2] = {grid, work};
    int old = 0, done = 0, conv = 0;
    const int total = steps + (step_convention ? 1 : 0);
    for (int it = 0; it < total; it++) {
        oracle_step3d(arith, nx, ny, nz, cx, cy, cz, buf[old], buf[1 - old]);
        old = 1 - old;
        done = it + 1;
        if (conv_every > 0) {
            int check = conv_phase ? (it % conv_every == 0) : ((it + 1) % conv_every == 0);
            if (check && oracle_max_abs_diff(n, buf[0], buf[1]) < conv_eps) {
                conv = 1;
                break;
            }
        }
    }
    if (old == 1) memcpy(grid, work, (size_t)n * sizeof(float));
    if (updates_done) *updates_done = done;
    if (converged) *converged = conv;
    return 0;
}
