/*
 * heat_oracle.h -- CPU restatement of the reference solver loop.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may build, load or call it, and only as the checker
 * or the timed CPU baseline.  The product (libheat_b200.so) never links it.
 *
 * Parity pin: this restatement is checked bit-for-bit against the UNMODIFIED
 * reference sources compiled from /root/reference by oracle/build_ref.py
 * (oracle/_ref/), see tests/golden/make_golden.py and tests/test_oracle.py.
 */
#ifndef HEAT_ORACLE_H
#define HEAT_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Arithmetic modes (same numbering as include/heat.h). */
enum {
    ORACLE_F32_FMA    = 0, /* cuda_heat.cu:155-161 as nvcc builds it by default (2 FFMA) */
    ORACLE_F32_NOFMA  = 1, /* cuda_heat.cu:155-161 with -fmad=false                      */
    ORACLE_MPI_F64MIX = 2  /* mpi_heat_improved_persistent_stat.c:168-174 (2.0 literal)   */
};

/* mpi_heat_improved_persistent_stat.c:315-321 == cuda_heat.cu:274-280, int32 wrap included. */
void oracle_inidat(int nx, int ny, float *u);

/* mpi_heat_improved_persistent_stat.c:326-341 == cuda_heat.cu:285-300. Returns 0 / -1. */
int oracle_prtdat(int nx, int ny, const float *u, const char *fnam);

/* One Jacobi update of rows 1..nx-2, cols 1..ny-2 of a dense nx*ny array.
 * Ring (row 0, row nx-1, col 0, col ny-1) of `out` is NOT written. */
void oracle_step(int arith, int nx, int ny, float cx, float cy,
                 const float *in, float *out);

/* max |a-b| over the whole array (float subtraction, as both references do). */
float oracle_max_abs_diff(int64_t n, const float *a, const float *b);

/* The whole solver loop.
 *   step_convention: 0 = run `steps` updates (cuda_heat.cu:204),
 *                    1 = run steps+1 updates (mpi_...stat.c:159, `it <= STEPS`)
 *   conv_every K (0 = off), conv_phase: 0 = check after update `it` when
 *   (it+1)%K==0 (mpi:157,236-238), 1 = when it%K==0 (cuda:219)
 *   converged <=> every point |old-new| < conv_eps   (mpi:245 / cuda:67)
 * grid: in = initial field, out = newest field.  work: scratch nx*ny floats.
 * updates_done = number of Jacobi updates executed (both programs print
 * updates_done-1 on convergence). */
int oracle_solve(int arith, int nx, int ny, int steps, float cx, float cy,
                 int conv_every, float conv_eps, int conv_phase, int step_convention,
                 float *grid, float *work, int *updates_done, int *converged);

/* Double-precision grids: PARITY UNPINNED (the reference has no double build; see heat_oracle.c).
 * fused: 0 = every operation rounded, 1 = the two multiply-adds contracted. */
void oracle_inidat_f64(int nx, int ny, double *u);
void oracle_step_f64(int fused, int nx, int ny, float cx, float cy, const double *in, double *out);
int oracle_solve_f64(int fused, int nx, int ny, int steps, float cx, float cy, int conv_every, float conv_eps,
                     int conv_phase, int step_convention, double *grid, double *work, int *updates_done, int *converged);

/* 3-D 7-point grids: PARITY UNPINNED (the reference is 2-D only; see heat_oracle.c). */
void oracle_inidat3d(int nx, int ny, int nz, float *u);
void oracle_step3d(int arith, int nx, int ny, int nz, float cx, float cy, float cz, const float *in, float *out);
int oracle_solve3d(int arith, int nx, int ny, int nz, int steps, float cx, float cy, float cz, int conv_every, float conv_eps,
                   int conv_phase, int step_convention, float *grid, float *work, int *updates_done, int *converged);

/* number of OpenMP threads the oracle will use (1 when built without OpenMP) */
int oracle_threads(void);

#ifdef __cplusplus
}
#endif
#endif
