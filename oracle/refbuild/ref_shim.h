/*
 * ref_shim.h -- force-included (-include) in front of the UNMODIFIED reference
 * sources by oracle/build_ref.py.  TEST INFRASTRUCTURE, written for this repo.
 *
 * It changes nothing the reference computes.  It only
 *   - lets STEPS / STEP be chosen at run time (-DSTEPS=heat_ref_steps_v, read
 *     from $HEAT_REF_STEPS / $HEAT_REF_STEP), so one binary per grid size is
 *     enough; NXPROB/NYPROB stay compile-time constants as in the reference;
 *   - routes prtdat's fopen/fprintf/fclose (M:326-341, C:285-300) through
 *     ref_shim.c, which can keep the text file ($HEAT_REF_TEXT=1), store the
 *     exact float bits of every "%6.1f" value ($HEAT_REF_BIN=1 -> <name>.f32,
 *     in prtdat order) or drop the dump altogether (timing runs).
 */
#ifndef HEAT_REF_SHIM_H
#define HEAT_REF_SHIM_H

#include <stdio.h>
#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifdef __cplusplus
extern "C" {
#endif
extern int heat_ref_steps_v;
extern int heat_ref_step_v;
FILE *heat_ref_fopen(const char *name, const char *mode);
int heat_ref_fprintf(FILE *fp, const char *fmt, ...);
int heat_ref_fclose(FILE *fp);
#ifdef __cplusplus
}
#endif

#define fopen heat_ref_fopen
#define fprintf heat_ref_fprintf
#define fclose heat_ref_fclose

#endif
