/*
 * ref_shim.c -- run-time side of ref_shim.h and of mpi_stub/mpi.h.
 * TEST INFRASTRUCTURE, written for this repo; linked next to the unmodified
 * reference object by oracle/build_ref.py.
 */
#define _POSIX_C_SOURCE 200809L
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* build_ref.py force-includes ref_shim.h into this file too: here the real libc calls are wanted */
#undef fopen
#undef fprintf
#undef fclose

int heat_ref_steps_v = 100;
int heat_ref_step_v = 20;
static int g_text = 0, g_bin = 0;

static int env_int(const char *k, int dflt)
{
    const char *v = getenv(k);
    return v && *v ? atoi(v) : dflt;
}

__attribute__((constructor)) static void heat_ref_init(void)
{
    heat_ref_steps_v = env_int("HEAT_REF_STEPS", 100);
    heat_ref_step_v = env_int("HEAT_REF_STEP", 20);
    g_text = env_int("HEAT_REF_TEXT", 0);
    g_bin = env_int("HEAT_REF_BIN", 0);
}

/* ---------------- prtdat capture ---------------- */
typedef struct {
    FILE *fp;
    char name[512];
    float *v;
    size_t n, cap;
} capture;
static capture g_cap[8];
static FILE *g_last_capture_fp; /* the dump being written: prtdat calls fprintf 2*nx*ny times */

static capture *find_capture(FILE *fp)
{
    for (int i = 0; i < 8; i++)
        if (g_cap[i].fp == fp && fp) return &g_cap[i];
    return NULL;
}

FILE *heat_ref_fopen(const char *name, const char *mode)
{
    FILE *fp = fopen(g_text ? name : "/dev/null", mode);
    if (!fp) return NULL;
    for (int i = 0; i < 8; i++)
        if (!g_cap[i].fp) {
            g_cap[i].fp = fp;
            snprintf(g_cap[i].name, sizeof g_cap[i].name, "%s", name);
            g_cap[i].v = NULL;
            g_cap[i].n = g_cap[i].cap = 0;
            g_last_capture_fp = fp;
            break;
        }
    return fp;
}

int heat_ref_fprintf(FILE *fp, const char *fmt, ...)
{
    va_list ap;
    int r = 0;
    if (fp == g_last_capture_fp && !g_bin && !g_text) return 0; /* timing runs: the dump is dropped */
    capture *c = find_capture(fp);
    if (!c) { /* not one of prtdat's files: plain fprintf */
        va_start(ap, fmt);
        r = vfprintf(fp, fmt, ap);
        va_end(ap);
        return r;
    }
    if (g_bin && strcmp(fmt, "%6.1f") == 0) {
        va_start(ap, fmt);
        double d = va_arg(ap, double); /* float promoted by the call: lossless */
        va_end(ap);
        if (c->n == c->cap) {
            c->cap = c->cap ? c->cap * 2 : (1u << 16);
            c->v = (float *)realloc(c->v, c->cap * sizeof(float));
            if (!c->v) abort();
        }
        c->v[c->n++] = (float)d;
    }
    if (g_text) {
        va_start(ap, fmt);
        r = vfprintf(fp, fmt, ap);
        va_end(ap);
    }
    return r;
}

int heat_ref_fclose(FILE *fp)
{
    capture *c = find_capture(fp);
    if (fp == g_last_capture_fp) g_last_capture_fp = NULL;
    if (c) {
        if (g_bin) {
            char out[600];
            snprintf(out, sizeof out, "%s.f32", c->name);
            FILE *b = fopen(out, "wb");
            if (b) {
                fwrite(c->v, sizeof(float), c->n, b);
                fclose(b);
            }
        }
        free(c->v);
        memset(c, 0, sizeof *c);
    }
    return fclose(fp);
}

/* ---------------- single-rank MPI ---------------- */
#ifndef HEAT_REF_NO_MPI
#include "mpi_stub/mpi.h"

static double now(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}
static double t_loop0 = 0, t_loop1 = 0;

int MPI_Init(int *argc, char ***argv) { (void)argc; (void)argv; return MPI_SUCCESS; }
int MPI_Finalize(void)
{
    /* loop-only time: first MPI_Startall (M:160) .. first MPI_Request_free (M:266) */
    if (t_loop0 > 0 && t_loop1 > 0)
        printf("HEAT_REF_LOOP_SECS %.6f\n", t_loop1 - t_loop0);
    return MPI_SUCCESS;
}
int MPI_Comm_size(MPI_Comm c, int *size) { (void)c; *size = 1; return MPI_SUCCESS; }
int MPI_Comm_rank(MPI_Comm c, int *rank) { (void)c; *rank = 0; return MPI_SUCCESS; }
int MPI_Dims_create(int nnodes, int ndims, int *dims)
{
    (void)nnodes;
    for (int i = 0; i < ndims; i++) dims[i] = 1;
    return MPI_SUCCESS;
}
int MPI_Cart_create(MPI_Comm c, int ndims, const int *dims, const int *periods, int reorder, MPI_Comm *out)
{
    (void)ndims; (void)dims; (void)periods; (void)reorder;
    *out = c;
    return MPI_SUCCESS;
}
int MPI_Cart_coords(MPI_Comm c, int rank, int maxdims, int *coords)
{
    (void)c; (void)rank;
    for (int i = 0; i < maxdims; i++) coords[i] = 0;
    return MPI_SUCCESS;
}
int MPI_Cart_shift(MPI_Comm c, int dir, int disp, int *src, int *dst)
{
    (void)c; (void)dir; (void)disp;
    *src = *dst = MPI_PROC_NULL; /* non-periodic 1x1 grid: no neighbours */
    return MPI_SUCCESS;
}
int MPI_Type_contiguous(int count, MPI_Datatype old, MPI_Datatype *newt) { (void)count; *newt = old; return MPI_SUCCESS; }
int MPI_Type_vector(int count, int blocklen, int stride, MPI_Datatype old, MPI_Datatype *newt)
{
    (void)count; (void)blocklen; (void)stride;
    *newt = old;
    return MPI_SUCCESS;
}
int MPI_Type_commit(MPI_Datatype *t) { (void)t; return MPI_SUCCESS; }
double MPI_Wtime(void) { return now(); }
static int only_null(int peer)
{
    if (peer != MPI_PROC_NULL) {
        fputs("mpi stub: message to a real rank in a 1-rank run\n", stderr);
        abort();
    }
    return MPI_SUCCESS;
}
int MPI_Send(const void *b, int n, MPI_Datatype t, int dest, int tag, MPI_Comm c) { (void)b; (void)n; (void)t; (void)tag; (void)c; return only_null(dest); }
int MPI_Recv(void *b, int n, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Status *s) { (void)b; (void)n; (void)t; (void)tag; (void)c; (void)s; return only_null(src); }
int MPI_Send_init(const void *b, int n, MPI_Datatype t, int dest, int tag, MPI_Comm c, MPI_Request *r) { (void)b; (void)n; (void)t; (void)tag; (void)c; *r = 0; return only_null(dest); }
int MPI_Recv_init(void *b, int n, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Request *r) { (void)b; (void)n; (void)t; (void)tag; (void)c; *r = 0; return only_null(src); }
int MPI_Startall(int count, MPI_Request *reqs)
{
    (void)count; (void)reqs;
    if (t_loop0 == 0) t_loop0 = now();
    return MPI_SUCCESS;
}
int MPI_Waitall(int count, MPI_Request *reqs, MPI_Status *sts) { (void)count; (void)reqs; (void)sts; return MPI_SUCCESS; }
int MPI_Request_free(MPI_Request *r)
{
    (void)r;
    if (t_loop1 == 0) t_loop1 = now();
    return MPI_SUCCESS;
}
int MPI_Allreduce(const void *in, void *out, int count, MPI_Datatype t, MPI_Op op, MPI_Comm c)
{
    (void)op; (void)c;
    memcpy(out, in, (size_t)count * (t == MPI_FLOAT ? sizeof(float) : sizeof(int)));
    return MPI_SUCCESS;
}
#endif
