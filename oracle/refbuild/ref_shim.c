/*
 * ref_shim.c -- run-time side of ref_shim.h and of mpi_stub/mpi.h.
 * TEST INFRASTRUCTURE, written for this repo; linked next to the unmodified
 * reference object by oracle/build_ref.py.
 */
#define _DEFAULT_SOURCE
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

/* ---------------- MPI stand-in: 1 rank, or N ranks as forked processes ----------------
 * $HEAT_REF_RANKS = N (default 1).  N > 1: MPI_Init forks N-1 children; every ordered pair of ranks
 * gets a byte ring in one shared anonymous mapping (created before the fork), through which
 * MPI_Send / persistent sends push [tag, bytes, packed payload] and MPI_Recv / MPI_Waitall pull
 * them in order.  That is enough for this program: every message names its source, and one
 * ordered pair never carries two interleaved streams (BEGIN rows, then UPDT halos, then END rows).
 * It is a shim, NOT a real MPI: no network, no unexpected-message queue, no wildcards -- but the
 * reference's own decomposition (MPI_Dims_create / MPI_Cart_*), its row_type / col_type halos,
 * persistent requests and the LAND allreduce run exactly as written. */
#ifndef HEAT_REF_NO_MPI
#include "mpi_stub/mpi.h"
#include <sched.h>
#include <stdatomic.h>
#include <stdint.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

static double now(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}
static double t_loop0 = 0, t_loop1 = 0;

#define MAX_RANKS 64
#define RING_BYTES ((uint64_t)2 << 20) /* per ordered pair (the largest message is one row or column); virtual until touched */

typedef struct {
    _Atomic uint64_t head; /* bytes ever written (producer) */
    char pad0[56];
    _Atomic uint64_t tail; /* bytes ever read (consumer) */
    char pad1[56];
} ring_ctl;

typedef struct {
    _Atomic int arrived;
    _Atomic int sense;
    char pad[56];
    int slot[MAX_RANKS]; /* allreduce operands */
    ring_ctl ctl[MAX_RANKS * MAX_RANKS];
} shared_area;

static shared_area *g_sh;
static char *g_rings;
static int g_rank = 0, g_size = 1, g_dims[2] = {1, 1}, g_sense = 0;
static pid_t g_child[MAX_RANKS], g_master;

static void die(const char *what)
{
    fprintf(stderr, "mpi stand-in (rank %d of %d): %s\n", g_rank, g_size, what);
    abort();
}

/* a rank that waits keeps an eye on the others, so that a crashed rank ends the run instead of hanging it */
static int g_reaped[MAX_RANKS];
static void check_ranks(void)
{
    if (g_rank == 0) {
        for (int r = 1; r < g_size; r++) {
            int st = 0;
            if (g_reaped[r] || waitpid(g_child[r], &st, WNOHANG) <= 0) continue;
            if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) die("a worker rank died");
            g_reaped[r] = 1;
        }
    } else if (getppid() != g_master) {
        _exit(3); /* rank 0 is gone (this process was re-parented) */
    }
}

static void relax(unsigned *spins)
{
    if (++*spins > 64) sched_yield();
    if ((*spins & 0xffff) == 0) check_ranks();
}

/* ---- datatypes: count blocks of blocklen floats, stride floats apart ---- */
typedef struct { int count, blocklen, stride; } dtype;
static dtype g_dt[64];
static int g_ndt = 0;
#define DT_BASE 16

static dtype dt_of(MPI_Datatype t)
{
    if (t == MPI_FLOAT || t == MPI_INT) { dtype d = {1, 1, 1}; return d; }
    if (t < DT_BASE || t >= DT_BASE + g_ndt) die("unknown datatype");
    return g_dt[t - DT_BASE];
}
static MPI_Datatype dt_new(int count, int blocklen, int stride, MPI_Datatype old)
{
    if (old != MPI_FLOAT) die("derived datatypes of MPI_FLOAT only");
    if (g_ndt == 64) die("too many datatypes");
    g_dt[g_ndt].count = count, g_dt[g_ndt].blocklen = blocklen, g_dt[g_ndt].stride = stride;
    return DT_BASE + g_ndt++;
}

/* ---- rings ---- */
static ring_ctl *ctl_of(int src, int dst) { return &g_sh->ctl[src * g_size + dst]; }
static char *ring_of(int src, int dst) { return g_rings + (uint64_t)(src * g_size + dst) * RING_BYTES; }

static void ring_put(char *ring, uint64_t *pos, const void *data, uint64_t n)
{
    const uint64_t off = *pos % RING_BYTES, first = n < RING_BYTES - off ? n : RING_BYTES - off;
    memcpy(ring + off, data, first);
    memcpy(ring, (const char *)data + first, n - first);
    *pos += n;
}
static void ring_get(const char *ring, uint64_t *pos, void *data, uint64_t n)
{
    const uint64_t off = *pos % RING_BYTES, first = n < RING_BYTES - off ? n : RING_BYTES - off;
    memcpy(data, ring + off, first);
    memcpy((char *)data + first, ring, n - first);
    *pos += n;
}

static void do_send(const void *buf, int count, MPI_Datatype t, int dest, int tag)
{
    if (dest == MPI_PROC_NULL) return;
    if (dest < 0 || dest >= g_size || dest == g_rank) die("send to a bad rank");
    if (count != 1) die("count != 1"); /* more would need the type's extent: the program always sends 1 */
    const dtype d = dt_of(t);
    const uint64_t bytes = (uint64_t)count * d.count * d.blocklen * 4, need = 8 + bytes;
    if (need > RING_BYTES) die("message larger than the ring");
    ring_ctl *c = ctl_of(g_rank, dest);
    char *ring = ring_of(g_rank, dest);
    uint64_t head = atomic_load_explicit(&c->head, memory_order_relaxed);
    unsigned spins = 0;
    while (head + need - atomic_load_explicit(&c->tail, memory_order_acquire) > RING_BYTES) relax(&spins);
    const int32_t hdr[2] = {tag, (int32_t)bytes};
    ring_put(ring, &head, hdr, 8);
    const float *f = (const float *)buf;
    for (int b = 0; b < d.count; b++) ring_put(ring, &head, f + (size_t)b * d.stride, (uint64_t)d.blocklen * 4);
    atomic_store_explicit(&c->head, head, memory_order_release);
}

static void do_recv(void *buf, int count, MPI_Datatype t, int src, int tag)
{
    if (src == MPI_PROC_NULL) return;
    if (src < 0 || src >= g_size || src == g_rank) die("receive from a bad rank");
    if (count != 1) die("count != 1");
    const dtype d = dt_of(t);
    const uint64_t bytes = (uint64_t)d.count * d.blocklen * 4;
    ring_ctl *c = ctl_of(src, g_rank);
    const char *ring = ring_of(src, g_rank);
    uint64_t tail = atomic_load_explicit(&c->tail, memory_order_relaxed);
    unsigned spins = 0;
    /* the producer publishes header and payload with one store of head */
    while (atomic_load_explicit(&c->head, memory_order_acquire) - tail < 8 + bytes) relax(&spins);
    int32_t hdr[2];
    ring_get(ring, &tail, hdr, 8);
    if (hdr[0] != tag || (uint64_t)hdr[1] != bytes) die("message does not match the receive (tag or size)");
    float *f = (float *)buf;
    for (int b = 0; b < d.count; b++) ring_get(ring, &tail, f + (size_t)b * d.stride, (uint64_t)d.blocklen * 4);
    atomic_store_explicit(&c->tail, tail, memory_order_release);
}

static void barrier(void)
{
    if (g_size == 1) return;
    g_sense = !g_sense;
    if (atomic_fetch_add(&g_sh->arrived, 1) == g_size - 1) {
        atomic_store(&g_sh->arrived, 0);
        atomic_store(&g_sh->sense, g_sense);
    } else {
        unsigned spins = 0;
        while (atomic_load(&g_sh->sense) != g_sense) relax(&spins);
    }
}

/* ---- the 20 entry points ---- */
int MPI_Init(int *argc, char ***argv)
{
    (void)argc; (void)argv;
    g_size = env_int("HEAT_REF_RANKS", 1);
    if (g_size < 1 || g_size > MAX_RANKS) die("HEAT_REF_RANKS out of range");
    if (g_size == 1) return MPI_SUCCESS;
    const size_t bytes = sizeof(shared_area) + (size_t)g_size * g_size * RING_BYTES;
    void *m = mmap(NULL, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (m == MAP_FAILED) die("mmap of the shared area failed");
    g_sh = (shared_area *)m; /* zero-filled: counters, rings and the barrier start at 0 */
    g_rings = (char *)m + sizeof(shared_area);
    fflush(NULL);
    g_master = getpid();
    for (int r = 1; r < g_size; r++) {
        const pid_t pid = fork();
        if (pid < 0) die("fork failed");
        if (pid == 0) {
            g_rank = r;
            break;
        }
        g_child[r] = pid;
    }
    return MPI_SUCCESS;
}

int MPI_Finalize(void)
{
    barrier();
    if (g_rank != 0) {
        fflush(NULL);
        _exit(0); /* a worker has nothing left to do (M:308 is its last statement) */
    }
    for (int r = 1; r < g_size; r++) {
        int st = 0;
        if (g_reaped[r]) continue;
        if (waitpid(g_child[r], &st, 0) < 0 || !WIFEXITED(st) || WEXITSTATUS(st) != 0) die("a worker rank failed");
    }
    /* loop-only time: first MPI_Startall (M:160) .. first MPI_Request_free (M:266), on rank 0 */
    if (t_loop0 > 0 && t_loop1 > 0)
        printf("HEAT_REF_LOOP_SECS %.6f\n", t_loop1 - t_loop0);
    return MPI_SUCCESS;
}
int MPI_Comm_size(MPI_Comm c, int *size) { (void)c; *size = g_size; return MPI_SUCCESS; }
int MPI_Comm_rank(MPI_Comm c, int *rank) { (void)c; *rank = g_rank; return MPI_SUCCESS; }
int MPI_Dims_create(int nnodes, int ndims, int *dims)
{
    /* the balanced factorisation, larger factor first, as MPI libraries return it for two free dimensions */
    if (ndims != 2 || dims[0] != 0 || dims[1] != 0) die("MPI_Dims_create: two free dimensions only");
    int b = 1;
    for (int f = 1; f * f <= nnodes; f++)
        if (nnodes % f == 0) b = f;
    dims[0] = nnodes / b;
    dims[1] = b;
    return MPI_SUCCESS;
}
int MPI_Cart_create(MPI_Comm c, int ndims, const int *dims, const int *periods, int reorder, MPI_Comm *out)
{
    (void)reorder;
    if (ndims != 2 || periods[0] || periods[1] || dims[0] * dims[1] != g_size) die("MPI_Cart_create: 2-D, non-periodic, all ranks");
    g_dims[0] = dims[0], g_dims[1] = dims[1];
    *out = c;
    return MPI_SUCCESS;
}
int MPI_Cart_coords(MPI_Comm c, int rank, int maxdims, int *coords)
{
    (void)c;
    if (maxdims < 2) die("MPI_Cart_coords: maxdims");
    coords[0] = rank / g_dims[1]; /* row-major rank order, as the standard defines it */
    coords[1] = rank % g_dims[1];
    return MPI_SUCCESS;
}
int MPI_Cart_shift(MPI_Comm c, int dir, int disp, int *src, int *dst)
{
    (void)c;
    int me[2] = {g_rank / g_dims[1], g_rank % g_dims[1]};
    int lo[2] = {me[0], me[1]}, hi[2] = {me[0], me[1]};
    lo[dir] -= disp;
    hi[dir] += disp;
    *src = (lo[dir] < 0 || lo[dir] >= g_dims[dir]) ? MPI_PROC_NULL : lo[0] * g_dims[1] + lo[1];
    *dst = (hi[dir] < 0 || hi[dir] >= g_dims[dir]) ? MPI_PROC_NULL : hi[0] * g_dims[1] + hi[1];
    return MPI_SUCCESS;
}
int MPI_Type_contiguous(int count, MPI_Datatype old, MPI_Datatype *newt) { *newt = dt_new(1, count, count, old); return MPI_SUCCESS; }
int MPI_Type_vector(int count, int blocklen, int stride, MPI_Datatype old, MPI_Datatype *newt)
{
    *newt = dt_new(count, blocklen, stride, old);
    return MPI_SUCCESS;
}
int MPI_Type_commit(MPI_Datatype *t) { (void)t; return MPI_SUCCESS; }
double MPI_Wtime(void) { return now(); }
int MPI_Send(const void *b, int n, MPI_Datatype t, int dest, int tag, MPI_Comm c) { (void)c; do_send(b, n, t, dest, tag); return MPI_SUCCESS; }
int MPI_Recv(void *b, int n, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Status *s)
{
    (void)c;
    do_recv(b, n, t, src, tag);
    if (s) s->MPI_SOURCE = src, s->MPI_TAG = tag, s->MPI_ERROR = MPI_SUCCESS;
    return MPI_SUCCESS;
}

/* persistent requests: a send is performed when it is started (its buffer is read then, which is what
 * the standard allows a started send to do), a receive when it is waited for */
typedef struct { int is_recv, active; void *buf; int count; MPI_Datatype t; int peer, tag; } preq;
static preq g_req[256];
static int g_nreq = 0;
static MPI_Request req_new(int is_recv, const void *b, int n, MPI_Datatype t, int peer, int tag)
{
    if (g_nreq == 256) die("too many persistent requests");
    preq *q = &g_req[g_nreq];
    q->is_recv = is_recv, q->active = 0, q->buf = (void *)b, q->count = n, q->t = t, q->peer = peer, q->tag = tag;
    return ++g_nreq; /* 0 stays the null request */
}
int MPI_Send_init(const void *b, int n, MPI_Datatype t, int dest, int tag, MPI_Comm c, MPI_Request *r) { (void)c; *r = req_new(0, b, n, t, dest, tag); return MPI_SUCCESS; }
int MPI_Recv_init(void *b, int n, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Request *r) { (void)c; *r = req_new(1, b, n, t, src, tag); return MPI_SUCCESS; }
int MPI_Startall(int count, MPI_Request *reqs)
{
    if (t_loop0 == 0) t_loop0 = now();
    for (int i = 0; i < count; i++) {
        if (reqs[i] < 1 || reqs[i] > g_nreq) die("MPI_Startall: bad request");
        preq *q = &g_req[reqs[i] - 1];
        if (q->active) die("MPI_Startall: request already active");
        q->active = 1;
        if (!q->is_recv) do_send(q->buf, q->count, q->t, q->peer, q->tag);
    }
    return MPI_SUCCESS;
}
int MPI_Waitall(int count, MPI_Request *reqs, MPI_Status *sts)
{
    (void)sts;
    for (int i = 0; i < count; i++) {
        if (reqs[i] < 1 || reqs[i] > g_nreq) die("MPI_Waitall: bad request");
        preq *q = &g_req[reqs[i] - 1];
        if (!q->active) continue;
        if (q->is_recv) do_recv(q->buf, q->count, q->t, q->peer, q->tag);
        q->active = 0;
    }
    return MPI_SUCCESS;
}
int MPI_Request_free(MPI_Request *r)
{
    (void)r;
    if (t_loop1 == 0) t_loop1 = now();
    return MPI_SUCCESS;
}
int MPI_Allreduce(const void *in, void *out, int count, MPI_Datatype t, MPI_Op op, MPI_Comm c)
{
    (void)c;
    if (g_size == 1) {
        memcpy(out, in, (size_t)count * (t == MPI_FLOAT ? sizeof(float) : sizeof(int)));
        return MPI_SUCCESS;
    }
    if (t != MPI_INT || op != MPI_LAND || count != 1) die("MPI_Allreduce: one MPI_INT with MPI_LAND only");
    g_sh->slot[g_rank] = *(const int *)in;
    barrier();
    int all = 1;
    for (int r = 0; r < g_size; r++) all = all && g_sh->slot[r];
    barrier(); /* nobody overwrites a slot before everybody has read it */
    *(int *)out = all;
    return MPI_SUCCESS;
}
#endif
