/*
 * heat.h -- C ABI of libheat_b200.so, the B200 (sm_100a) replacement for the
 * SOLVER LOOP of manospits/parallel_heat.
 *
 * The reference has no library interface: its solver loop is inline in main()
 *     M = mpi/mpi_heat_improved_persistent_stat.c   (loop M:129-269)
 *     C = cuda/cuda_heat.cu                         (loop C:176-242)
 * so every entry point below cites the region of those files it stands for.
 * Plain C, plain pointers and sizes; no CUDA or torch types in any signature
 * (streams and device pointers travel as void*).  Every function returns
 * HEAT_OK (0) or a negative HEAT_E_* code and never calls exit();
 * heat_last_error() gives the text for the calling thread.
 *
 * Grid layout is the reference's: dense row-major u[ix*ny + iy], iy fastest,
 * nx rows (M:320, C:279).  The ring (row 0, row nx-1, col 0, col ny-1) is never
 * updated (guards M:187, C:57).
 */
#ifndef HEAT_B200_H
#define HEAT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HEAT_OK            0
#define HEAT_E_INVALID    -1 /* bad argument                                      */
#define HEAT_E_CUDA       -2 /* a CUDA runtime call failed (see heat_last_error)  */
#define HEAT_E_NOMEM      -3 /* host or device allocation failed                  */
#define HEAT_E_STATE      -4 /* call not valid in the handle's current state      */
#define HEAT_E_PEER       -5 /* neighbour strip missing / peer access unavailable */
#define HEAT_E_TIMEOUT    -6 /* a neighbour strip did not arrive in time          */

/* Arithmetic of one update, selecting which reference build is reproduced bit for bit. */
#define HEAT_F32_FMA     0 /* C:155-161 as nvcc builds it by default: float, 2 FFMA       */
#define HEAT_F32_NOFMA   1 /* C:155-161 with -fmad=false: float, every op rounded         */
#define HEAT_MPI_F64MIX  2 /* M:168-174: float neighbour sums, rest in double (2.0 literal) */

/* Element type of the grid.  The reference hard-codes float; HEAT_GRID_F64 is its update with `double u`
 * (SURVEY 8(f) rank 4): all values and operations in double, cx/cy still the float constants promoted.
 * With a double grid `arith` only chooses the contraction: HEAT_F32_FMA = two fused multiply-adds, as nvcc builds
 * the CUDA source; the other two = every operation rounded, as gcc builds the MPI source.  Single updates on
 * one strip or row strips (temporal_block 1, nranks_y 1); data moves through the *_f64 entry points. */
#define HEAT_GRID_F32    0
#define HEAT_GRID_F64    1

/* 3-D grids (the reference is 2-D only): nz > 1 makes the grid nx*ny*nz, dense, iz fastest,
 * and the update the 7-point one, new = u + cx*((X+ + X-) - 2u) + cy*((Y+ + Y-) - 2u) + cz*((Z+ + Z-) - 2u),
 * evaluated left to right in the arithmetic `arith` names (F32_FMA: three fused multiply-adds); all six faces
 * are the ring.  nz must be a multiple of 4; float, single updates, one strip or strips of ix-planes (nranks_y 1,
 * temporal_block 1).  heat_solve, heat_upload, heat_download, heat_init_inidat, heat_run and the convergence
 * check work as in 2-D on the nx*ny*nz array. */

/* How many updates `steps` means, and when convergence is checked. */
#define HEAT_STEPS_CUDA  0 /* `steps` updates                 (C:204  i <  STEPS) */
#define HEAT_STEPS_MPI   1 /* `steps`+1 updates               (M:159  it <= STEPS) */
#define HEAT_PHASE_MPI   0 /* check after update it when (it+1)%K==0 (M:157,236-238) */
#define HEAT_PHASE_CUDA  1 /* check after update it when  it   %K==0 (C:219)         */

/* Run-time form of the reference's compile-time parameters
 * (NXPROB, NYPROB, STEPS, STEP, CONVERGE: M:7-21, C:7-23; struct Parms M:29-32). */
typedef struct heat_params {
    int   nx, ny;          /* NXPROB, NYPROB: rows, row length                         */
    int   steps;           /* STEPS                                                    */
    float cx, cy;          /* parms.cx (row +-1 term), parms.cy (col +-1 term); 0.1f    */
    int   conv_every;      /* STEP / CHECK_INTERVAL; 0 = built without -DCONVERGE      */
    float conv_eps;        /* converged <=> every |old-new| < conv_eps; 1e-3f (M:245,C:67) */
    int   conv_phase;      /* HEAT_PHASE_*                                             */
    int   step_convention; /* HEAT_STEPS_*                                             */
    int   arith;           /* HEAT_F32_FMA | HEAT_F32_NOFMA | HEAT_MPI_F64MIX          */
    int   temporal_block;  /* updates fused per pass over HBM: 1 to 4                  */
    int   device;          /* CUDA device ordinal of this strip                        */
    int   rank, nranks;    /* row-strip decomposition: this handle owns strip `rank`   */
    int   rank_y, nranks_y;/* column blocks on top of that (2-D blocks, M:51-75); 0, 1 = row strips only */
    int   grid_type;       /* HEAT_GRID_F32 (the reference's `float u`, M:46, C:170) | HEAT_GRID_F64       */
    int   nz;              /* 0 or 1: the reference's 2-D grid; > 1: 3-D grid u[ix][iy][iz], iz fastest    */
    float cz;              /* coefficient of the iz +-1 term of the 3-D update; 0.1f                       */
} heat_params;

typedef struct heat_handle heat_handle; /* opaque: owns device buffers, stream, graphs */

/* Fills *p with the reference's defaults: 20x20, steps 100, cx=cy=0.1, no convergence
 * check, conv_eps 1e-3, F32_FMA, CUDA conventions, 1 strip on device 0 (M:7-32, C:7-23). */
int heat_params_default(heat_params *p);

const char *heat_last_error(void);
/* 1 if a CUDA device is usable by this process, else 0 (never fails). */
int heat_device_count(void);

/* Environment variables read when a handle is created (tuning and debugging; defaults are the measured
 * best): HEAT_ROWS_PER_TILE (K1/K2 tile height, 8), HEAT_ROWS_PER_TILE_TB (K4 tile height, by row length),
 * HEAT_K1_VARIANT (vec = register-streaming K1, bulk = shared-memory staged K1b, auto = bulk for rows of
 * 16384 floats and more; default auto), HEAT_SERPENTINE (alternate sweep direction, 1), HEAT_GRAPHS (CUDA
 * graphs of 16 passes, 1), HEAT_RESIDENT (field resident in shared memory for grids that fit, 1),
 * HEAT_RESIDENT_G (updates per neighbour exchange of that kernel, 6), HEAT_TIMING (heat_solve prints its
 * phases to stderr, 0), HEAT_SOLVE_STREAMED / HEAT_SOLVE_BAND_ROWS (heat_solve's overlapped upload, see
 * there), HEAT_PROFILE_WAIT (count the edge CTAs' flag waits, see heat_debug_wait_stats, 0),
 * HEAT_COLS_FIRST (2-D blocks across processes: M > 0 hands the column-edge CTAs out first, M tile rows
 * each; 0 = grid order, the default -- bit-identical either way and not faster in any layout measured). */

/* ---- the reference's helpers, same arguments (M:315-341, C:274-300) ------------------ */
/* inidat: u[ix*ny+iy] = (float)(ix*(nx-ix-1)*iy*(ny-iy-1)) in wrapping int32 (M:320). */
int heat_inidat(int nx, int ny, float *u);
int heat_inidat_f64(int nx, int ny, double *u); /* same int product, same wrap, stored as double */
/* The 3-D twin: (float)(ix*(nx-ix-1)*iy*(ny-iy-1)*iz*(nz-iz-1)) in the same wrapping int arithmetic. */
int heat_inidat3d(int nx, int ny, int nz, float *u);
/* prtdat: "%6.1f" text, iy descending outer, ix ascending inner (M:326-341). */
int heat_prtdat(int nx, int ny, const float *u, const char *fnam);
/* Binary twin of prtdat: 16-byte header {"HEATB200", int32 nx, int32 ny} + nx*ny floats. */
int heat_prtdat_bin(int nx, int ny, const float *u, const char *fnam);
int heat_readdat_bin(const char *fnam, int *nx, int *ny, float *u, size_t capacity_floats);

/* ---- one-shot: the whole loop region with caller-owned HOST memory -------------------
 * Replaces C:176-242 (cudaMalloc, H2D to both buffers, step loop, D2H) and, on one strip,
 * M:129-269.  grid: in = initial field, out = newest field.  updates_done = Jacobi updates
 * executed; the reference programs print updates_done-1 when they converge (M:302, C:256).
 * Without a convergence check, and when steps is small against nx, the three phases overlap: the field
 * streams through the GPU in bands of rows (H2D of band i+1, all updates of band i -- time-skewed -- and
 * D2H of the rows that are final run at the same time; same bits as the plain loop).  Pass page-locked
 * memory (cudaHostAlloc / cudaHostRegister) to get that overlap; pageable memory works but copies
 * synchronously.  HEAT_SOLVE_STREAMED=0 switches the overlap off, HEAT_SOLVE_BAND_ROWS sets the band.
 * The two device buffers of the call are kept for the next heat_solve of the same size on the same device (one
 * pair at most; a call of another size frees it first) instead of paying cudaMalloc/cudaFree every time:
 * heat_trim() releases them, HEAT_SOLVE_KEEP_BUFFERS=0 turns that off. */
int heat_solve(const heat_params *p, float *grid, int *updates_done, int *converged);
/* The same for a double-precision grid (p->grid_type = HEAT_GRID_F64). */
int heat_solve_f64(const heat_params *p, double *grid, int *updates_done, int *converged);
/* Frees what heat_solve keeps between calls (nothing else holds device memory after its handle is destroyed). */
int heat_trim(void);

/* The same over the first `ngpus` devices of this process: one row strip per GPU, ghost rows moved by
 * peer stores over NVLink, one host thread.  Replaces M:51-297 (topology, scatter, halo requests, loop,
 * gather) inside one box.  p->rank/nranks/device are ignored. */
int heat_solve_gpus(const heat_params *p, int ngpus, float *grid, int *updates_done, int *converged);

/* ---- device-resident handle (what C keeps alive in dvc_grids[2], C:177-198) ---------- */
/* The field of a new handle is unspecified until heat_upload* / heat_init_inidat (a handle that is
 * run without either starts from a zero field).  Everything the time loop needs -- CUDA graphs of 16
 * passes, occupancy queries, scratch of the resident-field kernel -- is built when the field arrives,
 * not by the first heat_enqueue / heat_run. */
int heat_create(const heat_params *p, heat_handle **out);
int heat_destroy(heat_handle *h);
/* Use the caller's CUDA stream (a cudaStream_t passed as void*) instead of the handle's own. */
int heat_set_stream(heat_handle *h, void *cuda_stream);
/* Host -> device of the FULL nx*ny field into BOTH buffers (C:197-198); a strip takes its rows
 * (ghost rows included), a 2-D block its rows x columns window (ghost columns included).
 * Strips connected across processes: every strip of the job must upload the same number of times
 * (an upload is one epoch of the step flags); before the copy the call waits on the device until the
 * neighbours have finished the pass this strip finished last, so that none of their ghost-row stores
 * can land on top of the new field. */
int heat_upload(heat_handle *h, const float *full_grid);
int heat_upload_f64(heat_handle *h, const double *full_grid);   /* handles created with HEAT_GRID_F64 */
/* Same from a device pointer holding the full field (used when the field is already in HBM). */
int heat_upload_device(heat_handle *h, const void *d_full_grid);
/* inidat evaluated on the device straight into both buffers (bit-identical to heat_inidat). */
int heat_init_inidat(heat_handle *h);
/* Runs the time loop: `steps` interpreted per params.step_convention; stops early when the
 * convergence check (params.conv_every) succeeds.  Asynchronous work is complete on return.
 * Replaces C:204-238 / M:159-265.  May be called repeatedly; the update counter that drives
 * conv_phase continues across calls. */
int heat_run(heat_handle *h, int steps, int *updates_done, int *converged);
/* Enqueue `updates` plain updates (no convergence check) on the stream and return without
 * waiting: the timed region of a benchmark. */
int heat_enqueue(heat_handle *h, int updates);
/* Wait for everything enqueued on the handle's stream; HEAT_E_TIMEOUT if a neighbour strip
 * never arrived. */
int heat_sync(heat_handle *h);
/* Device time of the last heat_run's loop (CUDA events around the loop only). */
int heat_elapsed_ms(heat_handle *h, float *ms);
/* max |old-new| over this strip for the most recent update (runs one reduction pass). */
int heat_residual(heat_handle *h, float *residual);
/* Newest field -> host; a strip writes only the rows (a 2-D block: rows x columns) it owns into
 * full_grid, so every rank of a job may download into the same array. */
int heat_download(heat_handle *h, float *full_grid);
int heat_download_f64(heat_handle *h, double *full_grid);
/* The same for a rank that holds only its strip: local_rows = the strip's window, dense
 * local_rows x ny (global rows first_global_row ..., ghost rows included: see heat_device_view);
 * own_rows receives the rows the strip owns (heat_strip_rows), dense.  Row strips only:
 * HEAT_E_STATE for a handle with nranks_y > 1 (column blocks go through heat_upload/heat_download). */
int heat_upload_local(heat_handle *h, const float *local_rows);
int heat_download_local(heat_handle *h, float *own_rows);
/* Device pointer / pitch (in floats) / local row count of the newest (which=0) or older
 * (which=1) buffer, and the global row of local row 0.  For a 2-D block the pointer is the block's
 * first local column (its west ghost column when it has a west neighbour). */
int heat_device_view(heat_handle *h, int which, void **d_ptr, int64_t *pitch, int *local_rows, int *first_global_row);
/* Kernels launched by this handle so far (benchmarks report it). */
int heat_launch_count(heat_handle *h, int64_t *launches);

/* ---- stateless single update on caller-owned DEVICE memory ---------------------------
 * One Jacobi update of the interior of a dense nx*ny field (pitch = ny): the body of the
 * reference kernels C:140-163 (plain) and C:42-138 (with the convergence reduction fused;
 * d_residual_bits receives max|old-new| as float bits via atomicMax and must be zeroed by
 * the caller).  d_out's ring is left untouched.  stream: cudaStream_t as void*, may be NULL. */
int heat_step_device(const void *d_in, void *d_out, int nx, int ny, float cx, float cy, int arith, void *stream);
int heat_step_residual_device(const void *d_in, void *d_out, int nx, int ny, float cx, float cy, int arith,
                              void *d_residual_bits, void *stream);

/* Same update + residual as heat_step_residual_device, for callers that cannot allocate: an
 * application that gave all of the GPU's memory to its own pool, or a time loop that cannot
 * afford the device-wide synchronisation of cudaMalloc/cudaFree.  It has no counterpart in the
 * reference.  The reduction's working memory is a fixed array inside the library image
 * (HEAT_NOALLOC_SCRATCH_BYTES, held by every GPU that loads the library until the process
 * exits), so the call performs no allocation.  The caller passes the input field, the output
 * field (dense nx*ny, ring left untouched) and d_residual, one device float that receives
 * max|old-new|; nothing needs zeroing beforehand.  Limits and rules: ny <= HEAT_NOALLOC_MAX_NY,
 * larger rows are refused with HEAT_E_INVALID; all calls on one GPU share the array, so calls
 * that overlap on that GPU race -- issue them one after another; the first call on a GPU
 * looks up the array's address and should be made outside any stream capture.
 * The array that is used is the one of the device that is CURRENT in the calling thread: make the
 * device that owns d_in, d_out, d_residual and the stream current (cudaSetDevice) before the
 * call, as you would for a kernel launch of your own.  Two limits of the "no allocation" promise:
 * with CUDA's lazy module loading the FIRST call on a GPU makes the driver load the library's
 * code and this array onto that GPU, which takes device memory once -- every later call allocates
 * nothing; and the address found by that first call is kept for the life of the process, so it
 * does not survive cudaDeviceReset() -- do not reset a device between calls. */
#define HEAT_NOALLOC_MAX_BLOCKS    16384
#define HEAT_NOALLOC_SCRATCH_BYTES ((HEAT_NOALLOC_MAX_BLOCKS + 64) * 4)
#define HEAT_NOALLOC_MAX_NY        (HEAT_NOALLOC_MAX_BLOCKS * 256 + 2)
int heat_step_residual_noalloc(const void *d_in, void *d_out, int nx, int ny, float cx, float cy, int arith,
                               void *d_residual, void *stream);

/* ---- row strips across GPUs (replaces the 2-D block topology + persistent halo requests,
 *      M:51-84,130-155,160-161,177,263) --------------------------------------------------
 * Strip r of n owns global rows [row0, row0+rows).  Each strip keeps T = temporal_block ghost rows
 * per neighbour; in every pass (T updates) the update kernel stores the strip's first/last T owned
 * rows straight into the neighbour's ghost rows over NVLink and raises the neighbour's step flag;
 * only the CTAs that touch a neighbour's rows ever wait for a flag. */
int heat_strip_rows(int nx, int nranks, int rank, int *row0, int *rows);
/* Who moves the ghost rows: HEAT_HALO_P2P (default) = the update kernel stores them into the
 * connected neighbours; HEAT_HALO_EXTERNAL = the caller copies row 1 / row n-2 of the newest
 * buffer into the neighbours' ghost rows between updates (e.g. NCCL send/recv on the
 * handle's stream), and no connection is needed. */
#define HEAT_HALO_P2P      0
#define HEAT_HALO_EXTERNAL 1
int heat_set_halo_mode(heat_handle *h, int mode);
#define HEAT_IPC_BLOB_BYTES 256
/* Export this strip's buffers + flags for another PROCESS (CUDA IPC handles). */
int heat_ipc_export(heat_handle *h, void *blob, size_t blob_bytes);
/* Connect neighbours from blobs (NULL = no neighbour on that side == MPI_PROC_NULL, M:68). */
int heat_ipc_connect(heat_handle *h, const void *north_blob, const void *south_blob);
/* Connect neighbours living in THIS process (one host thread driving several GPUs, or several
 * strips on one GPU).  Such strips are ordered with CUDA events and must be driven together,
 * update by update, through heat_group_run / heat_group_enqueue; upload every strip first. */
int heat_connect_local(heat_handle *h, heat_handle *north, heat_handle *south);
int heat_group_run(heat_handle **strips, int n, int steps, int *updates_done, int *converged);
/* 2-D blocks -- the reference's own cartesian topology (MPI_Cart_create, M:51-69; col_type halos,
 * M:82-84,139-155): block (rank, rank_y) of nranks x nranks_y also has a west and an east neighbour and
 * keeps one ghost column towards each, as a dense array of one float per local row beside the field; its
 * first/last owned column is stored into the neighbour's array by the update kernel.  temporal_block must be 1.  Inside one NVSwitch box row
 * strips are the better choice (contiguous halos); blocks are for what comes after it. */
int heat_connect_local2d(heat_handle *h, heat_handle *north, heat_handle *south, heat_handle *west, heat_handle *east);
int heat_ipc_connect2d(heat_handle *h, const void *north_blob, const void *south_blob, const void *west_blob,
                       const void *east_blob);
int heat_group_enqueue(heat_handle **strips, int n, int updates);
/* Convergence across strips in OTHER processes: every conv_every updates heat_run hands the
 * local max|old-new| to this callback, which must return the maximum over all strips in
 * *value (one 4-byte max-allreduce; replaces MPI_Allreduce(LAND), M:255).  Return 0 on success. */
/* Diagnostics (handle created with HEAT_PROFILE_WAIT=1 in the environment, else HEAT_E_STATE): nanoseconds the
 * edge CTAs of this strip's update kernels spent waiting for each neighbour's step flag (sides: north, south,
 * west, east) and the number of such waits, since the previous call.  The evidence behind "interior tiles never
 * wait" (replaces reading MPI_Waitall time off a profiler, M:177). */
int heat_debug_wait_stats(heat_handle *h, unsigned long long *ns_per_side, unsigned long long *waits_per_side);
typedef int (*heat_allreduce_max_fn)(float *value, void *ctx);
int heat_set_allreduce_max(heat_handle *h, heat_allreduce_max_fn fn, void *ctx);

#ifdef __cplusplus
}
#endif
#endif /* HEAT_B200_H */
