// heat_api.cu -- the C ABI declared in include/heat.h: handle, buffers, stream, time loop.
// Host code only drives the device; there is NO CPU computation path: without a usable
// CUDA device every compute entry point fails with HEAT_E_CUDA.
#include "../../include/heat.h"
#include "heat_kernels.cuh"

#include <cuda.h>
#include <math.h>
#include <mutex>
#include <new>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

using namespace heatk;

// ------------------------------------------------------------------ errors
static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

#define CK(call)                                                                                        \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess)                                                                          \
            return fail(HEAT_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

extern "C" const char* heat_last_error(void) { return g_err; }

extern "C" int heat_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

extern "C" int heat_params_default(heat_params* p)
{
    if (!p) return fail(HEAT_E_INVALID, "heat_params_default: null");
    memset(p, 0, sizeof *p);
    p->nx = p->ny = 20; // NXPROB/NYPROB defaults (M:7-12, C:7-12)
    p->steps = 100;     // M:16-18
    p->cx = p->cy = 0.1f;
    p->conv_every = 0;
    p->conv_eps = 1e-3f;
    p->conv_phase = HEAT_PHASE_CUDA;
    p->step_convention = HEAT_STEPS_CUDA;
    p->arith = HEAT_F32_FMA;
    p->temporal_block = 1;
    p->device = 0;
    p->rank = 0;
    p->nranks = 1;
    p->rank_y = 0;
    p->nranks_y = 1;
    p->grid_type = HEAT_GRID_F32;
    p->nz = 1;
    p->cz = 0.1f;
    return HEAT_OK;
}

// ------------------------------------------------------------------ host helpers (M:315-341)
extern "C" int heat_inidat(int nx, int ny, float* u)
{
    if (nx < 1 || ny < 1 || !u) return fail(HEAT_E_INVALID, "heat_inidat: bad arguments");
    for (int ix = 0; ix < nx; ix++) {
        const uint32_t a = (uint32_t)ix * (uint32_t)(nx - ix - 1);
        float* row = u + (size_t)ix * ny;
        for (int iy = 0; iy < ny; iy++) row[iy] = (float)(int32_t)(a * (uint32_t)iy * (uint32_t)(ny - iy - 1));
    }
    return HEAT_OK;
}

extern "C" int heat_inidat_f64(int nx, int ny, double* u)
{
    if (nx < 1 || ny < 1 || !u) return fail(HEAT_E_INVALID, "heat_inidat_f64: bad arguments");
    for (int ix = 0; ix < nx; ix++) {
        const uint32_t a = (uint32_t)ix * (uint32_t)(nx - ix - 1);
        double* row = u + (size_t)ix * ny;
        for (int iy = 0; iy < ny; iy++) row[iy] = (double)(int32_t)(a * (uint32_t)iy * (uint32_t)(ny - iy - 1));
    }
    return HEAT_OK;
}

extern "C" int heat_inidat3d(int nx, int ny, int nz, float* u)
{
    if (nx < 1 || ny < 1 || nz < 1 || !u) return fail(HEAT_E_INVALID, "heat_inidat3d: bad arguments");
    for (int ix = 0; ix < nx; ix++)
        for (int iy = 0; iy < ny; iy++) {
            const uint32_t a = (uint32_t)ix * (uint32_t)(nx - ix - 1) * (uint32_t)iy * (uint32_t)(ny - iy - 1);
            float* row = u + ((size_t)ix * ny + iy) * nz;
            for (int iz = 0; iz < nz; iz++) row[iz] = (float)(int32_t)(a * (uint32_t)iz * (uint32_t)(nz - iz - 1));
        }
    return HEAT_OK;
}

// "%6.1f" of a float without printf (the text dump is 3 GB at 16384^2 and printf spends 0.7 us per value):
// 10*v is exact in double (24-bit significand times 10), so rounding it to an integer in the current rounding
// mode -- ties to even by default -- is exactly the decimal rounding glibc performs on the exact value; the
// sign is printed from the sign bit like printf does ("-0.0").  Huge values, inf and nan go to snprintf.
static inline char* put_6_1f(char* p, float v)
{
    const double x = (double)v * 10.0;
    if (!(fabs(x) < 1.0e15)) return p + snprintf(p, 48, "%6.1f", (double)v);
    uint64_t n = (uint64_t)fabs(nearbyint(x));
    char tmp[24];
    int len = 0;
    tmp[len++] = (char)('0' + n % 10);
    n /= 10;
    tmp[len++] = '.';
    do {
        tmp[len++] = (char)('0' + n % 10);
        n /= 10;
    } while (n);
    if (signbit(v)) tmp[len++] = '-';
    for (int pad = len; pad < 6; pad++) *p++ = ' ';
    while (len) *p++ = tmp[--len];
    return p;
}

extern "C" int heat_prtdat(int nx, int ny, const float* u, const char* fnam)
{
    if (nx < 1 || ny < 1 || !u || !fnam) return fail(HEAT_E_INVALID, "heat_prtdat: bad arguments");
    FILE* fp = fopen(fnam, "w");
    if (!fp) return fail(HEAT_E_INVALID, "heat_prtdat: cannot open %s", fnam);
    // One text row per iy (descending), nx values each (M:326-341): a text row is a COLUMN of u.  A band of
    // BAND text rows is formatted together so that u is read a cache line at a time, not one float per line.
    enum { BAND = 16, MAX_CHARS = 48 }; // "%6.1f" of -FLT_MAX is 42 characters, plus the separator
    char* line[BAND] = {nullptr};
    bool ok = true;
    for (int k = 0; k < BAND && k < ny && ok; k++) ok = (line[k] = (char*)malloc((size_t)nx * MAX_CHARS)) != nullptr;
    for (int hi = ny - 1; hi >= 0 && ok; hi -= BAND) {
        const int lo = hi - BAND + 1 > 0 ? hi - BAND + 1 : 0, nb = hi - lo + 1;
        char* w[BAND];
        for (int k = 0; k < nb; k++) w[k] = line[k];
        for (int ix = 0; ix < nx; ix++) {
            const float* src = u + (size_t)ix * ny + lo;
            const char sep = ix == nx - 1 ? '\n' : ' ';
            for (int k = 0; k < nb; k++) {
                w[k] = put_6_1f(w[k], src[k]);
                *w[k]++ = sep;
            }
        }
        for (int k = nb - 1; k >= 0 && ok; k--) ok = fwrite(line[k], 1, (size_t)(w[k] - line[k]), fp) == (size_t)(w[k] - line[k]);
    }
    for (int k = 0; k < BAND; k++) free(line[k]);
    ok = (fclose(fp) == 0) && ok;
    return ok ? HEAT_OK : fail(HEAT_E_INVALID, "heat_prtdat: write to %s failed (or out of memory)", fnam);
}

struct bin_header {
    char magic[8];
    int32_t nx, ny;
};

extern "C" int heat_prtdat_bin(int nx, int ny, const float* u, const char* fnam)
{
    if (nx < 1 || ny < 1 || !u || !fnam) return fail(HEAT_E_INVALID, "heat_prtdat_bin: bad arguments");
    FILE* fp = fopen(fnam, "wb");
    if (!fp) return fail(HEAT_E_INVALID, "heat_prtdat_bin: cannot open %s", fnam);
    bin_header h;
    memcpy(h.magic, "HEATB200", 8);
    h.nx = nx;
    h.ny = ny;
    const size_t n = (size_t)nx * ny;
    int ok = fwrite(&h, sizeof h, 1, fp) == 1 && fwrite(u, sizeof(float), n, fp) == n;
    ok = (fclose(fp) == 0) && ok;
    return ok ? HEAT_OK : fail(HEAT_E_INVALID, "heat_prtdat_bin: write to %s failed", fnam);
}

extern "C" int heat_readdat_bin(const char* fnam, int* nx, int* ny, float* u, size_t capacity_floats)
{
    if (!fnam || !nx || !ny) return fail(HEAT_E_INVALID, "heat_readdat_bin: bad arguments");
    FILE* fp = fopen(fnam, "rb");
    if (!fp) return fail(HEAT_E_INVALID, "heat_readdat_bin: cannot open %s", fnam);
    bin_header h;
    if (fread(&h, sizeof h, 1, fp) != 1 || memcmp(h.magic, "HEATB200", 8) != 0 || h.nx < 1 || h.ny < 1) {
        fclose(fp);
        return fail(HEAT_E_INVALID, "heat_readdat_bin: %s is not a HEATB200 dump", fnam);
    }
    *nx = h.nx;
    *ny = h.ny;
    int rc = HEAT_OK;
    if (u) {
        const size_t n = (size_t)h.nx * h.ny;
        if (n > capacity_floats)
            rc = fail(HEAT_E_INVALID, "heat_readdat_bin: buffer too small (%zu < %zu)", capacity_floats, n);
        else if (fread(u, sizeof(float), n, fp) != n)
            rc = fail(HEAT_E_INVALID, "heat_readdat_bin: %s is truncated", fnam);
    }
    fclose(fp);
    return rc;
}

// ------------------------------------------------------------------ strips
extern "C" int heat_strip_rows(int nx, int nranks, int rank, int* row0, int* rows)
{
    if (nx < 1 || nranks < 1 || rank < 0 || rank >= nranks || nranks > nx)
        return fail(HEAT_E_INVALID, "heat_strip_rows: nx=%d nranks=%d rank=%d", nx, nranks, rank);
    // balanced split; the reference assumes divisibility (M:72-73), this does not
    const int base = nx / nranks, rem = nx % nranks;
    const int r0 = rank * base + (rank < rem ? rank : rem);
    if (row0) *row0 = r0;
    if (rows) *rows = base + (rank < rem ? 1 : 0);
    return HEAT_OK;
}

// ------------------------------------------------------------------ handle
struct peer_side {
    float* buf[2] = {nullptr, nullptr};   // neighbour's two buffers (peer-mapped)
    unsigned long long* flag = nullptr;   // neighbour's flag slot that this strip writes
    int nloc = 0, nyloc = 0;
    int64_t pitch = 0;
    float* gcol = nullptr;                 // neighbour's dense ghost-column arrays (2-D blocks), see heat_handle::gcol
    void* ipc_base[3] = {nullptr, nullptr, nullptr}; // bases opened with cudaIpcOpenMemHandle
};

struct heat_handle {
    heat_params p;
    int row0 = 0, rows_own = 0; // owned global rows [row0, row0+rows_own)
    int lo = 0, nloc = 0;       // local array = global rows [lo, lo+nloc)
    int ghost = 1;              // ghost rows kept towards each neighbour (= temporal_block)
    int upd_lo = 1, upd_hi = 0; // local rows [upd_lo, upd_hi) are the ones this strip updates
    int col0 = 0, cols_own = 0; // owned global columns [col0, col0+cols_own) (2-D blocks; else all)
    int lo_c = 0, nyloc = 0;    // local array = global columns [lo_c, lo_c+nyloc): + one ghost column per neighbour
    int64_t pitch = 0;          // row pitch in FLOATS (4-byte units) whatever the element type: buf[k] + r*pitch is row r
    int nz = 1, ny3 = 0;        // 3-D grids: nz > 1 and ny3 = the caller's ny; a handle row is then one ix-plane of ny3*nz floats (p.ny)
    int es = 1;                 // floats per element: 1 = float grid, 2 = double grid (HEAT_GRID_F64)
    float* buf[2] = {nullptr, nullptr};
    int cur = 0;                // index of the newest buffer
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    unsigned* d_resid = nullptr; // max|old-new| as bits: 4 bytes used for float grids, 8 for double grids
    unsigned* h_resid = nullptr; // pinned, 8 bytes
    int* d_status = nullptr;     // mapped pinned: wait-kernel time-outs
    int* h_status = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_resid = nullptr;
    float last_ms = 0.0f;
    long long it = 0;            // updates since the last upload (drives conv_phase)
    long long sweeps = 0;        // passes over HBM so far (alternates the sweep direction)
    unsigned long long seq = 0;  // monotonic epoch for the strip flags
    int64_t launches = 0;
    int rows_per_tile = 0;
    int rows_per_tile_tb = 0;     // 0 = choose per launch
    int tb_per_sm = 0;            // resident CTAs per SM of this handle's K4 instantiation (queried once)
    int sm_count = 148;
    bool loaded = false;          // a field has been uploaded / initialised (else the first use zero-fills)
    bool keep_buffers = false;    // heat_solve's own handle: its two field buffers go to the one-entry cache when it is destroyed
    // neighbours: side 0 = north (lower rows), 1 = south, 2 = west (lower columns), 3 = east
    unsigned* d_edge_count = nullptr;    // [4] edge-CTA arrival counters of the fused strip ordering, per side
    unsigned long long* d_wait_prof = nullptr; // HEAT_PROFILE_WAIT=1: [8] ns waited / waits per side (heat_debug_wait_stats)
    strip_sync sync_args = {};           // filled per pass for flag-ordered strips, else all null
    col_peers col_args = {};             // 2-D blocks: where my first / last owned column goes
    unsigned long long* flags = nullptr; // [0] written by the north neighbour, [1] south, [2] west, [3] east
    // 2-D blocks: dense ghost columns [buffer][side: 0 west, 1 east][local row], GCOL_OFFSET bytes behind the
    // flags in the same allocation (so one IPC handle exports both)
    float* gcol = nullptr;
    peer_side peer[4];
    bool connected = false;
    // neighbours in this process are ordered with events instead of flags (no spinning kernels)
    heat_handle* local_nb[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_done[2] = {nullptr, nullptr}; // recorded after epoch e in ev_done[e & 1]
    bool serpentine = true;
    int k1_variant = 2; // HEAT_K1_VARIANT: 0 = vec (register streaming), 1 = bulk (shared-memory staged), 2 = auto (bulk for long rows)
    int cols_first = 0; // HEAT_COLS_FIRST = M: column-edge CTAs ahead of the rest, M tile rows each (2-D blocks across processes)
    // CUDA graphs of GRAPH_PASSES consecutive passes, one per (newest buffer, sweep direction, fused?)
    cudaGraphExec_t graphs[2][2][2] = {};
    cudaStream_t capture_stream = nullptr;
    bool use_graphs = true;
    // K5: the whole field resident in shared memory across many updates (small grids)
    bool use_resident = true;
    float* res_halo = nullptr;
    unsigned long long* res_flags = nullptr;
    unsigned long long res_epoch = 0;
    int res_ctas = 0, res_threads = 0, res_tpr = 0, res_ghost = 1;
    size_t res_smem = 0;
    bool external_halo = false; // the caller moves the ghost rows between updates
    heat_allreduce_max_fn allreduce = nullptr; // joins the residuals of strips in other processes
    void* allreduce_ctx = nullptr;
};

struct ipc_blob {
    char magic[8];
    cudaIpcMemHandle_t mem[3]; // buf0, buf1, flags
    int64_t offset[3];         // of the pointer inside the exported allocation
    int32_t nloc, rank, nranks, nyloc, rank_y, nranks_y;
    int64_t pitch;
};
static_assert(sizeof(ipc_blob) <= HEAT_IPC_BLOB_BYTES, "ipc blob too large");

static const size_t GCOL_OFFSET = 256; // bytes between the four flags and the dense ghost-column arrays

static int env_int(const char* k, int dflt)
{
    const char* v = getenv(k);
    return (v && *v) ? atoi(v) : dflt;
}

static int validate(const heat_params* p)
{
    if (!p) return fail(HEAT_E_INVALID, "params: null");
    if (p->nx < 1 || p->ny < 1) return fail(HEAT_E_INVALID, "params: nx=%d ny=%d", p->nx, p->ny);
    if (p->steps < 0) return fail(HEAT_E_INVALID, "params: steps=%d", p->steps);
    if (p->conv_every < 0) return fail(HEAT_E_INVALID, "params: conv_every=%d", p->conv_every);
    if (p->arith < HEAT_F32_FMA || p->arith > HEAT_MPI_F64MIX) return fail(HEAT_E_INVALID, "params: arith=%d", p->arith);
    if (p->temporal_block < 1 || p->temporal_block > 4)
        return fail(HEAT_E_INVALID, "params: temporal_block=%d (1 to 4)", p->temporal_block);
    if (p->nranks < 1 || p->rank < 0 || p->rank >= p->nranks)
        return fail(HEAT_E_INVALID, "params: rank=%d nranks=%d", p->rank, p->nranks);
    if (p->nranks > 1 && p->nx / p->nranks < (p->temporal_block > 2 ? p->temporal_block : 2))
        return fail(HEAT_E_INVALID, "params: %d rows cannot be cut into %d strips of >= %d rows", p->nx, p->nranks,
                    p->temporal_block > 2 ? p->temporal_block : 2);
    if (p->nranks_y < 1 || p->rank_y < 0 || p->rank_y >= p->nranks_y)
        return fail(HEAT_E_INVALID, "params: rank_y=%d nranks_y=%d", p->rank_y, p->nranks_y);
    if (p->nranks_y > 1 && (p->ny / p->nranks_y < 2 || p->temporal_block != 1))
        return fail(HEAT_E_INVALID, "params: column blocks need >= 2 columns each and temporal_block 1");
    if (p->grid_type != HEAT_GRID_F32 && p->grid_type != HEAT_GRID_F64) return fail(HEAT_E_INVALID, "params: grid_type=%d", p->grid_type);
    if (p->grid_type == HEAT_GRID_F64 && (p->temporal_block != 1 || p->nranks_y != 1))
        return fail(HEAT_E_INVALID, "params: double-precision grids run single updates on row strips (temporal_block 1, nranks_y 1)");
    if (p->nz < 0) return fail(HEAT_E_INVALID, "params: nz=%d", p->nz);
    if (p->nz > 1) {
        if (p->nz % 4 || p->nz < 4) return fail(HEAT_E_INVALID, "params: nz=%d (3-D grids need nz a multiple of 4)", p->nz);
        if ((long long)p->ny * p->nz > 0x7fffffffLL) return fail(HEAT_E_INVALID, "params: ny*nz=%lld does not fit an int", (long long)p->ny * p->nz);
        if (p->nranks_y != 1 || p->temporal_block != 1 || p->grid_type != HEAT_GRID_F32)
            return fail(HEAT_E_INVALID, "params: 3-D grids run single float updates on strips of planes (nranks_y 1, temporal_block 1)");
    }
    if (p->conv_phase != HEAT_PHASE_MPI && p->conv_phase != HEAT_PHASE_CUDA)
        return fail(HEAT_E_INVALID, "params: conv_phase=%d", p->conv_phase);
    if (p->step_convention != HEAT_STEPS_CUDA && p->step_convention != HEAT_STEPS_MPI)
        return fail(HEAT_E_INVALID, "params: step_convention=%d", p->step_convention);
    return HEAT_OK;
}

// heat_solve() allocates and frees both field buffers on every call: 9 + 16 ms at 65536^2, 5 of 12 ms at 8192^2, for a call
// whose useful part is bound by PCIe.  Its internal handle therefore hands the pair to a one-entry cache when it is
// destroyed, and the next heat_solve of the same size on the same device takes it from there.  At most one pair is held
// (a different size or device frees it first); heat_trim() releases it; HEAT_SOLVE_KEEP_BUFFERS=0 turns the cache off.
// Handles made by heat_create() never use it.
struct solve_buffer_cache {
    std::mutex m;
    int device = -1;
    size_t bytes = 0;
    float* buf[2] = {nullptr, nullptr};
};
static solve_buffer_cache g_solve_cache;

static void cache_drop_locked()
{
    solve_buffer_cache& c = g_solve_cache;
    if (!c.buf[0]) return;
    int cur = -1;
    cudaGetDevice(&cur);
    if (cur != c.device) cudaSetDevice(c.device);
    cudaFree(c.buf[0]);
    cudaFree(c.buf[1]);
    if (cur >= 0 && cur != c.device) cudaSetDevice(cur);
    c.buf[0] = c.buf[1] = nullptr;
    c.bytes = 0;
    c.device = -1;
    cudaGetLastError();
}

static bool cache_take(int device, size_t bytes, float** b0, float** b1)
{
    solve_buffer_cache& c = g_solve_cache;
    std::lock_guard<std::mutex> guard(c.m);
    if (!c.buf[0]) return false;
    if (c.device != device || c.bytes != bytes) {
        cache_drop_locked(); // make room: the caller is about to allocate another size
        return false;
    }
    *b0 = c.buf[0], *b1 = c.buf[1];
    c.buf[0] = c.buf[1] = nullptr;
    c.bytes = 0;
    c.device = -1;
    return true;
}

static void cache_put(int device, size_t bytes, float* b0, float* b1)
{
    solve_buffer_cache& c = g_solve_cache;
    std::lock_guard<std::mutex> guard(c.m);
    cache_drop_locked();
    c.device = device, c.bytes = bytes, c.buf[0] = b0, c.buf[1] = b1;
}

extern "C" int heat_trim(void)
{
    std::lock_guard<std::mutex> guard(g_solve_cache.m);
    cache_drop_locked();
    return HEAT_OK;
}

extern "C" int heat_destroy(heat_handle* h)
{
    if (!h) return HEAT_OK;
    cudaSetDevice(h->p.device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    for (int s = 0; s < 4; s++)
        for (int k = 0; k < 3; k++)
            if (h->peer[s].ipc_base[k]) cudaIpcCloseMemHandle(h->peer[s].ipc_base[k]);
    if (h->keep_buffers && h->buf[0] && h->buf[1]) {
        cache_put(h->p.device, (size_t)h->nloc * h->pitch * sizeof(float), h->buf[0], h->buf[1]);
    } else {
        if (h->buf[0]) cudaFree(h->buf[0]);
        if (h->buf[1]) cudaFree(h->buf[1]);
    }
    if (h->flags) cudaFree(h->flags);
    if (h->d_resid) cudaFree(h->d_resid);
    if (h->d_edge_count) cudaFree(h->d_edge_count);
    if (h->d_wait_prof) cudaFree(h->d_wait_prof);
    if (h->res_halo) cudaFree(h->res_halo);
    if (h->res_flags) cudaFree(h->res_flags);
    if (h->h_status) cudaFreeHost(h->h_status); // one pinned block: status word, then the residual words (h_resid)
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->ev_resid) cudaEventDestroy(h->ev_resid);
    for (int k = 0; k < 2; k++)
        if (h->ev_done[k]) cudaEventDestroy(h->ev_done[k]);
    for (auto& a : h->graphs)
        for (auto& b : a)
            for (auto& g : b)
                if (g) cudaGraphExecDestroy(g);
    if (h->capture_stream) cudaStreamDestroy(h->capture_stream);
    if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
    cudaGetLastError();
    delete h;
    return HEAT_OK;
}

static int create_handle(const heat_params* p_in, heat_handle** out, bool solve_cache);
extern "C" int heat_create(const heat_params* p_in, heat_handle** out) { return create_handle(p_in, out, false); }

static int create_handle(const heat_params* p_in, heat_handle** out, bool solve_cache)
{
    if (!out) return fail(HEAT_E_INVALID, "heat_create: null out");
    *out = nullptr;
    int rc = validate(p_in);
    if (rc) return rc;
    // 3-D grids: everything below (buffers, copies, streaming, residual pass) sees nx rows of ny*nz floats
    heat_params folded = *p_in;
    if (folded.nz > 1) folded.ny = p_in->ny * p_in->nz;
    const heat_params* p = &folded;
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (p->device < 0 || p->device >= ndev) return fail(HEAT_E_INVALID, "heat_create: device %d of %d", p->device, ndev);
    CK(cudaSetDevice(p->device));

    heat_handle* h = new (std::nothrow) heat_handle();
    if (!h) return fail(HEAT_E_NOMEM, "heat_create: out of host memory");
    h->p = *p;
    if (p_in->nz > 1) h->nz = p_in->nz, h->ny3 = p_in->ny;
    heat_strip_rows(p->nx, p->nranks, p->rank, &h->row0, &h->rows_own);
    // local array: owned rows plus one ghost row towards each existing neighbour
    // (T ghost rows when T updates are fused per pass: they are refreshed once per pass)
    h->ghost = p->temporal_block;
    h->lo = (p->rank > 0) ? h->row0 - h->ghost : 0;
    const int hi = (p->rank < p->nranks - 1) ? h->row0 + h->rows_own + h->ghost : p->nx;
    h->nloc = hi - h->lo;
    h->upd_lo = (p->rank > 0) ? h->ghost : 1;
    h->upd_hi = (p->rank < p->nranks - 1) ? h->nloc - h->ghost : h->nloc - 1;
    // 2-D blocks (the reference's topology, M:51-75): the same split along the columns, one ghost column
    // per neighbour; a single column block keeps all columns and the ring columns are its own
    heat_strip_rows(p->ny, p->nranks_y, p->rank_y, &h->col0, &h->cols_own);
    h->lo_c = (p->rank_y > 0) ? h->col0 - 1 : 0;
    const int hi_c = (p->rank_y < p->nranks_y - 1) ? h->col0 + h->cols_own + 1 : p->ny;
    h->nyloc = hi_c - h->lo_c;
    // rows padded to 128 bytes so that every row starts on a float4 / double2 (and a cache line)
    h->es = p->grid_type == HEAT_GRID_F64 ? 2 : 1;
    h->pitch = ((int64_t)h->nyloc * h->es + 31) / 32 * 32;

    const size_t bytes = (size_t)h->nloc * h->pitch * sizeof(float);
#define CKH(call)                                                                                      \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) {                                                                       \
            int c_ = (e_ == cudaErrorMemoryAllocation) ? HEAT_E_NOMEM : HEAT_E_CUDA;                   \
            fail(c_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__);      \
            heat_destroy(h);                                                                           \
            return c_;                                                                                 \
        }                                                                                              \
    } while (0)
    h->keep_buffers = solve_cache && p->nranks == 1 && p->nranks_y == 1 && env_int("HEAT_SOLVE_KEEP_BUFFERS", 1) != 0;
    if (!(h->keep_buffers && cache_take(p->device, bytes, &h->buf[0], &h->buf[1]))) {
        CKH(cudaMalloc((void**)&h->buf[0], bytes));
        CKH(cudaMalloc((void**)&h->buf[1], bytes));
    }
    const size_t flag_bytes = GCOL_OFFSET + (p->nranks_y > 1 ? 4 * (size_t)h->nloc * sizeof(float) : 0);
    CKH(cudaMalloc((void**)&h->flags, flag_bytes));
    if (p->nranks_y > 1) h->gcol = (float*)((char*)h->flags + GCOL_OFFSET);
    CKH(cudaMalloc((void**)&h->d_resid, 2 * sizeof(unsigned)));
    CKH(cudaMalloc((void**)&h->d_edge_count, 4 * sizeof(unsigned)));
    // one pinned, mapped block for the host-visible words (a pinned allocation costs 0.5-1 ms): [0] time-out status, [2..3] residual
    CKH(cudaHostAlloc((void**)&h->h_status, 4 * sizeof(unsigned), cudaHostAllocMapped));
    memset(h->h_status, 0, 4 * sizeof(unsigned));
    h->h_resid = (unsigned*)h->h_status + 2;
    CKH(cudaHostGetDevicePointer((void**)&h->d_status, h->h_status, 0));
    CKH(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    h->own_stream = true;
    // heat_upload / heat_init_inidat overwrite every row and column of the field; only the padding columns
    // [nyloc, pitch), which the kernels load and copy through, must hold something harmless
    if (h->pitch > (int64_t)h->nyloc * h->es)
        for (int k = 0; k < 2; k++)
            CKH(cudaMemset2DAsync(h->buf[k] + (size_t)h->nyloc * h->es, h->pitch * sizeof(float), 0,
                                  (size_t)(h->pitch - (int64_t)h->nyloc * h->es) * sizeof(float), h->nloc, h->stream));
    CKH(cudaMemsetAsync(h->flags, 0, flag_bytes, h->stream));
    CKH(cudaMemsetAsync(h->d_resid, 0, 2 * sizeof(unsigned), h->stream));
    CKH(cudaMemsetAsync(h->d_edge_count, 0, 4 * sizeof(unsigned), h->stream));
    if (env_int("HEAT_PROFILE_WAIT", 0) != 0) {
        CKH(cudaMalloc((void**)&h->d_wait_prof, 8 * sizeof(unsigned long long)));
        CKH(cudaMemsetAsync(h->d_wait_prof, 0, 8 * sizeof(unsigned long long), h->stream));
    }
    CKH(cudaEventCreate(&h->ev0));
    CKH(cudaEventCreate(&h->ev1));
    CKH(cudaEventCreateWithFlags(&h->ev_resid, cudaEventDisableTiming));
    CKH(cudaEventCreateWithFlags(&h->ev_done[0], cudaEventDisableTiming));
    CKH(cudaEventCreateWithFlags(&h->ev_done[1], cudaEventDisableTiming));
    CKH(cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, p->device));
    CKH(cudaStreamSynchronize(h->stream));
#undef CKH
    h->rows_per_tile = env_int("HEAT_ROWS_PER_TILE", 0);
    h->rows_per_tile_tb = env_int("HEAT_ROWS_PER_TILE_TB", 0);
    h->serpentine = env_int("HEAT_SERPENTINE", 1) != 0;
    {
        const char* v = getenv("HEAT_K1_VARIANT");
        h->k1_variant = !v || !*v || !strcmp(v, "auto") ? 2 : !strcmp(v, "bulk") ? 1 : 0;
#ifdef HEAT_BOUNDS_CHECK
        h->k1_variant = 0; // the checked build watches the loads of the register-streaming kernels
#endif
    }
    h->cols_first = env_int("HEAT_COLS_FIRST", 0);
    if (h->cols_first < 0) h->cols_first = 0;
#ifdef HEAT_BOUNDS_CHECK
    h->use_graphs = false; // the debug build synchronises around every pass
    h->use_resident = false;
#else
    h->use_graphs = env_int("HEAT_GRAPHS", 1) != 0;
    h->use_resident = env_int("HEAT_RESIDENT", 1) != 0;
#endif
    if (h->es == 2 || h->nz > 1) h->use_resident = false; // K5 keeps a 2-D float field in shared memory
    *out = h;
    return HEAT_OK;
}

extern "C" int heat_set_stream(heat_handle* h, void* cuda_stream)
{
    if (!h) return fail(HEAT_E_INVALID, "heat_set_stream: null handle");
    CK(cudaSetDevice(h->p.device));
    CK(cudaStreamSynchronize(h->stream));
    if (h->own_stream) {
        cudaStreamDestroy(h->stream);
        h->own_stream = false;
    }
    h->stream = (cudaStream_t)cuda_stream;
    return HEAT_OK;
}

// ------------------------------------------------------------------ launches
static int pick_rows_per_tile(const heat_handle* h, int threads)
{
    if (h->rows_per_tile > 0) return h->rows_per_tile;
    // Short tiles win (measured on B200, 8192^2: 8 rows 818 GLUP/s, 16 rows 778, 32 rows 754):
    // many small CTAs balance the 148 SMs better than the two halo rows per tile cost in L2 reads.
    (void)threads;
    return 8;
}

// how many CTAs touch the north / south neighbour (same predicate as edge_n / edge_s in the kernels)
static void count_edge_ctas(heat_handle* h, int rows, int rpt, unsigned grid_x, unsigned tiles)
{
    // tile t covers rows [t*rpt, min((t+1)*rpt, rows)) relative to upd_lo
    const unsigned g = (unsigned)h->ghost;
    unsigned tn = (g + rpt - 1) / rpt;                               // tiles with r0 < ghost
    const unsigned first_s = rows > (int)g ? (unsigned)(rows - g) / rpt : 0; // first tile with r1 > rows - ghost
    if (tn > tiles) tn = tiles;
    const unsigned ts = tiles - (first_s < tiles ? first_s : tiles);
    h->sync_args.count[0] = tn * grid_x;
    h->sync_args.count[1] = ts * grid_x;
    h->sync_args.count[2] = h->sync_args.count[3] = tiles; // one west-edge and one east-edge CTA per tile row
}

#ifndef HEAT_BULK_ROWS
#define HEAT_BULK_ROWS 8 // tile height of the shared-memory staged kernel
#endif
static const int64_t BULK_MIN_PITCH = 16384; // floats per row from which "auto" stages tiles in shared memory

template <int ARITH, bool RESID>
static cudaError_t launch_step_t(heat_handle* h, const float* in, float* out, float* peer_n, float* peer_s, int reverse)
{
    const int rows = h->upd_hi - h->upd_lo, ny = h->nyloc;
    if (rows < 1 || ny < 3) return cudaSuccess; // no interior: nothing to update (ring only)
#ifndef HEAT_K1_THREADS
#define HEAT_K1_THREADS 128
#endif
#ifndef HEAT_F64MIX_UNROLL
#define HEAT_F64MIX_UNROLL 4
#endif
    constexpr int THREADS = HEAT_K1_THREADS, UNROLL = ARITH == HEAT_ARITH_F64MIX ? HEAT_F64MIX_UNROLL : 4;
    int rpt = pick_rows_per_tile(h, THREADS);
    if (rpt < h->ghost) rpt = h->ghost;
    if ((rows + rpt - 1) / rpt > 65535) rpt = (rows + 65534) / 65535; // grid.y limit: taller tiles
    dim3 grid((unsigned)((h->pitch / 4 + THREADS - 1) / THREADS), (unsigned)((rows + rpt - 1) / rpt));
    count_edge_ctas(h, rows, rpt, grid.x, grid.y);
    // K1b, the shared-memory staged variant: measured faster than the register-streaming kernel where rows are long
    // (16384 floats and more), slower below; "auto" picks by row length.  Row strips and the residual are supported,
    // column blocks are not.
    // (MPI_F64MIX: measured slower staged, 680 vs 732 GLUP/s at 16384^2 -- its warps are busy converting, and the
    // staged form has fewer loads in flight per warp -- so "auto" keeps that mode on K1)
    // Flag-ordered strips: measured on 2 GPUs at 65536^2, three alternating runs of 21 repetitions each: K1 1704 GLUP/s
    // (repetitions 50.2-51.7 ms), K1b 1635 (49.9-55.4 ms), so "auto" stages only on a single strip.
    const bool bulk = h->p.nranks_y == 1 &&
                      (h->k1_variant == 1 || (h->k1_variant == 2 && h->pitch >= BULK_MIN_PITCH && ARITH != HEAT_ARITH_F64MIX &&
                                              h->p.nranks == 1));
    if (bulk) {
        constexpr int ROWS = HEAT_BULK_ROWS;
        if (h->ghost <= ROWS && (rows + ROWS - 1) / ROWS <= 65535) {
            dim3 g((unsigned)((h->pitch / 4 + THREADS - 1) / THREADS), (unsigned)((rows + ROWS - 1) / ROWS));
            count_edge_ctas(h, rows, ROWS, g.x, g.y);
            if (h->sync_args.counters)
                heat_step_bulk<ARITH, RESID, ROWS, THREADS, true, true><<<g, THREADS, 0, h->stream>>>(
                    in, out, h->upd_lo, h->upd_hi, ny, h->pitch, h->p.cx, h->p.cy, h->d_resid, peer_n, peer_s, h->ghost, reverse,
                    h->sync_args, h->d_status);
            else if (peer_n || peer_s)
                heat_step_bulk<ARITH, RESID, ROWS, THREADS, false, true><<<g, THREADS, 0, h->stream>>>(
                    in, out, h->upd_lo, h->upd_hi, ny, h->pitch, h->p.cx, h->p.cy, h->d_resid, peer_n, peer_s, h->ghost, reverse,
                    h->sync_args, h->d_status);
            else // a single strip's instantiation carries no peer-store code
                heat_step_bulk<ARITH, RESID, ROWS, THREADS, false, false><<<g, THREADS, 0, h->stream>>>(
                    in, out, h->upd_lo, h->upd_hi, ny, h->pitch, h->p.cx, h->p.cy, h->d_resid, nullptr, nullptr, h->ghost, reverse,
                    h->sync_args, h->d_status);
            h->launches++;
            return cudaGetLastError();
        }
    }
    if (h->p.nranks_y > 1) { // 2-D blocks: column neighbours too
        const int M = h->col_args.edge_first, xe = ((ny - 2) >> 2) / THREADS;
        if (M > 1 && h->sync_args.counters && cols_first_active((int)grid.x, (int)grid.y, xe))
            // column-edge CTAs team up M tile rows at a time (the two row-edge tile rows stay as they are)
            h->sync_args.count[2] = h->sync_args.count[3] = 2 + (grid.y - 2 + M - 1) / M;
        if (h->sync_args.counters)
            heat_step_vec<ARITH, RESID, THREADS, UNROLL, true, true><<<grid, THREADS, 0, h->stream>>>(
                in, out, h->upd_lo, h->upd_hi, ny, h->pitch, rpt, h->p.cx, h->p.cy, h->d_resid, peer_n, peer_s, h->ghost, reverse,
                h->sync_args, h->col_args);
        else
            heat_step_vec<ARITH, RESID, THREADS, UNROLL, false, true><<<grid, THREADS, 0, h->stream>>>(
                in, out, h->upd_lo, h->upd_hi, ny, h->pitch, rpt, h->p.cx, h->p.cy, h->d_resid, peer_n, peer_s, h->ghost, reverse,
                h->sync_args, h->col_args);
    } else if (h->sync_args.counters)
        heat_step_vec<ARITH, RESID, THREADS, UNROLL, true><<<grid, THREADS, 0, h->stream>>>(
            in, out, h->upd_lo, h->upd_hi, ny, h->pitch, rpt, h->p.cx, h->p.cy, h->d_resid, peer_n, peer_s, h->ghost, reverse,
            h->sync_args);
    else
        heat_step_vec<ARITH, RESID, THREADS, UNROLL, false><<<grid, THREADS, 0, h->stream>>>(
            in, out, h->upd_lo, h->upd_hi, ny, h->pitch, rpt, h->p.cx, h->p.cy, h->d_resid, peer_n, peer_s, h->ghost, reverse,
            h->sync_args);
    h->launches++;
    return cudaGetLastError();
}

// K4: T fused updates
template <int ARITH, int T>
static cudaError_t launch_tb_t(heat_handle* h, const float* in, float* out, float* peer_n, float* peer_s, int reverse)
{
    const int rows = h->upd_hi - h->upd_lo, ny = h->nyloc;
    if (rows < 1 || ny < 3) return cudaSuccess;
#ifndef HEAT_TB_WARPS
#define HEAT_TB_WARPS 1 // warps per CTA of the temporal-blocking kernel: 1 beat 2 and 4 by 3-5 % (finer balance)
#endif
    constexpr int WARPS = HEAT_TB_WARPS;
    const int64_t wcols = (h->pitch + 119) / 120;
    int rpt = h->rows_per_tile_tb;
    if (rpt <= 0) {
        // Tile height: a tile of R rows costs R + 3T row iterations (pipeline fill), so tall tiles
        // waste less -- but the resident CTAs should work on a compact band of rows (DRAM page / TLB
        // locality, and the next pass finds the end of this one in L2).  Measured on B200 (T=4):
        // best R = 48 / 96 / 128-192 / 128-256 rows at 8192^2 / 16384^2 / 32768^2 / 65536^2, i.e. the
        // band of all resident CTAs is about 2400 rows tall whatever the row length.  Modelling the
        // CTAs as rounds of `resident` (R = 168 at 8192^2: one exact round) was 12 % slower.
        // resident CTAs per SM of this instantiation on this handle's device: queried once per handle
        if (h->tb_per_sm < 1 &&
            (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&h->tb_per_sm, heat_step_tb<ARITH, T, WARPS, false>, WARPS * 32, 0) !=
                 cudaSuccess ||
             h->tb_per_sm < 1)) {
            cudaGetLastError();
            h->tb_per_sm = 6;
        }
        const int per_sm = h->tb_per_sm;
        const long long resident = (long long)per_sm * h->sm_count, colblocks = (wcols + WARPS - 1) / WARPS;
        // (multiples of 3: the kernel's row loop is unrolled by 3)
        rpt = (int)((2400 * colblocks / resident + 12) / 24 * 24);
        if (rpt < 24) rpt = 24;
        if (rpt > 264) rpt = 264;
        // small grids: at least two CTAs per SM before anything else
        while (rpt > 12 && colblocks * ((rows + rpt - 1) / rpt) < 2LL * h->sm_count) rpt -= 6;
    }
    if (rpt < T) rpt = T;
    if ((rows + rpt - 1) / rpt > 65535) rpt = (rows + 65534) / 65535;
    dim3 grid((unsigned)((wcols + WARPS - 1) / WARPS), (unsigned)((rows + rpt - 1) / rpt));
    count_edge_ctas(h, rows, rpt, grid.x, grid.y);
    if (h->sync_args.counters)
        heat_step_tb<ARITH, T, WARPS, true><<<grid, WARPS * 32, 0, h->stream>>>(in, out, h->nloc, h->upd_lo, h->upd_hi, ny, h->pitch,
                                                                               rpt, h->p.cx, h->p.cy, peer_n, peer_s, reverse, h->sync_args);
    else
        heat_step_tb<ARITH, T, WARPS, false><<<grid, WARPS * 32, 0, h->stream>>>(in, out, h->nloc, h->upd_lo, h->upd_hi, ny, h->pitch,
                                                                                rpt, h->p.cx, h->p.cy, peer_n, peer_s, reverse, h->sync_args);
    h->launches++;
    return cudaGetLastError();
}

static cudaError_t launch_tb(heat_handle* h, int T, const float* in, float* out, float* peer_n, float* peer_s, int reverse)
{
#define TB_CASE(A)                                                                      \
    (T == 2 ? launch_tb_t<A, 2>(h, in, out, peer_n, peer_s, reverse)                    \
            : T == 3 ? launch_tb_t<A, 3>(h, in, out, peer_n, peer_s, reverse) : launch_tb_t<A, 4>(h, in, out, peer_n, peer_s, reverse))
    switch (h->p.arith) {
    case HEAT_F32_FMA: return TB_CASE(HEAT_ARITH_F32_FMA);
    case HEAT_F32_NOFMA: return TB_CASE(HEAT_ARITH_F32_NOFMA);
    default: return TB_CASE(HEAT_ARITH_F64MIX);
    }
#undef TB_CASE
}

// double-precision grids: one kernel, contraction chosen by `arith` (HEAT_F32_FMA = contract like nvcc, else round every op)
template <bool FMA, bool RESID>
static cudaError_t launch_step_f64_t(heat_handle* h, const float* in, float* out, float* peer_n, float* peer_s, int reverse)
{
    const int rows = h->upd_hi - h->upd_lo, ny = h->nyloc;
    if (rows < 1 || ny < 3) return cudaSuccess;
    const int64_t epitch = h->pitch / 2;
    int rpt = pick_rows_per_tile(h, 128);
    if (rpt < h->ghost) rpt = h->ghost;
    if ((rows + rpt - 1) / rpt > 65535) rpt = (rows + 65534) / 65535;
    dim3 grid((unsigned)((epitch / 2 + 127) / 128), (unsigned)((rows + rpt - 1) / rpt));
    count_edge_ctas(h, rows, rpt, grid.x, grid.y);
    const double cx = (double)h->p.cx, cy = (double)h->p.cy; // struct Parms stays float (M:29-32): the constants are 0.1f widened
    if (h->sync_args.counters)
        heat_step_f64<FMA, RESID, true, true><<<grid, 128, 0, h->stream>>>((const double*)in, (double*)out, h->upd_lo, h->upd_hi, ny, epitch,
                                                                          rpt, cx, cy, (unsigned long long*)h->d_resid, (double*)peer_n,
                                                                          (double*)peer_s, h->ghost, reverse, h->sync_args);
    else if (peer_n || peer_s)
        heat_step_f64<FMA, RESID, false, true><<<grid, 128, 0, h->stream>>>((const double*)in, (double*)out, h->upd_lo, h->upd_hi, ny, epitch,
                                                                           rpt, cx, cy, (unsigned long long*)h->d_resid, (double*)peer_n,
                                                                           (double*)peer_s, h->ghost, reverse, h->sync_args);
    else // a single strip's instantiation carries no peer-store code
        heat_step_f64<FMA, RESID, false, false><<<grid, 128, 0, h->stream>>>((const double*)in, (double*)out, h->upd_lo, h->upd_hi, ny, epitch,
                                                                            rpt, cx, cy, (unsigned long long*)h->d_resid, nullptr, nullptr,
                                                                            h->ghost, reverse, h->sync_args);
    h->launches++;
    return cudaGetLastError();
}

// 3-D 7-point update: planes [upd_lo, upd_hi) of the handle's rows
template <int ARITH, bool RESID>
static cudaError_t launch_step_3d_t(heat_handle* h, const float* in, float* out, float* peer_n, float* peer_s)
{
    const int planes = h->upd_hi - h->upd_lo, ny = h->ny3, nz = h->nz;
    if (planes < 1 || ny < 3 || nz < 4) return cudaSuccess;
#ifndef HEAT_3D_TY
#define HEAT_3D_TY 2 // iy rows per thread.  Measured at 640^3 / 1024^3 (GLUP/s): TY 2 x 8 warps 631 / 627, 2 x 4 579 / 589,
#endif               // 4 x 4 557 / 566 (123 registers), 8 x 4 345 / 349 (190 registers): occupancy beats fewer halo loads
#ifndef HEAT_3D_WARPS
#define HEAT_3D_WARPS 8
#endif
    constexpr int TY = HEAT_3D_TY, WARPS = HEAT_3D_WARPS;
    // plane tiles: many more CTAs than fit at once (the tail of a 1.35-wave grid cost a quarter of the time), tiles
    // no shorter than 16 planes (two planes of halo per tile); HEAT_3D_PLANES overrides
    const long long ctas_xy = (long long)((nz / 4 + 31) / 32) * ((ny - 2 + TY * WARPS - 1) / (TY * WARPS));
    int ppt = planes;
    while (ppt > 16 && ctas_xy * ((planes + ppt - 1) / ppt) < 40LL * h->sm_count) ppt = (ppt + 1) / 2;
    if (env_int("HEAT_3D_PLANES", 0) > 0) ppt = env_int("HEAT_3D_PLANES", 0);
    if ((planes + ppt - 1) / ppt > 65535) ppt = (planes + 65534) / 65535;
    dim3 grid((unsigned)((nz / 4 + 31) / 32), (unsigned)((ny - 2 + TY * WARPS - 1) / (TY * WARPS)), (unsigned)((planes + ppt - 1) / ppt));
    // every CTA of the first / last plane tile touches the neighbour on that side (same predicate as in the kernel)
    h->sync_args.count[0] = h->sync_args.count[1] = grid.x * grid.y;
    if (h->sync_args.counters)
        heat_step_3d<ARITH, RESID, TY, WARPS, true, true><<<grid, 32 * WARPS, 0, h->stream>>>(
            in, out, h->upd_lo, h->upd_hi, ny, nz, h->pitch, ppt, h->p.cx, h->p.cy, h->p.cz, h->d_resid, peer_n, peer_s, h->sync_args);
    else if (peer_n || peer_s)
        heat_step_3d<ARITH, RESID, TY, WARPS, false, true><<<grid, 32 * WARPS, 0, h->stream>>>(
            in, out, h->upd_lo, h->upd_hi, ny, nz, h->pitch, ppt, h->p.cx, h->p.cy, h->p.cz, h->d_resid, peer_n, peer_s, h->sync_args);
    else
        heat_step_3d<ARITH, RESID, TY, WARPS, false, false><<<grid, 32 * WARPS, 0, h->stream>>>(
            in, out, h->upd_lo, h->upd_hi, ny, nz, h->pitch, ppt, h->p.cx, h->p.cy, h->p.cz, h->d_resid, nullptr, nullptr, h->sync_args);
    h->launches++;
    return cudaGetLastError();
}

static cudaError_t launch_step_3d(heat_handle* h, bool resid, const float* in, float* out, float* pn, float* ps)
{
    switch (h->p.arith * 2 + (resid ? 1 : 0)) {
    case 0: return launch_step_3d_t<HEAT_ARITH_F32_FMA, false>(h, in, out, pn, ps);
    case 1: return launch_step_3d_t<HEAT_ARITH_F32_FMA, true>(h, in, out, pn, ps);
    case 2: return launch_step_3d_t<HEAT_ARITH_F32_NOFMA, false>(h, in, out, pn, ps);
    case 3: return launch_step_3d_t<HEAT_ARITH_F32_NOFMA, true>(h, in, out, pn, ps);
    case 4: return launch_step_3d_t<HEAT_ARITH_F64MIX, false>(h, in, out, pn, ps);
    default: return launch_step_3d_t<HEAT_ARITH_F64MIX, true>(h, in, out, pn, ps);
    }
}

static cudaError_t launch_step(heat_handle* h, bool resid, const float* in, float* out, float* peer_n, float* peer_s,
                               int reverse)
{
    if (h->nz > 1) return launch_step_3d(h, resid, in, out, peer_n, peer_s);
    if (h->es == 2) {
        const bool fma = h->p.arith == HEAT_F32_FMA;
        return fma ? (resid ? launch_step_f64_t<true, true>(h, in, out, peer_n, peer_s, reverse)
                            : launch_step_f64_t<true, false>(h, in, out, peer_n, peer_s, reverse))
                   : (resid ? launch_step_f64_t<false, true>(h, in, out, peer_n, peer_s, reverse)
                            : launch_step_f64_t<false, false>(h, in, out, peer_n, peer_s, reverse));
    }
    switch (h->p.arith * 2 + (resid ? 1 : 0)) {
    case 0: return launch_step_t<HEAT_ARITH_F32_FMA, false>(h, in, out, peer_n, peer_s, reverse);
    case 1: return launch_step_t<HEAT_ARITH_F32_FMA, true>(h, in, out, peer_n, peer_s, reverse);
    case 2: return launch_step_t<HEAT_ARITH_F32_NOFMA, false>(h, in, out, peer_n, peer_s, reverse);
    case 3: return launch_step_t<HEAT_ARITH_F32_NOFMA, true>(h, in, out, peer_n, peer_s, reverse);
    case 4: return launch_step_t<HEAT_ARITH_F64MIX, false>(h, in, out, peer_n, peer_s, reverse);
    default: return launch_step_t<HEAT_ARITH_F64MIX, true>(h, in, out, peer_n, peer_s, reverse);
    }
}

static const unsigned long long WAIT_TIMEOUT_NS = 20ull * 1000000000ull;

// one update: [wait for neighbours] -> step kernel (also stores the ghost rows of the
// neighbours) -> [signal neighbours] -> swap
static int enqueue_update(heat_handle* h, bool resid, int fused = 1)
{
    // which neighbours exist: 0 north, 1 south (row blocks), 2 west, 3 east (column blocks)
    const bool has[4] = {!h->external_halo && h->p.rank > 0, !h->external_halo && h->p.rank < h->p.nranks - 1,
                         h->p.rank_y > 0, h->p.rank_y < h->p.nranks_y - 1};
    const bool strips = has[0] || has[1] || has[2] || has[3];
    if (strips && !h->connected)
        return fail(HEAT_E_PEER, "block (%d,%d) of %dx%d is not connected to its neighbours", h->p.rank, h->p.rank_y, h->p.nranks,
                    h->p.nranks_y);
    const bool local = h->local_nb[0] || h->local_nb[1] || h->local_nb[2] || h->local_nb[3];
    const float* in = h->buf[h->cur];
    float* out = h->buf[1 - h->cur];
    float *pn = nullptr, *ps = nullptr;
    h->col_args = col_peers{};
    h->col_args.edge_first = h->cols_first;
    if (strips) {
        if (local) {
            // neighbour's epoch `seq` (its previous update, or its upload) must be complete
            for (int side = 0; side < 4; side++)
                if (h->local_nb[side]) CK(cudaStreamWaitEvent(h->stream, h->local_nb[side]->ev_done[h->seq & 1], 0));
        }
        // my first owned row (local row 1) is the north neighbour's bottom ghost row, and vice versa
        if (has[0]) pn = h->peer[0].buf[1 - h->cur] + (int64_t)(h->peer[0].nloc - h->ghost) * h->pitch;
        if (has[1]) ps = h->peer[1].buf[1 - h->cur];
        // my first owned column is the west neighbour's last (ghost) column, my last its first
        // (dense arrays, one float per local row; column neighbours have my rows: [buffer][side][row])
        const int64_t n = h->nloc;
        if (has[2]) {
            h->col_args.w = h->peer[2].gcol + ((1 - h->cur) * 2 + 1) * n; // its EAST ghost, new buffer
            h->col_args.gw_in = h->gcol + (h->cur * 2 + 0) * n;
        }
        if (has[3]) {
            h->col_args.e = h->peer[3].gcol + ((1 - h->cur) * 2 + 0) * n; // its WEST ghost, new buffer
            h->col_args.ge_in = h->gcol + (h->cur * 2 + 1) * n;
        }
    }
    if (resid) CK(cudaMemsetAsync(h->d_resid, 0, 2 * sizeof(unsigned), h->stream));
    // flag-ordered strips (one process per GPU): the wait and the signal are fused into the pass
    h->sync_args = strip_sync{};
    if (strips && !local) {
        strip_sync& sy = h->sync_args;
        for (int side = 0; side < 4; side++) {
            sy.wait[side] = has[side] ? &h->flags[side] : nullptr;
            sy.sig[side] = has[side] ? h->peer[side].flag : nullptr;
        }
        sy.wait_value = h->seq;
        sy.sig_value = h->seq + 1;
        sy.counters = h->d_edge_count;
        sy.status = h->d_status;
        sy.timeout_ns = WAIT_TIMEOUT_NS;
        sy.prof = h->d_wait_prof;
    }
#ifdef HEAT_BOUNDS_CHECK
    {
        mem_window w = {};
        const size_t n = (size_t)h->nloc * h->pitch;
        w.lo[0] = in, w.hi[0] = in + n;
        w.lo[1] = out, w.hi[1] = out + n;
        if (pn) w.lo[2] = pn, w.hi[2] = pn + (size_t)h->ghost * h->pitch;
        if (ps) w.lo[3] = ps, w.hi[3] = ps + (size_t)h->ghost * h->pitch;
        // one window for the whole device: the debug build runs one pass at a time
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpyToSymbol(g_window, &w, sizeof w));
    }
#endif
    const int reverse = h->serpentine ? (int)(h->sweeps & 1) : 0;
    h->sweeps++;
    if (fused > 1) CK(launch_tb(h, fused, in, out, pn, ps, reverse));
    else CK(launch_step(h, resid, in, out, pn, ps, reverse));
    h->seq++;
    if (strips) {
        if (local) CK(cudaEventRecord(h->ev_done[h->seq & 1], h->stream));
    }
    CK(cudaGetLastError());
    h->cur = 1 - h->cur;
    h->it += fused;
    return HEAT_OK;
}

// ---- K5: many updates in one cooperative launch, field resident in shared memory
static const int RESIDENT_MIN_UPDATES = 4;

template <int ARITH>
static int resident_prepare_t(heat_handle* h)
{
    // one CTA per SM at most; every band at least G rows tall so that a ghost zone never reaches past a neighbour
    const int interior = h->nloc - 2;
    int G = env_int("HEAT_RESIDENT_G", 6); // measured at 1024^2: G=1 203, 2 313, 4 412, 6 439, 8 415 GLUP/s
    if (G < 1) G = 1;
    if (G > 16) G = 16;
    int ctas = interior / G;
    if (ctas > h->sm_count) ctas = h->sm_count;
    if (ctas < 1) ctas = 1;
    if (interior / ctas < G) G = interior / ctas; // a single short band
    const int ncg = (int)(h->pitch / 4);
    int tpr = (ncg + 31) / 32 * 32;
    if (tpr > 1024) tpr = 1024;
    int rgroups = 1024 / tpr;
    if (rgroups > interior / ctas + 2) rgroups = interior / ctas + 2;
    if (rgroups < 1) rgroups = 1;
    int dev_smem = 0, coop = 0, per_sm = 0;
    CK(cudaDeviceGetAttribute(&dev_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->p.device));
    CK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, h->p.device));
    size_t smem = 0;
    for (;; G--) { // fewer updates per exchange if the ghost zones do not fit
        smem = 2 * (size_t)(interior / ctas + 1 + 2 * G) * h->pitch * sizeof(float);
        if (smem <= (size_t)dev_smem || G == 1) break;
    }
    if (!coop || smem > (size_t)dev_smem) return 1; // does not fit: not an error
    CK(cudaFuncSetAttribute(heat_resident<ARITH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, heat_resident<ARITH>, tpr * rgroups, smem));
    if (per_sm < 1) return 1;
    CK(cudaMalloc((void**)&h->res_halo, (size_t)ctas * 4 * G * h->pitch * sizeof(float)));
    CK(cudaMalloc((void**)&h->res_flags, (size_t)ctas * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(h->res_flags, 0, (size_t)ctas * sizeof(unsigned long long), h->stream));
    h->res_ctas = ctas;
    h->res_threads = tpr * rgroups;
    h->res_tpr = tpr;
    h->res_ghost = G;
    h->res_smem = smem;
    return HEAT_OK;
}

// > 0: cannot be used for this handle; 0: ready; < 0: error
static int resident_prepare(heat_handle* h)
{
    if (h->res_ctas > 0) return HEAT_OK;
    if (h->res_ctas < 0 || !h->use_resident || h->p.nranks != 1 || h->p.nranks_y != 1 || h->nloc < 3 || h->p.ny < 3) return 1;
    int rc;
    switch (h->p.arith) {
    case HEAT_F32_FMA: rc = resident_prepare_t<HEAT_ARITH_F32_FMA>(h); break;
    case HEAT_F32_NOFMA: rc = resident_prepare_t<HEAT_ARITH_F32_NOFMA>(h); break;
    default: rc = resident_prepare_t<HEAT_ARITH_F64MIX>(h); break;
    }
    if (rc > 0) h->res_ctas = -1; // remember that it does not fit
    return rc;
}

static int launch_resident(heat_handle* h, int updates)
{
    resident_args a;
    const int new_cur = h->cur ^ (updates & 1);
    a.in = h->buf[h->cur];
    a.out_final = h->buf[new_cur];
    a.out_prev = h->buf[1 - new_cur];
    a.nrows = h->nloc;
    a.ny = h->p.ny;
    a.pitch = h->pitch;
    a.steps = updates;
    a.ghost = h->res_ghost;
    a.threads_per_row = h->res_tpr;
    a.cx = h->p.cx;
    a.cy = h->p.cy;
    a.halo = h->res_halo;
    a.flags = h->res_flags;
    a.epoch0 = h->res_epoch;
    a.status = h->d_status;
    a.timeout_ns = WAIT_TIMEOUT_NS;
    void* kargs[] = {&a};
    const void* fn = h->p.arith == HEAT_F32_FMA     ? (const void*)heat_resident<HEAT_ARITH_F32_FMA>
                     : h->p.arith == HEAT_F32_NOFMA ? (const void*)heat_resident<HEAT_ARITH_F32_NOFMA>
                                                    : (const void*)heat_resident<HEAT_ARITH_F64MIX>;
    CK(cudaLaunchCooperativeKernel(fn, dim3(h->res_ctas), dim3(h->res_threads), kargs, h->res_smem, h->stream));
    h->res_epoch += (unsigned long long)updates; // an upper bound of the exchanges made: flags stay monotonic
    h->cur = new_cur;
    h->it += updates;
    h->sweeps += updates;
    h->seq += (unsigned long long)updates;
    h->launches++;
    return HEAT_OK;
}

// GRAPH_PASSES consecutive passes as one CUDA graph (captured on a side stream, launched on the
// handle's stream).  Replaces per-kernel launches where they dominate: the loop of C:204-211 costs
// one launch per update, 2-4 us each in the reference's own Table 6.
static const int GRAPH_PASSES = 16;

// Captures GRAPH_PASSES passes starting from (newest buffer `cur`, sweep direction `dir`) and instantiates them.
static int build_pass_graph(heat_handle* h, int cur, int dir, int fused, cudaGraphExec_t* exec)
{
    if (!h->capture_stream) CK(cudaStreamCreateWithFlags(&h->capture_stream, cudaStreamNonBlocking));
    const cudaStream_t user = h->stream;
    const long long it0 = h->it, sweeps0 = h->sweeps;
    const int cur0 = h->cur;
    const unsigned long long seq0 = h->seq;
    const int64_t launches0 = h->launches;
    cudaGraph_t graph = nullptr;
    h->stream = h->capture_stream;
    h->cur = cur;
    h->sweeps = dir;
    cudaError_t e = cudaStreamBeginCapture(h->capture_stream, cudaStreamCaptureModeRelaxed);
    int rc = HEAT_OK;
    if (e == cudaSuccess) {
        for (int k = 0; k < GRAPH_PASSES && !rc; k++) rc = enqueue_update(h, false, fused);
        e = cudaStreamEndCapture(h->capture_stream, &graph);
    }
    h->stream = user;
    h->it = it0, h->sweeps = sweeps0, h->seq = seq0, h->launches = launches0, h->cur = cur0; // nothing ran yet
    if (rc) {
        if (graph) cudaGraphDestroy(graph);
        return rc;
    }
    if (e != cudaSuccess || !graph) return fail(HEAT_E_CUDA, "stream capture failed: %s", cudaGetErrorString(e));
    e = cudaGraphInstantiate(exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) return fail(HEAT_E_CUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(e));
    // the first launch of an executable graph uploads it to the device: do that now, not in the caller's loop
    e = cudaGraphUpload(*exec, user);
    if (e != cudaSuccess) cudaGetLastError(); // an optimisation only
    return HEAT_OK;
}

static bool graphs_apply(const heat_handle* h) { return h->p.nranks == 1 && h->p.nranks_y == 1 && h->use_graphs; }

// Everything a first heat_enqueue / heat_run would otherwise set up lazily, done when the field arrives
// (heat_upload*, heat_init_inidat): the pass graphs for both buffer parities and both sweep directions, the K4
// occupancy query, and the resident-field kernel's scratch.  A caller that times its first loop times the
// loop, not cudaStreamBeginCapture + cudaGraphInstantiate (round 1's bench did: 0.64 ms inside a 2 ms region).
static int prepare(heat_handle* h)
{
    if (h->p.nranks == 1 && h->p.nranks_y == 1 && h->use_resident) {
        int rc = resident_prepare(h);
        if (rc < 0) return rc;
    }
    // (a field that runs resident in shared memory never replays the pass graphs)
    if (!graphs_apply(h) || h->res_ctas > 0 || h->upd_hi - h->upd_lo < 1 || h->nyloc < 3) return HEAT_OK;
    const int fused = h->p.temporal_block;
    for (int cur = 0; cur < 2; cur++)
        for (int dir = 0; dir < (h->serpentine ? 2 : 1); dir++) {
            cudaGraphExec_t& exec = h->graphs[cur][dir][fused > 1 ? 1 : 0];
            if (exec) continue;
            int rc = build_pass_graph(h, cur, dir, fused, &exec);
            if (rc) return rc;
        }
    return HEAT_OK;
}

static int launch_pass_graph(heat_handle* h, int fused)
{
    const int dir = h->serpentine ? (int)(h->sweeps & 1) : 0;
    cudaGraphExec_t& exec = h->graphs[h->cur][dir][fused > 1 ? 1 : 0];
    if (!exec) { // only reached when prepare() was skipped (a handle used without an upload)
        int rc = build_pass_graph(h, h->cur, dir, fused, &exec);
        if (rc) return rc;
    }
    CK(cudaGraphLaunch(exec, h->stream));
    h->it += (long long)GRAPH_PASSES * fused;
    h->sweeps += GRAPH_PASSES;
    h->seq += GRAPH_PASSES;
    h->launches += GRAPH_PASSES;
    return HEAT_OK;
}

// A handle that is used before any field arrived starts from a zero field (heat_create does not pay for that).
static int ensure_loaded(heat_handle* h)
{
    if (h->loaded) return HEAT_OK;
    CK(cudaSetDevice(h->p.device));
    const size_t bytes = (size_t)h->nloc * h->pitch * sizeof(float);
    CK(cudaMemsetAsync(h->buf[0], 0, bytes, h->stream));
    CK(cudaMemsetAsync(h->buf[1], 0, bytes, h->stream));
    h->loaded = true;
    return HEAT_OK;
}

extern "C" int heat_group_enqueue(heat_handle** hs, int n, int updates)
{
    if (!hs || n < 1 || updates < 0) return fail(HEAT_E_INVALID, "heat_group_enqueue: bad arguments");
    for (int k = 0; k < n; k++) {
        if (!hs[k]) return fail(HEAT_E_INVALID, "heat_group_enqueue: null handle");
        int rc = ensure_loaded(hs[k]);
        if (rc) return rc;
    }
    // pass by pass across the strips: the event ordering between local strips relies on it
    const int T = hs[0]->p.temporal_block;
    for (int i = 0; i < updates;) {
        const int fused = (T > 1 && updates - i >= T) ? T : 1;
        // small grids: the field stays in shared memory for all the remaining updates of this stretch
        if (n == 1 && updates - i >= RESIDENT_MIN_UPDATES) {
            CK(cudaSetDevice(hs[0]->p.device));
            int rc = resident_prepare(hs[0]);
            if (rc < 0) return rc;
            if (rc == 0) {
                if (launch_resident(hs[0], updates - i) == HEAT_OK) break;
                // e.g. the GPU cannot co-schedule the CTAs right now: use ordinary passes from here on
                cudaGetLastError();
                hs[0]->res_ctas = -1;
            }
        }
        // a single strip replays a captured graph of GRAPH_PASSES passes while enough updates are left
        // (the launch-bound regime of small grids; state is unchanged after an even number of passes)
        if (n == 1 && hs[0]->p.nranks == 1 && hs[0]->p.nranks_y == 1 && hs[0]->use_graphs && updates - i >= GRAPH_PASSES * fused) {
            CK(cudaSetDevice(hs[0]->p.device));
            int rc = launch_pass_graph(hs[0], fused);
            if (rc) return rc;
            i += GRAPH_PASSES * fused;
            continue;
        }
        for (int k = 0; k < n; k++) {
            CK(cudaSetDevice(hs[k]->p.device));
            int rc = enqueue_update(hs[k], false, fused);
            if (rc) return rc;
        }
        i += fused;
    }
    return HEAT_OK;
}

extern "C" int heat_enqueue(heat_handle* h, int updates)
{
    if (!h) return fail(HEAT_E_INVALID, "heat_enqueue: null handle");
    if (h->local_nb[0] || h->local_nb[1] || h->local_nb[2] || h->local_nb[3])
        return fail(HEAT_E_STATE, "heat_enqueue: locally connected strips run through heat_group_*");
    return heat_group_enqueue(&h, 1, updates);
}

extern "C" int heat_sync(heat_handle* h)
{
    if (!h) return fail(HEAT_E_INVALID, "heat_sync: null handle");
    CK(cudaSetDevice(h->p.device));
    CK(cudaStreamSynchronize(h->stream));
    if (*h->h_status) return fail(HEAT_E_TIMEOUT, "strip %d: neighbour did not arrive within 20 s", h->p.rank);
    return HEAT_OK;
}

extern "C" int heat_set_allreduce_max(heat_handle* h, heat_allreduce_max_fn fn, void* ctx)
{
    if (!h) return fail(HEAT_E_INVALID, "heat_set_allreduce_max: null handle");
    h->allreduce = fn;
    h->allreduce_ctx = ctx;
    return HEAT_OK;
}

// max|old-new| as the float the convergence test compares (a double grid's maximum is rounded UP to float, so
// that `r < eps` can never hold for a double residual that is not below eps)
static float resid_value(const heat_handle* h)
{
    if (h->es == 2) {
        double d;
        memcpy(&d, h->h_resid, sizeof d);
        float f = (float)d;
        if ((double)f < d) f = nextafterf(f, INFINITY);
        return f;
    }
    float f;
    memcpy(&f, h->h_resid, sizeof f);
    return f;
}

// The time loop (C:204-238 / M:159-265) over one strip or over all strips of this process.
extern "C" int heat_group_run(heat_handle** hs, int n, int steps, int* updates_done, int* converged)
{
    if (!hs || n < 1 || steps < 0) return fail(HEAT_E_INVALID, "heat_run: bad arguments");
    for (int k = 0; k < n; k++)
        if (!hs[k]) return fail(HEAT_E_INVALID, "heat_run: null handle");
    for (int k = 0; k < n; k++) {
        int rc = ensure_loaded(hs[k]);
        if (rc) return rc;
    }
    heat_handle* h0 = hs[0];
    const long long total = (long long)steps + (h0->p.step_convention == HEAT_STEPS_MPI ? 1 : 0);
    const int K = h0->p.conv_every;
    long long done = 0;
    int conv = 0;
    for (int k = 0; k < n; k++) {
        CK(cudaSetDevice(hs[k]->p.device));
        CK(cudaEventRecord(hs[k]->ev0, hs[k]->stream));
    }
    while (done < total && !conv) {
        // how many plain updates come before the next checked one?
        long long plain = total - done;
        if (K > 0) {
            const long long ph = h0->p.conv_phase == HEAT_PHASE_CUDA ? h0->it % K : (h0->it + 1) % K;
            plain = (K - ph) % K; // 0: update number `it` is itself a checked one
            if (plain > total - done) plain = total - done;
        }
        if (plain > 0) { // fused passes and graphs are chosen inside; nothing for the host to decide
            int rc = heat_group_enqueue(hs, n, (int)plain);
            if (rc) return rc;
            done += plain;
            continue;
        }
        const bool check = true;
        for (int k = 0; k < n; k++) {
            CK(cudaSetDevice(hs[k]->p.device));
            int rc = enqueue_update(hs[k], check, 1);
            if (rc) return rc;
            if (check) {
                // one 4-byte read-back per K updates: the only host sync in the loop
                CK(cudaMemcpyAsync(hs[k]->h_resid, hs[k]->d_resid, 2 * sizeof(unsigned), cudaMemcpyDeviceToHost, hs[k]->stream));
                CK(cudaEventRecord(hs[k]->ev_resid, hs[k]->stream));
            }
        }
        done += 1;
        if (check) {
            float r = 0.0f;
            for (int k = 0; k < n; k++) {
                CK(cudaEventSynchronize(hs[k]->ev_resid));
                r = fmaxf(r, resid_value(hs[k]));
            }
            // strips in other processes: one small max-allreduce (replaces MPI_Allreduce(LAND), M:255)
            if (h0->allreduce && h0->allreduce(&r, h0->allreduce_ctx) != 0)
                return fail(HEAT_E_PEER, "heat_run: allreduce callback failed");
            if (r < h0->p.conv_eps) conv = 1; // M:245 / C:67
        }
    }
    float ms = 0.0f;
    for (int k = 0; k < n; k++) {
        CK(cudaSetDevice(hs[k]->p.device));
        CK(cudaEventRecord(hs[k]->ev1, hs[k]->stream));
    }
    for (int k = 0; k < n; k++) {
        int rc = heat_sync(hs[k]);
        if (rc) return rc;
        float t = 0.0f;
        CK(cudaEventElapsedTime(&t, hs[k]->ev0, hs[k]->ev1));
        ms = fmaxf(ms, t);
    }
    for (int k = 0; k < n; k++) hs[k]->last_ms = ms;
    if (updates_done) *updates_done = (int)done;
    if (converged) *converged = conv;
    return HEAT_OK;
}

extern "C" int heat_run(heat_handle* h, int steps, int* updates_done, int* converged)
{
    if (!h) return fail(HEAT_E_INVALID, "heat_run: null handle");
    if (h->local_nb[0] || h->local_nb[1] || h->local_nb[2] || h->local_nb[3])
        return fail(HEAT_E_STATE, "heat_run: locally connected strips run through heat_group_run");
    if ((h->p.nranks > 1 || h->p.nranks_y > 1) && h->p.conv_every > 0 && !h->allreduce)
        return fail(HEAT_E_STATE, "heat_run: strips in several processes need heat_set_allreduce_max for the convergence check");
    return heat_group_run(&h, 1, steps, updates_done, converged);
}

extern "C" int heat_elapsed_ms(heat_handle* h, float* ms)
{
    if (!h || !ms) return fail(HEAT_E_INVALID, "heat_elapsed_ms: bad arguments");
    *ms = h->last_ms;
    return HEAT_OK;
}

extern "C" int heat_launch_count(heat_handle* h, int64_t* launches)
{
    if (!h || !launches) return fail(HEAT_E_INVALID, "heat_launch_count: bad arguments");
    *launches = h->launches;
    return HEAT_OK;
}

// Diagnostics of the fused strip ordering (created with HEAT_PROFILE_WAIT=1 in the environment): how long the edge
// CTAs of this strip's update kernels have spun on each neighbour's flag since the last call, and how often.
extern "C" int heat_debug_wait_stats(heat_handle* h, unsigned long long* ns_per_side, unsigned long long* waits_per_side)
{
    if (!h || !ns_per_side || !waits_per_side) return fail(HEAT_E_INVALID, "heat_debug_wait_stats: bad arguments");
    if (!h->d_wait_prof) return fail(HEAT_E_STATE, "heat_debug_wait_stats: create the handle with HEAT_PROFILE_WAIT=1");
    CK(cudaSetDevice(h->p.device));
    unsigned long long v[8];
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaMemcpy(v, h->d_wait_prof, sizeof v, cudaMemcpyDeviceToHost));
    CK(cudaMemset(h->d_wait_prof, 0, sizeof v));
    for (int k = 0; k < 4; k++) ns_per_side[k] = v[k], waits_per_side[k] = v[4 + k];
    return HEAT_OK;
}

extern "C" int heat_residual(heat_handle* h, float* residual)
{
    if (!h || !residual) return fail(HEAT_E_INVALID, "heat_residual: bad arguments");
    CK(cudaSetDevice(h->p.device));
    CK(cudaMemsetAsync(h->d_resid, 0, 2 * sizeof(unsigned), h->stream));
    if (h->upd_hi > h->upd_lo && h->nyloc >= 3) {
        if (h->es == 2)
            heat_absdiff_max_f64<<<h->sm_count * 8, 256, 0, h->stream>>>((const double*)h->buf[0], (const double*)h->buf[1], h->upd_lo,
                                                                        h->upd_hi, h->nyloc, h->pitch / 2, (unsigned long long*)h->d_resid);
        else
            heat_absdiff_max<<<h->sm_count * 8, 256, 0, h->stream>>>(h->buf[0], h->buf[1], h->upd_lo, h->upd_hi, h->nyloc, h->pitch, h->d_resid);
        h->launches++;
        CK(cudaGetLastError());
    }
    CK(cudaMemcpyAsync(h->h_resid, h->d_resid, 2 * sizeof(unsigned), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    *residual = resid_value(h);
    return HEAT_OK;
}

// ------------------------------------------------------------------ data movement
extern "C" int heat_set_halo_mode(heat_handle* h, int mode)
{
    if (!h || (mode != HEAT_HALO_P2P && mode != HEAT_HALO_EXTERNAL)) return fail(HEAT_E_INVALID, "heat_set_halo_mode: bad arguments");
    h->external_halo = (mode == HEAT_HALO_EXTERNAL);
    return HEAT_OK;
}

// Flag-ordered strips (one process per GPU): a neighbour's last pass may still be storing into this strip's
// ghost rows when heat_sync() returns here -- a pass waits only for the neighbours' PREVIOUS pass.  Before new
// data is written over the field, wait on the device until every neighbour has finished the pass this strip
// finished last, so that no stale ghost-row store can land on top of the upload.
static bool has_side(const heat_handle* h, int side);
static int quiesce_neighbours(heat_handle* h)
{
    if (!h->connected || h->external_halo || h->seq == 0) return HEAT_OK;
    if (h->local_nb[0] || h->local_nb[1] || h->local_nb[2] || h->local_nb[3]) return HEAT_OK; // events order those
    bool any = false;
    const unsigned long long* f[4];
    for (int side = 0; side < 4; side++) {
        f[side] = has_side(h, side) ? &h->flags[side] : nullptr;
        any = any || f[side];
    }
    if (!any) return HEAT_OK;
    heat_wait<<<1, 32, 0, h->stream>>>(f[0], f[1], f[2], f[3], h->seq, WAIT_TIMEOUT_NS, h->d_status);
    h->launches++;
    CK(cudaGetLastError());
    return HEAT_OK;
}

static int after_upload(heat_handle* h)
{
    h->cur = 0;
    h->it = 0;
    h->loaded = true;
    // an upload counts as one epoch of the strip flags: neighbours may not store into this
    // strip's ghost rows before the upload has landed
    h->seq++;
    const bool has[4] = {h->p.rank > 0, h->p.rank < h->p.nranks - 1, h->p.rank_y > 0, h->p.rank_y < h->p.nranks_y - 1};
    if (h->gcol) { // the dense ghost-column arrays start as what the upload put into the field's ghost columns
        heat_gather_ghost_columns<<<(h->nloc + 255) / 256, 256, 0, h->stream>>>(h->buf[0], h->buf[1], h->nloc, h->nyloc, h->pitch, h->gcol);
        h->launches++;
        CK(cudaGetLastError());
    }
    if ((has[0] || has[1] || has[2] || has[3]) && h->connected && !h->external_halo) {
        if (h->local_nb[0] || h->local_nb[1] || h->local_nb[2] || h->local_nb[3]) {
            CK(cudaEventRecord(h->ev_done[h->seq & 1], h->stream));
        } else {
            heat_signal<<<1, 32, 0, h->stream>>>(has[0] ? h->peer[0].flag : nullptr, has[1] ? h->peer[1].flag : nullptr,
                                                 has[2] ? h->peer[2].flag : nullptr, has[3] ? h->peer[3].flag : nullptr, h->seq);
            h->launches++;
            CK(cudaGetLastError());
        }
    }
    int rc = prepare(h);
    if (rc) return rc;
    CK(cudaStreamSynchronize(h->stream));
    return HEAT_OK;
}

// src_first_row: the global row that src's first row holds (0 for a full field, h->lo for a strip's own window)
static int upload_from(heat_handle* h, const void* src, int src_first_row, cudaMemcpyKind kind)
{
    if (!h || !src) return fail(HEAT_E_INVALID, "heat_upload: bad arguments");
    if ((h->p.nranks > 1 || h->p.nranks_y > 1) && !h->connected && !h->external_halo)
        return fail(HEAT_E_STATE, "heat_upload: connect the strips first");
    CK(cudaSetDevice(h->p.device));
    int rc = quiesce_neighbours(h);
    if (rc) return rc;
    const size_t esz = (size_t)h->es * sizeof(float); // bytes per element
    const char* rows = (const char*)src + ((size_t)(h->lo - src_first_row) * h->p.ny + h->lo_c) * esz;
    // C:197-198: the initial field goes into both buffers -- over PCIe once, the second copy on the device
    CK(cudaMemcpy2DAsync(h->buf[0], h->pitch * sizeof(float), rows, (size_t)h->p.ny * esz, (size_t)h->nyloc * esz, h->nloc, kind,
                         h->stream));
    CK(cudaMemcpyAsync(h->buf[1], h->buf[0], (size_t)h->nloc * h->pitch * sizeof(float), cudaMemcpyDeviceToDevice, h->stream));
    return after_upload(h);
}

static int wrong_type(const heat_handle* h, int want_es, const char* fn)
{
    if (h && h->es != want_es)
        return fail(HEAT_E_STATE, "%s: the handle's grid is %s", fn, h->es == 2 ? "double (use the _f64 entry points)" : "float");
    return HEAT_OK;
}

extern "C" int heat_upload(heat_handle* h, const float* full_grid)
{
    int rc = wrong_type(h, 1, "heat_upload");
    return rc ? rc : upload_from(h, full_grid, 0, cudaMemcpyHostToDevice);
}
extern "C" int heat_upload_f64(heat_handle* h, const double* full_grid)
{
    int rc = wrong_type(h, 2, "heat_upload_f64");
    return rc ? rc : upload_from(h, full_grid, 0, cudaMemcpyHostToDevice);
}
extern "C" int heat_upload_device(heat_handle* h, const void* d_full_grid) { return upload_from(h, d_full_grid, 0, cudaMemcpyDeviceToDevice); }

extern "C" int heat_init_inidat(heat_handle* h)
{
    if (!h) return fail(HEAT_E_INVALID, "heat_init_inidat: null handle");
    if ((h->p.nranks > 1 || h->p.nranks_y > 1) && !h->connected && !h->external_halo)
        return fail(HEAT_E_STATE, "heat_init_inidat: connect the strips first");
    CK(cudaSetDevice(h->p.device));
    {
        int rc = quiesce_neighbours(h);
        if (rc) return rc;
    }
    if (h->nz > 1)
        heat_inidat3d_kernel<<<h->sm_count * 8, 256, 0, h->stream>>>(h->buf[0], h->buf[1], h->p.nx, h->ny3, h->nz, h->lo, h->nloc, h->pitch);
    else if (h->es == 2)
        heat_inidat_f64_kernel<<<h->sm_count * 8, 256, 0, h->stream>>>((double*)h->buf[0], (double*)h->buf[1], h->p.nx, h->p.ny, h->lo,
                                                                       h->nloc, h->pitch / 2, h->lo_c, h->nyloc);
    else
        heat_inidat_kernel<<<h->sm_count * 8, 256, 0, h->stream>>>(h->buf[0], h->buf[1], h->p.nx, h->p.ny, h->lo, h->nloc, h->pitch,
                                                                   h->lo_c, h->nyloc);
    h->launches++;
    CK(cudaGetLastError());
    return after_upload(h);
}

static int download_to(heat_handle* h, void* full_grid);
extern "C" int heat_download(heat_handle* h, float* full_grid)
{
    int rc = wrong_type(h, 1, "heat_download");
    return rc ? rc : download_to(h, full_grid);
}
extern "C" int heat_download_f64(heat_handle* h, double* full_grid)
{
    int rc = wrong_type(h, 2, "heat_download_f64");
    return rc ? rc : download_to(h, full_grid);
}

static int download_to(heat_handle* h, void* full_grid)
{
    if (!h || !full_grid) return fail(HEAT_E_INVALID, "heat_download: bad arguments");
    CK(cudaSetDevice(h->p.device));
    {
        int rc = ensure_loaded(h);
        if (rc) return rc;
    }
    // only the rows and columns this block owns (ghost rows/columns belong to the neighbours)
    const int first = h->row0 - h->lo, firstc = h->col0 - h->lo_c;
    const size_t esz = (size_t)h->es * sizeof(float);
    CK(cudaMemcpy2DAsync((char*)full_grid + ((size_t)h->row0 * h->p.ny + h->col0) * esz, (size_t)h->p.ny * esz,
                         h->buf[h->cur] + (size_t)first * h->pitch + (size_t)firstc * h->es, h->pitch * sizeof(float),
                         (size_t)h->cols_own * esz, h->rows_own, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return HEAT_OK;
}

// A strip's own window only: what a rank that never holds the full field passes in / gets back.
extern "C" int heat_upload_local(heat_handle* h, const float* local_rows)
{
    if (!h || !local_rows) return fail(HEAT_E_INVALID, "heat_upload_local: bad arguments");
    if (h->p.nranks_y > 1) return fail(HEAT_E_STATE, "heat_upload_local: row strips only (column blocks use heat_upload)");
    if (h->es != 1) return wrong_type(h, 1, "heat_upload_local");
    // same copy as heat_upload, from a source whose first row is the strip's first local row
    return upload_from(h, local_rows, h->lo, cudaMemcpyHostToDevice);
}

extern "C" int heat_download_local(heat_handle* h, float* own_rows)
{
    if (!h || !own_rows) return fail(HEAT_E_INVALID, "heat_download_local: bad arguments");
    if (h->p.nranks_y > 1) return fail(HEAT_E_STATE, "heat_download_local: row strips only (column blocks use heat_download)");
    if (h->es != 1) return wrong_type(h, 1, "heat_download_local");
    CK(cudaSetDevice(h->p.device));
    {
        int rc = ensure_loaded(h);
        if (rc) return rc;
    }
    const int first = h->row0 - h->lo;
    CK(cudaMemcpy2DAsync(own_rows, (size_t)h->p.ny * sizeof(float), h->buf[h->cur] + (size_t)first * h->pitch,
                         h->pitch * sizeof(float), (size_t)h->p.ny * sizeof(float), h->rows_own, cudaMemcpyDeviceToHost,
                         h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return HEAT_OK;
}

extern "C" int heat_device_view(heat_handle* h, int which, void** d_ptr, int64_t* pitch, int* local_rows, int* first_global_row)
{
    if (!h || (which != 0 && which != 1)) return fail(HEAT_E_INVALID, "heat_device_view: bad arguments");
    if (d_ptr) *d_ptr = h->buf[which == 0 ? h->cur : 1 - h->cur];
    if (pitch) *pitch = h->pitch / h->es;
    if (local_rows) *local_rows = h->nloc;
    if (first_global_row) *first_global_row = h->lo;
    return HEAT_OK;
}

// ------------------------------------------------------------------ one-shot (C:176-242)
// heat_solve() on a host field is bound by PCIe, not by the kernels: 8 bytes per point cross the bus once in each
// direction, against 8 bytes per point and UPDATE through HBM at ~100 times the bandwidth.  The reference does
// H2D, loop, D2H one after the other (C:197-198, 204-238, 242).  Here the three overlap: the field arrives in
// bands of rows on a copy stream; as soon as band i has landed, the compute stream runs ALL passes on it,
// pass p shifted up by p*s rows (time skewing: pass p of band i covers rows [i*B - p*s, (i+1)*B - p*s), so every
// row it reads already holds the level below, and nothing it overwrites is still needed); the rows that are final
// after band i leave on a second copy stream while band i+1 is still arriving.  Same kernels, same two buffers,
// same bits as the plain loop -- only the order in which (row, update) pairs are computed differs.
// Used when there is no convergence check (an early stop needs whole-field residuals) and the total skew is
// small against the field; HEAT_SOLVE_STREAMED=0 turns it off.
static bool solve_can_stream(const heat_handle* h, long long updates)
{
#ifdef HEAT_BOUNDS_CHECK
    return false; // the debug build's access windows are set per pass by enqueue_update
#endif
    if (env_int("HEAT_SOLVE_STREAMED", 1) == 0) return false;
    if (h->p.nranks != 1 || h->p.nranks_y != 1 || h->p.conv_every > 0 || updates < 1) return false;
    if (h->p.nx < 64 || h->p.ny < 3) return false;
    const int T = h->p.temporal_block;
    const long long passes = updates / T + updates % T;
    return passes * T * 4 <= h->p.nx; // the skew costs passes*T extra rows of pipeline: keep it under a quarter
}

static int solve_streamed(heat_handle* h, void* grid_, long long updates)
{
    const int nx = h->p.nx, ny = h->p.ny, T = h->p.temporal_block;
    const int nfused = (int)(T > 1 ? updates / T : 0), P = (int)(T > 1 ? updates / T + updates % T : updates);
    const int s = T;             // rows a pass is shifted against the one before: >= what a pass reads beyond its rows
    const long long S = (long long)P * s;
    const size_t row_bytes = (size_t)ny * h->es * sizeof(float);
    char* const grid = (char*)grid_;
    // band height: a sixteenth of the field, between 8 and 128 MiB (small enough to fill the pipeline quickly, large
    // enough that a pass over one band is several waves of CTAs); HEAT_SOLVE_BAND_ROWS overrides
    size_t band_bytes = (size_t)nx * row_bytes / 16;
    if (band_bytes < ((size_t)8 << 20)) band_bytes = (size_t)8 << 20;
    if (band_bytes > ((size_t)128 << 20)) band_bytes = (size_t)128 << 20;
    long long B = (long long)((band_bytes + row_bytes - 1) / row_bytes);
    B = env_int("HEAT_SOLVE_BAND_ROWS", (int)B);
    if (B < 2 * s) B = 2 * s;
    if (B < 8) B = 8;
    if (B > nx) B = nx;

    CK(cudaSetDevice(h->p.device));
    cudaStream_t in_s = nullptr, out_s = nullptr;
    cudaEvent_t ev_in = nullptr, ev_done = nullptr;
    int rc = HEAT_OK;
    const int upd_lo0 = h->upd_lo, upd_hi0 = h->upd_hi;
#define CKS(call)                                                                                                  \
    do {                                                                                                           \
        cudaError_t e_ = (call);                                                                                   \
        if (e_ != cudaSuccess && !rc)                                                                              \
            rc = fail(HEAT_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__);    \
    } while (0)
    CKS(cudaStreamCreateWithFlags(&in_s, cudaStreamNonBlocking));
    CKS(cudaStreamCreateWithFlags(&out_s, cudaStreamNonBlocking));
    CKS(cudaEventCreateWithFlags(&ev_in, cudaEventDisableTiming));
    CKS(cudaEventCreateWithFlags(&ev_done, cudaEventDisableTiming));
    CKS(cudaEventRecord(h->ev0, h->stream));
    float* const fin = h->buf[P & 1]; // pass p reads buf[(p-1)&1] and writes buf[p&1]
    long long sent = 0;               // rows [1, sent) are final and on their way to the host
    sent = 1;
    for (long long i = 0; !rc && sent < nx - 1; i++) {
        const long long b0 = i * B, b1 = (i + 1) * B < nx ? (i + 1) * B : nx;
        if (b0 < nx) { // a real band: rows [b0, b1) arrive in buf[0]
            CKS(cudaMemcpy2DAsync(h->buf[0] + b0 * h->pitch, h->pitch * sizeof(float), grid + b0 * row_bytes, row_bytes, row_bytes,
                                  (size_t)(b1 - b0), cudaMemcpyHostToDevice, in_s));
            CKS(cudaEventRecord(ev_in, in_s));
            CKS(cudaStreamWaitEvent(h->stream, ev_in, 0));
            // the ring rows never change: buf[1] needs its own copy of them (C:197-198 uploads the field twice)
            if (b0 == 0) CKS(cudaMemcpyAsync(h->buf[1], h->buf[0], row_bytes, cudaMemcpyDeviceToDevice, h->stream));
            if (b1 == nx)
                CKS(cudaMemcpyAsync(h->buf[1] + (size_t)(nx - 1) * h->pitch, h->buf[0] + (size_t)(nx - 1) * h->pitch, row_bytes,
                                    cudaMemcpyDeviceToDevice, h->stream));
        }
        for (int p = 1; p <= P && !rc; p++) {
            long long lo = b0 - (long long)p * s, hi = b0 + B - (long long)p * s;
            if (lo < 1) lo = 1;
            if (hi > nx - 1) hi = nx - 1;
            if (lo >= hi) continue;
            h->upd_lo = (int)lo;
            h->upd_hi = (int)hi;
            const float* in = h->buf[(p - 1) & 1];
            float* out = h->buf[p & 1];
            if (p <= nfused) CKS(launch_tb(h, T, in, out, nullptr, nullptr, 0));
            else CKS(launch_step(h, false, in, out, nullptr, nullptr, 0));
        }
        long long done = b0 + B - S; // rows below this are final now
        if (done > nx - 1) done = nx - 1;
        if (done > sent) {
            CKS(cudaEventRecord(ev_done, h->stream));
            CKS(cudaStreamWaitEvent(out_s, ev_done, 0));
            CKS(cudaMemcpy2DAsync(grid + sent * row_bytes, row_bytes, fin + sent * h->pitch, h->pitch * sizeof(float), row_bytes,
                                  (size_t)(done - sent), cudaMemcpyDeviceToHost, out_s));
            sent = done;
        }
    }
#undef CKS
    h->upd_lo = upd_lo0;
    h->upd_hi = upd_hi0;
    if (!rc && cudaEventRecord(h->ev1, h->stream) != cudaSuccess) rc = fail(HEAT_E_CUDA, "cudaEventRecord failed");
    // always drain the three streams before they go away (also after an error)
    cudaError_t e1 = in_s ? cudaStreamSynchronize(in_s) : cudaSuccess;
    cudaError_t e2 = cudaStreamSynchronize(h->stream);
    cudaError_t e3 = out_s ? cudaStreamSynchronize(out_s) : cudaSuccess;
    if (!rc && (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess))
        rc = fail(HEAT_E_CUDA, "heat_solve (streamed): %s", cudaGetErrorString(e1 != cudaSuccess ? e1 : e2 != cudaSuccess ? e2 : e3));
    if (!rc) {
        cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
        h->cur = P & 1;
        h->it += updates;
        h->sweeps += P;
        h->loaded = true;
    }
    if (ev_in) cudaEventDestroy(ev_in);
    if (ev_done) cudaEventDestroy(ev_done);
    if (in_s) cudaStreamDestroy(in_s);
    if (out_s) cudaStreamDestroy(out_s);
    cudaGetLastError();
    return rc;
}

static int solve_host(const heat_params* p, void* grid, int es, int* updates_done, int* converged);
extern "C" int heat_solve(const heat_params* p, float* grid, int* updates_done, int* converged)
{
    return solve_host(p, grid, 1, updates_done, converged);
}
extern "C" int heat_solve_f64(const heat_params* p, double* grid, int* updates_done, int* converged)
{
    return solve_host(p, grid, 2, updates_done, converged);
}

static int solve_host(const heat_params* p, void* grid, int es, int* updates_done, int* converged)
{
    if (!grid) return fail(HEAT_E_INVALID, "heat_solve: null grid");
    if (p && (p->grid_type == HEAT_GRID_F64) != (es == 2))
        return fail(HEAT_E_INVALID, "heat_solve: params.grid_type does not match the entry point (heat_solve = float, heat_solve_f64 = double)");
    if (p && (p->nranks != 1 || p->nranks_y != 1))
        return fail(HEAT_E_INVALID, "heat_solve: one strip only (use the handle API for strips)");
    const bool timing = env_int("HEAT_TIMING", 0) != 0;
    auto now = [] {
        struct timespec ts;
        clock_gettime(CLOCK_MONOTONIC, &ts);
        return ts.tv_sec + 1e-9 * ts.tv_nsec;
    };
    const double t0 = now();
    heat_handle* h = nullptr;
    int rc = create_handle(p, &h, true);
    if (rc) return rc;
    const double t1 = now();
    const long long total = (long long)p->steps + (p->step_convention == HEAT_STEPS_MPI ? 1 : 0);
    double t2, t3, t4;
    const bool streamed = solve_can_stream(h, total);
    if (streamed) {
        rc = solve_streamed(h, grid, total);
        if (!rc) {
            if (updates_done) *updates_done = (int)total;
            if (converged) *converged = 0;
        }
        t2 = t1, t3 = t4 = now();
    } else {
        rc = upload_from(h, grid, 0, cudaMemcpyHostToDevice);
        t2 = now();
        if (!rc) rc = heat_run(h, p->steps, updates_done, converged);
        t3 = now();
        if (!rc) rc = download_to(h, grid);
        t4 = now();
    }
    const float loop_ms = h->last_ms;
    char keep[sizeof g_err];
    memcpy(keep, g_err, sizeof keep); // heat_destroy must not hide the first error
    heat_destroy(h);
    if (rc) memcpy(g_err, keep, sizeof keep);
    if (timing)
        fprintf(stderr, "heat_solve%s: create %.3f ms, upload %.3f ms, run %.3f ms (device loop %.3f ms), download %.3f ms, destroy %.3f ms\n",
                streamed ? " (streamed: upload, loop and download overlap inside `run`)" : "", (t1 - t0) * 1e3, (t2 - t1) * 1e3,
                (t3 - t2) * 1e3, loop_ms, (t4 - t3) * 1e3, (now() - t4) * 1e3);
    return rc;
}

// One-shot over the first `ngpus` devices of this process: G row strips, one per GPU, one host thread
// (what src/heat_main.c --gpus G does).  Replaces the whole of M:51-297 for a single box.
extern "C" int heat_solve_gpus(const heat_params* p, int ngpus, float* grid, int* updates_done, int* converged)
{
    if (!p || !grid || ngpus < 1 || ngpus > 64) return fail(HEAT_E_INVALID, "heat_solve_gpus: bad arguments");
    if (ngpus == 1) return heat_solve(p, grid, updates_done, converged);
    heat_handle* h[64] = {};
    heat_params q = *p;
    q.nranks = ngpus;
    q.rank_y = 0;
    q.nranks_y = 1;
    int rc = HEAT_OK;
    for (int r = 0; r < ngpus && !rc; r++) {
        q.rank = r;
        q.device = r;
        rc = heat_create(&q, &h[r]);
    }
    for (int r = 0; r < ngpus && !rc; r++) rc = heat_connect_local(h[r], r > 0 ? h[r - 1] : nullptr, r < ngpus - 1 ? h[r + 1] : nullptr);
    for (int r = 0; r < ngpus && !rc; r++) rc = heat_upload(h[r], grid);
    if (!rc) rc = heat_group_run(h, ngpus, p->steps, updates_done, converged);
    for (int r = 0; r < ngpus && !rc; r++) rc = heat_download(h[r], grid);
    char keep[sizeof g_err];
    memcpy(keep, g_err, sizeof keep); // heat_destroy must not hide the first error
    for (int r = 0; r < ngpus; r++) heat_destroy(h[r]);
    if (rc) memcpy(g_err, keep, sizeof keep);
    return rc;
}

// ------------------------------------------------------------------ stateless updates (C:140-163, C:42-138)
template <bool RESID>
static int step_device(const void* d_in, void* d_out, int nx, int ny, float cx, float cy, int arith, unsigned* resid, void* stream)
{
    if (!d_in || !d_out || nx < 1 || ny < 1 || arith < 0 || arith > 2 || (RESID && !resid))
        return fail(HEAT_E_INVALID, "heat_step_device: bad arguments");
    if (nx < 3 || ny < 3) return HEAT_OK;
    int dev = 0, sms = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)); // a cached driver attribute: no device round trip
    dim3 grid((unsigned)((ny - 2 + 255) / 256), (unsigned)min(nx - 2, sms * 8));
    const float* in = (const float*)d_in;
    float* out = (float*)d_out;
    cudaStream_t st = (cudaStream_t)stream;
    switch (arith) {
    case 0: heat_step_generic<0, RESID><<<grid, 256, 0, st>>>(in, out, nx, ny, ny, ny, cx, cy, resid, nullptr, nullptr); break;
    case 1: heat_step_generic<1, RESID><<<grid, 256, 0, st>>>(in, out, nx, ny, ny, ny, cx, cy, resid, nullptr, nullptr); break;
    default: heat_step_generic<2, RESID><<<grid, 256, 0, st>>>(in, out, nx, ny, ny, ny, cx, cy, resid, nullptr, nullptr); break;
    }
    CK(cudaGetLastError());
    return HEAT_OK;
}

extern "C" int heat_step_device(const void* d_in, void* d_out, int nx, int ny, float cx, float cy, int arith, void* stream)
{
    return step_device<false>(d_in, d_out, nx, ny, cx, cy, arith, nullptr, stream);
}

extern "C" int heat_step_residual_device(const void* d_in, void* d_out, int nx, int ny, float cx, float cy, int arith,
                                         void* d_residual_bits, void* stream)
{
    return step_device<true>(d_in, d_out, nx, ny, cx, cy, arith, (unsigned*)d_residual_bits, stream);
}

// ------------------------------------------------------------------ neighbours
static int ptr_offset(const void* p, int64_t* off)
{
    // small cudaMalloc blocks are sub-allocated; an IPC handle exports the whole underlying
    // allocation, so the importer needs the offset.  The driver entry point is fetched at run
    // time so that the library has no link-time dependency on libcuda.so.1.
    typedef CUresult (*range_fn)(CUdeviceptr*, size_t*, CUdeviceptr);
    static range_fn get_range = nullptr;
    if (!get_range) {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        CK(cudaGetDriverEntryPoint("cuMemGetAddressRange", &f, cudaEnableDefault, &q));
        if (!f || q != cudaDriverEntryPointSuccess) return fail(HEAT_E_CUDA, "cuMemGetAddressRange not available");
        get_range = (range_fn)f;
    }
    CUdeviceptr base = 0;
    size_t size = 0;
    CUresult r = get_range(&base, &size, (CUdeviceptr)p);
    if (r != CUDA_SUCCESS) return fail(HEAT_E_CUDA, "cuMemGetAddressRange failed (%d)", (int)r);
    *off = (int64_t)((CUdeviceptr)p - base);
    return HEAT_OK;
}

extern "C" int heat_ipc_export(heat_handle* h, void* blob, size_t blob_bytes)
{
    if (!h || !blob || blob_bytes < sizeof(ipc_blob)) return fail(HEAT_E_INVALID, "heat_ipc_export: bad arguments");
    CK(cudaSetDevice(h->p.device));
    ipc_blob b;
    memset(&b, 0, sizeof b);
    memcpy(b.magic, "HEATIPC2", 8);
    void* ptrs[3] = {h->buf[0], h->buf[1], h->flags};
    for (int k = 0; k < 3; k++) {
        CK(cudaIpcGetMemHandle(&b.mem[k], ptrs[k]));
        int rc = ptr_offset(ptrs[k], &b.offset[k]);
        if (rc) return rc;
    }
    b.nloc = h->nloc;
    b.rank = h->p.rank;
    b.nranks = h->p.nranks;
    b.nyloc = h->nyloc;
    b.rank_y = h->p.rank_y;
    b.nranks_y = h->p.nranks_y;
    b.pitch = h->pitch;
    memset(blob, 0, blob_bytes);
    memcpy(blob, &b, sizeof b);
    return HEAT_OK;
}

// side: 0 north, 1 south, 2 west, 3 east.  Which block must sit there, and which of ITS flag slots is mine.
static void neighbour_of(const heat_handle* h, int side, int* rank, int* rank_y, int* their_slot)
{
    *rank = h->p.rank + (side == 0 ? -1 : side == 1 ? 1 : 0);
    *rank_y = h->p.rank_y + (side == 2 ? -1 : side == 3 ? 1 : 0);
    *their_slot = side ^ 1; // I am their south if they are my north, their east if they are my west
}

static bool has_side(const heat_handle* h, int side)
{
    return side == 0 ? h->p.rank > 0 : side == 1 ? h->p.rank < h->p.nranks - 1 : side == 2 ? h->p.rank_y > 0 : h->p.rank_y < h->p.nranks_y - 1;
}

static int open_side(heat_handle* h, int side, const void* blob)
{
    ipc_blob b;
    memcpy(&b, blob, sizeof b);
    if (memcmp(b.magic, "HEATIPC2", 8) != 0) return fail(HEAT_E_INVALID, "heat_ipc_connect: not an export blob");
    int want, want_y, slot;
    neighbour_of(h, side, &want, &want_y, &slot);
    // row neighbours share my columns (same pitch), column neighbours share my rows (same local row count)
    const bool shape_ok = side < 2 ? (b.pitch == h->pitch && b.nyloc == h->nyloc) : (b.nloc == h->nloc);
    if (b.rank != want || b.rank_y != want_y || b.nranks != h->p.nranks || b.nranks_y != h->p.nranks_y || !shape_ok)
        return fail(HEAT_E_PEER, "heat_ipc_connect: blob is block (%d,%d) of %dx%d, expected (%d,%d) of %dx%d with a matching shape",
                    b.rank, b.rank_y, b.nranks, b.nranks_y, want, want_y, h->p.nranks, h->p.nranks_y);
    void* base[3];
    for (int k = 0; k < 3; k++) {
        cudaError_t e = cudaIpcOpenMemHandle(&base[k], b.mem[k], cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return fail(HEAT_E_PEER, "cudaIpcOpenMemHandle failed: %s", cudaGetErrorString(e));
        h->peer[side].ipc_base[k] = base[k];
    }
    h->peer[side].buf[0] = (float*)((char*)base[0] + b.offset[0]);
    h->peer[side].buf[1] = (float*)((char*)base[1] + b.offset[1]);
    unsigned long long* f = (unsigned long long*)((char*)base[2] + b.offset[2]);
    h->peer[side].flag = &f[slot];
    h->peer[side].gcol = (float*)((char*)f + GCOL_OFFSET);
    h->peer[side].nloc = b.nloc;
    h->peer[side].nyloc = b.nyloc;
    h->peer[side].pitch = b.pitch;
    return HEAT_OK;
}

extern "C" int heat_ipc_connect2d(heat_handle* h, const void* north_blob, const void* south_blob, const void* west_blob,
                                  const void* east_blob)
{
    if (!h) return fail(HEAT_E_INVALID, "heat_ipc_connect: null handle");
    const void* blobs[4] = {north_blob, south_blob, west_blob, east_blob};
    for (int side = 0; side < 4; side++)
        if (has_side(h, side) != (blobs[side] != nullptr))
            return fail(HEAT_E_PEER, "heat_ipc_connect: block (%d,%d) of %dx%d %s a neighbour on side %d", h->p.rank, h->p.rank_y,
                        h->p.nranks, h->p.nranks_y, has_side(h, side) ? "needs" : "has no", side);
    CK(cudaSetDevice(h->p.device));
    for (int side = 0; side < 4; side++)
        if (blobs[side]) {
            int rc = open_side(h, side, blobs[side]);
            if (rc) return rc;
        }
    h->connected = true;
    return HEAT_OK;
}

extern "C" int heat_ipc_connect(heat_handle* h, const void* north_blob, const void* south_blob)
{
    return heat_ipc_connect2d(h, north_blob, south_blob, nullptr, nullptr);
}

extern "C" int heat_connect_local2d(heat_handle* h, heat_handle* north, heat_handle* south, heat_handle* west, heat_handle* east)
{
    if (!h) return fail(HEAT_E_INVALID, "heat_connect_local: null handle");
    heat_handle* nb[4] = {north, south, west, east};
    for (int side = 0; side < 4; side++)
        if (has_side(h, side) != (nb[side] != nullptr))
            return fail(HEAT_E_PEER, "heat_connect_local: block (%d,%d) of %dx%d %s a neighbour on side %d", h->p.rank, h->p.rank_y,
                        h->p.nranks, h->p.nranks_y, has_side(h, side) ? "needs" : "has no", side);
    CK(cudaSetDevice(h->p.device));
    for (int side = 0; side < 4; side++) {
        heat_handle* o = nb[side];
        if (!o) continue;
        int want, want_y, slot;
        neighbour_of(h, side, &want, &want_y, &slot);
        const bool shape_ok = side < 2 ? (o->pitch == h->pitch && o->nyloc == h->nyloc) : (o->nloc == h->nloc);
        if (o->p.rank != want || o->p.rank_y != want_y || o->p.nranks != h->p.nranks || o->p.nranks_y != h->p.nranks_y || !shape_ok)
            return fail(HEAT_E_PEER, "heat_connect_local: neighbour on side %d is block (%d,%d), not adjacent", side, o->p.rank,
                        o->p.rank_y);
        if (o->p.device != h->p.device) {
            int can = 0;
            CK(cudaDeviceCanAccessPeer(&can, h->p.device, o->p.device));
            if (!can) return fail(HEAT_E_PEER, "device %d cannot access device %d", h->p.device, o->p.device);
            cudaError_t e = cudaDeviceEnablePeerAccess(o->p.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                return fail(HEAT_E_PEER, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
            cudaGetLastError();
        }
        h->peer[side].buf[0] = o->buf[0];
        h->peer[side].buf[1] = o->buf[1];
        h->peer[side].flag = &o->flags[slot];
        h->peer[side].gcol = o->gcol;
        h->peer[side].nloc = o->nloc;
        h->peer[side].nyloc = o->nyloc;
        h->peer[side].pitch = o->pitch;
        h->local_nb[side] = o;
    }
    h->connected = true;
    return HEAT_OK;
}

extern "C" int heat_connect_local(heat_handle* h, heat_handle* north, heat_handle* south)
{
    return heat_connect_local2d(h, north, south, nullptr, nullptr);
}

// ------------------------------------------------------------------ update + residual without any allocation
// heat_step_residual_noalloc() is for callers that must not allocate: a host application that
// handed all of the GPU's memory to its own pool at start-up, or a solver loop with a hard
// latency budget that cannot afford the device-wide synchronisation a cudaMalloc/cudaFree
// implies.  Such a caller can still pass its input field, its output field and a 4-byte result
// slot, but it has nowhere to get the reduction's working memory from.  Here that working
// memory is a fixed array that is part of the library image, so the call allocates nothing.
//
// What to know before using it:
//   * The array belongs to the GPU, not to the call.  Every GPU that loads this library keeps
//     HEAT_NOALLOC_SCRATCH_BYTES of it until the process exits, and every call on that GPU uses
//     the same one.  Two calls that overlap on one GPU (two streams, two host threads) would
//     trample each other's partial results: run them one after another.
//   * The first call on a given GPU looks up the array's device address and remembers it.
//     Make that first call outside any stream capture.
//   * The array keeps whatever the previous call left in it.  The one word that has to start
//     at zero, the arrival counter, is cleared on the stream at the start of every call.
//   * Row length is limited to HEAT_NOALLOC_MAX_NY (one slot per thread block,
//     HEAT_NOALLOC_MAX_BLOCKS slots); longer rows are refused with HEAT_E_INVALID.
namespace noalloc {

constexpr int kThreads = 256;
constexpr int kSlots = HEAT_NOALLOC_MAX_BLOCKS; // one partial maximum per thread block
constexpr int kCounter = kSlots;                // the arrival counter sits behind the slots
constexpr int kWords = kSlots + 64;
static_assert(kWords * sizeof(unsigned) == HEAT_NOALLOC_SCRATCH_BYTES, "include/heat.h and the array disagree");
static_assert((long long)kSlots * kThreads + 2 == HEAT_NOALLOC_MAX_NY, "include/heat.h and the launch shape disagree");

// The working memory.  In host code this name is only a stand-in for the symbol, NOT a device
// address; host code gets the real address once per device from cudaGetSymbolAddress() below.
static __device__ __align__(256) unsigned scratch_words[kWords];

constexpr int kMaxDevices = 64;
static std::mutex lock;
static unsigned* scratch_addr[kMaxDevices]; // device address of the array, by device ordinal

// The update of heat_step_generic; each block leaves its maximum in its slot and the block
// that arrives last folds the slots into *d_residual.
template <int ARITH>
__global__ void __launch_bounds__(kThreads)
step_partials(const float* __restrict__ in, float* __restrict__ out, int nx, int ny, float cx, float cy, unsigned* scratch,
              float* d_residual)
{
    const int j = blockIdx.x * kThreads + threadIdx.x + 1;
    float m = 0.0f;
    if (j < ny - 1) {
        for (int i = blockIdx.y + 1; i < nx - 1; i += gridDim.y) {
            const float* c = in + (int64_t)i * ny + j;
            const float cv = c[0];
            const float o = upd<ARITH>(cv, c[-ny], c[ny], c[1], c[-1], cx, cy);
            out[(int64_t)i * ny + j] = o;
            m = fmaxf(m, fabsf(__fsub_rn(cv, o)));
        }
    }
    __shared__ unsigned warp_max[kThreads / 32];
    __shared__ bool last_block;
    const unsigned bits = __reduce_max_sync(0xffffffffu, __float_as_uint(m));
    if ((threadIdx.x & 31) == 0) warp_max[threadIdx.x >> 5] = bits;
    __syncthreads();
    const unsigned nblocks = gridDim.x * gridDim.y;
    if (threadIdx.x == 0) {
        unsigned v = 0;
        for (int w = 0; w < kThreads / 32; w++) v = max(v, warp_max[w]);
        scratch[blockIdx.y * gridDim.x + blockIdx.x] = v;
        __threadfence(); // the slot is visible before the ticket is taken
        last_block = (atomicAdd(&scratch[kCounter], 1u) == nblocks - 1);
    }
    __syncthreads();
    if (last_block) {
        unsigned v = 0;
        for (unsigned k = threadIdx.x; k < nblocks; k += kThreads) v = max(v, __ldcg(&scratch[k]));
        v = __reduce_max_sync(0xffffffffu, v);
        if ((threadIdx.x & 31) == 0) warp_max[threadIdx.x >> 5] = v;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned r = 0;
            for (int w = 0; w < kThreads / 32; w++) r = max(r, warp_max[w]);
            *d_residual = __uint_as_float(r);
            scratch[kCounter] = 0; // leave the counter clean for the next call as well
        }
    }
}

// Device address of the array on the CURRENT device; looked up on first use.
static int scratch_for_current_device(unsigned** out)
{
    int dev = -1;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(HEAT_E_INVALID, "heat_step_residual_noalloc: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> guard(lock);
    if (!scratch_addr[dev]) {
        unsigned* ptr = nullptr;
        CK(cudaGetSymbolAddress((void**)&ptr, scratch_words));
        if (!ptr) return fail(HEAT_E_CUDA, "heat_step_residual_noalloc: cudaGetSymbolAddress returned no address");
        scratch_addr[dev] = ptr;
    }
    *out = scratch_addr[dev];
    return HEAT_OK;
}

} // namespace noalloc

extern "C" int heat_step_residual_noalloc(const void* d_in, void* d_out, int nx, int ny, float cx, float cy, int arith,
                                          void* d_residual, void* stream)
{
    using namespace noalloc;
    if (!d_in || !d_out || !d_residual || nx < 1 || ny < 1 || arith < HEAT_F32_FMA || arith > HEAT_MPI_F64MIX)
        return fail(HEAT_E_INVALID, "heat_step_residual_noalloc: bad arguments");
    if (ny > HEAT_NOALLOC_MAX_NY)
        return fail(HEAT_E_INVALID, "heat_step_residual_noalloc: ny=%d is beyond the fixed working memory (HEAT_NOALLOC_MAX_NY=%d)",
                    ny, HEAT_NOALLOC_MAX_NY);
    unsigned* scratch = nullptr;
    int rc = scratch_for_current_device(&scratch);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    float* resid = (float*)d_residual;
    if (nx < 3 || ny < 3) { // no interior: nothing moves, the residual is zero
        CK(cudaMemsetAsync(resid, 0, sizeof(float), st));
        return HEAT_OK;
    }
    const int col_blocks = (ny - 2 + kThreads - 1) / kThreads;
    int row_blocks = kSlots / col_blocks;
    if (row_blocks > nx - 2) row_blocks = nx - 2;
    if (row_blocks > 65535) row_blocks = 65535;
    // the arrival counter must be zero when the kernel starts, whatever an earlier call left behind
    CK(cudaMemsetAsync(scratch + kCounter, 0, sizeof(unsigned), st));
    const dim3 grid((unsigned)col_blocks, (unsigned)row_blocks);
    const float* in = (const float*)d_in;
    float* out = (float*)d_out;
    switch (arith) {
    case HEAT_F32_FMA: step_partials<HEAT_ARITH_F32_FMA><<<grid, kThreads, 0, st>>>(in, out, nx, ny, cx, cy, scratch, resid); break;
    case HEAT_F32_NOFMA: step_partials<HEAT_ARITH_F32_NOFMA><<<grid, kThreads, 0, st>>>(in, out, nx, ny, cx, cy, scratch, resid); break;
    default: step_partials<HEAT_ARITH_F64MIX><<<grid, kThreads, 0, st>>>(in, out, nx, ny, cx, cy, scratch, resid); break;
    }
    CK(cudaGetLastError());
    return HEAT_OK;
}

// ------------------------------------------------------------------ debug build only
#ifdef HEAT_BOUNDS_CHECK
// number of out-of-window accesses the update kernels attempted on the current device so far
extern "C" int heat_debug_violations(unsigned* count)
{
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpyFromSymbol(count, g_violations, sizeof(unsigned)));
    return HEAT_OK;
}
#endif
