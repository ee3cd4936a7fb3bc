// heat_kernels.cuh -- sm_100a kernels of the solver loop.
//
// Reference regions replaced (M = mpi/mpi_heat_improved_persistent_stat.c,
// C = cuda/cuda_heat.cu; nothing here is derived from their code):
//   K0 heat_inidat_kernel     inidat                    M:315-321 / C:274-280
//   K1 heat_step_vec<..,0>    stencil update            M:162-234 / C:140-163
//   K2 heat_step_vec<..,1>    update + convergence      C:42-138 + C:32-40, M:236-261
//   K1g heat_step_generic     same update, any pitch/alignment (small or odd grids)
//   K3 (in K1/K2)             first/last owned row also stored into the neighbour
//                             strip's ghost row (replaces MPI halo requests M:130-155)
//   heat_signal/heat_wait     step flags between strips (replaces Startall/Waitall M:160-177,263)
//
// All three arithmetic modes are written with explicit rounding intrinsics, so the
// result does not depend on -fmad.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define HEAT_ARITH_F32_FMA 0
#define HEAT_ARITH_F32_NOFMA 1
#define HEAT_ARITH_F64MIX 2

#include <cuda/ptx>

namespace heatk {

// ------------------------------------------------------------------ arithmetic
// new = u + cx*((S+N) - 2u) + cy*((E+W) - 2u), evaluation order of M:168-174 / C:155-161.
template <int ARITH>
__device__ __forceinline__ float upd(float c, float n, float s, float e, float w, float cx, float cy)
{
    if constexpr (ARITH == HEAT_ARITH_F32_FMA) {
        // what nvcc makes of C:155-161 by default: 5 FADD + 2 FFMA (2u is exact)
        float t1 = __fsub_rn(__fadd_rn(s, n), __fadd_rn(c, c));
        float t2 = __fsub_rn(__fadd_rn(e, w), __fadd_rn(c, c));
        return __fmaf_rn(cy, t2, __fmaf_rn(cx, t1, c));
    } else if constexpr (ARITH == HEAT_ARITH_F32_NOFMA) {
        // C:155-161 with -fmad=false: every product and sum rounded to float
        float t1 = __fsub_rn(__fadd_rn(s, n), __fadd_rn(c, c));
        float t2 = __fsub_rn(__fadd_rn(e, w), __fadd_rn(c, c));
        return __fadd_rn(__fadd_rn(c, __fmul_rn(cx, t1)), __fmul_rn(cy, t2));
    } else {
        // M:168-174: float neighbour sums, then double (literal 2.0, promoted cx/cy), one final rounding.
        // The four float<->double conversions per point run on the XU pipe (16 lanes/clk/SM) and bound
        // this mode (ncu: XU 72 %, FP64 31 %).  Widening the centre value with integer operations
        // instead was tried and was slower (630 vs 730-760 GLUP/s at 8192^2).
        double cd = (double)c;
        double c2 = __dadd_rn(cd, cd);
        double t1 = __dsub_rn((double)__fadd_rn(s, n), c2);
        double t2 = __dsub_rn((double)__fadd_rn(e, w), c2);
        double r = __dadd_rn(__dadd_rn(cd, __dmul_rn((double)cx, t1)), __dmul_rn((double)cy, t2));
        return __double2float_rn(r);
    }
}

// ---- MPI_F64MIX, four points at a time -------------------------------------------------------------
// The scalar form above spends four conversions per point on the XU pipe (16 lanes/clk/SM): three
// float->double widenings and the final narrowing; that pipe, not HBM, bounds the mode (ncu: XU 72 %).
// Widening a NORMAL float is exact bit surgery -- sign kept, exponent re-biased by 896, mantissa
// shifted up 29 bits -- which the integer pipes do at four times the XU rate.  Zero, denormals, inf and
// nan do not fit that rule: ONE warp-uniform vote per float4 row decides whether every value to be
// widened is normal; if any lane holds a special value the whole warp takes the conversion
// instructions for that row.  HEAT_F64_WIDEN_INT = how many of the three widenings per point go the
// integer way when the vote allows it (0 = none: the scalar form; 1 = the centre value; 2 = centre and
// the N+S sum; 3 = all).  The final double->float rounding always stays a conversion instruction.
#ifndef HEAT_F64_WIDEN_INT
#define HEAT_F64_WIDEN_INT 1
#endif
// On the same fast path 2u can be had without the FP64 pipe as well: doubling a normal double is +1 on its
// exponent field (the widened float's exponent is at most 1150, far from overflow).  HEAT_F64_INT_DOUBLE.
#ifndef HEAT_F64_INT_DOUBLE
#define HEAT_F64_INT_DOUBLE 0
#endif

__device__ __forceinline__ double widen_normal(float v)
{
    const unsigned b = __float_as_uint(v);
    const unsigned hi = ((((b & 0x7fffffffu) >> 3) + 0x38000000u)) | (b & 0x80000000u);
    return __hiloint2double((int)hi, (int)(b << 29));
}
// (bits << 1) - 2^24 is below 0xfe000000 exactly for normal numbers (biased exponent 1..254)
__device__ __forceinline__ unsigned normal_key(float v) { return (__float_as_uint(v) << 1) - 0x01000000u; }

template <int NINT>
__device__ __forceinline__ float4 upd4_f64mix(const float4& c, const float4& n, const float4& s, float wl, float er, float cx,
                                              float cy)
{
    const float cv[4] = {c.x, c.y, c.z, c.w};
    const float sn[4] = {__fadd_rn(s.x, n.x), __fadd_rn(s.y, n.y), __fadd_rn(s.z, n.z), __fadd_rn(s.w, n.w)};
    const float ew[4] = {__fadd_rn(c.y, wl), __fadd_rn(c.z, c.x), __fadd_rn(c.w, c.y), __fadd_rn(er, c.z)};
    bool fast = false;
    if (NINT > 0) {
        unsigned k = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            k = max(k, normal_key(cv[i]));
            if (NINT > 1) k = max(k, normal_key(sn[i]));
            if (NINT > 2) k = max(k, normal_key(ew[i]));
        }
        fast = __all_sync(0xffffffffu, k < 0xfe000000u);
    }
    const double dcx = (double)cx, dcy = (double)cy;
    float o[4];
    if (fast) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const double cd = widen_normal(cv[i]);
#if HEAT_F64_INT_DOUBLE
            const double c2 = __hiloint2double(__double2hiint(cd) + 0x00100000, __double2loint(cd));
#else
            const double c2 = __dadd_rn(cd, cd);
#endif
            const double t1 = __dsub_rn(NINT > 1 ? widen_normal(sn[i]) : (double)sn[i], c2);
            const double t2 = __dsub_rn(NINT > 2 ? widen_normal(ew[i]) : (double)ew[i], c2);
            o[i] = __double2float_rn(__dadd_rn(__dadd_rn(cd, __dmul_rn(dcx, t1)), __dmul_rn(dcy, t2)));
        }
    } else {
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const double cd = (double)cv[i];
            const double c2 = __dadd_rn(cd, cd);
            const double t1 = __dsub_rn((double)sn[i], c2);
            const double t2 = __dsub_rn((double)ew[i], c2);
            o[i] = __double2float_rn(__dadd_rn(__dadd_rn(cd, __dmul_rn(dcx, t1)), __dmul_rn(dcy, t2)));
        }
    }
    return make_float4(o[0], o[1], o[2], o[3]);
}

// ---- float modes, four points at a time with the packed FP32 instructions of sm_100 ------------------
// FADD2 / FMUL2 / FFMA2 do two IEEE-rounded float operations per instruction on the register pairs a float4
// already sits in after LDG.128.  Same operations in the same order as upd<>, hence the same bits; only the
// four E+W sums stay scalar (their operands straddle the pairs): 16 floating-point instructions per float4
// instead of 28.  Measured (round 2, one B200): nothing for K1 (852 vs 852 GLUP/s at 8192^2, 848 vs 851 at
// 65536^2: the kernel waits for HBM, not for issue slots), but K4, which is FP32-issue-bound, gains: T=4 at
// 32768^2 2620 -> 2670 GLUP/s, 8192^2 2159 -> 2224 (F32_FMA).  So packed is the default in K4 (HEAT_PACKED_K4)
// and off in K1 (HEAT_PACKED_K1); F32_FMA only (see upd_quad).
#ifndef HEAT_PACKED_K1
#define HEAT_PACKED_K1 0
#endif
#ifndef HEAT_PACKED_K4
#define HEAT_PACKED_K4 1
#endif
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float a, float b)
{
    f32x2 r;
    asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ void upk2(f32x2 r, float& a, float& b) { asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(r)); }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

template <int ARITH>
__device__ __forceinline__ float4 upd4_packed(const float4& c, const float4& n, const float4& s, float wl, float er, float cx,
                                              float cy)
{
    const f32x2 c01 = pk2(c.x, c.y), c23 = pk2(c.z, c.w);
    const f32x2 cc01 = add2(c01, c01), cc23 = add2(c23, c23);
    const f32x2 a01 = sub2(add2(pk2(s.x, s.y), pk2(n.x, n.y)), cc01);
    const f32x2 a23 = sub2(add2(pk2(s.z, s.w), pk2(n.z, n.w)), cc23);
    // E + W: (c.y + wl, c.z + c.x, c.w + c.y, er + c.z)
    const f32x2 b01 = sub2(pk2(__fadd_rn(c.y, wl), __fadd_rn(c.z, c.x)), cc01);
    const f32x2 b23 = sub2(pk2(__fadd_rn(c.w, c.y), __fadd_rn(er, c.z)), cc23);
    const f32x2 cx2 = pk2(cx, cx), cy2 = pk2(cy, cy);
    f32x2 o01, o23;
    if constexpr (ARITH == HEAT_ARITH_F32_FMA) {
        o01 = fma2(cy2, b01, fma2(cx2, a01, c01));
        o23 = fma2(cy2, b23, fma2(cx2, a23, c23));
    } else {
        o01 = add2(add2(c01, mul2(cx2, a01)), mul2(cy2, b01));
        o23 = add2(add2(c23, mul2(cx2, a23)), mul2(cy2, b23));
    }
    float4 o;
    upk2(o01, o.x, o.y);
    upk2(o23, o.z, o.w);
    return o;
}

// One float4 of a row whose four columns are all interior, E/W neighbours across the float4 given (all lanes of
// the warp must call this together: the MPI_F64MIX form votes).
template <int ARITH, bool PACKED = (HEAT_PACKED_K1 != 0)>
__device__ __forceinline__ float4 upd_quad(const float4& c, const float4& n, const float4& s, float wl, float er, float cx, float cy)
{
    if constexpr (ARITH == HEAT_ARITH_F64MIX && HEAT_F64_WIDEN_INT > 0) {
        return upd4_f64mix<HEAT_F64_WIDEN_INT>(c, n, s, wl, er, cx, cy);
    } else if constexpr (ARITH == HEAT_ARITH_F32_FMA && PACKED) {
        // F32_FMA only: ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 (it never fuses the scalar .rn forms),
        // which would turn F32_NOFMA into F32_FMA -- caught by the parity tests, so F32_NOFMA stays scalar
        return upd4_packed<ARITH>(c, n, s, wl, er, cx, cy);
    } else {
        float4 o;
        o.x = upd<ARITH>(c.x, n.x, s.x, c.y, wl, cx, cy);
        o.y = upd<ARITH>(c.y, n.y, s.y, c.z, c.x, cx, cy);
        o.z = upd<ARITH>(c.z, n.z, s.z, c.w, c.y, cx, cy);
        o.w = upd<ARITH>(c.w, n.w, s.w, er, c.z, cx, cy);
        return o;
    }
}

#ifdef HEAT_BOUNDS_CHECK
// Debug build only (python -m parallel_heat_b200.build --debug -> libheat_b200_dbg.so): every
// access of the update kernels is checked against the buffer it must stay inside; a violation is
// counted and the access suppressed.  compute-sanitizer is closed on the GPU pool, this is its
// stand-in (tests/test_bounds_debug.py).  The host sets the windows before every launch.
struct mem_window {
    const float* lo[4]; // in, out, north peer rows, south peer rows
    const float* hi[4];
};
__device__ mem_window g_window;
__device__ unsigned g_violations;
__device__ __forceinline__ bool inside(const float* p, int n, int first, int last)
{
    for (int k = first; k <= last; k++)
        if (g_window.lo[k] && p >= g_window.lo[k] && p + n <= g_window.hi[k]) return true;
    atomicAdd(&g_violations, 1u);
    return false;
}
__device__ __forceinline__ float4 ld4(const float* p)
{
    return inside(p, 4, 0, 0) ? *reinterpret_cast<const float4*>(p) : make_float4(0.f, 0.f, 0.f, 0.f);
}
__device__ __forceinline__ float ld1(const float* p) { return inside(p, 1, 0, 0) ? *p : 0.0f; }
__device__ __forceinline__ void st4(float* p, float4 v)
{
    if (inside(p, 4, 1, 3)) *reinterpret_cast<float4*>(p) = v;
}
#else
// Cache policy of the field's loads and stores (build-time; 0 = plain).  Every byte is read once and written once
// per pass, so streaming hints (evict-first) are the candidates; measured in round 2, see DESIGN.md section 3.
#ifndef HEAT_LD_POLICY
#define HEAT_LD_POLICY 0 // 1 = ld.global.cs (evict first), 2 = ld.global.nc (read-only path)
#endif
#ifndef HEAT_ST_POLICY
#define HEAT_ST_POLICY 0 // 1 = st.global.cs (evict first), 2 = st.global.wt (write through)
#endif
__device__ __forceinline__ float4 ld4(const float* p)
{
#if HEAT_LD_POLICY == 1
    return __ldcs(reinterpret_cast<const float4*>(p));
#elif HEAT_LD_POLICY == 2
    return __ldg(reinterpret_cast<const float4*>(p));
#else
    return *reinterpret_cast<const float4*>(p);
#endif
}
__device__ __forceinline__ float ld1(const float* p) { return *p; }
__device__ __forceinline__ void st4(float* p, float4 v)
{
#if HEAT_ST_POLICY == 1
    __stcs(reinterpret_cast<float4*>(p), v);
#elif HEAT_ST_POLICY == 2
    __stwt(reinterpret_cast<float4*>(p), v);
#else
    *reinterpret_cast<float4*>(p) = v;
#endif
}
#endif

// ------------------------------------------------------------------ K0 inidat
// (float)(ix*(nx-ix-1)*iy*(ny-iy-1)) with the int32 wrap of the reference (M:320).
// Local array = global rows [row_lo, row_lo+nloc); writes both buffers (C:197-198).
__global__ void heat_inidat_kernel(float* __restrict__ a, float* __restrict__ b, int nx, int ny, int row_lo,
                                   int nloc, int64_t pitch, int col_lo, int nyloc)
{
    const int64_t total = (int64_t)nloc * pitch;
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < total; k += (int64_t)gridDim.x * blockDim.x) {
        const int lr = (int)(k / pitch);
        const int ly = (int)(k - (int64_t)lr * pitch);
        const int ix = row_lo + lr, iy = col_lo + ly;
        float v = 0.0f;
        if (ly < nyloc) {
            uint32_t p = (uint32_t)ix * (uint32_t)(nx - ix - 1) * (uint32_t)iy * (uint32_t)(ny - iy - 1);
            v = (float)(int32_t)p;
        }
        a[k] = v;
        b[k] = v;
    }
}

// ------------------------------------------------------------------ residual helpers
// non-negative floats order like their bit patterns, so max|d| is an integer max
__device__ __forceinline__ void block_max_to_global(float m, unsigned* gmax)
{
    unsigned bits = __reduce_max_sync(0xffffffffu, __float_as_uint(m));
    __shared__ unsigned smax[32];
    const int warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
    if ((threadIdx.x & 31) == 0) smax[warp] = bits;
    __syncthreads();
    if (warp == 0) {
        unsigned v = (threadIdx.x < nwarp) ? smax[threadIdx.x] : 0u;
        v = __reduce_max_sync(0xffffffffu, v);
        if (threadIdx.x == 0 && v != 0u) atomicMax(gmax, v);
    }
}

// ------------------------------------------------------------------ K1g generic
// One thread per point, lanes along iy (coalesced).  Any ny, any pitch, unaligned bases.
// Ring is not written.  Used for odd shapes and by the stateless entry points.
template <int ARITH, bool RESID>
__global__ void __launch_bounds__(256)
heat_step_generic(const float* __restrict__ in, float* __restrict__ out, int nrows, int ny, int64_t pitch_in,
                  int64_t pitch_out, float cx, float cy, unsigned* __restrict__ resid, float* peer_n, float* peer_s)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x + 1;
    float m = 0.0f;
    if (j < ny - 1) {
        for (int i = blockIdx.y + 1; i < nrows - 1; i += gridDim.y) {
            const float* c = in + (int64_t)i * pitch_in + j;
            const float cv = c[0];
            const float o = upd<ARITH>(cv, c[-pitch_in], c[pitch_in], c[1], c[-1], cx, cy);
            out[(int64_t)i * pitch_out + j] = o;
            if (i == 1 && peer_n) peer_n[j] = o;
            if (i == nrows - 2 && peer_s) peer_s[j] = o;
            if (RESID) m = fmaxf(m, fabsf(__fsub_rn(cv, o)));
        }
    }
    if (RESID) block_max_to_global(m, resid);
}

// ------------------------------------------------------------------ strip ordering inside the update kernels
// One process per GPU: neighbours are ordered by step flags in each other's memory (see
// heat_signal/heat_wait below for the stand-alone versions).  Fused into the update kernels so
// that a pass costs one launch: only the CTAs of the tiles that touch a neighbour's data
// (they read ghost rows and store the neighbour's ghost rows) wait for the neighbour's flag,
// interior tiles start at once; when the last of those edge CTAs is done, it raises the
// neighbour's flag.  All pointers null = single strip / event-ordered strips: no code runs.
// Sides: 0 = north (row above), 1 = south, 2 = west (column to the left), 3 = east.  Row strips use
// 0 and 1 only; 2-D blocks (the reference's own topology, M:51-75) use all four.
struct strip_sync {
    const unsigned long long* wait[4]; // my slots, each written by the neighbour on that side
    unsigned long long wait_value;     // epoch the neighbours must have completed
    unsigned long long* sig[4];        // the neighbours' slots for me
    unsigned long long sig_value;
    unsigned* counters;                // [4] edge CTAs done, per side
    unsigned count[4];                 // how many CTAs touch the neighbour on each side
    int* status;                       // raised when a wait times out
    unsigned long long timeout_ns;
    unsigned long long* prof;          // diagnostics (HEAT_PROFILE_WAIT=1), else null: [0..3] ns the edge CTAs spent
                                       // waiting per side, [4..7] number of waits per side
};

// Column neighbours of a 2-D block.  Ghost columns do not travel into the neighbour's field: a store per row
// into a column of a peer's field touches a new page of peer memory every row or two (the row pitch is tens of
// kilobytes), and the address translation of those stores, not their bytes, is what limits the pass (two column
// neighbours at 65536^2: 65536 stores per pass, 1.07 ms instead of 0.63).  Instead every block keeps, per buffer
// and side, a DENSE array with one float per local row; a block stores its first / last owned column into the
// neighbour's dense array of the new buffer and reads its own dense arrays of the old buffer in place of the
// ghost column of its field (which is never written again after the upload).
struct col_peers {
    float* w;          // west neighbour's dense EAST-ghost array (new buffer), indexed by local row
    float* e;          // east neighbour's dense WEST-ghost array (new buffer)
    const float* gw_in; // my dense WEST-ghost array of the buffer I read (values of local column 0)
    const float* ge_in; // my dense EAST-ghost array of the buffer I read (values of local column ny-1)
    int edge_first;    // HEAT_COLS_FIRST = M > 0: column-edge CTAs are handed out before the rest of the block and
                       // each covers M tile rows (see cols_first); 0 = plain grid order
};

__device__ __forceinline__ void spin_until(const unsigned long long* f, unsigned long long value,
                                           unsigned long long timeout_ns, int* status)
{
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    unsigned spins = 0;
    for (;;) {
        unsigned long long v;
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
        if (v >= value) return;
        if ((++spins & 255u) == 0) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t - t0 > timeout_ns) {
                *status = 1;
                __threadfence_system();
                return;
            }
        }
        __nanosleep(32);
    }
}

// all threads of the CTA call these (block barrier inside); the edge flags are CTA-uniform
__device__ __forceinline__ void strip_wait(const strip_sync& sy, bool edge_n, bool edge_s, bool edge_w = false, bool edge_e = false)
{
    const bool e[4] = {edge_n && sy.wait[0], edge_s && sy.wait[1], edge_w && sy.wait[2], edge_e && sy.wait[3]};
    if (e[0] || e[1] || e[2] || e[3]) {
        if (threadIdx.x == 0) {
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (e[k]) {
                    unsigned long long t0 = 0;
                    if (sy.prof) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
                    spin_until(sy.wait[k], sy.wait_value, sy.timeout_ns, sy.status);
                    if (sy.prof) {
                        unsigned long long t1;
                        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                        atomicAdd(&sy.prof[k], t1 - t0);
                        atomicAdd(&sy.prof[4 + k], 1ull);
                    }
                }
        }
        __syncthreads();
    }
}

__device__ __forceinline__ void strip_signal(const strip_sync& sy, bool edge_n, bool edge_s, bool edge_w = false, bool edge_e = false)
{
    const bool e[4] = {edge_n && sy.sig[0], edge_s && sy.sig[1], edge_w && sy.sig[2], edge_e && sy.sig[3]};
    if (e[0] || e[1] || e[2] || e[3]) {
        __syncthreads(); // every thread's stores (local and peer) are issued
        if (threadIdx.x == 0) {
            // Row neighbours: a handful of edge CTAs, each fences at system scope before it counts itself.
            // Column neighbours: one edge CTA per tile row (thousands); they fence at GPU scope only.  On every
            // side the CTA that arrives last fences at system scope BEHIND its ticket (fence-atomic ... atomic-fence
            // is the PTX model's release/acquire pattern), then releases the flag.
            if (e[0] || e[1]) __threadfence_system();
            else __threadfence();
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (e[k] && atomicAdd(&sy.counters[k], 1u) == sy.count[k] - 1) {
                    // last arrival: the fence behind the (relaxed) ticket makes this an acquire of every earlier
                    // arrival's fence + ticket, so their peer stores are ordered before the flag for the neighbour
                    __threadfence_system();
                    sy.counters[k] = 0;
                    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(sy.sig[k]), "l"(sy.sig_value) : "memory");
                }
        }
    }
}

// Which row tile a CTA works on.  Plain sweeps go top-down or (reverse) bottom-up.  With strip
// flags the two edge tiles are handed to the first CTAs of the grid in either direction: their
// system-scope fence + flag store (a few microseconds behind an NVLink round trip) then overlaps
// with the rest of the pass instead of trailing it, and the neighbour's flag is long there when
// the next pass starts.
template <bool SYNC>
__device__ __forceinline__ int tile_of_block(int by, int ntiles, int reverse)
{
    if (SYNC && ntiles > 2) {
        if (by == 0) return 0;
        if (by == 1) return ntiles - 1;
        return reverse ? (ntiles - by) : (by - 1);
    }
    return reverse ? (ntiles - 1 - by) : by;
}

// 2-D blocks with step flags, optional order (HEAT_COLS_FIRST=1): every tile row has one CTA next to the west
// and one next to the east neighbour, so in grid order the last of them finishes with the pass and the
// neighbours' flags travel while the next pass already waits for them.  This remaps (bx, by) so that after the
// two row-edge tile rows (by = 0, 1: unchanged) the 2*(gy-2) column-edge CTAs of the other tile rows come next,
// then everything else row by row.  A permutation of the grid; `xe` is the east-edge CTA column.
// Returns false when the grid is too small for it (the host applies the same test, see cols_first_active).
__host__ __device__ __forceinline__ bool cols_first_active(int gx, int gy, int xe) { return gy > 2 && gx > 2 && xe > 0; }
__device__ __forceinline__ bool cols_first(int& bx, int& by, int gx, int gy, int xe)
{
    if (!cols_first_active(gx, gy, xe)) return false;
    int b = by * gx + bx - 2 * gx;
    if (b < 0) return true;
    const int rows = gy - 2;
    if (b < 2 * rows) {
        by = 2 + (b >> 1);
        bx = (b & 1) ? xe : 0;
        return true;
    }
    b -= 2 * rows;
    const int w = gx - 2;
    by = 2 + b / w;
    int k = 1 + b % w; // 1 .. gx-2, stepping over xe
    if (k >= xe) k++;
    bx = k;
    return true;
}

// ------------------------------------------------------------------ K1/K2 vector row streaming
// Each thread owns 4 consecutive columns (one float4) and walks down `rows_per_tile`
// rows keeping the N / C / S rows in registers: per row one LDG.128, one STG.128, two
// shuffles for the E/W neighbours inside the warp and one predicated 4-byte load on lanes
// 0 and 31 for the neighbours across the warp edge.  Requirements: pitch % 4 == 0, base
// 16-byte aligned, columns [ny, pitch) are padding.  Ring columns (0 and ny-1) are written
// back unchanged: both buffers always hold the same ring, so this equals "not written".
// Rows [row_lo, row_hi) are updated (rows row_lo-1 and row_hi must exist: ring or ghost).
//   grid.x = ceil(pitch/4 / THREADS), grid.y = number of row tiles, block = THREADS
//   reverse != 0 walks the tiles bottom-up (alternate sweeps re-use what is left in L2).
#ifndef HEAT_F64MIX_THREADS_PER_SM
#define HEAT_F64MIX_THREADS_PER_SM 1024 // resident threads per SM asked of the compiler for MPI_F64MIX (64 registers)
#endif
template <int ARITH, bool RESID, int THREADS, int UNROLL, bool SYNC, bool COLS = false>
// the plain single-strip update is held to 48 registers (40 warps per SM), the variants with the
// residual or the strip flags to 56 (36 warps), the double-precision mode to 64
__global__ void __launch_bounds__(THREADS, (ARITH == HEAT_ARITH_F64MIX ? HEAT_F64MIX_THREADS_PER_SM : (SYNC || RESID || COLS) ? 1152 : 1280) / THREADS)
heat_step_vec(const float* __restrict__ in, float* __restrict__ out, int row_lo, int row_hi, int ny, int64_t pitch,
              int rows_per_tile, float cx, float cy, unsigned* __restrict__ resid, float* peer_n, float* peer_s,
              int ghost, int reverse, strip_sync sy, col_peers cp = col_peers{})
{
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int npitch = (int)pitch;
    int bx = (int)blockIdx.x, by = (int)blockIdx.y;
    // the CTA column that holds local column ny-2, the one next to the east neighbour (see edge_e below)
    const int xe = ((ny - 2) >> 2) / THREADS;
    int span = 1; // row tiles this CTA covers
    if (COLS && SYNC && cp.edge_first && cols_first(bx, by, (int)gridDim.x, (int)gridDim.y, xe)) {
        // Every column-edge CTA ends with a fence behind its peer stores (an NVLink round trip during which it
        // holds its slot).  With edge_first = M > 1 the column-edge CTAs of M consecutive tile rows team up: the
        // first does the rows of all M, the others leave at once -- M times fewer fences, and since these CTAs
        // were handed out first the long ones run beside the rest of the pass, not after it.
        if (cp.edge_first > 1 && by >= 2 && (bx == 0 || bx == xe)) {
            const int e = by - 2;
            if (e % cp.edge_first) return;
            span = min(cp.edge_first, (int)gridDim.y - 2 - e);
        }
    }
    int col = (bx * THREADS + threadIdx.x) * 4;

    const int ntiles = gridDim.y;
    int tile = tile_of_block<SYNC>(by, ntiles, reverse);
    if (COLS && SYNC && span > 1 && reverse) tile -= span - 1; // bottom-up sweeps number the tiles downwards
    const int r0 = row_lo + tile * rows_per_tile;
    const int r1 = min(r0 + span * rows_per_tile, row_hi);
    // tiles that touch a neighbour strip's data: they read ghost rows / store its ghost rows
    const bool edge_n = r0 < row_lo + ghost, edge_s = r1 > row_hi - ghost;
    // 2-D blocks: per tile row one CTA holds the column next to the west neighbour (local column 1) and one the
    // column next to the east neighbour (local column ny-2: it reads the ghost column ny-1 and stores into the
    // neighbour).  The latter is NOT always the last CTA of the row: the pitch is rounded up to 32 floats, so
    // a trailing CTA of padding exists whenever ny-2 is the last column of a CTA (ny = 1 mod 4*THREADS).
    const bool edge_w = COLS && bx == 0, edge_e = COLS && bx == xe;
    if (SYNC) strip_wait(sy, edge_n, edge_s, edge_w, edge_e);

    float m = 0.0f;
    // a whole warp beyond the row has nothing to do (warp-uniform, so the shuffles stay converged)
    const bool live = col < npitch;
    if (col - lane * 4 < npitch) {
    if (!live) col = npitch - 4; // keep the lane in bounds for loads; it never stores

    // which of my 4 columns are ring / padding (copied through, never updated)
    const bool kx = (col == 0) || (col >= ny - 1);
    const bool ky = (col + 1 >= ny - 1);
    const bool kz = (col + 2 >= ny - 1);
    const bool kw = (col + 3 >= ny - 1);
    // neighbours across the warp edge: lane 0 needs col-1, lane 31 needs col+4
    const bool edge_lane = (lane == 0 && col > 0) || (lane == 31 && col + 4 < npitch);
    const int edge_off = (lane == 0) ? -1 : 4;

    const float* p = in + (int64_t)(r0 - 1) * pitch + col;
    float* q = out + (int64_t)r0 * pitch + col;
    float4 n = ld4(p);
    float4 c = ld4(p + pitch);
    float ec = edge_lane ? ld1(p + pitch + edge_off) : 0.0f;
    p += 2 * pitch; // points at row r+1 (the S row of the first output row)

    // Almost every warp takes this branch: no ring or padding column among its 128 columns, a full
    // tile, no neighbour strip's rows to store.  Same arithmetic, none of the special cases.
    // (with column neighbours the float4 that ends at the last owned column is special too)
    const bool plain_warp = __all_sync(FULL, live && !(kx || ky || kz || kw) && !(COLS && col + 3 == ny - 2));
    const bool plain_tile = (r1 - r0 == span * rows_per_tile) && (rows_per_tile % UNROLL == 0) && !(peer_n && edge_n) &&
                            !(peer_s && edge_s);
    if (plain_warp && plain_tile) {
        const bool el = (lane == 0) || (lane == 31);
        for (int r = r0; r < r1; r += UNROLL) {
            float4 s[UNROLL];
            float es[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; u++) {
                s[u] = ld4(p + (int64_t)u * pitch);
                es[u] = el ? ld1(p + (int64_t)u * pitch + edge_off) : 0.0f;
            }
#pragma unroll
            for (int u = 0; u < UNROLL; u++) {
                float wl = __shfl_up_sync(FULL, c.w, 1);
                float er = __shfl_down_sync(FULL, c.x, 1);
                if (lane == 0) wl = ec;
                if (lane == 31) er = ec;
                const float4 o = upd_quad<ARITH>(c, n, s[u], wl, er, cx, cy);
                st4(q, o);
                if (RESID) {
                    m = fmaxf(m, fabsf(__fsub_rn(c.x, o.x)));
                    m = fmaxf(m, fabsf(__fsub_rn(c.y, o.y)));
                    m = fmaxf(m, fabsf(__fsub_rn(c.z, o.z)));
                    m = fmaxf(m, fabsf(__fsub_rn(c.w, o.w)));
                }
                n = c;
                c = s[u];
                ec = es[u];
                q += pitch;
            }
            p += (int64_t)UNROLL * pitch;
        }
    } else
    for (int r = r0; r < r1; r += UNROLL) {
        float4 s[UNROLL];
        float es[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            // rows past the tile are clamped to its last S row: loaded but never used for a store
            const int64_t off = (int64_t)min(u, r1 - 1 - r) * pitch;
            s[u] = ld4(p + off);
            es[u] = edge_lane ? ld1(p + off + edge_off) : 0.0f;
        }
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            if (r + u < r1) {
                if (COLS) {
                    // the ghost columns of row r+u come from the dense arrays, not from the field
                    if (cp.gw_in && col == 0) c.x = cp.gw_in[r + u];
                    if (cp.ge_in) {
                        const float g = (col <= ny - 1 && ny - 1 < col + 4) || (lane == 31 && col + 4 == ny - 1) ? cp.ge_in[r + u] : 0.0f;
                        const int j = ny - 1 - col;
                        if (j == 0) c.x = g;
                        if (j == 1) c.y = g;
                        if (j == 2) c.z = g;
                        if (j == 3) c.w = g;
                        if (lane == 31 && j == 4) ec = g;
                    }
                }
                float wl = __shfl_up_sync(FULL, c.w, 1);
                float er = __shfl_down_sync(FULL, c.x, 1);
                if (lane == 0) wl = ec;
                if (lane == 31) er = ec;
                float4 o;
                o.x = upd<ARITH>(c.x, n.x, s[u].x, c.y, wl, cx, cy);
                o.y = upd<ARITH>(c.y, n.y, s[u].y, c.z, c.x, cx, cy);
                o.z = upd<ARITH>(c.z, n.z, s[u].z, c.w, c.y, cx, cy);
                o.w = upd<ARITH>(c.w, n.w, s[u].w, er, c.z, cx, cy);
                if (kx) o.x = c.x;
                if (ky) o.y = c.y;
                if (kz) o.z = c.z;
                if (kw) o.w = c.w;
                if (live) {
                    if (COLS && ((cp.w && col == 0) || (cp.e && col <= ny - 1 && ny - 1 < col + 4))) {
                        // a ghost column lives in this float4: the neighbour stores it, I must not
                        const bool gw = cp.w != nullptr, ge = cp.e != nullptr;
                        if (!(gw && col == 0) && !(ge && col == ny - 1)) q[0] = o.x;
                        if (!(ge && col + 1 == ny - 1)) q[1] = o.y;
                        if (!(ge && col + 2 == ny - 1)) q[2] = o.z;
                        if (!(ge && col + 3 == ny - 1)) q[3] = o.w;
                        // my first owned column (local column 1) is the west neighbour's ghost column
                        if (gw && col == 0) cp.w[r + u] = o.y;
                    } else {
                        st4(q, o);
                    }
                    // my last owned column (local column ny-2) is the east neighbour's ghost column
                    if (COLS && cp.e && col <= ny - 2 && ny - 2 < col + 4) {
                        const int j = ny - 2 - col;
                        cp.e[r + u] = j == 0 ? o.x : j == 1 ? o.y : j == 2 ? o.z : o.w;
                    }
                    // K3: the first / last `ghost` updated rows are the neighbours' ghost rows
                    if (peer_n && r + u < row_lo + ghost) st4(peer_n + (int64_t)(r + u - row_lo) * pitch + col, o);
                    if (peer_s && r + u >= row_hi - ghost) st4(peer_s + (int64_t)(r + u - (row_hi - ghost)) * pitch + col, o);
                }
                if (RESID) {
                    m = fmaxf(m, fabsf(__fsub_rn(c.x, o.x)));
                    m = fmaxf(m, fabsf(__fsub_rn(c.y, o.y)));
                    m = fmaxf(m, fabsf(__fsub_rn(c.z, o.z)));
                    m = fmaxf(m, fabsf(__fsub_rn(c.w, o.w)));
                }
                n = c;
                c = s[u];
                ec = es[u];
                q += pitch;
            }
        }
        p += (int64_t)UNROLL * pitch;
    }
    } // warp inside the row
    if (RESID) block_max_to_global(live ? m : 0.0f, resid);
    if (SYNC) strip_signal(sy, edge_n, edge_s, edge_w, edge_e);
}

// ------------------------------------------------------------------ K1b: the same update from a shared-memory tile
// The alternative north_star names: a tile with its halo rows is staged in shared memory by the bulk-copy
// engine (cp.async.bulk, one copy per row, completion on an mbarrier; SASS: UBLKCP), then every thread computes
// its float4 column of the tile from shared memory (E/W neighbours by shuffle, across the warp edge from the
// tile).  One tile per CTA, no pipelining inside the CTA: the other CTAs of the SM (ROWS+2 rows of 2080 bytes
// each) cover the copy latency.  Same bytes over HBM as K1; what changes is who issues the loads: one thread
// issues ROWS+2 row copies of 2 KB instead of 128 threads issuing 16-byte loads.  Measured (round 2, one B200,
// 20-update repetitions): 8192^2 850 vs 852 GLUP/s for K1, 16384^2 862 vs 854, 65536^2 865 vs 851 -- it wins
// where rows are long, so HEAT_K1_VARIANT=auto (the default) takes it for rows of 16384 floats and more.
// Everything K1 does besides is here too: residual (K2), neighbour strips' ghost rows and step flags (K3).
template <int ARITH, bool RESID, int ROWS, int THREADS, bool SYNC, bool PEERS>
__global__ void __launch_bounds__(THREADS)
heat_step_bulk(const float* __restrict__ in, float* __restrict__ out, int row_lo, int row_hi, int ny, int64_t pitch, float cx,
               float cy, unsigned* __restrict__ resid, float* peer_n, float* peer_s, int ghost, int reverse, strip_sync sy,
               int* __restrict__ status)
{
    namespace ptx = cuda::ptx;
    constexpr int W = THREADS * 4, HALO = 4, SW = W + 2 * HALO; // staged row: 4 columns of halo on either side
    __shared__ __align__(128) float tile[(ROWS + 2) * SW];
    __shared__ __align__(8) uint64_t bar;
    const unsigned FULL = 0xffffffffu;
    const int npitch = (int)pitch, tid = threadIdx.x, lane = tid & 31;
    const int ntiles = gridDim.y;
    const int t = tile_of_block<SYNC>((int)blockIdx.y, ntiles, reverse);
    const int r0 = row_lo + t * ROWS, r1 = min(r0 + ROWS, row_hi), nrows = r1 - r0;
    const bool edge_n = r0 < row_lo + ghost, edge_s = r1 > row_hi - ghost;
    const int col0 = blockIdx.x * W;
    // staged columns [c_lo, c_hi) of rows r0-1 .. r1; column c sits at tile[row * SW + (c - (col0 - HALO))]
    const int c_lo = max(col0 - HALO, 0), c_hi = min(col0 + W + HALO, npitch);
    const unsigned row_bytes = (unsigned)(c_hi - c_lo) * 4u; // a multiple of 16: every bound is a multiple of 4 floats
    if (tid == 0) {
        ptx::mbarrier_init(&bar, 1);
        ptx::fence_proxy_async(ptx::space_shared); // the barrier's initialisation must be visible to the copy engine
    }
    if (SYNC) strip_wait(sy, edge_n, edge_s); // the ghost rows this tile reads are the neighbour's previous pass
    __syncthreads();
    if (tid == 0) {
        // the neighbour's ghost-row stores were observed through the generic proxy (the flag's ld.acquire); the row
        // copies below read through the async proxy, which needs its own fence to be ordered behind that
        if (SYNC && (edge_n || edge_s)) asm volatile("fence.proxy.async;" ::: "memory");
        ptx::mbarrier_arrive_expect_tx(ptx::sem_release, ptx::scope_cta, ptx::space_shared, &bar, row_bytes * (unsigned)(nrows + 2));
        for (int i = 0; i < nrows + 2; i++)
            ptx::cp_async_bulk(ptx::space_cluster, ptx::space_global, tile + i * SW + (c_lo - (col0 - HALO)),
                               in + (int64_t)(r0 - 1 + i) * pitch + c_lo, row_bytes, &bar);
    }
    // bounded wait: a mismatch between expected and copied bytes must end in a wrong answer, not in a hang
    unsigned tries = 0;
    while (!ptx::mbarrier_try_wait_parity(&bar, 0u)) {
        if (++tries > (1u << 22)) {
            if (tid == 0) *status = 2;
            break;
        }
    }
    float m = 0.0f;
    const int col = col0 + tid * 4;
    const bool live = col < npitch;
    if (col - lane * 4 < npitch) { // else the whole warp is beyond the row (warp-uniform: the shuffles stay converged)
        const bool kx = (col == 0) || (col >= ny - 1), ky = (col + 1 >= ny - 1), kz = (col + 2 >= ny - 1), kw = (col + 3 >= ny - 1);
        const float* c_row = tile + SW + HALO + tid * 4; // row r0, my first column
        float4 n = *reinterpret_cast<const float4*>(c_row - SW);
        float4 c = *reinterpret_cast<const float4*>(c_row);
        float* q = out + (int64_t)r0 * pitch + col;
        for (int i = 0; i < nrows; i++) {
            const float4 s = *reinterpret_cast<const float4*>(c_row + SW);
            float wl = __shfl_up_sync(FULL, c.w, 1);
            float er = __shfl_down_sync(FULL, c.x, 1);
            if (lane == 0) wl = c_row[-1];  // from the tile's halo columns; garbage only where the result is a ring / padding column
            if (lane == 31) er = c_row[4];
            float4 o = upd_quad<ARITH>(c, n, s, wl, er, cx, cy);
            if (kx) o.x = c.x;
            if (ky) o.y = c.y;
            if (kz) o.z = c.z;
            if (kw) o.w = c.w;
            if (live) {
                st4(q, o);
                // K3: the first / last `ghost` updated rows are the neighbours' ghost rows
                if (PEERS && peer_n && r0 + i < row_lo + ghost) st4(peer_n + (int64_t)(r0 + i - row_lo) * pitch + col, o);
                if (PEERS && peer_s && r0 + i >= row_hi - ghost) st4(peer_s + (int64_t)(r0 + i - (row_hi - ghost)) * pitch + col, o);
            }
            if (RESID) {
                m = fmaxf(m, fabsf(__fsub_rn(c.x, o.x)));
                m = fmaxf(m, fabsf(__fsub_rn(c.y, o.y)));
                m = fmaxf(m, fabsf(__fsub_rn(c.z, o.z)));
                m = fmaxf(m, fabsf(__fsub_rn(c.w, o.w)));
            }
            n = c;
            c = s;
            c_row += SW;
            q += pitch;
        }
    }
    if (RESID) block_max_to_global(live ? m : 0.0f, resid);
    if (SYNC) strip_signal(sy, edge_n, edge_s);
}

// ------------------------------------------------------------------ K4 temporal blocking
// T updates per pass over HBM, bit-identical to T single updates.  Pure register pipeline, no
// shared memory, no block barrier: a warp owns 128 consecutive columns (one float4 per lane) and
// streams down the rows; for every time level l < T it keeps the two previous rows of that level
// (A[l], B[l]).  When level-0 row g arrives, level 1 can produce row g-1, level 2 row g-2, ...,
// level T row g-T, which is stored.  E/W neighbours come from shuffles only, so each level loses
// one valid column at both ends of the warp: with T <= 4 the loss stays inside lanes 0 and 31,
// which are loaded but never stored -- warps overlap by 8 columns (column blocks of 120).
// Rows work the same way: a tile that stores rows [r0, r1) starts reading at r0-T and runs to
// r1+T.  Local rows 0 and nrows-1 (ring, or outermost ghost row of a strip) are copied through at
// every level; for a ghost row that makes rows < l wrong at level l, which is why a strip keeps T
// ghost rows and stores only [row_lo, row_hi).
//   grid.x = ceil(ceil(pitch/120) / WARPS), grid.y = row tiles, block = 32*WARPS
// Update of a float4 whose four columns and three rows are all interior: no ring fix-ups.
template <int ARITH>
__device__ __forceinline__ float4 upd4(const float4& c, const float4& n, const float4& s, float cx, float cy)
{
    const float wl = __shfl_up_sync(0xffffffffu, c.w, 1);
    const float er = __shfl_down_sync(0xffffffffu, c.x, 1);
    return upd_quad<ARITH, (HEAT_PACKED_K4 != 0)>(c, n, s, wl, er, cx, cy);
}

// Fast path of K4 for a warp whose 128 columns hold no ring/padding column and whose tile is far
// enough from local rows 0 and nrows-1 (and from a neighbour strip's rows): no row or column
// special cases, no peer stores, no address clamping.
// Every level keeps THREE consecutive rows of its input (S[l][0..2]); the levels are evaluated
// from the last to the first, each reading what the level below produced in the PREVIOUS
// iteration, so the T updates of one iteration are independent of each other (ILP = T) and the
// result of level l is written straight over the oldest row of level l+1.  With the row loop
// unrolled by 3 the slot indices are compile-time constants and no register is ever moved.
// Level l emits row g-(2l+1) of level l+1 in the iteration that loads level-0 row g, so the
// first 3T iterations (starting one row early, at r0-T-1, to make that a multiple of 3) only
// fill the pipeline and every later iteration stores one row.
//   requires r0-T-1 >= 0 and r1+2T+1 <= nrows-1 (all rows touched exist; none is a ring row)
// Three iterations (one turn of the slot rotation).  LMAX = number of levels that compute: while
// the pipeline fills, level l has nothing useful to do before iteration 3(l+1) (level l+1 first
// needs the OLDEST of three rows), so the fill groups run with LMAX = 0, 1, .. T-1: 18 instead of
// 48 level-steps for T=4.  (The drain has no such slack: the last store needs the row the level
// below emitted one iteration earlier.)
template <int ARITH, int T, bool STORE, int LMAX>
__device__ __forceinline__ void tb_fast_3(float4 (&S)[T][3], float4 (&P)[3], const float*& pl, float*& ps, int64_t pitch,
                                          bool stores, int& left, float cx, float cy)
{
#pragma unroll
    for (int ph = 0; ph < 3; ph++) {
#pragma unroll
        for (int l = T - 1; l >= 1; l--) {
            if (l < LMAX) {
                // rows of level l in age order: S[l][ph] (oldest), S[l][ph+1], S[l][ph+2] (newest)
                const float4 y = upd4<ARITH>(S[l][(ph + 1) % 3], S[l][ph], S[l][(ph + 2) % 3], cx, cy);
                if (l == T - 1) {
                    if (STORE) {
                        if (stores && left > 0) st4(ps, y);
                        ps += pitch;
                        left--;
                    }
                } else {
                    S[l + 1][ph] = y;
                }
            }
        }
        S[0][ph] = P[ph]; // the next level-0 row; its slot held the row three above
        P[ph] = ld4(pl);
        pl += pitch;
        if (LMAX >= 1) S[1][ph] = upd4<ARITH>(S[0][(ph + 2) % 3], S[0][(ph + 1) % 3], S[0][ph], cx, cy);
    }
}

template <int ARITH, int T>
__device__ __forceinline__ void tb_fast(const float* __restrict__ in, float* __restrict__ out, int64_t pitch, int col,
                                        bool stores, int r0, int r1, float cx, float cy)
{
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 S[T][3], P[3];
#pragma unroll
    for (int l = 0; l < T; l++) S[l][0] = S[l][1] = S[l][2] = zero;
    const float* pl = in + (int64_t)(r0 - T - 1) * pitch + col; // next level-0 row to load
    float* ps = out + (int64_t)r0 * pitch + col;                 // next row to store
#pragma unroll
    for (int j = 0; j < 3; j++) {
        P[j] = ld4(pl);
        pl += pitch;
    }
    int left = r1 - r0;
    // pipeline fill: 3T iterations, nothing to store yet, level l starts in group l+1
    tb_fast_3<ARITH, T, false, 0>(S, P, pl, ps, pitch, stores, left, cx, cy);
    tb_fast_3<ARITH, T, false, 1>(S, P, pl, ps, pitch, stores, left, cx, cy);
    if constexpr (T >= 3) tb_fast_3<ARITH, T, false, 2>(S, P, pl, ps, pitch, stores, left, cx, cy);
    if constexpr (T >= 4) tb_fast_3<ARITH, T, false, 3>(S, P, pl, ps, pitch, stores, left, cx, cy);
    // steady state: one stored row per iteration
    while (left > 0) tb_fast_3<ARITH, T, true, T>(S, P, pl, ps, pitch, stores, left, cx, cy);
}

template <int ARITH, int T, int WARPS, bool SYNC>
__global__ void __launch_bounds__(WARPS * 32, (ARITH == HEAT_ARITH_F64MIX ? 20 : 28) / WARPS) // <= 72 regs (F32): 7 CTAs of 4 warps
heat_step_tb(const float* __restrict__ in, float* __restrict__ out, int nrows, int row_lo, int row_hi, int ny,
             int64_t pitch, int rows_per_tile, float cx, float cy, float* peer_n, float* peer_s, int reverse,
             strip_sync sy)
{
    static_assert(T >= 2 && T <= 4, "the column halo of T levels must fit in one float4 lane");
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int npitch = (int)pitch;
    // The warps that hold a ring column (first and last warp column) and the first/last row tiles
    // take the generic path, about twice as slow: they are handed out first, so that they are not
    // what the end of the pass waits for.
    const int nwcol = gridDim.x * WARPS, wlin = blockIdx.x * WARPS + (threadIdx.x >> 5);
    const int wcol = (WARPS == 1 && nwcol > 2) ? (wlin == 0 ? 0 : wlin == 1 ? nwcol - 1 : wlin - 1) : wlin;
    const int c0 = wcol * 120 - 4;
    const int col = c0 + lane * 4;
    const bool inrange = col >= 0 && col < npitch;
    const bool stores = inrange && lane >= 1 && lane <= 30;

    const int ntiles = gridDim.y;
    const int tile = tile_of_block<true>((int)blockIdx.y, ntiles, reverse); // edge tiles first, always
    const int r0 = row_lo + tile * rows_per_tile;
    const int r1 = min(r0 + rows_per_tile, row_hi);
    const bool edge_n = r0 < row_lo + T, edge_s = r1 > row_hi - T;
    if (SYNC) strip_wait(sy, edge_n, edge_s);
    if (c0 + 4 < npitch) { // else: nothing to store in this warp (warp-uniform)

    const bool kx = (col <= 0) || (col >= ny - 1);
    const bool ky = (col + 1 >= ny - 1);
    const bool kz = (col + 2 >= ny - 1);
    const bool kw = (col + 3 >= ny - 1);

    // interior tile and no ring/padding/outside column in the whole warp: nothing special to do
    // (the extra margin also keeps such a tile clear of the rows a neighbour strip receives)
    const bool plain_rows = (r0 - T - 1 >= 1) && (r1 + 2 * T + 4 <= nrows - 1) && (!peer_n || r0 >= row_lo + T) &&
                            (!peer_s || r1 <= row_hi - T);
    const bool plain_warp = __all_sync(FULL, inrange && !(kx || ky || kz || kw));
    if (plain_rows && plain_warp) {
        tb_fast<ARITH, T>(in, out, pitch, col, stores, r0, r1, cx, cy);
    } else {
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 A[T], B[T];
#pragma unroll
    for (int l = 0; l < T; l++) A[l] = B[l] = zero;

    const float* p = in + (inrange ? col : 0);
    const int g0 = r0 - T, g1 = r1 + T;
    auto load_row = [&](int g) -> float4 {
        return (inrange && g >= 0 && g < nrows) ? ld4(p + (int64_t)g * pitch) : zero;
    };
    // two rows of look-ahead keep enough loads in flight (traffic is 1/T of the single-step kernel)
    float4 nx0 = load_row(g0), nx1 = load_row(g0 + 1);
    for (int g = g0; g < g1; g++) {
        float4 x = nx0;
        nx0 = nx1;
        nx1 = load_row(g + 2 < g1 ? g + 2 : g1 - 1);
#pragma unroll
        for (int l = 0; l < T; l++) {
            const int i = g - l - 1; // the row level l+1 can produce now
            float4 y = B[l];
            if (i > 0 && i < nrows - 1) {
                float wl = __shfl_up_sync(FULL, B[l].w, 1);
                float er = __shfl_down_sync(FULL, B[l].x, 1);
                float4 o;
                o.x = upd<ARITH>(B[l].x, A[l].x, x.x, B[l].y, wl, cx, cy);
                o.y = upd<ARITH>(B[l].y, A[l].y, x.y, B[l].z, B[l].x, cx, cy);
                o.z = upd<ARITH>(B[l].z, A[l].z, x.z, B[l].w, B[l].y, cx, cy);
                o.w = upd<ARITH>(B[l].w, A[l].w, x.w, er, B[l].z, cx, cy);
                if (!kx) y.x = o.x;
                if (!ky) y.y = o.y;
                if (!kz) y.z = o.z;
                if (!kw) y.w = o.w;
            }
            A[l] = B[l];
            B[l] = x;
            x = y;
        }
        const int orow = g - T;
        if (stores && orow >= r0 && orow < r1) {
            st4(out + (int64_t)orow * pitch + col, x);
            if (peer_n && orow < row_lo + T) st4(peer_n + (int64_t)(orow - row_lo) * pitch + col, x);
            if (peer_s && orow >= row_hi - T) st4(peer_s + (int64_t)(orow - (row_hi - T)) * pitch + col, x);
        }
    }
    } // generic path
    } // warp has columns to store
    if (SYNC) strip_signal(sy, edge_n, edge_s);
}

// ------------------------------------------------------------------ K5 resident field (small grids)
// For grids whose field fits in the shared memory of the whole chip (up to about 1700^2) a pass over
// HBM per update is not the cost -- the launch is (C:204-211 pays one per update; the reference's own
// Table 6 sits at 2-4 us per update).  One cooperative launch runs `steps` updates: CTA b (one per
// SM) keeps its band of rows plus G ghost rows each side in shared memory, in two copies (old /
// new).  It updates the band G times in a row -- the valid part of the ghost zone shrinks by one row
// per update, exactly like the temporal blocking of K4 -- and only then exchanges rows with its two
// neighbours: it publishes its first and last G owned rows in a small global scratch, raises its
// flag (st.release.gpu), waits for the neighbours' flags and copies their rows into its ghost zones.
// Only neighbours synchronise (about 3 us per exchange, hence G updates per exchange); the
// cooperative launch guarantees that all CTAs are resident.  Rows are streamed like in K1 (float4
// per lane, N/C/S in registers, E/W by shuffle), several row groups per CTA working in parallel.
// At the end the last two states go back to the two global buffers.
struct resident_args {
    const float* in;    // newest field (global, pitch-strided)
    float* out_final;   // receives the state after `steps` updates
    float* out_prev;    // receives the state after steps-1 updates (keeps heat_residual meaningful)
    int nrows, ny;
    int64_t pitch;
    int steps;
    int ghost;          // G: updates per exchange = ghost rows per side (<= rows of the shortest band)
    int threads_per_row; // threads that share one row (multiple of 32); blockDim.x / this = row groups
    float cx, cy;
    float* halo;                 // [P][2 sides][2 parities][G][pitch]
    unsigned long long* flags;   // [P] exchanges completed by CTA b (monotonic over launches)
    unsigned long long epoch0;
    int* status;
    unsigned long long timeout_ns;
};

__device__ __forceinline__ void spin_until_gpu(const unsigned long long* f, unsigned long long value,
                                               unsigned long long timeout_ns, int* status)
{
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    unsigned spins = 0;
    for (;;) {
        unsigned long long v;
        asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
        if (v >= value) return;
        if ((++spins & 1023u) == 0) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t - t0 > timeout_ns) {
                *status = 1;
                __threadfence_system();
                return;
            }
        }
    }
}

template <int ARITH>
__global__ void __launch_bounds__(1024, 1) heat_resident(resident_args a)
{
    extern __shared__ float4 resident_smem[];
    float* sm = reinterpret_cast<float*>(resident_smem);
    const unsigned FULL = 0xffffffffu;
    const int P = gridDim.x, b = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const int W = (int)a.pitch, G = a.ghost, ncg = W / 4;
    // balanced split of the interior rows 1..nrows-2 over the P CTAs
    const int interior = a.nrows - 2, base = interior / P, rem = interior % P;
    const int r_lo = 1 + b * base + min(b, rem), rb = base + (b < rem ? 1 : 0);
    // ghost zones: G rows towards a neighbouring CTA, the single (constant) ring row at the two ends
    const bool top_fixed = (b == 0), bot_fixed = (b == P - 1);
    const int gt = top_fixed ? 1 : G, gb = bot_fixed ? 1 : G;
    const int R = rb + gt + gb;            // local rows; owned rows are gt .. gt+rb-1
    const int copy = (base + 1 + 2 * G) * W; // floats per copy (tallest band); copy k starts at sm + k*copy

    // both copies start as the field: ring columns and padding are then never written again
    for (int k = tid; k < R * ncg; k += blockDim.x) {
        const int lr = k / ncg, c4 = k - lr * ncg;
        const float4 v = ld4(a.in + (int64_t)(r_lo - gt + lr) * a.pitch + c4 * 4);
        st4(sm + lr * W + c4 * 4, v);
        st4(sm + copy + lr * W + c4 * 4, v);
    }
    __syncthreads();

    const int tpr = a.threads_per_row, rgroups = blockDim.x / tpr, rg = tid / tpr, ct = tid - rg * tpr;
    float* const halo_mine = a.halo + (size_t)b * 4 * G * W;
    unsigned long long exchanges = a.epoch0;
    int s = 0;
    while (s < a.steps) {
        const int g = min(G, a.steps - s);
        for (int k = 0; k < g; k++) {
            const float* old = sm + ((s + k) & 1) * copy;
            float* nw = sm + (((s + k) & 1) ^ 1) * copy;
            // rows that can still be computed from valid neighbours: the zone shrinks towards the band
            const int lo = top_fixed ? 1 : k + 1, hi = bot_fixed ? R - 2 : R - 2 - k;
            const int chunk = (hi - lo + 1 + rgroups - 1) / rgroups;
            const int my_lo = lo + rg * chunk, my_hi = min(my_lo + chunk - 1, hi);
            if (my_lo <= my_hi) {
                for (int cg0 = (ct & ~31); cg0 < ncg; cg0 += tpr) {
                    int cg = cg0 + lane;
                    const bool live = cg < ncg;
                    if (!live) cg = ncg - 1;
                    const int col = cg * 4;
                    const bool kx = (col == 0) || (col >= a.ny - 1), ky = (col + 1 >= a.ny - 1), kz = (col + 2 >= a.ny - 1),
                               kw = (col + 3 >= a.ny - 1);
                    const bool e0 = lane == 0 && col > 0, e31 = lane == 31 && col + 4 < W;
                    float4 n = ld4(old + (my_lo - 1) * W + col), c = ld4(old + my_lo * W + col);
                    for (int i = my_lo; i <= my_hi; i++) {
                        const float4 sd = ld4(old + (i + 1) * W + col);
                        float wl = __shfl_up_sync(FULL, c.w, 1);
                        float er = __shfl_down_sync(FULL, c.x, 1);
                        if (e0) wl = old[i * W + col - 1];
                        if (e31) er = old[i * W + col + 4];
                        float4 o;
                        o.x = upd<ARITH>(c.x, n.x, sd.x, c.y, wl, a.cx, a.cy);
                        o.y = upd<ARITH>(c.y, n.y, sd.y, c.z, c.x, a.cx, a.cy);
                        o.z = upd<ARITH>(c.z, n.z, sd.z, c.w, c.y, a.cx, a.cy);
                        o.w = upd<ARITH>(c.w, n.w, sd.w, er, c.z, a.cx, a.cy);
                        if (kx) o.x = c.x;
                        if (ky) o.y = c.y;
                        if (kz) o.z = c.z;
                        if (kw) o.w = c.w;
                        if (live) st4(nw + i * W + col, o);
                        n = c;
                        c = sd;
                    }
                }
            }
            __syncthreads();
        }
        s += g;
        if (s == a.steps) break; // nothing left to exchange for
        // ---- exchange: my first / last G owned rows of the current state go to the neighbours
        float* cur = sm + (s & 1) * copy;
        const int par = (int)(exchanges & 1);
        for (int k = tid; k < G * ncg; k += blockDim.x) {
            const int lr = k / ncg, c4 = k - lr * ncg;
            if (!top_fixed)
                __stcg(reinterpret_cast<float4*>(halo_mine + ((0 * 2 + par) * G + lr) * W) + c4,
                       *reinterpret_cast<const float4*>(cur + (gt + lr) * W + c4 * 4));
            if (!bot_fixed)
                __stcg(reinterpret_cast<float4*>(halo_mine + ((1 * 2 + par) * G + lr) * W) + c4,
                       *reinterpret_cast<const float4*>(cur + (gt + rb - G + lr) * W + c4 * 4));
        }
        __syncthreads();
        exchanges++;
        if (tid == 0) {
            __threadfence();
            asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(a.flags + b), "l"(exchanges) : "memory");
            if (!top_fixed) spin_until_gpu(a.flags + b - 1, exchanges, a.timeout_ns, a.status);
            if (!bot_fixed) spin_until_gpu(a.flags + b + 1, exchanges, a.timeout_ns, a.status);
        }
        __syncthreads();
        // the north neighbour's last G rows are my top ghost zone, the south neighbour's first G rows my bottom one
        for (int k = tid; k < G * ncg; k += blockDim.x) {
            const int lr = k / ncg, c4 = k - lr * ncg;
            if (!top_fixed)
                st4(cur + lr * W + c4 * 4,
                    __ldcg(reinterpret_cast<const float4*>(a.halo + (((size_t)(b - 1) * 4 + 1 * 2 + par) * G + lr) * W) + c4));
            if (!bot_fixed)
                st4(cur + (gt + rb + lr) * W + c4 * 4,
                    __ldcg(reinterpret_cast<const float4*>(a.halo + (((size_t)(b + 1) * 4 + 0 * 2 + par) * G + lr) * W) + c4));
        }
        __syncthreads();
    }
    // ---- the last two states go back to HBM (owned rows only; ring rows never change)
    const float* fin = sm + (a.steps & 1) * copy;
    const float* prev = sm + ((a.steps & 1) ^ 1) * copy;
    for (int k = tid; k < rb * ncg; k += blockDim.x) {
        const int lr = gt + k / ncg, c4 = k - (lr - gt) * ncg;
        const int64_t g = (int64_t)(r_lo - gt + lr) * a.pitch + c4 * 4;
        st4(a.out_final + g, ld4(fin + lr * W + c4 * 4));
        if (a.out_prev) st4(a.out_prev + g, ld4(prev + lr * W + c4 * 4));
    }
}

// ------------------------------------------------------------------ double-precision grids
// The reference hard-codes `float` (M:46, C:170); this is its update with the element type widened: every value
// and every operation in double, cx/cy the float constants promoted (struct Parms stays float, M:29-32).
// FMA = true contracts the two multiply-adds like nvcc does for the CUDA source, false rounds every operation
// like gcc does for the MPI source.  16 B per point-update, so the HBM roofline is half the float one in LUP/s.
// Same structure as K1/K2/K3: lanes along iy, each thread owns 2 consecutive columns (one 16-byte access), N/C/S
// rows in registers, E/W by shuffle, first/last `ghost` rows also stored into the neighbour strips, step flags
// fused in.  Ring columns are written back unchanged, ring rows never written.
template <bool FMA>
__device__ __forceinline__ double upd64(double c, double n, double s, double e, double w, double cx, double cy)
{
    const double t1 = __dsub_rn(__dadd_rn(s, n), __dadd_rn(c, c));
    const double t2 = __dsub_rn(__dadd_rn(e, w), __dadd_rn(c, c));
    if (FMA) return __fma_rn(cy, t2, __fma_rn(cx, t1, c));
    return __dadd_rn(__dadd_rn(c, __dmul_rn(cx, t1)), __dmul_rn(cy, t2));
}

__global__ void heat_inidat_f64_kernel(double* __restrict__ a, double* __restrict__ b, int nx, int ny, int row_lo, int nloc,
                                       int64_t pitch, int col_lo, int nyloc)
{
    const int64_t total = (int64_t)nloc * pitch;
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < total; k += (int64_t)gridDim.x * blockDim.x) {
        const int lr = (int)(k / pitch);
        const int ly = (int)(k - (int64_t)lr * pitch);
        const int ix = row_lo + lr, iy = col_lo + ly;
        double v = 0.0;
        if (ly < nyloc) {
            uint32_t p = (uint32_t)ix * (uint32_t)(nx - ix - 1) * (uint32_t)iy * (uint32_t)(ny - iy - 1);
            v = (double)(int32_t)p; // the int product wraps before the conversion, whatever the element type (M:320)
        }
        a[k] = v;
        b[k] = v;
    }
}

__device__ __forceinline__ void block_max_to_global64(double m, unsigned long long* gmax)
{
    // non-negative doubles order like their bit patterns
    unsigned long long bits = (unsigned long long)__double_as_longlong(m);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, bits, o);
        bits = other > bits ? other : bits;
    }
    __shared__ unsigned long long smax64[32];
    const int warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
    if ((threadIdx.x & 31) == 0) smax64[warp] = bits;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long v = 0;
        for (int w = 0; w < nwarp; w++) v = smax64[w] > v ? smax64[w] : v;
        if (v != 0ull) atomicMax(gmax, v);
    }
}

template <bool FMA, bool RESID, bool SYNC, bool PEERS>
__global__ void __launch_bounds__(128, 8)
heat_step_f64(const double* __restrict__ in, double* __restrict__ out, int row_lo, int row_hi, int ny, int64_t pitch,
              int rows_per_tile, double cx, double cy, unsigned long long* __restrict__ resid, double* peer_n, double* peer_s,
              int ghost, int reverse, strip_sync sy)
{
    constexpr int UNROLL = 4;
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int npitch = (int)pitch;
    int col = ((int)blockIdx.x * 128 + threadIdx.x) * 2;
    const int ntiles = gridDim.y;
    const int tile = tile_of_block<SYNC>((int)blockIdx.y, ntiles, reverse);
    const int r0 = row_lo + tile * rows_per_tile;
    const int r1 = min(r0 + rows_per_tile, row_hi);
    const bool edge_n = r0 < row_lo + ghost, edge_s = r1 > row_hi - ghost;
    if (SYNC) strip_wait(sy, edge_n, edge_s);

    double m = 0.0;
    const bool live = col < npitch;
    if (col - lane * 2 < npitch) { // else the whole warp is beyond the row (warp-uniform)
        if (!live) col = npitch - 2; // keep the lane in bounds for loads; it never stores
        const bool kx = (col == 0) || (col >= ny - 1), ky = (col + 1 >= ny - 1); // ring / padding columns: copied through
        const bool edge_lane = (lane == 0 && col > 0) || (lane == 31 && col + 2 < npitch);
        const int edge_off = (lane == 0) ? -1 : 2;
        const double* p = in + (int64_t)(r0 - 1) * pitch + col;
        double* q = out + (int64_t)r0 * pitch + col;
        double2 n = *reinterpret_cast<const double2*>(p);
        double2 c = *reinterpret_cast<const double2*>(p + pitch);
        double ec = edge_lane ? p[pitch + edge_off] : 0.0;
        p += 2 * pitch;
        for (int r = r0; r < r1; r += UNROLL) {
            double2 s[UNROLL];
            double es[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; u++) {
                const int64_t off = (int64_t)min(u, r1 - 1 - r) * pitch; // rows past the tile: clamped, loaded, unused
                s[u] = *reinterpret_cast<const double2*>(p + off);
                es[u] = edge_lane ? p[off + edge_off] : 0.0;
            }
#pragma unroll
            for (int u = 0; u < UNROLL; u++) {
                if (r + u < r1) {
                    double wl = __shfl_up_sync(FULL, c.y, 1);
                    double er = __shfl_down_sync(FULL, c.x, 1);
                    if (lane == 0) wl = ec;
                    if (lane == 31) er = ec;
                    double2 o;
                    o.x = upd64<FMA>(c.x, n.x, s[u].x, c.y, wl, cx, cy);
                    o.y = upd64<FMA>(c.y, n.y, s[u].y, er, c.x, cx, cy);
                    if (kx) o.x = c.x;
                    if (ky) o.y = c.y;
                    if (live) {
                        *reinterpret_cast<double2*>(q) = o;
                        if (PEERS && peer_n && r + u < row_lo + ghost)
                            *reinterpret_cast<double2*>(peer_n + (int64_t)(r + u - row_lo) * pitch + col) = o;
                        if (PEERS && peer_s && r + u >= row_hi - ghost)
                            *reinterpret_cast<double2*>(peer_s + (int64_t)(r + u - (row_hi - ghost)) * pitch + col) = o;
                    }
                    if (RESID) {
                        m = fmax(m, fabs(__dsub_rn(c.x, o.x)));
                        m = fmax(m, fabs(__dsub_rn(c.y, o.y)));
                    }
                    n = c;
                    c = s[u];
                    ec = es[u];
                    q += pitch;
                }
            }
            p += (int64_t)UNROLL * pitch;
        }
    }
    if (RESID) block_max_to_global64(live ? m : 0.0, resid);
    if (SYNC) strip_signal(sy, edge_n, edge_s);
}

__global__ void __launch_bounds__(256)
heat_absdiff_max_f64(const double* __restrict__ a, const double* __restrict__ b, int row_lo, int row_hi, int ny, int64_t pitch,
                     unsigned long long* __restrict__ resid)
{
    double m = 0.0;
    const int64_t total = (int64_t)(row_hi - row_lo) * pitch;
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < total; k += (int64_t)gridDim.x * blockDim.x) {
        const int lr = (int)(k / pitch);
        const int iy = (int)(k - (int64_t)lr * pitch);
        if (iy < ny) {
            const int64_t idx = k + (int64_t)row_lo * pitch;
            m = fmax(m, fabs(__dsub_rn(a[idx], b[idx])));
        }
    }
    block_max_to_global64(m, resid);
}

// ------------------------------------------------------------------ 3-D, 7-point
// The reference is 2-D only; this is its update with a third axis: u[ix][iy][iz], iz fastest,
//   new = u + cx*((X+ + X-) - 2u) + cy*((Y+ + Y-) - 2u) + cz*((Z+ + Z-) - 2u),   evaluated left to right,
// in the same three arithmetics (F32_FMA: one more fused multiply-add on the chain; MPI_F64MIX: float
// neighbour sums, the rest in double).  All six faces are the ring.  A "row" of the handle is one ix-plane
// (ny*nz floats), so upload, download, the streamed one-shot and the residual pass are the 2-D ones.
// Register-blocked column streaming, one dimension up: lanes run along iz (one float4 each), a thread owns TY
// (2) consecutive iy rows of its four columns and streams along ix with the previous, current and next plane of
// those rows in registers.  Y neighbours inside the thread's rows are its own registers, the two rows beyond
// come with the plane (TY+2 loads for TY outputs: 4 for 2, the extra ones L1/L2 hits of the neighbouring warps), Z
// neighbours are shuffles (across the warp edge: one 4-byte load).  No shared memory.
template <int ARITH>
__device__ __forceinline__ float upd3(float c, float xs, float ys, float zs, float cx, float cy, float cz)
{
    // xs, ys, zs: the three float neighbour sums
    if constexpr (ARITH == HEAT_ARITH_F32_FMA) {
        const float c2 = __fadd_rn(c, c);
        return __fmaf_rn(cz, __fsub_rn(zs, c2), __fmaf_rn(cy, __fsub_rn(ys, c2), __fmaf_rn(cx, __fsub_rn(xs, c2), c)));
    } else if constexpr (ARITH == HEAT_ARITH_F32_NOFMA) {
        const float c2 = __fadd_rn(c, c);
        float r = __fadd_rn(c, __fmul_rn(cx, __fsub_rn(xs, c2)));
        r = __fadd_rn(r, __fmul_rn(cy, __fsub_rn(ys, c2)));
        return __fadd_rn(r, __fmul_rn(cz, __fsub_rn(zs, c2)));
    } else {
        const double cd = (double)c, c2 = __dadd_rn(cd, cd);
        double r = __dadd_rn(cd, __dmul_rn((double)cx, __dsub_rn((double)xs, c2)));
        r = __dadd_rn(r, __dmul_rn((double)cy, __dsub_rn((double)ys, c2)));
        r = __dadd_rn(r, __dmul_rn((double)cz, __dsub_rn((double)zs, c2)));
        return __double2float_rn(r);
    }
}

__global__ void heat_inidat3d_kernel(float* __restrict__ a, float* __restrict__ b, int nx, int ny, int nz, int plane_lo, int nplanes,
                                     int64_t pitch)
{
    const int64_t total = (int64_t)nplanes * pitch, plane = (int64_t)ny * nz;
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < total; k += (int64_t)gridDim.x * blockDim.x) {
        const int lx = (int)(k / pitch), ix = plane_lo + lx; // local array = global planes [plane_lo, plane_lo + nplanes)
        const int64_t in_plane = k - (int64_t)lx * pitch;
        float v = 0.0f;
        if (in_plane < plane) {
            const int iy = (int)(in_plane / nz), iz = (int)(in_plane - (int64_t)iy * nz);
            // the 2-D product of M:320 with a third factor pair, in the same wrapping int arithmetic
            uint32_t p = (uint32_t)ix * (uint32_t)(nx - ix - 1) * (uint32_t)iy * (uint32_t)(ny - iy - 1) * (uint32_t)iz *
                         (uint32_t)(nz - iz - 1);
            v = (float)(int32_t)p;
        }
        a[k] = v;
        b[k] = v;
    }
}

//   grid.x = ceil(nz/4 / 32), grid.y = ceil((ny-2) / (TY*WARPS)), grid.z = plane tiles; block = 32*WARPS
//   planes [plane_lo, plane_hi) are updated (planes plane_lo-1 and plane_hi must exist)
// Strips of planes across GPUs work like the 2-D row strips (K3): the first / last owned plane is also stored into
// the neighbour's ghost plane, the tiles that touch a neighbour wait for its flag and the last of them raises it.
#ifndef HEAT_3D_MINBLOCKS
#define HEAT_3D_MINBLOCKS 4 // resident CTAs per SM asked of the compiler: 4 = 64 registers, no spills.  Measured at 640^3 /
#endif                      // 1024^3: unconstrained (78 registers, 3 CTAs) 649 / 652 GLUP/s, 4: 716 / 709, 5 (48 registers, spills) 450 / 449

// PEERS = this strip has a neighbour to store ghost planes into (a single strip's instantiation carries none of that code)
template <int ARITH, bool RESID, int TY, int WARPS, bool SYNC, bool PEERS>
__global__ void __launch_bounds__(32 * WARPS, HEAT_3D_MINBLOCKS)
heat_step_3d(const float* __restrict__ in, float* __restrict__ out, int plane_lo, int plane_hi, int ny, int nz, int64_t pitch,
             int planes_per_tile, float cx, float cy, float cz, unsigned* __restrict__ resid, float* peer_n, float* peer_s,
             strip_sync sy)
{
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int col = ((int)blockIdx.x * 32 + lane) * 4;                    // my first iz
    const int y0 = 1 + ((int)blockIdx.y * WARPS + warp) * TY;         // my first iy (interior rows start at 1)
    const int tile = tile_of_block<SYNC>((int)blockIdx.z, (int)gridDim.z, 0); // with flags: the two edge tiles first
    const int x0 = plane_lo + tile * planes_per_tile, x1 = min(x0 + planes_per_tile, plane_hi);
    const bool edge_n = x0 == plane_lo, edge_s = x1 == plane_hi; // one ghost plane per neighbour (single updates)
    if (SYNC) strip_wait(sy, edge_n, edge_s);
    float m = 0.0f;
    const bool live = col < nz;
    if (y0 < ny - 1 && col - lane * 4 < nz) { // else this warp has no row / no column to update (warp-uniform)
        if (!live) col = nz - 4; // keep the lane in bounds for loads; it never stores
        const bool kx = (col == 0), kw = (col + 3 == nz - 1); // ring columns iz = 0 and nz-1: copied through
        const bool e0 = lane == 0 && col > 0, e31 = lane == 31 && col + 4 < nz;
        // rows y0-1 .. y0+TY of a plane; rows past ny-1 are clamped (loaded, never used for a store)
        int roff[TY + 2]; // offsets inside a plane: ny*nz fits an int (checked by heat_create)
#pragma unroll
        for (int k = 0; k < TY + 2; k++) roff[k] = min(y0 - 1 + k, ny - 1) * nz + col;
        float4 n[TY], c[TY + 2], s[TY + 2];
        const float* pn = in + (int64_t)(x0 - 1) * pitch;
#pragma unroll
        for (int k = 0; k < TY; k++) n[k] = ld4(pn + roff[k + 1]);
#pragma unroll
        for (int k = 0; k < TY + 2; k++) c[k] = ld4(pn + pitch + roff[k]);
        for (int x = x0; x < x1; x++) {
            const float* pc = in + (int64_t)x * pitch;
            const float* ps = pc + pitch;
#pragma unroll
            for (int k = 0; k < TY + 2; k++) s[k] = ld4(ps + roff[k]);
            float* q = out + (int64_t)x * pitch;
#pragma unroll
            for (int k = 1; k <= TY; k++) {
                const int y = y0 - 1 + k;
                float wl = __shfl_up_sync(FULL, c[k].w, 1);
                float er = __shfl_down_sync(FULL, c[k].x, 1);
                if (y < ny - 1) {
                    // the ring rows iy = 0 and ny-1 are written back unchanged by their neighbours' owner (like the
                    // ring columns below), so that `out` needs no copy of the field beforehand
                    if (y == 1 && live) st4(q + roff[0], c[0]);
                    if (y == ny - 2 && live) st4(q + roff[k + 1], c[k + 1]);
                    if (e0) wl = ld1(pc + roff[k] - 1);
                    if (e31) er = ld1(pc + roff[k] + 4);
                    const float4 cc = c[k], up = c[k - 1], dn = c[k + 1], xm = n[k - 1], xp = s[k];
                    float4 o;
                    o.x = upd3<ARITH>(cc.x, __fadd_rn(xp.x, xm.x), __fadd_rn(dn.x, up.x), __fadd_rn(cc.y, wl), cx, cy, cz);
                    o.y = upd3<ARITH>(cc.y, __fadd_rn(xp.y, xm.y), __fadd_rn(dn.y, up.y), __fadd_rn(cc.z, cc.x), cx, cy, cz);
                    o.z = upd3<ARITH>(cc.z, __fadd_rn(xp.z, xm.z), __fadd_rn(dn.z, up.z), __fadd_rn(cc.w, cc.y), cx, cy, cz);
                    o.w = upd3<ARITH>(cc.w, __fadd_rn(xp.w, xm.w), __fadd_rn(dn.w, up.w), __fadd_rn(er, cc.z), cx, cy, cz);
                    if (kx) o.x = cc.x;
                    if (kw) o.w = cc.w;
                    if (live) {
                        st4(q + roff[k], o);
                        if (PEERS && peer_n && x == plane_lo) st4(peer_n + roff[k], o);
                        if (PEERS && peer_s && x == plane_hi - 1) st4(peer_s + roff[k], o);
                    }
                    if (RESID) {
                        m = fmaxf(m, fabsf(__fsub_rn(cc.x, o.x)));
                        m = fmaxf(m, fabsf(__fsub_rn(cc.y, o.y)));
                        m = fmaxf(m, fabsf(__fsub_rn(cc.z, o.z)));
                        m = fmaxf(m, fabsf(__fsub_rn(cc.w, o.w)));
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < TY; k++) n[k] = c[k + 1];
#pragma unroll
            for (int k = 0; k < TY + 2; k++) c[k] = s[k];
        }
    }
    if (RESID) block_max_to_global(live ? m : 0.0f, resid);
    if (SYNC) strip_signal(sy, edge_n, edge_s);
}

// ------------------------------------------------------------------ residual of two buffers
// max |a-b| over the interior of a strip (heat_residual(): a separate pass, not on the hot path)
__global__ void __launch_bounds__(256)
heat_absdiff_max(const float* __restrict__ a, const float* __restrict__ b, int row_lo, int row_hi, int ny, int64_t pitch,
                 unsigned* __restrict__ resid)
{
    float m = 0.0f;
    const int64_t total = (int64_t)(row_hi - row_lo) * pitch;
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < total; k += (int64_t)gridDim.x * blockDim.x) {
        const int lr = (int)(k / pitch);
        const int iy = (int)(k - (int64_t)lr * pitch);
        if (iy < ny) {
            const int64_t idx = k + (int64_t)row_lo * pitch; // first updated row
            m = fmaxf(m, fabsf(__fsub_rn(a[idx], b[idx])));
        }
    }
    block_max_to_global(m, resid);
}

// 2-D blocks: after an upload both buffers hold the ghost columns in the field; copy them into the dense
// arrays the update kernel reads (gcol layout: [buffer][side: 0 west, 1 east][local row]).
__global__ void heat_gather_ghost_columns(const float* __restrict__ b0, const float* __restrict__ b1, int nloc, int ny,
                                          int64_t pitch, float* __restrict__ gcol)
{
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < nloc; r += gridDim.x * blockDim.x) {
        gcol[0 * (int64_t)nloc + r] = b0[(int64_t)r * pitch];
        gcol[1 * (int64_t)nloc + r] = b0[(int64_t)r * pitch + ny - 1];
        gcol[2 * (int64_t)nloc + r] = b1[(int64_t)r * pitch];
        gcol[3 * (int64_t)nloc + r] = b1[(int64_t)r * pitch + ny - 1];
    }
}

// ------------------------------------------------------------------ strip flags
// After update u completes, a strip writes u+1 into each neighbour's flag slot
// (system-scope release: its ghost-row stores are visible before the flag).
__global__ void heat_signal(unsigned long long* peer_flag_n, unsigned long long* peer_flag_s, unsigned long long* peer_flag_w,
                            unsigned long long* peer_flag_e, unsigned long long value)
{
    if (threadIdx.x == 0) {
        __threadfence_system();
        unsigned long long* f[4] = {peer_flag_n, peer_flag_s, peer_flag_w, peer_flag_e};
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (f[k]) asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(f[k]), "l"(value) : "memory");
    }
}

// Before update u+1 a strip waits until its neighbours have finished update u (also used before an upload
// overwrites a strip whose neighbours may still be storing ghost rows into it).
// Bounded: after timeout_ns the kernel gives up and raises *status (checked by the host).
__global__ void heat_wait(const unsigned long long* flag_n, const unsigned long long* flag_s, const unsigned long long* flag_w,
                          const unsigned long long* flag_e, unsigned long long value, unsigned long long timeout_ns, int* status)
{
    if (threadIdx.x != 0) return;
    const unsigned long long* f[4] = {flag_n, flag_s, flag_w, flag_e};
    for (int side = 0; side < 4; side++)
        if (f[side]) spin_until(f[side], value, timeout_ns, status);
}

} // namespace heatk
