// nbody_kernels.cuh -- sm_100a kernels of the direct-sum hot path.
//
// Reference computation being replaced (citations relative to the reference repo):
//   force accumulation  src/NaiveSimulation.cpp:12-41, src/MpiNaiveSimulation.cpp:15-125
//   integrator          src/NaiveSimulation.cpp:43-51, src/MpiNaiveSimulation.cpp:127-141
//   AoS record          include/Particle.hpp:53-59
//
// Data layout in HBM:
//   pos[2][n_slots]  float4 {x, y, z, mass}   global, double-buffered, every rank holds all slots
//   vel[n_local]     float4 {vx, vy, vz, -}   local slice
//   acc[n_local]     float4 {ax, ay, az, -}   local slice
//   partial[(grid + n_itiles) * TI] float4     scratch of the fast force kernel
// Slot s = r * slot_stride + l holds local particle l of rank r; slots past a rank's count are
// zero-mass padding parked far away, so they contribute exactly 0 to every sum.
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <curand_kernel.h>
#include <stdint.h>

namespace nbody {

constexpr int kThreads = 256;      // threads per CTA, all kernels
constexpr int kTileJ = 256;        // source slots per shared-memory tile (= kThreads: one load per thread)
constexpr int kFlushTiles = 16;     // source tiles per inner accumulation block of the fast kernel (4096 sources; power of two)
constexpr int kSlotAlign = 256;    // slot_stride granularity (= kTileJ, so slices start on tile edges)
constexpr float kPadCoord = 1.0e18f;  // padding slots sit at (kPadCoord, kPadCoord, kPadCoord), mass 0
constexpr int kMaxWorld = 16;

typedef unsigned long long u64;

// ---------------------------------------------------------------------------------------------
// packed FP32x2 arithmetic (sm_100: FADD2 / FMUL2 / FFMA2) and the approximate rsqrt (MUFU.RSQ)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ u64 pack2(float lo, float hi) {
    u64 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(u64 v, float &lo, float &hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ u64 sub2(u64 a, u64 b) {
    u64 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ u64 mul2(u64 a, u64 b) {
    u64 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
    u64 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ float rsqrt_approx(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// ---------------------------------------------------------------------------------------------
// Work plan of the fast force kernel.
//
// Targets are cut into tiles of 256*I targets (I targets per thread): n_main tiles of I_main, then
// up to two narrower tail tiles (I in {4,2,1}) for the remainder, so that fewer than 256 padded
// target rows are ever computed.  Each tile sweeps all n_slots sources, counted in granules of
// kGran sources.  A granule of a tile with I targets per thread costs cost(I); the flattened
// (tile, granule) sequence is cut into `grid` runs of equal COST, CTA c of a persistent grid
// taking the granules [first_gran(c), first_gran(c+1)).  Inside one tile such a run is a
// "segment"; every segment leaves one partial sum per target in scratch slot (c + t), and the
// finishing kernel adds the segments of a tile in ascending c, so the result does not depend on
// timing.  grid <= cost_total / max granule cost, hence no CTA is empty.
// ---------------------------------------------------------------------------------------------
constexpr int kGran = 16;          // sources per plan granule (8 packed pairs)
constexpr int kMaxTail = 2;

// which inner loop a force kernel runs
constexpr int kLoopUnit = 0;       // all masses equal: m_j dropped from the loop (11 FMA-pipe ops per pair)
constexpr int kLoopMassMul = 1;    // s = m_j / r^3 by an explicit multiply (12 ops per pair)

struct ForcePlan {
    int n_gran;                 // granules per target tile = n_slots / kGran
    int n_main;                 // target tiles of 256 * I_main targets
    int n_tail;                 // 0..kMaxTail narrower tiles after them
    int I_main;
    int tail_I[kMaxTail];
    int cost_main;
    int cost_tail[kMaxTail];
    int grid;                   // G
    long long cost_total;       // C
};

// relative cost of one granule of a tile with I targets per thread.  Narrow tiles run well below the
// rate of the I = 7 loop (fewer interactions per shared-memory read and per loop iteration); calibrated
// on B200 so that the CTAs holding tail tiles finish with the others (gpurun_out/r2c_ab.log)
__host__ __device__ __forceinline__ int plan_granule_cost(int I) { return I == 7 ? 73 : 10 * I + 7; }
__host__ __device__ __forceinline__ int plan_tiles(const ForcePlan &p) { return p.n_main + p.n_tail; }
__host__ __device__ __forceinline__ int plan_tile_I(const ForcePlan &p, int t) {
    return t < p.n_main ? p.I_main : (t == p.n_main ? p.tail_I[0] : p.tail_I[1]);
}
// first local target of tile t
__host__ __device__ __forceinline__ int plan_tile_target0(const ForcePlan &p, int t) {
    if (t <= p.n_main) return t * kThreads * p.I_main;
    return (p.n_main * p.I_main + p.tail_I[0]) * kThreads;      // second tail tile
}
// tile that holds local target l, and l's index inside it
__host__ __device__ __forceinline__ void plan_tile_of_target(const ForcePlan &p, int l, int &t, int &r) {
    const int TI = kThreads * p.I_main;
    if (l < p.n_main * TI) { t = l / TI; r = l - t * TI; return; }
    r = l - p.n_main * TI;
    t = p.n_main;
    if (p.n_tail == 2 && r >= kThreads * p.tail_I[0]) { r -= kThreads * p.tail_I[0]; t++; }
}
// cost position at which global granule g (= t * n_gran + granule within tile) starts; g may be the end
__host__ __device__ __forceinline__ long long plan_cost_start(const ForcePlan &p, long long g) {
    const long long main_gran = (long long)p.n_main * p.n_gran;
    if (g <= main_gran) return g * p.cost_main;
    long long c = main_gran * p.cost_main;
    g -= main_gran;
    if (g <= p.n_gran || p.n_tail < 2) return c + g * p.cost_tail[0];
    return c + (long long)p.n_gran * p.cost_tail[0] + (g - p.n_gran) * p.cost_tail[1];
}
// granule that contains cost position pos (pos == cost_total: the end)
__host__ __device__ __forceinline__ long long plan_gran_at(const ForcePlan &p, long long pos) {
    const long long main_gran = (long long)p.n_main * p.n_gran;
    const long long cm = main_gran * p.cost_main;
    if (pos < cm) return pos / p.cost_main;
    pos -= cm;
    if (p.n_tail == 0) return main_gran;
    const long long c0 = (long long)p.n_gran * p.cost_tail[0];
    if (pos < c0) return main_gran + pos / p.cost_tail[0];
    if (p.n_tail == 1) return main_gran + p.n_gran;
    pos -= c0;
    if (pos < (long long)p.n_gran * p.cost_tail[1]) return main_gran + p.n_gran + pos / p.cost_tail[1];
    return main_gran + 2ll * p.n_gran;
}
__host__ __device__ __forceinline__ long long plan_first_gran(const ForcePlan &p, int c) {
    return plan_gran_at(p, (long long)c * p.cost_total / p.grid);
}
// CTA that owns granule g: the largest c with first_gran(c) <= g
__host__ __device__ __forceinline__ int plan_owner(const ForcePlan &p, long long g) {
    return (int)((plan_cost_start(p, g + 1) * p.grid + p.cost_total - 1) / p.cost_total) - 1;
}
// scratch the plan needs, in float4
__host__ __device__ __forceinline__ size_t plan_partial_slots(const ForcePlan &p) {
    return ((size_t)p.grid + plan_tiles(p)) * kThreads * p.I_main;
}

// One source tile (all 128 pairs when kFullTile, else pairs [q0, q1), multiples of kGran / 2) against the I
// targets of this thread, packed two sources per instruction.  The whole-tile form has compile-time
// loop bounds on purpose: only then does ptxas keep the loop counter and the shared-memory addresses
// in uniform registers; with bounds that come out of the plan arithmetic it falls back to vector
// registers, whose reads compete with the FFMA2 operands (1.5 % slower, tools/sass_rf_model.py).  smem tile layout: A[q] = {x[2q], x[2q+1], y[2q], y[2q+1]},
// B[q] = {z[2q], z[2q+1], m[2q], m[2q+1]}.
//   kLoopUnit     m unused.  3 FADD2 + 3 FFMA2 + 2 MUFU.RSQ + 2 FMUL2 + 3 FFMA2 per source pair and target.
//   kLoopMassMul  one more FMUL2 (s = m_j / r^3).
// (Folding m_j into pre-scaled sources instead -- 11 operations for unequal masses too -- was measured 2.6 %
// slower: its differences become three-operand FFMA2s; tools/exp/notes/tried_and_not_kept.md.)
template <int I, bool kMaskSelf, int kLoop, bool kFullTile>
__device__ __forceinline__ void tile_interact(const float4 *__restrict__ sA, const float4 *__restrict__ sB,
                                              const u64 (&xi)[I], const u64 (&yi)[I], const u64 (&zi)[I],
                                              u64 eps2, u64 (&ax)[I], u64 (&ay)[I], u64 (&az)[I],
                                              int q0, int q1, int tile_slot0, int self_slot0) {
    // narrow tiles have too few independent chains per warp to cover the MUFU latency: unroll them further
    constexpr int kUnrollQ = (I >= 4) ? 2 : (I == 2 ? 4 : 8);
#pragma unroll kUnrollQ
    for (int q = kFullTile ? 0 : q0; q < (kFullTile ? kTileJ / 2 : q1); ++q) {
        const float4 a = sA[q];
        const float4 b = sB[q];
        const u64 xj = pack2(a.x, a.y), yj = pack2(a.z, a.w), zj = pack2(b.x, b.y), mj = pack2(b.z, b.w);
#pragma unroll
        for (int k = 0; k < I; ++k) {
            const u64 dx = sub2(xj, xi[k]), dy = sub2(yj, yi[k]), dz = sub2(zj, zi[k]);
            // How this body is written (order of the squares, of the accumulations and their operands, which lane's
            // MUFU comes first) changes nothing arithmetically worth mentioning, but it changes ptxas's schedule, and
            // the schedules differ by up to 4 % on the hardware (MUFU results consumed too early at the loop end).
            // This is the form that measured fastest for BOTH loops in a 93-variant sweep on B200
            // (profiles/r2_kernel_variant_matrix.md).
            u64 r2 = fma2(dx, dx, eps2);
            r2 = fma2(dy, dy, r2);
            r2 = fma2(dz, dz, r2);
            float r2lo, r2hi;
            unpack2(r2, r2lo, r2hi);
            const float rhi = rsqrt_approx(r2hi), rlo = rsqrt_approx(r2lo);      // high lane first
            const u64 rinv = pack2(rlo, rhi);
            const u64 rinv2 = mul2(rinv, rinv);
            u64 s = (kLoop == kLoopMassMul) ? mul2(rinv, mul2(mj, rinv2)) : mul2(rinv, rinv2);   // [m_j] / r^3
            if (kMaskSelf) {                       // the reference skips j == i by index
                float slo, shi;
                unpack2(s, slo, shi);
                const int j0 = tile_slot0 + 2 * q, self = self_slot0 + k * kThreads;
                slo = (j0 == self) ? 0.0f : slo;
                shi = (j0 + 1 == self) ? 0.0f : shi;
                s = pack2(slo, shi);
            }
            ax[k] = fma2(s, dx, ax[k]);
            ay[k] = fma2(s, dy, ay[k]);
            az[k] = fma2(s, dz, az[k]);
        }
    }
}

struct ForceArgs {
    const float4 *pos;     // current global positions, n_slots
    float4 *partial;       // scratch, plan_partial_slots()
    int n_local;           // targets owned by this rank
    int i_slot_begin;      // slot of local target 0 (multiple of kSlotAlign)
    float eps2;            // SimParams::min_r2
    int mask_self;         // skip j == i by index (needed only when the softening is too small, see below)
    ForcePlan plan;
    const unsigned *mass_info;   // {0 if all masses are equal, bits of that mass}; NULL: treat as unequal
};

// With softening the self term needs no mask: x_j - x_i is exactly 0 for j == i, so the term is
// 0 * s with s = [m_j] (eps^2)^-3/2 finite -- it adds exactly nothing, as the reference's skip does.  The
// masked tile variant costs 25 % more issue slots, and the I tiles that hold a tile's own targets sit
// in one CTA's range, which made that CTA the last to finish in small systems.  Below this softening
// (s could overflow to inf, 0 * inf = NaN; min_r2 = 0 is legal in the reference) the mask is kept.
constexpr float kMaskSelfBelowEps2 = 1.0e-12f;

constexpr int kBufFloats = (kTileJ / 2) * 4;     // floats per buffer of sA / sB

// One segment: sources [src_begin, src_end) (multiples of kGran) against the tile of 256*I targets that
// starts at local target `target0`; leaves the partial sums in out[0 .. 256*I).
template <int I, int kLoop>
__device__ __forceinline__ void force_segment(const ForceArgs &a, float4 (*sA)[kTileJ / 2], float4 (*sB)[kTileJ / 2],
                                              float (*sOuter)[kThreads], int target0, int src_begin, int src_end,
                                              float4 *__restrict__ out) {
    const int tid = threadIdx.x;
    const u64 eps2 = pack2(a.eps2, a.eps2);
    // where this thread parks its source particle inside a packed tile
    float *const stA0 = reinterpret_cast<float *>(&sA[0][tid >> 1]) + (tid & 1);
    float *const stB0 = reinterpret_cast<float *>(&sB[0][tid >> 1]) + (tid & 1);

    const int islot0 = a.i_slot_begin + target0;        // first target slot of this tile
    u64 xi[I], yi[I], zi[I];
    const int self_slot0 = islot0 + tid;                // target k of this thread sits in slot self_slot0 + 256 k
#pragma unroll
    for (int k = 0; k < I; ++k) {
        const int l = target0 + k * kThreads + tid;     // local target index
        const int slot = (l < a.n_local) ? islot0 + k * kThreads + tid : islot0;
        const float4 p = __ldcg(a.pos + slot);          // .cg: see resident_steps_kernel
        xi[k] = pack2(p.x, p.x);
        yi[k] = pack2(p.y, p.y);
        zi[k] = pack2(p.z, p.z);
    }
#pragma unroll
    for (int k = 0; k < 3 * I; ++k) sOuter[k][tid] = 0.0f;   // private column: no barrier needed

    const int diag0 = islot0 / kTileJ;                  // source tiles [diag0, diag0+I) hold the targets
    // The run is a partial first tile (head), whole tiles, and a partial last tile (tail).  Head and
    // tail are peeled out of the tile loop so that the loop body keeps compile-time bounds.
    const int jt_first = src_begin / kTileJ, jt_last = (src_end - 1) / kTileJ;
    const int hq0 = (src_begin - jt_first * kTileJ) >> 1;            // first pair of the first tile
    const int tq1 = (src_end - jt_last * kTileJ) >> 1;               // end pair of the last tile
    // (measured: the general loop gains 1.5 % from the peeled form, the uniform-mass loop loses 3 % -- it keeps
    // head and tail inside the loop, as run-time bounds of the same code)
    constexpr bool kPeel = (kLoop == kLoopMassMul);
    const bool has_head = kPeel && ((hq0 != 0) || (jt_first == jt_last && tq1 != kTileJ / 2));
    const bool has_tail = kPeel && (jt_first != jt_last) && (tq1 != kTileJ / 2);
    const int jtA = jt_first + (has_head ? 1 : 0), jtB = jt_last + 1 - (has_tail ? 1 : 0);   // whole tiles [jtA, jtB)
    int buf = 0;
    {
        const float4 p = __ldcg(a.pos + jt_first * kTileJ + tid);
        stA0[0] = p.x; stA0[2] = p.y; stB0[0] = p.z; stB0[2] = p.w;
    }
    __syncthreads();
    u64 ax[I], ay[I], az[I];
#pragma unroll
    for (int k = 0; k < I; ++k) ax[k] = ay[k] = az[k] = 0ull;
    if (has_head) {
        float4 nxt = make_float4(0.f, 0.f, 0.f, 0.f);
        const bool more = (jt_first < jt_last);
        if (more) nxt = __ldcg(a.pos + (jt_first + 1) * kTileJ + tid);
        const int hq1 = (jt_first == jt_last) ? tq1 : kTileJ / 2;
        if (a.mask_self && jt_first >= diag0 && jt_first < diag0 + I)
            tile_interact<I, true, kLoop, false>(sA[0], sB[0], xi, yi, zi, eps2, ax, ay, az, hq0, hq1, jt_first * kTileJ, self_slot0);
        else
            tile_interact<I, false, kLoop, false>(sA[0], sB[0], xi, yi, zi, eps2, ax, ay, az, hq0, hq1, jt_first * kTileJ, self_slot0);
        if (more) {
            float *dA = stA0 + kBufFloats, *dB = stB0 + kBufFloats;
            dA[0] = nxt.x; dA[2] = nxt.y; dB[0] = nxt.z; dB[2] = nxt.w;
        }
        __syncthreads();
        buf = 1;
    }
    for (int jt = jtA; jt < jtB; ++jt) {
        float4 nxt = make_float4(0.f, 0.f, 0.f, 0.f);
        const bool more = (jt < jt_last);
        if (more) nxt = __ldcg(a.pos + (jt + 1) * kTileJ + tid);

        const bool diag = a.mask_self && (jt >= diag0 && jt < diag0 + I);
        const int q0 = (jt == jt_first) ? hq0 : 0, q1 = (jt == jt_last) ? tq1 : kTileJ / 2;     // whole tile when peeled
        if (diag) tile_interact<I, true, kLoop, kPeel>(sA[buf], sB[buf], xi, yi, zi, eps2, ax, ay, az, q0, q1, jt * kTileJ, self_slot0);
        else tile_interact<I, false, kLoop, kPeel>(sA[buf], sB[buf], xi, yi, zi, eps2, ax, ay, az, q0, q1, jt * kTileJ, self_slot0);
        // blocked summation: every kFlushTiles tiles (and at the end of the run) the inner
        // accumulators are folded into the outer sums and restarted: 2048 terms per inner lane,
        // N/4096 terms per outer sum -- neither level adds a long sequence
        if (((jt - jtA) & (kFlushTiles - 1)) == kFlushTiles - 1) {
#pragma unroll
            for (int k = 0; k < I; ++k) {
                float lo, hi;
                unpack2(ax[k], lo, hi); sOuter[3 * k + 0][tid] += lo + hi;
                unpack2(ay[k], lo, hi); sOuter[3 * k + 1][tid] += lo + hi;
                unpack2(az[k], lo, hi); sOuter[3 * k + 2][tid] += lo + hi;
                ax[k] = ay[k] = az[k] = 0ull;
            }
        }
        if (more) {
            float *dA = stA0 + (buf ^ 1) * kBufFloats, *dB = stB0 + (buf ^ 1) * kBufFloats;
            dA[0] = nxt.x; dA[2] = nxt.y; dB[0] = nxt.z; dB[2] = nxt.w;
        }
        __syncthreads();
        buf ^= 1;
    }
    if (has_tail) {
        if (a.mask_self && jt_last >= diag0 && jt_last < diag0 + I)
            tile_interact<I, true, kLoop, false>(sA[buf], sB[buf], xi, yi, zi, eps2, ax, ay, az, 0, tq1, jt_last * kTileJ, self_slot0);
        else
            tile_interact<I, false, kLoop, false>(sA[buf], sB[buf], xi, yi, zi, eps2, ax, ay, az, 0, tq1, jt_last * kTileJ, self_slot0);
        __syncthreads();                                 // the next segment restages buffer 0
    }
#pragma unroll
    for (int k = 0; k < I; ++k) {                        // what is left in the inner accumulators
        float lo, hi;
        unpack2(ax[k], lo, hi); sOuter[3 * k + 0][tid] += lo + hi;
        unpack2(ay[k], lo, hi); sOuter[3 * k + 1][tid] += lo + hi;
        unpack2(az[k], lo, hi); sOuter[3 * k + 2][tid] += lo + hi;
    }
#pragma unroll
    for (int k = 0; k < I; ++k)
        out[k * kThreads + tid] = make_float4(sOuter[3 * k][tid], sOuter[3 * k + 1][tid], sOuter[3 * k + 2][tid], 0.0f);
}

// Fast force kernel: persistent CTAs, I targets per thread in registers, source tiles staged
// through double-buffered shared memory in the packed-pair layout, blocked summation (one
// inner accumulator pair per tile, flushed into an outer FP32 sum per tile).
// One CTA per SM for I >= 4: its eight warps stay in step through the per-tile barrier, so no
// warp is left running alone at the end (two independent CTAs per SM drift apart), and with two
// warps per scheduler consecutive instructions of a warp issue back to back, which keeps the
// operand-reuse cache effective (the register file, not the FMA pipe, is the tighter resource
// for FFMA2 with three distinct register pairs).
// kLoop selects the loop at compile time (separate kernels keep each one's register allocation
// tight); a launch whose loop does not match the device-side mass flag returns at once, so the
// host can enqueue the uniform-mass and the general kernel without knowing the flag.
template <int I, int kLoop>
__device__ __forceinline__ void force_phase(const ForceArgs &a, float4 (*sA)[kTileJ / 2], float4 (*sB)[kTileJ / 2],
                                            float (*sOuter)[kThreads]) {
    const int c = blockIdx.x;
    long long g = plan_first_gran(a.plan, c);
    const long long g_end = plan_first_gran(a.plan, c + 1);
    const size_t TI = (size_t)kThreads * I;
    while (g < g_end) {
        const int t = (int)(g / a.plan.n_gran);
        const int s0 = (int)(g - (long long)t * a.plan.n_gran);
        const long long left = g_end - g;
        const int run = (a.plan.n_gran - s0 < left) ? a.plan.n_gran - s0 : (int)left;
        g += run;
        const int It = plan_tile_I(a.plan, t);
        const int target0 = plan_tile_target0(a.plan, t);
        float4 *out = a.partial + (size_t)(c + t) * TI;
        const int sb = s0 * kGran, se = (s0 + run) * kGran;
        if (It == I) force_segment<I, kLoop>(a, sA, sB, sOuter, target0, sb, se, out);
        else if (I > 4 && It == 4) force_segment<(I > 4 ? 4 : 1), kLoop>(a, sA, sB, sOuter, target0, sb, se, out);
        else if (I > 2 && It == 2) force_segment<(I > 2 ? 2 : 1), kLoop>(a, sA, sB, sOuter, target0, sb, se, out);
        else force_segment<1, kLoop>(a, sA, sB, sOuter, target0, sb, se, out);
    }
}

template <int I, int kLoop>
__global__ void __launch_bounds__(kThreads, (I >= 4) ? 1 : 2) force_fast_kernel(const ForceArgs a) {
    __shared__ float4 sA[2][kTileJ / 2];
    __shared__ float4 sB[2][kTileJ / 2];
    __shared__ float sOuter[3 * I][kThreads];   // second level of the blocked sum, one column per thread

    const bool unit_mass = (a.mass_info != nullptr) && (a.mass_info[0] == 0u);   // grid-uniform
    if (unit_mass != (kLoop == kLoopUnit)) return;
    force_phase<I, kLoop>(a, sA, sB, sOuter);
}

// ---------------------------------------------------------------------------------------------
// Strict force kernel: the reference's exact operation order and rounding.
// One thread per target, sources in ascending global index, j == i skipped by index,
// every operation a separately rounded IEEE binary32 op (no FMA contraction):
//   r = p_j - p_i; r2 = ((rx*rx + ry*ry) + rz*rz) + eps2; inv_r = 1/sqrt(r2);
//   inv_r3 = (inv_r*inv_r)*inv_r; c = G*inv_r3; acc += (r*c)*m_j
// (src/NaiveSimulation.cpp:28-38).  Bit-identical to NaiveSimulation by SURVEY.md F8.
// ---------------------------------------------------------------------------------------------
struct StrictArgs {
    const float4 *pos;
    float4 *acc;        // n_local
    int n_local;
    int i_slot_begin;
    int world;
    int slot_stride;
    int counts[kMaxWorld];   // valid slots per rank
    float G, eps2;
};

__global__ void __launch_bounds__(kThreads) force_strict_kernel(const StrictArgs a) {
    __shared__ float4 tile[kTileJ];
    const int tid = threadIdx.x;
    const int l = blockIdx.x * kThreads + tid;
    const bool live = l < a.n_local;
    const int my_slot = a.i_slot_begin + (live ? l : 0);
    const float4 pi = a.pos[my_slot];
    float ax = 0.0f, ay = 0.0f, az = 0.0f;
    for (int r = 0; r < a.world; ++r) {
        const int base = r * a.slot_stride;
        const int cnt = a.counts[r];
        for (int j0 = 0; j0 < cnt; j0 += kTileJ) {
            __syncthreads();
            tile[tid] = a.pos[base + j0 + tid];     // slots up to slot_stride always exist
            __syncthreads();
            const int m = (cnt - j0 < kTileJ) ? cnt - j0 : kTileJ;
            for (int jj = 0; jj < m; ++jj) {
                if (base + j0 + jj == my_slot) continue;
                const float4 pj = tile[jj];
                const float rx = __fsub_rn(pj.x, pi.x), ry = __fsub_rn(pj.y, pi.y), rz = __fsub_rn(pj.z, pi.z);
                const float r2 = __fadd_rn(
                    __fadd_rn(__fadd_rn(__fmul_rn(rx, rx), __fmul_rn(ry, ry)), __fmul_rn(rz, rz)), a.eps2);
                const float inv_r = __fdiv_rn(1.0f, __fsqrt_rn(r2));
                const float inv_r3 = __fmul_rn(__fmul_rn(inv_r, inv_r), inv_r);
                const float cc = __fmul_rn(a.G, inv_r3);
                ax = __fadd_rn(ax, __fmul_rn(__fmul_rn(rx, cc), pj.w));
                ay = __fadd_rn(ay, __fmul_rn(__fmul_rn(ry, cc), pj.w));
                az = __fadd_rn(az, __fmul_rn(__fmul_rn(rz, cc), pj.w));
            }
        }
    }
    if (live) a.acc[l] = make_float4(ax, ay, az, 0.0f);
}

// ---------------------------------------------------------------------------------------------
// FP64 force kernel (optional variant, reported separately: the reference has no double).
// Same formula evaluated in double from the FP32 positions and masses, sources in ascending slot
// order, self skipped by index; the sum is rounded to FP32 once, at the end, because the state the
// reference keeps (include/Particle.hpp:53-59) is FP32.  Two targets per thread; source tiles are
// converted to double once per CTA on their way into shared memory.
// ---------------------------------------------------------------------------------------------
// 1/sqrt(x) in double from an FP32 seed: MUFU.RSQ on the FP32 value of x (relative error 2^-22),
// one Newton step y <- y (1.5 - 0.5 x y^2) in double (error -> 1.5 e^2 ~ 5e-14, far below the final
// rounding to FP32).  x must be a positive normal FP32-range number, which r^2 + eps^2 of FP32
// positions is; x = 0 (coincident particles, no softening) gives inf like the reference's 1/sqrt(0).
__device__ __forceinline__ double rsqrt_f64_from_f32_seed(double x, double half_x) {
    const double y = (double)rsqrt_approx(__double2float_rn(x));
    const double t = fma(-half_x, y * y, 1.5);
    return y * t;
}

template <int kTargets>
__global__ void __launch_bounds__(kThreads) force_fp64_kernel(const StrictArgs a) {
    __shared__ double4 tile[kTileJ];
    const int tid = threadIdx.x;
    int slot[kTargets];
    bool live[kTargets];
    double xi[kTargets], yi[kTargets], zi[kTargets], ax[kTargets], ay[kTargets], az[kTargets];
#pragma unroll
    for (int k = 0; k < kTargets; ++k) {
        const int l = blockIdx.x * (kTargets * kThreads) + k * kThreads + tid;
        live[k] = l < a.n_local;
        slot[k] = a.i_slot_begin + (live[k] ? l : 0);
        const float4 p = a.pos[slot[k]];
        xi[k] = p.x; yi[k] = p.y; zi[k] = p.z;
        ax[k] = ay[k] = az[k] = 0.0;
    }
    const double eps2 = (double)a.eps2;
    const int n_slots = a.world * a.slot_stride;
    for (int j0 = 0; j0 < n_slots; j0 += kTileJ) {
        __syncthreads();
        {
            const float4 q = a.pos[j0 + tid];       // padding has mass 0: contributes exactly 0
            tile[tid] = make_double4((double)q.x, (double)q.y, (double)q.z, (double)q.w);
        }
        __syncthreads();
#pragma unroll 2
        for (int jj = 0; jj < kTileJ; ++jj) {
            const double4 q = tile[jj];
#pragma unroll
            for (int k = 0; k < kTargets; ++k) {
                const double dx = q.x - xi[k], dy = q.y - yi[k], dz = q.z - zi[k];
                const double r2 = fma(dx, dx, fma(dy, dy, fma(dz, dz, eps2)));
                const double ri = rsqrt_f64_from_f32_seed(r2, 0.5 * r2);
                double s = (q.w * ri) * (ri * ri);
                s = (j0 + jj == slot[k]) ? 0.0 : s;     // the reference skips j == i by index
                ax[k] = fma(dx, s, ax[k]); ay[k] = fma(dy, s, ay[k]); az[k] = fma(dz, s, az[k]);
            }
        }
    }
    const double G = (double)a.G;
#pragma unroll
    for (int k = 0; k < kTargets; ++k) {
        const int l = blockIdx.x * (kTargets * kThreads) + k * kThreads + tid;
        if (live[k]) a.acc[l] = make_float4((float)(G * ax[k]), (float)(G * ay[k]), (float)(G * az[k]), 0.0f);
    }
}

// ---------------------------------------------------------------------------------------------
// Finishing kernel: (fast mode) add the segment partials of each target in ascending segment
// order and scale by G; then the reference's semi-implicit Euler step, each product and sum
// rounded separately (src/NaiveSimulation.cpp:47-50); then publish the new {x,y,z,m} to the
// NEXT position array of this rank and of every attached peer (NVLink stores).
// ---------------------------------------------------------------------------------------------
struct FinishArgs {
    const float4 *pos_cur;
    float4 *pos_next;
    float4 *vel;
    float4 *acc;
    const float4 *partial;
    int n_local;
    int i_slot_begin;
    int TI;
    float G, dt;
    ForcePlan plan;
    int n_peers;
    float4 *peer_next[kMaxWorld];
    const unsigned *mass_info;   // as ForceArgs::mass_info
    // cross-GPU hand-over, fused into this kernel: the CTA that finishes last raises this rank's
    // step counter in every peer's flag block once all stores above are system-visible
    unsigned *done_counter;      // own block; 0 between launches
    unsigned *peer_flags[kMaxWorld];
    unsigned signal_seq;
    int my_rank;
};

// The reference's semi-implicit Euler step, each product and sum rounded separately
// (src/NaiveSimulation.cpp:47-50): v += a*dt; x += v*dt.
__device__ __forceinline__ void integrate_one(float4 &p, float4 &v, const float4 acc, float dt) {
    v.x = __fadd_rn(v.x, __fmul_rn(acc.x, dt));
    v.y = __fadd_rn(v.y, __fmul_rn(acc.y, dt));
    v.z = __fadd_rn(v.z, __fmul_rn(acc.z, dt));
    p.x = __fadd_rn(p.x, __fmul_rn(v.x, dt));
    p.y = __fadd_rn(p.y, __fmul_rn(v.y, dt));
    p.z = __fadd_rn(p.z, __fmul_rn(v.z, dt));
}

// Sum of the segment partials of local target l, in ascending segment (CTA) order.  Small systems cut
// a tile into dozens of segments (N = 4096: 37 per tile); reading them one dependent load at a time
// made this the longest kernel of their step (14 us against 11 for the force kernel), so the loads are
// issued sixteen at a time and added afterwards, in the same order as before.
__device__ __forceinline__ void sum_segments(const float4 *__restrict__ partial, const ForcePlan &plan, int TI, int l,
                                             float &sx, float &sy, float &sz) {
    int t, r;
    plan_tile_of_target(plan, l, t, r);
    const int c0 = plan_owner(plan, (long long)t * plan.n_gran);
    const int c1 = plan_owner(plan, (long long)(t + 1) * plan.n_gran - 1);
    const float4 *src = partial + (size_t)(c0 + t) * TI + r;
    sx = sy = sz = 0.0f;
    constexpr int kBatch = 16;
    for (int c = c0; c <= c1; c += kBatch) {
        float4 v[kBatch];
        const int last = c1 - c;                       // loads past the last segment re-read it (and are dropped)
#pragma unroll
        for (int k = 0; k < kBatch; ++k) v[k] = __ldcg(src + (size_t)min(k, last) * TI);
#pragma unroll
        for (int k = 0; k < kBatch; ++k) {
            const bool live = k <= last;
            sx += live ? v[k].x : 0.0f; sy += live ? v[k].y : 0.0f; sz += live ? v[k].z : 0.0f;
        }
        src += (size_t)kBatch * TI;
    }
}

template <bool kFromPartials>
__global__ void __launch_bounds__(kThreads) finish_kernel(const FinishArgs a) {
    const int l = blockIdx.x * kThreads + threadIdx.x;
    if (l < a.n_local) {
        float4 acc;
        if (kFromPartials) {
            float sx, sy, sz;
            sum_segments(a.partial, a.plan, a.TI, l, sx, sy, sz);
            float scale = a.G;
            if (a.mass_info != nullptr && a.mass_info[0] == 0u) scale = a.G * __uint_as_float(a.mass_info[1]);
            acc = make_float4(scale * sx, scale * sy, scale * sz, 0.0f);
            a.acc[l] = acc;
        } else {
            acc = a.acc[l];
        }
        const int slot = a.i_slot_begin + l;
        float4 p = a.pos_cur[slot];
        float4 v = a.vel[l];
        integrate_one(p, v, acc, a.dt);
        a.vel[l] = v;
        a.pos_next[slot] = p;
        for (int q = 0; q < a.n_peers; ++q) a.peer_next[q][slot] = p;      // NVLink stores into each peer's array
    }
    if (a.n_peers > 0) {
        // the barrier orders the CTA's peer stores before thread 0, whose system-scope fence is
        // cumulative over them (one fence per CTA, not per thread); it then takes a ticket, and
        // whoever draws the last ticket knows every CTA is past its fence and signals the peers
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence_system();
            const unsigned ticket = atomicAdd(a.done_counter, 1u);
            if (ticket == gridDim.x - 1) {
                atomicExch(a.done_counter, 0u);
                __threadfence_system();
                for (int q = 0; q < a.n_peers; ++q)
                    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.peer_flags[q] + a.my_rank), "r"(a.signal_seq) : "memory");
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Resident stepping kernel for small systems on one GPU: n_steps whole steps in ONE cooperative
// launch.  Below N ~ 30 000 a step is tens of microseconds of arithmetic, and two dependent kernels
// per step cost about 12 us in launch, drain and first-touch latency whatever the size (measured:
// N = 1024 takes 12.3 us per step under CUDA-graph replay).  Here the force phase and the integrator
// phase of every step are separated by grid-wide barriers instead of kernel boundaries; the
// arithmetic, the plan, the partial-sum layout and the order of every sum are those of
// force_fast_kernel + finish_kernel, so the results are bit-identical to the two-kernel path.
// Positions and partials written by other CTAs inside the launch are read with ld.global.cg (L2),
// never through L1.
// ---------------------------------------------------------------------------------------------
struct ResidentArgs {
    float4 *pos[2];        // double-buffered global positions
    float4 *vel, *acc, *partial;
    int n;                 // particles (single GPU: all targets are local, slot == index)
    int cur;               // which position array is current on entry
    int n_steps;
    int mask_self;
    float eps2, G, dt;
    ForcePlan plan;
    const unsigned *mass_info;
};

template <int I, int kLoop>
__global__ void __launch_bounds__(kThreads, (I >= 4) ? 1 : 2) resident_steps_kernel(const ResidentArgs a) {
    __shared__ float4 sA[2][kTileJ / 2];
    __shared__ float4 sB[2][kTileJ / 2];
    __shared__ float sOuter[3 * I][kThreads];

    const bool unit_mass = (a.mass_info != nullptr) && (a.mass_info[0] == 0u);   // grid-uniform: all CTAs leave together
    if (unit_mass != (kLoop == kLoopUnit)) return;
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    const float scale = unit_mass ? a.G * __uint_as_float(a.mass_info[1]) : a.G;
    const int TI = kThreads * I;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n_chunks = (a.n + 31) >> 5;                 // the integrator phase works in 32-target chunks,
    int cur = a.cur;                                      // chunk k on warp (k / grid) of CTA (k % grid): every SM takes part
    for (int s = 0; s < a.n_steps; ++s) {
        ForceArgs fa;
        fa.pos = a.pos[cur]; fa.partial = a.partial; fa.n_local = a.n; fa.i_slot_begin = 0; fa.eps2 = a.eps2;
        fa.mask_self = a.mask_self; fa.plan = a.plan; fa.mass_info = a.mass_info;
        force_phase<I, kLoop>(fa, sA, sB, sOuter);
        grid.sync();                                      // every partial sum is written and visible
        const float4 *pos_cur = a.pos[cur];
        float4 *pos_next = a.pos[cur ^ 1];
        for (int k = warp * (int)gridDim.x + (int)blockIdx.x; k < n_chunks; k += (kThreads / 32) * (int)gridDim.x) {
            const int l = k * 32 + lane;
            if (l < a.n) {
                float sx, sy, sz;
                sum_segments(a.partial, a.plan, TI, l, sx, sy, sz);
                const float4 acc = make_float4(scale * sx, scale * sy, scale * sz, 0.0f);
                float4 p = __ldcg(pos_cur + l);
                float4 v = __ldcg(a.vel + l);
                integrate_one(p, v, acc, a.dt);
                a.acc[l] = acc;
                a.vel[l] = v;
                pos_next[l] = p;
            }
        }
        grid.sync();                                      // the new positions are complete before anyone reads them
        cur ^= 1;
    }
}

// Decides whether all real particles carry the same mass (bitwise): info[0] = 1 if not,
// info[1] = bits of the mass of the first real particle.  info[0] must be 0 on entry.
struct MassCheckArgs {
    const float4 *pos;
    unsigned *info;
    int world, slot_stride, first_slot;
    int counts[kMaxWorld];
};
__global__ void __launch_bounds__(kThreads) mass_check_kernel(const MassCheckArgs a) {
    const int k = blockIdx.x * kThreads + threadIdx.x;
    const float ref = a.pos[a.first_slot].w;
    if (k == 0) a.info[1] = __float_as_uint(ref);
    if (k >= a.world * a.slot_stride) return;
    const int r = k / a.slot_stride, l = k - r * a.slot_stride;
    if (l < a.counts[r] && __float_as_uint(a.pos[k].w) != __float_as_uint(ref)) a.info[0] = 1u;
}

// Copies the local slice of a position array to the same slots of every peer (the exchange that
// follows an upload when world > 1).
struct PublishArgs {
    const float4 *src;
    int slot_begin, n;
    int n_peers;
    float4 *peer[kMaxWorld];
};
__global__ void __launch_bounds__(kThreads) publish_kernel(const PublishArgs a) {
    const int k = blockIdx.x * kThreads + threadIdx.x;
    if (k >= a.n) return;
    const float4 p = a.src[a.slot_begin + k];
    for (int q = 0; q < a.n_peers; ++q) a.peer[q][a.slot_begin + k] = p;
}

// Cross-GPU barrier: each rank raises flag[rank] = seq in every peer's flag block once its
// stores are system-visible, then waits until every peer has raised seq in its own block.
struct SignalArgs {
    unsigned seq;
    int my_rank;
    int n_peers;
    unsigned *peer_flags[kMaxWorld];
};
__global__ void signal_kernel(const SignalArgs a) {
    const int q = threadIdx.x;
    if (q >= a.n_peers) return;
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.peer_flags[q] + a.my_rank), "r"(a.seq) : "memory");
}
struct WaitArgs {
    unsigned seq;
    int my_rank;
    int world;
    const unsigned *flags;   // own block: flags[r] = last barrier rank r has reached
    unsigned *fault;         // device alias of a page-locked host word: set to 1 + r if rank r never arrived
    unsigned long long timeout_ns;
};
__global__ void wait_kernel(const WaitArgs a) {
    const int r = threadIdx.x;
    if (r >= a.world || r == a.my_rank) return;
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    unsigned v;
    do {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a.flags + r) : "memory");
        if ((int)(v - a.seq) >= 0) break;
        __nanosleep(200);
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 > a.timeout_ns) {          // a peer died or fell out of step: give up instead of hanging the GPU
            *a.fault = 1u + (unsigned)r;          // host-mapped word: the host sees it without a stream operation
            __threadfence_system();
            break;
        }
    } while (true);
}

// ---------------------------------------------------------------------------------------------
// Boundary transposes: the reference's 40-byte AoS record <-> the SoA arrays above.
// A CTA moves 256 records through shared memory so that global traffic on the AoS side is
// fully coalesced (2560 consecutive floats).
// ---------------------------------------------------------------------------------------------
constexpr int kRecFloats = 10;   // include/Particle.hpp:53-59: 3 + 3 + 3 + 1 floats

__global__ void __launch_bounds__(kThreads) aos_to_soa_kernel(const float *__restrict__ aos, int n, int slot_begin,
                                                              float4 *__restrict__ pos, float4 *__restrict__ vel,
                                                              float4 *__restrict__ acc) {
    __shared__ float s[kThreads * kRecFloats];
    const int first = blockIdx.x * kThreads;
    const int cnt = min(kThreads, n - first);
    const float *src = aos + (size_t)first * kRecFloats;
    for (int k = threadIdx.x; k < cnt * kRecFloats; k += kThreads) s[k] = src[k];
    __syncthreads();
    if (threadIdx.x < cnt) {
        const float *r = s + threadIdx.x * kRecFloats;
        const int l = first + threadIdx.x;
        pos[slot_begin + l] = make_float4(r[0], r[1], r[2], r[9]);
        vel[l] = make_float4(r[3], r[4], r[5], 0.0f);
        acc[l] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
}

__global__ void __launch_bounds__(kThreads) soa_to_aos_kernel(float *__restrict__ aos, int n, int slot_begin,
                                                              const float4 *__restrict__ pos,
                                                              const float4 *__restrict__ vel,
                                                              const float4 *__restrict__ acc) {
    __shared__ float s[kThreads * kRecFloats];
    const int first = blockIdx.x * kThreads;
    const int cnt = min(kThreads, n - first);
    if (threadIdx.x < cnt) {
        const int l = first + threadIdx.x;
        const float4 p = pos[slot_begin + l], v = vel[l], a = acc[l];
        float *r = s + threadIdx.x * kRecFloats;
        r[0] = p.x; r[1] = p.y; r[2] = p.z;
        r[3] = v.x; r[4] = v.y; r[5] = v.z;
        r[6] = a.x; r[7] = a.y; r[8] = a.z;
        r[9] = p.w;
    }
    __syncthreads();
    float *dst = aos + (size_t)first * kRecFloats;
    for (int k = threadIdx.x; k < cnt * kRecFloats; k += kThreads) dst[k] = s[k];
}

// Positions of the local slice as packed float[n][3] -- the record layout of the reference's
// trajectory file (src/CheckpointManager.cpp:86-99) -- staged through shared memory so the
// 12-byte side is written coalesced.
__global__ void __launch_bounds__(kThreads) pack_positions_kernel(float *__restrict__ xyz, int n, int slot_begin,
                                                                  const float4 *__restrict__ pos) {
    __shared__ float s[kThreads * 3];
    const int first = blockIdx.x * kThreads;
    const int cnt = min(kThreads, n - first);
    if (threadIdx.x < cnt) {
        const float4 p = pos[slot_begin + first + threadIdx.x];
        s[3 * threadIdx.x + 0] = p.x; s[3 * threadIdx.x + 1] = p.y; s[3 * threadIdx.x + 2] = p.z;
    }
    __syncthreads();
    float *dst = xyz + (size_t)first * 3;
    for (int k = threadIdx.x; k < cnt * 3; k += kThreads) dst[k] = s[k];
}

// Fills position slots [begin, end) with zero-mass padding.
__global__ void pad_kernel(float4 *pos, int begin, int end) {
    const int k = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (k < end) pos[k] = make_float4(kPadCoord, kPadCoord, kPadCoord, 0.0f);
}

// ---------------------------------------------------------------------------------------------
// Energy diagnostic, FP64 (no counterpart in the reference).  One thread per local target i;
// potential over sources with a larger slot (slot order == global particle order; padding has
// mass 0).  Each CTA leaves {K, U/G} in out[2*blockIdx.x .. +1]; the host adds them in order.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) energy_kernel(const float4 *__restrict__ pos, const float4 *__restrict__ vel,
                                                          int n_local, int i_slot_begin, int n_slots, double eps2,
                                                          double *__restrict__ out) {
    __shared__ float4 tile[kTileJ];
    __shared__ double redk[kThreads / 32], redu[kThreads / 32];
    const int tid = threadIdx.x;
    const int l = blockIdx.x * kThreads + tid;
    const bool live = l < n_local;
    const int my_slot = i_slot_begin + (live ? l : 0);
    const float4 pi = pos[my_slot];
    const double xi = pi.x, yi = pi.y, zi = pi.z;
    double u = 0.0;
    const int first_tile = (i_slot_begin + blockIdx.x * kThreads) / kTileJ;   // tiles below hold no j > i
    for (int jt = first_tile; jt < n_slots / kTileJ; ++jt) {
        __syncthreads();
        tile[tid] = pos[jt * kTileJ + tid];
        __syncthreads();
        for (int jj = 0; jj < kTileJ; ++jj) {
            const float4 pj = tile[jj];
            const double dx = (double)pj.x - xi, dy = (double)pj.y - yi, dz = (double)pj.z - zi;
            const double r2 = dx * dx + dy * dy + dz * dz + eps2;
            const double term = (double)pj.w * rsqrt_f64_from_f32_seed(r2, 0.5 * r2);   // ~5e-14 relative
            u += (jt * kTileJ + jj > my_slot) ? term : 0.0;   // select, so a self term of inf is dropped
        }
    }
    double k = 0.0;
    if (live) {
        const float4 v = vel[l];
        k = 0.5 * (double)pi.w * ((double)v.x * v.x + (double)v.y * v.y + (double)v.z * v.z);
        u = -(double)pi.w * u;
    } else {
        u = 0.0;
    }
    for (int o = 16; o > 0; o >>= 1) {
        k += __shfl_down_sync(0xffffffffu, k, o);
        u += __shfl_down_sync(0xffffffffu, u, o);
    }
    if ((tid & 31) == 0) { redk[tid >> 5] = k; redu[tid >> 5] = u; }
    __syncthreads();
    if (tid == 0) {
        double K = 0.0, U = 0.0;
        for (int w = 0; w < kThreads / 32; ++w) { K += redk[w]; U += redu[w]; }
        out[2 * blockIdx.x] = K;
        out[2 * blockIdx.x + 1] = U;
    }
}

// ---------------------------------------------------------------------------------------------
// Input generators on the device.  Same distributions as the reference's
// generatePlummerSphere (src/Simulation.cpp:173-267) and generateGalaxyDisk (:88-136), drawn from
// a counter-based Philox stream keyed by (seed, global particle index) instead of the reference's
// sequential mt19937 -- so the particles are NOT the reference's particles (parity tests use the
// reference's own generators), but every rank can produce any particle independently and
// identically.  Every rank fills ALL position slots; velocities only for its own slice.
// ---------------------------------------------------------------------------------------------
struct GenArgs {
    float4 *pos;            // current global positions
    float4 *vel;            // local velocities
    double *vsum;           // per-CTA partial sums {vx, vy, vz} of ALL particles (bulk-velocity removal)
    int world, slot_stride, rank;
    int counts[kMaxWorld], displs[kMaxWorld];
    int n_total;
    unsigned long long seed;
    float a, total_mass, G;             // Plummer: scale radius, total mass
    float radius, thickness;            // disk
};

__device__ __forceinline__ float uni(curandStatePhilox4_32_10_t &st, float lo, float hi) {
    return lo + (hi - lo) * (1.0f - curand_uniform(&st));        // [lo, hi), like uniform_real_distribution
}
__device__ __forceinline__ float3 on_sphere(curandStatePhilox4_32_10_t &st, float r) {
    float x, y, s;                                                // Marsaglia, include/Random.hpp:51-67
    do { x = uni(st, -1.f, 1.f); y = uni(st, -1.f, 1.f); s = x * x + y * y; } while (s >= 1.0f);
    const float t = 2.0f * sqrtf(1.0f - s);
    return make_float3(r * x * t, r * y * t, r * (1.0f - 2.0f * s));
}
__device__ __forceinline__ void slot_of_global(const GenArgs &a, int g, int &rank, int &local) {
    rank = 0;
    for (int r = a.world - 1; r >= 0; --r)
        if (a.counts[r] > 0 && g >= a.displs[r]) { rank = r; break; }
    local = g - a.displs[rank];
}

template <int kKind>    // 0: Plummer sphere, 1: galaxy disk
__global__ void __launch_bounds__(kThreads) generate_kernel(const GenArgs a) {
    __shared__ double red[3][kThreads / 32];
    const int g = blockIdx.x * kThreads + threadIdx.x;
    float3 v = make_float3(0.f, 0.f, 0.f);
    float m = 0.f;
    if (g < a.n_total) {
        curandStatePhilox4_32_10_t st;
        curand_init(a.seed, (unsigned long long)g, 0, &st);
        float3 p;
        if (kKind == 0) {
            m = a.total_mass / (float)a.n_total;
            const float u = fminf(uni(st, 0.001f, 1.0f), 0.999999f);                 // Simulation.cpp:216 (kept below 1: r stays finite)
            const float r = a.a / sqrtf(powf(u, -2.0f / 3.0f) - 1.0f);               // :217
            p = on_sphere(st, r);
            float q, y;
            do {                                                                     // :229-242
                q = uni(st, 0.f, 1.f);
                y = uni(st, 0.f, 0.1f);
            } while (!(y < q * q * powf(1.0f - q * q, 3.5f)));
            const float vesc = sqrtf(2.0f * a.G * a.total_mass / sqrtf(r * r + a.a * a.a));   // :245-246
            v = on_sphere(st, q * vesc);
        } else {
            if (g == 0) {                                                            // :92-98 central mass
                p = make_float3(0.f, 0.f, 0.f);
                m = 1.0e6f;
            } else {
                const float r = fmaxf(a.radius * sqrtf(uni(st, 0.f, 1.f)), 1e-3f);   // :103
                const float th = uni(st, 0.f, 6.283185307179586f);
                const float y = a.thickness * 0.3f * curand_normal(&st);
                p = make_float3(r * cosf(th), y, r * sinf(th));
                const float vo = sqrtf(a.G * 1.0e6f / r), disp = 0.1f * vo;          // :112-113
                v = make_float3(-vo * sinf(th) + disp * curand_normal(&st), 0.5f * disp * curand_normal(&st),
                                vo * cosf(th) + disp * curand_normal(&st));
                m = uni(st, 0.5f, 2.0f);                                             // :121
            }
        }
        int r, l;
        slot_of_global(a, g, r, l);
        a.pos[r * a.slot_stride + l] = make_float4(p.x, p.y, p.z, m);
        if (r == a.rank) a.vel[l] = make_float4(v.x, v.y, v.z, 0.f);
    }
    // mass-weighted velocity sum of all particles, per CTA, in a fixed order
    double sx = (double)m * v.x, sy = (double)m * v.y, sz = (double)m * v.z;
    for (int o = 16; o > 0; o >>= 1) {
        sx += __shfl_down_sync(0xffffffffu, sx, o);
        sy += __shfl_down_sync(0xffffffffu, sy, o);
        sz += __shfl_down_sync(0xffffffffu, sz, o);
    }
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = sx; red[1][threadIdx.x >> 5] = sy; red[2][threadIdx.x >> 5] = sz; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double tx = 0, ty = 0, tz = 0;
        for (int w = 0; w < kThreads / 32; ++w) { tx += red[0][w]; ty += red[1][w]; tz += red[2][w]; }
        a.vsum[3 * blockIdx.x] = tx; a.vsum[3 * blockIdx.x + 1] = ty; a.vsum[3 * blockIdx.x + 2] = tz;
    }
}

// Plummer only: subtract the bulk velocity (Simulation.cpp:256-265), total momentum / total mass.
__global__ void __launch_bounds__(kThreads) remove_bulk_velocity_kernel(float4 *vel, int n_local, const double *vsum, int n_blocks,
                                                                        double total_mass) {
    __shared__ float bulk[3];
    if (threadIdx.x == 0) {
        double tx = 0, ty = 0, tz = 0;
        for (int b = 0; b < n_blocks; ++b) { tx += vsum[3 * b]; ty += vsum[3 * b + 1]; tz += vsum[3 * b + 2]; }
        bulk[0] = (float)(tx / total_mass); bulk[1] = (float)(ty / total_mass); bulk[2] = (float)(tz / total_mass);
    }
    __syncthreads();
    const int l = blockIdx.x * kThreads + threadIdx.x;
    if (l >= n_local) return;
    float4 v = vel[l];
    v.x -= bulk[0]; v.y -= bulk[1]; v.z -= bulk[2];
    vel[l] = v;
}

// ---------------------------------------------------------------------------------------------
// Headless frame of the device positions: what the reference's GLUT viewer draws
// -- every particle an orange sphere of radius 2 under one directional light, camera pulled back
// along -z and turned about x then y, 45-degree perspective, near 1, far 5000
// (src/renderers/GlutRenderer.cpp:19-30,98-141,148) -- as plain compute: no graphics API.
// Pass 1 splats each sphere's disc into a 64-bit z-buffer with atomicMin on (depth bits, slot), which
// makes the nearest-surface choice independent of thread order; pass 2 shades the winner per pixel.
// ---------------------------------------------------------------------------------------------
struct Camera {
    float cx, sx, cy, sy;     // cos/sin of the x and y rotation angles
    float distance;           // translation along -z
    float fx, fy;             // focal lengths in pixels: (h/2)/tan(fovy/2) both (square pixels)
    float radius;             // sphere radius in world units
    float znear, zfar;
    int width, height;
};

__device__ __forceinline__ void to_eye(const Camera &c, float x, float y, float z, float &ex, float &ey, float &ez) {
    const float x1 = x * c.cy + z * c.sy;          // glRotatef(angle_y, 0,1,0)
    const float z1 = -x * c.sy + z * c.cy;
    ex = x1;
    ey = y * c.cx - z1 * c.sx;                      // glRotatef(angle_x, 1,0,0)
    ez = y * c.sx + z1 * c.cx - c.distance;         // glTranslatef(0,0,-distance); the eye looks down -z
}

constexpr unsigned long long kEmptyPixel = 0xffffffffffffffffull;
constexpr int kMaxSplatRadius = 256;    // pixels; larger discs are clamped (a sphere that close fills the view)

__global__ void __launch_bounds__(kThreads) splat_kernel(const float4 *__restrict__ pos, int n_slots, int slot_stride,
                                                         const int *__restrict__ counts, Camera c,
                                                         unsigned long long *__restrict__ zbuf) {
    const int k = blockIdx.x * kThreads + threadIdx.x;
    if (k >= n_slots) return;
    const int r = k / slot_stride;
    if (k - r * slot_stride >= counts[r]) return;            // padding slot
    const float4 p = pos[k];
    float ex, ey, ez;
    to_eye(c, p.x, p.y, p.z, ex, ey, ez);
    const float depth = -ez;
    if (!(depth > c.znear) || !(depth < c.zfar)) return;     // centre outside the clip range (also drops NaN)
    const float px = c.width * 0.5f + c.fx * ex / depth;
    const float py = c.height * 0.5f - c.fy * ey / depth;
    float rp = c.fy * c.radius / depth;
    rp = fminf(rp, (float)kMaxSplatRadius);
    const int x0 = max(0, (int)floorf(px - rp)), x1 = min(c.width - 1, (int)ceilf(px + rp));
    const int y0 = max(0, (int)floorf(py - rp)), y1 = min(c.height - 1, (int)ceilf(py + rp));
    const float inv = 1.0f / fmaxf(rp, 0.5f);                 // sub-pixel spheres still light their own pixel
    for (int y = y0; y <= y1; ++y)
        for (int x = x0; x <= x1; ++x) {
            const float u = ((x + 0.5f) - px) * inv, v = (py - (y + 0.5f)) * inv;
            const float q = 1.0f - u * u - v * v;
            const bool centre = (x == (int)floorf(px)) && (y == (int)floorf(py));   // a sphere never vanishes between pixel centres
            if (q < 0.0f && !centre) continue;
            const float zs = depth - c.radius * sqrtf(fmaxf(q, 0.0f));   // depth of the sphere surface under this pixel
            const unsigned long long key = ((unsigned long long)__float_as_uint(fmaxf(zs, 0.0f)) << 32) | (unsigned)k;
            atomicMin(&zbuf[(size_t)y * c.width + x], key);
        }
}

__global__ void __launch_bounds__(kThreads) shade_kernel(const float4 *__restrict__ pos, Camera c,
                                                         const unsigned long long *__restrict__ zbuf,
                                                         unsigned char *__restrict__ rgb) {
    const int i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= c.width * c.height) return;
    const unsigned long long key = zbuf[i];
    unsigned char R = 0, G = 0, B = 0;                        // glClearColor(0,0,0,1)
    if (key != kEmptyPixel) {
        const int k = (int)(key & 0xffffffffu);
        const int x = i % c.width, y = i / c.width;
        const float4 p = pos[k];
        float ex, ey, ez;
        to_eye(c, p.x, p.y, p.z, ex, ey, ez);
        const float depth = -ez;
        const float px = c.width * 0.5f + c.fx * ex / depth, py = c.height * 0.5f - c.fy * ey / depth;
        const float rp = fminf(c.fy * c.radius / depth, (float)kMaxSplatRadius);
        const float inv = 1.0f / fmaxf(rp, 0.5f);
        const float u = ((x + 0.5f) - px) * inv, v = (py - (y + 0.5f)) * inv;
        const float w = sqrtf(fmaxf(0.0f, 1.0f - u * u - v * v));          // eye-space normal (u, v, w)
        // GL_LIGHT0 at (1,1,1,0) set under an identity modelview: direction fixed in eye space;
        // ambient 0.3 + default scene ambient 0.2, diffuse 1, GL_COLOR_MATERIAL ambient-and-diffuse
        const float ndotl = fmaxf(0.0f, (u + v + w) * 0.57735026919f);
        const float lit = fminf(1.0f, 0.5f + ndotl);
        R = (unsigned char)(255.0f * lit + 0.5f);                            // glColor3f(1, 0.5, 0)
        G = (unsigned char)(255.0f * fminf(1.0f, 0.5f * lit) + 0.5f);
        B = 0;
    }
    rgb[3 * i + 0] = R; rgb[3 * i + 1] = G; rgb[3 * i + 2] = B;
}

// ---------------------------------------------------------------------------------------------
// Helpers of the allocation-free entry point (nbody_cuda_accel_noalloc): copy the caller's n
// sources into a tile-padded staging area, and fold the fast kernel's segment partials into
// accelerations.  Both only ever touch memory handed to them.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) stage_positions_kernel(const float4 *__restrict__ src, int n,
                                                                   float4 *__restrict__ dst, int n_slots) {
    const int k = blockIdx.x * kThreads + threadIdx.x;
    if (k >= n_slots) return;
    dst[k] = (k < n) ? src[k] : make_float4(kPadCoord, kPadCoord, kPadCoord, 0.0f);   // padding is rewritten every call
}

__global__ void __launch_bounds__(kThreads) reduce_partials_kernel(const float4 *__restrict__ partial, int n, int TI,
                                                                   float G, ForcePlan plan, float4 *__restrict__ acc) {
    const int l = blockIdx.x * kThreads + threadIdx.x;
    if (l >= n) return;
    float sx, sy, sz;
    sum_segments(partial, plan, TI, l, sx, sy, sz);      // same order as finish_kernel: ascending segment
    acc[l] = make_float4(G * sx, G * sy, G * sz, 0.0f);
}

// ---------------------------------------------------------------------------------------------
// FMA-pipe micro-benchmark: dependent-free FFMA (scalar) or FFMA2 (packed) chains, used by
// bench.py to state the measured FP32 peak next to the nominal #SM x 128 x 2 x f.
// ---------------------------------------------------------------------------------------------
template <bool kPacked>
__global__ void __launch_bounds__(kThreads) fma_peak_kernel(float *out, int iters, float seed) {
    if (kPacked) {
        u64 acc[8];
        // operands derived from threadIdx so they live in registers, not in the constant bank
        const float off = (float)(threadIdx.x >> 16);
        const u64 m = pack2(seed + off, seed * 0.5f + off), b = pack2(1e-3f + off, 2e-3f + off);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = pack2((float)threadIdx.x + k, (float)k);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int k = 0; k < 8; ++k) acc[k] = fma2(acc[k], m, b);
        }
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) { float lo, hi; unpack2(acc[k], lo, hi); s += lo + hi; }
        if (s == 123.456f) out[0] = s;
    } else {
        float acc[16];
        const float off = (float)(threadIdx.x >> 16);
        const float m = seed + off, b = 1e-3f + off;
#pragma unroll
        for (int k = 0; k < 16; ++k) acc[k] = (float)threadIdx.x + k;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int k = 0; k < 16; ++k) acc[k] = fmaf(acc[k], m, b);
        }
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < 16; ++k) s += acc[k];
        if (s == 123.456f) out[0] = s;
    }
}

}  // namespace nbody
