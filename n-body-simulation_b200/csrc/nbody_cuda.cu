// nbody_cuda.cu -- implementation of the C ABI in include/nbody_cuda.h.
//
// Host-side plumbing around the kernels of nbody_kernels.cuh: device buffers, the slot layout,
// launches, peer mapping for the NVLink exchange, timing.  Built for sm_100a only; no CPU
// fallback exists anywhere in this file -- without a device every entry point fails.
#include "nbody_cuda.h"

#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include <unistd.h>

#include <mutex>
#include <string>
#include <vector>

#include "nbody_kernels.cuh"

using namespace nbody;

namespace {

thread_local std::string g_create_error;

struct PeerToken {              // what nbody_cuda_peer_export hands out (<= NBODY_PEER_HANDLE_BYTES)
    unsigned magic;
    int rank, world, device;
    long long pid;
    unsigned long long n_slots;
    unsigned long long block_bytes;
    cudaIpcMemHandle_t ipc;     // 64 bytes
};
static_assert(sizeof(PeerToken) <= NBODY_PEER_HANDLE_BYTES, "token too large");
constexpr unsigned kTokenMagic = 0x4e424f44u;   // "NBOD"

}  // namespace

struct nbody_cuda {
    int device = 0;
    int mode = NBODY_MODE_FAST;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    bool timed = false;

    // configuration
    bool configured = false;
    size_t n_total = 0;
    int rank = 0, world = 1;
    size_t i_begin = 0, n_local = 0;
    size_t slot_stride = 0, n_slots = 0;
    int counts[kMaxWorld] = {0};

    // shared block: [pos0 | pos1 | flags]
    void *block = nullptr;
    size_t block_bytes = 0;
    float4 *pos[2] = {nullptr, nullptr};
    unsigned *flags = nullptr;
    int cur = 0;
    bool mass_checked = false;   // mass_info (flags[40..41]) describes the uploaded masses
    int mass_state = -1;         // host copy of the flag: -1 not read yet, 0 unequal masses, 1 all equal
    unsigned barrier_seq = 0;
    unsigned long long barrier_timeout_ns = 30ull * 1000000000ull;
    // barrier fault word in page-locked, device-mapped host memory: the waiting kernel writes it, the
    // host reads it without touching the stream (so a faulted run is refused before more work is queued)
    volatile unsigned *host_fault = nullptr;
    unsigned *dev_fault = nullptr;

    // local state
    float4 *vel = nullptr, *acc = nullptr;
    size_t local_cap = 0;
    float4 *partial = nullptr;
    size_t partial_cap = 0;     // in float4
    void *d_aos = nullptr;
    size_t aos_cap = 0;         // bytes
    double *d_energy = nullptr;
    size_t energy_cap = 0;      // doubles

    // asynchronous trajectory records: two staging buffers on the device, filled on the compute stream
    // and drained to the caller's page-locked memory on a copy stream of their own
    cudaStream_t copy_stream = nullptr;
    float *rec_stage[2] = {nullptr, nullptr};
    size_t rec_cap = 0;                              // floats per staging buffer
    cudaEvent_t rec_packed[2] = {nullptr, nullptr};  // staging k holds the record (compute stream)
    cudaEvent_t rec_copied[2] = {nullptr, nullptr};  // staging k has left for the host (copy stream)
    bool rec_busy[2] = {false, false};
    unsigned rec_count = 0;

    // fast-kernel plan
    int targets_per_thread = 4;
    ForcePlan plan{};
    int force_occupancy = 2;
    int forced_I = 0;                  // NBODY_OPT_TARGETS_PER_THREAD (0: chosen by the plan)
    int resident_mode = 1;             // NBODY_OPT_RESIDENT_KERNEL: 0 never, 1 for small systems, 2 whenever possible
    bool resident_fits = false;        // the plan's grid is co-resident for the cooperative launch

    // peers
    bool attached[kMaxWorld] = {false};
    void *peer_block[kMaxWorld] = {nullptr};
    bool peer_is_ipc[kMaxWorld] = {false};
    int n_attached = 0;

    // CUDA graph of two consecutive steps (single GPU): replayed for long device-resident runs so
    // that small systems are not bound by launch latency
    cudaGraphExec_t step_graph = nullptr;
    float graph_G = 0.f, graph_dt = 0.f, graph_eps2 = 0.f;
    int graph_parity = 0, graph_mass_state = -1;
    long long graph_launches = 0;   // kernels in one replay
    bool capturing = false;

    // profiling
    bool profile_force = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> force_events;
    size_t force_events_used = 0;
    double force_ms_total = 0.0;
    int force_launches = 0;
    long long launches = 0;

    std::string err;
};

namespace {

int fail(nbody_cuda_t h, int code, const std::string &msg) {
    if (h) h->err = msg; else g_create_error = msg;
    return code;
}

#define CU_TRY(h, call)                                                                          \
    do {                                                                                         \
        cudaError_t e__ = (call);                                                                \
        if (e__ != cudaSuccess) {                                                                \
            cudaGetLastError();                                                                  \
            return fail((h), (e__ == cudaErrorMemoryAllocation) ? NBODY_E_NOMEM : NBODY_E_CUDA,  \
                        std::string(#call) + ": " + cudaGetErrorString(e__));                    \
        }                                                                                        \
    } while (0)

#define CHECK_LAUNCH(h, name)                                                                    \
    do {                                                                                         \
        cudaError_t e__ = cudaGetLastError();                                                    \
        if (e__ != cudaSuccess)                                                                  \
            return fail((h), NBODY_E_CUDA, std::string("launch of ") + name + ": " + cudaGetErrorString(e__)); \
        if ((h)->capturing) (h)->graph_launches++; else (h)->launches++;                         \
    } while (0)

size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// words of the 64-word flag block that follows the position arrays
// flags[0..15]: barrier counters (the barrier fault word lives in host-mapped memory, see nbody_cuda::host_fault)
constexpr int kMassInfoWord = 40; // flags[40]: masses differ, flags[41]: bits of the common mass
constexpr int kDoneWord = 48;     // flags[48]: CTA ticket counter of the finishing kernel

void slice_of(size_t n, int world, int r, size_t *begin, size_t *count) {
    // counts[r] = n/P + (r < n%P), displs = prefix sum: src/apps/headless_main.cpp:249-259
    const size_t base = n / (size_t)world, rem = n % (size_t)world;
    *begin = (size_t)r * base + ((size_t)r < rem ? (size_t)r : rem);
    *count = base + ((size_t)r < rem ? 1 : 0);
}

void drop_peers(nbody_cuda_t h) {
    for (int r = 0; r < kMaxWorld; ++r) {
        if (h->attached[r] && h->peer_is_ipc[r] && h->peer_block[r]) cudaIpcCloseMemHandle(h->peer_block[r]);
        h->attached[r] = false;
        h->peer_block[r] = nullptr;
        h->peer_is_ipc[r] = false;
    }
    h->n_attached = 0;
}

void drop_graph(nbody_cuda_t h) {
    if (h->step_graph) cudaGraphExecDestroy(h->step_graph);
    h->step_graph = nullptr;
}

void free_buffers(nbody_cuda_t h) {
    drop_peers(h);
    drop_graph(h);
    if (h->block) cudaFree(h->block);
    if (h->vel) cudaFree(h->vel);
    if (h->acc) cudaFree(h->acc);
    if (h->partial) cudaFree(h->partial);
    if (h->d_aos) cudaFree(h->d_aos);
    if (h->d_energy) cudaFree(h->d_energy);
    for (int k = 0; k < 2; ++k) {
        if (h->rec_stage[k]) cudaFree(h->rec_stage[k]);
        h->rec_stage[k] = nullptr; h->rec_busy[k] = false;
    }
    h->rec_cap = 0;
    h->block = nullptr; h->vel = h->acc = h->partial = nullptr; h->d_aos = nullptr; h->d_energy = nullptr;
    h->pos[0] = h->pos[1] = nullptr; h->flags = nullptr;
    h->block_bytes = h->local_cap = h->partial_cap = h->aos_cap = h->energy_cap = 0;
    h->configured = false;
}

// Targets-per-thread variants of the fast kernel (main tiles); narrower tail tiles come from
// {4, 2, 1} below the main width.  More targets per thread = fewer shared-memory reads per
// interaction and better operand reuse, until registers run out (I = 7: 200-odd registers).
struct Variant { int I; int ctas_per_sm; };
const Variant kVariants[] = {{7, 1}, {4, 1}, {2, 2}, {1, 2}};
constexpr int kNumVariants = (int)(sizeof(kVariants) / sizeof(kVariants[0]));

ForcePlan plan_for(size_t n_local, size_t n_slots, int sm_count, const Variant &v) {
    const size_t TI = (size_t)kThreads * v.I;
    ForcePlan p;
    memset(&p, 0, sizeof(p));
    p.I_main = v.I;
    p.n_gran = (int)(n_slots / kGran);
    p.n_main = (int)(n_local / TI);
    int blocks = (int)((n_local - (size_t)p.n_main * TI + kThreads - 1) / kThreads);    // 256-target blocks left over
    if (blocks >= v.I) { p.n_main++; blocks = 0; }
    for (int w = 4; w >= 1 && p.n_tail < kMaxTail; w >>= 1)       // blocks < I <= 7: at most two of {4,2,1}
        if (w < v.I && blocks >= w) { p.tail_I[p.n_tail++] = w; blocks -= w; }
    if (blocks > 0) {                                            // not representable (cannot happen for I in {7,4,2,1})
        p.n_main++; p.n_tail = 0; p.tail_I[0] = p.tail_I[1] = 0;
    }
    p.cost_main = plan_granule_cost(v.I);
    int max_cost = p.n_main > 0 ? p.cost_main : 0;
    p.cost_total = (long long)p.n_main * p.n_gran * p.cost_main;
    for (int k = 0; k < p.n_tail; ++k) {
        p.cost_tail[k] = plan_granule_cost(p.tail_I[k]);
        p.cost_total += (long long)p.n_gran * p.cost_tail[k];
        if (p.cost_tail[k] > max_cost) max_cost = p.cost_tail[k];
    }
    long long g = (long long)sm_count * v.ctas_per_sm;
    if (max_cost > 0 && g > p.cost_total / max_cost) g = p.cost_total / max_cost;   // every CTA gets at least one granule
    if (g < 1) g = 1;
    p.grid = (int)g;
    return p;
}

// Estimated duration of a plan in granule-cost units: the busiest CTA's share (plus one granule of
// imbalance) and what each of its segments costs besides the interactions -- target loads and the
// first source tile at L2 latency (~2000 cycles), clearing and writing the partial sums.
double plan_estimate(const ForcePlan &p, int sm_count) {
    const double share = (double)p.cost_total / p.grid;
    const int max_cost = p.n_main > 0 ? p.cost_main : p.cost_tail[0];
    const double per_tile = (double)p.n_gran * max_cost;
    const double segments = 1.0 + share / per_tile;
    const int resident = (p.grid + sm_count - 1) / sm_count;     // CTAs sharing an SM split its FMA pipe
    return (share + max_cost + segments * (10.0 + 14.0 * p.I_main)) * resident;
}

int pick_variant(size_t n_local, size_t n_slots, int sm_count, int forced_I, ForcePlan *plan) {
    double best = 0.0;
    int best_k = -1;
    for (int k = 0; k < kNumVariants; ++k) {
        if (forced_I > 0 && kVariants[k].I != forced_I) continue;
        const ForcePlan p = plan_for(n_local, n_slots, sm_count, kVariants[k]);
        const double t = plan_estimate(p, sm_count);
        if (best_k < 0 || t < best) { best = t; best_k = k; }
    }
    if (best_k < 0) return pick_variant(n_local, n_slots, sm_count, 0, plan);     // unknown forced width: choose freely
    *plan = plan_for(n_local, n_slots, sm_count, kVariants[best_k]);
    return best_k;
}

bool resident_grid_fits(nbody_cuda_t h);

// Largest system stepped by the resident (one cooperative launch for n_steps) kernel by default: above
// this the grid barriers cost as much as the kernel boundaries they replace (measured cross-over near N = 15 000).
constexpr size_t kResidentMaxN = 12288;

// Chooses the variant with the smallest estimated time and sizes the partial-sum scratch.
int make_plan(nbody_cuda_t h) {
    // Systems that the resident kernel steps (one GPU, N <= kResidentMaxN) measured faster with 1024-target
    // tiles than with 1792-target ones from N ~ 6000 up (N = 8192: 29.9 vs 32.9 us per step, N = 10 000:
    // 43.0 vs 47.2; profiles/r2_small_n_sweep.log): less to load, clear and write per segment, and every
    // CTA is in several segments there.  The cost estimate does not see that, so the width is capped here.
    int width = h->forced_I;
    if (width == 0 && h->world == 1 && h->n_total <= kResidentMaxN) width = (h->n_total > 3000) ? 4 : 0;
    int k = pick_variant(h->n_local, h->n_slots, h->sm_count, width, &h->plan);
    if (width == 4 && h->forced_I == 0) {                   // ... but never against a narrower choice of the estimate
        ForcePlan free_plan;
        const int kf = pick_variant(h->n_local, h->n_slots, h->sm_count, 0, &free_plan);
        if (kVariants[kf].I < 4) { k = kf; h->plan = free_plan; }
    }
    h->targets_per_thread = kVariants[k].I;
    h->force_occupancy = kVariants[k].ctas_per_sm;
    h->resident_fits = (h->world == 1) && resident_grid_fits(h);
    const size_t need = plan_partial_slots(h->plan);
    if (need > h->partial_cap) {
        if (h->partial) cudaFree(h->partial);
        h->partial = nullptr; h->partial_cap = 0;
        CU_TRY(h, cudaMalloc(&h->partial, need * sizeof(float4)));
        h->partial_cap = need;
    }
    return NBODY_OK;
}

int ensure_aos(nbody_cuda_t h, size_t bytes) {
    if (bytes <= h->aos_cap) return NBODY_OK;
    if (h->d_aos) cudaFree(h->d_aos);
    h->d_aos = nullptr; h->aos_cap = 0;
    CU_TRY(h, cudaMalloc(&h->d_aos, bytes));
    h->aos_cap = bytes;
    return NBODY_OK;
}

template <int kLoop>
void launch_fast_variant(int I, int grid, cudaStream_t st, const ForceArgs &a) {
    switch (I) {
        case 7: force_fast_kernel<7, kLoop><<<grid, kThreads, 0, st>>>(a); break;
        case 4: force_fast_kernel<4, kLoop><<<grid, kThreads, 0, st>>>(a); break;
        case 2: force_fast_kernel<2, kLoop><<<grid, kThreads, 0, st>>>(a); break;
        default: force_fast_kernel<1, kLoop><<<grid, kThreads, 0, st>>>(a); break;
    }
}

template <int I, int kLoop>
cudaError_t launch_resident_one(int grid, cudaStream_t st, ResidentArgs &a) {
    void *params[] = {(void *)&a};
    return cudaLaunchCooperativeKernel((const void *)resident_steps_kernel<I, kLoop>, dim3(grid), dim3(kThreads), params, 0, st);
}
template <int kLoop>
cudaError_t launch_resident_variant(int I, int grid, cudaStream_t st, ResidentArgs &a) {
    switch (I) {
        case 7: return launch_resident_one<7, kLoop>(grid, st, a);
        case 4: return launch_resident_one<4, kLoop>(grid, st, a);
        case 2: return launch_resident_one<2, kLoop>(grid, st, a);
        default: return launch_resident_one<1, kLoop>(grid, st, a);
    }
}
template <int I>
int resident_blocks_per_sm() {
    int a = 0, b = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, resident_steps_kernel<I, kLoopUnit>, kThreads, 0) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, resident_steps_kernel<I, kLoopMassMul>, kThreads, 0) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return a < b ? a : b;
}
// Can every CTA of the plan be resident at once (the grid barrier of the resident kernel needs it)?
bool resident_grid_fits(nbody_cuda_t h) {
    int coop = 0;
    if (cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, h->device) != cudaSuccess || !coop) { cudaGetLastError(); return false; }
    int per_sm = 0;
    switch (h->targets_per_thread) {
        case 7: per_sm = resident_blocks_per_sm<7>(); break;
        case 4: per_sm = resident_blocks_per_sm<4>(); break;
        case 2: per_sm = resident_blocks_per_sm<2>(); break;
        default: per_sm = resident_blocks_per_sm<1>(); break;
    }
    return (long long)per_sm * h->sm_count >= h->plan.grid;
}

// Masses never change on the device, so one look after each upload (and after the exchange that
// completes the replicated array) tells the fast kernel whether it may drop m_j from its loop.
int ensure_mass_info(nbody_cuda_t h) {
    if (h->mass_checked || h->mode != NBODY_MODE_FAST || h->n_total == 0) return NBODY_OK;
    unsigned *info = h->flags + kMassInfoWord;
    CU_TRY(h, cudaMemsetAsync(info, 0, 2 * sizeof(unsigned), h->stream));
    MassCheckArgs a;
    a.pos = h->pos[h->cur]; a.info = info; a.world = h->world; a.slot_stride = (int)h->slot_stride;
    a.first_slot = 0;
    for (int r = 0; r < kMaxWorld; ++r) a.counts[r] = h->counts[r];
    for (int r = 0; r < h->world; ++r)
        if (h->counts[r] > 0) { a.first_slot = (int)(r * h->slot_stride); break; }
    mass_check_kernel<<<(int)((h->n_slots + kThreads - 1) / kThreads), kThreads, 0, h->stream>>>(a);
    CHECK_LAUNCH(h, "mass_check_kernel");
    h->mass_checked = true;
    return NBODY_OK;
}

int launch_force(nbody_cuda_t h, float G, float min_r2) {
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (h->profile_force) {
        if (h->force_events_used == h->force_events.size()) {
            cudaEvent_t a, b;
            CU_TRY(h, cudaEventCreate(&a));
            CU_TRY(h, cudaEventCreate(&b));
            h->force_events.emplace_back(a, b);
        }
        e0 = h->force_events[h->force_events_used].first;
        e1 = h->force_events[h->force_events_used].second;
        h->force_events_used++;
        CU_TRY(h, cudaEventRecord(e0, h->stream));
    }
    if (h->mode == NBODY_MODE_STRICT || h->mode == NBODY_MODE_FP64) {
        StrictArgs a;
        a.pos = h->pos[h->cur]; a.acc = h->acc;
        a.n_local = (int)h->n_local; a.i_slot_begin = (int)(h->rank * h->slot_stride);
        a.world = h->world; a.slot_stride = (int)h->slot_stride;
        for (int r = 0; r < kMaxWorld; ++r) a.counts[r] = h->counts[r];
        a.G = G; a.eps2 = min_r2;
        if (h->mode == NBODY_MODE_FP64) {
            constexpr int kT = 2;                      // targets per thread
            const int grid = (int)((h->n_local + kT * kThreads - 1) / (kT * kThreads));
            force_fp64_kernel<kT><<<grid, kThreads, 0, h->stream>>>(a);
            CHECK_LAUNCH(h, "force_fp64_kernel");
        } else {
            const int grid = (int)((h->n_local + kThreads - 1) / kThreads);
            force_strict_kernel<<<grid, kThreads, 0, h->stream>>>(a);
            CHECK_LAUNCH(h, "force_strict_kernel");
        }
    } else {
        ForceArgs a;
        a.pos = h->pos[h->cur]; a.partial = h->partial;
        a.n_local = (int)h->n_local; a.i_slot_begin = (int)(h->rank * h->slot_stride);
        a.eps2 = min_r2; a.plan = h->plan;
        a.mask_self = !(min_r2 >= kMaskSelfBelowEps2);          // also masks for NaN / negative softening
        a.mass_info = h->flags + kMassInfoWord;
        // until the host has seen the mass flag (first synchronising call after an upload) both loops
        // are enqueued and the one that does not apply exits immediately
        if (h->mass_state != 0) {
            launch_fast_variant<kLoopUnit>(h->targets_per_thread, h->plan.grid, h->stream, a);
            CHECK_LAUNCH(h, "force_fast_kernel<uniform mass>");
        }
        if (h->mass_state != 1) {
            launch_fast_variant<kLoopMassMul>(h->targets_per_thread, h->plan.grid, h->stream, a);
            CHECK_LAUNCH(h, "force_fast_kernel<general>");
        }
    }
    if (h->profile_force) CU_TRY(h, cudaEventRecord(e1, h->stream));
    return NBODY_OK;
}

void fill_peer_ptrs(nbody_cuda_t h, int which_pos, float4 **out, int *n) {
    int k = 0;
    for (int r = 0; r < h->world; ++r) {
        if (r == h->rank || !h->attached[r]) continue;
        out[k++] = reinterpret_cast<float4 *>(h->peer_block[r]) + (size_t)which_pos * h->n_slots;
    }
    *n = k;
}


// After a stream synchronise: pick up the mass flag so that later steps enqueue only the loop
// that applies.
void learn_mass_state(nbody_cuda_t h) {
    if (h->mode != NBODY_MODE_FAST || !h->mass_checked || h->mass_state >= 0 || !h->flags) return;
    unsigned differ = 0;
    if (cudaMemcpy(&differ, h->flags + kMassInfoWord, sizeof(differ), cudaMemcpyDeviceToHost) == cudaSuccess)
        h->mass_state = differ ? 0 : 1;
    else
        cudaGetLastError();
}

// Reports a barrier that gave up (a peer never arrived).  Reads host memory only.
int report_fault(nbody_cuda_t h) {
    const unsigned f = h->host_fault ? *h->host_fault : 0u;
    if (f != 0)
        return fail(h, NBODY_E_CUDA, "peer exchange timed out waiting for rank " + std::to_string((int)f - 1) +
                                         " (ranks out of step, or a peer process died); positions from that step on are "
                                         "stale -- upload again (or reconfigure) to clear the fault");
    return NBODY_OK;
}

// Called after a stream synchronise.
int check_fault(nbody_cuda_t h) {
    learn_mass_state(h);
    return report_fault(h);
}

unsigned *peer_flags_of(nbody_cuda_t h, int r) {
    return reinterpret_cast<unsigned *>(reinterpret_cast<float4 *>(h->peer_block[r]) + 2 * h->n_slots);
}

// signal every peer (unless the finishing kernel of this step already did), then wait for every
// peer (device-side; nothing blocks the host)
int launch_barrier(nbody_cuda_t h, bool already_signalled) {
    if (!already_signalled) {
        h->barrier_seq++;
        SignalArgs s;
        s.seq = h->barrier_seq; s.my_rank = h->rank;
        int k = 0;
        for (int r = 0; r < h->world; ++r)
            if (r != h->rank) s.peer_flags[k++] = peer_flags_of(h, r);
        s.n_peers = k;
        signal_kernel<<<1, 32, 0, h->stream>>>(s);
        CHECK_LAUNCH(h, "signal_kernel");
    }
    WaitArgs w;
    w.seq = h->barrier_seq; w.my_rank = h->rank; w.world = h->world; w.flags = h->flags;
    w.fault = h->dev_fault; w.timeout_ns = h->barrier_timeout_ns;
    wait_kernel<<<1, 32, 0, h->stream>>>(w);
    CHECK_LAUNCH(h, "wait_kernel");
    return NBODY_OK;
}

int launch_finish(nbody_cuda_t h, float G, float dt, bool publish) {
    FinishArgs a;
    a.pos_cur = h->pos[h->cur]; a.pos_next = h->pos[h->cur ^ 1];
    a.vel = h->vel; a.acc = h->acc; a.partial = h->partial;
    a.n_local = (int)h->n_local; a.i_slot_begin = (int)(h->rank * h->slot_stride);
    a.TI = kThreads * h->targets_per_thread;
    a.G = G; a.dt = dt; a.plan = h->plan;
    a.n_peers = 0;
    a.mass_info = h->flags + kMassInfoWord;
    a.done_counter = h->flags + kDoneWord; a.signal_seq = 0; a.my_rank = h->rank;
    if (publish) fill_peer_ptrs(h, h->cur ^ 1, a.peer_next, &a.n_peers);
    if (a.n_peers > 0) {                      // this launch also raises our step counter at the peers
        a.signal_seq = ++h->barrier_seq;
        int k = 0;
        for (int r = 0; r < h->world; ++r)
            if (r != h->rank && h->attached[r]) a.peer_flags[k++] = peer_flags_of(h, r);
    }
    const int grid = (int)((h->n_local + kThreads - 1) / kThreads);
    if (h->mode != NBODY_MODE_FAST) finish_kernel<false><<<grid, kThreads, 0, h->stream>>>(a);
    else finish_kernel<true><<<grid, kThreads, 0, h->stream>>>(a);
    CHECK_LAUNCH(h, "finish_kernel");
    return NBODY_OK;
}

int one_step(nbody_cuda_t h, float G, float dt, float min_r2, bool exchange) {
    int rc;
    if ((rc = ensure_mass_info(h)) != NBODY_OK) return rc;
    bool signalled = false;
    if (h->n_local > 0) {
        if ((rc = launch_force(h, G, min_r2)) != NBODY_OK) return rc;
        if ((rc = launch_finish(h, G, dt, exchange)) != NBODY_OK) return rc;
        signalled = exchange && h->world > 1;       // the finishing kernel published and signalled
    }
    if (exchange && h->world > 1)
        if ((rc = launch_barrier(h, signalled)) != NBODY_OK) return rc;
    h->cur ^= 1;
    return NBODY_OK;
}

int all_peers_attached(nbody_cuda_t h) { return h->n_attached == h->world - 1; }

}  // namespace

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

size_t nbody_cuda_particle_bytes(void) { return kRecFloats * sizeof(float); }

const char *nbody_cuda_version(void) { return "nbody_cuda 0.1 sm_100a"; }

int nbody_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int nbody_cuda_create(nbody_cuda_t *out, int device, int mode) {
    if (!out) return fail(nullptr, NBODY_E_INVALID, "nbody_cuda_create: out is NULL");
    *out = nullptr;
    if (mode != NBODY_MODE_FAST && mode != NBODY_MODE_STRICT && mode != NBODY_MODE_FP64)
        return fail(nullptr, NBODY_E_INVALID, "nbody_cuda_create: unknown mode");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(nullptr, NBODY_E_CUDA,
                    std::string("nbody_cuda_create: no usable CUDA device (") +
                        (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                        "); this library has no CPU fallback");
    }
    if (device < 0 || device >= n) return fail(nullptr, NBODY_E_INVALID, "nbody_cuda_create: device ordinal out of range");
    nbody_cuda_t h = new nbody_cuda();
    h->device = device;
    h->mode = mode;
    if ((e = cudaSetDevice(device)) != cudaSuccess ||
        (e = cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaEventCreate(&h->ev_begin)) != cudaSuccess || (e = cudaEventCreate(&h->ev_end)) != cudaSuccess ||
        (e = cudaHostAlloc((void **)&h->host_fault, 64, cudaHostAllocMapped)) != cudaSuccess ||
        (e = cudaHostGetDevicePointer((void **)&h->dev_fault, (void *)h->host_fault, 0)) != cudaSuccess) {
        std::string msg = std::string("nbody_cuda_create: ") + cudaGetErrorString(e);
        cudaGetLastError();
        if (h->host_fault) cudaFreeHost((void *)h->host_fault);
        delete h;
        return fail(nullptr, NBODY_E_CUDA, msg);
    }
    *h->host_fault = 0u;
    *out = h;
    return NBODY_OK;
}

void nbody_cuda_destroy(nbody_cuda_t h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->copy_stream) cudaStreamSynchronize(h->copy_stream);
    free_buffers(h);
    for (auto &p : h->force_events) { cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
    if (h->ev_begin) cudaEventDestroy(h->ev_begin);
    if (h->ev_end) cudaEventDestroy(h->ev_end);
    for (int k = 0; k < 2; ++k) {
        if (h->rec_packed[k]) cudaEventDestroy(h->rec_packed[k]);
        if (h->rec_copied[k]) cudaEventDestroy(h->rec_copied[k]);
    }
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->host_fault) cudaFreeHost((void *)h->host_fault);
    cudaGetLastError();
    delete h;
}

const char *nbody_cuda_last_error(nbody_cuda_t h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int nbody_cuda_configure(nbody_cuda_t h, size_t n_total, int rank, int world) {
    if (!h) return NBODY_E_INVALID;
    if (world < 1 || world > kMaxWorld || rank < 0 || rank >= world)
        return fail(h, NBODY_E_INVALID, "nbody_cuda_configure: need 0 <= rank < world <= 16");
    if (n_total > ((size_t)1 << 30) - kSlotAlign) return fail(h, NBODY_E_INVALID, "nbody_cuda_configure: n_total above 2^30 - 256");
    CU_TRY(h, cudaSetDevice(h->device));
    if (h->configured && n_total == h->n_total && rank == h->rank && world == h->world) return NBODY_OK;

    CU_TRY(h, cudaStreamSynchronize(h->stream));
    if (h->copy_stream) CU_TRY(h, cudaStreamSynchronize(h->copy_stream));
    h->rec_busy[0] = h->rec_busy[1] = false;
    drop_peers(h);          // tokens describe the old block
    drop_graph(h);          // so does a captured step
    size_t begin, count;
    slice_of(n_total, world, rank, &begin, &count);
    size_t max_count = n_total / world + (n_total % world ? 1 : 0);
    const size_t stride = round_up(max_count > 0 ? max_count : 1, kSlotAlign);
    const size_t n_slots = stride * world;
    // 2 * n_slots (both position arrays) and every slot index must fit a 32-bit int
    if (n_slots > ((size_t)1 << 30) - kSlotAlign) return fail(h, NBODY_E_INVALID, "nbody_cuda_configure: too many slots (limit 2^30 - 256 including padding)");
    const size_t block_bytes = 2 * n_slots * sizeof(float4) + 256;

    if (block_bytes > h->block_bytes || n_slots != h->n_slots) {
        if (h->block) cudaFree(h->block);
        h->block = nullptr; h->block_bytes = 0;
        CU_TRY(h, cudaMalloc(&h->block, block_bytes));
        h->block_bytes = block_bytes;
    }
    h->pos[0] = reinterpret_cast<float4 *>(h->block);
    h->pos[1] = h->pos[0] + n_slots;
    h->flags = reinterpret_cast<unsigned *>(h->pos[1] + n_slots);
    if (count > h->local_cap) {
        if (h->vel) cudaFree(h->vel);
        if (h->acc) cudaFree(h->acc);
        h->vel = h->acc = nullptr; h->local_cap = 0;
        CU_TRY(h, cudaMalloc(&h->vel, count * sizeof(float4)));
        CU_TRY(h, cudaMalloc(&h->acc, count * sizeof(float4)));
        h->local_cap = count;
    }
    h->n_total = n_total; h->rank = rank; h->world = world;
    h->i_begin = begin; h->n_local = count; h->slot_stride = stride; h->n_slots = n_slots;
    for (int r = 0; r < kMaxWorld; ++r) h->counts[r] = 0;
    for (int r = 0; r < world; ++r) { size_t b, c; slice_of(n_total, world, r, &b, &c); h->counts[r] = (int)c; }
    h->cur = 0;
    h->barrier_seq = 0;
    h->mass_checked = false;
    h->mass_state = -1;
    *h->host_fault = 0u;        // the stream is idle: a new configuration starts unfaulted
    // everything is padding until particles arrive
    const int total = (int)(2 * n_slots);
    pad_kernel<<<(total + 255) / 256, 256, 0, h->stream>>>(h->pos[0], 0, total);
    CHECK_LAUNCH(h, "pad_kernel");
    CU_TRY(h, cudaMemsetAsync(h->flags, 0, 256, h->stream));
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    h->configured = true;
    if (h->mode == NBODY_MODE_FAST && count > 0) {
        int rc = make_plan(h);
        if (rc != NBODY_OK) { h->configured = false; return rc; }
    }
    return NBODY_OK;
}

int nbody_cuda_slice(nbody_cuda_t h, size_t *begin, size_t *count, size_t *slot_begin, size_t *slot_stride) {
    if (!h) return NBODY_E_INVALID;
    if (!h->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_slice: not configured");
    if (begin) *begin = h->i_begin;
    if (count) *count = h->n_local;
    if (slot_begin) *slot_begin = (size_t)h->rank * h->slot_stride;
    if (slot_stride) *slot_stride = h->slot_stride;
    return NBODY_OK;
}

int nbody_cuda_upload_aos(nbody_cuda_t h, const void *particles, size_t count) {
    if (!h) return NBODY_E_INVALID;
    if (!h->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_upload_aos: call nbody_cuda_configure first");
    if (count != h->n_local) return fail(h, NBODY_E_INVALID, "nbody_cuda_upload_aos: count differs from the configured slice");
    if (count && !particles) return fail(h, NBODY_E_INVALID, "nbody_cuda_upload_aos: particles is NULL");
    CU_TRY(h, cudaSetDevice(h->device));
    if (*h->host_fault != 0u) {                 // a barrier gave up earlier: new particles restart the run
        CU_TRY(h, cudaStreamSynchronize(h->stream));
        *h->host_fault = 0u;
    }
    if (count == 0) return NBODY_OK;
    const size_t bytes = count * kRecFloats * sizeof(float);
    int rc = ensure_aos(h, bytes);
    if (rc != NBODY_OK) return rc;
    CU_TRY(h, cudaMemcpyAsync(h->d_aos, particles, bytes, cudaMemcpyHostToDevice, h->stream));
    const int grid = (int)((count + kThreads - 1) / kThreads);
    aos_to_soa_kernel<<<grid, kThreads, 0, h->stream>>>(reinterpret_cast<const float *>(h->d_aos), (int)count,
                                                       (int)(h->rank * h->slot_stride), h->pos[h->cur], h->vel, h->acc);
    CHECK_LAUNCH(h, "aos_to_soa_kernel");
    h->mass_checked = false;
    h->mass_state = -1;
    return NBODY_OK;
}

int nbody_cuda_download_aos(nbody_cuda_t h, void *particles, size_t count) {
    if (!h) return NBODY_E_INVALID;
    if (!h->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_download_aos: not configured");
    if (count != h->n_local) return fail(h, NBODY_E_INVALID, "nbody_cuda_download_aos: count differs from the configured slice");
    if (count && !particles) return fail(h, NBODY_E_INVALID, "nbody_cuda_download_aos: particles is NULL");
    CU_TRY(h, cudaSetDevice(h->device));
    if (count == 0) return NBODY_OK;
    const size_t bytes = count * kRecFloats * sizeof(float);
    int rc = ensure_aos(h, bytes);
    if (rc != NBODY_OK) return rc;
    const int grid = (int)((count + kThreads - 1) / kThreads);
    soa_to_aos_kernel<<<grid, kThreads, 0, h->stream>>>(reinterpret_cast<float *>(h->d_aos), (int)count,
                                                       (int)(h->rank * h->slot_stride), h->pos[h->cur], h->vel, h->acc);
    CHECK_LAUNCH(h, "soa_to_aos_kernel");
    CU_TRY(h, cudaMemcpyAsync(particles, h->d_aos, bytes, cudaMemcpyDeviceToHost, h->stream));
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    return check_fault(h);
}

int nbody_cuda_download_positions(nbody_cuda_t h, float *xyz, size_t count) {
    if (!h) return NBODY_E_INVALID;
    if (!h->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_download_positions: not configured");
    if (count != h->n_local) return fail(h, NBODY_E_INVALID, "nbody_cuda_download_positions: count differs from the configured slice");
    if (count && !xyz) return fail(h, NBODY_E_INVALID, "nbody_cuda_download_positions: xyz is NULL");
    CU_TRY(h, cudaSetDevice(h->device));
    if (count == 0) return NBODY_OK;
    const size_t bytes = count * 3 * sizeof(float);
    int rc = ensure_aos(h, bytes);
    if (rc != NBODY_OK) return rc;
    pack_positions_kernel<<<(int)((count + kThreads - 1) / kThreads), kThreads, 0, h->stream>>>(
        reinterpret_cast<float *>(h->d_aos), (int)count, (int)(h->rank * h->slot_stride), h->pos[h->cur]);
    CHECK_LAUNCH(h, "pack_positions_kernel");
    CU_TRY(h, cudaMemcpyAsync(xyz, h->d_aos, bytes, cudaMemcpyDeviceToHost, h->stream));
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    return check_fault(h);
}

int nbody_cuda_download_positions_async(nbody_cuda_t h, float *xyz, size_t count) {
    if (!h) return NBODY_E_INVALID;
    if (!h->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_download_positions_async: not configured");
    if (count != h->n_local) return fail(h, NBODY_E_INVALID, "nbody_cuda_download_positions_async: count differs from the configured slice");
    if (count && !xyz) return fail(h, NBODY_E_INVALID, "nbody_cuda_download_positions_async: xyz is NULL");
    CU_TRY(h, cudaSetDevice(h->device));
    if (count == 0) return NBODY_OK;
    if (!h->copy_stream) {
        CU_TRY(h, cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
        for (int k = 0; k < 2; ++k) {
            CU_TRY(h, cudaEventCreateWithFlags(&h->rec_packed[k], cudaEventDisableTiming));
            CU_TRY(h, cudaEventCreateWithFlags(&h->rec_copied[k], cudaEventDisableTiming));
        }
    }
    const size_t floats = count * 3;
    if (floats > h->rec_cap) {
        CU_TRY(h, cudaStreamSynchronize(h->copy_stream));          // nothing may still be draining the old buffers
        for (int k = 0; k < 2; ++k) {
            if (h->rec_stage[k]) cudaFree(h->rec_stage[k]);
            h->rec_stage[k] = nullptr; h->rec_busy[k] = false;
        }
        h->rec_cap = 0;
        for (int k = 0; k < 2; ++k) CU_TRY(h, cudaMalloc(&h->rec_stage[k], floats * sizeof(float)));
        h->rec_cap = floats;
    }
    const int k = (int)(h->rec_count & 1u);
    // staging k may still be on its way to the host (the record before last): the pack waits for that copy,
    // on the device -- the host does not block
    if (h->rec_busy[k]) CU_TRY(h, cudaStreamWaitEvent(h->stream, h->rec_copied[k], 0));
    pack_positions_kernel<<<(int)((count + kThreads - 1) / kThreads), kThreads, 0, h->stream>>>(
        h->rec_stage[k], (int)count, (int)(h->rank * h->slot_stride), h->pos[h->cur]);
    CHECK_LAUNCH(h, "pack_positions_kernel");
    CU_TRY(h, cudaEventRecord(h->rec_packed[k], h->stream));
    CU_TRY(h, cudaStreamWaitEvent(h->copy_stream, h->rec_packed[k], 0));
    CU_TRY(h, cudaMemcpyAsync(xyz, h->rec_stage[k], floats * sizeof(float), cudaMemcpyDeviceToHost, h->copy_stream));
    CU_TRY(h, cudaEventRecord(h->rec_copied[k], h->copy_stream));
    h->rec_busy[k] = true;
    h->rec_count++;
    return NBODY_OK;
}

int nbody_cuda_download_positions_wait(nbody_cuda_t h) {
    if (!h) return NBODY_E_INVALID;
    CU_TRY(h, cudaSetDevice(h->device));
    for (int k = 0; k < 2; ++k) {
        if (!h->rec_busy[k]) continue;
        CU_TRY(h, cudaEventSynchronize(h->rec_copied[k]));
        h->rec_busy[k] = false;
    }
    return report_fault(h);
}

int nbody_cuda_generate(nbody_cuda_t h, int kind, float size_a, float size_b, float G, unsigned long long seed, float *min_r2_out) {
    if (!h) return NBODY_E_INVALID;
    if (!h->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_generate: call nbody_cuda_configure first");
    if (kind != NBODY_GEN_PLUMMER && kind != NBODY_GEN_DISK) return fail(h, NBODY_E_INVALID, "nbody_cuda_generate: unknown kind");
    CU_TRY(h, cudaSetDevice(h->device));
    const size_t n = h->n_total;
    const double pi = 3.14159265358979323846;
    float min_r2 = 0.f, total_mass = 0.f;
    if (kind == NBODY_GEN_PLUMMER) {          // softening: 15 % of the mean spacing inside 2.4 a (Simulation.cpp:186-195)
        total_mass = (size_b < 0.f) ? (float)n : size_b;
        const float reff = 2.4f * size_a;
        const float vol = (float)((4.0 / 3.0) * pi) * reff * reff * reff;
        const float soft = cbrtf(vol / (float)(n ? n : 1)) * 0.15f;
        min_r2 = soft * soft;
    } else {                                   // 10 % of the mean spacing of the disk volume (Simulation.cpp:126-131)
        const float vol = (float)pi * size_a * size_a * size_b;
        const float soft = cbrtf(vol / (float)(n ? n : 1)) * 0.1f;
        min_r2 = soft * soft;
    }
    if (min_r2_out) *min_r2_out = min_r2;
    if (n == 0) return NBODY_OK;
    const int blocks = (int)((n + kThreads - 1) / kThreads);
    if ((size_t)3 * blocks > h->energy_cap) {
        if (h->d_energy) cudaFree(h->d_energy);
        h->d_energy = nullptr; h->energy_cap = 0;
        CU_TRY(h, cudaMalloc(&h->d_energy, (size_t)3 * blocks * sizeof(double)));
        h->energy_cap = (size_t)3 * blocks;
    }
    GenArgs a;
    a.pos = h->pos[h->cur]; a.vel = h->vel; a.vsum = h->d_energy;
    a.world = h->world; a.slot_stride = (int)h->slot_stride; a.rank = h->rank; a.n_total = (int)n;
    int d = 0;
    for (int r = 0; r < kMaxWorld; ++r) { a.counts[r] = h->counts[r]; a.displs[r] = d; d += h->counts[r]; }
    a.seed = seed; a.a = size_a; a.total_mass = total_mass; a.G = G; a.radius = size_a; a.thickness = size_b;
    if (kind == NBODY_GEN_PLUMMER) generate_kernel<0><<<blocks, kThreads, 0, h->stream>>>(a);
    else generate_kernel<1><<<blocks, kThreads, 0, h->stream>>>(a);
    CHECK_LAUNCH(h, "generate_kernel");
    if (kind == NBODY_GEN_PLUMMER && h->n_local > 0) {
        remove_bulk_velocity_kernel<<<(int)((h->n_local + kThreads - 1) / kThreads), kThreads, 0, h->stream>>>(
            h->vel, (int)h->n_local, h->d_energy, blocks, (double)total_mass);
        CHECK_LAUNCH(h, "remove_bulk_velocity_kernel");
    }
    if (h->n_local > 0) CU_TRY(h, cudaMemsetAsync(h->acc, 0, h->n_local * sizeof(float4), h->stream));
    h->mass_checked = false;
    h->mass_state = -1;
    return NBODY_OK;
}

int nbody_cuda_render_frame(nbody_cuda_t h, int width, int height, float camera_distance, float angle_x_deg,
                            float angle_y_deg, unsigned char *rgb) {
    if (!h) return NBODY_E_INVALID;
    if (!h->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_render_frame: not configured");
    if (width < 1 || height < 1 || width > 16384 || height > 16384 || !rgb)
        return fail(h, NBODY_E_INVALID, "nbody_cuda_render_frame: need 1 <= width, height <= 16384 and a buffer");
    CU_TRY(h, cudaSetDevice(h->device));
    const size_t pixels = (size_t)width * height;
    const size_t bytes = pixels * (sizeof(unsigned long long) + 3) + kMaxWorld * sizeof(int) + 64;
    int rc = ensure_aos(h, bytes);                 // the boundary staging buffer doubles as frame memory
    if (rc != NBODY_OK) return rc;
    unsigned long long *zbuf = reinterpret_cast<unsigned long long *>(h->d_aos);
    int *d_counts = reinterpret_cast<int *>(zbuf + pixels);
    unsigned char *d_rgb = reinterpret_cast<unsigned char *>(d_counts + kMaxWorld);
    CU_TRY(h, cudaMemsetAsync(zbuf, 0xff, pixels * sizeof(unsigned long long), h->stream));
    CU_TRY(h, cudaMemcpyAsync(d_counts, h->counts, kMaxWorld * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    Camera c;
    const float deg = 0.017453292519943295f;
    c.cx = cosf(angle_x_deg * deg); c.sx = sinf(angle_x_deg * deg);
    c.cy = cosf(angle_y_deg * deg); c.sy = sinf(angle_y_deg * deg);
    c.distance = camera_distance;
    c.fy = 0.5f * height / tanf(0.5f * 45.0f * deg);          // gluPerspective(45, w/h, 1, 5000), GlutRenderer.cpp:148
    c.fx = c.fy;
    c.radius = 2.0f;                                           // glutSolidSphere(2, 8, 6), GlutRenderer.cpp:131
    c.znear = 1.0f; c.zfar = 5000.0f;
    c.width = width; c.height = height;
    if (h->n_total > 0) {
        splat_kernel<<<(int)((h->n_slots + kThreads - 1) / kThreads), kThreads, 0, h->stream>>>(
            h->pos[h->cur], (int)h->n_slots, (int)h->slot_stride, d_counts, c, zbuf);
        CHECK_LAUNCH(h, "splat_kernel");
    }
    shade_kernel<<<(int)((pixels + kThreads - 1) / kThreads), kThreads, 0, h->stream>>>(h->pos[h->cur], c, zbuf, d_rgb);
    CHECK_LAUNCH(h, "shade_kernel");
    CU_TRY(h, cudaMemcpyAsync(rgb, d_rgb, pixels * 3, cudaMemcpyDeviceToHost, h->stream));
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    return NBODY_OK;
}

int nbody_cuda_step(nbody_cuda_t h, float G, float dt, float min_r2, int n_steps) {
    if (!h) return NBODY_E_INVALID;
    if (!h->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_step: not configured");
    if (n_steps < 0) return fail(h, NBODY_E_INVALID, "nbody_cuda_step: n_steps < 0");
    if (h->world > 1 && !all_peers_attached(h))
        return fail(h, NBODY_E_STATE, "nbody_cuda_step: world > 1 needs every peer attached (or use nbody_cuda_step_local + an all-gather)");
    if (report_fault(h) != NBODY_OK) return NBODY_E_CUDA;      // an earlier exchange timed out: do not step on stale positions
    CU_TRY(h, cudaSetDevice(h->device));
    // small systems on one GPU: all n_steps in one cooperative launch (force and integrator phases
    // separated by grid barriers instead of kernel boundaries)
    const bool use_resident = h->world == 1 && h->mode == NBODY_MODE_FAST && !h->profile_force && h->n_local > 0 && n_steps > 0 &&
                              h->resident_fits && (h->resident_mode == 2 || (h->resident_mode == 1 && h->n_total <= kResidentMaxN));
    if (use_resident) {
        int rc = ensure_mass_info(h);
        if (rc != NBODY_OK) return rc;
        ResidentArgs a;
        a.pos[0] = h->pos[0]; a.pos[1] = h->pos[1]; a.vel = h->vel; a.acc = h->acc; a.partial = h->partial;
        a.n = (int)h->n_local; a.cur = h->cur; a.n_steps = n_steps;
        a.mask_self = !(min_r2 >= kMaskSelfBelowEps2);
        a.eps2 = min_r2; a.G = G; a.dt = dt; a.plan = h->plan; a.mass_info = h->flags + kMassInfoWord;
        CU_TRY(h, cudaEventRecord(h->ev_begin, h->stream));
        // as in launch_force: until the host knows the mass flag both loops are enqueued, one returns at once
        if (h->mass_state != 0) {
            CU_TRY(h, launch_resident_variant<kLoopUnit>(h->targets_per_thread, h->plan.grid, h->stream, a));
            h->launches++;
        }
        if (h->mass_state != 1) {
            CU_TRY(h, launch_resident_variant<kLoopMassMul>(h->targets_per_thread, h->plan.grid, h->stream, a));
            h->launches++;
        }
        h->cur ^= (n_steps & 1);
        CU_TRY(h, cudaEventRecord(h->ev_end, h->stream));
        h->timed = true;
        return NBODY_OK;
    }
    int done = 0;
    const bool use_graph = h->world == 1 && !h->profile_force && n_steps >= 4 && h->n_local > 0;
    if (use_graph) {
        int rc = ensure_mass_info(h);           // outside the capture: it runs once per upload
        if (rc != NBODY_OK) return rc;
        if (h->step_graph && (h->graph_G != G || h->graph_dt != dt || h->graph_eps2 != min_r2 || h->graph_parity != h->cur ||
                              h->graph_mass_state != h->mass_state))
            drop_graph(h);
        if (!h->step_graph) {
            // two steps = one full turn of the double-buffered positions, so the graph can be replayed
            cudaGraph_t g = nullptr;
            CU_TRY(h, cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
            h->capturing = true;
            h->graph_launches = 0;
            rc = one_step(h, G, dt, min_r2, true);
            if (rc == NBODY_OK) rc = one_step(h, G, dt, min_r2, true);
            h->capturing = false;
            cudaError_t e = cudaStreamEndCapture(h->stream, &g);
            if (rc != NBODY_OK) { if (g) cudaGraphDestroy(g); return rc; }
            CU_TRY(h, e);
            e = cudaGraphInstantiate(&h->step_graph, g, 0);
            cudaGraphDestroy(g);
            CU_TRY(h, e);
            h->graph_G = G; h->graph_dt = dt; h->graph_eps2 = min_r2; h->graph_parity = h->cur;
            h->graph_mass_state = h->mass_state;
        }
    }
    CU_TRY(h, cudaEventRecord(h->ev_begin, h->stream));
    if (use_graph) {
        for (; done + 2 <= n_steps; done += 2) {
            CU_TRY(h, cudaGraphLaunch(h->step_graph, h->stream));
            h->launches += h->graph_launches;
        }
    }
    for (; done < n_steps; ++done) {
        int rc = one_step(h, G, dt, min_r2, true);
        if (rc != NBODY_OK) return rc;
    }
    CU_TRY(h, cudaEventRecord(h->ev_end, h->stream));
    h->timed = true;
    return NBODY_OK;
}

int nbody_cuda_step_local(nbody_cuda_t h, float G, float dt, float min_r2) {
    if (!h) return NBODY_E_INVALID;
    if (!h->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_step_local: not configured");
    CU_TRY(h, cudaSetDevice(h->device));
    CU_TRY(h, cudaEventRecord(h->ev_begin, h->stream));
    int rc = one_step(h, G, dt, min_r2, false);
    if (rc != NBODY_OK) return rc;
    CU_TRY(h, cudaEventRecord(h->ev_end, h->stream));
    h->timed = true;
    return NBODY_OK;
}

int nbody_cuda_compute_step_aos(nbody_cuda_t h, void *particles, size_t n, float G, float dt, float min_r2) {
    if (!h) return NBODY_E_INVALID;
    int rc = nbody_cuda_configure(h, n, 0, 1);
    if (rc != NBODY_OK) return rc;
    if (n == 0) return NBODY_OK;
    if ((rc = nbody_cuda_upload_aos(h, particles, n)) != NBODY_OK) return rc;
    if ((rc = nbody_cuda_step(h, G, dt, min_r2, 1)) != NBODY_OK) return rc;
    return nbody_cuda_download_aos(h, particles, n);
}

int nbody_cuda_synchronize(nbody_cuda_t h) {
    if (!h) return NBODY_E_INVALID;
    CU_TRY(h, cudaSetDevice(h->device));
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    return check_fault(h);
}

int nbody_cuda_last_step_ms(nbody_cuda_t h, float *ms) {
    if (!h || !ms) return NBODY_E_INVALID;
    if (!h->timed) return fail(h, NBODY_E_STATE, "nbody_cuda_last_step_ms: no step has run");
    CU_TRY(h, cudaSetDevice(h->device));
    CU_TRY(h, cudaEventSynchronize(h->ev_end));
    CU_TRY(h, cudaEventElapsedTime(ms, h->ev_begin, h->ev_end));
    return NBODY_OK;
}

int nbody_cuda_profile_force(nbody_cuda_t h, int enable) {
    if (!h) return NBODY_E_INVALID;
    h->profile_force = enable != 0;
    if (enable > 1) {                       // reserve event pairs now, so that none is created between timed launches
        CU_TRY(h, cudaSetDevice(h->device));
        while (h->force_events.size() < (size_t)enable) {
            cudaEvent_t a, b;
            CU_TRY(h, cudaEventCreate(&a));
            CU_TRY(h, cudaEventCreate(&b));
            h->force_events.emplace_back(a, b);
        }
    }
    return NBODY_OK;
}

int nbody_cuda_force_time(nbody_cuda_t h, double *total_ms, int *launches, int reset) {
    if (!h) return NBODY_E_INVALID;
    CU_TRY(h, cudaSetDevice(h->device));
    CU_TRY(h, cudaStreamSynchronize(h->stream));
    for (size_t k = 0; k < h->force_events_used; ++k) {
        float ms = 0.f;
        CU_TRY(h, cudaEventElapsedTime(&ms, h->force_events[k].first, h->force_events[k].second));
        h->force_ms_total += ms;
        h->force_launches++;
    }
    h->force_events_used = 0;
    if (total_ms) *total_ms = h->force_ms_total;
    if (launches) *launches = h->force_launches;
    if (reset) { h->force_ms_total = 0.0; h->force_launches = 0; }
    return NBODY_OK;
}

int nbody_cuda_plan(nbody_cuda_t h, int *targets_per_thread, int *grid, int *target_tiles, int *source_tiles) {
    if (!h) return NBODY_E_INVALID;
    if (!h->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_plan: not configured");
    if (targets_per_thread) *targets_per_thread = h->targets_per_thread;
    if (grid) *grid = h->plan.grid;
    if (target_tiles) *target_tiles = plan_tiles(h->plan);
    if (source_tiles) *source_tiles = h->plan.n_gran * kGran / kTileJ;
    return NBODY_OK;
}

int nbody_cuda_plan_describe(size_t n_local, size_t n_slots, int sm_count, int targets_per_thread, long long *header,
                             long long *first_gran, size_t first_gran_len, int *tile_owner_first, int *tile_owner_last,
                             size_t tiles_len) {
    if (!header || n_local == 0 || n_slots == 0 || n_slots % kTileJ != 0 || sm_count < 1 || n_local > n_slots)
        return NBODY_E_INVALID;
    ForcePlan p;
    const int k = pick_variant(n_local, n_slots, sm_count, targets_per_thread, &p);
    header[0] = p.I_main; header[1] = p.n_main; header[2] = p.n_tail; header[3] = p.tail_I[0]; header[4] = p.tail_I[1];
    header[5] = p.grid; header[6] = p.n_gran; header[7] = p.cost_total; header[8] = kVariants[k].ctas_per_sm;
    header[9] = (long long)plan_partial_slots(p); header[10] = kGran; header[11] = plan_tiles(p);
    if (first_gran) {
        if (first_gran_len < (size_t)p.grid + 1) return NBODY_E_INVALID;
        for (int c = 0; c <= p.grid; ++c) first_gran[c] = plan_first_gran(p, c);
    }
    if (tile_owner_first && tile_owner_last) {
        if (tiles_len < (size_t)plan_tiles(p)) return NBODY_E_INVALID;
        for (int t = 0; t < plan_tiles(p); ++t) {
            tile_owner_first[t] = plan_owner(p, (long long)t * p.n_gran);
            tile_owner_last[t] = plan_owner(p, (long long)(t + 1) * p.n_gran - 1);
        }
    }
    return NBODY_OK;
}

int nbody_cuda_set_option(nbody_cuda_t h, int option, long long value) {
    if (!h) return NBODY_E_INVALID;
    switch (option) {
        case NBODY_OPT_TARGETS_PER_THREAD:
            if (value != 0 && value != 1 && value != 2 && value != 4 && value != 7)
                return fail(h, NBODY_E_INVALID, "nbody_cuda_set_option: targets per thread must be 0 (automatic), 1, 2, 4 or 7");
            h->forced_I = (int)value;
            if (h->configured && h->mode == NBODY_MODE_FAST && h->n_local > 0) {
                CU_TRY(h, cudaSetDevice(h->device));
                CU_TRY(h, cudaStreamSynchronize(h->stream));
                drop_graph(h);
                return make_plan(h);
            }
            return NBODY_OK;
        case NBODY_OPT_RESIDENT_KERNEL:
            if (value < 0 || value > 2) return fail(h, NBODY_E_INVALID, "nbody_cuda_set_option: resident kernel must be 0 (off), 1 (small systems) or 2 (always)");
            h->resident_mode = (int)value;
            return NBODY_OK;
        case NBODY_OPT_BARRIER_TIMEOUT_MS:
            if (value < 1) return fail(h, NBODY_E_INVALID, "nbody_cuda_set_option: barrier time-out must be at least 1 ms");
            h->barrier_timeout_ns = (unsigned long long)value * 1000000ull;
            return NBODY_OK;
        default:
            return fail(h, NBODY_E_INVALID, "nbody_cuda_set_option: unknown option");
    }
}

long long nbody_cuda_launch_count(nbody_cuda_t h) { return h ? h->launches : 0; }

int nbody_cuda_energy(nbody_cuda_t h, float G, float min_r2, double *kinetic, double *potential) {
    if (!h) return NBODY_E_INVALID;
    if (!h->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_energy: not configured");
    CU_TRY(h, cudaSetDevice(h->device));
    double K = 0.0, U = 0.0;
    if (h->n_local > 0) {
        const int grid = (int)((h->n_local + kThreads - 1) / kThreads);
        if ((size_t)2 * grid > h->energy_cap) {
            if (h->d_energy) cudaFree(h->d_energy);
            h->d_energy = nullptr; h->energy_cap = 0;
            CU_TRY(h, cudaMalloc(&h->d_energy, (size_t)2 * grid * sizeof(double)));
            h->energy_cap = (size_t)2 * grid;
        }
        energy_kernel<<<grid, kThreads, 0, h->stream>>>(h->pos[h->cur], h->vel, (int)h->n_local,
                                                        (int)(h->rank * h->slot_stride), (int)h->n_slots,
                                                        (double)min_r2, h->d_energy);
        CHECK_LAUNCH(h, "energy_kernel");
        std::vector<double> host((size_t)2 * grid);
        CU_TRY(h, cudaMemcpyAsync(host.data(), h->d_energy, host.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
        CU_TRY(h, cudaStreamSynchronize(h->stream));
        for (int b = 0; b < grid; ++b) { K += host[2 * b]; U += host[2 * b + 1]; }
    }
    if (kinetic) *kinetic = K;
    if (potential) *potential = (double)G * U;
    return NBODY_OK;
}

void *nbody_cuda_stream(nbody_cuda_t h) { return h ? (void *)h->stream : nullptr; }

void *nbody_cuda_positions(nbody_cuda_t h) { return (h && h->configured) ? (void *)h->pos[h->cur] : nullptr; }

int nbody_cuda_peer_export(nbody_cuda_t h, void *token_out) {
    if (!h || !token_out) return NBODY_E_INVALID;
    if (!h->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_peer_export: not configured");
    CU_TRY(h, cudaSetDevice(h->device));
    PeerToken t;
    memset(&t, 0, sizeof(t));
    t.magic = kTokenMagic; t.rank = h->rank; t.world = h->world; t.device = h->device;
    t.pid = (long long)getpid();
    t.n_slots = h->n_slots; t.block_bytes = h->block_bytes;
    CU_TRY(h, cudaIpcGetMemHandle(&t.ipc, h->block));
    memset(token_out, 0, NBODY_PEER_HANDLE_BYTES);
    memcpy(token_out, &t, sizeof(t));
    return NBODY_OK;
}

int nbody_cuda_peer_attach(nbody_cuda_t h, int peer_rank, const void *token) {
    if (!h || !token) return NBODY_E_INVALID;
    if (!h->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_peer_attach: not configured");
    PeerToken t;
    memcpy(&t, token, sizeof(t));
    if (t.magic != kTokenMagic) return fail(h, NBODY_E_INVALID, "nbody_cuda_peer_attach: not a peer token");
    if (peer_rank == h->rank || peer_rank < 0 || peer_rank >= h->world || t.rank != peer_rank || t.world != h->world)
        return fail(h, NBODY_E_INVALID, "nbody_cuda_peer_attach: token rank/world do not match this configuration");
    if (t.n_slots != h->n_slots) return fail(h, NBODY_E_INVALID, "nbody_cuda_peer_attach: peer has a different slot layout");
    if (t.pid == (long long)getpid())
        return fail(h, NBODY_E_INVALID, "nbody_cuda_peer_attach: peer lives in this process; use nbody_cuda_peer_attach_local");
    CU_TRY(h, cudaSetDevice(h->device));
    if (h->attached[peer_rank]) return fail(h, NBODY_E_STATE, "nbody_cuda_peer_attach: peer already attached");
    void *p = nullptr;
    CU_TRY(h, cudaIpcOpenMemHandle(&p, t.ipc, cudaIpcMemLazyEnablePeerAccess));
    h->peer_block[peer_rank] = p;
    h->peer_is_ipc[peer_rank] = true;
    h->attached[peer_rank] = true;
    h->n_attached++;
    return NBODY_OK;
}

int nbody_cuda_peer_attach_local(nbody_cuda_t h, int peer_rank, nbody_cuda_t peer) {
    if (!h || !peer) return NBODY_E_INVALID;
    if (!h->configured || !peer->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_peer_attach_local: not configured");
    if (peer_rank == h->rank || peer_rank < 0 || peer_rank >= h->world || peer->rank != peer_rank ||
        peer->world != h->world || peer->n_slots != h->n_slots)
        return fail(h, NBODY_E_INVALID, "nbody_cuda_peer_attach_local: peer does not match this configuration");
    if (peer->device == h->device)
        return fail(h, NBODY_E_INVALID, "nbody_cuda_peer_attach_local: peers must be on different GPUs (their kernels wait on one another)");
    CU_TRY(h, cudaSetDevice(h->device));
    if (h->attached[peer_rank]) return fail(h, NBODY_E_STATE, "nbody_cuda_peer_attach_local: peer already attached");
    int can = 0;
    CU_TRY(h, cudaDeviceCanAccessPeer(&can, h->device, peer->device));
    if (!can) return fail(h, NBODY_E_CUDA, "nbody_cuda_peer_attach_local: no peer access between the two devices");
    cudaError_t e = cudaDeviceEnablePeerAccess(peer->device, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CU_TRY(h, e);
    cudaGetLastError();
    h->peer_block[peer_rank] = peer->block;
    h->peer_is_ipc[peer_rank] = false;
    h->attached[peer_rank] = true;
    h->n_attached++;
    return NBODY_OK;
}

int nbody_cuda_peer_detach_all(nbody_cuda_t h) {
    if (!h) return NBODY_E_INVALID;
    CU_TRY(h, cudaSetDevice(h->device));
    if (h->stream) CU_TRY(h, cudaStreamSynchronize(h->stream));     // nothing of ours may still be storing into a peer
    drop_peers(h);
    return NBODY_OK;
}

int nbody_cuda_exchange(nbody_cuda_t h) {
    if (!h) return NBODY_E_INVALID;
    if (!h->configured) return fail(h, NBODY_E_STATE, "nbody_cuda_exchange: not configured");
    if (h->world == 1) return NBODY_OK;
    if (!all_peers_attached(h)) return fail(h, NBODY_E_STATE, "nbody_cuda_exchange: not every peer is attached");
    CU_TRY(h, cudaSetDevice(h->device));
    if (h->n_local > 0) {
        PublishArgs a;
        a.src = h->pos[h->cur]; a.slot_begin = (int)(h->rank * h->slot_stride); a.n = (int)h->n_local;
        fill_peer_ptrs(h, h->cur, a.peer, &a.n_peers);
        publish_kernel<<<(int)((h->n_local + kThreads - 1) / kThreads), kThreads, 0, h->stream>>>(a);
        CHECK_LAUNCH(h, "publish_kernel");
    }
    return launch_barrier(h, false);
}

int nbody_cuda_host_alloc(void **ptr, size_t bytes) {
    if (!ptr) return NBODY_E_INVALID;
    *ptr = nullptr;
    cudaError_t e = cudaHostAlloc(ptr, bytes ? bytes : 1, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(nullptr, NBODY_E_NOMEM, std::string("cudaHostAlloc: ") + cudaGetErrorString(e));
    }
    return NBODY_OK;
}

int nbody_cuda_host_free(void *ptr) {
    if (!ptr) return NBODY_OK;
    cudaError_t e = cudaFreeHost(ptr);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(nullptr, NBODY_E_CUDA, std::string("cudaFreeHost: ") + cudaGetErrorString(e));
    }
    return NBODY_OK;
}

int nbody_cuda_measure_fma_peak(nbody_cuda_t h, int packed, double *flops_per_second) {
    if (!h || !flops_per_second) return NBODY_E_INVALID;
    CU_TRY(h, cudaSetDevice(h->device));
    float *d = nullptr;
    CU_TRY(h, cudaMalloc(&d, 256));
    const int iters = 4096, grid = h->sm_count * 8;
    cudaEvent_t e0, e1;
    CU_TRY(h, cudaEventCreate(&e0));
    CU_TRY(h, cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        CU_TRY(h, cudaEventRecord(e0, h->stream));
        if (packed) fma_peak_kernel<true><<<grid, kThreads, 0, h->stream>>>(d, iters, 1.0001f);
        else fma_peak_kernel<false><<<grid, kThreads, 0, h->stream>>>(d, iters, 1.0001f);
        CHECK_LAUNCH(h, "fma_peak_kernel");
        CU_TRY(h, cudaEventRecord(e1, h->stream));
        CU_TRY(h, cudaEventSynchronize(e1));
        float ms = 0.f;
        CU_TRY(h, cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
    // per thread and iteration: 64 FMA instructions, x2 lanes when packed, 2 flop each
    const double flop = (double)grid * kThreads * iters * 64.0 * (packed ? 2.0 : 1.0) * 2.0;
    *flops_per_second = flop / (best * 1e-3);
    return NBODY_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Allocation-free accelerations.
//
// Who needs it: hosts that embed this library but may not let it touch the device allocator --
// an engine that handed the whole GPU to its own arena at start-up (a later cudaMalloc can fail
// or fragment it), a stepping loop that cannot afford the device-wide synchronisation cudaMalloc
// and cudaFree imply, a process already at its memory budget.  Every other entry point of this
// library grows device buffers on demand; this one never does.  Its working memory -- the
// tile-padded copy of the sources and the segment partials of the fast force kernel -- lives in
// the fixed array below, which every GPU that loads the library carries until the process ends.
//
// The array is shared by all calls and is never cleared: nothing in it has to start at zero,
// because each call rewrites the padding of the staging area and every partial it later reads
// before reading it.  For the same reason two calls whose kernels overlap on a GPU (different
// streams, different threads) corrupt each other: run them one after the other.
// ---------------------------------------------------------------------------------------------
namespace {

constexpr size_t kNoAllocMaxN = NBODY_NOALLOC_MAX_N;                     // sources == targets
constexpr size_t kNoAllocStageSlots = kNoAllocMaxN;                      // multiple of kTileJ
constexpr size_t kNoAllocPartialSlots = 384 * 1024;                      // float4; >= (grid + target tiles) * 256 * I
constexpr size_t kNoAllocSlots = kNoAllocStageSlots + kNoAllocPartialSlots;   // 458752 float4 = 7 MiB
static_assert(kNoAllocMaxN % kTileJ == 0, "staging area must be whole tiles");

static __device__ __align__(256) float4 g_noalloc_scratch[kNoAllocSlots];

constexpr int kNoAllocMaxDevices = 64;
struct NoAllocDevice {
    bool ready = false;
    float4 *scratch = nullptr;   // DEVICE address of g_noalloc_scratch on that GPU
    int sm_count = 0;
};
NoAllocDevice g_noalloc_dev[kNoAllocMaxDevices];
std::mutex g_noalloc_mutex;      // guards the table above (set-up only; the launches themselves are not serialised)

int noalloc_fail(char *err, size_t err_len, int code, const char *what, cudaError_t e = cudaSuccess) {
    if (err && err_len) {
        if (e != cudaSuccess) snprintf(err, err_len, "nbody_cuda_accel_noalloc: %s: %s", what, cudaGetErrorString(e));
        else snprintf(err, err_len, "nbody_cuda_accel_noalloc: %s", what);
    }
    if (e != cudaSuccess) cudaGetLastError();
    return code;
}

}  // namespace

extern "C" int nbody_cuda_accel_noalloc(const void *d_positions, void *d_accelerations, size_t n, float G, float min_r2,
                                        void *stream, char *err, size_t err_len) {
    if (err && err_len) err[0] = 0;
    if (n > kNoAllocMaxN) {
        if (err && err_len)
            snprintf(err, err_len, "nbody_cuda_accel_noalloc: n = %zu exceeds NBODY_NOALLOC_MAX_N = %zu (the built-in scratch "
                     "is fixed; use nbody_cuda_step on a handle for larger systems)", n, (size_t)kNoAllocMaxN);
        return NBODY_E_INVALID;
    }
    if (n == 0) return NBODY_OK;
    if (!d_positions || !d_accelerations) return noalloc_fail(err, err_len, NBODY_E_INVALID, "positions / accelerations is NULL");
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return noalloc_fail(err, err_len, NBODY_E_CUDA, "cudaGetDevice", e);
    if (dev < 0 || dev >= kNoAllocMaxDevices) return noalloc_fail(err, err_len, NBODY_E_INVALID, "device ordinal out of range");
    NoAllocDevice d;
    {
        // One-time set-up per device, under a lock: two host threads may make the first call at once.
        std::lock_guard<std::mutex> guard(g_noalloc_mutex);
        NoAllocDevice &slot = g_noalloc_dev[dev];
        if (!slot.ready) {
            // This device is the current one.  In host code the name g_noalloc_scratch is only a
            // stand-in for the symbol, not an address a kernel can use, so ask the runtime where the
            // array lives on THIS GPU and hand that pointer to the kernels.
            float4 *ptr = nullptr;
            e = cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch);
            if (e != cudaSuccess || !ptr) return noalloc_fail(err, err_len, NBODY_E_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch)", e);
            int sms = 0;
            e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            if (e != cudaSuccess || sms <= 0) return noalloc_fail(err, err_len, NBODY_E_CUDA, "cudaDeviceGetAttribute(multiProcessorCount)", e);
            slot.scratch = ptr;
            slot.sm_count = sms;
            slot.ready = true;
        }
        d = slot;
    }
    const size_t n_slots = (n + kTileJ - 1) / kTileJ * kTileJ;
    ForcePlan plan;                    // same choice as make_plan(), without a handle and without touching any buffer
    const int I = kVariants[pick_variant(n, n_slots, d.sm_count, 0, &plan)].I;
    const size_t need = plan_partial_slots(plan);
    if (need > kNoAllocPartialSlots)
        return noalloc_fail(err, err_len, NBODY_E_INVALID, "this GPU's SM count needs more partial-sum scratch than the built-in array holds");
    cudaStream_t st = (cudaStream_t)stream;
    float4 *stage = d.scratch;
    float4 *partial = d.scratch + kNoAllocStageSlots;

    stage_positions_kernel<<<(int)(n_slots / kThreads), kThreads, 0, st>>>((const float4 *)d_positions, (int)n, stage, (int)n_slots);
    if ((e = cudaGetLastError()) != cudaSuccess) return noalloc_fail(err, err_len, NBODY_E_CUDA, "launch of stage_positions_kernel", e);
    ForceArgs a;
    a.pos = stage; a.partial = partial; a.n_local = (int)n; a.i_slot_begin = 0; a.eps2 = min_r2; a.plan = plan;
    a.mask_self = !(min_r2 >= kMaskSelfBelowEps2);
    a.mass_info = nullptr;
    launch_fast_variant<kLoopMassMul>(I, plan.grid, st, a);      // mass_info is NULL: the general loop
    if ((e = cudaGetLastError()) != cudaSuccess) return noalloc_fail(err, err_len, NBODY_E_CUDA, "launch of force_fast_kernel", e);
    reduce_partials_kernel<<<(int)((n + kThreads - 1) / kThreads), kThreads, 0, st>>>(partial, (int)n, kThreads * I, G, plan,
                                                                                      (float4 *)d_accelerations);
    if ((e = cudaGetLastError()) != cudaSuccess) return noalloc_fail(err, err_len, NBODY_E_CUDA, "launch of reduce_partials_kernel", e);
    return NBODY_OK;
}
