/*
 * nbody_cuda.h -- C ABI of libnbody_cuda.so: the B200 (sm_100a) direct-sum N-body hot path.
 *
 * This is the drop-in boundary for ONE path of kalyniy/n-body-simulation: what a
 * `SimulationAlgorithm::computeStep` implementation needs in order to run
 * `NaiveSimulation` / `MpiNaiveSimulation`'s force accumulation + semi-implicit Euler update
 * on the GPU.  Plain pointers and sizes only; no C++ or torch types cross it.  Reference
 * citations are relative to the reference repository root.
 *
 * Conventions
 *   - every `int` function returns 0 on success and a negative NBODY_E_* code on failure; the
 *     text of the last failure is at nbody_cuda_last_error(handle) (NULL handle: the last
 *     failure of nbody_cuda_create on this thread).  Nothing throws, nothing calls exit().
 *   - every call on a handle makes the handle's device the calling thread's current CUDA device
 *     (cudaSetDevice) and leaves it so.
 *   - a handle is bound to ONE device and ONE host thread at a time (the reference calls
 *     computeStep from exactly one thread: src/Simulation.cpp:269-271).
 *   - `particles` is the reference's 40-byte array-of-structures record
 *     (include/Particle.hpp:53-59: position, velocity, acceleration, mass; offsets 0/12/24/36).
 *   - there is NO CPU fallback: without a usable CUDA device every call fails with
 *     NBODY_E_CUDA.
 */
#ifndef NBODY_CUDA_H
#define NBODY_CUDA_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nbody_cuda *nbody_cuda_t;

/* error codes */
#define NBODY_OK 0
#define NBODY_E_INVALID (-1) /* bad argument / call order */
#define NBODY_E_CUDA (-2)    /* a CUDA runtime call failed (message has the cudaError string) */
#define NBODY_E_NOMEM (-3)   /* device or pinned-host allocation failed */
#define NBODY_E_STATE (-4)   /* the handle is not configured for this call */

/* arithmetic modes */
#define NBODY_MODE_FAST 0   /* FP32, packed FFMA2 + MUFU.RSQ, blocked summation (the product)            */
#define NBODY_MODE_STRICT 1 /* FP32, the exact operation order and rounding of the SEQUENTIAL NaiveSimulation: bit-identical to it
                               for any world size; not to MpiNaiveSimulation, whose summation order depends on the rank count */
#define NBODY_MODE_FP64 2   /* forces summed in FP64 from the FP32 state, rounded to FP32 once; optional, no counterpart */

/* Size of the record this library reads and writes; equals sizeof(particle_t) = 40
 * (include/Particle.hpp:53-59). */
size_t nbody_cuda_particle_bytes(void);

/* Library / build identification, e.g. "nbody_cuda 0.1 sm_100a". */
const char *nbody_cuda_version(void);

/* Number of visible CUDA devices (0 when none or the driver is unusable). */
int nbody_cuda_device_count(void);

/* Replaces: constructing the algorithm object (src/apps/headless_main.cpp:38-69,
 * `std::make_unique<NaiveSimulation>()`).  `device` is a CUDA ordinal. */
int nbody_cuda_create(nbody_cuda_t *out, int device, int mode);
void nbody_cuda_destroy(nbody_cuda_t h);
const char *nbody_cuda_last_error(nbody_cuda_t h);

/* Declares how the particle index space is sharded.  `n_total` targets are cut into `world`
 * contiguous slices with the reference's own rule, counts[r] = n/P + (r < n%P)
 * (src/apps/headless_main.cpp:249-259); this handle owns slice `rank`.  Replaces the per-step
 * MPI_Allgather of counts in src/MpiNaiveSimulation.cpp:33-35 (static here).  (Re)allocates
 * device buffers when the sizes change.  world == 1: the whole system, as NaiveSimulation. */
int nbody_cuda_configure(nbody_cuda_t h, size_t n_total, int rank, int world);

/* Slice bookkeeping of the current configuration (any pointer may be NULL). */
int nbody_cuda_slice(nbody_cuda_t h, size_t *begin, size_t *count, size_t *slot_begin, size_t *slot_stride);

/* Host AoS -> device SoA for the LOCAL slice (count = slice count of `rank`).  Replaces the
 * MPI_Scatterv hand-over of src/apps/headless_main.cpp:264-269.  Uploads position, velocity and
 * mass; the incoming `acceleration` field is ignored (the generators leave it uninitialised,
 * src/Simulation.cpp:101,161,208, and computeStep overwrites it, src/NaiveSimulation.cpp:18-19).
 * With world > 1 the other ranks' positions must then be brought in by an exchange
 * (nbody_cuda_positions + a collective, or nbody_cuda_exchange once peers are attached). */
int nbody_cuda_upload_aos(nbody_cuda_t h, const void *particles, size_t count);

/* Device SoA -> host AoS for the local slice: writes position, velocity, acceleration; `mass`
 * bytes are rewritten with the values that were uploaded (never modified on the device).
 * Replaces the per-step MPI_Gatherv of src/apps/headless_main.cpp:368-373. */
int nbody_cuda_download_aos(nbody_cuda_t h, void *particles, size_t count);

/* Positions only, as packed float[count][3]: exactly one record of the reference's trajectory file
 * (src/CheckpointManager.cpp:86-99 builds the same array from the AoS vector), 12 instead of 40
 * bytes per particle over PCIe. */
int nbody_cuda_download_positions(nbody_cuda_t h, float *xyz, size_t count);

/* The same record WITHOUT stopping the run (the reference writes a record every step,
 * src/apps/headless_main.cpp:402-411, and its step loop waits for it).  _async packs the positions
 * into one of two device staging buffers on the handle's stream and starts the copy to `xyz` on a
 * separate copy stream, then returns: steps queued afterwards run while the record is still crossing
 * PCIe.  `xyz` should be page-locked (nbody_cuda_host_alloc; with pageable memory the copy is not
 * asynchronous) and must stay untouched until _wait returns.  At most two records are in flight --
 * a third _async call makes the DEVICE wait for the oldest copy, not the host.  _wait blocks until
 * every record requested so far has landed in host memory. */
int nbody_cuda_download_positions_async(nbody_cuda_t h, float *xyz, size_t count);
int nbody_cuda_download_positions_wait(nbody_cuda_t h);

/* Fills the configured system with a synthetic particle set generated ON the device, instead of
 * nbody_cuda_upload_aos: the distributions of NBodySimulation::generatePlummerSphere
 * (src/Simulation.cpp:173-267; size_a = scale radius, size_b = total mass, < 0 for one per particle)
 * or ::generateGalaxyDisk (:88-136; size_a = radius, size_b = thickness), drawn from a Philox stream
 * keyed by (seed, particle index) rather than the reference's mt19937 -- same statistics, not the
 * same particles.  Every rank of a sharded run calls it with the same arguments and gets the same
 * global system; no exchange is needed afterwards.  *min_r2_out receives the softening the
 * reference's generator would set (:131, :195). */
#define NBODY_GEN_PLUMMER 0
#define NBODY_GEN_DISK 1
int nbody_cuda_generate(nbody_cuda_t h, int kind, float size_a, float size_b, float G, unsigned long long seed, float *min_r2_out);

/* Headless frame of the device-resident positions, RGB8 rows top to bottom into `rgb`
 * (width*height*3 bytes, host): the picture src/renderers/GlutRenderer.cpp:98-141 draws -- one
 * orange sphere of radius 2 per particle, one directional light, camera at `camera_distance`
 * (default there: 800) turned by angle_x then angle_y degrees, gluPerspective(45, w/h, 1, 5000) --
 * computed by two CUDA kernels (z-buffered disc splat + per-pixel shading).  No graphics API is
 * involved.  Draws every particle of the replicated position array (all ranks' slices). */
int nbody_cuda_render_frame(nbody_cuda_t h, int width, int height, float camera_distance, float angle_x_deg,
                            float angle_y_deg, unsigned char *rgb);

/* n_steps x { force accumulation over all n_total sources for the local targets;
 * v += a*dt; x += v*dt }  -- src/NaiveSimulation.cpp:7-51 / src/MpiNaiveSimulation.cpp:10-141.
 * Device-resident: no host traffic.  With world > 1 this requires attached peers
 * (nbody_cuda_peer_attach): the integrator stores each new position into every peer's array
 * over NVLink and a device-side flag barrier closes the step (replaces the MPI_Sendrecv ring of
 * src/MpiNaiveSimulation.cpp:106-120).  Asynchronous w.r.t. the host; ordered on the handle's
 * stream. */
int nbody_cuda_step(nbody_cuda_t h, float G, float dt, float min_r2, int n_steps);

/* One step WITHOUT the exchange: afterwards the "current" position array holds the new
 * positions of the local slice only, and the caller completes it with an all-gather on
 * nbody_cuda_stream() (NCCL through torch.distributed), in place on nbody_cuda_positions(). */
int nbody_cuda_step_local(nbody_cuda_t h, float G, float dt, float min_r2);

/* The reference call itself: computeStep(std::vector<particle_t>&, const SimParams&)
 * (include/SimulationAlgorithm.h:18-22) for world == 1 -- upload `n` particles, one step,
 * download, synchronous.  Tolerates a different `n` on every call and n == 0. */
int nbody_cuda_compute_step_aos(nbody_cuda_t h, void *particles, size_t n, float G, float dt, float min_r2);

/* Blocks until everything queued on the handle's stream has finished. */
int nbody_cuda_synchronize(nbody_cuda_t h);

/* Device time, in milliseconds, of the most recent nbody_cuda_step / _step_local /
 * _compute_step_aos call (CUDA events on the handle's stream; synchronises). */
int nbody_cuda_last_step_ms(nbody_cuda_t h, float *ms);

/* Per-kernel timing of the force kernel: when enabled every force launch is bracketed by CUDA
 * events; _force_time returns the sum over launches since the last reset, and their number.
 * enable > 1 additionally creates that many event pairs up front (none is then created between
 * timed launches). */
int nbody_cuda_profile_force(nbody_cuda_t h, int enable);
int nbody_cuda_force_time(nbody_cuda_t h, double *total_ms, int *launches, int reset);

/* Work plan of the fast force kernel for the current configuration (no counterpart in the
 * reference): targets held per thread in the main target tiles, persistent grid size, number of
 * target tiles (main tiles plus up to two narrower tail tiles) and of 256-source tiles; the
 * tiles x sources rectangle is cut across the grid in 32-source granules of equal cost.  Any
 * pointer may be NULL. */
int nbody_cuda_plan(nbody_cuda_t h, int *targets_per_thread, int *grid, int *target_tiles, int *source_tiles);

/* The same plan as pure host arithmetic, for any slice shape and SM count -- needs no device, so
 * the plan's invariants (every granule owned by exactly one CTA, no empty CTA, the owner formula
 * the finishing kernel uses) are testable on a CPU-only machine.  n_slots is the padded source
 * count (multiple of 256), targets_per_thread 0 lets the plan choose.
 *   header[12]       I_main, n_main, n_tail, tail_I[0], tail_I[1], grid, granules per tile, total cost,
 *                    CTAs per SM, scratch float4s, sources per granule, tiles
 *   first_gran       [grid + 1] first global granule of each CTA (and the end), or NULL
 *   tile_owner_*     [tiles] first and last CTA that contribute to each target tile, or NULL */
int nbody_cuda_plan_describe(size_t n_local, size_t n_slots, int sm_count, int targets_per_thread, long long *header,
                             long long *first_gran, size_t first_gran_len, int *tile_owner_first, int *tile_owner_last,
                             size_t tiles_len);

/* Tuning and diagnostic switches of a handle (no counterpart in the reference).
 *   NBODY_OPT_TARGETS_PER_THREAD  0 (default): the plan chooses; 1, 2, 4 or 7 forces the main tile width.
 *   NBODY_OPT_RESIDENT_KERNEL     nbody_cuda_step on one GPU in fast mode can run all n_steps in ONE cooperative
 *                                 launch (force and integrator phases separated by grid-wide barriers): 1 (default)
 *                                 does so for systems of up to 12 288 particles, where the two kernel boundaries per
 *                                 step otherwise cost more than the arithmetic; 0 never; 2 whenever the grid fits.
 *                                 Results are bit-identical either way.
 *   NBODY_OPT_BARRIER_TIMEOUT_MS  how long the device-side wait for a peer's step counter may last before
 *                                 the run is reported as faulted (default 30 000); raise it for steps that
 *                                 take longer than that (strict / FP64 modes at very large N). */
#define NBODY_OPT_TARGETS_PER_THREAD 2
#define NBODY_OPT_BARRIER_TIMEOUT_MS 3
#define NBODY_OPT_RESIDENT_KERNEL 4
int nbody_cuda_set_option(nbody_cuda_t h, int option, long long value);

/* Number of kernels this library has launched on the handle since creation. */
long long nbody_cuda_launch_count(nbody_cuda_t h);

/* Energy diagnostic in FP64 (no counterpart in the reference, which has no energy code):
 * K = sum 1/2 m v^2 over the local slice; U = -G sum_{i in slice} sum_{j>i} m_i m_j / sqrt(r^2+min_r2)
 * over global indices, so that summing K and U over ranks gives the system totals. */
int nbody_cuda_energy(nbody_cuda_t h, float G, float min_r2, double *kinetic, double *potential);

/* --- multi-GPU plumbing (one process per GPU) -------------------------------------------- */

/* The handle's CUDA stream (a cudaStream_t), for torch.cuda.ExternalStream. */
void *nbody_cuda_stream(nbody_cuda_t h);

/* Device pointer to the CURRENT global position array: float4 {x,y,z,mass} per slot,
 * `world * slot_stride` slots; rank r's particles occupy slots [r*slot_stride, r*slot_stride+counts[r]),
 * the rest are zero-mass padding.  Valid until the next configure. */
void *nbody_cuda_positions(nbody_cuda_t h);

/* Peer-to-peer exchange set-up.  _peer_export writes an opaque NBODY_PEER_HANDLE_BYTES token
 * describing this handle's shared block; tokens are swapped between ranks out of band
 * (torch.distributed.all_gather_object) and each rank attaches every other rank's token.
 * Peers must be other processes on other GPUs of the same NVLink domain. */
#define NBODY_PEER_HANDLE_BYTES 128
int nbody_cuda_peer_export(nbody_cuda_t h, void *token_out);
int nbody_cuda_peer_attach(nbody_cuda_t h, int peer_rank, const void *token);
/* Same-process variant (one process driving several GPUs): attach another handle directly. */
int nbody_cuda_peer_attach_local(nbody_cuda_t h, int peer_rank, nbody_cuda_t peer);
/* Waits for this handle's stream and unmaps every attached peer block.  TEAR-DOWN ORDER for peers in
 * other processes: CUDA requires that an exported block is not freed while another process still
 * has it open, so before any rank destroys or reconfigures its handle (both free the block),
 *   1. every rank calls nbody_cuda_peer_detach_all,
 *   2. the ranks meet at a host barrier (MPI_Barrier / torch.distributed.barrier),
 *   3. only then nbody_cuda_destroy / nbody_cuda_configure.
 * (The reference's ranks have nothing to unmap: MPI_Sendrecv owns its buffers,
 * src/MpiNaiveSimulation.cpp:106-120.) */
int nbody_cuda_peer_detach_all(nbody_cuda_t h);

/* Publishes the local slice of the CURRENT position array to every attached peer and waits for
 * theirs (the exchange that must follow nbody_cuda_upload_aos when world > 1). */
int nbody_cuda_exchange(nbody_cuda_t h);

/* FP32 FMA-pipe micro-benchmark (no counterpart in the reference): dependency-free FFMA chains
 * (packed = 0) or FFMA2 chains (packed = 1) on every SM; reports achieved flop/s.  bench.py
 * prints it beside the nominal #SM x 128 lanes x 2 x clock so the roofline denominator is a
 * measured number. */
int nbody_cuda_measure_fma_peak(nbody_cuda_t h, int packed, double *flops_per_second);

/* Accelerations of n bodies on one another WITHOUT any allocation (this entry point has no
 * counterpart in the reference).  For hosts that must keep this library away from the device
 * allocator: an application that gave all GPU memory to its own arena at start-up, a stepping loop
 * that cannot tolerate the device-wide synchronisation of cudaMalloc / cudaFree, a process already
 * at its memory budget.  The call makes no cudaMalloc, cudaFree or other allocating call; its
 * working memory is a fixed 7 MiB array built into the library image (resident on every GPU that
 * loads the library, until exit), which is why n is limited to NBODY_NOALLOC_MAX_N -- larger n
 * is refused with NBODY_E_INVALID and a message, nothing is launched.
 *
 *   d_positions      device pointer, n x float4 {x, y, z, mass}       (read)
 *   d_accelerations  device pointer, n x float4 {ax, ay, az, 0}       (written)
 *   stream           a cudaStream_t, or NULL for the default stream; the call only enqueues work
 *   err, err_len     optional buffer that receives the failure text (may be NULL / 0)
 *
 * Same arithmetic as NBODY_MODE_FAST (a_i = G * sum_{j != i} m_j (p_j - p_i) / (r^2 + min_r2)^1.5).
 * Runs on the CURRENT device.  The first call on each device looks up where the built-in array
 * lives on that GPU (cudaGetSymbolAddress); make that first call before any stream capture begins.
 * All calls share the one array and it keeps whatever the previous call left (each call rewrites
 * everything it reads, so nothing needs clearing): calls whose kernels overlap on the same GPU
 * -- two streams, two threads -- race; serialise them.  d_accelerations may be the same buffer as
 * d_positions (the sources are copied into the built-in array before anything is written).  The
 * per-device look-up is remembered for the life of the process and is safe to trigger from several
 * host threads at once; it does not survive cudaDeviceReset -- after a reset, do not call this entry
 * point again in the same process. */
#define NBODY_NOALLOC_MAX_N 65536
int nbody_cuda_accel_noalloc(const void *d_positions, void *d_accelerations, size_t n, float G, float min_r2,
                             void *stream, char *err, size_t err_len);

/* --- host memory ------------------------------------------------------------------------- */

/* Page-locked host buffers for callers that want asynchronous, full-rate PCIe copies. */
int nbody_cuda_host_alloc(void **ptr, size_t bytes);
int nbody_cuda_host_free(void *ptr);

#ifdef __cplusplus
}
#endif
#endif /* NBODY_CUDA_H */
