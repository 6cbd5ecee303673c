/*
 * ref_shim.cpp -- C entry points around the UNMODIFIED reference classes.
 *
 * TEST INFRASTRUCTURE ONLY (same rule as nbody_oracle.c).  This file is
 * compiled TOGETHER WITH the reference's own sources, where they lie under
 * /root/reference, by oracle/Makefile into oracle/_ref/libnbody_ref.so.  No
 * reference source is copied into this repository; the .so is git-ignored and
 * travels to the GPU box as a built artefact.
 *
 * What it exposes:
 *   - layout probes (sizeof/offsetof of particle_t, SimParams, the checkpoint header)
 *   - the reference input generators behind a seed (Random.hpp:7 returns a
 *     mutable engine reference, so seeding it makes disk/Plummer/cluster
 *     inputs reproducible)
 *   - NaiveSimulation::computeStep          (src/NaiveSimulation.cpp:7-10)
 *   - MpiNaiveSimulation::computeStep       (src/MpiNaiveSimulation.cpp:10-13)
 *     with every MPI rank played by a host thread (oracle/mpi_shim/mpi.h)
 *   - NBodySimulation::step                 (src/Simulation.cpp:269-278)
 *   - CheckpointManager write/read          (src/CheckpointManager.cpp)
 */
#include <algorithm>
#include <barrier>
#include <chrono>
#include <cstddef>
#include <cstring>
#include <iostream>
#include <memory>
#include <sstream>
#include <thread>
#include <vector>

#include <mpi.h>  // oracle/mpi_shim/mpi.h

#include "CheckpointManager.h"
#include "MpiNaiveSimulation.h"
#include "NaiveSimulation.h"
#include "Random.hpp"
#include "Simulation.h"
#include "SimulationOutput.h"

// ---------------------------------------------------------------------------
// Thread-backed MPI stand-in
// ---------------------------------------------------------------------------
namespace {

struct Mailbox {
    const void *buf = nullptr;
    int count = 0;
    int type = 0;
    int dest = -1;
};

struct ShimWorld {
    int size = 1;
    std::unique_ptr<std::barrier<>> bar;
    std::vector<Mailbox> boxes;
    std::vector<unsigned char> gather;   // MPI_Allgather staging
};

ShimWorld *g_world = nullptr;
thread_local int t_rank = 0;

int type_bytes(MPI_Datatype t) { return t == MPI_BYTE ? 1 : 4; }

struct CoutSilencer {
    std::streambuf *old;
    std::ostringstream sink;
    CoutSilencer() : old(std::cout.rdbuf(sink.rdbuf())) {}
    ~CoutSilencer() { std::cout.rdbuf(old); }
};

}  // namespace

extern "C" {

int MPI_Comm_rank(MPI_Comm, int *rank) { *rank = t_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm, int *size) { *size = g_world ? g_world->size : 1; return MPI_SUCCESS; }
int MPI_Finalize(void) { return MPI_SUCCESS; }

int MPI_Allgather(const void *sendbuf, int sendcount, MPI_Datatype sendtype,
                  void *recvbuf, int, MPI_Datatype, MPI_Comm)
{
    const size_t bytes = (size_t)sendcount * type_bytes(sendtype);
    if (!g_world || g_world->size == 1) { std::memcpy(recvbuf, sendbuf, bytes); return MPI_SUCCESS; }
    ShimWorld &w = *g_world;
    if (t_rank == 0) w.gather.resize(bytes * w.size);
    w.bar->arrive_and_wait();
    std::memcpy(w.gather.data() + bytes * t_rank, sendbuf, bytes);
    w.bar->arrive_and_wait();
    std::memcpy(recvbuf, w.gather.data(), bytes * w.size);
    w.bar->arrive_and_wait();
    return MPI_SUCCESS;
}

int MPI_Sendrecv(const void *sendbuf, int sendcount, MPI_Datatype sendtype, int dest, int,
                 void *recvbuf, int recvcount, MPI_Datatype recvtype, int source, int,
                 MPI_Comm, MPI_Status *)
{
    if (!g_world || g_world->size == 1) {
        std::memcpy(recvbuf, sendbuf, (size_t)std::min(sendcount, recvcount) * type_bytes(sendtype));
        return MPI_SUCCESS;
    }
    ShimWorld &w = *g_world;
    w.boxes[t_rank] = Mailbox{sendbuf, sendcount, sendtype, dest};
    w.bar->arrive_and_wait();
    const Mailbox &m = w.boxes[source];
    const size_t have = (size_t)m.count * type_bytes(m.type);
    const size_t want = (size_t)recvcount * type_bytes(recvtype);
    std::memcpy(recvbuf, m.buf, std::min(have, want));
    w.bar->arrive_and_wait();
    return MPI_SUCCESS;
}

// ---------------------------------------------------------------------------
// Layout probes
// ---------------------------------------------------------------------------
int ref_sizeof_particle(void) { return (int)sizeof(particle_t); }
int ref_offsetof_velocity(void) { return (int)offsetof(particle_t, velocity); }
int ref_offsetof_acceleration(void) { return (int)offsetof(particle_t, acceleration); }
int ref_offsetof_mass(void) { return (int)offsetof(particle_t, mass); }
int ref_sizeof_simparams(void) { return (int)sizeof(SimParams); }
int ref_sizeof_output_header(void) { return (int)sizeof(SimulationOutputHeader); }
void ref_default_simparams(float *out3) { SimParams p; out3[0] = p.G; out3[1] = p.dt; out3[2] = p.min_r2; }

// ---------------------------------------------------------------------------
// Seeded generators (src/Simulation.cpp:41-267)
// ---------------------------------------------------------------------------
void ref_seed(unsigned s) { getRandomEngine().seed(s); }

static long copy_out(NBodySimulation &sim, void *out, long capacity, float *min_r2_out)
{
    const auto &v = sim.particles();
    if (min_r2_out) *min_r2_out = sim.params().min_r2;
    if ((long)v.size() > capacity) return -(long)v.size();
    std::memcpy(out, v.data(), v.size() * sizeof(particle_t));
    return (long)v.size();
}

static SimParams headless_params(float G)
{
    SimParams p;              // src/apps/headless_main.cpp:33-36
    p.G = G; p.dt = 0.5f; p.min_r2 = 2.0f;
    return p;
}

long ref_generate_disk(int n, float radius, float thickness, float G, void *out, long capacity, float *min_r2_out)
{
    CoutSilencer quiet;
    NBodySimulation sim(std::make_unique<NaiveSimulation>(), headless_params(G));
    sim.generateGalaxyDisk(n, radius, thickness);
    return copy_out(sim, out, capacity, min_r2_out);
}

long ref_generate_plummer(int n, float scale_radius, float total_mass, float G, void *out, long capacity, float *min_r2_out)
{
    CoutSilencer quiet;
    NBodySimulation sim(std::make_unique<NaiveSimulation>(), headless_params(G));
    sim.generatePlummerSphere(n, scale_radius, total_mass);
    return copy_out(sim, out, capacity, min_r2_out);
}

long ref_generate_clusters(int n_clusters, int per_cluster, float separation, void *out, long capacity)
{
    CoutSilencer quiet;
    NBodySimulation sim(std::make_unique<NaiveSimulation>(), headless_params(1.0f));
    sim.generateClusters(n_clusters, per_cluster, separation);
    return copy_out(sim, out, capacity, nullptr);
}

long ref_setup_solar(int W, int H, int D, float G, void *out, long capacity)
{
    CoutSilencer quiet;
    NBodySimulation sim(std::make_unique<NaiveSimulation>(), headless_params(G));
    sim.setupSolarSystem(W, H, D);
    return copy_out(sim, out, capacity, nullptr);
}

// ---------------------------------------------------------------------------
// The algorithms
// ---------------------------------------------------------------------------
double ref_naive_steps(void *particles, size_t n, float G, float dt, float min_r2, int steps)
{
    std::vector<particle_t> v(n);
    if (n) std::memcpy(v.data(), particles, n * sizeof(particle_t));
    SimParams p; p.G = G; p.dt = dt; p.min_r2 = min_r2;
    NaiveSimulation algo;
    auto t0 = std::chrono::steady_clock::now();
    for (int s = 0; s < steps; ++s) algo.computeStep(v, p);
    auto t1 = std::chrono::steady_clock::now();
    if (n) std::memcpy(particles, v.data(), n * sizeof(particle_t));
    return std::chrono::duration<double>(t1 - t0).count();
}

// Through the facade, as the headless loop does (headless_main.cpp:362-364):
// includes the render-buffer copy of Simulation.cpp:273-277.
double ref_facade_steps(void *particles, size_t n, float G, float dt, float min_r2, int steps)
{
    SimParams p; p.G = G; p.dt = dt; p.min_r2 = min_r2;
    NBodySimulation sim(std::make_unique<NaiveSimulation>(), p);
    sim.particles().resize(n);
    if (n) std::memcpy(sim.particles().data(), particles, n * sizeof(particle_t));
    auto t0 = std::chrono::steady_clock::now();
    for (int s = 0; s < steps; ++s) sim.step();
    auto t1 = std::chrono::steady_clock::now();
    if (n) std::memcpy(particles, sim.particles().data(), n * sizeof(particle_t));
    return std::chrono::duration<double>(t1 - t0).count();
}

// MpiNaiveSimulation with `nranks` thread-ranks.  Slices follow
// headless_main.cpp:249-259.  All ranks get the SAME SimParams (the reference
// app leaves min_r2 un-broadcast; that is an app bug, not part
// of the algorithm).  Returns the wall time of the step loop (max over ranks).
double ref_mpi_naive_steps(void *particles, size_t n, float G, float dt, float min_r2, int steps, int nranks)
{
    if (nranks < 1) nranks = 1;
    particle_t *all = static_cast<particle_t *>(particles);
    std::vector<size_t> counts(nranks), displs(nranks);
    const size_t base = n / nranks, rem = n % nranks;
    for (int r = 0; r < nranks; ++r) counts[r] = base + ((size_t)r < rem ? 1 : 0);
    displs[0] = 0;
    for (int r = 1; r < nranks; ++r) displs[r] = displs[r - 1] + counts[r - 1];

    ShimWorld world;
    world.size = nranks;
    world.bar = std::make_unique<std::barrier<>>(nranks);
    world.boxes.resize(nranks);
    g_world = &world;

    std::vector<double> secs(nranks, 0.0);
    auto body = [&](int r) {
        t_rank = r;
        std::vector<particle_t> local(all + displs[r], all + displs[r] + counts[r]);
        SimParams p; p.G = G; p.dt = dt; p.min_r2 = min_r2;
        MpiNaiveSimulation algo;
        world.bar->arrive_and_wait();
        auto t0 = std::chrono::steady_clock::now();
        for (int s = 0; s < steps; ++s) algo.computeStep(local, p);
        world.bar->arrive_and_wait();
        auto t1 = std::chrono::steady_clock::now();
        secs[r] = std::chrono::duration<double>(t1 - t0).count();
        if (counts[r]) std::memcpy(all + displs[r], local.data(), counts[r] * sizeof(particle_t));
    };
    std::vector<std::thread> threads;
    for (int r = 1; r < nranks; ++r) threads.emplace_back(body, r);
    body(0);
    for (auto &t : threads) t.join();
    t_rank = 0;
    g_world = nullptr;
    return *std::max_element(secs.begin(), secs.end());
}

// The reference's own definition of "seconds" for an MPI run (headless_main.cpp:344-421 with
// --disable-output): barrier; clock; steps x { sim.step()  -- computeStep on the local slice plus
// the render-buffer copy of Simulation.cpp:273-277 --  then MPI_Gatherv of every slice to rank 0 };
// barrier; clock.  The gather is modelled as what a shared-memory MPI does: every rank copies its
// 40-byte records into rank 0's buffer, rank 0 returns when all have arrived, the others return
// at once (no extra synchronisation is charged to the CPU side).  Returns rank 0's wall time.
double ref_mpi_headless_loop(void *particles, size_t n, float G, float dt, float min_r2, int steps, int nranks)
{
    if (nranks < 1) nranks = 1;
    particle_t *all = static_cast<particle_t *>(particles);
    std::vector<size_t> counts(nranks), displs(nranks);
    const size_t base = n / nranks, rem = n % nranks;
    for (int r = 0; r < nranks; ++r) counts[r] = base + ((size_t)r < rem ? 1 : 0);
    displs[0] = 0;
    for (int r = 1; r < nranks; ++r) displs[r] = displs[r - 1] + counts[r - 1];

    ShimWorld world;
    world.size = nranks;
    world.bar = std::make_unique<std::barrier<>>(nranks);
    world.boxes.resize(nranks);
    g_world = &world;
    CoutSilencer quiet;                              // once, on this thread: std::cout is shared by the thread-ranks
    std::barrier<> gathered(nranks);                 // phase completes when every rank's slice is in the buffer
    std::vector<particle_t> global_buffer(n);        // rank 0's receive buffer (headless_main.cpp:351-355)

    double secs0 = 0.0;
    auto body = [&](int r) {
        t_rank = r;
        SimParams p; p.G = G; p.dt = dt; p.min_r2 = min_r2;
        NBodySimulation sim(std::make_unique<MpiNaiveSimulation>(), p);
        sim.particles().assign(all + displs[r], all + displs[r] + counts[r]);
        world.bar->arrive_and_wait();
        auto t0 = std::chrono::steady_clock::now();
        for (int s = 0; s < steps; ++s) {
            sim.step();
            if (counts[r]) std::memcpy(global_buffer.data() + displs[r], sim.particles().data(), counts[r] * sizeof(particle_t));
            if (r == 0) gathered.arrive_and_wait();
            else (void)gathered.arrive();
        }
        world.bar->arrive_and_wait();
        auto t1 = std::chrono::steady_clock::now();
        if (r == 0) secs0 = std::chrono::duration<double>(t1 - t0).count();
    };
    std::vector<std::thread> threads;
    for (int r = 1; r < nranks; ++r) threads.emplace_back(body, r);
    body(0);
    for (auto &t : threads) t.join();
    t_rank = 0;
    g_world = nullptr;
    if (n) std::memcpy(all, global_buffer.data(), n * sizeof(particle_t));
    return secs0;
}

// ---------------------------------------------------------------------------
// CheckpointManager (trajectory file)
// ---------------------------------------------------------------------------
void ref_checkpoint_begin(const char *path, void *particles, size_t n, size_t target_steps)
{
    CheckpointManager *cm = CheckpointManager::getInstance();
    cm->setFilePath(path);
    SimulationOutputHeader h;
    h.n_particles = n; h.target_steps = target_steps; h.passed_steps = 0;
    cm->write_header(h);
    cm->write_masses(static_cast<particle_t *>(particles), (int)n);
}

void ref_checkpoint_write_step(const char *path, void *particles, size_t n)
{
    CheckpointManager *cm = CheckpointManager::getInstance();
    cm->setFilePath(path);
    cm->write_step(static_cast<particle_t *>(particles), (int)n);
}

void ref_checkpoint_read_header(const char *path, size_t *out3)
{
    CheckpointManager *cm = CheckpointManager::getInstance();
    cm->setFilePath(path);
    SimulationOutputHeader h = cm->read_header();
    out3[0] = h.n_particles; out3[1] = h.target_steps; out3[2] = h.passed_steps;
}

void ref_checkpoint_read_masses(const char *path, float *masses_out, size_t n)
{
    CheckpointManager *cm = CheckpointManager::getInstance();
    cm->setFilePath(path);
    cm->read_masses(masses_out, n);
}

size_t ref_checkpoint_read_step(const char *path, float *positions_out, size_t step_index)
{
    CheckpointManager *cm = CheckpointManager::getInstance();
    cm->setFilePath(path);
    return cm->read_step(positions_out, step_index);
}

}  // extern "C"
