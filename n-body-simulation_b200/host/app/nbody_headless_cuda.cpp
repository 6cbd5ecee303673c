// nbody_headless_cuda -- headless driver that runs the reference's own NBodySimulation facade,
// generators, CheckpointManager and PerformanceLogger (all compiled unmodified from the reference
// tree) with CudaNaiveSimulation plugged in where src/apps/headless_main.cpp:38-69 constructs
// NaiveSimulation / MpiNaiveSimulation.  Same observable behaviour as the reference's sequential
// build: same flags (headless_main.cpp:83-122), same defaults (G=1, dt=0.5, min_r2=2.0 unless the
// generator sets it, :33-36), same trajectory file (simulation_output.bin written every step
// unless --disable-output, :330-340, :402-411) and the same "N,steps,seconds,workers" log line
// (include/PerformanceLogger.hpp:25-33).
//
// Additional flags (not in the reference):
//   --plummer N          Plummer sphere input (the reference's MPI build uses it for --random, :191-195)
//   --seed S             seed the reference's generator engine (Random.hpp:7) for reproducible inputs
//   --mode fast|strict   arithmetic mode of the CUDA kernels (default fast)
//   --device D           CUDA ordinal of the first GPU
//   --gpus P             spread the particles over GPUs D .. D+P-1 (slices as headless_main.cpp:249-259)
//   --resident           keep the state on the GPU between steps (no per-step host round trip);
//                        positions come back only when a trajectory record is due
//   --checkpoint-every K write a trajectory record every K-th step only (default 1, as the reference)
//   --energy-every K     print the FP64 energy every K steps
//   --checkpoint FILE    trajectory file name (default simulation_output.bin)
//   --frame-every K      with --resident: write frame_<step>.ppm every K steps -- the viewer's picture
//                        (GlutRenderer.cpp:98-148) computed by CUDA kernels, no window, no graphics API
//   --frame-size WxH     frame size (default 800x600);  --camera D,AX,AY  distance and angles (default 800,0,0)
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <memory>
#include <string>
#include <vector>
#include <algorithm>

#include "CheckpointManager.h"
#include "CudaNaiveSimulation.h"
#include "PerformanceLogger.hpp"
#include "Random.hpp"
#include "Simulation.h"
#include "World.h"

namespace {

struct Options {
    std::string log_path = "output.txt";
    std::string traj_path = "simulation_output.bin";
    std::string hacc_dir;
    std::size_t steps = 1000;
    std::size_t disk_n = 0, plummer_n = 0;
    bool solar = false, write_output = true, resident = false, seeded = false;
    unsigned seed = 0;
    int device = 0, gpus = 1;
    std::size_t checkpoint_every = 1, energy_every = 0, frame_every = 0;
    int frame_w = 800, frame_h = 600;
    float cam_d = 800.f, cam_ax = 0.f, cam_ay = 0.f;
    CudaNaiveSimulation::Mode mode = CudaNaiveSimulation::Mode::Fast;
    SimParams params;
};

Options parse(int argc, char **argv) {
    Options o;
    o.params.G = 1.0f; o.params.dt = 0.5f; o.params.min_r2 = 2.0f;
    auto next = [&](int &i) -> std::string { return (i + 1 < argc) ? std::string(argv[++i]) : std::string(); };
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--random") o.disk_n = std::stoul(next(i));
        else if (a == "--plummer") o.plummer_n = std::stoul(next(i));
        else if (a == "--solar") o.solar = true;
        else if (a == "--hacc") o.hacc_dir = next(i);
        else if (a == "--steps") o.steps = std::stoul(next(i));
        else if (a == "--out") o.log_path = next(i);
        else if (a == "--dt") o.params.dt = std::stof(next(i));
        else if (a == "--G") o.params.G = std::stof(next(i));
        else if (a == "--disable-output") o.write_output = false;
        else if (a == "--seed") { o.seed = (unsigned)std::stoul(next(i)); o.seeded = true; }
        else if (a == "--mode") o.mode = (next(i) == "strict") ? CudaNaiveSimulation::Mode::Strict : CudaNaiveSimulation::Mode::Fast;
        else if (a == "--device") o.device = std::stoi(next(i));
        else if (a == "--gpus") o.gpus = std::max(1, std::stoi(next(i)));
        else if (a == "--resident") o.resident = true;
        else if (a == "--checkpoint-every") o.checkpoint_every = std::max<std::size_t>(1, std::stoul(next(i)));
        else if (a == "--energy-every") o.energy_every = std::stoul(next(i));
        else if (a == "--checkpoint") o.traj_path = next(i);
        else if (a == "--frame-every") o.frame_every = std::stoul(next(i));
        else if (a == "--frame-size") { const std::string v = next(i); std::sscanf(v.c_str(), "%dx%d", &o.frame_w, &o.frame_h); }
        else if (a == "--camera") { const std::string v = next(i); std::sscanf(v.c_str(), "%f,%f,%f", &o.cam_d, &o.cam_ax, &o.cam_ay); }
        // unknown flags are ignored, as in the reference
    }
    return o;
}

}  // namespace

int main(int argc, char **argv) {
    Options opt = parse(argc, argv);
    if (opt.seeded) getRandomEngine().seed(opt.seed);

    std::cout << "CudaNaiveSimulation (" << (opt.mode == CudaNaiveSimulation::Mode::Strict ? "strict" : "fast") << ")\n";
    std::vector<int> devices;
    for (int d = 0; d < opt.gpus; ++d) devices.push_back(opt.device + d);
    auto algo_owner = std::make_unique<CudaNaiveSimulation>(devices, opt.mode);
    CudaNaiveSimulation *algo = algo_owner.get();            // kept for the device-resident extension
    NBodySimulation sim(std::move(algo_owner), opt.params);  // the reference facade, unchanged
    sim.setDt(opt.params.dt);
    sim.setG(opt.params.G);

    if (!opt.hacc_dir.empty()) sim.loadHACC(opt.hacc_dir);
    else if (opt.plummer_n) sim.generatePlummerSphere((int)opt.plummer_n, 100.0f, (float)opt.plummer_n);
    else if (opt.disk_n) sim.generateGalaxyDisk((int)opt.disk_n, 300.0f, 50.0f);
    else if (opt.solar) sim.setupSolarSystem(WORLD_WIDTH, WORLD_HEIGHT, WORLD_DEPTH);
    else { std::cerr << "No initial condition specified (--hacc, --random, --plummer or --solar)\n"; return 1; }
    const std::size_t n = sim.particles().size();
    if (n == 0) { std::cerr << "No particles initialized\n"; return 1; }
    const SimParams params = sim.params();                   // generators may have set min_r2

    PerformanceLogger logger(opt.log_path, RunMode::Run);
    if (!logger.ok()) std::cerr << "Warning: could not open output log: " << opt.log_path << "\n";

    CheckpointManager *ckpt = CheckpointManager::getInstance();
    if (opt.write_output) {
        ckpt->setFilePath(opt.traj_path);
        SimulationOutputHeader header;
        header.n_particles = n; header.target_steps = opt.steps; header.passed_steps = 0;
        ckpt->write_header(header);
        ckpt->write_masses(sim.particles().data(), (int)n);
    }

    auto report_energy = [&](std::size_t step) {
        double k = 0, u = 0;
        algo->energy(params, k, u);
        std::printf("energy step %zu: K = %.10e  U = %.10e  E = %.10e\n", step, k, u, k + u);
    };

    const auto t0 = std::chrono::high_resolution_clock::now();
    if (opt.resident) algo->upload(sim.particles());
    if (opt.energy_every) { if (!opt.resident) algo->upload(sim.particles()); report_energy(0); }
    for (std::size_t step = 0; step < opt.steps; ++step) {
        const bool record = opt.write_output && ((step + 1) % opt.checkpoint_every == 0);
        if (opt.resident) {
            algo->steps(params, 1);
            if (record) algo->download(sim.particles());
        } else {
            sim.step();                                      // -> computeStep(particles, params), Simulation.cpp:271
        }
        if (record) {
            ckpt->write_step(sim.particles().data(), (int)n);
            if (step % 100 == 0) std::cout << "Step " << step << " / " << opt.steps << " completed\n";
        }
        if (opt.energy_every && (step + 1) % opt.energy_every == 0) report_energy(step + 1);
        if (opt.frame_every && opt.resident && (step + 1) % opt.frame_every == 0) {
            std::vector<unsigned char> rgb;
            algo->renderFrame(opt.frame_w, opt.frame_h, opt.cam_d, opt.cam_ax, opt.cam_ay, rgb);
            char name[64];
            std::snprintf(name, sizeof(name), "frame_%06zu.ppm", step + 1);
            if (FILE *f = std::fopen(name, "wb")) {
                std::fprintf(f, "P6\n%d %d\n255\n", opt.frame_w, opt.frame_h);
                std::fwrite(rgb.data(), 1, rgb.size(), f);
                std::fclose(f);
            }
        }
    }
    if (opt.resident) algo->download(sim.particles());       // also drains the GPU before the clock stops
    const auto t1 = std::chrono::high_resolution_clock::now();
    const double seconds = std::chrono::duration<double>(t1 - t0).count();

    logger.log(n, opt.steps, seconds, (std::size_t)opt.gpus);   // workers column, as the MPI build logs its rank count
    std::cout << "Simulation completed in " << seconds << " seconds\n";
    std::printf("%.4e pairwise interactions/s (N^2 * steps / seconds), %lld kernel launches\n",
                (double)n * (double)n * (double)opt.steps / seconds, algo->kernelLaunches());
    return 0;
}
