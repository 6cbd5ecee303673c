// GPU test of the C++ drop-in class, through the base-class pointer the reference's facade holds
// (include/Simulation.h:14,52).  Checker: the oracle's row-order restatement (oracle/nbody_oracle.c).
// Exit code 0 = pass.  Run by tests/test_gpu_cpp_host.py.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <memory>
#include <random>
#include <stdexcept>
#include <vector>

#include "CudaNaiveSimulation.h"

extern "C" void oracle_naive_step_rows_f32(void *p, size_t n, float G, float dt, float eps2, int steps);
extern "C" int nbody_cuda_device_count(void);

static std::vector<particle_t> make(size_t n, unsigned seed) {
    std::mt19937 rng(seed);
    std::uniform_real_distribution<float> u(-100.f, 100.f), m(0.5f, 2.f), v(-1.f, 1.f);
    std::vector<particle_t> p(n);
    for (auto &q : p) {
        q.position = {u(rng), u(rng), u(rng)};
        q.velocity = {v(rng), v(rng), v(rng)};
        q.acceleration = {7.f, 7.f, 7.f};    // garbage on purpose: must be overwritten
        q.mass = m(rng);
    }
    return p;
}

static int fails = 0;
#define EXPECT(c, ...) do { if (!(c)) { ++fails; std::printf("FAIL %s:%d: ", __FILE__, __LINE__); std::printf(__VA_ARGS__); std::printf("\n"); } } while (0)

int main() {
    SimParams params;              // reference defaults, then the headless overrides (headless_main.cpp:33-36)
    EXPECT(params.G == 1.0f && params.dt == 0.05f && params.min_r2 == 1e-8f, "SimParams defaults");
    params.dt = 0.5f; params.min_r2 = 2.0f;

    // strict mode through the abstract interface: bitwise equal to the reference order, sizes changing per call
    {
        std::unique_ptr<SimulationAlgorithm> algo = std::make_unique<CudaNaiveSimulation>(0, CudaNaiveSimulation::Mode::Strict);
        for (size_t n : {size_t(1000), size_t(0), size_t(1), size_t(4097), size_t(300)}) {
            auto p = make(n, 1234 + (unsigned)n), want = p;
            for (int s = 0; s < 3; ++s) algo->computeStep(p, params);
            oracle_naive_step_rows_f32(want.data(), n, params.G, params.dt, params.min_r2, 3);
            EXPECT(p.size() == n && (n == 0 || std::memcmp(p.data(), want.data(), n * sizeof(particle_t)) == 0),
                   "strict computeStep differs from the oracle at n=%zu", n);
        }
    }
    // fast mode: tolerance, device-resident extension, diagnostics
    {
        CudaNaiveSimulation fast(0, CudaNaiveSimulation::Mode::Fast);
        const size_t n = 20000;
        auto p = make(n, 99), want = p, q = p;
        fast.computeStep(p, params);
        oracle_naive_step_rows_f32(want.data(), n, params.G, params.dt, params.min_r2, 1);
        double worst = 0;
        for (size_t i = 0; i < n; ++i) {
            const float *a = &p[i].acceleration.x, *b = &want[i].acceleration.x;
            double num = 0, den = 0;
            for (int c = 0; c < 3; ++c) { num += double(a[c] - b[c]) * (a[c] - b[c]); den += double(b[c]) * b[c]; }
            worst = std::fmax(worst, std::sqrt(num / den));
            EXPECT(p[i].mass == q[i].mass, "mass changed");
        }
        EXPECT(worst < 2e-5, "fast acceleration error %.3g", worst);
        fast.upload(q);
        fast.steps(params, 1);
        std::vector<particle_t> r;
        fast.download(r);
        EXPECT(r.size() == n && std::memcmp(r.data(), p.data(), n * sizeof(particle_t)) == 0, "resident path differs from computeStep");
        double k = 0, u = 0;
        fast.energy(params, k, u);
        EXPECT(k > 0 && u < 0, "energy signs");
        EXPECT(fast.lastStepMilliseconds() > 0 && fast.kernelLaunches() > 0, "timers");
    }
    // several GPUs driven from this one thread: strict mode gives the same bits as one GPU
    {
        const int ndev = nbody_cuda_device_count();
        for (int world : {2, 4, 8}) {
            if (ndev < world) { std::printf("skip %d-GPU check (%d visible)\n", world, ndev); continue; }
            std::vector<int> devs;
            for (int d = 0; d < world; ++d) devs.push_back(d);
            for (auto mode : {CudaNaiveSimulation::Mode::Strict, CudaNaiveSimulation::Mode::Fast}) {
                std::unique_ptr<SimulationAlgorithm> algo = std::make_unique<CudaNaiveSimulation>(devs, mode);
                for (size_t n : {size_t(10007), size_t(3), size_t(65536)}) {
                    auto p = make(n, 777 + (unsigned)n), want = p;
                    for (int s = 0; s < 3; ++s) algo->computeStep(p, params);
                    oracle_naive_step_rows_f32(want.data(), n, params.G, params.dt, params.min_r2, 3);
                    if (mode == CudaNaiveSimulation::Mode::Strict) {
                        EXPECT(std::memcmp(p.data(), want.data(), n * sizeof(particle_t)) == 0, "%d-GPU strict differs at n=%zu", world, n);
                    } else {
                        double worst = 0;
                        for (size_t i = 0; i < n; ++i) {
                            const float *a = &p[i].position.x, *b = &want[i].position.x;
                            double num = 0, den = 0;
                            for (int c = 0; c < 3; ++c) { num += double(a[c] - b[c]) * (a[c] - b[c]); den += double(b[c]) * b[c]; }
                            worst = std::fmax(worst, std::sqrt(num / den));
                        }
                        EXPECT(worst < 1e-5, "%d-GPU fast position error %.3g at n=%zu", world, worst, n);
                    }
                }
                auto *cuda = static_cast<CudaNaiveSimulation *>(algo.get());
                auto p = make(40000, 5), q = p;
                cuda->upload(p);
                cuda->steps(params, 5);
                cuda->download(p);
                CudaNaiveSimulation one(0, mode);
                one.upload(q); one.steps(params, 5); one.download(q);
                if (mode == CudaNaiveSimulation::Mode::Strict)
                    EXPECT(std::memcmp(p.data(), q.data(), p.size() * sizeof(particle_t)) == 0, "%d-GPU resident run differs from 1 GPU", world);
                std::printf("%d GPUs %s: ok so far (%d failures)\n", world, mode == CudaNaiveSimulation::Mode::Strict ? "strict" : "fast", fails);
            }
        }
    }
    // errors are exceptions, not exits
    try {
        CudaNaiveSimulation bad(4096);
        EXPECT(false, "bad device accepted");
    } catch (const std::runtime_error &e) {
        EXPECT(std::strstr(e.what(), "device") != nullptr, "message: %s", e.what());
    }
    std::printf(fails ? "FAILED (%d)\n" : "PASS\n", fails);
    return fails ? 1 : 0;
}
