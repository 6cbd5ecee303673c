// CudaNaiveSimulation -- the direct-sum algorithm of NaiveSimulation (src/NaiveSimulation.cpp) on
// one or several B200s, behind the reference's plug-in interface (include/SimulationAlgorithm.h:18-22).
//
//   auto algo = std::make_unique<CudaNaiveSimulation>();          // as headless_main.cpp:57-62
//   NBodySimulation sim(std::move(algo), params);                 // unchanged reference facade
//   sim.step();                                                   // -> computeStep, in place
//
// computeStep keeps the reference's contract exactly: the caller's vector is the source of truth
// on entry and holds updated position, velocity AND acceleration on return; mass is untouched; the
// size may change between calls (headless_main.cpp:262) and may be 0.  Everything below the class
// goes through the C ABI in include/nbody_cuda.h; there is no CPU fallback -- construction throws
// std::runtime_error (the reference's own convention for unrecoverable set-up errors,
// src/DatasetLoader.cpp:12) when no CUDA device is usable.
//
// With several devices the particle index space is cut with the reference's slice rule
// (headless_main.cpp:249-259), every GPU keeps all positions and computes full rows for its slice,
// and the integrator kernel stores the new positions straight into every peer over NVLink -- the
// role MpiNaiveSimulation's ring plays (src/MpiNaiveSimulation.cpp:51-121) -- all driven from this
// one host thread, no MPI.  Strict mode gives the same bits for any device count.
#pragma once
#include <cstddef>
#include <string>
#include <vector>

#include "nbody_reference_iface.h"

struct nbody_cuda;   // opaque C handle (include/nbody_cuda.h)

class CudaNaiveSimulation : public SimulationAlgorithm {
public:
    enum class Mode { Fast = 0, Strict = 1 };   // Strict: bit-identical to NaiveSimulation

    explicit CudaNaiveSimulation(int device = 0, Mode mode = Mode::Fast);
    CudaNaiveSimulation(const std::vector<int> &devices, Mode mode);   // one slice per listed device
    ~CudaNaiveSimulation() override;
    CudaNaiveSimulation(const CudaNaiveSimulation &) = delete;
    CudaNaiveSimulation &operator=(const CudaNaiveSimulation &) = delete;

    // The reference call: one step, host vector in / host vector out.
    void computeStep(std::vector<particle_t> &particles, const SimParams &params) override;

    // Device-resident extension for long runs: upload once, step many times, download on demand
    // (what removes the per-step 40 B/particle round trip of computeStep).
    void upload(const std::vector<particle_t> &particles);
    void steps(const SimParams &params, int n_steps);
    void download(std::vector<particle_t> &particles);

    // Diagnostics
    double lastStepMilliseconds();                       // CUDA-event time of the last step call (max over devices)
    void energy(const SimParams &params, double &kinetic, double &potential);
    long long kernelLaunches() const;
    // Headless picture of the device positions with the reference viewer's camera and markers
    // (src/renderers/GlutRenderer.cpp:98-148), RGB8 rows top to bottom; CUDA kernels, no graphics API.
    void renderFrame(int width, int height, float camera_distance, float angle_x_deg, float angle_y_deg,
                     std::vector<unsigned char> &rgb);
    Mode mode() const { return mode_; }
    int deviceCount() const { return (int)handles_.size(); }

private:
    void check(nbody_cuda *h, int code, const char *what);
    void configure(std::size_t n);
    std::vector<nbody_cuda *> handles_;
    Mode mode_;
    std::size_t resident_n_ = 0;
    bool configured_ = false;
};
