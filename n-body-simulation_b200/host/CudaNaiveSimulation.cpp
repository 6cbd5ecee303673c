#include "CudaNaiveSimulation.h"

#include <algorithm>
#include <stdexcept>

#include "nbody_cuda.h"

CudaNaiveSimulation::CudaNaiveSimulation(int device, Mode mode) : CudaNaiveSimulation(std::vector<int>{device}, mode) {}

CudaNaiveSimulation::CudaNaiveSimulation(const std::vector<int> &devices, Mode mode) : mode_(mode) {
    if (devices.empty()) throw std::runtime_error("CudaNaiveSimulation: empty device list");
    for (int d : devices) {
        nbody_cuda_t h = nullptr;
        const int rc = nbody_cuda_create(&h, d, mode == Mode::Strict ? NBODY_MODE_STRICT : NBODY_MODE_FAST);
        if (rc != NBODY_OK) {
            const std::string msg = nbody_cuda_last_error(nullptr);
            for (auto *p : handles_) nbody_cuda_destroy(p);
            throw std::runtime_error("CudaNaiveSimulation: " + msg);
        }
        handles_.push_back(h);
    }
}

CudaNaiveSimulation::~CudaNaiveSimulation() {
    for (auto *h : handles_) nbody_cuda_synchronize(h);     // nobody may still be storing into a peer
    for (auto *h : handles_) nbody_cuda_destroy(h);
}

void CudaNaiveSimulation::check(nbody_cuda *h, int code, const char *what) {
    if (code != NBODY_OK)
        throw std::runtime_error(std::string("CudaNaiveSimulation::") + what + ": " + nbody_cuda_last_error(h));
}

// Declares the slices for n particles; with several devices also (re)connects every pair, since a
// size change moves the position arrays.
void CudaNaiveSimulation::configure(std::size_t n) {
    const int world = (int)handles_.size();
    if (configured_ && n == resident_n_) return;
    for (int r = 0; r < world; ++r) check(handles_[r], nbody_cuda_configure(handles_[r], n, r, world), "configure");
    for (int r = 0; r < world; ++r)
        for (int q = 0; q < world; ++q)
            if (q != r) check(handles_[r], nbody_cuda_peer_attach_local(handles_[r], q, handles_[q]), "attach");
    resident_n_ = n;
    configured_ = true;
}

void CudaNaiveSimulation::upload(const std::vector<particle_t> &particles) {
    configure(particles.size());
    const int world = (int)handles_.size();
    for (int r = 0; r < world; ++r) {
        std::size_t begin = 0, count = 0;
        check(handles_[r], nbody_cuda_slice(handles_[r], &begin, &count, nullptr, nullptr), "slice");
        check(handles_[r], nbody_cuda_upload_aos(handles_[r], particles.data() + begin, count), "upload");
    }
    for (auto *h : handles_) check(h, nbody_cuda_exchange(h), "exchange");      // no-op on one device
}

void CudaNaiveSimulation::steps(const SimParams &params, int n_steps) {
    if (handles_.size() == 1) {
        check(handles_[0], nbody_cuda_step(handles_[0], params.G, params.dt, params.min_r2, n_steps), "steps");
        return;
    }
    // step by step across the devices so that no device's queue runs ahead of the others' by more
    // than the launches of one step (each step ends in a cross-device barrier kernel)
    for (int s = 0; s < n_steps; ++s)
        for (auto *h : handles_) check(h, nbody_cuda_step(h, params.G, params.dt, params.min_r2, 1), "steps");
}

void CudaNaiveSimulation::download(std::vector<particle_t> &particles) {
    particles.resize(resident_n_);
    for (auto *h : handles_) {
        std::size_t begin = 0, count = 0;
        check(h, nbody_cuda_slice(h, &begin, &count, nullptr, nullptr), "slice");
        check(h, nbody_cuda_download_aos(h, particles.data() + begin, count), "download");
    }
}

void CudaNaiveSimulation::computeStep(std::vector<particle_t> &particles, const SimParams &params) {
    if (handles_.size() == 1) {
        check(handles_[0],
              nbody_cuda_compute_step_aos(handles_[0], particles.data(), particles.size(), params.G, params.dt, params.min_r2),
              "computeStep");
        resident_n_ = particles.size();
        configured_ = true;
        return;
    }
    if (particles.empty()) return;
    upload(particles);
    steps(params, 1);
    download(particles);
}

double CudaNaiveSimulation::lastStepMilliseconds() {
    float worst = 0.f;
    for (auto *h : handles_) {
        float ms = 0.f;
        check(h, nbody_cuda_last_step_ms(h, &ms), "lastStepMilliseconds");
        worst = std::max(worst, ms);
    }
    return worst;
}

void CudaNaiveSimulation::energy(const SimParams &params, double &kinetic, double &potential) {
    kinetic = potential = 0.0;
    for (auto *h : handles_) {
        double k = 0, u = 0;
        check(h, nbody_cuda_energy(h, params.G, params.min_r2, &k, &u), "energy");
        kinetic += k; potential += u;
    }
}

long long CudaNaiveSimulation::kernelLaunches() const {
    long long n = 0;
    for (auto *h : handles_) n += nbody_cuda_launch_count(h);
    return n;
}

void CudaNaiveSimulation::renderFrame(int width, int height, float camera_distance, float angle_x_deg, float angle_y_deg,
                                      std::vector<unsigned char> &rgb) {
    rgb.resize((std::size_t)std::max(width, 0) * std::max(height, 0) * 3);
    // every device holds all positions; the first one draws
    check(handles_[0], nbody_cuda_render_frame(handles_[0], width, height, camera_distance, angle_x_deg, angle_y_deg, rgb.data()),
          "renderFrame");
}
