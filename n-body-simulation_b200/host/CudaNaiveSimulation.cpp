#include "CudaNaiveSimulation.h"

#include <stdexcept>

#include "nbody_cuda.h"

CudaNaiveSimulation::CudaNaiveSimulation(int device, Mode mode) : mode_(mode) {
    nbody_cuda_t h = nullptr;
    const int rc = nbody_cuda_create(&h, device, mode == Mode::Strict ? NBODY_MODE_STRICT : NBODY_MODE_FAST);
    if (rc != NBODY_OK) throw std::runtime_error(std::string("CudaNaiveSimulation: ") + nbody_cuda_last_error(nullptr));
    handle_ = h;
}

CudaNaiveSimulation::~CudaNaiveSimulation() { nbody_cuda_destroy(handle_); }

void CudaNaiveSimulation::check(int code, const char *what) {
    if (code != NBODY_OK)
        throw std::runtime_error(std::string("CudaNaiveSimulation::") + what + ": " + nbody_cuda_last_error(handle_));
}

void CudaNaiveSimulation::computeStep(std::vector<particle_t> &particles, const SimParams &params) {
    check(nbody_cuda_compute_step_aos(handle_, particles.data(), particles.size(), params.G, params.dt, params.min_r2),
          "computeStep");
    resident_n_ = particles.size();
}

void CudaNaiveSimulation::upload(const std::vector<particle_t> &particles) {
    check(nbody_cuda_configure(handle_, particles.size(), 0, 1), "upload/configure");
    check(nbody_cuda_upload_aos(handle_, particles.data(), particles.size()), "upload");
    resident_n_ = particles.size();
}

void CudaNaiveSimulation::steps(const SimParams &params, int n_steps) {
    check(nbody_cuda_step(handle_, params.G, params.dt, params.min_r2, n_steps), "steps");
}

void CudaNaiveSimulation::download(std::vector<particle_t> &particles) {
    particles.resize(resident_n_);
    check(nbody_cuda_download_aos(handle_, particles.data(), particles.size()), "download");
}

double CudaNaiveSimulation::lastStepMilliseconds() {
    float ms = 0.f;
    check(nbody_cuda_last_step_ms(handle_, &ms), "lastStepMilliseconds");
    return ms;
}

void CudaNaiveSimulation::energy(const SimParams &params, double &kinetic, double &potential) {
    check(nbody_cuda_energy(handle_, params.G, params.min_r2, &kinetic, &potential), "energy");
}

long long CudaNaiveSimulation::kernelLaunches() const { return nbody_cuda_launch_count(handle_); }
