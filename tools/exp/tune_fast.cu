// Experiment (not product): times the PRODUCT force_fast_kernel<I> with alternative formulations
// of its inner tile loop (tile_variants.cuh, selected by -DVARIANT=k).
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#ifndef VARIANT
#define VARIANT 0
#endif
#if VARIANT != 0
#define NBODY_TILE_INTERACT_OVERRIDE "tile_variants.cuh"
#endif
#include "../../n-body-simulation_b200/csrc/nbody_kernels.cuh"
using namespace nbody;
#ifndef TI_I
#define TI_I 7
#endif
int main(int argc, char** argv) {
    const int n = argc > 1 ? atoi(argv[1]) : 262144;
    const int I = TI_I;
    int sm; cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0);
    const int n_slots = (n + 255) / 256 * 256;
    std::vector<float4> h(n_slots);
    srand(1);
    for (int k = 0; k < n_slots; ++k) h[k] = make_float4(rand() / (float)RAND_MAX * 100, rand() / (float)RAND_MAX * 100, rand() / (float)RAND_MAX * 100, 1.0f);
    float4 *pos, *partial; unsigned* info;
    cudaMalloc(&pos, n_slots * sizeof(float4));
    cudaMemcpy(pos, h.data(), n_slots * sizeof(float4), cudaMemcpyHostToDevice);
    const int TI = 256 * I;
    ForcePlan p; p.n_itiles = (n + TI - 1) / TI; p.n_jtiles = n_slots / 256; p.units = (long long)p.n_itiles * p.n_jtiles; p.grid = sm;
    cudaMalloc(&partial, (size_t)(p.grid + p.n_itiles) * TI * sizeof(float4));
    cudaMalloc(&info, 16);
    for (int unit = 1; unit >= 0; --unit) {
        unsigned hi[2] = {unit ? 0u : 1u, 0x3f800000u};
        cudaMemcpy(info, hi, 8, cudaMemcpyHostToDevice);
        ForceArgs a; a.pos = pos; a.partial = partial; a.n_local = n; a.i_slot_begin = 0; a.eps2 = 0.01f; a.plan = p; a.mass_info = info;
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        auto launch = [&]() { if (unit) force_fast_kernel<I, true><<<p.grid, 256>>>(a); else force_fast_kernel<I, false><<<p.grid, 256>>>(a); };
        launch(); cudaDeviceSynchronize();
        float best = 1e30f;
        for (int r = 0; r < 3; ++r) { cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms; }
        cudaFuncAttributes fa; if (unit) cudaFuncGetAttributes(&fa, force_fast_kernel<I, true>); else cudaFuncGetAttributes(&fa, force_fast_kernel<I, false>);
        std::vector<float4> out(8); cudaMemcpy(out.data(), partial, 8 * sizeof(float4), cudaMemcpyDeviceToHost);
        printf("variant %d I=%d %s n=%d regs=%d: %.3f ms  %.4e int/s  %.1f%% of 74.45 TF  (check %.6e)\n", VARIANT, I, unit ? "unit" : "general", n, fa.numRegs, best,
               (double)n * n / (best * 1e-3), (double)n * n / (best * 1e-3) * 20 / 74.45e12 * 100, out[3].x);
        cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) printf("ERR %s\n", cudaGetErrorString(e));
    }
    return 0;
}
