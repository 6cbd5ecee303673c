// Experiment (not product): does staging the source tiles with TMA bulk copies (cp.async.bulk ->
// UBLKCP, completion on an mbarrier) instead of LDG + STS change the force kernel's rate?
// Uniform-mass loop, I = 7 targets per thread, one CTA per SM, sources pre-packed in global memory
// in the shared-memory tile layout (4 KB per 256-source tile).
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pack2(float lo, float hi){ u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack2(u64 v, float& lo, float& hi){ asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 sub2(u64 a, u64 b){ u64 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b){ u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c){ u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float rsq(float x){ float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

__device__ __forceinline__ uint32_t smem_u32(const void* p){ return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count){ asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes){ asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity){
    asm volatile("{\n .reg .pred p;\n WAIT:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE;\n bra WAIT;\n DONE:\n}\n" :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar){
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

constexpr int I = 7, T = 256, TJ = 256;

template <bool kTMA>
__global__ void __launch_bounds__(T, 1) k_force(const float4* __restrict__ pos, const float4* __restrict__ packed, float4* __restrict__ out, int n, float eps2f)
{
    __shared__ alignas(128) float4 tile[2][TJ];     // [buf][A(128) | B(128)]
    __shared__ alignas(8) uint64_t full[2];
    const int tid = threadIdx.x;
    const int i0 = blockIdx.x * T * I;
    u64 xi[I], yi[I], zi[I], ax[I], ay[I], az[I];
    #pragma unroll
    for (int k=0;k<I;k++){ float4 p = pos[(i0 + k*T + tid) % n]; xi[k]=pack2(p.x,p.x); yi[k]=pack2(p.y,p.y); zi[k]=pack2(p.z,p.z); ax[k]=ay[k]=az[k]=0ull; }
    const u64 eps2 = pack2(eps2f, eps2f);
    const int ntiles = n / TJ;
    float *stA0 = reinterpret_cast<float*>(&tile[0][tid>>1]) + (tid&1);
    float *stB0 = reinterpret_cast<float*>(&tile[0][128 + (tid>>1)]) + (tid&1);
    if (kTMA) {
        if (tid == 0) { mbar_init(&full[0], 1); mbar_init(&full[1], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
        __syncthreads();
        if (tid == 0) {
            mbar_expect_tx(&full[0], 4096); bulk_g2s(&tile[0][0], packed, 4096, &full[0]);
            if (ntiles > 1) { mbar_expect_tx(&full[1], 4096); bulk_g2s(&tile[1][0], packed + TJ, 4096, &full[1]); }
        }
    } else {
        float4 p = pos[tid]; stA0[0]=p.x; stA0[2]=p.y; stB0[0]=p.z; stB0[2]=p.w;
        __syncthreads();
    }
    int buf = 0; uint32_t phase[2] = {0, 0};
    for (int jt=0; jt<ntiles; ++jt) {
        float4 nxt = make_float4(0,0,0,0);
        if (!kTMA && jt+1<ntiles) nxt = pos[(jt+1)*TJ + tid];
        if (kTMA) { mbar_wait(&full[buf], phase[buf]); phase[buf] ^= 1; }
        const float4* A = &tile[buf][0]; const float4* B = &tile[buf][128];
        #pragma unroll 2
        for (int q=0;q<TJ/2;q++){
            const float4 a=A[q], b=B[q];
            const u64 xj=pack2(a.x,a.y), yj=pack2(a.z,a.w), zj=pack2(b.x,b.y);
            #pragma unroll
            for (int k=0;k<I;k++){
                const u64 dx=sub2(xj,xi[k]), dy=sub2(yj,yi[k]), dz=sub2(zj,zi[k]);
                u64 r2=fma2(dx,dx,eps2); r2=fma2(dy,dy,r2); r2=fma2(dz,dz,r2);
                float lo,hi; unpack2(r2,lo,hi);
                const u64 ri=pack2(rsq(lo),rsq(hi));
                const u64 s=mul2(mul2(ri,ri),ri);
                ax[k]=fma2(dx,s,ax[k]); ay[k]=fma2(dy,s,ay[k]); az[k]=fma2(dz,s,az[k]);
            }
        }
        if (!kTMA && jt+1<ntiles){ float*dA=stA0+(buf^1)*TJ*4,*dB=stB0+(buf^1)*TJ*4; dA[0]=nxt.x; dA[2]=nxt.y; dB[0]=nxt.z; dB[2]=nxt.w; }
        __syncthreads();                       // everyone is done reading tile[buf]
        if (kTMA && tid == 0 && jt+2 < ntiles) { mbar_expect_tx(&full[buf], 4096); bulk_g2s(&tile[buf][0], packed + (size_t)(jt+2)*TJ, 4096, &full[buf]); }
        buf^=1;
    }
    #pragma unroll
    for (int k=0;k<I;k++){ float a0,a1,b0,b1,c0,c1; unpack2(ax[k],a0,a1); unpack2(ay[k],b0,b1); unpack2(az[k],c0,c1);
        out[(i0 + k*T + tid) % n] = make_float4(a0+a1, b0+b1, c0+c1, 0.f); }
}

template <bool kTMA> double run(const float4* pos, const float4* packed, float4* out, int n, int sm, float* check) {
    const int grid = sm * 2;
    cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_force<kTMA><<<grid,T>>>(pos,packed,out,n,0.01f); cudaDeviceSynchronize();
    float best=1e30f;
    for (int r=0;r<3;r++){ cudaEventRecord(e0); k_force<kTMA><<<grid,T>>>(pos,packed,out,n,0.01f); cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms,e0,e1); if (ms<best) best=ms; }
    float4 h; cudaMemcpy(&h, out + 5, sizeof(h), cudaMemcpyDeviceToHost); *check = h.x;
    cudaError_t e = cudaGetLastError(); if (e!=cudaSuccess) printf("ERR %s\n", cudaGetErrorString(e));
    return (double)grid*T*I*(double)n/(best*1e-3);
}
int main(int argc, char** argv){
    int n = argc>1 ? atoi(argv[1]) : 65536;
    int sm; cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0);
    std::vector<float4> h(n), pk(n);
    srand(1); for (auto& p: h){ p.x=rand()/(float)RAND_MAX*100; p.y=rand()/(float)RAND_MAX*100; p.z=rand()/(float)RAND_MAX*100; p.w=1.f; }
    for (int t=0;t<n/TJ;t++) for (int q=0;q<128;q++){ const float4 a=h[t*TJ+2*q], b=h[t*TJ+2*q+1];
        pk[t*TJ+q]=make_float4(a.x,b.x,a.y,b.y); pk[t*TJ+128+q]=make_float4(a.z,b.z,a.w,b.w); }
    float4 *pos,*packed,*out; cudaMalloc(&pos,n*sizeof(float4)); cudaMalloc(&packed,n*sizeof(float4)); cudaMalloc(&out,n*sizeof(float4));
    cudaMemcpy(pos,h.data(),n*sizeof(float4),cudaMemcpyHostToDevice); cudaMemcpy(packed,pk.data(),n*sizeof(float4),cudaMemcpyHostToDevice);
    float c0, c1;
    double r0 = run<false>(pos,packed,out,n,sm,&c0), r1 = run<true>(pos,packed,out,n,sm,&c1);
    printf("LDG+STS staging : %.4e int/s (%.1f%% of 74.45 TF)  check %.6e\n", r0, r0*20/74.45e12*100, c0);
    printf("TMA bulk staging: %.4e int/s (%.1f%% of 74.45 TF)  check %.6e\n", r1, r1*20/74.45e12*100, c1);
    return 0;
}
