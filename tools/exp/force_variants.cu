// Experiment harness (not product): throughput of force-kernel variants on one GPU.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o force_variants force_variants.cu
#include <cuda_runtime.h>
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

// simple (non-persistent) kernel: each CTA takes one i-tile and all j; for throughput experiments only
template <int I, int UNROLL, int THREADS, int MINB, int MODE>
__global__ void __launch_bounds__(THREADS, MINB) k_force(const float4* __restrict__ pos, float4* __restrict__ out, int n, float eps2f)
{
    constexpr int TJ = THREADS;
    __shared__ float4 sA[2][TJ/2], sB[2][TJ/2];
    const int tid = threadIdx.x;
    const int i0 = blockIdx.x * THREADS * I;
    float *stA0 = reinterpret_cast<float*>(&sA[0][tid>>1]) + (tid&1);
    float *stB0 = reinterpret_cast<float*>(&sB[0][tid>>1]) + (tid&1);
    constexpr int BF = (TJ/2)*4;
    float xs[I], ys[I], zs[I];
    u64 xi[I], yi[I], zi[I], ax[I], ay[I], az[I];
    #pragma unroll
    for (int k=0;k<I;k++){ float4 p = pos[(i0 + k*THREADS + tid) % n]; xs[k]=p.x; ys[k]=p.y; zs[k]=p.z;
        xi[k]=pack2(p.x,p.x); yi[k]=pack2(p.y,p.y); zi[k]=pack2(p.z,p.z); ax[k]=ay[k]=az[k]=0ull; }
    float fx[I], fy[I], fz[I];
    #pragma unroll
    for (int k=0;k<I;k++) fx[k]=fy[k]=fz[k]=0.f;
    const u64 eps2 = pack2(eps2f, eps2f);
    { float4 p = pos[tid]; stA0[0]=p.x; stA0[2]=p.y; stB0[0]=p.z; stB0[2]=p.w; }
    __syncthreads();
    int buf = 0;
    const int ntiles = n / TJ;
    for (int jt=0; jt<ntiles; ++jt) {
        float4 nxt = make_float4(0,0,0,0);
        if (jt+1<ntiles) nxt = pos[(jt+1)*TJ + tid];
        const float4* A = sA[buf]; const float4* B = sB[buf];
        if (MODE == 0 || MODE == 2) {            // packed j-pairs (MODE 2: unit mass)
            #pragma unroll UNROLL
            for (int q=0;q<TJ/2;q++){
                const float4 a=A[q], b=B[q];
                const u64 xj=pack2(a.x,a.y), yj=pack2(a.z,a.w), zj=pack2(b.x,b.y), mj=pack2(b.z,b.w);
                #pragma unroll
                for (int k=0;k<I;k++){
                    const u64 dx=sub2(xj,xi[k]), dy=sub2(yj,yi[k]), dz=sub2(zj,zi[k]);
                    u64 r2=fma2(dx,dx,eps2); r2=fma2(dy,dy,r2); r2=fma2(dz,dz,r2);
                    float lo,hi; unpack2(r2,lo,hi);
                    const u64 ri=pack2(rsq(lo),rsq(hi));
                    const u64 ri2=mul2(ri,ri);
                    const u64 s = (MODE == 2) ? mul2(ri2, ri) : mul2(mul2(mj,ri),ri2);
                    ax[k]=fma2(dx,s,ax[k]); ay[k]=fma2(dy,s,ay[k]); az[k]=fma2(dz,s,az[k]);
                }
            }
        } else {                    // scalar
            #pragma unroll UNROLL
            for (int q=0;q<TJ/2;q++){
                const float4 a=A[q], b=B[q];
                #pragma unroll
                for (int h=0;h<2;h++){
                    const float xj = h? a.y:a.x, yj = h? a.w:a.z, zj = h? b.y:b.x, mj = h? b.w:b.z;
                    #pragma unroll
                    for (int k=0;k<I;k++){
                        const float dx=xj-xs[k], dy=yj-ys[k], dz=zj-zs[k];
                        float r2=fmaf(dx,dx,eps2f); r2=fmaf(dy,dy,r2); r2=fmaf(dz,dz,r2);
                        const float ri=rsq(r2); const float ri2=ri*ri; const float s=(mj*ri)*ri2;
                        fx[k]=fmaf(dx,s,fx[k]); fy[k]=fmaf(dy,s,fy[k]); fz[k]=fmaf(dz,s,fz[k]);
                    }
                }
            }
        }
        if (jt+1<ntiles){ float*dA=stA0+(buf^1)*BF,*dB=stB0+(buf^1)*BF; dA[0]=nxt.x; dA[2]=nxt.y; dB[0]=nxt.z; dB[2]=nxt.w; }
        __syncthreads();
        buf^=1;
    }
    #pragma unroll
    for (int k=0;k<I;k++){
        float a0,a1,b0,b1,c0,c1; unpack2(ax[k],a0,a1); unpack2(ay[k],b0,b1); unpack2(az[k],c0,c1);
        out[(i0 + k*THREADS + tid) % n] = make_float4(a0+a1+fx[k], b0+b1+fy[k], c0+c1+fz[k], 0.f);
    }
}

static int g_dyn = 0;
template <int I, int UNROLL, int THREADS, int MINB, int MODE>
void run(const char* name, const float4* pos, float4* out, int n, int sm) {
    auto kern = k_force<I,UNROLL,THREADS,MINB,MODE>;
    if (g_dyn) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, g_dyn);
    int occ=0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, THREADS, g_dyn);
    cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, kern);
    // choose number of i-tiles = exactly sm*occ*w so that there is no wave quantisation: wrap indices mod n
    int grid = sm * occ * 2;
    cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    kern<<<grid,THREADS,g_dyn>>>(pos,out,n,0.01f);
    cudaDeviceSynchronize();
    float best=1e30f;
    for (int r=0;r<3;r++){ cudaEventRecord(e0); kern<<<grid,THREADS,g_dyn>>>(pos,out,n,0.01f); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms,e0,e1); if (ms<best) best=ms; }
    double inter = (double)grid*THREADS*I*(double)n;
    printf("%-28s I=%d U=%d T=%d minb=%d regs=%d occ=%d grid=%d  %.3f ms  %.3e int/s  %.1f TF20\n", name, I,UNROLL,THREADS,MINB, fa.numRegs, occ, grid, best, inter/(best*1e-3), inter/(best*1e-3)*20/1e12);
    cudaError_t e = cudaGetLastError(); if (e!=cudaSuccess) printf("ERR %s\n", cudaGetErrorString(e));
}

int main(int argc, char** argv){
    int n = argc>1 ? atoi(argv[1]) : 65536;
    int sm; cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0);
    std::vector<float4> h(n);
    srand(1); for (auto& p: h){ p.x=rand()/(float)RAND_MAX*100; p.y=rand()/(float)RAND_MAX*100; p.z=rand()/(float)RAND_MAX*100; p.w=1.f; }
    float4 *pos,*out; cudaMalloc(&pos,n*sizeof(float4)); cudaMalloc(&out,n*sizeof(float4));
    cudaMemcpy(pos,h.data(),n*sizeof(float4),cudaMemcpyHostToDevice);
    g_dyn = 0;
    run<7,2,256,1,2>("unit 2w/smsp", pos,out,n,sm);
    g_dyn = 120*1024;
    run<7,2,128,1,2>("unit 1w/smsp", pos,out,n,sm);
    run<7,4,128,1,2>("unit 1w/smsp", pos,out,n,sm);
    run<8,2,128,1,2>("unit 1w/smsp", pos,out,n,sm);
    run<10,2,128,1,2>("unit 1w/smsp", pos,out,n,sm);
    run<12,1,128,1,2>("unit 1w/smsp", pos,out,n,sm);
    run<12,2,128,1,2>("unit 1w/smsp", pos,out,n,sm);
    run<7,2,128,1,0>("general 1w/smsp", pos,out,n,sm);
    g_dyn = 0;
    return 0;
}
