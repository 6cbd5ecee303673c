// Alternative formulations of tile_interact for tools/exp/tune_fast.cu (selected by VARIANT).
// Same arithmetic as the product version; only statement order / grouping differs, to see how
// ptxas schedules them (operand reuse, MUFU spacing).
template <int I, bool kMaskSelf, bool kUnitMass>
__device__ __forceinline__ void tile_interact(const float4 *__restrict__ sA, const float4 *__restrict__ sB,
                                              const u64 (&xi)[I], const u64 (&yi)[I], const u64 (&zi)[I],
                                              u64 eps2, u64 (&ax)[I], u64 (&ay)[I], u64 (&az)[I],
                                              int tile_slot0, int self_slot0) {
#if VARIANT == 1 || VARIANT == 2 || VARIANT == 3 || VARIANT == 6
    // stage-major: every stage runs over all targets before the next stage starts
#if VARIANT == 3
#pragma unroll 1
#elif VARIANT == 6
#pragma unroll 4
#else
#pragma unroll 2
#endif
    for (int q = 0; q < kTileJ / 2; ++q) {
        const float4 a = sA[q];
        const float4 b = sB[q];
        const u64 xj = pack2(a.x, a.y), yj = pack2(a.z, a.w), zj = pack2(b.x, b.y), mj = pack2(b.z, b.w);
        u64 dx[I], dy[I], dz[I], r2[I], s[I];
#pragma unroll
        for (int k = 0; k < I; ++k) dx[k] = sub2(xj, xi[k]);
#pragma unroll
        for (int k = 0; k < I; ++k) dy[k] = sub2(yj, yi[k]);
#pragma unroll
        for (int k = 0; k < I; ++k) dz[k] = sub2(zj, zi[k]);
#pragma unroll
        for (int k = 0; k < I; ++k) r2[k] = fma2(dx[k], dx[k], eps2);
#pragma unroll
        for (int k = 0; k < I; ++k) r2[k] = fma2(dy[k], dy[k], r2[k]);
#pragma unroll
        for (int k = 0; k < I; ++k) r2[k] = fma2(dz[k], dz[k], r2[k]);
#pragma unroll
        for (int k = 0; k < I; ++k) {
            float lo, hi;
            unpack2(r2[k], lo, hi);
            r2[k] = pack2(rsqrt_approx(lo), rsqrt_approx(hi));
        }
#pragma unroll
        for (int k = 0; k < I; ++k) {
            const u64 rinv = r2[k];
            const u64 rinv2 = mul2(rinv, rinv);
            s[k] = kUnitMass ? mul2(rinv2, rinv) : mul2(mul2(mj, rinv), rinv2);
            if (kMaskSelf) {
                float slo, shi;
                unpack2(s[k], slo, shi);
                const int j0 = tile_slot0 + 2 * q, self = self_slot0 + k * kThreads;
                slo = (j0 == self) ? 0.0f : slo;
                shi = (j0 + 1 == self) ? 0.0f : shi;
                s[k] = pack2(slo, shi);
            }
        }
#pragma unroll
        for (int k = 0; k < I; ++k) {
#if VARIANT == 2
            ax[k] = fma2(s[k], dx[k], ax[k]);
            ay[k] = fma2(s[k], dy[k], ay[k]);
            az[k] = fma2(s[k], dz[k], az[k]);
#else
            ax[k] = fma2(dx[k], s[k], ax[k]);
            ay[k] = fma2(dy[k], s[k], ay[k]);
            az[k] = fma2(dz[k], s[k], az[k]);
#endif
        }
    }
#elif VARIANT == 4 || VARIANT == 5
    // software pipelined by hand: r^2 and rsqrt of pair q+1 are issued before the multiplies and
    // accumulations of pair q
    u64 dx[I], dy[I], dz[I], ri[I], mcur;
    {
        const float4 a = sA[0];
        const float4 b = sB[0];
        const u64 xj = pack2(a.x, a.y), yj = pack2(a.z, a.w), zj = pack2(b.x, b.y);
        mcur = pack2(b.z, b.w);
#pragma unroll
        for (int k = 0; k < I; ++k) {
            dx[k] = sub2(xj, xi[k]); dy[k] = sub2(yj, yi[k]); dz[k] = sub2(zj, zi[k]);
            u64 r2 = fma2(dx[k], dx[k], eps2); r2 = fma2(dy[k], dy[k], r2); r2 = fma2(dz[k], dz[k], r2);
            float lo, hi; unpack2(r2, lo, hi);
            ri[k] = pack2(rsqrt_approx(lo), rsqrt_approx(hi));
        }
    }
#if VARIANT == 5
#pragma unroll 2
#else
#pragma unroll 1
#endif
    for (int q = 0; q < kTileJ / 2; ++q) {
        const int qn = (q + 1 < kTileJ / 2) ? q + 1 : q;      // last iteration recomputes its own pair (discarded)
        const float4 a = sA[qn];
        const float4 b = sB[qn];
        const u64 xj = pack2(a.x, a.y), yj = pack2(a.z, a.w), zj = pack2(b.x, b.y), mnext = pack2(b.z, b.w);
        u64 ndx[I], ndy[I], ndz[I], nri[I];
#pragma unroll
        for (int k = 0; k < I; ++k) {
            ndx[k] = sub2(xj, xi[k]); ndy[k] = sub2(yj, yi[k]); ndz[k] = sub2(zj, zi[k]);
            u64 r2 = fma2(ndx[k], ndx[k], eps2); r2 = fma2(ndy[k], ndy[k], r2); r2 = fma2(ndz[k], ndz[k], r2);
            float lo, hi; unpack2(r2, lo, hi);
            nri[k] = pack2(rsqrt_approx(lo), rsqrt_approx(hi));
            // finish pair q for this target
            const u64 rinv = ri[k];
            const u64 rinv2 = mul2(rinv, rinv);
            u64 s = kUnitMass ? mul2(rinv2, rinv) : mul2(mul2(mcur, rinv), rinv2);
            if (kMaskSelf) {
                float slo, shi;
                unpack2(s, slo, shi);
                const int j0 = tile_slot0 + 2 * q, self = self_slot0 + k * kThreads;
                slo = (j0 == self) ? 0.0f : slo;
                shi = (j0 + 1 == self) ? 0.0f : shi;
                s = pack2(slo, shi);
            }
            ax[k] = fma2(dx[k], s, ax[k]);
            ay[k] = fma2(dy[k], s, ay[k]);
            az[k] = fma2(dz[k], s, az[k]);
        }
#pragma unroll
        for (int k = 0; k < I; ++k) { dx[k] = ndx[k]; dy[k] = ndy[k]; dz[k] = ndz[k]; ri[k] = nri[k]; }
        mcur = mnext;
    }
#elif VARIANT == 7 || VARIANT == 8 || VARIANT == 10 || VARIANT == 11
#if VARIANT == 7
#pragma unroll 4
#elif VARIANT == 8
#pragma unroll 3
#else
#pragma unroll 2
#endif
    for (int q = 0; q < kTileJ / 2; ++q) {
        const float4 a = sA[q];
        const float4 b = sB[q];
        const u64 xj = pack2(a.x, a.y), yj = pack2(a.z, a.w), zj = pack2(b.x, b.y), mj = pack2(b.z, b.w);
#pragma unroll
        for (int k = 0; k < I; ++k) {
            const u64 dx = sub2(xj, xi[k]), dy = sub2(yj, yi[k]), dz = sub2(zj, zi[k]);
#if VARIANT == 11
            u64 r2 = fma2(dz, dz, eps2);
            r2 = fma2(dy, dy, r2);
            r2 = fma2(dx, dx, r2);
#else
            u64 r2 = fma2(dx, dx, eps2);
            r2 = fma2(dy, dy, r2);
            r2 = fma2(dz, dz, r2);
#endif
            float r2lo, r2hi;
            unpack2(r2, r2lo, r2hi);
            const u64 rinv = pack2(rsqrt_approx(r2lo), rsqrt_approx(r2hi));
            const u64 rinv2 = mul2(rinv, rinv);
            u64 s = kUnitMass ? mul2(rinv2, rinv) : mul2(mul2(mj, rinv), rinv2);
            if (kMaskSelf) {
                float slo, shi;
                unpack2(s, slo, shi);
                const int j0 = tile_slot0 + 2 * q, self = self_slot0 + k * kThreads;
                slo = (j0 == self) ? 0.0f : slo;
                shi = (j0 + 1 == self) ? 0.0f : shi;
                s = pack2(slo, shi);
            }
#if VARIANT == 10
            az[k] = fma2(s, dz, az[k]);
            ay[k] = fma2(s, dy, ay[k]);
            ax[k] = fma2(s, dx, ax[k]);
#else
            ax[k] = fma2(dx, s, ax[k]);
            ay[k] = fma2(dy, s, ay[k]);
            az[k] = fma2(dz, s, az[k]);
#endif
        }
    }
#endif
}
