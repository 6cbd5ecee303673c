/*
 * nbody_oracle.c -- CPU restatement of the reference's direct-sum hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (the CUDA library, the
 * host classes, bench.py's GPU arm) may call into this file; only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * use it, and there only as the checker or the timed CPU baseline.
 *
 * Parity status: PINNED.  tests/test_oracle_vs_reference.py compares every
 * function below, bit for bit, with the reference's own NaiveSimulation /
 * MpiNaiveSimulation objects compiled unmodified into oracle/_ref/ (see
 * oracle/Makefile), and tests/golden/ holds vectors produced by that build.
 * The reference has no tests or golden vectors of its own.
 *
 * All citations are relative to /root/reference/.  Arithmetic is IEEE
 * binary32 with no contraction: build with -O2 -ffp-contract=off and no
 * -march / -ffast-math so the bits do not depend on the host.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* 40-byte AoS record: include/Particle.hpp:53-59 (three vector3_t then mass). */
typedef struct {
    float px, py, pz;
    float vx, vy, vz;
    float ax, ay, az;
    float m;
} oracle_particle;

int oracle_sizeof_particle(void) { return (int)sizeof(oracle_particle); }

/* ------------------------------------------------------------------------ */
/* A. The reference's literal loop order: half matrix j>i with the symmetric */
/*    update.  src/NaiveSimulation.cpp:12-41.                                */
/* ------------------------------------------------------------------------ */
void oracle_naive_accel_pairs_f32(oracle_particle *p, size_t n, float G, float eps2)
{
    for (size_t i = 0; i < n; ++i) {            /* NaiveSimulation.cpp:18-19 */
        p[i].ax = 0.0f; p[i].ay = 0.0f; p[i].az = 0.0f;
    }
    for (size_t i = 0; i < n; ++i) {            /* :21 */
        for (size_t j = i + 1; j < n; ++j) {    /* :25 */
            float rx = p[j].px - p[i].px;       /* :28  r = b.pos - a.pos */
            float ry = p[j].py - p[i].py;
            float rz = p[j].pz - p[i].pz;
            float r2 = rx * rx + ry * ry + rz * rz + eps2;     /* :29 left to right */
            float inv_r = 1.0f / sqrtf(r2);                    /* :31 */
            float inv_r3 = inv_r * inv_r * inv_r;              /* :32 */
            float c = G * inv_r3;                              /* :34 r * (G*inv_r3) */
            float ux = rx * c, uy = ry * c, uz = rz * c;
            float mi = p[i].m, mj = p[j].m;
            p[i].ax += ux * mj; p[i].ay += uy * mj; p[i].az += uz * mj;   /* :35,37 */
            p[j].ax -= ux * mi; p[j].ay -= uy * mi; p[j].az -= uz * mi;   /* :36,38 */
        }
    }
}

/* ------------------------------------------------------------------------ */
/* B. Row order: one accumulator per target i, sources j ascending, j != i.  */
/*    Bit-identical to A (checked in tests) because            */
/*    (p_i-p_j) == -(p_j-p_i) exactly and x -= y == x += (-y) in IEEE RN.    */
/*    This is the order the CUDA "strict" kernel follows, and it lets rows   */
/*    [i0,i1) be computed independently (threads, GPU slices).               */
/* ------------------------------------------------------------------------ */
void oracle_naive_accel_rows_f32(const oracle_particle *p, size_t n, size_t i0, size_t i1,
                                 float G, float eps2, float *acc_out /* 3*(i1-i0) */)
{
    for (size_t i = i0; i < i1; ++i) {
        float ax = 0.0f, ay = 0.0f, az = 0.0f;
        const float xi = p[i].px, yi = p[i].py, zi = p[i].pz;
        for (size_t j = 0; j < n; ++j) {
            if (j == i) continue;
            float rx = p[j].px - xi, ry = p[j].py - yi, rz = p[j].pz - zi;
            float r2 = rx * rx + ry * ry + rz * rz + eps2;
            float inv_r = 1.0f / sqrtf(r2);
            float inv_r3 = inv_r * inv_r * inv_r;
            float c = G * inv_r3;
            float mj = p[j].m;
            ax += (rx * c) * mj; ay += (ry * c) * mj; az += (rz * c) * mj;
        }
        acc_out[3 * (i - i0) + 0] = ax;
        acc_out[3 * (i - i0) + 1] = ay;
        acc_out[3 * (i - i0) + 2] = az;
    }
}

/* Semi-implicit Euler: src/NaiveSimulation.cpp:43-51 (v += a*dt; x += v*dt,  */
/* product and sum rounded separately).  Same as MpiNaiveSimulation.cpp:127-141. */
void oracle_integrate_f32(oracle_particle *p, size_t n, float dt)
{
    for (size_t i = 0; i < n; ++i) {
        p[i].vx += p[i].ax * dt; p[i].vy += p[i].ay * dt; p[i].vz += p[i].az * dt;
        p[i].px += p[i].vx * dt; p[i].py += p[i].vy * dt; p[i].pz += p[i].vz * dt;
    }
}

/* NaiveSimulation::computeStep (src/NaiveSimulation.cpp:7-10), literal order. */
void oracle_naive_step_pairs_f32(oracle_particle *p, size_t n, float G, float dt, float eps2, int steps)
{
    for (int s = 0; s < steps; ++s) {
        oracle_naive_accel_pairs_f32(p, n, G, eps2);
        oracle_integrate_f32(p, n, dt);
    }
}

/* Same step through the row-order restatement; rows are split over OpenMP   */
/* threads when built with -fopenmp (the result does not depend on the split). */
void oracle_naive_step_rows_f32(oracle_particle *p, size_t n, float G, float dt, float eps2, int steps)
{
    float *acc = (float *)malloc(sizeof(float) * 3 * (n ? n : 1));
    for (int s = 0; s < steps; ++s) {
        const long chunk = 64;
        const long nchunks = ((long)n + chunk - 1) / chunk;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1)
#endif
        for (long c = 0; c < nchunks; ++c) {
            size_t a = (size_t)(c * chunk), b = a + (size_t)chunk;
            if (b > n) b = n;
            oracle_naive_accel_rows_f32(p, n, a, b, G, eps2, acc + 3 * a);
        }
        for (size_t i = 0; i < n; ++i) {
            p[i].ax = acc[3 * i]; p[i].ay = acc[3 * i + 1]; p[i].az = acc[3 * i + 2];
        }
        oracle_integrate_f32(p, n, dt);
    }
    free(acc);
}

/* ------------------------------------------------------------------------ */
/* C. MpiNaiveSimulation's arithmetic for P ranks, restated without MPI.     */
/*    src/MpiNaiveSimulation.cpp:15-125: rank r owns a contiguous slice      */
/*    (counts as src/apps/headless_main.cpp:249-259), sums its own block     */
/*    first (j ascending, skip i, :62-81) and then the blocks of ranks       */
/*    r-1, r-2, ... (:56, :84-102).  Grouping differs from A/B:              */
/*    s = G*m_j*inv_r3; acc += s*r  (:76-79, :97-100).                        */
/* ------------------------------------------------------------------------ */
void oracle_slice_bounds(size_t n, int P, int r, size_t *begin, size_t *count)
{
    size_t base = n / (size_t)P, rem = n % (size_t)P;   /* headless_main.cpp:249-259 */
    size_t b = 0;
    for (int k = 0; k < r; ++k) b += base + ((size_t)k < rem ? 1 : 0);
    *begin = b;
    *count = base + ((size_t)r < rem ? 1 : 0);
}

void oracle_mpi_naive_accel_f32(oracle_particle *p, size_t n, int P, float G, float eps2)
{
    for (int r = 0; r < P; ++r) {
        size_t ib, ic;
        oracle_slice_bounds(n, P, r, &ib, &ic);
        for (size_t i = ib; i < ib + ic; ++i) {
            float ax = 0.0f, ay = 0.0f, az = 0.0f;
            for (int step = 0; step < P; ++step) {
                int owner = (r - step + P) % P;             /* :56 */
                size_t jb, jc;
                oracle_slice_bounds(n, P, owner, &jb, &jc);
                for (size_t j = jb; j < jb + jc; ++j) {
                    if (j == i) continue;                   /* :64 (own block only) */
                    float rx = p[j].px - p[i].px, ry = p[j].py - p[i].py, rz = p[j].pz - p[i].pz;
                    float r2 = rx * rx + ry * ry + rz * rz + eps2;
                    float inv_r = 1.0f / sqrtf(r2);
                    float inv_r3 = inv_r * inv_r * inv_r;
                    float s = G * p[j].m * inv_r3;          /* :76 */
                    ax += s * rx; ay += s * ry; az += s * rz;
                }
            }
            p[i].ax = ax; p[i].ay = ay; p[i].az = az;
        }
    }
}

void oracle_mpi_naive_step_f32(oracle_particle *p, size_t n, int P, float G, float dt, float eps2, int steps)
{
    /* positions of step s must all be read before any is advanced */
    for (int s = 0; s < steps; ++s) {
        oracle_mpi_naive_accel_f32(p, n, P, G, eps2);
        oracle_integrate_f32(p, n, dt);
    }
}

/* ------------------------------------------------------------------------ */
/* D. FP64 "truth": same formula evaluated in double from the float inputs.  */
/*    Not in the reference (it has no double anywhere); used   */
/*    to state how far the reference's own FP32 sum is from the exact sum,   */
/*    which bounds what a re-ordered FP32 kernel can be asked to match.      */
/* ------------------------------------------------------------------------ */
void oracle_accel_f64(const oracle_particle *p, size_t n, size_t i0, size_t i1,
                      double G, double eps2, double *acc_out /* 3*(i1-i0) */)
{
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
    for (long ii = (long)i0; ii < (long)i1; ++ii) {
        size_t i = (size_t)ii;
        double ax = 0.0, ay = 0.0, az = 0.0;
        const double xi = p[i].px, yi = p[i].py, zi = p[i].pz;
        for (size_t j = 0; j < n; ++j) {
            if (j == i) continue;
            double rx = (double)p[j].px - xi, ry = (double)p[j].py - yi, rz = (double)p[j].pz - zi;
            double r2 = rx * rx + ry * ry + rz * rz + eps2;
            double inv_r = 1.0 / sqrt(r2);
            double s = G * (double)p[j].m * inv_r * inv_r * inv_r;
            ax += s * rx; ay += s * ry; az += s * rz;
        }
        acc_out[3 * (i - i0) + 0] = ax;
        acc_out[3 * (i - i0) + 1] = ay;
        acc_out[3 * (i - i0) + 2] = az;
    }
}

/* FP64 trajectory: state arrays of doubles, same integrator as B.           */
void oracle_step_f64(double *pos, double *vel, double *acc, const double *mass, size_t n,
                     double G, double dt, double eps2, int steps)
{
    for (int s = 0; s < steps; ++s) {
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
        for (long ii = 0; ii < (long)n; ++ii) {
            size_t i = (size_t)ii;
            double ax = 0.0, ay = 0.0, az = 0.0;
            for (size_t j = 0; j < n; ++j) {
                if (j == i) continue;
                double rx = pos[3 * j] - pos[3 * i], ry = pos[3 * j + 1] - pos[3 * i + 1],
                       rz = pos[3 * j + 2] - pos[3 * i + 2];
                double r2 = rx * rx + ry * ry + rz * rz + eps2;
                double inv_r = 1.0 / sqrt(r2);
                double sc = G * mass[j] * inv_r * inv_r * inv_r;
                ax += sc * rx; ay += sc * ry; az += sc * rz;
            }
            acc[3 * i] = ax; acc[3 * i + 1] = ay; acc[3 * i + 2] = az;
        }
        for (size_t i = 0; i < 3 * n; ++i) { vel[i] += acc[i] * dt; pos[i] += vel[i] * dt; }
    }
}

/* Energy diagnostic (new: the reference has no energy code).  */
/* K = sum 1/2 m v^2 ; U = - sum_{i<j} G m_i m_j / sqrt(r^2 + eps2), the      */
/* potential whose gradient is the softened force of NaiveSimulation.cpp:29-34. */
void oracle_energy_f64(const oracle_particle *p, size_t n, double G, double eps2,
                       double *kinetic, double *potential)
{
    double K = 0.0, U = 0.0;
    for (size_t i = 0; i < n; ++i)
        K += 0.5 * (double)p[i].m *
             ((double)p[i].vx * p[i].vx + (double)p[i].vy * p[i].vy + (double)p[i].vz * p[i].vz);
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : U)
#endif
    for (long ii = 0; ii < (long)n; ++ii) {
        size_t i = (size_t)ii;
        double u = 0.0;
        for (size_t j = i + 1; j < n; ++j) {
            double rx = (double)p[j].px - p[i].px, ry = (double)p[j].py - p[i].py,
                   rz = (double)p[j].pz - p[i].pz;
            u += (double)p[j].m / sqrt(rx * rx + ry * ry + rz * rz + eps2);
        }
        U -= G * (double)p[i].m * u;
    }
    *kinetic = K;
    *potential = U;
}
