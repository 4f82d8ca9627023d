/* oracle/nbody_oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of the reference's all-pairs gravity + integrator path. It is the CHECKER for the
 * CUDA product path: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference`
 * leg may load it. The product library never links or calls it (there is no CPU fallback).
 *
 * Parity pin: this restatement is pinned (tests/test_oracle.py) against
 *   (1) the known-answer vectors produced from the unmodified reference sources (SURVEY.md §8c, KAT-1..3,
 *       committed under tests/golden/), and
 *   (2) oracle/_ref/libnbref.so — the reference's own Forces.cpp/System.cpp compiled unmodified —
 *       whenever that file is present.
 * The reference itself ships no tests or golden vectors for this path.
 *
 * Layout (same as the reference's Particles, simulator/core/Particles.hpp:9-16): posMass4[i] = (x,y,z,m),
 * vel4[i] = (vx,vy,vz,0), doubles, AoS. Units AU, Msun, yr. Compiled with -fno-fast-math -ffp-contract=off
 * so every operation below is the IEEE operation written.
 */
#include "nbody_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* simulator/core/Forces.cpp:14, System.cpp:42, cuda/Constants.cuh:6 */
static const double NBO_G = 39.47841760435743;
/* simulator/core/Forces.cpp:305-306 (eta, epsilon_min); cuda/Constants.cuh:7-8 */
static const double NBO_ETA = 0.01;
static const double NBO_EPS_MIN = 1e-4;
/* simulator/core/Forces.cpp:144,350: pairs with r^2 < 1e-10 are skipped */
static const double NBO_R2_CUT = 1e-10;
/* simulator/core/Forces.cpp:104,311: 64 x 64 tiles; the per-tile partial is summed first, then added */
#define NBO_TILE 64

void nbo_set_threads(int t) {
#ifdef _OPENMP
    if (t > 0) omp_set_num_threads(t);
#else
    (void)t;
#endif
}
int nbo_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* computeAverageMass — simulator/core/Forces.cpp:85-95 (returns 1.0 for n == 0, serial sum in index order) */
double nbo_average_mass(long n, const double* pm) {
    if (n == 0) return 1.0;
    double total = 0.0;
    for (long i = 0; i < n; ++i) total += pm[4 * i + 3];
    return total / (double)n;
}

/* calculatePairwiseAccelOpt — simulator/cuda/ForceKernels.cu:173-204: what forceMethod PAIRWISE_AVX2_FP32 runs on the
 * reference's GPU path (Plummer softening r^2 + 1e-10, no cut-off, self term included and equal to zero, float rsqrt
 * refined by one Newton step in double). The CPU twin (Forces.cpp:202-296) drops tile tails and keeps the unrefined
 * 12-bit rsqrt, so the GPU kernel is the parity target. Serial j order as in the kernel. */
static void plummer_row(const double* pm, long n, long i, double out[3]) {
    const double ix = pm[4 * i], iy = pm[4 * i + 1], iz = pm[4 * i + 2];
    double ax = 0.0, ay = 0.0, az = 0.0;
    for (long j = 0; j < n; ++j) {
        const double dx = pm[4 * j] - ix, dy = pm[4 * j + 1] - iy, dz = pm[4 * j + 2] - iz;
        const double r2 = dx * dx + dy * dy + dz * dz + 1e-10;                 /* :186 */
        if (r2 > 0.0) {                                                        /* :188 */
            const float rinv_f = 1.0f / sqrtf((float)r2);                      /* :189 rsqrtf((float)r2) */
            double rinv = (double)rinv_f;
            rinv = rinv * (1.5 - 0.5 * r2 * rinv * rinv);                      /* :191 Newton step */
            const double f = NBO_G * pm[4 * j + 3] * (rinv * rinv * rinv);     /* :193-194 */
            ax += f * dx; ay += f * dy; az += f * dz;                          /* :196-198 */
        }
    }
    out[0] = ax; out[1] = ay; out[2] = az;
}

/* One i-particle against the j-range [j0,j1): the inner loops of
 *   pairwiseAcceleration               simulator/core/Forces.cpp:126-160
 *   pairwiseAccelerationAdaptiveMutual simulator/core/Forces.cpp:332-377
 * Returns the tile partial (ax,ay,az) exactly as the reference forms it before adding to localAccel. */
static void tile_partial(const double* pm, long i, long j0, long j1, int adaptive, double avg_mass, double out[3]) {
    const double ix = pm[4 * i], iy = pm[4 * i + 1], iz = pm[4 * i + 2], im = pm[4 * i + 3];
    double ax = 0.0, ay = 0.0, az = 0.0;
    for (long j = j0; j < j1; ++j) {
        if (i == j) continue;                                   /* Forces.cpp:135,341 */
        const double dx = pm[4 * j] - ix, dy = pm[4 * j + 1] - iy, dz = pm[4 * j + 2] - iz;
        const double r2 = dx * dx + dy * dy + dz * dz;          /* Forces.cpp:142,348 */
        if (r2 < NBO_R2_CUT) continue;                          /* Forces.cpp:144,350 */
        const double dist = sqrt(r2);
        double factor;
        if (!adaptive) {
            factor = (NBO_G * pm[4 * j + 3]) / (dist * r2);     /* Forces.cpp:146-148 */
        } else {
            const double combined = im + pm[4 * j + 3];         /* Forces.cpp:353 */
            double eps = NBO_ETA * dist * cbrt(combined / (3.0 * avg_mass));   /* :354 */
            eps = eps > NBO_EPS_MIN ? eps : NBO_EPS_MIN;        /* :355 */
            const double s2 = r2 + eps * eps;                   /* :356 */
            const double s = sqrt(s2);                          /* :357 */
            factor = (NBO_G * pm[4 * j + 3]) / (s * s2);        /* :360 */
        }
        ax += factor * dx; ay += factor * dy; az += factor * dz; /* Forces.cpp:151-153,362-364 */
    }
    out[0] = ax; out[1] = ay; out[2] = az;
}

/* calculateForces dispatch — simulator/core/Forces.cpp:67-83. Only PAIRWISE (0) and ADAPTIVE_MUTUAL (2) are on
 * the hot path, PAIRWISE_AVX2_FP32 (1) follows the reference GPU kernel (plummer_row); BARNES_HUT returns non-zero. accel3 is n x (ax,ay,az) (Accelerations, Forces.hpp:5-37).
 * Each i is owned by one thread (as in the reference's omp-for over i-tiles) so the result does not depend on
 * the thread count. */
int nbo_forces(long n, const double* pm, int method, double* accel3) {
    if (method == NBO_FORCE_PAIRWISE_AVX2_FP32) {
        #pragma omp parallel for schedule(dynamic, 16)
        for (long i = 0; i < n; ++i) plummer_row(pm, n, i, accel3 + 3 * i);
        return 0;
    }
    if (method != NBO_FORCE_PAIRWISE && method != NBO_FORCE_ADAPTIVE_MUTUAL) return 1;
    const int adaptive = (method == NBO_FORCE_ADAPTIVE_MUTUAL);
    const double avg = adaptive ? nbo_average_mass(n, pm) : 0.0;   /* Forces.cpp:309 */
    #pragma omp parallel for schedule(dynamic, 1)
    for (long it = 0; it < n; it += NBO_TILE) {
        const long iend = it + NBO_TILE < n ? it + NBO_TILE : n;
        for (long i = it; i < iend; ++i) { accel3[3 * i] = accel3[3 * i + 1] = accel3[3 * i + 2] = 0.0; }
        for (long jt = 0; jt < n; jt += NBO_TILE) {
            const long jend = jt + NBO_TILE < n ? jt + NBO_TILE : n;
            for (long i = it; i < iend; ++i) {
                double part[3];
                tile_partial(pm, i, jt, jend, adaptive, avg, part);
                accel3[3 * i] += part[0]; accel3[3 * i + 1] += part[1]; accel3[3 * i + 2] += part[2];
            }
        }
    }
    return 0;
}

/* Sampled-subset arbiter for N too large for the full O(N^2) CPU pass (SURVEY.md §8c last row): the same
 * per-pair arithmetic as tile_partial, accumulated in long double, for the i-indices listed in idx. */
int nbo_forces_subset(long n, const double* pm, int method, long n_idx, const long* idx, double* accel3) {
    if (method == NBO_FORCE_PAIRWISE_AVX2_FP32) {
        #pragma omp parallel for schedule(dynamic, 1)
        for (long k = 0; k < n_idx; ++k) plummer_row(pm, n, idx[k], accel3 + 3 * k);
        return 0;
    }
    if (method != NBO_FORCE_PAIRWISE && method != NBO_FORCE_ADAPTIVE_MUTUAL) return 1;
    const int adaptive = (method == NBO_FORCE_ADAPTIVE_MUTUAL);
    const double avg = adaptive ? nbo_average_mass(n, pm) : 0.0;
    #pragma omp parallel for schedule(dynamic, 1)
    for (long k = 0; k < n_idx; ++k) {
        const long i = idx[k];
        const double ix = pm[4 * i], iy = pm[4 * i + 1], iz = pm[4 * i + 2], im = pm[4 * i + 3];
        long double ax = 0.0L, ay = 0.0L, az = 0.0L;
        for (long j = 0; j < n; ++j) {
            if (j == i) continue;
            const double dx = pm[4 * j] - ix, dy = pm[4 * j + 1] - iy, dz = pm[4 * j + 2] - iz;
            const double r2 = dx * dx + dy * dy + dz * dz;
            if (r2 < NBO_R2_CUT) continue;
            const double dist = sqrt(r2);
            double factor;
            if (!adaptive) factor = (NBO_G * pm[4 * j + 3]) / (dist * r2);
            else {
                double eps = NBO_ETA * dist * cbrt((im + pm[4 * j + 3]) / (3.0 * avg));
                eps = eps > NBO_EPS_MIN ? eps : NBO_EPS_MIN;
                const double s2 = r2 + eps * eps;
                factor = (NBO_G * pm[4 * j + 3]) / (sqrt(s2) * s2);
            }
            ax += (long double)(factor * dx); ay += (long double)(factor * dy); az += (long double)(factor * dz);
        }
        accel3[3 * k] = (double)ax; accel3[3 * k + 1] = (double)ay; accel3[3 * k + 2] = (double)az;
    }
    return 0;
}

/* centerOfMassVelocity — simulator/core/System.cpp:45-57 (serial, index order) */
void nbo_com_velocity(long n, const double* pm, const double* v, double vcm[3]) {
    double M = 0.0, x = 0.0, y = 0.0, z = 0.0;
    for (long i = 0; i < n; ++i) M += pm[4 * i + 3];
    for (long i = 0; i < n; ++i) {
        x += pm[4 * i + 3] * v[4 * i]; y += pm[4 * i + 3] * v[4 * i + 1]; z += pm[4 * i + 3] * v[4 * i + 2];
    }
    vcm[0] = x / M; vcm[1] = y / M; vcm[2] = z / M;
}

/* System::velocityVerlet — simulator/core/System.cpp:432-489 ; System::eulerIntegration — :385-429.
 * One call = `steps` calls of System::performIntegrationStep (:492-499), i.e. System::runBenchmark in CPU mode
 * (:168-194). In place. */
int nbo_step(long n, double* pm, double* v, int integ, int method, double dt, int steps) {
    if (integ != NBO_INTEG_EULER && integ != NBO_INTEG_VELOCITY_VERLET) return 1;
    double* a = (double*)malloc(sizeof(double) * 3 * (size_t)(n > 0 ? n : 1));
    double* a2 = (double*)malloc(sizeof(double) * 3 * (size_t)(n > 0 ? n : 1));
    if (!a || !a2) { free(a); free(a2); return 2; }
    int rc = 0;
    for (int s = 0; s < steps && rc == 0; ++s) {
        double vcm[3];
        nbo_com_velocity(n, pm, v, vcm);                                   /* :387,434 */
        for (long i = 0; i < n; ++i) { v[4*i] -= vcm[0]; v[4*i+1] -= vcm[1]; v[4*i+2] -= vcm[2]; } /* :392-400,438-446 */
        rc = nbo_forces(n, pm, method, a);                                 /* :403,449 */
        if (rc) break;
        if (integ == NBO_INTEG_EULER) {
            for (long i = 0; i < n; ++i)
                for (int c = 0; c < 3; ++c) {
                    pm[4*i+c] += v[4*i+c] * dt;                            /* :417-419 (old v) */
                    v[4*i+c]  += a[3*i+c] * dt;                            /* :422-424 */
                }
        } else {
            for (long i = 0; i < n; ++i)
                for (int c = 0; c < 3; ++c)
                    pm[4*i+c] += v[4*i+c] * dt + 0.5 * a[3*i+c] * (dt * dt);   /* :463-465 */
            rc = nbo_forces(n, pm, method, a2);                            /* :470 */
            if (rc) break;
            for (long i = 0; i < n; ++i)
                for (int c = 0; c < 3; ++c)
                    v[4*i+c] += 0.5 * (a[3*i+c] + a2[3*i+c]) * dt;         /* :484-486 */
        }
    }
    free(a); free(a2);
    return rc;
}

/* System::computeTotalEnergy — simulator/core/System.cpp:60-85: kinetic sum, then unsoftened pair potential
 * over i<j with the `dist > 0` guard. The pair sum is row-parallel here (each row summed in j order, rows
 * added in i order), which reassociates the reference's single running sum; the difference is O(1e-16) rel. */
double nbo_energy(long n, const double* pm, const double* v) {
    double kin = 0.0;
    for (long i = 0; i < n; ++i) {
        const double v2 = v[4*i]*v[4*i] + v[4*i+1]*v[4*i+1] + v[4*i+2]*v[4*i+2];
        kin += 0.5 * pm[4*i+3] * v2;
    }
    double* row = (double*)calloc((size_t)(n > 0 ? n : 1), sizeof(double));
    if (!row) return NAN;
    #pragma omp parallel for schedule(dynamic, 16)
    for (long i = 0; i < n; ++i) {
        double p = 0.0;
        for (long j = i + 1; j < n; ++j) {
            const double dx = pm[4*j] - pm[4*i], dy = pm[4*j+1] - pm[4*i+1], dz = pm[4*j+2] - pm[4*i+2];
            const double dist = sqrt(dx*dx + dy*dy + dz*dz);
            if (dist > 0) p -= NBO_G * pm[4*i+3] * pm[4*j+3] / dist;
        }
        row[i] = p;
    }
    double pot = 0.0;
    for (long i = 0; i < n; ++i) pot += row[i];
    free(row);
    return kin + pot;
}

/* Total linear momentum sum(m v): not a reference function; the drift criterion of BASELINE.json needs it. */
void nbo_momentum(long n, const double* pm, const double* v, double p[3]) {
    double x = 0.0, y = 0.0, z = 0.0;
    for (long i = 0; i < n; ++i) { x += pm[4*i+3]*v[4*i]; y += pm[4*i+3]*v[4*i+1]; z += pm[4*i+3]*v[4*i+2]; }
    p[0] = x; p[1] = y; p[2] = z;
}
