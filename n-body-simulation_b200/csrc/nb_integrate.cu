// nb_integrate.cu — O(N) kernels around the force pass: packing, fused drift / kick, reductions, snapshots.
//
// Replaces velocityVerletPositionKernel / velocityVerletVelocityKernel / eulerIntegrationKernel and the
// accel -> accelOld device copy of performGPUIntegrationStep (reference cuda/IntegrationKernels.cu:26-86,200-266),
// computeTotalEnergyKernel's reduction (:89-160) and the snapshot kernels (:163-197). Unlike the reference GPU path
// the centre-of-mass velocity is removed at the start of every step, as the CPU integrators do
// (core/System.cpp:45-57,392-400,438-446) — the CPU path is the parity target.
//
// Per step (Velocity Verlet):   drift  : v -= v_cm ; x += v dt + a dt^2/2 ; re-pack x into FP32 hi/lo tiles
//                               force  : part[s] = F(x) partial sums            (nb_force.cu)
//                               finish : a' = sum_s part[s] ; v += (a + a') dt/2 ; a = a' ; block sums of m v
//                               com    : v_cm numerator/denominator in fixed order (deterministic, no atomics)
// All reductions are fixed-order so a run is bit-reproducible.
#include "nb_common.cuh"

namespace nb {
namespace {

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block-wide sum of four values with warp shuffles; result valid in thread 0.
__device__ __forceinline__ void block_sum4(double v[4], double* sh /* [4][8] */) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int c = 0; c < 4; ++c) { v[c] = warp_sum(v[c]); if (lane == 0) sh[c * 8 + w] = v[c]; }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int c = 0; c < 4; ++c) { double s = 0.0; for (int k = 0; k < kThreads / 32; ++k) s += sh[c * 8 + k]; v[c] = s; }
    }
}

__device__ __forceinline__ void pack_one(float* __restrict__ pk, long i, const double4 p, double inv3mbar) {
    float* o = pk + (size_t)(i / kTJ) * kTileFloats + (i % kTJ);
    const float xh = (float)p.x, yh = (float)p.y, zh = (float)p.z;
    o[0 * kTJ] = xh; o[1 * kTJ] = yh; o[2 * kTJ] = zh;
    const double gmd = kG * p.w;
    const float gmh = (float)gmd;
    o[3 * kTJ] = gmh;
    o[kGmLoRow * kTJ] = (float)(gmd - (double)gmh);
    o[4 * kTJ] = (float)(p.x - (double)xh);
    o[5 * kTJ] = (float)(p.y - (double)yh);
    o[6 * kTJ] = (float)(p.z - (double)zh);
    o[7 * kTJ] = (float)(p.w * inv3mbar);
}

// Bounding boxes of the tile this CTA just packed (one CTA == one tile of kTJ particles) -> meta row of the tile:
//   [0..2] min, [3..5] max of the whole tile, [6] max |coordinate|,
//   [8 + 6 sb .. 8 + 6 sb + 5] min / max of sub-tile sb (64 consecutive particles), sb = 0..3.
__device__ __forceinline__ void tile_meta(float* __restrict__ pk, long tile, float x, float y, float z) {
    __shared__ float shm[7][kThreads / 32];
    float v[7] = {x, y, z, x, y, z, fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z)))};
#pragma unroll
    for (int c = 0; c < 7; ++c) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float w = __shfl_xor_sync(0xffffffffu, v[c], o);
            v[c] = c < 3 ? fminf(v[c], w) : fmaxf(v[c], w);
        }
        if ((threadIdx.x & 31) == 0) shm[c][threadIdx.x >> 5] = v[c];
    }
    __syncthreads();
    float* meta = pk + (size_t)tile * kTileFloats + kMetaRow * kTJ;
    if (threadIdx.x < 7) {
        const int c = threadIdx.x;
        float r = shm[c][0];
        for (int k = 1; k < kThreads / 32; ++k) r = c < 3 ? fminf(r, shm[c][k]) : fmaxf(r, shm[c][k]);
        meta[c] = r;
    } else if (threadIdx.x >= 32 && threadIdx.x < 32 + 24) {      // sub-tile sb = warps 2 sb and 2 sb + 1
        const int k = threadIdx.x - 32, sb = k / 6, c = k % 6;
        const float a = shm[c][2 * sb], b2 = shm[c][2 * sb + 1];
        meta[kSubBoxOffset + k] = c < 3 ? fminf(a, b2) : fmaxf(a, b2);
        (void)sb;
    }
}

// `first` and `count` are multiples of kTJ: CTA b packs tile first/kTJ + b.
__global__ void __launch_bounds__(kThreads)
pack_kernel(const double4* __restrict__ pos, float* __restrict__ pk, long first, long count, double inv3mbar) {
    const long i = first + (long)blockIdx.x * kThreads + threadIdx.x;
    const double4 p = pos[i];
    pack_one(pk, i, p, inv3mbar);
    tile_meta(pk, i / kTJ, (float)p.x, (float)p.y, (float)p.z);
}

// drift: one thread per owned particle; one CTA == one tile (A.i0 and A.ni are multiples of kTJ).
template <int INTEG>
__global__ void __launch_bounds__(kThreads)
drift_kernel(StepArrays A, double dt, int subtract_com, double inv3mbar) {
    const long i = A.i0 + (long)blockIdx.x * kThreads + threadIdx.x;
    double4 p = A.pos[i];
    const bool real = i < A.n;       // zero-mass pads stay at rest where they were put at upload
    if (real) {
        double4 v = A.vel[i];
        if (subtract_com) {
            const double M = A.red[3];
            const double inv = M != 0.0 ? 1.0 / M : 0.0;
            v.x -= A.red[0] * inv; v.y -= A.red[1] * inv; v.z -= A.red[2] * inv;
        }
        const double ax = A.acc[i], ay = A.acc[A.npad + i], az = A.acc[2 * A.npad + i];
        if (INTEG == NB_INTEG_VELOCITY_VERLET) {
            // cuda/IntegrationKernels.cu:36-38 ; core/System.cpp:463-465
            p.x += v.x * dt + 0.5 * ax * (dt * dt);
            p.y += v.y * dt + 0.5 * ay * (dt * dt);
            p.z += v.z * dt + 0.5 * az * (dt * dt);
            if (subtract_com) A.vel[i] = v;
        } else {
            // cuda/IntegrationKernels.cu:74-81 ; core/System.cpp:417-424: x uses the old v, then v += a dt
            p.x += v.x * dt; p.y += v.y * dt; p.z += v.z * dt;
            v.x += ax * dt; v.y += ay * dt; v.z += az * dt;
            A.vel[i] = v;
        }
        A.pos[i] = p;
        if (A.pk) pack_one(A.pk, i, p, inv3mbar);
    }
    if (A.pk) tile_meta(A.pk, i / kTJ, (float)p.x, (float)p.y, (float)p.z);
}

// finish: reduce the force partials in slot order; VV kick; per-block sums of (m vx, m vy, m vz, m).
template <bool KICK>
__global__ void __launch_bounds__(kThreads)
finish_kernel(StepArrays A, double dt, int S) {
    __shared__ double sh[32];
    const long k = (long)blockIdx.x * kThreads + threadIdx.x;
    double r[4] = {0.0, 0.0, 0.0, 0.0};
    if (k < A.ni && A.i0 + k < A.n) {
        const long i = A.i0 + k;
        double ax = 0.0, ay = 0.0, az = 0.0;
        for (int s = 0; s < S; ++s) {
            const double* q = A.part + (size_t)s * 3 * A.npad;
            ax += q[i]; ay += q[A.npad + i]; az += q[2 * A.npad + i];
        }
        double4 v = A.vel[i];
        if (KICK) {
            // cuda/IntegrationKernels.cu:55-57 ; core/System.cpp:484-486
            v.x += 0.5 * (A.acc[i] + ax) * dt;
            v.y += 0.5 * (A.acc[A.npad + i] + ay) * dt;
            v.z += 0.5 * (A.acc[2 * A.npad + i] + az) * dt;
            A.vel[i] = v;
        }
        A.acc[i] = ax; A.acc[A.npad + i] = ay; A.acc[2 * A.npad + i] = az;
        const double m = A.pos[i].w;
        r[0] = m * v.x; r[1] = m * v.y; r[2] = m * v.z; r[3] = m;
    }
    block_sum4(r, sh);
    if (threadIdx.x == 0) {
        double* o = A.red_part + (size_t)blockIdx.x * 4;
        o[0] = r[0]; o[1] = r[1]; o[2] = r[2]; o[3] = r[3];
    }
}

// Single-CTA fixed-order sum of the per-block partials -> red[0..3].
__global__ void __launch_bounds__(1024)
reduce4_kernel(const double* __restrict__ red_part, long nblocks, double* __restrict__ red) {
    __shared__ double sh[4][32];
    double v[4] = {0.0, 0.0, 0.0, 0.0};
    for (long b = threadIdx.x; b < nblocks; b += 1024) {
#pragma unroll
        for (int c = 0; c < 4; ++c) v[c] += red_part[b * 4 + c];
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int c = 0; c < 4; ++c) { v[c] = warp_sum(v[c]); if (lane == 0) sh[c][w] = v[c]; }
    __syncthreads();
    if (threadIdx.x < 4) { double s = 0.0; for (int k = 0; k < 32; ++k) s += sh[threadIdx.x][k]; red[threadIdx.x] = s; }
}

// Energy: block sums of (kinetic, m_i * pot_i, -, -) over all particles.
__global__ void __launch_bounds__(kThreads)
energy_partial_kernel(const double4* __restrict__ pos, const double4* __restrict__ vel, const double* __restrict__ pot,
                      int pot_splits, long npad, long first, long count, double* __restrict__ red_part) {
    __shared__ double sh[32];
    const long k = (long)blockIdx.x * kThreads + threadIdx.x;
    double r[4] = {0.0, 0.0, 0.0, 0.0};
    if (k < count) {
        const long i = first + k;
        const double4 p = pos[i], v = vel[i];
        r[0] = 0.5 * p.w * (v.x * v.x + v.y * v.y + v.z * v.z);   // core/System.cpp:66-70
        double phi = 0.0;
        for (int s = 0; s < pot_splits; ++s) phi += pot[(size_t)s * npad + i];      // partial potentials in split order
        r[1] = p.w * phi;
    }
    block_sum4(r, sh);
    if (threadIdx.x == 0) { double* o = red_part + (size_t)blockIdx.x * 4; o[0] = r[0]; o[1] = r[1]; o[2] = 0.0; o[3] = 0.0; }
}
__global__ void energy_final_kernel(const double* __restrict__ red, double* __restrict__ out3) {
    const double kin = red[0], potl = -kG * red[1];   // potential_fp64_kernel gives every unordered pair a total weight of 1
    out3[0] = kin + potl; out3[1] = kin; out3[2] = potl;
}

// Mass / momentum block sums straight from pos/vel (used after upload and by nb_momentum).
__global__ void __launch_bounds__(kThreads)
momentum_partial_kernel(const double4* __restrict__ pos, const double4* __restrict__ vel, long first, long count,
                        double* __restrict__ red_part) {
    __shared__ double sh[32];
    const long k = (long)blockIdx.x * kThreads + threadIdx.x;
    double r[4] = {0.0, 0.0, 0.0, 0.0};
    if (k < count) {
        const double4 p = pos[first + k];
        double4 v = make_double4(0, 0, 0, 0);
        if (vel) v = vel[first + k];
        r[0] = p.w * v.x; r[1] = p.w * v.y; r[2] = p.w * v.z; r[3] = p.w;
    }
    block_sum4(r, sh);
    if (threadIdx.x == 0) { double* o = red_part + (size_t)blockIdx.x * 4; o[0] = r[0]; o[1] = r[1]; o[2] = r[2]; o[3] = r[3]; }
}

// cuda/IntegrationKernels.cu:163-197: [i][6] = (x,y,z,vx,vy,vz); [i][2] = (id, mass)
__global__ void __launch_bounds__(kThreads)
snapshot_kernel(const double4* __restrict__ pos, const double4* __restrict__ vel, long n,
                double* __restrict__ particleData6, double* __restrict__ staticData2) {
    const long i = (long)blockIdx.x * kThreads + threadIdx.x;
    if (i >= n) return;
    const double4 p = pos[i], v = vel[i];
    double* o = particleData6 + (size_t)i * 6;
    o[0] = p.x; o[1] = p.y; o[2] = p.z; o[3] = v.x; o[4] = v.y; o[5] = v.z;
    if (staticData2) { staticData2[(size_t)i * 2] = (double)i; staticData2[(size_t)i * 2 + 1] = p.w; }
}

__global__ void __launch_bounds__(kThreads)
accel_export_kernel(const double* __restrict__ acc, long npad, long first, long count, double4* __restrict__ out) {
    const long k = (long)blockIdx.x * kThreads + threadIdx.x;
    if (k >= count) return;
    const long i = first + k;
    out[i] = make_double4(acc[i], acc[npad + i], acc[2 * npad + i], 0.0);   // cuda/ForceKernels.cu:94
}
__global__ void __launch_bounds__(kThreads)
accel_sum_export_kernel(const double* __restrict__ part, int S, long npad, long n, double4* __restrict__ out) {
    const long i = (long)blockIdx.x * kThreads + threadIdx.x;
    if (i >= n) return;
    double ax = 0.0, ay = 0.0, az = 0.0;
    for (int s = 0; s < S; ++s) {
        const double* q = part + (size_t)s * 3 * npad;
        ax += q[i]; ay += q[npad + i]; az += q[2 * npad + i];
    }
    out[i] = make_double4(ax, ay, az, 0.0);
}

inline unsigned blocks_for(long count) { return (unsigned)((count + kThreads - 1) / kThreads); }

} // namespace

int launch_pack(const StepArrays& A, long first, long count, double avg_mass, cudaStream_t st) {
    if (count <= 0 || !A.pk) return NB_OK;
    if (first % kTJ || count % kTJ) return fail(NB_ERR_INVALID, "pack: range [%ld,+%ld) is not tile aligned", first, count);
    const double inv3mbar = avg_mass > 0.0 ? 1.0 / (3.0 * avg_mass) : 0.0;
    pack_kernel<<<blocks_for(count), kThreads, 0, st>>>(A.pos, A.pk, first, count, inv3mbar);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

int launch_drift(const StepArrays& A, int integ, double dt, bool subtract_com, double avg_mass, cudaStream_t st) {
    if (A.ni <= 0) return NB_OK;
    const double inv3mbar = avg_mass > 0.0 ? 1.0 / (3.0 * avg_mass) : 0.0;
    if (integ == NB_INTEG_VELOCITY_VERLET)
        drift_kernel<NB_INTEG_VELOCITY_VERLET><<<blocks_for(A.ni), kThreads, 0, st>>>(A, dt, subtract_com ? 1 : 0, inv3mbar);
    else
        drift_kernel<NB_INTEG_EULER><<<blocks_for(A.ni), kThreads, 0, st>>>(A, dt, subtract_com ? 1 : 0, inv3mbar);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

int launch_finish(const StepArrays& A, int integ, bool kick, double dt, int S, cudaStream_t st) {
    if (A.ni <= 0) return NB_OK;
    (void)integ;
    if (kick) finish_kernel<true><<<blocks_for(A.ni), kThreads, 0, st>>>(A, dt, S);
    else      finish_kernel<false><<<blocks_for(A.ni), kThreads, 0, st>>>(A, dt, S);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

int launch_com_final(const StepArrays& A, cudaStream_t st) {
    reduce4_kernel<<<1, 1024, 0, st>>>(A.red_part, (long)blocks_for(A.ni), A.red);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

// Kinetic and potential sums over the particles [first, first+count) -> out4 = (sum 1/2 m v^2, sum m_i pot_i, 0, 0).
// A rank of a multi-GPU run sums its own slice; the four doubles are then all-reduced before launch_energy_final.
int launch_energy_partial(const StepArrays& A, const double* pot, int pot_splits, long first, long count, double* out4, cudaStream_t st) {
    const long nb = count > 0 ? (long)blocks_for(count) : 0;
    if (nb > 0) {
        energy_partial_kernel<<<(unsigned)nb, kThreads, 0, st>>>(A.pos, A.vel, pot, pot_splits, A.npad, first, count, A.red_part);
        NB_CUDA(cudaGetLastError());
    }
    reduce4_kernel<<<1, 1024, 0, st>>>(A.red_part, nb, out4);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}
int launch_energy_final(const double* in4, double* out3, cudaStream_t st) {
    energy_final_kernel<<<1, 1, 0, st>>>(in4, out3);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

int launch_snapshot(const double4* pos, const double4* vel, long n, double* particleData6, double* staticData2, cudaStream_t st) {
    if (n <= 0) return NB_OK;
    snapshot_kernel<<<blocks_for(n), kThreads, 0, st>>>(pos, vel, n, particleData6, staticData2);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

int launch_accel_export(const double* acc, long npad, long first, long count, double4* out, cudaStream_t st) {
    if (count <= 0) return NB_OK;
    accel_export_kernel<<<blocks_for(count), kThreads, 0, st>>>(acc, npad, first, count, out);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

int launch_accel_import_sum(const double* part, int S, long npad, long n, double4* out, cudaStream_t st) {
    if (n <= 0) return NB_OK;
    accel_sum_export_kernel<<<blocks_for(n), kThreads, 0, st>>>(part, S, npad, n, out);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

// red[0..3] = (sum m vx, sum m vy, sum m vz, sum m) over [first, first+count)
int launch_mass_sum(const double4* pos, long n, double* red_part, double* red, cudaStream_t st) {
    momentum_partial_kernel<<<blocks_for(n), kThreads, 0, st>>>(pos, nullptr, 0, n, red_part);
    NB_CUDA(cudaGetLastError());
    reduce4_kernel<<<1, 1024, 0, st>>>(red_part, (long)blocks_for(n), red);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

int launch_momentum(const double4* pos, const double4* vel, long first, long count, double* red_part, double* red, cudaStream_t st) {
    momentum_partial_kernel<<<blocks_for(count), kThreads, 0, st>>>(pos, vel, first, count, red_part);
    NB_CUDA(cudaGetLastError());
    reduce4_kernel<<<1, 1024, 0, st>>>(red_part, (long)blocks_for(count), red);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

// ---- device-side SPIRAL_GALAXY generator (core/Initializer.cpp:360-469) -------------------------------------------------
// Same distribution and the same formulas as the reference generator (equal-mass stars, uniform-area disc r = R sqrt(u), arm
// by index, random in-arm angle, uniform thickness, rotation curve v = v0 sqrt(r/R) plus a uniform velocity perturbation, random
// disc orientation), but every draw is a pure function of (seed, particle index, draw number) — a counter-based generator
// (splitmix64 finaliser) instead of one sequential std::mt19937 stream — so that the particles can be made where they are
// needed: on the device, by any number of threads or GPUs, identically. 16M particles take a fraction of a millisecond
// instead of seconds on the host plus 1 GB over PCIe.
__host__ __device__ inline unsigned long long nb_mix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
// uniform in [0, 1) with 53 random bits: draw k of particle i (i = 2^63 - 1: the per-system draws)
__host__ __device__ inline double nb_u01(unsigned long long seed, unsigned long long i, unsigned k) {
    return (double)(nb_mix64(nb_mix64(seed) ^ nb_mix64(i * 8ULL + k)) >> 11) * (1.0 / 9007199254740992.0);
}

struct SpiralParams { double mass, radius, thickness, perturbation, v0; int arms; double e1[3], e2[3], nrm[3]; };

__global__ void __launch_bounds__(kThreads)
spiral_init_kernel(unsigned long long seed, long n, SpiralParams P, double4* __restrict__ pos, double4* __restrict__ vel) {
    const long i = (long)blockIdx.x * kThreads + threadIdx.x;
    if (i >= n) return;
    const double kPi = 3.14159265358979323846;
    const double u = nb_u01(seed, i, 0);
    const double r = P.radius * sqrt(u);
    const double phi = (2.0 * kPi * (double)(i % P.arms)) / (double)P.arms + (r / P.radius) * 2.0 * kPi +
                       (nb_u01(seed, i, 1) * 2.0 - 1.0) * (kPi / 8.0);
    double sphi, cphi; sincos(phi, &sphi, &cphi);
    const double xd = r * cphi, yd = r * sphi, zd = (nb_u01(seed, i, 2) - 0.5) * P.thickness;
    const double sp = P.v0 * sqrt(r / P.radius);
    double x[3], v[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        x[c] = xd * P.e1[c] + yd * P.e2[c] + zd * P.nrm[c];
        v[c] = sp * (-sphi * P.e1[c] + cphi * P.e2[c]) + (nb_u01(seed, i, 3 + c) * 2.0 - 1.0) * P.perturbation;
    }
    pos[i] = make_double4(x[0], x[1], x[2], P.mass);
    vel[i] = make_double4(v[0], v[1], v[2], 0.0);
}

int launch_spiral_init(unsigned long long seed, long n, double total_mass, int arms, double radius, double thickness,
                       double perturbation, double4* pos, double4* vel, double* normal3, cudaStream_t st) {
    if (n <= 0 || arms <= 0 || !(radius > 0.0) || !(total_mass > 0.0))
        return fail(NB_ERR_INVALID, "spiral generator: n=%ld arms=%d radius=%g total mass=%g", n, arms, radius, total_mass);
    SpiralParams P;
    P.mass = total_mass / (double)n; P.radius = radius; P.thickness = thickness; P.perturbation = perturbation; P.arms = arms;
    P.v0 = sqrt(kG * total_mass / radius);
    const double kPi = 3.14159265358979323846;
    const unsigned long long sys = 0x7FFFFFFFFFFFFFFFULL;                  // the two per-system draws: disc orientation
    const double th = nb_u01(seed, sys, 0) * kPi, ph = nb_u01(seed, sys, 1) * 2.0 * kPi;
    P.nrm[0] = sin(th) * cos(ph); P.nrm[1] = sin(th) * sin(ph); P.nrm[2] = cos(th);
    if (fabs(P.nrm[0]) < 1e-6 && fabs(P.nrm[1]) < 1e-6) { P.e1[0] = 1; P.e1[1] = 0; P.e1[2] = 0; }
    else {
        const double len = sqrt(P.nrm[0] * P.nrm[0] + P.nrm[1] * P.nrm[1]);
        P.e1[0] = -P.nrm[1] / len; P.e1[1] = P.nrm[0] / len; P.e1[2] = 0.0;
    }
    P.e2[0] = P.nrm[1] * P.e1[2] - P.nrm[2] * P.e1[1];
    P.e2[1] = P.nrm[2] * P.e1[0] - P.nrm[0] * P.e1[2];
    P.e2[2] = P.nrm[0] * P.e1[1] - P.nrm[1] * P.e1[0];
    if (normal3) { normal3[0] = P.nrm[0]; normal3[1] = P.nrm[1]; normal3[2] = P.nrm[2]; }
    spiral_init_kernel<<<blocks_for(n), kThreads, 0, st>>>(seed, n, P, pos, vel);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

} // namespace nb
