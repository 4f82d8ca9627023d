// nb_sort.cu — spatial (Morton) ordering of the particles at upload.
//
// No counterpart in the reference (it processes particles in file/generator order). The force kernel decides per
// (i-block, j-tile) whether the cheap hi-only pair arithmetic is accurate enough from the tiles' bounding boxes;
// that only pays off when a tile of 256 consecutive particles is spatially compact, so the context keeps its state
// in Morton order and applies the inverse permutation whenever data leaves through the C ABI. Correctness never
// depends on the order (an unsorted system simply takes the accurate path for every tile).
#include "nb_common.cuh"

#include <cub/device/device_radix_sort.cuh>

namespace nb {
namespace {

__device__ __forceinline__ unsigned long long spread21(unsigned long long v) {
    v &= 0x1fffffULL;
    v = (v | (v << 32)) & 0x1f00000000ffffULL;
    v = (v | (v << 16)) & 0x1f0000ff0000ffULL;
    v = (v | (v << 8))  & 0x100f00f00f00f00fULL;
    v = (v | (v << 4))  & 0x10c30c30c30c30c3ULL;
    v = (v | (v << 2))  & 0x1249249249249249ULL;
    return v;
}

// per-block min/max of x,y,z -> part[block][6]
__global__ void __launch_bounds__(kThreads)
extent_partial_kernel(const double4* __restrict__ pos, long n, double* __restrict__ part) {
    __shared__ double sh[6][kThreads / 32];
    const long i = (long)blockIdx.x * kThreads + threadIdx.x;
    double v[6] = {1e300, 1e300, 1e300, -1e300, -1e300, -1e300};
    if (i < n) { const double4 p = pos[i]; v[0] = v[3] = p.x; v[1] = v[4] = p.y; v[2] = v[5] = p.z; }
#pragma unroll
    for (int c = 0; c < 6; ++c) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double w = __shfl_xor_sync(0xffffffffu, v[c], o);
            v[c] = c < 3 ? fmin(v[c], w) : fmax(v[c], w);
        }
        if ((threadIdx.x & 31) == 0) sh[c][threadIdx.x >> 5] = v[c];
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        const int c = threadIdx.x;
        double r = sh[c][0];
        for (int k = 1; k < kThreads / 32; ++k) r = c < 3 ? fmin(r, sh[c][k]) : fmax(r, sh[c][k]);
        part[(size_t)blockIdx.x * 6 + c] = r;
    }
}
__global__ void __launch_bounds__(1024)
extent_final_kernel(const double* __restrict__ part, long nblocks, double* __restrict__ box /* [8] */) {
    __shared__ double sh[6][32];
    double v[6] = {1e300, 1e300, 1e300, -1e300, -1e300, -1e300};
    for (long b = threadIdx.x; b < nblocks; b += 1024)
#pragma unroll
        for (int c = 0; c < 6; ++c) v[c] = c < 3 ? fmin(v[c], part[b * 6 + c]) : fmax(v[c], part[b * 6 + c]);
#pragma unroll
    for (int c = 0; c < 6; ++c) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double w = __shfl_xor_sync(0xffffffffu, v[c], o);
            v[c] = c < 3 ? fmin(v[c], w) : fmax(v[c], w);
        }
        if ((threadIdx.x & 31) == 0) sh[c][threadIdx.x >> 5] = v[c];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double r[6];
        for (int c = 0; c < 6; ++c) { r[c] = sh[c][0]; for (int k = 1; k < 32; ++k) r[c] = c < 3 ? fmin(r[c], sh[c][k]) : fmax(r[c], sh[c][k]); }
        double ext = fmax(r[3] - r[0], fmax(r[4] - r[1], r[5] - r[2]));
        if (!(ext > 0.0)) ext = 1.0;
        box[0] = r[0]; box[1] = r[1]; box[2] = r[2];
        box[3] = 2097151.0 / ext;                  // cells of equal size in all three dimensions (2^21 - 1 per axis)
    }
}
// 3-D Hilbert index from three 21-bit coordinates (Skilling's transpose algorithm, "Programming the Hilbert curve",
// AIP Conf. Proc. 707, 2004), bits interleaved into one 63-bit key. Unlike the Morton (Z) curve, consecutive keys are
// always adjacent cells, so a run of 64 / 128 / 256 / 1024 consecutive particles is a connected, compact blob and the
// bounding boxes the force kernel classifies tiles with overlap far less.
__device__ __forceinline__ unsigned long long hilbert63(unsigned int x, unsigned int y, unsigned int z) {
    unsigned int X[3] = {x, y, z};
    const unsigned int M = 1u << 20;
    for (unsigned int Q = M; Q > 1; Q >>= 1) {          // inverse undo
        const unsigned int P = Q - 1;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            if (X[i] & Q) X[0] ^= P;
            else { const unsigned int t = (X[0] ^ X[i]) & P; X[0] ^= t; X[i] ^= t; }
        }
    }
#pragma unroll
    for (int i = 1; i < 3; ++i) X[i] ^= X[i - 1];       // Gray encode
    unsigned int t = 0;
    for (unsigned int Q = M; Q > 1; Q >>= 1) if (X[2] & Q) t ^= Q - 1;
#pragma unroll
    for (int i = 0; i < 3; ++i) X[i] ^= t;
    // transpose -> key: bit b of X[0] is the most significant of each triple
    return (spread21(X[0]) << 2) | (spread21(X[1]) << 1) | spread21(X[2]);
}

__global__ void __launch_bounds__(kThreads)
morton_key_kernel(const double4* __restrict__ pos, long n, const double* __restrict__ box,
                  unsigned long long* __restrict__ keys, unsigned int* __restrict__ idx, int curve) {
    const long i = (long)blockIdx.x * kThreads + threadIdx.x;
    if (i >= n) return;
    const double4 p = pos[i];
    const double s = box[3];
    const unsigned long long kx = (unsigned long long)fmin(fmax((p.x - box[0]) * s, 0.0), 2097151.0);
    const unsigned long long ky = (unsigned long long)fmin(fmax((p.y - box[1]) * s, 0.0), 2097151.0);
    const unsigned long long kz = (unsigned long long)fmin(fmax((p.z - box[2]) * s, 0.0), 2097151.0);
    keys[i] = curve == 1 ? hilbert63((unsigned int)kx, (unsigned int)ky, (unsigned int)kz)
                         : (spread21(kx) | (spread21(ky) << 1) | (spread21(kz) << 2));
    idx[i] = (unsigned int)i;
}
__global__ void __launch_bounds__(kThreads)
iota_kernel(unsigned int* __restrict__ idx, long n) {
    const long i = (long)blockIdx.x * kThreads + threadIdx.x;
    if (i < n) idx[i] = (unsigned int)i;
}

// dst[s] = src[perm[s]] for s < n; pads (n <= s < npad) copy particle perm[n-1]'s position with zero mass / velocity
__global__ void __launch_bounds__(kThreads)
gather_state_kernel(const double4* __restrict__ src_pos, const double4* __restrict__ src_vel,
                    const unsigned int* __restrict__ perm, long n, long npad,
                    double4* __restrict__ pos, double4* __restrict__ vel) {
    const long s = (long)blockIdx.x * kThreads + threadIdx.x;
    if (s >= npad) return;
    if (s < n) {
        const unsigned int o = perm[s];
        pos[s] = src_pos[o];
        vel[s] = src_vel ? src_vel[o] : make_double4(0, 0, 0, 0);
    } else {
        double4 p = src_pos[perm[n - 1]];
        p.w = 0.0;
        pos[s] = p; vel[s] = make_double4(0, 0, 0, 0);
    }
}
// dst[perm[s]] = src[s]
__global__ void __launch_bounds__(kThreads)
scatter4_kernel(const double4* __restrict__ src, const unsigned int* __restrict__ perm, long first, long count,
                double4* __restrict__ dst) {
    const long k = (long)blockIdx.x * kThreads + threadIdx.x;
    if (k >= count) return;
    dst[perm[first + k]] = src[first + k];
}
__global__ void __launch_bounds__(kThreads)
scatter_acc_kernel(const double* __restrict__ acc, long npad, const unsigned int* __restrict__ perm, long first, long count,
                   double4* __restrict__ dst) {
    const long k = (long)blockIdx.x * kThreads + threadIdx.x;
    if (k >= count) return;
    const long s = first + k;
    dst[perm[s]] = make_double4(acc[s], acc[npad + s], acc[2 * npad + s], 0.0);   // cuda/ForceKernels.cu:94 layout
}
// cuda/IntegrationKernels.cu:163-197 in the caller's particle order: [i][6] = (x,y,z,vx,vy,vz); [i][2] = (id, mass)
__global__ void __launch_bounds__(kThreads)
snapshot_scatter_kernel(const double4* __restrict__ pos, const double4* __restrict__ vel, const unsigned int* __restrict__ perm,
                        long n, double* __restrict__ particleData6, double* __restrict__ staticData2) {
    const long s = (long)blockIdx.x * kThreads + threadIdx.x;
    if (s >= n) return;
    const size_t o = perm[s];
    const double4 p = pos[s], v = vel[s];
    double* q = particleData6 + o * 6;
    q[0] = p.x; q[1] = p.y; q[2] = p.z; q[3] = v.x; q[4] = v.y; q[5] = v.z;
    if (staticData2) { staticData2[o * 2] = (double)o; staticData2[o * 2 + 1] = p.w; }
}

// re-sort helpers: dst[s] = src[p2[s]] for the first n slots (pads keep their slots)
__global__ void __launch_bounds__(kThreads)
permute4_kernel(const double4* __restrict__ src, const unsigned int* __restrict__ p2, long n, double4* __restrict__ dst) {
    const long s = (long)blockIdx.x * kThreads + threadIdx.x;
    if (s < n) dst[s] = src[p2[s]];
}
__global__ void __launch_bounds__(kThreads)
permute_acc_kernel(const double* __restrict__ acc, long npad, const unsigned int* __restrict__ p2, long n, double* __restrict__ dst) {
    const long s = (long)blockIdx.x * kThreads + threadIdx.x;
    if (s >= n) return;
    const unsigned int o = p2[s];
    dst[s] = acc[o]; dst[npad + s] = acc[npad + o]; dst[2 * npad + s] = acc[2 * npad + o];
}
__global__ void __launch_bounds__(kThreads)
compose_perm_kernel(const unsigned int* __restrict__ perm, const unsigned int* __restrict__ p2, long n, unsigned int* __restrict__ dst) {
    const long s = (long)blockIdx.x * kThreads + threadIdx.x;
    if (s < n) dst[s] = perm[p2[s]];
}

inline unsigned blocks_for(long count) { return (unsigned)((count + kThreads - 1) / kThreads); }

} // namespace

int launch_permute4(const double4* src, const unsigned int* p2, long n, double4* dst, cudaStream_t st) {
    permute4_kernel<<<blocks_for(n), kThreads, 0, st>>>(src, p2, n, dst);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}
int launch_permute_acc(const double* acc, long npad, const unsigned int* p2, long n, double* dst, cudaStream_t st) {
    permute_acc_kernel<<<blocks_for(n), kThreads, 0, st>>>(acc, npad, p2, n, dst);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}
int launch_compose_perm(const unsigned int* perm, const unsigned int* p2, long n, unsigned int* dst, cudaStream_t st) {
    compose_perm_kernel<<<blocks_for(n), kThreads, 0, st>>>(perm, p2, n, dst);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

size_t sort_temp_bytes(long n) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                    (const unsigned int*)nullptr, (unsigned int*)nullptr, (int)n);
    return bytes;
}

// perm[s] = index (caller order) of the particle stored at sorted slot s. scratch: keys (2n u64), idx (n u32),
// extent partials (6 * blocks doubles + 8), cub temp.
int build_permutation(const double4* src_pos, long n, int spatial, SortScratch& S, unsigned int* perm, cudaStream_t st) {
    if (!spatial) {
        iota_kernel<<<blocks_for(n), kThreads, 0, st>>>(perm, n);
        NB_CUDA(cudaGetLastError());
        return NB_OK;
    }
    extent_partial_kernel<<<blocks_for(n), kThreads, 0, st>>>(src_pos, n, S.extent_part);
    NB_CUDA(cudaGetLastError());
    extent_final_kernel<<<1, 1024, 0, st>>>(S.extent_part, (long)blocks_for(n), S.box);
    NB_CUDA(cudaGetLastError());
    morton_key_kernel<<<blocks_for(n), kThreads, 0, st>>>(src_pos, n, S.box, S.keys_in, S.idx_in, spatial == 2 ? 0 : 1);
    NB_CUDA(cudaGetLastError());
    size_t bytes = S.cub_bytes;
    NB_CUDA(cub::DeviceRadixSort::SortPairs(S.cub_temp, bytes, S.keys_in, S.keys_out, S.idx_in, perm, (int)n, 0, 63, st));
    return NB_OK;
}

int launch_gather_state(const double4* src_pos, const double4* src_vel, const unsigned int* perm, long n, long npad,
                        double4* pos, double4* vel, cudaStream_t st) {
    gather_state_kernel<<<blocks_for(npad), kThreads, 0, st>>>(src_pos, src_vel, perm, n, npad, pos, vel);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}
int launch_scatter4(const double4* src, const unsigned int* perm, long first, long count, double4* dst, cudaStream_t st) {
    if (count <= 0) return NB_OK;
    scatter4_kernel<<<blocks_for(count), kThreads, 0, st>>>(src, perm, first, count, dst);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}
int launch_scatter_acc(const double* acc, long npad, const unsigned int* perm, long first, long count, double4* dst, cudaStream_t st) {
    if (count <= 0) return NB_OK;
    scatter_acc_kernel<<<blocks_for(count), kThreads, 0, st>>>(acc, npad, perm, first, count, dst);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}
int launch_snapshot_scatter(const double4* pos, const double4* vel, const unsigned int* perm, long n,
                            double* particleData6, double* staticData2, cudaStream_t st) {
    if (n <= 0) return NB_OK;
    snapshot_scatter_kernel<<<blocks_for(n), kThreads, 0, st>>>(pos, vel, perm, n, particleData6, staticData2);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

} // namespace nb
