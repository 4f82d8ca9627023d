// nb_noalloc.cu — nb_force_noalloc: one force evaluation that never allocates.
//
// Who this is for. Every other entry point of this library obtains device memory with cudaMalloc, either when a context
// is created or when its partial-sum buffer grows. Some users of an N-body force routine cannot afford that:
//   * a host application that grabbed (almost) all device memory for its own pool at start-up — a later cudaMalloc from a
//     plug-in simply fails with "out of memory";
//   * a per-frame or per-timestep loop with a latency budget (the reference's viewer loop calls its force step once per
//     frame): cudaMalloc/cudaFree synchronise the whole device and stall every stream of the process;
//   * code running where the owner of the GPU forbids allocation altogether after initialisation (fixed memory budgets
//     on shared machines, so that a job's footprint is known up front).
// For them this file offers ONE call that takes caller-owned device arrays in and out and keeps all of its intermediate
// data — the packed FP32 tiles and the per-split partial sums that the force kernel normally gets from the context — in a
// fixed block of device memory that is part of the library image itself (g_noalloc_scratch below). The price is a hard
// size limit (NB_NOALLOC_MAX_N particles) and ~7.6 MB of every GPU's memory that loads the library, held until exit.
//
// What to know before using it:
//   * The first call on each device looks up the device address of the block (cudaGetSymbolAddress) and remembers it;
//     make that first call during start-up, before any stream capture begins. The remembered address belongs to the
//     device's current CUDA context: a process that calls cudaDeviceReset afterwards must not use this entry point again
//     (the address is not looked up a second time; after a reset it may point anywhere).
//   * Every call on a device uses the same block and leaves its data behind. Nothing in the block needs to be zero
//     when a call starts: each call writes every word it later reads (tiles incl. padding, all S partial slots).
//     Two calls that overlap on the GPU (different streams, or different host threads) race on the block and
//     produce garbage — serialise them, e.g. by using one stream.
//   * No cudaMalloc, cudaFree, cudaMallocAsync, CUB or NCCL call happens here; error text goes to a fixed thread-local
//     buffer.
#include "nb_common.cuh"

#include <mutex>

// Capacity: NB_NOALLOC_MAX_N particles (32768) -> 1,310,720 bytes of packed tiles + 8 partial-sum slots of
// 3 x 32768 doubles (6,291,456 bytes) = 7,602,176 bytes per GPU.
constexpr long   kNoallocMaxN   = NB_NOALLOC_MAX_N;                 // multiple of nb::kAlign
constexpr int    kNoallocSplits = 8;                                // upper bound of j-splits (partial-sum slots)
constexpr size_t kTilesBytes    = (size_t)kNoallocMaxN * nb::kNC * sizeof(float);
constexpr size_t kPartBytes     = (size_t)kNoallocSplits * 3 * kNoallocMaxN * sizeof(double);
static_assert(kNoallocMaxN % nb::kAlign == 0, "capacity must be whole alignment units");
static_assert(kTilesBytes % 256 == 0, "partial sums must start 256-byte aligned");

// The scratch block (file scope). In HOST code this name is only a handle for cudaGetSymbolAddress — it is not a device
// pointer and must never be handed to a kernel directly.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kTilesBytes + kPartBytes];

namespace nb {
namespace {

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
unsigned char* g_scratch_addr[kMaxDevices] = {};       // per device: device address of g_noalloc_scratch on that device
int g_sm_count[kMaxDevices] = {};                      // per device: number of SMs (sizes the j-splits)

// caller-order double4 positions -> packed tiles (rows 0..7 and 9; the meta row 8 is not read by the variant-1 kernel)
__global__ void __launch_bounds__(kThreads)
noalloc_pack_kernel(const double4* __restrict__ pos, long n, float* __restrict__ pk, double inv3mbar) {
    const long i = (long)blockIdx.x * kThreads + threadIdx.x;        // grid covers exactly npad particles
    double4 p = pos[i < n ? i : n - 1];
    if (i >= n) p.w = 0.0;                                           // padding: massless copy of the last particle
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

// out[i] = (sum over the S partial slots, 0) in slot order — same layout as the reference's d_accel (cuda/ForceKernels.cu:94)
__global__ void __launch_bounds__(kThreads)
noalloc_reduce_kernel(const double* __restrict__ part, int S, long npad, long n, double4* __restrict__ out) {
    const long i = (long)blockIdx.x * kThreads + threadIdx.x;
    if (i >= n) return;
    double ax = 0.0, ay = 0.0, az = 0.0;
    for (int s = 0; s < S; ++s) {
        const double* q = part + (size_t)s * 3 * npad;
        ax += q[i]; ay += q[npad + i]; az += q[2 * npad + i];
    }
    out[i] = make_double4(ax, ay, az, 0.0);
}

// Device address of the scratch block on the CURRENT device (looked up once per device).
int scratch_for_current_device(unsigned char** out, int* sm_count) {
    int dev = 0;
    NB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(NB_ERR_INVALID, "nb_force_noalloc: device index %d not supported", dev);
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_scratch_addr[dev]) {
        int major = 0, sms = 0;
        NB_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        if (major != 10) return fail(NB_ERR_NO_DEVICE, "nb_force_noalloc: device %d is not sm_100 (this library contains sm_100a code only)", dev);
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_noalloc_scratch)", __FILE__, __LINE__);
        if (!p) return fail(NB_ERR_CUDA, "nb_force_noalloc: cudaGetSymbolAddress returned a null address");
        NB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        g_sm_count[dev] = sms > 0 ? sms : 1;
        g_scratch_addr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_scratch_addr[dev];
    *sm_count = g_sm_count[dev];
    return NB_OK;
}

} // namespace
} // namespace nb

using namespace nb;

extern "C" int nb_force_noalloc(int force_method, const double* d_posMass4, double* d_accel4, long n, double avg_mass, void* stream) {
    if (!d_posMass4 || !d_accel4) return fail(NB_ERR_INVALID, "nb_force_noalloc: NULL device pointer");
    if (n <= 0) return fail(NB_ERR_INVALID, "nb_force_noalloc: n=%ld", n);
    if (n > kNoallocMaxN)
        return fail(NB_ERR_INVALID, "nb_force_noalloc: n=%ld exceeds the built-in scratch capacity of %ld particles "
                                    "(NB_NOALLOC_MAX_N); use a context (nb_create / nb_launch_force) for larger systems", n, kNoallocMaxN);
    if (force_method == NB_FORCE_BARNES_HUT)
        return fail(NB_ERR_UNSUPPORTED, "nb_force_noalloc: forceMethod %d (BARNES_HUT) is not part of this library", force_method);
    if (force_method != NB_FORCE_PAIRWISE && force_method != NB_FORCE_PAIRWISE_AVX2_FP32 && force_method != NB_FORCE_ADAPTIVE_MUTUAL)
        return fail(NB_ERR_INVALID, "nb_force_noalloc: unknown forceMethod %d", force_method);
    const bool adaptive = (force_method == NB_FORCE_ADAPTIVE_MUTUAL);
    if (adaptive && !(avg_mass > 0.0))
        return fail(NB_ERR_INVALID, "nb_force_noalloc: ADAPTIVE_MUTUAL needs the mean mass of the system (avg_mass > 0)");

    unsigned char* base = nullptr;
    int sm_count = 0;
    int rc = scratch_for_current_device(&base, &sm_count); if (rc) return rc;

    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const long npad = (n + kAlign - 1) / kAlign * kAlign;
    const int ntiles = (int)(npad / kTJ);
    const int iblocks = ntiles / 2;                                  // two i-tiles per CTA
    int S = (sm_count * 2 * 4 + iblocks - 1) / iblocks;              // enough CTAs for a few full waves (2 resident per SM)
    if (S > kNoallocSplits) S = kNoallocSplits;
    if (S > ntiles) S = ntiles;

    const double inv3mbar = adaptive ? 1.0 / (3.0 * avg_mass) : 0.0;
    float* pk = reinterpret_cast<float*>(base);
    noalloc_pack_kernel<<<(unsigned)(npad / kThreads), kThreads, 0, st>>>(reinterpret_cast<const double4*>(d_posMass4), n, pk, inv3mbar);
    NB_CUDA(cudaGetLastError());
    double* part = reinterpret_cast<double*>(base + kTilesBytes);

    ForceLaunch L;
    L.it0 = 0; L.nit = ntiles; L.jt0 = 0; L.jcount = ntiles; L.ntiles = ntiles; L.npad = npad; L.s0 = 0; L.S = S; L.kappa = 3.0e38f; L.g0 = 1.f; L.stats = nullptr; L.stats_warps = 0;
    rc = launch_force_fp32(pk, part, L, force_method, /*variant=*/1, st); if (rc) return rc;

    noalloc_reduce_kernel<<<(unsigned)((n + kThreads - 1) / kThreads), kThreads, 0, st>>>(part, S, npad, n, reinterpret_cast<double4*>(d_accel4));
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}
