// nb_context.cu — the C ABI of include/nbody_b200.h: context, host<->device staging, step driver, multi-GPU exchange.
//
// Host-side counterpart of the reference GPU drivers runBenchmarkOnGPU (cuda/BenchmarkKernels.cu:15-126),
// runSimulationOnGPU (cuda/SimulationKernels.cu:22-286) and performGPUIntegrationStep
// (cuda/IntegrationKernels.cu:200-266), re-designed around a persistent context (no per-call cudaMalloc), error codes
// instead of exit(), and an i-particle partition with an NCCL all-gather of the packed positions per step.
#include "nb_common.cuh"

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <type_traits>
#include <dlfcn.h>
#include <vector>

namespace nb {

static thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
    return code;
}
int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
    const char* base = strrchr(file, '/'); base = base ? base + 1 : file;
    snprintf(g_err, sizeof(g_err), "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), base, line, what);
    cudaGetLastError();
    return e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? NB_ERR_NO_DEVICE : NB_ERR_CUDA;
}

// ---- NCCL, bound at run time (dlopen) so that the library loads on machines without it ------------------------------
namespace nccl {
typedef struct ncclComm* comm_t;
struct unique_id { char internal[128]; };
enum { kSuccess = 0 };
enum { kFloat32 = 7, kFloat64 = 8, kSum = 0 };      // ncclDataType_t / ncclRedOp_t values (nccl.h)
static void* lib = nullptr;
static int (*GetUniqueId)(unique_id*) = nullptr;
static int (*CommInitRank)(comm_t*, int, unique_id, int) = nullptr;
static int (*CommDestroy)(comm_t) = nullptr;
static int (*AllGather)(const void*, void*, size_t, int, comm_t, cudaStream_t) = nullptr;
static int (*AllReduce)(const void*, void*, size_t, int, int, comm_t, cudaStream_t) = nullptr;
static const char* (*GetErrorString)(int) = nullptr;
static int load() {
    if (lib) return NB_OK;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) { lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL); if (lib) break; }
    if (!lib) return fail(NB_ERR_NCCL, "cannot dlopen libnccl.so.2: %s", dlerror());
#define NB_SYM(var, name) *(void**)(&var) = dlsym(lib, name); if (!var) return fail(NB_ERR_NCCL, "NCCL symbol %s missing", name)
    NB_SYM(GetUniqueId, "ncclGetUniqueId"); NB_SYM(CommInitRank, "ncclCommInitRank"); NB_SYM(CommDestroy, "ncclCommDestroy");
    NB_SYM(AllGather, "ncclAllGather"); NB_SYM(AllReduce, "ncclAllReduce"); NB_SYM(GetErrorString, "ncclGetErrorString");
#undef NB_SYM
    return NB_OK;
}
} // namespace nccl
#define NB_NCCL(x) do { int _r = (x); if (_r != nccl::kSuccess) return nb::fail(NB_ERR_NCCL, "NCCL error %d (%s): %s", _r, nccl::GetErrorString ? nccl::GetErrorString(_r) : "?", #x); } while (0)

static long round_up(long v, long m) { return (v + m - 1) / m * m; }
// The one-shot drivers (nb_run_benchmark / nb_run_simulation / nb_run_visualization) send systems up to this size to the
// FP64 kernel even in FP32 mode; contexts made with nb_create do so only when the "small_n_fp64" option is set.
constexpr long kSmallSystem = 8192;

} // namespace nb

using namespace nb;

struct nb_ctx {
    int device = 0, precision = NB_PREC_FP32, sm_count = 148;
    long n = 0, npad = 0; int ntiles = 0;
    int integ = NB_INTEG_VELOCITY_VERLET, force = NB_FORCE_PAIRWISE, subtract_com = 1;
    int rank = 0, world = 1; long i0 = 0, ni = 0;
    cudaStream_t stream = nullptr, comm_stream = nullptr, near_stream = nullptr, far2_stream = nullptr;
    cudaEvent_t ev_force_begin = nullptr, ev_near_done = nullptr, ev_far2_done = nullptr;
    cudaEvent_t ev_packed = nullptr, ev_gathered = nullptr, ev_user = nullptr, tev0 = nullptr, tev1 = nullptr;
    static constexpr int kRing = 64;
    cudaEvent_t fev0[kRing] = {}, fev1[kRing] = {};   // ring of force-evaluation event pairs
    cudaEvent_t nev0[kRing] = {}, nev1[kRing] = {};   // the same for the near kernel alone (near_stream)
    long fev_count = 0;
    char* slab = nullptr; size_t slab_bytes = 0;      // one allocation behind all npad-sized arrays (alloc_arrays)
    double4 *pos = nullptr, *vel = nullptr, *tmp4 = nullptr, *stage_pos = nullptr, *stage_vel = nullptr;
    unsigned int* perm = nullptr;          // perm[sorted slot] = caller index
    double* snap = nullptr;                // snapshot staging [n][6] + [n][2], allocated by the first nb_snapshot
    SortScratch ss;
    int spatial_sort = 1; float kappa = 0.25f;
    int resort_interval = 16; long steps_since_sort = 0;   // re-establish the Morton order every so many steps
    unsigned int* perm2 = nullptr;
    RenderState render_state;
    int collect_stats = 0; unsigned long long* stats = nullptr;   // optional tile-classification counters (device, 4 words)
    double *acc = nullptr, *part = nullptr, *red_part = nullptr, *red = nullptr, *eout = nullptr;
    float* pk = nullptr;
    size_t part_bytes = 0;
    double avg_mass = 0.0;
    bool uploaded = false, have_accel = false, force_timed = false;
    nccl::comm_t comm = nullptr;
    long launches = 0;
    int force_variant = 0, j_splits = 0, near_splits = 0;
    // Small systems: one captured step (drift, FP64 force, finish, COM) replayed as a CUDA graph — see step_graph_for().
    bool step_graph_enabled = false; cudaGraphExec_t step_graph = nullptr;
    struct StepKey { double dt = 0, avg_mass = 0; int integ = -1, force = -1, sub = -1, S = 0; const void* part = nullptr; } step_key;
    long small_n_fp64 = 0;         // FP32-mode context, variant auto, single GPU: systems up to this size use the FP64 kernel
};

namespace {

void free_arrays(nb_ctx* c) {
    cudaFree(c->slab); c->slab = nullptr; c->slab_bytes = 0;            // every npad-sized array lives in this one block
    cudaFree(c->part); cudaFree(c->snap);
    cudaFree(c->stats); c->stats = nullptr; c->collect_stats = 0;       // sized by npad as well
    c->pos = c->vel = c->tmp4 = c->stage_pos = c->stage_vel = nullptr;
    c->acc = c->part = c->red_part = c->snap = nullptr; c->pk = nullptr; c->perm = nullptr; c->perm2 = nullptr; c->part_bytes = 0;
    c->ss = SortScratch();
    cudaGetLastError();
}

// Everything whose size depends on npad (the padded particle count).
// One cudaMalloc for all of it: a context is created and destroyed inside every nb_run_benchmark / nb_run_simulation
// call (the reference's per-call life cycle), and ~18 separate cudaMalloc / cudaFree pairs were 8 of the 15 ms that cost.
int alloc_arrays(nb_ctx* c) {
    const size_t np = (size_t)c->npad, n = (size_t)c->n;
    c->ss.cub_bytes = sort_temp_bytes((long)n);
    char* base = nullptr;
    size_t off = 0;
    auto carve = [&](auto*& ptr, size_t count) {
        using T = std::remove_reference_t<decltype(*ptr)>;
        ptr = base ? reinterpret_cast<T*>(base + off) : nullptr;
        off += (count * sizeof(T) + 255) / 256 * 256;
    };
    auto layout = [&]() {
        off = 0;
        carve(c->pos, np); carve(c->vel, np); carve(c->tmp4, np); carve(c->stage_pos, np); carve(c->stage_vel, np);
        carve(c->acc, np * 3); carve(c->red_part, (np / kThreads) * 4);
        carve(c->perm, np); carve(c->perm2, np);
        carve(c->ss.keys_in, n); carve(c->ss.keys_out, n); carve(c->ss.idx_in, n);
        carve(c->ss.extent_part, (np / kThreads) * 6); carve(c->ss.box, 8);
        unsigned char* cub = nullptr; carve(cub, c->ss.cub_bytes ? c->ss.cub_bytes : 16); c->ss.cub_temp = cub;
        if (c->precision == NB_PREC_FP32) carve(c->pk, np * kNC); else c->pk = nullptr;
    };
    layout();                                   // sizes only
    c->slab_bytes = off;
    NB_CUDA(cudaMalloc(&c->slab, c->slab_bytes));
    base = c->slab;
    layout();                                   // pointers
    if (c->pk) NB_CUDA(cudaMemsetAsync(c->pk, 0, np * kNC * sizeof(float), c->stream));
    NB_CUDA(cudaMemsetAsync(c->pos, 0, np * sizeof(double4), c->stream));
    NB_CUDA(cudaMemsetAsync(c->vel, 0, np * sizeof(double4), c->stream));
    NB_CUDA(cudaMemsetAsync(c->acc, 0, np * 3 * sizeof(double), c->stream));
    return NB_OK;
}

// Caller-order device arrays (n rows) -> Morton-ordered, padded master state + packed tiles.
int ingest(nb_ctx* c, const double4* src_pos, const double4* src_vel) {
    int rc = build_permutation(src_pos, c->n, c->spatial_sort, c->ss, c->perm, c->stream); if (rc) return rc;
    rc = launch_gather_state(src_pos, src_vel, c->perm, c->n, c->npad, c->pos, c->vel, c->stream); if (rc) return rc;
    c->launches += 5;
    return NB_OK;
}

StepArrays arrays(nb_ctx* c) {
    StepArrays A;
    A.pos = c->pos; A.vel = c->vel; A.acc = c->acc; A.pk = c->pk; A.part = c->part;
    A.red_part = c->red_part; A.red = c->red; A.npad = c->npad; A.n = c->n; A.i0 = c->i0; A.ni = c->ni;
    return A;
}

int use_device(nb_ctx* c) { NB_CUDA(cudaSetDevice(c->device)); return NB_OK; }

int check_force_method(int force) {
    if (force == NB_FORCE_PAIRWISE || force == NB_FORCE_PAIRWISE_AVX2_FP32 || force == NB_FORCE_ADAPTIVE_MUTUAL) return NB_OK;
    if (force == NB_FORCE_BARNES_HUT)
        return fail(NB_ERR_UNSUPPORTED, "forceMethod %d (BARNES_HUT) is reference-only: not part of the B200 all-pairs hot path", force);
    return fail(NB_ERR_INVALID, "unknown forceMethod %d", force);
}

// Small systems are latency-bound, not throughput-bound: the FP64 kernel is a single launch (no near/far split, no second
// stream), faster up to ~10^4 particles (profiles/r01_auto_benchmark_B200*.csv) and exact to 1e-12.
bool use_fp64_kernel(const nb_ctx* c) {
    return c->precision == NB_PREC_FP64 ||
           (c->world == 1 && c->force_variant == 0 && c->small_n_fp64 > 0 && c->n <= c->small_n_fp64);
}

int splits_for(nb_ctx* c, int iblocks, int jcount) {
    if (c->j_splits > 0) return c->j_splits < jcount ? c->j_splits : jcount;
    // CTAs of one launch differ a lot in cost (refined tiles, FP64 hand-offs, Hilbert-curve jumps), so the j-range is cut
    // much finer than the sm_count x 2 resident-CTA slots need and the block scheduler does the balancing: measured
    // (profiles/r01_kappa_splits_sweep.json) 64 splits beat the wave-filling choice by 6 % at N = 65536, 4.5 % at 262144
    // and 1.2 % at 1M. Aim for ~128 CTAs per slot, at least 4 j-tiles per CTA (FP32 kernels), at most 64 splits, and keep the
    // partial-sum slots (3 x npad doubles each) under 8 GB.
    const long slots = (long)c->sm_count * 2;
    long S = (128 * slots + iblocks - 1) / (iblocks > 0 ? iblocks : 1);
    if (S > 64) S = 64;
    // the far kernel's bulk-copy ring wants a few tiles per CTA; the FP64 kernel does not, and for a small system its
    // serial walk over the j-tiles at FP64 latency IS the step time, so there every tile may get its own CTA
    const int min_tiles = (use_fp64_kernel(c) && c->world == 1) ? 1 : 4;      // multi-GPU keeps the setting its tests ran with
    if (S > jcount / min_tiles) S = jcount / min_tiles;
    const long per_slot = 3L * c->npad * (long)sizeof(double);
    const long mem_cap = (8L << 30) / per_slot / 3;        // far A + far B + near may each take this many
    if (S > mem_cap) S = mem_cap;
    return S < 1 ? 1 : (int)S;
}

int ensure_part(nb_ctx* c, int S_total) {
    const size_t need = (size_t)S_total * 3 * (size_t)c->npad * sizeof(double);
    if (need <= c->part_bytes) return NB_OK;
    if (c->part) { NB_CUDA(cudaStreamSynchronize(c->stream)); NB_CUDA(cudaFree(c->part)); c->part = nullptr; c->part_bytes = 0; }
    NB_CUDA(cudaMalloc(&c->part, need));
    c->part_bytes = need;
    return NB_OK;
}

int launch_one_force(nb_ctx* c, const ForceLaunch& L, const double4* pos, const float* pk, double* part) {
    int rc;
    if (!use_fp64_kernel(c)) {
        const int variant = c->force_variant == 0 ? 3 : c->force_variant;
        rc = launch_force_fp32(pk, part, L, c->force, variant, c->stream);
    } else {
        rc = launch_force_fp64(pos, part, L, c->force, c->avg_mass, c->stream);
    }
    if (rc == NB_OK) c->launches++;
    return rc;
}

// part[0..S) = partial sums of F(x) for the owned i-slice; returns S. Multi-GPU: the own j-slice is processed while the
// all-gather of the remote slices is in flight on comm_stream. FP32 production path (variant 3): the FP64 near-field
// kernel runs on near_stream next to the FP32 far kernel (different pipes) and writes its own partial slots.
int eval_force(nb_ctx* c, int* S_out) {
    const int variant = c->force_variant == 0 ? 3 : c->force_variant;
    const bool fp64k = use_fp64_kernel(c);
    const bool split_near = (!fp64k && variant == 3);
    const int R = fp64k ? force_fp64_i_tiles(c->npad, c->force) : force_i_tiles_per_cta(variant);
    const int nit = (int)(c->ni / kTJ), it0 = (int)(c->i0 / kTJ);
    const int iblocks = nit / R;
    ForceLaunch L; L.it0 = it0; L.nit = nit; L.ntiles = c->ntiles; L.npad = c->npad;
    L.kappa = c->kappa > 0.f ? c->kappa : 3.0e38f;
    L.g0 = c->avg_mass > 0.0 ? (float)(kG * c->avg_mass) : 1.f;
    L.stats = c->collect_stats ? c->stats : nullptr;
    L.stats_warps = c->collect_stats ? 1 : 0;
    const int slot = (int)(c->fev_count % nb_ctx::kRing);
    NB_CUDA(cudaEventRecord(c->fev0[slot], c->stream));
    int S_far = 0, rc;
    // The near kernel runs one CTA per (owner = 128 i-particles, j-split). Two opposite costs (measured, N = 1M spiral / 262144
    // galaxy): without splits the kernel lasts as long as its heaviest owner (an owner straddling an empty stretch of the curve
    // has a huge box and tens of times the median number of near sub-tiles: 14.3 ms at 1M with one split, 6.8 with 37), with
    // many splits most CTAs find nothing and cost only their set-up (galaxy: 3.2 ms with 64 splits, 1.8 with one). About 128
    // j-tiles per CTA, at most 64 splits, and enough CTAs to give every SM a few.
    int S_near = 0;
    if (split_near) {
        const long owners = c->ni / 128;
        long want = c->ntiles / 128;
        const long fill = (8L * c->sm_count + owners - 1) / (owners > 0 ? owners : 1);
        if (want < fill) want = fill;
        if (want > 64) want = 64;
        if (want > c->ntiles / 2) want = c->ntiles / 2;
        S_near = want < 1 ? 1 : (int)want;
        if (c->near_splits > 0) S_near = c->near_splits < c->ntiles ? c->near_splits : c->ntiles;
    }
    if (c->world == 1) {
        L.jt0 = 0; L.jcount = c->ntiles; L.s0 = 0; L.S = splits_for(c, iblocks, L.jcount);
        S_far = L.S;
        rc = ensure_part(c, S_far + S_near); if (rc) return rc;
    } else {
        const int SA = splits_for(c, iblocks, nit);
        const int SB = splits_for(c, iblocks, c->ntiles - nit);
        S_far = SA + SB;
        rc = ensure_part(c, S_far + S_near); if (rc) return rc;
    }
    // Launch order: the far kernel first, so that its CTAs take the two slots of every SM for as long as it has CTAs left; the
    // near kernel (128-thread CTAs, low-priority stream) fills what the far kernel's last wave leaves idle.
    auto launch_near = [&]() -> int {
        // needs the packed tiles (this stream so far) and, multi-GPU, the gathered slices
        NB_CUDA(cudaStreamWaitEvent(c->near_stream, c->ev_force_begin, 0));
        if (c->world > 1) NB_CUDA(cudaStreamWaitEvent(c->near_stream, c->ev_gathered, 0));
        ForceLaunch N = L; N.jt0 = 0; N.jcount = c->ntiles; N.s0 = S_far; N.S = S_near; N.stats = nullptr;
        NB_CUDA(cudaEventRecord(c->nev0[slot], c->near_stream));
        int r2 = launch_force_near64(c->pk, c->part, N, c->force, c->near_stream); if (r2) return r2;
        c->launches++;
        NB_CUDA(cudaEventRecord(c->nev1[slot], c->near_stream));
        NB_CUDA(cudaEventRecord(c->ev_near_done, c->near_stream));
        return NB_OK;
    };
    if (split_near || c->world > 1) NB_CUDA(cudaEventRecord(c->ev_force_begin, c->stream));
    if (c->world == 1) {
        rc = launch_one_force(c, L, c->pos, c->pk, c->part); if (rc) return rc;
        if (split_near) { rc = launch_near(); if (rc) return rc; }
    } else {
        const int SA = splits_for(c, iblocks, nit);
        const int SB = S_far - SA;
        ForceLaunch A = L;
        A.jt0 = it0; A.jcount = nit; A.s0 = 0; A.S = SA;
        rc = launch_one_force(c, A, c->pos, c->pk, c->part); if (rc) return rc;
        if (split_near) { rc = launch_near(); if (rc) return rc; }
        // The pass over the remote j-tiles runs on a stream of its own (after the gather): its CTAs start while the last wave
        // of the own-slice pass drains instead of waiting for it (the two passes write different partial slots).
        NB_CUDA(cudaStreamWaitEvent(c->far2_stream, c->ev_force_begin, 0));
        NB_CUDA(cudaStreamWaitEvent(c->far2_stream, c->ev_gathered, 0));
        ForceLaunch B = L;
        B.jt0 = (it0 + nit) % c->ntiles; B.jcount = c->ntiles - nit; B.s0 = SA; B.S = SB;
        cudaStream_t main_stream = c->stream;
        c->stream = c->far2_stream;                      // launch_one_force launches on c->stream
        rc = launch_one_force(c, B, c->pos, c->pk, c->part);
        c->stream = main_stream;
        if (rc) return rc;
        NB_CUDA(cudaEventRecord(c->ev_far2_done, c->far2_stream));
        NB_CUDA(cudaStreamWaitEvent(c->stream, c->ev_far2_done, 0));
    }
    if (split_near) NB_CUDA(cudaStreamWaitEvent(c->stream, c->ev_near_done, 0));
    *S_out = S_far + S_near;
    NB_CUDA(cudaEventRecord(c->fev1[slot], c->stream));
    c->fev_count++;
    c->force_timed = true;
    return NB_OK;
}

// Multi-GPU exchange of the freshly packed own slice (in place inside the full packed array).
int exchange_positions(nb_ctx* c) {
    if (c->world == 1) return NB_OK;
    NB_CUDA(cudaEventRecord(c->ev_packed, c->stream));
    NB_CUDA(cudaStreamWaitEvent(c->comm_stream, c->ev_packed, 0));
    if (c->precision == NB_PREC_FP32) {
        const size_t count = (size_t)c->ni * kNC;                           // floats per rank
        NB_NCCL(nccl::AllGather(c->pk + (size_t)c->i0 * kNC, c->pk, count, nccl::kFloat32, c->comm, c->comm_stream));
    } else {
        const size_t count = (size_t)c->ni * 4;                             // doubles per rank
        NB_NCCL(nccl::AllGather((const double*)(c->pos + c->i0), (double*)c->pos, count, nccl::kFloat64, c->comm, c->comm_stream));
    }
    NB_CUDA(cudaEventRecord(c->ev_gathered, c->comm_stream));
    return NB_OK;
}

int reduce_com(nb_ctx* c) {
    StepArrays A = arrays(c);
    int rc = launch_com_final(A, c->stream); if (rc) return rc;
    c->launches++;
    if (c->world > 1) NB_NCCL(nccl::AllReduce(c->red, c->red, 4, nccl::kFloat64, nccl::kSum, c->comm, c->stream));
    return NB_OK;
}

// Multi-GPU: make every rank's copy of the (curve-ordered) master state complete before it leaves the device.
int gather_state_all(nb_ctx* c) {
    if (c->world == 1) return NB_OK;
    const size_t count = (size_t)c->ni * 4;
    NB_NCCL(nccl::AllGather((const double*)(c->pos + c->i0), (double*)c->pos, count, nccl::kFloat64, c->comm, c->stream));
    NB_NCCL(nccl::AllGather((const double*)(c->vel + c->i0), (double*)c->vel, count, nccl::kFloat64, c->comm, c->stream));
    return NB_OK;
}

// Particles move, so the space-filling-curve order established at upload decays and with it the compactness of the tiles
// (the result stays correct — more tiles simply take the FP64 near path). Every resort_interval steps the order is rebuilt
// from the current positions: x, v, a and the caller-order permutation are permuted consistently and the tiles re-packed.
// Multi-GPU: the slices are gathered first, every rank then sorts the identical full state.
int resort(nb_ctx* c) {
    if (!c->spatial_sort) return NB_OK;
    int rc;
    if (c->world > 1) {
        NB_CUDA(cudaStreamSynchronize(c->comm_stream));
        rc = gather_state_all(c); if (rc) return rc;
        for (int k = 0; k < 3; ++k)
            NB_NCCL(nccl::AllGather(c->acc + (size_t)k * c->npad + c->i0, c->acc + (size_t)k * c->npad, (size_t)c->ni,
                                    nccl::kFloat64, c->comm, c->stream));
    }
    rc = build_permutation(c->pos, c->n, c->spatial_sort, c->ss, c->perm2, c->stream); if (rc) return rc;
    const size_t nb4 = (size_t)c->n * sizeof(double4);
    rc = launch_permute4(c->pos, c->perm2, c->n, c->stage_pos, c->stream); if (rc) return rc;
    NB_CUDA(cudaMemcpyAsync(c->pos, c->stage_pos, nb4, cudaMemcpyDeviceToDevice, c->stream));
    rc = launch_permute4(c->vel, c->perm2, c->n, c->stage_vel, c->stream); if (rc) return rc;
    NB_CUDA(cudaMemcpyAsync(c->vel, c->stage_vel, nb4, cudaMemcpyDeviceToDevice, c->stream));
    double* tmp = reinterpret_cast<double*>(c->tmp4);                       // npad double4 >= 3 npad doubles
    NB_CUDA(cudaMemcpyAsync(tmp, c->acc, (size_t)c->npad * 3 * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
    rc = launch_permute_acc(tmp, c->npad, c->perm2, c->n, c->acc, c->stream); if (rc) return rc;
    rc = launch_compose_perm(c->perm, c->perm2, c->n, c->ss.idx_in, c->stream); if (rc) return rc;
    NB_CUDA(cudaMemcpyAsync(c->perm, c->ss.idx_in, (size_t)c->n * sizeof(unsigned int), cudaMemcpyDeviceToDevice, c->stream));
    StepArrays A = arrays(c);
    rc = launch_pack(A, 0, c->npad, c->avg_mass, c->stream); if (rc) return rc;
    c->launches += 9;
    c->steps_since_sort = 0;
    return NB_OK;
}

// ---- small systems: the step as a CUDA graph -------------------------------------------------------------------------
// Below ~10^4 particles a step is four dependent, microsecond-sized kernels and the cost is launch latency. The one-shot
// drivers therefore capture that sequence once (same pointers and scalars every step) and replay it; the periodic
// re-sort stays outside the graph. Only the single-GPU FP64-kernel route is captured: one stream, no events, no allocation.
int enqueue_small_step(nb_ctx* c, double dt, int S) {
    StepArrays A = arrays(c);
    int rc = launch_drift(A, c->integ, dt, c->subtract_com != 0, c->avg_mass, c->stream); if (rc) return rc;
    ForceLaunch L;
    L.it0 = 0; L.nit = (int)(c->ni / kTJ); L.jt0 = 0; L.jcount = c->ntiles; L.ntiles = c->ntiles; L.npad = c->npad;
    L.s0 = 0; L.S = S; L.g0 = 1.f; L.stats_warps = 0; L.stats = nullptr; L.kappa = 0.f;
    rc = launch_force_fp64(c->pos, c->part, L, c->force, c->avg_mass, c->stream); if (rc) return rc;
    rc = launch_finish(A, c->integ, c->integ == NB_INTEG_VELOCITY_VERLET, dt, S, c->stream); if (rc) return rc;
    return launch_com_final(A, c->stream);
}

void drop_step_graph(nb_ctx* c) {
    if (c->step_graph) { cudaGraphExecDestroy(c->step_graph); c->step_graph = nullptr; cudaGetLastError(); }
    c->step_key = nb_ctx::StepKey();
}

bool step_graph_applies(const nb_ctx* c) {
    return c->step_graph_enabled && c->world == 1 && use_fp64_kernel(c) && c->n <= 2 * kSmallSystem && !c->collect_stats;
}

// Returns the executable graph of one step for (dt, methods, mean mass, partial-slot buffer), building it when any of those changed.
int step_graph_for(nb_ctx* c, double dt, cudaGraphExec_t* out) {
    const int S = splits_for(c, (int)(c->ni / kTJ) / force_fp64_i_tiles(c->npad, c->force), c->ntiles);
    int rc = ensure_part(c, S); if (rc) return rc;                   // never inside the capture: it may allocate
    nb_ctx::StepKey k; k.dt = dt; k.avg_mass = c->avg_mass; k.integ = c->integ; k.force = c->force; k.sub = c->subtract_com; k.S = S; k.part = c->part;
    const nb_ctx::StepKey& o = c->step_key;
    if (c->step_graph && o.dt == k.dt && o.avg_mass == k.avg_mass && o.integ == k.integ && o.force == k.force && o.sub == k.sub &&
        o.S == k.S && o.part == k.part) { *out = c->step_graph; return NB_OK; }
    drop_step_graph(c);
    cudaGraph_t g = nullptr;
    NB_CUDA(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    rc = enqueue_small_step(c, dt, S);
    const cudaError_t e = cudaStreamEndCapture(c->stream, &g);        // always close the capture, also after a failed enqueue
    if (rc) { if (g) cudaGraphDestroy(g); cudaGetLastError(); return rc; }
    if (e != cudaSuccess) return cuda_fail(e, "cudaStreamEndCapture", __FILE__, __LINE__);
    const cudaError_t ei = cudaGraphInstantiate(&c->step_graph, g, 0);
    cudaGraphDestroy(g);
    if (ei != cudaSuccess) { c->step_graph = nullptr; return cuda_fail(ei, "cudaGraphInstantiate", __FILE__, __LINE__); }
    c->step_key = k;
    *out = c->step_graph;
    return NB_OK;
}

int compute_accel(nb_ctx* c, bool kick, double dt) {
    int S = 0;
    int rc = eval_force(c, &S); if (rc) return rc;
    StepArrays A = arrays(c);
    rc = launch_finish(A, c->integ, kick, dt, S, c->stream); if (rc) return rc;
    c->launches++;
    return reduce_com(c);
}

// ---- FILE_CSV runs: snapshots that do not stop the step loop ---------------------------------------------------------
// The reference takes its snapshots on a second stream and never synchronises the host inside the loop
// (cuda/SimulationKernels.cu:55-65,204-217). Here: at a report point the step stream copies the state (caller-order
// [n][6] rows through the permutation, plus the curve-order positions / velocities the energy needs) into one of kDepth
// slots and records an event; everything else — the O(N^2) pair potential, the reductions, the device-to-host copies into
// pinned memory — runs on a low-priority data stream while the next steps are already executing. The host only waits
// when it reuses a slot (the snapshot taken kDepth report points earlier) and once at the end.
struct SnapRing {
    static constexpr int kDepth = 3;
    nb_ctx* c = nullptr;
    cudaStream_t stream = nullptr;
    char* dev = nullptr; char* host = nullptr;
    double4 *dpos[kDepth] = {}, *dvel[kDepth] = {};
    double *d6[kDepth] = {}, *dpot[kDepth] = {}, *dred[kDepth] = {}, *dout[kDepth] = {};
    double *h6[kDepth] = {}, *he[kDepth] = {};
    cudaEvent_t ready[kDepth] = {}, done[kDepth] = {};
    double* dst6[kDepth] = {}; double* dstsys[kDepth] = {}; double when[kDepth] = {}; bool busy[kDepth] = {};
    long taken = 0;

    int init(nb_ctx* ctx) {
        c = ctx;
        const size_t np = (size_t)c->npad, n = (size_t)c->n;
        auto up = [](size_t b) { return (b + 255) / 256 * 256; };
        const size_t b_state = up(np * sizeof(double4)), b6 = up(n * 6 * sizeof(double)), b_pot = up(4 * np * sizeof(double)),
                     b_red = up((np / kThreads) * 4 * sizeof(double)), b_out = 256;
        const size_t per_dev = 2 * b_state + b6 + b_pot + b_red + b_out, per_host = b6 + 256;
        int lo = 0, hi = 0;
        NB_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        NB_CUDA(cudaStreamCreateWithPriority(&stream, cudaStreamNonBlocking, lo));
        NB_CUDA(cudaMalloc(&dev, per_dev * kDepth));
        NB_CUDA(cudaMallocHost(&host, per_host * kDepth));
        for (int k = 0; k < kDepth; ++k) {
            char* d = dev + per_dev * k;
            dpos[k] = reinterpret_cast<double4*>(d); d += b_state;
            dvel[k] = reinterpret_cast<double4*>(d); d += b_state;
            d6[k] = reinterpret_cast<double*>(d); d += b6;
            dpot[k] = reinterpret_cast<double*>(d); d += b_pot;
            dred[k] = reinterpret_cast<double*>(d); d += b_red;
            dout[k] = reinterpret_cast<double*>(d);
            h6[k] = reinterpret_cast<double*>(host + per_host * k);
            he[k] = reinterpret_cast<double*>(host + per_host * k + b6);
            NB_CUDA(cudaEventCreateWithFlags(&ready[k], cudaEventDisableTiming));
            NB_CUDA(cudaEventCreateWithFlags(&done[k], cudaEventDisableTiming));
        }
        return NB_OK;
    }
    // hand a finished slot to the caller's arrays (host side)
    int flush(int k) {
        if (!busy[k]) return NB_OK;
        NB_CUDA(cudaEventSynchronize(done[k]));
        memcpy(dst6[k], h6[k], (size_t)c->n * 6 * sizeof(double));
        dstsys[k][0] = when[k]; dstsys[k][1] = he[k][0];
        busy[k] = false;
        return NB_OK;
    }
    // snapshot of the state the step stream has reached; particleData6 / systemData2 are filled by a later flush()
    int take(double time, double* particleData6, double* systemData2) {
        const int k = (int)(taken % kDepth);
        int rc = flush(k); if (rc) return rc;
        const size_t np = (size_t)c->npad;
        rc = launch_snapshot_scatter(c->pos, c->vel, c->perm, c->n, d6[k], nullptr, c->stream); if (rc) return rc;
        NB_CUDA(cudaMemcpyAsync(dpos[k], c->pos, np * sizeof(double4), cudaMemcpyDeviceToDevice, c->stream));
        NB_CUDA(cudaMemcpyAsync(dvel[k], c->vel, np * sizeof(double4), cudaMemcpyDeviceToDevice, c->stream));
        NB_CUDA(cudaEventRecord(ready[k], c->stream));
        NB_CUDA(cudaStreamWaitEvent(stream, ready[k], 0));
        const int psplits = potential_splits(c->ntiles, c->sm_count);
        rc = launch_potential_fp64(dpos[k], dpot[k], 0, c->ntiles, c->ntiles, c->npad, psplits, stream); if (rc) return rc;
        StepArrays A = arrays(c);
        A.pos = dpos[k]; A.vel = dvel[k]; A.red_part = dred[k];
        rc = launch_energy_partial(A, dpot[k], psplits, 0, c->n, dout[k] + 4, stream); if (rc) return rc;
        rc = launch_energy_final(dout[k] + 4, dout[k], stream); if (rc) return rc;
        c->launches += 5;
        NB_CUDA(cudaMemcpyAsync(h6[k], d6[k], (size_t)c->n * 6 * sizeof(double), cudaMemcpyDeviceToHost, stream));
        NB_CUDA(cudaMemcpyAsync(he[k], dout[k], 3 * sizeof(double), cudaMemcpyDeviceToHost, stream));
        NB_CUDA(cudaEventRecord(done[k], stream));
        dst6[k] = particleData6; dstsys[k] = systemData2; when[k] = time; busy[k] = true;
        ++taken;
        return NB_OK;
    }
    int drain() {
        // oldest first, so that a failure leaves a prefix of complete snapshots
        for (long q = taken >= kDepth ? taken - kDepth : 0; q < taken; ++q) { int rc = flush((int)(q % kDepth)); if (rc) return rc; }
        return NB_OK;
    }
    ~SnapRing() {
        if (stream) { cudaStreamSynchronize(stream); cudaStreamDestroy(stream); }
        for (int k = 0; k < kDepth; ++k) { if (ready[k]) cudaEventDestroy(ready[k]); if (done[k]) cudaEventDestroy(done[k]); }
        cudaFree(dev); cudaFreeHost(host);
        cudaGetLastError();
    }
};

} // namespace

extern "C" {

const char* nb_last_error(void) { return g_err; }
int nb_abi_version(void) { return 1; }
int nb_built_sm(void) { return 100; }

int nb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int nb_partition_bounds(long n, int rank, int world, long* i_begin, long* i_count) {
    if (n <= 0 || world <= 0 || rank < 0 || rank >= world || !i_begin || !i_count)
        return fail(NB_ERR_INVALID, "nb_partition_bounds: bad arguments n=%ld rank=%d world=%d", n, rank, world);
    // Equal padded slices (the all-gather needs the same count on every rank); each a multiple of kAlign particles.
    const long per = round_up(round_up(n, world) / world, kAlign);
    long b = per * rank, e = b + per;
    if (b > n) b = n;
    if (e > n) e = n;
    *i_begin = b; *i_count = e - b;
    return NB_OK;
}

int nb_create(nb_ctx** out, long n, int device, int precision) {
    if (!out) return fail(NB_ERR_INVALID, "nb_create: out is NULL");
    *out = nullptr;
    if (n <= 0 || n > (1L << 30)) return fail(NB_ERR_INVALID, "nb_create: n=%ld out of range", n);
    if (precision != NB_PREC_FP32 && precision != NB_PREC_FP64) return fail(NB_ERR_INVALID, "nb_create: unknown precision %d", precision);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(NB_ERR_NO_DEVICE, "no CUDA device available (there is no CPU fallback)"); }
    if (device < 0) { NB_CUDA(cudaGetDevice(&device)); }
    if (device >= ndev) return fail(NB_ERR_INVALID, "nb_create: device %d of %d", device, ndev);
    NB_CUDA(cudaSetDevice(device));
    int cc_major = 0, cc_minor = 0, sm_count = 0;      // three attribute queries: cudaGetDeviceProperties costs milliseconds per call
    NB_CUDA(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, device));
    NB_CUDA(cudaDeviceGetAttribute(&cc_minor, cudaDevAttrComputeCapabilityMinor, device));
    NB_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device));
    if (cc_major != 10) return fail(NB_ERR_NO_DEVICE, "device %d is sm_%d%d; this library contains sm_100a code only", device, cc_major, cc_minor);
    nb_ctx* c = new nb_ctx();
    c->device = device; c->precision = precision; c->n = n;
    if (sm_count > 0) c->sm_count = sm_count;
    c->npad = round_up(n, kAlign); c->ntiles = (int)(c->npad / kTJ);
    c->i0 = 0; c->ni = c->npad;
    auto bail = [&](int rc) { nb_destroy(c); return rc; };
#define NB_TRY(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return bail(cuda_fail(_e, #x, __FILE__, __LINE__)); } while (0)
    int prio_lo = 0, prio_hi = 0;      // numerically greatest = lowest priority
    NB_TRY(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    NB_TRY(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prio_hi));
    NB_TRY(cudaStreamCreateWithPriority(&c->comm_stream, cudaStreamNonBlocking, prio_hi));
    NB_TRY(cudaStreamCreateWithPriority(&c->far2_stream, cudaStreamNonBlocking, prio_hi));
    NB_TRY(cudaStreamCreateWithPriority(&c->near_stream, cudaStreamNonBlocking, prio_lo));   // fills what the far kernel leaves
    NB_TRY(cudaEventCreateWithFlags(&c->ev_force_begin, cudaEventDisableTiming));
    NB_TRY(cudaEventCreateWithFlags(&c->ev_near_done, cudaEventDisableTiming));
    NB_TRY(cudaEventCreateWithFlags(&c->ev_far2_done, cudaEventDisableTiming));
    NB_TRY(cudaEventCreateWithFlags(&c->ev_packed, cudaEventDisableTiming));
    NB_TRY(cudaEventCreateWithFlags(&c->ev_gathered, cudaEventDisableTiming));
    NB_TRY(cudaEventCreateWithFlags(&c->ev_user, cudaEventDisableTiming));
    NB_TRY(cudaEventCreate(&c->tev0)); NB_TRY(cudaEventCreate(&c->tev1));
    for (int k = 0; k < nb_ctx::kRing; ++k) { NB_TRY(cudaEventCreate(&c->fev0[k])); NB_TRY(cudaEventCreate(&c->fev1[k])); }
    for (int k = 0; k < nb_ctx::kRing; ++k) { NB_TRY(cudaEventCreate(&c->nev0[k])); NB_TRY(cudaEventCreate(&c->nev1[k])); }
    NB_TRY(cudaMalloc(&c->red, 16 * sizeof(double)));
    NB_TRY(cudaMalloc(&c->eout, 16 * sizeof(double)));
    NB_TRY(cudaMemsetAsync(c->red, 0, 16 * sizeof(double), c->stream));
    { int rc = alloc_arrays(c); if (rc) return bail(rc); }
#undef NB_TRY
    *out = c;
    return NB_OK;
}

int nb_destroy(nb_ctx* c) {
    if (!c) return NB_OK;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->comm_stream) cudaStreamSynchronize(c->comm_stream);
    if (c->near_stream) cudaStreamSynchronize(c->near_stream);
    if (c->far2_stream) cudaStreamSynchronize(c->far2_stream);
    if (c->comm && nccl::CommDestroy) nccl::CommDestroy(c->comm);
    drop_step_graph(c);
    free_arrays(c);
    cudaFree(c->red); cudaFree(c->eout);
    if (c->ev_packed) cudaEventDestroy(c->ev_packed);
    if (c->ev_gathered) cudaEventDestroy(c->ev_gathered);
    if (c->ev_user) cudaEventDestroy(c->ev_user);
    if (c->tev0) cudaEventDestroy(c->tev0);
    if (c->tev1) cudaEventDestroy(c->tev1);
    for (int k = 0; k < nb_ctx::kRing; ++k) { if (c->fev0[k]) cudaEventDestroy(c->fev0[k]); if (c->fev1[k]) cudaEventDestroy(c->fev1[k]); }
    for (int k = 0; k < nb_ctx::kRing; ++k) { if (c->nev0[k]) cudaEventDestroy(c->nev0[k]); if (c->nev1[k]) cudaEventDestroy(c->nev1[k]); }
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->comm_stream) cudaStreamDestroy(c->comm_stream);
    if (c->near_stream) cudaStreamDestroy(c->near_stream);
    if (c->far2_stream) cudaStreamDestroy(c->far2_stream);
    if (c->ev_far2_done) cudaEventDestroy(c->ev_far2_done);
    if (c->ev_force_begin) cudaEventDestroy(c->ev_force_begin);
    if (c->ev_near_done) cudaEventDestroy(c->ev_near_done);
    cudaGetLastError();
    delete c;
    return NB_OK;
}

int nb_set_methods(nb_ctx* c, int integ, int force, int subtract_com) {
    if (!c) return fail(NB_ERR_INVALID, "nb_set_methods: ctx is NULL");
    if (integ != NB_INTEG_EULER && integ != NB_INTEG_VELOCITY_VERLET) return fail(NB_ERR_INVALID, "unknown integrationMethod %d", integ);
    int rc = check_force_method(force); if (rc) return rc;
    if (force != c->force) c->have_accel = false;
    c->integ = integ; c->force = force; c->subtract_com = subtract_com ? 1 : 0;
    return NB_OK;
}

int nb_set_option(nb_ctx* c, const char* key, long value) {
    if (!c || !key) return fail(NB_ERR_INVALID, "nb_set_option: NULL argument");
    if (!strcmp(key, "force_variant")) { if (value < 0 || value > 3) return fail(NB_ERR_INVALID, "force_variant must be 0..3"); c->force_variant = (int)value; return NB_OK; }
    if (!strcmp(key, "near_splits")) { if (value < 0 || value > 64) return fail(NB_ERR_INVALID, "near_splits must be 0..64"); c->near_splits = (int)value; return NB_OK; }
    if (!strcmp(key, "j_splits")) { if (value < 0 || value > 64) return fail(NB_ERR_INVALID, "j_splits must be 0..64"); c->j_splits = (int)value; return NB_OK; }
    if (!strcmp(key, "spatial_sort")) { if (value < 0 || value > 2) return fail(NB_ERR_INVALID, "spatial_sort must be 0..2"); c->spatial_sort = (int)value; return NB_OK; }
    if (!strcmp(key, "collect_stats")) {
        int rc = use_device(c); if (rc) return rc;
        // 8 counters + one word per (warp, j-split) of the far kernel (at most 64 splits)
        const size_t words = 8 + (size_t)(c->npad / 128) * 64;
        if (!c->stats) NB_CUDA(cudaMalloc(&c->stats, words * sizeof(unsigned long long)));
        NB_CUDA(cudaMemsetAsync(c->stats, 0, words * sizeof(unsigned long long), c->stream));
        c->collect_stats = value ? 1 : 0;
        return NB_OK;
    }
    if (!strcmp(key, "resort_interval")) { if (value < 0) return fail(NB_ERR_INVALID, "resort_interval must be >= 0"); c->resort_interval = (int)value; return NB_OK; }
    if (!strcmp(key, "step_graph")) { c->step_graph_enabled = value != 0; if (!value) drop_step_graph(c); return NB_OK; }
    if (!strcmp(key, "small_n_fp64")) { if (value < 0) return fail(NB_ERR_INVALID, "small_n_fp64 must be >= 0"); c->small_n_fp64 = value; return NB_OK; }
    if (!strcmp(key, "far_kappa_milli")) { if (value < 0) return fail(NB_ERR_INVALID, "far_kappa_milli must be >= 0"); c->kappa = (float)value * 1e-3f; return NB_OK; }
    return fail(NB_ERR_INVALID, "nb_set_option: unknown key '%s'", key);
}

int nb_comm_unique_id(void* id128) {
    if (!id128) return fail(NB_ERR_INVALID, "nb_comm_unique_id: NULL");
    int rc = nccl::load(); if (rc) return rc;
    nccl::unique_id id; memset(&id, 0, sizeof(id));
    NB_NCCL(nccl::GetUniqueId(&id));
    memcpy(id128, &id, 128);
    return NB_OK;
}

int nb_comm_init(nb_ctx* c, const void* id128, int rank, int world) {
    if (!c || !id128) return fail(NB_ERR_INVALID, "nb_comm_init: NULL argument");
    if (world < 1 || rank < 0 || rank >= world) return fail(NB_ERR_INVALID, "nb_comm_init: rank %d of %d", rank, world);
    if (c->uploaded) return fail(NB_ERR_INVALID, "nb_comm_init must precede nb_upload");
    if (c->comm) return fail(NB_ERR_INVALID, "nb_comm_init: this context already has a communicator");
    int rc = use_device(c); if (rc) return rc;
    c->rank = rank; c->world = world;
    if (world == 1) { c->i0 = 0; c->ni = c->npad; return NB_OK; }
    // equal, tile-aligned slices; the padded system grows to world * per
    const long per = round_up(round_up(c->n, world) / world, kAlign);
    const long npad = per * world;
    if (npad != c->npad) {             // equal slices need a larger padding: re-allocate everything sized by npad
        NB_CUDA(cudaStreamSynchronize(c->stream));
        free_arrays(c);
        c->npad = npad; c->ntiles = (int)(npad / kTJ);
        rc = alloc_arrays(c); if (rc) return rc;
    }
    c->i0 = per * rank; c->ni = per;
    rc = nccl::load(); if (rc) return rc;
    nccl::unique_id id; memcpy(&id, id128, 128);
    NB_NCCL(nccl::CommInitRank(&c->comm, world, id, rank));
    // NCCL connects its peers lazily, inside the first collective of each kind — seconds of host-side set-up that would
    // otherwise land in the first step (measured at N = 16M on two GPUs: the rank that ran it waited 5.7 s of a 53 s step for its
    // remote slices). Run one tiny all-gather and all-reduce now, on the streams the step uses, while nothing else is queued.
    // The all-gather is the step's own (same buffer, same size: NCCL picks protocol and channel count by size, and a small
    // message would leave the large-message path unconnected — at N = 16M on four GPUs one rank still lost 1.4 s of its first step
    // after a one-element warm-up); the arrays hold nothing yet (no upload so far, checked above).
    if (c->precision == NB_PREC_FP32)
        NB_NCCL(nccl::AllGather(c->pk + (size_t)c->i0 * kNC, c->pk, (size_t)c->ni * kNC, nccl::kFloat32, c->comm, c->comm_stream));
    else
        NB_NCCL(nccl::AllGather((const double*)(c->pos + c->i0), (double*)c->pos, (size_t)c->ni * 4, nccl::kFloat64, c->comm, c->comm_stream));
    NB_NCCL(nccl::AllReduce(c->red, c->red, 4, nccl::kFloat64, nccl::kSum, c->comm, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->comm_stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    NB_CUDA(cudaMemsetAsync(c->red, 0, 16 * sizeof(double), c->stream));
    return NB_OK;
}

int nb_upload(nb_ctx* c, const double* posMass4, const double* vel4) {
    if (!c || !posMass4 || !vel4) return fail(NB_ERR_INVALID, "nb_upload: NULL argument");
    int rc = use_device(c); if (rc) return rc;
    if (c->world == 1) {
        NB_CUDA(cudaMemcpyAsync(c->stage_pos, posMass4, (size_t)c->n * sizeof(double4), cudaMemcpyHostToDevice, c->stream));
        NB_CUDA(cudaMemcpyAsync(c->stage_vel, vel4, (size_t)c->n * sizeof(double4), cudaMemcpyHostToDevice, c->stream));
    } else {
        // Every rank was handed the whole system, but copies only its share of the rows over PCIe (caller order, chunk q =
        // ceil(n / world)); the chunks are then all-gathered over NVLink (64 B x n x (world-1)/world per rank at hundreds of GB/s
        // instead of 64 B x n per rank through the host's memory system, all ranks at once).
        const long q = (c->n + c->world - 1) / c->world;               // q * world <= npad (slices are >= q, rounded up)
        const long b = q * c->rank < c->n ? q * c->rank : c->n;
        const long e = b + q < c->n ? b + q : c->n;
        if (e > b) {
            NB_CUDA(cudaMemcpyAsync(c->stage_pos + b, posMass4 + 4 * b, (size_t)(e - b) * sizeof(double4), cudaMemcpyHostToDevice, c->stream));
            NB_CUDA(cudaMemcpyAsync(c->stage_vel + b, vel4 + 4 * b, (size_t)(e - b) * sizeof(double4), cudaMemcpyHostToDevice, c->stream));
        }
        NB_NCCL(nccl::AllGather((const double*)(c->stage_pos + q * c->rank), (double*)c->stage_pos, (size_t)q * 4, nccl::kFloat64, c->comm, c->stream));
        NB_NCCL(nccl::AllGather((const double*)(c->stage_vel + q * c->rank), (double*)c->stage_vel, (size_t)q * 4, nccl::kFloat64, c->comm, c->stream));
    }
    // Curve order and the gather into the master arrays need no mass statistics: enqueue them first, so that the device is
    // busy (copy, all-gather, key sort) while the host walks the masses.
    rc = ingest(c, c->stage_pos, c->stage_vel); if (rc) return rc;
    // computeAverageMass (core/Forces.cpp:85-95): serial sum in index order on the host, exactly as the reference.
    double total = 0.0;
    long bad = -1;
    for (long i = 0; i < c->n; ++i) { const double m = posMass4[4 * i + 3]; total += m; if (!(m >= 0.0) && bad < 0) bad = i; }
    // The FP32 far field works in a mass-scaled frame (k_j = sqrt(G0 / (G m_j))): a negative or NaN mass would turn into NaN
    // accelerations for everybody. The reference's arithmetic accepts such input; FP64 mode here does too.
    if (bad >= 0 && c->precision == NB_PREC_FP32) {
        NB_CUDA(cudaStreamSynchronize(c->stream));
        c->uploaded = false; c->have_accel = false;
        return fail(NB_ERR_INVALID, "nb_upload: particle %ld has mass %g; FP32 mode needs masses >= 0 (use NB_PREC_FP64)", bad, posMass4[4 * bad + 3]);
    }
    c->avg_mass = total / (double)c->n;
    StepArrays A = arrays(c);
    rc = launch_pack(A, 0, c->npad, c->avg_mass, c->stream); if (rc) return rc;
    c->launches++;
    c->uploaded = true; c->have_accel = false; c->steps_since_sort = 0;
    NB_CUDA(cudaStreamSynchronize(c->stream));      // the host buffers may be reused as soon as we return
    return NB_OK;
}

int nb_init_spiral(nb_ctx* c, unsigned long long seed, double total_mass, int arms, double radius, double thickness,
                   double perturbation, double* normal3) {
    if (!c) return fail(NB_ERR_INVALID, "nb_init_spiral: ctx is NULL");
    int rc = use_device(c); if (rc) return rc;
    // every rank generates the whole system (O(N), identical on every device by construction); no host arrays, no H2D
    rc = launch_spiral_init(seed, c->n, total_mass, arms, radius, thickness, perturbation, c->stage_pos, c->stage_vel, normal3, c->stream);
    if (rc) return rc;
    rc = ingest(c, c->stage_pos, c->stage_vel); if (rc) return rc;
    // computeAverageMass (core/Forces.cpp:85-95) of n equal masses, summed one by one like the reference does
    const double m = total_mass / (double)c->n;
    double total = 0.0;
    for (long i = 0; i < c->n; ++i) total += m;
    c->avg_mass = total / (double)c->n;
    StepArrays A = arrays(c);
    rc = launch_pack(A, 0, c->npad, c->avg_mass, c->stream); if (rc) return rc;
    c->launches += 2;
    c->uploaded = true; c->have_accel = false; c->steps_since_sort = 0;
    NB_CUDA(cudaStreamSynchronize(c->stream));
    return NB_OK;
}

int nb_download(nb_ctx* c, double* posMass4, double* vel4) {
    if (!c || !posMass4 || !vel4) return fail(NB_ERR_INVALID, "nb_download: NULL argument");
    if (!c->uploaded) return fail(NB_ERR_INVALID, "nb_download before nb_upload");
    int rc = use_device(c); if (rc) return rc;
    if (c->world > 1) NB_CUDA(cudaStreamSynchronize(c->comm_stream));
    rc = gather_state_all(c); if (rc) return rc;
    rc = launch_scatter4(c->pos, c->perm, 0, c->n, c->stage_pos, c->stream); if (rc) return rc;
    rc = launch_scatter4(c->vel, c->perm, 0, c->n, c->stage_vel, c->stream); if (rc) return rc;
    c->launches += 2;
    NB_CUDA(cudaMemcpyAsync(posMass4, c->stage_pos, (size_t)c->n * sizeof(double4), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaMemcpyAsync(vel4, c->stage_vel, (size_t)c->n * sizeof(double4), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    return NB_OK;
}

int nb_download_slice(nb_ctx* c, double* posMass4, double* vel4, unsigned int* index, long capacity, long* count) {
    if (!c || !posMass4 || !vel4 || !index || !count) return fail(NB_ERR_INVALID, "nb_download_slice: NULL argument");
    if (!c->uploaded) return fail(NB_ERR_INVALID, "nb_download_slice before nb_upload");
    int rc = use_device(c); if (rc) return rc;
    long b = c->i0, e = c->i0 + c->ni; if (e > c->n) e = c->n; if (b > e) b = e;      // the real particles of the own slice
    *count = e - b;
    if (capacity < e - b) return fail(NB_ERR_INVALID, "nb_download_slice: capacity %ld < %ld rows owned by this rank", capacity, e - b);
    if (e > b) {
        // the master arrays are in curve order: the own slice is contiguous, its caller indices are perm[b..e)
        NB_CUDA(cudaMemcpyAsync(posMass4, c->pos + b, (size_t)(e - b) * sizeof(double4), cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaMemcpyAsync(vel4, c->vel + b, (size_t)(e - b) * sizeof(double4), cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaMemcpyAsync(index, c->perm + b, (size_t)(e - b) * sizeof(unsigned int), cudaMemcpyDeviceToHost, c->stream));
    }
    NB_CUDA(cudaStreamSynchronize(c->stream));
    return NB_OK;
}

int nb_compute_accel(nb_ctx* c) {
    if (!c) return fail(NB_ERR_INVALID, "nb_compute_accel: ctx is NULL");
    if (!c->uploaded) return fail(NB_ERR_INVALID, "nb_compute_accel before nb_upload");
    int rc = use_device(c); if (rc) return rc;
    // every rank packed the whole system at upload, so no exchange is needed here; make the gather event "done"
    if (c->world > 1) NB_CUDA(cudaEventRecord(c->ev_gathered, c->comm_stream));
    rc = compute_accel(c, false, 0.0); if (rc) return rc;
    c->have_accel = true;
    return NB_OK;
}

int nb_get_accel(nb_ctx* c, double* accel4) {
    if (!c || !accel4) return fail(NB_ERR_INVALID, "nb_get_accel: NULL argument");
    if (!c->have_accel) return fail(NB_ERR_INVALID, "nb_get_accel before nb_compute_accel / nb_step");
    int rc = use_device(c); if (rc) return rc;
    if (c->world > 1) {
        for (int k = 0; k < 3; ++k)
            NB_NCCL(nccl::AllGather(c->acc + (size_t)k * c->npad + c->i0, c->acc + (size_t)k * c->npad, (size_t)c->ni,
                                    nccl::kFloat64, c->comm, c->stream));
    }
    rc = launch_scatter_acc(c->acc, c->npad, c->perm, 0, c->n, c->tmp4, c->stream); if (rc) return rc;
    c->launches++;
    NB_CUDA(cudaMemcpyAsync(accel4, c->tmp4, (size_t)c->n * sizeof(double4), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    return NB_OK;
}

int nb_step(nb_ctx* c, double dt, int nsteps) {
    if (!c) return fail(NB_ERR_INVALID, "nb_step: ctx is NULL");
    if (!c->uploaded) return fail(NB_ERR_INVALID, "nb_step before nb_upload");
    if (nsteps < 0) return fail(NB_ERR_INVALID, "nb_step: nsteps=%d", nsteps);
    int rc = use_device(c); if (rc) return rc;
    if (!c->have_accel) { rc = nb_compute_accel(c); if (rc) return rc; }
    const bool vv = (c->integ == NB_INTEG_VELOCITY_VERLET);
    if (nsteps > 0 && step_graph_applies(c)) {
        cudaGraphExec_t exec = nullptr;
        rc = step_graph_for(c, dt, &exec); if (rc) return rc;
        for (int s = 0; s < nsteps; ++s) {
            if (c->resort_interval > 0 && c->steps_since_sort >= c->resort_interval) { rc = resort(c); if (rc) return rc; }
            c->steps_since_sort++;
            NB_CUDA(cudaGraphLaunch(exec, c->stream));
            c->launches += 4;
        }
        return NB_OK;
    }
    for (int s = 0; s < nsteps; ++s) {
        if (c->resort_interval > 0 && c->steps_since_sort >= c->resort_interval) { rc = resort(c); if (rc) return rc; }
        c->steps_since_sort++;
        StepArrays A = arrays(c);
        rc = launch_drift(A, c->integ, dt, c->subtract_com != 0, c->avg_mass, c->stream); if (rc) return rc;
        c->launches++;
        rc = exchange_positions(c); if (rc) return rc;
        rc = compute_accel(c, vv, dt); if (rc) return rc;
    }
    return NB_OK;
}

int nb_sync(nb_ctx* c) {
    if (!c) return fail(NB_ERR_INVALID, "nb_sync: ctx is NULL");
    int rc = use_device(c); if (rc) return rc;
    NB_CUDA(cudaStreamSynchronize(c->comm_stream));
    NB_CUDA(cudaStreamSynchronize(c->near_stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    return NB_OK;
}

int nb_energy(nb_ctx* c, double out[3]) {
    if (!c || !out) return fail(NB_ERR_INVALID, "nb_energy: NULL argument");
    if (!c->uploaded) return fail(NB_ERR_INVALID, "nb_energy before nb_upload");
    int rc = use_device(c); if (rc) return rc;
    // Multi-GPU: collective. Every rank completes its copy of the positions, sums the pair potential of its own i-slice
    // against all j and its slice's kinetic energy, and the two sums are all-reduced (SURVEY 8e "small collectives").
    if (c->world > 1) { NB_CUDA(cudaStreamSynchronize(c->comm_stream)); rc = gather_state_all(c); if (rc) return rc; }
    double* pot = reinterpret_cast<double*>(c->tmp4);                  // npad double4 = room for four partial potentials
    const int psplits = potential_splits((int)(c->ni / kTJ), c->sm_count);
    rc = launch_potential_fp64(c->pos, pot, (int)(c->i0 / kTJ), (int)(c->ni / kTJ), c->ntiles, c->npad, psplits, c->stream); if (rc) return rc;
    StepArrays A = arrays(c);
    long b = c->i0, e = c->i0 + c->ni; if (e > c->n) e = c->n; if (b > e) b = e;
    rc = launch_energy_partial(A, pot, psplits, b, e - b, c->eout + 4, c->stream); if (rc) return rc;
    if (c->world > 1) NB_NCCL(nccl::AllReduce(c->eout + 4, c->eout + 4, 4, nccl::kFloat64, nccl::kSum, c->comm, c->stream));
    rc = launch_energy_final(c->eout + 4, c->eout, c->stream); if (rc) return rc;
    c->launches += 4;
    NB_CUDA(cudaMemcpyAsync(out, c->eout, 3 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    return NB_OK;
}

int nb_momentum(nb_ctx* c, double out[4]) {
    if (!c || !out) return fail(NB_ERR_INVALID, "nb_momentum: NULL argument");
    if (!c->uploaded) return fail(NB_ERR_INVALID, "nb_momentum before nb_upload");
    int rc = use_device(c); if (rc) return rc;
    long b = c->i0, e = c->i0 + c->ni; if (e > c->n) e = c->n; if (b > e) b = e;
    // red_part is shared with the step kernels: everything is on one stream, so this is ordered after them
    rc = launch_momentum(c->pos, c->vel, b, e - b > 0 ? e - b : 1, c->red_part, c->eout + 8, c->stream); if (rc) return rc;
    c->launches += 2;
    if (c->world > 1) NB_NCCL(nccl::AllReduce(c->eout + 8, c->eout + 8, 4, nccl::kFloat64, nccl::kSum, c->comm, c->stream));
    NB_CUDA(cudaMemcpyAsync(out, c->eout + 8, 4 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    return NB_OK;
}

int nb_average_mass(nb_ctx* c, double* out) {
    if (!c || !out) return fail(NB_ERR_INVALID, "nb_average_mass: NULL argument");
    if (!c->uploaded) return fail(NB_ERR_INVALID, "nb_average_mass before nb_upload");
    *out = c->avg_mass;
    return NB_OK;
}

int nb_snapshot(nb_ctx* c, double time, double* particleData6, double* systemData2, double* staticData2) {
    if (!c || !particleData6 || !systemData2) return fail(NB_ERR_INVALID, "nb_snapshot: NULL argument");
    double e[3];
    int rc = nb_energy(c, e); if (rc) return rc;       // multi-GPU: collective, and leaves every rank with the complete state
    systemData2[0] = time; systemData2[1] = e[0];
    // stage through persistent device scratch: [n][6] then [n][2] (the reference keeps the whole [T][n][6] history on the
    // device, cuda/SimulationKernels.cu:204-249; here only one snapshot is ever resident)
    if (!c->snap) NB_CUDA(cudaMalloc(&c->snap, (size_t)c->n * 8 * sizeof(double)));
    double* d6 = c->snap; double* d2 = d6 + (size_t)c->n * 6;
    rc = launch_snapshot_scatter(c->pos, c->vel, c->perm, c->n, d6, staticData2 ? d2 : nullptr, c->stream); if (rc) return rc;
    c->launches++;
    NB_CUDA(cudaMemcpyAsync(particleData6, d6, (size_t)c->n * 6 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (staticData2) NB_CUDA(cudaMemcpyAsync(staticData2, d2, (size_t)c->n * 2 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    return NB_OK;
}

int nb_render(nb_ctx* c, int width, int height, float rot_x, float rot_y, float distance, float translate_x, float translate_y,
              float scale, float point_size, int color_mode, const float* color4, const float* clear4, unsigned char* rgba8,
              float* info4) {
    if (!c || !rgba8) return fail(NB_ERR_INVALID, "nb_render: NULL argument");
    if (!c->uploaded) return fail(NB_ERR_INVALID, "nb_render before nb_upload");
    if (c->world > 1) return fail(NB_ERR_UNSUPPORTED, "nb_render needs the whole system on one device");
    if (color_mode != 0 && color_mode != 1) return fail(NB_ERR_INVALID, "nb_render: color_mode %d", color_mode);
    int rc = use_device(c); if (rc) return rc;
    const float white[4] = {1.0f, 1.0f, 1.0f, 0.9f};            // ParticleRenderer.cpp:36-39
    const float night[4] = {0.02f, 0.02f, 0.05f, 1.0f};         // VisualizationUtils.cpp:101
    rc = render_splat(c->pos, c->vel, c->perm, c->n, width, height, rot_x, rot_y, distance, translate_x, translate_y, scale,
                      point_size > 0.f ? point_size : 30.0f, color_mode, color4 ? color4 : white, clear4 ? clear4 : night,
                      c->render_state, info4, rgba8, c->stream);
    if (!rc) c->launches += 6;
    return rc;
}

long nb_launch_count(nb_ctx* c) { return c ? c->launches : 0; }

int nb_last_force_ms(nb_ctx* c, float* ms) {
    if (!c || !ms) return fail(NB_ERR_INVALID, "nb_last_force_ms: NULL argument");
    if (!c->force_timed || c->fev_count == 0) return fail(NB_ERR_INVALID, "no force evaluation has run yet");
    int rc = use_device(c); if (rc) return rc;
    const int slot = (int)((c->fev_count - 1) % nb_ctx::kRing);
    NB_CUDA(cudaEventSynchronize(c->fev1[slot]));
    NB_CUDA(cudaEventElapsedTime(ms, c->fev0[slot], c->fev1[slot]));
    return NB_OK;
}

int nb_force_timing(nb_ctx* c, int reset, float* mean_ms, int* evals) {
    if (!c || !mean_ms || !evals) return fail(NB_ERR_INVALID, "nb_force_timing: NULL argument");
    int rc = use_device(c); if (rc) return rc;
    NB_CUDA(cudaStreamSynchronize(c->stream));
    const long cnt = c->fev_count < nb_ctx::kRing ? c->fev_count : nb_ctx::kRing;
    double sum = 0.0;
    for (long k = 0; k < cnt; ++k) {
        const int slot = (int)((c->fev_count - 1 - k) % nb_ctx::kRing);
        float ms = 0.f;
        NB_CUDA(cudaEventElapsedTime(&ms, c->fev0[slot], c->fev1[slot]));
        sum += ms;
    }
    *mean_ms = cnt ? (float)(sum / (double)cnt) : 0.f;
    *evals = (int)cnt;
    if (reset) c->fev_count = 0;
    return NB_OK;
}

int nb_near_timing(nb_ctx* c, float* mean_ms, int* evals) {
    if (!c || !mean_ms || !evals) return fail(NB_ERR_INVALID, "nb_near_timing: NULL argument");
    int rc = use_device(c); if (rc) return rc;
    NB_CUDA(cudaStreamSynchronize(c->stream));
    NB_CUDA(cudaStreamSynchronize(c->near_stream));
    *mean_ms = 0.f; *evals = 0;
    const int variant = c->force_variant == 0 ? 3 : c->force_variant;
    if (use_fp64_kernel(c) || variant != 3) return NB_OK;          // no separate near kernel on these routes
    const long cnt = c->fev_count < nb_ctx::kRing ? c->fev_count : nb_ctx::kRing;
    double sum = 0.0;
    for (long k = 0; k < cnt; ++k) {
        const int slot = (int)((c->fev_count - 1 - k) % nb_ctx::kRing);
        float ms = 0.f;
        NB_CUDA(cudaEventElapsedTime(&ms, c->nev0[slot], c->nev1[slot]));
        sum += ms;
    }
    *mean_ms = cnt ? (float)(sum / (double)cnt) : 0.f;
    *evals = (int)cnt;
    return NB_OK;
}

int nb_get_order(nb_ctx* c, unsigned int* order) {
    if (!c || !order) return fail(NB_ERR_INVALID, "nb_get_order: NULL argument");
    if (!c->uploaded) return fail(NB_ERR_INVALID, "nb_get_order before nb_upload");
    int rc = use_device(c); if (rc) return rc;
    NB_CUDA(cudaMemcpyAsync(order, c->perm, (size_t)c->n * sizeof(unsigned int), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    return NB_OK;
}

int nb_tile_stats(nb_ctx* c, unsigned long long out[5]) {
    if (!c || !out) return fail(NB_ERR_INVALID, "nb_tile_stats: NULL argument");
    if (!c->stats) return fail(NB_ERR_INVALID, "nb_tile_stats: enable with nb_set_option(ctx, \"collect_stats\", 1) first");
    int rc = use_device(c); if (rc) return rc;
    NB_CUDA(cudaMemcpyAsync(out, c->stats, 5 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    return NB_OK;
}

// Per-warp FP64 sub-tile counts of the last force evaluation (collect_stats): out[k] for k < count; returns how many
// (warp, split) entries exist through *n_out.
int nb_tile_stats_warps(nb_ctx* c, unsigned long long* out, long cap, long* n_out) {
    if (!c || !out || !n_out) return fail(NB_ERR_INVALID, "nb_tile_stats_warps: NULL argument");
    if (!c->stats) return fail(NB_ERR_INVALID, "nb_tile_stats_warps: enable collect_stats first");
    int rc = use_device(c); if (rc) return rc;
    const long words = (long)(c->npad / 128) * 64;
    const long nn = words < cap ? words : cap;
    NB_CUDA(cudaMemcpyAsync(out, c->stats + 8, (size_t)nn * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(cudaStreamSynchronize(c->stream));
    *n_out = words;
    return NB_OK;
}

int nb_timer_start(nb_ctx* c) {
    if (!c) return fail(NB_ERR_INVALID, "nb_timer_start: ctx is NULL");
    int rc = use_device(c); if (rc) return rc;
    NB_CUDA(cudaEventRecord(c->tev0, c->stream));
    return NB_OK;
}

int nb_timer_stop(nb_ctx* c, float* ms) {
    if (!c || !ms) return fail(NB_ERR_INVALID, "nb_timer_stop: NULL argument");
    int rc = use_device(c); if (rc) return rc;
    NB_CUDA(cudaEventRecord(c->tev1, c->stream));
    NB_CUDA(cudaEventSynchronize(c->tev1));
    NB_CUDA(cudaStreamSynchronize(c->comm_stream));
    NB_CUDA(cudaEventElapsedTime(ms, c->tev0, c->tev1));
    return NB_OK;
}

int nb_run_benchmark(long n, double* posMass4, double* vel4, int integ, int force, double dt, int steps,
                     int block_size, int precision, int subtract_com, int device) {
    (void)block_size;
    nb_ctx* c = nullptr;
    int rc = nb_create(&c, n, device, precision); if (rc) return rc;
    c->small_n_fp64 = kSmallSystem; c->step_graph_enabled = true;
    rc = nb_set_methods(c, integ, force, subtract_com);
    if (!rc) rc = nb_upload(c, posMass4, vel4);
    if (!rc) rc = nb_compute_accel(c);                 // cuda/BenchmarkKernels.cu:68
    if (!rc) rc = nb_step(c, dt, steps);               // :75-88
    if (!rc) rc = nb_download(c, posMass4, vel4);      // :101-102
    nb_destroy(c);
    return rc;
}

int nb_run_simulation(long n, double* posMass4, double* vel4, int integ, int force, double dt, int steps, int step_freq,
                      double* h_particleData, double* h_systemData, double* h_staticData, long num_time_steps,
                      int precision, int subtract_com, int device) {
    if (!h_particleData || !h_systemData || !h_staticData) return fail(NB_ERR_INVALID, "nb_run_simulation: NULL output array");
    if (step_freq < 1) step_freq = 1;
    if (num_time_steps < 1) return fail(NB_ERR_INVALID, "nb_run_simulation: num_time_steps=%ld", num_time_steps);
    nb_ctx* c = nullptr;
    int rc = nb_create(&c, n, device, precision); if (rc) return rc;
    c->small_n_fp64 = kSmallSystem; c->step_graph_enabled = true;
    rc = nb_set_methods(c, integ, force, subtract_com);
    if (!rc) rc = nb_upload(c, posMass4, vel4);
    if (!rc) rc = nb_compute_accel(c);                                                     // cuda/SimulationKernels.cu:167
    if (!rc) rc = nb_snapshot(c, 0.0, h_particleData, h_systemData, h_staticData);          // :171-187 (snapshot 0)
    int counter = 0; long tidx = 1;
    {
        SnapRing ring;                                                                     // destroyed before the context
        if (!rc && steps >= step_freq && num_time_steps > 1) rc = ring.init(c);
        for (int step = 1; step <= steps && !rc; ++step) {                                 // :193-228
            rc = nb_step(c, dt, 1);
            if (!rc && ++counter >= step_freq) {
                if (tidx < num_time_steps)
                    rc = ring.take(step * dt, h_particleData + (size_t)tidx * n * 6, h_systemData + tidx * 2);
                ++tidx; counter = 0;
            }
        }
        if (!rc) rc = ring.drain();
    }
    if (!rc) rc = nb_download(c, posMass4, vel4);                                          // :246-249
    nb_destroy(c);
    return rc;
}

int nb_run_visualization(long n, double* posMass4, double* vel4, int integ, int force, double dt, int steps,
                         int physics_steps_per_frame, int width, int height, float scale, const char* frame_dir,
                         int frame_stride, unsigned char* last_rgba8, long* frames2, int precision, int subtract_com, int device) {
    if (width <= 0 || height <= 0 || width > 16384 || height > 16384)
        return fail(NB_ERR_INVALID, "nb_run_visualization: image %d x %d", width, height);
    if (steps < 0) return fail(NB_ERR_INVALID, "nb_run_visualization: steps=%d", steps);
    const int per_frame = physics_steps_per_frame > 0 ? physics_steps_per_frame : 1;     // VisualizationKernels.cu:159-160
    const int stride = frame_stride > 0 ? frame_stride : 1;
    std::vector<unsigned char> img((size_t)width * height * 4);
    nb_ctx* c = nullptr;
    int rc = nb_create(&c, n, device, precision); if (rc) return rc;
    c->small_n_fp64 = kSmallSystem; c->step_graph_enabled = true;
    rc = nb_set_methods(c, integ, force, subtract_com);
    if (!rc) rc = nb_upload(c, posMass4, vel4);
    if (!rc) rc = nb_compute_accel(c);                                                   // :118-133
    long rendered = 0, written = 0;
    int done = 0;
    while (!rc && done < steps) {                                                        // :163-250
        const int k = steps - done < per_frame ? steps - done : per_frame;
        rc = nb_step(c, dt, k);
        done += k;
        float info[4] = {0.f, 0.f, 0.f, 0.f};
        if (!rc) rc = nb_render(c, width, height, 0.2f, 0.9f, 5.0f, 0.0f, 0.0f, scale, 30.0f, 1, nullptr, nullptr, img.data(), info);
        if (rc) break;
        if (scale <= 0.f) scale = info[3];              // the scale found for the first frame stays for the run
        if (frame_dir && rendered % stride == 0) {
            char path[4096];
            snprintf(path, sizeof(path), "%s/frame_%06ld.ppm", frame_dir, written);
            FILE* f = fopen(path, "wb");
            if (!f) { rc = fail(NB_ERR_INVALID, "nb_run_visualization: cannot write %s", path); break; }
            fprintf(f, "P6\n%d %d\n255\n", width, height);
            std::vector<unsigned char> row((size_t)width * 3);
            for (int y = 0; y < height; ++y) {
                const unsigned char* src = img.data() + (size_t)y * width * 4;
                for (int x = 0; x < width; ++x) { row[3 * x] = src[4 * x]; row[3 * x + 1] = src[4 * x + 1]; row[3 * x + 2] = src[4 * x + 2]; }
                fwrite(row.data(), 1, row.size(), f);
            }
            fclose(f);
            ++written;
        }
        ++rendered;
    }
    if (!rc) rc = nb_download(c, posMass4, vel4);                                         // :259-261
    if (!rc && last_rgba8 && rendered > 0) memcpy(last_rgba8, img.data(), img.size());
    if (frames2) { frames2[0] = rendered; frames2[1] = written; }
    nb_destroy(c);
    return rc;
}

int nb_launch_force(nb_ctx* c, int force_method, const double* d_posMass4, double* d_accel4, double avg_mass, long n, void* stream) {
    if (!c || !d_posMass4 || !d_accel4) return fail(NB_ERR_INVALID, "nb_launch_force: NULL argument");
    if (n != c->n) return fail(NB_ERR_INVALID, "nb_launch_force: n=%ld but the context was created for %ld", n, c->n);
    if (c->world > 1) return fail(NB_ERR_UNSUPPORTED, "nb_launch_force is single-device");
    int rc = check_force_method(force_method); if (rc) return rc;
    rc = use_device(c); if (rc) return rc;
    cudaStream_t user = (cudaStream_t)stream;
    // Order the context stream after the caller's stream and vice versa with the context's own event. A NULL stream is the
    // legacy default stream: the event recorded on it orders this call after whatever the caller enqueued there.
    NB_CUDA(cudaEventRecord(c->ev_user, user)); NB_CUDA(cudaStreamWaitEvent(c->stream, c->ev_user, 0));
    NB_CUDA(cudaMemcpyAsync(c->stage_pos, d_posMass4, (size_t)n * sizeof(double4), cudaMemcpyDeviceToDevice, c->stream));
    // The context's particle state is replaced by these positions (velocities zero): a later nb_step needs a new nb_upload.
    c->uploaded = false; c->have_accel = false;
    const int saved_force = c->force;
    if (avg_mass > 0.0) c->avg_mass = avg_mass;
    else {
        rc = launch_mass_sum(c->stage_pos, n, c->red_part, c->eout + 12, c->stream); if (rc) return rc;
        double r4[4];
        NB_CUDA(cudaMemcpyAsync(r4, c->eout + 12, 4 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        NB_CUDA(cudaStreamSynchronize(c->stream));
        c->avg_mass = r4[3] / (double)n;
    }
    c->force = force_method;
    rc = ingest(c, c->stage_pos, nullptr);
    StepArrays A = arrays(c);
    if (!rc) rc = launch_pack(A, 0, c->npad, c->avg_mass, c->stream);
    int S = 0;
    if (!rc) rc = eval_force(c, &S);
    if (!rc) rc = launch_accel_import_sum(c->part, S, c->npad, n, c->tmp4, c->stream);
    if (!rc) rc = launch_scatter4(c->tmp4, c->perm, 0, n, reinterpret_cast<double4*>(d_accel4), c->stream);
    c->force = saved_force;
    if (rc) return rc;
    c->launches += 2;
    if (user) { NB_CUDA(cudaEventRecord(c->ev_user, c->stream)); NB_CUDA(cudaStreamWaitEvent(user, c->ev_user, 0)); }
    else NB_CUDA(cudaStreamSynchronize(c->stream));
    return NB_OK;
}

} // extern "C"
