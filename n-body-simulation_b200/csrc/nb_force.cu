// nb_force.cu — all-pairs force kernels for sm_100a.
//
// Replaces calculatePairwiseForceKernel / calculateAdaptiveForcesKernel (reference cuda/ForceKernels.cu:14-96,99-170)
// with the semantics of the CPU oracle (core/Forces.cpp:98-179,301-387): a_i = sum_{j != i, r^2 >= 1e-10} G m_j d / s^3.
//
// FP32 mode, production path (variant 3) = two kernels per evaluation:
//   * force_fp32_r4_kernel — FP32 far field: j-tiles of 256 particles arrive by cp.async.bulk into a 3-stage ring (mbarrier per
//     stage, UBLKCP / SYNCS in SASS) from this GPU's own packed-tile array; every tile is re-expressed relative to the CTA's
//     origin (or, close to the CTA, per warp and 64-particle sub-tile relative to the warp's origin) so that one float per
//     coordinate carries the separation; packed FP32x2 arithmetic (FFMA2 / FMUL2: one instruction = two j-particles), four
//     i-particles per thread, mass-scaled frame (11 packed operations + 1 MUFU.RSQ per interaction); FP32 partial sums are
//     flushed every 128 j into FP64 sums;
//   * force_near64_kernel — FP64 near field: the (warp, sub-tile) pairs the far kernel's box tests leave out (self pairs,
//     pairs under the cut-off, close encounters) on the two-float positions, one owner (128 i-particles) per CTA.
//   Variants 1 (force_fp32_ld_kernel: global two-float coordinates, plain loads; also behind nb_force_noalloc) and 2
//   (force_fp32_tma_kernel: CTA frame, FP32 two-float near field in one kernel) are kept as cross-checks.
// FP64 mode (force_fp64_kernel): tiles of (x, y, z, G m) doubles, 4 (PAIRWISE) or 2 (ADAPTIVE) i-particles per thread, branch-free
//   G m s^-3 from the MUFU.RSQ64H seed; potential_fp64_kernel: the pair potential of nb_energy, every unordered pair once.
// The j-range of every launch is split over gridDim.y; partial sums go to part[s][3][npad] and are reduced in slot order by the
// integrator kernel (deterministic, no atomics).
#include "nb_common.cuh"

#include <mutex>

namespace nb {

namespace {

// Force law of a kernel instance (template parameter MODE)
constexpr int kPairwise = 0;   // PAIRWISE: unsoftened, pairs with r^2 < 1e-10 skipped           (core/Forces.cpp:98-179)
constexpr int kPlummer  = 1;   // PAIRWISE_AVX2_FP32 as the reference GPU runs it: s^2 = r^2 + 1e-10, no cut-off (cuda/ForceKernels.cu:173-204)
constexpr int kAdaptive = 2;   // ADAPTIVE_MUTUAL                                              (core/Forces.cpp:301-387)
#define IS_ADAPTIVE(m) ((m) == kAdaptive)
#define IS_PLUMMER(m) ((m) == kPlummer)

__device__ __forceinline__ float rsq_approx(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float lg2_approx(float x) { float y; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float ex2_approx(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
// Bounded wait: a lost completion traps (launch error) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) { if (++spins > (1u << 26)) __trap(); }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ---- FP64 building blocks (FP64 kernel and the FP64 near-field path of the FP32 production kernel) --------------------
__device__ __forceinline__ double rsq64_seed(double x) { double y; asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x)); return y; }

// gm * s2^(-3/2) to full double precision without the branches of the library rsqrt(): seed y0 from MUFU.RSQ64H
// (~1e-6), e = 1 - s2 y0^2, then s2^(-3/2) = y0^3 (1 + 3/2 e + 15/8 e^2 + O(e^3)). 7 DP operations.
__device__ __forceinline__ double gm_inv_r3(double s2, double gm) {
    const double y0 = rsq64_seed(s2);
    const double t = y0 * y0;
    const double e = fma(s2, -t, 1.0);
    const double sc0 = (gm * y0) * t;
    const double c = e * fma(e, 1.875, 1.5);
    return fma(sc0, c, sc0);
}
// q^(2/3) for q >= 0: y0 ~ q^(-1/3) from the FP32 MUFU (lg2/ex2), one Newton step for y = q^(-1/3) in double
// (y1 = y0 (4 - q y0^3) / 3, error ~2e-12 relative), q^(2/3) = q y1. Replaces cbrt() of core/Forces.cpp:354.
__device__ __forceinline__ double pow23(double q) {
    const float qf = fmaxf((float)q, 1e-30f);
    const double y0 = (double)ex2_approx(lg2_approx(qf) * -0.3333333333333333f);
    const double a = q * y0;
    const double b = a * (y0 * y0);
    return a * fma(b, -0.3333333333333333, 1.3333333333333333);
}

// ------------------------------------------------------------------------------------------------------------------
// Register-resident i-particles: R per thread, every value duplicated into both halves of a register pair so that one
// packed instruction handles two j-particles.
template <int R, int MODE>
struct IState {
    float2 nxh[R], nyh[R], nzh[R];     // -(hi position)
    float2 nxl[R], nyl[R], nzl[R];     // -(lo position)
    float2 mq[R];                      // m_i / (3 mean mass) (ADAPTIVE_MUTUAL)
    float2 ax[R], ay[R], az[R];        // FP32 partial sums (.x: even j, .y: odd j)
    double dax[R], day[R], daz[R];     // FP64 running sums
};

// One j4 group (four consecutive j-particles = two packed pairs) against the R i-particles.
//   hi, lo : 3 rows (x,y,z) of kTJ floats each;  gm, mq : one row each.
// FAR (only used in the i-block-relative frame of variant 2): the hi parts alone carry d to ~3e-7 and no pair of the
// tile can be a self pair or inside the cut-off, so the lo adds and the compare/select are dropped:
// 12 instead of 18 FP32-pipe operations per interaction.
template <int R, int MODE, bool FAR>
__device__ __forceinline__ void interact4(IState<R, MODE>& S, const float* __restrict__ hi, const float* __restrict__ lo,
                                          const float* __restrict__ gmrow, const float* __restrict__ mqrow, int j4) {
    const float4 XH = reinterpret_cast<const float4*>(hi + 0 * kTJ)[j4];
    const float4 YH = reinterpret_cast<const float4*>(hi + 1 * kTJ)[j4];
    const float4 ZH = reinterpret_cast<const float4*>(hi + 2 * kTJ)[j4];
    const float4 GM = reinterpret_cast<const float4*>(gmrow)[j4];
    float4 XL = make_float4(0.f, 0.f, 0.f, 0.f), YL = XL, ZL = XL;
    if (!FAR) {
        XL = reinterpret_cast<const float4*>(lo + 0 * kTJ)[j4];
        YL = reinterpret_cast<const float4*>(lo + 1 * kTJ)[j4];
        ZL = reinterpret_cast<const float4*>(lo + 2 * kTJ)[j4];
    }
    float4 MQ = make_float4(0.f, 0.f, 0.f, 0.f);
    if (IS_ADAPTIVE(MODE)) MQ = reinterpret_cast<const float4*>(mqrow)[j4];
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const float2 xjh = p ? make_float2(XH.z, XH.w) : make_float2(XH.x, XH.y);
        const float2 yjh = p ? make_float2(YH.z, YH.w) : make_float2(YH.x, YH.y);
        const float2 zjh = p ? make_float2(ZH.z, ZH.w) : make_float2(ZH.x, ZH.y);
        const float2 gm  = p ? make_float2(GM.z, GM.w) : make_float2(GM.x, GM.y);
        const float2 xjl = p ? make_float2(XL.z, XL.w) : make_float2(XL.x, XL.y);
        const float2 yjl = p ? make_float2(YL.z, YL.w) : make_float2(YL.x, YL.y);
        const float2 zjl = p ? make_float2(ZL.z, ZL.w) : make_float2(ZL.x, ZL.y);
        const float2 mqj = p ? make_float2(MQ.z, MQ.w) : make_float2(MQ.x, MQ.y);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            // d = (xjh - xih) + (xjl - xil): the hi difference is exact for close pairs (Sterbenz), the lo difference
            // restores the bits the FP32 rounding of the coordinates dropped.
            float2 dx = __fadd2_rn(xjh, S.nxh[r]);
            float2 dy = __fadd2_rn(yjh, S.nyh[r]);
            float2 dz = __fadd2_rn(zjh, S.nzh[r]);
            if (!FAR) {
                dx = __fadd2_rn(dx, __fadd2_rn(xjl, S.nxl[r]));
                dy = __fadd2_rn(dy, __fadd2_rn(yjl, S.nyl[r]));
                dz = __fadd2_rn(dz, __fadd2_rn(zjl, S.nzl[r]));
            }
            float2 r2 = __fmul2_rn(dz, dz);
            r2 = __ffma2_rn(dy, dy, r2);
            r2 = __ffma2_rn(dx, dx, r2);
            float2 s2 = r2;
            if (IS_ADAPTIVE(MODE)) {
                // eps^2 = max(eta^2 r^2 q^(2/3), eps_min^2), q = (m_i+m_j)/(3 mbar)   (core/Forces.cpp:353-356)
                //  => s^2 = max(r^2 (1 + 1e-4 q^(2/3)), r^2 + 1e-8); q^(2/3) = 2^(2/3 log2 q) on the MUFU.
                const float2 q = __fadd2_rn(mqj, S.mq[r]);
                float2 lg; lg.x = lg2_approx(q.x); lg.y = lg2_approx(q.y);
                const float kLog2Eta2 = -13.287712379549449f;   // log2(eta^2) = log2(1e-4)
                const float2 e = __ffma2_rn(lg, make_float2(0.6666666666666666f, 0.6666666666666666f),
                                            make_float2(kLog2Eta2, kLog2Eta2));
                float2 w; w.x = ex2_approx(e.x); w.y = ex2_approx(e.y);
                const float2 sa = __ffma2_rn(r2, w, r2);
                const float2 sb = __fadd2_rn(r2, make_float2(1e-8f, 1e-8f));
                s2.x = fmaxf(sa.x, sb.x); s2.y = fmaxf(sa.y, sb.y);
            }
            // PAIRWISE_AVX2_FP32 (GPU twin): Plummer softening eps^2 = 1e-10 instead of a cut-off; the self term is d = 0
            // times a finite factor. Far tiles are >= 0.1 AU away (tile_is_far), where 1e-10 is below FP32 resolution.
            if (IS_PLUMMER(MODE) && !FAR) s2 = __fadd2_rn(r2, make_float2(1e-10f, 1e-10f));
            float2 ri; ri.x = rsq_approx(s2.x); ri.y = rsq_approx(s2.y);
            // cut-off (also removes the self term, for which r2 == 0 exactly): core/Forces.cpp:135,144
            if (!FAR && !IS_PLUMMER(MODE)) {
                ri.x = r2.x < 1e-10f ? 0.f : ri.x;
                ri.y = r2.y < 1e-10f ? 0.f : ri.y;
            }
            const float2 ri2 = __fmul2_rn(ri, ri);
            float2 sc = __fmul2_rn(gm, ri);
            sc = __fmul2_rn(sc, ri2);
            S.ax[r] = __ffma2_rn(sc, dx, S.ax[r]);
            S.ay[r] = __ffma2_rn(sc, dy, S.ay[r]);
            S.az[r] = __ffma2_rn(sc, dz, S.az[r]);
        }
    }
}

template <int R, int MODE, bool FAR>
__device__ __forceinline__ void tile_pass(IState<R, MODE>& S, const float* __restrict__ hi, const float* __restrict__ lo,
                                          const float* __restrict__ gmrow, const float* __restrict__ mqrow) {
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
#pragma unroll 2
        for (int j4 = half * (kTJ / 8); j4 < (half + 1) * (kTJ / 8); ++j4)
            interact4<R, MODE, FAR>(S, hi, lo, gmrow, mqrow, j4);
#pragma unroll
        for (int r = 0; r < R; ++r) {      // flush 128 j's of FP32 partial sums into the FP64 accumulators
            S.dax[r] += (double)(S.ax[r].x + S.ax[r].y);
            S.day[r] += (double)(S.ay[r].x + S.ay[r].y);
            S.daz[r] += (double)(S.az[r].x + S.az[r].y);
            S.ax[r] = S.ay[r] = S.az[r] = make_float2(0.f, 0.f);
        }
    }
}

template <int R, int MODE>
__device__ __forceinline__ void init_acc(IState<R, MODE>& S) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
        S.ax[r] = S.ay[r] = S.az[r] = make_float2(0.f, 0.f);
        S.dax[r] = S.day[r] = S.daz[r] = 0.0;
    }
}

template <int R, int MODE>
__device__ __forceinline__ void store_part(const IState<R, MODE>& S, double* __restrict__ part, const ForceLaunch& L,
                                           int itile_first, int t, int s) {
    double* o = part + (size_t)(L.s0 + s) * 3 * (size_t)L.npad;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const size_t i = (size_t)(itile_first + r) * kTJ + t;
        o[i] = S.dax[r]; o[(size_t)L.npad + i] = S.day[r]; o[2 * (size_t)L.npad + i] = S.daz[r];
    }
}

__device__ __forceinline__ int wrap_tile(int jt, int ntiles) { return jt >= ntiles ? jt - ntiles : jt; }

// ------------------------------------------------------------------------------------------------------------------
// Variant 1 (simple, always accurate): global hi+lo coordinates for every tile, strict cut-off; j-tiles staged with
// plain vector loads through registers, double-buffered, one barrier per tile. 18 FP32-pipe operations/interaction.
template <int R, int MODE>
__global__ void __launch_bounds__(kThreads, 2)
force_fp32_ld_kernel(const float* __restrict__ pk, double* __restrict__ part, const ForceLaunch L) {
    __shared__ __align__(128) float sm[2][kTileFloats];
    const int t = threadIdx.x, s = blockIdx.y;
    const int k0 = (int)(((long)L.jcount * s) / L.S), k1 = (int)(((long)L.jcount * (s + 1)) / L.S);
    const int itile_first = L.it0 + blockIdx.x * R;
    IState<R, MODE> S;
    init_acc<R, MODE>(S);
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const float* p = pk + (size_t)(itile_first + r) * kTileFloats + t;
        float v;
        v = -p[0 * kTJ]; S.nxh[r] = make_float2(v, v);
        v = -p[1 * kTJ]; S.nyh[r] = make_float2(v, v);
        v = -p[2 * kTJ]; S.nzh[r] = make_float2(v, v);
        v = -p[4 * kTJ]; S.nxl[r] = make_float2(v, v);
        v = -p[5 * kTJ]; S.nyl[r] = make_float2(v, v);
        v = -p[6 * kTJ]; S.nzl[r] = make_float2(v, v);
        v = IS_ADAPTIVE(MODE) ? p[7 * kTJ] : 0.f; S.mq[r] = make_float2(v, v);
    }
    const float4* gsrc = reinterpret_cast<const float4*>(pk);
    constexpr int kV = 8 * kTJ / 4;                // rows 0..7 of the tile (the meta row is not needed here)
    float4 g0 = make_float4(0, 0, 0, 0), g1 = g0;
    if (k0 < k1) {
        const float4* src = gsrc + (size_t)wrap_tile(L.jt0 + k0, L.ntiles) * (kTileFloats / 4);
        reinterpret_cast<float4*>(sm[0])[t] = src[t];
        reinterpret_cast<float4*>(sm[0])[t + kThreads] = src[t + kThreads];
    }
    static_assert(kV == 2 * kThreads, "two float4 per thread");
    __syncthreads();
    for (int k = k0; k < k1; ++k) {
        const int b = (k - k0) & 1;
        const bool more = (k + 1 < k1);
        if (more) {
            const float4* src = gsrc + (size_t)wrap_tile(L.jt0 + k + 1, L.ntiles) * (kTileFloats / 4);
            g0 = src[t]; g1 = src[t + kThreads];
        }
        tile_pass<R, MODE, false>(S, sm[b], sm[b] + 4 * kTJ, sm[b] + 3 * kTJ, sm[b] + 7 * kTJ);
        if (more) {
            reinterpret_cast<float4*>(sm[b ^ 1])[t] = g0;
            reinterpret_cast<float4*>(sm[b ^ 1])[t + kThreads] = g1;
        }
        __syncthreads();
    }
    store_part<R, MODE>(S, part, L, itile_first, t, s);
}

// ------------------------------------------------------------------------------------------------------------------
// Variant 2 (production):
//   * j-tiles arrive through the bulk-copy engine (cp.async.bulk -> UBLKCP) in a kStages-deep ring, each stage with an
//     mbarrier completed by the copy's byte count; one thread produces, all threads consume;
//   * every tile is re-expressed relative to the CTA's own origin O (centre of the bounding box of its i-particles):
//     thread t turns particle t of the raw tile into x' = (xh - O) + xl while the previous tile is being computed.
//     In that frame |x'| is the distance from the i-block, so for every tile whose box is further from the i-block than
//     kappa * (half-extent_i + half-extent_j) a single float per coordinate already carries d to ~3e-7 (FAR path);
//     only the few adjacent tiles need the two-float path (x', l') produced by an error-free TwoSum, and only they
//     can contain self pairs or pairs inside the 1e-5 AU cut-off.
constexpr int kStages = 3;
constexpr int kXfRows = 6;                         // x' y' z' and l'x l'y l'z

struct CtaFrame { float O[3]; float lo[3], hi[3]; float half; };

template <int R>
__device__ __forceinline__ void load_frame(CtaFrame& F, const float* __restrict__ pk, int itile_first) {
#pragma unroll
    for (int c = 0; c < 3; ++c) { F.lo[c] = 3.0e38f; F.hi[c] = -3.0e38f; }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const float* m = pk + (size_t)(itile_first + r) * kTileFloats + kMetaRow * kTJ;
#pragma unroll
        for (int c = 0; c < 3; ++c) { F.lo[c] = fminf(F.lo[c], __ldg(m + c)); F.hi[c] = fmaxf(F.hi[c], __ldg(m + 3 + c)); }
    }
    F.half = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) { F.O[c] = 0.5f * F.lo[c] + 0.5f * F.hi[c]; F.half = fmaxf(F.half, 0.5f * (F.hi[c] - F.lo[c])); }
}
// meta: the j-tile's meta row (shared memory). CTA-uniform.
__device__ __forceinline__ bool tile_is_far(const CtaFrame& F, const float* __restrict__ meta, float kappa, float min_gap2) {
    float g2 = 0.f, hj = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float g = fmaxf(0.f, fmaxf(meta[c] - F.hi[c], F.lo[c] - meta[3 + c]));
        g2 = fmaf(g, g, g2);
        hj = fmaxf(hj, 0.5f * (meta[3 + c] - meta[c]));
    }
    const float thr = kappa * (F.half + hj);
    return (g2 > thr * thr) && (g2 > min_gap2);
}
// (h, l) = two-float value of (xh + xl) - O: TwoSum(xh, -O) is error free, the low words are then folded in.
__device__ __forceinline__ void rel_two(float xh, float xl, float O, float& h, float& l) {
    const float b = -O;
    const float s = xh + b;
    const float bb = s - xh;
    const float e = (xh - (s - bb)) + (b - bb);
    const float u = e + xl;
    h = s + u;
    l = u - (h - s);
}
template <bool FAR>
__device__ __forceinline__ void transform_tile(const float* __restrict__ raw, float* __restrict__ T, const CtaFrame& F, int t) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float xh = raw[c * kTJ + t], xl = raw[(4 + c) * kTJ + t];
        if (FAR) {
            T[c * kTJ + t] = (xh - F.O[c]) + xl;
        } else {
            float h, l; rel_two(xh, xl, F.O[c], h, l);
            T[c * kTJ + t] = h; T[(3 + c) * kTJ + t] = l;
        }
    }
}

template <int R, int MODE>
__global__ void __launch_bounds__(kThreads, 2)
force_fp32_tma_kernel(const float* __restrict__ pk, double* __restrict__ part, const ForceLaunch L) {
    __shared__ __align__(128) float raw[kStages][kTileFloats];
    __shared__ __align__(16) float T[2][kXfRows * kTJ];
    __shared__ __align__(8) uint64_t full[kStages];
    const int t = threadIdx.x, s = blockIdx.y;
    const int k0 = (int)(((long)L.jcount * s) / L.S), k1 = (int)(((long)L.jcount * (s + 1)) / L.S);
    const int cnt = k1 - k0;
    const int itile_first = L.it0 + blockIdx.x * R;
    if (t == 0) {
#pragma unroll
        for (int b = 0; b < kStages; ++b) mbar_init(&full[b], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (t == 0) {
        const int pre = cnt < kStages ? cnt : kStages;
        for (int q = 0; q < pre; ++q) {
            mbar_expect_tx(&full[q], kTileBytes);
            bulk_g2s(raw[q], pk + (size_t)wrap_tile(L.jt0 + k0 + q, L.ntiles) * kTileFloats, kTileBytes, &full[q]);
        }
    }
    CtaFrame F;
    load_frame<R>(F, pk, itile_first);
    IState<R, MODE> S;
    init_acc<R, MODE>(S);
#pragma unroll
    for (int r = 0; r < R; ++r) {      // own i-particles in the CTA frame, as two-float values
        const float* p = pk + (size_t)(itile_first + r) * kTileFloats + t;
        float h, l;
        rel_two(p[0 * kTJ], p[4 * kTJ], F.O[0], h, l); S.nxh[r] = make_float2(-h, -h); S.nxl[r] = make_float2(-l, -l);
        rel_two(p[1 * kTJ], p[5 * kTJ], F.O[1], h, l); S.nyh[r] = make_float2(-h, -h); S.nyl[r] = make_float2(-l, -l);
        rel_two(p[2 * kTJ], p[6 * kTJ], F.O[2], h, l); S.nzh[r] = make_float2(-h, -h); S.nzl[r] = make_float2(-l, -l);
        const float v = IS_ADAPTIVE(MODE) ? p[7 * kTJ] : 0.f; S.mq[r] = make_float2(v, v);
    }
    bool far_cur = false;
    if (cnt > 0) {
        mbar_wait(&full[0], 0u);
        far_cur = tile_is_far(F, raw[0] + kMetaRow * kTJ, L.kappa, IS_PLUMMER(MODE) ? 1e-2f : 1e-8f);
        if (far_cur) transform_tile<true>(raw[0], T[0], F, t); else transform_tile<false>(raw[0], T[0], F, t);
    }
    __syncthreads();
    for (int q = 0; q < cnt; ++q) {
        const int b = q % kStages;
        bool far_next = false;
        if (q + 1 < cnt) {                 // re-express the next tile while this one is being computed
            const int bn = (q + 1) % kStages;
            mbar_wait(&full[bn], (uint32_t)(((q + 1) / kStages) & 1));
            far_next = tile_is_far(F, raw[bn] + kMetaRow * kTJ, L.kappa, IS_PLUMMER(MODE) ? 1e-2f : 1e-8f);
            float* Tn = T[(q + 1) & 1];
            if (far_next) transform_tile<true>(raw[bn], Tn, F, t); else transform_tile<false>(raw[bn], Tn, F, t);
        }
        const float* Tc = T[q & 1];
        if (far_cur) tile_pass<R, MODE, true>(S, Tc, Tc + 3 * kTJ, raw[b] + 3 * kTJ, raw[b] + 7 * kTJ);
        else         tile_pass<R, MODE, false>(S, Tc, Tc + 3 * kTJ, raw[b] + 3 * kTJ, raw[b] + 7 * kTJ);
        __syncthreads();                   // every warp is done with raw[b] and T[q&1]; T[(q+1)&1] is complete
        if (t == 0 && q + kStages < cnt) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(&full[b], kTileBytes);
            bulk_g2s(raw[b], pk + (size_t)wrap_tile(L.jt0 + k0 + q + kStages, L.ntiles) * kTileFloats, kTileBytes, &full[b]);
        }
        far_cur = far_next;
    }
    store_part<R, MODE>(S, part, L, itile_first, t, s);
}

// ------------------------------------------------------------------------------------------------------------------
// Variant 3 (production): FP32 far field with FOUR i-particles per thread + FP64 near field.
//
// * Thread t owns the four CONSECUTIVE particles 4t..4t+3 of the CTA's 1024 (Morton-ordered) i-particles, so a warp
//   owns 128 consecutive particles: a compact box of its own.
// * Far field (almost every tile): the CTA-wide frame. While tile q is computed, thread t re-expresses particle t of tile
//   q+1 relative to the CTA origin O (FP32 for the far loop, FP64 for the near field). One float per coordinate carries
//   the separation to < 7e-7 as long as the tile's box is further from the CTA's box than kappa (H_cta + h_tile).
// * Tiles that fail that test (the few tens of tiles around the CTA) are refined per WARP and per SUB-TILE of 64
//   j-particles: the warp switches its registers to its own frame (origin = centre of its 128 particles), re-expresses the
//   64 j-particles into a warp-private 1 KB buffer and runs the same far loop — unless the sub-tile's box is within
//   kappa (H_warp + h_sub) of the warp's box, in which case those 4 x 64 interactions per lane are evaluated in FP64
//   (two-float positions, two-float G m, branch-free G m s^-3, cut-off and self pairs handled here). Only ~1.5 k of the
//   N j-particles of an i-particle end up in the FP64 path, independent of N.
// * FP64 there is also what keeps the result accurate when two close neighbours pull in opposite directions (pair terms
//   100x larger than their sum): those terms are exact to ~1e-15 instead of ~1e-7.
// * FP32 partial sums are flushed every 128 j into FP64 sums kept in shared memory (thread-private slots).
#ifndef NB_FAR_UNROLL
#define NB_FAR_UNROLL 2
#endif
constexpr int kFarUnroll = NB_FAR_UNROLL;   // j4 groups per iteration of the far loop
constexpr int kR4 = 4;
constexpr int kFarFloats = 4 * kTJ;                         // rows k x', k y', k z', k  (PAIRWISE/PLUMMER)  |  x', y', z', -  (ADAPTIVE)
constexpr int kXfFloats = kFarFloats;
constexpr int kWarpXfFloats = 4 * kSubTile;                 // warp-private re-expressed sub-tile (same four rows)
constexpr size_t kR4SmemBytes = (size_t)kStages * kTileBytes + 2 * (size_t)kXfFloats * 4 + (size_t)kR4 * 3 * kTJ * 8 +
                                (size_t)(kThreads / 32) * kWarpXfFloats * 4 + kStages * 8 + 8;

template <int MODE>
struct IState4 {
    float2 nx[kR4], ny[kR4], nz[kR4];   // -(x'_i) duplicated in both halves (current frame)
    float2 mq[kR4];                     // m_i / (3 mean mass) (ADAPTIVE_MUTUAL)
    float2 ax[kR4], ay[kR4], az[kR4];   // FP32 partial sums
};

// far field: the four i-particles of the thread against one j4 group. Tf: re-expressed rows with stride STRIDE,
// raw: rows of the raw tile (stride kTJ) already offset to the first j of Tf.
//   PAIRWISE / PLUMMER: mass-scaled frame — Tf holds k_j x'_j and k_j = sqrt(G0 / (G m_j)); d' = k_j (x'_j - x'_i) is one
//   FFMA2 per coordinate and (G m_j / G0) d / r^3 = d' / |d'|^3, so the multiplication by G m_j disappears
//   (11 packed operations per pair). ADAPTIVE: plain frame, G m_j and m_j/(3 mbar) from the raw tile.
// The cost of a packed instruction is set by register-file bandwidth as much as by the FMA pipe: with three distinct
// 64-bit sources (the accumulating FFMA2s) it takes 3 cycles, not 2 (tools/sass_rf_model.py).
template <int MODE, int STRIDE>
__device__ __forceinline__ void far_j4(IState4<MODE>& S, const float* __restrict__ Tf, const float* __restrict__ raw, int j4) {
    const float4 X = reinterpret_cast<const float4*>(Tf + 0 * STRIDE)[j4];
    const float4 Y = reinterpret_cast<const float4*>(Tf + 1 * STRIDE)[j4];
    const float4 Z = reinterpret_cast<const float4*>(Tf + 2 * STRIDE)[j4];
    const float4 K = IS_ADAPTIVE(MODE) ? reinterpret_cast<const float4*>(raw + 3 * kTJ)[j4]       // G m_j
                                       : reinterpret_cast<const float4*>(Tf + 3 * STRIDE)[j4];    // k_j
    float4 MQ = make_float4(0.f, 0.f, 0.f, 0.f);
    if (IS_ADAPTIVE(MODE)) MQ = reinterpret_cast<const float4*>(raw + 7 * kTJ)[j4];
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const float2 xj = p ? make_float2(X.z, X.w) : make_float2(X.x, X.y);
        const float2 yj = p ? make_float2(Y.z, Y.w) : make_float2(Y.x, Y.y);
        const float2 zj = p ? make_float2(Z.z, Z.w) : make_float2(Z.x, Z.y);
        const float2 kj = p ? make_float2(K.z, K.w) : make_float2(K.x, K.y);
        const float2 mqj = p ? make_float2(MQ.z, MQ.w) : make_float2(MQ.x, MQ.y);
#pragma unroll
        for (int r = 0; r < kR4; ++r) {
            float2 dx, dy, dz;
            if (!IS_ADAPTIVE(MODE)) {
                dx = __ffma2_rn(S.nx[r], kj, xj); dy = __ffma2_rn(S.ny[r], kj, yj); dz = __ffma2_rn(S.nz[r], kj, zj);
            } else {
                dx = __fadd2_rn(xj, S.nx[r]); dy = __fadd2_rn(yj, S.ny[r]); dz = __fadd2_rn(zj, S.nz[r]);
            }
            float2 r2 = __fmul2_rn(dz, dz);
            r2 = __ffma2_rn(dy, dy, r2);
            r2 = __ffma2_rn(dx, dx, r2);
            float2 s2 = r2;
            if (IS_ADAPTIVE(MODE)) {
                const float2 q = __fadd2_rn(mqj, S.mq[r]);
                float2 lg; lg.x = lg2_approx(q.x); lg.y = lg2_approx(q.y);
                const float kLog2Eta2 = -13.287712379549449f;
                const float2 e = __ffma2_rn(lg, make_float2(0.6666666666666666f, 0.6666666666666666f), make_float2(kLog2Eta2, kLog2Eta2));
                float2 w; w.x = ex2_approx(e.x); w.y = ex2_approx(e.y);
                const float2 sa = __ffma2_rn(r2, w, r2);
                const float2 sb = __fadd2_rn(r2, make_float2(1e-8f, 1e-8f));
                s2.x = fmaxf(sa.x, sb.x); s2.y = fmaxf(sa.y, sb.y);
            }
            float2 ri; ri.x = rsq_approx(s2.x); ri.y = rsq_approx(s2.y);
            const float2 ri2 = __fmul2_rn(ri, ri);
            float2 sc;
            if (!IS_ADAPTIVE(MODE)) sc = __fmul2_rn(ri, ri2);
            else { sc = __fmul2_rn(kj, ri); sc = __fmul2_rn(sc, ri2); }
            S.ax[r] = __ffma2_rn(sc, dx, S.ax[r]); S.ay[r] = __ffma2_rn(sc, dy, S.ay[r]); S.az[r] = __ffma2_rn(sc, dz, S.az[r]);
        }
    }
}

// FP32 partial sums -> FP64 sums in shared memory. PAIRWISE / PLUMMER sums are in units of G0 (mass-scaled frame).
template <int MODE>
__device__ __forceinline__ void flush4(IState4<MODE>& S, double* __restrict__ dacc, int t, float g0) {
    const float unit = IS_ADAPTIVE(MODE) ? 1.f : g0;
#pragma unroll
    for (int r = 0; r < kR4; ++r) {
        double* d = dacc + (size_t)r * 3 * kTJ + t;
        d[0] += (double)((S.ax[r].x + S.ax[r].y) * unit);
        d[kTJ] += (double)((S.ay[r].x + S.ay[r].y) * unit);
        d[2 * kTJ] += (double)((S.az[r].x + S.az[r].y) * unit);
        S.ax[r] = S.ay[r] = S.az[r] = make_float2(0.f, 0.f);
    }
}

struct Box3 { float lo[3], hi[3]; };
// Squared gap between box a and box (blo, bhi), and the half-extent hb of the latter. Used by BOTH the far kernel and the
// FP64 near kernel to decide which of them owns a (warp, sub-tile) pair, so every operation is pinned (no contraction
// freedom): the two kernels must reach bit-identical decisions.
__device__ __forceinline__ float box_gap2(const Box3& a, const float* __restrict__ blo, const float* __restrict__ bhi, float& hb) {
    float g2 = 0.f; hb = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float g = fmaxf(0.f, fmaxf(__fsub_rn(blo[c], a.hi[c]), __fsub_rn(a.lo[c], bhi[c])));
        g2 = __fmaf_rn(g, g, g2);
        hb = fmaxf(hb, __fmul_rn(0.5f, __fsub_rn(bhi[c], blo[c])));
    }
    return g2;
}
// true: the pair of boxes is handled by the FP32 far path; false: refine / FP64
__device__ __forceinline__ bool boxes_far(const Box3& a, float half_a, const float* __restrict__ blo, const float* __restrict__ bhi,
                                          float kappa, float min_gap2) {
    float hb;
    const float g2 = box_gap2(a, blo, bhi, hb);
    const float thr = __fmul_rn(kappa, __fadd_rn(half_a, hb));
    return (g2 > __fmul_rn(thr, thr)) && (g2 > min_gap2);
}
// union box of the kR4 i-tiles of a CTA and its half-extent (from the tiles' meta rows)
__device__ __forceinline__ void cta_box(Box3& B, float& half, float* O, const float* __restrict__ pk, int itile_first, int ntile) {
#pragma unroll
    for (int c = 0; c < 3; ++c) { B.lo[c] = 3.0e38f; B.hi[c] = -3.0e38f; }
    for (int r = 0; r < ntile; ++r) {
        const float* m = pk + (size_t)(itile_first + r) * kTileFloats + kMetaRow * kTJ;
#pragma unroll
        for (int c = 0; c < 3; ++c) { B.lo[c] = fminf(B.lo[c], __ldg(m + c)); B.hi[c] = fmaxf(B.hi[c], __ldg(m + 3 + c)); }
    }
    half = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        O[c] = __fadd_rn(__fmul_rn(0.5f, B.lo[c]), __fmul_rn(0.5f, B.hi[c]));
        half = fmaxf(half, __fmul_rn(0.5f, __fsub_rn(B.hi[c], B.lo[c])));
    }
}
// box / centre / half-extent of the 128 consecutive particles a warp owns (lane: 4 particles, hi parts)
__device__ __forceinline__ void warp_box(Box3& W, float* Ow, float& Hw, const float* __restrict__ ip) {
    const float4 xh = *reinterpret_cast<const float4*>(ip + 0 * kTJ), yh = *reinterpret_cast<const float4*>(ip + 1 * kTJ),
                 zh = *reinterpret_cast<const float4*>(ip + 2 * kTJ);
    W.lo[0] = fminf(fminf(xh.x, xh.y), fminf(xh.z, xh.w)); W.hi[0] = fmaxf(fmaxf(xh.x, xh.y), fmaxf(xh.z, xh.w));
    W.lo[1] = fminf(fminf(yh.x, yh.y), fminf(yh.z, yh.w)); W.hi[1] = fmaxf(fmaxf(yh.x, yh.y), fmaxf(yh.z, yh.w));
    W.lo[2] = fminf(fminf(zh.x, zh.y), fminf(zh.z, zh.w)); W.hi[2] = fmaxf(fmaxf(zh.x, zh.y), fmaxf(zh.z, zh.w));
    Hw = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            W.lo[c] = fminf(W.lo[c], __shfl_xor_sync(0xffffffffu, W.lo[c], o));
            W.hi[c] = fmaxf(W.hi[c], __shfl_xor_sync(0xffffffffu, W.hi[c], o));
        }
        Ow[c] = __fadd_rn(__fmul_rn(0.5f, W.lo[c]), __fmul_rn(0.5f, W.hi[c]));
        Hw = fmaxf(Hw, __fmul_rn(0.5f, __fsub_rn(W.hi[c], W.lo[c])));
    }
}

// the thread's four particles in the frame with origin O: registers S.nx/ny/nz
template <int MODE>
__device__ __forceinline__ void load_i_frame(IState4<MODE>& S, const float* __restrict__ ip /* tile row 0 at the thread's first particle */,
                                             const float* __restrict__ O) {
    const float4 xh = *reinterpret_cast<const float4*>(ip + 0 * kTJ), xl = *reinterpret_cast<const float4*>(ip + 4 * kTJ);
    const float4 yh = *reinterpret_cast<const float4*>(ip + 1 * kTJ), yl = *reinterpret_cast<const float4*>(ip + 5 * kTJ);
    const float4 zh = *reinterpret_cast<const float4*>(ip + 2 * kTJ), zl = *reinterpret_cast<const float4*>(ip + 6 * kTJ);
    const float xs[4] = {-((xh.x - O[0]) + xl.x), -((xh.y - O[0]) + xl.y), -((xh.z - O[0]) + xl.z), -((xh.w - O[0]) + xl.w)};
    const float ys[4] = {-((yh.x - O[1]) + yl.x), -((yh.y - O[1]) + yl.y), -((yh.z - O[1]) + yl.z), -((yh.w - O[1]) + yl.w)};
    const float zs[4] = {-((zh.x - O[2]) + zl.x), -((zh.y - O[2]) + zl.y), -((zh.z - O[2]) + zl.z), -((zh.w - O[2]) + zl.w)};
#pragma unroll
    for (int r = 0; r < kR4; ++r) { S.nx[r] = make_float2(xs[r], xs[r]); S.ny[r] = make_float2(ys[r], ys[r]); S.nz[r] = make_float2(zs[r], zs[r]); }
}

// Register budget: 2 CTAs x 256 threads x 126 registers fill the SM's register file. Capping the kernel at 112 registers
// would leave room for one 128-thread near-kernel CTA beside them (FP64 pipe busy while the FMA pipe is), but the far loop
// itself then runs 3.5 % slower (measured: 435.5 instead of 420.3 ms per step at N = 1M) — more than the near kernel costs.
// Also tried and not kept: the warps of this kernel evaluating their own FP64 near field once the CTA's far loop is done (so
// that the FP64 pipe works while the SM's other warps keep the FMA pipe busy): 432.7 instead of 420.5 ms per step at N = 1M,
// 2.21 instead of 2.09 ms at N = 65536 — a CTA keeps its eight warp slots until its slowest warp has finished, and the near
// work per warp is heavy-tailed.
template <int MODE>
__global__ void __launch_bounds__(kThreads, 2)
force_fp32_r4_kernel(const float* __restrict__ pk, double* __restrict__ part, const ForceLaunch L) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* raw = reinterpret_cast<float*>(smem_raw);                                        // [kStages][kTileFloats]
    float* T = raw + kStages * kTileFloats;                                                 // [2][kXfFloats]
    double* dacc = reinterpret_cast<double*>(T + 2 * kXfFloats);                            // [kR4*3][kTJ]
    float* Tw = reinterpret_cast<float*>(dacc + kR4 * 3 * kTJ) + (threadIdx.x >> 5) * kWarpXfFloats;   // warp-private [4][kSubTile]
    uint64_t* full = reinterpret_cast<uint64_t*>(reinterpret_cast<float*>(dacc + kR4 * 3 * kTJ) + (kThreads / 32) * kWarpXfFloats);
    const int t = threadIdx.x, lane = t & 31, s = blockIdx.y;
    const int k0 = (int)(((long)L.jcount * s) / L.S), k1 = (int)(((long)L.jcount * (s + 1)) / L.S);
    const int cnt = k1 - k0;
    const int itile_first = L.it0 + blockIdx.x * kR4;
    if (t == 0) {
#pragma unroll
        for (int b = 0; b < kStages; ++b) mbar_init(&full[b], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
#pragma unroll
    for (int k = 0; k < kR4 * 3; ++k) dacc[k * kTJ + t] = 0.0;
    __syncthreads();
    if (t == 0) {
        const int pre = cnt < kStages ? cnt : kStages;
        for (int q = 0; q < pre; ++q) {
            mbar_expect_tx(&full[q], kTileBytes);
            bulk_g2s(raw + q * kTileFloats, pk + (size_t)wrap_tile(L.jt0 + k0 + q, L.ntiles) * kTileFloats, kTileBytes, &full[q]);
        }
    }
    CtaFrame F;                                // CTA-wide frame: union box of the four i-tiles
    Box3 Fbox;
    cta_box(Fbox, F.half, F.O, pk, itile_first, kR4);
    // the thread's four consecutive particles: tile itile_first + t/64, offset 4 (t % 64)
    const size_t i_first = (size_t)itile_first * kTJ + 4 * (size_t)t;
    const float* ip = pk + (i_first / kTJ) * kTileFloats + (i_first % kTJ);
    const float min_gap2 = IS_PLUMMER(MODE) ? 1e-2f : 1e-8f;
    const float inv_g0 = 1.f / L.g0;
    IState4<MODE> S;
    load_i_frame<MODE>(S, ip, F.O);
    {
        const float4 m4 = IS_ADAPTIVE(MODE) ? *reinterpret_cast<const float4*>(ip + 7 * kTJ) : make_float4(0.f, 0.f, 0.f, 0.f);
        S.mq[0] = make_float2(m4.x, m4.x); S.mq[1] = make_float2(m4.y, m4.y); S.mq[2] = make_float2(m4.z, m4.z); S.mq[3] = make_float2(m4.w, m4.w);
    }
#pragma unroll
    for (int r = 0; r < kR4; ++r) S.ax[r] = S.ay[r] = S.az[r] = make_float2(0.f, 0.f);
    // the warp's own box and frame (128 consecutive particles)
    Box3 Wbox; float Ow[3], Hw;
    warp_box(Wbox, Ow, Hw, ip);
    // re-express (thread t <-> particle t) tile q into buffer q & 1: FP32 rows in the CTA frame
    auto prepare = [&](int q) {
        const float* rw = raw + (q % kStages) * kTileFloats;
        float* Tf = T + (q & 1) * kXfFloats;
        float kj = 1.f;     // k_j = sqrt(G0 / (G m_j)); massless particles get a huge finite k, their r'^2 overflows to +inf,
                            // rsqrt gives 0 and they contribute exactly nothing
        if (!IS_ADAPTIVE(MODE)) { kj = rsq_approx(fmaxf(rw[3 * kTJ + t] * inv_g0, 1e-30f)); Tf[3 * kTJ + t] = kj; }
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float xh = rw[c * kTJ + t], xl = rw[(4 + c) * kTJ + t];
            Tf[c * kTJ + t] = ((xh - F.O[c]) + xl) * kj;
        }
    };
    int near_count = 0;                        // (warp, sub-tile) pairs left to the near kernel (diagnostics)
    if (cnt > 0) { mbar_wait(&full[0], 0u); prepare(0); }
    __syncthreads();
    for (int q = 0; q < cnt; ++q) {
        const int b = q % kStages;
        if (q + 1 < cnt) {                     // re-express the next tile while this one is being computed
            mbar_wait(&full[(q + 1) % kStages], (uint32_t)(((q + 1) / kStages) & 1));
            prepare(q + 1);
        }
        const float* rw = raw + b * kTileFloats;
        const float* meta = rw + kMetaRow * kTJ;
        const float* Tf = T + (q & 1) * kXfFloats;
        const bool far_cta = boxes_far(Fbox, F.half, meta, meta + 3, L.kappa, min_gap2);
        if (L.stats && t == 0) atomicAdd(L.stats + (far_cta ? 0 : 1), 1ULL);
        if (far_cta) {
            // ---- far tile: CTA frame, 64 j4 groups
#pragma unroll 1
            for (int half = 0; half < 2; ++half) {
#pragma unroll kFarUnroll
                for (int j4 = half * (kTJ / 8); j4 < (half + 1) * (kTJ / 8); ++j4) far_j4<MODE, kTJ>(S, Tf, rw, j4);
                flush4<MODE>(S, dacc, t, L.g0);
            }
        } else {
            // ---- tile close to the CTA: per warp, per sub-tile, in the warp's own frame
            load_i_frame<MODE>(S, ip, Ow);
#pragma unroll 1
            for (int sb = 0; sb < kTJ / kSubTile; ++sb) {
                const float* sbox = meta + kSubBoxOffset + 6 * sb;
                const bool far_w = boxes_far(Wbox, Hw, sbox, sbox + 3, L.kappa, min_gap2);
                const int j0 = sb * kSubTile;
                if (L.stats && lane == 0) atomicAdd(L.stats + (far_w ? 2 : 3), 1ULL);
                if (!far_w) ++near_count;
                if (far_w) {
                    __syncwarp();
#pragma unroll
                    for (int k = lane; k < kSubTile; k += 32) {          // two particles per lane
                        const int j = j0 + k;
                        float kj = 1.f;
                        if (!IS_ADAPTIVE(MODE)) { kj = rsq_approx(fmaxf(rw[3 * kTJ + j] * inv_g0, 1e-30f)); Tw[3 * kSubTile + k] = kj; }
#pragma unroll
                        for (int c = 0; c < 3; ++c) Tw[c * kSubTile + k] = ((rw[c * kTJ + j] - Ow[c]) + rw[(4 + c) * kTJ + j]) * kj;
                    }
                    __syncwarp();
#pragma unroll 2
                    for (int j4 = 0; j4 < kSubTile / 4; ++j4) far_j4<MODE, kSubTile>(S, Tw, rw + j0, j4);
                }   // else: this (warp, sub-tile) pair belongs to force_near64_kernel
            }
            flush4<MODE>(S, dacc, t, L.g0);
            load_i_frame<MODE>(S, ip, F.O);
        }
        __syncthreads();                       // all warps done with raw[b] and T[q&1]; tile q+1 is prepared
        if (t == 0 && q + kStages < cnt) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(&full[b], kTileBytes);
            bulk_g2s(raw + b * kTileFloats, pk + (size_t)wrap_tile(L.jt0 + k0 + q + kStages, L.ntiles) * kTileFloats, kTileBytes, &full[b]);
        }
    }
    if (L.stats && lane == 0) {
        atomicMax(L.stats + 4, (unsigned long long)near_count);
        if (L.stats_warps) L.stats[8 + ((size_t)s * gridDim.x + blockIdx.x) * (kThreads / 32) + (t >> 5)] = (unsigned long long)near_count;
    }
    double* o = part + (size_t)(L.s0 + s) * 3 * (size_t)L.npad + i_first;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        double4 v;
        v.x = dacc[(0 * 3 + c) * kTJ + t]; v.y = dacc[(1 * 3 + c) * kTJ + t]; v.z = dacc[(2 * 3 + c) * kTJ + t]; v.w = dacc[(3 * 3 + c) * kTJ + t];
        *reinterpret_cast<double4*>(o + (size_t)c * L.npad) = v;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// FP64 near field of the production path: every (warp of 128 i-particles, sub-tile of 64 j-particles) pair that the far
// kernel classifies as too close for FP32 (same pinned box tests, bit-identical decisions) is evaluated here in double
// on the two-float positions and two-float G m: self pairs, pairs under the cut-off and all close encounters live here.
// Warps are independent (no CTA barrier): a warp walks the meta rows of its j-range, skips what the far kernel takes and
// stages the rare near sub-tile into a warp-private 2 KB buffer. ~0.3 % (N = 1M) to ~4 % (N = 65 536) of all pairs;
// the kernel runs on its own stream next to the far kernel — the FP64 pipe is idle there.
// Launch shape: ONE far-kernel warp (128 i-particles) per CTA of 256 threads; thread = (i-particle, half of the sub-tile's 64
// j-particles). What limits this kernel is not the FP64 pipe but the longest chain: the near work of a warp is heavy-tailed
// (a warp whose 128 particles straddle an empty stretch of the curve has a huge box and tens of times the median number of
// near sub-tiles), and with one lane per 2-4 i-particles and all 64 j that chain ran for the whole kernel. Eight warps per owner
// and 32 interactions per thread and sub-tile cut it by 4-8x; every decision is CTA-uniform (one owner), so a sub-tile is staged
// once per CTA. Launched after the far kernel on a low-priority stream it also fills the SMs the far kernel's last wave
// leaves idle.
constexpr int kNearThreads = 256;
constexpr int kNearOwner = 128;                // i-particles per CTA = one far-kernel warp

template <int MODE>
__global__ void __launch_bounds__(kNearThreads, 4)
force_near64_kernel(const float* __restrict__ pk, double* __restrict__ part, const ForceLaunch L) {
    __shared__ __align__(16) double Js[4][kSubTile];                   // x', y', z' (owner frame), G m of the staged sub-tile
    __shared__ float Mqs[kSubTile];
    __shared__ int list[kNearThreads];                                 // near tiles of the current round of 256, in order
    __shared__ int wcount[kNearThreads / 32], wcount2[kNearThreads / 32];
    __shared__ int items[4 * kNearThreads];                            // (tile << 2) | sub-tile of the round's near items, in order
    __shared__ double red[3][kNearOwner];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5, s = blockIdx.y;
    const int il = t & (kNearOwner - 1), jh = t / kNearOwner;           // my i-particle, my half of every sub-tile
    const int k0 = (int)(((long)L.jcount * s) / L.S), k1 = (int)(((long)L.jcount * (s + 1)) / L.S);
    const size_t i_base = (size_t)L.it0 * kTJ + (size_t)blockIdx.x * kNearOwner;     // the owner's first particle
    const int itile = (int)(i_base / kTJ), off = (int)(i_base % kTJ);
    const int itile_first = L.it0 + (blockIdx.x / (2 * kR4)) * kR4;     // the far kernel's CTA (four i-tiles, eight warps)
    Box3 Fbox; float Fhalf, FO[3];
    cta_box(Fbox, Fhalf, FO, pk, itile_first, kR4);
    Box3 Wbox; float Ow[3], Hw;                                         // the owner's box, as the far kernel's warp computes it
    warp_box(Wbox, Ow, Hw, pk + (size_t)itile * kTileFloats + off + 4 * lane);
    const float min_gap2 = IS_PLUMMER(MODE) ? 1e-2f : 1e-8f;
    const float* q = pk + (size_t)itile * kTileFloats + off + il;
    const double xi = ((double)q[0 * kTJ] - (double)Ow[0]) + (double)q[4 * kTJ];
    const double yi = ((double)q[1 * kTJ] - (double)Ow[1]) + (double)q[5 * kTJ];
    const double zi = ((double)q[2 * kTJ] - (double)Ow[2]) + (double)q[6 * kTJ];
    const double mqi = IS_ADAPTIVE(MODE) ? (double)q[7 * kTJ] : 0.0;
    double ax = 0.0, ay = 0.0, az = 0.0;
    for (int base = k0; base < k1; base += kNearThreads) {
        // Level 1: 256 tiles per round against the far kernel's CTA box (its first-level test); ordered compaction.
        const int kk = base + t;
        bool refine = false;
        if (kk < k1) {
            const float* m = pk + (size_t)wrap_tile(L.jt0 + kk, L.ntiles) * kTileFloats + kMetaRow * kTJ;
            refine = !boxes_far(Fbox, Fhalf, m, m + 3, L.kappa, min_gap2);
        }
        const unsigned bal = __ballot_sync(0xffffffffu, refine);
        if (lane == 0) wcount[w] = __popc(bal);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int v = 0; v < kNearThreads / 32; ++v) { const int c = wcount[v]; if (v < w) before += c; total += c; }
        if (refine) list[before + __popc(bal & ((1u << lane) - 1u))] = kk;
        __syncthreads();
        // Level 2: thread e tests the four sub-tile boxes of surviving tile e against the owner's box — all tiles of the round
        // at once, so the meta rows arrive with one memory latency instead of one per tile; ordered compaction into items.
        int mask = 0, mytile = 0;
        if (t < total) {
            mytile = list[t];
            const float* meta = pk + (size_t)wrap_tile(L.jt0 + mytile, L.ntiles) * kTileFloats + kMetaRow * kTJ;
#pragma unroll
            for (int sb = 0; sb < kTJ / kSubTile; ++sb) {
                const float* sbox = meta + kSubBoxOffset + 6 * sb;
                if (!boxes_far(Wbox, Hw, sbox, sbox + 3, L.kappa, min_gap2)) mask |= 1 << sb;      // else: far kernel, warp frame
            }
        }
        const int cnt = __popc(mask);
        int incl = cnt;                                    // inclusive prefix sum over the warp
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
        if (lane == 31) wcount2[w] = incl;
        __syncthreads();
        int ibefore = 0, n_items = 0;
#pragma unroll
        for (int v = 0; v < kNearThreads / 32; ++v) { const int c = wcount2[v]; if (v < w) ibefore += c; n_items += c; }
        {
            int pos = ibefore + incl - cnt;
#pragma unroll
            for (int sb = 0; sb < kTJ / kSubTile; ++sb) if (mask & (1 << sb)) items[pos++] = (mytile << 2) | sb;
        }
        __syncthreads();
        // Items one after the other; the j-particles of item e+1 are fetched into registers while item e is computed.
        float pre[9];
        auto prefetch = [&](int e) {
            const int it = items[e];
            const float* pj = pk + (size_t)wrap_tile(L.jt0 + (it >> 2), L.ntiles) * kTileFloats + (it & 3) * kSubTile + t;
#pragma unroll
            for (int c = 0; c < 8; ++c) pre[c] = pj[c * kTJ];
            pre[8] = pj[kGmLoRow * kTJ];
        };
        if (n_items > 0 && t < kSubTile) prefetch(0);
        for (int e = 0; e < n_items; ++e) {
            __syncthreads();                                                                 // the previous item is done with
            if (t < kSubTile) {
                Js[0][t] = ((double)pre[0] - (double)Ow[0]) + (double)pre[4];
                Js[1][t] = ((double)pre[1] - (double)Ow[1]) + (double)pre[5];
                Js[2][t] = ((double)pre[2] - (double)Ow[2]) + (double)pre[6];
                Js[3][t] = (double)pre[3] + (double)pre[8];
                if (IS_ADAPTIVE(MODE)) Mqs[t] = pre[7];
            }
            __syncthreads();
            if (e + 1 < n_items && t < kSubTile) prefetch(e + 1);
            // the interactions of one sub-tile are summed on their own and then added to the running sums
            double sx = 0.0, sy = 0.0, sz = 0.0;
            const int j0 = jh * (kSubTile / 2);
#pragma unroll 4
            for (int j = j0; j < j0 + kSubTile / 2; ++j) {
                const double dx = Js[0][j] - xi, dy = Js[1][j] - yi, dz = Js[2][j] - zi, gm = Js[3][j];
                const double r2 = dx * dx + dy * dy + dz * dz;
                double s2 = r2;
                if (IS_ADAPTIVE(MODE)) {
                    const double wq = pow23(mqi + (double)Mqs[j]);
                    s2 = fmax(fma(r2 * wq, 1e-4, r2), r2 + 1e-8);
                }
                if (IS_PLUMMER(MODE)) s2 = r2 + 1e-10;                                          // cuda/ForceKernels.cu:186
                double sc = gm_inv_r3(s2, gm);
                if (!IS_PLUMMER(MODE)) sc = (__double_as_longlong(r2) < __double_as_longlong(kR2Cut)) ? 0.0 : sc;   // cut-off, self pair
                sx = fma(sc, dx, sx); sy = fma(sc, dy, sy); sz = fma(sc, dz, sz);
            }
            ax += sx; ay += sy; az += sz;
        }
        __syncthreads();                                                                     // list / items are rewritten next round
    }
    if (jh == 1) { red[0][il] = ax; red[1][il] = ay; red[2][il] = az; }
    __syncthreads();
    if (jh == 0) {
        double* o = part + (size_t)(L.s0 + s) * 3 * (size_t)L.npad + i_base + il;
        o[0] = ax + red[0][il]; o[(size_t)L.npad] = ay + red[1][il]; o[2 * (size_t)L.npad] = az + red[2][il];
    }
}

// ------------------------------------------------------------------------------------------------------------------
// FP64 mode. Tiles of 256 particles straight from the master array, staged as (x, y, z, G m) [+ m / (3 mbar) for
// ADAPTIVE_MUTUAL] so that nothing that depends on j alone is recomputed per pair; R i-particles per thread (4 for
// throughput, 2 for small systems where a CTA's serial walk over its j-range is the step time).
// Each tile's 256 interactions are summed on their own and then added to the running sums (one rounding per tile on
// the running sum instead of 256): with the particles in curve order consecutive j pull in the same direction, the running
// sum of a j-range is large compared with the final result, and every addition to it rounds at ITS magnitude.
#ifndef NB_FP64_UNROLL4
#define NB_FP64_UNROLL4 8
#endif
#ifndef NB_FP64_UNROLL2
#define NB_FP64_UNROLL2 4
#endif
#ifndef NB_FP64_BIG_R
#define NB_FP64_BIG_R 4
#endif
constexpr int kFp64Unroll2 = NB_FP64_UNROLL2;
constexpr int kFp64Unroll4 = NB_FP64_UNROLL4;      // j-particles per iteration of the inner loop with 4 i-particles per thread

#ifndef NB_FP64_MINB
#define NB_FP64_MINB 2
#endif
template <int MODE, int R>
__global__ void __launch_bounds__(kThreads, NB_FP64_MINB)
force_fp64_kernel(const double4* __restrict__ pos, double* __restrict__ part, const ForceLaunch L, double inv3mbar) {
    __shared__ __align__(16) double4 sm[2][kTJ];
    __shared__ double smq[IS_ADAPTIVE(MODE) ? 2 : 1][IS_ADAPTIVE(MODE) ? kTJ : 1];
    const int t = threadIdx.x, s = blockIdx.y;
    const int k0 = (int)(((long)L.jcount * s) / L.S), k1 = (int)(((long)L.jcount * (s + 1)) / L.S);
    const int itile_first = L.it0 + blockIdx.x * R;
    double xi[R], yi[R], zi[R], mqi[R], ax[R], ay[R], az[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const double4 p = pos[(size_t)(itile_first + r) * kTJ + t];
        xi[r] = p.x; yi[r] = p.y; zi[r] = p.z; mqi[r] = p.w * inv3mbar; ax[r] = ay[r] = az[r] = 0.0;
    }
    double4 g = make_double4(0, 0, 0, 0);
    if (k0 < k1) {
        g = pos[(size_t)wrap_tile(L.jt0 + k0, L.ntiles) * kTJ + t];
        if (IS_ADAPTIVE(MODE)) smq[0][t] = g.w * inv3mbar;
        g.w *= kG;
        sm[0][t] = g;
    }
    __syncthreads();
    for (int k = k0; k < k1; ++k) {
        const int b = (k - k0) & 1;
        const bool more = (k + 1 < k1);
        if (more) g = pos[(size_t)wrap_tile(L.jt0 + k + 1, L.ntiles) * kTJ + t];
        double tx[R], ty[R], tz[R];
#pragma unroll
        for (int r = 0; r < R; ++r) tx[r] = ty[r] = tz[r] = 0.0;
#pragma unroll (R >= 4 ? kFp64Unroll4 : kFp64Unroll2)
        for (int j = 0; j < kTJ; ++j) {
            const double4 pj = sm[b][j];                       // (x, y, z, G m)
            const double mqj = IS_ADAPTIVE(MODE) ? smq[IS_ADAPTIVE(MODE) ? b : 0][IS_ADAPTIVE(MODE) ? j : 0] : 0.0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const double dx = pj.x - xi[r], dy = pj.y - yi[r], dz = pj.z - zi[r];
                const double r2 = dx * dx + dy * dy + dz * dz;
                double s2 = r2;
                if (IS_ADAPTIVE(MODE)) {
                    // eps^2 = max(eta^2 r^2 q^(2/3), eps_min^2) (core/Forces.cpp:353-356)
                    const double w = pow23(mqi[r] + mqj);
                    s2 = fmax(fma(r2 * w, 1e-4, r2), r2 + 1e-8);
                }
                if (IS_PLUMMER(MODE)) s2 = r2 + 1e-10;                  // cuda/ForceKernels.cu:186
                double sc = gm_inv_r3(s2, pj.w);
                // r2 >= 0, so the integer order of the bit patterns is the floating-point order: the cut-off test
                // (core/Forces.cpp:144) runs on the integer pipe instead of the FP64 pipe.
                if (!IS_PLUMMER(MODE)) sc = (__double_as_longlong(r2) < __double_as_longlong(kR2Cut)) ? 0.0 : sc;
                tx[r] = fma(sc, dx, tx[r]); ty[r] = fma(sc, dy, ty[r]); tz[r] = fma(sc, dz, tz[r]);
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) { ax[r] += tx[r]; ay[r] += ty[r]; az[r] += tz[r]; }
        if (more) {
            if (IS_ADAPTIVE(MODE)) smq[IS_ADAPTIVE(MODE) ? b ^ 1 : 0][IS_ADAPTIVE(MODE) ? t : 0] = g.w * inv3mbar;
            g.w *= kG;
            sm[b ^ 1][t] = g;
        }
        __syncthreads();
    }
    double* o = part + (size_t)(L.s0 + s) * 3 * (size_t)L.npad;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const size_t i = (size_t)(itile_first + r) * kTJ + t;
        o[i] = ax[r]; o[(size_t)L.npad + i] = ay[r]; o[2 * (size_t)L.npad + i] = az[r];
    }
}

constexpr int kR64 = 2;   // i-particles per thread of the potential kernel

// Pair potential for computeTotalEnergy (core/System.cpp:72-82: every pair with r > 0, unsoftened), FP64, each unordered
// pair evaluated ONCE like the reference's kernel (cuda/IntegrationKernels.cu:128, `if (i < idx)`), but balanced: the
// particle in tile I sums m_j / r over its own tile with weight 1/2, over the next h = (T-1)/2 tiles of the circular tile
// order with weight 1 and — T even — over the antipodal tile I + T/2 with weight 1/2. Every unordered pair of particles then
// appears with total weight 1 (pairs in one tile and antipodal pairs from both sides with 1/2 each), every i-tile does the same
// T/2 + 1 tile passes, and a multi-GPU rank can do this for its own i-slice alone. pot[i] = sum_j w m_j / r_ij, so that
// E_pot = -G sum_i m_i pot[i]. The reciprocal root is the branch-free MUFU seed + second-order correction used by the force
// kernels (12 DP operations per pair); a tile's 256 terms are summed on their own before they join the running sum.
__global__ void __launch_bounds__(kThreads, 2)
potential_fp64_kernel(const double4* __restrict__ pos, double* __restrict__ pot, int it0, int ntiles, long npad) {
    __shared__ __align__(16) double4 sm[2][kTJ];
    const int t = threadIdx.x;
    const int itile_first = it0 + blockIdx.x * kR64;
    const int h = (ntiles - 1) / 2;
    const bool even = (ntiles % 2) == 0;
    const int span = kR64 + h + (even ? 1 : 0);            // tiles itile_first + 0 .. span-1 cover every i-tile's window
    double xi[kR64], yi[kR64], zi[kR64], acc[kR64];
#pragma unroll
    for (int r = 0; r < kR64; ++r) {
        const double4 p = pos[(size_t)(itile_first + r) * kTJ + t];
        xi[r] = p.x; yi[r] = p.y; zi[r] = p.z; acc[r] = 0.0;
    }
    // gridDim.y CTAs share one i-block's window (small systems: T/2 CTAs alone would leave most SMs empty); each writes its own
    // partial potential pot[blockIdx.y][i], summed in that order by the energy reduction
    const int kb = (int)(((long)span * blockIdx.y) / gridDim.y), ke = (int)(((long)span * (blockIdx.y + 1)) / gridDim.y);
    double4 g = make_double4(0, 0, 0, 0);
    sm[kb & 1][t] = pos[(size_t)((itile_first + kb) % ntiles) * kTJ + t];
    __syncthreads();
    for (int k = kb; k < ke; ++k) {
        const int b = k & 1;
        const bool more = (k + 1 < ke);
        if (more) g = pos[(size_t)((itile_first + k + 1) % ntiles) * kTJ + t];
        double w[kR64];
        bool any = false;
#pragma unroll
        for (int r = 0; r < kR64; ++r) {
            const int d = k - r;                             // distance of this tile from i-tile itile_first + r
            w[r] = d == 0 ? 0.5 : (d >= 1 && d <= h) ? 1.0 : (even && d == h + 1) ? 0.5 : 0.0;
            any = any || w[r] != 0.0;
        }
        if (any) {
            double ts[kR64];
#pragma unroll
            for (int r = 0; r < kR64; ++r) ts[r] = 0.0;
#pragma unroll 4
            for (int j = 0; j < kTJ; ++j) {
                const double4 pj = sm[b][j];
#pragma unroll
                for (int r = 0; r < kR64; ++r) {
                    const double dx = pj.x - xi[r], dy = pj.y - yi[r], dz = pj.z - zi[r];
                    const double r2 = dx * dx + dy * dy + dz * dz;
                    const double y0 = rsq64_seed(r2);
                    const double e = fma(r2, -(y0 * y0), 1.0);
                    const double c = e * fma(e, 0.375, 0.5);
                    double mri = pj.w * fma(y0, c, y0);
                    mri = __double_as_longlong(r2) < 0x0010000000000000LL ? 0.0 : mri;   // r > 0 only (self pair, coincident particles, pads;
                                                                                             // the MUFU seed flushes subnormal r^2 to zero)
                    ts[r] += mri;
                }
            }
#pragma unroll
            for (int r = 0; r < kR64; ++r) acc[r] = fma(w[r], ts[r], acc[r]);
        }
        if (more) sm[b ^ 1][t] = g;
        __syncthreads();
    }
#pragma unroll
    for (int r = 0; r < kR64; ++r) pot[(size_t)blockIdx.y * npad + (size_t)(itile_first + r) * kTJ + t] = acc[r];
}

template <typename K>
int regs_of(K kernel) {
    cudaFuncAttributes a;
    if (cudaFuncGetAttributes(&a, kernel) != cudaSuccess) { cudaGetLastError(); return -1; }
    return a.numRegs;
}

} // namespace

constexpr int kR32 = 2;   // i-particles per thread of variants 1 and 2

int force_i_tiles_per_cta(int variant) { return variant == 3 ? kR4 : kR32; }

template <int MODE>
int launch_fp32_mode(const float* pk, double* part, const ForceLaunch& L, int variant, cudaStream_t st) {
    if (variant == 3) {
        if (L.nit % kR4 != 0) return fail(NB_ERR_INVALID, "force: i-tile count %d not a multiple of %d", L.nit, kR4);
        // once per device and kernel instance; contexts on different host threads may get here at the same time
        static std::mutex attr_mutex;
        static bool attr_set[64] = {};
        int dev = 0; NB_CUDA(cudaGetDevice(&dev));
        {
            std::lock_guard<std::mutex> lock(attr_mutex);
            if (dev < 0 || dev >= 64 || !attr_set[dev]) {
                NB_CUDA(cudaFuncSetAttribute(force_fp32_r4_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kR4SmemBytes));
                if (dev >= 0 && dev < 64) attr_set[dev] = true;
            }
        }
        dim3 grid(L.nit / kR4, L.S), block(kThreads);
        force_fp32_r4_kernel<MODE><<<grid, block, kR4SmemBytes, st>>>(pk, part, L);
    } else {
        if (L.nit % kR32 != 0) return fail(NB_ERR_INVALID, "force: i-tile count %d not a multiple of %d", L.nit, kR32);
        dim3 grid(L.nit / kR32, L.S), block(kThreads);
        if (variant == 2) force_fp32_tma_kernel<kR32, MODE><<<grid, block, 0, st>>>(pk, part, L);
        else              force_fp32_ld_kernel<kR32, MODE><<<grid, block, 0, st>>>(pk, part, L);
    }
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

// FP64 near field of variant 3 (same grid decomposition and ForceLaunch as the far kernel, its own partial slots L.s0..)
int launch_force_near64(const float* pk, double* part, const ForceLaunch& L, int force_method, cudaStream_t st) {
    if (L.nit <= 0 || L.jcount <= 0) return NB_OK;
    if (L.nit % kR4 != 0) return fail(NB_ERR_INVALID, "force: i-tile count %d not a multiple of %d", L.nit, kR4);
    dim3 grid(L.nit * (kTJ / kNearOwner), L.S), block(kNearThreads);
    switch (force_method) {
        case kPairwise: force_near64_kernel<kPairwise><<<grid, block, 0, st>>>(pk, part, L); break;
        case kPlummer:  force_near64_kernel<kPlummer><<<grid, block, 0, st>>>(pk, part, L); break;
        case kAdaptive: force_near64_kernel<kAdaptive><<<grid, block, 0, st>>>(pk, part, L); break;
        default: return fail(NB_ERR_INVALID, "force: unknown method %d", force_method);
    }
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

// force_method: NB_FORCE_PAIRWISE (0), NB_FORCE_PAIRWISE_AVX2_FP32 (1), NB_FORCE_ADAPTIVE_MUTUAL (2) == kernel MODE
int launch_force_fp32(const float* pk, double* part, const ForceLaunch& L, int force_method, int variant, cudaStream_t st) {
    if (L.nit <= 0 || L.jcount <= 0) return NB_OK;
    switch (force_method) {
        case kPairwise: return launch_fp32_mode<kPairwise>(pk, part, L, variant, st);
        case kPlummer:  return launch_fp32_mode<kPlummer>(pk, part, L, variant, st);
        case kAdaptive: return launch_fp32_mode<kAdaptive>(pk, part, L, variant, st);
    }
    return fail(NB_ERR_INVALID, "force: unknown method %d", force_method);
}

// i-particles per thread of the FP64 kernel: 4 once the system fills the GPU (PAIRWISE: +5 %), 2 for small systems (more,
// shorter CTAs) and for ADAPTIVE_MUTUAL (4 leave too few registers for the q^(2/3) arithmetic: -5.5 %)
int force_fp64_i_tiles(long npad, int force_method) { return (npad >= 32768 && force_method != kAdaptive) ? NB_FP64_BIG_R : 2; }

template <int R>
static int launch_force_fp64_r(const double4* pos, double* part, const ForceLaunch& L, int force_method, double avg_mass, cudaStream_t st) {
    if (L.nit % R != 0) return fail(NB_ERR_INVALID, "force: i-tile count %d not a multiple of %d", L.nit, R);
    dim3 grid(L.nit / R, L.S), block(kThreads);
    const double inv3mbar = force_method == kAdaptive ? 1.0 / (3.0 * avg_mass) : 0.0;
    switch (force_method) {
        case kPairwise: force_fp64_kernel<kPairwise, R><<<grid, block, 0, st>>>(pos, part, L, inv3mbar); break;
        case kPlummer:  force_fp64_kernel<kPlummer, R><<<grid, block, 0, st>>>(pos, part, L, inv3mbar); break;
        case kAdaptive: force_fp64_kernel<kAdaptive, R><<<grid, block, 0, st>>>(pos, part, L, inv3mbar); break;
        default: return fail(NB_ERR_INVALID, "force: unknown method %d", force_method);
    }
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

int launch_force_fp64(const double4* pos, double* part, const ForceLaunch& L, int force_method, double avg_mass, cudaStream_t st) {
    if (L.nit <= 0 || L.jcount <= 0) return NB_OK;
    return force_fp64_i_tiles(L.npad, force_method) == 4 ? launch_force_fp64_r<4>(pos, part, L, force_method, avg_mass, st)
                                           : launch_force_fp64_r<2>(pos, part, L, force_method, avg_mass, st);
}

// Number of partial potentials per particle launch_potential_fp64 writes (pot must hold that many arrays of npad doubles, <= 4)
int potential_splits(int nit, int sm_count) {
    const int blocks = nit / kR64;
    int S = (4 * sm_count + blocks - 1) / (blocks > 0 ? blocks : 1);
    return S < 1 ? 1 : (S > 4 ? 4 : S);
}

int launch_potential_fp64(const double4* pos, double* pot, int it0, int nit, int ntiles, long npad, int splits, cudaStream_t st) {
    if (nit <= 0) return NB_OK;
    if (nit % kR64) return fail(NB_ERR_INVALID, "potential: i-tile count %d not a multiple of %d", nit, kR64);
    dim3 grid((unsigned)(nit / kR64), (unsigned)(splits < 1 ? 1 : splits)), block(kThreads);
    potential_fp64_kernel<<<grid, block, 0, st>>>(pos, pot, it0, ntiles, npad);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

int force_regs_info(char* buf, int len) {
    return snprintf(buf, len, "fp32_r4:%d fp32_r4_plummer:%d fp32_r4_adaptive:%d fp32_tma:%d fp32_ld:%d fp64:%d fp64_adaptive:%d",
                    regs_of(force_fp32_r4_kernel<kPairwise>), regs_of(force_fp32_r4_kernel<kPlummer>), regs_of(force_fp32_r4_kernel<kAdaptive>),
                    regs_of(force_fp32_tma_kernel<kR32, kPairwise>), regs_of(force_fp32_ld_kernel<kR32, kPairwise>),
                    regs_of(force_fp64_kernel<kPairwise, 4>), regs_of(force_fp64_kernel<kAdaptive, 4>));
}

} // namespace nb
