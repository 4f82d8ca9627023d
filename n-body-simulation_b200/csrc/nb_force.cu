// nb_force.cu — all-pairs force kernels for sm_100a.
//
// Replaces calculatePairwiseForceKernel / calculateAdaptiveForcesKernel (reference cuda/ForceKernels.cu:14-96,99-170)
// with the semantics of the CPU oracle (core/Forces.cpp:98-179,301-387): a_i = sum_{j != i, r^2 >= 1e-10} G m_j d / s^3.
//
// FP32 mode (force_fp32_kernel)
//   * packed FP32x2 arithmetic (FADD2/FMUL2/FFMA2, new on sm_100): one instruction works on TWO j-particles, so the
//     9+12 FP32 operations of an interaction cost ~10 issue slots instead of 21 and the MUFU/LDS/select instructions
//     issue in the slots that frees;
//   * two-float (hi+lo) positions: d = (xjh - xih) + (xjl - xil) keeps the relative error of d at ~1e-7 for any
//     separation (plain FP32 coordinates lose close pairs: error ~ eps*|x|/|d|);
//   * R i-particles per thread held in registers (duplicated into both halves of a register pair), j-tiles of 256
//     particles staged in shared memory either with plain loads (variant 1) or with cp.async.bulk + mbarrier
//     (variant 2, UBLKCP in SASS) into a 4-deep ring;
//   * FP32 accumulation over 128 j-particles (64 per half), then flushed into FP64 accumulators;
//   * j-range split over gridDim.y so that small N still fills 148 SMs; partial sums go to part[s][3][npad] and are
//     reduced in fixed order by the integrator kernel (deterministic, no atomics).
// FP64 mode (force_fp64_kernel): same tiling with double4 tiles and DFMA arithmetic.
#include "nb_common.cuh"

namespace nb {

namespace {

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

// ------------------------------------------------------------------------------------------------------------------
// Register-resident i-particles: R per thread, every value duplicated into both halves of a register pair so that one
// packed instruction handles two j-particles.
template <int R, bool ADAPTIVE>
struct IState {
    float2 nxh[R], nyh[R], nzh[R];     // -(hi position)
    float2 nxl[R], nyl[R], nzl[R];     // -(lo position)
    float2 mq[R];                      // m_i / (3 mean mass) (ADAPTIVE)
    float2 ax[R], ay[R], az[R];        // FP32 partial sums (.x: even j, .y: odd j)
    double dax[R], day[R], daz[R];     // FP64 running sums
};

// One j4 group (four consecutive j-particles = two packed pairs) against the R i-particles.
//   hi, lo : 3 rows (x,y,z) of kTJ floats each;  gm, mq : one row each.
// FAR (only used in the i-block-relative frame of variant 2): the hi parts alone carry d to ~3e-7 and no pair of the
// tile can be a self pair or inside the cut-off, so the lo adds and the compare/select are dropped:
// 12 instead of 18 FP32-pipe operations per interaction.
template <int R, bool ADAPTIVE, bool FAR>
__device__ __forceinline__ void interact4(IState<R, ADAPTIVE>& S, const float* __restrict__ hi, const float* __restrict__ lo,
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
    if (ADAPTIVE) MQ = reinterpret_cast<const float4*>(mqrow)[j4];
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
            if (ADAPTIVE) {
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
            float2 ri; ri.x = rsq_approx(s2.x); ri.y = rsq_approx(s2.y);
            // cut-off (also removes the self term, for which r2 == 0 exactly): core/Forces.cpp:135,144
            if (!FAR) {
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

template <int R, bool ADAPTIVE, bool FAR>
__device__ __forceinline__ void tile_pass(IState<R, ADAPTIVE>& S, const float* __restrict__ hi, const float* __restrict__ lo,
                                          const float* __restrict__ gmrow, const float* __restrict__ mqrow) {
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
#pragma unroll 2
        for (int j4 = half * (kTJ / 8); j4 < (half + 1) * (kTJ / 8); ++j4)
            interact4<R, ADAPTIVE, FAR>(S, hi, lo, gmrow, mqrow, j4);
#pragma unroll
        for (int r = 0; r < R; ++r) {      // flush 128 j's of FP32 partial sums into the FP64 accumulators
            S.dax[r] += (double)(S.ax[r].x + S.ax[r].y); S.ax[r] = make_float2(0.f, 0.f);
            S.day[r] += (double)(S.ay[r].x + S.ay[r].y); S.ay[r] = make_float2(0.f, 0.f);
            S.daz[r] += (double)(S.az[r].x + S.az[r].y); S.az[r] = make_float2(0.f, 0.f);
        }
    }
}

template <int R, bool ADAPTIVE>
__device__ __forceinline__ void init_acc(IState<R, ADAPTIVE>& S) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
        S.ax[r] = S.ay[r] = S.az[r] = make_float2(0.f, 0.f);
        S.dax[r] = S.day[r] = S.daz[r] = 0.0;
    }
}

template <int R, bool ADAPTIVE>
__device__ __forceinline__ void store_part(const IState<R, ADAPTIVE>& S, double* __restrict__ part, const ForceLaunch& L,
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
template <int R, bool ADAPTIVE>
__global__ void __launch_bounds__(kThreads, 2)
force_fp32_ld_kernel(const float* __restrict__ pk, double* __restrict__ part, const ForceLaunch L) {
    __shared__ __align__(128) float sm[2][kTileFloats];
    const int t = threadIdx.x, s = blockIdx.y;
    const int k0 = (int)(((long)L.jcount * s) / L.S), k1 = (int)(((long)L.jcount * (s + 1)) / L.S);
    const int itile_first = L.it0 + blockIdx.x * R;
    IState<R, ADAPTIVE> S;
    init_acc<R, ADAPTIVE>(S);
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
        v = ADAPTIVE ? p[7 * kTJ] : 0.f; S.mq[r] = make_float2(v, v);
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
        tile_pass<R, ADAPTIVE, false>(S, sm[b], sm[b] + 4 * kTJ, sm[b] + 3 * kTJ, sm[b] + 7 * kTJ);
        if (more) {
            reinterpret_cast<float4*>(sm[b ^ 1])[t] = g0;
            reinterpret_cast<float4*>(sm[b ^ 1])[t + kThreads] = g1;
        }
        __syncthreads();
    }
    store_part<R, ADAPTIVE>(S, part, L, itile_first, t, s);
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
__device__ __forceinline__ bool tile_is_far(const CtaFrame& F, const float* __restrict__ meta, float kappa) {
    float g2 = 0.f, hj = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float g = fmaxf(0.f, fmaxf(meta[c] - F.hi[c], F.lo[c] - meta[3 + c]));
        g2 = fmaf(g, g, g2);
        hj = fmaxf(hj, 0.5f * (meta[3 + c] - meta[c]));
    }
    const float thr = kappa * (F.half + hj);
    return (g2 > thr * thr) && (g2 > 1e-8f);
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

template <int R, bool ADAPTIVE>
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
    IState<R, ADAPTIVE> S;
    init_acc<R, ADAPTIVE>(S);
#pragma unroll
    for (int r = 0; r < R; ++r) {      // own i-particles in the CTA frame, as two-float values
        const float* p = pk + (size_t)(itile_first + r) * kTileFloats + t;
        float h, l;
        rel_two(p[0 * kTJ], p[4 * kTJ], F.O[0], h, l); S.nxh[r] = make_float2(-h, -h); S.nxl[r] = make_float2(-l, -l);
        rel_two(p[1 * kTJ], p[5 * kTJ], F.O[1], h, l); S.nyh[r] = make_float2(-h, -h); S.nyl[r] = make_float2(-l, -l);
        rel_two(p[2 * kTJ], p[6 * kTJ], F.O[2], h, l); S.nzh[r] = make_float2(-h, -h); S.nzl[r] = make_float2(-l, -l);
        const float v = ADAPTIVE ? p[7 * kTJ] : 0.f; S.mq[r] = make_float2(v, v);
    }
    bool far_cur = false;
    if (cnt > 0) {
        mbar_wait(&full[0], 0u);
        far_cur = tile_is_far(F, raw[0] + kMetaRow * kTJ, L.kappa);
        if (far_cur) transform_tile<true>(raw[0], T[0], F, t); else transform_tile<false>(raw[0], T[0], F, t);
    }
    __syncthreads();
    for (int q = 0; q < cnt; ++q) {
        const int b = q % kStages;
        bool far_next = false;
        if (q + 1 < cnt) {                 // re-express the next tile while this one is being computed
            const int bn = (q + 1) % kStages;
            mbar_wait(&full[bn], (uint32_t)(((q + 1) / kStages) & 1));
            far_next = tile_is_far(F, raw[bn] + kMetaRow * kTJ, L.kappa);
            float* Tn = T[(q + 1) & 1];
            if (far_next) transform_tile<true>(raw[bn], Tn, F, t); else transform_tile<false>(raw[bn], Tn, F, t);
        }
        const float* Tc = T[q & 1];
        if (far_cur) tile_pass<R, ADAPTIVE, true>(S, Tc, Tc + 3 * kTJ, raw[b] + 3 * kTJ, raw[b] + 7 * kTJ);
        else         tile_pass<R, ADAPTIVE, false>(S, Tc, Tc + 3 * kTJ, raw[b] + 3 * kTJ, raw[b] + 7 * kTJ);
        __syncthreads();                   // every warp is done with raw[b] and T[q&1]; T[(q+1)&1] is complete
        if (t == 0 && q + kStages < cnt) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(&full[b], kTileBytes);
            bulk_g2s(raw[b], pk + (size_t)wrap_tile(L.jt0 + k0 + q + kStages, L.ntiles) * kTileFloats, kTileBytes, &full[b]);
        }
        far_cur = far_next;
    }
    store_part<R, ADAPTIVE>(S, part, L, itile_first, t, s);
}

// ------------------------------------------------------------------------------------------------------------------
// FP64 mode. Tiles of 256 double4 (x,y,z,m) straight from the master array; R = 2 i-particles per thread.
constexpr int kR64 = 2;

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

template <bool ADAPTIVE>
__global__ void __launch_bounds__(kThreads, 2)
force_fp64_kernel(const double4* __restrict__ pos, double* __restrict__ part, const ForceLaunch L, double inv3mbar) {
    __shared__ __align__(16) double4 sm[2][kTJ];
    const int t = threadIdx.x, s = blockIdx.y;
    const int k0 = (int)(((long)L.jcount * s) / L.S), k1 = (int)(((long)L.jcount * (s + 1)) / L.S);
    const int itile_first = L.it0 + blockIdx.x * kR64;
    double xi[kR64], yi[kR64], zi[kR64], mqi[kR64], ax[kR64], ay[kR64], az[kR64];
#pragma unroll
    for (int r = 0; r < kR64; ++r) {
        const double4 p = pos[(size_t)(itile_first + r) * kTJ + t];
        xi[r] = p.x; yi[r] = p.y; zi[r] = p.z; mqi[r] = p.w * inv3mbar; ax[r] = ay[r] = az[r] = 0.0;
    }
    double4 g = make_double4(0, 0, 0, 0);
    if (k0 < k1) sm[0][t] = pos[(size_t)wrap_tile(L.jt0 + k0, L.ntiles) * kTJ + t];
    __syncthreads();
    for (int k = k0; k < k1; ++k) {
        const int b = (k - k0) & 1;
        const bool more = (k + 1 < k1);
        if (more) g = pos[(size_t)wrap_tile(L.jt0 + k + 1, L.ntiles) * kTJ + t];
#pragma unroll 4
        for (int j = 0; j < kTJ; ++j) {
            const double4 pj = sm[b][j];
            const double gm = kG * pj.w;
            const double mqj = ADAPTIVE ? pj.w * inv3mbar : 0.0;
#pragma unroll
            for (int r = 0; r < kR64; ++r) {
                const double dx = pj.x - xi[r], dy = pj.y - yi[r], dz = pj.z - zi[r];
                const double r2 = dx * dx + dy * dy + dz * dz;
                double s2 = r2;
                if (ADAPTIVE) {
                    // eps^2 = max(eta^2 r^2 q^(2/3), eps_min^2) (core/Forces.cpp:353-356)
                    const double w = pow23(mqi[r] + mqj);
                    s2 = fmax(fma(r2 * w, 1e-4, r2), r2 + 1e-8);
                }
                double sc = gm_inv_r3(s2, gm);
                // r2 >= 0, so the integer order of the bit patterns is the floating-point order: the cut-off test
                // (core/Forces.cpp:144) runs on the integer pipe instead of the FP64 pipe.
                sc = (__double_as_longlong(r2) < __double_as_longlong(kR2Cut)) ? 0.0 : sc;
                ax[r] = fma(sc, dx, ax[r]); ay[r] = fma(sc, dy, ay[r]); az[r] = fma(sc, dz, az[r]);
            }
        }
        if (more) sm[b ^ 1][t] = g;
        __syncthreads();
    }
    double* o = part + (size_t)(L.s0 + s) * 3 * (size_t)L.npad;
#pragma unroll
    for (int r = 0; r < kR64; ++r) {
        const size_t i = (size_t)(itile_first + r) * kTJ + t;
        o[i] = ax[r]; o[(size_t)L.npad + i] = ay[r]; o[2 * (size_t)L.npad + i] = az[r];
    }
}

// Per-particle potential sum_j m_j / r_ij over every pair with r > 0 (core/System.cpp:72-82), FP64.
__global__ void __launch_bounds__(kThreads, 2)
potential_fp64_kernel(const double4* __restrict__ pos, double* __restrict__ pot, int ntiles) {
    __shared__ __align__(16) double4 sm[2][kTJ];
    const int t = threadIdx.x;
    const int itile_first = blockIdx.x * kR64;
    double xi[kR64], yi[kR64], zi[kR64], acc[kR64];
#pragma unroll
    for (int r = 0; r < kR64; ++r) {
        const double4 p = pos[(size_t)(itile_first + r) * kTJ + t];
        xi[r] = p.x; yi[r] = p.y; zi[r] = p.z; acc[r] = 0.0;
    }
    double4 g = make_double4(0, 0, 0, 0);
    sm[0][t] = pos[t];
    __syncthreads();
    for (int k = 0; k < ntiles; ++k) {
        const int b = k & 1;
        const bool more = (k + 1 < ntiles);
        if (more) g = pos[(size_t)(k + 1) * kTJ + t];
#pragma unroll 4
        for (int j = 0; j < kTJ; ++j) {
            const double4 pj = sm[b][j];
#pragma unroll
            for (int r = 0; r < kR64; ++r) {
                const double dx = pj.x - xi[r], dy = pj.y - yi[r], dz = pj.z - zi[r];
                const double r2 = dx * dx + dy * dy + dz * dz;
                double ri = rsqrt(r2);
                ri = r2 > 0.0 ? ri : 0.0;
                acc[r] += pj.w * ri;
            }
        }
        if (more) sm[b ^ 1][t] = g;
        __syncthreads();
    }
#pragma unroll
    for (int r = 0; r < kR64; ++r) pot[(size_t)(itile_first + r) * kTJ + t] = acc[r];
}

template <typename K>
int regs_of(K kernel) {
    cudaFuncAttributes a;
    if (cudaFuncGetAttributes(&a, kernel) != cudaSuccess) { cudaGetLastError(); return -1; }
    return a.numRegs;
}

} // namespace

constexpr int kR32 = 2;   // i-particles per thread in FP32 mode

int launch_force_fp32(const float* pk, double* part, const ForceLaunch& L, bool adaptive, int variant, cudaStream_t st) {
    if (L.nit <= 0 || L.jcount <= 0) return NB_OK;
    if (L.nit % kR32 != 0) return fail(NB_ERR_INVALID, "force: i-tile count %d not a multiple of %d", L.nit, kR32);
    dim3 grid(L.nit / kR32, L.S), block(kThreads);
    if (variant == 2) {
        if (adaptive) force_fp32_tma_kernel<kR32, true><<<grid, block, 0, st>>>(pk, part, L);
        else          force_fp32_tma_kernel<kR32, false><<<grid, block, 0, st>>>(pk, part, L);
    } else {
        if (adaptive) force_fp32_ld_kernel<kR32, true><<<grid, block, 0, st>>>(pk, part, L);
        else          force_fp32_ld_kernel<kR32, false><<<grid, block, 0, st>>>(pk, part, L);
    }
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

int launch_force_fp64(const double4* pos, double* part, const ForceLaunch& L, bool adaptive, double avg_mass, cudaStream_t st) {
    if (L.nit <= 0 || L.jcount <= 0) return NB_OK;
    if (L.nit % kR64 != 0) return fail(NB_ERR_INVALID, "force: i-tile count %d not a multiple of %d", L.nit, kR64);
    dim3 grid(L.nit / kR64, L.S), block(kThreads);
    const double inv3mbar = adaptive ? 1.0 / (3.0 * avg_mass) : 0.0;
    if (adaptive) force_fp64_kernel<true><<<grid, block, 0, st>>>(pos, part, L, inv3mbar);
    else          force_fp64_kernel<false><<<grid, block, 0, st>>>(pos, part, L, inv3mbar);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

int launch_potential_fp64(const double4* pos, double* pot, long npad, int ntiles, cudaStream_t st) {
    dim3 grid((unsigned)(npad / (kTJ * kR64))), block(kThreads);
    potential_fp64_kernel<<<grid, block, 0, st>>>(pos, pot, ntiles);
    NB_CUDA(cudaGetLastError());
    return NB_OK;
}

int force_regs_info(char* buf, int len) {
    return snprintf(buf, len, "fp32_ld:%d fp32_ld_adaptive:%d fp32_tma:%d fp32_tma_adaptive:%d fp64:%d fp64_adaptive:%d",
                    regs_of(force_fp32_ld_kernel<kR32, false>), regs_of(force_fp32_ld_kernel<kR32, true>),
                    regs_of(force_fp32_tma_kernel<kR32, false>), regs_of(force_fp32_tma_kernel<kR32, true>),
                    regs_of(force_fp64_kernel<false>), regs_of(force_fp64_kernel<true>));
}

} // namespace nb
