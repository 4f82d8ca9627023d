// tools/pipe_bench.cu — measures the sm_100a issue/pipe rates the force kernel's roofline is built from:
// FFMA, packed FFMA2/FADD2 (FP32x2), MUFU.RSQ, FSETP/FSEL co-issue, DFMA, broadcast LDS.128, and the SM clock under
// an FP32 load. Prints one JSON object. Not part of the product library.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <string>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "CUDA %s at %d: %s\n", #x, __LINE__, cudaGetErrorString(e)); exit(2); } } while (0)

constexpr int ITERS = 2048;

__device__ __forceinline__ float rsq(float x) { float y; asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__global__ void __launch_bounds__(256) k_ffma(float* out, float a, float b) {
    float c[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) c[i] = threadIdx.x * 1e-3f + i;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int i = 0; i < 16; ++i) c[i] = fmaf(c[i], a, b);
    }
    float s = 0; for (int i = 0; i < 16; ++i) s += c[i];
    if (s == 12345.f) out[0] = s;
}
__global__ void __launch_bounds__(256) k_ffma2(float* out, float a, float b) {
    float2 c[16]; const float2 A = make_float2(a, a), B = make_float2(b, b);
#pragma unroll
    for (int i = 0; i < 16; ++i) c[i] = make_float2(threadIdx.x * 1e-3f + i, i);
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int i = 0; i < 16; ++i) c[i] = __ffma2_rn(c[i], A, B);
    }
    float s = 0; for (int i = 0; i < 16; ++i) s += c[i].x + c[i].y;
    if (s == 12345.f) out[0] = s;
}
__global__ void __launch_bounds__(256) k_fadd2(float* out, float a, float b) {
    float2 c[16]; const float2 A = make_float2(a, b);
#pragma unroll
    for (int i = 0; i < 16; ++i) c[i] = make_float2(threadIdx.x * 1e-3f + i, i);
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int i = 0; i < 16; ++i) c[i] = __fadd2_rn(c[i], A);
    }
    float s = 0; for (int i = 0; i < 16; ++i) s += c[i].x + c[i].y;
    if (s == 12345.f) out[0] = s;
}
// 8 FFMA2 + NSEL x (FSETP+FSEL) per unit: do the compare/select instructions issue in otherwise idle slots?
template <int NSEL>
__global__ void __launch_bounds__(256) k_ffma2_sel(float* out, float a, float b) {
    float2 c[8]; float d[8]; const float2 A = make_float2(a, a), B = make_float2(b, b);
#pragma unroll
    for (int i = 0; i < 8; ++i) { c[i] = make_float2(threadIdx.x * 1e-3f + i, i); d[i] = i + a; }
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 8; ++i) c[i] = __ffma2_rn(c[i], A, B);
#pragma unroll
            for (int i = 0; i < NSEL; ++i) d[i] = d[i] < c[i].x ? b : d[i];
        }
    }
    float s = 0; for (int i = 0; i < 8; ++i) s += c[i].x + c[i].y + d[i];
    if (s == 12345.f) out[0] = s;
}
// 6 FFMA2 + NMUFU x MUFU.RSQ per unit
template <int NMUFU>
__global__ void __launch_bounds__(256) k_ffma2_mufu(float* out, float a, float b) {
    float2 c[6]; float d[4]; const float2 A = make_float2(a, a), B = make_float2(b, b);
#pragma unroll
    for (int i = 0; i < 6; ++i) c[i] = make_float2(threadIdx.x * 1e-3f + i, i);
#pragma unroll
    for (int i = 0; i < 4; ++i) d[i] = 1.f + i + a;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 6; ++i) c[i] = __ffma2_rn(c[i], A, B);
#pragma unroll
            for (int i = 0; i < NMUFU; ++i) d[i] = rsq(d[i]);
        }
    }
    float s = 0; for (int i = 0; i < 6; ++i) s += c[i].x + c[i].y; for (int i = 0; i < 4; ++i) s += d[i];
    if (s == 12345.f) out[0] = s;
}
__global__ void __launch_bounds__(256) k_mufu(float* out, float a) {
    float d[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) d[i] = 1.f + i + a + threadIdx.x;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) d[i] = rsq(d[i]);
    }
    float s = 0; for (int i = 0; i < 8; ++i) s += d[i];
    if (s == 12345.f) out[0] = s;
}
__global__ void __launch_bounds__(256) k_dfma(double* out, double a, double b) {
    double c[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) c[i] = threadIdx.x * 1e-3 + i;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) c[i] = fma(c[i], a, b);
    }
    double s = 0; for (int i = 0; i < 8; ++i) s += c[i];
    if (s == 12345.0) out[0] = s;
}
// broadcast LDS.128 (all lanes read the same 16 bytes) + NF FFMA2 per load
template <int NF>
__global__ void __launch_bounds__(256) k_lds(float* out, float a, float b, int stride) {
    __shared__ __align__(16) float4 sm[512];
    for (int i = threadIdx.x; i < 512; i += 256) sm[i] = make_float4(i, a, b, 1.f);
    __syncthreads();
    float2 c[8]; const float2 B = make_float2(b, b);
#pragma unroll
    for (int i = 0; i < 8; ++i) c[i] = make_float2(threadIdx.x * 1e-3f + i, i);
    int idx = 0;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const float4 v = sm[(idx + u * stride) & 511];
#pragma unroll
            for (int i = 0; i < NF; ++i) c[i] = __ffma2_rn(c[i], make_float2(v.x, v.y), B);
            if (NF == 0) { c[u & 7].x += v.z; }
        }
        idx += 8 * stride;
    }
    float s = 0; for (int i = 0; i < 8; ++i) s += c[i].x + c[i].y;
    if (s == 12345.f) out[0] = s;
}
// 8 FFMA2 + ND DFMA per unit, independent chains: do the FP32x2 and the FP64 pipes run side by side in one instruction stream?
// (the far kernel leaves the FP64 pipe idle; this is the feasibility number for giving it part of the interactions)
template <int ND>
__global__ void __launch_bounds__(256) k_ffma2_dfma(float* out, float a, float b, double da, double db) {
    float2 c[8]; double d[8]; const float2 A = make_float2(a, a), B = make_float2(b, b);
#pragma unroll
    for (int i = 0; i < 8; ++i) { c[i] = make_float2(threadIdx.x * 1e-3f + i, i); d[i] = threadIdx.x * 1e-3 + i; }
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                c[i] = __ffma2_rn(c[i], A, B);
                if (i < ND) d[i] = fma(d[i], da, db);
            }
        }
    }
    float s = 0; for (int i = 0; i < 8; ++i) s += c[i].x + c[i].y + (float)d[i];
    if (s == 12345.f) out[0] = s;
}
__global__ void k_clock(unsigned long long* out, float a, float b) {
    unsigned long long t0, t1, c0, c1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    c0 = clock64();
    float c[8];
    for (int i = 0; i < 8; ++i) c[i] = threadIdx.x + i;
    for (int it = 0; it < 200000; ++it)
#pragma unroll
        for (int i = 0; i < 8; ++i) c[i] = fmaf(c[i], a, b);
    c1 = clock64();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    float s = 0; for (int i = 0; i < 8; ++i) s += c[i];
    if (blockIdx.x == 0 && threadIdx.x == 0) { out[0] = c1 - c0; out[1] = t1 - t0; out[2] = (unsigned long long)s; }
}

struct Result { std::string name; double warp_instr_per_clk_per_sm; double ms; };

int main() {
    int dev = 0; CK(cudaSetDevice(dev));
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, dev));
    const int sms = p.multiProcessorCount;
    float* out; CK(cudaMalloc(&out, 1024));
    unsigned long long* cl; CK(cudaMalloc(&cl, 64));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int grid = sms * 8;

    // SM clock under an FP32 load: run the clock kernel on all SMs
    k_clock<<<sms * 4, 256>>>(cl, 1.0001f, 0.5f); CK(cudaDeviceSynchronize());
    k_clock<<<sms * 4, 256>>>(cl, 1.0001f, 0.5f); CK(cudaDeviceSynchronize());
    unsigned long long h[3]; CK(cudaMemcpy(h, cl, 24, cudaMemcpyDeviceToHost));
    const double sm_ghz = (double)h[0] / (double)h[1];

    std::vector<Result> res;
    auto run = [&](const char* name, auto launch, double instr_per_thread) {
        launch(); CK(cudaDeviceSynchronize());
        launch(); CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0)); launch(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        const double warp_instr = instr_per_thread * (double)grid * 256.0 / 32.0;
        const double clks = ms * 1e-3 * sm_ghz * 1e9;
        res.push_back({name, warp_instr / clks / sms, ms});
    };
    run("ffma",            [&] { k_ffma<<<grid, 256>>>(out, 1.0001f, 0.5f); },               (double)ITERS * 64);
    run("ffma2",           [&] { k_ffma2<<<grid, 256>>>(out, 1.0001f, 0.5f); },              (double)ITERS * 64);
    run("fadd2",           [&] { k_fadd2<<<grid, 256>>>(out, 1.0001f, 0.5f); },              (double)ITERS * 64);
    run("mufu_rsq",        [&] { k_mufu<<<grid, 256>>>(out, 0.5f); },                        (double)ITERS * 64);
    run("dfma",            [&] { k_dfma<<<grid, 256>>>((double*)out, 1.0001, 0.5); },        (double)ITERS * 32);
    run("ffma2x8_sel0",    [&] { k_ffma2_sel<0><<<grid, 256>>>(out, 1.0001f, 0.5f); },       (double)ITERS * 64);
    run("ffma2x8_sel2",    [&] { k_ffma2_sel<2><<<grid, 256>>>(out, 1.0001f, 0.5f); },       (double)ITERS * 64);
    run("ffma2x8_sel4",    [&] { k_ffma2_sel<4><<<grid, 256>>>(out, 1.0001f, 0.5f); },       (double)ITERS * 64);
    run("ffma2x8_sel8",    [&] { k_ffma2_sel<8><<<grid, 256>>>(out, 1.0001f, 0.5f); },       (double)ITERS * 64);
    run("ffma2x6_mufu0",   [&] { k_ffma2_mufu<0><<<grid, 256>>>(out, 1.0001f, 0.5f); },      (double)ITERS * 48);
    run("ffma2x6_mufu1",   [&] { k_ffma2_mufu<1><<<grid, 256>>>(out, 1.0001f, 0.5f); },      (double)ITERS * 48);
    run("ffma2x6_mufu2",   [&] { k_ffma2_mufu<2><<<grid, 256>>>(out, 1.0001f, 0.5f); },      (double)ITERS * 48);
    run("ffma2x8_dfma0",   [&] { k_ffma2_dfma<0><<<grid, 256>>>(out, 1.0001f, 0.5f, 1.0001, 0.5); }, (double)ITERS * 64);
    run("ffma2x8_dfma2",   [&] { k_ffma2_dfma<2><<<grid, 256>>>(out, 1.0001f, 0.5f, 1.0001, 0.5); }, (double)ITERS * 64);
    run("ffma2x8_dfma4",   [&] { k_ffma2_dfma<4><<<grid, 256>>>(out, 1.0001f, 0.5f, 1.0001, 0.5); }, (double)ITERS * 64);
    run("ffma2x8_dfma6",   [&] { k_ffma2_dfma<6><<<grid, 256>>>(out, 1.0001f, 0.5f, 1.0001, 0.5); }, (double)ITERS * 64);
    run("ffma2x8_dfma8",   [&] { k_ffma2_dfma<8><<<grid, 256>>>(out, 1.0001f, 0.5f, 1.0001, 0.5); }, (double)ITERS * 64);
    run("lds128_bcast",    [&] { k_lds<0><<<grid, 256>>>(out, 1.0001f, 0.5f, 1); },          (double)ITERS * 8);
    run("lds128_ffma2x4",  [&] { k_lds<4><<<grid, 256>>>(out, 1.0001f, 0.5f, 1); },          (double)ITERS * 32);
    run("lds128_ffma2x8",  [&] { k_lds<8><<<grid, 256>>>(out, 1.0001f, 0.5f, 1); },          (double)ITERS * 64);

    printf("{\"gpu\": \"%s\", \"sms\": %d, \"sm_ghz_under_fp32_load\": %.4f, \"rates_warp_instr_per_clk_per_sm\": {", p.name, sms, sm_ghz);
    for (size_t i = 0; i < res.size(); ++i)
        printf("%s\"%s\": %.3f", i ? ", " : "", res[i].name.c_str(), res[i].warp_instr_per_clk_per_sm);
    printf("}, \"note\": \"rate counts the named (primary) instruction only: ffma2x8_selK = FFMA2 rate with K FSETP+FSEL pairs per 8 FFMA2; ffma2x6_mufuK = FFMA2 rate with K MUFU per 6 FFMA2; lds128_ffma2xK = FFMA2 rate with one broadcast LDS.128 per K FFMA2; ffma2x8_dfmaK = FFMA2 rate with K DFMA per 8 FFMA2\"}\n");
    return 0;
}
