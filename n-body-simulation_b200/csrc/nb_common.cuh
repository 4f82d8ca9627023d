// nb_common.cuh — shared declarations of libnbody_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>

#include "nbody_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ != 1000)
#error "libnbody_b200 kernels are written for sm_100a (B200) only"
#endif

namespace nb {

// Units and constants of the reference: cuda/Constants.cuh:6-8, core/Forces.cpp:14,305-306.
constexpr double kG        = 39.47841760435743;
constexpr double kEta      = 0.01;
constexpr double kEpsMin   = 1e-4;
constexpr double kR2Cut    = 1e-10;   // core/Forces.cpp:144,350 — pairs with r^2 < 1e-10 are skipped

// ---- packed j-tile layout (FP32 mode) ----------------------------------------------------------------------------
// Particles are grouped in tiles of kTJ. One tile is kNC contiguous rows of kTJ floats (8 KB), so a tile is ONE bulk
// copy into shared memory and a float4 shared load yields the same component of four consecutive j-particles:
//   row 0..2  xh yh zh   high parts of the position  (float)x
//   row 3     gm         (float)(G * m)
//   row 4..6  xl yl zl   low parts                   (float)(x - (double)xh)
//   row 7     mq         (float)(m / (3 * mean mass))   (ADAPTIVE_MUTUAL only)
//   row 8     meta       [0..2] bbox min, [3..5] bbox max of the tile's hi positions, [6] max |coordinate|,
//                        [8..31] min/max boxes of the four 64-particle sub-tiles; rest unused.
//                        Travels with the tile (bulk copy, all-gather) so the kernel can pick the cheap path per tile.
constexpr int kTJ = 256;                 // j-particles per tile == threads per CTA
//   row 9     gml        (float)(G m - (double)gm): low part of G m for the FP64 near-field path
constexpr int kNC = 10;                  // rows per tile
constexpr int kMetaRow = 8;
constexpr int kGmLoRow = 9;
constexpr int kSubTile = 64;              // j-particles per sub-tile (classification granularity of near tiles)
constexpr int kSubBoxOffset = 8;         // meta row: sub-tile boxes start here (4 x 6 floats)
constexpr int kTileFloats = kTJ * kNC;   // 2560 floats = 10240 bytes
constexpr int kTileBytes  = kTileFloats * 4;
constexpr int kAlign = 1024;             // particle counts are padded to a multiple of this (4 tiles)
constexpr int kThreads = 256;

struct ForceLaunch {          // one launch = i-tiles [it0, it0+nit) against the circular j-tile range
    int it0, nit;             // i-tiles (units of kTJ particles); nit must be a multiple of R
    int jt0, jcount;          // j-tiles jt0, jt0+1, ... (mod ntiles), jcount of them
    int ntiles;               // tiles in the whole padded system
    long npad;                // padded particle count (row stride of `part`)
    int s0, S;                // this launch writes partial slots s0 .. s0+S-1
    float g0;                 // mass scale G * (mean mass) of the mass-scaled far path (> 0)
    int stats_warps;           // != 0: stats has 8 + (warps x splits) words; per-warp near counts are stored from word 8
    unsigned long long* stats; // optional device counters [4]: far tiles, CTA-near tiles (per CTA), far / FP64 sub-tiles (per warp); or null
    float kappa;              // far-tile criterion: box gap > kappa * (|x|max_i + |x|max_j); huge = always accurate path
};

// Set thread-local error text; returns code.
int fail(int code, const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define NB_CUDA(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return nb::cuda_fail(_e, #x, __FILE__, __LINE__); } while (0)

// nb_force.cu
int launch_force_fp32(const float* pk, double* part, const ForceLaunch& L, int force_method, int variant, cudaStream_t st);
int launch_force_near64(const float* pk, double* part, const ForceLaunch& L, int force_method, cudaStream_t st);
int launch_force_fp64(const double4* pos, double* part, const ForceLaunch& L, int force_method, double avg_mass, cudaStream_t st);
int potential_splits(int nit, int sm_count);
int launch_potential_fp64(const double4* pos, double* pot, int it0, int nit, int ntiles, long npad, int splits, cudaStream_t st);
int force_regs_info(char* buf, int len);
int force_i_tiles_per_cta(int variant);
int force_fp64_i_tiles(long npad, int force_method);

// nb_integrate.cu
struct StepArrays {
    double4* pos; double4* vel;     // master state, AoS like the reference (core/Particles.hpp:9-16)
    double* acc;                    // current acceleration a = F(x), SoA [3][npad]
    float* pk;                      // packed FP32 tiles (may be null in FP64 mode)
    double* part;                   // force partials [S][3][npad]
    double* red_part;               // per-block partial sums [nblocks][4] (sum m vx, m vy, m vz, m)
    double* red;                    // final [4]
    long npad; long n; long i0; long ni;   // owned slice [i0, i0+ni) (tile aligned), n real particles
};
int launch_pack(const StepArrays& A, long first, long count, double avg_mass, cudaStream_t st);
int launch_drift(const StepArrays& A, int integ, double dt, bool subtract_com, double avg_mass, cudaStream_t st);
int launch_finish(const StepArrays& A, int integ, bool kick, double dt, int S, cudaStream_t st);
int launch_com_final(const StepArrays& A, cudaStream_t st);
int launch_energy_partial(const StepArrays& A, const double* pot, int pot_splits, long first, long count, double* out4, cudaStream_t st);
int launch_energy_final(const double* in4, double* out3, cudaStream_t st);
int launch_snapshot(const double4* pos, const double4* vel, long n, double* particleData6, double* staticData2, cudaStream_t st);
int launch_accel_export(const double* acc, long npad, long first, long count, double4* out, cudaStream_t st);
int launch_accel_import_sum(const double* part, int S, long npad, long n, double4* out, cudaStream_t st);
int launch_mass_sum(const double4* pos, long n, double* red_part, double* red, cudaStream_t st);
int launch_momentum(const double4* pos, const double4* vel, long first, long count, double* red_part, double* red, cudaStream_t st);
int launch_spiral_init(unsigned long long seed, long n, double total_mass, int arms, double radius, double thickness,
                       double perturbation, double4* pos, double4* vel, double* normal3, cudaStream_t st);

// nb_sort.cu
struct SortScratch {
    unsigned long long *keys_in = nullptr, *keys_out = nullptr;
    unsigned int* idx_in = nullptr;
    double *extent_part = nullptr, *box = nullptr;
    void* cub_temp = nullptr; size_t cub_bytes = 0;
};
size_t sort_temp_bytes(long n);
int build_permutation(const double4* src_pos, long n, int spatial /* 0 none, 1 Hilbert, 2 Morton */, SortScratch& S, unsigned int* perm, cudaStream_t st);
int launch_gather_state(const double4* src_pos, const double4* src_vel, const unsigned int* perm, long n, long npad,
                        double4* pos, double4* vel, cudaStream_t st);
int launch_permute4(const double4* src, const unsigned int* p2, long n, double4* dst, cudaStream_t st);
int launch_permute_acc(const double* acc, long npad, const unsigned int* p2, long n, double* dst, cudaStream_t st);
int launch_compose_perm(const unsigned int* perm, const unsigned int* p2, long n, unsigned int* dst, cudaStream_t st);
int launch_scatter4(const double4* src, const unsigned int* perm, long first, long count, double4* dst, cudaStream_t st);
int launch_scatter_acc(const double* acc, long npad, const unsigned int* perm, long first, long count, double4* dst, cudaStream_t st);
int launch_snapshot_scatter(const double4* pos, const double4* vel, const unsigned int* perm, long n,
                            double* particleData6, double* staticData2, cudaStream_t st);


// nb_render.cu
struct RenderState {
    bool have_mass = false; float min_mass = 1.f, max_mass = 1.f; bool have_vel = false; float smoothed = 0.f;
    // device scratch kept across frames (grown on demand, never shrunk): [0] what depends on n and the image size,
    // [1] what depends on the number of fragments. Released with the context (its device current).
    void* buf[2] = {nullptr, nullptr}; size_t cap[2] = {0, 0};
    ~RenderState() { cudaFree(buf[0]); cudaFree(buf[1]); cudaGetLastError(); }
};
int render_splat(const double4* pos, const double4* vel, const unsigned int* perm, long n, int W, int H,
                 float rot_x, float rot_y, float distance, float tx, float ty, float scale, float point_size, int color_mode,
                 const float color4[4], const float clear4[4], RenderState& st, float* range_out4, unsigned char* h_rgba,
                 cudaStream_t stream);

} // namespace nb
