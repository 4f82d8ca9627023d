/* include/nbody_b200.h — C ABI of libnbody_b200.so
 *
 * B200-native (sm_100a) replacement for ONE hot path of tusca99/N-body-simulation: the all-pairs gravity force
 * evaluation (PAIRWISE, ADAPTIVE_MUTUAL) fused with the Velocity-Verlet / Euler step, i.e. what the reference runs
 * behind config.json executionMode:"GPU". Plain C: pointers, integers and doubles only; no C++/torch types.
 * Every function returns NB_OK (0) or an NB_ERR_* code; nb_last_error() gives the message for the calling thread.
 * Nothing here ever exits the process, throws across the boundary, or falls back to a CPU implementation.
 *
 * Citations are file:line into the reference tree (simulator/...). INTEGRATION.md shows the reference-side binding.
 *
 * Particle layout is the reference's (core/Particles.hpp:9-16): AoS double4, posMass4[i] = (x,y,z,m),
 * vel4[i] = (vx,vy,vz,0); units AU, Msun, yr; G = 39.47841760435743 (cuda/Constants.cuh:6).
 */
#ifndef NBODY_B200_H
#define NBODY_B200_H

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes -------------------------------------------------------------------------------------------- */
#define NB_OK               0
#define NB_ERR_INVALID      1   /* bad argument / call order                                  */
#define NB_ERR_CUDA         2   /* CUDA runtime error (replaces gpuErrchk's exit(), cuda/CudaUtils.cu:5-10) */
#define NB_ERR_NCCL         3
#define NB_ERR_UNSUPPORTED  4   /* reference-only method (BARNES_HUT, VISUALIZATION)                         */
#define NB_ERR_NO_DEVICE    5   /* replaces the std::runtime_error of cuda/SimulationKernels.cu:35-50         */

/* ---- enums: numeric values are the reference's enum-class values ------------------------------------------------ */
/* IntegrationMethod, core/System.hpp:15-18 */
#define NB_INTEG_EULER            0
#define NB_INTEG_VELOCITY_VERLET  1
/* ForceMethod, core/Forces.hpp:40-46 */
#define NB_FORCE_PAIRWISE            0
#define NB_FORCE_PAIRWISE_AVX2_FP32  1   /* semantics of the reference's GPU kernel calculatePairwiseAccelOpt
                                          * (cuda/ForceKernels.cu:173-204): Plummer softening r^2 + 1e-10, no cut-off */
#define NB_FORCE_ADAPTIVE_MUTUAL     2
#define NB_FORCE_BARNES_HUT          3   /* reference-only: NB_ERR_UNSUPPORTED */
/* Arithmetic of the pair loop (new, optional; the reference GPU path is FP64 only) */
#define NB_PREC_FP32   0   /* two-float positions; FP32 pair math for far tiles, FP64 for adjacent tiles, FP64 sums; <=1e-5 */
#define NB_PREC_FP64   1   /* FP64 pair math throughout; <=1e-12                                                    */

typedef struct nb_ctx nb_ctx;

/* Message of the last failing call on this thread ("" if none). Never NULL. */
const char* nb_last_error(void);
/* Library/ABI version and the SM architecture the kernels were built for (100). */
int nb_abi_version(void);
int nb_built_sm(void);
/* Number of CUDA devices visible (0 when none); NB_ERR_NO_DEVICE semantics are left to the caller. */
int nb_device_count(void);

/* ---- context: device memory that persists across calls -----------------------------------------------------------
 * Replaces the per-call cudaMalloc/cudaFree of cuda/BenchmarkKernels.cu:28-37,105-113 and
 * cuda/SimulationKernels.cu:82-117,303-323. One context = one device = one host thread at a time. */
int nb_create(nb_ctx** out, long n, int device, int precision);
int nb_destroy(nb_ctx* ctx);

/* Methods used by nb_compute_accel / nb_step. subtract_com != 0 removes the centre-of-mass velocity at the start of
 * every step like the reference CPU integrators (core/System.cpp:45-57,392-400,438-446; the parity target);
 * 0 reproduces the reference GPU drivers, which never do. Default after nb_create: VELOCITY_VERLET, PAIRWISE, 1. */
int nb_set_methods(nb_ctx* ctx, int integration_method, int force_method, int subtract_com);

/* Multi-GPU: this context owns the i-particles [n*rank/world, n*(rank+1)/world) (rounded to whole j-tiles) and
 * all-gathers packed positions every step. No counterpart in the reference (single GPU).
 * nb_partition_bounds is pure host arithmetic and needs no device. */
int nb_partition_bounds(long n, int rank, int world, long* i_begin, long* i_count);
int nb_comm_unique_id(void* id128);                               /* 128-byte ncclUniqueId, rank 0 creates it */
int nb_comm_init(nb_ctx* ctx, const void* id128, int rank, int world);

/* Host -> device of the whole system (cuda/BenchmarkKernels.cu:40-41). Buffers may be pinned or pageable. On the device
 * the particles are kept in Hilbert-curve order so that a tile of 256 consecutive particles is spatially compact;
 * every entry point that returns per-particle data applies the inverse permutation, so callers only ever see their
 * own order. Multi-GPU: collective; all ranks pass identical arrays holding the whole system, each rank copies only its
 * share of the rows to its device (rows [ceil(n/world)*rank, ...)) and the shares are all-gathered over NVLink.
 * FP32-mode contexts return NB_ERR_INVALID for a negative or NaN mass (the far field works in a mass-scaled frame);
 * FP64-mode contexts take any mass, like the reference's arithmetic. */
int nb_upload(nb_ctx* ctx, const double* posMass4, const double* vel4);
/* SPIRAL_GALAXY initial conditions made on the device instead of nb_upload (Initializer::initSpiralGalaxy,
 * core/Initializer.cpp:360-469: same distribution, formulas and parameters — total mass, arms, radius, thickness, velocity
 * perturbation — as the reference generator). Every draw is a function of (seed, particle index, draw number), so the result
 * does not depend on the number of GPUs or threads; it is NOT the std::mt19937 sequence of the host generator. Particle i of
 * the caller's order is what nb_download returns in row i. normal3 (may be NULL) receives the disc's unit normal.
 * Multi-GPU: every rank calls it with the same arguments. */
int nb_init_spiral(nb_ctx* ctx, unsigned long long seed, double total_mass, int arms, double radius, double thickness,
                   double perturbation, double* normal3);
/* Device -> host of the final state (cuda/BenchmarkKernels.cu:101-102), all n rows, caller order. Multi-GPU: collective
 * (the slices are all-gathered first), every rank receives the complete state. */
int nb_download(nb_ctx* ctx, double* posMass4, double* vel4);

/* Multi-GPU egress without the gather: only the particles this rank owns, in the device's curve order. posMass4 / vel4 /
 * index receive *count rows (capacity rows must be available; the slice length of nb_partition_bounds suffices); index[k] is
 * the caller's index of row k. Not collective; the ranks' index sets are disjoint and cover 0..n-1. Single GPU: all n rows. */
int nb_download_slice(nb_ctx* ctx, double* posMass4, double* vel4, unsigned int* index, long capacity, long* count);

/* a = F(x) for the current positions: the initial launchForceKernel of cuda/BenchmarkKernels.cu:68 /
 * cuda/ForceKernels.cu:208-232. Must precede the first nb_step after an upload (nb_step does it if missing). */
int nb_compute_accel(nb_ctx* ctx);
/* Parity hook: copy accelerations to the host as n x double4 (ax,ay,az,0), the layout of the reference's d_accel
 * (cuda/ForceKernels.cu:94), caller order. Multi-GPU: collective, every rank receives all rows. */
int nb_get_accel(nb_ctx* ctx, double* accel4);

/* nsteps x performGPUIntegrationStep (cuda/IntegrationKernels.cu:200-266). Asynchronous: returns after enqueueing;
 * nb_sync / nb_download / nb_energy wait. */
int nb_step(nb_ctx* ctx, double dt, int nsteps);
int nb_sync(nb_ctx* ctx);

/* out[0]=E, out[1]=kinetic, out[2]=potential with the semantics of System::computeTotalEnergy
 * (core/System.cpp:60-85: unsoftened, every pair with r>0). O(N^2) in FP64. Multi-GPU: collective — each rank sums the
 * pair potential of its own slice against the whole system, the kinetic and potential sums are all-reduced, every rank
 * receives the same three numbers. */
int nb_energy(nb_ctx* ctx, double out[3]);
/* out = (sum m vx, sum m vy, sum m vz, sum m). */
int nb_momentum(nb_ctx* ctx, double out[4]);
/* Mean mass as computeAverageMass (core/Forces.cpp:85-95). */
int nb_average_mass(nb_ctx* ctx, double* out);

/* FILE_CSV sink (cuda/IntegrationKernels.cu:163-197; output/OutputData.hpp:14-31,387): writes one snapshot row.
 * particleData6: n x (x,y,z,vx,vy,vz); systemData2: (time, energy) — energy is computed here; staticData2 (may be
 * NULL): n x (id, mass). Host pointers. Multi-GPU: collective; every rank's buffers receive the whole system. */
int nb_snapshot(nb_ctx* ctx, double time, double* particleData6, double* systemData2, double* staticData2);

/* Headless image of the current state: the arithmetic of the reference viewer's shaders and fixed-function state
 * (visualization/ShaderManager.cpp:147-384, CameraController.cpp:35-102, ParticleRenderer.cpp:233-306,414-465,
 * visualization/cuda/VisualizationKernels.cu:28-49) evaluated by CUDA kernels into an RGBA8 image (rows top-down), since the
 * target machine has no graphics stack. Camera: rot_x, rot_y, distance, translate_x/y (defaults of the reference: 0.2, 0.9,
 * 5, 0, 0); scale <= 0: 2 / max |r| (core/System.cpp:236-248); point_size <= 0: 30; color_mode 0 UNIFORM | 1
 * VELOCITY_MAGNITUDE; color4 / clear4 NULL: (1,1,1,0.9) / (0.02,0.02,0.05,1). The mass range is taken at the first call
 * and the colour-ramp maximum speed is the 0.9/0.1 moving average over calls, as in the reference. info4 (may be NULL)
 * receives (min mass, max mass, max speed used, scale used). */
int nb_render(nb_ctx* ctx, int width, int height, float rot_x, float rot_y, float distance, float translate_x, float translate_y,
              float scale, float point_size, int color_mode, const float* color4, const float* clear4, unsigned char* rgba8,
              float* info4);

/* Kernel launches enqueued by this context since creation (force + integrator + reduction kernels). */
long nb_launch_count(nb_ctx* ctx);
/* Duration in ms of the most recent force-kernel launch(es) of one force evaluation, measured with CUDA events on the
 * context's stream (valid after nb_sync). */
int nb_last_force_ms(nb_ctx* ctx, float* ms);
/* Mean duration (ms) of the force evaluations since the last reset (at most the 64 most recent) and how many were
 * averaged; waits for the stream. reset != 0 starts a new measurement window. */
int nb_force_timing(nb_ctx* ctx, int reset, float* mean_ms, int* evals);
/* As nb_force_timing, for the FP64 near-field kernel alone (its own stream; it shares the SMs with the far kernel, so this is
 * its span, not its cost). Call before nb_force_timing(reset=1). evals = 0 on routes without a separate near kernel. */
int nb_near_timing(nb_ctx* ctx, float* mean_ms, int* evals);
/* order[k] = caller index of the particle currently in slot k of the device's space-filling-curve order (n entries, host
 * pointer). Multi-GPU: rank r owns slots [r*per, (r+1)*per), per = nb_partition_bounds' slice length. Lets a checker pick
 * the particles on either side of a slice boundary. */
int nb_get_order(nb_ctx* ctx, unsigned int* order);
/* Diagnostics of the production FP32 kernel (after nb_set_option(ctx, "collect_stats", 1), which also resets them):
 * out[0] j-tiles taken through the CTA-frame far path, out[1] j-tiles refined per warp (both counted once per CTA),
 * out[2] 64-particle sub-tiles through the warp-frame far path, out[3] sub-tiles evaluated in FP64 (once per warp),
 * out[4] the largest number of FP64 sub-tiles any single warp (of one j-split) has. */
int nb_tile_stats(nb_ctx* ctx, unsigned long long out[5]);
/* Per-(warp, j-split) FP64 sub-tile counts of the last evaluation (zero-filled beyond the launched grid). */
int nb_tile_stats_warps(nb_ctx* ctx, unsigned long long* out, long cap, long* n_out);
/* CUDA-event stopwatch on the context's stream (torch.cuda.Event only sees torch's own streams):
 * nb_timer_start records the start event; nb_timer_stop records the stop event, waits for it and returns the device
 * time in ms of everything the context enqueued in between. */
int nb_timer_start(nb_ctx* ctx);
int nb_timer_stop(nb_ctx* ctx, float* ms);
/* Tuning knobs (optional; set before nb_upload):
 *   "force_variant"    0 auto (= 3)
 *                      | 1 plain-load tiles, global two-float coordinates for every tile (simple cross-check)
 *                      | 2 bulk-copy (TMA) tiles in the i-block's frame, 2 i-particles/thread, FP32 two-float near field
 *                      | 3 as 2 with 4 i-particles/thread and the near field (adjacent tiles) in FP64 (production)
 *   "j_splits"         0 auto | 1..64 CTAs sharing one i-block's j-range
 *   "near_splits"      0 auto | 1..64 j-splits of the FP64 near-field kernel (FP32 mode)
 *   "spatial_sort"     1 (default) order the particles along a 3-D Hilbert curve at upload | 2 Morton (Z) curve
 *                      | 0 keep the caller's order
 *   "collect_stats"    1 = count tile classifications (see nb_tile_stats); costs a few atomics per tile
 *   "resort_interval"  re-establish the curve order from the current positions every so many steps (default 16; 0 never);
 *                      results stay correct without it, but tiles that lose their compactness take the slower near path
 *   "far_kappa_milli"  1000*kappa of the far-tile criterion box gap > kappa*(half-extent_i + half-extent_j)
 *                      (default 250); 0 = every tile through the near-field path
 *   "small_n_fp64"     FP32-mode context, force_variant 0, one GPU: systems of at most this many particles are evaluated
 *                      by the FP64 kernel (one launch, no near/far split: faster below ~10^4 particles, where a step is
 *                      latency-bound, and exact to 1e-12). Default 0 for nb_create contexts; the whole-run entry points
 *                      below use 8192.
 *   "step_graph"       1 = systems of at most 16384 particles on the FP64-kernel route replay one captured step (drift,
 *                      force, finish, COM) as a CUDA graph instead of four launches; bit-identical results. Default 0
 *                      for nb_create contexts; the whole-run entry points below switch it on. */
int nb_set_option(nb_ctx* ctx, const char* key, long value);

/* ---- whole-run entry points with HOST buffers: what the reference's host code binds ------------------------------ */
/* runBenchmarkOnGPU(Particles&, IntegrationMethod, ForceMethod, dt, steps, BLOCK_SIZE)
 * (cuda/BenchmarkKernels.cu:15-126, cuda/BenchmarkKernels.h:12-15). In place on posMass4/vel4. block_size is accepted
 * and ignored (the launch configuration is fixed by the kernels). device<0: current device. */
int nb_run_benchmark(long n, double* posMass4, double* vel4, int integration_method, int force_method,
                     double dt, int steps, int block_size, int precision, int subtract_com, int device);
/* runSimulationOnGPU(Particles&, method, forceMethod, dt, steps, stepFreq, OutputData&)
 * (cuda/SimulationKernels.cu:22-286): snapshot 0 is the initial state, snapshot k is taken every step_freq steps;
 * h_particleData [T][n][6], h_systemData [T][2], h_staticData [n][2] are the OutputData host arrays
 * (output/OutputData.hpp:14-31); num_time_steps = T = steps/step_freq + 1 (core/System.cpp:107). */
int nb_run_simulation(long n, double* posMass4, double* vel4, int integration_method, int force_method,
                      double dt, int steps, int step_freq,
                      double* h_particleData, double* h_systemData, double* h_staticData, long num_time_steps,
                      int precision, int subtract_com, int device);

/* runVisualizationOnGPU(particles, method, forceMethod, dt, steps, stepFreq, physicsStepsPerFrame)
 * (visualization/cuda/VisualizationKernels.cu:52-300) without a window: the same loop — initial force, then per frame
 * physics_steps_per_frame integrator steps (<= 0: 1, as the reference) followed by one image of the state — with the
 * frame rendered by nb_render's CUDA kernels (reference camera defaults, VELOCITY_MAGNITUDE colours, 30 px points,
 * max-speed moving average across frames) instead of OpenGL, which is unavailable on the target machine.
 * scale > 0: world-to-view factor (the reference's GPU driver uses 1e-3, VisualizationUtils.cpp:19); scale <= 0:
 * 2 / max |r| of the initial state, kept for the whole run (the reference's CPU viewer, core/System.cpp:236-248).
 * frame_dir != NULL: every frame_stride-th frame (<= 0: 1) is written as binary PPM frame_dir/frame_NNNNNN.ppm (the
 * directory must exist); NULL: frames are rendered and dropped. last_rgba8 (may be NULL): width*height*4 bytes, receives
 * the last frame. frames2 (may be NULL) receives (frames rendered, frames written). The final state is written back into
 * posMass4 / vel4 like the reference does (:259-261). */
int nb_run_visualization(long n, double* posMass4, double* vel4, int integration_method, int force_method,
                         double dt, int steps, int physics_steps_per_frame, int width, int height, float scale,
                         const char* frame_dir, int frame_stride, unsigned char* last_rgba8, long* frames2,
                         int precision, int subtract_com, int device);

/* ---- device-pointer entry points: the inner boundary of the reference drivers ------------------------------------ */
/* launchForceKernel(method, blocks, blockSize, smem, stream, d_posMass, d_accel, avgMass, n)
 * (cuda/ForceKernels.cu:208-232). d_posMass4 / d_accel4 are caller-owned device double4 arrays; the context provides
 * the packed-tile scratch. avg_mass <= 0: computed on the device. stream: cudaStream_t; the call is ordered after the work already
 * enqueued on it and the stream waits for the result; NULL = the legacy default stream, and the call then also blocks the host until the result is there.
 * The call replaces the context's particle state: nb_step afterwards needs a fresh nb_upload. Masses are not inspected
 * here (they are on the device): in an FP32-mode context they must be >= 0. */
int nb_launch_force(nb_ctx* ctx, int force_method, const double* d_posMass4, double* d_accel4,
                    double avg_mass, long n, void* stream);

/* ---- allocation-free force evaluation ---------------------------------------------------------------------------------
 * No counterpart in the reference (its drivers cudaMalloc on every call). One force evaluation a = F(x) on caller-owned
 * device arrays (n x double4 in, n x double4 (ax,ay,az,0) out, caller's particle order) that performs no device or
 * pinned allocation at all: the packed tiles and partial sums live in a fixed ~7.6 MB block of device memory built into
 * the library (held by every GPU that loads it, until the process exits). For callers that must not allocate after
 * start-up: plug-ins inside an application whose own pool already owns the device memory (cudaMalloc would fail),
 * per-frame / per-step loops that cannot take cudaMalloc's device-wide synchronisation, jobs with a fixed, declared
 * memory footprint. Limits and rules:
 *   - n <= NB_NOALLOC_MAX_N, otherwise NB_ERR_INVALID with an explanatory message;
 *   - PAIRWISE, PAIRWISE_AVX2_FP32 and ADAPTIVE_MUTUAL; ADAPTIVE_MUTUAL needs avg_mass > 0 (the mean mass of the system);
 *   - FP32 pair math on two-float positions for every pair, FP64 sums; there is no FP64 near field as in a context, so the
 *     error grows with the cancellation among a particle's nearest neighbours: 2.3e-6 (spiral, 20000), 3.7e-6 (galaxy +
 *     planets: 4501 ADAPTIVE_MUTUAL; 27001: 2.5e-6 PAIRWISE, 3.5e-6 ADAPTIVE_MUTUAL), 2.1e-6 (12000 particles 6e4 AU from
 *     the origin) — the largest relative error per particle seen in tests/test_zz_no_alloc.py, whose bound is 1e-5;
 *   - asynchronous on `stream` (cudaStream_t; NULL = default stream); the device current at the call is used;
 *   - the first call on each device resolves the block's device address; issue it during start-up, before any stream
 *     capture begins; the address is kept for the life of the process, so do not use this entry point after a
 *     cudaDeviceReset;
 *   - all calls on a device share the block: calls that overlap on the GPU race — keep them on one stream. */
#define NB_NOALLOC_MAX_N 32768
int nb_force_noalloc(int force_method, const double* d_posMass4, double* d_accel4, long n, double avg_mass, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NBODY_B200_H */
