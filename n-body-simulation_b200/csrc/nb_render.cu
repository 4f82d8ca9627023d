// nb_render.cu — headless CUDA point-splat of the reference viewer's shader arithmetic.
//
// The reference draws the particles with OpenGL (GLFW window, GL 3.3 point sprites); there is no graphics stack on the
// target machine, so the same arithmetic is evaluated by CUDA kernels into a device RGBA8 buffer:
//   * VBO update          visualization/cuda/VisualizationKernels.cu:28-49   (double -> float, positions x scale)
//   * camera matrices     visualization/CameraController.cpp:35-102          (perspective 45 deg, rotY * rotX, translate)
//   * vertex shader       visualization/ShaderManager.cpp:213-286            (log-mass point size, perspective scaling)
//   * fragment shader     visualization/ShaderManager.cpp:288-384            (velocity colour ramp, mass brightness, edge alpha)
//   * sprite texture      visualization/ShaderManager.cpp:147-211            (alpha = 1 inside 0.6, smoothstep to 1.0)
//   * fixed function      visualization/ParticleRenderer.cpp:414-465         (clear, SRC_ALPHA / ONE_MINUS_SRC_ALPHA blend,
//                                                                             depth test LEQUAL, points in buffer order)
//   * mass range / max-speed EMA 0.9/0.1   visualization/ParticleRenderer.cpp:233-306
// What OpenGL leaves to the implementation is fixed as follows (and restated identically by the oracle): a point sprite
// of size s covers the pixels whose centres lie in [xw - s/2, xw + s/2) x [yw - s/2, yw + s/2); s is clamped to
// [1, 1024] and not rounded; points whose centre is outside the clip volume are culled; the sprite texture is sampled
// analytically (no mip-mapping, no 8-bit texel quantisation); depth is a float; the colour buffer is RGBA8 and every
// blend result is rounded to 8 bits, as a UNORM8 framebuffer does. The coordinate axes (GL_LINES) are not drawn.
//
// Order-dependent blending (points are blended in buffer order, each subject to the depth test) is reproduced exactly:
// fragments are generated per particle, sorted by (pixel, particle index) with a radix sort, and each pixel then walks
// its fragments sequentially.
#include "nb_common.cuh"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

namespace nb {
namespace {

struct RenderParams {
    int W, H;
    float V[16], P[16];          // column-major, as handed to glUniformMatrix4fv(..., GL_FALSE, ...)
    float scale, point_size, min_mass, max_mass, max_velocity;
    int color_mode;              // 0 UNIFORM, 1 VELOCITY_MAGNITUDE (ParticleRenderer.h:59-62)
    float color[4];
    float clear[4];
};

struct Vert {                    // per particle, after the vertex stage
    float xw, yw, zw, size;      // window coordinates, depth in [0,1], sprite size in pixels (0: culled)
    float r, g, b, a;            // fragment colour before the sprite alpha
    float nmass;                 // normalised log-mass (edge width / opacity)
    int x0, x1, y0, y1;          // covered pixel range [x0, x1) x [y0, y1)
};

__host__ __device__ inline float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
__host__ __device__ inline float smoothstepf(float e0, float e1, float x) {
    const float t = clampf((x - e0) / (e1 - e0), 0.f, 1.f);
    return t * t * (3.f - 2.f * t);
}
__host__ __device__ inline float mixf(float a, float b, float t) { return a * (1.f - t) + b * t; }

// ShaderManager.cpp:299-343
__device__ inline void velocity_to_color(float vx, float vy, float vz, float mass, const RenderParams& p, float rgb[3]) {
    const float speed = sqrtf(vx * vx + vy * vy + vz * vz);
    const float vmax = fmaxf(p.max_velocity, 0.00001f);
    float ns = clampf(speed / vmax, 0.f, 1.f);
    ns = smoothstepf(0.f, 1.f, ns);
    float c0[3], c1[3], t;
    if (ns < 0.2f)      { t = ns / 0.2f;          c0[0] = 0.f; c0[1] = 0.2f; c0[2] = 0.8f; c1[0] = 0.f;  c1[1] = 0.7f; c1[2] = 0.9f; }
    else if (ns < 0.5f) { t = (ns - 0.2f) / 0.3f; c0[0] = 0.f; c0[1] = 0.7f; c0[2] = 0.9f; c1[0] = 0.8f; c1[1] = 0.8f; c1[2] = 0.f; }
    else if (ns < 0.8f) { t = (ns - 0.5f) / 0.3f; c0[0] = 0.8f; c0[1] = 0.8f; c0[2] = 0.f; c1[0] = 1.f;  c1[1] = 0.5f; c1[2] = 0.f; }
    else                { t = (ns - 0.8f) / 0.2f; c0[0] = 1.f; c0[1] = 0.5f; c0[2] = 0.f; c1[0] = 1.f;  c1[1] = 0.f;  c1[2] = 0.f; }
    float nm = 0.f;
    if (mass > 0.f) nm = clampf((logf(mass) - logf(p.min_mass)) / (logf(p.max_mass) - logf(p.min_mass)), 0.f, 1.f);
    const float boost = 1.f + 0.8f * nm;
#pragma unroll
    for (int k = 0; k < 3; ++k) rgb[k] = mixf(c0[k], c1[k], t) * boost;
}

__global__ void __launch_bounds__(kThreads)
vertex_kernel(const double4* __restrict__ pos, const double4* __restrict__ vel, const unsigned int* __restrict__ perm, long n,
              RenderParams p, Vert* __restrict__ verts, unsigned int* __restrict__ counts) {
    const long s = (long)blockIdx.x * kThreads + threadIdx.x;
    if (s >= n) return;
    const unsigned int i = perm[s];          // the reference draws in ITS buffer order: vertex i = caller's particle i
    const double4 q = pos[s], w = vel[s];
    // VisualizationKernels.cu:36-47
    const float px = (float)(q.x * (double)p.scale), py = (float)(q.y * (double)p.scale), pz = (float)(q.z * (double)p.scale);
    const float mass = (float)q.w;
    const float vx = (float)w.x, vy = (float)w.y, vz = (float)w.z;
    float vp[4], cp[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) vp[r] = fmaf(p.V[0 + r], px, fmaf(p.V[4 + r], py, fmaf(p.V[8 + r], pz, p.V[12 + r])));
#pragma unroll
    for (int r = 0; r < 4; ++r) cp[r] = fmaf(p.P[0 + r], vp[0], fmaf(p.P[4 + r], vp[1], fmaf(p.P[8 + r], vp[2], p.P[12 + r] * vp[3])));
    // point size, ShaderManager.cpp:233-282
    const float min_size = 2.f, max_size = 15.f;
    float size;
    if (mass <= 0.f) size = min_size * p.point_size / 1.f;
    else {
        const float l10 = logf(10.f);
        const float lm = logf(fmaxf(mass, 1e-20f)) / l10, lmin = logf(fmaxf(p.min_mass, 1e-20f)) / l10,
                    lmax = logf(fmaxf(p.max_mass, p.min_mass * 1.01f)) / l10;
        const float range = lmax - lmin;
        if (range < 1e-10f) size = (min_size + max_size) / 2.f * p.point_size / 30.f;
        else {
            const float np = clampf((lm - lmin) / range, 0.f, 1.f);
            const float base = min_size + np * (max_size - min_size);
            const float dcam = fmaxf(sqrtf(vp[0] * vp[0] + vp[1] * vp[1] + vp[2] * vp[2]), 0.1f);
            size = base * 8.f / dcam;
            size = size * (p.point_size / 30.f);
            size = fminf(size, p.point_size * 0.8f);
        }
    }
    Vert o;
    o.size = 0.f; o.x0 = o.x1 = o.y0 = o.y1 = 0;
    const float cw = cp[3];
    const bool visible = cw > 0.f && fabsf(cp[0]) <= cw && fabsf(cp[1]) <= cw && fabsf(cp[2]) <= cw;
    if (visible) {
        const float ndx = cp[0] / cw, ndy = cp[1] / cw, ndz = cp[2] / cw;
        o.xw = (ndx * 0.5f + 0.5f) * (float)p.W;
        o.yw = (ndy * 0.5f + 0.5f) * (float)p.H;
        o.zw = ndz * 0.5f + 0.5f;
        o.size = clampf(size, 1.f, 1024.f);
        const float h = 0.5f * o.size;
        o.x0 = max(0, (int)ceilf(o.xw - h - 0.5f)); o.x1 = min(p.W, (int)ceilf(o.xw + h - 0.5f));
        o.y0 = max(0, (int)ceilf(o.yw - h - 0.5f)); o.y1 = min(p.H, (int)ceilf(o.yw + h - 0.5f));
        if (o.x1 < o.x0) o.x1 = o.x0;
        if (o.y1 < o.y0) o.y1 = o.y0;
    }
    // fragment colour (constant over the sprite), ShaderManager.cpp:358-365
    float rgb[3] = {p.color[0], p.color[1], p.color[2]};
    if (p.color_mode == 1) velocity_to_color(vx, vy, vz, mass, p, rgb);
    o.r = rgb[0]; o.g = rgb[1]; o.b = rgb[2]; o.a = p.color[3];
    o.nmass = 0.f;
    if (mass > 0.f && p.max_mass > p.min_mass)
        o.nmass = clampf((logf(mass) - logf(p.min_mass)) / (logf(p.max_mass) - logf(p.min_mass)), 0.f, 1.f);
    verts[i] = o;
    counts[i] = (unsigned int)((o.x1 - o.x0) * (o.y1 - o.y0));
}

// sprite alpha of one fragment, ShaderManager.cpp:346-383 with the texture of :147-211 evaluated analytically; < 0: discard
__device__ inline float fragment_alpha(const Vert& v, int ix, int iy) {
    const float s = 0.5f + ((float)ix + 0.5f - v.xw) / v.size;
    const float t = 0.5f - ((float)iy + 0.5f - v.yw) / v.size;
    const float dx = s - 0.5f, dy = t - 0.5f;
    const float dist = sqrtf(dx * dx + dy * dy) * 2.f;
    if (dist > 1.f) return -1.f;
    const float tex_a = dist <= 0.6f ? 1.f : (dist < 1.f ? smoothstepf(1.f, 0.6f, dist) : 0.f);
    const float edge_w = mixf(0.3f, 0.1f, v.nmass);
    float a = smoothstepf(1.f, 1.f - edge_w, dist) * tex_a;
    a = mixf(a, a * 1.3f, v.nmass);
    return clampf(a, 0.f, 1.f);
}

__global__ void __launch_bounds__(kThreads)
fragment_kernel(const Vert* __restrict__ verts, const unsigned long long* __restrict__ offs, long n, int W,
                unsigned long long* __restrict__ keys, float* __restrict__ alphas) {
    const long i = (long)blockIdx.x * kThreads + threadIdx.x;
    if (i >= n) return;
    const Vert v = verts[i];
    unsigned long long o = offs[i];
    for (int iy = v.y0; iy < v.y1; ++iy)
        for (int ix = v.x0; ix < v.x1; ++ix) {
            keys[o] = ((unsigned long long)((unsigned)iy * (unsigned)W + (unsigned)ix) << 32) | (unsigned long long)i;
            alphas[o] = fragment_alpha(v, ix, iy);
            ++o;
        }
}

__device__ inline float q8(float x) { return rintf(clampf(x, 0.f, 1.f) * 255.f) / 255.f; }

// one thread per pixel: its fragments are contiguous in the sorted list and already in particle order
__global__ void __launch_bounds__(kThreads)
resolve_kernel(const unsigned long long* __restrict__ keys, const float* __restrict__ alphas, unsigned long long nfrag,
               const Vert* __restrict__ verts, RenderParams p, unsigned char* __restrict__ rgba) {
    const int pix = blockIdx.x * kThreads + threadIdx.x;
    if (pix >= p.W * p.H) return;
    // lower bound of (pix << 32)
    unsigned long long lo = 0, hi = nfrag;
    const unsigned long long want = (unsigned long long)pix << 32;
    while (lo < hi) { const unsigned long long mid = (lo + hi) >> 1; if (keys[mid] < want) lo = mid + 1; else hi = mid; }
    float r = q8(p.clear[0]), g = q8(p.clear[1]), b = q8(p.clear[2]), a = q8(p.clear[3]), depth = 1.f;
    for (unsigned long long f = lo; f < nfrag && (keys[f] >> 32) == (unsigned long long)pix; ++f) {
        const float fa = alphas[f];
        if (fa < 0.f) continue;                                  // discarded fragment: no colour, no depth write
        const Vert& v = verts[(unsigned int)(keys[f] & 0xffffffffu)];
        if (!(v.zw <= depth)) continue;                          // GL_LEQUAL
        depth = v.zw;
        r = q8(v.r * fa + r * (1.f - fa)); g = q8(v.g * fa + g * (1.f - fa)); b = q8(v.b * fa + b * (1.f - fa));
        a = q8(fa * fa + a * (1.f - fa));
    }
    const int iy = pix / p.W, ix = pix % p.W;
    unsigned char* o = rgba + ((size_t)(p.H - 1 - iy) * p.W + ix) * 4;      // image rows top-down, GL rows bottom-up
    o[0] = (unsigned char)rintf(r * 255.f); o[1] = (unsigned char)rintf(g * 255.f);
    o[2] = (unsigned char)rintf(b * 255.f); o[3] = (unsigned char)rintf(a * 255.f);
}

// mass range (positive masses) and max speed, ParticleRenderer.cpp:240-289: block partials then a single-block pass
__global__ void __launch_bounds__(kThreads)
range_partial_kernel(const double4* __restrict__ pos, const double4* __restrict__ vel, long n, float* __restrict__ part) {
    __shared__ float sh[4][kThreads / 32];
    const long i = (long)blockIdx.x * kThreads + threadIdx.x;
    float mn = 3.4e38f, mx = -3.4e38f, sp = 0.f, md = 0.f;
    if (i < n) {
        const float m = (float)pos[i].w;
        if (m > 0.f) { mn = m; mx = m; }
        const float vx = (float)vel[i].x, vy = (float)vel[i].y, vz = (float)vel[i].z;
        sp = sqrtf(vx * vx + vy * vy + vz * vz);
        md = (float)sqrt(pos[i].x * pos[i].x + pos[i].y * pos[i].y + pos[i].z * pos[i].z);      // core/System.cpp:238-243
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o)); mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        sp = fmaxf(sp, __shfl_xor_sync(0xffffffffu, sp, o)); md = fmaxf(md, __shfl_xor_sync(0xffffffffu, md, o));
    }
    if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = mn; sh[1][threadIdx.x >> 5] = mx; sh[2][threadIdx.x >> 5] = sp; sh[3][threadIdx.x >> 5] = md; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < kThreads / 32; ++k) { mn = fminf(mn, sh[0][k]); mx = fmaxf(mx, sh[1][k]); sp = fmaxf(sp, sh[2][k]); md = fmaxf(md, sh[3][k]); }
        part[blockIdx.x * 4 + 0] = mn; part[blockIdx.x * 4 + 1] = mx; part[blockIdx.x * 4 + 2] = sp; part[blockIdx.x * 4 + 3] = md;
    }
}
__global__ void range_final_kernel(const float* __restrict__ part, int nblocks, float* __restrict__ out) {
    float mn = 3.4e38f, mx = -3.4e38f, sp = 0.f, md = 0.f;
    for (int k = 0; k < nblocks; ++k) { mn = fminf(mn, part[k * 4]); mx = fmaxf(mx, part[k * 4 + 1]); sp = fmaxf(sp, part[k * 4 + 2]); md = fmaxf(md, part[k * 4 + 3]); }
    out[0] = mn; out[1] = mx; out[2] = sp; out[3] = md;
}

inline unsigned blocks_for(long count) { return (unsigned)((count + kThreads - 1) / kThreads); }

} // namespace

// CameraController.cpp:35-102
void camera_matrices(float rot_x, float rot_y, float distance, float tx, float ty, float aspect, float* V, float* P) {
    const float fovy = 45.0f, nz = 0.1f, fz = 1000.0f;
    const float f = 1.0f / tanf(fovy * 0.5f * 3.14159f / 180.0f);
    for (int i = 0; i < 16; ++i) P[i] = 0.f;
    P[0] = f / aspect; P[5] = f; P[10] = (fz + nz) / (nz - fz); P[11] = -1.0f; P[14] = (2.0f * fz * nz) / (nz - fz);
    const float cx = cosf(rot_x), sx = sinf(rot_x), cy = cosf(rot_y), sy = sinf(rot_y);
    const float rx[16] = {1, 0, 0, 0, 0, cx, -sx, 0, 0, sx, cx, 0, 0, 0, 0, 1};
    const float ry[16] = {cy, 0, sy, 0, 0, 1, 0, 0, -sy, 0, cy, 0, 0, 0, 0, 1};
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) { float t = 0; for (int k = 0; k < 4; ++k) t += ry[i * 4 + k] * rx[k * 4 + j]; V[i * 4 + j] = t; }
    V[12] = tx; V[13] = ty; V[14] = -distance;
}


int render_splat(const double4* pos, const double4* vel, const unsigned int* perm, long n, int W, int H,
                 float rot_x, float rot_y, float distance, float tx, float ty, float scale, float point_size, int color_mode,
                 const float color4[4], const float clear4[4], RenderState& st, float* range_out3, unsigned char* h_rgba,
                 cudaStream_t stream) {
    if (W <= 0 || H <= 0 || (long)W * H > (1L << 26)) return fail(NB_ERR_INVALID, "nb_render: image size %dx%d", W, H);
    // Scratch that survives between frames: a render loop (nb_run_visualization) calls this once per frame between
    // asynchronous steps, and a cudaMalloc / cudaFree pair per buffer would synchronise the device against them every time.
    auto reserve = [&](int which, size_t bytes) -> int {
        if (bytes <= st.cap[which]) return NB_OK;
        NB_CUDA(cudaStreamSynchronize(stream));
        cudaFree(st.buf[which]); st.buf[which] = nullptr; st.cap[which] = 0;
        const size_t want = bytes + bytes / 4;                 // head room: fragment counts drift from frame to frame
        NB_CUDA(cudaMalloc(&st.buf[which], want));
        st.cap[which] = want;
        return NB_OK;
    };
    auto up = [](size_t b) { return (b + 255) / 256 * 256; };
    // mass range once, max-speed EMA every call (ParticleRenderer.cpp:233-306)
    const int nb = (int)blocks_for(n);
    size_t scan_bytes = 0;
    NB_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (unsigned int*)nullptr, (unsigned long long*)nullptr, (int)n, stream));
    const size_t b_part = up((size_t)nb * 4 * sizeof(float) + 4 * sizeof(float)), b_verts = up((size_t)n * sizeof(Vert)),
                 b_counts = up((size_t)n * sizeof(unsigned int)), b_offs = up((size_t)(n + 1) * sizeof(unsigned long long)),
                 b_img = up((size_t)W * H * 4), b_scan = up(scan_bytes ? scan_bytes : 16);
    { int rc0 = reserve(0, b_part + b_verts + b_counts + b_offs + b_img + b_scan); if (rc0) return rc0; }
    char* base0 = static_cast<char*>(st.buf[0]);
    float* d_part = reinterpret_cast<float*>(base0);
    float* d_out = d_part + (size_t)nb * 4;
    range_partial_kernel<<<nb, kThreads, 0, stream>>>(pos, vel, n, d_part);
    range_final_kernel<<<1, 1, 0, stream>>>(d_part, nb, d_out);
    float r3[4];
    NB_CUDA(cudaMemcpyAsync(r3, d_out, sizeof(r3), cudaMemcpyDeviceToHost, stream));
    NB_CUDA(cudaStreamSynchronize(stream));
    if (!st.have_mass) {
        if (r3[0] < r3[1]) { st.min_mass = r3[0]; st.max_mass = r3[1]; }
        else { st.min_mass = r3[0] < 3e38f ? r3[0] : 1.f; st.max_mass = st.min_mass * 10.0f; }      // :255-262 fallback
        st.have_mass = true;
    }
    if (st.smoothed == 0.0f || !st.have_vel) { st.smoothed = r3[2]; st.have_vel = true; }
    else st.smoothed = st.smoothed * 0.9f + r3[2] * 0.1f;
    RenderParams p;
    p.W = W; p.H = H;
    camera_matrices(rot_x, rot_y, distance, tx, ty, (float)W / (float)H, p.V, p.P);
    if (!(scale > 0.f)) scale = r3[3] > 0.f ? 2.0f / r3[3] : 1.f;      // core/System.cpp:245-248: targetViewSize / maxDistance
    p.scale = scale; p.point_size = point_size; p.min_mass = st.min_mass; p.max_mass = st.max_mass;
    p.max_velocity = fmaxf(st.smoothed, 0.001f); p.color_mode = color_mode;
    for (int k = 0; k < 4; ++k) { p.color[k] = color4[k]; p.clear[k] = clear4[k]; }
    if (range_out3) { range_out3[0] = p.min_mass; range_out3[1] = p.max_mass; range_out3[2] = p.max_velocity; range_out3[3] = p.scale; }

    Vert* verts = reinterpret_cast<Vert*>(base0 + b_part);
    unsigned int* counts = reinterpret_cast<unsigned int*>(base0 + b_part + b_verts);
    unsigned long long* offs = reinterpret_cast<unsigned long long*>(base0 + b_part + b_verts + b_counts);
    unsigned char* d_img = reinterpret_cast<unsigned char*>(base0 + b_part + b_verts + b_counts + b_offs);
    void* scan_tmp = base0 + b_part + b_verts + b_counts + b_offs + b_img;
    vertex_kernel<<<blocks_for(n), kThreads, 0, stream>>>(pos, vel, perm, n, p, verts, counts);
    NB_CUDA(cudaGetLastError());
    size_t tb = b_scan;
    NB_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp, tb, counts, offs, (int)n, stream));
    unsigned long long last_off = 0; unsigned int last_cnt = 0;
    NB_CUDA(cudaMemcpyAsync(&last_off, offs + (n - 1), sizeof(last_off), cudaMemcpyDeviceToHost, stream));
    NB_CUDA(cudaMemcpyAsync(&last_cnt, counts + (n - 1), sizeof(last_cnt), cudaMemcpyDeviceToHost, stream));
    NB_CUDA(cudaStreamSynchronize(stream));
    const unsigned long long nfrag = last_off + last_cnt;
    if (nfrag > (1ULL << 30)) return fail(NB_ERR_INVALID, "nb_render: %llu fragments (sprites too large for this image); reduce point_size", nfrag);
    unsigned long long *keys = nullptr, *keys2 = nullptr; float *al = nullptr, *al2 = nullptr;
    if (nfrag > 0) {
        size_t sort_bytes = 0;
        NB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, keys, keys2, al, al2, (int)nfrag, 0, 64, stream));
        const size_t b_keys = up(nfrag * 8), b_al = up(nfrag * 4), b_sort = up(sort_bytes ? sort_bytes : 16);
        { int rc1 = reserve(1, 2 * b_keys + 2 * b_al + b_sort); if (rc1) return rc1; }
        char* base1 = static_cast<char*>(st.buf[1]);
        keys = reinterpret_cast<unsigned long long*>(base1); keys2 = reinterpret_cast<unsigned long long*>(base1 + b_keys);
        al = reinterpret_cast<float*>(base1 + 2 * b_keys); al2 = reinterpret_cast<float*>(base1 + 2 * b_keys + b_al);
        void* sort_tmp = base1 + 2 * b_keys + 2 * b_al;
        fragment_kernel<<<blocks_for(n), kThreads, 0, stream>>>(verts, offs, n, W, keys, al);
        NB_CUDA(cudaGetLastError());
        size_t sb = b_sort;
        NB_CUDA(cub::DeviceRadixSort::SortPairs(sort_tmp, sb, keys, keys2, al, al2, (int)nfrag, 0, 64, stream));
    }
    resolve_kernel<<<blocks_for((long)W * H), kThreads, 0, stream>>>(keys2, al2, nfrag, verts, p, d_img);
    NB_CUDA(cudaGetLastError());
    NB_CUDA(cudaMemcpyAsync(h_rgba, d_img, (size_t)W * H * 4, cudaMemcpyDeviceToHost, stream));
    NB_CUDA(cudaStreamSynchronize(stream));
    return NB_OK;
}

} // namespace nb
