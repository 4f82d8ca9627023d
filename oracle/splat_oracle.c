/* oracle/splat_oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * Sequential plain-C restatement of the reference viewer's per-particle / per-fragment arithmetic, used to check the CUDA
 * point-splat (csrc/nb_render.cu). Parity unpinned: the reference evaluates this arithmetic in GLSL on an OpenGL device,
 * which is unavailable on this box, so there is no reference output to pin against; this file restates the shader and
 * fixed-function arithmetic line by line instead:
 *   VBO update            simulator/visualization/cuda/VisualizationKernels.cu:28-49
 *   camera matrices       simulator/visualization/CameraController.cpp:35-102
 *   vertex shader         simulator/visualization/ShaderManager.cpp:213-286
 *   fragment shader       simulator/visualization/ShaderManager.cpp:288-384
 *   sprite texture        simulator/visualization/ShaderManager.cpp:147-211 (sampled analytically)
 *   blend / depth state   simulator/visualization/ParticleRenderer.cpp:414-465
 *   mass range, max speed simulator/visualization/ParticleRenderer.cpp:233-306; scale core/System.cpp:236-248
 * Implementation-defined GL behaviour is fixed exactly as documented at the top of csrc/nb_render.cu.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

static float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
static float smoothstepf(float e0, float e1, float x) { float t = clampf((x - e0) / (e1 - e0), 0.f, 1.f); return t * t * (3.f - 2.f * t); }
static float mixf(float a, float b, float t) { return a * (1.f - t) + b * t; }
static float q8(float x) { return rintf(clampf(x, 0.f, 1.f) * 255.f) / 255.f; }

/* out4 = (min positive mass, max positive mass, max speed, max distance from the origin) */
void nbo_splat_ranges(long n, const double* pm, const double* v, float out4[4]) {
    float mn = 3.4e38f, mx = -3.4e38f, sp = 0.f, md = 0.f;
    for (long i = 0; i < n; ++i) {
        const float m = (float)pm[4 * i + 3];
        if (m > 0.f) { mn = fminf(mn, m); mx = fmaxf(mx, m); }
        const float vx = (float)v[4 * i], vy = (float)v[4 * i + 1], vz = (float)v[4 * i + 2];
        sp = fmaxf(sp, sqrtf(vx * vx + vy * vy + vz * vz));
        md = fmaxf(md, (float)sqrt(pm[4 * i] * pm[4 * i] + pm[4 * i + 1] * pm[4 * i + 1] + pm[4 * i + 2] * pm[4 * i + 2]));
    }
    out4[0] = mn; out4[1] = mx; out4[2] = sp; out4[3] = md;
}

static void camera(float rot_x, float rot_y, float distance, float tx, float ty, float aspect, float* V, float* P) {
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

int nbo_splat(long n, const double* pm, const double* vel, int W, int H, float rot_x, float rot_y, float distance, float tx,
              float ty, float scale, float point_size, int color_mode, const float color4[4], const float clear4[4],
              float min_mass, float max_mass, float max_velocity, unsigned char* rgba8) {
    float V[16], P[16];
    camera(rot_x, rot_y, distance, tx, ty, (float)W / (float)H, V, P);
    float* col = (float*)malloc(sizeof(float) * 4 * (size_t)W * H);
    float* dep = (float*)malloc(sizeof(float) * (size_t)W * H);
    if (!col || !dep) { free(col); free(dep); return 1; }
    for (long k = 0; k < (long)W * H; ++k) { for (int c = 0; c < 4; ++c) col[4 * k + c] = q8(clear4[c]); dep[k] = 1.f; }
    for (long i = 0; i < n; ++i) {                                   /* glDrawArrays(GL_POINTS, 0, n): buffer order */
        const float px = (float)(pm[4 * i] * (double)scale), py = (float)(pm[4 * i + 1] * (double)scale), pz = (float)(pm[4 * i + 2] * (double)scale);
        const float mass = (float)pm[4 * i + 3];
        const float vx = (float)vel[4 * i], vy = (float)vel[4 * i + 1], vz = (float)vel[4 * i + 2];
        float vp[4], cp[4];
        for (int r = 0; r < 4; ++r) vp[r] = fmaf(V[0 + r], px, fmaf(V[4 + r], py, fmaf(V[8 + r], pz, V[12 + r])));
        for (int r = 0; r < 4; ++r) cp[r] = fmaf(P[0 + r], vp[0], fmaf(P[4 + r], vp[1], fmaf(P[8 + r], vp[2], P[12 + r] * vp[3])));
        /* gl_PointSize: ShaderManager.cpp:233-282 */
        const float min_size = 2.f, max_size = 15.f;
        float size;
        if (mass <= 0.f) size = min_size * point_size / 1.f;
        else {
            const float l10 = logf(10.f);
            const float lm = logf(fmaxf(mass, 1e-20f)) / l10, lmin = logf(fmaxf(min_mass, 1e-20f)) / l10,
                        lmax = logf(fmaxf(max_mass, min_mass * 1.01f)) / l10;
            const float range = lmax - lmin;
            if (range < 1e-10f) size = (min_size + max_size) / 2.f * point_size / 30.f;
            else {
                const float np = clampf((lm - lmin) / range, 0.f, 1.f);
                const float base = min_size + np * (max_size - min_size);
                const float dcam = fmaxf(sqrtf(vp[0] * vp[0] + vp[1] * vp[1] + vp[2] * vp[2]), 0.1f);
                size = base * 8.f / dcam;
                size = size * (point_size / 30.f);
                size = fminf(size, point_size * 0.8f);
            }
        }
        const float cw = cp[3];
        if (!(cw > 0.f && fabsf(cp[0]) <= cw && fabsf(cp[1]) <= cw && fabsf(cp[2]) <= cw)) continue;   /* point clipped by its centre */
        const float xw = (cp[0] / cw * 0.5f + 0.5f) * (float)W, yw = (cp[1] / cw * 0.5f + 0.5f) * (float)H, zw = cp[2] / cw * 0.5f + 0.5f;
        size = clampf(size, 1.f, 1024.f);
        /* colour: ShaderManager.cpp:299-343,358-365 */
        float rgb[3] = {color4[0], color4[1], color4[2]};
        float nmass_c = 0.f;
        if (color_mode == 1) {
            const float speed = sqrtf(vx * vx + vy * vy + vz * vz);
            float ns = clampf(speed / fmaxf(max_velocity, 0.00001f), 0.f, 1.f);
            ns = smoothstepf(0.f, 1.f, ns);
            float c0[3], c1[3], t;
            if (ns < 0.2f)      { t = ns / 0.2f;          c0[0] = 0.f; c0[1] = 0.2f; c0[2] = 0.8f; c1[0] = 0.f;  c1[1] = 0.7f; c1[2] = 0.9f; }
            else if (ns < 0.5f) { t = (ns - 0.2f) / 0.3f; c0[0] = 0.f; c0[1] = 0.7f; c0[2] = 0.9f; c1[0] = 0.8f; c1[1] = 0.8f; c1[2] = 0.f; }
            else if (ns < 0.8f) { t = (ns - 0.5f) / 0.3f; c0[0] = 0.8f; c0[1] = 0.8f; c0[2] = 0.f; c1[0] = 1.f;  c1[1] = 0.5f; c1[2] = 0.f; }
            else                { t = (ns - 0.8f) / 0.2f; c0[0] = 1.f; c0[1] = 0.5f; c0[2] = 0.f; c1[0] = 1.f;  c1[1] = 0.f;  c1[2] = 0.f; }
            if (mass > 0.f) nmass_c = clampf((logf(mass) - logf(min_mass)) / (logf(max_mass) - logf(min_mass)), 0.f, 1.f);
            for (int k = 0; k < 3; ++k) rgb[k] = mixf(c0[k], c1[k], t) * (1.f + 0.8f * nmass_c);
        }
        float nmass = 0.f;                                                                /* :367-371 */
        if (mass > 0.f && max_mass > min_mass) nmass = clampf((logf(mass) - logf(min_mass)) / (logf(max_mass) - logf(min_mass)), 0.f, 1.f);
        const float h = 0.5f * size;
        int x0 = (int)ceilf(xw - h - 0.5f), x1 = (int)ceilf(xw + h - 0.5f), y0 = (int)ceilf(yw - h - 0.5f), y1 = (int)ceilf(yw + h - 0.5f);
        if (x0 < 0) x0 = 0; if (y0 < 0) y0 = 0; if (x1 > W) x1 = W; if (y1 > H) y1 = H;
        for (int iy = y0; iy < y1; ++iy)
            for (int ix = x0; ix < x1; ++ix) {
                const float s = 0.5f + ((float)ix + 0.5f - xw) / size, t = 0.5f - ((float)iy + 0.5f - yw) / size;   /* gl_PointCoord */
                const float dx = s - 0.5f, dy = t - 0.5f;
                const float dist = sqrtf(dx * dx + dy * dy) * 2.f;
                if (dist > 1.f) continue;                                                 /* discard */
                const float tex_a = dist <= 0.6f ? 1.f : (dist < 1.f ? smoothstepf(1.f, 0.6f, dist) : 0.f);
                const float edge_w = mixf(0.3f, 0.1f, nmass);
                float a = smoothstepf(1.f, 1.f - edge_w, dist) * tex_a;
                a = mixf(a, a * 1.3f, nmass);
                a = clampf(a, 0.f, 1.f);
                const long k = (long)iy * W + ix;
                if (!(zw <= dep[k])) continue;                                            /* GL_LEQUAL */
                dep[k] = zw;
                for (int c = 0; c < 3; ++c) col[4 * k + c] = q8(rgb[c] * a + col[4 * k + c] * (1.f - a));
                col[4 * k + 3] = q8(a * a + col[4 * k + 3] * (1.f - a));
            }
    }
    for (int iy = 0; iy < H; ++iy)
        for (int ix = 0; ix < W; ++ix)
            for (int c = 0; c < 4; ++c)
                rgba8[((size_t)(H - 1 - iy) * W + ix) * 4 + c] = (unsigned char)rintf(col[4 * ((long)iy * W + ix) + c] * 255.f);
    free(col); free(dep);
    return 0;
}
