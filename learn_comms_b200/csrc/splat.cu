// Headless picture of the crowd: what render::draw shows (enginee/src/render.cpp:364-415) — robots as
// discs coloured by comm_vis(c), the message -> colour maps of render.cpp:87-172, with their black outline, walls as
// squares, on the renderer's background — splatted by CUDA kernels into an RGBA8 buffer.  No graphics API is involved
// (there is none on this box): the orthographic top-down camera of render.cpp:352-355 / Camera.cpp:83-93 is two
// multiplications per pixel.  Agents of equal height z-fight in the reference; here the one later in the table wins.
#include "../../include/learn_comms.h"
#include "crowd.cuh"
#include <cuda_runtime.h>
#include <math.h>
#include <vector>

int lc_set_error(int code, const char* msg);  // capi.cu
struct lc_sim;
// capi.cu: device-side views of a handle's state for the renderer (engine-id order tables on the simulation's stream)
int lc_render_state(lc_sim* s, float2** pos, float** c3, int* n, const lc::WallRect** walls, int* n_walls, cudaStream_t* st);

namespace lcv {

// render.cpp:87-96 xyz2: message -> (L, A, B); glm::vec3::b aliases .z
__host__ __device__ inline void xyz2(const float c[3], float lab[3]) {
    const float pi = 3.14159265358979323846264338327950288f;
    const float rx = cosf(2 * pi / 3), ry = sinf(2 * pi / 3), gx = cosf(4 * pi / 3), gy = sinf(4 * pi / 3), bx = cosf(2 * pi), by = sinf(pi);
    lab[1] = rx * c[0] + gx * c[1] + bx * c[2];
    lab[2] = ry * c[0] + gy * c[1] + by * c[2];
    lab[0] = sqrtf(sqrtf(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]) / sqrtf(3.f));
}
// render.cpp:98-110 Luv2rgb
__host__ __device__ inline void luv2rgb(const float luv[3], float rgb[3]) {
    rgb[0] = +0.000f * luv[1] + 1.140f * luv[2] + luv[0];
    rgb[1] = -0.395f * luv[1] - 0.581f * luv[2] + luv[0];
    rgb[2] = +2.032f * luv[1] + 0.000f * luv[2] + luv[0];
    for (int k = 0; k < 3; ++k) rgb[k] = fmaxf(0.f, fminf(1.f, rgb[k]));
}
// render.cpp:114-155 Lab2rgb (easyrgb.com formulas, double intermediates as in the reference)
__host__ __device__ inline double maybe_cube(double a) {
    const double cube = a * a * a;
    return cube > 0.008856 ? cube : (a - 16. / 116.) / 7.787;
}
__host__ __device__ inline double finish_rgb(double a) {
    if (a > 0.0031308) a = 1.055 * pow(a, 1. / 2.4) - 0.055; else a *= 12.92;
    return pow(fmax(0., fmin(1., a)), 2.2);
}
__host__ __device__ inline void lab2rgb(const float lab_in[3], float rgb[3]) {
    const float L = lab_in[0] * 100.f, A = lab_in[1] * 128.f, B = lab_in[2] * 128.f;
    float X[3];
    X[0] = (float)((L + 16.) / 116.);
    X[1] = (float)(A / 500. + X[0]);
    X[2] = (float)(X[0] - B / 200.);
    X[0] = (float)(0.95047 * maybe_cube(X[0]));
    X[1] = (float)(1.00000 * maybe_cube(X[1]));
    X[2] = (float)(1.08883 * maybe_cube(X[2]));
    // glm::vec3 * glm::mat3 (row vector times matrix whose COLUMNS are the braces of render.cpp:134-138)
    const float m[3][3] = {{3.2406f, -1.5372f, -0.4986f}, {-0.9689f, 1.8758f, 0.0415f}, {0.0557f, -0.2040f, 1.0570f}};
    for (int k = 0; k < 3; ++k) rgb[k] = (float)finish_rgb((double)(X[0] * m[k][0] + X[1] * m[k][1] + X[2] * m[k][2]));
}
// render.cpp:30,300-309: VisF { XYZ2LAB, X2RG, Y2RG, Z2RG, XYZ2LUV }
__host__ __device__ inline void comm_vis(int vis_f, const float c[3], float rgb[3]) {
    if (vis_f >= 1 && vis_f <= 3) {  // render.cpp:157-170: one channel, red below zero, green above
        const float v = c[vis_f - 1];
        rgb[0] = v < 0 ? -v : 0.f; rgb[1] = v < 0 ? 0.f : v; rgb[2] = 0.f;
        return;
    }
    float lab[3];
    xyz2(c, lab);
    if (vis_f == 4) luv2rgb(lab, rgb); else lab2rgb(lab, rgb);
}
__host__ __device__ inline unsigned pack_rgba(const float rgb[3]) {
    unsigned out = 0xff000000u;
    for (int k = 0; k < 3; ++k) {
        const float v = fmaxf(0.f, fminf(1.f, rgb[k]));  // the framebuffer clamps (NaN -> 0)
        out |= (unsigned)(v * 255.f + 0.5f) << (8 * k);
    }
    return out;
}

struct View { int w, h; float size; };  // Camera.cpp:86-90: ortho window [-size, size]^2, size = 25 cam_dist / 2
__host__ __device__ inline float px_x(const View& v, int u) { return -v.size + ((float)u + 0.5f) * (2.f * v.size / (float)v.w); }
__host__ __device__ inline float px_z(const View& v, int r) { return -v.size + ((float)r + 0.5f) * (2.f * v.size / (float)v.h); }

constexpr float BODY_R = 0.5f * 0.75f * 0.3f;  // CylinderMesh radius .5 scaled by (1 - outline_percent) s, demo.cpp:540-541,559
constexpr float RING_R = 0.5f * 0.3f;          // the outline cylinder: scaled back by 1 / (1 - outline_percent), demo.cpp:598
constexpr unsigned WALL_RGBA = 0xff808080u;    // the wall texture (res/outline.png) is not part of this repository: mid grey
constexpr unsigned BACKGROUND = 0xff000000u | (227u << 16) | (246u << 8) | 253u;  // render.cpp:365

__global__ void k_splat_walls(View v, const lc::WallRect* walls, int n_walls, unsigned long long* key) {
    const int u = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y * blockDim.y + threadIdx.y;
    if (u >= v.w || r >= v.h) return;
    const float x = px_x(v, u), z = px_z(v, r);
    unsigned long long k = 0ull;
    for (int q = 0; q < n_walls; ++q)
        if (fabsf(x - walls[q].ox) <= walls[q].w / 2 && fabsf(z - walls[q].oy) <= walls[q].h / 2) k = (1ull << 32) | WALL_RGBA;
    key[(size_t)r * v.w + u] = k;
}
// one thread per agent (engine-id order): colour from its message, then the pixels of its disc; the larger key wins a pixel
__global__ void k_splat_agents(View v, const float2* pos, const float* c3, int n, int vis_f, unsigned long long* key) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float2 p = pos[i];
    if (!(fabsf(p.x) <= v.size + RING_R) || !(fabsf(p.y) <= v.size + RING_R)) return;
    const float c[3] = {c3[3 * i], c3[3 * i + 1], c3[3 * i + 2]};
    float rgb[3];
    comm_vis(vis_f, c, rgb);
    const unsigned body = pack_rgba(rgb);
    const float sx = (float)v.w / (2.f * v.size), sz = (float)v.h / (2.f * v.size);
    const int u0 = max((int)floorf((p.x - RING_R + v.size) * sx - 0.5f), 0), u1 = min((int)ceilf((p.x + RING_R + v.size) * sx - 0.5f), v.w - 1);
    const int r0 = max((int)floorf((p.y - RING_R + v.size) * sz - 0.5f), 0), r1 = min((int)ceilf((p.y + RING_R + v.size) * sz - 0.5f), v.h - 1);
    for (int r = r0; r <= r1; ++r)
        for (int u = u0; u <= u1; ++u) {
            const float dx = px_x(v, u) - p.x, dz = px_z(v, r) - p.y, d2 = dx * dx + dz * dz;
            if (d2 > RING_R * RING_R) continue;
            const unsigned rgba = d2 <= BODY_R * BODY_R ? body : 0xff000000u;
            atomicMax(&key[(size_t)r * v.w + u], ((unsigned long long)(i + 2) << 32) | rgba);
        }
}
__global__ void k_splat_resolve(View v, const unsigned long long* key, unsigned* rgba) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= v.w * v.h) return;
    const unsigned long long k = key[t];
    rgba[t] = k ? (unsigned)(k & 0xffffffffull) : BACKGROUND;
}

}  // namespace lcv

extern "C" int32_t lc_render(lc_sim* sim, int32_t vis_f, float cam_dist, int32_t width, int32_t height, uint8_t* rgba) {
    if (!sim || !rgba || width < 1 || height < 1 || width > 8192 || height > 8192 || !(cam_dist > 0.f) || vis_f < 0 || vis_f > 4)
        return lc_set_error(LC_ERR_ARG, "lc_render: bad argument");
    float2* pos = nullptr; float* c3 = nullptr; const lc::WallRect* walls = nullptr;
    int n = 0, n_walls = 0;
    cudaStream_t st = nullptr;
    int rc = lc_render_state(sim, &pos, &c3, &n, &walls, &n_walls, &st);
    if (rc) return rc;
    const lcv::View v{width, height, 25.f * cam_dist / 2.f};
    unsigned long long* key = nullptr;
    unsigned* out = nullptr;
    const size_t px = (size_t)width * height;
    if (cudaMalloc((void**)&key, px * 8) != cudaSuccess || cudaMalloc((void**)&out, px * 4) != cudaSuccess) {
        cudaFree(key);
        return lc_set_error(LC_ERR_CUDA, "lc_render: cudaMalloc failed");
    }
    lcv::k_splat_walls<<<dim3((width + 15) / 16, (height + 15) / 16), dim3(16, 16), 0, st>>>(v, walls, n_walls, key);
    lcv::k_splat_agents<<<(n + 127) / 128, 128, 0, st>>>(v, pos, c3, n, vis_f, key);
    lcv::k_splat_resolve<<<(unsigned)((px + 255) / 256), 256, 0, st>>>(v, key, out);
    cudaError_t e = cudaMemcpyAsync(rgba, out, px * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(key);
    cudaFree(out);
    if (e != cudaSuccess) return lc_set_error(LC_ERR_CUDA, cudaGetErrorString(e));
    return LC_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// lc_render_static: the same picture without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that
// forbid it).  The per-pixel winner keys lc_render keeps in a cudaMalloc'ed buffer, and its device-side RGBA staging, live
// in file-scope __device__ arrays of LC_RENDER_STATIC_MAX_PIXELS entries (12 MiB on every GPU that loads the library, held
// until exit).  Host code may not hand the arrays' NAMES to a kernel — on the host they are stand-ins, not device
// addresses — so their device addresses are looked up once per device with cudaGetSymbolAddress, with that device current.
// Every call shares the arrays: k_splat_walls rewrites all width x height keys of the call before the agents are splatted
// (nothing of the previous call survives in the part that is used), and calls that overlap on the GPU race.
namespace lcv {
static __device__ __align__(256) unsigned long long g_static_key[LC_RENDER_STATIC_MAX_PIXELS];
static __device__ __align__(256) unsigned g_static_rgba[LC_RENDER_STATIC_MAX_PIXELS];
constexpr int STATIC_MAX_DEVICES = 64;
struct StaticScratch { unsigned long long* key; unsigned* rgba; };
static StaticScratch g_static_ptr[STATIC_MAX_DEVICES];   // zero-initialised: not looked up yet
}  // namespace lcv

extern "C" int32_t lc_render_static(lc_sim* sim, int32_t vis_f, float cam_dist, int32_t width, int32_t height, uint8_t* rgba) {
    if (!sim || !rgba || width < 1 || height < 1 || !(cam_dist > 0.f) || vis_f < 0 || vis_f > 4)
        return lc_set_error(LC_ERR_ARG, "lc_render_static: bad argument");
    if ((long long)width * (long long)height > (long long)LC_RENDER_STATIC_MAX_PIXELS)
        return lc_set_error(LC_ERR_CAPACITY, "lc_render_static: width x height exceeds LC_RENDER_STATIC_MAX_PIXELS (1048576 pixels: the "
                                             "static scratch of this entry point); use lc_render for larger pictures");
    float2* pos = nullptr; float* c3 = nullptr; const lc::WallRect* walls = nullptr;
    int n = 0, n_walls = 0;
    cudaStream_t st = nullptr;
    int rc = lc_render_state(sim, &pos, &c3, &n, &walls, &n_walls, &st);   // makes the handle's device current; allocates nothing
    if (rc) return rc;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= lcv::STATIC_MAX_DEVICES)
        return lc_set_error(LC_ERR_CUDA, "lc_render_static: cannot identify the current device");
    lcv::StaticScratch& sc = lcv::g_static_ptr[dev];
    if (!sc.key) {   // once per device, with that device current (a handle is confined to one host thread at a time;
                     // two threads racing here on one device would both store the same two addresses)
        unsigned long long* k = nullptr;
        unsigned* o = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&k, lcv::g_static_key);
        if (e == cudaSuccess) e = cudaGetSymbolAddress((void**)&o, lcv::g_static_rgba);
        if (e != cudaSuccess || !k || !o) return lc_set_error(LC_ERR_CUDA, e != cudaSuccess ? cudaGetErrorString(e) : "lc_render_static: null symbol address");
        sc.rgba = o;
        sc.key = k;
    }
    const lcv::View v{width, height, 25.f * cam_dist / 2.f};
    const size_t px = (size_t)width * height;
    lcv::k_splat_walls<<<dim3((width + 15) / 16, (height + 15) / 16), dim3(16, 16), 0, st>>>(v, walls, n_walls, sc.key);   // rewrites every key in use
    lcv::k_splat_agents<<<(n + 127) / 128, 128, 0, st>>>(v, pos, c3, n, vis_f, sc.key);
    lcv::k_splat_resolve<<<(unsigned)((px + 255) / 256), 256, 0, st>>>(v, sc.key, sc.rgba);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(rgba, sc.rgba, px * 4, cudaMemcpyDefault, st);   // rgba: host or device memory
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return lc_set_error(LC_ERR_CUDA, cudaGetErrorString(e));
    return LC_OK;
}
