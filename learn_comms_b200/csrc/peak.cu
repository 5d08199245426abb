// FP32 CUDA-core throughput probe (bench.py's roofline.fp32 denominator).  The library may not contract a * b + c into
// one FFMA (parity with the reference's SSE2 build), so its own ceiling is one flop per issued FADD / FMUL / lane; the
// FFMA figure is the hardware's nominal two.  Independent chains, no memory traffic.
#include "../../include/learn_comms.h"
#include "lc_math.cuh"
#include <cuda_runtime.h>

int lc_set_error(int code, const char* msg);  // capi.cu

namespace {

constexpr int CHAINS = 8, ITERS = 4096;

// MODE 0: scalar FMUL, FADD (2 flop per chain step); 1: packed mul.rn.f32x2, add.rn.f32x2 as in the pair filter (4 flop);
// 2: FFMA, FFMA (4 flop)
template <int MODE>
__global__ void __launch_bounds__(256) k_fp32_peak(float* out, float a, float b) {
    float2 x[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) x[k] = make_float2((float)(threadIdx.x + k) * 1e-3f, (float)(threadIdx.x + 2 * k) * 1e-3f);
    const float2 a2 = make_float2(a, a), b2 = make_float2(b, b);
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < CHAINS; ++k) {
            if (MODE == 2) { x[k].x = __fmaf_rn(x[k].x, a, b); x[k].x = __fmaf_rn(x[k].x, a, b); }
            // (ptxas contracts a packed mul feeding a packed add into FFMA2 even under -fmad=false: even chains only
            // multiply, odd chains only add, so no sum ever consumes a product)
            else if (MODE == 1) { if (k & 1) { x[k] = lcm::add2(x[k], b2); x[k] = lcm::add2(x[k], a2); } else { x[k] = lcm::mul2(x[k], a2); x[k] = lcm::mul2(x[k], b2); } }
            else { x[k].x = __fmul_rn(x[k].x, a); x[k].x = __fadd_rn(x[k].x, b); }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < CHAINS; ++k) s += x[k].x + x[k].y;
    if (s == 123.456f) out[0] = s;  // never true: keeps the chains alive
}

}  // namespace

extern "C" int32_t lc_fp32_peak(int32_t device, int32_t repeats, float* tflops /*[3]: scalar mul+add, packed f32x2 mul+add, FFMA*/) {
    if (cudaSetDevice(device) != cudaSuccess) return lc_set_error(LC_ERR_CUDA, "lc_fp32_peak: no such CUDA device");
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return lc_set_error(LC_ERR_CUDA, "lc_fp32_peak: cudaGetDeviceProperties failed");
    float* d = nullptr;
    if (cudaMalloc(&d, 4) != cudaSuccess) return lc_set_error(LC_ERR_CUDA, "lc_fp32_peak: cudaMalloc failed");
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = prop.multiProcessorCount * 8, threads = 256;
    float best[3] = {0.f, 0.f, 0.f};
    for (int v = 0; v < 3; ++v)
        for (int r = 0; r < (repeats > 0 ? repeats : 5) + 1; ++r) {
            cudaEventRecord(e0);
            if (v == 0) k_fp32_peak<0><<<blocks, threads>>>(d, 0.999f, 1e-3f);
            else if (v == 1) k_fp32_peak<1><<<blocks, threads>>>(d, 0.999f, 1e-3f);
            else k_fp32_peak<2><<<blocks, threads>>>(d, 0.999f, 1e-3f);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            // per chain and iteration: mul + add = 2 flop without contraction, 2 FFMA = 4 flop with
            const double flop = (double)blocks * threads * ITERS * CHAINS * (v == 0 ? 2.0 : 4.0);
            const float tf = ms > 0.f ? (float)(flop / (ms * 1e-3) / 1e12) : 0.f;
            if (r > 0 && tf > best[v]) best[v] = tf;  // the first launch warms up
        }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    if (cudaGetLastError() != cudaSuccess) return lc_set_error(LC_ERR_CUDA, "lc_fp32_peak: kernel failed");
    if (tflops) { tflops[0] = best[0]; tflops[1] = best[1]; tflops[2] = best[2]; }
    return LC_OK;
}
