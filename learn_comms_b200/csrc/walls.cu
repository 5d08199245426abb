// Static-wall geometry on the device (walls.cuh) behind the C ABI.  lc_wall_geometry is the parity entry point for the
// geometry of SURVEY §8 rows a10 / a14 / a16: it evaluates every wall function on caller-supplied records so the GPU
// suite can hold the device code against the reference's own bit patterns (tests/golden/wall_geometry.npz) before the
// functions are wired into k_force / k_collide for walled scenes.
#include "../../include/learn_comms.h"
#include "walls.cuh"
#include <cuda_runtime.h>
#include <cstdio>

int lc_set_error(int code, const char* msg);  // capi.cu

namespace {

// one record per thread: 36 B in, 48 B out; the records are independent, so this is a plain grid-stride map
__global__ void __launch_bounds__(256) k_wall_geometry(int32_t n, const float* __restrict__ rec, uint32_t* __restrict__ out) {
    for (int32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        float p[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) p[i] = rec[(size_t)k * 9 + i];
        uint32_t w[12];
        lcw::geometry_record(p, w);
        uint4* o = reinterpret_cast<uint4*>(out + (size_t)k * 12);  // 48 B per record: three 16 B stores
        o[0] = make_uint4(w[0], w[1], w[2], w[3]);
        o[1] = make_uint4(w[4], w[5], w[6], w[7]);
        o[2] = make_uint4(w[8], w[9], w[10], w[11]);
    }
}

// LMP::lookahead for one agent per thread; the wall list is small (<= 60 tiles in the shipped scenes) and read by every
// thread, so it is staged in shared memory once per block
__global__ void __launch_bounds__(256) k_lookahead(int32_t n_walls, const lc_rect* __restrict__ walls, float agent_r, int32_t n,
                                                   const lc_look_record* __restrict__ rec, lc_look_result* __restrict__ out) {
    extern __shared__ lcw::Wall s_walls[];
    for (int k = threadIdx.x; k < n_walls; k += blockDim.x) s_walls[k] = {walls[k].ox, walls[k].oy, walls[k].w, walls[k].h};
    __syncthreads();
    for (int32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        lc_look_record r = rec[k];
        int nd = r.num_done;
        float lgx = r.local_goal[0], lgy = r.local_goal[1];
        lcw::lookahead(s_walls, n_walls, agent_r, r.pos[0], r.pos[1], &r.plan[0][0], r.n_plan, nd, lgx, lgy);
        out[k].num_done = nd;
        out[k].local_goal[0] = lgx;
        out[k].local_goal[1] = lgy;
    }
}

int cuda_fail(const char* what, cudaError_t e) {
    char buf[256];
    snprintf(buf, sizeof buf, "walls: %s failed: %s", what, cudaGetErrorString(e));
    (void)cudaGetLastError();  // a refused call must not show up as the next launch's error
    return lc_set_error(LC_ERR_CUDA, buf);
}

}  // namespace

extern "C" int32_t lc_wall_geometry(int32_t n_records, const float* records, uint32_t* words, int32_t device) {
    if (n_records < 0 || (n_records > 0 && (!records || !words)))
        return lc_set_error(LC_ERR_ARG, "lc_wall_geometry: bad arguments");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return cuda_fail("cudaSetDevice", e);
    if (n_records == 0) return LC_OK;
    float* d_rec = nullptr;
    uint32_t* d_out = nullptr;
    size_t in_bytes = sizeof(float) * 9 * (size_t)n_records, out_bytes = sizeof(uint32_t) * 12 * (size_t)n_records;
    if ((e = cudaMalloc((void**)&d_rec, in_bytes)) != cudaSuccess) return cuda_fail("cudaMalloc", e);
    if ((e = cudaMalloc((void**)&d_out, out_bytes)) != cudaSuccess) {
        cudaFree(d_rec);
        return cuda_fail("cudaMalloc", e);
    }
    int rc = LC_OK;
    if ((e = cudaMemcpy(d_rec, records, in_bytes, cudaMemcpyHostToDevice)) != cudaSuccess) rc = cuda_fail("cudaMemcpy", e);
    if (rc == LC_OK) {
        int blocks = (n_records + 255) / 256;
        if (blocks > 148 * 8) blocks = 148 * 8;  // 148 SMs x 8 resident blocks of 256 threads
        k_wall_geometry<<<blocks, 256>>>(n_records, d_rec, d_out);
        if ((e = cudaGetLastError()) != cudaSuccess) rc = cuda_fail("k_wall_geometry launch", e);
    }
    if (rc == LC_OK && (e = cudaMemcpy(words, d_out, out_bytes, cudaMemcpyDeviceToHost)) != cudaSuccess)
        rc = cuda_fail("cudaMemcpy", e);
    cudaFree(d_rec);
    cudaFree(d_out);
    return rc;
}

extern "C" int32_t lc_lookahead(int32_t n_walls, const lc_rect* walls, float agent_radius, int32_t n_records,
                                const lc_look_record* records, lc_look_result* results, int32_t device) {
    if (n_walls < 0 || n_walls > 2048 || (n_walls > 0 && !walls) || n_records < 0 || (n_records > 0 && (!records || !results)))
        return lc_set_error(LC_ERR_ARG, "lc_lookahead: bad arguments");
    for (int32_t k = 0; k < n_records; ++k)
        if (records[k].n_plan < 1 || records[k].n_plan > LC_LOOK_MAX_WAYPOINTS || records[k].num_done < 0)
            return lc_set_error(LC_ERR_ARG, "lc_lookahead: n_plan must be 1..8 and num_done >= 0");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return cuda_fail("cudaSetDevice", e);
    if (n_records == 0) return LC_OK;
    lc_rect* d_walls = nullptr;
    lc_look_record* d_rec = nullptr;
    lc_look_result* d_out = nullptr;
    int rc = LC_OK;
    if ((e = cudaMalloc((void**)&d_walls, sizeof(lc_rect) * (size_t)(n_walls ? n_walls : 1))) != cudaSuccess ||
        (e = cudaMalloc((void**)&d_rec, sizeof(lc_look_record) * (size_t)n_records)) != cudaSuccess ||
        (e = cudaMalloc((void**)&d_out, sizeof(lc_look_result) * (size_t)n_records)) != cudaSuccess)
        rc = cuda_fail("cudaMalloc", e);
    if (rc == LC_OK && n_walls &&
        (e = cudaMemcpy(d_walls, walls, sizeof(lc_rect) * (size_t)n_walls, cudaMemcpyHostToDevice)) != cudaSuccess)
        rc = cuda_fail("cudaMemcpy", e);
    if (rc == LC_OK &&
        (e = cudaMemcpy(d_rec, records, sizeof(lc_look_record) * (size_t)n_records, cudaMemcpyHostToDevice)) != cudaSuccess)
        rc = cuda_fail("cudaMemcpy", e);
    if (rc == LC_OK) {
        int blocks = (n_records + 255) / 256;
        if (blocks > 148 * 8) blocks = 148 * 8;
        k_lookahead<<<blocks, 256, sizeof(lcw::Wall) * (size_t)n_walls>>>(n_walls, d_walls, agent_radius, n_records, d_rec, d_out);
        if ((e = cudaGetLastError()) != cudaSuccess) rc = cuda_fail("k_lookahead launch", e);
    }
    if (rc == LC_OK &&
        (e = cudaMemcpy(results, d_out, sizeof(lc_look_result) * (size_t)n_records, cudaMemcpyDeviceToHost)) != cudaSuccess)
        rc = cuda_fail("cudaMemcpy", e);
    cudaFree(d_walls);
    cudaFree(d_rec);
    cudaFree(d_out);
    return rc;
}
