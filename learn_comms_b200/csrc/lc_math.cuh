// Device math shared by the crowd and batch kernels.  Everything here is IEEE binary32 with round-to-nearest
// and NO fused multiply-add (the reference CPU build has no -march, so it runs plain SSE2 mul+add;
// enginee/CMakeLists.txt:68-80).  The library is compiled with -fmad=false -prec-div=true -prec-sqrt=true
// -ftz=false; the explicit _rn intrinsics below document the intent at the places where contraction would
// change a neighbour-list decision.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace lcm {

#define LC_FLT_MAX 3.402823466e+38f
#define LC_FLT_MIN 1.175494351e-38f
#define LC_DT (1.f / 60.f)

__device__ __forceinline__ float dot2(float ax, float ay, float bx, float by) {
    return __fadd_rn(__fmul_rn(ax, bx), __fmul_rn(ay, by));
}
// glm::dot(vec3, vec3) with both y components zero: (x*x' + 0*0) + z*z'
__device__ __forceinline__ float dot3_xz(float ax, float az, float bx, float bz) {
    return __fadd_rn(__fadd_rn(__fmul_rn(ax, bx), 0.0f), __fmul_rn(az, bz));
}
__device__ __forceinline__ float len2(float x, float y) { return __fsqrt_rn(dot2(x, y, x, y)); }
__device__ __forceinline__ float glm_min(float a, float b) { return (b < a) ? b : a; }
__device__ __forceinline__ float glm_max(float a, float b) { return (a < b) ? b : a; }
__device__ __forceinline__ float glm_clamp(float x, float lo, float hi) { return glm_min(glm_max(x, lo), hi); }

// minstd_rand0 (Seeder.cpp:7; LMP.cpp:261-263): x <- a*x mod (2^31 - 1)
__host__ __device__ __forceinline__ uint32_t lcg_mul(uint32_t x, uint32_t a) {
    uint64_t p = (uint64_t)x * a;
    p = (p & 0x7fffffffull) + (p >> 31);
    p = (p & 0x7fffffffull) + (p >> 31);
    if (p >= 0x7fffffffull) p -= 0x7fffffffull;
    return (uint32_t)p;
}
// libstdc++ generate_canonical<float,24> over minstd_rand0 + uniform_real_distribution<float>(a, b)
__device__ __forceinline__ float lcg_uniform(uint32_t x, float a, float b) {
    float sum = __uint2float_rn(x - 1u);
    float ret = __fdiv_rn(sum, 2147483648.0f);
    if (ret >= 1.0f) ret = __uint_as_float(0x3f7fffffu);
    return __fadd_rn(__fmul_rn(ret, __fsub_rn(b, a)), a);
}

// Packed binary32 pairs (sm_100 FADD2 / FMUL2): two IEEE round-to-nearest operations per issued instruction, each lane
// bit-identical to the scalar __fadd_rn / __fmul_rn.  No packed FMA is used anywhere (parity needs mul and add rounded
// separately).  The pair loop is issue-bound, so halving its FP instruction count is what this buys.
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
    float2 r;
    asm("{.reg .b64 ra, rb, rc; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; add.rn.f32x2 rc, ra, rb; mov.b64 {%0,%1}, rc;}"
        : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return r;
}
__device__ __forceinline__ float2 sub2(float2 a, float2 b) {
    float2 r;
    asm("{.reg .b64 ra, rb, rc; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; sub.rn.f32x2 rc, ra, rb; mov.b64 {%0,%1}, rc;}"
        : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return r;
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
    float2 r;
    asm("{.reg .b64 ra, rb, rc; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; mul.rn.f32x2 rc, ra, rb; mov.b64 {%0,%1}, rc;}"
        : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return r;
}

// Feature-set variants of comm::x (comm.cpp:22-28: the #defines COSINE_REL, PROX, SOFTSIGN_GOAL; HEAD builds with all
// three).  Bits of lc_desc::flags / lc_batch_desc::flags (learn_comms.h LC_FLAG_FEAT_*).
#define LCM_FEAT_PLAIN 32u        /* neither COSINE_REL nor NORM_REL: x = [c_j, rel_v.x, rel_v.y, rel_p.x, rel_p.y, g] (comm.cpp:189) */
#define LCM_FEAT_NO_PROX 64u      /* PROX undefined: mag_rp stays |rel_p| (comm.cpp:172-174) */
#define LCM_FEAT_NO_SOFTSIGN 128u /* SOFTSIGN_GOAL undefined: g = goal_dist (comm.cpp:178-180) */

// comm.cpp:144-192 x(): the 8 inputs of M_cv from the partner's message c_j, the relative velocity dv = v_j - v_i and
// position dp = p_j - p_i expressed in the agent's (right, facing) frame, and its goal distance.
__device__ __forceinline__ void comm_features(uint32_t flags, float rightx, float righty, float fx, float fy, float dvx, float dvy,
                                              float dpx, float dpy, float goal_dist, float cjx, float cjy, float cjz, float x[8]) {
    float rvx = dot2(rightx, righty, dvx, dvy), rvy = dot2(fx, fy, dvx, dvy);
    float rpx = dot2(rightx, righty, dpx, dpy), rpy = dot2(fx, fy, dpx, dpy);
    float g = goal_dist;
    if (!(flags & LCM_FEAT_NO_SOFTSIGN)) g = __fdiv_rn(g, __fadd_rn(1.f, g));
    x[0] = cjx; x[1] = cjy; x[2] = cjz;
    if (flags & LCM_FEAT_PLAIN) {
        x[3] = rvx; x[4] = rvy; x[5] = rpx; x[6] = rpy; x[7] = g;
        return;
    }
    const float mag_rv = len2(rvx, rvy);
    if (mag_rv > 0.f) { rvx = __fdiv_rn(rvx, mag_rv); rvy = __fdiv_rn(rvy, mag_rv); }
    float mag_rp = len2(rpx, rpy);
    if (mag_rp > 0.f) {
        rpx = __fdiv_rn(rpx, mag_rp); rpy = __fdiv_rn(rpy, mag_rp);
        if (!(flags & LCM_FEAT_NO_PROX)) mag_rp = __fdiv_rn(1.f, mag_rp);
    }
    x[3] = rvy; x[4] = mag_rv; x[5] = rpy; x[6] = mag_rp; x[7] = g;
}

// order-preserving float <-> uint encoding for atomicMin/atomicMax reductions
__device__ __forceinline__ uint32_t f2ord(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__device__ __forceinline__ float ord2f(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// LMP.cpp:13-55 ttc / ttc_(Circ,Circ) -> BoundVolume.cpp:102-122 Circ::intersect in the relative frame.
// Returns true and t when the pair interacts (ttc <= 5, LMP.cpp:297-300).  c_out = |m|^2 - R^2.
__device__ __forceinline__ bool ttc_pair(float mx, float my, float dvx, float dvy, float RR, float& t, float& c_out) {
    float a = dot2(dvx, dvy, dvx, dvy);
    float b = dot2(mx, my, dvx, dvy);
    float c = __fsub_rn(dot2(mx, my, mx, my), RR);
    c_out = c;
    if (c > 0.f && b > 0.f) return false;
    float disc = __fsub_rn(__fmul_rn(b, b), __fmul_rn(a, c));
    if (disc < 0.f) return false;
    float tt = __fdiv_rn(__fsub_rn(-b, __fsqrt_rn(disc)), a);
    if (tt < 0.f) tt = 0.f;
    if (!(tt >= 0.f)) return false;   // NaN -> FLT_MAX (LMP.cpp:31-34)
    if (tt > 5.0f) return false;      // (double)ttc > 5.0
    t = tt;
    return true;
}

// LMP.cpp:138-158 ttc_forces + :115-127 ttc_forces_ (f64 magnitude, quirk Q11)
__device__ __forceinline__ void ttc_force(float pix, float piy, float vix, float viy, float pjx, float pjy,
                                          float vjx, float vjy, float t, float& fx, float& fy) {
    float dx = __fsub_rn(__fadd_rn(pix, __fmul_rn(vix, t)), __fadd_rn(pjx, __fmul_rn(vjx, t)));
    float dy = __fsub_rn(__fadd_rn(piy, __fmul_rn(viy, t)), __fadd_rn(pjy, __fmul_rn(vjy, t)));
    float len = dot2(dx, dy, dx, dy);
    if (len > 0.f) {
        float s = __fsqrt_rn(len);
        dx = __fdiv_rn(dx, s);
        dy = __fdiv_rn(dy, s);
    }
    double td = (double)t;
    double mag = __ddiv_rn(__dsub_rn(5.0, td), __dadd_rn(td, 0.001));
    mag = mag > 20.0 ? 20.0 : mag;
    fx = __double2float_rn(__dmul_rn(mag, (double)dx));
    fy = __double2float_rn(__dmul_rn(mag, (double)dy));
}

// BVH.cpp:106-120 leaf predicate (BoundVolume.cpp:203-207) and push-out vector (BoundVolume.cpp:209-211)
__device__ __forceinline__ bool circ_circ_hit(float qx, float qy, float bx, float by, float r, float RR,
                                              float& ax, float& ay) {
    float dx = __fsub_rn(qx, bx), dy = __fsub_rn(qy, by);
    if (!(dot2(dx, dy, dx, dy) < RR)) return false;
    float ux = __fsub_rn(bx, qx), uy = __fsub_rn(by, qy);
    float inv = __fdiv_rn(1.0f, __fsqrt_rn(dot2(ux, uy, ux, uy)));
    float onqx = __fadd_rn(__fmul_rn(__fmul_rn(ux, inv), r), qx);
    float onqy = __fadd_rn(__fmul_rn(__fmul_rn(uy, inv), r), qy);
    float wx = __fsub_rn(onqx, bx), wy = __fsub_rn(onqy, by);
    float inv2 = __fdiv_rn(1.0f, __fsqrt_rn(dot2(wx, wy, wx, wy)));
    float onbx = __fadd_rn(__fmul_rn(__fmul_rn(wx, inv2), r), bx);
    float onby = __fadd_rn(__fmul_rn(__fmul_rn(wy, inv2), r), by);
    ax = __fsub_rn(onbx, onqx);
    ay = __fsub_rn(onby, onqy);
    return true;
}

// physics.cpp:23-36 push_away (acc is dead after this point in the frame)
__device__ __forceinline__ void push_away(float& px, float& py, float& hx, float& hy, float ax, float ay) {
    float l2 = dot3_xz(ax, ay, ax, ay);
    if (l2 > 0.f) {
        float inv = __fdiv_rn(1.0f, __fsqrt_rn(l2));
        px = __fadd_rn(px, __fadd_rn(ax, __fmul_rn(0.001f, __fmul_rn(ax, inv))));
        py = __fadd_rn(py, __fadd_rn(ay, __fmul_rn(0.001f, __fmul_rn(ay, inv))));
    }
    float d = dot3_xz(ax, ay, hx, hy);
    hx = __fadd_rn(hx, __fmul_rn(ax, d));
    hy = __fadd_rn(hy, __fmul_rn(ay, d));
}

}  // namespace lcm
