// Batched rollouts for sm_100a: thousands of independent small scenarios / PSO particles in ONE launch.
// Replaces pso_opt.py:236-267 (one comm_norender process per (particle, scene) + a file round trip per evaluation).
//
// One thread block per scenario, one thread per agent (n <= 256).  The whole rollout — every frame of
// main.cpp:98-150 until demo::run reports all_done — runs inside the block: the neighbour-visible state
// (pos, vel, c) lives in shared memory (double-buffered, 80 B/agent), everything else in registers; global
// memory is touched once at the start and once at the end.  Two __syncthreads per frame.  Blocks retire as their
// scenario finishes (600-2700 frames), so the 148 SMs stay full until the tail.
//
// Arithmetic is the same IEEE binary32 operation sequence as the single-crowd kernels (crowd.cu k_force /
// k_collide) and the oracle's brute-force mode: all pairs in ascending id order (LMP.cpp:288-304), so the force
// sum, the comm-partner arg-max (comm.cpp:201-220) and the push-out order need no sorting.
#include "../../include/learn_comms.h"
#include "lc_math.cuh"
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <math_constants.h>
#include <string>
#include <vector>

namespace lcbatch {
using namespace lcm;

constexpr int MAX_N = 256;
constexpr int MAX_WARPS = MAX_N / 32;
constexpr int MAX_NEAR = 8;

struct ScenarioCtl {
    uint32_t frame;
    int finished;
    int total_num_done;
    float run_avg, M2, run_std, avg_velocity;
    int error;
    unsigned long long overlap2;   // ordered overlapping pairs at frame start, summed over all frames
    unsigned long long coll_hits;  // ordered (q, bv) hits of the collision pass, summed over all frames
    unsigned long long pair_tests; // TTC pair evaluations (throughput accounting)
};

struct Arrays {
    int S, n;
    float4* pv;      // [S][n] pos.x pos.z vel.x vel.z          Dynamics.h:4,6
    float2* vfh;     // vel_forehalf                              Dynamics.h:7
    float2* goal;    // final_goal                                Agent.h:11
    float2* lgoal;   // local_goal                                Agent.h:12
    float2* facing;  // CommComp::facing                          CommComp.h:17
    float4* c4;      // c[0..2], goal_dist                        CommComp.h:14, Agent.h:19
    int4* meta;      // flags (bit0 num_done, bit1 overhead_counted), private LCG state, min_time bits, partner (diag)
    int* nbr_cnt;    // [S][n] |{j : ttc_ij <= 5}| of the last executed frame (diag)
    float* m_cv;     // [S][24]
    float* m_f;      // [S][6]
    float* max_duration;  // [S]
    ScenarioCtl* ctl;     // [S]
};

struct Consts {
    float r, RR, near_c;
    uint32_t lcg_stride;  // 16807^n: one frame of the scenario's global stream
    uint32_t flags;
};

// demo.cpp:655-663 update_runs (Welford, all float)
__device__ __forceinline__ void update_runs(ScenarioCtl& c, float overhead) {
    float ftotal = (float)(++c.total_num_done);
    float delta = __fsub_rn(overhead, c.run_avg);
    c.run_avg = __fadd_rn(c.run_avg, __fdiv_rn(delta, ftotal));
    float delta2 = __fsub_rn(overhead, c.run_avg);
    c.M2 = __fadd_rn(c.M2, __fmul_rn(delta, delta2));
    c.run_std = __fsqrt_rn(__fdiv_rn(c.M2, ftotal));
}

// physics.cpp:72-76 with vel_count == 0 forever (quirk Q6): the running "average" is the speed of the last not-done
// agent in table order unless a non-finite value poisoned the chain (NaN * 0).
__device__ __forceinline__ void fold_avg_velocity(ScenarioCtl& c, const unsigned* ndm, const float* spd, const unsigned* nfm,
                                                  int nwarps) {
    int top = -1, nonfinite = 0;
    for (int w = 0; w < nwarps; ++w) {
        if (ndm[w]) top = w;
        nonfinite += __popc(nfm[w]);
    }
    if (top < 0) return;
    const float sp = spd[top];
    const int others = nonfinite - ((sp <= LC_FLT_MAX) ? 0 : 1);
    const bool poisoned = others > 0 || !(fabsf(c.avg_velocity) <= LC_FLT_MAX);
    c.avg_velocity = poisoned ? CUDART_NAN_F : sp;
}

__global__ void __launch_bounds__(MAX_N) k_batch_rollout(Arrays A, Consts K, int max_frames) {
    extern __shared__ float4 s_dyn[];
    const int n = A.n;
    const int nt = blockDim.x;  // n rounded up to a whole warp
    float4* s_pv0 = s_dyn;            // frame-start (pos, vel), what neighbours read in the pair loop
    float4* s_pv1 = s_dyn + nt;       // integrated, not yet pushed (pos, vel): what the collision pass reads
    float4* s_c0 = s_dyn + 2 * nt;    // c, double-buffered by frame parity
    float4* s_c1 = s_dyn + 3 * nt;
    float4* s_pq = s_dyn + 4 * nt;    // pair-interleaved copy of s_pv0: (x0,x1,z0,z1),(vx0,vx1,vz0,vz1) per agent pair
    int* s_near = reinterpret_cast<int*>(s_dyn + 5 * nt);  // [MAX_NEAR][nt]
    __shared__ float s_w[32];                 // M_cv (24) | M_f (6)
    __shared__ float s_mt[MAX_N];             // min_time (constant per agent, Agent.h:17)
    __shared__ float s_ft[MAX_N];             // timeout finish times (demo.cpp:693-697)
    __shared__ unsigned s_mask[2][3][MAX_WARPS];  // per-warp ballots by frame parity: newly done, not done, needs replan
    __shared__ unsigned s_ndm[MAX_WARPS], s_nfm[MAX_WARPS];
    __shared__ float s_spd[MAX_WARPS];
    __shared__ ScenarioCtl s_ctl;

    const int sc = blockIdx.x;
    const int i = threadIdx.x, lane = i & 31, wid = i >> 5, nwarps = nt >> 5;
    const bool valid = i < n;
    const size_t g = (size_t)sc * n + (valid ? i : 0);

    // ---- load: global -> registers / shared ----
    float4 me = A.pv[g];
    float2 vfh = A.vfh[g];
    const float2 goal = A.goal[g];
    float2 lgoal = A.lgoal[g];
    float2 facing = A.facing[g];
    float4 c4 = A.c4[g];
    int4 mt = A.meta[g];
    int last_cnt = 0;
    float* const my_pq = reinterpret_cast<float*>(s_pq + 2 * (i >> 1)) + (i & 1);
    if (valid) {
        s_pv0[i] = me;
        my_pq[0] = me.x; my_pq[2] = me.y; my_pq[4] = me.z; my_pq[6] = me.w;
        s_c0[i] = c4;
        s_mt[i] = __int_as_float(mt.z);
    } else {
        my_pq[0] = my_pq[2] = my_pq[4] = my_pq[6] = 0.f;  // padding of an odd last pair: finite, never selected
    }
    if (i < 24) s_w[i] = A.m_cv[(size_t)sc * 24 + i];
    if (i >= 24 && i < 30) s_w[i] = A.m_f[(size_t)sc * 6 + (i - 24)];
    if (i == 0) s_ctl = A.ctl[sc];
    if (i < MAX_WARPS) { s_ndm[i] = 0u; s_nfm[i] = 0u; s_spd[i] = 0.f; }
    __syncthreads();
    const float max_duration = A.max_duration[sc];
    const uint32_t frame0 = s_ctl.frame;
    bool finished = s_ctl.finished != 0;
    const bool comm = !(K.flags & LC_FLAG_NO_COMM);
    unsigned long long overlap2 = 0ull, coll_hits = 0ull, pair_tests = 0ull;
    int err = 0;
    int executed = 0;

    for (int it = 0; it < max_frames && !finished; ++it) {
        const uint32_t frame = frame0 + (uint32_t)it + 1u;
        const int par = it & 1;
        const float4* __restrict__ pvA = s_pv0;
        const float4* cA = par ? s_c1 : s_c0;
        float4* cB = par ? s_c0 : s_c1;
        const float px = me.x, py = me.y, vx = me.z, vy = me.w;
        const float fake_time = __fmul_rn(LC_DT, (float)frame);
        const bool timeout = fake_time > max_duration;

        // ---- S0: demo::run (demo.cpp:665-702), block-wide facts through per-warp ballots ----
        const bool done_before = mt.x & 1;
        const float gd = len2(__fsub_rn(goal.x, px), __fsub_rn(goal.y, py));  // ai.cpp:154
        const bool f_new = valid && done_before && !(mt.x & 2);
        const bool f_nd = valid && !done_before;
        const bool f_rp = valid && done_before && gd > 1.f;                   // ai.cpp:161-163 (wall-free)
        const unsigned m_new = __ballot_sync(0xffffffffu, f_new);
        const unsigned m_nd = __ballot_sync(0xffffffffu, f_nd);
        const unsigned m_rp = __ballot_sync(0xffffffffu, f_rp);
        if (lane == 0) { s_mask[par][0][wid] = m_new; s_mask[par][1][wid] = m_nd; s_mask[par][2][wid] = m_rp; }
        if (timeout && f_nd)  // demo.cpp:693-697
            s_ft[i] = __fsub_rn(__fadd_rn(__fmul_rn(2.f, gd), fake_time), __int_as_float(mt.z));
        if (f_new) mt.x |= 2;
        __syncthreads();  // B1: ballots + last frame's final positions (s_pv0) + last frame's c (cA) are visible
        int nnd = 0, rp_below = 0;
        for (int w = 0; w < nwarps; ++w) {
            nnd += __popc(s_mask[par][1][w]);
            if (w < wid) rp_below += __popc(s_mask[par][2][w]);
        }
        rp_below += __popc(m_rp & ((1u << lane) - 1u));
        const bool all_done = nnd == 0 || timeout;
        if (i == 0) {
            // avg_velocity of the previous frame, then this frame's finishers in table order
            fold_avg_velocity(s_ctl, s_ndm, s_spd, s_nfm, nwarps);
            for (int w = 0; w < nwarps; ++w)
                for (unsigned m = s_mask[par][0][w]; m; m &= m - 1u)
                    update_runs(s_ctl, __fsub_rn(fake_time, s_mt[32 * w + __ffs(m) - 1]));
            if (timeout)
                for (int w = 0; w < nwarps; ++w)
                    for (unsigned m = s_mask[par][1][w]; m; m &= m - 1u) update_runs(s_ctl, s_ft[32 * w + __ffs(m) - 1]);
            s_ctl.frame = frame;
            if (all_done) s_ctl.finished = 1;
        }

        float npx = px, npy = py, nvx = vx, nvy = vy;
        int n_near = 0;
        bool not_done_after = false;
        if (valid) {
            // ---- all pairs, ascending id: comm partner arg-max (comm.cpp:201-220, quirk Q1), TTC forces
            //      (LMP.cpp:288-304), near list of the collision pass ----
            float best = LC_FLT_MIN;
            int bj = -1;
            float ttcFx = 0.f, ttcFy = 0.f;
            int cnt = 0, n_overlap = 0;
            // Phase 2 (rare): one pair that survived the packed filter, exact scalar arithmetic in the oracle's order
            // (LMP.cpp:37-55 / BoundVolume.cpp:102-122 in the relative frame), near list, force on a hit.
            auto slow = [&](int j) {
                const float4 q = pvA[j];
                const float mx = __fsub_rn(q.x, px), my = __fsub_rn(q.y, py);
                const float dvx = __fsub_rn(q.z, vx), dvy = __fsub_rn(q.w, vy);
                const float b = dot2(mx, my, dvx, dvy);
                const float c = __fsub_rn(dot2(mx, my, mx, my), K.RR);
                if (c < K.near_c) {
                    if (n_near < MAX_NEAR) s_near[n_near * nt + i] = j;
                    ++n_near;
                    if (c < 0.f) ++n_overlap;  // BoundVolume.cpp:203-207 on frame-start positions
                }
                if (c > 0.f && b > 0.f) return;
                const float a = dot2(dvx, dvy, dvx, dvy);
                const float disc = __fsub_rn(__fmul_rn(b, b), __fmul_rn(a, c));
                if (disc < 0.f) return;
                float t = __fdiv_rn(__fsub_rn(-b, __fsqrt_rn(disc)), a);
                if (t < 0.f) t = 0.f;
                if (!(t >= 0.f)) return;  // NaN -> FLT_MAX (LMP.cpp:31-34)
                if (t > 5.0f) return;     // (double)ttc > 5.0 (LMP.cpp:297-300)
                float fx, fy;
                ttc_force(px, py, vx, vy, q.x, q.y, q.z, q.w, t, fx, fy);
                ttcFx = __fadd_rn(ttcFx, fx);
                ttcFy = __fadd_rn(ttcFy, fy);
                ++cnt;
            };
            // Phase 1: every lane walks the same j = 0..n-1 (uniform trip count, no divergence), two candidates per
            // iteration from the pair-interleaved copy with packed FADD2 / FMUL2 (sums that consume products stay scalar:
            // ptxas would contract them into FFMA2, see crowd.cu).  Survivors of the conservative filter are collected in
            // a 32-bit mask per chunk and revisited in ascending j, so the force is still summed in table order.
            const float2 px2 = make_float2(px, px), py2 = make_float2(py, py), vx2 = make_float2(vx, vx), vy2 = make_float2(vy, vy);
            const float2 RR2 = make_float2(K.RR, K.RR);
            for (int jb = 0; jb < n; jb += 32) {
                unsigned m = 0u;
                const int je = min(32, n - jb);
                const float4* __restrict__ pp = s_pq + jb;
#pragma unroll 2
                for (int jj = 0; jj < je; jj += 2, pp += 2) {
                    const float4 P4 = pp[0], V4 = pp[1];
                    const float2 mx = sub2(make_float2(P4.x, P4.y), px2), my = sub2(make_float2(P4.z, P4.w), py2);
                    const float2 dvx = sub2(make_float2(V4.x, V4.y), vx2), dvy = sub2(make_float2(V4.z, V4.w), vy2);
                    const float2 xx = mul2(mx, mx), yy2 = mul2(my, my), xd = mul2(mx, dvx), yd = mul2(my, dvy);
                    const float2 mm = make_float2(__fadd_rn(xx.x, yy2.x), __fadd_rn(xx.y, yy2.y));
                    const float2 bq = make_float2(__fadd_rn(xd.x, yd.x), __fadd_rn(xd.y, yd.y));
                    const bool ok1 = jj + 1 < je;
                    // comm.cpp:201-220: running max, strict '>', table order (the self pair has mm == 0 < FLT_MIN)
                    if (mm.x > best) { best = mm.x; bj = jb + jj; }
                    if (ok1 && mm.y > best) { best = mm.y; bj = jb + jj + 1; }
                    const float2 c = sub2(mm, RR2);
                    const float2 dxx = mul2(dvx, dvx), dyy = mul2(dvy, dvy);
                    const float2 a = make_float2(__fadd_rn(dxx.x, dyy.x), __fadd_rn(dxx.y, dyy.y));
                    const float2 bb = mul2(bq, bq), ac = mul2(a, c);
                    const float2 disc = make_float2(__fsub_rn(bb.x, ac.x), __fsub_rn(bb.y, ac.y));
                    // conservative "t > 5" filter (same derivation as crowd.cu scan_cells / DESIGN.md)
                    const float2 a5 = mul2(a, make_float2(5.f, 5.f));
                    const float2 u = make_float2(__fsub_rn(-bq.x, a5.x), __fsub_rn(-bq.y, a5.y));
                    const float2 uu = mul2(u, u), mb = mul2(bb, make_float2(1e-4f, 1e-4f));
                    // near_c > 0 (the collision reach exceeds the contact distance), so beyond it c > 0 holds and the tests
                    // on the sign of c drop out of the general form (crowd.cu filter2); comparison polarity kept, so a NaN
                    // anywhere still makes the pair a candidate
                    const bool far0 = u.x > 0.f && uu.x > __fadd_rn(disc.x, mb.x);
                    const bool far1 = u.y > 0.f && uu.y > __fadd_rn(disc.y, mb.y);
                    const bool cand0 = c.x < K.near_c || !(bq.x > 0.f || disc.x < 0.f || far0);
                    const bool cand1 = ok1 && (c.y < K.near_c || !(bq.y > 0.f || disc.y < 0.f || far1));
                    m |= ((unsigned)cand0 | ((unsigned)cand1 << 1)) << jj;
                }
                if ((unsigned)(i - jb) < 32u) m &= ~(1u << (i - jb));  // the self pair (m = dv = 0) always survives
                while (m) {
                    const int j = jb + __ffs(m) - 1;
                    m &= m - 1u;
                    slow(j);
                }
            }
            last_cnt = cnt;
            overlap2 += (unsigned long long)n_overlap;
            pair_tests += (unsigned long long)(n - 1);

            float forcex = 0.f, forcey = 0.f;
            // ---- S1: comm::run (comm.cpp:267-305) ----
            if (comm) {
                if (bj < 0) { err = LC_ERR_STATE; bj = i; }  // comm.cpp:273-285: null partner is UB in the reference
                mt.w = bj;
                const float4 pj = pvA[bj];
                const float4 cj = cA[bj];
                // comm.cpp:144-192 x(): COSINE_REL + PROX + SOFTSIGN_GOAL; partner (x, y) of a vec3 = (x, 0): quirk Q2
                const float rightx = facing.y, righty = -facing.x;
                const float dvx = __fsub_rn(pj.z, vx), dvy = __fsub_rn(0.f, vy);
                float rvx = dot2(rightx, righty, dvx, dvy), rvy = dot2(facing.x, facing.y, dvx, dvy);
                const float mag_rv = len2(rvx, rvy);
                if (mag_rv > 0.f) { rvx = __fdiv_rn(rvx, mag_rv); rvy = __fdiv_rn(rvy, mag_rv); }
                const float dpx = __fsub_rn(pj.x, px), dpy = __fsub_rn(0.f, py);
                float rpx = dot2(rightx, righty, dpx, dpy), rpy = dot2(facing.x, facing.y, dpx, dpy);
                float mag_rp = len2(rpx, rpy);
                if (mag_rp > 0.f) {
                    rpx = __fdiv_rn(rpx, mag_rp); rpy = __fdiv_rn(rpy, mag_rp);
                    mag_rp = __fdiv_rn(1.f, mag_rp);
                }
                float gg = c4.w;  // goal_dist of the previous frame (ai.cpp:154 runs after comm::run)
                gg = __fdiv_rn(gg, __fadd_rn(1.f, gg));
                const float x[8] = {cj.x, cj.y, cj.z, rvy, mag_rv, rpy, mag_rp, gg};
                float cn[3];
#pragma unroll
                for (int r = 0; r < 3; ++r) {  // comm.cpp:285 M_cv * x (Eigen 3.3 unrolled redux order, see oracle/canon.c)
                    const float* m = &s_w[8 * r];
                    float p0 = __fmul_rn(m[0], x[0]), p1 = __fmul_rn(m[1], x[1]), p2 = __fmul_rn(m[2], x[2]),
                          p3 = __fmul_rn(m[3], x[3]), p4 = __fmul_rn(m[4], x[4]), p5 = __fmul_rn(m[5], x[5]),
                          p6 = __fmul_rn(m[6], x[6]), p7 = __fmul_rn(m[7], x[7]);
                    float cp = __fadd_rn(__fadd_rn(__fadd_rn(p0, p1), __fadd_rn(p2, p3)),
                                         __fadd_rn(__fadd_rn(p4, p5), __fadd_rn(p6, p7)));
                    cn[r] = __fadd_rn(__fdiv_rn(cp, __fadd_rn(1.f, fabsf(__fmul_rn(2.f, cp)))), 0.5f);  // CommComp.h:20-27
                }
                float fb[2];
#pragma unroll
                for (int r = 0; r < 2; ++r) {
                    const float* m = &s_w[24 + 3 * r];
                    float fp = __fadd_rn(__fmul_rn(m[0], cj.x), __fadd_rn(__fmul_rn(m[1], cj.y), __fmul_rn(m[2], cj.z)));
                    fb[r] = glm_clamp(fp, -1.f, 1.f);  // CommComp.h:30-32
                }
                // pass 2, comm.cpp:288-304
                const float diffx = __fadd_rn(__fmul_rn(rightx, fb[0]), __fmul_rn(facing.x, fb[1]));
                const float diffy = __fadd_rn(__fmul_rn(righty, fb[0]), __fmul_rn(facing.y, fb[1]));
                if (!done_before) { forcex = __fadd_rn(forcex, diffx); forcey = __fadd_rn(forcey, diffy); }
                const float speed = __fsqrt_rn(dot3_xz(vx, vy, vx, vy));
                const float tlx = __fsub_rn(lgoal.x, px), tly = __fsub_rn(lgoal.y, py);
                if (speed > 0.f && len2(tlx, tly) > 0.f) {
                    const float inv = __fdiv_rn(1.0f, __fsqrt_rn(dot2(tlx, tly, tlx, tly)));
                    facing = make_float2(__fmul_rn(tlx, inv), __fmul_rn(tly, inv));
                }
                c4.x = cn[0]; c4.y = cn[1]; c4.z = cn[2];  // swap()
            }

            // ---- S2: ai::update_agents (ai.cpp:153-172) ----
            c4.w = gd;
            int num_done = mt.x & 1;
            if (f_rp && rp_below < 2) num_done = 0;  // replan gate: the first two in table order (ai.cpp:151-168) -> GMP::plan_one
            // LMP::lookahead (LMP.cpp:84-113) with plan = [final_goal]
            if (num_done < 1) {
                const float ox = __fsub_rn(px, goal.x), oy = __fsub_rn(py, goal.y);
                if (dot2(ox, oy, ox, oy) < 1.f) num_done = 1; else lgoal = goal;
            } else {
                lgoal = goal;
            }
            mt.x = (mt.x & ~1) | num_done;
            not_done_after = num_done == 0;
            // LMP.cpp:261-273: one draw for both components
            const float eps = lcg_uniform((uint32_t)mt.y, -.01f, .01f);
            mt.y = (int)lcg_mul((uint32_t)mt.y, K.lcg_stride);
            const float dgx = __fsub_rn(__fadd_rn(lgoal.x, eps), px), dgy = __fsub_rn(__fadd_rn(lgoal.y, eps), py);
            const float dl = glm_max(1.0f, len2(dgx, dgy));
            const float gvx = __fmul_rn(1.0f, __fdiv_rn(dgx, dl)), gvy = __fmul_rn(1.0f, __fdiv_rn(dgy, dl));
            const float goalFx = __fmul_rn(2.0f, __fsub_rn(gvx, vx)), goalFy = __fmul_rn(2.0f, __fsub_rn(gvy, vy));
            forcex = __fadd_rn(forcex, __fadd_rn(goalFx, ttcFx));  // ai.cpp:172
            forcey = __fadd_rn(forcey, __fadd_rn(goalFy, ttcFy));

            // ---- S3a: physics::simulate leap-frog (physics.cpp:41-51), mass == 1 ----
            float hx = __fadd_rn(vfh.x, __fmul_rn(LC_DT, forcex)), hy = __fadd_rn(vfh.y, __fmul_rn(LC_DT, forcey));
            const float h2 = dot3_xz(hx, hy, hx, hy);
            if (h2 > 1.f) {
                const float inv = __fdiv_rn(1.0f, __fsqrt_rn(h2));
                hx = __fmul_rn(hx, inv); hy = __fmul_rn(hy, inv);
            }
            nvx = __fadd_rn(__fdiv_rn(hx, 2.f), __fdiv_rn(vfh.x, 2.f));
            nvy = __fadd_rn(__fdiv_rn(hy, 2.f), __fdiv_rn(vfh.y, 2.f));
            npx = __fadd_rn(px, __fmul_rn(hx, LC_DT));
            npy = __fadd_rn(py, __fmul_rn(hy, LC_DT));
            vfh = make_float2(hx, hy);
            s_pv1[i] = make_float4(npx, npy, nvx, nvy);
            cB[i] = c4;
        }
        const int any_near = __syncthreads_or(n_near > 0);  // B2: integrated positions + new c are visible

        // ---- S3b: collision push-out (physics.cpp:58-69 + BVH.cpp:106-120), Jacobi, ascending id ----
        float fx_ = npx, fy_ = npy;
        if (valid && any_near && n_near > 0) {
            int cnt = 0;
            if (n_near <= MAX_NEAR) {
                for (int k = 0; k < n_near; ++k) {
                    const float4 o = s_pv1[s_near[k * nt + i]];
                    float ax, ay;
                    if (circ_circ_hit(npx, npy, o.x, o.y, K.r, K.RR, ax, ay)) { push_away(fx_, fy_, vfh.x, vfh.y, ax, ay); ++cnt; }
                }
            } else {
                for (int j = 0; j < n; ++j) {
                    if (j == i) continue;
                    const float4 o = s_pv1[j];
                    float ax, ay;
                    if (circ_circ_hit(npx, npy, o.x, o.y, K.r, K.RR, ax, ay)) { push_away(fx_, fy_, vfh.x, vfh.y, ax, ay); ++cnt; }
                }
            }
            coll_hits += (unsigned long long)cnt;
        }
        me = make_float4(fx_, fy_, nvx, nvy);
        if (valid) {  // read again only after the next B1 (or the final barrier)
            s_pv0[i] = me;
            my_pq[0] = me.x; my_pq[2] = me.y; my_pq[4] = me.z; my_pq[6] = me.w;
        }

        // physics.cpp:72-76 inputs for the fold at the next B1
        const float speed = __fsqrt_rn(dot3_xz(nvx, nvy, nvx, nvy));
        const unsigned ndm = __ballot_sync(0xffffffffu, not_done_after);
        const unsigned nfm = __ballot_sync(0xffffffffu, not_done_after && !(speed <= LC_FLT_MAX));
        if (lane == 0) { s_ndm[wid] = ndm; s_nfm[wid] = nfm; }
        if (ndm && lane == 31 - __clz(ndm)) s_spd[wid] = speed;
        finished = all_done;
        ++executed;
    }
    __syncthreads();
    if (i == 0) {
        if (executed) fold_avg_velocity(s_ctl, s_ndm, s_spd, s_nfm, nwarps);
        A.ctl[sc].frame = s_ctl.frame;
        A.ctl[sc].finished = s_ctl.finished;
        A.ctl[sc].total_num_done = s_ctl.total_num_done;
        A.ctl[sc].run_avg = s_ctl.run_avg;
        A.ctl[sc].M2 = s_ctl.M2;
        A.ctl[sc].run_std = s_ctl.run_std;
        A.ctl[sc].avg_velocity = s_ctl.avg_velocity;
    }
    // ---- store: registers -> global ----
    if (valid) {
        A.pv[g] = me;
        A.vfh[g] = vfh;
        A.lgoal[g] = lgoal;
        A.facing[g] = facing;
        A.c4[g] = c4;
        A.meta[g] = mt;
        A.nbr_cnt[g] = last_cnt;
        if (overlap2) atomicAdd(&A.ctl[sc].overlap2, overlap2);
        if (coll_hits) atomicAdd(&A.ctl[sc].coll_hits, coll_hits);
        atomicAdd(&A.ctl[sc].pair_tests, pair_tests);
        if (err) atomicCAS(&A.ctl[sc].error, 0, err);
    }
}

}  // namespace lcbatch

// ---------------------------------------------------------------------------------------------------------------------
int lc_set_error(int code, const char* msg);  // capi.cu: text returned by lc_last_error()
namespace {
int bfail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    return lc_set_error(code, buf);
}
}  // namespace

#define BCK(call)                                                                                                  \
    do {                                                                                                           \
        cudaError_t e_ = (call);                                                                                   \
        if (e_ != cudaSuccess)                                                                                     \
            return bfail(LC_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

struct lc_batch {
    lcbatch::Arrays a{};
    lcbatch::Consts k{};
    int device = 0;
    cudaStream_t st = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float last_ms = 0.f;
    long long launches = 0;
    std::vector<void*> allocs;
    // host copies of the initial state (lc_batch_reset) and a pinned mirror of the per-scenario control blocks
    std::vector<float4> h_pv, h_c4;
    std::vector<float2> h_goal, h_facing;
    std::vector<int4> h_meta;
    lcbatch::ScenarioCtl* h_ctl = nullptr;
    std::vector<int> order;  // lc_batch_desc::table_order: order[i] = caller's agent index of engine agent i (empty = identity)
    // index of engine agent g (scenario-major) in the caller's tables
    size_t ext(size_t g) const {
        if (order.empty()) return g;
        const size_t n = (size_t)a.n;
        return (g / n) * n + (size_t)order[g % n];
    }
};

static uint32_t b_lcg_pow(uint32_t a, uint64_t e) {
    uint32_t r = 1;
    while (e) {
        if (e & 1) r = lcm::lcg_mul(r, a);
        a = lcm::lcg_mul(a, a);
        e >>= 1;
    }
    return r;
}

template <class T> static cudaError_t balloc(lc_batch* b, T** p, size_t count) {
    cudaError_t e = cudaMalloc((void**)p, sizeof(T) * (count ? count : 1));
    if (e == cudaSuccess) b->allocs.push_back(*p);
    return e;
}

static int batch_upload_initial(lc_batch* b) {
    const size_t N = (size_t)b->a.S * b->a.n;
    BCK(cudaMemcpyAsync(b->a.pv, b->h_pv.data(), sizeof(float4) * N, cudaMemcpyHostToDevice, b->st));
    BCK(cudaMemsetAsync(b->a.vfh, 0, sizeof(float2) * N, b->st));
    BCK(cudaMemcpyAsync(b->a.goal, b->h_goal.data(), sizeof(float2) * N, cudaMemcpyHostToDevice, b->st));
    BCK(cudaMemsetAsync(b->a.lgoal, 0, sizeof(float2) * N, b->st));
    BCK(cudaMemcpyAsync(b->a.facing, b->h_facing.data(), sizeof(float2) * N, cudaMemcpyHostToDevice, b->st));
    BCK(cudaMemcpyAsync(b->a.c4, b->h_c4.data(), sizeof(float4) * N, cudaMemcpyHostToDevice, b->st));
    BCK(cudaMemcpyAsync(b->a.meta, b->h_meta.data(), sizeof(int4) * N, cudaMemcpyHostToDevice, b->st));
    BCK(cudaMemsetAsync(b->a.nbr_cnt, 0, sizeof(int) * N, b->st));
    BCK(cudaMemsetAsync(b->a.ctl, 0, sizeof(lcbatch::ScenarioCtl) * (size_t)b->a.S, b->st));
    BCK(cudaStreamSynchronize(b->st));
    return LC_OK;
}

extern "C" {

void lc_batch_destroy(lc_batch* b) {
    if (!b) return;
    cudaSetDevice(b->device);
    if (b->st) cudaStreamSynchronize(b->st);
    for (void* p : b->allocs) cudaFree(p);
    if (b->h_ctl) cudaFreeHost(b->h_ctl);
    if (b->ev0) cudaEventDestroy(b->ev0);
    if (b->ev1) cudaEventDestroy(b->ev1);
    if (b->st) cudaStreamDestroy(b->st);
    delete b;
}

int32_t lc_batch_create(const lc_batch_desc* d, lc_batch** out) {
    if (!d || !out) return bfail(LC_ERR_ARG, "lc_batch_create: null argument");
    *out = nullptr;
    if (d->n_scenarios < 1) return bfail(LC_ERR_ARG, "lc_batch_create: n_scenarios must be >= 1");
    if (d->n_agents < 2 || d->n_agents > lcbatch::MAX_N)
        return bfail(LC_ERR_ARG, "lc_batch_create: n_agents must be in [2, %d] (use lc_create for larger crowds)", lcbatch::MAX_N);
    if (!d->pos || !d->goal || !d->max_duration || !d->rng_state) return bfail(LC_ERR_ARG, "lc_batch_create: pos, goal, max_duration and rng_state are required");
    if (!(d->radius > 0.f)) return bfail(LC_ERR_ARG, "lc_batch_create: radius must be positive");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return bfail(LC_ERR_CUDA, "lc_batch_create: no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(e));
    if (d->device < 0 || d->device >= ndev) return bfail(LC_ERR_ARG, "lc_batch_create: device %d out of range", d->device);
    BCK(cudaSetDevice(d->device));
    lc_batch* b = new lc_batch;
    b->device = d->device;
    const int S = b->a.S = d->n_scenarios, n = b->a.n = d->n_agents;
    if (d->table_order) {
        std::vector<char> seen((size_t)n, 0);
        for (int i = 0; i < n; ++i) {
            const int e = d->table_order[i];
            if (e < 0 || e >= n || seen[e]) { delete b; return bfail(LC_ERR_ARG, "lc_batch_create: table_order is not a permutation of 0..n-1"); }
            seen[e] = 1;
        }
        b->order.assign(d->table_order, d->table_order + n);
    }
    const size_t N = (size_t)S * n;
#define BCKB(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { bfail(LC_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); lc_batch_destroy(b); return LC_ERR_CUDA; } } while (0)
    BCKB(cudaStreamCreateWithFlags(&b->st, cudaStreamNonBlocking));
    BCKB(cudaEventCreate(&b->ev0));
    BCKB(cudaEventCreate(&b->ev1));
    BCKB(balloc(b, &b->a.pv, N));
    BCKB(balloc(b, &b->a.vfh, N));
    BCKB(balloc(b, &b->a.goal, N));
    BCKB(balloc(b, &b->a.lgoal, N));
    BCKB(balloc(b, &b->a.facing, N));
    BCKB(balloc(b, &b->a.c4, N));
    BCKB(balloc(b, &b->a.meta, N));
    BCKB(balloc(b, &b->a.nbr_cnt, N));
    BCKB(balloc(b, &b->a.m_cv, (size_t)S * 24));
    BCKB(balloc(b, &b->a.m_f, (size_t)S * 6));
    BCKB(balloc(b, &b->a.max_duration, (size_t)S));
    BCKB(balloc(b, &b->a.ctl, (size_t)S));
    BCKB(cudaMallocHost((void**)&b->h_ctl, sizeof(lcbatch::ScenarioCtl) * (size_t)S));
    // constants shared by all scenarios (same derivation as lc_create)
    b->k.r = d->radius;
    const float R = d->radius + d->radius;  // LMP.cpp:53 i._r + j._r
    b->k.RR = R * R;                        // BoundVolume.cpp:106
    const float reach = R + 2.f * (1.f / 60.f) * 1.01f + 0.002f;
    b->k.near_c = reach * reach - b->k.RR;
    b->k.lcg_stride = b_lcg_pow(16807u, (uint64_t)n);
    b->k.flags = d->flags;
    // initial component values: demo.cpp:563-581 create_robo, ai.cpp:107-127 ai::init, physics.cpp:15-21 prewarm
    b->h_pv.resize(N); b->h_c4.resize(N); b->h_goal.resize(N); b->h_facing.resize(N); b->h_meta.resize(N);
    for (int s = 0; s < S; ++s) {
        uint32_t x = d->rng_state[s];
        for (int i = 0; i < n; ++i) {
            const size_t g = (size_t)s * n + i, e = b->ext(g);  // engine agent i = table position; e = the caller's index
            const float px = d->pos[2 * e], py = d->pos[2 * e + 1], gx = d->goal[2 * e], gy = d->goal[2 * e + 1];
            const float dx = gx - px, dy = gy - py;
            const float len = sqrtf(dx * dx + dy * dy);  // glm::length (ai.cpp:116)
            float mtime = -1.0f;                         // ai.cpp:121-126
            mtime += len;
            x = lcm::lcg_mul(x, 16807u);                 // agent i draws element (frame-1)*n + i + 1 of the scenario's stream
            b->h_pv[g] = make_float4(px, py, 0.f, 0.f);
            b->h_goal[g] = make_float2(gx, gy);
            b->h_facing[g] = make_float2(dx, dy);        // un-normalised (demo.cpp:580)
            b->h_c4[g] = make_float4(0.f, 0.f, 0.f, len);
            int mbits;
            memcpy(&mbits, &mtime, 4);
            b->h_meta[g] = make_int4(0, (int)x, mbits, -1);
        }
    }
    int rc = batch_upload_initial(b);
    if (rc) { lc_batch_destroy(b); return rc; }
    std::vector<float> zeros((size_t)S * 24, 0.f);
    BCKB(cudaMemcpy(b->a.m_cv, d->m_cv ? d->m_cv : zeros.data(), sizeof(float) * (size_t)S * 24, cudaMemcpyHostToDevice));
    BCKB(cudaMemcpy(b->a.m_f, d->m_f ? d->m_f : zeros.data(), sizeof(float) * (size_t)S * 6, cudaMemcpyHostToDevice));
    BCKB(cudaMemcpy(b->a.max_duration, d->max_duration, sizeof(float) * (size_t)S, cudaMemcpyHostToDevice));
#undef BCKB
    *out = b;
    return LC_OK;
}

#define BATCH_ENTER(b)                                                 \
    if (!(b)) return bfail(LC_ERR_ARG, "%s: null handle", __func__);   \
    BCK(cudaSetDevice((b)->device))

int32_t lc_batch_reset(lc_batch* b, const float* m_cv, const float* m_f) {
    BATCH_ENTER(b);
    int rc = batch_upload_initial(b);
    if (rc) return rc;
    if (m_cv) BCK(cudaMemcpy(b->a.m_cv, m_cv, sizeof(float) * (size_t)b->a.S * 24, cudaMemcpyHostToDevice));
    if (m_f) BCK(cudaMemcpy(b->a.m_f, m_f, sizeof(float) * (size_t)b->a.S * 6, cudaMemcpyHostToDevice));
    return LC_OK;
}

int32_t lc_batch_run(lc_batch* b, int32_t max_frames, float* avg_velocity, float* group_time, int32_t* n_done,
                     int32_t* frames) {
    BATCH_ENTER(b);
    if (max_frames < 0) return bfail(LC_ERR_ARG, "lc_batch_run: max_frames < 0");
    const int nt = (b->a.n + 31) & ~31;
    const size_t smem = sizeof(float4) * 5 * nt + sizeof(int) * lcbatch::MAX_NEAR * nt;
    BCK(cudaEventRecord(b->ev0, b->st));
    lcbatch::k_batch_rollout<<<b->a.S, nt, smem, b->st>>>(b->a, b->k, max_frames);
    BCK(cudaEventRecord(b->ev1, b->st));
    b->launches += 1;
    BCK(cudaMemcpyAsync(b->h_ctl, b->a.ctl, sizeof(lcbatch::ScenarioCtl) * (size_t)b->a.S, cudaMemcpyDeviceToHost, b->st));
    BCK(cudaStreamSynchronize(b->st));
    BCK(cudaGetLastError());
    BCK(cudaEventElapsedTime(&b->last_ms, b->ev0, b->ev1));
    int first_err = 0, first_err_s = -1;
    for (int s = 0; s < b->a.S; ++s) {
        const lcbatch::ScenarioCtl& c = b->h_ctl[s];
        if (avg_velocity) avg_velocity[s] = c.avg_velocity;
        if (group_time) group_time[s] = c.run_avg + c.run_std * 3;  // comm.cpp:139
        if (n_done) n_done[s] = c.total_num_done;
        if (frames) frames[s] = (int32_t)c.frame;
        if (c.error && !first_err) { first_err = c.error; first_err_s = s; }
    }
    if (first_err) return bfail(first_err, "scenario %d: device-side error %d (an agent without comm partner, comm.cpp:273-285)", first_err_s, first_err);
    return LC_OK;
}

int32_t lc_batch_finished(lc_batch* b, int32_t* n_finished) {
    BATCH_ENTER(b);
    BCK(cudaMemcpyAsync(b->h_ctl, b->a.ctl, sizeof(lcbatch::ScenarioCtl) * (size_t)b->a.S, cudaMemcpyDeviceToHost, b->st));
    BCK(cudaStreamSynchronize(b->st));
    int k = 0;
    for (int s = 0; s < b->a.S; ++s) k += b->h_ctl[s].finished != 0;
    if (n_finished) *n_finished = k;
    return LC_OK;
}

int32_t lc_batch_get_state(lc_batch* b, float* pos, float* vel, float* vfh, float* c, float* facing, float* lgoal,
                           float* goal_dist, int32_t* num_done) {
    BATCH_ENTER(b);
    const size_t N = (size_t)b->a.S * b->a.n;
    std::vector<float4> pv(N), c4(N);
    std::vector<float2> f2(N);
    std::vector<int4> mt(N);
    BCK(cudaStreamSynchronize(b->st));
    if (pos || vel) {
        BCK(cudaMemcpy(pv.data(), b->a.pv, sizeof(float4) * N, cudaMemcpyDeviceToHost));
        for (size_t g = 0; g < N; ++g) {
            const size_t e = b->ext(g);
            if (pos) { pos[2 * e] = pv[g].x; pos[2 * e + 1] = pv[g].y; }
            if (vel) { vel[2 * e] = pv[g].z; vel[2 * e + 1] = pv[g].w; }
        }
    }
    for (int t = 0; t < 3; ++t) {
        float* dst = t == 0 ? vfh : t == 1 ? facing : lgoal;
        const float2* src = t == 0 ? b->a.vfh : t == 1 ? b->a.facing : b->a.lgoal;
        if (!dst) continue;
        BCK(cudaMemcpy(f2.data(), src, sizeof(float2) * N, cudaMemcpyDeviceToHost));
        for (size_t g = 0; g < N; ++g) { const size_t e = b->ext(g); dst[2 * e] = f2[g].x; dst[2 * e + 1] = f2[g].y; }
    }
    if (c || goal_dist) {
        BCK(cudaMemcpy(c4.data(), b->a.c4, sizeof(float4) * N, cudaMemcpyDeviceToHost));
        for (size_t g = 0; g < N; ++g) {
            const size_t e = b->ext(g);
            if (c) { c[3 * e] = c4[g].x; c[3 * e + 1] = c4[g].y; c[3 * e + 2] = c4[g].z; }
            if (goal_dist) goal_dist[e] = c4[g].w;
        }
    }
    if (num_done) {
        BCK(cudaMemcpy(mt.data(), b->a.meta, sizeof(int4) * N, cudaMemcpyDeviceToHost));
        for (size_t g = 0; g < N; ++g) num_done[b->ext(g)] = mt[g].x & 1;
    }
    return LC_OK;
}

int32_t lc_batch_get_diag(lc_batch* b, int32_t* partner, int32_t* neighbor_count, int64_t* overlap_pairs,
                          int64_t* collision_hits, int64_t* pair_tests) {
    BATCH_ENTER(b);
    const size_t N = (size_t)b->a.S * b->a.n;
    BCK(cudaStreamSynchronize(b->st));
    if (partner) {
        std::vector<int4> mt(N);
        BCK(cudaMemcpy(mt.data(), b->a.meta, sizeof(int4) * N, cudaMemcpyDeviceToHost));
        for (size_t g = 0; g < N; ++g) {
            const int pj = mt[g].w;  // engine index within the scenario
            partner[b->ext(g)] = (pj >= 0 && !b->order.empty()) ? b->order[pj] : pj;
        }
    }
    if (neighbor_count) {
        std::vector<int> cnt(N);
        BCK(cudaMemcpy(cnt.data(), b->a.nbr_cnt, sizeof(int) * N, cudaMemcpyDeviceToHost));
        for (size_t g = 0; g < N; ++g) neighbor_count[b->ext(g)] = cnt[g];
    }
    BCK(cudaMemcpy(b->h_ctl, b->a.ctl, sizeof(lcbatch::ScenarioCtl) * (size_t)b->a.S, cudaMemcpyDeviceToHost));
    for (int s = 0; s < b->a.S; ++s) {
        if (overlap_pairs) overlap_pairs[s] = (int64_t)(b->h_ctl[s].overlap2 / 2);
        if (collision_hits) collision_hits[s] = (int64_t)b->h_ctl[s].coll_hits;
        if (pair_tests) pair_tests[s] = (int64_t)b->h_ctl[s].pair_tests;
    }
    return LC_OK;
}

int32_t lc_batch_timer(lc_batch* b, float* milliseconds) {
    BATCH_ENTER(b);
    if (milliseconds) *milliseconds = b->last_ms;
    return LC_OK;
}

int64_t lc_batch_kernel_launches(lc_batch* b) { return b ? b->launches : 0; }

}  // extern "C"
