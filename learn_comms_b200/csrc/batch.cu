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
#include "walls.cuh"
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <math_constants.h>
#include <map>
#include <string>
#include <vector>

namespace lcbatch {
using namespace lcm;

constexpr int MAX_N = 256;
constexpr int MAX_WARPS = MAX_N / 32;
constexpr int MAX_NEAR = 8;

constexpr int WP_CAP = 32;      // waypoints per agent (walled scenarios)
constexpr int MAX_WALLS = 64;   // walls of the layout a batch shares (the catalogue's largest scene has 60)

struct ScenarioCtl {
    uint32_t frame;
    int finished;
    int total_num_done;
    float run_avg, M2, run_std, avg_velocity;
    int error;
    unsigned long long overlap2;   // ordered overlapping pairs at frame start, summed over all frames
    unsigned long long coll_hits;  // ordered (q, bv) hits of the collision pass, summed over all frames
    unsigned long long pair_tests; // TTC pair evaluations (throughput accounting)
    // walled scenarios: the block stops in front of a frame whose replan gate fired (ai.cpp:152-168) and waits for the
    // host planner; wr1 / wr2 = the two lowest requesting agents, req = where they replan from
    int pause, served, wr1, wr2;
    float req[4];
    unsigned long long draws;      // minstd_rand0 draws consumed so far (plan-less agents do not draw, LMP.cpp:267-270)
    unsigned long long wall_hits;
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
    // walled scenarios (one wall layout per batch); per-agent arrays [S][n]
    int n_walls, n_bvh;
    const lcw::Wall* walls;      // [n_walls] creation order
    const int* wall_order;       // [n_walls] statics order (LMP.cpp:307)
    const lc_bvh_node* bvh;      // [n_bvh] static BVH, pre-order (lc_static_bvh)
    float2* plan;                // [S][n][WP_CAP]  Agent::plan
    int* n_plan;                 // Agent::plan->size(); 0 = no plan
    int* wnd;                    // Agent::num_done (waypoints passed)
    int* lost;                   // Agent::lost_frames; -1 = "replanned by the host for the coming frame"
    uint32_t* rng0;              // [S] stream state before the scenario's first draw
    uint32_t* frame_target;      // [S] lc_batch_run stops a scenario in front of this frame
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

__device__ __forceinline__ uint32_t lcg_pow_dev(unsigned long long e) {  // 16807^e mod (2^31 - 1)
    uint32_t r = 1u, a = 16807u;
    while (e) {
        if (e & 1ull) r = lcg_mul(r, a);
        a = lcg_mul(a, a);
        e >>= 1;
    }
    return r;
}

#ifndef LC_BATCH_UNROLL
#define LC_BATCH_UNROLL 2
#endif
constexpr int PAIR_UNROLL = LC_BATCH_UNROLL;  // slot pairs per iteration of the pair loop (#pragma unroll does not expand macros)
template <bool WALLS>
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
    __shared__ lcw::Wall s_walls[WALLS ? MAX_WALLS : 1];
    __shared__ int s_worder[WALLS ? MAX_WALLS : 1];
    __shared__ lc_bvh_node s_bvh[WALLS ? 2 * MAX_WALLS : 1];
    __shared__ unsigned s_req[MAX_WARPS], s_plm[MAX_WARPS];  // per-warp ballots: needs a replan / has a plan

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
    // walled scenario: waypoint list (global, by agent), counters in registers, the shared wall layout in shared memory
    int np_ = 0, nd = 0, lost = 0;
    const float2* myplan = nullptr;
    if constexpr (WALLS) {
        if (valid) { np_ = A.n_plan[g]; nd = A.wnd[g]; lost = A.lost[g]; }
        myplan = A.plan + g * WP_CAP;
        for (int k = i; k < A.n_walls; k += nt) { s_walls[k] = A.walls[k]; s_worder[k] = A.wall_order[k]; }
        for (int k = i; k < A.n_bvh; k += nt) s_bvh[k] = A.bvh[k];
    }
    __syncthreads();
    const float max_duration = A.max_duration[sc];
    const uint32_t frame0 = s_ctl.frame;
    bool finished = s_ctl.finished != 0;
    bool served = WALLS && s_ctl.served != 0;
    int paused = 0, wr1 = -1, wr2 = -1;
    unsigned long long draws = s_ctl.draws, wall_hits = 0ull;
    const uint32_t rng0 = WALLS ? A.rng0[sc] : 0u;
    const uint32_t target = WALLS ? A.frame_target[sc] : 0u;
    const bool comm = !(K.flags & LC_FLAG_NO_COMM);
    unsigned long long overlap2 = 0ull, coll_hits = 0ull, pair_tests = 0ull;
    int err = 0;
    int executed = 0;

    for (int it = 0; (WALLS || it < max_frames) && !finished; ++it) {
        const uint32_t frame = frame0 + (uint32_t)it + 1u;
        if (WALLS && frame0 + (uint32_t)it >= target) break;
        const int par = it & 1;
        const float4* __restrict__ pvA = s_pv0;
        const float4* cA = par ? s_c1 : s_c0;
        float4* cB = par ? s_c0 : s_c1;
        const float px = me.x, py = me.y, vx = me.z, vy = me.w;
        const float fake_time = __fmul_rn(LC_DT, (float)frame);
        const bool timeout = fake_time > max_duration;

        // ---- walled: the replan gate of this frame (ai.cpp:152-163), BEFORE anything of the frame has side effects.
        // The two lowest requesting agents (= the first two in table order, ai.cpp:151-152,164) are planned by the host
        // (GMP::plan_one needs the roadmap); the block stops here and is resumed by the next launch. ----
        int lost_next = 0;
        if constexpr (WALLS) {
            bool need = false;
            if (valid) {
                if (lost < 0) {
                    lost_next = 0;  // replanned for this frame: lost_frames = 0 (ai.cpp:167), no second request
                } else {
                    const bool has = np_ > 0, done = has && nd == np_;
                    const float gd0 = len2(__fsub_rn(goal.x, px), __fsub_rn(goal.y, py));
                    bool visible = false;
                    if (has && nd < np_) {
                        const float2 t = myplan[nd];
                        visible = true;
                        for (int k = 0; k < A.n_walls && visible; ++k) visible = lcw::cspace_line_of_sight(s_walls[k], K.r, px, py, t.x, t.y);
                    }
                    lost_next = visible ? 0 : lost + 1;
                    need = !has || (done && gd0 > 1.f) || (!done && !visible && lost_next >= 360);
                }
            }
            const unsigned m_req = __ballot_sync(0xffffffffu, need);
            const unsigned m_pl = __ballot_sync(0xffffffffu, valid && np_ > 0);
            if (lane == 0) { s_req[wid] = m_req; s_plm[wid] = m_pl; }
            __syncthreads();  // B0
            if (!served) {
                wr1 = wr2 = -1;
                for (int w = 0; w < nwarps && wr2 < 0; ++w)
                    for (unsigned m = s_req[w]; m && wr2 < 0; m &= m - 1u) {
                        const int id = 32 * w + __ffs(m) - 1;
                        if (wr1 < 0) wr1 = id; else wr2 = id;
                    }
                if (wr1 >= 0) {
                    if (i == wr1) { s_ctl.req[0] = px; s_ctl.req[1] = py; }
                    if (i == wr2) { s_ctl.req[2] = px; s_ctl.req[3] = py; }
                    paused = 1;
                    break;
                }
            }
            served = false;
        }

        // ---- S0: demo::run (demo.cpp:665-702), block-wide facts through per-warp ballots ----
        const bool done_before = mt.x & 1;
        const float gd = len2(__fsub_rn(goal.x, px), __fsub_rn(goal.y, py));  // ai.cpp:154
        const bool f_new = valid && done_before && !(mt.x & 2);
        const bool f_nd = valid && !done_before;
        const bool f_rp = !WALLS && valid && done_before && gd > 1.f;         // ai.cpp:161-163 (wall-free; walled: the gate above)
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
#pragma unroll PAIR_UNROLL
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
            if constexpr (WALLS) {  // LMP.cpp:306-319: the statics loop, after the agents, in the BoundVolume* table's order
                for (int k = 0; k < A.n_walls; ++k) {
                    float fx, fy;
                    if (lcw::wall_force(px, py, K.r, vx, vy, s_walls[s_worder[k]], fx, fy)) {
                        ttcFx = __fadd_rn(ttcFx, fx);
                        ttcFy = __fadd_rn(ttcFy, fy);
                    }
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
                const float dpx = __fsub_rn(pj.x, px), dpy = __fsub_rn(0.f, py);
                float x[8];  // goal_dist is last frame's (ai.cpp:154 runs after comm::run)
                comm_features(K.flags, rightx, righty, facing.x, facing.y, dvx, dvy, dpx, dpy, c4.w, cj.x, cj.y, cj.z, x);
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
            int num_done;         // 1 = Agent::done() after this frame's lookahead
            bool planned = true;  // Agent::has_plan()
            float eps = 0.f;
            if constexpr (!WALLS) {
                num_done = mt.x & 1;
                if (f_rp && rp_below < 2) num_done = 0;  // replan gate: the first two in table order (ai.cpp:151-168) -> GMP::plan_one
                // LMP::lookahead (LMP.cpp:84-113) with plan = [final_goal]
                if (num_done < 1) {
                    const float ox = __fsub_rn(px, goal.x), oy = __fsub_rn(py, goal.y);
                    if (dot2(ox, oy, ox, oy) < 1.f) num_done = 1; else lgoal = goal;
                } else {
                    lgoal = goal;
                }
                // LMP.cpp:261-273: one draw for both components
                eps = lcg_uniform((uint32_t)mt.y, -.01f, .01f);
                mt.y = (int)lcg_mul((uint32_t)mt.y, K.lcg_stride);
            } else {
                lost = lost_next;  // ai.cpp:161
                planned = np_ > 0;
                if (planned) {
                    lcw::lookahead(s_walls, A.n_walls, K.r, px, py, reinterpret_cast<const float*>(myplan), np_, nd, lgoal.x, lgoal.y);
                    // this agent draws element draws + (planned agents before it in the table) + 1 of the scenario's stream
                    int rank = __popc(s_plm[wid] & ((1u << lane) - 1u));
                    for (int w = 0; w < wid; ++w) rank += __popc(s_plm[w]);
                    eps = lcg_uniform(lcg_mul(rng0, lcg_pow_dev(draws + (unsigned long long)rank + 1ull)), -.01f, .01f);
                } else {
                    lgoal = make_float2(px, py);  // LMP.cpp:267-270: no plan, no draw, goal_vel = 0
                }
                num_done = planned && nd == np_;
            }
            mt.x = (mt.x & ~1) | num_done;
            not_done_after = num_done == 0;
            const float dgx = __fsub_rn(__fadd_rn(lgoal.x, eps), px), dgy = __fsub_rn(__fadd_rn(lgoal.y, eps), py);
            const float dl = glm_max(1.0f, len2(dgx, dgy));
            float gvx = __fmul_rn(1.0f, __fdiv_rn(dgx, dl)), gvy = __fmul_rn(1.0f, __fdiv_rn(dgy, dl));
            if (WALLS && !planned) gvx = gvy = 0.f;
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
        if constexpr (WALLS) {
            // physics.cpp:60,67-69: static_bvh->query at the integrated position, pushed out after the agent hits;
            // BVH::query_(Circ*) (BVH.cpp:106-138): depth first, left before right, single-point inner test (quirk Q8)
            if (valid && A.n_bvh > 0) {
                int stack[32], sp = 0;
                stack[sp++] = 0;
                while (sp > 0) {
                    const lc_bvh_node nd_ = s_bvh[stack[--sp]];
                    float nx, ny;
                    if (nd_.right < 0) {
                        if (lcw::circ_wall_hit(npx, npy, K.r, s_walls[nd_.left], nx, ny)) { push_away(fx_, fy_, vfh.x, vfh.y, nx, ny); ++wall_hits; }
                    } else if (lcw::circ_wall_hit(npx, npy, K.r, lcw::Wall{nd_.ox, nd_.oy, nd_.w, nd_.h}, nx, ny) && sp + 2 <= 32) {
                        stack[sp++] = nd_.right;
                        stack[sp++] = nd_.left;
                    }
                }
            }
            int npl = 0;  // draws this frame consumed = agents with a plan (block-uniform)
            for (int w = 0; w < nwarps; ++w) npl += __popc(s_plm[w]);
            draws += (unsigned long long)npl;
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
    if constexpr (WALLS) {
        if (valid) {
            A.wnd[g] = nd;
            A.lost[g] = lost;
            if (wall_hits) atomicAdd(&A.ctl[sc].wall_hits, wall_hits);
        }
        if (i == 0) {
            A.ctl[sc].pause = paused;
            A.ctl[sc].served = 0;
            A.ctl[sc].wr1 = wr1; A.ctl[sc].wr2 = wr2;
            A.ctl[sc].req[0] = s_ctl.req[0]; A.ctl[sc].req[1] = s_ctl.req[1]; A.ctl[sc].req[2] = s_ctl.req[2]; A.ctl[sc].req[3] = s_ctl.req[3];
            A.ctl[sc].draws = draws;
        }
    }
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

// Replans served by the host (GMP::plan_one, GMP.cpp:91-110): new waypoint list, num_done = 0, "replanned" marker for
// lost_frames; the scenario may run again.
struct ReplanUpdate { int scenario, agent, n_waypoints, pad; float wp[2 * WP_CAP]; };
__global__ void k_batch_apply_replans(Arrays A, const ReplanUpdate* up, int count) {
    const int u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= count) return;
    const ReplanUpdate& r = up[u];
    const size_t g = (size_t)r.scenario * A.n + r.agent;
    for (int k = 0; k < r.n_waypoints; ++k) A.plan[g * WP_CAP + k] = make_float2(r.wp[2 * k], r.wp[2 * k + 1]);
    A.n_plan[g] = r.n_waypoints;
    A.wnd[g] = 0;
    A.lost[g] = -1;
    A.ctl[r.scenario].served = 1;
    A.ctl[r.scenario].pause = 0;
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
    std::vector<float2> h_goal2, h_facing;
    std::vector<int4> h_meta;
    lcbatch::ScenarioCtl* h_ctl = nullptr;
    // walled batches: one wall layout; scenarios with identical start / goal tables form a group that shares its initial
    // plans and a template planner (state after ai::init's planning loop); a scenario gets its own copy of the roadmap at
    // its first replan (the reference's roadmap is per process and stateful)
    int n_walls = 0;
    std::vector<lc_rect> walls;
    std::vector<int> group_of;            // [S]
    std::vector<lc_planner*> tmpl;        // [G]
    std::vector<lc_planner*> planner;     // [S] lazily cloned
    std::vector<float> plan0;             // [G][n][WP_CAP][2]
    std::vector<int> nplan0;              // [G][n]
    std::vector<float> h_goal;            // [S][n][2] by engine index
    std::vector<uint32_t> h_rng0;         // [S]
    lcbatch::ReplanUpdate* d_updates = nullptr;
    size_t updates_cap = 0;
    long long replans = 0;
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
    BCK(cudaMemcpyAsync(b->a.goal, b->h_goal2.data(), sizeof(float2) * N, cudaMemcpyHostToDevice, b->st));
    BCK(cudaMemsetAsync(b->a.lgoal, 0, sizeof(float2) * N, b->st));
    BCK(cudaMemcpyAsync(b->a.facing, b->h_facing.data(), sizeof(float2) * N, cudaMemcpyHostToDevice, b->st));
    BCK(cudaMemcpyAsync(b->a.c4, b->h_c4.data(), sizeof(float4) * N, cudaMemcpyHostToDevice, b->st));
    BCK(cudaMemcpyAsync(b->a.meta, b->h_meta.data(), sizeof(int4) * N, cudaMemcpyHostToDevice, b->st));
    BCK(cudaMemsetAsync(b->a.nbr_cnt, 0, sizeof(int) * N, b->st));
    BCK(cudaMemsetAsync(b->a.ctl, 0, sizeof(lcbatch::ScenarioCtl) * (size_t)b->a.S, b->st));
    if (b->n_walls > 0) {  // initial plans of every scenario = its group's; counters at zero; roadmaps back to the templates
        const size_t n = (size_t)b->a.n;
        std::vector<float> plan((size_t)b->a.S * n * lcbatch::WP_CAP * 2);
        std::vector<int> nplan((size_t)b->a.S * n);
        for (int s = 0; s < b->a.S; ++s) {
            const size_t gi = (size_t)b->group_of[s];
            std::copy(b->plan0.begin() + gi * n * lcbatch::WP_CAP * 2, b->plan0.begin() + (gi + 1) * n * lcbatch::WP_CAP * 2,
                      plan.begin() + (size_t)s * n * lcbatch::WP_CAP * 2);
            std::copy(b->nplan0.begin() + gi * n, b->nplan0.begin() + (gi + 1) * n, nplan.begin() + (size_t)s * n);
        }
        BCK(cudaMemcpyAsync(b->a.plan, plan.data(), sizeof(float) * plan.size(), cudaMemcpyHostToDevice, b->st));
        BCK(cudaMemcpyAsync(b->a.n_plan, nplan.data(), sizeof(int) * nplan.size(), cudaMemcpyHostToDevice, b->st));
        BCK(cudaMemsetAsync(b->a.wnd, 0, sizeof(int) * N, b->st));
        BCK(cudaMemsetAsync(b->a.lost, 0, sizeof(int) * N, b->st));
        BCK(cudaStreamSynchronize(b->st));
        for (lc_planner*& p : b->planner) { if (p) lc_planner_destroy(p); p = nullptr; }
        b->replans = 0;
    }
    BCK(cudaStreamSynchronize(b->st));
    return LC_OK;
}

extern "C" {

void lc_batch_destroy(lc_batch* b) {
    if (!b) return;
    cudaSetDevice(b->device);
    if (b->st) cudaStreamSynchronize(b->st);
    for (void* p : b->allocs) cudaFree(p);
    for (lc_planner* p : b->planner) if (p) lc_planner_destroy(p);
    for (lc_planner* p : b->tmpl) if (p) lc_planner_destroy(p);
    if (b->d_updates) cudaFree(b->d_updates);
    if (b->h_ctl) cudaFreeHost(b->h_ctl);
    if (b->ev0) cudaEventDestroy(b->ev0);
    if (b->ev1) cudaEventDestroy(b->ev1);
    if (b->st) cudaStreamDestroy(b->st);
    delete b;
}

static int32_t batch_create_impl(const lc_batch_desc* d, lc_batch** out) {
    if (!d || !out) return bfail(LC_ERR_ARG, "lc_batch_create: null argument");
    *out = nullptr;
    if (d->n_scenarios < 1) return bfail(LC_ERR_ARG, "lc_batch_create: n_scenarios must be >= 1");
    if (d->n_agents < 2 || d->n_agents > lcbatch::MAX_N)
        return bfail(LC_ERR_ARG, "lc_batch_create: n_agents must be in [2, %d] (use lc_create for larger crowds)", lcbatch::MAX_N);
    if (!d->pos || !d->goal || !d->max_duration || !d->rng_state) return bfail(LC_ERR_ARG, "lc_batch_create: pos, goal, max_duration and rng_state are required");
    if (!(d->radius > 0.f)) return bfail(LC_ERR_ARG, "lc_batch_create: radius must be positive");
    if (d->n_walls < 0 || d->n_walls > lcbatch::MAX_WALLS || (d->n_walls > 0 && !d->walls))
        return bfail(LC_ERR_ARG, "lc_batch_create: n_walls must be in [0, %d] with the rectangles given", lcbatch::MAX_WALLS);
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
    // walled batch (ai::init, ai.cpp:24-129): static BVH, one planner per group of identical scenarios, initial plans in
    // table order (engine index = table position), min_time = plan length minus the untimed last metre
    std::vector<float> g_min_time;  // [G][n]
    if (d->n_walls > 0) {
        const int nw = b->n_walls = d->n_walls;
        b->walls.assign(d->walls, d->walls + nw);
        std::vector<int32_t> worder((size_t)nw);
        std::vector<char> seen((size_t)nw, 0);
        for (int k = 0; k < nw; ++k) {
            worder[k] = d->wall_order ? d->wall_order[k] : k;
            if (worder[k] < 0 || worder[k] >= nw || seen[worder[k]]) { lc_batch_destroy(b); return bfail(LC_ERR_ARG, "lc_batch_create: wall_order is not a permutation"); }
            seen[worder[k]] = 1;
        }
        std::vector<lc_bvh_node> bvh((size_t)2 * nw);
        int32_t n_bvh = 0;
        if (lc_static_bvh(nw, d->walls, worder.data(), bvh.data(), (int32_t)bvh.size(), &n_bvh) != LC_OK) { lc_batch_destroy(b); return bfail(LC_ERR_ARG, "lc_batch_create: static BVH construction failed"); }
        b->group_of.assign((size_t)S, 0);
        b->planner.assign((size_t)S, nullptr);
        std::map<std::string, int> groups;
        for (int s = 0; s < S; ++s) {
            std::string key((const char*)(d->pos + (size_t)2 * s * n), sizeof(float) * 2 * (size_t)n);
            key.append((const char*)(d->goal + (size_t)2 * s * n), sizeof(float) * 2 * (size_t)n);
            auto it = groups.find(key);
            if (it != groups.end()) { b->group_of[s] = it->second; continue; }
            const int gi = (int)b->tmpl.size();
            groups.emplace(std::move(key), gi);
            b->group_of[s] = gi;
            lc_planner* pl = nullptr;
            if (lc_planner_create(nw, d->walls, d->radius, &pl) != LC_OK) { lc_batch_destroy(b); return LC_ERR_CAPACITY; }
            b->tmpl.push_back(pl);
            b->plan0.resize((size_t)(gi + 1) * n * lcbatch::WP_CAP * 2, 0.f);
            b->nplan0.resize((size_t)(gi + 1) * n, 0);
            g_min_time.resize((size_t)(gi + 1) * n, 0.f);
            for (int i = 0; i < n; ++i) {
                const size_t e = b->ext((size_t)s * n + i);
                int32_t k = 0;
                const int rc = lc_planner_plan(pl, d->pos + 2 * e, d->goal + 2 * e, lcbatch::WP_CAP,
                                               &b->plan0[((size_t)gi * n + i) * lcbatch::WP_CAP * 2], &k, &g_min_time[(size_t)gi * n + i]);
                if (rc) { lc_batch_destroy(b); return bfail(rc, "lc_batch_create: the plan of agent %d of scenario %d has %d waypoints (capacity %d)", i, s, k, lcbatch::WP_CAP); }
                b->nplan0[(size_t)gi * n + i] = k;
            }
        }
        lcw::Wall* dw = nullptr; int* dwo = nullptr; lc_bvh_node* dbvh = nullptr;
        BCKB(balloc(b, &dw, (size_t)nw)); BCKB(balloc(b, &dwo, (size_t)nw)); BCKB(balloc(b, &dbvh, (size_t)n_bvh));
        static_assert(sizeof(lcw::Wall) == sizeof(lc_rect), "wall layout shared with the C ABI");
        BCKB(cudaMemcpy(dw, d->walls, sizeof(lc_rect) * (size_t)nw, cudaMemcpyHostToDevice));
        BCKB(cudaMemcpy(dwo, worder.data(), sizeof(int) * (size_t)nw, cudaMemcpyHostToDevice));
        BCKB(cudaMemcpy(dbvh, bvh.data(), sizeof(lc_bvh_node) * (size_t)n_bvh, cudaMemcpyHostToDevice));
        b->a.n_walls = nw; b->a.n_bvh = n_bvh; b->a.walls = dw; b->a.wall_order = dwo; b->a.bvh = dbvh;
        BCKB(balloc(b, &b->a.plan, N * lcbatch::WP_CAP)); BCKB(balloc(b, &b->a.n_plan, N)); BCKB(balloc(b, &b->a.wnd, N));
        BCKB(balloc(b, &b->a.lost, N)); BCKB(balloc(b, &b->a.rng0, (size_t)S)); BCKB(balloc(b, &b->a.frame_target, (size_t)S));
        b->h_rng0.assign(d->rng_state, d->rng_state + S);
        BCKB(cudaMemcpy(b->a.rng0, d->rng_state, sizeof(uint32_t) * (size_t)S, cudaMemcpyHostToDevice));
        b->h_goal.resize((size_t)2 * N);
    }
    // initial component values: demo.cpp:563-581 create_robo, ai.cpp:107-127 ai::init, physics.cpp:15-21 prewarm
    b->h_pv.resize(N); b->h_c4.resize(N); b->h_goal2.resize(N); b->h_facing.resize(N); b->h_meta.resize(N);
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
            b->h_goal2[g] = make_float2(gx, gy);
            if (b->n_walls > 0) {
                b->h_goal[2 * g] = gx; b->h_goal[2 * g + 1] = gy;
                mtime = g_min_time[(size_t)b->group_of[s] * n + i];  // length of the planned path instead of the bee line
            }
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

static int32_t batch_run_impl(lc_batch* b, int32_t max_frames, float* avg_velocity, float* group_time, int32_t* n_done,
                              int32_t* frames) {
    BATCH_ENTER(b);
    if (max_frames < 0) return bfail(LC_ERR_ARG, "lc_batch_run: max_frames < 0");
    const int nt = (b->a.n + 31) & ~31;
    const size_t smem = sizeof(float4) * 5 * nt + sizeof(int) * lcbatch::MAX_NEAR * nt;
    const int S = b->a.S, n = b->a.n;
    if (b->n_walls == 0) {
        BCK(cudaEventRecord(b->ev0, b->st));
        lcbatch::k_batch_rollout<false><<<S, nt, smem, b->st>>>(b->a, b->k, max_frames);
        BCK(cudaEventRecord(b->ev1, b->st));
        b->launches += 1;
        BCK(cudaMemcpyAsync(b->h_ctl, b->a.ctl, sizeof(lcbatch::ScenarioCtl) * (size_t)S, cudaMemcpyDeviceToHost, b->st));
        BCK(cudaStreamSynchronize(b->st));
        BCK(cudaGetLastError());
        BCK(cudaEventElapsedTime(&b->last_ms, b->ev0, b->ev1));
    } else {
        // Walled batch: every block runs until its scenario is done, reaches its frame budget, or stops in front of a frame
        // whose replan gate fired; the host plans for the waiting scenarios (each on its own copy of the roadmap, created at
        // its first replan) and launches again.  Device time = sum over the launches.
        BCK(cudaMemcpyAsync(b->h_ctl, b->a.ctl, sizeof(lcbatch::ScenarioCtl) * (size_t)S, cudaMemcpyDeviceToHost, b->st));
        BCK(cudaStreamSynchronize(b->st));
        std::vector<uint32_t> target((size_t)S);
        for (int s = 0; s < S; ++s) {
            const uint64_t t = (uint64_t)b->h_ctl[s].frame + (uint64_t)max_frames;
            target[s] = t > 0xffffffffull ? 0xffffffffu : (uint32_t)t;
        }
        BCK(cudaMemcpyAsync(b->a.frame_target, target.data(), sizeof(uint32_t) * (size_t)S, cudaMemcpyHostToDevice, b->st));
        b->last_ms = 0.f;
        std::vector<lcbatch::ReplanUpdate> ups;
        for (int round = 0;; ++round) {
            BCK(cudaEventRecord(b->ev0, b->st));
            lcbatch::k_batch_rollout<true><<<S, nt, smem, b->st>>>(b->a, b->k, max_frames);
            BCK(cudaEventRecord(b->ev1, b->st));
            b->launches += 1;
            BCK(cudaMemcpyAsync(b->h_ctl, b->a.ctl, sizeof(lcbatch::ScenarioCtl) * (size_t)S, cudaMemcpyDeviceToHost, b->st));
            BCK(cudaStreamSynchronize(b->st));
            BCK(cudaGetLastError());
            float ms = 0.f;
            BCK(cudaEventElapsedTime(&ms, b->ev0, b->ev1));
            b->last_ms += ms;
            ups.clear();
            for (int s = 0; s < S; ++s) {
                const lcbatch::ScenarioCtl& c = b->h_ctl[s];
                if (!c.pause || c.finished || c.frame >= target[s]) continue;
                if (!b->planner[s] && lc_planner_clone(b->tmpl[(size_t)b->group_of[s]], &b->planner[s]) != LC_OK) return LC_ERR_CAPACITY;
                const int ids[2] = {c.wr1, c.wr2};
                for (int q = 0; q < 2; ++q) {
                    if (ids[q] < 0 || ids[q] >= n) continue;
                    lcbatch::ReplanUpdate u;
                    u.scenario = s; u.agent = ids[q]; u.pad = 0;
                    int32_t k = 0;
                    const int rc = lc_planner_plan(b->planner[s], &c.req[2 * q], &b->h_goal[2 * ((size_t)s * n + ids[q])], lcbatch::WP_CAP, u.wp, &k, nullptr);
                    if (rc) return bfail(rc, "lc_batch_run: a replan of scenario %d produced %d waypoints (capacity %d)", s, k, lcbatch::WP_CAP);
                    u.n_waypoints = k;
                    ups.push_back(u);
                }
            }
            if (ups.empty()) break;
            b->replans += (long long)ups.size();
            if (ups.size() > b->updates_cap) {
                if (b->d_updates) cudaFree(b->d_updates);
                b->updates_cap = ups.size() * 2;
                BCK(cudaMalloc((void**)&b->d_updates, sizeof(lcbatch::ReplanUpdate) * b->updates_cap));
            }
            BCK(cudaMemcpyAsync(b->d_updates, ups.data(), sizeof(lcbatch::ReplanUpdate) * ups.size(), cudaMemcpyHostToDevice, b->st));
            lcbatch::k_batch_apply_replans<<<(unsigned)((ups.size() + 127) / 128), 128, 0, b->st>>>(b->a, b->d_updates, (int)ups.size());
            b->launches += 1;
            BCK(cudaStreamSynchronize(b->st));  // `ups` is the copy source
        }
    }
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

// no exception crosses the C ABI (learn_comms.h): allocation failures of the host containers become LC_ERR_CAPACITY
int32_t lc_batch_create(const lc_batch_desc* d, lc_batch** out) {
    try {
        return batch_create_impl(d, out);
    } catch (...) {
        if (out) *out = nullptr;
        return bfail(LC_ERR_CAPACITY, "lc_batch_create: host allocation failed");
    }
}
int32_t lc_batch_run(lc_batch* b, int32_t max_frames, float* avg_velocity, float* group_time, int32_t* n_done, int32_t* frames) {
    try {
        return batch_run_impl(b, max_frames, avg_velocity, group_time, n_done, frames);
    } catch (...) {
        return bfail(LC_ERR_CAPACITY, "lc_batch_run: host allocation failed while serving replans");
    }
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
        if (b->n_walls > 0) {  // walled: Agent::num_done counts the waypoints passed
            std::vector<int> wnd(N);
            BCK(cudaMemcpy(wnd.data(), b->a.wnd, sizeof(int) * N, cudaMemcpyDeviceToHost));
            for (size_t g = 0; g < N; ++g) num_done[b->ext(g)] = wnd[g];
        }
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

int32_t lc_batch_get_walled(lc_batch* b, int32_t* n_plan, int32_t* lost_frames, int64_t* wall_hits, int64_t* replans) {
    BATCH_ENTER(b);
    if (b->n_walls == 0) return bfail(LC_ERR_ARG, "lc_batch_get_walled: not a walled batch");
    const size_t N = (size_t)b->a.S * b->a.n;
    BCK(cudaStreamSynchronize(b->st));
    std::vector<int> tmp(N);
    for (int t = 0; t < 2; ++t) {
        int32_t* dst = t == 0 ? n_plan : lost_frames;
        if (!dst) continue;
        BCK(cudaMemcpy(tmp.data(), t == 0 ? b->a.n_plan : b->a.lost, sizeof(int) * N, cudaMemcpyDeviceToHost));
        for (size_t g = 0; g < N; ++g) dst[b->ext(g)] = tmp[g];
    }
    if (wall_hits) {
        BCK(cudaMemcpy(b->h_ctl, b->a.ctl, sizeof(lcbatch::ScenarioCtl) * (size_t)b->a.S, cudaMemcpyDeviceToHost));
        for (int s = 0; s < b->a.S; ++s) wall_hits[s] = (int64_t)b->h_ctl[s].wall_hits;
    }
    if (replans) *replans = b->replans;
    return LC_OK;
}

int32_t lc_batch_timer(lc_batch* b, float* milliseconds) {
    BATCH_ENTER(b);
    if (milliseconds) *milliseconds = b->last_ms;
    return LC_OK;
}

int64_t lc_batch_kernel_launches(lc_batch* b) { return b ? b->launches : 0; }

}  // extern "C"
