// Single-crowd frame kernels for sm_100a.  One frame of the reference loop (enginee/src/main.cpp:98-150)
//   S0 demo::run -> S1 comm::run -> S2 ai::update_agents (LMP::calc_sum_force) -> S3 physics::simulate -> S4 all_sync
// is eight launches (k_refine on a side stream, beside k_pairs):
//   k_frame_begin  S0 scoring + latch; re-sort this frame or keep the slots (skin)? grid / roles of the two state copies (1 block)
//   k_bin          re-sort frames: cell id + rank per agent (counting sort, phase 1); directional extremes  (n threads)
//   k_scan_hull    re-sort frames: exclusive scan of the cell histogram; candidate-pruning polygon          (tiles + 1 block)
//   k_reorder      re-sort frames: scatter the whole SoA state into cell order, stage-1 candidates          (n threads)
//   k_refine       second pruning stage of the comm-partner candidates (64 - 256 directions)                (1 block)
//   k_pairs        S2a: pair filter over the forward half of the disc-shaped cell stencil, every unordered pair
//                  once; survivors go into both agents' pair lists                                  (n threads)
//   k_force        S1 + S2b + S3a fused: partner arg-max, message update, goal + TTC (+ wall) forces from the pair list,
//                  lookahead, leap-frog integrate; emits near-lists for the collision pass          (n threads)
//   k_collide      S3b collision push-out (Jacobi; walls through the static BVH) + avg_velocity + next frame's
//                  reductions, stage-1 candidates, replan gate                                      (n threads)
// Seven frames in eight keep the slot order of the last re-sort (k_bin / k_scan_hull / k_reorder return at once).
// FP32 CUDA cores only: the path is a gather / pair-test loop, not a dense contraction.
#include "crowd.cuh"
#include "lc_math.cuh"
#include "walls.cuh"
#include <math_constants.h>

namespace lc {
using namespace lcm;

// roles of the two state copies in the frame in flight (crowd.cuh Buffers)
__device__ __forceinline__ const AgentArrays& set_in(const Buffers& b) { return b.set[b.ctl->in_set]; }       // frame-start state, read
__device__ __forceinline__ const AgentArrays& set_out(const Buffers& b) { return b.set[b.ctl->in_set ^ 1]; }  // this frame's result, written
__device__ __forceinline__ const AgentArrays& set_src(const Buffers& b) { return b.set[b.ctl->src_set]; }     // last frame's result (input of a re-sort)
__device__ __forceinline__ const AgentArrays& set_live(const Buffers& b) { return b.set[b.ctl->live]; }       // state after the last executed frame

// Domain mode: the input of a frame's counting sort is (a) last frame's local slots, read in place from `cur` — an
// owned agent that stays on this rank is kept, one that left or was a ghost is skipped — followed by (b) the records the
// left and (c) the right neighbour's k_collide stored into this rank's mailbox (migrants and ghosts).  No unpack copy and
// no self copy: only boundary agents ever travel.
struct DDRec {
    const float4* base;  // plane 0 of the region; plane p is at base + p * cap
    int cap, idx;
    __device__ __forceinline__ float4 plane(int p) const { return base[(size_t)p * cap + idx]; }
};
// i >= recv[1]: which neighbour record input index i is
__device__ __forceinline__ DDRec dd_locate(const Params* P, const Buffers& b, int i) {
    const DomainDev* dd = b.dd;
    const int par = b.ctl->frame & 1u;  // k_dd_sync_begin already advanced the frame counter
    const int n_prev = dd->recv[1], n_l = dd->recv[2];
    int reg;
    DDRec r;
    if (i < n_prev + n_l) { reg = 2; r.idx = i - n_prev; }   // from the left neighbour
    else { reg = 0; r.idx = i - n_prev - n_l; }              // from the right neighbour
    r.cap = dd->lay.cap_halo;
    r.base = reinterpret_cast<const float4*>(dd->mbox[P->dd_rank] + dd->lay.region[par][reg]);
    return r;
}
// position of input record i; false = the slot is not part of this frame's local set
__device__ __forceinline__ bool input_pos(const Params* P, const Buffers& b, int i, float2& p) {
    if (b.dd) {
        if (i >= b.dd->recv[1]) { const float4 r0 = dd_locate(P, b, i).plane(0); p = make_float2(r0.x, r0.y); return true; }
        if (!b.dd->keep[i]) return false;
    }
    const float4 q = set_src(b).pvf[i];
    p = make_float2(q.x, q.y);
    return true;
}

__constant__ float c_dirx[HULL_K];
__constant__ float c_diry[HULL_K];
__constant__ float c_dir2x[HULL_K2];
__constant__ float c_dir2y[HULL_K2];

// Programmatic dependent launch (opt-in, LC_PDL=1): every frame kernel is launched with
// programmaticStreamSerializationAllowed, lets its successor's blocks take the slots its own blocks free up
// (launch_dependents), and touches no memory before its predecessor is complete and flushed (wait).  Meant to hide the
// launch gaps between the kernels of a frame; inside the 16-frame graph it measured SLOWER (plaza125k 75.0 vs 64.0 us
// per frame, plaza1m 279 vs 278 us), so it is off by default.  Without the launch attribute both instructions are no-ops.
__device__ __forceinline__ void stamp(Ctl* ctl, int which) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    ctl->ts[which] = t;
}
// entry stamp of a frame kernel (thread 0 of block 0)
__device__ __forceinline__ void stamp_entry(Ctl* ctl, int which) {
    if (blockIdx.x == 0 && threadIdx.x == 0) stamp(ctl, which);
}
__device__ __forceinline__ void pdl_enter() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

// demo.cpp:655-663 update_runs (Welford, all float)
__device__ __forceinline__ void update_runs(Ctl* ctl, float overhead, float* fin_log) {
    float ftotal = (float)(++ctl->total_num_done);
    float delta = __fsub_rn(overhead, ctl->run_avg);
    ctl->run_avg = __fadd_rn(ctl->run_avg, __fdiv_rn(delta, ftotal));
    float delta2 = __fsub_rn(overhead, ctl->run_avg);
    ctl->M2 = __fadd_rn(ctl->M2, __fmul_rn(delta, delta2));
    ctl->run_std = __fsqrt_rn(__fdiv_rn(ctl->M2, ftotal));
    if (fin_log) {  // demo.cpp:662: the reference prints "t:%f u:%f s:%f" per finisher; kept for the drop-in's stdout
        float* l = fin_log + 3 * (size_t)(ctl->total_num_done - 1);
        l[0] = overhead; l[1] = ctl->run_avg; l[2] = ctl->run_std;
    }
}

// physics.cpp:72-76 with vel_count == 0 forever (quirk Q6): avg = (avg*0 + |vel|) / (0 + 1) per not-done agent in
// table order, i.e. the speed of the last one unless a non-finite value poisoned the chain.
__device__ __forceinline__ void fold_avg_velocity(Ctl* ctl) {
    unsigned long long lv = ctl->lastvel;
    if (lv != 0ull) {
        float sp = __uint_as_float((uint32_t)(lv & 0xffffffffull));
        int others = ctl->nonfinite - ((sp <= LC_FLT_MAX) ? 0 : 1);
        bool poisoned = others > 0 || !(fabsf(ctl->avg_velocity) <= LC_FLT_MAX);
        ctl->avg_velocity = poisoned ? CUDART_NAN_F : sp;
        ctl->lastvel = 0ull;
        ctl->nonfinite = 0;
    }
}

// Fold the per-block partials written by k_collide / k_init_reduce into Ctl.  Called by every thread of ONE block.
__device__ void consume_block_reductions(Ctl* ctl, const Buffers& b) {
    __shared__ uint32_t s_r[5], s_dm;
    __shared__ unsigned long long s_lv;
    const int nb = ctl->n_blkred;
    if (threadIdx.x == 0) { s_r[0] = s_r[1] = 0xffffffffu; s_r[2] = s_r[3] = s_r[4] = 0u; s_lv = 0ull; s_dm = 0u; }
    __syncthreads();
    if (nb > 0) {
        uint32_t mnx = 0xffffffffu, mny = 0xffffffffu, mxx = 0u, mxy = 0u, sp = 0u, dm = 0u;
        unsigned long long lv = 0ull;
        for (int i = threadIdx.x; i < nb; i += blockDim.x) {
            const BlockRed r = b.blkred[i];
            mnx = min(mnx, r.minx); mny = min(mny, r.miny); mxx = max(mxx, r.maxx); mxy = max(mxy, r.maxy);
            sp = max(sp, r.vmax);
            dm = max(dm, r.dmax);
            lv = r.lastvel > lv ? r.lastvel : lv;
        }
        dm = __reduce_max_sync(0xffffffffu, dm);
        if ((threadIdx.x & 31) == 0 && dm) atomicMax(&s_dm, dm);
        // warp first: 1024 same-address shared atomics per value serialise (they were most of this kernel's time)
        mnx = __reduce_min_sync(0xffffffffu, mnx); mny = __reduce_min_sync(0xffffffffu, mny);
        mxx = __reduce_max_sync(0xffffffffu, mxx); mxy = __reduce_max_sync(0xffffffffu, mxy);
        sp = __reduce_max_sync(0xffffffffu, sp);
        const uint32_t hi = __reduce_max_sync(0xffffffffu, (uint32_t)(lv >> 32));
        if ((threadIdx.x & 31) == 0) {
            atomicMin(&s_r[0], mnx); atomicMin(&s_r[1], mny);
            atomicMax(&s_r[2], mxx); atomicMax(&s_r[3], mxy); atomicMax(&s_r[4], sp);
        }
        if (lv && (uint32_t)(lv >> 32) == hi) atomicMax(&s_lv, lv);  // ids are unique: one lane per warp
    }
    __syncthreads();
    if (threadIdx.x == 0 && nb > 0) {
        ctl->minx = min(ctl->minx, s_r[0]); ctl->miny = min(ctl->miny, s_r[1]);
        ctl->maxx = max(ctl->maxx, s_r[2]); ctl->maxy = max(ctl->maxy, s_r[3]);
        ctl->vmax = max(ctl->vmax, s_r[4]);
        ctl->dmax = max(ctl->dmax, s_dm);
        ctl->lastvel = s_lv > ctl->lastvel ? s_lv : ctl->lastvel;
        ctl->n_blkred = 0;
    }
    __syncthreads();
}

__device__ void grid_from_reductions(const Params* P, Ctl* ctl) {
    float minx = ord2f(ctl->minx), miny = ord2f(ctl->miny), maxx = ord2f(ctl->maxx), maxy = ord2f(ctl->maxy);
    float vmax = __uint_as_float(ctl->vmax);
    if (!(vmax <= 1.0e3f)) { vmax = 1.0e3f; if (!ctl->error) ctl->error = -3; }
    if (!(maxx >= minx) || !(maxy >= miny) || !(maxx - minx <= 1e9f) || !(maxy - miny <= 1e9f)) {
        minx = miny = 0.f; maxx = maxy = 1.f;
        if (!ctl->error) ctl->error = -3;
    }
    // ttc <= 5 implies |m| <= R + 5*|dv| <= R + 10*vmax (LMP.cpp:297-300); slack covers float rounding
    float cutoff = P->R + 10.f * vmax * 1.001f + 0.01f;
    // the same cell block also has to contain every agent that can be in contact after this frame's move (near list
    // of the collision pass): both move at most dt, so the reach is R + 2 dt (+ slack), which exceeds the TTC cutoff
    // when the crowd is (almost) at rest
    cutoff = fmaxf(cutoff, P->near_reach);
    int st = P->stencil_req < 1 ? 1 : P->stencil_req;
    // Cells are CELL_XSUB times narrower in x than in y: a row of the cell block is one contiguous slot range whatever the
    // cell width, so narrow cells trim the overshoot at both ends of every row without adding rows (= range setups).
    float h = cutoff / (float)st, hx = h / (float)CELL_XSUB;
    float ext = fmaxf(maxx - minx, maxy - miny);
    int nx, ny;
    for (;;) {
        nx = (int)floorf((maxx - minx) / hx) + 1;
        ny = (int)floorf((maxy - miny) / h) + 1;
        if ((long long)nx * ny <= CELL_CAP && nx <= 65535 && ny <= 65535) break;
        if (hx < h) hx *= 2.f; else { h *= 2.f; hx = h; }
    }
    ctl->x0 = minx; ctl->y0 = miny; ctl->h = h; ctl->hx = hx; ctl->nx = nx; ctl->ny = ny; ctl->ncells = nx * ny;
    ctl->cutoff = cutoff;
    ctl->stencil = (int)ceilf(cutoff / h);
    ctl->stencil_x = (int)ceilf(cutoff / hx);
    ctl->cx = 0.5f * (minx + maxx); ctl->cy = 0.5f * (miny + maxy);
    ctl->delta = 1e-5f * ext + 1e-4f;
}

// S0 proper: demo::run (demo.cpp:665-702) for frame `frame` on a (shared-memory) copy of the control block — the agents
// that became done in the last frame are counted in table order (Welford), the time-out penalty folds every agent that
// is still walking, the all_done latch is set.  cnt = done-events waiting in b.events, nnd = agents not done.  Called by
// every thread of ONE block of 1024 threads.
__device__ void demo_run_score(const Params* __restrict__ P, const Buffers& b, Ctl* ctl, uint32_t frame, int cnt, int nnd, bool dd,
                               int2* s_ev, float* s_ft, int* s_flag) {
    const int tid = threadIdx.x;
    const int n = P->n_ids;
    const float fake_time = __fmul_rn(LC_DT, (float)frame);
    // newly done agents, table order (demo.cpp:669-676)
    if (cnt > 0 && cnt <= EV_SORT_CAP) {
        int m = 1;
        while (m < cnt) m <<= 1;
        for (int i = tid; i < m; i += blockDim.x) s_ev[i] = i < cnt ? b.events[i] : make_int2(0x7fffffff, 0);
        __syncthreads();
        for (int k = 2; k <= m; k <<= 1)
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int i = tid; i < m; i += blockDim.x) {
                    int ixj = i ^ j;
                    if (ixj > i) {
                        int2 a = s_ev[i], c = s_ev[ixj];
                        bool up = (i & k) == 0;
                        if ((a.x > c.x) == up) { s_ev[i] = c; s_ev[ixj] = a; }
                    }
                }
                __syncthreads();
            }
        if (tid == 0)
            for (int i = 0; i < cnt; ++i) update_runs(ctl, __fsub_rn(fake_time, __int_as_float(s_ev[i].y)), b.fin_log);
    } else if (cnt > EV_SORT_CAP) {
        for (int i = tid; i < n; i += blockDim.x) b.ev_by_id[i] = CUDART_NAN_F;
        __syncthreads();
        for (int i = tid; i < cnt; i += blockDim.x) b.ev_by_id[b.events[i].x] = __int_as_float(b.events[i].y);
        __syncthreads();
        for (int base = 0; base < n; base += blockDim.x) {
            int i = base + tid;
            float v = i < n ? b.ev_by_id[i] : CUDART_NAN_F;
            s_ft[tid] = v;
            s_flag[tid] = (v == v);
            __syncthreads();
            if (tid == 0)
                for (int k = 0; k < blockDim.x; ++k)
                    if (s_flag[k]) update_runs(ctl, __fsub_rn(fake_time, s_ft[k]), b.fin_log);
            __syncthreads();
        }
    }
    __syncthreads();
    bool all_done = nnd == 0;
    if (fake_time > P->max_duration && dd) {
        // domain mode: k_dd_pack of every rank scattered the finish times of its owned, not-done agents into the
        // scoring rank's id-indexed table (NaN = none); fold them in id order there.  Other ranks only latch.
        if (P->dd_rank == 0) {
            float* tab = reinterpret_cast<float*>(b.dd->mbox[0] + b.dd->lay.ev_by_id);
            for (int base = 0; base < n; base += blockDim.x) {
                const int id = base + tid;
                const float v = id < n ? tab[id] : CUDART_NAN_F;
                s_ft[tid] = v;
                s_flag[tid] = (v == v);
                __syncthreads();
                if (tid == 0)
                    for (int k = 0; k < blockDim.x; ++k)
                        if (s_flag[k]) update_runs(ctl, s_ft[k], b.fin_log);
                __syncthreads();
            }
        }
        all_done = true;
    } else if (fake_time > P->max_duration) {  // demo.cpp:691-700
        for (int base = 0; base < n; base += blockDim.x) {
            int id = base + tid;
            int flag = 0;
            float ft = 0.f;
            if (id < n) {
                int s = b.slot_of_id[id];
                const AgentArrays& last = b.set[ctl->live];
                    int4 mt = last.meta[s];
                if (!(mt.y & 1)) {
                    float2 g = last.goal[s]; const float4 pf = last.pvf[s]; const float2 p = make_float2(pf.x, pf.y);
                    float l = len2(__fsub_rn(g.x, p.x), __fsub_rn(g.y, p.y));
                    ft = __fsub_rn(__fadd_rn(__fmul_rn(2.f, l), fake_time), __int_as_float(mt.w));
                    flag = 1;
                }
            }
            s_ft[tid] = ft;
            s_flag[tid] = flag;
            __syncthreads();
            if (tid == 0)
                for (int k = 0; k < blockDim.x; ++k)
                    if (s_flag[k]) update_runs(ctl, s_ft[k], b.fin_log);
            __syncthreads();
        }
        all_done = true;
    }
    if (tid == 0 && all_done) ctl->finished = 1;
}

// ---------------------------------------------------------------------------------------------------------
// S0: demo::run (demo.cpp:665-702)
__device__ __forceinline__ void frame_begin_body(const Params* __restrict__ P, const Buffers& b) {
    // The control block is worked on in shared memory and written back at the end: thread 0's serial part (Welford
    // updates, grid set-up, resets) is a chain of ~60 dependent accesses, each an L2 round trip on the global copy.
    // Nothing else runs on the stream while this kernel does (k_refine's side stream joined before k_force).
    Ctl* const gctl = b.ctl;
    __shared__ Ctl s_ctl;
    __shared__ int2 s_ev[EV_SORT_CAP];
    __shared__ float s_ft[1024];
    __shared__ int s_flag[1024];
    const int tid = threadIdx.x;
    static_assert(sizeof(Ctl) % 4 == 0, "Ctl is copied word-wise");
    for (int i = tid; i < (int)(sizeof(Ctl) / 4); i += blockDim.x)
        reinterpret_cast<uint32_t*>(&s_ctl)[i] = reinterpret_cast<const uint32_t*>(gctl)[i];
    __syncthreads();
    Ctl* ctl = &s_ctl;
    if (tid == 0) stamp(ctl, TS_BEGIN);
    const int n = P->n_ids;  // everything below that walks "all agents" is indexed by global id
    const bool dd = P->dd_world > 1;
    consume_block_reductions(ctl, b);
    if (tid == 0) fold_avg_velocity(ctl);
    __syncthreads();
    const int was_finished = ctl->finished;
    const int cnt = ctl->ev_count;
    const int nnd = ctl->nnd_acc;
    const uint32_t frame = ctl->frame + 1;
    __syncthreads();
    if (was_finished && !(P->flags & 2u)) {
        if (tid == 0) gctl->halt = 1;  // nothing else of the shared copy is kept: the block partials stay queued for lc_get_scores
        return;
    }
    if (P->n_walls > 0 && ctl->wr1 != 0x7fffffff && !ctl->w_served) {
        // walled scene: some agent needs GMP::plan_one before this frame's update_agents (ai.cpp:162-168).  The planner
        // lives on the host (planner.cpp): stop here — this and every later frame of the graph replay is a no-op and the
        // frame counter does not advance — until lc_step has uploaded the new waypoint lists.
        if (tid == 0) gctl->halt = 2;
        return;
    }
    if (!was_finished) demo_run_score(P, b, ctl, frame, cnt, nnd, dd, s_ev, s_ft, s_flag);
    if (tid == 0) {
        ctl->frame = frame;
        ctl->ev_count = 0;
        ctl->n_not_done = nnd;
        ctl->nnd_acc = 0;
        const float vm = __uint_as_float(ctl->vmax);
        {   // re-sort this frame, or keep last re-sort's slot order and search a skin wider (crowd.cuh)
            const float moved = ctl->skin + __uint_as_float(ctl->dmax);   // NaN / inf (a non-finite state) forces the re-sort
            // (the candidate list k_collide compiled for this frame allowed for CAND_MARGIN of movement in its frame)
            const bool rebin = P->rebin_every <= 1 || ctl->force_rebin || ctl->since + 1 >= P->rebin_every || !(moved <= 0.25f * ctl->h) ||
                               !(__uint_as_float(ctl->dmax) <= CAND_MARGIN);
            if (rebin) {
                grid_from_reductions(P, ctl);
                ctl->skin = 0.f;
                ctl->since = 0;
            } else {
                // same cells, same slots; the cutoff follows this frame's speeds as it does on a re-sort frame
                float vmax = vm;
                if (!(vmax <= 1.0e3f)) { vmax = 1.0e3f; if (!ctl->error) ctl->error = -3; }
                ctl->cutoff = fmaxf(P->R + 10.f * vmax * 1.001f + 0.01f, P->near_reach);
                ctl->skin = moved;
                ctl->since += 1;
                ctl->stencil = (int)ceilf((ctl->cutoff + 2.f * moved) / ctl->h);     // clamps of the rare full-block walks
                ctl->stencil_x = (int)ceilf((ctl->cutoff + 2.f * moved) / ctl->hx);
            }
            ctl->rebin = rebin;
            ctl->n_resorts += rebin ? 1ull : 0ull;
            ctl->force_rebin = 0;
            ctl->dmax = 0u;
            // roles of the two state copies (crowd.cuh Buffers): a re-sort scatters last frame's result into the other copy
            // and the frame reads that; otherwise last frame's result is read in place
            const int last = ctl->live;
            ctl->src_set = last;
            ctl->in_set = rebin ? last ^ 1 : last;
            ctl->live = ctl->in_set ^ 1;
        }
        ctl->minx = ctl->miny = 0xffffffffu;
        ctl->maxx = ctl->maxy = 0u;
        ctl->vmax = 0u;
        ctl->n_cand = ctl->rebin ? 0 : ctl->n_cand_next;   // re-sort frames: k_reorder compiles the list
        ctl->n_cand_next = 0;
        {   // k_bin's gate radius from last frame's directional extremes (0.1 m covers the agents' and the centre's motion)
            float rho = CUDART_INF_F;
            for (int k = 0; k < HULL_K; ++k) rho = fminf(rho, ctl->ext_val[k]);
            rho -= 0.1f;
            ctl->gate_rho2 = rho > 0.f ? rho * rho : -1.f;
        }
        if (!dd) {  // the gate k_collide evaluated on the positions this frame starts from (domain mode: dd_global_body merges it)
            ctl->replan1 = ctl->replan1n; ctl->replan2 = ctl->replan2n;
            ctl->replan1n = ctl->replan2n = 0x7fffffff;
        }
        if (dd) {  // the ghost band was sized for the speed bound (domain.cuh); a wider cutoff would miss neighbours
            if (!(P->R + 10.f * vm * 1.001f + 0.01f + P->near_reach <= P->dd_halo) && !ctl->error) ctl->error = -5;
        }
        ctl->overlap2 = 0ull;
        ctl->coll_hits = 0ull;
        if (P->n_walls > 0) {  // this frame's k_collide collects the next frame's replan requests
            ctl->wr1 = ctl->wr2 = 0x7fffffff;
            ctl->w_served = 0;
            ctl->draws += (unsigned long long)ctl->drawn_last;  // plan-less agents do not draw (LMP.cpp:267-270)
            ctl->drawn_last = ctl->n_planned;
        }
    }
    if (tid < HULL_K) ctl->ext[tid] = 0ull;
    __syncthreads();
    for (int i = tid; i < (int)(sizeof(Ctl) / 4); i += blockDim.x)
        reinterpret_cast<uint32_t*>(gctl)[i] = reinterpret_cast<const uint32_t*>(&s_ctl)[i];
}

__global__ void __launch_bounds__(1024) k_frame_begin(const Params* __restrict__ P, Buffers b) {
    pdl_enter();
    frame_begin_body(P, b);
}

// ---------------------------------------------------------------------------------------------------------
// Uniform grid, phase 1: cell id + arrival rank per agent; directional extremes for candidate pruning.
__device__ __forceinline__ int cell_of(const Ctl* ctl, float x, float y) {
    int ix = (int)floorf((x - ctl->x0) / ctl->hx), iy = (int)floorf((y - ctl->y0) / ctl->h);
    ix = min(max(ix, 0), ctl->nx - 1);
    iy = min(max(iy, 0), ctl->ny - 1);
    return iy * ctl->nx + ix;
}

__global__ void __launch_bounds__(1024) k_bin(const Params* __restrict__ P, Buffers b) {
    pdl_enter();
    const Ctl* ctl = b.ctl;
    if (ctl->halt) return;
    stamp_entry(b.ctl, TS_BIN);
    if (!ctl->rebin) return;  // slot order and cell_start of the last re-sort stay in force
    const int n = P->n_in;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned lane = threadIdx.x & 31;
    float2 p = make_float2(0.f, 0.f);
    int cell = -1;
    bool present = false;
    if (i < n) {
        present = input_pos(P, b, i, p);
        if (present) cell = cell_of(ctl, p.x, p.y);
    }
    {   // Counting sort, phase 1.  The agents arrive in last frame's cell order, so equal cells form runs of
        // consecutive lanes: one atomicAdd per run instead of one per agent (same-address atomics serialise in L2).
        const int prev = __shfl_up_sync(0xffffffffu, cell, 1);
        const bool head = lane == 0 || prev != cell;
        const unsigned heads = __ballot_sync(0xffffffffu, head);
        const unsigned below = heads & (0xffffffffu >> (31 - lane));   // heads at or below this lane
        const int my_head = 31 - __clz(below);
        const unsigned above = lane == 31 ? 0u : (heads >> (lane + 1)); // heads above this lane
        const int run_end = above ? (int)lane + __ffs(above) : 32;      // exclusive, valid on the head lane
        int base = 0;
        if (head && cell >= 0) base = atomicAdd(&b.cell_count[cell], run_end - (int)lane);
        base = __shfl_sync(0xffffffffu, base, my_head);
        if (i < n) b.cellrank[i] = make_int2(cell, base + ((int)lane - my_head));
    }
    if (P->flags & 1u) return;  // no comm: no partner search
    // directional extremes of this block (stage-1 polygon for the comm-partner candidates); any maximiser is a valid vertex
    __shared__ unsigned long long s_key[32][HULL_K];
    const float rx = p.x - ctl->cx, ry = p.y - ctl->cy;
    // Gate: last frame's extreme in direction k had projection ext_val[k]; agents move a few cm per frame, so only agents
    // at least min_k ext_val[k] - margin away from the centre can be this frame's extremes.  Warps whose agents are all
    // inside that circle (3/4 of a square crowd) skip the 16 warp reductions.  If the guess is ever wrong the polygon just
    // gets a vertex that is not extreme, or an edge less — any polygon through agents keeps the pruning exact.
    const bool ok = present && rx == rx && ry == ry && (rx * rx + ry * ry >= ctl->gate_rho2);
    if (__any_sync(0xffffffffu, ok)) {
#pragma unroll
        for (int k = 0; k < HULL_K; ++k) {
            const float v = rx * c_dirx[k] + ry * c_diry[k];
            const uint32_t ov = ok ? f2ord(v) : 0u;
            const uint32_t best = __reduce_max_sync(0xffffffffu, ov);
            const unsigned who = __ballot_sync(0xffffffffu, ov == best);
            if (lane == (unsigned)(__ffs(who) - 1))
                s_key[threadIdx.x >> 5][k] = best ? (((unsigned long long)best << 32) | (uint32_t)i) : 0ull;
        }
    } else if (lane < HULL_K) {
        s_key[threadIdx.x >> 5][lane] = 0ull;
    }
    __syncthreads();
    if (threadIdx.x < HULL_K) {
        unsigned long long m = 0ull;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) {
            const unsigned long long v = s_key[w][threadIdx.x];
            m = v > m ? v : m;
        }
        b.blkext[(size_t)blockIdx.x * HULL_K + threadIdx.x] = m;
    }
}

// Exclusive scan of the cell histogram: one 4096-cell tile per block, tile offsets by look-back over the aggregates the
// earlier tiles publish (a tile only waits for lower-numbered tiles, and blocks are dispatched in index order, so the
// spin cannot deadlock).
// The extra last block folds k_bin's per-block directional extremes and builds the stage-1 pruning polygon.
__global__ void __launch_bounds__(1024) k_scan_hull(const Params* __restrict__ P, Buffers b) {
    pdl_enter();
    Ctl* ctl = b.ctl;
    if (ctl->halt) return;
    stamp_entry(ctl, TS_SCAN);
    if (!ctl->rebin) return;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (blockIdx.x == gridDim.x - 1) {
        if (P->flags & 1u) return;
        __shared__ unsigned long long s_ext[HULL_K];
        if (tid < HULL_K) s_ext[tid] = 0ull;
        __syncthreads();
        const int nblk = (P->n_in + 1023) / 1024;
        unsigned long long m = 0ull;
        for (int q = tid / HULL_K; q < nblk; q += 1024 / HULL_K) {
            const unsigned long long v = b.blkext[(size_t)q * HULL_K + (tid % HULL_K)];
            m = v > m ? v : m;
        }
        if (m) atomicMax(&s_ext[tid % HULL_K], m);
        __syncthreads();
        // polygon through consecutive distinct extreme agents (CCW); inward unit normals.  One thread per edge (the
        // dependent position loads and the f64 sqrt used to sit in one thread's serial loop), thread 0 compacts.
        __shared__ float s_e[HULL_K][4];
        __shared__ int s_ev[HULL_K];
        if (tid < HULL_K) {
            const int k = tid;
            ctl->ext_val[k] = s_ext[k] ? ord2f((uint32_t)(s_ext[k] >> 32)) : -CUDART_INF_F;  // next frame's k_bin gate
            int ok = 0;
            const unsigned long long ka = s_ext[k], kb = s_ext[(k + 1) % HULL_K];
            if (ka != 0ull && kb != 0ull) {
                const int a = (int)(ka & 0xffffffffull), c = (int)(kb & 0xffffffffull);
                if (a != c) {
                    float2 pa, pc;
                    input_pos(P, b, a, pa);  // extremes are present records by construction
                    input_pos(P, b, c, pc);
                    const double ux = (double)pc.x - pa.x, uy = (double)pc.y - pa.y, l = sqrt(ux * ux + uy * uy);
                    if (l > 0.0) {
                        s_e[k][0] = pa.x; s_e[k][1] = pa.y;
                        s_e[k][2] = (float)(-uy / l); s_e[k][3] = (float)(ux / l);
                        ok = 1;
                    }
                }
            }
            s_ev[k] = ok;
        }
        __syncthreads();
        if (tid == 0) {
            int ne = 0;
            for (int k = 0; k < HULL_K; ++k) {
                if (!s_ev[k]) continue;
                ctl->ex[ne] = s_e[k][0]; ctl->ey[ne] = s_e[k][1];
                ctl->enx[ne] = s_e[k][2]; ctl->eny[ne] = s_e[k][3];
                ++ne;
            }
            ctl->n_edges = ne;
            // Inner box: centre +- t (half extents of the bounding box), t the largest factor that keeps all four corners
            // inside every edge by D = delta + the most anybody can move before the next re-sort is forced (0.25 h) + the
            // margin of the early list (k_collide).  Empty when the polygon is degenerate.
            const float hxw = ctl->cx - ctl->x0, hyw = ctl->cy - ctl->y0;
            const float D = ctl->delta + 0.25f * ctl->h + 2.f * CAND_MARGIN;
            float t = ne >= 3 ? 1.f : 0.f;
            for (int e = 0; e < ne; ++e) {
                const float clear = (ctl->cx - ctl->ex[e]) * ctl->enx[e] + (ctl->cy - ctl->ey[e]) * ctl->eny[e] - D;
                const float span = hxw * fabsf(ctl->enx[e]) + hyw * fabsf(ctl->eny[e]);
                t = fminf(t, span > 0.f ? clear / span : 0.f);
            }
            t = t * 0.999f - 1e-3f;
            const bool ok = t > 0.f && hxw > 0.f && hyw > 0.f;   // (NaN anywhere: no box)
            ctl->ibx0 = ok ? ctl->cx - t * hxw : 1.f; ctl->ibx1 = ok ? ctl->cx + t * hxw : 0.f;
            ctl->iby0 = ok ? ctl->cy - t * hyw : 1.f; ctl->iby1 = ok ? ctl->cy + t * hyw : 0.f;
        }
        return;
    }
    const int nc = ctl->ncells;
    const int t = blockIdx.x;
    if (t * SCAN_TILE > nc) return;  // index nc itself must be written (== n): tiles cover [0, nc]
    __shared__ int s_warp[32];
    __shared__ int s_off;
    const unsigned long long epoch = (unsigned long long)ctl->frame << 32;
    const int base = t * SCAN_TILE + tid * 4;
    const int4 v = *reinterpret_cast<const int4*>(b.cell_count + base);  // zero beyond ncells
    // the histogram is dead once read: leave it cleared for the next frame's k_bin (cells beyond this frame's grid were
    // never touched and are still zero from lc_create) — saves a 4 MB memset node per frame
    *reinterpret_cast<int4*>(b.cell_count + base) = make_int4(0, 0, 0, 0);
    const int sum = v.x + v.y + v.z + v.w;
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
    }
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int w = s_warp[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int u = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += u;
        }
        s_warp[lane] = wi - w;
        if (lane == 31) atomicExch(&b.tile_agg[t], epoch | (unsigned long long)(uint32_t)wi);  // publish this tile's total
        // look-back: sum the totals of all earlier tiles of this frame
        int off = 0;
        for (int q = lane; q < t; q += 32) {
            unsigned long long a;
            do { a = *(volatile unsigned long long*)&b.tile_agg[q]; } while ((a & 0xffffffff00000000ull) != epoch);
            off += (int)(uint32_t)a;
        }
        off = __reduce_add_sync(0xffffffffu, off);
        if (lane == 0) s_off = off;
    }
    __syncthreads();
    const int excl = s_off + s_warp[wid] + incl - sum;
    int4 o4;
    o4.x = excl; o4.y = excl + v.x; o4.z = o4.y + v.y; o4.w = o4.z + v.z;
    *reinterpret_cast<int4*>(b.cell_start + base) = o4;
}

// Stage 1 of the comm-partner pruning: is (x, z) inside the polygon of extreme agents, by more than `delta`?  The polygon's
// vertices are where the extreme agents stood at the last re-sort; eroded by what anybody can have moved since, it still
// lies inside the hull of the current positions.  The box test first: the box lies inside the polygon eroded by
// more than any delta in use, and holds all but a few per cent of a compact crowd.  NaN-safe: a non-finite position is
// never "inside".
__device__ __forceinline__ bool candidate_inside(const Ctl* ctl, float x, float y, float delta) {
    if (x > ctl->ibx0 && x < ctl->ibx1 && y > ctl->iby0 && y < ctl->iby1) return true;
    const int ne = ctl->n_edges;
    bool inside = ne >= 3;
    for (int e = 0; e < ne && inside; ++e) {
        const float d = (x - ctl->ex[e]) * ctl->enx[e] + (y - ctl->ey[e]) * ctl->eny[e];
        if (!(d >= delta)) inside = false;
    }
    return inside;
}

// Uniform grid, phase 2 (re-sort frames only): scatter the state into cell order (counting sort) and compile the stage-1
// list of comm-partner candidates.
__device__ __forceinline__ void dd_route(const Params* P, float x, int role[3]);  // domain.cuh
__global__ void __launch_bounds__(256) k_reorder(const Params* __restrict__ P, Buffers b) {
    pdl_enter();
    Ctl* ctl = b.ctl;
    if (ctl->halt) return;
    stamp_entry(ctl, TS_REORDER);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (!ctl->rebin) {
        // the agents keep their slots, last frame's output IS this frame's input, and k_collide compiled the candidates.
        // Decomposed crowd: the neighbours' k_collide refreshed the records of our ghosts (dd_push_body); copy each over its
        // ghost's slot.  The id -> slot table holds the ghosts' slots since the re-sort.
        if (!b.dd) return;
        const DomainDev* dd = b.dd;
        const int n_l = dd->recv[2], n_r = dd->recv[0];
        if (i >= n_l + n_r) return;
        DDRec r;
        r.cap = dd->lay.cap_halo;
        r.idx = i < n_l ? i : i - n_l;
        r.base = reinterpret_cast<const float4*>(dd->mbox[P->dd_rank] + dd->lay.region[ctl->frame & 1u][i < n_l ? 2 : 0]);
        const float4 r0 = r.plane(0), r1 = r.plane(1), r2 = r.plane(2), r4 = r.plane(4);
        const int id = __float_as_int(r4.x);
        const AgentArrays& dst = set_in(b);
        const int s = (id >= 0 && id < P->n_ids) ? b.slot_of_id[id] : -1;
        if (s < 0 || s >= P->n || dst.meta[s].x != id || !(dst.meta[s].y & META_GHOST)) {  // the two ranks disagree on the frozen set
            atomicCAS(&ctl->error, 0, -5);
            return;
        }
        dst.pvf[s] = r0;
        {
            float* q = reinterpret_cast<float*>(dst.pq + 2 * (s >> 1)) + (s & 1);
            q[0] = r0.x; q[2] = r0.y; q[4] = r0.z; q[6] = r0.w;
        }
        dst.vfh[s] = make_float2(r1.x, r1.y);
        dst.goal[s] = make_float2(r1.z, r1.w);
        dst.lgoal[s] = make_float2(r2.x, r2.y);
        dst.facing[s] = make_float2(r2.z, r2.w);
        dst.c4[s] = r.plane(3);
        dst.meta[s] = make_int4(id, __float_as_int(r4.y), __float_as_int(r4.z), __float_as_int(r4.w));
        return;
    }
    const int n = P->n_in;
    if (i >= n) return;
    const AgentArrays& src = set_src(b);
    const AgentArrays& dst = set_in(b);   // this frame reads what is scattered here
    const int2 cr = b.cellrank[i];
    if (cr.x < 0) return;  // domain mode: a slot that is not part of this frame's local set (left, or was a ghost)
    const int s = b.cell_start[cr.x] + cr.y;
    b.slot_cell[s] = cr.x;
    float2 p;
    int4 mt;
    if (b.dd && i >= b.dd->recv[1]) {  // records of the neighbours' pack: (pos.xy vel.xy | vfh goal | lgoal facing | c4 | meta)
        const DDRec r = dd_locate(P, b, i);
        const float4 r0 = r.plane(0), r1 = r.plane(1), r2 = r.plane(2), r4 = r.plane(4);
        p = make_float2(r0.x, r0.y);
        mt = make_int4(__float_as_int(r4.x), __float_as_int(r4.y), __float_as_int(r4.z), __float_as_int(r4.w));
        dst.pvf[s] = r0;
        {
            float* q = reinterpret_cast<float*>(dst.pq + 2 * (s >> 1)) + (s & 1);
            q[0] = r0.x; q[2] = r0.y; q[4] = r0.z; q[6] = r0.w;
        }
        dst.vfh[s] = make_float2(r1.x, r1.y);
        dst.goal[s] = make_float2(r1.z, r1.w);
        dst.lgoal[s] = make_float2(r2.x, r2.y);
        dst.facing[s] = make_float2(r2.z, r2.w);
        dst.c4[s] = r.plane(3);
        dst.meta[s] = mt;
    } else {
        const float4 pf = src.pvf[i];
        p = make_float2(pf.x, pf.y);
        mt = src.meta[i];
        if (b.dd && b.dd->keep[i] == 2) mt.y |= META_GHOST;  // crossed into the adjacent strip last frame: its ghost from now on
        dst.pvf[s] = pf;
        {   // pair-interleaved copy for the packed pair loop: slots (2k, 2k+1) -> (x0, x1, z0, z1), (vx0, vx1, vz0, vz1)
            float* q = reinterpret_cast<float*>(dst.pq + 2 * (s >> 1)) + (s & 1);
            q[0] = pf.x; q[2] = pf.y; q[4] = pf.z; q[6] = pf.w;
        }
        dst.vfh[s] = src.vfh[i];
        dst.goal[s] = src.goal[i];
        dst.lgoal[s] = src.lgoal[i];
        dst.facing[s] = src.facing[i];
        dst.c4[s] = src.c4[i];
        dst.meta[s] = mt;
    }
    if (b.dd) {
        // The local set is now frozen until the next re-sort.  An owned agent inside a neighbour's ghost band is that
        // neighbour's ghost for the whole period (the neighbour reached the same verdict on the same position, either when it
        // routed the record to us or when it kept the agent as a ghost): give it a fixed index in the neighbour's region.
        int2 si = make_int2(-1, -1);
        if (!(mt.y & META_GHOST)) {
            int role[3];
            dd_route(P, p.x, role);
            if (role[0]) si.x = atomicAdd(&b.dd->rcount[0], 1);
            if (role[2]) si.y = atomicAdd(&b.dd->rcount[2], 1);
        } else {
            b.slot_of_id[mt.x] = s;   // where the owner's refreshes of this ghost go
        }
        b.dd->send_idx[s] = si;
    }
    // (the id -> slot table of the owned agents and the wall-free replan gate of ai.cpp:151-168 are k_collide's business:
    // both are a matter of the final state of the previous frame)
    if (!(P->flags & 1u) && !(mt.y & META_GHOST)) {  // ghosts are published as candidates by their owner (dd_global_body)
        // keep every agent that is not inside (by delta) the polygon of extreme agents: only those can be
        // the arg-max of comm.cpp:201-220 for any query (proof in DESIGN.md)
        if (!candidate_inside(ctl, p.x, p.y, ctl->delta)) {
            int k = atomicAdd(&ctl->n_cand, 1);
            b.cand[k] = make_float4(p.x, p.y, __int_as_float(mt.x), __int_as_float(s));
        }
    }
}

// Second pruning stage (one block): the stage-1 survivors (a band along the crowd's outline, ~sqrt(n) agents) are
// tested against the polygon through THEIR HULL_K2 directional extremes.  Any convex polygon whose vertices are
// agents lies inside the true hull, so the margin argument of stage 1 applies unchanged; only the
// number of survivors depends on the polygon.  Typically leaves the hull vertices plus a few near-hull agents.
// Directional extremes of the stage-1 candidates, D directions per warp in one pass: the running maxima are D independent
// chains (one direction per pass left every warp waiting on a single dependent chain: 256 directions cost 14 us more
// than 64 that way, 6.6 us this way).  Any maximiser is a valid polygon vertex: the first lane that holds the maximum.
template <int D>
__device__ __forceinline__ void refine_search(const Buffers& b, int nc, bool staged, const float2* s_cd, float cx, float cy, int kstep,
                                              unsigned long long* s_ext) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t bv[D];
    int bc[D];
    float ddx[D], ddy[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
        bv[d] = 0u; bc[d] = 0;
        ddx[d] = c_dir2x[(wid + 32 * d) * kstep]; ddy[d] = c_dir2y[(wid + 32 * d) * kstep];
    }
    for (int c = lane; c < nc; c += 32) {
        float2 q;
        if (staged) q = s_cd[c]; else { const float4 cd = b.cand[c]; q = make_float2(cd.x, cd.y); }
        const float rx = q.x - cx, ry = q.y - cy;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const uint32_t ov = f2ord(rx * ddx[d] + ry * ddy[d]);
            if (ov > bv[d]) { bv[d] = ov; bc[d] = c; }
        }
    }
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const uint32_t best = __reduce_max_sync(0xffffffffu, bv[d]);
        const unsigned who = __ballot_sync(0xffffffffu, bv[d] == best);
        const int cc = __shfl_sync(0xffffffffu, bc[d], __ffs(who) - 1);
        if (lane == 0) s_ext[wid + 32 * d] = ((unsigned long long)best << 32) | (uint32_t)cc;
    }
}

constexpr int REFINE_SMEM = 2048;  // candidates staged in shared memory (typical: a few hundred); beyond that they are read from global
__device__ __forceinline__ void refine_body(const Params* __restrict__ P, const Buffers& b) {
    Ctl* ctl = b.ctl;
    const int nc = ctl->n_cand;
    if (nc > REFINE_MAX || nc < 8) {  // degenerate outline (or tiny crowd): keep the stage-1 list
        if (threadIdx.x == 0) { ctl->cand_sel = 0; ctl->n_cand2 = nc; }
        return;
    }
    __shared__ unsigned long long s_ext[HULL_K2];
    __shared__ float s_ex[HULL_K2], s_ey[HULL_K2], s_enx[HULL_K2], s_eny[HULL_K2];
    __shared__ float2 s_cd[REFINE_SMEM];
    __shared__ int s_ne, s_out;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const float cx = ctl->cx, cy = ctl->cy;
    const bool staged = nc <= REFINE_SMEM;
    if (tid == 0) { s_ne = 0; s_out = 0; }
    if (staged)
        for (int c = tid; c < nc; c += blockDim.x) { const float4 cd = b.cand[c]; s_cd[c] = make_float2(cd.x, cd.y); }
    __syncthreads();
    // How many directions?  The search costs K2 x nc / 1024 steps of this one block; every survivor it removes saves work in
    // the third stage of ALL blocks of k_force.  On one GPU the kernel runs beside k_pairs and is never the longer of the
    // two (plaza, 1 M agents: 64 / 128 / 256 directions leave 195 / 90 / 90 candidates, k_force 87 / 77 / 74 us, k_refine
    // 15 / 19 / 28 us beside 101 us of k_pairs; corridor with its 1 400 collinear outline agents and the 1 000-agent circle,
    // all of them hull vertices: nothing to gain, and more directions make k_refine the longer one — profiles/ab_r2y_*,
    // bench_r2z_*).
    // Both kernels take about 1e-4 us per unit — direction x candidate here, agent there — so the search stays the
    // shorter one while K2 x nc is well below the agent count.
    const long long budget = (long long)P->n / 2;
    const int K2 = P->hull_k2 > 0 ? P->hull_k2 : (256ll * nc <= budget ? 256 : 128ll * nc <= budget ? 128 : 64);
    const int kstep = HULL_K2 / K2;   // the K2 directions are every kstep-th of the table's
    // warp `wid` finds the extremes of directions wid, wid + 32, ... in ONE pass over the candidates (refine_search)
    if (K2 == 256) refine_search<8>(b, nc, staged, s_cd, cx, cy, kstep, s_ext);
    else if (K2 == 128) refine_search<4>(b, nc, staged, s_cd, cx, cy, kstep, s_ext);
    else refine_search<2>(b, nc, staged, s_cd, cx, cy, kstep, s_ext);
    __syncthreads();
    if (tid < K2) {  // edge from extreme k to extreme k+1 (CCW), inward unit normal
        const int a = (int)(s_ext[tid] & 0xffffffffull), c = (int)(s_ext[(tid + 1) % K2] & 0xffffffffull);
        if (a != c) {
            const float4 pa = b.cand[a], pc = b.cand[c];
            const double ux = (double)pc.x - pa.x, uy = (double)pc.y - pa.y, l = sqrt(ux * ux + uy * uy);
            if (l > 0.0) {
                const int e = atomicAdd(&s_ne, 1);
                s_ex[e] = pa.x; s_ey[e] = pa.y;
                s_enx[e] = (float)(-uy / l); s_eny[e] = (float)(ux / l);
            }
        }
    }
    __syncthreads();
    const int ne = s_ne;
    const float delta = ctl->delta;
    for (int c = tid; c < nc; c += blockDim.x) {
        float2 q;
        if (staged) q = s_cd[c]; else { const float4 cd = b.cand[c]; q = make_float2(cd.x, cd.y); }
        bool inside = ne >= 3;
        for (int e = 0; e < ne && inside; ++e) {
            const float d = (q.x - s_ex[e]) * s_enx[e] + (q.y - s_ey[e]) * s_eny[e];
            if (!(d >= delta)) inside = false;
        }
        if (!inside) b.cand2[atomicAdd(&s_out, 1)] = b.cand[c];
    }
    __syncthreads();
    if (tid == 0) { ctl->cand_sel = 1; ctl->n_cand2 = s_out; }
}

__device__ void dd_global_body(const Params* __restrict__ P, const Buffers& b);

// k_refine; in domain mode the same block then does the frame's all-rank exchange (comm-partner candidates, scoring
// inputs, replan gate, latch) and S0 (domain.cuh) — beside k_pairs, off the frame's critical path
__global__ void __launch_bounds__(1024) k_refine(const Params* __restrict__ P, Buffers b) {
    pdl_enter();
    Ctl* ctl = b.ctl;
    if (ctl->halt) return;
    stamp_entry(ctl, TS_REFINE);
    if (!(P->flags & 1u)) refine_body(P, b);
    if (b.dd) {
        __syncthreads();
        dd_global_body(P, b);
    }
    __syncthreads();
    stamp_entry(ctl, TS_REFINE_END);
}

// ---------------------------------------------------------------------------------------------------------
// S1 + S2 + S3a fused, one thread per agent in cell order.

// One pair (s, j), from scratch: LMP.cpp:37-55 / BoundVolume.cpp:102-122 in the relative frame (same operations as
// lc_math.cuh ttc_pair and, bit for bit, as the packed lanes of the filter below), the near list of the collision pass,
// and the TTC force on a hit (LMP.cpp:288-304).
template <bool FIRST, class F>
__device__ __forceinline__ void pair_eval(const Buffers& b, int s, int j, float px, float py, float vx, float vy, float RR,
                                          float near_c, int& n_near, int& n_overlap, F&& f) {
    const float4 q = __ldg(&set_in(b).pvf[j]);
    const float mx = __fsub_rn(q.x, px), my = __fsub_rn(q.y, py), dvx = __fsub_rn(q.z, vx), dvy = __fsub_rn(q.w, vy);
    const float bq = dot2(mx, my, dvx, dvy), c = __fsub_rn(dot2(mx, my, mx, my), RR);
    const float a = dot2(dvx, dvy, dvx, dvy);
    const float disc = __fsub_rn(__fmul_rn(bq, bq), __fmul_rn(a, c));
    if (FIRST && c < near_c) {
        if (n_near < MAX_NEAR) b.near_idx[(size_t)s * MAX_NEAR + n_near] = j;
        ++n_near;
        if (c < 0.f) ++n_overlap;  // BoundVolume.cpp:203-207 on frame-start positions
    }
    if ((c > 0.f && bq > 0.f) || disc < 0.f) return;
    float t = __fdiv_rn(__fsub_rn(-bq, __fsqrt_rn(disc)), a);
    if (t < 0.f) t = 0.f;
    if (!(t >= 0.f)) return;  // NaN -> FLT_MAX (LMP.cpp:31-34)
    if (t > 5.0f) return;     // (double)ttc > 5.0 (LMP.cpp:297-300)
    float fx, fy;
    ttc_force(px, py, vx, vy, q.x, q.y, q.z, q.w, t, fx, fy);
    f(j, fx, fy);
}

// The packed filter on one slot pair: cand = "may interact (ttc <= 5) or within reach of the collision pass".
// Straight-line on purpose: with 32 lanes x 2 candidates some lane always needs the whole filter, so "early outs" only
// added branches (the branchy version ran 77 instructions per iteration at 21 lanes); what survives is rare (~0.6
// interacting neighbours per agent in the plaza) and costs one predictable branch per iteration.
template <bool FIRST>
__device__ __forceinline__ void pair_filter2(const float4 P4, const float4 V4, const float2 px2, const float2 py2, const float2 vx2,
                                             const float2 vy2, const float2 RR2, const float near_c, bool& cand0, bool& cand1) {
    const float2 mx = sub2(make_float2(P4.x, P4.y), px2), my = sub2(make_float2(P4.z, P4.w), py2);
    const float2 dvx = sub2(make_float2(V4.x, V4.y), vx2), dvy = sub2(make_float2(V4.z, V4.w), vy2);
    // ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 even under -fmad=false, which would round once instead
    // of twice: every sum that consumes a product is therefore a SCALAR __fadd_rn / __fsub_rn (never contracted);
    // packed adds only ever consume loads or other sums.  tests/test_oracle_cpu.py checks the SASS for FFMA2.
    const float2 xx = mul2(mx, mx), yy2 = mul2(my, my), xd = mul2(mx, dvx), yd = mul2(my, dvy);
    const float2 bq = make_float2(__fadd_rn(xd.x, yd.x), __fadd_rn(xd.y, yd.y));
    const float2 c = sub2(make_float2(__fadd_rn(xx.x, yy2.x), __fadd_rn(xx.y, yy2.y)), RR2);
    const float2 dxx = mul2(dvx, dvx), dyy = mul2(dvy, dvy);
    const float2 a = make_float2(__fadd_rn(dxx.x, dyy.x), __fadd_rn(dxx.y, dyy.y));
    const float2 bb = mul2(bq, bq), ac = mul2(a, c);
    const float2 disc = make_float2(__fsub_rn(bb.x, ac.x), __fsub_rn(bb.y, ac.y));
    // Conservative "t > 5" filter, so that sqrt + division only run for near-certain neighbours.  With c >= 0 and
    // b <= 0: t = (-b - sqrt(disc)) / a > 5  <=>  u = -b - 5a > sqrt(disc).  Requiring u > 0 and
    // u^2 > disc + 1e-4 b^2 leaves u - sqrt(disc) > 5e-5 |b|, two orders of magnitude above the rounding error of
    // the reference's float evaluation of t, so no pair with float ttc <= 5 is ever filtered.
    const float2 a5 = mul2(a, make_float2(5.f, 5.f));
    const float2 u = make_float2(__fsub_rn(-bq.x, a5.x), __fsub_rn(-bq.y, a5.y));  // scalar: keeps the SASS free of FFMA2
    const float2 uu = mul2(u, u), mb = mul2(bb, make_float2(1e-4f, 1e-4f));
    if (FIRST) {
        // near_c > 0 (the collision reach exceeds the contact distance), so beyond it c > 0 holds and the tests on the
        // sign of c drop out: miss = b > 0 || disc < 0 || far.  Comparisons keep the polarity of the general form
        // below, so a NaN anywhere still makes the pair a candidate.
        const bool far0 = u.x > 0.f && uu.x > __fadd_rn(disc.x, mb.x);
        const bool far1 = u.y > 0.f && uu.y > __fadd_rn(disc.y, mb.y);
        cand0 = c.x < near_c || !(bq.x > 0.f || disc.x < 0.f || far0);
        cand1 = c.y < near_c || !(bq.y > 0.f || disc.y < 0.f || far1);
    } else {
        const bool far0 = c.x >= 0.f && u.x > 0.f && uu.x > __fadd_rn(disc.x, mb.x);
        const bool far1 = c.y >= 0.f && u.y > 0.f && uu.y > __fadd_rn(disc.y, mb.y);
        cand0 = !((c.x > 0.f && bq.x > 0.f) || disc.x < 0.f || far0);
        cand1 = !((c.y > 0.f && bq.y > 0.f) || disc.y < 0.f || far1);
    }
}

// Pair filter over the cell block of agent s: note(j) for every j in the block that MAY interact with s (ttc <= 5) or
// is within reach of the collision pass (FIRST).  Cells of one grid row are contiguous in the sorted arrays, so each
// row of the block is ONE contiguous slot range.  HALF: only partners that come after s in slot order (the rest of
// its own row and the rows above) — the filter is exactly symmetric in (s, j): m and dv only change sign, every
// product and sum is the same float, so one test serves both agents (k_pairs).
template <bool HALF, bool FIRST, class N>
__device__ __forceinline__ void scan_cells(const Buffers& b, const Ctl* ctl, int s, float px, float py, float vx, float vy,
                                           float RR, float near_c, N&& note, long long* tested = nullptr) {
    const int nx = ctl->nx, ny = ctl->ny;
    // Disc-shaped stencil around the agent's CURRENT position: a row of cells whose nearest edge is gap_y away can only hold
    // partners within sqrt(reach^2 - gap_y^2) in x, reach = cutoff + skin (the cells hold the agents where they stood at
    // the last re-sort; skin bounds how far anybody has moved since).  Inflated by 1e-3 so that float rounding here can
    // never exclude a cell that cell_of (the same monotonic expression, applied to a position inside the disc) put an
    // agent into.
    const float cut = ctl->cutoff * 1.001f + 1e-3f + ctl->skin, cut2 = cut * cut;
    const float inv_h = 1.f / ctl->hx, inv_hy = 1.f / ctl->h;
    // HALF: every unordered pair is tested by the agent that comes first in SLOT order, i.e. in the order of the cells of
    // the last re-sort: rows from the agent's sorted-into row upwards, in that row the slots after its own
    const int row_b = HALF ? b.slot_cell[s] / nx : 0;
    const float yl = (py - cut - ctl->y0) * inv_hy, yh = (py + cut - ctl->y0) * inv_hy;
    const int row_lo = HALF ? row_b : min(max((int)floorf(yl - (1e-3f + 1e-6f * fabsf(yl))), 0), ny - 1);
    const int row_hi = min(max((int)floorf(yh + (1e-3f + 1e-6f * fabsf(yh))), HALF ? row_b : 0), ny - 1);
    const float2 px2 = make_float2(px, px), py2 = make_float2(py, py), vx2 = make_float2(vx, vx), vy2 = make_float2(vy, vy);
    const float2 RR2 = make_float2(RR, RR);
    const float4* __restrict__ pq = set_in(b).pq;
    const float4* __restrict__ pvf = set_in(b).pvf;
    // scalar version of the filter for the odd slot at either end of a row range
    auto single = [&](int j) {
        const float4 q = __ldg(&pvf[j]);
        const float mx = __fsub_rn(q.x, px), my = __fsub_rn(q.y, py), dvx = __fsub_rn(q.z, vx), dvy = __fsub_rn(q.w, vy);
        const float bq = dot2(mx, my, dvx, dvy), c = __fsub_rn(dot2(mx, my, mx, my), RR);
        if (bq > 0.f && c >= near_c) return;
        const float a = dot2(dvx, dvy, dvx, dvy);
        const float bb = __fmul_rn(bq, bq), disc = __fsub_rn(bb, __fmul_rn(a, c));
        const float u = __fsub_rn(-bq, __fmul_rn(a, 5.f));
        const bool far = c >= 0.f && u > 0.f && __fmul_rn(u, u) > __fadd_rn(disc, __fmul_rn(bb, 1e-4f));
        const bool miss = (c > 0.f && bq > 0.f) || disc < 0.f || far;
        if (!miss || (FIRST && c < near_c)) note(j);
    };
    auto filter2 = [&](const float4 P4, const float4 V4, bool& cand0, bool& cand1) {
        pair_filter2<FIRST>(P4, V4, px2, py2, vx2, vy2, RR2, near_c, cand0, cand1);
    };
    for (int yy = row_lo; yy <= row_hi; ++yy) {
        // distance from the agent to the row; the outermost rows also hold whatever was clamped into them, so they reach to
        // infinity on their outer side
        const float ylo = ctl->y0 + (float)yy * ctl->h;
        const float below = yy == 0 ? 0.f : ylo - py, above = yy == ny - 1 ? 0.f : py - (ylo + ctl->h);
        const float gap = fmaxf(fmaxf(below, above), 0.f);
        const float rem = cut2 - gap * gap;
        if (!(rem > 0.f)) continue;
        // approximate, but widened beyond every rounding: rsqrt is good to 2 ulp, the product with 1/h to < 1e-6 of its
        // value, and the range only has to CONTAIN the cells cell_of() can have put a partner inside the disc into
        const float reach = rem * rsqrtf(rem) * 1.0002f + 1e-4f;
        const float xl = (px - reach - ctl->x0) * inv_h, xh = (px + reach - ctl->x0) * inv_h;
        const int x_lo = min(max((int)floorf(xl - (1e-3f + 1e-6f * fabsf(xl))), 0), nx - 1);
        const int x_hi = min(max((int)floorf(xh + (1e-3f + 1e-6f * fabsf(xh))), 0), nx - 1);
        // in its own row the partners after s start at slot s + 1
        int j0 = (HALF && yy == row_b) ? s + 1 : b.cell_start[yy * nx + x_lo], j1 = b.cell_start[yy * nx + x_hi + 1];
        if (j0 >= j1) continue;
        if (tested) { *tested += j1 - j0; continue; }  // counting mode (lc_count_pair_tests): the ranges, not the filter
        if (j0 & 1) single(j0++);
        if ((j1 & 1) && j1 > j0) single(--j1);
        // two candidates per packed operation: slots (j, j + 1), j even, are stored as (x0, x1, z0, z1), (vx0, vx1, vz0, vz1);
        // two such slot pairs per iteration (loop overhead was a fifth of the instructions at one)
        int p = j0 >> 1;
        const int pe = j1 >> 1;
        for (; p + 1 < pe; p += 2) {
            const float4 A0 = __ldg(&pq[2 * p]), B0 = __ldg(&pq[2 * p + 1]), A1 = __ldg(&pq[2 * p + 2]), B1 = __ldg(&pq[2 * p + 3]);
            bool c0, c1, c2, c3;
            filter2(A0, B0, c0, c1);
            filter2(A1, B1, c2, c3);
            if (c0 | c1 | c2 | c3) {
                if (c0) note(2 * p);
                if (c1) note(2 * p + 1);
                if (c2) note(2 * p + 2);
                if (c3) note(2 * p + 3);
            }
        }
        if (p < pe) {
            const float4 A0 = __ldg(&pq[2 * p]), B0 = __ldg(&pq[2 * p + 1]);
            bool c0, c1;
            filter2(A0, B0, c0, c1);
            if (c0) note(2 * p);
            if (c1) note(2 * p + 1);
        }
    }
}

// EXPERIMENT (LC_PAIRS_FLAT=1, off by default): k_pairs' walk over the HALF block of agent s as ONE loop over all its rows.
// Idea: in scan_cells every row is a loop of its own and the lanes of a warp have different trip counts per row (22 of 32
// lanes active, profiles/prof_k_pairs_r2j_raw.csv); with the row ranges set up first, in lock step, one loop could run
// through them back to back.  Measured on a B200 (profiles/ab_r2k_*): SLOWER, 130 vs 105 us at 1 M agents, 19 lanes
// active.  The imbalance is not between rows but between agents: the half block of an agent at the bottom of its row is
// up to twice that of one at the top, so the totals differ as much as the per-row counts do (a model of both loops over
// the plaza gives 0.82 vs 0.84 lane efficiency), and widening the ranges to whole groups of four costs 10 % more tests.
// Kept for the record and for finer stencils, where the rows are shorter.
constexpr int FLAT_ROWS = 3;
template <class N>
__device__ __forceinline__ bool scan_pairs_flat(const Buffers& b, const Ctl* ctl, int s, float px, float py, float vx, float vy,
                                                float RR, float near_c, N&& note, long long* tested = nullptr) {
    const int nx = ctl->nx, ny = ctl->ny;
    const float cut = ctl->cutoff * 1.001f + 1e-3f + ctl->skin, cut2 = cut * cut;   // as in scan_cells
    const float inv_h = 1.f / ctl->hx, inv_hy = 1.f / ctl->h;
    const int row_b = b.slot_cell[s] / nx;
    const float yh = (py + cut - ctl->y0) * inv_hy;
    const int row_hi = min(max((int)floorf(yh + (1e-3f + 1e-6f * fabsf(yh))), row_b), ny - 1);
    if (row_hi - row_b >= FLAT_ROWS) return false;
    int lo[FLAT_ROWS], hi[FLAT_ROWS];
#pragma unroll
    for (int r = 0; r < FLAT_ROWS; ++r) {
        const int yy = min(row_b + r, ny - 1);
        const float ylo = ctl->y0 + (float)yy * ctl->h;
        const float below = yy == 0 ? 0.f : ylo - py, above = yy == ny - 1 ? 0.f : py - (ylo + ctl->h);
        const float gap = fmaxf(fmaxf(below, above), 0.f);
        const float rem = cut2 - gap * gap;
        const bool live = row_b + r <= row_hi && rem > 0.f;
        const float remc = live ? rem : 1.f;
        const float reach = remc * rsqrtf(remc) * 1.0002f + 1e-4f;
        const float xl = (px - reach - ctl->x0) * inv_h, xh = (px + reach - ctl->x0) * inv_h;
        const int x_lo = min(max((int)floorf(xl - (1e-3f + 1e-6f * fabsf(xl))), 0), nx - 1);
        const int x_hi = min(max((int)floorf(xh + (1e-3f + 1e-6f * fabsf(xh))), 0), nx - 1);
        const int j0 = r == 0 ? s + 1 : b.cell_start[yy * nx + x_lo], j1 = b.cell_start[yy * nx + x_hi + 1];
        lo[r] = live && j0 < j1 ? j0 : 0;
        hi[r] = live && j0 < j1 ? j1 : 0;
    }
    if (tested) {  // counting mode: the filter evaluations the loop below performs
#pragma unroll
        for (int r = 0; r < FLAT_ROWS; ++r) *tested += 4 * ((((hi[r] + 1) >> 1) - (lo[r] >> 1) + 1) >> 1);
        return true;
    }
    const float2 px2 = make_float2(px, px), py2 = make_float2(py, py), vx2 = make_float2(vx, vx), vy2 = make_float2(vy, vy);
    const float2 RR2 = make_float2(RR, RR);
    const float4* __restrict__ pq = set_in(b).pq;
    int r = 0, jl = lo[0], jh = hi[0], p = jl >> 1, pe = (jh + 1) >> 1;
    for (;;) {
        while (p >= pe) {  // next row (register selects: the arrays are never indexed dynamically)
            if (++r >= FLAT_ROWS) return true;
            jl = r == 1 ? lo[1] : lo[2];
            jh = r == 1 ? hi[1] : hi[2];
            p = jl >> 1; pe = (jh + 1) >> 1;
        }
        // slot pairs p and p + 1 (the second may lie past the range, even past the last agent: the arrays are padded and
        // the range check below drops whatever it yields)
        const float4 A0 = __ldg(&pq[2 * p]), B0 = __ldg(&pq[2 * p + 1]), A1 = __ldg(&pq[2 * p + 2]), B1 = __ldg(&pq[2 * p + 3]);
        bool c0, c1, c2, c3;
        pair_filter2<true>(A0, B0, px2, py2, vx2, vy2, RR2, near_c, c0, c1);
        pair_filter2<true>(A1, B1, px2, py2, vx2, vy2, RR2, near_c, c2, c3);
        if (c0 | c1 | c2 | c3) {
            const int j = 2 * p;
            if (c0 && j >= jl && j < jh) note(j);
            if (c1 && j + 1 >= jl && j + 1 < jh) note(j + 1);
            if (c2 && j + 2 >= jl && j + 2 < jh) note(j + 2);
            if (c3 && j + 3 >= jl && j + 3 < jh) note(j + 3);
        }
        p += 2;
    }
}

// Every interacting neighbour of agent s, evaluated on the spot over its whole cell block.  The rare paths only: an
// agent whose pair list overflowed, and the per-neighbour re-scan when the hit buffer overflowed.
template <bool FIRST, class F>
__device__ __forceinline__ void scan_neighbours(const Buffers& b, const Ctl* ctl, int s, float px, float py, float vx, float vy,
                                             float RR, float near_c, int& n_near, int& n_overlap, F&& f) {
    scan_cells<false, FIRST>(b, ctl, s, px, py, vx, vy, RR, near_c, [&](int j) {
        if (j != s) pair_eval<FIRST>(b, s, j, px, py, vx, vy, RR, near_c, n_near, n_overlap, f);
    });
}

// S2a: pair filter, each unordered pair tested ONCE (by the agent that comes first in cell order) and recorded in the
// pair lists of both; k_force evaluates the lists.  Halves the pair tests of a per-agent scan of the whole disc.
// EXPERIMENT (LC_PAIRS_STAGED=1, off by default): north_star's "stage each grid cell's agents in shared memory".  The 128
// agents of a block are neighbours in slot order, so the slot ranges their half blocks cover overlap almost entirely: per
// row offset (own row, the two rows above) the block stages the UNION of its threads' ranges — pair-interleaved (pos, vel)
// records, ~450 slots = 7 KB — with coalesced loads, and every thread then runs the per-row walk of scan_cells over shared
// memory instead of through L1.  Blocks that straddle two grid rows, see more than three rows or would need more than
// STAGE_PAIRS slot pairs take the ordinary path (block-uniform decision).  Same tests, same order, same notes.
// Measured on a B200: see DESIGN.md section 5 (profiles/ab_r2v_*, prof_k_pairs_r2v_staged_raw.csv).
#ifndef LC_PAIRS_STAGED
#define LC_PAIRS_STAGED 0
#endif
constexpr int STAGE_PAIRS = 352;   // slot pairs a block stages (704 agents, 11 KB)
template <class N>
__device__ __forceinline__ bool pairs_staged(const Buffers& b, const Ctl* ctl, int s, bool valid, float4 me, float RR, float near_c, N&& note) {
    __shared__ float4 s_pq[2 * STAGE_PAIRS];
    __shared__ int s_lo[3], s_hi[3], s_rowmin, s_rowmax, s_bad;
    const float px = me.x, py = me.y, vx = me.z, vy = me.w;
    const int nx = ctl->nx, ny = ctl->ny;
    if (threadIdx.x < 3) { s_lo[threadIdx.x] = 0x7fffffff; s_hi[threadIdx.x] = 0; }
    if (threadIdx.x == 0) { s_rowmin = 0x7fffffff; s_rowmax = -1; s_bad = 0; }
    __syncthreads();
    int lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    int row_b = 0x7fffffff;
    bool bad = false;
    if (valid) {
        const float cut = ctl->cutoff * 1.001f + 1e-3f + ctl->skin, cut2 = cut * cut;   // as in scan_cells
        const float inv_h = 1.f / ctl->hx, inv_hy = 1.f / ctl->h;
        row_b = b.slot_cell[s] / nx;
        const float yh = (py + cut - ctl->y0) * inv_hy;
        const int row_hi = min(max((int)floorf(yh + (1e-3f + 1e-6f * fabsf(yh))), row_b), ny - 1);
        bad = row_hi - row_b >= 3;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int yy = min(row_b + r, ny - 1);
            const float ylo = ctl->y0 + (float)yy * ctl->h;
            const float below = yy == 0 ? 0.f : ylo - py, above = yy == ny - 1 ? 0.f : py - (ylo + ctl->h);
            const float gap = fmaxf(fmaxf(below, above), 0.f);
            const float rem = cut2 - gap * gap;
            const bool live = row_b + r <= row_hi && rem > 0.f;
            const float remc = live ? rem : 1.f;
            const float reach = remc * rsqrtf(remc) * 1.0002f + 1e-4f;
            const float xl = (px - reach - ctl->x0) * inv_h, xh = (px + reach - ctl->x0) * inv_h;
            const int x_lo = min(max((int)floorf(xl - (1e-3f + 1e-6f * fabsf(xl))), 0), nx - 1);
            const int x_hi = min(max((int)floorf(xh + (1e-3f + 1e-6f * fabsf(xh))), 0), nx - 1);
            const int j0 = r == 0 ? s + 1 : b.cell_start[yy * nx + x_lo], j1 = b.cell_start[yy * nx + x_hi + 1];
            if (live && j0 < j1) { lo[r] = j0; hi[r] = j1; }
        }
    }
    {   // union of the block's ranges per row offset; do all its agents sit in one grid row?
        const unsigned full = 0xffffffffu;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int l = __reduce_min_sync(full, hi[r] > lo[r] ? lo[r] : 0x7fffffff), h = __reduce_max_sync(full, hi[r]);
            if ((threadIdx.x & 31) == 0) { if (l != 0x7fffffff) atomicMin(&s_lo[r], l); if (h) atomicMax(&s_hi[r], h); }
        }
        const int rmin = __reduce_min_sync(full, row_b), rmax = __reduce_max_sync(full, valid ? row_b : -1);
        const int anybad = __any_sync(full, bad);
        if ((threadIdx.x & 31) == 0) { atomicMin(&s_rowmin, rmin); atomicMax(&s_rowmax, rmax); if (anybad) s_bad = 1; }
    }
    __syncthreads();
    int base[3], plo[3];
    int total = 0;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const int L = s_lo[r], H = s_hi[r];
        plo[r] = L == 0x7fffffff ? 0 : L >> 1;
        const int cnt = L == 0x7fffffff ? 0 : ((H + 1) >> 1) - plo[r];
        base[r] = total;
        total += cnt;
    }
    if (s_bad || s_rowmin != s_rowmax || total > STAGE_PAIRS) return false;   // block-uniform
    const float4* __restrict__ pq = set_in(b).pq;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const int cnt = (r < 2 ? base[r + 1] : total) - base[r];
        for (int k = threadIdx.x; k < 2 * cnt; k += blockDim.x) s_pq[2 * base[r] + k] = __ldg(&pq[2 * plo[r] + k]);
    }
    __syncthreads();
    if (!valid) return true;
    const float2 px2 = make_float2(px, px), py2 = make_float2(py, py), vx2 = make_float2(vx, vx), vy2 = make_float2(vy, vy);
    const float2 RR2 = make_float2(RR, RR);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        int j0 = lo[r], j1 = hi[r];
        if (j0 >= j1) continue;
        const float4* sp = s_pq + 2 * (base[r] - plo[r]);   // sp[2 p], sp[2 p + 1] = slot pair p
        auto single = [&](int j) {   // the scalar filter of scan_cells on one half of a slot pair
            const float4 P4 = sp[2 * (j >> 1)], V4 = sp[2 * (j >> 1) + 1];
            const bool odd = j & 1;
            const float qx = odd ? P4.y : P4.x, qy = odd ? P4.w : P4.z, qvx = odd ? V4.y : V4.x, qvy = odd ? V4.w : V4.z;
            const float mx = __fsub_rn(qx, px), my = __fsub_rn(qy, py), dvx = __fsub_rn(qvx, vx), dvy = __fsub_rn(qvy, vy);
            const float bq = dot2(mx, my, dvx, dvy), c = __fsub_rn(dot2(mx, my, mx, my), RR);
            if (bq > 0.f && c >= near_c) return;
            const float a = dot2(dvx, dvy, dvx, dvy);
            const float bb = __fmul_rn(bq, bq), disc = __fsub_rn(bb, __fmul_rn(a, c));
            const float u = __fsub_rn(-bq, __fmul_rn(a, 5.f));
            const bool far = c >= 0.f && u > 0.f && __fmul_rn(u, u) > __fadd_rn(disc, __fmul_rn(bb, 1e-4f));
            const bool miss = (c > 0.f && bq > 0.f) || disc < 0.f || far;
            if (!miss || c < near_c) note(j);
        };
        if (j0 & 1) single(j0++);
        if ((j1 & 1) && j1 > j0) single(--j1);
        int p = j0 >> 1;
        const int pe = j1 >> 1;
        for (; p + 1 < pe; p += 2) {
            const float4 A0 = sp[2 * p], B0 = sp[2 * p + 1], A1 = sp[2 * p + 2], B1 = sp[2 * p + 3];
            bool c0, c1, c2, c3;
            pair_filter2<true>(A0, B0, px2, py2, vx2, vy2, RR2, near_c, c0, c1);
            pair_filter2<true>(A1, B1, px2, py2, vx2, vy2, RR2, near_c, c2, c3);
            if (c0 | c1 | c2 | c3) {
                if (c0) note(2 * p);
                if (c1) note(2 * p + 1);
                if (c2) note(2 * p + 2);
                if (c3) note(2 * p + 3);
            }
        }
        if (p < pe) {
            const float4 A0 = sp[2 * p], B0 = sp[2 * p + 1];
            bool c0, c1;
            pair_filter2<true>(A0, B0, px2, py2, vx2, vy2, RR2, near_c, c0, c1);
            if (c0) note(2 * p);
            if (c1) note(2 * p + 1);
        }
    }
    return true;
}

#ifndef LC_PAIR_BLOCKS
#define LC_PAIR_BLOCKS 12
#endif
#ifndef LC_PAIRS_FLAT
#define LC_PAIRS_FLAT 0   // 1: scan_pairs_flat — measured and lost (130 vs 105 us at 1 M agents: profiles/ab_r2k_*.json, prof_k_pairs_r2k_flat_raw.csv)
#endif
struct PairNote {   // record the pair in the lists of both agents; entries past PAIR_CAP are only counted: k_force then scans the whole disc
    const Buffers& b; int s; size_t stride;
    __device__ __forceinline__ void operator()(int j) const {
        int k = atomicAdd(&b.pair_cnt[s], 1);
        if (k < PAIR_CAP) b.pair_list[k * stride + s] = j;
        k = atomicAdd(&b.pair_cnt[j], 1);
        if (k < PAIR_CAP) b.pair_list[k * stride + j] = s;
    }
};
// The same, deferred: survivors are parked in a small per-thread queue in shared memory and recorded when the queue is
// full and at the end of the walk.  A survivor costs two atomics whose results are needed for the list stores — an L2
// round trip in the middle of the pair loop — and although only ~1 % of the tests survive, with 32 lanes x 4 candidates
// per iteration SOME lane has one in most iterations, so the whole warp sat through that round trip again and again
// (long-scoreboard was the kernel's top stall, and staging the candidates in shared memory did not move it).  Flushing
// puts a thread's atomics in flight together, one count for its own list, and the warp waits once.
#ifndef LC_NOTE_QUEUE
#define LC_NOTE_QUEUE 4   // 0: record every survivor on the spot (PairNote)
#endif
struct QueuedNote {
    const Buffers& b; int s; size_t stride;
    int* q;    // this thread's queue: entry k at q[k * 128]
    int nq;
    __device__ __forceinline__ void flush() {
        constexpr int Q = LC_NOTE_QUEUE > 0 ? LC_NOTE_QUEUE : 1;
        if (nq == 0) return;
        const int base = atomicAdd(&b.pair_cnt[s], nq);
        int js[Q], kj[Q];
#pragma unroll
        for (int k = 0; k < Q; ++k)
            if (k < nq) { js[k] = q[k * 128]; kj[k] = atomicAdd(&b.pair_cnt[js[k]], 1); }
#pragma unroll
        for (int k = 0; k < Q; ++k)
            if (k < nq) {
                if (base + k < PAIR_CAP) b.pair_list[(size_t)(base + k) * stride + s] = js[k];
                if (kj[k] < PAIR_CAP) b.pair_list[(size_t)kj[k] * stride + js[k]] = s;
            }
        nq = 0;
    }
    __device__ __forceinline__ void operator()(int j) {
        constexpr int Q = LC_NOTE_QUEUE > 0 ? LC_NOTE_QUEUE : 1;
        if (nq == Q) flush();
        q[nq * 128] = j;
        ++nq;
    }
};
__global__ void __launch_bounds__(128, LC_PAIR_BLOCKS) k_pairs(const Params* __restrict__ P, Buffers b) {
    pdl_enter();
    const Ctl* ctl = b.ctl;
    if (ctl->halt) return;
    stamp_entry(b.ctl, TS_PAIRS);
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
#if LC_PAIRS_STAGED
    {
        if ((int)(blockIdx.x * blockDim.x) >= P->n) return;
        const bool valid = s < P->n;
        const float4 me0 = valid ? set_in(b).pvf[s] : make_float4(0.f, 0.f, 0.f, 0.f);
        if (pairs_staged(b, ctl, s, valid, me0, P->RR, P->near_c, PairNote{b, s, (size_t)P->cap})) return;
    }
#endif
    if (s >= P->n) return;
    const float4 me = set_in(b).pvf[s];
#if LC_NOTE_QUEUE > 0
    __shared__ int s_queue[LC_NOTE_QUEUE][128];
    QueuedNote note{b, s, (size_t)P->cap, &s_queue[0][threadIdx.x], 0};
    scan_cells<true, true>(b, ctl, s, me.x, me.y, me.z, me.w, P->RR, P->near_c, note);
    note.flush();
#else
    const PairNote note{b, s, (size_t)P->cap};
#if LC_PAIRS_FLAT
    if (scan_pairs_flat(b, ctl, s, me.x, me.y, me.z, me.w, P->RR, P->near_c, note)) return;
#endif
    // more rows than scan_pairs_flat handles (a finer stencil was requested): the per-row walk
    scan_cells<true, true>(b, ctl, s, me.x, me.y, me.z, me.w, P->RR, P->near_c, note);
#endif
}

// Diagnostic twin of k_pairs: how many pair-filter evaluations the last frame's k_pairs performed (the same row ranges,
// nothing evaluated or written).  Feeds the FP32 leg of bench.py's roofline.
__global__ void __launch_bounds__(128) k_pair_count(const Params* __restrict__ P, Buffers b, unsigned long long* out) {
    const Ctl* ctl = b.ctl;
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    long long tested = 0;
    if (s < P->n) {
        const float4 me = set_in(b).pvf[s];
#if LC_PAIRS_FLAT
        if (!scan_pairs_flat(b, ctl, s, me.x, me.y, me.z, me.w, P->RR, P->near_c, [](int) {}, &tested))
#endif
        scan_cells<true, true>(b, ctl, s, me.x, me.y, me.z, me.w, P->RR, P->near_c, [](int) {}, &tested);
    }
    for (int o = 16; o > 0; o >>= 1) tested += __shfl_xor_sync(0xffffffffu, tested, o);
    if ((threadIdx.x & 31) == 0 && tested) atomicAdd(out, (unsigned long long)tested);
}
void launch_pair_count(const Params* dP, int n, const Buffers& b, unsigned long long* out, cudaStream_t st, long long* launches) {
    k_pair_count<<<(n + 127) / 128, 128, 0, st>>>(dP, b, out);
    if (launches) *launches += 1;
}

#ifndef LC_FORCE_BLOCKS
#define LC_FORCE_BLOCKS 10
#endif

// minstd_rand0 skip-ahead: 16807^e mod (2^31 - 1)
__device__ __forceinline__ uint32_t lcg_pow_dev(unsigned long long e) {
    uint32_t r = 1u, a = 16807u;
    while (e) {
        if (e & 1ull) r = lcg_mul(r, a);
        a = lcg_mul(a, a);
        e >>= 1;
    }
    return r;
}

// walls of the scene for one block: shared memory when they fit, else straight from global (L1-resident: every thread
// walks the same list)
__device__ __forceinline__ const lcw::Wall* stage_walls(const Params* P, const Buffers& b, lcw::Wall* s_walls) {
    const int nw = P->n_walls;
    if (nw > WALL_SMEM) return reinterpret_cast<const lcw::Wall*>(b.w.walls);
    for (int k = threadIdx.x; k < nw; k += blockDim.x) {
        const WallRect w = b.w.walls[k];
        s_walls[k] = lcw::Wall{w.ox, w.oy, w.w, w.h};
    }
    __syncthreads();
    return s_walls;
}

// ai.cpp:152-163 for the COMING frame, evaluated on the final (synced) position of this one: sight line to the next
// waypoint, lost_frames, need_replan; requests are reduced to the two lowest ids (= the first two in table order, the
// only ones the reference serves per frame, ai.cpp:151-152,164).
__device__ __forceinline__ void walled_gate(const Params* P, const Buffers& b, const lcw::Wall* walls, int id, float px, float py,
                                            float gx, float gy) {
    const WalledArrays& W = b.w;
    const int np = W.n_plan[id], nd = W.num_done[id];
    const bool has = np > 0, done = has && nd == np;
    const float gd = len2(__fsub_rn(gx, px), __fsub_rn(gy, py));
    bool visible = false;
    if (has && nd < np) {
        const float2 t = W.plan[(size_t)id * WP_CAP + nd];
        visible = true;
        for (int k = 0; k < P->n_walls && visible; ++k) visible = lcw::cspace_line_of_sight(walls[k], P->r, px, py, t.x, t.y);
    }
    const int lost = visible ? 0 : W.lost[id] + 1;
    W.lost_next[id] = lost;
    const bool need = !has || (done && gd > 1.f) || (!done && !visible && lost >= 360);
    if (need) {
        W.req_pos[id] = make_float2(px, py);
        const int old = atomicMin(&b.ctl->wr1, id);
        const int loser = old < id ? id : old;
        if (loser != 0x7fffffff) atomicMin(&b.ctl->wr2, loser);
    }
}

template <bool WALLS>
__global__ void __launch_bounds__(128, LC_FORCE_BLOCKS) k_force(const Params* __restrict__ P, Buffers b) {
    pdl_enter();
    Ctl* ctl = b.ctl;
    if (ctl->halt) return;
    const AgentArrays& in = set_in(b);
    const AgentArrays& out = set_out(b);
    stamp_entry(ctl, TS_FORCE);
    __shared__ lcw::Wall s_walls[WALLS ? WALL_SMEM : 1];
    const lcw::Wall* walls = nullptr;
    if constexpr (WALLS) walls = stage_walls(P, b, s_walls);
    const int n = P->n;
    if ((int)(blockIdx.x * blockDim.x) >= n) return;  // domain mode launches for the capacity, not the count
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    int not_done_after = 0;
    // The agent's own state is only loaded after the candidate pruning and the pair list (register pressure); ask L2 for
    // it now so that those loads do not each pay a DRAM round trip on the block's critical path.
    auto pf = [](const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); };
    int np_early = 0, j_first = 0;  // pair count and first list entry: loads issued now, consumed after the pruning
    if (s < n) {
        np_early = b.pair_cnt[s];
        j_first = b.pair_list[s];   // (stale or uninitialised when the count is 0: clamped before use)
        pf(&in.meta[s]); pf(&in.goal[s]); pf(&in.lgoal[s]); pf(&in.facing[s]); pf(&in.c4[s]); pf(&in.vfh[s]);
    }  // (k_force 81 -> 77 us at 1 M agents)
    // Third pruning stage of the comm-partner candidates, per block: the 128 agents of a block are neighbours in cell
    // order, so they share a small bounding box B.  Candidate k can be the arg-max of comm.cpp:201-220 for some query
    // in B only if max_{q in B} |q - k|^2 is not below L = max_k' min_{q in B} |q - k'|^2 (1e-4 relative slack >> the
    // float rounding of length2).  Survivors are staged in shared memory; typically a handful remain.
    __shared__ float4 s_cand[CAND_SMEM];
    __shared__ uint32_t s_bb[5];
    __shared__ int s_nc;
    const float4* __restrict__ cand = nullptr;   // the global list, when the block's survivors do not fit into shared memory
    bool staged = false;
    int ncand = 0;
    if (!(P->flags & 1u)) {
        const int nc = ctl->n_cand2;
        const float4* __restrict__ gc = ctl->cand_sel ? b.cand2 : b.cand;
        if (threadIdx.x == 0) { s_bb[0] = s_bb[1] = 0xffffffffu; s_bb[2] = s_bb[3] = 0u; s_bb[4] = 0u; s_nc = 0; }
        __syncthreads();
        const bool valid = s < n;
        const float4 me0 = valid ? in.pvf[s] : make_float4(0.f, 0.f, 0.f, 0.f);
        const uint32_t mnx = __reduce_min_sync(0xffffffffu, valid ? f2ord(me0.x) : 0xffffffffu);
        const uint32_t mxx = __reduce_max_sync(0xffffffffu, valid ? f2ord(me0.x) : 0u);
        const uint32_t mny = __reduce_min_sync(0xffffffffu, valid ? f2ord(me0.y) : 0xffffffffu);
        const uint32_t mxy = __reduce_max_sync(0xffffffffu, valid ? f2ord(me0.y) : 0u);
        if ((threadIdx.x & 31) == 0) {
            atomicMin(&s_bb[0], mnx); atomicMin(&s_bb[1], mny); atomicMax(&s_bb[2], mxx); atomicMax(&s_bb[3], mxy);
        }
        __syncthreads();
        if (valid && np_early > 0) {  // the first list entry has arrived by now: get its partner on the way
            const int jc = min(max(j_first, 0), n - 1);
            pf(&in.pvf[jc]); pf(&in.meta[jc]);
        }
        const float bx0 = ord2f(s_bb[0]), by0 = ord2f(s_bb[1]), bx1 = ord2f(s_bb[2]), by1 = ord2f(s_bb[3]);
        float lmax = 0.f;
        for (int k = threadIdx.x; k < nc; k += blockDim.x) {
            const float4 cd = __ldg(&gc[k]);
            const float dx = fmaxf(fmaxf(bx0 - cd.x, cd.x - bx1), 0.f), dy = fmaxf(fmaxf(by0 - cd.y, cd.y - by1), 0.f);
            const float d2 = dx * dx + dy * dy;
            lmax = d2 > lmax || d2 != d2 ? d2 : lmax;  // NaN propagates: then nothing is pruned
        }
        // d2 >= 0 (or NaN, whose bits sort above every finite value)
        const uint32_t wl = __reduce_max_sync(0xffffffffu, __float_as_uint(lmax));
        if ((threadIdx.x & 31) == 0 && wl) atomicMax(&s_bb[4], wl);
        __syncthreads();
        const float L = __uint_as_float(s_bb[4]);
        for (int k = threadIdx.x; k < nc; k += blockDim.x) {
            const float4 cd = __ldg(&gc[k]);
            const float dx = fmaxf(fabsf(cd.x - bx0), fabsf(cd.x - bx1)), dy = fmaxf(fabsf(cd.y - by0), fabsf(cd.y - by1));
            const float d2 = (dx * dx + dy * dy) * 1.0001f;
            if (!(d2 < L)) {
                const int q = atomicAdd(&s_nc, 1);
                if (q < CAND_SMEM) s_cand[q] = cd;
            }
        }
        __syncthreads();
        if (s_nc <= CAND_SMEM) { staged = true; ncand = s_nc; } else { cand = gc; ncand = nc; }
    }
    if (s < n) {
        const float4 me = in.pvf[s];
        const float px = me.x, py = me.y, vx = me.z, vy = me.w;
        const int id = in.meta[s].x;

        // The pair loop runs first, while only (pos, vel, id) are live, so that it gets by with few registers; the
        // rest of the agent's state is loaded after it.  Order of the float operations is unaffected.
        // LMP.cpp:288-304: TTC forces, summed in table (id) order
        int hid[MAX_HITS];
        float hfx[MAX_HITS], hfy[MAX_HITS];
        int cnt = 0, n_near = 0, n_overlap = 0;
        auto hit = [&](int j, float fx, float fy) {
            if (cnt < MAX_HITS) { hid[cnt] = __ldg(&in.meta[j].x); hfx[cnt] = fx; hfy[cnt] = fy; }
            ++cnt;
        };
        const int np = np_early;
        if (np <= PAIR_CAP) {  // survivors of k_pairs' filter (in no particular order: hits and near list are sorted by id)
            const size_t stride = (size_t)P->cap;
            for (int k = 0; k < np; ++k)
                pair_eval<true>(b, s, k == 0 ? j_first : b.pair_list[k * stride + s], px, py, vx, vy, P->RR, P->near_c, n_near,
                                n_overlap, hit);
        } else {
            scan_neighbours<true>(b, ctl, s, px, py, vx, vy, P->RR, P->near_c, n_near, n_overlap, hit);
        }
        float ttcFx = 0.f, ttcFy = 0.f;
        int* lst = ((P->flags & 4u) && !(in.meta[s].y & META_GHOST)) ? b.nbr_list + (size_t)id * b.list_cap : nullptr;
        if (cnt <= MAX_HITS) {
            int last = -1;
            for (int k = 0; k < cnt; ++k) {
                int bi = 0, bid = 0x7fffffff;
                for (int q = 0; q < cnt; ++q) {
                    int v = hid[q];
                    if (v > last && v < bid) { bid = v; bi = q; }
                }
                ttcFx = __fadd_rn(ttcFx, hfx[bi]);
                ttcFy = __fadd_rn(ttcFy, hfy[bi]);
                last = bid;
                if (lst && k < b.list_cap) lst[k] = bid;
            }
        } else {  // rare: more interacting neighbours than the register buffer holds — re-scan per neighbour
            atomicAdd(&ctl->hit_overflow, 1);
            int last = -1, dummy0 = 0, dummy1 = 0;
            for (int k = 0; k < cnt; ++k) {
                int bid = 0x7fffffff;
                float bfx = 0.f, bfy = 0.f;
                scan_neighbours<false>(b, ctl, s, px, py, vx, vy, P->RR, P->near_c, dummy0, dummy1,
                                       [&](int j, float fx, float fy) {
                                           int v = __ldg(&in.meta[j].x);
                                           if (v > last && v < bid) { bid = v; bfx = fx; bfy = fy; }
                                       });
                ttcFx = __fadd_rn(ttcFx, bfx);
                ttcFy = __fadd_rn(ttcFy, bfy);
                last = bid;
                if (lst && k < b.list_cap) lst[k] = bid;
            }
        }
        if constexpr (WALLS) {  // LMP.cpp:306-319: the statics loop, after the agents, in the BoundVolume* table's order
            for (int k = 0; k < P->n_walls; ++k) {
                float fx, fy;
                if (lcw::wall_force(px, py, P->r, vx, vy, walls[b.w.wall_order[k]], fx, fy)) {
                    ttcFx = __fadd_rn(ttcFx, fx);
                    ttcFy = __fadd_rn(ttcFy, fy);
                }
            }
        }
        int4 mt = in.meta[s];
        const bool owned = !(mt.y & META_GHOST);  // ghosts are advanced too (the collision pass needs their integrated
                                                  // position) but never counted: their owner does that
        b.diag_cnt[s] = cnt;  // slot-indexed (coalesced); lc_get_neighbor_counts translates to ids (k_get_diag)
        if (n_near > MAX_NEAR) atomicAdd(&ctl->near_overflow, 1);
        b.near_cnt[s] = n_near;
        if (n_overlap && owned) atomicAdd(&ctl->overlap2, (unsigned long long)n_overlap);

        const float2 goal = in.goal[s];
        float2 lgoal = in.lgoal[s];
        float2 facing = in.facing[s];
        float4 c4 = in.c4[s];
        const bool done_before = mt.y & 1;
        float forcex = 0.f, forcey = 0.f;

        // ---- S1: comm::run (comm.cpp:267-305) ----
        if (!(P->flags & 1u)) {
            // comm.cpp:201-220 select_nearest == arg-max of squared distance (quirk Q1), strict '>', first in table order
            float best = LC_FLT_MIN;
            int best_id = -1, best_slot = s;
            // Most of the ~65 candidates a block keeps are far from being this agent's arg-max.  A fused-multiply-add
            // estimate of the squared distance (within 2 ulp of the reference's mul/mul/add value) against a running
            // threshold 1e-5 below the largest estimate seen so far rejects them in 6 instructions; only candidates
            // within that margin go through the reference's own arithmetic, predicate and tie rule.  NaN-safe polarity:
            // a non-finite estimate is never rejected here.
            float thr = 0.f;
            auto consider = [&](float cx, float cy, int cid, int cslot) {
                const float dx = __fsub_rn(cx, px), dy = __fsub_rn(cy, py);
                const float est = __fmaf_rn(dx, dx, __fmul_rn(dy, dy));
                if (est < thr) return;
                if (cid == id) return;
                const float d2 = dot2(dx, dy, dx, dy);
                if (d2 > best || (d2 == best && best_id >= 0 && cid < best_id)) {
                    best = d2; best_id = cid; best_slot = cslot;
                }
                const float t = __fmul_rn(est, 0.99999f);
                if (t > thr) thr = t;
            };
            // (A packed version of this loop — two candidates per FADD2 / FMUL2 — made no difference on a B200,
            // profiles/ab_r2n_*.json: the kernel waits on memory and barriers here, not on issue slots.)
            if (staged) {
                for (int k = 0; k < ncand; ++k) {
                    const float4 cd = s_cand[k];
                    consider(cd.x, cd.y, __float_as_int(cd.z), __float_as_int(cd.w));
                }
            } else {
                for (int k = 0; k < ncand; ++k) {
                    const float4 cd = __ldg(&cand[k]);
                    consider(cd.x, cd.y, __float_as_int(cd.z), __float_as_int(cd.w));
                }
            }
            if (best_id < 0) atomicCAS(&ctl->error, 0, -3);  // comm.cpp:273-285: null partner is UB in the reference
            b.diag_partner[s] = best_id;  // slot-indexed; lc_get_partners translates to ids (k_get_diag)
            const float4 pj = __ldg(&in.pvf[best_slot]);
            const float4 cj = __ldg(&in.c4[best_slot]);
            // comm.cpp:144-192 x(): COSINE_REL + PROX + SOFTSIGN_GOAL; partner (x, y) of a vec3 = (x, 0): quirk Q2
            const float rightx = facing.y, righty = -facing.x;
            const float dvx = __fsub_rn(pj.z, vx), dvy = __fsub_rn(0.f, vy);
            const float dpx = __fsub_rn(pj.x, px), dpy = __fsub_rn(0.f, py);
            float x[8];  // goal_dist is last frame's (ai.cpp:154 runs after comm::run)
            comm_features(P->flags, rightx, righty, facing.x, facing.y, dvx, dvy, dpx, dpy, c4.w, cj.x, cj.y, cj.z, x);
            float cn[3];
#pragma unroll
            for (int r = 0; r < 3; ++r) {  // comm.cpp:285 M_cv * x (Eigen 3.3 unrolled redux order, see oracle/canon.c)
                const float* m = &P->m_cv[8 * r];
                float p0 = __fmul_rn(m[0], x[0]), p1 = __fmul_rn(m[1], x[1]), p2 = __fmul_rn(m[2], x[2]),
                      p3 = __fmul_rn(m[3], x[3]), p4 = __fmul_rn(m[4], x[4]), p5 = __fmul_rn(m[5], x[5]),
                      p6 = __fmul_rn(m[6], x[6]), p7 = __fmul_rn(m[7], x[7]);
                float cp = __fadd_rn(__fadd_rn(__fadd_rn(p0, p1), __fadd_rn(p2, p3)),
                                     __fadd_rn(__fadd_rn(p4, p5), __fadd_rn(p6, p7)));
                // CommComp.h:20-27 buf_in: x / (1 + |2x|) + 0.5
                cn[r] = __fadd_rn(__fdiv_rn(cp, __fadd_rn(1.f, fabsf(__fmul_rn(2.f, cp)))), 0.5f);
            }
            float fb[2];
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                const float* m = &P->m_f[3 * r];
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
        const float gd = len2(__fsub_rn(goal.x, px), __fsub_rn(goal.y, py));
        c4.w = gd;
        int num_done;       // 1 = Agent::done() after this frame's lookahead (Agent.h:24-26)
        bool planned = true;  // Agent::has_plan()
        float eps = 0.f;
        if constexpr (!WALLS) {
            num_done = mt.y & 1;
            if (num_done && gd > 1.f && (id == ctl->replan1 || id == ctl->replan2)) num_done = 0;  // GMP::plan_one
            // LMP::lookahead (LMP.cpp:84-113) with plan = [final_goal]
            if (num_done < 1) {
                const float ox = __fsub_rn(px, goal.x), oy = __fsub_rn(py, goal.y);
                if (dot2(ox, oy, ox, oy) < 1.f) num_done = 1; else lgoal = goal;
            } else {
                lgoal = goal;
            }
            // LMP.cpp:261-273: one draw for both components
            eps = lcg_uniform((uint32_t)mt.z, -.01f, .01f);
            mt.z = (int)lcg_mul((uint32_t)mt.z, P->lcg_stride);
        } else {
            // Walled scene: the waypoint list lives by engine id; replans of this frame (ai.cpp:164-168) were served by the
            // host before the frame started (new list, num_done = 0, lost_frames = 0), S0 / S1 above still saw the old
            // done() through the flag bit, as demo::run and comm::run do in the reference.
            const WalledArrays& W = b.w;
            const int np = W.n_plan[id];
            int nd = W.num_done[id];
            W.lost[id] = W.lost_next[id];  // ai.cpp:161 (the gate evaluated it on this frame's start position)
            planned = np > 0;
            if (planned) {
                lcw::lookahead(walls, P->n_walls, P->r, px, py, reinterpret_cast<const float*>(W.plan + (size_t)id * WP_CAP), np, nd,
                               lgoal.x, lgoal.y);  // LMP.cpp:84-113
                // agent k of the table draws element draws + (planned agents before it) + 1 of the global stream
                const unsigned long long e = ctl->draws + (unsigned long long)(id - W.noplan_before[id]) + 1ull;
                eps = lcg_uniform(lcg_mul(P->rng0, lcg_pow_dev(e)), -.01f, .01f);
            } else {
                lgoal = make_float2(px, py);  // LMP.cpp:267-270: no plan, no draw, goal_vel = 0
            }
            W.num_done[id] = nd;
            num_done = planned && nd == np;
        }
        if (num_done == 1 && !(mt.y & 2)) {  // will be counted by demo::run at the start of the next frame
            mt.y |= 2;
            if (owned) {
                int k = atomicAdd(&ctl->ev_count, 1);
                b.events[k] = make_int2(id, mt.w);
            }
        }
        mt.y = (mt.y & ~1) | num_done;
        not_done_after = num_done == 0 && owned;
        const float dgx = __fsub_rn(__fadd_rn(lgoal.x, eps), px), dgy = __fsub_rn(__fadd_rn(lgoal.y, eps), py);
        const float dl = glm_max(1.0f, len2(dgx, dgy));
        float gvx = __fmul_rn(1.0f, __fdiv_rn(dgx, dl)), gvy = __fmul_rn(1.0f, __fdiv_rn(dgy, dl));
        if (WALLS && !planned) gvx = gvy = 0.f;
        const float goalFx = __fmul_rn(2.0f, __fsub_rn(gvx, vx)), goalFy = __fmul_rn(2.0f, __fsub_rn(gvy, vy));

        forcex = __fadd_rn(forcex, __fadd_rn(goalFx, ttcFx));  // ai.cpp:172
        forcey = __fadd_rn(forcey, __fadd_rn(goalFy, ttcFy));

        // ---- S3a: physics::simulate leap-frog (physics.cpp:41-51), mass == 1 ----
        const float2 vbh = in.vfh[s];
        float hx = __fadd_rn(vbh.x, __fmul_rn(LC_DT, forcex)), hy = __fadd_rn(vbh.y, __fmul_rn(LC_DT, forcey));
        const float h2 = dot3_xz(hx, hy, hx, hy);
        if (h2 > 1.f) {
            const float inv = __fdiv_rn(1.0f, __fsqrt_rn(h2));
            hx = __fmul_rn(hx, inv); hy = __fmul_rn(hy, inv);
        }
        const float nvx = __fadd_rn(__fdiv_rn(hx, 2.f), __fdiv_rn(vbh.x, 2.f));
        const float nvy = __fadd_rn(__fdiv_rn(hy, 2.f), __fdiv_rn(vbh.y, 2.f));
        const float npx = __fadd_rn(px, __fmul_rn(hx, LC_DT)), npy = __fadd_rn(py, __fmul_rn(hy, LC_DT));
        out.pv[s] = make_float4(npx, npy, nvx, nvy);
        out.vfh[s] = make_float2(hx, hy);
        out.goal[s] = goal;
        out.lgoal[s] = lgoal;
        out.facing[s] = facing;
        out.c4[s] = c4;
        out.meta[s] = mt;
    }
    // demo.cpp:667-669: all_done of the next frame needs the number of agents that are not done after lookahead
    const int nnd = __syncthreads_count(not_done_after);
    if (threadIdx.x == 0 && nnd) atomicAdd(&ctl->nnd_acc, nnd);
}

// ---------------------------------------------------------------------------------------------------------
// block-level reductions feeding the next frame's grid (bounding box of final positions, max speed)
// Per-block partials instead of global atomics: one hot cache line shared by ~10^4 warps serialises in a single
// L2 slice.  Every block stores one BlockRed; the next k_frame_begin (one block) folds them.
__device__ __forceinline__ void block_reduce_store(const Buffers& b, bool valid, float x, float y, float speed,
                                                   uint32_t lv_id, float disp = 0.f) {
    __shared__ uint32_t s_r[5], s_dmax;
    __shared__ unsigned long long s_lv;
    if (threadIdx.x == 0) { s_r[0] = s_r[1] = 0xffffffffu; s_r[2] = s_r[3] = s_r[4] = 0u; s_lv = 0ull; s_dmax = 0u; }
    __syncthreads();
    const uint32_t mnx = __reduce_min_sync(0xffffffffu, valid ? f2ord(x) : 0xffffffffu);
    const uint32_t mxx = __reduce_max_sync(0xffffffffu, valid ? f2ord(x) : 0u);
    const uint32_t mny = __reduce_min_sync(0xffffffffu, valid ? f2ord(y) : 0xffffffffu);
    const uint32_t mxy = __reduce_max_sync(0xffffffffu, valid ? f2ord(y) : 0u);
    // speed >= 0 (or NaN, which sorts above everything)
    const uint32_t sp = __reduce_max_sync(0xffffffffu, valid ? __float_as_uint(speed) : 0u);
    // highest-id not-done agent (physics.cpp:72-76 with vel_count == 0: the last one processed wins)
    const uint32_t top = __reduce_max_sync(0xffffffffu, lv_id);
    // displacement >= 0 (or NaN, which sorts above everything and forces a re-sort)
    const uint32_t dm = __reduce_max_sync(0xffffffffu, __float_as_uint(disp));
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&s_r[0], mnx); atomicMin(&s_r[1], mny);
        atomicMax(&s_r[2], mxx); atomicMax(&s_r[3], mxy); atomicMax(&s_r[4], sp);
        atomicMax(&s_dmax, dm);
    }
    if (top != 0u && lv_id == top) atomicMax(&s_lv, ((unsigned long long)top << 32) | __float_as_uint(speed));
    __syncthreads();
    if (threadIdx.x == 0) {
        BlockRed r;
        r.minx = s_r[0]; r.miny = s_r[1]; r.maxx = s_r[2]; r.maxy = s_r[3]; r.vmax = s_r[4]; r.dmax = s_dmax; r.lastvel = s_lv;
        b.blkred[blockIdx.x] = r;
        if (blockIdx.x == 0) b.ctl->n_blkred = (int)gridDim.x;
    }
}

template <int NT>
__device__ __forceinline__ void dd_pack_body(const Params* __restrict__ P, const Buffers& b);  // domain.cuh
__device__ __forceinline__ void dd_push_body(const Params* __restrict__ P, const Buffers& b);

// S3b: collision push-out (physics.cpp:58-69 + BVH.cpp:106-120), Jacobi: every agent tests the others at their
// integrated, not yet pushed positions; hits are applied in table (id) order.  Then physics.cpp:72-76.
template <bool DD, bool WALLS>
#ifndef LC_COLLIDE_BLOCKS
#define LC_COLLIDE_BLOCKS 10
#endif
__global__ void __launch_bounds__(128, LC_COLLIDE_BLOCKS) k_collide(const Params* __restrict__ P, Buffers b) {
    pdl_enter();
    Ctl* ctl = b.ctl;
    if (ctl->halt) return;
    const AgentArrays& in = set_in(b);
    const AgentArrays& out = set_out(b);
    stamp_entry(ctl, TS_COLLIDE);
    __shared__ lcw::Wall s_walls[WALLS ? WALL_SMEM : 1];
    const lcw::Wall* walls = nullptr;
    if constexpr (WALLS) walls = stage_walls(P, b, s_walls);
    const int n = P->n;
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = s < n;
    float px = 0.f, py = 0.f, speed = 0.f, disp = 0.f;
    uint32_t lv_id = 0u;
    if (valid) {
        const float4 me = out.pv[s];
        px = me.x; py = me.y;
        const float qx = me.x, qy = me.y;
        const int4 mt = out.meta[s];
        int nn = b.near_cnt[s];
        if (nn > 0) {
            float2 h = out.vfh[s];
            int hid[MAX_NEAR];
            float hax[MAX_NEAR], hay[MAX_NEAR];
            int cnt = 0;
            if (nn <= MAX_NEAR) {
                for (int k = 0; k < nn; ++k) {
                    const int j = b.near_idx[(size_t)s * MAX_NEAR + k];
                    const float4 o = __ldg(&out.pv[j]);
                    float ax, ay;
                    if (circ_circ_hit(qx, qy, o.x, o.y, P->r, P->RR, ax, ay)) {
                        hid[cnt] = __ldg(&out.meta[j].x); hax[cnt] = ax; hay[cnt] = ay; ++cnt;
                    }
                }
                int last = -1;
                for (int k = 0; k < cnt; ++k) {
                    int bi = 0, bid = 0x7fffffff;
                    for (int q = 0; q < cnt; ++q) {
                        int v = hid[q];
                        if (v > last && v < bid) { bid = v; bi = q; }
                    }
                    push_away(px, py, h.x, h.y, hax[bi], hay[bi]);
                    last = bid;
                }
            } else {
                // near-list overflow (more than MAX_NEAR agents within ~0.34 m): walk the cell block, one hit per pass
                const int nx = ctl->nx, ny = ctl->ny, st = ctl->stencil, stx = ctl->stencil_x;
                const float4 old = in.pvf[s];  // frame-start position decides the cell block
                int ix = min(max((int)floorf((old.x - ctl->x0) / ctl->hx), 0), nx - 1);
                int iy = min(max((int)floorf((old.y - ctl->y0) / ctl->h), 0), ny - 1);
                int last = -1;
                for (;;) {
                    int bid = 0x7fffffff;
                    float bax = 0.f, bay = 0.f;
                    for (int yy = max(iy - st, 0); yy <= min(iy + st, ny - 1); ++yy) {
                        const int j0 = b.cell_start[yy * nx + max(ix - stx, 0)], j1 = b.cell_start[yy * nx + min(ix + stx, nx - 1) + 1];
                        for (int j = j0; j < j1; ++j) {
                            if (j == s) continue;
                            const float4 o = __ldg(&out.pv[j]);
                            float ax, ay;
                            if (circ_circ_hit(qx, qy, o.x, o.y, P->r, P->RR, ax, ay)) {
                                int v = __ldg(&out.meta[j].x);
                                if (v > last && v < bid) { bid = v; bax = ax; bay = ay; }
                            }
                        }
                    }
                    if (bid == 0x7fffffff) break;
                    push_away(px, py, h.x, h.y, bax, bay);
                    last = bid;
                    ++cnt;
                }
            }
            if (cnt) {
                out.vfh[s] = h;
                if (!(mt.y & META_GHOST)) atomicAdd(&ctl->coll_hits, (unsigned long long)cnt);
            }
        }
        if constexpr (WALLS) {
            // physics.cpp:60,67-69: static_bvh->query at the integrated position, pushed out after the agent hits.
            // BVH::query_(Circ*) (BVH.cpp:106-138): depth first, left before right; inner nodes are entered through the
            // same single-point test as the leaves (quirk Q8), so hits can be missed exactly as the reference misses them.
            float2 h = out.vfh[s];
            int wcnt = 0, sp = 0;
            int stack[48];
            stack[sp++] = 0;
            while (sp > 0) {
                const BvhNode nd = b.w.bvh[stack[--sp]];
                float nx, ny;
                if (nd.right < 0) {
                    if (lcw::circ_wall_hit(qx, qy, P->r, walls[nd.left], nx, ny)) { push_away(px, py, h.x, h.y, nx, ny); ++wcnt; }
                } else if (lcw::circ_wall_hit(qx, qy, P->r, lcw::Wall{nd.ox, nd.oy, nd.w, nd.h}, nx, ny) && sp + 2 <= 48) {
                    stack[sp++] = nd.right;
                    stack[sp++] = nd.left;
                }
            }
            if (wcnt) {
                out.vfh[s] = h;
                atomicAdd(&ctl->wall_hits, (unsigned long long)wcnt);
            }
        }
        // k_force consumed the pair list: next frame's k_pairs counts from zero again.  (Here, not in k_force right after its
        // load: there the store cost 30 us at 1 M agents, profiles/ab_r2n_*.json.)
        b.pair_cnt[s] = 0;
        out.pvf[s] = make_float4(px, py, me.z, me.w);  // the state all_sync publishes = the next frame's frame-start state
        {   // ... and its pair-interleaved copy for the next frame's pair loop
            float* q = reinterpret_cast<float*>(out.pq + 2 * (s >> 1)) + (s & 1);
            q[0] = px; q[2] = py; q[4] = me.z; q[6] = me.w;
        }
        {
            const float ex = px - qx, ey = py - qy;
            disp = LC_DT * 1.0001f + sqrtf(ex * ex + ey * ey);
        }
        if constexpr (WALLS) {  // replan gate of the coming frame (ai.cpp:152-163) on the position all_sync publishes
            const float2 g = out.goal[s];
            walled_gate(P, b, walls, mt.x, px, py, g.x, g.y);
        }
        // The id -> slot table is a scattered 4-byte store per agent (a 32-byte sector each).  Its only reader inside the
        // frame loop is the time-out walk of k_frame_begin (demo.cpp:691-700, id order); lc_set_state rebuilds it in
        // k_snapshot.  So it is written only by the frame before the time-out one (and every frame after it).
        if (!DD && __fmul_rn(LC_DT, (float)(ctl->frame + 1u)) > P->max_duration) b.slot_of_id[mt.x] = s;
        if (!DD) {
            // ai.cpp:154,161-163 for the coming frame: need_replan = done && goal_dist > 1 on the position it starts from
            // (wall-free: the lost_frames clause cannot fire); the two lowest ids are served (ai.cpp:151-152,164)
            if (mt.y & 1) {
                const float2 g = out.goal[s];
                if (len2(__fsub_rn(g.x, px), __fsub_rn(g.y, py)) > 1.f) {
                    const int old = atomicMin(&ctl->replan1n, mt.x);
                    const int loser = old < mt.x ? mt.x : old;
                    if (loser != 0x7fffffff) atomicMin(&ctl->replan2n, loser);
                }
            }
            // stage-1 comm-partner candidates of the coming frame, if it keeps the slots (k_frame_begin re-sorts after all
            // when this frame turns out to have moved somebody by more than CAND_MARGIN)
            if (!(P->flags & 1u) && ctl->since + 1 < P->rebin_every &&
                !candidate_inside(ctl, px, py, ctl->delta + ctl->skin + CAND_MARGIN)) {
                const int k = atomicAdd(&ctl->n_cand_next, 1);
                b.cand[k] = make_float4(px, py, __int_as_float(mt.x), __int_as_float(s));
            }
        } else if (!(P->flags & 1u) && !ctl->rebin_next && !(mt.y & META_GHOST) &&
                   !candidate_inside(ctl, px, py, ctl->delta + ctl->skin + CAND_MARGIN)) {
            // decomposed crowd: the same early list, over the owned agents (ghosts are published by their owner)
            const int k = atomicAdd(&ctl->n_cand_next, 1);
            b.cand[k] = make_float4(px, py, __int_as_float(mt.x), __int_as_float(s));
        }
        speed = __fsqrt_rn(dot3_xz(me.z, me.w, me.z, me.w));
        if (!(mt.y & 1) && !(mt.y & META_GHOST)) {  // physics.cpp:72-76
            lv_id = (uint32_t)(mt.x + 1);
            // a non-finite speed anywhere but at the last agent poisons the running average (NaN * 0): counted here,
            // resolved in fold_avg_velocity
            if (!(speed <= LC_FLT_MAX)) atomicAdd(&ctl->nonfinite, 1);
        }
    }
    // how far the agent got this frame at most: the integration step (|vel_forehalf| <= 1 after the clamp) plus the push-out
    block_reduce_store(b, valid && px == px && py == py, px, py, speed, lv_id, valid ? disp : 0.f);
    // domain mode: route every owned agent's record of the NEXT frame to its owner / ghost hosts (reads back what this
    // thread just stored: posf, vfh)
    if (DD) {
        if (ctl->rebin_next) dd_pack_body<128>(P, b);
        else dd_push_body(P, b);
    }
}

// reductions of an arbitrary state (after lc_create / lc_set_state)
__global__ void __launch_bounds__(128) k_init_reduce(const Params* __restrict__ P, Buffers b) {
    const AgentArrays& live = set_live(b);
    Ctl* ctl = b.ctl;
    const int n = P->n;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = i < n;
    float x = 0.f, y = 0.f, speed = 0.f;
    int nd = 0;
    if (valid) {
        const float4 pv = live.pvf[i];
        x = pv.x; y = pv.y;
        speed = __fsqrt_rn(dot3_xz(pv.z, pv.w, pv.z, pv.w));
        const int4 mt = live.meta[i];
        nd = !(mt.y & 1);
        b.slot_of_id[mt.x] = i;
        if (!nd && P->dd_world <= 1) {  // replan gate of the coming frame (later frames: k_collide)
            const float2 g = live.goal[i];
            if (len2(__fsub_rn(g.x, x), __fsub_rn(g.y, y)) > 1.f) {
                const int old = atomicMin(&ctl->replan1n, mt.x);
                const int loser = old < mt.x ? mt.x : old;
                if (loser != 0x7fffffff) atomicMin(&ctl->replan2n, loser);
            }
        }
    }
    const int cnt = __syncthreads_count(nd);
    if (threadIdx.x == 0 && cnt) atomicAdd(&ctl->nnd_acc, cnt);
    block_reduce_store(b, valid && x == x && y == y, x, y, speed, 0u);
}

// replan gate of the coming frame from an arbitrary state (after lc_create / lc_set_state); later frames: k_collide's tail
__global__ void __launch_bounds__(128) k_walled_gate(const Params* __restrict__ P, Buffers b) {
    const AgentArrays& live = set_live(b);
    __shared__ lcw::Wall s_walls[WALL_SMEM];
    const lcw::Wall* walls = stage_walls(P, b, s_walls);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P->n) return;
    const float4 p = live.pvf[i];
    const float2 g = live.goal[i];
    walled_gate(P, b, walls, live.meta[i].x, p.x, p.y, g.x, g.y);
}

__global__ void __launch_bounds__(1024) k_finalize(const Params* __restrict__ P, Buffers b) {
    // lc_get_scores between frames: fold the last frame's lastvel partial, but leave the bbox/vmax partials for the
    // next k_frame_begin (consume is idempotent on those: min/max into Ctl, which frame_begin resets after use)
    consume_block_reductions(b.ctl, b);
    if (threadIdx.x == 0) fold_avg_velocity(b.ctl);
}

// state in id order for the host (POOL.all_sync view, Pool.cpp:69-91)
__global__ void __launch_bounds__(256) k_get_state(const Params* __restrict__ P, Buffers b, float2* pos, float2* vel,
                                                    float2* vfh, float* c3, float2* facing, float2* lgoal,
                                                    float* goal_dist, int* num_done, int* owned) {
    const AgentArrays& live = set_live(b);
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= P->n) return;
    const int4 mt = live.meta[s];
    if (mt.y & META_GHOST) return;  // the owner reports it
    const int id = b.ext_of ? b.ext_of[mt.x] : mt.x;  // the caller's agent index
    if (owned) owned[id] = 1;
    const float4 pv = live.pvf[s];
    const float4 c4 = live.c4[s];
    if (pos) pos[id] = make_float2(pv.x, pv.y);
    if (vel) vel[id] = make_float2(pv.z, pv.w);
    if (vfh) vfh[id] = live.vfh[s];
    if (c3) { c3[3 * id] = c4.x; c3[3 * id + 1] = c4.y; c3[3 * id + 2] = c4.z; }
    if (facing) facing[id] = live.facing[s];
    if (lgoal) lgoal[id] = live.lgoal[s];
    if (goal_dist) goal_dist[id] = c4.w;
    if (num_done) num_done[id] = P->n_walls > 0 ? b.w.num_done[mt.x] : (mt.y & 1);  // Agent::num_done (waypoints passed)
}

// Diagnostics of the last frame (comm partner, |{j : ttc_ij <= 5}|) are kept per slot by k_force — an id-indexed store
// is a 32-byte sector per agent and frame — and brought into id order only when the host asks for them.  Ids this rank
// does not own keep whatever the table held.
__global__ void __launch_bounds__(256) k_get_diag(const Params* __restrict__ P, Buffers b) {
    const AgentArrays& live = set_live(b);
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= P->n) return;
    const int4 mt = live.meta[s];
    if (mt.y & META_GHOST) return;
    b.partner_id[mt.x] = b.diag_partner[s];
    b.nbr_cnt[mt.x] = b.diag_cnt[s];
}

// domain mode, host-facing shard: (id, pos, vel) of the agents this rank owns, compacted (order arbitrary)
__global__ void __launch_bounds__(256) k_get_owned(const Params* __restrict__ P, Buffers b, int* ids, float2* pos, float2* vel,
                                                    int* count) {
    const AgentArrays& live = set_live(b);
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    const int4 mt = s < P->n ? live.meta[s] : make_int4(0, META_GHOST, 0, 0);
    const bool own = !(mt.y & META_GHOST);
    const unsigned m = __ballot_sync(0xffffffffu, own);
    int base = 0;
    if ((threadIdx.x & 31) == 0 && m) base = atomicAdd(count, __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (own) {
        const int k = base + __popc(m & ((1u << (threadIdx.x & 31)) - 1u));
        const float4 pv = live.pvf[s];
        ids[k] = b.ext_of ? b.ext_of[mt.x] : mt.x;
        pos[k] = make_float2(pv.x, pv.y);
        vel[k] = make_float2(pv.z, pv.w);
    }
}
void launch_get_owned(const Params* dP, int n, const Buffers& b, int* ids, float2* pos, float2* vel, int* count,
                      cudaStream_t st, long long* launches) {
    k_get_owned<<<(n + 255) / 256, 256, 0, st>>>(dP, b, ids, pos, vel, count);
    if (launches) *launches += 1;
}

// per-agent private LCG streams: agent id draws x0 * 16807^((frame-1)*n + id + 1) in frame `frame` (Seeder.cpp, LMP.cpp:261-263)
__global__ void __launch_bounds__(256) k_set_rng(const Params* __restrict__ P, Buffers b, uint32_t base) {
    const AgentArrays& live = set_live(b);
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= P->n) return;
    int4 mt = live.meta[s];
    uint32_t e = (uint32_t)mt.x + 1u, a = 16807u, x = base;
    while (e) {
        if (e & 1u) x = lcg_mul(x, a);
        a = lcg_mul(a, a);
        e >>= 1;
    }
    mt.z = (int)x;
    live.meta[s] = mt;
}

// state upload in id order (slot == id): demo.cpp:563-581 create_robo / teacher forcing
__global__ void __launch_bounds__(256) k_set_state(const Params* __restrict__ P, Buffers b, const float2* pos,
                                                    const float2* vel, const float2* vfh, const float* c3,
                                                    const float2* facing, const float2* lgoal, const float* goal_dist,
                                                    const int* num_done, const float2* goal, const float* min_time,
                                                    int init, const int* ids) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P->n) return;
    // bring the agent with id i to slot i
    const AgentArrays& live = set_live(b);
    const AgentArrays& snap = b.set[b.ctl->live ^ 1];  // k_snapshot's copy of the state before the call
    const int so = init ? i : b.slot_of_id[i];
    float4 pv = init ? make_float4(0, 0, 0, 0) : snap.pvf[so];
    float2 p = init ? make_float2(0, 0) : make_float2(pv.x, pv.y);
    float2 h = init ? make_float2(0, 0) : snap.vfh[so];
    float2 g = init ? make_float2(0, 0) : snap.goal[so];
    float2 lg = init ? make_float2(0, 0) : snap.lgoal[so];
    float2 fc = init ? make_float2(0, 0) : snap.facing[so];
    float4 c4 = init ? make_float4(0, 0, 0, 0) : snap.c4[so];
    int4 mt = init ? make_int4(ids ? ids[i] : i, 0, 0, 0) : snap.meta[so];
    // host tables: at creation (init) they were gathered by the host in local order; afterwards they are indexed by the
    // caller's agent index
    const int e = (init || !b.ext_of) ? i : b.ext_of[i];
    if (pos) p = pos[e];
    if (vel) { pv.z = vel[e].x; pv.w = vel[e].y; }
    if (vfh) h = vfh[e];
    if (c3) { c4.x = c3[3 * e]; c4.y = c3[3 * e + 1]; c4.z = c3[3 * e + 2]; }
    if (facing) fc = facing[e];
    if (lgoal) lg = lgoal[e];
    if (goal_dist) c4.w = goal_dist[e];
    if (num_done) {
        if (P->n_walls > 0 && !init) {  // walled: the waypoint counter itself; the flag bit is Agent::done()
            const int np = b.w.n_plan[mt.x];
            b.w.num_done[mt.x] = num_done[e];
            mt.y = (mt.y & ~1) | ((np > 0 && num_done[e] == np) ? 1 : 0);
        } else {
            mt.y = (mt.y & ~1) | (num_done[e] & 1);
        }
    }
    if (goal) g = goal[e];
    if (min_time) mt.w = __float_as_int(min_time[e]);
    live.pvf[i] = make_float4(p.x, p.y, pv.z, pv.w);
    live.vfh[i] = h;
    live.goal[i] = g;
    live.lgoal[i] = lg;
    live.facing[i] = fc;
    live.c4[i] = c4;
    live.meta[i] = mt;
    if (!init) {  // slot == id from here on: the per-slot diagnostics follow (launch_set_state ran k_get_diag first)
        b.diag_partner[i] = b.partner_id[i];
        b.diag_cnt[i] = b.nbr_cnt[i];
    }
}

// snapshot cur -> srt (posf folded into pv) so k_set_state can permute in place
__global__ void __launch_bounds__(256) k_snapshot(const Params* __restrict__ P, Buffers b) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P->n) return;
    const AgentArrays& live = set_live(b);
    const AgentArrays& snap = b.set[b.ctl->live ^ 1];
    snap.pvf[i] = live.pvf[i];
    snap.vfh[i] = live.vfh[i];
    snap.goal[i] = live.goal[i];
    snap.lgoal[i] = live.lgoal[i];
    snap.facing[i] = live.facing[i];
    snap.c4[i] = live.c4[i];
    const int4 mt = live.meta[i];
    snap.meta[i] = mt;
    b.slot_of_id[mt.x] = i;  // k_reorder does not maintain the table (see there)
}

__global__ void __launch_bounds__(1024) k_reset_reductions(Buffers b) {
    Ctl* ctl = b.ctl;
    consume_block_reductions(ctl, b);  // keeps the avg_velocity of the last executed frame
    if (threadIdx.x == 0) {
        fold_avg_velocity(ctl);
        ctl->minx = ctl->miny = 0xffffffffu;
        ctl->maxx = ctl->maxy = 0u;
        ctl->vmax = 0u;
        ctl->nnd_acc = 0;
        ctl->force_rebin = 1;              // arbitrary new state: the next frame re-sorts the grid
        ctl->dmax = 0u;
        ctl->n_cand_next = 0;
        ctl->rebin_next = 1;               // decomposed crowds: the first frame routes and sorts
        ctl->replan1n = ctl->replan2n = 0x7fffffff;  // k_init_reduce re-evaluates the gate on the new state
        ctl->wr1 = ctl->wr2 = 0x7fffffff;  // walled scenes: k_walled_gate re-evaluates the requests on the new state
        ctl->w_served = 0;
        for (int k = 0; k < HULL_K; ++k) ctl->ext_val[k] = -CUDART_INF_F;  // arbitrary new state: no gate next frame
    }
}

#include "domain.cuh"

// ---------------------------------------------------------------------------------------------------------
static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

DomainLayout domain_layout(int world, int cap_self, int cap_halo, int n_ids) {
    DomainLayout L{};
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
    L.cap_self = cap_self; L.cap_halo = cap_halo;
    for (int p = 0; p < 2; ++p) {
        L.nblob[p] = take(sizeof(NeighbourBlob) * 3);
        L.nflag[p] = take(sizeof(unsigned long long) * 3);
        L.blob[p] = take(sizeof(DomainBlob) * DD_MAX_WORLD);
        L.cflag[p] = take(sizeof(unsigned long long) * DD_MAX_WORLD);
        for (int r = 0; r < 3; ++r) L.region[p][r] = r == 1 ? 0 : take(sizeof(float4) * DD_PLANES * (size_t)cap_halo);  // stayers do not travel
        L.events[p] = take(sizeof(int2) * (size_t)DD_MAX_WORLD * DD_EV_CAP);
        L.cand[p] = take(sizeof(float4) * 2 * (size_t)DD_MAX_WORLD * DD_CAND_CAP);
    }
    L.ev_by_id = take(sizeof(float) * (size_t)n_ids);
    L.total = off;
    (void)world;
    return L;
}

void upload_dirs() {
    float dx[HULL_K], dy[HULL_K];
    for (int k = 0; k < HULL_K; ++k) {
        double th = 2.0 * 3.14159265358979323846 * k / HULL_K;
        dx[k] = (float)cos(th);
        dy[k] = (float)sin(th);
    }
    cudaMemcpyToSymbol(c_dirx, dx, sizeof dx);
    cudaMemcpyToSymbol(c_diry, dy, sizeof dy);
    float ex[HULL_K2], ey[HULL_K2];
    for (int k = 0; k < HULL_K2; ++k) {
        double th = 2.0 * 3.14159265358979323846 * k / HULL_K2;
        ex[k] = (float)cos(th);
        ey[k] = (float)sin(th);
    }
    cudaMemcpyToSymbol(c_dir2x, ex, sizeof ex);
    cudaMemcpyToSymbol(c_dir2y, ey, sizeof ey);
}

// Kernel launch, optionally with the programmatic-dependent-launch attribute (pdl_enter): LC_PDL=1 in the environment.
static bool pdl_enabled() {
    static const bool on = [] { const char* e = getenv("LC_PDL"); return e && e[0] == '1'; }();
    return on;
}
template <class... KArgs, class... Args>
static void launch(void (*kernel)(KArgs...), unsigned grid, unsigned block, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

void launch_frame(Params* dP, int n, const Buffers& b, cudaStream_t st, long long* launches, cudaEvent_t* ev,
                  const SideStream* side, bool walled) {
    // n = launch width (the capacity in domain mode; kernels read the live count from Params::n on the device)
    // ev (optional, FRAME_KERNELS + 1 events): per-kernel device timing for bench.py's roofline leg
    const bool dd = b.dd != nullptr;
    const Params* cP = dP;
    if (ev) cudaEventRecord(ev[0], st);
    if (dd) {  // boundary exchange over peer memory (domain.cuh): the records were packed by the previous k_collide
        launch(k_dd_sync_begin, 1, 1024, st, dP, b);   // publish + wait + merge + k_frame_begin's work
    } else {
        launch(k_frame_begin, 1, 1024, st, cP, b);
    }
    if (ev) cudaEventRecord(ev[1], st);
    launch(k_bin, cdiv(n, 1024), 1024, st, cP, b);
    if (ev) cudaEventRecord(ev[2], st);
    launch(k_scan_hull, SCAN_TILES + 1, 1024, st, cP, b);
    if (ev) cudaEventRecord(ev[3], st);
    launch(k_reorder, cdiv(n, 256), 256, st, cP, b);
    if (ev) cudaEventRecord(ev[4], st);
    if (side && !ev) {  // (the profiled path stays serial so that the per-kernel times mean something)
        cudaEventRecord(side->fork, st);
        cudaStreamWaitEvent(side->st, side->fork, 0);
        k_refine<<<1, 1024, 0, side->st>>>(dP, b);  // + candidate all-gather in domain mode
        cudaEventRecord(side->join, side->st);
        launch(k_pairs, cdiv(n, 128), 128, st, cP, b);
        cudaStreamWaitEvent(st, side->join, 0);
        // two predecessors: a plain (full) dependency on both
        if (walled) k_force<true><<<cdiv(n, 128), 128, 0, st>>>(dP, b);
        else k_force<false><<<cdiv(n, 128), 128, 0, st>>>(dP, b);
    } else {
        launch(k_refine, 1, 1024, st, cP, b);
        if (ev) cudaEventRecord(ev[5], st);
        launch(k_pairs, cdiv(n, 128), 128, st, cP, b);
        if (ev) cudaEventRecord(ev[6], st);
        if (walled) launch(k_force<true>, cdiv(n, 128), 128, st, cP, b);
        else launch(k_force<false>, cdiv(n, 128), 128, st, cP, b);
    }
    if (ev) cudaEventRecord(ev[7], st);
    if (dd) launch(k_collide<true, false>, cdiv(n, 128), 128, st, cP, b);
    else if (walled) launch(k_collide<false, true>, cdiv(n, 128), 128, st, cP, b);
    else launch(k_collide<false, false>, cdiv(n, 128), 128, st, cP, b);
    if (ev) cudaEventRecord(ev[8], st);
    if (launches) *launches += dd ? FRAME_KERNELS + DD_KERNELS : FRAME_KERNELS;
}

__global__ void __launch_bounds__(256) k_to_external(const int* __restrict__ ext_of, const int* __restrict__ src, int* __restrict__ dst,
                                                     int rows, int width, int values) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (long long)rows * width) return;
    const int i = (int)(t / width), k = (int)(t % width);
    int v = src[t];
    if (values && v >= 0) v = ext_of[v];
    dst[(size_t)ext_of[i] * width + k] = v;
}
void launch_to_external(const Buffers& b, const int* src, int* dst, int rows, int width, int values, cudaStream_t st, long long* launches) {
    const long long tot = (long long)rows * width;
    k_to_external<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(b.ext_of, src, dst, rows, width, values);
    if (launches) *launches += 1;
}

void launch_get_diag(const Params* dP, int n, const Buffers& b, cudaStream_t st, long long* launches) {
    k_get_diag<<<cdiv(n, 256), 256, 0, st>>>(dP, b);
    if (launches) *launches += 1;
}

void launch_dd_pack(const Params* dP, int n, const Buffers& b, cudaStream_t st, long long* launches) {
    k_dd_pack<<<cdiv(n, 256), 256, 0, st>>>(dP, b);  // before the first frame only; afterwards k_collide packs
    if (launches) *launches += 1;
}

void launch_walled_gate(const Params* dP, int n, const Buffers& b, cudaStream_t st, long long* launches) {
    k_walled_gate<<<cdiv(n, 128), 128, 0, st>>>(dP, b);
    if (launches) *launches += 1;
}

void launch_init_reduce(const Params* dP, int n, const Buffers& b, cudaStream_t st, long long* launches) {
    k_reset_reductions<<<1, 1024, 0, st>>>(b);
    k_init_reduce<<<cdiv(n, 128), 128, 0, st>>>(dP, b);
    if (launches) *launches += 2;
}

__global__ void __launch_bounds__(1024) k_consume_only(Buffers b) { consume_block_reductions(b.ctl, b); }

void launch_finalize(const Params* dP, const Buffers& b, cudaStream_t st, long long* launches) {
    // domain mode: avg_velocity of the last frame is a cross-rank quantity; only fold the block partials into Ctl
    // (lc_dd_get_partials reads them, the host merges the ranks) and leave the rest to the next k_dd_sync_begin
    if (b.dd) k_consume_only<<<1, 1024, 0, st>>>(b);
    else k_finalize<<<1, 1024, 0, st>>>(dP, b);
    if (launches) *launches += 1;
}

void launch_get_state(const Params* dP, int n, const Buffers& b, float2* pos, float2* vel, float2* vfh, float* c3,
                      float2* facing, float2* lgoal, float* goal_dist, int* num_done, cudaStream_t st,
                      long long* launches, int* owned) {
    k_get_state<<<cdiv(n, 256), 256, 0, st>>>(dP, b, pos, vel, vfh, c3, facing, lgoal, goal_dist, num_done, owned);
    if (launches) *launches += 1;
}

void launch_set_state(const Params* dP, int n, const Buffers& b, const float2* pos, const float2* vel,
                      const float2* vfh, const float* c3, const float2* facing, const float2* lgoal,
                      const float* goal_dist, const int* num_done, const float2* goal, const float* min_time, int init,
                      cudaStream_t st, long long* launches, const int* ids) {
    if (!init) {
        k_get_diag<<<cdiv(n, 256), 256, 0, st>>>(dP, b);  // per-slot diagnostics -> id order, before the slots change
        k_snapshot<<<cdiv(n, 256), 256, 0, st>>>(dP, b);
        if (launches) *launches += 2;
    }
    k_set_state<<<cdiv(n, 256), 256, 0, st>>>(dP, b, pos, vel, vfh, c3, facing, lgoal, goal_dist, num_done, goal,
                                              min_time, init, ids);
    if (launches) *launches += 1;
}

void launch_set_rng(const Params* dP, int n, const Buffers& b, uint32_t base, cudaStream_t st, long long* launches) {
    k_set_rng<<<cdiv(n, 256), 256, 0, st>>>(dP, b, base);
    if (launches) *launches += 1;
}

}  // namespace lc
