// Spatial domain decomposition of ONE crowd over several GPUs (BASELINE.json configs[4]): x strips, one rank per
// GPU, boundary agents exchanged every frame by kernels that store straight into the neighbours' memory over
// NVLink (peer mappings of one mailbox allocation per rank) — no host round trip, no NCCL call on the data path.
// Included at the end of crowd.cu (same translation unit: it reuses the block reductions).
//
// Frame f on every rank (all inside the CUDA graph):
//   (pack)        tail of the PREVIOUS frame's k_collide (dd_pack_body; k_dd_pack only before the very first frame): an
//                 owned agent that stays on this rank is merely marked (keep[slot] = 1: next frame's sort reads it in
//                 place); one that migrates to an adjacent strip, or lies within `dd_halo` of one, is stored as an
//                 80-byte record into that neighbour's mailbox.  Only boundary agents ever travel.  At the time-out
//                 frame the finish times of demo.cpp:691-700 go into the scoring rank's id-indexed table.
//   (frame start) k_dd_sync_begin, one block: tells the two ADJACENT ranks how many records they got and bounds of what
//                 those can be (max speed, y extent), waits for THEIR two flags only, sets this frame's record counts and
//                 grid.  Nothing global is needed to start a frame: the grid of a rank is a matter of the records it
//                 holds, and any cutoff that covers them yields a candidate superset, i.e. the same result.
//   k_bin, k_scan_hull, k_reorder sort stayers (from `cur`) and received records (from the mailbox, in place) into cell order.
//   (global)      dd_global_body, the tail of k_refine on the side stream BESIDE k_pairs: every rank publishes its
//                 comm-partner candidates (comm.cpp:201-220 is a GLOBAL arg-max) and its scoring inputs (not-done count,
//                 avg_velocity inputs, replan requests, done-events) to all ranks, waits for all of theirs — the only
//                 all-rank wait of the frame, off the critical path — merges, and runs S0 (demo::run: Welford on rank 0,
//                 all_done latch, replan gate).  Its consumers are k_force (partner, replan gate) and the next frame
//                 (latch), both after the join.
//   k_force, k_collide: ghosts are advanced too — the collision pass needs their integrated positions — but never
//                 counted; their next state always comes from the owner.
// Why the result is bit-identical to the single-GPU run: an owned agent sees every neighbour within the interaction
// cutoff (ghost band = cutoff at the speed bound + collision reach), forces and hits are summed in id order, the
// comm partner is the global arg-max over the union of the per-rank hull survivors, and the replan gate, done count
// and scoring inputs are merged across ranks before they are used.  tests/test_gpu_domain.py checks it, bench.py's
// state_hash shows it at every N.
//
// Mailboxes are double-buffered by frame parity: no rank can pass k_force of frame f before every rank has reached the
// global exchange of frame f, so the ranks are never more than a frame apart and parity f + 2 is only written after
// parity f has been consumed.
#pragma once

__device__ __forceinline__ unsigned long long dd_now_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ unsigned long long dd_ld_flag(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void dd_st_flag(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// spin until *p == want; false on time-out (a peer died or the hosts enqueued different frame counts)
__device__ __forceinline__ bool dd_wait_flag(const unsigned long long* p, unsigned long long want, unsigned long long timeout_ns) {
    // Poll back to back; look at the clock (and yield for a moment) only every 32 polls.
    const unsigned long long t0 = dd_now_ns();
    for (unsigned it = 1;; ++it) {
        if (dd_ld_flag(p) == want) return true;
        if ((it & 31u) == 0u) {
            if (dd_now_ns() - t0 > timeout_ns) return false;
            __nanosleep(32);
        }
    }
}

__device__ __forceinline__ bool dd_active(const Params* P, const Ctl* ctl) {
    return !(ctl->finished && !(P->flags & 2u));  // frames after the all_done frame are no-ops on every rank alike
}

// which of (rank-1, self, rank+1) gets an owned agent at x, and in which role (1 = owner, 2 = ghost)
__device__ __forceinline__ void dd_route(const Params* P, float x, int role[3]) {
    const int me = P->dd_rank, w = P->dd_world;
    int own = me;
    if (x < P->dd_xb[me] && me > 0) own = me - 1;
    else if (!(x < P->dd_xb[me + 1]) && me < w - 1) own = me + 1;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int q = me - 1 + k;
        role[k] = 0;
        if (q < 0 || q >= w) continue;
        if (q == own) role[k] = 1;
        else if (x >= P->dd_xb[q] - P->dd_halo && x < P->dd_xb[q + 1] + P->dd_halo) role[k] = 2;
    }
}

// What every owned agent contributes to the coming frame's global exchange, whichever way its record travels
__device__ __forceinline__ void dd_owned_tail(const Params* __restrict__ P, const Buffers& b, const int4 mt, const float2 p, const float2 g,
                                              const uint32_t frame) {
    Ctl* ctl = b.ctl;
    DomainDev* dd = b.dd;
    // replan gate (ai.cpp:151-168): done && goal_dist > 1, the two lowest ids over ALL ranks; this rank's two lowest here
    if (mt.y & 1) {
        const float gd = len2(__fsub_rn(g.x, p.x), __fsub_rn(g.y, p.y));
        if (gd > 1.f) {
            const int old = atomicMin(&dd->replan[0], mt.x);
            const int loser = old < mt.x ? mt.x : old;
            if (loser != 0x7fffffff) atomicMin(&dd->replan[1], loser);
        }
    }
    // demo.cpp:691-700 at the time-out frame: finish time of every not-done agent, folded in id order by the scoring rank
    if (!ctl->finished && !(mt.y & 1)) {
        const float fake_time = __fmul_rn(LC_DT, (float)frame);
        if (fake_time > P->max_duration) {
            const float l = len2(__fsub_rn(g.x, p.x), __fsub_rn(g.y, p.y));
            reinterpret_cast<float*>(dd->mbox[0] + dd->lay.ev_by_id)[mt.x] =
                __fsub_rn(__fadd_rn(__fmul_rn(2.f, l), fake_time), __int_as_float(mt.w));
        }
    }
}

// Body of the pack step, for a block of NT threads.  Runs at the END of k_collide (the agent's final position of frame f
// is what decides its owner and ghost roles in frame f + 1), and as a kernel of its own before the very first frame.
template <int NT>
__device__ __forceinline__ void dd_pack_body(const Params* __restrict__ P, const Buffers& b) {
    Ctl* ctl = b.ctl;
    DomainDev* dd = b.dd;
    if (!dd_active(P, ctl)) return;
    const int n = P->n;
    if ((int)(blockIdx.x * blockDim.x) >= n) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int me = P->dd_rank;
    const AgentArrays& cur = set_live(b);   // the frame that just ended wrote it (before the first frame: the initial state)
    const uint32_t frame = ctl->frame + 1u;
    const int par = frame & 1u;
    // Slot allocation in the neighbours' regions keeps the agents' order (ballot prefix inside the warp, warp totals, one
    // global atomicAdd per block and destination); stayers are only counted.
    __shared__ int s_wcnt[3][NT / 32], s_base[3];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int role[3] = {0, 0, 0}, rank_in_blk[3] = {0, 0, 0};
    int4 mt = make_int4(0, 0, 0, 0);
    float2 p = make_float2(0.f, 0.f), g = make_float2(0.f, 0.f);
    bool owned = false;
    if (i < n) {
        mt = cur.meta[i];
        owned = !(mt.y & META_GHOST);
        if (owned) {
            const float4 pf = cur.pvf[i];
            p = make_float2(pf.x, pf.y);
            g = cur.goal[i];
            dd_route(P, p.x, role);
        }
        // stays in the local set: as owner, or — having just crossed into the adjacent strip while still inside our ghost
        // band — as a ghost of its new owner (who only learns of it next frame).  Next frame's sort reads it from `cur`.
        // (`cur` itself is left alone: until the next frame runs, this rank is still the one that reports the agent)
        dd->keep[i] = (unsigned char)role[1];
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const unsigned m = __ballot_sync(0xffffffffu, role[k] != 0);
        rank_in_blk[k] = __popc(m & ((1u << lane) - 1u));
        if (lane == 0) s_wcnt[k][wid] = __popc(m);
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        int tot = 0;
        for (int w = 0; w < NT / 32; ++w) { const int c = s_wcnt[threadIdx.x][w]; s_wcnt[threadIdx.x][w] = tot; tot += c; }
        s_base[threadIdx.x] = tot ? atomicAdd(&dd->count[threadIdx.x], tot) : 0;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 3; ++k) rank_in_blk[k] += s_wcnt[k][wid];
    if (!owned) return;
    if (role[0] | role[2]) {  // a boundary agent: its record goes to the adjacent rank(s), as owner (migration) or as ghost
        const float4 pv = cur.pvf[i];
        const float2 vfh = cur.vfh[i], lg = cur.lgoal[i], fc = cur.facing[i];
        const float4 c4 = cur.c4[i];
        const float4 r0 = make_float4(p.x, p.y, pv.z, pv.w), r1 = make_float4(vfh.x, vfh.y, g.x, g.y),
                     r2 = make_float4(lg.x, lg.y, fc.x, fc.y);
#pragma unroll
        for (int k = 0; k < 3; k += 2) {
            if (!role[k]) continue;
            const int q = me - 1 + k;
            // region of the receiver: 0 = from ITS right neighbour (we are that when we send left, k == 0),
            // 2 = from its left neighbour (we send right, k == 2)
            const int cap = dd->lay.cap_halo;
            const int slot = s_base[k] + rank_in_blk[k];
            if (slot >= cap) { atomicCAS(&ctl->error, 0, -5); continue; }
            float4* dst = reinterpret_cast<float4*>(dd->mbox[q] + dd->lay.region[par][k]);
            int4 m = mt;
            m.y = role[k] == 2 ? (mt.y | META_GHOST) : (mt.y & ~META_GHOST);
            dst[slot] = r0;
            dst[(size_t)cap + slot] = r1;
            dst[(size_t)2 * cap + slot] = r2;
            dst[(size_t)3 * cap + slot] = c4;
            dst[(size_t)4 * cap + slot] = make_float4(__int_as_float(m.x), __int_as_float(m.y), __int_as_float(m.z), __int_as_float(m.w));
        }
    }
    dd_owned_tail(P, b, mt, p, g, frame);
}

// Between re-sorts (k_collide's tail when the next frame keeps the slots): nothing is routed — owners, ghosts and slots are
// frozen — the record of every owned agent that is a neighbour's ghost is stored at ITS index of that neighbour's region,
// from where the neighbour copies it over the ghost's slot (k_reorder).  Same wire format as the pack, no counting.
__device__ __forceinline__ void dd_push_body(const Params* __restrict__ P, const Buffers& b) {
    Ctl* ctl = b.ctl;
    DomainDev* dd = b.dd;
    if (!dd_active(P, ctl)) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P->n) return;
    const AgentArrays& cur = set_live(b);
    const int4 mt = cur.meta[i];
    if (mt.y & META_GHOST) return;
    const uint32_t frame = ctl->frame + 1u;
    const int par = frame & 1u, me = P->dd_rank;
    const float4 pv = cur.pvf[i];
    const float2 g = cur.goal[i];
    const int2 si = dd->send_idx[i];
    if ((si.x & si.y) != -1) {
        const float2 vfh = cur.vfh[i], lg = cur.lgoal[i], fc = cur.facing[i];
        const float4 c4 = cur.c4[i];
        const int cap = dd->lay.cap_halo;
#pragma unroll
        for (int k = 0; k < 3; k += 2) {
            const int slot = k == 0 ? si.x : si.y;
            if (slot < 0) continue;
            if (slot >= cap) { atomicCAS(&ctl->error, 0, -5); continue; }
            float4* dst = reinterpret_cast<float4*>(dd->mbox[me - 1 + k] + dd->lay.region[par][k]);
            dst[slot] = pv;
            dst[(size_t)cap + slot] = make_float4(vfh.x, vfh.y, g.x, g.y);
            dst[(size_t)2 * cap + slot] = make_float4(lg.x, lg.y, fc.x, fc.y);
            dst[(size_t)3 * cap + slot] = c4;
            dst[(size_t)4 * cap + slot] = make_float4(__int_as_float(mt.x), __int_as_float(mt.y | META_GHOST), __int_as_float(mt.z), __int_as_float(mt.w));
        }
    }
    dd_owned_tail(P, b, mt, make_float2(pv.x, pv.y), g, frame);
}

__global__ void __launch_bounds__(256) k_dd_pack(const Params* __restrict__ P, Buffers b) {
    pdl_enter();
    dd_pack_body<256>(P, b);
}

// Frame start: neighbour handshake + this frame's record counts and grid, ONE single-block kernel.
__global__ void __launch_bounds__(1024) k_dd_sync_begin(Params* P, Buffers b) {
    pdl_enter();
    Ctl* const gctl = b.ctl;
    DomainDev* dd = b.dd;
    const int me = P->dd_rank, w = P->dd_world, tid = threadIdx.x;
    __shared__ Ctl s_ctl;
    for (int i = tid; i < (int)(sizeof(Ctl) / 4); i += blockDim.x)
        reinterpret_cast<uint32_t*>(&s_ctl)[i] = reinterpret_cast<const uint32_t*>(gctl)[i];
    __syncthreads();
    Ctl* ctl = &s_ctl;
    if (tid == 0) stamp(ctl, TS_BEGIN);
    if (!dd_active(P, ctl)) {  // latched: this and every later frame is a no-op on every rank alike
        if (tid == 0) gctl->halt = 1;
        return;
    }
    const uint32_t frame = ctl->frame + 1u;
    const int par = frame & 1u;
    consume_block_reductions(ctl, b);  // bounds of last frame's local agents (owned and ghosts): max speed, box, lastvel
    // ---- tell the adjacent ranks what they got from us ----
    if (tid < 3 && tid != 1) {
        const int q = me - 1 + tid;
        if (q >= 0 && q < w) {
            NeighbourBlob nb;
            nb.n_sent = ctl->rebin_next ? dd->count[tid] : dd->rcount[tid];   // routed records / refreshed ghosts
            nb.vmax = ctl->vmax; nb.miny = ctl->miny; nb.maxy = ctl->maxy;
            nb.error = ctl->error;
            nb.dmax = ctl->dmax;
            nb.pad[0] = nb.pad[1] = 0;
            // slot in the receiver: 0 = from its right neighbour (we send left), 2 = from its left neighbour (we send right)
            *(reinterpret_cast<NeighbourBlob*>(dd->mbox[q] + dd->lay.nblob[par]) + tid) = nb;
            __threadfence_system();
            dd_st_flag(reinterpret_cast<unsigned long long*>(dd->mbox[q] + dd->lay.nflag[par]) + tid, (unsigned long long)frame);
        }
    }
    // ---- wait for the adjacent ranks' word on what we got ----
    __shared__ int s_ok;
    if (tid == 0) { s_ok = 1; stamp(ctl, TS_NBR_WAIT0); }
    __syncthreads();
    if (tid < 3 && tid != 1) {
        const int q = me - 1 + tid;   // the rank whose records sit in region (2 - tid): left neighbour -> slot 2, right -> slot 0
        if (q >= 0 && q < w) {
            const unsigned long long* fl = reinterpret_cast<const unsigned long long*>(dd->mbox[me] + dd->lay.nflag[par]) + (2 - tid);
            // after a time-out every later wait gives up at once: a broken run ends in milliseconds, not minutes
            if (!dd_wait_flag(fl, (unsigned long long)frame, ctl->error == -2 ? 0ull : dd->spin_timeout_ns)) s_ok = 0;
        }
    }
    __syncthreads();
    __threadfence_system();
    if (tid == 0) {
        stamp(ctl, TS_NBR_WAIT1);
        const NeighbourBlob* nbs = reinterpret_cast<const NeighbourBlob*>(dd->mbox[me] + dd->lay.nblob[par]);
        int err = ctl->error;
        if (!s_ok && !err) err = -2;
        int n_l = 0, n_r = 0;
        uint32_t vmax = ctl->vmax, miny = ctl->miny, maxy = ctl->maxy, dmax = ctl->dmax;
        if (s_ok) {
            if (me > 0) { const NeighbourBlob nb = nbs[2]; n_l = nb.n_sent; vmax = max(vmax, nb.vmax); miny = min(miny, nb.miny); maxy = max(maxy, nb.maxy); dmax = max(dmax, nb.dmax); if (nb.error && !err) err = nb.error; }
            if (me < w - 1) { const NeighbourBlob nb = nbs[0]; n_r = nb.n_sent; vmax = max(vmax, nb.vmax); miny = min(miny, nb.miny); maxy = max(maxy, nb.maxy); dmax = max(dmax, nb.dmax); if (nb.error && !err) err = nb.error; }
        }
        const float vm_now = __uint_as_float(vmax);
        ctl->dmax_last = __uint_as_float(dmax);
        if (!ctl->rebin_next) {
            // ---- the slots stay: the neighbours refreshed our ghosts in place (k_reorder copies the records over them) ----
            if (n_l > dd->lay.cap_halo || n_r > dd->lay.cap_halo || n_l < 0 || n_r < 0) { if (!err) err = -5; n_l = n_r = 0; }
            dd->recv[0] = n_r; dd->recv[2] = n_l;
            ctl->error = err;
            float vmaxf = vm_now;
            if (!(vmaxf <= 1.0e3f)) { vmaxf = 1.0e3f; if (!ctl->error) ctl->error = -3; }
            // the ghosts' true movement is their owner's to report: the skin grows by the largest displacement either side saw
            const float moved = ctl->skin + __uint_as_float(dmax);
            ctl->cutoff = fmaxf(P->R + 10.f * vmaxf * 1.001f + 0.01f, P->near_reach);
            ctl->skin = moved;
            ctl->since += 1;
            ctl->stencil = (int)ceilf((ctl->cutoff + 2.f * moved) / ctl->h);
            ctl->stencil_x = (int)ceilf((ctl->cutoff + 2.f * moved) / ctl->hx);
            // the band was widened by 2 * skin_allow for exactly this (lc_dd_create); dd_global_body asks for a re-sort at
            // half of it, so only a single frame that moves somebody by metres ends up here
            if (!(moved <= dd->skin_allow) && !ctl->error) ctl->error = -5;
            ctl->frame = frame;
            ctl->rebin = 0;
            ctl->dmax = 0u;
            const int last = ctl->live;
            ctl->src_set = last; ctl->in_set = last; ctl->live = last ^ 1;
            ctl->n_cand = ctl->n_cand_next;   // compiled by k_collide
            ctl->n_cand_next = 0;
            ctl->minx = ctl->miny = 0xffffffffu;
            ctl->maxx = ctl->maxy = 0u;
            ctl->vmax = 0u;
            if (!(P->R + 10.f * vm_now * 1.001f + 0.01f + P->near_reach + 2.f * dd->skin_allow <= P->dd_halo) && !ctl->error) ctl->error = -5;
            ctl->overlap2 = 0ull;
            ctl->coll_hits = 0ull;
        } else {
        const int n_prev = P->n, n_stay = dd->count[1];
        int n_in = n_prev + n_l + n_r, n_new = n_stay + n_l + n_r;
        if (n_in > P->cap || n_l > dd->lay.cap_halo || n_r > dd->lay.cap_halo || n_l < 0 || n_r < 0) {
            if (!err) err = -5;
            n_in = n_new = 0;
        }
        dd->recv[0] = n_r; dd->recv[1] = n_prev; dd->recv[2] = n_l;
        dd->count[0] = dd->count[1] = dd->count[2] = 0;  // consumed
        P->n = n_new;
        P->n_in = n_in;
        ctl->error = err;
        // Grid of the new local set: every record here lies in the ghost-extended strip; its y extent and speed are bounded
        // by the senders' local bounds.  (A box that is too small would still be correct — cell_of clamps, and clamping
        // never increases a cell distance — a box that is too large only wastes cells.)
        ctl->vmax = vmax; ctl->miny = miny; ctl->maxy = maxy;
        {
            const bool any = ctl->minx <= ctl->maxx;
            float lo = any ? ord2f(ctl->minx) : 3.0e38f, hi = any ? ord2f(ctl->maxx) : -3.0e38f;
            // what the neighbours sent lies in the ghost band on their side of the cut (ghosts) or just inside ours (migrants)
            if (me > 0 && n_l > 0) lo = fminf(lo, P->dd_xb[me] - P->dd_halo - 0.25f);
            if (me < w - 1 && n_r > 0) hi = fmaxf(hi, P->dd_xb[me + 1] + P->dd_halo + 0.25f);
            // ... and nothing local lies beyond the ghost-extended strip
            if (me > 0) lo = fmaxf(lo, P->dd_xb[me] - P->dd_halo - 0.25f);
            if (me < w - 1) hi = fminf(hi, P->dd_xb[me + 1] + P->dd_halo + 0.25f);
            if (!(hi >= lo)) lo = hi = (me > 0 ? P->dd_xb[me] : (me < w - 1 ? P->dd_xb[me + 1] : 0.f));
            ctl->minx = f2ord(lo); ctl->maxx = f2ord(hi);
        }
        const float vm = __uint_as_float(ctl->vmax);
        ctl->frame = frame;
        grid_from_reductions(P, ctl);
        ctl->rebin = 1;   // ownership and ghost roles are re-decided: a full re-sort of stayers and received records
        ctl->n_resorts += 1ull;
        ctl->skin = 0.f; ctl->since = 0; ctl->force_rebin = 0; ctl->dmax = 0u;
        dd->rcount[0] = dd->rcount[1] = dd->rcount[2] = 0;   // k_reorder rebuilds the refresh lists
        ctl->n_cand_next = 0;
        ctl->src_set = ctl->live;          // roles of the two state copies (crowd.cuh Buffers)
        ctl->in_set = ctl->live ^ 1;
        ctl->live = ctl->in_set ^ 1;       // (unchanged: a re-sort frame writes its result back into the copy it scattered from)
        ctl->minx = ctl->miny = 0xffffffffu;
        ctl->maxx = ctl->maxy = 0u;
        ctl->vmax = 0u;
        ctl->n_cand = 0;
        {   // k_bin's gate radius from last frame's directional extremes
            float rho = CUDART_INF_F;
            for (int k = 0; k < HULL_K; ++k) rho = fminf(rho, ctl->ext_val[k]);
            rho -= 0.1f;
            ctl->gate_rho2 = rho > 0.f ? rho * rho : -1.f;
        }
        // the ghost band was sized for the speed bound (capi.cu); a wider cutoff would miss neighbours
        if (!(P->R + 10.f * vm * 1.001f + 0.01f + P->near_reach + 2.f * dd->skin_allow <= P->dd_halo) && !ctl->error) ctl->error = -5;
        ctl->overlap2 = 0ull;
        ctl->coll_hits = 0ull;
        }
    }
    if (tid < HULL_K) ctl->ext[tid] = 0ull;
    if (tid == 0) stamp(ctl, TS_BEGIN_END);
    __syncthreads();
    // everything S0 works on (ev_count, nnd_acc, lastvel, nonfinite, the Welford state) is handed on untouched to
    // dd_global_body, the only other writer of the control block, which runs strictly later in this frame
    for (int i = tid; i < (int)(sizeof(Ctl) / 4); i += blockDim.x)
        reinterpret_cast<uint32_t*>(gctl)[i] = reinterpret_cast<const uint32_t*>(&s_ctl)[i];
}

// The frame's one all-rank exchange, in the tail of k_refine on the side stream.  Publishes this rank's comm-partner
// candidates (with the payload comm::x reads from the partner) and scoring inputs, waits for everybody's, appends their
// candidates behind the local arrays ("annex" slots >= cap of srt.pv / srt.c4), merges the scoring inputs and runs S0.
__device__ void dd_global_body(const Params* __restrict__ P, const Buffers& b) {
    Ctl* const gctl = b.ctl;
    DomainDev* dd = b.dd;
    const int me = P->dd_rank, w = P->dd_world, tid = threadIdx.x;
    const uint32_t frame = gctl->frame;  // k_dd_sync_begin already advanced it
    const int par = frame & 1u;
    const bool comm = !(P->flags & 1u);
    float4* list = gctl->cand_sel ? b.cand2 : b.cand;
    const int nloc = comm ? gctl->n_cand2 : 0;
    const int nsend = min(nloc, DD_CAND_CAP);
    const int nev = min(gctl->ev_count, DD_EV_CAP);
    if (tid == 0 && (nloc > DD_CAND_CAP || gctl->ev_count > DD_EV_CAP)) atomicCAS(&gctl->error, 0, -5);
    // ---- publish ----
    for (int k = tid; k < nsend; k += blockDim.x) {
        const float4 cd = list[k];
        const int slot = __float_as_int(cd.w);
        const float4 pj = set_in(b).pvf[slot], cj = set_in(b).c4[slot];
        const float4 A = make_float4(cd.x, cd.y, cd.z, pj.z), B = make_float4(cj.x, cj.y, cj.z, 0.f);
        for (int q = 0; q < w; ++q) {
            if (q == me) continue;
            float4* dst = reinterpret_cast<float4*>(dd->mbox[q] + dd->lay.cand[par]) + (size_t)me * 2 * DD_CAND_CAP;
            dst[2 * k] = A;
            dst[2 * k + 1] = B;
        }
    }
    {   // done-events of the frame that just ended go to the scoring rank (rank 0), which folds them in id order
        int2* evdst = reinterpret_cast<int2*>(dd->mbox[0] + dd->lay.events[par]) + (size_t)me * DD_EV_CAP;
        if (me != 0) for (int k = tid; k < nev; k += blockDim.x) evdst[k] = b.events[k];
    }
    __shared__ DomainBlob s_blob;
    if (tid == 0) {
        s_blob.nnd = gctl->nnd_acc;
        s_blob.nonfinite = gctl->nonfinite;
        s_blob.lastvel = gctl->lastvel;
        s_blob.replan[0] = dd->replan[0]; s_blob.replan[1] = dd->replan[1];
        s_blob.n_events = nev;
        s_blob.n_cand = nsend;
        s_blob.error = gctl->error;
        // (the frame in flight will add its own displacement, unknown yet: allow for four times the last one)
        // and nothing is known about it in the very first frame
        s_blob.want_rebin = frame < 2u || !(gctl->skin + 4.f * gctl->dmax_last <= 0.5f * dd->skin_allow);
    }
    __syncthreads();
    constexpr int WORDS = sizeof(DomainBlob) / 4;
    for (int k = tid; k < w * WORDS; k += blockDim.x) {
        const int q = k / WORDS, o = k % WORDS;
        reinterpret_cast<int*>(reinterpret_cast<DomainBlob*>(dd->mbox[q] + dd->lay.blob[par]) + me)[o] = reinterpret_cast<const int*>(&s_blob)[o];
    }
    __threadfence_system();
    __syncthreads();
    if (tid < w) dd_st_flag(reinterpret_cast<unsigned long long*>(dd->mbox[tid] + dd->lay.cflag[par]) + me, (unsigned long long)frame);
    // ---- wait for every rank ----
    __shared__ int s_off[DD_MAX_WORLD + 1], s_evoff[DD_MAX_WORLD + 1];
    __shared__ int s_ok;
    if (tid == 0) { s_ok = 1; stamp(gctl, TS_ALL_WAIT0); }
    __syncthreads();
    if (tid < w) {
        const unsigned long long* fl = reinterpret_cast<const unsigned long long*>(dd->mbox[me] + dd->lay.cflag[par]) + tid;
        if (!dd_wait_flag(fl, (unsigned long long)frame, gctl->error == -2 ? 0ull : dd->spin_timeout_ns)) s_ok = 0;
    }
    __syncthreads();
    __threadfence_system();
    if (tid == 0) stamp(gctl, TS_ALL_WAIT1);
    // ---- merge ----
    __shared__ Ctl s_ctl;   // S0 works on a copy; only the scoring fields go back (k_pairs is reading the grid fields meanwhile)
    for (int i = tid; i < (int)(sizeof(Ctl) / 4); i += blockDim.x)
        reinterpret_cast<uint32_t*>(&s_ctl)[i] = reinterpret_cast<const uint32_t*>(gctl)[i];
    __syncthreads();
    Ctl* ctl = &s_ctl;
    const DomainBlob* blobs = reinterpret_cast<const DomainBlob*>(dd->mbox[me] + dd->lay.blob[par]);
    if (tid == 0) {
        int nnd = 0, nonfinite = 0, err = ctl->error, want = 0;
        if (!s_ok && !err) err = -2;
        unsigned long long lastvel = 0ull;
        int r1 = 0x7fffffff, r2 = 0x7fffffff;
        int off = 0, evoff = 0;
        for (int q = 0; q < w; ++q) {
            const DomainBlob bl = s_ok ? blobs[q] : s_blob;
            nnd += bl.nnd;
            nonfinite += bl.nonfinite;
            lastvel = bl.lastvel > lastvel ? bl.lastvel : lastvel;
            for (int k = 0; k < 2; ++k) {  // two smallest distinct ids of the union
                const int v = bl.replan[k];
                if (v < r1) { r2 = r1; r1 = v; } else if (v > r1 && v < r2) r2 = v;
            }
            if (bl.error && !err) err = bl.error;
            want |= bl.want_rebin;
            s_off[q] = off;
            off += (q == me || !s_ok) ? 0 : min(bl.n_cand, DD_CAND_CAP);
            s_evoff[q] = evoff;
            evoff += (me == 0 && s_ok) ? min(bl.n_events, DD_EV_CAP) : 0;
        }
        s_off[w] = off;
        s_evoff[w] = evoff;
        ctl->nnd_acc = nnd;
        ctl->nonfinite = nonfinite;
        ctl->lastvel = lastvel;
        ctl->replan1 = r1; ctl->replan2 = r2;
        ctl->ev_count = evoff;
        ctl->error = err;
        dd->replan[0] = dd->replan[1] = 0x7fffffff;
        if (comm) gctl->n_cand2 = nloc + off;   // k_force's candidate list: local survivors + everybody else's
        // Does the NEXT frame re-sort?  By the frame count, or because somebody's skin is half used up: every rank sees the
        // same blobs and the same counter, so every rank decides the same — this frame's k_collide routes records if so and
        // only refreshes the ghosts if not.
        gctl->rebin_next = (P->rebin_every <= 1 || ctl->since + 1 >= P->rebin_every || want || !s_ok) ? 1 : 0;
    }
    __syncthreads();
    if (comm) {
        const int cap = P->cap;
        for (int q = 0; q < w; ++q) {
            const int cnt = s_off[q + 1] - s_off[q];
            const float4* src = reinterpret_cast<const float4*>(dd->mbox[me] + dd->lay.cand[par]) + (size_t)q * 2 * DD_CAND_CAP;
            for (int k = tid; k < cnt; k += blockDim.x) {
                const float4 A = src[2 * k], B = src[2 * k + 1];
                const int annex = cap + s_off[q] + k;
                set_in(b).pvf[annex] = make_float4(A.x, A.y, A.w, 0.f);  // comm::x reads pos.x and vel.x of the partner only (quirk Q2)
                set_in(b).c4[annex] = B;
                list[nloc + s_off[q] + k] = make_float4(A.x, A.y, A.z, __int_as_float(annex));
            }
        }
    }
    if (me == 0) {  // gather every rank's done-events (own ones are already in b.events[0 .. own count)); S0 sorts them by id
        const int2* evsrc = reinterpret_cast<const int2*>(dd->mbox[0] + dd->lay.events[par]);
        // other ranks' events go behind our own: shift by our own count
        const int own = s_evoff[1] - s_evoff[0];
        for (int q = 1; q < w; ++q) {
            const int cnt = s_evoff[q + 1] - s_evoff[q];
            for (int k = tid; k < cnt; k += blockDim.x) b.events[s_evoff[q] + k] = evsrc[(size_t)q * DD_EV_CAP + k];
        }
        (void)own;
    }
    __threadfence();
    __syncthreads();
    // ---- S0: demo::run (demo.cpp:665-702) with the merged inputs, exactly as k_frame_begin does it on one GPU ----
    __shared__ int2 s_ev[EV_SORT_CAP];
    __shared__ float s_ft[1024];
    __shared__ int s_flag[1024];
    if (tid == 0) fold_avg_velocity(ctl);
    __syncthreads();
    const int was_finished = ctl->finished;
    const int cnt = ctl->ev_count, nnd = ctl->nnd_acc;
    __syncthreads();
    if (!was_finished) demo_run_score(P, b, ctl, frame, cnt, nnd, true, s_ev, s_ft, s_flag);
    __syncthreads();
    if (tid == 0) {  // only the scoring fields go back
        gctl->total_num_done = ctl->total_num_done;
        gctl->run_avg = ctl->run_avg; gctl->M2 = ctl->M2; gctl->run_std = ctl->run_std;
        gctl->avg_velocity = ctl->avg_velocity;
        gctl->lastvel = ctl->lastvel; gctl->nonfinite = ctl->nonfinite;
        gctl->finished = ctl->finished;
        gctl->n_not_done = nnd;
        gctl->replan1 = ctl->replan1; gctl->replan2 = ctl->replan2;
        gctl->ev_count = 0;   // k_force of this frame collects the next frame's events and not-done count
        gctl->nnd_acc = 0;
        gctl->error = ctl->error;
    }
}

