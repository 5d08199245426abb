// Spatial domain decomposition of ONE crowd over several GPUs (BASELINE.json configs[4]): x strips, one rank per
// GPU, boundary agents exchanged every frame by kernels that store straight into the neighbours' memory over
// NVLink (peer mappings of one mailbox allocation per rank) — no host round trip, no NCCL call on the data path.
// Included at the end of crowd.cu (same translation unit: it reuses the block reductions).
//
// Frame f on every rank (all inside the CUDA graph):
//   (pack)        owned agents -> records in the mailbox of their new owner (stay / migrate to an adjacent strip) and,
//                 as ghosts, of every adjacent rank whose strip they are within `dd_halo` of; at the time-out frame
//                 the finish times of demo.cpp:691-700 are scattered into the scoring rank's id-indexed table.
//                 Done by the tail of the PREVIOUS frame's k_collide (dd_pack_body; one pass over the state and one
//                 launch less per frame); kernel k_dd_pack only runs once, before the first frame
//   (publish)     folds the per-block reductions of the previous frame, writes this rank's DomainBlob (record counts,
//                 max speed, not-done count, avg_velocity inputs, replan gate, done-events) into every peer, then
//                 releases flag[rank] = f with a system-scope fence
//   (wait)        spins until every peer's flag reads f, merges the blobs into Ctl (GLOBAL max speed, not-done count,
//                 replan gate, events on rank 0) and sets Params::n to the number of local records
//                 (publish + wait + k_frame_begin's work are ONE single-block kernel, k_dd_sync_begin)
//   k_bin, k_scan_hull, k_reorder read the records in place (owners and ghosts mixed; meta bit2 tells them apart) and sort
//                 them into cell order like a single-GPU crowd; there is no unpack copy
//   (candidates)  dd_cand_body, the tail of k_refine: publishes this rank's surviving comm-partner candidates (with the partner payload x() needs:
//                 pos, vel.x, c) to every peer, waits for theirs and appends them (comm.cpp:201-220 is a GLOBAL arg-max)
//   k_force, k_collide: ghosts are advanced too — the collision pass needs their integrated positions — but never
//                 counted; their next state always comes from the owner.
// Why the result is bit-identical to the single-GPU run: an owned agent sees every neighbour within the interaction
// cutoff (ghost band = cutoff at the speed bound + collision reach), forces and hits are summed in id order, the
// comm partner is the global arg-max over the union of the per-rank hull survivors, and the replan gate, max speed,
// done count and scoring inputs are merged across ranks before they are used.  tests/test_gpu_domain.py checks it.
//
// Mailboxes are double-buffered by frame parity: a peer can be at most one exchange ahead (it needs our flag of
// frame f before it can start frame f + 1's pack), so parity f + 2 is only written after we consumed parity f.
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
    // Poll back to back; look at the clock (and yield for a moment) only every 32 polls.  A 200 ns sleep per poll added
    // up to half a microsecond of wake-up delay to each of the two waits of a frame (8 GPUs: 9.81e9 -> 9.88e9 with 32 ns).
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

// which of (rank-1, self, rank+1) gets a record of an owned agent at x, and in which role (1 = owner, 2 = ghost)
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
    const uint32_t frame = ctl->frame + 1u;
    const int par = frame & 1u;
    // Slot allocation keeps the agents' order (ballot prefix inside the warp, warp totals, one global atomicAdd per block
    // and destination): the records stay in last frame's cell order, which k_bin's run-aggregated counting relies on.
    __shared__ int s_wcnt[3][NT / 32], s_base[3];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int role[3] = {0, 0, 0}, rank_in_blk[3] = {0, 0, 0};
    int4 mt = make_int4(0, 0, 0, 0);
    float2 p = make_float2(0.f, 0.f), g = make_float2(0.f, 0.f);
    bool owned = false;
    if (i < n) {
        mt = b.cur.meta[i];
        owned = !(mt.y & META_GHOST);
        if (owned) {
            p = b.posf[i];
            g = b.cur.goal[i];
            dd_route(P, p.x, role);
        }
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
    const float4 pv = b.cur.pv[i];
    const float2 vfh = b.cur.vfh[i], lg = b.cur.lgoal[i], fc = b.cur.facing[i];
    const float4 c4 = b.cur.c4[i];
    const float4 r0 = make_float4(p.x, p.y, pv.z, pv.w), r1 = make_float4(vfh.x, vfh.y, g.x, g.y),
                 r2 = make_float4(lg.x, lg.y, fc.x, fc.y);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        if (!role[k]) continue;
        const int q = me - 1 + k;
        // region of the receiver: 0 = from ITS right neighbour (we are that when we send left, k == 0), 1 = itself,
        // 2 = from its left neighbour (we send right, k == 2)
        const int reg = k;
        const int cap = reg == 1 ? dd->lay.cap_self : dd->lay.cap_halo;
        const int slot = s_base[k] + rank_in_blk[k];
        if (slot >= cap) { atomicCAS(&ctl->error, 0, -5); continue; }
        float4* dst = reinterpret_cast<float4*>(dd->mbox[q] + dd->lay.region[par][reg]);
        int4 m = mt;
        m.y = role[k] == 2 ? (mt.y | META_GHOST) : (mt.y & ~META_GHOST);
        dst[slot] = r0;
        dst[(size_t)cap + slot] = r1;
        dst[(size_t)2 * cap + slot] = r2;
        dst[(size_t)3 * cap + slot] = c4;
        dst[(size_t)4 * cap + slot] = make_float4(__int_as_float(m.x), __int_as_float(m.y), __int_as_float(m.z), __int_as_float(m.w));
    }
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

__global__ void __launch_bounds__(256) k_dd_pack(const Params* __restrict__ P, Buffers b) {
    pdl_enter();
    dd_pack_body<256>(P, b);
}

// publish + wait + merge + S0 (k_frame_begin's work) in ONE single-block kernel: three launch gaps less per frame
__global__ void __launch_bounds__(1024) k_dd_sync_begin(Params* P, Buffers b) {
    pdl_enter();
    Ctl* ctl = b.ctl;
    DomainDev* dd = b.dd;
    if (!dd_active(P, ctl)) { frame_begin_body(P, b); return; }  // latches halt, like every later frame of a finished run
    const int me = P->dd_rank, w = P->dd_world, tid = threadIdx.x;
    const uint32_t frame = ctl->frame + 1u;
    const int par = frame & 1u;
    consume_block_reductions(ctl, b);  // max speed / lastvel of the frame that just ended (owned agents)
    __shared__ DomainBlob s_blob;
    const int nev = min(ctl->ev_count, DD_EV_CAP);
    if (tid == 0) {
        if (ctl->ev_count > DD_EV_CAP && !ctl->error) ctl->error = -5;
        s_blob.n_sent[0] = dd->count[0]; s_blob.n_sent[1] = dd->count[1]; s_blob.n_sent[2] = dd->count[2];
        s_blob.vmax = ctl->vmax;
        s_blob.bbox[0] = ctl->minx; s_blob.bbox[1] = ctl->miny; s_blob.bbox[2] = ctl->maxx; s_blob.bbox[3] = ctl->maxy;
        s_blob.nnd = ctl->nnd_acc;
        s_blob.nonfinite = ctl->nonfinite;
        s_blob.lastvel = ctl->lastvel;
        s_blob.replan[0] = dd->replan[0]; s_blob.replan[1] = dd->replan[1];
        s_blob.n_events = nev;
        s_blob.error = ctl->error;
        s_blob.pad[0] = s_blob.pad[1] = 0;
    }
    // done-events of the frame that just ended go to the scoring rank (rank 0), which folds them in id order
    int2* evdst = reinterpret_cast<int2*>(dd->mbox[0] + dd->lay.events[par]) + (size_t)me * DD_EV_CAP;
    for (int k = tid; k < nev; k += blockDim.x) evdst[k] = b.events[k];
    __syncthreads();
    constexpr int WORDS = sizeof(DomainBlob) / 4;
    for (int k = tid; k < w * WORDS; k += blockDim.x) {
        const int q = k / WORDS, o = k % WORDS;
        reinterpret_cast<int*>(reinterpret_cast<DomainBlob*>(dd->mbox[q] + dd->lay.blob[par]) + me)[o] = reinterpret_cast<const int*>(&s_blob)[o];
    }
    __threadfence_system();
    __syncthreads();
    if (tid < w) dd_st_flag(reinterpret_cast<unsigned long long*>(dd->mbox[tid] + dd->lay.flag[par]) + me, (unsigned long long)frame);

    // ---- wait for every peer's blob of this frame and merge ----
    __shared__ int s_ok;
    __shared__ int s_evoff[DD_MAX_WORLD + 1];
    if (tid == 0) s_ok = 1;
    __syncthreads();
    if (tid < w) {
        const unsigned long long* fl = reinterpret_cast<const unsigned long long*>(dd->mbox[me] + dd->lay.flag[par]) + tid;
        // after a time-out every later wait gives up at once: a broken run ends in milliseconds, not minutes
        if (!dd_wait_flag(fl, (unsigned long long)frame, ctl->error == -2 ? 0ull : dd->spin_timeout_ns)) s_ok = 0;
    }
    __syncthreads();
    __threadfence_system();
    const DomainBlob* blobs = reinterpret_cast<const DomainBlob*>(dd->mbox[me] + dd->lay.blob[par]);
    if (tid == 0) {
        if (!s_ok && !ctl->error) ctl->error = -2;
        uint32_t vmax = 0u, bb0 = 0xffffffffu, bb1 = 0xffffffffu, bb2 = 0u, bb3 = 0u;
        int nnd = 0, nonfinite = 0, err = ctl->error;
        unsigned long long lastvel = 0ull;
        int r1 = 0x7fffffff, r2 = 0x7fffffff;
        int evoff = 0;
        for (int q = 0; q < w; ++q) {
            const DomainBlob bl = blobs[q];
            vmax = max(vmax, bl.vmax);
            bb0 = min(bb0, bl.bbox[0]); bb1 = min(bb1, bl.bbox[1]); bb2 = max(bb2, bl.bbox[2]); bb3 = max(bb3, bl.bbox[3]);
            nnd += bl.nnd;
            nonfinite += bl.nonfinite;
            lastvel = bl.lastvel > lastvel ? bl.lastvel : lastvel;
            for (int k = 0; k < 2; ++k) {  // two smallest distinct ids of the union
                const int v = bl.replan[k];
                if (v < r1) { r2 = r1; r1 = v; } else if (v > r1 && v < r2) r2 = v;
            }
            if (bl.error && !err) err = bl.error;
            s_evoff[q] = evoff;
            evoff += (me == 0) ? bl.n_events : 0;
        }
        s_evoff[w] = evoff;
        // records addressed to this rank: own stayers, then what the left / right neighbour sent (owners and ghosts mixed)
        const int n_self = blobs[me].n_sent[1];
        const int n_l = me > 0 ? blobs[me - 1].n_sent[2] : 0;
        const int n_r = me < w - 1 ? blobs[me + 1].n_sent[0] : 0;
        int total = n_self + n_l + n_r;
        if (total > P->cap || n_self > dd->lay.cap_self || n_l > dd->lay.cap_halo || n_r > dd->lay.cap_halo) {
            if (!err) err = -5;
            total = 0;
        }
        dd->recv[0] = n_r; dd->recv[1] = n_self; dd->recv[2] = n_l;
        dd->count[0] = dd->count[1] = dd->count[2] = 0;  // the publish step above has consumed them
        P->n = total;
        ctl->vmax = vmax;
        ctl->nnd_acc = nnd;
        ctl->nonfinite = nonfinite;
        ctl->lastvel = lastvel;
        ctl->replan1 = r1; ctl->replan2 = r2;
        ctl->ev_count = evoff;
        ctl->error = err;
        // Grid extent of the new local set: every record routed here lies in the ghost-extended strip, and inside the
        // crowd's global box in y.  (A box that is too small would still be correct — cell_of clamps, and clamping never
        // increases a cell distance — a box that is too large only wastes cells.)
        {
            float gx0 = ord2f(bb0), gx1 = ord2f(bb2);
            const float sx0 = P->dd_xb[me] - P->dd_halo - 0.25f, sx1 = P->dd_xb[me + 1] + P->dd_halo + 0.25f;
            if (bb0 <= bb2) { gx0 = fmaxf(gx0, sx0); gx1 = fminf(gx1, sx1); if (!(gx1 >= gx0)) gx1 = gx0; }
            ctl->minx = bb0 <= bb2 ? f2ord(gx0) : bb0; ctl->maxx = bb0 <= bb2 ? f2ord(gx1) : bb2;
            ctl->miny = bb1; ctl->maxy = bb3;
        }
        dd->replan[0] = dd->replan[1] = 0x7fffffff;
    }
    __syncthreads();
    if (me == 0) {  // gather every rank's done-events; k_frame_begin sorts them by id before the Welford update
        const int2* evsrc = reinterpret_cast<const int2*>(dd->mbox[0] + dd->lay.events[par]);
        for (int q = 0; q < w; ++q) {
            const int cnt = s_evoff[q + 1] - s_evoff[q];
            for (int k = tid; k < cnt; k += blockDim.x) b.events[s_evoff[q] + k] = evsrc[(size_t)q * DD_EV_CAP + k];
        }
    }
    __threadfence();
    __syncthreads();
    frame_begin_body(P, b);  // S0: scoring + latch + this frame's grid (demo.cpp:665-702), as on a single GPU
}

// Comm-partner candidates are a global matter (comm.cpp:201-220 scans ALL agents).  Each rank's list after k_refine
// holds the hull survivors of its own agents; publish them with the payload comm::x reads from the partner and
// append everybody else's behind the local arrays ("annex" slots >= cap of srt.pv / srt.c4).
__device__ void dd_cand_body(const Params* __restrict__ P, const Buffers& b) {
    Ctl* ctl = b.ctl;
    DomainDev* dd = b.dd;
    const int me = P->dd_rank, w = P->dd_world, tid = threadIdx.x;
    const uint32_t frame = ctl->frame;  // k_frame_begin already advanced it
    const int par = frame & 1u;
    float4* list = ctl->cand_sel ? b.cand2 : b.cand;
    const int nloc = ctl->n_cand2;
    const int nsend = min(nloc, DD_CAND_CAP);
    if (tid == 0 && nloc > DD_CAND_CAP) atomicCAS(&ctl->error, 0, -5);
    for (int k = tid; k < nsend; k += blockDim.x) {
        const float4 cd = list[k];
        const int slot = __float_as_int(cd.w);
        const float4 pj = b.srt.pv[slot], cj = b.srt.c4[slot];
        const float4 A = make_float4(cd.x, cd.y, cd.z, pj.z), B = make_float4(cj.x, cj.y, cj.z, 0.f);
        for (int q = 0; q < w; ++q) {
            if (q == me) continue;
            float4* dst = reinterpret_cast<float4*>(dd->mbox[q] + dd->lay.cand[par]) + (size_t)me * 2 * DD_CAND_CAP;
            dst[2 * k] = A;
            dst[2 * k + 1] = B;
        }
    }
    __threadfence_system();
    __syncthreads();
    if (tid < w && tid != me) {
        reinterpret_cast<int*>(dd->mbox[tid] + dd->lay.ncand[par])[me] = nsend;
        __threadfence_system();
        dd_st_flag(reinterpret_cast<unsigned long long*>(dd->mbox[tid] + dd->lay.cflag[par]) + me, (unsigned long long)frame);
    }
    __shared__ int s_off[DD_MAX_WORLD + 1];
    __shared__ int s_ok;
    if (tid == 0) s_ok = 1;
    __syncthreads();
    if (tid < w && tid != me) {
        const unsigned long long* fl = reinterpret_cast<const unsigned long long*>(dd->mbox[me] + dd->lay.cflag[par]) + tid;
        if (!dd_wait_flag(fl, (unsigned long long)frame, ctl->error == -2 ? 0ull : dd->spin_timeout_ns)) s_ok = 0;
    }
    __syncthreads();
    __threadfence_system();
    if (tid == 0) {
        if (!s_ok) atomicCAS(&ctl->error, 0, -2);
        const int* nc = reinterpret_cast<const int*>(dd->mbox[me] + dd->lay.ncand[par]);
        int off = 0;
        for (int q = 0; q < w; ++q) {
            s_off[q] = off;
            off += (q == me || !s_ok) ? 0 : min(nc[q], DD_CAND_CAP);
        }
        s_off[w] = off;
        ctl->n_cand2 = nloc + off;
    }
    __syncthreads();
    const int cap = P->cap;
    for (int q = 0; q < w; ++q) {
        const int cnt = s_off[q + 1] - s_off[q];
        const float4* src = reinterpret_cast<const float4*>(dd->mbox[me] + dd->lay.cand[par]) + (size_t)q * 2 * DD_CAND_CAP;
        for (int k = tid; k < cnt; k += blockDim.x) {
            const float4 A = src[2 * k], B = src[2 * k + 1];
            const int annex = cap + s_off[q] + k;
            b.srt.pv[annex] = make_float4(A.x, A.y, A.w, 0.f);  // comm::x reads pos.x and vel.x of the partner only (quirk Q2)
            b.srt.c4[annex] = B;
            list[nloc + s_off[q] + k] = make_float4(A.x, A.y, A.z, __int_as_float(annex));
        }
    }
}

