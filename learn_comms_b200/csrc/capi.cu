// C ABI of liblearncomms.so (include/learn_comms.h): handle management, host<->device plumbing, CUDA-graph
// replay of the frame.  No compute happens on the host: every entry point that advances or reads the
// simulation fails with LC_ERR_CUDA when no CUDA device is usable.
#include "../../include/learn_comms.h"
#include "crowd.cuh"
#include "lc_math.cuh"
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <string>
#include <vector>

namespace {

thread_local std::string g_err;

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CK(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess)                                                                           \
            return fail(LC_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

uint32_t lcg_pow(uint32_t a, uint64_t e) {
    uint32_t r = 1;
    while (e) {
        if (e & 1) r = lcm::lcg_mul(r, a);
        a = lcm::lcg_mul(a, a);
        e >>= 1;
    }
    return r;
}

}  // namespace

// shared with batch.cu / domain.cu: one error text per host thread, whichever entry point failed last
int lc_set_error(int code, const char* msg) {
    g_err = msg;
    return code;
}

namespace {

template <class T> cudaError_t dalloc(T** p, size_t count) { return cudaMalloc((void**)p, sizeof(T) * (count ? count : 1)); }

}  // namespace

struct lc_sim {
    lc::Params hp{};
    lc::Params* dP = nullptr;
    lc::Buffers b{};
    cudaStream_t st = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    lc::SideStream side{nullptr, nullptr, nullptr};  // k_refine next to k_pairs (crowd.cuh)
    cudaGraphExec_t gexec[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};  // 1, 2, 4, 8, 16 frames
    long long launches = 0;
    int device = 0;
    int n = 0;           // launch width: slot capacity of the local arrays (== n_ids for a single-GPU crowd)
    int n_ids = 0;       // global id space: size of every id-indexed table and of the host-facing arrays
    uint32_t rng0 = 1;   // stream state before frame 1's first draw
    // spatial domain decomposition (lc_dd_*): mailbox other ranks store into, device-side exchange state
    int dd_rank = 0, dd_world = 1;
    char* mailbox = nullptr;
    lc::DomainLayout lay{};
    lc::DomainDev* ddev = nullptr;
    int* owned_dev = nullptr;
    std::vector<void*> ipc_opened;
    bool dd_packed = false;  // the records of the first frame are in the mailboxes (k_dd_pack ran); later frames: k_collide
    // staging (device) for state exchange, pinned host mirror of Ctl
    float* stage = nullptr;
    size_t stage_floats = 0;
    // lc_exchange_state: second staging area and copy stream, so that the upload of the next forced state runs beside the
    // download of the current one (and beside whatever frame is still executing)
    float* stage_in = nullptr;
    cudaStream_t st_in = nullptr;
    cudaEvent_t ev_in = nullptr;
    lc::Ctl* hctl = nullptr;
    // walled scenes: host planner (stateful roadmap, planner.cpp), host mirrors indexed by engine id
    lc_planner* planner = nullptr;
    std::vector<float> h_goal;   // [n_ids][2]
    std::vector<int> h_nplan;    // [n_ids]
    long long replans = 0;
    std::vector<int> h_order;    // lc_desc::table_order (empty = ascending)
    int hp_ext(size_t id) const { return h_order.empty() ? (int)id : h_order[id]; }
    int* ext_of = nullptr;     // device copy of lc_desc::table_order (nullptr = ascending)
    int* xlate = nullptr;      // staging for diagnostics translated to the caller's agent indices
    std::vector<void*> allocs;
};

// lc_step(n) replays n frames as the binary decomposition of n over these graphs: 20 frames = 16 + 4, two launches
// (with only 1- and 16-frame graphs it was 16 + 1 + 1 + 1 + 1: five, whose host-side skew between the ranks of a
// decomposed crowd ended up in the peer-flag waits)
static_assert(LC_TRACE_STAMPS == lc::TS_COUNT, "trace layout shared with the C ABI");
static const int GRAPH_LEVELS = 5;
static const int GRAPH_BATCH = 1 << (GRAPH_LEVELS - 1);

static int alloc_arrays(lc_sim* s, lc::AgentArrays& a, int n) {
    CK(dalloc(&a.pvf, n)); s->allocs.push_back(a.pvf);
    CK(dalloc(&a.pq, (size_t)n + 4)); s->allocs.push_back(a.pq);  // 2 float4 per slot PAIR, padded to a whole pair + one pair of over-read (scan_pairs_flat)
    CK(cudaMemset(a.pq, 0, sizeof(float4) * ((size_t)n + 4)));
    CK(dalloc(&a.pv, n)); s->allocs.push_back(a.pv);
    CK(dalloc(&a.vfh, n)); s->allocs.push_back(a.vfh);
    CK(dalloc(&a.goal, n)); s->allocs.push_back(a.goal);
    CK(dalloc(&a.lgoal, n)); s->allocs.push_back(a.lgoal);
    CK(dalloc(&a.facing, n)); s->allocs.push_back(a.facing);
    CK(dalloc(&a.c4, n)); s->allocs.push_back(a.c4);
    CK(dalloc(&a.meta, n)); s->allocs.push_back(a.meta);
    return 0;
}

static int build_graphs(lc_sim* s) {
    for (int g = 0; g < GRAPH_LEVELS; ++g) {
        int frames = 1 << g;
        cudaGraph_t graph;
        CK(cudaStreamBeginCapture(s->st, cudaStreamCaptureModeThreadLocal));
        for (int f = 0; f < frames; ++f) lc::launch_frame(s->dP, s->n, s->b, s->st, nullptr, nullptr, &s->side, s->hp.n_walls > 0);
        CK(cudaStreamEndCapture(s->st, &graph));
        CK(cudaGraphInstantiate(&s->gexec[g], graph, 0));
        CK(cudaGraphDestroy(graph));
        // have the executable graph resident before its first launch: without this the first cudaGraphLaunch of each size
        // pays the upload inside the caller's timed region (≈0.3 ms for the 16-frame graph)
        CK(cudaGraphUpload(s->gexec[g], s->st));
    }
    return 0;
}

static int upload_params(lc_sim* s) {
    CK(cudaMemcpyAsync(s->dP, &s->hp, sizeof(lc::Params), cudaMemcpyHostToDevice, s->st));
    return 0;
}
// only the parameter matrices: Params::n is rewritten on the device every frame in domain mode
static int upload_weights(lc_sim* s) {
    const size_t off = offsetof(lc::Params, m_cv);
    CK(cudaMemcpyAsync((char*)s->dP + off, (const char*)&s->hp + off, sizeof(lc::Params) - off, cudaMemcpyHostToDevice, s->st));
    return 0;
}

static int read_ctl(lc_sim* s) {
    CK(cudaMemcpyAsync(s->hctl, s->b.ctl, sizeof(lc::Ctl), cudaMemcpyDeviceToHost, s->st));
    CK(cudaStreamSynchronize(s->st));
    return 0;
}

extern "C" {

const char* lc_last_error(void) { return g_err.c_str(); }
const char* lc_version(void) { return "learn_comms_b200 0.1 (sm_100a)"; }

void lc_destroy(lc_sim* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    if (s->st) cudaStreamSynchronize(s->st);
    for (auto& g : s->gexec) if (g) cudaGraphExecDestroy(g);
    if (s->side.st) cudaStreamDestroy(s->side.st);
    if (s->side.fork) cudaEventDestroy(s->side.fork);
    if (s->side.join) cudaEventDestroy(s->side.join);
    if (s->planner) lc_planner_destroy(s->planner);
    for (void* p : s->ipc_opened) cudaIpcCloseMemHandle(p);
    for (void* p : s->allocs) cudaFree(p);
    if (s->mailbox) cudaFree(s->mailbox);
    if (s->dP) cudaFree(s->dP);
    if (s->stage) cudaFree(s->stage);
    if (s->stage_in) cudaFree(s->stage_in);
    if (s->st_in) cudaStreamDestroy(s->st_in);
    if (s->ev_in) cudaEventDestroy(s->ev_in);
    if (s->hctl) cudaFreeHost(s->hctl);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->st) cudaStreamDestroy(s->st);
    delete s;
}

}  // extern "C"

// Walled scene set-up (ai::init, ai.cpp:24-129): static BVH, configuration space + roadmap (host planner), one plan per
// agent in table order — engine ids ARE table positions — and min_time = plan length minus the untimed last metre.
template <class Ext>
static int setup_walls(lc_sim* s, const lc_desc* d, Ext ext, std::vector<float>& min_time) {
    const int n = d->n, nw = d->n_walls;
    std::vector<int32_t> worder((size_t)nw);
    std::vector<char> seen((size_t)nw, 0);
    for (int k = 0; k < nw; ++k) {
        worder[k] = d->wall_order ? d->wall_order[k] : k;
        if (worder[k] < 0 || worder[k] >= nw || seen[worder[k]]) return fail(LC_ERR_ARG, "lc_create: wall_order is not a permutation of 0..n_walls-1");
        seen[worder[k]] = 1;
    }
    std::vector<lc_bvh_node> bvh((size_t)2 * nw);
    int32_t n_bvh = 0;
    int rc = lc_static_bvh(nw, d->walls, worder.data(), bvh.data(), (int32_t)bvh.size(), &n_bvh);
    if (rc) return fail(rc, "lc_create: static BVH construction failed");
    rc = lc_planner_create(nw, d->walls, d->radius, &s->planner);
    if (rc) return rc;
    std::vector<float> plan((size_t)n * lc::WP_CAP * 2, 0.f);
    std::vector<int> noplan((size_t)n, 0);
    s->h_nplan.assign((size_t)n, 0);
    s->h_goal.resize((size_t)2 * n);
    min_time.assign((size_t)n, 0.f);
    int missing = 0;
    for (int id = 0; id < n; ++id) {
        const int e = ext(id);
        const float start[2] = {d->pos[2 * e], d->pos[2 * e + 1]}, goal[2] = {d->goal[2 * e], d->goal[2 * e + 1]};
        s->h_goal[2 * id] = goal[0]; s->h_goal[2 * id + 1] = goal[1];
        int32_t k = 0;
        rc = lc_planner_plan(s->planner, start, goal, lc::WP_CAP, &plan[(size_t)id * lc::WP_CAP * 2], &k, &min_time[id]);
        if (rc) return fail(rc, "lc_create: the plan of agent %d has %d waypoints (capacity %d)", e, k, lc::WP_CAP);
        s->h_nplan[id] = k;
        noplan[id] = missing;
        if (k == 0) ++missing;
    }
    lc::WalledArrays& W = s->b.w;
    lc::WallRect* dw = nullptr; int* dwo = nullptr; lc::BvhNode* dbvh = nullptr;
    static_assert(sizeof(lc::WallRect) == sizeof(lc_rect) && sizeof(lc::BvhNode) == sizeof(lc_bvh_node), "layouts shared with the C ABI");
#define WCK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(LC_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); } while (0)
#define WAL(ptr, count) do { WCK(dalloc(&ptr, count)); s->allocs.push_back(ptr); } while (0)
    WAL(dw, nw); WAL(dwo, nw); WAL(dbvh, n_bvh);
    WAL(W.plan, (size_t)n * lc::WP_CAP); WAL(W.n_plan, n); WAL(W.num_done, n); WAL(W.lost, n); WAL(W.lost_next, n);
    WAL(W.noplan_before, n); WAL(W.req_pos, n);
    WCK(cudaMemcpy(dw, d->walls, sizeof(lc_rect) * (size_t)nw, cudaMemcpyHostToDevice));
    WCK(cudaMemcpy(dwo, worder.data(), sizeof(int) * (size_t)nw, cudaMemcpyHostToDevice));
    WCK(cudaMemcpy(dbvh, bvh.data(), sizeof(lc_bvh_node) * (size_t)n_bvh, cudaMemcpyHostToDevice));
    WCK(cudaMemcpy(W.plan, plan.data(), sizeof(float) * plan.size(), cudaMemcpyHostToDevice));
    WCK(cudaMemcpy(W.n_plan, s->h_nplan.data(), sizeof(int) * (size_t)n, cudaMemcpyHostToDevice));
    WCK(cudaMemcpy(W.noplan_before, noplan.data(), sizeof(int) * (size_t)n, cudaMemcpyHostToDevice));
    WCK(cudaMemset(W.num_done, 0, sizeof(int) * (size_t)n));
    WCK(cudaMemset(W.lost, 0, sizeof(int) * (size_t)n));
    WCK(cudaMemset(W.lost_next, 0, sizeof(int) * (size_t)n));
    WCK(cudaMemset(W.req_pos, 0, sizeof(float2) * (size_t)n));
    W.walls = dw; W.wall_order = dwo; W.bvh = dbvh;
    s->hp.n_bvh = n_bvh;
    // control block: draws per frame, no request pending
    lc::Ctl c0;
    memset(&c0, 0, sizeof c0);
    c0.wr1 = c0.wr2 = 0x7fffffff;
    c0.n_planned = n - missing;
    WCK(cudaMemcpy(s->b.ctl, &c0, sizeof c0, cudaMemcpyHostToDevice));
#undef WAL
#undef WCK
    return 0;
}

// lc_create (dom == nullptr) and lc_dd_create share everything but the choice of the initially local agents
static int create_impl(const lc_desc* d, const lc_domain_desc* dom, lc_sim** out) {
    if (!d || !out) return fail(LC_ERR_ARG, "lc_create: null argument");
    *out = nullptr;
    if (d->n < 2) return fail(LC_ERR_ARG, "lc_create: n must be >= 2 (the reference dereferences a null comm partner for n == 1, comm.cpp:273-285)");
    if (!d->pos || !d->goal) return fail(LC_ERR_ARG, "lc_create: pos/goal are required");
    if (d->n_walls < 0 || (d->n_walls > 0 && !d->walls)) return fail(LC_ERR_ARG, "lc_create: n_walls > 0 needs the wall rectangles");
    if (d->n_walls > 0 && dom) return fail(LC_ERR_ARG, "lc_dd_create: walled scenes run on one GPU (replans go through the host planner between frames)");
    if (!(d->radius > 0.f)) return fail(LC_ERR_ARG, "lc_create: radius must be positive");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(LC_ERR_CUDA, "lc_create: no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(e));
    if (d->device < 0 || d->device >= ndev) return fail(LC_ERR_ARG, "lc_create: device %d out of range", d->device);
    CK(cudaSetDevice(d->device));
    const int n_ids = d->n;
    // table order (quirk Q4): engine id = position in the reference's table iteration order
    const int32_t* order = d->table_order;
    if (order) {
        std::vector<char> seen((size_t)n_ids, 0);
        for (int i = 0; i < n_ids; ++i) {
            if (order[i] < 0 || order[i] >= n_ids || seen[order[i]]) return fail(LC_ERR_ARG, "lc_create: table_order is not a permutation of 0..n-1");
            seen[order[i]] = 1;
        }
    }
    auto ext = [order](int i) { return order ? order[i] : i; };
    // ---- domain decomposition: strip bounds, the initially owned subset, capacities ----
    const int world = dom ? dom->world : 1, rank = dom ? dom->rank : 0;
    std::vector<int> ids;
    std::vector<float> xb(lc::DD_MAX_WORLD + 1, 0.f);
    float halo = 0.f;
    int cap = n_ids, annex = 0, cap_halo = 0, dd_every = 1;
    float dd_skin_allow = 0.f;
    if (dom) {
        if (world < 2 || world > lc::DD_MAX_WORLD || rank < 0 || rank >= world)
            return fail(LC_ERR_ARG, "lc_dd_create: need 2 <= world <= %d and 0 <= rank < world", lc::DD_MAX_WORLD);
        if (d->flags & LC_FLAG_DEBUG_LISTS) return fail(LC_ERR_ARG, "lc_dd_create: LC_FLAG_DEBUG_LISTS is single-GPU only");
        if (n_ids < 2 * world) return fail(LC_ERR_ARG, "lc_dd_create: %d agents cannot be cut into %d strips", n_ids, world);
        const float R = d->radius + d->radius;
        const float reach = R + 2.f * (1.f / 60.f) * 1.01f + 0.002f + 0.002f;
        // |vel| <= 1.045 always (unit clamp, physics.cpp:45-48, then at most 9 % from push_away, physics.cpp:32-34);
        // the band covers the TTC cutoff at 1.05 plus the collision reach.  k_frame_begin checks the live cutoff against it.
        // Between grid re-sorts (every dd_every frames) owners, ghosts and slots are frozen and only the ghosts' records are
        // refreshed (domain.cuh): the band is widened by twice the movement allowed for in that time.
        {
            const char* e = getenv("LC_REBIN_EVERY");
            dd_every = e ? atoi(e) : 8;
            if (dd_every < 1) dd_every = 1;
            if (dd_every > 64) dd_every = 64;
        }
        dd_skin_allow = dd_every > 1 ? 0.05f * (float)dd_every + 0.1f : 0.f;
        halo = R + 10.f * 1.05f * 1.001f + 0.01f + reach + 0.05f + 2.f * dd_skin_allow;
        xb[0] = -3.0e38f; xb[world] = 3.0e38f;
        if (dom->x_bounds) {
            for (int r = 1; r < world; ++r) xb[r] = dom->x_bounds[r];
        } else {  // balanced strips (host.cpp lc_dd_strips: same on every rank, a function of the global table only)
            std::vector<float> xs(n_ids);
            for (int i = 0; i < n_ids; ++i) xs[i] = d->pos[2 * i];
            if (lc_dd_strips(n_ids, xs.data(), 1, world, halo, xb.data()) != LC_OK) return LC_ERR_ARG;
        }
        for (int r = 1; r + 1 < world; ++r)
            if (!(xb[r + 1] - xb[r] >= 1.05f * halo))
                return fail(LC_ERR_ARG, "lc_dd_create: strip %d is %.2f m wide; strips must be wider than the %.2f m ghost band", r, xb[r + 1] - xb[r], halo);
        for (int r = 1; r < world; ++r)
            if (!(xb[r] > xb[r - 1])) return fail(LC_ERR_ARG, "lc_dd_create: x_bounds must increase");
        for (int i = 0; i < n_ids; ++i) {
            const float x = d->pos[2 * ext(i)];
            if (x >= xb[rank] && x < xb[rank + 1]) ids.push_back(i);
        }
        cap_halo = dom->halo_capacity > 0 ? dom->halo_capacity : std::max(65536, n_ids / world / 2);
        cap = dom->capacity > 0 ? dom->capacity : (int)std::min<long long>((long long)n_ids + 2 * cap_halo, (long long)(n_ids / world) * 3 / 2 + 2 * cap_halo + 4096);
        if ((int)ids.size() > cap) return fail(LC_ERR_CAPACITY, "lc_dd_create: %d initial agents exceed the capacity %d", (int)ids.size(), cap);
        annex = world * lc::DD_CAND_CAP;
    }
    lc_sim* s = new lc_sim;
    s->device = d->device;
    s->n_ids = n_ids;
    s->dd_rank = rank; s->dd_world = world;
    const int n = s->n = cap;           // slot capacity == launch width
    const int n_loc = dom ? (int)ids.size() : n_ids;
    int rc = 0;
    auto bail = [&](int code) { lc_destroy(s); return code; };
#define CKB(x) do { rc = (x); if (rc) return bail(rc); } while (0)
#define CKC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { fail(LC_ERR_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); return bail(LC_ERR_CUDA); } } while (0)
    CKC(cudaStreamCreateWithFlags(&s->st, cudaStreamNonBlocking));
    CKC(cudaEventCreate(&s->ev0));
    CKC(cudaEventCreate(&s->ev1));
    {
        int lo = 0, hi = 0;
        CKC(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CKC(cudaStreamCreateWithPriority(&s->side.st, cudaStreamNonBlocking, hi));  // one block: let it in ahead of k_pairs
        CKC(cudaEventCreateWithFlags(&s->side.fork, cudaEventDisableTiming));
        CKC(cudaEventCreateWithFlags(&s->side.join, cudaEventDisableTiming));
    }
    CKB(alloc_arrays(s, s->b.set[0], n + annex));  // annex: comm-partner payload of other ranks' candidates (dd_global_body)
    CKB(alloc_arrays(s, s->b.set[1], n + annex));
    lc::Buffers& b = s->b;
    CKC(dalloc(&b.cellrank, n)); s->allocs.push_back(b.cellrank);
    CKC(dalloc(&b.slot_cell, n)); s->allocs.push_back(b.slot_cell);
    CKC(cudaMemset(b.slot_cell, 0, sizeof(int) * (size_t)n));
    CKC(dalloc(&b.cell_count, lc::CELL_ALLOC)); s->allocs.push_back(b.cell_count);
    CKC(cudaMemset(b.cell_count, 0, sizeof(int) * lc::CELL_ALLOC));  // k_scan_hull leaves it cleared after every frame
    CKC(dalloc(&b.cell_start, lc::CELL_ALLOC)); s->allocs.push_back(b.cell_start);
    CKC(dalloc(&b.tile_agg, lc::SCAN_TILES)); s->allocs.push_back(b.tile_agg);
    CKC(cudaMemset(b.tile_agg, 0, sizeof(unsigned long long) * lc::SCAN_TILES));
    CKC(dalloc(&b.slot_of_id, n_ids)); s->allocs.push_back(b.slot_of_id);
    CKC(dalloc(&b.cand, n + annex)); s->allocs.push_back(b.cand);
    CKC(dalloc(&b.cand2, lc::REFINE_MAX + annex)); s->allocs.push_back(b.cand2);
    CKC(dalloc(&b.blkred, (size_t)(n + 127) / 128)); s->allocs.push_back(b.blkred);
    CKC(dalloc(&b.blkext, (size_t)((n + 1023) / 1024) * lc::HULL_K)); s->allocs.push_back(b.blkext);
    CKC(dalloc(&b.near_cnt, n)); s->allocs.push_back(b.near_cnt);
    CKC(dalloc(&b.near_idx, (size_t)n * lc::MAX_NEAR)); s->allocs.push_back(b.near_idx);
    CKC(dalloc(&b.pair_cnt, n)); s->allocs.push_back(b.pair_cnt);
    CKC(dalloc(&b.pair_list, (size_t)n * lc::PAIR_CAP)); s->allocs.push_back(b.pair_list);
    CKC(cudaMemset(b.pair_cnt, 0, sizeof(int) * n));
    CKC(dalloc(&b.events, (size_t)std::max(n, world * lc::DD_EV_CAP))); s->allocs.push_back(b.events);
    CKC(dalloc(&b.ev_by_id, n_ids)); s->allocs.push_back(b.ev_by_id);
    CKC(dalloc(&b.partner_id, n_ids)); s->allocs.push_back(b.partner_id);
    CKC(dalloc(&b.nbr_cnt, n_ids)); s->allocs.push_back(b.nbr_cnt);
    CKC(dalloc(&b.diag_partner, n)); s->allocs.push_back(b.diag_partner);
    CKC(dalloc(&b.diag_cnt, n)); s->allocs.push_back(b.diag_cnt);
    CKC(cudaMemset(b.diag_partner, 0xff, sizeof(int) * n));
    CKC(cudaMemset(b.diag_cnt, 0, sizeof(int) * n));
    b.list_cap = 0;
    b.nbr_list = nullptr;
    if (d->flags & LC_FLAG_DEBUG_LISTS) {
        b.list_cap = n_ids - 1 < 256 ? n_ids - 1 : 256;
        CKC(dalloc(&b.nbr_list, (size_t)n_ids * b.list_cap)); s->allocs.push_back(b.nbr_list);
        CKC(cudaMemset(b.nbr_list, 0xff, sizeof(int) * (size_t)n_ids * b.list_cap));
    }
    CKC(dalloc(&b.fin_log, (size_t)3 * n_ids)); s->allocs.push_back(b.fin_log);
    CKC(dalloc(&b.ctl, 1)); s->allocs.push_back(b.ctl);
    CKC(cudaMemset(b.ctl, 0, sizeof(lc::Ctl)));
    CKC(cudaMemset(b.partner_id, 0xff, sizeof(int) * n_ids));
    CKC(cudaMemset(b.nbr_cnt, 0, sizeof(int) * n_ids));
    CKC(cudaMemset(b.near_cnt, 0, sizeof(int) * n));
    CKC(cudaMalloc((void**)&s->dP, sizeof(lc::Params)));
    CKC(cudaMallocHost((void**)&s->hctl, sizeof(lc::Ctl)));
    s->stage_floats = std::max((size_t)n_ids * 16, (size_t)8 + 5 * (size_t)n);  // lc_get_state tables / lc_dd_get_shard
    CKC(cudaMalloc((void**)&s->stage, sizeof(float) * s->stage_floats));
    b.ext_of = nullptr;
    if (order) s->h_order.assign(order, order + n_ids);
    if (order) {
        CKC(dalloc(&s->ext_of, n_ids)); s->allocs.push_back(s->ext_of);
        CKC(cudaMemcpy(s->ext_of, order, sizeof(int) * (size_t)n_ids, cudaMemcpyHostToDevice));
        CKC(dalloc(&s->xlate, (size_t)n_ids * std::max(1, b.list_cap))); s->allocs.push_back(s->xlate);
        b.ext_of = s->ext_of;
    }
    b.dd = nullptr;
    if (dom) {
        s->lay = lc::domain_layout(world, n, cap_halo, n_ids);
        CKC(cudaMalloc((void**)&s->mailbox, s->lay.total));
        CKC(cudaMemset(s->mailbox, 0, s->lay.total));
        CKC(cudaMemset(s->mailbox + s->lay.ev_by_id, 0xff, sizeof(float) * (size_t)n_ids));  // NaN = "no finish time"
        CKC(dalloc(&s->ddev, 1)); s->allocs.push_back(s->ddev);
        CKC(dalloc(&s->owned_dev, n_ids)); s->allocs.push_back(s->owned_dev);
        lc::DomainDev hd{};
        hd.lay = s->lay;
        hd.mbox[rank] = s->mailbox;
        hd.replan[0] = hd.replan[1] = 0x7fffffff;
        {   // per-slot "stays on this rank" marks written by the pack, read by the next frame's sort
            unsigned char* keep = nullptr;
            CKC(cudaMalloc((void**)&keep, (size_t)n)); s->allocs.push_back(keep);
            CKC(cudaMemset(keep, 0, (size_t)n));
            hd.keep = keep;
            int2* si = nullptr;
            CKC(cudaMalloc((void**)&si, sizeof(int2) * (size_t)n)); s->allocs.push_back(si);
            CKC(cudaMemset(si, 0xff, sizeof(int2) * (size_t)n));
            hd.send_idx = si;
            hd.skin_allow = dd_skin_allow;
        }
        {   // how long a rank waits for a peer's flag before it reports LC_ERR_CUDA instead of hanging the GPU
            const char* e = getenv("LC_DD_TIMEOUT_MS");
            const unsigned long long ms = e ? strtoull(e, nullptr, 10) : 10000ull;
            hd.spin_timeout_ns = ms * 1000ull * 1000ull;
        }
        CKC(cudaMemcpy(s->ddev, &hd, sizeof hd, cudaMemcpyHostToDevice));
        b.dd = s->ddev;
    }

    lc::Params& P = s->hp;
    P.n = n_loc;
    P.n_in = n_loc;
    P.n_ids = n_ids;
    P.cap = n;
    P.dd_rank = rank; P.dd_world = world;
    for (int r = 0; r <= lc::DD_MAX_WORLD; ++r) P.dd_xb[r] = xb[r];
    P.dd_halo = halo;
    P.r = d->radius;
    P.R = d->radius + d->radius;       // LMP.cpp:53 i._r + j._r
    P.RR = P.R * P.R;                  // BoundVolume.cpp:106 _r * _r
    {   // pairs that can overlap after this frame's move: both agents travel <= dt (|vel_forehalf| <= 1 after the clamp)
        float reach = P.R + 2.f * (1.f / 60.f) * 1.01f + 0.002f;
        P.near_c = reach * reach - P.RR;
        P.near_reach = reach + 0.002f;
    }
    P.max_duration = d->max_duration;
    P.lcg_stride = lcg_pow(16807u, (uint64_t)n_ids);
    P.flags = d->flags;
    P.stencil_req = d->stencil > 0 ? d->stencil : 2;
    {   // grid re-sort period (crowd.cuh: skin); a decomposed crowd also freezes owners and ghosts in between (domain.cuh)
        const char* e = getenv("LC_REBIN_EVERY");
        const int k = e ? atoi(e) : 8;
        P.rebin_every = dom ? dd_every : (k < 1 ? 1 : k);
    }
    {   // directions of the second comm-partner pruning stage: 0 = chosen per frame by the length of the stage-1 list
        const char* e = getenv("LC_HULL_K2");
        const int k = e ? atoi(e) : 0;
        P.hull_k2 = (k == 64 || k == 128 || k == 256) ? k : 0;
    }
    P.n_walls = d->n_walls;
    P.n_bvh = 0;
    P.rng0 = d->rng_state;
    memset(P.m_cv, 0, sizeof P.m_cv);
    memset(P.m_f, 0, sizeof P.m_f);
    if (d->m_cv) memcpy(P.m_cv, d->m_cv, sizeof P.m_cv);
    if (d->m_f) memcpy(P.m_f, d->m_f, sizeof P.m_f);
    std::vector<float> w_min_time;  // walled scenes: min_time by engine id (ai.cpp:118-126)
    memset(&b.w, 0, sizeof b.w);
    if (d->n_walls > 0) CKB(setup_walls(s, d, ext, w_min_time));
    CKB(upload_params(s));
    lc::upload_dirs();

    // initial component values: demo.cpp:563-581 create_robo, ai.cpp:107-127 ai::init, physics.cpp:15-21 prewarm
    {
        const size_t m = (size_t)n_loc;
        std::vector<float> h(m * 13 + 1);
        float* hpos = h.data();             // [m][2]
        float* hgoal = h.data() + 2 * m;    // [m][2]
        float* facing = h.data() + 4 * m;   // [m][2]
        float* gd = h.data() + 6 * m;       // [m]
        float* mt = h.data() + 7 * m;       // [m]
        float* hvel = h.data() + 8 * m;     // [m][2]  (domain mode: optional developed-flow start)
        float* hvfh = h.data() + 10 * m;    // [m][2]
        int* hid = reinterpret_cast<int*>(h.data() + 12 * m);  // [m]
        const bool has_v = dom && dom->vel && dom->vel_forehalf;
        for (size_t k = 0; k < m; ++k) {
            const int id = dom ? ids[k] : (int)k;   // engine id (table position)
            const int i = ext(id);                   // the caller's agent index
            hpos[2 * k] = d->pos[2 * i]; hpos[2 * k + 1] = d->pos[2 * i + 1];
            hgoal[2 * k] = d->goal[2 * i]; hgoal[2 * k + 1] = d->goal[2 * i + 1];
            float dx = d->goal[2 * i] - d->pos[2 * i], dy = d->goal[2 * i + 1] - d->pos[2 * i + 1];
            facing[2 * k] = dx; facing[2 * k + 1] = dy;      // un-normalised (demo.cpp:580)
            float len = sqrtf(dx * dx + dy * dy);            // glm::length (ai.cpp:116)
            gd[k] = len;
            float mn = -1.0f;                                 // ai.cpp:121-126
            mn += len;
            mt[k] = w_min_time.empty() ? mn : w_min_time[id];  // walled: length of the planned path instead of the bee line
            hid[k] = id;
            if (has_v) {
                hvel[2 * k] = dom->vel[2 * i]; hvel[2 * k + 1] = dom->vel[2 * i + 1];
                hvfh[2 * k] = dom->vel_forehalf[2 * i]; hvfh[2 * k + 1] = dom->vel_forehalf[2 * i + 1];
            }
        }
        float* dpos = s->stage;
        float* dgoal = s->stage + 2 * m;
        float* dfacing = s->stage + 4 * m;
        float* dgd = s->stage + 6 * m;
        float* dmt = s->stage + 7 * m;
        float* dvel = s->stage + 8 * m;
        float* dvfh = s->stage + 10 * m;
        int* did = reinterpret_cast<int*>(s->stage + 12 * m);
        CKC(cudaMemcpyAsync(s->stage, h.data(), sizeof(float) * 13 * m, cudaMemcpyHostToDevice, s->st));
        lc::launch_set_state(s->dP, n, b, (const float2*)dpos, has_v ? (const float2*)dvel : nullptr,
                             has_v ? (const float2*)dvfh : nullptr, nullptr, (const float2*)dfacing, nullptr,
                             dgd, nullptr, (const float2*)dgoal, dmt, 1, s->st, &s->launches, dom ? did : nullptr);
        CKC(cudaStreamSynchronize(s->st));  // h goes out of scope
    }
    s->rng0 = d->rng_state;
    lc::launch_set_rng(s->dP, n, b, s->rng0, s->st, &s->launches);
    lc::launch_init_reduce(s->dP, n, b, s->st, &s->launches);
    if (d->n_walls > 0) lc::launch_walled_gate(s->dP, n, b, s->st, &s->launches);  // replan requests of frame 1
    CKC(cudaStreamSynchronize(s->st));
    CKC(cudaGetLastError());
    CKB(build_graphs(s));
    *out = s;
    return LC_OK;
#undef CKB
#undef CKC
}

// no exception crosses the C ABI (learn_comms.h): allocation failures of the host containers become LC_ERR_CAPACITY
static int create_guarded(const lc_desc* d, const lc_domain_desc* dom, lc_sim** out) {
    try {
        return create_impl(d, dom, out);
    } catch (const std::exception& e) {
        if (out) *out = nullptr;
        return fail(LC_ERR_CAPACITY, "lc_create: host allocation failed (%s)", e.what());
    } catch (...) {
        if (out) *out = nullptr;
        return fail(LC_ERR_CAPACITY, "lc_create: host allocation failed");
    }
}

extern "C" {
int32_t lc_create(const lc_desc* d, lc_sim** out) { return create_guarded(d, nullptr, out); }
int32_t lc_dd_create(const lc_desc* d, const lc_domain_desc* dom, lc_sim** out) {
    if (!dom) return fail(LC_ERR_ARG, "lc_dd_create: null domain description");
    return create_guarded(d, dom, out);
}
}  // extern "C"

#define SIM_ENTER(s)                                                   \
    if (!(s)) return fail(LC_ERR_ARG, "%s: null handle", __func__);    \
    CK(cudaSetDevice((s)->device))

// the status a device-side error maps to (Ctl::error holds the lc_status of the first one), from the host copy of Ctl
static int device_error(lc_sim* s) {
    const int e = s->hctl->error;
    if (!e) return LC_OK;
    const char* what = e == LC_ERR_STATE ? "non-finite state or an agent without comm partner, comm.cpp:273-285"
                     : e == LC_ERR_CAPACITY ? "a fixed-capacity exchange / candidate / event buffer overflowed"
                     : e == LC_ERR_CUDA ? "a peer rank did not arrive (spin time-out)" : "unexpected";
    return fail(e, "device-side error %d at frame %u (%s)", e, s->hctl->frame, what);
}
static int check_device_error(lc_sim* s) {
    int rc = read_ctl(s);
    if (rc) return rc;
    return device_error(s);
}

extern "C" {

int32_t lc_set_weights(lc_sim* s, const float* m_cv, const float* m_f) {
    SIM_ENTER(s);
    if (m_cv) memcpy(s->hp.m_cv, m_cv, sizeof s->hp.m_cv);
    if (m_f) memcpy(s->hp.m_f, m_f, sizeof s->hp.m_f);
    CK(cudaStreamSynchronize(s->st));  // hp is the copy source; keep it stable
    return upload_weights(s);
}

// domain mode: the first frame's records are packed by a kernel of its own (every later frame's by the previous k_collide)
static void ensure_packed(lc_sim* s) {
    if (s->ddev && !s->dd_packed) {
        lc::launch_dd_pack(s->dP, s->n, s->b, s->st, &s->launches);
        s->dd_packed = true;
    }
}

static int enqueue_frames(lc_sim* s, int nframes) {
    int left = nframes;
    for (int g = GRAPH_LEVELS - 1; g >= 0; --g) {
        const int frames = 1 << g;
        while (left >= frames) {
            CK(cudaGraphLaunch(s->gexec[g], s->st));
            left -= frames;
            s->launches += (long long)(lc::FRAME_KERNELS + (s->ddev ? lc::DD_KERNELS : 0)) * frames;
        }
    }
    return LC_OK;
}

// Walled scenes: GMP::plan_one for the (at most two) agents the paused frame's gate selected (ai.cpp:151-168), in table
// order, from the position the agent asked from; new waypoint list, num_done = 0, lost_frames = 0 (GMP.cpp:109, ai.cpp:167).
static int serve_replans(lc_sim* s) {
    const lc::WalledArrays& W = s->b.w;
    const int ids[2] = {s->hctl->wr1, s->hctl->wr2};
    bool plan_set_changed = false;
    int served = 0;
    for (int q = 0; q < 2; ++q) {
        const int id = ids[q];
        if (id == 0x7fffffff) continue;
        if (id < 0 || id >= s->n_ids) return fail(LC_ERR_STATE, "lc_step: replan request for agent %d out of range", id);
        float start[2];
        CK(cudaMemcpy(start, W.req_pos + id, sizeof start, cudaMemcpyDeviceToHost));
        float wp[2 * lc::WP_CAP];
        int32_t k = 0;
        int rc = lc_planner_plan(s->planner, start, &s->h_goal[2 * (size_t)id], lc::WP_CAP, wp, &k, nullptr);
        if (rc) return fail(rc, "lc_step: a replan produced %d waypoints (capacity %d)", k, lc::WP_CAP);
        const int zero = 0;
        if (k > 0) CK(cudaMemcpy(W.plan + (size_t)id * lc::WP_CAP, wp, sizeof(float) * 2 * (size_t)k, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(W.n_plan + id, &k, sizeof(int), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(W.num_done + id, &zero, sizeof(int), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(W.lost_next + id, &zero, sizeof(int), cudaMemcpyHostToDevice));
        if ((s->h_nplan[id] == 0) != (k == 0)) plan_set_changed = true;
        s->h_nplan[id] = k;
        ++served;
    }
    int n_planned = s->hctl->n_planned;
    if (plan_set_changed) {  // who draws from the global stream changed (LMP.cpp:259-270)
        std::vector<int> noplan((size_t)s->n_ids);
        int missing = 0;
        for (int i = 0; i < s->n_ids; ++i) { noplan[i] = missing; missing += s->h_nplan[i] == 0; }
        CK(cudaMemcpy(W.noplan_before, noplan.data(), sizeof(int) * noplan.size(), cudaMemcpyHostToDevice));
        n_planned = s->n_ids - missing;
    }
    s->replans += served;
    const int one = 1, zero = 0;
    char* c = (char*)s->b.ctl;
    CK(cudaMemcpy(c + offsetof(lc::Ctl, n_planned), &n_planned, sizeof(int), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c + offsetof(lc::Ctl, w_served), &one, sizeof(int), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c + offsetof(lc::Ctl, halt), &zero, sizeof(int), cudaMemcpyHostToDevice));
    return LC_OK;
}

int32_t lc_step(lc_sim* s, int32_t nframes) {
    SIM_ENTER(s);
    if (nframes < 0) return fail(LC_ERR_ARG, "lc_step: nframes < 0");
    if (nframes > 0) ensure_packed(s);
    if (s->hp.n_walls == 0) return enqueue_frames(s, nframes);
    // Walled scene: frames pause on the device whenever the replan gate fires; serve and resume until the requested
    // number of frames has executed (or demo::run reported all_done: later frames are no-ops, as without walls).
    int rc = read_ctl(s);
    if (rc) return rc;
    const uint32_t target = s->hctl->frame + (uint32_t)nframes;
    for (int guard = 0; s->hctl->frame < target; ++guard) {
        if (s->hctl->halt == 2) {
            rc = serve_replans(s);
            if (rc) return rc;
        } else if (s->hctl->halt == 1 || guard > 4 * nframes + 8) {
            break;  // finished (frames after all_done do not execute), or no progress: never spin
        }
        rc = enqueue_frames(s, (int)(target - s->hctl->frame));
        if (rc) return rc;
        rc = read_ctl(s);
        if (rc) return rc;
        if (s->hctl->error) return device_error(s);
    }
    return LC_OK;
}

int32_t lc_sync(lc_sim* s) {
    SIM_ENTER(s);
    CK(cudaStreamSynchronize(s->st));
    CK(cudaGetLastError());
    return LC_OK;
}

int32_t lc_finished(lc_sim* s, int32_t* finished, uint32_t* frame) {
    SIM_ENTER(s);
    int rc = read_ctl(s);
    if (rc) return rc;
    if (finished) *finished = s->hctl->finished;
    if (frame) *frame = s->hctl->frame;
    return device_error(s);  // a host that only steps and polls still sees Ctl::error
}

int32_t lc_run_to_completion(lc_sim* s, int32_t* frames) {
    SIM_ENTER(s);
    // main.cpp:98 while (!all_done): frames after the all_done frame are device-side no-ops, so stepping in
    // batches and polling the latch executes exactly the reference's frame count.
    for (;;) {
        int rc = lc_step(s, 4 * GRAPH_BATCH);
        if (rc) return rc;
        rc = check_device_error(s);
        if (rc) return rc;
        if (s->hctl->finished) break;
    }
    if (frames) *frames = (int32_t)s->hctl->frame;
    return LC_OK;
}

int32_t lc_get_state(lc_sim* s, float* pos, float* vel, float* vfh, float* c, float* facing, float* lgoal,
                     float* goal_dist, int32_t* num_done) {
    SIM_ENTER(s);
    const size_t n = (size_t)s->n_ids;
    float* st = s->stage;
    // float2 tables first: every offset is a multiple of 2 n floats, i.e. 8-byte aligned for odd n too
    float2* dpos = (float2*)st; float2* dvel = (float2*)(st + 2 * n); float2* dvfh = (float2*)(st + 4 * n);
    float2* dfac = (float2*)(st + 6 * n); float2* dlg = (float2*)(st + 8 * n); float* dc = st + 10 * n;
    float* dgd = st + 13 * n; int* dnd = (int*)(st + 14 * n);
    if (s->dd_world > 1) {  // domain mode: entries of agents owned by other ranks read as zero; lc_dd_get_owned tells which
        CK(cudaMemsetAsync(st, 0, sizeof(float) * s->stage_floats, s->st));
        CK(cudaMemsetAsync(s->owned_dev, 0, sizeof(int) * n, s->st));
    }
    lc::launch_get_state(s->dP, s->n, s->b, pos ? dpos : nullptr, vel ? dvel : nullptr, vfh ? dvfh : nullptr,
                         c ? dc : nullptr, facing ? dfac : nullptr, lgoal ? dlg : nullptr, goal_dist ? dgd : nullptr,
                         num_done ? dnd : nullptr, s->st, &s->launches, s->owned_dev);
    if (pos) CK(cudaMemcpyAsync(pos, dpos, sizeof(float) * 2 * n, cudaMemcpyDeviceToHost, s->st));
    if (vel) CK(cudaMemcpyAsync(vel, dvel, sizeof(float) * 2 * n, cudaMemcpyDeviceToHost, s->st));
    if (vfh) CK(cudaMemcpyAsync(vfh, dvfh, sizeof(float) * 2 * n, cudaMemcpyDeviceToHost, s->st));
    if (c) CK(cudaMemcpyAsync(c, dc, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost, s->st));
    if (facing) CK(cudaMemcpyAsync(facing, dfac, sizeof(float) * 2 * n, cudaMemcpyDeviceToHost, s->st));
    if (lgoal) CK(cudaMemcpyAsync(lgoal, dlg, sizeof(float) * 2 * n, cudaMemcpyDeviceToHost, s->st));
    if (goal_dist) CK(cudaMemcpyAsync(goal_dist, dgd, sizeof(float) * n, cudaMemcpyDeviceToHost, s->st));
    if (num_done) CK(cudaMemcpyAsync(num_done, dnd, sizeof(int) * n, cudaMemcpyDeviceToHost, s->st));
    return check_device_error(s);  // synchronises; the tables are delivered either way
}

int32_t lc_set_state(lc_sim* s, const float* pos, const float* vel, const float* vfh, const float* c,
                     const float* facing, const float* lgoal, const float* goal_dist, const int32_t* num_done) {
    SIM_ENTER(s);
    if (s->dd_world > 1) return fail(LC_ERR_STATE, "lc_set_state: not available for a domain-decomposed crowd (pass the initial tables to lc_dd_create)");
    const size_t n = (size_t)s->n;
    float* st = s->stage;
    // float2 tables first: every offset is a multiple of 2 n floats, i.e. 8-byte aligned for odd n too
    float2* dpos = (float2*)st; float2* dvel = (float2*)(st + 2 * n); float2* dvfh = (float2*)(st + 4 * n);
    float2* dfac = (float2*)(st + 6 * n); float2* dlg = (float2*)(st + 8 * n); float* dc = st + 10 * n;
    float* dgd = st + 13 * n; int* dnd = (int*)(st + 14 * n);
    if (pos) CK(cudaMemcpyAsync(dpos, pos, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, s->st));
    if (vel) CK(cudaMemcpyAsync(dvel, vel, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, s->st));
    if (vfh) CK(cudaMemcpyAsync(dvfh, vfh, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, s->st));
    if (c) CK(cudaMemcpyAsync(dc, c, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, s->st));
    if (facing) CK(cudaMemcpyAsync(dfac, facing, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, s->st));
    if (lgoal) CK(cudaMemcpyAsync(dlg, lgoal, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, s->st));
    if (goal_dist) CK(cudaMemcpyAsync(dgd, goal_dist, sizeof(float) * n, cudaMemcpyHostToDevice, s->st));
    if (num_done) CK(cudaMemcpyAsync(dnd, num_done, sizeof(int) * n, cudaMemcpyHostToDevice, s->st));
    lc::launch_set_state(s->dP, s->n, s->b, pos ? dpos : nullptr, vel ? dvel : nullptr, vfh ? dvfh : nullptr,
                         c ? dc : nullptr, facing ? dfac : nullptr, lgoal ? dlg : nullptr, goal_dist ? dgd : nullptr,
                         num_done ? dnd : nullptr, nullptr, nullptr, 0, s->st, &s->launches);
    lc::launch_init_reduce(s->dP, s->n, s->b, s->st, &s->launches);
    if (s->hp.n_walls > 0) lc::launch_walled_gate(s->dP, s->n, s->b, s->st, &s->launches);  // requests of the coming frame
    return LC_OK;
}

// lc_get_state(out...) followed by lc_set_state(in...), as one call: the upload of the `in` tables starts at once on a copy
// stream of its own — beside the frame that may still be executing and beside the download of the `out` tables (the two
// directions of the link are independent) — and is applied once the current state has been read out.
int32_t lc_exchange_state(lc_sim* s, float* o_pos, float* o_vel, float* o_vfh, float* o_c, float* o_facing, float* o_lgoal,
                          float* o_goal_dist, int32_t* o_num_done, const float* pos, const float* vel, const float* vfh,
                          const float* c, const float* facing, const float* lgoal, const float* goal_dist,
                          const int32_t* num_done) {
    SIM_ENTER(s);
    if (s->dd_world > 1) return fail(LC_ERR_STATE, "lc_exchange_state: not available for a domain-decomposed crowd");
    const size_t n = (size_t)s->n;
    if (!s->stage_in) {
        CK(cudaMalloc((void**)&s->stage_in, sizeof(float) * 16 * n));
        CK(cudaStreamCreateWithFlags(&s->st_in, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&s->ev_in, cudaEventDisableTiming));
    }
    // ---- upload (copy stream): nothing on the compute stream reads stage_in until the event below ----
    {
        float* st = s->stage_in;
        float2* dpos = (float2*)st; float2* dvel = (float2*)(st + 2 * n); float2* dvfh = (float2*)(st + 4 * n);
        float2* dfac = (float2*)(st + 6 * n); float2* dlg = (float2*)(st + 8 * n); float* dc = st + 10 * n;
        float* dgd = st + 13 * n; int* dnd = (int*)(st + 14 * n);
        if (pos) CK(cudaMemcpyAsync(dpos, pos, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, s->st_in));
        if (vel) CK(cudaMemcpyAsync(dvel, vel, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, s->st_in));
        if (vfh) CK(cudaMemcpyAsync(dvfh, vfh, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, s->st_in));
        if (c) CK(cudaMemcpyAsync(dc, c, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, s->st_in));
        if (facing) CK(cudaMemcpyAsync(dfac, facing, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, s->st_in));
        if (lgoal) CK(cudaMemcpyAsync(dlg, lgoal, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, s->st_in));
        if (goal_dist) CK(cudaMemcpyAsync(dgd, goal_dist, sizeof(float) * n, cudaMemcpyHostToDevice, s->st_in));
        if (num_done) CK(cudaMemcpyAsync(dnd, num_done, sizeof(int) * n, cudaMemcpyHostToDevice, s->st_in));
        CK(cudaEventRecord(s->ev_in, s->st_in));
    }
    // ---- download (compute stream): gather into id order, then the copies ----
    {
        float* st = s->stage;
        float2* dpos = (float2*)st; float2* dvel = (float2*)(st + 2 * n); float2* dvfh = (float2*)(st + 4 * n);
        float2* dfac = (float2*)(st + 6 * n); float2* dlg = (float2*)(st + 8 * n); float* dc = st + 10 * n;
        float* dgd = st + 13 * n; int* dnd = (int*)(st + 14 * n);
        lc::launch_get_state(s->dP, s->n, s->b, o_pos ? dpos : nullptr, o_vel ? dvel : nullptr, o_vfh ? dvfh : nullptr,
                             o_c ? dc : nullptr, o_facing ? dfac : nullptr, o_lgoal ? dlg : nullptr, o_goal_dist ? dgd : nullptr,
                             o_num_done ? dnd : nullptr, s->st, &s->launches, s->owned_dev);
        if (o_pos) CK(cudaMemcpyAsync(o_pos, dpos, sizeof(float) * 2 * n, cudaMemcpyDeviceToHost, s->st));
        if (o_vel) CK(cudaMemcpyAsync(o_vel, dvel, sizeof(float) * 2 * n, cudaMemcpyDeviceToHost, s->st));
        if (o_vfh) CK(cudaMemcpyAsync(o_vfh, dvfh, sizeof(float) * 2 * n, cudaMemcpyDeviceToHost, s->st));
        if (o_c) CK(cudaMemcpyAsync(o_c, dc, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost, s->st));
        if (o_facing) CK(cudaMemcpyAsync(o_facing, dfac, sizeof(float) * 2 * n, cudaMemcpyDeviceToHost, s->st));
        if (o_lgoal) CK(cudaMemcpyAsync(o_lgoal, dlg, sizeof(float) * 2 * n, cudaMemcpyDeviceToHost, s->st));
        if (o_goal_dist) CK(cudaMemcpyAsync(o_goal_dist, dgd, sizeof(float) * n, cudaMemcpyDeviceToHost, s->st));
        if (o_num_done) CK(cudaMemcpyAsync(o_num_done, dnd, sizeof(int) * n, cudaMemcpyDeviceToHost, s->st));
        // ---- apply the new state (it has been arriving on the copy stream all the while) ----
        CK(cudaStreamWaitEvent(s->st, s->ev_in, 0));
        float* si = s->stage_in;
        lc::launch_set_state(s->dP, s->n, s->b, pos ? (const float2*)si : nullptr, vel ? (const float2*)(si + 2 * n) : nullptr,
                             vfh ? (const float2*)(si + 4 * n) : nullptr, c ? si + 10 * n : nullptr,
                             facing ? (const float2*)(si + 6 * n) : nullptr, lgoal ? (const float2*)(si + 8 * n) : nullptr,
                             goal_dist ? si + 13 * n : nullptr, num_done ? (const int*)(si + 14 * n) : nullptr, nullptr, nullptr, 0,
                             s->st, &s->launches);
        lc::launch_init_reduce(s->dP, s->n, s->b, s->st, &s->launches);
        if (s->hp.n_walls > 0) lc::launch_walled_gate(s->dP, s->n, s->b, s->st, &s->launches);
    }
    return check_device_error(s);  // synchronises: the out tables are complete, the new state is in place
}

// stream state before the first draw of the NEXT frame; every frame consumes n draws (LMP.cpp:261-263)
int32_t lc_get_rng(lc_sim* s, uint32_t* rng_state) {
    SIM_ENTER(s);
    int rc = read_ctl(s);
    if (rc) return rc;
    if (s->hp.n_walls > 0)  // plan-less agents do not draw (LMP.cpp:267-270): the device counts the draws
        *rng_state = lcm::lcg_mul(s->hp.rng0, lcg_pow(16807u, (uint64_t)s->hctl->draws + (uint64_t)s->hctl->drawn_last));
    else
        *rng_state = lcm::lcg_mul(s->rng0, lcg_pow(16807u, (uint64_t)s->hctl->frame * (uint64_t)s->n_ids));
    return LC_OK;
}

int32_t lc_set_rng(lc_sim* s, uint32_t rng_state) {
    SIM_ENTER(s);
    if (s->ddev) return fail(LC_ERR_STATE, "lc_set_rng: not available on a domain-decomposed crowd (the next frame's records are already packed)");
    int rc = read_ctl(s);
    if (rc) return rc;
    if (s->hp.n_walls > 0) {  // rebase the stream so that the next draw is the first after rng_state
        const uint64_t used = (uint64_t)s->hctl->draws + (uint64_t)s->hctl->drawn_last;
        s->hp.rng0 = lcm::lcg_mul(rng_state, lcg_pow(lcg_pow(16807u, used), 2147483647ull - 2ull));
        CK(cudaStreamSynchronize(s->st));
        CK(cudaMemcpy((char*)s->dP + offsetof(lc::Params, rng0), &s->hp.rng0, sizeof(uint32_t), cudaMemcpyHostToDevice));
        return LC_OK;
    }
    // rng0 is defined relative to frame 0; rebase so that lc_get_rng stays consistent
    uint64_t consumed = (uint64_t)s->hctl->frame * (uint64_t)s->n_ids;
    // inverse of 16807^consumed mod (2^31-1): a^(m-2) is the modular inverse (m prime)
    uint32_t inv = lcg_pow(lcg_pow(16807u, consumed), 2147483647ull - 2ull);
    s->rng0 = lcm::lcg_mul(rng_state, inv);
    lc::launch_set_rng(s->dP, s->n, s->b, rng_state, s->st, &s->launches);
    return LC_OK;
}

int32_t lc_get_scores(lc_sim* s, float* avg_velocity, float* group_time, float* run_avg, float* run_std,
                      int32_t* total_num_done) {
    SIM_ENTER(s);
    lc::launch_finalize(s->dP, s->b, s->st, &s->launches);
    int rc = read_ctl(s);
    if (rc) return rc;
    const lc::Ctl& c = *s->hctl;
    if (avg_velocity) *avg_velocity = c.avg_velocity;
    if (group_time) *group_time = c.run_avg + c.run_std * 3;  // comm.cpp:139
    if (run_avg) *run_avg = c.run_avg;
    if (run_std) *run_std = c.run_std;
    if (total_num_done) *total_num_done = c.total_num_done;
    if (c.error) return fail(c.error, "device-side error %d at frame %u", c.error, c.frame);
    return LC_OK;
}

int32_t lc_get_partners(lc_sim* s, int32_t* partner) {
    SIM_ENTER(s);
    lc::launch_get_diag(s->dP, s->n, s->b, s->st, &s->launches);
    const int* src = s->b.partner_id;
    if (s->ext_of) { lc::launch_to_external(s->b, src, s->xlate, s->n_ids, 1, 1, s->st, &s->launches); src = s->xlate; }
    CK(cudaMemcpyAsync(partner, src, sizeof(int) * (size_t)s->n_ids, cudaMemcpyDeviceToHost, s->st));
    CK(cudaStreamSynchronize(s->st));
    return LC_OK;
}

int32_t lc_get_neighbor_counts(lc_sim* s, int32_t* count) {
    SIM_ENTER(s);
    lc::launch_get_diag(s->dP, s->n, s->b, s->st, &s->launches);
    const int* src = s->b.nbr_cnt;
    if (s->ext_of) { lc::launch_to_external(s->b, src, s->xlate, s->n_ids, 1, 0, s->st, &s->launches); src = s->xlate; }
    CK(cudaMemcpyAsync(count, src, sizeof(int) * (size_t)s->n_ids, cudaMemcpyDeviceToHost, s->st));
    CK(cudaStreamSynchronize(s->st));
    return LC_OK;
}

int32_t lc_get_neighbor_lists(lc_sim* s, int32_t* idx, int32_t cap) {
    SIM_ENTER(s);
    if (!s->b.nbr_list) return fail(LC_ERR_ARG, "lc_get_neighbor_lists: create the simulation with LC_FLAG_DEBUG_LISTS");
    if (cap != s->b.list_cap) return fail(LC_ERR_ARG, "lc_get_neighbor_lists: cap_per_agent must be %d", s->b.list_cap);
    const int* src = s->b.nbr_list;   // rows and entries in table order (the order LMP.cpp:288-304 accumulates in)
    if (s->ext_of) { lc::launch_to_external(s->b, src, s->xlate, s->n_ids, cap, 1, s->st, &s->launches); src = s->xlate; }
    CK(cudaMemcpyAsync(idx, src, sizeof(int) * (size_t)s->n_ids * cap, cudaMemcpyDeviceToHost, s->st));
    CK(cudaStreamSynchronize(s->st));
    return LC_OK;
}

int32_t lc_get_counters(lc_sim* s, int64_t* overlap_pairs, int64_t* collision_hits, int64_t* candidates) {
    SIM_ENTER(s);
    int rc = read_ctl(s);
    if (rc) return rc;
    if (overlap_pairs) *overlap_pairs = (int64_t)(s->hctl->overlap2 / 2);
    if (collision_hits) *collision_hits = (int64_t)s->hctl->coll_hits;
    if (candidates) *candidates = s->hctl->n_cand2;
    return LC_OK;
}

int32_t lc_get_resorts(lc_sim* s, int64_t* resorts) {
    SIM_ENTER(s);
    int rc = read_ctl(s);
    if (rc) return rc;
    if (resorts) *resorts = (int64_t)s->hctl->n_resorts;
    return LC_OK;
}

int32_t lc_count_pair_tests(lc_sim* s, int64_t* pair_tests) {
    SIM_ENTER(s);
    if (!pair_tests) return fail(LC_ERR_ARG, "lc_count_pair_tests: null argument");
    unsigned long long* d = (unsigned long long*)s->stage;
    CK(cudaMemsetAsync(d, 0, sizeof(unsigned long long), s->st));
    lc::launch_pair_count(s->dP, s->n, s->b, d, s->st, &s->launches);
    unsigned long long h = 0;
    CK(cudaMemcpyAsync(&h, d, sizeof h, cudaMemcpyDeviceToHost, s->st));
    CK(cudaStreamSynchronize(s->st));
    *pair_tests = (int64_t)h;
    return LC_OK;
}

int32_t lc_get_finish_log(lc_sim* s, float* log, int32_t cap, int32_t* count) {
    SIM_ENTER(s);
    int rc = read_ctl(s);
    if (rc) return rc;
    const int k = s->hctl->total_num_done;
    if (count) *count = k;
    if (log && k > 0) {
        if (k > cap) return fail(LC_ERR_CAPACITY, "lc_get_finish_log: %d finishers, the buffer holds %d", k, cap);
        CK(cudaMemcpy(log, s->b.fin_log, sizeof(float) * 3 * (size_t)k, cudaMemcpyDeviceToHost));
    }
    return LC_OK;
}

int32_t lc_get_walled(lc_sim* s, int32_t* n_plan, int32_t* lost_frames, int64_t* replans, int64_t* wall_hits) {
    SIM_ENTER(s);
    if (s->hp.n_walls == 0) return fail(LC_ERR_ARG, "lc_get_walled: not a walled scene");
    int rc = read_ctl(s);
    if (rc) return rc;
    const size_t n = (size_t)s->n_ids;
    std::vector<int> tmp(n);
    for (int t = 0; t < 2; ++t) {
        int32_t* dst = t == 0 ? n_plan : lost_frames;
        if (!dst) continue;
        CK(cudaMemcpy(tmp.data(), t == 0 ? s->b.w.n_plan : s->b.w.lost, sizeof(int) * n, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < n; ++i) dst[s->hp_ext(i)] = tmp[i];
    }
    if (replans) *replans = s->replans;
    if (wall_hits) *wall_hits = (int64_t)s->hctl->wall_hits;
    return LC_OK;
}

int32_t lc_get_plan(lc_sim* s, int32_t agent, float* waypoints, int32_t cap, int32_t* n_waypoints) {
    SIM_ENTER(s);
    if (s->hp.n_walls == 0 || agent < 0 || agent >= s->n_ids || !n_waypoints) return fail(LC_ERR_ARG, "lc_get_plan: bad argument");
    int id = agent;
    if (!s->h_order.empty()) { id = -1; for (int i = 0; i < s->n_ids; ++i) if (s->h_order[i] == agent) { id = i; break; } }
    CK(cudaStreamSynchronize(s->st));
    const int k = s->h_nplan[id];
    *n_waypoints = k;
    if (k > cap) return fail(LC_ERR_CAPACITY, "lc_get_plan: %d waypoints, the buffer holds %d", k, cap);
    if (waypoints && k > 0) CK(cudaMemcpy(waypoints, s->b.w.plan + (size_t)id * lc::WP_CAP, sizeof(float) * 2 * (size_t)k, cudaMemcpyDeviceToHost));
    return LC_OK;
}

int32_t lc_get_trace(lc_sim* s, uint64_t* ts) {
    SIM_ENTER(s);
    if (!ts) return fail(LC_ERR_ARG, "lc_get_trace: null argument");
    int rc = read_ctl(s);
    if (rc) return rc;
    for (int k = 0; k < LC_TRACE_STAMPS; ++k) ts[k] = (uint64_t)s->hctl->ts[k];
    return LC_OK;
}

}  // extern "C"

// splat.cu: the tables lc_render draws from, in engine-id order on the simulation's stream (staging buffer)
int lc_render_state(lc_sim* s, float2** pos, float** c3, int* n, const lc::WallRect** walls, int* n_walls, cudaStream_t* st) {
    SIM_ENTER(s);
    if (s->dd_world > 1) return fail(LC_ERR_ARG, "lc_render: single-GPU crowds only");
    const size_t m = (size_t)s->n_ids;
    float2* dpos = (float2*)s->stage;
    float* dc = s->stage + 10 * m;
    // k_get_state writes by the caller's agent index; the picture only needs SOME fixed order, and "later in the table wins"
    // is defined on that index
    lc::launch_get_state(s->dP, s->n, s->b, dpos, nullptr, nullptr, dc, nullptr, nullptr, nullptr, nullptr, s->st, &s->launches, s->owned_dev);
    *pos = dpos; *c3 = dc; *n = s->n_ids; *walls = s->b.w.walls; *n_walls = s->hp.n_walls; *st = s->st;
    s->launches += 3;
    return LC_OK;
}

extern "C" {

int32_t lc_timer_start(lc_sim* s) {
    SIM_ENTER(s);
    CK(cudaEventRecord(s->ev0, s->st));
    return LC_OK;
}

int32_t lc_timer_stop(lc_sim* s, float* ms) {
    SIM_ENTER(s);
    CK(cudaEventRecord(s->ev1, s->st));
    CK(cudaEventSynchronize(s->ev1));
    CK(cudaEventElapsedTime(ms, s->ev0, s->ev1));
    return LC_OK;
}

int64_t lc_kernel_launches(lc_sim* s) { return s ? s->launches : 0; }

int32_t lc_step_profiled(lc_sim* s, int32_t nframes, float* ms_per_kernel) {
    SIM_ENTER(s);
    if (nframes < 0 || !ms_per_kernel) return fail(LC_ERR_ARG, "lc_step_profiled: bad argument");
    if (s->hp.n_walls > 0) return fail(LC_ERR_ARG, "lc_step_profiled: wall-free crowds only (walled frames pause for host replans)");
    constexpr int K = lc::FRAME_KERNELS;
    if (nframes > 0) ensure_packed(s);
    cudaEvent_t ev[K + 1];
    for (auto& e : ev) CK(cudaEventCreate(&e));
    double acc[K] = {0};
    for (int f = 0; f < nframes; ++f) {
        lc::launch_frame(s->dP, s->n, s->b, s->st, &s->launches, ev);
        CK(cudaEventSynchronize(ev[K]));
        for (int k = 0; k < K; ++k) {
            float ms = 0.f;
            CK(cudaEventElapsedTime(&ms, ev[k], ev[k + 1]));
            acc[k] += ms;
        }
    }
    for (int k = 0; k < K; ++k) ms_per_kernel[k] = nframes ? (float)(acc[k] / nframes) : 0.f;
    for (auto& e : ev) cudaEventDestroy(e);
    CK(cudaGetLastError());
    return LC_OK;
}


// ---- spatial domain decomposition: wiring the ranks' mailboxes together ----
int32_t lc_dd_mailbox(lc_sim* s, void** dev_ptr, uint64_t* bytes) {
    SIM_ENTER(s);
    if (!s->mailbox) return fail(LC_ERR_ARG, "lc_dd_mailbox: not a domain-decomposed crowd");
    if (dev_ptr) *dev_ptr = s->mailbox;
    if (bytes) *bytes = (uint64_t)s->lay.total;
    return LC_OK;
}

int32_t lc_dd_ipc_handle(lc_sim* s, void* handle64) {
    SIM_ENTER(s);
    if (!s->mailbox || !handle64) return fail(LC_ERR_ARG, "lc_dd_ipc_handle: bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, s->mailbox));
    memcpy(handle64, &h, 64);
    return LC_OK;
}

static int dd_set_peer(lc_sim* s, int peer_rank, void* ptr) {
    if (peer_rank < 0 || peer_rank >= s->dd_world || peer_rank == s->dd_rank) return fail(LC_ERR_ARG, "lc_dd_connect: bad peer rank %d", peer_rank);
    CK(cudaStreamSynchronize(s->st));
    CK(cudaMemcpy((char*)s->ddev + offsetof(lc::DomainDev, mbox) + sizeof(char*) * peer_rank, &ptr, sizeof(char*), cudaMemcpyHostToDevice));
    return LC_OK;
}

int32_t lc_dd_connect(lc_sim* s, int32_t peer_rank, void* peer_mailbox, int32_t peer_device) {
    SIM_ENTER(s);
    if (!s->mailbox || !peer_mailbox) return fail(LC_ERR_ARG, "lc_dd_connect: bad argument");
    if (peer_device != s->device) {
        int can = 0;
        CK(cudaDeviceCanAccessPeer(&can, s->device, peer_device));
        if (!can) return fail(LC_ERR_CUDA, "lc_dd_connect: device %d cannot access device %d (no NVLink / P2P path)", s->device, peer_device);
        cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else if (e != cudaSuccess) return fail(LC_ERR_CUDA, "cudaDeviceEnablePeerAccess failed: %s", cudaGetErrorString(e));
    }
    return dd_set_peer(s, peer_rank, peer_mailbox);
}

int32_t lc_dd_connect_ipc(lc_sim* s, int32_t peer_rank, const void* handle64) {
    SIM_ENTER(s);
    if (!s->mailbox || !handle64) return fail(LC_ERR_ARG, "lc_dd_connect_ipc: bad argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void* ptr = nullptr;
    CK(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    s->ipc_opened.push_back(ptr);
    return dd_set_peer(s, peer_rank, ptr);
}

int32_t lc_dd_get_owned(lc_sim* s, int32_t* owned) {
    SIM_ENTER(s);
    if (!s->owned_dev || !owned) return fail(LC_ERR_ARG, "lc_dd_get_owned: bad argument");
    const size_t n = (size_t)s->n_ids;
    CK(cudaMemsetAsync(s->owned_dev, 0, sizeof(int) * n, s->st));
    lc::launch_get_state(s->dP, s->n, s->b, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, s->st,
                         &s->launches, s->owned_dev);
    CK(cudaMemcpyAsync(owned, s->owned_dev, sizeof(int) * n, cudaMemcpyDeviceToHost, s->st));
    CK(cudaStreamSynchronize(s->st));
    return LC_OK;
}

int32_t lc_dd_get_shard(lc_sim* s, int32_t* ids, float* pos, float* vel, int32_t cap, int32_t* count) {
    SIM_ENTER(s);
    if (s->dd_world <= 1 || !ids || !count) return fail(LC_ERR_ARG, "lc_dd_get_shard: bad argument");
    // staging: [count (16 B) | pos (2 cap) | vel (2 cap) | ids (cap)]; the rank never holds more than s->n agents
    const size_t m = (size_t)s->n;
    int* dcount = (int*)s->stage;
    float2* dpos = (float2*)(dcount + 4);   // 16-byte header, then the float2 tables, ids last: aligned for odd m
    float2* dvel = dpos + m;
    int* dids = (int*)(dvel + m);
    if ((4 + 5 * m) > s->stage_floats) return fail(LC_ERR_CAPACITY, "lc_dd_get_shard: staging too small");
    CK(cudaMemsetAsync(dcount, 0, sizeof(int), s->st));
    lc::launch_get_owned(s->dP, s->n, s->b, dids, dpos, dvel, dcount, s->st, &s->launches);
    CK(cudaMemcpyAsync(count, dcount, sizeof(int), cudaMemcpyDeviceToHost, s->st));
    CK(cudaStreamSynchronize(s->st));
    const size_t k = (size_t)*count;
    if ((int32_t)k > cap) return fail(LC_ERR_CAPACITY, "lc_dd_get_shard: %d owned agents, buffers hold %d", (int)k, cap);
    CK(cudaMemcpyAsync(ids, dids, sizeof(int) * k, cudaMemcpyDeviceToHost, s->st));
    if (pos) CK(cudaMemcpyAsync(pos, dpos, sizeof(float2) * k, cudaMemcpyDeviceToHost, s->st));
    if (vel) CK(cudaMemcpyAsync(vel, dvel, sizeof(float2) * k, cudaMemcpyDeviceToHost, s->st));
    return check_device_error(s);
}

int32_t lc_dd_get_partials(lc_sim* s, uint64_t* lastvel, int32_t* nonfinite, float* avg_velocity_before) {
    SIM_ENTER(s);
    if (s->dd_world <= 1) return fail(LC_ERR_ARG, "lc_dd_get_partials: not a domain-decomposed crowd");
    lc::launch_finalize(s->dP, s->b, s->st, &s->launches);
    int rc = read_ctl(s);
    if (rc) return rc;
    if (lastvel) *lastvel = (uint64_t)s->hctl->lastvel;
    if (nonfinite) *nonfinite = s->hctl->nonfinite;
    if (avg_velocity_before) *avg_velocity_before = s->hctl->avg_velocity;
    return LC_OK;
}

int32_t lc_dd_counts(lc_sim* s, int32_t* n_local, int32_t* capacity) {
    SIM_ENTER(s);
    lc::Params hp;
    CK(cudaMemcpyAsync(&hp, s->dP, sizeof hp, cudaMemcpyDeviceToHost, s->st));
    CK(cudaStreamSynchronize(s->st));
    if (n_local) *n_local = hp.n;
    if (capacity) *capacity = s->n;
    return LC_OK;
}

}  // extern "C"
