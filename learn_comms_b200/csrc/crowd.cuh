// Single-crowd engine: data layout shared by crowd.cu (kernels) and capi.cu (C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace lc {

constexpr int HULL_K = 16;        // directional extremes used to prune comm-partner candidates (stage 1, all agents)
constexpr int HULL_K2 = 256;      // stage 2, over the stage-1 survivors only: UP TO this many directions (refine_body picks 64 / 128 / 256
                                  // by the length of the stage-1 list; Params::hull_k2 / env LC_HULL_K2 force a value)
constexpr int CAND_SMEM = 128;     // per-block candidate list of k_force (stage 3); falls back to the global list beyond
constexpr int REFINE_MAX = 1 << 16; // stage 2 is skipped when stage 1 leaves more than this many candidates
constexpr int PAIR_CAP = 32;      // per-agent list of pair-filter survivors (k_pairs -> k_force); beyond it k_force scans the whole disc
constexpr int MAX_HITS = 48;      // per-thread buffer of interacting neighbours (sorted by id before summation)
constexpr int MAX_NEAR = 8;       // per-agent list of neighbours close enough to collide this frame
constexpr int CELL_CAP = 1 << 20; // max grid cells (cell sides grow if exceeded)
#ifndef LC_CELL_XSUB
#define LC_CELL_XSUB 4
#endif
constexpr int CELL_XSUB = LC_CELL_XSUB;  // cells per cell height along x (grid_from_reductions)
constexpr int SCAN_TILE = 4096;   // cells scanned per block of k_scan_hull
constexpr int SCAN_TILES = CELL_CAP / SCAN_TILE + 1;  // tiles cover indices [0, ncells] inclusive
constexpr int CELL_ALLOC = SCAN_TILES * SCAN_TILE;    // allocation of cell_count / cell_start
constexpr int EV_SORT_CAP = 2048; // events sorted in shared memory; above this the per-id path is used
constexpr float CAND_MARGIN = 0.1f; // displacement per frame the early candidate list allows for; a frame that moved somebody further re-sorts

// One copy of the per-agent state, structure of arrays.  All arrays have n elements.
//   pv     (pos.x, pos.z, vel.x, vel.z)          Dynamics.h:4,6   — integrated, not yet pushed out (k_force -> k_collide)
//   vfh    vel_forehalf.xz                       Dynamics.h:7
//   goal   final_goal                            Agent.h:11
//   lgoal  local_goal                            Agent.h:12
//   facing CommComp::facing                      CommComp.h:17
//   c4     (c[0], c[1], c[2], goal_dist)         CommComp.h:14, Agent.h:19
//   meta   (id, flags, rng, min_time bits)       Agent.h:14,17-18; flags bit0 = num_done, bit1 = overhead_counted
struct AgentArrays {
    float4* pvf;    // (pos.x, pos.z, vel.x, vel.z) at the END of the frame that wrote the set (after the push-out) = the
                    // frame-start state of the next frame: what all_sync publishes (Pool.cpp:69-91)
    float4* pq;     // [2 * ceil(n/2)] pair-interleaved copy of pvf: (x0,x1,z0,z1), (vx0,vx1,vz0,vz1) per slot pair
    float4* pv;
    float2* vfh;
    float2* goal;
    float2* lgoal;
    float2* facing;
    float4* c4;
    int4* meta;
};

// per-block partial reductions of k_collide / k_init_reduce (128-thread blocks), folded by the next k_frame_begin
struct BlockRed {
    uint32_t minx, miny, maxx, maxy, vmax, dmax;   // dmax: float bits of the largest displacement of the frame (>= 0)
    unsigned long long lastvel;
};

enum { TS_BEGIN = 0, TS_BIN, TS_SCAN, TS_REORDER, TS_REFINE, TS_PAIRS, TS_FORCE, TS_COLLIDE, TS_NBR_WAIT0, TS_NBR_WAIT1,
       TS_ALL_WAIT0, TS_ALL_WAIT1, TS_REFINE_END, TS_BEGIN_END, TS_COUNT = 16 };
struct Ctl {
    // frame / latch (main.cpp:98-150)
    uint32_t frame;
    int finished;   // demo::run returned all_done in some frame <= frame
    int halt;       // frames after the all_done frame are no-ops
    int error;      // lc_status of the first device-side error
    // uniform grid: geometry of the last re-sort, cutoff of the current frame
    float x0, y0, h, hx, cutoff;   // cell height (y) and width (x)
    int nx, ny, ncells, stencil, stencil_x;
    // Skin: the agents are re-sorted into cells only every Params::rebin_every frames.  In between, slot order and
    // cell_start stay as they are; a search from an agent's CURRENT position reaches cutoff + skin, where skin bounds how
    // far anybody has moved since the re-sort, so the candidate set is still a superset and the exact predicate decides.
    int in_set, src_set, live;   // Buffers::set roles of the frame in flight (crowd.cuh Buffers)
    int rebin;         // this frame re-sorts (k_bin / k_scan_hull / scatter in k_reorder); else the slots stay and the copies swap roles
    int rebin_next;    // decomposed crowds: the NEXT frame re-sorts — the same decision on every rank (dd_global_body), known before
                       // this frame's k_collide chooses between routing records (dd_pack_body) and refreshing ghosts (dd_push_body)
    int since;         // frames since the last re-sort
    int force_rebin;   // the state was replaced (lc_create / lc_set_state)
    float skin;        // sum of the per-frame maximum displacements since the last re-sort, at this frame's start
    uint32_t dmax;     // largest displacement of the frame in flight (float bits; consumed by the next frame_begin)
    float dmax_last;   // decomposed crowds: largest displacement of the last executed frame, over this rank and its neighbours
    // comm-partner candidate pruning
    float cx, cy, delta;
    unsigned long long ext[HULL_K];
    float ext_val[HULL_K];   // last frame's extreme projection per direction (k_scan_hull) -> gate of the next k_bin
    float gate_rho2;
    int n_edges;
    float ex[HULL_K], ey[HULL_K], enx[HULL_K], eny[HULL_K];
    int n_cand;
    int n_cand2, cand_sel;   // list k_force scans: cand_sel ? cand2 : cand, n_cand2 entries
    // Between re-sorts the stage-1 list of a frame is compiled by the tail of the PREVIOUS frame's k_collide (it holds the
    // final positions in registers; a pass of its own over all agents cost 11 us at 1 M): entries wait here until the next
    // k_frame_begin, which adopts them unless it decides to re-sort after all (then k_reorder compiles the list).
    int n_cand_next;
    float ibx0, iby0, ibx1, iby1;   // axis-aligned box well inside the stage-1 polygon (k_scan_hull): agents in it skip the edge tests
    // reductions over the final positions / velocities of the frame (consumed by the next frame_begin)
    uint32_t minx, miny, maxx, maxy, vmax;
    // scoring: demo.cpp:650-702, physics.cpp:72-76
    int total_num_done;
    float run_avg, M2, run_std, avg_velocity;
    unsigned long long lastvel;  // ((id + 1) << 32) | speed bits of the highest-id not-done agent this frame
    int nonfinite;               // some earlier not-done agent had a non-finite speed (sticky NaN of physics.cpp:75)
    int n_not_done, nnd_acc;
    int n_blkred;                // BlockRed entries waiting to be folded
    int ev_count;
    int replan1, replan2;        // ai.cpp:151-168: first two agents (table order) that need a replan
    int replan1n, replan2n;      // ... of the coming frame, collected by k_collide (k_init_reduce after lc_create / lc_set_state)
    // counters (diagnostics)
    unsigned long long overlap2, coll_hits;
    unsigned long long n_resorts;   // frames that re-sorted the grid (lc_get_resorts)
    int near_overflow, hit_overflow;
    // walled scenes (walls.cuh): replans are served by the host planner between frames.  The tail of k_collide evaluates
    // the gate of ai.cpp:152-168 for the NEXT frame and leaves the two lowest requesting ids here; k_frame_begin then
    // pauses the graph replay (halt = 2, frame not advanced) until the host has uploaded the new waypoint lists.
    int wr1, wr2;                // requesting ids (0x7fffffff = none)
    int w_served;                // the host has served wr1 / wr2 for the coming frame
    int n_planned;               // agents with a non-empty plan = minstd_rand0 draws per frame (LMP.cpp:259-263)
    int drawn_last;              // draws the last executed frame consumed
    unsigned long long draws;    // draws consumed BEFORE the last executed frame (its agent k draws element draws + k + 1)
    unsigned long long wall_hits, replans;
    // device-clock timeline of the last executed frame (globaltimer, ns): thread 0 of block 0 of every frame kernel stamps its
    // entry; the two peer waits of a decomposed crowd stamp begin and end.  lc_get_trace reads it (bench.py's timeline).
    unsigned long long ts[TS_COUNT];
};

constexpr int DD_MAX_WORLD = 8;     // ranks of one spatially decomposed crowd (one B200 box)
// Per-rank, per-frame capacities of the all-rank exchange; beyond them the frame raises LC_ERR_CAPACITY (-5) on every rank
// (tests/test_gpu_domain.py drives both).  4x the round-1 values: 8 ranks x 16384 x 32 B = 4 MB of mailbox per parity.
constexpr int DD_EV_CAP = 16384;     // done-events one rank can forward to the scoring rank per frame
constexpr int DD_CAND_CAP = 16384;   // comm-partner candidates one rank can publish per frame
constexpr int META_GHOST = 4;        // meta.y bit2: halo copy of an agent owned by a neighbouring rank

struct Params {
    int n;          // agents in the local (cell-sorted) arrays this frame; in domain mode (dd_world > 1) rewritten on the device every frame
    int n_in;       // input records of this frame's counting sort (k_bin / k_reorder): == n for a single-GPU crowd; domain mode:
                    // last frame's local slots (stayers are read in place, leavers and old ghosts are skipped) + the records the
                    // neighbours sent
    int n_ids;      // size of the global id space (id-indexed tables, LCG stride); == n for a single-GPU crowd
    int cap;        // capacity of the slot-indexed arrays (launch width); == n for a single-GPU crowd
    // spatial domain decomposition in x strips: rank r owns xb[r] <= pos.x < xb[r+1]
    int dd_rank, dd_world;
    float dd_xb[DD_MAX_WORLD + 1];
    float dd_halo;  // ghost band width: interaction cutoff at the speed bound + collision reach (domain.cuh)
    float r, R, RR, near_c, near_reach;
    float max_duration;
    uint32_t lcg_stride;  // 16807^n mod (2^31-1): advances an agent's private stream by one frame
    uint32_t flags;
    int stencil_req;
    int hull_k2;          // directions of the second pruning stage (0 = by the length of the stage-1 list)
    int rebin_every;      // re-sort the grid every this many frames (1 = every frame: decomposed crowds, whose local set changes every frame)
    int n_walls, n_bvh;   // static walls (0 = wall-free: the kernels' WALLS = false instantiations)
    uint32_t rng0;        // walled: stream state before frame 1's first draw (draw k of the run = rng0 * 16807^k)
    float m_cv[24];
    float m_f[6];
};

constexpr int WP_CAP = 32;        // waypoints per agent in walled scenes (the catalogue's plans have <= ~10)
constexpr int WALL_SMEM = 128;    // walls staged in shared memory per block; beyond that they are read from global
struct WallRect { float ox, oy, w, h; };                      // == lcw::Wall
struct BvhNode { float ox, oy, w, h; int left, right; };     // == lc_bvh_node
// Walled-scene state, indexed by ENGINE ID (not slot: nothing here is reordered with the grid)
struct WalledArrays {
    const WallRect* walls;   // [n_walls] creation order
    const int* wall_order;   // [n_walls] statics order (LMP.cpp:307 loop)
    const BvhNode* bvh;      // [n_bvh] static BVH, pre-order (lc_static_bvh)
    float2* plan;            // [n_ids][WP_CAP]  Agent::plan (Agent.h:14)
    int* n_plan;             // [n_ids]  0 = no plan (unreachable goal)
    int* num_done;           // [n_ids]  Agent::num_done (Agent.h:13)
    int* lost;               // [n_ids]  Agent::lost_frames after the last executed update_agents (Agent.h:28)
    int* lost_next;          // [n_ids]  ... as the coming frame's gate will leave it (0 for an agent the host replanned)
    int* noplan_before;      // [n_ids]  plan-less agents with a smaller id (they do not draw, LMP.cpp:267-270)
    float2* req_pos;         // [n_ids]  position a requesting agent replans from (GMP.cpp:96: a.start)
};

struct Buffers {
    // Two copies of the per-agent state.  A frame reads one (frame-start state: Ctl::in_set) and writes the other, slot for
    // slot, so between grid re-sorts the roles simply alternate and nothing is copied; a re-sort frame first scatters the
    // last frame's output (Ctl::src_set) into the other copy and reads that.  Ctl::live = the copy holding the state after
    // the last executed frame (what lc_get_state reads).
    AgentArrays set[2];
    int2* cellrank;    // (cell, rank within cell) per agent, cur order
    int* slot_cell;    // [n] by slot: the cell the agent was sorted into at the last re-sort
    int* cell_count;   // [CELL_ALLOC]
    int* cell_start;   // [CELL_ALLOC]
    unsigned long long* tile_agg; // [SCAN_TILES] (frame << 32) | tile total, published by k_scan_hull
    int* slot_of_id;   // inverse permutation
    float4* cand;      // comm-partner candidates: (x, z, id bits, slot bits)
    BlockRed* blkred;  // [ceil(n / 128)]
    unsigned long long* blkext; // [ceil(n / 1024)][HULL_K] per-block directional extremes of k_bin
    float4* cand2;     // [REFINE_MAX] after the second pruning stage
    int* near_cnt;     // [n]
    int* near_idx;     // [n][MAX_NEAR]
    int* pair_cnt;     // [n] by slot: filter survivors recorded for the agent this frame (may exceed PAIR_CAP)
    int* pair_list;    // [PAIR_CAP][n] by slot: partner slots, entry k of all agents contiguous
    int2* events;      // (id, min_time bits) of agents that became done this frame
    float* ev_by_id;   // [n] per-id event path (NaN = none)
    int* partner_id;   // [n] by id (diagnostic; filled from diag_partner by k_get_diag when the host asks)
    int* nbr_cnt;      // [n] by id (diagnostic; filled from diag_cnt by k_get_diag)
    int* diag_partner; // [cap] by slot: comm partner id of the last frame (k_force)
    int* diag_cnt;     // [cap] by slot: interacting neighbours of the last frame (k_force)
    int* nbr_list;     // [n][list_cap] by id (LC_FLAG_DEBUG_LISTS)
    int list_cap;
    // Table order (quirk Q4): the engine's agent ids are POSITIONS in the reference's table iteration order, so that
    // "ascending id" everywhere in the kernels is the reference's loop order.  ext_of[id] = the caller's agent index;
    // nullptr = identity.  Only the host-facing kernels (get / set state, diagnostics, shards) translate.
    const int* ext_of;
    WalledArrays w;    // valid when Params::n_walls > 0
    float* fin_log;    // [n_ids][3] (overhead, run_avg, run_std) per counted finisher, in Welford order (demo.cpp:655-663)
    Ctl* ctl;
    struct DomainDev* dd;  // device-side exchange state (domain.cuh); nullptr for a single-GPU crowd
};

// ---- spatial domain decomposition over NVLink peer memory (domain.cuh) ----
// Agent record on the wire: 5 x 16 B planes (posf.xy vel.xy | vfh goal | lgoal facing | c4 | meta)
constexpr int DD_PLANES = 5;
// Frame start needs the two ADJACENT ranks only: how many records they sent, and bounds (max speed, y extent) of the
// agents those records can be — the grid of a rank is a matter of the records it holds.  One per (neighbour, parity).
struct NeighbourBlob {
    int n_sent;                   // records the sender put into this rank's region
    uint32_t vmax;                // max speed bits over the sender's local agents
    uint32_t miny, maxy;          // order-encoded y extent of the sender's local agents
    int error;
    uint32_t dmax;                // largest displacement bits of the sender's last frame (the skin covers the ghosts' true movement)
    int pad[2];
};
// Everything that is global (scoring, latch, replan gate, comm-partner candidates) travels later in the frame, in the
// side-stream kernel beside k_pairs.  One per (sender, parity), written into EVERY peer's mailbox.
struct DomainBlob {
    int nnd;                      // owned agents that are not done (demo.cpp:667-669)
    int nonfinite;
    unsigned long long lastvel;   // physics.cpp:72-76: ((id + 1) << 32) | speed bits of the highest-id owned not-done agent
    int replan[2];                // two lowest owned ids that need a replan (ai.cpp:151-168)
    int n_events;                 // done-events forwarded to the scoring rank (rank 0)
    int n_cand;                   // comm-partner candidates published this frame
    int error;
    int want_rebin;               // this rank's skin has used up half of what the ghost band allows for: re-sort next frame
};
static_assert(sizeof(DomainBlob) % 8 == 0 && sizeof(NeighbourBlob) % 8 == 0, "blobs are copied word-wise");
struct DomainLayout {             // byte offsets inside one rank's mailbox allocation; identical on every rank
    size_t nblob[2], nflag[2], blob[2], cflag[2], region[2][3], events[2], cand[2], ev_by_id, total;   // region[.][1] is unused
    int cap_self, cap_halo;
};
struct DomainDev {
    char* mbox[DD_MAX_WORLD];     // mailbox base of every rank (own allocation + peer mappings)
    DomainLayout lay;
    int count[3];                 // this frame's pack: records sent to rank-1, agents that stay, records sent to rank+1
    int recv[3];                  // this frame's input: records from the right, last frame's local slots, records from the left
    int replan[2];
    unsigned char* keep;          // [cap] by last frame's slot: 0 = not part of the next frame's local set, 1 = stays as owned agent,
                                  // 2 = stays as ghost of its new owner; 1 / 2 are read in place from `cur` by the next frame's sort
    // Between re-sorts the local set is frozen: every rank keeps its owned agents and its ghosts in their slots, and the tail
    // of k_collide only REFRESHES the ghosts (dd_push_body): the record of an owned boundary agent goes to the same index of
    // the neighbour's region every frame, and the neighbour copies it over its ghost's slot (k_reorder, looked up by id).
    int2* send_idx;               // [cap] by slot: record index in the left (x) / right (y) neighbour's region, -1 = not its ghost
    int rcount[3];                // records of this rank's refresh lists (left, -, right), built by the re-sort frame's k_reorder
    float skin_allow;             // movement since the last re-sort the ghost band was widened for (lc_dd_create)
    unsigned long long spin_timeout_ns;
};
DomainLayout domain_layout(int world, int cap_self, int cap_halo, int n_ids);

void upload_dirs();
constexpr int FRAME_KERNELS = 8;  // kernels per frame, in launch order: frame_begin, bin, scan_hull, reorder, refine, pairs, force, collide
constexpr int DD_KERNELS = 0;     // domain mode: dd_sync_begin replaces frame_begin, the pack rides in k_collide, the candidate exchange in k_refine
void launch_get_diag(const Params* dP, int n, const Buffers& b, cudaStream_t st, long long* launches);
// dst[ext_of[i] * width + k] = (values ? ext_of[src[i * width + k]] (negative entries kept) : src[...]) for i < rows
void launch_to_external(const Buffers& b, const int* src, int* dst, int rows, int width, int values, cudaStream_t st, long long* launches);
void launch_dd_pack(const Params* dP, int n, const Buffers& b, cudaStream_t st, long long* launches);  // once, before the first frame
// Side stream of a frame: k_refine (one block: second pruning stage and, in domain mode, the candidate exchange with the
// peers) does not feed k_pairs, so the two run side by side (fork after k_reorder, join before k_force).
struct SideStream { cudaStream_t st; cudaEvent_t fork, join; };
void launch_frame(Params* dP, int n, const Buffers& b, cudaStream_t st, long long* launches,
                  cudaEvent_t* ev = nullptr, const SideStream* side = nullptr, bool walled = false);
void launch_pair_count(const Params* dP, int n, const Buffers& b, unsigned long long* out, cudaStream_t st, long long* launches);
void launch_init_reduce(const Params* dP, int n, const Buffers& b, cudaStream_t st, long long* launches);
void launch_walled_gate(const Params* dP, int n, const Buffers& b, cudaStream_t st, long long* launches);  // gate of the coming frame
void launch_finalize(const Params* dP, const Buffers& b, cudaStream_t st, long long* launches);
void launch_get_state(const Params* dP, int n, const Buffers& b, float2* pos, float2* vel, float2* vfh, float* c3,
                      float2* facing, float2* lgoal, float* goal_dist, int* num_done, cudaStream_t st,
                      long long* launches, int* owned = nullptr);
void launch_set_state(const Params* dP, int n, const Buffers& b, const float2* pos, const float2* vel,
                      const float2* vfh, const float* c3, const float2* facing, const float2* lgoal,
                      const float* goal_dist, const int* num_done, const float2* goal, const float* min_time, int init,
                      cudaStream_t st, long long* launches, const int* ids = nullptr);
void launch_get_owned(const Params* dP, int n, const Buffers& b, int* ids, float2* pos, float2* vel, int* count,
                      cudaStream_t st, long long* launches);
void launch_set_rng(const Params* dP, int n, const Buffers& b, uint32_t base, cudaStream_t st, long long* launches);

}  // namespace lc
