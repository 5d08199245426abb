/* learn_comms.h — C ABI of the B200-native learn-comms hot path (liblearncomms.so).
 *
 * The reference (daltonhildreth/learn-comms) has no FFI: the path is reached through the process-global
 * system functions that enginee/src/main.cpp:98-150 calls every frame.  Each entry point below names the
 * reference interface it replaces.  Plain pointers and sizes only; no exceptions cross this boundary;
 * every function returns 0 on success or a negative lc_status, with lc_last_error() giving the text.
 * A handle is confined to one host thread at a time; distinct handles may be used concurrently.
 * There is no CPU fallback: without a CUDA device every compute entry point fails with LC_ERR_CUDA.
 *
 * Array conventions: agent arrays are in agent-id (creation) order; vec2 arrays are [n][2] = (x, z).
 */
#ifndef LEARN_COMMS_H
#define LEARN_COMMS_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    LC_OK = 0,
    LC_ERR_ARG = -1,      /* bad argument (reference: assert / exit(1), io.cpp:42-45) */
    LC_ERR_CUDA = -2,     /* CUDA runtime error / no device */
    LC_ERR_STATE = -3,    /* simulation reached a state the reference treats as UB (e.g. no comm partner, comm.cpp:273-285) */
    LC_ERR_IO = -4,       /* config / result file error (comm.cpp:78,92) */
    LC_ERR_CAPACITY = -5  /* an internal fixed-capacity buffer overflowed */
} lc_status;

/* flags for lc_desc.flags */
#define LC_FLAG_NO_COMM 1u        /* nocomm_* executables: comm::run is compiled out (main.cpp:130-132) */
#define LC_FLAG_FREE_RUN 2u       /* keep stepping the dynamics after demo::run reported all_done (fixed-step benchmarks) */
#define LC_FLAG_DEBUG_LISTS 4u    /* record per-agent neighbour lists {j : ttc_ij <= 5} each frame (parity tests) */

/* feature-set variants of comm::x (the #defines at comm.cpp:22-28; HEAD builds with COSINE_REL + PROX + SOFTSIGN_GOAL = no bit set) */
#define LC_FLAG_FEAT_PLAIN 32u        /* neither COSINE_REL nor NORM_REL: x = [c_j, rel_v.x, rel_v.y, rel_p.x, rel_p.y, g] (comm.cpp:130,189).
                                       * M_cv columns in that order.  HEAD cannot be built this way (its parser does not compile without
                                       * COSINE_REL): this is the feature set of the legacy "3 5" config files, see lc_read_config_legacy */
#define LC_FLAG_FEAT_NO_PROX 64u      /* PROX undefined (comm.cpp:172-174) */
#define LC_FLAG_FEAT_NO_SOFTSIGN 128u /* SOFTSIGN_GOAL undefined (comm.cpp:178-180) */

typedef struct lc_rect { float ox, oy, w, h; } lc_rect; /* BoundVolume.h Rect: origin (centre), width, height */

/* Scene description: replaces demo::init + ai::init + physics::init/prewarm (demo.cpp:611-648, ai.cpp:24-129,
 * physics.cpp:13-21).  All pointers are HOST pointers, copied during the call. */
typedef struct lc_desc {
    int32_t n;              /* robots, >= 2 (comm.cpp:273-285: n == 1 is UB in the reference) */
    const float* pos;       /* [n][2] initial positions   (Scene::pos_of, demo.cpp:630-632) */
    const float* goal;      /* [n][2] final goals         (Scene::goal_of, demo.cpp:572) */
    float radius;           /* robot radius, reference 0.15 (demo.cpp:540,552) */
    float max_duration;     /* Scene::max_duration, seconds of simulated time */
    uint32_t rng_state;     /* minstd_rand0 state before the first per-frame draw (Seeder.cpp:15-18; LMP.cpp:261-263) */
    const float* m_cv;      /* [3][8] row-major comm matrix M_cv (comm.cpp:124-131); NULL = zeros */
    const float* m_f;       /* [2][3] row-major force matrix M_f (comm.cpp:112-121); NULL = zeros */
    uint32_t flags;
    int32_t device;         /* CUDA device ordinal */
    int32_t n_walls;        /* static walls (demo.cpp:515-537 create_wall); 0 = wall-free scene.  Walled scenes: the global
                             * planner (Cspace2d + PRM + A*, ai.cpp:41-127) runs on the host at creation and for every
                             * replan (ai.cpp:162-168), so lc_step is synchronous for them; single GPU only */
    const lc_rect* walls;   /* [n_walls] in creation order */
    int32_t stencil;        /* grid cells per interaction cutoff (0 = default 2): cell side = cutoff / stencil */
    const int32_t* table_order; /* [n] permutation: table_order[k] = index of the k-th agent the reference's component-table
                                 * loops visit (quirk Q4: std::unordered_map order, Pool.h:24, Pool.cpp:156-160).  Decides
                                 * which minstd_rand0 draw an agent gets, the order forces / hits / Welford updates are
                                 * summed in, arg-max ties and the two replan slots.  NULL = ascending index.  lc_table_order
                                 * returns the reference's own order; the comm_norender drop-in uses it. */
    const int32_t* wall_order;  /* [n_walls] order in which the statics loop (LMP.cpp:307-319) visits the walls and in which
                                 * ai::init hands them to the static BVH (ai.cpp:28-36); NULL = creation order */
} lc_desc;

typedef struct lc_sim lc_sim;

/* ---- single crowd: one simulation resident on one GPU ---- */
int32_t lc_create(const lc_desc* desc, lc_sim** out);
void lc_destroy(lc_sim* sim);
/* comm::init's parameter load (comm.cpp:80-131) from memory */
int32_t lc_set_weights(lc_sim* sim, const float* m_cv, const float* m_f);
/* nframes iterations of the frame loop main.cpp:98-150: demo::run -> comm::run -> ai::update_agents ->
 * physics::simulate -> POOL.all_sync.  Asynchronous; frames after all_done are no-ops unless LC_FLAG_FREE_RUN. */
int32_t lc_step(lc_sim* sim, int32_t nframes);
/* the whole while(!all_done) loop (main.cpp:98); returns the number of frames executed */
int32_t lc_run_to_completion(lc_sim* sim, int32_t* frames);
int32_t lc_sync(lc_sim* sim);
int32_t lc_finished(lc_sim* sim, int32_t* finished, uint32_t* frame);

/* state exchange with the host engine's component tables (Dynamics.h, Agent.h, CommComp.h); HOST pointers, any may be NULL.
 * get = what POOL.all_sync publishes (Pool.cpp:69-91); set = teacher forcing / a host engine that owns the state.
 * Teacher forcing (lc_set_state, lc_exchange_state, lc_set_rng) is for single-GPU crowds only: on a lc_dd_* handle these
 * return LC_ERR_STATE (the initial tables go to lc_dd_create; the host reads a decomposed crowd through lc_dd_get_shard). */
int32_t lc_get_state(lc_sim* sim, float* pos, float* vel, float* vel_forehalf, float* c, float* facing,
                     float* local_goal, float* goal_dist, int32_t* num_done);
int32_t lc_set_state(lc_sim* sim, const float* pos, const float* vel, const float* vel_forehalf, const float* c,
                     const float* facing, const float* local_goal, const float* goal_dist, const int32_t* num_done);
/* lc_get_state(out...) followed by lc_set_state(in...) in one call, for teacher forcing from a recorded trajectory: the `in`
 * tables are uploaded on a copy stream of their own while the frame still executing finishes and while the `out` tables
 * travel the other way, and applied once the current state has been read out.  Null tables as in the two calls.
 * Single-GPU crowds only. */
int32_t lc_exchange_state(lc_sim* sim, float* out_pos, float* out_vel, float* out_vel_forehalf, float* out_c, float* out_facing,
                          float* out_local_goal, float* out_goal_dist, int32_t* out_num_done, const float* pos, const float* vel,
                          const float* vel_forehalf, const float* c, const float* facing, const float* local_goal,
                          const float* goal_dist, const int32_t* num_done);
int32_t lc_set_rng(lc_sim* sim, uint32_t rng_state); /* Seeder::seed (Seeder.cpp:5-20); like lc_set_state, LC_ERR_STATE on a lc_dd_* handle */
int32_t lc_get_rng(lc_sim* sim, uint32_t* rng_state);

/* comm::terminate's two numbers (comm.cpp:136-142) + demo's running statistics (demo.cpp:650-663) */
int32_t lc_get_scores(lc_sim* sim, float* avg_velocity, float* group_time, float* run_avg, float* run_std,
                      int32_t* total_num_done);

/* Headless picture of the crowd as render::draw shows it (render.cpp:364-415; no graphics API: CUDA kernels into a buffer):
 * top-down orthographic window of 25 cam_dist metres (Camera.cpp:86-90, render.cpp:354), robots as discs coloured by
 * comm_vis(c) (render.cpp:87-172; vis_f = render::VisF: 0 XYZ2LAB, 1 X2RG, 2 Y2RG, 3 Z2RG, 4 XYZ2LUV) with their black outline,
 * walls as grey squares, background of render.cpp:365.  rgba: HOST buffer [height][width][4], row 0 = far side (-z). */
int32_t lc_render(lc_sim* sim, int32_t vis_f, float cam_dist, int32_t width, int32_t height, uint8_t* rgba);
/* lc_render without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments and same picture as lc_render; the per-pixel scratch lc_render allocates per call lives in static device
 * arrays of LC_RENDER_STATIC_MAX_PIXELS entries (12 MiB on every GPU that loads the library, held until exit), so
 * width * height above that limit is refused with LC_ERR_CAPACITY.  rgba may be host or device memory.  The call performs
 * no cudaMalloc / cudaFree and calls no allocating library routine.  All calls (of all handles on one GPU) share the
 * scratch: every call rewrites the part it uses before reading it, but calls that overlap on the GPU race — serialise
 * them.  The call returns after the picture has arrived in rgba. */
#define LC_RENDER_STATIC_MAX_PIXELS (1 << 20)
int32_t lc_render_static(lc_sim* sim, int32_t vis_f, float cam_dist, int32_t width, int32_t height, uint8_t* rgba);

/* demo::update_runs' per-finisher log (demo.cpp:662 prints "t:%f u:%f s:%f" for each): [count][3] = overhead, running mean,
 * running sigma, in the order the reference's Welford update saw the finishers */
int32_t lc_get_finish_log(lc_sim* sim, float* log /*[cap][3]*/, int32_t cap, int32_t* count);
/* walled scenes: Agent::plan->size() and Agent::lost_frames per agent (Agent.h:14,28), replans served so far (ai.cpp:164-168),
 * wall push-outs so far (physics.cpp:67-69); any pointer may be NULL */
int32_t lc_get_walled(lc_sim* sim, int32_t* n_plan, int32_t* lost_frames, int64_t* replans, int64_t* wall_hits);
int32_t lc_get_plan(lc_sim* sim, int32_t agent, float* waypoints /*[cap][2]*/, int32_t cap, int32_t* n_waypoints);

/* diagnostics of the last executed frame (parity tests) */
int32_t lc_get_partners(lc_sim* sim, int32_t* partner);         /* [n] select_nearest result, comm.cpp:201-220 */
int32_t lc_get_neighbor_counts(lc_sim* sim, int32_t* count);    /* [n] |{j : ttc_ij <= 5}|, LMP.cpp:297-300 */
int32_t lc_get_neighbor_lists(lc_sim* sim, int32_t* idx, int32_t cap_per_agent); /* [n][cap], ascending id; needs LC_FLAG_DEBUG_LISTS */
int32_t lc_get_counters(lc_sim* sim, int64_t* overlap_pairs, int64_t* collision_hits, int64_t* candidates);
/* how many of the frames executed so far re-sorted the uniform grid (the others kept slots, owners and ghosts: env
 * LC_REBIN_EVERY, default 8; a frame that moved somebody too far re-sorts early) */
int32_t lc_get_resorts(lc_sim* sim, int64_t* resorts);

/* pair-filter evaluations k_pairs performed in the last executed frame (recounted from that frame's grid; diagnostic for
 * the FP32 leg of the roofline: the reference evaluates n (n - 1) pairs per frame, LMP.cpp:288-304) */
int32_t lc_count_pair_tests(lc_sim* sim, int64_t* pair_tests);
/* measured FP32 throughput of the device's CUDA cores, TFLOP/s, best of `repeats`: [0] independent scalar FMUL / FADD
 * chains without contraction (the arithmetic this library is restricted to for bit parity), [1] the same as packed
 * mul.rn.f32x2 / add.rn.f32x2 (what the pair filter issues), [2] FFMA chains (the hardware's nominal peak) */
int32_t lc_fp32_peak(int32_t device, int32_t repeats, float* tflops /*[3]*/);
/* device-clock timeline of the last executed frame (ns, %globaltimer): entry of k_frame_begin, k_bin, k_scan_hull, k_reorder,
 * k_refine, k_pairs, k_force, k_collide [0..7]; a decomposed crowd also stamps its two peer waits: neighbour handshake begin /
 * end [8, 9], all-rank exchange begin / end [10, 11], end of k_refine [12], end of the frame-start kernel [13] */
#define LC_TRACE_STAMPS 16
int32_t lc_get_trace(lc_sim* sim, uint64_t* ts /*[LC_TRACE_STAMPS]*/);
/* timing on the simulation's own stream (CUDA events) */
int32_t lc_timer_start(lc_sim* sim);
int32_t lc_timer_stop(lc_sim* sim, float* milliseconds);
int64_t lc_kernel_launches(lc_sim* sim); /* kernels launched by this handle so far */
/* nframes frames launched kernel by kernel (no graph) with a CUDA event between launches; ms_per_kernel[8] receives
 * the mean device time of k_frame_begin, k_bin, k_scan_hull, k_reorder, k_refine, k_pairs, k_force, k_collide (roofline accounting) */
#define LC_FRAME_KERNELS 8
int32_t lc_step_profiled(lc_sim* sim, int32_t nframes, float* ms_per_kernel);

/* ---- one crowd across several GPUs: spatial domain decomposition in x strips, one handle (rank) per GPU ----
 * Replaces nothing in the reference (it is single-threaded, main.cpp:98-150); this is how BASELINE.json's 1M-agent
 * plaza runs on 2/4/8 B200.  Every rank is created from the SAME global scene description and keeps the agents of its
 * strip; each frame the kernels store boundary agents straight into the neighbours' mailboxes over NVLink (no host
 * round trip).  All ranks must be stepped by the same number of frames.  Results are bit-identical to lc_create's. */
typedef struct lc_domain_desc {
    int32_t rank, world;        /* 2 <= world <= 8 */
    const float* x_bounds;      /* [world + 1] strip boundaries (entries 1..world-1 are used); NULL = lc_dd_strips (local counts equalised) */
    int32_t capacity;           /* local slot capacity (owned + ghost agents); 0 = 1.5 n/world + 2 halo_capacity */
    int32_t halo_capacity;      /* records one neighbour may send per frame; 0 = max(65536, n/world/2) */
    const float* vel;           /* optional [n][2] initial vel and vel_forehalf (developed-flow benchmarks); NULL = at rest */
    const float* vel_forehalf;
} lc_domain_desc;
int32_t lc_dd_create(const lc_desc* global_desc, const lc_domain_desc* dom, lc_sim** out);
/* the strips lc_dd_create picks when x_bounds is NULL (host code, no GPU needed): cuts x_bounds[1 .. world-1] such that the
 * LOCAL agent counts — own strip plus the ghost bands of width ghost_band on its inner sides — are equal; x[i * stride] = pos.x
 * of agent i; x_bounds[0] / [world] = -/+ 3e38.  No counterpart in the reference (single-threaded). */
int32_t lc_dd_strips(int32_t n, const float* x, int32_t stride, int32_t world, float ghost_band, float* x_bounds /*[world + 1]*/);
/* wiring: every rank needs a device pointer to every other rank's mailbox.  Same process: lc_dd_mailbox + lc_dd_connect
 * (enables peer access).  One process per GPU: lc_dd_ipc_handle (64-byte cudaIpcMemHandle_t, exchanged by the host's
 * own means, e.g. torch.distributed.all_gather) + lc_dd_connect_ipc. */
int32_t lc_dd_mailbox(lc_sim* sim, void** dev_ptr, uint64_t* bytes);
int32_t lc_dd_connect(lc_sim* sim, int32_t peer_rank, void* peer_mailbox, int32_t peer_device);
int32_t lc_dd_ipc_handle(lc_sim* sim, void* handle64);
int32_t lc_dd_connect_ipc(lc_sim* sim, int32_t peer_rank, const void* handle64);
/* lc_get_state on a rank returns global-size arrays with zeros for agents it does not own; owned[id] = 1 marks its own */
int32_t lc_dd_get_owned(lc_sim* sim, int32_t* owned /*[n]*/);
/* the rank's own shard, compact: ids[k], pos[k][2], vel[k][2] for the *count agents it owns this frame (order arbitrary).
 * What a host engine that consumes the tables every frame reads instead of the global-size lc_get_state arrays. */
int32_t lc_dd_get_shard(lc_sim* sim, int32_t* ids, float* pos, float* vel, int32_t cap, int32_t* count);
/* lc_get_scores on a rank: Welford statistics are complete on rank 0 (the scoring rank); avg_velocity is the value
 * after the previous frame.  The last frame's physics.cpp:72-76 inputs are per-rank partials: the speed of the
 * highest-id not-done agent is ((lastvel >> 32) - 1 = id, low 32 bits = speed) of the rank with the largest lastvel. */
int32_t lc_dd_get_partials(lc_sim* sim, uint64_t* lastvel, int32_t* nonfinite, float* avg_velocity_before);
int32_t lc_dd_counts(lc_sim* sim, int32_t* n_local /*owned + ghosts this frame*/, int32_t* capacity);

/* ---- batched rollouts: many independent small scenarios / PSO particles in one launch ----
 * replaces pso_opt.py:236-267 simulate_set/simulate (one comm_norender process per (particle, scene)). */
typedef struct lc_batch lc_batch;
typedef struct lc_batch_desc {
    int32_t n_scenarios;
    int32_t n_agents;         /* agents per scenario (<= 256), identical for all scenarios */
    const float* pos;         /* [S][n][2] */
    const float* goal;        /* [S][n][2] */
    const float* m_cv;        /* [S][3][8] per-scenario comm matrices */
    const float* m_f;         /* [S][2][3] */
    const float* max_duration;/* [S] */
    const uint32_t* rng_state;/* [S] */
    float radius;
    uint32_t flags;
    int32_t device;
    const int32_t* table_order; /* [n_agents] as in lc_desc, shared by all scenarios; NULL = ascending index */
    int32_t n_walls;            /* one wall layout shared by all scenarios of the batch (<= 64 walls); 0 = wall-free.  Replans */
    const lc_rect* walls;       /* (ai.cpp:162-168) are served by the host planner: lc_batch_run relaunches the rollout kernel */
    const int32_t* wall_order;  /* until no scenario waits for one.  wall_order as in lc_desc */
} lc_batch_desc;
int32_t lc_batch_create(const lc_batch_desc* desc, lc_batch** out);
void lc_batch_destroy(lc_batch* b);
/* Advance every unfinished scenario by up to max_frames frames of main.cpp:98-150 (pass INT32_MAX to run each to its own
 * all_done, i.e. the whole `while (!all_done)` loop of every comm_norender process pso_opt.py would have spawned).
 * Resumable: state stays on the device between calls.  Outputs are HOST arrays [S] (any may be NULL):
 * comm::terminate's two numbers (comm.cpp:136-142), demo's total_num_done and the frames executed so far. */
int32_t lc_batch_run(lc_batch* b, int32_t max_frames, float* avg_velocity, float* group_time, int32_t* n_done,
                     int32_t* frames);
/* back to the initial state (frame 0) with new per-scenario parameters (NULL keeps the old ones): the next PSO iteration */
int32_t lc_batch_reset(lc_batch* b, const float* m_cv /*[S][24]*/, const float* m_f /*[S][6]*/);
int32_t lc_batch_finished(lc_batch* b, int32_t* n_finished);
/* component tables of all scenarios, [S][n]... in the conventions of lc_get_state; HOST pointers, any may be NULL */
int32_t lc_batch_get_state(lc_batch* b, float* pos, float* vel, float* vel_forehalf, float* c, float* facing,
                           float* local_goal, float* goal_dist, int32_t* num_done);
/* diagnostics (parity tests): partner / neighbour count [S][n] of the last executed frame; per-scenario totals over all
 * executed frames of overlapping pairs at frame start, collision-pass hits and TTC pair evaluations */
int32_t lc_batch_get_diag(lc_batch* b, int32_t* partner, int32_t* neighbor_count, int64_t* overlap_pairs,
                          int64_t* collision_hits, int64_t* pair_tests);
/* walled batches: Agent::plan->size() and lost_frames [S][n], wall push-outs per scenario [S], replans served (total) */
int32_t lc_batch_get_walled(lc_batch* b, int32_t* n_plan, int32_t* lost_frames, int64_t* wall_hits, int64_t* replans);
int32_t lc_batch_timer(lc_batch* b, float* milliseconds); /* device time of the last lc_batch_run (CUDA events) */
int64_t lc_batch_kernel_launches(lc_batch* b);

/* ---- files: comm::init / comm::terminate formats (comm.cpp:74-142, pso_opt.py:17-41) ---- */
int32_t lc_read_config(const char* path, float* m_cv /*[24]*/, float* m_f /*[6]*/);
int32_t lc_write_result(const char* path, float avg_velocity, float group_time);
/* The legacy parameter files at the reference's root (comms_good.config, comms_82.config, comms_g.config, comms_hand.config;
 * written by optimizer.py:20-27): a "3 5" header, then a 3 x 5 matrix of an older one-channel model — row 0 =
 * [M_c | M_vx M_vy M_px M_py] (plain relative velocity / position components, no goal input), rows 1-2 column 0 = M_f
 * (2 x 1); optimizer.py:66-73 only ever varies those seven entries.  HEAD's parser aborts on them (comm.cpp:80-121).
 * Embedded into HEAD's three-channel layout with the plain feature set: m_cv[3][8] / m_f[2][3] for LC_FLAG_FEAT_PLAIN
 * (channel 0 carries the model, channels 1-2 and the goal input have zero weights).  *flags receives LC_FLAG_FEAT_PLAIN. */
int32_t lc_read_config_legacy(const char* path, float* m_cv /*[24]*/, float* m_f /*[6]*/, uint32_t* flags);

/* ---- iteration order of the reference's component tables (quirk Q4; Pool.h:24, Pool.cpp:7-16,156-160) for a scene
 * with n_agents robots followed by n_walls walls (demo.cpp:628-635): agent_order[n_agents], wall_order[n_walls] (either may
 * be NULL when its count is 0).  Ascending when the scene exceeds the reference's 65 535 entities (Entity.h:12). ---- */
int32_t lc_table_order(int32_t n_agents, int32_t n_walls, int32_t* agent_order, int32_t* wall_order);

/* ---- static BVH over the walls as ai::init builds it (ai.cpp:16-18,36; BVH.cpp:9-49,144-251), flattened in pre-order.
 * wall_order = the statics order (lc_table_order's wall_order; NULL = creation order).  Inner node: box + child indices;
 * leaf: right = -1, left = creation index of the wall.  2 n_walls - 1 nodes.  lc_create builds it itself; exported for the
 * parity tests (pinned node for node against the reference's own BVH) and for hosts that want to inspect it. ---- */
typedef struct lc_bvh_node { float ox, oy, w, h; int32_t left, right; } lc_bvh_node;
int32_t lc_static_bvh(int32_t n_walls, const lc_rect* walls, const int32_t* wall_order, lc_bvh_node* nodes, int32_t cap, int32_t* n_nodes);

/* ---- scene catalogue (demo.cpp:141-511 make_scene + :611-635 the create_robo / create_wall loops): scenes 0-18 and the
 * variable-size doorway 900 + k; returns the agent count or a negative status.  lc_scene serves the wall-free scenes (and
 * refuses the others: the walls are part of the scene), lc_scene_full all of them. ---- */
int32_t lc_scene(uint32_t scene, uint64_t seed, float* pos, float* goal, int32_t cap, float* max_duration,
                 uint32_t* rng_state_after);
int32_t lc_scene_full(uint32_t scene, uint64_t seed, float* pos, float* goal, int32_t cap, float* max_duration,
                      uint32_t* rng_state_after, lc_rect* walls, int32_t wall_cap, int32_t* n_walls);

/* ---- static-wall geometry (walled scenes; not yet part of lc_step) ----
 * Evaluates the device restatement of the reference's wall functions on n_records HOST records of 9 floats
 * {wall ox, oy, w, h; a.x, a.y; b.x, b.y; r} and writes 12 words per record:
 *   f32 Rect::intersect(a, b) (BoundVolume.cpp:160-188) | i32 Rect::line_of_sight(a, b) (:131-155) |
 *   i32 Circ(wall origin, r)::line_of_sight(a, b) (:76-98) | f32 LMP::ttc(Circ(a, r), b, wall, 0) (LMP.cpp:13-34,64-82) |
 *   i32 circ_rect_collider (BoundVolume.cpp:196-201) | f32 x2 push vector of BVH::query_ (BVH.cpp:121-130) |
 *   i32 Cspace2d::line_of_sight(a, b) (PRM.cpp:133-173) | i32 wall interacts, f32 x2 its force (LMP.cpp:307-319,128-136) | 0 */
int32_t lc_wall_geometry(int32_t n_records, const float* records, uint32_t* words, int32_t device);

/* LMP::lookahead (LMP.cpp:84-113) on the device, one agent per record: advances num_done along the waypoint list (target
 * reached within 1 m; later waypoints skipped while visible through the c-space of `walls`) and returns the new
 * local_goal.  Parity entry point like lc_wall_geometry; plans here are capped at LC_LOOK_MAX_WAYPOINTS. */
#define LC_LOOK_MAX_WAYPOINTS 8
typedef struct lc_look_record {
    float pos[2];        /* BoundVolume origin of the agent */
    float local_goal[2]; /* Agent::local_goal before the call */
    int32_t num_done;    /* Agent::num_done before the call, >= 0 */
    int32_t n_plan;      /* 1..LC_LOOK_MAX_WAYPOINTS */
    float plan[LC_LOOK_MAX_WAYPOINTS][2];
} lc_look_record;
typedef struct lc_look_result { int32_t num_done; float local_goal[2]; } lc_look_result;
int32_t lc_lookahead(int32_t n_walls, const lc_rect* walls, float agent_radius, int32_t n_records,
                     const lc_look_record* records, lc_look_result* results, int32_t device);

/* ---- host-side global planner for walled scenes: Cspace2d + PRM roadmap + A* as ai::init sets them up
 * (ai.cpp:41-63: threshold 5*sqrt(2), no bin samples, eight points per inflated wall; PRM.cpp:8-105; GMP.cpp:7-110).
 * The roadmap is stateful exactly as the reference's shared std_prm is: plan calls must come in the reference's order
 * (ai::init's agent loop, then the replans of ai.cpp:161-166) for equal-cost ties to resolve identically. ---- */
typedef struct lc_planner lc_planner;
int32_t lc_planner_create(int32_t n_walls, const lc_rect* walls, float agent_radius, lc_planner** out);
void lc_planner_destroy(lc_planner* p);
int32_t lc_planner_clone(const lc_planner* p, lc_planner** out);  /* same roadmap state (ids handed out, tie order), independent from then on */
int32_t lc_planner_roadmap_size(lc_planner* p, int32_t* vertices, int32_t* directed_edges);
/* GMP::plan_one for one agent: waypoints [n][2] from the first after `start` to `goal` inclusive (n = 0: unreachable);
 * min_time = path length - 1 (ai.cpp:118-126).  LC_ERR_CAPACITY (n still reported) when n > cap. */
int32_t lc_planner_plan(lc_planner* p, const float* start, const float* goal, int32_t cap, float* waypoints,
                        int32_t* n_waypoints, float* min_time);

const char* lc_last_error(void);
const char* lc_version(void);

#ifdef __cplusplus
}
#endif
#endif
