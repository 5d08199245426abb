/* O-canon — CPU restatement of learn-comms' per-frame loop.  TEST INFRASTRUCTURE ONLY.
 *
 * This header and canon.c are the parity oracle for the CUDA path.  They are never linked
 * into, imported by, or executed from the product library (liblearncomms.so); only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load them.
 *
 * The algorithm restated is the reference frame loop enginee/src/main.cpp:98-150 with
 * canonical (ascending agent index) table order, 32-bit ids and a Jacobi collision pass
 * (SURVEY.md §8c "O-canon").  Every function in canon.c cites the reference file:line it
 * follows.  Parity status: the reference ships no tests or golden vectors
 * (enginee/test/PURPOSE), so this oracle is pinned against the reference's own translation
 * units compiled here (oracle/_ref, see oracle/ref/) — see DESIGN.md §Oracle for the measured
 * O-canon vs O-ref gap.
 */
#ifndef LEARN_COMMS_ORACLE_CANON_H
#define LEARN_COMMS_ORACLE_CANON_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct canon_sim canon_sim;

/* axis-aligned rectangular wall (demo.cpp:515-537 create_wall; BoundVolume.h Rect) */
typedef struct { float ox, oy, w, h; } canon_rect;

typedef struct {
    int32_t n;               /* robots */
    const float* pos;        /* [n][2] initial (x,z) */
    const float* goal;       /* [n][2] final_goal */
    float radius;            /* 0.15 (demo.cpp:540,552) */
    float max_duration;      /* Scene::max_duration */
    uint32_t rng_state;      /* minstd_rand0 state BEFORE the first per-frame draw */
    const float* m_cv;       /* [3][8] row-major  (comm.cpp:124-131) */
    const float* m_f;        /* [2][3] row-major  (comm.cpp:112-121) */
    int32_t comm_enabled;    /* 0 = nocomm_norender flavour (main.cpp:130-132) */
    int32_t use_grid;        /* 0 = brute-force all pairs (reference), 1 = uniform-grid search (same result) */
    int32_t n_walls;         /* static Rect obstacles (0 in wall-free scenes) */
    const canon_rect* walls;
    /* table iteration orders (quirk Q4): table_order[k] = index of the k-th agent every component-table loop visits,
     * wall_order[k] = index of the k-th wall the statics loop (LMP.cpp:307) and the static push-out visit; NULL = ascending */
    const int32_t* table_order;
    const int32_t* wall_order;
    /* global planner (GMP::plan_one, GMP.cpp:91-110), required when n_walls > 0: fills waypoints[cap][2] with the path from
     * the first waypoint after `start` to `goal` inclusive and returns their number (0 = unreachable, < 0 = error).  Called
     * once per agent in table order at creation (ai::init, ai.cpp:107-127) and for every replan (ai.cpp:161-166).  The
     * roadmap behind it is stateful in the reference (shared std_prm), so it is NOT restated here: tests bind the host
     * planner that tests/test_planner.py pins bit for bit against the reference's own Cspace2d / PRM / GMP. */
    int32_t (*plan_fn)(void* user, int32_t agent, const float* start, const float* goal, int32_t cap, float* waypoints);
    void* plan_user;
    /* optional flattened static BVH (BVH.cpp:9-49,144-251 over the walls): n_bvh nodes of 6 floats + 2 ints laid out as
     * canon_bvh_node; NULL = every wall is tested with the exact leaf predicate in wall_order */
    int32_t n_bvh;
    const void* bvh;
    /* feature-set variant of comm::x (comm.cpp:22-28): 0 = HEAD (COSINE_REL + PROX + SOFTSIGN_GOAL); bit 32 = plain
     * (x = [c_j, rel_v, rel_p, g], comm.cpp:189), 64 = PROX undefined, 128 = SOFTSIGN_GOAL undefined */
    uint32_t features;
} canon_desc;

/* one node of the static BVH: aabb (Rect), children (left, right >= 0) or leaf (right < 0, left = wall index) */
typedef struct { float ox, oy, w, h; int32_t left, right; } canon_bvh_node;
#define CANON_MAX_PLAN 64

canon_sim* canon_create(const canon_desc* d);
void canon_destroy(canon_sim* s);

/* run `nframes` frames of main.cpp:98-150 (S0..S4).  Stops early once demo::run reported
 * all_done (the frame in which it is reported still executes S1..S4, as in the reference).
 * Returns the number of frames executed. */
int32_t canon_step(canon_sim* s, int32_t nframes);
/* run exactly nframes without the all_done latch ever stopping the dynamics (fixed-step benchmarks);
 * scoring still latches at the first all_done. */
int32_t canon_step_fixed(canon_sim* s, int32_t nframes);
int32_t canon_finished(const canon_sim* s);
uint32_t canon_frame(const canon_sim* s);

/* state access (teacher forcing).  Arrays are [n][2] except c [n][3]; any pointer may be NULL. */
void canon_get_state(const canon_sim* s, float* pos, float* vel, float* vel_forehalf, float* c,
                     float* facing, float* local_goal, float* goal_dist, int32_t* num_done);
void canon_set_state(canon_sim* s, const float* pos, const float* vel, const float* vel_forehalf,
                     const float* c, const float* facing, const float* local_goal,
                     const float* goal_dist, const int32_t* num_done);
void canon_get_rng(const canon_sim* s, uint32_t* state);
void canon_set_rng(canon_sim* s, uint32_t state);

/* scores: comm.cpp:136-142 (avg_velocity, run_avg_time + 3*run_std_time) + demo.cpp:650-663 */
void canon_get_scores(const canon_sim* s, float* avg_velocity, float* group_time, float* run_avg,
                      float* run_std, int32_t* total_num_done);

/* diagnostics of the LAST executed frame */
void canon_get_partners(const canon_sim* s, int32_t* partner);               /* [n] comm.cpp:201-220 */
void canon_get_neighbor_counts(const canon_sim* s, int32_t* cnt);            /* [n] |{j : ttc_ij <= 5}| */
/* neighbour lists, CSR; returns total; pass idx=NULL to query the size */
int64_t canon_get_neighbor_lists(const canon_sim* s, int64_t* row_ptr, int32_t* idx);
int64_t canon_overlap_pairs(const canon_sim* s);   /* unordered pairs overlapping at frame start */
int64_t canon_collision_hits(const canon_sim* s);  /* ordered (q,bv) hits in physics::simulate */
void canon_record_neighbors(canon_sim* s, int32_t on);
/* iterate every component-table loop in `order` (permutation of 0..n-1; NULL keeps it) and, if gauss_seidel, let the
 * collision pass see already-processed agents at their integrated positions (reference semantics) */
int32_t canon_set_order(canon_sim* s, const int32_t* order, int32_t gauss_seidel);

/* RNG restatement (Seeder.cpp:5-20 + libstdc++ minstd_rand0 / uniform_real_distribution<float>) */
uint32_t canon_lcg_next(uint32_t x);
uint32_t canon_lcg_seed(uint64_t seed);
float canon_uniform(uint32_t x_after_draw, float a, float b);

/* static-wall geometry, pinned against the reference's BoundVolume.cpp / LMP.cpp /
 * PRM.cpp through oracle/ref/geom_main.cpp.  2-vectors are float[2] = (x, z). */
float canon_rect_intersect(const canon_rect* r, const float* bo, const float* v);              /* BoundVolume.cpp:160-188 */
int32_t canon_rect_line_of_sight(const canon_rect* r, const float* a, const float* b);        /* BoundVolume.cpp:131-155 */
int32_t canon_circ_line_of_sight(const float* o, float r, const float* start, const float* end); /* BoundVolume.cpp:76-98 */
float canon_ttc_circ_wall(const float* o, float r, const float* iv, const canon_rect* j);     /* LMP.cpp:13-34,64-82 */
int32_t canon_wall_force(const float* o, float r, const float* vel, const canon_rect* w, float* f); /* LMP.cpp:307-319,128-136 */
int32_t canon_circ_wall_hit(const float* o, float r, const canon_rect* w, float* away);       /* BVH.cpp:121-130 */
int32_t canon_cspace_line_of_sight(const canon_rect* walls, int32_t n_walls, float agent_r, const float* a,
                                   const float* b);                                           /* PRM.cpp:133-156,169-173 */

void canon_lookahead(const canon_rect* walls, int32_t n_walls, float agent_r, const float* pos, const float* plan,
                     int32_t n_plan, int32_t* num_done, float* local_goal);                   /* LMP.cpp:84-113 */

/* scene catalogue (demo.cpp:141-511): fills pos/goal; returns n (0 = unsupported or walled: use canon_scene_full).
 * rng_state is advanced by construction-time draws (scenes 2 and 12). */
int32_t canon_scene(uint32_t scene, uint32_t* rng_state, float* pos, float* goal, float* max_duration,
                    int32_t cap);
/* the whole catalogue incl. walls (demo.cpp:161-511, create_wall :515-537: square Rects of side wall_scale) */
int32_t canon_scene_full(uint32_t scene, uint32_t* rng_state, float* pos, float* goal, float* max_duration, int32_t cap,
                         canon_rect* walls, int32_t* n_walls, int32_t wall_cap);
/* walled-scene diagnostics of the last frame: plan of an agent, lost_frames, replans / wall hits so far */
int32_t canon_get_plan(const canon_sim* s, int32_t agent, float* waypoints, int32_t cap);
void canon_get_walled(const canon_sim* s, int32_t* n_plan, int32_t* lost_frames, int64_t* replans, int64_t* wall_hits);

/* the renderer's message -> colour maps (render.cpp:87-172; vis_f = render::VisF: 0 XYZ2LAB, 1 X2RG, 2 Y2RG, 3 Z2RG,
 * 4 XYZ2LUV) and a CPU restatement of the top-down picture render::draw produces (render.cpp:352-355,364-415;
 * Camera.cpp:83-93; demo.cpp:540-541,559,598): checker of lc_render.  render.cpp itself needs an OpenGL context and cannot
 * be built here, so this half of the oracle is unpinned.  rgba = [height][width][4] bytes. */
void canon_comm_vis(int32_t vis_f, const float* c, float* rgb);
void canon_render(int32_t n, const float* pos, const float* c, int32_t n_walls, const canon_rect* walls, int32_t vis_f,
                  float cam_dist, int32_t width, int32_t height, uint8_t* rgba);

/* config / result files (comm.cpp:74-142, pso_opt.py:17-41) */
int32_t canon_read_config(const char* path, float* m_cv /*[24]*/, float* m_f /*[6]*/);
int32_t canon_write_result(const char* path, float avg_velocity, float group_time);

#ifdef __cplusplus
}
#endif
#endif
