/* O-canon — CPU restatement of learn-comms' per-frame loop.  TEST INFRASTRUCTURE ONLY.
 * See canon.h for the rules (never linked into the product).  Compile with
 *   gcc -O2 -ffp-contract=off -fno-fast-math  (no FMA contraction: the reference's CMake sets no -march / -O,
 *   enginee/CMakeLists.txt:68-80, so it executes plain SSE2 mul+add).
 * All arithmetic is IEEE binary32 unless marked f64.  "table order" = ascending agent index.
 */
#include "canon.h"
#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---------- glm restatement (glm @ d214fbaa, 0.9.9 API) ---------- */
typedef struct { float x, y; } v2;
static inline v2 V2(float x, float y) { v2 r = {x, y}; return r; }
static inline v2 add2(v2 a, v2 b) { return V2(a.x + b.x, a.y + b.y); }
static inline v2 sub2(v2 a, v2 b) { return V2(a.x - b.x, a.y - b.y); }
static inline v2 mul2s(v2 a, float s) { return V2(a.x * s, a.y * s); }
static inline v2 div2s(v2 a, float s) { return V2(a.x / s, a.y / s); }
static inline float dot2(v2 a, v2 b) { return a.x * b.x + a.y * b.y; }       /* tmp.x + tmp.y */
static inline float len2_2(v2 a) { return dot2(a, a); }                        /* gtx/norm length2 */
static inline float len_2(v2 a) { return sqrtf(dot2(a, a)); }
static inline v2 normalize2(v2 a) { return mul2s(a, 1.0f / sqrtf(dot2(a, a))); } /* v * inversesqrt(dot) */
static inline float glm_min(float a, float b) { return (b < a) ? b : a; }
static inline float glm_max(float a, float b) { return (a < b) ? b : a; }
static inline float glm_clamp(float x, float lo, float hi) { return glm_min(glm_max(x, lo), hi); }
/* vec3 with y == 0: dot = (x*x' + 0*0) + z*z' */
static inline float dot3_xz(float ax, float az, float bx, float bz) { return (ax * bx + 0.0f * 0.0f) + az * bz; }

/* ---------- RNG: Seeder.cpp:5-20; libstdc++ minstd_rand0 + uniform_real_distribution<float> ---------- */
uint32_t canon_lcg_next(uint32_t x) { return (uint32_t)(((uint64_t)x * 16807u) % 2147483647u); }
uint32_t canon_lcg_seed(uint64_t seed) {
    /* linear_congruential_engine::seed(s): if (c % m == 0 && s % m == 0) x = 1 else x = s % m */
    uint64_t x = seed % 2147483647u;
    return x == 0 ? 1u : (uint32_t)x;
}
/* generate_canonical<float,24>: k = 1 call; sum = float(x - min); tmp = float(2147483646.0L);
 * ret = sum / tmp, clamped below 1; then u * (b - a) + a   (bits/random.tcc) */
float canon_uniform(uint32_t x, float a, float b) {
    float sum = (float)(x - 1u);
    float tmp = (float)2147483646.0L;
    float ret = sum / tmp;
    if (ret >= 1.0f) ret = nextafterf(1.0f, 0.0f);
    return ret * (b - a) + a;
}

/* ---------- state ---------- */
struct canon_sim {
    int32_t n;
    float r;
    float max_duration;
    uint32_t rng;
    float m_cv[24], m_f[6];
    int comm_enabled, use_grid;
    uint32_t features;
    int32_t n_walls;
    canon_rect* walls;
    /* Dynamics.h:2-11 (live subset), Agent.h:6-33, CommComp.h:13-17 */
    v2 *pos, *vel, *vfh, *force;
    v2 *goal, *local_goal, *facing;
    float (*c)[3], (*c_buf)[3];
    v2* f_buf;
    float *goal_dist, *min_time;
    int32_t* num_done;
    uint8_t* counted;
    uint32_t* lost;
    /* demo.cpp:650-653 */
    int32_t total_num_done;
    float run_avg, M2, run_std;
    float avg_velocity; /* physics.cpp:10 */
    uint32_t frame;
    int finished;
    /* diagnostics */
    int32_t *partner, *nbr_cnt;
    int record_nbrs;
    int64_t* nbr_ptr;
    int32_t* nbr_idx;
    int64_t nbr_cap, nbr_total;
    int64_t overlap_pairs, coll_hits;
    /* table iteration order (default ascending index); reference = libstdc++ unordered_map order, quirk Q4 */
    int32_t* order;
    int gauss_seidel; /* physics pass sees already-processed agents at their integrated positions (reference, §3.4) */
    /* walled scenes: per-agent waypoint lists (Agent::plan, Agent.h:14), planner callback, wall visiting order, static BVH */
    float* plan;       /* [n][CANON_MAX_PLAN][2] */
    int32_t* n_plan;   /* [n]; 0 = no plan (unreachable goal): has_plan() false, Agent.h:20-22 */
    int32_t* wall_order;
    int32_t (*plan_fn)(void*, int32_t, const float*, const float*, int32_t, float*);
    void* plan_user;
    int32_t n_bvh;
    canon_bvh_node* bvh;
    int64_t replans, wall_hits;
    /* scratch */
    v2* pos_new;
    /* grid scratch */
    int32_t *cell_start, *cell_items;
    int32_t gcap;
};

/* Agent.h:20-26: has_plan() = plan non-empty; done() = has_plan() && num_done == plan->size().  Wall-free scenes:
 * GMP::plan_one yields plan = [final_goal] (GMP.cpp:91-110). */
static inline int has_plan(const canon_sim* s, int i) { return s->n_plan[i] > 0; }
static inline int is_done(const canon_sim* s, int i) { return has_plan(s, i) && s->num_done[i] == s->n_plan[i]; }
static inline float* plan_of(const canon_sim* s, int i) { return s->plan + (size_t)i * CANON_MAX_PLAN * 2; }

/* GMP::plan_one (GMP.cpp:91-110): new waypoint list from the agent's current position, num_done = 0 */
static int plan_one(canon_sim* s, int i) {
    float* wp = plan_of(s, i);
    if (s->n_walls > 0) {
        float start[2] = {s->pos[i].x, s->pos[i].y}, goal[2] = {s->goal[i].x, s->goal[i].y};
        int32_t k = s->plan_fn(s->plan_user, i, start, goal, CANON_MAX_PLAN, wp);
        if (k < 0 || k > CANON_MAX_PLAN) { fprintf(stderr, "canon: planner failed for agent %d (%d)\n", i, k); abort(); }
        s->n_plan[i] = k;
    } else {
        wp[0] = s->goal[i].x; wp[1] = s->goal[i].y;
        s->n_plan[i] = 1;
    }
    s->num_done[i] = 0;
    return s->n_plan[i];
}

canon_sim* canon_create(const canon_desc* d) {
    if (d->n_walls > 0 && !d->plan_fn) return NULL; /* walled rollouts need the global planner (canon.h) */
    canon_sim* s = (canon_sim*)calloc(1, sizeof(*s));
    int n = d->n;
    s->n = n;
    s->r = d->radius;
    s->max_duration = d->max_duration;
    s->rng = d->rng_state;
    s->comm_enabled = d->comm_enabled;
    s->use_grid = d->use_grid;
    s->features = d->features;
    if (d->m_cv) memcpy(s->m_cv, d->m_cv, sizeof(s->m_cv));
    if (d->m_f) memcpy(s->m_f, d->m_f, sizeof(s->m_f));
    s->n_walls = d->n_walls;
    if (d->n_walls > 0) {
        s->walls = (canon_rect*)malloc(sizeof(canon_rect) * (size_t)d->n_walls);
        memcpy(s->walls, d->walls, sizeof(canon_rect) * (size_t)d->n_walls);
    }
#define AL(p, T) s->p = (T*)calloc((size_t)n + 1, sizeof(T))
    AL(pos, v2); AL(vel, v2); AL(vfh, v2); AL(force, v2); AL(goal, v2); AL(local_goal, v2); AL(facing, v2);
    AL(f_buf, v2); AL(goal_dist, float); AL(min_time, float); AL(num_done, int32_t); AL(counted, uint8_t);
    AL(lost, uint32_t); AL(partner, int32_t); AL(nbr_cnt, int32_t); AL(pos_new, v2);
#undef AL
#define AL2(p) s->p = (int32_t*)calloc((size_t)n + 1, sizeof(int32_t))
    AL2(order);
    AL2(n_plan);
    for (int i = 0; i < n; ++i) s->order[i] = d->table_order ? d->table_order[i] : i;
    s->plan = (float*)calloc((size_t)(n + 1) * CANON_MAX_PLAN * 2, sizeof(float));
    s->plan_fn = d->plan_fn;
    s->plan_user = d->plan_user;
    if (d->n_walls > 0) {
        s->wall_order = (int32_t*)malloc(sizeof(int32_t) * (size_t)d->n_walls);
        for (int k = 0; k < d->n_walls; ++k) s->wall_order[k] = d->wall_order ? d->wall_order[k] : k;
        if (d->n_bvh > 0 && d->bvh) {
            s->n_bvh = d->n_bvh;
            s->bvh = (canon_bvh_node*)malloc(sizeof(canon_bvh_node) * (size_t)d->n_bvh);
            memcpy(s->bvh, d->bvh, sizeof(canon_bvh_node) * (size_t)d->n_bvh);
        }
    }
    s->c = calloc((size_t)n + 1, sizeof(float[3]));
    s->c_buf = calloc((size_t)n + 1, sizeof(float[3]));
    for (int i = 0; i < n; ++i) {
        /* demo.cpp:563-581 create_robo; physics.cpp:15-21 prewarm (all zero) */
        s->pos[i] = V2(d->pos[2 * i], d->pos[2 * i + 1]);
        s->goal[i] = V2(d->goal[2 * i], d->goal[2 * i + 1]);
        s->facing[i] = sub2(s->goal[i], s->pos[i]); /* un-normalised, demo.cpp:580 */
    }
    for (int oi = 0; oi < n; ++oi) { /* ai.cpp:107-127 ai::init, POOL.for_<Agent> = table order (the planner is stateful) */
        int i = s->order[oi];
        s->local_goal[i] = V2(0.f, 0.f); /* ai.cpp:112 */
        plan_one(s, i);                  /* ai.cpp:115 */
        s->goal_dist[i] = len_2(sub2(s->goal[i], s->pos[i])); /* ai.cpp:116 */
        float mt = -1.0f;                /* ai.cpp:121-126: path length minus the untimed last metre */
        v2 last = s->pos[i];
        const float* wp = plan_of(s, i);
        for (int k = 0; k < s->n_plan[i]; ++k) {
            v2 p = V2(wp[2 * k], wp[2 * k + 1]);
            mt += len_2(sub2(p, last));
            last = p;
        }
        s->min_time[i] = mt;
    }
    return s;
}

void canon_destroy(canon_sim* s) {
    if (!s) return;
    free(s->pos); free(s->vel); free(s->vfh); free(s->force); free(s->goal); free(s->local_goal);
    free(s->facing); free(s->c); free(s->c_buf); free(s->f_buf); free(s->goal_dist); free(s->min_time);
    free(s->num_done); free(s->counted); free(s->lost); free(s->partner); free(s->nbr_cnt);
    free(s->nbr_ptr); free(s->nbr_idx); free(s->pos_new); free(s->cell_start); free(s->cell_items);
    free(s->walls); free(s->order); free(s->plan); free(s->n_plan); free(s->wall_order); free(s->bvh);
    free(s);
}

/* ---------- S0: demo::run / update_runs (demo.cpp:655-702) ---------- */
static void update_runs(canon_sim* s, float overhead) { /* demo.cpp:655-663 */
    float ftotal = (float)(++s->total_num_done);
    float delta = overhead - s->run_avg;
    s->run_avg += delta / ftotal;
    float delta2 = overhead - s->run_avg;
    s->M2 += delta * delta2;
    s->run_std = sqrtf(s->M2 / ftotal);
}

static int demo_run(canon_sim* s, uint32_t frame_count) { /* demo.cpp:665-702 */
    int all_done = 1;
    float dt = 1.f / 60.f;
    float fake_time = dt * (float)frame_count;
    for (int k = 0; k < s->n; ++k) {
        int i = s->order[k];
        if (!is_done(s, i)) {
            all_done = 0;
        } else if (!s->counted[i]) {
            update_runs(s, fake_time - s->min_time[i]);
            s->counted[i] = 1;
        }
    }
    if (fake_time > s->max_duration) { /* demo.cpp:691-700 */
        for (int k = 0; k < s->n; ++k) {
            int i = s->order[k];
            if (!is_done(s, i)) {
                float finish_time = 2 * len_2(sub2(s->goal[i], s->pos[i])) + fake_time - s->min_time[i];
                update_runs(s, finish_time);
            }
        }
        all_done = 1;
    }
    return all_done;
}

/* ---------- S1: comm::run (comm.cpp:144-305, CommComp.h:18-38) ---------- */

/* comm.cpp:201-220 select_nearest — returns the FARTHEST other agent (quirk Q1): running max of
 * length2(pos_k.xz - pos_i.xz), strict '>', initialised to FLT_MIN, table order. */
static int farthest_brute(const canon_sim* s, int i, const int32_t* cand, int ncand) {
    v2 p_i = s->pos[i];
    float max_dist2 = FLT_MIN;
    int best = -1;
    if (cand) {
        for (int q = 0; q < ncand; ++q) {
            int k = cand[q];
            if (k == i) continue;
            float dist2 = len2_2(sub2(s->pos[k], p_i));
            if (dist2 > max_dist2) { max_dist2 = dist2; best = k; }
        }
    } else {
        for (int q = 0; q < s->n; ++q) {
            int k = s->order[q];
            if (k == i) continue;
            float dist2 = len2_2(sub2(s->pos[k], p_i));
            if (dist2 > max_dist2) { max_dist2 = dist2; best = k; }
        }
    }
    return best;
}

/* Exact pruning for large N (not in the reference, which is O(N^2)): the farthest point from any query is
 * a convex-hull vertex; agents strictly inside (by margin delta) the polygon spanned by the K directional
 * extreme agents can never win even under float rounding of length2 (margin argument in DESIGN.md).  The
 * survivors are scanned in table order with the reference's own predicate, so the result equals brute force. */
#define HULL_K 64
static int hull_candidates(const canon_sim* s, int32_t* cand) {
    int n = s->n;
    double lox = 1e300, hix = -1e300, loy = 1e300, hiy = -1e300;
    for (int i = 0; i < n; ++i) {
        double x = s->pos[i].x, y = s->pos[i].y;
        if (!(x == x) || !(y == y)) continue;
        if (x < lox) lox = x; if (x > hix) hix = x;
        if (y < loy) loy = y; if (y > hiy) hiy = y;
    }
    double cx = 0.5 * (lox + hix), cy = 0.5 * (loy + hiy);
    double extent = fmax(hix - lox, hiy - loy);
    double delta = 1e-5 * extent + 1e-4;
    int ext[HULL_K];
    for (int k = 0; k < HULL_K; ++k) {
        double th = 2.0 * 3.14159265358979323846 * k / HULL_K, dx = cos(th), dy = sin(th), best = -1e300;
        int bi = -1;
        for (int i = 0; i < n; ++i) {
            double v = (s->pos[i].x - cx) * dx + (s->pos[i].y - cy) * dy;
            if (v > best) { best = v; bi = i; }
        }
        ext[k] = bi;
    }
    /* edges between consecutive distinct extremes (CCW) */
    double ex[HULL_K], ey[HULL_K], enx[HULL_K], eny[HULL_K];
    int ne = 0;
    for (int k = 0; k < HULL_K; ++k) {
        int a = ext[k], b = ext[(k + 1) % HULL_K];
        if (a < 0 || b < 0 || a == b) continue;
        double ax = s->pos[a].x, ay = s->pos[a].y, bx = s->pos[b].x, by = s->pos[b].y;
        double ux = bx - ax, uy = by - ay, l = sqrt(ux * ux + uy * uy);
        if (l == 0) continue;
        ex[ne] = ax; ey[ne] = ay; enx[ne] = -uy / l; eny[ne] = ux / l; /* inward normal of a CCW polygon */
        ++ne;
    }
    int nc = 0;
    for (int i = 0; i < n; ++i) {
        int inside = ne >= 3;
        double x = s->pos[i].x, y = s->pos[i].y;
        if (!(x == x) || !(y == y)) inside = 0;
        for (int e = 0; e < ne && inside; ++e)
            if ((x - ex[e]) * enx[e] + (y - ey[e]) * eny[e] < delta) inside = 0;
        if (!inside) cand[nc++] = i;
    }
    return nc;
}

static void comm_run(canon_sim* s) { /* comm.cpp:267-305 */
    int n = s->n;
    int32_t* cand = NULL;
    int ncand = 0;
    if (s->use_grid && n > 2048) {
        cand = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
        ncand = hull_candidates(s, cand);
    }
    for (int oi = 0; oi < n; ++oi) { /* pass 1, comm.cpp:268-286 */
        int i = s->order[oi];
        int j = farthest_brute(s, i, cand, ncand);
        s->partner[i] = j;
        if (j < 0) { /* quirk Q3: the reference dereferences a null CommComp here (UB) */
            fprintf(stderr, "canon: agent %d has no comm partner (reference UB, comm.cpp:273-285)\n", i);
            abort();
        }
        v2 p_j = V2(s->pos[j].x, 0.f); /* vec3 -> vec2 keeps (x, y): quirk Q2, comm.cpp:276-277 */
        v2 v_j = V2(s->vel[j].x, 0.f);
        /* comm.cpp:144-192 x() with COSINE_REL + PROX + SOFTSIGN_GOAL */
        v2 facing = s->facing[i];
        v2 right = V2(facing.y, -facing.x); /* CommComp.h:38 */
        v2 diff_v = sub2(v_j, s->vel[i]);
        v2 rel_v = V2(right.x * diff_v.x + right.y * diff_v.y, facing.x * diff_v.x + facing.y * diff_v.y);
        const int plain = (s->features & 32u) != 0; /* neither COSINE_REL nor NORM_REL (comm.cpp:158-163,167-176) */
        float mag_rv = 0.f;
        if (!plain) {
            mag_rv = len_2(rel_v);
            if (mag_rv > 0) rel_v = div2s(rel_v, mag_rv);
        }
        v2 diff_p = sub2(p_j, s->pos[i]);
        v2 rel_p = V2(right.x * diff_p.x + right.y * diff_p.y, facing.x * diff_p.x + facing.y * diff_p.y);
        float mag_rp = 0.f;
        if (!plain) {
            mag_rp = len_2(rel_p);
            if (mag_rp > 0) {
                rel_p = div2s(rel_p, mag_rp);
                if (!(s->features & 64u)) mag_rp = 1 / mag_rp; /* PROX */
            }
        }
        float g = s->goal_dist[i];
        if (!(s->features & 128u)) g /= 1 + g; /* SOFTSIGN_GOAL */
        float x[8] = {s->c[j][0], s->c[j][1], s->c[j][2], rel_v.y, mag_rv, rel_p.y, mag_rp, g};
        if (plain) { x[3] = rel_v.x; x[4] = rel_v.y; x[5] = rel_p.x; x[6] = rel_p.y; } /* comm.cpp:189 */
        /* comm.cpp:285 M_cv * x_i, M_f * c_j.  Eigen 3.3 coeff-based lazy product, non-vectorised redux:
         * pairwise tree ((p0+p1)+(p2+p3))+((p4+p5)+(p6+p7)); depth 3: p0+(p1+p2).  Eigen is not vendored
         * (lib/init.sh:58-59) so this order is restated from the 3.3 sources' redux_novec_unroller —
         * "parity unpinned at this boundary" (SURVEY.md §8c). */
        for (int r = 0; r < 3; ++r) {
            const float* m = &s->m_cv[8 * r];
            float p[8];
            for (int q = 0; q < 8; ++q) p[q] = m[q] * x[q];
            float cp = ((p[0] + p[1]) + (p[2] + p[3])) + ((p[4] + p[5]) + (p[6] + p[7]));
            s->c_buf[i][r] = cp / (1.f + fabsf(2.f * cp)) + 0.5f; /* CommComp.h:20-27 */
        }
        float fp[2];
        for (int r = 0; r < 2; ++r) {
            const float* m = &s->m_f[3 * r];
            fp[r] = m[0] * s->c[j][0] + (m[1] * s->c[j][1] + m[2] * s->c[j][2]);
        }
        s->f_buf[i] = V2(glm_clamp(fp[0], -1.f, 1.f), glm_clamp(fp[1], -1.f, 1.f)); /* CommComp.h:30-32 */
    }
    free(cand);
    for (int oi = 0; oi < n; ++oi) { /* pass 2, comm.cpp:288-304 */
        int i = s->order[oi];
        v2 facing = s->facing[i];
        v2 right = V2(facing.y, -facing.x);
        v2 fb = s->f_buf[i];
        v2 diff = V2(right.x * fb.x + facing.x * fb.y, right.y * fb.x + facing.y * fb.y);
        if (!is_done(s, i)) s->force[i] = add2(s->force[i], diff);
        float speed = sqrtf(dot3_xz(s->vel[i].x, s->vel[i].y, s->vel[i].x, s->vel[i].y));
        v2 to_lg = sub2(s->local_goal[i], s->pos[i]); /* a->start == synced position (Pool.cpp:86-88) */
        if (speed > 0 && len_2(to_lg) > 0) s->facing[i] = normalize2(to_lg);
        s->c[i][0] = s->c_buf[i][0]; s->c[i][1] = s->c_buf[i][1]; s->c[i][2] = s->c_buf[i][2]; /* swap() */
        s->f_buf[i] = V2(0.f, 0.f);
    }
}

/* ---------- uniform grid (search acceleration only; results identical to brute force) ---------- */
typedef struct { float x0, y0, h; int nx, ny; } grid_t;
static grid_t g_grid;

static float max_speed(const canon_sim* s) {
    float m = 0.f;
    for (int i = 0; i < s->n; ++i) {
        float v = len_2(s->vel[i]);
        if (v > m) m = v;
    }
    return m;
}

static inline int cell_of(const grid_t* g, v2 p, int* cx, int* cy) {
    int ix = (int)floorf((p.x - g->x0) / g->h), iy = (int)floorf((p.y - g->y0) / g->h);
    if (ix < 0) ix = 0; if (ix >= g->nx) ix = g->nx - 1;
    if (iy < 0) iy = 0; if (iy >= g->ny) iy = g->ny - 1;
    *cx = ix; *cy = iy;
    return iy * g->nx + ix;
}

static void grid_build(canon_sim* s, float cutoff) {
    grid_t* g = &g_grid;
    float lox = FLT_MAX, loy = FLT_MAX, hix = -FLT_MAX, hiy = -FLT_MAX;
    for (int i = 0; i < s->n; ++i) {
        v2 p = s->pos[i];
        if (p.x < lox) lox = p.x; if (p.x > hix) hix = p.x;
        if (p.y < loy) loy = p.y; if (p.y > hiy) hiy = p.y;
    }
    float ext = fmaxf(hix - lox, hiy - loy);
    g->h = fmaxf(cutoff, ext / 2048.f);
    g->x0 = lox; g->y0 = loy;
    g->nx = (int)floorf((hix - lox) / g->h) + 1;
    g->ny = (int)floorf((hiy - loy) / g->h) + 1;
    int nc = g->nx * g->ny;
    if (nc + 1 > s->gcap) {
        free(s->cell_start);
        s->cell_start = (int32_t*)malloc(sizeof(int32_t) * (size_t)(nc + 1));
        s->gcap = nc + 1;
    }
    if (!s->cell_items) s->cell_items = (int32_t*)malloc(sizeof(int32_t) * (size_t)s->n);
    memset(s->cell_start, 0, sizeof(int32_t) * (size_t)(nc + 1));
    int cx, cy;
    for (int i = 0; i < s->n; ++i) s->cell_start[cell_of(g, s->pos[i], &cx, &cy) + 1]++;
    for (int c = 0; c < nc; ++c) s->cell_start[c + 1] += s->cell_start[c];
    int32_t* fill = (int32_t*)malloc(sizeof(int32_t) * (size_t)nc);
    memcpy(fill, s->cell_start, sizeof(int32_t) * (size_t)nc);
    for (int i = 0; i < s->n; ++i) s->cell_items[fill[cell_of(g, s->pos[i], &cx, &cy)]++] = i; /* ascending id per cell */
    free(fill);
}

/* ---------- S2: ai::update_agents + LMP::calc_sum_force (ai.cpp:132-174, LMP.cpp:13-158,244-335) ---------- */

/* LMP.cpp:13-55 ttc / ttc_(Circ,Circ) + BoundVolume.cpp:102-122 Circ::intersect */
static inline float ttc_circ_circ(v2 oi, v2 vi, v2 oj, v2 vj, float R) {
    v2 d = sub2(vj, vi);
    v2 m = sub2(oj, oi);
    float a = dot2(d, d);
    float b = dot2(m, d);
    float c = dot2(m, m) - R * R;
    float t;
    if (c > 0.f && b > 0.f) return FLT_MAX;
    float discrim = b * b - a * c;
    if (discrim < 0.f) return FLT_MAX;
    t = (-b - sqrtf(discrim)) / a;
    if (t < 0.f) t = 0.f;
    if (t < 0 || isnan(t)) return FLT_MAX; /* LMP.cpp:31-34 */
    return t;
}

/* LMP.cpp:115-127 ttc_forces_ (f64 magnitude, quirk Q11) */
static inline v2 ttc_forces_(double ttc, v2 dir) {
    float len = len2_2(dir);
    if (len > 0) dir = div2s(dir, sqrtf(len));
    double t_h = 5.0;
    double mag = 0;
    if (ttc >= 0 && ttc <= t_h) mag = (t_h - ttc) / (ttc + 0.001);
    mag = mag > 20 ? 20 : mag;
    return V2((float)(mag * (double)dir.x), (float)(mag * (double)dir.y));
}

typedef struct { int32_t id; v2 f; } hit_t;
static int hit_cmp(const void* a, const void* b) { return ((const hit_t*)a)->id - ((const hit_t*)b)->id; }

static void record_nbr(canon_sim* s, int32_t j) {
    if (s->nbr_total == s->nbr_cap) {
        s->nbr_cap = s->nbr_cap ? s->nbr_cap * 2 : 1024;
        s->nbr_idx = (int32_t*)realloc(s->nbr_idx, sizeof(int32_t) * (size_t)s->nbr_cap);
    }
    s->nbr_idx[s->nbr_total++] = j;
}

static void update_agents(canon_sim* s) {
    int n = s->n;
    float R = s->r + s->r;
    int replanned = 0;
    const int limit = 2;
    hit_t* hits = NULL;
    int hcap = 0;
    s->overlap_pairs = 0;
    s->nbr_total = 0;
    if (s->record_nbrs && !s->nbr_ptr) s->nbr_ptr = (int64_t*)calloc((size_t)n + 1, sizeof(int64_t));
    int stencil = 1;
    if (s->use_grid) {
        float cutoff = R + 10.f * max_speed(s) * 1.001f + 0.01f;
        /* the collision pass reuses this grid with a +-1 cell block: two agents each move <= dt per frame */
        cutoff = fmaxf(cutoff, R + 2.f * (1.f / 60.f) * 1.01f + 0.004f);
        grid_build(s, cutoff);
        stencil = (int)ceilf(cutoff / g_grid.h);
    }
    for (int oi = 0; oi < n; ++oi) {
        int i = s->order[oi];
        /* ai.cpp:153-168 */
        s->goal_dist[i] = len_2(sub2(s->goal[i], s->pos[i]));
        v2 o = s->pos[i];
        const float opos[2] = {o.x, o.y};
        int next = s->num_done[i];
        int visible_next = 0;
        if (has_plan(s, i) && next < s->n_plan[i]) /* Cspace2d::line_of_sight (PRM.cpp:169-173): trivially true without walls */
            visible_next = canon_cspace_line_of_sight(s->walls, s->n_walls, s->r, opos, plan_of(s, i) + 2 * next);
        s->lost[i] = visible_next ? 0 : s->lost[i] + 1;
        int need_replan = !has_plan(s, i) || (is_done(s, i) && s->goal_dist[i] > 1.f) ||
                          (!is_done(s, i) && !visible_next && s->lost[i] >= 360);
        if (replanned < limit && need_replan) {
            plan_one(s, i); /* GMP::plan_one from the current (synced) position; num_done = 0 (GMP.cpp:91-110) */
            ++replanned;
            ++s->replans;
            s->lost[i] = 0;
        }
        /* LMP::calc_sum_force, LMP.cpp:244-273 */
        v2 goal_vel;
        if (has_plan(s, i)) {
            float lg[2] = {s->local_goal[i].x, s->local_goal[i].y};
            int32_t nd = s->num_done[i];
            canon_lookahead(s->walls, s->n_walls, s->r, opos, plan_of(s, i), s->n_plan[i], &nd, lg); /* LMP.cpp:84-113 */
            s->num_done[i] = nd;
            s->local_goal[i] = V2(lg[0], lg[1]);
            s->rng = canon_lcg_next(s->rng); /* one draw, used for both components (LMP.cpp:261-263) */
            float eps = canon_uniform(s->rng, -.01f, .01f);
            v2 fuzzy = add2(s->local_goal[i], V2(eps, eps));
            v2 diff = sub2(fuzzy, o);
            goal_vel = mul2s(div2s(diff, glm_max(1.0f, len_2(diff))), 1.0f);
        } else { /* LMP.cpp:267-270: no plan, no draw */
            s->local_goal[i] = o;
            goal_vel = V2(0.f, 0.f);
        }
        v2 vi = s->vel[i];
        v2 goal_F = mul2s(sub2(goal_vel, vi), 2.0f);
        v2 ttc_F = V2(0.f, 0.f);
        int cnt = 0;
        if (s->record_nbrs) s->nbr_ptr[i] = s->nbr_total;
#define PAIR(j)                                                                           \
    do {                                                                                  \
        v2 oj = s->pos[j], vj = s->vel[j];                                                \
        v2 mm = sub2(oj, o);                                                              \
        if (dot2(mm, mm) < R * R) s->overlap_pairs++; /* BoundVolume.cpp:203-207 */       \
        double ttc = (double)ttc_circ_circ(o, vi, oj, vj, R);                             \
        if (ttc > 5.0) break;                                                             \
        float tf = (float)ttc;                                                            \
        v2 dir = sub2(add2(o, V2(vi.x * tf, vi.y * tf)), add2(oj, V2(vj.x * tf, vj.y * tf))); \
        fcontrib = ttc_forces_(tf, dir);                                                  \
        hit = 1;                                                                          \
    } while (0)
        if (!s->use_grid) {
            for (int oj = 0; oj < n; ++oj) { /* LMP.cpp:288-304 */
                int j = s->order[oj];
                if (j == i) continue;
                v2 fcontrib; int hit = 0;
                PAIR(j);
                if (hit) { ttc_F = add2(ttc_F, fcontrib); ++cnt; if (s->record_nbrs) record_nbr(s, j); }
            }
        } else {
            int cx, cy, nh = 0;
            cell_of(&g_grid, o, &cx, &cy);
            for (int yy = cy - stencil; yy <= cy + stencil; ++yy) {
                if (yy < 0 || yy >= g_grid.ny) continue;
                for (int xx = cx - stencil; xx <= cx + stencil; ++xx) {
                    if (xx < 0 || xx >= g_grid.nx) continue;
                    int c = yy * g_grid.nx + xx;
                    for (int q = s->cell_start[c]; q < s->cell_start[c + 1]; ++q) {
                        int j = s->cell_items[q];
                        if (j == i) continue;
                        v2 fcontrib; int hit = 0;
                        PAIR(j);
                        if (hit) {
                            if (nh == hcap) { hcap = hcap ? 2 * hcap : 256; hits = (hit_t*)realloc(hits, sizeof(hit_t) * (size_t)hcap); }
                            hits[nh].id = j; hits[nh].f = fcontrib; ++nh;
                        }
                    }
                }
            }
            qsort(hits, (size_t)nh, sizeof(hit_t), hit_cmp); /* accumulate in table order */
            for (int q = 0; q < nh; ++q) { ttc_F = add2(ttc_F, hits[q].f); if (s->record_nbrs) record_nbr(s, hits[q].id); }
            cnt = nh;
        }
#undef PAIR
        s->nbr_cnt[i] = cnt;
        for (int k = 0; k < s->n_walls; ++k) { /* LMP.cpp:307-319: the statics loop, in the BoundVolume* table's order */
            float f[2];
            const float vel2[2] = {vi.x, vi.y};
            if (canon_wall_force(opos, s->r, vel2, &s->walls[s->wall_order[k]], f)) ttc_F = add2(ttc_F, V2(f[0], f[1]));
        }
        v2 f2d = add2(goal_F, ttc_F);
        s->force[i] = add2(s->force[i], f2d); /* ai.cpp:172 */
    }
    if (s->record_nbrs) s->nbr_ptr[n] = s->nbr_total;
    s->overlap_pairs /= 2;
    free(hits);
}

/* ---------- S3: physics::simulate (physics.cpp:23-78), Jacobi collision pass ---------- */
static inline void push_away(v2* pos, v2* vfh, v2 away) { /* physics.cpp:23-36 (acc is dead afterwards) */
    float l2 = dot3_xz(away.x, away.y, away.x, away.y);
    if (l2 > 0.f) {
        float inv = 1.0f / sqrtf(l2);
        pos->x += away.x + 0.001f * (away.x * inv);
        pos->y += away.y + 0.001f * (away.y * inv);
    }
    float d = dot3_xz(away.x, away.y, vfh->x, vfh->y);
    vfh->x += away.x * d;
    vfh->y += away.y * d;
}

/* BVH.cpp:106-120 leaf predicate + normal; BoundVolume.cpp:203-211 */
static inline int circ_circ_hit(v2 pq, v2 pb, float rq, float rb, v2* away) {
    float r = rq + rb;
    v2 diff = sub2(pq, pb);
    if (!(len2_2(diff) < r * r)) return 0;
    v2 on_q = add2(mul2s(normalize2(sub2(pb, pq)), rq), pq);    /* closest_circ_point(bv->_o, q) */
    v2 on_bv = add2(mul2s(normalize2(sub2(on_q, pb)), rb), pb); /* closest_circ_point(on_q, bv_circ) */
    *away = sub2(on_bv, on_q);
    return 1;
}

static void physics_simulate(canon_sim* s, float dt) {
    int n = s->n;
    /* phase A: leap-frog + clamp + move (physics.cpp:41-51) */
    for (int i = 0; i < n; ++i) {
        v2 vbh = s->vfh[i];
        v2 acc = V2(s->force[i].x / 1.f, s->force[i].y / 1.f);
        v2 vfh = V2(vbh.x + dt * acc.x, vbh.y + dt * acc.y);
        if (dot3_xz(vfh.x, vfh.y, vfh.x, vfh.y) > 1.f) {
            float inv = 1.0f / sqrtf(dot3_xz(vfh.x, vfh.y, vfh.x, vfh.y));
            vfh = V2(vfh.x * inv, vfh.y * inv);
        }
        s->vel[i] = V2(vfh.x / 2.f + vbh.x / 2.f, vfh.y / 2.f + vbh.y / 2.f);
        s->pos_new[i] = V2(s->pos[i].x + vfh.x * dt, s->pos[i].y + vfh.y * dt);
        s->vfh[i] = vfh;
        s->force[i] = V2(0.f, 0.f);
    }
    /* phase B: every agent queries all others at their integrated (pre-push) positions, table order
     * (the reference is Gauss-Seidel in hash-map order through a stale BVH: quirks Q4/Q8, gap measured vs O-ref) */
    s->coll_hits = 0;
    float r = s->r;
    for (int oi = 0; oi < n; ++oi) {
        int i = s->order[oi];
        v2 p = s->pos_new[i], vfh = s->vfh[i];
        v2 pq = p;
        if (!s->use_grid) {
            for (int oj = 0; oj < n; ++oj) {
                int j = s->order[oj];
                v2 away;
                if (j == i) continue;
                /* Gauss-Seidel (reference, physics.cpp:56 + BVH.cpp:109): agents later in the table still sit at their
                 * frame-start position (s->pos[j] is only overwritten once j has been processed) */
                v2 pj = (s->gauss_seidel && oj > oi) ? s->pos[j] : s->pos_new[j];
                if (circ_circ_hit(pq, pj, r, r, &away)) { push_away(&p, &vfh, away); s->coll_hits++; }
            }
        } else {
            int cx, cy, nh = 0;
            hit_t hits[64];
            cell_of(&g_grid, s->pos[i], &cx, &cy); /* grid of start-of-frame positions; moves <= dt */
            for (int yy = cy - 1; yy <= cy + 1; ++yy) {
                if (yy < 0 || yy >= g_grid.ny) continue;
                for (int xx = cx - 1; xx <= cx + 1; ++xx) {
                    if (xx < 0 || xx >= g_grid.nx) continue;
                    int c = yy * g_grid.nx + xx;
                    for (int q = s->cell_start[c]; q < s->cell_start[c + 1]; ++q) {
                        int j = s->cell_items[q];
                        v2 away;
                        if (j == i) continue;
                        if (circ_circ_hit(pq, s->pos_new[j], r, r, &away)) {
                            if (nh == 64) { fprintf(stderr, "canon: >64 simultaneous collisions\n"); abort(); }
                            hits[nh].id = j; hits[nh].f = away; ++nh;
                        }
                    }
                }
            }
            qsort(hits, (size_t)nh, sizeof(hit_t), hit_cmp);
            for (int q = 0; q < nh; ++q) { push_away(&p, &vfh, hits[q].f); s->coll_hits++; }
        }
        if (s->n_walls > 0) { /* physics.cpp:60,67-69: static_bvh->query at the integrated position, after the agent hits */
            const float q[2] = {pq.x, pq.y};
            float away[2];
            if (s->bvh) { /* BVH::query_(Circ*) (BVH.cpp:106-138): depth-first, left before right, approximate inner test */
                int32_t stack[64], sp = 0;
                stack[sp++] = 0;
                while (sp > 0) {
                    const canon_bvh_node* nd = &s->bvh[stack[--sp]];
                    if (nd->right < 0) {
                        if (canon_circ_wall_hit(q, r, &s->walls[nd->left], away)) { push_away(&p, &vfh, V2(away[0], away[1])); s->wall_hits++; }
                    } else {
                        canon_rect box = {nd->ox, nd->oy, nd->w, nd->h};
                        float unused[2];
                        if (canon_circ_wall_hit(q, r, &box, unused)) { stack[sp++] = nd->right; stack[sp++] = nd->left; }
                    }
                }
            } else {
                for (int k = 0; k < s->n_walls; ++k)
                    if (canon_circ_wall_hit(q, r, &s->walls[s->wall_order[k]], away)) { push_away(&p, &vfh, V2(away[0], away[1])); s->wall_hits++; }
            }
        }
        s->pos[i] = p;
        s->vfh[i] = vfh;
        if (!is_done(s, i)) { /* physics.cpp:72-76; vel_count is never incremented (quirk Q6) */
            float v = (float)0;
            s->avg_velocity = (s->avg_velocity * v + sqrtf(dot3_xz(s->vel[i].x, s->vel[i].y, s->vel[i].x, s->vel[i].y))) / (v + 1.f);
        }
    }
}

/* ---------- frame loop: main.cpp:98-150 ---------- */
static void one_frame(canon_sim* s, int free_run) {
    ++s->frame;
    int all_done = 0;
    if (!s->finished) all_done = demo_run(s, s->frame); /* S0 */
    if (s->comm_enabled) comm_run(s);                   /* S1  (#ifndef NO_COMM) */
    update_agents(s);                                   /* S2 */
    physics_simulate(s, 1.f / 60.f);                    /* S3; S4 all_sync is the identity in SoA */
    if (all_done) s->finished = 1;
    (void)free_run;
}

int32_t canon_step(canon_sim* s, int32_t nframes) {
    int32_t k = 0;
    while (k < nframes && !s->finished) { one_frame(s, 0); ++k; }
    return k;
}
int32_t canon_step_fixed(canon_sim* s, int32_t nframes) {
    for (int32_t k = 0; k < nframes; ++k) one_frame(s, 1);
    return nframes;
}
int32_t canon_finished(const canon_sim* s) { return s->finished; }
uint32_t canon_frame(const canon_sim* s) { return s->frame; }

/* ---------- accessors ---------- */
void canon_get_state(const canon_sim* s, float* pos, float* vel, float* vfh, float* c, float* facing,
                     float* local_goal, float* goal_dist, int32_t* num_done) {
    for (int i = 0; i < s->n; ++i) {
        if (pos) { pos[2 * i] = s->pos[i].x; pos[2 * i + 1] = s->pos[i].y; }
        if (vel) { vel[2 * i] = s->vel[i].x; vel[2 * i + 1] = s->vel[i].y; }
        if (vfh) { vfh[2 * i] = s->vfh[i].x; vfh[2 * i + 1] = s->vfh[i].y; }
        if (c) { c[3 * i] = s->c[i][0]; c[3 * i + 1] = s->c[i][1]; c[3 * i + 2] = s->c[i][2]; }
        if (facing) { facing[2 * i] = s->facing[i].x; facing[2 * i + 1] = s->facing[i].y; }
        if (local_goal) { local_goal[2 * i] = s->local_goal[i].x; local_goal[2 * i + 1] = s->local_goal[i].y; }
        if (goal_dist) goal_dist[i] = s->goal_dist[i];
        if (num_done) num_done[i] = s->num_done[i];
    }
}
void canon_set_state(canon_sim* s, const float* pos, const float* vel, const float* vfh, const float* c,
                     const float* facing, const float* local_goal, const float* goal_dist,
                     const int32_t* num_done) {
    for (int i = 0; i < s->n; ++i) {
        if (pos) s->pos[i] = V2(pos[2 * i], pos[2 * i + 1]);
        if (vel) s->vel[i] = V2(vel[2 * i], vel[2 * i + 1]);
        if (vfh) s->vfh[i] = V2(vfh[2 * i], vfh[2 * i + 1]);
        if (c) { s->c[i][0] = c[3 * i]; s->c[i][1] = c[3 * i + 1]; s->c[i][2] = c[3 * i + 2]; }
        if (facing) s->facing[i] = V2(facing[2 * i], facing[2 * i + 1]);
        if (local_goal) s->local_goal[i] = V2(local_goal[2 * i], local_goal[2 * i + 1]);
        if (goal_dist) s->goal_dist[i] = goal_dist[i];
        if (num_done) s->num_done[i] = num_done[i];
    }
}
void canon_get_rng(const canon_sim* s, uint32_t* state) { *state = s->rng; }
void canon_set_rng(canon_sim* s, uint32_t state) { s->rng = state; }

void canon_get_scores(const canon_sim* s, float* avg_velocity, float* group_time, float* run_avg,
                      float* run_std, int32_t* total_num_done) {
    if (avg_velocity) *avg_velocity = s->avg_velocity;
    if (group_time) *group_time = s->run_avg + s->run_std * 3; /* comm.cpp:139 */
    if (run_avg) *run_avg = s->run_avg;
    if (run_std) *run_std = s->run_std;
    if (total_num_done) *total_num_done = s->total_num_done;
}
void canon_get_partners(const canon_sim* s, int32_t* partner) { memcpy(partner, s->partner, sizeof(int32_t) * (size_t)s->n); }
void canon_get_neighbor_counts(const canon_sim* s, int32_t* cnt) { memcpy(cnt, s->nbr_cnt, sizeof(int32_t) * (size_t)s->n); }
int64_t canon_get_neighbor_lists(const canon_sim* s, int64_t* row_ptr, int32_t* idx) {
    if (!s->nbr_ptr) return -1;
    /* rows were recorded in table order (nbr_ptr[i] = where agent i's row starts); hand out a proper CSR by agent index,
     * each row still in the order LMP.cpp:288-304 visited the neighbours */
    int64_t at = 0;
    for (int i = 0; i < s->n; ++i) {
        if (row_ptr) row_ptr[i] = at;
        if (idx) memcpy(idx + at, s->nbr_idx + s->nbr_ptr[i], sizeof(int32_t) * (size_t)s->nbr_cnt[i]);
        at += s->nbr_cnt[i];
    }
    if (row_ptr) row_ptr[s->n] = at;
    return s->nbr_total;
}
int32_t canon_get_plan(const canon_sim* s, int32_t agent, float* waypoints, int32_t cap) {
    int32_t k = s->n_plan[agent];
    if (waypoints) memcpy(waypoints, plan_of(s, agent), sizeof(float) * 2 * (size_t)(k < cap ? k : cap));
    return k;
}
void canon_get_walled(const canon_sim* s, int32_t* n_plan, int32_t* lost_frames, int64_t* replans, int64_t* wall_hits) {
    for (int i = 0; i < s->n; ++i) {
        if (n_plan) n_plan[i] = s->n_plan[i];
        if (lost_frames) lost_frames[i] = (int32_t)s->lost[i];
    }
    if (replans) *replans = s->replans;
    if (wall_hits) *wall_hits = s->wall_hits;
}
int64_t canon_overlap_pairs(const canon_sim* s) { return s->overlap_pairs; }
int64_t canon_collision_hits(const canon_sim* s) { return s->coll_hits; }
void canon_record_neighbors(canon_sim* s, int32_t on) { s->record_nbrs = on; }
/* table iteration order (a permutation of 0..n-1) for every loop the reference runs over a component table; brute-force
 * mode only.  With the reference's own unordered_map order (dumped by oracle/_ref/oref) and gauss_seidel = 1 the
 * restatement follows the reference operation for operation. */
int32_t canon_set_order(canon_sim* s, const int32_t* order, int32_t gauss_seidel) {
    if (s->use_grid) return -1;
    if (order) memcpy(s->order, order, sizeof(int32_t) * (size_t)s->n);
    s->gauss_seidel = gauss_seidel;
    return 0;
}

/* ---------- static walls: the geometry behind SURVEY §8 rows a10 / a14 / a16 (walled scenes, §8f-2) ----------
 * Pinned bit for bit against the reference's own BoundVolume.cpp / LMP.cpp / PRM.cpp (oracle/ref/geom_main.cpp,
 * tests/golden/wall_geometry.npz).  Not yet wired into the frame loop: canon_create refuses n_walls > 0 rollouts
 * because the planner (PRM + A*, GMP.cpp) is not restated. */

/* BoundVolume.cpp:160-188 Rect::intersect */
float canon_rect_intersect(const canon_rect* r, const float* bo, const float* v) {
    float tmin = 0.0f, tmax = FLT_MAX;
    float rmin[2] = {r->ox - r->w / 2, r->oy - r->h / 2};
    float rmax[2] = {r->ox + r->w / 2, r->oy + r->h / 2};
    for (int i = 0; i < 2; ++i) {
        if (fabs(v[i]) < 0.0000001f) {
            if (bo[i] < rmin[i] || bo[i] > rmax[i]) return FLT_MAX;
        } else {
            float near = (rmin[i] - bo[i]) / v[i];
            float far = (rmax[i] - bo[i]) / v[i];
            if (near > far) { float t = near; near = far; far = t; }
            tmin = (tmin < near) ? near : tmin; /* std::max(tmin, near) */
            tmax = (far < tmax) ? far : tmax;   /* std::min(tmax, far) */
            if (tmin > tmax) return FLT_MAX;
        }
    }
    return tmin;
}

/* BoundVolume.cpp:131-155 Rect::line_of_sight (1 = segment a-b misses the rect) */
int32_t canon_rect_line_of_sight(const canon_rect* r, const float* a, const float* b) {
    v2 rmin = V2(r->ox - r->w / 2, r->oy - r->h / 2);
    v2 rmax = V2(r->ox + r->w / 2, r->oy + r->h / 2);
    v2 e = sub2(rmax, rmin);
    v2 m = sub2(sub2(add2(V2(a[0], a[1]), V2(b[0], b[1])), rmin), rmax);
    v2 d = sub2(V2(b[0], b[1]), V2(a[0], a[1]));
    float adx = (float)fabs(d.x);
    if ((float)fabs(m.x) > e.x + adx) return 1;
    float ady = (float)fabs(d.y);
    if ((float)fabs(m.y) > e.y + ady) return 1;
    adx += 0.0000001f;
    ady += 0.0000001f;
    if ((float)fabs(m.x * d.y - m.y * d.x) > e.x * ady + e.y * adx) return 1;
    return 0;
}

/* BoundVolume.cpp:76-98 Circ::line_of_sight */
int32_t canon_circ_line_of_sight(const float* o, float r, const float* start, const float* end) {
    v2 d = sub2(V2(end[0], end[1]), V2(start[0], start[1]));
    v2 m = sub2(V2(start[0], start[1]), V2(o[0], o[1]));
    float b = dot2(m, d);
    float c = dot2(m, m) - r * r;
    if (c > 0.f && b > 0.f) return 1;
    float a = dot2(d, d);
    float discrim = b * b - a * c;
    if (discrim < 0.f) return 1;
    float t = (-b - sqrtf(discrim)) / a;
    if (t * t > a) return 1;
    return 0;
}

/* BoundVolume.cpp:102-122 Circ::intersect on its own (ttc_circ_circ above folds the LMP::ttc wrapper in) */
static inline float circ_intersect(v2 co, float r, v2 o, v2 d) {
    v2 m = sub2(o, co);
    float a = dot2(d, d);
    float b = dot2(m, d);
    float c = dot2(m, m) - r * r;
    if (c > 0.f && b > 0.f) return FLT_MAX;
    float discrim = b * b - a * c;
    if (discrim < 0.f) return FLT_MAX;
    float t = (-b - sqrtf(discrim)) / a;
    if (t < 0.f) return 0.f;
    return t;
}

/* LMP.cpp:13-34 ttc (Circ vs Rect branch) + :64-82 ttc_(Circ,Rect): min over the two extruded rects and four corner
 * discs; jv = 0 for a wall (LMP.cpp:314-315) */
float canon_ttc_circ_wall(const float* o, float r, const float* iv, const canon_rect* j) {
    canon_rect h = {j->ox, j->oy, 2 * r + j->w, j->h};
    canon_rect v = {j->ox, j->oy, j->w, 2 * r + j->h};
    v2 corner[4] = {V2(j->ox + j->w / 2, j->oy + j->h / 2), V2(j->ox - j->w / 2, j->oy + j->h / 2),
                    V2(j->ox + j->w / 2, j->oy - j->h / 2), V2(j->ox - j->w / 2, j->oy - j->h / 2)};
    float dv[2] = {iv[0] - 0.f, iv[1] - 0.f};
    float t_min = FLT_MAX, t;
    t = canon_rect_intersect(&h, o, dv); if (t < t_min) t_min = t;
    t = canon_rect_intersect(&v, o, dv); if (t < t_min) t_min = t;
    for (int k = 0; k < 4; ++k) {
        t = circ_intersect(corner[k], r, V2(o[0], o[1]), V2(dv[0], dv[1]));
        if (t < t_min) t_min = t;
    }
    if (t_min < 0 || isnan(t_min)) return FLT_MAX;
    return t_min;
}

/* LMP.cpp:307-319: one iteration of the statics loop of calc_sum_force (distance gate on length2 > 5 -- quirk Q10 --
 * then ttc > 5, then ttc_forces(d, bv, bbv, ttc) :128-136).  Returns 1 and the force if the wall interacts. */
int32_t canon_wall_force(const float* o, float r, const float* vel, const canon_rect* w, float* f) {
    f[0] = f[1] = 0.f;
    if (len2_2(sub2(V2(w->ox, w->oy), V2(o[0], o[1]))) > 5.0) return 0;
    double ttc = canon_ttc_circ_wall(o, r, vel, w);
    if (ttc > 5.0) return 0;
    float tf = (float)ttc;
    v2 dir = sub2(add2(V2(o[0], o[1]), V2(vel[0] * tf, vel[1] * tf)), V2(w->ox, w->oy));
    v2 F = ttc_forces_(tf, dir);
    f[0] = F.x; f[1] = F.y;
    return 1;
}

/* BoundVolume.cpp:209-211 closest_circ_point */
static inline v2 closest_circ_point(v2 o, v2 co, float r) { return add2(mul2s(normalize2(sub2(o, co)), r), co); }

/* BoundVolume.cpp:213-230 closest_aabb_point: clamp into the rect; a point already inside goes to the nearer edge pair */
static inline v2 closest_aabb_point(v2 o, const canon_rect* w) {
    v2 closest = o;
    closest.x = glm_clamp(closest.x, w->ox - w->w / 2, w->ox + w->w / 2);
    closest.y = glm_clamp(closest.y, w->oy - w->h / 2, w->oy + w->h / 2);
    if (closest.x == o.x && closest.y == o.y) {
        v2 in = sub2(o, V2(w->ox, w->oy));
        if (fabsf(in.x) > fabsf(in.y)) closest.x = w->ox + (in.x > 0 ? +w->w : -w->w) / 2;
        else closest.y = w->oy + (in.y > 0 ? +w->h : -w->h) / 2;
    }
    return closest;
}

/* BVH.cpp:121-130 leaf normal for a Rect leaf: on_bv - on_q */
static inline v2 circ_wall_normal(v2 qo, float r, const canon_rect* w) {
    v2 on_q = closest_circ_point(V2(w->ox, w->oy), qo, r);
    return sub2(closest_aabb_point(on_q, w), on_q);
}

/* BoundVolume.cpp:125-127 Rect::collides, :196-201 circ_rect_collider, BVH.cpp:121-130.  Returns 1 and the push
 * vector on contact. */
int32_t canon_circ_wall_hit(const float* o, float r, const canon_rect* w, float* away) {
    v2 qo = V2(o[0], o[1]), ro = V2(w->ox, w->oy);
    v2 on_q = closest_circ_point(ro, qo, r);
    int q_has_ro = len2_2(sub2(ro, qo)) <= r * r;
    int r_has_onq = fabs(on_q.x - w->ox) <= w->w / 2 && fabs(on_q.y - w->oy) <= w->h / 2;
    away[0] = away[1] = 0.f;
    if (!(q_has_ro || r_has_onq)) return 0;
    v2 n = circ_wall_normal(qo, r, w);
    away[0] = n.x;
    away[1] = n.y;
    return 1;
}

/* PRM.cpp:133-156,169-173 Cspace2d: the agent's bounding square (2r x 2r) summed into every wall
 * (Rect::minkowski_sum_(Rect*), BoundVolume.cpp:55-58), line of sight = every inflated wall is missed */
int32_t canon_cspace_line_of_sight(const canon_rect* walls, int32_t n_walls, float agent_r, const float* a,
                                   const float* b) {
    for (int32_t k = 0; k < n_walls; ++k) {
        canon_rect infl = {walls[k].ox, walls[k].oy, 2 * agent_r + walls[k].w, 2 * agent_r + walls[k].h};
        if (!canon_rect_line_of_sight(&infl, a, b)) return 0;
    }
    return 1;
}

/* LMP.cpp:84-113 lookahead with a multi-waypoint plan.  plan = [n_plan][2], n_plan >= 1; num_done / local_goal in, out */
void canon_lookahead(const canon_rect* walls, int32_t n_walls, float agent_r, const float* pos, const float* plan,
                     int32_t n_plan, int32_t* num_done, float* local_goal) {
    size_t target = (size_t)*num_done;
    if (target < (size_t)n_plan) {
        v2 d = sub2(V2(pos[0], pos[1]), V2(plan[2 * target], plan[2 * target + 1]));
        int at_target = len2_2(d) < 1.f; /* ai::at_dist, ai.cpp:14 */
        if (at_target) ++target;
        size_t next = target + 1;
        if (target < (size_t)n_plan) { local_goal[0] = plan[2 * target]; local_goal[1] = plan[2 * target + 1]; }
        while (1) {
            if (!(next < (size_t)n_plan)) break;
            if (!canon_cspace_line_of_sight(walls, n_walls, agent_r, pos, plan + 2 * next)) break;
            ++target;
            local_goal[0] = plan[2 * target]; local_goal[1] = plan[2 * target + 1];
            next = target + 1;
        }
        *num_done = (int32_t)target;
    } else {
        local_goal[0] = plan[2 * (n_plan - 1)]; local_goal[1] = plan[2 * (n_plan - 1) + 1];
    }
}

/* the record layout of oracle/ref/geom_main.cpp (9 floats in, 12 words out), so one call checks a whole fixture */
void canon_wall_geometry_batch(int32_t K, const float* in, void* out) {
    for (int32_t k = 0; k < K; ++k) {
        const float* p = in + (size_t)k * 9;
        union { float f; int32_t i; } w[12];
        canon_rect rect = {p[0], p[1], p[2], p[3]};
        const float* a = p + 4;
        const float* b = p + 6;
        float r = p[8], away[2], F[2];
        memset(w, 0, sizeof(w));
        w[0].f = canon_rect_intersect(&rect, a, b);
        w[1].i = canon_rect_line_of_sight(&rect, a, b);
        w[2].i = canon_circ_line_of_sight(p, r, a, b);
        w[3].f = canon_ttc_circ_wall(a, r, b, &rect);
        w[4].i = canon_circ_wall_hit(a, r, &rect, away);
        {   /* the harness evaluates the push vector whether or not the collider fires */
            v2 nrm = circ_wall_normal(V2(a[0], a[1]), r, &rect);
            away[0] = nrm.x;
            away[1] = nrm.y;
        }
        w[5].f = away[0];
        w[6].f = away[1];
        w[7].i = canon_cspace_line_of_sight(&rect, 1, r, a, b);
        w[8].i = canon_wall_force(a, r, b, &rect, F);
        w[9].f = F[0];
        w[10].f = F[1];
        memcpy((char*)out + (size_t)k * 48, w, 48);
    }
}

/* ---------- scene catalogue, wall-free scenes only (demo.cpp:48-104,141-206,356-373) ---------- */
static const float GLM_PI_F = 3.14159265358979323846264338327950288f; /* glm::pi<float>() */

static v2 radial(unsigned i, unsigned num, float radius) { /* demo.cpp:48-54 */
    float radian = (float)i / (float)num * 2.f * GLM_PI_F;
    return V2(radius * cosf(radian), radius * sinf(radian));
}
typedef struct { v2 facing, head, spacing; unsigned cols, n; } regiment;
static v2 right_face(regiment r) { /* demo.cpp:63-65: cross(vec3(facing,0), (0,0,1)).xy */
    return V2(r.facing.y * 1.f - 0.f * 0.f, 0.f * 0.f - 1.f * r.facing.x);
}
static regiment center(regiment r) { /* demo.cpp:66-70 */
    float half = (r.spacing.x * (float)(r.cols - 1)) / 2.f;
    r.head = sub2(r.head, mul2s(right_face(r), half));
    return r;
}
static v2 regimented(const regiment* regs, unsigned i) { /* demo.cpp:72-95 */
    unsigned current = 0, last = regs[0].n, j = i;
    while (j >= last) { j -= last; last = regs[++current].n; }
    unsigned col = j % regs[current].cols, row = j / regs[current].cols;
    float u = regs[current].spacing.x * (float)col, v = regs[current].spacing.y * (float)row;
    v2 x = mul2s(right_face(regs[current]), u);
    v2 y = mul2s(V2(-regs[current].facing.x, -regs[current].facing.y), v);
    return add2(add2(x, y), regs[current].head);
}
static int sgnf(float v) { return (0.f < v) - (v < 0.f); }
static v2 axial_goal(v2 pos) { /* demo.cpp:36-41 */
    int axis = sgnf(fabsf(pos.x) - fabsf(pos.y));
    axis = (axis == 0 ? 1 : axis);
    float faxis = (float)axis;
    return V2(-faxis * pos.x, faxis * pos.y);
}

/* demo.cpp:106-124 wall layouts */
static v2 hole_wall(unsigned i, float gap, float scale) { /* make_hole_wall */
    float x = 0;
    float side = ((i % 2) == 0 ? 1.f : -1.f);
    float height = (float)(i / 2);
    float y = side * (scale * (height + .5f) + gap / 2.f);
    return V2(x, y);
}
static v2 hall_wall(unsigned i, float gap, float scale) { /* make_hall; (i + 2 % 4) / 4 == (i + 2) / 4 */
    float side = ((i % 2) == 0 ? 1.f : -1.f);
    float width = (float)((i + 2 % 4) / 4);
    float xside = (i % 4 < 2 ? 1.f : -1.f);
    float x = xside * scale * width;
    float y = side * .5f * (gap + scale);
    return V2(x, y);
}
static const float SCENE6_WALLS[12][2] = {{-3, -3}, {-3, 3}, {3, -3}, {3, 3}, {4, 4}, {4, -4}, {6, -5}, {6, 5}, {-4, 4}, {-6, 5}, {-4, -4}, {-6, -5}};
static const float SCENE11_WALLS[32][2] = {
    {0, 8}, {2, 8}, {4, 8}, {8, 8}, {-6, 8}, {-4, 8}, {-2, 8}, {-8, 0}, {-8, 2}, {-8, 4}, {-8, 6}, {-8, 8}, {-8, -6}, {-8, -4}, {-8, -2},
    {-7.4f, -8}, {1.4f, -8}, {3.4f, -8}, {5.4f, -8}, {-5.4f, -8}, {-3.4f, -8}, {-1.4f, -8}, {8, 0}, {8, 2}, {8, 4}, {8, 6}, {7.4f, -8},
    {8, -6}, {8, -4}, {8, -2}, {6, 3}, {4, 3}};
static const float SCENE18_Y[20] = {1, -1, 3, -3, 6, -5, 8, -8, 10, -10, 12, -12, 14, -14, 16, -16, 18, -18, 20, -20};

int32_t canon_scene_full(uint32_t scene, uint32_t* rng_state, float* pos, float* goal, float* max_duration, int32_t cap,
                         canon_rect* walls, int32_t* n_walls_out, int32_t wall_cap) {
    int32_t n = 0, nw = 0;
    float md = 0.f, ws = 0.f;
    enum { G_OPPOSITE, G_AXIAL, G_MIRROR_X, G_TRANSLATE, G_RANDOM } gk = G_OPPOSITE;
    enum { P_RADIAL, P_REGIMENT, P_LIST, P_RANDOM } pk = P_RADIAL;
    float radius = 0.f;
    v2 translate = V2(0.f, 0.f);
    v2 list[4];
    regiment regs[4];
    memset(regs, 0, sizeof regs);
#define REG2(a, b) do { regs[0] = center(a); regs[1] = center(b); pk = P_REGIMENT; } while (0)
    switch (scene) { /* demo.cpp:141-511 make_scene */
    case 0: n = 60; md = 45.f; radius = 10.f; break;
    case 1: n = 90; md = 70.f; radius = 13.f; break;
    case 2: n = 60; nw = 60; ws = 0.3f; md = 40.f; radius = 10.f; break;
    case 3: n = 60; nw = 1; ws = 4.0f; md = 40.f; radius = 10.f; break;
    case 4: case 5: case 6: { /* demo.cpp:96-104 make_intersecting_bots(14, 7, 6.f) */
        n = 56; md = scene == 5 ? 70.f : 40.f; gk = scene == 5 ? G_OPPOSITE : G_AXIAL;
        if (scene == 6) { nw = 12; ws = 2.0f; }
        float d = 6.f;
        regiment a = {{1, 0}, {-d, 0}, {1.f, 1.f}, 7, 14}, b = {{-1, 0}, {d, 0}, {1.f, 1.f}, 7, 14};
        regiment c = {{0, 1}, {0, -d}, {1.f, 1.f}, 7, 14}, e = {{0, -1}, {0, d}, {1.f, 1.f}, 7, 14};
        regs[0] = center(a); regs[1] = center(b); regs[2] = center(c); regs[3] = center(e);
        pk = P_REGIMENT;
        break;
    }
    case 7: n = 2; nw = 30; ws = 2.0f; md = 70.f; gk = G_MIRROR_X; pk = P_LIST; list[0] = V2(-3.f, 0); list[1] = V2(3.f, 0); break;
    case 8: { n = 4; nw = 30; ws = 2.0f; md = 30.f; gk = G_MIRROR_X;
        regiment a = {{1, 0}, {-3, 0}, {1.f, 1.f}, 2, 2}, b = {{-1, 0}, {3, 0}, {1.f, 1.f}, 2, 2}; REG2(a, b); break; }
    case 9: { n = 40; nw = 30; ws = 2.0f; md = 100.f; gk = G_MIRROR_X;
        regiment a = {{1, 0}, {-3, 0}, {1.f, 1.f}, 5, 20}, b = {{-1, 0}, {3, 0}, {1.f, 1.f}, 5, 20}; REG2(a, b); break; }
    case 10: { n = 40; nw = 30; ws = 2.0f; md = 100.f; gk = G_MIRROR_X;
        regiment a = {{1, 0}, {-3, 0}, {1.f, 1.f}, 5, 5}, b = {{-1, 0}, {3, 0}, {1.f, 1.f}, 5, 35}; REG2(a, b); break; }
    case 11: { n = 70; nw = 32; ws = 2.0f; md = 100.0f; gk = G_TRANSLATE; translate = V2(-20.f, 0.f);
        regiment a = {{0, 1}, {1, 6}, {2.f, 1.5f}, 3, 6}, b = {{-1, 0}, {-4, -2}, {1.f, 1.f}, 8, 64}; REG2(a, b); break; }
    case 12: n = 200; md = 30.f; pk = P_RANDOM; gk = G_RANDOM; break;
    case 13: { n = 40; nw = 18; ws = 2.0f; md = 90.0f; gk = G_MIRROR_X;
        regiment a = {{1, 0}, {-3, 0}, {.5f, .7f}, 5, 20}, b = {{-1, 0}, {3, 0}, {.5f, .7f}, 5, 20}; REG2(a, b); break; }
    case 14: n = 2; md = 30.f; gk = G_MIRROR_X; pk = P_LIST; list[0] = V2(-5.f, .1f); list[1] = V2(3.f, 0); break;
    case 15: n = 3; md = 40.f; gk = G_MIRROR_X; pk = P_LIST; list[0] = V2(-5.f, .1f); list[1] = V2(3.f, 0); list[2] = V2(7.f, 0); break;
    case 16: n = 4; md = 40.f; gk = G_MIRROR_X; pk = P_LIST;
        list[0] = V2(-5.f, .1f); list[1] = V2(3.f, 0); list[2] = V2(7.f, 0); list[3] = V2(-2.f, .1f); break;
    case 17: { n = 4; md = 45.f; gk = G_MIRROR_X;
        regiment a = {{1, 0}, {-5, .1f}, {1.f, 1.f}, 2, 2}, b = {{-1, 0}, {3, 0}, {1.f, 1.f}, 2, 2}; REG2(a, b); break; }
    case 18: { n = 65; nw = 20; ws = 2.0f; md = 60.f; gk = G_TRANSLATE; translate = V2(-15.f, 0.f);
        regiment a = {{-1, 0}, {3, 3}, {1.f, 1.f}, 5, 5}, b = {{-1, 0}, {3, -6}, {1.f, 1.f}, 10, 60}; REG2(a, b); break; }
    default:
        if (scene > 100 && scene >= 900) { /* demo.cpp:490-508: variable-size doorway, num_robos = 40 + (scn - 900) */
            n = 40 + (int32_t)(scene - 900); nw = 30; ws = 2.0f; md = 100.f; gk = G_MIRROR_X;
            unsigned even_size = (unsigned)n / 2, odd_size = even_size + (unsigned)n % 2;
            regiment a = {{1, 0}, {-3, 0}, {1.f, 1.f}, 5, odd_size}, b = {{-1, 0}, {3, 0}, {1.f, 1.f}, 5, even_size};
            REG2(a, b);
        } else return 0; /* 19..100: the reference leaves num_robos uninitialised */
    }
#undef REG2
    if (n > cap) return -n;
    if (nw > 0 && (!walls || nw > wall_cap)) return 0;
    *max_duration = md;
    uint32_t x = *rng_state;
    for (int32_t i = 0; i < n; ++i) {
        v2 p, g;
        if (pk == P_RADIAL) p = radial((unsigned)i, (unsigned)n, radius);
        else if (pk == P_REGIMENT) p = regimented(regs, (unsigned)i);
        else if (pk == P_LIST) p = list[i];
        else { /* scene 12, demo.cpp:358-371: glm::vec2(inside(gen), inside(gen)); g++ evaluates the constructor
                  arguments right to left, so the first draw lands in .y (checked against O-ref) */
            x = canon_lcg_next(x); p.y = canon_uniform(x, -10.f, 10.f);
            x = canon_lcg_next(x); p.x = canon_uniform(x, -10.f, 10.f);
        }
        switch (gk) {
        case G_OPPOSITE: g = V2(-p.x, -p.y); break;
        case G_AXIAL: g = axial_goal(p); break;
        case G_MIRROR_X: g = V2(-p.x, p.y); break;
        case G_TRANSLATE: g = add2(p, translate); break;
        default:
            x = canon_lcg_next(x); g.y = canon_uniform(x, -10.f, 10.f);
            x = canon_lcg_next(x); g.x = canon_uniform(x, -10.f, 10.f);
        }
        pos[2 * i] = p.x; pos[2 * i + 1] = p.y;
        goal[2 * i] = g.x; goal[2 * i + 1] = g.y;
    }
    for (int32_t i = 0; i < nw; ++i) { /* demo.cpp:633-635 create_wall(wall_pos_of(i)): square Rect of side wall_scale */
        v2 p;
        if (scene == 2) { /* demo.cpp:170-174: two draws per wall, right-to-left argument evaluation as above */
            x = canon_lcg_next(x); p.y = canon_uniform(x, -7.f, 7.f);
            x = canon_lcg_next(x); p.x = canon_uniform(x, -7.f, 7.f);
        } else if (scene == 3) p = V2(0.f, 0.f);
        else if (scene == 6) p = V2(SCENE6_WALLS[i][0], SCENE6_WALLS[i][1]);
        else if (scene == 11) p = V2(SCENE11_WALLS[i][0], SCENE11_WALLS[i][1]);
        else if (scene == 13) p = hall_wall((unsigned)i, 3.f, ws);
        else if (scene == 18) p = V2(0.f, SCENE18_Y[i]);
        else p = hole_wall((unsigned)i, 1.f, ws);
        walls[i].ox = p.x; walls[i].oy = p.y; walls[i].w = ws; walls[i].h = ws;
    }
    *rng_state = x;
    if (n_walls_out) *n_walls_out = nw;
    return n;
}

int32_t canon_scene(uint32_t scene, uint32_t* rng_state, float* pos, float* goal, float* max_duration, int32_t cap) {
    canon_rect walls[64];
    int32_t nw = 0;
    uint32_t x = *rng_state;
    float md = 0.f;
    int32_t n = canon_scene_full(scene, &x, pos, goal, &md, cap, walls, &nw, 64);
    if (n <= 0 || nw > 0) return n < 0 ? n : 0; /* walled scenes: canon_scene_full */
    *rng_state = x;
    *max_duration = md;
    return n;
}

/* ---------- render.cpp:87-172 comm_vis colour maps + the picture of render::draw (unpinned: needs OpenGL to build) ---------- */
static void vis_xyz2(const float* c, float* lab) { /* render.cpp:87-96; vec3::b is .z */
    float pi = GLM_PI_F;
    float rx = cosf(2 * pi / 3), ry = sinf(2 * pi / 3), gx = cosf(4 * pi / 3), gy = sinf(4 * pi / 3), bx = cosf(2 * pi), by = sinf(pi);
    lab[1] = rx * c[0] + gx * c[1] + bx * c[2];
    lab[2] = ry * c[0] + gy * c[1] + by * c[2];
    lab[0] = sqrtf(sqrtf(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]) / sqrtf(3.f));
}
static double vis_maybe_cube(double a) { double cube = a * a * a; return cube > 0.008856 ? cube : (a - 16. / 116.) / 7.787; }
static double vis_finish(double a) {
    if (a > 0.0031308) a = 1.055 * pow(a, 1. / 2.4) - 0.055; else a *= 12.92;
    return pow(fmax(0., fmin(1., a)), 2.2);
}
void canon_comm_vis(int32_t vis_f, const float* c, float* rgb) {
    if (vis_f >= 1 && vis_f <= 3) { /* render.cpp:157-170 */
        float v = c[vis_f - 1];
        rgb[0] = v < 0 ? -v : 0.f; rgb[1] = v < 0 ? 0.f : v; rgb[2] = 0.f;
        return;
    }
    float lab[3];
    vis_xyz2(c, lab);
    if (vis_f == 4) { /* render.cpp:98-112 Luv2rgb */
        float r[3] = {0.000f * lab[1] + 1.140f * lab[2], -0.395f * lab[1] - 0.581f * lab[2], 2.032f * lab[1] + 0.000f * lab[2]};
        for (int k = 0; k < 3; ++k) rgb[k] = fmaxf(0.f, fminf(1.f, r[k] + lab[0]));
        return;
    }
    /* render.cpp:114-155 Lab2rgb */
    float L = lab[0] * 100.f, A = lab[1] * 128.f, B = lab[2] * 128.f, X[3];
    X[0] = (float)((L + 16.) / 116.);
    X[1] = (float)(A / 500. + X[0]);
    X[2] = (float)(X[0] - B / 200.);
    X[0] = (float)(0.95047 * vis_maybe_cube(X[0]));
    X[1] = (float)(1.00000 * vis_maybe_cube(X[1]));
    X[2] = (float)(1.08883 * vis_maybe_cube(X[2]));
    static const float m[3][3] = {{3.2406f, -1.5372f, -0.4986f}, {-0.9689f, 1.8758f, 0.0415f}, {0.0557f, -0.2040f, 1.0570f}};
    for (int k = 0; k < 3; ++k) rgb[k] = (float)vis_finish((double)(X[0] * m[k][0] + X[1] * m[k][1] + X[2] * m[k][2]));
}
static uint32_t vis_pack(const float* rgb) {
    uint32_t out = 0xff000000u;
    for (int k = 0; k < 3; ++k) {
        float v = fmaxf(0.f, fminf(1.f, rgb[k]));
        out |= (uint32_t)(v * 255.f + 0.5f) << (8 * k);
    }
    return out;
}
void canon_render(int32_t n, const float* pos, const float* c, int32_t n_walls, const canon_rect* walls, int32_t vis_f,
                  float cam_dist, int32_t width, int32_t height, uint8_t* rgba) {
    const float size = 25.f * cam_dist / 2.f; /* Camera.cpp:86-90 */
    const float body_r = 0.5f * 0.75f * 0.3f, ring_r = 0.5f * 0.3f;
    uint32_t* img = (uint32_t*)rgba;
    for (int r = 0; r < height; ++r)
        for (int u = 0; u < width; ++u) {
            float x = -size + ((float)u + 0.5f) * (2.f * size / (float)width), z = -size + ((float)r + 0.5f) * (2.f * size / (float)height);
            uint32_t px = 0xff000000u | (227u << 16) | (246u << 8) | 253u; /* render.cpp:365 */
            for (int q = 0; q < n_walls; ++q)
                if (fabsf(x - walls[q].ox) <= walls[q].w / 2 && fabsf(z - walls[q].oy) <= walls[q].h / 2) px = 0xff808080u;
            for (int i = 0; i < n; ++i) { /* later in the table wins */
                float dx = x - pos[2 * i], dz = z - pos[2 * i + 1], d2 = dx * dx + dz * dz;
                if (d2 > ring_r * ring_r) continue;
                if (d2 <= body_r * body_r) { float rgb[3]; canon_comm_vis(vis_f, c + 3 * i, rgb); px = vis_pack(rgb); }
                else px = 0xff000000u;
            }
            img[(size_t)r * width + u] = px;
        }
}

/* ---------- comms.config / comms.result (comm.cpp:74-142; pso_opt.py:17-41) ---------- */
int32_t canon_read_config(const char* path, float* m_cv, float* m_f) {
    FILE* f = fopen(path, "r");
    if (!f) return -1;
    char line[4096];
    for (int row = 0; row < 5; ++row) {
        if (!fgets(line, sizeof line, f)) { fclose(f); return -2; } /* reference: stof("") throws */
        char* p = line;
        int ncol = row < 3 ? 8 : 3;
        for (int col = 0; col < ncol; ++col) {
            /* std::getline(sl, item, ' ') + std::stof(item): single-space separated tokens */
            char* e = p;
            while (*e && *e != ' ' && *e != '\n') ++e;
            char save = *e;
            *e = 0;
            char* endp;
            float v = strtof(p, &endp);
            if (endp == p) { fclose(f); return -3; }
            if (row < 3) m_cv[8 * row + col] = v; else m_f[3 * (row - 3) + col] = v;
            *e = save;
            p = (*e == ' ') ? e + 1 : e;
        }
    }
    fclose(f);
    return 0;
}

int32_t canon_write_result(const char* path, float avg_velocity, float group_time) {
    FILE* f = fopen(path, "w");
    if (!f) return -1;
    fprintf(f, "%g\n%g\n", (double)avg_velocity, (double)group_time); /* ostream << float: %g, precision 6 */
    fclose(f);
    return 0;
}
