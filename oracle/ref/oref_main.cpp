// O-ref driver: runs the reference's own systems (compiled unmodified from /root/reference) through the frame loop of
// enginee/src/main.cpp:70-154 and dumps the per-frame component state, so that the O-canon restatement can be pinned
// against the real code.  TEST INFRASTRUCTURE ONLY; this file is ours, everything it calls is the reference's.
//
//   oref <scene> <seed> <data_dir> <max_frames> <dump_every> <dump_file|-> [time]
//
// data_dir must hold comms.config (comm::init, comm.cpp:74-121); the reference writes agent_paths.csv and
// comms.result there.  The dump is little-endian binary:
//   int32 magic 0x4f524546, n, n_walls, frames_dumped_capacity
//   int32 order[n]            iteration order of POOL.for_<Agent> as entity ids (the unordered_map order, quirk Q4)
//   float pos0[n][2], goal[n][2], min_time[n]          state after ai::init + prewarm + all_sync, entity-id order
//   int32 wall_order[n_walls]  iteration order of the walls in POOL.for_<BoundVolume*> (0-based wall index);
//   float walls[n_walls][4]    Rect origin x, z, width, height in creation order (demo.cpp:633-635)
//   per dumped frame: int32 frame; float pos[n][2], vel[n][2], vfh[n][2], c[n][3], facing[n][2], local_goal[n][2],
//                     goal_dist[n]; int32 num_done[n]; float run_avg, run_std, avg_velocity; int32 total_num_done, all_done
//   trailer: int32 -1, frames_executed; float avg_velocity, group_time, run_avg, run_std; int32 total_num_done
#include "Pool.h"
#include "ai/ai.h"
#include "ai/physics.h"
#include "comms/comm.h"
#include "demo/demo.h"
#include "util/Seeder.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace demo {
extern int total_num_done;  // demo.cpp:650 (namespace-scope, external linkage)
}

static FILE* g_out = nullptr;
template <typename T> static void put(const T* p, size_t n) {
    if (g_out) fwrite(p, sizeof(T), n, g_out);
}
static void put_i(int32_t v) { put(&v, 1); }
static void put_f(float v) { put(&v, 1); }

struct Row {
    float pos[2], vel[2], vfh[2], c[3], facing[2], lgoal[2], goal_dist, goal[2], min_time;
    int32_t num_done;
};

// component state by entity id (robots are entities 1..n, demo.cpp:629-631)
static std::vector<Row> snapshot(unsigned n) {
    std::vector<Row> rows(n);
    POOL.for_<Agent>([&](Agent& a, Entity& e) {
        if (e.id < 1 || e.id > n) return;
        Row& r = rows[e.id - 1];
        Dynamics& d = *POOL.get<Dynamics>(e);
        CommComp* cp = POOL.get<CommComp>(e);  // absent in the nocomm flavour
        r.pos[0] = d.pos.x; r.pos[1] = d.pos.z;
        r.vel[0] = d.vel.x; r.vel[1] = d.vel.z;
        r.vfh[0] = d.vel_forehalf.x; r.vfh[1] = d.vel_forehalf.z;
        for (int k = 0; k < 3; ++k) r.c[k] = cp ? cp->c[k] : 0.f;
        r.facing[0] = cp ? cp->facing.x : 0.f; r.facing[1] = cp ? cp->facing.y : 0.f;
        r.lgoal[0] = a.local_goal.x; r.lgoal[1] = a.local_goal.y;
        r.goal_dist = a.goal_dist;
        r.goal[0] = a.final_goal.x; r.goal[1] = a.final_goal.y;
        r.min_time = a.min_time;
        r.num_done = a.num_done;
    });
    return rows;
}

int main(int argc, char** argv) {
    if (argc < 7) {
        fprintf(stderr, "usage: oref <scene> <seed> <data_dir> <max_frames> <dump_every> <dump_file|-> [time]\n");
        return 2;
    }
    const unsigned scene = (unsigned)atoi(argv[1]);
    const uint64_t seed = strtoull(argv[2], nullptr, 10);
    const std::string data_dir = argv[3];
    const unsigned max_frames = (unsigned)atoi(argv[4]);
    const unsigned dump_every = (unsigned)atoi(argv[5]);
    if (strcmp(argv[6], "-") != 0) g_out = fopen(argv[6], "wb");
    const bool timing = argc > 7 && !strcmp(argv[7], "time");

    // main.cpp:69-95
    Seeder s;
    s.seed(seed);
    demo::init(scene, data_dir);
    // Optional: start every robot from the positions in $OREF_POS_FILE (float32 [n][2], entity-id order) instead of the
    // scene's own — the 1-ulp-perturbation ensembles that calibrate the rollout tolerance (tools/make_golden_ensemble.py).
    // Goals keep the scene's values; every mirror of the position that demo::create_robo / POOL.all_sync maintain
    // (Dynamics.pos, Agent.start, BoundVolume._o, Transform, CommComp.facing = goal - start, demo.cpp:563-581) follows.
    if (const char* pf = getenv("OREF_POS_FILE")) {
        FILE* f = fopen(pf, "rb");
        if (!f) { fprintf(stderr, "oref: cannot open %s\n", pf); return 2; }
        std::vector<float> xy;
        float v[2];
        while (fread(v, sizeof(float), 2, f) == 2) { xy.push_back(v[0]); xy.push_back(v[1]); }
        fclose(f);
        const size_t m = xy.size() / 2;
        POOL.for_<Agent>([&](Agent& a, Entity& e) {
            if (e.id < 1 || e.id > m) return;
            const glm::vec2 p(xy[2 * (e.id - 1)], xy[2 * (e.id - 1) + 1]);
            Dynamics& d = *POOL.get<Dynamics>(e);
            d.pos = glm::vec3(p.x, 0, p.y);
            a.start = p;
            (*POOL.get<BoundVolume*>(e))->_o = p;
            POOL.get<Transform>(e)->set_pos(d.pos);
#ifndef NO_COMM
            POOL.get<CommComp>(e)->facing = a.final_goal - a.start;
#endif
        });
    }
    comm::init(data_dir);
    ai::init();
    physics::init();
    unsigned total_frames = 0;
    bool all_done = false;
    physics::prewarm(1.f / 60.f);
    POOL.all_sync();

    unsigned n = 0, n_other = 0;
    std::vector<int32_t> order;
    POOL.for_<Agent>([&](Agent&, Entity& e) { order.push_back((int32_t)e.id); ++n; });
    POOL.for_<BoundVolume*>([&](BoundVolume*&, Entity&) { ++n_other; });
    put_i(0x4f524546); put_i((int32_t)n); put_i((int32_t)(n_other - n)); put_i((int32_t)max_frames);
    put(order.data(), order.size());
    {
        auto rows = snapshot(n);
        for (auto& r : rows) put(r.pos, 2);
        for (auto& r : rows) put(r.goal, 2);
        for (auto& r : rows) put_f(r.min_time);
        std::vector<int32_t> worder;
        std::vector<float> wrect(4 * (size_t)(n_other - n), 0.f);
        POOL.for_<BoundVolume*>([&](BoundVolume*& bv, Entity& e) {
            if (e.id <= n) return;
            worder.push_back((int32_t)(e.id - n - 1));
            Rect* r = static_cast<Rect*>(bv);
            float* w = &wrect[4 * (size_t)(e.id - n - 1)];
            w[0] = r->_o.x; w[1] = r->_o.y; w[2] = r->_w; w[3] = r->_h;
        });
        put(worder.data(), worder.size());
        put(wrect.data(), wrect.size());
    }

    auto t0 = std::chrono::steady_clock::now();
    // main.cpp:98-150
    while (!all_done && total_frames < max_frames) {
        ++total_frames;
        all_done = demo::run(1.f / 60.f, (double)total_frames / 60.0, total_frames);
#ifndef NO_COMM
        comm::run();
#endif
        ai::update_agents();
        physics::simulate(1.f / 60.f);
        POOL.all_sync();
        if (g_out && dump_every && (total_frames % dump_every == 0 || all_done)) {
            auto rows = snapshot(n);
            put_i((int32_t)total_frames);
            for (auto& r : rows) put(r.pos, 2);
            for (auto& r : rows) put(r.vel, 2);
            for (auto& r : rows) put(r.vfh, 2);
            for (auto& r : rows) put(r.c, 3);
            for (auto& r : rows) put(r.facing, 2);
            for (auto& r : rows) put(r.lgoal, 2);
            for (auto& r : rows) put_f(r.goal_dist);
            for (auto& r : rows) put_i(r.num_done);
            put_f(demo::run_avg_time); put_f(demo::run_std_time); put_f(physics::avg_velocity);
            put_i(demo::total_num_done); put_i(all_done ? 1 : 0);
        }
    }
    double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    put_i(-1); put_i((int32_t)total_frames);
    put_f(physics::avg_velocity); put_f(demo::run_avg_time + demo::run_std_time * 3);
    put_f(demo::run_avg_time); put_f(demo::run_std_time); put_i(demo::total_num_done);
    if (g_out) fclose(g_out);
    if (timing) fprintf(stderr, "OREF_TIME n=%u frames=%u seconds=%.6f agent_steps_per_s=%.1f\n", n, total_frames, secs,
                        (double)n * total_frames / secs);
    // main.cpp:152-154
    demo::terminate();
    comm::terminate();
    ai::terminate();
    return 0;
}
