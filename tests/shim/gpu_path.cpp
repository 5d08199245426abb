// Source-level drop-in, compiled and run: the reference engine's own main loop
// (enginee/src/main.cpp:70-154) with its own demo / comm / ai / physics translation units linked in UNMODIFIED, where the
// three per-frame hot-path calls
//     comm::run();  ai::update_agents();  physics::simulate(1/60);  POOL.all_sync();      (main.cpp:131-144)
// are replaced by ONE call into liblearncomms.so (lc_step) followed by a write-back of the component tables the rest of
// the engine reads (Dynamics, Agent, CommComp, BoundVolume: what those four calls leave behind).  Everything else — scene
// construction (demo::init), config parsing (comm::init), the first plans and min_time (ai::init), done detection, Welford
// statistics and the time-out penalty (demo::run), the result file (comm::terminate) — is the reference's own code running
// on the GPU-produced state.  TEST INFRASTRUCTURE: built by oracle/ref/Makefile into oracle/_ref/oref_gpu_shim where
// /root/reference is present; tests/test_shim.py runs it on the GPU box.
//
//   oref_gpu_shim <scene> <seed> <data_dir>        (data_dir holds comms.config; comms.result / agent_paths.csv are written there)
#include "../../include/learn_comms.h"
#include "Pool.h"
#include "ai/ai.h"
#include "ai/physics.h"
#include "comms/comm.h"
#include "demo/demo.h"
#include "util/Seeder.h"
#include <cstdio>
#include <cstdlib>
#include <sstream>
#include <string>
#include <vector>

namespace {

struct GpuPath {
    lc_sim* sim = nullptr;
    unsigned n = 0, n_walls = 0;
    std::vector<float> pos, vel, vfh, c, facing, lgoal, gdist;
    std::vector<int32_t> num_done;
    long long replans_seen = 0;

    [[noreturn]] static void die(const char* what) {
        fprintf(stderr, "gpu path: %s: %s\n", what, lc_last_error());
        exit(EXIT_FAILURE);
    }

    // Build the GPU simulation from the component tables demo::init + ai::init filled (entity ids: robots 1..n, walls after)
    void attach(const std::string& data_dir, float max_duration) {
        std::vector<float> p0, goal;
        std::vector<lc_rect> walls;
        POOL.for_<Agent>([&](Agent&, Entity&) { ++n; });
        p0.resize(2 * n); goal.resize(2 * n);
        POOL.for_<Agent>([&](Agent& a, Entity& e) {
            p0[2 * (e.id - 1)] = a.start.x; p0[2 * (e.id - 1) + 1] = a.start.y;
            goal[2 * (e.id - 1)] = a.final_goal.x; goal[2 * (e.id - 1) + 1] = a.final_goal.y;
        });
        unsigned total = 0;
        POOL.for_<BoundVolume*>([&](BoundVolume*&, Entity&) { ++total; });
        n_walls = total - n;
        walls.resize(n_walls ? n_walls : 1);
        POOL.for_<BoundVolume*>([&](BoundVolume*& bv, Entity& e) {
            if (e.id <= n) return;
            Rect* r = static_cast<Rect*>(bv);
            walls[e.id - n - 1] = lc_rect{r->_o.x, r->_o.y, r->_w, r->_h};
        });
        float m_cv[24] = {0}, m_f[6] = {0};
#ifndef NO_COMM
        if (lc_read_config((data_dir + "/comms.config").c_str(), m_cv, m_f) != LC_OK) die("lc_read_config");
#endif
        // the engine's table iteration order is the GPU's summation / draw order (quirk Q4)
        std::vector<int32_t> order(n), worder(n_walls ? n_walls : 1);
        if (lc_table_order((int32_t)n, (int32_t)n_walls, order.data(), worder.data()) != LC_OK) die("lc_table_order");
        // state of the global minstd_rand0 after scene construction: the stream the per-frame draws continue (LMP.cpp:261-263)
        std::stringstream ss;
        ss << Seeder().gen();
        unsigned long rng = 1;
        ss >> rng;
        lc_desc d = {};
        d.n = (int32_t)n; d.pos = p0.data(); d.goal = goal.data(); d.radius = 0.15f; d.max_duration = max_duration;
        d.rng_state = (uint32_t)rng; d.m_cv = m_cv; d.m_f = m_f; d.device = 0;
#ifdef NO_COMM
        d.flags = LC_FLAG_NO_COMM;
#endif
        d.n_walls = (int32_t)n_walls; d.walls = walls.data(); d.table_order = order.data(); d.wall_order = worder.data();
        if (lc_create(&d, &sim) != LC_OK) die("lc_create");
        pos.resize(2 * n); vel.resize(2 * n); vfh.resize(2 * n); c.resize(3 * n); facing.resize(2 * n); lgoal.resize(2 * n);
        gdist.resize(n); num_done.resize(n);
    }

    // comm::run + ai::update_agents + physics::simulate + POOL.all_sync, on the GPU; then the tables as those calls leave them
    void frame() {
        if (lc_step(sim, 1) != LC_OK) die("lc_step");
        if (lc_get_state(sim, pos.data(), vel.data(), vfh.data(), c.data(), facing.data(), lgoal.data(), gdist.data(), num_done.data()) != LC_OK)
            die("lc_get_state");
        bool plans_changed = false;
        if (n_walls) {  // replans (ai.cpp:164-168) swap an agent's waypoint list: mirror it, demo::run's done() reads its size
            int64_t rp = 0;
            if (lc_get_walled(sim, nullptr, nullptr, &rp, nullptr) != LC_OK) die("lc_get_walled");
            plans_changed = rp != replans_seen;
            replans_seen = rp;
        }
        POOL.for_<Agent>([&](Agent& a, Entity& e) {
            const unsigned i = e.id - 1;
            Dynamics& d = *POOL.get<Dynamics>(e);
            d.pos = glm::vec3(pos[2 * i], 0, pos[2 * i + 1]);
            d.vel = glm::vec3(vel[2 * i], 0, vel[2 * i + 1]);
            d.vel_forehalf = glm::vec3(vfh[2 * i], 0, vfh[2 * i + 1]);
            d.force = glm::vec3(0);
            a.start = glm::vec2(pos[2 * i], pos[2 * i + 1]);             // all_sync, Pool.cpp:86-88
            (*POOL.get<BoundVolume*>(e))->_o = a.start;                   // Pool.cpp:83-85
            POOL.get<Transform>(e)->set_pos(d.pos);                       // Pool.cpp:77-79
            a.local_goal = glm::vec2(lgoal[2 * i], lgoal[2 * i + 1]);
            a.goal_dist = gdist[i];
            a.num_done = num_done[i];
            if (plans_changed) {
                float wp[2 * 64];
                int32_t k = 0;
                if (lc_get_plan(sim, (int32_t)i, wp, 64, &k) != LC_OK) die("lc_get_plan");
                a.plan->clear();
                for (int q = 0; q < k; ++q) a.plan->push_back(glm::vec2(wp[2 * q], wp[2 * q + 1]));
            }
#ifndef NO_COMM
            CommComp& cc = *POOL.get<CommComp>(e);
            for (int k = 0; k < 3; ++k) cc.c[k] = c[3 * i + k];
            cc.facing = glm::vec2(facing[2 * i], facing[2 * i + 1]);
#endif
        });
    }

    void finish() {
        float av = 0.f;
        lc_get_scores(sim, &av, nullptr, nullptr, nullptr, nullptr);
        physics::avg_velocity = av;   // physics.cpp:72-76's running value (quirk Q6), computed where the speeds are
        lc_destroy(sim);
    }
};

}  // namespace

int main(int argc, char** argv) {
    if (argc < 4) {
        fprintf(stderr, "usage: oref_gpu_shim <scene> <seed> <data_dir>\n");
        return 2;
    }
    const unsigned scene = (unsigned)atoi(argv[1]);
    const uint64_t seed = strtoull(argv[2], nullptr, 10);
    const std::string data_dir = argv[3];
    // main.cpp:69-95, the reference's own set-up
    Seeder s;
    s.seed(seed);
    demo::init(scene, data_dir);
    comm::init(data_dir);
    ai::init();
    physics::init();
    physics::prewarm(1.f / 60.f);
    POOL.all_sync();
    // Scene::max_duration is private to demo.cpp; the product's catalogue carries the same table (checked against O-ref)
    float max_duration = 0.f;
    {
        std::vector<float> p(2 * 4096), g(2 * 4096);
        std::vector<lc_rect> w(64);
        int32_t nw = 0;
        if (lc_scene_full(scene, seed, p.data(), g.data(), 4096, &max_duration, nullptr, w.data(), 64, &nw) <= 0) {
            fprintf(stderr, "scene %u is not in the catalogue\n", scene);
            return EXIT_FAILURE;
        }
    }
    GpuPath gpu;
    gpu.attach(data_dir, max_duration);
    // main.cpp:98-150 with the hot path swapped
    unsigned total_frames = 0;
    bool all_done = false;
    while (!all_done) {
        ++total_frames;
        all_done = demo::run(1.f / 60.f, (double)total_frames / 60.0, total_frames);
        gpu.frame();
    }
    gpu.finish();
    fprintf(stderr, "\nFRAMES: %u\n", total_frames);
    // main.cpp:152-154
    demo::terminate();
    comm::terminate();
    ai::terminate();
    return 0;
}
