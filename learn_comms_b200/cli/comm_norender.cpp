// comm_norender / nocomm_norender — drop-in for the reference executables that pso_opt.py:245-267 launches:
//     <build>/bin/comm_norender <scene> --seed <u64> --data <dir>
// Same argument grammar as enginee/src/main.cpp:53-69 + io.cpp:63-125 (options --help, --record u, --data dir,
// --split n, --seed u; one positional `scene`; misuse -> message + usage on stderr, exit 1), same files:
// every catalogue scene, walled ones included; reads <project>data/<dir>/comms.config (comm.cpp:74-131), writes <project>data/<dir>/comms.result (comm.cpp:136-142)
// and <project>data/<dir>/agent_paths.csv (demo.cpp:678-684; LEARN_COMMS_NO_PATHS=1 skips it and runs the rollout in one call).
// The frame loop itself (main.cpp:98-150) runs on the GPU through the C ABI; there is no CPU fallback: without a CUDA
// device the process exits 1 with the library's error text.  Build: tools/build_cli.sh (g++ + liblearncomms.so).
// PROJECT_DIR: compile-time in the reference (CMake); here -DLC_PROJECT_DIR or the LEARN_COMMS_PROJECT_DIR variable.
#include "../../include/learn_comms.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

#ifndef LC_PROJECT_DIR
#define LC_PROJECT_DIR "./"
#endif

static std::string g_usage;
[[noreturn]] static void usage_error(const char* msg) {  // io.cpp:42-45
    fprintf(stderr, "%s\n%s\n", msg, g_usage.c_str());
    exit(EXIT_FAILURE);
}
[[noreturn]] static void lib_error(const char* what) {
    fprintf(stderr, "%s: %s\n", what, lc_last_error());
    exit(EXIT_FAILURE);
}

int main(int argc, char** argv) {
    std::string data_dir;
    unsigned long long seed = 0;
    long scene = -1;
    bool seen_scene = false, seen[5] = {false, false, false, false, false};
    const char* names[5] = {"data", "help", "record", "seed", "split"};  // std::map order of the reference's usage line
    const char* args[5] = {"dir", "", "u", "u", "n"};
    g_usage = std::string("Usage: ") + argv[0];
    for (int k = 0; k < 5; ++k) g_usage += std::string(" [--") + names[k] + (args[k][0] ? std::string(" ") + args[k] : "") + "]";
    g_usage += " scene";
    for (int i = 1; i < argc; ++i) {
        std::string a(argv[i]);
        int k = -1;
        if (a.compare(0, 2, "--") == 0)
            for (int q = 0; q < 5; ++q)
                if (a.substr(2) == names[q]) k = q;
        if (k >= 0) {
            if (seen[k]) usage_error("no double options");
            seen[k] = true;
            if (k == 1) { printf("%s\n", g_usage.c_str()); return EXIT_SUCCESS; }
            if (i + 1 >= argc) usage_error("no arg to read");
            std::string v(argv[++i]);
            if (v.find('-') != std::string::npos) usage_error("bad arg");
            try {
                if (k == 0) data_dir = v;
                else if (k == 3) seed = std::stoull(v);
                else (void)std::stoi(v);  // --record / --split: accepted and ignored (no renderer on this path)
            } catch (...) { usage_error("bad arg"); }
        } else {
            if (seen_scene) usage_error("no further positionals");
            if (a.find("--") != std::string::npos) usage_error("unknown option");
            try { scene = std::stoi(a); } catch (...) { usage_error("bad arg"); }
            seen_scene = true;
        }
    }
    if (!seen_scene) usage_error("missing required positional");

    const char* env = getenv("LEARN_COMMS_PROJECT_DIR");
    std::string project = env ? env : LC_PROJECT_DIR;
    if (!project.empty() && project.back() != '/') project += '/';
    data_dir = project + "data/" + data_dir;  // main.cpp:68

    // demo::init (scene catalogue) + Seeder::seed (main.cpp:69-70)
    std::vector<float> pos(2 * 4096), goal(2 * 4096);
    std::vector<lc_rect> walls(64);
    float max_duration = 0.f;
    uint32_t rng = 0;
    int32_t n_walls = 0;
    int n = lc_scene_full((uint32_t)scene, seed, pos.data(), goal.data(), 4096, &max_duration, &rng, walls.data(), 64, &n_walls);
    if (n <= 0) {
        fprintf(stderr, "scene %ld is not in the catalogue (0-18 and 900+k, demo.cpp:141-511)\n", scene);
        return EXIT_FAILURE;
    }
    float m_cv[24] = {0}, m_f[6] = {0};
    uint32_t flags = 0;
#ifdef NO_COMM
    flags |= LC_FLAG_NO_COMM;  // main.cpp:130-132: comm::run compiled out; comm::init still runs but reads nothing
#else
    if (lc_read_config((data_dir + "/comms.config").c_str(), m_cv, m_f) != LC_OK) {
        // not HEAD's 5 x 8 format: one of the legacy "3 5" files (optimizer.py:20-27; comms_good.config ...)?  HEAD's own
        // parser aborts on those; the drop-in runs them with the plain feature set they were written for
        uint32_t feat = 0;
        if (lc_read_config_legacy((data_dir + "/comms.config").c_str(), m_cv, m_f, &feat) == LC_OK) {
            flags |= feat;
        } else {
            fprintf(stderr, "gg! Failed to open file %s/comms.config\n", data_dir.c_str());  // io.cpp:130
            return EXIT_FAILURE;  // the reference dereferences an empty optional / throws from stof here
        }
    }
#endif
    lc_desc d;
    memset(&d, 0, sizeof d);
    d.n = n; d.pos = pos.data(); d.goal = goal.data(); d.radius = 0.15f; d.max_duration = max_duration;
    d.rng_state = rng; d.m_cv = m_cv; d.m_f = m_f; d.flags = flags; d.device = 0;
    // the reference's component-table iteration order (quirk Q4): which RNG draw an agent gets, summation and Welford order
    std::vector<int32_t> order((size_t)n), wall_order((size_t)(n_walls > 0 ? n_walls : 1));
    if (lc_table_order(n, n_walls, order.data(), wall_order.data()) != LC_OK) lib_error("lc_table_order");
    d.table_order = order.data();
    d.n_walls = n_walls; d.walls = walls.data(); d.wall_order = wall_order.data();
    lc_sim* sim = nullptr;
    if (lc_create(&d, &sim) != LC_OK) lib_error("lc_create");
    int32_t frames = 0;
    if (getenv("LEARN_COMMS_NO_PATHS")) {
        if (lc_run_to_completion(sim, &frames) != LC_OK) lib_error("lc_run_to_completion");
    } else {
        // demo.cpp:616,678-684: one row per agent and frame, written by demo::run BEFORE the frame is simulated:
        //   id,groupID(0),pos.x,pos.z,vel.x,vel.z,radius,dt*(frame-1)     (default ostream float formatting)
        // id: the ECS entity id = creation index + 1 (demo.cpp:628-631); rows of a frame come in the reference's table
        // iteration order (POOL.for_<Agent>, Pool.cpp:156-160).
        std::ofstream paths(data_dir + "/agent_paths.csv", std::ios::binary);
        if (!paths) { fprintf(stderr, "cannot write %s/agent_paths.csv\n", data_dir.c_str()); return EXIT_FAILURE; }
        std::vector<float> p(2 * (size_t)n), v(2 * (size_t)n);
        const float dt = static_cast<float>(1.0 / 60.0);
        for (;;) {
            if (lc_get_state(sim, p.data(), v.data(), nullptr, nullptr, nullptr, nullptr, nullptr, nullptr) != LC_OK) lib_error("lc_get_state");
            const float t = dt * static_cast<float>(frames);
            for (int k = 0; k < n; ++k) {
                const int i = order[k];
                paths << (i + 1) << ",0," << p[2 * i] << "," << p[2 * i + 1] << "," << v[2 * i] << "," << v[2 * i + 1] << ","
                      << d.radius << "," << t << "\n";
            }
            if (lc_step(sim, 1) != LC_OK) lib_error("lc_step");
            ++frames;
            int32_t fin = 0;
            if (lc_finished(sim, &fin, nullptr) != LC_OK) lib_error("lc_finished");
            if (fin) break;
        }
    }
    float avg_velocity = 0.f, group_time = 0.f, run_avg = 0.f, run_std = 0.f;
    int32_t total = 0;
    int rc = lc_get_scores(sim, &avg_velocity, &group_time, &run_avg, &run_std, &total);
    if (rc != LC_OK && rc != LC_ERR_STATE) lib_error("lc_get_scores");
    fprintf(stderr, "\nFRAMES: %d\nNUM DONE: %d\n", frames, total);  // main.cpp:118-119, demo.cpp:686
    {   // demo.cpp:662: one "t:%f u:%f s:%f" line per finisher, in the order the Welford update saw them
        std::vector<float> log(3 * (size_t)(total > 0 ? total : 1));
        int32_t k = 0;
        if (lc_get_finish_log(sim, log.data(), total, &k) != LC_OK) lib_error("lc_get_finish_log");
        for (int i = 0; i < k; ++i) printf("t:%f u:%f s:%f\n", log[3 * i], log[3 * i + 1], log[3 * i + 2]);
    }
    if (lc_write_result((data_dir + "/comms.result").c_str(), avg_velocity, group_time) != LC_OK) {
        fprintf(stderr, "cannot write %s/comms.result\n", data_dir.c_str());
        return EXIT_FAILURE;
    }
    lc_destroy(sim);
    return EXIT_SUCCESS;
}
