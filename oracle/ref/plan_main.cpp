// O-ref planner harness: runs the reference's own configuration space, roadmap and A* (compiled unmodified from
// /root/reference: ai/PRM.cpp, ai/GMP.cpp, util/Graph.h) for caller-supplied walls and (start, goal) pairs, the way
// ai::init does (ai.cpp:41-63,107-127), so that the product's host-side planner (lc_plan_paths) can be pinned.
// TEST INFRASTRUCTURE ONLY; this file is ours, everything it calls is the reference's.
//
//   oref_plan < in.bin > out.bin
//
// in : int32 n_walls; float walls[n_walls][4] {ox, oy, w, h}; float agent_radius; int32 n_agents;
//      float start[n_agents][2]; float goal[n_agents][2]
// out: int32 roadmap_vertices, roadmap_directed_edges; then per agent: int32 n_waypoints; float min_time;
//      float waypoints[n][2]
#include "ai/Agent.h"
#include "ai/BoundVolume.h"
#include "ai/GMP.h"
#include "ai/PRM.h"
#include "ai/ai.h"
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <memory>
#include <vector>

template <class T> static bool get(T* p, size_t n) { return fread(p, sizeof(T), n, stdin) == n; }
template <class T> static void put(const T* p, size_t n) { fwrite(p, sizeof(T), n, stdout); }

int main() {
    int32_t n_walls = 0, n_agents = 0;
    float radius = 0;
    if (!get(&n_walls, 1)) return 1;
    std::vector<float> walls(static_cast<size_t>(n_walls) * 4);
    if (n_walls && !get(walls.data(), walls.size())) return 1;
    if (!get(&radius, 1) || !get(&n_agents, 1)) return 1;
    std::vector<float> start(static_cast<size_t>(n_agents) * 2), goal(start.size());
    if (!get(start.data(), start.size()) || !get(goal.data(), goal.size())) return 1;

    std::vector<BoundVolume*> obs;
    for (int32_t k = 0; k < n_walls; ++k)
        obs.push_back(new Rect(glm::vec2(walls[4 * k], walls[4 * k + 1]), walls[4 * k + 2], walls[4 * k + 3]));
    Circ agent_bv(glm::vec2(start[0], start[1]), radius);

    // ai.cpp:25,41-63
    static const float root2 = static_cast<float>(sqrt(2));
    std::unique_ptr<Cspace2d> cs(new Cspace2d(obs, &agent_bv));
    Cspace2d* std_cspace = cs.get();
    glm::vec2 center_2d(0, 0);
    glm::vec2 dim(25 / 2, 25 / 2);
    dim *= 1;
    PRM* std_prm = new PRM(std::move(cs), 5.f * root2, 0.f, glm::vec2(2.5f, 2.5f), 0, center_2d - dim, center_2d + dim, 1.f);

    int32_t nv = static_cast<int32_t>(std_prm->_roadmap->vertex_num());
    int32_t ne = 0;
    std_prm->_roadmap->for_edge([&](NodeId, NodeId) { ++ne; });
    put(&nv, 1);
    put(&ne, 1);

    for (int32_t i = 0; i < n_agents; ++i) {  // ai.cpp:107-127
        Agent a;
        a.start = glm::vec2(start[2 * i], start[2 * i + 1]);
        a.final_goal = glm::vec2(goal[2 * i], goal[2 * i + 1]);
        a.num_done = 0;
        a.cspace = std_cspace;
        a.prm = std_prm;
        a.local_goal = glm::vec2(0, 0);
        a.plan = nullptr;
        GMP::plan_one(a);
        a.min_time = -ai::at_dist;
        glm::vec2 last = a.start;
        for (const auto& p : *a.plan) {
            a.min_time += glm::length(p - last);
            last = p;
        }
        int32_t n = static_cast<int32_t>(a.plan->size());
        put(&n, 1);
        put(&a.min_time, 1);
        for (const auto& p : *a.plan) {
            float xy[2] = {p.x, p.y};
            put(xy, 2);
        }
        delete a.plan;
    }
    return 0;
}
