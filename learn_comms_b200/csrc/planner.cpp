// Host-side global planner for walled scenes: configuration space, roadmap and A* that give every agent
// its waypoint list before the first frame and at every replan (ai.cpp:41-63,107-127,161-166).  The reference plans on
// the CPU at init time too; the per-frame consumers of the plan (lookahead, sight lines, wall forces, push-out) are the
// device functions in walls.cuh.
//
// Which of several equally short paths A* returns depends on the order in which ties surface, and the reference lets
// libstdc++'s hash containers decide that (Graph.h:29-31 keeps vertices and adjacency in unordered_map / unordered_set,
// GMP.cpp:26 a std::priority_queue).  Collinear roadmap points make such ties the rule, not the exception, so the
// roadmap here is kept in the same standard containers and mutated by the same sequence of insertions and erasures:
// that, not a re-implementation of their internals, is what makes the waypoint lists come out identical.
#include "../../include/learn_comms.h"
#include "walls.cuh"
#include <cmath>
#include <cstring>
#include <queue>
#include <tuple>
#include <unordered_map>
#include <unordered_set>
#include <vector>

int lc_set_error(int code, const char* msg);  // capi.cu

struct lc_planner {
    struct P { float x, y; };
    std::vector<lcw::Wall> walls;  // as given
    float agent_r = 0.f;
    std::unordered_map<uint32_t, P> vertex;                        // Graph::_vertices
    std::unordered_map<uint32_t, std::unordered_set<uint32_t>> adj;  // Graph::_edges
    uint32_t next_id = 0;

    static float dist(P a, P b) {  // glm::distance -> length(b - a) -> sqrt(dot)
        float dx = b.x - a.x, dy = b.y - a.y;
        return sqrtf(dx * dx + dy * dy);
    }
    // Cspace2d::line_of_sight (PRM.cpp:169-173) over the walls inflated by the agent's bounding square (:133-156)
    bool sight(P a, P b) const {
        for (const lcw::Wall& w : walls)
            if (!lcw::cspace_line_of_sight(w, agent_r, a.x, a.y, b.x, b.y)) return false;
        return true;
    }
    // Cspace2d::collides (PRM.cpp:160-167) with Rect::collides (BoundVolume.cpp:125-127)
    bool blocked(P p) const {
        for (const lcw::Wall& w : walls) {
            float iw = 2 * agent_r + w.w, ih = 2 * agent_r + w.h;
            if (fabsf(p.x - w.ox) <= iw / 2 && fabsf(p.y - w.oy) <= ih / 2) return true;
        }
        return false;
    }
    uint32_t add_vertex(P p) {
        vertex[next_id] = p;
        return next_id++;
    }
    void del_vertex(uint32_t v) {  // Graph.h:47-57
        vertex.erase(v);
        auto it = adj.find(v);
        if (it != adj.end())
            for (uint32_t u : it->second) adj[u].erase(v);
        adj.erase(v);
    }
    // GMP.cpp:85-91: temporary start / goal vertices see every vertex in line of sight, whatever the distance
    void connect_to_all(uint32_t v) {
        for (const auto& kv : vertex) {
            uint32_t u = kv.first;
            if (v != u && sight(vertex[v], vertex[u])) {
                adj[v].insert(u);
                adj[u].insert(v);
            }
        }
    }

    // PRM.cpp:40-105 with ai.cpp:53-63's parameters: bin_samp = 0, so the roadmap is the eight points 0.3 m outside
    // each inflated wall (corners and face midpoints) that lie in free space ...
    void sample() {
        for (const lcw::Wall& w : walls) {
            float iw = 2 * agent_r + w.w, ih = 2 * agent_r + w.h;
            P h = {(float)((double)(w.ox + iw / 2) + 0.3), (float)((double)(w.oy + ih / 2) + 0.3)};
            P l = {(float)((double)(w.ox - iw / 2) - 0.3), (float)((double)(w.oy - ih / 2) - 0.3)};
            float mx = (h.x + l.x) * .5f, my = (h.y + l.y) * .5f;
            const P pts[8] = {{h.x, h.y}, {h.x, l.y}, {mx, h.y}, {mx, l.y}, {l.x, h.y}, {l.x, l.y}, {h.x, my}, {l.x, my}};
            for (const P& p : pts)
                if (!blocked(p)) add_vertex(p);
        }
    }
    // ... joined when closer than 5 sqrt(2) m and in line of sight (PRM.cpp:8-38)
    void connect() {
        const float threshold = 5.f * static_cast<float>(sqrt(2));
        for (const auto& kv : vertex) {
            uint32_t v = kv.first;
            P from = kv.second;
            std::unordered_set<uint32_t> near;
            for (const auto& ku : vertex)
                if (ku.first != v && dist(from, ku.second) < threshold) near.insert(ku.first);
            for (uint32_t u : near)
                if (sight(vertex[v], vertex[u])) adj[v].insert(u);
        }
    }

    // GMP.cpp:7-78 find_path_astar with e = 1: waypoints from the first after s to g inclusive; empty if unreachable
    std::vector<P> astar(uint32_t s, uint32_t g) {
        typedef std::tuple<P, float, uint32_t> Item;
        auto worse = [](Item a, Item b) { return std::get<1>(a) > std::get<1>(b); };
        std::priority_queue<Item, std::vector<Item>, decltype(worse)> open(worse);
        std::unordered_map<uint32_t, uint32_t> parent;
        std::unordered_map<uint32_t, float> g_cost;
        const P goal = vertex[g];
        open.push(Item(vertex[s], 0.f, s));
        parent[s] = s;
        g_cost[s] = 0.f;
        while (!open.empty()) {
            uint32_t cur = std::get<2>(open.top());
            open.pop();
            P cur_p = vertex[cur];
            if (cur == g) {
                std::vector<P> path;
                for (uint32_t v = g; v != s; v = parent[v]) path.insert(path.begin(), vertex[v]);
                return path;
            }
            auto e = adj.find(cur);
            if (e == adj.end()) continue;
            for (uint32_t nb : e->second) {
                P nb_p = vertex[nb];
                float alt = g_cost[cur] + dist(nb_p, cur_p);
                if (!g_cost.count(nb) || alt < g_cost[nb]) {
                    g_cost[nb] = alt;
                    open.push(Item(nb_p, alt + 1.f * dist(goal, nb_p), nb));
                    parent[nb] = cur;
                }
            }
        }
        return {};
    }
};

extern "C" {

int32_t lc_planner_create(int32_t n_walls, const lc_rect* walls, float agent_radius, lc_planner** out) {
    if (!out || n_walls < 0 || (n_walls > 0 && !walls) || !(agent_radius > 0.f))
        return lc_set_error(LC_ERR_ARG, "lc_planner_create: bad arguments");
    lc_planner* p = nullptr;
    try {  // no exception crosses the C ABI
        p = new lc_planner();
        p->agent_r = agent_radius;
        for (int32_t k = 0; k < n_walls; ++k) p->walls.push_back({walls[k].ox, walls[k].oy, walls[k].w, walls[k].h});
        p->sample();
        p->connect();
    } catch (...) {
        delete p;
        *out = nullptr;
        return lc_set_error(LC_ERR_CAPACITY, "lc_planner_create: host allocation failed");
    }
    *out = p;
    return LC_OK;
}

void lc_planner_destroy(lc_planner* p) { delete p; }

// Independent copy with the same roadmap STATE (vertex ids handed out so far, container layouts): many scenarios that share
// a wall layout and an initial planning history (a PSO batch) need one roadmap each only from their first replan on.
// libstdc++ copies an unordered container node by node in iteration order into the same number of buckets, so the copy
// iterates — and breaks A* ties — exactly like the original (tests/test_planner.py holds a clone against a fresh planner).
int32_t lc_planner_clone(const lc_planner* p, lc_planner** out) {
    if (!p || !out) return lc_set_error(LC_ERR_ARG, "lc_planner_clone: null argument");
    try {
        *out = new lc_planner(*p);
    } catch (...) {
        *out = nullptr;
        return lc_set_error(LC_ERR_CAPACITY, "lc_planner_clone: host allocation failed");
    }
    return LC_OK;
}

int32_t lc_planner_roadmap_size(lc_planner* p, int32_t* vertices, int32_t* directed_edges) {
    if (!p) return lc_set_error(LC_ERR_ARG, "lc_planner_roadmap_size: null planner");
    size_t e = 0;
    for (const auto& kv : p->adj) e += kv.second.size();
    if (vertices) *vertices = (int32_t)p->vertex.size();
    if (directed_edges) *directed_edges = (int32_t)e;
    return LC_OK;
}

int32_t lc_planner_plan(lc_planner* p, const float* start, const float* goal, int32_t cap, float* waypoints,
                        int32_t* n_waypoints, float* min_time) {
    if (!p || !start || !goal || cap < 0 || (cap > 0 && !waypoints) || !n_waypoints)
        return lc_set_error(LC_ERR_ARG, "lc_planner_plan: bad arguments");
    // GMP.cpp:93-110 plan_one: start and goal join the roadmap for the duration of the search
    std::vector<lc_planner::P> path;
    try {
        uint32_t s = p->add_vertex({start[0], start[1]});
        uint32_t g = p->add_vertex({goal[0], goal[1]});
        p->connect_to_all(s);
        p->connect_to_all(g);
        path = p->astar(s, g);
        p->del_vertex(s);
        p->del_vertex(g);
    } catch (...) {
        return lc_set_error(LC_ERR_CAPACITY, "lc_planner_plan: host allocation failed");
    }
    *n_waypoints = (int32_t)path.size();
    if (min_time) {  // ai.cpp:118-126: path length minus the untimed last metre (ai::at_dist)
        float t = -1.f;
        lc_planner::P last = {start[0], start[1]};
        for (const lc_planner::P& q : path) {
            float dx = q.x - last.x, dy = q.y - last.y;
            t += sqrtf(dx * dx + dy * dy);
            last = q;
        }
        *min_time = t;
    }
    if ((int32_t)path.size() > cap) return lc_set_error(LC_ERR_CAPACITY, "lc_planner_plan: waypoint buffer too small");
    for (size_t k = 0; k < path.size(); ++k) {
        waypoints[2 * k] = path[k].x;
        waypoints[2 * k + 1] = path[k].y;
    }
    return LC_OK;
}

}  // extern "C"
