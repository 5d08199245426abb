// Host-side pieces of the C ABI that never touch the GPU: the comms.config / comms.result file formats
// (comm.cpp:74-142, pso_opt.py:17-41) and the wall-free part of the scene catalogue (demo.cpp:141-511).
// Compiled with -ffp-contract=off: initial positions must be bit-identical to the reference's float math.
#include "../../include/learn_comms.h"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <functional>
#include <sstream>
#include <algorithm>
#include <limits>
#include <string>
#include <unordered_map>
#include <vector>

namespace {

struct Vec2 {
    float x, y;
};
inline Vec2 operator+(Vec2 a, Vec2 b) { return {a.x + b.x, a.y + b.y}; }
inline Vec2 operator-(Vec2 a, Vec2 b) { return {a.x - b.x, a.y - b.y}; }
inline Vec2 operator*(Vec2 a, float s) { return {a.x * s, a.y * s}; }
inline Vec2 operator-(Vec2 a) { return {-a.x, -a.y}; }

// libstdc++ std::minstd_rand0 + std::uniform_real_distribution<float> (Seeder.cpp:5-20), restated so the
// catalogue does not depend on the C++ runtime's distribution algorithm staying the same.
struct Lcg {
    uint32_t x;
    explicit Lcg(uint64_t seed) {
        uint64_t s = seed % 2147483647ull;
        x = s == 0 ? 1u : (uint32_t)s;
    }
    float uniform(float a, float b) {
        x = (uint32_t)(((uint64_t)x * 16807ull) % 2147483647ull);
        float ret = (float)(x - 1u) / 2147483648.0f;
        if (ret >= 1.0f) ret = std::nextafter(1.0f, 0.0f);
        return ret * (b - a) + a;
    }
};

struct Regiment {  // demo.cpp:56-62
    Vec2 facing, head, spacing;
    unsigned cols, n;
};
Vec2 right_face(const Regiment& r) {  // cross((facing, 0), (0, 0, 1)).xy, demo.cpp:63-65
    return {r.facing.y * 1.f - 0.f * 0.f, 0.f * 0.f - 1.f * r.facing.x};
}
Regiment centered(Regiment r) {  // demo.cpp:66-70
    float half = (r.spacing.x * static_cast<float>(r.cols - 1)) / 2.f;
    r.head = r.head - right_face(r) * half;
    return r;
}
Vec2 regiment_pos(const Regiment* regs, unsigned i) {  // demo.cpp:77-94
    unsigned k = 0, last = regs[0].n, j = i;
    while (j >= last) {
        j -= last;
        last = regs[++k].n;
    }
    float u = regs[k].spacing.x * static_cast<float>(j % regs[k].cols);
    float v = regs[k].spacing.y * static_cast<float>(j / regs[k].cols);
    return right_face(regs[k]) * u + (-regs[k].facing) * v + regs[k].head;
}
Vec2 radial_pos(unsigned i, unsigned num, float radius) {  // demo.cpp:48-54
    const float pi = 3.14159265358979323846264338327950288f;
    float radian = static_cast<float>(i) / static_cast<float>(num) * 2.f * pi;
    return Vec2{std::cos(radian), std::sin(radian)} * radius;
}
Vec2 axial_goal(Vec2 p) {  // demo.cpp:36-41
    float d = std::fabs(p.x) - std::fabs(p.y);
    int axis = (0.f < d) - (d < 0.f);
    if (axis == 0) axis = 1;
    float f = static_cast<float>(axis);
    return {-f * p.x, f * p.y};
}

}  // namespace

extern "C" {

int32_t lc_scene(uint32_t scene, uint64_t seed, float* pos, float* goal, int32_t cap, float* max_duration,
                 uint32_t* rng_state_after) {
    int32_t n;
    float md;
    switch (scene) {
    case 0: n = 60; md = 45.f; break;   // demo.cpp:145-151
    case 1: n = 90; md = 70.f; break;   // demo.cpp:153-159
    case 4: n = 56; md = 40.f; break;   // demo.cpp:189-195
    case 5: n = 56; md = 70.f; break;   // demo.cpp:197-203
    case 12: n = 200; md = 30.f; break; // demo.cpp:356-373
    default: return LC_ERR_ARG;         // walled scenes: SURVEY §8f-2, not built yet
    }
    if (cap < n || !pos || !goal) return LC_ERR_ARG;
    Lcg rng(seed);
    Regiment regs[4] = {centered({{1, 0}, {-6.f, 0}, {1.f, 1.f}, 7, 14}), centered({{-1, 0}, {6.f, 0}, {1.f, 1.f}, 7, 14}),
                        centered({{0, 1}, {0, -6.f}, {1.f, 1.f}, 7, 14}), centered({{0, -1}, {0, 6.f}, {1.f, 1.f}, 7, 14})};
    for (int32_t i = 0; i < n; ++i) {
        Vec2 p{}, g{};
        if (scene == 0) { p = radial_pos((unsigned)i, 60, 10.f); g = -p; }
        else if (scene == 1) { p = radial_pos((unsigned)i, 90, 13.f); g = -p; }
        else if (scene == 4) { p = regiment_pos(regs, (unsigned)i); g = axial_goal(p); }
        else if (scene == 5) { p = regiment_pos(regs, (unsigned)i); g = -p; }
        else {
            // glm::vec2(inside(gen), inside(gen)): g++ evaluates the arguments right to left (verified with O-ref)
            p.y = rng.uniform(-10.f, 10.f); p.x = rng.uniform(-10.f, 10.f);
            g.y = rng.uniform(-10.f, 10.f); g.x = rng.uniform(-10.f, 10.f);
        }
        pos[2 * i] = p.x; pos[2 * i + 1] = p.y;
        goal[2 * i] = g.x; goal[2 * i + 1] = g.y;
    }
    if (max_duration) *max_duration = md;
    if (rng_state_after) *rng_state_after = rng.x;
    return n;
}

// Iteration order of the reference's component tables (quirk Q4).  Every DataTable is a
// std::unordered_map<uint16_t, ...> whose keys 1, 2, 3, ... are emplaced in ascending order and never erased
// (Pool.cpp:7-16), and every loop of the frame (comm.cpp:268,288; ai.cpp:153; LMP.cpp:288; physics.cpp:39; demo.cpp:669,692)
// walks such a table from begin() to end().  Which order that is belongs to libstdc++ (bucket growth policy, insertion at
// the head of the node list), so the same standard container is put through the same insertions here and read back;
// nothing of its internals is restated.  agent_order[k] = 0-based index of the k-th agent visited; wall_order[k] =
// 0-based index of the k-th wall visited by the statics loop (LMP.cpp:307: walls are keys n+1 .. n+W of the BoundVolume*
// table, demo.cpp:628-635).  More than 65 535 entities do not fit the reference's uint16_t ids (Entity.h:12): ascending.
int32_t lc_table_order(int32_t n_agents, int32_t n_walls, int32_t* agent_order, int32_t* wall_order) {
    if (n_agents < 0 || n_walls < 0 || (n_agents > 0 && !agent_order) || (n_walls > 0 && !wall_order)) return LC_ERR_ARG;
    if ((int64_t)n_agents + n_walls > 65535) {
        for (int32_t i = 0; i < n_agents; ++i) agent_order[i] = i;
        for (int32_t i = 0; i < n_walls; ++i) wall_order[i] = i;
        return LC_OK;
    }
    try {
        std::unordered_map<uint16_t, char> agents, volumes;
        for (int32_t k = 1; k <= n_agents; ++k) agents.emplace(std::piecewise_construct, std::forward_as_tuple((uint16_t)k), std::forward_as_tuple(0));
        int32_t m = 0;
        for (const auto& kv : agents) agent_order[m++] = (int32_t)kv.first - 1;
        if (n_walls > 0) {
            for (int32_t k = 1; k <= n_agents + n_walls; ++k) volumes.emplace(std::piecewise_construct, std::forward_as_tuple((uint16_t)k), std::forward_as_tuple(0));
            m = 0;
            for (const auto& kv : volumes)
                if ((int32_t)kv.first > n_agents) wall_order[m++] = (int32_t)kv.first - 1 - n_agents;
        }
    } catch (...) {
        return LC_ERR_CAPACITY;
    }
    return LC_OK;
}

// ---- static BVH over the walls (ai.cpp:16-18,36: built once by ai::init from the statics in table order) ----
// physics::simulate pushes an agent out of every wall static_bvh->query reports (physics.cpp:60,67-69).  That query is
// NOT "every wall the disc touches": inner nodes are entered through circ_rect_collider on the node's box, which tests a
// single point of the disc (BoundVolume.cpp:196-201, quirk Q8), so a wall near the END of a long row of tiles can be
// missed, and the order of the reported hits (= order of the push-outs) is the tree's depth-first leaf order.  Both are a
// function of the tree alone — walls never move — so the tree is rebuilt here exactly as BVH.cpp:9-49,144-251 does: two
// index tables sorted by origin x / z with std::sort (ties between tiles of one row are decided by libstdc++'s
// introsort on the same input sequence, hence the same container algorithm here), tight box per node, split of the
// longer axis into floor(n/2) | rest while the other table is stably partitioned.  Nodes are emitted in pre-order.
namespace {
struct BvhIdx { unsigned obj, oth; };
struct BvhBuild {
    const lc_rect* walls;        // creation order
    const int32_t* objs;         // statics order -> creation index
    std::vector<lc_bvh_node>* out;
    int32_t build(const std::vector<BvhIdx>& sx, const std::vector<BvhIdx>& sz) {
        const int32_t me = (int32_t)out->size();
        out->push_back(lc_bvh_node{0.f, 0.f, 0.f, 0.f, -1, -1});
        if (sx.size() == 1) {  // leaf: the wall itself
            (*out)[me].left = objs[sx[0].obj];
            return me;
        }
        float min_x = std::numeric_limits<float>::max(), min_z = min_x, max_x = -min_x, max_z = -min_x;
        for (const BvhIdx& k : sx) {
            const lc_rect& r = walls[objs[k.obj]];
            const float hx = r.w / 2, hz = r.h / 2;
            min_x = std::min(min_x, r.ox - hx); max_x = std::max(max_x, r.ox + hx);
            min_z = std::min(min_z, r.oy - hz); max_z = std::max(max_z, r.oy + hz);
        }
        const float dx = max_x - min_x, dz = max_z - min_z;
        (*out)[me].ox = min_x + dx / 2; (*out)[me].oy = min_z + dz / 2; (*out)[me].w = dx; (*out)[me].h = dz;
        std::vector<BvhIdx> xl, zl, xr, zr;
        if (dx > dz) split(sx, sz, xl, zl, xr, zr); else split(sz, sx, zl, xl, zr, xr);
        const int32_t l = build(xl, zl), r = build(xr, zr);
        (*out)[me].left = l; (*out)[me].right = r;
        return me;
    }
    // a = table of the split axis, b = the other one; each entry's .oth is its position in the other table
    static void split(const std::vector<BvhIdx>& a, const std::vector<BvhIdx>& b, std::vector<BvhIdx>& al, std::vector<BvhIdx>& bl,
                      std::vector<BvhIdx>& ar, std::vector<BvhIdx>& br) {
        const unsigned half = static_cast<unsigned>(static_cast<float>(a.size()) / 2.f);
        const unsigned rest = static_cast<unsigned>(static_cast<float>(a.size()) / 2.f + .5f);
        al.assign(half, BvhIdx{0, 0}); bl.assign(half, BvhIdx{0, 0}); ar.assign(rest, BvhIdx{0, 0}); br.assign(rest, BvhIdx{0, 0});
        for (size_t i = 0; i < a.size(); ++i) (i < half ? al[i] : ar[i - half]) = BvhIdx{a[i].obj, ~0u};
        unsigned nl = 0, nr = 0;
        for (const BvhIdx& e : b) {
            if (e.oth < half) { bl[nl] = BvhIdx{e.obj, e.oth}; al[e.oth].oth = nl++; }
            else { br[nr] = BvhIdx{e.obj, e.oth - half}; ar[e.oth - half].oth = nr++; }
        }
    }
};
}  // namespace

int32_t lc_static_bvh(int32_t n_walls, const lc_rect* walls, const int32_t* wall_order, lc_bvh_node* nodes, int32_t cap, int32_t* n_nodes) {
    if (n_walls < 1 || !walls || !nodes || !n_nodes) return LC_ERR_ARG;
    try {
        std::vector<int32_t> objs((size_t)n_walls);
        for (int32_t k = 0; k < n_walls; ++k) objs[k] = wall_order ? wall_order[k] : k;
        std::vector<BvhIdx> sx((size_t)n_walls), sz((size_t)n_walls);
        for (unsigned i = 0; i < (unsigned)n_walls; ++i) sx[i] = sz[i] = BvhIdx{i, 0};
        std::sort(sx.begin(), sx.end(), [&](BvhIdx a, BvhIdx b) { return walls[objs[a.obj]].ox < walls[objs[b.obj]].ox; });
        for (size_t i = 0; i < sx.size(); ++i) sz[sx[i].obj].oth = (unsigned)i;
        std::sort(sz.begin(), sz.end(), [&](BvhIdx a, BvhIdx b) { return walls[objs[a.obj]].oy < walls[objs[b.obj]].oy; });
        for (size_t i = 0; i < sz.size(); ++i) sx[sz[i].oth].oth = (unsigned)i;
        std::vector<lc_bvh_node> out;
        BvhBuild bb{walls, objs.data(), &out};
        bb.build(sx, sz);
        *n_nodes = (int32_t)out.size();
        if ((int32_t)out.size() > cap) return LC_ERR_CAPACITY;
        std::copy(out.begin(), out.end(), nodes);
    } catch (...) {
        return LC_ERR_CAPACITY;
    }
    return LC_OK;
}

// comm.cpp:80-121: 3 rows of 8 tokens (M_c | M_vy M_s M_py M_d M_g), then 2 rows whose first 3 tokens are M_f.
// Tokens are separated by single spaces (std::getline(sl, item, ' ')) and parsed with std::stof.
int32_t lc_read_config(const char* path, float* m_cv, float* m_f) {
    if (!path || !m_cv || !m_f) return LC_ERR_ARG;
    std::ifstream in(path);
    if (!in) return LC_ERR_IO;  // the reference dereferences an empty optional here (comm.cpp:78)
    std::string line;
    for (int row = 0; row < 5; ++row) {
        if (!std::getline(in, line)) line.clear();
        std::stringstream sl(line);
        std::string item;
        const int ncol = row < 3 ? 8 : 3;
        for (int col = 0; col < ncol; ++col) {
            std::getline(sl, item, ' ');
            float v;
            try {
                v = std::stof(item);
            } catch (...) {
                return LC_ERR_IO;  // reference: uncaught std::invalid_argument aborts the process
            }
            if (row < 3) m_cv[8 * row + col] = v; else m_f[3 * (row - 3) + col] = v;
        }
    }
    return LC_OK;
}

// comm.cpp:136-142: default ostream formatting (6 significant digits), avg_velocity then mean + 3 sigma
int32_t lc_write_result(const char* path, float avg_velocity, float group_time) {
    if (!path) return LC_ERR_ARG;
    std::ofstream out(path);
    if (!out) return LC_ERR_IO;
    out << avg_velocity << "\n" << group_time << "\n";
    return out.good() ? LC_OK : LC_ERR_IO;
}

}  // extern "C"
