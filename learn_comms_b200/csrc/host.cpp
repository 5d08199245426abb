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

// ---- scene catalogue, demo.cpp:141-511 make_scene ----
namespace {

enum class Goal { opposite, axial, mirror_x, translate, random };
enum class Place { radial, regiments, list, random };
enum class Walls { none, random_posts, centre, table, hole, hall, column };

struct SceneSpec {
    int n = 0, n_walls = 0;
    float wall_scale = 0.f, max_duration = 0.f, radius = 0.f;
    Goal goal = Goal::opposite;
    Place place = Place::radial;
    Walls walls = Walls::none;
    Vec2 shift{0.f, 0.f};
    Regiment regs[4];
    Vec2 list[4];
    const float (*wall_table)[2] = nullptr;
};

// demo.cpp:214-231 (scene 6), :312-353 (scene 11), :461-486 (scene 18)
const float kCurbs[12][2] = {{-3, -3}, {-3, 3}, {3, -3}, {3, 3}, {4, 4}, {4, -4}, {6, -5}, {6, 5}, {-4, 4}, {-6, 5}, {-4, -4}, {-6, -5}};
const float kRoom[32][2] = {{0, 8},     {2, 8},     {4, 8},     {8, 8},      {-6, 8},     {-4, 8},     {-2, 8},  {-8, 0},
                            {-8, 2},    {-8, 4},    {-8, 6},    {-8, 8},     {-8, -6},    {-8, -4},    {-8, -2}, {-7.4f, -8},
                            {1.4f, -8}, {3.4f, -8}, {5.4f, -8}, {-5.4f, -8}, {-3.4f, -8}, {-1.4f, -8}, {8, 0},   {8, 2},
                            {8, 4},     {8, 6},     {7.4f, -8}, {8, -6},     {8, -4},     {8, -2},     {6, 3},   {4, 3}};
const float kTwoDoorY[20] = {1, -1, 3, -3, 6, -5, 8, -8, 10, -10, 12, -12, 14, -14, 16, -16, 18, -18, 20, -20};

void two_regiments(SceneSpec& s, Regiment a, Regiment b) {
    s.place = Place::regiments;
    s.regs[0] = centered(a);
    s.regs[1] = centered(b);
}
void doorway(SceneSpec& s) {  // demo.cpp:106-114 make_hole_wall(1.f, 2.f): two columns of 2 m tiles with a 1 m gap at the origin
    s.n_walls = 30; s.wall_scale = 2.0f; s.walls = Walls::hole; s.goal = Goal::mirror_x;
}

bool scene_spec(uint32_t scene, SceneSpec& s) {
    switch (scene) {
    case 0: s.n = 60; s.max_duration = 45.f; s.radius = 10.f; return true;   // demo.cpp:145-151
    case 1: s.n = 90; s.max_duration = 70.f; s.radius = 13.f; return true;   // :153-159
    case 2: s.n = 60; s.max_duration = 40.f; s.radius = 10.f; s.n_walls = 60; s.wall_scale = 0.3f; s.walls = Walls::random_posts; return true;  // :161-175
    case 3: s.n = 60; s.max_duration = 40.f; s.radius = 10.f; s.n_walls = 1; s.wall_scale = 4.0f; s.walls = Walls::centre; return true;        // :177-187
    case 4: case 5: case 6:  // :189-232, make_intersecting_bots(14, 7, 6.f) :96-104
        s.n = 56; s.max_duration = scene == 5 ? 70.f : 40.f; s.goal = scene == 5 ? Goal::opposite : Goal::axial;
        s.place = Place::regiments;
        s.regs[0] = centered({{1, 0}, {-6.f, 0}, {1.f, 1.f}, 7, 14}); s.regs[1] = centered({{-1, 0}, {6.f, 0}, {1.f, 1.f}, 7, 14});
        s.regs[2] = centered({{0, 1}, {0, -6.f}, {1.f, 1.f}, 7, 14}); s.regs[3] = centered({{0, -1}, {0, 6.f}, {1.f, 1.f}, 7, 14});
        if (scene == 6) { s.n_walls = 12; s.wall_scale = 2.0f; s.walls = Walls::table; s.wall_table = kCurbs; }
        return true;
    case 7: s.n = 2; s.max_duration = 70.f; doorway(s); s.place = Place::list; s.list[0] = {-3.f, 0}; s.list[1] = {3.f, 0}; return true;  // :234-246
    case 8: s.n = 4; s.max_duration = 30.f; doorway(s); two_regiments(s, {{1, 0}, {-3, 0}, {1.f, 1.f}, 2, 2}, {{-1, 0}, {3, 0}, {1.f, 1.f}, 2, 2}); return true;
    case 9: s.n = 40; s.max_duration = 100.f; doorway(s); two_regiments(s, {{1, 0}, {-3, 0}, {1.f, 1.f}, 5, 20}, {{-1, 0}, {3, 0}, {1.f, 1.f}, 5, 20}); return true;
    case 10: s.n = 40; s.max_duration = 100.f; doorway(s); two_regiments(s, {{1, 0}, {-3, 0}, {1.f, 1.f}, 5, 5}, {{-1, 0}, {3, 0}, {1.f, 1.f}, 5, 35}); return true;
    case 11:  // escaping room, :300-355
        s.n = 70; s.max_duration = 100.0f; s.n_walls = 32; s.wall_scale = 2.0f; s.walls = Walls::table; s.wall_table = kRoom;
        s.goal = Goal::translate; s.shift = {-20.f, 0.f};
        two_regiments(s, {{0, 1}, {1, 6}, {2.f, 1.5f}, 3, 6}, {{-1, 0}, {-4, -2}, {1.f, 1.f}, 8, 64});
        return true;
    case 12: s.n = 200; s.max_duration = 30.f; s.place = Place::random; s.goal = Goal::random; return true;  // :358-373
    case 13:  // two formations in a hall, :375-390
        s.n = 40; s.max_duration = 90.0f; s.n_walls = 18; s.wall_scale = 2.0f; s.walls = Walls::hall; s.goal = Goal::mirror_x;
        two_regiments(s, {{1, 0}, {-3, 0}, {.5f, .7f}, 5, 20}, {{-1, 0}, {3, 0}, {.5f, .7f}, 5, 20});
        return true;
    case 14: s.n = 2; s.max_duration = 30.f; s.goal = Goal::mirror_x; s.place = Place::list; s.list[0] = {-5.f, .1f}; s.list[1] = {3.f, 0}; return true;
    case 15: s.n = 3; s.max_duration = 40.f; s.goal = Goal::mirror_x; s.place = Place::list;
        s.list[0] = {-5.f, .1f}; s.list[1] = {3.f, 0}; s.list[2] = {7.f, 0}; return true;
    case 16: s.n = 4; s.max_duration = 40.f; s.goal = Goal::mirror_x; s.place = Place::list;
        s.list[0] = {-5.f, .1f}; s.list[1] = {3.f, 0}; s.list[2] = {7.f, 0}; s.list[3] = {-2.f, .1f}; return true;
    case 17: s.n = 4; s.max_duration = 45.f; s.goal = Goal::mirror_x;
        two_regiments(s, {{1, 0}, {-5, .1f}, {1.f, 1.f}, 2, 2}, {{-1, 0}, {3, 0}, {1.f, 1.f}, 2, 2}); return true;
    case 18:  // two-door, :448-488
        s.n = 65; s.max_duration = 60.f; s.n_walls = 20; s.wall_scale = 2.0f; s.walls = Walls::column;
        s.goal = Goal::translate; s.shift = {-15.f, 0.f};
        two_regiments(s, {{-1, 0}, {3, 3}, {1.f, 1.f}, 5, 5}, {{-1, 0}, {3, -6}, {1.f, 1.f}, 10, 60});
        return true;
    default:
        if (scene >= 900 && scene < 900 + 3000) {  // :490-508: scene 9 with 40 + (scn - 900) robots (the reference needs scn > 100)
            s.n = 40 + (int)(scene - 900); s.max_duration = 100.f; doorway(s);
            const unsigned even = (unsigned)s.n / 2, odd = even + (unsigned)s.n % 2;
            two_regiments(s, {{1, 0}, {-3, 0}, {1.f, 1.f}, 5, odd}, {{-1, 0}, {3, 0}, {1.f, 1.f}, 5, even});
            return true;
        }
        return false;  // 19 .. 899: the reference leaves Scene::num_robos uninitialised
    }
}

}  // namespace

extern "C" {

// demo::init's make_scene + the create_robo / create_wall loops (demo.cpp:611-635): robots first, then walls; scenes 2
// and 12 draw from the seeded global stream while they are built.
int32_t lc_scene_full(uint32_t scene, uint64_t seed, float* pos, float* goal, int32_t cap, float* max_duration,
                      uint32_t* rng_state_after, lc_rect* walls, int32_t wall_cap, int32_t* n_walls) {
    SceneSpec sp;
    if (!scene_spec(scene, sp)) return LC_ERR_ARG;
    if (cap < sp.n || !pos || !goal) return LC_ERR_ARG;
    if (sp.n_walls > 0 && (!walls || wall_cap < sp.n_walls)) return LC_ERR_ARG;
    Lcg rng(seed);
    for (int32_t i = 0; i < sp.n; ++i) {
        Vec2 p{}, g{};
        switch (sp.place) {
        case Place::radial: p = radial_pos((unsigned)i, (unsigned)sp.n, sp.radius); break;
        case Place::regiments: p = regiment_pos(sp.regs, (unsigned)i); break;
        case Place::list: p = sp.list[i]; break;
        case Place::random:  // glm::vec2(inside(gen), inside(gen)): g++ evaluates the arguments right to left (verified with O-ref)
            p.y = rng.uniform(-10.f, 10.f); p.x = rng.uniform(-10.f, 10.f); break;
        }
        switch (sp.goal) {
        case Goal::opposite: g = -p; break;                       // demo.cpp:29
        case Goal::axial: g = axial_goal(p); break;               // :36-41
        case Goal::mirror_x: g = {-p.x, p.y}; break;              // :33-35
        case Goal::translate: g = p + sp.shift; break;            // :30-32
        case Goal::random: g.y = rng.uniform(-10.f, 10.f); g.x = rng.uniform(-10.f, 10.f); break;
        }
        pos[2 * i] = p.x; pos[2 * i + 1] = p.y;
        goal[2 * i] = g.x; goal[2 * i + 1] = g.y;
    }
    for (int32_t i = 0; i < sp.n_walls; ++i) {  // create_wall: square Rect of side wall_scale (demo.cpp:44-46,523)
        Vec2 p{};
        const float side = (i % 2) == 0 ? 1.f : -1.f;
        switch (sp.walls) {
        case Walls::random_posts: p.y = rng.uniform(-7.f, 7.f); p.x = rng.uniform(-7.f, 7.f); break;  // demo.cpp:170-174
        case Walls::centre: break;
        case Walls::table: p = {sp.wall_table[i][0], sp.wall_table[i][1]}; break;
        case Walls::hole: p = {0.f, side * (sp.wall_scale * (static_cast<float>(i / 2) + .5f) + 1.f / 2.f)}; break;  // :106-114
        case Walls::hall:  // make_hall(3.f, scale), :115-124; (i + 2 % 4) / 4 parses as (i + 2) / 4
            p = {(i % 4 < 2 ? 1.f : -1.f) * sp.wall_scale * static_cast<float>((i + 2 % 4) / 4), side * .5f * (3.f + sp.wall_scale)};
            break;
        case Walls::column: p = {0.f, kTwoDoorY[i]}; break;
        case Walls::none: break;
        }
        walls[i] = lc_rect{p.x, p.y, sp.wall_scale, sp.wall_scale};
    }
    if (max_duration) *max_duration = sp.max_duration;
    if (rng_state_after) *rng_state_after = rng.x;
    if (n_walls) *n_walls = sp.n_walls;
    return sp.n;
}

// the wall-free scenes through the original entry point; walled ones need lc_scene_full (the walls are part of the scene)
int32_t lc_scene(uint32_t scene, uint64_t seed, float* pos, float* goal, int32_t cap, float* max_duration,
                 uint32_t* rng_state_after) {
    SceneSpec sp;
    if (!scene_spec(scene, sp) || sp.n_walls > 0) return LC_ERR_ARG;
    return lc_scene_full(scene, seed, pos, goal, cap, max_duration, rng_state_after, nullptr, 0, nullptr);
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

// optimizer.py:20-27's format (the four comms_*.config files at the reference's root): "3 5", then 3 rows of 5
int32_t lc_read_config_legacy(const char* path, float* m_cv, float* m_f, uint32_t* flags) {
    if (!path || !m_cv || !m_f) return LC_ERR_ARG;
    std::ifstream in(path);
    if (!in) return LC_ERR_IO;
    int rows = 0, cols = 0;
    if (!(in >> rows >> cols) || rows != 3 || cols != 5) return LC_ERR_IO;
    float m[15];
    for (float& v : m)
        if (!(in >> v)) return LC_ERR_IO;
    for (int k = 0; k < 24; ++k) m_cv[k] = 0.f;
    for (int k = 0; k < 6; ++k) m_f[k] = 0.f;
    m_cv[0] = m[0];                              // M_c (1 x 1) -> channel 0's self weight
    m_cv[3] = m[1]; m_cv[4] = m[2];              // M_vx, M_vy
    m_cv[5] = m[3]; m_cv[6] = m[4];              // M_px, M_py
    m_f[0] = m[5]; m_f[3] = m[10];               // M_f (2 x 1) -> column 0
    if (flags) *flags = LC_FLAG_FEAT_PLAIN;
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

// lc_dd_create's default strips.  What a rank's kernels work on is its LOCAL set — the agents it owns plus the ghosts of its
// one (edge strips) or two (interior strips) bands — and the ranks run in lock step through the neighbour waits, so it is
// the local counts that are equalised: start from the n/world quantiles of the x coordinates, then shift every cut by half
// the difference of the loads either side of it, a few sweeps.  Host code, same result on every rank.
int lc_set_error(int code, const char* msg);  // capi.cu: text returned by lc_last_error()
extern "C" int32_t lc_dd_strips(int32_t n, const float* x, int32_t stride, int32_t world, float ghost_band, float* x_bounds) {
    if (!x || !x_bounds || stride < 1 || world < 1 || n < 2 * world || !(ghost_band >= 0.f))
        return lc_set_error(LC_ERR_ARG, "lc_dd_strips: need x, x_bounds, stride >= 1, world >= 1, n >= 2 world, ghost_band >= 0");
    std::vector<float> xs((size_t)n);
    for (int i = 0; i < n; ++i) xs[i] = x[(size_t)i * stride];
    std::sort(xs.begin(), xs.end());
    std::vector<long long> cut((size_t)world + 1);   // cut[r] = agents left of strip r
    for (int r = 0; r <= world; ++r) cut[r] = (long long)r * n / world;
    auto cut_x = [&](long long k) { return 0.5f * (xs[k - 1] + xs[k]); };
    auto load = [&](int r) {   // agents in the ghost-extended strip r
        const float lo = r == 0 ? -3.0e38f : cut_x(cut[r]) - ghost_band, hi = r == world - 1 ? 3.0e38f : cut_x(cut[r + 1]) + ghost_band;
        return (long long)(std::lower_bound(xs.begin(), xs.end(), hi) - std::lower_bound(xs.begin(), xs.end(), lo));
    };
    for (int sweep = 0; sweep < 12; ++sweep)
        for (int r = 1; r < world; ++r) {
            const long long shift = (load(r - 1) - load(r)) / 2;   // > 0: the left strip is heavier, the cut moves left
            cut[r] = std::min(std::max(cut[r] - shift, cut[r - 1] + 1), cut[r + 1] - 1);
        }
    x_bounds[0] = -3.0e38f; x_bounds[world] = 3.0e38f;
    for (int r = 1; r < world; ++r) x_bounds[r] = cut_x(std::min<long long>(std::max<long long>(cut[r], 1), n - 1));
    return LC_OK;
}
