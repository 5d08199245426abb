// O-ref lookahead harness: the reference's own LMP::lookahead (ai/LMP.cpp:84-113, compiled unmodified from
// /root/reference) over its own Cspace2d, on caller-supplied walls, waypoint lists and agent positions, so that the
// restatements (oracle/canon.c canon_lookahead, learn_comms_b200/csrc/walls.cuh lcw::lookahead) can be pinned.
// TEST INFRASTRUCTURE ONLY; this file is ours, everything it calls is the reference's.
//
//   oref_look < in.bin > out.bin
//
// in : int32 n_walls; float walls[n_walls][4]; float agent_radius; int32 K; then K records
//      { float pos[2]; float local_goal[2]; int32 num_done; int32 n_plan; float plan[8][2] }
// out: K records { int32 num_done; float local_goal[2] }
#include "ai/Agent.h"
#include "ai/BoundVolume.h"
#include "ai/LMP.h"
#include "ai/PRM.h"
#include <cstdint>
#include <cstdio>
#include <vector>

struct Rec { float pos[2], lg[2]; int32_t num_done, n_plan; float plan[8][2]; };

int main() {
    int32_t n_walls = 0, K = 0;
    float radius = 0;
    if (fread(&n_walls, 4, 1, stdin) != 1) return 1;
    std::vector<float> walls(static_cast<size_t>(n_walls) * 4);
    if (n_walls && fread(walls.data(), 4, walls.size(), stdin) != walls.size()) return 1;
    if (fread(&radius, 4, 1, stdin) != 1 || fread(&K, 4, 1, stdin) != 1) return 1;
    std::vector<Rec> recs(static_cast<size_t>(K));
    if (K && fread(recs.data(), sizeof(Rec), recs.size(), stdin) != recs.size()) return 1;

    std::vector<BoundVolume*> obs;
    for (int32_t k = 0; k < n_walls; ++k)
        obs.push_back(new Rect(glm::vec2(walls[4 * k], walls[4 * k + 1]), walls[4 * k + 2], walls[4 * k + 3]));
    Circ proto(glm::vec2(0, 0), radius);
    Cspace2d cspace(obs, &proto);

    for (const Rec& r : recs) {
        Agent a;
        a.cspace = &cspace;
        a.prm = nullptr;
        a.num_done = r.num_done;
        a.local_goal = glm::vec2(r.lg[0], r.lg[1]);
        std::vector<glm::vec2> plan;
        for (int32_t k = 0; k < r.n_plan; ++k) plan.push_back(glm::vec2(r.plan[k][0], r.plan[k][1]));
        a.plan = &plan;
        Circ bv(glm::vec2(r.pos[0], r.pos[1]), radius);
        glm::vec2 lg = LMP::lookahead(a, bv);
        int32_t nd = a.num_done;
        float out[2] = {lg.x, lg.y};
        fwrite(&nd, 4, 1, stdout);
        fwrite(out, 4, 2, stdout);
    }
    return 0;
}
