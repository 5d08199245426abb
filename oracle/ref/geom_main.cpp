// O-ref geometry harness: evaluates the reference's own static-wall geometry (compiled unmodified from /root/reference:
// ai/BoundVolume.cpp, ai/LMP.cpp, ai/PRM.cpp) on caller-supplied inputs, so that O-canon's wall functions
// (oracle/canon.c "static walls") can be pinned bit for bit.  TEST INFRASTRUCTURE ONLY; this file is ours, everything
// it calls is the reference's.
//
//   oref_geom < in.bin > out.bin
//
// in : int32 K, then K records of 9 floats  { rect ox, oy, w, h;  a.x, a.y;  b.x, b.y;  r }
// out: K records of 12 words:
//   f32 Rect(o,w,h).intersect(a, b)                      BoundVolume.cpp:160-188   (b read as a direction)
//   i32 Rect(o,w,h).line_of_sight(a, b)                  BoundVolume.cpp:131-155
//   i32 Circ(o, r).line_of_sight(a, b)                   BoundVolume.cpp:76-98
//   f32 LMP::ttc(Circ(a,r), b, Rect(o,w,h), 0)           LMP.cpp:13-34,64-82       (b read as the agent velocity)
//   i32 circ_rect_collider(Circ(a,r), Rect)              BoundVolume.cpp:196-201
//   f32 x2  closest_aabb_point(on_q, Rect) - on_q, on_q = closest_circ_point(Rect._o, Circ(a,r))   BVH.cpp:121-130
//   i32 Cspace2d({Rect}, Circ(a,r)).line_of_sight(a, b)  PRM.cpp:133-156,169-173
//   i32 interacts; f32 x2 force                          LMP.cpp:307-319 statics-loop body with d.pos = a, d.vel = b
//   i32 0 (pad)
#include "ai/BoundVolume.h"
#include "ai/Dynamics.h"
#include "ai/LMP.h"
#include "ai/PRM.h"
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

union Word { float f; int32_t i; };

int main() {
    int32_t K = 0;
    if (fread(&K, 4, 1, stdin) != 1) return 1;
    std::vector<float> in(static_cast<size_t>(K) * 9);
    if (fread(in.data(), 4, in.size(), stdin) != in.size()) return 1;
    for (int32_t k = 0; k < K; ++k) {
        const float* p = &in[static_cast<size_t>(k) * 9];
        Rect rect(glm::vec2(p[0], p[1]), p[2], p[3]);
        glm::vec2 a(p[4], p[5]), b(p[6], p[7]);
        float r = p[8];
        Circ disc(glm::vec2(p[0], p[1]), r);
        Circ agent(a, r);
        Word w[12];
        memset(w, 0, sizeof(w));
        w[0].f = rect.intersect(a, b);
        w[1].i = rect.line_of_sight(a, b);
        w[2].i = disc.line_of_sight(a, b);
        w[3].f = LMP::ttc(agent, b, rect, glm::vec2(0));
        w[4].i = circ_rect_collider(&agent, &rect);
        glm::vec2 on_q = closest_circ_point(rect._o, &agent);
        glm::vec2 away = closest_aabb_point(on_q, &rect) - on_q;
        w[5].f = away.x;
        w[6].f = away.y;
        {
            std::vector<BoundVolume*> obs{&rect};
            Cspace2d cs(obs, &agent);
            w[7].i = cs.line_of_sight(a, b);
        }
        {   // LMP.cpp:307-319, one iteration
            Dynamics d;
            memset(&d, 0, sizeof(d));
            d.pos = glm::vec3(a.x, 0, a.y);
            d.vel = glm::vec3(b.x, 0, b.y);
            const double TTC_THRESHOLD = 5.0;  // LMP.cpp:115
            glm::vec2 F(0);
            int hit = 0;
            if (!(glm::length2(rect._o - glm::vec2(d.pos.x, d.pos.z)) > TTC_THRESHOLD)) {
                double ttc = LMP::ttc(agent, glm::vec2(d.vel.x, d.vel.z), rect, glm::vec2(0));
                if (!(ttc > TTC_THRESHOLD)) {
                    F = LMP::ttc_forces(d, agent, rect, static_cast<float>(ttc));
                    hit = 1;
                }
            }
            w[8].i = hit;
            w[9].f = F.x;
            w[10].f = F.y;
        }
        fwrite(w, 4, 12, stdout);
    }
    return 0;
}
