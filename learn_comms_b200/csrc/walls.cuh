// Static-wall geometry (first piece of the walled scenes, §8f-2): what an agent disc
// needs to know about an axis-aligned wall tile.  Plain binary32 operators only: the library is compiled with
// -fmad=false -prec-div=true -prec-sqrt=true -ftz=false, so every * + / sqrtf below is one IEEE round-to-nearest
// operation, in the reference's order.  The functions are __host__ __device__ so that the CPU suite can compile this
// very header with g++ and hold it against the reference's bit patterns (tests/test_oracle_cpu.py); the product only
// ever runs them on the device.
#pragma once
#include <math.h>
#include <stdint.h>
#ifdef __CUDACC__
#define LCW_HD __host__ __device__ __forceinline__
#else
#define LCW_HD static inline
#endif

namespace lcw {

#define LCW_FLT_MAX 3.402823466e+38f

struct Wall { float ox, oy, w, h; };  // demo.cpp:515-537 create_wall -> Rect(origin, wall_scale, wall_scale)

LCW_HD float dot2(float ax, float ay, float bx, float by) { return ax * bx + ay * by; }
LCW_HD float glm_min(float a, float b) { return (b < a) ? b : a; }
LCW_HD float glm_max(float a, float b) { return (a < b) ? b : a; }
LCW_HD float glm_clamp(float x, float lo, float hi) { return glm_min(glm_max(x, lo), hi); }

// BoundVolume.cpp:160-188 Rect::intersect: slab test of the ray o + v t, FLT_MAX on a miss
LCW_HD float rect_intersect(Wall r, float ox, float oy, float vx, float vy) {
    float tmin = 0.0f, tmax = LCW_FLT_MAX;
    const float lo[2] = {r.ox - r.w / 2, r.oy - r.h / 2}, hi[2] = {r.ox + r.w / 2, r.oy + r.h / 2};
    const float o[2] = {ox, oy}, v[2] = {vx, vy};
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        if (fabsf(v[i]) < 0.0000001f) {
            if (o[i] < lo[i] || o[i] > hi[i]) return LCW_FLT_MAX;
        } else {
            float t_near = (lo[i] - o[i]) / v[i];
            float t_far = (hi[i] - o[i]) / v[i];
            if (t_near > t_far) { float t = t_near; t_near = t_far; t_far = t; }
            tmin = (tmin < t_near) ? t_near : tmin;  // std::max(tmin, near)
            tmax = (t_far < tmax) ? t_far : tmax;    // std::min(tmax, far)
            if (tmin > tmax) return LCW_FLT_MAX;
        }
    }
    return tmin;
}

// BoundVolume.cpp:131-155 Rect::line_of_sight: true when the segment a-b misses the rect (separating axes)
LCW_HD bool rect_line_of_sight(Wall r, float ax, float ay, float bx, float by) {
    float lox = r.ox - r.w / 2, loy = r.oy - r.h / 2, hix = r.ox + r.w / 2, hiy = r.oy + r.h / 2;
    float ex = hix - lox, ey = hiy - loy;
    float mx = ((ax + bx) - lox) - hix, my = ((ay + by) - loy) - hiy;
    float dx = bx - ax, dy = by - ay;
    float adx = fabsf(dx);
    if (fabsf(mx) > ex + adx) return true;
    float ady = fabsf(dy);
    if (fabsf(my) > ey + ady) return true;
    adx += 0.0000001f;
    ady += 0.0000001f;
    if (fabsf(mx * dy - my * dx) > ex * ady + ey * adx) return true;
    return false;
}

// BoundVolume.cpp:76-98 Circ::line_of_sight
LCW_HD bool circ_line_of_sight(float cx, float cy, float r, float sx, float sy, float ex, float ey) {
    float dx = ex - sx, dy = ey - sy;
    float mx = sx - cx, my = sy - cy;
    float b = dot2(mx, my, dx, dy);
    float c = dot2(mx, my, mx, my) - r * r;
    if (c > 0.f && b > 0.f) return true;
    float a = dot2(dx, dy, dx, dy);
    float discrim = b * b - a * c;
    if (discrim < 0.f) return true;
    float t = (-b - sqrtf(discrim)) / a;
    if (t * t > a) return true;
    return false;
}

// BoundVolume.cpp:102-122 Circ::intersect
LCW_HD float circ_intersect(float cx, float cy, float r, float ox, float oy, float dx, float dy) {
    float mx = ox - cx, my = oy - cy;
    float a = dot2(dx, dy, dx, dy);
    float b = dot2(mx, my, dx, dy);
    float c = dot2(mx, my, mx, my) - r * r;
    if (c > 0.f && b > 0.f) return LCW_FLT_MAX;
    float discrim = b * b - a * c;
    if (discrim < 0.f) return LCW_FLT_MAX;
    float t = (-b - sqrtf(discrim)) / a;
    if (t < 0.f) return 0.f;
    return t;
}

// LMP.cpp:13-34 ttc + :64-82 ttc_(Circ,Rect): earliest hit over the wall extruded by r along each axis and the four
// corner discs; the wall does not move (LMP.cpp:314-315).  Negative or NaN -> FLT_MAX.
LCW_HD float ttc_circ_wall(float ox, float oy, float r, float vx, float vy, Wall j) {
    Wall h = {j.ox, j.oy, 2 * r + j.w, j.h};
    Wall v = {j.ox, j.oy, j.w, 2 * r + j.h};
    float dvx = vx - 0.f, dvy = vy - 0.f;
    float t_min = LCW_FLT_MAX, t;
    t = rect_intersect(h, ox, oy, dvx, dvy); if (t < t_min) t_min = t;
    t = rect_intersect(v, ox, oy, dvx, dvy); if (t < t_min) t_min = t;
    t = circ_intersect(j.ox + j.w / 2, j.oy + j.h / 2, r, ox, oy, dvx, dvy); if (t < t_min) t_min = t;
    t = circ_intersect(j.ox - j.w / 2, j.oy + j.h / 2, r, ox, oy, dvx, dvy); if (t < t_min) t_min = t;
    t = circ_intersect(j.ox + j.w / 2, j.oy - j.h / 2, r, ox, oy, dvx, dvy); if (t < t_min) t_min = t;
    t = circ_intersect(j.ox - j.w / 2, j.oy - j.h / 2, r, ox, oy, dvx, dvy); if (t < t_min) t_min = t;
    if (!(t_min >= 0.f)) return LCW_FLT_MAX;
    return t_min;
}

// LMP.cpp:307-319, one iteration of the statics loop of calc_sum_force: squared-distance gate against 5 (quirk Q10),
// ttc <= 5, then ttc_forces(d, bv, bbv, ttc) (:128-136) with the f64 magnitude of ttc_forces_ (:115-127, quirk Q11)
LCW_HD bool wall_force(float ox, float oy, float r, float vx, float vy, Wall w, float& fx, float& fy) {
    fx = fy = 0.f;
    float gx = w.ox - ox, gy = w.oy - oy;
    if ((double)dot2(gx, gy, gx, gy) > 5.0) return false;
    float t = ttc_circ_wall(ox, oy, r, vx, vy, w);
    if ((double)t > 5.0) return false;
    float dx = (ox + vx * t) - w.ox, dy = (oy + vy * t) - w.oy;
    float len = dot2(dx, dy, dx, dy);
    if (len > 0) {
        float s = sqrtf(len);
        dx = dx / s;
        dy = dy / s;
    }
    double td = (double)t, mag = 0;
    if (td >= 0 && td <= 5.0) mag = (5.0 - td) / (td + 0.001);
    mag = mag > 20 ? 20 : mag;
    fx = (float)(mag * (double)dx);
    fy = (float)(mag * (double)dy);
    return true;
}

// BoundVolume.cpp:209-230 closest_circ_point / closest_aabb_point and BVH.cpp:121-130: the push vector on_bv - on_q
LCW_HD void circ_wall_normal(float qx, float qy, float r, Wall w, float& onqx, float& onqy, float& nx, float& ny) {
    float ux = w.ox - qx, uy = w.oy - qy;
    float inv = 1.0f / sqrtf(dot2(ux, uy, ux, uy));
    onqx = (ux * inv) * r + qx;
    onqy = (uy * inv) * r + qy;
    float cx = glm_clamp(onqx, w.ox - w.w / 2, w.ox + w.w / 2);
    float cy = glm_clamp(onqy, w.oy - w.h / 2, w.oy + w.h / 2);
    if (cx == onqx && cy == onqy) {  // inside the wall: go to the nearer pair of edges
        float ix = onqx - w.ox, iy = onqy - w.oy;
        if (fabsf(ix) > fabsf(iy)) cx = w.ox + (ix > 0 ? +w.w : -w.w) / 2;
        else cy = w.oy + (iy > 0 ? +w.h : -w.h) / 2;
    }
    nx = cx - onqx;
    ny = cy - onqy;
}

// BoundVolume.cpp:125-127 Rect::collides, :196-201 circ_rect_collider: contact test of BVH::query_(Circ*) on a Rect leaf
LCW_HD bool circ_wall_hit(float qx, float qy, float r, Wall w, float& nx, float& ny) {
    float onqx, onqy;
    circ_wall_normal(qx, qy, r, w, onqx, onqy, nx, ny);
    float ux = w.ox - qx, uy = w.oy - qy;
    bool q_has_origin = dot2(ux, uy, ux, uy) <= r * r;
    bool wall_has_onq = fabsf(onqx - w.ox) <= w.w / 2 && fabsf(onqy - w.oy) <= w.h / 2;
    return q_has_origin || wall_has_onq;
}

// PRM.cpp:133-156 Cspace2d (agent's 2r x 2r bounding square summed into the wall, BoundVolume.cpp:55-58) and
// :169-173 line_of_sight for ONE wall; the caller ANDs over walls
LCW_HD bool cspace_line_of_sight(Wall w, float agent_r, float ax, float ay, float bx, float by) {
    Wall infl = {w.ox, w.oy, 2 * agent_r + w.w, 2 * agent_r + w.h};
    return rect_line_of_sight(infl, ax, ay, bx, by);
}

// LMP.cpp:84-113 lookahead: advance along the waypoint list.  The current target counts as reached within 1 m
// (squared distance against ai::at_dist = 1); further waypoints are skipped while the one after the target is in line
// of sight through the configuration space.  plan = [n_plan][2]; num_done and local_goal are updated in place
// (local_goal keeps its old value when the reached target was the last waypoint, as in the reference).
LCW_HD void lookahead(const Wall* walls, int n_walls, float agent_r, float px, float py, const float* plan, int n_plan,
                      int& num_done, float& lgx, float& lgy) {
    int target = num_done;
    if (target < n_plan) {
        float dx = px - plan[2 * target], dy = py - plan[2 * target + 1];
        if (dot2(dx, dy, dx, dy) < 1.f) ++target;
        if (target < n_plan) {
            lgx = plan[2 * target];
            lgy = plan[2 * target + 1];
        }
        for (int next = target + 1; next < n_plan; next = target + 1) {
            bool visible = true;
            for (int k = 0; k < n_walls && visible; ++k)
                visible = cspace_line_of_sight(walls[k], agent_r, px, py, plan[2 * next], plan[2 * next + 1]);
            if (!visible) break;
            ++target;
            lgx = plan[2 * target];
            lgy = plan[2 * target + 1];
        }
        num_done = target;
    } else {
        lgx = plan[2 * (n_plan - 1)];
        lgy = plan[2 * (n_plan - 1) + 1];
    }
}

// One record of the parity fixture (oracle/ref/geom_main.cpp): 9 floats {wall; a; b; r} -> 12 words
LCW_HD void geometry_record(const float* p, uint32_t* out) {
    union { float f; int32_t i; uint32_t u; } w[12];
    for (int k = 0; k < 12; ++k) w[k].u = 0;
    Wall rect = {p[0], p[1], p[2], p[3]};
    float ax = p[4], ay = p[5], bx = p[6], by = p[7], r = p[8];
    w[0].f = rect_intersect(rect, ax, ay, bx, by);
    w[1].i = rect_line_of_sight(rect, ax, ay, bx, by);
    w[2].i = circ_line_of_sight(rect.ox, rect.oy, r, ax, ay, bx, by);
    w[3].f = ttc_circ_wall(ax, ay, r, bx, by, rect);
    float nx, ny, fx, fy;
    w[4].i = circ_wall_hit(ax, ay, r, rect, nx, ny);
    w[5].f = nx;
    w[6].f = ny;
    w[7].i = cspace_line_of_sight(rect, r, ax, ay, bx, by);
    w[8].i = wall_force(ax, ay, r, bx, by, rect, fx, fy);
    w[9].f = fx;
    w[10].f = fy;
    for (int k = 0; k < 12; ++k) out[k] = w[k].u;
}

}  // namespace lcw
