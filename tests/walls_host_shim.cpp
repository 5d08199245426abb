// TEST INFRASTRUCTURE ONLY.  Compiles the product's device header learn_comms_b200/csrc/walls.cuh for the HOST (its
// functions are __host__ __device__, plain IEEE binary32 operators) so that the CPU suite can hold the very source the
// GPU runs against the reference's bit patterns without a GPU.  Never built into or loaded by liblearncomms.so.
#include "../learn_comms_b200/csrc/walls.cuh"

extern "C" void walls_host_batch(int32_t n, const float* rec, uint32_t* out) {
    for (int32_t k = 0; k < n; ++k) lcw::geometry_record(rec + (size_t)k * 9, out + (size_t)k * 12);
}

// records of tests/_oref.py LOOK_REC / LOOK_OUT
struct LookRec { float pos[2], lg[2]; int32_t num_done, n_plan; float plan[8][2]; };
struct LookOut { int32_t num_done; float lg[2]; };
extern "C" void walls_host_lookahead(int32_t n_walls, const float* walls, float r, int32_t n, const LookRec* rec, LookOut* out) {
    for (int32_t k = 0; k < n; ++k) {
        int nd = rec[k].num_done;
        float lgx = rec[k].lg[0], lgy = rec[k].lg[1];
        lcw::lookahead(reinterpret_cast<const lcw::Wall*>(walls), n_walls, r, rec[k].pos[0], rec[k].pos[1], &rec[k].plan[0][0],
                       rec[k].n_plan, nd, lgx, lgy);
        out[k].num_done = nd;
        out[k].lg[0] = lgx;
        out[k].lg[1] = lgy;
    }
}
