// TEST INFRASTRUCTURE ONLY.  Compiles the product's device header learn_comms_b200/csrc/walls.cuh for the HOST (its
// functions are __host__ __device__, plain IEEE binary32 operators) so that the CPU suite can hold the very source the
// GPU runs against the reference's bit patterns without a GPU.  Never built into or loaded by liblearncomms.so.
#include "../learn_comms_b200/csrc/walls.cuh"

extern "C" void walls_host_batch(int32_t n, const float* rec, uint32_t* out) {
    for (int32_t k = 0; k < n; ++k) lcw::geometry_record(rec + (size_t)k * 9, out + (size_t)k * 12);
}
