// Batched rollouts (placeholder until the block-per-scenario kernel lands in this file).
#include "../../include/learn_comms.h"
extern "C" {
int32_t lc_batch_create(const lc_batch_desc*, lc_batch** out) { if (out) *out = nullptr; return LC_ERR_ARG; }
void lc_batch_destroy(lc_batch*) {}
int32_t lc_batch_run(lc_batch*, int32_t, float*, float*, int32_t*, int32_t*) { return LC_ERR_ARG; }
int32_t lc_batch_get_state(lc_batch*, float*, float*, float*) { return LC_ERR_ARG; }
int32_t lc_batch_timer(lc_batch*, float*) { return LC_ERR_ARG; }
}
