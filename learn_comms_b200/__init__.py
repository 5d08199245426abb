"""learn_comms_b200 — B200-native (sm_100a) implementation of learn-comms' per-frame collision-avoidance + comms loop.

The product is liblearncomms.so (hand-written CUDA behind the C ABI in include/learn_comms.h); this package is the thin
ctypes binding used by the tests, bench.py and the batched PSO evaluator.  There is no CPU fallback.
"""
from .api import Sim, DomainSim, Batch, LcError, lib, scene, scene_full, read_config_legacy, FEAT_PLAIN, FEAT_NO_PROX, FEAT_NO_SOFTSIGN, table_order, static_bvh, dd_strips, fp32_peak, wall_geometry, lookahead, Planner, read_config, write_result, FLAG_NO_COMM, FLAG_FREE_RUN, FLAG_DEBUG_LISTS  # noqa: F401
