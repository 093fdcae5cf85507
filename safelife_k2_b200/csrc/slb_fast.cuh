// slb_fast.cuh — bit-sliced fast path (stub until the kernels land: everything goes generic).
#pragma once
#include "slb_common.cuh"
namespace slb {
inline bool fast_shape(int, int) { return false; }
inline void launch_advance_fast(const uint16_t *, uint16_t *, int, int, int, const float *, uint64_t,
                                uint64_t, uint64_t *, int, cudaStream_t) {}
inline void launch_env_rollout_fast(const slb_pool &, int, uint16_t *, uint16_t *, slb_env_state *,
                                    const uint8_t *, int, int, uint64_t, uint64_t, int32_t *,
                                    uint8_t *, int, cudaStream_t) {}
inline void launch_side_effect_fast(const uint16_t *, const uint16_t *, int, int, int, const int32_t *,
                                    int, const float *, uint64_t, uint64_t, uint64_t *, int, uint16_t *,
                                    uint16_t *, int32_t *, cudaStream_t) {}
}  // namespace slb
