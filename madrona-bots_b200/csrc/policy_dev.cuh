/* policy_dev.cuh -- device helpers shared by the policy kernels (policy_kernels.cu, policy_fused.cu). */
#pragma once
#include <stdint.h>
#include "sim_prims.cuh"

namespace mbk {
using namespace mbp;

constexpr int kObsDim = 69;      /* learn/util.py:23-29: depth 32 | health 1 | position 2 | semantic 32 | surrounding 2 */

/* column c of construct_obs for observation row `row` */
__device__ __forceinline__ float obs_element(const uint8_t *__restrict__ depth, const float *__restrict__ health,
                                             const float *__restrict__ position, const int8_t *__restrict__ semantic,
                                             const float *__restrict__ surrounding, size_t row, uint32_t c)
{
    if (c < 32u) return (float)depth[row * 32 + c];
    if (c == 32u) return health[row];
    if (c < 35u) return position[row * 2 + (c - 33u)];
    if (c < 67u) return (float)semantic[row * 32 + (c - 35u)];
    return surrounding[row * 2 + (c - 67u)];
}

/* The agent that owns observation row `row`, as a shard-invariant 64-bit id: (GLOBAL world id << 10) | slot. */
__device__ __forceinline__ unsigned long long agent_id(const uint32_t *__restrict__ owner, uint32_t world_off, uint32_t row)
{
    const uint32_t o = owner[row];
    return ((unsigned long long)(world_off + (o >> 10)) << 10) | (unsigned long long)(o & 1023u);
}

/* u = uniform(split(draw_key, agent id)); action = first k with u * total < w0 + ... + wk (sums in index order);
 * stores the one-hot int32 row `row` of the action export and returns the pick. Keyed on the agent's (global world,
 * slot), not on the row number: the same agent draws the same number however the worlds are sharded over GPUs. */
__device__ __forceinline__ int sample_and_write(const float (&w)[6], Key draw_key, unsigned long long id, uint32_t row,
                                                int32_t *__restrict__ action)
{
    const float u = key_sample_uniform(key_split(draw_key, (uint32_t)id, (uint32_t)(id >> 32)));
    float total = w[0];
#pragma unroll
    for (int k = 1; k < 6; ++k) total = total + w[k];
    const float thr = u * total;
    int pick = 5;
    float c = 0.f;
    bool done = false;
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        c = c + w[k];
        if (!done && thr < c) { pick = k; done = true; }
    }
    int2 *ap = reinterpret_cast<int2 *>(action + (size_t)row * 6);
    ap[0] = make_int2(pick == 0, pick == 1);
    ap[1] = make_int2(pick == 2, pick == 3);
    ap[2] = make_int2(pick == 4, pick == 5);
    return pick;
}

} // namespace mbk
