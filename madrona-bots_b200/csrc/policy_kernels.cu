/* policy_kernels.cu -- the two memory-bound ends of policy inference on the simulator's exports
 * (reference: learn/util.py:23-29 construct_obs, learn/training_loop.py:57-83,136).
 *
 *   pack_obs_kernel   construct_obs in one pass: the five observation exports of a row range are
 *                     converted and concatenated into the (n, 69) float32 matrix the species networks
 *                     take -- [depth u8 x32 | health | position x2 | semantic i8 x32 | surrounding x2] --
 *                     instead of five dtype-conversion kernels plus a torch.cat.
 *   sample_kernel     Categorical sampling + the one-hot write-back: one thread per agent draws
 *                     u ~ U[0,1) from a counter-based Threefry stream keyed (seed, draw, row), picks the
 *                     action by inverse CDF over the six weights, and stores the int32 one-hot row
 *                     straight into the action export (training_loop.py:82-83,136).
 * The dense layers in between are plain library GEMMs (cuBLAS through torch); nothing here is one.
 */
#include <cuda_runtime.h>
#include <stdint.h>
#include "sim_dev.h"
#include "sim_prims.cuh"
#include "policy_dev.cuh"

namespace mbk {
using namespace mbp;

__global__ void __launch_bounds__(256) pack_obs_kernel(const uint8_t *__restrict__ depth, const float *__restrict__ health,
                                                       const float *__restrict__ position, const int8_t *__restrict__ semantic,
                                                       const float *__restrict__ surrounding, uint32_t row_begin,
                                                       uint32_t rows, float *__restrict__ out)
{
    const size_t total = (size_t)rows * kObsDim;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const uint32_t r = (uint32_t)(idx / kObsDim), c = (uint32_t)(idx - (size_t)r * kObsDim);
        const size_t row = (size_t)row_begin + r;
        out[idx] = obs_element(depth, health, position, semantic, surrounding, row, c);
    }
}

/* weights: (rows, 6) non-negative float32 (probabilities need not be normalised).
 * u = uniform(split(split(init(seed), draw), agent id)); action = first k with u * total < w0 + ... + wk. */
__global__ void __launch_bounds__(256) sample_kernel(const float *__restrict__ weights, int32_t *__restrict__ action,
                                                     const uint32_t *__restrict__ owner, uint32_t world_off,
                                                     uint32_t row_begin, uint32_t rows, uint32_t seed, uint32_t draw,
                                                     int32_t *__restrict__ out_index)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    const float2 *wp = reinterpret_cast<const float2 *>(weights + (size_t)r * 6);
    const float2 w01 = wp[0], w23 = wp[1], w45 = wp[2];
    const float w[6] = {w01.x, w01.y, w23.x, w23.y, w45.x, w45.y};
    const uint32_t row = row_begin + r;
    const int pick = sample_and_write(w, key_split(key_init(seed), draw), agent_id(owner, world_off, row), row, action);
    if (out_index) out_index[r] = pick;
}

void launch_pack_obs(const ObsSet &o, int is_prev, uint32_t row_begin, uint32_t rows, float *out, cudaStream_t s)
{
    if (rows == 0) return;
    const size_t total = (size_t)rows * kObsDim;
    size_t blocks = (total + 255) / 256;
    if (blocks > 148u * 32u) blocks = 148u * 32u;
    if (is_prev)
        pack_obs_kernel<<<(unsigned)blocks, 256, 0, s>>>(o.prev_depth, o.prev_health, o.prev_position, o.prev_semantic,
                                                         o.prev_surrounding, row_begin, rows, out);
    else
        pack_obs_kernel<<<(unsigned)blocks, 256, 0, s>>>(o.depth, o.health, o.position, o.semantic, o.surrounding,
                                                         row_begin, rows, out);
}

void launch_sample(const ObsSet &o, uint32_t row_begin, uint32_t rows, const float *weights, uint32_t seed, uint32_t draw,
                   int32_t *out_index, cudaStream_t s)
{
    if (rows == 0) return;
    sample_kernel<<<(rows + 255) / 256, 256, 0, s>>>(weights, o.action, o.owner, o.worldIdOffset, row_begin, rows, seed, draw, out_index);
}

} // namespace mbk
