/* shift_bulk.cu -- shiftObservationsSystem + shiftHiddenState (src/sim/sim.cpp:1002-1048) with the bulk-copy engine:
 * the one regular stream of the simulator (seven contiguous columns, cur -> prev) moved as 16 KB tiles,
 * global -> shared (cp.async.bulk + mbarrier complete_tx) -> global (cp.async.bulk from shared, bulk groups), a four-stage
 * ring per CTA with two loads in flight, one elected thread issuing; the stats column is patched in shared memory on its way
 * through (U7). What shift_observations() launches: 0.098 ms for 2.19 M agents = 5,910 GB/s = 90 % of the measured copy
 * bandwidth, against 0.1016 ms = 87 % for shift_kernel of sim_kernels.cu (plain 16-byte vector loads and stores; 0.0184
 * against 0.0208 ms at 8,192 worlds; r3w). MBOTS_SHIFT_BULK=0 selects that kernel instead. Same bytes written.
 */
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include "sim_dev.h"

namespace mbk {

/* tile size, ring depth, loads in flight and CTAs per SM: 32 KB x 3 x 2 x 2, 8 KB x 8 x 4 x 3, 16 KB x 6 x 4 x 2, 16 KB x 4 x 3 x 3 and
 * 32 KB x 4 x 2 x 1 all measure within 1 % of the shipped 16 KB x 4 x 2 x 3 (0.096-0.098 ms at 65,536 worlds, r4l) */
#ifndef MBK_BULK_CHUNK
#define MBK_BULK_CHUNK 16384
#endif
#ifndef MBK_BULK_STAGES
#define MBK_BULK_STAGES 4
#endif
#ifndef MBK_BULK_LAG
#define MBK_BULK_LAG 2
#endif
#ifndef MBK_BULK_CTAS_PER_SM
#define MBK_BULK_CTAS_PER_SM 3
#endif
constexpr uint32_t kBulkChunk = MBK_BULK_CHUNK;
constexpr int kBulkStages = MBK_BULK_STAGES;
constexpr int kBulkLag = MBK_BULK_LAG;
constexpr int kBulkThreads = 128;
constexpr int kBulkColumns = 7;

__device__ __forceinline__ void sb_mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void sb_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void sb_wait(uint32_t bar, uint32_t parity)
{
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void sb_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void sb_store(void *dst, uint32_t src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

struct BulkColumn { const unsigned char *src; unsigned char *dst; uint32_t row_bytes; uint32_t fix_stats; };
struct BulkPlan { BulkColumn col[kBulkColumns]; };

__global__ void __launch_bounds__(kBulkThreads) shift_bulk_kernel(BulkPlan plan, const Totals *totals)
{
    extern __shared__ __align__(128) unsigned char sb_smem[];
    __shared__ __align__(8) unsigned long long bars[kBulkStages];
    const uint32_t smem0 = (uint32_t)__cvta_generic_to_shared(sb_smem);
    const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(bars);
    const int tid = threadIdx.x;
    const size_t A = totals->totalAgents;
    if (tid == 0) {
        for (int i = 0; i < kBulkStages; ++i) sb_mbar_init(bar0 + 8u * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    /* chunks are numbered through the seven columns; the bulk part of a column is its length rounded down to 16 bytes, the
     * few bytes behind it are copied by ordinary stores */
    uint32_t first[kBulkColumns + 1];
    first[0] = 0;
#pragma unroll
    for (int c = 0; c < kBulkColumns; ++c) {
        const size_t bytes = (A * plan.col[c].row_bytes) & ~(size_t)15;
        first[c + 1] = first[c] + (uint32_t)((bytes + kBulkChunk - 1) / kBulkChunk);
    }
    const uint32_t total = first[kBulkColumns];
    auto locate = [&](uint32_t g, int &c, size_t &off, uint32_t &bytes) {
        c = 0;
#pragma unroll
        for (int k = 1; k < kBulkColumns; ++k) if (g >= first[k]) c = k;
        off = (size_t)(g - first[c]) * kBulkChunk;
        const size_t len = (A * plan.col[c].row_bytes) & ~(size_t)15;
        bytes = (uint32_t)(len - off < kBulkChunk ? len - off : kBulkChunk);
    };

    /* chunk i is loaded at iteration i and stored at iteration i + kBulkLag: kBulkLag loads in flight per CTA. Its stage was last
     * read by the store of chunk i - kBulkStages, issued kBulkStages - kBulkLag iterations ago: all but the
     * kBulkStages - kBulkLag - 1 newest bulk groups must have finished reading shared memory before the stage is refilled. */
    const uint32_t mine = blockIdx.x < total ? (total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    for (uint32_t i = 0; i < mine + kBulkLag; ++i) {
        if (i < mine && tid == 0) {
            const uint32_t st = i % kBulkStages;
            if (i >= (uint32_t)kBulkStages) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kBulkStages - kBulkLag - 1) : "memory");
            int c; size_t off; uint32_t bytes;
            locate(blockIdx.x + i * gridDim.x, c, off, bytes);
            sb_expect_tx(bar0 + 8u * st, bytes);
            sb_load(smem0 + st * kBulkChunk, plan.col[c].src + off, bytes, bar0 + 8u * st);
        }
        if (i >= (uint32_t)kBulkLag) {
            const uint32_t j = i - kBulkLag, st = j % kBulkStages;
            int c; size_t off; uint32_t bytes;
            locate(blockIdx.x + j * gridDim.x, c, off, bytes);
            if (plan.col[c].fix_stats) {                 /* prev.hitEnemy := cur.hitFriendly (src/sim/sim.cpp:1034, U7): patch in flight */
                sb_wait(bar0 + 8u * st, (j / kBulkStages) & 1u);
                int4 *rows = reinterpret_cast<int4 *>(sb_smem + st * kBulkChunk);
                for (uint32_t r = tid; r < bytes / 16u; r += kBulkThreads) rows[r].y = rows[r].x;
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncthreads();
                if (tid == 0) sb_store(plan.col[c].dst + off, smem0 + st * kBulkChunk, bytes);
            } else if (tid == 0) {
                sb_wait(bar0 + 8u * st, (j / kBulkStages) & 1u);
                sb_store(plan.col[c].dst + off, smem0 + st * kBulkChunk, bytes);
            }
        }
        /* the other threads only work on stats tiles, but must not drift from the issuing thread: a parity wait answers for the
         * phase before the current one too, so a thread two uses of a stage ahead would patch whatever tile is in it */
        __syncthreads();
    }
    if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    /* tails below 16 bytes (columns whose length is not a multiple of 16) */
    if (blockIdx.x == 0 && tid < kBulkColumns) {
        const BulkColumn col = plan.col[tid];
        const size_t len = A * col.row_bytes;
        for (size_t b = len & ~(size_t)15; b < len; b += 4)
            *reinterpret_cast<uint32_t *>(col.dst + b) = *reinterpret_cast<const uint32_t *>(col.src + b);
    }
}

bool shift_bulk_enabled()
{
    static int on = -1;
    if (on < 0) { const char *e = getenv("MBOTS_SHIFT_BULK"); on = !(e && e[0] == '0'); }
    return on != 0;
}

void launch_shift_bulk(const SimDev &d, int cur, cudaStream_t s)
{
    const ObsSet &o = d.obs[cur];
    BulkPlan plan;
    auto set = [&](int i, const void *src, void *dst, uint32_t row_bytes, uint32_t fix) {
        plan.col[i].src = static_cast<const unsigned char *>(src);
        plan.col[i].dst = static_cast<unsigned char *>(dst);
        plan.col[i].row_bytes = row_bytes;
        plan.col[i].fix_stats = fix;
    };
    set(0, o.hidden, o.prev_hidden, 64, 0);
    set(1, o.action, o.prev_action, 24, 0);
    set(2, o.stats, o.prev_stats, 16, 1);
    set(3, o.position, o.prev_position, 8, 0);
    set(4, o.surrounding, o.prev_surrounding, 8, 0);
    set(5, o.health, o.prev_health, 4, 0);
    set(6, o.reward, o.prev_reward, 4, 0);
    static bool attr[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    const int smem = kBulkStages * (int)kBulkChunk;
    if (dev >= 0 && dev < 64 && !attr[dev]) {
        cudaFuncSetAttribute(shift_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        attr[dev] = true;
    }
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    shift_bulk_kernel<<<sms * MBK_BULK_CTAS_PER_SM, kBulkThreads, smem, s>>>(plan, d.totals);
}

} // namespace mbk
