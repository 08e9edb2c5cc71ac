/* sim_kernels.cu -- the per-step hot path of madrona-bots as hand-written sm_100a kernels.
 *
 *   step_kernel        one WARP per world (four worlds per CTA, no block barriers). Restates the Step task graph up to
 *                      the world sort (src/sim/sim.cpp:1061-1132): chunk reset, food spawn, actionSystem, healthSync,
 *                      surroundings, species tracker, speciesInfoSync + respawn, and the compaction the reference gets
 *                      from SortArchetypeNode<Agent, WorldID>. World state lives in the warp's shared memory for the
 *                      duration; every cross-agent effect (shots, chunk counters, food, births) is a shared-memory
 *                      atomic, a ballot + popc rank or a warp reduction. Actions come from the action export or straight
 *                      from the caller's device matrix (StepOpts).
 *   scan_fused_kernel  the only cross-world dependency: exclusive scan of the per-(world, species) row counts in
 *                      species-major order -> first row of each (world, species) segment. Replaces the global radix
 *                      sort by SpeciesObservation (src/sim/sim.cpp:1148-1149). One launch with a grid barrier
 *                      (tile_sum_kernel + scan_kernel when the grid cannot be resident at once).
 *   export_kernel      one warp per world (or 2-8 single-warp CTAs per world for small launches). Gathers every column
 *                      that travels with an agent from its old row to its new row (the "sort"), writes the observation
 *                      exports (updateObservations :687-717, updateSensorOutputIdx :736-789, rewardSystem :840-983) and
 *                      runs the 32-ray sensor over the world's agents staged in shared memory
 *                      (M/src/mw/device/bvh_raycast.cpp:193-305; two-stage angular cull + exact ray / sphere tests instead
 *                      of a TLAS). With StepOpts::fuseShift it also leaves prev := cur, i.e. the shift.
 *   shift_kernel       shiftObservationsSystem + shiftHiddenState (src/sim/sim.cpp:1002-1048) as a call of its own.
 *
 * Compiled with -fmad=false: results must be bit-identical to the CPU oracle.
 */
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include "sim_dev.h"
#include "sim_prims.cuh"

namespace mbk {
using namespace mbp;

#define FULL 0xffffffffu

/* -DMBK_CHECKED (lib/variants/libmbots_checked.so, `make checked`): every shared-memory and row index the kernels form is
 * asserted in range (device assert -> the next CUDA call fails). Stands in for compute-sanitizer's memcheck, which is
 * closed on this GPU pool; the parity suite is run once against this build (tools/gpu_checked.sh). */
#if defined(MBK_CHECKED)
#include <assert.h>
#define MBK_CHECK(cond) assert(cond)
#else
#define MBK_CHECK(cond) ((void)0)
#endif

/* ------------------------------------------------------------------------- */
/* Options of a fused rollout step (mbots_rollout_step_async). actSrc: the caller's action matrix -- rows below
 * actRows are read from it instead of the action export (the step takes its actions from there, the export's carry
 * writes them into the new rows), so no separate action-write pass runs. fuseShift: the export kernel leaves the
 * prev_* columns as shiftObservationsSystem + shiftHiddenState (src/sim/sim.cpp:1002-1048) would right after the
 * step -- prev := cur, written in the same pass -- so no separate shift pass runs and the old prev rows are never read. */
struct StepOpts {
    const int2 *actSrc;
    unsigned long long actRows;
    int fuseShift;
};

__device__ __forceinline__ const int2 *action_row(const StepOpts &o, const int32_t *exported, int row)
{
    const bool ext = o.actSrc != nullptr && (unsigned long long)row < o.actRows;
    return (ext ? o.actSrc : reinterpret_cast<const int2 *>(exported)) + (size_t)row * 3;
}

/* ------------------------------------------------------------------------- */
/* small block-level helpers                                                   */

__device__ __forceinline__ int comp(const int4 &v, int k)
{
    return k == 0 ? v.x : k == 1 ? v.y : k == 2 ? v.z : v.w;
}

/* ------------------------------------------------------------------------- */
/* init: Sim::Sim + initWorld + initializeChunks (src/sim/sim.cpp:1232-1256, 233-300) */

__global__ void init_kernel(SimDev d)
{
    const uint32_t w = blockIdx.x;
    const uint32_t s = threadIdx.x;
    const Key key = key_split(key_init(d.seed), 0u, d.worldIdOffset + w);
    const size_t base = (size_t)w * d.cap;
    if (s < d.initAgents) {
        const float x = key_sample_uniform(key_split(key, 2u * s)) * kWorldX;
        const float y = key_sample_uniform(key_split(key, 2u * s + 1u)) * kWorldY;
        d.pose[base + s] = make_float4(x, y, 1.f, 0.f);
        d.health[base + s] = 100;
        d.species[base + s] = (uint8_t)((s % kNumSpecies) + 1u);
        d.target[base + s] = (int16_t)-1;
        d.scratch[base + s] = make_float4(__int_as_float(-1), 0.f, 0.f, __int_as_float(0));
    }
    if (s < (uint32_t)kNumChunks) d.food[(size_t)w * kNumChunks + s] = make_uint2(0u, 0u);
    if (s == 0) {
        d.worldCount[w] = (int32_t)d.initAgents;
        d.rngKey[w] = make_uint2(key.a, key.b);
        d.rngCount[w] = 2u * d.initAgents;
        d.numFood[w] = 0;
        int4 c;
        const int q = (int)(d.initAgents / kNumSpecies), r = (int)(d.initAgents % kNumSpecies);
        c.x = q + (r > 0); c.y = q + (r > 1); c.z = q + (r > 2); c.w = q;
        d.speciesLive[w] = c;
        d.speciesCount[w] = make_int4(0, 0, 0, 0);
        d.speciesReward[w] = make_float4(0.f, 0.f, 0.f, 0.f);
        d.speciesHealth[w] = make_int4(0, 0, 0, 0);
        WorldStats z = {0, 0, 0, 0, 0, 0, 0, 0};
        d.wstats[w] = z;
    }
}

/* ------------------------------------------------------------------------- */
/* step_kernel: one WARP per world, kStepWarps independent worlds per CTA.     */
/* No block barriers. Agents are taken 32 at a time (one or two passes for a   */
/* normal world); what an agent needs in a later phase is parked in this       */
/* warp's shared memory, cross-agent effects are shared-memory atomics, and    */
/* every count / rank is a ballot + popc or a warp reduction.                  */

#ifndef MBK_STEP_WARPS
#define MBK_STEP_WARPS 4
#endif
constexpr int kStepWarps = MBK_STEP_WARPS;           /* worlds (warps) per CTA */

/* per-warp shared memory, sized for `slots` >= the largest world before the step */
struct StepLayout { uint32_t chunkAgents, chunkSpeed, food, pose, accum, oldrow, info, spec, total; };
__host__ __device__ inline StepLayout step_layout(uint32_t slots)
{
    StepLayout L;
    uint32_t o = 0;
    L.pose = o; o += slots * 16;                        /* float4 x, y, rot.w, rot.z (after the move) */
    L.food = o; o += kNumChunks * 8;
    L.chunkAgents = o; o += kNumChunks * 4;
    L.chunkSpeed = o; o += kNumChunks * 4;
    L.accum = o; o += slots * 4;                        /* HealthAccumulator */
    L.oldrow = o; o += slots * 4;
    L.info = o; o += slots * 4;
    L.spec = o; o += (slots + 15u) & ~15u;
    L.total = (o + 15u) & ~15u;
    return L;
}

size_t step_smem_bytes(uint32_t slots) { return (size_t)step_layout(slots).total * kStepWarps; }

/* info word of an agent: action bits 0-5 | stat bits 6-9 (hitFriendly, hitEnemy, ateFood, reproduced) |
 * species 10-12 | finder target + 1 in 13-23 */
__device__ __forceinline__ uint32_t food_cell(const uint2 &f, int p)
{
    return p < 4 ? (f.x >> (8 * p)) & 0xffu : f.y & 0xffu;
}

/* packages of chunk record `f` that hold food at `cell` */
__device__ __forceinline__ uint32_t food_match(const uint2 &f, uint32_t cell)
{
    const uint32_t has = (f.y >> 8) & 0x1fu;
    uint32_t m = 0;
#pragma unroll
    for (int p = 0; p < kMaxFoodPackages; ++p)
        if (food_cell(f, p) == cell) m |= 1u << p;
    return m & has;
}

__device__ __forceinline__ int warp_sum(int v) { return __reduce_add_sync(FULL, v); }

#ifndef MBK_STEP_MIN_BLOCKS
#define MBK_STEP_MIN_BLOCKS (32 / MBK_STEP_WARPS)
#endif
__global__ void __launch_bounds__(kStepWarps * 32, MBK_STEP_MIN_BLOCKS) step_kernel(SimDev d, int cur, uint32_t slots, StepOpts opts)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t w = blockIdx.x * kStepWarps + warp;
    if (w >= d.numWorlds) return;                       /* whole warp */
    const StepLayout L = step_layout(slots);
    unsigned char *my = smem_raw + (size_t)L.total * warp;
    float4 *spose = reinterpret_cast<float4 *>(my + L.pose);
    uint2 *food = reinterpret_cast<uint2 *>(my + L.food);
    uint32_t *chunkAgents = reinterpret_cast<uint32_t *>(my + L.chunkAgents);
    uint32_t *chunkSpeed = reinterpret_cast<uint32_t *>(my + L.chunkSpeed);
    int *accum = reinterpret_cast<int *>(my + L.accum);
    int *soldrow = reinterpret_cast<int *>(my + L.oldrow);
    uint32_t *sinfo = reinterpret_cast<uint32_t *>(my + L.info);
    uint8_t *spec = my + L.spec;

    const uint32_t cap = d.cap;
    const size_t base = (size_t)w * cap;
    const ObsSet &from = d.obs[cur];
    const unsigned lt = (1u << lane) - 1u;

    /* ---- P0: stage the world: state, the action row of every agent, food; reset the chunk counters ---- */
    /* Two DRAM round trips for a world of up to 64 agents: (1) the state of slots lane and lane + 32 (requested
     * before the live count is known -- slots below the capacity always exist), the per-world scalars and the
     * food records; (2) the action rows, whose addresses come from (1). Larger worlds take further passes. */
    const bool has_b = 32u + (uint32_t)lane < cap;
    const size_t ia = base + lane, ib = has_b ? base + 32 + lane : ia;   /* never form an address outside the world's slots */
    float4 pose_a = d.pose[ia], pose_b = d.pose[ib];
    int hp_a = d.health[ia], hp_b = d.health[ib];
    uint32_t sp_a = d.species[ia], sp_b = d.species[ib];
    int tgt_a = d.target[ia], tgt_b = d.target[ib];
    int row_a = d.row[ia], row_b = d.row[ib];
    const int n0 = d.worldCount[w];
    MBK_CHECK(n0 >= 0 && (uint32_t)n0 <= slots && slots <= cap);
    const uint2 kk = d.rngKey[w];
    const uint32_t rng0 = d.rngCount[w];
    int numFood = d.numFood[w];
    static_assert(kNumChunks == 48, "two loads per lane cover the chunk records");
    const uint2 food0 = d.food[(size_t)w * kNumChunks + lane];
    const uint2 food1 = d.food[(size_t)w * kNumChunks + 32 + (lane & 15)];
    auto park = [&](int s, const float4 &pose, int hp, uint32_t sp, int tgt, int row, const int2 &a0, const int2 &a1,
                    const int2 &a2) {
        const uint32_t act = (a0.x != 0) | ((a0.y != 0) << 1) | ((a1.x != 0) << 2) | ((a1.y != 0) << 3) |
                             ((a2.x != 0) << 4) | ((a2.y != 0) << 5);
        MBK_CHECK(s >= 0 && (uint32_t)s < slots && sp >= 1u && sp <= 4u && tgt >= -1 && tgt < n0);
        MBK_CHECK(row >= 0 && (uint32_t)row < d.numWorlds * d.cap);
        spose[s] = pose;
        accum[s] = hp;
        soldrow[s] = row;
        spec[s] = (uint8_t)sp;
        sinfo[s] = act | (sp << 10) | ((uint32_t)(tgt + 1) << 13);
    };
    {
        const bool on_a = lane < n0, on_b = has_b && 32 + lane < n0;
        const int2 zero = make_int2(0, 0);
        int2 a0 = zero, a1 = zero, a2 = zero, b0 = zero, b1 = zero, b2 = zero;
        if (on_a) {
            const int2 *ap = action_row(opts, from.action, row_a);
            a0 = ap[0]; a1 = ap[1]; a2 = ap[2];
        }
        if (on_b) {
            const int2 *bp = action_row(opts, from.action, row_b);
            b0 = bp[0]; b1 = bp[1]; b2 = bp[2];
        }
        if (on_a) park(lane, pose_a, hp_a, sp_a, tgt_a, row_a, a0, a1, a2);
        if (on_b) park(32 + lane, pose_b, hp_b, sp_b, tgt_b, row_b, b0, b1, b2);
    }
    for (int s = 64 + lane; s < n0; s += 32) {
        const int row = d.row[base + s];
        const int2 *ap = action_row(opts, from.action, row);
        park(s, d.pose[base + s], d.health[base + s], d.species[base + s], d.target[base + s], row, ap[0], ap[1], ap[2]);
    }
    chunkAgents[lane] = 0;                              /* resetChunkInfoSystem (:390-396) */
    chunkSpeed[lane] = 0;
    food[lane] = food0;
    if (lane < kNumChunks - 32) {
        chunkAgents[32 + lane] = 0;
        chunkSpeed[32 + lane] = 0;
        food[32 + lane] = food1;
    }
    __syncwarp();

    /* ---- addFoodSystem + addFoodToChunk (src/sim/sim.cpp:308-387). The stream is counter-based: draw k is
     * split_i(key, count + k) whatever happened before it. Nine steps in ten use draw 0 only; otherwise the
     * warp pre-splits 32 draws (at most 2 + 3*7 = 23 are used), one per lane, and every lane walks the same
     * sequential logic, fetching draw k by shuffle. ---- */
    const Key wkey{kk.x, kk.y};
    uint32_t foodDraws = 1;
    bool foodDirty = false;
    if (key_sample_i32(key_split(wkey, rng0), 0, 10) == 0) {
        const Key mine = key_split(wkey, rng0 + (uint32_t)lane);
        auto draw = [&](uint32_t k, int lo, int hi) {
            return key_sample_i32(Key{__shfl_sync(FULL, mine.a, (int)k), __shfl_sync(FULL, mine.b, (int)k)}, lo, hi);
        };
        uint32_t k = 1;
        uint32_t nadd = (uint32_t)draw(k++, 1, 3);
        const uint32_t allowed = (uint32_t)kTotalAllowedFood - (uint32_t)numFood;   /* unsigned, U12 */
        nadd = nadd < allowed ? nadd : allowed;
        for (uint32_t i = 0; i < nadd; ++i) {
            const uint32_t cx = (uint32_t)draw(k, 0, kChunksX);
            const uint32_t cy = (uint32_t)draw(k + 1, 0, kChunksY);
            k += 4;                                     /* + two discarded draws (:311-312) */
            uint2 f = food[cx + cy * kChunksX];
            const uint32_t mask = (f.y >> 8) & 0x1fu;
            if (mask != 0x1fu) {
                const uint32_t p = (uint32_t)__ffs((int)(~mask & 0x1fu)) - 1u;   /* first empty package */
                const uint32_t rx = (uint32_t)draw(k, 0, kChunkWidth);
                const uint32_t ry = (uint32_t)draw(k + 1, 0, kChunkWidth);
                k += 3;                                 /* + yaw of the render-only food entity (:340) */
                const uint32_t cell = (rx << 4) | ry;
                if (p < 4u) f.x = (f.x & ~(0xffu << (8 * p))) | (cell << (8 * p));
                else f.y = (f.y & ~0xffu) | cell;
                f.y |= 1u << (8 + p);
                __syncwarp();
                if (lane == 0) food[cx + cy * kChunksX] = f;
                __syncwarp();
                numFood += 1;
                foodDirty = true;
            }
        }
        foodDraws = k;
    }
    const uint32_t rngBase = rng0 + foodDraws;

    /* ---- P1: actionSystem (:419-502) ---- */
    for (int s = lane; s < n0; s += 32) {
        uint32_t info = sinfo[s];
        const int tgt = (int)((info >> 13) & 0x7ffu) - 1;
        const uint32_t sp = (info >> 10) & 7u;
        if ((info & 16u) && tgt >= 0) {                 /* shoot with last frame's finder hit (:434-454) */
            atomicAdd(&accum[tgt], -50);
            info |= (spec[tgt] == sp) ? (1u << 6) : (2u << 6);
            info |= 1u << 24;                           /* a shot that hit (statistics) */
        }
        float4 pose = spose[s];
        Rot rot{pose.z, pose.w};
        const float len = move_agent(pose.x, pose.y, rot, (info & 1u) != 0, (info & 2u) != 0, (info & 4u) != 0, (info & 8u) != 0);
        pose.z = rot.w; pose.w = rot.z;
        const int c = chunk_of(pose.x, pose.y);
        MBK_CHECK(c >= 0 && c < kNumChunks);
        atomicAdd(&chunkAgents[c], 1u);
        atomicAdd(&chunkSpeed[c], (uint32_t)(len * 2.f));
        spose[s] = pose;
        sinfo[s] = info;
    }
    __syncwarp();

    /* ---- P2: healthSync (:505-581), updateSurroundingObservation (:583-654), speciesTrackerUpdate (:719-734),
     * and the compaction the reference gets from SortArchetypeNode<Agent, WorldID>: survivors keep their
     * order, then newborns in parent order (U11). Passes run in slot order. ---- */
    const int room = (int)cap - n0;                     /* makeAgent fails once the slot table is full (D1) */
    int nLive = 0, nBreed = 0, eaten = 0, shots = 0;
    int cnt[4] = {0, 0, 0, 0}, hsum[4] = {0, 0, 0, 0};
    for (int s0 = 0; s0 < n0; s0 += 32) {
        const int s = s0 + lane;
        const bool active = s < n0;
        uint32_t info = 0;
        float4 pose = make_float4(0.f, 0.f, 1.f, 0.f);
        int hp = 0, oldrow = -1;
        if (active) { info = sinfo[s]; pose = spose[s]; hp = accum[s]; oldrow = soldrow[s]; }
        const int sp = (int)((info >> 10) & 7u);
        const int tgt = (int)((info >> 13) & 0x7ffu) - 1;
        uint32_t stat = (info >> 6) & 15u;
        shots += __popc(__ballot_sync(FULL, (info >> 24) & 1u));

        /* food: the first package of the agent's chunk at the agent's cell that still holds food. Agents that
         * want the same cell take the matching packages in slot order (FoodPackage::consume CAS race,
         * sim.inl:76-99 -> lowest slot wins, U11). Rare: skipped unless some agent stands on food. */
        const int mychunk = active ? chunk_of(pose.x, pose.y) : 0;
        const uint32_t mycell = cell_of(pose.x, pose.y);
        const uint32_t mm = active ? food_match(food[mychunk], mycell) : 0u;
        bool ate = false;
        if (__any_sync(FULL, mm != 0u)) {
            const uint32_t key = mm ? ((uint32_t)mychunk << 8) | mycell : 0x80000000u | (uint32_t)lane;
            const unsigned group = __match_any_sync(FULL, key);
            const int r = __popc(group & lt);           /* wanting agents of lower slot at the same cell */
            if (mm && r < __popc(mm)) {
                uint32_t rest = mm;
                for (int i = 0; i < r; ++i) rest &= rest - 1u;
                const uint32_t p = (uint32_t)__ffs((int)rest) - 1u;
                atomicAnd(&food[mychunk].y, ~(1u << (8 + p)));
                ate = true;
            }
            const int e = __popc(__ballot_sync(FULL, ate));
            numFood -= e;
            eaten += e;
            foodDirty = true;
            __syncwarp();                               /* later passes see the packages gone */
        }
        if (ate) { hp += 20; stat |= 4u; }
        const bool breeds = active && (info & 32u) && tgt >= 0 && spec[tgt] == sp && hp > 10;
        if (breeds) { hp -= 40; stat |= 8u; }
        const bool live = active && hp > 0;
        const unsigned bb = __ballot_sync(FULL, breeds), bl = __ballot_sync(FULL, live);
        const int bidx = nBreed + __popc(bb & lt), lidx = nLive + __popc(bl & lt);
        const bool childOk = breeds && bidx < room;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const bool mine = sp == k + 1;
            cnt[k] += __popc(__ballot_sync(FULL, mine && live)) + __popc(__ballot_sync(FULL, mine && childOk));
            hsum[k] += warp_sum(mine ? (live ? hp : 0) + (childOk ? 50 : 0) : 0);
        }
        float sn = 0.f, ss = 0.f;
        if (live || childOk) surroundings(pose.x, pose.y, chunkAgents, chunkSpeed, sn, ss);
        __syncwarp();                                   /* every lane has read its parked record */
        if (live) {
            const size_t o = base + (size_t)lidx;
            MBK_CHECK(lidx >= 0 && (uint32_t)lidx < cap && mychunk >= 0 && mychunk < kNumChunks);
            d.pose[o] = pose;
            d.health[o] = hp;
            d.species[o] = (uint8_t)sp;
            d.scratch[o] = make_float4(__int_as_float(oldrow), sn, ss, __int_as_float((int)stat));
        }
        if (childOk) {                                  /* park the newborn: its slot needs the final live count */
            MBK_CHECK(bidx >= 0 && (uint32_t)bidx < slots);
            spose[bidx] = make_float4(pose.x, pose.y, sn, ss);
            sinfo[bidx] = (uint32_t)sp;
        }
        nBreed += __popc(bb);
        nLive += __popc(bl);
    }
    const int bornOk = nBreed < room ? nBreed : room;
    __syncwarp();
    for (int b = lane; b < bornOk; b += 32) {           /* a newborn shares its parent's position and surroundings */
        const float4 c = spose[b];
        const size_t o = base + (size_t)(nLive + b);
        MBK_CHECK((uint32_t)(nLive + b) < cap);
        d.pose[o] = make_float4(c.x, c.y, 1.f, 0.f);
        d.health[o] = 50;
        d.species[o] = (uint8_t)sinfo[b];
        d.scratch[o] = make_float4(__int_as_float(-1), c.z, c.w, __int_as_float(0));
    }

    /* ---- P3: speciesInfoSync + respawn (:791-838) ---- */
    const uint32_t q = d.initAgents / kNumSpecies;
    uint32_t need[4], pre[4];
    uint32_t needTotal = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        need[k] = (uint32_t)cnt[k] < q ? q - (uint32_t)cnt[k] : 0u;
        pre[k] = needTotal;
        needTotal += need[k];
    }
    const int size1 = n0 + bornOk;                      /* table size before respawn (dead rows still count) */
    const int room2 = (int)cap - size1;
    const uint32_t respOk = (int)needTotal < room2 ? needTotal : (uint32_t)room2;
    for (uint32_t s = (uint32_t)lane; s < needTotal; s += 32u) {
        int k = 0;
        if (s >= pre[1] + need[1]) k = (s >= pre[2] + need[2]) ? 3 : 2;
        else if (s >= pre[0] + need[0]) k = 1;
        const float x = key_sample_uniform(key_split(wkey, rngBase + 2u * s)) * kWorldX;
        const float y = key_sample_uniform(key_split(wkey, rngBase + 2u * s + 1u)) * kWorldY;
        if (s < respOk) {
            const size_t o = base + (size_t)(nLive + bornOk) + s;
            MBK_CHECK((uint32_t)(nLive + bornOk) + s < cap);
            d.pose[o] = make_float4(x, y, 1.f, 0.f);
            d.health[o] = 100;
            d.species[o] = (uint8_t)(k + 1);
            d.scratch[o] = make_float4(__int_as_float(-1), 0.f, 0.f, __int_as_float(0));
        }
    }
    /* the species reward terms cost three IEEE divisions each: lane k computes species k, lane 0 collects */
    float rmine;
    {
        const int k = lane & 3;
        const int ck = k == 0 ? cnt[0] : k == 1 ? cnt[1] : k == 2 ? cnt[2] : cnt[3];
        const int hk = k == 0 ? hsum[0] : k == 1 ? hsum[1] : k == 2 ? hsum[2] : hsum[3];
        rmine = species_reward((uint32_t)ck, (uint32_t)hk, d.initAgents);
    }
    const float4 rw = make_float4(__shfl_sync(FULL, rmine, 0), __shfl_sync(FULL, rmine, 1), __shfl_sync(FULL, rmine, 2),
                                  __shfl_sync(FULL, rmine, 3));
    if (lane == 0) {
        int4 liveRows;
        int *lr = &liveRows.x;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t ok = respOk > pre[k] ? (respOk - pre[k] < need[k] ? respOk - pre[k] : need[k]) : 0u;
            lr[k] = cnt[k] + (int)ok;
        }
        const int count1 = nLive + bornOk + (int)respOk;
        d.speciesLive[w] = liveRows;
        d.speciesCount[w] = make_int4(cnt[0], cnt[1], cnt[2], cnt[3]);
        d.speciesHealth[w] = make_int4(hsum[0], hsum[1], hsum[2], hsum[3]);
        d.speciesReward[w] = rw;
        d.worldCount[w] = count1;
        d.rngCount[w] = rngBase + 2u * needTotal;
        d.numFood[w] = numFood;
        const uint32_t deaths = (uint32_t)(n0 - nLive);
        const uint32_t overflow = (uint32_t)(nBreed - bornOk) + (needTotal - respOk);
        /* cumulative event counters: fire-and-forget reductions, nothing waits on them */
        WorldStats *ws = d.wstats + w;
        if (bornOk) atomicAdd(&ws->births, (uint32_t)bornOk);
        if (deaths) atomicAdd(&ws->deaths, deaths);
        if (respOk) atomicAdd(&ws->respawns, respOk);
        if (overflow) atomicAdd(&ws->overflow, overflow);
        if (eaten) atomicAdd(&ws->eaten, (uint32_t)eaten);
        if (shots) atomicAdd(&ws->shots, (uint32_t)shots);
    }
    if (foodDirty) {
        __syncwarp();
        for (int c = lane; c < kNumChunks; c += 32) d.food[(size_t)w * kNumChunks + c] = food[c];
    }
}

/* ------------------------------------------------------------------------- */
/* scan_kernel: rows are ordered (species, world, slot). One CTA. */

static_assert(kScanTile == 1024, "one thread per world of a tile");

__device__ __forceinline__ int warp_incl_scan(int v)
{
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(FULL, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

/* tile_sum_kernel: per-tile partial sums of the row scan, one CTA per tile of 1,024 worlds. (Accumulating
 * them with atomics from step_kernel put every concurrently running world on the same five addresses; the L2
 * serialises same-address atomics, and that backlog -- not DRAM -- set step_kernel's duration.) */
__global__ void __launch_bounds__(kScanTile) tile_sum_kernel(SimDev d, int cur)
{
    __shared__ int red[32][5];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int w = blockIdx.x * kScanTile + tid;
    int4 c = make_int4(0, 0, 0, 0);
    if (w < (int)d.numWorlds) c = d.speciesLive[w];
    int v[5] = {c.x, c.y, c.z, c.w, c.x + c.y + c.z + c.w};
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const int r = k == 4 ? __reduce_max_sync(FULL, v[k]) : __reduce_add_sync(FULL, v[k]);
        if (lane == 0) red[warp][k] = r;
    }
    __syncthreads();
    if (tid < 5) {
        int a = 0;
        for (int i = 0; i < kScanTile / 32; ++i) a = tid == 4 ? max(a, red[i][tid]) : a + red[i][tid];
        TileSum *ts = d.tileSum[cur] + blockIdx.x;
        if (tid == 4) ts->maxCount = a; else ts->c[tid] = a;
    }
}

/* One CTA per tile of 1,024 worlds, no inter-CTA dependency: every CTA reduces the per-tile partial sums
 * (tile_sum_kernel) to the species totals and to the prefix of its own tile, then scans its worlds. */
__global__ void __launch_bounds__(kScanTile) scan_kernel(SimDev d, int cur, int count_step)
{
    __shared__ int red[32][9];
    __shared__ int fin[9];
    __shared__ int wsum[32][5];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int W = (int)d.numWorlds, tile = blockIdx.x, nt = gridDim.x;
    const TileSum *ts = d.tileSum[cur];

    int v9[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};           /* totals[4], prefix of this tile[4], max count */
    for (int t = tid; t < nt; t += kScanTile) {
        const TileSum x = ts[t];
#pragma unroll
        for (int k = 0; k < 4; ++k) { v9[k] += x.c[k]; if (t < tile) v9[4 + k] += x.c[k]; }
        v9[8] = max(v9[8], x.maxCount);
    }
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        int v = v9[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const int u = __shfl_xor_sync(FULL, v, o);
            v = k == 8 ? max(v, u) : v + u;
        }
        if (lane == 0) red[warp][k] = v;
    }
    const int w = tile * kScanTile + tid;
    int4 c = make_int4(0, 0, 0, 0);
    if (w < W) c = d.speciesLive[w];
    int v[5] = {c.x, c.y, c.z, c.w, c.x + c.y + c.z + c.w};
    int inc[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        inc[k] = warp_incl_scan(v[k]);
        if (lane == 31) wsum[warp][k] = inc[k];
    }
    __syncthreads();
    if (tid < 9) {
        int a = 0;
        for (int i = 0; i < kScanTile / 32; ++i) a = tid == 8 ? max(a, red[i][tid]) : a + red[i][tid];
        fin[tid] = a;
    }
    __syncthreads();
    int base[5];
    int spBase = 0, agentsBefore = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        base[k] = spBase + fin[4 + k];                 /* species block start + rows of earlier tiles */
        spBase += fin[k];
        agentsBefore += fin[4 + k];
    }
    base[4] = agentsBefore;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        int b = base[k];
        for (int i = 0; i < warp; ++i) b += wsum[i][k];
        base[k] = b + inc[k] - v[k];
    }
    if (w < W) {
        d.rowBase[cur ^ 1][w] = make_int4(base[0], base[1], base[2], base[3]);
        d.worldOffset[w] = base[4];
    }
    if (tile == 0) {
        if (tid == 0) {
            for (int k = 0; k < 4; ++k) d.totals->speciesTotal[k] = (uint32_t)fin[k];
            d.totals->totalAgents = (uint32_t)spBase;
            d.totals->maxWorldCount = (uint32_t)fin[8];
            if (count_step) { d.totals->agentSteps += (unsigned long long)spBase; d.totals->steps += 1ull; }
        }
    }
}

/* scan_fused_kernel: tile sums and row scan in ONE launch. Every CTA publishes the sums of its tile, then waits until
 * all tiles have (a grid-wide barrier on a global counter: valid because the launcher only uses this kernel when every CTA
 * of the grid is resident at once), then proceeds exactly as scan_kernel. The last CTA through resets the counters, so a
 * captured graph can replay the launch with the same arguments. */
__global__ void __launch_bounds__(kScanTile) scan_fused_kernel(SimDev d, int cur, int count_step, unsigned int *sync2)
{
    __shared__ int red[32][9];
    __shared__ int fin[9];
    __shared__ int wsum[32][5];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int W = (int)d.numWorlds, tile = blockIdx.x, nt = gridDim.x;
    TileSum *ts = d.tileSum[cur];
    const int w = tile * kScanTile + tid;
    int4 c = make_int4(0, 0, 0, 0);
    if (w < W) c = d.speciesLive[w];
    int v[5] = {c.x, c.y, c.z, c.w, c.x + c.y + c.z + c.w};
    int inc[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        inc[k] = warp_incl_scan(v[k]);
        if (lane == 31) wsum[warp][k] = inc[k];
        const int mx = k == 4 ? __reduce_max_sync(FULL, v[k]) : 0;
        if (lane == 0 && k == 4) red[warp][0] = mx;
    }
    __syncthreads();
    if (tid < 5) {                                      /* this tile's sums (count of each species, largest world) */
        int a = 0;
        for (int i = 0; i < kScanTile / 32; ++i) a = tid == 4 ? max(a, red[i][0]) : a + wsum[i][tid];
        if (tid == 4) ts[tile].maxCount = a; else ts[tile].c[tid] = a;
    }
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        atomicAdd(&sync2[0], 1u);
        while (atomicAdd(&sync2[0], 0u) < (unsigned)nt) __nanosleep(64);
        __threadfence();
    }
    __syncthreads();
    int v9[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};           /* totals[4], prefix of this tile[4], max count */
    for (int t = tid; t < nt; t += kScanTile) {
        /* written by other CTAs of this launch: read through L2, never from this SM's L1 */
        const int4 xc = __ldcg(reinterpret_cast<const int4 *>(ts + t));
        const int4 xm = __ldcg(reinterpret_cast<const int4 *>(ts + t) + 1);
        const int xcv[4] = {xc.x, xc.y, xc.z, xc.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) { v9[k] += xcv[k]; if (t < tile) v9[4 + k] += xcv[k]; }
        v9[8] = max(v9[8], xm.x);
    }
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        int u = v9[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const int y = __shfl_xor_sync(FULL, u, o);
            u = k == 8 ? max(u, y) : u + y;
        }
        if (lane == 0) red[warp][k] = u;
    }
    __syncthreads();
    if (tid < 9) {
        int a = 0;
        for (int i = 0; i < kScanTile / 32; ++i) a = tid == 8 ? max(a, red[i][tid]) : a + red[i][tid];
        fin[tid] = a;
    }
    __syncthreads();
    int base[5];
    int spBase = 0, agentsBefore = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        base[k] = spBase + fin[4 + k];
        spBase += fin[k];
        agentsBefore += fin[4 + k];
    }
    base[4] = agentsBefore;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        int b = base[k];
        for (int i = 0; i < warp; ++i) b += wsum[i][k];
        base[k] = b + inc[k] - v[k];
    }
    if (w < W) {
        d.rowBase[cur ^ 1][w] = make_int4(base[0], base[1], base[2], base[3]);
        d.worldOffset[w] = base[4];
    }
    if (tid == 0) {
        if (tile == 0) {
            for (int k = 0; k < 4; ++k) d.totals->speciesTotal[k] = (uint32_t)fin[k];
            d.totals->totalAgents = (uint32_t)spBase;
            d.totals->maxWorldCount = (uint32_t)fin[8];
            if (count_step) { d.totals->agentSteps += (unsigned long long)spBase; d.totals->steps += 1ull; }
        }
        if (atomicAdd(&sync2[1], 1u) == (unsigned)nt - 1u) { sync2[0] = 0u; sync2[1] = 0u; __threadfence(); }
    }
}

/* ------------------------------------------------------------------------- */
/* export_kernel: one WARP per world, kExportWarps independent worlds per CTA    */
/* (one: worlds differ in population, and a CTA of four held its slot until    */
/* the slowest world was done -- 10 % of the warp slots idled).                */
/* No block barriers: every phase of a world is synchronised with __syncwarp,   */
/* so warps drift apart and hide each other's memory latency.                   */
/*                                                                             */
/* Two instantiations share the code:                                          */
/*   FAST  worlds of at most `slots` <= 64 agents (every world of a normal     */
/*         rollout). Nearest hits are 32-bit keys -- (t bits - bits(1.0f)) in   */
/*         26 bits, target slot in 6 -- so one native shared-memory atomicMin   */
/*         keeps them, and the keys of a tile of 32 observers (4 KB) fit beside */
/*         the agent arrays with 32 such CTAs resident per SM.                  */
/*   full  any world up to the capacity (<= 1024): 64-bit keys, 8 observers per */
/*         tile. Launched after the FAST kernel for the worlds it skipped, and  */
/*         alone when the host has no fresh bound on the population.            */

#ifndef MBK_EXPORT_WARPS
#define MBK_EXPORT_WARPS 1
#endif
constexpr int kExportWarps = MBK_EXPORT_WARPS;      /* worlds (warps) per CTA */
#ifndef MBK_EXPORT_SM_WARPS
#define MBK_EXPORT_SM_WARPS 32                       /* resident warps per SM the FAST kernel is built for (<= 64 registers) */
#endif
#ifndef MBK_EXPORT_FAST_TILE
#define MBK_EXPORT_FAST_TILE 32                      /* observers per sensor tile of the FAST kernel: 4 KB of keys */
#endif
constexpr int kExportFastMinBlocks = MBK_EXPORT_SM_WARPS / kExportWarps;
constexpr int kQueue = 64;           /* pending pairs / pending exact tests of a warp: < 32 carried over, + <= 32 pushed */
constexpr int kFastSlots = 64;       /* 6 slot bits in a 32-bit key */
constexpr int kFullTile = 8;

template <bool FAST> struct Sense;
template <> struct Sense<true> {
    typedef uint32_t Key;
    static constexpr int kJBits = 6;
    static constexpr uint32_t kMiss = 0xffffffffu;
    static __device__ __forceinline__ Key make(float t, int j) { return ((__float_as_uint(t) - 0x3f800000u) << 6) | (uint32_t)j; }
    static __device__ __forceinline__ int slot(Key k) { return (int)(k & 63u); }
    static __device__ __forceinline__ float t(Key k) { return __uint_as_float((k >> 6) + 0x3f800000u); }
};
template <> struct Sense<false> {
    typedef unsigned long long Key;
    static constexpr int kJBits = 10;
    static constexpr unsigned long long kMiss = ~0ull;
    static __device__ __forceinline__ Key make(float t, int j) { return ((Key)__float_as_uint(t) << 32) | (Key)(uint32_t)j; }
    static __device__ __forceinline__ int slot(Key k) { return (int)(uint32_t)k; }
    static __device__ __forceinline__ float t(Key k) { return __uint_as_float((uint32_t)(k >> 32)); }
};

/* per-warp shared memory: agent arrays sized for `slots` agents + the keys of one sensor tile */
struct ExportLayout {
    uint32_t best, pos, head, queue, pairq, candq, newrow, oldrow, theta, spec, total;
};
__host__ __device__ inline ExportLayout export_layout(uint32_t slots, uint32_t tile, uint32_t key_bytes)
{
    ExportLayout L;
    uint32_t o = 0;
    L.best = o; o += tile * kSensor * key_bytes;        /* 16-byte aligned first */
    L.pos = o; o += slots * 8;                          /* float2 */
    L.head = o; o += slots * 8;                         /* float2 unit heading (cull only) */
    L.queue = o; o += kQueue * 4;
    L.pairq = o; o += kQueue * 4;
    L.candq = o; o += kQueue * 4;                       /* 128 16-bit (observer, target) codes */
    L.newrow = o; o += slots * 4;
    /* old rows are dead once the columns are carried, the queues idle until then: share the bytes when they fit
     * (FAST: 28 single-warp CTAs of 56 slots then fit one SM's shared memory, not 26) */
    if (slots <= (uint32_t)kQueue) L.oldrow = L.queue; else { L.oldrow = o; o += slots * 4; }
    L.theta = o; o += slots * 4;
    L.spec = o; o += slots < 64u ? 64u : (slots + 15u) & ~15u;   /* FAST reads the 0xff sentinel at [63] */
    L.total = (o + 15u) & ~15u;
    return L;
}

size_t export_smem_bytes(uint32_t slots, uint32_t tile, bool fast)
{
    return (size_t)export_layout(slots, tile, fast ? 4u : 8u).total * kExportWarps;
}

/* Copy one column of every agent of the world from its old row to its new row; lanes run over
 * (agent, part) so that consecutive lanes touch consecutive bytes. V = bytes per lane access,
 * PARTS = row bytes / sizeof(V). All loads of a batch are issued before its stores. A record
 * created this step (old row < 0) is zero-filled (DEVIATIONS.md D3). */
template <typename V, int PARTS, int UNROLL>
__device__ __forceinline__ void carry_column(void *dst, const void *src, const int *newrow, const int *oldrow,
                                             int n, int lane, void *dst2 = nullptr, const void *ext = nullptr,
                                             unsigned long long ext_rows = 0)
{
    /* dst2: a second column that receives the same rows (fused shift: prev := cur); ext: rows below ext_rows are read
     * from this matrix instead of `src` (the caller's action matrix) */
    const int tasks = n * PARTS;
    V *dp = reinterpret_cast<V *>(dst);
    V *dp2 = reinterpret_cast<V *>(dst2);
    const V *sp = reinterpret_cast<const V *>(src);
    const V *ep = reinterpret_cast<const V *>(ext);
    for (int t0 = lane; t0 < tasks; t0 += 32 * UNROLL) {
        V v[UNROLL];
        uint32_t di[UNROLL];                            /* rows * PARTS < 2^32: mbots_create bounds the row capacity */
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const int t = t0 + 32 * u;
            v[u] = V();
            di[u] = 0;
            if (t < tasks) {
                const int a = t / PARTS, p = t - a * PARTS;
                const int o = oldrow[a];
                di[u] = (uint32_t)newrow[a] * PARTS + p;
                if (o >= 0) v[u] = (ep != nullptr && (unsigned long long)o < ext_rows ? ep : sp)[(size_t)o * PARTS + p];
            }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u)
            if (t0 + 32 * u < tasks) {
                dp[di[u]] = v[u];
                if (dp2 != nullptr) dp2[di[u]] = v[u];
            }
    }
}

/* One world, one warp -- or, for small launches, `nsplit` independent single-warp CTAs per world: every one of them
 * stages the whole world in its own shared memory (cheap: one pose per agent), then exports, carries and senses only its
 * share [n * split / nsplit, n * (split + 1) / nsplit) of the agents. No synchronisation between the parts; a step of a
 * small launch is then not bounded by one warp walking a whole world's sensor pass.
 * `my` = this warp's shared memory (export_layout). */
template <bool FAST, bool SPLIT, bool FUSED>
__device__ __forceinline__ void export_world(const SimDev &d, int cur, int with_sensors, uint32_t slots, uint32_t tile,
                                             uint32_t w, int n, unsigned char *my, int split, int nsplit, const StepOpts &opts)
{
    /* SPLIT and FUSED are compile-time: the common launch (whole world per warp) carries none of the part bookkeeping,
     * and the fused-shift variant none of the old prev-row loads (64 registers are tight) */
    typedef Sense<FAST> S;
    typedef typename S::Key Key;
    const int lane = threadIdx.x & 31;
    const ExportLayout L = export_layout(slots, tile, sizeof(Key));
    Key *best = reinterpret_cast<Key *>(my + L.best);
    float2 *pos = reinterpret_cast<float2 *>(my + L.pos);
    float2 *head = reinterpret_cast<float2 *>(my + L.head);
    uint32_t *queue = reinterpret_cast<uint32_t *>(my + L.queue);
    uint32_t *pairq = reinterpret_cast<uint32_t *>(my + L.pairq);
    uint16_t *candq = reinterpret_cast<uint16_t *>(my + L.candq);
    int *newrow = reinterpret_cast<int *>(my + L.newrow);
    int *oldrow = reinterpret_cast<int *>(my + L.oldrow);
    float *theta = reinterpret_cast<float *>(my + L.theta);
    uint8_t *spec = my + L.spec;

    const size_t base = (size_t)w * d.cap;
    const ObsSet &from = d.obs[cur];
    const ObsSet &to = d.obs[cur ^ 1];
    const int4 rb = d.rowBase[cur ^ 1][w];
    const int off = d.worldOffset[w];
    const float4 srew = d.speciesReward[w];
    const unsigned lt = (1u << lane) - 1u;
    const int a_lo = SPLIT ? (n * split) / nsplit : 0, a_hi = SPLIT ? (n * (split + 1)) / nsplit : n;   /* this part's agents */

    if (FAST && lane == 0) spec[63] = 0xffu;            /* species byte a missed ray decodes to (n <= 63) */
    /* ---- per agent (one lane each): new row, observation exports, small carried columns ---- */
    int seen[4] = {0, 0, 0, 0};                         /* agents of each species in lower passes */
    for (int s0 = 0; s0 < n; s0 += 32) {
        const int s = s0 + lane;
        const bool on = s < n;
        float4 p = make_float4(0.f, 0.f, 1.f, 0.f), sc = make_float4(0.f, 0.f, 0.f, 0.f);
        int sp = 0, hp = 0;
        if (on) {
            p = d.pose[base + s];
            sc = d.scratch[base + s];
            sp = d.species[base + s];
            hp = d.health[base + s];
        }
        /* new row = first row of (world, species) + rank among lower slots of the species */
        int rank = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const unsigned bsp = __ballot_sync(FULL, on && sp == k + 1);
            if (sp == k + 1) rank = seen[k] + __popc(bsp & lt);
            seen[k] += __popc(bsp);
        }
        if (on) {
            const int row = comp(rb, sp - 1) + rank;
            const int old = __float_as_int(sc.x);
            const bool mine = !SPLIT || (s >= a_lo && s < a_hi);
            MBK_CHECK((uint32_t)s < slots && sp >= 1 && sp <= 4 && row >= 0 && (uint32_t)row < d.numWorlds * d.cap);
            MBK_CHECK(old >= -1 && (old < 0 || (uint32_t)old < d.numWorlds * d.cap));
            /* carried small columns: issue the loads first */
            int4 pst = make_int4(0, 0, 0, 0);
            float2 ppos = make_float2(0.f, 0.f), psur = make_float2(0.f, 0.f);
            float prew = 0.f, phl = 0.f;
            /* the wide columns this agent's record is carried from: into L2 now, a few hundred instructions before the carry
             * loads them (they were written a step ago and have left the L2 at full size; the carry is where a world's warp
             * waits longest). 8,192 worlds: export 0.0890 -> 0.0868 ms; 65,536: 0.4807 -> 0.4776 (r4c). */
            if (old >= 0 && mine) {
                asm volatile("prefetch.global.L2 [%0];" ::"l"(from.hidden + (size_t)old * 16));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(from.semantic + (size_t)old * 32));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(from.depth + (size_t)old * 32));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(action_row(opts, from.action, old)));
            }
            if (old >= 0 && mine && !FUSED) {
                pst = reinterpret_cast<const int4 *>(from.prev_stats)[old];
                ppos = reinterpret_cast<const float2 *>(from.prev_position)[old];
                psur = reinterpret_cast<const float2 *>(from.prev_surrounding)[old];
                prew = from.prev_reward[old];
                phl = from.prev_health[old];
            }
            pos[s] = make_float2(p.x, p.y);
            newrow[s] = row;
            oldrow[s] = old;
            theta[s] = view_yaw(Rot{p.z, p.w});
            const Dir hd = cull_heading(Rot{p.z, p.w});
            head[s] = make_float2(hd.x, hd.y);
            spec[s] = (uint8_t)sp;
            if (mine) {
                /* updateObservations (:687-717), rewardSystem (:943-956), SensorOutputIndex (:783-784) */
                const uint32_t st = (uint32_t)__float_as_int(sc.w);
                reinterpret_cast<float2 *>(to.position)[row] = make_float2(p.x, p.y);
                to.health[row] = (float)hp;
                reinterpret_cast<float2 *>(to.surrounding)[row] = make_float2(sc.y, sc.z);
                reinterpret_cast<int4 *>(to.stats)[row] = make_int4(st & 1u, (st >> 1) & 1u, (st >> 2) & 1u, (st >> 3) & 1u);
                const float sterm = sp == 1 ? srew.x : sp == 2 ? srew.y : sp == 3 ? srew.z : srew.w;
                const float rew = with_sensors ? agent_reward(sterm, hp, (st & 4u) != 0, (st & 8u) != 0, (st & 2u) != 0) : 0.f;
                to.reward[row] = rew;
                if (FUSED) {                            /* shiftObservationsSystem (:1018-1036) incl. the U7 field mix-up */
                    pst = make_int4(st & 1u, st & 1u, (st >> 2) & 1u, (st >> 3) & 1u);
                    ppos = make_float2(p.x, p.y);
                    psur = make_float2(sc.y, sc.z);
                    prew = rew;
                    phl = (float)hp;
                }
                d.sensorIndex[off + s] = row;
                d.row[base + s] = row;
                to.owner[row] = (w << 10) | (uint32_t)s;
                if (!with_sensors) d.target[base + s] = (int16_t)-1;
                reinterpret_cast<int4 *>(to.prev_stats)[row] = pst;
                reinterpret_cast<float2 *>(to.prev_position)[row] = ppos;
                reinterpret_cast<float2 *>(to.prev_surrounding)[row] = psur;
                to.prev_reward[row] = prew;
                to.prev_health[row] = phl;
            }
        }
    }
    __syncwarp();

    /* ---- the sort: the wide columns of the observation record follow their agent ---- */
    const int na = a_hi - a_lo;
    const int *nr = newrow + a_lo, *orow = oldrow + a_lo;
    if (FUSED) {                                        /* shiftHiddenState (:1039-1048), prev_action (:1025-1030): one read, two writes */
        carry_column<float4, 4, 3>(to.hidden, from.hidden, nr, orow, na, lane, to.prev_hidden);
        carry_column<int2, 3, 4>(to.action, from.action, nr, orow, na, lane, to.prev_action, opts.actSrc, opts.actRows);
    } else {
        carry_column<float4, 4, 3>(to.hidden, from.hidden, nr, orow, na, lane);
        carry_column<float4, 4, 3>(to.prev_hidden, from.prev_hidden, nr, orow, na, lane);
        carry_column<int2, 3, 4>(to.action, from.action, nr, orow, na, lane, nullptr, opts.actSrc, opts.actRows);
        carry_column<int2, 3, 4>(to.prev_action, from.prev_action, nr, orow, na, lane);
    }
    /* updateSensorOutputIdx (:774-780): last frame's sensor rows become the prev rows */
    carry_column<int4, 2, 3>(to.prev_semantic, from.semantic, nr, orow, na, lane);
    carry_column<int4, 2, 3>(to.prev_depth, from.depth, nr, orow, na, lane);

    if (!with_sensors) return;

    /* ---- sensor (M/src/mw/device/bvh_raycast.cpp:193-305). Observers in tiles of `tile` (FAST: the whole world):
     *  1. cull. Within a tile, blocks of <= 32 observers are spread over the lanes as an (observer x target)
     *     grid -- 32 x 1, or for the last few observers 8 x 4 ... 1 x 32 -- so a lane keeps its observer in
     *     registers and steps through the targets with no index arithmetic. Per pair: conservative candidate
     *     rays of one fan (pair_run, ~0.27 per pair). Pairs with candidates are ballot-compacted into a pair queue.
     *  2. expand. 32 queued pairs at a time: a warp scan of their candidate counts, then every lane of a
     *     32-test chunk finds the pair that owns its slot by binary search over the scan (shuffles only, no
     *     divergence) and queues (observer, target, ray).
     *  3. exact test, 32 queued candidates at full lane occupancy (make_ray + ray/sphere, the oracle's
     *     arithmetic): nearest accepted hit kept with a shared-memory atomicMin on (t bits, target slot) --
     *     ties go to the lowest slot, and the order in which candidates are tested does not matter.
     *  4. one lane per (observer, 4 rays): decode, write 4 depth + 4 semantic bytes, finder hit */
    constexpr int JB = S::kJBits;
    for (int tile0 = a_lo; tile0 < a_hi; tile0 += (int)tile) {
        const int m = min((int)tile, a_hi - tile0);
        {
            uint4 *b4 = reinterpret_cast<uint4 *>(best);
            const int vecs = m * kSensor * (int)sizeof(Key) / 16;
            for (int idx = lane; idx < vecs; idx += 32) b4[idx] = make_uint4(~0u, ~0u, ~0u, ~0u);
        }
        __syncwarp();

        int qn = 0;                                     /* pending tests, warp-uniform, < 32 between chunks */
        auto test = [&](int ol, int j, int r) {
            const float2 qi = pos[tile0 + ol], qj = pos[j];
            const Ray ray = make_ray(theta[tile0 + ol], r);
            float t;
            MBK_CHECK(ol >= 0 && ol < m && j >= 0 && j < n && r >= 0 && r < kSensor);
            if (ray_sphere(ray, qi.x - qj.x, qi.y - qj.y, t) && t > kNearSphere) {
                MBK_CHECK(!FAST || (t >= 1.0f && t < 256.0f));
                atomicMin(&best[ol * kSensor + r], S::make(t, j));
            }
        };
        auto run_tests = [&](int first, int count) {
            if (lane < count) {
                const uint32_t e = queue[first + lane];
                test((int)(e >> (5 + JB)), (int)((e >> 5) & ((1u << JB) - 1u)), (int)(e & 31u));
            }
        };
        /* 32 pairs with candidates, one per lane. The first candidate of every pair is tested right here, at full
         * lane occupancy and with no queue traffic (most pairs have exactly one); the others are expanded into
         * the test queue. `flush`: also run a last partial chunk of queued tests. */
        auto expand = [&](uint32_t info, bool valid, bool flush) {
            if (valid) test((int)(info >> (17 + JB)), (int)((info >> 17) & ((1u << JB) - 1u)), (int)(info & 31u));
            const int c = valid ? (int)((info >> 5) & 63u) - 1 : 0;
            int inc = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(FULL, inc, o);
                if (lane >= o) inc += u;
            }
            const int total = __shfl_sync(FULL, inc, 31);
            const int exc = inc - c;
            for (int q0 = 0; q0 < total || (flush && qn > 0); q0 += 32) {
                const int q = q0 + lane;
                int own = 0;                            /* smallest lane with inc > q */
#pragma unroll
                for (int st = 16; st > 0; st >>= 1) {
                    const int v = __shfl_sync(FULL, inc, own + st - 1);
                    if (v <= q) own += st;
                }
                const uint32_t oi = __shfl_sync(FULL, info, own);
                const int r = (int)(oi & 31u) + q - __shfl_sync(FULL, exc, own) + 1;   /* candidate 0 was tested above */
                MBK_CHECK(q >= total || (qn + lane < kQueue && r >= 0 && r < kSensor));
                if (q < total) queue[qn + lane] = ((oi >> 17) << 5) | (uint32_t)r;
                qn += max(0, min(32, total - q0));
                __syncwarp();
                const bool last = q0 + 32 >= total;
                if (qn >= 32 || (flush && last)) {
                    const int cnt = min(32, qn);
                    qn -= cnt;
                    run_tests(qn, cnt);
                    __syncwarp();
                }
            }
        };

        int pqn = 0;                                    /* pending pairs with candidates, warp-uniform, < 32 between calls */
        /* stage 2 of the cull: 32 pairs that passed the wedge test, one per lane -> candidate spans */
        auto spans = [&](uint32_t code, bool valid) {
            const int ol = (int)(code >> JB), j = (int)(code & ((1u << JB) - 1u));
            const float2 pi = pos[tile0 + ol], hd = head[tile0 + ol], pj = pos[j];
            const RayRun rr = pair_run(pi.x - pj.x, pi.y - pj.y, Dir{hd.x, hd.y});
            /* first ray 0-4 | rays 5-10 | target 17.. (JB bits) | observer of the tile */
            const uint32_t info = (uint32_t)(rr.lo & 31) | ((uint32_t)rr.count << 5) | (code << 17);
            const bool has = valid && rr.count > 0;
            const unsigned bal = __ballot_sync(FULL, has);
            MBK_CHECK(pqn + __popc(bal & lt) < kQueue && (!valid || (rr.lo >= 0 && rr.lo + rr.count <= kSensor)));
            if (has) pairq[pqn + __popc(bal & lt)] = info;
            pqn += __popc(bal);
            __syncwarp();
            if (pqn >= 32) {
                pqn -= 32;
                expand(pairq[pqn + lane], true, false);
                __syncwarp();
            }
        };

        int cqn = 0;                                    /* pairs that passed stage 1, warp-uniform, < 32 between iterations */
        for (int sub0 = 0, ms = 0; sub0 < m; sub0 += ms) {
            /* a block of 2^lg observers costs 2^lg / 32 of a full one whatever part of it is occupied: take 16 (or 8) and leave
             * the rest to the next block rather than round 17-27 (9-13) observers up -- half a world is 16-20 observers */
            const int rem = m - sub0;
            const int lg = rem >= 28 ? 5 : rem >= 14 ? 4 : rem > 4 ? 3 : rem > 2 ? 2 : rem > 1 ? 1 : 0;
            ms = min(1 << lg, rem);
            const int il = lane & ((1 << lg) - 1), jl = lane >> lg, tpi = 32 >> lg;
            const bool on_i = il < ms;
            const int ol = sub0 + (on_i ? il : 0);      /* observer within the tile */
            const float2 pi = pos[tile0 + ol], hd = head[tile0 + ol];
            const int jend = on_i ? n : 0;
            /* two targets per trip: both position loads and both wedge tests are in flight together, and the queue
             * bookkeeping (counter, branch) is paid once; the queue of 16-bit codes holds < 32 + 64 entries */
            for (int j0 = 0; j0 < n; j0 += 2 * tpi) {
                const int ja = j0 + jl, jb = ja + tpi;
                const float2 pa = pos[min(ja, n - 1)], pb = pos[min(jb, n - 1)];
                /* stage 1: is the target anywhere near the two fans? (pair_maybe, sim_prims.cuh) */
                const bool ma = pair_maybe(pi.x - pa.x, pi.y - pa.y, Dir{hd.x, hd.y}) && ja < jend;
                const bool mb = pair_maybe(pi.x - pb.x, pi.y - pb.y, Dir{hd.x, hd.y}) && jb < jend;
                const unsigned ba = __ballot_sync(FULL, ma), bb = __ballot_sync(FULL, mb);
                const int ca = __popc(ba);
                MBK_CHECK(cqn + ca + __popc(bb) <= 2 * kQueue);
                if (ma) candq[cqn + __popc(ba & lt)] = (uint16_t)(((uint32_t)ol << JB) | (uint32_t)ja);
                if (mb) candq[cqn + ca + __popc(bb & lt)] = (uint16_t)(((uint32_t)ol << JB) | (uint32_t)jb);
                cqn += ca + __popc(bb);
                while (cqn >= 32) {
                    __syncwarp();
                    cqn -= 32;
                    spans((uint32_t)candq[cqn + lane], true);
                }
            }
        }
        __syncwarp();
        spans(lane < cqn ? (uint32_t)candq[lane] : 0u, lane < cqn);
        expand(lane < pqn ? pairq[lane] : 0u, lane < pqn, true);
        __syncwarp();                                   /* the decode below reads keys other lanes wrote with atomicMin */

        for (int idx = lane; idx < m * 8; idx += 32) {
            const int ol = idx >> 3, quad = idx & 7, i = tile0 + ol;
            const int row = newrow[i];
            uint32_t dep = 0, sem = 0;
            if (FAST) {
                /* A 32-bit key decodes without a hit / miss branch: the miss key 0xffffffff carries slot 63, whose
                 * species byte is the 0xff sentinel (FAST worlds hold at most 63 agents), and its time field is
                 * 255.99998f, i.e. depth 255 -- exactly what a miss exports. */
                const uint4 v = reinterpret_cast<const uint4 *>(best)[idx];
                const uint32_t key[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint32_t dv = (uint32_t)(uint8_t)(int)S::t((Key)key[k]);
                    const uint32_t sv = (uint32_t)spec[key[k] & 63u];
                    dep |= dv << (8 * k);
                    sem |= sv << (8 * k);
                }
                if (quad == 0) dep = (dep & ~0xffu) | ((uint32_t)row & 0xffu);          /* "sanity" overwrite (U9) */
                if (quad == kFinderRay / 4) {
                    const uint32_t fk = key[kFinderRay % 4];
                    d.target[base + i] = (int16_t)(fk != 0xffffffffu ? (int)(fk & 63u) : -1);
                }
            } else {
                Key key[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) key[k] = best[idx * 4 + k];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const bool hit = key[k] != S::kMiss;
                    const int slot = hit ? S::slot(key[k]) : -1;
                    uint32_t dv = hit ? (uint32_t)(uint8_t)(int)S::t(key[k]) : 255u;
                    const uint32_t sv = hit ? (uint32_t)spec[slot] : 0xffu;
                    if (quad == 0 && k == 0) dv = (uint32_t)row & 0xffu;      /* "sanity" overwrite (U9) */
                    if (quad == kFinderRay / 4 && k == kFinderRay % 4) d.target[base + i] = (int16_t)slot;
                    dep |= dv << (8 * k);
                    sem |= sv << (8 * k);
                }
            }
            reinterpret_cast<uint32_t *>(to.depth)[(size_t)row * 8 + quad] = dep;
            reinterpret_cast<uint32_t *>(to.semantic)[(size_t)row * 8 + quad] = sem;
        }
        __syncwarp();
    }
}

/* skip: FAST -> worlds with more than `slots` agents are left to the full kernel;
 *       full -> worlds with at most `skip_upto` agents were done by the FAST kernel (grid-stride over the
 *               worlds with a small grid, and nothing at all when no world is that large). */
template <bool FAST, bool SPLIT, bool FUSED>
__global__ void __launch_bounds__(kExportWarps * 32, FAST ? kExportFastMinBlocks : 16 / kExportWarps)
export_kernel(SimDev d, int cur, int with_sensors, uint32_t slots, uint32_t tile, int skip_upto, int split_log2, StepOpts opts)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5;
    if (!FAST && skip_upto >= 0 && (int)d.totals->maxWorldCount <= skip_upto) return;
    unsigned char *my = smem_raw + (size_t)export_layout(slots, tile, sizeof(typename Sense<FAST>::Key)).total * warp;
    if (FAST) {
        const uint32_t unit = blockIdx.x * kExportWarps + warp;  /* one (world, part) per warp */
        const uint32_t w = unit >> split_log2;
        if (w >= d.numWorlds) return;               /* whole warp: no block-level barrier anywhere */
        const int n = d.worldCount[w];
        if (n > (int)slots || n > 63) return;       /* slot 63 is the miss sentinel of the 32-bit keys */
        export_world<FAST, SPLIT, FUSED>(d, cur, with_sensors, slots, tile, w, n, my, (int)(unit & ((1u << split_log2) - 1u)),
                                         1 << split_log2, opts);
    } else {
        for (uint32_t w = blockIdx.x * kExportWarps + warp; w < d.numWorlds; w += gridDim.x * kExportWarps) {
            const int n = d.worldCount[w];
            if (n <= skip_upto) continue;
            export_world<FAST, false, FUSED>(d, cur, with_sensors, slots, tile, w, n, my, 0, 1, opts);
            __syncwarp();
        }
    }
}

/* ------------------------------------------------------------------------- */
/* shift_kernel: cur -> prev for every row below totalAgents                   */

#ifndef MBK_SHIFT_BATCH
#define MBK_SHIFT_BATCH 4
#endif
#ifndef MBK_SHIFT_CS
#define MBK_SHIFT_CS 0
#endif
template <typename V> __device__ __forceinline__ V shift_ld(const V *p)
{
#if MBK_SHIFT_CS
    return __ldcs(p);
#else
    return *p;
#endif
}
template <typename V> __device__ __forceinline__ void shift_st(V *p, V v)
{
#if MBK_SHIFT_CS
    __stcs(p, v);
#else
    *p = v;
#endif
}
/* Plain column copy with MBK_SHIFT_BATCH independent loads in flight per thread before the first store (the
 * columns never alias; without the batching every store waits on its own load and an SM keeps only ~32 KB in
 * flight). The last, partial batch is predicated, not serialised. */
template <typename V, typename F>
__device__ __forceinline__ void shift_column(V *__restrict__ p, const V *__restrict__ s, size_t n, size_t t0,
                                             size_t stride, F fix)
{
    for (size_t i = t0; i < n; i += MBK_SHIFT_BATCH * stride) {
        V v[MBK_SHIFT_BATCH];
#pragma unroll
        for (int u = 0; u < MBK_SHIFT_BATCH; ++u)
            if (i + u * stride < n) v[u] = shift_ld(s + i + u * stride);
#pragma unroll
        for (int u = 0; u < MBK_SHIFT_BATCH; ++u)
            if (i + u * stride < n) shift_st(p + i + u * stride, fix(v[u]));
    }
}

#ifndef MBK_SHIFT_MIN_BLOCKS
#define MBK_SHIFT_MIN_BLOCKS 4
#endif
#ifndef MBK_SHIFT_CTAS_PER_SM
#define MBK_SHIFT_CTAS_PER_SM 4
#endif
__global__ void __launch_bounds__(256, MBK_SHIFT_MIN_BLOCKS) shift_kernel(SimDev d, int cur)
{
    const ObsSet &o = d.obs[cur];
    const size_t A = d.totals->totalAgents;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t t0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    auto same = [](auto v) { return v; };
    /* hidden: 4 x 16 B per row */
    shift_column(reinterpret_cast<float4 *>(o.prev_hidden), reinterpret_cast<const float4 *>(o.hidden), A * 4, t0, stride, same);
    /* action: 24 B per row = 3 x 8 B */
    shift_column(reinterpret_cast<int2 *>(o.prev_action), reinterpret_cast<const int2 *>(o.action), A * 3, t0, stride, same);
    /* stats, with the reference's field mix-up (src/sim/sim.cpp:1034, U7) */
    shift_column(reinterpret_cast<int4 *>(o.prev_stats), reinterpret_cast<const int4 *>(o.stats), A, t0, stride,
                 [](int4 v) { return make_int4(v.x, v.x, v.z, v.w); });
    shift_column(reinterpret_cast<float2 *>(o.prev_position), reinterpret_cast<const float2 *>(o.position), A, t0, stride, same);
    shift_column(reinterpret_cast<float2 *>(o.prev_surrounding), reinterpret_cast<const float2 *>(o.surrounding), A, t0, stride, same);
    shift_column(o.prev_health, o.health, A, t0, stride, same);
    shift_column(o.prev_reward, o.reward, A, t0, stride, same);
}

/* ------------------------------------------------------------------------- */
/* action_write_kernel: the caller's action rows -> the action export, row count   */
/* resolved ON THE DEVICE (totals->totalAgents), so that a host that keeps its     */
/* actions (or its policy) on the GPU never has to wait for a step to learn how    */
/* many rows to write (learn/training_loop.py:136 writes them through the view).   */

__global__ void __launch_bounds__(256) action_write_kernel(SimDev d, int cur, const int2 *__restrict__ src, unsigned long long src_rows)
{
    unsigned long long rows = d.totals->totalAgents;
    if (rows > src_rows) rows = src_rows;
    int2 *dst = reinterpret_cast<int2 *>(d.obs[cur].action);
    const size_t n = (size_t)rows * 3, stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += 4 * stride) {
        int2 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (i + u * stride < n) v[u] = src[i + u * stride];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (i + u * stride < n) dst[i + u * stride] = v[u];
    }
}

void launch_action_write(const SimDev &d, int cur, const void *src, unsigned long long src_rows, uint32_t rows_upper, cudaStream_t s)
{
    size_t vecs = (size_t)rows_upper * 3;
    size_t blocks = (vecs + 256 * 4 - 1) / (256 * 4);
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks == 0) blocks = 1;
    action_write_kernel<<<(unsigned)blocks, 256, 0, s>>>(d, cur, reinterpret_cast<const int2 *>(src), src_rows);
}

/* ------------------------------------------------------------------------- */
/* episode statistics and debug gathers (not on the hot path)                  */

/* out[16] must be zero on entry; every CTA adds its share */
__global__ void __launch_bounds__(256) stats_reduce_kernel(SimDev d, long long *out)
{
    __shared__ long long acc[14];
    if (threadIdx.x < 14) acc[threadIdx.x] = 0;
    __syncthreads();
    long long l[14];
    for (int k = 0; k < 14; ++k) l[k] = 0;
    for (uint32_t w = blockIdx.x * blockDim.x + threadIdx.x; w < d.numWorlds; w += gridDim.x * blockDim.x) {
        const int4 c = d.speciesCount[w], h = d.speciesHealth[w];
        const WorldStats ws = d.wstats[w];
        l[0] += c.x; l[1] += c.y; l[2] += c.z; l[3] += c.w;
        l[4] += h.x; l[5] += h.y; l[6] += h.z; l[7] += h.w;
        l[8] += ws.births; l[9] += ws.deaths; l[10] += ws.respawns; l[11] += ws.overflow;
        l[12] += ws.eaten; l[13] += ws.shots;
    }
    for (int k = 0; k < 14; ++k) {
        long long v = l[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(reinterpret_cast<unsigned long long *>(&acc[k]), (unsigned long long)v);
    }
    __syncthreads();
    if (threadIdx.x < 14 && acc[threadIdx.x])
        atomicAdd(reinterpret_cast<unsigned long long *>(&out[2 + threadIdx.x]), (unsigned long long)acc[threadIdx.x]);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        out[0] = (long long)d.totals->agentSteps;
        out[1] = (long long)d.totals->steps;
    }
}

/* what: 0 pose (A,4) float, 1 health, 2 species, 3 target (A,1) int32; agent order */
__global__ void debug_gather_kernel(SimDev d, int what, void *out)
{
    const uint32_t w = blockIdx.x;
    const int n = d.worldCount[w], off = d.worldOffset[w];
    const size_t base = (size_t)w * d.cap;
    for (int s = threadIdx.x; s < n; s += blockDim.x) {
        if (what == 0) reinterpret_cast<float4 *>(out)[off + s] = d.pose[base + s];
        else if (what == 1) reinterpret_cast<int32_t *>(out)[off + s] = d.health[base + s];
        else if (what == 2) reinterpret_cast<int32_t *>(out)[off + s] = d.species[base + s];
        else reinterpret_cast<int32_t *>(out)[off + s] = d.target[base + s];
    }
}

/* ------------------------------------------------------------------------- */
/* launchers                                                                   */

void launch_init(const SimDev &d, cudaStream_t s)
{
    uint32_t threads = d.cap < 64 ? 64 : d.cap;
    init_kernel<<<d.numWorlds, threads, 0, s>>>(d);
}

/* Dynamic shared memory above 48 KB needs an opt-in per kernel: done once per handle for the largest geometry the
 * handle can ask for (its capacity), outside any stream capture. */
void prepare_kernels(const SimDev &d)
{
    static size_t granted[64][2];                       /* per device: never shrink what an earlier handle asked for */
    int dev = 0;
    cudaGetDevice(&dev);
    dev &= 63;
    const size_t s1 = step_smem_bytes(d.cap), s2 = export_smem_bytes(d.cap, kFullTile, false);
    if (s1 > 48 * 1024 && s1 > granted[dev][0]) {
        cudaFuncSetAttribute(step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s1);
        granted[dev][0] = s1;
    }
    if (s2 > 48 * 1024 && s2 > granted[dev][1]) {
        cudaFuncSetAttribute(export_kernel<false, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s2);
        cudaFuncSetAttribute(export_kernel<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s2);
        granted[dev][1] = s2;
    }
}

static StepOpts make_opts(const void *act_src, unsigned long long act_rows, int fuse_shift)
{
    StepOpts o;
    o.actSrc = reinterpret_cast<const int2 *>(act_src);
    o.actRows = act_src ? act_rows : 0ull;
    o.fuseShift = fuse_shift;
    return o;
}

void launch_step(const SimDev &d, int cur, uint32_t slots, const void *act_src, unsigned long long act_rows, cudaStream_t s)
{
    /* slots: a bound on the live agents of any world before the step; the capacity always is one */
    if (slots == 0 || slots > d.cap) slots = d.cap;
    const size_t smem = step_smem_bytes(slots);
    step_kernel<<<(d.numWorlds + kStepWarps - 1) / kStepWarps, kStepWarps * 32, smem, s>>>(d, cur, slots, make_opts(act_src, act_rows, 0));
}

/* Returns the kernels launched: one (scan_fused_kernel) while every tile's CTA fits on the GPU at once -- its grid-wide
 * barrier needs that (SM count x occupancy, queried per device: >= 151,552 worlds on a B200) -- otherwise tile sums and
 * row scan as two. */
int launch_scan(const SimDev &d, int cur, int count_step, cudaStream_t s)
{
    const unsigned tiles = (d.numWorlds + kScanTile - 1) / kScanTile;
    static int fused = -1;
    static unsigned resident[64];                       /* per device: CTAs of scan_fused_kernel the GPU holds at once */
    if (fused < 0) { const char *e = getenv("MBOTS_SCAN_FUSED"); fused = !(e && e[0] == '0'); }
    int dev = 0;
    cudaGetDevice(&dev);
    dev &= 63;
    if (resident[dev] == 0) {
        int sms = 0, per_sm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, scan_fused_kernel, kScanTile, 0);
        resident[dev] = (unsigned)(sms > 0 && per_sm > 0 ? sms * per_sm : 1);
    }
    if (fused && tiles <= resident[dev] && d.scanSync != nullptr) {
        scan_fused_kernel<<<tiles, kScanTile, 0, s>>>(d, cur, count_step, d.scanSync);
        return 1;
    }
    tile_sum_kernel<<<tiles, kScanTile, 0, s>>>(d, cur);
    scan_kernel<<<tiles, kScanTile, 0, s>>>(d, cur, count_step);
    return 2;
}

template <bool FAST, bool SPLIT, bool FUSED>
static void launch_export_inst(const SimDev &d, int cur, int with_sensors, uint32_t slots, uint32_t tile, int skip_upto,
                               int split_log2, const StepOpts &opts, uint32_t grid, size_t smem, cudaStream_t s)
{
    export_kernel<FAST, SPLIT, FUSED><<<grid, kExportWarps * 32, smem, s>>>(d, cur, with_sensors, slots, tile, skip_upto, split_log2, opts);
}

template <bool FAST>
static void launch_export_variant(const SimDev &d, int cur, int with_sensors, uint32_t slots, uint32_t tile, int skip_upto,
                                  int split_log2, const StepOpts &opts, cudaStream_t s)
{
    const size_t smem = export_smem_bytes(slots, tile, FAST);
    uint32_t grid = ((d.numWorlds << split_log2) + kExportWarps - 1) / kExportWarps;
    if (!FAST && skip_upto >= 0 && grid > 148u * 4u) grid = 148u * 4u;   /* clean-up pass: few worlds, if any */
    const bool split = FAST && split_log2 > 0, fused = opts.fuseShift != 0;
    if (split) {
        if (fused) launch_export_inst<FAST, FAST, true>(d, cur, with_sensors, slots, tile, skip_upto, split_log2, opts, grid, smem, s);
        else launch_export_inst<FAST, FAST, false>(d, cur, with_sensors, slots, tile, skip_upto, split_log2, opts, grid, smem, s);
    } else {
        if (fused) launch_export_inst<FAST, false, true>(d, cur, with_sensors, slots, tile, skip_upto, 0, opts, grid, smem, s);
        else launch_export_inst<FAST, false, false>(d, cur, with_sensors, slots, tile, skip_upto, 0, opts, grid, smem, s);
    }
}

/* Parts per world of the FAST export kernel (log2): a full GPU wants ~32 resident single-warp CTAs per SM; below that a
 * world is split over 2, 4 or 8 CTAs so that the launch is not bounded by one warp walking a whole world's sensor pass.
 * MBOTS_EXPORT_SPLIT_LOG2 overrides (0..3). */
static int export_split_log2(uint32_t numWorlds)
{
    static int forced = -2;
    if (forced == -2) {
        const char *e = getenv("MBOTS_EXPORT_SPLIT_LOG2");
        forced = e ? atoi(e) : -1;
        if (forced < 0 || forced > 3) forced = -1;
    }
    if (forced >= 0) return forced;
    const uint32_t per_sm = numWorlds / 148u;
    return per_sm >= 64u ? 0 : per_sm >= 24u ? 1 : per_sm >= 8u ? 2 : 3;
}

/* bound: the largest world the caller expects after this step (0 = unknown). Worlds above it are still
 * handled -- by the full kernel -- so a wrong bound costs time, never correctness. Returns the launches made. */
int launch_export(const SimDev &d, int cur, int with_sensors, uint32_t bound, const void *act_src, unsigned long long act_rows,
                  int fuse_shift, cudaStream_t s, cudaStream_t s_full)
{
    /* s_full: the stream of the full-capacity clean-up launch. The two variants touch disjoint worlds, so under a graph
     * capture the caller forks it onto a side stream and the (normally empty) clean-up pass leaves the critical path. */
    const StepOpts opts = make_opts(act_src, act_rows, fuse_shift);
    uint32_t fast_slots = (bound + 7u) & ~7u;
    /* a bound beyond what the fast variant can take (64 slots, worlds of up to 63 agents) only means that the largest worlds
     * go to the full variant -- never that all of them do */
    if (fast_slots > (uint32_t)kFastSlots) fast_slots = (uint32_t)kFastSlots;
    if (fast_slots > d.cap) fast_slots = d.cap & ~7u;
    if (bound == 0 || fast_slots < 8u) {
        launch_export_variant<false>(d, cur, with_sensors, d.cap, kFullTile, -1, 0, opts, s);
        return 1;
    }
    /* 32 observers per sensor tile: 56 slots + 4 KB of keys = 6.1 KB per single-warp CTA, 32 of them per SM.
     * (One tile for the whole world -- 40 observers, 26-28 CTAs per SM -- measured 1 % slower.) */
    uint32_t tile = (uint32_t)MBK_EXPORT_FAST_TILE;
    if (tile > fast_slots) tile = fast_slots;
    launch_export_variant<true>(d, cur, with_sensors, fast_slots, tile, 0, export_split_log2(d.numWorlds), opts, s);
    launch_export_variant<false>(d, cur, with_sensors, d.cap, kFullTile, (int)(fast_slots < 63u ? fast_slots : 63u), 0, opts,
                                 s_full ? s_full : s);
    return 2;
}

void launch_shift(const SimDev &d, int cur, uint32_t rows_upper, cudaStream_t s)
{
    /* enough CTAs to cover the hidden-state column at one 16-B vector per thread, capped */
    size_t vecs = (size_t)rows_upper * 4;
    size_t blocks = (vecs + 255) / 256;
    if (blocks > 148 * MBK_SHIFT_CTAS_PER_SM) blocks = 148 * MBK_SHIFT_CTAS_PER_SM;
    if (blocks == 0) blocks = 1;
    shift_kernel<<<(unsigned)blocks, 256, 0, s>>>(d, cur);
}

void launch_stats_reduce(const SimDev &d, long long *out16, cudaStream_t s)
{
    cudaMemsetAsync(out16, 0, 16 * sizeof(long long), s);
    unsigned blocks = (d.numWorlds + 255) / 256;
    if (blocks > 148u * 4u) blocks = 148u * 4u;
    stats_reduce_kernel<<<blocks, 256, 0, s>>>(d, out16);
}

/* Consistency of a restored state (mbots_load_state): flags[0] |= 1 a world count above the capacity, 2 a row outside
 * [0, totalAgents), 4 a finder target outside [-1, count), 8 a species outside 1..4; flags[1] += live agents. */
__global__ void validate_state_kernel(SimDev d, uint32_t total_agents, unsigned int *flags)
{
    const uint32_t w = blockIdx.x;
    const int n = d.worldCount[w];
    if (n < 0 || n > (int)d.cap) { if (threadIdx.x == 0) atomicOr(&flags[0], 1u); return; }
    if (threadIdx.x == 0) atomicAdd(&flags[1], (unsigned)n);
    const size_t base = (size_t)w * d.cap;
    for (int s = threadIdx.x; s < n; s += blockDim.x) {
        const int row = d.row[base + s], tgt = d.target[base + s], sp = d.species[base + s];
        if (row < 0 || (uint32_t)row >= total_agents) atomicOr(&flags[0], 2u);
        if (tgt < -1 || tgt >= n) atomicOr(&flags[0], 4u);
        if (sp < 1 || sp > kNumSpecies) atomicOr(&flags[0], 8u);
    }
}

void launch_validate_state(const SimDev &d, uint32_t total_agents, unsigned int *flags2, cudaStream_t s)
{
    cudaMemsetAsync(flags2, 0, 2 * sizeof(unsigned int), s);
    validate_state_kernel<<<d.numWorlds, 64, 0, s>>>(d, total_agents, flags2);
}

/* Debug aid for the pointer-lifetime rule of include/mbots.h (row exports ping-pong between two sets, so a view fetched
 * before a step points at the DEAD set afterwards): poison the two caller-writable columns of the dead set after a step,
 * and before the next one count the words that no longer hold the poison -- a write through a stale view. */
constexpr uint32_t kPoison = 0xDEADA110u;
__global__ void __launch_bounds__(256) poison_kernel(SimDev d, int set, size_t rows, int check, unsigned int *bad)
{
    uint32_t *act = reinterpret_cast<uint32_t *>(d.obs[set].action);
    uint32_t *hid = reinterpret_cast<uint32_t *>(d.obs[set].hidden);
    const size_t na = rows * 6, nh = rows * 16, stride = (size_t)gridDim.x * blockDim.x;
    unsigned int mine = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < na + nh; i += stride) {
        uint32_t *p = i < na ? act + i : hid + (i - na);
        if (check) mine += *p != kPoison; else *p = kPoison;
    }
    if (check && mine) atomicAdd(bad, mine);
}

void launch_poison(const SimDev &d, int set, size_t rows, int check, unsigned int *bad, cudaStream_t s)
{
    if (check) cudaMemsetAsync(bad, 0, sizeof(unsigned int), s);
    poison_kernel<<<148 * 4, 256, 0, s>>>(d, set, rows, check, bad);
}

void launch_debug_gather(const SimDev &d, int what, void *out, cudaStream_t s)
{
    debug_gather_kernel<<<d.numWorlds, 128, 0, s>>>(d, what, out);
}

} // namespace mbk
