/* sim_dev.h -- device-side data layout of one simulator handle.
 *
 * Replaces the reference's generic ECS tables (Agent / AgentObservationArchetype /
 * ChunkInfo / SpeciesInfoArchetype, src/sim/types.hpp:187-252) with a fixed
 * structure-of-arrays:
 *
 *   agent state      world-major: index = world * cap + slot, slots [0, count) live
 *   observation rows species-major: row = rowBase[world][species] + rank in world,
 *                    the order the reference's global sort by SpeciesObservation
 *                    produces (src/sim/sim.cpp:1148-1149). Two complete sets of row
 *                    columns; each step gathers from set `cur` and writes set
 *                    `cur^1` (that IS the sort: every column travels with its agent).
 */
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace mbk {

/* Row-indexed exports of one set (AgentObservationArchetype columns + sensor rows). */
struct ObsSet {
    int32_t *action, *prev_action;        /* 6 / row */
    float *hidden, *prev_hidden;          /* 16 */
    float *reward, *prev_reward;          /* 1 */
    float *position, *prev_position;      /* 2 */
    float *health, *prev_health;          /* 1 */
    float *surrounding, *prev_surrounding;/* 2 */
    int32_t *stats, *prev_stats;          /* 4 */
    int8_t *semantic, *prev_semantic;     /* 32 */
    uint8_t *depth, *prev_depth;          /* 32 */
    /* row -> (local world << 10) | slot of the agent that owns the row (written by the export kernel): the identity
     * the policy's action sampling is keyed on, so that an agent's draw does not depend on how worlds are sharded */
    uint32_t *owner;
    uint32_t worldIdOffset;               /* global id of local world 0 (copy of SimDev::worldIdOffset) */
};

/* Per-world cumulative event counters (episode statistics). */
struct WorldStats { uint32_t births, deaths, respawns, overflow, eaten, shots, pad0, pad1; };

/* Partial sums of the row scan over one tile of kScanTile consecutive worlds. */
constexpr int kScanTile = 1024;
struct TileSum { int c[4]; int maxCount; int pad[3]; };

struct Totals {
    uint32_t totalAgents;     /* SimBridge::totalNumAgents */
    uint32_t maxWorldCount;   /* most agents in any world */
    unsigned long long agentSteps, steps;
    uint32_t speciesTotal[4]; /* rows per species block */
};

struct SimDev {
    uint32_t numWorlds, cap, initAgents, worldIdOffset, seed;

    /* agent state, world-major [numWorlds * cap] */
    float4 *pose;             /* x, y, rot.w, rot.z */
    int32_t *health;
    uint8_t *species;         /* 1..4 */
    int16_t *target;          /* finder hit of the last sensor pass: slot in own world, -1 none */
    float4 *scratch;          /* step -> export hand-off: old row (int bits), surr.x, surr.y, stat bits */
    int32_t *row;             /* export -> step hand-off: the agent's observation row in the current set */

    /* per world */
    int32_t *worldCount;      /* live agents */
    int32_t *worldOffset;     /* exclusive scan of worldCount (Manager::agentWorldOffsets) */
    uint2 *rngKey;            /* split_i(initKey(seed), 0, global world id) */
    uint32_t *rngCount;       /* draws consumed */
    int32_t *numFood;         /* Sim::currentNumFood */
    uint2 *food;              /* [numWorlds * 48]: byte p<5 = (x<<4|y) of package p, byte 5 = has-food mask */
    int4 *speciesLive;        /* rows per (world, species) after respawn: input of the row scan */
    TileSum *tileSum[2];      /* per-tile sums of speciesLive (tile_sum_kernel), indexed by the parity of the step */
    int4 *rowBase[2];         /* first row of (world, species) in obs set 0 / 1 */
    int4 *speciesCount;       /* export SpeciesCount (tracker counts, before respawn) */
    float4 *speciesReward;    /* SpeciesReward */
    int4 *speciesHealth;      /* tracker health sums of the last step */
    WorldStats *wstats;

    int32_t *sensorIndex;     /* agent order (worldOffset + slot) -> row */
    int32_t *done;            /* zeros (U10) */
    int32_t *reset;           /* zeros, ignored (U10) */
    Totals *totals;
    unsigned int *scanSync;   /* two counters of scan_fused_kernel's grid barrier (zero between launches) */

    ObsSet obs[2];
};

/* launchers (sim_kernels.cu) */
void prepare_kernels(const SimDev &d);
void launch_init(const SimDev &d, cudaStream_t s);
void launch_step(const SimDev &d, int cur, uint32_t slots, const void *act_src, unsigned long long act_rows, cudaStream_t s);
int launch_scan(const SimDev &d, int cur, int count_step, cudaStream_t s);
int launch_export(const SimDev &d, int cur, int with_sensors, uint32_t bound, const void *act_src, unsigned long long act_rows,
                  int fuse_shift, cudaStream_t s, cudaStream_t s_full = nullptr);
void launch_shift(const SimDev &d, int cur, uint32_t rows_upper, cudaStream_t s);
void launch_action_write(const SimDev &d, int cur, const void *src, unsigned long long src_rows, uint32_t rows_upper, cudaStream_t s);
void launch_stats_reduce(const SimDev &d, long long *out16, cudaStream_t s);
void launch_debug_gather(const SimDev &d, int what, void *out, cudaStream_t s);
void launch_poison(const SimDev &d, int set, size_t rows, int check, unsigned int *bad, cudaStream_t s);
void launch_validate_state(const SimDev &d, uint32_t total_agents, unsigned int *flags2, cudaStream_t s);
/* launchers (policy_kernels.cu) */
void launch_pack_obs(const ObsSet &o, int is_prev, uint32_t row_begin, uint32_t rows, float *out, cudaStream_t s);
void launch_sample(const ObsSet &o, uint32_t row_begin, uint32_t rows, const float *weights, uint32_t seed, uint32_t draw,
                   int32_t *out_index, cudaStream_t s);
size_t step_smem_bytes(uint32_t slots);
size_t export_smem_bytes(uint32_t slots, uint32_t tile, bool fast);

} // namespace mbk
