/* mbots.h -- C-ABI of the B200-native madrona-bots step library (libmbots.so).
 *
 * This is the drop-in boundary for the per-step hot path of llGuy/madrona-bots:
 * every entry point replaces one member of the reference's `Manager`
 * (src/entry/mgr.hpp:28-62, src/entry/mgr.cpp) that the nanobind module
 * `madrona_bots.SimManager` binds (src/entry/entry.cpp:16-44).
 *
 * Conventions
 *   - plain C, no exceptions cross the boundary; every call returns 0 on
 *     success and a negative MBOTS_E* code otherwise; mbots_last_error() gives
 *     the message of the last failure on the calling thread.
 *   - the library OWNS all device memory of a handle; pointers returned by
 *     mbots_tensor() stay valid until the next mbots_step*() call on that
 *     handle (row-indexed exports are double-buffered: a step re-packs every
 *     agent's row, exactly as the reference's per-step archetype sorts do, and
 *     the caller is already required to re-fetch tensors every step because
 *     their leading dimension changes -- src/entry/mgr.cpp:199-422).
 *   - one host thread per handle; calls on a handle are serialised by the
 *     caller. One handle per GPU per process is the intended use.
 *   - `stream` arguments are CUDA stream handles (cudaStream_t / CUstream)
 *     passed as void*; NULL = the legacy default stream.
 */
#ifndef MBOTS_H
#define MBOTS_H

#include <stdint.h>
#include "mbots_policy.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mbots_sim mbots_sim;

/* Same order and meaning as the reference's ExportID (src/sim/sim.hpp:18-55). */
enum mbots_export_id {
    MBOTS_EXPORT_RESET = 0,            /* int32  (numWorlds, 1)  accepted, ignored (U10) */
    MBOTS_EXPORT_ACTION,               /* int32  (A, 6)   caller writes  */
    MBOTS_EXPORT_PREV_ACTION,          /* int32  (A, 6) */
    MBOTS_EXPORT_HIDDEN_STATE,         /* float32 (A, 16) caller writes  */
    MBOTS_EXPORT_PREV_HIDDEN_STATE,    /* float32 (A, 16) */
    MBOTS_EXPORT_REWARD,               /* float32 (A, 1) */
    MBOTS_EXPORT_PREV_REWARD,          /* float32 (A, 1) */
    MBOTS_EXPORT_DONE,                 /* int32  (A, 1)   always 0 (U10), agent order */
    MBOTS_EXPORT_POSITION,             /* float32 (A, 2) */
    MBOTS_EXPORT_PREV_POSITION,        /* float32 (A, 2) */
    MBOTS_EXPORT_HEALTH,               /* float32 (A, 1)  numeric value (U4) */
    MBOTS_EXPORT_PREV_HEALTH,          /* float32 (A, 1) */
    MBOTS_EXPORT_SURROUNDING,          /* float32 (A, 2) */
    MBOTS_EXPORT_PREV_SURROUNDING,     /* float32 (A, 2) */
    MBOTS_EXPORT_SENSOR_SEMANTIC,      /* int8   (A, 32) */
    MBOTS_EXPORT_SENSOR_DEPTH,         /* uint8  (A, 32) */
    MBOTS_EXPORT_PREV_SENSOR_SEMANTIC, /* int8   (A, 32) */
    MBOTS_EXPORT_PREV_SENSOR_DEPTH,    /* uint8  (A, 32) */
    MBOTS_EXPORT_STATS,                /* int32  (A, 4) */
    MBOTS_EXPORT_PREV_STATS,           /* int32  (A, 4) */
    MBOTS_EXPORT_SENSOR_INDEX,         /* int32  (A, 1)  agent (world, slot) order -> obs row */
    MBOTS_EXPORT_SPECIES_COUNT,        /* int32  (numWorlds, 4) */
    MBOTS_NUM_EXPORTS,

    /* Not part of the reference: internal state exposed for parity audits.
     * "agent order" = (world, slot), the order of the reference's Agent table
     * after its per-step world sort (src/sim/sim.cpp:1129-1130). */
    MBOTS_DEBUG_WORLD_COUNT = 100,     /* int32  (numWorlds, 1) live agents per world */
    MBOTS_DEBUG_WORLD_OFFSET,          /* int32  (numWorlds, 1) exclusive scan of the above */
    MBOTS_DEBUG_RNG_COUNT,             /* int32  (numWorlds, 1) draws consumed so far */
    MBOTS_DEBUG_NUM_FOOD,              /* int32  (numWorlds, 1) */
    MBOTS_DEBUG_FOOD,                  /* uint8  (numWorlds*48, 8): cells[5] (x<<4|y), mask, 0, 0 */
    MBOTS_DEBUG_POSE,                  /* float32 (A, 4) x, y, quat.w, quat.z; agent order */
    MBOTS_DEBUG_AGENT_HEALTH,          /* int32  (A, 1) agent order */
    MBOTS_DEBUG_AGENT_SPECIES,         /* int32  (A, 1) agent order */
    MBOTS_DEBUG_AGENT_TARGET,          /* int32  (A, 1) finder hit as slot in own world, -1 none */
    MBOTS_DEBUG_ROW_BASE,              /* int32  (numWorlds, 4) first obs row of (world, species) */
    MBOTS_DEBUG_ROW_OWNER              /* int32  (A, 1) observation row -> (local world << 10) | slot of the agent that owns it */
};

enum mbots_dtype {
    MBOTS_I32 = 0,
    MBOTS_F32 = 1,
    MBOTS_I8 = 2,
    MBOTS_U8 = 3
};

enum mbots_error {
    MBOTS_OK = 0,
    MBOTS_EINVAL = -1,   /* bad argument */
    MBOTS_ECUDA = -2,    /* a CUDA call failed */
    MBOTS_ENOMEM = -3,
    MBOTS_ESTATE = -4    /* call not valid in this state */
};

/* Sizes fixed by the reference (src/entry/mgr.cpp:104-113,132-133). */
#define MBOTS_NUM_SPECIES 4
#define MBOTS_HIDDEN_DIM 16
#define MBOTS_SENSOR_SIZE 32
#define MBOTS_CHUNKS_X 8
#define MBOTS_CHUNKS_Y 6
#define MBOTS_NUM_CHUNKS 48
#define MBOTS_FOOD_CAP 30
#define MBOTS_NUM_STATS 16

typedef struct mbots_config {
    int32_t gpu_id;                    /* Manager::Config::gpuID */
    uint32_t num_worlds;               /* Manager::Config::numWorlds (worlds owned by THIS handle) */
    uint32_t rand_seed;                /* Manager::Config::randSeed */
    uint32_t init_num_agents_per_world;/* Manager::Config::initNumAgentsPerWorld */
    /* Extensions (0 = default): */
    uint32_t max_agents_per_world;     /* per-world slot capacity; default max(128, 4*init) rounded to 32 */
    uint32_t world_id_offset;          /* global id of local world 0 (multi-GPU sharding); RNG streams
                                          are keyed on the GLOBAL id so results are shard-invariant */
} mbots_config;

/* Manager::Manager (src/entry/mgr.cpp:176-179 -> Impl::make :98-172) as bound by
 * SimManager.__init__ (src/entry/entry.cpp:17-32). */
int mbots_create(int32_t gpu_id, uint32_t num_worlds, uint32_t rand_seed,
                 uint32_t init_num_agents_per_world, mbots_sim **out);
int mbots_create_ex(const mbots_config *cfg, mbots_sim **out);
int mbots_destroy(mbots_sim *sim);

/* Manager::step (src/entry/mgr.cpp:189-192 -> Impl::step :51-63): step graph + sensor
 * graph, then blocks until the device is done (cuda_exec.cpp:2977-2981). */
int mbots_step(mbots_sim *sim, void *stream);
/* Same work, enqueued only (no host synchronisation). mbots_num_agents() and the
 * dims reported by mbots_tensor() are stale until mbots_sync(). */
int mbots_step_async(mbots_sim *sim, void *stream);
/* Blocks until the last enqueued step is complete (exports written, counts on the host). Work queued behind that
 * step on the stream -- e.g. an asynchronous shift -- is not waited for: step_async, shift_observations_async,
 * sync lets the shift run while the host prepares the next step. */
int mbots_sync(mbots_sim *sim, void *stream);

/* One rollout step as ONE launch (a cached CUDA graph; the reference launches one graph per step too,
 * M/src/mw/cuda_exec.cpp:2977-2981): [write the action rows] -> step + sensor -> [shift observations].
 *   actions_dev  device pointer to (action_rows, 6) int32 rows in observation-row order, or NULL to leave the
 *                action export as the caller wrote it through its view. min(live agents, action_rows) rows are
 *                written; the live count is read ON THE DEVICE, so the host may enqueue steps without knowing it.
 *                The matrix is read when the step RUNS, not when it is enqueued: keep it alive and unchanged until then
 *                (mbots_sync, or an event of the caller's recorded behind the call).
 *   shift        non-zero: Manager::shiftObservations of the new rows follows the step.
 * Nothing is waited for. Several steps may be in flight (the host runs ahead of the device); launch geometry is
 * then sized from the newest population that has reached the host plus slack, and worlds that outgrow it take the
 * slower general-capacity kernels -- never a wrong result. mbots_sync() waits for the last enqueued step.
 * mbots_step / mbots_step_async use the same graphs (without the action write and the shift). MBOTS_GRAPHS=0 in the
 * environment falls back to plain stream launches; so does mbots_profile(sim, 1) (events sit between the kernels). */
int mbots_rollout_step_async(mbots_sim *sim, const int32_t *actions_dev, uint64_t action_rows, int32_t shift, void *stream);
/* The action write alone (asynchronous, row count resolved on the device as above). */
int mbots_write_actions(mbots_sim *sim, const int32_t *actions_dev, uint64_t action_rows, void *stream);
/* Pointer lifetime -- a deliberate deviation from the reference, whose export base pointers never move
 * (M/src/mw/cuda_exec.cpp:3075; SURVEY.md 8(b) "fixed addresses"). Here every row-indexed export lives in one of TWO
 * row sets: a step gathers each agent's record from its old row in one set and writes its new row in the other (that
 * gather IS the species-major sort, and it cannot be done in place: births and deaths anywhere shift every later row).
 * Keeping the two caller-writable exports (action, hidden state) at one address would cost a copy-back of 88 B read +
 * 88 B written per agent and step (+0.07 ms at 2.2 M agents, ~10 % of the step); the choice made is to keep the
 * ping-pong and make its one hazard detectable: a view fetched before a step points at the DEAD set after it, and a
 * write through it is lost. The reference's own callers are unaffected -- they re-fetch every tensor after every step
 * because the leading dimension changes (learn/training_loop.py:47-50,136-137). Debug mode (this call, or
 * MBOTS_POISON_DEAD_SET=1 in the environment): after every step the dead set's action and hidden-state columns are
 * filled with a poison pattern, and the next step call first verifies the pattern: a write through a stale view makes
 * it fail with MBOTS_ESTATE and a message. Costs two small kernels and a host wait per step; off by default. */
int mbots_debug_stale_views(mbots_sim *sim, int32_t enable);

/* Number of step graphs launched by this handle since creation (bench bookkeeping). */
int mbots_graph_launch_count(mbots_sim *sim, uint64_t *out);

/* Manager::shiftObservations (src/entry/mgr.cpp:194-197). Blocking. */
int mbots_shift_observations(mbots_sim *sim, void *stream);
int mbots_shift_observations_async(mbots_sim *sim, void *stream);

/* SimBridge::totalNumAgents (src/sim/sim.hpp:73-77; written by bridgeSyncSystem,
 * src/sim/sim.cpp:985-999). */
int mbots_num_agents(mbots_sim *sim, uint32_t *out);

/* Manager::*Tensor (src/entry/mgr.cpp:199-422 -> Impl::exportTensor :70-76): device
 * pointer, dims and dtype of one export. `id` is an mbots_export_id. */
int mbots_tensor(mbots_sim *sim, int32_t id, void **dev_ptr, int64_t dims[2], int32_t *dtype);

/* First row of every species block of the current exports, and the total: out[k] .. out[k + 1] are the rows of species
 * k + 1 (rows are species-major, src/sim/sim.cpp:1148-1149). Exact -- the cumulative sum of species_count_tensor the
 * reference's loop uses (learn/training_loop.py:43-45) is taken before the step's respawns -- and free: the totals
 * arrive on the host with every step. MBOTS_ESTATE while asynchronous steps are pending. */
int mbots_species_blocks(mbots_sim *sim, uint32_t out[5]);

/* Rows every agent-indexed export is allocated for (num_worlds x per-world capacity): the buffers behind
 * mbots_tensor() stay valid, at fixed addresses, for this many rows, so a binding may build one view per
 * export and re-slice it to the live count after each step instead of re-wrapping the pointer. */
int mbots_row_capacity(mbots_sim *sim, uint64_t *out);

/* Manager::setAction (src/entry/mgr.cpp:251-272): write one row of the action export. */
int mbots_set_action(mbots_sim *sim, uint32_t agent_idx, int32_t forward, int32_t backward,
                     int32_t rotate_left, int32_t rotate_right, int32_t shoot, int32_t breed);

/* Manager::agentOffsetForWorld (src/entry/mgr.cpp:274-277). */
int mbots_agent_offset_for_world(mbots_sim *sim, uint32_t world_idx, uint32_t *out);

/* Host-buffer helpers for callers that keep their data on the host (bench e2e):
 * copy `bytes` between a host buffer and the start of export `id` on `stream`. */
int mbots_copy_from_host(mbots_sim *sim, int32_t id, const void *host, uint64_t bytes, void *stream);
int mbots_copy_to_host(mbots_sim *sim, int32_t id, void *host, uint64_t bytes, void *stream);

/* Cumulative episode statistics of this handle since creation, as 16 int64:
 *  [0] agent_steps (sum over steps of live agents)  [1] steps
 *  [2..5] per-species live count (last step)        [6..9] per-species health sum (last step)
 *  [10] births (cumulative) [11] deaths [12] respawns [13] overflow (dropped spawns)
 *  [14] food eaten [15] shots that hit
 * These are the scalars a multi-GPU driver all-reduces (NCCL, sum). */
int mbots_episode_stats(mbots_sim *sim, int64_t out[MBOTS_NUM_STATS], void *stream);

/* Number of kernels this handle has launched since creation (bench bookkeeping). */
int mbots_launch_count(mbots_sim *sim, uint64_t *out);

/* Average device time in milliseconds of each of the step's kernels over the calls since
 * the last reset (CUDA events on the launch stream; enabled by mbots_profile(sim, 1)).
 * out[0] = step kernel, out[1] = row scan, out[2] = export+sensor kernel, out[3] = shift. */
int mbots_profile(mbots_sim *sim, int32_t enable);
int mbots_profile_read(mbots_sim *sim, float out_ms[4], uint32_t out_calls[4]);

/* Simulator state checkpoint / restore. The reference has no such facility (its
 * learn/ckpt.py saves network weights only); this is SURVEY section 8(f) item 3: dump / reload every
 * persistent array of the handle (agent SoA, food, RNG keys + draw counters, observation rows, episode
 * counters) so that a rollout can be replayed bit-exactly from any step. The blob is host memory owned
 * by the caller; mbots_state_bytes gives its size for the CURRENT population. Restoring requires a
 * handle created with the same num_worlds, capacity and init_num_agents_per_world (MBOTS_EINVAL
 * otherwise). All three synchronise the stream; MBOTS_ESTATE while asynchronous steps are pending. */
int mbots_state_bytes(mbots_sim *sim, uint64_t *out);
int mbots_save_state(mbots_sim *sim, void *host, uint64_t bytes, void *stream);
int mbots_load_state(mbots_sim *sim, const void *host, uint64_t bytes, void *stream);

/* Policy-inference ends, for callers that run the species networks on the
 * exports. Rows [row_begin, row_end) of the CURRENT row set; both are asynchronous on `stream`.
 *
 * mbots_pack_observations: learn/util.py:23-29 construct_obs in one kernel -- writes the (n, 69) float32
 *   matrix [depth u8 x32 | health | position x2 | semantic i8 x32 | surrounding x2] of the rows (is_prev
 *   selects the prev_* exports) to the device buffer `out`.
 * mbots_sample_actions: learn/training_loop.py:57,77-83,136 -- Categorical sample + one-hot write-back.
 *   `weights` is a device (n, 6) float32 matrix of non-negative action weights (softmax of the logits; need
 *   not be normalised). Row r draws u = uniform(split(split(init(seed), draw), id, id >> 32)) with the
 *   simulator's Threefry (M/include/madrona/rand.inl), id = (GLOBAL world id << 10) | slot of the agent that owns the
 *   row, and takes the first k with u * sum(w) < w0 + .. + wk (sums in index order, fp32, no FMA), then stores the
 *   int32 one-hot row into the action export. The picked index also goes to out_index[r] when that is not NULL.
 *   Deterministic in (seed, draw, agent): an agent's draw does not depend on how the worlds are sharded over handles. */
int mbots_pack_observations(mbots_sim *sim, int32_t is_prev, uint32_t row_begin, uint32_t row_end, float *out,
                            void *stream);
int mbots_sample_actions(mbots_sim *sim, uint32_t row_begin, uint32_t row_end, const float *weights, uint32_t seed,
                         uint32_t draw, int32_t *out_index, void *stream);

/* The whole acting path of learn/training_loop.py:52-83,136 for rows [row_begin, row_end) as ONE kernel:
 * construct_obs -> species network (mbots_policy.h program over a device blob of bf16 weights) -> softmax ->
 * Categorical sample (same stream as mbots_sample_actions: keyed seed, draw, agent) -> one-hot row in the action
 * export. Tensor-core products (bf16 in, fp32 accumulate); activations never leave the SM between layers.
 * program->layout picks the kernel: MBOTS_LAYOUT_SW128 = tcgen05.mma + tensor memory (policy_umma.cu),
 * MBOTS_LAYOUT_ROWPAD = mma.sync (policy_fused.cu); the blob must be laid out accordingly (mbots_policy.h).
 * sample = 0 skips sampling and the action write (use out_logits). out_logits: device (n, 6) float32 or NULL;
 * out_index: device (n) int32 or NULL. Asynchronous on `stream`. MBOTS_EINVAL for a malformed program. */
int mbots_policy_act(mbots_sim *sim, uint32_t row_begin, uint32_t row_end, const void *blob, uint64_t blob_bytes,
                     const mbots_policy_program *program, uint32_t seed, uint32_t draw, int32_t sample,
                     float *out_logits, int32_t *out_index, void *stream);

/* The same kernel for the whole block of ONE species (0..3) of the current row set, with the row range resolved ON THE
 * DEVICE from the totals the last step's row scan wrote: the host needs neither mbots_species_blocks() nor a wait for the
 * step, so a rollout with the policy in the loop -- learn/training_loop.py:36-83,135-137: act, step, shift -- can be
 * enqueued several steps ahead (mbots_rollout_step_async with actions = NULL reads the rows this call wrote). Valid while
 * asynchronous steps are pending. MBOTS_LAYOUT_SW128 programs only. out_logits (rows, 6) / out_index (rows) are indexed by
 * OBSERVATION ROW here (size them for mbots_row_capacity or the population bound you keep), not by row within the block. */
int mbots_policy_act_species(mbots_sim *sim, int32_t species, const void *blob, uint64_t blob_bytes,
                             const mbots_policy_program *program, uint32_t seed, uint32_t draw, int32_t sample,
                             float *out_logits, int32_t *out_index, void *stream);

/* All four species in ONE launch of that kernel (network k acts on the block of species k + 1; row ranges resolved on the
 * device as above): the persistent CTAs walk the tiles of all four blocks in one pass, so the launch ramps up and drains once
 * instead of four times -- what matters when a GPU holds few worlds. Same draws and results as four
 * mbots_policy_act_species calls. MBOTS_EINVAL if the four networks' bias vectors together exceed the kernel's shared-memory
 * copy (6,144 floats): use the per-species call then. */
int mbots_policy_act_all(mbots_sim *sim, const void *const blobs[4], const uint64_t blob_bytes[4],
                         const mbots_policy_program *const programs[4], uint32_t seed, uint32_t draw, int32_t sample,
                         float *out_logits, int32_t *out_index, void *stream);

/* NO COUNTERPART IN THE REFERENCE -- added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * The step's row scan (what orders the observation rows species-major, src/sim/sim.cpp:1148-1149) as a stand-alone
 * call that ALLOCATES NOTHING: no cudaMalloc / cudaFree / allocating library routine, no handle. Its scratch memory
 * (one 32-byte record per tile of 1,024 worlds) is a file-scope __device__ array of the library, 8 KiB, held by every
 * GPU that loads the library until exit -- hence the limit num_worlds <= MBOTS_ROW_SCAN_STATIC_MAX_WORLDS; more is
 * rejected with MBOTS_EINVAL and a message (use a simulator handle for larger populations).
 *
 *   species_rows  in   device int32 (num_worlds, 4): rows of (world, species k + 1); 16-byte aligned
 *   row_base      out  device int32 (num_worlds, 4): first row of (world, species) when rows are ordered
 *                      (species, world): rows of earlier species in all worlds + rows of this species in earlier
 *                      worlds -- what MBOTS_DEBUG_ROW_BASE shows of a handle; 16-byte aligned
 *   world_offset  out  device int32 (num_worlds) or NULL: exclusive scan of the per-world sums
 *                      (Manager::agentWorldOffsets; MBOTS_DEBUG_WORLD_OFFSET of a handle)
 *   totals        out  device uint32 (6) or NULL: rows per species [0..3], all rows [4], largest world [5]
 *
 * Sums are int32: the total must stay below 2^31. All buffers must be device memory of ONE GPU (the call makes that
 * GPU current for its duration); `stream` must belong to it. Asynchronous on `stream`: two kernel launches.
 * Every call on a GPU shares the one scratch array, which keeps the last call's contents; nothing in it needs to
 * start at zero (each call overwrites what it reads), but calls that OVERLAP on the GPU (different streams, no
 * dependency) race on it -- serialise them. The first call on a GPU looks up the array's device address
 * (cudaGetSymbolAddress); make that call outside stream capture. */
#define MBOTS_ROW_SCAN_STATIC_MAX_WORLDS 262144u
int mbots_row_scan_static(const int32_t *species_rows, uint32_t num_worlds, int32_t *row_base, int32_t *world_offset,
                          uint32_t *totals, void *stream);

const char *mbots_last_error(void);
const char *mbots_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MBOTS_H */
