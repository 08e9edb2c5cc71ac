/* mbots_api.cu -- the C-ABI of include/mbots.h over the kernels of sim_kernels.cu.
 *
 * Host-side counterpart of the reference's Manager (src/entry/mgr.cpp): owns the device
 * memory of one simulator, launches the step, and hands out device pointers + dims for
 * the exports. No CPU fallback: every entry point either drives the GPU or fails.
 */
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include <map>
#include <tuple>
#include <new>

#include "../../include/mbots.h"
#include "../../include/mbots_policy.h"
#include "sim_dev.h"

using namespace mbk;

static thread_local std::string g_err;

static int fail(int code, const char *what, cudaError_t ce = cudaSuccess)
{
    char buf[512];
    if (ce != cudaSuccess) snprintf(buf, sizeof buf, "%s: %s", what, cudaGetErrorString(ce));
    else snprintf(buf, sizeof buf, "%s", what);
    g_err = buf;
    return code;
}

namespace mbk { bool shift_bulk_enabled(); void launch_shift_bulk(const SimDev &d, int cur, cudaStream_t s); }   /* shift_bulk.cu */

/* row_scan_static.cu reports through the same mbots_last_error() */
namespace mbk { int set_error(int code, const char *what, cudaError_t ce) { return fail(code, what, ce); } }

#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(MBOTS_ECUDA, #call, e_); } while (0)

enum { K_STEP = 0, K_SCAN = 1, K_EXPORT = 2, K_SHIFT = 3, K_COUNT = 4 };

/* One step's kernel chain as a CUDA graph (the reference launches one graph per step too,
 * M/src/mw/cuda_exec.cpp:2977-2981): [action write] -> step -> tile sums -> row scan -> export (fast, full) ->
 * 40-byte copy of the totals -> event -> [shift]. Everything a launch bakes in is part of the key. */
struct ChainKey {
    int cur, withShift;
    uint32_t stepSlots, bound, rowsUpper;
    const void *actions;
    unsigned long long actionRows;
    bool operator<(const ChainKey &o) const
    {
        return std::tie(cur, withShift, stepSlots, bound, rowsUpper, actions, actionRows) <
               std::tie(o.cur, o.withShift, o.stepSlots, o.bound, o.rowsUpper, o.actions, o.actionRows);
    }
};

struct mbots_sim {
    mbots_config cfg;
    SimDev d;
    int cur;                       /* obs set that holds the current exports */
    int device;
    std::vector<void *> allocs;
    Totals *hTotals;               /* pinned mirror of d.totals */
    long long *dStats, *hStats;    /* 16 x int64 */
    void *dbg;                     /* scratch for debug gathers: rows * 16 B */
    uint32_t totalAgents;          /* host copy, valid after a sync */
    bool dirty;                    /* async work enqueued since the last sync */
    uint64_t launches;
    /* per-kernel timing */
    bool profile;
    cudaEvent_t ev[5], evShift[2];
    cudaEvent_t evTotals;          /* recorded behind the step's last kernel and the copy of the totals */
    bool shiftPending;             /* evShift recorded, elapsed time not read yet */
    bool stepPending;              /* ev[0..3] recorded, elapsed times not read yet */
    double ms[K_COUNT];
    uint32_t calls[K_COUNT];
    /* step graphs */
    bool useGraphs;
    cudaStream_t capStream;        /* capture happens here: the caller's stream may be the legacy default stream */
    cudaStream_t sideStream[2];    /* forked branches of the captured chain */
    cudaEvent_t evFork, evJoin[2];
    struct StepGraph { cudaGraphExec_t exec; int kernels; };
    std::map<ChainKey, StepGraph> graphs;
    uint64_t graphLaunches;
    /* stale-view detector (MBOTS_POISON_DEAD_SET=1 or mbots_debug_stale_views) */
    bool poison, poisoned;
};

namespace {

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev)
    {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

template <typename T>
int dev_alloc(mbots_sim *s, T **out, size_t count)
{
    void *p = nullptr;
    size_t bytes = count * sizeof(T);
    if (bytes == 0) bytes = 16;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) return fail(e == cudaErrorMemoryAllocation ? MBOTS_ENOMEM : MBOTS_ECUDA, "cudaMalloc", e);
    e = cudaMemset(p, 0, bytes);
    if (e != cudaSuccess) return fail(MBOTS_ECUDA, "cudaMemset", e);
    s->allocs.push_back(p);
    *out = static_cast<T *>(p);
    return MBOTS_OK;
}

#define ALLOC(ptr, count) do { int rc_ = dev_alloc(s, &(ptr), (count)); if (rc_ != MBOTS_OK) return rc_; } while (0)

int alloc_obs(mbots_sim *s, ObsSet &o, size_t rows)
{
    ALLOC(o.action, rows * 6); ALLOC(o.prev_action, rows * 6);
    ALLOC(o.hidden, rows * 16); ALLOC(o.prev_hidden, rows * 16);
    ALLOC(o.reward, rows); ALLOC(o.prev_reward, rows);
    ALLOC(o.position, rows * 2); ALLOC(o.prev_position, rows * 2);
    ALLOC(o.health, rows); ALLOC(o.prev_health, rows);
    ALLOC(o.surrounding, rows * 2); ALLOC(o.prev_surrounding, rows * 2);
    ALLOC(o.stats, rows * 4); ALLOC(o.prev_stats, rows * 4);
    ALLOC(o.semantic, rows * 32); ALLOC(o.prev_semantic, rows * 32);
    ALLOC(o.depth, rows * 32); ALLOC(o.prev_depth, rows * 32);
    ALLOC(o.owner, rows);
    o.worldIdOffset = s->d.worldIdOffset;
    return MBOTS_OK;
}

int alloc_all(mbots_sim *s)
{
    SimDev &d = s->d;
    const size_t W = d.numWorlds, rows = W * d.cap;
    ALLOC(d.pose, rows); ALLOC(d.health, rows); ALLOC(d.species, rows); ALLOC(d.target, rows);
    ALLOC(d.scratch, rows); ALLOC(d.row, rows);
    ALLOC(d.worldCount, W); ALLOC(d.worldOffset, W); ALLOC(d.rngKey, W); ALLOC(d.rngCount, W);
    ALLOC(d.numFood, W); ALLOC(d.food, W * 48);
    ALLOC(d.speciesLive, W); ALLOC(d.rowBase[0], W); ALLOC(d.rowBase[1], W);
    const size_t tiles = (W + kScanTile - 1) / kScanTile;
    ALLOC(d.tileSum[0], tiles); ALLOC(d.tileSum[1], tiles);
    ALLOC(d.speciesCount, W); ALLOC(d.speciesReward, W); ALLOC(d.speciesHealth, W); ALLOC(d.wstats, W);
    ALLOC(d.sensorIndex, rows); ALLOC(d.done, rows); ALLOC(d.reset, W);
    ALLOC(d.totals, 1);
    ALLOC(d.scanSync, 4);
    int rc = alloc_obs(s, d.obs[0], rows);
    if (rc != MBOTS_OK) return rc;
    rc = alloc_obs(s, d.obs[1], rows);
    if (rc != MBOTS_OK) return rc;
    ALLOC(s->dStats, 16);
    {
        float4 *tmp = nullptr;
        ALLOC(tmp, rows);
        s->dbg = tmp;
    }
    return MBOTS_OK;
}

void free_all(mbots_sim *s)
{
    for (void *p : s->allocs) cudaFree(p);
    s->allocs.clear();
    if (s->hTotals) cudaFreeHost(s->hTotals);
    if (s->hStats) cudaFreeHost(s->hStats);
    for (auto &e : s->ev) if (e) cudaEventDestroy(e);
    for (auto &e : s->evShift) if (e) cudaEventDestroy(e);
    if (s->evTotals) cudaEventDestroy(s->evTotals);
    for (auto &kv : s->graphs) cudaGraphExecDestroy(kv.second.exec);
    s->graphs.clear();
    if (s->capStream) cudaStreamDestroy(s->capStream);
    for (auto &q : s->sideStream) if (q) cudaStreamDestroy(q);
    if (s->evFork) cudaEventDestroy(s->evFork);
    for (auto &e : s->evJoin) if (e) cudaEventDestroy(e);
}

/* Kernel timing events are read lazily so that profiling never turns an asynchronous call into a blocking one:
 * `wait` = false reads them only if they have completed (the read is retried later), true blocks for them. */
void resolve_shift_timing(mbots_sim *s, bool wait = true)
{
    if (!s->shiftPending) return;
    if (!wait && cudaEventQuery(s->evShift[1]) != cudaSuccess) return;
    s->shiftPending = false;
    float a = 0;
    if (cudaEventSynchronize(s->evShift[1]) == cudaSuccess &&
        cudaEventElapsedTime(&a, s->evShift[0], s->evShift[1]) == cudaSuccess) {
        s->ms[K_SHIFT] += a; s->calls[K_SHIFT]++;
    }
}

void resolve_step_timing(mbots_sim *s)
{
    if (!s->stepPending) return;
    s->stepPending = false;
    float a = 0, b = 0, c = 0;
    if (cudaEventSynchronize(s->ev[3]) != cudaSuccess) return;
    cudaEventElapsedTime(&a, s->ev[0], s->ev[1]);
    cudaEventElapsedTime(&b, s->ev[1], s->ev[2]);
    cudaEventElapsedTime(&c, s->ev[2], s->ev[3]);
    s->ms[K_STEP] += a; s->ms[K_SCAN] += b; s->ms[K_EXPORT] += c;
    s->calls[K_STEP]++; s->calls[K_SCAN]++; s->calls[K_EXPORT]++;
}

/* Wait for the last enqueued step (its exports and the host copy of the totals), not for work queued behind it:
 * a shift enqueued before the sync keeps running while the host prepares the next step. */
int finish(mbots_sim *s, cudaStream_t st)
{
    (void)st;
    CU(cudaEventSynchronize(s->evTotals));
    resolve_step_timing(s);                 /* complete by now */
    resolve_shift_timing(s, false);         /* a shift queued behind the step may still be running */
    s->totalAgents = s->hTotals->totalAgents;
    s->dirty = false;
    return MBOTS_OK;
}

int enqueue_repack(mbots_sim *s, cudaStream_t st, int with_sensors, int count_step, uint32_t bound)
{
    const int nscan = launch_scan(s->d, s->cur, count_step, st);
    if (s->profile && with_sensors) cudaEventRecord(s->ev[2], st);
    const int nexp = launch_export(s->d, s->cur, with_sensors, bound, nullptr, 0, 0, st);
    if (s->profile && with_sensors) cudaEventRecord(s->ev[3], st);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(s->hTotals, s->d.totals, sizeof(Totals), cudaMemcpyDeviceToHost, st));
    CU(cudaEventRecord(s->evTotals, st));
    s->cur ^= 1;
    s->launches += (uint64_t)(nscan + nexp);            /* row scan (one or two kernels), export variant(s) */
    s->dirty = true;
    return MBOTS_OK;
}

uint32_t pow2_at_least(uint32_t v)
{
    uint32_t p = 256;
    while (p < v && p < 0x80000000u) p <<= 1;
    return p;
}

/* Launch geometry of the next step. With the host in step with the device (no step enqueued since the last sync)
 * it is sized from the exact totals; while steps are in flight, from the newest totals that have landed in the
 * pinned mirror plus slack. The bounds only size shared memory, grids and the fast / full split of the export:
 * a world that outgrows them is still handled (step: capacity-sized when unsure; export: the full variant). */
ChainKey chain_key(mbots_sim *s, const void *actions, unsigned long long action_rows, int with_shift)
{
    ChainKey k;
    uint32_t need = s->hTotals->maxWorldCount;
    if (need < s->d.initAgents) need = s->d.initAgents;
    if (need < 32u) need = 32u;
    uint32_t rows = s->hTotals->totalAgents;
    if (s->dirty) {
        need += 3u;                                     /* the largest world moves by a few agents per step; keep 32 CTAs per SM */
        rows += rows / 4u + 1024u;
    }
    const uint64_t cap_rows = (uint64_t)s->d.numWorlds * s->d.cap;
    if (rows > cap_rows) rows = (uint32_t)cap_rows;
    k.cur = s->cur;
    k.withShift = with_shift;
    k.stepSlots = s->dirty ? s->d.cap : ((need + 31u) & ~31u);
    if (k.stepSlots > s->d.cap) k.stepSlots = s->d.cap;
    k.bound = ((need + 8u) + 7u) & ~7u;                 /* launch_export rounds to 8 anyway */
    k.rowsUpper = pow2_at_least(rows);
    k.actions = actions;
    k.actionRows = actions ? action_rows : 0ull;
    return k;
}

/* the chain itself, on `st` (a real stream, or the capture stream). Returns the kernels launched. */
int record_chain(mbots_sim *s, const ChainKey &k, cudaStream_t st, bool timing, bool external_event, int *launches)
{
    /* Fused rollout step: the step reads its actions straight from the caller's matrix and the export's carry writes them
     * into the new rows (no action-write pass); with a shift requested the export leaves prev := cur (no shift pass). */
    int n = 0;
    if (timing) cudaEventRecord(s->ev[0], st);
    launch_step(s->d, k.cur, k.stepSlots, k.actions, k.actionRows, st);
    if (timing) cudaEventRecord(s->ev[1], st);
    n += 1 + launch_scan(s->d, k.cur, 1, st);
    if (timing) cudaEventRecord(s->ev[2], st);
    if (external_event) {
        /* under capture: the totals are final once the scan has run, and the clean-up export touches other worlds than the
         * fast one -- both fork off onto side streams and join before the step's completion event:
         *     step -> scan -> { export (fast) | export (full capacity) | 40-byte copy of the totals } -> event */
        CU(cudaEventRecord(s->evFork, st));
        CU(cudaStreamWaitEvent(s->sideStream[0], s->evFork, 0));
        CU(cudaStreamWaitEvent(s->sideStream[1], s->evFork, 0));
        CU(cudaMemcpyAsync(s->hTotals, s->d.totals, sizeof(Totals), cudaMemcpyDeviceToHost, s->sideStream[1]));
        n += launch_export(s->d, k.cur, 1, k.bound, k.actions, k.actionRows, k.withShift, st, s->sideStream[0]);
        CU(cudaGetLastError());
        CU(cudaEventRecord(s->evJoin[0], s->sideStream[0]));
        CU(cudaEventRecord(s->evJoin[1], s->sideStream[1]));
        CU(cudaStreamWaitEvent(st, s->evJoin[0], 0));
        CU(cudaStreamWaitEvent(st, s->evJoin[1], 0));
        CU(cudaEventRecordWithFlags(s->evTotals, st, cudaEventRecordExternal));
    } else {
        n += launch_export(s->d, k.cur, 1, k.bound, k.actions, k.actionRows, k.withShift, st);
        if (timing) cudaEventRecord(s->ev[3], st);
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(s->hTotals, s->d.totals, sizeof(Totals), cudaMemcpyDeviceToHost, st));
        CU(cudaEventRecord(s->evTotals, st));
    }
    *launches = n;
    return MBOTS_OK;
}



/* Debug mode: did the caller write the action / hidden-state export through a view it fetched BEFORE the last step?
 * Those views point at the dead row set, whose two writable columns were poisoned after that step. */
int check_stale_views(mbots_sim *s, cudaStream_t st)
{
    if (!s->poison || !s->poisoned) return MBOTS_OK;
    unsigned int *bad = s->d.scanSync + 2;
    launch_poison(s->d, s->cur ^ 1, (size_t)s->d.numWorlds * s->d.cap, 1, bad, st);
    unsigned int h = 0;
    CU(cudaMemcpyAsync(&h, bad, sizeof h, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    s->poisoned = false;
    if (h != 0) {
        char msg[200];
        snprintf(msg, sizeof msg, "%u words of the action / hidden-state export were written through a view fetched before "
                 "the previous step (row exports are re-fetched after every step: include/mbots.h)", h);
        return fail(MBOTS_ESTATE, msg);
    }
    return MBOTS_OK;
}

int enqueue_step(mbots_sim *s, cudaStream_t st, const void *actions = nullptr, unsigned long long action_rows = 0,
                 int with_shift = 0)
{
    int rc0 = check_stale_views(s, st);
    if (rc0 != MBOTS_OK) return rc0;
    const ChainKey k = chain_key(s, actions, action_rows, with_shift);
    int n = 0;
    if (s->profile || !s->useGraphs) {
        /* per-kernel timing wants events between the kernels: plain stream launches */
        if (s->profile) resolve_step_timing(s);
        int rc = record_chain(s, k, st, s->profile, false, &n);
        if (rc != MBOTS_OK) return rc;
        if (s->profile) s->stepPending = true;
    } else {
        auto it = s->graphs.find(k);
        if (it == s->graphs.end()) {
            if (s->graphs.size() >= 256) {              /* geometry keeps changing: start over rather than grow */
                for (auto &kv : s->graphs) cudaGraphExecDestroy(kv.second.exec);
                s->graphs.clear();
            }
            cudaGraph_t g = nullptr;
            CU(cudaStreamBeginCapture(s->capStream, cudaStreamCaptureModeThreadLocal));
            int rc = record_chain(s, k, s->capStream, false, true, &n);
            cudaError_t ce = cudaStreamEndCapture(s->capStream, &g);
            if (rc != MBOTS_OK) { if (g) cudaGraphDestroy(g); return rc; }
            if (ce != cudaSuccess) return fail(MBOTS_ECUDA, "cudaStreamEndCapture", ce);
            cudaGraphExec_t ex = nullptr;
            ce = cudaGraphInstantiate(&ex, g, 0);
            cudaGraphDestroy(g);
            if (ce != cudaSuccess) return fail(MBOTS_ECUDA, "cudaGraphInstantiate", ce);
            it = s->graphs.emplace(k, mbots_sim::StepGraph{ex, n}).first;
        }
        CU(cudaGraphLaunch(it->second.exec, st));
        n = it->second.kernels;
        s->graphLaunches += 1;
    }
    s->cur ^= 1;
    s->launches += (uint64_t)n;
    s->dirty = true;
    if (s->poison) {
        launch_poison(s->d, s->cur ^ 1, (size_t)s->d.numWorlds * s->d.cap, 0, nullptr, st);
        s->poisoned = true;
    }
    return MBOTS_OK;
}

} // namespace

extern "C" {

int mbots_create_ex(const mbots_config *cfg, mbots_sim **out)
{
    if (!cfg || !out) return fail(MBOTS_EINVAL, "null argument");
    *out = nullptr;
    if (cfg->num_worlds == 0) return fail(MBOTS_EINVAL, "num_worlds must be > 0");
    uint32_t cap = cfg->max_agents_per_world;
    if (cap == 0) {
        cap = 4u * cfg->init_num_agents_per_world;
        if (cap < 128u) cap = 128u;
    }
    cap = (cap + 31u) & ~31u;
    if (cfg->num_worlds > (1u << 22)) return fail(MBOTS_EINVAL, "num_worlds must be <= 4,194,304 per handle");
    if (cap < cfg->init_num_agents_per_world || cap > 1024u)
        return fail(MBOTS_EINVAL, "per-world capacity must hold the initial agents and be <= 1024");
    if ((uint64_t)cfg->num_worlds * cap > 0x1fffffffull) return fail(MBOTS_EINVAL, "num_worlds * capacity exceeds 2^29 rows");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return fail(MBOTS_ECUDA, "no CUDA device (this library has no CPU path)", e);
    if (cfg->gpu_id < 0 || cfg->gpu_id >= ndev) return fail(MBOTS_EINVAL, "gpu_id out of range");

    mbots_sim *s = new (std::nothrow) mbots_sim();
    if (!s) return fail(MBOTS_ENOMEM, "host allocation failed");
    s->cfg = *cfg;
    s->cfg.max_agents_per_world = cap;
    s->device = cfg->gpu_id;
    memset(&s->d, 0, sizeof(s->d));
    s->d.numWorlds = cfg->num_worlds; s->d.cap = cap; s->d.initAgents = cfg->init_num_agents_per_world;
    s->d.worldIdOffset = cfg->world_id_offset; s->d.seed = cfg->rand_seed;
    DeviceGuard g(s->device);
    int rc = g.ok ? MBOTS_OK : fail(MBOTS_ECUDA, "cudaSetDevice");
    if (rc == MBOTS_OK) rc = alloc_all(s);
    if (rc == MBOTS_OK) {
        cudaError_t e1 = cudaMallocHost((void **)&s->hTotals, sizeof(Totals));
        cudaError_t e2 = cudaMallocHost((void **)&s->hStats, 16 * sizeof(long long));
        if (e1 != cudaSuccess || e2 != cudaSuccess) rc = fail(MBOTS_ENOMEM, "cudaMallocHost");
        else memset(s->hTotals, 0, sizeof(Totals));
    }
    if (rc == MBOTS_OK) {
        for (auto &ev : s->ev) if (cudaEventCreate(&ev) != cudaSuccess) rc = fail(MBOTS_ECUDA, "cudaEventCreate");
        for (auto &ev : s->evShift) if (cudaEventCreate(&ev) != cudaSuccess) rc = fail(MBOTS_ECUDA, "cudaEventCreate");
        if (cudaEventCreateWithFlags(&s->evTotals, cudaEventDisableTiming) != cudaSuccess) rc = fail(MBOTS_ECUDA, "cudaEventCreate");
        if (cudaStreamCreateWithFlags(&s->capStream, cudaStreamNonBlocking) != cudaSuccess) rc = fail(MBOTS_ECUDA, "cudaStreamCreate");
        for (auto &q : s->sideStream)
            if (cudaStreamCreateWithFlags(&q, cudaStreamNonBlocking) != cudaSuccess) rc = fail(MBOTS_ECUDA, "cudaStreamCreate");
        if (cudaEventCreateWithFlags(&s->evFork, cudaEventDisableTiming) != cudaSuccess) rc = fail(MBOTS_ECUDA, "cudaEventCreate");
        for (auto &e : s->evJoin)
            if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) rc = fail(MBOTS_ECUDA, "cudaEventCreate");
        const char *g = getenv("MBOTS_GRAPHS");
        s->useGraphs = !(g && g[0] == '0');
        const char *pz = getenv("MBOTS_POISON_DEAD_SET");
        s->poison = pz && pz[0] == '1';
    }
    if (rc == MBOTS_OK) {
        prepare_kernels(s->d);                  /* shared-memory opt-in for the largest geometry, once, outside any capture */
        /* Sim::Sim for every world, then the first row assignment (no sensor pass: the
         * reference's tensors are only meaningful after the first step()). */
        launch_init(s->d, 0);
        s->cur = 0;
        s->launches = 1;
        rc = enqueue_repack(s, 0, 0, 0, 0);
        if (rc == MBOTS_OK) rc = finish(s, 0);
    }
    if (rc != MBOTS_OK) { free_all(s); delete s; return rc; }
    *out = s;
    return MBOTS_OK;
}

int mbots_create(int32_t gpu_id, uint32_t num_worlds, uint32_t rand_seed,
                 uint32_t init_num_agents_per_world, mbots_sim **out)
{
    mbots_config c;
    memset(&c, 0, sizeof c);
    c.gpu_id = gpu_id; c.num_worlds = num_worlds; c.rand_seed = rand_seed;
    c.init_num_agents_per_world = init_num_agents_per_world;
    return mbots_create_ex(&c, out);
}

int mbots_destroy(mbots_sim *s)
{
    if (!s) return fail(MBOTS_EINVAL, "null handle");
    DeviceGuard g(s->device);
    cudaDeviceSynchronize();
    free_all(s);
    delete s;
    return MBOTS_OK;
}

int mbots_step_async(mbots_sim *s, void *stream)
{
    if (!s) return fail(MBOTS_EINVAL, "null handle");
    DeviceGuard g(s->device);
    return enqueue_step(s, (cudaStream_t)stream);
}

int mbots_rollout_step_async(mbots_sim *s, const int32_t *actions_dev, uint64_t action_rows, int32_t shift, void *stream)
{
    if (!s) return fail(MBOTS_EINVAL, "null handle");
    DeviceGuard g(s->device);
    return enqueue_step(s, (cudaStream_t)stream, actions_dev, action_rows, shift != 0);
}

int mbots_write_actions(mbots_sim *s, const int32_t *actions_dev, uint64_t action_rows, void *stream)
{
    if (!s || !actions_dev) return fail(MBOTS_EINVAL, "null argument");
    DeviceGuard g(s->device);
    uint32_t rows = s->hTotals->totalAgents;
    if (s->dirty) rows += rows / 4u + 1024u;
    launch_action_write(s->d, s->cur, actions_dev, action_rows, rows, (cudaStream_t)stream);
    CU(cudaGetLastError());
    s->launches += 1;
    return MBOTS_OK;
}

int mbots_debug_stale_views(mbots_sim *s, int32_t enable)
{
    if (!s) return fail(MBOTS_EINVAL, "null handle");
    s->poison = enable != 0;
    s->poisoned = false;
    return MBOTS_OK;
}

int mbots_graph_launch_count(mbots_sim *s, uint64_t *out)
{
    if (!s || !out) return fail(MBOTS_EINVAL, "null argument");
    *out = s->graphLaunches;
    return MBOTS_OK;
}

int mbots_sync(mbots_sim *s, void *stream)
{
    if (!s) return fail(MBOTS_EINVAL, "null handle");
    DeviceGuard g(s->device);
    return finish(s, (cudaStream_t)stream);
}

int mbots_step(mbots_sim *s, void *stream)
{
    if (!s) return fail(MBOTS_EINVAL, "null handle");
    DeviceGuard g(s->device);
    int rc = enqueue_step(s, (cudaStream_t)stream);
    if (rc != MBOTS_OK) return rc;
    return finish(s, (cudaStream_t)stream);
}

int mbots_shift_observations_async(mbots_sim *s, void *stream)
{
    if (!s) return fail(MBOTS_EINVAL, "null handle");
    DeviceGuard g(s->device);
    cudaStream_t st = (cudaStream_t)stream;
    if (s->profile) { resolve_shift_timing(s); cudaEventRecord(s->evShift[0], st); }
    /* grid sized from the capacity bound when the host count is stale */
    if (mbk::shift_bulk_enabled()) mbk::launch_shift_bulk(s->d, s->cur, st);
    else launch_shift(s->d, s->cur, s->dirty ? s->d.numWorlds * s->d.cap : s->totalAgents, st);
    if (s->profile) { cudaEventRecord(s->evShift[1], st); s->shiftPending = true; }
    CU(cudaGetLastError());
    s->launches += 1;
    return MBOTS_OK;
}

int mbots_shift_observations(mbots_sim *s, void *stream)
{
    int rc = mbots_shift_observations_async(s, stream);
    if (rc != MBOTS_OK) return rc;
    DeviceGuard g(s->device);
    CU(cudaStreamSynchronize((cudaStream_t)stream));
    return MBOTS_OK;
}

int mbots_num_agents(mbots_sim *s, uint32_t *out)
{
    if (!s || !out) return fail(MBOTS_EINVAL, "null argument");
    *out = s->totalAgents;
    return MBOTS_OK;
}

int mbots_species_blocks(mbots_sim *s, uint32_t out[5])
{
    if (!s || !out) return fail(MBOTS_EINVAL, "null argument");
    if (s->dirty) return fail(MBOTS_ESTATE, "asynchronous work pending: call mbots_sync first");
    uint32_t a = 0;
    for (int k = 0; k < 4; ++k) { out[k] = a; a += s->hTotals->speciesTotal[k]; }
    out[4] = a;
    return MBOTS_OK;
}

int mbots_row_capacity(mbots_sim *s, uint64_t *out)
{
    if (!s || !out) return fail(MBOTS_EINVAL, "null argument");
    *out = (uint64_t)s->d.numWorlds * s->d.cap;
    return MBOTS_OK;
}

int mbots_tensor(mbots_sim *s, int32_t id, void **dev_ptr, int64_t dims[2], int32_t *dtype)
{
    if (!s || !dev_ptr || !dims || !dtype) return fail(MBOTS_EINVAL, "null argument");
    const ObsSet &o = s->d.obs[s->cur];
    const int64_t A = s->totalAgents, W = s->d.numWorlds;
    auto set = [&](void *p, int64_t d0, int64_t d1, int32_t dt) {
        *dev_ptr = p; dims[0] = d0; dims[1] = d1; *dtype = dt; return (int)MBOTS_OK;
    };
    auto gather = [&](int what, int64_t cols, int32_t dt) {
        DeviceGuard g(s->device);
        launch_debug_gather(s->d, what, s->dbg, 0);
        s->launches += 1;
        cudaError_t e = cudaStreamSynchronize(0);
        if (e != cudaSuccess) return fail(MBOTS_ECUDA, "debug gather", e);
        return set(s->dbg, A, cols, dt);
    };
    switch (id) {
    case MBOTS_EXPORT_RESET: return set(s->d.reset, W, 1, MBOTS_I32);
    case MBOTS_EXPORT_ACTION: return set(o.action, A, 6, MBOTS_I32);
    case MBOTS_EXPORT_PREV_ACTION: return set(o.prev_action, A, 6, MBOTS_I32);
    case MBOTS_EXPORT_HIDDEN_STATE: return set(o.hidden, A, 16, MBOTS_F32);
    case MBOTS_EXPORT_PREV_HIDDEN_STATE: return set(o.prev_hidden, A, 16, MBOTS_F32);
    case MBOTS_EXPORT_REWARD: return set(o.reward, A, 1, MBOTS_F32);
    case MBOTS_EXPORT_PREV_REWARD: return set(o.prev_reward, A, 1, MBOTS_F32);
    case MBOTS_EXPORT_DONE: return set(s->d.done, A, 1, MBOTS_I32);
    case MBOTS_EXPORT_POSITION: return set(o.position, A, 2, MBOTS_F32);
    case MBOTS_EXPORT_PREV_POSITION: return set(o.prev_position, A, 2, MBOTS_F32);
    case MBOTS_EXPORT_HEALTH: return set(o.health, A, 1, MBOTS_F32);
    case MBOTS_EXPORT_PREV_HEALTH: return set(o.prev_health, A, 1, MBOTS_F32);
    case MBOTS_EXPORT_SURROUNDING: return set(o.surrounding, A, 2, MBOTS_F32);
    case MBOTS_EXPORT_PREV_SURROUNDING: return set(o.prev_surrounding, A, 2, MBOTS_F32);
    case MBOTS_EXPORT_SENSOR_SEMANTIC: return set(o.semantic, A, 32, MBOTS_I8);
    case MBOTS_EXPORT_SENSOR_DEPTH: return set(o.depth, A, 32, MBOTS_U8);
    case MBOTS_EXPORT_PREV_SENSOR_SEMANTIC: return set(o.prev_semantic, A, 32, MBOTS_I8);
    case MBOTS_EXPORT_PREV_SENSOR_DEPTH: return set(o.prev_depth, A, 32, MBOTS_U8);
    case MBOTS_EXPORT_STATS: return set(o.stats, A, 4, MBOTS_I32);
    case MBOTS_EXPORT_PREV_STATS: return set(o.prev_stats, A, 4, MBOTS_I32);
    case MBOTS_EXPORT_SENSOR_INDEX: return set(s->d.sensorIndex, A, 1, MBOTS_I32);
    case MBOTS_EXPORT_SPECIES_COUNT: return set(s->d.speciesCount, W, 4, MBOTS_I32);
    case MBOTS_DEBUG_WORLD_COUNT: return set(s->d.worldCount, W, 1, MBOTS_I32);
    case MBOTS_DEBUG_WORLD_OFFSET: return set(s->d.worldOffset, W, 1, MBOTS_I32);
    case MBOTS_DEBUG_RNG_COUNT: return set(s->d.rngCount, W, 1, MBOTS_I32);
    case MBOTS_DEBUG_NUM_FOOD: return set(s->d.numFood, W, 1, MBOTS_I32);
    case MBOTS_DEBUG_FOOD: return set(s->d.food, W * 48, 8, MBOTS_U8);
    case MBOTS_DEBUG_ROW_BASE: return set(s->d.rowBase[s->cur], W, 4, MBOTS_I32);
    case MBOTS_DEBUG_ROW_OWNER: return set(o.owner, A, 1, MBOTS_I32);
    case MBOTS_DEBUG_POSE: return gather(0, 4, MBOTS_F32);
    case MBOTS_DEBUG_AGENT_HEALTH: return gather(1, 1, MBOTS_I32);
    case MBOTS_DEBUG_AGENT_SPECIES: return gather(2, 1, MBOTS_I32);
    case MBOTS_DEBUG_AGENT_TARGET: return gather(3, 1, MBOTS_I32);
    default: return fail(MBOTS_EINVAL, "unknown export id");
    }
}

int mbots_set_action(mbots_sim *s, uint32_t agent_idx, int32_t forward, int32_t backward,
                     int32_t rotate_left, int32_t rotate_right, int32_t shoot, int32_t breed)
{
    if (!s) return fail(MBOTS_EINVAL, "null handle");
    if (agent_idx >= s->totalAgents) return fail(MBOTS_EINVAL, "agent index out of range");
    DeviceGuard g(s->device);
    const int32_t a[6] = {forward, backward, rotate_left, rotate_right, shoot, breed};
    CU(cudaMemcpy(s->d.obs[s->cur].action + (size_t)agent_idx * 6, a, sizeof a, cudaMemcpyHostToDevice));
    return MBOTS_OK;
}

int mbots_agent_offset_for_world(mbots_sim *s, uint32_t world_idx, uint32_t *out)
{
    if (!s || !out) return fail(MBOTS_EINVAL, "null argument");
    if (world_idx >= s->d.numWorlds) return fail(MBOTS_EINVAL, "world index out of range");
    DeviceGuard g(s->device);
    int32_t v = 0;
    CU(cudaMemcpy(&v, s->d.worldOffset + world_idx, sizeof v, cudaMemcpyDeviceToHost));
    *out = (uint32_t)v;
    return MBOTS_OK;
}

static int export_bytes(mbots_sim *s, int32_t id, void **p, uint64_t *bytes)
{
    int64_t dims[2];
    int32_t dt;
    int rc = mbots_tensor(s, id, p, dims, &dt);
    if (rc != MBOTS_OK) return rc;
    const uint64_t es = (dt == MBOTS_I32 || dt == MBOTS_F32) ? 4u : 1u;
    *bytes = (uint64_t)dims[0] * (uint64_t)dims[1] * es;
    return MBOTS_OK;
}

int mbots_copy_from_host(mbots_sim *s, int32_t id, const void *host, uint64_t bytes, void *stream)
{
    if (!s || !host) return fail(MBOTS_EINVAL, "null argument");
    void *p; uint64_t have;
    int rc = export_bytes(s, id, &p, &have);
    if (rc != MBOTS_OK) return rc;
    if (bytes > have) return fail(MBOTS_EINVAL, "copy larger than the export");
    DeviceGuard g(s->device);
    CU(cudaMemcpyAsync(p, host, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return MBOTS_OK;
}

int mbots_copy_to_host(mbots_sim *s, int32_t id, void *host, uint64_t bytes, void *stream)
{
    if (!s || !host) return fail(MBOTS_EINVAL, "null argument");
    void *p; uint64_t have;
    int rc = export_bytes(s, id, &p, &have);
    if (rc != MBOTS_OK) return rc;
    if (bytes > have) return fail(MBOTS_EINVAL, "copy larger than the export");
    DeviceGuard g(s->device);
    CU(cudaMemcpyAsync(host, p, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    return MBOTS_OK;
}

int mbots_episode_stats(mbots_sim *s, int64_t out[MBOTS_NUM_STATS], void *stream)
{
    if (!s || !out) return fail(MBOTS_EINVAL, "null argument");
    DeviceGuard g(s->device);
    cudaStream_t st = (cudaStream_t)stream;
    launch_stats_reduce(s->d, s->dStats, st);
    s->launches += 1;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(s->hStats, s->dStats, 16 * sizeof(long long), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    for (int i = 0; i < MBOTS_NUM_STATS; ++i) out[i] = (int64_t)s->hStats[i];
    return MBOTS_OK;
}

int mbots_launch_count(mbots_sim *s, uint64_t *out)
{
    if (!s || !out) return fail(MBOTS_EINVAL, "null argument");
    *out = s->launches;
    return MBOTS_OK;
}

int mbots_profile(mbots_sim *s, int32_t enable)
{
    if (!s) return fail(MBOTS_EINVAL, "null handle");
    resolve_step_timing(s);
    resolve_shift_timing(s);
    s->profile = enable != 0;
    for (int i = 0; i < K_COUNT; ++i) { s->ms[i] = 0; s->calls[i] = 0; }
    return MBOTS_OK;
}

int mbots_profile_read(mbots_sim *s, float out_ms[4], uint32_t out_calls[4])
{
    if (!s || !out_ms || !out_calls) return fail(MBOTS_EINVAL, "null argument");
    resolve_step_timing(s);
    resolve_shift_timing(s);
    for (int i = 0; i < K_COUNT; ++i) {
        out_calls[i] = s->calls[i];
        out_ms[i] = s->calls[i] ? (float)(s->ms[i] / s->calls[i]) : 0.f;
    }
    return MBOTS_OK;
}

const char *mbots_last_error(void) { return g_err.c_str(); }
const char *mbots_version(void) { return "mbots-b200 0.1 (sm_100a)"; }

} /* extern "C" */

/* ---- simulator state checkpoint / restore (the reference has none) ------------
 * A checkpoint is a self-describing byte blob: header, then every persistent array of the handle. Agent
 * state travels whole (numWorlds x capacity slots), observation rows are cut at the live row count, and
 * the blob is always written as if obs set 0 were current, so a restore lands in a fixed configuration. */
namespace {

struct CkptHeader {
    char magic[8];                 /* "MBOTSCK1" */
    uint32_t numWorlds, cap, initAgents, worldIdOffset, seed, totalAgents;
    uint64_t bytes;                /* whole blob */
    Totals totals;
};

struct Segment { void *ptr; size_t bytes; };

/* the persistent arrays, in blob order; `set` = obs set to read from / write to */
void ckpt_segments(mbots_sim *s, int set, size_t A, std::vector<Segment> &out)
{
    SimDev &d = s->d;
    const size_t W = d.numWorlds, slots = W * d.cap;
    auto add = [&](void *p, size_t b) { out.push_back(Segment{p, b}); };
    add(d.pose, slots * 16); add(d.health, slots * 4); add(d.species, slots); add(d.target, slots * 2);
    add(d.scratch, slots * 16); add(d.row, slots * 4);
    add(d.worldCount, W * 4); add(d.worldOffset, W * 4); add(d.rngKey, W * 8); add(d.rngCount, W * 4);
    add(d.numFood, W * 4); add(d.food, W * 48 * 8); add(d.speciesLive, W * 16); add(d.rowBase[set], W * 16);
    add(d.speciesCount, W * 16); add(d.speciesReward, W * 16); add(d.speciesHealth, W * 16);
    add(d.wstats, W * sizeof(WorldStats));
    add(d.sensorIndex, A * 4);
    ObsSet &o = d.obs[set];
    add(o.action, A * 24); add(o.prev_action, A * 24); add(o.hidden, A * 64); add(o.prev_hidden, A * 64);
    add(o.reward, A * 4); add(o.prev_reward, A * 4); add(o.position, A * 8); add(o.prev_position, A * 8);
    add(o.health, A * 4); add(o.prev_health, A * 4); add(o.surrounding, A * 8); add(o.prev_surrounding, A * 8);
    add(o.stats, A * 16); add(o.prev_stats, A * 16); add(o.semantic, A * 32); add(o.prev_semantic, A * 32);
    add(o.depth, A * 32); add(o.prev_depth, A * 32);
    add(o.owner, A * 4);
}

uint64_t ckpt_total(const std::vector<Segment> &segs)
{
    uint64_t b = sizeof(CkptHeader);
    for (const Segment &g : segs) b += (g.bytes + 15u) & ~(uint64_t)15u;
    return b;
}

} // namespace

extern "C" {

int mbots_state_bytes(mbots_sim *s, uint64_t *out)
{
    if (!s || !out) return fail(MBOTS_EINVAL, "null argument");
    if (s->dirty) return fail(MBOTS_ESTATE, "asynchronous work pending: call mbots_sync first");
    std::vector<Segment> segs;
    ckpt_segments(s, s->cur, s->totalAgents, segs);
    *out = ckpt_total(segs);
    return MBOTS_OK;
}

int mbots_save_state(mbots_sim *s, void *host, uint64_t bytes, void *stream)
{
    if (!s || !host) return fail(MBOTS_EINVAL, "null argument");
    if (s->dirty) return fail(MBOTS_ESTATE, "asynchronous work pending: call mbots_sync first");
    std::vector<Segment> segs;
    ckpt_segments(s, s->cur, s->totalAgents, segs);
    const uint64_t need = ckpt_total(segs);
    if (bytes < need) return fail(MBOTS_EINVAL, "buffer smaller than mbots_state_bytes()");
    DeviceGuard g(s->device);
    cudaStream_t st = (cudaStream_t)stream;
    CkptHeader h;
    memset(&h, 0, sizeof h);
    memcpy(h.magic, "MBOTSCK1", 8);
    h.numWorlds = s->d.numWorlds; h.cap = s->d.cap; h.initAgents = s->d.initAgents;
    h.worldIdOffset = s->d.worldIdOffset; h.seed = s->d.seed; h.totalAgents = s->totalAgents;
    h.bytes = need;
    CU(cudaMemcpyAsync(&h.totals, s->d.totals, sizeof(Totals), cudaMemcpyDeviceToHost, st));
    unsigned char *p = static_cast<unsigned char *>(host) + sizeof(CkptHeader);
    for (const Segment &sg : segs) {
        if (sg.bytes) CU(cudaMemcpyAsync(p, sg.ptr, sg.bytes, cudaMemcpyDeviceToHost, st));
        p += (sg.bytes + 15u) & ~(size_t)15u;
    }
    CU(cudaStreamSynchronize(st));
    memcpy(host, &h, sizeof h);
    return MBOTS_OK;
}

int mbots_load_state(mbots_sim *s, const void *host, uint64_t bytes, void *stream)
{
    if (!s || !host) return fail(MBOTS_EINVAL, "null argument");
    if (bytes < sizeof(CkptHeader)) return fail(MBOTS_EINVAL, "not a checkpoint (too short)");
    CkptHeader h;
    memcpy(&h, host, sizeof h);
    if (memcmp(h.magic, "MBOTSCK1", 8) != 0) return fail(MBOTS_EINVAL, "not a checkpoint (bad magic)");
    if (h.numWorlds != s->d.numWorlds || h.cap != s->d.cap || h.initAgents != s->d.initAgents)
        return fail(MBOTS_EINVAL, "checkpoint was taken with a different num_worlds / capacity / init agents");
    if ((uint64_t)h.totalAgents > (uint64_t)s->d.numWorlds * s->d.cap) return fail(MBOTS_EINVAL, "corrupt checkpoint");
    if (s->dirty) return fail(MBOTS_ESTATE, "asynchronous work pending: call mbots_sync first");
    DeviceGuard g(s->device);
    cudaStream_t st = (cudaStream_t)stream;
    CU(cudaDeviceSynchronize());                /* nothing on any stream (a caller's copy stream, a shift) may still touch the arrays */
    std::vector<Segment> segs;
    ckpt_segments(s, 0, h.totalAgents, segs);
    if (ckpt_total(segs) != h.bytes || bytes < h.bytes) return fail(MBOTS_EINVAL, "checkpoint size mismatch");
    const unsigned char *p = static_cast<const unsigned char *>(host) + sizeof(CkptHeader);
    for (const Segment &sg : segs) {
        if (sg.bytes) CU(cudaMemcpyAsync(sg.ptr, p, sg.bytes, cudaMemcpyHostToDevice, st));
        p += (sg.bytes + 15u) & ~(size_t)15u;
    }
    const size_t tiles = (s->d.numWorlds + kScanTile - 1) / kScanTile;
    CU(cudaMemsetAsync(s->d.tileSum[0], 0, tiles * sizeof(TileSum), st));
    CU(cudaMemsetAsync(s->d.tileSum[1], 0, tiles * sizeof(TileSum), st));
    CU(cudaMemcpyAsync(s->d.totals, &h.totals, sizeof(Totals), cudaMemcpyHostToDevice, st));
    /* a truncated or corrupt blob must not make the next step index out of bounds: check what the kernels trust */
    launch_validate_state(s->d, h.totalAgents, reinterpret_cast<unsigned int *>(s->dStats), st);
    CU(cudaMemcpyAsync(s->hStats, s->dStats, 2 * sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    {
        const unsigned int *fl = reinterpret_cast<const unsigned int *>(s->hStats);
        if (fl[0] != 0 || fl[1] != h.totalAgents || h.totals.totalAgents != h.totalAgents) {
            /* leave the handle in a defined state: a fresh simulator of the same configuration */
            launch_init(s->d, st);
            s->cur = 0;
            enqueue_repack(s, st, 0, 0, 0);
            finish(s, st);
            char msg[160];
            snprintf(msg, sizeof msg, "corrupt checkpoint (flags %u, live agents %u vs header %u): handle re-initialised", fl[0], fl[1],
                     h.totalAgents);
            return fail(MBOTS_EINVAL, msg);
        }
    }
    s->d.worldIdOffset = h.worldIdOffset; s->d.seed = h.seed;
    s->d.obs[0].worldIdOffset = s->d.obs[1].worldIdOffset = h.worldIdOffset;
    for (auto &kv : s->graphs) cudaGraphExecDestroy(kv.second.exec);      /* the SimDev is baked into every kernel node */
    s->graphs.clear();
    s->cur = 0;
    *s->hTotals = h.totals;
    s->totalAgents = h.totalAgents;
    s->dirty = false;
    return MBOTS_OK;
}

} /* extern "C" */

/* ---- policy inference ends ---------------------------------------------------- */
extern "C" {

int mbots_pack_observations(mbots_sim *s, int32_t is_prev, uint32_t row_begin, uint32_t row_end, float *out, void *stream)
{
    if (!s || (!out && row_end > row_begin)) return fail(MBOTS_EINVAL, "null argument");
    if (row_begin > row_end || row_end > s->totalAgents) return fail(MBOTS_EINVAL, "row range outside [0, num_agents]");
    DeviceGuard g(s->device);
    launch_pack_obs(s->d.obs[s->cur], is_prev != 0, row_begin, row_end - row_begin, out, (cudaStream_t)stream);
    CU(cudaGetLastError());
    s->launches += row_end > row_begin ? 1 : 0;
    return MBOTS_OK;
}

int mbots_sample_actions(mbots_sim *s, uint32_t row_begin, uint32_t row_end, const float *weights, uint32_t seed,
                         uint32_t draw, int32_t *out_index, void *stream)
{
    if (!s || (!weights && row_end > row_begin)) return fail(MBOTS_EINVAL, "null argument");
    if (row_begin > row_end || row_end > s->totalAgents) return fail(MBOTS_EINVAL, "row range outside [0, num_agents]");
    DeviceGuard g(s->device);
    launch_sample(s->d.obs[s->cur], row_begin, row_end - row_begin, weights, seed, draw, out_index, (cudaStream_t)stream);
    CU(cudaGetLastError());
    s->launches += row_end > row_begin ? 1 : 0;
    return MBOTS_OK;
}

} /* extern "C" */

/* ---- fused acting kernel (policy_fused.cu) ------------------------------------------------------------------ */
namespace mbk {
const char *launch_policy_fused(const ObsSet &o, uint32_t row_begin, uint32_t rows, const void *blob, uint64_t blob_bytes,
                                const mbots_policy_program &prog, uint32_t seed, uint32_t draw, int do_sample,
                                float *out_logits, int32_t *out_index, cudaStream_t s);
const char *launch_policy_umma(const ObsSet &o, uint32_t row_begin, uint32_t rows, const void *blob, uint64_t blob_bytes,
                               const mbots_policy_program &prog, uint32_t seed, uint32_t draw, int do_sample,
                               float *out_logits, int32_t *out_index, cudaStream_t s);
}

namespace mbk {
const char *launch_policy_umma_species(const ObsSet &o, const Totals *totals, int species0, int n, const void *const *blobs,
                                       const uint64_t *blob_bytes, const mbots_policy_program *const *progs, uint32_t seed,
                                       uint32_t draw, int do_sample, float *out_logits, int32_t *out_index, cudaStream_t s);
}

static int policy_act_device_rows(mbots_sim *s, int species0, int n, const void *const *blobs, const uint64_t *blob_bytes,
                                  const mbots_policy_program *const *programs, uint32_t seed, uint32_t draw, int32_t sample,
                                  float *out_logits, int32_t *out_index, void *stream)
{
    for (int k = 0; k < n; ++k) {
        if (!blobs[k] || !programs[k]) return fail(MBOTS_EINVAL, "null argument");
        if (programs[k]->layout != MBOTS_LAYOUT_SW128)
            return fail(MBOTS_EINVAL, "device-resolved row ranges take MBOTS_LAYOUT_SW128 programs only");
    }
    DeviceGuard g(s->device);
    const char *err = launch_policy_umma_species(s->d.obs[s->cur], s->d.totals, species0, n, blobs, blob_bytes, programs, seed, draw,
                                                 sample != 0, out_logits, out_index, (cudaStream_t)stream);
    if (err) return fail(MBOTS_EINVAL, err);
    CU(cudaGetLastError());
    s->launches += 1;
    return MBOTS_OK;
}

extern "C" int mbots_policy_act_species(mbots_sim *s, int32_t species, const void *blob, uint64_t blob_bytes,
                                        const mbots_policy_program *program, uint32_t seed, uint32_t draw, int32_t sample,
                                        float *out_logits, int32_t *out_index, void *stream)
{
    if (!s || !blob || !program) return fail(MBOTS_EINVAL, "null argument");
    if (species < 0 || species > 3) return fail(MBOTS_EINVAL, "species must be 0..3");
    return policy_act_device_rows(s, species, 1, &blob, &blob_bytes, &program, seed, draw, sample, out_logits, out_index, stream);
}

extern "C" int mbots_policy_act_all(mbots_sim *s, const void *const blobs[4], const uint64_t blob_bytes[4],
                                    const mbots_policy_program *const programs[4], uint32_t seed, uint32_t draw, int32_t sample,
                                    float *out_logits, int32_t *out_index, void *stream)
{
    if (!s || !blobs || !blob_bytes || !programs) return fail(MBOTS_EINVAL, "null argument");
    return policy_act_device_rows(s, 0, 4, blobs, blob_bytes, programs, seed, draw, sample, out_logits, out_index, stream);
}

extern "C" int mbots_policy_act(mbots_sim *s, uint32_t row_begin, uint32_t row_end, const void *blob, uint64_t blob_bytes,
                                const mbots_policy_program *program, uint32_t seed, uint32_t draw, int32_t sample,
                                float *out_logits, int32_t *out_index, void *stream)
{
    if (!s || !blob || !program) return fail(MBOTS_EINVAL, "null argument");
    if (row_begin > row_end || row_end > s->totalAgents) return fail(MBOTS_EINVAL, "row range outside [0, num_agents]");
    DeviceGuard g(s->device);
    if (program->layout != MBOTS_LAYOUT_ROWPAD && program->layout != MBOTS_LAYOUT_SW128) return fail(MBOTS_EINVAL, "unknown blob layout");
    const auto launch = program->layout == MBOTS_LAYOUT_SW128 ? launch_policy_umma : launch_policy_fused;
    const char *err = launch(s->d.obs[s->cur], row_begin, row_end - row_begin, blob, blob_bytes, *program, seed, draw,
                             sample != 0, out_logits, out_index, (cudaStream_t)stream);
    if (err) return fail(MBOTS_EINVAL, err);
    CU(cudaGetLastError());
    s->launches += row_end > row_begin ? 1 : 0;
    return MBOTS_OK;
}
