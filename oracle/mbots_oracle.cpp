/* mbots_oracle.cpp -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the per-step hot path of llGuy/madrona-bots: the Step,
 * Sensor and ShiftObservations task graphs of src/sim/sim.cpp plus the 1-D
 * sensor raycast of external/madrona (M/). It is the CHECKER the CUDA library
 * is compared against; nothing under madrona-bots_b200/ may include, link or
 * call it. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it.
 *
 * Parity status: PINNED to the reference's own source text. The reference has no CPU backend and
 * does not compile at its pinned submodule commit, and its tests hold no
 * golden vectors for the step -- but its system bodies are plain C++ once an `Engine` exists:
 * oracle/refmock/ compiles src/sim/sim.cpp, external/madrona/src/render/ecs_system.cpp and
 * raySphereIntersect / traceRayTLAS / bvhRaycastEntry of external/madrona/src/mw/device/
 * bvh_raycast.cpp FROM WHERE THEY LIE (`make -C oracle refsim` -> oracle/_ref/librefsim.so) and
 * runs one world through the reference's own task graphs under a host mock of the ECS runtime.
 * tests/test_reference_bodies.py requires every export of this oracle to equal that world's, bit
 * for bit, at every step (positions, rotations, health, food, RNG stream, rewards, surroundings,
 * statistics, 32-ray depth rows, finder targets, prev columns), under the rules U5, U6, U8, U19,
 * D2, D3 of DEVIATIONS.md -- each switchable, with the literal text's differential tested too.
 * tests/golden/refsim_rollouts.npz holds digests of such worlds for boxes without /root/reference.
 * Also pinned: the RNG and quaternion arithmetic, by building this same file against the
 * reference's own rand.hpp / math.hpp (-DMBO_REF_PRIMS -> oracle/_ref/) with bit-identical
 * trajectories, and the known answers of M/tests/rand.cpp:131-141 and M/tests/math.cpp:23-48.
 * Still unpinned: the semantic sensor channel (no source for it exists at the pinned commit) and
 * the cross-world row order (global atomics in the reference; canonical rule U11 here).
 *
 * One world is advanced by one thread; worlds are independent, so a small
 * std::thread pool parallelises over worlds (no OpenMP runtime dependency). All arithmetic is IEEE single precision, compiled
 * with -ffp-contract=off.
 */
#include <cstdint>
#include <cstring>
#include <cmath>
#include <vector>
#include <algorithm>
#include <thread>
#include <mutex>
#include <condition_variable>
#include <atomic>
#include <functional>

#include "oracle_prims.h"
#include "../include/mbots_detmath.h"
#include "../include/mbots.h"

namespace mbo {

/* Minimal persistent thread pool: parallel_for over [0, n) in contiguous blocks. */
class Pool {
public:
    explicit Pool(int n) { resize(n); }
    ~Pool() { stop(); }
    int size() const { return nthreads_; }
    void resize(int n)
    {
        stop();
        nthreads_ = std::max(1, n);
        quit_ = false;
        for (int i = 1; i < nthreads_; ++i) workers_.emplace_back([this, i, e = epoch_] { loop(i, e); });
    }
    void parallel_for(int64_t n, const std::function<void(int64_t)> &fn)
    {
        if (nthreads_ == 1 || n < 2) { for (int64_t i = 0; i < n; ++i) fn(i); return; }
        {
            std::lock_guard<std::mutex> lk(m_);
            fn_ = &fn; n_ = n; pending_ = nthreads_ - 1; ++epoch_;
        }
        cv_.notify_all();
        run_block(0);
        std::unique_lock<std::mutex> lk(m_);
        done_.wait(lk, [this] { return pending_ == 0; });
    }
private:
    void run_block(int t)
    {
        int64_t per = (n_ + nthreads_ - 1) / nthreads_;
        int64_t b = per * t, e = std::min<int64_t>(n_, b + per);
        for (int64_t i = b; i < e; ++i) (*fn_)(i);
    }
    void loop(int t, uint64_t seen)   /* seen: the epoch at creation (a resized pool must not replay old work) */
    {
        for (;;) {
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [&] { return quit_ || epoch_ != seen; });
                if (quit_) return;
                seen = epoch_;
            }
            run_block(t);
            {
                std::lock_guard<std::mutex> lk(m_);
                if (--pending_ == 0) done_.notify_one();
            }
        }
    }
    void stop()
    {
        { std::lock_guard<std::mutex> lk(m_); quit_ = true; }
        cv_.notify_all();
        for (auto &w : workers_) w.join();
        workers_.clear();
    }
    int nthreads_ = 1;
    std::vector<std::thread> workers_;
    std::mutex m_;
    std::condition_variable cv_, done_;
    const std::function<void(int64_t)> *fn_ = nullptr;
    int64_t n_ = 0;
    int pending_ = 0;
    uint64_t epoch_ = 0;
    bool quit_ = false;
};

static Pool &pool()
{
    static Pool p((int)std::max(1u, std::thread::hardware_concurrency()));
    return p;
}

constexpr uint32_t kNumSpecies = 4;     /* src/sim/types.hpp:12 */
constexpr uint32_t kHidden = 16;        /* src/sim/types.hpp:13 */
constexpr uint32_t kChunkWidth = 16;    /* src/sim/types.hpp:78 */
constexpr uint32_t kMaxFoodPackages = 5;/* src/sim/types.hpp:79 */
constexpr uint32_t kChunksX = 8, kChunksY = 6;    /* src/entry/mgr.cpp:106-107 */
constexpr float kCellDim = 1.f;                    /* src/entry/mgr.cpp:108 */
constexpr uint32_t kTotalAllowedFood = 30;         /* src/entry/mgr.cpp:111 */
constexpr uint32_t kSensor = 32;                   /* src/entry/entry.cpp:27 */
constexpr float kNearSphere = 1.1f;                /* src/entry/mgr.cpp:133 */
constexpr float kPi = 3.14159265358979323846264338327950288f; /* M/include/madrona/math.hpp:26 */

struct FoodPackage { uint32_t numFood; uint8_t x, y; };          /* types.hpp:59-69 */
struct ChunkInfo {                                                /* types.hpp:76-103 */
    uint32_t numAgents, totalSpeed;
    FoodPackage pkg[kMaxFoodPackages];
};

struct Action { int32_t forward, backward, rotateLeft, rotateRight, shoot, breed; }; /* types.hpp:22-33 */

struct Agent {                 /* archetype Agent, types.hpp:187-223 (columns the step touches) */
    Vec3 pos;
    Quat rot;
    uint32_t species;          /* 1..4 */
    int32_t health, accum;
    Action action;
    float surr[2];
    int32_t stats[4];          /* hitFriendly, hitEnemy, ateFood, reproduced */
    int32_t target;            /* last sensor pass: finder hit as slot in this world, -1 = none */
    int64_t obs_row;           /* row of this agent's observation record; -1 = created this step */
    bool dead;
};

struct World {
    Rng rng;
    std::vector<Agent> agents;
    ChunkInfo chunks[kChunksX * kChunksY];
    int32_t currentNumFood = 0;
    uint32_t trackCount[kNumSpecies] = {0, 0, 0, 0};
    uint32_t trackHealth[kNumSpecies] = {0, 0, 0, 0};
    uint32_t counts[kNumSpecies] = {0, 0, 0, 0};
    float rewards[kNumSpecies] = {0, 0, 0, 0};
    uint32_t lastHealth[kNumSpecies] = {0, 0, 0, 0};
    /* per-step bookkeeping for the statistics vector */
    int64_t births = 0, deaths = 0, respawns = 0, overflow = 0, eaten = 0, shots = 0;
};

/* One set of row-indexed exports (AgentObservationArchetype columns, types.hpp:228-252,
 * plus the raycast output rows, which U19 aligns with the observation rows). */
struct ObsSet {
    std::vector<int32_t> action, prev_action;       /* 6 / row */
    std::vector<float> hidden, prev_hidden;         /* 16 */
    std::vector<float> reward, prev_reward;         /* 1 */
    std::vector<float> position, prev_position;     /* 2 */
    std::vector<float> health, prev_health;         /* 1  (U4: numeric float) */
    std::vector<float> surrounding, prev_surrounding; /* 2 */
    std::vector<int32_t> stats, prev_stats;         /* 4 */
    std::vector<int8_t> semantic, prev_semantic;    /* 32 */
    std::vector<uint8_t> depth, prev_depth;         /* 32 */
    void resize(size_t rows)
    {
        action.assign(rows * 6, 0); prev_action.assign(rows * 6, 0);
        hidden.assign(rows * kHidden, 0.f); prev_hidden.assign(rows * kHidden, 0.f);
        reward.assign(rows, 0.f); prev_reward.assign(rows, 0.f);
        position.assign(rows * 2, 0.f); prev_position.assign(rows * 2, 0.f);
        health.assign(rows, 0.f); prev_health.assign(rows, 0.f);
        surrounding.assign(rows * 2, 0.f); prev_surrounding.assign(rows * 2, 0.f);
        stats.assign(rows * 4, 0); prev_stats.assign(rows * 4, 0);
        semantic.assign(rows * kSensor, 0); prev_semantic.assign(rows * kSensor, 0);
        depth.assign(rows * kSensor, 0); prev_depth.assign(rows * kSensor, 0);
    }
};

struct Sim {
    uint32_t numWorlds, seed, initAgents, cap, worldIdOffset;
    std::vector<World> worlds;
    ObsSet obs[2];
    int cur = 0;
    uint32_t totalAgents = 0;
    std::vector<int32_t> speciesCount;   /* numWorlds x 4 (SpeciesCount export) */
    std::vector<int32_t> rowBase;        /* numWorlds x 4 */
    std::vector<int32_t> worldOffset, worldCount;
    std::vector<int32_t> sensorIndex, done, reset;
    int64_t statAgentSteps = 0, statSteps = 0;
    /* debug dumps (filled on demand) */
    std::vector<int32_t> dbgI32;
    std::vector<float> dbgF32;
    std::vector<uint8_t> dbgU8;
};

/* ------------------------------------------------------------------------- */
/* world helpers (src/sim/sim.inl:39-73)                                      */

static inline int32_t chunkIndex(float cx, float cy)
{
    int32_t x = (int32_t)cx, y = (int32_t)cy;
    if (x < 0 || y < 0 || x >= (int32_t)kChunksX || y >= (int32_t)kChunksY) return -1;
    return x + y * (int32_t)kChunksX;
}

/* Vector2 / float is "multiply by 1.f / o" in the reference (math.inl:129-134). */
static inline float vdiv(float v, float o) { float inv = 1.f / o; return v * inv; }

/* makeAgent (src/sim/sim.cpp:196-231). Returns false when the world's slot table
 * is full (capacity is a property of the fixed-size restatement, DEVIATIONS.md D1). */
static bool makeAgent(World &w, uint32_t cap, Vec3 pos, uint32_t species, int32_t health)
{
    if (w.agents.size() >= cap) { w.overflow++; return false; }
    Agent a{};
    a.pos = pos;
    a.rot = quat_angle_axis_z(0.f);
    a.species = species;
    a.health = health;
    a.accum = health;
    a.target = -1;
    a.obs_row = -1;
    a.dead = false;
    w.agents.push_back(a);
    return true;
}

/* Sim::Sim + initWorld + initializeChunks (src/sim/sim.cpp:1232-1256, 233-300) */
static void initWorld(Sim &s, uint32_t wi)
{
    World &w = s.worlds[wi];
    w.rng.k = key_split(key_init(s.seed), 0u, s.worldIdOffset + wi);
    w.rng.count = 0;
    const float limx = (float)(kChunksX * kChunkWidth) * kCellDim;
    const float limy = (float)(kChunksY * kChunkWidth) * kCellDim;
    for (uint32_t i = 0; i < s.initAgents; ++i) {
        uint32_t species = (i % kNumSpecies) + 1;
        float x = w.rng.sample_uniform() * limx;
        float y = w.rng.sample_uniform() * limy;
        makeAgent(w, s.cap, Vec3{x, y, 1.f}, species, 100);
    }
    for (auto &c : w.chunks) {
        c.numAgents = 0; c.totalSpeed = 0;
        for (auto &p : c.pkg) { p.numFood = 0; p.x = 0; p.y = 0; }
    }
    w.currentNumFood = 0;
}

/* addFoodToChunk (src/sim/sim.cpp:308-361). The two draws at :311-312 are
 * shadowed and unused but still advance the stream (U12); the top-up branch
 * at :354-357 cannot fire with kMaxFoodPerPackage == 1. */
static bool addFoodToChunk(World &w, ChunkInfo &c)
{
    (void)w.rng.sample_i32(0, kChunkWidth);
    (void)w.rng.sample_i32(0, kChunkWidth);
    for (uint32_t i = 0; i < kMaxFoodPackages; ++i) {
        FoodPackage &p = c.pkg[i];
        if (p.numFood == 0) {
            uint32_t rx = (uint32_t)w.rng.sample_i32(0, kChunkWidth);
            uint32_t ry = (uint32_t)w.rng.sample_i32(0, kChunkWidth);
            p.x = (uint8_t)rx;
            p.y = (uint8_t)ry;
            p.numFood = 1;
            (void)w.rng.sample_uniform(); /* yaw of the render-only food entity (:340) */
            return true;
        }
    }
    return false;
}

/* addFoodSystem (src/sim/sim.cpp:363-387) */
static void addFoodSystem(World &w)
{
    if (w.rng.sample_i32(0, 10) == 0) {
        uint32_t n = (uint32_t)w.rng.sample_i32(1, 3);
        uint32_t allowed = kTotalAllowedFood - (uint32_t)w.currentNumFood; /* unsigned, U12 */
        n = std::min(n, allowed);
        for (uint32_t i = 0; i < n; ++i) {
            uint32_t cx = (uint32_t)w.rng.sample_i32(0, kChunksX);
            uint32_t cy = (uint32_t)w.rng.sample_i32(0, kChunksY);
            if (addFoodToChunk(w, w.chunks[cx + cy * kChunksX])) w.currentNumFood += 1;
        }
    }
}

/* actionSystem (src/sim/sim.cpp:419-502). `a.action` was loaded from the
 * agent's observation row by the caller (:430). */
static void actionSystem(World &w, Agent &a)
{
    if (a.action.shoot) {
        if (a.target >= 0) {
            Agent &t = w.agents[(size_t)a.target];
            t.accum += -50;
            w.shots++;
            if (t.species == a.species) a.stats[0] = 1; else a.stats[1] = 1;
        }
    }
    if (a.action.rotateLeft) a.rot = quat_mul(a.rot, quat_angle_axis_z(0.1f));
    else if (a.action.rotateRight) a.rot = quat_mul(a.rot, quat_angle_axis_z(-0.1f));

    Vec3 old = a.pos;
    Vec3 view = quat_rotate(a.rot, Vec3{1.f, 0.f, 0.f});
    view = vec_normalize(view);
    if (a.action.forward) { a.pos.x += view.x; a.pos.y += view.y; a.pos.z += view.z; }
    else if (a.action.backward) { a.pos.x -= view.x; a.pos.y -= view.y; a.pos.z -= view.z; }

    const float limx = kCellDim * (float)kChunkWidth * (float)kChunksX;
    const float limy = kCellDim * (float)kChunkWidth * (float)kChunksY;
    a.pos.x = std::min(limx - 1.0f, std::max(0.f, a.pos.x));
    a.pos.y = std::min(limy - 1.0f, std::max(0.f, a.pos.y));

    Vec3 d{a.pos.x - old.x, a.pos.y - old.y, a.pos.z - old.z};
    float len = vec_length(d);

    /* getChunkCoord (sim.inl:39-48) + getChunkIndex */
    float ccx = std::floor(vdiv(vdiv(a.pos.x, kCellDim), (float)kChunkWidth));
    float ccy = std::floor(vdiv(vdiv(a.pos.y, kCellDim), (float)kChunkWidth));
    int32_t ci = chunkIndex(ccx, ccy);
    ChunkInfo &c = w.chunks[ci];
    c.numAgents += 1;
    c.totalSpeed += (uint32_t)(len * 2.f);
}

/* healthSync (src/sim/sim.cpp:505-581); FoodPackage::consume (sim.inl:76-99).
 * Agents run in slot order, which realises "lowest slot wins" for the food CAS
 * race (U11). Newborns are appended and not visited by this loop. */
static void healthSync(World &w, uint32_t cap, size_t idx)
{
    Agent &a = w.agents[idx];
    a.health = a.accum;
    {
        float cpx = vdiv(a.pos.x, kCellDim), cpy = vdiv(a.pos.y, kCellDim);
        float chx = vdiv(cpx, (float)kChunkWidth), chy = vdiv(cpy, (float)kChunkWidth);
        uint8_t x = (uint8_t)((float)kChunkWidth * (chx - std::floor(chx)));
        uint8_t y = (uint8_t)((float)kChunkWidth * (chy - std::floor(chy)));
        int32_t ci = chunkIndex(chx, chy);
        ChunkInfo &c = w.chunks[ci];
        for (uint32_t i = 0; i < kMaxFoodPackages; ++i) {
            FoodPackage &p = c.pkg[i];
            if (p.x == x && p.y == y) {
                if (p.numFood != 0) {
                    p.numFood -= 1;
                    if (p.numFood == 0) w.currentNumFood -= 1;
                    a.health += 20;
                    a.stats[2] = 1;
                    w.eaten++;
                    break;
                }
            }
        }
    }
    if (a.action.breed && a.health > 10) {
        if (a.target >= 0) {
            if (w.agents[(size_t)a.target].species == a.species) {
                a.health -= 40;
                Vec3 p = a.pos;
                uint32_t sp = a.species;
                /* push_back may move `a`; finish with the index */
                if (makeAgent(w, cap, p, sp, 50)) w.births++;
                w.agents[idx].stats[3] = 1;
            }
        }
    }
    Agent &b = w.agents[idx];
    if (b.health <= 0) { b.dead = true; w.deaths++; }
    b.accum = b.health;
}

/* updateSurroundingObservation (src/sim/sim.cpp:583-654) */
static void surroundings(World &w, Agent &a)
{
    float cpx = vdiv(a.pos.x, kCellDim) - (float)kChunkWidth * 0.5f;
    float cpy = vdiv(a.pos.y, kCellDim) - (float)kChunkWidth * 0.5f;
    float chx = vdiv(cpx, (float)kChunkWidth), chy = vdiv(cpy, (float)kChunkWidth);
    float fx = std::floor(chx), fy = std::floor(chy), cx = std::ceil(chx), cy = std::ceil(chy);
    int32_t i00 = chunkIndex(fx, fy), i10 = chunkIndex(cx, fy), i01 = chunkIndex(fx, cy), i11 = chunkIndex(cx, cy);
    float xi = chx - fx, yi = chy - fy;
    auto na = [&](int32_t i) { return i < 0 ? 0.f : (float)w.chunks[i].numAgents; };
    auto ts = [&](int32_t i) { return i < 0 ? 0.f : (float)w.chunks[i].totalSpeed; };
    float n00 = na(i00), n10 = na(i10), n01 = na(i01), n11 = na(i11);
    float s00 = ts(i00), s10 = ts(i10), s01 = ts(i01), s11 = ts(i11);
    float nx0 = xi * n10 + (1.f - xi) * n00;
    float nx1 = xi * n11 + (1.f - xi) * n01;
    float sx0 = xi * s10 + (1.f - xi) * s00;
    float sx1 = xi * s11 + (1.f - xi) * s01;
    a.surr[0] = yi * nx1 + (1.f - yi) * nx0;
    a.surr[1] = yi * sx1 + (1.f - yi) * sx0;
}

/* speciesInfoSync (src/sim/sim.cpp:791-838) */
static void speciesInfoSync(Sim &s, World &w)
{
    const float limx = (float)(kChunksX * kChunkWidth) * kCellDim;
    const float limy = (float)(kChunksY * kChunkWidth) * kCellDim;
    uint32_t perSpecies = s.initAgents / kNumSpecies;
    for (uint32_t i = 0; i < kNumSpecies; ++i) {
        uint32_t count = w.trackCount[i];
        uint32_t total = w.trackHealth[i];
        float avg = (float)total / (float)count;
        if (count == 0) avg = 0;
        w.counts[i] = count;
        w.lastHealth[i] = total;
        w.rewards[i] = (float)count / (float)s.initAgents + avg / 100.0f - 2.f;
        w.trackCount[i] = 0;
        w.trackHealth[i] = 0;
        if (count < perSpecies) {
            for (uint32_t e = count; e < perSpecies; ++e) {
                float x = w.rng.sample_uniform() * limx;
                float y = w.rng.sample_uniform() * limy;
                if (makeAgent(w, s.cap, Vec3{x, y, 1.f}, i + 1, 100)) w.respawns++;
            }
        }
    }
}

/* Step graph nodes resetChunkInfo .. speciesInfoSync for one world
 * (src/sim/sim.cpp:1061-1127), then the world-sort compaction (:1129-1132). */
static void stepWorld(Sim &s, uint32_t wi)
{
    World &w = s.worlds[wi];
    const ObsSet &o = s.obs[s.cur];
    for (auto &c : w.chunks) { c.numAgents = 0; c.totalSpeed = 0; }   /* :390-396 */
    addFoodSystem(w);
    const size_t n0 = w.agents.size();
    for (size_t i = 0; i < n0; ++i) {
        Agent &a = w.agents[i];
        const int32_t *act = &o.action[(size_t)a.obs_row * 6];
        a.action = Action{act[0], act[1], act[2], act[3], act[4], act[5]};
        actionSystem(w, a);
    }
    for (size_t i = 0; i < n0; ++i) healthSync(w, s.cap, i);
    for (auto &a : w.agents) if (!a.dead) surroundings(w, a);
    for (auto &a : w.agents) if (!a.dead) {                           /* :719-734 */
        w.trackCount[a.species - 1] += 1;
        w.trackHealth[a.species - 1] += (uint32_t)a.health;
    }
    speciesInfoSync(s, w);
    /* SortArchetypeNode<Agent, WorldID>: stable, drops destroyed rows (U11) */
    w.agents.erase(std::remove_if(w.agents.begin(), w.agents.end(), [](const Agent &a) { return a.dead; }),
                   w.agents.end());
}

/* ------------------------------------------------------------------------- */
/* Sensor graph for one agent: viewTransformUpdate (M/src/render/ecs_system.cpp:154-186)
 * + bvhRaycastEntry / traceRayTLAS / raySphereIntersect (M/src/mw/device/bvh_raycast.cpp).
 * The BVH is replaced by an exhaustive nearest-hit search (U8). */
struct RayHit { bool hit; float t; int32_t slot; };

static inline RayHit traceRay(const World &w, size_t self, Vec3 o, Vec3 d)
{
    const float eps = 0.00001f;                      /* :102-114 */
    if (d.x == 0.f) d.x += eps;
    if (d.y == 0.f) d.y += eps;
    if (d.z == 0.f) d.z += eps;
    RayHit best{false, 0.f, -1};
    for (size_t j = 0; j < w.agents.size(); ++j) {
        (void)self; /* the own sphere is rejected arithmetically (near root = -1 < 1.1) */
        const Agent &b = w.agents[j];
        /* instance: position.xy, zOffset = pos.z, radius = scale.x = 1 (:166-171) */
        Vec3 oc{o.x - b.pos.x, o.y - b.pos.y, o.z - b.pos.z};
        float a = d.x * d.x + d.y * d.y + d.z * d.z;
        float bq = 2.0f * (oc.x * d.x + oc.y * d.y + oc.z * d.z);
        float c = (oc.x * oc.x + oc.y * oc.y + oc.z * oc.z) - 1.f * 1.f;
        float disc = bq * bq - 4 * a * c;
        if (disc < 0) continue;
        float t = (-bq - sqrtf(disc)) / (2.0f * a);
        if (t > kNearSphere) {
            if (!best.hit || t < best.t) { best.hit = true; best.t = t; best.slot = (int32_t)j; }
        }
    }
    return best;
}

static void senseAgent(Sim &s, World &w, size_t idx, ObsSet &o, int64_t row)
{
    Agent &a = w.agents[idx];
    Vec3 rx = quat_rotate(a.rot, Vec3{1.f, 0.f, 0.f});
    float theta = mb_atan2f(rx.y, rx.x);             /* ecs_system.cpp:178 (atan2f, DEVIATIONS.md D2) */
    const float nf = (float)(3 * kSensor / 4), nb = (float)(kSensor / 4);
    Vec3 origin{a.pos.x, a.pos.y, a.pos.z};
    uint8_t *depth = &o.depth[(size_t)row * kSensor];
    int8_t *sem = &o.semantic[(size_t)row * kSensor];
    a.target = -1;
    for (uint32_t i = 0; i < kSensor; ++i) {
        float rt = theta;
        if (i < 3 * kSensor / 4) {
            rt = (rt - kPi * 0.25f) + (float)i * (kPi * 0.5f / (nf - 1.f));
        } else {
            rt += kPi;
            rt = (rt - kPi * 0.25f) + (float)(i - 3 * kSensor / 4) * (kPi * 0.5f / (nb - 1.f));
        }
        float sn, cs;
        mb_sincosf(rt, &sn, &cs);                    /* bvh_raycast.cpp:250 (cosf/sinf, D2) */
        RayHit h = traceRay(w, idx, origin, Vec3{cs, sn, 0.0f});
        if (h.hit) {
            depth[i] = (uint8_t)h.t;
            sem[i] = (int8_t)w.agents[(size_t)h.slot].species;
            if (i == (3 * kSensor / 4) / 2) a.target = h.slot;
        } else {
            depth[i] = 255;
            sem[i] = -1;
        }
        if (i == 0) depth[0] = (uint8_t)(uint32_t)row;  /* "sanity" overwrite, :296-299 (U9) */
    }
    (void)s;
}

/* ------------------------------------------------------------------------- */
/* Row assignment: observation rows are ordered (species, world, slot) -- the
 * stable global sort by SpeciesObservation of src/sim/sim.cpp:1148-1149 applied
 * to records created in world/slot order (U11). */
static void assignRows(Sim &s)
{
    const uint32_t W = s.numWorlds;
    std::vector<int64_t> perSpecies(kNumSpecies, 0);
    std::vector<int32_t> cnt((size_t)W * kNumSpecies, 0);
    for (uint32_t wi = 0; wi < W; ++wi)
        for (auto &a : s.worlds[wi].agents) cnt[(size_t)wi * 4 + (a.species - 1)]++;
    int64_t base = 0;
    for (uint32_t sp = 0; sp < kNumSpecies; ++sp)
        for (uint32_t wi = 0; wi < W; ++wi) {
            s.rowBase[(size_t)wi * 4 + sp] = (int32_t)base;
            base += cnt[(size_t)wi * 4 + sp];
        }
    s.totalAgents = (uint32_t)base;
    int32_t off = 0;
    for (uint32_t wi = 0; wi < W; ++wi) {
        s.worldOffset[wi] = off;
        s.worldCount[wi] = (int32_t)s.worlds[wi].agents.size();
        off += s.worldCount[wi];
    }
}

/* updateObservations (:687-717), the species sort carrying every column of the
 * record (:1148-1149), updateSensorOutputIdx (:736-789), rewardSystem (:840-983),
 * then the sensor graph, for one world. Reads set `from`, writes set `to`. */
static void exportWorld(Sim &s, uint32_t wi, bool withSensors)
{
    World &w = s.worlds[wi];
    const ObsSet &from = s.obs[s.cur];
    ObsSet &to = s.obs[s.cur ^ 1];
    int32_t next[kNumSpecies];
    for (uint32_t sp = 0; sp < kNumSpecies; ++sp) next[sp] = s.rowBase[(size_t)wi * 4 + sp];
    std::vector<int64_t> rows(w.agents.size());
    for (size_t i = 0; i < w.agents.size(); ++i) {
        Agent &a = w.agents[i];
        const int64_t row = next[a.species - 1]++;
        const int64_t old = a.obs_row;
        rows[i] = row;
        /* columns the sort carries along with the record */
        if (old >= 0) {
            memcpy(&to.action[row * 6], &from.action[old * 6], 24);
            memcpy(&to.hidden[row * kHidden], &from.hidden[old * kHidden], 64);
            memcpy(&to.prev_action[row * 6], &from.prev_action[old * 6], 24);
            memcpy(&to.prev_hidden[row * kHidden], &from.prev_hidden[old * kHidden], 64);
            to.prev_reward[row] = from.prev_reward[old];
            memcpy(&to.prev_position[row * 2], &from.prev_position[old * 2], 8);
            to.prev_health[row] = from.prev_health[old];
            memcpy(&to.prev_surrounding[row * 2], &from.prev_surrounding[old * 2], 8);
            memcpy(&to.prev_stats[row * 4], &from.prev_stats[old * 4], 16);
            /* updateSensorOutputIdx: prev sensor rows <- last frame's rows (:774-780) */
            memcpy(&to.prev_semantic[row * kSensor], &from.semantic[old * kSensor], kSensor);
            memcpy(&to.prev_depth[row * kSensor], &from.depth[old * kSensor], kSensor);
        } else { /* record created this step: defined as zero (DEVIATIONS.md D3) */
            memset(&to.action[row * 6], 0, 24);
            memset(&to.hidden[row * kHidden], 0, 64);
            memset(&to.prev_action[row * 6], 0, 24);
            memset(&to.prev_hidden[row * kHidden], 0, 64);
            to.prev_reward[row] = 0.f;
            memset(&to.prev_position[row * 2], 0, 8);
            to.prev_health[row] = 0.f;
            memset(&to.prev_surrounding[row * 2], 0, 8);
            memset(&to.prev_stats[row * 4], 0, 16);
            memset(&to.prev_semantic[row * kSensor], 0, kSensor);
            memset(&to.prev_depth[row * kSensor], 0, kSensor);
        }
        /* updateObservations */
        to.position[row * 2 + 0] = a.pos.x;
        to.position[row * 2 + 1] = a.pos.y;
        to.health[row] = (float)a.health;
        to.surrounding[row * 2 + 0] = a.surr[0];
        to.surrounding[row * 2 + 1] = a.surr[1];
        for (int k = 0; k < 4; ++k) to.stats[row * 4 + k] = a.stats[k];
        /* rewardSystem, reward setting 8 (:943-956) with the U5 index fix */
        float r = w.rewards[a.species - 1] + (float)a.health / 100.f - 0.5f;
        if (a.stats[2]) r += 10.f;
        if (a.stats[3]) r += 10.f;
        if (a.stats[1]) r += 15.f;
        to.reward[row] = withSensors ? r : 0.f;
        a.stats[0] = a.stats[1] = a.stats[2] = a.stats[3] = 0;
        a.obs_row = row;
        s.sensorIndex[(size_t)s.worldOffset[wi] + i] = (int32_t)row;
    }
    if (withSensors)
        for (size_t i = 0; i < w.agents.size(); ++i) senseAgent(s, w, i, to, rows[i]);
    for (uint32_t sp = 0; sp < kNumSpecies; ++sp) s.speciesCount[(size_t)wi * 4 + sp] = (int32_t)w.counts[sp];
}

static void repack(Sim &s, bool withSensors)
{
    assignRows(s);
    pool().parallel_for(s.numWorlds, [&](int64_t wi) { exportWorld(s, (uint32_t)wi, withSensors); });
    s.cur ^= 1;
}

static void step(Sim &s)
{
    pool().parallel_for(s.numWorlds, [&](int64_t wi) { stepWorld(s, (uint32_t)wi); });
    repack(s, true);
    s.statSteps += 1;
    s.statAgentSteps += s.totalAgents;
}

/* shiftObservationsSystem + shiftHiddenState (src/sim/sim.cpp:1002-1048) */
static void shift(Sim &s)
{
    ObsSet &o = s.obs[s.cur];
    const size_t A = s.totalAgents;
    memcpy(o.prev_position.data(), o.position.data(), A * 8);
    memcpy(o.prev_health.data(), o.health.data(), A * 4);
    memcpy(o.prev_surrounding.data(), o.surrounding.data(), A * 8);
    memcpy(o.prev_reward.data(), o.reward.data(), A * 4);
    memcpy(o.prev_action.data(), o.action.data(), A * 24);
    for (size_t r = 0; r < A; ++r) {
        o.prev_stats[r * 4 + 0] = o.stats[r * 4 + 0];
        o.prev_stats[r * 4 + 1] = o.stats[r * 4 + 0];   /* [sic] :1034 (U7) */
        o.prev_stats[r * 4 + 2] = o.stats[r * 4 + 2];
        o.prev_stats[r * 4 + 3] = o.stats[r * 4 + 3];
    }
    memcpy(o.prev_hidden.data(), o.hidden.data(), A * 64);
}

} // namespace mbo

/* ------------------------------------------------------------------------- */
/* C interface, mirroring include/mbots.h with host pointers.                  */
using namespace mbo;

extern "C" {

const char *mbo_flavour(void) { return prims_flavour(); }

int mbo_create(uint32_t num_worlds, uint32_t seed, uint32_t init_agents, uint32_t cap,
               uint32_t world_id_offset, void **out)
{
    if (!out || num_worlds == 0) return MBOTS_EINVAL;
    if (cap == 0) {
        cap = std::max<uint32_t>(128u, 4u * init_agents);
        cap = (cap + 31u) & ~31u;
    }
    if (cap < init_agents) return MBOTS_EINVAL;
    Sim *s = new Sim();
    s->numWorlds = num_worlds; s->seed = seed; s->initAgents = init_agents; s->cap = cap;
    s->worldIdOffset = world_id_offset;
    s->worlds.resize(num_worlds);
    size_t rows = (size_t)num_worlds * cap;
    s->obs[0].resize(rows); s->obs[1].resize(rows);
    s->speciesCount.assign((size_t)num_worlds * 4, 0);
    s->rowBase.assign((size_t)num_worlds * 4, 0);
    s->worldOffset.assign(num_worlds, 0); s->worldCount.assign(num_worlds, 0);
    s->sensorIndex.assign(rows, 0); s->done.assign(rows, 0); s->reset.assign(num_worlds, 0);
    pool().parallel_for(num_worlds, [&](int64_t wi) { initWorld(*s, (uint32_t)wi); });
    repack(*s, false);
    *out = s;
    return MBOTS_OK;
}

int mbo_destroy(void *h) { delete (Sim *)h; return MBOTS_OK; }
int mbo_step(void *h) { step(*(Sim *)h); return MBOTS_OK; }
int mbo_shift_observations(void *h) { shift(*(Sim *)h); return MBOTS_OK; }
int mbo_num_agents(void *h, uint32_t *out) { *out = ((Sim *)h)->totalAgents; return MBOTS_OK; }
int mbo_set_threads(int n) { pool().resize(n); return pool().size(); }
int mbo_max_threads(void) { return pool().size(); }

int mbo_tensor(void *h, int32_t id, void **ptr, int64_t dims[2], int32_t *dtype)
{
    Sim &s = *(Sim *)h;
    ObsSet &o = s.obs[s.cur];
    const int64_t A = s.totalAgents, W = s.numWorlds;
    auto set = [&](void *p, int64_t d0, int64_t d1, int32_t dt) { *ptr = p; dims[0] = d0; dims[1] = d1; *dtype = dt; return MBOTS_OK; };
    switch (id) {
    case MBOTS_EXPORT_RESET: return set(s.reset.data(), W, 1, MBOTS_I32);
    case MBOTS_EXPORT_ACTION: return set(o.action.data(), A, 6, MBOTS_I32);
    case MBOTS_EXPORT_PREV_ACTION: return set(o.prev_action.data(), A, 6, MBOTS_I32);
    case MBOTS_EXPORT_HIDDEN_STATE: return set(o.hidden.data(), A, 16, MBOTS_F32);
    case MBOTS_EXPORT_PREV_HIDDEN_STATE: return set(o.prev_hidden.data(), A, 16, MBOTS_F32);
    case MBOTS_EXPORT_REWARD: return set(o.reward.data(), A, 1, MBOTS_F32);
    case MBOTS_EXPORT_PREV_REWARD: return set(o.prev_reward.data(), A, 1, MBOTS_F32);
    case MBOTS_EXPORT_DONE: return set(s.done.data(), A, 1, MBOTS_I32);
    case MBOTS_EXPORT_POSITION: return set(o.position.data(), A, 2, MBOTS_F32);
    case MBOTS_EXPORT_PREV_POSITION: return set(o.prev_position.data(), A, 2, MBOTS_F32);
    case MBOTS_EXPORT_HEALTH: return set(o.health.data(), A, 1, MBOTS_F32);
    case MBOTS_EXPORT_PREV_HEALTH: return set(o.prev_health.data(), A, 1, MBOTS_F32);
    case MBOTS_EXPORT_SURROUNDING: return set(o.surrounding.data(), A, 2, MBOTS_F32);
    case MBOTS_EXPORT_PREV_SURROUNDING: return set(o.prev_surrounding.data(), A, 2, MBOTS_F32);
    case MBOTS_EXPORT_SENSOR_SEMANTIC: return set(o.semantic.data(), A, 32, MBOTS_I8);
    case MBOTS_EXPORT_SENSOR_DEPTH: return set(o.depth.data(), A, 32, MBOTS_U8);
    case MBOTS_EXPORT_PREV_SENSOR_SEMANTIC: return set(o.prev_semantic.data(), A, 32, MBOTS_I8);
    case MBOTS_EXPORT_PREV_SENSOR_DEPTH: return set(o.prev_depth.data(), A, 32, MBOTS_U8);
    case MBOTS_EXPORT_STATS: return set(o.stats.data(), A, 4, MBOTS_I32);
    case MBOTS_EXPORT_PREV_STATS: return set(o.prev_stats.data(), A, 4, MBOTS_I32);
    case MBOTS_EXPORT_SENSOR_INDEX: return set(s.sensorIndex.data(), A, 1, MBOTS_I32);
    case MBOTS_EXPORT_SPECIES_COUNT: return set(s.speciesCount.data(), W, 4, MBOTS_I32);
    case MBOTS_DEBUG_WORLD_COUNT: return set(s.worldCount.data(), W, 1, MBOTS_I32);
    case MBOTS_DEBUG_WORLD_OFFSET: return set(s.worldOffset.data(), W, 1, MBOTS_I32);
    case MBOTS_DEBUG_ROW_BASE: return set(s.rowBase.data(), W, 4, MBOTS_I32);
    case MBOTS_DEBUG_RNG_COUNT:
        s.dbgI32.resize(W);
        for (int64_t i = 0; i < W; ++i) s.dbgI32[i] = (int32_t)s.worlds[i].rng.count;
        return set(s.dbgI32.data(), W, 1, MBOTS_I32);
    case MBOTS_DEBUG_NUM_FOOD:
        s.dbgI32.resize(W);
        for (int64_t i = 0; i < W; ++i) s.dbgI32[i] = s.worlds[i].currentNumFood;
        return set(s.dbgI32.data(), W, 1, MBOTS_I32);
    case MBOTS_DEBUG_FOOD:
        s.dbgU8.assign((size_t)W * 48 * 8, 0);
        for (int64_t i = 0; i < W; ++i)
            for (int c = 0; c < 48; ++c) {
                uint8_t *d = &s.dbgU8[((size_t)i * 48 + c) * 8];
                uint8_t mask = 0;
                for (int p = 0; p < 5; ++p) {
                    const FoodPackage &fp = s.worlds[i].chunks[c].pkg[p];
                    d[p] = (uint8_t)((fp.x << 4) | fp.y);
                    if (fp.numFood) mask |= (uint8_t)(1u << p);
                }
                d[5] = mask;
            }
        return set(s.dbgU8.data(), W * 48, 8, MBOTS_U8);
    case MBOTS_DEBUG_POSE:
        s.dbgF32.resize((size_t)A * 4);
        for (int64_t i = 0; i < W; ++i)
            for (size_t k = 0; k < s.worlds[i].agents.size(); ++k) {
                const Agent &a = s.worlds[i].agents[k];
                float *d = &s.dbgF32[((size_t)s.worldOffset[i] + k) * 4];
                d[0] = a.pos.x; d[1] = a.pos.y; d[2] = a.rot.w; d[3] = a.rot.z;
            }
        return set(s.dbgF32.data(), A, 4, MBOTS_F32);
    case MBOTS_DEBUG_AGENT_HEALTH:
    case MBOTS_DEBUG_AGENT_SPECIES:
    case MBOTS_DEBUG_AGENT_TARGET:
        s.dbgI32.resize((size_t)A);
        for (int64_t i = 0; i < W; ++i)
            for (size_t k = 0; k < s.worlds[i].agents.size(); ++k) {
                const Agent &a = s.worlds[i].agents[k];
                s.dbgI32[(size_t)s.worldOffset[i] + k] =
                    id == MBOTS_DEBUG_AGENT_HEALTH ? a.health
                    : id == MBOTS_DEBUG_AGENT_SPECIES ? (int32_t)a.species : a.target;
            }
        return set(s.dbgI32.data(), A, 1, MBOTS_I32);
    default: return MBOTS_EINVAL;
    }
}

int mbo_episode_stats(void *h, int64_t out[MBOTS_NUM_STATS])
{
    Sim &s = *(Sim *)h;
    memset(out, 0, sizeof(int64_t) * MBOTS_NUM_STATS);
    out[0] = s.statAgentSteps; out[1] = s.statSteps;
    for (auto &w : s.worlds) {
        for (int k = 0; k < 4; ++k) { out[2 + k] += w.counts[k]; out[6 + k] += w.lastHealth[k]; }
        out[10] += w.births; out[11] += w.deaths; out[12] += w.respawns;
        out[13] += w.overflow; out[14] += w.eaten; out[15] += w.shots;
    }
    return MBOTS_OK;
}

/* Full-precision views for tests of the quaternion restatement: (x,y,z,w) untouched
 * components must stay (+-0). Returns max |q.x|,|q.y|,|pos.z - 1| over all agents. */
float mbo_debug_planarity(void *h)
{
    Sim &s = *(Sim *)h;
    float m = 0.f;
    for (auto &w : s.worlds)
        for (auto &a : w.agents) m = std::max(m, std::max(std::fabs(a.rot.x), std::max(std::fabs(a.rot.y), std::fabs(a.pos.z - 1.f))));
    return m;
}

/* primitive-level entry points (KATs and ref-header cross-checks) */
void mbo_prim_split(uint32_t a, uint32_t b, uint32_t idx, uint32_t idx_upper, uint32_t out[2])
{
    Key k = key_split(Key{a, b}, idx, idx_upper);
    out[0] = k.a; out[1] = k.b;
}
void mbo_prim_init_key(uint32_t seed, uint32_t out[2])
{
    Key k = key_init(seed);
    out[0] = k.a; out[1] = k.b;
}
int32_t mbo_prim_sample_i32(uint32_t a, uint32_t b, int32_t lo, int32_t hi) { return key_sample_i32(Key{a, b}, lo, hi); }
float mbo_prim_sample_uniform(uint32_t a, uint32_t b) { return key_sample_uniform(Key{a, b}); }
void mbo_prim_angle_axis_z(float angle, float out[4])
{
    Quat q = quat_angle_axis_z(angle);
    out[0] = q.w; out[1] = q.x; out[2] = q.y; out[3] = q.z;
}
void mbo_prim_quat_mul(const float a[4], const float b[4], float out[4])
{
    Quat q = quat_mul(Quat{a[0], a[1], a[2], a[3]}, Quat{b[0], b[1], b[2], b[3]});
    out[0] = q.w; out[1] = q.x; out[2] = q.y; out[3] = q.z;
}
void mbo_prim_rotate(const float q[4], const float v[3], float out[3])
{
    Vec3 r = quat_rotate(Quat{q[0], q[1], q[2], q[3]}, Vec3{v[0], v[1], v[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void mbo_prim_normalize(const float v[3], float out[3])
{
    Vec3 r = vec_normalize(Vec3{v[0], v[1], v[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void mbo_detmath_sincos(const float *x, float *s, float *c, int64_t n)
{
    for (int64_t i = 0; i < n; ++i) mb_sincosf(x[i], &s[i], &c[i]);
}
void mbo_detmath_atan2(const float *y, const float *x, float *out, int64_t n)
{
    for (int64_t i = 0; i < n; ++i) out[i] = mb_atan2f(y[i], x[i]);
}

} /* extern "C" */
