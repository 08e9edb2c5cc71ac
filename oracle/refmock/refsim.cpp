/* refsim.cpp -- ONE world of the reference, run from the reference's own source text under a host mock.
 *
 * TEST INFRASTRUCTURE (oracle/): built by `make -C oracle refsim` into oracle/_ref/librefsim.so, used only by
 * tests/test_reference_bodies.py to pin oracle/mbots_oracle.cpp to the reference. Never linked by the product.
 *
 * What is compiled here FROM WHERE IT LIES under /root/reference (nothing of it is in this repository):
 *   src/sim/sim.cpp (+ sim.hpp, sim.inl, types.hpp)     every system body, Sim::Sim / initWorld, and the
 *                                                       task graphs setupStepTasks / setupShiftObservationsTasks
 *                                                       (via oracle/_ref/gen/ref_sim.inc, see extract_ref.py)
 *   external/madrona/src/render/ecs_system.cpp          instanceTransformUpdate, viewTransformUpdate, the view /
 *                                                       renderable bookkeeping, RenderingSystem::setupTasks
 *   external/madrona/src/mw/device/bvh_raycast.cpp      raySphereIntersect, traceRayTLAS, bvhRaycastEntry
 *   external/madrona/src/mw/device/include/madrona/bvh.hpp   QBVHNode (construct / convertToAABB), BVHParams
 *   external/madrona/include/madrona/{math,rand,sync,ecs}.hpp  arithmetic, RNG, atomics, Entity / Loc
 * What the mock supplies: tables, entity bookkeeping, node execution (mock_ecs.hpp), a 4-wide QBVH built over the
 * instances with the reference's own QBVHNode::construct (the TLAS *build* kernels are out of scope),
 * and the CUDA thread indices bvhRaycastEntry reads.
 *
 * Interventions, each a rule of DEVIATIONS.md and each switchable through `flags` so the differential of the
 * LITERAL text is testable:
 *   MBR_FIX_U5   rewardSystem indexes rewards[speciesID] 1-based (sim.cpp:943): the row is shifted by one float
 *                around that node so the text reads the value the writers stored for the species
 *   MBR_FIX_U6   a dying agent's observation record dies with it (sim.cpp:574-578 leaks it)
 *   MBR_FIX_U19  sensor rows follow the observation rows (ecs_system.cpp:184 resets rowIDX to the agent row)
 *   MBR_FIX_U8   traceRayTLAS keeps the LAST accepted hit in traversal order, not the nearest (bvh_raycast.cpp:173-180).
 *                With the flag the reference's kernel text is run once per agent instance (a one-leaf tree each time)
 *                and the nearest of those single-instance answers is kept per ray: depth byte = minimum, finder =
 *                minimum float t, ties to the lowest agent row. Same text, order-independent result.
 *   MBR_FIX_D3   the prev sensor rows of a record created this step are zero (updateSensorOutputIdx copies them from
 *                the row its zero-initialised SensorOutputIndex names, sim.cpp:771-780: whatever agent sat in row 0)
 *   MBR_DETMATH (compile time)  cosf / sinf / atan2f of the reference text resolve to include/mbots_detmath.h (D2)
 */
#include <cmath>
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>
#include <memory>

#include "../../include/mbots_detmath.h"

#include <madrona/math.hpp>
#include <madrona/rand.hpp>
#include <madrona/sync.hpp>
#include "mock_ecs.hpp"

#ifdef MBR_DETMATH
static inline float mbr_cosf(float x) { float s, c; mb_sincosf(x, &s, &c); return c; }
static inline float mbr_sinf(float x) { float s, c; mb_sincosf(x, &s, &c); return s; }
#define cosf mbr_cosf
#define sinf mbr_sinf
#define atan2f mb_atan2f
#endif

/* ---- the reference's simulation -------------------------------------------------------------- */
#include "ref_sim.inc"
#undef LOG

/* ---- the reference's render-side ECS systems (wrapped whole in #ifdef MADRONA_GPU_MODE; every madrona
 * header is already included above, so the macro only opens this file's own text) ---------------- */
#define MADRONA_GPU_MODE 1
#include "ecs_system.cpp"
#undef MADRONA_GPU_MODE
#undef LOG

/* ---- the reference's raycast kernel ---------------------------------------------------------- */
#define __device__
#define __global__
#include <madrona/bvh.hpp>
namespace {
struct MockDim { uint32_t x, y, z; };
thread_local MockDim blockIdx, threadIdx, gridDim, blockDim;
inline void __syncwarp() {}
}
extern "C" { madrona::BVHParams bvhParams; }
#define LOG(...) madrona::mwGPU::HostPrint::log(__VA_ARGS__)
using namespace madrona;
#include "ref_raycast.inc"
#undef LOG

#ifdef MBR_DETMATH
#undef cosf
#undef sinf
#undef atan2f
#endif

/* ---- driver ---------------------------------------------------------------------------------- */
namespace {

enum : uint32_t { MBR_FIX_U5 = 1u, MBR_FIX_U6 = 2u, MBR_FIX_U19 = 4u, MBR_FIX_U8 = 8u, MBR_FIX_D3 = 16u };

using namespace mbots;
namespace mr = madrona::render;

struct RefWorld {
    StateManager sm;
    TaskGraphManager graphs;
    mbots::Sim::Config cfg;
    mbots::SimBridge bridge;
    void *simStorage = nullptr;
    std::unique_ptr<Engine> ctx;
    uint32_t flags = 0;
    std::vector<QBVHNode> nodes;
    alignas(16) unsigned char internalRaw[sizeof(BVHInternalData)];   /* no default constructor (atomics) */
    BVHInternalData &internal() { return *reinterpret_cast<BVHInternalData *>(internalRaw); }
    std::vector<mr::FinderOutput> finderByRow;
    uint64_t literalRays = 0, nearestRays = 0;      /* U8 bookkeeping of the last sensor pass */

    Sim &sim() { return *static_cast<Sim *>(simStorage); }
};

void run_graph(RefWorld &w, TaskGraphID id)
{
    TaskGraphBuilder &g = w.graphs.graphs[(size_t)id];
    const void *reward_fn = reinterpret_cast<const void *>(+mbots::rewardSystem);
    for (MockNode &n : g.nodes) {
        const bool u5 = (w.flags & MBR_FIX_U5) && n.fn == reward_fn;
        float *r = nullptr;
        if (u5) {       /* rewards[k] -> rewards[k + 1] for the duration of the node (one float of slack follows the row) */
            r = w.ctx->get<SpeciesReward>(w.sim().speciesInfoTracker).rewards;
            memmove(r + 1, r, 4 * sizeof(float));
        }
        n.run(*w.ctx);
        if (u5) memmove(r, r + 1, 4 * sizeof(float));
    }
}

/* A 4-wide QBVH over the instances in table order, nodes made by the reference's QBVHNode::construct. */
void build_tlas(RefWorld &w, const mr::TLBVHNode *aabbs, int32_t n)
{
    constexpr int Wd = QBVHNode::NodeWidth;
    struct Item { math::AABB box; int32_t idx; };
    std::vector<Item> level;
    for (int32_t i = 0; i < n; ++i) level.push_back(Item{aabbs[i].aabb, -(i + 1)});
    std::vector<std::vector<QBVHNode>> levels;
    /* bottom-up; node indices are assigned afterwards (root first), so children are recorded as (level, slot) */
    struct Pending { std::vector<Item> kids; };
    std::vector<std::vector<Pending>> plevels;
    while (true) {
        std::vector<Pending> ps;
        std::vector<Item> up;
        for (size_t i = 0; i < level.size(); i += Wd) {
            Pending p;
            for (size_t k = i; k < std::min(level.size(), i + (size_t)Wd); ++k) p.kids.push_back(level[k]);
            math::AABB box = p.kids[0].box;
            for (const Item &it : p.kids) box = math::AABB::merge(box, it.box);
            up.push_back(Item{box, (int32_t)ps.size()});          /* idx = slot in this level, fixed up below */
            ps.push_back(std::move(p));
        }
        plevels.push_back(std::move(ps));
        if (up.size() <= 1) break;
        level = std::move(up);
    }
    /* flatten: root level last in plevels -> index 0 */
    std::vector<int32_t> base(plevels.size());
    int32_t total = 0;
    for (int l = (int)plevels.size() - 1; l >= 0; --l) { base[l] = total; total += (int32_t)plevels[l].size(); }
    w.nodes.assign((size_t)std::max(total, 1), QBVHNode{});
    for (int l = 0; l < (int)plevels.size(); ++l) {
        for (size_t s = 0; s < plevels[l].size(); ++s) {
            const Pending &p = plevels[l][s];
            math::AABB boxes[Wd];
            QBVHNode::NodeIndexT idx[Wd];
            for (size_t k = 0; k < p.kids.size(); ++k) {
                boxes[k] = p.kids[k].box;
                idx[k] = (QBVHNode::NodeIndexT)(l == 0 ? p.kids[k].idx : base[l - 1] + p.kids[k].idx + 1);
            }
            w.nodes[(size_t)base[l] + s] = QBVHNode::construct((uint32_t)p.kids.size(), boxes, idx);
        }
    }
    if (n == 0) w.nodes[0].numChildren = 0;
}

void run_sensor(RefWorld &w)
{
    StateManager &sm = w.sm;
    run_graph(w, TaskGraphID::Sensor);          /* RenderingSystem::setupTasks: transforms, morton codes, the three sorts */
    MockTable &cams = sm.table<mr::RenderCameraArchetype>();
    MockTable &inst = sm.table<mr::RenderableArchetype>();
    auto *views = sm.getArchetypeComponent<mr::RenderCameraArchetype, mr::PerspectiveCameraData>();
    auto *instances = sm.getArchetypeComponent<mr::RenderableArchetype, mr::InstanceData>();
    auto *aabbs = sm.getArchetypeComponent<mr::RenderableArchetype, mr::TLBVHNode>();
    if (w.flags & MBR_FIX_U19) {
        /* sim.cpp:783 set rowIDX to the observation row; the vendored viewTransformUpdate (ecs_system.cpp:184)
         * put the agent row back. Restore the sim's value. */
        MockTable &agents = sm.table<Agent>();
        for (int32_t r = 0; r < agents.numRows; ++r) {
            if (agents.dead[r]) continue;
            const Entity e = agents.rowEntity[r];
            const mr::RenderCamera &cam = sm.getUnsafe<mr::RenderCamera>(e);
            sm.getUnsafe<mr::PerspectiveCameraData>(cam.cameraEntity).rowIDX = sm.getUnsafe<SensorOutputIndex>(e).idx;
        }
    }
    build_tlas(w, aabbs, inst.numRows);
    memset(w.internalRaw, 0, sizeof w.internalRaw);
    w.internal().traversalNodes = w.nodes.data();
    /* the kernel indexes the per-world arrays with the view's worldIDX (bvh_raycast.cpp:128-129,215): give it arrays
     * long enough for this world's id, with this world's entries filled in */
    const int32_t ninst = inst.numRows, nviews = cams.numRows;
    const size_t wid = (size_t)w.ctx->worldID().idx;
    std::vector<int32_t> zeros(wid + 1, 0), inst_counts(wid + 1, 0), view_counts(wid + 1, 0), one_counts(wid + 1, 1);
    inst_counts[wid] = ninst;
    view_counts[wid] = nviews;
    memset(&bvhParams, 0, sizeof bvhParams);
    bvhParams.numWorlds = (uint32_t)wid + 1;
    bvhParams.instances = instances;
    bvhParams.views = views;
    bvhParams.aabbs = aabbs;
    bvhParams.instanceOffsets = zeros.data();
    bvhParams.instanceCounts = inst_counts.data();
    bvhParams.viewOffsets = zeros.data();
    bvhParams.viewCounts = view_counts.data();
    bvhParams.internalData = &w.internal();
    bvhParams.renderOutput = sm.getArchetypeComponent<mr::RaycastOutputArchetype, mr::DepthOutputBuffer>();
    w.finderByRow.assign((size_t)kMockRows, mr::FinderOutput{Entity::none(), 255.f});
    bvhParams.finderOutput = w.finderByRow.data();
    bvhParams.renderOutputResolution = 32;          /* src/entry/mgr.cpp:132 */
    bvhParams.nearSphere = 1.1f;                    /* src/entry/mgr.cpp:133 */
    /* launch shape of M/src/mw/cuda_exec.cpp:2533-2551: blocks of 128 threads = 4 views; here one block per 4 views */
    blockDim = MockDim{128, 1, 1};
    gridDim = MockDim{(uint32_t)std::max(1, (nviews + 3) / 4), 1, 1};
    auto launch = [&]() {
        for (uint32_t b = 0; b < gridDim.x; ++b)
            for (uint32_t t = 0; t < 128; ++t) {
                blockIdx = MockDim{b, 0, 0};
                threadIdx = MockDim{t, 0, 0};
                if (nviews > 0) bvhRaycastEntry();
            }
    };
    if (!(w.flags & MBR_FIX_U8)) {
        launch();
    } else {
        uint8_t *depth = reinterpret_cast<uint8_t *>(bvhParams.renderOutput);
        std::vector<uint8_t> best_depth((size_t)kMockRows * 32, 255);
        std::vector<mr::FinderOutput> best_finder((size_t)kMockRows, mr::FinderOutput{Entity::none(), 255.f});
        std::vector<int32_t> best_row((size_t)kMockRows, 0x7fffffff);
        std::vector<uint8_t> scratch((size_t)kMockRows * 32);
        bvhParams.renderOutput = scratch.data();
        bvhParams.instanceCounts = one_counts.data();
        bool any = false;
        for (int32_t k = 0; k < ninst; ++k) {
            if (instances[k].objectIDX != 0) continue;           /* only agents are ever hit (bvh_raycast.cpp:164) */
            any = true;
            bvhParams.instances = instances + k;
            build_tlas(w, aabbs + k, 1);
            w.internal().traversalNodes = w.nodes.data();
            std::fill(w.finderByRow.begin(), w.finderByRow.end(), mr::FinderOutput{Entity::none(), 255.f});
            launch();
            const int32_t agent_row = sm.getLoc(instances[k].owner).row;
            for (int32_t c = 0; c < nviews; ++c) {
                const uint32_t row = views[c].rowIDX;
                for (int p = 0; p < 32; ++p)
                    best_depth[(size_t)row * 32 + p] = std::min(best_depth[(size_t)row * 32 + p], scratch[(size_t)row * 32 + p]);
                const mr::FinderOutput &fo = w.finderByRow[row];
                if (fo.hitEntity != Entity::none()) {
                    mr::FinderOutput &bf = best_finder[row];
                    if (bf.hitEntity == Entity::none() || fo.depth < bf.depth || (fo.depth == bf.depth && agent_row < best_row[row])) {
                        bf = fo;
                        best_row[row] = agent_row;
                    }
                }
            }
        }
        if (!any) {                                             /* no agent instance: every ray misses */
            bvhParams.instanceCounts = zeros.data();
            build_tlas(w, aabbs, 0);
            w.internal().traversalNodes = w.nodes.data();
            launch();
            for (int32_t c = 0; c < nviews; ++c)
                memcpy(&best_depth[(size_t)views[c].rowIDX * 32], &scratch[(size_t)views[c].rowIDX * 32], 32);
        }
        for (int32_t c = 0; c < nviews; ++c) {
            const uint32_t row = views[c].rowIDX;
            memcpy(depth + (size_t)row * 32, &best_depth[(size_t)row * 32], 32);
            depth[(size_t)row * 32] = (uint8_t)row;              /* the pixel-0 stamp is the same in every pass */
            w.finderByRow[row] = best_finder[row];
        }
    }
    /* the finder record is written at rowIDX and read back through the camera entity's row (render/ecs.inl:5-26) */
    auto *finder = reinterpret_cast<uint8_t *>(sm.getArchetypeComponent<mr::RaycastOutputArchetype, mr::FinderOutputBuffer>());
    for (int32_t c = 0; c < cams.numRows; ++c)
        memcpy(finder + sizeof(mr::FinderOutput) * (size_t)c, &w.finderByRow[views[c].rowIDX], sizeof(mr::FinderOutput));
}

} // namespace

extern "C" {

void *mbr_create(uint32_t seed, uint32_t world_id, uint32_t init_agents, uint32_t flags)
{
    RefWorld *w = new RefWorld();
    w->flags = flags;
    mwGPU::mockState() = &w->sm;
    /* Manager::Impl::make, src/entry/mgr.cpp:104-113,132-133 */
    w->cfg.initRandKey = rand::initKey(seed);
    w->cfg.numChunksX = 8;
    w->cfg.numChunksY = 6;
    w->cfg.cellDim = 1.f;
    w->cfg.renderBridge = nullptr;
    w->cfg.simBridge = &w->bridge;
    w->cfg.totalAllowedFood = 30;
    w->cfg.initNumAgentsPerWorld = init_agents;
    ECSRegistry reg(&w->sm);
    Sim::registerTypes(reg, w->cfg);
    Sim::setupTasks(w->graphs, w->cfg);
    if (flags & MBR_FIX_U6) {
        StateManager *sm = &w->sm;
        const uint32_t agent_arch = sm->archetypeId<Agent>();
        sm->onDestroy = [sm, agent_arch](Entity, Loc l) {
            if (l.archetype == agent_arch) sm->destroyEntityNow(sm->getUnsafe<AgentObservationBridge>(l).obsEntity);
        };
    }
    w->simStorage = aligned_alloc(64, (sizeof(Sim) + 63) & ~(size_t)63);
    w->ctx.reset(new Engine(static_cast<WorldBase *>(static_cast<Sim *>(w->simStorage)), &w->sm, (int32_t)world_id));
    Sim::WorldInit wi;
    new (w->simStorage) Sim(*w->ctx, w->cfg, wi);
    run_graph(*w, TaskGraphID::Init);
    {   /* before the first sensor pass the finder records are uninitialised memory in the reference; here: no hit */
        auto *finder = reinterpret_cast<uint8_t *>(w->sm.getArchetypeComponent<mr::RaycastOutputArchetype, mr::FinderOutputBuffer>());
        const mr::FinderOutput none{Entity::none(), 255.f};
        for (int32_t r = 0; r < w->sm.table<mr::RaycastOutputArchetype>().numRows; ++r)
            memcpy(finder + sizeof(mr::FinderOutput) * (size_t)r, &none, sizeof none);
    }
    return w;
}

void mbr_destroy(void *h)
{
    RefWorld *w = static_cast<RefWorld *>(h);
    free(w->simStorage);
    delete w;
}

/* Manager::Impl::step, src/entry/mgr.cpp:51-63: step graph, then sensor graph */
void mbr_step(void *h)
{
    RefWorld *w = static_cast<RefWorld *>(h);
    mwGPU::mockState() = &w->sm;
    const int32_t first_new = (int32_t)w->sm.ents.size();
    run_graph(*w, TaskGraphID::Step);
    if (w->flags & MBR_FIX_D3) {
        StateManager &sm = w->sm;
        MockTable &agents = sm.table<Agent>();
        auto *pd = reinterpret_cast<uint8_t *>(sm.getArchetypeComponent<mr::RaycastOutputArchetype, mr::PrevDepthOutputBuffer>());
        auto *ps = reinterpret_cast<uint8_t *>(sm.getArchetypeComponent<mr::RaycastOutputArchetype, mr::PrevSemanticOutputBuffer>());
        for (int32_t r = 0; r < agents.numRows; ++r) {
            if (agents.rowEntity[r].id < first_new) continue;
            const uint32_t row = sm.getUnsafe<SensorOutputIndex>(agents.rowEntity[r]).idx;
            memset(pd + 32 * (size_t)row, 0, 32);
            memset(ps + 32 * (size_t)row, 0, 32);
        }
    }
    run_sensor(*w);
}

void mbr_shift(void *h)
{
    RefWorld *w = static_cast<RefWorld *>(h);
    mwGPU::mockState() = &w->sm;
    run_graph(*w, TaskGraphID::ShiftObservations);
}

int32_t mbr_num_agents(void *h) { return static_cast<RefWorld *>(h)->sm.table<Agent>().numRows; }
int32_t mbr_num_obs_rows(void *h) { return static_cast<RefWorld *>(h)->sm.table<AgentObservationArchetype>().numRows; }

/* write the (rows, 6) int32 action matrix / (rows, 16) hidden matrix into the observation table (what Python does
 * through action_tensor / hidden_state_tensor, learn/training_loop.py:136-137) */
void mbr_set_actions(void *h, const int32_t *a, int32_t rows)
{
    RefWorld *w = static_cast<RefWorld *>(h);
    Action *col = w->sm.getArchetypeComponent<AgentObservationArchetype, Action>();
    memcpy(col, a, sizeof(Action) * (size_t)rows);
}
void mbr_set_hidden(void *h, const float *v, int32_t rows)
{
    RefWorld *w = static_cast<RefWorld *>(h);
    HiddenState *col = w->sm.getArchetypeComponent<AgentObservationArchetype, HiddenState>();
    memcpy(col, v, sizeof(HiddenState) * (size_t)rows);
}

/* Agent-table order. out: (n, 8) float = x, y, z, quat w, x, y, z, 0 ; ints: (n, 4) = health, species, finder target
 * (row of the hit agent in the Agent table, -1 none), observation row */
int32_t mbr_agents(void *h, float *pose8, int32_t *ints4)
{
    RefWorld *w = static_cast<RefWorld *>(h);
    StateManager &sm = w->sm;
    MockTable &t = sm.table<Agent>();
    const uint32_t agent_arch = sm.archetypeId<Agent>();
    auto *finder = reinterpret_cast<uint8_t *>(sm.getArchetypeComponent<mr::RaycastOutputArchetype, mr::FinderOutputBuffer>());
    for (int32_t r = 0; r < t.numRows; ++r) {
        const Loc l{agent_arch, r};
        const auto &p = sm.getUnsafe<base::Position>(l);
        const auto &q = sm.getUnsafe<base::Rotation>(l);
        float *o = pose8 + 8 * (size_t)r;
        o[0] = p.x; o[1] = p.y; o[2] = p.z; o[3] = q.w; o[4] = q.x; o[5] = q.y; o[6] = q.z; o[7] = 0.f;
        int32_t *i4 = ints4 + 4 * (size_t)r;
        i4[0] = sm.getUnsafe<Health>(l).v;
        i4[1] = (int32_t)sm.getUnsafe<Species>(l).speciesID;
        const mr::RenderCamera &cam = sm.getUnsafe<mr::RenderCamera>(l);
        mr::FinderOutput fo;
        memcpy(&fo, finder + sizeof(mr::FinderOutput) * (size_t)sm.getLoc(cam.cameraEntity).row, sizeof fo);
        i4[2] = fo.hitEntity == Entity::none() ? -1 : sm.getLoc(fo.hitEntity).row;
        i4[3] = sm.getLoc(sm.getUnsafe<AgentObservationBridge>(l).obsEntity).row;
    }
    return t.numRows;
}

/* Observation-table order, `rows` rows. `which`: 0 current, 1 prev. Layout of one row of out_f (32 floats):
 * pos.x pos.y health surr.x surr.y reward species hidden[16] ; out_i (10 ints): action[6] stats[4] ;
 * out_b (64 bytes): depth[32] semantic[32] */
int32_t mbr_obs(void *h, int32_t which, float *out_f, int32_t *out_i, uint8_t *out_b)
{
    RefWorld *w = static_cast<RefWorld *>(h);
    StateManager &sm = w->sm;
    MockTable &t = sm.table<AgentObservationArchetype>();
    const uint32_t arch = sm.archetypeId<AgentObservationArchetype>();
    auto *depth = reinterpret_cast<uint8_t *>(sm.getArchetypeComponent<mr::RaycastOutputArchetype, mr::DepthOutputBuffer>());
    auto *sem = reinterpret_cast<uint8_t *>(sm.getArchetypeComponent<mr::RaycastOutputArchetype, mr::SemanticOutputBuffer>());
    auto *pdepth = reinterpret_cast<uint8_t *>(sm.getArchetypeComponent<mr::RaycastOutputArchetype, mr::PrevDepthOutputBuffer>());
    auto *psem = reinterpret_cast<uint8_t *>(sm.getArchetypeComponent<mr::RaycastOutputArchetype, mr::PrevSemanticOutputBuffer>());
    for (int32_t r = 0; r < t.numRows; ++r) {
        const Loc l{arch, r};
        float *f = out_f + 32 * (size_t)r;
        int32_t *iv = out_i + 10 * (size_t)r;
        uint8_t *b = out_b + 64 * (size_t)r;
        if (which == 0) {
            const auto &p = sm.getUnsafe<PositionObservation>(l);
            const auto &s = sm.getUnsafe<SurroundingObservation>(l);
            f[0] = p.pos.x; f[1] = p.pos.y; f[2] = (float)sm.getUnsafe<HealthObservation>(l).v;
            f[3] = s.presenceHeuristic; f[4] = s.movementHeuristic; f[5] = sm.getUnsafe<Reward>(l).v;
            f[6] = (float)sm.getUnsafe<SpeciesObservation>(l).speciesID;
            memcpy(f + 7, sm.getUnsafe<HiddenState>(l).values, 64);
            memcpy(iv, &sm.getUnsafe<Action>(l), 24);
            memcpy(iv + 6, &sm.getUnsafe<StatsObservation>(l), 16);
            memcpy(b, depth + 32 * (size_t)r, 32); memcpy(b + 32, sem + 32 * (size_t)r, 32);
        } else {
            const auto &p = sm.getUnsafe<PrevPositionObservation>(l);
            const auto &s = sm.getUnsafe<PrevSurroundingObservation>(l);
            f[0] = p.pos.x; f[1] = p.pos.y; f[2] = (float)sm.getUnsafe<PrevHealthObservation>(l).v;
            f[3] = s.presenceHeuristic; f[4] = s.movementHeuristic; f[5] = sm.getUnsafe<PrevReward>(l).v;
            f[6] = (float)sm.getUnsafe<PrevSpeciesObservation>(l).speciesID;
            memcpy(f + 7, sm.getUnsafe<PrevHiddenState>(l).values, 64);
            memcpy(iv, &sm.getUnsafe<PrevAction>(l), 24);
            memcpy(iv + 6, &sm.getUnsafe<PrevStatsObservation>(l), 16);
            memcpy(b, pdepth + 32 * (size_t)r, 32); memcpy(b + 32, psem + 32 * (size_t)r, 32);
        }
        for (int k = 23; k < 32; ++k) f[k] = 0.f;
    }
    return t.numRows;
}

/* per world: out_i[0..3] SpeciesCount, [4] currentNumFood, [5] RNG draws consumed, [6] agent rows, [7] obs rows;
 * out_f[0..3] SpeciesReward; food: (48, 8) uint8 in the layout of MBOTS_DEBUG_FOOD (cells x<<4|y of packages 0-4, mask);
 * chunks: (48, 2) uint32 numAgents, totalSpeed */
void mbr_world(void *h, int32_t *out_i, float *out_f, uint8_t *food, uint32_t *chunks)
{
    RefWorld *w = static_cast<RefWorld *>(h);
    StateManager &sm = w->sm;
    Sim &sim = w->sim();
    const SpeciesCount &sc = sm.getUnsafe<SpeciesCount>(sim.speciesInfoTracker);
    const SpeciesReward &sr = sm.getUnsafe<SpeciesReward>(sim.speciesInfoTracker);
    for (int k = 0; k < 4; ++k) { out_i[k] = (int32_t)sc.counts[k]; out_f[k] = sr.rewards[k]; }
    out_i[4] = sim.currentNumFood.load_relaxed();
    struct RngBits { RandKey k; uint32_t count; } rb;       /* RNG's private members, M/include/madrona/rand.hpp:50-53 */
    static_assert(sizeof(RNG) == sizeof(RngBits), "RNG layout");
    memcpy(&rb, &sim.rng, sizeof rb);
    out_i[5] = (int32_t)rb.count;
    out_i[6] = sm.table<Agent>().numRows;
    out_i[7] = sm.table<AgentObservationArchetype>().numRows;
    for (int c = 0; c < 48; ++c) {
        Loc l = sim.chunksLoc;
        l.row += c;
        ChunkInfo &ci = sm.getUnsafe<ChunkInfo>(l);
        uint8_t *f = food + 8 * (size_t)c;
        memset(f, 0, 8);
        for (int p = 0; p < 5; ++p) {
            const bool has = ci.foodPackages[p].numFood.load_relaxed() > 0;
            f[p] = has ? (uint8_t)((ci.foodPackages[p].x << 4) | ci.foodPackages[p].y) : 0;
            if (has) f[5] |= (uint8_t)(1u << p);
        }
        chunks[2 * c] = ci.numAgents.load_relaxed();
        chunks[2 * c + 1] = ci.totalSpeed.load_relaxed();
    }
}

/* ---- free functions of the reference that need no world (VERDICT: "start with the functions that need no mock") */
int32_t mbr_ray_sphere(const float *centre3, float radius, const float *origin3, const float *dir3, float *t_out)
{
    const SphereIntersection s = raySphereIntersect(math::Vector3{centre3[0], centre3[1], centre3[2]}, radius,
                                                    math::Vector3{origin3[0], origin3[1], origin3[2]},
                                                    math::Vector3{dir3[0], dir3[1], dir3[2]});
    *t_out = s.tHitClosest;
    return s.intersects ? 1 : 0;
}

} /* extern "C" */
