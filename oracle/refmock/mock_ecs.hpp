/* mock_ecs.hpp -- a host stand-in for the slice of Madrona's MWCUDA runtime that the reference's
 * src/sim/sim.cpp and external/madrona/src/render/ecs_system.cpp are written against.
 *
 * TEST INFRASTRUCTURE (oracle/): nothing under madrona-bots_b200/ includes or links this.
 *
 * Purpose (VERDICT r1 "What's missing" #3): the reference's step cannot be built (SURVEY.md section 0
 * R1/R2), but its SYSTEM BODIES are plain C++ once `Engine`, the ECS tables and the task-graph
 * builder exist. This header provides exactly that, so that oracle/refmock/refsim.cpp can
 * `#include` the reference's own source files from where they lie under /root/reference and run
 * one world through the reference's own task graphs (setupStepTasks / RenderingSystem::setupTasks /
 * setupShiftObservationsTasks), in the reference's own node order. Nothing of the reference is
 * copied here; what IS restated is runtime semantics, each with the file:line it follows
 * (M/ = external/madrona/):
 *
 *   tables              one table per archetype, column per component, rows appended by
 *                       makeEntityNow (M/src/mw/device/state.cpp:488-527); addresses are stable
 *                       (the reference maps virtual memory; here columns are allocated at capacity)
 *   destroyEntity       marks the row (WorldID = -1), data stays readable until the next sort
 *                       (state.cpp:583-601); ctx.get() does not check generations in release
 *                       builds (M/src/mw/device/include/madrona/state.inl:287-307)
 *   ParallelForNode     invocation count taken at node start, deleted rows skipped
 *                       (M/src/mw/device/include/madrona/taskgraph.inl:188-230,272-281); rows run
 *                       sequentially in row order = the canonical order of SURVEY U11
 *   SortArchetypeNode   stable; by WorldID = compaction of deleted rows (one world per mock), by a
 *                       component = stable sort on its first 32-bit word + compaction
 *                       (M/src/mw/device/sort_archetype.cpp:1259-1350,1443-1454)
 *   singletons          one-row archetypes (M/include/madrona/registry.inl)
 *
 * One MockWorld = one world of the reference (worlds are independent; the global species sort
 * only interleaves worlds, which the canonical (species, world, slot) export order undoes).
 */
#pragma once
#include <cassert>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <typeindex>
#include <typeinfo>
#include <unordered_map>
#include <vector>
#include <algorithm>
#include <type_traits>

#include <madrona/ecs.hpp>      /* the reference's own Entity / Loc / Archetype<> / WorldID / WorldBase */
#include <madrona/sync.hpp>
#include <madrona/math.hpp>
#include <madrona/span.hpp>       /* the reference's own Span */

namespace madrona {

struct TaskGraphNodeID { uint32_t id; };

constexpr int32_t kMockRows = 8192;          /* capacity of every table (rows); asserts on overflow */

struct MockColumn {
    size_t elem = 0;
    uint8_t *data = nullptr;
};

struct MockTable {
    std::vector<std::type_index> comps;
    std::unordered_map<std::type_index, MockColumn> cols;
    std::vector<Entity> rowEntity;
    std::vector<uint8_t> dead;
    int32_t numRows = 0;
    int32_t worldOffset = 0, worldCount = 0;       /* one world: offset 0, count = rows after the last sort */
    bool has(std::type_index t) const { return cols.find(t) != cols.end(); }
};

struct MockEntitySlot { Loc loc; uint32_t gen; bool live; };

template <typename A> struct ArchetypeComponents;
template <typename... Ts> struct ArchetypeComponents<Archetype<Ts...>> {
    static void each(const std::function<void(std::type_index, size_t)> &f) { (f(std::type_index(typeid(Ts)), sizeof(Ts)), ...); }
};

class StateManager {
public:
    std::vector<MockTable *> tables;                              /* index = Loc::archetype */
    std::unordered_map<std::type_index, uint32_t> archetypeIds;
    std::unordered_map<std::type_index, size_t> sizeOverride;     /* registerComponent<T>(bytes) */
    std::vector<MockEntitySlot> ents;
    /* interventions the test harness can switch (each is a SURVEY U-rule, see refsim.cpp) */
    std::function<void(Entity, Loc)> onDestroy;

    ~StateManager()
    {
        for (MockTable *t : tables) {
            for (auto &kv : t->cols) free(kv.second.data);
            delete t;
        }
    }

    template <typename A> uint32_t archetypeId()
    {
        const std::type_index key(typeid(A));
        auto it = archetypeIds.find(key);
        if (it != archetypeIds.end()) return it->second;
        MockTable *t = new MockTable();
        ArchetypeComponents<typename A::Base>::each([&](std::type_index ti, size_t sz) {
            auto ov = sizeOverride.find(ti);
            if (ov != sizeOverride.end()) sz = ov->second;
            MockColumn c;
            c.elem = sz;
            c.data = (uint8_t *)aligned_alloc(64, ((sz * (size_t)kMockRows + 256) + 63) & ~(size_t)63);
            memset(c.data, 0, sz * (size_t)kMockRows + 256);
            t->comps.push_back(ti);
            t->cols.emplace(ti, c);
        });
        t->rowEntity.assign(kMockRows, Entity::none());
        t->dead.assign(kMockRows, 0);
        const uint32_t id = (uint32_t)tables.size();
        tables.push_back(t);
        archetypeIds.emplace(key, id);
        return id;
    }
    template <typename A> MockTable &table() { return *tables[archetypeId<A>()]; }

    int32_t appendRows(MockTable &t, int32_t count)
    {
        const int32_t row = t.numRows;
        assert(row + count <= kMockRows && "mock table capacity");
        t.numRows += count;
        for (auto &kv : t.cols)                                   /* a fresh record is zero (DEVIATIONS D3) */
            memset(kv.second.data + kv.second.elem * (size_t)row, 0, kv.second.elem * (size_t)count);
        for (int32_t r = row; r < row + count; ++r) { t.rowEntity[r] = Entity::none(); t.dead[r] = 0; }
        return row;
    }

    Entity makeEntityNow(uint32_t archetype)                      /* state.cpp:488-527 */
    {
        MockTable &t = *tables[archetype];
        const int32_t row = appendRows(t, 1);
        Entity e{0u, (int32_t)ents.size()};
        ents.push_back(MockEntitySlot{Loc{archetype, row}, 0u, true});
        t.rowEntity[row] = e;
        return e;
    }
    Loc makeStationary(uint32_t archetype, uint32_t count)        /* state.cpp:557-581 */
    {
        MockTable &t = *tables[archetype];
        return Loc{archetype, appendRows(t, (int32_t)count)};
    }
    void destroyEntityNow(Entity e)                               /* state.cpp:583-601 */
    {
        MockEntitySlot &s = ents[e.id];
        if (!s.live) return;
        s.live = false;
        s.gen++;
        tables[s.loc.archetype]->dead[s.loc.row] = 1;
        if (onDestroy) onDestroy(e, s.loc);
    }
    Loc getLoc(Entity e) const { return ents[e.id].loc; }

    template <typename C> C &getUnsafe(Loc l)                     /* state.inl:296-307 */
    {
        MockTable &t = *tables[l.archetype];
        auto it = t.cols.find(std::type_index(typeid(C)));
        assert(it != t.cols.end() && "component not in archetype");
        return *reinterpret_cast<C *>(it->second.data + it->second.elem * (size_t)l.row);
    }
    template <typename C> C &getUnsafe(Entity e) { return getUnsafe<C>(ents[e.id].loc); }

    template <typename A, typename C> C *getArchetypeComponent()
    {
        MockTable &t = table<A>();
        auto it = t.cols.find(std::type_index(typeid(C)));
        assert(it != t.cols.end());
        return reinterpret_cast<C *>(it->second.data);
    }
    template <typename A> uint32_t getArchetypeNumRows() { return (uint32_t)table<A>().numRows; }
    template <typename A> int32_t *getArchetypeWorldOffsets() { return &table<A>().worldOffset; }
    template <typename A> int32_t *getArchetypeWorldCounts() { return &table<A>().worldCount; }

    /* stable reorder of the live rows by `key(row)`; deleted rows drop out (sort_archetype.cpp:1259-1417) */
    void sortTable(MockTable &t, uint32_t archetype, const std::function<uint32_t(int32_t)> &key)
    {
        std::vector<int32_t> order;
        for (int32_t r = 0; r < t.numRows; ++r)
            if (!t.dead[r]) order.push_back(r);
        std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return key(a) < key(b); });
        for (auto &kv : t.cols) {
            MockColumn &c = kv.second;
            std::vector<uint8_t> tmp(c.elem * order.size());
            for (size_t i = 0; i < order.size(); ++i) memcpy(tmp.data() + c.elem * i, c.data + c.elem * (size_t)order[i], c.elem);
            memcpy(c.data, tmp.data(), tmp.size());
        }
        std::vector<Entity> ents_new(order.size());
        for (size_t i = 0; i < order.size(); ++i) ents_new[i] = t.rowEntity[order[i]];
        for (size_t i = 0; i < order.size(); ++i) {
            t.rowEntity[i] = ents_new[i];
            t.dead[i] = 0;
            if (ents_new[i].id >= 0 && ents_new[i] != Entity::none()) ents[ents_new[i].id].loc = Loc{archetype, (int32_t)i};
        }
        t.numRows = (int32_t)order.size();
        t.worldCount = t.numRows;
    }
};

namespace mwGPU {
inline StateManager *&mockState() { static thread_local StateManager *s = nullptr; return s; }
inline StateManager *getStateManager() { return mockState(); }
}

template <typename S> struct MockSingletonArchetype : Archetype<S> {};

class ECSRegistry {
public:
    explicit ECSRegistry(StateManager *s) : sm(s) {}
    template <typename C> void registerComponent(uint32_t bytes = 0)
    {
        if (bytes) sm->sizeOverride[std::type_index(typeid(C))] = bytes;
    }
    template <typename A> void registerArchetype() { (void)sm->archetypeId<A>(); }
    template <typename S> void registerSingleton()
    {
        const uint32_t id = sm->archetypeId<MockSingletonArchetype<S>>();
        if (sm->tables[id]->numRows == 0) sm->appendRows(*sm->tables[id], 1);     /* one row per world */
    }
    template <typename S> void exportSingleton(uint32_t) {}
    template <typename A, typename C> void exportColumn(uint32_t) {}
    StateManager *sm;
};

/* Context: the members of M/src/mw/device/include/madrona/context.inl:15-100 the path uses */
class Context {
public:
    Context(WorldBase *data, StateManager *sm, int32_t world) : data_(data), sm_(sm), world_id_{world} {}
    template <typename A> Entity makeEntity() { return sm_->makeEntityNow(sm_->archetypeId<A>()); }
    template <typename A> Loc makeStationary(uint32_t n) { return sm_->makeStationary(sm_->archetypeId<A>(), n); }
    void destroyEntity(Entity e) { sm_->destroyEntityNow(e); }
    Loc loc(Entity e) const { return sm_->getLoc(e); }
    template <typename C> C &get(Entity e) { return sm_->getUnsafe<C>(e); }
    template <typename C> C &get(Loc l) { return sm_->getUnsafe<C>(l); }
    template <typename S> S &singleton() { return sm_->getUnsafe<S>(Loc{sm_->archetypeId<MockSingletonArchetype<S>>(), 0}); }
    WorldID worldID() const { return world_id_; }
    WorldBase *data_;
    StateManager *sm_;
    WorldID world_id_;
};

template <typename ContextT, typename DataT> class CustomContext : public Context {
public:
    using Context::Context;
    DataT &data() { return *static_cast<DataT *>(data_); }
};

/* ---- task graphs ---------------------------------------------------------------------------- */
struct MockNode {
    const char *kind;                      /* "parallel_for", "sort", "reset_tmp", "recycle" */
    const void *fn;                        /* the system's function address (parallel_for) */
    std::function<void(Context &)> run;
};

class TaskGraphBuilder {
public:
    template <typename NodeT> TaskGraphNodeID addToGraph(Span<const TaskGraphNodeID>)
    {
        NodeT::add(*this);
        return TaskGraphNodeID{(uint32_t)nodes.size() - 1u};
    }
    std::vector<MockNode> nodes;
};
struct TaskGraph { using Builder = TaskGraphBuilder; using NodeID = TaskGraphNodeID; };

class TaskGraphManager {
public:
    template <typename EnumT> TaskGraphBuilder &init(EnumT id)
    {
        const size_t i = (size_t)id;
        if (graphs.size() <= i) graphs.resize(i + 1);
        return graphs[i];
    }
    std::vector<TaskGraphBuilder> graphs;
};

namespace mock_detail {
template <typename C> struct Arg {
    static C &get(MockTable &t, int32_t row)
    {
        MockColumn &c = t.cols.find(std::type_index(typeid(C)))->second;
        return *reinterpret_cast<C *>(c.data + c.elem * (size_t)row);
    }
};
template <> struct Arg<Entity> {
    static Entity &get(MockTable &t, int32_t row) { return t.rowEntity[row]; }
};
template <typename C> bool tableHas(const MockTable &t)
{
    if (std::is_same<C, Entity>::value) return true;
    return t.has(std::type_index(typeid(C)));
}
}

/* taskgraph.inl:188-230: one invocation per row of every archetype holding all the components */
template <typename ContextT, auto Fn, typename... ComponentTs> struct ParallelForNode {
    static void add(TaskGraphBuilder &b)
    {
        MockNode n;
        n.kind = "parallel_for";
        n.fn = reinterpret_cast<const void *>(+Fn);
        n.run = [](Context &base) {
            ContextT &ctx = static_cast<ContextT &>(base);
            StateManager *sm = base.sm_;
            const size_t ntables = sm->tables.size();
            for (size_t ti = 0; ti < ntables; ++ti) {
                MockTable &t = *sm->tables[ti];
                if (!(mock_detail::tableHas<ComponentTs>(t) && ...)) continue;
                const int32_t rows = t.numRows;                 /* invocation count taken at node start (:272-281) */
                for (int32_t r = 0; r < rows; ++r) {
                    if (t.dead[r]) continue;                    /* WorldID == -1 rows are skipped (:208-210) */
                    Fn(ctx, mock_detail::Arg<ComponentTs>::get(t, r)...);
                }
            }
        };
        b.nodes.push_back(std::move(n));
    }
};

template <typename ArchetypeT, typename ComponentT> struct SortArchetypeNode {
    static void add(TaskGraphBuilder &b)
    {
        MockNode n;
        n.kind = "sort";
        n.fn = nullptr;
        n.run = [](Context &base) {
            StateManager *sm = base.sm_;
            const uint32_t id = sm->archetypeId<ArchetypeT>();
            MockTable &t = *sm->tables[id];
            if (std::is_same<ComponentT, WorldID>::value) {
                sm->sortTable(t, id, [](int32_t) { return 0u; });
            } else {
                MockColumn &c = t.cols.find(std::type_index(typeid(ComponentT)))->second;
                sm->sortTable(t, id, [&](int32_t r) { uint32_t k; memcpy(&k, c.data + c.elem * (size_t)r, 4); return k; });
            }
        };
        b.nodes.push_back(std::move(n));
    }
};

struct ResetTmpAllocNode {
    static void add(TaskGraphBuilder &b) { b.nodes.push_back(MockNode{"reset_tmp", nullptr, [](Context &) {}}); }
};
struct RecycleEntitiesNode {
    static void add(TaskGraphBuilder &b) { b.nodes.push_back(MockNode{"recycle", nullptr, [](Context &) {}}); }
};

} // namespace madrona
