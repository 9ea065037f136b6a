// TEST INFRASTRUCTURE (oracle) -- not part of the product path.
//
// "mini-madrona": the smallest stand-in for the un-vendored Madrona engine
// (CMakeLists.txt:4-9 of the reference pulls it from MADRONA_EXTERNAL_DIR)
// that lets the reference's own sim.cpp / utils.cpp / level_gen.cpp compile
// unmodified from /root/reference/src.  Only what those three files touch is
// provided (SURVEY.md section 8c lists the surface).
//
// Executor semantics assumed (the real CPU executor is not available here;
// stated in DESIGN.md): task-graph nodes run in insertion order per world,
// ParallelForNode visits archetype rows in creation order, tables are
// zero-filled, destroyEntity is applied at the next CompactArchetypeNode.
#pragma once

#include <cstdint>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <cassert>
#include <cfloat>
#include <cmath>
#include <math.h>
#include <array>
#include <bit>
#include <vector>
#include <tuple>
#include <type_traits>
#include <initializer_list>
#include <functional>
#include <memory>
#include <utility>
#include <string>
#include <algorithm>

#include <brt/types.hpp>
#include <brt/math.hpp>
#include <brt/rand.hpp>

#define MADRONA_UNREACHABLE() __builtin_unreachable()
#define MADRONA_UNROLL
#define MADRONA_ALWAYS_INLINE __attribute__((always_inline))
#define MADRONA_BUILD_MWGPU_ENTRY(...)

using std::signbit;

namespace madrona {

using u64 = uint64_t; using i64 = int64_t;
using u32 = uint32_t; using i32 = int32_t;
using u16 = uint16_t; using i16 = int16_t;
using u8 = uint8_t;   using i8 = int8_t;
using f32 = float;
using CountT = int32_t;

using brt::operator""_u32;
using brt::operator""_u64;
using brt::operator""_i32;
using brt::operator""_i64;

// ---------------------------------------------------------------- math / rand
namespace math {
using namespace brt;
constexpr inline Vector3 up { 0, 0, 1 };
constexpr inline Vector3 fwd { 0, 1, 0 };
constexpr inline Vector3 right { 1, 0, 0 };
}

using RandKey = brt::RandKey;
using RNG = brt::RNG;
namespace rand = brt::rand;

// ---------------------------------------------------------------- ids
struct Entity {
    uint32_t gen;
    int32_t id;
    static constexpr Entity none() { return Entity { 0xFFFFFFFFu, -1 }; }
    friend bool operator==(Entity a, Entity b)
    { return a.gen == b.gen && a.id == b.id; }
    friend bool operator!=(Entity a, Entity b) { return !(a == b); }
};

struct Loc {
    uint32_t archetype;
    int32_t row;
};

struct WorldID { int32_t idx; };

template <typename... Ts> struct Archetype {
    using Components = std::tuple<Ts...>;
};

using brt::Span;

inline void * rawAlloc(size_t n) { return calloc(1, n ? n : 1); }
inline void rawDealloc(void *p) { free(p); }

// ---------------------------------------------------------------- sync
namespace sync {
enum memory_order { relaxed, acquire, release, acq_rel, seq_cst };
}

class AtomicU32Ref {
public:
    AtomicU32Ref(uint32_t &r) : r_(&r) {}
    uint32_t fetch_add_relaxed(uint32_t v) { uint32_t o = *r_; *r_ += v; return o; }
    template <sync::memory_order> void store(uint32_t v) { *r_ = v; }
    template <sync::memory_order> uint32_t load() const { return *r_; }
    template <sync::memory_order> uint32_t fetch_or(uint32_t v)
    { uint32_t o = *r_; *r_ |= v; return o; }
private:
    uint32_t *r_;
};

template <typename T> class AtomicRef {
public:
    AtomicRef(T &r) : r_(&r) {}
    T fetch_add_relaxed(T v) { T o = *r_; *r_ += v; return o; }
private:
    T *r_;
};

// ---------------------------------------------------------------- type ids
namespace shim {

uint32_t allocTypeID();

template <typename T> struct TypeID {
    static uint32_t get() { static uint32_t id = allocTypeID(); return id; }
};

struct ColumnDesc {
    uint32_t typeID;
    uint32_t numBytes;
};

struct ArchetypeDesc {
    std::vector<ColumnDesc> columns; // component columns, in archetype order
    bool registered = false;
};

struct Table {
    std::vector<char *> cols;        // one zero-filled buffer per component
    std::vector<uint32_t> colBytes;
    std::vector<Entity> entities;    // row -> entity handle
    std::vector<uint8_t> dead;       // row marked by destroyEntity
    int32_t numRows = 0;
    int32_t capacity = 0;
};

struct EntitySlot {
    uint32_t gen;
    uint32_t archetype;
    int32_t row;     // -1 when free
};

struct ExportDesc {
    bool valid = false;
    bool singleton = false;
    uint32_t archetype = 0;
    uint32_t typeID = 0;
    uint32_t numBytes = 0;
};

struct GlobalRegistry {
    std::vector<ArchetypeDesc> archetypes;     // indexed by archetype TypeID
    std::vector<uint32_t> componentBytes;       // indexed by component TypeID
    std::vector<uint32_t> singletonBytes;       // indexed by TypeID (0 = none)
    std::vector<ExportDesc> exports;
};

GlobalRegistry & globalRegistry();

struct WorldState {
    int32_t worldIdx;
    std::vector<Table> tables;           // indexed by archetype TypeID
    std::vector<char *> singletons;      // indexed by TypeID
    std::vector<EntitySlot> slots;
    std::vector<int32_t> freeSlots;
    void *userData = nullptr;            // Sim *

    void init(int32_t world_idx);
    Table & table(uint32_t arch);
    int colIndex(uint32_t arch, uint32_t type_id) const;
    Entity makeEntity(uint32_t arch);
    Loc makeTemporary(uint32_t arch);
    void destroyEntity(Entity e);
    void compact(uint32_t arch);
    void clearTable(uint32_t arch);
    bool entityValid(Entity e) const;
};

template <typename Tuple> struct TupleColumns;
template <typename... Ts> struct TupleColumns<std::tuple<Ts...>> {
    static void fill(std::vector<ColumnDesc> &out) {
        (out.push_back(ColumnDesc { TypeID<Ts>::get(), (uint32_t)sizeof(Ts) }), ...);
    }
};

template <typename T>
inline void ensureSize(std::vector<T> &v, size_t n) { if (v.size() < n) v.resize(n); }

} // namespace shim

// ---------------------------------------------------------------- registry
class ECSRegistry {
public:
    template <typename T> void registerComponent()
    {
        auto &g = shim::globalRegistry();
        uint32_t id = shim::TypeID<T>::get();
        shim::ensureSize(g.componentBytes, id + 1);
        g.componentBytes[id] = sizeof(T);
    }

    template <typename A> void registerArchetype()
    {
        auto &g = shim::globalRegistry();
        uint32_t id = shim::TypeID<A>::get();
        shim::ensureSize(g.archetypes, id + 1);
        auto &d = g.archetypes[id];
        if (d.registered) return;
        shim::TupleColumns<typename A::Components>::fill(d.columns);
        d.registered = true;
    }

    template <typename T> void registerSingleton()
    {
        auto &g = shim::globalRegistry();
        uint32_t id = shim::TypeID<T>::get();
        shim::ensureSize(g.singletonBytes, id + 1);
        g.singletonBytes[id] = sizeof(T);
    }

    template <typename A, typename C, typename SlotT> void exportColumn(SlotT slot)
    {
        auto &g = shim::globalRegistry();
        shim::ensureSize(g.exports, (size_t)slot + 1);
        auto &e = g.exports[(size_t)slot];
        e.valid = true; e.singleton = false;
        e.archetype = shim::TypeID<A>::get();
        e.typeID = shim::TypeID<C>::get();
        e.numBytes = sizeof(C);
    }

    template <typename T, typename SlotT> void exportSingleton(SlotT slot)
    {
        auto &g = shim::globalRegistry();
        shim::ensureSize(g.exports, (size_t)slot + 1);
        auto &e = g.exports[(size_t)slot];
        e.valid = true; e.singleton = true;
        e.typeID = shim::TypeID<T>::get();
        e.numBytes = sizeof(T);
    }
};

// ---------------------------------------------------------------- context
template <typename T> class ResultRef {
public:
    ResultRef(T *p) : p_(p) {}
    ResultRef(std::nullptr_t) : p_(nullptr) {}
    bool valid() const { return p_ != nullptr; }
    T & value() { return *p_; }
    const T & value() const { return *p_; }
private:
    T *p_;
};

class Context {
public:
    Context(shim::WorldState *w) : w_(w) {}

    WorldID worldID() const { return WorldID { w_->worldIdx }; }

    template <typename A> Entity makeEntity()
    { return w_->makeEntity(shim::TypeID<A>::get()); }

    template <typename A> Loc makeTemporary()
    { return w_->makeTemporary(shim::TypeID<A>::get()); }

    void destroyEntity(Entity e) { w_->destroyEntity(e); }

    template <typename T> T & get(Entity e)
    {
        assert(w_->entityValid(e));
        const shim::EntitySlot &s = w_->slots[e.id];
        int col = w_->colIndex(s.archetype, shim::TypeID<T>::get());
        assert(col >= 0);
        return ((T *)w_->tables[s.archetype].cols[col])[s.row];
    }

    template <typename T> ResultRef<T> getSafe(Entity e)
    {
        if (!w_->entityValid(e)) return ResultRef<T>(nullptr);
        const shim::EntitySlot &s = w_->slots[e.id];
        int col = w_->colIndex(s.archetype, shim::TypeID<T>::get());
        if (col < 0) return ResultRef<T>(nullptr);
        return ResultRef<T>(&((T *)w_->tables[s.archetype].cols[col])[s.row]);
    }

    template <typename T> ResultRef<T> getCheck(Entity e) { return getSafe<T>(e); }

    // col_idx counts the implicit entity / world-id columns first (madrona
    // user columns start at 2).
    template <typename T> T & getDirect(int col_idx, Loc loc)
    {
        return ((T *)w_->tables[loc.archetype].cols[col_idx - 2])[loc.row];
    }

    template <typename T> T & singleton()
    {
        uint32_t id = shim::TypeID<T>::get();
        assert(id < w_->singletons.size() && w_->singletons[id]);
        return *(T *)w_->singletons[id];
    }

    shim::WorldState * shimWorld() { return w_; }

protected:
    shim::WorldState *w_;
};

struct WorldBase {
    WorldBase(Context &) {}
};

template <typename ContextT, typename DataT>
class CustomContext : public Context {
public:
    using Context::Context;
    DataT & data() { return *(DataT *)w_->userData; }
};

// ---------------------------------------------------------------- task graph
struct TaskGraphNodeID { int32_t id; };

namespace shim {

struct GraphNode {
    std::function<void(WorldState &)> run;
};

template <typename T> struct IsEntity : std::false_type {};
template <> struct IsEntity<Entity> : std::true_type {};

// Does archetype `arch` hold every component of the query?
template <typename... Cs>
inline bool archetypeMatches(const WorldState &w, uint32_t arch)
{
    bool ok = true;
    ((ok = ok && (IsEntity<Cs>::value ||
                  w.colIndex(arch, TypeID<Cs>::get()) >= 0)), ...);
    return ok;
}

template <typename C>
inline C * columnPtr(WorldState &w, uint32_t arch)
{
    Table &t = w.tables[arch];
    if constexpr (IsEntity<C>::value) {
        return t.entities.data();
    } else {
        return (C *)t.cols[w.colIndex(arch, TypeID<C>::get())];
    }
}

} // namespace shim

template <typename ContextT, auto Fn, typename... Cs>
struct ParallelForNode {
    static void run(shim::WorldState &w)
    {
        auto &g = shim::globalRegistry();
        ContextT ctx(&w);
        for (uint32_t arch = 0; arch < (uint32_t)g.archetypes.size(); arch++) {
            if (!g.archetypes[arch].registered) continue;
            if (arch >= w.tables.size()) continue;
            if (!shim::archetypeMatches<Cs...>(w, arch)) continue;
            shim::Table &t = w.tables[arch];
            int32_t n = t.numRows;
            std::tuple<Cs *...> ptrs { shim::columnPtr<Cs>(w, arch)... };
            for (int32_t i = 0; i < n; i++) {
                if (t.dead[i]) continue;
                std::apply([&](auto *...p) { Fn(ctx, p[i]...); }, ptrs);
            }
        }
    }
};

// Singleton queries: madrona exposes singletons as one-row archetypes, so
// ParallelForNode<Engine, fn, SingletonT> runs once per world.
template <typename ContextT, auto Fn, typename S>
struct SingletonForNode {
    static void run(shim::WorldState &w)
    {
        ContextT ctx(&w);
        Fn(ctx, ctx.template singleton<S>());
    }
};

template <typename A> struct CompactArchetypeNode {
    static void run(shim::WorldState &w) { w.compact(shim::TypeID<A>::get()); }
};
template <typename A> struct ClearTmpNode {
    static void run(shim::WorldState &w) { w.clearTable(shim::TypeID<A>::get()); }
};
struct ResetTmpAllocNode { static void run(shim::WorldState &) {} };

class TaskGraphBuilder {
public:
    template <typename NodeT>
    TaskGraphNodeID addToGraph(Span<const TaskGraphNodeID>)
    {
        nodes_.push_back(shim::GraphNode { dispatch<NodeT>() });
        return TaskGraphNodeID { (int32_t)nodes_.size() - 1 };
    }

    void run(shim::WorldState &w) { for (auto &n : nodes_) n.run(w); }
    size_t numNodes() const { return nodes_.size(); }

private:
    template <typename NodeT> struct Dispatch {
        static std::function<void(shim::WorldState &)> make()
        { return &NodeT::run; }
    };

    // A ParallelForNode over exactly one component that is a registered
    // singleton (and belongs to no archetype) iterates the singleton.
    template <typename ContextT, auto Fn, typename C>
    struct Dispatch<ParallelForNode<ContextT, Fn, C>> {
        static std::function<void(shim::WorldState &)> make()
        {
            return [](shim::WorldState &w) {
                auto &g = shim::globalRegistry();
                uint32_t id = shim::TypeID<C>::get();
                if (id < g.singletonBytes.size() && g.singletonBytes[id] != 0) {
                    SingletonForNode<ContextT, Fn, C>::run(w);
                } else {
                    ParallelForNode<ContextT, Fn, C>::run(w);
                }
            };
        }
    };

    template <typename NodeT>
    static std::function<void(shim::WorldState &)> dispatch()
    { return Dispatch<NodeT>::make(); }

    std::vector<shim::GraphNode> nodes_;
};

class TaskGraphManager {
public:
    template <typename IDT> TaskGraphBuilder & init(IDT id)
    {
        shim::ensureSize(graphs_, (size_t)id + 1);
        return graphs_[(size_t)id];
    }
    TaskGraphBuilder & graph(size_t i) { return graphs_[i]; }
private:
    std::vector<TaskGraphBuilder> graphs_;
};

// ---------------------------------------------------------------- components
namespace base {
struct Position : math::Vector3 {
    Position() = default;
    Position(math::Vector3 v) : Vector3(v) {}
};
struct Rotation : math::Quat {
    Rotation() = default;
    Rotation(math::Quat q) : Quat(q) {}
};
struct Scale : math::Diag3x3 {
    Scale() = default;
    Scale(math::Diag3x3 d) : Diag3x3(d) {}
};
struct ObjectID { int32_t idx; };
inline void registerTypes(ECSRegistry &r)
{
    r.registerComponent<Position>();
    r.registerComponent<Rotation>();
    r.registerComponent<Scale>();
    r.registerComponent<ObjectID>();
}
}

namespace phys {
struct ExternalForce : math::Vector3 {
    ExternalForce() = default;
    ExternalForce(math::Vector3 v) : Vector3(v) {}
};
struct ExternalTorque : math::Vector3 {
    ExternalTorque() = default;
    ExternalTorque(math::Vector3 v) : Vector3(v) {}
};
struct Velocity { math::Vector3 linear; math::Vector3 angular; };
struct PhysicsSystem {
    static void registerTypes(ECSRegistry &r)
    {
        r.registerComponent<ExternalForce>();
        r.registerComponent<ExternalTorque>();
        r.registerComponent<Velocity>();
    }
};
}

// ---------------------------------------------------------------- navmesh
// POD stand-in. samplePoint is only reachable with SimFlags::NavmeshSpawn or
// Task::Turret, neither of which is inside the configs.
struct Navmesh {
    math::Vector3 *vertices;
    uint32_t *triIndices;
    uint32_t *triAdjacency;
    float *triSampleCDF;
    uint32_t numVerts;
    uint32_t numTris;

    math::Vector3 samplePoint(RandKey) const { abort(); }
    void getTriangleVertices(uint32_t tri, math::Vector3 *a, math::Vector3 *b,
                             math::Vector3 *c) const
    {
        *a = vertices[triIndices[3 * tri]];
        *b = vertices[triIndices[3 * tri + 1]];
        *c = vertices[triIndices[3 * tri + 2]];
    }
};

} // namespace madrona
