// TEST INFRASTRUCTURE (oracle shim): runtime of the mini-madrona ECS used to
// run the reference's own sim code.  See madrona/shim_core.hpp.
#include "madrona/shim_core.hpp"

namespace madrona::shim {

uint32_t allocTypeID()
{
    static uint32_t next = 0;
    return next++;
}

GlobalRegistry & globalRegistry()
{
    static GlobalRegistry g;
    return g;
}

static constexpr int32_t kInitialCapacity = 64;

void WorldState::init(int32_t world_idx)
{
    auto &g = globalRegistry();
    worldIdx = world_idx;

    tables.resize(g.archetypes.size());
    for (size_t a = 0; a < g.archetypes.size(); a++) {
        const ArchetypeDesc &d = g.archetypes[a];
        if (!d.registered) continue;
        Table &t = tables[a];
        t.capacity = kInitialCapacity;
        t.numRows = 0;
        t.cols.resize(d.columns.size());
        t.colBytes.resize(d.columns.size());
        for (size_t c = 0; c < d.columns.size(); c++) {
            t.colBytes[c] = d.columns[c].numBytes;
            t.cols[c] = (char *)calloc((size_t)t.capacity, d.columns[c].numBytes);
        }
        t.entities.assign((size_t)t.capacity, Entity::none());
        t.dead.assign((size_t)t.capacity, 0);
    }

    singletons.assign(g.singletonBytes.size(), nullptr);
    for (size_t s = 0; s < g.singletonBytes.size(); s++) {
        if (g.singletonBytes[s] != 0) {
            singletons[s] = (char *)calloc(1, g.singletonBytes[s]);
        }
    }
}

Table & WorldState::table(uint32_t arch) { return tables[arch]; }

int WorldState::colIndex(uint32_t arch, uint32_t type_id) const
{
    const auto &cols = globalRegistry().archetypes[arch].columns;
    for (size_t i = 0; i < cols.size(); i++) {
        if (cols[i].typeID == type_id) return (int)i;
    }
    return -1;
}

static int32_t appendRow(Table &t)
{
    if (t.numRows == t.capacity) {
        // Growing moves column storage.  The reference never holds a
        // reference into a table across a makeEntity on that same table.
        int32_t new_cap = t.capacity * 2;
        for (size_t c = 0; c < t.cols.size(); c++) {
            char *n = (char *)calloc((size_t)new_cap, t.colBytes[c]);
            memcpy(n, t.cols[c], (size_t)t.capacity * t.colBytes[c]);
            free(t.cols[c]);
            t.cols[c] = n;
        }
        t.entities.resize((size_t)new_cap, Entity::none());
        t.dead.resize((size_t)new_cap, 0);
        t.capacity = new_cap;
    }
    int32_t row = t.numRows++;
    for (size_t c = 0; c < t.cols.size(); c++) {
        memset(t.cols[c] + (size_t)row * t.colBytes[c], 0, t.colBytes[c]);
    }
    t.dead[row] = 0;
    return row;
}

Entity WorldState::makeEntity(uint32_t arch)
{
    Table &t = tables[arch];
    int32_t row = appendRow(t);

    int32_t slot_idx;
    if (!freeSlots.empty()) {
        slot_idx = freeSlots.back();
        freeSlots.pop_back();
    } else {
        slot_idx = (int32_t)slots.size();
        slots.push_back(EntitySlot { 0, 0, -1 });
    }

    EntitySlot &s = slots[slot_idx];
    s.archetype = arch;
    s.row = row;

    Entity e { s.gen, slot_idx };
    t.entities[row] = e;
    return e;
}

Loc WorldState::makeTemporary(uint32_t arch)
{
    Table &t = tables[arch];
    int32_t row = appendRow(t);
    t.entities[row] = Entity::none();
    return Loc { arch, row };
}

bool WorldState::entityValid(Entity e) const
{
    if (e.id < 0 || e.id >= (int32_t)slots.size()) return false;
    const EntitySlot &s = slots[e.id];
    return s.gen == e.gen && s.row >= 0;
}

void WorldState::destroyEntity(Entity e)
{
    if (!entityValid(e)) return;
    EntitySlot &s = slots[e.id];
    tables[s.archetype].dead[s.row] = 1;
    // The handle dies now (getCheck / getSafe fail from here on); the row is
    // reclaimed by the next CompactArchetypeNode.
    s.gen += 1;
    s.row = -1;
    freeSlots.push_back(e.id);
}

void WorldState::compact(uint32_t arch)
{
    Table &t = tables[arch];
    int32_t out = 0;
    for (int32_t i = 0; i < t.numRows; i++) {
        if (t.dead[i]) continue;
        if (out != i) {
            for (size_t c = 0; c < t.cols.size(); c++) {
                memmove(t.cols[c] + (size_t)out * t.colBytes[c],
                        t.cols[c] + (size_t)i * t.colBytes[c], t.colBytes[c]);
            }
            t.entities[out] = t.entities[i];
            Entity e = t.entities[out];
            if (e.id >= 0) slots[e.id].row = out;
        }
        t.dead[out] = 0;
        out++;
    }
    for (int32_t i = out; i < t.numRows; i++) t.dead[i] = 0;
    t.numRows = out;
}

void WorldState::clearTable(uint32_t arch)
{
    Table &t = tables[arch];
    t.numRows = 0;
}

}
