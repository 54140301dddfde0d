// Host-side world builder (see topology.h).  Behavioural contract = reference
// SimulatorCPP/ObjectCreator.cpp:193-641; the implementation is organised around "donor" tables
// (which already-visited neighbour can lend a corner point / a side edge) instead of the
// reference's pointer juggling, but visits cells in the same row-major order and keeps the same
// "id 0 reads as unset" rule (Boxel.cpp:6-14), so the numbering is identical.
#include "topology.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <unordered_map>

namespace mevo {

namespace {

constexpr double kCell = 0.1;                       // ObjectCreator.h:36
constexpr double kMass = 1.0;                       // ObjectCreator.h:38
const double kRigidMain = 300000000.0 / 3.5;        // ObjectCreator.h:40-47
const double kRigidDiag = kRigidMain / 2.0;
const double kActMain = 300000000.0 / 5;
const double kActDiag = kActMain / 2.0;
const double kSoftMain = 300000000.0 / 6;
const double kSoftDiag = kSoftMain / 4.0;

enum Corner { TL = 0, TR = 1, BR = 2, BL = 3 };
enum Side { S_TOP = 0, S_RIGHT = 1, S_BOT = 2, S_LEFT = 3 };

struct Donor { int dir; int idx; };  // neighbour direction, and which of ITS corners/sides to take

// For each of my corners: (neighbour, neighbour's corner) in the reference's override order --
// later entries win (ObjectCreator.cpp:333-393).
const Donor kCornerDonors[4][3] = {
    /* TL */ {{TOP, BL}, {LEFT, TR}, {TOP_LEFT, BR}},
    /* TR */ {{TOP, BR}, {RIGHT, TL}, {TOP_RIGHT, BL}},
    /* BR */ {{BOT, TR}, {RIGHT, BL}, {BOT_RIGHT, TL}},
    /* BL */ {{BOT, TL}, {LEFT, BR}, {BOT_LEFT, TR}},
};
// x/y offsets (in cells, y down) of each corner relative to the cell's top-left
const int kCornerDx[4] = {0, 1, 1, 0};
const int kCornerDy[4] = {0, 0, 1, 1};
// end points of each side as corners, in creation order (ObjectCreator.cpp:400-423)
const int kSideA[4] = {TL, TR, BL, TL};
const int kSideB[4] = {TR, BR, BR, BL};
// outward reference direction of each side (ObjectCreator.cpp:617-625)
const double kSideNx[4] = {0.0, 1.0, 0.0, -1.0};
const double kSideNy[4] = {1.0, 0.0, -1.0, 0.0};
const int kSideMissingNeighbour[4] = {TOP, RIGHT, BOT, LEFT};

const int kNdx[8] = {0, 0, -1, 1, -1, 1, -1, 1};
const int kNdy[8] = {-1, 1, 0, 0, -1, -1, 1, 1};

// ---- built-object cache.  Design loops and batched front ends add the same terrain object to thousands of worlds, each time after
// a different robot: everything about the object except its global point / edge offsets is the same.  A successful build of an
// object of >= kCacheMinCells cells that is not the first of its world (with offset 0 the "id 0 reads as unset" rule changes what
// is shared, see the header comment) is kept, keyed by its complete input, and instantiated by shifting ids.
constexpr int kCacheMinCells = 32;
constexpr size_t kCacheMaxEntries = 128;

struct ObjectTemplate {
    int P0 = 0, E0 = 0;
    std::vector<double> pos, el0, ek;
    std::vector<uint8_t> fixed, on_surface;
    std::vector<int> ea, eb;
    VoxelObject obj;
};

std::mutex g_cache_mu;
std::unordered_map<std::string, std::shared_ptr<const ObjectTemplate>> g_cache;

std::string cache_key(const double* grid, int w, int h, const double* conn, int k, double x, double y, bool is_robot) {
    std::string key;
    key.reserve(64 + (size_t)w * h + 8 * (size_t)k);
    auto put = [&](const void* p, size_t n) { key.append(static_cast<const char*>(p), n); };
    const int hdr[4] = {w, h, k, is_robot ? 1 : 0};
    put(hdr, sizeof(hdr)); put(&x, sizeof(x)); put(&y, sizeof(y));
    for (int i = 0; i < w * h; i++) { const signed char c = (signed char)grid[i]; key.push_back((char)c); if ((double)c != grid[i]) return std::string(); }
    for (int i = 0; i < 2 * k; i++) { const int v = (int)conn[i]; put(&v, sizeof(v)); if ((double)v != conn[i]) return std::string(); }
    return key;
}

}  // namespace

void World::init(int w, int h) {
    world_w = w;
    world_h = h;
    world_grid.assign((size_t)(w > 0 ? w : 0) * (size_t)(h > 0 ? h : 0), 0);
}

int World::find_object(const std::string& name) const {
    auto it = name_to_index.find(name);
    return it == name_to_index.end() ? -1 : it->second;
}

bool World::add_object(const double* grid, int w, int h, const double* conn, int k,
                       const std::string& name, double x, double y, bool is_robot) {
    const int n_cells = w * h;
    const int P0 = num_points(), E0 = num_edges();
    const double sx = x * kCell, sy = y * kCell;          // :197
    const double max_y = sy + h * kCell;                  // :201
    auto cell = [w](int cx, int cy) { return cy * w + cx; };

    // -- bounds / overlap against the world occupancy grid (:204-214)
    for (int cy = 0; cy < h; cy++)
        for (int cx = 0; cx < w; cx++) {
            if (grid[cell(cx, cy)] == 0) continue;
            const int wx = (int)(cx + x), wy = (int)(world_h - h + cy - y);
            if (wx < 0 || wx >= world_w || wy < 0 || wy >= world_h) {
                std::printf("Error: Could not read in object - object has cells outside of defined grid area.\n");
                return false;
            }
            if (world_grid[(size_t)wy * world_w + wx] != 0) {
                std::printf("Error: Could not read in object - object has overlapping cells with other objects.\n");
                return false;
            }
        }

    auto mark_world_grid = [&]() {
        for (int cy = 0; cy < h; cy++)
            for (int cx = 0; cx < w; cx++)
                if (grid[cell(cx, cy)] != 0)
                    world_grid[(size_t)((int)(world_h - h + cy - y)) * world_w + (int)(cx + x)] = (int)grid[cell(cx, cy)];
    };
    const bool cacheable = n_cells >= kCacheMinCells && P0 > 0 && E0 > 0;
    std::string key;
    if (cacheable) {
        key = cache_key(grid, w, h, conn, k, x, y, is_robot);
        std::shared_ptr<const ObjectTemplate> t;
        if (!key.empty()) {
            std::lock_guard<std::mutex> lock(g_cache_mu);
            auto it = g_cache.find(key);
            if (it != g_cache.end()) t = it->second;
        }
        if (t) {
            // a known-good build: what can still refuse it is the name (same message and order as below)
            if (name_to_index.count(name)) {
                std::printf("Error Could not read in %s %s - duplicate name.\n", is_robot ? "robot" : "object", name.c_str());
                return false;
            }
            const int dp = P0 - t->P0, de = E0 - t->E0;
            name_to_index[name] = (int)objects.size();
            pos.insert(pos.end(), t->pos.begin(), t->pos.end());
            mass.insert(mass.end(), t->fixed.size(), kMass);
            fixed.insert(fixed.end(), t->fixed.begin(), t->fixed.end());
            const size_t ne = t->ea.size();
            ea.reserve(ea.size() + ne); eb.reserve(eb.size() + ne);
            for (size_t i = 0; i < ne; i++) { ea.push_back(t->ea[i] + dp); eb.push_back(t->eb[i] + dp); }
            el0.insert(el0.end(), t->el0.begin(), t->el0.end());
            ek.insert(ek.end(), t->ek.begin(), t->ek.end());
            on_surface.insert(on_surface.end(), t->on_surface.begin(), t->on_surface.end());
            VoxelObject obj = t->obj;
            obj.name = name;
            obj.min_point += dp; obj.max_point += dp; obj.min_edge += de; obj.max_edge += de;
            for (Boxel& b : obj.boxels)
                for (int q = 0; q < 4; q++) { b.pt[q] += dp; b.ed[q] += de; }
            mark_world_grid();
            objects.push_back(std::move(obj));
            return true;
        }
    }

    std::vector<Boxel> cells(n_cells);
    for (int i = 0; i < n_cells; i++) {
        cells[i].cell_type = (int)grid[i];
        cells[i].grid_index = i;
    }

    // -- cardinal neighbour flags from the connection list (:244-271)
    for (int i = 0; i < k; i++) {
        int b1 = (int)conn[i], b2 = (int)conn[k + i];
        if (b1 > b2) { int t = b1; b1 = b2; b2 = t; }
        if (b1 < 0 || b2 >= n_cells) continue;   // the reference would read out of bounds here
        if (b2 - b1 == 1 && w != 1) { cells[b1].nei[RIGHT] = true; cells[b2].nei[LEFT] = true; }
        if (b2 - b1 == w) { cells[b1].nei[BOT] = true; cells[b2].nei[TOP] = true; }
    }

    std::vector<double> new_pos;       // interleaved
    struct NewEdge { int a, b; double l0, k; };
    std::vector<NewEdge> new_edges;
    auto make_point = [&](double px, double py) {
        new_pos.push_back(px); new_pos.push_back(py);
        return (int)(new_pos.size() / 2) + P0 - 1;
    };
    auto make_edge = [&](int a, int b, double l0, double kk) {
        new_edges.push_back({a, b, l0, kk});
        return (int)new_edges.size() + E0 - 1;
    };

    // -- points and shared side edges, row-major scan (:275-430)
    for (int cy = 0; cy < h; cy++)
        for (int cx = 0; cx < w; cx++) {
            Boxel& cur = cells[cell(cx, cy)];
            if (cur.cell_type == CELL_EMPTY) continue;
            const Boxel* nb[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
            for (int d = TOP; d <= RIGHT; d++)
                if (cur.nei[d]) nb[d] = &cells[cell(cx + kNdx[d], cy + kNdy[d])];
            // a diagonal neighbour exists when reachable through two connected cardinal hops (:304-322)
            auto via = [&](int first, int second) { return nb[first] && nb[first]->nei[second]; };
            if (via(TOP, LEFT) || via(LEFT, TOP)) nb[TOP_LEFT] = &cells[cell(cx - 1, cy - 1)];
            if (via(TOP, RIGHT) || via(RIGHT, TOP)) nb[TOP_RIGHT] = &cells[cell(cx + 1, cy - 1)];
            if (via(BOT, LEFT) || via(LEFT, BOT)) nb[BOT_LEFT] = &cells[cell(cx - 1, cy + 1)];
            if (via(BOT, RIGHT) || via(RIGHT, BOT)) nb[BOT_RIGHT] = &cells[cell(cx + 1, cy + 1)];
            for (int d = TOP_LEFT; d <= BOT_RIGHT; d++)
                if (nb[d]) cur.nei[d] = true;

            const int corner_order[4] = {TL, TR, BL, BR};   // creation order of :333-393
            for (int c : corner_order) {
                int id = 0;
                for (const Donor& dn : kCornerDonors[c])
                    if (nb[dn.dir] && nb[dn.dir]->pt[dn.idx] != 0) id = nb[dn.dir]->pt[dn.idx];
                if (id == 0)
                    id = make_point(sx + (cx + kCornerDx[c]) * kCell, max_y - (cy + kCornerDy[c]) * kCell);
                cur.pt[c] = id;
            }
            // sides in the reference's order top, bot, left, right.  Only TOP and LEFT/RIGHT can
            // actually borrow (the donor for BOT lies later in the scan, and :405-406 is dead code).
            const int order[4] = {S_TOP, S_BOT, S_LEFT, S_RIGHT};
            for (int s : order) {
                int id = 0;
                if (s == S_TOP && nb[TOP] && nb[TOP]->ed[S_BOT] != 0) id = nb[TOP]->ed[S_BOT];
                if (s == S_LEFT && nb[LEFT] && nb[LEFT]->ed[S_RIGHT] != 0) id = nb[LEFT]->ed[S_RIGHT];
                if (s == S_RIGHT && nb[RIGHT] && nb[RIGHT]->ed[S_LEFT] != 0) id = nb[RIGHT]->ed[S_LEFT];
                if (id == 0) id = make_edge(cur.pt[kSideA[s]], cur.pt[kSideB[s]], std::fabs(kCell), kRigidMain);
                cur.ed[s] = id;
            }
        }

    // -- per-material constants and the two diagonals of every voxel (:432-494)
    const double diag = std::sqrt(kCell * kCell + kCell * kCell);
    for (int i = 0; i < n_cells; i++) {
        const Boxel& cur = cells[i];
        double k_main = 0, k_diag = 0;
        switch (cur.cell_type) {
            case CELL_RIGID: case CELL_FIXED: k_diag = kRigidDiag; break;
            case CELL_SOFT: k_main = kSoftMain; k_diag = kSoftDiag; break;
            case CELL_ACT_H: case CELL_ACT_V: k_main = kActMain; k_diag = kActDiag; break;
            default: continue;
        }
        if (k_main != 0)
            for (int s = 0; s < 4; s++) new_edges[cur.ed[s] - E0].k = k_main;
        make_edge(cur.pt[TR], cur.pt[BL], diag, k_diag);
        make_edge(cur.pt[TL], cur.pt[BR], diag, k_diag);
    }

    std::vector<uint8_t> new_fixed(new_pos.size() / 2, 0);
    for (const Boxel& c : cells)
        if (c.cell_type == CELL_FIXED)
            for (int q = 0; q < 4; q++) new_fixed[c.pt[q] - P0] = 1;

    // -- exactly one connected component over the cardinal flags (:19-61, :512-532)
    {
        std::vector<int> comp(n_cells, -1);
        int n_comp = 0;
        std::vector<int> stack;
        for (int i = 0; i < n_cells; i++) {
            if (comp[i] != -1 || cells[i].cell_type == CELL_EMPTY) continue;
            stack.assign(1, i);
            while (!stack.empty()) {
                int c = stack.back(); stack.pop_back();
                if (comp[c] != -1 || cells[c].cell_type == CELL_EMPTY) continue;
                comp[c] = n_comp;
                for (int d = TOP; d <= RIGHT; d++)
                    if (cells[c].nei[d]) {
                        int nx = c % w + kNdx[d], ny = c / w + kNdy[d];
                        if (nx >= 0 && nx < w && ny >= 0 && ny < h) stack.push_back(cell(nx, ny));
                    }
            }
            n_comp++;
        }
        if (n_comp > 1) {
            std::printf("Error: Could not read in object - object must be connected.\n");
            return false;
        }
        // a grid without voxels: the reference reports success and adds nothing (ObjectCreator.cpp:193-641 falls through all its
        // loops, no SimObject is registered), so a caller sees `true` and an unchanged world
        if (n_comp == 0) return true;
    }
    if (name_to_index.count(name)) {
        std::printf("Error Could not read in %s %s - duplicate name.\n", is_robot ? "robot" : "object", name.c_str());
        return false;
    }

    // -- commit
    name_to_index[name] = (int)objects.size();
    VoxelObject obj;
    obj.name = name;
    obj.is_robot = is_robot;
    obj.grid_w = w; obj.grid_h = h;
    obj.world_x = (int)x; obj.world_y = (int)y;
    obj.min_point = P0;
    obj.min_edge = E0;
    pos.insert(pos.end(), new_pos.begin(), new_pos.end());
    mass.insert(mass.end(), new_pos.size() / 2, kMass);
    fixed.insert(fixed.end(), new_fixed.begin(), new_fixed.end());
    for (const NewEdge& e : new_edges) {
        ea.push_back(e.a); eb.push_back(e.b); el0.push_back(e.l0); ek.push_back(e.k); on_surface.push_back(0);
    }
    obj.max_point = num_points() - 1;
    obj.max_edge = num_edges() - 1;

    mark_world_grid();

    obj.all_fixed = true;
    for (const Boxel& c : cells) {
        if (c.cell_type == CELL_EMPTY) continue;
        obj.boxels.push_back(c);
        if (c.cell_type != CELL_FIXED) obj.all_fixed = false;
        if (c.cell_type == CELL_RIGID) obj.n_rigid++;
    }

    // -- surface + outward orientation of surface edges (SimObject.cpp:235-307, ObjectCreator.cpp:603-632)
    for (size_t j = 0; j < obj.boxels.size(); j++) {
        Boxel& b = obj.boxels[j];
        for (int s = 0; s < 4; s++) {
            if (b.nei[kSideMissingNeighbour[s]]) continue;
            b.on_surface = true;
            const int e = b.ed[s];
            on_surface[e] = 1;
            const double dx = pos[2 * eb[e]] - pos[2 * ea[e]], dy = pos[2 * eb[e] + 1] - pos[2 * ea[e] + 1];
            const double nx = -dy, ny = dx;
            if (kSideNx[s] * nx + kSideNy[s] * ny < 0) { int t = ea[e]; ea[e] = eb[e]; eb[e] = t; }
        }
        if (b.on_surface) obj.surface_boxels.push_back((int)j);
        if (is_robot && (b.cell_type == CELL_ACT_H || b.cell_type == CELL_ACT_V)) {  // Robot.cpp:9-22
            obj.actuator_grid_to_boxel[b.grid_index] = (int)j;
            obj.actuator_grid_index.push_back(b.grid_index);
        }
    }
    if (cacheable && !key.empty()) {
        auto t = std::make_shared<ObjectTemplate>();
        t->P0 = P0; t->E0 = E0;
        t->pos.assign(pos.begin() + 2 * (size_t)P0, pos.end());
        t->fixed.assign(fixed.begin() + P0, fixed.end());
        t->ea.assign(ea.begin() + E0, ea.end()); t->eb.assign(eb.begin() + E0, eb.end());
        t->el0.assign(el0.begin() + E0, el0.end()); t->ek.assign(ek.begin() + E0, ek.end());
        t->on_surface.assign(on_surface.begin() + E0, on_surface.end());
        t->obj = obj;
        std::lock_guard<std::mutex> lock(g_cache_mu);
        if (g_cache.size() >= kCacheMaxEntries) g_cache.clear();
        g_cache.emplace(std::move(key), std::move(t));
    }
    objects.push_back(std::move(obj));
    return true;
}

}  // namespace mevo
