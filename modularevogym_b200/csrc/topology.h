// Host-side world builder: voxel grids -> point masses, springs, boxels, surfaces.
//
// Produces, array for array, what the reference's ObjectCreator + SimObject::compute_surface +
// edge-orientation pass + Robot::init produce (reference evogym/simulator/SimulatorCPP/
// ObjectCreator.cpp:193-641, SimObject.cpp:235-307, Environment.cpp:111-116, Robot.cpp:9-22), so
// global point / edge numbering, spring constants, fixed flags, surface flags and surface-edge
// orientation are identical and everything downstream (actions, read-back, contact normals) lines
// up with the reference engine.  The bounding-box tree is NOT built: the device broad phase uses a
// uniform grid + all-pairs test that yields the same leaf-pair set.
#pragma once

#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace mevo {

enum CellType : int { CELL_EMPTY = 0, CELL_RIGID, CELL_SOFT, CELL_ACT_H, CELL_ACT_V, CELL_FIXED };
enum Dir : int { TOP = 0, BOT, LEFT, RIGHT, TOP_LEFT, TOP_RIGHT, BOT_LEFT, BOT_RIGHT };

// One voxel ("boxel" in the reference, Boxel.h:12-59).
struct Boxel {
    int cell_type = CELL_EMPTY;
    int grid_index = 0;          // y_top * grid_w + x inside the object's own grid
    int pt[4] = {0, 0, 0, 0};    // TL, TR, BR, BL   (Boxel.cpp:24)
    int ed[4] = {0, 0, 0, 0};    // top, right, bot, left (Boxel.cpp:25)
    bool nei[8] = {false, false, false, false, false, false, false, false};
    bool on_surface = false;     // at least one cardinal neighbour missing -> leaf of the broad phase
};

struct VoxelObject {
    std::string name;
    bool is_robot = false;
    int min_point = 0, max_point = -1;   // inclusive global point range (ObjectCreator.cpp:589-593)
    int min_edge = 0, max_edge = -1;     // inclusive global edge range
    int grid_w = 0, grid_h = 0;
    int world_x = 0, world_y = 0;        // bottom-left voxel in the world grid
    std::vector<Boxel> boxels;           // non-empty cells, row-major from the top row
    std::vector<int> surface_boxels;     // ascending indices into boxels
    std::vector<int> actuator_grid_index;        // Robot::init order
    std::map<int, int> actuator_grid_to_boxel;   // grid_index -> index into boxels
    bool all_fixed = false;              // every voxel is CELL_FIXED: never moves, prunable
    int n_rigid = 0;
};

// One environment's world: everything Sim/Environment own for it on the host.
struct World {
    int world_w = 0, world_h = 0;
    std::vector<int> world_grid;             // occupancy, row 0 = top (ObjectCreator.cpp:102-110)
    std::vector<double> pos;                 // 2P interleaved x,y (initial positions)
    std::vector<double> mass;                // P
    std::vector<uint8_t> fixed;              // P
    std::vector<int> ea, eb;                 // E (after surface orientation)
    std::vector<double> ek, el0;             // E
    std::vector<uint8_t> on_surface;         // E
    std::vector<VoxelObject> objects;
    std::map<std::string, int> name_to_index;

    void init(int w, int h);
    int num_points() const { return (int)mass.size(); }
    int num_edges() const { return (int)ea.size(); }

    // grid: row-major h x w, row 0 = top row.  conn: 2 x k (conn[i], conn[k+i]).  Returns true on
    // success; on refusal prints the reference's own "Error: ..." line to stdout and returns false
    // leaving the world untouched (ObjectCreator.cpp:173-191,526-541).
    bool add_object(const double* grid, int w, int h, const double* conn, int k,
                    const std::string& name, double x, double y, bool is_robot);

    int find_object(const std::string& name) const;
};

}  // namespace mevo
