// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// C-ABI shim over the UNMODIFIED reference engine (evogym/simulator/SimulatorCPP/*.cpp,
// compiled where they lie under /root/reference by oracle/Makefile into
// oracle/_ref/libmevo_ref.so).  It exposes exactly what the parity tests and the CPU
// baseline need:
//   * build a world through the reference's own public entry points
//     (Sim::init / read_robot_from_array / read_object_from_array, SimulatorCPP/Sim.cpp:26-99),
//   * step one substep (Environment::step, SimulatorCPP/Environment.cpp:166-205) or one
//     full step (Sim::step, SimulatorCPP/Sim.cpp:105-115),
//   * dump / overwrite the full dynamic state (teacher forcing), dump topology,
//   * list the leaf-leaf collision candidates in the reference's own descent order
//     (PhysicsEngine::resolve_object_collisions, SimulatorCPP/PhysicsEngine.cpp:752-817).
//
// Private members are reached with the usual test-only `#define private public`; the
// reference translation units themselves are compiled untouched.

#include <cstring>
#include <string>
#include <vector>
#include <map>
#include <set>
#include <unordered_set>
#include <fstream>
#include <iostream>
#include <sstream>
#include <tuple>
#include <cfloat>
#include <cmath>
#include <Eigen/Dense>
#include <Eigen/Core>

#define private public
#include "Sim.h"
#undef private

namespace {

struct RefSim {
    Sim sim;
    std::vector<std::string> names;
};

// The Sim copy-assigns Environment/ObjectCreator in its constructor, which leaves
// PhysicsEngine holding pointers into a dead temporary... except that the reference
// does exactly this and then re-inits: Sim::Sim() calls environment.init() on the member
// (Sim.cpp:11-13), so pointers are valid as long as the Sim object itself never moves.
// We therefore always heap-allocate RefSim and never copy it.

inline Environment& env(void* h) { return static_cast<RefSim*>(h)->sim.environment; }
inline Sim& sim(void* h) { return static_cast<RefSim*>(h)->sim; }

template <class M>
void copy_out(const M& m, double* out) {
    for (int j = 0; j < m.cols(); j++)
        for (int i = 0; i < m.rows(); i++) out[j * m.rows() + i] = m(i, j);
}

struct SilenceStdout {
    std::streambuf* old;
    std::ostringstream sink;
    SilenceStdout() : old(std::cout.rdbuf(sink.rdbuf())) {}
    ~SilenceStdout() { std::cout.rdbuf(old); }
};

void descend(SimObject* oa, BBTreeNode* a, SimObject* ob, BBTreeNode* b, int ia, int ib,
             std::vector<int>& out) {
    // Same control flow as PhysicsEngine::resolve_object_collisions (PhysicsEngine.cpp:752-817),
    // recording leaf-leaf hits instead of resolving them.
    if (!a->bbox.in(&b->bbox)) return;
    if (a->is_leaf && b->is_leaf) {
        out.push_back(ia); out.push_back(a->boxel_index);
        out.push_back(ib); out.push_back(b->boxel_index);
        return;
    }
    if (a->is_leaf) {
        descend(oa, a, ob, &ob->nodes[b->a_index], ia, ib, out);
        descend(oa, a, ob, &ob->nodes[b->b_index], ia, ib, out);
        return;
    }
    if (b->is_leaf) {
        descend(oa, &oa->nodes[a->a_index], ob, b, ia, ib, out);
        descend(oa, &oa->nodes[a->b_index], ob, b, ia, ib, out);
        return;
    }
    if (a->bbox.area() < b->bbox.area()) {
        descend(oa, a, ob, &ob->nodes[b->a_index], ia, ib, out);
        descend(oa, a, ob, &ob->nodes[b->b_index], ia, ib, out);
    } else {
        descend(oa, &oa->nodes[a->a_index], ob, b, ia, ib, out);
        descend(oa, &oa->nodes[a->b_index], ob, b, ia, ib, out);
    }
}

}  // namespace

extern "C" {

void* mref_create(int world_w, int world_h) {
    RefSim* r = new RefSim();
    r->sim.init(world_w, world_h);
    return r;
}

void mref_destroy(void* h) { delete static_cast<RefSim*>(h); }

// grid: row-major h x w, row 0 = TOP row (what EvoSim passes after np.flipud, evogym/sim.py:67).
// conn: 2 x k, conn[0*k + i] / conn[1*k + i] = the two voxel indices of connection i.
int mref_add_object(void* h, const double* grid, int w, int hgt, const double* conn, int k,
                    const char* name, double x, double y, int is_robot, int quiet) {
    Matrix<double, Dynamic, Dynamic> g(hgt, w);
    for (int r = 0; r < hgt; r++)
        for (int c = 0; c < w; c++) g(r, c) = grid[r * w + c];
    Matrix<double, 2, Dynamic> cn(2, k);
    for (int i = 0; i < k; i++) { cn(0, i) = conn[i]; cn(1, i) = conn[k + i]; }
    bool ok;
    if (quiet) {
        SilenceStdout s;
        ok = is_robot ? sim(h).read_robot_from_array(g, cn, name, x, y)
                      : sim(h).read_object_from_array(g, cn, name, x, y);
    } else {
        ok = is_robot ? sim(h).read_robot_from_array(g, cn, name, x, y)
                      : sim(h).read_object_from_array(g, cn, name, x, y);
    }
    if (ok) static_cast<RefSim*>(h)->names.push_back(name);
    return ok ? 1 : 0;
}

// Sim::read_object_from_file / read_robot_from_file (Sim.cpp:30-48 -> ObjectCreator.cpp:118-171): the reference's own text loader
int mref_add_object_from_file(void* h, const char* file, const char* name, double x, double y, int is_robot, int quiet) {
    bool ok;
    if (quiet) {
        SilenceStdout s;
        ok = is_robot ? sim(h).read_robot_from_file(file, name, x, y) : sim(h).read_object_from_file(file, name, x, y);
    } else {
        ok = is_robot ? sim(h).read_robot_from_file(file, name, x, y) : sim(h).read_object_from_file(file, name, x, y);
    }
    if (ok) static_cast<RefSim*>(h)->names.push_back(name);
    return ok ? 1 : 0;
}

int mref_num_points(void* h) { return env(h).get_num_points(); }
int mref_num_edges(void* h) { return env(h).get_num_edges(); }
int mref_num_objects(void* h) { return (int)env(h).get_objects()->size(); }

void mref_get_state(void* h, double* pos, double* vel, double* pos_last, double* vel_true) {
    Environment& e = env(h);
    if (pos) copy_out(e.points_pos, pos);
    if (vel) copy_out(e.points_vel, vel);
    if (pos_last) copy_out(e.points_pos_last, pos_last);
    if (vel_true) copy_out(e.points_vel_true, vel_true);
}

void mref_set_state(void* h, const double* pos, const double* vel, const double* pos_last,
                    const double* vel_true) {
    Environment& e = env(h);
    int n = e.get_num_points();
    for (int j = 0; j < n; j++)
        for (int i = 0; i < 2; i++) {
            if (pos) e.points_pos(i, j) = pos[2 * j + i];
            if (vel) e.points_vel(i, j) = vel[2 * j + i];
            if (pos_last) e.points_pos_last(i, j) = pos_last[2 * j + i];
            if (vel_true) e.points_vel_true(i, j) = vel_true[2 * j + i];
        }
}

void mref_get_point_consts(void* h, double* mass, unsigned char* fixed) {
    Environment& e = env(h);
    int n = e.get_num_points();
    for (int j = 0; j < n; j++) {
        if (mass) mass[j] = e.points_mass(0, j);
        if (fixed) fixed[j] = e.points_fixed(0, j) ? 1 : 0;
    }
}

void mref_get_edges(void* h, int* a, int* b, double* k, double* l0, double* leq, double* goal,
                    unsigned char* on_surface) {
    Environment& e = env(h);
    int n = e.get_num_edges();
    for (int i = 0; i < n; i++) {
        if (a) a[i] = e.a_index[i];
        if (b) b[i] = e.b_index[i];
        if (k) k[i] = e.spring_const[i];
        if (l0) l0[i] = e.init_length_eq[i];
        if (leq) leq[i] = e.length_eq[i];
        if (goal) goal[i] = e.length_eq_goal[i];
        if (on_surface) on_surface[i] = e.edges[i].isOnSurface ? 1 : 0;
    }
}

void mref_set_edge_state(void* h, const double* leq, const double* goal) {
    Environment& e = env(h);
    int n = e.get_num_edges();
    for (int i = 0; i < n; i++) {
        if (leq) e.length_eq[i] = leq[i];
        if (goal) e.length_eq_goal[i] = goal[i];
    }
}

// info[0..6] = is_robot, min_point_index, max_point_index, n_boxels, grid_w, grid_h, n_surface_boxels
void mref_object_info(void* h, int obj, int* info) {
    SimObject* o = env(h).get_objects()->at(obj);
    info[0] = o->is_robot ? 1 : 0;
    info[1] = o->min_point_index;
    info[2] = o->max_point_index;
    info[3] = (int)o->boxels.size();
    info[4] = o->grid_w;
    info[5] = o->grid_h;
    info[6] = (int)o->surface_boxels_index.size();
}

// per boxel 18 ints: cell_type, grid_index, TL,TR,BR,BL points, top,right,bot,left edges, 8 neighbour flags
void mref_object_boxels(void* h, int obj, int* out) {
    SimObject* o = env(h).get_objects()->at(obj);
    for (size_t j = 0; j < o->boxels.size(); j++) {
        Boxel& b = o->boxels[j];
        int* r = out + 18 * j;
        r[0] = b.cell_type; r[1] = b.grid_index;
        r[2] = b.point_top_left_index; r[3] = b.point_top_right_index;
        r[4] = b.point_bot_right_index; r[5] = b.point_bot_left_index;
        r[6] = b.edge_top_index; r[7] = b.edge_right_index;
        r[8] = b.edge_bot_index; r[9] = b.edge_left_index;
        for (int n = 0; n < 8; n++) r[10 + n] = b.neighbors[n] ? 1 : 0;
    }
}

// surface boxel indices in the reference's own (unordered_set iteration) order
void mref_object_surface_boxels(void* h, int obj, int* out) {
    SimObject* o = env(h).get_objects()->at(obj);
    for (int j = 0; j < o->surface_boxels_index.size(); j++) out[j] = (int)o->surface_boxels_index[j];
}

const char* mref_object_name(void* h, int obj) {
    RefSim* r = static_cast<RefSim*>(h);
    return (obj >= 0 && obj < (int)r->names.size()) ? r->names[obj].c_str() : "";
}

int mref_num_actuators(void* h, const char* robot) {
    SilenceStdout s;
    Robot* r = env(h).get_robot(robot);
    return r ? (int)r->actuator_indicies.size() : -1;
}

void mref_actuator_indices(void* h, const char* robot, int* out) {
    Robot* r = env(h).get_robot(robot);
    if (!r) return;
    for (int i = 0; i < r->actuator_indicies.size(); i++) out[i] = r->actuator_indicies[i];
}

// action: n x 2 row-major (grid_index, value) -- what EvoSim.set_action stacks (evogym/sim.py:141)
void mref_set_action(void* h, const char* robot, const double* action, int n) {
    MatrixXd a(n, 2);
    for (int i = 0; i < n; i++) { a(i, 0) = action[2 * i]; a(i, 1) = action[2 * i + 1]; }
    sim(h).set_action(robot, a);
}

int mref_substep(void* h) { return env(h).step() ? 1 : 0; }
int mref_step(void* h) { return sim(h).step() ? 1 : 0; }
int mref_steps(void* h, int n) {
    int unstable = 0;
    for (int i = 0; i < n; i++) unstable += sim(h).step() ? 1 : 0;
    return unstable;
}
int mref_get_time(void* h) { return sim(h).get_time(); }
void mref_revert(void* h, long t) { SilenceStdout s; sim(h).revert(t); }
void mref_translate_object(void* h, double x, double y, const char* name) {
    sim(h).translate_object(x, y, name);
}
double mref_object_orientation_at_time(void* h, long t, const char* name) {
    return sim(h).object_orientation_at_time(t, name);
}

// returns number of columns written (0 if the time/name is unknown)
int mref_pos_at_time(void* h, long t, double* out, int max_cols, int vel) {
    Environment& e = env(h);
    if (e.history.count(t) <= 0) return 0;
    const Matrix<double, 2, Dynamic>& m = vel ? e.history[t].points_vel : e.history[t].points_pos;
    int n = (int)m.cols() < max_cols ? (int)m.cols() : max_cols;
    for (int j = 0; j < n; j++) { out[2 * j] = m(0, j); out[2 * j + 1] = m(1, j); }
    return n;
}

// Individual phases of Environment::step, for phase-by-phase parity.
void mref_phase_collisions(void* h) { env(h).physics_handler.resolve_collisions(); }
void mref_phase_actuators(void* h) { env(h).physics_handler.update_actuators(); }
void mref_phase_rk4(void* h) { env(h).physics_handler.step_rk4(0.0001); }
void mref_phase_constraints(void* h) { env(h).physics_handler.resolve_edge_constraints(); }
void mref_phase_true_vel(void* h) { env(h).physics_handler.update_true_vel(0.0001); }
int mref_phase_self_collide(void* h) { return env(h).physics_handler.is_any_robot_self_colliding() ? 1 : 0; }

void mref_get_collision_forces(void* h, double* out) {
    PhysicsEngine& p = env(h).physics_handler;
    int n = env(h).get_num_points();
    if (p.collision_forces.cols() != n) { std::memset(out, 0, sizeof(double) * 2 * n); return; }
    copy_out(p.collision_forces, out);
}

// Leaf-leaf candidates for the CURRENT positions, in the reference's descent order.
// Each candidate = 4 ints (object_a, boxel_a, object_b, boxel_b).  Returns the count; writes at
// most max_cands of them.
int mref_candidates(void* h, int* out, int max_cands) {
    Environment& e = env(h);
    std::vector<SimObject*>* objs = e.get_objects();
    for (size_t i = 0; i < objs->size(); i++) objs->at(i)->recompute_bbs(e.points_pos);
    std::vector<int> rec;
    for (size_t i = 0; i < objs->size(); i++)
        for (size_t j = i; j < objs->size(); j++)
            descend(objs->at(i), &objs->at(i)->nodes[objs->at(i)->tree_root], objs->at(j),
                    &objs->at(j)->nodes[objs->at(j)->tree_root], (int)i, (int)j, rec);
    int n = (int)rec.size() / 4;
    int w = n < max_cands ? n : max_cands;
    if (out) std::memcpy(out, rec.data(), sizeof(int) * 4 * w);
    return n;
}

}  // extern "C"
