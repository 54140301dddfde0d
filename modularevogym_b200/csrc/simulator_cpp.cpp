// pybind11 module `simulator_cpp` -- the reference's Python boundary (SimulatorCPP/PythonBindings.cpp:14-59)
// over the C ABI of include/mevo_b200.h.  Same class and method names, argument meaning and error behaviour,
// so `evogym.sim.EvoSim(Sim)` (evogym/sim.py:14,18) and `evogym.viewer` (evogym/viewer.py:6) import and run
// unchanged.  `Sim` is a one-environment view of a batch; `BatchSim` exposes N environments per GPU.
//
// Host-side C++ only: no CUDA here, everything goes through libmevo_b200.so.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "mevo_b200.h"

namespace py = pybind11;
using darray = py::array_t<double, py::array::c_style | py::array::forcecast>;

namespace {

int default_device() {
    const char* e = std::getenv("MEVO_DEVICE");
    return e ? std::atoi(e) : 0;
}

[[noreturn]] void fail() { throw std::runtime_error(mevo_last_error()); }

py::array_t<double> zeros_2x4() {   // what the reference hands back for an unknown time / name (Environment.cpp:267-272)
    py::array_t<double> a({2, 4});
    std::fill(a.mutable_data(), a.mutable_data() + 8, 0.0);
    return a;
}

struct Snapshot {                    // SimulatorCPP/Snapshot.h:11-29
    std::vector<double> pos, vel;    // (P, 2) row-major
};

// columns [c0, c1] of a (P,2) row-major buffer as a (2, n) array (the reference's column-major 2 x n matrices)
py::array_t<double> as_2xn(const std::vector<double>& buf, int c0, int c1) {
    const int n = c1 - c0 + 1;
    py::array_t<double> a({2, n});
    auto r = a.mutable_unchecked<2>();
    for (int j = 0; j < n; j++) {
        r(0, j) = buf[2 * (c0 + j)];
        r(1, j) = buf[2 * (c0 + j) + 1];
    }
    return a;
}

// ------------------------------------------------------------------------------------------------
class Sim {
public:
    Sim() : batch_(nullptr), env_(-1) {}
    ~Sim() { if (batch_) mevo_batch_destroy(batch_); }
    Sim(const Sim&) = delete;
    Sim& operator=(const Sim&) = delete;

    mevo_batch* batch_handle() const { return batch_; }
    int env_index() const { return env_; }

    static void get_version() { std::cout << "Using Evolution Gym Simulator v2.2.5 (mevo_b200 CUDA engine)" << "\n"; }

    void init(int x, int y) {                                  // Sim::init, Sim.cpp:26-28
        if (!batch_) {
            batch_ = mevo_batch_create(default_device());
            if (!batch_) fail();
        }
        if (env_ >= 0) throw std::runtime_error("Sim.init called twice");
        env_ = mevo_batch_begin_env(batch_, x, y);
        if (env_ < 0) fail();
    }

    bool read_object_from_array(darray grid, darray connections, const std::string& name, double x, double y) {
        return add(grid, connections, name, x, y, false);
    }
    bool read_robot_from_array(darray grid, darray connections, const std::string& name, double x, double y) {
        return add(grid, connections, name, x, y, true);
    }
    bool read_object_from_file(const std::string& file, const std::string& name, double x, double y) {
        return add_file(file, name, x, y, false);
    }
    bool read_robot_from_file(const std::string& file, const std::string& name, double x, double y) {
        return add_file(file, name, x, y, true);
    }

    void set_action(const std::string& robot, darray action) {  // Sim.cpp:101-103; rows = (actuator grid index, value)
        need_env();
        if (action.ndim() != 2 || action.shape(1) != 2) throw std::invalid_argument("action must be an (n, 2) array");
        if (mevo_env_set_action(batch_, env_, robot.c_str(), action.data(), (int)action.shape(0)) < 0) fail();
    }

    bool step() {                                               // Sim.cpp:105-115: true = unstable, time not advanced
        need_env();
        unsigned char flag = 0;
        if (mevo_batch_step(batch_, 1, 0, &flag) < 0) fail();
        if (flag) return true;
        save_snapshot(get_time());
        return false;
    }

    void revert(long t) {                                       // Sim.cpp:126-130 + Environment.cpp:247-263
        need_env();
        auto it = history_.find(t);
        if (it == history_.end()) {
            std::cout << "Error: Could not revert simulation - no save found for time argument.\n";
            return;
        }
        const Snapshot& s = it->second;
        if (mevo_env_set_state(batch_, env_, s.pos.data(), s.vel.data(), s.pos.data(), s.vel.data()) < 0) fail();
        if (mevo_env_set_time(batch_, env_, t) < 0) fail();
        history_.erase(std::next(it), history_.end());
    }

    int get_time() {
        need_env();
        return (int)mevo_env_get_time(batch_, env_);
    }

    py::array_t<double> pos_at_time(long t) { return whole(t, false); }
    py::array_t<double> vel_at_time(long t) { return whole(t, true); }
    py::array_t<double> object_pos_at_time(long t, const std::string& name) { return part(t, name, false); }
    py::array_t<double> object_vel_at_time(long t, const std::string& name) { return part(t, name, true); }

    // Environment::object_orientation_at_time (Environment.cpp:317-380): magnitude-weighted mean angle between the
    // centred point vectors at time t and at time 0, flipped to [0, 2 pi) by the sign of the summed cross product
    double object_orientation_at_time(long t, const std::string& name) {
        need_env();
        auto now = history_.find(t), then = history_.find(0);
        int c0, c1;
        if (now == history_.end() || then == history_.end() || !range(name, c0, c1)) return 0;
        const int n = c1 - c0 + 1;
        const std::vector<double>&pn = now->second.pos, &po = then->second.pos;
        // centroid in the order of the reference's rowwise().sum() (Eigen packetwise redux: column 0, then four columns at a
        // time as (c1 + c2) + (c3 + c4), then one by one) -- see orientation_kernel in kernels.cuh for why the order matters
        auto rowsum = [&](const std::vector<double>& p, int comp) {
            double acc = p[2 * c0 + comp];
            int i = 1;
            const int size4 = (n - 1) & ~3;
            auto at = [&](int k) { return p[2 * (c0 + k) + comp]; };
            for (; i < size4; i += 4) acc += (at(i) + at(i + 1)) + (at(i + 2) + at(i + 3));
            for (; i < n; i++) acc += at(i);
            return acc;
        };
        const double cnx = rowsum(pn, 0) / n, cny = rowsum(pn, 1) / n, cox = rowsum(po, 0) / n, coy = rowsum(po, 1) / n;
        double weighted = 0, weights = 0, cross = 0;
        for (int j = c0; j <= c1; j++) {
            double ax = pn[2 * j] - cnx, ay = pn[2 * j + 1] - cny, bx = po[2 * j] - cox, by = po[2 * j + 1] - coy;
            const double ma = std::sqrt(ax * ax + ay * ay), mb = std::sqrt(bx * bx + by * by);
            if (ax * ax + ay * ay > 0) { ax /= ma; ay /= ma; }
            if (bx * bx + by * by > 0) { bx /= mb; by /= mb; }
            double c = ax * bx + ay * by;
            if (ma * mb < 0.0000001) c = 0.99999;
            if (c > 1.0) c = 0.99999;
            if (c < -1.0) c = -0.99999;
            weighted += std::acos(c) * mb;
            weights += mb;
            cross += ax * by - ay * bx;
        }
        double angle = weighted / weights;
        if (cross > 0) angle = 2 * 3.1415926 - angle;
        return angle;
    }

    void translate_object(double x, double y, const std::string& name) {   // Sim.cpp:148-151
        need_env();
        if (mevo_env_translate_object(batch_, env_, name.c_str(), x, y) < 0) fail();
        save_snapshot(get_time());
    }

    py::array_t<int> get_indices_of_actuators(const std::string& robot) {  // (1, A) int32, Robot.cpp:9-22 order
        need_env();
        std::vector<int> buf(4096);
        const int n = mevo_env_actuator_indices(batch_, env_, robot.c_str(), buf.data(), (int)buf.size());
        if (n < 0) fail();
        py::array_t<int> a({1, n});
        for (int i = 0; i < n; i++) a.mutable_at(0, i) = buf[i];
        return a;
    }

    mevo_batch* handle() { return batch_; }

private:
    void need_env() const { if (!batch_ || env_ < 0) throw std::runtime_error("Sim.init(width, height) has not been called"); }

    bool add(const darray& grid, const darray& conn, const std::string& name, double x, double y, bool robot) {
        need_env();
        if (grid.ndim() != 2) throw std::invalid_argument("grid must be a 2-d array");
        int k = 0;
        if (conn.size() > 0) {
            if (conn.ndim() != 2 || conn.shape(0) != 2) throw std::invalid_argument("connections must be a (2, k) array");
            k = (int)conn.shape(1);
        }
        const int rc = mevo_batch_add_object(batch_, env_, grid.data(), (int)grid.shape(1), (int)grid.shape(0), k ? conn.data() : nullptr, k,
                                             name.c_str(), x, y, robot ? 1 : 0);
        if (rc < 0) fail();
        if (rc == 1) save_snapshot(0);                          // Sim.cpp:70-71,94-96
        return rc == 1;
    }

    // text format of ObjectCreator::read_object_from_file (ObjectCreator.cpp:118-171):
    // width height, width*height cell types (row 0 = top), then index pairs until end of file
    bool add_file(const std::string& file, const std::string& name, double x, double y, bool robot) {
        need_env();
        std::ifstream f(file);
        if (!f.is_open() || !f.good()) {
            std::cout << "Error: Could not read in object - file '" << file << "' not found.\n";
            return false;
        }
        int w = 0, h = 0;
        f >> w >> h;
        if (w <= 0 || h <= 0) return false;
        std::vector<double> grid((size_t)w * h, 0.0);
        for (auto& v : grid) { int t = 0; f >> t; v = t; }
        std::vector<double> c0, c1;
        int a, b;
        while (f >> a >> b) { c0.push_back(a); c1.push_back(b); }
        std::vector<double> conn(c0);
        conn.insert(conn.end(), c1.begin(), c1.end());
        const int rc = mevo_batch_add_object(batch_, env_, grid.data(), w, h, conn.empty() ? nullptr : conn.data(), (int)c0.size(),
                                             name.c_str(), x, y, robot ? 1 : 0);
        if (rc < 0) fail();
        if (rc == 1) save_snapshot(0);
        return rc == 1;
    }

    void save_snapshot(long t) {                                // Environment::save_snapshot, Environment.cpp:239-245
        const int P = mevo_env_num_points(batch_, env_);
        Snapshot s;
        s.pos.resize(2 * (size_t)P);
        s.vel.resize(2 * (size_t)P);
        if (mevo_env_get_state(batch_, env_, s.pos.data(), s.vel.data(), nullptr, nullptr) < 0) fail();
        history_[t] = std::move(s);
    }

    bool range(const std::string& name, int& c0, int& c1) {
        const int o = mevo_env_object_index(batch_, env_, name.c_str());
        if (o < 0) return false;
        int info[8];
        if (mevo_env_object_info(batch_, env_, o, info) < 0) return false;
        c0 = info[1]; c1 = info[2];
        return true;
    }

    py::array_t<double> whole(long t, bool vel) {
        need_env();
        auto it = history_.find(t);
        if (it == history_.end()) return zeros_2x4();
        const std::vector<double>& b = vel ? it->second.vel : it->second.pos;
        return as_2xn(b, 0, (int)b.size() / 2 - 1);
    }

    py::array_t<double> part(long t, const std::string& name, bool vel) {
        need_env();
        auto it = history_.find(t);
        int c0, c1;
        if (it == history_.end() || !range(name, c0, c1)) return zeros_2x4();
        return as_2xn(vel ? it->second.vel : it->second.pos, c0, c1);
    }

    mevo_batch* batch_;
    int env_;
    std::map<long, Snapshot> history_;                          // Environment.h:57
};

// ------------------------------------------------------------------------------------------------
// Viewer / Camera: the reference draws through OpenGL (SimulatorCPP/Interface.cpp, Camera.cpp).  There is no
// graphics stack on the B200 boxes, so the classes exist (evogym/viewer.py imports them) and say so when used.
// Camera / Viewer: the reference renders with OpenGL (Camera.cpp, Interface.cpp), which is unavailable on this box.  An image
// camera (`Camera(True)`, evogym/viewer.py:393-410) is rendered headlessly by the engine's CUDA rasteriser (csrc/render.cu)
// into the buffer get_image hands out; a screen camera (`Camera(False)`) has no window to draw to and renders nothing.
class Camera {
public:
    explicit Camera(bool is_renderable) : renderable_(is_renderable) {}
    void set_pos(double x, double y) { px_ = x; py_ = y; }
    void set_size(double x, double y) { sx_ = x; sy_ = y; }
    void set_resolution(double x, double y) {
        rw_ = (int)x; rh_ = (int)y;
        if (renderable_) image_.assign((size_t)std::max(0, rw_) * std::max(0, rh_) * 3, 0);
    }
    int get_resolution_width() const { return rw_; }
    int get_resolution_height() const { return rh_; }
    py::memoryview get_image() {                                 // like the reference: a view of 3 * w * h bytes, row 0 = top
        if (!renderable_) throw std::runtime_error("Camera.get_image: this camera renders to a window, which is unavailable on this box");
        return py::memoryview::from_memory(image_.data(), (py::ssize_t)image_.size());
    }
    bool renderable_;
    double px_ = 0, py_ = 0, sx_ = 1, sy_ = 1;
    int rw_ = 0, rh_ = 0;
    std::vector<unsigned char> image_;
};

class Viewer {
public:
    explicit Viewer(Sim* sim) : sim_(sim) {}
    void init() {}                                               // the reference opens the GLFW context here
    void render(Camera& cam, bool hide_background, bool hide_grid, bool hide_edges, bool hide_boxels, bool dont_clear);
    void show_debug_window() {}
    void hide_debug_window() {}
    std::vector<int> get_debug_window_pos() { return {0, 0}; }
private:
    Sim* sim_;
};

void Viewer::render(Camera& cam, bool hide_background, bool hide_grid, bool hide_edges, bool hide_boxels, bool dont_clear) {
    if (!cam.renderable_) return;                                // screen camera: no window on this box
    if (!sim_ || !sim_->batch_handle() || sim_->env_index() < 0) throw std::runtime_error("Viewer.render: the Sim has no world yet");
    if (cam.image_.empty()) throw std::runtime_error("Viewer.render: set the camera's resolution first");
    const int flags = (hide_background ? 1 : 0) | (hide_grid ? 2 : 0) | (hide_edges ? 4 : 0) | (hide_boxels ? 8 : 0) | (dont_clear ? 16 : 0);
    if (mevo_env_render(sim_->batch_handle(), sim_->env_index(), cam.px_, cam.py_, cam.sx_, cam.sy_, cam.rw_, cam.rh_, flags,
                        cam.image_.data()) < 0)
        fail();
}

// ------------------------------------------------------------------------------------------------
// BatchSim: N independent environments on one GPU (what SubprocVecEnv + N reference Sims provide, in one process).
class BatchSim {
public:
    explicit BatchSim(int device) {
        b_ = mevo_batch_create(device < 0 ? default_device() : device);
        if (!b_) fail();
    }
    ~BatchSim() { if (b_) mevo_batch_destroy(b_); }
    BatchSim(const BatchSim&) = delete;
    BatchSim& operator=(const BatchSim&) = delete;

    int begin_env(int w, int h) { return chk(mevo_batch_begin_env(b_, w, h)); }
    bool add_object(int env, darray grid, darray conn, const std::string& name, double x, double y, bool robot) {
        if (grid.ndim() != 2) throw std::invalid_argument("grid must be a 2-d array");
        const int k = conn.size() > 0 ? (int)conn.shape(1) : 0;
        return chk(mevo_batch_add_object(b_, env, grid.data(), (int)grid.shape(1), (int)grid.shape(0), k ? conn.data() : nullptr, k,
                                         name.c_str(), x, y, robot ? 1 : 0)) == 1;
    }
    void finalize() { chk(mevo_batch_finalize(b_)); }
    int num_envs() const { return mevo_batch_num_envs(b_); }
    int action_stride() const { return mevo_batch_action_stride(b_); }
    py::array_t<int> actuator_indices(int env, const std::string& robot) {
        std::vector<int> buf(4096);
        const int n = chk(mevo_env_actuator_indices(b_, env, robot.c_str(), buf.data(), (int)buf.size()));
        py::array_t<int> a(n);
        for (int i = 0; i < n; i++) a.mutable_at(i) = buf[i];
        return a;
    }
    void set_actions(darray values) {                           // (n_envs, stride) target length ratios
        if (values.ndim() != 2 || values.shape(0) != num_envs()) throw std::invalid_argument("actions must be (n_envs, stride)");
        chk(mevo_batch_set_actions(b_, values.data(), (int)values.shape(1)));
    }
    void set_actions_device(std::uintptr_t ptr, int stride) { chk(mevo_batch_set_actions_device(b_, reinterpret_cast<const double*>(ptr), stride)); }
    void bind_actions_device(std::uintptr_t ptr, int stride) { chk(mevo_batch_bind_actions_device(b_, reinterpret_cast<const double*>(ptr), stride)); }
    void set_auto_reset(bool on) { chk(mevo_batch_set_auto_reset(b_, on ? 1 : 0)); }
    py::array_t<unsigned char> step(int n_steps) {              // per-env unstable flags of the last step
        py::array_t<unsigned char> flags(num_envs());
        chk(mevo_batch_step(b_, n_steps, 1, flags.mutable_data()));
        return flags;
    }
    void step_async(int n_steps) { chk(mevo_batch_step_async(b_, n_steps, 1)); }
    void sync() { chk(mevo_batch_sync(b_)); }
    void reset(py::object env_ids) {
        if (env_ids.is_none()) { chk(mevo_batch_reset(b_, nullptr, 0)); return; }
        auto ids = py::cast<py::array_t<int, py::array::c_style | py::array::forcecast>>(env_ids);
        chk(mevo_batch_reset(b_, ids.data(), (int)ids.size()));
    }
    int get_time(int env) { return (int)mevo_env_get_time(b_, env); }
    py::array_t<double> pos(int env) { return state(env, 0); }
    py::array_t<double> vel(int env) { return state(env, 1); }
    py::array_t<double> object_com(int obj) {                   // (n_envs, 2) centre of mass of object `obj`
        py::array_t<double> a({num_envs(), 2});
        chk(mevo_batch_object_com(b_, obj, a.mutable_data()));
        return a;
    }
    void object_com_device(int obj, std::uintptr_t ptr) { chk(mevo_batch_object_com_device(b_, obj, reinterpret_cast<double*>(ptr))); }
    // (n_envs, width) rows [vel com, relative x, relative y, zeros]: what Walker-v0 observes (evogym/envs/walk.py:58-61)
    py::array_t<double> object_obs(int obj) {
        const int w = mevo_batch_object_obs_width(b_, obj);
        if (w < 0) throw std::runtime_error(mevo_last_error());
        py::array_t<double> a({(py::ssize_t)mevo_batch_num_envs(b_), (py::ssize_t)w});
        chk(mevo_batch_object_obs(b_, obj, a.mutable_data(), w));
        return a;
    }
    void object_obs_device(int obj, std::uintptr_t ptr, int stride) {
        chk(mevo_batch_object_obs_device(b_, obj, reinterpret_cast<double*>(ptr), stride));
    }
    py::array_t<double> object_orientation(int obj) {           // (n_envs,) Sim::object_orientation_at_time, now vs snapshot 0
        py::array_t<double> a(num_envs());
        chk(mevo_batch_object_orientation(b_, obj, a.mutable_data()));
        return a;
    }
    void object_orientation_device(int obj, std::uintptr_t ptr) { chk(mevo_batch_object_orientation_device(b_, obj, reinterpret_cast<double*>(ptr))); }

    // ---- task layer (include/mevo_b200.h "task layer"): the gym environments either side of Sim::step
    void task_configure(int task, const std::string& robot, py::object package, int max_episode_steps) {
        const std::string pk = package.is_none() ? std::string() : py::cast<std::string>(package);
        chk(mevo_batch_task_configure(b_, task, robot.c_str(), pk.empty() ? nullptr : pk.c_str(), max_episode_steps));
    }
    py::dict task_dims() {
        int d[5];
        chk(mevo_batch_task_dims(b_, d));
        py::dict out;
        out["flat"] = d[0]; out["modular"] = d[1]; out["other"] = d[2]; out["max_voxels"] = d[3]; out["action_stride"] = d[4];
        return out;
    }
    // BidirectionalWalker-v0: goal lists (one sequence of x positions in voxels per environment, starting at env_first)
    void task_set_goals(const std::vector<std::vector<int>>& goals, int env_first) {
        size_t stride = 1;
        for (const auto& g : goals) stride = std::max(stride, g.size());
        std::vector<int> counts(goals.size()), table(goals.size() * stride, 0);
        for (size_t i = 0; i < goals.size(); i++) {
            counts[i] = (int)goals[i].size();
            std::copy(goals[i].begin(), goals[i].end(), table.begin() + i * stride);
        }
        chk(mevo_batch_task_set_goals(b_, env_first, (int)goals.size(), counts.data(), table.data(), (int)stride));
    }
    void task_bind_flag_planes(std::uintptr_t device_planes) { chk(mevo_batch_task_bind_flag_planes(b_, reinterpret_cast<void*>(device_planes))); }
    void task_draw_goals(unsigned long long seed) { chk(mevo_batch_task_draw_goals(b_, seed)); }
    py::dict task_get_goals(int env) {
        int count = 0, current = 0, buf[64];
        chk(mevo_batch_task_get_goals(b_, env, &count, &current, buf, 64));
        py::dict out;
        out["goals"] = std::vector<int>(buf, buf + count); out["current"] = current;
        return out;
    }
    py::dict task_env_tables(int env) {
        int d[5];
        chk(mevo_batch_task_dims(b_, d));
        py::array_t<int> idx({d[3], 4}), am(d[3]), om(d[3]);
        std::fill(idx.mutable_data(), idx.mutable_data() + idx.size(), -1);
        std::fill(am.mutable_data(), am.mutable_data() + am.size(), 1);
        std::fill(om.mutable_data(), om.mutable_data() + om.size(), 1);
        const int corners = chk(mevo_batch_task_env_tables(b_, env, idx.mutable_data(), am.mutable_data(), om.mutable_data(), d[3]));
        py::dict out;
        out["obs_index_array"] = idx; out["act_mask"] = am; out["obs_mask"] = om; out["voxel_corner_size"] = corners;
        return out;
    }
    void task_apply_policy(darray actions, bool per_voxel) {
        if (actions.ndim() != 2 || actions.shape(0) != num_envs()) throw std::invalid_argument("actions must be (n_envs, width)");
        chk(mevo_batch_task_apply_policy(b_, actions.data(), (int)actions.shape(1), per_voxel ? 1 : 0));
    }
    void task_apply_policy_device(std::uintptr_t ptr, int stride, bool per_voxel, bool is_f32) {
        chk(mevo_batch_task_apply_policy_device(b_, reinterpret_cast<const void*>(ptr), stride, per_voxel ? 1 : 0, is_f32 ? 1 : 0));
    }
    py::dict task_outputs(bool reset_done) { return task_io(true, reset_done); }
    py::dict task_reset() { return task_io(false, false); }
    void task_outputs_device(std::uintptr_t reward, std::uintptr_t done, std::uintptr_t flat, int flat_stride, std::uintptr_t modular,
                             int modular_stride, std::uintptr_t other, int other_stride, bool is_f32, bool reset_done) {
        chk(mevo_batch_task_outputs_device(b_, reinterpret_cast<void*>(reward), reinterpret_cast<unsigned char*>(done),
                                           reinterpret_cast<void*>(flat), flat_stride, reinterpret_cast<void*>(modular), modular_stride,
                                           reinterpret_cast<void*>(other), other_stride, is_f32 ? 1 : 0, reset_done ? 1 : 0));
    }
    void task_reset_device(std::uintptr_t flat, int flat_stride, std::uintptr_t modular, int modular_stride, std::uintptr_t other,
                           int other_stride, bool is_f32) {
        chk(mevo_batch_task_reset_device(b_, reinterpret_cast<void*>(flat), flat_stride, reinterpret_cast<void*>(modular), modular_stride,
                                         reinterpret_cast<void*>(other), other_stride, is_f32 ? 1 : 0));
    }
    std::uintptr_t cuda_stream() { return reinterpret_cast<std::uintptr_t>(mevo_batch_cuda_stream(b_)); }
    long long unstable_events() { return mevo_batch_unstable_events(b_); }

private:
    static int chk(int rc) { if (rc < 0) fail(); return rc; }
    py::dict task_io(bool step_outputs, bool reset_done) {
        int d[5];
        chk(mevo_batch_task_dims(b_, d));
        const py::ssize_t n = num_envs();
        py::array_t<double> flat({n, (py::ssize_t)d[0]}), modular({n, (py::ssize_t)d[1]}), other({n, (py::ssize_t)d[2]});
        py::dict out;
        if (step_outputs) {
            py::array_t<double> reward(n);
            py::array_t<unsigned char> done(n);
            chk(mevo_batch_task_outputs(b_, reward.mutable_data(), done.mutable_data(), flat.mutable_data(), d[0], modular.mutable_data(), d[1],
                                        other.mutable_data(), d[2], reset_done ? 1 : 0));
            out["reward"] = reward; out["done"] = done;
        } else {
            chk(mevo_batch_task_reset(b_, flat.mutable_data(), d[0], modular.mutable_data(), d[1], other.mutable_data(), d[2]));
        }
        out["flat"] = flat; out["modular"] = modular; out["other"] = other;
        return out;
    }
    py::array_t<double> state(int env, int which) {
        const int P = chk(mevo_env_num_points(b_, env));
        std::vector<double> buf(2 * (size_t)P);
        chk(mevo_env_get_state(b_, env, which == 0 ? buf.data() : nullptr, which == 1 ? buf.data() : nullptr, nullptr, nullptr));
        return as_2xn(buf, 0, P - 1);
    }
    mevo_batch* b_ = nullptr;
};

}  // namespace

PYBIND11_MODULE(simulator_cpp, m) {
    m.doc() = "EvoGym simulator API (Sim, Viewer, Camera) on the mevo_b200 CUDA engine, plus BatchSim";

    py::class_<Viewer>(m, "Viewer")
        .def(py::init<Sim*>())
        .def("init", &Viewer::init)
        .def("render", &Viewer::render, py::arg("camera"), py::arg("hide_background") = false, py::arg("hide_grid") = false,
             py::arg("hide_edges") = false, py::arg("hide_boxels") = false, py::arg("dont_clear") = false)
        .def("show_debug_window", &Viewer::show_debug_window)
        .def("hide_debug_window", &Viewer::hide_debug_window)
        .def("get_debug_window_pos", &Viewer::get_debug_window_pos);

    py::class_<Sim>(m, "Sim")
        .def(py::init<>())
        .def("init", &Sim::init)
        .def_static("get_version", &Sim::get_version)
        .def("read_object_from_file", &Sim::read_object_from_file)
        .def("read_robot_from_file", &Sim::read_robot_from_file)
        .def("read_object_from_array", &Sim::read_object_from_array)
        .def("read_robot_from_array", &Sim::read_robot_from_array)
        .def("step", &Sim::step)
        .def("set_action", &Sim::set_action)
        .def("revert", &Sim::revert)
        .def("get_time", &Sim::get_time)
        .def("pos_at_time", &Sim::pos_at_time)
        .def("vel_at_time", &Sim::vel_at_time)
        .def("object_pos_at_time", &Sim::object_pos_at_time)
        .def("object_vel_at_time", &Sim::object_vel_at_time)
        .def("object_orientation_at_time", &Sim::object_orientation_at_time)
        .def("translate_object", &Sim::translate_object)
        .def("get_indices_of_actuators", &Sim::get_indices_of_actuators);

    py::class_<Camera>(m, "Camera")
        .def(py::init<bool>())
        .def("set_pos", &Camera::set_pos, py::arg("x"), py::arg("y"))
        .def("set_size", &Camera::set_size, py::arg("x"), py::arg("y"))
        .def("set_resolution", &Camera::set_resolution, py::arg("x"), py::arg("y"))
        .def("get_resolution_width", &Camera::get_resolution_width)
        .def("get_resolution_height", &Camera::get_resolution_height)
        .def("get_image", &Camera::get_image);

    py::class_<BatchSim>(m, "BatchSim")
        .def(py::init<int>(), py::arg("device") = -1)
        .def("begin_env", &BatchSim::begin_env)
        .def("add_object", &BatchSim::add_object, py::arg("env"), py::arg("grid"), py::arg("connections"), py::arg("name"),
             py::arg("x"), py::arg("y"), py::arg("is_robot"))
        .def("finalize", &BatchSim::finalize)
        .def_property_readonly("num_envs", &BatchSim::num_envs)
        .def_property_readonly("action_stride", &BatchSim::action_stride)
        .def("actuator_indices", &BatchSim::actuator_indices)
        .def("set_actions", &BatchSim::set_actions)
        .def("set_actions_device", &BatchSim::set_actions_device)
        .def("bind_actions_device", &BatchSim::bind_actions_device)
        .def("set_auto_reset", &BatchSim::set_auto_reset)
        .def("step", &BatchSim::step, py::arg("n_steps") = 1)
        .def("step_async", &BatchSim::step_async, py::arg("n_steps") = 1)
        .def("sync", &BatchSim::sync)
        .def("reset", &BatchSim::reset, py::arg("env_ids") = py::none())
        .def("get_time", &BatchSim::get_time)
        .def("pos", &BatchSim::pos)
        .def("vel", &BatchSim::vel)
        .def("object_com", &BatchSim::object_com)
        .def("object_com_device", &BatchSim::object_com_device)
        .def("object_obs", &BatchSim::object_obs)
        .def("object_obs_device", &BatchSim::object_obs_device)
        .def("object_orientation", &BatchSim::object_orientation)
        .def("object_orientation_device", &BatchSim::object_orientation_device)
        .def("task_configure", &BatchSim::task_configure, py::arg("task"), py::arg("robot") = "robot", py::arg("package") = py::none(),
             py::arg("max_episode_steps") = 0)
        .def("task_dims", &BatchSim::task_dims)
        .def("task_env_tables", &BatchSim::task_env_tables)
        .def("task_apply_policy", &BatchSim::task_apply_policy, py::arg("actions"), py::arg("per_voxel") = false)
        .def("task_apply_policy_device", &BatchSim::task_apply_policy_device)
        .def("task_outputs", &BatchSim::task_outputs, py::arg("reset_done") = true)
        .def("task_outputs_device", &BatchSim::task_outputs_device)
        .def("task_set_goals", &BatchSim::task_set_goals, py::arg("goals"), py::arg("env_first") = 0)
        .def("task_bind_flag_planes", &BatchSim::task_bind_flag_planes)
        .def("task_draw_goals", &BatchSim::task_draw_goals)
        .def("task_get_goals", &BatchSim::task_get_goals)
        .def("task_reset", &BatchSim::task_reset)
        .def("task_reset_device", &BatchSim::task_reset_device)
        .def_property_readonly("cuda_stream", &BatchSim::cuda_stream)
        .def_property_readonly("unstable_events", &BatchSim::unstable_events);
}
