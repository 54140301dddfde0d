// extern "C" surface of libmevo_b200.so (declared in include/mevo_b200.h).
// Every entry point catches C++ exceptions and converts them to MEVO_ERR + mevo_last_error().
#include "../../include/mevo_b200.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <exception>
#include <stdexcept>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "engine.h"

using mevo::Batch;

struct mevo_batch {
    Batch impl;
    explicit mevo_batch(int device) : impl(device) {}
};

struct mevo_world {
    mevo::World impl;
};

namespace {
thread_local std::string g_err;

template <class F>
int guard(F&& f) {
    try {
        return f();
    } catch (const std::exception& ex) {
        g_err = ex.what();
    } catch (...) {
        g_err = "unknown C++ exception";
    }
    return MEVO_ERR;
}

void fill_object_info(const mevo::World& w, int obj, bool is_static, int* info) {
    if (obj < 0 || obj >= (int)w.objects.size()) throw std::out_of_range("object index out of range");
    const mevo::VoxelObject& o = w.objects[obj];
    info[0] = o.is_robot ? 1 : 0;
    info[1] = o.min_point; info[2] = o.max_point;
    info[3] = (int)o.boxels.size();
    info[4] = o.grid_w; info[5] = o.grid_h;
    info[6] = (int)o.surface_boxels.size();
    info[7] = is_static ? 1 : 0;
}

int fill_object_boxels(const mevo::World& w, int obj, int* out) {
    if (obj < 0 || obj >= (int)w.objects.size()) throw std::out_of_range("object index out of range");
    const mevo::VoxelObject& o = w.objects[obj];
    for (size_t j = 0; j < o.boxels.size(); j++) {
        const mevo::Boxel& bx = o.boxels[j];
        int* r = out + 18 * j;
        r[0] = bx.cell_type; r[1] = bx.grid_index;
        for (int q = 0; q < 4; q++) { r[2 + q] = bx.pt[q]; r[6 + q] = bx.ed[q]; }
        for (int q = 0; q < 8; q++) r[10 + q] = bx.nei[q] ? 1 : 0;
    }
    return (int)o.boxels.size();
}

int fill_surface_boxels(const mevo::World& w, int obj, int* out) {
    if (obj < 0 || obj >= (int)w.objects.size()) throw std::out_of_range("object index out of range");
    const mevo::VoxelObject& o = w.objects[obj];
    for (size_t j = 0; j < o.surface_boxels.size(); j++) out[j] = o.surface_boxels[j];
    return (int)o.surface_boxels.size();
}

int fill_actuator_indices(const mevo::World& w, const char* robot, int* out, int cap) {
    int i = w.find_object(robot);
    if (i < 0 || !w.objects[i].is_robot) throw std::invalid_argument(std::string(robot) + " is not a robot");
    const auto& idx = w.objects[i].actuator_grid_index;
    for (size_t j = 0; j < idx.size() && (int)j < cap; j++) out[j] = idx[j];
    return (int)idx.size();
}

void fill_topology(const mevo::World& w, double* pos0, double* mass, unsigned char* fixed, int* edge_a, int* edge_b,
                   double* edge_k, double* edge_l0, unsigned char* edge_on_surface) {
    const size_t P = w.num_points(), E = w.num_edges();
    if (pos0) std::memcpy(pos0, w.pos.data(), sizeof(double) * 2 * P);
    if (mass) std::memcpy(mass, w.mass.data(), sizeof(double) * P);
    if (fixed) std::memcpy(fixed, w.fixed.data(), P);
    if (edge_a) std::memcpy(edge_a, w.ea.data(), sizeof(int) * E);
    if (edge_b) std::memcpy(edge_b, w.eb.data(), sizeof(int) * E);
    if (edge_k) std::memcpy(edge_k, w.ek.data(), sizeof(double) * E);
    if (edge_l0) std::memcpy(edge_l0, w.el0.data(), sizeof(double) * E);
    if (edge_on_surface) std::memcpy(edge_on_surface, w.on_surface.data(), E);
}
}  // namespace

extern "C" {

int mevo_abi_version(void) { return MEVO_ABI_VERSION; }
const char* mevo_last_error(void) { return g_err.c_str(); }

int mevo_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { g_err = std::string("CUDA unusable: ") + cudaGetErrorString(e); return MEVO_ERR; }
    return n;
}

mevo_batch* mevo_batch_create(int device) {
    try {
        int n = 0;
        cudaError_t e = cudaGetDeviceCount(&n);
        if (e != cudaSuccess || n <= 0)
            throw std::runtime_error(std::string("mevo_b200 needs a CUDA device (sm_100a); none usable: ") +
                                     (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                                     " -- there is no CPU fallback");
        if (device < 0 || device >= n) throw std::runtime_error("device index out of range");
        return new mevo_batch(device);
    } catch (const std::exception& ex) {
        g_err = ex.what();
    } catch (...) {
        g_err = "unknown C++ exception";
    }
    return nullptr;
}

void mevo_batch_destroy(mevo_batch* b) { delete b; }

int mevo_batch_begin_env(mevo_batch* b, int world_w, int world_h) {
    return guard([&] { return b->impl.begin_env(world_w, world_h); });
}

int mevo_batch_add_object(mevo_batch* b, int env, const double* grid, int w, int h, const double* conn, int k,
                          const char* name, double x, double y, int is_robot) {
    return guard([&] {
        if (!grid || w <= 0 || h <= 0 || k < 0 || (k > 0 && !conn) || !name) throw std::invalid_argument("bad add_object arguments");
        return b->impl.add_object(env, grid, w, h, conn, k, name, x, y, is_robot != 0) ? 1 : 0;
    });
}

int mevo_batch_add_robot_envs(mevo_batch* b, int n_envs, int world_w, int world_h, int n_terrain, const double* terrain_grids,
                              const int* terrain_wh, const double* terrain_conn, const int* terrain_k, const char* terrain_names,
                              const double* terrain_xy, const double* robot_grids, const int* robot_wh, const double* robot_conn,
                              const int* robot_k, const char* robot_name, double rx, double ry) {
    return guard([&] {
        if (n_envs < 0 || n_terrain < 0 || !robot_grids || !robot_wh || !robot_k || !robot_name ||
            (n_terrain > 0 && (!terrain_grids || !terrain_wh || !terrain_k || !terrain_names || !terrain_xy)))
            throw std::invalid_argument("bad add_robot_envs arguments");
        std::vector<std::string> names;
        {
            std::string all = n_terrain > 0 ? terrain_names : "";
            size_t at = 0;
            for (int t = 0; t < n_terrain; t++) {
                const size_t nl = all.find('\n', at);
                names.push_back(all.substr(at, nl == std::string::npos ? std::string::npos : nl - at));
                if (names.back().empty()) throw std::invalid_argument("add_robot_envs: fewer terrain names than objects");
                at = nl == std::string::npos ? all.size() : nl + 1;
            }
        }
        // environments are independent host worlds: create them, then fill them on all host cores
        std::vector<size_t> g_off(n_envs + 1, 0), c_off(n_envs + 1, 0);
        for (int i = 0; i < n_envs; i++) {
            const int w = robot_wh[2 * i], h = robot_wh[2 * i + 1], k = robot_k[i];
            if (w <= 0 || h <= 0 || k < 0 || (k > 0 && !robot_conn)) throw std::invalid_argument("bad robot in add_robot_envs");
            g_off[i + 1] = g_off[i] + (size_t)w * h;
            c_off[i + 1] = c_off[i] + 2 * (size_t)k;
        }
        int first = -1;
        for (int i = 0; i < n_envs; i++) {
            const int env = b->impl.begin_env(std::max(world_w, (int)rx + robot_wh[2 * i]), std::max(world_h, (int)ry + robot_wh[2 * i + 1]));
            if (first < 0) first = env;
        }
        std::mutex err_mu;
        std::string err;
        auto fill = [&](int lo, int hi) {
            for (int i = lo; i < hi; i++) {
                std::string why;
                try {
                    const int env = first + i;
                    if (!b->impl.add_object(env, robot_grids + g_off[i], robot_wh[2 * i], robot_wh[2 * i + 1], robot_conn + c_off[i], robot_k[i],
                                            robot_name, rx, ry, true))
                        why = "add_robot_envs: the robot of environment " + std::to_string(i) + " was refused";
                    const double* tg = terrain_grids;
                    const double* tc = terrain_conn;
                    for (int t = 0; t < n_terrain && why.empty(); t++) {
                        const int tw = terrain_wh[2 * t], th = terrain_wh[2 * t + 1], tk = terrain_k[t];
                        if (!b->impl.add_object(env, tg, tw, th, tc, tk, names[t], terrain_xy[2 * t], terrain_xy[2 * t + 1], false))
                            why = "add_robot_envs: object " + names[t] + " of environment " + std::to_string(i) + " was refused";
                        tg += (size_t)tw * th;
                        tc += 2 * (size_t)tk;
                    }
                } catch (const std::exception& e) {
                    why = e.what();
                }
                if (!why.empty()) {
                    std::lock_guard<std::mutex> lock(err_mu);
                    if (err.empty()) err = why;
                    return;
                }
            }
        };
        const int n_thr = std::max(1, std::min({(int)std::thread::hardware_concurrency(), 16, n_envs / 64}));
        if (n_thr <= 1) {
            fill(0, n_envs);
        } else {
            std::vector<std::thread> pool;
            for (int t = 0; t < n_thr; t++)
                pool.emplace_back(fill, (int)((long long)n_envs * t / n_thr), (int)((long long)n_envs * (t + 1) / n_thr));
            for (auto& th : pool) th.join();
        }
        if (!err.empty()) throw std::runtime_error(err);
        return first < 0 ? b->impl.num_envs() : first;
    });
}

int mevo_batch_finalize(mevo_batch* b) {
    return guard([&] { b->impl.finalize(); return 0; });
}

int mevo_batch_finalize_epoch(const mevo_batch* b) {
    return b ? b->impl.finalize_epoch() : MEVO_ERR;
}

int mevo_batch_num_envs(const mevo_batch* b) { return b->impl.num_envs(); }
int mevo_env_num_points(const mevo_batch* b, int env) { return guard([&] { return b->impl.world(env).num_points(); }); }
int mevo_env_num_edges(const mevo_batch* b, int env) { return guard([&] { return b->impl.world(env).num_edges(); }); }
int mevo_env_num_objects(const mevo_batch* b, int env) { return guard([&] { return (int)b->impl.world(env).objects.size(); }); }

int mevo_env_object_index(const mevo_batch* b, int env, const char* name) {
    return guard([&] {
        int i = b->impl.world(env).find_object(name);
        if (i < 0) throw std::invalid_argument(std::string("no object named ") + name);
        return i;
    });
}

int mevo_env_object_info(const mevo_batch* b, int env, int obj, int* info) {
    return guard([&] {
        const mevo::World& w = b->impl.world(env);
        if (obj < 0 || obj >= (int)w.objects.size()) throw std::out_of_range("object index out of range");
        const mevo::VoxelObject& o = w.objects[obj];
        fill_object_info(w, obj, o.all_fixed && !o.is_robot && !b->impl.env_host(env).obj_force_dynamic[obj], info);
        return 0;
    });
}

int mevo_env_object_boxels(const mevo_batch* b, int env, int obj, int* out) {
    return guard([&] { return fill_object_boxels(b->impl.world(env), obj, out); });
}

int mevo_env_object_surface_boxels(const mevo_batch* b, int env, int obj, int* out) {
    return guard([&] { return fill_surface_boxels(b->impl.world(env), obj, out); });
}

int mevo_env_actuator_indices(const mevo_batch* b, int env, const char* robot, int* out, int cap) {
    return guard([&] { return fill_actuator_indices(b->impl.world(env), robot, out, cap); });
}

int mevo_env_get_topology(const mevo_batch* b, int env, double* pos0, double* mass, unsigned char* fixed, int* edge_a,
                          int* edge_b, double* edge_k, double* edge_l0, unsigned char* edge_on_surface) {
    return guard([&] {
        fill_topology(b->impl.world(env), pos0, mass, fixed, edge_a, edge_b, edge_k, edge_l0, edge_on_surface);
        return 0;
    });
}

/* ---- host-only world builder (no GPU needed) ------------------------------------------------ */
mevo_world* mevo_world_create(int world_w, int world_h) {
    try {
        mevo_world* w = new mevo_world();
        w->impl.init(world_w, world_h);
        return w;
    } catch (const std::exception& ex) {
        g_err = ex.what();
    }
    return nullptr;
}

void mevo_world_destroy(mevo_world* w) { delete w; }

int mevo_world_add_object(mevo_world* w, const double* grid, int gw, int gh, const double* conn, int k, const char* name,
                          double x, double y, int is_robot) {
    return guard([&] {
        if (!grid || gw <= 0 || gh <= 0 || k < 0 || (k > 0 && !conn) || !name) throw std::invalid_argument("bad add_object arguments");
        return w->impl.add_object(grid, gw, gh, conn, k, name, x, y, is_robot != 0) ? 1 : 0;
    });
}

int mevo_world_num_points(const mevo_world* w) { return w->impl.num_points(); }
int mevo_world_num_edges(const mevo_world* w) { return w->impl.num_edges(); }
int mevo_world_num_objects(const mevo_world* w) { return (int)w->impl.objects.size(); }

int mevo_world_object_info(const mevo_world* w, int obj, int* info) {
    return guard([&] {
        if (obj < 0 || obj >= (int)w->impl.objects.size()) throw std::out_of_range("object index out of range");
        const mevo::VoxelObject& o = w->impl.objects[obj];
        fill_object_info(w->impl, obj, o.all_fixed && !o.is_robot, info);
        return 0;
    });
}

int mevo_world_object_boxels(const mevo_world* w, int obj, int* out) {
    return guard([&] { return fill_object_boxels(w->impl, obj, out); });
}

int mevo_world_object_surface_boxels(const mevo_world* w, int obj, int* out) {
    return guard([&] { return fill_surface_boxels(w->impl, obj, out); });
}

int mevo_world_actuator_indices(const mevo_world* w, const char* robot, int* out, int cap) {
    return guard([&] { return fill_actuator_indices(w->impl, robot, out, cap); });
}

int mevo_world_get_topology(const mevo_world* w, double* pos0, double* mass, unsigned char* fixed, int* edge_a, int* edge_b,
                            double* edge_k, double* edge_l0, unsigned char* edge_on_surface) {
    return guard([&] {
        fill_topology(w->impl, pos0, mass, fixed, edge_a, edge_b, edge_k, edge_l0, edge_on_surface);
        return 0;
    });
}

int mevo_env_set_action(mevo_batch* b, int env, const char* robot, const double* action, int n) {
    return guard([&] { return b->impl.env_set_action(env, robot, action, n) ? 1 : 0; });
}

int mevo_batch_action_stride(const mevo_batch* b) { return b->impl.act_stride(); }

int mevo_batch_set_actions(mevo_batch* b, const double* host_values, int stride) {
    return guard([&] { b->impl.set_actions(host_values, stride); return 0; });
}

int mevo_batch_set_actions_device(mevo_batch* b, const double* device_values, int stride) {
    return guard([&] { b->impl.set_actions_device(device_values, stride); return 0; });
}

int mevo_batch_bind_actions_device(mevo_batch* b, const double* device_values, int stride) {
    return guard([&] { b->impl.bind_actions_device(device_values, stride); return 0; });
}

int mevo_batch_set_auto_reset(mevo_batch* b, int on) {
    return guard([&] { b->impl.set_auto_reset(on != 0); return 0; });
}

long long mevo_batch_unstable_events(mevo_batch* b) {
    long long v = MEVO_ERR;
    guard([&] { v = (long long)b->impl.unstable_events(); return 0; });
    return v;
}

int mevo_batch_step(mevo_batch* b, int n_steps, int apply_actions, unsigned char* unstable_out) {
    return guard([&] { return b->impl.step(n_steps, apply_actions != 0, unstable_out, 30); });
}

int mevo_batch_step_substeps(mevo_batch* b, int n_steps, int n_substeps, int apply_actions, unsigned char* unstable_out) {
    return guard([&] { return b->impl.step(n_steps, apply_actions != 0, unstable_out, n_substeps); });
}

int mevo_batch_step_async(mevo_batch* b, int n_steps, int apply_actions) {
    return guard([&] { b->impl.step_async(n_steps, apply_actions != 0, 30); return 0; });
}

int mevo_batch_step_static_scratch(mevo_batch* b, int n_steps, const double* device_actions, int stride, unsigned char* device_unstable_out) {
    return guard([&] { b->impl.step_static_scratch(n_steps, device_actions, stride, device_unstable_out); return 0; });
}

int mevo_batch_sync(mevo_batch* b) {
    return guard([&] { b->impl.sync(); return 0; });
}

int mevo_batch_last_launches(const mevo_batch* b) { return b->impl.last_stats().launches; }

long mevo_env_get_time(mevo_batch* b, int env) {
    long t = MEVO_ERR;
    guard([&] { t = b->impl.get_time(env); return 0; });
    return t;
}

int mevo_env_set_time(mevo_batch* b, int env, long t) {
    return guard([&] { b->impl.set_time(env, t); return 0; });
}

int mevo_env_get_state(mevo_batch* b, int env, double* pos, double* vel, double* pos_last, double* vel_true) {
    return guard([&] { b->impl.get_state(env, pos, vel, pos_last, vel_true); return 0; });
}

int mevo_env_set_state(mevo_batch* b, int env, const double* pos, const double* vel, const double* pos_last,
                       const double* vel_true) {
    return guard([&] { b->impl.set_state(env, pos, vel, pos_last, vel_true); return 0; });
}

int mevo_env_get_goal(mevo_batch* b, int env, double* goal) {
    return guard([&] { b->impl.get_goal(env, goal); return 0; });
}

int mevo_env_set_goal(mevo_batch* b, int env, const double* goal) {
    return guard([&] { b->impl.set_goal(env, goal); return 0; });
}

int mevo_env_translate_object(mevo_batch* b, int env, const char* name, double dx, double dy) {
    return guard([&] { b->impl.translate_object(env, name, dx, dy); return 0; });
}

int mevo_batch_save_reset_state(mevo_batch* b) {
    return guard([&] { b->impl.save_reset_state(); return 0; });
}

int mevo_batch_reset(mevo_batch* b, const int* env_ids, int n) {
    return guard([&] {
        if (!env_ids) b->impl.reset_all();
        else b->impl.reset_envs(env_ids, n);
        return 0;
    });
}

int mevo_batch_save_snapshot(mevo_batch* b, int slot) {
    return guard([&] { b->impl.save_snapshot(slot); return 0; });
}

int mevo_batch_revert(mevo_batch* b, const int* env_ids, int n, int slot) {
    return guard([&] { b->impl.revert_envs(env_ids, n, slot); return 0; });
}

long mevo_env_read_snapshot(mevo_batch* b, int env, int slot, double* pos, double* vel) {
    long t = MEVO_ERR;
    const int rc = guard([&] { t = b->impl.read_snapshot(env, slot, pos, vel); return 0; });
    return rc < 0 ? (long)MEVO_ERR - 1 : t;
}

int mevo_batch_object_com(mevo_batch* b, int obj, double* out) {
    return guard([&] { b->impl.object_com(obj, out); return 0; });
}

int mevo_batch_fetch_step_results(mevo_batch* b, unsigned char* unstable_out, int obj, double* com_out) {
    return guard([&] { return b->impl.fetch_step_results(unstable_out, obj, com_out); });
}

int mevo_batch_object_com_device(mevo_batch* b, int obj, double* device_out) {
    return guard([&] {
        if (!device_out) throw std::invalid_argument("object_com_device: null output");
        b->impl.object_com_device(obj, device_out);
        return 0;
    });
}

int mevo_batch_object_obs_width(mevo_batch* b, int obj) {
    int w = -1;
    const int rc = guard([&] { w = b->impl.object_obs_width(obj); return 0; });
    return rc ? -1 : w;
}

int mevo_batch_object_obs(mevo_batch* b, int obj, double* out, int stride) {
    return guard([&] {
        if (!out || stride < 2) throw std::invalid_argument("object_obs: null output or stride < 2");
        b->impl.object_obs(obj, out, stride);
        return 0;
    });
}

int mevo_batch_object_orientation(mevo_batch* b, int obj, double* out) {
    return guard([&] {
        if (!out) throw std::invalid_argument("object_orientation: null output");
        b->impl.object_orientation(obj, out);
        return 0;
    });
}

int mevo_batch_object_orientation_device(mevo_batch* b, int obj, double* device_out) {
    return guard([&] {
        if (!device_out) throw std::invalid_argument("object_orientation_device: null output");
        b->impl.object_orientation_device(obj, device_out);
        return 0;
    });
}

static std::vector<std::string> split_names(const char* csv) {
    std::vector<std::string> out;
    std::string cur;
    for (const char* c = csv ? csv : ""; ; c++) {
        if (*c == ',' || *c == 0) { if (!cur.empty()) out.push_back(cur); cur.clear(); if (*c == 0) break; }
        else cur.push_back(*c);
    }
    return out;
}

int mevo_batch_sight_obs(mevo_batch* b, int obj, const char* terrain_csv, int sight_dist, double sight_range, int ceil, double* out,
                         int stride) {
    return guard([&] {
        if (!out) throw std::invalid_argument("sight_obs: null output");
        b->impl.sight_obs(obj, split_names(terrain_csv), sight_dist, sight_range, ceil != 0, out, stride);
        return 0;
    });
}

int mevo_batch_sight_obs_device(mevo_batch* b, int obj, const char* terrain_csv, int sight_dist, double sight_range, int ceil,
                                void* device_out, int stride, int is_f32) {
    return guard([&] {
        if (!device_out) throw std::invalid_argument("sight_obs_device: null output");
        b->impl.sight_obs_device(obj, split_names(terrain_csv), sight_dist, sight_range, ceil != 0, device_out, stride, is_f32 != 0);
        return 0;
    });
}

int mevo_batch_object_obs_device(mevo_batch* b, int obj, double* device_out, int stride) {
    return guard([&] {
        if (!device_out || stride < 2) throw std::invalid_argument("object_obs_device: null output or stride < 2");
        b->impl.object_obs_device(obj, device_out, stride);
        return 0;
    });
}

// ---- task layer ---------------------------------------------------------------------------------

int mevo_batch_task_configure(mevo_batch* b, int task, const char* robot, const char* package, int max_episode_steps) {
    return guard([&] {
        if (!robot) throw std::invalid_argument("task_configure: robot name is null");
        b->impl.task_configure(task, robot, package ? package : "", max_episode_steps);
        return 0;
    });
}

int mevo_batch_task_dims(mevo_batch* b, int* dims) {
    return guard([&] {
        if (!dims) throw std::invalid_argument("task_dims: null output");
        b->impl.task_dims(dims);
        return 0;
    });
}

int mevo_batch_task_env_tables(mevo_batch* b, int env, int* index_array, int* act_mask, int* obs_mask, int cap_voxels) {
    return guard([&] { return b->impl.task_env_tables(env, index_array, act_mask, obs_mask, cap_voxels); });
}

int mevo_modular_corner_index(const int* body, int h, int w, int* out) {
    return guard([&] {
        if (!body || !out || h <= 0 || w <= 0) throw std::invalid_argument("modular_corner_index: bad arguments");
        std::vector<int> grid(body, body + (size_t)h * w), idx;
        const int loc = mevo::modular_corner_index(grid, h, w, idx);
        std::copy(idx.begin(), idx.end(), out);
        return loc;
    });
}

int mevo_batch_task_apply_policy(mevo_batch* b, const double* host_actions, int stride, int per_voxel) {
    return guard([&] {
        if (!host_actions) throw std::invalid_argument("task_apply_policy: null actions");
        b->impl.task_apply_policy(host_actions, stride, per_voxel != 0);
        return 0;
    });
}

int mevo_batch_task_apply_policy_device(mevo_batch* b, const void* device_actions, int stride, int per_voxel, int is_f32) {
    return guard([&] {
        if (!device_actions) throw std::invalid_argument("task_apply_policy_device: null actions");
        b->impl.task_apply_policy_device(device_actions, stride, per_voxel != 0, is_f32 != 0);
        return 0;
    });
}

int mevo_batch_task_outputs(mevo_batch* b, double* reward, unsigned char* done, double* flat, int flat_stride, double* modular,
                            int modular_stride, double* other, int other_stride, int reset_done) {
    return guard([&] {
        b->impl.task_outputs(reward, done, flat, flat_stride, modular, modular_stride, other, other_stride, reset_done != 0);
        return 0;
    });
}

int mevo_batch_task_outputs_device(mevo_batch* b, void* reward, unsigned char* done, void* flat, int flat_stride, void* modular,
                                   int modular_stride, void* other, int other_stride, int is_f32, int reset_done) {
    return guard([&] {
        b->impl.task_outputs_device(reward, done, flat, flat_stride, modular, modular_stride, other, other_stride, is_f32 != 0,
                                    reset_done != 0);
        return 0;
    });
}

int mevo_batch_task_reset(mevo_batch* b, double* flat, int flat_stride, double* modular, int modular_stride, double* other,
                          int other_stride) {
    return guard([&] {
        b->impl.task_reset(flat, flat_stride, modular, modular_stride, other, other_stride);
        return 0;
    });
}

int mevo_batch_task_reset_device(mevo_batch* b, void* flat, int flat_stride, void* modular, int modular_stride, void* other,
                                 int other_stride, int is_f32) {
    return guard([&] {
        b->impl.task_reset_device(flat, flat_stride, modular, modular_stride, other, other_stride, is_f32 != 0);
        return 0;
    });
}

int mevo_batch_task_launch_sources(mevo_batch* b, int* out) {
    return guard([&] {
        if (!out) throw std::runtime_error("mevo_batch_task_launch_sources: null argument");
        b->impl.task_launch_sources(out);
        return 0;
    });
}

int mevo_batch_task_bind_flag_planes(mevo_batch* b, void* device_planes) {
    return guard([&] {
        b->impl.task_bind_flag_planes(device_planes);
        return 0;
    });
}

int mevo_batch_task_set_goals(mevo_batch* b, int env_first, int n_envs, const int* counts, const int* goals, int stride) {
    return guard([&] {
        if (n_envs > 0 && (!counts || !goals)) throw std::runtime_error("mevo_batch_task_set_goals: null argument");
        b->impl.task_set_goals(env_first, n_envs, counts, goals, stride);
        return 0;
    });
}

int mevo_batch_task_draw_goals(mevo_batch* b, unsigned long long seed) {
    return guard([&] {
        b->impl.task_draw_goals(seed);
        return 0;
    });
}

int mevo_batch_task_get_goals(mevo_batch* b, int env, int* count, int* current, int* goals, int cap) {
    return guard([&] {
        if (!count || !current || (cap > 0 && !goals)) throw std::runtime_error("mevo_batch_task_get_goals: null argument");
        const int r = b->impl.task_get_goals(env, goals, cap);
        *count = r & 0xffff; *current = r >> 16;
        return 0;
    });
}

int mevo_env_render(mevo_batch* b, int env, double cam_x, double cam_y, double cam_w, double cam_h, int res_w, int res_h, int flags,
                    unsigned char* out) {
    return guard([&] {
        if (!out) throw std::invalid_argument("render: null output");
        b->impl.render(env, cam_x, cam_y, cam_w, cam_h, res_w, res_h, flags, out);
        return 0;
    });
}

int mevo_env_render_device(mevo_batch* b, int env, double cam_x, double cam_y, double cam_w, double cam_h, int res_w, int res_h, int flags,
                           unsigned char* device_out) {
    return guard([&] {
        if (!device_out) throw std::invalid_argument("render_device: null output");
        b->impl.render_device(env, cam_x, cam_y, cam_w, cam_h, res_w, res_h, flags, device_out);
        return 0;
    });
}

int mevo_batch_capture_begin(mevo_batch* b) { return guard([&] { b->impl.capture_begin(); return 0; }); }
int mevo_batch_capture_end(mevo_batch* b) { return guard([&] { return b->impl.capture_end(); }); }
int mevo_batch_graph_launch(mevo_batch* b, int graph) { return guard([&] { b->impl.graph_launch(graph); return 0; }); }
int mevo_batch_graph_destroy(mevo_batch* b, int graph) { return guard([&] { b->impl.graph_destroy(graph); return 0; }); }

double* mevo_batch_actions_device_ptr(mevo_batch* b) { return b->impl.actions_device_ptr(); }
const unsigned char* mevo_batch_unstable_device_ptr(mevo_batch* b) { return b->impl.unstable_device_ptr(); }
void* mevo_batch_cuda_stream(mevo_batch* b) { return b->impl.stream(); }
long long mevo_batch_device_bytes(const mevo_batch* b) { return (long long)b->impl.device_bytes(); }

int mevo_batch_enable_candidate_dump(mevo_batch* b, int cap_per_env) {
    return guard([&] { b->impl.enable_candidate_dump(cap_per_env); return 0; });
}

int mevo_env_read_candidates(mevo_batch* b, int env, int* out, int cap) {
    return guard([&] { return b->impl.read_candidates(env, out, cap); });
}

int mevo_batch_enable_force_dump(mevo_batch* b) {
    return guard([&] { b->impl.enable_force_dump(); return 0; });
}

int mevo_env_read_collision_forces(mevo_batch* b, int env, double* out) {
    return guard([&] { b->impl.read_collision_forces(env, out); return 0; });
}

int mevo_env_error_flags(mevo_batch* b, int env) {
    return guard([&] { return b->impl.error_flags(env); });
}

}  // extern "C"
