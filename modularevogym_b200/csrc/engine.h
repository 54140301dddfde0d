// Batched soft-body engine: host-side owner of the device buffers (see engine.cu for the kernels).
//
// One Batch = N independent environments (each what the reference calls one `Sim`,
// SimulatorCPP/Sim.h:12-58) resident in HBM on one GPU.  No torch types, no exceptions across the
// C-ABI (c_api.cpp catches and converts to status codes).
#pragma once

#include <cstdint>
#include <string>
#include <vector>

#include "topology.h"

namespace mevo {

struct EnvHost {
    World world;
    // full-state host mirror in the reference's global numbering
    std::vector<double> pos, vel, pos_last, vel_true;   // 2P interleaved
    std::vector<double> goal;                            // E  (length_eq_goal)
    std::vector<uint8_t> obj_force_dynamic;              // per object: simulate even if all-fixed
    // packing results
    std::vector<int> point_local;    // global point -> dynamic local id, or -(static local id)-1
    std::vector<int> edge_local;     // global edge  -> dynamic local id, or -1
    std::vector<int> dyn_points;     // dynamic local id -> global point
    std::vector<int> dyn_edges;      // dynamic local id -> global edge
    std::vector<int> always_pairs;   // dynamic leaf pairs (a, b, a, b, ...) that share a point: candidates in every substep
    std::vector<int> act_slot_obj;   // action slot -> object index
    std::vector<int> act_slot_boxel; // action slot -> boxel index in that object
    int design_id = -1, static_id = -1;
    size_t state_off = 0;            // doubles into the state buffer
    int n_act = 0;
    int cls = 0;                     // size class (index into kGroupSizes)
    long sim_time = 0;
    bool host_fresh = true;          // host mirror newer than device?
};

struct StepStats {
    int launches = 0;                // kernel launches issued by the last step() call
};

class Batch {
public:
    explicit Batch(int device);
    ~Batch();
    Batch(const Batch&) = delete;
    Batch& operator=(const Batch&) = delete;

    int device() const { return device_; }
    int num_envs() const { return (int)envs_.size(); }
    bool finalized() const { return finalized_; }
    int finalize_epoch() const { return finalize_epoch_; }      // bumped by every finalize(): device pointers and graphs of older epochs are stale

    int begin_env(int world_w, int world_h);
    bool add_object(int env, const double* grid, int w, int h, const double* conn, int k,
                    const std::string& name, double x, double y, bool is_robot);
    void finalize();   // (re)pack every environment and upload; current state is preserved

    const World& world(int env) const { return envs_.at(env).world; }
    const EnvHost& env_host(int env) const { return envs_.at(env); }
    int act_stride() const { return act_stride_; }
    int num_actuators(int env) const { return envs_.at(env).n_act; }

    // dense actions: values[env * stride + slot]; slot order = robots in load order, actuators in
    // Robot::init order.  Applied by the next step(apply_actions = true).
    void set_actions(const double* host_values, int stride);
    void set_actions_device(const double* dev_values, int stride);
    double* actions_device_ptr() { return d_act_; }
    // zero-copy: subsequent steps read actions straight from this device buffer (nullptr = internal buffer)
    void bind_actions_device(const double* dev_values, int stride);
    // unstable environments are reverted to snapshot 0 inside the step kernel (flag still reported)
    void set_auto_reset(bool on) { auto_reset_ = on; }
    unsigned long long unstable_events();
    // reference semantics (PhysicsEngine::update_actuator_goals): partial (grid_index, value) list
    bool env_set_action(int env, const std::string& robot, const double* action_n_by_2, int n);

    // n_steps env-steps of n_substeps substeps each.  Returns the number of environments that
    // reported "unstable" in the LAST of the n_steps; per-env flags land in unstable_host (may be
    // null).  An unstable env-step does not advance that env's time (Sim.cpp:105-115).
    int step(int n_steps, bool apply_actions, uint8_t* unstable_host, int n_substeps = 30);
    void step_async(int n_steps, bool apply_actions, int n_substeps = 30);   // no host sync
    void sync();
    // allocation-free variant of step_async (no counterpart in the reference; engine.cu: g_static_scratch)
    void step_static_scratch(int n_steps, const double* dev_actions, int stride, uint8_t* dev_unstable_out);
    const uint8_t* unstable_device_ptr() const { return d_unstable_; }
    const int* time_device_ptr() const { return d_time_; }

    // full state access in the reference's numbering, (P,2) row-major
    void get_state(int env, double* pos, double* vel, double* pos_last, double* vel_true);
    void set_state(int env, const double* pos, const double* vel, const double* pos_last, const double* vel_true);
    void get_goal(int env, double* goal);
    void set_goal(int env, const double* goal);
    long get_time(int env);
    void set_time(int env, long t);
    void translate_object(int env, const std::string& name, double dx, double dy);

    // snapshot 0 of every env (state right after finalize, or after save_reset_state())
    void save_reset_state();
    void reset_envs(const int* env_ids, int n);     // Environment::revert_to_snapshot(0) semantics
    void reset_all();
    // batched snapshot history: slot 0 = the reset snapshot, slots 1 .. 7 the caller's (device-resident copies)
    void save_snapshot(int slot);                                   // Environment::save_snapshot of every environment
    void revert_envs(const int* env_ids, int n, int slot);          // Environment::revert_to_snapshot; env_ids nullptr = all
    long read_snapshot(int env, int slot, double* pos, double* vel); // Sim::pos_at_time / vel_at_time; returns the slot's time or -1

    // packed robot positions for the learner side: COM (x, y) of object `obj_index` per env
    void object_com(int obj_index, double* out_host_2n);
    void object_com_device(int obj_index, double* out_device_2n);   // async on stream()
    // flags of the last step + COM of `obj_index`, one synchronisation; returns the number of unstable environments
    int fetch_step_results(uint8_t* unstable_host, int obj_index, double* com_host_2n);
    // Environment::object_orientation_at_time of object `obj_index` per env: current state against snapshot 0
    void object_orientation(int obj_index, double* out_host_n);
    void object_orientation_device(int obj_index, double* out_device_n);   // async on stream()
    // out[env * row_stride]; state_like: a device buffer laid out like the state (the task layer's saved snapshot), nullptr = the live state
    void object_orientation_strided_device(int obj_index, void* out, int row_stride, bool is_f32, const double* state_like = nullptr);
    // Walker-style observation rows [vel com (2), relative x (n), relative y (n), zero padding] (evogym/envs/walk.py:58-61)
    int object_obs_width(int obj_index) const;                       // 2 + 2 * (largest point count of the object)
    void object_obs_device(int obj_index, double* out_device, int stride);   // async on stream()
    void object_obs(int obj_index, double* out_host, int stride);

    // ---- task layer (task.cu): what the reference's gym envs compute either side of Sim::step, on the device.
    // task: 0 Walker-v0 (walk.py WalkingFlat), 1 Pusher-v0 (manipulate.py PushSmallRect), 2 Thrower-v0 (ThrowSmallRect)
    void task_configure(int task, const std::string& robot, const std::string& package, int max_episode_steps);
    void task_dims(int* dims /* flat width, modular width, other width, max voxels, action stride */);
    // per env host tables: obs_index_array (n_vox x 4: TL,TR,BL,BR), act_mask, obs_mask; returns voxel_corner_size
    int task_env_tables(int env, int* index_array, int* act_mask, int* obs_mask, int cap_voxels);
    // policy output (one value per actuator slot, or per voxel in modular mode) -> +1, clip [0.6, 1.6] -> action buffer
    void task_apply_policy_device(const void* actions, int stride, bool per_voxel, bool is_f32);
    void task_apply_policy(const double* host_actions, int stride, bool per_voxel);
    // after step(): reward, done flags (bit 0 done, bit 1 truncated by the step limit, bit 2 unstable), observations of the
    // resulting state; reset_done: environments that are done are reverted to snapshot 0 first (vec-env semantics)
    void task_outputs_device(void* reward, unsigned char* done, void* flat, int flat_stride, void* modular, int modular_stride,
                             void* other, int other_stride, bool is_f32, bool reset_done);
    void task_outputs(double* reward, unsigned char* done, double* flat, int flat_stride, double* modular, int modular_stride,
                      double* other, int other_stride, bool reset_done);
    void task_reset_device(void* flat, int flat_stride, void* modular, int modular_stride, void* other, int other_stride, bool is_f32);
    void task_reset(double* flat, int flat_stride, double* modular, int modular_stride, double* other, int other_stride);
    // BidirectionalWalker-v0 (multi_goal.py:91-205): per-environment goal lists -- from the host, or drawn on the device
    void task_launch_sources(int* out);                 // Catcher-v0: whose launch position each environment runs now (n ints)
    void task_bind_flag_planes(void* device_planes);    // 3 x n bytes: done, truncated, unstable as 0 / 1 (nullptr: off)
    void task_set_goals(int env_first, int n_envs, const int* counts, const int* goals, int stride);
    void task_draw_goals(unsigned long long seed);
    int task_get_goals(int env, int* out, int cap);      // returns count | current << 16

    // get_floor_obs / get_ceil_obs (evogym/envs/base.py:269-372) of object `obj_index` against the named terrain objects
    void sight_obs_device(int obj_index, const std::vector<std::string>& terrain, int sight_dist, double sight_range, bool ceil,
                          void* out, int stride, bool is_f32);                       // async on stream()
    // as above into a column block of a wider row: `width` columns written (zero padded), rows `row_stride` elements apart
    void sight_obs_strided_device(int obj_index, const std::vector<std::string>& terrain, int sight_dist, double sight_range, bool ceil,
                                  void* out, int width, int row_stride, bool is_f32, const int* env_off = nullptr,   // env_off: device, + elements per env
                                  const double* state_like = nullptr);
    void sight_obs(int obj_index, const std::vector<std::string>& terrain, int sight_dist, double sight_range, bool ceil,
                   double* out_host, int stride);

    // headless rasteriser (render.cu): Interface::render of one environment into an RGB8 image (res_h x res_w x 3, row 0 = top).
    // flags: 1 hide_background, 2 hide_grid, 4 hide_edges, 8 hide_boxels, 16 dont_clear
    void render_device(int env, double cam_x, double cam_y, double cam_w, double cam_h, int res_w, int res_h, int flags,
                       unsigned char* out_device);
    void render(int env, double cam_x, double cam_y, double cam_w, double cam_h, int res_w, int res_h, int flags, unsigned char* out_host);

    // ---- CUDA graphs: everything enqueued on the batch's stream between capture_begin() and capture_end() (policy actions, the
    // step launches of all size classes with their fork/join, task outputs, observations ...) becomes one graph; graph_launch
    // replays it with one driver call.  The device pointers passed to the captured calls are baked in: use persistent buffers.
    // Every captured entry point must have run once before (first calls configure kernels and allocate tables).
    void capture_begin();
    int capture_end();                 // returns a graph id
    void graph_launch(int id);
    void graph_destroy(int id);

    // debug / test hooks: candidate pairs of the first substep of the next step() call
    void enable_candidate_dump(int cap_per_env);
    int read_candidates(int env, int* out_pairs /* cap x 4: kind_a, leaf_a, kind_b, leaf_b */, int cap);
    int error_flags(int env);
    // collision_forces of the first substep of subsequent step calls (PhysicsEngine.cpp:715-751), (P, 2) in global numbering
    void enable_force_dump();
    void read_collision_forces(int env, double* out_2P);
    void leaf_info(int env, int dynamic, int leaf, int* obj, int* boxel) const;

    StepStats last_stats() const { return stats_; }
    void* stream() const { return stream_; }
    size_t device_bytes() const { return device_bytes_; }

private:
    struct Impl;
    struct Task;
    struct Sight;
    void task_build();
    void task_free();
    void sight_free();
    void upload_state(int env);
    void download_state(int env);
    void download_all();
    void bind_object_ranges(int obj_index);
    void pack_env(int env, std::vector<uint8_t>& design_blob, std::vector<uint8_t>& static_blob);

    int device_;
    bool finalized_ = false;
    std::vector<EnvHost> envs_;
    int act_stride_ = 0;
    StepStats stats_;
    size_t device_bytes_ = 0;

    // device buffers
    uint8_t* d_design_ = nullptr;
    uint8_t* d_static_ = nullptr;
    double* d_state_ = nullptr;
    double* d_state0_ = nullptr;
    double* d_act_ = nullptr;
    void* d_envdesc_ = nullptr;
    int* d_class_envs_ = nullptr;
    uint8_t* d_unstable_ = nullptr;
    int* d_time_ = nullptr;
    int* d_err_ = nullptr;
    unsigned long long* d_events_ = nullptr;
    const double* d_act_ext_ = nullptr;
    int act_ext_stride_ = 0;
    bool auto_reset_ = false;
    unsigned int* d_cand_ = nullptr;
    int* d_cand_n_ = nullptr;
    int cand_cap_ = 0;
    size_t state_doubles_ = 0;
    void* stream_ = nullptr;
    Impl* impl_ = nullptr;
    std::vector<void*> graphs_;        // cudaGraphExec_t, index = graph id (nullptr after destroy)
    bool capturing_ = false;
    Task* task_ = nullptr;
    Sight* sight_ = nullptr;
    int finalize_epoch_ = 0;
    bool force_dump_ = false;
};

// transform_to_modular_obs (evogym/envs/base.py:427-466), host only: corner indices per cell (h*w*4: TL,TR,BL,BR), returns loc_idx
int modular_corner_index(const std::vector<int>& body, int h, int w, std::vector<int>& out);

}  // namespace mevo
