// Task layer: what the reference's gym environments compute either side of Sim::step, batched on the device
// (SURVEY 8f-1 / 8f-2).  One warp per environment, one launch per env-step, no host round trip:
//
//   policy_actions_kernel   BenchmarkBase.step (evogym/envs/base.py:384-391: action + 1) -> EvoGymBase.step
//                           (base.py:40-45: clip to [0.6, 1.6]); in modular mode the action is one value per
//                           voxel and the actuator voxels are picked in row-major order (walk.py:41-42
//                           `action[~act_mask]`).
//   task_outputs_kernel     reward / done / observation of
//                             WalkingFlat    evogym/envs/walk.py:35-84        (Walker-v0)
//                             PushSmallRect  evogym/envs/manipulate.py:346-398 (Pusher-v0)
//                             ThrowSmallRect manipulate.py:535-600             (Thrower-v0)
//                           flat observation (walk.py:58-61, manipulate.py:22-36) and ModularEvoGym's per-voxel
//                           observation (base.py:427-506 transform_to_modular_obs / modular_ob_wrapper_padding,
//                           base.py:746-753 get_modular_obs); then what the vec-env worker does on `done`
//                           (examples/modular_envs/wrappers/subproc_vec_env.py: reset and return the reset
//                           observation) and gym's TimeLimit.
//
// Unit convention of the reference, kept operation for operation: EvoSim getters divide engine metres by
// VOXEL_SIZE = 0.1 (evogym/sim.py:188-231), BenchmarkBase multiplies by 0.1 again (base.py:393-421).
#include <cuda_runtime.h>

#include <algorithm>
#include <cstddef>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "device_layout.h"
#include "engine.h"

namespace mevo {

#define CUDA_CHECK(expr)                                                                          \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess)                                                                    \
            throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " + \
                                     __FILE__ + ":" + std::to_string(__LINE__) + " (" #expr ")");   \
    } while (0)

namespace {

constexpr double kVoxel = 0.1;          // EvoSim.VOXEL_SIZE, BenchmarkBase.VOXEL_SIZE

struct TaskEnvRec {                      // 32 bytes per environment
    int r_start, r_count;                // robot: first dynamic-local point, point count
    int p_start, p_count;                // package (count 0: task has none)
    int Pd;                              // dynamic points of the env (state section length)
    int vox_off, n_vox;                  // rows of the modular table
    int n_act;
};

constexpr double kPi = 3.14159265358979323846;       // math.pi

struct TaskParams {
    const double* state;
    double* state_rw;
    const double* state0;
    double* saved;                       // the state of the last SAVED snapshot (Sim.cpp:105-115: history[sim_time]), same layout as `state`
    const EnvDesc* envs;
    const TaskEnvRec* tenv;
    const short4* vox_idx;               // per modular row: TL, TR, BL, BR point index inside the robot, -1 = empty voxel
    const unsigned char* vox_mat;        // per modular row: voxel type
    double* prev;                        // per env: robot com x, y, package com x, y at the last observation
    int* episode_len;
    unsigned char* unstable;             // flags of the last step launch (cleared for environments reset here)
    int* sim_time;
    void* reward; unsigned char* done;
    unsigned char* flag_planes;          // optional: 3 x n_envs bytes (0 / 1) -- done, truncated, unstable -- next to the bit field in `done`
    void* flat; int flat_stride;
    void* modular; int modular_stride;
    void* other; int other_stride;
    int n_envs, task, max_episode_steps, reset_done, other_width;
    // orientation rule (tasks 6, 8, 10): orientation of the robot after this step (nullptr: no orientation rule) and at
    // the last observation -- an unstable step saved no snapshot, the environment then reads the old one (Sim.cpp:105-115)
    int goal_axis;                       // 0: no goal, 1: robot com x, 2: robot com y beyond `goal` voxels ends the episode with `goal_bonus`
    double goal, goal_bonus;
    double fall_y;                       // > 0: a robot whose centre of mass is below this height [m] has fallen off: reward - 3, done
    const double* ort_now;
    double* ort_prev;
    double ort_band;
    int* flips;                          // Flipper-v0: whole turns so far (flip.py:34, 78-85)
    double* span_prev;                   // Wingspan / Height maximizer: the robot's extent at the last observation
    int flat_prefix;                     // values in front of the relative positions in the flat observation
    // BidirectionalWalker-v0 (multi_goal.py:21-205): per environment kGoalStride ints -- [0] number of goals, [1] index of the current
    // goal, [2] lists drawn on the device so far, [3 ...] the goals' x positions in voxels
    int* goals;
    unsigned long long goal_seed;
    // Catcher-v0 (manipulate.py:815: every reset re-draws where the package is launched from): an episode restarted on the device
    // takes over the launch position -- static geometry (the two pegs) and the package's snapshot-0 points -- of a uniformly drawn
    // environment of the batch, i.e. one of the batch's own draws of the reference's random.randint pair
    EnvDesc* envs_rw;                    // nullptr: off
    const unsigned long long* static0;   // static_off of every environment as built
    int* draws;                          // per environment: resamples so far
    int* source;                         // per environment: whose launch position it runs now
};

constexpr int kGoalStride = 64, kMaxGoals = kGoalStride - 3;

// BiWalk.set_random_goals(20, 50, 100) (multi_goal.py:115-129) for an episode the DEVICE starts (the vec-env's reset of a finished
// episode): positions uniform in [20, 50) until the path from voxel 35 through them is at least 100 voxels long.  The reference
// draws from Python's global, unseeded `random`; here a counter-based generator keyed on (seed, environment, list number) gives
// identically distributed lists -- hosts that want Python's stream hand their own lists over with mevo_batch_task_set_goals.
__device__ __forceinline__ unsigned long long splitmix64(unsigned long long& x) {
    unsigned long long z = (x += 0x9e3779b97f4a7c15ull);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
__device__ void draw_goals(int* g, int env, unsigned long long seed) {
    unsigned long long x = seed ^ (0xd1342543de82ef95ull * (unsigned long long)(env + 1)) ^ (0x2545f4914f6cdd1dull * (unsigned long long)(g[2] + 1));
    g[2] += 1;
    int curr = 35, dist = 0, n = 0;
    while (dist < 100 && n < kMaxGoals) {
        const int next = 20 + (int)__umul64hi(splitmix64(x), 30ull);
        dist += abs(curr - next);
        curr = next;
        g[3 + n++] = next;
    }
    g[0] = n; g[1] = 0;
}

struct ObjMeans { double ax, ay, bx, by, cx, cy, dx, dy; };

// means over one object's points of: a = (x / 0.1) * 0.1 (BenchmarkBase.object_pos_at_time), b = x / 0.1
// (EvoSim.object_pos_at_time), c = v / 0.1, d = (v / 0.1) * 0.1.
__device__ __forceinline__ ObjMeans object_means(const double2* pos, const double2* vel, int start, int count, int lane) {
    ObjMeans m = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = lane; i < count; i += 32) {
        const double2 q = pos[start + i], v = vel[start + i];
        const double qx = q.x / kVoxel, qy = q.y / kVoxel, vx = v.x / kVoxel, vy = v.y / kVoxel;
        m.ax += qx * kVoxel; m.ay += qy * kVoxel; m.bx += qx; m.by += qy;
        m.cx += vx; m.cy += vy; m.dx += vx * kVoxel; m.dy += vy * kVoxel;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m.ax += __shfl_xor_sync(0xffffffffu, m.ax, o); m.ay += __shfl_xor_sync(0xffffffffu, m.ay, o);
        m.bx += __shfl_xor_sync(0xffffffffu, m.bx, o); m.by += __shfl_xor_sync(0xffffffffu, m.by, o);
        m.cx += __shfl_xor_sync(0xffffffffu, m.cx, o); m.cy += __shfl_xor_sync(0xffffffffu, m.cy, o);
        m.dx += __shfl_xor_sync(0xffffffffu, m.dx, o); m.dy += __shfl_xor_sync(0xffffffffu, m.dy, o);
    }
    if (count > 0) {
        const double n = (double)count;
        m.ax /= n; m.ay /= n; m.bx /= n; m.by /= n; m.cx /= n; m.cy /= n; m.dx /= n; m.dy /= n;
    }
    return m;
}

// extent of one object along x (axis 0) or y (axis 1) in the units BenchmarkBase.object_pos_at_time returns
__device__ __forceinline__ double object_span(const double2* pos, int start, int count, int lane, int axis) {
    double lo = 1e300, hi = -1e300;
    for (int i = lane; i < count; i += 32) {
        const double2 q = pos[start + i];
        const double v = ((axis ? q.y : q.x) / kVoxel) * kVoxel;
        lo = fmin(lo, v); hi = fmax(hi, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    return hi - lo;
}

// Signed area of the convex hull of one object's points as ShapeBase computes it (evogym/envs/change_shape.py:45-82): the gift
// wrapping there is restated statement for statement -- start at the first point of smallest x, `>=` in the turn test, the running
// `endpoint` scan in point order, the shoelace sum in hull order without an absolute value -- because its sign and its choice among
// collinear points follow from exactly that order.  Sequential by nature: one lane per environment.
__device__ double hull_area_signed(const double2* pos, int start, int count) {
    auto pt = [&](int j) { const double2 q = pos[start + j]; return make_double2((q.x / kVoxel) * kVoxel, (q.y / kVoxel) * kVoxel); };
    if (count <= 0) return 0.0;
    int first = 0;
    double min_x = pt(0).x;
    for (int j = 1; j < count; j++) { const double x = pt(j).x; if (x < min_x) { min_x = x; first = j; } }
    const double2 p0 = pt(first);
    double2 on_hull = p0;
    double area = 0.0;
    for (int i = 0; i < count; i++) {                    // (the reference's list has `count` slots)
        double2 endpoint = pt(0);
        for (int j = 1; j < count; j++) {
            const double2 sj = pt(j);
            // CCW(p1 = S[j], p2 = P[i], p3 = endpoint)
            const bool ccw = (endpoint.y - sj.y) * (on_hull.x - sj.x) >= (on_hull.y - sj.y) * (endpoint.x - sj.x);
            if ((endpoint.x == on_hull.x && endpoint.y == on_hull.y) || !ccw) endpoint = sj;
        }
        area += (on_hull.x * endpoint.y - endpoint.x * on_hull.y);
        on_hull = endpoint;
        if (endpoint.x == p0.x && endpoint.y == p0.y) break;
    }
    return 0.5 * area;
}

template <typename OutT>
__global__ void __launch_bounds__(128) task_outputs_kernel(TaskParams p) {
    const int env = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (env >= p.n_envs) return;
    const TaskEnvRec te = p.tenv[env];
    const size_t soff = p.envs[env].state_off;
    double2* pos = reinterpret_cast<double2*>(p.state_rw + soff);
    double2* vel = pos + te.Pd;

    const bool unstable = p.unstable[env] != 0;
    const double ri_x = p.prev[4 * env], ri_y = p.prev[4 * env + 1], pi_x = p.prev[4 * env + 2], pi_y = p.prev[4 * env + 3];
    ObjMeans R = object_means(pos, vel, te.r_start, te.r_count, lane);
    ObjMeans P = object_means(pos, vel, te.p_start, te.p_count, lane);
    // Sim::step that reports "unstable" saves no snapshot and does not advance time (Sim.cpp:105-115): the
    // environment then reads the same snapshot before and after the step.
    const double rf_x = unstable ? ri_x : R.ax, rf_y = unstable ? ri_y : R.ay;
    const double pf_x = unstable ? pi_x : P.ax, pf_y = unstable ? pi_y : P.ay;

    double reward;
    bool done = unstable;
    int goal_shown = 0;                                  // BidirectionalWalker-v0: the goal index the observation reports
    bool goals_done = false;
    if (p.task == 29) {                                  // multi_goal.py:41-58, 73-84 (GoalBase.get_reward over WalkToX.evaluate)
        int* g = p.goals + (size_t)env * kGoalStride;
        const int count = g[0];
        int cur = g[1];
        goal_shown = cur < count ? cur : count - 1;      // BiWalk.step builds the observation BEFORE get_reward moves on (:143-157)
        reward = 0;
        if (p.reward || p.done) {
            bool reached = true;
            while (reached && cur != count) {
                const double gx = g[3 + cur] * 0.1;
                reward = fabs(gx - ri_x) - fabs(gx - rf_x);
                reached = fabs(gx - rf_x) < 2 * 0.1;     // within two voxels: on to the next goal, evaluated on the same step
                if (reached) cur += 1;
            }
            goals_done = cur == count;
            __syncwarp();
            if (lane == 0) g[1] = cur;
        }
    } else if (p.task == 0 || (p.task >= 4 && p.task <= 8) || p.task == 10 || p.task == 12 || p.task == 18 || p.task == 19 || p.task == 30 || p.task == 32) {   // walk.py:70-84, 180-183, 295-298; traverse.py:20-26 (StairsBase.get_reward)
        reward = rf_x - ri_x;
    } else if (p.task == 3 || p.task == 11) {            // climb.py:96-99, 190-193
        reward = rf_y - ri_y;
    } else if (p.task == 13) {                           // climb.py:288-291
        reward = (rf_y - ri_y) + (rf_x - ri_x) * 0.2;
    } else if (p.task == 9) {                            // jump.py:72-75
        reward = (rf_y - ri_y) * 10 - fabs(rf_x - ri_x) * 5;
    } else if (p.task == 16) {                           // balance.py:47-53 (Balancer-v0)
        reward = fabs(17 * kVoxel - ri_x) - fabs(17 * kVoxel - rf_x);
        reward += (fabs(5.0 * kVoxel - ri_y) - fabs(5.0 * kVoxel - rf_y));
    } else if (p.task == 17) {                           // balance.py:171-177 (Balancer-v1)
        reward = fabs(17.5 * kVoxel - ri_x) - fabs(17.5 * kVoxel - rf_x);
        reward += (fabs(6 * kVoxel - ri_y) - fabs(6 * kVoxel - rf_y));
    } else if (p.task == 27 || p.task == 28) {           // change_shape.py:152-164, 253-265: growth / shrinkage of the convex hull's area
        double area = 0.0;
        if (lane == 0) area = hull_area_signed(pos, te.r_start, te.r_count);
        area = __shfl_sync(0xffffffffu, area, 0);
        const double area_final = unstable ? p.span_prev[env] : area;
        reward = p.task == 27 ? (area_final - p.span_prev[env]) * 10 : (p.span_prev[env] - area_final) * 10;
    } else if (p.task == 25 || p.task == 26) {           // change_shape.py:354-367, 456-469: growth of the robot's extent in x / y
        const double span = object_span(pos, te.r_start, te.r_count, lane, p.task == 26);
        reward = (unstable ? p.span_prev[env] : span) - p.span_prev[env];
    } else if (p.task == 23) {                           // manipulate.py:926-945 (BeamToppler-v0; the "package" is the beam)
        reward = fabs(pi_x - ri_x) - fabs(pf_x - rf_x);
        reward += fabs(pf_x - pi_x) * 1.0;
        reward += fabs(pf_y - pi_y) * 3.0;
        reward += (pi_y - pf_y) * 10;
    } else if (p.task == 31) {                           // manipulate.py:737-753 (Catcher-v0, thresh_height 5 voxels)
        reward = fabs(ri_x - pi_x) - fabs(rf_x - pf_x);
        if (pf_y < 5.0 * kVoxel) reward += 10 * (pf_y - pi_y);
    } else if (p.task == 24) {                           // manipulate.py:1088-1104 (BeamSlider-v0)
        reward = fabs(pi_x - ri_x) - fabs(pf_x - rf_x);
        reward += (pf_x - pi_x) * 1.0;
    } else if (p.task == 22) {                           // flip.py:78-90 (Flipper-v0): change of the unwrapped orientation
        const double ort_1 = p.ort_prev[env], ort_2 = (unstable || !p.ort_now) ? ort_1 : p.ort_now[env];
        int n = p.flips[env];
        const double flattened_1 = n * 2 * kPi + ort_1;
        if (ort_1 < kPi / 3 && ort_2 > 5 * kPi / 3) n -= 1;
        if (ort_1 > 5 * kPi / 3 && ort_2 < kPi / 3) n += 1;
        const double flattened_2 = n * 2 * kPi + ort_2;
        reward = flattened_2 - flattened_1;
        if (lane == 0) p.flips[env] = n;
    } else if (p.task == 21) {                           // manipulate.py:1232-1250 (Lifter-v0)
        reward = (pf_y - pi_y) * 10;
        const double goal = 5.5 * kVoxel;
        reward += (fabs(goal - pi_x) - fabs(goal - pf_x)) * 10;
        if (rf_y < 3 * kVoxel) reward += 20 * (rf_y - ri_y);
    } else if (p.task == 20) {                           // manipulate.py:230-250 (Carrier-v1; its package term is overwritten there, :239-240)
        reward = (fabs(40 * kVoxel - ri_x) - fabs(40 * kVoxel - rf_x));
        const double thresh = pf_x > 20 * kVoxel ? 4.0 * kVoxel : 6.0 * kVoxel;
        if (pf_y < thresh) reward += 10 * (pf_y - pi_y);
    } else if (p.task == 14) {                           // manipulate.py:111-127 (Carrier-v0, thresh_height 3 voxels)
        reward = (pf_x - pi_x) * 0.5;
        reward += (rf_x - ri_x) * 0.5;
        if (pf_y < 3.0 * kVoxel) reward += 10 * (pf_y - pi_y);
    } else if (p.task == 1 || p.task == 15) {            // manipulate.py:38-58 (Pusher-v0, Pusher-v1: thresh_height 0)
        reward = (pf_x - pi_x) * 0.75;
        reward += (rf_x - ri_x) * 0.5;
        reward += fabs(ri_x - pi_x) - fabs(rf_x - pf_x);
        if (pf_y < 0.0 * kVoxel) reward += 10 * (pf_y - pi_y);
    } else {                                             // manipulate.py:545-562
        double motion_penalty = (ri_x - rf_x) * 0.25;
        if (rf_x < 0) motion_penalty *= -1;
        reward = pf_x - pi_x + motion_penalty;
    }
    if (unstable) reward -= 3.0;
    if (goals_done) { done = true; reward += 1.0; }      // multi_goal.py:171-173
    // goal met (walk.py:79-82, 190-193; manipulate.py:390-393; climb.py:107-109, 201-203; traverse.py:120-123, 222-225, 321-324, 420-423)
    if (p.goal_axis == 1 && rf_x > p.goal * kVoxel) { done = true; reward += p.goal_bonus; }
    if (p.goal_axis == 2 && rf_y > p.goal * kVoxel) { done = true; reward += p.goal_bonus; }
    if ((p.task == 23 || p.task == 24) && pf_x < 2 * kVoxel) { reward += 1.0; done = true; }   // manipulate.py:942-944, 996-998 (task_complete)
    if (p.task == 22 && rf_x < 1 * kVoxel) done = true;                          // flip.py:97-98: walked off the left end
    if (p.fall_y > 0 && rf_y < p.fall_y) { reward -= 3.0; done = true; }        // traverse.py:609-613, 734-738
    if (p.ort_now && p.task != 22) {
        // DownStepper-v0 / ObstacleTraverser-v0 / Hurdler-v0: a robot on its back ends the episode (traverse.py:226-228, 325-327, 514-516)
        const double ort = unstable ? p.ort_prev[env] : p.ort_now[env];
        // (UpStepper-v0 in flat mode: the band is pi / 12 wider on both sides, traverse.py:124-129)
        if (ort > kPi / 2 - p.ort_band && ort < 3 * kPi / 2 + p.ort_band) { done = true; reward -= 3.0; }
    }
    if (p.task == 5 && rf_x > 60 * kVoxel) { done = true; reward += 1.0; }        // walk.py:190-193
    int flags = (done ? 1 : 0) | (unstable ? 4 : 0);
    bool was_reset = false;
    int ep = p.episode_len[env] + 1;
    if (p.max_episode_steps > 0 && ep >= p.max_episode_steps) {   // gym TimeLimit: truncated = not already done
        if (!done) flags |= 2;
        flags |= 1;
    }
    if ((flags & 1) && p.reset_done) {
        // Environment::revert_to_snapshot(0) (Environment.cpp:247-263): pos, vel <- snapshot 0; pos_last <- pos;
        // vel_true <- vel; length_eq_goal untouched.
        const double2* s0 = reinterpret_cast<const double2*>(p.state0 + soff);
        for (int i = lane; i < te.Pd; i += 32) {
            const double2 q = s0[i], v = s0[te.Pd + i];
            pos[i] = q; vel[i] = v; pos[2 * te.Pd + i] = q; pos[3 * te.Pd + i] = v;
        }
        __syncwarp();
        if (lane == 0) { p.sim_time[env] = 0; p.unstable[env] = 0; if (p.flips) p.flips[env] = 0; }
        if (p.envs_rw) {                                 // Catcher-v0: the launch position of a drawn environment (see TaskParams)
            int j = 0;
            if (lane == 0) {
                unsigned long long x = p.goal_seed ^ (0xd1342543de82ef95ull * (unsigned long long)(env + 1)) ^
                                       (0x2545f4914f6cdd1dull * (unsigned long long)(p.draws[env] + 1));
                p.draws[env] += 1;
                j = (int)__umul64hi(splitmix64(x), (unsigned long long)p.n_envs);
                p.envs_rw[env].static_off = p.static0[j];
                p.source[env] = j;
            }
            j = __shfl_sync(0xffffffffu, j, 0);
            const TaskEnvRec tj = p.tenv[j];
            const double2* sj = reinterpret_cast<const double2*>(p.state0 + p.envs[j].state_off);
            for (int i = lane; i < te.p_count && i < tj.p_count; i += 32) {
                const double2 q = sj[tj.p_start + i], v = sj[tj.Pd + tj.p_start + i];
                pos[te.p_start + i] = q; vel[te.p_start + i] = v; pos[2 * te.Pd + te.p_start + i] = q; pos[3 * te.Pd + te.p_start + i] = v;
            }
            __syncwarp();
        }
        if (p.task == 29) {                              // BiWalk.reset (multi_goal.py:178-184): first goal of a new list
            if (lane == 0) draw_goals(p.goals + (size_t)env * kGoalStride, env, p.goal_seed);
            __syncwarp();
            goal_shown = 0;
        }
        was_reset = true;
        ep = 0;
        R = object_means(pos, vel, te.r_start, te.r_count, lane);
        P = object_means(pos, vel, te.p_start, te.p_count, lane);
    }
    // What the environment reads from here on is the last SAVED snapshot.  A step that went through (or a reset) has saved one:
    // the live positions and velocities.  An unstable step has not (Sim.cpp:105-115): the live state is partially advanced, the
    // reference's getters -- observation, and the next step's "before" values -- still see the snapshot of the step before.
    double2* opos = reinterpret_cast<double2*>(p.saved + soff);
    double2* ovel = opos + te.Pd;
    if (!unstable || was_reset) {
        for (int i = lane; i < 2 * te.Pd; i += 32) opos[i] = pos[i];
        __syncwarp();
    } else {
        R = object_means(opos, ovel, te.r_start, te.r_count, lane);
        P = object_means(opos, ovel, te.p_start, te.p_count, lane);
    }
    if (p.span_prev) {                                   // the shape measure of that snapshot
        if (p.task >= 27) {
            if (lane == 0) p.span_prev[env] = hull_area_signed(opos, te.r_start, te.r_count);
        } else {
            const double span = object_span(opos, te.r_start, te.r_count, lane, p.task == 26);
            if (lane == 0) p.span_prev[env] = span;
        }
    }
    if (lane == 0) {
        p.episode_len[env] = ep;
        p.prev[4 * env] = R.ax; p.prev[4 * env + 1] = R.ay; p.prev[4 * env + 2] = P.ax; p.prev[4 * env + 3] = P.ay;
        if (p.reward) reinterpret_cast<OutT*>(p.reward)[env] = (OutT)reward;
        if (p.done) p.done[env] = (unsigned char)flags;
        if (p.flag_planes && p.done) {
            p.flag_planes[env] = flags & 1; p.flag_planes[p.n_envs + env] = (flags >> 1) & 1; p.flag_planes[2 * p.n_envs + env] = (flags >> 2) & 1;
        }
    }

    // ---- observations of the (possibly reset) state
    double oth[6];
    oth[2] = oth[3] = oth[4] = oth[5] = 0.0;
    if (p.task == 0 || (p.task >= 3 && p.task <= 13) || p.task == 18 || p.task == 19 || p.task == 30 || p.task == 32) {  // get_vel_com_obs (base.py:214-226 * VOXEL_SIZE, base.py:405-409)
        oth[0] = R.cx * kVoxel; oth[1] = R.cy * kVoxel;  // task 4: orientation and floor sight follow in columns 2.. (host launches)
    } else if (p.task == 29) {                           // get_vel_com_obs, as Walker-v0
        oth[0] = R.cx * kVoxel; oth[1] = R.cy * kVoxel;
    } else if (p.task == 16 || p.task == 17 || p.task == 22 || (p.task >= 25 && p.task <= 28)) {   // Balancer, Flipper: the orientation alone, written by the host's launch; ShapeChange: [0, 0] (base.py:666-667)
        oth[0] = 0.0; oth[1] = 0.0;
    } else if (p.task == 23 || p.task == 24 || p.task == 31) {   // get_obs_topple / get_obs_catch (manipulate.py:726-735) + vel com of robot and beam (base.py:644-650, 696-705); column 6: host
        oth[0] = P.ax - R.ax; oth[1] = P.ay - R.ay; oth[2] = R.cx * kVoxel; oth[3] = R.cy * kVoxel; oth[4] = P.cx * kVoxel; oth[5] = P.cy * kVoxel;
    } else {                                             // PackageBase.get_obs (manipulate.py:22-36)
        oth[0] = R.dx; oth[1] = R.dy; oth[2] = P.ax - R.ax; oth[3] = P.ay - R.ay; oth[4] = P.dx; oth[5] = P.dy;
    }
    const int ow = p.other_width;
    if (p.other) {
        OutT* row = reinterpret_cast<OutT*>(p.other) + (size_t)env * p.other_stride;
        if (lane < p.other_stride) row[lane] = lane < ow ? (OutT)oth[lane < 6 ? lane : 0] : (OutT)0;
        for (int i = 32 + lane; i < p.other_stride; i += 32) row[i] = (OutT)0;
    }
    if (p.flat) {                                        // walk.py:58-61 / manipulate.py:363-367
        OutT* row = reinterpret_cast<OutT*>(p.flat) + (size_t)env * p.flat_stride;
        const int fp = p.flat_prefix;
        if (lane < fp && lane < p.flat_stride) row[lane] = (OutT)oth[lane < 6 ? lane : 0];
        const int n = te.r_count;
        for (int i = lane; i < n; i += 32) {             // get_relative_pos_obs (base.py:228-240 * VOXEL_SIZE)
            const double2 q = opos[te.r_start + i];
            if (fp + i < p.flat_stride) row[fp + i] = (OutT)((q.x / kVoxel - R.bx) * kVoxel);
            if (fp + n + i < p.flat_stride) row[fp + n + i] = (OutT)((q.y / kVoxel - R.by) * kVoxel);
        }
        int tail = 0;
        if (p.task == 29) {                              // [current_goal, x_goal, x_goal - com x] (multi_goal.py:86-88, 146-157)
            tail = 3;
            const double gx = p.goals[(size_t)env * kGoalStride + 3 + goal_shown] * 0.1;
            const double com_x = (unstable && !was_reset) ? ri_x : R.ax;
            const double v = lane == 0 ? (double)goal_shown : lane == 1 ? gx : gx - com_x;
            if (lane < 3 && fp + 2 * n + lane < p.flat_stride) row[fp + 2 * n + lane] = (OutT)v;
        }
        for (int i = fp + 2 * n + tail + lane; i < p.flat_stride; i += 32) row[i] = (OutT)0;
    }
    if (p.modular) {                                     // base.py:480-506: 9 values per voxel, non-empty voxels first
        OutT* row = reinterpret_cast<OutT*>(p.modular) + (size_t)env * p.modular_stride;
        const int total = 9 * te.n_vox;
        for (int i = lane; i < p.modular_stride; i += 32) {
            double v = 0.0;
            if (i < total) {
                const int j = i / 9, c = i - 9 * j;
                const short4 id = p.vox_idx[te.vox_off + j];
                if (c == 8) {
                    v = (double)p.vox_mat[te.vox_off + j];
                } else if (id.x >= 0) {
                    const int k = c & 3;
                    const int pt = k == 0 ? id.x : k == 1 ? id.y : k == 2 ? id.z : id.w;
                    const double2 q = opos[te.r_start + pt];
                    v = c < 4 ? (q.x / kVoxel - R.bx) * kVoxel : (q.y / kVoxel - R.by) * kVoxel;
                }
            }
            row[i] = (OutT)v;
        }
    }
}

constexpr int kSightCols = 64;            // sight blocks up to this many columns (sight_dist <= 31) are computed in one pass

struct SightEnvRec {                     // 48 bytes per environment
    int o_start, o_count;                // observed object: dynamic-local point range
    int Pd;
    int st_off, st_count;                // static terrain points (double2 units into the shared table)
    int n_dyn;
    int dyn_start[3], dyn_count[3];      // terrain objects that are simulated (mixed / moved terrain): ranges in the state
};

// get_floor_obs / get_ceil_obs (evogym/envs/base.py:269-372, scaled by VOXEL_SIZE in BenchmarkBase, base.py:411-415): for
// each of the 2 * sight_dist + 1 voxel columns around the object's centre of mass, the distance to the highest terrain
// point below it (floor) or the lowest above it (ceil), clipped to [0, sight_range].  One warp per environment; the
// comparisons are the reference's expressions in the reference's (voxel) units.
template <typename OutT>
__global__ void __launch_bounds__(128) sight_kernel(const double* state, const EnvDesc* envs, const SightEnvRec* rec, const double2* st_pts,
                                                    int n_envs, int sight_dist, double sight_range, int ceil, OutT* out, int stride,
                                                    int row_stride, const int* env_off) {
    const int env = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (env >= n_envs) return;
    const SightEnvRec r = rec[env];
    const double2* pos = reinterpret_cast<const double2*>(state + envs[env].state_off);
    double cx = 0.0, cy = 0.0;
    for (int i = lane; i < r.o_count; i += 32) { const double2 q = pos[r.o_start + i]; cx += q.x / kVoxel; cy += q.y / kVoxel; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { cx += __shfl_xor_sync(0xffffffffu, cx, o); cy += __shfl_xor_sync(0xffffffffu, cy, o); }
    if (r.o_count > 0) { cx /= r.o_count; cy /= r.o_count; }
    const double x_lo = cx - (sight_dist + 0.5), x_hi = cx + (sight_dist + 0.5);
    const int n_cols = 2 * sight_dist + 1;
    OutT* row = out + (size_t)env * row_stride + (env_off ? env_off[env] : 0);   // env_off: the block's column depends on the robot (flat rows)
    if (n_cols <= kSightCols) {
        // ONE pass over the terrain points.  The columns are the open intervals (cx + i - 0.5, cx + i + 0.5): neighbours share their
        // end point, so a point lies in at most one of them -- found from the rounded offset and confirmed (against its two neighbours
        // too: x - cx rounds) with the reference's own comparisons.  Per column the highest (floor) / lowest (ceiling) y is kept as an
        // order-preserving integer key in shared memory; a maximum does not depend on the visiting order, so the result is the
        // column-by-column scan's, bit for bit.
        __shared__ unsigned long long s_key[4][kSightCols];
        unsigned long long* key = s_key[(threadIdx.x >> 5) & 3];
        const unsigned long long none = ceil ? ~0ull : 0ull;
        for (int c = lane; c < n_cols; c += 32) key[c] = none;
        __syncwarp();
        auto visit = [&](double2 q) {
            const double x = q.x / kVoxel, y = q.y / kVoxel;
            if (!(x > x_lo) || !(x < x_hi)) return;
            if (ceil ? !(y > cy) : !(y < cy)) return;
            const int g = (int)floor(x - cx + 0.5);
            for (int i = g - 1; i <= g + 1; i++) {
                if (i < -sight_dist || i > sight_dist) continue;
                const double c_lo = cx + (i - 0.5), c_hi = cx + (i + 0.5);
                if (!(x > c_lo) || !(x < c_hi)) continue;
                const long long b = __double_as_longlong(y);
                const unsigned long long k = b < 0 ? ~(unsigned long long)b : (unsigned long long)b | 0x8000000000000000ull;
                if (ceil) atomicMin(&key[i + sight_dist], k); else atomicMax(&key[i + sight_dist], k);
                break;
            }
        };
        for (int k = lane; k < r.st_count; k += 32) visit(st_pts[r.st_off + k]);
        for (int s = 0; s < r.n_dyn; s++)
            for (int k = lane; k < r.dyn_count[s]; k += 32) visit(pos[r.dyn_start[s] + k]);
        __syncwarp();
        for (int c = lane; c < n_cols && c < stride; c += 32) {
            const unsigned long long k = key[c];
            const bool any = k != none;
            const double best = __longlong_as_double((k & 0x8000000000000000ull) ? (long long)(k & 0x7fffffffffffffffull) : (long long)~k);
            double elev = any ? best : (ceil ? (0.0 + sight_range) + cy : (0.0 - sight_range) + cy);
            elev = ceil ? elev - cy : cy - elev;
            elev = fmin(fmax(elev, 0.0), sight_range);
            row[c] = (OutT)(elev * kVoxel);
        }
    } else {                                             // wider than any environment of the reference looks: column by column
        for (int c = 0; c < n_cols && c < stride; c++) {
            const int i = c - sight_dist;
            const double c_lo = cx + (i - 0.5), c_hi = cx + (i + 0.5);
            double best = 0.0;
            bool any = false;
            auto visit = [&](double2 q) {
                const double x = q.x / kVoxel, y = q.y / kVoxel;
                if (!(x > x_lo) || !(x < x_hi)) return;
                if (ceil ? !(y > cy) : !(y < cy)) return;
                if (!(x > c_lo) || !(x < c_hi)) return;
                if (!any) { best = y; any = true; }
                else best = ceil ? fmin(best, y) : fmax(best, y);
            };
            for (int k = lane; k < r.st_count; k += 32) visit(st_pts[r.st_off + k]);
            for (int s = 0; s < r.n_dyn; s++)
                for (int k = lane; k < r.dyn_count[s]; k += 32) visit(pos[r.dyn_start[s] + k]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double ob = __shfl_xor_sync(0xffffffffu, best, o);
                const bool oa = __shfl_xor_sync(0xffffffffu, any ? 1 : 0, o) != 0;
                if (oa) { best = any ? (ceil ? fmin(best, ob) : fmax(best, ob)) : ob; any = true; }
            }
            double elev = any ? best : (ceil ? (0.0 + sight_range) + cy : (0.0 - sight_range) + cy);
            elev = ceil ? elev - cy : cy - elev;
            elev = fmin(fmax(elev, 0.0), sight_range);
            if (lane == 0) row[c] = (OutT)(elev * kVoxel);
        }
    }
    for (int c = n_cols + lane; c < stride; c += 32) row[c] = (OutT)0;
}

__global__ void source_identity_kernel(int* source, int n_envs) {
    const int env = blockIdx.x * blockDim.x + threadIdx.x;
    if (env < n_envs) source[env] = env;
}

__global__ void draw_goals_kernel(int* goals, int n_envs, unsigned long long seed) {
    const int env = blockIdx.x * blockDim.x + threadIdx.x;
    if (env < n_envs) draw_goals(goals + (size_t)env * kGoalStride, env, seed);
}

template <typename InT>
__global__ void policy_actions_kernel(const InT* a, int in_stride, int per_voxel, const TaskEnvRec* tenv, const short* act_vox,
                                      double* act, int act_stride, int n_envs) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_envs * act_stride) return;
    const int env = idx / act_stride, slot = idx - env * act_stride;
    double out = 1.0;
    if (slot < tenv[env].n_act) {
        const int col = per_voxel ? act_vox[idx] : slot;
        // the reference adds 1 and clips in the policy's dtype (numpy keeps float32 under python scalars)
        InT v = a[(size_t)env * in_stride + col] + (InT)1;
        v = v < (InT)0.6 ? (InT)0.6 : v;
        v = v > (InT)1.6 ? (InT)1.6 : v;
        out = (double)v;
    }
    act[idx] = out;
}

// ---- what distinguishes the tasks on the HOST side (the reward formulas are in task_outputs_kernel, by task id) ----------------------
struct SightPart { const char* terrain; int dist; bool ceil; int col; };      // terrain: object names separated by ','; nullptr: none
struct TaskDesc {
    const char* env;                 // the reference's environment id (documentation)
    int goal_axis;                   // 0: no goal, 1: robot com x, 2: robot com y beyond `goal` voxels ends the episode with `bonus`
    double goal, bonus;
    bool ort_rule;                   // a robot on its back ends the episode with -3 (traverse.py:226-228, 325-327, 514-516, 615-617)
    double fall_y;                   // > 0: a centre of mass below this height [m] ends the episode with -3
    bool package;                    // the task has a second object (package / beam)
    int other_width;                 // width of the "other" observation
    int kernel_width;                // ... of which the task kernel writes the first values itself; the rest comes from the launches below
    int flat_prefix;                 // values in front of the relative positions in the flat observation
    int ort_col;                     // column of an orientation value in "other" (-1: none) ...
    bool ort_of_package;             // ... of the package instead of the robot ...
    bool ort_in_flat;                // ... also at the same column of the flat observation
    SightPart sight[2];              // floor / ceiling sight blocks of "other"
};
constexpr const char* kPlatforms5 = "platform_1,platform_2,platform_3,platform_4,platform_5";
constexpr const char* kPlatforms7 = "platform_1,platform_2,platform_3,platform_4,platform_5,platform_6,platform_7";
// BlockSoup.terrain_list (traverse.py:815-823): the pit and the 27 loose blocks in it
constexpr const char* kBlockSoup = "ground,block_0,block_1,block_2,block_3,block_4,block_5,block_6,block_7,block_8,block_9,block_10,block_11,block_12,block_13,block_14,block_15,block_16,block_17,block_18,block_19,block_20,block_21,block_22,block_23,block_24,block_25,block_26";
#define MEVO_NO_SIGHT {{nullptr, 0, false, 0}, {nullptr, 0, false, 0}}
#define MEVO_FLOOR5(names) {{names, 5, false, 3}, {nullptr, 0, false, 0}}
const TaskDesc kTasks[] = {
    /*  0 */ {"Walker-v0", 1, 99, 1, false, 0, false, 2, 2, 2, -1, false, false, MEVO_NO_SIGHT},
    /*  1 */ {"Pusher-v0", 1, 99, 1, false, 0, true, 6, 6, 6, -1, false, false, MEVO_NO_SIGHT},
    /*  2 */ {"Thrower-v0", 0, 0, 0, false, 0, true, 6, 6, 6, -1, false, false, MEVO_NO_SIGHT},
    /*  3 */ {"Climber-v0", 2, 86, 1, false, 0, false, 2, 2, 2, -1, false, false, MEVO_NO_SIGHT},
    /*  4 */ {"ObstacleTraverser-v1", 1, 59, 2, false, 0, false, 14, 2, 2, 2, false, false, MEVO_FLOOR5("ground")},
    /*  5 */ {"BridgeWalker-v0", 1, 60, 1, false, 0, false, 3, 3, 3, 2, false, true, MEVO_NO_SIGHT},
    /*  6 */ {"ObstacleTraverser-v0", 1, 79, 2, true, 0, false, 14, 2, 2, 2, false, false, MEVO_FLOOR5("ground")},
    /*  7 */ {"UpStepper-v0", 1, 69, 2, false, 0, false, 14, 2, 2, 2, false, false, MEVO_FLOOR5("ground")},
    /*  8 */ {"DownStepper-v0", 1, 74, 2, true, 0, false, 14, 2, 2, 2, false, false, MEVO_FLOOR5("ground")},
    /*  9 */ {"Jumper-v0", 0, 0, 0, false, 0, false, 7, 2, 2, -1, false, false, {{"ground", 2, false, 2}, {nullptr, 0, false, 0}}},
    /* 10 */ {"Hurdler-v0", 0, 0, 0, true, 0, false, 14, 2, 2, 2, false, false, MEVO_FLOOR5("ground")},
    /* 11 */ {"Climber-v1", 2, 65, 1, false, 0, false, 2, 2, 2, -1, false, false, MEVO_NO_SIGHT},
    /* 12 */ {"CaveCrawler-v0", 1, 69, 1, false, 0, false, 24, 2, 2, -1, false, false, {{"terrain", 5, false, 2}, {"terrain", 5, true, 13}}},
    /* 13 */ {"Climber-v2", 0, 0, 0, false, 0, false, 10, 2, 2, 2, false, false, {{"pipe", 3, true, 3}, {nullptr, 0, false, 0}}},
    /* 14 */ {"Carrier-v0", 1, 99, 1, false, 0, true, 6, 6, 6, -1, false, false, MEVO_NO_SIGHT},
    /* 15 */ {"Pusher-v1", 1, 69, 1, false, 0, true, 6, 6, 6, -1, false, false, MEVO_NO_SIGHT},
    /* 16 */ {"Balancer-v0", 0, 0, 0, false, 0, false, 1, 1, 1, 0, false, true, MEVO_NO_SIGHT},
    /* 17 */ {"Balancer-v1", 0, 0, 0, false, 0, false, 1, 1, 1, 0, false, true, MEVO_NO_SIGHT},
    /* 18 */ {"GapJumper-v0", 0, 0, 0, false, 4 * kVoxel, false, 14, 2, 2, 2, false, false, MEVO_FLOOR5(kPlatforms5)},
    /* 19 */ {"PlatformJumper-v0", 0, 0, 0, true, 0.4, false, 14, 2, 2, 2, false, false, MEVO_FLOOR5(kPlatforms7)},
    /* 20 */ {"Carrier-v1", 0, 0, 0, false, 0, true, 6, 6, 6, -1, false, false, MEVO_NO_SIGHT},
    /* 21 */ {"Lifter-v0", 0, 0, 0, false, 0, true, 7, 7, 7, 6, true, true, MEVO_NO_SIGHT},
    /* 22 */ {"Flipper-v0", 0, 0, 0, false, 0, false, 1, 1, 1, 0, false, true, MEVO_NO_SIGHT},
    /* 23 */ {"BeamToppler-v0", 0, 0, 0, false, 0, true, 7, 7, 7, 6, true, true, MEVO_NO_SIGHT},
    /* 24 */ {"BeamSlider-v0", 0, 0, 0, false, 0, true, 7, 7, 7, 6, true, true, MEVO_NO_SIGHT},
    /* 25 */ {"WingspanMazimizer-v0", 0, 0, 0, false, 0, false, 2, 2, 0, -1, false, false, MEVO_NO_SIGHT},
    /* 26 */ {"HeightMaximizer-v0", 0, 0, 0, false, 0, false, 2, 2, 0, -1, false, false, MEVO_NO_SIGHT},
    /* 27 */ {"AreaMaximizer-v0", 0, 0, 0, false, 0, false, 2, 2, 0, -1, false, false, MEVO_NO_SIGHT},
    /* 28 */ {"AreaMinimizer-v0", 0, 0, 0, false, 0, false, 2, 2, 0, -1, false, false, MEVO_NO_SIGHT},
    /* 29 */ {"BidirectionalWalker-v0", 0, 0, 0, false, 0, false, 2, 2, 2, -1, false, false, MEVO_NO_SIGHT},
    /* 30 */ {"Traverser-v0", 1, 59, 2, false, 0, false, 14, 2, 2, 2, false, false, MEVO_FLOOR5(kBlockSoup)},
    /* 31 */ {"Catcher-v0", 0, 0, 0, false, 0, true, 7, 7, 7, 6, true, true, MEVO_NO_SIGHT},
    /* 32 */ {"UpStepper-v0 (flat mode)", 1, 69, 2, true, 0, false, 14, 2, 2, 2, false, false, MEVO_FLOOR5("ground")},
};
constexpr int kNumTasks = (int)(sizeof(kTasks) / sizeof(kTasks[0]));
constexpr int kTaskFlipper = 22, kTaskWingspan = 25, kTaskAreaMin = 28, kTaskBiWalk = 29, kTaskCatcher = 31, kTaskUpStepperFlat = 32;

std::vector<std::string> split_names(const char* names) {
    std::vector<std::string> out;
    std::string cur;
    for (const char* c = names; ; c++) {
        if (*c == ',' || *c == 0) { out.push_back(cur); cur.clear(); if (*c == 0) break; }
        else cur.push_back(*c);
    }
    return out;
}

}  // namespace

struct Batch::Task {
    int task = -1;
    std::string robot, package;
    int max_episode_steps = 0;
    int epoch = -1;                          // finalize epoch the tables were built for
    int flat_width = 0, modular_width = 0, other_width = 0, max_vox = 0;
    int robot_obj = -1;
    int package_obj = -1;                // object index of the package (the same in every environment), -1: none
    std::vector<TaskEnvRec> rec;
    std::vector<short4> vox_idx;
    std::vector<unsigned char> vox_mat;
    std::vector<short> act_vox;
    std::vector<std::vector<int>> index_array;   // per env: n_vox x 4 in grid order (transform_to_modular_obs)
    std::vector<std::vector<int>> act_mask, obs_mask;
    std::vector<int> corner_size;
    TaskEnvRec* d_rec = nullptr;
    short4* d_vox_idx = nullptr;
    unsigned char* d_vox_mat = nullptr;
    short* d_act_vox = nullptr;
    double* d_prev = nullptr;
    double* d_ort_now = nullptr;         // orientation rule of tasks 6, 8: after the step / at the last observation
    double* d_ort_prev = nullptr;
    int* d_flips = nullptr;              // Flipper-v0
    double* d_span_prev = nullptr;       // Wingspan / Height maximizer
    int* d_source = nullptr;             // Catcher-v0: environment whose launch position each environment runs now
    unsigned long long* d_static0 = nullptr;     // Catcher-v0: static_off of every environment as built
    double* d_saved = nullptr;           // positions and velocities of the last saved snapshot of every environment (layout of the state)
    unsigned char* d_flag_planes = nullptr;      // caller's buffer (mevo_batch_task_bind_flag_planes), not owned
    int* d_flat_tail = nullptr;          // sighted tasks: 2 x robot points per environment = how far the sight blocks move right in the flat row
    int flat_fixed = 0;                  // values in front of the relative positions in the flat row
    int* d_goals = nullptr;              // BidirectionalWalker-v0: kGoalStride ints per environment (TaskParams::goals)
    unsigned long long goal_seed = 0x6a09e667f3bcc909ull;
    int* d_episode = nullptr;
    void* d_scratch = nullptr;               // host-facing outputs staged here
    size_t scratch_bytes = 0;
    void free_device() {
        cudaFree(d_rec); cudaFree(d_vox_idx); cudaFree(d_vox_mat); cudaFree(d_act_vox); cudaFree(d_prev); cudaFree(d_ort_now); cudaFree(d_ort_prev); cudaFree(d_flips); cudaFree(d_span_prev); cudaFree(d_goals); cudaFree(d_flat_tail); cudaFree(d_saved); cudaFree(d_static0); cudaFree(d_source); cudaFree(d_episode);
        cudaFree(d_scratch);
        d_rec = nullptr; d_vox_idx = nullptr; d_vox_mat = nullptr; d_act_vox = nullptr; d_prev = nullptr; d_ort_now = nullptr; d_ort_prev = nullptr; d_flips = nullptr; d_span_prev = nullptr; d_goals = nullptr; d_flat_tail = nullptr; d_saved = nullptr; d_static0 = nullptr; d_source = nullptr; d_episode = nullptr;
        d_scratch = nullptr; scratch_bytes = 0;
    }
};

void Batch::task_free() {
    if (!task_) return;
    task_->free_device();
    delete task_;
    task_ = nullptr;
}

// transform_to_modular_obs (evogym/envs/base.py:427-466) on a voxel grid, row 0 = top: for each cell the indices
// (into the object's point list) of its TL, TR, BL, BR corners, -1 for empty cells; returns loc_idx.
int modular_corner_index(const std::vector<int>& body, int h, int w, std::vector<int>& out) {
    out.assign((size_t)h * w * 4, -1);
    auto at = [&](int y, int x, int r, int c) -> int& { return out[((size_t)(y * w + x) * 2 + r) * 2 + c]; };
    int loc = 0;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            if (body[y * w + x] == 0) continue;
            const bool up = y - 1 >= 0 && body[(y - 1) * w + x] != 0;
            const bool left = x - 1 >= 0 && body[y * w + x - 1] != 0;
            const bool right = x + 1 < w && body[y * w + x + 1] != 0;
            const bool up_right = x + 1 < w && y - 1 >= 0 && body[(y - 1) * w + x + 1] != 0;
            if (up) { at(y, x, 0, 0) = at(y - 1, x, 1, 0); at(y, x, 0, 1) = at(y - 1, x, 1, 1); }
            if (left) { at(y, x, 0, 0) = at(y, x - 1, 0, 1); at(y, x, 1, 0) = at(y, x - 1, 1, 1); }
            if (!(up || left)) at(y, x, 0, 0) = loc++;
            if (!up) {
                if (right && up_right) at(y, x, 0, 1) = at(y - 1, x + 1, 1, 0);
                else at(y, x, 0, 1) = loc++;
            }
            if (!left) at(y, x, 1, 0) = loc++;
            at(y, x, 1, 1) = loc++;
        }
    return loc;
}

void Batch::task_configure(int task, const std::string& robot, const std::string& package, int max_episode_steps) {
    if (task < 0 || task >= kNumTasks) throw std::runtime_error("task_configure: unknown task id");
    const bool needs_package = kTasks[task].package;
    if (needs_package && package.empty()) throw std::runtime_error("task_configure: this task needs a package object");
    if (!task_) task_ = new Task();
    task_->task = task;
    task_->robot = robot;
    task_->package = needs_package ? package : std::string();
    task_->max_episode_steps = max_episode_steps;
    task_->epoch = -1;
    task_->d_flag_planes = nullptr;
}

void Batch::task_build() {
    if (!task_ || task_->task < 0) throw std::runtime_error("task layer used before mevo_batch_task_configure");
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) finalize();
    if (task_->epoch == finalize_epoch_) return;
    Task& T = *task_;
    const int n = num_envs();
    T.free_device();
    T.rec.assign(n, TaskEnvRec());
    T.vox_idx.clear(); T.vox_mat.clear();
    T.act_vox.assign((size_t)n * std::max(1, act_stride_), 0);
    T.index_array.assign(n, {}); T.act_mask.assign(n, {}); T.obs_mask.assign(n, {}); T.corner_size.assign(n, 0);
    const TaskDesc& D = kTasks[T.task];
    T.other_width = D.other_width;
    T.robot_obj = -1; T.package_obj = -1;
    T.flat_width = T.other_width; T.modular_width = 0; T.max_vox = 0;
    // Flat rows of the sighted environments (traverse.py:96-102, 396-401, walk.py:282-287, climb.py:272-277, jump.py:60-64): the
    // values of "other" with the relative positions inserted in front of the first sight block
    int first_sight = -1;
    for (const SightPart& sp : D.sight) if (sp.terrain && (first_sight < 0 || sp.col < first_sight)) first_sight = sp.col;
    T.flat_fixed = first_sight >= 0 ? first_sight : D.flat_prefix;
    std::vector<int> flat_tail(n, 0);
    for (int e = 0; e < n; e++) {
        const EnvHost& E = envs_[e];
        const int ro = E.world.find_object(T.robot);
        if (ro < 0) throw std::runtime_error("task: environment " + std::to_string(e) + " has no object named " + T.robot);
        const VoxelObject& rb = E.world.objects[ro];
        if (T.robot_obj < 0) T.robot_obj = ro;
        else if (T.robot_obj != ro) throw std::runtime_error("task: the robot must be the same object index in every environment");
        TaskEnvRec& r = T.rec[e];
        r.Pd = (int)E.dyn_points.size();
        r.r_start = E.point_local[rb.min_point];
        r.r_count = rb.max_point - rb.min_point + 1;
        if (r.r_start < 0) throw std::runtime_error("task: the robot is a static object");
        if (!T.package.empty()) {
            const int po = E.world.find_object(T.package);
            if (po < 0) throw std::runtime_error("task: environment " + std::to_string(e) + " has no object named " + T.package);
            if (T.package_obj < 0) T.package_obj = po;
            else if (T.package_obj != po) throw std::runtime_error("task: the package must be the same object index in every environment");
            const VoxelObject& pk = E.world.objects[po];
            r.p_start = E.point_local[pk.min_point];
            r.p_count = pk.max_point - pk.min_point + 1;
            if (r.p_start < 0) throw std::runtime_error("task: the package is a static object");
        }
        // voxel grid of the robot, row 0 = top
        const int w = rb.grid_w, h = rb.grid_h;
        std::vector<int> body((size_t)w * h, 0);
        for (const Boxel& b : rb.boxels) body[b.grid_index] = b.cell_type;
        std::vector<int> corner;
        T.corner_size[e] = modular_corner_index(body, h, w, corner);
        T.index_array[e] = corner;
        if (T.corner_size[e] > r.r_count)
            throw std::runtime_error("task: modular corner numbering exceeds the robot's point count (env " + std::to_string(e) + ")");
        r.vox_off = (int)T.vox_idx.size();
        r.n_vox = w * h;
        // modular_ob_wrapper_padding (base.py:480-506): non-empty voxels in grid order, then the empty ones
        for (int pass = 0; pass < 2; pass++)
            for (int v = 0; v < w * h; v++) {
                if ((body[v] != 0) != (pass == 0)) continue;
                short4 id;
                id.x = (short)corner[4 * v]; id.y = (short)corner[4 * v + 1]; id.z = (short)corner[4 * v + 2]; id.w = (short)corner[4 * v + 3];
                T.vox_idx.push_back(id);
                T.vox_mat.push_back((unsigned char)body[v]);
            }
        // get_act_mask (base.py:536-545): 0 on actuator voxels; get_obs_mask (:547-557): 0 per non-empty voxel, then 1 per empty
        T.act_mask[e].resize(w * h);
        int k = 0, n_full = 0;
        for (int v = 0; v < w * h; v++) {
            const bool act = body[v] == CELL_ACT_H || body[v] == CELL_ACT_V;
            T.act_mask[e][v] = act ? 0 : 1;
            if (act && k < act_stride_) T.act_vox[(size_t)e * act_stride_ + k++] = (short)v;
            n_full += body[v] != 0;
        }
        T.obs_mask[e].assign(w * h, 1);
        std::fill(T.obs_mask[e].begin(), T.obs_mask[e].begin() + n_full, 0);
        r.n_act = std::min(k, E.n_act);
        T.flat_width = std::max(T.flat_width, first_sight >= 0 ? D.other_width + 2 * r.r_count
                                                                 : D.flat_prefix + 2 * r.r_count + (T.task == kTaskBiWalk ? 3 : 0));
        flat_tail[e] = 2 * r.r_count;
        T.modular_width = std::max(T.modular_width, 9 * w * h);
        T.max_vox = std::max(T.max_vox, w * h);
    }
    auto up = [&](auto*& d, const auto& v) {
        using Elem = typename std::remove_reference<decltype(v)>::type::value_type;
        CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&d), sizeof(Elem) * std::max<size_t>(1, v.size())));
        if (!v.empty()) CUDA_CHECK(cudaMemcpy(d, v.data(), sizeof(Elem) * v.size(), cudaMemcpyHostToDevice));
    };
    up(T.d_flat_tail, flat_tail);
    up(T.d_rec, T.rec); up(T.d_vox_idx, T.vox_idx); up(T.d_vox_mat, T.vox_mat); up(T.d_act_vox, T.act_vox);
    CUDA_CHECK(cudaMalloc(&T.d_saved, sizeof(double) * (state_doubles_ + 2)));
    CUDA_CHECK(cudaMemcpy(T.d_saved, d_state_, sizeof(double) * state_doubles_, cudaMemcpyDeviceToDevice));
    CUDA_CHECK(cudaMalloc(&T.d_prev, sizeof(double) * 4 * n));
    CUDA_CHECK(cudaMalloc(&T.d_episode, sizeof(int) * n));
    CUDA_CHECK(cudaMemset(T.d_prev, 0, sizeof(double) * 4 * n));
    CUDA_CHECK(cudaMalloc(&T.d_ort_now, sizeof(double) * n));
    CUDA_CHECK(cudaMalloc(&T.d_ort_prev, sizeof(double) * n));
    CUDA_CHECK(cudaMemset(T.d_ort_prev, 0, sizeof(double) * n));
    CUDA_CHECK(cudaMalloc(&T.d_flips, sizeof(int) * n));
    CUDA_CHECK(cudaMemset(T.d_flips, 0, sizeof(int) * n));
    CUDA_CHECK(cudaMalloc(&T.d_span_prev, sizeof(double) * n));
    CUDA_CHECK(cudaMemset(T.d_span_prev, 0, sizeof(double) * n));
    CUDA_CHECK(cudaMemset(T.d_episode, 0, sizeof(int) * n));
    if (T.task == kTaskCatcher) {                        // the launch positions the batch was built with
        std::vector<EnvDesc> desc(n);
        CUDA_CHECK(cudaMemcpy(desc.data(), d_envdesc_, sizeof(EnvDesc) * n, cudaMemcpyDeviceToHost));
        std::vector<unsigned long long> off(n);
        for (int e = 0; e < n; e++) off[e] = desc[e].static_off;
        CUDA_CHECK(cudaMalloc(&T.d_static0, sizeof(unsigned long long) * n));
        CUDA_CHECK(cudaMemcpy(T.d_static0, off.data(), sizeof(unsigned long long) * n, cudaMemcpyHostToDevice));
        std::vector<int> self(n);
        for (int e = 0; e < n; e++) self[e] = e;
        CUDA_CHECK(cudaMalloc(&T.d_source, sizeof(int) * n));
        CUDA_CHECK(cudaMemcpy(T.d_source, self.data(), sizeof(int) * n, cudaMemcpyHostToDevice));
    }
    if (T.task == kTaskBiWalk) {                         // every environment starts with a drawn list (BiWalk.__init__, multi_goal.py:108)
        CUDA_CHECK(cudaMalloc(&T.d_goals, sizeof(int) * kGoalStride * n));
        CUDA_CHECK(cudaMemset(T.d_goals, 0, sizeof(int) * kGoalStride * n));
        draw_goals_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream_>>>(T.d_goals, n, T.goal_seed);
        CUDA_CHECK(cudaGetLastError());
    }
    T.epoch = finalize_epoch_;
}

// BidirectionalWalker-v0: the goal lists.  `goals` holds `stride` ints per environment of which the first counts[i] are used
// (x positions in voxels, the argument of WalkToX, multi_goal.py:64-71); the current-goal index of those environments returns to 0.
void Batch::task_set_goals(int env_first, int n_envs, const int* counts, const int* goals, int stride) {
    task_build();
    Task& T = *task_;
    if (T.task != kTaskBiWalk) throw std::runtime_error("task_set_goals: the configured task has no goal list");
    if (env_first < 0 || n_envs < 0 || env_first + n_envs > num_envs()) throw std::runtime_error("task_set_goals: environment range out of bounds");
    std::vector<int> host((size_t)n_envs * kGoalStride, 0);
    for (int i = 0; i < n_envs; i++) {
        if (counts[i] < 1) throw std::runtime_error("task_set_goals: an environment needs at least one goal (multi_goal.py:28-29)");
        if (counts[i] > kMaxGoals || counts[i] > stride)
            throw std::runtime_error("task_set_goals: at most " + std::to_string(kMaxGoals) + " goals per environment");
        host[(size_t)i * kGoalStride] = counts[i];
        for (int k = 0; k < counts[i]; k++) host[(size_t)i * kGoalStride + 3 + k] = goals[(size_t)i * stride + k];
    }
    if (n_envs == 0) return;
    cudaStream_t s = (cudaStream_t)stream_;
    CUDA_CHECK(cudaStreamSynchronize(s));
    // the per-environment draw counters ([2]) are kept: read, merge, write back
    std::vector<int> cur((size_t)n_envs * kGoalStride);
    CUDA_CHECK(cudaMemcpy(cur.data(), T.d_goals + (size_t)env_first * kGoalStride, sizeof(int) * cur.size(), cudaMemcpyDeviceToHost));
    for (int i = 0; i < n_envs; i++) host[(size_t)i * kGoalStride + 2] = cur[(size_t)i * kGoalStride + 2];
    CUDA_CHECK(cudaMemcpy(T.d_goals + (size_t)env_first * kGoalStride, host.data(), sizeof(int) * host.size(), cudaMemcpyHostToDevice));
}

// New lists for every environment, drawn on the device from `seed` (and the seed later resets of finished episodes draw from).
void Batch::task_draw_goals(unsigned long long seed) {
    task_build();
    Task& T = *task_;
    if (T.task != kTaskBiWalk) throw std::runtime_error("task_draw_goals: the configured task has no goal list");
    T.goal_seed = seed;
    draw_goals_kernel<<<(num_envs() + 127) / 128, 128, 0, (cudaStream_t)stream_>>>(T.d_goals, num_envs(), seed);
    CUDA_CHECK(cudaGetLastError());
}

// The done byte of task_outputs as three byte planes of 0 / 1 (done, truncated, unstable; n_envs bytes each) in a device buffer of the
// caller -- what a torch.bool tensor of shape (3, n) is -- so that a vec-env front needs no kernels of its own to split the bit field.
void Batch::task_bind_flag_planes(void* device_planes) {
    if (!task_) throw std::runtime_error("task layer used before mevo_batch_task_configure");
    task_->d_flag_planes = static_cast<unsigned char*>(device_planes);
}

// Catcher-v0: out[env] = the environment whose launch position (pegs, package start) `env` runs now -- itself until the task kernel
// restarts its episode.  The host-side world of `env` (mevo_env_get_state's rows of static points, the rasteriser) keeps showing the
// pegs where the environment was BUILT; the simulation uses those of out[env].
void Batch::task_launch_sources(int* out) {
    task_build();
    Task& T = *task_;
    if (!T.d_source) throw std::runtime_error("task_launch_sources: the configured task does not re-draw launch positions");
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    CUDA_CHECK(cudaMemcpy(out, T.d_source, sizeof(int) * num_envs(), cudaMemcpyDeviceToHost));
}

// (count, current index, goals ...) of one environment, for hosts and tests
int Batch::task_get_goals(int env, int* out, int cap) {
    task_build();
    Task& T = *task_;
    if (T.task != kTaskBiWalk) throw std::runtime_error("task_get_goals: the configured task has no goal list");
    if (env < 0 || env >= num_envs()) throw std::runtime_error("task_get_goals: environment out of range");
    int host[kGoalStride];
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    CUDA_CHECK(cudaMemcpy(host, T.d_goals + (size_t)env * kGoalStride, sizeof(host), cudaMemcpyDeviceToHost));
    for (int k = 0; k < host[0] && k < cap; k++) out[k] = host[3 + k];
    return host[0] | (host[1] << 16);
}

void Batch::task_dims(int* dims) {
    task_build();
    dims[0] = task_->flat_width; dims[1] = task_->modular_width; dims[2] = task_->other_width; dims[3] = task_->max_vox;
    dims[4] = act_stride_;
}

int Batch::task_env_tables(int env, int* index_array, int* act_mask, int* obs_mask, int cap_voxels) {
    task_build();
    const Task& T = *task_;
    const int nv = T.rec.at(env).n_vox;
    if (cap_voxels < nv) throw std::runtime_error("task_env_tables: buffers too small");
    if (index_array) std::copy(T.index_array[env].begin(), T.index_array[env].end(), index_array);
    if (act_mask) std::copy(T.act_mask[env].begin(), T.act_mask[env].end(), act_mask);
    if (obs_mask) std::copy(T.obs_mask[env].begin(), T.obs_mask[env].end(), obs_mask);
    return T.corner_size[env];
}

void Batch::task_apply_policy_device(const void* actions, int stride, bool per_voxel, bool is_f32) {
    task_build();
    const int n = num_envs();
    if (stride < (per_voxel ? task_->max_vox : act_stride_)) throw std::runtime_error("task_apply_policy: stride too small");
    d_act_ext_ = nullptr;                    // the step reads the internal action buffer
    const int total = n * act_stride_;
    if (total == 0) return;
    const int grid = (total + 255) / 256;
    if (is_f32)
        policy_actions_kernel<float><<<grid, 256, 0, (cudaStream_t)stream_>>>(static_cast<const float*>(actions), stride, per_voxel ? 1 : 0,
                                                                             task_->d_rec, task_->d_act_vox, d_act_, act_stride_, n);
    else
        policy_actions_kernel<double><<<grid, 256, 0, (cudaStream_t)stream_>>>(static_cast<const double*>(actions), stride, per_voxel ? 1 : 0,
                                                                              task_->d_rec, task_->d_act_vox, d_act_, act_stride_, n);
    CUDA_CHECK(cudaGetLastError());
}

void Batch::task_apply_policy(const double* host_actions, int stride, bool per_voxel) {
    task_build();
    const size_t bytes = sizeof(double) * (size_t)stride * num_envs();
    double* d = nullptr;
    CUDA_CHECK(cudaMalloc(&d, bytes));
    try {
        CUDA_CHECK(cudaMemcpyAsync(d, host_actions, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream_));
        task_apply_policy_device(d, stride, per_voxel, false);
        CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    } catch (...) { cudaFree(d); throw; }
    cudaFree(d);
}

void Batch::task_outputs_device(void* reward, unsigned char* done, void* flat, int flat_stride, void* modular, int modular_stride,
                                void* other, int other_stride, bool is_f32, bool reset_done) {
    task_build();
    Task& T = *task_;
    TaskParams p;
    p.state = d_state_; p.state_rw = d_state_; p.state0 = d_state0_; p.saved = T.d_saved;
    p.envs = reinterpret_cast<const EnvDesc*>(d_envdesc_);
    p.tenv = T.d_rec; p.vox_idx = T.d_vox_idx; p.vox_mat = T.d_vox_mat;
    p.prev = T.d_prev; p.episode_len = T.d_episode; p.unstable = d_unstable_; p.sim_time = d_time_;
    p.reward = reward; p.done = done; p.flag_planes = T.d_flag_planes;
    p.flat = flat; p.flat_stride = flat_stride; p.modular = modular; p.modular_stride = modular_stride;
    p.other = other; p.other_stride = other_stride;
    p.n_envs = num_envs(); p.task = T.task; p.max_episode_steps = T.max_episode_steps; p.reset_done = reset_done ? 1 : 0;
    const TaskDesc& D = kTasks[T.task];
    const bool ort_pair = D.ort_rule || T.task == kTaskFlipper;        // the kernel needs the orientation after the step and at the last observation
    p.other_width = D.kernel_width;
    p.flat_prefix = T.flat_fixed;
    p.goal_axis = D.goal_axis; p.goal = D.goal; p.goal_bonus = D.bonus;
    p.fall_y = D.fall_y;
    p.ort_now = nullptr; p.ort_prev = T.d_ort_prev;
    p.ort_band = T.task == kTaskUpStepperFlat ? kPi / 12 : 0.0;
    p.flips = T.task == kTaskFlipper ? T.d_flips : nullptr;
    p.span_prev = T.task >= kTaskWingspan && T.task <= kTaskAreaMin ? T.d_span_prev : nullptr;
    p.goals = T.d_goals; p.goal_seed = T.goal_seed;
    const bool resample = T.task == kTaskCatcher && T.d_static0 && reset_done;
    p.envs_rw = resample ? reinterpret_cast<EnvDesc*>(d_envdesc_) : nullptr; p.static0 = T.d_static0; p.draws = T.d_flips; p.source = T.d_source;          // the ShapeChange tasks (25 ... 28)
    if (other && other_stride < D.other_width)
        throw std::runtime_error("task_outputs: the 'other' observation of this task is " + std::to_string(D.other_width) + " wide");
    const bool sighted = D.sight[0].terrain || D.sight[1].terrain;
    if (flat && flat_stride < (sighted ? T.flat_width : T.flat_fixed))
        throw std::runtime_error(sighted ? "task_outputs: the flat observation of this task is " + std::to_string(T.flat_width) + " wide (mevo_batch_task_dims)"
                                         : std::string("task_outputs: the flat observation is narrower than its fixed part"));
    if (ort_pair && (reward || done)) {
        object_orientation_strided_device(T.robot_obj, T.d_ort_now, 1, false);
        p.ort_now = T.d_ort_now;
    }
    const int grid = (p.n_envs + 3) / 4;
    if (is_f32) task_outputs_kernel<float><<<grid, 128, 0, (cudaStream_t)stream_>>>(p);
    else task_outputs_kernel<double><<<grid, 128, 0, (cudaStream_t)stream_>>>(p);
    CUDA_CHECK(cudaGetLastError());
    // what the kernel left open in "other" (and the flat observation), on the (possibly reset) state: get_ort_obs (base.py:258-267),
    // get_floor_obs / get_ceil_obs (base.py:269-372) as get_other_obs lays them out (base.py:652-744)
    const size_t el = is_f32 ? sizeof(float) : sizeof(double);
    if (D.ort_col >= 0) {
        const int obj = D.ort_of_package ? T.package_obj : T.robot_obj;
        if (other) object_orientation_strided_device(obj, static_cast<char*>(other) + D.ort_col * el, other_stride, is_f32, T.d_saved);
        if (flat && (D.ort_in_flat || (sighted && D.ort_col < T.flat_fixed))) object_orientation_strided_device(obj, static_cast<char*>(flat) + D.ort_col * el, flat_stride, is_f32, T.d_saved);
    }
    if (ort_pair) object_orientation_strided_device(T.robot_obj, T.d_ort_prev, 1, false, T.d_saved);
    for (const SightPart& sp : D.sight) {
        if (sp.terrain && other)
            sight_obs_strided_device(T.robot_obj, split_names(sp.terrain), sp.dist, 5.0, sp.ceil, static_cast<char*>(other) + sp.col * el,
                                     2 * sp.dist + 1, other_stride, is_f32, nullptr, T.d_saved);
        if (sp.terrain && flat)                          // same block, behind this environment's relative positions
            sight_obs_strided_device(T.robot_obj, split_names(sp.terrain), sp.dist, 5.0, sp.ceil, static_cast<char*>(flat) + sp.col * el,
                                     2 * sp.dist + 1, flat_stride, is_f32, T.d_flat_tail, T.d_saved);
    }
    if (reset_done) for (auto& e : envs_) e.host_fresh = false;
}

void Batch::task_outputs(double* reward, unsigned char* done, double* flat, int flat_stride, double* modular, int modular_stride,
                         double* other, int other_stride, bool reset_done) {
    task_build();
    Task& T = *task_;
    const size_t n = num_envs();
    const size_t b_rew = sizeof(double) * n, b_flat = flat ? sizeof(double) * n * flat_stride : 0,
                 b_mod = modular ? sizeof(double) * n * modular_stride : 0, b_oth = other ? sizeof(double) * n * other_stride : 0;
    const size_t need = b_rew + b_flat + b_mod + b_oth + n;
    if (need > T.scratch_bytes) {
        cudaFree(T.d_scratch); T.d_scratch = nullptr; T.scratch_bytes = 0;
        CUDA_CHECK(cudaMalloc(&T.d_scratch, need));
        T.scratch_bytes = need;
    }
    char* base = static_cast<char*>(T.d_scratch);
    double* d_rew = reinterpret_cast<double*>(base);
    double* d_flat = reinterpret_cast<double*>(base + b_rew);
    double* d_mod = reinterpret_cast<double*>(base + b_rew + b_flat);
    double* d_oth = reinterpret_cast<double*>(base + b_rew + b_flat + b_mod);
    unsigned char* d_done = reinterpret_cast<unsigned char*>(base + b_rew + b_flat + b_mod + b_oth);
    // (no reward and no done flags wanted: the reset observation -- the kernel then leaves the tasks' episode state alone)
    task_outputs_device(reward ? d_rew : nullptr, done ? d_done : nullptr, flat ? d_flat : nullptr, flat_stride, modular ? d_mod : nullptr, modular_stride,
                        other ? d_oth : nullptr, other_stride, false, reset_done);
    cudaStream_t s = (cudaStream_t)stream_;
    if (reward) CUDA_CHECK(cudaMemcpyAsync(reward, d_rew, b_rew, cudaMemcpyDeviceToHost, s));
    if (done) CUDA_CHECK(cudaMemcpyAsync(done, d_done, n, cudaMemcpyDeviceToHost, s));
    if (flat) CUDA_CHECK(cudaMemcpyAsync(flat, d_flat, b_flat, cudaMemcpyDeviceToHost, s));
    if (modular) CUDA_CHECK(cudaMemcpyAsync(modular, d_mod, b_mod, cudaMemcpyDeviceToHost, s));
    if (other) CUDA_CHECK(cudaMemcpyAsync(other, d_oth, b_oth, cudaMemcpyDeviceToHost, s));
    CUDA_CHECK(cudaStreamSynchronize(s));
}

// env.reset() of every environment (EvoGymBase.reset, base.py:47-51 -> Sim::revert(0)): state <- snapshot 0, episode
// counters cleared, the "previous" centres of mass refreshed; observations via task_outputs(reset_done = false)
// would add a step to the counters, so the reset observation is produced here.
void Batch::task_reset_device(void* flat, int flat_stride, void* modular, int modular_stride, void* other, int other_stride, bool is_f32) {
    task_build();
    reset_all();
    Task& T = *task_;
    CUDA_CHECK(cudaMemsetAsync(T.d_episode, 0xff, sizeof(int) * num_envs(), (cudaStream_t)stream_));   // -1: the outputs pass counts one
    CUDA_CHECK(cudaMemsetAsync(T.d_flips, 0, sizeof(int) * num_envs(), (cudaStream_t)stream_));
    if (T.d_goals)                                       // current goal <- 0; the lists stay (new ones: task_set_goals / task_draw_goals first)
        CUDA_CHECK(cudaMemset2DAsync(T.d_goals + 1, sizeof(int) * kGoalStride, 0, sizeof(int), num_envs(), (cudaStream_t)stream_));
    if (T.d_static0)                                     // Catcher-v0: snapshot 0 is the launch position the environment was built with
        CUDA_CHECK(cudaMemcpy2DAsync(static_cast<char*>(d_envdesc_) + offsetof(EnvDesc, static_off), sizeof(EnvDesc), T.d_static0,
                                     sizeof(unsigned long long), sizeof(unsigned long long), num_envs(), cudaMemcpyDeviceToDevice,
                                     (cudaStream_t)stream_));
    if (T.d_source) source_identity_kernel<<<(num_envs() + 127) / 128, 128, 0, (cudaStream_t)stream_>>>(T.d_source, num_envs());
    task_outputs_device(nullptr, nullptr, flat, flat_stride, modular, modular_stride, other, other_stride, is_f32, false);
}

void Batch::task_reset(double* flat, int flat_stride, double* modular, int modular_stride, double* other, int other_stride) {
    task_build();
    reset_all();
    Task& T = *task_;
    CUDA_CHECK(cudaMemsetAsync(T.d_episode, 0xff, sizeof(int) * num_envs(), (cudaStream_t)stream_));
    CUDA_CHECK(cudaMemsetAsync(T.d_flips, 0, sizeof(int) * num_envs(), (cudaStream_t)stream_));
    if (T.d_goals)                                       // current goal <- 0; the lists stay (new ones: task_set_goals / task_draw_goals first)
        CUDA_CHECK(cudaMemset2DAsync(T.d_goals + 1, sizeof(int) * kGoalStride, 0, sizeof(int), num_envs(), (cudaStream_t)stream_));
    if (T.d_static0)                                     // Catcher-v0: snapshot 0 is the launch position the environment was built with
        CUDA_CHECK(cudaMemcpy2DAsync(static_cast<char*>(d_envdesc_) + offsetof(EnvDesc, static_off), sizeof(EnvDesc), T.d_static0,
                                     sizeof(unsigned long long), sizeof(unsigned long long), num_envs(), cudaMemcpyDeviceToDevice,
                                     (cudaStream_t)stream_));
    if (T.d_source) source_identity_kernel<<<(num_envs() + 127) / 128, 128, 0, (cudaStream_t)stream_>>>(T.d_source, num_envs());
    task_outputs(nullptr, nullptr, flat, flat_stride, modular, modular_stride, other, other_stride, false);
}

// ---- terrain sight observations -----------------------------------------------------------------

struct Batch::Sight {
    std::string key;                     // object index + terrain names the tables were built for
    int epoch = -1;
    SightEnvRec* d_rec = nullptr;
    double2* d_pts = nullptr;
    void free_device() { cudaFree(d_rec); cudaFree(d_pts); d_rec = nullptr; d_pts = nullptr; }
};

void Batch::sight_free() {
    if (!sight_) return;
    sight_->free_device();
    delete sight_;
    sight_ = nullptr;
}

void Batch::sight_obs_device(int obj_index, const std::vector<std::string>& terrain, int sight_dist, double sight_range, bool ceil,
                             void* out, int stride, bool is_f32) {
    sight_obs_strided_device(obj_index, terrain, sight_dist, sight_range, ceil, out, stride, stride, is_f32);
}

void Batch::sight_obs_strided_device(int obj_index, const std::vector<std::string>& terrain, int sight_dist, double sight_range, bool ceil,
                                     void* out, int stride, int row_stride, bool is_f32, const int* env_off, const double* state_like) {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) finalize();
    if (sight_dist < 0 || sight_dist > 1000) throw std::runtime_error("sight_obs: bad sight_dist");
    if (stride < 2 * sight_dist + 1) throw std::runtime_error("sight_obs: stride smaller than 2 * sight_dist + 1");
    std::string key = std::to_string(obj_index);
    for (const std::string& t : terrain) key += "|" + t;
    if (!sight_) sight_ = new Sight();
    Sight& S = *sight_;
    const int n = num_envs();
    if (S.key != key || S.epoch != finalize_epoch_) {
        S.free_device();
        std::vector<SightEnvRec> rec(n);
        std::vector<double2> pts;
        std::vector<std::pair<std::vector<double>, int>> seen;      // static terrain point lists already stored -> offset
        for (int e = 0; e < n; e++) {
            const EnvHost& E = envs_[e];
            SightEnvRec& r = rec[e];
            std::memset(&r, 0, sizeof(r));
            r.Pd = (int)E.dyn_points.size();
            if (obj_index < 0 || obj_index >= (int)E.world.objects.size()) throw std::runtime_error("sight_obs: no such object");
            const VoxelObject& ob = E.world.objects[obj_index];
            r.o_start = E.point_local[ob.min_point];
            r.o_count = ob.max_point - ob.min_point + 1;
            if (r.o_start < 0) throw std::runtime_error("sight_obs: the observed object is static");
            std::vector<double> stat;
            std::vector<std::pair<int, int>> dyn;                  // simulated terrain objects: (first point, count) in the state
            for (const std::string& t : terrain) {
                const int ti = E.world.find_object(t);
                if (ti < 0) continue;                              // the reference would raise on an unknown name; envs never pass one
                const VoxelObject& tb = E.world.objects[ti];
                const int l0 = E.point_local[tb.min_point];
                if (l0 >= 0) {
                    dyn.emplace_back(l0, tb.max_point - tb.min_point + 1);
                } else {
                    for (int p = tb.min_point; p <= tb.max_point; p++) { stat.push_back(E.pos[2 * p]); stat.push_back(E.pos[2 * p + 1]); }
                }
            }
            // the kernel takes a maximum / minimum over all terrain points, so neighbouring ranges are one range (the 27 blocks of
            // Traverser-v0 follow each other in the state)
            std::sort(dyn.begin(), dyn.end());
            for (const auto& d : dyn) {
                if (r.n_dyn > 0 && r.dyn_start[r.n_dyn - 1] + r.dyn_count[r.n_dyn - 1] == d.first) { r.dyn_count[r.n_dyn - 1] += d.second; continue; }
                if (r.n_dyn >= 3) throw std::runtime_error("sight_obs: more than 3 separate ranges of simulated terrain points");
                r.dyn_start[r.n_dyn] = d.first; r.dyn_count[r.n_dyn] = d.second; r.n_dyn++;
            }
            int off = -1;
            for (const auto& sv : seen) if (sv.first == stat) { off = sv.second; break; }
            if (off < 0) {
                off = (int)pts.size();
                for (size_t k = 0; k + 1 < stat.size(); k += 2) pts.push_back(make_double2(stat[k], stat[k + 1]));
                seen.emplace_back(stat, off);
            }
            r.st_off = off; r.st_count = (int)stat.size() / 2;
        }
        CUDA_CHECK(cudaMalloc(&S.d_rec, sizeof(SightEnvRec) * n));
        CUDA_CHECK(cudaMemcpy(S.d_rec, rec.data(), sizeof(SightEnvRec) * n, cudaMemcpyHostToDevice));
        CUDA_CHECK(cudaMalloc(&S.d_pts, sizeof(double2) * std::max<size_t>(1, pts.size())));
        if (!pts.empty()) CUDA_CHECK(cudaMemcpy(S.d_pts, pts.data(), sizeof(double2) * pts.size(), cudaMemcpyHostToDevice));
        S.key = key; S.epoch = finalize_epoch_;
    }
    const int grid = (n + 3) / 4;
    if (is_f32)
        sight_kernel<float><<<grid, 128, 0, (cudaStream_t)stream_>>>(state_like ? state_like : d_state_, reinterpret_cast<const EnvDesc*>(d_envdesc_), S.d_rec, S.d_pts, n,
                                                                    sight_dist, sight_range, ceil ? 1 : 0, static_cast<float*>(out), stride, row_stride, env_off);
    else
        sight_kernel<double><<<grid, 128, 0, (cudaStream_t)stream_>>>(state_like ? state_like : d_state_, reinterpret_cast<const EnvDesc*>(d_envdesc_), S.d_rec, S.d_pts, n,
                                                                     sight_dist, sight_range, ceil ? 1 : 0, static_cast<double*>(out), stride, row_stride, env_off);
    CUDA_CHECK(cudaGetLastError());
}

void Batch::sight_obs(int obj_index, const std::vector<std::string>& terrain, int sight_dist, double sight_range, bool ceil, double* out_host,
                      int stride) {
    const size_t bytes = sizeof(double) * (size_t)stride * num_envs();
    double* d = nullptr;
    CUDA_CHECK(cudaSetDevice(device_));
    CUDA_CHECK(cudaMalloc(&d, bytes));
    try {
        sight_obs_device(obj_index, terrain, sight_dist, sight_range, ceil, d, stride, false);
        CUDA_CHECK(cudaMemcpyAsync(out_host, d, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream_));
        CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    } catch (...) { cudaFree(d); throw; }
    cudaFree(d);
}

}  // namespace mevo
