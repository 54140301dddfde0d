/* mevo_b200 -- C ABI of the B200-native batched soft-body engine.
 *
 * Drop-in boundary for ONE hot path of ModularEvoGym / EvoGym: the substep loop
 * Sim::step -> Environment::step -> PhysicsEngine::* (reference
 * evogym/simulator/SimulatorCPP/Sim.cpp:105-115, Environment.cpp:166-205, PhysicsEngine.cpp).
 * The reference exposes that path through the pybind11 class `Sim`
 * (SimulatorCPP/PythonBindings.cpp:26-44); each entry point below names the `Sim` / `Environment`
 * member it replaces.  The pybind11 module `simulator_cpp` shipped in modularevogym_b200/csrc is a
 * thin C++ layer over exactly these functions, and tests bind them with ctypes.
 *
 * Conventions
 *   - plain pointers and sizes only; no C++ or torch types; no exception crosses this boundary.
 *   - every function returning `int` returns >= 0 on success and MEVO_ERR (-1) on failure;
 *     mevo_last_error() then describes the failure (thread-local string).
 *   - one `mevo_batch` owns N independent environments on one GPU and one CUDA stream.  An
 *     "environment" is what the reference calls one `Sim`.
 *   - point arrays are (P, 2) row-major doubles x0,y0,x1,y1,... in the reference's global point
 *     numbering (the memory order of its column-major 2 x P Eigen matrices); engine units (metres,
 *     voxel = 0.1), i.e. what `Sim::pos_at_time` returns BEFORE `EvoSim` divides by VOXEL_SIZE.
 *   - the library needs a CUDA device (sm_100a); there is no CPU fallback: creation fails loudly.
 */
#ifndef MEVO_B200_H
#define MEVO_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define MEVO_ERR (-1)
#define MEVO_ABI_VERSION 1

typedef struct mevo_batch mevo_batch;

int mevo_abi_version(void);
const char* mevo_last_error(void);
int mevo_device_count(void);                       /* CUDA devices visible; MEVO_ERR if CUDA is unusable */

/* ---- lifetime ------------------------------------------------------------------------------- */
mevo_batch* mevo_batch_create(int device);         /* NULL on failure */
void mevo_batch_destroy(mevo_batch* b);

/* ---- world construction (host) -------------------------------------------------------------- */
/* Sim::init(x, y)  (Sim.cpp:26-28): opens a new environment with a world grid of w x h voxels.
 * Returns the environment id. */
int mevo_batch_begin_env(mevo_batch* b, int world_w, int world_h);

/* Sim::read_robot_from_array / Sim::read_object_from_array  (Sim.cpp:50-99).
 *   grid : h x w row-major voxel types (0 empty, 1 rigid, 2 soft, 3 h-act, 4 v-act, 5 fixed), row 0 = TOP
 *   conn : 2 x k, conn[i] and conn[k + i] are the two voxel indices (into grid) of connection i
 *   x, y : position of the bottom-left voxel in the world grid, in voxels
 * Returns 1 when the object was added, 0 when refused like the reference refuses it (message
 * "Error: ..." on stdout, world unchanged), MEVO_ERR on API misuse. */
int mevo_batch_add_object(mevo_batch* b, int env, const double* grid, int w, int h, const double* conn,
                          int k, const char* name, double x, double y, int is_robot);

/* One call for what a design loop does per environment, n_envs times (env.update(body, connections), evogym/envs/walk.py:105-120:
 * a new world from the environment's world file + the new robot): begin_env, the robot `robot_name` at voxel (rx, ry), then the
 * n_terrain terrain objects -- the same for every environment, given in the order EvoSim loads them (robots first, the others
 * alphabetical, evogym/sim.py:52-62; none of them may be a robot).  World size per environment: (world_w, world_h) grown to hold
 * the robot, like EvoWorld.add_object does.
 *   terrain_grids / terrain_conn : the objects' grids (h x w) and connection lists (2 x k) back to back; terrain_wh = w, h per object,
 *   terrain_k = k per object, terrain_xy = x, y per object, terrain_names = the names separated by '\n'
 *   robot_grids / robot_conn / robot_wh / robot_k : the same per environment
 * Returns the id of the first new environment (ids are consecutive), MEVO_ERR when an object is refused (the message names the
 * environment; environments added before it stay). */
int mevo_batch_add_robot_envs(mevo_batch* b, int n_envs, int world_w, int world_h, int n_terrain, const double* terrain_grids,
                              const int* terrain_wh, const double* terrain_conn, const int* terrain_k, const char* terrain_names,
                              const double* terrain_xy, const double* robot_grids, const int* robot_wh, const double* robot_conn,
                              const int* robot_k, const char* robot_name, double rx, double ry);

/* Packs all environments into HBM (struct-of-arrays blobs) and uploads.  Called implicitly by the
 * first step.  May be called again after adding environments/objects; state is preserved. */
int mevo_batch_finalize(mevo_batch* b);
/* Number of finalize calls so far (explicit, implicit at the first step, or the re-pack that mevo_env_translate_object does when
 * it moves an all-FIXED object).  Device pointers handed out and CUDA graphs captured in an earlier epoch are stale: a caller
 * that caches them compares epochs (modularevogym_b200/vec_env.py re-captures its graph). */
int mevo_batch_finalize_epoch(const mevo_batch* b);

/* ---- queries --------------------------------------------------------------------------------- */
int mevo_batch_num_envs(const mevo_batch* b);
int mevo_env_num_points(const mevo_batch* b, int env);
int mevo_env_num_edges(const mevo_batch* b, int env);
int mevo_env_num_objects(const mevo_batch* b, int env);
int mevo_env_object_index(const mevo_batch* b, int env, const char* name);      /* MEVO_ERR if unknown */
/* info[0..7] = is_robot, min_point_index, max_point_index, n_boxels, grid_w, grid_h, n_surface_boxels,
 * is_static (all-FIXED object pruned from integration) */
int mevo_env_object_info(const mevo_batch* b, int env, int obj, int* info);
/* per boxel 18 ints: cell_type, grid_index, TL,TR,BR,BL point ids, top,right,bot,left edge ids, 8 neighbour flags */
int mevo_env_object_boxels(const mevo_batch* b, int env, int obj, int* out);
int mevo_env_object_surface_boxels(const mevo_batch* b, int env, int obj, int* out);
/* Sim::get_actuator_indices (Sim.cpp:161-163): grid indices of the robot's actuators; returns the count */
int mevo_env_actuator_indices(const mevo_batch* b, int env, const char* robot, int* out, int cap);
/* topology arrays in global numbering; any pointer may be NULL */
int mevo_env_get_topology(const mevo_batch* b, int env, double* pos0, double* mass, unsigned char* fixed,
                          int* edge_a, int* edge_b, double* edge_k, double* edge_l0, unsigned char* edge_on_surface);

/* ---- host-only world builder ----------------------------------------------------------------- */
/* The same voxel-grid -> points/springs/boxels builder the batch uses (replaces ObjectCreator::
 * read_object_from_array + SimObject::compute_surface, ObjectCreator.cpp:193-641, SimObject.cpp:235-307),
 * usable without a GPU: topology checks, design tools, CPU tests of the host logic. */
typedef struct mevo_world mevo_world;
mevo_world* mevo_world_create(int world_w, int world_h);
void mevo_world_destroy(mevo_world* w);
int mevo_world_add_object(mevo_world* w, const double* grid, int gw, int gh, const double* conn, int k,
                          const char* name, double x, double y, int is_robot);
int mevo_world_num_points(const mevo_world* w);
int mevo_world_num_edges(const mevo_world* w);
int mevo_world_num_objects(const mevo_world* w);
int mevo_world_object_info(const mevo_world* w, int obj, int* info);
int mevo_world_object_boxels(const mevo_world* w, int obj, int* out);
int mevo_world_object_surface_boxels(const mevo_world* w, int obj, int* out);
int mevo_world_actuator_indices(const mevo_world* w, const char* robot, int* out, int cap);
int mevo_world_get_topology(const mevo_world* w, double* pos0, double* mass, unsigned char* fixed,
                            int* edge_a, int* edge_b, double* edge_k, double* edge_l0, unsigned char* edge_on_surface);

/* ---- actions --------------------------------------------------------------------------------- */
/* Sim::set_action (Sim.cpp:101-103 -> Environment::set_robot_action -> PhysicsEngine::update_actuator_goals):
 * action is n x 2 row-major (actuator grid_index, target length ratio).  Only the listed actuators' edges
 * are updated, exactly like the reference.  Returns 1, or 0 when refused (unknown robot / actuator). */
int mevo_env_set_action(mevo_batch* b, int env, const char* robot, const double* action, int n);
/* Batched form: values[env * stride + slot]; slot = actuators of the env's robots in load order, each
 * robot's actuators in mevo_env_actuator_indices order.  Consumed by the next step with apply_actions = 1. */
int mevo_batch_action_stride(const mevo_batch* b);            /* max actuators per env (after finalize) */
int mevo_batch_set_actions(mevo_batch* b, const double* host_values, int stride);
int mevo_batch_set_actions_device(mevo_batch* b, const double* device_values, int stride);
/* Zero-copy form: subsequent steps read values[env * stride + slot] straight from this DEVICE buffer
 * (it must stay valid until those steps have run; NULL returns to the internal buffer). */
int mevo_batch_bind_actions_device(mevo_batch* b, const double* device_values, int stride);

/* ---- stepping -------------------------------------------------------------------------------- */
/* Sim::step (Sim.cpp:105-115) for every environment, n_steps times, each 30 substeps of
 * Environment::step.  An environment whose substep reports "unstable" stops there: its time does not
 * advance (as in the reference) and its flag is set.  unstable_out (host, n_envs bytes) may be NULL.
 * Returns the number of unstable environments.  apply_actions: 1 = derive edge goals from the batched
 * action buffer on the device, 0 = keep the goals as last set by mevo_env_set_action. */
int mevo_batch_step(mevo_batch* b, int n_steps, int apply_actions, unsigned char* unstable_out);
/* as above with an explicit substep count (tests step one substep at a time = Environment::step) */
int mevo_batch_step_substeps(mevo_batch* b, int n_steps, int n_substeps, int apply_actions, unsigned char* unstable_out);
/* enqueue only; pair with mevo_batch_sync */
int mevo_batch_step_async(mevo_batch* b, int n_steps, int apply_actions);
int mevo_batch_sync(mevo_batch* b);
/* Allocation-free stepping.  NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * mevo_batch_step_async(b, n_steps, device_actions != NULL) on an already finalized batch, but the call itself allocates
 * nothing -- no cudaMalloc / cudaFree, no stream or event creation, no allocating library routine: the size classes run back to
 * back on the batch's stream and the group kernel's neighbour-list scratch is a file-scope `static __device__` array of
 * MEVO_STATIC_SCRATCH_BYTES (32 MiB, about 1000 environments of 20x20 robots) that every GPU loading the library holds until
 * the process exits.  A batch whose environments need more scratch than that is refused with MEVO_ERR and a message naming
 * both sizes; so is a batch that has not been finalized (finalize allocates, this call never does).
 *   device_actions     : caller's DEVICE buffer, values[env * stride + slot] (stride >= mevo_batch_action_stride);
 *                        NULL = keep the edge goals as last set (apply_actions = 0)
 *   device_unstable_out: caller's DEVICE buffer of n_envs bytes receiving the per-environment "unstable" flags;
 *                        NULL = the batch's own buffer (mevo_batch_unstable_device_ptr)
 * The state stays in the batch (mevo_env_get_state ...).  Enqueues on the batch's stream and does not sync.  All calls, of
 * all batches on one GPU, share the one array, which keeps the last call's contents (nothing in it needs to start at zero:
 * every launch rebuilds the lists it reads): calls that OVERLAP on one GPU -- two batches stepping concurrently through this
 * entry point -- race on it; serialise them or use mevo_batch_step_async. */
#define MEVO_STATIC_SCRATCH_BYTES (32u << 20)
int mevo_batch_step_static_scratch(mevo_batch* b, int n_steps, const double* device_actions, int stride,
                                   unsigned char* device_unstable_out);
int mevo_batch_last_launches(const mevo_batch* b);            /* kernels launched by the last step call */
/* on = 1: an environment whose step reports "unstable" is reverted to snapshot 0 (Sim::revert(0),
 * what the reference's envs do through reset() after the -3 reward, evogym/envs/walk.py:75-78) inside
 * the step kernel; its flag is still reported for that step.  Default 0 = reference Sim::step behaviour. */
int mevo_batch_set_auto_reset(mevo_batch* b, int on);
long long mevo_batch_unstable_events(mevo_batch* b);          /* unstable env-steps since creation */

/* ---- state ----------------------------------------------------------------------------------- */
/* Sim::get_time (Sim.cpp:135-137) */
long mevo_env_get_time(mevo_batch* b, int env);
int mevo_env_set_time(mevo_batch* b, int env, long t);
/* current pos / vel / pos_last / vel_true (Environment.h:28-34); any pointer may be NULL */
int mevo_env_get_state(mevo_batch* b, int env, double* pos, double* vel, double* pos_last, double* vel_true);
/* overwrite the state (Sim::revert = pos, vel, pos_last := pos, vel_true := vel; Environment.cpp:247-263) */
int mevo_env_set_state(mevo_batch* b, int env, const double* pos, const double* vel, const double* pos_last,
                       const double* vel_true);
/* length_eq_goal over all edges (Environment.h:41) */
int mevo_env_get_goal(mevo_batch* b, int env, double* goal);
int mevo_env_set_goal(mevo_batch* b, int env, const double* goal);
/* Sim::translate_object (Sim.cpp:148-151) without the snapshot (the caller owns history) */
int mevo_env_translate_object(mevo_batch* b, int env, const char* name, double dx, double dy);

/* batched reset = Sim::revert(0) for the listed environments (goals are NOT reset, like the reference) */
int mevo_batch_save_reset_state(mevo_batch* b);
int mevo_batch_reset(mevo_batch* b, const int* env_ids, int n);    /* env_ids NULL = all */
/* Batched snapshot history: Environment::save_snapshot / revert_to_snapshot(t) / Sim::revert (Environment.cpp:239-263,
 * Sim.cpp:126-133) for every environment at once, in a small fixed number of DEVICE-resident slots instead of the reference's
 * unbounded per-step std::map (a batch of 65 536 environments cannot keep every step).  Slot 0 is the reset snapshot
 * (mevo_batch_save_reset_state / mevo_batch_reset); slots 1 .. MEVO_SNAPSHOT_SLOTS - 1 are the caller's.  A snapshot holds pos,
 * vel and the environment's time, like the reference's; revert restores pos, vel, sets pos_last = pos and vel_true = vel, sets
 * the time to the snapshot's and leaves length_eq_goal alone ("goals are not reverted").  Reverting to an empty slot prints
 * the reference's "Error: Could not revert simulation - no save found for time argument." and does nothing.
 * mevo_batch_finalize drops all slots except 0 (the packing changes).  The pybind `Sim` keeps the reference's full per-step
 * history for its single environment on the host. */
#define MEVO_SNAPSHOT_SLOTS 8
int mevo_batch_save_snapshot(mevo_batch* b, int slot);
int mevo_batch_revert(mevo_batch* b, const int* env_ids, int n, int slot);     /* env_ids NULL = all environments */
/* Sim::pos_at_time / vel_at_time of one environment from a slot: (P, 2) doubles each, global numbering, either may be NULL.
 * Returns the time stored with the snapshot, -1 for an empty slot, < -1 on error. */
long mevo_env_read_snapshot(mevo_batch* b, int env, int slot, double* pos, double* vel);
/* centre of mass (mean of point positions) of object `obj` in every env: out = n_envs x 2 doubles (host) */
int mevo_batch_object_com(mevo_batch* b, int obj, double* out);
/* What a rollout loop reads after mevo_batch_step_async, in one host round trip: the unstable flags of that step (n_envs bytes)
 * and the centre of mass of object `obj` (n_envs x 2 doubles); either pointer may be NULL.  Returns the number of unstable
 * environments. */
int mevo_batch_fetch_step_results(mevo_batch* b, unsigned char* unstable_out, int obj, double* com_out);
/* same, into a caller-owned DEVICE buffer (n_envs x 2 doubles), enqueued on the batch's stream, no sync */
int mevo_batch_object_com_device(mevo_batch* b, int obj, double* device_out);
/* Sim::object_orientation_at_time (Environment.cpp:317-380) of object `obj` in every env, current state against
 * snapshot 0 (what get_ort_obs feeds the BridgeWalker / ObstacleTraverser / Climber-v2 observations, evogym/envs/base.py:
 * 258-267): out = n_envs doubles, radians in [0, 2 pi) */
int mevo_batch_object_orientation(mevo_batch* b, int obj, double* out);
int mevo_batch_object_orientation_device(mevo_batch* b, int obj, double* device_out);     /* device, on the stream, no sync */
/* get_floor_obs (ceil = 0) / get_ceil_obs (ceil = 1) of object `obj` in every env (evogym/envs/base.py:269-372, times
 * VOXEL_SIZE as BenchmarkBase returns them, base.py:411-415): 2 * sight_dist + 1 distances to the terrain below / above the
 * object's centre of mass, clipped to [0, sight_range].  terrain_csv: comma-separated names of the terrain objects (static or
 * simulated).  out = n_envs x stride, zero padded. */
int mevo_batch_sight_obs(mevo_batch* b, int obj, const char* terrain_csv, int sight_dist, double sight_range, int ceil, double* out,
                         int stride);
int mevo_batch_sight_obs_device(mevo_batch* b, int obj, const char* terrain_csv, int sight_dist, double sight_range, int ceil,
                                void* device_out, int stride, int is_f32);             /* device, on the stream, no sync */
/* Walker-v0 observation of object `obj` in every env (evogym/envs/walk.py:58-61 = get_vel_com_obs + get_relative_pos_obs,
 * evogym/envs/base.py:214-240), simulator units: row = [mean vel x, mean vel y, x_i - com_x (i < n), y_i - com_y
 * (i < n), zeros up to `stride`].  width = 2 + 2 * (largest point count of `obj` over the envs). */
int mevo_batch_object_obs_width(mevo_batch* b, int obj);
int mevo_batch_object_obs(mevo_batch* b, int obj, double* out, int stride);                 /* host, n_envs x stride */
int mevo_batch_object_obs_device(mevo_batch* b, int obj, double* device_out, int stride);   /* device, on the stream, no sync */

/* ---- task layer: the gym environments either side of Sim::step, batched on the device ----------------- */
/* What evogym/envs/walk.py:35-84 (WalkingFlat, Walker-v0), evogym/envs/manipulate.py:346-398 (PushSmallRect,
 * Pusher-v0) and manipulate.py:535-600 (ThrowSmallRect, Thrower-v0) compute around `EvoGymBase.step`:
 * policy action + 1 clipped to [0.6, 1.6] (base.py:384-391, 40-45), reward, done, the flat observation, and
 * ModularEvoGym's per-voxel observation (base.py:427-506, 746-753); plus what the vec-env worker does when an
 * environment is done (reset, return the reset observation) and gym's TimeLimit.  Values are in the units the
 * reference's environments see.  All *_device entry points enqueue on the batch's stream and do not sync. */
#define MEVO_TASK_WALKER 0
#define MEVO_TASK_PUSHER 1
#define MEVO_TASK_THROWER 2
#define MEVO_TASK_CLIMBER 3             /* Climb0, evogym/envs/climb.py:38-112 (Climber-v0): reward = rise of the centre of mass */
#define MEVO_TASK_OBSTACLE_TRAVERSER 4  /* WalkingBumpy2, evogym/envs/traverse.py:350-426 (ObstacleTraverser-v1): "other" observation =
                                           [vel com 2, orientation 1, get_floor_obs("robot", ["ground"], 5) 11]; the flat observation
                                           = [vel com, orientation, relative positions, floor sight] (traverse.py:396-401): in the flat rows
                                           of all sighted tasks the sight blocks follow each environment's own relative positions */
#define MEVO_TASK_BRIDGE_WALKER 5       /* SoftBridge, evogym/envs/walk.py:122-232 (BridgeWalker-v0): Walker's reward, goal at x > 60 voxels;
                                           "other" = [vel com 2, orientation 1]; flat = [vel com, orientation, relative positions] */
/* the other environments of traverse.py's StairsBase family: reward, "other" observation and flat observation as
 * MEVO_TASK_OBSTACLE_TRAVERSER; goal column 79 / 69 / 74.  ObstacleTraverser-v0 and DownStepper-v0 also end the episode with
 * reward - 3 when the robot lies on its back (orientation in (pi/2, 3pi/2), traverse.py:226-228, 325-327).  UpStepper-v0 has
 * that rule (with a wider band) in flat mode only (traverse.py:124-129): MEVO_TASK_UP_STEPPER is the modular-mode environment
 * (no such rule), MEVO_TASK_UP_STEPPER_FLAT the flat-mode one. */
#define MEVO_TASK_OBSTACLE_TRAVERSER_V0 6   /* WalkingBumpy, traverse.py:251-347 */
#define MEVO_TASK_UP_STEPPER 7              /* StepsUp, traverse.py:51-149 */
#define MEVO_TASK_DOWN_STEPPER 8            /* StepsDown, traverse.py:152-248 */
#define MEVO_TASK_JUMPER 9                  /* StationaryJump, evogym/envs/jump.py:14-118 (Jumper-v0, the environment of examples/CuCo):
                                               reward = 10 * rise - 5 * |drift| of the centre of mass, no goal;
                                               "other" = [vel com 2, get_floor_obs("robot", ["ground"], 2) 5]; modular mode */
#define MEVO_TASK_HURDLER 10                /* VerticalBarrier, traverse.py:446-536 (Hurdler-v0): StairsBase reward and observation, no goal,
                                               the orientation rule of MEVO_TASK_DOWN_STEPPER (traverse.py:514-516); modular mode */
#define MEVO_TASK_CLIMBER_V1 11             /* Climb1, evogym/envs/climb.py:131-231 (Climber-v1): MEVO_TASK_CLIMBER with the goal at y > 65 voxels */
#define MEVO_TASK_CAVE_CRAWLER 12           /* Duck, evogym/envs/walk.py:235-345 (CaveCrawler-v0): Walker's reward, goal at x > 69 voxels;
                                               "other" = [vel com 2, floor sight 11, ceiling sight 11] over "terrain"; modular mode */
#define MEVO_TASK_CLIMBER_V2 13             /* Climb2, evogym/envs/climb.py:226-320 (Climber-v2): reward = rise + 0.2 * advance, no goal;
                                               "other" = [vel com 2, orientation 1, ceiling sight 7 over "pipe"]; modular mode */
#define MEVO_TASK_CARRIER 14                /* CarrySmallRect, manipulate.py:83-199 (Carrier-v0): reward = half the advance of package and robot,
                                               10 x the package's fall below 3 voxels; goal at robot x > 99; observations as MEVO_TASK_PUSHER */
#define MEVO_TASK_PUSHER_V1 15              /* PushSmallRectOnOppositeSide, manipulate.py:419-516 (Pusher-v1): MEVO_TASK_PUSHER with the goal at x > 69 */
#define MEVO_TASK_BALANCER 16               /* Balance, evogym/envs/balance.py:14-135 (Balancer-v0): reward = approach of the centre of mass to
                                               (17, 5) voxels; "other" = [orientation]; flat = [orientation, relative positions] */
#define MEVO_TASK_BALANCER_V1 17            /* BalanceJump, balance.py:138-259 (Balancer-v1): the same towards (17.5, 6) */
#define MEVO_TASK_GAP_JUMPER 18             /* Gaps, traverse.py:663-782 (GapJumper-v0): StairsBase reward and observation (floor sight over the five
                                               platforms); a robot whose centre of mass drops below 4 voxels has fallen: reward - 3, done */
#define MEVO_TASK_PLATFORM_JUMPER 19        /* FloatingPlatform, traverse.py:539-660 (PlatformJumper-v0): the same over seven platforms, fall below
                                               0.4 m, and the orientation rule of MEVO_TASK_DOWN_STEPPER */
#define MEVO_TASK_CARRIER_V1 20             /* CarrySmallRectToTable, manipulate.py:202-315 (Carrier-v1): reward = the robot's approach to x = 40 voxels
                                               (the package term is overwritten in the reference, :239-240, and so here) + 10 x the package's
                                               fall below 6 voxels (4 past x = 20); no goal; observations as MEVO_TASK_PUSHER */
#define MEVO_TASK_LIFTER 21                 /* LiftSmallRect, manipulate.py:1207-1340 (Lifter-v0): reward = 10 x the package's rise + 10 x its approach
                                               to x = 5.5 voxels + 20 x the robot's fall below 3 voxels; observations as MEVO_TASK_PUSHER followed
                                               by the package's orientation (7 values) */
#define MEVO_TASK_FLIPPER 22                /* Flipping, evogym/envs/flip.py:15-140 (Flipper-v0): reward = change of the robot's unwrapped orientation
                                               (whole turns are counted per environment); done when the centre of mass passes x < 1 voxel;
                                               "other" = [orientation]; flat = [orientation, relative positions] */
#define MEVO_TASK_BEAM_TOPPLER 23           /* ToppleBeam, manipulate.py:887-1046 (BeamToppler-v0): `package` names the beam; reward from the beam's
                                               approach, displacement and fall; done + 1 when the beam's centre of mass is left of 2 voxels;
                                               "other" = [beam - robot com 2, vel com robot 2, vel com beam 2, beam orientation 1] */
#define MEVO_TASK_BEAM_SLIDER 24            /* SlideBeam, manipulate.py:1049-1204 (BeamSlider-v0): the same observation, reward = approach + advance */
#define MEVO_TASK_WINGSPAN 25               /* MaximizeXShape, evogym/envs/change_shape.py:287-385 (WingspanMazimizer-v0, the reference's spelling):
                                               reward = growth of the robot's extent in x; "other" = [0, 0]; flat = relative positions only */
#define MEVO_TASK_HEIGHT 26                 /* MaximizeYShape, change_shape.py:388-487 (HeightMaximizer-v0): the same in y */
#define MEVO_TASK_AREA_MAX 27               /* MaximizeShape, change_shape.py:85-183 (AreaMaximizer-v0): reward = 10 x growth of the signed area of the
                                               robot's convex hull, computed by the reference's own gift wrapping (one lane per environment) */
#define MEVO_TASK_AREA_MIN 28               /* MinimizeShape, change_shape.py:186-284 (AreaMinimizer-v0): 10 x its shrinkage */
#define MEVO_TASK_BIDIRECTIONAL_WALKER 29   /* BiWalk, multi_goal.py:91-205 (BidirectionalWalker-v0): walk to a list of x positions in turn; the reward is the
                                               approach to the current goal (+1 after the last), the flat observation ends with [goal index, x goal,
                                               x goal - com x].  The reference has no modular mode for it; goal lists: mevo_batch_task_set_goals */
#define MEVO_TASK_TRAVERSER 30               /* BlockSoup, traverse.py:784-911 (Traverser-v0): x progress across a pit filled with 27 loose rigid blocks;
                                               com x beyond voxel 59 ends the episode with + 2; "other" as task 18 with the pit and all blocks as terrain */
#define MEVO_TASK_CATCHER 31                 /* CatchSmallRect, manipulate.py:630-886 (Catcher-v0): a package tumbles down from two pegs; reward = approach of
                                               robot and package in x + 10 x the package's fall below 5 voxels; observations laid out as task 23's */
#define MEVO_TASK_UP_STEPPER_FLAT 32         /* StepsUp with mode=None, traverse.py:51-149: as task 7 plus "orientation in (pi/2 - pi/12, 3pi/2 + pi/12)
                                               ends the episode with - 3" (traverse.py:124-129) */
#define MEVO_DONE 1            /* done byte, bit 0: episode over (goal, unstable or step limit) */
#define MEVO_DONE_TRUNCATED 2  /* bit 1: over only because of max_episode_steps (gym TimeLimit.truncated) */
#define MEVO_DONE_UNSTABLE 4   /* bit 2: Sim::step reported an unstable simulation (reward already carries the -3) */
/* package: used by MEVO_TASK_PUSHER / _THROWER / _CARRIER / _CARRIER_V1 / _PUSHER_V1 / _LIFTER / _BEAM_* only (NULL otherwise).  max_episode_steps: 0 = no limit (gym registration: Walker-v0 500). */
int mevo_batch_task_configure(mevo_batch* b, int task, const char* robot, const char* package, int max_episode_steps);
/* dims[0..4] = flat observation width, modular observation width (9 x voxels), "other" observation width,
 * largest voxel count of a robot grid, action stride */
int mevo_batch_task_dims(mevo_batch* b, int* dims);
/* per environment, host: obs_index_array (voxels x 4: TL, TR, BL, BR point index, -1 = empty voxel; base.py:427-466),
 * act_mask (base.py:536-545), obs_mask (base.py:547-557); any pointer may be NULL; returns voxel_corner_size */
int mevo_batch_task_env_tables(mevo_batch* b, int env, int* index_array, int* act_mask, int* obs_mask, int cap_voxels);
/* transform_to_modular_obs on a bare h x w voxel grid (row 0 = top), host only, no GPU needed; out: h*w*4 ints */
int mevo_modular_corner_index(const int* body, int h, int w, int* out);
/* policy outputs -> action buffer of the next step.  per_voxel = 0: actions[env * stride + slot] (flat mode);
 * per_voxel = 1: one value per voxel of the robot grid, actuator voxels picked in row-major order (walk.py:41-42) */
int mevo_batch_task_apply_policy(mevo_batch* b, const double* host_actions, int stride, int per_voxel);
int mevo_batch_task_apply_policy_device(mevo_batch* b, const void* device_actions, int stride, int per_voxel, int is_f32);
/* after mevo_batch_step*: reward (n_envs), done bytes (n_envs), observations (n_envs x stride, zero padded); any
 * output may be NULL.  reset_done = 1: environments that are done are reverted to snapshot 0 first and their
 * observation is the reset observation (vec-env semantics). */
/* NOTE on unstable steps: Sim::step leaves the partially advanced state live, saves no snapshot and does not advance time; the
 * environment's getters read the last SAVED snapshot (Sim.cpp:105-115, history[sim_time]).  The task layer keeps that snapshot's
 * positions and velocities per environment (written by every call for the environments whose step went through, or that it reset):
 * reward, done, the observation returned for an unstable step and the "before" values of the steps after it are the reference's
 * also with reset_done = 0, while the live state (mevo_env_get_state, the next mevo_batch_step) stays partially advanced as in
 * the reference.  With reset_done = 1 an unstable environment is reverted and reports the reset observation, like the
 * reference's vec-env worker. */
int mevo_batch_task_outputs(mevo_batch* b, double* reward, unsigned char* done, double* flat, int flat_stride, double* modular,
                            int modular_stride, double* other, int other_stride, int reset_done);
int mevo_batch_task_outputs_device(mevo_batch* b, void* reward, unsigned char* done, void* flat, int flat_stride, void* modular,
                                   int modular_stride, void* other, int other_stride, int is_f32, int reset_done);
/* env.reset() of every environment (EvoGymBase.reset, base.py:47-51): state <- snapshot 0, observation returned */
int mevo_batch_task_reset(mevo_batch* b, double* flat, int flat_stride, double* modular, int modular_stride, double* other,
                          int other_stride);
int mevo_batch_task_reset_device(mevo_batch* b, void* flat, int flat_stride, void* modular, int modular_stride, void* other,
                                 int other_stride, int is_f32);

/* Catcher-v0 (task 31): CatchSmallRect.reset re-draws where the package is launched from and rebuilds its world
 * (manipulate.py:646-671, 815).  Environments are BUILT with one draw each (the host's); when mevo_batch_task_outputs(reset_done = 1)
 * restarts an episode it gives the environment the launch position -- the two static pegs and the package's snapshot-0 points -- of
 * a uniformly drawn environment of the batch, i.e. one of the batch's own draws.  out[env] (n_envs ints) = the environment whose
 * launch position `env` runs now.  Host-side views of `env`'s world (the static rows of mevo_env_get_state, mevo_env_render) keep
 * showing the pegs where it was built; mevo_batch_task_reset returns every environment to its own. */
int mevo_batch_task_launch_sources(mevo_batch* b, int* out);

/* Optional second form of the done byte: `device_planes` points to 3 * n_envs bytes on the device; every later
 * mevo_batch_task_outputs[_device] call that writes `done` also writes done / truncated / unstable as 0 or 1 into the three planes
 * (the layout of a bool tensor of shape (3, n): what SubprocVecEnv's `done` array and the TimeLimit / unstable infos are).  The
 * buffer belongs to the caller and must outlive the calls; NULL switches the planes off.  Cleared by mevo_batch_task_configure. */
int mevo_batch_task_bind_flag_planes(mevo_batch* b, void* device_planes);

/* BidirectionalWalker-v0 (evogym/envs/multi_goal.py:91-205, task 29): each environment walks to a list of x positions in turn
 * (WalkToX, multi_goal.py:62-88).  The reference draws a new list in __init__ and in every reset() from Python's global `random`
 * (set_random_goals(20, 50, 100), multi_goal.py:115-129).  Here every environment gets a list when the task tables are built and a
 * new one whenever mevo_batch_task_outputs(reset_done = 1) restarts its episode, drawn on the device by the same rule from a
 * counter-based generator (identically distributed lists, not Python's stream); mevo_batch_task_reset keeps the lists and returns
 * to their first goal, so a host that wants its own lists (or Python's stream) sets them before it resets.
 *   set_goals: `goals` holds `stride` ints per environment, the first counts[i] (1 ... 61) are that environment's x goals in voxels
 *   draw_goals: new lists for all environments from `seed` (also the seed of later device draws)
 *   get_goals: list of one environment (at most `cap` values), its length and the index of the current goal */
int mevo_batch_task_set_goals(mevo_batch* b, int env_first, int n_envs, const int* counts, const int* goals, int stride);
int mevo_batch_task_draw_goals(mevo_batch* b, unsigned long long seed);
int mevo_batch_task_get_goals(mevo_batch* b, int env, int* count, int* current, int* goals, int cap);

/* ---- headless rasteriser ------------------------------------------------------------------------------- */
/* Interface::render (SimulatorCPP/Interface.cpp:55-135: background, voxel grid lines, one quad per voxel coloured by cell
 * type / actuation, surface edges) of one environment's current state, rasterised by a CUDA kernel into an RGB8 image of
 * res_h x res_w x 3 bytes, row 0 = top -- the buffer Camera::get_image hands to EvoViewer.render('img')
 * (evogym/viewer.py:239-250).  Camera: centre (cam_x, cam_y) and full view size (cam_w, cam_h) in engine units (metres), as
 * Camera::set_pos / set_size receive them.  The reference's OpenGL path itself is unavailable on this box. */
#define MEVO_RENDER_HIDE_BACKGROUND 1
#define MEVO_RENDER_HIDE_GRID 2
#define MEVO_RENDER_HIDE_EDGES 4
#define MEVO_RENDER_HIDE_BOXELS 8
#define MEVO_RENDER_DONT_CLEAR 16
int mevo_env_render(mevo_batch* b, int env, double cam_x, double cam_y, double cam_w, double cam_h, int res_w, int res_h, int flags,
                    unsigned char* out);
int mevo_env_render_device(mevo_batch* b, int env, double cam_x, double cam_y, double cam_w, double cam_h, int res_w, int res_h,
                           int flags, unsigned char* device_out);

/* ---- CUDA graphs --------------------------------------------------------------------------------------- */
/* Everything the *_device / *_async entry points enqueue on the batch's stream between capture_begin and capture_end
 * (policy actions, the step launches of all size classes, task outputs, observations) is recorded into one CUDA graph
 * instead of being executed; mevo_batch_graph_launch replays it with a single driver call.  Device pointers passed to the
 * captured calls are baked into the graph (use persistent buffers); every captured entry point must have been called once
 * before the capture (first calls configure kernels and build tables); mevo_batch_finalize invalidates all graphs.
 * capture_end returns the graph id. */
int mevo_batch_capture_begin(mevo_batch* b);
int mevo_batch_capture_end(mevo_batch* b);
int mevo_batch_graph_launch(mevo_batch* b, int graph);
int mevo_batch_graph_destroy(mevo_batch* b, int graph);

/* device pointers for zero-copy consumers (valid until the next finalize) */
double* mevo_batch_actions_device_ptr(mevo_batch* b);
const unsigned char* mevo_batch_unstable_device_ptr(mevo_batch* b);
void* mevo_batch_cuda_stream(mevo_batch* b);
long long mevo_batch_device_bytes(const mevo_batch* b);

/* ---- test hooks ------------------------------------------------------------------------------ */
/* record the compacted collision-candidate list of the first substep of subsequent step calls */
int mevo_batch_enable_candidate_dump(mevo_batch* b, int cap_per_env);
/* out: cap x 4 ints (object_a, boxel_a, object_b, boxel_b); returns the number of candidates found.
 * Pairs of two all-fixed objects are never generated (they cannot move anything); pairs inside one
 * object are listed once (a < b) and applied twice, which is what the reference's ordered descent does. */
int mevo_env_read_candidates(mevo_batch* b, int env, int* out, int cap);
/* record collision_forces (PhysicsEngine.cpp:715-751: the contact response accumulated by resolve_collisions) of the first
 * substep of subsequent step calls; read as (P, 2) doubles in global numbering.  Points of pruned all-FIXED objects read 0:
 * the engine does not accumulate forces on points that can never move. */
int mevo_batch_enable_force_dump(mevo_batch* b);
int mevo_env_read_collision_forces(mevo_batch* b, int env, double* out);
int mevo_env_error_flags(mevo_batch* b, int env);             /* always 0 since the candidate buffers are drained in rounds; kept for ABI stability */

#ifdef __cplusplus
}
#endif
#endif /* MEVO_B200_H */
