/* TEST INFRASTRUCTURE ONLY -- plain-C restatement of the reference's substep loop.
 *
 * Not part of the product: only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline leg may
 * load this library.  Parity status: PINNED -- every function here is checked against the
 * unmodified reference engine compiled from /root/reference (oracle/_ref/libmevo_ref.so) by
 * tests/test_oracle_vs_ref.py, and against the committed fixtures in tests/golden/ generated from
 * that same library (tests/golden/make_golden.py).
 *
 * All point arrays are (P, 2) row-major doubles: x0,y0,x1,y1,... (the memory order of the
 * reference's column-major 2 x P Eigen matrices, SimulatorCPP/Environment.h:28-34).
 */
#ifndef MEVO_ORACLE_H
#define MEVO_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mo_env mo_env;

/* boxels: n_boxels x 12 ints = cell_type, grid_index, TL, TR, BR, BL point ids, top, right, bot,
 * left edge ids, object index, is_surface.  Boxels are listed object by object, in the reference's
 * boxel order.  obj_info: n_obj x 4 ints = is_robot, first_boxel, n_boxels, grid_w. */
mo_env* mo_create(int P, int E, int n_obj, int n_boxels,
                  const double* pos0, const double* mass, const unsigned char* fixed,
                  const int* ea, const int* eb, const double* ek, const double* el0,
                  const unsigned char* on_surface, const int* boxels, const int* obj_info);
void mo_destroy(mo_env* e);

void mo_get_state(const mo_env* e, double* pos, double* vel, double* pos_last, double* vel_true);
void mo_set_state(mo_env* e, const double* pos, const double* vel, const double* pos_last,
                  const double* vel_true);
void mo_get_edge_state(const mo_env* e, double* leq, double* goal);
void mo_set_edge_state(mo_env* e, const double* leq, const double* goal);
void mo_get_collision_forces(const mo_env* e, double* out);

/* PhysicsEngine::update_actuator_goals (PhysicsEngine.cpp:38-97): boxel_ids are GLOBAL boxel
 * indices (row in `boxels`) of the actuators named by the action, in action order. */
void mo_set_action(mo_env* e, const int* boxel_ids, const double* values, int n);

/* phases of Environment::step (Environment.cpp:166-205) */
int  mo_resolve_collisions(mo_env* e, int* cand_out, int max_cand); /* returns #candidate leaf pairs */
void mo_update_actuators(mo_env* e);
void mo_step_rk4(mo_env* e, double dt);
void mo_resolve_edge_constraints(mo_env* e);
void mo_update_true_vel(mo_env* e, double dt);
int  mo_is_any_robot_self_colliding(mo_env* e);

int  mo_substep(mo_env* e);               /* Environment::step; 1 = unstable */
int  mo_step(mo_env* e, int n_substeps);  /* Sim::step's loop (Sim.cpp:105-115); 1 = unstable (aborted) */
long mo_contact_count(const mo_env* e);   /* contacts resolved by the last mo_resolve_collisions */

#ifdef __cplusplus
}
#endif
#endif
