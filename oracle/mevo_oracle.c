/* TEST INFRASTRUCTURE ONLY -- plain-C restatement of the reference's substep loop.
 * See mevo_oracle.h for the parity status (PINNED against oracle/_ref and tests/golden).
 *
 * Every arithmetic expression below keeps the reference's operand order and association; the
 * file is compiled with -ffp-contract=off so no product/sum pair is fused.  Citations are into
 * /root/reference/evogym/simulator/SimulatorCPP/.
 *
 * One deliberate difference: collision candidates are enumerated in ascending (leaf a, leaf b)
 * order instead of the bounding-box-tree descent order (PhysicsEngine.cpp:752-817).  The candidate
 * SET is identical (internal tree boxes are exact unions of their leaves, SimObject.cpp:210-218),
 * only the order in which contact forces are summed into collision_forces differs.
 */
#include "mevo_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* PhysicsEngine.h:17-27 */
static const double viscous_drag = 0.1;
static const double gravity = 110;
static const double collision_const_ground = 2000000.0 * 7;
static const double collision_const_obj = 300000.0 * 7;
static const double collision_vel_damping = 140.0;
static const double collision_base_dist_additive = 0.005;
static const double dynamic_friction_const = 0.2;
static const double friction_const = 150.0 * 8;

enum { CELL_EMPTY = 0, CELL_RIGID, CELL_SOFT, CELL_ACT_H, CELL_ACT_V, CELL_FIXED };
enum { B_TYPE = 0, B_GRID, B_TL, B_TR, B_BR, B_BL, B_ET, B_ER, B_EB, B_EL, B_OBJ, B_SURF, B_STRIDE };

struct mo_env {
    int P, E, n_obj, n_boxels, n_leaves;
    double *pos, *vel, *pos_last, *vel_true, *mass;
    unsigned char* fixed;
    int *ea, *eb;
    double *ek, *el0, *leq, *goal;
    unsigned char* on_surface;
    int* boxels;   /* n_boxels x B_STRIDE */
    int* obj_info; /* n_obj x 4: is_robot, first_boxel, n_boxels, grid_w */
    int* leaves;   /* global boxel index of each surface boxel, ascending */
    double* bb;    /* n_leaves x 4: minx, miny, maxx, maxy -- refit at the start of each substep */
    double* cf;    /* collision_forces (P,2) */
    double *scratch; /* RK4 + constraint temporaries */
    long contacts;
};

static void* dup_mem(const void* src, size_t bytes) {
    void* p = malloc(bytes ? bytes : 1);
    if (src && bytes) memcpy(p, src, bytes);
    return p;
}

mo_env* mo_create(int P, int E, int n_obj, int n_boxels, const double* pos0, const double* mass,
                  const unsigned char* fixed, const int* ea, const int* eb, const double* ek,
                  const double* el0, const unsigned char* on_surface, const int* boxels,
                  const int* obj_info) {
    mo_env* e = (mo_env*)calloc(1, sizeof(mo_env));
    e->P = P; e->E = E; e->n_obj = n_obj; e->n_boxels = n_boxels;
    e->pos = (double*)dup_mem(pos0, sizeof(double) * 2 * P);
    e->pos_last = (double*)dup_mem(pos0, sizeof(double) * 2 * P);      /* Environment.cpp:57-58 */
    e->vel = (double*)calloc(2 * P + 1, sizeof(double));
    e->vel_true = (double*)calloc(2 * P + 1, sizeof(double));
    e->mass = (double*)dup_mem(mass, sizeof(double) * P);
    e->fixed = (unsigned char*)dup_mem(fixed, P);
    e->ea = (int*)dup_mem(ea, sizeof(int) * E);
    e->eb = (int*)dup_mem(eb, sizeof(int) * E);
    e->ek = (double*)dup_mem(ek, sizeof(double) * E);
    e->el0 = (double*)dup_mem(el0, sizeof(double) * E);
    e->leq = (double*)dup_mem(el0, sizeof(double) * E);                /* Environment.cpp:104-106 */
    e->goal = (double*)dup_mem(el0, sizeof(double) * E);
    e->on_surface = (unsigned char*)dup_mem(on_surface, E);
    e->boxels = (int*)dup_mem(boxels, sizeof(int) * B_STRIDE * n_boxels);
    e->obj_info = (int*)dup_mem(obj_info, sizeof(int) * 4 * n_obj);
    e->leaves = (int*)malloc(sizeof(int) * (n_boxels + 1));
    e->n_leaves = 0;
    for (int b = 0; b < n_boxels; b++)
        if (e->boxels[b * B_STRIDE + B_SURF]) e->leaves[e->n_leaves++] = b;
    e->bb = (double*)calloc(4 * e->n_leaves + 1, sizeof(double));
    e->cf = (double*)calloc(2 * P + 1, sizeof(double));
    e->scratch = (double*)calloc(4 * P * 7 + 6 * E + 8, sizeof(double));
    return e;
}

void mo_destroy(mo_env* e) {
    if (!e) return;
    free(e->pos); free(e->vel); free(e->pos_last); free(e->vel_true); free(e->mass); free(e->fixed);
    free(e->ea); free(e->eb); free(e->ek); free(e->el0); free(e->leq); free(e->goal);
    free(e->on_surface); free(e->boxels); free(e->obj_info); free(e->leaves); free(e->bb);
    free(e->cf); free(e->scratch); free(e);
}

void mo_get_state(const mo_env* e, double* pos, double* vel, double* pos_last, double* vel_true) {
    size_t n = sizeof(double) * 2 * e->P;
    if (pos) memcpy(pos, e->pos, n);
    if (vel) memcpy(vel, e->vel, n);
    if (pos_last) memcpy(pos_last, e->pos_last, n);
    if (vel_true) memcpy(vel_true, e->vel_true, n);
}
void mo_set_state(mo_env* e, const double* pos, const double* vel, const double* pos_last,
                  const double* vel_true) {
    size_t n = sizeof(double) * 2 * e->P;
    if (pos) memcpy(e->pos, pos, n);
    if (vel) memcpy(e->vel, vel, n);
    if (pos_last) memcpy(e->pos_last, pos_last, n);
    if (vel_true) memcpy(e->vel_true, vel_true, n);
}
void mo_get_edge_state(const mo_env* e, double* leq, double* goal) {
    if (leq) memcpy(leq, e->leq, sizeof(double) * e->E);
    if (goal) memcpy(goal, e->goal, sizeof(double) * e->E);
}
void mo_set_edge_state(mo_env* e, const double* leq, const double* goal) {
    if (leq) memcpy(e->leq, leq, sizeof(double) * e->E);
    if (goal) memcpy(e->goal, goal, sizeof(double) * e->E);
}
void mo_get_collision_forces(const mo_env* e, double* out) { memcpy(out, e->cf, sizeof(double) * 2 * e->P); }
long mo_contact_count(const mo_env* e) { return e->contacts; }

/* ------------------------------------------------------------------------------------------ */
/* PhysicsEngine::update_actuator_goals, PhysicsEngine.cpp:38-97 */
void mo_set_action(mo_env* e, const int* boxel_ids, const double* values, int n) {
    int* count = (int*)calloc(e->E + 1, sizeof(int));
    unsigned char* touched = (unsigned char*)calloc(e->E + 1, 1);
    for (int i = 0; i < n; i++) {                                   /* :43-69 zero the touched goals */
        const int* b = e->boxels + boxel_ids[i] * B_STRIDE;
        int e0 = -1, e1 = -1;
        if (b[B_TYPE] == CELL_ACT_H) { e0 = b[B_ET]; e1 = b[B_EB]; }
        if (b[B_TYPE] == CELL_ACT_V) { e0 = b[B_EL]; e1 = b[B_ER]; }
        if (e0 < 0) continue;
        e->goal[e0] = 0; e->goal[e1] = 0;
        count[e0] = 0; count[e1] = 0;
        touched[e0] = 1; touched[e1] = 1;
    }
    for (int i = 0; i < n; i++) {                                   /* :71-92 accumulate */
        const int* b = e->boxels + boxel_ids[i] * B_STRIDE;
        int e0 = -1, e1 = -1;
        if (b[B_TYPE] == CELL_ACT_H) { e0 = b[B_ET]; e1 = b[B_EB]; }
        if (b[B_TYPE] == CELL_ACT_V) { e0 = b[B_EL]; e1 = b[B_ER]; }
        if (e0 < 0) continue;
        e->goal[e0] += values[i]; e->goal[e1] += values[i];
        count[e0] += 1; count[e1] += 1;
    }
    for (int k = 0; k < e->E; k++)                                  /* :94-96 */
        if (touched[k]) e->goal[k] *= e->el0[k] / count[k];
    free(count); free(touched);
}

/* PhysicsEngine::update_actuators, PhysicsEngine.cpp:99-119 */
void mo_update_actuators(mo_env* e) {
    const double convergence_factor = 0.006;
    for (int k = 0; k < e->E; k++) {
        int a = e->ea[k], b = e->eb[k];
        double dx = e->pos[2 * a] - e->pos[2 * b];
        double dy = e->pos[2 * a + 1] - e->pos[2 * b + 1];
        double dist = sqrt(dx * dx + dy * dy);
        double additive = (e->goal[k] - dist) * convergence_factor;
        e->leq[k] = dist + additive;
    }
}

/* PhysicsEngine::get_pos_vel_slope, PhysicsEngine.cpp:487-600.
 * sp, sv: stage positions / velocities (P,2).  out_dp = stage velocity, out_dv = force / mass. */
static void slope(const mo_env* e, const double* sp, const double* sv, double* out_dp, double* out_dv,
                  double* F, double* ef) {
    const int P = e->P, E = e->E;
    for (int i = 0; i < 2 * P; i++) F[i] = 0;                                       /* :494 */
    for (int k = 0; k < E; k++) {                                                   /* :498-528 */
        int a = e->ea[k], b = e->eb[k];
        double dx = sp[2 * a] - sp[2 * b];
        double dy = sp[2 * a + 1] - sp[2 * b + 1];
        double dist = sqrt(dx * dx + dy * dy);
        double mag = (dist - e->leq[k]) * e->ek[k] / dist;
        ef[2 * k] = dx * mag;
        ef[2 * k + 1] = dy * mag;
    }
    for (int k = 0; k < E; k++) { int a = e->ea[k]; F[2 * a] -= ef[2 * k]; F[2 * a + 1] -= ef[2 * k + 1]; } /* :530 */
    for (int k = 0; k < E; k++) { int b = e->eb[k]; F[2 * b] += ef[2 * k]; F[2 * b + 1] += ef[2 * k + 1]; } /* :531 */
    for (int i = 0; i < 2 * P; i++) F[i] += sv[i] * (-viscous_drag);                /* :540 */
    for (int p = 0; p < P; p++) F[2 * p + 1] += e->mass[p] * (-gravity);            /* :549 */
    for (int i = 0; i < 2 * P; i++) F[i] += e->cf[i];                               /* :558 */
    for (int p = 0; p < P; p++) {                                                   /* :569-577 */
        double y = sp[2 * p + 1];
        int hit = y < 0;
        double normal_force = (hit ? y : 0) * (-collision_const_ground);
        F[2 * p + 1] += normal_force;
        double friction_temp = hit ? tanh(sv[2 * p] * normal_force * dynamic_friction_const) : 0;
        F[2 * p] += (-10 * dynamic_friction_const) * (normal_force * friction_temp);
    }
    for (int i = 0; i < 2 * P; i++) out_dp[i] = sv[i];                              /* :592 */
    for (int p = 0; p < P; p++) {                                                   /* :597 */
        out_dv[2 * p] = F[2 * p] / e->mass[p];
        out_dv[2 * p + 1] = F[2 * p + 1] / e->mass[p];
    }
}

/* PhysicsEngine::step_rk4, PhysicsEngine.cpp:427-485 */
void mo_step_rk4(mo_env* e, double dt) {
    const int n = 2 * e->P;
    double* s = e->scratch;
    double *k1p = s, *k1v = s + n, *k2p = s + 2 * n, *k2v = s + 3 * n, *k3p = s + 4 * n, *k3v = s + 5 * n;
    double *k4p = s + 6 * n, *k4v = s + 7 * n, *tp = s + 8 * n, *tv = s + 9 * n, *F = s + 10 * n;
    double* ef = s + 14 * n; /* 2E doubles; scratch holds 14n + 6E */
    slope(e, e->pos, e->vel, k1p, k1v, F, ef);
    for (int i = 0; i < n; i++) { k1p[i] *= dt; k1v[i] *= dt; }
    for (int i = 0; i < n; i++) { tp[i] = e->pos[i] + (k1p[i] * 0.5); tv[i] = e->vel[i] + (k1v[i] * 0.5); }
    slope(e, tp, tv, k2p, k2v, F, ef);
    for (int i = 0; i < n; i++) { k2p[i] *= dt; k2v[i] *= dt; }
    for (int i = 0; i < n; i++) { tp[i] = e->pos[i] + (k2p[i] * 0.5); tv[i] = e->vel[i] + (k2v[i] * 0.5); }
    slope(e, tp, tv, k3p, k3v, F, ef);
    for (int i = 0; i < n; i++) { k3p[i] *= dt; k3v[i] *= dt; }
    for (int i = 0; i < n; i++) { tp[i] = e->pos[i] + k3p[i]; tv[i] = e->vel[i] + k3v[i]; }
    slope(e, tp, tv, k4p, k4v, F, ef);
    for (int i = 0; i < n; i++) { k4p[i] *= dt; k4v[i] *= dt; }
    for (int i = 0; i < n; i++) {                                                   /* :476-484 */
        double dp = (k1p[i] + 2 * k2p[i] + 2 * k3p[i] + k4p[i]) / 6;
        double dv = (k1v[i] + 2 * k2v[i] + 2 * k3v[i] + k4v[i]) / 6;
        if (e->fixed[i / 2]) { dp = 0.0f; dv = 0.0f; }
        e->pos[i] += dp;
        e->vel[i] += dv;
    }
}

/* PhysicsEngine::resolve_edge_constraints, PhysicsEngine.cpp:608-713 */
void mo_resolve_edge_constraints(mo_env* e) {
    const double thresh_diff = 0.25, thresh_diff_rigid = 0.03;
    const int E = e->E;
    double* corr = e->scratch; /* 2E */
    for (int k = 0; k < E; k++) {                                                   /* :613-634 */
        int a = e->ea[k], b = e->eb[k];
        double dx = e->pos[2 * a] - e->pos[2 * b];
        double dy = e->pos[2 * a + 1] - e->pos[2 * b + 1];
        double dist = sqrt(dx * dx + dy * dy);
        double over_mag = (dist - e->el0[k] * (1 + thresh_diff)) / (dist * 2);
        double under_mag = (dist - e->el0[k] * (1 - thresh_diff)) / (dist * 2);
        double ox = dx * over_mag, oy = dy * over_mag;
        double ux = dx * under_mag, uy = dy * under_mag;
        double stress = dist / e->el0[k];
        if (!(stress > (1 + thresh_diff))) { ox = 0.0f; oy = 0.0f; }
        if (!(stress < (1 - thresh_diff))) { ux = 0.0f; uy = 0.0f; }
        corr[2 * k] = (ox + ux) * 0.8;
        corr[2 * k + 1] = (oy + uy) * 0.8;
    }
    for (int k = 0; k < E; k++) {                                                   /* :670 */
        int a = e->ea[k];
        e->pos[2 * a] -= e->fixed[a] ? 0.0 : corr[2 * k];
        e->pos[2 * a + 1] -= e->fixed[a] ? 0.0 : corr[2 * k + 1];
    }
    for (int k = 0; k < E; k++) {                                                   /* :671 */
        int b = e->eb[k];
        e->pos[2 * b] += e->fixed[b] ? 0.0 : corr[2 * k];
        e->pos[2 * b + 1] += e->fixed[b] ? 0.0 : corr[2 * k + 1];
    }
    for (int j = 0; j < e->n_boxels; j++) {                                         /* :675-710 */
        const int* bx = e->boxels + j * B_STRIDE;
        if (bx[B_TYPE] != CELL_RIGID) continue;
        for (int q = 0; q < 4; q++) {
            int k = bx[B_ET + q]; /* top, right, bot, left (Boxel.cpp:25) */
            int a = e->ea[k], b = e->eb[k];
            double dx = e->pos[2 * a] - e->pos[2 * b];
            double dy = e->pos[2 * a + 1] - e->pos[2 * b + 1];
            double dist = sqrt(dx * dx + dy * dy);
            double over_mag = (dist - e->el0[k] * (1 + thresh_diff_rigid)) / (dist * 2);
            double under_mag = (dist - e->el0[k] * (1 - thresh_diff_rigid)) / (dist * 2);
            double ox = dx * over_mag, oy = dy * over_mag;
            double ux = dx * under_mag, uy = dy * under_mag;
            double stress = dist / e->el0[k];
            if (stress < (1 + thresh_diff_rigid)) { ox = 0; oy = 0; }
            if (stress > (1 - thresh_diff_rigid)) { ux = 0; uy = 0; }
            e->pos[2 * a] -= (ox + ux) * 0.5;
            e->pos[2 * a + 1] -= (oy + uy) * 0.5;
            e->pos[2 * b] += (ox + ux) * 0.5;
            e->pos[2 * b + 1] += (oy + uy) * 0.5;
        }
    }
}

/* PhysicsEngine::update_true_vel, PhysicsEngine.cpp:602-607 */
void mo_update_true_vel(mo_env* e, double dt) {
    for (int i = 0; i < 2 * e->P; i++) {
        e->vel_true[i] = (e->pos[i] - e->pos_last[i]) / dt;
        e->pos_last[i] = e->pos[i];
    }
}

/* ------------------------------------------------------------------------------------------ */
/* collisions */

/* Boxel::get_bounding_box, Boxel.cpp:31-35 -- leaf refit of SimObject::recompute_bbs */
static void refit(mo_env* e) {
    for (int l = 0; l < e->n_leaves; l++) {
        const int* bx = e->boxels + e->leaves[l] * B_STRIDE;
        double minx = e->pos[2 * bx[B_TL]], maxx = minx;
        double miny = e->pos[2 * bx[B_TL] + 1], maxy = miny;
        for (int q = 1; q < 4; q++) {
            double x = e->pos[2 * bx[B_TL + q]], y = e->pos[2 * bx[B_TL + q] + 1];
            if (x < minx) minx = x;
            if (x > maxx) maxx = x;
            if (y < miny) miny = y;
            if (y > maxy) maxy = y;
        }
        double* bb = e->bb + 4 * l;
        bb[0] = minx; bb[1] = miny; bb[2] = maxx; bb[3] = maxy;
    }
}

/* BoundingBox::in, main.h:64-72 (F_ERROR_TOL = 0) */
static int bb_in(const double* a, const double* b) {
    if (a[0] > b[2] || a[2] < b[0]) return 0;
    if (a[1] > b[3] || a[3] < b[1]) return 0;
    return 1;
}

static double cross2(double ax, double ay, double bx, double by) { return ax * by - ay * bx; } /* Boxel.cpp:37-39 */

/* Boxel::point_in_boxel, Boxel.cpp:41-65 */
static int point_in_boxel(const mo_env* e, const int* bx, double px, double py) {
    const double* pos = e->pos;
    double tlx = pos[2 * bx[B_TL]] - px, tly = pos[2 * bx[B_TL] + 1] - py;
    double trx = pos[2 * bx[B_TR]] - px, try_ = pos[2 * bx[B_TR] + 1] - py;
    double blx = pos[2 * bx[B_BL]] - px, bly = pos[2 * bx[B_BL] + 1] - py;
    double brx = pos[2 * bx[B_BR]] - px, bry = pos[2 * bx[B_BR] + 1] - py;
    double a = cross2(tlx, tly, trx, try_);
    double b = cross2(trx, try_, brx, bry);
    double c = cross2(brx, bry, tlx, tly);
    if (c * a >= 0.0 && c * b >= 0.0) return 1;
    double d = cross2(tlx, tly, blx, bly);
    double ee = cross2(blx, bly, brx, bry);
    return (c * d >= 0.0 && c * ee >= 0.0);
}

/* PhysicsEngine::get_intersecting, PhysicsEngine.cpp:1094-1121 */
static int get_intersecting(const mo_env* e, int e1, int e2) {
    const double* pos = e->pos;
    int a1 = e->ea[e1], b1 = e->eb[e1], a2 = e->ea[e2], b2 = e->eb[e2];
    double d1x = pos[2 * b1] - pos[2 * a1], d1y = pos[2 * b1 + 1] - pos[2 * a1 + 1];
    double d2x = pos[2 * b2] - pos[2 * a2], d2y = pos[2 * b2 + 1] - pos[2 * a2 + 1];
    double det = -d1x * d2y + d2x * d1y;
    if (fabs(det) < 0.0001) return 0;
    double x_diff = pos[2 * a2] - pos[2 * a1];
    double y_diff = pos[2 * a2 + 1] - pos[2 * a1 + 1];
    double t = (-d2y * x_diff + d2x * y_diff) / det;
    double s = (-d1y * x_diff + d1x * y_diff) / det;
    return (0.0 <= t && t <= 1 && 0.0 <= s && s <= 1);
}

/* PhysicsEngine::dist_point_edge, PhysicsEngine.cpp:1123-1135 */
static double dist_point_edge(const mo_env* e, int p, int k) {
    const double* pos = e->pos;
    int a = e->ea[k], b = e->eb[k];
    double bx = pos[2 * b] - pos[2 * a], by = pos[2 * b + 1] - pos[2 * a + 1];
    double base_length = sqrt(bx * bx + by * by);
    if (base_length < 0.000001) return 1000.0;
    double sx = pos[2 * p] - pos[2 * a], sy = pos[2 * p + 1] - pos[2 * a + 1];
    double cross = fabs(bx * sy - sx * by);
    return cross / base_length;
}

/* MatrixBase::normalized (Eigen/src/Core/Dot.h:124-134): v / sqrt(v.v) when v.v > 0 */
static void normalized(double* x, double* y) {
    double z = (*x) * (*x) + (*y) * (*y);
    if (z > 0) { double s = sqrt(z); *x = *x / s; *y = *y / s; }
}

/* PhysicsEngine::resolve_point_boxel_collisions, PhysicsEngine.cpp:910-1092 */
static void resolve_point_boxel(mo_env* e, int p, const int* mainb, const int* refb) {
    int main_edges[8];
    int nm = 0;
    if (p == mainb[B_TL]) { main_edges[nm++] = mainb[B_ET]; main_edges[nm++] = mainb[B_EL]; }
    if (p == mainb[B_TR]) { main_edges[nm++] = mainb[B_ET]; main_edges[nm++] = mainb[B_ER]; }
    if (p == mainb[B_BL]) { main_edges[nm++] = mainb[B_EB]; main_edges[nm++] = mainb[B_EL]; }
    if (p == mainb[B_BR]) { main_edges[nm++] = mainb[B_EB]; main_edges[nm++] = mainb[B_ER]; }

    int best = -1;
    double min_dist = 0;
    for (int i = 0; i < 4; i++) {                                                   /* :943-960 */
        int re = refb[B_ET + i];
        if (!e->on_surface[re]) continue;
        int viable = 0;
        for (int j = 0; j < nm; j++)
            if (get_intersecting(e, main_edges[j], re)) { viable = 1; break; }
        if (!viable) continue;
        double d = dist_point_edge(e, p, re);
        if (best < 0 || d < min_dist) { best = re; min_dist = d; }                  /* :1026-1037 */
    }
    if (best < 0) return;
    e->contacts++;

    const double* pos = e->pos;
    const double* vt = e->vel_true;
    int ap = e->ea[best], bp = e->eb[best];
    /* Edge::get_normal, Edge.cpp:16-22 */
    double sx = pos[2 * bp] - pos[2 * ap], sy = pos[2 * bp + 1] - pos[2 * ap + 1];
    double nx = -sy, ny = sx;
    normalized(&nx, &ny);
    double nfx = nx * collision_const_obj * (collision_base_dist_additive + min_dist);   /* :1041 */
    double nfy = ny * collision_const_obj * (collision_base_dist_additive + min_dist);
    double ux = nx, uy = ny;                                                        /* :1044-1045 */
    normalized(&ux, &uy);
    double dp = ux * vt[2 * p] + uy * vt[2 * p + 1];
    double da = ux * vt[2 * ap] + uy * vt[2 * ap + 1];
    double db = ux * vt[2 * bp] + uy * vt[2 * bp + 1];
    double v1x = ux * (-1) * dp * collision_vel_damping, v1y = uy * (-1) * dp * collision_vel_damping; /* :1051-1053 */
    double v2x = ux * (1) * da * collision_vel_damping, v2y = uy * (1) * da * collision_vel_damping;
    double v3x = ux * (1) * db * collision_vel_damping, v3y = uy * (1) * db * collision_vel_damping;
    e->cf[2 * p] += nfx + v1x;            e->cf[2 * p + 1] += nfy + v1y;            /* :1057-1059 */
    e->cf[2 * ap] -= nfx / 2 + v2x / 2;   e->cf[2 * ap + 1] -= nfy / 2 + v2y / 2;
    e->cf[2 * bp] -= nfx / 2 + v3x / 2;   e->cf[2 * bp + 1] -= nfy / 2 + v3y / 2;

    double tx0 = nx, ty0 = ny;                                                      /* :1065-1066 */
    normalized(&tx0, &ty0);
    double tx = -ty0, ty = tx0;
    double proj = tx * vt[2 * p] + ty * vt[2 * p + 1];
    double nfm = sqrt(nfx * nfx + nfy * nfy);
    if (proj < 0) { proj *= -1; tx *= -1; ty *= -1; }
    double th = tanh(proj / (dynamic_friction_const * nfm));
    double ffx = friction_const * tx * (-1) * dynamic_friction_const * nfm * th;    /* :1077 */
    double ffy = friction_const * ty * (-1) * dynamic_friction_const * nfm * th;
    e->cf[2 * p] += ffx;          e->cf[2 * p + 1] += ffy;                          /* :1080-1082 */
    e->cf[2 * ap] -= ffx / 2;     e->cf[2 * ap + 1] -= ffy / 2;
    e->cf[2 * bp] -= ffx / 2;     e->cf[2 * bp + 1] -= ffy / 2;
}

/* PhysicsEngine::resolve_boxel_collisions, PhysicsEngine.cpp:891-908 */
static void resolve_boxel(mo_env* e, const int* mainb, const int* refb) {
    for (int i = 0; i < 4; i++) {
        int p = mainb[B_TL + i];
        int nb = 0;
        for (int j = 0; j < 4; j++)
            if (p == refb[B_TL + j]) nb = 1;
        if (nb) continue;
        if (point_in_boxel(e, refb, e->pos[2 * p], e->pos[2 * p + 1])) resolve_point_boxel(e, p, mainb, refb);
    }
}

/* PhysicsEngine::resolve_collisions, PhysicsEngine.cpp:715-751, with the dual-tree descent
 * (:752-817) replaced by the equivalent all-pairs AABB test in ascending leaf order.
 * Object pairs i <= j; for i == j every ORDERED leaf pair (incl. (a,a)) is visited, so each self
 * pair is resolved twice in each direction -- as the reference's descent does. */
int mo_resolve_collisions(mo_env* e, int* cand_out, int max_cand) {
    refit(e);
    memset(e->cf, 0, sizeof(double) * 2 * e->P);
    e->contacts = 0;
    int n = 0;
    for (int la = 0; la < e->n_leaves; la++) {
        const int* ba = e->boxels + e->leaves[la] * B_STRIDE;
        for (int lb = 0; lb < e->n_leaves; lb++) {
            const int* bb = e->boxels + e->leaves[lb] * B_STRIDE;
            if (ba[B_OBJ] > bb[B_OBJ]) continue;
            if (!bb_in(e->bb + 4 * la, e->bb + 4 * lb)) continue;
            if (cand_out && n < max_cand) { cand_out[2 * n] = e->leaves[la]; cand_out[2 * n + 1] = e->leaves[lb]; }
            n++;
            resolve_boxel(e, ba, bb);                                               /* :791-792 */
            resolve_boxel(e, bb, ba);
        }
    }
    return n;
}

/* PhysicsEngine::is_any_robot_self_colliding + count_object_self_collisions,
 * PhysicsEngine.cpp:819-889.  Uses the boxes refit at the START of the substep (e->bb). */
int mo_is_any_robot_self_colliding(mo_env* e) {
    for (int o = 0; o < e->n_obj; o++) {
        const int* info = e->obj_info + 4 * o;
        if (!info[0]) continue;
        int w = info[3];
        int count = 0, n_surface = 0;
        for (int la = 0; la < e->n_leaves; la++) {
            const int* ba = e->boxels + e->leaves[la] * B_STRIDE;
            if (ba[B_OBJ] != o) continue;
            n_surface++;
            for (int lb = 0; lb < e->n_leaves; lb++) {
                const int* bb = e->boxels + e->leaves[lb] * B_STRIDE;
                if (bb[B_OBJ] != o) continue;
                if (!bb_in(e->bb + 4 * la, e->bb + 4 * lb)) continue;
                if (ba[B_GRID] == bb[B_GRID]) continue;
                int x1 = ba[B_GRID] % w, y1 = ba[B_GRID] / w;
                int x2 = bb[B_GRID] % w, y2 = bb[B_GRID] / w;
                if (abs(x1 - x2) < 1.001 && abs(y1 - y2) < 1.001) continue;
                count++;
            }
        }
        if (count * 1.0 > n_surface) return 1;
    }
    return 0;
}

/* Environment::step, Environment.cpp:166-205 */
int mo_substep(mo_env* e) {
    mo_resolve_collisions(e, 0, 0);
    mo_update_actuators(e);
    mo_step_rk4(e, 0.0001);
    mo_resolve_edge_constraints(e);
    mo_update_true_vel(e, 0.0001);
    return mo_is_any_robot_self_colliding(e);
}

/* the substep loop of Sim::step, Sim.cpp:105-115 (the caller owns sim_time and the snapshot) */
int mo_step(mo_env* e, int n_substeps) {
    for (int i = 0; i < n_substeps; i++)
        if (mo_substep(e)) return 1;
    return 0;
}
