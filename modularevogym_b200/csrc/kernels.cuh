// Shared device code of the batched soft-body engine's step kernels (wkernel.cuh: one warp per environment,
// gkernel.cuh: one CTA per environment) plus the small observation-side kernels.
//
// An environment is owned by one warp / CTA for a whole env-step: its dynamic state is staged from HBM into shared
// memory once, the 30 substeps of Sim::step (reference SimulatorCPP/Sim.cpp:105-115 -> Environment::step,
// Environment.cpp:166-205) run out of shared memory, and the state is written back once.  No tensor cores: the path is a
// gather / sqrt / div / scatter chain in FP64, bounded by latency and FP64 issue, not a contraction.
//
// Arithmetic contract: every expression keeps the reference's operand order and association, and
// this translation unit is compiled with --fmad=false, so each product and sum is rounded
// separately exactly like the reference's x86-64 build (no FMA).  IEEE sqrt / div are the CUDA
// defaults for double.  Spring forces and constraint corrections are summed per point through a
// gather list sorted "a-sides by edge id, then b-sides by edge id", which is the order the
// reference's sequential Eigen indexed-view updates produce (PhysicsEngine.cpp:530-531,670-671),
// so those sums are bit-identical.  Contact forces are summed in candidate-list order (the
// reference uses its bounding-box-tree descent order) -> ulp-level differences, see DESIGN.md.
#pragma once

#include <cuda_runtime.h>

#include "device_layout.h"

namespace mevo {

struct StepParams {
    const uint8_t* design;
    const uint8_t* statics;
    double* state;
    const double* act;          // dense actions, act[env * act_stride + slot]
    int act_stride;
    const double* state0;       // reset snapshot (same layout as state; pos and vel sections used)
    int auto_reset;             // 1: an environment that reports "unstable" is reverted to snapshot 0 in place
    unsigned long long* unstable_events;
    const EnvDesc* envs;
    const int* class_envs;
    int n_class_envs;
    int n_steps;
    int n_substeps;
    int apply_actions;
    int cta_sync;               // warp kernel: CTA barrier once per substep (keeps a CTA's warps in one code region)
    int envs_per_cta;           // warp kernel: environments one CTA takes (<= its warp count)
    uint8_t* unstable;
    int* sim_time;
    int* err;
    unsigned int* dbg_cand;     // optional: candidate codes of the first substep
    int* dbg_cand_n;
    int dbg_cand_cap;
    double* dbg_cf;             // optional: collision_forces after the first substep's collision phase (state-block offsets)
    unsigned int* scratch;      // group kernel: neighbour lists (EnvDesc::scratch_off16)
    // group kernel: an env-step may be issued as several launches of consecutive substep ranges ("chunks", Batch::step_async).
    // chunk_first == 0: this launch continues a step -- an environment whose earlier chunk reported "unstable" is left alone;
    // chunk_last == 0: more chunks follow -- the step is not counted yet
    int chunk_first, chunk_last;
    // warp kernel, item form (wkernel.cuh): work[0] = item counter, work[1 + g] = chunks CTA-group g has completed (zeroed by
    // the host before the launch); the env-step is n_chunks chunks of n_substeps / n_chunks substeps.  nullptr: one item per CTA
    int* work;
    int n_chunks;
};

// PhysicsEngine.h:17-27
#define MEVO_VISCOUS_DRAG 0.1
#define MEVO_GRAVITY 110.0
#define MEVO_COLLISION_CONST_GROUND (2000000.0 * 7)
#define MEVO_COLLISION_CONST_OBJ (300000.0 * 7)
#define MEVO_COLLISION_VEL_DAMPING 140.0
#define MEVO_COLLISION_BASE_DIST 0.005
#define MEVO_DYN_FRICTION 0.2
#define MEVO_FRICTION_CONST (150.0 * 8)
#define MEVO_DT 0.0001
#define MEVO_CELL 0.1

// a / b for a compile-time constant b, bit-identical to the IEEE quotient in three instructions instead of the
// ~25 of a general double division (Markstein's theorem: with rb = RN(1 / b) and a faithful q = RN(a * rb), the
// remainder r = a - q b is exact in one fma and RN(q + r rb) is the correctly rounded a / b; also checked against
// the hardware quotient on 8e8 random operands, tools/check_const_div.c).  Zero, denormal-range, huge and non-finite
// operands take the ordinary division.
__device__ __noinline__ double div_slow(double a, double b) { return a / b; }
__device__ __forceinline__ double div_const(double a, const double b) {
    const double rb = 1.0 / b;                          // folded at compile time
    const double m = fabs(a);
    if (!(m > 1e-290 && m < 1e290)) return div_slow(a, b);
    const double q = a * rb;
    const double r = __fma_rn(-q, b, a);
    return __fma_rn(r, rb, q);
}

__device__ __forceinline__ bool aabb_overlap(const double4& a, const double4& b) {
    // BoundingBox::in, main.h:64-72 (closed intervals, F_ERROR_TOL = 0); (x,y,z,w) = (minx,miny,maxx,maxy)
    // (no short-circuit: one predicate chain instead of four branches; a NaN compares false as before)
    return !(a.x > b.z) & !(a.z < b.x) & !(a.y > b.w) & !(a.w < b.y);
}

__device__ __forceinline__ double4 ldg_d4(const double4* p) {
    const double2 lo = __ldg(reinterpret_cast<const double2*>(p));
    const double2 hi = __ldg(reinterpret_cast<const double2*>(p) + 1);
    return make_double4(lo.x, lo.y, hi.x, hi.y);
}

// necessary condition of Boxel::point_in_boxel: both triangles lie inside the box of the four corners
__device__ __forceinline__ bool point_in_box(const double2 p, const double4& b) {
    return (p.x >= b.x) & (p.x <= b.z) & (p.y >= b.y) & (p.y <= b.w);
}

// A LeafRec held in registers: fields picked with shifts, never with a dynamic array index (which would put the
// record in local memory and every access on an L1 / L2 round trip).
struct LeafRegs {
    unsigned long long pt, ea, eb;     // 4 x u16 each
    unsigned int misc;                 // surf_mask | obj << 8 | gx << 16
    __device__ __forceinline__ int point(int k) const { return (int)((pt >> (16 * k)) & 0xffffu); }
    __device__ __forceinline__ int edge_a(int k) const { return (int)((ea >> (16 * k)) & 0xffffu); }
    __device__ __forceinline__ int edge_b(int k) const { return (int)((eb >> (16 * k)) & 0xffffu); }
    __device__ __forceinline__ bool on_surface(int k) const { return (misc >> k) & 1u; }
};

__device__ __forceinline__ void load_leaf(LeafRegs& out, const LeafRec* dyn_leaf, const LeafRec* st_leaf, unsigned id, bool is_static) {
    uint4 d0, d1;
    if (is_static) {
        const uint4* g = reinterpret_cast<const uint4*>(st_leaf + id);
        d0 = __ldg(g); d1 = __ldg(g + 1);
    } else {
        const uint4* g = reinterpret_cast<const uint4*>(dyn_leaf + id);
        d0 = g[0]; d1 = g[1];
    }
    out.pt = (unsigned long long)d0.x | ((unsigned long long)d0.y << 32);
    out.ea = (unsigned long long)d0.z | ((unsigned long long)d0.w << 32);
    out.eb = (unsigned long long)d1.x | ((unsigned long long)d1.y << 32);
    out.misc = d1.z;
}

__device__ __forceinline__ double cross2(double ax, double ay, double bx, double by) {
    return ax * by - ay * bx;   // Boxel.cpp:37-39
}

// MatrixBase::normalized (Eigen/src/Core/Dot.h:124-134)
__device__ __forceinline__ void normalize2(double& x, double& y) {
    double z = x * x + y * y;
    if (z > 0 && z != 1.0) { double s = sqrt(z); x = x / s; y = y / s; }     // z == 1: s == 1 and x / 1 == x
}

struct PointSrc {
    const double2* dyn_pos;     // shared memory
    const double2* dyn_vt;      // shared memory
    const double2* st_pos;      // global (read-only)
    __device__ __forceinline__ double2 pos(int id, bool is_static) const {
        return is_static ? __ldg(st_pos + id) : dyn_pos[id];
    }
    __device__ __forceinline__ double2 vt(int id, bool is_static) const {
        return is_static ? make_double2(0.0, 0.0) : dyn_vt[id];
    }
};

// PhysicsEngine::get_intersecting, PhysicsEngine.cpp:1094-1121: t = nt / det and s = ns / det both in [0, 1].
// Decided without the two divisions: for det > 0 that is 0 <= n <= det, for det < 0 it is det <= n <= 0.  The rounded
// quotient of the reference can only disagree with this when n / det lies within one rounding error of 0 or 1 (n
// the neighbouring double of det, or a quotient that underflows) -- a measure-zero set of inputs, like the box
// prefilter's (exact shortcuts).
__device__ __forceinline__ bool segments_intersect(double2 a1, double2 b1, double2 a2, double2 b2) {
    double d1x = b1.x - a1.x, d1y = b1.y - a1.y;
    double d2x = b2.x - a2.x, d2y = b2.y - a2.y;
    double det = -d1x * d2y + d2x * d1y;
    if (fabs(det) < 0.0001) return false;
    double x_diff = a2.x - a1.x;
    double y_diff = a2.y - a1.y;
    double nt = -d2y * x_diff + d2x * y_diff;
    double ns = -d1y * x_diff + d1x * y_diff;
    if (det < 0) { nt = -nt; ns = -ns; det = -det; }
    return (0.0 <= nt && nt <= det && 0.0 <= ns && ns <= det);
}

// PhysicsEngine::dist_point_edge, PhysicsEngine.cpp:1123-1135
__device__ __forceinline__ double dist_point_segment(double2 p, double2 a, double2 b) {
    double bx = b.x - a.x, by = b.y - a.y;
    double base_length = sqrt(bx * bx + by * by);
    if (base_length < 0.000001) return 1000.0;
    double sx = p.x - a.x, sy = p.y - a.y;
    double cross = fabs(bx * sy - sx * by);
    return cross / base_length;
}

struct Contact {
    int p, a, b;                 // point ids (namespace given by the flags)
    bool p_static, ab_static;
    double fpx, fpy;             // normal_force + vel_additive1        (added to p)
    double fax, fay;             // normal_force/2 + vel_additive2/2    (subtracted from a)
    double fbx, fby;             // normal_force/2 + vel_additive3/2    (subtracted from b)
    double frx, fry;             // friction_force (added to p; halves subtracted from a and b)
};

// Boxel::point_in_boxel (Boxel.cpp:41-65) followed by
// PhysicsEngine::resolve_point_boxel_collisions (PhysicsEngine.cpp:910-1092).
// corner = index of p in main's TL,TR,BR,BL list.  Returns false when no contact results.
__device__ __forceinline__ bool resolve_point_vs_boxel(const PointSrc& src, const LeafRegs& mainb, bool main_static,
                                                       const LeafRegs& refb, bool ref_static, int corner, Contact& c) {
    // corners shared with refb (PhysicsEngine.cpp:895-901) never reach this function: the host lists only the
    // non-shared corners of point-sharing leaf pairs, and no other pair shares a point
    const int p = mainb.point(corner);
    const double2 pp = src.pos(p, main_static);
    const double2 rtl = src.pos(refb.point(0), ref_static), rtr = src.pos(refb.point(1), ref_static);
    const double2 rbr = src.pos(refb.point(2), ref_static), rbl = src.pos(refb.point(3), ref_static);
    {
        double tlx = rtl.x - pp.x, tly = rtl.y - pp.y;
        double trx = rtr.x - pp.x, try_ = rtr.y - pp.y;
        double blx = rbl.x - pp.x, bly = rbl.y - pp.y;
        double brx = rbr.x - pp.x, bry = rbr.y - pp.y;
        double a = cross2(tlx, tly, trx, try_);
        double b = cross2(trx, try_, brx, bry);
        double cc = cross2(brx, bry, tlx, tly);
        bool inside = (cc * a >= 0.0 && cc * b >= 0.0);
        if (!inside) {
            double d = cross2(tlx, tly, blx, bly);
            double e = cross2(blx, bly, brx, bry);
            inside = (cc * d >= 0.0 && cc * e >= 0.0);
        }
        if (!inside) return false;
    }
    // the two main-boxel sides incident to corner p (PhysicsEngine.cpp:915-938): sides are top,right,bot,left
    const int s0 = (corner == 0 || corner == 1) ? 0 : 2;      // top for TL/TR, bot for BR/BL
    const int s1 = (corner == 0 || corner == 3) ? 3 : 1;      // left for TL/BL, right for TR/BR
    const double2 m0a = src.pos(mainb.edge_a(s0), main_static), m0b = src.pos(mainb.edge_b(s0), main_static);
    const double2 m1a = src.pos(mainb.edge_a(s1), main_static), m1b = src.pos(mainb.edge_b(s1), main_static);

    int best = -1;
    double min_dist = 0.0;
    double2 best_a = make_double2(0, 0), best_b = make_double2(0, 0);
#pragma unroll 1
    for (int s = 0; s < 4; s++) {                                              // :943-960
        if (!refb.on_surface(s)) continue;
        const double2 ra = src.pos(refb.edge_a(s), ref_static), rb = src.pos(refb.edge_b(s), ref_static);
        bool viable = segments_intersect(m0a, m0b, ra, rb);
        if (!viable) viable = segments_intersect(m1a, m1b, ra, rb);
        if (!viable) continue;
        double d = dist_point_segment(pp, ra, rb);
        if (best < 0 || d < min_dist) { best = s; min_dist = d; best_a = ra; best_b = rb; }   // :1026-1037
    }
    if (best < 0) return false;

    c.p = p; c.a = refb.edge_a(best); c.b = refb.edge_b(best);
    c.p_static = main_static; c.ab_static = ref_static;
    // Edge::get_normal, Edge.cpp:16-22
    double sx = best_b.x - best_a.x, sy = best_b.y - best_a.y;
    double nx = -sy, ny = sx;
    normalize2(nx, ny);
    double nfx = nx * MEVO_COLLISION_CONST_OBJ * (MEVO_COLLISION_BASE_DIST + min_dist);     // :1041
    double nfy = ny * MEVO_COLLISION_CONST_OBJ * (MEVO_COLLISION_BASE_DIST + min_dist);
    double ux = nx, uy = ny;                                                               // :1044-1045
    normalize2(ux, uy);
    const double2 vp = src.vt(c.p, main_static), va = src.vt(c.a, ref_static), vb = src.vt(c.b, ref_static);
    double dp = ux * vp.x + uy * vp.y;
    double da = ux * va.x + uy * va.y;
    double db = ux * vb.x + uy * vb.y;
    double v1x = ux * (-1.0) * dp * MEVO_COLLISION_VEL_DAMPING, v1y = uy * (-1.0) * dp * MEVO_COLLISION_VEL_DAMPING;
    double v2x = ux * (1.0) * da * MEVO_COLLISION_VEL_DAMPING, v2y = uy * (1.0) * da * MEVO_COLLISION_VEL_DAMPING;
    double v3x = ux * (1.0) * db * MEVO_COLLISION_VEL_DAMPING, v3y = uy * (1.0) * db * MEVO_COLLISION_VEL_DAMPING;
    c.fpx = nfx + v1x;           c.fpy = nfy + v1y;                                         // :1057-1059
    c.fax = nfx / 2 + v2x / 2;   c.fay = nfy / 2 + v2y / 2;
    c.fbx = nfx / 2 + v3x / 2;   c.fby = nfy / 2 + v3y / 2;
    // :1065-1066 normalises the normal once more for the tangent: same input as (ux, uy), same result
    double tx = -uy, ty = ux;
    double proj = tx * vp.x + ty * vp.y;
    double nfm = sqrt(nfx * nfx + nfy * nfy);
    if (proj < 0) { proj *= -1.0; tx *= -1.0; ty *= -1.0; }
    double th = tanh(proj / (MEVO_DYN_FRICTION * nfm));
    c.frx = MEVO_FRICTION_CONST * tx * (-1.0) * MEVO_DYN_FRICTION * nfm * th;               // :1077
    c.fry = MEVO_FRICTION_CONST * ty * (-1.0) * MEVO_DYN_FRICTION * nfm * th;
    return true;
}

// collision_forces += contact, in the reference's statement order (PhysicsEngine.cpp:1057-1082).
// Forces on static (all-fixed, pruned) points are dropped: they can never move anything.
__device__ __forceinline__ void apply_contact(double2* cf, const Contact& c, int reps) {
    // p, a and b are three different points, so each one's running sum can stay in a register: same additions in
    // the same order per point as the statement-by-statement read-modify-write
    double2 vp = make_double2(0.0, 0.0), va = vp, vb = vp;
    if (!c.p_static) vp = cf[c.p];
    if (!c.ab_static) { va = cf[c.a]; vb = cf[c.b]; }
    const double hx = c.frx / 2, hy = c.fry / 2;
    for (int r = 0; r < reps; r++) {
        vp.x += c.fpx; vp.y += c.fpy;
        va.x -= c.fax; va.y -= c.fay;
        vb.x -= c.fbx; vb.y -= c.fby;
        vp.x += c.frx; vp.y += c.fry;
        va.x -= hx; va.y -= hy;
        vb.x -= hx; vb.y -= hy;
    }
    if (!c.p_static) cf[c.p] = vp;
    if (!c.ab_static) { cf[c.a] = va; cf[c.b] = vb; }
}

// COM of a contiguous dynamic-point range per env (observation side of the path)
__global__ void com_kernel(const double* state, const EnvDesc* envs, const int* ranges /* n_envs x 3: start, count, Pd */,
                           int n_envs, double* out) {
    const int env = blockIdx.x * blockDim.x + threadIdx.x;
    if (env >= n_envs) return;
    const int start = ranges[3 * env], count = ranges[3 * env + 1];
    const double2* pos = reinterpret_cast<const double2*>(state + envs[env].state_off);
    double sx = 0.0, sy = 0.0;
    for (int i = 0; i < count; i++) { const double2 q = pos[start + i]; sx += q.x; sy += q.y; }
    out[2 * env] = count ? sx / count : 0.0;
    out[2 * env + 1] = count ? sy / count : 0.0;
}

// Walker-style observation of one object per env (evogym/envs/walk.py:58-61 = get_vel_com_obs + get_relative_pos_obs,
// evogym/envs/base.py:214-240): row = [mean vel x, mean vel y, x_i - com_x (i < n), y_i - com_y (i < n), 0 ...].
// One warp per environment.
__global__ void obs_kernel(const double* state, const EnvDesc* envs, const int* ranges /* n_envs x 3: start, count, Pd */,
                           int n_envs, double* out, int stride) {
    const int env = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (env >= n_envs) return;
    const int start = ranges[3 * env], count = ranges[3 * env + 1], Pd = ranges[3 * env + 2];
    const double2* pos = reinterpret_cast<const double2*>(state + envs[env].state_off);
    const double2* vel = pos + Pd;
    double sx = 0.0, sy = 0.0, vx = 0.0, vy = 0.0;
    for (int i = lane; i < count; i += 32) {
        const double2 q = pos[start + i], v = vel[start + i];
        sx += q.x; sy += q.y; vx += v.x; vy += v.y;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sx += __shfl_xor_sync(0xffffffffu, sx, o); sy += __shfl_xor_sync(0xffffffffu, sy, o);
        vx += __shfl_xor_sync(0xffffffffu, vx, o); vy += __shfl_xor_sync(0xffffffffu, vy, o);
    }
    double* row = out + (size_t)env * stride;
    const double cx = count ? sx / count : 0.0, cy = count ? sy / count : 0.0;
    if (lane == 0 && stride >= 2) { row[0] = count ? vx / count : 0.0; row[1] = count ? vy / count : 0.0; }
    for (int i = lane; i < count; i += 32) {
        const double2 q = pos[start + i];
        if (2 + i < stride) row[2 + i] = q.x - cx;
        if (2 + count + i < stride) row[2 + count + i] = q.y - cy;
    }
    for (int i = 2 + 2 * count + lane; i < stride; i += 32) row[i] = 0.0;
}

// Environment::object_orientation_at_time (Environment.cpp:317-380) of one object per env, current state against
// snapshot 0: magnitude-weighted mean angle between the centred point vectors, reflected to [0, 2 pi) by the sign of the
// summed cross product (the reference's pi is the literal 3.1415926).  One warp per environment.
template <typename OutT>
__global__ void orientation_kernel(const double* state, const double* state0, const EnvDesc* envs,
                                   const int* ranges /* n_envs x 3: start, count, Pd */, int n_envs, OutT* out, int out_stride) {
    const int env = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (env >= n_envs) return;
    const int start = ranges[3 * env], count = ranges[3 * env + 1];
    const double2* pn = reinterpret_cast<const double2*>(state + envs[env].state_off) + start;
    const double2* po = reinterpret_cast<const double2*>(state0 + envs[env].state_off) + start;
    // The centroids are summed by one lane in the order of the reference's `rowwise().sum()` (Eigen 3.3.90 packetwise redux,
    // Eigen/src/Core/PartialReduxEvaluator.h:107-129: column 0, then columns four at a time as (c1 + c2) + (c3 + c4), then
    // the rest one by one): another order moves a centroid by an ulp, which can carry a normalised dot product across 1.0
    // and with it flip the reference's clamp to 0.99999 -- a 1e-3 change of the angle near zero rotation.
    double cnx = 0.0, cny = 0.0, cox = 0.0, coy = 0.0;
    if (lane == 0 && count > 0) {
        cnx = pn[0].x; cny = pn[0].y; cox = po[0].x; coy = po[0].y;
        int i = 1;
        const int size4 = (count - 1) & ~3;
        for (; i < size4; i += 4) {
            cnx += (pn[i].x + pn[i + 1].x) + (pn[i + 2].x + pn[i + 3].x); cny += (pn[i].y + pn[i + 1].y) + (pn[i + 2].y + pn[i + 3].y);
            cox += (po[i].x + po[i + 1].x) + (po[i + 2].x + po[i + 3].x); coy += (po[i].y + po[i + 1].y) + (po[i + 2].y + po[i + 3].y);
        }
        for (; i < count; i++) { cnx += pn[i].x; cny += pn[i].y; cox += po[i].x; coy += po[i].y; }
    }
    cnx = __shfl_sync(0xffffffffu, cnx, 0); cny = __shfl_sync(0xffffffffu, cny, 0);
    cox = __shfl_sync(0xffffffffu, cox, 0); coy = __shfl_sync(0xffffffffu, coy, 0);
    if (count == 0) { if (lane == 0) out[(size_t)env * out_stride] = (OutT)0; return; }
    cnx /= count; cny /= count; cox /= count; coy /= count;
    double weighted = 0.0, weights = 0.0, cross = 0.0;
    for (int i = lane; i < count; i += 32) {
        double ax = pn[i].x - cnx, ay = pn[i].y - cny, bx = po[i].x - cox, by = po[i].y - coy;
        const double ma = sqrt(ax * ax + ay * ay), mb = sqrt(bx * bx + by * by);
        if (ma > 0.0) { ax /= ma; ay /= ma; }           // Eigen normalized(): left as is when the norm is 0
        if (mb > 0.0) { bx /= mb; by /= mb; }
        double c = ax * bx + ay * by;
        if (ma * mb < 0.0000001) c = 0.99999;
        if (c > 1.0) c = 0.99999;
        if (c < -1.0) c = -0.99999;
        weighted += acos(c) * mb;
        weights += mb;
        cross += ax * by - ay * bx;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        weighted += __shfl_xor_sync(0xffffffffu, weighted, o); weights += __shfl_xor_sync(0xffffffffu, weights, o);
        cross += __shfl_xor_sync(0xffffffffu, cross, o);
    }
    double angle = weighted / weights;
    if (cross > 0.0) angle = 2 * 3.1415926 - angle;
    if (lane == 0) out[(size_t)env * out_stride] = (OutT)angle;
}

}  // namespace mevo
