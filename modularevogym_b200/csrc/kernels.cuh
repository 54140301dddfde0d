// sm_100a kernels of the batched soft-body engine.
//
// One GROUP of G threads (G = 32 .. 512, one warp for 5x5 robots) owns one environment for a whole
// env-step: its dynamic state is staged from HBM into shared memory once, the 30 substeps of
// Sim::step (reference SimulatorCPP/Sim.cpp:105-115 -> Environment::step, Environment.cpp:166-205)
// run out of shared memory, and the state is written back once.  No tensor cores: the path is a
// gather / sqrt / div / scatter chain in FP64, bounded by FP64 issue and latency, not a contraction.
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
    int smem_per_env;
    int n_steps;
    int n_substeps;
    int apply_actions;
    int cta_sync;               // 1: CTA barrier once per substep (multi-environment CTAs)
    int envs_per_cta;           // warp kernel: environments one CTA takes (<= its warp count)
    uint8_t* unstable;
    int* sim_time;
    int* err;
    unsigned int* dbg_cand;     // optional: candidate codes of the first substep
    int* dbg_cand_n;
    int dbg_cand_cap;
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

template <int G, int EPC>
__device__ __forceinline__ void group_sync(int grp) {
    if (EPC == 1) {
        __syncthreads();
    } else if (G == 32) {
        __syncwarp();
    } else {
        asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "n"(G) : "memory");
    }
}

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

// group-wide OR of a per-thread predicate (uniform result); `slot` is a shared-memory word of the group
template <int G, int EPC>
__device__ __forceinline__ bool group_any(int grp, bool pred, volatile int* slot) {
    if (G == 32) return __any_sync(0xffffffffu, pred) != 0;
    if ((threadIdx.x % G) == 0) *slot = 0;
    group_sync<G, EPC>(grp);
    if (__any_sync(0xffffffffu, pred) && (threadIdx.x & 31) == 0) *slot = 1;
    group_sync<G, EPC>(grp);
    const bool r = *slot != 0;
    group_sync<G, EPC>(grp);
    return r;
}

// group-wide OR of a per-thread bit mask (uniform result)
template <int G, int EPC>
__device__ __forceinline__ int group_or(int grp, int bits, volatile int* slot) {
    if (G == 32) return __reduce_or_sync(0xffffffffu, (unsigned)bits);
    if ((threadIdx.x % G) == 0) *slot = 0;
    group_sync<G, EPC>(grp);
    const unsigned w = __reduce_or_sync(0xffffffffu, (unsigned)bits);
    if (w && (threadIdx.x & 31) == 0) atomicOr((int*)slot, (int)w);
    group_sync<G, EPC>(grp);
    const int r = *slot;
    group_sync<G, EPC>(grp);
    return r;
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

#ifdef MEVO_PROF
// development only (-DMEVO_PROF): cycles of the group kernel by phase, taken by the first thread of every group
// [0] K2a boxes, [1] K2b broad phase, [2] K3a filter, [3] K3b resolve + contacts, [4] K4 RK4, [5] K5 limiters, [6] stage in / out
__device__ unsigned long long g_gprof[8];
#define MEVO_GPROF(k)                                                                      \
    do {                                                                                   \
        if (t == 0) {                                                                      \
            const long long _n = clock64();                                                \
            atomicAdd(&g_gprof[gp_phase], (unsigned long long)(_n - gp_t));                \
            gp_t = _n; gp_phase = (k);                                                     \
        }                                                                                  \
    } while (0)
#else
#define MEVO_GPROF(k)
#endif

template <int G, int EPC>
__global__ void __launch_bounds__(G* EPC, ((G >= 64 && G * EPC < 512) ? 512 / (G * EPC) : 1)) step_kernel(const StepParams prm) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int NW = G / 32;
    const int grp = threadIdx.x / G;
    const int t = threadIdx.x % G;
    const int lane = threadIdx.x & 31;
    const int warp = t >> 5;
    int slot = blockIdx.x * EPC + grp;
    // a group beyond the end of the list shadows the last environment (same work, nothing written back) so that
    // every warp of the CTA reaches the per-substep CTA barrier
    const bool ghost = slot >= prm.n_class_envs;
    if (ghost) slot = prm.n_class_envs - 1;
    const int env = prm.class_envs[slot];
    const EnvDesc ed = prm.envs[env];
    const uint8_t* db = prm.design + ed.design_off;
    const DesignHeader H = *reinterpret_cast<const DesignHeader*>(db);
    const int Pd = H.Pd, Ed = H.Ed, n_leaf = H.n_leaf;
    const bool has_static = ed.static_off != ~0ull;
    const uint8_t* sb = has_static ? prm.statics + ed.static_off : nullptr;
    StaticHeader SH;
    if (has_static) SH = *reinterpret_cast<const StaticHeader*>(sb);
    else { SH.Ps = 0; SH.n_leaf = 0; SH.gx0 = SH.gy0 = 0; SH.gw = SH.gh = 0; SH.o_pos = SH.o_leaf = SH.o_aabb = SH.o_grid = 0; }
    const double2* st_pos = reinterpret_cast<const double2*>(sb + SH.o_pos);
    const LeafRec* st_leaf = reinterpret_cast<const LeafRec*>(sb + SH.o_leaf);
    const double4* st_aabb = reinterpret_cast<const double4*>(sb + SH.o_aabb);
    const int16_t* st_grid = reinterpret_cast<const int16_t*>(sb + SH.o_grid);

    unsigned char* sm = smem_raw + (size_t)grp * prm.smem_per_env;
    constexpr bool GT = (G == kGlobalTablesGroup);        // read-only tables stay in the design blob (L1-cached loads)
    const SmemLayout L = smem_layout(Pd, Ed, n_leaf, H.cand_cap, GT);
    double2* pos = reinterpret_cast<double2*>(sm + L.o_pos);
    double2* vel = reinterpret_cast<double2*>(sm + L.o_vel);
    double2* last = reinterpret_cast<double2*>(sm + L.o_last);
    double2* vt = reinterpret_cast<double2*>(sm + L.o_vt);
    double2* cf = reinterpret_cast<double2*>(sm + L.o_cf);
    double2* spos = reinterpret_cast<double2*>(sm + L.o_spos);
    double2* svel = reinterpret_cast<double2*>(sm + L.o_svel);
    double2* accp = reinterpret_cast<double2*>(sm + L.o_accp);
    double2* accv = reinterpret_cast<double2*>(sm + L.o_accv);
    const double* mass = GT ? reinterpret_cast<const double*>(db + H.o_mass) : reinterpret_cast<const double*>(sm + L.o_mass);
    const double* ek = GT ? reinterpret_cast<const double*>(db + H.o_ek) : reinterpret_cast<const double*>(sm + L.o_ek);
    const double* el0 = GT ? reinterpret_cast<const double*>(db + H.o_el0) : reinterpret_cast<const double*>(sm + L.o_el0);
    double* leq = reinterpret_cast<double*>(sm + L.o_leq);
    double* goal = reinterpret_cast<double*>(sm + L.o_goal);
    double2* ef = reinterpret_cast<double2*>(sm + L.o_ef);
    double4* aabb = reinterpret_cast<double4*>(sm + L.o_aabb);
    LeafRec* leaf = reinterpret_cast<LeafRec*>(sm + L.o_leaf);
    const uint16_t* ea = GT ? reinterpret_cast<const uint16_t*>(db + H.o_ea) : reinterpret_cast<const uint16_t*>(sm + L.o_ea);
    const uint16_t* eb = GT ? reinterpret_cast<const uint16_t*>(db + H.o_eb) : reinterpret_cast<const uint16_t*>(sm + L.o_eb);
    const uint16_t* csro = GT ? reinterpret_cast<const uint16_t*>(db + H.o_csro) : reinterpret_cast<const uint16_t*>(sm + L.o_csro);
    const uint16_t* csre = GT ? reinterpret_cast<const uint16_t*>(db + H.o_csre) : reinterpret_cast<const uint16_t*>(sm + L.o_csre);
    const uint8_t* fixedp = GT ? (db + H.o_fixed) : reinterpret_cast<const uint8_t*>(sm + L.o_fixed);
    unsigned int* cand = reinterpret_cast<unsigned int*>(sm + L.o_cand);
    unsigned int* surv = reinterpret_cast<unsigned int*>(sm + L.o_surv);
    const uint8_t* eflag = GT ? (db + H.o_eflag) : reinterpret_cast<const uint8_t*>(sm + L.o_eflag);
    volatile int* misc = reinterpret_cast<volatile int*>(sm + L.o_misc);
    const int seg_cap = H.cand_cap / NW;

    // ---------------------------------------------------------------- stage in (coalesced, once per step)
    double* gstate = prm.state + ed.state_off;
    {
        const double2* g2 = reinterpret_cast<const double2*>(gstate);
        for (int i = t; i < Pd; i += G) {
            pos[i] = g2[i];
            vel[i] = g2[Pd + i];
            last[i] = g2[2 * Pd + i];
            vt[i] = g2[3 * Pd + i];
            if (!GT) {
                const_cast<double*>(mass)[i] = __ldg(reinterpret_cast<const double*>(db + H.o_mass) + i);
                const_cast<uint8_t*>(fixedp)[i] = __ldg(db + H.o_fixed + i);
            }
        }
        const double* ggoal = gstate + 8 * (size_t)Pd;
        const int16_t* eact = reinterpret_cast<const int16_t*>(db + H.o_eact);
        const double* act = prm.act + (size_t)env * prm.act_stride;
        for (int e = t; e < Ed; e += G) {
            const double l0 = __ldg(reinterpret_cast<const double*>(db + H.o_el0) + e);
            if (!GT) {
                const_cast<double*>(ek)[e] = __ldg(reinterpret_cast<const double*>(db + H.o_ek) + e);
                const_cast<double*>(el0)[e] = l0;
                const_cast<uint8_t*>(eflag)[e] = __ldg(db + H.o_eflag + e);
                const_cast<uint16_t*>(ea)[e] = __ldg(reinterpret_cast<const uint16_t*>(db + H.o_ea) + e);
                const_cast<uint16_t*>(eb)[e] = __ldg(reinterpret_cast<const uint16_t*>(db + H.o_eb) + e);
            }
            double g = ggoal[e];
            if (prm.apply_actions) {
                // K1: PhysicsEngine::update_actuator_goals (PhysicsEngine.cpp:38-97) as a per-edge gather:
                // goal = (sum of the <= 2 actuations driving this edge) * (init_len / count)
                int a0 = eact[2 * e], a1 = eact[2 * e + 1];
                if (a0 >= 0) {
                    double sum = 0.0 + act[a0];
                    int cnt = 1;
                    if (a1 >= 0) { sum += act[a1]; cnt = 2; }
                    g = sum * (l0 / cnt);
                }
            }
            goal[e] = g;
        }
        if (!GT) {
            for (int i = t; i < Pd + 1; i += G) const_cast<uint16_t*>(csro)[i] = __ldg(reinterpret_cast<const uint16_t*>(db + H.o_csro) + i);
            for (int i = t; i < 2 * Ed; i += G) const_cast<uint16_t*>(csre)[i] = __ldg(reinterpret_cast<const uint16_t*>(db + H.o_csre) + i);
        }
        const uint4* gl = reinterpret_cast<const uint4*>(db + H.o_leaf);
        uint4* sl = reinterpret_cast<uint4*>(leaf);
        for (int i = t; i < 2 * n_leaf; i += G) sl[i] = __ldg(gl + i);
        if (t < kMiscInts) misc[t] = 0;
    }
    group_sync<G, EPC>(grp);

    const ObjRec* objs = reinterpret_cast<const ObjRec*>(db + H.o_obj);
    const unsigned int* pairs = reinterpret_cast<const unsigned int*>(db + H.o_pairs);
    const unsigned int* items = reinterpret_cast<const unsigned int*>(db + H.o_items);
    const uint16_t* rlev = reinterpret_cast<const uint16_t*>(db + H.o_rlev);
    const uint16_t* rvis = reinterpret_cast<const uint16_t*>(db + H.o_rvis);
    PointSrc src;
    src.dyn_pos = pos; src.dyn_vt = vt; src.st_pos = st_pos;

    int unstable = 0;
    int completed_steps = 0;
    bool overflow = false;

#ifdef MEVO_PROF
    long long gp_t = clock64();
    int gp_phase = 6;
#endif
    for (int step = 0; step < prm.n_steps; step++) {
        for (int sub = 0; sub < prm.n_substeps; sub++) {
            // Warps of one CTA are kept within one substep of each other: they then run the same code at about
            // the same time and share its instruction-cache lines (the loop body is far larger than the I-cache).
            if (EPC > 1 && prm.cta_sync) __syncthreads();
            if (unstable) continue;                       // Sim::step aborted (Sim.cpp:107-110); only keeps the barrier
            // ======================================================= K2a: leaf AABB refit (a-4)
            MEVO_GPROF(0);
            for (int l = t; l < n_leaf; l += G) {
                const LeafRec r = leaf[l];
                double2 q = pos[r.pt[0]];
                double minx = q.x, maxx = q.x, miny = q.y, maxy = q.y;
#pragma unroll
                for (int c = 1; c < 4; c++) {
                    q = pos[r.pt[c]];
                    minx = fmin(minx, q.x); maxx = fmax(maxx, q.x);
                    miny = fmin(miny, q.y); maxy = fmax(maxy, q.y);
                }
                aabb[l] = make_double4(minx, miny, maxx, maxy);
            }
            for (int i = t; i < Pd; i += G) cf[i] = make_double2(0.0, 0.0);
            if (t < 8) misc[16 + t] = 0;
            group_sync<G, EPC>(grp);

            if (EPC > 1 && prm.cta_sync > 1) __syncthreads();
            // ======================================================= K2b: broad phase + compaction (a-4, a-12)
            MEVO_GPROF(1);
            int cnt = 0;     // warp-uniform number of candidates in this warp's segment
            unsigned int* myseg = cand + warp * seg_cap;
            // (i) dynamic-vs-dynamic leaf pairs from the precomputed pair table (each unordered pair once)
            //     four pairs per thread in flight: the table is read from L2
            for (int base = 0; base < H.n_pairs; base += 4 * G) {
                unsigned int code[4];
                bool hit[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int i = base + u * G + t;
                    code[u] = __ldg(pairs + (i < H.n_pairs ? i : base));
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int i = base + u * G + t;
                    const int a = code[u] & kCodeLeafMask, b = (code[u] >> 12) & kCodeLeafMask;
                    hit[u] = (i < H.n_pairs) & aabb_overlap(aabb[a], aabb[b]);
                    if (hit[u] && (code[u] & kCodeCount))          // ordered pairs (a,b) and (b,a) both count (a-12)
                        atomicAdd((int*)&misc[16 + (code[u] >> kCodeObjShift)], 2);
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    if (base + u * G >= H.n_pairs) break;
                    const unsigned m = __ballot_sync(0xffffffffu, hit[u]);
                    if (hit[u]) {
                        const int at = cnt + __popc(m & ((1u << lane) - 1u));
                        if (at < seg_cap) myseg[at] = code[u] & (kCodeLeafMask | (kCodeLeafMask << 12) | kCodeSameObj);
                        else overflow = true;
                    }
                    cnt += __popc(m);
                }
            }
            // (ii) dynamic leaves vs the static uniform grid (cell = one voxel); the window is widened
            //      by one cell each way, then the exact closed-interval test decides
            if (has_static) {
                for (int base = 0; base < n_leaf; base += G) {
                    const int l = base + t;
                    int cx0 = 0, cy0 = 0, wx = 0, n_cells = 0;
                    double4 bb = make_double4(0, 0, 0, 0);
                    if (l < n_leaf) {
                        bb = aabb[l];
                        // static leaf in lattice cell c spans [c, c+1] * 0.1 (to ~1e-15): the closed-interval test
                        // can only succeed for ceil(min*10 - 1) <= c <= floor(max*10); 1e-6 cells of slack, then
                        // the exact test below decides
                        cx0 = max((int)ceil(bb.x * 10.0 - 1.000001) - SH.gx0, 0);
                        int cx1 = min((int)floor(bb.z * 10.0 + 0.000001) - SH.gx0, SH.gw - 1);
                        cy0 = max((int)ceil(bb.y * 10.0 - 1.000001) - SH.gy0, 0);
                        int cy1 = min((int)floor(bb.w * 10.0 + 0.000001) - SH.gy0, SH.gh - 1);
                        wx = cx1 - cx0 + 1;
                        const int wy = cy1 - cy0 + 1;
                        n_cells = (wx > 0 && wy > 0) ? wx * wy : 0;
                    }
                    const int max_cells = __reduce_max_sync(0xffffffffu, n_cells);
                    for (int it = 0; it < max_cells; it++) {
                        bool hit = false;
                        int s = -1;
                        if (it < n_cells) {
                            const int cx = cx0 + it % wx, cy = cy0 + it / wx;
                            s = __ldg(st_grid + cy * SH.gw + cx);
                            if (s >= 0) hit = aabb_overlap(bb, ldg_d4(st_aabb + s));
                        }
                        const unsigned m = __ballot_sync(0xffffffffu, hit);
                        if (hit) {
                            const int at = cnt + __popc(m & ((1u << lane) - 1u));
                            if (at < seg_cap) myseg[at] = (unsigned)l | ((unsigned)s << 12) | kCodeStatic;
                            else overflow = true;
                        }
                        cnt += __popc(m);
                    }
                }
            }
            if (cnt > seg_cap) cnt = seg_cap;
            if (lane == 0) misc[warp] = cnt;
            group_sync<G, EPC>(grp);

            if (prm.dbg_cand && step == 0 && sub == 0 && !ghost) {
                int off = 0;
                for (int w = 0; w < NW; w++) {
                    const int c = misc[w];
                    for (int i = t; i < c; i += G)
                        if (off + i < prm.dbg_cand_cap) prm.dbg_cand[(size_t)env * prm.dbg_cand_cap + off + i] = cand[w * seg_cap + i];
                    off += c;
                }
                if (t == 0) prm.dbg_cand_n[env] = off;
            }

            if (EPC > 1 && prm.cta_sync > 1) __syncthreads();
            // ======================================================= K3a: narrow-phase filter
            MEVO_GPROF(2);
            // A corner can only lie inside a boxel (Boxel::point_in_boxel) if it lies inside the boxel's
            // bounding box, so every (corner, boxel) item is first tested against the box (4 compares) and
            // only the survivors -- real or near penetrations, a handful per substep -- reach the resolver.
            // Survivors are collected in rounds of at most seg_cap per warp (one round in practice; a collapsed
            // robot with hundreds of simultaneous penetrations takes several) so the list can never overflow.
            unsigned int* mysurv = surv + warp * seg_cap;
            int next_ic = warp;          // next 32-wide chunk of the always-items owned by this warp
            int next_cc = 0;             // next 32-wide chunk of this warp's own candidate corners
            __syncwarp();
            bool more;
            do {
            int scnt = 0;
            while (scnt + 32 <= seg_cap) {
                bool hit = false;
                unsigned int code = 0;
                if (next_ic * 32 < H.n_items) {
                    // (i) corners of leaf pairs that share a point (always candidates, never tested in K2b)
                    const int i = next_ic * 32 + lane;
                    next_ic += NW;
                    if (i < H.n_items) {
                        code = __ldg(items + i);
                        const int p = leaf[code & kCodeLeafMask].pt[(code >> kItemCornerShift) & 3];
                        hit = point_in_box(pos[p], aabb[(code >> 12) & kCodeLeafMask]);
                    }
                } else if (next_cc * 32 < cnt * 8) {
                    // (ii) the 8 corners of this warp's own candidates (no point is shared by these pairs)
                    const int i = next_cc * 32 + lane;
                    next_cc++;
                    if (i < cnt * 8) {
                        const unsigned int cc = myseg[i >> 3];
                        const unsigned a = cc & kCodeLeafMask, b = (cc >> 12) & kCodeLeafMask;
                        const bool b_static = (cc & kCodeStatic) != 0;
                        const int dir = (i >> 2) & 1, corner = i & 3;
                        const unsigned same = (cc & kCodeSameObj) ? kItemSameObj : 0u;
                        if (dir == 0) {
                            code = a | (b << 12) | ((unsigned)corner << kItemCornerShift) | (b_static ? kItemRefStatic : 0u) | same;
                            const double4 bb = b_static ? ldg_d4(st_aabb + b) : aabb[b];
                            hit = point_in_box(pos[leaf[a].pt[corner]], bb);
                        } else {
                            code = b | (a << 12) | ((unsigned)corner << kItemCornerShift) | (b_static ? kItemMainStatic : 0u) | same;
                            const double2 pp = b_static ? __ldg(st_pos + __ldg(&st_leaf[b].pt[corner])) : pos[leaf[b].pt[corner]];
                            hit = point_in_box(pp, aabb[a]);
                        }
                    }
                } else {
                    break;
                }
                const unsigned m = __ballot_sync(0xffffffffu, hit);
                if (hit) mysurv[scnt + __popc(m & ((1u << lane) - 1u))] = code;
                scnt += __popc(m);
            }
            more = (next_ic * 32 < H.n_items) || (next_cc * 32 < cnt * 8);
            if (lane == 0) misc[48 + warp] = scnt;
            group_sync<G, EPC>(grp);

            // ======================================================= K3b: resolve + contact response (a-5, a-6)
            MEVO_GPROF(3);
            int n_surv = 0;
#pragma unroll
            for (int w = 0; w < NW; w++) n_surv += misc[48 + w];
            for (int base = 0; base < n_surv; base += G) {
                int i = base + t;
                bool has = false;
                Contact c;
                int reps = 1;
                if (i < n_surv) {
                    int w = 0;
                    while (i >= misc[48 + w]) { i -= misc[48 + w]; w++; }
                    const unsigned code = surv[w * seg_cap + i];
                    const bool m_static = (code & kItemMainStatic) != 0, r_static = (code & kItemRefStatic) != 0;
                    reps = (code & kItemSameObj) ? 2 : 1;
                    LeafRegs lm, lr;
                    load_leaf(lm, leaf, st_leaf, code & kCodeLeafMask, m_static);
                    load_leaf(lr, leaf, st_leaf, (code >> 12) & kCodeLeafMask, r_static);
                    has = resolve_point_vs_boxel(src, lm, m_static, lr, r_static, (code >> kItemCornerShift) & 3, c);
                }
                // deterministic, race-free accumulation: warps in order, lanes in order
                unsigned m = __ballot_sync(0xffffffffu, has);
                if (NW == 1) {
                    while (m) {
                        const int l = __ffs(m) - 1;
                        if (lane == l) apply_contact(cf, c, reps);
                        m &= m - 1;
                        __syncwarp();
                    }
                } else {
                    if (lane == 0) misc[32 + warp] = (int)m;
                    group_sync<G, EPC>(grp);
                    for (int w2 = 0; w2 < NW; w2++) {
                        unsigned m2 = (unsigned)misc[32 + w2];
                        if (m2 == 0) continue;           // uniform across the group
                        if (warp == w2) {
                            while (m2) {
                                const int l = __ffs(m2) - 1;
                                if (lane == l) apply_contact(cf, c, reps);
                                m2 &= m2 - 1;
                                __syncwarp();
                            }
                        }
                        group_sync<G, EPC>(grp);
                    }
                    group_sync<G, EPC>(grp);             // masks are rewritten in the next round
                }
            }
            if (NW == 1) __syncwarp(); else group_sync<G, EPC>(grp);
            } while (group_any<G, EPC>(grp, more, misc + 25));
            // K6a: unstable test uses the boxes of the START of this substep (a-12)
            int sub_unstable = 0;
            for (int o = 0; o < H.n_obj; o++) {
                const ObjRec orc = objs[o];
                if (orc.is_robot && misc[16 + o] > (int)orc.n_surface) sub_unstable = 1;
            }
            group_sync<G, EPC>(grp);

            if (EPC > 1 && prm.cta_sync > 1) __syncthreads();
            // ======================================================= K4: actuators + RK4 (a-7, a-8, a-9)
            MEVO_GPROF(4);
            // (PhysicsEngine::update_actuators is done inside the first stage's edge pass: same positions, same distance)
#pragma unroll 1
            for (int s = 0; s < 4; s++) {
                if (EPC > 1 && prm.cta_sync > 1) __syncthreads();
                const double2* sp = (s == 0) ? pos : spos;
                const double2* sv = (s == 0) ? vel : svel;
                for (int e = t; e < Ed; e += G) {
                    const double2 pa = sp[ea[e]], pb = sp[eb[e]];
                    const double dx = pa.x - pb.x, dy = pa.y - pb.y;
                    const double dist = sqrt(dx * dx + dy * dy);
                    double q;
                    if (s == 0) {
                        const double additive = (goal[e] - dist) * 0.006;
                        q = dist + additive;
                        leq[e] = q;                       // written and read by the same thread (same striding)
                    } else {
                        q = leq[e];
                    }
                    const double mag = (dist - q) * ek[e] / dist;
                    ef[e] = make_double2(dx * mag, dy * mag);
                }
                group_sync<G, EPC>(grp);
                for (int p = t; p < Pd; p += G) {
                    const double2 v = sv[p];
                    const double2 x = sp[p];
                    double Fx = 0.0, Fy = 0.0;
                    const int j1 = csro[p + 1];
                    for (int j = csro[p]; j < j1; j++) {
                        const unsigned ent = csre[j];
                        const double2 f = ef[ent >> 1];
                        if (ent & 1u) { Fx += f.x; Fy += f.y; } else { Fx -= f.x; Fy -= f.y; }
                    }
                    Fx += v.x * (-MEVO_VISCOUS_DRAG);
                    Fy += v.y * (-MEVO_VISCOUS_DRAG);
                    const double m = mass[p];
                    Fy += m * (-MEVO_GRAVITY);
                    const double2 c = cf[p];
                    Fx += c.x; Fy += c.y;
                    {   // analytic ground plane y = 0 (PhysicsEngine.cpp:566-577)
                        const bool hit = x.y < 0;
                        const double normal_force = (hit ? x.y : 0.0) * (-MEVO_COLLISION_CONST_GROUND);
                        Fy += normal_force;
                        const double friction_temp = hit ? tanh(v.x * normal_force * MEVO_DYN_FRICTION) : 0.0;
                        Fx += (-10 * MEVO_DYN_FRICTION) * (normal_force * friction_temp);
                    }
                    const double ax = (m == 1.0) ? Fx : Fx / m;    // x / 1.0 == x exactly
                    const double ay = (m == 1.0) ? Fy : Fy / m;
                    const double kpx = v.x * MEVO_DT, kpy = v.y * MEVO_DT;
                    const double kvx = ax * MEVO_DT, kvy = ay * MEVO_DT;
                    if (s == 0) {
                        accp[p] = make_double2(kpx, kpy);
                        accv[p] = make_double2(kvx, kvy);
                    } else if (s < 3) {
                        double2 q = accp[p]; q.x = q.x + 2 * kpx; q.y = q.y + 2 * kpy; accp[p] = q;
                        q = accv[p]; q.x = q.x + 2 * kvx; q.y = q.y + 2 * kvy; accv[p] = q;
                    }
                    if (s < 2) {
                        const double2 x0 = pos[p], v0 = vel[p];
                        spos[p] = make_double2(x0.x + (kpx * 0.5), x0.y + (kpy * 0.5));
                        svel[p] = make_double2(v0.x + (kvx * 0.5), v0.y + (kvy * 0.5));
                    } else if (s == 2) {
                        const double2 x0 = pos[p], v0 = vel[p];
                        spos[p] = make_double2(x0.x + kpx, x0.y + kpy);
                        svel[p] = make_double2(v0.x + kvx, v0.y + kvy);
                    } else {
                        double2 qp = accp[p], qv = accv[p];
                        double dpx = div_const(qp.x + kpx, 6.0), dpy = div_const(qp.y + kpy, 6.0);
                        double dvx = div_const(qv.x + kvx, 6.0), dvy = div_const(qv.y + kvy, 6.0);
                        if (fixedp[p]) { dpx = 0.0; dpy = 0.0; dvx = 0.0; dvy = 0.0; }
                        double2 x0 = pos[p], v0 = vel[p];
                        x0.x += dpx; x0.y += dpy; v0.x += dvx; v0.y += dvy;
                        pos[p] = x0; vel[p] = v0;
                    }
                }
                group_sync<G, EPC>(grp);
            }

            if (EPC > 1 && prm.cta_sync > 1) __syncthreads();
            // ======================================================= K5: edge constraints + true velocity (a-10, a-11)
            MEVO_GPROF(5);
            // Fast path: the limiters only act beyond +-25 % strain (+-3 % on the sides of rigid voxels).  Squared
            // lengths decide (no sqrt), with a 1e-7 relative guard band so that the decision never depends on the
            // rounding of dist / L0; inside the band every correction of the reference is exactly zero and the
            // positions are left untouched, which is what the reference computes.  Otherwise: the exact passes.
            int viol = 0;
            for (int e = t; e < Ed; e += G) {
                const double2 pa = pos[ea[e]], pb = pos[eb[e]];
                const double dx = pa.x - pb.x, dy = pa.y - pb.y;
                const double d2 = dx * dx + dy * dy;
                const double l0 = el0[e];
                const double l2 = l0 * l0;
                if (!(d2 > l2 * 0.5625001 && d2 < l2 * 1.5624999)) viol |= 1;
                if ((eflag[e] & kEdgeRigidCell) && !(d2 > l2 * 0.9409001 && d2 < l2 * 1.0608999)) viol |= 2;
            }
            viol = group_or<G, EPC>(grp, viol, misc + 24);
            if (viol) {
            if (viol & 1) {
            for (int e = t; e < Ed; e += G) {
                const double2 pa = pos[ea[e]], pb = pos[eb[e]];
                const double dx = pa.x - pb.x, dy = pa.y - pb.y;
                const double d2 = dx * dx + dy * dy;
                const double l0 = el0[e];
                const double l2 = l0 * l0;
                double2 c = make_double2(0.0, 0.0);       // inside the guarded band the correction is exactly zero
                if (!(d2 > l2 * 0.5625001 && d2 < l2 * 1.5624999)) {
                    const double dist = sqrt(d2);
                    const double over_mag = (dist - l0 * (1 + 0.25)) / (dist * 2);
                    const double under_mag = (dist - l0 * (1 - 0.25)) / (dist * 2);
                    double ox = dx * over_mag, oy = dy * over_mag;
                    double ux = dx * under_mag, uy = dy * under_mag;
                    const double stress = dist / l0;
                    if (!(stress > (1 + 0.25))) { ox = 0.0; oy = 0.0; }
                    if (!(stress < (1 - 0.25))) { ux = 0.0; uy = 0.0; }
                    c = make_double2((ox + ux) * 0.8, (oy + uy) * 0.8);
                }
                ef[e] = c;
            }
            group_sync<G, EPC>(grp);
            for (int p = t; p < Pd; p += G) {
                if (fixedp[p]) continue;                  // fixed end points receive "0.0" (PhysicsEngine.cpp:670-671)
                double2 x = pos[p];
                const int j1 = csro[p + 1];
                for (int j = csro[p]; j < j1; j++) {
                    const unsigned ent = csre[j];
                    const double2 c = ef[ent >> 1];
                    if (ent & 1u) { x.x += c.x; x.y += c.y; } else { x.x -= c.x; x.y -= c.y; }
                }
                pos[p] = x;
            }
            group_sync<G, EPC>(grp);
            }   // Jacobi pass
            // ordered rigid-cell pass (PhysicsEngine.cpp:675-710), scheduled in point-disjoint levels that keep the reference's
            // order between visits sharing a point.  Levels are narrow (a dozen visits), so the group's first warp walks them
            // alone with warp-level synchronisation; the other warps wait at one barrier instead of one per level.
            if (viol & 2) {
                if (warp == 0) {
                    for (int lv = 0; lv < H.n_rlev; lv++) {
                        const int i1 = __ldg(rlev + lv + 1);
                        for (int i = __ldg(rlev + lv) + lane; i < i1; i += 32) {
                            const int e = __ldg(rvis + i);
                            const int a = ea[e], b = eb[e];
                            double2 pa = pos[a], pb = pos[b];
                            const double dx = pa.x - pb.x, dy = pa.y - pb.y;
                            const double d2 = dx * dx + dy * dy;
                            const double l0 = el0[e];
                            const double l2 = l0 * l0;
                            if (d2 > l2 * 0.9409001 && d2 < l2 * 1.0608999) continue;     // inside the guarded band: a no-op
                            const double dist = sqrt(d2);
                            const double over_mag = (dist - l0 * (1 + 0.03)) / (dist * 2);
                            const double under_mag = (dist - l0 * (1 - 0.03)) / (dist * 2);
                            double ox = dx * over_mag, oy = dy * over_mag;
                            double ux = dx * under_mag, uy = dy * under_mag;
                            const double stress = dist / l0;
                            if (stress < (1 + 0.03)) { ox = 0.0; oy = 0.0; }
                            if (stress > (1 - 0.03)) { ux = 0.0; uy = 0.0; }
                            const double cx = (ox + ux) * 0.5, cy = (oy + uy) * 0.5;
                            pa.x -= cx; pa.y -= cy; pb.x += cx; pb.y += cy;
                            pos[a] = pa; pos[b] = pb;
                        }
                        __syncwarp();
                    }
                }
                group_sync<G, EPC>(grp);
            }
            }   // exact passes
            for (int p = t; p < Pd; p += G) {             // PhysicsEngine::update_true_vel (:602-607)
                const double2 x = pos[p], l = last[p];
                vt[p] = make_double2(div_const(x.x - l.x, MEVO_DT), div_const(x.y - l.y, MEVO_DT));
                last[p] = x;
            }
            group_sync<G, EPC>(grp);

            if (sub_unstable) unstable = 1;               // Sim::step aborts the remaining substeps (Sim.cpp:107-110)
        }
        if (!unstable) completed_steps++;
    }

    MEVO_GPROF(6);
    // ---------------------------------------------------------------- stage out
    if (!ghost) {
        double2* g2 = reinterpret_cast<double2*>(gstate);
        if (unstable && prm.auto_reset) {
            // Sim::revert(0) in place (Environment.cpp:247-263): pos, vel <- snapshot; pos_last = pos; vel_true = vel
            const double2* s0 = reinterpret_cast<const double2*>(prm.state0 + ed.state_off);
            for (int i = t; i < Pd; i += G) {
                const double2 p0 = s0[i], v0 = s0[Pd + i];
                g2[i] = p0; g2[Pd + i] = v0; g2[2 * Pd + i] = p0; g2[3 * Pd + i] = v0;
            }
        } else {
            for (int i = t; i < Pd; i += G) {
                g2[i] = pos[i];
                g2[Pd + i] = vel[i];
                g2[2 * Pd + i] = last[i];
                g2[3 * Pd + i] = vt[i];
            }
        }
        double* ggoal = gstate + 8 * (size_t)Pd;
        for (int e = t; e < Ed; e += G) ggoal[e] = goal[e];
        if (overflow) atomicOr(prm.err + env, 1);
        if (t == 0) {
            prm.unstable[env] = (uint8_t)unstable;
            if (unstable && prm.auto_reset) prm.sim_time[env] = 0;
            else prm.sim_time[env] += completed_steps;
            if (unstable) atomicAdd(prm.unstable_events, 1ull);
        }
    }
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
