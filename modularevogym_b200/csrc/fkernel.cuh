// CTA-per-group-of-environments step kernel for small worlds (one 5x5-class robot, optionally a small free
// object): the Walker-v0 / Pusher-v0 / Thrower-v0 configurations of BASELINE.json.
//
// The environments of one CTA (8-16 of them, picked by the host so that their totals fill the CTA's capacity) are
// laid out in shared memory as ONE flat mass-spring system: flat point ids, flat edge ids, every index
// pre-translated at stage-in.  The integrator phases (length_eq, 4 RK stages, strain-band test, true velocity)
// then run over the flat arrays with all threads of the CTA -- no per-environment padding (a 36-point robot no
// longer costs two 32-wide passes), equal work per warp, and one code region at a time for the whole CTA, which
// is what the instruction cache wants (the substep body is far larger than it).  The phases that are ordered
// per environment (contacts, Jacobi limiter, rigid-cell visits) are done by one warp per environment.
//
// Same arithmetic, bit for bit, as step_kernel (kernels.cuh) and wstep_kernel (wkernel.cuh); tests run all three.
#pragma once

#include "wkernel.cuh"

namespace mevo {

#ifdef MEVO_PROF
// development only (-DMEVO_PROF): cycle counts per phase, printed when a batch is destroyed
__device__ unsigned long long g_fprof[16];
#define MEVO_PROF_ADD(i, v) atomicAdd(&g_fprof[i], (unsigned long long)(v))
#define MEVO_PROF_MAX(i, v) atomicMax(&g_fprof[i], (unsigned long long)(v))
#else
#define MEVO_PROF_ADD(i, v)
#define MEVO_PROF_MAX(i, v)
#endif

struct CtaDesc {            // host-built: which environments one CTA owns
    int first;              // index into the class's environment list
    int n_env;
};

struct FlatEnv {            // per environment slot, in shared memory (104 bytes)
    int p0, nP, e0, nE, l0, nL, pr0, nPr, it0, nIt, v0, nV;     // ranges in the flat arrays
    int gx0, gy0, gw, gh;                                       // terrain grid (gw = 0: none)
    int env, dead;                                              // global id; 1 = Sim::step aborted (unstable)
    const int16_t* st_grid;
    const double4* st_aabb;
    const double2* st_pos;
    const LeafRec* st_leaf;
    const LeafRec* g_leaf;                                      // full dynamic leaf records (resolver only)
    double* gstate;
};
static_assert(sizeof(FlatEnv) == 120, "FlatEnv layout");

// flat edge word: a (10 bits) | b (10) << 10 | rigid-cell << 20 | type8 << 21, type8 = env slot * 16 + local type
constexpr int kFeB = 10, kFeRigid = 20, kFeType = 21;

template <int NT, int PMAX, int EMAX, int LMAX, int PRMAX, int ITMAX, int VMAX, int NEMAX>
struct FlatSmem {
    static constexpr int NW = NT / 32;
    static constexpr int kCandCap = 64, kSurvCap = 64;
    static constexpr int kZero = EMAX;                          // all-zero entry of ef
    static constexpr int o_pos = 0;
    static constexpr int o_spos = o_pos + 16 * PMAX;
    static constexpr int o_vel = o_spos + 16 * PMAX;
    static constexpr int o_vt = o_vel + 16 * PMAX;
    static constexpr int o_cf = o_vt + 16 * PMAX;
    // uint4 per point: 8 x u16 gather slots, in the reference's summation order (PhysicsEngine.cpp:530-531): the
    // edges whose a end this point is (F -= f) from slot 0 up, the edges whose b end it is (F += f) right-aligned to
    // slot 7.  Slots 0-2 subtract, slots 3-7 add; an a side that lands in slot 3 or 4 carries bit 15 = "negate"
    // (x + (-f) == x - f exactly).  Low 15 bits = flat edge id, empty slot = kZero.
    static constexpr int o_csr = o_cf + 16 * PMAX;
    static constexpr int o_ef = o_csr + 16 * PMAX;              // double2 per edge: the force on its b end (a end: minus)
    static constexpr int o_goal = o_ef + 16 * (EMAX + 1);
    static constexpr int o_etab = o_goal + 8 * EMAX;            // double2 per (env slot, local type)
    static constexpr int o_aabb = o_etab + 16 * 16 * NEMAX;
    static constexpr int o_lpt = o_aabb + 32 * LMAX;            // uint2 per leaf: 4 x u16 flat corner point ids
    static constexpr int o_eidx = o_lpt + 8 * LMAX;             // u32 per edge (flat edge word)
    static constexpr int o_cand = o_eidx + 4 * EMAX;            // per warp
    static constexpr int o_surv = o_cand + 4 * kCandCap * NW;   // per warp
    static constexpr int o_pairs = o_surv + 4 * kSurvCap * NW;  // u16 pair codes (leaf ids local to the environment)
    static constexpr int o_items = o_pairs + 2 * PRMAX;
    static constexpr int o_rord = o_items + 2 * ITMAX;          // u16 flat edge id per rigid-cell visit
    static constexpr int o_fixed = o_rord + 2 * VMAX;           // u8 per point
    static constexpr int o_env = (o_fixed + PMAX + 7) & ~7;     // FlatEnv per slot
    static constexpr int o_cnt = o_env + 120 * NEMAX;           // int[8] self-collision counts per slot
    static constexpr int o_nsurf = o_cnt + 32 * NEMAX;          // int[8] n_surface (robots) or INT_MAX per slot
    static constexpr int o_tot = o_nsurf + 32 * NEMAX;          // int[8]: n_env, totP, totE, totL
    static constexpr int kBytes = (o_tot + 32 + 15) & ~15;
};

// =================================================================================== collisions of one environment
// K2b broad phase, K3 narrow phase + contact response, K6 unstable count, by ONE warp (SURVEY 8a-4,5,6,12); the leaf
// boxes were refit by the whole CTA.  Same producer / consumer rounds as warp_collide (wkernel.cuh), on the flat
// arrays.  Returns 1 when PhysicsEngine::is_any_robot_self_colliding would.
template <class S>
__device__ __noinline__ int flat_collide_env(unsigned char* sm, int es, unsigned int* dbg_row, int* dbg_n, int dbg_cap) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const FlatEnv* E = reinterpret_cast<const FlatEnv*>(sm + S::o_env) + es;
    const int p0 = E->p0, l0 = E->l0, n_leaf = E->nL, n_pairs = E->nPr, n_items = E->nIt;
    const double2* pos = reinterpret_cast<const double2*>(sm + S::o_pos);           // flat
    const double4* aabb = reinterpret_cast<const double4*>(sm + S::o_aabb) + l0;     // this environment's leaves
    const uint2* lpt = reinterpret_cast<const uint2*>(sm + S::o_lpt) + l0;
    unsigned int* cand = reinterpret_cast<unsigned int*>(sm + S::o_cand) + warp * S::kCandCap;
    unsigned int* surv = reinterpret_cast<unsigned int*>(sm + S::o_surv) + warp * S::kSurvCap;
    const uint16_t* pairs = reinterpret_cast<const uint16_t*>(sm + S::o_pairs) + E->pr0;
    const uint16_t* items = reinterpret_cast<const uint16_t*>(sm + S::o_items) + E->it0;
    volatile int* cnts = reinterpret_cast<volatile int*>(sm + S::o_cnt) + 8 * es;
    const int* nsurf = reinterpret_cast<const int*>(sm + S::o_nsurf) + 8 * es;

    auto corner_pos = [&](unsigned leaf, unsigned corner) -> double2 {      // flat point id of a leaf corner
        const uint2 w = lpt[leaf];
        const unsigned v = (corner & 2u) ? w.y : w.x;
        return pos[(v >> ((corner & 1u) * 16)) & 0xffffu];
    };

    // ---- K2a: leaf boxes (a-4), zero collision_forces and the self-collision counts
    {
        double4* bb = reinterpret_cast<double4*>(sm + S::o_aabb) + l0;
        if (lane < n_leaf) {
            const uint2 w = lpt[lane];
            double2 q = pos[w.x & 0xffffu];
            double minx = q.x, maxx = q.x, miny = q.y, maxy = q.y;
            q = pos[w.x >> 16];
            minx = fmin(minx, q.x); maxx = fmax(maxx, q.x); miny = fmin(miny, q.y); maxy = fmax(maxy, q.y);
            q = pos[w.y & 0xffffu];
            minx = fmin(minx, q.x); maxx = fmax(maxx, q.x); miny = fmin(miny, q.y); maxy = fmax(maxy, q.y);
            q = pos[w.y >> 16];
            minx = fmin(minx, q.x); maxx = fmax(maxx, q.x); miny = fmin(miny, q.y); maxy = fmax(maxy, q.y);
            bb[lane] = make_double4(minx, miny, maxx, maxy);
        }
        double2* cfz = reinterpret_cast<double2*>(sm + S::o_cf) + p0;
        for (int i = lane; i < E->nP; i += 32) cfz[i] = make_double2(0.0, 0.0);
        if (lane < 8) cnts[lane] = 0;
        __syncwarp();
    }

    // ---- terrain window of this lane's leaf
    const int gw = E->gw;
    const int16_t* st_grid = E->st_grid;
    const double4* st_aabb = E->st_aabb;
    int cx0 = 0, cy0 = 0, wx = 1, n_cells = 0, max_cells = 0;
    double4 mybb = make_double4(0, 0, 0, 0);
    if (gw > 0) {
        if (lane < n_leaf) {
            mybb = aabb[lane];
            // static leaf in lattice cell c spans [c, c+1] * 0.1: only ceil(min*10 - 1) <= c <= floor(max*10) can
            // pass the closed-interval test; 1e-6 cells of slack, the exact test decides
            cx0 = max((int)ceil(mybb.x * 10.0 - 1.000001) - E->gx0, 0);
            const int cx1 = min((int)floor(mybb.z * 10.0 + 0.000001) - E->gx0, gw - 1);
            cy0 = max((int)ceil(mybb.y * 10.0 - 1.000001) - E->gy0, 0);
            const int cy1 = min((int)floor(mybb.w * 10.0 + 0.000001) - E->gy0, E->gh - 1);
            wx = cx1 - cx0 + 1;
            const int wy = cy1 - cy0 + 1;
            n_cells = (wx > 0 && wy > 0) ? wx * wy : 0;
            if (n_cells == 0) wx = 1;
        }
        max_cells = __reduce_max_sync(0xffffffffu, n_cells);
    }

    int pair_base = 0, cell_it = 0, next_ic = 0, dbg_off = 0;
    bool producing = true;
    while (producing) {
        // ---- produce candidates (32-bit codes of kernels.cuh, leaf ids local to the environment)
        int cnt = 0;
        while (cnt + 32 <= S::kCandCap) {
            bool hit = false;
            unsigned int code = 0;
            if (pair_base < n_pairs) {
                const int i = pair_base + lane;
                pair_base += 32;
                if (i < n_pairs) {
                    const unsigned pc = pairs[i];
                    const unsigned a = pc & 31u, b = (pc >> 5) & 31u;
                    hit = aabb_overlap(aabb[a], aabb[b]);
                    if (hit && (pc & kPair16Count)) atomicAdd((int*)&cnts[pc >> kPair16ObjShift], 2);   // both ordered pairs (a-12)
                    code = a | (b << 12) | ((pc & kPair16Same) ? kCodeSameObj : 0u);
                }
            } else if (cell_it < max_cells) {
                const int it = cell_it++;
                if (it < n_cells) {
                    const int cx = cx0 + it % wx, cy = cy0 + it / wx;
                    const int sl = __ldg(st_grid + cy * gw + cx);
                    if (sl >= 0) {
                        hit = aabb_overlap(mybb, ldg_d4(st_aabb + sl));
                        code = (unsigned)lane | ((unsigned)sl << 12) | kCodeStatic;
                    }
                }
            } else {
                producing = false;
                break;
            }
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (hit) cand[cnt + __popc(m & ((1u << lane) - 1u))] = code;
            cnt += __popc(m);
        }
        __syncwarp();
        if (dbg_row) {
            for (int i = lane; i < cnt; i += 32)
                if (dbg_off + i < dbg_cap) dbg_row[dbg_off + i] = cand[i];
            dbg_off += cnt;
            if (lane == 0) *dbg_n = dbg_off;
        }
        // ---- consume: prefilter into survivors, resolve, apply
        int next_cc = 0;
        bool more = true;
        while (more) {
            int scnt = 0;
            while (scnt + 32 <= S::kSurvCap) {
                bool hit = false;
                unsigned int code = 0;
                if (next_ic * 32 < n_items) {
                    // non-shared corners of leaf pairs that share a point
                    const int i = next_ic * 32 + lane;
                    next_ic++;
                    if (i < n_items) {
                        const unsigned ic = items[i];
                        const unsigned mn = ic & 31u, rf = (ic >> 5) & 31u, corner = (ic >> 10) & 3u;
                        code = mn | (rf << 12) | (corner << kItemCornerShift) | kItemSameObj;
                        hit = point_in_box(corner_pos(mn, corner), aabb[rf]);
                    }
                } else if (next_cc * 32 < cnt * 8) {
                    // the 8 corners of each candidate pair (these pairs share no point)
                    const int i = next_cc * 32 + lane;
                    next_cc++;
                    if (i < cnt * 8) {
                        const unsigned int cc = cand[i >> 3];
                        const unsigned a = cc & kCodeLeafMask, b = (cc >> 12) & kCodeLeafMask;
                        const bool b_static = (cc & kCodeStatic) != 0;
                        const unsigned dir = (i >> 2) & 1, corner = i & 3;
                        const unsigned same = (cc & kCodeSameObj) ? kItemSameObj : 0u;
                        if (dir == 0) {
                            code = a | (b << 12) | (corner << kItemCornerShift) | (b_static ? kItemRefStatic : 0u) | same;
                            const double4 bb = b_static ? ldg_d4(st_aabb + b) : aabb[b];
                            hit = point_in_box(corner_pos(a, corner), bb);
                        } else {
                            code = b | (a << 12) | (corner << kItemCornerShift) | (b_static ? kItemMainStatic : 0u) | same;
                            const double2 pp = b_static ? __ldg(E->st_pos + __ldg(&E->st_leaf[b].pt[corner])) : corner_pos(b, corner);
                            hit = point_in_box(pp, aabb[a]);
                        }
                    }
                } else {
                    more = false;
                    break;
                }
                const unsigned m = __ballot_sync(0xffffffffu, hit);
                if (hit) surv[scnt + __popc(m & ((1u << lane) - 1u))] = code;
                scnt += __popc(m);
            }
            __syncwarp();
            if (scnt > 0) {
                // the resolver works with the environment's own (dynamic-local) point ids
                PointSrc src;
                src.dyn_pos = pos + p0;
                src.dyn_vt = reinterpret_cast<const double2*>(sm + S::o_vt) + p0;
                src.st_pos = E->st_pos;
                double2* cf = reinterpret_cast<double2*>(sm + S::o_cf) + p0;
                for (int base = 0; base < scnt; base += 32) {
                    const int i = base + lane;
                    bool has = false;
                    Contact c;
                    int reps = 1;
                    if (i < scnt) {
                        const unsigned code = surv[i];
                        const bool m_static = (code & kItemMainStatic) != 0, r_static = (code & kItemRefStatic) != 0;
                        reps = (code & kItemSameObj) ? 2 : 1;
                        LeafRec lm, lr;
                        load_leaf(lm, E->g_leaf, E->st_leaf, code & kCodeLeafMask, m_static);
                        load_leaf(lr, E->g_leaf, E->st_leaf, (code >> 12) & kCodeLeafMask, r_static);
                        has = resolve_point_vs_boxel(src, lm, m_static, lr, r_static, (code >> kItemCornerShift) & 3, c);
                    }
                    unsigned m = __ballot_sync(0xffffffffu, has);
                    while (m) {                           // deterministic, race-free: lanes in order
                        const int l = __ffs(m) - 1;
                        if (lane == l) apply_contact(cf, c, reps);
                        m &= m - 1;
                        __syncwarp();
                    }
                }
            }
            __syncwarp();
        }
    }
    // K6a: unstable test uses the boxes of the START of this substep (a-12)
    return __any_sync(0xffffffffu, lane < 8 && cnts[lane] > nsurf[lane]);
}

// F += the 8 gather slots of one point record (FlatSmem::o_csr), in slot order
__device__ __forceinline__ void flat_gather(const double2* ef, const uint4 rec, double& Fx, double& Fy) {
    double2 f[4];
    f[0] = ef[rec.x & 0xffffu]; f[1] = ef[rec.x >> 16]; f[2] = ef[rec.y & 0xffffu];
    f[3] = ef[(rec.y >> 16) & 0x7fffu];
    Fx -= f[0].x; Fy -= f[0].y;
    Fx -= f[1].x; Fy -= f[1].y;
    Fx -= f[2].x; Fy -= f[2].y;
    {
        const int m = (int)(rec.y & 0x80000000u);                        // bit 15 of slot 3
        Fx += __hiloint2double(__double2hiint(f[3].x) ^ m, __double2loint(f[3].x));
        Fy += __hiloint2double(__double2hiint(f[3].y) ^ m, __double2loint(f[3].y));
    }
    f[0] = ef[rec.z & 0x7fffu]; f[1] = ef[rec.z >> 16]; f[2] = ef[rec.w & 0xffffu]; f[3] = ef[rec.w >> 16];
    {
        const int m = (int)((rec.z & 0x8000u) << 16);                    // bit 15 of slot 4
        Fx += __hiloint2double(__double2hiint(f[0].x) ^ m, __double2loint(f[0].x));
        Fy += __hiloint2double(__double2hiint(f[0].y) ^ m, __double2loint(f[0].y));
    }
    Fx += f[1].x; Fy += f[1].y;
    Fx += f[2].x; Fy += f[2].y;
    Fx += f[3].x; Fy += f[3].y;
}

// =================================================================================== limiters + true velocity of one environment
// PhysicsEngine::resolve_edge_constraints + update_true_vel (SURVEY 8a-10,11) on the new positions in spos, by ONE
// warp.  Band test on squared lengths (no sqrt), guarded by 1e-7 so the decision never depends on the rounding of
// dist / L0: inside the band every correction of the reference is exactly zero.  Outside: the Jacobi limiter
// (PhysicsEngine.cpp:613-671), then the rigid-cell visits (:675-710) jumping from one violating visit to the next.
template <class S>
__device__ __noinline__ void flat_constrain_env(unsigned char* sm, int es, bool first_substep) {
    const int lane = threadIdx.x & 31;
    const FlatEnv* E = reinterpret_cast<const FlatEnv*>(sm + S::o_env) + es;
    double2* pos = reinterpret_cast<double2*>(sm + S::o_pos);
    double2* spos = reinterpret_cast<double2*>(sm + S::o_spos);
    double2* ef = reinterpret_cast<double2*>(sm + S::o_ef);
    const double2* etab = reinterpret_cast<const double2*>(sm + S::o_etab);
    const uint32_t* eidx = reinterpret_cast<const uint32_t*>(sm + S::o_eidx);
    const int e0 = E->e0, e1 = e0 + E->nE;
    const int p0 = E->p0, p1 = p0 + E->nP;
    bool jviol = false, rviol = false;
    for (int e = e0 + lane; e < e1; e += 32) {
        const uint32_t ix = eidx[e];
        const double2 pa = spos[ix & 0x3ff], pb = spos[(ix >> kFeB) & 0x3ff];
        const double dx = pa.x - pb.x, dy = pa.y - pb.y;
        const double d2 = dx * dx + dy * dy;
        const double l0 = etab[ix >> kFeType].y;
        const double l2 = l0 * l0;
        if (!(d2 > l2 * 0.5625001 && d2 < l2 * 1.5624999)) jviol = true;
        if (((ix >> kFeRigid) & 1u) && !(d2 > l2 * 0.9409001 && d2 < l2 * 1.0608999)) rviol = true;
    }
    const bool jacobi = __any_sync(0xffffffffu, jviol);
    if (jacobi) {
        for (int e = e0 + lane; e < e1; e += 32) {
            const uint32_t ix = eidx[e];
            const double2 pa = spos[ix & 0x3ff], pb = spos[(ix >> kFeB) & 0x3ff];
            const double dx = pa.x - pb.x, dy = pa.y - pb.y;
            const double dist = sqrt(dx * dx + dy * dy);
            const double l0 = etab[ix >> kFeType].y;
            const double over_mag = (dist - l0 * (1 + 0.25)) / (dist * 2);
            const double under_mag = (dist - l0 * (1 - 0.25)) / (dist * 2);
            double ox = dx * over_mag, oy = dy * over_mag;
            double ux = dx * under_mag, uy = dy * under_mag;
            const double stress = dist / l0;
            if (!(stress > (1 + 0.25))) { ox = 0.0; oy = 0.0; }
            if (!(stress < (1 - 0.25))) { ux = 0.0; uy = 0.0; }
            ef[e] = make_double2((ox + ux) * 0.8, (oy + uy) * 0.8);
        }
        __syncwarp();
        const uint4* csr = reinterpret_cast<const uint4*>(sm + S::o_csr);
        const uint8_t* fixed = reinterpret_cast<const uint8_t*>(sm + S::o_fixed);
        for (int p = p0 + lane; p < p1; p += 32) {
            if (fixed[p]) continue;                   // fixed end points receive "0.0" (PhysicsEngine.cpp:670-671)
            double2 x = spos[p];
            flat_gather(ef, csr[p], x.x, x.y);
            spos[p] = x;
        }
        __syncwarp();
    }
    const int n_vis = E->nV;
    if (n_vis > 0 && (jacobi || __any_sync(0xffffffffu, rviol))) {
        const uint16_t* rord = reinterpret_cast<const uint16_t*>(sm + S::o_rord) + E->v0;
        int start = 0;
#ifdef MEVO_PROF
        const long long tr0 = clock64();
        int iters = 0;
#endif
        while (start < n_vis) {
            int first = -1;
            for (int c = start & ~31; c < n_vis; c += 32) {
                const int i = c + lane;
                bool v = false;
                if (i >= start && i < n_vis) {
                    const uint32_t ix = eidx[rord[i]];
                    const double2 pa = spos[ix & 0x3ff], pb = spos[(ix >> kFeB) & 0x3ff];
                    const double dx = pa.x - pb.x, dy = pa.y - pb.y;
                    const double d2 = dx * dx + dy * dy;
                    const double l0 = etab[ix >> kFeType].y;
                    const double l2 = l0 * l0;
                    v = !(d2 > l2 * 0.9409001 && d2 < l2 * 1.0608999);
                }
                const unsigned m = __ballot_sync(0xffffffffu, v);
                if (m) { first = c + __ffs(m) - 1; break; }
            }
            if (first < 0) break;
            if (lane == (first & 31)) {
                const uint32_t ix = eidx[rord[first]];
                rigid_visit_exact(spos, ix & 0x3ff, (ix >> kFeB) & 0x3ff, etab[ix >> kFeType].y);
            }
            __syncwarp();
            start = first + 1;
#ifdef MEVO_PROF
            iters++;
#endif
        }
#ifdef MEVO_PROF
        if (lane == 0) { MEVO_PROF_ADD(5, iters); MEVO_PROF_MAX(6, iters); MEVO_PROF_ADD(8, clock64() - tr0); MEVO_PROF_ADD(9, 1); if (jacobi) MEVO_PROF_ADD(10, 1); }
#endif
    }
    // PhysicsEngine::update_true_vel (:602-607).  pos_last == pos at the start of every substep except possibly the
    // first of a launch, where it is read from HBM
    double2* vt = reinterpret_cast<double2*>(sm + S::o_vt);
    const double2* g_last = reinterpret_cast<const double2*>(E->gstate) + 2 * E->nP - p0;
    for (int p = p0 + lane; p < p1; p += 32) {
        const double2 x = spos[p];
        const double2 l = first_substep ? g_last[p] : pos[p];
        vt[p] = make_double2((x.x - l.x) / MEVO_DT, (x.y - l.y) / MEVO_DT);
        pos[p] = x;
    }
    __syncwarp();
}

// spring force of one flat edge for stage positions sp (PhysicsEngine.cpp:498-528)
__device__ __forceinline__ double2 flat_spring_force(const double2* sp, uint32_t ix, double k, double leq) {
    const double2 pa = sp[ix & 0x3ff], pb = sp[(ix >> kFeB) & 0x3ff];
    const double dx = pa.x - pb.x, dy = pa.y - pb.y;
    const double dist = sqrt(dx * dx + dy * dy);
    const double mag = (dist - leq) * k / dist;
    return make_double2(dx * mag, dy * mag);
}

// =================================================================================== the kernel
template <int NT, int PMAX, int EMAX, int LMAX, int PRMAX, int ITMAX, int VMAX, int NEMAX, int MINB>
__global__ void __launch_bounds__(NT, MINB) fstep_kernel(const StepParams prm, const CtaDesc* ctas) {
    using S = FlatSmem<NT, PMAX, EMAX, LMAX, PRMAX, ITMAX, VMAX, NEMAX>;
    constexpr int NW = NT / 32;
    extern __shared__ __align__(16) unsigned char sm[];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const CtaDesc cd = ctas[blockIdx.x];
    const int n_env = cd.n_env;
    FlatEnv* envs = reinterpret_cast<FlatEnv*>(sm + S::o_env);
    int* tot = reinterpret_cast<int*>(sm + S::o_tot);
    double2* pos = reinterpret_cast<double2*>(sm + S::o_pos);
    double2* spos = reinterpret_cast<double2*>(sm + S::o_spos);
    double2* vel = reinterpret_cast<double2*>(sm + S::o_vel);
    double2* vt = reinterpret_cast<double2*>(sm + S::o_vt);
    double2* cf = reinterpret_cast<double2*>(sm + S::o_cf);
    uint4* csr = reinterpret_cast<uint4*>(sm + S::o_csr);
    double2* ef = reinterpret_cast<double2*>(sm + S::o_ef);
    double* goal = reinterpret_cast<double*>(sm + S::o_goal);
    double2* etab = reinterpret_cast<double2*>(sm + S::o_etab);
    double4* aabb = reinterpret_cast<double4*>(sm + S::o_aabb);
    uint2* lpt = reinterpret_cast<uint2*>(sm + S::o_lpt);
    uint32_t* eidx = reinterpret_cast<uint32_t*>(sm + S::o_eidx);
    uint8_t* fixed = reinterpret_cast<uint8_t*>(sm + S::o_fixed);

    // ---------------------------------------------------------------- stage in (once per launch)
    if (t < n_env) {
        const int env = prm.class_envs[cd.first + t];
        const EnvDesc ed = prm.envs[env];
        const uint8_t* db = prm.design + ed.design_off;
        const DesignHeader* Hp = reinterpret_cast<const DesignHeader*>(db);
        FlatEnv& E = envs[t];
        E.nP = Hp->Pd; E.nE = Hp->Ed; E.nL = Hp->n_leaf; E.nPr = Hp->n_pairs; E.nIt = Hp->n_items; E.nV = Hp->n_rvis;
        E.env = env; E.dead = 0;
        E.g_leaf = reinterpret_cast<const LeafRec*>(db + Hp->o_leaf);
        E.gstate = prm.state + ed.state_off;
        if (ed.static_off != ~0ull) {
            const uint8_t* sb = prm.statics + ed.static_off;
            const StaticHeader* SHp = reinterpret_cast<const StaticHeader*>(sb);
            E.gx0 = SHp->gx0; E.gy0 = SHp->gy0; E.gw = SHp->gw; E.gh = SHp->gh;
            E.st_grid = reinterpret_cast<const int16_t*>(sb + SHp->o_grid);
            E.st_aabb = reinterpret_cast<const double4*>(sb + SHp->o_aabb);
            E.st_pos = reinterpret_cast<const double2*>(sb + SHp->o_pos);
            E.st_leaf = reinterpret_cast<const LeafRec*>(sb + SHp->o_leaf);
        } else {
            E.gx0 = E.gy0 = E.gw = E.gh = 0;
            E.st_grid = nullptr; E.st_aabb = nullptr; E.st_pos = nullptr; E.st_leaf = nullptr;
        }
    }
    __syncthreads();
    if (t == 0) {
        int p = 0, e = 0, l = 0, pr = 0, it = 0, v = 0;
        for (int s = 0; s < n_env; s++) {
            FlatEnv& E = envs[s];
            E.p0 = p; p += E.nP;
            E.e0 = e; e += E.nE;
            E.l0 = l; l += E.nL;
            E.pr0 = pr; pr += E.nPr;
            E.it0 = it; it += E.nIt;
            E.v0 = v; v += E.nV;
        }
        tot[0] = n_env; tot[1] = p; tot[2] = e; tot[3] = l;
        ef[S::kZero] = make_double2(0.0, 0.0);
    }
    __syncthreads();
    for (int es = warp; es < n_env; es += NW) {
        const FlatEnv& E = envs[es];
        const EnvDesc ed = prm.envs[E.env];
        const uint8_t* db = prm.design + ed.design_off;
        const DesignHeader* Hp = reinterpret_cast<const DesignHeader*>(db);
        const int Pd = E.nP, Ed = E.nE, p0 = E.p0, e0 = E.e0;
        const double2* g2 = reinterpret_cast<const double2*>(E.gstate);
        for (int i = lane; i < Pd; i += 32) {
            pos[p0 + i] = g2[i];
            vel[p0 + i] = g2[Pd + i];
            vt[p0 + i] = g2[3 * Pd + i];
            fixed[p0 + i] = __ldg(db + Hp->o_fixed + i);
            // 8 x (edge << 1 | side), a sides first -> gather record (see FlatSmem::o_csr); host-checked: <= 5 a sides,
            // <= 5 b sides, <= 8 together
            const uint4 rec = __ldg(reinterpret_cast<const uint4*>(db + Hp->o_csr8) + i);
            const unsigned w[4] = {rec.x, rec.y, rec.z, rec.w};
            unsigned long long wa = (unsigned long long)S::kZero * 0x0001000100010001ull, wb = wa;
            int nb = 0;
#pragma unroll
            for (int q = 0; q < 8; q++) {
                const unsigned ent = (w[q >> 1] >> ((q & 1) * 16)) & 0xffffu;
                if (ent != 0xffffu && (ent & 1u)) nb++;
            }
            int ia = 0, ib = 8 - nb;
#pragma unroll
            for (int q = 0; q < 8; q++) {
                const unsigned ent = (w[q >> 1] >> ((q & 1) * 16)) & 0xffffu;
                if (ent != 0xffffu) {
                    const bool bside = (ent & 1u) != 0;
                    const int slot = bside ? ib++ : ia++;
                    unsigned long long id = (ent >> 1) + e0;
                    if (!bside && slot >= 3) id |= 0x8000ull;      // an a side in an adding slot: negate
                    if (slot < 4) wa = (wa & ~(0xffffull << (16 * slot))) | (id << (16 * slot));
                    else wb = (wb & ~(0xffffull << (16 * (slot - 4)))) | (id << (16 * (slot - 4)));
                }
            }
            csr[p0 + i] = make_uint4((unsigned)wa, (unsigned)(wa >> 32), (unsigned)wb, (unsigned)(wb >> 32));
        }
        const double* ggoal = E.gstate + 8 * (size_t)Pd;
        const int16_t* eact = reinterpret_cast<const int16_t*>(db + Hp->o_eact);
        const double* act = prm.act + (size_t)E.env * prm.act_stride;
        for (int e = lane; e < Ed; e += 32) {
            double g = ggoal[e];
            if (prm.apply_actions) {
                // K1: PhysicsEngine::update_actuator_goals (PhysicsEngine.cpp:38-97) as a per-edge gather
                const int a0 = eact[2 * e], a1 = eact[2 * e + 1];
                if (a0 >= 0) {
                    const double l0 = __ldg(reinterpret_cast<const double*>(db + Hp->o_el0) + e);
                    double sum = 0.0 + act[a0];
                    int cnt = 1;
                    if (a1 >= 0) { sum += act[a1]; cnt = 2; }
                    g = sum * (l0 / cnt);
                }
            }
            goal[e0 + e] = g;
            const uint32_t ix = __ldg(reinterpret_cast<const uint32_t*>(db + Hp->o_eidx) + e);   // a | b << 8 | type << 16 | rigid << 20
            eidx[e0 + e] = ((ix & 0xff) + p0) | ((((ix >> 8) & 0xff) + p0) << kFeB) | (((ix >> kEidxRigidBit) & 1u) << kFeRigid) |
                           (((unsigned)es * 16u + ((ix >> kEidxTypeShift) & 15u)) << kFeType);
        }
        if (lane < 16) etab[es * 16 + lane] = lane < Hp->n_types ? __ldg(reinterpret_cast<const double2*>(db + Hp->o_etab) + lane)
                                                                 : make_double2(0.0, 1.0);
        if (lane < E.nL) {
            const LeafRec r = E.g_leaf[lane];
            lpt[E.l0 + lane] = make_uint2((unsigned)(r.pt[0] + p0) | ((unsigned)(r.pt[1] + p0) << 16),
                                          (unsigned)(r.pt[2] + p0) | ((unsigned)(r.pt[3] + p0) << 16));
        }
        uint16_t* pairs = reinterpret_cast<uint16_t*>(sm + S::o_pairs) + E.pr0;
        uint16_t* items = reinterpret_cast<uint16_t*>(sm + S::o_items) + E.it0;
        uint16_t* rord = reinterpret_cast<uint16_t*>(sm + S::o_rord) + E.v0;
        for (int i = lane; i < E.nPr; i += 32) pairs[i] = __ldg(reinterpret_cast<const uint16_t*>(db + Hp->o_pairs16) + i);
        for (int i = lane; i < E.nIt; i += 32) items[i] = __ldg(reinterpret_cast<const uint16_t*>(db + Hp->o_items16) + i);
        for (int i = lane; i < E.nV; i += 32) rord[i] = (uint16_t)(__ldg(db + Hp->o_rord8 + i) + e0);
        if (lane < 8) {
            int ns = 0x7fffffff;                        // never "unstable" unless a robot (PhysicsEngine.cpp:819-834)
            if (lane < Hp->n_obj) {
                const ObjRec orc = reinterpret_cast<const ObjRec*>(db + Hp->o_obj)[lane];
                if (orc.is_robot) ns = orc.n_surface;
            }
            reinterpret_cast<int*>(sm + S::o_nsurf)[8 * es + lane] = ns;
        }
    }
    __syncthreads();
    const int totP = tot[1], totE = tot[2];
    // Thread t owns point t and edges t, t + NT, t + 2 NT for the whole launch: their Runge-Kutta accumulators, stage
    // velocity and length_eq live in registers.
    static_assert(PMAX == NT && EMAX == 3 * NT, "one point and three edges per thread");
    const bool has_p = t < totP;
    const int p = has_p ? t : 0;
    const bool fix_p = fixed[p] != 0;
    // edges of this warp's lanes that exist: warp-uniform count (0..3), per-lane validity below
    const int wtrips = min(max((totE - (t & ~31) + NT - 1) / NT, 0), 3);
    const bool ve0 = t < totE, ve1 = t + NT < totE, ve2 = t + 2 * NT < totE;
    const int e0i = ve0 ? t : 0, e1i = ve1 ? t + NT : e0i, e2i = ve2 ? t + 2 * NT : e0i;
    bool first_substep = true;

    for (int step = 0; step < prm.n_steps; step++) {
        for (int sub = 0; sub < prm.n_substeps; sub++) {
#ifdef MEVO_PROF
            const long long tp0 = clock64();
#endif
            // ======================================================= K2/K3/K6: collisions, one warp per environment
            for (int es = warp; es < n_env; es += NW) {
                FlatEnv& E = envs[es];
                if (E.dead) continue;                     // Sim::step aborted (Sim.cpp:107-110)
                const bool dump = prm.dbg_cand != nullptr && step == 0 && sub == 0;
                const int u = flat_collide_env<S>(sm, es, dump ? prm.dbg_cand + (size_t)E.env * prm.dbg_cand_cap : nullptr,
                                                  dump ? prm.dbg_cand_n + E.env : nullptr, prm.dbg_cand_cap);
                if (u && lane == 0) E.dead = -(step + 1);  // dying: finishes this substep, then its state is frozen
            }
#ifdef MEVO_PROF
            const long long tp1 = clock64();
#endif
            __syncthreads();
#ifdef MEVO_PROF
            const long long tp2 = clock64();
#endif

            // ======================================================= K4: length_eq + RK4 over the flat system (a-7,8,9)
            // (registers of this block are not live across the per-environment calls)
            const uint32_t ix0 = eidx[e0i], ix1 = eidx[e1i], ix2 = eidx[e2i];
            double2 svel_r = make_double2(0.0, 0.0), accp_r = svel_r, accv_r = svel_r;
            double leq0 = 0.0, leq1 = 0.0, leq2 = 0.0;
            {
                // PhysicsEngine::update_actuators: length_eq of this thread's edges
                auto length_eq = [&](uint32_t ix, int e) {
                    const double2 a = pos[ix & 0x3ff], b = pos[(ix >> kFeB) & 0x3ff];
                    const double dx = a.x - b.x, dy = a.y - b.y;
                    const double d = sqrt(dx * dx + dy * dy);
                    const double add = (goal[e] - d) * 0.006;
                    return d + add;
                };
                if (wtrips > 0) leq0 = length_eq(ix0, e0i);
                if (wtrips > 1) leq1 = length_eq(ix1, e1i);
                if (wtrips > 2) leq2 = length_eq(ix2, e2i);
            }
#pragma unroll 1
            for (int s = 0; s < 4; s++) {
                const double2* sp = (s == 0) ? pos : spos;
                // spring pass: ef[e] = force on the b end of edge e; edges 0 and 1 of a thread in flight together
                if (wtrips > 1) {
                    const double2 f0 = flat_spring_force(sp, ix0, etab[ix0 >> kFeType].x, leq0);
                    const double2 f1 = flat_spring_force(sp, ix1, etab[ix1 >> kFeType].x, leq1);
                    ef[e0i] = f0;
                    if (ve1) ef[e1i] = f1;
                } else if (wtrips > 0 && ve0) {
                    ef[e0i] = flat_spring_force(sp, ix0, etab[ix0 >> kFeType].x, leq0);
                }
                if (wtrips > 2 && ve2) ef[e2i] = flat_spring_force(sp, ix2, etab[ix2 >> kFeType].x, leq2);
                __syncthreads();
                if (has_p) {
                    const double2 v = (s == 0) ? vel[p] : svel_r;
                    double Fx = 0.0, Fy = 0.0;
                    {
                        // F -= f over the edges whose a end this is, in edge order, then F += f over those whose b end
                        // it is (PhysicsEngine.cpp:530-531); empty slots read zero
                        const uint4 rec = csr[p];
                        flat_gather(ef, rec, Fx, Fy);
                    }
                    Fx += v.x * (-MEVO_VISCOUS_DRAG);
                    Fy += v.y * (-MEVO_VISCOUS_DRAG);
                    Fy += 1.0 * (-MEVO_GRAVITY);            // every point mass is 1.0 (host-checked), F / m == F
                    const double2 c = cf[p];
                    Fx += c.x; Fy += c.y;
                    const double y = sp[p].y;
                    if (y < 0) {                            // analytic ground plane y = 0 (PhysicsEngine.cpp:566-577)
                        const double normal_force = y * (-MEVO_COLLISION_CONST_GROUND);
                        Fy += normal_force;
                        const double friction_temp = ground_friction_tanh(v.x * normal_force * MEVO_DYN_FRICTION);
                        Fx += (-10 * MEVO_DYN_FRICTION) * (normal_force * friction_temp);
                    }
                    const double kpx = v.x * MEVO_DT, kpy = v.y * MEVO_DT;
                    const double kvx = Fx * MEVO_DT, kvy = Fy * MEVO_DT;
                    const double2 x0 = pos[p], v0 = vel[p];
                    if (s == 0) {
                        accp_r = make_double2(kpx, kpy);
                        accv_r = make_double2(kvx, kvy);
                        spos[p] = make_double2(x0.x + (kpx * 0.5), x0.y + (kpy * 0.5));
                        svel_r = make_double2(v0.x + (kvx * 0.5), v0.y + (kvy * 0.5));
                    } else if (s < 3) {
                        accp_r.x = accp_r.x + 2 * kpx; accp_r.y = accp_r.y + 2 * kpy;
                        accv_r.x = accv_r.x + 2 * kvx; accv_r.y = accv_r.y + 2 * kvy;
                        const double h = (s == 1) ? 0.5 : 1.0;      // x * 1.0 == x exactly
                        spos[p] = make_double2(x0.x + (kpx * h), x0.y + (kpy * h));
                        svel_r = make_double2(v0.x + (kvx * h), v0.y + (kvy * h));
                    } else {
                        double dpx = (accp_r.x + kpx) / 6, dpy = (accp_r.y + kpy) / 6;
                        double dvx = (accv_r.x + kvx) / 6, dvy = (accv_r.y + kvy) / 6;
                        if (fix_p) { dpx = 0.0; dpy = 0.0; dvx = 0.0; dvy = 0.0; }
                        // the new position goes to spos: pos keeps the substep's start position (= pos_last)
                        spos[p] = make_double2(x0.x + dpx, x0.y + dpy);
                        vel[p] = make_double2(v0.x + dvx, v0.y + dvy);
                    }
                }
                __syncthreads();
            }

#ifdef MEVO_PROF
            const long long tp3 = clock64();
#endif
            // ======================================================= K5: limiters + true velocity, one warp per environment
            // (the same warp that does the environment's collisions of the next substep: no CTA barrier in between)
            for (int es = warp; es < n_env; es += NW) {
                if (envs[es].dead > 0) continue;
                flat_constrain_env<S>(sm, es, first_substep);
            }
            first_substep = false;
#ifdef MEVO_PROF
            if (lane == 0) {
                const long long tp4 = clock64();
                MEVO_PROF_ADD(0, tp1 - tp0); MEVO_PROF_ADD(1, tp2 - tp1); MEVO_PROF_ADD(2, tp3 - tp2); MEVO_PROF_ADD(3, tp4 - tp3);
                MEVO_PROF_ADD(4, 1); MEVO_PROF_MAX(11, tp1 - tp0); MEVO_PROF_MAX(12, tp4 - tp3);
            }
#endif
            // an environment that reported "unstable" in this substep is written back now and frozen
            for (int es = warp; es < n_env; es += NW) {
                FlatEnv& E = envs[es];
                if (E.dead >= 0) continue;
                const int Pd = E.nP, p0 = E.p0;
                double2* o2 = reinterpret_cast<double2*>(E.gstate);
                if (prm.auto_reset) {
                    // Sim::revert(0) in place (Environment.cpp:247-263)
                    const double2* s0 = reinterpret_cast<const double2*>(prm.state0 + (E.gstate - prm.state));
                    for (int i = lane; i < Pd; i += 32) {
                        const double2 q0 = s0[i], v0 = s0[Pd + i];
                        o2[i] = q0; o2[Pd + i] = v0; o2[2 * Pd + i] = q0; o2[3 * Pd + i] = v0;
                    }
                } else {
                    for (int i = lane; i < Pd; i += 32) {
                        const double2 x = pos[p0 + i];
                        o2[i] = x; o2[Pd + i] = vel[p0 + i]; o2[2 * Pd + i] = x; o2[3 * Pd + i] = vt[p0 + i];
                    }
                }
                if (lane == 0) {
                    const int steps_done = -E.dead - 1;
                    prm.unstable[E.env] = 1;
                    if (prm.auto_reset) prm.sim_time[E.env] = 0;
                    else prm.sim_time[E.env] += steps_done;
                    atomicAdd(prm.unstable_events, 1ull);
                    E.dead = 1;
                }
                __syncwarp();
            }
        }
    }

    // ---------------------------------------------------------------- stage out
    __syncthreads();
    for (int es = warp; es < n_env; es += NW) {
        const FlatEnv& E = envs[es];
        const int Pd = E.nP, Ed = E.nE, p0 = E.p0;
        double* ggoal = E.gstate + 8 * (size_t)Pd;
        for (int e = lane; e < Ed; e += 32) ggoal[e] = goal[E.e0 + e];
        if (E.dead) continue;
        double2* o2 = reinterpret_cast<double2*>(E.gstate);
        if (!first_substep) {
            for (int i = lane; i < Pd; i += 32) {
                const double2 x = pos[p0 + i];
                o2[i] = x; o2[Pd + i] = vel[p0 + i]; o2[2 * Pd + i] = x; o2[3 * Pd + i] = vt[p0 + i];
            }
        }
        if (lane == 0) {
            prm.unstable[E.env] = 0;
            prm.sim_time[E.env] += prm.n_steps;
        }
    }
}

}  // namespace mevo
