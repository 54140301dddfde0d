// Warp-per-environment step kernel for small robots (<= 64 dynamic points, <= ECAP dynamic edges,
// <= 32 surface leaves): the Walker-v0 / Pusher-v0 / Thrower-v0 5x5 configurations.
//
// Same arithmetic as step_kernel (kernels.cuh) -- the two are interchangeable bit for bit -- but laid out
// for the instruction front end, which is what bounds this path on B200 (ncu: `no_inst` and fixed-latency
// `wait` stalls dominate; FP64 pipe < 15 %, DRAM ~ 0):
//   * compile-time shared-memory layout (template capacities) -> no pointer registers, one base per warp,
//   * the spring pass handles two edges per lane at a time (two independent sqrt / div chains in flight),
//   * per-point gather lists are one 16-byte record (8 x u16) instead of a CSR walk,
//   * rarely taken paths (contact resolver, tanh of the analytic ground, exact strain limiters) are
//     out-of-line so the hot loop body stays small,
//   * every warp of a CTA owns one environment; a CTA barrier per substep keeps the warps in the same code
//     region so they share instruction-cache lines.
#pragma once

#include "kernels.cuh"

namespace mevo {

template <int PCAP, int ECAP, int LCAP>
struct WarpSmem {
    static constexpr int kCandCap = 64;            // overlapping pairs per round (two 32-wide chunks)
    static constexpr int kSurvCap = 64;            // survivors per round (two 32-wide chunks)
    static constexpr int o_pos = 0;
    static constexpr int o_spos = o_pos + 16 * PCAP;
    static constexpr int o_vel = o_spos + 16 * PCAP;
    static constexpr int o_svel = o_vel + 16 * PCAP;
    static constexpr int o_vt = o_svel + 16 * PCAP;
    static constexpr int o_cf = o_vt + 16 * PCAP;
    static constexpr int o_accp = o_cf + 16 * PCAP;
    static constexpr int o_accv = o_accp + 16 * PCAP;
    static constexpr int o_csr = o_accv + 16 * PCAP;          // uint4 per point: 8 x (edge << 1 | side)
    static constexpr int o_ef = o_csr + 16 * PCAP;            // double2 per edge, [ECAP] = (0, 0)
    static constexpr int o_ekl = o_ef + 16 * (ECAP + 1);      // double2 per edge: (spring constant, length_eq)
    static constexpr int o_l0g = o_ekl + 16 * ECAP;           // double2 per edge: (init_length_eq, length_eq_goal)
    static constexpr int o_aabb = o_l0g + 16 * ECAP;
    static constexpr int o_leaf = o_aabb + 32 * LCAP;
    static constexpr int o_eidx = o_leaf + 32 * LCAP;         // u32 per edge: a | b << 8 | flags << 16
    static constexpr int o_cand = o_eidx + 4 * ECAP;
    static constexpr int o_surv = o_cand + 4 * kCandCap;
    static constexpr int o_misc = o_surv + 4 * kSurvCap;      // [0..7] self-collision counts per object
    static constexpr int kBytes = (o_misc + 4 * 8 + 15) & ~15;
};

// tanh of the analytic ground plane's friction term (PhysicsEngine.cpp:573-577); out of line: a robot standing
// on voxel terrain never gets below y = 0
__device__ __noinline__ double ground_friction_tanh(double x) { return tanh(x); }

// exact strain limiters of PhysicsEngine::resolve_edge_constraints (PhysicsEngine.cpp:608-713) for one warp;
// out of line: runs only in substeps where some edge left the +-25 % / +-3 % band
__device__ __noinline__ void warp_exact_constraints(double2* pos, double2* ef, const double2* l0g, const uint32_t* eidx,
                                                    const uint4* csr, int Pd, int Ed, unsigned fixmask, int ecap,
                                                    const uint16_t* rlev, const uint16_t* rvis, int n_rlev) {
    const int lane = threadIdx.x & 31;
    for (int e = lane; e < Ed; e += 32) {
        const uint32_t ix = eidx[e];
        const double2 pa = pos[ix & 0xff], pb = pos[(ix >> 8) & 0xff];
        const double dx = pa.x - pb.x, dy = pa.y - pb.y;
        const double dist = sqrt(dx * dx + dy * dy);
        const double l0 = l0g[e].x;
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
    for (int p = lane, k = 0; p < Pd; p += 32, k++) {
        if ((fixmask >> k) & 1u) continue;            // fixed end points receive "0.0" (PhysicsEngine.cpp:670-671)
        double2 x = pos[p];
        const uint4 rec = csr[p];
        const unsigned w[4] = {rec.x, rec.y, rec.z, rec.w};
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const unsigned ent = (w[j >> 1] >> ((j & 1) * 16)) & 0xffffu;
            if (ent == 0xffffu) continue;
            const double2 c = ef[ent >> 1];
            if (ent & 1u) { x.x += c.x; x.y += c.y; } else { x.x -= c.x; x.y -= c.y; }
        }
        pos[p] = x;
    }
    __syncwarp();
    // ordered rigid-cell pass, scheduled in point-disjoint levels (PhysicsEngine.cpp:675-710)
    for (int lv = 0; lv < n_rlev; lv++) {
        const int i1 = __ldg(rlev + lv + 1);
        for (int i = __ldg(rlev + lv) + lane; i < i1; i += 32) {
            const int e = __ldg(rvis + i);
            const uint32_t ix = eidx[e];
            const int a = ix & 0xff, b = (ix >> 8) & 0xff;
            double2 pa = pos[a], pb = pos[b];
            const double dx = pa.x - pb.x, dy = pa.y - pb.y;
            const double d2 = dx * dx + dy * dy;
            const double l0 = l0g[e].x;
            // cheap exact skip: inside the +-3 % band (guarded) every term of the reference is zero
            const double l2 = l0 * l0;
            if (d2 > l2 * 0.9409001 && d2 < l2 * 1.0608999) continue;
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

// spring force of one edge for stage positions sp (PhysicsEngine.cpp:498-528)
__device__ __forceinline__ double2 spring_force(const double2* sp, uint32_t ix, double2 kl) {
    const double2 pa = sp[ix & 0xff], pb = sp[(ix >> 8) & 0xff];
    const double dx = pa.x - pb.x, dy = pa.y - pb.y;
    const double dist = sqrt(dx * dx + dy * dy);
    const double mag = (dist - kl.y) * kl.x / dist;
    return make_double2(dx * mag, dy * mag);
}

template <int PCAP, int ECAP, int LCAP, int EPC, int MINB>
__global__ void __launch_bounds__(32 * EPC, MINB) wstep_kernel(const StepParams prm) {
    using S = WarpSmem<PCAP, ECAP, LCAP>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int grp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    int slot = blockIdx.x * EPC + grp;
    const bool ghost = slot >= prm.n_class_envs;      // shadows the last environment, writes nothing back
    if (ghost) slot = prm.n_class_envs - 1;
    const int env = prm.class_envs[slot];
    const EnvDesc ed = prm.envs[env];
    const uint8_t* db = prm.design + ed.design_off;
    const DesignHeader* Hp = reinterpret_cast<const DesignHeader*>(db);
    const int Pd = Hp->Pd, Ed = Hp->Ed, n_leaf = Hp->n_leaf;
    const int n_pairs = Hp->n_pairs, n_items = Hp->n_items, n_obj = Hp->n_obj;
    const bool has_static = ed.static_off != ~0ull;
    const uint8_t* sb = prm.statics + (has_static ? ed.static_off : 0);
    const StaticHeader* SHp = reinterpret_cast<const StaticHeader*>(sb);
    const double2* st_pos = reinterpret_cast<const double2*>(sb + (has_static ? SHp->o_pos : 0));
    const LeafRec* st_leaf = reinterpret_cast<const LeafRec*>(sb + (has_static ? SHp->o_leaf : 0));
    const double4* st_aabb = reinterpret_cast<const double4*>(sb + (has_static ? SHp->o_aabb : 0));

    unsigned char* sm = smem_raw + grp * S::kBytes;
    double2* pos = reinterpret_cast<double2*>(sm + S::o_pos);
    double2* spos = reinterpret_cast<double2*>(sm + S::o_spos);
    double2* vel = reinterpret_cast<double2*>(sm + S::o_vel);
    double2* svel = reinterpret_cast<double2*>(sm + S::o_svel);
    double2* vt = reinterpret_cast<double2*>(sm + S::o_vt);
    double2* cf = reinterpret_cast<double2*>(sm + S::o_cf);
    double2* accp = reinterpret_cast<double2*>(sm + S::o_accp);
    double2* accv = reinterpret_cast<double2*>(sm + S::o_accv);
    uint4* csr = reinterpret_cast<uint4*>(sm + S::o_csr);
    double2* ef = reinterpret_cast<double2*>(sm + S::o_ef);
    double2* ekl = reinterpret_cast<double2*>(sm + S::o_ekl);
    double2* l0g = reinterpret_cast<double2*>(sm + S::o_l0g);
    double4* aabb = reinterpret_cast<double4*>(sm + S::o_aabb);
    LeafRec* leaf = reinterpret_cast<LeafRec*>(sm + S::o_leaf);
    uint32_t* eidx = reinterpret_cast<uint32_t*>(sm + S::o_eidx);
    unsigned int* cand = reinterpret_cast<unsigned int*>(sm + S::o_cand);
    unsigned int* surv = reinterpret_cast<unsigned int*>(sm + S::o_surv);
    volatile int* misc = reinterpret_cast<volatile int*>(sm + S::o_misc);

    // ---------------------------------------------------------------- stage in (once per launch)
    double* gstate = prm.state + ed.state_off;
    const double2* g2 = reinterpret_cast<const double2*>(gstate);
    unsigned fixmask = 0;                               // bit k: point lane + 32 k is fixed
    for (int i = lane, k = 0; i < Pd; i += 32, k++) {
        pos[i] = g2[i];
        vel[i] = g2[Pd + i];
        vt[i] = g2[3 * Pd + i];
        csr[i] = __ldg(reinterpret_cast<const uint4*>(db + Hp->o_csr8) + i);
        if (__ldg(db + Hp->o_fixed + i)) fixmask |= 1u << k;
    }
    {
        const double* ggoal = gstate + 8 * (size_t)Pd;
        const int16_t* eact = reinterpret_cast<const int16_t*>(db + Hp->o_eact);
        const double* act = prm.act + (size_t)env * prm.act_stride;
        for (int e = lane; e < Ed; e += 32) {
            const double l0 = __ldg(reinterpret_cast<const double*>(db + Hp->o_el0) + e);
            double g = ggoal[e];
            if (prm.apply_actions) {
                // K1: PhysicsEngine::update_actuator_goals (PhysicsEngine.cpp:38-97) as a per-edge gather
                const int a0 = eact[2 * e], a1 = eact[2 * e + 1];
                if (a0 >= 0) {
                    double sum = 0.0 + act[a0];
                    int cnt = 1;
                    if (a1 >= 0) { sum += act[a1]; cnt = 2; }
                    g = sum * (l0 / cnt);
                }
            }
            l0g[e] = make_double2(l0, g);
            ekl[e] = make_double2(__ldg(reinterpret_cast<const double*>(db + Hp->o_ek) + e), 0.0);
            eidx[e] = __ldg(reinterpret_cast<const uint32_t*>(db + Hp->o_eidx) + e);
        }
        if (lane == 0) ef[ECAP] = make_double2(0.0, 0.0);
        const uint4* gl = reinterpret_cast<const uint4*>(db + Hp->o_leaf);
        uint4* sl = reinterpret_cast<uint4*>(leaf);
        for (int i = lane; i < 2 * n_leaf; i += 32) sl[i] = __ldg(gl + i);
    }
    __syncwarp();

    const unsigned int* pairs = reinterpret_cast<const unsigned int*>(db + Hp->o_pairs);
    const unsigned int* items = reinterpret_cast<const unsigned int*>(db + Hp->o_items);
    PointSrc src;
    src.dyn_pos = pos; src.dyn_vt = vt; src.st_pos = st_pos;

    int unstable = 0;
    int completed_steps = 0;
    bool first_substep = true;

    for (int step = 0; step < prm.n_steps; step++) {
        for (int sub = 0; sub < prm.n_substeps; sub++) {
            if (EPC > 1) __syncthreads();                 // keeps the CTA's warps in the same code region (I-cache)
            if (unstable) continue;                       // Sim::step aborted (Sim.cpp:107-110)

            // ======================================================= K2a: leaf boxes (a-4), zero collision_forces
            if (lane < n_leaf) {
                const LeafRec r = leaf[lane];
                double2 q = pos[r.pt[0]];
                double minx = q.x, maxx = q.x, miny = q.y, maxy = q.y;
#pragma unroll
                for (int c = 1; c < 4; c++) {
                    q = pos[r.pt[c]];
                    minx = fmin(minx, q.x); maxx = fmax(maxx, q.x);
                    miny = fmin(miny, q.y); maxy = fmax(maxy, q.y);
                }
                aabb[lane] = make_double4(minx, miny, maxx, maxy);
            }
            for (int i = lane; i < Pd; i += 32) cf[i] = make_double2(0.0, 0.0);
            if (lane < 8) misc[lane] = 0;
            __syncwarp();

            // ======================================================= K2b + K3: broad phase, narrow phase, contacts
            // Producer: tested leaf pairs (pair table) and terrain cells under each leaf -> overlapping pairs are
            //           compacted (ballot + popc) into a small candidate buffer.
            // Consumer: every (corner, boxel) item of the candidates, plus the non-shared corners of leaf pairs
            //           that share a point (always candidates), passes the box prefilter -- a necessary condition
            //           of Boxel::point_in_boxel -- into a survivor buffer; survivors go through the out-of-line
            //           resolver and their contacts are applied in list order.
            // Both buffers are drained whenever they fill, so no configuration can overflow them.
            {
                int gx0 = 0, gy0 = 0, gw = 0, gh = 0;
                const int16_t* st_grid = nullptr;
                int cx0 = 0, cy0 = 0, wx = 1, n_cells = 0, max_cells = 0;
                double4 mybb = make_double4(0, 0, 0, 0);
                if (has_static) {
                    gx0 = SHp->gx0; gy0 = SHp->gy0; gw = SHp->gw; gh = SHp->gh;
                    st_grid = reinterpret_cast<const int16_t*>(sb + SHp->o_grid);
                    if (lane < n_leaf) {
                        mybb = aabb[lane];
                        // static leaf in lattice cell c spans [c, c+1] * 0.1: only ceil(min*10 - 1) <= c <= floor(max*10)
                        // can pass the closed-interval test; 1e-6 cells of slack, the exact test decides
                        cx0 = max((int)ceil(mybb.x * 10.0 - 1.000001) - gx0, 0);
                        const int cx1 = min((int)floor(mybb.z * 10.0 + 0.000001) - gx0, gw - 1);
                        cy0 = max((int)ceil(mybb.y * 10.0 - 1.000001) - gy0, 0);
                        const int cy1 = min((int)floor(mybb.w * 10.0 + 0.000001) - gy0, gh - 1);
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
                    // ---- produce candidates
                    int cnt = 0;
                    while (cnt + 32 <= S::kCandCap) {
                        bool hit = false;
                        unsigned int code = 0;
                        if (pair_base < n_pairs) {
                            const int i = pair_base + lane;
                            pair_base += 32;
                            if (i < n_pairs) {
                                code = __ldg(pairs + i);
                                hit = aabb_overlap(aabb[code & kCodeLeafMask], aabb[(code >> 12) & kCodeLeafMask]);
                                if (hit && (code & kCodeCount)) atomicAdd((int*)&misc[code >> kCodeObjShift], 2);   // both ordered pairs (a-12)
                                code &= kCodeLeafMask | (kCodeLeafMask << 12) | kCodeSameObj;
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
                    if (prm.dbg_cand && step == 0 && sub == 0 && !ghost) {
                        for (int i = lane; i < cnt; i += 32)
                            if (dbg_off + i < prm.dbg_cand_cap) prm.dbg_cand[(size_t)env * prm.dbg_cand_cap + dbg_off + i] = cand[i];
                        dbg_off += cnt;
                        if (lane == 0) prm.dbg_cand_n[env] = dbg_off;
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
                                    code = __ldg(items + i);
                                    const int p = leaf[code & kCodeLeafMask].pt[(code >> kItemCornerShift) & 3];
                                    hit = point_in_box(pos[p], aabb[(code >> 12) & kCodeLeafMask]);
                                }
                            } else if (next_cc * 32 < cnt * 8) {
                                // the 8 corners of each candidate pair (these pairs share no point)
                                const int i = next_cc * 32 + lane;
                                next_cc++;
                                if (i < cnt * 8) {
                                    const unsigned int cc = cand[i >> 3];
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
                                more = false;
                                break;
                            }
                            const unsigned m = __ballot_sync(0xffffffffu, hit);
                            if (hit) surv[scnt + __popc(m & ((1u << lane) - 1u))] = code;
                            scnt += __popc(m);
                        }
                        __syncwarp();
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
                                load_leaf(lm, leaf, st_leaf, code & kCodeLeafMask, m_static);
                                load_leaf(lr, leaf, st_leaf, (code >> 12) & kCodeLeafMask, r_static);
                                has = resolve_point_vs_boxel(src, lm, m_static, lr, r_static, (code >> kItemCornerShift) & 3, c);
                            }
                            unsigned m = __ballot_sync(0xffffffffu, has);
                            while (m) {                   // deterministic, race-free: lanes in order
                                const int l = __ffs(m) - 1;
                                if (lane == l) apply_contact(cf, c, reps);
                                m &= m - 1;
                                __syncwarp();
                            }
                        }
                        __syncwarp();
                    }
                }
            }
            // K6a: unstable test uses the boxes of the START of this substep (a-12)
            int sub_unstable = 0;
            {
                const ObjRec* objs = reinterpret_cast<const ObjRec*>(db + Hp->o_obj);
                for (int o = 0; o < n_obj; o++) {
                    const ObjRec orc = objs[o];
                    if (orc.is_robot && misc[o] > (int)orc.n_surface) sub_unstable = 1;
                }
            }

            // ======================================================= K4: actuators + RK4 (a-7, a-8, a-9)
            // two edges per lane per trip: e and e + 32 (the second clamped and its stores predicated)
            for (int e0 = lane; e0 < Ed; e0 += 64) {
                const int e1 = e0 + 32;
                const bool v1 = e1 < Ed;
                const int e1c = v1 ? e1 : e0;
                const uint32_t i0 = eidx[e0], i1 = eidx[e1c];
                const double2 a0 = pos[i0 & 0xff], b0 = pos[(i0 >> 8) & 0xff];
                const double2 a1 = pos[i1 & 0xff], b1 = pos[(i1 >> 8) & 0xff];
                const double dx0 = a0.x - b0.x, dy0 = a0.y - b0.y;
                const double dx1 = a1.x - b1.x, dy1 = a1.y - b1.y;
                const double d0 = sqrt(dx0 * dx0 + dy0 * dy0);
                const double d1 = sqrt(dx1 * dx1 + dy1 * dy1);
                const double add0 = (l0g[e0].y - d0) * 0.006;
                const double add1 = (l0g[e1c].y - d1) * 0.006;
                ekl[e0].y = d0 + add0;
                if (v1) ekl[e1].y = d1 + add1;
            }
            // length_eq[e] is written and read by the same lane: no barrier needed
#pragma unroll 1
            for (int s = 0; s < 4; s++) {
                const double2* sp = (s == 0) ? pos : spos;
                const double2* sv = (s == 0) ? vel : svel;
                for (int e0 = lane; e0 < Ed; e0 += 64) {
                    const int e1 = e0 + 32;
                    const bool v1 = e1 < Ed;
                    const int e1c = v1 ? e1 : e0;
                    const double2 f0 = spring_force(sp, eidx[e0], ekl[e0]);
                    const double2 f1 = spring_force(sp, eidx[e1c], ekl[e1c]);
                    ef[e0] = f0;
                    if (v1) ef[e1] = f1;
                }
                __syncwarp();
                for (int p = lane, k = 0; p < Pd; p += 32, k++) {
                    const double2 v = sv[p];
                    const double2 x = sp[p];
                    double Fx = 0.0, Fy = 0.0;
                    {
                        const uint4 rec = csr[p];
                        const unsigned w[4] = {rec.x, rec.y, rec.z, rec.w};
                        double2 f[8];
#pragma unroll
                        for (int j = 0; j < 8; j++) {
                            const unsigned ent = (w[j >> 1] >> ((j & 1) * 16)) & 0xffffu;
                            f[j] = ef[min(ent >> 1, (unsigned)ECAP)];          // empty slot -> the zero entry
                            if (!(ent & 1u)) { f[j].x = -f[j].x; f[j].y = -f[j].y; }
                        }
#pragma unroll
                        for (int j = 0; j < 8; j++) { Fx += f[j].x; Fy += f[j].y; }
                    }
                    Fx += v.x * (-MEVO_VISCOUS_DRAG);
                    Fy += v.y * (-MEVO_VISCOUS_DRAG);
                    Fy += 1.0 * (-MEVO_GRAVITY);            // every point mass is 1.0 (host-checked), F / m == F
                    const double2 c = cf[p];
                    Fx += c.x; Fy += c.y;
                    if (x.y < 0) {                          // analytic ground plane y = 0 (PhysicsEngine.cpp:566-577)
                        const double normal_force = x.y * (-MEVO_COLLISION_CONST_GROUND);
                        Fy += normal_force;
                        const double friction_temp = ground_friction_tanh(v.x * normal_force * MEVO_DYN_FRICTION);
                        Fx += (-10 * MEVO_DYN_FRICTION) * (normal_force * friction_temp);
                    }
                    const double kpx = v.x * MEVO_DT, kpy = v.y * MEVO_DT;
                    const double kvx = Fx * MEVO_DT, kvy = Fy * MEVO_DT;
                    const double2 x0 = pos[p], v0 = vel[p];
                    if (s == 0) {
                        accp[p] = make_double2(kpx, kpy);
                        accv[p] = make_double2(kvx, kvy);
                        spos[p] = make_double2(x0.x + (kpx * 0.5), x0.y + (kpy * 0.5));
                        svel[p] = make_double2(v0.x + (kvx * 0.5), v0.y + (kvy * 0.5));
                    } else if (s < 3) {
                        double2 q = accp[p]; q.x = q.x + 2 * kpx; q.y = q.y + 2 * kpy; accp[p] = q;
                        q = accv[p]; q.x = q.x + 2 * kvx; q.y = q.y + 2 * kvy; accv[p] = q;
                        const double h = (s == 1) ? 0.5 : 1.0;      // x * 1.0 == x exactly
                        spos[p] = make_double2(x0.x + (kpx * h), x0.y + (kpy * h));
                        svel[p] = make_double2(v0.x + (kvx * h), v0.y + (kvy * h));
                    } else {
                        const double2 qp = accp[p], qv = accv[p];
                        double dpx = (qp.x + kpx) / 6, dpy = (qp.y + kpy) / 6;
                        double dvx = (qv.x + kvx) / 6, dvy = (qv.y + kvy) / 6;
                        if ((fixmask >> k) & 1u) { dpx = 0.0; dpy = 0.0; dvx = 0.0; dvy = 0.0; }
                        // the new position goes to spos: pos keeps the substep's start position (= pos_last)
                        spos[p] = make_double2(x0.x + dpx, x0.y + dpy);
                        vel[p] = make_double2(v0.x + dvx, v0.y + dvy);
                    }
                }
                __syncwarp();
            }

            // ======================================================= K5: edge constraints + true velocity (a-10, a-11)
            // strain-band fast path on squared lengths (no sqrt): inside the guarded band every correction of the
            // reference is exactly zero; otherwise the exact, ordered passes run out of line
            {
                bool viol = false;
                for (int e = lane; e < Ed; e += 32) {
                    const uint32_t ix = eidx[e];
                    const double2 pa = spos[ix & 0xff], pb = spos[(ix >> 8) & 0xff];
                    const double dx = pa.x - pb.x, dy = pa.y - pb.y;
                    const double d2 = dx * dx + dy * dy;
                    const double l0 = l0g[e].x;
                    const double l2 = l0 * l0;
                    const bool rigid = (ix & (kEdgeRigidCell << 16)) != 0;
                    const double lo = rigid ? 0.9409001 : 0.5625001, hi = rigid ? 1.0608999 : 1.5624999;
                    if (!(d2 > l2 * lo && d2 < l2 * hi)) viol = true;
                }
                if (__any_sync(0xffffffffu, viol))
                    warp_exact_constraints(spos, ef, l0g, eidx, csr, Pd, Ed, fixmask, ECAP,
                                           reinterpret_cast<const uint16_t*>(db + Hp->o_rlev),
                                           reinterpret_cast<const uint16_t*>(db + Hp->o_rvis), Hp->n_rlev);
            }
            for (int p = lane; p < Pd; p += 32) {        // PhysicsEngine::update_true_vel (:602-607)
                const double2 x = spos[p];
                // pos_last == pos at the start of every substep except possibly the first of a launch
                const double2 l = first_substep ? g2[2 * Pd + p] : pos[p];
                vt[p] = make_double2((x.x - l.x) / MEVO_DT, (x.y - l.y) / MEVO_DT);
                pos[p] = x;
            }
            first_substep = false;
            __syncwarp();
            if (sub_unstable) unstable = 1;               // Sim::step aborts the remaining substeps (Sim.cpp:107-110)
        }
        if (!unstable) completed_steps++;
    }

    // ---------------------------------------------------------------- stage out
    if (!ghost) {
        double2* o2 = reinterpret_cast<double2*>(gstate);
        if (unstable && prm.auto_reset) {
            // Sim::revert(0) in place (Environment.cpp:247-263)
            const double2* s0 = reinterpret_cast<const double2*>(prm.state0 + ed.state_off);
            for (int i = lane; i < Pd; i += 32) {
                const double2 p0 = s0[i], v0 = s0[Pd + i];
                o2[i] = p0; o2[Pd + i] = v0; o2[2 * Pd + i] = p0; o2[3 * Pd + i] = v0;
            }
        } else if (!first_substep) {
            for (int i = lane; i < Pd; i += 32) {
                const double2 x = pos[i];
                o2[i] = x;
                o2[Pd + i] = vel[i];
                o2[2 * Pd + i] = x;                        // pos_last = pos after update_true_vel
                o2[3 * Pd + i] = vt[i];
            }
        }
        double* ggoal = gstate + 8 * (size_t)Pd;
        for (int e = lane; e < Ed; e += 32) ggoal[e] = l0g[e].y;
        if (lane == 0) {
            prm.unstable[env] = (uint8_t)unstable;
            if (unstable && prm.auto_reset) prm.sim_time[env] = 0;
            else prm.sim_time[env] += completed_steps;
            if (unstable) atomicAdd(prm.unstable_events, 1ull);
        }
    }
}

}  // namespace mevo
