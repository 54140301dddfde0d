// Warp-per-environment step kernel for small worlds (one 5x5-class robot, optionally a small free object):
// the Walker-v0 / Pusher-v0 / Thrower-v0 configurations of BASELINE.json.
//
// Same arithmetic as gstep_kernel (gkernel.cuh) -- the two agree bit for bit and tests run both -- but laid out
// for what bounds this path on B200 (ncu: DRAM ~ 0, FP64 pipe < 20 %, issue slots < 40 %): each environment is one
// long dependent chain (FP64 sqrt / div sequences, shared-memory round trips, ballots), shared memory allows 16 of
// them per SM, the substep body is larger than the instruction cache, and with the barrier that keeps the warps in
// step a substep lasts as long as its slowest warp.  Hence
//   * everything a substep touches lives in shared memory in compact form (16-bit pair / item codes, 16-byte
//     gather records, an edge-type table instead of per-edge constants): no global load inside the loop,
//   * compile-time layout (template capacities): one base register per warp, constant offsets,
//   * every spring force is stored with both signs, so the per-point gather is 8 x (LDS.128 + 2 DADD),
//   * independent work per lane is kept in flight: three edges in the spring pass (length_eq is relaxed inside
//     the first stage's pass), two pairs / items in the collision trips; box tests are branch-free,
//   * strain limiters: squared-length band test (no sqrt); the Jacobi limiter evaluates only the edges outside the
//     band; the ordered rigid-cell pass evaluates 32 visits at once and re-evaluates only what the applied visit
//     touched,
//   * the contact resolver keeps its leaf records in registers, decides segment crossings without dividing, and
//     accumulates a contact's forces in registers,
//   * division by a compile-time constant is three instructions and bit-identical (div_const, kernels.cuh),
//   * one environment per warp, 16 warps = one CTA per SM; a CTA barrier every substep keeps them in the same code
//     region so they share instruction-cache lines; the host sizes CTAs for equally full waves, or issues the step as
//     (CTA-group, chunk) work items that the CTAs draw from a counter (wstep_kernel below),
//   * stage-in by the bulk-copy engine: the design blob holds the tables in the form shared memory wants them.
#pragma once

#include "kernels.cuh"

namespace mevo {

// ---- bulk copies global -> shared memory (the TMA unit's 1-D form) completing on an mbarrier
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// dst, src 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(__cvta_generic_to_global(src)), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra MBAR_DONE;\n"
        "bra MBAR_WAIT;\n"
        "MBAR_DONE:\n"
        "}" ::"r"(bar),
        "r"(parity)
        : "memory");
}

template <int PCAP, int ECAP, int LCAP, int PAIRCAP, int ITEMCAP, int VISCAP>
struct WarpSmem {
    static constexpr int kCandCap = 64;            // overlapping pairs per round (two 32-wide chunks)
    static constexpr int kSurvCap = 64;            // prefilter survivors per round
    static constexpr int kZero = 2 * ECAP;         // index of the all-zero entry of ef2
    static constexpr int o_pos = 0;                // double2 per point ...
    static constexpr int o_spos = o_pos + 16 * PCAP;
    static constexpr int o_vel = o_spos + 16 * PCAP;
    static constexpr int o_svel = o_vel + 16 * PCAP;
    static constexpr int o_vt = o_svel + 16 * PCAP;
    static constexpr int o_cf = o_vt + 16 * PCAP;
    static constexpr int o_accp = o_cf + 16 * PCAP;
    static constexpr int o_accv = o_accp + 16 * PCAP;
    static constexpr int o_csr = o_accv + 16 * PCAP;           // uint4 per point: 8 x u16 index into ef2 (edge << 1 | side)
    static constexpr int o_ef2 = o_csr + 16 * PCAP;            // double2 x 2 per edge: [2e] = -f (a side), [2e+1] = +f (b side)
    static constexpr int o_leq = o_ef2 + 16 * (2 * ECAP + 1);  // double per edge: length_eq
    static constexpr int o_goal = o_leq + 8 * ECAP;            // double per edge: length_eq_goal
    static constexpr int o_etab = o_goal + 8 * ECAP;           // double2 x 16: (spring constant, init_length_eq) per edge type
    static constexpr int o_aabb = o_etab + 16 * 16;            // double4 per leaf
    static constexpr int o_eidx = o_aabb + 32 * LCAP;          // u32 per edge: a | b << 8 | type << 16 | rigid-cell << 20
    static constexpr int o_lpt = o_eidx + 4 * ECAP;            // u32 per leaf: the 4 corner point ids (TL,TR,BR,BL), u8 each
    static constexpr int o_cand = o_lpt + 4 * LCAP;
    static constexpr int o_surv = o_cand + 4 * kCandCap;
    static constexpr int o_pairs = o_surv + 4 * kSurvCap;      // u16 pair codes
    static constexpr int o_items = o_pairs + 2 * PAIRCAP;      // u16 item codes
    static constexpr int o_rord = o_items + 2 * ITEMCAP;       // u8 edge id per rigid-cell visit, reference order
    static constexpr int o_misc = (o_rord + VISCAP + 3) & ~3;  // int[8] self-collision counts, int[8] n_surface (robots) or INT_MAX
    static constexpr int o_hdr = (o_misc + 4 * 16 + 7) & ~7;   // WarpHdr: per-environment scalars and pointers
    static constexpr int kBytes = (o_hdr + 96 + 15) & ~15;
};

// Per-environment scalars the phases need, kept in shared memory instead of registers (the phases are separate
// out-of-line functions so that each gets the whole register budget; broadcast LDS is cheap).
struct WarpHdr {                                   // 96 bytes
    int Pd, Ed, n_leaf, n_pairs, n_items, n_vis;
    int gx0, gy0, gw, gh;                          // terrain grid (gw = 0: no static terrain)
    const int16_t* st_grid;
    const double4* st_aabb;
    const double2* st_pos;
    const LeafRec* st_leaf;
    const LeafRec* g_leaf;                         // full dynamic leaf records (resolver only)
    const double2* g_last;                         // pos_last in HBM: read in the first substep of a launch only
};
static_assert(sizeof(WarpHdr) == 88, "WarpHdr layout");

#ifdef MEVO_PROF
// development only (-DMEVO_PROF): per phase, the sum over warps of their own time and the sum over CTA-substeps of
// the slowest warp's time x warps -- their ratio is what the per-substep barrier costs
__device__ unsigned long long g_wprof[16];
__device__ unsigned long long g_whist[3][16][2];     // [collide by survivors / collide by candidates / constrain by rigid moves][bucket][cycles, count]
#endif

// tanh of the analytic ground plane's friction term (PhysicsEngine.cpp:573-577); out of line: a robot standing
// on voxel terrain never gets below y = 0
__device__ __noinline__ double ground_friction_tanh(double x) { return tanh(x); }

// Jacobi half of PhysicsEngine::resolve_edge_constraints (PhysicsEngine.cpp:613-671) for one warp; out of line:
// runs only in substeps where some edge left the +-25 % band.  `list` holds the n_list edges outside the (guarded)
// band: every other edge's correction is exactly zero, so only those are evaluated (sqrt + three divisions each).
// list == nullptr: all edges.
__device__ __noinline__ void warp_jacobi_limiter(double2* pos, double2* ef2, const uint32_t* eidx, const double2* etab,
                                                 const uint4* csr, int Pd, int Ed, unsigned fixmask, int zero_idx,
                                                 const unsigned int* list, int n_list) {
    const int lane = threadIdx.x & 31;
    if (list) {
        for (int i = lane; i < 2 * Ed; i += 32) ef2[i] = make_double2(0.0, 0.0);
        __syncwarp();
    } else {
        n_list = Ed;
    }
    for (int i = lane; i < n_list; i += 32) {
        const int e = list ? (int)list[i] : i;
        const uint32_t ix = eidx[e];
        const double2 pa = pos[ix & 0xff], pb = pos[(ix >> 8) & 0xff];
        const double dx = pa.x - pb.x, dy = pa.y - pb.y;
        const double dist = sqrt(dx * dx + dy * dy);
        const double l0 = etab[(ix >> kEidxTypeShift) & 15].y;
        const double over_mag = (dist - l0 * (1 + 0.25)) / (dist * 2);
        const double under_mag = (dist - l0 * (1 - 0.25)) / (dist * 2);
        double ox = dx * over_mag, oy = dy * over_mag;
        double ux = dx * under_mag, uy = dy * under_mag;
        const double stress = dist / l0;
        if (!(stress > (1 + 0.25))) { ox = 0.0; oy = 0.0; }
        if (!(stress < (1 - 0.25))) { ux = 0.0; uy = 0.0; }
        const double cx = (ox + ux) * 0.8, cy = (oy + uy) * 0.8;
        ef2[2 * e] = make_double2(-cx, -cy);
        ef2[2 * e + 1] = make_double2(cx, cy);
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
            const double2 c = ef2[ent];
            x.x += c.x; x.y += c.y;
        }
        pos[p] = x;
    }
    __syncwarp();
}

// The +-3 % limiter of one rigid-cell visit (PhysicsEngine.cpp:675-710): does the visit move its end points, and by
// what.  The reference applies pa -= c, pb += c on every visit; outside (0.97, 1.03) c is what is computed here,
// inside it is zero.
__device__ __forceinline__ bool rigid_visit_eval(const double2* pos, int a, int b, double l0, double& cx, double& cy) {
    const double2 pa = pos[a], pb = pos[b];
    const double dx = pa.x - pb.x, dy = pa.y - pb.y;
    const double dist = sqrt(dx * dx + dy * dy);
    const double over_mag = (dist - l0 * (1 + 0.03)) / (dist * 2);
    const double under_mag = (dist - l0 * (1 - 0.03)) / (dist * 2);
    double ox = dx * over_mag, oy = dy * over_mag;
    double ux = dx * under_mag, uy = dy * under_mag;
    const double stress = dist / l0;
    const bool no_over = stress < (1 + 0.03), no_under = stress > (1 - 0.03);
    if (no_over) { ox = 0.0; oy = 0.0; }
    if (no_under) { ux = 0.0; uy = 0.0; }
    cx = (ox + ux) * 0.5; cy = (oy + uy) * 0.5;
    return !(no_over && no_under);
}

// Spring force of one edge for stage positions sp (PhysicsEngine.cpp:498-528).  q is the edge's length_eq; in the
// first stage of a substep it comes in as length_eq_goal and leaves as the relaxed length_eq
// (PhysicsEngine::update_actuators: the stage's distance is the substep's start distance).
__device__ __forceinline__ double2 spring_force(const double2* sp, uint32_t ix, double k, double& q, bool first_stage) {
    const double2 pa = sp[ix & 0xff], pb = sp[(ix >> 8) & 0xff];
    const double dx = pa.x - pb.x, dy = pa.y - pb.y;
    const double dist = sqrt(dx * dx + dy * dy);
    if (first_stage) {
        const double add = (q - dist) * 0.006;
        q = dist + add;
    }
    const double mag = (dist - q) * k / dist;
    return make_double2(dx * mag, dy * mag);
}

__device__ __forceinline__ double flip_sign(double x) {       // -x on the integer pipe (exact)
    return __hiloint2double(__double2hiint(x) ^ 0x80000000, __double2loint(x));
}


// =================================================================================== phase 1: collisions
// K2a leaf boxes, K2b broad phase, K3 narrow phase + contact response, K6 unstable count (SURVEY 8a-4,5,6,12).
// Returns 1 when PhysicsEngine::is_any_robot_self_colliding would (boxes of the START of the substep).
//
// Producer: tested leaf pairs (pair table) and terrain cells under each leaf -> overlapping pairs are compacted
//           (ballot + popc) into a small candidate buffer.
// Consumer: every (corner, boxel) item of the candidates, plus the non-shared corners of leaf pairs that share a
//           point (always candidates), passes the box prefilter -- a necessary condition of
//           Boxel::point_in_boxel -- into a survivor buffer; survivors go through the out-of-line resolver and
//           their contacts are applied in list order.
// Both buffers are drained whenever they fill, so no configuration can overflow them.
template <class S>
__device__ __noinline__ int warp_collide(unsigned char* sm, unsigned int* dbg_row, int* dbg_n, int dbg_cap, double2* dbg_cf) {
    const int lane = threadIdx.x & 31;
    const WarpHdr* H = reinterpret_cast<const WarpHdr*>(sm + S::o_hdr);
    double2* pos = reinterpret_cast<double2*>(sm + S::o_pos);
    double2* vt = reinterpret_cast<double2*>(sm + S::o_vt);
    double2* cf = reinterpret_cast<double2*>(sm + S::o_cf);
    double4* aabb = reinterpret_cast<double4*>(sm + S::o_aabb);
    const uint32_t* lpt = reinterpret_cast<const uint32_t*>(sm + S::o_lpt);
    unsigned int* cand = reinterpret_cast<unsigned int*>(sm + S::o_cand);
    unsigned int* surv = reinterpret_cast<unsigned int*>(sm + S::o_surv);
    const uint16_t* pairs = reinterpret_cast<const uint16_t*>(sm + S::o_pairs);
    const uint16_t* items = reinterpret_cast<const uint16_t*>(sm + S::o_items);
    volatile int* misc = reinterpret_cast<volatile int*>(sm + S::o_misc);
    const int Pd = H->Pd, n_leaf = H->n_leaf, n_pairs = H->n_pairs, n_items = H->n_items;

    // ---- K2a: leaf boxes (a-4), zero collision_forces
    if (lane < n_leaf) {
        const uint32_t lp = lpt[lane];
        double2 q = pos[lp & 0xff];
        double minx = q.x, maxx = q.x, miny = q.y, maxy = q.y;
#pragma unroll
        for (int c = 1; c < 4; c++) {
            q = pos[(lp >> (8 * c)) & 0xff];
            minx = fmin(minx, q.x); maxx = fmax(maxx, q.x);
            miny = fmin(miny, q.y); maxy = fmax(maxy, q.y);
        }
        aabb[lane] = make_double4(minx, miny, maxx, maxy);
    }
    for (int i = lane; i < Pd; i += 32) cf[i] = make_double2(0.0, 0.0);
    if (lane < 8) misc[lane] = 0;
    __syncwarp();

    // ---- terrain window of this lane's leaf
    const int gw = H->gw;
    const int16_t* st_grid = H->st_grid;
    const double4* st_aabb = H->st_aabb;
    int cx0 = 0, cy0 = 0, wx = 1, n_cells = 0, max_cells = 0;
    double4 mybb = make_double4(0, 0, 0, 0);
    if (gw > 0) {
        if (lane < n_leaf) {
            mybb = aabb[lane];
            // static leaf in lattice cell c spans [c, c+1] * 0.1: only ceil(min*10 - 1) <= c <= floor(max*10) can
            // pass the closed-interval test; 1e-6 cells of slack, the exact test decides
            cx0 = max((int)ceil(mybb.x * 10.0 - 1.000001) - H->gx0, 0);
            const int cx1 = min((int)floor(mybb.z * 10.0 + 0.000001) - H->gx0, gw - 1);
            cy0 = max((int)ceil(mybb.y * 10.0 - 1.000001) - H->gy0, 0);
            const int cy1 = min((int)floor(mybb.w * 10.0 + 0.000001) - H->gy0, H->gh - 1);
            wx = cx1 - cx0 + 1;
            const int wy = cy1 - cy0 + 1;
            n_cells = (wx > 0 && wy > 0) ? wx * wy : 0;
            if (n_cells == 0) wx = 1;
        }
        max_cells = __reduce_max_sync(0xffffffffu, n_cells);
    }

    int pair_base = 0, cell_it = 0, next_ic = 0, dbg_off = 0;
#ifdef MEVO_PROF
    const long long tc0 = clock64();
    int prof_cand = 0, prof_surv = 0;
#endif
    bool producing = true;
    while (producing) {
        // ---- produce candidates (32-bit codes of kernels.cuh): tested leaf pairs first, then terrain cells
        int cnt = 0;
        while (pair_base < n_pairs && cnt + 64 <= S::kCandCap) {      // two pairs per lane in flight
            const int i0 = pair_base + lane, i1 = i0 + 32;
            const unsigned pc0 = pairs[i0 < n_pairs ? i0 : pair_base], pc1 = pairs[i1 < n_pairs ? i1 : pair_base];
            pair_base += 64;
            const unsigned a0 = pc0 & 31u, b0 = (pc0 >> 5) & 31u, a1 = pc1 & 31u, b1 = (pc1 >> 5) & 31u;
            const bool hit0 = (i0 < n_pairs) & aabb_overlap(aabb[a0], aabb[b0]);
            const bool hit1 = (i1 < n_pairs) & aabb_overlap(aabb[a1], aabb[b1]);
            const unsigned m0 = __ballot_sync(0xffffffffu, hit0), m1 = __ballot_sync(0xffffffffu, hit1);
            if (m0 | m1) {
                if (hit0 && (pc0 & kPair16Count)) atomicAdd((int*)&misc[pc0 >> kPair16ObjShift], 2);   // both ordered pairs (a-12)
                if (hit1 && (pc1 & kPair16Count)) atomicAdd((int*)&misc[pc1 >> kPair16ObjShift], 2);
                if (hit0) cand[cnt + __popc(m0 & ((1u << lane) - 1u))] = a0 | (b0 << 12) | ((pc0 & kPair16Same) ? kCodeSameObj : 0u);
                cnt += __popc(m0);
                if (hit1) cand[cnt + __popc(m1 & ((1u << lane) - 1u))] = a1 | (b1 << 12) | ((pc1 & kPair16Same) ? kCodeSameObj : 0u);
                cnt += __popc(m1);
            }
        }
        while (pair_base < n_pairs && cnt + 32 <= S::kCandCap) {
            const int i = pair_base + lane;              // lanes past the end repeat a valid pair: no branch
            const unsigned pc = pairs[i < n_pairs ? i : pair_base];
            pair_base += 32;
            const unsigned a = pc & 31u, b = (pc >> 5) & 31u;
            const bool hit = (i < n_pairs) & aabb_overlap(aabb[a], aabb[b]);
            if (hit && (pc & kPair16Count)) atomicAdd((int*)&misc[pc >> kPair16ObjShift], 2);   // both ordered pairs (a-12)
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (hit) cand[cnt + __popc(m & ((1u << lane) - 1u))] = a | (b << 12) | ((pc & kPair16Same) ? kCodeSameObj : 0u);
            cnt += __popc(m);
        }
        if (pair_base >= n_pairs) {
            while (cell_it < max_cells && cnt + 32 <= S::kCandCap) {
                const int it = cell_it++;
                bool hit = false;
                unsigned int code = 0;
                if (it < n_cells) {
                    const int cx = cx0 + it % wx, cy = cy0 + it / wx;
                    const int sl = __ldg(st_grid + cy * gw + cx);
                    if (sl >= 0) {
                        hit = aabb_overlap(mybb, ldg_d4(st_aabb + sl));
                        code = (unsigned)lane | ((unsigned)sl << 12) | kCodeStatic;
                    }
                }
                const unsigned m = __ballot_sync(0xffffffffu, hit);
                if (hit) cand[cnt + __popc(m & ((1u << lane) - 1u))] = code;
                cnt += __popc(m);
            }
            if (cell_it >= max_cells) producing = false;
        }
        __syncwarp();
#ifdef MEVO_PROF
        prof_cand += cnt;
#endif
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
            // non-shared corners of leaf pairs that share a point (always candidates) first ...
            while (next_ic * 32 < n_items && scnt + 64 <= S::kSurvCap) {      // two items per lane in flight
                const int i0 = next_ic * 32 + lane, i1 = i0 + 32;
                const unsigned ic0 = items[i0 < n_items ? i0 : next_ic * 32], ic1 = items[i1 < n_items ? i1 : next_ic * 32];
                next_ic += 2;
                const unsigned mn0 = ic0 & 31u, rf0 = (ic0 >> 5) & 31u, cr0 = (ic0 >> 10) & 3u;
                const unsigned mn1 = ic1 & 31u, rf1 = (ic1 >> 5) & 31u, cr1 = (ic1 >> 10) & 3u;
                const bool hit0 = (i0 < n_items) & point_in_box(pos[(lpt[mn0] >> (8 * cr0)) & 0xff], aabb[rf0]);
                const bool hit1 = (i1 < n_items) & point_in_box(pos[(lpt[mn1] >> (8 * cr1)) & 0xff], aabb[rf1]);
                const unsigned m0 = __ballot_sync(0xffffffffu, hit0), m1 = __ballot_sync(0xffffffffu, hit1);
                if (m0 | m1) {
                    if (hit0) surv[scnt + __popc(m0 & ((1u << lane) - 1u))] = mn0 | (rf0 << 12) | (cr0 << kItemCornerShift) | kItemSameObj;
                    scnt += __popc(m0);
                    if (hit1) surv[scnt + __popc(m1 & ((1u << lane) - 1u))] = mn1 | (rf1 << 12) | (cr1 << kItemCornerShift) | kItemSameObj;
                    scnt += __popc(m1);
                }
            }
            while (next_ic * 32 < n_items && scnt + 32 <= S::kSurvCap) {
                const int i = next_ic * 32 + lane;
                const unsigned ic = items[i < n_items ? i : next_ic * 32];
                next_ic++;
                const unsigned mn = ic & 31u, rf = (ic >> 5) & 31u, corner = (ic >> 10) & 3u;
                const bool hit = (i < n_items) & point_in_box(pos[(lpt[mn] >> (8 * corner)) & 0xff], aabb[rf]);
                const unsigned m = __ballot_sync(0xffffffffu, hit);
                if (hit) surv[scnt + __popc(m & ((1u << lane) - 1u))] = mn | (rf << 12) | (corner << kItemCornerShift) | kItemSameObj;
                scnt += __popc(m);
            }
            // ... then the 8 corners of each candidate pair (these pairs share no point)
            if (next_ic * 32 >= n_items) {
                while (next_cc * 32 < cnt * 8 && scnt + 32 <= S::kSurvCap) {
                    const int i = next_cc * 32 + lane;
                    next_cc++;
                    bool hit = false;
                    unsigned int code = 0;
                    if (i < cnt * 8) {
                        const unsigned int cc = cand[i >> 3];
                        const unsigned a = cc & kCodeLeafMask, b = (cc >> 12) & kCodeLeafMask;
                        const bool b_static = (cc & kCodeStatic) != 0;
                        const int dir = (i >> 2) & 1, corner = i & 3;
                        const unsigned same = (cc & kCodeSameObj) ? kItemSameObj : 0u;
                        if (dir == 0) {
                            code = a | (b << 12) | ((unsigned)corner << kItemCornerShift) | (b_static ? kItemRefStatic : 0u) | same;
                            const double4 bb = b_static ? ldg_d4(st_aabb + b) : aabb[b];
                            hit = point_in_box(pos[(lpt[a] >> (8 * corner)) & 0xff], bb);
                        } else {
                            code = b | (a << 12) | ((unsigned)corner << kItemCornerShift) | (b_static ? kItemMainStatic : 0u) | same;
                            const double2 pp = b_static ? __ldg(H->st_pos + __ldg(&H->st_leaf[b].pt[corner]))
                                                        : pos[(lpt[b] >> (8 * corner)) & 0xff];
                            hit = point_in_box(pp, aabb[a]);
                        }
                    }
                    const unsigned m = __ballot_sync(0xffffffffu, hit);
                    if (hit) surv[scnt + __popc(m & ((1u << lane) - 1u))] = code;
                    scnt += __popc(m);
                }
                if (next_cc * 32 >= cnt * 8) more = false;
            }
            __syncwarp();
#ifdef MEVO_PROF
            prof_surv += scnt;
#endif
            if (scnt > 0) {
                PointSrc src;
                src.dyn_pos = pos; src.dyn_vt = vt; src.st_pos = H->st_pos;
                for (int base = 0; base < scnt; base += 32) {
                    const int i = base + lane;
                    bool has = false;
                    Contact c;
                    int reps = 1;
                    if (i < scnt) {
                        const unsigned code = surv[i];
                        const bool m_static = (code & kItemMainStatic) != 0, r_static = (code & kItemRefStatic) != 0;
                        reps = (code & kItemSameObj) ? 2 : 1;
                        LeafRegs lm, lr;
                        load_leaf(lm, H->g_leaf, H->st_leaf, code & kCodeLeafMask, m_static);
                        load_leaf(lr, H->g_leaf, H->st_leaf, (code >> 12) & kCodeLeafMask, r_static);
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
#ifdef MEVO_PROF
    if (lane == 0) {
        const unsigned long long d = clock64() - tc0;
        const int bs = min(prof_surv / 4, 15), bc = min(prof_cand / 4, 15);
        atomicAdd(&g_whist[0][bs][0], d); atomicAdd(&g_whist[0][bs][1], 1ull);
        atomicAdd(&g_whist[1][bc][0], d); atomicAdd(&g_whist[1][bc][1], 1ull);
    }
#endif
    if (dbg_cf)                                        // test hook: collision_forces of this substep
        for (int i = lane; i < Pd; i += 32) dbg_cf[i] = cf[i];
    // K6a: unstable test uses the boxes of the START of this substep (a-12)
    return __any_sync(0xffffffffu, lane < 8 && misc[lane] > misc[8 + lane]);
}

// =================================================================================== phase 2: actuators + RK4
// PhysicsEngine::update_actuators + step_rk4 + get_pos_vel_slope (SURVEY 8a-7,8,9).  On return spos holds the new
// positions (pos still holds the substep's start positions = pos_last) and vel the new velocities.
template <class S>
__device__ __noinline__ void warp_rk4(unsigned char* sm, unsigned fixmask) {
    const int lane = threadIdx.x & 31;
    const WarpHdr* H = reinterpret_cast<const WarpHdr*>(sm + S::o_hdr);
    double2* pos = reinterpret_cast<double2*>(sm + S::o_pos);
    double2* spos = reinterpret_cast<double2*>(sm + S::o_spos);
    double2* vel = reinterpret_cast<double2*>(sm + S::o_vel);
    double2* svel = reinterpret_cast<double2*>(sm + S::o_svel);
    const double2* cf = reinterpret_cast<const double2*>(sm + S::o_cf);
    double2* accp = reinterpret_cast<double2*>(sm + S::o_accp);
    double2* accv = reinterpret_cast<double2*>(sm + S::o_accv);
    const uint4* csr = reinterpret_cast<const uint4*>(sm + S::o_csr);
    double2* ef2 = reinterpret_cast<double2*>(sm + S::o_ef2);
    double* leq = reinterpret_cast<double*>(sm + S::o_leq);
    const double* goal = reinterpret_cast<const double*>(sm + S::o_goal);
    const double2* etab = reinterpret_cast<const double2*>(sm + S::o_etab);
    const uint32_t* eidx = reinterpret_cast<const uint32_t*>(sm + S::o_eidx);
    const int Pd = H->Pd, Ed = H->Ed;

#pragma unroll 1
    for (int s = 0; s < 4; s++) {
        const double2* sp = (s == 0) ? pos : spos;
        const double2* sv = (s == 0) ? vel : svel;
        // Spring pass: three edges per lane in flight (e, e + 32, e + 64: three independent sqrt / div chains); lanes
        // past the end recompute a valid edge and store nothing.  Stage 0 also does PhysicsEngine::update_actuators:
        // its distance is the one length_eq is relaxed from (same positions, same expression).
        for (int base = 0; base < Ed; base += 96) {
            const int e0 = base + lane, e1 = e0 + 32, e2 = e0 + 64;
            const bool v0 = e0 < Ed, v1 = e1 < Ed, v2 = e2 < Ed;
            const int c0 = v0 ? e0 : base, c1 = v1 ? e1 : c0, c2 = v2 ? e2 : c0;
            const uint32_t i0 = eidx[c0], i1 = eidx[c1], i2 = eidx[c2];
            double q0, q1, q2;                          // stage 0: length_eq_goal; later stages: length_eq
            if (s == 0) { q0 = goal[c0]; q1 = goal[c1]; q2 = goal[c2]; }
            else { q0 = leq[c0]; q1 = leq[c1]; q2 = leq[c2]; }
            const double2 f0 = spring_force(sp, i0, etab[(i0 >> kEidxTypeShift) & 15].x, q0, s == 0);
            const double2 f1 = spring_force(sp, i1, etab[(i1 >> kEidxTypeShift) & 15].x, q1, s == 0);
            const double2 f2 = spring_force(sp, i2, etab[(i2 >> kEidxTypeShift) & 15].x, q2, s == 0);
            if (v0) {
                ef2[2 * e0] = make_double2(flip_sign(f0.x), flip_sign(f0.y));
                ef2[2 * e0 + 1] = f0;
            }
            if (v1) {
                ef2[2 * e1] = make_double2(flip_sign(f1.x), flip_sign(f1.y));
                ef2[2 * e1 + 1] = f1;
            }
            if (v2) {
                ef2[2 * e2] = make_double2(flip_sign(f2.x), flip_sign(f2.y));
                ef2[2 * e2 + 1] = f2;
            }
            if (s == 0) {
                if (v0) leq[e0] = q0;
                if (v1) leq[e1] = q1;
                if (v2) leq[e2] = q2;
            }
        }
        __syncwarp();
        for (int p = lane, k = 0; p < Pd; p += 32, k++) {
            const double2 v = sv[p];
            double Fx = 0.0, Fy = 0.0;
            {
                // F -= f over the a sides in edge order, then F += f over the b sides (PhysicsEngine.cpp:530-531): the
                // record lists them in that order, the sign is in the stored value, empty slots read zero
                const uint4 rec = csr[p];
                const unsigned w[4] = {rec.x, rec.y, rec.z, rec.w};
#pragma unroll
                for (int half = 0; half < 2; half++) {
                    double2 f[4];
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const int q = 4 * half + j;
                        const unsigned ent = (w[q >> 1] >> ((q & 1) * 16)) & 0xffffu;
                        f[j] = ef2[ent];
                    }
#pragma unroll
                    for (int j = 0; j < 4; j++) { Fx += f[j].x; Fy += f[j].y; }
                }
            }
            Fx += v.x * (-MEVO_VISCOUS_DRAG);
            Fy += v.y * (-MEVO_VISCOUS_DRAG);
            Fy += 1.0 * (-MEVO_GRAVITY);                    // every point mass is 1.0 (host-checked), F / m == F
            const double2 c = cf[p];
            Fx += c.x; Fy += c.y;
            const double y = sp[p].y;
            if (y < 0) {                                    // analytic ground plane y = 0 (PhysicsEngine.cpp:566-577)
                const double normal_force = y * (-MEVO_COLLISION_CONST_GROUND);
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
                double dpx = div_const(qp.x + kpx, 6.0), dpy = div_const(qp.y + kpy, 6.0);
                double dvx = div_const(qv.x + kvx, 6.0), dvy = div_const(qv.y + kvy, 6.0);
                if ((fixmask >> k) & 1u) { dpx = 0.0; dpy = 0.0; dvx = 0.0; dvy = 0.0; }
                // the new position goes to spos: pos keeps the substep's start position (= pos_last)
                spos[p] = make_double2(x0.x + dpx, x0.y + dpy);
                vel[p] = make_double2(v0.x + dvx, v0.y + dvy);
            }
        }
        __syncwarp();
    }
}

// =================================================================================== phase 3: limiters + true velocity
// PhysicsEngine::resolve_edge_constraints + update_true_vel (SURVEY 8a-10,11) on the new positions in spos.
// Band test on squared lengths (no sqrt), guarded by 1e-7 so the decision never depends on the rounding of
// dist / L0: inside the band every correction of the reference is exactly zero.
template <class S>
__device__ __noinline__ void warp_constrain(unsigned char* sm, unsigned fixmask, bool first_substep) {
    const int lane = threadIdx.x & 31;
    const WarpHdr* H = reinterpret_cast<const WarpHdr*>(sm + S::o_hdr);
    double2* pos = reinterpret_cast<double2*>(sm + S::o_pos);
    double2* spos = reinterpret_cast<double2*>(sm + S::o_spos);
    double2* vt = reinterpret_cast<double2*>(sm + S::o_vt);
    const uint4* csr = reinterpret_cast<const uint4*>(sm + S::o_csr);
    double2* ef2 = reinterpret_cast<double2*>(sm + S::o_ef2);
    const double2* etab = reinterpret_cast<const double2*>(sm + S::o_etab);
    const uint32_t* eidx = reinterpret_cast<const uint32_t*>(sm + S::o_eidx);
    const uint8_t* rord = reinterpret_cast<const uint8_t*>(sm + S::o_rord);
    const int Pd = H->Pd, Ed = H->Ed, n_vis = H->n_vis;

#ifdef MEVO_PROF
    const long long tk0 = clock64();
    int prof_moves = 0;
#endif
    // edges outside the +-25 % band are listed (candidate + survivor buffers are free in this phase)
    unsigned int* jlist = reinterpret_cast<unsigned int*>(sm + S::o_cand);
    constexpr int kJCap = S::kCandCap + S::kSurvCap;
    static_assert(S::o_surv == S::o_cand + 4 * S::kCandCap, "the two buffers are contiguous");
    int nj = 0;
    bool rviol = false;
    for (int e0 = 0; e0 < Ed; e0 += 128) {                // four edges per lane in flight, then their ballots
        bool jv[4];
#pragma unroll
        for (int t = 0; t < 4; t++) {
            const int e = e0 + 32 * t + lane;
            const bool valid = e < Ed;
            const uint32_t ix = eidx[valid ? e : e0];
            const double2 pa = spos[ix & 0xff], pb = spos[(ix >> 8) & 0xff];
            const double dx = pa.x - pb.x, dy = pa.y - pb.y;
            const double d2 = dx * dx + dy * dy;
            const double l0 = etab[(ix >> kEidxTypeShift) & 15].y;
            const double l2 = l0 * l0;
            jv[t] = valid && !(d2 > l2 * 0.5625001 && d2 < l2 * 1.5624999);
            if (valid && ((ix >> kEidxRigidBit) & 1u) && !(d2 > l2 * 0.9409001 && d2 < l2 * 1.0608999)) rviol = true;
        }
#pragma unroll
        for (int t = 0; t < 4; t++) {
            const unsigned m = __ballot_sync(0xffffffffu, jv[t]);
            if (m) {
                const int at = nj + __popc(m & ((1u << lane) - 1u));
                if (jv[t] && at < kJCap) jlist[at] = (unsigned int)(e0 + 32 * t + lane);
                nj += __popc(m);
            }
        }
    }
    const bool anyj = nj > 0;
    if (anyj) {
        __syncwarp();
        warp_jacobi_limiter(spos, ef2, eidx, etab, csr, Pd, Ed, fixmask, S::kZero, nj <= kJCap ? jlist : nullptr, nj);
    }
    if (n_vis > 0 && (anyj || __any_sync(0xffffffffu, rviol))) {
        // Ordered rigid-cell pass (PhysicsEngine.cpp:675-710).  A visit inside the +-3 % band changes nothing and a
        // visit's outcome depends only on its two end points, so: every lane evaluates its own visit from the current
        // positions (32 visits at once), the first one that moves its points is applied, and only the later visits
        // that share a point with it are evaluated again.  Visits are retired strictly in the reference's order.
        for (int c0 = 0; c0 < n_vis; c0 += 32) {
            const int i = c0 + lane;
            bool done = i >= n_vis, fresh = false, moves = false;
            int a = 0, b = 0;
            double l0 = 1.0, cx = 0.0, cy = 0.0;
            if (!done) {
                const uint32_t ix = eidx[rord[i]];
                a = ix & 0xff; b = (ix >> 8) & 0xff;
                l0 = etab[(ix >> kEidxTypeShift) & 15].y;
            }
            while (true) {
                if (!done && !fresh) { moves = rigid_visit_eval(spos, a, b, l0, cx, cy); fresh = true; }
                const unsigned m = __ballot_sync(0xffffffffu, !done && moves);
                if (!m) break;
                const int f = __ffs(m) - 1;
#ifdef MEVO_PROF
                prof_moves++;
#endif
                if (lane == f) {
                    double2 pa = spos[a], pb = spos[b];
                    pa.x -= cx; pa.y -= cy; pb.x += cx; pb.y += cy;
                    spos[a] = pa; spos[b] = pb;
                }
                __syncwarp();
                const int fa = __shfl_sync(0xffffffffu, a, f), fb = __shfl_sync(0xffffffffu, b, f);
                if (lane <= f) done = true;
                else if (a == fa || a == fb || b == fa || b == fb) fresh = false;
            }
        }
    }
#ifdef MEVO_PROF
    if (lane == 0) {
        const int bk = (anyj ? 8 : 0) + min(prof_moves, 7);
        atomicAdd(&g_whist[2][bk][0], (unsigned long long)(clock64() - tk0)); atomicAdd(&g_whist[2][bk][1], 1ull);
    }
#endif
    for (int p = lane; p < Pd; p += 32) {                // PhysicsEngine::update_true_vel (:602-607)
        const double2 x = spos[p];
        // pos_last == pos at the start of every substep except possibly the first of a launch
        const double2 l = first_substep ? __ldcg(H->g_last + p) : pos[p];
        vt[p] = make_double2(div_const(x.x - l.x, MEVO_DT), div_const(x.y - l.y, MEVO_DT));
        pos[p] = x;
    }
    __syncwarp();
}

// =================================================================================== the kernel
// parameters of the work item a CTA is on (shared memory)
struct WCtl {
    int item;       // CTA-group: environments item * envs_per_cta ... of the class list, one per warp
    int chunk;      // which chunk of the env-step
    int nsub;       // substeps of the item
    int first;      // the item starts an env-step (else it continues one: an environment whose earlier chunk aborted is left alone)
    int last;       // it ends the step (the step is counted)
    int apply;      // actions -> goals at stage in
};

template <int PCAP, int ECAP, int LCAP, int PAIRCAP, int ITEMCAP, int VISCAP, int EPC, int MINB, bool ITEMS>
__device__ __forceinline__ void wstep_item(const StepParams& prm, unsigned char* smem_raw, const volatile WCtl* ctl) {
    using S = WarpSmem<PCAP, ECAP, LCAP, PAIRCAP, ITEMCAP, VISCAP>;
    const int grp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    // A CTA takes prm.envs_per_cta (<= EPC) consecutive entries of the class list; the host picks that number so that
    // the launch is a whole number of equally full waves.  Spare warps only keep the barrier count.
    const int slot = (ITEMS ? ctl->item : (int)blockIdx.x) * prm.envs_per_cta + grp;
    if (grp >= prm.envs_per_cta || slot >= prm.n_class_envs) {                 // spare warps only keep the barrier count
        if (EPC > 1 && prm.cta_sync)
            for (int i = prm.n_steps * (ITEMS ? ctl->nsub : prm.n_substeps); i > 0; i--) __syncthreads();
        return;
    }
    const int env = prm.class_envs[slot];
    if (ITEMS) {
        // an item that continues a step leaves an environment alone whose earlier chunk aborted the step (Sim.cpp:107-110)
        if (!ctl->first && __ldcg(prm.unstable + env)) {
            if (EPC > 1 && prm.cta_sync)
                for (int i = prm.n_steps * ctl->nsub; i > 0; i--) __syncthreads();
            return;
        }
    }
    unsigned char* sm = smem_raw + grp * S::kBytes;
    double* gstate;
    unsigned fixmask = 0;                               // bit k: point lane + 32 k is fixed

    // ---------------------------------------------------------------- stage in (once per launch / item)
    // The design blob holds the constant tables in the form shared memory wants them, the state block holds pos | vel | pos_last |
    // vel_true | goals back to back: lane 0 hands every array to the bulk-copy engine (cp.async.bulk, one mbarrier per warp) and
    // the warp waits once, instead of one dependent L2 round trip per table.  What is computed stays in lanes: actions -> goals.
    {
        const EnvDesc ed = prm.envs[env];
        const uint8_t* db = prm.design + ed.design_off;
        const DesignHeader* Hp = reinterpret_cast<const DesignHeader*>(db);
        const int4 h0 = __ldg(reinterpret_cast<const int4*>(Hp));              // Pd, Ed, n_leaf, n_obj
        const int4 h1 = __ldg(reinterpret_cast<const int4*>(Hp) + 1);          // n_pairs, n_rlev, n_rvis, n_act
        const int Pd = h0.x, Ed = h0.y, n_leaf = h0.z, n_obj = h0.w;
        const int n_pairs = h1.x, n_vis = h1.z;
        const int n_items = Hp->n_items, n_types = Hp->n_types;
        gstate = prm.state + ed.state_off;
        const double2* g2 = reinterpret_cast<const double2*>(gstate);
        const uint32_t bar = smem_u32(sm + S::o_hdr + 88);
        const int apply = ITEMS ? ctl->apply : prm.apply_actions;
        if (lane == 0) {
            mbar_init(bar, 1);
            asm volatile("fence.proxy.async;" ::: "memory");                   // the barrier, and what an earlier chunk's CTA
                                                                               // released through the generic proxy
            auto up16 = [](int n) { return (uint32_t)((n + 15) & ~15); };
            const uint32_t b_pts = 16u * Pd, b_goal = up16(8 * Ed), b_eidx = up16(4 * Ed), b_etab = 16u * n_types;
            const uint32_t b_lpt = up16(4 * n_leaf), b_pairs = up16(2 * n_pairs), b_items = up16(2 * n_items), b_rord = up16(n_vis);
            mbar_expect_tx(bar, 4 * b_pts + b_goal + b_eidx + b_etab + b_lpt + b_pairs + b_items + b_rord);
            if (b_pts) {                                 // (a world of all-FIXED pieces has nothing to simulate)
                bulk_g2s(smem_u32(sm + S::o_pos), g2, b_pts, bar);
                bulk_g2s(smem_u32(sm + S::o_vel), g2 + Pd, b_pts, bar);
                bulk_g2s(smem_u32(sm + S::o_vt), g2 + 3 * Pd, b_pts, bar);
                bulk_g2s(smem_u32(sm + S::o_csr), db + Hp->o_csr8, b_pts, bar);
            }
            if (b_goal) bulk_g2s(smem_u32(sm + S::o_goal), gstate + 8 * (size_t)Pd, b_goal, bar);
            if (b_eidx) bulk_g2s(smem_u32(sm + S::o_eidx), db + Hp->o_eidx, b_eidx, bar);
            if (b_etab) bulk_g2s(smem_u32(sm + S::o_etab), db + Hp->o_etab, b_etab, bar);
            if (b_lpt) bulk_g2s(smem_u32(sm + S::o_lpt), db + Hp->o_lpt8, b_lpt, bar);
            if (b_pairs) bulk_g2s(smem_u32(sm + S::o_pairs), db + Hp->o_pairs16, b_pairs, bar);
            if (b_items) bulk_g2s(smem_u32(sm + S::o_items), db + Hp->o_items16, b_items, bar);
            if (b_rord) bulk_g2s(smem_u32(sm + S::o_rord), db + Hp->o_rord8, b_rord, bar);
            WarpHdr* H = reinterpret_cast<WarpHdr*>(sm + S::o_hdr);
            H->Pd = Pd; H->Ed = Ed; H->n_leaf = n_leaf; H->n_pairs = n_pairs; H->n_items = n_items; H->n_vis = n_vis;
            H->g_leaf = reinterpret_cast<const LeafRec*>(db + Hp->o_leaf);
            H->g_last = g2 + 2 * Pd;
            if (ed.static_off != ~0ull) {
                const uint8_t* sb = prm.statics + ed.static_off;
                const StaticHeader* SHp = reinterpret_cast<const StaticHeader*>(sb);
                H->gx0 = SHp->gx0; H->gy0 = SHp->gy0; H->gw = SHp->gw; H->gh = SHp->gh;
                H->st_grid = reinterpret_cast<const int16_t*>(sb + SHp->o_grid);
                H->st_aabb = reinterpret_cast<const double4*>(sb + SHp->o_aabb);
                H->st_pos = reinterpret_cast<const double2*>(sb + SHp->o_pos);
                H->st_leaf = reinterpret_cast<const LeafRec*>(sb + SHp->o_leaf);
            } else {
                H->gx0 = H->gy0 = H->gw = H->gh = 0;
                H->st_grid = nullptr; H->st_aabb = nullptr; H->st_pos = nullptr; H->st_leaf = nullptr;
            }
            reinterpret_cast<double2*>(sm + S::o_ef2)[S::kZero] = make_double2(0.0, 0.0);
        }
        fixmask = ((Hp->fixed_lo >> lane) & 1u) | (((Hp->fixed_hi >> lane) & 1u) << 1);
        volatile int* misc = reinterpret_cast<volatile int*>(sm + S::o_misc);
        if (lane < 8) {
            int ns = 0x7fffffff;                        // never "unstable" unless a robot (PhysicsEngine.cpp:819-834)
            if (lane < n_obj) {
                const ObjRec orc = reinterpret_cast<const ObjRec*>(db + Hp->o_obj)[lane];
                if (orc.is_robot) ns = orc.n_surface;
            }
            misc[8 + lane] = ns;
        }
        // K1: PhysicsEngine::update_actuator_goals (PhysicsEngine.cpp:38-97) as a per-edge gather, loads in flight under the copies
        constexpr int kEPL = (ECAP + 31) / 32;
        double gnew[kEPL];
        unsigned gset = 0;
        if (apply) {
            const int16_t* eact = reinterpret_cast<const int16_t*>(db + Hp->o_eact);
            const double* act = prm.act + (size_t)env * prm.act_stride;
#pragma unroll
            for (int q = 0; q < kEPL; q++) {
                const int e = lane + 32 * q;
                gnew[q] = 0.0;
                if (e < Ed) {
                    const int a0 = eact[2 * e], a1 = eact[2 * e + 1];
                    if (a0 >= 0) {
                        const double l0 = __ldg(reinterpret_cast<const double*>(db + Hp->o_el0) + e);
                        double sum = 0.0 + act[a0];
                        int cnt = 1;
                        if (a1 >= 0) { sum += act[a1]; cnt = 2; }
                        gnew[q] = sum * (l0 / cnt);
                        gset |= 1u << q;
                    }
                }
            }
        }
        __syncwarp();                                   // lane 0 has initialised the barrier
        mbar_wait(bar, 0);
        // empty gather slots (0xffff) point at the all-zero entry of ef2 from here on
        uint4* csr = reinterpret_cast<uint4*>(sm + S::o_csr);
        for (int i = lane; i < Pd; i += 32) {
            uint4 rec = csr[i];
            unsigned* w = reinterpret_cast<unsigned*>(&rec);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                if ((w[q] & 0xffffu) == 0xffffu) w[q] = (w[q] & 0xffff0000u) | (unsigned)S::kZero;
                if ((w[q] >> 16) == 0xffffu) w[q] = (w[q] & 0xffffu) | ((unsigned)S::kZero << 16);
            }
            csr[i] = rec;
        }
        if (apply) {
            double* goal = reinterpret_cast<double*>(sm + S::o_goal);
#pragma unroll
            for (int q = 0; q < kEPL; q++)
                if ((gset >> q) & 1u) goal[lane + 32 * q] = gnew[q];
        }
    }
    __syncwarp();

    int unstable = 0;
    int completed_steps = 0;
    bool first_substep = true;
    const bool dbg = prm.dbg_cand != nullptr || prm.dbg_cf != nullptr;

    for (int step = 0; step < prm.n_steps; step++) {
        for (int sub = 0; sub < (ITEMS ? ctl->nsub : prm.n_substeps); sub++) {
            // keeps the CTA's warps in the same code region: they share instruction-cache lines
            if (EPC > 1 && prm.cta_sync) __syncthreads();
            if (unstable) continue;                       // Sim::step aborted (Sim.cpp:107-110)
            const bool dump = dbg && step == 0 && sub == 0;
#ifdef MEVO_PROF
            const long long tq0 = clock64();
#endif
            const int sub_unstable = warp_collide<S>(sm, (dump && prm.dbg_cand) ? prm.dbg_cand + (size_t)env * prm.dbg_cand_cap : nullptr,
                                                     (dump && prm.dbg_cand) ? prm.dbg_cand_n + env : nullptr, prm.dbg_cand_cap,
                                                     (dump && prm.dbg_cf) ? reinterpret_cast<double2*>(prm.dbg_cf + (gstate - prm.state)) : nullptr);
#ifdef MEVO_PROF
            const long long tq1 = clock64();
#endif
            warp_rk4<S>(sm, fixmask);
#ifdef MEVO_PROF
            const long long tq2 = clock64();
#endif
            warp_constrain<S>(sm, fixmask, first_substep);
#ifdef MEVO_PROF
            {
                __shared__ unsigned int pmax[4];
                const long long tq3 = clock64();
                const unsigned int d[4] = {(unsigned)(tq1 - tq0), (unsigned)(tq2 - tq1), (unsigned)(tq3 - tq2), (unsigned)(tq3 - tq0)};
                if (threadIdx.x < 4) pmax[threadIdx.x] = 0;
                __syncthreads();
                if (lane == 0) for (int q = 0; q < 4; q++) { atomicMax(&pmax[q], d[q]); atomicAdd(&g_wprof[q], (unsigned long long)d[q]); }
                __syncthreads();
                if (threadIdx.x < 4) atomicAdd(&g_wprof[4 + threadIdx.x], (unsigned long long)pmax[threadIdx.x] * prm.envs_per_cta);
                __syncthreads();
            }
#endif
            first_substep = false;
            if (sub_unstable) unstable = 1;               // Sim::step aborts the remaining substeps (Sim.cpp:107-110)
        }
        if (!unstable) completed_steps++;
    }

    // ---------------------------------------------------------------- stage out
    {
        const WarpHdr* H = reinterpret_cast<const WarpHdr*>(sm + S::o_hdr);
        const int Pd = H->Pd, Ed = H->Ed;
        double2* o2 = reinterpret_cast<double2*>(gstate);
        if (unstable && prm.auto_reset) {
            // Sim::revert(0) in place (Environment.cpp:247-263)
            const double2* s0 = reinterpret_cast<const double2*>(prm.state0 + (gstate - prm.state));
            for (int i = lane; i < Pd; i += 32) {
                const double2 p0 = s0[i], v0 = s0[Pd + i];
                o2[i] = p0; o2[Pd + i] = v0; o2[2 * Pd + i] = p0; o2[3 * Pd + i] = v0;
            }
        } else if (!first_substep) {
            const double2* pos = reinterpret_cast<const double2*>(sm + S::o_pos);
            const double2* vel = reinterpret_cast<const double2*>(sm + S::o_vel);
            const double2* vt = reinterpret_cast<const double2*>(sm + S::o_vt);
            for (int i = lane; i < Pd; i += 32) {
                const double2 x = pos[i];
                o2[i] = x;
                o2[Pd + i] = vel[i];
                o2[2 * Pd + i] = x;                        // pos_last = pos after update_true_vel
                o2[3 * Pd + i] = vt[i];
            }
        }
        const double* goal = reinterpret_cast<const double*>(sm + S::o_goal);
        double* ggoal = gstate + 8 * (size_t)Pd;
        for (int e = lane; e < Ed; e += 32) ggoal[e] = goal[e];
        if (lane == 0) {
            prm.unstable[env] = (uint8_t)unstable;
            if (unstable && prm.auto_reset) prm.sim_time[env] = 0;
            else if (ITEMS ? ctl->last : prm.chunk_last) prm.sim_time[env] += completed_steps;
            if (unstable) atomicAdd(prm.unstable_events, 1ull);
        }
    }
}

// The kernel.  ITEMS == false: CTA b takes CTA-group b for a whole launch's worth (prm.n_substeps substeps, chunk flags of prm).
// ITEMS == true: the grid has one CTA per (CTA-group, chunk) item -- chunks being consecutive substep ranges of ONE env-step --
// and a CTA DRAWS its item from a counter when it starts: item i = chunk i / n_groups of group i % n_groups.  A chunk waits until
// the group's previous chunk (drawn n_groups items earlier, by a CTA that is running or done -- so the wait cannot deadlock) has
// published its state.  The hardware's CTA scheduler is the work queue: an SM takes the next item the moment a CTA retires.
// With 4096 walkers on 148 SMs a single launch is 1.73 waves of 16-warp CTAs (or two waves of 14-warp CTAs); four chunks make
// that 6.92 rounds of a quarter step.  The item's parameters sit in shared memory (WCtl): the body is at its register cap and
// reads them where it needs them.  (A persistent loop around the body was measured 3 % slower: what the loop keeps live.)
template <int PCAP, int ECAP, int LCAP, int PAIRCAP, int ITEMCAP, int VISCAP, int EPC, int MINB, bool ITEMS>
__global__ void __launch_bounds__(32 * EPC, MINB) wstep_kernel(const StepParams prm) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    if (!ITEMS) {
        wstep_item<PCAP, ECAP, LCAP, PAIRCAP, ITEMCAP, VISCAP, EPC, MINB, false>(prm, smem_raw, nullptr);
        return;
    }
    __shared__ WCtl ctl;
    if (threadIdx.x == 0) {
        const int n_groups = (prm.n_class_envs + prm.envs_per_cta - 1) / prm.envs_per_cta;
        const int i = atomicAdd(prm.work, 1);
        const int k = i / n_groups, g = i - k * n_groups;
        if (k > 0) {
            const int* done = prm.work + 1 + g;
            for (;;) {
                int v;
                asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(done) : "memory");
                if (v >= k) break;
                __nanosleep(256);
            }
        }
        const int s0 = prm.n_substeps * k / prm.n_chunks, s1 = prm.n_substeps * (k + 1) / prm.n_chunks;
        ctl.item = g;
        ctl.chunk = k;
        ctl.nsub = s1 - s0;
        ctl.first = k == 0;
        ctl.last = k == prm.n_chunks - 1;
        ctl.apply = prm.apply_actions && k == 0;
    }
    __syncthreads();
    wstep_item<PCAP, ECAP, LCAP, PAIRCAP, ITEMCAP, VISCAP, EPC, MINB, true>(prm, smem_raw, &ctl);
    __syncthreads();                                           // every warp has staged out
    if (threadIdx.x == 0) {
        __threadfence();
        asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(prm.work + 1 + ctl.item), "r"(ctl.chunk + 1) : "memory");
    }
}

}  // namespace mevo
