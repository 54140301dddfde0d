// Group-per-environment step kernel for everything larger than the warp kernel's capacity classes: 7x7 ... 20x20 robots,
// simulated (non-FIXED) terrain, several free objects -- BASELINE.json configs[3] and configs[4].
//
// One CTA of G = 64 ... 512 threads owns one environment for a whole launch (n_steps x 30 substeps of
// Sim::step, reference SimulatorCPP/Sim.cpp:105-115 -> Environment::step, Environment.cpp:166-205).  Same arithmetic,
// statement for statement, as wstep_kernel (wkernel.cuh); the exactness contract is in kernels.cuh / DESIGN.md.
//
// What bounds this path on B200 (ncu, profiles/r02k_gstep256_*; round 1: r01h_step512_*): not HBM (the state is on chip) and not the FP64 pipe
// (11 % busy) but a long chain of short dependent phases separated by CTA barriers, run by too few warps to hide their
// latency.  The layout below is built around that:
//   * shared memory holds only what is gathered across threads every stage (positions, stage positions, contact forces,
//     per-edge forces, length_eq, boxes) plus compact tables (one u32 per edge, one 16-byte gather record per point, 8 bytes
//     per leaf): a 20x20 robot needs ~90 KB instead of 187 KB, so TWO environments of 256 threads are resident per SM (four
//     of 128 threads for 15x15, eight of 64 for 10x10) and one's serial phases (ordered rigid-cell pass, contact rounds)
//     overlap the others' parallel ones.  What only its owner thread touches stays out of shared memory: velocities are
//     read from / written to the state block in HBM once per substep and live in registers in between, like the RK4
//     accumulators and stage velocities; true velocities (read by the contact resolver only) and length_eq_goal (read once
//     per substep) stay in the state block too;
//   * broad phase by temporal coherence: the all-pairs leaf table (23 k pairs on a 20x20 robot) is scanned once per
//     launch -- and again only when points have moved more than half the margin since -- into a per-warp neighbour list
//     of the pairs whose boxes are within kListMargin of each other; every substep tests only the listed pairs exactly.
//     The candidate SET is still exactly "all leaf pairs whose boxes overlap" (the list is a superset by construction, the
//     displacement bound is checked every substep), which is what the reference's tree descent produces;
//   * candidate and survivor buffers are drained in rounds: no configuration can overflow them;
//   * contacts are applied by all threads at once, in rounds: a contact is applied when it is the lowest-numbered
//     unapplied contact on each of its points, so every point still receives its contacts in list order (deterministic,
//     same sums as a serial walk) without one barrier per warp;
//   * U edges per thread in flight in the spring / limiter passes (independent sqrt / div chains);
//   * CTA-wide flags travel on barriers that exist anyway (cta_or: three rotating slots, one barrier per use): a
//     contact-free substep of a robot inside its strain limits takes 12 barriers;
//   * a launch may cover only a range of an env-step's substeps (StepParams::chunk_first / chunk_last): the host issues a step
//     as several such launches on two streams per class where one launch would leave a partly filled last wave of CTAs
//     (Batch::step_async).
#pragma once

#include "kernels.cuh"

namespace mevo {

constexpr double kListMargin = 0.03;     // neighbour-list margin [m]; a voxel is 0.1
constexpr double kListSlack = 1e-6;      // >> rounding of the box arithmetic (positions are O(10), ulp ~ 1e-15)
constexpr int kGMiscInts = 96;
// misc words: [0..7] self-collision counts per object (a-12), [8..10] rotating slots of cta_or, [12..13] largest point
// displacement of a substep (high word of the double, by substep parity), [16..31] per-warp neighbour-list lengths
// (-1: list overflowed, this warp scans its table range), [32..63] per-warp survivor counts (two round parities),
// [64] candidate-dump cursor, [66..81] per-warp item-list lengths, [84..87] development counters (-DMEVO_PROF)
constexpr int kGmSelf = 0, kGmOr = 8, kGmDisp = 12, kGmList = 16, kGmSurv = 32, kGmDump = 64, kGmIList = 66, kGmProf = 84;

constexpr int kEidxBShift = 10, kEidxTypeShift10 = 20, kEidxRigidBit10 = 26;

constexpr int kGHdrBytes = 320;
struct GHdr {                                 // first kGHdrBytes of the CTA's shared memory
    int Pd, Ed, n_leaf, n_pairs, n_items, n_rlev, n_obj, seg_cap;
    int lseg_cap, iseg_cap, gx0, gy0, gw, gh, env, pad0;
    int o_pos, o_spos, o_cf, o_ef, o_leq, o_aabb, o_csr, o_etab;
    int o_btab, o_eidx, o_lpt, o_rsched, o_rlev, o_cand, o_surv, o_misc;
    const int16_t* st_grid;
    const double4* st_aabb;
    const double2* st_pos;
    const LeafRec* st_leaf;
    const LeafRec* g_leaf;                    // full dynamic leaf records (resolver only)
    const double2* g_last;                    // pos_last in HBM: read in the first substep of a launch only
    const unsigned int* pairs;
    const unsigned int* items;
    const ObjRec* objs;
    const double* ggoal;                      // length_eq_goal of this environment (state block)
    const uint4* csrx;                        // nullptr, or gather slots 8..15 per point (design blob; 0xffff = empty)
    unsigned int* list;                       // neighbour list scratch of this environment: tested pairs ...
    unsigned int* ilist;                      // ... and always-candidate items
    double* gstate;                           // state block of this environment
    const double* gstate0;                    // its reset snapshot
};
static_assert(sizeof(GHdr) <= kGHdrBytes, "GHdr must fit its slot");

struct GLayout {
    int o_pos, o_spos, o_cf, o_ef, o_leq, o_aabb, o_csr, o_etab, o_btab, o_eidx, o_lpt, o_rsched, o_rlev, o_cand, o_surv, o_misc, total;
};

MEVO_HD GLayout g_layout(int Pd, int Ed, int n_leaf, int n_types, int cand_cap, int n_rvis, int n_rlev) {
    GLayout L;
    int o = kGHdrBytes;
    L.o_pos = o;   o += 16 * Pd;
    L.o_spos = o;  o += 16 * Pd;
    L.o_cf = o;    o += 16 * Pd;
    L.o_ef = o;    o += 16 * (Ed + 1);              // [Ed] stays zero: empty gather slots point at it.  The first 4 * Pd bytes
                                                    // double as the contact-claim words while contacts are applied (ef is dead then)
    L.o_leq = o;   o += align16(8 * Ed);
    L.o_aabb = o;  o += 32 * n_leaf;
    L.o_csr = o;   o += 16 * Pd;
    L.o_etab = o;  o += 16 * n_types;               // (spring constant, init_length_eq) per edge type
    L.o_btab = o;  o += 32 * n_types;               // squared-length bands per edge type: Jacobi lo, hi; rigid lo, hi
    L.o_eidx = o;  o += align16(4 * Ed);
    L.o_lpt = o;   o += align16(8 * n_leaf);
    L.o_rsched = o; o += align16(4 * n_rvis);       // packed edge word of every rigid-cell visit, level order
    L.o_rlev = o;  o += align16(2 * (n_rlev + 3));  // level offsets into rsched (+ two padding entries for the prefetch)
    L.o_cand = o;  o += align16(4 * cand_cap);
    L.o_surv = o;  o += align16(4 * cand_cap);
    L.o_misc = o;  o += 4 * kGMiscInts;
    L.total = align16(o);
    return L;
}

extern __shared__ __align__(16) unsigned char gsm_raw[];

// CTA-wide OR of a per-thread bit mask on ONE barrier.  Three slots rotate with the (uniform) call counter r: call r
// ORs into slot r % 3 and clears slot (r + 1) % 3, which was last read after the barrier of call r - 2 -- and every thread
// has passed the barrier of call r - 1 since.  Whatever else the caller published to shared memory before the call is
// visible after it.
__device__ __forceinline__ int cta_or(volatile int* slots, int bits, int& r) {
    const int s = r % 3;
    const unsigned w = __reduce_or_sync(0xffffffffu, (unsigned)bits);
    if ((threadIdx.x & 31) == 0 && w) atomicOr((int*)&slots[s], (int)w);
    if (threadIdx.x == 0) slots[(s + 1) % 3] = 0;
    __syncthreads();
    r++;
    return slots[s];
}

// BoundingBox::in (main.h:64-72) with box a grown by m on every side
__device__ __forceinline__ bool aabb_overlap_margin(const double4& a, const double4& b, double m) {
    return !(a.x - m > b.z) & !(a.z + m < b.x) & !(a.y - m > b.w) & !(a.w + m < b.y);
}

// a point inside box b grown by m on every side (necessary for point_in_box after either has moved by <= m / 2)
__device__ __forceinline__ bool point_in_box_margin(const double2 p, const double4& b, double m) {
    return (p.x >= b.x - m) & (p.x <= b.z + m) & (p.y >= b.y - m) & (p.y <= b.w + m);
}

__device__ __forceinline__ unsigned leaf_corner(const uint2 lp, unsigned corner) {      // TL, TR, BR, BL point id of a leaf
    return (((corner & 2u) ? lp.y : lp.x) >> ((corner & 1u) * 16)) & 0xffffu;
}

__device__ __forceinline__ double2 spring_force10(const double2* sp, uint32_t ix, double k, double& q, bool first_stage) {
    const double2 pa = sp[ix & 0x3ff], pb = sp[(ix >> kEidxBShift) & 0x3ff];
    const double dx = pa.x - pb.x, dy = pa.y - pb.y;
    const double dist = sqrt(dx * dx + dy * dy);
    if (first_stage) {                                  // PhysicsEngine::update_actuators (PhysicsEngine.cpp:99-119)
        const double add = (q - dist) * 0.006;
        q = dist + add;
    }
    const double mag = (dist - q) * k / dist;           // PhysicsEngine.cpp:498-528
    return make_double2(dx * mag, dy * mag);
}

__device__ __forceinline__ double flip_if(double x, unsigned neg) {      // neg ? -x : x, on the integer pipe (exact)
    return __hiloint2double(__double2hiint(x) ^ (int)(neg << 31), __double2loint(x));
}

// Gather slots 8..15 of point p (designs with unconnected neighbouring voxels: a side edge per voxel, up to 12 springs on
// a point).  Continues the a-sides-then-b-sides list of the first record in the same order; read from the design blob.
__device__ __noinline__ double2 g_gather_ext(const uint4* csrx, const double2* ef, int p, double2 acc) {
    const uint4 rec = __ldg(csrx + p);
    const unsigned w[4] = {rec.x, rec.y, rec.z, rec.w};
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const unsigned ent = (w[j >> 1] >> ((j & 1) * 16)) & 0xffffu;
        if (ent == 0xffffu) break;                                      // the list is dense: nothing follows an empty slot
        const double2 c = ef[ent >> 1];
        const unsigned neg = (ent & 1u) ^ 1u;
        acc.x += flip_if(c.x, neg); acc.y += flip_if(c.y, neg);
    }
    return acc;
}

#ifdef MEVO_PROF
// development only (-DMEVO_PROF): cycles of thread 0 of every CTA by phase
// [0] boxes + list upkeep [1] broad phase + filter [2] resolve + contacts [3] RK4 [4] band test [5] Jacobi [6] rigid pass [7] true velocity
__device__ unsigned long long g_g2prof[8];
#define MEVO_GPROF(k)                                                                             \
    do {                                                                                          \
        if (threadIdx.x == 0) {                                                                   \
            long long* _st = reinterpret_cast<long long*>(gsm_raw + reinterpret_cast<const GHdr*>(gsm_raw)->o_misc + 4 * kGmProf); \
            const long long _n = clock64();                                                       \
            atomicAdd(&g_g2prof[(int)_st[1]], (unsigned long long)(_n - _st[0]));                 \
            _st[0] = _n; _st[1] = (k);                                                            \
        }                                                                                         \
    } while (0)
// [0] substeps [1] list rebuilds [2] survivors [3] contact batches [4] contact rounds [5] substeps with Jacobi [6] substeps with
// rigid pass [7] rigid visits that moved their points
__device__ unsigned long long g_g2cnt[8];
#define MEVO_GCOUNT(k, v) do { if (threadIdx.x == 0) atomicAdd(&g_g2cnt[k], (unsigned long long)(v)); } while (0)
#define MEVO_GCOUNT_ANY(k, v) atomicAdd(&g_g2cnt[k], (unsigned long long)(v))
#else
#define MEVO_GPROF(k)
#define MEVO_GCOUNT(k, v)
#define MEVO_GCOUNT_ANY(k, v)
#endif

// =================================================================================== phase 1: collisions
// K2a leaf boxes, neighbour list upkeep, K2b broad phase, K3 narrow phase + contact response, K6 unstable count
// (SURVEY 8a-4,5,6,12).  Returns (r << 1) | unstable, unstable as PhysicsEngine::is_any_robot_self_colliding
// (boxes of the START of the substep).
// (MINB is part of every phase's signature only so that each build of the kernel gets its own copy, compiled to its own
// register budget: ptxas compiles a shared out-of-line function once, to the tightest budget of its callers.)
template <int G, int MINB>
__device__ __noinline__ int g_collide(int rebuild, int r, unsigned int* dbg_row, int* dbg_n, int dbg_cap, double2* dbg_cf) {
    constexpr int NW = G / 32;
    unsigned char* sm = gsm_raw;
    const GHdr* H = reinterpret_cast<const GHdr*>(sm);
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const double2* pos = reinterpret_cast<const double2*>(sm + H->o_pos);
    const double2* vt = reinterpret_cast<const double2*>(H->gstate) + 3 * H->Pd;   // vel_true: state block (written by update_true_vel)
    double2* cf = reinterpret_cast<double2*>(sm + H->o_cf);
    double4* aabb = reinterpret_cast<double4*>(sm + H->o_aabb);
    const uint2* lpt = reinterpret_cast<const uint2*>(sm + H->o_lpt);
    int* claim = reinterpret_cast<int*>(sm + H->o_ef);        // aliases ef, which is dead between a substep's start and its RK4
    volatile int* misc = reinterpret_cast<volatile int*>(sm + H->o_misc);
    const int Pd = H->Pd, n_leaf = H->n_leaf, n_pairs = H->n_pairs, n_items = H->n_items, seg_cap = H->seg_cap;
    unsigned int* cand = reinterpret_cast<unsigned int*>(sm + H->o_cand) + warp * seg_cap;
    unsigned int* surv_all = reinterpret_cast<unsigned int*>(sm + H->o_surv);
    unsigned int* surv = surv_all + warp * seg_cap;

    MEVO_GPROF(0);
    MEVO_GCOUNT(0, 1);
    MEVO_GCOUNT(1, rebuild);
    // ---- K2a: leaf boxes (a-4), zero collision_forces
    for (int l = t; l < n_leaf; l += G) {
        const uint2 lp = lpt[l];
        double2 q = pos[lp.x & 0xffffu];
        double minx = q.x, maxx = q.x, miny = q.y, maxy = q.y;
        q = pos[lp.x >> 16];
        minx = fmin(minx, q.x); maxx = fmax(maxx, q.x); miny = fmin(miny, q.y); maxy = fmax(maxy, q.y);
        q = pos[lp.y & 0xffffu];
        minx = fmin(minx, q.x); maxx = fmax(maxx, q.x); miny = fmin(miny, q.y); maxy = fmax(maxy, q.y);
        q = pos[lp.y >> 16];
        minx = fmin(minx, q.x); maxx = fmax(maxx, q.x); miny = fmin(miny, q.y); maxy = fmax(maxy, q.y);
        aabb[l] = make_double4(minx, miny, maxx, maxy);
    }
    for (int i = t; i < Pd; i += G) { cf[i] = make_double2(0.0, 0.0); claim[i] = 0x7fffffff; }
    if (t < 8) misc[kGmSelf + t] = 0;
    __syncthreads();

    // ---- neighbour list of this warp's range of the pair table (see the header comment)
    const int R = ((n_pairs + NW * 32 - 1) / (NW * 32)) * 32;
    const int p0 = min(n_pairs, warp * R), p1 = min(n_pairs, p0 + R);
    const unsigned int* pairs = H->pairs;
    unsigned int* lseg = H->list + (size_t)warp * H->lseg_cap;
    if (rebuild) {
        int lcnt = 0;
        bool ok = true;
        for (int base = p0; base < p1; base += 64) {                     // two pairs per lane in flight
            const int i0 = base + lane, i1 = i0 + 32;
            const unsigned c0 = __ldg(pairs + (i0 < p1 ? i0 : base)), c1 = __ldg(pairs + (i1 < p1 ? i1 : base));
            const bool h0 = (i0 < p1) & aabb_overlap_margin(aabb[c0 & kCodeLeafMask], aabb[(c0 >> 12) & kCodeLeafMask], kListMargin);
            const bool h1 = (i1 < p1) & aabb_overlap_margin(aabb[c1 & kCodeLeafMask], aabb[(c1 >> 12) & kCodeLeafMask], kListMargin);
            const unsigned m0 = __ballot_sync(0xffffffffu, h0), m1 = __ballot_sync(0xffffffffu, h1);
            if (m0 | m1) {
                const int n0 = __popc(m0), at0 = lcnt + __popc(m0 & ((1u << lane) - 1u)), at1 = lcnt + n0 + __popc(m1 & ((1u << lane) - 1u));
                if (lcnt + n0 + __popc(m1) <= H->lseg_cap) {
                    if (h0) lseg[at0] = c0;
                    if (h1) lseg[at1] = c1;
                    lcnt += n0 + __popc(m1);
                } else {
                    ok = false;
                }
            }
        }
        if (lane == 0) misc[kGmList + warp] = ok ? lcnt : -1;
        // always-candidate items (non-shared corners of leaf pairs that share a point) whose corner is within the margin of
        // the other leaf's box; this warp's chunks of the item table can never exceed its list segment
        unsigned int* iseg = H->ilist + (size_t)warp * H->iseg_cap;
        int icnt = 0;
        for (int c = warp; c * 32 < n_items; c += 2 * NW) {              // two chunks in flight
            const int i0 = c * 32 + lane, i1 = i0 + NW * 32;
            const unsigned ic0 = __ldg(H->items + (i0 < n_items ? i0 : 0)), ic1 = __ldg(H->items + (i1 < n_items ? i1 : 0));
            const unsigned q0 = leaf_corner(lpt[ic0 & kCodeLeafMask], (ic0 >> kItemCornerShift) & 3u);
            const unsigned q1 = leaf_corner(lpt[ic1 & kCodeLeafMask], (ic1 >> kItemCornerShift) & 3u);
            const bool h0 = (i0 < n_items) & point_in_box_margin(pos[q0], aabb[(ic0 >> 12) & kCodeLeafMask], kListMargin);
            const bool h1 = (i1 < n_items) & point_in_box_margin(pos[q1], aabb[(ic1 >> 12) & kCodeLeafMask], kListMargin);
            const unsigned m0 = __ballot_sync(0xffffffffu, h0), m1 = __ballot_sync(0xffffffffu, h1);
            if (h0) iseg[icnt + __popc(m0 & ((1u << lane) - 1u))] = ic0;
            icnt += __popc(m0);
            if (h1) iseg[icnt + __popc(m1 & ((1u << lane) - 1u))] = ic1;
            icnt += __popc(m1);
        }
        if (lane == 0) misc[kGmIList + warp] = icnt;
        __syncwarp();                                                    // list entries are read by other lanes of this warp
    }
    const int lcnt = misc[kGmList + warp];
    const unsigned int* psrc = lcnt >= 0 ? lseg : pairs + p0;             // overflowed list: the table range itself, every substep
    const int pn = lcnt >= 0 ? lcnt : p1 - p0;

    MEVO_GPROF(1);
    // ---- K2b + K3 in rounds: produce candidates into this warp's segment, filter their corners into survivors, resolve
    const int gw = H->gw;
    int pi = 0;                                   // cursor in psrc
    int sc = warp, sit = 0, max_cells = -1;       // terrain: leaf chunk (leaves sc * 32 + lane), cell iteration
    int cx0 = 0, cy0 = 0, wx = 1, n_cells = 0;
    double4 mybb = make_double4(0, 0, 0, 0);
    const unsigned int* iseg = H->ilist + (size_t)warp * H->iseg_cap;
    const int icnt = misc[kGmIList + warp];
    int ii = 0;                                   // cursor in this warp's item list
    int rc = 0;                                   // survivor round counter (uniform)
    int flags;
    do {
        int cnt = 0;
        while (pi < pn && cnt + 64 <= seg_cap) {                          // two pairs per lane in flight
            const int i0 = pi + lane, i1 = i0 + 32;
            const unsigned c0 = psrc[i0 < pn ? i0 : pi], c1 = psrc[i1 < pn ? i1 : pi];
            pi += 64;
            const unsigned a0 = c0 & kCodeLeafMask, b0 = (c0 >> 12) & kCodeLeafMask, a1 = c1 & kCodeLeafMask, b1 = (c1 >> 12) & kCodeLeafMask;
            const bool h0 = (i0 < pn) & aabb_overlap(aabb[a0], aabb[b0]);
            const bool h1 = (i1 < pn) & aabb_overlap(aabb[a1], aabb[b1]);
            const unsigned m0 = __ballot_sync(0xffffffffu, h0), m1 = __ballot_sync(0xffffffffu, h1);
            if (m0 | m1) {
                // ordered pairs (a,b) and (b,a) both count (a-12, PhysicsEngine.cpp:842-889)
                if (h0 && (c0 & kCodeCount)) atomicAdd((int*)&misc[kGmSelf + (c0 >> kCodeObjShift)], 2);
                if (h1 && (c1 & kCodeCount)) atomicAdd((int*)&misc[kGmSelf + (c1 >> kCodeObjShift)], 2);
                if (h0) cand[cnt + __popc(m0 & ((1u << lane) - 1u))] = c0 & (kCodeLeafMask | (kCodeLeafMask << 12) | kCodeSameObj);
                cnt += __popc(m0);
                if (h1) cand[cnt + __popc(m1 & ((1u << lane) - 1u))] = c1 & (kCodeLeafMask | (kCodeLeafMask << 12) | kCodeSameObj);
                cnt += __popc(m1);
            }
        }
        if (pi >= pn && gw > 0) {
            // dynamic leaves vs the static voxel grid (cell = one voxel): window widened by 1e-6 cells, exact test decides
            while (sc * 32 < n_leaf) {
                if (max_cells < 0) {
                    const int l = sc * 32 + lane;
                    n_cells = 0; wx = 1;
                    if (l < n_leaf) {
                        mybb = aabb[l];
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
                    sit = 0;
                }
                while (sit < max_cells && cnt + 64 <= seg_cap) {           // two cells per lane in flight
                    const int it0 = sit, it1 = sit + 1;
                    sit += 2;
                    int s0 = -1, s1 = -1;
                    if (it0 < n_cells) s0 = __ldg(H->st_grid + (cy0 + it0 / wx) * gw + cx0 + it0 % wx);
                    if (it1 < n_cells) s1 = __ldg(H->st_grid + (cy0 + it1 / wx) * gw + cx0 + it1 % wx);
                    bool h0 = false, h1 = false;
                    if (s0 >= 0) h0 = aabb_overlap(mybb, ldg_d4(H->st_aabb + s0));
                    if (s1 >= 0) h1 = aabb_overlap(mybb, ldg_d4(H->st_aabb + s1));
                    const unsigned m0 = __ballot_sync(0xffffffffu, h0), m1 = __ballot_sync(0xffffffffu, h1);
                    if (m0 | m1) {
                        const unsigned me = (unsigned)(sc * 32 + lane) | kCodeStatic;
                        if (h0) cand[cnt + __popc(m0 & ((1u << lane) - 1u))] = me | ((unsigned)s0 << 12);
                        cnt += __popc(m0);
                        if (h1) cand[cnt + __popc(m1 & ((1u << lane) - 1u))] = me | ((unsigned)s1 << 12);
                        cnt += __popc(m1);
                    }
                }
                if (sit < max_cells) break;                               // segment full: next round
                sc += NW; max_cells = -1;
            }
        }
        const bool prod_done = pi >= pn && (gw <= 0 || sc * 32 >= n_leaf);
        __syncwarp();
        if (dbg_row) {                                                    // test hook: candidate SET of this substep
            int off = 0;
            if (lane == 0) off = atomicAdd((int*)&misc[kGmDump], cnt);
            off = __shfl_sync(0xffffffffu, off, 0);
            for (int i = lane; i < cnt; i += 32)
                if (off + i < dbg_cap) dbg_row[off + i] = cand[i];
        }
        // ---- consume: box prefilter of every (corner, boxel) item -- a necessary condition of Boxel::point_in_boxel --
        //      into this warp's survivor segment, in rounds
        int next_cc = 0;
        do {
            int scnt = 0;
            // listed always-candidate items (see the list upkeep above), two chunks in flight
            while (ii < icnt && scnt + 64 <= seg_cap) {
                const int i0 = ii + lane, i1 = i0 + 32;
                const unsigned ic0 = iseg[i0 < icnt ? i0 : ii], ic1 = iseg[i1 < icnt ? i1 : ii];
                ii += 64;
                const unsigned q0 = leaf_corner(lpt[ic0 & kCodeLeafMask], (ic0 >> kItemCornerShift) & 3u);
                const unsigned q1 = leaf_corner(lpt[ic1 & kCodeLeafMask], (ic1 >> kItemCornerShift) & 3u);
                const bool h0 = (i0 < icnt) & point_in_box(pos[q0], aabb[(ic0 >> 12) & kCodeLeafMask]);
                const bool h1 = (i1 < icnt) & point_in_box(pos[q1], aabb[(ic1 >> 12) & kCodeLeafMask]);
                const unsigned m0 = __ballot_sync(0xffffffffu, h0), m1 = __ballot_sync(0xffffffffu, h1);
                if (m0 | m1) {
                    if (h0) surv[scnt + __popc(m0 & ((1u << lane) - 1u))] = ic0;
                    scnt += __popc(m0);
                    if (h1) surv[scnt + __popc(m1 & ((1u << lane) - 1u))] = ic1;
                    scnt += __popc(m1);
                }
            }
            const bool items_done = ii >= icnt;
            if (items_done) {
                // ... then the 8 corners of each candidate pair of this round (these pairs share no point)
                while (next_cc * 32 < cnt * 8 && scnt + 32 <= seg_cap) {
                    const int i = next_cc * 32 + lane;
                    next_cc++;
                    bool hit = false;
                    unsigned int code = 0;
                    if (i < cnt * 8) {
                        const unsigned int cc = cand[i >> 3];
                        const unsigned a = cc & kCodeLeafMask, b = (cc >> 12) & kCodeLeafMask;
                        const bool b_static = (cc & kCodeStatic) != 0;
                        const unsigned dir = (i >> 2) & 1, corner = i & 3;
                        const unsigned same = (cc & kCodeSameObj) ? kItemSameObj : 0u;
                        if (dir == 0) {
                            code = a | (b << 12) | (corner << kItemCornerShift) | (b_static ? kItemRefStatic : 0u) | same;
                            const double4 bb = b_static ? ldg_d4(H->st_aabb + b) : aabb[b];
                            hit = point_in_box(pos[leaf_corner(lpt[a], corner)], bb);
                        } else {
                            code = b | (a << 12) | (corner << kItemCornerShift) | (b_static ? kItemMainStatic : 0u) | same;
                            double2 pp;
                            if (b_static) {
                                pp = __ldg(H->st_pos + __ldg(&H->st_leaf[b].pt[corner]));
                            } else {
                                pp = pos[leaf_corner(lpt[b], corner)];
                            }
                            hit = point_in_box(pp, aabb[a]);
                        }
                    }
                    const unsigned m = __ballot_sync(0xffffffffu, hit);
                    if (hit) surv[scnt + __popc(m & ((1u << lane) - 1u))] = code;
                    scnt += __popc(m);
                }
            }
            const bool cons_done = items_done && next_cc * 32 >= cnt * 8;
            volatile int* scount = misc + kGmSurv + (rc & 1) * 16;         // two parities: a fast warp may start the next round
            if (lane == 0) scount[warp] = scnt;
            flags = cta_or(misc + kGmOr, (cons_done ? 0 : 1) | (prod_done ? 0 : 2), r);
            rc++;
            int n_surv = 0;
#pragma unroll
            for (int w = 0; w < NW; w++) n_surv += scount[w];
            if (n_surv > 0) {
                MEVO_GPROF(2);
                MEVO_GCOUNT(2, n_surv);
                // ---- K3b: resolve survivors, apply their contacts (a-5, a-6)
                PointSrc src;
                src.dyn_pos = pos; src.dyn_vt = vt; src.st_pos = H->st_pos;
                for (int base = 0; base < n_surv; base += G) {
                    const int gi = base + t;
                    bool pending = false;
                    Contact c;
                    int reps = 1;
                    if (gi < n_surv) {
                        int i = gi, w = 0;
                        while (i >= scount[w]) { i -= scount[w]; w++; }
                        const unsigned code = surv_all[w * seg_cap + i];
                        const bool m_static = (code & kItemMainStatic) != 0, r_static = (code & kItemRefStatic) != 0;
                        reps = (code & kItemSameObj) ? 2 : 1;
                        LeafRegs lm, lr;
                        load_leaf(lm, H->g_leaf, H->st_leaf, code & kCodeLeafMask, m_static);
                        load_leaf(lr, H->g_leaf, H->st_leaf, (code >> 12) & kCodeLeafMask, r_static);
                        pending = resolve_point_vs_boxel(src, lm, m_static, lr, r_static, (code >> kItemCornerShift) & 3, c);
                    }
                    // Contacts are applied in rounds: the lowest-numbered unapplied contact on each of its points goes
                    // first, so every point receives its contacts in list order -- the sums of a serial walk of the list.
                    MEVO_GCOUNT(3, 1);
                    while (cta_or(misc + kGmOr, pending ? 1 : 0, r)) {
                        MEVO_GCOUNT(4, 1);
                        if (pending) {
                            if (!c.p_static) atomicMin(&claim[c.p], gi);
                            if (!c.ab_static) { atomicMin(&claim[c.a], gi); atomicMin(&claim[c.b], gi); }
                        }
                        __syncthreads();
                        if (pending) {
                            const bool mine = (c.p_static || claim[c.p] == gi) && (c.ab_static || (claim[c.a] == gi && claim[c.b] == gi));
                            if (mine) {
                                apply_contact(cf, c, reps);
                                if (!c.p_static) claim[c.p] = 0x7fffffff;
                                if (!c.ab_static) { claim[c.a] = 0x7fffffff; claim[c.b] = 0x7fffffff; }
                                pending = false;
                            }
                        }
                    }
                }
                MEVO_GPROF(1);
            }
        } while (flags & 1);
    } while (flags & 2);

    if (dbg_row && t == 0) { *dbg_n = misc[kGmDump]; misc[kGmDump] = 0; }
    if (dbg_cf)
        for (int i = t; i < Pd; i += G) dbg_cf[i] = cf[i];
    // K6a: unstable test uses the boxes of the START of this substep (a-12)
    int sub_unstable = 0;
    for (int o = 0; o < H->n_obj; o++) {
        const ObjRec orc = H->objs[o];
        if (orc.is_robot && misc[kGmSelf + o] > (int)orc.n_surface) sub_unstable = 1;
    }
    return (r << 1) | sub_unstable;
}

// =================================================================================== phase 2: actuators + RK4
// PhysicsEngine::update_actuators + step_rk4 + get_pos_vel_slope (SURVEY 8a-7,8,9).  On return spos holds the new
// positions (pos still holds the substep's start positions = pos_last) and the state block the new velocities.
template <int G, int MINB, int PPT, bool XD>
__device__ __noinline__ void g_rk4(unsigned fixmask) {
    constexpr int U = (G == 128) ? 4 : 3;            // edges per thread in flight: independent sqrt / div chains
    unsigned char* sm = gsm_raw;
    const GHdr* H = reinterpret_cast<const GHdr*>(sm);
    const int t = threadIdx.x;
    const double2* pos = reinterpret_cast<const double2*>(sm + H->o_pos);
    double2* spos = reinterpret_cast<double2*>(sm + H->o_spos);
    const double2* cf = reinterpret_cast<const double2*>(sm + H->o_cf);
    double2* ef = reinterpret_cast<double2*>(sm + H->o_ef);
    double* leq = reinterpret_cast<double*>(sm + H->o_leq);
    const uint4* csr = reinterpret_cast<const uint4*>(sm + H->o_csr);
    const double2* etab = reinterpret_cast<const double2*>(sm + H->o_etab);
    const uint32_t* eidx = reinterpret_cast<const uint32_t*>(sm + H->o_eidx);
    const double* ggoal = H->ggoal;
    const int Pd = H->Pd, Ed = H->Ed;
    MEVO_GPROF(3);
    double2* gvel = reinterpret_cast<double2*>(H->gstate) + Pd;       // velocities: state block, touched by the owner thread only
    double2 v0s[PPT], sv[PPT], ap[PPT], av[PPT];     // velocity, stage velocity and RK4 accumulators of this thread's points
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        sv[k] = make_double2(0, 0); ap[k] = sv[k]; av[k] = sv[k];
        v0s[k] = (t + k * G < Pd) ? gvel[t + k * G] : sv[k];         // in flight while the first spring pass runs
    }

#pragma unroll 1
    for (int s = 0; s < 4; s++) {
        const double2* sp = (s == 0) ? pos : spos;
        for (int base = 0; base < Ed; base += U * G) {
            int e[U];
            uint32_t ix[U];
            double q[U];
            double2 f[U];
#pragma unroll
            for (int u = 0; u < U; u++) {
                e[u] = base + u * G + t;
                const int c = e[u] < Ed ? e[u] : Ed - 1;            // threads past the end recompute a valid edge, store nothing
                ix[u] = eidx[c];
                q[u] = (s == 0) ? ggoal[c] : leq[c];                 // stage 0: length_eq_goal in, relaxed length_eq out
            }
#pragma unroll
            for (int u = 0; u < U; u++) f[u] = spring_force10(sp, ix[u], etab[(ix[u] >> kEidxTypeShift10) & 63].x, q[u], s == 0);
#pragma unroll
            for (int u = 0; u < U; u++)
                if (e[u] < Ed) {
                    ef[e[u]] = f[u];
                    if (s == 0) leq[e[u]] = q[u];
                }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            const int p = t + k * G;
            if (p < Pd) {
                const double2 v = (s == 0) ? v0s[k] : sv[k];
                double Fx = 0.0, Fy = 0.0;
                {
                    // F -= f over the a sides in edge order, then F += f over the b sides (PhysicsEngine.cpp:530-531): the
                    // record lists (edge << 1 | side) in that order; empty slots point at the all-zero entry ef[Ed]
                    const uint4 rec = csr[p];
                    const unsigned w[4] = {rec.x, rec.y, rec.z, rec.w};
#pragma unroll
                    for (int half = 0; half < 2; half++) {
                        double2 g[4];
                        unsigned sg[4];
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            const int qd = 4 * half + j;
                            const unsigned ent = (w[qd >> 1] >> ((qd & 1) * 16)) & 0xffffu;
                            g[j] = ef[ent >> 1];
                            sg[j] = (ent & 1u) ^ 1u;                    // a side: subtract
                        }
#pragma unroll
                        for (int j = 0; j < 4; j++) { Fx += flip_if(g[j].x, sg[j]); Fy += flip_if(g[j].y, sg[j]); }
                    }
                    if (XD && H->csrx) { const double2 a = g_gather_ext(H->csrx, ef, p, make_double2(Fx, Fy)); Fx = a.x; Fy = a.y; }
                }
                Fx += v.x * (-MEVO_VISCOUS_DRAG);
                Fy += v.y * (-MEVO_VISCOUS_DRAG);
                Fy += 1.0 * (-MEVO_GRAVITY);                            // every point mass is 1.0 (host-checked), F / m == F
                const double2 c = cf[p];
                Fx += c.x; Fy += c.y;
                const double y = sp[p].y;
                if (y < 0) {                                            // analytic ground plane y = 0 (PhysicsEngine.cpp:566-577)
                    const double normal_force = y * (-MEVO_COLLISION_CONST_GROUND);
                    Fy += normal_force;
                    const double friction_temp = tanh(v.x * normal_force * MEVO_DYN_FRICTION);
                    Fx += (-10 * MEVO_DYN_FRICTION) * (normal_force * friction_temp);
                }
                const double kpx = v.x * MEVO_DT, kpy = v.y * MEVO_DT;
                const double kvx = Fx * MEVO_DT, kvy = Fy * MEVO_DT;
                const double2 x0 = pos[p], v0 = v0s[k];
                if (s == 0) {
                    ap[k] = make_double2(kpx, kpy);
                    av[k] = make_double2(kvx, kvy);
                    spos[p] = make_double2(x0.x + (kpx * 0.5), x0.y + (kpy * 0.5));
                    sv[k] = make_double2(v0.x + (kvx * 0.5), v0.y + (kvy * 0.5));
                } else if (s < 3) {
                    ap[k].x = ap[k].x + 2 * kpx; ap[k].y = ap[k].y + 2 * kpy;
                    av[k].x = av[k].x + 2 * kvx; av[k].y = av[k].y + 2 * kvy;
                    const double h = (s == 1) ? 0.5 : 1.0;              // x * 1.0 == x exactly
                    spos[p] = make_double2(x0.x + (kpx * h), x0.y + (kpy * h));
                    sv[k] = make_double2(v0.x + (kvx * h), v0.y + (kvy * h));
                } else {
                    double dpx = div_const(ap[k].x + kpx, 6.0), dpy = div_const(ap[k].y + kpy, 6.0);
                    double dvx = div_const(av[k].x + kvx, 6.0), dvy = div_const(av[k].y + kvy, 6.0);
                    if ((fixmask >> k) & 1u) { dpx = 0.0; dpy = 0.0; dvx = 0.0; dvy = 0.0; }
                    // the new position goes to spos: pos keeps the substep's start position (= pos_last)
                    spos[p] = make_double2(x0.x + dpx, x0.y + dpy);
                    gvel[p] = make_double2(v0.x + dvx, v0.y + dvy);
                }
            }
        }
        __syncthreads();
    }
}

// =================================================================================== phase 3: limiters + true velocity
// PhysicsEngine::resolve_edge_constraints + update_true_vel (SURVEY 8a-10,11) on the new positions in spos.  Band test on
// squared lengths (no sqrt), guarded by 1e-7 so the decision never depends on the rounding of dist / L0: inside the band
// every correction of the reference is exactly zero.  Returns r.
template <int G, int MINB, int PPT, bool XD>
__device__ __noinline__ int g_constrain(unsigned fixmask, int first_substep, int r, int ss) {
    constexpr int U = (G == 128) ? 4 : 3;
    unsigned char* sm = gsm_raw;
    const GHdr* H = reinterpret_cast<const GHdr*>(sm);
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    double2* pos = reinterpret_cast<double2*>(sm + H->o_pos);
    double2* spos = reinterpret_cast<double2*>(sm + H->o_spos);
    double2* vt = reinterpret_cast<double2*>(H->gstate) + 3 * H->Pd;               // vel_true: state block
    double2* ef = reinterpret_cast<double2*>(sm + H->o_ef);
    const uint4* csr = reinterpret_cast<const uint4*>(sm + H->o_csr);
    const double2* etab = reinterpret_cast<const double2*>(sm + H->o_etab);
    const uint32_t* eidx = reinterpret_cast<const uint32_t*>(sm + H->o_eidx);
    volatile int* misc = reinterpret_cast<volatile int*>(sm + H->o_misc);
    const int Pd = H->Pd, Ed = H->Ed;

    MEVO_GPROF(4);
    const double4* btab = reinterpret_cast<const double4*>(sm + H->o_btab);
    int viol = 0;
    for (int base = 0; base < Ed; base += U * G) {
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int e = base + u * G + t;
            const bool valid = e < Ed;
            const uint32_t ix = eidx[valid ? e : Ed - 1];
            const double2 pa = spos[ix & 0x3ff], pb = spos[(ix >> kEidxBShift) & 0x3ff];
            const double dx = pa.x - pb.x, dy = pa.y - pb.y;
            const double d2 = dx * dx + dy * dy;
            const double4 bd = btab[(ix >> kEidxTypeShift10) & 63];      // l0^2 * {0.5625001, 1.5624999, 0.9409001, 1.0608999}
            if (valid && !(d2 > bd.x && d2 < bd.y)) viol |= 1;
            if (valid && ((ix >> kEidxRigidBit10) & 1u) && !(d2 > bd.z && d2 < bd.w)) viol |= 2;
        }
    }
    viol = cta_or(misc + kGmOr, viol, r);
    if (viol & 1) {
        MEVO_GPROF(5);
        MEVO_GCOUNT(5, 1);
        // Jacobi half (PhysicsEngine.cpp:613-671): corrections from the pre-correction positions, summed per point in edge order
        for (int e = t; e < Ed; e += G) {
            const uint32_t ix = eidx[e];
            const double2 pa = spos[ix & 0x3ff], pb = spos[(ix >> kEidxBShift) & 0x3ff];
            const double dx = pa.x - pb.x, dy = pa.y - pb.y;
            const double d2 = dx * dx + dy * dy;
            const double l0 = etab[(ix >> kEidxTypeShift10) & 63].y;
            const double l2 = l0 * l0;
            double2 c = make_double2(0.0, 0.0);               // inside the guarded band the correction is exactly zero
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
        __syncthreads();
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            const int p = t + k * G;
            if (p < Pd && !((fixmask >> k) & 1u)) {           // fixed end points receive "0.0" (PhysicsEngine.cpp:670-671)
                double2 x = spos[p];
                const uint4 rec = csr[p];
                const unsigned w[4] = {rec.x, rec.y, rec.z, rec.w};
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const unsigned ent = (w[j >> 1] >> ((j & 1) * 16)) & 0xffffu;
                    const double2 c = ef[ent >> 1];
                    const unsigned neg = (ent & 1u) ^ 1u;
                    x.x += flip_if(c.x, neg); x.y += flip_if(c.y, neg);
                }
                if (XD && H->csrx) x = g_gather_ext(H->csrx, ef, p, x);
                spos[p] = x;
            }
        }
        __syncthreads();
    }
    if (viol & 2) {
        MEVO_GPROF(6);
        MEVO_GCOUNT(6, 1);
        // ordered rigid-cell pass (PhysicsEngine.cpp:675-710), scheduled in point-disjoint levels that keep the reference's
        // order between visits sharing a point; the first warp walks the (narrow) levels with warp-level synchronisation
        if (warp == 0) {
            // the schedule (packed edge word of every visit, level by level) sits in shared memory and the next level's words and
            // band constants are fetched while the current level computes: what remains on the level-to-level chain is the
            // position loads, the band test and, for a visit that moves its points, one sqrt -> divide sequence
            const uint32_t* rsched = reinterpret_cast<const uint32_t*>(sm + H->o_rsched);
            const uint16_t* rl = reinterpret_cast<const uint16_t*>(sm + H->o_rlev);
            const int n_rlev = H->n_rlev;
            auto visit = [&](uint32_t ix, double l0, double lo, double hi) {
                const int a = ix & 0x3ff, b = (ix >> kEidxBShift) & 0x3ff;
                double2 pa = spos[a], pb = spos[b];
                const double dx = pa.x - pb.x, dy = pa.y - pb.y;
                const double d2 = dx * dx + dy * dy;
                if (d2 > lo && d2 < hi) return;                                   // inside the guarded band: a no-op
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
                spos[a] = pa; spos[b] = pb;
                MEVO_GCOUNT_ANY(7, 1);
            };
            int beg = rl[0], end = rl[1];
            bool have = beg + lane < end;
            uint32_t ix = have ? rsched[beg + lane] : 0u;
            double l0 = etab[(ix >> kEidxTypeShift10) & 63].y;
            double4 bd = btab[(ix >> kEidxTypeShift10) & 63];
            for (int lv = 0; lv < n_rlev; lv++) {
                const int nbeg = end, nend = rl[lv + 2];                           // rl is padded: the last levels read "empty"
                const bool nhave = nbeg + lane < nend;
                const uint32_t nix = nhave ? rsched[nbeg + lane] : 0u;
                const double nl0 = etab[(nix >> kEidxTypeShift10) & 63].y;
                const double4 nbd = btab[(nix >> kEidxTypeShift10) & 63];
                if (have) visit(ix, l0, bd.z, bd.w);
                for (int i = beg + lane + 32; i < end; i += 32) {                  // levels wider than a warp (rare)
                    const uint32_t jx = rsched[i];
                    const double4 jb = btab[(jx >> kEidxTypeShift10) & 63];
                    visit(jx, etab[(jx >> kEidxTypeShift10) & 63].y, jb.z, jb.w);
                }
                __syncwarp();
                beg = nbeg; end = nend; have = nhave; ix = nix; l0 = nl0; bd = nbd;
            }
        }
        __syncthreads();
    }
    MEVO_GPROF(7);
    // PhysicsEngine::update_true_vel (:602-607) + largest displacement of this substep (neighbour-list validity)
    unsigned dmax = 0;
    if (t == 0) misc[kGmDisp + ((ss + 1) & 1)] = 0;
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        const int p = t + k * G;
        if (p < Pd) {
            const double2 x = spos[p], x0 = pos[p];
            // pos_last == pos at the start of every substep except possibly the first of a launch
            const double2 l = first_substep ? H->g_last[p] : x0;
            vt[p] = make_double2(div_const(x.x - l.x, MEVO_DT), div_const(x.y - l.y, MEVO_DT));
            pos[p] = x;
            // upper bound of max(|dx|, |dy|) as the high word of the double, + 1 (NaN / inf stay huge: the list is rebuilt)
            const double d = fmax(fabs(x.x - x0.x), fabs(x.y - x0.y));
            const unsigned hi = (d == d) ? (unsigned)__double2hiint(d) + 1u : 0x7ff00000u;
            dmax = max(dmax, hi);
        }
    }
    dmax = __reduce_max_sync(0xffffffffu, dmax);
    if (lane == 0 && dmax) atomicMax((int*)&misc[kGmDisp + (ss & 1)], (int)dmax);
    __syncthreads();
    return r;
}

// =================================================================================== the kernel
// XD: some design of the launch carries gather slots 8..15 (GHdr::csrx); false compiles the extension gather out
template <int G, int MINB, int PPT, bool XD>
__global__ void __launch_bounds__(G, MINB) gstep_kernel(const StepParams prm) {
    static_assert(G % 32 == 0 && G >= 64 && G <= 512, "group size");
    unsigned char* sm = gsm_raw;
    const int t = threadIdx.x;
    const int env = prm.class_envs[blockIdx.x];
    if (!prm.chunk_first && prm.unstable[env]) return;  // Sim::step aborted in an earlier chunk of this step (Sim.cpp:107-110)
    const EnvDesc ed = prm.envs[env];
    const uint8_t* db = prm.design + ed.design_off;
    const DesignHeader* DH = reinterpret_cast<const DesignHeader*>(db);
    const int Pd = DH->Pd, Ed = DH->Ed, n_leaf = DH->n_leaf;
    double* gstate = prm.state + ed.state_off;
    const GLayout L = g_layout(Pd, Ed, n_leaf, DH->n_types, DH->cand_cap, DH->n_rvis, DH->n_rlev);
    unsigned fixmask = 0;                               // bit k: point t + k * G is fixed

    // ---------------------------------------------------------------- stage in (once per launch)
    {
        if (t == 0) {
            GHdr* H = reinterpret_cast<GHdr*>(sm);
            H->Pd = Pd; H->Ed = Ed; H->n_leaf = n_leaf; H->n_pairs = DH->n_pairs; H->n_items = DH->n_items;
            H->n_rlev = DH->n_rlev; H->n_obj = DH->n_obj;
            H->seg_cap = DH->cand_cap / (G / 32);
            H->lseg_cap = DH->list_cap / (G / 32);
            H->iseg_cap = DH->ilist_cap / (G / 32);
            H->o_pos = L.o_pos; H->o_spos = L.o_spos; H->o_cf = L.o_cf; H->o_ef = L.o_ef;
            H->o_leq = L.o_leq; H->o_aabb = L.o_aabb; H->o_csr = L.o_csr; H->o_etab = L.o_etab; H->o_btab = L.o_btab; H->o_eidx = L.o_eidx;
            H->o_lpt = L.o_lpt; H->o_rsched = L.o_rsched; H->o_rlev = L.o_rlev; H->o_cand = L.o_cand; H->o_surv = L.o_surv; H->o_misc = L.o_misc;
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
            H->g_leaf = reinterpret_cast<const LeafRec*>(db + DH->o_leaf);
            H->g_last = reinterpret_cast<const double2*>(gstate) + 2 * Pd;
            H->pairs = reinterpret_cast<const unsigned int*>(db + DH->o_pairs);
            H->items = reinterpret_cast<const unsigned int*>(db + DH->o_items);
            H->objs = reinterpret_cast<const ObjRec*>(db + DH->o_obj);
            H->ggoal = gstate + 8 * (size_t)Pd;
            H->csrx = DH->o_csrx ? reinterpret_cast<const uint4*>(db + DH->o_csrx) : nullptr;
            H->list = prm.scratch + (size_t)ed.scratch_off16 * 4;
            H->ilist = H->list + DH->list_cap;
            H->gstate = gstate;
            H->gstate0 = prm.state0 + ed.state_off;
            H->env = env;
        }
        const double2* g2 = reinterpret_cast<const double2*>(gstate);
        double2* pos = reinterpret_cast<double2*>(sm + L.o_pos);
        uint4* csr = reinterpret_cast<uint4*>(sm + L.o_csr);
        const unsigned zero_ent = (unsigned)Ed << 1;
        for (int i = t; i < Pd; i += G) {
            pos[i] = g2[i];
            uint4 rec = __ldg(reinterpret_cast<const uint4*>(db + DH->o_csr8) + i);
            unsigned* w = reinterpret_cast<unsigned*>(&rec);
#pragma unroll
            for (int q = 0; q < 4; q++) {                   // empty gather slots (0xffff) point at the all-zero entry ef[Ed]
                if ((w[q] & 0xffffu) == 0xffffu) w[q] = (w[q] & 0xffff0000u) | zero_ent;
                if ((w[q] >> 16) == 0xffffu) w[q] = (w[q] & 0xffffu) | (zero_ent << 16);
            }
            csr[i] = rec;
        }
#pragma unroll
        for (int k = 0; k < PPT; k++)
            if (t + k * G < Pd && __ldg(db + DH->o_fixed + t + k * G)) fixmask |= 1u << k;
        uint32_t* eidx = reinterpret_cast<uint32_t*>(sm + L.o_eidx);
        double* ggoal = gstate + 8 * (size_t)Pd;
        const int16_t* eact = reinterpret_cast<const int16_t*>(db + DH->o_eact);
        const double* act = prm.act + (size_t)env * prm.act_stride;
        for (int e = t; e < Ed; e += G) {
            eidx[e] = __ldg(reinterpret_cast<const uint32_t*>(db + DH->o_eidx10) + e);
            if (prm.apply_actions) {
                // K1: PhysicsEngine::update_actuator_goals (PhysicsEngine.cpp:38-97) as a per-edge gather, written to the
                // state block in place: goal = (sum of the <= 2 actuations driving this edge) * (init_len / count).  The
                // spring pass reads ggoal[e] with the same thread (e % G), so program order is all the ordering it needs.
                const int a0 = eact[2 * e], a1 = eact[2 * e + 1];
                if (a0 >= 0) {
                    const double l0 = __ldg(reinterpret_cast<const double*>(db + DH->o_el0) + e);
                    double sum = 0.0 + act[a0];
                    int cnt = 1;
                    if (a1 >= 0) { sum += act[a1]; cnt = 2; }
                    ggoal[e] = sum * (l0 / cnt);
                }
            }
        }
        double2* etab = reinterpret_cast<double2*>(sm + L.o_etab);
        double4* btab = reinterpret_cast<double4*>(sm + L.o_btab);
        for (int i = t; i < DH->n_types; i += G) {
            const double2 kl = __ldg(reinterpret_cast<const double2*>(db + DH->o_etab) + i);
            etab[i] = kl;
            const double l2 = kl.y * kl.y;              // the limiters' guarded bands on squared lengths, as they are evaluated per edge
            btab[i] = make_double4(l2 * 0.5625001, l2 * 1.5624999, l2 * 0.9409001, l2 * 1.0608999);
        }
        {   // rigid-cell visit schedule: packed edge word per visit, level offsets (padded with two copies of the end)
            uint32_t* rsched = reinterpret_cast<uint32_t*>(sm + L.o_rsched);
            uint16_t* rl = reinterpret_cast<uint16_t*>(sm + L.o_rlev);
            const uint16_t* g_rvis = reinterpret_cast<const uint16_t*>(db + DH->o_rvis);
            const uint16_t* g_rlev = reinterpret_cast<const uint16_t*>(db + DH->o_rlev);
            for (int i = t; i < DH->n_rvis; i += G)
                rsched[i] = __ldg(reinterpret_cast<const uint32_t*>(db + DH->o_eidx10) + __ldg(g_rvis + i));
            for (int i = t; i < DH->n_rlev + 3; i += G) rl[i] = __ldg(g_rlev + min(i, DH->n_rlev));
        }
        if (t == 0) reinterpret_cast<double2*>(sm + L.o_ef)[Ed] = make_double2(0.0, 0.0);
        uint2* lpt = reinterpret_cast<uint2*>(sm + L.o_lpt);
        for (int l = t; l < n_leaf; l += G) lpt[l] = __ldg(reinterpret_cast<const uint2*>(db + DH->o_lpt16) + l);
        volatile int* misc = reinterpret_cast<volatile int*>(sm + L.o_misc);
        for (int i = t; i < kGMiscInts; i += G) misc[i] = 0;
    }
    __syncthreads();

    int unstable = 0, completed_steps = 0, first_substep = 1;
    int r = 0, ss = 0;                                  // cta_or call counter, executed-substep counter (uniform)
    double D = 0.0;                                     // bound of the displacement of any point since the list was built
    // (everything else the phases need is in the shared-memory header: the loop below keeps a dozen registers live across
    // the out-of-line calls)
#ifdef MEVO_PROF
    if (threadIdx.x == 0) {
        long long* st = reinterpret_cast<long long*>(sm + L.o_misc + 4 * kGmProf);
        st[0] = clock64(); st[1] = 0;
    }
#endif

    for (int step = 0; step < prm.n_steps; step++) {
        for (int sub = 0; sub < prm.n_substeps; sub++) {
            if (unstable) continue;                       // Sim::step aborted (Sim.cpp:107-110)
            const bool dump = step == 0 && sub == 0;
            const int rebuild = first_substep || !(2.0 * D <= kListMargin - kListSlack);
            unsigned int* dbg_row = nullptr;
            int* dbg_n = nullptr;
            double2* dbg_cf = nullptr;
            if (dump && (prm.dbg_cand || prm.dbg_cf)) {
                const GHdr* H = reinterpret_cast<const GHdr*>(sm);
                if (prm.dbg_cand) { dbg_row = prm.dbg_cand + (size_t)H->env * prm.dbg_cand_cap; dbg_n = prm.dbg_cand_n + H->env; }
                if (prm.dbg_cf) dbg_cf = reinterpret_cast<double2*>(prm.dbg_cf + (H->gstate - prm.state));
            }
            const int cr = g_collide<G, MINB>(rebuild, r, dbg_row, dbg_n, prm.dbg_cand_cap, dbg_cf);
            r = cr >> 1;
            if (rebuild) D = 0.0;
            g_rk4<G, MINB, PPT, XD>(fixmask);
            r = g_constrain<G, MINB, PPT, XD>(fixmask, first_substep, r, ss);
            {
                const GHdr* H = reinterpret_cast<const GHdr*>(sm);
                const volatile int* misc = reinterpret_cast<const volatile int*>(sm + H->o_misc);
                D += __hiloint2double(misc[kGmDisp + (ss & 1)], 0);
            }
            ss++;
            first_substep = 0;
            if (cr & 1) unstable = 1;                     // Sim::step aborts the remaining substeps (Sim.cpp:107-110)
        }
        if (!unstable) completed_steps++;
    }

    // ---------------------------------------------------------------- stage out (goals were written at stage in, velocities every substep)
    {
        const GHdr* H = reinterpret_cast<const GHdr*>(sm);
        const int tt = threadIdx.x, n_pts = H->Pd, e_id = H->env;
        double2* o2 = reinterpret_cast<double2*>(H->gstate);
        if (unstable && prm.auto_reset) {
            // Sim::revert(0) in place (Environment.cpp:247-263): pos, vel <- snapshot; pos_last = pos; vel_true = vel
            const double2* s0 = reinterpret_cast<const double2*>(H->gstate0);
            for (int i = tt; i < n_pts; i += G) {
                const double2 p0 = s0[i], v0 = s0[n_pts + i];
                o2[i] = p0; o2[n_pts + i] = v0; o2[2 * n_pts + i] = p0; o2[3 * n_pts + i] = v0;
            }
        } else if (!first_substep) {
            const double2* pos = reinterpret_cast<const double2*>(sm + H->o_pos);
            for (int i = tt; i < n_pts; i += G) {          // vel and vel_true are in place already
                const double2 x = pos[i];
                o2[i] = x;
                o2[2 * n_pts + i] = x;                     // pos_last = pos after update_true_vel
            }
        }
        if (tt == 0) {
            prm.unstable[e_id] = (uint8_t)unstable;
            if (unstable && prm.auto_reset) prm.sim_time[e_id] = 0;
            else if (prm.chunk_last) prm.sim_time[e_id] += completed_steps;
            if (unstable) atomicAdd(prm.unstable_events, 1ull);
        }
    }
}

}  // namespace mevo
