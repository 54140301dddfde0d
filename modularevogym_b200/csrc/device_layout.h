// Byte layouts shared by the host packer (engine.cu: Batch::pack_env) and the kernels.
//
// Three kinds of blobs live in HBM, all 16-byte aligned, addressed by byte offset:
//   * design blob  -- read-only topology of the DYNAMIC part of one environment (robot(s) and any
//                     object that can move): spring constants, end points, per-point gather lists,
//                     surface leaves, leaf-pair table, rigid-cell visit schedule.  Identical designs
//                     are stored once and shared by all environments that use them.
//   * static blob  -- read-only collision geometry of the all-FIXED objects (terrain): points,
//                     leaves, precomputed AABBs and a dense uniform grid (cell = one voxel) mapping
//                     world cell -> leaf.  Shared by all environments with the same terrain.
//   * state block  -- per environment, read+written once per step: pos, vel, pos_last, vel_true as
//                     arrays of double2 over the dynamic points, then length_eq_goal over the
//                     dynamic edges.
#pragma once

#include <cstdint>

#if defined(__CUDACC__)
#define MEVO_HD __host__ __device__ __forceinline__
#else
#define MEVO_HD inline
#endif

namespace mevo {

constexpr int kMaxWarpsPerGroup = 16;    // group sizes 32 .. 512
constexpr int kMaxDynObjects = 8;        // dynamic objects per environment (3-bit field in pair codes)
constexpr int kMaxLeaves = 4095;         // 12-bit leaf ids in pair / candidate codes

struct alignas(16) LeafRec {  // 32 bytes; one surface boxel
    uint16_t pt[4];          // TL, TR, BR, BL point ids (dynamic-local or static-local)
    uint16_t ea[4];          // a end point of the top, right, bot, left edge (oriented, Edge.h:14-15)
    uint16_t eb[4];          // b end point
    uint8_t surf_mask;       // bit s: edge s isOnSurface (Edge.h:25)
    uint8_t obj;             // dynamic object slot (dynamic leaves only)
    int16_t gx, gy;          // cell coordinates inside the owning object's grid (x, y from the top)
    uint16_t pad;
};
static_assert(sizeof(LeafRec) == 32, "LeafRec must be 32 bytes");

struct ObjRec {              // 8 bytes; one dynamic object
    uint8_t is_robot;
    uint8_t pad;
    uint16_t n_surface;      // surface_boxels_index.size() (PhysicsEngine.cpp:830)
    uint16_t leaf_start, leaf_count;
};
static_assert(sizeof(ObjRec) == 8, "ObjRec must be 8 bytes");

// pair-table / candidate code bits
constexpr uint32_t kCodeLeafMask = 0xfffu;
constexpr uint32_t kCodeSameObj = 1u << 24;   // both leaves in the same object: every contact counts twice
constexpr uint32_t kCodeCount = 1u << 25;     // pair table: robot self pair with non-adjacent cells (a-12)
constexpr uint32_t kCodeStatic = 1u << 25;    // candidate list: leaf b is a static leaf
constexpr int kCodeObjShift = 26;

// narrow-phase item code: one corner of leaf `main` against leaf `ref`
//   bits 0-11 main leaf, 12-23 ref leaf, 24-25 corner (TL,TR,BR,BL), 26 main is static, 27 ref is static,
//   28 both leaves in one object (every contact counts twice)
constexpr uint32_t kItemMainStatic = 1u << 26;
constexpr uint32_t kItemRefStatic = 1u << 27;
constexpr uint32_t kItemSameObj = 1u << 28;
constexpr int kItemCornerShift = 24;

// per-edge flags (design blob / shared memory)
constexpr uint8_t kEdgeRigidCell = 1;         // side of a CELL_RIGID voxel: visited by the ordered rigid pass

struct DesignHeader {        // 128 bytes
    int Pd, Ed, n_leaf, n_obj, n_pairs, n_rlev, n_rvis, n_act;
    int cand_cap, o_mass, o_ek, o_el0, o_ea, o_eb, o_eact, o_fixed;
    int o_csro, o_csre, o_leaf, o_obj, o_pairs, o_rlev, o_rvis, total;
    int n_items, o_items, o_eflag;
    // warp kernel (wkernel.cuh): compact copies that fit its shared-memory layout
    int max_deg, unit_mass, o_csr8, o_eidx;   // 8-slot gather records; a | b << 8 | type << 16 | rigid-cell << 20
    int n_types, o_etab;                      // distinct (spring constant, init_length_eq) pairs, double2 each
    int o_pairs16, o_items16, o_rord8;        // 16-bit pair / item codes, rigid-cell visits (u8 edge ids) in reference order
    int warp_ok;                              // every compact encoding above is valid for this design
    int reserved[3];
};
static_assert(sizeof(DesignHeader) == 160, "DesignHeader must be 160 bytes");

// 16-bit codes of the warp kernel (<= 32 leaves): pair = a | b << 5 | same-object << 10 | count << 11 | object << 12
//                                                 item = main | ref << 5 | corner << 10 | same-object << 12
constexpr uint32_t kPair16Same = 1u << 10, kPair16Count = 1u << 11;
constexpr int kPair16ObjShift = 12;
constexpr uint32_t kItem16Same = 1u << 12;
constexpr int kEidxTypeShift = 16, kEidxRigidBit = 20;

struct StaticHeader {        // 64 bytes
    int Ps, n_leaf, gx0, gy0, gw, gh, o_pos, o_leaf;
    int o_aabb, o_grid, total, reserved[5];
};
static_assert(sizeof(StaticHeader) == 64, "StaticHeader must be 64 bytes");

struct EnvDesc {             // 32 bytes
    unsigned long long design_off;   // bytes into the design buffer
    unsigned long long static_off;   // bytes into the static buffer, ~0ull = no static geometry
    unsigned long long state_off;    // doubles into the state buffer
    int act_off;                     // doubles into the action buffer
    int pad;
};

MEVO_HD int align16(int x) { return (x + 15) & ~15; }

// shared-memory carve-up of one environment (byte offsets from the group's base)
struct SmemLayout {
    int o_pos, o_vel, o_last, o_vt, o_cf, o_spos, o_svel, o_accp, o_accv;
    int o_mass, o_ek, o_el0, o_leq, o_goal, o_ef, o_aabb, o_leaf;
    int o_ea, o_eb, o_csro, o_csre, o_fixed, o_eflag, o_cand, o_surv, o_misc, total;
};

constexpr int kMiscInts = 64;   // [0..15] per-warp candidate counts, [16..23] self-collision counts,
                                // [24] constraint-violation flag, [32..47] per-warp contact masks,
                                // [48..63] per-warp survivor counts

// Group sizes whose read-only tables (mass, fixed flags, spring constants, rest lengths, end points, gather lists, edge flags:
// 9 B per point + 25 B per edge) are read from the design blob through L1 instead of being copied to shared memory: for the
// 256-thread class that is the difference between one and two CTAs per SM.
constexpr int kGlobalTablesGroup = 256;
MEVO_HD bool tables_in_global(int group_size) { return group_size == kGlobalTablesGroup; }

MEVO_HD SmemLayout smem_layout(int Pd, int Ed, int n_leaf, int cand_cap, bool global_tables = false) {
    SmemLayout L;
    int o = 0;
    const int tp = global_tables ? 0 : Pd, te = global_tables ? 0 : Ed;      // table sizes in shared memory
    L.o_pos = o;  o += 16 * Pd;
    L.o_vel = o;  o += 16 * Pd;
    L.o_last = o; o += 16 * Pd;
    L.o_vt = o;   o += 16 * Pd;
    L.o_cf = o;   o += 16 * Pd;
    L.o_spos = o; o += 16 * Pd;
    L.o_svel = o; o += 16 * Pd;
    L.o_accp = o; o += 16 * Pd;
    L.o_accv = o; o += 16 * Pd;
    L.o_mass = o; o += align16(8 * tp);
    L.o_ek = o;   o += align16(8 * te);
    L.o_el0 = o;  o += align16(8 * te);
    L.o_leq = o;  o += align16(8 * Ed);
    L.o_goal = o; o += align16(8 * Ed);
    L.o_ef = o;   o += 16 * Ed;
    L.o_aabb = o; o += 32 * n_leaf;
    L.o_leaf = o; o += 32 * n_leaf;
    L.o_ea = o;   o += align16(2 * te);
    L.o_eb = o;   o += align16(2 * te);
    L.o_csro = o; o += global_tables ? 0 : align16(2 * (Pd + 1));
    L.o_csre = o; o += align16(4 * te);
    L.o_fixed = o; o += align16(tp);
    L.o_eflag = o; o += align16(te);
    L.o_cand = o; o += align16(4 * cand_cap);
    L.o_surv = o; o += align16(4 * cand_cap);
    L.o_misc = o; o += 4 * kMiscInts;
    L.total = align16(o);
    return L;
}

}  // namespace mevo
