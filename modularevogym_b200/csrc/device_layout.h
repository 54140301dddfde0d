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

constexpr int kMaxDynObjects = 8;        // robots per environment (self-collision counters), and dynamic objects the warp kernels take
constexpr int kMaxPairObjects = 64;      // dynamic objects per environment (6-bit object field of the pair codes, kCodeObjShift)
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

struct DesignHeader {        // 160 bytes
    int Pd, Ed, n_leaf, n_obj, n_pairs, n_rlev, n_rvis, n_act;
    int cand_cap;                             // group kernel: candidate / survivor buffer entries (64 per warp, drained in rounds)
    int list_cap;                             // group kernel: neighbour-list entries in the environment's scratch block
    int o_el0, o_eact, o_fixed, o_leaf, o_obj, o_pairs;
    int o_rlev, o_rvis, n_items, o_items, total;
    // compact tables shared by both kernels
    int max_deg, unit_mass, o_csr8;           // 8-slot gather records (edge << 1 | side, u16 each, 0xffff = empty)
    int n_types, o_etab;                      // distinct (spring constant, init_length_eq) pairs, double2 each
    // warp kernel (wkernel.cuh): encodings that fit its shared-memory layout
    int o_eidx;                               // u32 per edge: a | b << 8 | type << 16 | rigid-cell << 20
    int o_pairs16, o_items16, o_rord8;        // 16-bit pair / item codes, rigid-cell visits (u8 edge ids) in reference order
    int warp_ok;                              // every compact encoding above is valid for this design
    // group kernel (gkernel.cuh)
    int o_eidx10;                             // u32 per edge: a | b << 10 | type << 20 | rigid-cell << 26
    int o_lpt16;                              // 8 bytes per leaf: TL, TR, BR, BL point ids, u16 each
    int ilist_cap;                            // neighbour-list entries for the always-candidate items (follows the pair list)
    int o_csrx;                               // 0, or (max_deg > 8: unshared duplicate side edges) a second 8-slot record per point
    // warp kernel: tables in the exact form its shared memory holds them, staged in by bulk copies
    int o_lpt8;                               // u32 per leaf: TL | TR << 8 | BR << 16 | BL << 24 point ids
    unsigned int fixed_lo, fixed_hi;          // bit p: simulated point p (< 64) is fixed
    int reserved[2];
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
    int scratch_off16;               // 16-byte units into the scratch buffer (group kernel: neighbour list)
};

MEVO_HD int align16(int x) { return (x + 15) & ~15; }

}  // namespace mevo
