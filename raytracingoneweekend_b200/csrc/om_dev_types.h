// Layout of the frozen scene as the kernels see it.  Shared by the host packer (om_scene.cpp)
// and the device code (om_device.cuh).  Everything is addressed in 16-byte units so that every
// access is one LDS.128 / LDG.128.
//
//  hot image  (staged into shared memory by TMA bulk copy when it fits; else read from global/L2)
//      BVH nodes 112 B (four children per node) | sphere world->local rows 48 B | cube world->local rows 48 B |
//      triangle 64 B | infinite plane 32 B | parallelogram 64 B | always-test refs 4 B |
//      marched sphere 16 B | marched box 32 B | marched torus 144 B | dyn marched 112 B + its program, 4 B per word | dyn traced 112 B |
//      sphere material type 1 B
//  cold image (global memory; touched once per hit, not per test)
//      sphere local->world rows 48 B | cube local->world rows 48 B | dyn traced local->world rows 48 B | material+bloom 32 B per object
#pragma once
#include <stdint.h>

namespace omd {

// primitive type codes in the reference's scan order (hits.rs:370-371): on equal t the LATER
// object of the linear scan wins (`root > t_max` rejects only strictly larger roots,
// traced.rs:51,108,178,282), so a larger ref beats a smaller one.
// PT_TDYN (om_dyn_traced, the reference's traced_objects) is scanned after the parallelograms (hits.rs:280-285): among the traced
// types it has the largest code, so it wins ties like a later object.  (Marched codes never meet traced ones in a comparison.)
enum : uint32_t { PT_SPHERE = 0, PT_CUBE = 1, PT_TRI = 2, PT_PLANE = 3, PT_PARA = 4, PT_MSPHERE = 5, PT_MBOX = 6, PT_MTORUS = 7, PT_MDYN = 8, PT_TDYN = 9 };
constexpr uint32_t PT_COUNT = 10;
constexpr uint32_t REF_NONE = 0xFFFFFFFFu;
#if defined(__CUDACC__)
#define OMD_HD __host__ __device__ __forceinline__
#else
#define OMD_HD inline
#endif
// ref = type code (4 bits, at 27..30) | index in its list (27 bits).  Bit 31 stays clear: a BVH leaf is stored as ~ref and must read
// back negative (om_bvh.h), and refs compare as (type, index) pairs for the tie rule above.
constexpr uint32_t REF_INDEX_BITS = 27u, REF_MAX_OBJECTS = (1u << REF_INDEX_BITS) - 1u;
OMD_HD uint32_t make_ref(uint32_t type, uint32_t index) { return (type << REF_INDEX_BITS) | index; }
OMD_HD uint32_t ref_type(uint32_t ref) { return ref >> REF_INDEX_BITS; }
OMD_HD uint32_t ref_index(uint32_t ref) { return ref & REF_MAX_OBJECTS; }
static_assert(PT_COUNT <= 16u, "type codes have 4 bits");

// sizes in 16-byte units
constexpr uint32_t U_NODE = 4 /* binary node: distance BVH over the marched objects */, U_NODE4 = 7 /* traversal node, omb::Node4 */, U_XF = 3, U_BARY = 4, U_PLANE = 2, U_MSPHERE = 1, U_MBOX = 2, U_MTORUS = 9, U_MDYN = 7, U_MAT = 2,
                   U_TDYN = 7 /* world->local rows 0-2, {A,B,C,D}, {E,F,G,H}, {I,J,clip_lo.x,clip_lo.y}, {clip_lo.z,clip_hi.xyz} */;

struct SceneHeader {
    uint32_t n[PT_COUNT];  // object count per PT_* type
    uint32_t n_nodes;      // BVH nodes, four children each (0 = no BVH primitives)
    uint32_t n_large;      // refs in the always-test list (infinite planes, scene-sized objects)
    uint32_t n_bvh_prims;  // primitives reachable through the BVH
    uint32_t bvh_depth;    // levels of nodes
    uint32_t stack_need;   // most entries a traversal can hold on its stack (omb::Wide4::stack_need)
    // hot image offsets, 16-byte units
    uint32_t off_nodes, off_hot[PT_COUNT], off_large;
    uint32_t hot_units;    // total size of the hot image
    // cold image offsets, 16-byte units
    uint32_t off_sph_l2w, off_cube_l2w, off_mat;
    uint32_t off_tdyn_l2w;  // dyn traced objects: local->world rows 0-2 (U_XF units each)
    uint32_t mat_base[PT_COUNT];  // first material record of each type
    uint32_t cold_units;
    // padded union of the marched objects' bounds: a ray that misses it cannot hit a marched
    // object and cannot start inside one.  lo > hi when the scene has no marched object.
    float march_lo[3], march_hi[3];
    uint32_t n_marched;
    // Candidate grid over the marched objects (0 cells = none): cell (ix,iy,iz) of the box starting at mgrid_lo with
    // edge 1/mgrid_inv_h lists, in the reference's order, every marched object that can attain min_j |sdf_j(p)| at some
    // point p of the cell; all others are provably farther than some listed one everywhere in the cell, so skipping
    // them leaves both the minimum and the first object attaining it (hits.rs:301-311) unchanged.
    // cold image: cells = uint2 {first, count} (2 per unit) at off_mgrid_cells, lists = j (4 per unit) at off_mgrid_lists
    float mgrid_lo[3], mgrid_inv_h;
    uint32_t mgrid_dim[3], off_mgrid_cells, off_mgrid_lists;
    // Distance BVH over the marched objects (0 nodes = none), used where the grid does not reach: nodes as in the traced
    // BVH (64 B, both children's boxes, child < 0 = ~j) at off_mbvh_nodes in the cold image; objects too large to bound
    // usefully (j, 4 per unit) at off_mbvh_always.  |sdf_j(p)| >= mbvh_f * dist(p, box_j) - mbvh_margin * (1 + |p|).
    uint32_t n_mbvh_nodes, off_mbvh_nodes, n_mbvh_always, off_mbvh_always;
    float mbvh_f, mbvh_margin;
    // != 0: the BVH boxes carry only the distance-independent padding; the traversal widens each box by
    // dyn_pad * (largest distance from the ray origin to a point of the box), see om_scene.cpp
    float dyn_pad;
    // with dyn_pad: one float4 per wide node at off_node_far in the hot image, {centre, radius} of a ball around the node's
    // children - |origin - centre| + radius bounds the distance from a ray origin to any point of their boxes
    uint32_t off_node_far;
    // dyn marched object i (PT_MDYN): world->local rows 0-2, local->world rows 0-2, {world_f, first word, word count, 0};
    // the programs' words, 4 per unit, start at off_dyn_code
    uint32_t off_dyn_code;
    uint32_t off_sph_mat;  // hot image: one byte per sphere = its material type (shade class of a hit, om_wavefront.cuh)
};

}  // namespace omd
