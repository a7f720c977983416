// Flattened BVH over the traced primitives' world bounding boxes.  The reference has no BVH: it
// has Bounded::build_world_bounding_box (bounding_box.rs:69-73; traced.rs:65-70,208-219,301-304)
// feeding a screen-space grid for primary rays only (camera_hash.rs) and a linear scan for every
// other ray (hits.rs:270-286).  This builder takes the same per-object world AABBs (tightened:
// exact ellipsoid extents, the cube's real +-0.5 corners, the triangle/parallelogram vertices)
// and produces a binary tree that serves every ray.
#pragma once
#include <stdint.h>

#include <vector>

namespace omb {

struct Aabb {
    float lo[3], hi[3];
};

// 64-byte node: both children's boxes + child references (one shared-memory fetch per visit).
// child >= 0: index of an inner node; child < 0: leaf, ~child is the primitive reference
// (omd::make_ref: type code << 27 | index, always below 2^31 so that ~ref is negative).  An empty child has an inverted box and is never entered.
struct alignas(16) Node {
    float lmin[3], lmax[3], rmin[3], rmax[3];
    int32_t left, right;
    uint32_t pad0, pad1;
};
static_assert(sizeof(Node) == 64, "node must be 64 bytes");

struct BuildResult {
    std::vector<Node> nodes;  // nodes[0] is the root when !nodes.empty()
    uint32_t depth = 0;
};

// The traversal's node: the binary tree collapsed to four children per node (boxes stored per axis, low planes and high
// planes 16 bytes apart, so that a ray fetches its near and far planes by address from the sign of its direction).
// child >= 0: inner node index; child < 0: leaf, ~child is the primitive reference; an empty slot has an inverted
// box (lo = +INF, hi = -INF) that no ray enters.
struct alignas(16) Node4 {
    float lox[4], hix[4], loy[4], hiy[4], loz[4], hiz[4];
    int32_t child[4];
};
static_assert(sizeof(Node4) == 112, "wide node must be 112 bytes");

struct Wide4 {
    std::vector<Node4> nodes;  // nodes[0] is the root when !nodes.empty()
    uint32_t depth = 0;        // levels of wide nodes
    uint32_t stack_need = 0;   // most entries a traversal can have on its stack: max over root-to-node paths of sum (children - 1)
};
// Collapses the binary tree: a node's children are the binary node's two children, then repeatedly the inner child
// with the largest box is replaced by its own two children until there are `width` (2..4) of them.
void collapse4(const BuildResult& in, Wide4& out, uint32_t width = 4);

// refs[i] / boxes[i] describe primitive i.  Binned SAH, one primitive per leaf.
void build(const std::vector<uint32_t>& refs, const std::vector<Aabb>& boxes, BuildResult& out);

// The same tree shape contract built on the current CUDA device (om_lbvh.cu): Morton order + Karras' hierarchy.
// false = not built, `err` says why (a CUDA error, fewer than 2 primitives, a non-finite box, or a tree deeper
// than max_depth, which the traversal stacks could not hold).
bool build_gpu(const std::vector<uint32_t>& refs, const std::vector<Aabb>& boxes, BuildResult& out, uint32_t max_depth, char* err, size_t errlen);

}  // namespace omb
