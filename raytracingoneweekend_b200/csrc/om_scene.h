// FrozenHittableList (hits.rs:63-69) as device data: HittableList::freeze packs the eight
// object lists, their materials and a BVH into the two images described in om_dev_types.h.
#pragma once
#include <string>
#include <vector>

#include "om_dev_types.h"
#include "om_host.h"

namespace oms {

struct FrozenScene {
    omd::SceneHeader hdr;
    std::vector<float> hot;   // hdr.hot_units * 4 floats
    std::vector<float> cold;  // hdr.cold_units * 4 floats
    uint64_t n_total = 0;
};

// flags: bit 0 = put every bounded primitive in the BVH even when it is scene-sized;
//        bit 1 / bit 2 = force / forbid per-ray box padding (SceneHeader::dyn_pad; default: chosen from the scene)
//        bit 3 = no distance BVH over the marched objects (SceneHeader::mbvh_*)
//        bit 4 = build the BVH of the traced primitives on the current CUDA device (om_lbvh.cu)
// cam (may be NULL) only widens the numerical margins of the boxes; it never changes a result.
// Returns an empty string, or what went wrong (only the device BVH build, flag bit 4, can fail).
std::string freeze(const omh::SceneLists& lists, const om_camera* cam, uint32_t flags, FrozenScene& out);

}  // namespace oms
