// Fast build of the kernels: the same source with FMA contraction on.
#define OM_EXACT 0
#define OM_NS om_fast
#include "om_device.cuh"
namespace OM_NS {
#include "om_wavefront.cuh"
#include "om_pool.cuh"
}
#include "om_launch.inl"
