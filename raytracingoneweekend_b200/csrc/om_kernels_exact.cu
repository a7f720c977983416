// Parity build of the kernels: compiled with -fmad=false (no FMA contraction); div and sqrt are
// IEEE by nvcc's defaults (-prec-div=true -prec-sqrt=true), denormals are kept (-ftz=false).
#define OM_EXACT 1
#define OM_NS om_exact
#include "om_device.cuh"
namespace OM_NS {
#include "om_wavefront.cuh"
#include "om_pool.cuh"
}
#include "om_launch.inl"
