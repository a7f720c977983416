// Small f32 vector/matrix layer shared by the host mirror (constructors, scene generators) and
// the CUDA kernels.  The association order of every expression follows the reference's
// operators (src/math/vec3.rs, vec4.rs, mat3x3.rs, mat4x4.rs) so that a build without FMA
// contraction reproduces the reference arithmetic bit for bit; the fast kernels compile the very
// same expressions with contraction on.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define OM_HD __host__ __device__ __forceinline__
#else
#define OM_HD inline
#endif

namespace om {

struct V3 {
    float x, y, z;
};
OM_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
OM_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
OM_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
OM_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }  // vec3.rs:194 a+(-b), same bits
OM_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
OM_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
OM_HD V3 operator*(float s, V3 a) { return v3(a.x * s, a.y * s, a.z * s); }
OM_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }  // vec3.rs:29-31
OM_HD float length_squared(V3 a) { return dot(a, a); }
OM_HD float length(V3 a) { return sqrtf(dot(a, a)); }
OM_HD V3 div(V3 a, float s) { return a * (1.f / s); }  // vec3.rs:236-240: multiply by the reciprocal
OM_HD V3 unit(V3 a) { return div(a, length(a)); }      // vec3.rs:38-40
OM_HD V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
OM_HD V3 vabs(V3 a) { return v3(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }
OM_HD float max_val(V3 a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }  // vec3.rs:47-49
OM_HD float min_val(V3 a) { return fminf(a.x, fminf(a.y, a.z)); }
OM_HD float comp(V3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

}  // namespace om
