// Host stand-in for the handful of OptiX *device* intrinsics the reference's programs call
// (editable_gauss_refl/cuda/csrc/shaders.cu, forward_pass.cu, backward_pass.cu).  TEST INFRASTRUCTURE ONLY.
// The acceleration structure is replaced by a brute-force loop over the instances (the BVH only enumerates
// candidates); everything the programs compute is the reference's own source, compiled in place.
#pragma once
#include <cstdint>

typedef unsigned long long OptixTraversableHandle;
typedef unsigned int OptixVisibilityMask;
enum { OPTIX_RAY_FLAG_NONE = 0 };

struct RefInstance {      // what OptiX keeps per instance (bvh_wrapper.cu:33-59)
    float4 l2w[3];        // object-to-world rows (3x4)
    float4 w2l[3];        // world-to-object rows (3x4)
    float3 center;        // bounding sphere of the transformed unit box (pure accelerator)
    float radius2;
    unsigned visible;
};

struct RefThreadState {
    uint3 launch_index, launch_dim;
    uint32_t *payload[8];
    float3 obj_origin, obj_direction;
    uint32_t instance_index;
};
extern thread_local RefThreadState ref_ts;
extern const RefInstance *ref_instances;
extern uint32_t ref_num_instances;

extern "C" void __intersection__gaussian();

inline uint3 optixGetLaunchIndex() { return ref_ts.launch_index; }
inline uint3 optixGetLaunchDimensions() { return ref_ts.launch_dim; }
inline float3 optixGetObjectRayOrigin() { return ref_ts.obj_origin; }
inline float3 optixGetObjectRayDirection() { return ref_ts.obj_direction; }
inline unsigned int optixGetInstanceIndex() { return ref_ts.instance_index; }
#define REF_PAYLOAD(n)                                                              \
    inline unsigned int optixGetPayload_##n() { return *ref_ts.payload[n]; }       \
    inline void optixSetPayload_##n(unsigned int v) { *ref_ts.payload[n] = v; }
REF_PAYLOAD(0) REF_PAYLOAD(1) REF_PAYLOAD(2) REF_PAYLOAD(3) REF_PAYLOAD(4) REF_PAYLOAD(5) REF_PAYLOAD(6) REF_PAYLOAD(7)

inline OptixTraversableHandle optixGetInstanceTraversableFromIAS(OptixTraversableHandle, unsigned int id) { return id; }
inline const float4 *optixGetInstanceInverseTransformFromHandle(OptixTraversableHandle h) { return ref_instances[h].w2l; }
inline const float4 *optixGetInstanceTransformFromHandle(OptixTraversableHandle h) { return ref_instances[h].l2w; }

// optixTraverse: present to the intersection program every visible instance whose object-space unit box
// [-1,1]^3 (bvh_wrapper.h:79) is crossed by the ray segment [tmin, tmax].
void optixTraverse(OptixTraversableHandle, float3 o, float3 d, float tmin, float tmax, float, OptixVisibilityMask,
                   unsigned, unsigned, unsigned, unsigned, uint32_t &p0, uint32_t &p1, uint32_t &p2, uint32_t &p3,
                   uint32_t &p4, uint32_t &p5, uint32_t &p6, uint32_t &p7);
