// cl_ref_driver.cpp -- TEST INFRASTRUCTURE.  Host driver around the reference's OpenCL kernel
// `rayTracer` (RayTrace/resources/shaders/rayTracer.cl:111-202) compiled as C++ through cl_shim.h:
// one call of the kernel function per work item, exactly the NDRange of
// MainState::executeRayTracerOpenCL (states/MainState.cpp:858-869: global size = pixelCount).
// The kernel text is included from a scratch copy made by oracle/build_ref.py whose ONLY edit is the
// OpenCL vector-literal syntax `(float4)(a,b,c,d)` -> `make_float4(a,b,c,d)` (a C++ cast would keep
// only `d`).  Nothing of the reference is written into this repository.
#include "cl_shim.h"

#include <cstdint>
#ifdef _OPENMP
#include <omp.h>
#endif

thread_local size_t cl_shim_global_id = 0;

#include "raytracer_cl.inc"

extern "C" {

// Arguments = the kernel's (rayTracer.cl:111-114).  output: 4 ints per ray.
int cl_ref_run(size_t n_rays, int* output, int numSpheres, const float* sphereOrigins, const float* sphereRadius,
               const float* sphereColours, int numCubes, const float* cubeVertices, const float* cubeColours,
               const float* rayOrigins, const float* rayDir, int threads) {
    if (!output || !rayOrigins || !rayDir) return 1;
    const float4 dir = make_float4(rayDir[0], rayDir[1], rayDir[2], rayDir[3]);
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 256) num_threads(threads)
#endif
    for (long long i = 0; i < (long long)n_rays; ++i) {
        cl_shim_global_id = (size_t)i;
        rayTracer(output, numSpheres, (float4*)sphereOrigins, (float*)sphereRadius, (float4*)sphereColours, numCubes,
                  (float4*)cubeVertices, (float4*)cubeColours, (float4*)rayOrigins, dir);
    }
    return 0;
}

}  // extern "C"
