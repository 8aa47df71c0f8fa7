// pt_lbvh.h — device builder of the FAST-mode wide BVH: an alternative to the host builder of the
// reference (BVH_Create / BVH_BuildStructure, Controleur/PathTracer_BVH.cpp:12-356: single-threaded, 336-byte swaps)
// followed by the host collapse of pt_wide_build.cu. See pt_lbvh.cu.
#pragma once

#include <cstdint>
#include <string>

#include <cuda_runtime.h>

#include "pt_device.cuh"
#include "pt_wide.cuh"

namespace ptc
{
struct DeviceWideBvh  // device allocations (cudaMalloc), owned by the caller after a successful build
{
    DevNode8 *nodes = nullptr;
    DevTri *tris8 = nullptr;         // 24 records per node
    uint32_t *tri8Index = nullptr;   // 24 per node: index into the caller's triangle array, kNoTriangle = unused
    uint32_t nNodes = 0;
    uint32_t depth = 0;              // levels of wide nodes
    float msBuild = 0.f;             // device time of the whole build (CUDA events)
};

// verts: 9 floats per triangle (S1, S2, S3 xyz) on the device; tris: the nT intersection records in the same order.
// Builds a Morton-order LBVH over the triangles, collapses it with the shared dynamic programme and emits the wide
// nodes. false + message on failure (nothing is left allocated).
bool build_wide_bvh_device(cudaStream_t stream, const float *verts, const DevTri *tris, uint32_t nT, DeviceWideBvh &out, std::string &err);
}  // namespace ptc
