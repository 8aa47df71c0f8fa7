// pt_wide_build.h — host builder of the FAST-mode 8-wide compressed BVH (see pt_wide_build.cu, pt_wide.cuh).
#pragma once

#include <cstdint>
#include <string>
#include <vector>

#include "pathtracer_cuda.h"
#include "pt_wide.cuh"

namespace ptc
{
struct WideBvh
{
    std::vector<DevNode8> nodes;     // nodes[0] is the root
    std::vector<uint32_t> triOrder;  // 24 records per node (slot-major, 3 per slot) -> index in the BVH-ordered triangle array, kNoTriangle = unused
    uint32_t depth = 0;              // levels of wide nodes
};

// false + message on malformed input. An empty scene gives an empty result.
bool build_wide_bvh(const ptc_scene_view *scene, WideBvh &out, std::string &err);
}  // namespace ptc
