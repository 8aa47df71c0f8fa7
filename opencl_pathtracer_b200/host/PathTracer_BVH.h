// PathTracer_BVH.h — host BVH construction with the reference's interface.
//
// Same entry point and same output contract as reference Controleur/PathTracer_BVH.h:11 /
// Controleur/PathTracer_BVH.cpp:12-37: fills globalVars.bvh (depth-first numbering, left child =
// parent + 1), bvhSize, bvhMaxDepth, and permutes globalVars.triangulation into BVH order.
// The node numbering and the triangle permutation are IDENTICAL to the reference builder's (checked
// against the reference's own PathTracer_BVH.cpp in tests/test_bvh_parity.py), because the closest-hit
// tie-break of the kernel is traversal order.
#ifndef PATHTRACER_B200_BVH_H
#define PATHTRACER_B200_BVH_H

#include "PathTracer_Types.h"

namespace PathTracerNS
{
void BVH_Create(GlobalVars &globalVars);

// Diagnostics, reference Controleur/PathTracer_BVH.cpp:362-401 (same meaning, plain struct instead of 7 refs).
struct BVHCharacteristics
{
    uint maxLeafSize = 0, minLeafSize = 0xffffffffu, maxDepth = 0, minDepth = 0xffffffffu, nNodes = 1, nLeafs = 0;
};
BVHCharacteristics BVH_GetCharacteristics(const Node *bvh);
}  // namespace PathTracerNS

#endif
