// PathTracer_ProceduralImporter.h — deterministic synthetic scenes behind the reference's importer API.
//
// The reference has no sample scene; its importers need the Maya / LumenRT SDKs. This
// importer generates the benchmark and parity scenes of BASELINE.json from a name and a seed, following
// the Maya importer's data conventions so that the kernel sees what it would see from Maya
// (reference Maya/PathTracer_MayaImporter.cpp):
//   * points carry w = 1, vectors w = 0 (:29-33 of the importer header);
//   * Triangle.N = normalize(cross(s2-s1, s3-s1)) computed BEFORE the vertices are sorted, with the host
//     cross() returning w = 1, so N.w = 1 (:959);
//   * vertices (and their normals / UVs) stored in lexicographic order (:961-1021);
//   * vertex normals shorter than 0.5 replaced by N, others normalised (:1031-1037);
//   * loadSky = false gives six 1x1 sky faces over one (0,0,0,0) texel (:703-719);
//   * z is up (the Maya importer permutes xyz -> zxy).
//
// Scenes ("name" or "name:param"):
//   cornell            30 triangles, diffuse, one point light under the ceiling        (configs C1, C2, C5)
//   cornell_specular   same, tall block glass, short block varnished
//   gallery            ~1.5 k triangles exercising every kernel branch: all material types, textures,
//                      sky cube map, directional + point + spot light, water
//   atrium             ~262 k triangles, diffuse + varnished + glass, 2 point + 1 directional (config C3)
//   atrium:k           the same at 1/k of the tessellation (k = 2, 4, 8 ...) for fast tests
//   terrain            ~10 M triangles: 64x64 flattened instances of a displaced rock over a displaced
//                      terrain, 16 textures 1024^2, sun + sky cube map                      (config C4)
//   terrain:g          g x g instances and a proportionally smaller terrain / textures
#ifndef PATHTRACER_B200_PROCEDURAL_IMPORTER_H
#define PATHTRACER_B200_PROCEDURAL_IMPORTER_H

#include <string>

#include "PathTracer_Importer.h"

namespace PathTracerNS
{
class PathTracerProceduralImporter : public PathTracerImporter
{
public:
    explicit PathTracerProceduralImporter(const std::string &scene, uint seed = 20261104u) : sceneName(scene), seed(seed) {}
    void Import(uint image_width, uint image_height, bool loadSky) override;

    std::string sceneName;
    uint seed;
};

// Maya-convention triangle constructor (see above); exposed for tests.
void Triangle_Create(Triangle *This, Float4 const &s1, Float4 const &s2, Float4 const &s3, Float2 const &uvp1,
                     Float2 const &uvp2, Float2 const &uvp3, Float2 const &uvn1, Float2 const &uvn2, Float2 const &uvn3,
                     Float4 const &n1, Float4 const &n2, Float4 const &n3, uint positiveMatIndex, uint negativeMatIndex,
                     uint id);
bool Triangle_isValid(Float4 const &s1, Float4 const &s2, Float4 const &s3);
}  // namespace PathTracerNS

#endif
