// PathTracer_FileImporter.h — the ".pth" scene files of the reference, with a defined layout.
//
// Counterpart of reference Controleur/PathTracer_FileImporter.{h,cpp}:15-146: an importer that reads a scene from
// three files and an Export() that writes the current scene into them, so that scenes can enter the path tracer
// without Maya / LumenRT and a scene exported once can be re-rendered. The reference dumps its structs raw (MSVC x64
// layout, a host pointer inside Material) into a folder hard-coded at build time. Here:
//   * the folder is an argument;
//   * the byte layout is the reference's, which include/pathtracer_structs.h pins with static_asserts (Triangle 336 B,
//     Light 64 B, Material 48 B, Texture 12 B, Sky 92 B, Float4 16 B), so files written by a reference build load here
//     and vice versa; little-endian, no padding between arrays;
//   * Material::textureName (a char* in the reference, meaningless in a file) is written as 0 and ignored on read;
//   * short or missing files are errors (the reference ignores fread's result).
//
//   sizes.pth        cameraDirection, cameraRight, cameraUp, cameraPosition (4 x Float4), then triangulationSize,
//                    lightsSize, materiauxSize, texturesSize, texturesDataSize (5 x uint32)          = 84 bytes
//   pointers.pth     Triangle[triangulationSize], Light[lightsSize], Material[materiauxSize],
//                    Texture[texturesSize], Sky
//   textureData.pth  Uchar4[texturesDataSize]
#ifndef PATHTRACER_B200_FILEIMPORTER_H
#define PATHTRACER_B200_FILEIMPORTER_H

#include <string>

#include "PathTracer_Importer.h"

namespace PathTracerNS
{
class PathTracerFileImporter : public PathTracerImporter
{
public:
    explicit PathTracerFileImporter(const std::string &folder);

    // reference :67-146. Throws std::runtime_error naming the files that cannot be read (as the reference does) or
    // that are shorter than sizes.pth announces. loadSky = false resets the sky to a 1x1 black cube as the reference does.
    void Import(uint image_width, uint image_height, bool loadSky) override;

    // reference :15-63, after Initialize(globalVars): writes the scene the aliases point to. Throws on I/O failure.
    void Export() const;

    std::string fileSizesPath, filePointersPath, fileTextureDataPath;
};
}  // namespace PathTracerNS

#endif
