// PathTracer_FileImporter.cpp — see PathTracer_FileImporter.h. Format and call sequence follow reference
// Controleur/PathTracer_FileImporter.cpp:15-146; the code is this repository's own (portable stdio, checked reads).
#include "PathTracer_FileImporter.h"

#include <cstdio>
#include <cstring>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <vector>

namespace PathTracerNS
{
namespace
{
struct FileCloser
{
    void operator()(FILE *f) const { if (f) std::fclose(f); }
};
typedef std::unique_ptr<FILE, FileCloser> File;

std::string joinPath(const std::string &folder, const char *name)
{
    if (folder.empty()) return name;
    const char last = folder[folder.size() - 1];
    return (last == '/' || last == '\\') ? folder + name : folder + "/" + name;
}

void readExact(FILE *f, void *dst, size_t size, size_t count, const std::string &path, const char *what)
{
    if (count == 0) return;
    if (std::fread(dst, size, count, f) != count)
    {
        std::ostringstream m;
        m << path << " is shorter than announced: " << count << " x " << what << " (" << size << " bytes each)";
        throw std::runtime_error(m.str());
    }
}

void writeExact(FILE *f, const void *src, size_t size, size_t count, const std::string &path)
{
    if (count == 0) return;
    if (std::fwrite(src, size, count, f) != count) throw std::runtime_error("Fail to write " + path);
}
}  // namespace

PathTracerFileImporter::PathTracerFileImporter(const std::string &folder)
    : fileSizesPath(joinPath(folder, "sizes.pth")),
      filePointersPath(joinPath(folder, "pointers.pth")),
      fileTextureDataPath(joinPath(folder, "textureData.pth"))
{
}

void PathTracerFileImporter::Export() const
{
    File sizes(std::fopen(fileSizesPath.c_str(), "wb")), ptrs(std::fopen(filePointersPath.c_str(), "wb")),
        tex(std::fopen(fileTextureDataPath.c_str(), "wb"));
    if (!sizes || !ptrs || !tex)
    {
        std::ostringstream m;
        m << "Fail to open the files to export :";
        if (!sizes) m << "\n\t" << fileSizesPath;
        if (!ptrs) m << "\n\t" << filePointersPath;
        if (!tex) m << "\n\t" << fileTextureDataPath;
        throw std::runtime_error(m.str());
    }
    // the reference's field order: direction, right, up, position (:27-30)
    writeExact(sizes.get(), ptr__global__cameraDirection, sizeof(Float4), 1, fileSizesPath);
    writeExact(sizes.get(), ptr__global__cameraRight, sizeof(Float4), 1, fileSizesPath);
    writeExact(sizes.get(), ptr__global__cameraUp, sizeof(Float4), 1, fileSizesPath);
    writeExact(sizes.get(), ptr__global__cameraPosition, sizeof(Float4), 1, fileSizesPath);
    const uint counts[5] = {*ptr__global__triangulationSize, *ptr__global__lightsSize, *ptr__global__materiauxSize, *ptr__global__texturesSize,
                            *ptr__global__texturesDataSize};
    writeExact(sizes.get(), counts, sizeof(uint), 5, fileSizesPath);

    writeExact(ptrs.get(), *ptr__global__triangulation, sizeof(Triangle), counts[0], filePointersPath);
    writeExact(ptrs.get(), *ptr__global__lights, sizeof(Light), counts[1], filePointersPath);
    {
        // Material::textureName is a host pointer: never written
        std::vector<Material> m(*ptr__global__materiaux, *ptr__global__materiaux + counts[2]);
        for (size_t i = 0; i < m.size(); i++) std::memset(&m[i].textureName, 0, sizeof(m[i].textureName));
        writeExact(ptrs.get(), m.data(), sizeof(Material), counts[2], filePointersPath);
    }
    writeExact(ptrs.get(), *ptr__global__textures, sizeof(Texture), counts[3], filePointersPath);
    writeExact(ptrs.get(), ptr__global__sky, sizeof(Sky), 1, filePointersPath);

    writeExact(tex.get(), *ptr__global__texturesData, sizeof(Uchar4), counts[4], fileTextureDataPath);
    if (std::fflush(sizes.get()) || std::fflush(ptrs.get()) || std::fflush(tex.get())) throw std::runtime_error("Fail to write the exported scene");
}

void PathTracerFileImporter::Import(uint image_width, uint image_height, bool loadSky)
{
    File sizes(std::fopen(fileSizesPath.c_str(), "rb")), ptrs(std::fopen(filePointersPath.c_str(), "rb")),
        tex(std::fopen(fileTextureDataPath.c_str(), "rb"));
    if (!sizes || !ptrs || !tex)
    {
        std::ostringstream m;  // message shape of the reference (:80-87)
        m << "Fail to read the files to import : \n";
        if (!sizes) m << "\t" << fileSizesPath << "\n";
        if (!ptrs) m << "\t" << filePointersPath << "\n";
        if (!tex) m << "\t" << fileTextureDataPath << "\n";
        throw std::runtime_error(m.str());
    }
    readExact(sizes.get(), ptr__global__cameraDirection, sizeof(Float4), 1, fileSizesPath, "Float4");
    readExact(sizes.get(), ptr__global__cameraRight, sizeof(Float4), 1, fileSizesPath, "Float4");
    readExact(sizes.get(), ptr__global__cameraUp, sizeof(Float4), 1, fileSizesPath, "Float4");
    readExact(sizes.get(), ptr__global__cameraPosition, sizeof(Float4), 1, fileSizesPath, "Float4");
    uint counts[5];
    readExact(sizes.get(), counts, sizeof(uint), 5, fileSizesPath, "uint");
    *ptr__global__triangulationSize = counts[0];
    *ptr__global__lightsSize = counts[1];
    *ptr__global__materiauxSize = counts[2];
    *ptr__global__texturesSize = counts[3];
    *ptr__global__texturesDataSize = counts[4];
    *ptr__global__imageWidth = image_width;
    *ptr__global__imageHeight = image_height;
    *ptr__global__imageSize = image_width * image_height;

    // (the arrays are owned by GlobalVars and released by PathTracer_Clear, as in the reference; at least one element each
    // so that zero-length arrays stay valid pointers)
    *ptr__global__triangulation = new Triangle[counts[0] ? counts[0] : 1];
    *ptr__global__lights = new Light[counts[1] ? counts[1] : 1];
    *ptr__global__materiaux = new Material[counts[2] ? counts[2] : 1];
    *ptr__global__textures = new Texture[counts[3] ? counts[3] : 1];
    *ptr__global__texturesData = new Uchar4[counts[4] ? counts[4] : 1];
    readExact(ptrs.get(), *ptr__global__triangulation, sizeof(Triangle), counts[0], filePointersPath, "Triangle");
    readExact(ptrs.get(), *ptr__global__lights, sizeof(Light), counts[1], filePointersPath, "Light");
    readExact(ptrs.get(), *ptr__global__materiaux, sizeof(Material), counts[2], filePointersPath, "Material");
    readExact(ptrs.get(), *ptr__global__textures, sizeof(Texture), counts[3], filePointersPath, "Texture");
    readExact(ptrs.get(), ptr__global__sky, sizeof(Sky), 1, filePointersPath, "Sky");
    readExact(tex.get(), *ptr__global__texturesData, sizeof(Uchar4), counts[4], fileTextureDataPath, "Uchar4");
    for (uint i = 0; i < counts[2]; i++) std::memset(&(*ptr__global__materiaux)[i].textureName, 0, sizeof((*ptr__global__materiaux)[i].textureName));

    if (!loadSky)
    {
        // :128-143: a black 1x1 cube. The reference also cuts texturesDataSize to 1, which would leave every textured material
        // pointing past the uploaded texels; here texel 0 is blackened and the data kept whenever a scene texture needs it.
        Sky &sky = *ptr__global__sky;
        sky.cosRotationAngle = 1;
        sky.sinRotationAngle = 0;
        sky.exposantFactorX = 0;
        sky.exposantFactorY = 0;
        sky.groundScale = 1;
        for (int i = 0; i < 6; i++) { sky.skyTextures[i].height = 1; sky.skyTextures[i].width = 1; sky.skyTextures[i].offset = 0; }
        if (counts[3] == 0)
        {
            *ptr__global__texturesDataSize = 1;
            (*ptr__global__texturesData)[0] = Uchar4(0, 0, 0, 0);
        }
        else
        {
            // keep the scene textures; the sky's single texel must not alias one of theirs: append it
            Uchar4 *grown = new Uchar4[(size_t)counts[4] + 1];
            std::memcpy(grown, *ptr__global__texturesData, sizeof(Uchar4) * counts[4]);
            grown[counts[4]] = Uchar4(0, 0, 0, 0);
            delete[] *ptr__global__texturesData;
            *ptr__global__texturesData = grown;
            *ptr__global__texturesDataSize = counts[4] + 1;
            for (int i = 0; i < 6; i++) sky.skyTextures[i].offset = counts[4];
        }
    }
}
}  // namespace PathTracerNS
