#include "PathTracer_Dialog.h"

#include <cstdio>
#include <cstring>

namespace PathTracerNS
{
std::string PathTracerDialog::exportFolderPath = "./";

std::vector<unsigned char> ConvertRGBAToBMPBuffer(RGBAColor const *imageColor, float const *imageRay, int width, int height)
{
    std::vector<unsigned char> out;
    if (!imageColor || width <= 0 || height <= 0) return out;
    const int scan = width * 3;
    const int psw = (scan + 3) & ~3;
    out.assign((size_t)psw * height, 0);
    size_t pos = 0;
    for (int y = 0; y < height; y++)
    {
        unsigned char *row = out.data() + (size_t)y * psw;
        for (int x = 0; x < width; x++, pos++)
        {
            const float n = imageRay ? imageRay[pos] : 1.f;
            int r, g, b;
            if (imageColor[pos].x < 0) { r = 255; g = 0; b = 0; }
            else
            {
                float fr = imageColor[pos].x * 255.f / n, fg = imageColor[pos].y * 255.f / n, fb = imageColor[pos].z * 255.f / n;
                r = (int)(255.f < fr ? 255.f : fr);
                g = (int)(255.f < fg ? 255.f : fg);
                b = (int)(255.f < fb ? 255.f : fb);
            }
            row[3 * x + 0] = (unsigned char)b;
            row[3 * x + 1] = (unsigned char)g;
            row[3 * x + 2] = (unsigned char)r;
        }
    }
    return out;
}

bool SaveBMP(const std::vector<unsigned char> &buffer, int width, int height, const std::string &path)
{
    unsigned char hdr[54];
    std::memset(hdr, 0, sizeof(hdr));
    auto put32 = [&](int off, unsigned v) { hdr[off] = v & 255; hdr[off + 1] = (v >> 8) & 255; hdr[off + 2] = (v >> 16) & 255; hdr[off + 3] = (v >> 24) & 255; };
    hdr[0] = 'B'; hdr[1] = 'M';
    put32(2, 54u + (unsigned)buffer.size());
    put32(10, 0x36);
    put32(14, 40);
    put32(18, (unsigned)width);
    put32(22, (unsigned)height);
    hdr[26] = 1;
    hdr[28] = 24;
    put32(38, 0x0ec4);
    put32(42, 0x0ec4);
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    bool ok = std::fwrite(hdr, 1, 54, f) == 54 && std::fwrite(buffer.data(), 1, buffer.size(), f) == buffer.size();
    std::fclose(f);
    return ok;
}

bool PathTracerDialog::PaintWindow(RGBAColor const *imageColor, float const *imageRay)
{
    if (saveRenderedImages || imageIndex == 500)
    {
        char name[32];
        std::snprintf(name, sizeof(name), "%03u.bmp", imageIndex);
        std::vector<unsigned char> buf = ConvertRGBAToBMPBuffer(imageColor, imageRay, (int)pathTracerWidth, (int)pathTracerHeight);
        if (!SaveBMP(buf, (int)pathTracerWidth, (int)pathTracerHeight, exportFolderPath + name))
        {
            std::fprintf(stderr, "WARNING : unable to save rendered picture in the folder \"%s\". Aborting saving.\n", exportFolderPath.c_str());
            saveRenderedImages = false;
        }
    }
    imageIndex++;
    return true;
}
}  // namespace PathTracerNS
