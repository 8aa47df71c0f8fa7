// PathTracer_Dialog.h — headless stand-in for the reference's display classes.
//
// Same class name and member functions as reference Alone/PathTracer_Dialog.h:16-60, so the controller code
// is unchanged. PaintWindow() does what the stand-alone reference does (Alone/PathTracer_Dialog.cpp:160-184):
// when saveRenderedImages is set, the running sum / count image is converted to 8 bit and written as
// NNN.bmp into exportFolderPath — with a portable writer instead of Win32 CreateFile/WriteFile.
#ifndef PATHTRACER_B200_DIALOG_H
#define PATHTRACER_B200_DIALOG_H

#include <string>
#include <vector>

#include "PathTracer_Types.h"

namespace PathTracerNS
{
// reference Alone/PathTracer_bitmap.cpp:237-286 and :146-201 (24-bit BMP, bottom-up rows as the reference
// writes them: row 0 of the image is the first row in the file)
std::vector<unsigned char> ConvertRGBAToBMPBuffer(RGBAColor const *imageColor, float const *imageRay, int width, int height);
bool SaveBMP(const std::vector<unsigned char> &buffer, int width, int height, const std::string &path);

class PathTracerDialog
{
public:
    PathTracerDialog() : pathTracerWidth(0), pathTracerHeight(0), imageIndex(0), saveRenderedImages(false) {}

    bool PaintWindow(RGBAColor const *imageColor, float const *imageRay);

    void SetWidth(uint w) { pathTracerWidth = w; }
    void SetHeight(uint h) { pathTracerHeight = h; }
    void TestPaintWindow() {}
    void SetNumberOfTriangles(uint) {}
    void SetSizeOfBVH(uint) {}
    void SetMaxDepthOfBVH(uint) {}
    void SetRaysPerPixel(uint) {}
    void SetPixelPerSeconds(uint) {}
    void SetRaysPerSeconds(uint) {}
    void UpdateElapsedTime() {}
    void Create() {}
    void OnClose() {}

    uint pathTracerWidth;
    uint pathTracerHeight;
    uint imageIndex;
    bool saveRenderedImages;

    static std::string exportFolderPath;  // the reference hard-codes a Windows folder (PathTracer_PreProc.h:9-12)
};
}  // namespace PathTracerNS

#endif
