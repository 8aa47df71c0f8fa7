// main.cpp — headless stand-alone front-end, the counterpart of reference Alone/main.cpp:9-34
// (same call sequence: PathTracer_SetImporter + PathTracer_Main; parameters come from the command line
// instead of being hard-coded, and the importer is the procedural one).
//
// usage: pathtracer_alone [scene] [width] [height] [numImages] [rayMaxDepth] [sampler 0|1|2] [mode 0|1] [save 0|1] [outdir] [exportdir]
//   environment: PATHTRACER_DEVICES = n (the first n CUDA devices) or "all": one context over several GPUs (ptc_create_multi)
//   scene      a procedural scene name (cornell, gallery, atrium[:n], terrain[:n]) or file:<folder> = the .pth triple
//              in <folder> (PathTracer_FileImporter.h, what the reference's PathTracerFileImporter reads)
//   exportdir  if given, the imported scene is written there as sizes.pth / pointers.pth / textureData.pth before
//              rendering (the reference's exportScene flag, Controleur/PathTracer.cpp:67-71)
#include <cstdlib>
#include <iostream>

#include <cstring>

#include "PathTracer.h"
#include "PathTracer_FileImporter.h"
#include "PathTracer_ProceduralImporter.h"

int main(int argc, char **argv)
{
    const char *scene = argc > 1 ? argv[1] : "cornell";
    unsigned image_width = argc > 2 ? std::atoi(argv[2]) : 1920;
    unsigned image_height = argc > 3 ? std::atoi(argv[3]) : 1080;
    unsigned numImageToRender = argc > 4 ? std::atoi(argv[4]) : 1;
    unsigned rayMaxDepth = argc > 5 ? std::atoi(argv[5]) : 10;
    PathTracerNS::Sampler sampler = (PathTracerNS::Sampler)(argc > 6 ? std::atoi(argv[6]) : 0);
    PathTracerNS::CUDA_Options opt;
    opt.mode = argc > 7 ? std::atoi(argv[7]) : 0;
    bool saveRenderedImages = argc > 8 ? std::atoi(argv[8]) != 0 : true;
    if (argc > 9) PathTracerNS::PathTracerDialog::exportFolderPath = argv[9];
    opt.readBackEvery = saveRenderedImages ? numImageToRender : 1;  // one picture at the end
    if (const char *d = std::getenv("PATHTRACER_DEVICES")) opt.nDevices = std::strcmp(d, "all") == 0 ? -1 : std::atoi(d);
    PathTracerNS::CUDA_SetOptions(opt);

    const bool exportScene = argc > 10;
    if (exportScene) PathTracerNS::PathTracer_SetSceneFolder(argv[10]);
    if (std::strncmp(scene, "file:", 5) == 0) PathTracerNS::PathTracer_SetImporter(new PathTracerNS::PathTracerFileImporter(scene + 5));
    else PathTracerNS::PathTracer_SetImporter(new PathTracerNS::PathTracerProceduralImporter(scene));
    bool success = PathTracerNS::PathTracer_Main(image_width, image_height, numImageToRender, saveRenderedImages, true, exportScene, sampler,
                                                 rayMaxDepth, false, false);
    return success ? 0 : 1;
}
