// PathTracer_CUDA.h — C++ shims with the signatures of the reference's device boundary.
//
// Drop-in for reference Controleur/PathTracer_OpenCL.h:17-19: the three functions PathTracer_Main calls
// (Controleur/PathTracer.cpp:73-89) keep their exact signatures and semantics; they are backed by the C ABI
// of include/pathtracer_cuda.h instead of an OpenCL context/queue. Errors surface as std::runtime_error,
// as OpenCL_ErrorHandling does (Controleur/PathTracer_OpenCL.cpp:407-486).
#ifndef PATHTRACER_B200_CUDA_SHIMS_H
#define PATHTRACER_B200_CUDA_SHIMS_H

#include "PathTracer_Types.h"

namespace PathTracerNS
{
void CUDA_SetupContext(GlobalVars &globalVars, Sampler sampler);
void CUDA_InitializeMemory(GlobalVars &globalVars);
void CUDA_RunKernel(GlobalVars &globalVars, bool (*UpdateWindowFunc)(void), uint numImagesToRender, double *pathTracingTime,
                    double *memoryTime, double *displayTime);

// Knobs the reference does not have (all optional; defaults reproduce the reference loop: PARITY arithmetic,
// read-back + callback after every iteration, device 0).
struct CUDA_Options
{
    int device = 0;
    // Several GPUs of the box behind the same three calls (ptc_create_multi): nDevices > 1 selects devices[0 .. nDevices); -1 = every
    // CUDA device present. The devices render iteration ranges (or pixel stripes when a read-back interval holds fewer iterations
    // than there are devices); CUDA_RunKernel's read-backs sum them with an NCCL reduce.
    int nDevices = 0;
    int devices[16] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15};
    int mode = 0;                 // ptc_mode
    uint readBackEvery = 1;       // iterations between read-back + UpdateWindowFunc (the last iteration always reads back)
    uint iterationsPerBatch = 0;  // 0 = auto
    uint firstIteration = 0;      // iteration range offset (multi-GPU sharding by iteration)
    uint flags = 0;               // PTC_FLAG_*
};
void CUDA_SetOptions(const CUDA_Options &options);
const CUDA_Options &CUDA_GetOptions();
}  // namespace PathTracerNS

#endif
