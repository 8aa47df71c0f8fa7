"""B200-native hot path of adjerbetian/OpenCL_PathTracer: wavefront path tracing kernels for sm_100a behind the
reference's own device boundary (see include/pathtracer_cuda.h)."""
from ._native import (FLAG_BVH2, FLAG_GPU_BVH, FLAG_NO_GRAPH, FLAG_NO_VARIANCE, FLAG_RUSSIAN_ROULETTE, FLAG_STAGE_TIMING, FLAG_TEXTURE_BILINEAR, FLAG_WORK_COUNTERS, MODE_FAST, MODE_PARITY, SAMPLER_JITTERED, SAMPLER_RANDOM,
                      SAMPLER_UNIFORM)
from .pathtracer import PathTracer_Main, PathTracerCUDA, PathTracerError, run_default_drop_in
from .scene import Scene, area_light

__all__ = ["PathTracerCUDA", "PathTracerError", "PathTracer_Main", "run_default_drop_in", "Scene", "area_light", "MODE_PARITY", "MODE_FAST", "SAMPLER_JITTERED",
           "SAMPLER_RANDOM", "SAMPLER_UNIFORM", "FLAG_STAGE_TIMING", "FLAG_NO_VARIANCE", "FLAG_NO_GRAPH", "FLAG_BVH2", "FLAG_GPU_BVH", "FLAG_RUSSIAN_ROULETTE", "FLAG_TEXTURE_BILINEAR", "FLAG_WORK_COUNTERS"]
