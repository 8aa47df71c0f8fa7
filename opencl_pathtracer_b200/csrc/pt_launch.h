// pt_launch.h — host-callable launchers of the wavefront stages (defined in pt_kernels.cu).
#pragma once

#include "pt_device.cuh"

namespace ptc
{
struct LaunchCfg
{
    int traverseGrid;  // persistent extend CTAs (multiple of the SM count)
    int connectGrid;   // persistent connect CTAs
    int streamGrid;    // grid of the streaming stages (generate / shade / accumulate)
    uint32_t refillMin;  // idle lanes of a warp that trigger a refill from the work cursor (1..32)
    int stackDepth;      // traversal stack entries per thread (BVH depth + 1, at most kStackDepth)
    bool wide;           // FAST mode: traverse the 8-wide compressed BVH (stackDepth then counts 8-byte entries)
    bool partition;      // k_classify partitions every bounce's queue by class; connect and shade take classes from positions (decided per
                         // scene at upload: pays where shade is instruction bound, costs where it is DRAM bound — k_shade in pt_kernels.cu)
    bool partitionConnect;  // the same pass for the connect stage alone (shade keeps the bounce's own queue): groups of 32 LIT paths per staging
    bool count;          // per-ray work counters + the two 5000-bin histograms (PARITY always; FAST with PTC_FLAG_WORK_COUNTERS)
};

size_t traverse_shared_bytes(bool wide, int stackDepth, bool any);  // dynamic shared memory of one traversal CTA (stacks + ray staging)
int traverse_ctas_per_sm(bool fast, bool wide, int stackDepth, bool any);  // resident CTAs per SM of the extend (any = false) / connect kernel

void launch_reset(cudaStream_t s, Counters *cn, uint32_t nPaths);
void launch_generate(cudaStream_t s, const LaunchCfg &lc, const FrameParams &fp, const Camera &cam, const PathState &st);
// `-D SUPER_SAMPLING`: drops paths by the reference's variance criterion and builds the queue of bounce 0 (launch_reset with 0 paths first)
void launch_supersample(cudaStream_t s, const LaunchCfg &lc, const FrameParams &fp, const PathState &st, const ImageDev &img,
                        const float *x2inv, uint32_t x2invSize, uint32_t *queue0, Counters *cn);
void launch_extend(cudaStream_t s, const LaunchCfg &lc, bool fast, const SceneDev &sc, const PathState &st, const uint32_t *queue,
                   Counters *cn, int bounce);
void launch_connect(cudaStream_t s, const LaunchCfg &lc, bool fast, const SceneDev &sc, const PathState &st, const uint32_t *queue,
                    Counters *cn, int bounce);
// re-writes the queue of the bounce into PathState::qpart with every 1024-entry tile partitioned lit / other (after extend: it reads
// the flags extend left)
void launch_classify(cudaStream_t s, const LaunchCfg &lc, const PathState &st, const uint32_t *queue, Counters *cn, int bounce);
// shades the bounce's queue — qin, or PathState::qpart when LaunchCfg::partition — in queue order and appends the survivors to qout
void launch_shade(cudaStream_t s, const LaunchCfg &lc, bool fast, const SceneDev &sc, const FrameParams &fp, const PathState &st,
                  const uint32_t *qin, uint32_t *qout, Counters *cn, int bounce);
void launch_accumulate(cudaStream_t s, const LaunchCfg &lc, const FrameParams &fp, const PathState &st, const ImageDev &img,
                       Counters *cn);
// batch graphs: host stubs of k_generate(fp, cam, st) and k_accumulate(fp, st, img, cn), to find their nodes in a captured graph
const void *generate_kernel();
const void *accumulate_kernel(uint32_t sampler);
// parity probes
void launch_probe_primary(cudaStream_t s, const FrameParams &fp, const PathState &st, int32_t *hit, float2 *stOut, float *dist2,
                          float4 *dir, int32_t *seedAfter, uint2 *counts);
void launch_probe_rays(cudaStream_t s, uint32_t n, const float4 *origin, const float4 *direction, const PathState &st);
// roofline denominators (ptc_measure_read_bandwidth)
void launch_read_stream(cudaStream_t s, int grid, const void *buf, size_t bytes, uint32_t repeats, uint32_t *sink);
// display epilogue: sum / count -> clamp -> 8 bit (reference Alone/PathTracer_bitmap.cpp:259-279)
void launch_rgb8(cudaStream_t s, const ImageDev &img, const FrameParams &fp, uint8_t *rgb);
void launch_clear(cudaStream_t s, const LaunchCfg &lc, const ImageDev &img, uint32_t nPixels, Counters *cn);
void launch_add_image(cudaStream_t s, const ImageDev &img, uint32_t nPixels, const float4 *color, const float *rayNb);

}  // namespace ptc
