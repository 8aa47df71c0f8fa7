/* pathtracer_cuda.h — PathTracer_CUDA: the C ABI of the B200-native path-tracing hot path.
 *
 * Drop-in seam: these entry points replace the three calls PathTracer_Main makes into the
 * reference's OpenCL layer (reference Controleur/PathTracer_OpenCL.h:17-19, called from
 * Controleur/PathTracer.cpp:73-89):
 *
 *     OpenCL_SetupContext(globalVars, sampler)      ->  ptc_create
 *     OpenCL_InitializeMemory(globalVars)           ->  ptc_upload_scene
 *     OpenCL_RunKernel(globalVars, cb, n, &t,&t,&t) ->  { ptc_render; ptc_read_image; cb } x n,
 *                                                       ptc_read_stats, ptc_destroy
 *
 * C++ shims with the reference's exact signatures (CUDA_SetupContext / CUDA_InitializeMemory /
 * CUDA_RunKernel) live in opencl_pathtracer_b200/host/PathTracer_CUDA.h; INTEGRATION.md shows the
 * three-line change in PathTracer.cpp.
 *
 * Conventions: every function returns a ptc_status (0 = success); no exceptions, no STL and no
 * CUDA/torch types cross this boundary; host arrays are borrowed only for the duration of the call.
 * A context is owned by one host thread at a time. There is NO CPU fallback: without a CUDA device
 * ptc_create fails with PTC_ERR_NO_DEVICE.
 */
#ifndef PATHTRACER_CUDA_H
#define PATHTRACER_CUDA_H

#include <stddef.h>
#include <stdint.h>

#include "pathtracer_structs.h"

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define PTC_API __declspec(dllexport)
#else
#define PTC_API __attribute__((visibility("default")))
#endif

#define PTC_MAX_INTERSECTION_NUMBER 5000u /* reference Controleur/PathTracer_PreProc.h:17 */
#define PTC_BVH_MAX_DEPTH 30u             /* reference Controleur/PathTracer_PreProc.h:18, kernel header.cl:16 */
#define PTC_MAX_LIGHT_SIZE 30u            /* reference Controleur/PathTracer_PreProc.h:19 */

typedef struct ptc_ctx ptc_ctx;

typedef enum ptc_status
{
    PTC_OK = 0,
    PTC_ERR_INVALID = 1,     /* bad argument / inconsistent scene (message says which) */
    PTC_ERR_CUDA = 2,        /* a CUDA runtime call failed (message carries cudaGetErrorString) */
    PTC_ERR_NO_DEVICE = 3,   /* no usable CUDA device: this library has no CPU path */
    PTC_ERR_STATE = 4,       /* call order violated (e.g. render before upload) */
    PTC_ERR_LIMIT = 5,       /* reference limits exceeded: BVH depth >= 30 or lights >= 30
                                (reference Controleur/PathTracer.cpp:54-65 throws for the same) */
    PTC_ERR_UNSUPPORTED = 6  /* input outside the documented contract (see DESIGN.md "w lanes") */
} ptc_status;

/* reference Controleur/PathTracer_Structs.h:130-135 (enum Sampler), same numbering */
typedef enum ptc_sampler { PTC_SAMPLER_JITTERED = 0, PTC_SAMPLER_RANDOM = 1, PTC_SAMPLER_UNIFORM = 2 } ptc_sampler;

/* PARITY: reference traversal order, culling quirks and IEEE arithmetic (bit-exact hits).
 * FAST:   same estimator, free traversal order / correct culling / FMA (checked by image RMSE). */
typedef enum ptc_mode { PTC_MODE_PARITY = 0, PTC_MODE_FAST = 1 } ptc_mode;

enum
{
    PTC_FLAG_STAGE_TIMING = 1u << 0, /* bracket every stage launch with CUDA events (ptc_read_counters) */
    PTC_FLAG_NO_VARIANCE = 1u << 1,  /* do not maintain imageV (only the dead supersampling path reads it) */
    PTC_FLAG_NO_GRAPH = 1u << 2,     /* launch stages directly instead of replaying a captured CUDA graph */
    PTC_FLAG_BVH2 = 1u << 3,         /* FAST mode: traverse the reference BVH2 itself instead of the 8-wide compressed BVH
                                        that ptc_upload_scene collapses it into (PARITY always walks the BVH2) */
    PTC_FLAG_GPU_BVH = 1u << 4,      /* FAST mode: ignore the caller's BVH (ptc_scene_view.bvh may be NULL / bvh_size 0, so
                                        BVH_Create need not run at all) and build the 8-wide BVH on the device from the
                                        triangles in the order given (csrc/pt_lbvh.cu). Hits are reported as indices into
                                        that order. Not valid in PARITY mode or together with PTC_FLAG_BVH2. */
    PTC_FLAG_RUSSIAN_ROULETTE = 1u << 5,/* FAST mode: after MIN_REFLECTION_NUMBER = 5 bounces a path survives with probability
                                        c = max throughput / (bounce - 5) and its throughput is divided by c (unbiased). The
                                        reference has this stage disabled and commented out (Kernel/PathTracer_FullKernel_header.cl:12,
                                        PathTracer_FullKernel.cl:1306-1314); the random streams then differ from the reference's. */
    PTC_FLAG_WORK_COUNTERS = 1u << 7, /* FAST mode: keep the reference's per-ray work counters (Ray3D::numIntersectedBBx / numIntersectedTri,
                                        Kernel/PathTracer_FullKernel_header.cl:95-96) and the two 5000-bin histograms they feed
                                        (PathTracer_FullKernel.cl:1319-1331). PARITY always keeps them (they are the reference's, bit for
                                        bit). In FAST they describe the 8-wide walk, depend on the warp a ray shares and are compared with
                                        nothing; without this flag FAST drops them (two registers and two atomics per ray in the
                                        traversal kernels): ptc_read_stats then returns rayDepths and two all-zero work histograms.
                                        Global totals (ptc_counters.box_tests / triangle_tests) and the parity probes' per-ray counts
                                        (ptc_trace_primary / ptc_trace_rays) are always kept. */
    PTC_FLAG_TEXTURE_BILINEAR = 1u << 6 /* FAST mode: surface textures are sampled with bilinear interpolation of the four texels
                                        around (u, v) instead of the reference's nearest texel
                                        (Texture_GetPixelColorValue, Kernel/PathTracer_FullKernel_header.cl:430-459); same wrap
                                        and alpha convention. The sky keeps the reference's lookup. Un-oracled. */
};

/* Replaces the `-D` build options of reference Controleur/PathTracer_OpenCL.cpp:292-314: all run-time here. */
typedef struct ptc_config
{
    uint32_t struct_size;          /* = sizeof(ptc_config), for forward compatibility */
    int32_t device;                /* CUDA device ordinal */
    uint32_t image_width;          /* IMAGE_WIDTH */
    uint32_t image_height;         /* IMAGE_HEIGHT */
    uint32_t ray_max_depth;        /* MAX_REFLECTION_NUMBER */
    uint32_t sampler;              /* ptc_sampler */
    uint32_t mode;                 /* ptc_mode */
    uint32_t iterations_per_batch; /* samples per pixel kept in flight per wavefront pass; 0 = auto (>= 16M paths, 112 B each) */
    uint32_t flags;                /* PTC_FLAG_* */
    uint32_t reserved;
    void *stream;                  /* cudaStream_t to launch on; NULL = the library creates its own */
} ptc_config;

/* What OpenCL_InitializeMemory reads out of GlobalVars (reference Controleur/PathTracer_OpenCL.cpp:165-183). */
typedef struct ptc_scene_view
{
    uint32_t struct_size;
    uint32_t bvh_size;
    uint32_t triangulation_size;
    uint32_t lights_size;
    uint32_t materiaux_size;
    uint32_t textures_size;
    uint32_t textures_data_size;
    uint32_t reserved;
    const ptc_node *bvh;               /* bvh[0] is the root */
    const ptc_triangle *triangulation; /* in BVH order (after BVH_Create permuted it) */
    const ptc_light *lights;
    const ptc_material *materiaux;
    const ptc_texture *textures;
    const ptc_uchar4 *textures_data;   /* sky faces and scene textures, one flat array */
    const ptc_sky *sky;
    ptc_float4 camera_position;
    ptc_float4 camera_direction;
    ptc_float4 camera_right;
    ptc_float4 camera_up;
} ptc_scene_view;

/* Work and time counters, cumulative since ptc_create / ptc_reset_accumulation. */
typedef struct ptc_counters
{
    uint64_t paths;             /* (pixel, iteration) samples finished */
    uint64_t extension_rays;    /* closest-hit traversals  (BVH_IntersectRay calls) */
    uint64_t shadow_rays;       /* any-hit traversals      (BVH_IntersectShadowRay calls) */
    uint64_t box_tests;         /* BoundingBox_Intersects calls, all rays */
    uint64_t triangle_tests;    /* Triangle_Intersects calls, all rays */
    uint64_t kernel_launches;   /* kernels launched by this library */
    double ms_generate;         /* per-stage device time, only with PTC_FLAG_STAGE_TIMING */
    double ms_extend;
    double ms_connect;
    double ms_shade;
    double ms_accumulate;
    uint64_t launches_extend;
    uint64_t launches_connect;
    uint64_t scene_bytes_nodes;     /* device bytes of the flattened BVH */
    uint64_t scene_bytes_triangles; /* device bytes of the intersection records */
    uint64_t ext_box_tests;         /* share of box_tests / triangle_tests spent by closest-hit traversals */
    uint64_t ext_triangle_tests;
    /* CHECKED build only (libpathtracer_cuda.so compiled with -DPTC_CHECKED, ptc_version() ends in "checked"): bit c set = bounds check c
       of the kernels failed (queue entries, node / triangle / material / texture / texel indices, stack levels, target pixels:
       csrc/pt_device.cuh FaultCode); code / value / bound of the first failure. Always 0 in the normal build. */
    uint32_t faults, fault_code, fault_value, fault_bound;
} ptc_counters;

/* Device pointers of the accumulators, for zero-copy collectives (NCCL reduce of per-GPU images). */
typedef struct ptc_device_buffers
{
    void *image_color;  /* float4[W*H]  running sum      (reference global__imageColor) */
    void *image_ray_nb; /* float [W*H]  running count    (reference global__imageRayNb) */
    void *image_v;      /* float4[W*H]  variance accumulator (reference global__imageV) or NULL */
    void *ray_depths;   /* uint32[ray_max_depth+1] */
    void *ray_bbx;      /* uint32[5000] */
    void *ray_tri;      /* uint32[5000] */
    void *stream;       /* cudaStream_t all of the above are written on */
    void *rgb8;         /* uint8[3*W*H] display frame, filled by ptc_convert_rgb8 / ptc_read_rgb8 */
} ptc_device_buffers;

/* OpenCL_SetupContext: pick the device, allocate image/statistics accumulators (zeroed — the reference
 * never zeroes them) and the wavefront state. */
PTC_API int ptc_create(const ptc_config *config, ptc_ctx **out_ctx);

/* Multi-GPU: ONE context that drives `n_devices` CUDA devices of this box from the caller's single host thread —
 * so the reference's own caller (PathTracer_Main -> OpenCL_RunKernel, Controleur/PathTracer.cpp:73-89) scales to the whole box
 * through the same three calls. `config->device` is ignored, `config->stream` must be NULL. Every ptc_* call accepts the context:
 *   ptc_upload_scene   flattens and uploads once, then copies the device scene to the other devices (peer copies over NVLink);
 *   ptc_render(f, n)   n >= n_devices: device g renders the g-th share of the iterations [f, f + n) (shares differ by at most one,
 *                      earlier devices take the extras) over the whole frame; n < n_devices (a progressive frame): every device
 *                      renders all n iterations of its interleaved 8-row pixel stripes. The random stream of a path depends on
 *                      (x, y, iteration, W, H) only (Kernel/PathTracer_FullKernel_header.cl:255-264): either way exactly the
 *                      paths of a one-device render are traced, each once;
 *   ptc_read_image / ptc_read_stats / ptc_get_device_buffers   first sum the devices' accumulators into the first device
 *                      (ncclReduce of imageColor, imageRayNb and the three uint32 histograms; NCCL is bound at run time with dlopen,
 *                      PTC_MULTI_REDUCE=p2p selects peer copies + an add kernel instead) and clear the other devices';
 *   ptc_read_rgb8 / ptc_convert_rgb8   after stripes-only renders: every device converts its own rows, one uint8 reduce composes
 *                      the frame (3 bytes per pixel over NVLink instead of 20); otherwise as ptc_read_image + conversion;
 *   ptc_read_counters  sums over the devices; the parity probes and ptc_add_image act on the first device.
 * imageV is not maintained (not linearly summable): ptc_read_variance and ptc_set_supersampling refuse; ptc_set_pixel_stripes
 * refuses (the context shards by itself). Sums are made in another order than one device makes them: counts and histograms are
 * exactly those of a one-device render, colours agree to an RMSE of 1e-4 of the mean.
 * n_devices == 1 is ptc_create on devices[0]. */
PTC_API int ptc_create_multi(const ptc_config *config, const int32_t *devices, uint32_t n_devices, ptc_ctx **out_ctx);
/* Devices of a context (1 for ptc_create); writes up to `capacity` ordinals, returns the count. */
PTC_API int ptc_context_devices(const ptc_ctx *ctx, int32_t *devices, uint32_t capacity);
/* "nccl", "p2p" (peer copies + add kernel) or "none" (one device): how a context sums its devices' accumulators. */
PTC_API const char *ptc_multi_reduce_kind(const ptc_ctx *ctx);

/* OpenCL_InitializeMemory: validate, flatten and upload the scene; set the camera. May be called again
 * with a new scene (accumulators are kept; call ptc_reset_accumulation to start over). */
PTC_API int ptc_upload_scene(ptc_ctx *ctx, const ptc_scene_view *scene);

/* New (the reference bakes the camera once): position, direction, right, up. */
PTC_API int ptc_set_camera(ptc_ctx *ctx, const float camera[16]);

/* Zero imageColor / imageRayNb / imageV / the three histograms and the counters. */
PTC_API int ptc_reset_accumulation(ptc_ctx *ctx);

/* The reference's `-D SUPER_SAMPLING` build option (globalVars.superSampling, Controleur/PathTracer_OpenCL.cpp:311): from
 * iteration 6 on Kernel_Main drops a path before tracing it when one more draw of its random stream exceeds
 * 100 * max_rgb(imageV / n) / X2inv[(uint) n] + 0.05, n = the pixel's sample count (Kernel/PathTracer_FullKernel.cl:1152-1172,
 * 1219-1222): pixels whose running variance is small stop receiving samples (imageRayNb then differs per pixel; the caller
 * divides by it as always). The reference ships the table X2inv as an EMPTY file (Kernel/X2inv.cl), so its option cannot
 * even compile; here the caller supplies it: x2inv[count] (indices beyond the end read the last entry), or x2inv = NULL
 * for the 95 % chi-square quantiles of n = 0..count-1 degrees of freedom. count = 0 switches it off (the default).
 * PARITY then renders one iteration per wavefront pass — the reference's sequential loop, decision for decision; FAST
 * keeps its passes and decides on the statistics at the start of each. Needs imageV (no PTC_FLAG_NO_VARIANCE). */
PTC_API int ptc_set_supersampling(ptc_ctx *ctx, const float *x2inv, uint32_t count);

/* The body of OpenCL_RunKernel's loop for iterations [first, first+n): enqueue only, returns without
 * waiting. Iteration numbers select the random streams (header.cl:255-264), so disjoint ranges rendered
 * by different contexts / GPUs sum to exactly the paths a single run would trace. */
PTC_API int ptc_render(ptc_ctx *ctx, uint32_t first_iteration, uint32_t n_iterations);

PTC_API int ptc_synchronize(ptc_ctx *ctx);

/* Pixel shard for progressive rendering on several GPUs (a 1-spp frame cannot be split by iterations). The
 * context renders only the image rows y with (y / stripe_rows) % world == rank — interleaved stripes, for load balance — of
 * every following ptc_render; the random stream of a path depends on (x, y, iteration, W, H) alone
 * (Kernel/PathTracer_FullKernel_header.cl:255-264), so each shard traces exactly the paths the whole-frame render traces for
 * its rows. Accumulators of rows outside the shard stay as they are (zero after ptc_reset_accumulation), and
 * ptc_convert_rgb8 writes 0 there, so the display frames of all shards SUM to the whole frame (one uint8 reduce over NCCL).
 * stripe_rows = 0 returns to the whole frame. The parity probes (ptc_trace_primary, ptc_render_single) refuse a shard. */
PTC_API int ptc_set_pixel_stripes(ptc_ctx *ctx, uint32_t stripe_rows, uint32_t rank, uint32_t world);

/* clEnqueueReadBuffer of imageColor / imageRayNb (reference PathTracer_OpenCL.cpp:97-98); waits for
 * pending renders. Either pointer may be NULL. */
PTC_API int ptc_read_image(ptc_ctx *ctx, float *image_color, float *image_ray_nb);
PTC_API int ptc_read_variance(ptc_ctx *ctx, float *image_v);

/* rayDepths[depth+1], rayIntersectedBBx[5000], rayIntersectedTri[5000] (reference :110-112). */
PTC_API int ptc_read_stats(ptc_ctx *ctx, uint32_t *ray_depths, uint32_t *ray_bbx, uint32_t *ray_tri);

/* Display epilogue on the device: sum/count -> clamp -> 8-bit, restating
 * reference Alone/PathTracer_bitmap.cpp:259-279 (negative red marks the pixel pure red). 3 bytes per
 * pixel, R,G,B, row-major, no padding. */
PTC_API int ptc_read_rgb8(ptc_ctx *ctx, uint8_t *rgb);

/* Resume / multi-GPU merge: add host accumulators into the device ones. */
PTC_API int ptc_add_image(ptc_ctx *ctx, const float *image_color, const float *image_ray_nb);

PTC_API int ptc_read_counters(ptc_ctx *ctx, ptc_counters *out);
PTC_API int ptc_get_device_buffers(ptc_ctx *ctx, ptc_device_buffers *out);

/* The device half of ptc_read_rgb8: sum / count -> clamp -> RGB8 into the context's display buffer
 * (ptc_device_buffers.rgb8), enqueued on the context's stream, no copy to the host. */
PTC_API int ptc_convert_rgb8(ptc_ctx *ctx);

/* Parity probe: generate + one closest-hit traversal for every pixel of one iteration, without touching
 * the accumulators. hit = triangle index in BVH order or -1; st = (s,t); dist2 = squared hit distance;
 * dir = 4 floats per pixel; seed_after = RNG state after the sampler; counts = (boxes, triangles) tested.
 * Any pointer may be NULL. */
PTC_API int ptc_trace_primary(ptc_ctx *ctx, uint32_t iteration, int32_t *hit, float *st, float *dist2, float *dir,
                              int32_t *seed_after, uint32_t *counts);

/* Parity probe: closest hits of n caller-supplied rays (origin, direction: 4 floats per ray; the direction is normalised
 * over its four lanes as Ray3D_Create does, reference Kernel/PathTracer_FullKernel_header.cl:286-305) through the extend
 * stage = BVH_IntersectRay (Kernel/PathTracer_FullKernel.cl:620-702). Outputs as ptc_trace_primary; any may be NULL.
 * n must not exceed the context's path slots (image pixels x iterations_per_batch). Does not touch the accumulators. */
PTC_API int ptc_trace_rays(ptc_ctx *ctx, uint32_t n, const float *origin, const float *direction, int32_t *hit, float *st,
                           float *dist2, float *dir, uint32_t *counts);

/* Parity probe: per-pixel radiance of exactly one iteration (what Kernel_Main adds to imageColor). */
PTC_API int ptc_render_single(ptc_ctx *ctx, uint32_t iteration, float *radiance);

/* Parity probe: the state of every path of one iteration after `bounces` bounces (0 = the camera rays): ray origin and direction
 * of bounce `bounces` (4 floats per pixel each), the hit of bounce `bounces` - 1 as (bits of the triangle index in BVH order or
 * -1, s, t, ray parameter) and the path's meta word (bits 7..0 = reflectionId, i.e. how many bounces the path really took: a
 * path that ended earlier keeps its last state). Any pointer may be NULL. Does not touch the accumulators. */
PTC_API int ptc_path_probe(ptc_ctx *ctx, uint32_t iteration, uint32_t bounces, float *origin, float *direction, float *hit, uint32_t *meta);

/* ptc_trace_rays without device allocation. NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * ptc_trace_rays allocates and frees seven device buffers per call; this entry point runs the same kernels with ALL of its
 * scratch — staged rays, the extend kernel's path-state records and queue counters, staged outputs — inside one file-scope
 * `static __device__` array of PTC_STATIC_SCRATCH_BYTES (opencl_pathtracer_b200/csrc/pt_static.cu), which every GPU that loads
 * the library holds until the process exits. It calls no cudaMalloc / cudaFree / allocating library routine.
 *   - n > PTC_STATIC_MAX_RAYS is refused with PTC_ERR_LIMIT (ptc_last_error names the limit); split the batch or use ptc_trace_rays.
 *   - origin, direction and the outputs may be host OR device pointers (copied with cudaMemcpyDefault on the context's stream);
 *     meaning and layout as ptc_trace_rays (no `dir` output); any output may be NULL. Results are bit-identical to ptc_trace_rays.
 *   - Every call on a device shares the array, whatever the context; it keeps the last call's contents and the entry point clears
 *     what must start at zero. Calls that overlap on the GPU (two host threads, two contexts with their own streams) RACE: serialise
 *     them. The call returns after the context's stream has drained. Does not touch the accumulators. */
#define PTC_STATIC_MAX_RAYS 32768u
#define PTC_STATIC_SCRATCH_BYTES (4u << 20)
PTC_API int ptc_trace_rays_static(ptc_ctx *ctx, uint32_t n, const float *origin, const float *direction, int32_t *hit, float *st,
                                  float *dist2, uint32_t *counts);

/* Host utility: the 8-wide compressed BVH that ptc_upload_scene collapses the reference BVH2 into in FAST mode
 * (opencl_pathtracer_b200/csrc/pt_wide_build.cu), exposed for inspection and structural tests. nodes = n_nodes records of
 * 80 bytes (DevNode8, csrc/pt_wide.cuh); tri_order has n_triangles = 24 * n_nodes entries: entry 24 * node + 3 * slot + k =
 * index in the BVH-ordered triangulation of triangle k of the leaf child in `slot`, 0xffffffff where there is none;
 * depth = levels of wide nodes. Needs no CUDA device. Release with ptc_wide_bvh_free. */
typedef struct ptc_wide_bvh
{
    void *nodes;
    uint32_t *tri_order;
    uint32_t n_nodes;
    uint32_t n_triangles;
    uint32_t depth;
    uint32_t reserved;
} ptc_wide_bvh;
PTC_API int ptc_wide_bvh_build(const ptc_scene_view *scene, ptc_wide_bvh *out);
/* The wide BVH the context traverses (after ptc_upload_scene, FAST mode without PTC_FLAG_BVH2), read back from the device:
 * the host-built one or, with PTC_FLAG_GPU_BVH, the one built on the device. reserved = device build time in microseconds
 * (0 for the host build). Release with ptc_wide_bvh_free. */
PTC_API int ptc_wide_bvh_read(ptc_ctx *ctx, ptc_wide_bvh *out);
PTC_API void ptc_wide_bvh_free(ptc_wide_bvh *bvh);

/* Measurement utility (roofline denominators of bench.py): read-only streaming of a `bytes`-sized device buffer, `repeats`
 * passes, with the 256-bit loads the traversal kernels use, L1 bypassed. A buffer that fits the L2 (e.g. 32 MB) gives the L2
 * hit bandwidth the node / triangle fetches of an L2-resident scene are bounded by; a buffer far larger than the L2 gives the
 * HBM read bandwidth. Result in GB/s (1e9 bytes). */
PTC_API int ptc_measure_read_bandwidth(ptc_ctx *ctx, size_t bytes, uint32_t repeats, double *gb_per_s);

/* Pinned host memory for read-backs inside timed regions. */
PTC_API void *ptc_host_alloc(size_t bytes);
PTC_API void ptc_host_free(void *p);

/* Message of the last failure on this context (or of the last failed ptc_create when ctx is NULL). */
PTC_API const char *ptc_last_error(const ptc_ctx *ctx);
PTC_API const char *ptc_version(void);
PTC_API int ptc_device_count(void);

/* Releases everything OpenCL_RunKernel releases at its end (reference PathTracer_OpenCL.cpp:120-139). */
PTC_API void ptc_destroy(ptc_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* PATHTRACER_CUDA_H */
