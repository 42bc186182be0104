/* scn_render.h — C ABI of the wavefront batching stage around the medium query.
 *
 * The reference traces one path at a time (integrators/path_tracer/PathTracer.cpp:17-204) and calls
 * Medium::sampleDistance once per bounce (:68).  This stage keeps every live path of a tile range in SoA buffers and, per bounce,
 * (1) packs the medium queries, (2) calls scn_sample_distance_batch (include/scn.h), (3) shades the results.  It covers a RESTRICTED
 * scene, enough for the BASELINE.json 3D render configs:
 *   - pinhole camera (cameras/PinholeCamera.cpp:26-32, 70-86), box pixel filter;
 *   - one axis-aligned cube with a "forward" BSDF as the medium container (primitives/Cube.cpp), analytic slab test;
 *   - the GPIS medium with phase function "brdf" over a Lambert BSDF (phasefunctions/BRDFPhaseFunction.cpp:27-44,
 *     bsdfs/LambertBsdf.cpp): cosine-weighted direction about the sampled gradient, weight = albedo;
 *   - a constant infinite-sphere emitter (primitives/InfiniteSphere.cpp), gathered when a path leaves the cube;
 *   - Russian roulette as PathTracer.cpp:146-152, integrator max_bounces.
 * Next-event estimation (TraceBase.cpp:346-420) is NOT done: with a uniform environment the estimator without it has the same
 * expectation, so images agree with the reference's in the mean, not sample by sample (the reference's per-tile sampler streams
 * cannot be reproduced on a wavefront; this stage draws from a counter-based hash of (pixel, sample, bounce, dimension)).
 *
 * Tiles are the reference's 16 x 16 tiles in diceTiles order (row-major, integrators/path_tracer/PathTraceIntegrator.cpp:26-41);
 * a GPU owns a contiguous tile range and the framebuffer slice of exactly those tiles.
 */
#ifndef SCN_RENDER_H_
#define SCN_RENDER_H_

#include "scn.h"

#ifdef __cplusplus
extern "C" {
#endif

#define SCN_WF_TILE 16

typedef struct scn_wf_scene {
    float cam_pos[3], cam_look_at[3], cam_up[3]; /* "transform": position / look_at / up  */
    float fov_deg;                               /* "fov"                                  */
    uint32_t res_x, res_y;                       /* "resolution"                           */
    float cube_center[3], cube_half[3];          /* cube "transform": position, scale / 2  */
    float emission;                              /* infinite_sphere "emission" (grey)      */
    float albedo;                                /* lambert "albedo" of the brdf phase     */
    int32_t max_bounces;                         /* integrator "max_bounces"               */
    uint32_t scene_seed;                         /* TraceableScene seed (-s)               */
} scn_wf_scene;

/* Per-path buffers of one wavefront (all device pointers, n elements each, caller-owned). */
typedef struct scn_wf_paths {
    scn_rays rays;          /* mutable view: the stage writes the arrays the const pointers alias      */
    scn_state state;
    scn_samples samples;
    float *throughput;      /* path throughput (grey)                                                  */
    float *radiance;        /* radiance carried by the path once it has ended                          */
    uint8_t *alive;         /* 1 while the path still needs medium queries                             */
} scn_wf_paths;

/* Number of tiles of an image (diceTiles). */
uint32_t scn_wf_tile_count(uint32_t res_x, uint32_t res_y);

/* Camera rays for tiles [tile_first, tile_first + tile_count) x samples [spp_first, spp_first + spp_count): path
 * i = (local_pixel * spp_count + s), local_pixel = (tile - tile_first) * 256 + in-tile row-major offset.  Paths whose pixel lies
 * outside the image (edge tiles) or whose ray misses the cube are finished at once.  Replaces the camera part of
 * PathTracer::traceSample (:31-59) and the first scene intersection. */
int scn_wf_raygen(const scn_wf_scene *scene, const scn_medium *m, uint32_t tile_first, uint32_t tile_count, uint32_t spp_first, uint32_t spp_count,
                  const scn_wf_paths *paths, size_t n, void *stream);

/* After scn_sample_distance_batch on the wavefront: terminate / scatter every path (PathTracer.cpp:69-169 restricted as above)
 * and write the next bounce's rays in place.  `bounce` is the index of the bounce just queried. */
int scn_wf_shade(const scn_wf_scene *scene, const scn_medium *m, const scn_wf_paths *paths, size_t n, int32_t bounce, uint32_t spp_first, uint32_t spp_count,
                 uint32_t tile_first, void *stream);

/* framebuffer[local_pixel] += sum over the spp_count samples of the pixel (fixed order: deterministic).  Replaces
 * OutputBuffer::addSample (PathTraceIntegrator.cpp:159). */
int scn_wf_accumulate(const float *radiance, float *framebuffer, size_t n_pixels, uint32_t spp_count, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* SCN_RENDER_H_ */
