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
 * With scene.nee == 0 next-event estimation (TraceBase.cpp:346-420) is NOT done: with a uniform environment the estimator
 * without it has the same expectation.  With scene.nee == 1 the stage runs the reference's estimator (light + phase sampling with
 * MIS at every medium scatter, mirror / conductor phase functions over the 1D-noise NEE entry points, one quad emitter), in RGB.
 * Either way images agree with the reference's in the mean, not sample by sample (the reference's per-tile sampler streams
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
    /* ---- next-event estimation stage; everything below is ignored while nee == 0 ---- */
    int32_t nee;                                 /* 1: the reference's estimator — volumeEstimateDirect at every medium scatter
                                                    (TraceBase.cpp:346-420, 506-518, 539-563), emitters seen directly only by
                                                    unscattered paths (TraceBase.cpp:629-637); RGB                                 */
    int32_t phase_type;                          /* SCN_WF_PHASE_*: bsdf of the "brdf" phase function                             */
    float albedo_rgb[3];                         /* bsdf "albedo"                                                                 */
    float eta[3], k[3];                          /* conductor: complex IOR (bsdfs/ComplexIorData.hpp), 0/0 = perfect mirror       */
    float env_rgb[3];                            /* infinite_sphere "emission"                                                    */
    int32_t env_sample;                          /* infinite_sphere "sample"                                                      */
    int32_t has_quad;                            /* one quad emitter (primitives/Quad.cpp), null bsdf                             */
    int32_t quad_first;                          /* 1 when the quad precedes the infinite_sphere in "primitives" (light order)    */
    float quad_base[3], quad_edge0[3], quad_edge1[3]; /* Quad::prepareForRender (:296-316)                                        */
    float quad_rgb[3];                           /* quad "emission"                                                               */
} scn_wf_scene;

#define SCN_WF_PHASE_LAMBERT 0
#define SCN_WF_PHASE_MIRROR 1
#define SCN_WF_PHASE_CONDUCTOR 2

/* Per-path buffers of one wavefront (all device pointers, n elements each, caller-owned). */
typedef struct scn_wf_paths {
    scn_rays rays;          /* mutable view: the stage writes the arrays the const pointers alias      */
    scn_state state;
    scn_samples samples;
    float *throughput;      /* path throughput (grey)                                                  */
    float *radiance;        /* radiance carried by the path once it has ended                          */
    uint8_t *alive;         /* 1 while the path still needs medium queries                             */
    /* nee stage only (NULL otherwise): green / blue planes (throughput / radiance above are the red ones) and the
     * "has scattered in the medium" flag (wasSpecular == false, PathTracer.cpp:58, TraceBase.cpp:544) */
    float *throughput_gb[2];
    float *radiance_gb[2];
    uint8_t *scattered;
} scn_wf_paths;

/* Shadow rays of one half of volumeSampleDirect (TraceBase.cpp:444-456): the light-sampled ray (volumeLightSample) or the
 * phase-sampled ray (volumePhaseSample).  scn_wf_nee_setup fills rays / state (a copy of the path's MediumState with the segment
 * index advanced, lastAniso replaced by neeGrad for specular phase functions: TraceBase.cpp:372-376, 543-546) and the unshadowed
 * contribution; the caller runs scn_transmittance_batch(rays, state) -> T; scn_wf_nee_gather adds contrib * T to the radiance. */
typedef struct scn_wf_shadow {
    scn_rays rays;
    scn_state state;
    float *T;
    float *contrib[3];
} scn_wf_shadow;

/* Number of tiles of an image (diceTiles). */
uint32_t scn_wf_tile_count(uint32_t res_x, uint32_t res_y);

/* Camera rays for tiles [tile_first, tile_first + tile_count) x samples [spp_first, spp_first + spp_count): path
 * i = (local_pixel * spp_count + s), local_pixel = (tile - tile_first) * 256 + in-tile row-major offset.  Paths whose pixel lies
 * outside the image (edge tiles) or whose ray misses the cube are finished at once.  Replaces the camera part of
 * PathTracer::traceSample (:31-59) and the first scene intersection. */
int scn_wf_raygen(const scn_wf_scene *scene, const scn_medium *m, uint32_t tile_first, uint32_t tile_count, uint32_t spp_first, uint32_t spp_count,
                  const scn_wf_paths *paths, size_t n, void *stream);

/* The same for tiles tile_first + k * tile_stride, k in [0, tile_count): round-robin tile ownership across GPUs (SURVEY.md 8(e):
 * interleaved ranges balance the load when contiguous ranges leave some GPUs with background-only tiles).
 * local_pixel = k * 256 + in-tile offset. */
int scn_wf_raygen_strided(const scn_wf_scene *scene, const scn_medium *m, uint32_t tile_first, uint32_t tile_stride, uint32_t tile_count,
                          uint32_t spp_first, uint32_t spp_count, const scn_wf_paths *paths, size_t n, void *stream);

/* After scn_sample_distance_batch on the wavefront: terminate / scatter every path (PathTracer.cpp:69-169 restricted as above)
 * and write the next bounce's rays in place.  `bounce` is the index of the bounce just queried. */
int scn_wf_shade(const scn_wf_scene *scene, const scn_medium *m, const scn_wf_paths *paths, size_t n, int32_t bounce, uint32_t spp_first, uint32_t spp_count,
                 uint32_t tile_first, void *stream);

/* framebuffer[local_pixel] += sum over the spp_count samples of the pixel (fixed order: deterministic).  Replaces
 * OutputBuffer::addSample (PathTraceIntegrator.cpp:159). */
int scn_wf_accumulate(const float *radiance, float *framebuffer, size_t n_pixels, uint32_t spp_count, void *stream);

/* volumeEstimateDirect for every path that scattered at `bounce` (TraceBase.cpp:346-420, 444-518): light choice
 * (chooseLight over approximateRadiance), Quad / InfiniteSphere sampleDirect and directPdf, phase eval / sample / pdf of the
 * Lambert, mirror and conductor bsdfs — for the specular ones through neePDF / neeGrad of the medium's 1D noise
 * (bsdfs/ConductorBsdf.cpp:58-140, bsdfs/MirrorBsdf.cpp:40-110) — and the power heuristic.  Call between
 * scn_sample_distance_batch and scn_wf_shade. */
int scn_wf_nee_setup(const scn_wf_scene *scene, const scn_medium *m, const scn_wf_paths *paths, const scn_wf_shadow *light_rays,
                     const scn_wf_shadow *phase_rays, size_t n, int32_t bounce, void *stream);
int scn_wf_nee_gather(const scn_wf_paths *paths, const scn_wf_shadow *light_rays, const scn_wf_shadow *phase_rays, size_t n, void *stream);
/* The sample clamp of PathTraceIntegrator::renderTile (PathTraceIntegrator.cpp:148-156) on the finished paths of an nee (RGB)
 * wavefront: a sample with a NaN, an infinite component or a component above 65536 becomes black.  Call before scn_wf_accumulate. */
int scn_wf_clamp_samples(const scn_wf_paths *paths, size_t n, void *stream);

/* The whole render loop of one GPU's tiles in ONE call: for every chunk of spp_chunk samples per pixel
 *   raygen -> max_bounces x { scn_sample_distance_batch, [nee: setup + two scn_transmittance_batch], shade (+ gather) } -> clamp -> accumulate
 * enqueued on `stream` without a host synchronisation.  Replaces PathTraceIntegrator::renderTile
 * (integrators/path_tracer/PathTraceIntegrator.cpp:135-163) over PathTracer::traceSample (PathTracer.cpp:62-169).  The per-path
 * buffers must hold tile_count * 256 * spp_chunk paths; framebuffer[0] (red or grey) and, with scene.nee, [1] / [2] (green, blue)
 * hold tile_count * 256 running sums.  light_rays / phase_rays may be NULL when scene.nee == 0. */
int scn_wf_render(const scn_wf_scene *scene, const scn_medium *m, uint32_t tile_first, uint32_t tile_stride, uint32_t tile_count, uint32_t spp_first,
                  uint32_t spp_count, uint32_t spp_chunk, const scn_wf_paths *paths, const scn_wf_shadow *light_rays, const scn_wf_shadow *phase_rays,
                  float *const framebuffer[3], void *stream);

#ifdef __cplusplus
}
#endif
#endif /* SCN_RENDER_H_ */
