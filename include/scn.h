/* scn.h — C ABI of the B200-native sparse-convolution-noise GPIS medium.
 *
 * Drop-in boundary for ONE path of PeterZs/sparse-conv-gpis-tungsten: the ray-marched GPIS medium
 * query.  Every entry point names the reference interface it replaces (paths relative to the
 * reference's src/core/).  Plain pointers and sizes only; no C++/torch types cross this boundary.
 *
 * Conventions
 *   - every function returns 0 on success, a negative SCN_E* code otherwise; scn_last_error()
 *     returns the message of the calling thread's last failure.  No exception crosses the ABI.
 *   - batch buffers are caller-owned DEVICE pointers (struct-of-arrays, one element per query);
 *     the *_host variants take HOST pointers and do the host<->device copies themselves, through
 *     a staging arena owned by the medium (grown on first use, reused afterwards).  Host buffers
 *     SHOULD be page-locked (cudaHostAlloc / cudaHostRegister): pageable buffers give the same
 *     results, but the driver stages every copy and the upload / kernel / download stages of the
 *     pipeline no longer overlap (bench.py reports both: e2e.value pinned, e2e.pageable).
 *   - a wavefront holds at most 2^30 queries (SCN_EINVAL above, before anything is enqueued).
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  Calls on
 *     distinct streams are re-entrant; a scn_medium is immutable after creation.
 *   - optional pointers may be NULL where stated.
 */
#ifndef SCN_H_
#define SCN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SCN_ABI_VERSION 5

/* error codes */
#define SCN_OK 0
#define SCN_EINVAL (-1)      /* bad argument / bad descriptor                                  */
#define SCN_EJSON (-2)       /* JSON could not be parsed or holds an invalid enum string        */
#define SCN_EUNSUPPORTED (-3)/* valid reference configuration that this library does not cover  */
#define SCN_ECUDA (-4)       /* CUDA runtime error (no device, launch failure, ...)             */
#define SCN_ENOMEM (-5)

/* GPCorrelationContext, math/GaussianProcess.hpp:26-31 (same order) */
#define SCN_CTX_GLOBAL 0
#define SCN_CTX_RENEWAL_PLUS 1
#define SCN_CTX_RENEWAL 2
#define SCN_CTX_NONE 3

/* SparseConv1DSamplingScheme, media/Medium.hpp (same order) */
#define SCN_SCHEME_UNI 0
#define SCN_SCHEME_NEE 1
#define SCN_SCHEME_MIS 2

/* MeanFunctionFactory strings, math/GaussianProcessFactory.cpp:14-22 (covered subset) */
#define SCN_MEAN_HOMOGENEOUS 0
#define SCN_MEAN_SPHERICAL 1
#define SCN_MEAN_LINEAR 2
#define SCN_MEAN_PROCEDURAL 3 /* "procedural": ProceduralMean over an SDF function or a procedural scalar field, finite-difference gradient */

/* CovarianceFunctionFactory strings, math/GaussianProcessFactory.cpp:25-38 (covered subset) */
#define SCN_COV_SQUARED_EXPONENTIAL 0
#define SCN_COV_PROC_NONSTATIONARY 1 /* "proc_nonstationary" wrapping a squared_exponential "cov" */
/* alternative splat kernels (math/GPFunctions.cpp:1020-1082, 1127-1214).  They have no second derivative, no 1D family and
 * never initialise the world<->isotropic transforms, so the reference only defines them for world-space 3D sampling
 * without Renewal conditioning; scn_medium_create answers SCN_EUNSUPPORTED for anything else. */
#define SCN_COV_MATERN 2             /* "matern": "sigma" "v" (0.5 | 1.5 | 2.5) "lengthScale" "aniso" "localScale" */
#define SCN_COV_GABOR_ANISO 3        /* "gabor_aniso": "sigma" "a_inv" "f_inv" "omega" "localScale"               */
#define SCN_COV_GABOR_ISO 4          /* "gabor_iso": "sigma" "a_inv" "f_inv" "localScale"                         */

/* ProceduralScalar / ProceduralVector, math/GPFunctions.hpp:407-427,604-692,693-781 */
#define SCN_FIELD_NONE 0
#define SCN_FIELD_CONSTANT 1
#define SCN_FIELD_BOTTOM_TOP 2
#define SCN_FIELD_LEFT_RIGHT 3
#define SCN_FIELD_FRONT_BACK 4
#define SCN_FIELD_BOTTOM_TOP_LEFT_RIGHT 5
#define SCN_FIELD_SANDSTONE 7   /* "noise":"sandstone": three nested fbm over simplex noise (math/GPFunctions.cpp:70-77, 104-111; math/SdfFunctions.cpp:200-296) */
#define SCN_FIELD_RUST 8        /* "noise":"rust" (math/GPFunctions.cpp:78-83, 112-117) */
#define SCN_FIELD_REGULAR_GRID 6 /* "type":"regular_grid" (RegularGridScalar / RegularGridVector, math/GPFunctions.hpp:429-578) */

/* RegularGrid<ElemType>::interp (math/GPFunctions.hpp:241-306; strings of grids/Grid.cpp:6-26) */
#define SCN_INTERP_POINT 0
#define SCN_INTERP_LINEAR 1
#define SCN_INTERP_QUADRATIC 2


/* RegularGrid<double> / RegularGrid<Vec3d> (math/GPFunctions.hpp:241-330): a dense res^3 array over `bounds`, element
 * (x res + y) res + z (:255-258), looked up by point / trilinear / triquadratic interpolation in double (:284-306).
 * `values` is a HOST pointer to res^3 doubles (scalar field) or 3 res^3 doubles (vector field, xyz interleaved like the reference's
 * "-values.bin" file); scn_medium_create copies it to the device, the caller keeps ownership — except for grids that
 * scn_desc_from_json_at loaded from disk (`owned` = 1), which scn_desc_release frees. */
typedef struct scn_grid_desc {
    const double *values;
    double bounds_min[3], bounds_max[3]; /* "bounds_min" "bounds_max" */
    uint32_t res;                        /* "res"                      */
    int32_t interp;                      /* "interpolate": SCN_INTERP_*, default linear */
    int32_t owned, _pad;
} scn_grid_desc;

/* procedural scalar/vector field: "type":"constant"|"noise"|"regular_grid" (math/GPFunctions.cpp:53-138, GPFunctions.hpp:429-578) */
typedef struct scn_field_desc {
    int32_t type;        /* SCN_FIELD_*                                              */
    int32_t _pad;
    double val[3];       /* constant: "val" (scalar fields use val[0])               */
    double min, max, start, end;      /* noise ramps: "min" "max" "start" "end"      */
    double min2, max2, start2, end2;  /* bottom_top_left_right second ramp           */
    scn_grid_desc grid;  /* regular_grid: "file" -> grid JSON + binary values (appended in ABI version 5) */
} scn_field_desc;

/* SdfFunctions::Function (math/SdfFunctions.hpp:12-18) for "type":"procedural" means given by "func" */
#define SCN_SDF_FIELD 0        /* no "func": the mean is the ProceduralScalar "f" (constant | noise | regular_grid)          */
#define SCN_SDF_KNOB_INNER 1   /* "knob_inner": sphere of radius 0.75 in the 0.8-scaled frame (math/SdfFunctions.cpp:148-157) */
#define SCN_SDF_TWO_SPHERES 2  /* "two_spheres" (math/SdfFunctions.cpp:181-183)                                               */
#define SCN_SDF_PLANE 3        /* "plane": p.y (math/SdfFunctions.hpp:53)                                                     */
#define SCN_SDF_KNOB 4         /* "knob": the material-test knob (math/SdfFunctions.cpp:100-145): sphere, cut-outs, torus etch, cone base */
#define SCN_SDF_KNOB_OUTER 5   /* "knob_outer" (math/SdfFunctions.cpp:159-179)                                                */

/* mean function: math/GPFunctions.hpp:867-1005 (homogeneous, spherical, linear), :1049-1087 + math/GPFunctions.cpp:250-296 (procedural) */
typedef struct scn_mean_desc {
    int32_t type;        /* SCN_MEAN_*                                               */
    float offset;        /* homogeneous, procedural: "offset"                        */
    double center[3];    /* spherical: "center"                                      */
    float radius;        /* spherical: "radius"                                      */
    float scale;         /* linear, procedural: "scale"                              */
    double ref[3];       /* linear: "reference_point"                                */
    double dir[3];       /* linear: "direction" (normalised at parse time)           */
    float min;           /* linear, procedural: "min" (default -FLT_MAX)             */
    int32_t sdf_func;    /* procedural: SCN_SDF_* ("func"; SCN_SDF_FIELD when the mean is the field "f")                         */
    /* procedural (appended in ABI version 5): the INVERSE of "transform" (Mat4f::invert, math/Mat4f.hpp:75-101), rows 0-2 of the 4 x 4
     * matrix row-major — ProceduralMean::mean evaluates f(invConfigTransform.transformPoint(Vec3f(p))) * scale + offset, floored at min */
    float inv_transform[12];
    scn_field_desc f;    /* procedural: "f" (a scalar field: a regular_grid here is a tabulated signed-distance mean)             */
} scn_mean_desc;

/* Flattened JSON of one "sparse_conv_noise" medium and its "gaussian_process".
 * Keys and defaults: media/SparseConvolutionNoiseMedium.cpp:17-35,57-73,
 * media/GaussianProcessMedium.cpp:97-126, media/Medium.cpp:28-38, math/GaussianProcess.cpp:172-190,
 * math/GPFunctions.cpp:654-679, math/GPFunctions.hpp:1134-1137,1481-1484,2199-2205. */
typedef struct scn_medium_desc {
    /* sparse_conv_noise */
    float step_size;                  /* "step_size"               default 0.01       */
    uint32_t min_step;                /* "min_step"                default 8          */
    uint32_t seed;                    /* "seed"                    default 0          */
    float impulse_density;            /* "impulse_density"         default 3 (truncated to uint per cell) */
    int32_t single_realization;       /* "single_realization"      default false      */
    int32_t isotropic_3d_sampling;    /* "isotropic_3D_sampling"   default false      */
    int32_t sampling_1d;              /* "1D_sampling"             default false      */
    int32_t sampling_scheme_1d;       /* "1D_sampling_scheme"      SCN_SCHEME_*       */
    int32_t correlation_xy;           /* "1D_gradient_correlationXY" default false    */
    int32_t surf_vol_phase_separate;  /* "surf_vol_phase_separate" default false      */
    float surf_vol_phase_amp_thresh;  /* "surf_vol_phase_amp_thresh" default 0        */
    /* gaussian_process medium base */
    float sigma_a, sigma_s, density;  /* "sigma_a" "sigma_s" "density" (0, 0, 1)      */
    int32_t correlation_context;      /* "correlation_context" SCN_CTX_* (mandatory)  */
    int32_t max_bounces;              /* "max_bounces"             default 1024       */
    /* gaussian process */
    int32_t gp_id;                    /* "id"                      default 0          */
    int32_t has_mean_additional;      /* "csg": right GP's mean (math/GaussianProcess.cpp:25-39) */
    int32_t id_additional;
    int32_t _pad0;
    scn_mean_desc mean;
    scn_mean_desc mean_additional;
    /* covariance */
    int32_t cov_type;                 /* SCN_COV_*                                    */
    int32_t use_aniso_mtx;            /* "useAnisoMtx"             default false      */
    float sigma;                      /* "sigma"                   default 1          */
    float length_scale;               /* "lengthScale"             default 1          */
    float aniso[3];                   /* "aniso"                   default [1,1,1]    */
    float kernel_scale;               /* "localScale"              default 3          */
    float aniso_mtx[9];               /* "anisoMtx" row-major                         */
    /* proc_nonstationary */
    int32_t multires;                 /* "multiResolutionGrid"     default false      */
    scn_field_desc ls;                /* "ls"  (vector field; lateral scale)          */
    scn_field_desc var;               /* "var" (scalar field; variance)               */
    /* matern / gabor (appended in ABI version 2) */
    float matern_v;                   /* "v"                       default 1          */
    float gabor_a_inv;                /* "a_inv"                   default 1          */
    float gabor_f_inv;                /* "f_inv"                   default 1          */
    float gabor_omega[3];             /* "omega" (normalised at parse time; the reference leaves it uninitialised when absent) */
    /* proc_nonstationary rotation anisotropy (appended in ABI version 3): "aniso" scalar field, angle = field(p) * pi/2 about z,
     * axis scales (1.5, 1/1.5, 1) hard-coded in the reference (math/GPFunctions.cpp:1678-1690, GPFunctions.hpp:2372).
     * 3D sampling only: the 1D path reads StationaryCovariance::_local_to_world, whose off-diagonal is never initialised. */
    int32_t _pad1;
    scn_field_desc aniso_field;
} scn_medium_desc;

typedef struct scn_medium scn_medium; /* opaque; replaces the SparseConvolutionNoiseMedium object */

/* Rays of one wavefront.  Replaces the (Ray, sampler.next1D(), state.info.pixelSampleSegment)
 * arguments of Medium::sampleDistance / transmittance (media/Medium.hpp:104-111, math/Ray.hpp:24-41,
 * samplerecords/MediumSample.hpp:13-17).  `u` is the one path-sampler draw of the march
 * (media/SparseConvolutionNoiseMedium.cpp:129). */
typedef struct scn_rays {
    const float *ox, *oy, *oz;   /* ray.pos()                                   */
    const float *dx, *dy, *dz;   /* ray.dir()                                   */
    const float *tmin, *tmax;    /* ray.nearT(), ray.farT() (+inf allowed)      */
    const float *u;              /* march jitter in [0,1)                       */
    const uint32_t *px, *py, *spp, *segment; /* info.pixelSampleSegment (x,y,sample,bounce); may all be NULL when single_realization */
    uint32_t scene_seed;         /* info.sceneSeed                              */
} scn_rays;

/* Medium::MediumState (media/Medium.hpp:59-88) + the conditioning coefficients the reference keeps in
 * state.gpContext (math/SparseConvolutionNoise.hpp:7-21).  Read and written in place by
 * scn_sample_distance_batch; read only by scn_transmittance_batch. */
typedef struct scn_state {
    uint8_t *first_scatter;
    float *last_val;
    float *last_gx, *last_gy, *last_gz;  /* lastAniso */
    int32_t *bounce;
    int32_t *last_gp_id;
    float *info_t;                       /* state.info.t */
    uint8_t *scheme;                     /* state.sparseConv1DSamplingScheme */
    float *ctx[7];                       /* value_scale, gradient_scale xyz, ray_origin xyz */
} scn_state;

/* MediumSample (samplerecords/MediumSample.hpp:20-36) as SoA. */
typedef struct scn_samples {
    uint8_t *ok;            /* return value of sampleDistance                                  */
    uint8_t *exited;
    float *t, *continued_t;
    float *gx, *gy, *gz;    /* sample.aniso                                                    */
    int32_t *gp_id;
    uint8_t *scheme;
    float *weight;          /* sigmaS/sigmaT (colour fields are host-side)                      */
    uint32_t *n_evals;      /* optional: noise evaluations spent on this query (bench counter)  */
} scn_samples;

/* what to evaluate in scn_eval_batch */
#define SCN_EVAL_NOISE 1     /* evaluateNoise3D(p, dir, seed, ., false)  -> val, gx, gy, gz (math/SparseConvolutionNoise.cpp:101-124);
                                1D media: evaluateNoise1DNormalized(...) -> val, 0, 0, gz (iso-ray space, :322-360) */
#define SCN_EVAL_VALUE 2     /* evaluateValue  (math/SparseConvolutionNoise.cpp:73-89)  -> val, gp_id */
#define SCN_EVAL_GRADIENT 4  /* evaluateGradient (math/SparseConvolutionNoise.cpp:92-99) -> gx, gy, gz */

typedef struct scn_eval_in {
    const float *px, *py, *pz;       /* query point (world)                                       */
    const float *dx, *dy, *dz;       /* ray direction (world)                                     */
    const float *t;                  /* optional: t argument (1D: tSegment selector of evaluateGradient) */
    const float *info_t;             /* optional: info.t                                          */
    const uint32_t *pix_x, *pix_y, *pix_spp, *pix_seg; /* optional pixelSampleSegment             */
    uint32_t scene_seed;
    uint32_t noise_seed;             /* SCN_EVAL_NOISE: the explicit `seed` argument              */
    const float *ctx[7];             /* optional conditioning coefficients (all or none)          */
} scn_eval_in;

typedef struct scn_eval_out {
    float *val;          /* may be NULL unless requested */
    float *gx, *gy, *gz;
    int32_t *gp_id;      /* optional */
} scn_eval_out;

/* 1D next-event estimation (math/SparseConvolutionNoise.cpp:601-743) */
typedef struct scn_nee_in {
    const float *dx, *dy, *dz;       /* rayDir                                  */
    const float *nx, *ny, *nz;       /* normal requested by the light sample    */
    const float *px, *py, *pz;       /* vertex position                         */
    const float *t_segment;          /* neePDF only                             */
    const float *info_t;
    const uint32_t *pix_x, *pix_y, *pix_spp, *pix_seg;
    uint32_t scene_seed;
    const float *ctx[7];
} scn_nee_in;

const char *scn_last_error(void);
int scn_abi_version(void);

/* JSON -> descriptor.  Accepts the reference's medium JSON unchanged
 * (media/SparseConvolutionNoiseMedium.cpp:57-73 and the fromJson chain below it).
 * Unknown keys (phase_function, transmittance, name, ...) are ignored like the reference does. */
int scn_desc_default(scn_medium_desc *desc);
int scn_desc_from_json(const char *medium_json, scn_medium_desc *desc);
/* The same with the directory that relative resource paths ("file" of a regular_grid field, then "path" inside the grid JSON relative
 * to that file) resolve against, like Scene::fetchResource does against the scene file's directory (io/Scene.cpp, load_grid:
 * math/GPFunctions.hpp:386-401).  scn_desc_from_json uses the current directory.  Grids are read into malloc'ed host arrays owned by
 * the descriptor: call scn_desc_release when the descriptor is no longer needed (after scn_medium_create; a no-op otherwise). */
int scn_desc_from_json_at(const char *medium_json, const char *base_dir, scn_medium_desc *desc);
void scn_desc_release(scn_medium_desc *desc);

/* MediumFactory "sparse_conv_noise" + prepareForRender (media/MediumFactory.cpp:13-22,
 * media/GaussianProcessMedium.cpp:152-158).  `device` is the CUDA ordinal that owns the medium. */
int scn_medium_create(const scn_medium_desc *desc, int device, scn_medium **out);
void scn_medium_destroy(scn_medium *m);

/* Which kernel family serves a medium (no reference counterpart: the reference has one code path).  Configurations outside the
 * tuned kernels' coverage run on the one-thread-per-ray generic kernels — correct, golden-checked, several times slower — and
 * this call is how a caller sees that instead of discovering it in a profile. */
#define SCN_KERNEL_GENERIC 0       /* k_sample_distance / k_transmittance / k_eval_noise: any configuration                     */
#define SCN_KERNEL_MARCH3D 1       /* k_march: 3D sampling, squared-exponential kernel, shared-memory impulse cache             */
#define SCN_KERNEL_MARCH1D 2       /* k_cond1d -> k_march1d -> k_finish1d: 1D sampling, register cell window                    */
#define SCN_KERNEL_FAST_NOISE3D 3  /* k_fast_noise3d: SCN_EVAL_NOISE of stationary 3D media (the roofline kernel)               */
typedef struct scn_medium_info_t {
    int32_t query_kernel;          /* SCN_KERNEL_* serving scn_sample_distance_batch / scn_transmittance_batch                  */
    int32_t eval_kernel;           /* SCN_KERNEL_* serving scn_eval_batch(SCN_EVAL_NOISE)                                        */
    int32_t march_block;           /* threads per block of the query kernel (0: generic)                                        */
    int32_t cache_entries;         /* k_march: impulse-cache capacity per lane and level                                        */
    uint32_t smem_bytes;           /* k_march: dynamic shared memory per block                                                  */
    int32_t device, sm_count;
    float kernel_radius;           /* splattingKernelRadius of the sampling space = lattice cell size                           */
} scn_medium_info_t;
int scn_medium_info(const scn_medium *m, scn_medium_info_t *out);

/* SparseConvolutionNoiseRealization::evaluateNoise3D / evaluateValue / evaluateGradient */
int scn_eval_batch(const scn_medium *m, const scn_eval_in *in, const scn_eval_out *out, size_t n, int what, void *stream);
int scn_eval_batch_host(scn_medium *m, const scn_eval_in *in, const scn_eval_out *out, size_t n, int what);

/* SparseConvolutionNoiseRealization::conditioning (math/SparseConvolutionNoise.cpp:431-595):
 * p = ray origin, target value/gradient -> the 7 coefficients written to ctx_out[0..6] */
int scn_conditioning_batch(const scn_medium *m, const scn_eval_in *in, const float *target_val, const float *target_gx,
                           const float *target_gy, const float *target_gz, float *const ctx_out[7], size_t n, void *stream);

/* Medium::sampleDistance (media/GaussianProcessMedium.cpp:221-341 over
 * media/SparseConvolutionNoiseMedium.cpp:93-183) */
int scn_sample_distance_batch(const scn_medium *m, const scn_rays *rays, const scn_state *state, const scn_samples *out,
                              size_t n, void *stream);
int scn_sample_distance_batch_host(scn_medium *m, const scn_rays *rays, const scn_state *state, const scn_samples *out, size_t n);

/* NO COUNTERPART IN THE REFERENCE — added for callers that cannot allocate device memory
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same contract and same results as
 * scn_sample_distance_batch (device buffers, state updated in place), for the media whose query kernel needs per-ray scratch
 * memory (1D-sampling media, scn_medium_info().query_kernel == SCN_KERNEL_MARCH1D): the 52 bytes of scratch per ray that
 * scn_sample_distance_batch takes from the medium's stream-ordered pool live here in ONE file-scope device array of
 * SCN_STATIC_SCRATCH_MAX_RAYS rays (3.25 MiB, held by every GPU that loads the library until the process exits).  The call performs
 * no cudaMalloc / cudaFree / pool allocation; it enqueues a clear of the part of the array it uses, then the kernels.
 *   n > SCN_STATIC_SCRATCH_MAX_RAYS        -> SCN_EINVAL (split the wavefront or call scn_sample_distance_batch)
 *   a medium whose kernel uses no scratch  -> SCN_EUNSUPPORTED (scn_sample_distance_batch allocates nothing for it either)
 * Every call on a device shares the one array: calls that can overlap on the GPU (different streams, different media, different
 * host threads) RACE — serialise them on one stream or with events.  The array keeps the previous call's contents between calls. */
#define SCN_STATIC_SCRATCH_MAX_RAYS 65536
int scn_sample_distance_batch_static(const scn_medium *m, const scn_rays *rays, const scn_state *state, const scn_samples *out,
                                     size_t n, void *stream);

/* Medium::transmittance (media/GaussianProcessMedium.cpp:343-393): T_out[i] = 1 (no crossing) or 0 */
int scn_transmittance_batch(const scn_medium *m, const scn_rays *rays, const scn_state *state, float *T_out, size_t n, void *stream);
int scn_transmittance_batch_host(scn_medium *m, const scn_rays *rays, const scn_state *state, float *T_out, size_t n);

/* ((GPContextSparseConvNoise*)sample.ctxt)->noise.neePDF / neeGrad
 * (bsdfs/ConductorBsdf.cpp:72-73,96-97,106-107; bsdfs/MirrorBsdf.cpp:45,69,79) */
int scn_nee_pdf_batch(const scn_medium *m, const scn_nee_in *in, float *pdf_out, size_t n, void *stream);
int scn_nee_grad_batch(const scn_medium *m, const scn_nee_in *in, float *gx, float *gy, float *gz, size_t n, void *stream);

/* integer primitives exposed for the bit-exactness tests (math/MathUtil.hpp:179-224,
 * sampling/UniformSampler.hpp:41-75): device-side hash of n key quadruples and PCG streams. */
int scn_debug_hash4(const uint32_t *a, const uint32_t *b, const uint32_t *c, const uint32_t *d, uint32_t *out, size_t n, void *stream);
int scn_debug_pcg_stream(const uint32_t *seed32, float *out, size_t n, int draws, void *stream);
/* per-cell impulses exactly as cell3D generates them: out[(i*count + k)*4 + {0,1,2,3}] = x, y, z, w */
int scn_debug_cell_impulses(const uint32_t *ci, const uint32_t *cj, const uint32_t *ck, uint32_t seed, uint32_t count,
                            float *out, size_t n, void *stream);

/* test hook: on != 0 makes the query entry points of this medium use the generic one-thread-per-ray kernels even where a tuned
 * kernel covers the configuration (the tests compare the two families on the same wavefront) */
int scn_debug_force_generic(scn_medium *m, int on);

/* per point: how many of the 27*N impulses generated by evaluateNoise3D pass the rejection test |p - p_i|^2 < 1
 * (math/SparseConvolutionNoise.cpp:389) — stationary 3D media; proves the accepted-impulse SETS match the CPU's. */
int scn_debug_accepted_counts(const scn_medium *m, const float *px, const float *py, const float *pz, const float *dx, const float *dy,
                              const float *dz, uint32_t seed, uint32_t *out, size_t n, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* SCN_H_ */
