/* oracle/scn_oracle.h — TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C, scalar CPU restatement of the reference's GPIS medium-query path, used only as the
 * checker by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg.  It is never linked
 * into, imported by, or called from the product library (sparse_conv_gpis_tungsten_b200/csrc).
 *
 * Parity status: PINNED.  Every function here is checked against the UNMODIFIED reference compiled
 * from /root/reference (oracle/_ref/libscn_ref.so, see oracle/Makefile `ref`) by
 * tests/test_oracle_vs_ref.py, and against golden vectors generated from that library
 * (tests/golden/, generator tests/golden/make_golden.py).  The reference ships no tests of its own.
 *
 * The entry points deliberately use the same array-of-structs signatures as oracle/ref_harness.cpp
 * (ref_* <-> orc_*), so one test body can drive either.
 */
#ifndef SCN_ORACLE_H_
#define SCN_ORACLE_H_

#include <stddef.h>
#include <stdint.h>
#include "../include/scn.h" /* the POD descriptor only */

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_medium orc_medium;

orc_medium *orc_medium_create(const scn_medium_desc *desc);
void orc_medium_destroy(orc_medium *m);
const char *orc_last_error(void);
int orc_hw_threads(void);

uint32_t orc_xxhash32_1(uint32_t a);
uint32_t orc_xxhash32_2(uint32_t a, uint32_t b);
uint32_t orc_xxhash32_3(uint32_t a, uint32_t b, uint32_t c);
uint32_t orc_xxhash32_4(uint32_t a, uint32_t b, uint32_t c, uint32_t d);
void orc_pcg_stream(uint64_t state, int n, float *out);
void orc_pcg_next3d(uint64_t state, float *out);
void orc_box_muller(uint64_t state, int npairs, double *out);
/* impulses of one 3D cell exactly as cell3D draws them: out[4k..4k+3] = x, y, z, w */
void orc_cell_impulses(uint32_t ci, uint32_t cj, uint32_t ck, uint32_t seed, uint32_t count, float *out);
void orc_multires_level(int k, float *scale, int *additional_seed);

int orc_eval_noise3d(orc_medium *m, size_t n, const float *p, const float *d, uint32_t seed, float *out4, int nthreads);
/* per point: how many of the 27*N generated impulses pass the rejection test inside evaluateNoise3D (both levels when multi-resolution) */
int orc_accepted_counts(orc_medium *m, size_t n, const float *p, const float *d, uint32_t seed, uint32_t *out, int nthreads);
int orc_eval_noise1d(orc_medium *m, size_t n, const float *p, const float *d, uint32_t seed, float *out4, int nthreads);
int orc_eval(orc_medium *m, size_t n, const float *p, const float *d, const float *t, const uint32_t *pix, uint32_t sceneSeed,
             const float *info_t, const float *coeffs, float *out_val, float *out_grad, int *out_gpid, int nthreads);
int orc_conditioning(orc_medium *m, size_t n, const float *p, const float *d, const float *target_val, const float *target_grad,
                     const uint32_t *pix, uint32_t sceneSeed, const float *info_t, float *out_coeffs, int nthreads);
int orc_sample_distance(orc_medium *m, size_t n, const float *o, const float *d, const float *tmin, const float *tmax, const float *u,
                        const uint32_t *pix, uint32_t sceneSeed,
                        uint8_t *first_scatter, float *last_val, double *last_grad, int *bounce, int *last_gp_id, float *info_t,
                        uint8_t *ok, uint8_t *exited, float *t_out, float *cont_t, double *grad, int *gp_id, uint8_t *scheme,
                        float *coeffs, float *weight, int nthreads);
int orc_transmittance(orc_medium *m, size_t n, const float *o, const float *d, const float *tmin, const float *tmax, const float *u,
                      const uint32_t *pix, uint32_t sceneSeed,
                      const uint8_t *first_scatter, const float *last_val, const double *last_grad, const int *bounce,
                      const int *last_gp_id, const float *info_t, float *T_out, int nthreads);
int orc_nee(orc_medium *m, size_t n, const float *ray_dir, const float *normal, const float *p, const float *t_segment,
            const uint32_t *pix, uint32_t sceneSeed, const float *info_t, const float *coeffs,
            float *out_pdf, float *out_grad, int nthreads);
/* number of noise evaluations (evaluateValue/evaluateGradient calls) made by the last
 * orc_sample_distance / orc_transmittance call of this thread (bench bookkeeping) */
uint64_t orc_last_eval_count(void);

#ifdef __cplusplus
}
#endif
#endif
