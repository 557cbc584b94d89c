/* TEST INFRASTRUCTURE ONLY (oracle/): the GSL random-number entry points named by the whitening
 * pre-step (/root/reference/demod_binary.c:872,990-991,1033).  They are only reached with -W;
 * see oracle/shim_gsl.c for what is and is not restated. */
#ifndef BRP_ORACLE_SHIM_GSL_RNG_H
#define BRP_ORACLE_SHIM_GSL_RNG_H
#ifdef __cplusplus
extern "C" {
#endif
typedef struct { int id; } gsl_rng_type;
typedef struct { unsigned long s1, s2, s3; } gsl_rng;
extern const gsl_rng_type *gsl_rng_taus2;
gsl_rng *gsl_rng_alloc(const gsl_rng_type *T);
void gsl_rng_set(const gsl_rng *r, unsigned long seed);
unsigned long gsl_rng_get(const gsl_rng *r);
void gsl_rng_free(gsl_rng *r);
#ifdef __cplusplus
}
#endif
#endif
