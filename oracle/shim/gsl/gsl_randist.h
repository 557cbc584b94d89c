/* TEST INFRASTRUCTURE ONLY (oracle/): see gsl_rng.h. */
#ifndef BRP_ORACLE_SHIM_GSL_RANDIST_H
#define BRP_ORACLE_SHIM_GSL_RANDIST_H
#include <gsl/gsl_rng.h>
#ifdef __cplusplus
extern "C" {
#endif
double gsl_ran_gaussian_ziggurat(const gsl_rng *r, double sigma);
#ifdef __cplusplus
}
#endif
#endif
