/* TEST INFRASTRUCTURE ONLY (oracle/): prototypes of the two GSL chi-square tail functions used by
 * /root/reference/demod_binary.c:1161-1165,1281,1517-1545.  Implemented in oracle/shim_gsl.c from the
 * closed form for an even number of degrees of freedom (the only case the reference uses). */
#ifndef BRP_ORACLE_SHIM_GSL_CDF_H
#define BRP_ORACLE_SHIM_GSL_CDF_H
#ifdef __cplusplus
extern "C" {
#endif
double gsl_cdf_chisq_Q(double x, double nu);
double gsl_cdf_chisq_Qinv(double Q, double nu);
#ifdef __cplusplus
}
#endif
#endif
