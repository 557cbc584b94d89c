/* TEST INFRASTRUCTURE ONLY (oracle/): the three GSL helpers the reference's hot-path
 * translation units use (gsl_pow_2: demod_binary_fft_fftw.c:107; GSL_MIN_*: demod_binary.c:1097-1099).
 * GSL itself is not installed here; these are the documented definitions. */
#ifndef BRP_ORACLE_SHIM_GSL_MATH_H
#define BRP_ORACLE_SHIM_GSL_MATH_H
static inline double gsl_pow_2(const double x) { return x * x; }
#define GSL_MIN_INT(a, b) ((int)(a) < (int)(b) ? (int)(a) : (int)(b))
#define GSL_MIN_DBL(a, b) ((double)(a) < (double)(b) ? (double)(a) : (double)(b))
#endif
