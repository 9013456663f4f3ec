/* TEST INFRASTRUCTURE (oracle): stand-in for <gsl/gsl_cdf.h>, only so that the reference's
 * lib/msprime.c compiles unmodified for the input generator (oracle/ancestry_driver.c).
 * Nothing in the product path may include this file. */
#ifndef MSPB200_ORACLE_GSL_CDF_H
#define MSPB200_ORACLE_GSL_CDF_H
#ifdef __cplusplus
extern "C" {
#endif
double gsl_cdf_beta_P(double x, double a, double b);
double gsl_cdf_beta_Pinv(double p, double a, double b);
#ifdef __cplusplus
}
#endif
#endif
