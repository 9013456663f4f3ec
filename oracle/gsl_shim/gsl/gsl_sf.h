/* TEST INFRASTRUCTURE (oracle): stand-in for <gsl/gsl_sf.h>, only so that the reference's
 * lib/msprime.c compiles unmodified for the input generator (oracle/ancestry_driver.c).
 * Nothing in the product path may include this file. */
#ifndef MSPB200_ORACLE_GSL_SF_H
#define MSPB200_ORACLE_GSL_SF_H
#ifdef __cplusplus
extern "C" {
#endif
double gsl_sf_choose(unsigned int n, unsigned int m);
double gsl_sf_lnchoose(unsigned int n, unsigned int m);
double gsl_sf_lnbeta(double a, double b);
double gsl_sf_beta_inc(double a, double b, double x);
#ifdef __cplusplus
}
#endif
#endif
