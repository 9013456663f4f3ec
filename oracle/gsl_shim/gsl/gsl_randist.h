/* TEST INFRASTRUCTURE (oracle): stand-in for <gsl/gsl_randist.h>; see gsl_rng.h. */
#ifndef MSPB200_ORACLE_GSL_RANDIST_H
#define MSPB200_ORACLE_GSL_RANDIST_H
#include <gsl/gsl_rng.h>
#ifdef __cplusplus
extern "C" {
#endif
unsigned int gsl_ran_poisson(gsl_rng *r, double mu);
double gsl_ran_flat(gsl_rng *r, double a, double b);
double gsl_ran_gamma_int(gsl_rng *r, unsigned int a);
unsigned int gsl_ran_binomial(gsl_rng *r, double p, unsigned int n);
#ifdef __cplusplus
}
#endif
#endif
