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
/* the ancestry simulator (lib/msprime.c) only; defined in gsl_shim_ancestry.c */
double gsl_ran_exponential(gsl_rng *r, double mu);
unsigned int gsl_ran_geometric(gsl_rng *r, double p);
void gsl_ran_multinomial(gsl_rng *r, size_t K, unsigned int N, const double p[], unsigned int n[]);
double gsl_ran_beta(gsl_rng *r, double a, double b);
#ifdef __cplusplus
}
#endif
#endif
