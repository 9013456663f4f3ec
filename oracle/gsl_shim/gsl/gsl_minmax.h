/* TEST INFRASTRUCTURE (oracle): stand-in for <gsl/gsl_minmax.h>; see gsl_rng.h. */
#ifndef MSPB200_ORACLE_GSL_MINMAX_H
#define MSPB200_ORACLE_GSL_MINMAX_H
#define GSL_MAX(a, b) ((a) > (b) ? (a) : (b))
#define GSL_MIN(a, b) ((a) < (b) ? (a) : (b))
#endif
