/* TEST INFRASTRUCTURE (oracle): stand-in for <gsl/gsl_math.h>; see gsl_rng.h.
 * <float.h> is included because the reference's lib/util.c relies on it
 * transitively. */
#ifndef MSPB200_ORACLE_GSL_MATH_H
#define MSPB200_ORACLE_GSL_MATH_H
#include <math.h>
#include <float.h>
#include <gsl/gsl_minmax.h>
#ifdef __cplusplus
extern "C" {
#endif
int gsl_fcmp(double x1, double x2, double epsilon);
int gsl_isnan(const double x);
#ifndef GSL_POSINF
#define GSL_POSINF INFINITY
#define GSL_NEGINF (-INFINITY)
#define GSL_NAN NAN
#endif
#ifdef __cplusplus
}
#endif
#endif
