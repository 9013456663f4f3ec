/* TEST INFRASTRUCTURE (oracle): stand-in for <gsl/gsl_statistics_int.h>, only so that the reference's
 * lib/msprime.c compiles unmodified for the input generator (oracle/ancestry_driver.c).
 * Nothing in the product path may include this file. */
#ifndef MSPB200_ORACLE_GSL_STATISTICS_INT_H
#define MSPB200_ORACLE_GSL_STATISTICS_INT_H
#ifdef __cplusplus
extern "C" {
#endif
/* nothing on the Hudson path uses it */
#ifdef __cplusplus
}
#endif
#endif
