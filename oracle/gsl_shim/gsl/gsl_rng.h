/* TEST INFRASTRUCTURE (oracle): minimal stand-in for <gsl/gsl_rng.h>.
 *
 * GSL is not vendored by the reference and is absent from this image
 * (SURVEY.md F2/F3).  This header declares just enough of the gsl_rng interface
 * for the reference's lib/mutgen.c, lib/util.c and lib/msprime.h to compile
 * unmodified.  Two generator types exist:
 *   - gsl_rng_mt19937 : MT19937 as GSL seeds and scales it (free-running);
 *   - gsl_rng_replay  : pops successive doubles from a caller-supplied array
 *                       (the "replay harness" of BASELINE.json north_star).
 * Nothing in the product path may include this file.
 */
#ifndef MSPB200_ORACLE_GSL_RNG_H
#define MSPB200_ORACLE_GSL_RNG_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    const char *name;
    int kind; /* 0 = mt19937, 1 = replay */
} gsl_rng_type;

typedef struct gsl_rng_struct {
    const gsl_rng_type *type;
    /* mt19937 */
    uint32_t mt[624];
    int mti;
    /* replay */
    const double *stream;
    uint64_t stream_len;
    uint64_t stream_pos;
    int exhausted; /* set when the replay stream ran dry (draws return 0.5) */
    uint64_t draws; /* total get_double calls, both kinds */
} gsl_rng;

extern const gsl_rng_type *gsl_rng_mt19937;
extern const gsl_rng_type *gsl_rng_replay;
extern const gsl_rng_type *gsl_rng_default;

gsl_rng *gsl_rng_alloc(const gsl_rng_type *T);
void gsl_rng_free(gsl_rng *r);
void gsl_rng_set(gsl_rng *r, unsigned long int seed);
void gsl_rng_replay_set(gsl_rng *r, const double *stream, uint64_t n);
double gsl_rng_uniform(gsl_rng *r);
double gsl_rng_uniform_pos(gsl_rng *r);
unsigned long int gsl_rng_get(gsl_rng *r);
unsigned long int gsl_rng_uniform_int(gsl_rng *r, unsigned long int n); /* gsl_shim_ancestry.c */

#ifdef __cplusplus
}
#endif
#endif
