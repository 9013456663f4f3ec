/* TEST INFRASTRUCTURE (oracle): the few further GSL routines the reference's ancestry simulator
 * (lib/msprime.c) calls, so that it compiles unmodified for the INPUT GENERATOR of the parity tests
 * (oracle/ancestry_driver.c, `make ancestry`).  The mutation path never reaches this file.
 *
 * GSL is a dependency of the reference that is absent from /root/reference and from this image
 * (SURVEY.md F2/F3).  The routines the standard (Hudson) coalescent draws from are restated from GSL's
 * published algorithms:
 *   gsl_rng_uniform_int   rng/rng.c      : scale = range / n on the raw 32-bit output, rejection above n
 *   gsl_ran_exponential   randist/exponential.c : -mu * log1p(-u)
 *   gsl_ran_geometric     randist/geometric.c   : floor(log(u) / log(1 - p)) + 1 with u in (0, 1)
 *   gsl_ran_multinomial   randist/multinomial.c : conditional binomials
 * The special functions behind the multiple-merger coalescents (beta / Dirac) are provided through lgamma
 * where that is exact enough for a generator (choose, lnchoose, lnbeta); the regularised incomplete beta
 * function, its inverse and the beta variate are NOT provided: they stop the process with a message, and
 * the driver only ever selects the Hudson model.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <gsl/gsl_cdf.h>
#include <gsl/gsl_math.h>
#include <gsl/gsl_randist.h>
#include <gsl/gsl_rng.h>
#include <gsl/gsl_sf.h>

static void
not_provided(const char *what)
{
    fprintf(stderr, "oracle GSL shim: %s is not provided (only the Hudson model is supported)\n", what);
    abort();
}

unsigned long int
gsl_rng_uniform_int(gsl_rng *r, unsigned long int n)
{
    /* mt19937: min 0, max 0xffffffff */
    const unsigned long int range = 0xffffffffUL;
    unsigned long int scale, k;

    if (n == 0 || n > range) {
        not_provided("gsl_rng_uniform_int with n outside [1, 2^32 - 1]");
    }
    if (r->type->kind != 0) {
        /* replay stream: one double per draw */
        k = (unsigned long int) (gsl_rng_uniform(r) * (double) n);
        return k < n ? k : n - 1;
    }
    scale = range / n;
    do {
        k = gsl_rng_get(r) / scale;
    } while (k >= n);
    return k;
}

double
gsl_ran_exponential(gsl_rng *r, double mu)
{
    double u = gsl_rng_uniform(r);
    return -mu * log1p(-u);
}

unsigned int
gsl_ran_geometric(gsl_rng *r, double p)
{
    double u = gsl_rng_uniform_pos(r);
    if (p == 1) {
        return 1;
    }
    return (unsigned int) (log(u) / log(1 - p) + 1);
}

void
gsl_ran_multinomial(gsl_rng *r, size_t K, unsigned int N, const double p[], unsigned int n[])
{
    size_t k;
    double norm = 0.0, sum_p = 0.0;
    unsigned int sum_n = 0;

    for (k = 0; k < K; k++) {
        norm += p[k];
    }
    for (k = 0; k < K; k++) {
        if (p[k] > 0.0) {
            n[k] = gsl_ran_binomial(r, p[k] / (norm - sum_p), N - sum_n);
        } else {
            n[k] = 0;
        }
        sum_p += p[k];
        sum_n += n[k];
    }
}

int
gsl_isnan(const double x)
{
    return isnan(x) ? 1 : 0;
}

double
gsl_sf_lnchoose(unsigned int n, unsigned int m)
{
    if (m > n) {
        not_provided("gsl_sf_lnchoose with m > n");
    }
    return lgamma(n + 1.0) - lgamma(m + 1.0) - lgamma(n - m + 1.0);
}

double
gsl_sf_choose(unsigned int n, unsigned int m)
{
    double prod = 1.0;
    unsigned int k;

    if (m > n) {
        not_provided("gsl_sf_choose with m > n");
    }
    if (m * 2 > n) {
        m = n - m;
    }
    for (k = 1; k <= m; k++) {
        prod *= (double) (n - m + k) / (double) k;
    }
    return prod;
}

double
gsl_sf_lnbeta(double a, double b)
{
    return lgamma(a) + lgamma(b) - lgamma(a + b);
}

double
gsl_sf_beta_inc(double a, double b, double x)
{
    (void) a; (void) b; (void) x;
    not_provided("gsl_sf_beta_inc");
    return 0;
}

double
gsl_cdf_beta_P(double x, double a, double b)
{
    (void) a; (void) b; (void) x;
    not_provided("gsl_cdf_beta_P");
    return 0;
}

double
gsl_cdf_beta_Pinv(double p, double a, double b)
{
    (void) a; (void) b; (void) p;
    not_provided("gsl_cdf_beta_Pinv");
    return 0;
}

double
gsl_ran_beta(gsl_rng *r, double a, double b)
{
    (void) r; (void) a; (void) b;
    not_provided("gsl_ran_beta");
    return 0;
}
