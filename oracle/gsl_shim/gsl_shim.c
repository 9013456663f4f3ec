/* TEST INFRASTRUCTURE (oracle).  Not part of the product path.
 *
 * Bodies for the handful of GSL entry points that the reference's mutation
 * path calls:
 *   gsl_ran_poisson   <- lib/mutgen.c:1120
 *   gsl_ran_flat      <- lib/util.c:731 (msp_gsl_ran_flat)
 *   gsl_fcmp          <- lib/util.c:375 (doubles_almost_equal)
 *   gsl_rng_alloc/set <- msprime/_msprimemodule.c:265,248 (default type = mt19937)
 *
 * GSL itself is a third-party dependency of the reference; it is not vendored,
 * not version-pinned and absent from this image.  The algorithms below restate
 * GSL 2.x's *published* methods (Matsumoto-Nishimura MT19937 with the 2002
 * initialiser; Knuth's product method for Poisson with the gamma/binomial
 * reduction for mu > 10; Ahrens' rejection for integer-order gamma; the
 * Kachitvichyanukul-Schmeiser BTPE binomial; Knuth's relative float compare).
 * They could not be checked against a real libgsl on this box: PARITY UNPINNED
 * at the GSL boundary for the mu > 10 branch.  Bit-exactness claims elsewhere
 * in this repository are therefore "reference mutgen.c + this shim".
 * tests/test_gsl_shim.py holds what can be held without libgsl: MT19937 bit for
 * bit against numpy's and its published check value, the distributions against
 * scipy's exact ones.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <gsl/gsl_rng.h>
#include <gsl/gsl_randist.h>
#include <gsl/gsl_math.h>

static const gsl_rng_type mt_type = { "mt19937", 0 };
static const gsl_rng_type replay_type = { "replay", 1 };
const gsl_rng_type *gsl_rng_mt19937 = &mt_type;
const gsl_rng_type *gsl_rng_replay = &replay_type;
const gsl_rng_type *gsl_rng_default = &mt_type;

/* ---------------- MT19937 ---------------- */

#define MT_N 624
#define MT_M 397

static void
mt_seed(gsl_rng *r, unsigned long int s)
{
    int i;
    if (s == 0) {
        s = 4357; /* GSL's default seed */
    }
    r->mt[0] = (uint32_t)(s & 0xffffffffUL);
    for (i = 1; i < MT_N; i++) {
        r->mt[i] = (uint32_t)(1812433253UL * (r->mt[i - 1] ^ (r->mt[i - 1] >> 30)) + (uint32_t) i);
    }
    r->mti = MT_N;
}

static uint32_t
mt_next(gsl_rng *r)
{
    uint32_t y;
    if (r->mti >= MT_N) {
        int kk;
        for (kk = 0; kk < MT_N; kk++) {
            uint32_t a = r->mt[kk];
            uint32_t b = r->mt[(kk + 1) % MT_N];
            y = (a & 0x80000000UL) | (b & 0x7fffffffUL);
            r->mt[kk] = r->mt[(kk + MT_M) % MT_N] ^ (y >> 1) ^ ((y & 1U) ? 0x9908b0dfUL : 0U);
        }
        r->mti = 0;
    }
    y = r->mt[r->mti++];
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9d2c5680UL;
    y ^= (y << 15) & 0xefc60000UL;
    y ^= (y >> 18);
    return y;
}

/* ---------------- generator plumbing ---------------- */

gsl_rng *
gsl_rng_alloc(const gsl_rng_type *T)
{
    gsl_rng *r = calloc(1, sizeof(*r));
    if (r != NULL) {
        r->type = T;
        if (T->kind == 0) {
            mt_seed(r, 0);
        }
    }
    return r;
}

void
gsl_rng_free(gsl_rng *r)
{
    free(r);
}

void
gsl_rng_set(gsl_rng *r, unsigned long int seed)
{
    if (r->type->kind == 0) {
        mt_seed(r, seed);
    }
    r->draws = 0;
}

void
gsl_rng_replay_set(gsl_rng *r, const double *stream, uint64_t n)
{
    r->stream = stream;
    r->stream_len = n;
    r->stream_pos = 0;
    r->exhausted = 0;
    r->draws = 0;
}

unsigned long int
gsl_rng_get(gsl_rng *r)
{
    r->draws++;
    if (r->type->kind == 0) {
        return mt_next(r);
    }
    return (unsigned long int) (gsl_rng_uniform(r) * 4294967296.0);
}

double
gsl_rng_uniform(gsl_rng *r)
{
    r->draws++;
    if (r->type->kind == 0) {
        return mt_next(r) / 4294967296.0;
    }
    if (r->stream_pos >= r->stream_len) {
        r->exhausted = 1;
        return 0.5;
    }
    return r->stream[r->stream_pos++];
}

double
gsl_rng_uniform_pos(gsl_rng *r)
{
    double x;
    do {
        x = gsl_rng_uniform(r);
    } while (x == 0);
    return x;
}

/* ---------------- distributions ---------------- */

double
gsl_ran_flat(gsl_rng *r, double a, double b)
{
    double u = gsl_rng_uniform(r);
    return a * (1 - u) + b * u;
}

/* Ahrens' rejection from a Cauchy envelope (as reported in Knuth vol 2), a > 1. */
static double
gamma_large(gsl_rng *r, double a)
{
    double sqa = sqrt(2 * a - 1);
    double x, y, v;
    do {
        do {
            y = tan(M_PI * gsl_rng_uniform(r));
            x = sqa * y + a - 1;
        } while (x <= 0);
        v = gsl_rng_uniform(r);
    } while (v > (1 + y * y) * exp((a - 1) * log(x / (a - 1)) - sqa * y));
    return x;
}

double
gsl_ran_gamma_int(gsl_rng *r, unsigned int a)
{
    if (a < 12) {
        unsigned int i;
        double prod = 1;
        for (i = 0; i < a; i++) {
            prod *= gsl_rng_uniform_pos(r);
        }
        return -log(prod);
    }
    return gamma_large(r, (double) a);
}

static double
pow_uint(double x, unsigned int n)
{
    double value = 1.0;
    do {
        if (n & 1) {
            value *= x;
        }
        n >>= 1;
        x *= x;
    } while (n);
    return value;
}

/* Tail of the Stirling series used by BTPE's final acceptance test. */
static double
stirling_tail(double y1)
{
    double y2 = y1 * y1;
    return (13860.0 - (462.0 - (132.0 - (99.0 - 140.0 / y2) / y2) / y2) / y2) / y1 / 166320.0;
}

#define BTPE_SMALL_MEAN 14
#define BTPE_BINV_CUTOFF 110
#define BTPE_FAR_FROM_MEAN 20

unsigned int
gsl_ran_binomial(gsl_rng *r, double p, unsigned int n)
{
    int ix = 0;
    int flipped = 0;
    double q, s, np;

    if (n == 0) {
        return 0;
    }
    if (p > 0.5) {
        p = 1.0 - p;
        flipped = 1;
    }
    q = 1 - p;
    s = p / q;
    np = n * p;

    if (np < BTPE_SMALL_MEAN) {
        /* inverse-CDF walk from x = 0 */
        double f0 = pow_uint(q, n);
        int done = 0;
        while (!done) {
            double f = f0;
            double u = gsl_rng_uniform(r);
            for (ix = 0; ix <= BTPE_BINV_CUTOFF; ++ix) {
                if (u < f) {
                    done = 1;
                    break;
                }
                u -= f;
                f *= s * (n - ix) / (ix + 1);
            }
        }
    } else {
        double ffm = np + p;
        int m = (int) ffm;
        double xm = m + 0.5;
        double npq = np * q;
        double p1 = floor(2.195 * sqrt(npq) - 4.6 * q) + 0.5;
        double xl = xm - p1;
        double xr = xm + p1;
        double c = 0.134 + 20.5 / (15.3 + m);
        double p2 = p1 * (1.0 + c + c);
        double al = (ffm - xl) / (ffm - xl * p);
        double lambda_l = al * (1.0 + 0.5 * al);
        double ar = (xr - ffm) / (xr * q);
        double lambda_r = ar * (1.0 + 0.5 * ar);
        double p3 = p2 + c / lambda_l;
        double p4 = p3 + c / lambda_r;

        for (;;) {
            double var, accept;
            double u = gsl_rng_uniform(r) * p4;
            double v = gsl_rng_uniform(r);
            int k;

            if (u <= p1) {
                /* triangle: always accepted */
                ix = (int) (xm - p1 * v + u);
                break;
            } else if (u <= p2) {
                /* parallelograms */
                double x = xl + (u - p1) / c;
                v = v * c + 1.0 - fabs(x - xm) / p1;
                if (v > 1.0 || v <= 0.0) {
                    continue;
                }
                ix = (int) x;
            } else if (u <= p3) {
                /* left exponential tail */
                ix = (int) (xl + log(v) / lambda_l);
                if (ix < 0) {
                    continue;
                }
                v *= ((u - p2) * lambda_l);
            } else {
                /* right exponential tail */
                ix = (int) (xr - log(v) / lambda_r);
                if (ix > (double) n) {
                    continue;
                }
                v *= ((u - p3) * lambda_r);
            }

            /* accept iff v <= f(ix)/f(m) */
            k = abs(ix - m);
            if (k <= BTPE_FAR_FROM_MEAN) {
                double g = (n + 1) * s;
                double f = 1.0;
                int i;
                var = v;
                if (m < ix) {
                    for (i = m + 1; i <= ix; i++) {
                        f *= (g / i - s);
                    }
                } else if (m > ix) {
                    for (i = ix + 1; i <= m; i++) {
                        f /= (g / i - s);
                    }
                }
                accept = f;
            } else {
                var = log(v);
                if (k < npq / 2 - 1) {
                    double amaxp
                        = k / npq * ((k * (k / 3.0 + 0.625) + (1.0 / 6.0)) / npq + 0.5);
                    double ynorm = -(k * k / (2.0 * npq));
                    if (var < ynorm - amaxp) {
                        break;
                    }
                    if (var > ynorm + amaxp) {
                        continue;
                    }
                }
                {
                    double x1 = ix + 1.0;
                    double w1 = n - ix + 1.0;
                    double f1 = m + 1.0;
                    double z1 = n + 1.0 - m;
                    accept = xm * log(f1 / x1) + (n - m + 0.5) * log(z1 / w1)
                             + (ix - m) * log(w1 * p / (x1 * q)) + stirling_tail(f1)
                             + stirling_tail(z1) - stirling_tail(x1) - stirling_tail(w1);
                }
            }
            if (var <= accept) {
                break;
            }
        }
    }
    return flipped ? (n - (unsigned int) ix) : (unsigned int) ix;
}

/* Knuth's product-of-uniforms method; for mu > 10 peel off a gamma-distributed
 * chunk of the unit-rate process (m = 7/8 mu arrivals) and recurse, finishing
 * with a binomial when the chunk overshoots.  Consumes >= 1 uniform even when
 * mu <= 0 (this matters for the replay stream: SURVEY.md section 3.2 step 5). */
unsigned int
gsl_ran_poisson(gsl_rng *r, double mu)
{
    double emu;
    double prod = 1.0;
    unsigned int k = 0;

    while (mu > 10) {
        unsigned int m = (unsigned int) (mu * (7.0 / 8.0));
        double X = gsl_ran_gamma_int(r, m);
        if (X >= mu) {
            return k + gsl_ran_binomial(r, mu / X, m - 1);
        }
        k += m;
        mu -= X;
    }
    emu = exp(-mu);
    do {
        prod *= gsl_rng_uniform(r);
        k++;
    } while (prod > emu);
    return k - 1;
}

/* Knuth's relative comparison (Seminumerical Algorithms 4.2.2). */
int
gsl_fcmp(double x1, double x2, double epsilon)
{
    int exponent;
    double delta, difference;
    double mx = (fabs(x1) > fabs(x2)) ? x1 : x2;
    frexp(mx, &exponent);
    delta = ldexp(epsilon, exponent);
    difference = x1 - x2;
    if (difference > delta) {
        return 1;
    }
    if (difference < -delta) {
        return -1;
    }
    return 0;
}
