// msprime_b200: replay-harness random source (sm_100a).
//
// In replay mode the device consumes a caller-supplied array of uniforms in
// exactly the order in which the reference consumes its generator
// (lib/mutgen.c:1103-1151 for placement, :1174-1215 for alleles), so that the
// resulting tables can be compared bit for bit with the reference run on the
// same array.  The number of uniforms a Poisson draw consumes depends on the
// draw itself, so the stream offsets of the (edge, rate interval) pairs are
// resolved by ONE sequential walker (replay_resolve_kernel); everything after
// that indexes the stream by offset and runs in parallel.
//
// The Poisson walker restates GSL 2.x's published algorithm, which is what the
// reference links against: Knuth's product method, reduced for mu > 10 by
// integer-order gamma variates (Ahrens) and finished by a BTPE binomial.
#pragma once
#include "common.cuh"

namespace mspb {

struct ReplayStream {
    const double *u;
    uint64_t len;
    uint64_t pos;
    int exhausted;
    MSPB_DEVFN double next()
    {
        if (pos >= len) {
            exhausted = 1;
            pos++;
            return 0.5;
        }
        return u[pos++];
    }
    MSPB_DEVFN double next_pos()
    {
        double x;
        do {
            x = next();
        } while (x == 0.0 && !exhausted);
        return x;
    }
};

MSPB_DEVFN double replay_gamma_large(ReplayStream &s, double a)
{
    const double sqa = sqrt(2.0 * a - 1.0);
    double x, y, v;
    do {
        do {
            y = tan(3.14159265358979323846 * s.next());
            x = sqa * y + a - 1.0;
        } while (x <= 0.0 && !s.exhausted);
        v = s.next();
    } while (v > (1.0 + y * y) * exp((a - 1.0) * log(x / (a - 1.0)) - sqa * y) && !s.exhausted);
    return x;
}

MSPB_DEVFN double replay_gamma_int(ReplayStream &s, unsigned int a)
{
    if (a < 12) {
        double prod = 1.0;
        for (unsigned int i = 0; i < a; i++) {
            prod *= s.next_pos();
        }
        return -log(prod);
    }
    return replay_gamma_large(s, (double) a);
}

MSPB_DEVFN double replay_pow_uint(double x, unsigned int n)
{
    double value = 1.0;
    do {
        if (n & 1u) {
            value *= x;
        }
        n >>= 1;
        x *= x;
    } while (n);
    return value;
}

MSPB_DEVFN double replay_stirling(double y1)
{
    const double y2 = y1 * y1;
    return (13860.0 - (462.0 - (132.0 - (99.0 - 140.0 / y2) / y2) / y2) / y2) / y1 / 166320.0;
}

MSPB_DEVFN unsigned int replay_binomial(ReplayStream &st, double p, unsigned int n)
{
    int ix = 0;
    bool flipped = false;
    if (n == 0) {
        return 0;
    }
    if (p > 0.5) {
        p = 1.0 - p;
        flipped = true;
    }
    const double q = 1.0 - p;
    const double s = p / q;
    const double np = n * p;
    if (np < 14.0) {
        const double f0 = replay_pow_uint(q, n);
        bool done = false;
        while (!done && !st.exhausted) {
            double f = f0;
            double u = st.next();
            for (ix = 0; ix <= 110; ++ix) {
                if (u < f) {
                    done = true;
                    break;
                }
                u -= f;
                f *= s * (n - ix) / (ix + 1);
            }
        }
    } else {
        const double ffm = np + p;
        const int m = (int) ffm;
        const double xm = m + 0.5;
        const double npq = np * q;
        const double p1 = floor(2.195 * sqrt(npq) - 4.6 * q) + 0.5;
        const double xl = xm - p1;
        const double xr = xm + p1;
        const double c = 0.134 + 20.5 / (15.3 + m);
        const double p2 = p1 * (1.0 + c + c);
        const double al = (ffm - xl) / (ffm - xl * p);
        const double lambda_l = al * (1.0 + 0.5 * al);
        const double ar = (xr - ffm) / (xr * q);
        const double lambda_r = ar * (1.0 + 0.5 * ar);
        const double p3 = p2 + c / lambda_l;
        const double p4 = p3 + c / lambda_r;
        while (!st.exhausted) {
            double var, accept;
            const double u = st.next() * p4;
            double v = st.next();
            if (u <= p1) {
                ix = (int) (xm - p1 * v + u);
                break;
            } else if (u <= p2) {
                const double x = xl + (u - p1) / c;
                v = v * c + 1.0 - fabs(x - xm) / p1;
                if (v > 1.0 || v <= 0.0) {
                    continue;
                }
                ix = (int) x;
            } else if (u <= p3) {
                ix = (int) (xl + log(v) / lambda_l);
                if (ix < 0) {
                    continue;
                }
                v *= ((u - p2) * lambda_l);
            } else {
                ix = (int) (xr - log(v) / lambda_r);
                if (ix > (double) n) {
                    continue;
                }
                v *= ((u - p3) * lambda_r);
            }
            const int k = abs(ix - m);
            if (k <= 20) {
                const double g = (n + 1) * s;
                double f = 1.0;
                var = v;
                if (m < ix) {
                    for (int i = m + 1; i <= ix; i++) {
                        f *= (g / i - s);
                    }
                } else if (m > ix) {
                    for (int i = ix + 1; i <= m; i++) {
                        f /= (g / i - s);
                    }
                }
                accept = f;
            } else {
                var = log(v);
                if (k < npq / 2 - 1) {
                    const double amaxp
                        = k / npq * ((k * (k / 3.0 + 0.625) + (1.0 / 6.0)) / npq + 0.5);
                    const double ynorm = -(k * k / (2.0 * npq));
                    if (var < ynorm - amaxp) {
                        break;
                    }
                    if (var > ynorm + amaxp) {
                        continue;
                    }
                }
                const double x1 = ix + 1.0;
                const double w1 = n - ix + 1.0;
                const double f1 = m + 1.0;
                const double z1 = n + 1.0 - m;
                accept = xm * log(f1 / x1) + (n - m + 0.5) * log(z1 / w1)
                         + (ix - m) * log(w1 * p / (x1 * q)) + replay_stirling(f1)
                         + replay_stirling(z1) - replay_stirling(x1) - replay_stirling(w1);
            }
            if (var <= accept) {
                break;
            }
        }
    }
    return flipped ? (n - (unsigned int) ix) : (unsigned int) ix;
}

// Consumes >= 1 uniform even when mu <= 0, as the reference's call does.
MSPB_DEVFN unsigned int replay_poisson(ReplayStream &s, double mu)
{
    double prod = 1.0;
    unsigned int k = 0;
    while (mu > 10.0) {
        const unsigned int m = (unsigned int) (mu * (7.0 / 8.0));
        const double X = replay_gamma_int(s, m);
        if (X >= mu) {
            return k + replay_binomial(s, mu / X, m - 1);
        }
        k += m;
        mu -= X;
        if (s.exhausted) {
            return k;
        }
    }
    const double emu = exp(-mu);
    do {
        prod *= s.next();
        k++;
    } while (prod > emu && !s.exhausted);
    return k - 1;
}

} // namespace mspb
