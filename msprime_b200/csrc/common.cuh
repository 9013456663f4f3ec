// msprime_b200: shared device helpers (sm_100a).
//
// Counter-based random numbers (Philox4x32-10), order-preserving float keys and
// the Poisson samplers of the native stream.  The replay-harness generator that
// walks a pre-drawn uniform stream in the reference's order lives in replay.cuh.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#include <map>
#include <string>
#include <vector>

#define MSPB_DEVFN __device__ __forceinline__

namespace mspb {

// ---------------------------------------------------------------- launch log (host)
// Counts every kernel this library launches and, when timing is on, brackets each
// launch with CUDA events on the launching stream so that bench.py can report the
// average duration of the dominant kernel live (no profiler attached).
struct LaunchLog {
    struct Rec {
        const char *name;
        cudaEvent_t a, b;
    };
    struct Stat {
        uint64_t launches = 0;
        double ms = 0;
    };
    uint64_t count = 0;
    bool timing = false;
    cudaStream_t stream = nullptr;
    std::vector<Rec> pending;
    std::vector<cudaEvent_t> pool;
    std::map<std::string, Stat> stats;
    std::vector<std::string> names; // stable iteration order for the C ABI

    cudaEvent_t get_event()
    {
        if (!pool.empty()) {
            cudaEvent_t e = pool.back();
            pool.pop_back();
            return e;
        }
        cudaEvent_t e = nullptr;
        cudaEventCreate(&e);
        return e;
    }
    void begin(const char *name)
    {
        count++;
        if (timing) {
            Rec r { name, get_event(), get_event() };
            cudaEventRecord(r.a, stream);
            pending.push_back(r);
        }
    }
    void end()
    {
        if (timing) {
            cudaEventRecord(pending.back().b, stream);
        }
    }
    // call after the stream has been synchronised
    void collect()
    {
        for (const Rec &r : pending) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
                Stat &s = stats[r.name];
                if (s.launches == 0) {
                    bool known = false;
                    for (const auto &n : names) {
                        known = known || n == r.name;
                    }
                    if (!known) {
                        names.push_back(r.name);
                    }
                }
                s.launches++;
                s.ms += ms;
            }
            pool.push_back(r.a);
            pool.push_back(r.b);
        }
        pending.clear();
        cudaGetLastError();
    }
    void reset()
    {
        collect();
        stats.clear();
        names.clear();
    }
    void release()
    {
        collect();
        for (cudaEvent_t e : pool) {
            cudaEventDestroy(e);
        }
        pool.clear();
    }
};

#define MSPB_LAUNCH(log, name, ...)                                                                                    \
    do {                                                                                                               \
        (log).begin(name);                                                                                             \
        __VA_ARGS__;                                                                                                   \
        (log).end();                                                                                                   \
    } while (0)

// ---------------------------------------------------------------- Philox4x32-10
struct Philox {
    uint32_t k0, k1;
    MSPB_DEVFN Philox(uint64_t seed) : k0((uint32_t) seed), k1((uint32_t) (seed >> 32)) {}
    MSPB_DEVFN uint4 operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const
    {
        uint32_t a = k0, b = k1;
#pragma unroll
        for (int r = 0; r < 10; r++) {
            const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
            const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
            c0 = hi1 ^ c1 ^ a;
            c1 = lo1;
            c2 = hi0 ^ c3 ^ b;
            c3 = lo0;
            a += 0x9E3779B9u;
            b += 0xBB67AE85u;
        }
        return make_uint4(c0, c1, c2, c3);
    }
};

// 53-bit uniform in [0, 1) from two 32-bit words
MSPB_DEVFN double u53(uint32_t hi, uint32_t lo)
{
    const uint64_t v = ((uint64_t) (hi >> 5) << 26) | (uint64_t) (lo >> 6);
    return (double) v * (1.0 / 9007199254740992.0);
}

// purpose tags of the native stream: counter = (id_lo, id_hi, index, tag)
enum : uint32_t { TAG_COUNT = 0x10000000u, TAG_PLACE = 0x20000000u, TAG_ALLELE = 0x30000000u };

// Sequential uniforms for one (id, tag): used by the Poisson rejection loop.
struct PhiloxSeq {
    Philox ph;
    uint32_t id_lo, id_hi, tag, idx;
    uint4 buf;
    int have;
    MSPB_DEVFN PhiloxSeq(uint64_t seed, uint64_t id, uint32_t tag_)
        : ph(seed), id_lo((uint32_t) id), id_hi((uint32_t) (id >> 32)), tag(tag_), idx(0), have(0)
    {
    }
    MSPB_DEVFN double next()
    {
        if (have == 0) {
            buf = ph(id_lo, id_hi, idx++, tag);
            have = 1;
            return u53(buf.x, buf.y);
        }
        have = 0;
        return u53(buf.z, buf.w);
    }
};

// ---------------------------------------------------------------- float keys
// Monotone map double -> uint64 (negative values and -0.0 handled).
MSPB_DEVFN uint64_t ordered_bits(double x)
{
    uint64_t b = (uint64_t) __double_as_longlong(x + 0.0);
    return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}

// ---------------------------------------------------------------- Poisson (native)
// mu < 12: inversion by sequential search (one uniform).  Otherwise Hoermann's
// PTRS transformed rejection (Insurance: Mathematics and Economics 12, 1993).
constexpr double POISSON_SMALL_MU = 12.0;

// mu < POISSON_SMALL_MU: inversion by sequential search on one uniform
MSPB_DEVFN uint32_t poisson_small(double u, double mu)
{
    double p = exp(-mu), F = p;
    uint32_t k = 0;
    while (u > F && k < 200) {
        k++;
        p *= mu / (double) k;
        F += p;
    }
    return k;
}

MSPB_DEVFN uint32_t poisson_native(PhiloxSeq &rng, double mu)
{
    if (!(mu > 0.0)) {
        return 0;
    }
    if (mu < POISSON_SMALL_MU) {
        return poisson_small(rng.next(), mu);
    }
    const double slam = sqrt(mu), loglam = log(mu);
    const double b = 0.931 + 2.53 * slam;
    const double a = -0.059 + 0.02483 * b;
    const double invalpha = 1.1239 + 1.1328 / (b - 3.4);
    const double vr = 0.9277 - 3.6224 / (b - 2.0);
    for (int it = 0; it < 10000; it++) {
        const double U = rng.next() - 0.5;
        const double V = rng.next();
        const double us = 0.5 - fabs(U);
        const double kf = floor((2.0 * a / us + b) * U + mu + 0.43);
        if (us >= 0.07 && V <= vr) {
            return (uint32_t) kf;
        }
        if (kf < 0.0 || (us < 0.013 && V > us)) {
            continue;
        }
        if (log(V) + log(invalpha) - log(a / (us * us) + b) <= -mu + kf * loglam - lgamma(kf + 1.0)) {
            return (uint32_t) kf;
        }
    }
    return (uint32_t) mu;
}

// ---------------------------------------------------------------- rate map
// Device view of the rate map plus cumulative mass at the breakpoints.  In the
// discrete genome the mass of an interval sits on the integers it contains
// (lib/mutgen.c:1079-1085), so cum[] counts ceil() differences.
struct RateMapView {
    const double *pos; // size + 1
    const double *rate; // size + 1 (rate[size] = 0)
    const double *cum; // size + 1
    uint32_t size;
    int discrete;

    // largest i with pos[i] <= x (lib/rate_map.c:137-146)
    MSPB_DEVFN uint32_t index(double x) const
    {
        uint32_t lo = 0, hi = size + 1;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if (pos[mid] <= x) {
                lo = mid + 1;
            } else {
                hi = mid;
            }
        }
        return lo == 0 ? 0 : lo - 1;
    }
    MSPB_DEVFN double coord(double x) const { return discrete ? ceil(x) : x; }
    // mass of [0, x)
    MSPB_DEVFN double mass_upto(double x) const
    {
        if (size == 1) {
            return rate[0] * coord(x);
        }
        const uint32_t i = index(x);
        return cum[i] + rate[i] * (coord(x) - coord(pos[i]));
    }
    // position at cumulative mass m, restricted to the edge [left, right)
    MSPB_DEVFN double position_at(double m, double left, double right) const
    {
        uint32_t i = 0;
        if (size > 1) {
            // last i with cum[i] <= m, skipping zero-mass intervals
            uint32_t lo = 0, hi = size;
            while (lo < hi) {
                const uint32_t mid = (lo + hi) >> 1;
                if (cum[mid + 1] <= m) {
                    lo = mid + 1;
                } else {
                    hi = mid;
                }
            }
            i = lo < size ? lo : size - 1;
        }
        const double a = coord(fmax(left, pos[i]));
        const double b = coord(fmin(right, pos[i + 1]));
        double x = coord(pos[i]) + (m - cum[i]) / rate[i];
        if (discrete) {
            x = floor(x);
            if (x > b - 1.0) {
                x = b - 1.0;
            }
            if (x < a) {
                x = a;
            }
        } else {
            if (!(x < b)) {
                x = nextafter(b, -1.7976931348623157e308);
            }
            if (x < a) {
                x = a;
            }
        }
        return x;
    }
};

} // namespace mspb

namespace mspb {

// ---------------------------------------------------------------- decoupled look-back
// Single-pass prefix sums over tiles (Merrill & Garland): a tile publishes its
// aggregate, then sums the aggregates of the tiles before it until it meets one
// whose inclusive prefix is already known.  T is a small POD with operator+ and a
// multiple of 4 bytes in size.  Flags: 0 = nothing yet, 1 = aggregate, 2 = inclusive.
template <typename T>
struct Lookback {
    uint32_t *flag;
    T *agg;
    T *incl;
};

template <typename T>
MSPB_DEVFN T ld_cg_pod(const T *p)
{
    static_assert(sizeof(T) % 4 == 0, "POD of 32-bit words");
    T v;
    uint32_t *w = reinterpret_cast<uint32_t *>(&v);
    const uint32_t *s = reinterpret_cast<const uint32_t *>(p);
#pragma unroll
    for (int i = 0; i < (int) (sizeof(T) / 4); i++) {
        w[i] = __ldcg(s + i);
    }
    return v;
}

template <typename T>
MSPB_DEVFN void st_cg_pod(T *p, const T &v)
{
    const uint32_t *w = reinterpret_cast<const uint32_t *>(&v);
    uint32_t *d = reinterpret_cast<uint32_t *>(p);
#pragma unroll
    for (int i = 0; i < (int) (sizeof(T) / 4); i++) {
        __stcg(d + i, w[i]);
    }
}

template <typename T>
MSPB_DEVFN T shfl_xor_pod(const T &v, int mask)
{
    T r;
    const uint32_t *w = reinterpret_cast<const uint32_t *>(&v);
    uint32_t *o = reinterpret_cast<uint32_t *>(&r);
#pragma unroll
    for (int i = 0; i < (int) (sizeof(T) / 4); i++) {
        o[i] = __shfl_xor_sync(0xffffffffu, w[i], mask);
    }
    return r;
}

// Called by every thread of the block with the block's aggregate (thread 0's value is
// used); returns the sum over all earlier tiles.  *total (optional, device memory)
// receives the grand total from the last tile.  `bcast` is one T in shared memory.
template <typename T>
MSPB_DEVFN T lookback_exclusive(const Lookback<T> &lb, uint32_t tile, uint32_t num_tiles, const T &aggregate, T *bcast,
    T *total)
{
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        T prefix = T();
        if (tile > 0) {
            if (lane == 0) {
                st_cg_pod(&lb.agg[tile], aggregate);
                __threadfence();
                *((volatile uint32_t *) &lb.flag[tile]) = 1u;
            }
            int64_t base = (int64_t) tile - 1;
            for (;;) {
                const int64_t p = base - lane;
                uint32_t f = 2u;
                if (p >= 0) {
                    do {
                        f = *((volatile uint32_t *) &lb.flag[p]);
                    } while (f == 0u);
                }
                __threadfence();
                const uint32_t incl_mask = __ballot_sync(0xffffffffu, f == 2u);
                const int stop = incl_mask != 0u ? (__ffs(incl_mask) - 1) : 32;
                T v = T();
                if (p >= 0 && lane <= stop) {
                    v = f == 2u ? ld_cg_pod(&lb.incl[p]) : ld_cg_pod(&lb.agg[p]);
                }
#pragma unroll
                for (int s = 16; s > 0; s >>= 1) {
                    v = v + shfl_xor_pod(v, s);
                }
                prefix = prefix + v;
                if (incl_mask != 0u) {
                    break;
                }
                base -= 32;
            }
        }
        if (lane == 0) {
            const T inclusive = prefix + aggregate;
            st_cg_pod(&lb.incl[tile], inclusive);
            __threadfence();
            *((volatile uint32_t *) &lb.flag[tile]) = 2u;
            *bcast = prefix;
            if (total != nullptr && tile + 1 == num_tiles) {
                *total = inclusive;
            }
        }
    }
    __syncthreads();
    const T r = *bcast;
    __syncthreads();
    return r;
}

// ---------------------------------------------------------------- records
// One per kept site ("marker"), kept mutation and new mutation, indexed by slot.
struct __align__(16) Rec {
    double time;
    int32_t node;
    uint32_t ord; // new mutations: global ordinal (counter of the allele draws)
};

MSPB_DEVFN double from_ordered_bits(uint64_t b)
{
    b = (b & 0x8000000000000000ull) ? (b & 0x7fffffffffffffffull) : ~b;
    return __longlong_as_double((long long) b);
}

} // namespace mspb
