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
enum : uint32_t {
    TAG_COUNT = 0x10000000u,
    TAG_PLACE = 0x20000000u,
    TAG_ALLELE = 0x30000000u,
    TAG_REDRAW = 0x40000000u // + round: position redraws after a collision (continuous genome)
};

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

// 1 / k for the inversion loop (a double division costs ~30 instructions, and the warp
// iterates as long as its slowest lane)
__constant__ double POISSON_RCP[64] = { 0.0, 1.0 / 1, 1.0 / 2, 1.0 / 3, 1.0 / 4, 1.0 / 5, 1.0 / 6, 1.0 / 7, 1.0 / 8,
    1.0 / 9, 1.0 / 10, 1.0 / 11, 1.0 / 12, 1.0 / 13, 1.0 / 14, 1.0 / 15, 1.0 / 16, 1.0 / 17, 1.0 / 18, 1.0 / 19,
    1.0 / 20, 1.0 / 21, 1.0 / 22, 1.0 / 23, 1.0 / 24, 1.0 / 25, 1.0 / 26, 1.0 / 27, 1.0 / 28, 1.0 / 29, 1.0 / 30,
    1.0 / 31, 1.0 / 32, 1.0 / 33, 1.0 / 34, 1.0 / 35, 1.0 / 36, 1.0 / 37, 1.0 / 38, 1.0 / 39, 1.0 / 40, 1.0 / 41,
    1.0 / 42, 1.0 / 43, 1.0 / 44, 1.0 / 45, 1.0 / 46, 1.0 / 47, 1.0 / 48, 1.0 / 49, 1.0 / 50, 1.0 / 51, 1.0 / 52,
    1.0 / 53, 1.0 / 54, 1.0 / 55, 1.0 / 56, 1.0 / 57, 1.0 / 58, 1.0 / 59, 1.0 / 60, 1.0 / 61, 1.0 / 62, 1.0 / 63 };

// mu < POISSON_SMALL_MU: inversion by sequential search on one uniform
MSPB_DEVFN uint32_t poisson_small(double u, double mu)
{
    double p = exp(-mu), F = p;
    uint32_t k = 0;
    while (u > F && k < 200) {
        k++;
        p *= mu * (k < 64 ? POISSON_RCP[k] : 1.0 / (double) k);
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
            return kf < 4294967295.0 ? (uint32_t) kf : 0xffffffffu; // (K1 rejects totals beyond 31 bits)
        }
        if (kf < 0.0 || (us < 0.013 && V > us)) {
            continue;
        }
        if (log(V) + log(invalpha) - log(a / (us * us) + b) <= -mu + kf * loglam - lgamma(kf + 1.0)) {
            return kf < 4294967295.0 ? (uint32_t) kf : 0xffffffffu;
        }
    }
    return mu < 4294967295.0 ? (uint32_t) mu : 0xffffffffu;
}

// ---------------------------------------------------------------- rate map
// Device view of the rate map plus cumulative mass at the breakpoints.  In the
// discrete genome the mass of an interval sits on the integers it contains
// (lib/mutgen.c:1079-1085), so cum[] counts ceil() differences.
struct RateMapView {
    const double *pos; // size + 1
    const double *rate; // size + 1 (rate[size] = 0)
    const double *cum; // size + 1
    const double *inv_rate; // size + 1: 1 / rate, 0 where the rate is 0 (no mass: never selected)
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
        if (size == 1) {
            // one interval [0, L): the general path below with i = 0, cum[0] = 0 and coord(pos[0]) = 0
            // (bit for bit the same arithmetic, without its loads of pos / cum)
            const double a = coord(left), b = coord(right);
            double x = m * inv_rate[0];
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
        double x = coord(pos[i]) + (m - cum[i]) * inv_rate[i]; // (the clamps below absorb the rounding)
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
// whose inclusive prefix is already known.
//
// A tile's state is HALVES 16-byte words, each written and read with ONE 128-bit
// access, with the status in the top two bits of its last 32-bit lane (0 = nothing yet,
// 1 = aggregate, 2 = inclusive prefix), so value and status always travel together and
// no fence is needed.  The chain advances by at most one window of tiles per L2 round
// trip, so the window is wide (LB_WINDOW predecessors per round) -- with a 32-tile
// window the propagation rate, not the memory system, bounded every scan kernel here.
struct __align__(16) LbWord {
    uint32_t w[4];
};

constexpr int LB_PER_LANE = 4;
constexpr int LB_WINDOW = 32 * LB_PER_LANE;
constexpr uint32_t LB_AGG = 1u, LB_INCL = 2u;

MSPB_DEVFN LbWord lb_load(const LbWord *p)
{
    LbWord v;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3])
                 : "l"(p)
                 : "memory");
    return v;
}

MSPB_DEVFN void lb_store(LbWord *p, const LbWord &v)
{
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.w[0]), "r"(v.w[1]), "r"(v.w[2]),
                 "r"(v.w[3])
                 : "memory");
}

// Payload of a tile: HALVES x 4 lanes of 32 bits, summed lane-wise as 4 x u32 (WIDE64 =
// false) or 2 x u64 (true).  The last 32-bit lane of every half must stay below 2^30.
template <int HALVES, bool WIDE64>
struct LbVal {
    uint32_t w[HALVES][4];
    MSPB_DEVFN void clear()
    {
#pragma unroll
        for (int h = 0; h < HALVES; h++) {
            w[h][0] = w[h][1] = w[h][2] = w[h][3] = 0;
        }
    }
    MSPB_DEVFN void add(const LbVal &o)
    {
#pragma unroll
        for (int h = 0; h < HALVES; h++) {
            if (WIDE64) {
#pragma unroll
                for (int q = 0; q < 2; q++) {
                    const uint64_t a = ((uint64_t) w[h][2 * q + 1] << 32) | w[h][2 * q];
                    const uint64_t b = ((uint64_t) o.w[h][2 * q + 1] << 32) | o.w[h][2 * q];
                    const uint64_t c = a + b;
                    w[h][2 * q] = (uint32_t) c;
                    w[h][2 * q + 1] = (uint32_t) (c >> 32);
                }
            } else {
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    w[h][q] += o.w[h][q];
                }
            }
        }
    }
};

// Look-back by ONE full warp (lane = threadIdx.x & 31) with the tile's aggregate (the same
// value on every lane); returns the sum over all earlier tiles on every lane.  `state`
// holds num_tiles x HALVES words, zeroed before the launch.  The rest of the block may do
// independent work meanwhile.
template <int HALVES, bool WIDE64>
MSPB_DEVFN LbVal<HALVES, WIDE64> lookback_warp(LbWord *state, uint32_t tile, const LbVal<HALVES, WIDE64> &aggregate)
{
    typedef LbVal<HALVES, WIDE64> V;
    const int lane = threadIdx.x & 31;
    V prefix;
    prefix.clear();
    if (tile > 0) {
#pragma unroll
        for (int h = 0; h < HALVES; h++) {
            if (lane == h) {
                LbWord me;
                me.w[0] = aggregate.w[h][0];
                me.w[1] = aggregate.w[h][1];
                me.w[2] = aggregate.w[h][2];
                me.w[3] = aggregate.w[h][3] | (LB_AGG << 30);
                lb_store(&state[(uint64_t) tile * HALVES + h], me);
            }
        }
        int64_t base = (int64_t) tile - 1;
        for (;;) {
            V part;
            part.clear();
            int stop = LB_WINDOW; // offset of the nearest predecessor with an inclusive prefix
#pragma unroll
            for (int r = 0; r < LB_PER_LANE; r++) {
                if (stop != LB_WINDOW) {
                    break; // (warp-uniform) an inclusive prefix was found among nearer tiles
                }
                const int64_t p = base - (r * 32 + lane);
                V v;
                v.clear();
                uint32_t flag = LB_INCL; // before the first tile: an inclusive prefix of zero
                if (p >= 0) {
                    for (;;) {
                        bool ok = true;
                        flag = 0;
#pragma unroll
                        for (int h = 0; h < HALVES; h++) {
                            const LbWord x = lb_load(&state[(uint64_t) p * HALVES + h]);
                            const uint32_t f = x.w[3] >> 30;
                            ok = ok && f != 0 && (h == 0 || f == flag);
                            flag = f;
                            v.w[h][0] = x.w[0];
                            v.w[h][1] = x.w[1];
                            v.w[h][2] = x.w[2];
                            v.w[h][3] = x.w[3] & 0x3fffffffu;
                        }
                        if (ok) {
                            break;
                        }
                    }
                }
                const uint32_t incl = __ballot_sync(0xffffffffu, flag == LB_INCL);
                if (stop == LB_WINDOW && incl != 0u) {
                    stop = r * 32 + (__ffs(incl) - 1);
                }
                if (r * 32 + lane <= stop) {
                    part.add(v);
                }
            }
#pragma unroll
            for (int sft = 16; sft > 0; sft >>= 1) {
                V o;
#pragma unroll
                for (int h = 0; h < HALVES; h++) {
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        o.w[h][q] = __shfl_xor_sync(0xffffffffu, part.w[h][q], sft);
                    }
                }
                part.add(o);
            }
            prefix.add(part);
            if (stop < LB_WINDOW) {
                break;
            }
            base -= LB_WINDOW;
        }
    }
    V inclusive = prefix;
    inclusive.add(aggregate);
#pragma unroll
    for (int h = 0; h < HALVES; h++) {
        if (lane == h) {
            LbWord me;
            me.w[0] = inclusive.w[h][0];
            me.w[1] = inclusive.w[h][1];
            me.w[2] = inclusive.w[h][2];
            me.w[3] = inclusive.w[h][3] | (LB_INCL << 30);
            lb_store(&state[(uint64_t) tile * HALVES + h], me);
        }
    }
    return prefix;
}

// Called by every thread of the block with the block's aggregate; warp 0 looks back and
// the exclusive prefix is handed to every thread through `bcast` (one LbVal in shared
// memory).
template <int HALVES, bool WIDE64>
MSPB_DEVFN LbVal<HALVES, WIDE64> lookback_prefix(LbWord *state, uint32_t tile, const LbVal<HALVES, WIDE64> &aggregate,
    LbVal<HALVES, WIDE64> *bcast)
{
    typedef LbVal<HALVES, WIDE64> V;
    if (threadIdx.x < 32) {
        const V prefix = lookback_warp<HALVES, WIDE64>(state, tile, aggregate);
        if (threadIdx.x == 0) {
            *bcast = prefix;
        }
    }
    __syncthreads();
    const V r = *bcast;
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
