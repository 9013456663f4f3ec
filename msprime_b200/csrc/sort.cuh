// msprime_b200: device scan and LSD radix sort primitives (sm_100a).
//
// The sort replaces the reference's AVL site map (lib/avl.c, lib/mutgen.c:858-883,
// 1133) and per-site qsort (lib/mutgen.c:62-95): all mutation records are ordered
// once by a 64-bit key (position, coarse time) with a 32-bit payload (record id).
//
// Layout: keys and payloads are separate SoA arrays, ping-ponged between passes.
// One launch builds the digit histograms of every pass; then each pass is ONE
// kernel ("onesweep"): a tile ranks its keys with warp match-any, publishes its
// per-digit counts, obtains its global offsets by decoupled look-back over the
// preceding tiles, stages the tile in shared memory in sorted order and writes
// contiguous runs.  Digit widths are chosen per call (<= 8 bits) so that only the
// bits that vary are sorted.
#pragma once
#include "common.cuh"

namespace mspb {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

// ---------------------------------------------------------------- block helpers
template <typename T>
MSPB_DEVFN T warp_inclusive_scan(T v)
{
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const T n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) {
            v += n;
        }
    }
    return v;
}

// exclusive scan of one value per thread across a 256-thread block; returns the
// exclusive prefix and writes the block total to *total
template <typename T>
MSPB_DEVFN T block_exclusive_scan(T v, T *smem /* >= 9 entries */, T *total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const T inc = warp_inclusive_scan(v);
    if (lane == 31) {
        smem[warp] = inc;
    }
    __syncthreads();
    if (warp == 0) {
        T w = lane < (SCAN_THREADS / 32) ? smem[lane] : T(0);
        const T winc = warp_inclusive_scan(w);
        if (lane < (SCAN_THREADS / 32)) {
            smem[lane] = winc - w;
        }
        if (lane == (SCAN_THREADS / 32) - 1) {
            smem[8] = winc;
        }
    }
    __syncthreads();
    const T base = smem[warp];
    *total = smem[8];
    __syncthreads();
    return base + inc - v;
}

// ---------------------------------------------------------------- exclusive scan
// out[i] = sum_{j<i} f(j); *total = sum of all.  Reduce-then-scan: tile sums, one
// spine block, then a rescan of each tile with its base.
template <typename T, typename F>
__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_sums_kernel(uint64_t n, F f, T *tile_sums)
{
    __shared__ T sm[9];
    const uint64_t base = (uint64_t) blockIdx.x * SCAN_TILE;
    T acc = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const uint64_t i = base + (uint64_t) k * SCAN_THREADS + threadIdx.x;
        if (i < n) {
            acc += f(i);
        }
    }
    T total;
    block_exclusive_scan(acc, sm, &total);
    if (threadIdx.x == 0) {
        tile_sums[blockIdx.x] = total;
    }
}

template <typename T>
__global__ void __launch_bounds__(SCAN_THREADS) scan_spine_kernel(uint32_t num_tiles, T *tile_sums, T *total_out)
{
    __shared__ T sm[9];
    T carry = 0;
    for (uint32_t start = 0; start < num_tiles; start += SCAN_THREADS) {
        const uint32_t i = start + threadIdx.x;
        const T v = i < num_tiles ? tile_sums[i] : T(0);
        T total;
        const T ex = block_exclusive_scan(v, sm, &total);
        if (i < num_tiles) {
            tile_sums[i] = carry + ex;
        }
        carry += total;
    }
    if (threadIdx.x == 0 && total_out != nullptr) {
        *total_out = carry;
    }
}

template <typename T, typename F>
__global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(uint64_t n, F f, const T *tile_sums, T *out)
{
    __shared__ T sm[9];
    // blocked arrangement: thread t owns SCAN_ITEMS consecutive elements
    const uint64_t base = (uint64_t) blockIdx.x * SCAN_TILE + (uint64_t) threadIdx.x * SCAN_ITEMS;
    T v[SCAN_ITEMS];
    T acc = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const uint64_t i = base + k;
        v[k] = i < n ? f(i) : T(0);
        acc += v[k];
    }
    T total;
    T run = tile_sums[blockIdx.x] + block_exclusive_scan(acc, sm, &total);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const uint64_t i = base + k;
        if (i < n) {
            out[i] = run;
        }
        run += v[k];
    }
}

// tile_sums must hold ceil(n / SCAN_TILE) entries.  Launches 3 kernels (0 if n == 0,
// in which case *total_out is set by a memset).
template <typename T, typename F>
inline void device_exclusive_scan(LaunchLog &log, uint64_t n, F f, T *out, T *tile_sums, T *total_out, cudaStream_t st)
{
    if (n == 0) {
        if (total_out != nullptr) {
            cudaMemsetAsync(total_out, 0, sizeof(T), st);
        }
        return;
    }
    const uint32_t tiles = (uint32_t) ((n + SCAN_TILE - 1) / SCAN_TILE);
    MSPB_LAUNCH(log, "scan_tile_sums", scan_tile_sums_kernel<T, F><<<tiles, SCAN_THREADS, 0, st>>>(n, f, tile_sums));
    MSPB_LAUNCH(log, "scan_spine", scan_spine_kernel<T><<<1, SCAN_THREADS, 0, st>>>(tiles, tile_sums, total_out));
    MSPB_LAUNCH(log, "scan_apply", scan_apply_kernel<T, F><<<tiles, SCAN_THREADS, 0, st>>>(n, f, tile_sums, out));
}

inline uint64_t scan_num_tiles(uint64_t n)
{
    return (n + SCAN_TILE - 1) / SCAN_TILE;
}

// ---------------------------------------------------------------- radix sort
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_DIGIT_BITS = 8;
constexpr int RS_BINS = 1 << RS_DIGIT_BITS;
constexpr int RS_MAX_PASSES = 8;
// look-back word: 2 flag bits + 30-bit count
constexpr uint32_t RS_FLAG_AGG = 1u << 30, RS_FLAG_INCL = 2u << 30, RS_COUNT_MASK = (1u << 30) - 1u;
constexpr uint64_t RS_MAX_N = (1ull << 30) - 1ull;

struct RadixPlan {
    int passes;
    int shift[RS_MAX_PASSES];
    int bits[RS_MAX_PASSES];
};

// Digits covering the set bits of `varying` (bit i set = bit i differs between keys),
// at most 8 bits wide, as few passes as possible and as even as possible.
inline RadixPlan radix_plan(uint64_t varying)
{
    RadixPlan p;
    p.passes = 0;
    int pos = 0;
    while (pos < 64) {
        while (pos < 64 && !((varying >> pos) & 1ull)) {
            pos++;
        }
        if (pos >= 64) {
            break;
        }
        // contiguous stretch [pos, end) up to the highest varying bit
        int end = 64;
        while (end > pos && !((varying >> (end - 1)) & 1ull)) {
            end--;
        }
        // find a run of constant bits >= 8 wide to split on (skipping it saves a pass)
        int gap_start = -1;
        for (int i = pos, run = 0; i < end; i++) {
            if (!((varying >> i) & 1ull)) {
                run++;
                if (run == 8) {
                    gap_start = i - 7;
                    break;
                }
            } else {
                run = 0;
            }
        }
        const int stretch_end = gap_start >= 0 ? gap_start : end;
        const int width = stretch_end - pos;
        const int np = (width + 7) / 8;
        for (int k = 0; k < np; k++) {
            const int lo = pos + (int) ((long long) width * k / np);
            const int hi = pos + (int) ((long long) width * (k + 1) / np);
            p.shift[p.passes] = lo;
            p.bits[p.passes] = hi - lo;
            p.passes++;
        }
        pos = stretch_end;
    }
    return p;
}

struct RadixPlanDev {
    int passes;
    int shift[RS_MAX_PASSES];
    uint32_t mask[RS_MAX_PASSES];
};

template <typename KeyT>
MSPB_DEVFN uint32_t rs_digit(KeyT k, int shift, uint32_t mask)
{
    return (uint32_t) (k >> shift) & mask;
}

// OR / AND of all keys (which bits vary)
template <typename KeyT>
__global__ void __launch_bounds__(256) rs_key_bits_kernel(const KeyT *__restrict__ keys, uint64_t n, uint64_t *or_and)
{
    uint64_t o = 0, a = ~0ull;
    for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
        const uint64_t k = (uint64_t) keys[i];
        o |= k;
        a &= k;
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        o |= __shfl_xor_sync(0xffffffffu, o, s);
        a &= __shfl_xor_sync(0xffffffffu, a, s);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicOr((unsigned long long *) &or_and[0], (unsigned long long) o);
        atomicAnd((unsigned long long *) &or_and[1], (unsigned long long) a);
    }
}

// Global digit histograms of every pass in one read of the keys, hist[pass * RS_BINS +
// digit]; the last block to finish turns each pass's histogram into its exclusive prefix
// (where each digit's run starts in the output).  hist[RS_MAX_PASSES * RS_BINS] is the
// ticket counter.
template <typename KeyT>
__global__ void __launch_bounds__(RS_THREADS) rs_global_hist_kernel(const KeyT *__restrict__ keys, uint64_t n,
    RadixPlanDev plan, uint32_t *__restrict__ hist)
{
    __shared__ uint32_t sh[RS_MAX_PASSES * RS_BINS];
    __shared__ uint32_t scan_tmp[9];
    __shared__ bool is_last;
    for (int i = threadIdx.x; i < plan.passes * RS_BINS; i += RS_THREADS) {
        sh[i] = 0;
    }
    __syncthreads();
    // 16-byte loads, two in flight per thread; the tail (n not a multiple of PER) goes key by key
    constexpr int PER = 16 / (int) sizeof(KeyT);
    const uint64_t nvec = n / PER;
    const uint4 *vec = reinterpret_cast<const uint4 *>(keys);
    const uint64_t stride = (uint64_t) gridDim.x * RS_THREADS;
    for (uint64_t v = (uint64_t) blockIdx.x * RS_THREADS + threadIdx.x; v < nvec; v += 2 * stride) {
        const bool two = v + stride < nvec;
        const uint4 a = vec[v];
        const uint4 b = two ? vec[v + stride] : make_uint4(0u, 0u, 0u, 0u);
        KeyT k[2 * PER];
        if (sizeof(KeyT) == 4) {
            k[0] = (KeyT) a.x, k[1] = (KeyT) a.y, k[2] = (KeyT) a.z, k[3] = (KeyT) a.w;
            k[4] = (KeyT) b.x, k[5] = (KeyT) b.y, k[6] = (KeyT) b.z, k[7] = (KeyT) b.w;
        } else {
            k[0] = (KeyT) ((uint64_t) a.x | ((uint64_t) a.y << 32));
            k[1] = (KeyT) ((uint64_t) a.z | ((uint64_t) a.w << 32));
            k[2] = (KeyT) ((uint64_t) b.x | ((uint64_t) b.y << 32));
            k[3] = (KeyT) ((uint64_t) b.z | ((uint64_t) b.w << 32));
        }
        for (int p = 0; p < plan.passes; p++) {
            const int sft = plan.shift[p];
            const uint32_t msk = plan.mask[p];
            uint32_t *h = sh + p * RS_BINS;
#pragma unroll
            for (int q = 0; q < PER; q++) {
                atomicAdd(&h[rs_digit(k[q], sft, msk)], 1u);
            }
            if (two) {
#pragma unroll
                for (int q = PER; q < 2 * PER; q++) {
                    atomicAdd(&h[rs_digit(k[q], sft, msk)], 1u);
                }
            }
        }
    }
    if (blockIdx.x == 0) {
        for (uint64_t i = nvec * PER + threadIdx.x; i < n; i += RS_THREADS) {
            const KeyT k = keys[i];
            for (int p = 0; p < plan.passes; p++) {
                atomicAdd(&sh[p * RS_BINS + rs_digit(k, plan.shift[p], plan.mask[p])], 1u);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < plan.passes * RS_BINS; i += RS_THREADS) {
        const uint32_t c = sh[i];
        if (c != 0) {
            atomicAdd(&hist[i], c);
        }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        is_last = atomicAdd(&hist[RS_MAX_PASSES * RS_BINS], 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        for (int p = 0; p < plan.passes; p++) {
            volatile uint32_t *h = hist + p * RS_BINS;
            uint32_t tot;
            const uint32_t ex = block_exclusive_scan<uint32_t>(h[threadIdx.x], scan_tmp, &tot);
            h[threadIdx.x] = ex;
        }
    }
}

__global__ void __launch_bounds__(256) rs_iota_kernel(uint32_t *vals, uint64_t n)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        vals[i] = (uint32_t) i;
    }
}

// Keys per thread and per tile: 16 x 256 for 32-bit keys, 8 x 256 for 64-bit keys (so that
// four blocks fit one SM's shared memory either way).
template <typename KeyT>
struct RsTraits {
    static constexpr int ITEMS = sizeof(KeyT) == 4 ? 16 : 8;
    static constexpr int TILE = RS_THREADS * ITEMS;
    static constexpr size_t SMEM = (size_t) RS_WARPS * RS_BINS * 4 + 2 * RS_BINS * 4 + 16 * 4
                                   + (size_t) TILE * (sizeof(KeyT) + 4);
};

// One tile of one pass.  Warp w owns the contiguous chunk [w*32*ITEMS, (w+1)*32*ITEMS) of
// the tile and iteration k covers 32 consecutive keys, so (warp, iteration, lane) order is
// index order and ranks accumulated in that order make the pass stable.  Keys and
// payloads are staged in shared memory in sorted order and written as contiguous runs.
// vals_in == nullptr means the identity payload (first pass).  FULL tiles carry no bounds
// checks; NBITS >= the digit width (one ballot per bit).
template <typename KeyT, int NBITS, bool FULL>
MSPB_DEVFN void rs_onesweep_tile(const KeyT *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
    KeyT *__restrict__ keys_out, uint32_t *__restrict__ vals_out, uint64_t n, int shift, uint32_t mask,
    const uint32_t *__restrict__ global_excl, volatile uint32_t *tile_state, unsigned char *rs_smem)
{
    constexpr int ITEMS = RsTraits<KeyT>::ITEMS;
    constexpr int TILE = RsTraits<KeyT>::TILE;
    KeyT *skeys = reinterpret_cast<KeyT *>(rs_smem);
    uint32_t *svals = reinterpret_cast<uint32_t *>(skeys + TILE);
    uint32_t(*cnt)[RS_BINS] = reinterpret_cast<uint32_t(*)[RS_BINS]>(svals + TILE);
    uint32_t *local_start = reinterpret_cast<uint32_t *>(cnt + RS_WARPS);
    uint32_t *global_base = local_start + RS_BINS;
    uint32_t *scan_tmp = global_base + RS_BINS;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tile = blockIdx.x;
#pragma unroll
    for (int w = 0; w < RS_WARPS; w++) {
        cnt[w][threadIdx.x] = 0;
    }
    __syncthreads();
    const uint64_t tile_base = (uint64_t) tile * TILE;
    const uint32_t tile_n = FULL ? (uint32_t) TILE : (uint32_t) (n - tile_base);
    const uint32_t wbase = (uint32_t) warp * (32 * ITEMS) + lane; // this lane's first key inside the tile
    const KeyT *kin = keys_in + tile_base + wbase;
    KeyT key[ITEMS];
    uint32_t pos2[ITEMS / 2]; // ranks inside the tile (< 4096), two per register
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        key[k] = (FULL || wbase + k * 32 < tile_n) ? kin[k * 32] : (KeyT) 0;
    }
#pragma unroll
    for (int k = 0; k < ITEMS / 2; k++) {
        pos2[k] = 0;
    }
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        const bool valid = FULL || wbase + k * 32 < tile_n;
        const uint32_t d = rs_digit(key[k], shift, mask);
        // lanes holding the same digit, from one ballot per digit bit (MATCH.ANY runs one
        // round per distinct value here and is slower)
        uint32_t peers = FULL ? 0xffffffffu : __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int b = 0; b < NBITS; b++) {
            const bool bit = (d & (1u << b)) != 0u;
            const uint32_t bal = __ballot_sync(0xffffffffu, bit);
            peers &= bal ^ (bit ? 0u : 0xffffffffu);
        }
        // the lowest lane of each group bumps the warp's digit counter for the whole group
        const int leader = __ffs(FULL ? peers : (peers | (valid ? 0u : (1u << lane)))) - 1;
        uint32_t prev = 0;
        if (valid && lane == leader) {
            prev = atomicAdd(&cnt[warp][d], (uint32_t) __popc(peers));
        }
        prev = __shfl_sync(0xffffffffu, prev, leader);
        pos2[k >> 1] |= (prev + (uint32_t) __popc(peers & lt)) << (16 * (k & 1));
    }
    // payloads: issued here so that their latency hides behind the scans below
    uint32_t val[ITEMS];
    if (vals_in != nullptr) {
        const uint32_t *vin = vals_in + tile_base + wbase;
#pragma unroll
        for (int k = 0; k < ITEMS; k++) {
            val[k] = (FULL || wbase + k * 32 < tile_n) ? vin[k * 32] : 0u;
        }
    } else {
#pragma unroll
        for (int k = 0; k < ITEMS; k++) {
            val[k] = (uint32_t) tile_base + wbase + k * 32;
        }
    }
    __syncthreads();
    // thread b owns digit b
    uint32_t tile_count = 0;
    {
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) {
            const uint32_t c = cnt[w][threadIdx.x];
            cnt[w][threadIdx.x] = tile_count;
            tile_count += c;
        }
    }
    // publish, then look back for the number of keys with this digit in earlier tiles
    volatile uint32_t *mine = tile_state + (uint64_t) tile * RS_BINS + threadIdx.x;
    uint32_t prefix = 0;
    if (tile == 0) {
        *mine = RS_FLAG_INCL | tile_count;
    } else {
        *mine = RS_FLAG_AGG | tile_count;
    }
    // exclusive scan over the 256 digits within the tile (the global one is precomputed)
    uint32_t tot;
    const uint32_t ls = block_exclusive_scan<uint32_t>(tile_count, scan_tmp, &tot);
    const uint32_t gs = global_excl[threadIdx.x];
    local_start[threadIdx.x] = ls;
    __syncthreads();
    // stage keys and payloads in sorted order (overlaps the look-back latency of other warps)
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        if (FULL || wbase + k * 32 < tile_n) {
            const uint32_t d = rs_digit(key[k], shift, mask);
            const uint32_t q = ((pos2[k >> 1] >> (16 * (k & 1))) & 0xffffu) + local_start[d] + cnt[warp][d];
            skeys[q] = key[k];
            svals[q] = val[k];
        }
    }
    if (tile > 0) {
        // this digit's counts in earlier tiles, nearest first, until one carries its inclusive
        // prefix.  RS_LOOK predecessors are read at once (independent loads, one L2 round trip)
        // and consumed in order up to the first that is not published yet.
        constexpr int RS_LOOK = 4;
        int64_t p = (int64_t) tile - 1; // nearest predecessor not consumed yet
        for (bool done = false; !done;) {
            uint32_t v[RS_LOOK];
#pragma unroll
            for (int w = 0; w < RS_LOOK; w++) {
                // before the first tile: an inclusive prefix of zero
                v[w] = p - w >= 0 ? (uint32_t) tile_state[(uint64_t) (p - w) * RS_BINS + threadIdx.x] : (2u << 30);
            }
            int used = 0;
#pragma unroll
            for (int w = 0; w < RS_LOOK; w++) {
                if (!done && used == w && (v[w] & ~RS_COUNT_MASK) != 0u) {
                    prefix += v[w] & RS_COUNT_MASK;
                    used = w + 1;
                    done = (v[w] & RS_FLAG_INCL) != 0u;
                }
            }
            p -= used;
        }
        *mine = RS_FLAG_INCL | (prefix + tile_count);
    }
    global_base[threadIdx.x] = gs + prefix - ls; // dst = global_base[d] + local position
    __syncthreads();
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        const uint32_t j = (uint32_t) k * RS_THREADS + threadIdx.x;
        if (FULL || j < tile_n) {
            const KeyT kk = skeys[j];
            const uint32_t dst = global_base[rs_digit(kk, shift, mask)] + j;
            keys_out[dst] = kk;
            vals_out[dst] = svals[j];
        }
    }
}

template <typename KeyT, int NBITS>
__global__ void __launch_bounds__(RS_THREADS, 4) rs_onesweep_kernel(const KeyT *__restrict__ keys_in,
    const uint32_t *__restrict__ vals_in, KeyT *__restrict__ keys_out, uint32_t *__restrict__ vals_out, uint64_t n,
    int shift, uint32_t mask, const uint32_t *__restrict__ global_excl, volatile uint32_t *tile_state)
{
    extern __shared__ __align__(16) unsigned char rs_smem[];
    if ((uint64_t) (blockIdx.x + 1) * RsTraits<KeyT>::TILE <= n) {
        rs_onesweep_tile<KeyT, NBITS, true>(keys_in, vals_in, keys_out, vals_out, n, shift, mask, global_excl,
            tile_state, rs_smem);
    } else {
        rs_onesweep_tile<KeyT, NBITS, false>(keys_in, vals_in, keys_out, vals_out, n, shift, mask, global_excl,
            tile_state, rs_smem);
    }
}

template <typename KeyT>
struct RadixSortBuffers {
    KeyT *keys[2];
    uint32_t *vals[2];
    uint32_t *hist; // RS_MAX_PASSES * RS_BINS
    uint32_t *tile_state; // RS_MAX_PASSES * num_tiles * RS_BINS
    uint64_t *or_and; // 2 (device)
    uint64_t *or_and_host; // 2 (pinned)
    uint64_t *or_and_host_dev; // the same two words as the device addresses them (nullptr: copy with the DMA engine)
};

// A few words from device memory into pinned host memory WITHOUT the copy engine: a cudaMemcpyAsync of 16 bytes
// queues behind whatever that engine is already moving for other streams of the process (the 100 MB column copies
// of another context's fetch: 13 ms measured), a store from a kernel does not.
__global__ void rs_words_to_host_kernel(const uint32_t *__restrict__ src, uint32_t *dst_host, uint32_t words)
{
    for (uint32_t i = threadIdx.x; i < words; i += blockDim.x) {
        dst_host[i] = src[i];
    }
    __threadfence_system();
}

template <typename KeyT>
inline uint32_t rs_num_tiles(uint64_t n)
{
    return (uint32_t) ((n + RsTraits<KeyT>::TILE - 1) / RsTraits<KeyT>::TILE);
}

// words of look-back state for any key width (64-bit keys have the smaller tiles)
inline uint64_t rs_state_words(uint64_t n)
{
    return (uint64_t) RS_MAX_PASSES * rs_num_tiles<uint64_t>(n > 0 ? n : 1) * RS_BINS;
}

// Sorts n (key, val) pairs held in buf.keys[0] (payload: buf.vals[0], or the identity
// 0..n-1 when identity_vals); returns the index (0 or 1) of the buffers holding the
// result.  `varying` names the key bits that may differ; pass 0 to have them measured
// (one reduction kernel and one stream synchronisation).  n must be <= RS_MAX_N.
template <typename KeyT>
inline int device_radix_sort(RadixSortBuffers<KeyT> &buf, uint64_t n, uint64_t varying, bool identity_vals,
    cudaStream_t st, LaunchLog &log, int *passes_out)
{
    if (passes_out != nullptr) {
        *passes_out = 0;
    }
    if (n <= 1) {
        if (n == 1 && identity_vals) {
            cudaMemsetAsync(buf.vals[0], 0, sizeof(uint32_t), st);
        }
        return 0;
    }
    if (varying == 0) {
        cudaMemsetAsync(buf.or_and, 0, sizeof(uint64_t), st); // OR of all keys
        cudaMemsetAsync(buf.or_and + 1, 0xff, sizeof(uint64_t), st); // AND of all keys
        MSPB_LAUNCH(log, "rs_key_bits", rs_key_bits_kernel<KeyT><<<148 * 4, 256, 0, st>>>(buf.keys[0], n, buf.or_and));
        if (buf.or_and_host_dev != nullptr) {
            rs_words_to_host_kernel<<<1, 32, 0, st>>>(reinterpret_cast<const uint32_t *>(buf.or_and),
                reinterpret_cast<uint32_t *>(buf.or_and_host_dev), 4u);
        } else {
            cudaMemcpyAsync(buf.or_and_host, buf.or_and, 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, st);
        }
        cudaStreamSynchronize(st);
        varying = buf.or_and_host[0] & ~buf.or_and_host[1];
    }
    const RadixPlan plan = radix_plan(varying);
    const uint32_t tiles = rs_num_tiles<KeyT>(n);
    if (plan.passes == 0) {
        // all keys equal: the input order is the sorted order
        if (identity_vals) {
            MSPB_LAUNCH(log, "rs_iota", rs_iota_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, st>>>(buf.vals[0], n));
        }
        return 0;
    }
    RadixPlanDev pd;
    pd.passes = plan.passes;
    for (int p = 0; p < plan.passes; p++) {
        pd.shift[p] = plan.shift[p];
        pd.mask[p] = (1u << plan.bits[p]) - 1u;
    }
    cudaMemsetAsync(buf.hist, 0, ((size_t) RS_MAX_PASSES * RS_BINS + 1) * sizeof(uint32_t), st);
    cudaMemsetAsync(buf.tile_state, 0, (size_t) plan.passes * tiles * RS_BINS * sizeof(uint32_t), st);
    const unsigned hist_blocks = tiles < 148u * 8u ? tiles : 148u * 8u;
    MSPB_LAUNCH(log, "rs_global_hist",
        rs_global_hist_kernel<KeyT><<<hist_blocks, RS_THREADS, 0, st>>>(buf.keys[0], n, pd, buf.hist));
    int cur = 0;
    static bool smem_opt_in = false; // per key type (this is a template)
    if (!smem_opt_in) {
        const int bytes = (int) RsTraits<KeyT>::SMEM;
        cudaFuncSetAttribute(rs_onesweep_kernel<KeyT, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        cudaFuncSetAttribute(rs_onesweep_kernel<KeyT, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        cudaFuncSetAttribute(rs_onesweep_kernel<KeyT, 7>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        cudaFuncSetAttribute(rs_onesweep_kernel<KeyT, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        smem_opt_in = true;
    }
    for (int p = 0; p < plan.passes; p++) {
        const uint32_t *vin = (p == 0 && identity_vals) ? nullptr : buf.vals[cur];
        uint32_t *state = buf.tile_state + (size_t) p * tiles * RS_BINS;
#define MSPB_RS_PASS(NB)                                                                                             \
    MSPB_LAUNCH(log, "rs_onesweep",                                                                                  \
        (rs_onesweep_kernel<KeyT, NB><<<tiles, RS_THREADS, RsTraits<KeyT>::SMEM, st>>>(buf.keys[cur], vin,           \
            buf.keys[cur ^ 1], buf.vals[cur ^ 1], n, pd.shift[p], pd.mask[p], buf.hist + p * RS_BINS, state)))
        // digits narrower than 5 bits run through the 5-ballot instantiation (extra bits are zero)
        switch (plan.bits[p]) {
        case 8:
            MSPB_RS_PASS(8);
            break;
        case 7:
            MSPB_RS_PASS(7);
            break;
        case 6:
            MSPB_RS_PASS(6);
            break;
        default:
            MSPB_RS_PASS(5);
            break;
        }
#undef MSPB_RS_PASS
        cur ^= 1;
    }
    if (passes_out != nullptr) {
        *passes_out = plan.passes;
    }
    return cur;
}

} // namespace mspb
