// msprime_b200: device scan and LSD radix sort primitives (sm_100a).
//
// The sort replaces the reference's AVL site map (lib/avl.c, lib/mutgen.c:858-883,
// 1133) and per-site qsort (lib/mutgen.c:62-95): all mutation records are ordered
// once by a 64-bit key (position, coarse time) with a 32-bit payload (record id).
//
// Layout: keys and payloads are separate SoA arrays, ping-ponged between passes.
// Each pass handles one 8-bit digit with three launches: per-tile digit histogram,
// one exclusive scan of the (digit-major) histogram matrix, and a stable scatter
// that ranks keys with warp match-any.  Passes whose digit is constant over all
// keys are skipped (the varying-bit mask comes from one OR/AND reduction).
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
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;
constexpr int RS_BINS = 256;

__global__ void __launch_bounds__(256) rs_key_bits_kernel(const uint64_t *__restrict__ keys, uint64_t n, uint64_t *or_and)
{
    uint64_t o = 0, a = ~0ull;
    for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
        const uint64_t k = keys[i];
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

// tile_hist is digit-major: tile_hist[d * num_tiles + tile]
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const uint64_t *__restrict__ keys, uint64_t n, int shift,
    uint32_t num_tiles, uint32_t *__restrict__ tile_hist)
{
    __shared__ uint32_t hist[RS_BINS];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const uint64_t base = (uint64_t) blockIdx.x * RS_TILE;
#pragma unroll
    for (int k = 0; k < RS_ITEMS; k++) {
        const uint64_t i = base + (uint64_t) k * RS_THREADS + threadIdx.x;
        const bool valid = i < n;
        const uint32_t d = valid ? (uint32_t) ((keys[i] >> shift) & 0xffu) : 0xffffffffu;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        if (valid && (peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0) {
            atomicAdd(&hist[d], (uint32_t) __popc(peers));
        }
    }
    __syncthreads();
    tile_hist[(uint64_t) threadIdx.x * num_tiles + blockIdx.x] = hist[threadIdx.x];
}

// Stable scatter.  Warp w of a tile owns the contiguous chunk [w*512, (w+1)*512) of
// the tile; iteration k covers 32 consecutive keys, so element order is (warp,
// iteration, lane) and ranks are accumulated in that order.
__global__ void __launch_bounds__(RS_THREADS) rs_scatter_kernel(const uint64_t *__restrict__ keys_in,
    const uint32_t *__restrict__ vals_in, uint64_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out, uint64_t n,
    int shift, uint32_t num_tiles, const uint32_t *__restrict__ tile_base)
{
    __shared__ uint32_t cnt[RS_WARPS][RS_BINS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int w = 0; w < RS_WARPS; w++) {
        cnt[w][threadIdx.x] = 0;
    }
    __syncthreads();
    const uint64_t base = (uint64_t) blockIdx.x * RS_TILE + (uint64_t) warp * (32 * RS_ITEMS);
    uint64_t key[RS_ITEMS];
    uint32_t rank[RS_ITEMS];
#pragma unroll
    for (int k = 0; k < RS_ITEMS; k++) {
        const uint64_t i = base + (uint64_t) k * 32 + lane;
        key[k] = i < n ? keys_in[i] : ~0ull;
    }
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int k = 0; k < RS_ITEMS; k++) {
        const uint64_t i = base + (uint64_t) k * 32 + lane;
        const bool valid = i < n;
        const uint32_t d = valid ? (uint32_t) ((key[k] >> shift) & 0xffu) : 0xffffffffu;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const uint32_t before = __popc(peers & lt);
        uint32_t prev = 0;
        if (valid) {
            prev = cnt[warp][d];
        }
        __syncwarp();
        if (valid && before == 0) {
            cnt[warp][d] = prev + (uint32_t) __popc(peers);
        }
        __syncwarp();
        rank[k] = prev + before;
    }
    __syncthreads();
    {
        // thread t owns digit t: turn per-warp counts into global offsets
        uint32_t run = tile_base[(uint64_t) threadIdx.x * num_tiles + blockIdx.x];
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) {
            const uint32_t c = cnt[w][threadIdx.x];
            cnt[w][threadIdx.x] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < RS_ITEMS; k++) {
        const uint64_t i = base + (uint64_t) k * 32 + lane;
        if (i < n) {
            const uint32_t d = (uint32_t) ((key[k] >> shift) & 0xffu);
            const uint32_t dst = cnt[warp][d] + rank[k];
            keys_out[dst] = key[k];
            vals_out[dst] = vals_in[i];
        }
    }
}

struct RadixSortBuffers {
    uint64_t *keys[2];
    uint32_t *vals[2];
    uint32_t *tile_hist; // RS_BINS * num_tiles
    uint32_t *scan_tiles; // scan_num_tiles(RS_BINS * num_tiles)
    uint64_t *or_and; // 2 (device)
    uint64_t *or_and_host; // 2 (pinned)
};

inline uint32_t rs_num_tiles(uint64_t n)
{
    return (uint32_t) ((n + RS_TILE - 1) / RS_TILE);
}

// Sorts n (key, val) pairs held in buf.keys[0] / buf.vals[0]; returns the index
// (0 or 1) of the buffer holding the result.  *launches is incremented by the
// number of kernels launched.  Synchronises the stream once (to read the bit mask).
inline int device_radix_sort(RadixSortBuffers &buf, uint64_t n, cudaStream_t st, LaunchLog &log, int *passes_out)
{
    if (passes_out != nullptr) {
        *passes_out = 0;
    }
    if (n <= 1) {
        return 0;
    }
    const uint64_t init[2] = { 0ull, ~0ull };
    cudaMemcpyAsync(buf.or_and, init, sizeof(init), cudaMemcpyHostToDevice, st);
    MSPB_LAUNCH(log, "rs_key_bits", rs_key_bits_kernel<<<148 * 4, 256, 0, st>>>(buf.keys[0], n, buf.or_and));
    cudaMemcpyAsync(buf.or_and_host, buf.or_and, 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, st);
    cudaStreamSynchronize(st);
    const uint64_t varying = buf.or_and_host[0] & ~buf.or_and_host[1];
    const uint32_t tiles = rs_num_tiles(n);
    int cur = 0;
    for (int pass = 0; pass < 8; pass++) {
        const int shift = pass * 8;
        if (((varying >> shift) & 0xffull) == 0) {
            continue;
        }
        MSPB_LAUNCH(log, "rs_hist",
            rs_hist_kernel<<<tiles, RS_THREADS, 0, st>>>(buf.keys[cur], n, shift, tiles, buf.tile_hist));
        const uint32_t *th = buf.tile_hist;
        device_exclusive_scan<uint32_t>(
            log, (uint64_t) RS_BINS * tiles, [th] __device__(uint64_t i) { return th[i]; }, buf.tile_hist,
            buf.scan_tiles, (uint32_t *) nullptr, st);
        MSPB_LAUNCH(log, "rs_scatter",
            rs_scatter_kernel<<<tiles, RS_THREADS, 0, st>>>(
                buf.keys[cur], buf.vals[cur], buf.keys[cur ^ 1], buf.vals[cur ^ 1], n, shift, tiles, buf.tile_hist));
        cur ^= 1;
        if (passes_out != nullptr) {
            (*passes_out)++;
        }
    }
    return cur;
}

} // namespace mspb
