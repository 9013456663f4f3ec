// msprime_b200: C ABI and host orchestration (see include/msprime_b200.h).
//
// The host side only validates parameters, sizes buffers and launches kernels; all
// table work happens on the device.  There is no CPU fallback: without a usable GPU
// mspb_ctx_create fails with MSPB_ERR_NO_DEVICE.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <cfloat>
#include <map>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "../../include/msprime_b200.h"
#include "kernels.cuh"
#include "consumers.cuh"

using namespace mspb;

#define TSK_E(code) ((code) ^ (1 << 13))

namespace {

template <typename T>
struct DBuf {
    T *p = nullptr;
    size_t cap = 0;
    uint64_t gen = 0; // bumped whenever the buffer is (re)allocated, i.e. its contents are lost
    int ensure(size_t n)
    {
        if (n <= cap && p != nullptr) {
            return 0;
        }
        gen++;
        if (p != nullptr) {
            cudaFree(p);
            p = nullptr;
        }
        size_t want = n + n / 8 + 64;
        if (cudaMalloc((void **) &p, want * sizeof(T)) != cudaSuccess) {
            cap = 0;
            p = nullptr;
            cudaGetLastError();
            return MSPB_ERR_NO_MEMORY;
        }
        cap = want;
        return 0;
    }
    void release()
    {
        if (p != nullptr) {
            cudaFree(p);
        }
        p = nullptr;
        cap = 0;
    }
};

struct HostTotals {
    uint64_t k1_total; // new mutations (low half) | those inside the slice (high half)
    Counts3 p1; // sites, mutation rows, allele draws
    uint32_t collisions, dense;
    uint32_t k4_next, k4_pad; // K4's work counter: warps claim 32 sites of the work list at a time
    Bytes4 bytes; // ragged column sizes
    unsigned long long first_bad; // validation: (row << 16 | code index), min wins
    int err;
    uint32_t fractional_sites; // kept site positions that are not integers < 2^32
    uint32_t kslice[4]; // kept sites [0], [1]) and kept mutations [2], [3]) inside the genome slice
    uint32_t kept_unsorted; // kept site positions are not increasing (sliced calls need that)
    ReplayTotals replay;
};

} // namespace

struct mspb_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[MSPB_NUM_PHASES + 1] = {};
    // host copies of the rate map / matrix model whose device copies are current (upload_params)
    bool cache_map_valid = false, cache_map_discrete = false, cache_model_valid = false;
    std::vector<double> cache_map_pos, cache_map_rate, cache_model_root, cache_model_matrix;
    std::vector<uint64_t> cache_model_off;
    std::vector<char> cache_model_bytes;
    DBuf<ChainFree> chain_free; // K5's tile prefixes without a chain (emit.cuh)
    cudaEvent_t ev_p1 = nullptr, ev_cf = nullptr; // site rows are final / the prefix table is ready
    cudaEvent_t ev_async0 = nullptr, ev_async1 = nullptr; // around the stages of mspb_generate_begin_device
    bool tables_pending = false; // between mspb_ctx_load_tables_begin and _end
    cudaStream_t copy_stream = nullptr; // load_tables: the columns K0's sorts do not read arrive here
    cudaEvent_t ev_keys = nullptr, ev_copied = nullptr;
    cudaEvent_t ev_k1 = nullptr; // the K1 totals have reached the host (no timing: cheap to wait on)
    float phase_ms[MSPB_NUM_PHASES] = {};
    LaunchLog log;
    bool tables_loaded = false, generated = false;
    // tables
    double L = 0;
    uint64_t N = 0, E = 0, S0 = 0, M0 = 0;
    bool have_mut_time = false, kept_fractional = false;
    DBuf<double> node_time, edge_left, edge_right;
    DBuf<ChildIv> ci;
    DBuf<int32_t> edge_parent, edge_child;
    DBuf<uint32_t> child_start;
    DBuf<double> k_site_pos, k_mut_time;
    DBuf<uint64_t> k_site_anc_off, k_site_md_off, k_mut_der_off, k_mut_md_off;
    DBuf<char> k_site_anc, k_site_md, k_mut_der, k_mut_md;
    DBuf<int32_t> k_mut_site, k_mut_node, k_site_allele, k_mut_allele;
    // call parameters on the device
    DBuf<double> map_pos, map_rate, map_cum, map_inv, model_root, model_matrix, replay;
    DBuf<char> model_alleles;
    DBuf<uint64_t> model_allele_off;
    uint64_t kept_anc_bytes = 0, kept_smd_bytes = 0, kept_der_bytes = 0, kept_mmd_bytes = 0;
    // work
    DBuf<uint32_t> goff, loff; // K1: offsets of each edge's mutations inside its K1 tile
    DBuf<uint64_t> k1_tiles; // K1: tile bases (count | in-slice count << 32), total appended
    // edges overlapping the slice of the last clipped call (index over the loaded tables)
    DBuf<uint32_t> slice_edges, slice_scan_tiles;
    double slice_cached_left = 0, slice_cached_right = 0;
    uint64_t slice_num_edges = 0;
    bool slice_cache_valid = false;
    // device-side exchange of a sharded call: send[8] | recv[MAX_WORLD][8] | bases1[16] | bases2[16]
    DBuf<uint64_t> xchg;
    uint64_t *xchg_host = nullptr; // pinned mirror of send[8] + bases1[16] + bases2[16]
    DBuf<Rec> rec;
    DBuf<uint64_t> keys0, keys1; // 64-bit keys (ping-pong)
    DBuf<uint32_t> k32_0, k32_1; // 32-bit keys (ping-pong)
    DBuf<uint32_t> vals0, vals1, tile_hist, rs_scan_tiles;
    DBuf<uint64_t> or_and;
    DBuf<LbWord> lb_state; // look-back scan state, shared by K1, P1 and K5 (stream ordered)
    DBuf<uint32_t> site_scan, mut_scan, draw_scan, multi_list;
    DBuf<int32_t> parent, state, site_state;
    DBuf<uint32_t> rp_edge, rp_interval, rp_count, rp_out_off;
    DBuf<uint64_t> rp_stream_off;
    DBuf<uint64_t> replay_set; // replay harness, continuous genome: the positions taken so far (hash set)
    DBuf<char> dev_totals; // HostTotals mirror
    HostTotals *host = nullptr; // pinned
    uint64_t *or_and_host = nullptr; // pinned
    // the same pinned words as kernels address them: small read-backs are stored by a kernel, not copied by the
    // DMA engine, where they would queue behind the column copies of other contexts (sort.cuh)
    HostTotals *host_dev = nullptr;
    uint64_t *or_and_host_dev = nullptr;
    // outputs
    DBuf<double> o_site_pos, o_mut_time;
    DBuf<char> o_site_anc, o_site_md, o_mut_der, o_mut_md;
    DBuf<uint64_t> o_site_anc_off, o_site_md_off, o_mut_der_off, o_mut_md_off;
    DBuf<int32_t> o_mut_site, o_mut_node, o_mut_parent;
    mspb_counts_t counts = {};
    SortedView last_sv = {};
    // two-phase calls (multi-GPU): everything K5 needs, kept between _begin and _end
    struct {
        bool valid = false, defer = false;
        bool defer_resolve = false; // SLiM in a two-phase call: K4 needs the slice's id base, so it runs in _end
        uint32_t fixed_len = 0; // matrix model: the common length of all alleles, or 0
        int exchange_stage = 0; // 2: mspb_generate_end_device ran, mspb_generate_finish_device is due
        bool async = false; // mspb_generate_begin_device: the host has not seen P1's totals yet
        // ... and both exchanges run BESIDE K4, on the second stream, because the byte counts follow from
        // P1's totals in closed form (ChainFree): K5 then starts with every base known and nothing is shifted
        bool early = false;
        bool cf_ready = false; // the ChainFree table of this call is already enqueued (early exchange)
        bool key32 = false; // ... and whether a "dense" flag would mean a second sort (KEY32_POS)
        float begin_ms = 0; // device time of the _begin stage (async: read at _finish)
        SortedView sv;
        ResolveView rv;
        KeptView kept;
        ModelView model;
        CallView call;
        OutView out;
        uint64_t R = 0, S0 = 0, Mnew = 0, M = 0, replay_draws = 0, start_allele = 0, replay_len = 0;
        int model_kind = 0;
        bool replay = false;
    } pending;
    // leading bytes of the metadata offset columns known to be zero (and for which buffer)
    const void *md_zero_site_ptr = nullptr, *md_zero_mut_ptr = nullptr;
    uint64_t md_zero_site_bytes = 0, md_zero_mut_bytes = 0;
    uint64_t last_fetch_bytes = 0; // device-to-host bytes of the last mspb_fetch
    // K2: one block per 256 edges (dense) or one thread per mutation (sparse); chosen from the last call's density
    int k2_force = 0; // 0: by density; 1: dense; 2: sparse (mspb_ctx_force_placement, tests)
    uint64_t last_num_work = 0, last_new = 0;
    // sample-count step functions of the loaded tree sequence (consumers.cuh), built on request
    struct {
        bool valid = false;
        uint32_t num_samples = 0, num_x = 0, levels = 0;
        uint64_t steps = 0; // entries of fx / fv in use
        DBuf<uint32_t> is_sample, foff, fcnt, samples; // samples: the sample nodes in increasing order
        DBuf<uint64_t> xs;
        uint32_t *fx = nullptr, *fv = nullptr;
        size_t fcap = 0;
        double build_ms = 0;
    } sidx;
    DBuf<uint32_t> q_below, q_own; // per mutation row of the resident result
    DBuf<unsigned long long> q_afs;
    DBuf<uint32_t> q_row_begin, q_row_end, q_allele, q_walk_id; // variant decoding
    DBuf<double> q_tmax;
    DBuf<int32_t> q_geno;
};

static const char *const PHASE_NAMES[MSPB_NUM_PHASES] = { "params+kept", "K1 count+scan", "K2 place", "K3 sort",
    "P1 sites", "K4 resolve", "K5 emit", "dense re-sort" };

extern "C" const char *mspb_phase_name(int phase)
{
    return (phase >= 0 && phase < MSPB_NUM_PHASES) ? PHASE_NAMES[phase] : "";
}

extern "C" const char *mspb_version(void)
{
    return MSPB_VERSION;
}

extern "C" int mspb_cuda_runtime_version(void)
{
    int v = 0;
    if (cudaRuntimeGetVersion(&v) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return v;
}

extern "C" const char *mspb_strerror(int err)
{
    switch (err) {
        case 0:
            return "ok";
        case MSPB_ERR_GENERIC:
            return "Generic error; please file a bug report";
        case MSPB_ERR_NO_MEMORY:
            return "Out of memory";
        case MSPB_ERR_BAD_PARAM_VALUE:
            return "Bad parameter value provided";
        case MSPB_ERR_DUPLICATE_SITE_POSITION:
            return "Duplicate site positions.";
        case MSPB_ERR_BAD_RATE_VALUE:
            return "Rates must be non-negative and finite";
        case MSPB_ERR_INCOMPATIBLE_MUTATION_MAP_LENGTH:
            return "Incompatible mutation map: sequence length differs from that of the tree sequence.";
        case MSPB_ERR_INSUFFICIENT_INTERVALS:
            return "At least one interval must be specified.";
        case MSPB_ERR_INTERVAL_MAP_START_NON_ZERO:
            return "The first interval must start with zero";
        case MSPB_ERR_NONFINITE_INTERVAL_POSITION:
            return "Interval positions must be finite.";
        case MSPB_ERR_INTERVAL_POSITIONS_UNSORTED:
            return "Interval positions must be listed in increasing order";
        case MSPB_ERR_UNKNOWN_ALLELE:
            return "Existing allele(s) incompatible with mutation model alphabet.";
        case MSPB_ERR_INSUFFICIENT_ALLELES:
            return "Must have at least two alleles.";
        case MSPB_ERR_BAD_ROOT_PROBABILITIES:
            return "Root probabilities must be nonnegative and sum to one.";
        case MSPB_ERR_BAD_TRANSITION_MATRIX:
            return "Each row of the transition matrix must be nonnegative and sum to one.";
        case MSPB_ERR_BAD_SLIM_PARAMETERS:
            return "SLiM mutation IDs and mutation type IDs must be nonnegative.";
        case MSPB_ERR_MUTATION_ID_OVERFLOW:
            return "Mutation ID overflow.";
        case MSPB_ERR_UNKNOWN_TIME_NOT_SUPPORTED:
            return "Kept mutations must have known mutation times (which can be added using "
                   "compute_mutation_times).";
        case MSPB_ERR_CUDA:
            return "CUDA runtime error";
        case MSPB_ERR_NO_DEVICE:
            return "No usable CUDA device (this library has no CPU fallback)";
        case MSPB_ERR_TABLES_NOT_LOADED:
            return "Tables have not been loaded into the context";
        case MSPB_ERR_REPLAY_STREAM_EXHAUSTED:
            return "Replay stream exhausted";
        case MSPB_ERR_REPLAY_REDRAW:
            return "Replay stream needs a redraw (value >= 1, x >= hi or a position collision)";
        case MSPB_ERR_TOO_MANY_ROWS:
            return "Too many rows for 32-bit ids";
        case MSPB_ERR_UNSORTED_MUTATIONS:
            return "Kept mutations must be grouped by site in site order";
        case MSPB_ERR_UNSORTED_SITES:
            return "A genome-sliced call needs the kept sites sorted by position and the kept mutations by site";
        case MSPB_ERR_OFFSET_OVERFLOW:
            return "A ragged column may exceed 4 GiB: 32-bit offsets cannot be used";
        case MSPB_ERR_NOT_GENERATED:
            return "No result is resident: call mspb_generate first";
        case MSPB_ERR_KEPT_MUTATIONS_UNSORTED:
            return "Existing mutations of a site are not listed oldest first and a new mutation falls on that site "
                   "(the reference's result is undefined for such tables)";
        case MSPB_ERR_RETRY_WITH_HOST:
            return "A rank of this sharded call needs the host between its stages: make the call again with "
                   "mspb_generate_begin";
        case MSPB_ERR_NOALLOC_TOO_LARGE:
            return "Input exceeds the fixed scratch of the allocation-free entry point "
                   "(more than MSPB_NOALLOC_MAX_RECORDS records)";
        default:
            break;
    }
    if (err == TSK_E(-200)) {
        return "Bad offset provided in input array. (TSK_ERR_BAD_OFFSET)";
    }
    if (err < 0 && !(err & (1 << 13))) {
        return "tskit table integrity error (code = err ^ 8192)";
    }
    return "Unknown error";
}

#define CUDA_OK(call)                                                                                                  \
    do {                                                                                                               \
        cudaError_t e__ = (call);                                                                                      \
        if (e__ != cudaSuccess) {                                                                                      \
            fprintf(stderr, "msprime_b200: CUDA error %s at %s:%d\n", cudaGetErrorString(e__), __FILE__, __LINE__);    \
            return MSPB_ERR_CUDA;                                                                                      \
        }                                                                                                              \
    } while (0)

#define TRY(call)                                                                                                      \
    do {                                                                                                               \
        int r__ = (call);                                                                                              \
        if (r__ != 0) {                                                                                                \
            return r__;                                                                                                \
        }                                                                                                              \
    } while (0)

static inline unsigned grid_for(uint64_t n, unsigned threads)
{
    return (unsigned) ((n + threads - 1) / threads);
}

extern "C" void *mspb_pinned_alloc(uint64_t bytes);
extern "C" void mspb_pinned_free(void *p);

// ---------------------------------------------------------------- uploads from pageable memory
// A cudaMemcpyAsync from pageable memory is staged by the driver on one thread at ~10 GB/s (25 ms for the
// 272 MB of a bench-size tree sequence, against 5 ms of DMA).  While a StageScope lives on the calling thread,
// large columns that are NOT page-locked are copied into blocks of the pinned pool by a few host threads first
// and go to the device from there; the scope returns the blocks to the pool when it ends, which must be after
// the streams have been drained (declare it BEFORE the StreamDrain of the function).
struct StageScope {
    std::vector<void *> blocks;
    StageScope *outer;
    static StageScope *&current()
    {
        static thread_local StageScope *cur = nullptr;
        return cur;
    }
    StageScope() : outer(current())
    {
        current() = this;
    }
    ~StageScope()
    {
        for (void *b : blocks) {
            mspb_pinned_free(b);
        }
        current() = outer;
    }
};

constexpr size_t STAGE_MIN_BYTES = 4u << 20, STAGE_CHUNK_BYTES = 2u << 20;
constexpr unsigned STAGE_MAX_THREADS = 8; // (1 / 2 / 4 / 8 / 12 / 16 threads: 47 / 34 / 29 / 27 / 26 / 28 ms per bench-size one-shot call)

// the source as the device will read it: `src` itself when it is page-locked (or small, or no scope is
// open), else a pinned copy of it
static const void *staged_source(const void *src, size_t bytes)
{
    StageScope *scope = StageScope::current();
    if (scope == nullptr || bytes < STAGE_MIN_BYTES) {
        return src;
    }
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, src) != cudaSuccess) {
        cudaGetLastError();
        return src;
    }
    if (attr.type != cudaMemoryTypeUnregistered) {
        return src;
    }
    char *stage = (char *) mspb_pinned_alloc(bytes);
    if (stage == nullptr) {
        return src;
    }
    scope->blocks.push_back(stage);
    unsigned threads = std::thread::hardware_concurrency();
    threads = threads == 0 ? 1 : (threads > STAGE_MAX_THREADS ? STAGE_MAX_THREADS : threads);
    const size_t chunks = (bytes + STAGE_CHUNK_BYTES - 1) / STAGE_CHUNK_BYTES;
    if ((size_t) threads > chunks) {
        threads = (unsigned) chunks;
    }
    auto work = [=](unsigned k) {
        for (size_t c = k; c < chunks; c += threads) {
            const size_t off = c * STAGE_CHUNK_BYTES;
            const size_t len = off + STAGE_CHUNK_BYTES <= bytes ? STAGE_CHUNK_BYTES : bytes - off;
            memcpy(stage + off, (const char *) src + off, len);
        }
    };
    std::vector<std::thread> pool;
    for (unsigned k = 1; k < threads; k++) {
        pool.emplace_back(work, k);
    }
    work(0);
    for (std::thread &t : pool) {
        t.join();
    }
    return stage;
}

template <typename T>
static int upload(DBuf<T> &dst, const T *src, size_t n, cudaStream_t st)
{
    TRY(dst.ensure(n > 0 ? n : 1));
    if (n > 0) {
        CUDA_OK(cudaMemcpyAsync(dst.p, staged_source(src, n * sizeof(T)), n * sizeof(T), cudaMemcpyHostToDevice, st));
    }
    return 0;
}

extern "C" int mspb_ctx_create(int device, mspb_ctx_t **out)
{
    int count = 0;
    if (out == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    *out = nullptr;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0 || device < 0 || device >= count) {
        cudaGetLastError();
        return MSPB_ERR_NO_DEVICE;
    }
    if (cudaSetDevice(device) != cudaSuccess) {
        cudaGetLastError();
        return MSPB_ERR_NO_DEVICE;
    }
    mspb_ctx *ctx = new (std::nothrow) mspb_ctx();
    if (ctx == nullptr) {
        return MSPB_ERR_NO_MEMORY;
    }
    ctx->device = device;
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete ctx;
        return MSPB_ERR_CUDA;
    }
    bool events_ok = true;
    for (auto &e : ctx->ev) {
        events_ok = events_ok && cudaEventCreate(&e) == cudaSuccess;
    }
    for (cudaEvent_t *e : { &ctx->ev_k1, &ctx->ev_keys, &ctx->ev_p1, &ctx->ev_cf, &ctx->ev_copied }) {
        events_ok = events_ok && cudaEventCreateWithFlags(e, cudaEventDisableTiming) == cudaSuccess;
    }
    for (cudaEvent_t *e : { &ctx->ev_async0, &ctx->ev_async1 }) {
        events_ok = events_ok && cudaEventCreate(e) == cudaSuccess;
    }
    if (!events_ok || cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess) {
        cudaGetLastError();
        mspb_ctx_destroy(ctx);
        return MSPB_ERR_CUDA;
    }
    ctx->log.stream = ctx->stream;
    if (cudaMallocHost((void **) &ctx->host, sizeof(HostTotals)) != cudaSuccess
        || cudaMallocHost((void **) &ctx->or_and_host, 2 * sizeof(uint64_t)) != cudaSuccess
        || ctx->dev_totals.ensure(sizeof(HostTotals)) != 0 || ctx->or_and.ensure(2) != 0) {
        mspb_ctx_destroy(ctx);
        return MSPB_ERR_NO_MEMORY;
    }
    if (cudaHostGetDevicePointer((void **) &ctx->host_dev, ctx->host, 0) != cudaSuccess
        || cudaHostGetDevicePointer((void **) &ctx->or_and_host_dev, ctx->or_and_host, 0) != cudaSuccess) {
        cudaGetLastError();
        ctx->host_dev = nullptr; // (no unified addressing: fall back to copies)
        ctx->or_and_host_dev = nullptr;
    }
    *out = ctx;
    return 0;
}

// dt (device totals) -> ctx->host, in stream order
static int totals_to_host(mspb_ctx *ctx)
{
    static_assert(sizeof(HostTotals) % 4 == 0, "copied as 32-bit words");
    if (ctx->host_dev != nullptr) {
        rs_words_to_host_kernel<<<1, 64, 0, ctx->stream>>>(reinterpret_cast<const uint32_t *>(ctx->dev_totals.p),
            reinterpret_cast<uint32_t *>(ctx->host_dev), (uint32_t) (sizeof(HostTotals) / 4));
        ctx->log.count++;
        CUDA_OK(cudaGetLastError());
    } else {
        CUDA_OK(cudaMemcpyAsync(ctx->host, ctx->dev_totals.p, sizeof(HostTotals), cudaMemcpyDeviceToHost, ctx->stream));
    }
    return 0;
}

extern "C" void mspb_ctx_destroy(mspb_ctx_t *ctx)
{
    if (ctx == nullptr) {
        return;
    }
    cudaSetDevice(ctx->device);
    if (ctx->stream != nullptr) {
        cudaStreamSynchronize(ctx->stream);
    }
    if (ctx->copy_stream != nullptr) {
        cudaStreamSynchronize(ctx->copy_stream);
    }
#define REL(b) ctx->b.release()
    REL(node_time); REL(edge_left); REL(edge_right); REL(ci);
    REL(edge_parent); REL(edge_child); REL(child_start);
    REL(k_site_pos); REL(k_mut_time); REL(k_site_anc_off); REL(k_site_md_off); REL(k_mut_der_off);
    REL(k_mut_md_off); REL(k_site_anc); REL(k_site_md); REL(k_mut_der); REL(k_mut_md);
    REL(k_mut_site); REL(k_mut_node); REL(k_site_allele); REL(k_mut_allele);
    REL(map_pos); REL(map_rate); REL(map_cum); REL(map_inv); REL(model_root); REL(model_matrix); REL(replay);
    REL(model_alleles); REL(model_allele_off); REL(chain_free);
    REL(goff); REL(loff); REL(k1_tiles); REL(xchg); REL(slice_edges); REL(slice_scan_tiles); REL(rec); REL(keys0); REL(keys1); REL(k32_0); REL(k32_1); REL(vals0); REL(vals1);
    REL(tile_hist); REL(rs_scan_tiles); REL(or_and); REL(lb_state);
    REL(site_scan); REL(mut_scan); REL(draw_scan); REL(multi_list); REL(parent); REL(state); REL(site_state);
    REL(rp_edge); REL(rp_interval); REL(rp_count); REL(rp_out_off); REL(rp_stream_off); REL(replay_set); REL(dev_totals);
    REL(o_site_pos); REL(o_mut_time); REL(o_site_anc); REL(o_site_md); REL(o_mut_der); REL(o_mut_md);
    REL(o_site_anc_off); REL(o_site_md_off); REL(o_mut_der_off); REL(o_mut_md_off);
    REL(o_mut_site); REL(o_mut_node); REL(o_mut_parent);
    REL(sidx.is_sample); REL(sidx.foff); REL(sidx.fcnt); REL(sidx.samples); REL(sidx.xs); REL(q_below); REL(q_own); REL(q_afs);
    REL(q_row_begin); REL(q_row_end); REL(q_allele); REL(q_walk_id); REL(q_tmax); REL(q_geno);
#undef REL
    if (ctx->sidx.fx != nullptr) {
        cudaFree(ctx->sidx.fx);
    }
    if (ctx->sidx.fv != nullptr) {
        cudaFree(ctx->sidx.fv);
    }
    ctx->log.release();
    if (ctx->host != nullptr) {
        cudaFreeHost(ctx->host);
    }
    if (ctx->or_and_host != nullptr) {
        cudaFreeHost(ctx->or_and_host);
    }
    if (ctx->xchg_host != nullptr) {
        cudaFreeHost(ctx->xchg_host);
    }
    for (auto &e : ctx->ev) {
        if (e != nullptr) {
            cudaEventDestroy(e);
        }
    }
    for (cudaEvent_t e : { ctx->ev_k1, ctx->ev_keys, ctx->ev_copied, ctx->ev_p1, ctx->ev_cf, ctx->ev_async0,
             ctx->ev_async1 }) {
        if (e != nullptr) {
            cudaEventDestroy(e);
        }
    }
    if (ctx->copy_stream != nullptr) {
        cudaStreamDestroy(ctx->copy_stream);
    }
    if (ctx->stream != nullptr) {
        cudaStreamDestroy(ctx->stream);
    }
    delete ctx;
}

extern "C" uint64_t mspb_kernel_launches(const mspb_ctx_t *ctx)
{
    return ctx == nullptr ? 0 : ctx->log.count;
}

extern "C" void *mspb_ctx_stream(const mspb_ctx_t *ctx)
{
    return ctx == nullptr ? nullptr : (void *) ctx->stream;
}

extern "C" int mspb_kernel_timing(mspb_ctx_t *ctx, int enable)
{
    if (ctx == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    ctx->log.reset();
    ctx->log.timing = enable != 0;
    return 0;
}

extern "C" int mspb_kernel_time(const mspb_ctx_t *ctx, int index, const char **name, uint64_t *launches, double *ms)
{
    if (ctx == nullptr || index < 0 || (size_t) index >= ctx->log.names.size()) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    const std::string &n = ctx->log.names[(size_t) index];
    const auto it = ctx->log.stats.find(n);
    if (name != nullptr) {
        *name = n.c_str();
    }
    if (launches != nullptr) {
        *launches = it->second.launches;
    }
    if (ms != nullptr) {
        *ms = it->second.ms;
    }
    return 0;
}

extern "C" int mspb_phase_ms(const mspb_ctx_t *ctx, float *ms, int n)
{
    if (ctx == nullptr || ms == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    for (int i = 0; i < n && i < MSPB_NUM_PHASES; i++) {
        ms[i] = ctx->phase_ms[i];
    }
    return 0;
}

// ---------------------------------------------------------------- table validation
// The subset of tskit's check_integrity(options=0) that the reference runs before
// generating (lib/mutgen.c:1316; tskit tables.c:10397-10470, 10519-10531, 10565-10598).
// The lowest (table, row) wins so the reported error matches a serial scan.
namespace {

__device__ __forceinline__ void report_bad(unsigned long long *first_bad, int table, uint64_t row, int code)
{
    const unsigned long long v
        = ((unsigned long long) table << 60) | ((unsigned long long) row << 16) | (unsigned long long) (-code);
    atomicMin(first_bad, v);
}

__global__ void validate_nodes_kernel(uint64_t n, const double *time, unsigned long long *first_bad)
{
    const uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n && !isfinite(time[j])) {
        report_bad(first_bad, 0, j, -210);
    }
}

__global__ void validate_edges_kernel(TreeSeqView ts, unsigned long long *first_bad)
{
    const uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= ts.num_edges) {
        return;
    }
    const int32_t p = ts.edge_parent[j], c = ts.edge_child[j];
    const double l = ts.edge_left[j], r = ts.edge_right[j];
    int code = 0;
    if (p == -1) {
        code = -300;
    } else if (p < 0 || (uint64_t) p >= ts.num_nodes) {
        code = -202;
    } else if (c == -1) {
        code = -301;
    } else if (c < 0 || (uint64_t) c >= ts.num_nodes) {
        code = -202;
    } else if (!(isfinite(l) && isfinite(r))) {
        code = -211;
    } else if (l < 0) {
        code = -310;
    } else if (r > ts.sequence_length) {
        code = -309;
    } else if (l >= r) {
        code = -307;
    } else if (ts.node_time[c] >= ts.node_time[p]) {
        code = -306;
    }
    if (code != 0) {
        report_bad(first_bad, 1, j, code);
    }
}

__global__ void validate_sites_kernel(
    uint64_t n, const double *pos, double L, unsigned long long *first_bad, uint32_t *fractional)
{
    const uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) {
        return;
    }
    if (!isfinite(pos[j]) || pos[j] < 0 || pos[j] >= L) {
        report_bad(first_bad, 2, j, -402);
    } else if (pos[j] != floor(pos[j]) || pos[j] >= 4294967296.0) {
        atomicAdd(fractional, 1u);
    }
}

__global__ void validate_mutations_kernel(uint64_t n, const int32_t *site, const int32_t *node, const double *time,
    uint64_t num_sites, uint64_t num_nodes, const double *node_time, unsigned long long *first_bad)
{
    const uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) {
        return;
    }
    int code = 0;
    if (site[j] < 0 || (uint64_t) site[j] >= num_sites) {
        code = -205;
    } else if (node[j] < 0 || (uint64_t) node[j] >= num_nodes) {
        code = -202;
    } else if (time != nullptr) {
        const double t = time[j];
        if ((uint64_t) __double_as_longlong(t) != 0x7FF874736B697421ull) {
            if (!isfinite(t)) {
                code = -210;
            } else if (t < node_time[node[j]]) {
                code = -506;
            }
        }
    }
    if (code != 0) {
        report_bad(first_bad, 3, j, code);
    }
}

TreeSeqView make_ts_view(const mspb_ctx *ctx)
{
    TreeSeqView v;
    v.num_nodes = ctx->N;
    v.num_edges = ctx->E;
    v.sequence_length = ctx->L;
    v.node_time = ctx->node_time.p;
    v.edge_left = ctx->edge_left.p;
    v.edge_right = ctx->edge_right.p;
    v.edge_parent = ctx->edge_parent.p;
    v.edge_child = ctx->edge_child.p;
    v.child_start = ctx->child_start.p;
    v.ci = ctx->ci.p;
    return v;
}

int ensure_sort_buffers(mspb_ctx *ctx, uint64_t n, bool k32, bool k64)
{
    if (k64) {
        TRY(ctx->keys0.ensure(n + 1));
        TRY(ctx->keys1.ensure(n + 1));
    }
    if (k32) {
        TRY(ctx->k32_0.ensure(n + 1));
        TRY(ctx->k32_1.ensure(n + 1));
    }
    TRY(ctx->vals0.ensure(n + 1));
    TRY(ctx->vals1.ensure(n + 1));
    TRY(ctx->tile_hist.ensure((size_t) RS_MAX_PASSES * RS_BINS + 1));
    TRY(ctx->rs_scan_tiles.ensure(rs_state_words(n)));
    return 0;
}

RadixSortBuffers<uint64_t> sort_buffers64(mspb_ctx *ctx)
{
    RadixSortBuffers<uint64_t> b;
    b.keys[0] = ctx->keys0.p;
    b.keys[1] = ctx->keys1.p;
    b.vals[0] = ctx->vals0.p;
    b.vals[1] = ctx->vals1.p;
    b.hist = ctx->tile_hist.p;
    b.tile_state = ctx->rs_scan_tiles.p;
    b.or_and = ctx->or_and.p;
    b.or_and_host = ctx->or_and_host;
    b.or_and_host_dev = ctx->or_and_host_dev;
    return b;
}

RadixSortBuffers<uint32_t> sort_buffers32(mspb_ctx *ctx)
{
    RadixSortBuffers<uint32_t> b;
    b.keys[0] = ctx->k32_0.p;
    b.keys[1] = ctx->k32_1.p;
    b.vals[0] = ctx->vals0.p;
    b.vals[1] = ctx->vals1.p;
    b.hist = ctx->tile_hist.p;
    b.tile_state = ctx->rs_scan_tiles.p;
    b.or_and = ctx->or_and.p;
    b.or_and_host = ctx->or_and_host;
    b.or_and_host_dev = ctx->or_and_host_dev;
    return b;
}

// look-back state for `tiles` tiles of `halves` 16-byte words each, cleared on the stream
int lookback_state(mspb_ctx *ctx, uint64_t tiles, int halves, LbWord **lb)
{
    TRY(ctx->lb_state.ensure((tiles + 1) * halves));
    CUDA_OK(cudaMemsetAsync(ctx->lb_state.p, 0, (tiles + 1) * halves * sizeof(LbWord), ctx->stream));
    *lb = ctx->lb_state.p;
    return 0;
}

} // namespace

namespace {

struct StreamDrain {
    // whatever way a loader returns, no copy from the caller's buffers is left in flight
    cudaStream_t a, b;
    ~StreamDrain()
    {
        cudaStreamSynchronize(a);
        cudaStreamSynchronize(b);
    }
};

int tables_prepare(mspb_ctx_t *ctx, const mspb_tables_t *t)
{
    ctx->tables_loaded = false;
    ctx->tables_pending = false;
    ctx->generated = false;
    ctx->sidx.valid = false;
    if (!std::isfinite(t->sequence_length) || t->sequence_length <= 0) {
        /* mutgen_alloc rejects L <= 0 before anything else (lib/mutgen.c:760) */
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (t->num_nodes > 0x7fffffffull || t->num_edges > 0x7fffffffull || t->num_sites > 0x7fffffffull
        || t->num_mutations > 0x7fffffffull) {
        return MSPB_ERR_TOO_MANY_ROWS;
    }
    ctx->kept_anc_bytes = ctx->kept_smd_bytes = ctx->kept_der_bytes = ctx->kept_mmd_bytes = 0;
    ctx->L = t->sequence_length;
    ctx->N = t->num_nodes;
    ctx->E = t->num_edges;
    ctx->S0 = t->num_sites;
    ctx->M0 = t->num_mutations;
    return 0;
}

// A ragged column's offsets must start at 0 and never decrease (tskit's check_offsets, tables.c:
// TSK_ERR_BAD_OFFSET, applied by tsk_*_table_set_columns before the reference's mutgen ever sees a table):
// the kernels take lengths as differences of neighbours and would read past the column otherwise.
bool offsets_ok(const uint64_t *off, uint64_t rows)
{
    if (off == nullptr) {
        return true; // absent metadata column: all rows empty
    }
    uint64_t bad = off[0] != 0 ? 1 : 0;
    for (uint64_t j = 0; j < rows; j++) {
        bad |= off[j + 1] < off[j] ? 1 : 0;
    }
    return bad == 0;
}

// existing sites and mutations (read by keep=True calls), on stream cs
int tables_upload_kept(mspb_ctx_t *ctx, const mspb_tables_t *t, cudaStream_t cs)
{
    static const uint64_t zero_off[1] = { 0 };
    if (ctx->S0 > 0
        && (t->site_position == nullptr || t->site_ancestral_state_offset == nullptr
            || (t->site_ancestral_state == nullptr && t->site_ancestral_state_offset[ctx->S0] != 0))) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (ctx->M0 > 0
        && (t->mutation_site == nullptr || t->mutation_node == nullptr || t->mutation_derived_state_offset == nullptr
            || (t->mutation_derived_state == nullptr && t->mutation_derived_state_offset[ctx->M0] != 0))) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if ((ctx->S0 > 0
            && !(offsets_ok(t->site_ancestral_state_offset, ctx->S0) && offsets_ok(t->site_metadata_offset, ctx->S0)))
        || (ctx->M0 > 0
            && !(offsets_ok(t->mutation_derived_state_offset, ctx->M0)
                && offsets_ok(t->mutation_metadata_offset, ctx->M0)))) {
        return TSK_E(-200); // TSK_ERR_BAD_OFFSET
    }
    if (ctx->S0 > 0) {
        const uint64_t *ao = t->site_ancestral_state_offset, *mo = t->site_metadata_offset;
        TRY(upload(ctx->k_site_pos, t->site_position, ctx->S0, cs));
        TRY(upload(ctx->k_site_anc_off, ao, ctx->S0 + 1, cs));
        TRY(upload(ctx->k_site_anc, t->site_ancestral_state, (size_t) ao[ctx->S0], cs));
        ctx->kept_anc_bytes = ao[ctx->S0];
        ctx->kept_smd_bytes = mo != nullptr ? mo[ctx->S0] : 0;
        if (mo != nullptr) {
            TRY(upload(ctx->k_site_md_off, mo, ctx->S0 + 1, cs));
            TRY(upload(ctx->k_site_md, t->site_metadata, (size_t) mo[ctx->S0], cs));
        } else {
            TRY(ctx->k_site_md_off.ensure(ctx->S0 + 1));
            CUDA_OK(cudaMemsetAsync(ctx->k_site_md_off.p, 0, (ctx->S0 + 1) * sizeof(uint64_t), cs));
            TRY(ctx->k_site_md.ensure(1));
        }
    }
    if (ctx->M0 > 0) {
        const uint64_t *dof = t->mutation_derived_state_offset, *mo = t->mutation_metadata_offset;
        TRY(upload(ctx->k_mut_site, t->mutation_site, ctx->M0, cs));
        TRY(upload(ctx->k_mut_node, t->mutation_node, ctx->M0, cs));
        ctx->have_mut_time = t->mutation_time != nullptr;
        if (ctx->have_mut_time) {
            TRY(upload(ctx->k_mut_time, t->mutation_time, ctx->M0, cs));
        }
        TRY(upload(ctx->k_mut_der_off, dof, ctx->M0 + 1, cs));
        TRY(upload(ctx->k_mut_der, t->mutation_derived_state, (size_t) dof[ctx->M0], cs));
        ctx->kept_der_bytes = dof[ctx->M0];
        ctx->kept_mmd_bytes = mo != nullptr ? mo[ctx->M0] : 0;
        if (mo != nullptr) {
            TRY(upload(ctx->k_mut_md_off, mo, ctx->M0 + 1, cs));
            TRY(upload(ctx->k_mut_md, t->mutation_metadata, (size_t) mo[ctx->M0], cs));
        } else {
            TRY(ctx->k_mut_md_off.ensure(ctx->M0 + 1));
            CUDA_OK(cudaMemsetAsync(ctx->k_mut_md_off.p, 0, (ctx->M0 + 1) * sizeof(uint64_t), cs));
            TRY(ctx->k_mut_md.ensure(1));
        }
    }
    (void) zero_off;
    return 0;
}

// K0, first half: the stable sort by left, then by child, reads only edge_left and edge_child
// (any values sort; the rows are validated before the permutation uses them)
int k0_sort(mspb_ctx_t *ctx, RadixSortBuffers<uint64_t> *sb_out, int *cur_out)
{
    cudaStream_t st = ctx->stream;
    int k0_cur = 0;
    RadixSortBuffers<uint64_t> k0_sb;
    memset(&k0_sb, 0, sizeof(k0_sb));
    if (ctx->E > 0) {
        TreeSeqView tsk = make_ts_view(ctx);
        TRY(ensure_sort_buffers(ctx, ctx->E, false, true));
        k0_sb = sort_buffers64(ctx);
        MSPB_LAUNCH(ctx->log, "k0_child_keys", k0_child_keys_kernel<<<grid_for(ctx->E, 256), 256, 0, st>>>(tsk, k0_sb.keys[0], k0_sb.vals[0]));
        int cur = device_radix_sort<uint64_t>(k0_sb, ctx->E, 0, false, st, ctx->log, nullptr);
        if (cur != 0) {
            std::swap(k0_sb.keys[0], k0_sb.keys[1]);
            std::swap(k0_sb.vals[0], k0_sb.vals[1]);
        }
        MSPB_LAUNCH(ctx->log, "k0_child_keys2", k0_child_keys2_kernel<<<grid_for(ctx->E, 256), 256, 0, st>>>(tsk, k0_sb.vals[0], k0_sb.keys[0]));
        k0_cur = device_radix_sort<uint64_t>(k0_sb, ctx->E, 0, false, st, ctx->log, nullptr);
    }
    *sb_out = k0_sb;
    *cur_out = k0_cur;
    return 0;
}

// validation of every row, then K0's second half
int tables_finish(mspb_ctx_t *ctx, RadixSortBuffers<uint64_t> k0_sb, int k0_cur)
{
    cudaStream_t st = ctx->stream;
    // validation
    HostTotals *dt = (HostTotals *) ctx->dev_totals.p;
    CUDA_OK(cudaMemsetAsync(&dt->first_bad, 0xff, sizeof(unsigned long long), st));
    CUDA_OK(cudaMemsetAsync(&dt->fractional_sites, 0, sizeof(uint32_t), st));
    TreeSeqView ts = make_ts_view(ctx);
    if (ctx->N > 0) {
        MSPB_LAUNCH(ctx->log, "validate_nodes", validate_nodes_kernel<<<grid_for(ctx->N, 256), 256, 0, st>>>(ctx->N, ctx->node_time.p, &dt->first_bad));
    }
    if (ctx->E > 0) {
        MSPB_LAUNCH(ctx->log, "validate_edges", validate_edges_kernel<<<grid_for(ctx->E, 256), 256, 0, st>>>(ts, &dt->first_bad));
    }
    if (ctx->S0 > 0) {
        MSPB_LAUNCH(ctx->log, "validate_sites", validate_sites_kernel<<<grid_for(ctx->S0, 256), 256, 0, st>>>(
            ctx->S0, ctx->k_site_pos.p, ctx->L, &dt->first_bad, &dt->fractional_sites));
    }
    if (ctx->M0 > 0) {
        MSPB_LAUNCH(ctx->log, "validate_mutations", validate_mutations_kernel<<<grid_for(ctx->M0, 256), 256, 0, st>>>(ctx->M0, ctx->k_mut_site.p, ctx->k_mut_node.p,
            ctx->have_mut_time ? ctx->k_mut_time.p : nullptr, ctx->S0, ctx->N, ctx->node_time.p, &dt->first_bad));
    }
    TRY(totals_to_host(ctx));
    CUDA_OK(cudaStreamSynchronize(st));
    ctx->kept_fractional = ctx->host->fractional_sites > 0;
    if (ctx->host->first_bad != ~0ull) {
        const int code = -(int) (ctx->host->first_bad & 0xffffull);
        return TSK_E(code);
    }

    // K0, second half: the 32-byte entries in (child, left) order, then the parent links
    if (ctx->E == 0) {
        CUDA_OK(cudaMemsetAsync(ctx->child_start.p, 0, (ctx->N + 2) * sizeof(uint32_t), st));
    } else {
        MSPB_LAUNCH(ctx->log, "k0_permute", k0_permute_kernel<<<grid_for(ctx->E, 256), 256, 0, st>>>(
            ts, k0_sb.vals[k0_cur], ctx->ci.p, ctx->child_start.p));
        // MSPB_K0_MAX_SPAN (tests only) lowers the span-count limit so that the walk's range
        // recovery path runs on small inputs
        uint32_t max_count = 256;
        if (const char *env = getenv("MSPB_K0_MAX_SPAN")) {
            const long v = strtol(env, nullptr, 10);
            if (v >= 1 && v <= 256) {
                max_count = (uint32_t) v;
            }
        }
        MSPB_LAUNCH(ctx->log, "k0_link", k0_link_kernel<<<grid_for(ctx->E, 256), 256, 0, st>>>(ts, ctx->ci.p, max_count));
    }
    CUDA_OK(cudaStreamSynchronize(st));
    CUDA_OK(cudaGetLastError());
    ctx->log.collect();
    ctx->tables_loaded = true;
    ctx->slice_cache_valid = false;
    return 0;
}

} // namespace

extern "C" int mspb_ctx_load_tables(mspb_ctx_t *ctx, const mspb_tables_t *t)
{
    if (ctx == nullptr || t == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    TRY(tables_prepare(ctx, t));
    StageScope staging; // (ends after the drain below)
    StreamDrain drain { st, ctx->copy_stream };
    // The two columns K0's sorts read go first, on the main stream; everything else follows on
    // the copy stream, so that with pinned sources the sorts run under the rest of the upload.
    cudaStream_t cs = ctx->copy_stream;
    TRY(upload(ctx->edge_left, t->edge_left, ctx->E, st));
    TRY(upload(ctx->edge_child, t->edge_child, ctx->E, st));
    CUDA_OK(cudaEventRecord(ctx->ev_keys, st));
    CUDA_OK(cudaStreamWaitEvent(cs, ctx->ev_keys, 0)); // one copy engine direction: keep the order
    TRY(upload(ctx->edge_right, t->edge_right, ctx->E, cs));
    TRY(upload(ctx->edge_parent, t->edge_parent, ctx->E, cs));
    TRY(upload(ctx->node_time, t->node_time, ctx->N, cs));
    TRY(tables_upload_kept(ctx, t, cs));
    TRY(ctx->child_start.ensure(ctx->N + 2));
    TRY(ctx->ci.ensure(ctx->E + 1));
    CUDA_OK(cudaEventRecord(ctx->ev_copied, cs));
    RadixSortBuffers<uint64_t> k0_sb;
    int k0_cur = 0;
    TRY(k0_sort(ctx, &k0_sb, &k0_cur));
    CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_copied, 0));
    return tables_finish(ctx, k0_sb, k0_cur);
}

// Sharded upload of the replicated tables (multi-GPU): this rank copies its 1 / world share of
// the five big columns; the caller all-gathers the shares in place on the context's stream
// (NVLink instead of `world` copies of everything over PCIe); mspb_ctx_load_tables_end then
// validates and builds the index.
extern "C" int mspb_ctx_load_tables_begin(mspb_ctx_t *ctx, const mspb_tables_t *t, uint32_t rank, uint32_t world,
    mspb_shard_t *shards)
{
    if (ctx == nullptr || t == nullptr || shards == nullptr || world == 0 || rank >= world) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    TRY(tables_prepare(ctx, t));
    StageScope staging; // (ends after the drain below)
    StreamDrain drain { st, ctx->copy_stream };
#define SHARD(k, buf, src, n, T)                                                                                     \
    do {                                                                                                             \
        const uint64_t per = ((uint64_t) (n) + world - 1) / world;                                                   \
        TRY(ctx->buf.ensure((size_t) (per * world > 0 ? per * world : 1)));                                          \
        const uint64_t lo = (uint64_t) rank * per;                                                                   \
        const uint64_t hi = lo + per < (uint64_t) (n) ? lo + per : (uint64_t) (n);                                   \
        if (hi > lo) {                                                                                               \
            CUDA_OK(cudaMemcpyAsync(ctx->buf.p + lo, staged_source((src) + lo, (hi - lo) * sizeof(T)),              \
                (hi - lo) * sizeof(T), cudaMemcpyHostToDevice, st));                                                 \
        }                                                                                                            \
        shards[k].ptr = ctx->buf.p;                                                                                  \
        shards[k].bytes_per_rank = per * sizeof(T);                                                                  \
    } while (0)
    SHARD(0, node_time, t->node_time, ctx->N, double);
    SHARD(1, edge_left, t->edge_left, ctx->E, double);
    SHARD(2, edge_right, t->edge_right, ctx->E, double);
    SHARD(3, edge_parent, t->edge_parent, ctx->E, int32_t);
    SHARD(4, edge_child, t->edge_child, ctx->E, int32_t);
#undef SHARD
    TRY(tables_upload_kept(ctx, t, st));
    TRY(ctx->child_start.ensure(ctx->N + 2));
    TRY(ctx->ci.ensure(ctx->E + 1));
    ctx->tables_pending = true;
    return 0;
}

extern "C" int mspb_ctx_load_tables_end(mspb_ctx_t *ctx)
{
    if (ctx == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (!ctx->tables_pending) {
        return MSPB_ERR_TABLES_NOT_LOADED;
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    ctx->tables_pending = false;
    RadixSortBuffers<uint64_t> k0_sb;
    int k0_cur = 0;
    TRY(k0_sort(ctx, &k0_sb, &k0_cur));
    return tables_finish(ctx, k0_sb, k0_cur);
}

// ---------------------------------------------------------------- generate
namespace {

// Knuth's relative comparison, as the reference's doubles_almost_equal uses it
// (lib/util.c:372-376 via gsl_fcmp).
bool almost_equal(double a, double b, double eps)
{
    if (std::fabs(a) < eps && std::fabs(b) < eps) {
        return true;
    }
    int exponent;
    const double mx = (std::fabs(a) > std::fabs(b)) ? a : b;
    std::frexp(mx, &exponent);
    const double delta = std::ldexp(eps, exponent);
    const double diff = a - b;
    return !(diff > delta) && !(diff < -delta);
}

bool valid_probabilities(uint64_t n, const double *p)
{
    double sum = 0, lo = 0;
    for (uint64_t j = 0; j < n; j++) {
        sum += p[j];
        lo = p[j] < lo ? p[j] : lo;
    }
    return almost_equal(sum, 1.0, 1e-12) && lo >= 0.0;
}

// lib/mutgen.c:501-524
int check_model(const mspb_params_t *p)
{
    if (p->model_kind == MSPB_MODEL_INFINITE_ALLELES) {
        return 0;
    }
    if (p->model_kind == MSPB_MODEL_SLIM) {
        /* slim_mutator_check_validity, lib/mutgen.c:326-343 */
        return ((int64_t) p->start_allele < 0 || p->slim_mutation_type_id < 0) ? MSPB_ERR_BAD_SLIM_PARAMETERS : 0;
    }
    if (p->model_kind != MSPB_MODEL_MATRIX) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (p->num_alleles < 2) {
        return MSPB_ERR_INSUFFICIENT_ALLELES;
    }
    if (p->alleles == nullptr || p->allele_offset == nullptr || p->root_distribution == nullptr
        || p->transition_matrix == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (p->allele_offset[0] != 0) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    for (uint64_t i = 0; i < p->num_alleles; i++) {
        if (p->allele_offset[i + 1] < p->allele_offset[i]) {
            return MSPB_ERR_BAD_PARAM_VALUE; // allele lengths are differences of neighbouring offsets
        }
    }
    if (!valid_probabilities(p->num_alleles, p->root_distribution)) {
        return MSPB_ERR_BAD_ROOT_PROBABILITIES;
    }
    for (uint64_t i = 0; i < p->num_alleles; i++) {
        if (!valid_probabilities(p->num_alleles, p->transition_matrix + i * p->num_alleles)) {
            return MSPB_ERR_BAD_TRANSITION_MATRIX;
        }
    }
    return 0;
}

// lib/rate_map.c:45-100 and lib/mutgen.c:796-813
int check_rate_map(const mspb_params_t *p, double L)
{
    if (p->map_size < 1 || p->map_position == nullptr || p->map_rate == nullptr) {
        return MSPB_ERR_INSUFFICIENT_INTERVALS;
    }
    if (p->map_position[0] != 0.0) {
        return MSPB_ERR_INTERVAL_MAP_START_NON_ZERO;
    }
    for (uint64_t j = 0; j <= p->map_size; j++) {
        if (!std::isfinite(p->map_position[j])) {
            return MSPB_ERR_NONFINITE_INTERVAL_POSITION;
        }
        if (j < p->map_size) {
            if (p->map_position[j] >= p->map_position[j + 1]) {
                return MSPB_ERR_INTERVAL_POSITIONS_UNSORTED;
            }
            if (!std::isfinite(p->map_rate[j]) || p->map_rate[j] < 0) {
                return MSPB_ERR_BAD_RATE_VALUE;
            }
        }
    }
    if (p->map_position[p->map_size] != L) {
        return MSPB_ERR_INCOMPATIBLE_MUTATION_MAP_LENGTH;
    }
    return 0;
}

int upload_params(mspb_ctx *ctx, const mspb_params_t *p, RateMapView *map, ModelView *model)
{
    cudaStream_t st = ctx->stream;
    const uint64_t R = p->map_size;
    const bool discrete = p->discrete_genome != 0;
    // Repeated calls usually come with the same rate map (and model, below): the device copies
    // of the previous call are reused when the host arrays compare equal, which saves three
    // small copies and a stream wait per overlay.
    const bool same_map = ctx->cache_map_valid && ctx->cache_map_discrete == discrete
                          && ctx->cache_map_pos.size() == R + 1 && ctx->cache_map_rate.size() == R
                          && memcmp(ctx->cache_map_pos.data(), p->map_position, (R + 1) * sizeof(double)) == 0
                          && (R == 0 || memcmp(ctx->cache_map_rate.data(), p->map_rate, R * sizeof(double)) == 0);
    if (!same_map) {
        ctx->cache_map_valid = false;
        std::vector<double> rate(R + 1), cum(R + 1), inv(R + 1);
        double sum = 0;
        for (uint64_t j = 0; j <= R; j++) {
            cum[j] = sum;
            rate[j] = j < R ? p->map_rate[j] : 0.0;
            inv[j] = rate[j] > 0 ? 1.0 / rate[j] : 0.0;
            if (j < R) {
                const double a = discrete ? std::ceil(p->map_position[j]) : p->map_position[j];
                const double b = discrete ? std::ceil(p->map_position[j + 1]) : p->map_position[j + 1];
                sum += (b - a) * rate[j];
            }
        }
        TRY(upload(ctx->map_pos, p->map_position, R + 1, st));
        TRY(upload(ctx->map_rate, rate.data(), R + 1, st));
        TRY(upload(ctx->map_cum, cum.data(), R + 1, st));
        TRY(upload(ctx->map_inv, inv.data(), R + 1, st));
        // the vectors must outlive the async copies
        CUDA_OK(cudaStreamSynchronize(st));
        ctx->cache_map_pos.assign(p->map_position, p->map_position + R + 1);
        ctx->cache_map_rate.assign(p->map_rate, p->map_rate + R);
        ctx->cache_map_discrete = discrete;
        ctx->cache_map_valid = true;
    }
    map->pos = ctx->map_pos.p;
    map->rate = ctx->map_rate.p;
    map->cum = ctx->map_cum.p;
    map->inv_rate = ctx->map_inv.p;
    map->size = (uint32_t) R;
    map->discrete = discrete ? 1 : 0;

    model->kind = p->model_kind;
    model->start_allele = p->start_allele;
    model->slim_type = p->slim_mutation_type_id;
    model->slim_generation = p->slim_generation;
    model->num_alleles = 0;
    model->alleles = nullptr;
    model->allele_off = nullptr;
    model->root = nullptr;
    model->matrix = nullptr;
    if (p->model_kind == MSPB_MODEL_MATRIX) {
        const uint64_t A = p->num_alleles;
        const size_t nb = (size_t) p->allele_offset[A];
        const bool same_model = ctx->cache_model_valid && ctx->cache_model_off.size() == A + 1
                                && ctx->cache_model_bytes.size() == nb
                                && memcmp(ctx->cache_model_off.data(), p->allele_offset, (A + 1) * sizeof(uint64_t)) == 0
                                && (nb == 0 || memcmp(ctx->cache_model_bytes.data(), p->alleles, nb) == 0)
                                && memcmp(ctx->cache_model_root.data(), p->root_distribution, A * sizeof(double)) == 0
                                && memcmp(ctx->cache_model_matrix.data(), p->transition_matrix, A * A * sizeof(double))
                                       == 0;
        if (!same_model) {
            ctx->cache_model_valid = false;
            TRY(upload(ctx->model_alleles, p->alleles, nb, st));
            TRY(upload(ctx->model_allele_off, p->allele_offset, A + 1, st));
            TRY(upload(ctx->model_root, p->root_distribution, A, st));
            TRY(upload(ctx->model_matrix, p->transition_matrix, A * A, st));
            CUDA_OK(cudaStreamSynchronize(st)); // pageable sources: the caller may free them after the call
            ctx->cache_model_off.assign(p->allele_offset, p->allele_offset + A + 1);
            ctx->cache_model_bytes.assign(p->alleles, p->alleles + nb);
            ctx->cache_model_root.assign(p->root_distribution, p->root_distribution + A);
            ctx->cache_model_matrix.assign(p->transition_matrix, p->transition_matrix + A * A);
            ctx->cache_model_valid = true;
        }
        model->num_alleles = (uint32_t) A;
        model->alleles = ctx->model_alleles.p;
        model->allele_off = ctx->model_allele_off.p;
        model->root = ctx->model_root.p;
        model->matrix = ctx->model_matrix.p;
    }
    return 0;
}

// View of the kept rows [site_lo, site_lo + S0) and [mut_lo, mut_lo + M0) (the whole tables
// unless the call owns a genome slice).  Offset columns keep indexing the full byte arrays.
KeptView make_kept_view(const mspb_ctx *ctx, bool keep, uint64_t site_lo, uint64_t S0, uint64_t mut_lo, uint64_t M0)
{
    KeptView k;
    memset(&k, 0, sizeof(k));
    if (!keep) {
        return k;
    }
    k.num_sites = S0;
    k.num_mutations = M0;
    k.site_id_offset = (int32_t) site_lo;
    k.site_position = ctx->k_site_pos.p + site_lo;
    k.site_anc_off = ctx->k_site_anc_off.p + site_lo;
    k.site_md_off = ctx->k_site_md_off.p + site_lo;
    k.site_anc = ctx->k_site_anc.p;
    k.site_md = ctx->k_site_md.p;
    k.mut_site = ctx->k_mut_site.p + mut_lo;
    k.mut_node = ctx->k_mut_node.p + mut_lo;
    k.mut_time = ctx->have_mut_time ? ctx->k_mut_time.p + mut_lo : nullptr;
    k.mut_der_off = ctx->k_mut_der_off.p + mut_lo;
    k.mut_md_off = ctx->k_mut_md_off.p + mut_lo;
    k.mut_der = ctx->k_mut_der.p;
    k.mut_md = ctx->k_mut_md.p;
    k.site_allele = ctx->k_site_allele.p;
    k.mut_allele = ctx->k_mut_allele.p;
    return k;
}

// Sliced calls with kept rows: the rows inside [lo, hi) form one contiguous range when the
// site table is sorted by position (tskit's own requirement for a tree sequence).
__global__ void kept_slice_kernel(uint64_t S0, const double *pos, uint64_t M0, const int32_t *mut_site, double lo,
    double hi, uint32_t *out4)
{
    if (blockIdx.x != 0 || threadIdx.x != 0) {
        return;
    }
    uint64_t a = 0, b = S0;
    while (a < b) { // first site with position >= lo
        const uint64_t m = (a + b) >> 1;
        if (pos[m] < lo) {
            a = m + 1;
        } else {
            b = m;
        }
    }
    const uint64_t site_lo = a;
    b = S0;
    while (a < b) { // first site with position >= hi
        const uint64_t m = (a + b) >> 1;
        if (pos[m] < hi) {
            a = m + 1;
        } else {
            b = m;
        }
    }
    const uint64_t site_hi = a;
    uint64_t c = 0, d = M0;
    while (c < d) {
        const uint64_t m = (c + d) >> 1;
        if ((uint64_t) mut_site[m] < site_lo) {
            c = m + 1;
        } else {
            d = m;
        }
    }
    const uint64_t mut_lo = c;
    d = M0;
    while (c < d) {
        const uint64_t m = (c + d) >> 1;
        if ((uint64_t) mut_site[m] < site_hi) {
            c = m + 1;
        } else {
            d = m;
        }
    }
    out4[0] = (uint32_t) site_lo;
    out4[1] = (uint32_t) site_hi;
    out4[2] = (uint32_t) mut_lo;
    out4[3] = (uint32_t) c;
}

__global__ void kept_sorted_check_kernel(uint64_t S0, const double *pos, uint64_t M0, const int32_t *mut_site,
    uint32_t *unsorted)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 < S0 && !(pos[i] < pos[i + 1])) {
        *unsorted = 1u;
    }
    if (i + 1 < M0 && mut_site[i] > mut_site[i + 1]) {
        *unsorted = 1u;
    }
}

} // namespace

namespace {

uint32_t digits_u64(uint64_t v)
{
    uint32_t n = 1;
    while (v >= 10) {
        v /= 10;
        n++;
    }
    return n;
}

// longest state a new mutation or new site can get under this model
uint64_t max_new_state_len(const mspb_params_t *p, uint64_t Mnew)
{
    if (p->model_kind == MSPB_MODEL_MATRIX) {
        uint64_t m = 0;
        for (uint64_t a = 0; a < p->num_alleles; a++) {
            const uint64_t len = p->allele_offset[a + 1] - p->allele_offset[a];
            m = len > m ? len : m;
        }
        return m;
    }
    if (p->model_kind == MSPB_MODEL_SLIM) {
        return 20; /* first guess only: the sizing pass of K5 gives the exact totals */
    }
    const uint64_t last = p->start_allele + 2 * Mnew + 1;
    return last < p->start_allele ? 20 : digits_u64(last); /* 20 digits if the counter wraps */
}

struct StageB {
    SortedView sv;
    ResolveView rv;
};

// one thread per mutation instead of one block per 256 edges: when the edges are mostly empty (sharded calls;
// measured cross-over in DESIGN.md section 6)
bool k2_sparse_pays(uint64_t new_mutations, uint64_t edges)
{
    return new_mutations * 5 < edges * 4; // fewer than 0.8 mutations per edge
}

// K5 on the state saved by run_sorted_stages
int launch_emit(mspb_ctx *ctx)
{
    cudaStream_t st = ctx->stream;
    HostTotals *dt = (HostTotals *) ctx->dev_totals.p;
    const uint64_t R = ctx->pending.R;
    const uint32_t k5_tiles = (uint32_t) ((R + K5_TILE - 1) / K5_TILE);
    LbWord *lb4 = nullptr;
    TRY(lookback_state(ctx, k5_tiles, 2, &lb4));
    const bool kept_rows = ctx->pending.sv.has_kept != 0, off32 = ctx->pending.out.off32 != 0;
    const int model_kind = ctx->pending.model.kind;
    // Without kept rows, infinite-alleles and equal-length matrix states let every tile compute
    // its byte prefix on its own (ChainFree): the table is filled on the second stream, beside
    // K4, from the site rows P1 left.
    ctx->pending.out.chain_free = nullptr;
    // (Single calls, and sharded calls whose exchanges run beside K4 (pending.early: the table is enqueued
    // already).  Any other two-phase call learns its allele base after K4 has finished, so the table's ~30 us
    // would sit on the critical path, which is more than the chain costs for a slice's tile count.)
    if (ctx->pending.cf_ready) {
        ctx->pending.out.chain_free = ctx->chain_free.p;
    } else if (!kept_rows && !ctx->pending.out.dry_run && R > 0 && !ctx->pending.defer
        && (model_kind == MSPB_MODEL_INFINITE_ALLELES || (model_kind == MSPB_MODEL_MATRIX && ctx->pending.fixed_len > 0))) {
        TRY(ctx->chain_free.ensure(1));
        cudaStream_t cs = ctx->copy_stream;
        CUDA_OK(cudaStreamWaitEvent(cs, ctx->ev_p1, 0));
        k5_prefix_table_kernel<<<1, 32, 0, cs>>>(ctx->pending.sv.site_row, &dt->p1, R, ctx->pending.model.start_allele,
            ctx->pending.out.dev_bases, model_kind == MSPB_MODEL_MATRIX ? ctx->pending.fixed_len : 0u, &dt->dense,
            ctx->chain_free.p);
        ctx->log.count++;
        CUDA_OK(cudaEventRecord(ctx->ev_cf, cs));
        CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_cf, 0));
        ctx->pending.out.chain_free = ctx->chain_free.p;
    }
#define MSPB_K5(M, K, O)                                                                                             \
    MSPB_LAUNCH(ctx->log, "k5_emit",                                                                                 \
        (k5_emit_kernel<M, K, O><<<k5_tiles, K5_THREADS, 0, st>>>(ctx->pending.sv, ctx->pending.rv,                \
            ctx->pending.kept, ctx->pending.model, ctx->pending.call, ctx->pending.out, lb4, k5_tiles, &dt->p1,      \
            &dt->bytes, &dt->dense, &dt->collisions, &dt->err)))
#define MSPB_K5_MODEL(M)                                                                                             \
    if (kept_rows) {                                                                                                 \
        if (off32) {                                                                                                 \
            MSPB_K5(M, true, true);                                                                                  \
        } else {                                                                                                     \
            MSPB_K5(M, true, false);                                                                                 \
        }                                                                                                            \
    } else if (off32) {                                                                                              \
        MSPB_K5(M, false, true);                                                                                     \
    } else {                                                                                                         \
        MSPB_K5(M, false, false);                                                                                    \
    }
    if (model_kind == 0) {
        MSPB_K5_MODEL(0)
    } else if (model_kind == 1) {
        MSPB_K5_MODEL(1)
    } else {
        MSPB_K5_MODEL(2)
    }
#undef MSPB_K5_MODEL
#undef MSPB_K5
    return 0;
}

// K4 on the state saved by run_sorted_stages
int launch_resolve(mspb_ctx *ctx)
{
    HostTotals *dt = (HostTotals *) ctx->dev_totals.p;
    const TreeSeqView ts = make_ts_view(ctx);
    MSPB_LAUNCH(ctx->log, "k4_resolve",
        k4_resolve_kernel<<<148 * 12, 128, 0, ctx->stream>>>(ts, ctx->pending.kept, ctx->pending.model,
            ctx->pending.call, ctx->pending.sv, ctx->pending.rv, ctx->multi_list.p, &dt->k4_next, &dt->p1, &dt->dense,
            &dt->collisions, &dt->err));
    return 0;
}

// SLiM: state and metadata stack along parent chains, so their totals are not bounded up front:
// K5 first runs as a sizing pass (lengths and scan only), then for real.  Leaves the totals on
// the host (waits twice).
int slim_size_and_emit(mspb_ctx *ctx)
{
    cudaStream_t st = ctx->stream;
    HostTotals *ht = ctx->host;
    ctx->pending.out.dry_run = 1;
    TRY(launch_emit(ctx));
    ctx->pending.out.dry_run = 0;
    TRY(totals_to_host(ctx));
    CUDA_OK(cudaStreamSynchronize(st));
    if (ht->err != 0) {
        return ht->err;
    }
    if (ctx->pending.out.off32 && (ht->bytes.der >= 0xffffffffull || ht->bytes.mmd >= 0xffffffffull)) {
        return MSPB_ERR_OFFSET_OVERFLOW;
    }
    TRY(ctx->o_mut_der.ensure(ht->bytes.der + 1));
    TRY(ctx->o_mut_md.ensure(ht->bytes.mmd + 1));
    ctx->pending.out.mut_der = ctx->o_mut_der.p;
    ctx->pending.out.mut_md = ctx->o_mut_md.p;
    TRY(launch_emit(ctx));
    TRY(totals_to_host(ctx));
    return 0;
}

// sort -> P1 -> K4 -> K5 for one key mode; everything is enqueued, nothing is awaited
int run_sorted_stages(mspb_ctx *ctx, const mspb_params_t *p, const TreeSeqView &ts, const KeptView &kept,
    const ModelView &model, CallView call, uint64_t R, uint64_t S0, uint64_t M0, const OutView &out, bool first_sort)
{
    cudaStream_t st = ctx->stream;
    HostTotals *dt = (HostTotals *) ctx->dev_totals.p;
    const bool replay = p->replay_stream != nullptr;
    SortedView sv;
    memset(&sv, 0, sizeof(sv));
    sv.n = R;
    sv.S0 = S0;
    sv.M0 = M0;
    sv.rec = ctx->rec.p;
    sv.has_kept = (S0 + M0) > 0 ? 1 : 0;
    sv.root_draw = p->model_kind != MSPB_MODEL_SLIM ? 1 : 0;
    sv.site_row = ctx->site_scan.p;
    sv.mut_scan = ctx->mut_scan.p;
    sv.draw_scan = ctx->draw_scan.p;
    sv.key.mode = call.key_mode;
    sv.key.k32 = nullptr;
    sv.key.k64 = nullptr;

    // ---- K3: stable sort by position
    if (first_sort && ctx->log.timing) {
        CUDA_OK(cudaEventRecord(ctx->ev[3], st));
    }
    if (call.key_mode == KEY32_POS) {
        RadixSortBuffers<uint32_t> sb = sort_buffers32(ctx);
        uint64_t top = (uint64_t) std::ceil(ctx->L) - 1;
        uint64_t varying = 1;
        while (varying <= top) {
            varying <<= 1;
        }
        varying -= 1;
        const int cur = device_radix_sort<uint32_t>(sb, R, varying != 0 ? varying : 1, true, st, ctx->log, nullptr);
        sv.key.k32 = sb.keys[cur];
        sv.val = sb.vals[cur];
    } else {
        RadixSortBuffers<uint64_t> sb = sort_buffers64(ctx);
        // (identity payload on the first sort only: re-sorts after a redraw carry the slots along)
        const int cur = device_radix_sort<uint64_t>(
            sb, R, 0, call.key_mode == KEY64_POS && first_sort, st, ctx->log, nullptr);
        sv.key.k64 = sb.keys[cur];
        sv.val = sb.vals[cur];
    }

    // ---- P1: site boundaries, row and draw numbering
    if (first_sort && ctx->log.timing) {
        CUDA_OK(cudaEventRecord(ctx->ev[4], st));
    }
    const uint32_t p1_tiles = (uint32_t) ((R + P1_TILE - 1) / P1_TILE);
    LbWord *lb3 = nullptr;
    TRY(lookback_state(ctx, p1_tiles, 1, &lb3));
    if (!first_sort) {
        CUDA_OK(cudaMemsetAsync(&dt->dense, 0, sizeof(uint32_t), st)); // (the call's totals start at zero)
    }
#define MSPB_P1(MODE, K)                                                                                             \
    MSPB_LAUNCH(ctx->log, "p1_sites",                                                                                \
        (p1_sites_kernel<MODE, K><<<p1_tiles, SCAN_THREADS, 0, st>>>(sv, call.discrete, lb3, p1_tiles, &dt->p1,      \
            ctx->multi_list.p, &dt->collisions, &dt->dense, &dt->err)))
#define MSPB_P1_MODE(MODE)                                                                                           \
    if (sv.has_kept) {                                                                                               \
        MSPB_P1(MODE, true);                                                                                         \
    } else {                                                                                                         \
        MSPB_P1(MODE, false);                                                                                        \
    }
    if (call.key_mode == KEY32_POS) {
        MSPB_P1_MODE(KEY32_POS)
    } else if (call.key_mode == KEY64_POS) {
        MSPB_P1_MODE(KEY64_POS)
    } else {
        MSPB_P1_MODE(KEY64_POS_TIME)
    }
#undef MSPB_P1_MODE
#undef MSPB_P1

    CUDA_OK(cudaEventRecord(ctx->ev_p1, st)); // the site rows are final from here on
    // ---- K4: sites with several records
    if (first_sort && ctx->log.timing) {
        CUDA_OK(cudaEventRecord(ctx->ev[5], st));
    }
    ResolveView rv;
    rv.parent = ctx->parent.p;
    rv.state = ctx->state.p;
    rv.site_state = ctx->site_state.p;
    rv.replay = nullptr;
    rv.replay_len = 0;
    if (replay) {
        const uint64_t used = ctx->host->replay.draws;
        rv.replay = ctx->replay.p + used;
        rv.replay_len = p->replay_len - used;
    }
    ctx->pending.sv = sv;
    ctx->pending.rv = rv;
    ctx->pending.kept = kept;
    ctx->pending.model = model;
    ctx->pending.call = call;
    ctx->pending.out = out;
    ctx->pending.R = R;
    if (!ctx->pending.defer_resolve) {
        TRY(launch_resolve(ctx));
    }

    // ---- K5: byte offsets + rows
    if (first_sort && ctx->log.timing) {
        CUDA_OK(cudaEventRecord(ctx->ev[6], st));
    }
    if (!ctx->pending.defer) {
        TRY(launch_emit(ctx));
    }
    if (first_sort && ctx->log.timing) {
        CUDA_OK(cudaEventRecord(ctx->ev[7], st));
    }
    ctx->last_sv = sv;
    return 0;
}

} // namespace

static int generate_impl(mspb_ctx_t *ctx, const mspb_params_t *p, mspb_counts_t *counts_out, bool defer,
    bool async_begin = false)
{
    if (ctx == nullptr || p == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    ctx->pending.valid = false;
    ctx->pending.async = false;
    ctx->pending.early = false;
    ctx->pending.cf_ready = false;
    ctx->pending.defer = defer;
    ctx->pending.defer_resolve = false;
    ctx->pending.R = 0;
    if (!ctx->tables_loaded) {
        return MSPB_ERR_TABLES_NOT_LOADED;
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    ctx->generated = false;
    // same order of checks as the reference call sequence: model, rate map, time window
    TRY(check_model(p));
    TRY(check_rate_map(p, ctx->L));
    if (p->end_time < p->start_time) {
        return MSPB_ERR_BAD_PARAM_VALUE; /* lib/mutgen.c:786 */
    }
    const bool replay = p->replay_stream != nullptr;
    const bool in_slice = !(p->slice_left == 0.0 && (p->slice_right == 0.0 || p->slice_right >= ctx->L));
    if (in_slice && (replay || !(p->slice_left < p->slice_right) || p->slice_left < 0)) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    // two ways to own a slice: draw whole edges and keep what falls inside (slice_stream == 0:
    // the union of slices is exactly the unsliced result), or cut the edges to the slice and
    // draw on the slice's own Philox stream (slice_stream != 0: work scales with the slice)
    const bool clipped = in_slice && p->slice_stream != 0;
    const bool sliced = in_slice && !clipped;
    const bool keep = p->keep != 0;
    const bool discrete = p->discrete_genome != 0;
    uint64_t S0 = keep ? ctx->S0 : 0, M0 = keep ? ctx->M0 : 0, site_lo = 0, mut_lo = 0;
    HostTotals *dt = (HostTotals *) ctx->dev_totals.p;
    HostTotals *ht = ctx->host;

    if (async_begin && (replay || p->model_kind == MSPB_MODEL_SLIM)) {
        // the replay harness and the SLiM model need the host between the stages: mspb_generate_begin
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    CUDA_OK(cudaEventRecord(ctx->ev[0], st));
    if (async_begin) {
        CUDA_OK(cudaEventRecord(ctx->ev_async0, st));
    }
    RateMapView map;
    ModelView model;
    TRY(upload_params(ctx, p, &map, &model));
    CUDA_OK(cudaMemsetAsync(dt, 0, sizeof(HostTotals), st));
    TreeSeqView ts = make_ts_view(ctx);
    CallView call;
    call.start_time = p->start_time;
    call.end_time = p->end_time;
    call.slice_left = p->slice_left;
    call.slice_right = in_slice ? p->slice_right : ctx->L;
    call.discrete = discrete ? 1 : 0;
    call.sliced = sliced ? 1 : 0;
    call.clipped = clipped ? 1 : 0;
    call.stream = clipped ? p->slice_stream : 0u;
    call.seed = p->seed;
    call.key_mode = (discrete && ctx->L <= 4294967296.0 && !(keep && ctx->kept_fractional)) ? KEY32_POS : KEY64_POS;
    call.edge_list = nullptr;
    call.num_work = ctx->E;
    if (clipped && ctx->E > 0) {
        if (!ctx->slice_cache_valid || ctx->slice_cached_left != call.slice_left
            || ctx->slice_cached_right != call.slice_right) {
            EdgeInSlice f;
            f.left = ts.edge_left;
            f.right = ts.edge_right;
            f.slice_left = call.slice_left;
            f.slice_right = call.slice_right;
            TRY(ctx->goff.ensure(ctx->E + 2));
            TRY(ctx->slice_edges.ensure(ctx->E + 1));
            TRY(ctx->slice_scan_tiles.ensure(scan_num_tiles(ctx->E) + 2));
            uint32_t *total = ctx->slice_scan_tiles.p + scan_num_tiles(ctx->E) + 1;
            device_exclusive_scan<uint32_t, EdgeInSlice>(
                ctx->log, ctx->E, f, ctx->goff.p, ctx->slice_scan_tiles.p, total, st);
            MSPB_LAUNCH(ctx->log, "slice_edges_scatter",
                slice_edges_scatter_kernel<<<grid_for(ctx->E, 256), 256, 0, st>>>(
                    f, ctx->E, ctx->goff.p, ctx->slice_edges.p));
            uint32_t n_in = 0;
            CUDA_OK(cudaMemcpyAsync(&n_in, total, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
            CUDA_OK(cudaStreamSynchronize(st));
            ctx->slice_num_edges = n_in;
            ctx->slice_cached_left = call.slice_left;
            ctx->slice_cached_right = call.slice_right;
            ctx->slice_cache_valid = true;
        }
        call.edge_list = ctx->slice_edges.p;
        call.num_work = ctx->slice_num_edges;
    }
    if (in_slice && S0 + M0 > 0) {
        // which kept rows this slice owns (read back with the K1 totals)
        const uint64_t mx = S0 > M0 ? S0 : M0;
        MSPB_LAUNCH(ctx->log, "kept_sorted_check",
            kept_sorted_check_kernel<<<grid_for(mx, 256), 256, 0, st>>>(
                S0, ctx->k_site_pos.p, M0, ctx->k_mut_site.p, &dt->kept_unsorted));
        MSPB_LAUNCH(ctx->log, "kept_slice",
            kept_slice_kernel<<<1, 32, 0, st>>>(S0, ctx->k_site_pos.p, M0, ctx->k_mut_site.p, call.slice_left,
                call.slice_right, dt->kslice));
    }

    // ---- K1: how many new mutations, and where they go
    // (the events between the phases are recorded while kernel timing is on only: nine records per overlay are
    // not free on a 1.7 ms call)
    const bool phase_events = ctx->log.timing;
    if (phase_events) {
        CUDA_OK(cudaEventRecord(ctx->ev[1], st));
    }
    uint64_t Mnew = 0, num_pairs = 0;
    uint64_t spec_rec = 0, spec_key = 0; // generations of the buffers of a K2 launched ahead of the K1 read-back
    size_t spec_cap = 0;
    uint64_t spec_sparse_n = 0; // > 0: the early launch was the sparse kernel and covered this many slots
    ReplayPairs pairs;
    memset(&pairs, 0, sizeof(pairs));
    bool serial_placed = false; // replay harness, continuous genome: replay_resolve wrote the records itself
    if (!replay) {
        TRY(ctx->goff.ensure(ctx->E + 2));
        if (sliced) {
            TRY(ctx->loff.ensure(ctx->E + 2));
        }
        if (call.num_work > 0) {
            const uint32_t tiles = (uint32_t) ((call.num_work + K1_TILE - 1) / K1_TILE);
            TRY(ctx->k1_tiles.ensure(tiles + 1));
            MSPB_LAUNCH(ctx->log, "k1_count",
                k1_count_kernel<<<tiles, SCAN_THREADS, 0, st>>>(
                    ts, map, call, ctx->goff.p, sliced ? ctx->loff.p : nullptr, ctx->k1_tiles.p, &dt->err));
            MSPB_LAUNCH(ctx->log, "k1_spine",
                k1_spine_kernel<<<1, 1024, 0, st>>>(ctx->k1_tiles.p, tiles, &dt->k1_total, &dt->err));
        }
        TRY(totals_to_host(ctx));
        CUDA_OK(cudaEventRecord(ctx->ev_k1, st));
        // K2 does not need the total to be launched (its grid follows the edges): when the
        // record and key buffers of earlier calls are there, it goes out now and the host's
        // wait for the total overlaps it.  The kernel checks the total against the capacity.
        if (!sliced && S0 + M0 == 0 && call.num_work > 0) {
            spec_rec = ctx->rec.gen;
            spec_key = call.key_mode == KEY32_POS ? ctx->k32_0.gen : ctx->keys0.gen;
            const size_t key_cap = call.key_mode == KEY32_POS ? ctx->k32_0.cap : ctx->keys0.cap;
            spec_cap = ctx->rec.cap < key_cap ? ctx->rec.cap : key_cap;
            if (ctx->rec.p != nullptr && spec_cap > 1) {
                spec_cap -= 1;
                KeyView skeys;
                skeys.mode = call.key_mode;
                skeys.k32 = ctx->k32_0.p;
                skeys.k64 = ctx->keys0.p;
                // (the density of this call is not known yet: the previous call over the same edges decides)
                const bool sparse = ctx->k2_force == 2
                                    || (ctx->k2_force == 0 && ctx->last_num_work == call.num_work
                                        && k2_sparse_pays(ctx->last_new, call.num_work));
                if (sparse) {
                    spec_sparse_n = ctx->last_new + ctx->last_new / 4 + 4096;
                    if (ctx->k2_force == 2 && ctx->last_num_work != call.num_work) {
                        spec_sparse_n = 4096; // (forced without a previous call: the second launch does the rest)
                    }
                    spec_sparse_n = spec_sparse_n < spec_cap ? spec_sparse_n : spec_cap;
                    MSPB_LAUNCH(ctx->log, "k2_place_sparse",
                        k2_place_sparse_kernel<<<grid_for(spec_sparse_n, 256), 256, 0, st>>>(ts, map, call, ctx->goff.p,
                            ctx->k1_tiles.p, 0u, ctx->rec.p, skeys, &dt->k1_total, spec_cap, &dt->err, 0));
                } else {
                    MSPB_LAUNCH(ctx->log, "k2_place",
                        k2_place_kernel<<<grid_for(call.num_work, K2_EDGES), K2_EDGES, 0, st>>>(ts, map, call,
                            ctx->goff.p, nullptr, ctx->k1_tiles.p, 0u, ctx->rec.p, skeys, &dt->k1_total, spec_cap,
                            &dt->err));
                }
            } else {
                spec_cap = 0;
            }
        }
        CUDA_OK(cudaEventSynchronize(ctx->ev_k1)); // the copy, not the kernel launched after it
        if (ht->err != 0) {
            // more new mutations than 31-bit ids can name (K1 / the spine refused the count)
            CUDA_OK(cudaStreamSynchronize(st));
            ctx->log.collect();
            return ht->err;
        }
        Mnew = sliced ? (ht->k1_total >> 32) : (ht->k1_total & 0xffffffffull);
    } else {
        const uint64_t cap = p->replay_len / 2 + 1;
        TRY(upload(ctx->replay, p->replay_stream, p->replay_len, st));
        TRY(ctx->rp_edge.ensure(cap));
        TRY(ctx->rp_interval.ensure(cap));
        TRY(ctx->rp_count.ensure(cap));
        TRY(ctx->rp_out_off.ensure(cap));
        TRY(ctx->rp_stream_off.ensure(cap));
        pairs.edge = ctx->rp_edge.p;
        pairs.interval = ctx->rp_interval.p;
        pairs.count = ctx->rp_count.p;
        pairs.out_off = ctx->rp_out_off.p;
        pairs.stream_off = ctx->rp_stream_off.p;
        pairs.capacity = cap;
        // Continuous genome: the walker places the mutations itself, because a position that is
        // taken is drawn again from the stream (lib/mutgen.c:1127-1134).  Records and keys are
        // sized from the stream length here (a mutation takes at least two uniforms).
        ReplaySerial sp;
        memset(&sp, 0, sizeof(sp));
        if (!discrete) {
            const uint64_t bound = S0 + M0 + cap;
            if (bound > RS_MAX_N) {
                return MSPB_ERR_TOO_MANY_ROWS;
            }
            uint64_t hcap = 1024;
            while (hcap < 2 * (S0 + cap) + 2) {
                hcap <<= 1;
            }
            TRY(ctx->rec.ensure(bound + 1));
            TRY(ensure_sort_buffers(ctx, bound, false, true));
            TRY(ctx->replay_set.ensure(hcap));
            CUDA_OK(cudaMemsetAsync(ctx->replay_set.p, 0, hcap * sizeof(uint64_t), st));
            sp.set = ctx->replay_set.p;
            sp.mask = hcap - 1;
            sp.kept_position = ctx->k_site_pos.p;
            sp.num_kept_sites = S0;
            sp.rec = ctx->rec.p;
            sp.k64 = ctx->keys0.p;
            sp.base = (uint32_t) (S0 + M0);
            sp.max_new = cap;
            serial_placed = true;
        }
        MSPB_LAUNCH(ctx->log, "replay_resolve",
            replay_resolve_kernel<<<1, 32, 0, st>>>(ts, map, call, ctx->replay.p, p->replay_len, pairs, sp,
                &dt->replay, &dt->err));
        TRY(totals_to_host(ctx));
        CUDA_OK(cudaStreamSynchronize(st));
        if (ht->err != 0) {
            return ht->err;
        }
        if (ht->replay.exhausted) {
            return MSPB_ERR_REPLAY_STREAM_EXHAUSTED;
        }
        Mnew = ht->replay.num_mutations;
        num_pairs = ht->replay.num_pairs;
    }
    if (in_slice && S0 + M0 > 0) {
        if (ht->kept_unsorted != 0) {
            return MSPB_ERR_UNSORTED_SITES;
        }
        site_lo = ht->kslice[0];
        S0 = ht->kslice[1] - ht->kslice[0];
        mut_lo = ht->kslice[2];
        M0 = ht->kslice[3] - ht->kslice[2];
    }
    if (keep && p->model_kind == MSPB_MODEL_MATRIX) {
        TRY(ctx->k_site_allele.ensure(S0 + 1));
        TRY(ctx->k_mut_allele.ensure(M0 + 1));
        if (S0 > 0) {
            MSPB_LAUNCH(ctx->log, "kept_allele_index",
                kept_allele_index_kernel<<<grid_for(S0, 256), 256, 0, st>>>(
                    model, S0, ctx->k_site_anc.p, ctx->k_site_anc_off.p + site_lo, ctx->k_site_allele.p));
        }
        if (M0 > 0) {
            MSPB_LAUNCH(ctx->log, "kept_allele_index",
                kept_allele_index_kernel<<<grid_for(M0, 256), 256, 0, st>>>(
                    model, M0, ctx->k_mut_der.p, ctx->k_mut_der_off.p + mut_lo, ctx->k_mut_allele.p));
        }
    }
    KeptView kept = make_kept_view(ctx, keep, site_lo, S0, mut_lo, M0);
    const uint64_t R = S0 + M0 + Mnew;
    if (R > RS_MAX_N) {
        return MSPB_ERR_TOO_MANY_ROWS;
    }

    // ---- buffers: rows are bounded before the sort (M is exact, S <= S0 + Mnew)
    if (phase_events) {
        CUDA_OK(cudaEventRecord(ctx->ev[2], st));
    }
    const uint64_t M = M0 + Mnew, Scap = S0 + Mnew;
    // (two-phase calls learn their allele-number base only later: size for any uint64)
    const uint64_t maxlen
        = (defer && p->model_kind == MSPB_MODEL_INFINITE_ALLELES) ? 20 : max_new_state_len(p, Mnew);
    const uint64_t der_cap = (keep ? ctx->kept_der_bytes : 0) + Mnew * maxlen;
    const uint64_t anc_cap = (keep ? ctx->kept_anc_bytes : 0) + Mnew * maxlen;
    const bool slim = p->model_kind == MSPB_MODEL_SLIM;
    // (a slice of a two-phase call learns the first id it may use from the other ranks, and the SLiM
    // model's ids feed its state lengths: K4 and K5 then run in mspb_generate_end)
    const bool slim_two_phase = slim && defer;
    ctx->pending.defer_resolve = slim_two_phase;
    if (slim && Mnew > 0 && p->start_allele + Mnew - 1 >= (uint64_t) INT64_MAX) {
        return MSPB_ERR_MUTATION_ID_OVERFLOW; /* lib/mutgen.c:386-389 */
    }
    const uint64_t smd_cap = keep ? ctx->kept_smd_bytes : 0,
                   mmd_cap = (keep ? ctx->kept_mmd_bytes : 0) + (slim ? Mnew * SLIM_MD_SIZE : 0);
    TRY(ctx->rec.ensure(R + 1));
    TRY(ensure_sort_buffers(ctx, R, call.key_mode == KEY32_POS, call.key_mode != KEY32_POS));
    TRY(ctx->site_scan.ensure(R + 1));
    TRY(ctx->multi_list.ensure(R / 2 + 2));
    TRY(ctx->parent.ensure(R + 1));
    TRY(ctx->state.ensure(R + 1));
    TRY(ctx->site_state.ensure(R + 1));
    if (S0 + M0 > 0) {
        TRY(ctx->mut_scan.ensure(R + 1));
        TRY(ctx->draw_scan.ensure(R + 1));
    }
    TRY(ctx->o_site_pos.ensure(Scap + 1));
    TRY(ctx->o_site_anc.ensure(anc_cap + 1));
    TRY(ctx->o_site_anc_off.ensure(Scap + 2));
    TRY(ctx->o_site_md.ensure(smd_cap + 1));
    TRY(ctx->o_site_md_off.ensure(Scap + 2));
    TRY(ctx->o_mut_site.ensure(M + 1));
    TRY(ctx->o_mut_node.ensure(M + 1));
    TRY(ctx->o_mut_parent.ensure(M + 1));
    TRY(ctx->o_mut_time.ensure(M + 1));
    TRY(ctx->o_mut_der.ensure(der_cap + 1));
    TRY(ctx->o_mut_der_off.ensure(M + 2));
    TRY(ctx->o_mut_md.ensure(mmd_cap + 1));
    TRY(ctx->o_mut_md_off.ensure(M + 2));
    const bool off32 = p->offsets32 != 0;
    if (off32 && (der_cap >= 0xffffffffull || anc_cap >= 0xffffffffull || smd_cap >= 0xffffffffull
                     || mmd_cap >= 0xffffffffull)) {
        return MSPB_ERR_OFFSET_OVERFLOW;
    }
    const uint64_t ow = off32 ? 4 : 8;
    OutView out;
    out.off32 = off32 ? 1 : 0;
    out.write_md_off = (smd_cap + mmd_cap) > 0 ? 1 : 0;
    if (out.write_md_off) {
        ctx->md_zero_site_bytes = ctx->md_zero_mut_bytes = 0;
    } else {
        // metadata offsets are all zero: clear the columns only if they are not zero already
        if (ctx->md_zero_site_ptr != ctx->o_site_md_off.p || ctx->md_zero_site_bytes < (Scap + 2) * ow) {
            CUDA_OK(cudaMemsetAsync(ctx->o_site_md_off.p, 0, ctx->o_site_md_off.cap * sizeof(uint64_t), st));
            ctx->md_zero_site_ptr = ctx->o_site_md_off.p;
            ctx->md_zero_site_bytes = ctx->o_site_md_off.cap * sizeof(uint64_t);
        }
        if (ctx->md_zero_mut_ptr != ctx->o_mut_md_off.p || ctx->md_zero_mut_bytes < (M + 2) * ow) {
            CUDA_OK(cudaMemsetAsync(ctx->o_mut_md_off.p, 0, ctx->o_mut_md_off.cap * sizeof(uint64_t), st));
            ctx->md_zero_mut_ptr = ctx->o_mut_md_off.p;
            ctx->md_zero_mut_bytes = ctx->o_mut_md_off.cap * sizeof(uint64_t);
        }
    }
    out.site_position = ctx->o_site_pos.p;
    out.site_anc = ctx->o_site_anc.p;
    out.site_anc_off = ctx->o_site_anc_off.p;
    out.site_md = ctx->o_site_md.p;
    out.site_md_off = ctx->o_site_md_off.p;
    out.mut_site = ctx->o_mut_site.p;
    out.mut_node = ctx->o_mut_node.p;
    out.mut_parent = ctx->o_mut_parent.p;
    out.mut_time = ctx->o_mut_time.p;
    out.mut_der = ctx->o_mut_der.p;
    out.mut_der_off = ctx->o_mut_der_off.p;
    out.mut_md = ctx->o_mut_md.p;
    out.mut_md_off = ctx->o_mut_md_off.p;
    out.site_base = 0;
    out.mut_base = 0;
    out.dev_bases = nullptr;
    out.dev_byte_bases = nullptr;
    out.dry_run = 0;
    out.chain_free = nullptr;
    // matrix models whose alleles all have one (non-zero) length: state bytes = rows x length
    ctx->pending.fixed_len = 0;
    if (p->model_kind == MSPB_MODEL_MATRIX && p->num_alleles > 0) {
        const uint64_t len0 = p->allele_offset[1] - p->allele_offset[0];
        bool same = len0 > 0 && len0 < 256;
        for (uint64_t a = 1; a < p->num_alleles && same; a++) {
            same = p->allele_offset[a + 1] - p->allele_offset[a] == len0;
        }
        ctx->pending.fixed_len = same ? (uint32_t) len0 : 0u;
    }
    if (slim) {
        ctx->pending.defer = true; // K5 runs below, after its sizing pass
    }

    // ---- K2: records (kept markers, kept mutations, new mutations) and their sort keys
    KeyView keys;
    keys.mode = call.key_mode;
    keys.k32 = ctx->k32_0.p;
    keys.k64 = ctx->keys0.p;
    if (S0 + M0 > 0) {
        MSPB_LAUNCH(ctx->log, "fill_kept",
            fill_kept_kernel<<<grid_for(S0 + M0, 256), 256, 0, st>>>(kept, call, ctx->rec.p, keys, &dt->err));
    }
    if (Mnew > 0) {
        const bool placed = spec_cap > 0 && Mnew <= spec_cap && spec_rec == ctx->rec.gen
                            && spec_key == (call.key_mode == KEY32_POS ? ctx->k32_0.gen : ctx->keys0.gen);
        if (placed || serial_placed) {
            // the early launch had room and the buffers did not move (or the replay walker placed
            // the mutations itself): the records are in place
            if (placed && spec_sparse_n > 0 && Mnew > spec_sparse_n) {
                // ... up to the slots the sparse launch was sized for (from the previous call's count)
                MSPB_LAUNCH(ctx->log, "k2_place_sparse",
                    k2_place_sparse_kernel<<<grid_for(Mnew - spec_sparse_n, 256), 256, 0, st>>>(ts, map, call,
                        ctx->goff.p, ctx->k1_tiles.p, 0u, ctx->rec.p, keys, nullptr, 0, nullptr, spec_sparse_n));
            }
        } else if (!replay) {
            const bool sparse = !sliced
                                && (ctx->k2_force == 2 || (ctx->k2_force == 0 && k2_sparse_pays(Mnew, call.num_work)));
            if (sparse) {
                MSPB_LAUNCH(ctx->log, "k2_place_sparse",
                    k2_place_sparse_kernel<<<grid_for(Mnew, 256), 256, 0, st>>>(ts, map, call, ctx->goff.p,
                        ctx->k1_tiles.p, (uint32_t) (S0 + M0), ctx->rec.p, keys, nullptr, 0, nullptr, 0));
            } else {
                MSPB_LAUNCH(ctx->log, "k2_place",
                    k2_place_kernel<<<grid_for(call.num_work, K2_EDGES), K2_EDGES, 0, st>>>(ts, map, call, ctx->goff.p,
                        sliced ? ctx->loff.p : nullptr, ctx->k1_tiles.p, (uint32_t) (S0 + M0), ctx->rec.p, keys,
                        nullptr, 0, nullptr));
            }
        } else {
            MSPB_LAUNCH(ctx->log, "replay_place",
                replay_place_kernel<<<grid_for(num_pairs, 128), 128, 0, st>>>(ts, map, call, ctx->replay.p, pairs,
                    num_pairs, (uint32_t) (S0 + M0), ctx->rec.p, keys, &dt->err));
        }
    }
    if (!replay) {
        ctx->last_num_work = call.num_work;
        ctx->last_new = Mnew;
    }
    if (R == 0) {
        // no sites and no mutations: only the leading zero offsets exist
        CUDA_OK(cudaMemsetAsync(out.site_anc_off, 0, sizeof(uint64_t), st));
        CUDA_OK(cudaMemsetAsync(out.site_md_off, 0, sizeof(uint64_t), st));
        CUDA_OK(cudaMemsetAsync(out.mut_der_off, 0, sizeof(uint64_t), st));
        CUDA_OK(cudaMemsetAsync(out.mut_md_off, 0, sizeof(uint64_t), st));
        ctx->md_zero_site_bytes = ctx->md_zero_mut_bytes = 0;
        for (int e = 3; e <= 7 && phase_events; e++) {
            CUDA_OK(cudaEventRecord(ctx->ev[e], st));
        }
    } else {
        TRY(run_sorted_stages(ctx, p, ts, kept, model, call, R, S0, M0, out, true));
    }
    if (async_begin) {
        // mspb_generate_begin_device: K3 .. K4 are enqueued and nothing is awaited.  What the host would
        // look at here (an error, a dense genome's second sort, collisions to draw again) travels with the
        // exchanged counts instead (xchg_pack_counts_kernel) and mspb_generate_finish_device asks for the
        // call to be made again through mspb_generate_begin when any rank has raised it.
        CUDA_OK(cudaEventRecord(ctx->ev_async1, st));
        mspb_counts_t &ca = ctx->counts;
        memset(&ca, 0, sizeof(ca));
        ca.num_mutations = M;
        ca.num_new_mutations = Mnew;
        ca.offset_bytes = ow;
        ctx->generated = false;
        ctx->pending.valid = true;
        ctx->pending.async = true;
        ctx->pending.early = S0 + M0 == 0 && R > 0
                             && (p->model_kind == MSPB_MODEL_INFINITE_ALLELES
                                 || (p->model_kind == MSPB_MODEL_MATRIX && ctx->pending.fixed_len > 0));
        ctx->pending.key32 = call.key_mode == KEY32_POS;
        ctx->pending.S0 = S0;
        ctx->pending.Mnew = Mnew;
        ctx->pending.M = M;
        ctx->pending.start_allele = p->start_allele;
        ctx->pending.model_kind = p->model_kind;
        if (counts_out != nullptr) {
            *counts_out = ca;
        }
        return 0;
    }
    TRY(totals_to_host(ctx));
    CUDA_OK(cudaStreamSynchronize(st));
    CUDA_OK(cudaGetLastError());
    if (R > 0 && ht->err == 0 && ht->dense != 0 && call.key_mode == KEY32_POS) {
        // dense genome: many mutations per site.  Sort again with time in the key so that
        // the per-site ordering step starts from a nearly sorted run.
        TRY(ensure_sort_buffers(ctx, R, false, true));
        RadixSortBuffers<uint64_t> sb = sort_buffers64(ctx);
        // the sorted values live in vals0 or vals1; write the new pairs into the other one
        uint32_t *val_in = ctx->last_sv.val;
        uint32_t *val_tmp = val_in == ctx->vals0.p ? ctx->vals1.p : ctx->vals0.p;
        MSPB_LAUNCH(ctx->log, "rekey_pos_time",
            rekey_pos_time_kernel<<<grid_for(R, 256), 256, 0, st>>>(ctx->last_sv, sb.keys[0], val_tmp));
        if (val_tmp != sb.vals[0]) {
            CUDA_OK(cudaMemcpyAsync(sb.vals[0], val_tmp, R * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
        }
        call.key_mode = KEY64_POS_TIME;
        CUDA_OK(cudaMemsetAsync(&dt->p1, 0, sizeof(Counts3), st));
        CUDA_OK(cudaMemsetAsync(&dt->k4_next, 0, sizeof(uint32_t), st));
        CUDA_OK(cudaMemsetAsync(&dt->collisions, 0, sizeof(uint32_t), st));
        TRY(run_sorted_stages(ctx, p, ts, kept, model, call, R, S0, M0, out, false));
        TRY(totals_to_host(ctx));
    }
    // continuous genome, native stream: new mutations on occupied positions draw again
    // (lib/mutgen.c:1127-1134) until every position is unique.  (The replay harness and
    // invariant slices report the collision instead: the former would need the reference's
    // exact stream order, the latter could move a record out of the slice.)
    for (uint32_t round = 1; R > 0 && ht->err == 0 && ht->collisions > 0 && !replay && !sliced
                             && call.key_mode == KEY64_POS && round <= 32;
         round++) {
        RadixSortBuffers<uint64_t> sb = sort_buffers64(ctx);
        uint64_t *key_in = ctx->last_sv.key.k64;
        uint64_t *key_tmp = key_in == ctx->keys0.p ? ctx->keys1.p : ctx->keys0.p;
        uint32_t *val_in = ctx->last_sv.val;
        uint32_t *val_tmp = val_in == ctx->vals0.p ? ctx->vals1.p : ctx->vals0.p;
        MSPB_LAUNCH(ctx->log, "redraw_positions",
            redraw_positions_kernel<<<grid_for(R, 256), 256, 0, st>>>(ctx->last_sv, ts, map, call, ctx->goff.p,
                ctx->k1_tiles.p, round, key_tmp, val_tmp));
        if (key_tmp != sb.keys[0]) {
            CUDA_OK(cudaMemcpyAsync(sb.keys[0], key_tmp, R * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
        }
        if (val_tmp != sb.vals[0]) {
            CUDA_OK(cudaMemcpyAsync(sb.vals[0], val_tmp, R * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
        }
        CUDA_OK(cudaMemsetAsync(&dt->p1, 0, sizeof(Counts3), st));
        CUDA_OK(cudaMemsetAsync(&dt->k4_next, 0, sizeof(uint32_t), st));
        CUDA_OK(cudaMemsetAsync(&dt->collisions, 0, sizeof(uint32_t), st));
        TRY(run_sorted_stages(ctx, p, ts, kept, model, call, R, S0, M0, out, false));
        TRY(totals_to_host(ctx));
        CUDA_OK(cudaStreamSynchronize(st));
    }
    CUDA_OK(cudaEventRecord(ctx->ev[8], st));
    CUDA_OK(cudaStreamSynchronize(st));
    CUDA_OK(cudaGetLastError());
    ctx->log.collect();
    if (ht->err != 0) {
        return ht->err;
    }
    if (ht->collisions > 0) {
        return MSPB_ERR_REPLAY_REDRAW;
    }
    if (slim && R > 0 && !slim_two_phase) {
        TRY(slim_size_and_emit(ctx));
        CUDA_OK(cudaEventRecord(ctx->ev[8], st));
        CUDA_OK(cudaStreamSynchronize(st));
        CUDA_OK(cudaGetLastError());
        ctx->log.collect();
        if (ht->err != 0) {
            return ht->err;
        }
    }
    if (slim && !slim_two_phase) {
        defer = false;
    }
    const uint64_t draws = ht->p1.d;
    if (replay && p->model_kind == MSPB_MODEL_MATRIX && ht->replay.draws + draws > p->replay_len) {
        return MSPB_ERR_REPLAY_STREAM_EXHAUSTED;
    }
    const uint64_t S = ht->p1.s;
    float total = 0;
    for (int i = 0; i < MSPB_NUM_PHASES; i++) {
        ctx->phase_ms[i] = 0.0f;
        if (phase_events) {
            cudaEventElapsedTime(&ctx->phase_ms[i], ctx->ev[i], ctx->ev[i + 1]);
        }
    }
    cudaEventElapsedTime(&total, ctx->ev[0], ctx->ev[8]);
    mspb_counts_t &c = ctx->counts;
    memset(&c, 0, sizeof(c));
    c.num_sites = S;
    c.num_mutations = M;
    c.num_new_mutations = Mnew;
    c.ancestral_state_bytes = ht->bytes.anc;
    c.site_metadata_bytes = ht->bytes.smd;
    c.derived_state_bytes = ht->bytes.der;
    c.mutation_metadata_bytes = ht->bytes.mmd;
    c.num_new_sites = S - S0;
    c.next_allele = p->start_allele + (p->model_kind != MSPB_MODEL_MATRIX ? draws : 0);
    c.draws = replay ? ht->replay.draws + (p->model_kind == MSPB_MODEL_MATRIX ? draws : 0) : 0;
    c.device_ms = total;
    c.offset_bytes = ow;
    ctx->generated = !defer;
    if (defer) {
        ctx->pending.valid = true;
        ctx->pending.S0 = S0;
        ctx->pending.Mnew = Mnew;
        ctx->pending.M = M;
        ctx->pending.start_allele = p->start_allele;
        ctx->pending.model_kind = p->model_kind;
    }
    if (counts_out != nullptr) {
        *counts_out = c;
    }
    return 0;
}

extern "C" int mspb_ctx_force_placement(mspb_ctx_t *ctx, int mode)
{
    if (ctx == nullptr || mode < 0 || mode > 2) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    ctx->k2_force = mode;
    return 0;
}

extern "C" int mspb_generate(mspb_ctx_t *ctx, const mspb_params_t *p, mspb_counts_t *counts_out)
{
    return generate_impl(ctx, p, counts_out, false);
}

extern "C" int mspb_generate_begin(mspb_ctx_t *ctx, const mspb_params_t *p, mspb_counts_t *counts_out)
{
    return generate_impl(ctx, p, counts_out, true);
}

extern "C" int mspb_generate_begin_device(mspb_ctx_t *ctx, const mspb_params_t *p)
{
    return generate_impl(ctx, p, nullptr, true, true);
}

extern "C" int mspb_generate_end(mspb_ctx_t *ctx, uint64_t site_base, uint64_t mutation_base, uint64_t allele_base,
    mspb_counts_t *counts_out)
{
    if (ctx == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (!ctx->pending.valid) {
        return MSPB_ERR_NOT_GENERATED;
    }
    if (ctx->pending.async) {
        return MSPB_ERR_BAD_PARAM_VALUE; // the host does not hold this call's counts: mspb_generate_end_device
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    HostTotals *dt = (HostTotals *) ctx->dev_totals.p;
    HostTotals *ht = ctx->host;
    mspb_counts_t &c = ctx->counts;
    if (site_base + c.num_sites > 0x7fffffffull || mutation_base + c.num_mutations > 0x7fffffffull) {
        return MSPB_ERR_TOO_MANY_ROWS;
    }
    ctx->pending.valid = false;
    ctx->pending.out.site_base = (uint32_t) site_base;
    ctx->pending.out.mut_base = (uint32_t) mutation_base;
    ctx->pending.model.start_allele = allele_base;
    cudaEvent_t e0 = ctx->ev[0], e1 = ctx->ev[1];
    CUDA_OK(cudaEventRecord(e0, st));
    if (ctx->pending.defer_resolve) {
        // SLiM: the ids of this slice start at allele_base (lib/mutgen.c:386-389 for the overflow rule)
        if (ctx->pending.Mnew > 0 && allele_base + ctx->pending.Mnew - 1 >= (uint64_t) INT64_MAX) {
            return MSPB_ERR_MUTATION_ID_OVERFLOW;
        }
        if (ctx->pending.R > 0) {
            CUDA_OK(cudaMemsetAsync(&dt->k4_next, 0, sizeof(uint32_t), st));
            TRY(launch_resolve(ctx));
            TRY(slim_size_and_emit(ctx));
        }
    } else if (ctx->pending.R > 0) {
        TRY(launch_emit(ctx));
    }
    TRY(totals_to_host(ctx));
    CUDA_OK(cudaEventRecord(e1, st));
    CUDA_OK(cudaStreamSynchronize(st));
    CUDA_OK(cudaGetLastError());
    ctx->log.collect();
    if (ht->err != 0) {
        return ht->err;
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    c.ancestral_state_bytes = ht->bytes.anc;
    c.site_metadata_bytes = ht->bytes.smd;
    c.derived_state_bytes = ht->bytes.der;
    c.mutation_metadata_bytes = ht->bytes.mmd;
    if (ctx->pending.model_kind != MSPB_MODEL_MATRIX) {
        c.next_allele = allele_base + (c.next_allele - ctx->pending.start_allele); // allele number / SLiM id counter
    }
    c.device_ms += ms;
    ctx->generated = true;
    if (counts_out != nullptr) {
        *counts_out = c;
    }
    return 0;
}

// adds a constant to every entry of an offset column (uint32 or uint64)
namespace {
__global__ void shift_offsets_kernel(void *col, uint64_t n, uint64_t delta, int off32)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        if (off32) {
            static_cast<uint32_t *>(col)[i] += (uint32_t) delta;
        } else {
            static_cast<uint64_t *>(col)[i] += delta;
        }
    }
}
} // namespace

extern "C" int mspb_shift_offsets(mspb_ctx_t *ctx, uint64_t ancestral_state_base, uint64_t site_metadata_base,
    uint64_t derived_state_base, uint64_t mutation_metadata_base)
{
    if (ctx == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (!ctx->generated) {
        return MSPB_ERR_NOT_GENERATED;
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const mspb_counts_t &c = ctx->counts;
    const int off32 = c.offset_bytes == 4;
    const uint64_t S1 = c.num_sites + 1, M1 = c.num_mutations + 1;
    if (off32
        && (ancestral_state_base + c.ancestral_state_bytes > 0xffffffffull
               || site_metadata_base + c.site_metadata_bytes > 0xffffffffull
               || derived_state_base + c.derived_state_bytes > 0xffffffffull
               || mutation_metadata_base + c.mutation_metadata_bytes > 0xffffffffull)) {
        return MSPB_ERR_OFFSET_OVERFLOW;
    }
    if (ancestral_state_base != 0) {
        MSPB_LAUNCH(ctx->log, "shift_offsets",
            shift_offsets_kernel<<<grid_for(S1, 256), 256, 0, st>>>(ctx->o_site_anc_off.p, S1, ancestral_state_base, off32));
    }
    if (site_metadata_base != 0) {
        MSPB_LAUNCH(ctx->log, "shift_offsets",
            shift_offsets_kernel<<<grid_for(S1, 256), 256, 0, st>>>(ctx->o_site_md_off.p, S1, site_metadata_base, off32));
        ctx->md_zero_site_bytes = 0;
    }
    if (derived_state_base != 0) {
        MSPB_LAUNCH(ctx->log, "shift_offsets",
            shift_offsets_kernel<<<grid_for(M1, 256), 256, 0, st>>>(ctx->o_mut_der_off.p, M1, derived_state_base, off32));
    }
    if (mutation_metadata_base != 0) {
        MSPB_LAUNCH(ctx->log, "shift_offsets",
            shift_offsets_kernel<<<grid_for(M1, 256), 256, 0, st>>>(ctx->o_mut_md_off.p, M1, mutation_metadata_base, off32));
        ctx->md_zero_mut_bytes = 0;
    }
    CUDA_OK(cudaStreamSynchronize(st));
    CUDA_OK(cudaGetLastError());
    ctx->log.collect();
    return 0;
}

// ---------------------------------------------------------------- device-side exchange
// A sharded call whose counts never visit the host between its stages: the caller
// all-gathers send -> recv on the context's stream (NCCL through torch.distributed, or any
// device copy), K5 and the offset shift read their bases from device memory.
namespace {
constexpr uint64_t XCHG_WORDS = MSPB_EXCHANGE_WORDS, XCHG_MAX_WORLD = MSPB_EXCHANGE_MAX_WORLD;
constexpr uint64_t XCHG_RECV = XCHG_WORDS, XCHG_B1 = XCHG_RECV + XCHG_MAX_WORLD * XCHG_WORDS,
                   XCHG_B2 = XCHG_B1 + 2 * XCHG_WORDS, XCHG_TOTAL = XCHG_B2 + 2 * XCHG_WORDS;

int ensure_exchange(mspb_ctx *ctx)
{
    if (ctx->xchg.p == nullptr) {
        TRY(ctx->xchg.ensure(XCHG_TOTAL));
        CUDA_OK(cudaMemsetAsync(ctx->xchg.p, 0, XCHG_TOTAL * sizeof(uint64_t), ctx->stream));
    }
    if (ctx->xchg_host == nullptr) {
        CUDA_OK(cudaMallocHost((void **) &ctx->xchg_host, 5 * XCHG_WORDS * sizeof(uint64_t)));
    }
    return 0;
}

// bases[j] = sum over ranks below `rank` of recv[r][j]; bases[WORDS + j] = sum over all ranks
__global__ void xchg_bases_kernel(const uint64_t *recv, uint32_t rank, uint32_t world, uint64_t *bases)
{
    const uint32_t j = threadIdx.x;
    if (j < XCHG_WORDS) {
        uint64_t ex = 0, tot = 0;
        for (uint32_t r = 0; r < world; r++) {
            const uint64_t v = recv[r * XCHG_WORDS + j];
            ex += r < rank ? v : 0ull;
            tot += v;
        }
        bases[j] = ex;
        bases[XCHG_WORDS + j] = tot;
    }
}

// mspb_generate_begin_device: this rank's counts straight from the device totals.  send[4] asks every rank to
// make the call again with the host in the loop (an error, a dense genome's second sort, positions to draw again).
__global__ void xchg_pack_counts_kernel(const HostTotals *dt, uint64_t num_mutations, uint64_t num_new, int numbered,
    int key32, uint64_t *send)
{
    send[0] = dt->p1.s;
    send[1] = num_mutations;
    send[2] = numbered ? dt->p1.d : 0ull;
    send[3] = num_new;
    send[4] = (dt->err != 0 || (dt->dense != 0u && key32) || dt->collisions != 0u) ? 1ull : 0ull;
    send[5] = send[6] = send[7] = 0;
}

// early exchange: this rank's byte counts from the ChainFree table (before any byte is written).  send[4]: the
// closed form does not apply (allele numbers wrap around 2^64, or the genome is dense): back to the host-paced call
__global__ void xchg_pack_bytes_cf_kernel(const ChainFree *cf, const Counts3 *p1, uint64_t n, uint64_t *send)
{
    uint64_t der = 0, anc = 0;
    if (cf->enabled != 0u) {
        cf->prefix(n, p1->s, &der, &anc);
    }
    send[0] = anc;
    send[1] = 0;
    send[2] = der;
    send[3] = 0;
    send[4] = cf->enabled != 0u ? 0ull : 1ull;
    send[5] = send[6] = send[7] = 0;
}

__global__ void xchg_pack_bytes_kernel(const HostTotals *dt, uint64_t *send)
{
    send[0] = dt->bytes.anc;
    send[1] = dt->bytes.smd;
    send[2] = dt->bytes.der;
    send[3] = dt->bytes.mmd;
    send[4] = send[5] = send[6] = send[7] = 0;
}

// (n_dev: the number of rows when only the device knows it; n then bounds the grid)
__global__ void shift_offsets_dev_kernel(void *col, uint64_t n, const uint64_t *delta, int off32,
    const uint32_t *n_dev = nullptr)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t d = *delta;
    if (n_dev != nullptr) {
        n = (uint64_t) *n_dev + 1;
    }
    if (i < n && d != 0) {
        if (off32) {
            static_cast<uint32_t *>(col)[i] += (uint32_t) d;
        } else {
            static_cast<uint64_t *>(col)[i] += d;
        }
    }
}
} // namespace

extern "C" int mspb_exchange_buffers(mspb_ctx_t *ctx, void **send, void **recv)
{
    if (ctx == nullptr || send == nullptr || recv == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    TRY(ensure_exchange(ctx));
    *send = ctx->xchg.p;
    *recv = ctx->xchg.p + XCHG_RECV;
    return 0;
}

/* the stream the caller's next all-gather goes on: the context's own, or its second stream while the exchanges
 * of a call begun with mspb_generate_begin_device run beside K4 */
extern "C" void *mspb_exchange_stream(const mspb_ctx_t *ctx)
{
    if (ctx == nullptr) {
        return nullptr;
    }
    return (void *) ((ctx->pending.async && ctx->pending.early) ? ctx->copy_stream : ctx->stream);
}

/* after mspb_generate_begin: puts this rank's (sites, mutations, allele numbers used, new
 * mutations) into the send buffer, on the stream */
extern "C" int mspb_exchange_post_counts(mspb_ctx_t *ctx)
{
    if (ctx == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (!ctx->pending.valid) {
        return MSPB_ERR_NOT_GENERATED;
    }
    if (ctx->pending.defer_resolve) {
        // the SLiM model sizes its columns from the id base, which the host must know: mspb_generate_end
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    TRY(ensure_exchange(ctx));
    if (ctx->pending.async) {
        cudaStream_t ps = ctx->stream;
        if (ctx->pending.early) {
            // the counts are final once P1 is through: this exchange and the next run beside K4
            ps = ctx->copy_stream;
            CUDA_OK(cudaStreamWaitEvent(ps, ctx->ev_p1, 0));
        }
        xchg_pack_counts_kernel<<<1, 1, 0, ps>>>((const HostTotals *) ctx->dev_totals.p, ctx->pending.M,
            ctx->pending.Mnew, ctx->pending.model_kind != MSPB_MODEL_MATRIX ? 1 : 0, ctx->pending.key32 ? 1 : 0,
            ctx->xchg.p);
        ctx->log.count++;
        CUDA_OK(cudaGetLastError());
        return 0;
    }
    const mspb_counts_t &c = ctx->counts;
    uint64_t *h = ctx->xchg_host;
    memset(h, 0, XCHG_WORDS * sizeof(uint64_t));
    h[0] = c.num_sites;
    h[1] = c.num_mutations;
    h[2] = c.next_allele - ctx->pending.start_allele;
    h[3] = c.num_new_mutations;
    CUDA_OK(cudaMemcpyAsync(ctx->xchg.p, h, XCHG_WORDS * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream));
    return 0;
}

/* recv holds every rank's counts: bases on the device, K5 with them, byte totals into send.
 * Nothing is awaited. */
extern "C" int mspb_generate_end_device(mspb_ctx_t *ctx, uint32_t rank, uint32_t world)
{
    if (ctx == nullptr || world == 0 || world > XCHG_MAX_WORLD || rank >= world) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (!ctx->pending.valid) {
        return MSPB_ERR_NOT_GENERATED;
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    TRY(ensure_exchange(ctx));
    cudaStream_t st = ctx->stream;
    HostTotals *dt = (HostTotals *) ctx->dev_totals.p;
    uint64_t *x = ctx->xchg.p;
    ctx->pending.valid = false;
    CUDA_OK(cudaEventRecord(ctx->ev[0], st));
    if (ctx->pending.early) {
        // beside K4, on the second stream: bases of the first exchange, the ChainFree table with this slice's
        // allele base, this slice's byte counts from it.  K5 is launched by _finish_device.
        cudaStream_t xs = ctx->copy_stream;
        TRY(ctx->chain_free.ensure(1));
        xchg_bases_kernel<<<1, 32, 0, xs>>>(x + XCHG_RECV, rank, world, x + XCHG_B1);
        k5_prefix_table_kernel<<<1, 32, 0, xs>>>(ctx->pending.sv.site_row, &dt->p1, ctx->pending.R,
            ctx->pending.start_allele, x + XCHG_B1,
            ctx->pending.model_kind == MSPB_MODEL_MATRIX ? ctx->pending.fixed_len : 0u, &dt->dense, ctx->chain_free.p);
        xchg_pack_bytes_cf_kernel<<<1, 1, 0, xs>>>(ctx->chain_free.p, &dt->p1, ctx->pending.R, x);
        ctx->log.count += 3;
        CUDA_OK(cudaGetLastError());
        ctx->pending.cf_ready = true;
        ctx->pending.exchange_stage = 2;
        return 0;
    }
    MSPB_LAUNCH(ctx->log, "xchg_bases", xchg_bases_kernel<<<1, 32, 0, st>>>(x + XCHG_RECV, rank, world, x + XCHG_B1));
    ctx->pending.out.site_base = 0;
    ctx->pending.out.mut_base = 0;
    ctx->pending.out.dev_bases = x + XCHG_B1;
    ctx->pending.model.start_allele = ctx->pending.start_allele;
    if (ctx->pending.R > 0) {
        TRY(launch_emit(ctx));
    }
    ctx->pending.out.dev_bases = nullptr;
    MSPB_LAUNCH(ctx->log, "xchg_pack_bytes", xchg_pack_bytes_kernel<<<1, 1, 0, st>>>(dt, x));
    ctx->pending.exchange_stage = 2;
    return 0;
}

/* recv holds every rank's byte totals: offsets shifted on the device, then the one wait
 * of the exchange.  counts_out: this rank's counts (next_allele is the global one);
 * totals_out[8]: global sites, mutations, allele numbers used, new mutations, ancestral /
 * site-metadata / derived / mutation-metadata bytes; bases_out[8]: this rank's bases in
 * the same order. */
extern "C" int mspb_generate_finish_device(mspb_ctx_t *ctx, uint32_t rank, uint32_t world, mspb_counts_t *counts_out,
    uint64_t *totals_out, uint64_t *bases_out)
{
    if (ctx == nullptr || world == 0 || world > XCHG_MAX_WORLD || rank >= world) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (ctx->pending.exchange_stage != 2) {
        return MSPB_ERR_NOT_GENERATED;
    }
    ctx->pending.exchange_stage = 0;
    CUDA_OK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    HostTotals *ht = ctx->host;
    uint64_t *x = ctx->xchg.p;
    mspb_counts_t &c = ctx->counts;
    const int off32 = c.offset_bytes == 4;
    const bool async = ctx->pending.async;
    ctx->pending.async = false;
    // (after mspb_generate_begin_device the number of sites is still on the device: the grid covers every
    // record and the kernel reads the count)
    const uint64_t S1 = (async ? ctx->pending.R : c.num_sites) + 1, M1 = c.num_mutations + 1;
    const uint32_t *s_dev = async ? &((const HostTotals *) ctx->dev_totals.p)->p1.s : nullptr;
    const bool early = ctx->pending.early;
    ctx->pending.early = false;
    if (early) {
        // both exchanges are through (second stream): K5 writes ids, allele numbers and byte offsets with their
        // global bases, and nothing is shifted afterwards
        cudaStream_t xs = ctx->copy_stream;
        xchg_bases_kernel<<<1, 32, 0, xs>>>(x + XCHG_RECV, rank, world, x + XCHG_B2);
        ctx->log.count++;
        CUDA_OK(cudaEventRecord(ctx->ev_cf, xs));
        CUDA_OK(cudaStreamWaitEvent(st, ctx->ev_cf, 0));
        ctx->pending.out.site_base = 0;
        ctx->pending.out.mut_base = 0;
        ctx->pending.out.dev_bases = x + XCHG_B1;
        ctx->pending.out.dev_byte_bases = x + XCHG_B2;
        ctx->pending.model.start_allele = ctx->pending.start_allele;
        TRY(launch_emit(ctx));
        ctx->pending.out.dev_bases = nullptr;
        ctx->pending.out.dev_byte_bases = nullptr;
        ctx->pending.cf_ready = false;
    } else {
    MSPB_LAUNCH(ctx->log, "xchg_bases", xchg_bases_kernel<<<1, 32, 0, st>>>(x + XCHG_RECV, rank, world, x + XCHG_B2));
    MSPB_LAUNCH(ctx->log, "shift_offsets",
        shift_offsets_dev_kernel<<<grid_for(S1, 256), 256, 0, st>>>(ctx->o_site_anc_off.p, S1, x + XCHG_B2 + 0, off32,
            s_dev));
    MSPB_LAUNCH(ctx->log, "shift_offsets",
        shift_offsets_dev_kernel<<<grid_for(M1, 256), 256, 0, st>>>(ctx->o_mut_der_off.p, M1, x + XCHG_B2 + 2, off32));
    if (ctx->pending.out.write_md_off) {
        MSPB_LAUNCH(ctx->log, "shift_offsets",
            shift_offsets_dev_kernel<<<grid_for(S1, 256), 256, 0, st>>>(ctx->o_site_md_off.p, S1, x + XCHG_B2 + 1, off32,
                s_dev));
        MSPB_LAUNCH(ctx->log, "shift_offsets",
            shift_offsets_dev_kernel<<<grid_for(M1, 256), 256, 0, st>>>(ctx->o_mut_md_off.p, M1, x + XCHG_B2 + 3, off32));
        ctx->md_zero_site_bytes = ctx->md_zero_mut_bytes = 0;
    }
    }
    uint64_t *h = ctx->xchg_host;
    CUDA_OK(cudaMemcpyAsync(h + XCHG_WORDS, x + XCHG_B1, 4 * XCHG_WORDS * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
    TRY(totals_to_host(ctx));
    CUDA_OK(cudaEventRecord(ctx->ev[1], st));
    CUDA_OK(cudaStreamSynchronize(st));
    CUDA_OK(cudaGetLastError());
    ctx->log.collect();
    const uint64_t *b1 = h + XCHG_WORDS, *t1 = b1 + XCHG_WORDS, *b2 = t1 + XCHG_WORDS, *t2 = b2 + XCHG_WORDS;
    if (async) {
        if (t1[4] != 0 || t2[4] != 0) {
            // some rank (every rank reads the same sum) met an error, a dense genome or taken positions in
            // its _begin stage: the call is made again through mspb_generate_begin, which deals with those
            return MSPB_ERR_RETRY_WITH_HOST;
        }
        float begin_ms = 0;
        cudaEventElapsedTime(&begin_ms, ctx->ev_async0, ctx->ev_async1);
        c.device_ms = begin_ms;
        c.num_sites = ht->p1.s;
        c.num_new_sites = c.num_sites - ctx->pending.S0;
        c.next_allele = ctx->pending.start_allele + (ctx->pending.model_kind != MSPB_MODEL_MATRIX ? ht->p1.d : 0);
    }
    if (ht->err != 0) {
        return ht->err;
    }
    if (t1[0] > 0x7fffffffull || t1[1] > 0x7fffffffull) {
        return MSPB_ERR_TOO_MANY_ROWS;
    }
    if (off32 && (t2[0] > 0xffffffffull || t2[1] > 0xffffffffull || t2[2] > 0xffffffffull || t2[3] > 0xffffffffull)) {
        return MSPB_ERR_OFFSET_OVERFLOW;
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
    c.ancestral_state_bytes = ht->bytes.anc;
    c.site_metadata_bytes = ht->bytes.smd;
    c.derived_state_bytes = ht->bytes.der;
    c.mutation_metadata_bytes = ht->bytes.mmd;
    if (ctx->pending.model_kind == MSPB_MODEL_INFINITE_ALLELES) {
        c.next_allele = ctx->pending.start_allele + t1[2];
    }
    c.device_ms += ms;
    ctx->generated = true;
    if (counts_out != nullptr) {
        *counts_out = c;
    }
    for (int j = 0; j < 4; j++) {
        if (totals_out != nullptr) {
            totals_out[j] = t1[j];
            totals_out[4 + j] = t2[j];
        }
        if (bases_out != nullptr) {
            bases_out[j] = b1[j];
            bases_out[4 + j] = b2[j];
        }
    }
    return 0;
}

extern "C" int mspb_device_columns(mspb_ctx_t *ctx, mspb_columns_t *d)
{
    if (ctx == nullptr || d == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (!ctx->generated) {
        return MSPB_ERR_NOT_GENERATED;
    }
    d->site_position = ctx->o_site_pos.p;
    d->site_ancestral_state = ctx->o_site_anc.p;
    d->site_ancestral_state_offset = ctx->o_site_anc_off.p;
    d->site_metadata = ctx->o_site_md.p;
    d->site_metadata_offset = ctx->o_site_md_off.p;
    d->mutation_site = ctx->o_mut_site.p;
    d->mutation_node = ctx->o_mut_node.p;
    d->mutation_parent = ctx->o_mut_parent.p;
    d->mutation_time = ctx->o_mut_time.p;
    d->mutation_derived_state = ctx->o_mut_der.p;
    d->mutation_derived_state_offset = ctx->o_mut_der_off.p;
    d->mutation_metadata = ctx->o_mut_md.p;
    d->mutation_metadata_offset = ctx->o_mut_md_off.p;
    return 0;
}

extern "C" int mspb_fetch(mspb_ctx_t *ctx, const mspb_columns_t *h)
{
    if (ctx == nullptr || h == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (!ctx->generated) {
        return MSPB_ERR_NOT_GENERATED;
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const mspb_counts_t &c = ctx->counts;
    const uint64_t S = c.num_sites, M = c.num_mutations, ow = c.offset_bytes;
    ctx->last_fetch_bytes = 0;
#define D2H(dst, src, bytes)                                                                                           \
    do {                                                                                                               \
        if ((dst) != nullptr && (bytes) > 0) {                                                                         \
            CUDA_OK(cudaMemcpyAsync((dst), (src), (bytes), cudaMemcpyDeviceToHost, st));                               \
            ctx->last_fetch_bytes += (bytes);                                                                          \
        }                                                                                                              \
    } while (0)
    D2H(h->site_position, ctx->o_site_pos.p, S * sizeof(double));
    D2H(h->site_ancestral_state, ctx->o_site_anc.p, c.ancestral_state_bytes);
    D2H(h->site_ancestral_state_offset, ctx->o_site_anc_off.p, (S + 1) * ow);
    D2H(h->site_metadata, ctx->o_site_md.p, c.site_metadata_bytes);
    D2H(h->mutation_site, ctx->o_mut_site.p, M * sizeof(int32_t));
    D2H(h->mutation_node, ctx->o_mut_node.p, M * sizeof(int32_t));
    D2H(h->mutation_parent, ctx->o_mut_parent.p, M * sizeof(int32_t));
    D2H(h->mutation_time, ctx->o_mut_time.p, M * sizeof(double));
    D2H(h->mutation_derived_state, ctx->o_mut_der.p, c.derived_state_bytes);
    D2H(h->mutation_derived_state_offset, ctx->o_mut_der_off.p, (M + 1) * ow);
    D2H(h->mutation_metadata, ctx->o_mut_md.p, c.mutation_metadata_bytes);
    // The metadata offset columns are all zero unless rows were kept with metadata or the model
    // writes some (SLiM), and the context tracks that (md_zero_*): such a column is not worth
    // 8 bytes per row of PCIe traffic.  The host fills it while the other copies are in flight.
    const bool site_md_zero = ctx->md_zero_site_ptr == ctx->o_site_md_off.p && ctx->md_zero_site_bytes >= (S + 1) * ow
                              && c.site_metadata_bytes == 0;
    const bool mut_md_zero = ctx->md_zero_mut_ptr == ctx->o_mut_md_off.p && ctx->md_zero_mut_bytes >= (M + 1) * ow
                             && c.mutation_metadata_bytes == 0;
    if (!site_md_zero) {
        D2H(h->site_metadata_offset, ctx->o_site_md_off.p, (S + 1) * ow);
    }
    if (!mut_md_zero) {
        D2H(h->mutation_metadata_offset, ctx->o_mut_md_off.p, (M + 1) * ow);
    }
#undef D2H
    if (site_md_zero && h->site_metadata_offset != nullptr) {
        memset(h->site_metadata_offset, 0, (S + 1) * ow);
    }
    if (mut_md_zero && h->mutation_metadata_offset != nullptr) {
        memset(h->mutation_metadata_offset, 0, (M + 1) * ow);
    }
    CUDA_OK(cudaStreamSynchronize(st));
    return 0;
}

extern "C" uint64_t mspb_last_fetch_bytes(const mspb_ctx_t *ctx)
{
    return ctx != nullptr ? ctx->last_fetch_bytes : 0;
}

namespace {
bool pinned_pool_owns(void *p);
void host_column_free(void *p)
{
    if (p == nullptr) {
        return;
    }
    if (pinned_pool_owns(p)) {
        mspb_pinned_free(p);
    } else {
        free(p);
    }
}
} // namespace

extern "C" void mspb_columns_free(mspb_columns_t *c)
{
    if (c == nullptr) {
        return;
    }
    host_column_free(c->site_position);
    host_column_free(c->site_ancestral_state);
    host_column_free(c->site_ancestral_state_offset);
    host_column_free(c->site_metadata);
    host_column_free(c->site_metadata_offset);
    host_column_free(c->mutation_site);
    host_column_free(c->mutation_node);
    host_column_free(c->mutation_parent);
    host_column_free(c->mutation_time);
    host_column_free(c->mutation_derived_state);
    host_column_free(c->mutation_derived_state_offset);
    host_column_free(c->mutation_metadata);
    host_column_free(c->mutation_metadata_offset);
    memset(c, 0, sizeof(*c));
}

// The one-shot call keeps ONE context per device between calls (the buffers of a bench-size overlay are several
// GB of device memory: allocating and freeing them took most of a call) and serialises the calls on it;
// mspb_oneshot_release gives the memory back.  The output columns come from the pinned pool (a download into
// fresh pageable memory runs at a tenth of the link's rate); mspb_columns_free returns them to it.
namespace {
constexpr int ONESHOT_MAX_DEVICES = 64;
std::mutex g_oneshot_mutex[ONESHOT_MAX_DEVICES];
mspb_ctx_t *g_oneshot_ctx[ONESHOT_MAX_DEVICES] = {};

void *host_column(uint64_t bytes)
{
    void *p = mspb_pinned_alloc(bytes);
    return p != nullptr ? p : malloc((size_t) bytes);
}
} // namespace

extern "C" void mspb_oneshot_release(int device)
{
    for (int d = 0; d < ONESHOT_MAX_DEVICES; d++) {
        if (device >= 0 && d != device) {
            continue;
        }
        std::lock_guard<std::mutex> lock(g_oneshot_mutex[d]);
        if (g_oneshot_ctx[d] != nullptr) {
            mspb_ctx_destroy(g_oneshot_ctx[d]);
            g_oneshot_ctx[d] = nullptr;
        }
    }
}

extern "C" int mspb_sim_mutations(int device, const mspb_tables_t *tables, const mspb_params_t *params,
    mspb_counts_t *counts, mspb_columns_t *out)
{
    mspb_counts_t c;
    if (tables == nullptr || params == nullptr || out == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    memset(out, 0, sizeof(*out));
    if (device < 0 || device >= ONESHOT_MAX_DEVICES) {
        return MSPB_ERR_NO_DEVICE;
    }
    std::lock_guard<std::mutex> lock(g_oneshot_mutex[device]);
    int ret = 0;
    if (g_oneshot_ctx[device] == nullptr) {
        ret = mspb_ctx_create(device, &g_oneshot_ctx[device]);
        if (ret != 0) {
            g_oneshot_ctx[device] = nullptr;
            return ret;
        }
    }
    mspb_ctx_t *ctx = g_oneshot_ctx[device];
    ret = mspb_ctx_load_tables(ctx, tables);
    if (ret == 0) {
        ret = mspb_generate(ctx, params, &c);
    }
    if (ret == 0) {
        const uint64_t S = c.num_sites, M = c.num_mutations;
        out->site_position = (double *) host_column((S + 1) * sizeof(double));
        out->site_ancestral_state = (char *) host_column(c.ancestral_state_bytes + 1);
        out->site_ancestral_state_offset = host_column((S + 1) * sizeof(uint64_t));
        out->site_metadata = (char *) host_column(c.site_metadata_bytes + 1);
        out->site_metadata_offset = host_column((S + 1) * sizeof(uint64_t));
        out->mutation_site = (int32_t *) host_column((M + 1) * sizeof(int32_t));
        out->mutation_node = (int32_t *) host_column((M + 1) * sizeof(int32_t));
        out->mutation_parent = (int32_t *) host_column((M + 1) * sizeof(int32_t));
        out->mutation_time = (double *) host_column((M + 1) * sizeof(double));
        out->mutation_derived_state = (char *) host_column(c.derived_state_bytes + 1);
        out->mutation_derived_state_offset = host_column((M + 1) * sizeof(uint64_t));
        out->mutation_metadata = (char *) host_column(c.mutation_metadata_bytes + 1);
        out->mutation_metadata_offset = host_column((M + 1) * sizeof(uint64_t));
        if (!out->site_position || !out->site_ancestral_state || !out->site_ancestral_state_offset
            || !out->site_metadata || !out->site_metadata_offset || !out->mutation_site || !out->mutation_node
            || !out->mutation_parent || !out->mutation_time || !out->mutation_derived_state
            || !out->mutation_derived_state_offset || !out->mutation_metadata || !out->mutation_metadata_offset) {
            ret = MSPB_ERR_NO_MEMORY;
        } else {
            ret = mspb_fetch(ctx, out);
        }
        if (ret != 0) {
            mspb_columns_free(out);
        } else if (counts != nullptr) {
            *counts = c;
        }
    }
    if (ret != 0) {
        // whatever went wrong, the next call starts from a fresh context
        mspb_ctx_destroy(ctx);
        g_oneshot_ctx[device] = nullptr;
    }
    return ret;
}

// ---------------------------------------------------------------- diagnostics
namespace {

template <typename KeyT>
int selftest_sort_impl(mspb_ctx *ctx, uint64_t n, uint64_t key_mask, uint64_t seed, bool measure_bits, int *passes)
{
    cudaStream_t st = ctx->stream;
    std::vector<KeyT> h(n > 0 ? n : 1);
    uint64_t x = seed;
    for (uint64_t i = 0; i < n; i++) {
        x += 0x9E3779B97F4A7C15ull;
        uint64_t z = x;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        h[i] = (KeyT) (z & key_mask);
    }
    DBuf<KeyT> k0, k1;
    DBuf<uint32_t> v0, v1, hist, state;
    int ret = 0;
    if (k0.ensure(n + 1) || k1.ensure(n + 1) || v0.ensure(n + 1) || v1.ensure(n + 1)
        || hist.ensure((size_t) RS_MAX_PASSES * RS_BINS + 1) || state.ensure(rs_state_words(n))) {
        ret = MSPB_ERR_NO_MEMORY;
    }
    if (ret == 0) {
        cudaMemcpyAsync(k0.p, h.data(), n * sizeof(KeyT), cudaMemcpyHostToDevice, st);
        RadixSortBuffers<KeyT> b;
        b.keys[0] = k0.p;
        b.keys[1] = k1.p;
        b.vals[0] = v0.p;
        b.vals[1] = v1.p;
        b.hist = hist.p;
        b.tile_state = state.p;
        b.or_and = ctx->or_and.p;
        b.or_and_host = ctx->or_and_host;
        b.or_and_host_dev = ctx->or_and_host_dev;
        const int cur = device_radix_sort<KeyT>(b, n, measure_bits ? 0 : (key_mask != 0 ? key_mask : 1), true, st, ctx->log,
            passes);
        std::vector<KeyT> ok(n > 0 ? n : 1);
        std::vector<uint32_t> ov(n > 0 ? n : 1);
        cudaMemcpyAsync(ok.data(), b.keys[cur], n * sizeof(KeyT), cudaMemcpyDeviceToHost, st);
        cudaMemcpyAsync(ov.data(), b.vals[cur], n * sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
        if (cudaStreamSynchronize(st) != cudaSuccess || cudaGetLastError() != cudaSuccess) {
            ret = MSPB_ERR_CUDA;
        }
        ctx->log.collect();
        for (uint64_t i = 0; i < n && ret == 0; i++) {
            if (ov[i] >= n || h[ov[i]] != ok[i]) {
                ret = MSPB_ERR_GENERIC; // payload does not travel with its key
            } else if (i > 0 && (ok[i - 1] > ok[i] || (ok[i - 1] == ok[i] && ov[i - 1] >= ov[i]))) {
                ret = MSPB_ERR_GENERIC; // not sorted, or not stable
            }
        }
    }
    k0.release();
    k1.release();
    v0.release();
    v1.release();
    hist.release();
    state.release();
    return ret;
}

} // namespace

extern "C" int mspb_selftest_sort(mspb_ctx_t *ctx, uint64_t n, int key_bytes, uint64_t key_mask, uint64_t seed,
    int measure_bits, int *passes)
{
    if (ctx == nullptr || n > RS_MAX_N || (key_bytes != 4 && key_bytes != 8)) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    if (key_bytes == 4) {
        return selftest_sort_impl<uint32_t>(ctx, n, key_mask & 0xffffffffull, seed, measure_bits != 0, passes);
    }
    return selftest_sort_impl<uint64_t>(ctx, n, key_mask, seed, measure_bits != 0, passes);
}

// ---------------------------------------------------------------- pinned host buffers
// A download into pageable memory is staged by the driver at a few GB/s (0.22 s for the 0.74 GB of a bench-size
// result, against 13 ms of DMA into pinned memory), and pinning memory per call costs as much.  Callers that
// want their result columns in buffers of their own lifetime take them from this pool: blocks are pinned once
// and handed out again after mspb_pinned_free (the Python layer frees a block when the numpy array over it is
// collected), so a steady caller allocates nothing.
namespace {

std::mutex g_pinned_mutex;
std::multimap<size_t, void *> g_pinned_free; // capacity -> block
std::map<void *, size_t> g_pinned_live; // block -> capacity
size_t g_pinned_cached_bytes = 0;
uint64_t g_pinned_stats[4] = { 0, 0, 0, 0 }; // blocks pinned, blocks reused, bytes pinned so far, blocks unpinned
constexpr size_t PINNED_CACHE_LIMIT = 16ull << 30; // free blocks kept pinned at most

} // namespace

namespace {
bool pinned_pool_owns(void *p)
{
    std::lock_guard<std::mutex> lock(g_pinned_mutex);
    return g_pinned_live.find(p) != g_pinned_live.end();
}
} // namespace

extern "C" void *mspb_pinned_alloc(uint64_t bytes)
{
    // size classes an eighth of an octave apart (at least 1 MiB): the columns of successive results differ by
    // a fraction of a percent in size, and must find the blocks of the results before them
    size_t want = (size_t) 1 << 20;
    const size_t need = (size_t) (bytes > 0 ? bytes : 1);
    while (want < need) {
        size_t step = want >> 3;
        size_t next = want;
        for (int k = 0; k < 8 && next < need; k++) {
            next += step;
        }
        if (next >= need) {
            want = next;
            break;
        }
        want <<= 1;
    }
    {
        std::lock_guard<std::mutex> lock(g_pinned_mutex);
        auto it = g_pinned_free.lower_bound(want);
        if (it != g_pinned_free.end() && it->first <= want + want / 4) {
            void *p = it->second;
            g_pinned_live[p] = it->first;
            g_pinned_cached_bytes -= it->first;
            g_pinned_free.erase(it);
            g_pinned_stats[1]++;
            return p;
        }
    }
    void *p = nullptr;
    if (cudaHostAlloc(&p, want, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(g_pinned_mutex);
    g_pinned_live[p] = want;
    g_pinned_stats[0]++;
    g_pinned_stats[2] += want;
    return p;
}

// blocks pinned, blocks reused from the pool, bytes pinned so far, blocks unpinned again
extern "C" void mspb_pinned_stats(uint64_t out[4])
{
    std::lock_guard<std::mutex> lock(g_pinned_mutex);
    for (int k = 0; k < 4; k++) {
        out[k] = g_pinned_stats[k];
    }
}

extern "C" void mspb_pinned_free(void *p)
{
    if (p == nullptr) {
        return;
    }
    size_t cap = 0;
    {
        std::lock_guard<std::mutex> lock(g_pinned_mutex);
        auto it = g_pinned_live.find(p);
        if (it == g_pinned_live.end()) {
            return; // not one of ours
        }
        cap = it->second;
        g_pinned_live.erase(it);
        if (g_pinned_cached_bytes + cap <= PINNED_CACHE_LIMIT) {
            g_pinned_free.emplace(cap, p);
            g_pinned_cached_bytes += cap;
            return;
        }
    }
    {
        std::lock_guard<std::mutex> lock(g_pinned_mutex);
        g_pinned_stats[3]++;
    }
    cudaFreeHost(p);
}

// ---------------------------------------------------------------- consumers of the tables
namespace {

// room for `need` steps in fx / fv, keeping what is there
int sidx_reserve(mspb_ctx *ctx, size_t need)
{
    if (need <= ctx->sidx.fcap) {
        return 0;
    }
    const size_t cap = need + need / 2 + 1024;
    uint32_t *nx = nullptr, *nv = nullptr;
    if (cudaMalloc((void **) &nx, cap * sizeof(uint32_t)) != cudaSuccess
        || cudaMalloc((void **) &nv, cap * sizeof(uint32_t)) != cudaSuccess) {
        cudaGetLastError();
        if (nx != nullptr) {
            cudaFree(nx);
        }
        return MSPB_ERR_NO_MEMORY;
    }
    if (ctx->sidx.steps > 0) {
        CUDA_OK(cudaMemcpyAsync(nx, ctx->sidx.fx, ctx->sidx.steps * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
        CUDA_OK(cudaMemcpyAsync(nv, ctx->sidx.fv, ctx->sidx.steps * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
        CUDA_OK(cudaStreamSynchronize(ctx->stream));
    }
    if (ctx->sidx.fx != nullptr) {
        cudaFree(ctx->sidx.fx);
        cudaFree(ctx->sidx.fv);
    }
    ctx->sidx.fx = nx;
    ctx->sidx.fv = nv;
    ctx->sidx.fcap = cap;
    return 0;
}

SampleFnView sidx_view(const mspb_ctx *ctx)
{
    SampleFnView v;
    v.num_nodes = ctx->N;
    v.is_sample = ctx->sidx.is_sample.p;
    v.foff = ctx->sidx.foff.p;
    v.fcnt = ctx->sidx.fcnt.p;
    v.fx = ctx->sidx.fx;
    v.fv = ctx->sidx.fv;
    v.xs = ctx->sidx.xs.p;
    v.num_x = ctx->sidx.num_x;
    return v;
}

uint64_t bits_mask(uint64_t max_value)
{
    uint64_t m = 1;
    while (m <= max_value && m != 0) {
        m <<= 1;
    }
    return m - 1; // (m == 0 after a wrap: all ones)
}

struct SidxTemps {
    DBuf<uint32_t> level, lrank, rrank, eorder, count, first, offset, delta, before, kept_before, tile_sums, small;
    ~SidxTemps()
    {
        level.release(); lrank.release(); rrank.release(); eorder.release(); count.release(); first.release();
        offset.release(); delta.release(); before.release(); kept_before.release(); tile_sums.release(); small.release();
    }
};

int build_sample_index(mspb_ctx *ctx, const uint32_t *node_flags)
{
    cudaStream_t st = ctx->stream;
    const uint64_t N = ctx->N, E = ctx->E;
    auto &si = ctx->sidx;
    si.valid = false;
    si.steps = 0;
    si.levels = 0;
    SidxTemps t;
    cudaEvent_t e0 = ctx->ev[0], e1 = ctx->ev[1];
    CUDA_OK(cudaEventRecord(e0, st));
    TRY(si.is_sample.ensure(N + 1));
    TRY(si.foff.ensure(N + 1));
    TRY(si.fcnt.ensure(N + 1));
    TRY(t.small.ensure(8));
    TRY(t.level.ensure(N + 1));
    CUDA_OK(cudaMemsetAsync(t.small.p, 0, 8 * sizeof(uint32_t), st));
    CUDA_OK(cudaMemsetAsync(si.fcnt.p, 0, (N + 1) * sizeof(uint32_t), st));
    CUDA_OK(cudaMemsetAsync(si.foff.p, 0, (N + 1) * sizeof(uint32_t), st));
    CUDA_OK(cudaMemsetAsync(t.level.p, 0, (N + 1) * sizeof(uint32_t), st));
    // node flags arrive in t.level's neighbour (a temporary of N words)
    TRY(t.lrank.ensure((N > E ? N : E) + 1));
    if (N > 0) {
        CUDA_OK(cudaMemcpyAsync(t.lrank.p, node_flags, N * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
        MSPB_LAUNCH(ctx->log, "sf_flags",
            sf_flags_kernel<<<grid_for(N, 256), 256, 0, st>>>(N, t.lrank.p, si.is_sample.p, t.small.p));
    }
    uint32_t h_small[8] = { 0 };
    CUDA_OK(cudaMemcpyAsync(h_small, t.small.p, sizeof(h_small), cudaMemcpyDeviceToHost, st));
    CUDA_OK(cudaStreamSynchronize(st)); // (also: the caller's flags are no longer read)
    si.num_samples = h_small[0];
    si.num_x = 0;
    // the sample nodes in increasing order (tskit's default sample list)
    TRY(si.samples.ensure((size_t) si.num_samples + 1));
    if (N > 0) {
        TRY(t.before.ensure(N + 1));
        TRY(t.tile_sums.ensure(scan_num_tiles(N) + 2));
        device_exclusive_scan<uint32_t, SfIsSample>(ctx->log, N, SfIsSample { si.is_sample.p }, t.before.p, t.tile_sums.p,
            nullptr, st);
        MSPB_LAUNCH(ctx->log, "sf_sample_list",
            sf_sample_list_kernel<<<grid_for(N, 256), 256, 0, st>>>(N, si.is_sample.p, t.before.p, si.samples.p));
    }
    TRY(si.xs.ensure(2 * E + 2));
    TRY(sidx_reserve(ctx, 5 * E + 1024));
    if (E > 0) {
        const TreeSeqView ts = make_ts_view(ctx);
        // ---- the distinct edge end points, ascending
        TRY(ensure_sort_buffers(ctx, 2 * E, false, true));
        RadixSortBuffers<uint64_t> sb = sort_buffers64(ctx);
        MSPB_LAUNCH(ctx->log, "sf_endpoint_keys",
            sf_endpoint_keys_kernel<<<grid_for(E, 256), 256, 0, st>>>(ts, sb.keys[0]));
        int cur = device_radix_sort<uint64_t>(sb, 2 * E, 0, true, st, ctx->log, nullptr);
        TRY(t.before.ensure(2 * E + 1));
        TRY(t.tile_sums.ensure(scan_num_tiles(2 * E) + 2));
        SfHead head { sb.keys[cur] };
        device_exclusive_scan<uint32_t, SfHead>(ctx->log, 2 * E, head, t.before.p, t.tile_sums.p, t.small.p + 1, st);
        MSPB_LAUNCH(ctx->log, "sf_unique",
            sf_unique_kernel<<<grid_for(2 * E, 256), 256, 0, st>>>(2 * E, sb.keys[cur], t.before.p, si.xs.p));
        CUDA_OK(cudaMemcpyAsync(h_small, t.small.p, sizeof(h_small), cudaMemcpyDeviceToHost, st));
        CUDA_OK(cudaStreamSynchronize(st));
        si.num_x = h_small[1];
        TRY(t.rrank.ensure(E + 1));
        MSPB_LAUNCH(ctx->log, "sf_edge_ranks",
            sf_edge_ranks_kernel<<<grid_for(E, 256), 256, 0, st>>>(ts, si.xs.p, si.num_x, t.lrank.p, t.rrank.p));
        // ---- levels: relax until nothing changes (one round per level of the deepest node at most)
        uint32_t rounds = 0;
        for (;;) {
            CUDA_OK(cudaMemsetAsync(t.small.p + 2, 0, sizeof(uint32_t), st));
            // a few rounds per wait: the flag is only ever set
            for (int k = 0; k < 4; k++) {
                MSPB_LAUNCH(ctx->log, "sf_relax_levels",
                    sf_relax_levels_kernel<<<grid_for(E, 256), 256, 0, st>>>(ts, t.level.p, t.small.p + 2));
            }
            rounds += 4;
            CUDA_OK(cudaMemcpyAsync(h_small, t.small.p, sizeof(h_small), cudaMemcpyDeviceToHost, st));
            CUDA_OK(cudaStreamSynchronize(st));
            if (h_small[2] == 0u) {
                break;
            }
            if (rounds > N + 8) {
                return MSPB_ERR_GENERIC; // cannot happen for edges whose parents are older than their children
            }
        }
        // ---- edges by (level of the parent, parent).  The deepest level is read from the levels themselves: the
        // in-place rounds above can climb several levels at once, so their number is no bound (with `rounds + 2`
        // buckets the deepest levels of a 2e6-leaf Hudson tree sequence shared the last bucket and their nodes were
        // summed before their children were final)
        CUDA_OK(cudaMemsetAsync(t.small.p + 5, 0, sizeof(uint32_t), st));
        MSPB_LAUNCH(ctx->log, "sf_max_level",
            sf_max_level_kernel<<<grid_for(N, 256), 256, 0, st>>>(N, t.level.p, t.small.p + 5));
        CUDA_OK(cudaMemcpyAsync(h_small, t.small.p, sizeof(h_small), cudaMemcpyDeviceToHost, st));
        CUDA_OK(cudaStreamSynchronize(st));
        const uint32_t max_levels = h_small[5] + 2;
        DBuf<uint32_t> level_count;
        TRY(level_count.ensure(max_levels + 1));
        struct Rel {
            DBuf<uint32_t> &b;
            ~Rel() { b.release(); }
        } rel { level_count };
        CUDA_OK(cudaMemsetAsync(level_count.p, 0, (max_levels + 1) * sizeof(uint32_t), st));
        TRY(ensure_sort_buffers(ctx, E, false, true));
        sb = sort_buffers64(ctx);
        MSPB_LAUNCH(ctx->log, "sf_edge_level_keys",
            sf_edge_level_keys_kernel<<<grid_for(E, 256), 256, 0, st>>>(ts, t.level.p, sb.keys[0], level_count.p, max_levels));
        cur = device_radix_sort<uint64_t>(sb, E, ((uint64_t) bits_mask(max_levels) << 32) | bits_mask(N), true, st,
            ctx->log, nullptr);
        TRY(t.eorder.ensure(E + 1));
        CUDA_OK(cudaMemcpyAsync(t.eorder.p, sb.vals[cur], E * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
        std::vector<uint32_t> h_count(max_levels + 1);
        CUDA_OK(cudaMemcpyAsync(h_count.data(), level_count.p, (max_levels + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        CUDA_OK(cudaStreamSynchronize(st));
        // ---- one level at a time (level 0 has no child edges: those nodes are constant)
        const uint64_t key_bits = ((uint64_t) bits_mask(N) << 32) | bits_mask(si.num_x);
        uint64_t a = h_count[0]; // (no edge has a parent of level 0)
        for (uint32_t lv = 1; lv <= max_levels; lv++) {
            const uint64_t n_e = h_count[lv];
            if (n_e == 0) {
                continue;
            }
            si.levels = lv + 1;
            SfLevelEdges le { ts, t.eorder.p + a, t.lrank.p, t.rrank.p };
            a += n_e;
            TRY(t.count.ensure(n_e + 1));
            TRY(t.first.ensure(n_e + 1));
            TRY(t.offset.ensure(n_e + 1));
            TRY(t.tile_sums.ensure(scan_num_tiles(n_e) + 2));
            MSPB_LAUNCH(ctx->log, "sf_count_events",
                sf_count_events_kernel<<<grid_for(n_e, 256), 256, 0, st>>>(le, n_e, sidx_view(ctx), t.count.p, t.first.p));
            device_exclusive_scan<uint32_t, SfArray>(ctx->log, n_e, SfArray { t.count.p }, t.offset.p, t.tile_sums.p,
                t.small.p + 3, st);
            CUDA_OK(cudaMemcpyAsync(h_small, t.small.p, sizeof(h_small), cudaMemcpyDeviceToHost, st));
            CUDA_OK(cudaStreamSynchronize(st));
            const uint64_t T = h_small[3];
            if (T > RS_MAX_N) {
                return MSPB_ERR_TOO_MANY_ROWS;
            }
            TRY(ensure_sort_buffers(ctx, T, false, true));
            sb = sort_buffers64(ctx);
            TRY(t.delta.ensure(T + 1));
            TRY(t.before.ensure(T + 1));
            TRY(t.kept_before.ensure(T + 1));
            TRY(t.tile_sums.ensure(scan_num_tiles(T) + 2));
            MSPB_LAUNCH(ctx->log, "sf_emit_events",
                sf_emit_events_kernel<<<grid_for(n_e, 256), 256, 0, st>>>(le, n_e, sidx_view(ctx), t.count.p, t.first.p,
                    t.offset.p, sb.keys[0], t.delta.p));
            cur = device_radix_sort<uint64_t>(sb, T, key_bits, true, st, ctx->log, nullptr);
            device_exclusive_scan<uint32_t, SfGather>(ctx->log, T, SfGather { t.delta.p, sb.vals[cur] }, t.before.p,
                t.tile_sums.p, nullptr, st);
            device_exclusive_scan<uint32_t, SfKeep>(ctx->log, T, SfKeep { sb.keys[cur], T }, t.kept_before.p,
                t.tile_sums.p, t.small.p + 4, st);
            CUDA_OK(cudaMemcpyAsync(h_small, t.small.p, sizeof(h_small), cudaMemcpyDeviceToHost, st));
            CUDA_OK(cudaStreamSynchronize(st));
            const uint64_t K = h_small[4];
            if (si.steps + K > 0xfffffff0ull) {
                return MSPB_ERR_TOO_MANY_ROWS;
            }
            TRY(sidx_reserve(ctx, si.steps + K));
            MSPB_LAUNCH(ctx->log, "sf_store_steps",
                sf_store_steps_kernel<<<grid_for(T, 256), 256, 0, st>>>(T, sb.keys[cur], t.delta.p, sb.vals[cur],
                    t.before.p, t.kept_before.p, si.is_sample.p, (uint32_t) si.steps, si.fx, si.fv, si.foff.p,
                    si.fcnt.p));
            si.steps += K;
        }
    }
    CUDA_OK(cudaEventRecord(e1, st));
    CUDA_OK(cudaStreamSynchronize(st));
    CUDA_OK(cudaGetLastError());
    ctx->log.collect();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    si.build_ms = ms;
    si.valid = true;
    return 0;
}

// below / own of every mutation row of the resident result
int query_below(mspb_ctx *ctx)
{
    cudaStream_t st = ctx->stream;
    if (!ctx->generated) {
        return MSPB_ERR_NOT_GENERATED;
    }
    if (!ctx->sidx.valid) {
        return MSPB_ERR_TABLES_NOT_LOADED;
    }
    if (ctx->pending.out.site_base != 0 || ctx->pending.out.mut_base != 0 || ctx->pending.out.dev_bases != nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE; // rows of one genome slice of a sharded call: ids are not local
    }
    const uint64_t M = ctx->counts.num_mutations;
    TRY(ctx->q_below.ensure(M + 1));
    TRY(ctx->q_own.ensure(M + 1));
    if (M > 0) {
        MSPB_LAUNCH(ctx->log, "sf_below",
            sf_below_kernel<<<grid_for(M, 256), 256, 0, st>>>(M, ctx->o_mut_site.p, ctx->o_mut_node.p, ctx->o_site_pos.p,
                sidx_view(ctx), ctx->q_below.p, ctx->q_own.p));
        MSPB_LAUNCH(ctx->log, "sf_own",
            sf_own_kernel<<<grid_for(M, 256), 256, 0, st>>>(M, ctx->o_mut_parent.p, ctx->q_below.p, ctx->q_own.p));
    }
    return 0;
}

} // namespace

extern "C" int mspb_ctx_build_sample_index(mspb_ctx_t *ctx, const uint32_t *node_flags, uint64_t *num_samples,
    uint64_t *num_levels, uint64_t *num_steps, double *build_ms)
{
    if (ctx == nullptr || (node_flags == nullptr && ctx->N > 0)) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (!ctx->tables_loaded) {
        return MSPB_ERR_TABLES_NOT_LOADED;
    }
    if (ctx->pending.valid) {
        return MSPB_ERR_BAD_PARAM_VALUE; // between mspb_generate_begin and _end: the sort buffers are in use
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    const int ret = build_sample_index(ctx, node_flags);
    if (ret != 0) {
        cudaStreamSynchronize(ctx->stream);
        ctx->log.collect();
        return ret;
    }
    if (num_samples != nullptr) {
        *num_samples = ctx->sidx.num_samples;
    }
    if (num_levels != nullptr) {
        *num_levels = ctx->sidx.levels;
    }
    if (num_steps != nullptr) {
        *num_steps = ctx->sidx.steps;
    }
    if (build_ms != nullptr) {
        *build_ms = ctx->sidx.build_ms;
    }
    return 0;
}

extern "C" int mspb_samples_below(mspb_ctx_t *ctx, uint32_t *below)
{
    if (ctx == nullptr || below == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    TRY(query_below(ctx));
    const uint64_t M = ctx->counts.num_mutations;
    if (M > 0) {
        CUDA_OK(cudaMemcpyAsync(below, ctx->q_below.p, M * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    }
    CUDA_OK(cudaStreamSynchronize(ctx->stream));
    CUDA_OK(cudaGetLastError());
    ctx->log.collect();
    return 0;
}

extern "C" int mspb_decode_genotypes(mspb_ctx_t *ctx, uint64_t site_begin, uint64_t site_end, int32_t *genotypes,
    double *device_ms)
{
    if (ctx == nullptr || genotypes == nullptr || site_end < site_begin) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (!ctx->generated) {
        return MSPB_ERR_NOT_GENERATED;
    }
    if (!ctx->sidx.valid) {
        return MSPB_ERR_TABLES_NOT_LOADED;
    }
    if (site_end > ctx->counts.num_sites || ctx->pending.out.site_base != 0 || ctx->pending.out.mut_base != 0
        || ctx->pending.out.dev_bases != nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const uint64_t ns = site_end - site_begin, n = ctx->sidx.num_samples, M = ctx->counts.num_mutations;
    if (ns == 0 || n == 0) {
        return 0;
    }
    if (ns > 65535ull || ns * n > RS_MAX_N * 16ull) {
        return MSPB_ERR_TOO_MANY_ROWS; // decode a range of sites at a time
    }
    TRY(ctx->q_row_begin.ensure(ns + 1));
    TRY(ctx->q_row_end.ensure(ns + 1));
    TRY(ctx->q_tmax.ensure(ns + 1));
    TRY(ctx->q_allele.ensure(M + 1));
    TRY(ctx->q_walk_id.ensure(M + 1));
    TRY(ctx->q_geno.ensure(ns * n + 1));
    SiteRows sr { ctx->q_row_begin.p, ctx->q_row_end.p, ctx->q_tmax.p, ctx->q_allele.p, ctx->q_walk_id.p };
    const TreeSeqView ts = make_ts_view(ctx);
    cudaEvent_t e0 = ctx->ev[0], e1 = ctx->ev[1];
    CUDA_OK(cudaEventRecord(e0, st));
    if (ctx->counts.offset_bytes == 4) {
        MSPB_LAUNCH(ctx->log, "sf_site_alleles",
            sf_site_alleles_kernel<true><<<grid_for(ns, 128), 128, 0, st>>>(ts, site_begin, ns, M, ctx->o_mut_site.p,
                ctx->o_mut_node.p, ctx->o_mut_der.p, ctx->o_mut_der_off.p, ctx->o_site_anc.p, ctx->o_site_anc_off.p, sr));
    } else {
        MSPB_LAUNCH(ctx->log, "sf_site_alleles",
            sf_site_alleles_kernel<false><<<grid_for(ns, 128), 128, 0, st>>>(ts, site_begin, ns, M, ctx->o_mut_site.p,
                ctx->o_mut_node.p, ctx->o_mut_der.p, ctx->o_mut_der_off.p, ctx->o_site_anc.p, ctx->o_site_anc_off.p, sr));
    }
    const dim3 grid(grid_for(n, 256), (unsigned) ns);
    MSPB_LAUNCH(ctx->log, "sf_decode",
        sf_decode_kernel<<<grid, 256, 0, st>>>(ts, site_begin, ns, ctx->o_site_pos.p, ctx->sidx.samples.p, n, sr,
            ctx->q_geno.p));
    CUDA_OK(cudaEventRecord(e1, st));
    CUDA_OK(cudaMemcpyAsync(genotypes, ctx->q_geno.p, ns * n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CUDA_OK(cudaStreamSynchronize(st));
    CUDA_OK(cudaGetLastError());
    ctx->log.collect();
    if (device_ms != nullptr) {
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        *device_ms = ms;
    }
    return 0;
}

extern "C" int mspb_site_afs(mspb_ctx_t *ctx, uint64_t *afs, uint64_t afs_len, double *device_ms)
{
    if (ctx == nullptr || afs == nullptr) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    CUDA_OK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (ctx->sidx.valid && afs_len != (uint64_t) ctx->sidx.num_samples + 1) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    cudaEvent_t e0 = ctx->ev[0], e1 = ctx->ev[1];
    CUDA_OK(cudaEventRecord(e0, st));
    TRY(query_below(ctx));
    const uint64_t M = ctx->counts.num_mutations;
    TRY(ctx->q_afs.ensure(afs_len + 1));
    CUDA_OK(cudaMemsetAsync(ctx->q_afs.p, 0, afs_len * sizeof(unsigned long long), st));
    if (M > 0) {
        const unsigned afs_grid = grid_for(M, 256) < 148u * 8u ? grid_for(M, 256) : 148u * 8u;
        if (ctx->counts.offset_bytes == 4) {
            MSPB_LAUNCH(ctx->log, "sf_afs",
                sf_afs_kernel<true><<<afs_grid, 256, 0, st>>>(M, ctx->o_mut_site.p, ctx->q_own.p, ctx->o_mut_der.p,
                    ctx->o_mut_der_off.p, ctx->o_site_anc.p, ctx->o_site_anc_off.p, ctx->sidx.num_samples, ctx->q_afs.p));
        } else {
            MSPB_LAUNCH(ctx->log, "sf_afs",
                sf_afs_kernel<false><<<afs_grid, 256, 0, st>>>(M, ctx->o_mut_site.p, ctx->q_own.p, ctx->o_mut_der.p,
                    ctx->o_mut_der_off.p, ctx->o_site_anc.p, ctx->o_site_anc_off.p, ctx->sidx.num_samples, ctx->q_afs.p));
        }
    }
    CUDA_OK(cudaEventRecord(e1, st));
    CUDA_OK(cudaMemcpyAsync(afs, ctx->q_afs.p, afs_len * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
    CUDA_OK(cudaStreamSynchronize(st));
    CUDA_OK(cudaGetLastError());
    ctx->log.collect();
    if (device_ms != nullptr) {
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        *device_ms = ms;
    }
    return 0;
}

// ---------------------------------------------------------------- allocation-free entry point
// mspb_site_rows_noalloc (include/msprime_b200.h): the site numbering step (a variant of
// p1_sites_kernel for 32-bit position keys without kept rows) for callers that cannot call
// cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid it.
// The cross-tile look-back state, which the regular path keeps in a context buffer from
// cudaMalloc, lives here in a file-scope device array of fixed size: every GPU that loads the
// library holds these 256.3 KB until the process exits.  Every call on a device shares the
// array and finds the previous call's contents in it, so the part a call uses is cleared on
// the caller's stream before the kernel; two calls that overlap on one GPU (different
// streams, no ordering between them) race on it.  No cudaMalloc, cudaFree, cudaMallocHost or
// other allocating routine is called on this path.
namespace {

constexpr uint64_t NOALLOC_MAX_RECORDS = MSPB_NOALLOC_MAX_RECORDS;
constexpr uint64_t NOALLOC_MAX_TILES = (NOALLOC_MAX_RECORDS + P1_TILE - 1) / P1_TILE; // 16384
constexpr size_t NOALLOC_HEADER_BYTES = 256; // word 0: number of sites
constexpr size_t NOALLOC_SCRATCH_BYTES = NOALLOC_HEADER_BYTES + (NOALLOC_MAX_TILES + 1) * sizeof(LbWord);

// In HOST code this name is only a stand-in for the symbol, not a device address: the address
// each device gives it comes from cudaGetSymbolAddress (noalloc_scratch_address below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
void *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

// device address of the scratch array on the CURRENT device, looked up once per device
int noalloc_scratch_address(int device, unsigned char **out)
{
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    void *p = __atomic_load_n(&g_noalloc_ptr[device], __ATOMIC_ACQUIRE);
    if (p == nullptr) {
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr) {
            fprintf(stderr, "msprime_b200: cudaGetSymbolAddress failed on device %d: %s\n", device,
                cudaGetErrorString(e));
            cudaGetLastError();
            return MSPB_ERR_CUDA;
        }
        // (two threads racing here store the same address)
        __atomic_store_n(&g_noalloc_ptr[device], p, __ATOMIC_RELEASE);
    }
    *out = static_cast<unsigned char *>(p);
    return 0;
}

// Site row of every record of a position-sorted array of 32-bit keys: the number of distinct
// positions before it (p1_sites_kernel<KEY32_POS, false> without the work list, the draw
// numbering and the input checks).  Thread t owns 16 consecutive records; one block scan and
// one decoupled look-back across tiles whose state is `lb` (zeroed before the launch).
__global__ void __launch_bounds__(SCAN_THREADS, 4) site_rows_noalloc_kernel(const uint32_t *__restrict__ keys,
    uint64_t n, uint32_t *__restrict__ site_row, LbWord *lb, uint32_t num_tiles, uint64_t *num_sites)
{
    __shared__ uint32_t sm[9];
    __shared__ LbVal<1, false> bcast;
    const uint64_t tile_base = (uint64_t) blockIdx.x * P1_TILE;
    const uint64_t i0 = tile_base + (uint64_t) threadIdx.x * P1_ITEMS;
    const uint64_t last = n - 1;
    const bool full = tile_base + P1_TILE <= n;
    uint32_t pk[P1_ITEMS + 1]; // pk[k + 1]: key of record i0 + k
    if (full) {
        load16<uint32_t>(keys + i0, pk + 1);
    } else {
#pragma unroll
        for (int k = 0; k < P1_ITEMS; k++) {
            pk[k + 1] = keys[i0 + k < last ? i0 + k : last];
        }
    }
    pk[0] = keys[i0 > 0 ? (i0 - 1 < last ? i0 - 1 : last) : 0];
    uint32_t heads = 0, acc = 0; // bit k: record i0 + k opens a site
#pragma unroll
    for (int k = 0; k < P1_ITEMS; k++) {
        const uint64_t i = i0 + k;
        if (i < n && (i == 0 || pk[k + 1] != pk[k])) {
            heads |= 1u << k;
            acc++;
        }
    }
    uint32_t block_total;
    uint32_t run = block_exclusive_scan<uint32_t>(acc, sm, &block_total);
    LbVal<1, false> agg;
    agg.w[0][0] = block_total;
    agg.w[0][1] = agg.w[0][2] = agg.w[0][3] = 0;
    const LbVal<1, false> pw = lookback_prefix<1, false>(lb, blockIdx.x, agg, &bcast);
    if (blockIdx.x + 1 == num_tiles && threadIdx.x == 0) {
        *num_sites = (uint64_t) pw.w[0][0] + block_total;
    }
    uint32_t site[P1_ITEMS];
    run += pw.w[0][0];
#pragma unroll
    for (int k = 0; k < P1_ITEMS; k++) {
        run += (heads >> k) & 1u;
        site[k] = run - 1u; // inclusive count of heads: the record's own site row is count - 1
    }
    if (full) {
        store16(site_row + i0, site);
    } else {
#pragma unroll
        for (int k = 0; k < P1_ITEMS; k++) {
            if (i0 + k < n) {
                site_row[i0 + k] = site[k];
            }
        }
    }
}

} // namespace

extern "C" int mspb_site_rows_noalloc(int device, void *stream, const uint32_t *sorted_positions, uint64_t n,
    uint32_t *site_row, uint64_t *num_sites)
{
    if (num_sites == nullptr || (n > 0 && (sorted_positions == nullptr || site_row == nullptr))) {
        return MSPB_ERR_BAD_PARAM_VALUE;
    }
    if (n > NOALLOC_MAX_RECORDS) {
        return MSPB_ERR_NOALLOC_TOO_LARGE; // the scratch array is fixed: no larger input can be served
    }
    if (n > 0 && ((reinterpret_cast<uintptr_t>(sorted_positions) | reinterpret_cast<uintptr_t>(site_row)) & 15u) != 0) {
        return MSPB_ERR_BAD_PARAM_VALUE; // full tiles move 16 bytes at a time
    }
    int previous = -1;
    CUDA_OK(cudaGetDevice(&previous));
    if (previous != device) {
        CUDA_OK(cudaSetDevice(device));
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned char *scratch = nullptr;
    int ret = noalloc_scratch_address(device, &scratch); // set-up: once per device, that device current
    if (ret == 0) {
        if (n == 0) {
            ret = cudaMemsetAsync(num_sites, 0, sizeof(uint64_t), st) == cudaSuccess ? 0 : MSPB_ERR_CUDA;
        } else {
            const uint32_t tiles = (uint32_t) ((n + P1_TILE - 1) / P1_TILE);
            LbWord *lb = reinterpret_cast<LbWord *>(scratch + NOALLOC_HEADER_BYTES);
            // the look-back state must start at zero; the array keeps the last call's contents
            if (cudaMemsetAsync(lb, 0, ((size_t) tiles + 1) * sizeof(LbWord), st) != cudaSuccess) {
                ret = MSPB_ERR_CUDA;
            } else {
                site_rows_noalloc_kernel<<<tiles, SCAN_THREADS, 0, st>>>(sorted_positions, n, site_row, lb, tiles,
                    num_sites);
                ret = cudaPeekAtLastError() == cudaSuccess ? 0 : MSPB_ERR_CUDA;
            }
        }
        if (ret == MSPB_ERR_CUDA) {
            fprintf(stderr, "msprime_b200: mspb_site_rows_noalloc: %s\n", cudaGetErrorString(cudaGetLastError()));
        }
    }
    if (previous != device) {
        cudaSetDevice(previous);
    }
    return ret;
}
