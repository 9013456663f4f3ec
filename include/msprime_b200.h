/* msprime_b200 -- C ABI of the B200-native mutation overlay.
 *
 * This is the drop-in boundary for ONE path of msprime: the C call behind
 * msprime.sim_mutations.  Everything is extern "C", plain pointers and sizes, no
 * torch or CUDA types.  Each entry point names the reference interface it
 * replaces (paths relative to the msprime source tree):
 *
 *   mspb_sim_mutations      <- msprime_sim_mutations, msprime/_msprimemodule.c:3047-3120
 *                              (= mutgen_alloc + mutgen_set_rate_map +
 *                              mutgen_set_time_interval + mutgen_generate,
 *                              lib/mutgen.c:742-821,1304-1350)
 *   mspb_ctx_load_tables    <- LightweightTableCollection.fromdict, the table hand-off
 *                              in msprime/mutations.py:1686-1687
 *   mspb_generate           <- mutgen_generate, lib/mutgen.c:1304
 *   mspb_fetch              <- LightweightTableCollection.asdict, msprime/mutations.py:1699
 *   mspb_strerror           <- msp_strerror, lib/util.c:344-359
 *
 * Status codes follow lib/util.h:48-136 (MSP_ERR_*); tskit integrity errors are
 * folded in with bit 13 flipped exactly as msp_set_tsk_error does (lib/util.c:337).
 * Codes <= -1000 are specific to this implementation (CUDA failures, replay limits).
 *
 * Table columns use the tskit layout (tskit_lwt_interface.h): float64 positions and
 * times, int32 ids, ragged columns as a byte array plus an offset array of
 * num_rows + 1 entries.  Offsets are uint64 here (tsk_size_t, tskit core.h:111).
 */
#ifndef MSPRIME_B200_H
#define MSPRIME_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MSPB_VERSION "0.1.0"

/* msprime numbering (lib/util.h) */
#define MSPB_ERR_GENERIC (-1)
#define MSPB_ERR_NO_MEMORY (-2)
#define MSPB_ERR_BAD_PARAM_VALUE (-4)
#define MSPB_ERR_DUPLICATE_SITE_POSITION (-21)
#define MSPB_ERR_BAD_RATE_VALUE (-43)
#define MSPB_ERR_INCOMPATIBLE_MUTATION_MAP_LENGTH (-44)
#define MSPB_ERR_INSUFFICIENT_INTERVALS (-45)
#define MSPB_ERR_INTERVAL_MAP_START_NON_ZERO (-46)
#define MSPB_ERR_NONFINITE_INTERVAL_POSITION (-47)
#define MSPB_ERR_INTERVAL_POSITIONS_UNSORTED (-48)
#define MSPB_ERR_UNKNOWN_ALLELE (-51)
#define MSPB_ERR_INSUFFICIENT_ALLELES (-52)
#define MSPB_ERR_BAD_ROOT_PROBABILITIES (-53)
#define MSPB_ERR_BAD_TRANSITION_MATRIX (-54)
#define MSPB_ERR_BAD_SLIM_PARAMETERS (-56)
#define MSPB_ERR_MUTATION_ID_OVERFLOW (-57)
#define MSPB_ERR_UNKNOWN_TIME_NOT_SUPPORTED (-69)
/* implementation specific */
#define MSPB_ERR_CUDA (-1000)
#define MSPB_ERR_NO_DEVICE (-1001)
#define MSPB_ERR_TABLES_NOT_LOADED (-1002)
#define MSPB_ERR_REPLAY_STREAM_EXHAUSTED (-1003)
#define MSPB_ERR_REPLAY_REDRAW (-1004) /* stream would need a redraw (x >= hi or collision) */
#define MSPB_ERR_TOO_MANY_ROWS (-1005) /* more than 2^31 - 1 sites or mutations */
#define MSPB_ERR_UNSORTED_MUTATIONS (-1006) /* kept mutations not grouped by site */
#define MSPB_ERR_NOT_GENERATED (-1007)
#define MSPB_ERR_UNSORTED_SITES (-1009) /* sliced call: kept sites must be sorted by position */
#define MSPB_ERR_OFFSET_OVERFLOW (-1008) /* offsets32 asked for, but a column may reach 4 GiB */
#define MSPB_ERR_RETRY_WITH_HOST (-1011) /* mspb_generate_finish_device after mspb_generate_begin_device: some rank needs the host between the stages; make the call again with mspb_generate_begin */
#define MSPB_ERR_KEPT_MUTATIONS_UNSORTED (-1012) /* keep: existing mutations of a site are not listed oldest first and a new mutation falls on that site; the reference's within-site comparator (lib/mutgen.c:36-49) is not a total order there (its debug build asserts at lib/mutgen.c:1200) */
#define MSPB_ERR_NOALLOC_TOO_LARGE (-1010) /* mspb_site_rows_noalloc: more than MSPB_NOALLOC_MAX_RECORDS records */

/* the reference's three C models (lib/mutgen.c:184-237, 257-413, 431-481) */
enum { MSPB_MODEL_MATRIX = 0, MSPB_MODEL_INFINITE_ALLELES = 1, MSPB_MODEL_SLIM = 2 };

/* Input tables (host memory, borrowed for the duration of the call). */
typedef struct {
    double sequence_length;
    uint64_t num_nodes;
    const double *node_time;
    uint64_t num_edges;
    const double *edge_left;
    const double *edge_right;
    const int32_t *edge_parent;
    const int32_t *edge_child;
    /* existing sites / mutations; read only when keep != 0 */
    uint64_t num_sites;
    const double *site_position;
    const char *site_ancestral_state;
    const uint64_t *site_ancestral_state_offset;
    const char *site_metadata;
    const uint64_t *site_metadata_offset;
    uint64_t num_mutations;
    const int32_t *mutation_site;
    const int32_t *mutation_node;
    const double *mutation_time; /* NULL = all unknown */
    const char *mutation_derived_state;
    const uint64_t *mutation_derived_state_offset;
    const char *mutation_metadata;
    const uint64_t *mutation_metadata_offset;
} mspb_tables_t;

/* Call parameters: rate map, model, flags, window, random source, genome slice. */
typedef struct {
    /* rate map: map_size intervals, map_size + 1 positions (lib/rate_map.c:45) */
    uint64_t map_size;
    const double *map_position;
    const double *map_rate;
    /* model (lib/msprime.h:440-470) */
    int32_t model_kind;
    uint64_t num_alleles;
    const char *alleles;
    const uint64_t *allele_offset; /* num_alleles + 1 */
    const double *root_distribution; /* num_alleles */
    const double *transition_matrix; /* num_alleles^2, row major */
    uint64_t start_allele; /* infinite alleles: next_allele on entry; SLiM: next_mutation_id (int64 >= 0) */
    /* flags and window (MSP_DISCRETE_SITES / MSP_KEEP_SITES, lib/msprime.h:66-68) */
    int32_t discrete_genome;
    int32_t keep;
    double start_time;
    double end_time;
    /* random source: the native stream is Philox4x32-10 keyed by seed; a non-NULL
     * replay_stream switches to the replay harness, where uniforms are consumed
     * from the array in the reference's order (single slice only). */
    uint64_t seed;
    const double *replay_stream;
    uint64_t replay_len;
    /* genome slice [slice_left, slice_right) owned by this call; 0,0 = whole genome.
     * The native stream is keyed by (edge, ordinal), so the union of the slices'
     * outputs is identical to the single-slice output. */
    double slice_left;
    double slice_right;
    /* nonzero: the offset columns of the result are uint32 instead of uint64, as tskit's
     * dict protocol exports them when they fit (tskit_lwt_interface.h:1634-1635) */
    int32_t offsets32;
    /* How a genome slice is drawn.  0: every edge is drawn whole and the mutations outside
     * the slice are dropped, so the union of the slices is bit-identical to the unsliced
     * call (every rank pays for all edges).  Nonzero: edges are cut to the slice and drawn on
     * this Philox stream (use rank + 1), so work scales with the slice; the result then
     * depends on the cuts, with the same distribution. */
    uint32_t slice_stream;
    /* SLiM model (slim_mutation_model_alloc, lib/mutgen.c:552-589): new sites have the empty
     * ancestral state; a new mutation's derived state is its parent's state, a comma and the
     * next mutation id, its metadata the parent's metadata plus a 17-byte MutationMetadataRec
     * with origin generation = slim_generation - (int32) time.  Single-call only (not with
     * mspb_generate_begin). */
    int32_t slim_mutation_type_id;
    int32_t slim_generation;
} mspb_params_t;

/* Row and byte counts of the result of the last mspb_generate. */
typedef struct {
    uint64_t num_sites;
    uint64_t num_mutations;
    uint64_t num_new_mutations;
    uint64_t ancestral_state_bytes;
    uint64_t site_metadata_bytes;
    uint64_t derived_state_bytes;
    uint64_t mutation_metadata_bytes;
    uint64_t next_allele; /* infinite alleles: counter after the call */
    uint64_t draws; /* replay: uniforms consumed */
    uint64_t num_new_sites;
    double device_ms; /* CUDA-event time of the generate step */
    uint64_t offset_bytes; /* width of one offset entry in the result: 8, or 4 with offsets32 */
} mspb_counts_t;

/* Output columns; caller-allocated to the sizes in mspb_counts_t (offset arrays
 * are num_rows + 1 entries of counts.offset_bytes bytes: uint64, or uint32 with
 * params.offsets32).  Host pointers for mspb_fetch; mspb_device_columns fills the
 * same struct with device pointers owned by the context. */
typedef struct {
    double *site_position;
    char *site_ancestral_state;
    void *site_ancestral_state_offset;
    char *site_metadata;
    void *site_metadata_offset;
    int32_t *mutation_site;
    int32_t *mutation_node;
    int32_t *mutation_parent;
    double *mutation_time;
    char *mutation_derived_state;
    void *mutation_derived_state_offset;
    char *mutation_metadata;
    void *mutation_metadata_offset;
} mspb_columns_t;

typedef struct mspb_ctx mspb_ctx_t;

/* Context on CUDA device `device`; fails with MSPB_ERR_NO_DEVICE when there is no
 * usable GPU (there is no CPU fallback). */
int mspb_ctx_create(int device, mspb_ctx_t **ctx);
void mspb_ctx_destroy(mspb_ctx_t *ctx);
/* Validates (tskit check_integrity(0) subset, lib/mutgen.c:1316; and, for the existing site and
 * mutation rows, what tsk_*_table_set_columns checks before the reference ever sees them: offset
 * columns start at 0 and never decrease, else TSK_ERR_BAD_OFFSET with bit 13 flipped), uploads the
 * tables and builds the child-sorted edge index.  Reusable across mspb_generate calls. */
int mspb_ctx_load_tables(mspb_ctx_t *ctx, const mspb_tables_t *tables);

/* The same with the upload shared between the `world` GPUs of a genome-sharded job (the tables
 * are replicated: every rank needs all of them).  _begin copies only this rank's 1 / world
 * share of the five large columns (node time, edge left / right / parent / child; kept rows
 * are copied whole) and describes the device buffers: buffer k is world * bytes_per_rank
 * bytes long and this rank's share sits at rank * bytes_per_rank.  The caller all-gathers
 * every buffer IN PLACE on the context's stream (ncclAllGather / torch.distributed
 * .all_gather_into_tensor over NVLink) and then calls _end, which validates the rows and
 * builds the index.  With world == 1 the pair is equivalent to mspb_ctx_load_tables. */
typedef struct {
    void *ptr; /* device memory owned by the context */
    uint64_t bytes_per_rank;
} mspb_shard_t;
#define MSPB_NUM_TABLE_SHARDS 5
int mspb_ctx_load_tables_begin(mspb_ctx_t *ctx, const mspb_tables_t *tables, uint32_t rank, uint32_t world,
    mspb_shard_t shards[MSPB_NUM_TABLE_SHARDS]);
int mspb_ctx_load_tables_end(mspb_ctx_t *ctx);
/* Runs the overlay on the device; the result stays resident until the next call. */
int mspb_generate(mspb_ctx_t *ctx, const mspb_params_t *params, mspb_counts_t *counts);
/* Two-phase form for genome-sharded calls (one context per GPU, params.slice_left/right
 * set): _begin runs everything up to the row numbering and reports this slice's
 * num_sites, num_mutations and next_allele - start_allele (allele numbers used); the
 * caller exchanges those between ranks; _end writes the tables with the slice's global
 * site-id, mutation-id and allele-number bases.  Byte offsets stay slice-local until
 * mspb_shift_offsets adds the bases obtained from a second exchange of the byte counts. */
int mspb_generate_begin(mspb_ctx_t *ctx, const mspb_params_t *params, mspb_counts_t *counts);
int mspb_generate_end(mspb_ctx_t *ctx, uint64_t site_base, uint64_t mutation_base, uint64_t allele_base,
    mspb_counts_t *counts);
int mspb_shift_offsets(mspb_ctx_t *ctx, uint64_t ancestral_state_base, uint64_t site_metadata_base,
    uint64_t derived_state_base, uint64_t mutation_metadata_base);
/* The same exchange without host round trips (the two small
 * all-gathers): the counts stay on the device.  `send` (MSPB_EXCHANGE_WORDS uint64) and
 * `recv` (world x MSPB_EXCHANGE_WORDS uint64, rank-major) are device buffers owned by the
 * context; the caller all-gathers send -> recv on the context's stream (NCCL through
 * torch.distributed, or a device copy) where marked:
 *     mspb_generate_begin; mspb_exchange_post_counts;      <all-gather>
 *     mspb_generate_end_device;                             <all-gather>
 *     mspb_generate_finish_device                           (the only wait)
 * _end_device derives the slice's site / mutation / allele-number bases from recv on the
 * device and writes the tables with them; _finish_device shifts the byte offsets the same
 * way and returns this rank's counts, the global totals[8] and this rank's bases[8], both
 * ordered (sites, mutations, allele numbers used, new mutations, ancestral-state bytes,
 * site-metadata bytes, derived-state bytes, mutation-metadata bytes). */
#define MSPB_EXCHANGE_WORDS 8
#define MSPB_EXCHANGE_MAX_WORLD 64
int mspb_exchange_buffers(mspb_ctx_t *ctx, void **send, void **recv);
int mspb_exchange_post_counts(mspb_ctx_t *ctx);
int mspb_generate_end_device(mspb_ctx_t *ctx, uint32_t rank, uint32_t world);
int mspb_generate_finish_device(mspb_ctx_t *ctx, uint32_t rank, uint32_t world, mspb_counts_t *counts,
    uint64_t *totals, uint64_t *bases);
/* mspb_generate_begin without its closing wait (no counterpart in the reference: lib/mutgen.c has one address space):
 * K1 .. K4 are enqueued and the call returns; mspb_exchange_post_counts then takes this rank's counts from the
 * device, so that the host waits once per overlay, in mspb_generate_finish_device, instead of once per stage.
 * What the host would have looked at between the stages (an error, the second sort of a dense genome, positions
 * to draw again) is exchanged with the counts: when any rank raises it, every rank's _finish_device returns
 * MSPB_ERR_RETRY_WITH_HOST and the caller makes the call again through mspb_generate_begin.  Not for the replay
 * harness or the SLiM model (MSPB_ERR_BAD_PARAM_VALUE); mspb_generate_end refuses a call begun this way. */
int mspb_generate_begin_device(mspb_ctx_t *ctx, const mspb_params_t *params);
/* The stream (cudaStream_t) the caller's next all-gather of send -> recv goes on.  Normally the context's own
 * (mspb_ctx_stream).  After mspb_generate_begin_device of a call without kept rows under the infinite-alleles model
 * or a matrix model with equally long alleles it is the context's second stream: the counts are final once the
 * sites are numbered and the byte counts follow from them in closed form, so both exchanges run BESIDE the parent
 * search (K4), the table writer (K5) starts with every base known, and no offset is shifted afterwards. */
void *mspb_exchange_stream(const mspb_ctx_t *ctx);
/* Copies the resident result to host buffers. */
int mspb_fetch(mspb_ctx_t *ctx, const mspb_columns_t *host_columns);
/* Bytes the last mspb_fetch moved over the bus (columns known to be all zero, i.e. the metadata
 * offsets of a result without metadata, are filled in by the host instead of being copied). */
uint64_t mspb_last_fetch_bytes(const mspb_ctx_t *ctx);
/* Device pointers of the resident result (valid until the next mspb_generate). */
int mspb_device_columns(mspb_ctx_t *ctx, mspb_columns_t *device_columns);
/* Which placement kernel (K2) the next calls use: 0 = chosen from the mutations per edge (the default: one
 * block per 256 edges when edges carry a mutation or more each, one thread per mutation when they are mostly
 * empty, as in the clipped slices of a sharded call), 1 = always the former, 2 = always the latter.  Both write
 * the same records; the switch exists for tests and measurements.  No counterpart in the reference. */
int mspb_ctx_force_placement(mspb_ctx_t *ctx, int mode);
/* Number of kernels this library launched since the context was created. */
uint64_t mspb_kernel_launches(const mspb_ctx_t *ctx);
/* The CUDA stream (cudaStream_t) every kernel and copy of this context runs on, so
 * a caller can bracket calls with its own events on that stream. */
void *mspb_ctx_stream(const mspb_ctx_t *ctx);
/* Per-kernel timing: enable != 0 brackets every launch with CUDA events from now on
 * (and clears the accumulated table); mspb_kernel_time reads entry `index` of the
 * table (kernel name, launches, summed milliseconds) and fails past the end. */
int mspb_kernel_timing(mspb_ctx_t *ctx, int enable);
int mspb_kernel_time(const mspb_ctx_t *ctx, int index, const char **name, uint64_t *launches, double *ms);
/* Per-phase device milliseconds of the last mspb_generate (n <= MSPB_NUM_PHASES); recorded while
 * mspb_kernel_timing is enabled, zeros otherwise. */
#define MSPB_NUM_PHASES 8
int mspb_phase_ms(const mspb_ctx_t *ctx, float *ms, int n);
const char *mspb_phase_name(int phase);

/* Diagnostic: sorts n pseudo-random keys (key_bytes 4 or 8, masked with key_mask) with
 * the library's device radix sort and checks order, stability and payload on the host.
 * measure_bits != 0 lets the sort measure the varying bits itself.  0 = pass. */
int mspb_selftest_sort(mspb_ctx_t *ctx, uint64_t n, int key_bytes, uint64_t key_mask, uint64_t seed,
    int measure_bits, int *passes);

/* One-shot form with the semantics of _msprime.sim_mutations: host tables in, library-owned host columns
 * out (page-locked blocks of the library's pool where pinning succeeds, malloc otherwise: release them with
 * mspb_columns_free, never with free).  Calls on one device are serialised and share one context, which stays
 * alive between calls with its device buffers (a bench-size overlay holds several GB; allocating them per call
 * took most of the call): mspb_oneshot_release(device) frees it (device < 0: every device).  Tables in pageable
 * memory are staged through the pinned pool by a few host threads (mspb_ctx_load_tables does the same). */
int mspb_sim_mutations(int device, const mspb_tables_t *tables, const mspb_params_t *params,
    mspb_counts_t *counts, mspb_columns_t *out);
void mspb_oneshot_release(int device);
void mspb_columns_free(mspb_columns_t *cols);

/* Pinned host memory from a pool (no reference counterpart): result columns copied into pinned buffers arrive at
 * the link's rate, into pageable ones at a fraction of it.  Blocks are pinned once and reused after
 * mspb_pinned_free; NULL when the allocation fails.  Thread safe. */
void *mspb_pinned_alloc(uint64_t bytes);
void mspb_pinned_free(void *block);
/* pool counters: blocks pinned, blocks reused, bytes pinned so far, blocks unpinned again */
void mspb_pinned_stats(uint64_t out[4]);

/* Consumers of the resident tables on the device: what the reference's users compute
 * next from the tables this path writes (verification.py:6008-6013).
 *   mspb_ctx_build_sample_index  <- the num_samples bookkeeping of tskit's tree sweep (tsk_tree_t.num_samples,
 *                                   git-submodules/tskit/c/tskit/trees.c), as per-node step functions over the
 *                                   genome, built once per loaded tree sequence.  node_flags: tskit nodes.flags
 *                                   (host, num_nodes entries; bit 0 = TSK_NODE_IS_SAMPLE).
 *   mspb_samples_below           <- tsk_tree_get_num_samples(tree covering the site, mutation's node) for every
 *                                   mutation row of the result of the last mspb_generate (host, num_mutations).
 *   mspb_site_afs                <- tsk_treeseq_allele_frequency_spectrum with TSK_STAT_SITE | TSK_STAT_POLARISED,
 *                                   one sample set (all samples), no span normalisation (trees.c:3816, 3527-3540):
 *                                   afs[k] = derived alleles (byte strings other than the site's ancestral state)
 *                                   carried by exactly k samples, 0 < k < num_samples; afs_len = num_samples + 1.
 *   mspb_decode_genotypes        <- tsk_variant_decode (git-submodules/tskit/c/tskit/genotypes.c:474) for the sites
 *                                   [site_begin, site_end) of the result: genotypes[(site - site_begin) * num_samples
 *                                   + j] = allele index of sample j (sample nodes in increasing order; allele 0 is
 *                                   the ancestral state, derived states follow in order of first appearance);
 *                                   host buffer of (site_end - site_begin) * num_samples int32; at most 65535 sites
 *                                   per call.
 * Results of unsliced calls only (a genome slice of a sharded call carries global row ids). */
int mspb_ctx_build_sample_index(mspb_ctx_t *ctx, const uint32_t *node_flags, uint64_t *num_samples,
    uint64_t *num_levels, uint64_t *num_steps, double *build_ms);
int mspb_samples_below(mspb_ctx_t *ctx, uint32_t *below);
int mspb_site_afs(mspb_ctx_t *ctx, uint64_t *afs, uint64_t afs_len, double *device_ms);
int mspb_decode_genotypes(mspb_ctx_t *ctx, uint64_t site_begin, uint64_t site_end, int32_t *genotypes,
    double *device_ms);

/* Allocation-free site numbering.  This entry point has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  It is the site-numbering step of the overlay
 * (the kernel behind mspb_generate that turns position-sorted records into site rows) on caller-owned
 * device memory: sorted_positions[n] (uint32 integer positions in non-decreasing order) in, site_row[n]
 * (the number of distinct positions before each record = the `site` column of the mutation table) and
 * *num_sites out; all three are DEVICE pointers, the arrays 16-byte aligned.  The work is enqueued on
 * `stream` (a cudaStream_t, NULL = the default stream) of CUDA device `device` and the call returns
 * without waiting.  It calls no cudaMalloc / cudaFree / cudaMallocHost and no allocating library routine:
 * the kernel's cross-tile scratch is a file-scope __device__ array of 256.3 KB that every GPU loading the
 * library holds until exit, which bounds the input: n > MSPB_NOALLOC_MAX_RECORDS (2^26) is refused with
 * MSPB_ERR_NOALLOC_TOO_LARGE.  All calls on a device share that array and it keeps the last call's
 * contents (the call clears what it uses, on `stream`, first): calls that overlap on one GPU race; order
 * them on one stream or with events.  The first call on a device looks the array's address up
 * (cudaGetSymbolAddress) and should be made outside stream capture. */
#define MSPB_NOALLOC_MAX_RECORDS (1ull << 26)
int mspb_site_rows_noalloc(int device, void *stream, const uint32_t *sorted_positions, uint64_t n,
    uint32_t *site_row, uint64_t *num_sites);

const char *mspb_strerror(int err);
const char *mspb_version(void);
/* CUDA runtime the library was built against, as cudaRuntimeGetVersion reports it (12090 = 12.9); works
 * without a device.  Recorded in provenance where the reference records its GSL version
 * (msprime/provenance.py:66-69). */
int mspb_cuda_runtime_version(void);

#ifdef __cplusplus
}
#endif
#endif
