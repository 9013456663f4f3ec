// msprime_b200: the mutation-overlay kernels (sm_100a).
//
// Pipeline (DESIGN.md has the data layout and the byte counts):
//   K0  child-sorted edge interval index              (once per tree sequence)
//   K1  per-edge Poisson count, scanned in the same   <- lib/mutgen.c:1103-1120
//       kernel by decoupled look-back
//   K2  position / time placement, one thread per     <- lib/mutgen.c:1121-1151
//       mutation (edges staged in shared memory)
//   K3  stable radix sort by position (sort.cuh)      <- AVL map, lib/mutgen.c:858-883
//   P1  site boundaries + row / draw numbering        <- site order, lib/mutgen.c:1027-1074
//   K4  sites with several records: time order,       <- lib/mutgen.c:36-95, 1160-1224
//       parent search, allele chain
//   K5  byte offsets (look-back) + table emission     <- lib/mutgen.c:1027-1074
//
// Records: one per kept site ("marker"), kept mutation and new mutation, in that
// slot order, so a record's kind follows from its slot index and slot order is the
// final tie-break (kept before new, kept by row id, new by creation order).
// A record is 16 bytes (Rec: time, node, ordinal); its position lives in the sort key.
#pragma once
#include "common.cuh"
#include "replay.cuh"
#include "sort.cuh"

namespace mspb {

// error word: first error wins
MSPB_DEVFN void set_error(int *err, int code)
{
    atomicCAS(err, 0, code);
}

struct TreeSeqView {
    uint64_t num_nodes, num_edges;
    double sequence_length;
    const double *node_time;
    const double *edge_left, *edge_right;
    const int32_t *edge_parent, *edge_child;
    // child-sorted interval index (K0): the edges of child c are ci[child_start[c] .. child_start[c+1]),
    // one 32-byte entry per edge; an entry also names the range of its parent's entries, so a
    // step of the parent walk is a binary search of 32-byte entries and nothing else
    const uint32_t *child_start;
    const struct ChildIv *ci;
};

struct __align__(32) ChildIv {
    double left, right;
    double parent_time;
    // The parent.  parent_id identifies the node: child_start[p] when p has entries of its own,
    // CI_NODE | p when it is a child nowhere.  parent_span names the parent's entries that
    // OVERLAP [left, right) (k0_link_kernel): (first - child_start[p]) << 8 | count, so a step
    // of the walk usually probes one entry; CI_SPAN_UNKNOWN when that does not fit (the walk
    // then recovers the parent's full range from child_start).
    uint32_t parent_id, parent_span;
};
constexpr uint32_t CI_NODE = 0x80000000u;
constexpr uint32_t CI_SPAN_UNKNOWN = 0xffffffffu;

struct KeptView {
    uint64_t num_sites, num_mutations; // 0 when keep is off
    int32_t site_id_offset; // input id of site 0 of this view (genome-sliced calls see a sub-range)
    const double *site_position;
    const uint64_t *site_anc_off, *site_md_off;
    const char *site_anc, *site_md;
    const int32_t *mut_site, *mut_node;
    const double *mut_time;
    const uint64_t *mut_der_off, *mut_md_off;
    const char *mut_der, *mut_md;
    const int32_t *site_allele; // model allele index of the ancestral state, -1 unknown
    const int32_t *mut_allele; // model allele index of the derived state, -1 unknown
};

struct ModelView {
    int kind;
    uint32_t num_alleles;
    const char *alleles;
    const uint64_t *allele_off;
    const double *root; // num_alleles
    const double *matrix; // num_alleles^2
    uint64_t start_allele; // infinite alleles: next allele; SLiM: next mutation id
    int32_t slim_type, slim_generation; // SLiM model (kind 2), lib/mutgen.c:552-589
};
constexpr int SLIM_MD_SIZE = 17; // MutationMetadataRec, lib/mutgen.c:262-290

// How positions are packed into sort keys.
enum KeyMode : int {
    KEY32_POS = 0, // uint32 key = integer position (discrete genome below 2^32)
    KEY64_POS = 1, // uint64 key = ordered_bits(position): exact, any genome
    KEY64_POS_TIME = 2 // uint64 key = integer position << 32 | coarse descending time
};

struct CallView {
    double start_time, end_time;
    double slice_left, slice_right;
    int discrete;
    int sliced; // invariant slicing: draw every edge whole, keep what falls inside the slice
    int clipped; // clipped slicing: edges are cut to the slice and drawn on stream `stream`
    uint32_t stream; // Philox stream of a clipped slice (0 otherwise)
    uint64_t seed;
    int key_mode;
    // clipped slices: K1 and K2 run over the edges that overlap the slice only
    // (edge_list[0 .. num_work), ascending edge ids); otherwise over all num_work = E edges
    const uint32_t *edge_list;
    uint64_t num_work;
};

// keys of the records, before or after the sort
struct KeyView {
    uint32_t *k32;
    uint64_t *k64;
    int mode;
    MSPB_DEVFN void store(uint64_t i, double pos, double time, bool marker) const
    {
        if (mode == KEY32_POS) {
            k32[i] = (uint32_t) pos;
        } else if (mode == KEY64_POS) {
            k64[i] = ordered_bits(pos);
        } else {
            const uint64_t t = marker ? 0ull : ((~ordered_bits(time)) >> 32);
            k64[i] = ((uint64_t) pos << 32) | t;
        }
    }
    MSPB_DEVFN bool same_pos(uint64_t i, uint64_t j) const
    {
        if (mode == KEY32_POS) {
            return k32[i] == k32[j];
        }
        const uint64_t m = mode == KEY64_POS ? ~0ull : 0xffffffff00000000ull;
        return ((k64[i] ^ k64[j]) & m) == 0;
    }
    MSPB_DEVFN bool same_key(uint64_t i, uint64_t j) const
    {
        return mode == KEY32_POS ? k32[i] == k32[j] : k64[i] == k64[j];
    }
    MSPB_DEVFN double position(uint64_t i) const
    {
        if (mode == KEY32_POS) {
            return (double) k32[i];
        }
        if (mode == KEY64_POS) {
            return from_ordered_bits(k64[i]);
        }
        return (double) (k64[i] >> 32);
    }
};

// ---------------------------------------------------------------- K0
__global__ void k0_child_keys_kernel(TreeSeqView ts, uint64_t *key_left, uint32_t *val)
{
    const uint64_t e = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (e < ts.num_edges) {
        key_left[e] = ordered_bits(ts.edge_left[e]);
        val[e] = (uint32_t) e;
    }
}

__global__ void k0_child_keys2_kernel(TreeSeqView ts, const uint32_t *val, uint64_t *key_child)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < ts.num_edges) {
        key_child[i] = (uint64_t) (uint32_t) ts.edge_child[val[i]];
    }
}

__global__ void k0_permute_kernel(TreeSeqView ts, const uint32_t *perm, ChildIv *ci, uint32_t *child_start)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < ts.num_edges) {
        const uint32_t e = perm[i];
        ChildIv iv;
        iv.left = ts.edge_left[e];
        iv.right = ts.edge_right[e];
        const uint32_t c = (uint32_t) ts.edge_child[e];
        const uint32_t p = (uint32_t) ts.edge_parent[e];
        iv.parent_time = ts.node_time[p];
        iv.parent_id = p; // k0_link_kernel replaces the pair by (identity, overlapping entries)
        iv.parent_span = 0;
        ci[i] = iv;
        // CSR starts: child c begins at the first i whose child is >= c
        const uint32_t prev = i == 0 ? 0u : (uint32_t) ts.edge_child[perm[i - 1]] + 1u;
        for (uint32_t k = (i == 0 ? 0u : prev); k <= c; k++) {
            child_start[k] = (uint32_t) i;
        }
        if (i == ts.num_edges - 1) {
            for (uint64_t k = (uint64_t) c + 1; k <= ts.num_nodes; k++) {
                child_start[k] = (uint32_t) ts.num_edges;
            }
        }
    }
}

// max_count: spans with more overlapping entries are stored as CI_SPAN_UNKNOWN (256 in
// production; tests lower it to drive the recovery path of the walk)
__global__ void k0_link_kernel(TreeSeqView ts, ChildIv *ci, uint32_t max_count)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < ts.num_edges) {
        const uint32_t p = ci[i].parent_id;
        const uint32_t a = ts.child_start[p], b = ts.child_start[p + 1];
        if (a >= b) {
            ci[i].parent_id = CI_NODE | p;
            return;
        }
        const double l = ci[i].left, r = ci[i].right;
        // first entry of p that ends after l, first entry of p that starts at or after r
        uint32_t lo = a, hi = b;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if (ci[mid].right <= l) {
                lo = mid + 1;
            } else {
                hi = mid;
            }
        }
        const uint32_t f = lo;
        hi = b;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if (ci[mid].left < r) {
                lo = mid + 1;
            } else {
                hi = mid;
            }
        }
        const uint32_t count = lo - f;
        ci[i].parent_id = a;
        ci[i].parent_span = (f - a < 0xffffffu && count < max_count) ? ((f - a) << 8 | count) : CI_SPAN_UNKNOWN;
    }
}

// a node as the parent walk names it: its identity and the range of its entries that can
// cover the walk's position
struct WalkNode {
    uint32_t id, lo, hi;
};

MSPB_DEVFN uint32_t walk_id(const TreeSeqView &ts, int32_t u)
{
    const uint32_t a = ts.child_start[u], b = ts.child_start[u + 1];
    return a < b ? a : (CI_NODE | (uint32_t) u);
}

MSPB_DEVFN WalkNode walk_node(const TreeSeqView &ts, int32_t u)
{
    WalkNode w;
    const uint32_t a = ts.child_start[u], b = ts.child_start[u + 1];
    w.id = a < b ? a : (CI_NODE | (uint32_t) u);
    w.lo = a;
    w.hi = b;
    return w;
}

// Steps from node w to its parent at position x in the local tree: false when w is a root
// there.  Intervals of one child are disjoint and sorted.
MSPB_DEVFN bool parent_at(const TreeSeqView &ts, WalkNode *w, double x, double *time)
{
    uint32_t lo = w->lo, hi = w->hi;
    ChildIv iv;
    bool found = false;
    if (hi - lo == 2) {
        // both candidates in one round trip (they share a 64-byte segment or are adjacent)
        const ChildIv i0 = ts.ci[lo], i1 = ts.ci[lo + 1];
        iv = x < i0.right ? i0 : i1;
        found = x >= iv.left && x < iv.right;
    } else {
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            iv = ts.ci[mid];
            if (x < iv.left) {
                hi = mid;
            } else if (x >= iv.right) {
                lo = mid + 1;
            } else {
                found = true;
                break;
            }
        }
    }
    if (!found) {
        return false;
    }
    *time = iv.parent_time;
    w->id = iv.parent_id;
    if (iv.parent_id & CI_NODE) {
        w->lo = w->hi = 0;
    } else if (iv.parent_span != CI_SPAN_UNKNOWN) {
        w->lo = iv.parent_id + (iv.parent_span >> 8);
        w->hi = w->lo + (iv.parent_span & 0xffu);
    } else {
        // the node whose entries start at parent_id: the last u with child_start[u] <= parent_id
        uint64_t a = 0, b = ts.num_nodes; // invariant: child_start[a] <= id < child_start[b]
        while (b - a > 1) {
            const uint64_t mid = (a + b) >> 1;
            if (ts.child_start[mid] <= iv.parent_id) {
                a = mid;
            } else {
                b = mid;
            }
        }
        w->lo = iv.parent_id;
        w->hi = ts.child_start[a + 1];
    }
    return true;
}


// ---------------------------------------------------------------- K1 (native)
struct EdgeDraw {
    double branch_start, branch_end, mass_lo, mass, left, right;
    int32_t child;
};

MSPB_DEVFN EdgeDraw edge_draw(const TreeSeqView &ts, const RateMapView &map, const CallView &call, uint64_t e)
{
    EdgeDraw d;
    d.left = ts.edge_left[e];
    d.right = ts.edge_right[e];
    if (call.clipped) {
        d.left = fmax(d.left, call.slice_left);
        d.right = fmin(d.right, call.slice_right);
        if (!(d.left < d.right)) {
            // outside this call's slice: nothing to draw, skip the node-time gathers
            d.child = 0;
            d.branch_start = d.branch_end = d.mass_lo = d.mass = 0.0;
            return d;
        }
    }
    d.child = ts.edge_child[e];
    const double tc = ts.node_time[d.child];
    const double tp = ts.node_time[ts.edge_parent[e]];
    d.branch_start = fmax(call.start_time, tc);
    d.branch_end = fmin(call.end_time, tp);
    d.mass_lo = map.mass_upto(d.left);
    d.mass = d.left < d.right ? map.mass_upto(d.right) - d.mass_lo : 0.0;
    return d;
}

MSPB_DEVFN void place_one(const Philox &ph, const RateMapView &map, const CallView &call, const EdgeDraw &d,
    uint64_t e, uint32_t k, double *pos, double *time)
{
    const uint4 r = ph((uint32_t) e, call.stream, k, TAG_PLACE);
    const double up = u53(r.x, r.y), ut = u53(r.z, r.w);
    *pos = map.position_at(d.mass_lo + up * d.mass, d.left, d.right);
    double t = d.branch_start + ut * (d.branch_end - d.branch_start);
    if (!(t < d.branch_end)) {
        t = nextafter(d.branch_end, -1.7976931348623157e308);
    }
    *time = t;
}

// how many of an edge's n mutations fall inside the call's genome slice
MSPB_DEVFN uint32_t k1_in_slice(const Philox &ph, const RateMapView &map, const CallView &call, const EdgeDraw &d,
    uint64_t e, uint32_t n)
{
    uint32_t in = 0;
    if (call.sliced && n > 0 && d.left < call.slice_right && d.right > call.slice_left) {
        for (uint32_t q = 0; q < n; q++) {
            double x, t;
            place_one(ph, map, call, d, e, q, &x, &t);
            in += (x >= call.slice_left && x < call.slice_right) ? 1u : 0u;
        }
    }
    return in;
}

constexpr int K1_ITEMS = 4;
constexpr int K1_TILE = SCAN_THREADS * K1_ITEMS;
constexpr int K1_TILE_SHIFT = 10;
static_assert((1 << K1_TILE_SHIFT) == K1_TILE, "K1 tile size");

// the rest of edge_draw once the edge's columns and node times are in registers
MSPB_DEVFN EdgeDraw edge_draw_from(const RateMapView &map, const CallView &call, double left, double right,
    int32_t child, double tc, double tp)
{
    EdgeDraw d;
    d.left = left;
    d.right = right;
    d.child = child;
    d.branch_start = fmax(call.start_time, tc);
    d.branch_end = fmin(call.end_time, tp);
    d.mass_lo = map.mass_upto(d.left);
    d.mass = d.left < d.right ? map.mass_upto(d.right) - d.mass_lo : 0.0;
    return d;
}

// count[e] = Poisson(branch length x rate mass), and how many of those fall in this
// call's genome slice (low / high half of one 64-bit count).  Written as offsets INSIDE
// the tile (goff[e], loff[e]) plus one total per tile: tiles are independent, the tile
// bases come from k1_spine_kernel, and K2 adds them when it reads.
__global__ void __launch_bounds__(SCAN_THREADS, 4) k1_count_kernel(TreeSeqView ts, RateMapView map, CallView call,
    uint32_t *__restrict__ goff, uint32_t *__restrict__ loff, uint64_t *__restrict__ tile_tot, int *err)
{
    __shared__ uint64_t sm[9];
    __shared__ uint64_t s_c[K1_TILE];
    __shared__ uint16_t s_big[K1_TILE]; // edges whose mean needs the rejection sampler
    __shared__ uint32_t s_nbig;
    __shared__ unsigned long long s_exact; // the tile's count in 64 bits (only summed when it may be huge)
    const uint64_t tile_base = (uint64_t) blockIdx.x * K1_TILE;
    const uint64_t last = call.num_work - 1;
    const Philox ph(call.seed);
    if (threadIdx.x == 0) {
        s_nbig = 0;
        s_exact = 0ull;
    }
    __syncthreads();
    // striped: consecutive threads take consecutive edges (coalesced column reads).  All
    // column loads of the thread's edges are issued first, then all node-time gathers.
    double mu[K1_ITEMS];
    uint32_t eid[K1_ITEMS]; // edge ids of the thread's work items
    {
        double left[K1_ITEMS], right[K1_ITEMS], tc[K1_ITEMS], tp[K1_ITEMS];
        int32_t child[K1_ITEMS], parent[K1_ITEMS];
#pragma unroll
        for (int k = 0; k < K1_ITEMS; k++) {
            const uint64_t w = tile_base + (uint32_t) k * SCAN_THREADS + threadIdx.x;
            const uint64_t wc = w < last ? w : last;
            const uint64_t ec = call.edge_list != nullptr ? call.edge_list[wc] : wc;
            eid[k] = (uint32_t) ec;
            left[k] = ts.edge_left[ec];
            right[k] = ts.edge_right[ec];
            child[k] = ts.edge_child[ec];
            parent[k] = ts.edge_parent[ec];
        }
#pragma unroll
        for (int k = 0; k < K1_ITEMS; k++) {
            if (call.clipped) {
                left[k] = fmax(left[k], call.slice_left);
                right[k] = fmin(right[k], call.slice_right);
            }
            // outside this call's slice: nothing to draw, skip the node-time gathers
            const bool inside = !call.clipped || left[k] < right[k];
            tc[k] = inside ? ts.node_time[child[k]] : 0.0;
            tp[k] = inside ? ts.node_time[parent[k]] : 0.0;
        }
#pragma unroll
        for (int k = 0; k < K1_ITEMS; k++) {
            const uint64_t w = tile_base + (uint32_t) k * SCAN_THREADS + threadIdx.x;
            mu[k] = 0.0;
            if (w <= last && left[k] < right[k]) {
                const EdgeDraw d = edge_draw_from(map, call, left[k], right[k], child[k], tc[k], tp[k]);
                mu[k] = (d.branch_end - d.branch_start) * d.mass;
            }
        }
    }
    // the common small means are drawn here by inversion; the rare large ones are set
    // aside so that their long rejection loop runs on full warps afterwards
#pragma unroll
    for (int k = 0; k < K1_ITEMS; k++) {
        const uint32_t j = (uint32_t) k * SCAN_THREADS + threadIdx.x;
        const uint64_t e = eid[k];
        uint64_t c = 0;
        if (mu[k] >= POISSON_SMALL_MU) {
            s_big[atomicAdd(&s_nbig, 1u)] = (uint16_t) j;
        } else if (mu[k] > 0.0) {
            PhiloxSeq rng(call.seed, e | ((uint64_t) call.stream << 32), TAG_COUNT);
            const uint32_t n = poisson_small(rng.next(), mu[k]);
            c = n;
            if (call.sliced && n > 0) {
                c |= (uint64_t) k1_in_slice(ph, map, call, edge_draw(ts, map, call, e), e, n) << 32;
            }
        }
        s_c[j] = c;
    }
    __syncthreads();
    for (uint32_t b = threadIdx.x; b < s_nbig; b += SCAN_THREADS) {
        const uint32_t j = s_big[b];
        const uint64_t e = call.edge_list != nullptr ? call.edge_list[tile_base + j] : tile_base + j;
        const EdgeDraw d = edge_draw(ts, map, call, e);
        const double m = (d.branch_end - d.branch_start) * d.mass;
        PhiloxSeq rng(call.seed, e | ((uint64_t) call.stream << 32), TAG_COUNT);
        const uint32_t n = poisson_native(rng, m);
        s_c[j] = (uint64_t) n | ((uint64_t) k1_in_slice(ph, map, call, d, e, n) << 32);
    }
    __syncthreads();
    // blocked: thread t scans the K1_ITEMS consecutive edges it owns
    uint64_t c[K1_ITEMS];
    uint64_t acc = 0, low = 0;
#pragma unroll
    for (int k = 0; k < K1_ITEMS; k++) {
        c[k] = s_c[threadIdx.x * K1_ITEMS + k];
        acc += c[k];
        low += c[k] & 0xffffffffull;
    }
    // Counts are packed (new mutations | those inside the slice << 32) and ids are int32: a tile
    // whose count does not fit 31 bits would carry into the upper half and cannot belong to a
    // valid result anyway (MSPB_ERR_TOO_MANY_ROWS).  Every thread below 2^21 means the tile is
    // below 2^29; otherwise the tile is summed exactly.
    if (__syncthreads_or(low >= (1ull << 21) ? 1 : 0)) {
        atomicAdd(&s_exact, (unsigned long long) low);
        __syncthreads();
        if (s_exact > 0x7fffffffull) {
            if (threadIdx.x == 0) {
                set_error(err, -1005);
            }
            acc = 0; // keep the packed sums in range; the call fails
#pragma unroll
            for (int k = 0; k < K1_ITEMS; k++) {
                c[k] = 0;
            }
        }
    }
    uint64_t block_total;
    uint64_t run = block_exclusive_scan<uint64_t>(acc, sm, &block_total);
    if (threadIdx.x == 0) {
        tile_tot[blockIdx.x] = block_total;
    }
#pragma unroll
    for (int k = 0; k < K1_ITEMS; k++) {
        s_c[threadIdx.x * K1_ITEMS + k] = run;
        run += c[k];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < K1_ITEMS; k++) {
        const uint32_t j = (uint32_t) k * SCAN_THREADS + threadIdx.x;
        const uint64_t w = tile_base + j;
        if (w <= last) {
            const uint64_t v = s_c[j];
            goff[w] = (uint32_t) v;
            if (loff != nullptr) {
                loff[w] = (uint32_t) (v >> 32);
            }
        }
    }
}

// tile_tot[t] (count | in-slice << 32) -> exclusive prefix in place, with the grand total
// appended at tile_tot[num_tiles] and copied to *total.  One block; each thread owns a
// contiguous run of tiles, so it is one pass and one block scan.
__global__ void __launch_bounds__(1024) k1_spine_kernel(uint64_t *tile_tot, uint32_t num_tiles, uint64_t *total,
    int *err)
{
    __shared__ uint64_t warp_sum[32];
    __shared__ unsigned long long s_exact;
    if (threadIdx.x == 0) {
        s_exact = 0ull;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t per = (num_tiles + 1023u) / 1024u;
    const uint32_t lo = threadIdx.x * per, hi = lo + per < num_tiles ? lo + per : num_tiles;
    // (a thread's few tiles are loaded together: up to 16 independent loads in flight)
    uint64_t mine[16];
    uint64_t sum = 0, low = 0;
    if (per <= 16) {
#pragma unroll
        for (uint32_t q = 0; q < 16; q++) {
            mine[q] = lo + q < hi ? tile_tot[lo + q] : 0ull;
        }
#pragma unroll
        for (uint32_t q = 0; q < 16; q++) {
            sum += mine[q];
            low += mine[q] & 0xffffffffull;
        }
    } else {
        for (uint32_t i = lo; i < hi; i++) {
            const uint64_t v = tile_tot[i];
            sum += v;
            low += v & 0xffffffffull;
        }
    }
    // the grand total must fit 31 bits (ids are int32; the packed halves must not carry): 1024
    // threads below 2^21 each are safe, otherwise sum exactly
    if (__syncthreads_or(low >= (1ull << 21) ? 1 : 0)) {
        atomicAdd(&s_exact, (unsigned long long) low);
        __syncthreads();
        if (s_exact > 0x7fffffffull) {
            if (threadIdx.x == 0) {
                set_error(err, -1005);
            }
            sum = 0; // the call fails; keep what K2's guard reads in range
#pragma unroll
            for (uint32_t q = 0; q < 16; q++) {
                mine[q] = 0;
            }
        }
    }
    const uint64_t inc = warp_inclusive_scan(sum);
    if (lane == 31) {
        warp_sum[warp] = inc;
    }
    __syncthreads();
    if (warp == 0) {
        const uint64_t w = warp_sum[lane];
        warp_sum[lane] = warp_inclusive_scan(w) - w;
    }
    __syncthreads();
    uint64_t run = warp_sum[warp] + inc - sum;
    if (per <= 16) {
#pragma unroll
        for (uint32_t q = 0; q < 16; q++) {
            if (lo + q < hi) {
                tile_tot[lo + q] = run;
            }
            run += mine[q];
        }
    } else {
        for (uint32_t i = lo; i < hi; i++) {
            const uint64_t v = tile_tot[i];
            tile_tot[i] = run;
            run += v;
        }
    }
    if (threadIdx.x == 1023) {
        tile_tot[num_tiles] = run;
        *total = run;
    }
}

// global offset of edge e's first mutation (e == E: the total), from the tile-local
// offsets and the tile bases; `hi` selects the in-slice counts
MSPB_DEVFN uint32_t k1_offset(const uint32_t *__restrict__ local, const uint64_t *__restrict__ tile_base, uint64_t e,
    uint64_t E, bool hi)
{
    const uint64_t t = e >> K1_TILE_SHIFT;
    const uint64_t b = e < E ? tile_base[t] : tile_base[(E + K1_TILE - 1) >> K1_TILE_SHIFT];
    return (e < E ? local[e] : 0u) + (uint32_t) (hi ? b >> 32 : b);
}

// ---------------------------------------------------------------- slice edge list
// Edges overlapping [slice_left, slice_right), as an ascending list of edge ids (an index
// over the loaded tables, rebuilt only when the slice changes).
struct EdgeInSlice {
    const double *left, *right;
    double slice_left, slice_right;
    MSPB_DEVFN uint32_t operator()(uint64_t e) const
    {
        return (left[e] < slice_right && right[e] > slice_left) ? 1u : 0u;
    }
};

__global__ void __launch_bounds__(256) slice_edges_scatter_kernel(EdgeInSlice f, uint64_t num_edges,
    const uint32_t *__restrict__ pos, uint32_t *__restrict__ list)
{
    const uint64_t e = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (e < num_edges && f(e) != 0u) {
        list[pos[e]] = (uint32_t) e;
    }
}

// ---------------------------------------------------------------- K2 (native)
// One block per 256 edges: their draw parameters are staged in shared memory, then
// each thread places one mutation (slot j of the tile's contiguous slot range), so
// record and key writes are coalesced whatever the per-edge counts are.
constexpr int K2_EDGES = 256;

__global__ void __launch_bounds__(K2_EDGES) k2_place_kernel(TreeSeqView ts, RateMapView map, CallView call,
    const uint32_t *__restrict__ goff, const uint32_t *__restrict__ loff, const uint64_t *__restrict__ tile_base,
    uint32_t base, Rec *__restrict__ rec, KeyView keys, const uint64_t *__restrict__ k1_total, uint64_t capacity,
    const int *err)
{
    // Launched before the host has read the K1 total (so that the read overlaps this kernel):
    // the buffers then have the capacity of earlier calls, and a total that does not fit (or a
    // count that K1 refused) makes the whole grid return; the host launches again after growing them.
    if (k1_total != nullptr && ((*k1_total & 0xffffffffull) + base > capacity || *err != 0)) {
        return;
    }
    __shared__ EdgeDraw sd[K2_EDGES];
    __shared__ uint32_t sid[K2_EDGES]; // edge ids (the Philox key of a placement)
    __shared__ uint8_t s_owner[2][K2_EDGES];
    static_assert(K2_EDGES <= 256, "edge index inside the block fits one byte");
    __shared__ uint32_t soff[K2_EDGES + 1];
    __shared__ uint32_t sm[9];
    const uint64_t E = call.num_work; // work items: all edges, or the slice's edge list
    const uint64_t ebase = (uint64_t) blockIdx.x * K2_EDGES;
    const uint64_t e = ebase + threadIdx.x;
    soff[threadIdx.x] = k1_offset(goff, tile_base, e < E ? e : E, E, false);
    if (threadIdx.x == 0) {
        soff[K2_EDGES] = k1_offset(goff, tile_base, ebase + K2_EDGES < E ? ebase + K2_EDGES : E, E, false);
    }
    __syncthreads();
    const uint32_t s0 = soff[0], s1 = soff[K2_EDGES];
    if (s0 == s1) {
        return;
    }
    if (e < E) {
        const uint32_t id = call.edge_list != nullptr ? call.edge_list[e] : (uint32_t) e;
        sid[threadIdx.x] = id;
        if (soff[threadIdx.x + 1] > soff[threadIdx.x]) {
            sd[threadIdx.x] = edge_draw(ts, map, call, id);
        }
    }
    __syncthreads();
    const Philox ph(call.seed);
    uint32_t out = base + (call.sliced ? k1_offset(loff, tile_base, ebase, E, true) : s0);
    const uint32_t my0 = soff[threadIdx.x], my1 = soff[threadIdx.x + 1];
    for (uint32_t j0 = s0, round = 0; j0 < s1; j0 += K2_EDGES, round++) {
        // which edge owns each of this round's 256 slots: every edge writes its own index into
        // the slots it covers (as many stores as there are slots; double buffered by round)
        uint8_t *owner = s_owner[round & 1u];
        {
            const uint32_t a = my0 > j0 ? my0 : j0, b = my1 < j0 + K2_EDGES ? my1 : j0 + K2_EDGES;
            for (uint32_t q = a; q < b; q++) {
                owner[q - j0] = (uint8_t) threadIdx.x;
            }
        }
        __syncthreads();
        const uint32_t j = j0 + threadIdx.x;
        const bool valid = j < s1;
        double x = 0, t = 0;
        int32_t child = 0;
        bool in = false;
        if (valid) {
            const uint32_t lo = owner[threadIdx.x];
            const EdgeDraw &d = sd[lo];
            place_one(ph, map, call, d, sid[lo], j - soff[lo], &x, &t);
            child = d.child;
            in = !call.sliced || (x >= call.slice_left && x < call.slice_right);
        }
        uint32_t slot;
        if (call.sliced) {
            uint32_t tot;
            slot = out + block_exclusive_scan<uint32_t>(in ? 1u : 0u, sm, &tot);
            out += tot;
        } else {
            slot = out + threadIdx.x;
            out += K2_EDGES;
        }
        if (in) {
            Rec r;
            r.time = t;
            r.node = child;
            r.ord = j;
            rec[slot] = r;
            keys.store(slot, x, t, false);
        }
    }
}

// ---------------------------------------------------------------- K2, one thread per mutation (native)
// k2_place_kernel gives a block 256 EDGES and lets each thread place one of their mutations per round: right
// when edges carry a mutation or more each (config 5 on one GPU: 1.5), wasteful when they are sparse -- the
// clipped slices of a sharded call keep every edge that crosses the slice but 1 / N of its mass (0.4 mutations
// per edge at 8 GPUs: a block's one round runs with a fifth of its lanes).  Here the grid follows the MUTATIONS:
// thread j finds the edge that owns slot j by two binary searches over the offsets K1 left (tile bases, then the
// tile's local offsets: a warp's 32 slots fall into the same tile or two, so the probes coalesce) and places it.
// Same slots, same Philox keys, same records as k2_place_kernel; not for invariant slices (their slots follow
// an in-slice count).  n_grid: the slots this launch covers; a total beyond it (the launch was sized from the
// previous call) leaves the rest to a second launch by the host.
__global__ void __launch_bounds__(256) k2_place_sparse_kernel(TreeSeqView ts, RateMapView map, CallView call,
    const uint32_t *__restrict__ goff, const uint64_t *__restrict__ tile_base, uint32_t base, Rec *__restrict__ rec,
    KeyView keys, const uint64_t *__restrict__ k1_total, uint64_t capacity, const int *err, uint64_t first_slot)
{
    const uint64_t E = call.num_work;
    const uint32_t num_tiles = (uint32_t) ((E + K1_TILE - 1) >> K1_TILE_SHIFT);
    const uint64_t total = k1_total != nullptr ? (*k1_total & 0xffffffffull) : (tile_base[num_tiles] & 0xffffffffull);
    if (k1_total != nullptr && (total + base > capacity || *err != 0)) {
        return; // launched ahead of the K1 read-back: see k2_place_kernel
    }
    const uint64_t j0 = first_slot + (uint64_t) blockIdx.x * blockDim.x;
    if (j0 >= total) {
        return;
    }
    // the last tile whose base is <= the block's first slot: one search per block ...
    __shared__ uint32_t s_lo;
    if (threadIdx.x == 0) {
        uint32_t lo = 0, hi = num_tiles; // invariant: base(lo) <= j0, base(hi) > j0 (hi == num_tiles: the total)
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if ((tile_base[mid] & 0xffffffffull) <= j0) {
                lo = mid;
            } else {
                hi = mid;
            }
        }
        s_lo = lo;
    }
    __syncthreads();
    const uint64_t j = j0 + threadIdx.x;
    if (j >= total) {
        return;
    }
    // ... and a step or two from there for each slot (a tile holds hundreds of mutations unless it is nearly empty)
    uint32_t lo = s_lo;
    while (lo + 1 < num_tiles && (tile_base[lo + 1] & 0xffffffffull) <= j) {
        lo++;
    }
    const uint32_t in_tile = (uint32_t) (j - (tile_base[lo] & 0xffffffffull));
    // the last edge of that tile whose local offset is <= in_tile
    const uint64_t e0 = (uint64_t) lo << K1_TILE_SHIFT;
    uint32_t a = 0, b = (uint32_t) (E - e0 < (uint64_t) K1_TILE ? E - e0 : (uint64_t) K1_TILE);
    while (b - a > 1) {
        const uint32_t mid = (a + b) >> 1;
        if (goff[e0 + mid] <= in_tile) {
            a = mid;
        } else {
            b = mid;
        }
    }
    const uint64_t e = e0 + a;
    const uint32_t id = call.edge_list != nullptr ? call.edge_list[e] : (uint32_t) e;
    const EdgeDraw d = edge_draw(ts, map, call, id);
    double x, t;
    place_one(Philox(call.seed), map, call, d, id, in_tile - goff[e], &x, &t);
    Rec r;
    r.time = t;
    r.node = d.child;
    r.ord = (uint32_t) j;
    const uint64_t slot = (uint64_t) base + j;
    rec[slot] = r;
    keys.store(slot, x, t, false);
}

// ---------------------------------------------------------------- K1/K2 (replay)
// Nonempty (edge, rate interval) pairs found by the sequential stream walker.
struct ReplayPairs {
    uint32_t *edge;
    uint32_t *interval;
    uint32_t *count;
    uint32_t *out_off; // mutations before this pair
    uint64_t *stream_off; // index of the pair's first position uniform
    uint64_t capacity;
};

struct ReplayTotals {
    unsigned long long num_pairs, num_mutations, draws;
    int exhausted;
};

// Continuous genome under the replay harness: a position that is already taken (by a kept site
// or by an earlier new mutation) is drawn again from the NEXT uniform of the stream, exactly as
// the reference's loop does (lib/mutgen.c:1127-1134), and so is a flat() value that rounds up to
// its upper bound (lib/util.c:717-735).  Either makes the number of uniforms a mutation consumes
// depend on everything placed before it, so the sequential walker places the mutations itself:
// it keeps the taken positions in an open-addressing hash set in device memory (`set`, zeroed by
// the host, capacity a power of two at least twice the number of positions it can hold) and
// writes records and keys directly.  set == nullptr: discrete genome, placement stays parallel.
struct ReplaySerial {
    uint64_t *set;
    uint64_t mask; // capacity - 1
    const double *kept_position; // positions of the kept sites (taken from the start)
    uint64_t num_kept_sites;
    Rec *rec;
    uint64_t *k64;
    uint32_t base; // slot of the first new mutation
    uint64_t max_new; // room for new mutations in rec / k64
};

MSPB_DEVFN uint64_t replay_hash(uint64_t k)
{
    k ^= k >> 33;
    k *= 0xff51afd7ed558ccdull;
    k ^= k >> 33;
    k *= 0xc4ceb9fe1a85ec53ull;
    k ^= k >> 33;
    return k;
}

// true when key was already in the set; inserts it otherwise (keys are never 0: positions are >= 0)
MSPB_DEVFN bool replay_set_test_and_insert(const ReplaySerial &sp, uint64_t key)
{
    uint64_t h = replay_hash(key) & sp.mask;
    for (;;) {
        const uint64_t v = sp.set[h];
        if (v == key) {
            return true;
        }
        if (v == 0ull) {
            sp.set[h] = key;
            return false;
        }
        h = (h + 1) & sp.mask;
    }
}

MSPB_DEVFN bool replay_set_contains(const ReplaySerial &sp, uint64_t key)
{
    uint64_t h = replay_hash(key) & sp.mask;
    for (;;) {
        const uint64_t v = sp.set[h];
        if (v == key) {
            return true;
        }
        if (v == 0ull) {
            return false;
        }
        h = (h + 1) & sp.mask;
    }
}

// msp_gsl_ran_flat (lib/util.c:717-735) on the stream: no draw when lo == hi, again while x >= hi
MSPB_DEVFN double replay_flat_serial(ReplayStream &s, double lo, double hi)
{
    if (lo == hi) {
        return lo;
    }
    double x;
    do {
        const double u = s.next();
        x = lo * (1 - u) + hi * u;
    } while (x >= hi && !s.exhausted);
    return x;
}

// One thread walks the stream in the reference's order: edges in table order, rate
// intervals left to right, Poisson first, then two uniforms per mutation.
__global__ void replay_resolve_kernel(TreeSeqView ts, RateMapView map, CallView call, const double *stream,
    uint64_t stream_len, ReplayPairs pairs, ReplaySerial sp, ReplayTotals *totals, int *err)
{
    if (blockIdx.x != 0 || threadIdx.x != 0) {
        return;
    }
    ReplayStream s { stream, stream_len, 0, 0 };
    uint64_t np = 0, nm = 0;
    const bool serial = sp.set != nullptr;
    if (serial) {
        // two kept sites at one position are P1's to report (MSP_ERR_DUPLICATE_SITE_POSITION)
        for (uint64_t j = 0; j < sp.num_kept_sites; j++) {
            replay_set_test_and_insert(sp, ordered_bits(sp.kept_position[j]));
        }
    }
    for (uint64_t e = 0; e < ts.num_edges && !s.exhausted; e++) {
        double left = ts.edge_left[e];
        const double edge_right = ts.edge_right[e];
        const int32_t child = ts.edge_child[e];
        const double branch_start = fmax(call.start_time, ts.node_time[child]);
        const double branch_end = fmin(call.end_time, ts.node_time[ts.edge_parent[e]]);
        const double branch_length = branch_end - branch_start;
        uint32_t idx = map.index(left);
        double right = 0;
        while (right != edge_right && !s.exhausted) {
            right = fmin(edge_right, map.pos[idx + 1]);
            const double site_left = map.coord(left), site_right = map.coord(right);
            const double mu = branch_length * (site_right - site_left) * map.rate[idx];
            const unsigned int n = replay_poisson(s, mu);
            if (n > 0) {
                if (np >= pairs.capacity || nm + n > 0x7fffffffull || (serial && nm + n > sp.max_new)) {
                    if (serial && nm + n > sp.max_new && nm + n <= 0x7fffffffull) {
                        s.exhausted = 1; // every mutation takes two uniforms or more: the stream cannot cover them
                        break;
                    }
                    set_error(err, -1005);
                    s.exhausted = 1;
                    break;
                }
                pairs.edge[np] = (uint32_t) e;
                pairs.interval[np] = idx;
                pairs.count[np] = n;
                pairs.out_off[np] = (uint32_t) nm;
                pairs.stream_off[np] = s.pos;
                np++;
                if (!serial) {
                    nm += n;
                    s.pos += 2ull * n;
                    if (s.pos > s.len) {
                        s.exhausted = 1;
                    }
                } else {
                    for (unsigned int k = 0; k < n && !s.exhausted; k++, nm++) {
                        double x;
                        uint64_t key;
                        do {
                            x = replay_flat_serial(s, site_left, site_right);
                            key = ordered_bits(x);
                        } while (replay_set_contains(sp, key) && !s.exhausted);
                        const double t = replay_flat_serial(s, branch_start, branch_end);
                        replay_set_test_and_insert(sp, key);
                        Rec r;
                        r.time = t;
                        r.node = child;
                        r.ord = (uint32_t) nm;
                        sp.rec[sp.base + nm] = r;
                        sp.k64[sp.base + nm] = key;
                    }
                }
            }
            left = right;
            idx++;
        }
    }
    totals->num_pairs = np;
    totals->num_mutations = nm;
    totals->draws = s.pos;
    totals->exhausted = s.exhausted;
}

// flat(lo, hi) of lib/util.c:717-735 on a given uniform; a value that the reference
// would redraw (x >= hi) cannot be replayed in parallel and is reported.
MSPB_DEVFN double replay_flat(double lo, double hi, double u, int *err)
{
    const double x = lo * (1 - u) + hi * u;
    if (x >= hi) {
        set_error(err, -1004);
    }
    return x;
}

__global__ void __launch_bounds__(128) replay_place_kernel(TreeSeqView ts, RateMapView map, CallView call,
    const double *__restrict__ stream, ReplayPairs pairs, uint64_t num_pairs, uint32_t base, Rec *__restrict__ rec,
    KeyView keys, int *err)
{
    const uint64_t p = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= num_pairs) {
        return;
    }
    const uint64_t e = pairs.edge[p];
    const uint32_t idx = pairs.interval[p];
    const double left = fmax(ts.edge_left[e], map.pos[idx]);
    const double right = fmin(ts.edge_right[e], map.pos[idx + 1]);
    const double site_left = map.coord(left), site_right = map.coord(right);
    const int32_t child = ts.edge_child[e];
    const double branch_start = fmax(call.start_time, ts.node_time[child]);
    const double branch_end = fmin(call.end_time, ts.node_time[ts.edge_parent[e]]);
    const double *u = stream + pairs.stream_off[p];
    const uint32_t n = pairs.count[p];
    uint32_t slot = base + pairs.out_off[p];
    for (uint32_t k = 0; k < n; k++, slot++) {
        double x = replay_flat(site_left, site_right, u[2 * k], err);
        if (call.discrete) {
            x = floor(x);
        }
        const double t = replay_flat(branch_start, branch_end, u[2 * k + 1], err);
        Rec r;
        r.time = t;
        r.node = child;
        r.ord = slot - base;
        rec[slot] = r;
        keys.store(slot, x, t, false);
    }
}

// ---------------------------------------------------------------- kept records
// slots [0, S0): site markers; [S0, S0 + M0): kept mutations (lib/mutgen.c:974-1025)
__global__ void __launch_bounds__(256) fill_kept_kernel(KeptView kept, CallView call, Rec *__restrict__ rec,
    KeyView keys, int *err)
{
    const uint64_t r = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t S0 = kept.num_sites, M0 = kept.num_mutations;
    if (r >= S0 + M0) {
        return;
    }
    double x, t = 0.0;
    int32_t node = -1;
    if (r < S0) {
        x = kept.site_position[r];
    } else {
        const uint64_t m = r - S0;
        x = kept.site_position[kept.mut_site[m] - kept.site_id_offset];
        t = kept.mut_time == nullptr ? nan("") : kept.mut_time[m];
        node = kept.mut_node[m];
        if (isnan(t)) {
            set_error(err, -69); // MSP_ERR_UNKNOWN_TIME_NOT_SUPPORTED
            t = 0.0;
        }
        if (m > 0 && kept.mut_site[m - 1] > kept.mut_site[m]) {
            set_error(err, -1006);
        }
    }
    Rec q;
    q.time = t;
    q.node = node;
    q.ord = 0;
    rec[r] = q;
    keys.store(r, x, t, r < S0);
}

// model allele index of each kept state (exact byte comparison, lib/mutgen.c:184-198)
__global__ void __launch_bounds__(256) kept_allele_index_kernel(ModelView model, uint64_t n, const char *bytes,
    const uint64_t *off, int32_t *out)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) {
        return;
    }
    const uint64_t a = off[i], len = off[i + 1] - a;
    int32_t found = -1;
    for (uint32_t j = 0; j < model.num_alleles && found < 0; j++) {
        const uint64_t b = model.allele_off[j], blen = model.allele_off[j + 1] - b;
        if (blen == len) {
            bool same = true;
            for (uint64_t k = 0; k < len && same; k++) {
                same = bytes[a + k] == model.alleles[b + k];
            }
            if (same) {
                found = (int32_t) j;
            }
        }
    }
    out[i] = found;
}

// ---------------------------------------------------------------- after the sort
struct SortedView {
    uint64_t n; // records
    uint64_t S0, M0; // kept sites / kept mutations (slot ranges)
    KeyView key; // sorted keys
    uint32_t *val; // slot of the record at each sorted index (K4 reorders inside sites)
    const Rec *rec;
    int has_kept; // S0 + M0 > 0
    int root_draw; // a new site takes one draw / allele number for its root state (not the SLiM model)
    // P1 results
    uint32_t *site_row; // row of the site that record i belongs to
    uint32_t *mut_scan; // mutation rows before i (has_kept only; otherwise it is i)
    uint32_t *draw_scan; // root + transition draws before i (has_kept only)

    MSPB_DEVFN bool head(uint64_t i) const { return i == 0 || !key.same_pos(i, i - 1); }
    MSPB_DEVFN bool is_mut(uint64_t i) const { return !has_kept || val[i] >= S0; }
    MSPB_DEVFN bool is_new(uint64_t i) const { return !has_kept || val[i] >= S0 + M0; }
    // a site with exactly one record: nothing to order, no parent to find
    MSPB_DEVFN bool single(uint64_t i) const { return head(i) && (i + 1 >= n || head(i + 1)); }
    MSPB_DEVFN uint32_t mut_row(uint64_t i) const { return has_kept ? mut_scan[i] : (uint32_t) i; }
    MSPB_DEVFN uint32_t draws_before(uint64_t i) const
    {
        if (has_kept) {
            return draw_scan[i];
        }
        return root_draw ? (uint32_t) i + site_row[i] + 1u - (head(i) ? 1u : 0u) : (uint32_t) i;
    }
};

struct Counts3 {
    uint32_t s, m, d; // sites, mutation rows, allele draws
    uint32_t u; // sites with more than one record (K4's work list)
    MSPB_DEVFN Counts3 operator+(const Counts3 &o) const
    {
        Counts3 r;
        r.s = s + o.s;
        r.m = m + o.m;
        r.d = d + o.d;
        r.u = u + o.u;
        return r;
    }
};

constexpr int P1_ITEMS = 16;
constexpr int P1_TILE = SCAN_THREADS * P1_ITEMS;

template <int MODE>
struct P1Key {
    typedef uint64_t type;
    static MSPB_DEVFN const uint64_t *ptr(const KeyView &k) { return k.k64; }
    static MSPB_DEVFN bool same_pos(uint64_t a, uint64_t b)
    {
        return MODE == KEY64_POS ? a == b : (a >> 32) == (b >> 32);
    }
};

template <>
struct P1Key<KEY32_POS> {
    typedef uint32_t type;
    static MSPB_DEVFN const uint32_t *ptr(const KeyView &k) { return k.k32; }
    static MSPB_DEVFN bool same_pos(uint32_t a, uint32_t b) { return a == b; }
};

// 16 consecutive 4- or 8-byte values starting at a 64-byte aligned index, as 16-byte loads
template <typename T>
MSPB_DEVFN void load16(const T *__restrict__ src, T *dst)
{
    constexpr int PER = 16 / (int) sizeof(T);
    const uint4 *v = reinterpret_cast<const uint4 *>(src);
#pragma unroll
    for (int q = 0; q < P1_ITEMS / PER; q++) {
        const uint4 w = v[q];
        if (sizeof(T) == 4) {
            dst[4 * q] = (T) w.x;
            dst[4 * q + 1] = (T) w.y;
            dst[4 * q + 2] = (T) w.z;
            dst[4 * q + 3] = (T) w.w;
        } else {
            dst[2 * q] = (T) ((uint64_t) w.x | ((uint64_t) w.y << 32));
            dst[2 * q + 1] = (T) ((uint64_t) w.z | ((uint64_t) w.w << 32));
        }
    }
}

MSPB_DEVFN void store16(uint32_t *dst, const uint32_t *v)
{
    uint4 *o = reinterpret_cast<uint4 *>(dst);
#pragma unroll
    for (int q = 0; q < P1_ITEMS / 4; q++) {
        o[q] = make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
    }
}

// per-record counts packed for one scan: sites | mutation rows << 16 | draws << 32 | multi << 48
MSPB_DEVFN uint64_t p1_counts(uint32_t f)
{
    return (uint64_t) (f & 1u) | ((uint64_t) ((f >> 1) & 1u) << 16) | ((uint64_t) ((f >> 2) & 3u) << 32)
           | ((uint64_t) ((f >> 4) & 1u) << 48);
}

// Site boundaries and numbering in one pass: site row of every record, mutation rows
// and allele draws before it, plus the input checks that need the sorted order (two
// kept sites at one position, lib/mutgen.c:875-879; a new mutation sharing a position
// in the continuous genome, which the reference redraws, lib/mutgen.c:1127-1134).
// Thread t owns 16 consecutive records, read and written with 16-byte accesses and kept
// in registers throughout (no shared-memory transposes); one block scan and one look-back
// across tiles give the numbering.
template <int MODE, bool KEPT>
__global__ void __launch_bounds__(SCAN_THREADS, (MODE == KEY32_POS && !KEPT) ? 4 : 2) p1_sites_kernel(SortedView sv, int discrete, LbWord *lb,
    uint32_t num_tiles, Counts3 *total, uint32_t *__restrict__ multi_list, uint32_t *collisions, uint32_t *dense,
    int *err)
{
    typedef typename P1Key<MODE>::type K;
    __shared__ uint64_t sm[9];
    __shared__ LbVal<1, false> bcast;
    const K *__restrict__ keys = P1Key<MODE>::ptr(sv.key);
    const uint64_t tile_base = (uint64_t) blockIdx.x * P1_TILE;
    const uint64_t i0 = tile_base + (uint64_t) threadIdx.x * P1_ITEMS;
    const uint64_t last = sv.n - 1;
    const bool full = tile_base + P1_TILE <= sv.n;
    K pk[P1_ITEMS + 2]; // pk[k + 1]: key of record i0 + k
    uint32_t slot[KEPT ? P1_ITEMS : 1];
    if (full) {
        load16<K>(keys + i0, pk + 1);
        if (KEPT) {
            load16<uint32_t>(sv.val + i0, slot);
        }
    } else {
#pragma unroll
        for (int k = 0; k < P1_ITEMS; k++) {
            const uint64_t ic = i0 + k < last ? i0 + k : last;
            pk[k + 1] = keys[ic];
            if (KEPT) {
                slot[k] = sv.val[ic];
            }
        }
    }
    pk[0] = keys[i0 > 0 ? (i0 - 1 < last ? i0 - 1 : last) : 0];
    pk[P1_ITEMS + 1] = keys[i0 + P1_ITEMS < last ? i0 + P1_ITEMS : last];
    // flags: head | mut << 1 | draws << 2 (2 bits) | multi << 4
    uint32_t flw[P1_ITEMS / 4] = { 0u, 0u, 0u, 0u }; // one byte per record
    uint64_t acc = 0;
#pragma unroll
    for (int k = 0; k < P1_ITEMS; k++) {
        const uint64_t i = i0 + k;
        uint32_t f = 0;
        if (i < sv.n) {
            const bool h = i == 0 || !P1Key<MODE>::same_pos(pk[k + 1], pk[k]);
            const bool mu = !KEPT || slot[KEPT ? k : 0] >= sv.S0;
            const bool nw = !KEPT || slot[KEPT ? k : 0] >= sv.S0 + sv.M0;
            const uint32_t d = (nw ? 1u : 0u) + ((h && mu && sv.root_draw) ? 1u : 0u);
            const bool multi = h && i + 1 < sv.n && P1Key<MODE>::same_pos(pk[k + 2], pk[k + 1]);
            f = (h ? 1u : 0u) | ((mu ? 1u : 0u) << 1) | (d << 2) | ((multi ? 1u : 0u) << 4);
            if (!h) {
                if (KEPT && !mu) {
                    set_error(err, -21); // MSP_ERR_DUPLICATE_SITE_POSITION
                }
                if (!discrete && (nw || sv.is_new(i - 1))) {
                    atomicAdd(collisions, 1u);
                }
                if (MODE == KEY32_POS && i >= 32 && sv.key.same_pos(i, i - 32)) {
                    *dense = 1u;
                }
            }
        }
        flw[k >> 2] |= f << (8 * (k & 3));
        acc += p1_counts(f);
    }
    uint64_t block_total;
    uint64_t run = block_exclusive_scan<uint64_t>(acc, sm, &block_total);
    LbVal<1, false> agg;
    agg.w[0][0] = (uint32_t) (block_total & 0xffffull); // sites
    agg.w[0][1] = (uint32_t) ((block_total >> 16) & 0xffffull); // mutation rows
    agg.w[0][2] = (uint32_t) ((block_total >> 32) & 0xffffull); // draws
    agg.w[0][3] = (uint32_t) (block_total >> 48); // multi-record sites (< 2^30: carries the status bits)
    const LbVal<1, false> pw = lookback_prefix<1, false>(lb, blockIdx.x, agg, &bcast);
    Counts3 pre;
    pre.s = pw.w[0][0];
    pre.m = pw.w[0][1];
    pre.d = pw.w[0][2];
    pre.u = pw.w[0][3];
    if (blockIdx.x + 1 == num_tiles && threadIdx.x == 0) {
        Counts3 all;
        all.s = pre.s + agg.w[0][0];
        all.m = pre.m + agg.w[0][1];
        all.d = pre.d + agg.w[0][2];
        all.u = pre.u + agg.w[0][3];
        *total = all;
    }
    uint32_t site[P1_ITEMS], mut[KEPT ? P1_ITEMS : 1], draw[KEPT ? P1_ITEMS : 1];
#pragma unroll
    for (int k = 0; k < P1_ITEMS; k++) {
        const uint32_t f = (flw[k >> 2] >> (8 * (k & 3))) & 0xffu;
        // site: inclusive count (the record's own site row is count - 1); others exclusive
        site[k] = pre.s + (uint32_t) (run & 0xffffull) + (f & 1u) - 1u;
        if (KEPT) {
            mut[k] = pre.m + (uint32_t) ((run >> 16) & 0xffffull);
            draw[k] = pre.d + (uint32_t) ((run >> 32) & 0xffffull);
        }
        if ((f >> 4) & 1u) {
            multi_list[pre.u + (uint32_t) (run >> 48)] = (uint32_t) (i0 + k);
        }
        run += p1_counts(f);
    }
    if (full) {
        store16(sv.site_row + i0, site);
        if (KEPT) {
            store16(sv.mut_scan + i0, mut);
            store16(sv.draw_scan + i0, draw);
        }
    } else {
#pragma unroll
        for (int k = 0; k < P1_ITEMS; k++) {
            if (i0 + k < sv.n) {
                sv.site_row[i0 + k] = site[k];
                if (KEPT) {
                    sv.mut_scan[i0 + k] = mut[k];
                    sv.draw_scan[i0 + k] = draw[k];
                }
            }
        }
    }
}

// ---------------------------------------------------------------- decimal text
// decimal length of v (snprintf "%" PRIu64, lib/mutgen.c:431-447); allele numbers are
// almost always below 2^32, where 32-bit compares and divisions are much cheaper
MSPB_DEVFN uint32_t num_digits32(uint32_t v)
{
    return v < 100000u ? (v < 100u ? (v < 10u ? 1u : 2u) : (v < 1000u ? 3u : (v < 10000u ? 4u : 5u)))
                       : (v < 10000000u ? (v < 1000000u ? 6u : 7u)
                                        : (v < 100000000u ? 8u : (v < 1000000000u ? 9u : 10u)));
}

MSPB_DEVFN uint32_t num_digits(uint64_t v)
{
    if (v <= 0xffffffffull) {
        return num_digits32((uint32_t) v);
    }
    uint32_t n = 9; // v >= 2^32 > 10^9: the low nine digits are full
    v /= 1000000000ull;
    while (v > 0xffffffffull) {
        v /= 1000000000ull;
        n += 9;
    }
    return n + num_digits32((uint32_t) v);
}

MSPB_DEVFN void write_digits(char *dst, uint64_t v, uint32_t n)
{
    uint32_t k = n;
    while (v > 0xffffffffull) {
        uint32_t low = (uint32_t) (v % 1000000000ull);
        v /= 1000000000ull;
        for (int d = 0; d < 9; d++) {
            dst[--k] = (char) ('0' + (low % 10u));
            low /= 10u;
        }
    }
    uint32_t w = (uint32_t) v;
    while (k > 0) {
        dst[--k] = (char) ('0' + (w % 10u));
        w /= 10u;
    }
}

MSPB_DEVFN void copy_bytes(char *dst, const char *src, uint64_t n)
{
    for (uint64_t k = 0; k < n; k++) {
        dst[k] = src[k];
    }
}

// ---------------------------------------------------------------- K4
// The reference's within-site order (lib/mutgen.c:36-49) on two records of one site:
// marker first; kept vs kept by row; otherwise time descending; ties by slot.
MSPB_DEVFN bool rec_before(const SortedView &sv, uint32_t a, uint32_t b)
{
    const bool am = a < sv.S0, bm = b < sv.S0;
    if (am || bm) {
        return am && !bm ? true : (bm && !am ? false : a < b);
    }
    const bool ak = a < sv.S0 + sv.M0, bk = b < sv.S0 + sv.M0;
    if (ak && bk) {
        return a < b;
    }
    const double ta = sv.rec[a].time, tb = sv.rec[b].time;
    if (ta != tb) {
        return ta > tb;
    }
    return a < b;
}

// Per-record results of the per-site step (written for sites with several records only).
struct ResolveView {
    int32_t *parent; // sorted index of the parent mutation, or -1
    int32_t *state; // matrix: allele index (new and kept; -1 = unknown kept allele)
    int32_t *site_state; // matrix: allele index of a new site's ancestral state (at the head)
    const double *replay; // phase-B uniforms (replay mode), else nullptr
    uint64_t replay_len;
};

MSPB_DEVFN uint32_t select_prob(double u, uint32_t n, const double *probs)
{
    // first i < n-1 with u < running sum, else n-1 (lib/util.c:400-424)
    uint32_t i = 0;
    double sum = 0.0;
    while (i + 1 < n) {
        sum += probs[i];
        if (u < sum) {
            break;
        }
        i++;
    }
    return i;
}

MSPB_DEVFN double phase_b_uniform(const ResolveView &rv, uint64_t idx, int *err)
{
    if (idx >= rv.replay_len) {
        set_error(err, -1003);
        return 0.5;
    }
    const double u = rv.replay[idx];
    if (u >= 1.0) {
        set_error(err, -1004);
    }
    return u;
}

// uniforms of the allele draws of the new mutation with ordinal `ord`:
// root (if it opens a new site) and transition
MSPB_DEVFN void allele_uniforms(const Philox &ph, uint32_t stream, uint32_t ord, double *u_root, double *u_trans)
{
    const uint4 w = ph(ord, stream, 0u, TAG_ALLELE);
    *u_trans = u53(w.x, w.y);
    *u_root = u53(w.z, w.w);
}

// One thread per site that has more than one record (P1's work list, so that all lanes
// of a warp are busy even when such sites are rare): put the records in the
// reference's order, then for each mutation find the nearest ancestor-or-self node
// already holding a mutation at this site (lib/mutgen.c:1185-1216) and draw its state.
// The walk uses the child-sorted interval index instead of a tree sweep and stops once
// the node is older than every earlier mutation's node.
MSPB_DEVFN void resolve_site(const TreeSeqView &ts, const KeptView &kept, const ModelView &model, const CallView &call,
    const SortedView &sv, const ResolveView &rv, uint64_t i, int *err);

__global__ void __launch_bounds__(128, 12) k4_resolve_kernel(TreeSeqView ts, KeptView kept, ModelView model, CallView call,
    SortedView sv, ResolveView rv, const uint32_t *__restrict__ multi_list, uint32_t *next, const Counts3 *p1_total,
    const uint32_t *dense, const uint32_t *collisions, int *err)
{
    if (*dense != 0u && sv.key.mode == KEY32_POS) {
        return; // the host re-sorts with time in the key and runs this step again
    }
    if (*collisions != 0u) {
        // continuous genome: some new mutation sits on an occupied position.  The reference draws
        // such a position again BEFORE any allele is read (lib/mutgen.c:1127-1134), so nothing
        // may be resolved (or fail on an allele the model does not know) until positions are
        // unique: the host redraws and runs this step again, or reports the collision.
        return;
    }
    // The walks differ a lot in length, so the work list is handed out dynamically: the grid is
    // as large as the device holds at once, and a warp claims the next 32 sites whenever it is
    // through with its last (`next` is zeroed with the call's totals).
    const uint32_t n_multi = p1_total->u;
    const uint32_t lane = threadIdx.x & 31u;
    for (;;) {
        uint32_t first = 0;
        if (lane == 0) {
            first = atomicAdd(next, 32u);
        }
        first = __shfl_sync(0xffffffffu, first, 0);
        if (first >= n_multi) {
            break;
        }
        if (first + lane < n_multi) {
            resolve_site(ts, kept, model, call, sv, rv, multi_list[first + lane], err);
        }
        __syncwarp();
    }
}

MSPB_DEVFN void resolve_site(const TreeSeqView &ts, const KeptView &kept, const ModelView &model, const CallView &call,
    const SortedView &sv, const ResolveView &rv, uint64_t i, int *err)
{
    uint64_t end = i + 2;
    while (end < sv.n && !sv.head(end)) {
        end++;
    }
    if (sv.has_kept) {
        // Existing mutations of this site that are NOT listed oldest first, with a new mutation on the site: the
        // reference's comparator (existing rows by id, anything else by time, lib/mutgen.c:36-49) is then not a
        // total order, what its qsort returns is undefined and its debug build stops at lib/mutgen.c:1200.
        // Refused with a code of its own.  Records arrive in slot order (= row order for existing rows) or, after
        // a dense genome's second sort, oldest first: either way a pair out of order shows between neighbours.
        const uint64_t KM0 = sv.S0 + sv.M0;
        bool any_new = false, out_of_order = false;
        uint32_t prev = 0xffffffffu;
        double prev_t = 0.0;
        for (uint64_t a = i; a < end; a++) {
            const uint32_t r = sv.val[a];
            if (r >= KM0) {
                any_new = true;
            } else if (r >= sv.S0) {
                const double t = sv.rec[r].time;
                if (prev != 0xffffffffu && ((prev < r && prev_t < t) || (prev > r && t < prev_t))) {
                    out_of_order = true;
                }
                prev = r;
                prev_t = t;
            }
        }
        if (any_new && out_of_order) {
            set_error(err, -1012); // MSPB_ERR_KEPT_MUTATIONS_UNSORTED
        }
    }
    // insertion sort of val[i .. end) (runs are short, or nearly sorted in KEY64_POS_TIME mode)
    for (uint64_t a = i + 1; a < end; a++) {
        const uint32_t v = sv.val[a];
        uint64_t b = a;
        while (b > i && rec_before(sv, v, sv.val[b - 1])) {
            sv.val[b] = sv.val[b - 1];
            b--;
        }
        sv.val[b] = v;
    }
    const uint64_t S0 = sv.S0, KM = sv.S0 + sv.M0;
    const uint32_t r0 = sv.val[i];
    const bool kept_site = sv.has_kept && r0 < S0;
    const uint64_t first = i + (kept_site ? 1 : 0);
    const bool matrix = model.kind == 0;
    const Philox ph(call.seed);
    const uint32_t draw0 = sv.draws_before(i);

    int32_t anc = -1;
    if (matrix) {
        if (kept_site) {
            anc = kept.site_allele[r0];
        } else {
            double u, ut;
            if (rv.replay != nullptr) {
                u = phase_b_uniform(rv, draw0, err);
            } else {
                allele_uniforms(ph, call.stream, sv.rec[r0].ord, &u, &ut);
            }
            anc = (int32_t) select_prob(u, model.num_alleles, model.root);
            rv.site_state[i] = anc;
        }
    }
    const double x = sv.key.position(i);
    double tmax = -INFINITY; // oldest node among the mutations seen so far
    // draws (matrix: uniforms; infinite alleles: allele numbers) before record j, redone
    // here because the reordering above may have interleaved kept and new mutations
    uint32_t cur = draw0;
    constexpr int SEEN = 8; // nodes of the site's first records stay in registers
    uint32_t seen[SEEN];
#pragma unroll
    for (int t = 0; t < SEEN; t++) {
        seen[t] = CI_SPAN_UNKNOWN; // no node has this identity
    }
    for (uint64_t j = first; j < end; j++) {
        const uint32_t r = sv.val[j];
        const bool is_new = !sv.has_kept || r >= KM;
        // a new site's first record also takes the root draw
        const uint32_t opens = (j == i && sv.root_draw) ? 1u : 0u;
        if (sv.has_kept) {
            sv.draw_scan[j] = cur;
        }
        const int32_t node = sv.rec[r].node;
        int32_t par = -1;
        const double node_t = ts.node_time[node];
        const WalkNode self = walk_node(ts, node);
#pragma unroll
        for (int t = 0; t < SEEN; t++) {
            if ((uint64_t) t == j - first) {
                seen[t] = self.id;
            }
        }
        if (j > first) {
            WalkNode u = self;
            double tu = node_t;
            bool at_self = true;
            // A strict ancestor of `node` is older than this mutation, so once the walk has left
            // `node` it can only meet earlier mutations whose node is older than this mutation.
            // (New mutations only: a kept mutation's time is not checked against its parent node.)
            const double mut_t = is_new ? sv.rec[r].time : -INFINITY;
            for (;;) {
                if (tu > tmax || (!at_self && tmax <= mut_t)) {
                    break;
                }
                // the latest earlier record on node u
#pragma unroll
                for (int t = 0; t < SEEN; t++) {
                    if ((uint64_t) t < j - first && seen[t] == u.id) {
                        par = (int32_t) (first + t);
                    }
                }
                for (uint64_t q = j; q-- > first + SEEN;) {
                    if (walk_id(ts, sv.rec[sv.val[q]].node) == u.id) {
                        par = (int32_t) q;
                        break;
                    }
                }
                if (par >= 0 || !parent_at(ts, &u, x, &tu)) {
                    break;
                }
                at_self = false;
            }
        }
        rv.parent[j] = par;
        if (matrix) {
            int32_t st;
            if (!is_new) {
                st = kept.mut_allele[r - S0];
            } else {
                const int32_t pa = par < 0 ? anc : rv.state[par];
                double u, ur;
                if (rv.replay != nullptr) {
                    u = phase_b_uniform(rv, (uint64_t) cur + opens, err);
                } else {
                    allele_uniforms(ph, call.stream, sv.rec[r].ord, &ur, &u);
                }
                if (pa < 0) {
                    set_error(err, -51); // MSP_ERR_UNKNOWN_ALLELE
                    st = 0;
                } else {
                    st = (int32_t) select_prob(u, model.num_alleles, model.matrix + (uint64_t) pa * model.num_alleles);
                }
            }
            rv.state[j] = st;
        } else if (model.kind == 2) {
            // SLiM: lengths of the stacked derived state and metadata (lib/mutgen.c:355-413); K5
            // materialises them by walking the same parent chain
            int32_t dl, ml;
            if (!is_new) {
                const uint64_t km = r - S0;
                dl = (int32_t) (kept.mut_der_off[km + 1] - kept.mut_der_off[km]);
                ml = (int32_t) (kept.mut_md_off[km + 1] - kept.mut_md_off[km]);
            } else {
                int32_t bl = 0, bm = 0;
                if (par >= 0) {
                    bl = rv.state[par];
                    bm = rv.site_state[par];
                } else if (kept_site) {
                    bl = (int32_t) (kept.site_anc_off[r0 + 1] - kept.site_anc_off[r0]);
                    bm = (int32_t) (kept.site_md_off[r0 + 1] - kept.site_md_off[r0]);
                }
                dl = bl + (bl > 0 ? 1 : 0) + (int32_t) num_digits(model.start_allele + cur);
                ml = bm + SLIM_MD_SIZE;
            }
            rv.state[j] = dl;
            rv.site_state[j] = ml;
        }
        cur += (is_new ? 1u : 0u) + opens;
        tmax = fmax(tmax, node_t);
    }
}

// ---------------------------------------------------------------- K5

// Byte prefixes of K5's tiles without a chain across tiles, for calls without kept rows whose
// state lengths follow from a record's index alone.  Infinite alleles: record i of site row
// s (rows count the heads before it) writes the number start + i + s + 1, and its site, if i
// is the head, start + i + s; both grow with i, so "how many records / heads write at most k
// digits" is one binary search per k (k5_prefix_table_kernel) and a tile's prefix is a sum over
// at most 20 digit classes.  Matrix models whose alleles all have one length: rows x length.
struct ChainFree {
    uint32_t enabled; // 0: K5 chains its tiles with the look-back as usual
    uint32_t fixed_len; // > 0: every state is this long (matrix model)
    uint64_t der_upto[21]; // [k]: records whose derived number has at most k digits
    uint64_t anc_upto[21]; // [k]: heads whose ancestral number has at most k digits
    MSPB_DEVFN void prefix(uint64_t records, uint64_t heads, uint64_t *der, uint64_t *anc) const
    {
        if (fixed_len > 0) {
            *der = records * fixed_len;
            *anc = heads * fixed_len;
            return;
        }
        uint64_t d = 0, a = 0, dp = 0, ap = 0;
        for (uint32_t k = 1; k <= 20 && (dp < records || ap < heads); k++) {
            const uint64_t dk = der_upto[k] < records ? der_upto[k] : records;
            const uint64_t ak = anc_upto[k] < heads ? anc_upto[k] : heads;
            d += (uint64_t) k * (dk - dp);
            a += (uint64_t) k * (ak - ap);
            dp = dk;
            ap = ak;
        }
        *der = d;
        *anc = a;
    }
};

struct OutView {
    double *site_position;
    char *site_anc;
    void *site_anc_off;
    char *site_md;
    void *site_md_off;
    int32_t *mut_site, *mut_node, *mut_parent;
    double *mut_time;
    char *mut_der;
    void *mut_der_off;
    char *mut_md;
    void *mut_md_off;
    uint32_t site_base, mut_base; // added to site ids / parent ids (genome slices of a sharded call)
    // sharded calls that exchange their counts on the device: site / mutation / allele-number
    // bases are read from here (3 words) and added to the values above
    const uint64_t *dev_bases;
    // ... and, when the byte counts were exchanged BEFORE the tables are written, the bytes of the slices before
    // this one (ancestral, site metadata, derived, mutation metadata), added to the offset columns only
    const uint64_t *dev_byte_bases;
    int off32; // offset columns are uint32 (else uint64)
    int write_md_off; // 0: the metadata offset columns are known to be all zero already
    int dry_run; // K5 computes the byte totals only (sizing pass of the SLiM model)
    const ChainFree *chain_free; // tile prefixes without the look-back (nullptr: not applicable)
    MSPB_DEVFN void store_off(void *col, uint64_t row, uint64_t v) const
    {
        if (off32) {
            static_cast<uint32_t *>(col)[row] = (uint32_t) v;
        } else {
            static_cast<uint64_t *>(col)[row] = v;
        }
    }
};

} // namespace mspb

#include "emit.cuh"

namespace mspb {

// Continuous genome: a new mutation that landed on an occupied position (a kept site, or an
// earlier new mutation) draws its position again, as the reference does
// (lib/mutgen.c:1127-1134).  Runs on the position-sorted records: colliders are the new
// records that are not the first of their position.  The record's edge is found from its
// ordinal (K2 numbered the mutations edge by edge); the new position comes from the same
// edge's span on a Philox counter of its own (TAG_REDRAW + round).
__global__ void __launch_bounds__(256) redraw_positions_kernel(SortedView sv, TreeSeqView ts, RateMapView map,
    CallView call, const uint32_t *__restrict__ goff, const uint64_t *__restrict__ tile_base, uint32_t round,
    uint64_t *__restrict__ key_out, uint32_t *__restrict__ val_out)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sv.n) {
        return;
    }
    const uint32_t r = sv.val[i];
    uint64_t key = sv.key.k64[i];
    if (i > 0 && sv.is_new(i) && sv.key.same_pos(i, i - 1)) {
        const uint32_t j = sv.rec[r].ord;
        // largest work item w with offset(w) <= j and at least one mutation
        const uint64_t E = call.num_work;
        uint64_t lo = 0, hi = E;
        while (hi - lo > 1) {
            const uint64_t mid = (lo + hi) >> 1;
            if (k1_offset(goff, tile_base, mid, E, false) <= j) {
                lo = mid;
            } else {
                hi = mid;
            }
        }
        const uint64_t e = call.edge_list != nullptr ? call.edge_list[lo] : lo;
        const uint32_t k = j - k1_offset(goff, tile_base, lo, E, false);
        const EdgeDraw d = edge_draw(ts, map, call, e);
        const Philox ph(call.seed);
        const uint4 q = ph((uint32_t) e, call.stream, k, TAG_REDRAW + round);
        const double x = map.position_at(d.mass_lo + u53(q.x, q.y) * d.mass, d.left, d.right);
        key = ordered_bits(x);
    }
    key_out[i] = key;
    val_out[i] = r;
}

// Dense-genome fallback: rebuild keys as position << 32 | coarse descending time from the
// position-sorted records, so that the next sort also orders sites internally.
__global__ void __launch_bounds__(256) rekey_pos_time_kernel(SortedView sv, uint64_t *__restrict__ key_out,
    uint32_t *__restrict__ val_out)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sv.n) {
        return;
    }
    const uint32_t r = sv.val[i];
    const bool marker = sv.has_kept && r < sv.S0;
    const uint64_t t = marker ? 0ull : ((~ordered_bits(sv.rec[r].time)) >> 32);
    key_out[i] = ((uint64_t) sv.key.k32[i] << 32) | t;
    val_out[i] = r;
}

} // namespace mspb
