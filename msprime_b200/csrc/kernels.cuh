// msprime_b200: the mutation-overlay kernels (sm_100a).
//
// Pipeline (DESIGN.md has the data layout and the byte counts):
//   K0  child-sorted edge interval index            (once per tree sequence)
//   K1  per-edge Poisson count + exclusive scan     <- lib/mutgen.c:1103-1120
//   K2  position / time placement                   <- lib/mutgen.c:1121-1151
//   K3  radix sort by (position, time desc)         <- AVL map + per-site qsort
//   K4  per-site parent search + allele assignment  <- lib/mutgen.c:1160-1224
//   K5  site / mutation table emission              <- lib/mutgen.c:1027-1074
//
// Records: one per kept site ("marker"), kept mutation and new mutation, in that
// slot order, so a record's kind follows from its slot index and slot order is the
// final tie-break (kept before new, kept by row id, new by creation order).
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
    // child-sorted interval index (K0): edges of child c are ci_*[child_start[c] .. child_start[c+1])
    const uint32_t *child_start;
    const double *ci_left, *ci_right;
    const int32_t *ci_parent;
};

struct KeptView {
    uint64_t num_sites, num_mutations; // 0 when keep is off
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
    uint64_t start_allele;
};

struct CallView {
    double start_time, end_time;
    double slice_left, slice_right;
    int discrete;
    int sliced;
    uint64_t seed;
    int pos_shift; // key = (uint64(pos) << 32 | coarse time) when >= 0; else key = ordered_bits(pos)
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

__global__ void k0_permute_kernel(TreeSeqView ts, const uint32_t *perm, double *ci_left, double *ci_right,
    int32_t *ci_parent, uint32_t *child_start)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < ts.num_edges) {
        const uint32_t e = perm[i];
        ci_left[i] = ts.edge_left[e];
        ci_right[i] = ts.edge_right[e];
        ci_parent[i] = ts.edge_parent[e];
        // CSR starts: child c begins at the first i whose child is >= c
        const uint32_t c = (uint32_t) ts.edge_child[e];
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

// parent of node u at position x in the local tree, or -1
MSPB_DEVFN int32_t parent_at(const TreeSeqView &ts, int32_t u, double x)
{
    uint32_t lo = ts.child_start[u], hi = ts.child_start[u + 1];
    const uint32_t first = lo;
    // last interval with left <= x
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (ts.ci_left[mid] <= x) {
            lo = mid + 1;
        } else {
            hi = mid;
        }
    }
    if (lo == first) {
        return -1;
    }
    lo--;
    return x < ts.ci_right[lo] ? ts.ci_parent[lo] : -1;
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
    d.child = ts.edge_child[e];
    const double tc = ts.node_time[d.child];
    const double tp = ts.node_time[ts.edge_parent[e]];
    d.branch_start = fmax(call.start_time, tc);
    d.branch_end = fmin(call.end_time, tp);
    d.mass_lo = map.mass_upto(d.left);
    d.mass = map.mass_upto(d.right) - d.mass_lo;
    return d;
}

MSPB_DEVFN void place_one(const Philox &ph, const RateMapView &map, const EdgeDraw &d, uint64_t e, uint32_t k,
    double *pos, double *time)
{
    const uint4 r = ph((uint32_t) e, (uint32_t) (e >> 32), k, TAG_PLACE);
    const double up = u53(r.x, r.y), ut = u53(r.z, r.w);
    *pos = map.position_at(d.mass_lo + up * d.mass, d.left, d.right);
    double t = d.branch_start + ut * (d.branch_end - d.branch_start);
    if (!(t < d.branch_end)) {
        t = nextafter(d.branch_end, -1.7976931348623157e308);
    }
    *time = t;
}

// count[e] = Poisson(branch length x rate mass); local[e] = how many fall in the slice
__global__ void __launch_bounds__(256) k1_count_kernel(TreeSeqView ts, RateMapView map, CallView call,
    uint32_t *__restrict__ count, uint32_t *__restrict__ local)
{
    const uint64_t e = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= ts.num_edges) {
        return;
    }
    const EdgeDraw d = edge_draw(ts, map, call, e);
    const double mu = (d.branch_end - d.branch_start) * d.mass;
    PhiloxSeq rng(call.seed, e, TAG_COUNT);
    const uint32_t n = poisson_native(rng, mu);
    count[e] = n;
    if (call.sliced) {
        uint32_t in = 0;
        if (n > 0 && d.left < call.slice_right && d.right > call.slice_left) {
            const Philox ph(call.seed);
            for (uint32_t k = 0; k < n; k++) {
                double x, t;
                place_one(ph, map, d, e, k, &x, &t);
                in += (x >= call.slice_left && x < call.slice_right) ? 1u : 0u;
            }
        }
        local[e] = in;
    }
}

MSPB_DEVFN uint64_t make_key(const CallView &call, double pos, double time, bool marker)
{
    if (call.pos_shift >= 0) {
        const uint64_t t = marker ? 0ull : ((~ordered_bits(time)) >> 32);
        return ((uint64_t) pos << 32) | t;
    }
    return ordered_bits(pos);
}

// ---------------------------------------------------------------- K2 (native)
// goff: exclusive scan of count (global ordinal); loff: exclusive scan of the in-slice
// counts (== goff when not sliced).  Records go to slots base + loff[e] + j.
__global__ void __launch_bounds__(256) k2_place_kernel(TreeSeqView ts, RateMapView map, CallView call,
    const uint32_t *__restrict__ count, const uint32_t *__restrict__ goff, const uint32_t *__restrict__ loff,
    uint32_t base, double *__restrict__ rec_pos, double *__restrict__ rec_time, int32_t *__restrict__ rec_node,
    uint32_t *__restrict__ rec_ord, uint64_t *__restrict__ key, uint32_t *__restrict__ val)
{
    const uint64_t e = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= ts.num_edges) {
        return;
    }
    const uint32_t n = count[e];
    if (n == 0) {
        return;
    }
    const EdgeDraw d = edge_draw(ts, map, call, e);
    const Philox ph(call.seed);
    uint32_t slot = base + loff[e];
    const uint32_t g0 = goff[e];
    for (uint32_t k = 0; k < n; k++) {
        double x, t;
        place_one(ph, map, d, e, k, &x, &t);
        if (call.sliced && !(x >= call.slice_left && x < call.slice_right)) {
            continue;
        }
        rec_pos[slot] = x;
        rec_time[slot] = t;
        rec_node[slot] = d.child;
        rec_ord[slot - base] = g0 + k;
        key[slot] = make_key(call, x, t, false);
        val[slot] = slot;
        slot++;
    }
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

// One thread walks the stream in the reference's order: edges in table order, rate
// intervals left to right, Poisson first, then two uniforms per mutation.
__global__ void replay_resolve_kernel(TreeSeqView ts, RateMapView map, CallView call, const double *stream,
    uint64_t stream_len, ReplayPairs pairs, ReplayTotals *totals, int *err)
{
    if (blockIdx.x != 0 || threadIdx.x != 0) {
        return;
    }
    ReplayStream s { stream, stream_len, 0, 0 };
    uint64_t np = 0, nm = 0;
    for (uint64_t e = 0; e < ts.num_edges && !s.exhausted; e++) {
        double left = ts.edge_left[e];
        const double edge_right = ts.edge_right[e];
        const double branch_start = fmax(call.start_time, ts.node_time[ts.edge_child[e]]);
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
                if (np >= pairs.capacity || nm + n > 0x7fffffffull) {
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
                nm += n;
                s.pos += 2ull * n;
                if (s.pos > s.len) {
                    s.exhausted = 1;
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
    const double *__restrict__ stream, ReplayPairs pairs, uint64_t num_pairs, uint32_t base,
    double *__restrict__ rec_pos, double *__restrict__ rec_time, int32_t *__restrict__ rec_node,
    uint32_t *__restrict__ rec_ord, uint64_t *__restrict__ key, uint32_t *__restrict__ val, int *err)
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
        rec_pos[slot] = x;
        rec_time[slot] = t;
        rec_node[slot] = child;
        rec_ord[slot - base] = slot - base;
        key[slot] = make_key(call, x, t, false);
        val[slot] = slot;
    }
}

// ---------------------------------------------------------------- kept records
// slots [0, S0): site markers; [S0, S0 + M0): kept mutations (lib/mutgen.c:974-1025)
__global__ void __launch_bounds__(256) fill_kept_kernel(KeptView kept, CallView call, double *__restrict__ rec_pos,
    double *__restrict__ rec_time, int32_t *__restrict__ rec_node, uint64_t *__restrict__ key,
    uint32_t *__restrict__ val, int *err)
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
        x = kept.site_position[kept.mut_site[m]];
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
    rec_pos[r] = x;
    rec_time[r] = t;
    rec_node[r] = node;
    key[r] = make_key(call, x, t, r < S0);
    val[r] = (uint32_t) r;
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
    const uint64_t *key;
    uint32_t *val; // slot of the record at each sorted index
    const double *rec_pos, *rec_time;
    const int32_t *rec_node;
    const uint32_t *rec_ord;
    uint64_t pos_mask; // bits of the key that hold the position
};

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
    const double ta = sv.rec_time[a], tb = sv.rec_time[b];
    if (ta != tb) {
        return ta > tb;
    }
    return a < b;
}

// Runs of equal keys are ordered only coarsely by the sort; finish them exactly.
__global__ void __launch_bounds__(256) finalize_runs_kernel(SortedView sv)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 >= sv.n) {
        return;
    }
    const uint64_t k = sv.key[i];
    if (sv.key[i + 1] != k || (i > 0 && sv.key[i - 1] == k)) {
        return;
    }
    uint64_t end = i + 2;
    while (end < sv.n && sv.key[end] == k) {
        end++;
    }
    // insertion sort of val[i .. end)
    for (uint64_t a = i + 1; a < end; a++) {
        const uint32_t v = sv.val[a];
        uint64_t b = a;
        while (b > i && rec_before(sv, v, sv.val[b - 1])) {
            sv.val[b] = sv.val[b - 1];
            b--;
        }
        sv.val[b] = v;
    }
}

MSPB_DEVFN bool is_head(const SortedView &sv, uint64_t i)
{
    return i == 0 || ((sv.key[i] ^ sv.key[i - 1]) & sv.pos_mask) != 0;
}

// ---------------------------------------------------------------- K4
// Per-record results of the per-site step, indexed by sorted position.
struct ResolveView {
    const uint32_t *site_scan; // sites before i
    const uint32_t *mut_scan; // mutation rows before i
    const uint32_t *draw_scan; // phase-B draws (or new alleles) before i
    int32_t *parent; // sorted index of the parent mutation, or -1
    int32_t *state; // matrix: allele index (new and kept; -1 = unknown kept allele)
    int32_t *site_state; // matrix: allele index of a new site's ancestral state (by sorted head index)
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

// One thread per site: root state for a new site, then for each mutation in order the
// nearest ancestor-or-self node already holding a mutation at this site gives the parent
// (lib/mutgen.c:1185-1216).  The walk uses the child-sorted interval index instead of a
// tree sweep and stops once the node is older than every earlier mutation's node.
__global__ void __launch_bounds__(128) k4_resolve_kernel(TreeSeqView ts, KeptView kept, ModelView model, CallView call,
    SortedView sv, ResolveView rv, int *err)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sv.n || !is_head(sv, i)) {
        return;
    }
    const uint64_t S0 = sv.S0, KM = sv.S0 + sv.M0;
    const uint32_t r0 = sv.val[i];
    const bool kept_site = r0 < S0;
    const uint64_t first = i + (kept_site ? 1 : 0);
    const bool matrix = model.kind == 0;
    const Philox ph(call.seed);

    int32_t anc = -1;
    if (matrix) {
        if (kept_site) {
            anc = kept.site_allele[r0];
        } else {
            double u;
            if (rv.replay != nullptr) {
                u = phase_b_uniform(rv, rv.draw_scan[i], err);
            } else {
                const uint32_t g = sv.rec_ord[r0 - KM];
                const uint4 w = ph(g, 0u, 0u, TAG_ALLELE);
                u = u53(w.z, w.w);
            }
            anc = (int32_t) select_prob(u, model.num_alleles, model.root);
            rv.site_state[i] = anc;
        }
    }
    const double x = sv.rec_pos[r0];
    double tmax = -INFINITY; // oldest node among the mutations seen so far
    for (uint64_t j = first; j < sv.n && (j == i || !is_head(sv, j)); j++) {
        const uint32_t r = sv.val[j];
        const int32_t node = sv.rec_node[r];
        int32_t par = -1;
        if (j > first) {
            int32_t u = node;
            while (u != -1) {
                if (ts.node_time[u] > tmax) {
                    break;
                }
                for (uint64_t q = j; q-- > first;) {
                    if (sv.rec_node[sv.val[q]] == u) {
                        par = (int32_t) q;
                        break;
                    }
                }
                if (par >= 0) {
                    break;
                }
                u = parent_at(ts, u, x);
            }
        }
        rv.parent[j] = par;
        if (matrix) {
            int32_t st;
            if (r < KM) {
                st = kept.mut_allele[r - S0];
            } else {
                const int32_t pa = par < 0 ? anc : rv.state[par];
                double u;
                if (rv.replay != nullptr) {
                    u = phase_b_uniform(rv, (uint64_t) rv.draw_scan[j] + ((j == i) ? 1u : 0u), err);
                } else {
                    const uint32_t g = sv.rec_ord[r - KM];
                    const uint4 w = ph(g, 0u, 0u, TAG_ALLELE);
                    u = u53(w.x, w.y);
                }
                if (pa < 0) {
                    set_error(err, -51); // MSP_ERR_UNKNOWN_ALLELE
                    st = 0;
                } else {
                    st = (int32_t) select_prob(u, model.num_alleles, model.matrix + (uint64_t) pa * model.num_alleles);
                }
            }
            rv.state[j] = st;
        }
        tmax = fmax(tmax, ts.node_time[node]);
    }
}

// ---------------------------------------------------------------- K5
MSPB_DEVFN uint32_t num_digits(uint64_t v)
{
    uint32_t n = 1;
    while (v >= 10) {
        v /= 10;
        n++;
    }
    return n;
}

MSPB_DEVFN void write_digits(char *dst, uint64_t v, uint32_t n)
{
    for (uint32_t k = n; k-- > 0;) {
        dst[k] = (char) ('0' + (v % 10));
        v /= 10;
    }
}

struct EmitView {
    SortedView sv;
    ResolveView rv;
    KeptView kept;
    ModelView model;

    MSPB_DEVFN bool head(uint64_t i) const { return is_head(sv, i); }
    MSPB_DEVFN bool is_mut(uint64_t i) const { return sv.val[i] >= sv.S0; }
    MSPB_DEVFN bool is_new(uint64_t i) const { return sv.val[i] >= sv.S0 + sv.M0; }
    MSPB_DEVFN bool new_site(uint64_t i) const { return head(i) && is_mut(i); }
    // infinite alleles: the number given to record i's root / mutation
    MSPB_DEVFN uint64_t root_number(uint64_t i) const { return model.start_allele + rv.draw_scan[i]; }
    MSPB_DEVFN uint64_t mut_number(uint64_t i) const
    {
        return model.start_allele + rv.draw_scan[i] + (new_site(i) ? 1u : 0u);
    }
    MSPB_DEVFN uint64_t der_len(uint64_t i) const
    {
        if (!is_mut(i)) {
            return 0;
        }
        const uint32_t r = sv.val[i];
        if (!is_new(i)) {
            const uint64_t m = r - sv.S0;
            return kept.mut_der_off[m + 1] - kept.mut_der_off[m];
        }
        if (model.kind == 0) {
            const int32_t s = rv.state[i];
            return model.allele_off[s + 1] - model.allele_off[s];
        }
        return num_digits(mut_number(i));
    }
    MSPB_DEVFN uint64_t mut_md_len(uint64_t i) const
    {
        if (!is_mut(i) || is_new(i)) {
            return 0;
        }
        const uint64_t m = sv.val[i] - sv.S0;
        return kept.mut_md_off[m + 1] - kept.mut_md_off[m];
    }
    MSPB_DEVFN uint64_t anc_len(uint64_t i) const
    {
        if (!head(i)) {
            return 0;
        }
        const uint32_t r = sv.val[i];
        if (r < sv.S0) {
            return kept.site_anc_off[r + 1] - kept.site_anc_off[r];
        }
        if (model.kind == 0) {
            const int32_t s = rv.site_state[i];
            return model.allele_off[s + 1] - model.allele_off[s];
        }
        return num_digits(root_number(i));
    }
    MSPB_DEVFN uint64_t site_md_len(uint64_t i) const
    {
        if (!head(i) || sv.val[i] >= sv.S0) {
            return 0;
        }
        const uint32_t r = sv.val[i];
        return kept.site_md_off[r + 1] - kept.site_md_off[r];
    }
};

struct OutView {
    double *site_position;
    char *site_anc;
    uint64_t *site_anc_off;
    char *site_md;
    uint64_t *site_md_off;
    int32_t *mut_site, *mut_node, *mut_parent;
    double *mut_time;
    char *mut_der;
    uint64_t *mut_der_off;
    char *mut_md;
    uint64_t *mut_md_off;
};

MSPB_DEVFN void copy_bytes(char *dst, const char *src, uint64_t n)
{
    for (uint64_t k = 0; k < n; k++) {
        dst[k] = src[k];
    }
}

// byte offsets come from four scans over the sorted records (der, mut md, anc, site md)
__global__ void __launch_bounds__(256) k5_emit_kernel(EmitView ev, OutView out, const uint64_t *__restrict__ der_scan,
    const uint64_t *__restrict__ mmd_scan, const uint64_t *__restrict__ anc_scan, const uint64_t *__restrict__ smd_scan)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const SortedView &sv = ev.sv;
    if (i >= sv.n) {
        return;
    }
    const uint32_t r = sv.val[i];
    const bool head = ev.head(i);
    const uint32_t site_row = ev.rv.site_scan[i] + (head ? 1u : 0u) - 1u;
    if (head) {
        out.site_position[site_row] = sv.rec_pos[r];
        const uint64_t ao = anc_scan[i], mo = smd_scan[i];
        out.site_anc_off[site_row] = ao;
        out.site_md_off[site_row] = mo;
        if (r < sv.S0) {
            copy_bytes(out.site_anc + ao, ev.kept.site_anc + ev.kept.site_anc_off[r], ev.anc_len(i));
            copy_bytes(out.site_md + mo, ev.kept.site_md + ev.kept.site_md_off[r], ev.site_md_len(i));
        } else if (ev.model.kind == 0) {
            const int32_t s = ev.rv.site_state[i];
            copy_bytes(out.site_anc + ao, ev.model.alleles + ev.model.allele_off[s], ev.anc_len(i));
        } else {
            const uint64_t v = ev.root_number(i);
            write_digits(out.site_anc + ao, v, num_digits(v));
        }
    }
    if (r >= sv.S0) {
        const uint32_t row = ev.rv.mut_scan[i];
        const int32_t par = ev.rv.parent[i];
        out.mut_site[row] = (int32_t) site_row;
        out.mut_node[row] = sv.rec_node[r];
        out.mut_time[row] = sv.rec_time[r];
        out.mut_parent[row] = par < 0 ? -1 : (int32_t) ev.rv.mut_scan[par];
        const uint64_t d = der_scan[i], m = mmd_scan[i];
        out.mut_der_off[row] = d;
        out.mut_md_off[row] = m;
        if (!ev.is_new(i)) {
            const uint64_t km = r - sv.S0;
            copy_bytes(out.mut_der + d, ev.kept.mut_der + ev.kept.mut_der_off[km], ev.der_len(i));
            copy_bytes(out.mut_md + m, ev.kept.mut_md + ev.kept.mut_md_off[km], ev.mut_md_len(i));
        } else if (ev.model.kind == 0) {
            const int32_t s = ev.rv.state[i];
            copy_bytes(out.mut_der + d, ev.model.alleles + ev.model.allele_off[s], ev.der_len(i));
        } else {
            const uint64_t v = ev.mut_number(i);
            write_digits(out.mut_der + d, v, num_digits(v));
        }
    }
}

// After the sort: two kept sites at one position are an input error
// (lib/mutgen.c:875-879); in the continuous genome a new mutation must not share a
// position with any other record (the reference redraws, lib/mutgen.c:1127-1134).
__global__ void __launch_bounds__(256) post_sort_check_kernel(SortedView sv, int discrete, uint32_t *collisions, int *err)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sv.n || i == 0 || is_head(sv, i)) {
        return;
    }
    if (sv.val[i] < sv.S0) {
        set_error(err, -21); // MSP_ERR_DUPLICATE_SITE_POSITION
    }
    if (!discrete && (sv.val[i] >= sv.S0 + sv.M0 || sv.val[i - 1] >= sv.S0 + sv.M0)) {
        atomicAdd(collisions, 1u);
    }
}

} // namespace mspb
