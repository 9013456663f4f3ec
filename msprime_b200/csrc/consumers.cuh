// msprime_b200: the first consumers of the overlay's tables, on the device (sm_100a).
//
// what the reference's users do next with the tables this path writes
// (verification.py:6008-6013): the number of samples below a mutation and the site allele
// frequency spectrum, which tskit computes with a left-to-right sweep over the trees
//   tsk_treeseq_allele_frequency_spectrum   git-submodules/tskit/c/tskit/trees.c:3816
//   (TSK_STAT_SITE | TSK_STAT_POLARISED, one sample set, not span normalised; trees.c:3527-3540)
//   tsk_variant_decode                      git-submodules/tskit/c/tskit/genotypes.c:474
// maintaining num_samples[u] along the path to the root at every edge insertion and removal:
// E x depth DEPENDENT updates, which one GPU thread would take a minute over at bench size.
//
// Here below(u, x), the number of samples under node u in the tree covering x, is kept as a
// STEP FUNCTION of x for every node:
//   f_u = [u is a sample] + sum over edges (u, c, l, r) of f_c restricted to [l, r)
// so the functions are built children first.  A LEVEL is the set of nodes whose children are all
// done (level(u) = 1 + max level of its children); one level is a handful of data-parallel
// launches: every edge into a node of the level contributes the changes of its child's function
// inside [l, r) (the value at l, the child's own steps, minus the value at r), the level's events
// are sorted by (node, x) with the library's radix sort, and ONE running sum over the sorted
// events gives all the functions of the level (the events of a node add up to zero, so the sum
// needs no segment resets).  The index depends on nodes and edges only: it is built once per
// loaded tree sequence, and a query is a binary search in the node's function.
//
// Coordinates are the RANKS of the distinct edge end points (every step of every function sits
// on one), so a key is (node << 32 | rank) for any genome, discrete or continuous.
#pragma once

namespace mspb {

struct SampleFnView {
    uint64_t num_nodes;
    const uint32_t *is_sample; // 0 / 1 per node
    const uint32_t *foff, *fcnt; // per node: its steps are fx / fv[foff .. foff + fcnt)
    const uint32_t *fx; // rank of the coordinate where the function changes
    const uint32_t *fv; // value from there on
    const uint64_t *xs; // ordered_bits of the distinct edge end points, ascending
    uint32_t num_x;

    // number of steps of node u at ranks <= r (inclusive) or < r
    MSPB_DEVFN uint32_t steps_upto(uint32_t u, uint32_t r, bool inclusive) const
    {
        const uint32_t a = foff[u];
        uint32_t lo = 0, hi = fcnt[u];
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            const uint32_t v = fx[a + mid];
            if (inclusive ? v <= r : v < r) {
                lo = mid + 1;
            } else {
                hi = mid;
            }
        }
        return lo;
    }
    // f_u at rank r
    MSPB_DEVFN uint32_t value_at(uint32_t u, uint32_t r) const
    {
        const uint32_t k = steps_upto(u, r, true);
        return k > 0 ? fv[foff[u] + k - 1] : is_sample[u];
    }
    // rank of the last end point <= x, or -1 when x lies left of every edge
    MSPB_DEVFN int64_t rank_of(double x) const
    {
        const uint64_t key = ordered_bits(x);
        uint32_t lo = 0, hi = num_x;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if (xs[mid] <= key) {
                lo = mid + 1;
            } else {
                hi = mid;
            }
        }
        return (int64_t) lo - 1;
    }
};

// ---------------------------------------------------------------- build
__global__ void __launch_bounds__(256) sf_flags_kernel(uint64_t n, const uint32_t *__restrict__ flags,
    uint32_t *__restrict__ is_sample, uint32_t *num_samples)
{
    const uint64_t u = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t s = (u < n && (flags[u] & 1u) != 0u) ? 1u : 0u; // TSK_NODE_IS_SAMPLE
    if (u < n) {
        is_sample[u] = s;
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, s != 0u);
    if ((threadIdx.x & 31u) == 0u && bal != 0u) {
        atomicAdd(num_samples, (uint32_t) __popc(bal));
    }
}

__global__ void __launch_bounds__(256) sf_endpoint_keys_kernel(TreeSeqView ts, uint64_t *__restrict__ keys)
{
    const uint64_t e = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (e < ts.num_edges) {
        keys[2 * e] = ordered_bits(ts.edge_left[e]);
        keys[2 * e + 1] = ordered_bits(ts.edge_right[e]);
    }
}

struct SfHead {
    const uint64_t *keys;
    MSPB_DEVFN uint32_t operator()(uint64_t i) const { return (i == 0 || keys[i] != keys[i - 1]) ? 1u : 0u; }
};

__global__ void __launch_bounds__(256) sf_unique_kernel(uint64_t n, const uint64_t *__restrict__ keys,
    const uint32_t *__restrict__ pos, uint64_t *__restrict__ xs)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && (i == 0 || keys[i] != keys[i - 1])) {
        xs[pos[i]] = keys[i];
    }
}

// ranks of each edge's end points among the distinct coordinates (they are in the list)
__global__ void __launch_bounds__(256) sf_edge_ranks_kernel(TreeSeqView ts, const uint64_t *__restrict__ xs,
    uint32_t num_x, uint32_t *__restrict__ lrank, uint32_t *__restrict__ rrank)
{
    const uint64_t e = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= ts.num_edges) {
        return;
    }
    for (int side = 0; side < 2; side++) {
        const uint64_t key = ordered_bits(side == 0 ? ts.edge_left[e] : ts.edge_right[e]);
        uint32_t lo = 0, hi = num_x;
        while (lo < hi) { // first rank with xs[rank] >= key
            const uint32_t mid = (lo + hi) >> 1;
            if (xs[mid] < key) {
                lo = mid + 1;
            } else {
                hi = mid;
            }
        }
        (side == 0 ? lrank : rrank)[e] = lo;
    }
}

// one round of level[parent] = max(level[parent], level[child] + 1)
__global__ void __launch_bounds__(256) sf_relax_levels_kernel(TreeSeqView ts, uint32_t *level, uint32_t *changed)
{
    const uint64_t e = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= ts.num_edges) {
        return;
    }
    const uint32_t want = level[ts.edge_child[e]] + 1u;
    const int32_t p = ts.edge_parent[e];
    if (level[p] < want) {
        atomicMax(&level[p], want);
        *changed = 1u;
    }
}

// The deepest level.  (A round of the relaxation above works in place, so a node can climb several levels in one
// round when its descendants' updates happen to land first: the number of rounds does NOT bound the levels.)
__global__ void __launch_bounds__(256) sf_max_level_kernel(uint64_t n, const uint32_t *__restrict__ level, uint32_t *out)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t v = i < n ? level[i] : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const uint32_t w = __shfl_xor_sync(0xffffffffu, v, o);
        v = w > v ? w : v;
    }
    if ((threadIdx.x & 31) == 0 && v > 0u) {
        atomicMax(out, v);
    }
}

// sort key of an edge: (level of its parent, parent), and how many edges every level has
__global__ void __launch_bounds__(256) sf_edge_level_keys_kernel(TreeSeqView ts, const uint32_t *__restrict__ level,
    uint64_t *__restrict__ keys, uint32_t *__restrict__ level_count, uint32_t max_levels)
{
    const uint64_t e = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= ts.num_edges) {
        return;
    }
    const uint32_t p = (uint32_t) ts.edge_parent[e];
    const uint32_t lv = level[p];
    keys[e] = ((uint64_t) lv << 32) | p;
    atomicAdd(&level_count[lv < max_levels ? lv : max_levels - 1], 1u);
}

struct SfLevelEdges {
    TreeSeqView ts;
    const uint32_t *order; // edge ids of the level, grouped by parent
    const uint32_t *lrank, *rrank;
};

// events an edge of the level contributes: the value at its left end, its child's steps strictly
// inside (l, r), and minus the value at its right end
__global__ void __launch_bounds__(256) sf_count_events_kernel(SfLevelEdges le, uint64_t n, SampleFnView fn,
    uint32_t *__restrict__ count, uint32_t *__restrict__ first_step)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) {
        return;
    }
    const uint32_t e = le.order[i];
    const uint32_t c = (uint32_t) le.ts.edge_child[e];
    const uint32_t lo = fn.steps_upto(c, le.lrank[e], true);
    const uint32_t hi = fn.steps_upto(c, le.rrank[e], false);
    first_step[i] = lo;
    count[i] = 2u + (hi > lo ? hi - lo : 0u);
}

struct SfArray {
    const uint32_t *a;
    MSPB_DEVFN uint32_t operator()(uint64_t i) const { return a[i]; }
};

__global__ void __launch_bounds__(256) sf_emit_events_kernel(SfLevelEdges le, uint64_t n, SampleFnView fn,
    const uint32_t *__restrict__ count, const uint32_t *__restrict__ first_step, const uint32_t *__restrict__ offset,
    uint64_t *__restrict__ keys, uint32_t *__restrict__ delta)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) {
        return;
    }
    const uint32_t e = le.order[i];
    const uint32_t c = (uint32_t) le.ts.edge_child[e];
    const uint64_t p = (uint64_t) (uint32_t) le.ts.edge_parent[e] << 32;
    const uint32_t a = fn.foff[c], lo = first_step[i], inner = count[i] - 2u;
    uint32_t cur = lo > 0 ? fn.fv[a + lo - 1] : fn.is_sample[c];
    uint64_t o = offset[i];
    keys[o] = p | le.lrank[e];
    delta[o] = cur;
    o++;
    for (uint32_t t = 0; t < inner; t++, o++) {
        const uint32_t v = fn.fv[a + lo + t];
        keys[o] = p | fn.fx[a + lo + t];
        delta[o] = v - cur; // (two's complement: the running sum below wraps the same way)
        cur = v;
    }
    keys[o] = p | le.rrank[e];
    delta[o] = 0u - cur;
}

struct SfGather {
    const uint32_t *delta, *perm;
    MSPB_DEVFN uint32_t operator()(uint64_t i) const { return delta[perm[i]]; }
};

struct SfKeep {
    const uint64_t *keys;
    uint64_t n;
    MSPB_DEVFN uint32_t operator()(uint64_t i) const { return (i + 1 == n || keys[i] != keys[i + 1]) ? 1u : 0u; }
};

// the level's functions: the last event of every (node, x) carries the value from x on
__global__ void __launch_bounds__(256) sf_store_steps_kernel(uint64_t n, const uint64_t *__restrict__ keys,
    const uint32_t *__restrict__ delta, const uint32_t *__restrict__ perm, const uint32_t *__restrict__ before,
    const uint32_t *__restrict__ kept_before, const uint32_t *__restrict__ is_sample, uint32_t base,
    uint32_t *__restrict__ fx, uint32_t *__restrict__ fv, uint32_t *__restrict__ foff, uint32_t *__restrict__ fcnt)
{
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) {
        return;
    }
    const uint64_t key = keys[i];
    const uint32_t u = (uint32_t) (key >> 32);
    if (i == 0 || (uint32_t) (keys[i - 1] >> 32) != u) {
        foff[u] = base + kept_before[i]; // the node's first kept step is at or after its first event
    }
    if (i + 1 == n || keys[i + 1] != key) {
        const uint32_t at = base + kept_before[i];
        fx[at] = (uint32_t) key;
        fv[at] = is_sample[u] + before[i] + delta[perm[i]];
        atomicAdd(&fcnt[u], 1u);
    }
}

// ---------------------------------------------------------------- queries on the resident result
template <bool OFF32>
MSPB_DEVFN uint64_t sf_load_off(const void *col, uint64_t row)
{
    return OFF32 ? (uint64_t) static_cast<const uint32_t *>(col)[row] : static_cast<const uint64_t *>(col)[row];
}

// below[m] = samples under mutation m's node in the tree covering its site; own[m] starts equal
__global__ void __launch_bounds__(256) sf_below_kernel(uint64_t num_mutations, const int32_t *__restrict__ mut_site,
    const int32_t *__restrict__ mut_node, const double *__restrict__ site_position, SampleFnView fn,
    uint32_t *__restrict__ below, uint32_t *__restrict__ own)
{
    const uint64_t m = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= num_mutations) {
        return;
    }
    const uint32_t u = (uint32_t) mut_node[m];
    const int64_t r = fn.rank_of(site_position[mut_site[m]]);
    const uint32_t b = r < 0 ? fn.is_sample[u] : fn.value_at(u, (uint32_t) r);
    below[m] = b;
    own[m] = b;
}

// a mutation takes the samples below it from its parent mutation (or from the ancestral state)
__global__ void __launch_bounds__(256) sf_own_kernel(uint64_t num_mutations, const int32_t *__restrict__ mut_parent,
    const uint32_t *__restrict__ below, uint32_t *__restrict__ own)
{
    const uint64_t m = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (m < num_mutations && mut_parent[m] >= 0) {
        atomicSub(&own[mut_parent[m]], below[m]);
    }
}

constexpr int AFS_SMEM_BINS = 4096;

// One thread per site (the first of its mutation rows): every derived state that differs from the
// ancestral one and is carried by some but not all samples adds one to afs[carriers]
// (trees.c:3527-3540).  States are compared as byte strings; equal states of one site are summed.
template <bool OFF32>
__global__ void __launch_bounds__(256) sf_afs_kernel(uint64_t num_mutations, const int32_t *__restrict__ mut_site,
    const uint32_t *__restrict__ own, const char *__restrict__ der, const void *__restrict__ der_off,
    const char *__restrict__ anc, const void *__restrict__ anc_off, uint32_t num_samples,
    unsigned long long *__restrict__ afs)
{
    __shared__ uint32_t bins[AFS_SMEM_BINS];
    for (int b = threadIdx.x; b < AFS_SMEM_BINS; b += blockDim.x) {
        bins[b] = 0u;
    }
    __syncthreads();
    // (a few blocks per SM walk the rows with a grid stride: every block flushes its bins once)
    for (uint64_t m0 = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; m0 < num_mutations;
         m0 += (uint64_t) gridDim.x * blockDim.x) {
        if (m0 != 0 && mut_site[m0 - 1] == mut_site[m0]) {
            continue;
        }
        const int32_t s = mut_site[m0];
        uint64_t m1 = m0 + 1;
        while (m1 < num_mutations && mut_site[m1] == s) {
            m1++;
        }
        const uint64_t a0 = sf_load_off<OFF32>(anc_off, (uint64_t) s), alen = sf_load_off<OFF32>(anc_off, (uint64_t) s + 1) - a0;
        for (uint64_t m = m0; m < m1; m++) {
            const uint64_t d0 = sf_load_off<OFF32>(der_off, m), dlen = sf_load_off<OFF32>(der_off, m + 1) - d0;
            // the first row of its state speaks for all rows with that state
            bool first = true;
            for (uint64_t q = m0; q < m && first; q++) {
                const uint64_t q0 = sf_load_off<OFF32>(der_off, q);
                if (sf_load_off<OFF32>(der_off, q + 1) - q0 == dlen) {
                    bool same = true;
                    for (uint64_t k = 0; k < dlen && same; k++) {
                        same = der[q0 + k] == der[d0 + k];
                    }
                    first = !same;
                }
            }
            if (!first) {
                continue;
            }
            bool ancestral = alen == dlen;
            for (uint64_t k = 0; k < dlen && ancestral; k++) {
                ancestral = anc[a0 + k] == der[d0 + k];
            }
            if (ancestral) {
                continue;
            }
            uint32_t carriers = own[m];
            for (uint64_t q = m + 1; q < m1; q++) {
                const uint64_t q0 = sf_load_off<OFF32>(der_off, q);
                if (sf_load_off<OFF32>(der_off, q + 1) - q0 == dlen) {
                    bool same = true;
                    for (uint64_t k = 0; k < dlen && same; k++) {
                        same = der[q0 + k] == der[d0 + k];
                    }
                    if (same) {
                        carriers += own[q];
                    }
                }
            }
            if (carriers > 0u && carriers < num_samples) {
                if (carriers < (uint32_t) AFS_SMEM_BINS) {
                    atomicAdd(&bins[carriers], 1u);
                } else {
                    atomicAdd(&afs[carriers], 1ull);
                }
            }
        }
    }
    __syncthreads();
    for (int b = threadIdx.x; b < AFS_SMEM_BINS; b += blockDim.x) {
        if (bins[b] != 0u && (uint32_t) b <= num_samples) {
            atomicAdd(&afs[b], (unsigned long long) bins[b]);
        }
    }
}

// ---------------------------------------------------------------- variant decoding
// tsk_variant_decode (git-submodules/tskit/c/tskit/genotypes.c:474): the alleles of a site are its
// ancestral state followed by the derived states in order of first appearance; every sample carries
// the allele of the nearest mutation of the site on its path to the root (the last such row of the
// table when a node holds several), allele 0 when there is none.

struct SfIsSample {
    const uint32_t *is_sample;
    MSPB_DEVFN uint32_t operator()(uint64_t u) const { return is_sample[u]; }
};

__global__ void __launch_bounds__(256) sf_sample_list_kernel(uint64_t n, const uint32_t *__restrict__ is_sample,
    const uint32_t *__restrict__ pos, uint32_t *__restrict__ samples)
{
    const uint64_t u = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (u < n && is_sample[u] != 0u) {
        samples[pos[u]] = (uint32_t) u;
    }
}

struct SiteRows {
    uint32_t *row_begin, *row_end; // mutation rows of site (site_begin + k)
    double *tmax; // oldest node among its mutations
    uint32_t *allele; // per mutation row: index of its derived state in the site's allele list
    uint32_t *walk_id; // per mutation row: its node as the walk names it
};

// one thread per site of the range: its rows, the oldest mutated node, the allele list
template <bool OFF32>
__global__ void __launch_bounds__(128) sf_site_alleles_kernel(TreeSeqView ts, uint64_t site_begin, uint64_t num_sites,
    uint64_t num_mutations, const int32_t *__restrict__ mut_site, const int32_t *__restrict__ mut_node,
    const char *__restrict__ der, const void *__restrict__ der_off, const char *__restrict__ anc,
    const void *__restrict__ anc_off, SiteRows out)
{
    const uint64_t k = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= num_sites) {
        return;
    }
    const int64_t s = (int64_t) (site_begin + k);
    uint64_t lo = 0, hi = num_mutations; // first row with mut_site >= s
    while (lo < hi) {
        const uint64_t mid = (lo + hi) >> 1;
        if ((int64_t) mut_site[mid] < s) {
            lo = mid + 1;
        } else {
            hi = mid;
        }
    }
    const uint64_t m0 = lo;
    uint64_t m1 = m0;
    double tmax = -INFINITY;
    while (m1 < num_mutations && (int64_t) mut_site[m1] == s) {
        tmax = fmax(tmax, ts.node_time[mut_node[m1]]);
        out.walk_id[m1] = walk_id(ts, mut_node[m1]);
        m1++;
    }
    out.row_begin[k] = (uint32_t) m0;
    out.row_end[k] = (uint32_t) m1;
    out.tmax[k] = tmax;
    const uint64_t a0 = sf_load_off<OFF32>(anc_off, (uint64_t) s), alen = sf_load_off<OFF32>(anc_off, (uint64_t) s + 1) - a0;
    uint32_t distinct = 1; // the ancestral state is allele 0
    for (uint64_t m = m0; m < m1; m++) {
        const uint64_t d0 = sf_load_off<OFF32>(der_off, m), dlen = sf_load_off<OFF32>(der_off, m + 1) - d0;
        bool same = alen == dlen;
        for (uint64_t q = 0; q < dlen && same; q++) {
            same = anc[a0 + q] == der[d0 + q];
        }
        uint32_t idx = 0;
        if (!same) {
            // an earlier row with the same state names the allele, else it is a new one
            bool found = false;
            for (uint64_t e = m0; e < m && !found; e++) {
                const uint64_t e0 = sf_load_off<OFF32>(der_off, e);
                if (sf_load_off<OFF32>(der_off, e + 1) - e0 == dlen) {
                    bool eq = true;
                    for (uint64_t q = 0; q < dlen && eq; q++) {
                        eq = der[e0 + q] == der[d0 + q];
                    }
                    if (eq) {
                        idx = out.allele[e];
                        found = true;
                    }
                }
            }
            if (!found) {
                idx = distinct++;
            }
        }
        out.allele[m] = idx;
    }
}

// one thread per (site, sample): up from the sample until a node of the site's mutations
__global__ void __launch_bounds__(256) sf_decode_kernel(TreeSeqView ts, uint64_t site_begin, uint64_t num_sites,
    const double *__restrict__ site_position, const uint32_t *__restrict__ samples, uint64_t num_samples, SiteRows sr,
    int32_t *__restrict__ genotypes)
{
    const uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t k = blockIdx.y;
    if (j >= num_samples || k >= num_sites) {
        return;
    }
    const uint32_t m0 = sr.row_begin[k], m1 = sr.row_end[k];
    int32_t g = 0;
    if (m1 > m0) {
        const double x = site_position[site_begin + k], tmax = sr.tmax[k];
        const uint32_t u = samples[j];
        WalkNode w = walk_node(ts, (int32_t) u);
        double tu = ts.node_time[u];
        for (;;) {
            if (tu > tmax) {
                break; // older than every mutated node of the site
            }
            int64_t hit = -1;
            for (uint32_t m = m0; m < m1; m++) {
                if (sr.walk_id[m] == w.id) {
                    hit = m; // the last row on this node wins
                }
            }
            if (hit >= 0) {
                g = (int32_t) sr.allele[hit];
                break;
            }
            if (!parent_at(ts, &w, x, &tu)) {
                break;
            }
        }
    }
    genotypes[k * num_samples + j] = g;
}

} // namespace mspb
