// msprime_b200: K5, the table writer (sm_100a).  Included by kernels.cuh.
//
// Replaces mutgen_populate_tables (lib/mutgen.c:1027-1074) and the allele strings of
// infinite_alleles_make_allele (lib/mutgen.c:431-447) / the matrix model's aliasing of its
// allele table (lib/mutgen.c:184-237).  One pass over the position-sorted records writes
// every fixed-width column coalesced and assembles the ragged state bytes in shared memory;
// byte offsets come from a tile scan chained across tiles by the look-back in common.cuh.
//
// The kernel is specialised on <model, kept rows present, offset width>: the common case
// (no kept rows) carries no trace of the kept-row handling, and every record is classified
// once, the result carried in registers between the length pass and the write pass.
#pragma once

namespace mspb {

constexpr int K5_ITEMS = 4;
constexpr int K5_TILE = SCAN_THREADS * K5_ITEMS;
constexpr int K5_STAGE = 12 * 1024; // bytes of one ragged column staged per tile
// 8 worker warps + 1 warp that only chains the tile's byte counts to the earlier tiles'
// (decoupled look-back), so that no worker ever waits for the chain before the very end
constexpr int K5_THREADS = SCAN_THREADS + 32;

constexpr uint32_t EM_HEAD = 1u, EM_MUT = 2u, EM_NEW = 4u, EM_SINGLE = 8u;

// two ASCII digits of p < 100, tens digit in the low byte
MSPB_DEVFN uint32_t two_digits(uint32_t p)
{
    const uint32_t t = (p * 205u) >> 11; // p / 10 for p < 1029
    return 0x3030u + t + ((p - t * 10u) << 8);
}

// the eight decimal digits of v < 10^8, zero padded, most significant digit in byte 0
MSPB_DEVFN uint64_t eight_digits(uint32_t v)
{
    const uint32_t a = v / 10000u, b = v - a * 10000u;
    const uint32_t a1 = a / 100u, a0 = a - a1 * 100u;
    const uint32_t b1 = b / 100u, b0 = b - b1 * 100u;
    const uint32_t lo = two_digits(a1) | (two_digits(a0) << 16);
    const uint32_t hi = two_digits(b1) | (two_digits(b0) << 16);
    return (uint64_t) lo | ((uint64_t) hi << 32);
}

// decimal text of v, n = num_digits(v) bytes (snprintf "%" PRIu64, lib/mutgen.c:431-447)
MSPB_DEVFN void write_number(char *dst, uint64_t v, uint32_t n)
{
    if (v < 100000000ull) {
        const uint64_t d = eight_digits((uint32_t) v) >> (8u * (8u - n));
        const uint32_t lo = (uint32_t) d, hi = (uint32_t) (d >> 32);
#pragma unroll
        for (uint32_t b = 0; b < 4; b++) {
            if (b < n) {
                dst[b] = (char) (lo >> (8u * b));
            }
        }
#pragma unroll
        for (uint32_t b = 0; b < 4; b++) {
            if (b + 4 < n) {
                dst[b + 4] = (char) (hi >> (8u * b));
            }
        }
    } else {
        write_digits(dst, v, n);
    }
}

MSPB_DEVFN void copy_bytes32(char *dst, const char *src, uint32_t n)
{
    for (uint32_t k = 0; k < n; k++) {
        dst[k] = src[k];
    }
}

template <bool OFF32>
MSPB_DEVFN void store_offset(void *col, uint64_t row, uint64_t v)
{
    if (OFF32) {
        static_cast<uint32_t *>(col)[row] = (uint32_t) v;
    } else {
        static_cast<uint64_t *>(col)[row] = v;
    }
}

// Tile-local exclusive offsets of one length column held in shared memory (striped by
// record), returning the tile total.  Thread t scans the K5_ITEMS consecutive records it
// owns in blocked order; lens[] is overwritten with the offsets.
template <typename SumT>
MSPB_DEVFN uint64_t tile_offsets(uint32_t *lens, uint64_t *scan_tmp)
{
    static_assert(K5_ITEMS == 4, "one 16-byte shared-memory access per thread");
    uint4 *mine = reinterpret_cast<uint4 *>(lens) + threadIdx.x;
    const uint4 l = *mine;
    const SumT sum = (SumT) l.x + l.y + l.z + l.w;
    SumT total;
    const SumT run = block_exclusive_scan<SumT>(sum, reinterpret_cast<SumT *>(scan_tmp), &total);
    uint4 o;
    o.x = (uint32_t) run;
    o.y = (uint32_t) (run + l.x);
    o.z = (uint32_t) (run + l.x + l.y);
    o.w = (uint32_t) (run + l.x + l.y + l.z);
    *mine = o;
    return (uint64_t) total;
}

// Copies a tile's staged bytes (stage[0 .. total), 16-byte aligned) to global memory
// at byte offset g0 with aligned 16-byte stores.  Source and destination differ by
// g0 mod 16, so each 16-byte chunk is assembled from five shared-memory words with funnel
// shifts; the (at most two) partial chunks at the ends go byte by byte.
MSPB_DEVFN void flush_stage(const char *stage, char *dst_base, uint64_t g0, uint32_t total)
{
    const uint32_t shift = (uint32_t) (g0 & 15ull);
    const uint32_t span = shift + total;
    char *aligned = dst_base + (g0 - shift);
    const uint32_t *words = reinterpret_cast<const uint32_t *>(stage);
    const uint32_t rot = 8u * ((16u - shift) & 3u); // bit offset of a chunk's first byte in its first word
    for (uint32_t lo = threadIdx.x * 16u; lo < span; lo += SCAN_THREADS * 16u) {
        if (lo >= shift && lo + 16u <= span) {
            const uint32_t w0 = (lo - shift) >> 2;
            const uint32_t x0 = words[w0], x1 = words[w0 + 1], x2 = words[w0 + 2], x3 = words[w0 + 3],
                           x4 = words[w0 + 4];
            uint4 o;
            o.x = __funnelshift_r(x0, x1, rot);
            o.y = __funnelshift_r(x1, x2, rot);
            o.z = __funnelshift_r(x2, x3, rot);
            o.w = __funnelshift_r(x3, x4, rot);
            *reinterpret_cast<uint4 *>(aligned + lo) = o;
        } else {
            const uint32_t hi = lo + 16u < span ? lo + 16u : span;
            for (uint32_t b = lo < shift ? shift : lo; b < hi; b++) {
                aligned[b] = stage[b - shift];
            }
        }
    }
}

// Both state columns of a tile in one scan (lengths packed as derived | ancestral << 32);
// used when there are no kept rows, where a tile's state bytes are far below 2^32.
MSPB_DEVFN void tile_offsets_pair(uint32_t *der, uint32_t *anc, uint64_t *scan_tmp, uint64_t *der_total,
    uint64_t *anc_total)
{
    static_assert(K5_ITEMS == 4, "one 16-byte shared-memory access per thread");
    uint4 *md = reinterpret_cast<uint4 *>(der) + threadIdx.x;
    uint4 *ma = reinterpret_cast<uint4 *>(anc) + threadIdx.x;
    const uint4 d = *md, a = *ma;
    const uint64_t p0 = (uint64_t) d.x | ((uint64_t) a.x << 32), p1 = (uint64_t) d.y | ((uint64_t) a.y << 32);
    const uint64_t p2 = (uint64_t) d.z | ((uint64_t) a.z << 32), p3 = (uint64_t) d.w | ((uint64_t) a.w << 32);
    uint64_t total;
    const uint64_t r0 = block_exclusive_scan<uint64_t>(p0 + p1 + p2 + p3, scan_tmp, &total);
    const uint64_t r1 = r0 + p0, r2 = r1 + p1, r3 = r2 + p2;
    uint4 od, oa;
    od.x = (uint32_t) r0;
    od.y = (uint32_t) r1;
    od.z = (uint32_t) r2;
    od.w = (uint32_t) r3;
    oa.x = (uint32_t) (r0 >> 32);
    oa.y = (uint32_t) (r1 >> 32);
    oa.z = (uint32_t) (r2 >> 32);
    oa.w = (uint32_t) (r3 >> 32);
    *md = od;
    *ma = oa;
    *der_total = total & 0xffffffffull;
    *anc_total = total >> 32;
}

struct Bytes4 {
    uint64_t der, mmd, anc, smd;
    MSPB_DEVFN Bytes4 operator+(const Bytes4 &o) const
    {
        Bytes4 r;
        r.der = der + o.der;
        r.mmd = mmd + o.mmd;
        r.anc = anc + o.anc;
        r.smd = smd + o.smd;
        return r;
    }
};

// The SLiM model's stacked state of new mutation i (lib/mutgen.c:355-413): the parent's
// derived state, a comma if that is not empty, the mutation id; and the parent's metadata plus
// one 17-byte record.  Written back to front along the parent chain: new ancestors contribute
// their own id / record, the first kept ancestor (or the kept site, when the chain ends without
// one: lib/mutgen.c:1190-1194) contributes its whole state and metadata.
MSPB_DEVFN void slim_write_state(const SortedView &sv, const ResolveView &rv, const KeptView &kept,
    const ModelView &model, uint64_t i, bool single, char *der_end, char *md_end)
{
    uint64_t cur = i;
    for (;;) {
        const uint32_t r = sv.val[cur];
        const uint64_t id = model.start_allele + (sv.has_kept ? sv.draw_scan[cur] : (uint32_t) cur);
        const uint32_t n = num_digits(id);
        der_end -= n;
        write_digits(der_end, id, n);
        {
            const int32_t type_id = model.slim_type, subpop = -1;
            const int32_t origin = model.slim_generation - (int32_t) sv.rec[r].time;
            const float sel = 0.0f;
            const int8_t nuc = -1;
            char rec17[SLIM_MD_SIZE];
            memcpy(rec17, &type_id, 4);
            memcpy(rec17 + 4, &sel, 4);
            memcpy(rec17 + 8, &subpop, 4);
            memcpy(rec17 + 12, &origin, 4);
            memcpy(rec17 + 16, &nuc, 1);
            md_end -= SLIM_MD_SIZE;
            for (int b = 0; b < SLIM_MD_SIZE; b++) {
                md_end[b] = rec17[b];
            }
        }
        const int32_t p = single ? -1 : rv.parent[cur];
        const char *bs = nullptr, *bm = nullptr; // the base that ends the chain, if it is reached
        uint64_t bl = 0, bml = 0;
        bool ends = false;
        if (p < 0) {
            ends = true;
            if (sv.has_kept) {
                uint64_t h = cur; // this site's first record: a kept site's marker
                while (h > 0 && sv.site_row[h - 1] == sv.site_row[h]) {
                    h--;
                }
                const uint32_t r0 = sv.val[h];
                if (r0 < sv.S0) {
                    bs = kept.site_anc + kept.site_anc_off[r0];
                    bl = kept.site_anc_off[r0 + 1] - kept.site_anc_off[r0];
                    bm = kept.site_md + kept.site_md_off[r0];
                    bml = kept.site_md_off[r0 + 1] - kept.site_md_off[r0];
                }
            }
        } else {
            const uint32_t rp = sv.val[p];
            if (sv.has_kept && rp < sv.S0 + sv.M0) {
                const uint64_t km = rp - sv.S0;
                ends = true;
                bs = kept.mut_der + kept.mut_der_off[km];
                bl = kept.mut_der_off[km + 1] - kept.mut_der_off[km];
                bm = kept.mut_md + kept.mut_md_off[km];
                bml = kept.mut_md_off[km + 1] - kept.mut_md_off[km];
            }
        }
        if (ends) {
            if (bl > 0) {
                *--der_end = ',';
                copy_bytes(der_end - bl, bs, bl);
            }
            copy_bytes(md_end - bml, bm, bml);
            return;
        }
        *--der_end = ','; // a new parent's state is never empty
        cur = (uint64_t) p;
    }
}

// Fills the ChainFree table for an infinite-alleles call without kept rows: lane k finds the
// first record whose derived (ancestral) number has more than k digits.  One warp; runs beside
// K4 on the context's second stream.
__global__ void k5_prefix_table_kernel(const uint32_t *__restrict__ site_row, const Counts3 *p1_total, uint64_t n,
    uint64_t start_allele, const uint64_t *dev_bases, uint32_t fixed_len, const uint32_t *dense, ChainFree *cf)
{
    const uint32_t k = threadIdx.x;
    if (dev_bases != nullptr) {
        start_allele += dev_bases[2];
    }
    if (k == 0) {
        // numbers must not wrap around 2^64 inside the call (they do in the reference, whose
        // counter is a uint64 too: the look-back path reproduces that)
        cf->enabled = (*dense == 0u && (fixed_len > 0 || start_allele <= ~0ull - (2 * n + 2))) ? 1u : 0u;
        cf->fixed_len = fixed_len;
        cf->der_upto[0] = cf->anc_upto[0] = 0;
    }
    if (k < 1 || k > 20 || fixed_len > 0) {
        return;
    }
    uint64_t pw = 1;
    bool all = false; // 10^k does not fit 64 bits: every number has at most k digits
    for (uint32_t j = 0; j < k; j++) {
        if (pw > ~0ull / 10ull) {
            all = true;
            break;
        }
        pw *= 10ull;
    }
    const uint64_t S = p1_total->s;
    for (int which = 0; which < 2; which++) {
        // first record i with start + i + site_row[i] + (derived ? 1 : 0) >= 10^k
        const uint64_t add = which == 0 ? 1ull : 0ull;
        uint64_t lo = 0, hi = n;
        while (!all && lo < hi) {
            const uint64_t mid = (lo + hi) >> 1;
            if (start_allele + mid + site_row[mid] + add >= pw) {
                hi = mid;
            } else {
                lo = mid + 1;
            }
        }
        const uint64_t i = all ? n : lo;
        if (which == 0) {
            cf->der_upto[k] = i;
        } else {
            // heads before record i
            uint64_t h = S;
            if (i < n) {
                const bool head = i == 0 || site_row[i - 1] != site_row[i];
                h = (uint64_t) site_row[i] + (head ? 0u : 1u);
            }
            cf->anc_upto[k] = h;
        }
    }
}

// MODEL: 0 matrix, 1 infinite alleles, 2 SLiM.  KEPT: kept site / mutation rows are present.
template <int MODEL, bool KEPT, bool OFF32>
__global__ void __launch_bounds__(K5_THREADS, 4) k5_emit_kernel(SortedView sv, ResolveView rv, KeptView kept,
    ModelView model, CallView call, OutView out, LbWord *lb, uint32_t num_tiles, const Counts3 *p1_total,
    Bytes4 *total, const uint32_t *dense, const uint32_t *collisions, int *err)
{
    __shared__ uint64_t sm[9];
    __shared__ LbVal<2, true> bcast, s_agg;
    __shared__ __align__(16) uint32_t s_der[K5_TILE];
    __shared__ __align__(16) uint32_t s_anc[K5_TILE];
    // metadata columns carry bytes when rows are kept or the model writes metadata (SLiM)
    constexpr bool MD = KEPT || MODEL == 2;
    __shared__ __align__(16) uint32_t s_mmd[MD ? K5_TILE : 4];
    __shared__ __align__(16) uint32_t s_smd[MD ? K5_TILE : 4];
    __shared__ __align__(16) char st_der[K5_STAGE + 32];
    __shared__ __align__(16) char st_anc[K5_STAGE + 32];
    if ((*dense != 0u && sv.key.mode == KEY32_POS) || *collisions != 0u) {
        return; // the host sorts again (dense genome) or redraws the colliding positions first
    }
    if (out.dev_bases != nullptr) {
        out.site_base += (uint32_t) out.dev_bases[0];
        out.mut_base += (uint32_t) out.dev_bases[1];
        model.start_allele += out.dev_bases[2];
    }
    // without kept rows the tile's byte prefix may follow from its first record's index alone
    const bool chain_free = !KEPT && MODEL != 2 && out.chain_free != nullptr && out.chain_free->enabled != 0u;
    if (threadIdx.x >= SCAN_THREADS) {
        // ---- the look-back warp: meets the workers at their barriers, otherwise only chains
        __syncthreads(); // lengths are in shared memory
        for (int b = 0; b < (MD ? 12 : 3); b++) {
            __syncthreads(); // the barriers of the tile scans
        }
        __syncthreads(); // offsets and the tile's totals are visible
        const LbVal<2, true> aw = s_agg;
        if (!chain_free) {
            const LbVal<2, true> pw = lookback_warp<2, true>(lb, blockIdx.x, aw);
            if (threadIdx.x == SCAN_THREADS) {
                bcast = pw;
            }
        }
        const uint64_t der = (uint64_t) aw.w[0][0] | ((uint64_t) aw.w[0][1] << 32);
        const uint64_t anc = (uint64_t) aw.w[1][0] | ((uint64_t) aw.w[1][1] << 32);
        const bool no_md = (aw.w[0][2] | aw.w[0][3] | aw.w[1][2] | aw.w[1][3]) == 0u;
        if (!out.dry_run && !(der <= K5_STAGE && anc <= K5_STAGE && no_md)) {
            __syncthreads(); // unstaged tiles: the workers wait for the prefix before writing
        }
        __syncthreads(); // rows and bytes are done
        return;
    }
    const uint64_t base = (uint64_t) blockIdx.x * K5_TILE;
    const uint64_t last = sv.n - 1;
    // ---- all loads of the tile up front: slots and the site rows around every record
    // (P1's site rows give the site boundaries: no key compares here), then the record
    // gathers, one random sector each
    uint32_t meta[K5_ITEMS]; // EM_* | derived length << 8 | ancestral length << 16 (infinite alleles)
    uint32_t va[K5_ITEMS]; // infinite alleles: draws before the record; matrix: derived allele index
    uint32_t vb[K5_ITEMS]; // matrix: ancestral allele index of a new site
    uint32_t srow[K5_ITEMS];
    uint32_t pos32[K5_ITEMS]; // 32-bit position keys are read with everything else (wider ones later)
    Rec rcs[K5_ITEMS];
    uint32_t heads_before = 0; // thread 0: site rows opened before the tile's first record
    {
        uint32_t slot[K5_ITEMS], before[K5_ITEMS], after[K5_ITEMS];
#pragma unroll
        for (int k = 0; k < K5_ITEMS; k++) {
            const uint64_t i = base + (uint32_t) k * SCAN_THREADS + threadIdx.x;
            const uint64_t ic = i < last ? i : last;
            slot[k] = sv.val[ic];
            pos32[k] = sv.key.mode == KEY32_POS ? sv.key.k32[ic] : 0u;
            srow[k] = sv.site_row[ic];
            before[k] = sv.site_row[ic > 0 ? ic - 1 : 0];
            after[k] = sv.site_row[ic < last ? ic + 1 : last];
        }
#pragma unroll
        for (int k = 0; k < K5_ITEMS; k++) {
            rcs[k] = sv.rec[slot[k]];
        }
#pragma unroll
        for (int k = 0; k < K5_ITEMS; k++) {
            const uint64_t i = base + (uint32_t) k * SCAN_THREADS + threadIdx.x;
            const bool head = i == 0 || before[k] != srow[k];
            const bool single = head && (i >= last || after[k] != srow[k]);
            const bool is_mut = !KEPT || slot[k] >= sv.S0;
            const bool is_new = !KEPT || slot[k] >= sv.S0 + sv.M0;
            meta[k] = i < sv.n ? (head ? EM_HEAD : 0u) | (is_mut ? EM_MUT : 0u) | (is_new ? EM_NEW : 0u)
                                     | (single ? EM_SINGLE : 0u)
                               : 0u;
            va[k] = slot[k];
        }
        heads_before = srow[0] + ((base == 0 || before[0] != srow[0]) ? 0u : 1u);
    }
    // ---- lengths to shared memory
#pragma unroll
    for (int k = 0; k < K5_ITEMS; k++) {
        const uint32_t j = (uint32_t) k * SCAN_THREADS + threadIdx.x;
        const uint64_t i = base + j;
        const uint32_t m = meta[k], r = va[k];
        const bool head = (m & EM_HEAD) != 0u;
        uint32_t d = 0, a = 0, md = 0, sd = 0;
        vb[k] = 0;
        if (m & EM_NEW) {
            // a head that is a mutation opens a new site (kept sites start with their marker)
            if (MODEL == 2) {
                // SLiM: the id counts new mutations only; a new site's ancestral state is empty
                const uint32_t db = KEPT ? sv.draw_scan[i] : (uint32_t) i;
                va[k] = db;
                if (m & EM_SINGLE) {
                    d = num_digits(model.start_allele + db);
                    md = SLIM_MD_SIZE;
                } else {
                    d = (uint32_t) rv.state[i];
                    md = (uint32_t) rv.site_state[i];
                }
            } else if (MODEL == 1) {
                const uint32_t db = KEPT ? sv.draw_scan[i] : (uint32_t) i + srow[k] + 1u - (head ? 1u : 0u);
                const uint64_t root_number = model.start_allele + db;
                va[k] = db;
                d = num_digits(root_number + (head ? 1u : 0u));
                if (head) {
                    a = num_digits(root_number);
                }
                meta[k] = m | (d << 8) | (a << 16);
            } else {
                int32_t s, ss = 0;
                if (m & EM_SINGLE) {
                    // a new site with one new mutation: root draw, then one transition
                    double ur, ut;
                    if (rv.replay != nullptr) {
                        const uint64_t db = sv.draws_before(i);
                        ur = phase_b_uniform(rv, db, err);
                        ut = phase_b_uniform(rv, db + 1, err);
                    } else {
                        allele_uniforms(Philox(call.seed), call.stream, rcs[k].ord, &ur, &ut);
                    }
                    ss = (int32_t) select_prob(ur, model.num_alleles, model.root);
                    s = (int32_t) select_prob(ut, model.num_alleles, model.matrix + (uint64_t) ss * model.num_alleles);
                } else {
                    s = rv.state[i];
                    if (head) {
                        ss = rv.site_state[i];
                    }
                }
                va[k] = (uint32_t) s;
                vb[k] = (uint32_t) ss;
                d = (uint32_t) (model.allele_off[s + 1] - model.allele_off[s]);
                if (head) {
                    a = (uint32_t) (model.allele_off[ss + 1] - model.allele_off[ss]);
                }
            }
        } else if (KEPT && m != 0u) {
            if (!(m & EM_MUT)) {
                a = (uint32_t) (kept.site_anc_off[r + 1] - kept.site_anc_off[r]);
                sd = (uint32_t) (kept.site_md_off[r + 1] - kept.site_md_off[r]);
            } else {
                const uint64_t km = r - sv.S0;
                d = (uint32_t) (kept.mut_der_off[km + 1] - kept.mut_der_off[km]);
                md = (uint32_t) (kept.mut_md_off[km + 1] - kept.mut_md_off[km]);
            }
        }
        s_der[j] = d;
        s_anc[j] = a;
        if (MD) {
            s_mmd[j] = md;
            s_smd[j] = sd;
        }
    }
    __syncthreads();
    // ---- offsets inside the tile
    Bytes4 agg = Bytes4();
    if (MD) {
        agg.der = tile_offsets<uint64_t>(s_der, sm);
        agg.anc = tile_offsets<uint64_t>(s_anc, sm);
        agg.mmd = tile_offsets<uint64_t>(s_mmd, sm);
        agg.smd = tile_offsets<uint64_t>(s_smd, sm);
    } else {
        tile_offsets_pair(s_der, s_anc, sm, &agg.der, &agg.anc);
    }
    LbVal<2, true> aw; // half 0: derived, mutation metadata; half 1: ancestral, site metadata
    aw.w[0][0] = (uint32_t) agg.der;
    aw.w[0][1] = (uint32_t) (agg.der >> 32);
    aw.w[0][2] = (uint32_t) agg.mmd;
    aw.w[0][3] = (uint32_t) (agg.mmd >> 32);
    aw.w[1][0] = (uint32_t) agg.anc;
    aw.w[1][1] = (uint32_t) (agg.anc >> 32);
    aw.w[1][2] = (uint32_t) agg.smd;
    aw.w[1][3] = (uint32_t) (agg.smd >> 32);
    if (threadIdx.x == 0) {
        s_agg = aw;
        if (chain_free) {
            // what the look-back warp would hand over at the end is known now
            uint64_t pd, pa;
            out.chain_free->prefix(base, heads_before, &pd, &pa);
            LbVal<2, true> pw;
            pw.w[0][0] = (uint32_t) pd;
            pw.w[0][1] = (uint32_t) (pd >> 32);
            pw.w[0][2] = pw.w[0][3] = 0u;
            pw.w[1][0] = (uint32_t) pa;
            pw.w[1][1] = (uint32_t) (pa >> 32);
            pw.w[1][2] = pw.w[1][3] = 0u;
            bcast = pw;
        }
    }
    __syncthreads(); // offsets are read striped below, by other threads than wrote them; totals out
    // ---- offsets across tiles.  When the tile's state bytes fit the stage (always, unless
    // kept rows carry long strings) nothing below needs the prefix until the very end:
    // the look-back warp chains the tiles while the workers write rows and assemble bytes.
    // (a tile of long states does not fit the stage whatever the model: 1024 alleles of 13 bytes or more,
    // i.e. allele numbers from 10^12 on or a matrix model with long alleles, go straight to global memory)
    const bool staged = agg.der <= K5_STAGE && agg.anc <= K5_STAGE && agg.mmd == 0 && agg.smd == 0;
    Bytes4 pre = Bytes4();
    if (out.dry_run) {
        // sizing pass (models whose state lengths are not bounded up front): totals only
        __syncthreads();
        if (blockIdx.x + 1 == num_tiles && threadIdx.x == 0) {
            const LbVal<2, true> pw = bcast;
            pre.der = (uint64_t) pw.w[0][0] | ((uint64_t) pw.w[0][1] << 32);
            pre.mmd = (uint64_t) pw.w[0][2] | ((uint64_t) pw.w[0][3] << 32);
            pre.anc = (uint64_t) pw.w[1][0] | ((uint64_t) pw.w[1][1] << 32);
            pre.smd = (uint64_t) pw.w[1][2] | ((uint64_t) pw.w[1][3] << 32);
            *total = pre + agg;
        }
        return;
    }
    if (!staged) {
        __syncthreads(); // (the look-back warp has the prefix)
        const LbVal<2, true> pw = bcast;
        pre.der = (uint64_t) pw.w[0][0] | ((uint64_t) pw.w[0][1] << 32);
        pre.mmd = (uint64_t) pw.w[0][2] | ((uint64_t) pw.w[0][3] << 32);
        pre.anc = (uint64_t) pw.w[1][0] | ((uint64_t) pw.w[1][1] << 32);
        pre.smd = (uint64_t) pw.w[1][2] | ((uint64_t) pw.w[1][3] << 32);
    }
    // ---- rows and bytes (the remaining loads, positions and parents, all in flight first)
    double posv[K5_ITEMS];
    int32_t parv[K5_ITEMS];
#pragma unroll
    for (int k = 0; k < K5_ITEMS; k++) {
        const uint64_t i = base + (uint32_t) k * SCAN_THREADS + threadIdx.x;
        const uint32_t m = meta[k];
        posv[k] = sv.key.mode == KEY32_POS ? (double) pos32[k] : ((m & EM_HEAD) ? sv.key.position(i) : 0.0);
        parv[k] = ((m & EM_MUT) && !(m & EM_SINGLE)) ? rv.parent[i] : -1;
    }
    // (the loop is instantiated twice, for the stage in shared memory and for the table in global memory:
    // with one pointer for both the state bytes went out as generic 64-bit stores, ST.E.U8)
    auto emit_rows = [&](char *der_base, char *anc_base) {
#pragma unroll
    for (int k = 0; k < K5_ITEMS; k++) {
        const uint32_t j = (uint32_t) k * SCAN_THREADS + threadIdx.x;
        const uint64_t i = base + j;
        const uint32_t m = meta[k];
        const uint32_t site_row = srow[k];
        if (m & EM_HEAD) {
            char *anc_dst = anc_base + s_anc[j];
            out.site_position[site_row] = posv[k];
            if (KEPT && !(m & EM_MUT)) {
                const uint32_t r = va[k];
                const uint64_t a0 = kept.site_anc_off[r], m0 = kept.site_md_off[r];
                copy_bytes(anc_dst, kept.site_anc + a0, kept.site_anc_off[r + 1] - a0);
                copy_bytes(out.site_md + pre.smd + s_smd[j], kept.site_md + m0, kept.site_md_off[r + 1] - m0);
            } else if (MODEL == 0) {
                const uint64_t a0 = model.allele_off[vb[k]];
                copy_bytes32(anc_dst, model.alleles + a0, (uint32_t) (model.allele_off[vb[k] + 1] - a0));
            } else if (MODEL == 1) {
                write_number(anc_dst, model.start_allele + va[k], (m >> 16) & 0xffu);
            }
        }
        if (m & EM_MUT) {
            const uint32_t row = KEPT ? sv.mut_scan[i] : (uint32_t) i;
            const int32_t p = parv[k];
            const int32_t par = p < 0 ? -1 : (int32_t) ((KEPT ? sv.mut_scan[p] : (uint32_t) p) + out.mut_base);
            char *der_dst = der_base + s_der[j];
            out.mut_site[row] = (int32_t) (site_row + out.site_base);
            out.mut_node[row] = rcs[k].node;
            out.mut_time[row] = rcs[k].time;
            out.mut_parent[row] = par;
            if (KEPT && !(m & EM_NEW)) {
                const uint64_t km = va[k] - sv.S0;
                const uint64_t d0 = kept.mut_der_off[km], m0 = kept.mut_md_off[km];
                copy_bytes(der_dst, kept.mut_der + d0, kept.mut_der_off[km + 1] - d0);
                copy_bytes(out.mut_md + pre.mmd + s_mmd[j], kept.mut_md + m0, kept.mut_md_off[km + 1] - m0);
            } else if (MODEL == 0) {
                const uint64_t d0 = model.allele_off[va[k]];
                copy_bytes32(der_dst, model.alleles + d0, (uint32_t) (model.allele_off[va[k] + 1] - d0));
            } else if (MODEL == 1) {
                write_number(der_dst, model.start_allele + va[k] + ((m & EM_HEAD) ? 1u : 0u), (m >> 8) & 0xffu);
            } else {
                const uint32_t dl = (m & EM_SINGLE) ? num_digits(model.start_allele + va[k]) : (uint32_t) rv.state[i];
                const uint32_t ml = (m & EM_SINGLE) ? (uint32_t) SLIM_MD_SIZE : (uint32_t) rv.site_state[i];
                slim_write_state(sv, rv, kept, model, i, (m & EM_SINGLE) != 0u, der_dst + dl,
                    out.mut_md + pre.mmd + s_mmd[j] + ml);
            }
        }
    }
    };
    if (staged) {
        emit_rows(st_der, st_anc);
    } else {
        emit_rows(out.mut_der + pre.der, out.site_anc + pre.anc);
    }
    __syncthreads();
    if (staged) {
        const LbVal<2, true> pw = bcast;
        pre.der = (uint64_t) pw.w[0][0] | ((uint64_t) pw.w[0][1] << 32);
        pre.mmd = (uint64_t) pw.w[0][2] | ((uint64_t) pw.w[0][3] << 32);
        pre.anc = (uint64_t) pw.w[1][0] | ((uint64_t) pw.w[1][1] << 32);
        pre.smd = (uint64_t) pw.w[1][2] | ((uint64_t) pw.w[1][3] << 32);
    }
    // ---- the offset columns, which are what needed the prefix (and, in a sharded call that has exchanged its
    // byte counts before this kernel, the bytes of the slices before this one: ob)
    Bytes4 ob = Bytes4();
    if (out.dev_byte_bases != nullptr) {
        ob.anc = out.dev_byte_bases[0];
        ob.smd = out.dev_byte_bases[1];
        ob.der = out.dev_byte_bases[2];
        ob.mmd = out.dev_byte_bases[3];
    }
#pragma unroll
    for (int k = 0; k < K5_ITEMS; k++) {
        const uint32_t j = (uint32_t) k * SCAN_THREADS + threadIdx.x;
        const uint32_t m = meta[k];
        if (m & EM_HEAD) {
            store_offset<OFF32>(out.site_anc_off, srow[k], ob.anc + pre.anc + s_anc[j]);
            if (out.write_md_off) {
                store_offset<OFF32>(out.site_md_off, srow[k], ob.smd + pre.smd + (MD ? s_smd[j] : 0u));
            }
        }
        if (m & EM_MUT) {
            const uint32_t row = KEPT ? sv.mut_scan[base + j] : (uint32_t) (base + j);
            store_offset<OFF32>(out.mut_der_off, row, ob.der + pre.der + s_der[j]);
            if (out.write_md_off) {
                store_offset<OFF32>(out.mut_md_off, row, ob.mmd + pre.mmd + (MD ? s_mmd[j] : 0u));
            }
        }
    }
    if (blockIdx.x + 1 == num_tiles && threadIdx.x == 0) {
        *total = pre + agg;
        store_offset<OFF32>(out.site_anc_off, p1_total->s, ob.anc + pre.anc + agg.anc);
        store_offset<OFF32>(out.mut_der_off, p1_total->m, ob.der + pre.der + agg.der);
        if (out.write_md_off) {
            store_offset<OFF32>(out.site_md_off, p1_total->s, ob.smd + pre.smd + agg.smd);
            store_offset<OFF32>(out.mut_md_off, p1_total->m, ob.mmd + pre.mmd + agg.mmd);
        }
    }
    if (staged) {
        flush_stage(st_der, out.mut_der, pre.der, (uint32_t) agg.der);
        flush_stage(st_anc, out.site_anc, pre.anc, (uint32_t) agg.anc);
    }
}

} // namespace mspb
