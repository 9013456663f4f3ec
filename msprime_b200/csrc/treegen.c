/* msprime_b200: synthetic tree-sequence generator (benchmark / test INPUTS only).
 *
 * The reference's benchmark inputs come from msprime.sim_ancestry, which needs GSL
 * and the Hudson simulator (lib/msprime.c); neither exists in this image or on the
 * GPU box.  This file builds coalescent-with-recombination tree
 * sequences with the SMC' left-to-right construction instead: a Kingman tree at
 * position 0, then one subtree-prune-and-regraft (SPR) move per recombination
 * breakpoint.  The prune point is uniform on the tree's total branch length; the
 * regraft time has the exact coalescent hazard k(s) / (2 Ne) of the lineages that
 * remain; the regraft lineage is uniform among those alive at that time when at
 * most DFS_LIMIT lineages are alive, and otherwise the ancestor (alive at that
 * time) of a uniformly chosen leaf -- an O(depth) draw that favours large clades
 * slightly; it is only used near the leaves, where each lineage is touched less
 * than once per genome on average.
 *
 * Cost per SPR move is O(log n): branch lengths live in a Fenwick tree over node
 * slots, and the internal node times live in quantile buckets with a per-bucket
 * (count, sum) so that the integral of k(s) over a whole bucket is O(1).
 *
 * Output follows tskit's table requirements (tables.c:10474-10497): nodes numbered
 * so that id order is time order (samples first), edges sorted by (parent, child,
 * left), parent strictly older than child.  This is not part of the product path.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    uint64_t num_nodes;
    uint64_t num_samples;
    uint64_t num_edges;
    uint64_t num_trees;
    double *node_time;
    double *edge_left;
    double *edge_right;
    int32_t *edge_parent;
    int32_t *edge_child;
} treegen_result_t;

/* ---------------------------------------------------------------- random numbers */
typedef struct {
    uint64_t s[4];
} rng_t;

static uint64_t
splitmix(uint64_t *x)
{
    uint64_t z = (*x += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

static void
rng_seed(rng_t *r, uint64_t seed)
{
    for (int i = 0; i < 4; i++) {
        r->s[i] = splitmix(&seed);
    }
}

static inline uint64_t
rotl(uint64_t x, int k)
{
    return (x << k) | (x >> (64 - k));
}

static inline uint64_t
rng_next(rng_t *r) /* xoshiro256** */
{
    uint64_t *s = r->s;
    const uint64_t result = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
    s[2] ^= s[0];
    s[3] ^= s[1];
    s[1] ^= s[2];
    s[0] ^= s[3];
    s[2] ^= t;
    s[3] = rotl(s[3], 45);
    return result;
}

static inline double
rng_uniform(rng_t *r) /* [0, 1) */
{
    return (double) (rng_next(r) >> 11) * (1.0 / 9007199254740992.0);
}

static inline double
rng_exp(rng_t *r) /* Exp(1) */
{
    return -log(1.0 - rng_uniform(r));
}

static inline uint64_t
rng_below(rng_t *r, uint64_t n)
{
    return (uint64_t) (rng_uniform(r) * (double) n) % n;
}

/* ---------------------------------------------------------------- Fenwick tree of doubles */
typedef struct {
    uint64_t n, top; /* top = largest power of two <= n */
    double *tree; /* 1-based partial sums */
    double *val;
} fenwick_t;

static int
fenwick_init(fenwick_t *f, uint64_t n)
{
    f->n = n;
    f->top = 1;
    while (f->top * 2 <= n) {
        f->top *= 2;
    }
    f->tree = calloc(n + 1, sizeof(double));
    f->val = calloc(n + 1, sizeof(double));
    return (f->tree == NULL || f->val == NULL) ? -1 : 0;
}

/* bulk build from val[] */
static void
fenwick_build(fenwick_t *f)
{
    for (uint64_t i = 1; i <= f->n; i++) {
        f->tree[i] = f->val[i - 1];
    }
    for (uint64_t i = 1; i <= f->n; i++) {
        const uint64_t j = i + (i & (~i + 1));
        if (j <= f->n) {
            f->tree[j] += f->tree[i];
        }
    }
}

static inline void
fenwick_set(fenwick_t *f, uint64_t idx, double v)
{
    const double d = v - f->val[idx];
    f->val[idx] = v;
    for (uint64_t i = idx + 1; i <= f->n; i += i & (~i + 1)) {
        f->tree[i] += d;
    }
}

static inline double
fenwick_total(const fenwick_t *f)
{
    double s = 0;
    for (uint64_t i = f->n; i > 0; i -= i & (~i + 1)) {
        s += f->tree[i];
    }
    return s;
}

/* smallest index whose inclusive prefix sum exceeds x */
static inline uint64_t
fenwick_find(const fenwick_t *f, double x)
{
    uint64_t pos = 0;
    for (uint64_t step = f->top; step > 0; step >>= 1) {
        const uint64_t nxt = pos + step;
        if (nxt <= f->n && f->tree[nxt] <= x) {
            pos = nxt;
            x -= f->tree[nxt];
        }
    }
    return pos < f->n ? pos : f->n - 1;
}

/* ---------------------------------------------------------------- time buckets */
typedef struct {
    double *t; /* sorted */
    uint32_t n, cap;
    double sum;
} bucket_t;

typedef struct {
    uint32_t nb;
    double *lo; /* nb + 1 boundaries; lo[0] = 0, lo[nb] = +inf */
    bucket_t *b;
} timeset_t;

static uint32_t
timeset_bucket(const timeset_t *ts, double t)
{
    uint32_t lo = 0, hi = ts->nb; /* last bucket with lo[b] <= t */
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (ts->lo[mid] <= t) {
            lo = mid;
        } else {
            hi = mid;
        }
    }
    return lo;
}

/* index of the first entry > t in bucket */
static inline uint32_t
bucket_upper(const bucket_t *b, double t)
{
    uint32_t lo = 0, hi = b->n;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (b->t[mid] <= t) {
            lo = mid + 1;
        } else {
            hi = mid;
        }
    }
    return lo;
}

static int
timeset_insert(timeset_t *ts, double t)
{
    bucket_t *b = &ts->b[timeset_bucket(ts, t)];
    if (b->n == b->cap) {
        const uint32_t cap = b->cap < 16 ? 16 : b->cap * 2;
        double *p = realloc(b->t, cap * sizeof(double));
        if (p == NULL) {
            return -1;
        }
        b->t = p;
        b->cap = cap;
    }
    const uint32_t i = bucket_upper(b, t);
    memmove(b->t + i + 1, b->t + i, (b->n - i) * sizeof(double));
    b->t[i] = t;
    b->n++;
    b->sum += t;
    return 0;
}

static void
timeset_remove(timeset_t *ts, double t)
{
    bucket_t *b = &ts->b[timeset_bucket(ts, t)];
    uint32_t i = bucket_upper(b, t);
    if (i == 0 || b->t[i - 1] != t) {
        return; /* cannot happen */
    }
    i--;
    memmove(b->t + i, b->t + i + 1, (b->n - i - 1) * sizeof(double));
    b->n--;
    b->sum -= t;
}

/* ---------------------------------------------------------------- growable edge output */
typedef struct {
    uint64_t n, cap;
    double *left, *right;
    uint32_t *parent, *child; /* output node ids (creation order) */
} edgebuf_t;

static int
edgebuf_push(edgebuf_t *e, double l, double r, uint32_t p, uint32_t c)
{
    if (e->n == e->cap) {
        const uint64_t cap = e->cap == 0 ? (1u << 16) : e->cap * 2;
        double *nl = realloc(e->left, cap * sizeof(double));
        if (nl != NULL) {
            e->left = nl;
        }
        double *nr = realloc(e->right, cap * sizeof(double));
        if (nr != NULL) {
            e->right = nr;
        }
        uint32_t *np = realloc(e->parent, cap * sizeof(uint32_t));
        if (np != NULL) {
            e->parent = np;
        }
        uint32_t *nc = realloc(e->child, cap * sizeof(uint32_t));
        if (nc != NULL) {
            e->child = nc;
        }
        if (nl == NULL || nr == NULL || np == NULL || nc == NULL) {
            return -1;
        }
        e->cap = cap;
    }
    e->left[e->n] = l;
    e->right[e->n] = r;
    e->parent[e->n] = p;
    e->child[e->n] = c;
    e->n++;
    return 0;
}

typedef struct {
    uint64_t n, cap;
    double *t;
} timebuf_t;

static int
timebuf_push(timebuf_t *b, double t)
{
    if (b->n == b->cap) {
        const uint64_t cap = b->cap == 0 ? (1u << 16) : b->cap * 2;
        double *p = realloc(b->t, cap * sizeof(double));
        if (p == NULL) {
            return -1;
        }
        b->t = p;
        b->cap = cap;
    }
    b->t[b->n++] = t;
    return 0;
}

/* ---------------------------------------------------------------- sorting helpers */
typedef struct {
    double t;
    uint32_t id;
} timed_t;

static int
cmp_timed(const void *a, const void *b)
{
    const timed_t *x = a, *y = b;
    if (x->t != y->t) {
        return x->t < y->t ? -1 : 1;
    }
    return x->id < y->id ? -1 : (x->id > y->id);
}

static int
cmp_double(const void *a, const void *b)
{
    const double x = *(const double *) a, y = *(const double *) b;
    return x < y ? -1 : (x > y);
}

/* stable LSD radix sort of idx[] by key[idx], 16-bit digits over `bits` bits */
static int
radix_sort_index(uint64_t n, const uint64_t *key, uint32_t *idx, int bits)
{
    uint32_t *tmp = malloc(n * sizeof(uint32_t));
    uint64_t *count = malloc(65537 * sizeof(uint64_t));
    if (tmp == NULL || count == NULL) {
        free(tmp);
        free(count);
        return -1;
    }
    uint32_t *src = idx, *dst = tmp;
    for (int shift = 0; shift < bits; shift += 16) {
        memset(count, 0, 65537 * sizeof(uint64_t));
        for (uint64_t i = 0; i < n; i++) {
            count[((key[src[i]] >> shift) & 0xffff) + 1]++;
        }
        for (int d = 0; d < 65536; d++) {
            count[d + 1] += count[d];
        }
        for (uint64_t i = 0; i < n; i++) {
            dst[count[(key[src[i]] >> shift) & 0xffff]++] = src[i];
        }
        uint32_t *t = src;
        src = dst;
        dst = t;
    }
    if (src != idx) {
        memcpy(idx, src, n * sizeof(uint32_t));
    }
    free(tmp);
    free(count);
    return 0;
}

/* ---------------------------------------------------------------- the generator */
#define DFS_LIMIT 2048
#define NIL (-1)

typedef struct {
    int32_t *par;
    int32_t (*ch)[2];
    double *tm;
    double *eleft;
    uint32_t *outid;
} forest_t;

static inline void
replace_child(forest_t *f, int32_t p, int32_t old, int32_t new_)
{
    if (p >= 0) {
        f->ch[p][f->ch[p][0] == old ? 0 : 1] = new_;
    }
}

void
treegen_free(treegen_result_t *r)
{
    if (r != NULL) {
        free(r->node_time);
        free(r->edge_left);
        free(r->edge_right);
        free(r->edge_parent);
        free(r->edge_child);
        memset(r, 0, sizeof(*r));
    }
}

/* n sample nodes, genome [0, L), recombination rate r per unit length per generation,
 * pairwise coalescence rate 1 / two_ne.  Returns 0, or -1 on allocation failure,
 * -2 on bad arguments. */
int
treegen_smc(uint64_t n, double L, double r, double two_ne, uint64_t seed, int discrete, treegen_result_t *out)
{
    int ret = -1;
    rng_t rng;
    forest_t f;
    fenwick_t fw;
    timeset_t ts;
    edgebuf_t eb;
    timebuf_t tb;
    int32_t *stack = NULL, *live = NULL;
    double *sorted = NULL;
    timed_t *order = NULL;
    uint32_t *newid = NULL, *idx = NULL;
    uint64_t *key = NULL;
    uint8_t *used = NULL;
    uint64_t num_trees = 1;

    memset(&f, 0, sizeof(f));
    memset(&fw, 0, sizeof(fw));
    memset(&ts, 0, sizeof(ts));
    memset(&eb, 0, sizeof(eb));
    memset(&tb, 0, sizeof(tb));
    memset(out, 0, sizeof(*out));
    if (n < 2 || n > 0x3fffffffull || !(L > 0) || !(two_ne > 0) || r < 0) {
        return -2;
    }
    const uint64_t slots = 2 * n - 1;
    rng_seed(&rng, seed);
    f.par = malloc(slots * sizeof(int32_t));
    f.ch = malloc(slots * sizeof(*f.ch));
    f.tm = calloc(slots, sizeof(double));
    f.eleft = calloc(slots, sizeof(double));
    f.outid = malloc(slots * sizeof(uint32_t));
    live = malloc(n * sizeof(int32_t));
    stack = malloc((2 * DFS_LIMIT + 64) * sizeof(int32_t));
    if (!f.par || !f.ch || !f.tm || !f.eleft || !f.outid || !live || !stack || fenwick_init(&fw, slots) != 0) {
        goto done;
    }
    /* Kingman tree */
    for (uint64_t i = 0; i < slots; i++) {
        f.par[i] = NIL;
        f.ch[i][0] = f.ch[i][1] = NIL;
        f.outid[i] = (uint32_t) i;
    }
    for (uint64_t i = 0; i < n; i++) {
        live[i] = (int32_t) i;
        if (timebuf_push(&tb, 0.0) != 0) {
            goto done;
        }
    }
    {
        double t = 0;
        uint64_t nxt = n;
        for (uint64_t k = n; k > 1; k--) {
            t += rng_exp(&rng) * two_ne / ((double) k * (double) (k - 1) / 2.0);
            const uint64_t i = rng_below(&rng, k);
            uint64_t j = rng_below(&rng, k - 1);
            if (j >= i) {
                j++;
            }
            const int32_t a = live[i], b = live[j];
            f.par[a] = f.par[b] = (int32_t) nxt;
            f.ch[nxt][0] = a;
            f.ch[nxt][1] = b;
            f.tm[nxt] = t;
            if (timebuf_push(&tb, t) != 0) {
                goto done;
            }
            /* remove i and j (larger index first), append nxt */
            const uint64_t hi = i > j ? i : j, lo = i > j ? j : i;
            live[hi] = live[k - 1];
            live[lo] = live[k - 2];
            live[k - 2] = (int32_t) nxt;
            nxt++;
        }
    }
    int32_t root = (int32_t) (slots - 1);
    for (uint64_t i = 0; i < slots; i++) {
        fw.val[i] = f.par[i] >= 0 ? f.tm[f.par[i]] - f.tm[i] : 0.0;
    }
    fenwick_build(&fw);
    /* time buckets: quantiles of the initial internal node times (already sorted) */
    {
        const uint64_t ni = n - 1;
        uint32_t nb = (uint32_t) (ni / 256);
        nb = nb < 1 ? 1 : (nb > 8192 ? 8192 : nb);
        ts.nb = nb;
        ts.lo = malloc((nb + 1) * sizeof(double));
        ts.b = calloc(nb, sizeof(bucket_t));
        if (!ts.lo || !ts.b) {
            goto done;
        }
        ts.lo[0] = 0.0;
        for (uint32_t b = 1; b < nb; b++) {
            ts.lo[b] = f.tm[n + (uint64_t) b * ni / nb];
        }
        ts.lo[nb] = INFINITY;
        for (uint64_t i = n; i < slots; i++) {
            if (timeset_insert(&ts, f.tm[i]) != 0) {
                goto done;
            }
        }
    }

#define CLOSE_EDGE(c, x)                                                                                               \
    do {                                                                                                               \
        if (f.par[c] >= 0 && f.eleft[c] < (x)) {                                                                       \
            if (edgebuf_push(&eb, f.eleft[c], (x), f.outid[f.par[c]], f.outid[c]) != 0) {                             \
                goto done;                                                                                             \
            }                                                                                                          \
        }                                                                                                              \
    } while (0)

    double x = 0.0, last_bp = 0.0;
    while (r > 0) {
        const double total = fenwick_total(&fw);
        if (!(total > 0)) {
            break;
        }
        x += rng_exp(&rng) / (r * total);
        const double bp = discrete ? ceil(x) : x;
        if (bp >= L) {
            break;
        }
        if (bp > last_bp) {
            num_trees++;
            last_bp = bp;
        }
        /* prune point */
        int32_t u = (int32_t) fenwick_find(&fw, rng_uniform(&rng) * total);
        if (f.par[u] < 0) {
            continue; /* numerical edge: landed on the root's zero weight */
        }
        const int32_t p = f.par[u];
        double t = f.tm[u] + rng_uniform(&rng) * (f.tm[p] - f.tm[u]);
        const int32_t s = f.ch[p][0] == u ? f.ch[p][1] : f.ch[p][0];
        const int32_t g = f.par[p];
        CLOSE_EDGE(u, bp);
        CLOSE_EDGE(s, bp);
        CLOSE_EDGE(p, bp);
        f.par[s] = g;
        replace_child(&f, g, p, s);
        f.eleft[s] = bp;
        f.par[u] = NIL;
        f.par[p] = NIL;
        if (g < 0) {
            root = s;
        }
        fenwick_set(&fw, (uint64_t) u, 0.0);
        fenwick_set(&fw, (uint64_t) p, 0.0);
        fenwick_set(&fw, (uint64_t) s, g >= 0 ? f.tm[g] - f.tm[s] : 0.0);
        timeset_remove(&ts, f.tm[p]);

        /* regraft time: integral of k(s) from t to T equals an Exp(1) x two_ne draw,
         * with k(s) = n - 1 - #{internal node times <= s} */
        double area = rng_exp(&rng) * two_ne;
        double T = 0;
        uint64_t kT = 0;
        {
            uint32_t b = timeset_bucket(&ts, t);
            uint64_t below = 0; /* internal node times <= current point */
            for (uint32_t q = 0; q < b; q++) {
                below += ts.b[q].n;
            }
            const bucket_t *bk = &ts.b[b];
            uint32_t i = bucket_upper(bk, t);
            below += i;
            double cur = t;
            int found = 0;
            for (;;) {
                /* inside bucket b, from cur, entries i.. */
                for (; i < bk->n; i++) {
                    const double k = (double) (n - 1 - below);
                    const double w = bk->t[i] - cur;
                    if (k * w >= area) {
                        T = cur + area / k;
                        found = 1;
                        break;
                    }
                    area -= k * w;
                    cur = bk->t[i];
                    below++;
                }
                if (found) {
                    break;
                }
                /* to the end of this bucket */
                {
                    const double k = (double) (n - 1 - below);
                    const double hi = ts.lo[b + 1];
                    if (b + 1 == ts.nb || k * (hi - cur) >= area) {
                        T = cur + area / k;
                        break;
                    }
                    area -= k * (hi - cur);
                    cur = hi;
                }
                /* whole buckets in O(1) each */
                b++;
                while (b + 1 < ts.nb) {
                    const bucket_t *w = &ts.b[b];
                    const double k = (double) (n - 1 - below);
                    const double hi = ts.lo[b + 1];
                    const double a = k * (hi - cur) - (double) w->n * hi + w->sum;
                    if (a >= area) {
                        break;
                    }
                    area -= a;
                    below += w->n;
                    cur = hi;
                    b++;
                }
                bk = &ts.b[b];
                i = 0;
            }
            /* lineages alive at T */
            {
                uint32_t bT = timeset_bucket(&ts, T);
                uint64_t cnt = 0;
                for (uint32_t q = 0; q < bT; q++) {
                    cnt += ts.b[q].n;
                }
                cnt += bucket_upper(&ts.b[bT], T);
                kT = n - 1 - cnt;
            }
        }
        /* regraft lineage c: alive at T in the main tree */
        int32_t c = NIL;
        if (kT <= 1 || T >= f.tm[root]) {
            c = root;
        } else if (kT <= DFS_LIMIT) {
            /* enumerate the lineages crossing T from the root down */
            uint64_t sp = 0, found = 0;
            const uint64_t want = rng_below(&rng, kT);
            stack[sp++] = root;
            int32_t lastc = NIL;
            while (sp > 0 && c == NIL) {
                const int32_t v = stack[--sp];
                if (f.tm[v] <= T) {
                    lastc = v;
                    if (found == want) {
                        c = v;
                    }
                    found++;
                } else {
                    stack[sp++] = f.ch[v][0];
                    stack[sp++] = f.ch[v][1];
                    if (sp > 2 * DFS_LIMIT + 32) {
                        break; /* cannot happen: at most kT open lineages */
                    }
                }
            }
            if (c == NIL) {
                c = lastc != NIL ? lastc : root;
            }
        } else {
            for (int tries = 0; tries < 1000 && c == NIL; tries++) {
                int32_t v = (int32_t) rng_below(&rng, n);
                while (f.par[v] >= 0 && f.tm[f.par[v]] <= T) {
                    v = f.par[v];
                }
                if (v != u && f.par[v] >= 0) {
                    c = v;
                } else if (v != u && v == root) {
                    c = v;
                }
            }
            if (c == NIL) {
                c = root;
            }
        }
        if (T <= f.tm[c]) {
            T = nextafter(f.tm[c], INFINITY);
        }
        if (T <= t) {
            T = nextafter(t, INFINITY);
        }
        /* regraft: the freed slot p becomes the new node q above c */
        const int32_t q = p;
        const int32_t pc = f.par[c];
        if (pc >= 0 && T >= f.tm[pc]) {
            T = 0.5 * (fmax(f.tm[c], t) + f.tm[pc]);
        }
        CLOSE_EDGE(c, bp);
        f.tm[q] = T;
        f.outid[q] = (uint32_t) tb.n;
        if (timebuf_push(&tb, T) != 0 || timeset_insert(&ts, T) != 0) {
            goto done;
        }
        f.par[q] = pc;
        replace_child(&f, pc, c, q);
        if (pc < 0) {
            root = q;
        }
        f.eleft[q] = bp;
        f.par[c] = q;
        f.eleft[c] = bp;
        f.par[u] = q;
        f.eleft[u] = bp;
        f.ch[q][0] = c;
        f.ch[q][1] = u;
        fenwick_set(&fw, (uint64_t) c, T - f.tm[c]);
        fenwick_set(&fw, (uint64_t) u, T - f.tm[u]);
        fenwick_set(&fw, (uint64_t) q, pc >= 0 ? f.tm[pc] - T : 0.0);
    }
    for (uint64_t v = 0; v < slots; v++) {
        CLOSE_EDGE(v, L);
    }
#undef CLOSE_EDGE

    /* ---- finalise: drop unused nodes, number by time, sort edges */
    {
        const uint64_t NO = tb.n, E = eb.n;
        used = calloc(NO, 1);
        order = malloc(NO * sizeof(timed_t));
        newid = malloc(NO * sizeof(uint32_t));
        idx = malloc((E + 1) * sizeof(uint32_t));
        key = malloc((E + 1) * sizeof(uint64_t));
        if (!used || !order || !newid || !idx || !key || NO > 0x7fffffffull || E > 0x7fffffffull) {
            goto done;
        }
        for (uint64_t i = 0; i < n; i++) {
            used[i] = 1;
        }
        for (uint64_t e = 0; e < E; e++) {
            used[eb.parent[e]] = 1;
            used[eb.child[e]] = 1;
        }
        uint64_t m = 0;
        for (uint64_t i = n; i < NO; i++) {
            if (used[i]) {
                order[m].t = tb.t[i];
                order[m].id = (uint32_t) i;
                m++;
            }
        }
        qsort(order, m, sizeof(timed_t), cmp_timed);
        const uint64_t N = n + m;
        out->node_time = malloc(N * sizeof(double));
        out->edge_left = malloc((E + 1) * sizeof(double));
        out->edge_right = malloc((E + 1) * sizeof(double));
        out->edge_parent = malloc((E + 1) * sizeof(int32_t));
        out->edge_child = malloc((E + 1) * sizeof(int32_t));
        if (!out->node_time || !out->edge_left || !out->edge_right || !out->edge_parent || !out->edge_child) {
            goto done;
        }
        for (uint64_t i = 0; i < n; i++) {
            newid[i] = (uint32_t) i;
            out->node_time[i] = 0.0;
        }
        for (uint64_t i = 0; i < m; i++) {
            newid[order[i].id] = (uint32_t) (n + i);
            out->node_time[n + i] = order[i].t;
        }
        int bits = 1;
        while ((1ull << bits) < N) {
            bits++;
        }
        for (uint64_t e = 0; e < E; e++) {
            key[e] = ((uint64_t) newid[eb.parent[e]] << bits) | newid[eb.child[e]];
            idx[e] = (uint32_t) e;
        }
        /* edges were emitted in order of increasing right end, hence increasing left
         * within one (parent, child) pair: a stable sort keeps that order */
        if (radix_sort_index(E, key, idx, 2 * bits) != 0) {
            goto done;
        }
        for (uint64_t i = 0; i < E; i++) {
            const uint32_t e = idx[i];
            out->edge_left[i] = eb.left[e];
            out->edge_right[i] = eb.right[e];
            out->edge_parent[i] = (int32_t) newid[eb.parent[e]];
            out->edge_child[i] = (int32_t) newid[eb.child[e]];
        }
        out->num_nodes = N;
        out->num_samples = n;
        out->num_edges = E;
        out->num_trees = num_trees;
    }
    ret = 0;
done:
    (void) cmp_double;
    (void) sorted;
    free(f.par);
    free(f.ch);
    free(f.tm);
    free(f.eleft);
    free(f.outid);
    free(fw.tree);
    free(fw.val);
    if (ts.b != NULL) {
        for (uint32_t b = 0; b < ts.nb; b++) {
            free(ts.b[b].t);
        }
    }
    free(ts.b);
    free(ts.lo);
    free(eb.left);
    free(eb.right);
    free(eb.parent);
    free(eb.child);
    free(tb.t);
    free(stack);
    free(live);
    free(order);
    free(newid);
    free(idx);
    free(key);
    free(used);
    if (ret != 0) {
        treegen_free(out);
    }
    return ret;
}
