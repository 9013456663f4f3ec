/* TEST INFRASTRUCTURE (oracle).  Not part of the product path.
 *
 * CPU restatement of the reference mutation overlay, written for this project
 * in plain C99 on flat column arrays (no tskit, no AVL tree, no linked lists).
 * It follows the reference's *draw order* exactly so that, given the same
 * generator (oracle/gsl_shim: MT19937 or a replay stream), it produces the
 * same site and mutation tables bit for bit as /root/reference/lib/mutgen.c.
 * tests/test_oracle.py pins that equality against oracle/_ref (the compiled
 * reference) and against the committed fixtures in tests/golden/.
 *
 * Reference map (file:line in /root/reference):
 *   place_mutations()  <- lib/mutgen.c:1076-1158  (edge x rate-interval loop)
 *   flat()             <- lib/util.c:717-735      (msp_gsl_ran_flat)
 *   map_index()        <- lib/rate_map.c:137-146  (rate_map_get_index)
 *   check_rate_map()   <- lib/rate_map.c:45-100, lib/mutgen.c:796-813
 *   load_kept()        <- lib/mutgen.c:974-1025   (mutgen_initialise_sites)
 *   record order       <- lib/mutgen.c:36-49      (cmp_mutationp)
 *   resolve_site()     <- lib/mutgen.c:1160-1224  (mutgen_choose_alleles)
 *   sweep in run()     <- lib/mutgen.c:1226-1302  (mutgen_apply_mutations)
 *   select_prob()      <- lib/util.c:400-424      (probability_list_select)
 *   matrix checks      <- lib/mutgen.c:148-160,501-524
 *   infinite alleles   <- lib/mutgen.c:431-481
 *   emit()             <- lib/mutgen.c:1027-1074  (mutgen_populate_tables)
 *   basic integrity    <- tskit tables.c:10397-10470 (edge checks, options 0)
 */
#include <math.h>
#include <stdbool.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <gsl/gsl_rng.h>
#include <gsl/gsl_randist.h>
#include <gsl/gsl_math.h>

#include "oracle_api.h"

/* msprime error numbering (lib/util.h:48-136); tskit codes have bit 13 flipped
 * (lib/util.c:337-342). */
#define E_NO_MEMORY (-2)
#define E_BAD_PARAM_VALUE (-4)
#define E_DUPLICATE_SITE_POSITION (-21)
#define E_BAD_RATE_VALUE (-43)
#define E_INCOMPATIBLE_MUTATION_MAP_LENGTH (-44)
#define E_INSUFFICIENT_INTERVALS (-45)
#define E_INTERVAL_MAP_START_NON_ZERO (-46)
#define E_NONFINITE_INTERVAL_POSITION (-47)
#define E_INTERVAL_POSITIONS_UNSORTED (-48)
#define E_UNKNOWN_ALLELE (-51)
#define E_INSUFFICIENT_ALLELES (-52)
#define E_BAD_ROOT_PROBABILITIES (-53)
#define E_BAD_SLIM_PARAMETERS (-56)
#define E_MUTATION_ID_OVERFLOW (-57)
#define E_BAD_TRANSITION_MATRIX (-54)
#define E_UNKNOWN_TIME_NOT_SUPPORTED (-69)
#define TSK_E(code) ((code) ^ (1 << 13))
#define TSK_NODE_OUT_OF_BOUNDS (-202)
#define TSK_SITE_OUT_OF_BOUNDS (-205)
#define TSK_TIME_NONFINITE (-210)
#define TSK_GENOME_COORDS_NONFINITE (-211)
#define TSK_NULL_PARENT (-300)
#define TSK_NULL_CHILD (-301)
#define TSK_BAD_NODE_TIME_ORDERING (-306)
#define TSK_BAD_EDGE_INTERVAL (-307)
#define TSK_RIGHT_GREATER_SEQ_LENGTH (-309)
#define TSK_LEFT_LESS_ZERO (-310)
#define TSK_BAD_SITE_POSITION (-402)
#define TSK_MUTATION_TIME_YOUNGER_THAN_NODE (-506)
#define TSK_BAD_SEQUENCE_LENGTH (-701)

const char *
orc_name(void)
{
    return "mutgen_port.c (project restatement) + gsl shim";
}

/* One mutation, kept or new. */
typedef struct {
    double position;
    double time;
    int32_t node;
    int32_t kept_id; /* row in the input table, or -1 for a new mutation */
    uint64_t seq; /* creation order, tie-break only */
    int64_t parent; /* index into the sorted record array, or -1 */
    int32_t out_id;
    /* derived state: pointer + length (aliases input, model or arena storage) */
    const char *state;
    uint64_t state_len;
    const char *meta;
    uint64_t meta_len;
} rec_t;

/* One site, kept or new. */
typedef struct {
    double position;
    int32_t kept_id; /* row in the input site table, or -1 */
    const char *anc;
    uint64_t anc_len;
    const char *meta;
    uint64_t meta_len;
    uint64_t first, count; /* range in the sorted record array */
} srec_t;

typedef struct {
    const orc_input_t *in;
    gsl_rng *rng;
    rec_t *recs;
    uint64_t num_recs, cap_recs;
    /* open-addressing set of occupied positions (continuous genome only) */
    uint64_t *pos_set;
    uint64_t pos_cap, pos_count;
    /* arena for infinite-alleles strings */
    char *arena;
    uint64_t arena_len, arena_cap;
    uint64_t next_allele;
    uint64_t seq;
    /* SLiM model: every new state and metadata string is its own allocation */
    char **slim_strings;
    uint64_t num_slim_strings, cap_slim_strings;
} port_t;

static double
flat(gsl_rng *rng, double lo, double hi)
{
    double x;
    if (lo == hi) {
        return lo; /* no draw */
    }
    do {
        x = gsl_ran_flat(rng, lo, hi);
    } while (x >= hi);
    return x;
}

static size_t
select_prob(double u, size_t n, const double *probs)
{
    size_t i = 0;
    double sum = 0.0;
    if (n == 0) {
        return 0;
    }
    while (i < n - 1) {
        sum += probs[i];
        if (u < sum) {
            break;
        }
        i++;
    }
    return i;
}

static bool
almost_equal(double a, double b, double eps)
{
    return (fabs(a) < eps && fabs(b) < eps) || gsl_fcmp(a, b, eps) == 0;
}

static bool
valid_probs(size_t n, const double *p)
{
    double sum = 0, lo = 0;
    size_t j;
    for (j = 0; j < n; j++) {
        sum += p[j];
        lo = p[j] < lo ? p[j] : lo;
    }
    return almost_equal(sum, 1.0, 1e-12) && lo >= 0.0;
}

static int
check_model(const orc_input_t *in)
{
    size_t i;
    if (in->model_kind == ORC_MODEL_SLIM) {
        /* lib/mutgen.c:326-343 */
        return ((int64_t) in->start_allele < 0 || in->slim_mutation_type_id < 0) ? E_BAD_SLIM_PARAMETERS : 0;
    }
    if (in->model_kind != ORC_MODEL_MATRIX) {
        return 0;
    }
    if (in->num_alleles < 2) {
        return E_INSUFFICIENT_ALLELES;
    }
    if (!valid_probs(in->num_alleles, in->root_distribution)) {
        return E_BAD_ROOT_PROBABILITIES;
    }
    for (i = 0; i < in->num_alleles; i++) {
        if (!valid_probs(in->num_alleles, in->transition_matrix + i * in->num_alleles)) {
            return E_BAD_TRANSITION_MATRIX;
        }
    }
    return 0;
}

static int
check_rate_map(const orc_input_t *in)
{
    uint64_t j;
    if (in->map_size < 1) {
        return E_INSUFFICIENT_INTERVALS;
    }
    if (in->map_position[0] != 0.0) {
        return E_INTERVAL_MAP_START_NON_ZERO;
    }
    for (j = 0; j <= in->map_size; j++) {
        if (!isfinite(in->map_position[j])) {
            return E_NONFINITE_INTERVAL_POSITION;
        }
        if (j < in->map_size) {
            if (in->map_position[j] >= in->map_position[j + 1]) {
                return E_INTERVAL_POSITIONS_UNSORTED;
            }
            if (!isfinite(in->map_rate[j]) || in->map_rate[j] < 0) {
                return E_BAD_RATE_VALUE;
            }
        }
    }
    if (in->map_position[in->map_size] != in->sequence_length) {
        return E_INCOMPATIBLE_MUTATION_MAP_LENGTH;
    }
    return 0;
}

/* largest i with position[i] <= x */
static uint64_t
map_index(const orc_input_t *in, double x)
{
    uint64_t lo = 0, hi = in->map_size + 1;
    while (lo < hi) { /* lower bound: first i with position[i] >= x */
        uint64_t mid = lo + (hi - lo) / 2;
        if (in->map_position[mid] < x) {
            lo = mid + 1;
        } else {
            hi = mid;
        }
    }
    if (lo > in->map_size) {
        lo = in->map_size;
    }
    if (in->map_position[lo] > x) {
        lo--;
    }
    return lo;
}

static int
check_tables(const orc_input_t *in)
{
    uint64_t j;
    const double L = in->sequence_length;
    if (!isfinite(L) || L <= 0) {
        return TSK_E(TSK_BAD_SEQUENCE_LENGTH);
    }
    for (j = 0; j < in->num_nodes; j++) {
        if (!isfinite(in->node_time[j])) {
            return TSK_E(TSK_TIME_NONFINITE);
        }
    }
    for (j = 0; j < in->num_edges; j++) {
        int32_t p = in->edge_parent[j], c = in->edge_child[j];
        double l = in->edge_left[j], r = in->edge_right[j];
        if (p == -1) {
            return TSK_E(TSK_NULL_PARENT);
        }
        if (p < 0 || (uint64_t) p >= in->num_nodes) {
            return TSK_E(TSK_NODE_OUT_OF_BOUNDS);
        }
        if (c == -1) {
            return TSK_E(TSK_NULL_CHILD);
        }
        if (c < 0 || (uint64_t) c >= in->num_nodes) {
            return TSK_E(TSK_NODE_OUT_OF_BOUNDS);
        }
        if (!(isfinite(l) && isfinite(r))) {
            return TSK_E(TSK_GENOME_COORDS_NONFINITE);
        }
        if (l < 0) {
            return TSK_E(TSK_LEFT_LESS_ZERO);
        }
        if (r > L) {
            return TSK_E(TSK_RIGHT_GREATER_SEQ_LENGTH);
        }
        if (l >= r) {
            return TSK_E(TSK_BAD_EDGE_INTERVAL);
        }
        if (in->node_time[c] >= in->node_time[p]) {
            return TSK_E(TSK_BAD_NODE_TIME_ORDERING);
        }
    }
    for (j = 0; j < in->num_sites; j++) {
        double x = in->site_position[j];
        if (!isfinite(x) || x < 0 || x >= L) {
            return TSK_E(TSK_BAD_SITE_POSITION);
        }
    }
    for (j = 0; j < in->num_mutations; j++) {
        int32_t s = in->mutation_site[j], u = in->mutation_node[j];
        if (s < 0 || (uint64_t) s >= in->num_sites) {
            return TSK_E(TSK_SITE_OUT_OF_BOUNDS);
        }
        if (u < 0 || (uint64_t) u >= in->num_nodes) {
            return TSK_E(TSK_NODE_OUT_OF_BOUNDS);
        }
        if (in->mutation_time != NULL) {
            double t = in->mutation_time[j];
            uint64_t bits;
            memcpy(&bits, &t, 8);
            if (bits != 0x7FF874736B697421ULL) { /* TSK_UNKNOWN_TIME */
                if (!isfinite(t)) {
                    return TSK_E(TSK_TIME_NONFINITE);
                }
                if (t < in->node_time[u]) {
                    return TSK_E(TSK_MUTATION_TIME_YOUNGER_THAN_NODE);
                }
            }
        }
    }
    return 0;
}

/* ---------- occupied-position set (stands in for the reference's AVL map) ---------- */

static uint64_t
dbl_bits(double x)
{
    uint64_t b;
    if (x == 0.0) {
        x = 0.0; /* fold -0.0 */
    }
    memcpy(&b, &x, 8);
    return b;
}

static uint64_t
mix64(uint64_t z)
{
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
    return z ^ (z >> 31);
}

#define POS_EMPTY 0xFFFFFFFFFFFFFFFFULL /* a NaN pattern: never a valid position */

static int
pos_set_grow(port_t *self, uint64_t cap)
{
    uint64_t *old = self->pos_set;
    uint64_t old_cap = self->pos_cap, j;
    self->pos_set = malloc(cap * sizeof(uint64_t));
    if (self->pos_set == NULL) {
        self->pos_set = old;
        return E_NO_MEMORY;
    }
    memset(self->pos_set, 0xff, cap * sizeof(uint64_t));
    self->pos_cap = cap;
    for (j = 0; j < old_cap; j++) {
        if (old[j] != POS_EMPTY) {
            uint64_t h = mix64(old[j]) & (cap - 1);
            while (self->pos_set[h] != POS_EMPTY) {
                h = (h + 1) & (cap - 1);
            }
            self->pos_set[h] = old[j];
        }
    }
    free(old);
    return 0;
}

static bool
pos_set_contains(port_t *self, double x)
{
    uint64_t key = dbl_bits(x), h;
    if (self->pos_cap == 0) {
        return false;
    }
    h = mix64(key) & (self->pos_cap - 1);
    while (self->pos_set[h] != POS_EMPTY) {
        if (self->pos_set[h] == key) {
            return true;
        }
        h = (h + 1) & (self->pos_cap - 1);
    }
    return false;
}

/* returns 1 if inserted, 0 if already present, < 0 on error */
static int
pos_set_insert(port_t *self, double x)
{
    uint64_t key = dbl_bits(x), h;
    if ((self->pos_count + 1) * 2 > self->pos_cap) {
        int ret = pos_set_grow(self, self->pos_cap == 0 ? 1024 : self->pos_cap * 2);
        if (ret != 0) {
            return ret;
        }
    }
    h = mix64(key) & (self->pos_cap - 1);
    while (self->pos_set[h] != POS_EMPTY) {
        if (self->pos_set[h] == key) {
            return 0;
        }
        h = (h + 1) & (self->pos_cap - 1);
    }
    self->pos_set[h] = key;
    self->pos_count++;
    return 1;
}

static rec_t *
push_rec(port_t *self)
{
    rec_t *r;
    if (self->num_recs == self->cap_recs) {
        uint64_t cap = self->cap_recs == 0 ? 1024 : self->cap_recs * 2;
        rec_t *p = realloc(self->recs, cap * sizeof(rec_t));
        if (p == NULL) {
            return NULL;
        }
        self->recs = p;
        self->cap_recs = cap;
    }
    r = &self->recs[self->num_recs++];
    memset(r, 0, sizeof(*r));
    r->kept_id = -1;
    r->parent = -1;
    r->seq = self->seq++;
    return r;
}

/* Existing sites and mutations (keep=True). */
static int
load_kept(port_t *self, srec_t **sites_out, uint64_t *num_sites_out)
{
    const orc_input_t *in = self->in;
    uint64_t s, j = 0;
    srec_t *sites = calloc(in->num_sites > 0 ? in->num_sites : 1, sizeof(srec_t));
    if (sites == NULL) {
        return E_NO_MEMORY;
    }
    *sites_out = sites;
    *num_sites_out = in->num_sites;
    for (s = 0; s < in->num_sites; s++) {
        int ins = pos_set_insert(self, in->site_position[s]);
        if (ins < 0) {
            return ins;
        }
        if (ins == 0) {
            return E_DUPLICATE_SITE_POSITION;
        }
        sites[s].position = in->site_position[s];
        sites[s].kept_id = (int32_t) s;
        sites[s].anc = in->site_ancestral_state + in->site_ancestral_state_offset[s];
        sites[s].anc_len
            = in->site_ancestral_state_offset[s + 1] - in->site_ancestral_state_offset[s];
        sites[s].meta = in->site_metadata + in->site_metadata_offset[s];
        sites[s].meta_len = in->site_metadata_offset[s + 1] - in->site_metadata_offset[s];
        while (j < in->num_mutations && in->mutation_site[j] == (int32_t) s) {
            rec_t *r;
            double t = in->mutation_time == NULL ? NAN : in->mutation_time[j];
            if (isnan(t)) {
                return E_UNKNOWN_TIME_NOT_SUPPORTED;
            }
            r = push_rec(self);
            if (r == NULL) {
                return E_NO_MEMORY;
            }
            r->position = in->site_position[s];
            r->time = t;
            r->node = in->mutation_node[j];
            r->kept_id = (int32_t) j;
            r->state = in->mutation_derived_state + in->mutation_derived_state_offset[j];
            r->state_len
                = in->mutation_derived_state_offset[j + 1] - in->mutation_derived_state_offset[j];
            r->meta = in->mutation_metadata + in->mutation_metadata_offset[j];
            r->meta_len = in->mutation_metadata_offset[j + 1] - in->mutation_metadata_offset[j];
            j++;
        }
    }
    return 0;
}

/* Loop A: edges in table order, rate intervals left to right; per (edge, interval)
 * one Poisson draw, then position and time for each mutation. */
static int
place_mutations(port_t *self)
{
    const orc_input_t *in = self->in;
    const bool discrete = in->discrete_genome != 0;
    uint64_t j, k, n, idx;

    for (j = 0; j < in->num_edges; j++) {
        double left = in->edge_left[j];
        const double edge_right = in->edge_right[j];
        const int32_t child = in->edge_child[j];
        const double t_child = in->node_time[child];
        const double t_parent = in->node_time[in->edge_parent[j]];
        const double branch_start = in->start_time > t_child ? in->start_time : t_child;
        const double branch_end = in->end_time < t_parent ? in->end_time : t_parent;
        const double branch_length = branch_end - branch_start;
        double right = 0;

        idx = map_index(in, left);
        while (right != edge_right) {
            double site_left, site_right, mu;
            right = edge_right < in->map_position[idx + 1] ? edge_right
                                                           : in->map_position[idx + 1];
            site_left = discrete ? ceil(left) : left;
            site_right = discrete ? ceil(right) : right;
            mu = branch_length * (site_right - site_left) * in->map_rate[idx];
            n = gsl_ran_poisson(self->rng, mu);
            for (k = 0; k < n; k++) {
                double position, time;
                rec_t *r;
                bool taken;
                do {
                    position = flat(self->rng, site_left, site_right);
                    if (discrete) {
                        position = floor(position);
                    }
                    taken = !discrete && pos_set_contains(self, position);
                } while (taken);
                time = flat(self->rng, branch_start, branch_end);
                if (!discrete) {
                    int ins = pos_set_insert(self, position);
                    if (ins < 0) {
                        return ins;
                    }
                }
                r = push_rec(self);
                if (r == NULL) {
                    return E_NO_MEMORY;
                }
                r->position = position;
                r->time = time;
                r->node = child;
            }
            left = right;
            idx++;
        }
    }
    return 0;
}

/* Site order, then the reference's within-site order: kept vs kept by row id,
 * anything involving a new mutation by time descending.  Ties that the reference
 * leaves to qsort (measure zero) fall back to: kept first, then creation order. */
static int
cmp_rec(const void *a, const void *b)
{
    const rec_t *x = a, *y = b;
    if (x->position != y->position) {
        return (x->position > y->position) - (x->position < y->position);
    }
    if (x->kept_id >= 0 && y->kept_id >= 0) {
        return (x->kept_id > y->kept_id) - (x->kept_id < y->kept_id);
    }
    if (x->time != y->time) {
        return (y->time > x->time) - (y->time < x->time);
    }
    if ((x->kept_id >= 0) != (y->kept_id >= 0)) {
        return x->kept_id >= 0 ? -1 : 1;
    }
    return (x->seq > y->seq) - (x->seq < y->seq);
}

static int
cmp_srec(const void *a, const void *b)
{
    const srec_t *x = a, *y = b;
    return (x->position > y->position) - (x->position < y->position);
}

static int
make_allele(port_t *self, const char **dest, uint64_t *len)
{
    char tmp[32];
    int n = snprintf(tmp, sizeof(tmp), "%llu", (unsigned long long) self->next_allele);
    if (self->arena_len + (uint64_t) n > self->arena_cap) {
        /* the arena must not move: states alias it.  Chain fixed blocks instead. */
        return E_NO_MEMORY;
    }
    memcpy(self->arena + self->arena_len, tmp, (size_t) n);
    *dest = self->arena + self->arena_len;
    *len = (uint64_t) n;
    self->arena_len += (uint64_t) n;
    self->next_allele++;
    return 0;
}

/* SLiM model (lib/mutgen.c:257-413): the derived state is the parent's state, a comma if
 * that is not empty, and the next mutation id in decimal; the metadata is the parent's
 * metadata plus one 17-byte MutationMetadataRec (type id, selection coefficient 0, subpopulation
 * -1, origin generation = slim_generation - (int32) time, nucleotide -1). */
#define SLIM_MD_SIZE 17

static char *
slim_alloc(port_t *self, uint64_t n)
{
    char *b;
    if (self->num_slim_strings == self->cap_slim_strings) {
        uint64_t cap = self->cap_slim_strings == 0 ? 1024 : 2 * self->cap_slim_strings;
        char **g = realloc(self->slim_strings, cap * sizeof(*g));
        if (g == NULL) {
            return NULL;
        }
        self->slim_strings = g;
        self->cap_slim_strings = cap;
    }
    b = malloc(n > 0 ? n : 1);
    if (b != NULL) {
        self->slim_strings[self->num_slim_strings++] = b;
    }
    return b;
}

static int
slim_transition(port_t *self, const char *pa, uint64_t palen, const char *pm, uint64_t pmlen, rec_t *mut)
{
    const orc_input_t *in = self->in;
    char digits[32], *buff;
    int n = snprintf(digits, sizeof(digits), "%lld", (long long) (int64_t) self->next_allele);
    uint64_t len = palen + (palen > 0 ? 1 : 0) + (uint64_t) n;
    int32_t type_id = in->slim_mutation_type_id, subpop = -1;
    int32_t origin = in->slim_generation - (int32_t) mut->time;
    float sel = 0;
    int8_t nuc = -1;

    buff = slim_alloc(self, len);
    if (buff == NULL) {
        return E_NO_MEMORY;
    }
    if (palen > 0) {
        memcpy(buff, pa, palen);
        buff[palen] = ',';
    }
    memcpy(buff + len - (uint64_t) n, digits, (size_t) n);
    if ((int64_t) self->next_allele == INT64_MAX) {
        return E_MUTATION_ID_OVERFLOW;
    }
    self->next_allele++;
    mut->state = buff;
    mut->state_len = len;
    buff = slim_alloc(self, pmlen + SLIM_MD_SIZE);
    if (buff == NULL) {
        return E_NO_MEMORY;
    }
    if (pmlen > 0) {
        memcpy(buff, pm, pmlen);
    }
    memcpy(buff + pmlen, &type_id, 4);
    memcpy(buff + pmlen + 4, &sel, 4);
    memcpy(buff + pmlen + 8, &subpop, 4);
    memcpy(buff + pmlen + 12, &origin, 4);
    memcpy(buff + pmlen + 16, &nuc, 1);
    mut->meta = buff;
    mut->meta_len = pmlen + SLIM_MD_SIZE;
    return 0;
}

static int
allele_index(const orc_input_t *in, const char *a, uint64_t len)
{
    uint64_t j;
    for (j = 0; j < in->num_alleles; j++) {
        uint64_t alen = in->allele_offset[j + 1] - in->allele_offset[j];
        if (alen == len && memcmp(a, in->alleles + in->allele_offset[j], len) == 0) {
            return (int) j;
        }
    }
    return -1;
}

/* Per-site step of loop B: root state for a new site, then for every mutation in
 * order the nearest ancestor-or-self node that already carries a mutation at this
 * site gives the parent; new mutations draw their derived state. */
static int
resolve_site(port_t *self, srec_t *site, const int32_t *parent, int64_t *bottom)
{
    const orc_input_t *in = self->in;
    rec_t *recs = self->recs + site->first;
    uint64_t k;
    int ret = 0;

    if (site->kept_id < 0) {
        if (in->model_kind == ORC_MODEL_MATRIX) {
            double u = flat(self->rng, 0.0, 1.0);
            size_t j = select_prob(u, in->num_alleles, in->root_distribution);
            site->anc = in->alleles + in->allele_offset[j];
            site->anc_len = in->allele_offset[j + 1] - in->allele_offset[j];
        } else if (in->model_kind == ORC_MODEL_SLIM) {
            site->anc = NULL; /* lib/mutgen.c:346-353: the empty string, no draw */
            site->anc_len = 0;
        } else {
            ret = make_allele(self, &site->anc, &site->anc_len);
            if (ret != 0) {
                return ret;
            }
        }
    }
    for (k = 0; k < site->count; k++) {
        rec_t *mut = &recs[k];
        const char *pa, *pmd;
        uint64_t palen, pmdlen;
        int32_t u = mut->node;
        while (u != -1 && bottom[u] < 0) {
            u = parent[u];
        }
        if (u == -1) {
            pa = site->anc;
            palen = site->anc_len;
            pmd = site->meta; /* lib/mutgen.c:1190-1194: the SITE's metadata stands in */
            pmdlen = site->meta_len;
        } else {
            rec_t *pm = &self->recs[bottom[u]];
            mut->parent = bottom[u];
            pa = pm->state;
            palen = pm->state_len;
            pmd = pm->meta;
            pmdlen = pm->meta_len;
        }
        if (mut->kept_id < 0) {
            if (in->model_kind == ORC_MODEL_MATRIX) {
                double v = flat(self->rng, 0.0, 1.0);
                int pi = allele_index(in, pa, palen);
                size_t j;
                if (pi < 0) {
                    ret = E_UNKNOWN_ALLELE;
                    break;
                }
                j = select_prob(v, in->num_alleles,
                    in->transition_matrix + (uint64_t) pi * in->num_alleles);
                mut->state = in->alleles + in->allele_offset[j];
                mut->state_len = in->allele_offset[j + 1] - in->allele_offset[j];
            } else if (in->model_kind == ORC_MODEL_SLIM) {
                ret = slim_transition(self, pa, palen, pmd, pmdlen, mut);
                if (ret != 0) {
                    break;
                }
            } else {
                ret = make_allele(self, &mut->state, &mut->state_len);
                if (ret != 0) {
                    break;
                }
            }
        }
        bottom[mut->node] = (int64_t) (site->first + k);
    }
    for (k = 0; k < site->count; k++) {
        bottom[recs[k].node] = -1;
    }
    return ret;
}

typedef struct {
    double x;
    uint64_t e;
} ev_t;

static int
cmp_ev(const void *a, const void *b)
{
    const ev_t *x = a, *y = b;
    if (x->x != y->x) {
        return (x->x > y->x) - (x->x < y->x);
    }
    return (x->e > y->e) - (x->e < y->e);
}

static void *
col(uint64_t n, size_t sz)
{
    return calloc(n > 0 ? n : 1, sz);
}

/* Dense ids in position order; new sites without mutations cannot occur here
 * (a new site is only created by a mutation) and kept sites always survive. */
static int
emit(port_t *self, srec_t *sites, uint64_t num_sites, orc_output_t *out)
{
    uint64_t s, k, M = self->num_recs, anc_bytes = 0, smd_bytes = 0, der_bytes = 0,
                   mmd_bytes = 0, row = 0;
    for (s = 0; s < num_sites; s++) {
        anc_bytes += sites[s].anc_len;
        smd_bytes += sites[s].meta_len;
    }
    for (k = 0; k < M; k++) {
        der_bytes += self->recs[k].state_len;
        mmd_bytes += self->recs[k].meta_len;
    }
    out->num_sites = num_sites;
    out->num_mutations = M;
    out->site_position = col(num_sites, 8);
    out->site_ancestral_state = col(anc_bytes, 1);
    out->site_ancestral_state_offset = col(num_sites + 1, 8);
    out->site_metadata = col(smd_bytes, 1);
    out->site_metadata_offset = col(num_sites + 1, 8);
    out->mutation_site = col(M, 4);
    out->mutation_node = col(M, 4);
    out->mutation_parent = col(M, 4);
    out->mutation_time = col(M, 8);
    out->mutation_derived_state = col(der_bytes, 1);
    out->mutation_derived_state_offset = col(M + 1, 8);
    out->mutation_metadata = col(mmd_bytes, 1);
    out->mutation_metadata_offset = col(M + 1, 8);
    if (!out->site_position || !out->site_ancestral_state || !out->site_ancestral_state_offset
        || !out->site_metadata || !out->site_metadata_offset || !out->mutation_site
        || !out->mutation_node || !out->mutation_parent || !out->mutation_time
        || !out->mutation_derived_state || !out->mutation_derived_state_offset
        || !out->mutation_metadata || !out->mutation_metadata_offset) {
        return E_NO_MEMORY;
    }
    anc_bytes = smd_bytes = der_bytes = mmd_bytes = 0;
    for (s = 0; s < num_sites; s++) {
        srec_t *site = &sites[s];
        out->site_position[s] = site->position;
        memcpy(out->site_ancestral_state + anc_bytes, site->anc, site->anc_len);
        anc_bytes += site->anc_len;
        out->site_ancestral_state_offset[s + 1] = anc_bytes;
        if (site->meta_len > 0) {
            memcpy(out->site_metadata + smd_bytes, site->meta, site->meta_len);
        }
        smd_bytes += site->meta_len;
        out->site_metadata_offset[s + 1] = smd_bytes;
        for (k = 0; k < site->count; k++) {
            rec_t *m = &self->recs[site->first + k];
            m->out_id = (int32_t) row;
            out->mutation_site[row] = (int32_t) s;
            out->mutation_node[row] = m->node;
            out->mutation_parent[row] = m->parent < 0 ? -1 : self->recs[m->parent].out_id;
            out->mutation_time[row] = m->time;
            if (m->state_len > 0) {
                memcpy(out->mutation_derived_state + der_bytes, m->state, m->state_len);
            }
            der_bytes += m->state_len;
            out->mutation_derived_state_offset[row + 1] = der_bytes;
            if (m->meta_len > 0) {
                memcpy(out->mutation_metadata + mmd_bytes, m->meta, m->meta_len);
            }
            mmd_bytes += m->meta_len;
            out->mutation_metadata_offset[row + 1] = mmd_bytes;
            row++;
        }
    }
    return 0;
}

static double
now_seconds(void)
{
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return (double) t.tv_sec + 1e-9 * (double) t.tv_nsec;
}

static int
generate(port_t *self, orc_output_t *out)
{
    const orc_input_t *in = self->in;
    const uint64_t E = in->num_edges, N = in->num_nodes;
    srec_t *kept_sites = NULL, *sites = NULL;
    uint64_t num_kept_sites = 0, num_sites = 0, cap_sites, k, s, ks;
    ev_t *ins = NULL, *rem = NULL;
    int32_t *parent = NULL;
    int64_t *bottom = NULL;
    uint64_t tj, tk, cur;
    double left, right;
    int ret;

    ret = check_tables(in);
    if (ret != 0) {
        goto out;
    }
    if (in->keep) {
        ret = load_kept(self, &kept_sites, &num_kept_sites);
        if (ret != 0) {
            goto out;
        }
    }
    ret = place_mutations(self);
    if (ret != 0) {
        goto out;
    }
    /* order mutations; group into sites; merge with the kept sites */
    qsort(self->recs, self->num_recs, sizeof(rec_t), cmp_rec);
    if (num_kept_sites > 0) {
        qsort(kept_sites, num_kept_sites, sizeof(srec_t), cmp_srec);
    }
    cap_sites = num_kept_sites + self->num_recs + 1;
    sites = calloc(cap_sites, sizeof(srec_t));
    if (sites == NULL) {
        ret = E_NO_MEMORY;
        goto out;
    }
    k = 0;
    ks = 0;
    while (k < self->num_recs || ks < num_kept_sites) {
        srec_t *site = &sites[num_sites++];
        bool take_kept = ks < num_kept_sites
                         && (k >= self->num_recs
                             || kept_sites[ks].position <= self->recs[k].position);
        if (take_kept) {
            *site = kept_sites[ks++];
        } else {
            memset(site, 0, sizeof(*site));
            site->position = self->recs[k].position;
            site->kept_id = -1;
        }
        site->first = k;
        while (k < self->num_recs && self->recs[k].position == site->position) {
            k++;
        }
        site->count = k - site->first;
    }
    /* storage for infinite-alleles strings: at most one per site plus one per mutation */
    self->arena_cap = 21 * (num_sites + self->num_recs + 1);
    self->arena = malloc(self->arena_cap);
    ins = malloc((E > 0 ? E : 1) * sizeof(ev_t));
    rem = malloc((E > 0 ? E : 1) * sizeof(ev_t));
    parent = malloc((N > 0 ? N : 1) * sizeof(int32_t));
    bottom = malloc((N > 0 ? N : 1) * sizeof(int64_t));
    if (!self->arena || !ins || !rem || !parent || !bottom) {
        ret = E_NO_MEMORY;
        goto out;
    }
    for (k = 0; k < E; k++) {
        ins[k].x = in->edge_left[k];
        ins[k].e = k;
        rem[k].x = in->edge_right[k];
        rem[k].e = k;
    }
    qsort(ins, E, sizeof(ev_t), cmp_ev);
    qsort(rem, E, sizeof(ev_t), cmp_ev);
    for (k = 0; k < N; k++) {
        parent[k] = -1;
        bottom[k] = -1;
    }
    /* loop B: sweep the trees left to right, resolving the sites under each tree */
    tj = tk = 0;
    cur = 0;
    left = 0;
    while (tj < E || left < in->sequence_length) {
        while (tk < E && rem[tk].x == left) {
            parent[in->edge_child[rem[tk].e]] = -1;
            tk++;
        }
        while (tj < E && ins[tj].x == left) {
            parent[in->edge_child[ins[tj].e]] = in->edge_parent[ins[tj].e];
            tj++;
        }
        right = in->sequence_length;
        if (tj < E && ins[tj].x < right) {
            right = ins[tj].x;
        }
        if (tk < E && rem[tk].x < right) {
            right = rem[tk].x;
        }
        while (cur < num_sites && sites[cur].position < right) {
            ret = resolve_site(self, &sites[cur], parent, bottom);
            if (ret != 0) {
                goto out;
            }
            cur++;
        }
        left = right;
    }
    ret = emit(self, sites, num_sites, out);
    s = 0;
    (void) s;
out:
    free(kept_sites);
    free(sites);
    free(ins);
    free(rem);
    free(parent);
    free(bottom);
    return ret;
}

int
orc_sim_mutations(const orc_input_t *in, orc_output_t *out)
{
    port_t self;
    double t0;
    int ret;

    memset(out, 0, sizeof(*out));
    memset(&self, 0, sizeof(self));
    self.in = in;
    self.next_allele = in->start_allele;
    self.rng = gsl_rng_alloc(in->rng_kind == ORC_RNG_REPLAY ? gsl_rng_replay : gsl_rng_mt19937);
    if (self.rng == NULL) {
        ret = E_NO_MEMORY;
        goto out;
    }
    if (in->rng_kind == ORC_RNG_REPLAY) {
        gsl_rng_replay_set(self.rng, in->stream, in->stream_len);
    } else {
        gsl_rng_set(self.rng, in->seed);
    }
    /* same order of checks as the driver around the reference:
     * model, mutgen_alloc, rate map, time interval, then generate */
    ret = check_model(in);
    if (ret != 0) {
        goto out;
    }
    if (in->sequence_length <= 0) {
        ret = E_BAD_PARAM_VALUE;
        goto out;
    }
    ret = check_rate_map(in);
    if (ret != 0) {
        goto out;
    }
    if (in->end_time < in->start_time) {
        ret = E_BAD_PARAM_VALUE;
        goto out;
    }
    t0 = now_seconds();
    ret = generate(&self, out);
    out->seconds = now_seconds() - t0;
    out->next_allele = self.next_allele;
out:
    if (self.rng != NULL) {
        out->draws = self.rng->draws;
        out->stream_exhausted = self.rng->exhausted;
        gsl_rng_free(self.rng);
    }
    free(self.recs);
    free(self.pos_set);
    free(self.arena);
    {
        uint64_t q;
        for (q = 0; q < self.num_slim_strings; q++) {
            free(self.slim_strings[q]);
        }
        free(self.slim_strings);
    }
    if (ret != 0) {
        int32_t ex = out->stream_exhausted;
        uint64_t dr = out->draws;
        orc_free_output(out);
        out->stream_exhausted = ex;
        out->draws = dr;
    }
    out->status = ret;
    return ret;
}

void
orc_free_output(orc_output_t *out)
{
    free(out->site_position);
    free(out->site_ancestral_state);
    free(out->site_ancestral_state_offset);
    free(out->site_metadata);
    free(out->site_metadata_offset);
    free(out->mutation_site);
    free(out->mutation_node);
    free(out->mutation_parent);
    free(out->mutation_time);
    free(out->mutation_derived_state);
    free(out->mutation_derived_state_offset);
    free(out->mutation_metadata);
    free(out->mutation_metadata_offset);
    memset(out, 0, sizeof(*out));
}
