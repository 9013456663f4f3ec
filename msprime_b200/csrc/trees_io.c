/*
 * `.trees` (kastore) reader / writer for the overlay's file hand-off: include/msprime_b200_io.h.
 *
 * Host only, plain C.  The container layout follows kastore's file format
 * (git-submodules/tskit/c/subprojects/kastore/kastore.c:128-337: header fields, descriptor
 * fields, packing rules), the schema checks follow tsk_table_collection_load
 * (git-submodules/tskit/c/tskit/tables.c:120-270, 11588-11720).  Not a port: the reader maps the
 * file instead of reading it item by item, so the columns go to the GPU without an
 * intermediate copy, and the writer streams borrowed arrays with one pass of writev-sized
 * chunks.
 */
#define _GNU_SOURCE
#include <errno.h>
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "../../include/msprime_b200_io.h"

#define HEADER_SIZE 64u
#define DESCRIPTOR_SIZE 64u
#define ARRAY_ALIGN 8u
#define FILE_VERSION_MAJOR 1u
#define FILE_VERSION_MINOR 0u
static const char MAGIC[8] = { '\211', 'K', 'A', 'S', '\r', '\n', '\032', '\n' };

#define TREES_FORMAT_NAME "tskit.trees"
#define TREES_VERSION_MAJOR 12u

static const uint64_t TYPE_SIZE[MSPB_KAS_NUM_TYPES] = { 1, 1, 2, 2, 4, 4, 8, 8, 4, 8 };

typedef struct {
    uint64_t key_start, key_len, array_start, array_len;
    int32_t type;
} entry_t;

struct mspb_store {
    int fd;
    const unsigned char *map; /* NULL for a zero-item store */
    uint64_t size;
    uint64_t num_items;
    entry_t *items;
    /* widened offset columns handed out by mspb_store_tables */
    void *owned[8];
    int num_owned;
};

static uint64_t
rd64(const unsigned char *p)
{
    uint64_t v;
    memcpy(&v, p, 8);
    return v;
}

/* a + b*c with overflow reported as "does not fit" */
static int
span_fits(uint64_t start, uint64_t count, uint64_t width, uint64_t limit)
{
    if (width != 0 && count > UINT64_MAX / width) {
        return 0;
    }
    const uint64_t bytes = count * width;
    return start <= limit && bytes <= limit - start;
}

static int
key_cmp(const char *a, uint64_t alen, const char *b, uint64_t blen)
{
    const uint64_t n = alen < blen ? alen : blen;
    const int c = memcmp(a, b, (size_t) n);
    if (c != 0) {
        return c;
    }
    return (alen > blen) - (alen < blen);
}

void
mspb_store_close(mspb_store_t *s)
{
    if (s == NULL) {
        return;
    }
    for (int j = 0; j < s->num_owned; j++) {
        free(s->owned[j]);
    }
    if (s->map != NULL) {
        munmap((void *) s->map, (size_t) s->size);
    }
    if (s->fd >= 0) {
        close(s->fd);
    }
    free(s->items);
    free(s);
}

int
mspb_store_open(const char *path, mspb_store_t **out)
{
    int ret = MSPB_KAS_ERR_BAD_FILE_FORMAT;
    struct stat st;
    mspb_store_t *s = calloc(1, sizeof(*s));

    *out = NULL;
    if (s == NULL) {
        return MSPB_IO_KAS(MSPB_KAS_ERR_NO_MEMORY);
    }
    s->fd = open(path, O_RDONLY);
    if (s->fd < 0 || fstat(s->fd, &st) != 0) {
        ret = MSPB_KAS_ERR_IO;
        goto out;
    }
    if (st.st_size == 0) {
        ret = MSPB_KAS_ERR_EOF; /* kastore_read_header: nothing to read */
        goto out;
    }
    if ((uint64_t) st.st_size < HEADER_SIZE) {
        goto out;
    }
    s->size = (uint64_t) st.st_size;
    s->map = mmap(NULL, (size_t) s->size, PROT_READ, MAP_PRIVATE, s->fd, 0);
    if (s->map == MAP_FAILED) {
        s->map = NULL;
        ret = MSPB_KAS_ERR_IO;
        goto out;
    }
    /* header: magic[8] version_major u16 version_minor u16 num_items u32 file_size u64 */
    if (memcmp(s->map, MAGIC, 8) != 0) {
        goto out;
    }
    uint16_t major;
    uint32_t n32;
    memcpy(&major, s->map + 8, 2);
    memcpy(&n32, s->map + 12, 4);
    const uint64_t file_size = rd64(s->map + 16);
    if (major < FILE_VERSION_MAJOR) {
        ret = MSPB_KAS_ERR_VERSION_TOO_OLD;
        goto out;
    }
    if (major > FILE_VERSION_MAJOR) {
        ret = MSPB_KAS_ERR_VERSION_TOO_NEW;
        goto out;
    }
    /* a shorter file than its header claims is truncated; trailing bytes are tolerated only
     * as far as kastore does (it reads file_size bytes and never looks further) */
    if (file_size < HEADER_SIZE || file_size > s->size) {
        goto out;
    }
    s->num_items = n32;
    if (s->num_items == 0) {
        ret = file_size == HEADER_SIZE ? 0 : MSPB_KAS_ERR_BAD_FILE_FORMAT;
        goto out;
    }
    if (!span_fits(HEADER_SIZE, s->num_items, DESCRIPTOR_SIZE, file_size)) {
        goto out;
    }
    s->items = calloc((size_t) s->num_items, sizeof(*s->items));
    if (s->items == NULL) {
        ret = MSPB_KAS_ERR_NO_MEMORY;
        goto out;
    }
    /* descriptor: type u8, 7 reserved, key_start, key_len, array_start, array_len (u64 each) */
    uint64_t offset = HEADER_SIZE + s->num_items * DESCRIPTOR_SIZE;
    for (uint64_t j = 0; j < s->num_items; j++) {
        const unsigned char *d = s->map + HEADER_SIZE + j * DESCRIPTOR_SIZE;
        entry_t *e = &s->items[j];
        e->type = d[0];
        e->key_start = rd64(d + 8);
        e->key_len = rd64(d + 16);
        e->array_start = rd64(d + 24);
        e->array_len = rd64(d + 32);
        if (e->type >= MSPB_KAS_NUM_TYPES) {
            ret = MSPB_KAS_ERR_BAD_TYPE;
            goto out;
        }
        if (!span_fits(e->key_start, e->key_len, 1, file_size)
            || !span_fits(e->array_start, e->array_len, TYPE_SIZE[e->type], file_size)) {
            goto out;
        }
        /* keys are packed back to back right after the descriptors */
        if (e->key_start != offset) {
            goto out;
        }
        offset += e->key_len;
    }
    for (uint64_t j = 0; j < s->num_items; j++) {
        const entry_t *e = &s->items[j];
        offset = (offset + ARRAY_ALIGN - 1) / ARRAY_ALIGN * ARRAY_ALIGN;
        if (e->array_start != offset) {
            goto out;
        }
        offset += e->array_len * TYPE_SIZE[e->type];
        if (e->key_len == 0) {
            goto out;
        }
        /* kastore looks keys up by bsearch, so they must be strictly increasing */
        if (j > 0) {
            const entry_t *p = &s->items[j - 1];
            if (key_cmp((const char *) s->map + p->key_start, p->key_len, (const char *) s->map + e->key_start,
                    e->key_len) >= 0) {
                goto out;
            }
        }
    }
    if (offset != file_size) {
        goto out;
    }
    ret = 0;
out:
    if (ret != 0) {
        mspb_store_close(s);
        return MSPB_IO_KAS(ret);
    }
    *out = s;
    return 0;
}

uint64_t
mspb_store_num_items(const mspb_store_t *s)
{
    return s->num_items;
}

static void
fill_item(const mspb_store_t *s, const entry_t *e, mspb_store_item_t *item)
{
    item->key = (const char *) s->map + e->key_start;
    item->key_len = e->key_len;
    item->type = e->type;
    item->array = s->map + e->array_start;
    item->array_len = e->array_len;
}

int
mspb_store_item(const mspb_store_t *s, uint64_t index, mspb_store_item_t *item)
{
    if (index >= s->num_items) {
        return MSPB_IO_KAS(MSPB_KAS_ERR_KEY_NOT_FOUND);
    }
    fill_item(s, &s->items[index], item);
    return 0;
}

int
mspb_store_get(const mspb_store_t *s, const char *key, mspb_store_item_t *item)
{
    const uint64_t klen = strlen(key);
    uint64_t lo = 0, hi = s->num_items;

    if (klen == 0) {
        return MSPB_IO_KAS(MSPB_KAS_ERR_EMPTY_KEY);
    }
    while (lo < hi) {
        const uint64_t mid = lo + (hi - lo) / 2;
        const entry_t *e = &s->items[mid];
        const int c = key_cmp(key, klen, (const char *) s->map + e->key_start, e->key_len);
        if (c == 0) {
            fill_item(s, e, item);
            return 0;
        }
        if (c < 0) {
            hi = mid;
        } else {
            lo = mid + 1;
        }
    }
    return MSPB_IO_KAS(MSPB_KAS_ERR_KEY_NOT_FOUND);
}

/* ---------------------------------------------------------------- tskit.trees schema */

#define NOT_FOUND MSPB_IO_KAS(MSPB_KAS_ERR_KEY_NOT_FOUND)
#define ROWS_UNSET UINT64_MAX

/* read_table_cols (tables.c:100-146): a fixed-width column; all columns of a table agree on
 * the row count */
static int
get_col(const mspb_store_t *s, const char *key, int type, int optional, uint64_t *num_rows, const void **dest)
{
    mspb_store_item_t it;
    int ret = mspb_store_get(s, key, &it);

    *dest = NULL;
    if (ret == NOT_FOUND) {
        return optional ? 0 : MSPB_IO_TSK(MSPB_TSK_ERR_REQUIRED_COL_NOT_FOUND);
    }
    if (ret != 0) {
        return ret;
    }
    if (*num_rows == ROWS_UNSET) {
        *num_rows = it.array_len;
    } else if (*num_rows != it.array_len) {
        return MSPB_IO_TSK(MSPB_TSK_ERR_FILE_FORMAT);
    }
    if (it.type != type) {
        return MSPB_IO_TSK(MSPB_TSK_ERR_BAD_COLUMN_TYPE);
    }
    *dest = it.array;
    return 0;
}

/* read_table_ragged_cols (tables.c:166-262): data + "<name>_offset" (uint32 or uint64) */
static int
get_ragged(mspb_store_t *s, const char *key, uint64_t *num_rows, const char **data, const uint64_t **offset)
{
    char off_key[96];
    mspb_store_item_t d, o;
    int ret;

    *data = NULL;
    *offset = NULL;
    snprintf(off_key, sizeof(off_key), "%s_offset", key);
    ret = mspb_store_get(s, key, &d);
    const int have_data = ret == 0;
    if (ret != 0 && ret != NOT_FOUND) {
        return ret;
    }
    if (have_data && d.type != MSPB_KAS_UINT8) {
        return MSPB_IO_TSK(MSPB_TSK_ERR_BAD_COLUMN_TYPE);
    }
    if (!have_data) {
        return MSPB_IO_TSK(MSPB_TSK_ERR_REQUIRED_COL_NOT_FOUND);
    }
    ret = mspb_store_get(s, off_key, &o);
    if (ret == NOT_FOUND) {
        return MSPB_IO_TSK(MSPB_TSK_ERR_BOTH_COLUMNS_REQUIRED);
    }
    if (ret != 0) {
        return ret;
    }
    if (o.array_len == 0) {
        return MSPB_IO_TSK(MSPB_TSK_ERR_FILE_FORMAT);
    }
    if (*num_rows == ROWS_UNSET) {
        *num_rows = o.array_len - 1;
    } else if (*num_rows != o.array_len - 1) {
        return MSPB_IO_TSK(MSPB_TSK_ERR_FILE_FORMAT);
    }
    if (o.type == MSPB_KAS_UINT64) {
        *offset = (const uint64_t *) o.array;
    } else if (o.type == MSPB_KAS_UINT32) {
        if (s->num_owned >= (int) (sizeof(s->owned) / sizeof(s->owned[0]))) {
            return MSPB_IO_KAS(MSPB_KAS_ERR_GENERIC);
        }
        uint64_t *wide = malloc((size_t) o.array_len * sizeof(*wide));
        if (wide == NULL) {
            return MSPB_IO_KAS(MSPB_KAS_ERR_NO_MEMORY);
        }
        const uint32_t *narrow = (const uint32_t *) o.array;
        for (uint64_t j = 0; j < o.array_len; j++) {
            wide[j] = narrow[j];
        }
        s->owned[s->num_owned++] = wide;
        *offset = wide;
    } else {
        return MSPB_IO_TSK(MSPB_TSK_ERR_BAD_COLUMN_TYPE);
    }
    if ((*offset)[*num_rows] != d.array_len) {
        return MSPB_IO_TSK(MSPB_TSK_ERR_BAD_OFFSET);
    }
    *data = (const char *) d.array;
    return 0;
}

int
mspb_store_tables(mspb_store_t *s, mspb_tables_t *t, const char **time_units, uint64_t *time_units_len)
{
    mspb_store_item_t it;
    int ret;
    const void *col;
    uint64_t n;

    memset(t, 0, sizeof(*t));
    for (int j = 0; j < s->num_owned; j++) {
        free(s->owned[j]);
    }
    s->num_owned = 0;
    /* format/name, format/version, sequence_length: tables.c:11588-11634 */
    ret = mspb_store_get(s, "format/name", &it);
    if (ret != 0) {
        goto out;
    }
    if (it.type != MSPB_KAS_INT8) {
        ret = MSPB_IO_KAS(MSPB_KAS_ERR_TYPE_MISMATCH);
        goto out;
    }
    if (it.array_len != strlen(TREES_FORMAT_NAME) || memcmp(it.array, TREES_FORMAT_NAME, it.array_len) != 0) {
        ret = MSPB_IO_TSK(MSPB_TSK_ERR_FILE_FORMAT);
        goto out;
    }
    ret = mspb_store_get(s, "format/version", &it);
    if (ret != 0) {
        goto out;
    }
    if (it.type != MSPB_KAS_UINT32) {
        ret = MSPB_IO_KAS(MSPB_KAS_ERR_TYPE_MISMATCH);
        goto out;
    }
    if (it.array_len != 2) {
        ret = MSPB_IO_TSK(MSPB_TSK_ERR_FILE_FORMAT);
        goto out;
    }
    const uint32_t major = ((const uint32_t *) it.array)[0];
    if (major < TREES_VERSION_MAJOR) {
        ret = MSPB_IO_TSK(MSPB_TSK_ERR_FILE_VERSION_TOO_OLD);
        goto out;
    }
    if (major > TREES_VERSION_MAJOR) {
        ret = MSPB_IO_TSK(MSPB_TSK_ERR_FILE_VERSION_TOO_NEW);
        goto out;
    }
    ret = mspb_store_get(s, "sequence_length", &it);
    if (ret != 0) {
        goto out;
    }
    if (it.type != MSPB_KAS_FLOAT64) {
        ret = MSPB_IO_KAS(MSPB_KAS_ERR_TYPE_MISMATCH);
        goto out;
    }
    if (it.array_len != 1) {
        ret = MSPB_IO_TSK(MSPB_TSK_ERR_FILE_FORMAT);
        goto out;
    }
    t->sequence_length = ((const double *) it.array)[0];
    if (!(t->sequence_length > 0.0)) {
        ret = MSPB_IO_TSK(MSPB_TSK_ERR_BAD_SEQUENCE_LENGTH);
        goto out;
    }
    if (time_units != NULL) {
        *time_units = NULL;
        *time_units_len = 0;
        ret = mspb_store_get(s, "time_units", &it);
        if (ret == 0) {
            if (it.type != MSPB_KAS_INT8) {
                ret = MSPB_IO_KAS(MSPB_KAS_ERR_TYPE_MISMATCH);
                goto out;
            }
            *time_units = (const char *) it.array;
            *time_units_len = it.array_len;
        } else if (ret != NOT_FOUND) {
            goto out;
        }
    }

    /* nodes: tables.c:2583-2586 (only the time column is read by the overlay, the others
     * are checked for presence and shape as the reference's loader would) */
    n = ROWS_UNSET;
    if ((ret = get_col(s, "nodes/time", MSPB_KAS_FLOAT64, 0, &n, &col)) != 0) {
        goto out;
    }
    t->node_time = (const double *) col;
    if ((ret = get_col(s, "nodes/flags", MSPB_KAS_UINT32, 0, &n, &col)) != 0) {
        goto out;
    }
    t->num_nodes = n;
    /* edges: tables.c:3289-3292 */
    n = ROWS_UNSET;
    if ((ret = get_col(s, "edges/left", MSPB_KAS_FLOAT64, 0, &n, &col)) != 0) {
        goto out;
    }
    t->edge_left = (const double *) col;
    if ((ret = get_col(s, "edges/right", MSPB_KAS_FLOAT64, 0, &n, &col)) != 0) {
        goto out;
    }
    t->edge_right = (const double *) col;
    if ((ret = get_col(s, "edges/parent", MSPB_KAS_INT32, 0, &n, &col)) != 0) {
        goto out;
    }
    t->edge_parent = (const int32_t *) col;
    if ((ret = get_col(s, "edges/child", MSPB_KAS_INT32, 0, &n, &col)) != 0) {
        goto out;
    }
    t->edge_child = (const int32_t *) col;
    t->num_edges = n;
    /* sites: tables.c:4030-4041 */
    n = ROWS_UNSET;
    if ((ret = get_col(s, "sites/position", MSPB_KAS_FLOAT64, 0, &n, &col)) != 0) {
        goto out;
    }
    t->site_position = (const double *) col;
    if ((ret = get_ragged(s, "sites/ancestral_state", &n, &t->site_ancestral_state, &t->site_ancestral_state_offset))
        != 0) {
        goto out;
    }
    if ((ret = get_ragged(s, "sites/metadata", &n, &t->site_metadata, &t->site_metadata_offset)) != 0) {
        goto out;
    }
    t->num_sites = n;
    /* mutations: tables.c:4840-4851; time is optional (absent = all unknown) */
    n = ROWS_UNSET;
    if ((ret = get_col(s, "mutations/site", MSPB_KAS_INT32, 0, &n, &col)) != 0) {
        goto out;
    }
    t->mutation_site = (const int32_t *) col;
    if ((ret = get_col(s, "mutations/node", MSPB_KAS_INT32, 0, &n, &col)) != 0) {
        goto out;
    }
    t->mutation_node = (const int32_t *) col;
    if ((ret = get_col(s, "mutations/parent", MSPB_KAS_INT32, 0, &n, &col)) != 0) {
        goto out;
    }
    if ((ret = get_col(s, "mutations/time", MSPB_KAS_FLOAT64, 1, &n, &col)) != 0) {
        goto out;
    }
    t->mutation_time = (const double *) col;
    if ((ret = get_ragged(
             s, "mutations/derived_state", &n, &t->mutation_derived_state, &t->mutation_derived_state_offset))
        != 0) {
        goto out;
    }
    if ((ret = get_ragged(s, "mutations/metadata", &n, &t->mutation_metadata, &t->mutation_metadata_offset)) != 0) {
        goto out;
    }
    t->num_mutations = n;
out:
    /* tables.c:11706: a missing key of the format block is a missing required column */
    if (ret == NOT_FOUND) {
        ret = MSPB_IO_TSK(MSPB_TSK_ERR_REQUIRED_COL_NOT_FOUND);
    }
    return ret;
}

/* ---------------------------------------------------------------- writer */

typedef struct {
    const mspb_store_item_t *item;
    uint64_t key_start, array_start;
} out_entry_t;

static int
out_cmp(const void *a, const void *b)
{
    const out_entry_t *x = a, *y = b;
    return key_cmp(x->item->key, x->item->key_len, y->item->key, y->item->key_len);
}

static int
write_all(int fd, const void *buf, uint64_t n)
{
    const unsigned char *p = buf;
    while (n > 0) {
        /* large arrays go out in 256 MiB pieces (write(2) is capped near 2 GiB) */
        const size_t chunk = n > (1u << 28) ? (1u << 28) : (size_t) n;
        const ssize_t w = write(fd, p, chunk);
        if (w < 0) {
            if (errno == EINTR) {
                continue;
            }
            return MSPB_KAS_ERR_IO;
        }
        p += w;
        n -= (uint64_t) w;
    }
    return 0;
}

int
mspb_store_write(const char *path, const mspb_store_item_t *items, uint64_t num_items)
{
    int ret = 0;
    int fd = -1;
    out_entry_t *order = NULL;
    unsigned char *front = NULL;
    static const unsigned char ZERO[ARRAY_ALIGN] = { 0 };

    if (num_items > UINT32_MAX) {
        return MSPB_IO_KAS(MSPB_KAS_ERR_GENERIC);
    }
    order = malloc((size_t) (num_items ? num_items : 1) * sizeof(*order));
    if (order == NULL) {
        return MSPB_IO_KAS(MSPB_KAS_ERR_NO_MEMORY);
    }
    for (uint64_t j = 0; j < num_items; j++) {
        if (items[j].key_len == 0) {
            ret = MSPB_KAS_ERR_EMPTY_KEY;
            goto out;
        }
        if (items[j].type < 0 || items[j].type >= MSPB_KAS_NUM_TYPES) {
            ret = MSPB_KAS_ERR_BAD_TYPE;
            goto out;
        }
        order[j].item = &items[j];
    }
    qsort(order, (size_t) num_items, sizeof(*order), out_cmp);
    for (uint64_t j = 1; j < num_items; j++) {
        if (out_cmp(&order[j - 1], &order[j]) == 0) {
            ret = MSPB_KAS_ERR_DUPLICATE_KEY;
            goto out;
        }
    }
    /* layout: header, descriptors, keys, then every array on an 8-byte boundary */
    uint64_t offset = HEADER_SIZE + num_items * DESCRIPTOR_SIZE;
    for (uint64_t j = 0; j < num_items; j++) {
        order[j].key_start = offset;
        offset += order[j].item->key_len;
    }
    const uint64_t front_size = offset;
    for (uint64_t j = 0; j < num_items; j++) {
        offset = (offset + ARRAY_ALIGN - 1) / ARRAY_ALIGN * ARRAY_ALIGN;
        order[j].array_start = offset;
        offset += order[j].item->array_len * TYPE_SIZE[order[j].item->type];
    }
    const uint64_t file_size = offset;

    front = calloc(1, (size_t) front_size);
    if (front == NULL) {
        ret = MSPB_KAS_ERR_NO_MEMORY;
        goto out;
    }
    const uint16_t major = FILE_VERSION_MAJOR, minor = FILE_VERSION_MINOR;
    const uint32_t n32 = (uint32_t) num_items;
    memcpy(front, MAGIC, 8);
    memcpy(front + 8, &major, 2);
    memcpy(front + 10, &minor, 2);
    memcpy(front + 12, &n32, 4);
    memcpy(front + 16, &file_size, 8);
    for (uint64_t j = 0; j < num_items; j++) {
        unsigned char *d = front + HEADER_SIZE + j * DESCRIPTOR_SIZE;
        const mspb_store_item_t *it = order[j].item;
        d[0] = (unsigned char) it->type;
        memcpy(d + 8, &order[j].key_start, 8);
        memcpy(d + 16, &it->key_len, 8);
        memcpy(d + 24, &order[j].array_start, 8);
        memcpy(d + 32, &it->array_len, 8);
        memcpy(front + order[j].key_start, it->key, (size_t) it->key_len);
    }
    fd = open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) {
        ret = MSPB_KAS_ERR_IO;
        goto out;
    }
    if ((ret = write_all(fd, front, front_size)) != 0) {
        goto out;
    }
    offset = front_size;
    for (uint64_t j = 0; j < num_items; j++) {
        const mspb_store_item_t *it = order[j].item;
        if ((ret = write_all(fd, ZERO, order[j].array_start - offset)) != 0) {
            goto out;
        }
        const uint64_t bytes = it->array_len * TYPE_SIZE[it->type];
        if ((ret = write_all(fd, it->array, bytes)) != 0) {
            goto out;
        }
        offset = order[j].array_start + bytes;
    }
out:
    if (fd >= 0) {
        if (close(fd) != 0 && ret == 0) {
            ret = MSPB_KAS_ERR_IO;
        }
        if (ret != 0) {
            unlink(path); /* as tsk_table_collection_dump: no half-written file is left behind */
        }
    }
    free(front);
    free(order);
    return ret == 0 ? 0 : MSPB_IO_KAS(ret);
}

const char *
mspb_io_strerror(int err)
{
    if (err == 0) {
        return "Normal exit condition.";
    }
    const int tsk = err ^ (1 << 13);
    switch (tsk) {
        case MSPB_TSK_ERR_IO:
            return strerror(errno);
        case MSPB_TSK_ERR_FILE_FORMAT:
            return "File format error.";
        case MSPB_TSK_ERR_FILE_VERSION_TOO_OLD:
            return "File format version is too old.";
        case MSPB_TSK_ERR_FILE_VERSION_TOO_NEW:
            return "File format version is too new.";
        case MSPB_TSK_ERR_REQUIRED_COL_NOT_FOUND:
            return "A required column was not found in the file.";
        case MSPB_TSK_ERR_BOTH_COLUMNS_REQUIRED:
            return "Both columns in a related pair must be provided.";
        case MSPB_TSK_ERR_BAD_COLUMN_TYPE:
            return "An incompatible type for a column was found in the file.";
        case MSPB_TSK_ERR_BAD_OFFSET:
            return "Bad offset provided in input array.";
        case MSPB_TSK_ERR_BAD_SEQUENCE_LENGTH:
            return "Sequence length must be > 0.";
        default:
            break;
    }
    switch (tsk ^ (1 << 14)) {
        case MSPB_KAS_ERR_GENERIC:
            return "Generic error; please file a bug report.";
        case MSPB_KAS_ERR_NO_MEMORY:
            return "Out of memory.";
        case MSPB_KAS_ERR_BAD_FILE_FORMAT:
            return "File not in kastore format.";
        case MSPB_KAS_ERR_VERSION_TOO_OLD:
            return "File format version is too old.";
        case MSPB_KAS_ERR_VERSION_TOO_NEW:
            return "File format version is too new.";
        case MSPB_KAS_ERR_BAD_TYPE:
            return "Unknown data type.";
        case MSPB_KAS_ERR_EMPTY_KEY:
            return "Keys cannot be empty.";
        case MSPB_KAS_ERR_DUPLICATE_KEY:
            return "Duplicate keys are not permitted.";
        case MSPB_KAS_ERR_KEY_NOT_FOUND:
            return "Key not found.";
        case MSPB_KAS_ERR_TYPE_MISMATCH:
            return "Mismatch between requested and stored types for array.";
        case MSPB_KAS_ERR_EOF:
            return "End of file.";
        default:
            break;
    }
    return "Unknown error.";
}
