/*
 * msprime_b200 file hand-off: the `.trees` container on either side of the overlay
 * (SURVEY.md section 8 f3).  C ABI, host only, no CUDA or torch types.
 *
 * A `.trees` file is a kastore: a 64-byte header, one 64-byte descriptor per item, the
 * keys packed back to back, then the arrays, each 8-byte aligned, items ordered by key.
 * What each entry point replaces in the reference:
 *
 *   mspb_store_open        <- kastore_open(..., "r", 0) + kastore_read_header /
 *                             kastore_read_descriptors,
 *                             git-submodules/tskit/c/subprojects/kastore/kastore.c:154-337, 547
 *   mspb_store_get         <- kastore_gets, kastore.c:686-760
 *   mspb_store_write       <- kastore_puts x n + kastore_close (sort by key, pack, write),
 *                             kastore.c:200-220, 464-484
 *   mspb_store_tables      <- the node / edge / site / mutation part of
 *                             tsk_table_collection_load, git-submodules/tskit/c/tskit/tables.c:11588-11720
 *                             (format/name, format/version, sequence_length checks; uint32
 *                             offset columns widened as cast_offset_array does, tables.c:245)
 *
 * The reader maps the file; arrays handed out point INTO the mapping (zero copy: they go
 * straight to mspb_ctx_load_tables) and stay valid until mspb_store_close.
 *
 * Errors: every entry point returns 0 or a negative code in the numbering the reference
 * stack (msprime <- tskit <- kastore) would hand to Python: see MSPB_IO_KAS / MSPB_IO_TSK.
 */
#ifndef MSPRIME_B200_IO_H
#define MSPRIME_B200_IO_H

#include <stdint.h>

#include "msprime_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* kastore.h:35-93 */
#define MSPB_KAS_ERR_GENERIC (-1)
#define MSPB_KAS_ERR_IO (-2)
#define MSPB_KAS_ERR_NO_MEMORY (-4)
#define MSPB_KAS_ERR_BAD_FILE_FORMAT (-5)
#define MSPB_KAS_ERR_VERSION_TOO_OLD (-6)
#define MSPB_KAS_ERR_VERSION_TOO_NEW (-7)
#define MSPB_KAS_ERR_BAD_TYPE (-8)
#define MSPB_KAS_ERR_EMPTY_KEY (-9)
#define MSPB_KAS_ERR_DUPLICATE_KEY (-10)
#define MSPB_KAS_ERR_KEY_NOT_FOUND (-11)
#define MSPB_KAS_ERR_TYPE_MISMATCH (-13)
#define MSPB_KAS_ERR_EOF (-14)
/* Values RETURNED: a kastore code e travels as the reference stack would report it,
 * tsk_set_kas_error (core.c:712-722: IO -> TSK_ERR_IO, else bit 14 flipped) and then
 * msp_set_tsk_error (lib/util.c:337-342: bit 13 flipped); a tskit code t as t ^ (1 << 13). */
#define MSPB_IO_KAS(e) ((e) == MSPB_KAS_ERR_IO ? MSPB_IO_TSK(MSPB_TSK_ERR_IO) : (((e) ^ (1 << 14)) ^ (1 << 13)))
#define MSPB_IO_TSK(t) ((t) ^ (1 << 13))
/* schema level (tskit core.h:263-568) */
#define MSPB_TSK_ERR_IO (-3)
#define MSPB_TSK_ERR_FILE_FORMAT (-100)
#define MSPB_TSK_ERR_FILE_VERSION_TOO_OLD (-101)
#define MSPB_TSK_ERR_FILE_VERSION_TOO_NEW (-102)
#define MSPB_TSK_ERR_REQUIRED_COL_NOT_FOUND (-103)
#define MSPB_TSK_ERR_BOTH_COLUMNS_REQUIRED (-104)
#define MSPB_TSK_ERR_BAD_COLUMN_TYPE (-105)
#define MSPB_TSK_ERR_BAD_OFFSET (-200)
#define MSPB_TSK_ERR_BAD_SEQUENCE_LENGTH (-701)

/* array element types (kastore.h:108-117) */
enum {
    MSPB_KAS_INT8 = 0, MSPB_KAS_UINT8 = 1, MSPB_KAS_INT16 = 2, MSPB_KAS_UINT16 = 3, MSPB_KAS_INT32 = 4,
    MSPB_KAS_UINT32 = 5, MSPB_KAS_INT64 = 6, MSPB_KAS_UINT64 = 7, MSPB_KAS_FLOAT32 = 8, MSPB_KAS_FLOAT64 = 9,
    MSPB_KAS_NUM_TYPES = 10
};

typedef struct {
    const char *key; /* not NUL terminated inside a mapped store: use key_len */
    uint64_t key_len;
    int32_t type;
    const void *array;
    uint64_t array_len; /* elements */
} mspb_store_item_t;

typedef struct mspb_store mspb_store_t;

/* Opens and validates a kastore file (read only, memory mapped). */
int mspb_store_open(const char *path, mspb_store_t **store);
void mspb_store_close(mspb_store_t *store);
uint64_t mspb_store_num_items(const mspb_store_t *store);
/* Item `index` in key order. */
int mspb_store_item(const mspb_store_t *store, uint64_t index, mspb_store_item_t *item);
/* Item by key (binary search over the sorted keys). */
int mspb_store_get(const mspb_store_t *store, const char *key, mspb_store_item_t *item);

/* Checks the tskit.trees schema and points `tables` at the columns the overlay reads
 * (kept sites and mutations included).  uint32 offset columns are widened into buffers
 * owned by the store.  `time_units` (optional) receives a pointer/length into the file. */
int mspb_store_tables(mspb_store_t *store, mspb_tables_t *tables, const char **time_units,
    uint64_t *time_units_len);

/* Writes a kastore file holding `items` (any order; keys must be unique and non-empty). */
int mspb_store_write(const char *path, const mspb_store_item_t *items, uint64_t num_items);

const char *mspb_io_strerror(int err);

#ifdef __cplusplus
}
#endif

#endif
