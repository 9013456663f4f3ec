/* TEST INFRASTRUCTURE (oracle).  Not part of the product path.
 *
 * Driver that exposes the UNMODIFIED reference mutation generator
 * (/root/reference/lib/mutgen.c, compiled where it lies by oracle/Makefile)
 * through the flat interface of oracle_api.h.  It does what
 * msprime/_msprimemodule.c:3047-3120 (msprime_sim_mutations) does: build the
 * model, mutgen_alloc, set the rate map and time interval, mutgen_generate with
 * the MSP_DISCRETE_SITES / MSP_KEEP_SITES flags, then read the tables back.
 * No reference source is copied here; this file only calls its public C API.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "msprime.h"
#include "oracle_api.h"

const char *
orc_name(void)
{
    return "reference lib/mutgen.c + gsl shim";
}

static void *
dup_bytes(const void *src, size_t n)
{
    void *p = malloc(n > 0 ? n : 1);
    if (p != NULL && n > 0) {
        memcpy(p, src, n);
    }
    return p;
}

static const char EMPTY[1] = { 0 };
#define NZ(p) ((p) != NULL ? (p) : EMPTY)

static int
load_tables(const orc_input_t *in, tsk_table_collection_t *tables)
{
    int ret;
    tsk_flags_t *flags = calloc(in->num_nodes > 0 ? in->num_nodes : 1, sizeof(*flags));

    if (flags == NULL) {
        return MSP_ERR_NO_MEMORY;
    }
    tables->sequence_length = in->sequence_length;
    ret = tsk_node_table_set_columns(
        &tables->nodes, in->num_nodes, flags, in->node_time, NULL, NULL, NULL, NULL);
    free(flags);
    if (ret != 0) {
        return msp_set_tsk_error(ret);
    }
    ret = tsk_edge_table_set_columns(&tables->edges, in->num_edges, in->edge_left,
        in->edge_right, in->edge_parent, in->edge_child, NULL, NULL);
    if (ret != 0) {
        return msp_set_tsk_error(ret);
    }
    if (in->num_sites > 0) {
        ret = tsk_site_table_set_columns(&tables->sites, in->num_sites, in->site_position,
            NZ(in->site_ancestral_state), (const tsk_size_t *) in->site_ancestral_state_offset,
            NZ(in->site_metadata), (const tsk_size_t *) in->site_metadata_offset);
        if (ret != 0) {
            return msp_set_tsk_error(ret);
        }
    }
    if (in->num_mutations > 0) {
        ret = tsk_mutation_table_set_columns(&tables->mutations, in->num_mutations,
            in->mutation_site, in->mutation_node, NULL, in->mutation_time,
            NZ(in->mutation_derived_state),
            (const tsk_size_t *) in->mutation_derived_state_offset, NZ(in->mutation_metadata),
            (const tsk_size_t *) in->mutation_metadata_offset);
        if (ret != 0) {
            return msp_set_tsk_error(ret);
        }
    }
    return 0;
}

static int
build_model(const orc_input_t *in, mutation_model_t *model)
{
    int ret;
    size_t j;

    if (in->model_kind == ORC_MODEL_INFINITE_ALLELES) {
        return infinite_alleles_mutation_model_alloc(model, in->start_allele, 0);
    } else if (in->model_kind == ORC_MODEL_SLIM) {
        return slim_mutation_model_alloc(model, in->slim_mutation_type_id, (int64_t) in->start_allele,
            in->slim_generation, 0);
    } else {
        char **alleles = calloc(in->num_alleles + 1, sizeof(*alleles));
        size_t *lengths = calloc(in->num_alleles + 1, sizeof(*lengths));
        if (alleles == NULL || lengths == NULL) {
            free(alleles);
            free(lengths);
            return MSP_ERR_NO_MEMORY;
        }
        for (j = 0; j < in->num_alleles; j++) {
            alleles[j] = (char *) (uintptr_t) (in->alleles + in->allele_offset[j]);
            lengths[j] = (size_t) (in->allele_offset[j + 1] - in->allele_offset[j]);
        }
        ret = matrix_mutation_model_alloc(model, in->num_alleles, alleles, lengths,
            (double *) (uintptr_t) in->root_distribution,
            (double *) (uintptr_t) in->transition_matrix);
        free(alleles);
        free(lengths);
        return ret;
    }
}

static int
export_tables(tsk_table_collection_t *tables, orc_output_t *out)
{
    tsk_site_table_t *s = &tables->sites;
    tsk_mutation_table_t *m = &tables->mutations;

    out->num_sites = s->num_rows;
    out->site_position = dup_bytes(s->position, s->num_rows * sizeof(double));
    out->site_ancestral_state = dup_bytes(s->ancestral_state, s->ancestral_state_length);
    out->site_ancestral_state_offset
        = dup_bytes(s->ancestral_state_offset, (s->num_rows + 1) * sizeof(uint64_t));
    out->site_metadata = dup_bytes(s->metadata, s->metadata_length);
    out->site_metadata_offset
        = dup_bytes(s->metadata_offset, (s->num_rows + 1) * sizeof(uint64_t));
    out->num_mutations = m->num_rows;
    out->mutation_site = dup_bytes(m->site, m->num_rows * sizeof(int32_t));
    out->mutation_node = dup_bytes(m->node, m->num_rows * sizeof(int32_t));
    out->mutation_parent = dup_bytes(m->parent, m->num_rows * sizeof(int32_t));
    out->mutation_time = dup_bytes(m->time, m->num_rows * sizeof(double));
    out->mutation_derived_state = dup_bytes(m->derived_state, m->derived_state_length);
    out->mutation_derived_state_offset
        = dup_bytes(m->derived_state_offset, (m->num_rows + 1) * sizeof(uint64_t));
    out->mutation_metadata = dup_bytes(m->metadata, m->metadata_length);
    out->mutation_metadata_offset
        = dup_bytes(m->metadata_offset, (m->num_rows + 1) * sizeof(uint64_t));
    return 0;
}

int
orc_sim_mutations(const orc_input_t *in, orc_output_t *out)
{
    int ret;
    int flags = 0;
    tsk_table_collection_t tables;
    mutation_model_t model;
    mutgen_t mutgen;
    gsl_rng *rng = NULL;
    int have_tables = 0, have_model = 0, have_mutgen = 0;
    struct timespec t0, t1;

    memset(out, 0, sizeof(*out));
    memset(&model, 0, sizeof(model));
    memset(&mutgen, 0, sizeof(mutgen));

    rng = gsl_rng_alloc(in->rng_kind == ORC_RNG_REPLAY ? gsl_rng_replay : gsl_rng_mt19937);
    if (rng == NULL) {
        ret = MSP_ERR_NO_MEMORY;
        goto out;
    }
    if (in->rng_kind == ORC_RNG_REPLAY) {
        gsl_rng_replay_set(rng, in->stream, in->stream_len);
    } else {
        gsl_rng_set(rng, in->seed);
    }
    ret = tsk_table_collection_init(&tables, 0);
    if (ret != 0) {
        ret = msp_set_tsk_error(ret);
        goto out;
    }
    have_tables = 1;
    ret = load_tables(in, &tables);
    if (ret != 0) {
        goto out;
    }
    ret = build_model(in, &model);
    have_model = 1;
    if (ret != 0) {
        goto out;
    }
    ret = mutgen_alloc(&mutgen, rng, &tables, &model, 0);
    have_mutgen = 1;
    if (ret != 0) {
        goto out;
    }
    ret = mutgen_set_rate_map(&mutgen, in->map_size, (double *) (uintptr_t) in->map_position,
        (double *) (uintptr_t) in->map_rate);
    if (ret != 0) {
        goto out;
    }
    ret = mutgen_set_time_interval(&mutgen, in->start_time, in->end_time);
    if (ret != 0) {
        goto out;
    }
    if (in->discrete_genome) {
        flags |= MSP_DISCRETE_SITES;
    }
    if (in->keep) {
        flags |= MSP_KEEP_SITES;
    }
    clock_gettime(CLOCK_MONOTONIC, &t0);
    ret = mutgen_generate(&mutgen, flags);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    out->seconds = (double) (t1.tv_sec - t0.tv_sec) + 1e-9 * (double) (t1.tv_nsec - t0.tv_nsec);
    if (ret != 0) {
        goto out;
    }
    if (in->model_kind == ORC_MODEL_INFINITE_ALLELES) {
        out->next_allele = model.params.infinite_alleles.next_allele;
    } else if (in->model_kind == ORC_MODEL_SLIM) {
        out->next_allele = (uint64_t) model.params.slim_mutator.next_mutation_id;
    }
    ret = export_tables(&tables, out);
out:
    if (rng != NULL) {
        out->draws = rng->draws;
        out->stream_exhausted = rng->exhausted;
    }
    if (have_mutgen) {
        mutgen_free(&mutgen);
    }
    if (have_model) {
        mutation_model_free(&model);
    }
    if (have_tables) {
        tsk_table_collection_free(&tables);
    }
    if (rng != NULL) {
        gsl_rng_free(rng);
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

/* ---- `.trees` files through the reference's own tskit C (tsk_table_collection_dump /
 * _load / _equals, git-submodules/tskit/c/tskit/tables.c): the checker for the product's
 * file reader / writer (msprime_b200/csrc/trees_io.c).  Reference build only. */

/* Writes the tree sequence of `in` (kept sites / mutations included, sample flags on the
 * first num_samples nodes, indexes built, one provenance row) as a .trees file. */
int
orc_trees_dump(const orc_input_t *in, uint64_t num_samples, const char *time_units, const char *path)
{
    int ret;
    uint64_t j;
    tsk_table_collection_t tables;
    static const char *TS = "2026-01-01T00:00:00";
    static const char *REC = "{\"software\": \"oracle\"}";

    ret = tsk_table_collection_init(&tables, 0);
    if (ret != 0) {
        return msp_set_tsk_error(ret);
    }
    ret = load_tables(in, &tables);
    if (ret != 0) {
        goto out;
    }
    for (j = 0; j < num_samples && j < tables.nodes.num_rows; j++) {
        tables.nodes.flags[j] = TSK_NODE_IS_SAMPLE;
    }
    ret = tsk_table_collection_set_time_units(&tables, time_units, strlen(time_units));
    if (ret != 0) {
        ret = msp_set_tsk_error(ret);
        goto out;
    }
    ret = tsk_provenance_table_add_row(&tables.provenances, TS, strlen(TS), REC, strlen(REC));
    if (ret < 0) {
        ret = msp_set_tsk_error(ret);
        goto out;
    }
    ret = tsk_table_collection_build_index(&tables, 0);
    if (ret != 0) {
        ret = msp_set_tsk_error(ret);
        goto out;
    }
    /* the input columns carry no parents (mutgen recomputes them): make the file a valid
     * tree sequence, as a file written by tskit after sim_mutations would be */
    ret = tsk_table_collection_compute_mutation_parents(&tables, 0);
    if (ret != 0) {
        ret = msp_set_tsk_error(ret);
        goto out;
    }
    ret = tsk_table_collection_dump(&tables, path, 0);
    if (ret != 0) {
        ret = msp_set_tsk_error(ret);
    }
out:
    tsk_table_collection_free(&tables);
    return ret;
}

/* Loads a .trees file with the reference and hands its site / mutation tables back. */
int
orc_trees_load(const char *path, orc_output_t *out)
{
    int ret;
    tsk_table_collection_t tables;

    memset(out, 0, sizeof(*out));
    ret = tsk_table_collection_load(&tables, path, 0);
    if (ret != 0) {
        ret = msp_set_tsk_error(ret);
        tsk_table_collection_free(&tables);
        out->status = ret;
        return ret;
    }
    ret = export_tables(&tables, out);
    tsk_table_collection_free(&tables);
    out->status = ret;
    return ret;
}

/* 1 when the reference finds the two files' table collections equal (every table, the
 * metadata, time units and provenance; tsk_table_collection_equals with options 0),
 * 0 when not, negative on a load error (`which` says which file). */
int
orc_trees_equal(const char *path_a, const char *path_b, int32_t *which)
{
    int ret;
    tsk_table_collection_t a, b;

    *which = 0;
    ret = tsk_table_collection_load(&a, path_a, 0);
    if (ret != 0) {
        tsk_table_collection_free(&a);
        return msp_set_tsk_error(ret);
    }
    *which = 1;
    ret = tsk_table_collection_load(&b, path_b, 0);
    if (ret != 0) {
        tsk_table_collection_free(&a);
        tsk_table_collection_free(&b);
        return msp_set_tsk_error(ret);
    }
    /* also require what equals() does not look at: the loaded trees must pass the
     * reference's integrity check with indexes present */
    ret = (int) tsk_table_collection_check_integrity(&b, TSK_CHECK_TREES);
    if (ret < 0) {
        ret = msp_set_tsk_error(ret);
    } else {
        ret = tsk_table_collection_equals(&a, &b, 0) ? 1 : 0;
    }
    tsk_table_collection_free(&a);
    tsk_table_collection_free(&b);
    return ret;
}

/* Reference load + dump of a file (what `tskit.load(x).dump(y)` does). */
int
orc_trees_copy(const char *path_in, const char *path_out)
{
    int ret;
    tsk_table_collection_t tables;

    ret = tsk_table_collection_load(&tables, path_in, 0);
    if (ret == 0) {
        ret = tsk_table_collection_dump(&tables, path_out, 0);
    }
    tsk_table_collection_free(&tables);
    return ret != 0 ? msp_set_tsk_error(ret) : 0;
}
