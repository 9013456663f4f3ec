/* TEST INFRASTRUCTURE (oracle): input generator.  Runs the reference's OWN ancestry simulator
 * (lib/msprime.c, standard Hudson coalescent with recombination, compiled unmodified from where it lies
 * under /root/reference by `make ancestry`) and hands back the node and edge tables, so that the parity
 * tests of the mutation path can run on the tree sequences the reference itself would pass to
 * sim_mutations (route 2) instead of on this project's own generator.
 *
 * The calls below are what msprime/ancestry.py `_parse_sim_ancestry` + `Simulator.__init__` make for
 * sim_ancestry(samples=n, sequence_length=L, recombination_rate=r, population_size=N, ploidy=p,
 * random_seed=s): sample nodes at time 0 in population 0, one population of constant size, a flat
 * recombination map, the Hudson model, run to completion, finalise the tables
 * (lib/tests/test_ancestry.c runs the library the same way).  Nothing in the product path links this.
 */
#include <float.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <gsl/gsl_rng.h>

#include "msprime.h"

typedef struct {
    msp_t msp;
    tsk_table_collection_t tables;
    gsl_rng *rng;
    int have_msp;
} ancestry_run_t;

void
mspb_ref_ancestry_free(void *handle)
{
    ancestry_run_t *run = (ancestry_run_t *) handle;

    if (run == NULL) {
        return;
    }
    if (run->have_msp) {
        msp_free(&run->msp);
    }
    tsk_table_collection_free(&run->tables);
    if (run->rng != NULL) {
        gsl_rng_free(run->rng);
    }
    free(run);
}

/* Returns 0 or the library's (negative) error code; *events receives the number of events the run took. */
int
mspb_ref_ancestry_run(uint32_t num_sample_nodes, double sequence_length, double recombination_rate,
    double population_size, int ploidy, int discrete_genome, unsigned long seed, uint64_t *num_nodes,
    uint64_t *num_edges, void **handle)
{
    int ret;
    uint32_t j;
    ancestry_run_t *run = calloc(1, sizeof(*run));

    *handle = NULL;
    if (run == NULL) {
        return -2;
    }
    run->rng = gsl_rng_alloc(gsl_rng_mt19937);
    gsl_rng_set(run->rng, seed);
    ret = tsk_table_collection_init(&run->tables, 0);
    if (ret != 0) {
        goto out;
    }
    run->tables.sequence_length = sequence_length;
    for (j = 0; j < num_sample_nodes; j++) {
        ret = (int) tsk_node_table_add_row(
            &run->tables.nodes, TSK_NODE_IS_SAMPLE, 0.0, 0, TSK_NULL, NULL, 0);
        if (ret < 0) {
            goto out;
        }
    }
    ret = (int) tsk_population_table_add_row(&run->tables.populations, NULL, 0);
    if (ret < 0) {
        goto out;
    }
    ret = msp_alloc(&run->msp, &run->tables, run->rng);
    run->have_msp = 1;
    if (ret != 0) {
        goto out;
    }
    if ((ret = msp_set_discrete_genome(&run->msp, discrete_genome != 0)) != 0
        || (ret = msp_set_ploidy(&run->msp, ploidy)) != 0
        || (ret = msp_set_recombination_rate(&run->msp, recombination_rate)) != 0
        || (ret = msp_set_population_configuration(&run->msp, 0, population_size, 0.0, true)) != 0
        || (ret = msp_set_simulation_model_hudson(&run->msp)) != 0
        || (ret = msp_initialise(&run->msp)) != 0) {
        goto out;
    }
    ret = msp_run(&run->msp, DBL_MAX, UINT32_MAX);
    while (ret == MSP_EXIT_MAX_EVENTS) {
        ret = msp_run(&run->msp, DBL_MAX, UINT32_MAX);
    }
    if (ret != 0) {
        goto out;
    }
    ret = msp_finalise_tables(&run->msp);
    if (ret != 0) {
        goto out;
    }
    *num_nodes = run->tables.nodes.num_rows;
    *num_edges = run->tables.edges.num_rows;
    *handle = run;
    return 0;
out:
    mspb_ref_ancestry_free(run);
    return ret;
}

void
mspb_ref_ancestry_copy(void *handle, uint32_t *flags, double *time, double *left, double *right,
    int32_t *parent, int32_t *child)
{
    ancestry_run_t *run = (ancestry_run_t *) handle;
    tsk_size_t n = run->tables.nodes.num_rows, e = run->tables.edges.num_rows;

    memcpy(flags, run->tables.nodes.flags, n * sizeof(*flags));
    memcpy(time, run->tables.nodes.time, n * sizeof(*time));
    memcpy(left, run->tables.edges.left, e * sizeof(*left));
    memcpy(right, run->tables.edges.right, e * sizeof(*right));
    memcpy(parent, run->tables.edges.parent, e * sizeof(*parent));
    memcpy(child, run->tables.edges.child, e * sizeof(*child));
}
