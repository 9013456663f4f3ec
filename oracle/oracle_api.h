/* TEST INFRASTRUCTURE (oracle).  Not part of the product path.
 *
 * One flat C interface implemented twice:
 *   - oracle/ref_driver.c  : thin driver around the UNMODIFIED reference
 *                            lib/mutgen.c (built into oracle/_ref/, only in the
 *                            container that has /root/reference);
 *   - oracle/mutgen_port.c : this project's CPU restatement of the same
 *                            algorithm (plain C, no tskit, travels as source).
 * Both draw from the GSL shim in oracle/gsl_shim (MT19937 or a replay stream).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
 * arm may load these libraries.
 */
#ifndef MSPB200_ORACLE_API_H
#define MSPB200_ORACLE_API_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_MODEL_MATRIX = 0, ORC_MODEL_INFINITE_ALLELES = 1, ORC_MODEL_SLIM = 2 };
enum { ORC_RNG_MT19937 = 0, ORC_RNG_REPLAY = 1 };

typedef struct {
    /* tree sequence (tskit column layout; ragged offsets are uint64, n+1 long) */
    double sequence_length;
    uint64_t num_nodes;
    const double *node_time;
    uint64_t num_edges;
    const double *edge_left;
    const double *edge_right;
    const int32_t *edge_parent;
    const int32_t *edge_child;
    /* existing sites and mutations (only read when keep != 0) */
    uint64_t num_sites;
    const double *site_position;
    const char *site_ancestral_state;
    const uint64_t *site_ancestral_state_offset;
    const char *site_metadata;
    const uint64_t *site_metadata_offset;
    uint64_t num_mutations;
    const int32_t *mutation_site;
    const int32_t *mutation_node;
    const double *mutation_time;
    const char *mutation_derived_state;
    const uint64_t *mutation_derived_state_offset;
    const char *mutation_metadata;
    const uint64_t *mutation_metadata_offset;
    /* rate map: map_size intervals, map_size + 1 positions */
    uint64_t map_size;
    const double *map_position;
    const double *map_rate;
    /* model */
    int32_t model_kind;
    uint64_t num_alleles;
    const char *alleles;
    const uint64_t *allele_offset;
    const double *root_distribution;
    const double *transition_matrix;
    uint64_t start_allele; /* infinite alleles: value of next_allele on entry */
    /* call parameters */
    int32_t discrete_genome;
    int32_t keep;
    double start_time;
    double end_time;
    /* random source */
    int32_t rng_kind;
    uint32_t seed;
    const double *stream;
    uint64_t stream_len;
    /* SLiM model (lib/mutgen.c:552-589): start_allele is next_mutation_id on entry */
    int32_t slim_mutation_type_id;
    int32_t slim_generation;
} orc_input_t;

typedef struct {
    int32_t status; /* 0 or a negative MSP_ERR_ code (lib/util.h:48-136 numbering) */
    int32_t stream_exhausted;
    uint64_t draws; /* uniforms consumed */
    uint64_t next_allele; /* infinite alleles / SLiM: allele or mutation-id counter after the call */
    double seconds; /* wall time of the generate step alone */
    uint64_t num_sites;
    double *site_position;
    char *site_ancestral_state;
    uint64_t *site_ancestral_state_offset;
    char *site_metadata;
    uint64_t *site_metadata_offset;
    uint64_t num_mutations;
    int32_t *mutation_site;
    int32_t *mutation_node;
    int32_t *mutation_parent;
    double *mutation_time;
    char *mutation_derived_state;
    uint64_t *mutation_derived_state_offset;
    char *mutation_metadata;
    uint64_t *mutation_metadata_offset;
} orc_output_t;

/* Runs one sim_mutations-equivalent call.  Fills *out (malloc'd columns);
 * release with orc_free_output.  Returns out->status. */
int orc_sim_mutations(const orc_input_t *in, orc_output_t *out);
void orc_free_output(orc_output_t *out);
const char *orc_name(void);

#ifdef __cplusplus
}
#endif
#endif
