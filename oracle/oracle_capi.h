/* TEST INFRASTRUCTURE ONLY. C interface shared by the two CPU oracles:
 *
 *   oracle/_ref/libepik_ref_{dna,aa}.so   the reference's own epik/src/epik/place.cpp, compiled
 *                                         UNMODIFIED from /root/reference over the i2l shim (ref_capi.cpp)
 *   oracle/libepik_port.so                an independent plain restatement of the same path (port.cpp)
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * these libraries. Nothing under epik_b200/ links, includes or dlopens anything from oracle/.
 *
 * PARITY STATUS: the accumulate / threshold-default / top-k / LWR steps are pinned by the reference's
 * own place.cpp (and by the KAT-1 vectors of SURVEY.md Appendix D). The k-mer codec, the database
 * lookup and the load-time filter come from phylo42/i2l @ 43805ff, which is absent from
 * /root/reference: for those steps parity is UNPINNED.
 */
#ifndef EPIK_ORACLE_CAPI_H
#define EPIK_ORACLE_CAPI_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_entry
{
    uint32_t branch; /* post-order node id */
    float score;     /* log10 probability */
} orc_entry;

typedef struct orc_db_desc
{
    uint32_t alphabet;   /* 0 = DNA (2 bits/letter), 1 = amino acids (5 bits/letter) */
    uint32_t kmer_size;
    float omega;
    uint64_t n_keys;
    const uint32_t* keys;               /* [n_keys] k-mer codes, any order, unique */
    const uint64_t* offsets;            /* [n_keys + 1] into entries */
    const orc_entry* entries;           /* [offsets[n_keys]] */
    uint32_t n_nodes;
    const double* branch_length;        /* [n_nodes] by post-order id */
    const uint64_t* subtree_num_nodes;  /* [n_nodes] db.tree_index()[i].subtree_num_nodes */
    const double* subtree_total_length; /* [n_nodes] db.tree_index()[i].subtree_total_length */
} orc_db_desc;

/* Which alphabet this library handles: 0 DNA, 1 AA, 2 both (the port). */
int orc_alphabet(void);
/* "reference" for _ref, "port" for the restatement. */
const char* orc_kind(void);

/* Returns NULL on error (message via orc_last_error). */
void* orc_create(const orc_db_desc* desc, uint64_t keep_at_most, double keep_factor, uint64_t max_threads);
void orc_destroy(void* handle);
const char* orc_last_error(void);

/* Places n sequences given as packed bytes (sequence i = seq[seq_off[i] .. seq_off[i+1])).
 * The records are cut into batches of `batch_size` (main.cpp:214 default 2000) and each batch goes
 * through placer::place (identical sequences are grouped per batch, place.cpp:207).
 * Outputs are per INPUT record, row-major [n][keep_at_most]; rows are filled up to n_place[i].
 * group_rep[i] = index of the first record of the batch with the same sequence (its "nm" group).
 * seconds = wall time spent inside placer::place only (timed like main.cpp:343-345).
 * Any output pointer except n_place may be NULL. Returns 0 on success. */
int orc_place(void* handle, const char* seq, const uint64_t* seq_off, uint32_t n,
              uint32_t num_threads, uint32_t batch_size,
              uint32_t* n_place, uint32_t* branch, float* score, double* lwr, uint64_t* count,
              double* distal, double* pendant, uint32_t* group_rep, double* seconds);

/* The k-mer stream of one sequence, one record per (window, resolved key) in the order place.cpp
 * consumes them: pos = window start, key = code, group = number of keys of the window (1 = exact),
 * hit_len = posting-list length or -1 when the key is absent. Returns the number of records
 * (may exceed cap; only cap are written). */
uint64_t orc_kmers(void* handle, const char* seq, uint64_t len, uint64_t cap,
                   uint32_t* pos, uint32_t* key, uint32_t* group, int64_t* hit_len);

/* float score_threshold(omega, k) and its log10, as the placer holds them (place.cpp:87-88). */
void orc_thresholds(void* handle, float* threshold, float* log10_threshold);

#ifdef __cplusplus
}
#endif

#endif
