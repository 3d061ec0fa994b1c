/* epik_b200 — C ABI of the B200 (sm_100a) placement hot path of EPIK.
 *
 * EPIK has no plugin/FFI interface of its own: the seam this library sits behind is the C++ class
 * epik::placer (reference epik/include/epik/place.h:81-140) and the i2l::phylo_kmer_db it reads
 * (call sites epik/src/epik/place.cpp:87,113-114,294,300,311; epik/src/epik/main.cpp:277-295).
 * Every entry point below names the reference interface it stands in for. The ABI is plain C:
 * pointers and sizes only, no C++ types, no exceptions, no torch types. Errors come back as an
 * epk_status plus a thread-local message (epk_last_error), the way main.cpp:384-388 turns a
 * std::runtime_error into "Error: ..." and a non-zero exit code.
 *
 * There is no CPU fallback: every call that computes needs a CUDA device and fails with
 * EPK_E_CUDA when there is none.
 */
#ifndef EPIK_B200_H
#define EPIK_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EPK_ABI_VERSION 1
#define EPK_IPC_HANDLE_BYTES 64 /* a CUDA IPC memory handle */

typedef enum epk_status
{
    EPK_OK = 0,
    EPK_E_INVALID = 1,      /* bad argument / malformed database description            */
    EPK_E_CUDA = 2,         /* CUDA runtime error or no device                          */
    EPK_E_NOMEM = 3,        /* host or device allocation failed                         */
    EPK_E_SHORT_SEQ = 4,    /* a sequence shorter than k-1 letters: the reference dies  */
                            /* on it (size_t wrap -> vector::reserve, place.cpp:291)    */
    EPK_E_UNSUPPORTED = 5   /* shape outside what the kernels were built for            */
} epk_status;

enum { EPK_DNA = 0, EPK_AA = 1 };

/* One posting: i2l::pkdb_value {branch, score} (place.cpp:358 structured binding;
 * main.cpp:257 uses its sizeof as the --max-ram accounting unit, 8 bytes). */
typedef struct epk_entry
{
    uint32_t branch; /* post-order node id */
    float score;     /* log10 probability  */
} epk_entry;

/* What i2l::phylo_kmer_db holds after i2l::load (main.cpp:277), flattened to CSR:
 * keys[i] owns entries[offsets[i] .. offsets[i+1]). All pointers are HOST memory and are only
 * read during epk_db_create. */
typedef struct epk_db_desc
{
    uint32_t alphabet;                  /* EPK_DNA (2 bits/letter) or EPK_AA (5 bits/letter)      */
    uint32_t kmer_size;                 /* db.kmer_size()                                          */
    float omega;                        /* db.omega()                                              */
    uint64_t n_keys;
    const uint32_t* keys;               /* [n_keys] unique k-mer codes (phylo_kmer::key_type)      */
    const uint64_t* offsets;            /* [n_keys + 1]                                            */
    const epk_entry* entries;           /* [offsets[n_keys]]; branches distinct within one list    */
    const float* filter_value;          /* [n_keys] or NULL; larger = more informative. Needed     */
                                        /* only when a load filter with mu < 1 / max_entries is    */
                                        /* given (the order i2l stores k-mers in on disk)          */
    uint32_t n_nodes;                   /* tree.get_node_count() (place.cpp:99,166)                */
    const double* branch_length;        /* [n_nodes] get_by_postorder_id(i)->get_branch_length()   */
    const uint64_t* subtree_num_nodes;  /* [n_nodes] db.tree_index()[i] (place.cpp:113)            */
    const double* subtree_total_length; /* [n_nodes] db.tree_index()[i] (place.cpp:114)            */
} epk_db_desc;

/* The three filter arguments of i2l::load(file, mu, omega, max_entries) (main.cpp:277; --mu,
 * --omega, --max-ram of epik.py:40-55 / main.cpp:215-218,252-265). */
typedef struct epk_load_filter
{
    double mu;            /* keep the mu most informative fraction of the entries, in [0, 1]      */
    float omega;          /* user omega: entries scoring below log10((omega/|alphabet|)^k) go    */
    uint64_t max_entries; /* --max-ram / sizeof(pkdb_value); UINT64_MAX = unlimited              */
} epk_load_filter;

typedef struct epk_db_info
{
    uint32_t alphabet, kmer_size, n_nodes, table_kind; /* table_kind: 0 direct-address, 1 hash    */
    float omega;                 /* omega in force after the load filter (db.omega())             */
    float threshold;             /* i2l::score_threshold(omega, k)        (place.cpp:87)          */
    float log10_threshold;       /* std::log10(threshold)                 (place.cpp:88)          */
    uint64_t n_keys_loaded;
    uint64_t n_entries_loaded;   /* db.get_num_entries_loaded()           (main.cpp:290)          */
    uint64_t n_entries_total;    /* db.get_num_entries_total()            (main.cpp:291)          */
    uint64_t table_slots;        /* slots of the HBM key table                                    */
    uint64_t device_bytes;       /* HBM held by this database                                     */
} epk_db_info;

typedef struct epk_db epk_db;
typedef struct epk_placer epk_placer;

/* Options of a placer beyond the reference's constructor arguments. Zero-initialise for defaults. */
typedef struct epk_placer_opts
{
    uint32_t variant;        /* 0 = default: 16 warps per SM on small trees, one warp per query   */
                             /* at C2's size, teams of warps per query on large trees;            */
                             /* 1 = always one warp per query, 8 per SM                           */
    uint32_t chunk_reads;    /* reads per pipelined H2D/kernel/D2H chunk of epk_place; 0 = auto   */
    uint32_t reserved[6];
} epk_placer_opts;

/* Results of placing n sequences, row-major [n][keep_at_most]; row i is filled up to n_place[i],
 * best placement first (place.cpp:153-156 order). These are the fields of epik::placement
 * (place.h:45-56). distal/pendant may be NULL (they are table look-ups by branch,
 * place.cpp:98-125,434-437). For epk_place all pointers are HOST memory (pinned is faster:
 * epk_host_alloc); for epk_place_device they are DEVICE memory and distal/pendant must be NULL. */
typedef struct epk_result
{
    uint32_t* n_place; /* [n]                                                                      */
    uint32_t* branch;  /* [n*K] placement.branch_id                                                */
    float* score;      /* [n*K] placement.score        (jplace "likelihood")                       */
    double* lwr;       /* [n*K] placement.weight_ratio (jplace "like_weight_ratio")                */
    uint32_t* count;   /* [n*K] placement.count                                                    */
    double* distal;    /* [n*K] or NULL                                                            */
    double* pendant;   /* [n*K] or NULL                                                            */
} epk_result;

/* ABI version of the loaded library (EPK_ABI_VERSION it was built with). */
uint32_t epk_abi_version(void);

/* Message of the last failing call on this thread ("" if none). */
const char* epk_last_error(void);

/* Number of CUDA devices visible (0 when there is none or the driver is missing). */
int epk_device_count(void);

/* Stands in for i2l::load(db_file, mu, omega, max_entries) (main.cpp:277) once the file has been
 * parsed: applies the load filter (NULL = keep everything) ON THE DEVICE, flattens the map into an
 * HBM-resident key table (direct-address when alphabet_bits*k <= 26, open addressing otherwise)
 * plus 16-byte aligned (branch, score) posting lists, and precomputes the distal / pendant tables
 * of placer::placer (place.cpp:98-125). `device` is the CUDA ordinal. */
epk_status epk_db_create(const epk_db_desc* desc, const epk_load_filter* filter, int device, epk_db** out);
epk_status epk_db_get_info(const epk_db* db, epk_db_info* info);
void epk_db_destroy(epk_db* db);

/* epik::placer::placer(db, tree, keep_at_most, keep_factor, num_threads) (place.h:94-95,
 * place.cpp:83-126). The tree comes with the database description; there are no host threads to
 * size scratch for. The database must outlive the placer (place.h:91-93). */
epk_status epk_placer_create(const epk_db* db, uint64_t keep_at_most, double keep_factor,
                             const epk_placer_opts* opts, epk_placer** out);
void epk_placer_destroy(epk_placer* placer);

/* epik::placer::place(seq_records, num_threads) (place.h:103, place.cpp:201-275) for n sequences
 * given as packed bytes: sequence i = seq[seq_off[i] .. seq_off[i+1]). HOST buffers; the call
 * streams them to the device in chunks (pinned staging, copies overlapped with the kernels) and
 * returns when `out` is complete. Results are per INPUT record: identical sequences get identical
 * rows, which is what the reference's per-batch grouping (place.cpp:207) amounts to; use
 * epk_group_batch for the grouping itself. */
epk_status epk_place(epk_placer* placer, const char* seq, const uint64_t* seq_off, uint32_t n,
                     const epk_result* out);

/* epk_place over several GPUs from ONE host process (the reference is one process, one FASTA stream, one jplace:
 * main.cpp:332-366): placers[g] sits on its own device over its own replica of the database; the pass is cut into
 * n_placers contiguous slices, each placed by one placer on one host thread, and the rows land in `out` in input
 * order - exactly what epk_place on one device leaves there. The placers must share keep_at_most / keep_factor.
 * Returns the most serious status of the slices (EPK_E_SHORT_SEQ / EPK_E_UNSUPPORTED leave every other row complete). */
epk_status epk_place_many(epk_placer* const* placers, uint32_t n_placers, const char* seq, const uint64_t* seq_off,
                          uint32_t n, const epk_result* out);

/* Same work with everything already resident in HBM: DEVICE pointers, asynchronous on `stream`
 * (a cudaStream_t, NULL = the legacy default stream). A placer is a single-stream object: its work cursor
 * and counters are shared by all its launches, so calls on one placer (this one, the partial / finalize /
 * ceiling calls below, epk_place) must be ordered - the same stream, or a synchronisation in between;
 * use one placer per concurrent stream (placers share the database). */
epk_status epk_place_device(epk_placer* placer, const char* d_seq, const uint64_t* d_seq_off, uint32_t n,
                            const epk_result* d_out, void* stream);

/* epk_place_device for callers that cannot have memory allocated under them (CUDA graph capture, tight serving loops,
 * harnesses that forbid it). NO COUNTERPART IN THE REFERENCE. The call
 * allocates nothing - no cudaMalloc / cudaFree, no allocating library routine: it runs the accumulators-in-HBM form
 * of the placement kernel with its scratch (launch counters + one accumulator row per warp) in a static device array of
 * EPK_NOALLOC_SCRATCH_BYTES that every GPU which loads the library holds until exit. Same arguments and results as
 * epk_place_device (DEVICE pointers, asynchronous on `stream`); scores, branches and counts are the same to the bit,
 * weight ratios to 1e-9 (the sweep adds the likelihood mass in another order). Limits and rules:
 *   - trees of more than EPK_NOALLOC_MAX_NODES nodes are refused with EPK_E_UNSUPPORTED;
 *   - every call on a device shares the one array; the call clears what it uses on `stream` first, so calls in
 *     sequence are independent, but calls that OVERLAP on a GPU (two streams, two placers, two threads) race;
 *   - the work counters of the placer are not updated (epk_placer_counters does not see this call); rows of
 *     sequences shorter than k - 1 letters come back with n_place = 0, no status reports them;
 *   - the first call on a device resolves the array's device address (once per device; make it before capturing);
 *   - several times slower per read than epk_place_device where that one keeps its accumulators in shared memory. */
#define EPK_NOALLOC_SCRATCH_BYTES (32u << 20)
#define EPK_NOALLOC_MAX_NODES 65536u
epk_status epk_place_device_noalloc(epk_placer* placer, const char* d_seq, const uint64_t* d_seq_off, uint32_t n,
                                    const epk_result* d_out, void* stream);

/* group_by_sequence_content (place.cpp:73-81) for one batch: group_rep[i] = index of the first
 * record of [0, n) with the same sequence. HOST buffers. */
epk_status epk_group_batch(const char* seq, const uint64_t* seq_off, uint32_t n, uint32_t* group_rep);

/* The same for a whole pass of n records cut into batches of batch_size (main.cpp:332-344: sequences merge only
 * within one --batch-size batch): group_rep[i] = index WITHIN ITS BATCH of the first record of that batch with the same
 * sequence; the batches are spread over n_threads host threads. */
epk_status epk_group_batches(const char* seq, const uint64_t* seq_off, uint32_t n, uint32_t batch_size, uint32_t n_threads,
                             uint32_t* group_rep);

/* The k-mer stream of one sequence in the order query_kmers (place.cpp:278-316) consumes it, one
 * record per (window, resolved key): pos = window start, key = code, group = number of keys of
 * the window (1 = unambiguous), hit_len = posting-list length or -1 when the key is absent.
 * Computed ON THE DEVICE by a one-thread kernel through classify_window / db_search (the placement kernel's
 * ambiguous-window path uses the same two; its exact k-mers come from a rolling per-lane encoder, whose
 * output is exposed by epk_placer_set_trace below). Returns the
 * number of records through *n_out (only `cap` are written). HOST buffers. */
epk_status epk_kmers(const epk_db* db, const char* seq, uint64_t len, uint64_t cap,
                     uint32_t* pos, uint32_t* key, uint32_t* group, int64_t* hit_len, uint64_t* n_out);

/* The same evidence from the placement kernel itself: while d_trace (DEVICE memory, 4 uint32 per query of the
 * largest batch to come; NULL switches it off) is set, every epk_place_device call leaves per query
 *   { keys looked up, keys found, postings of the keys found, checksum }
 * where the keys are those of query_kmers (place.cpp:294-313): one per exact k-mer, one per resolution of a
 * window with a single ambiguous letter; checksum = sum modulo 2^32 over them of epk_trace_term(window
 * position, code, keys of the window, list length or 0). These are the rolling encoder and the probes of the
 * hot kernel, not a second implementation: a dropped, duplicated or miscoded window changes the record.
 * Queries that are not placed (too short, unsupported) leave zeros. */
epk_status epk_placer_set_trace(epk_placer* placer, uint32_t* d_trace);
uint32_t epk_trace_term(uint32_t pos, uint32_t key, uint32_t group, uint32_t len);

/* Work counters of the last epk_place / epk_place_device call on this placer (for roofline
 * arithmetic): windows looked up, hits, posting entries gathered, kernels launched. */
typedef struct epk_counters
{
    uint64_t n_reads, n_windows, n_hits, n_entries, n_launches;
    uint64_t n_short;       /* sequences shorter than k - 1 letters (left with n_place = 0)            */
    uint64_t n_unsupported; /* sequences left with n_place = 0 because a launch form does not exist for */
                            /* them: only in database-sharded mode (more than 65535 windows, a tree too */
                            /* large for shared memory, ambiguous windows without shard views). The      */
                            /* fused path places everything: accumulators that do not fit shared memory  */
                            /* live in HBM (slower, same results)                                        */
} epk_counters;
epk_status epk_placer_counters(epk_placer* placer, epk_counters* out);

/* ---- database-sharded mode (used when a database exceeds one GPU's HBM) -----------
 *
 * The reference has one phylo_kmer_db in one address space (main.cpp:277); when it does not fit one
 * HBM the keys are split over the G GPUs of a group by a hash range. Every GPU runs the encode /
 * probe / accumulate phases (place.cpp:278-316,349-371) of EVERY query against its shard and leaves a
 * COMPACT partial row per query  { u32 flags, u32 postings, u32 touched, 0 | bitmap of the touched branches |
 * touched branches before each 128-branch step | f32 S of the touched branches | u16 C of the touched branches }
 * in a slot of epk_partial_row_bytes (room for every branch; only the used prefix of each part is written and
 * read, so the bytes exchanged scale with the branches a shard touched) in the memory of the GPU that owns the
 * query; the owner adds the G rows in shard order and finishes the query (place.cpp:417-422, 164-199, 238-267). Transport between the two calls
 * is the caller's: an NCCL all-to-all of the row buffers, or row buffers that ARE the owners' memory
 * (CUDA IPC / peer pointers over NVLink), in which case the partial launch is the exchange.
 * Scores agree with the reference within float rounding of a re-associated sum (1e-5 relative), not
 * bit for bit. Queries with more than 65535 windows are reported through epk_counters.n_unsupported,
 * n_place = 0; so are queries with ambiguous windows (place.cpp:373-415 needs the global key order) unless
 * the owner has been given the views of all the shards - see "windows with one ambiguous letter" below. */

/* epk_db_create keeping only the keys of shard `shard_index` of `shard_count`; the load filter's
 * informative prefix (mu / max_entries) is still taken over the whole description. */
epk_status epk_db_create_shard(const epk_db_desc* desc, const epk_load_filter* filter, int device,
                               uint32_t shard_index, uint32_t shard_count, epk_db** out);
/* The shard a k-mer code belongs to. */
uint32_t epk_shard_of_key(uint32_t key, uint32_t shard_count);

/* Bytes of one partial-row slot for this placer's tree (Np = n_nodes rounded up to 128): 16 + Np / 8 + Np / 32 + 6 * Np,
 * rounded up to 16. */
uint64_t epk_partial_row_bytes(const epk_placer* placer);

/* Phases 1-3 of placer::place_seq for n queries against this placer's (shard) database. Query i with
 * owner_first[g] <= i < owner_first[g+1] leaves its row at d_rows[g] + (i - owner_first[g]) * row_bytes.
 * owner_first [n_owners + 1] and d_rows [n_owners] are HOST arrays; d_rows[g] are DEVICE pointers
 * (16-byte aligned; local or peer memory). d_seq / d_seq_off are DEVICE memory. Asynchronous on `stream`. */
epk_status epk_place_partial_device(epk_placer* placer, const char* d_seq, const uint64_t* d_seq_off, uint32_t n,
                                    uint32_t n_owners, const uint32_t* owner_first, void* const* d_rows,
                                    void* stream);

/* The owner's side: row i of shard g is at d_rows + g * shard_stride_bytes + i * row_bytes. Adds the
 * n_shards rows of each of the n queries and writes their placements to d_out (DEVICE memory, rows
 * [0, n)). d_seq_off[0 .. n] are the offsets of these n queries (only their differences are used). */
epk_status epk_place_finalize_device(epk_placer* placer, const void* d_rows, uint32_t n_shards,
                                     uint64_t shard_stride_bytes, const uint64_t* d_seq_off, uint32_t n,
                                     const epk_result* d_out, void* stream);

/* ---- pooled database: the shards read in place over NVLink ---------------------------------------------
 *
 * The other way to hold a database that exceeds one HBM, for ONE process that drives the GPUs (as the reference is
 * one process, main.cpp:332-366): shards[g] = shard g of n_shards (epk_db_create_shard / epk_db_create_synthetic,
 * one per GPU; several on one GPU work too) are mapped one after the other into one range of virtual addresses that
 * every GPU of the process reads (CUDA virtual memory management; the GPUs must be peers), and out[g] is a database
 * on shards[g]'s device whose key table points into that range. To a placer out[g] is a WHOLE database: every entry
 * point above takes it, the fused kernel gathers each posting list from whichever HBM holds it (1 / n_shards local,
 * the rest over NVLink), and the results are those of the unsharded database to the bit. Place with one placer per
 * out[g] and epk_place_many. The shards are copied: destroy them once the pool exists (until then each GPU holds its
 * shard twice). Needs the direct-address key table (alphabet bits x k <= 26) and at most 2^33 postings in total.
 * The reference has nothing like it (one phylo_kmer_db in one address space, main.cpp:277). */
epk_status epk_db_pool(epk_db* const* shards, uint32_t n_shards, epk_db** out /* [n_shards] */);

/* The same pool with one PROCESS per GPU (torchrun, MPI): a shard's memory travels as a POSIX file descriptor, its key
 * table as a CUDA IPC handle, and every process builds the common address range for itself.
 *   1. epk_db_pool_share(shard, &fd, &bytes, table_handle): the shard's postings are copied into exportable memory;
 *      send {fd (over a unix socket, SCM_RIGHTS), bytes, table_handle [EPK_IPC_HANDLE_BYTES]} to every peer;
 *   2. epk_db_pool_join(shard, n_shards, fds, bytes, table_handles, &pooled): fds / bytes / table_handles are indexed by
 *      shard (the entry of the caller's own shard index is ignored for fds and table_handles); `pooled` is the whole
 *      database on the caller's device. Keep every shard alive until all the processes have joined (a barrier), then
 *      close the descriptors and destroy the shards;
 *   3. the processes agree on the score range of all the shards (epk_db_get_score_range of each shard: min of lo, max
 *      of hi, AND of fast_ok, sums of the counts) and each sets it on its pooled database
 *      (epk_db_pool_set_score_range) before placing: it decides which arithmetic the sweep may use.
 * Works with several processes on one GPU too (tests). */
epk_status epk_db_pool_share(epk_db* shard, int* fd_out, uint64_t* bytes_out, unsigned char* table_handle_out);
epk_status epk_db_pool_join(epk_db* shard, uint32_t n_shards, const int* fds, const uint64_t* bytes,
                            const unsigned char* table_handles /* [n_shards][EPK_IPC_HANDLE_BYTES] */, epk_db** out);
epk_status epk_db_get_score_range(const epk_db* db, float* lo, float* hi, uint32_t* fast_ok, uint64_t* entries_loaded,
                                  uint64_t* keys_loaded);
epk_status epk_db_pool_set_score_range(epk_db* pooled, float lo, float hi, uint32_t fast_ok, uint64_t entries_loaded,
                                       uint64_t keys_loaded);

/* ---- windows with one ambiguous letter in database-sharded mode -----------------------------------------
 *
 * place.cpp:373-415 lets the FIRST ambiguous key (window order, then resolution order) that names a branch
 * decide that branch's contribution - an order over all keys, which no shard has. The owner of a query has
 * it: with the views of all the shards set (epk_placer_set_shard_views) and the letters passed along
 * (epk_place_finalize_seq_device), the finalising launch replays the ambiguous windows of its flagged queries
 * itself, looking every resolved key up in the one shard that holds it (its own HBM, or a peer's key table
 * and postings mapped over NVLink with epk_ipc_export / epk_ipc_import on the addresses of the view), and adds
 * the contributions to the query's row of shard 0 before the rows are summed and swept. Without views such
 * queries stay unplaced (n_place = 0, counted in n_unsupported) as before. */
typedef struct epk_db_view
{
    uint64_t table;       /* device address of the key table (base of its allocation)              */
    uint64_t postings;    /* device address of the postings  (base of its allocation)              */
    uint64_t hash_mask;   /* open addressing: slots - 1                                            */
    uint32_t hash_shift;
    uint32_t table_kind;  /* 0 direct-address, 1 open addressing                                   */
} epk_db_view;

/* The addresses of this database on its own device. */
epk_status epk_db_get_view(const epk_db* db, epk_db_view* out);

/* views[g] = shard g of the group (this placer's own shard included), with addresses valid on THIS placer's
 * device; n_shards as passed to epk_place_finalize_*; n_shards = 0 forgets them. The views are copied. */
epk_status epk_placer_set_shard_views(epk_placer* placer, uint32_t n_shards, const epk_db_view* views);

/* epk_place_finalize_device with the letters of the n queries: d_seq + d_seq_off[i] is query i (the same
 * absolute offsets and base the partial launches were given). d_seq = NULL behaves like
 * epk_place_finalize_device. The rows of flagged queries are modified in place. */
epk_status epk_place_finalize_seq_device(epk_placer* placer, void* d_rows, uint32_t n_shards,
                                         uint64_t shard_stride_bytes, const char* d_seq, const uint64_t* d_seq_off,
                                         uint32_t n, const epk_result* d_out, void* stream);

/* Peer memory for the fused exchange: with one process per GPU, each owner allocates its slab with
 * epk_device_alloc, exports it (64-byte CUDA IPC handle, sent to the peers by any means), the peers
 * import it and pass the mapped address as d_rows[owner] of epk_place_partial_device: the partial
 * launch then stores its rows straight into the owner's HBM over NVLink. */
epk_status epk_device_alloc(int device, uint64_t bytes, void** out);
void epk_device_free(int device, void* p);
epk_status epk_ipc_export(const void* d_ptr, unsigned char* handle_out /* [EPK_IPC_HANDLE_BYTES] */);
epk_status epk_ipc_import(int device, const unsigned char* handle, void** out);
void epk_ipc_close(int device, void* p);

/* Measuring stick for the roofline ("measured random-gather ceiling"): runs the same
 * encode / probe / posting-gather instruction stream as epk_place_device for the queries with up to
 * 255 windows, with the accumulator updates replaced by an XOR into registers and no sweep, so that
 * its duration is what the memory system allows for this access pattern. d_checksum [n] (DEVICE)
 * receives one word per query; epk_placer_counters reports the postings gathered. */
epk_status epk_gather_ceiling_device(epk_placer* placer, const char* d_seq, const uint64_t* d_seq_off, uint32_t n,
                                     uint32_t* d_checksum, void* stream);

/* A second, independent measuring stick: n_probes uniformly random k-mer codes looked up in this database's
 * key table, every hit's posting list gathered, by a kernel with no accumulators, no shared-memory budget and
 * full occupancy (method 0 / 1: 128-bit loads, 4 / 8 lists in flight per warp, 32 warps per SM; method 2: bulk
 * asynchronous copies of whole lists into shared memory, completion on an mbarrier). What the memory system
 * gives the path's access pattern, rather than what the placement kernel's design asks of it. Synchronous;
 * `reps` timed launches after one warm-up. bytes = 16 * n_probes + 8 * n_entries (SURVEY.md 8d's accounting). */
typedef struct epk_microbench
{
    double ms_per_launch;
    uint64_t n_probes, n_hits, n_entries, bytes, checksum;
} epk_microbench;
epk_status epk_gather_microbench(const epk_db* db, uint64_t n_probes, uint64_t seed, uint32_t method,
                                 uint32_t reps, epk_microbench* out);

/* Test / bench infrastructure: a synthetic database generated ON THE DEVICE, a pure function of (seed, key, position
 * in the list) that epik_b200/synth.py restates in numpy for any subset of keys (the reference is fed exactly the
 * lists a set of reads looks up). BASELINE config 5 - k = 12, ~20 k branches, ~4 G postings = 32 GB - exists only
 * this way: numpy would need minutes and > 100 GB of host memory for it. The entries are generated one key range
 * at a time into the chunk buffer the build streams every database through (raw and packed never coexist in HBM).
 *   key present    iff  stream(3, key) < presence_threshold
 *   list length     =   length_table[stream(4, key) >> (64 - length_table_bits)]   (each in [1, n_nodes / 8])
 *   branches        =   (stream(7, key) % n_nodes + running sum of strides stream(6, key << 24 ^ j) % 7 + 1) % n_nodes
 *   scores          =   min(float(score_lo + (-0.3 - score_lo) * unit(stream(8, key << 24 ^ j))), -0.3f)
 * (stream / unit: epik_b200/csrc/synth_gen.cuh). Keys are the codes 0 .. |alphabet|^k (or key_space_limit). */
typedef struct epk_synth_desc
{
    uint32_t alphabet, kmer_size;
    float omega;
    uint32_t n_nodes;
    const double* branch_length;        /* [n_nodes] HOST, as in epk_db_desc */
    const uint64_t* subtree_num_nodes;
    const double* subtree_total_length;
    uint64_t seed;
    uint64_t presence_threshold;
    const uint32_t* length_table;       /* [1 << length_table_bits] HOST */
    uint32_t length_table_bits;
    double score_lo;
    uint64_t key_space_limit;           /* 0 = the whole code space */
} epk_synth_desc;
epk_status epk_db_create_synthetic(const epk_synth_desc* desc, const epk_load_filter* filter, int device,
                                   uint32_t shard_index, uint32_t shard_count, epk_db** out);

/* Pinned host memory for the buffers of epk_place. */
void* epk_host_alloc(size_t bytes);
void epk_host_free(void* p);

#ifdef __cplusplus
}
#endif

#endif /* EPIK_B200_H */
