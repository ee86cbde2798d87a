/*
 * lzani_b200.h -- C ABI of the B200-native LZ-ANI pair-comparison hot path.
 *
 * This is the drop-in boundary for the ONE path this repository accelerates: what
 * CLZMatcher::do_matching (reference src/lz_matcher.cpp:172-277) does with its private
 * CParser objects (reference src/parser.h:243-253):
 *
 *     CParser parser(params);                       -> lzani_create()
 *     parser.prepare_reference(seq_view(ref), ..)   \
 *     parser.prepare_data(seq_view(qry), ..)         |  lzani_set_sequences() once, then
 *     parser.parse()                                 |-> lzani_run() / lzani_run_rows()
 *     results_t r = parser.calc_stats()             /     (batch form: one call, all pairs)
 *     vector<region_t> v = parser.get_parsing()     -> lzani_run_rows_regions() (--out-alignment)
 * and, past the parse, the pairing + --out-filter arithmetic of store_results
 * (lz_matcher.cpp:404-463)                          -> lzani_filter_pairs()
 *
 * Plain pointers and sizes only; no C++/torch types.  All work is done by hand-written
 * sm_100a CUDA kernels; there is NO CPU fallback: every entry point fails with
 * LZANI_ERR_CUDA when no usable device is present.
 *
 * Ids are indices into the array handed to lzani_set_sequences() (the reference uses the
 * post-reorder reservoir index, lz_matcher.cpp:78-85).  "reference r / query q" follow the
 * reference's direction convention: out[(r,q)] is the parse of q against the
 * index of r.
 */
#ifndef LZANI_B200_H
#define LZANI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LZANI_ABI_VERSION 1

/* The eight parse parameters (reference src/params.h:38-45, defaults 11 7 40 40 35 15 7 3). */
typedef struct lzani_params_t {
    int32_t mal; /* min_anchor_len     --mal */
    int32_t msl; /* min_seed_len       --msl */
    int32_t mrd; /* max_dist_in_ref    --mrd */
    int32_t mqd; /* max_dist_in_query  --mqd */
    int32_t reg; /* min_region_len     --reg */
    int32_t aw;  /* approx_window      --aw  */
    int32_t am;  /* approx_mismatches  --am  */
    int32_t ar;  /* approx_run_len     --ar  */
} lzani_params_t;

/* == results_t (reference src/defs.h:48-65): nt_match, nt_mismatch, num_alns. */
typedef struct lzani_result_t {
    int32_t sym_in_matches;
    int32_t sym_in_literals;
    int32_t no_components;
} lzani_result_t;

/* == region_t (reference src/defs.h:67-153), as returned by CParser::get_parsing(). */
typedef struct lzani_region_t {
    int32_t ref_start, ref_end;
    int32_t seq_start, seq_end;
    int32_t num_matches, num_mismatches;
} lzani_region_t;

/* Thresholds of --out-filter (params.h:98-99, set_output_filter :200-209) in the order the
 * reference tests them (lz_matcher.cpp:451-462); 0 = not set. */
typedef struct lzani_out_filter_t {
    double gani, ani, tani, qcov, rcov;
} lzani_out_filter_t;

/* One unordered pair (a < b) that survives the filter: both directions' results_t, as
 * store_results pairs them up (lz_matcher.cpp:404-418).  ab = parse of b against the index of a
 * (`q->results`), ba = parse of a against the index of b (`p->results`).  lines bit 0: the output
 * line "reference a, query b" passes, bit 1: the line "reference b, query a" passes. */
typedef struct lzani_pair_t {
    uint32_t a, b;
    lzani_result_t ab, ba;
    uint32_t lines;
} lzani_pair_t;

/* Device timings of the last lzani_run*() call (CUDA events on the library's stream). */
typedef struct lzani_timing_t {
    double h2d_ms;      /* pair list (+ sequences if not yet resident) host->device */
    double pack_ms;     /* K1: base-6 triples -> bit-plane records               */
    double index_ms;    /* K2: anchor presence bitmap + anchor table build        */
    double parse_ms;    /* K3: the parse stage (staged kernels + warp-per-pair walk) */
    double d2h_ms;      /* results device->host                                   */
    double total_ms;    /* first event to last event                              */
    uint64_t n_pairs;   /* ordered pairs processed                                */
    uint64_t q_bases;   /* sum of query lengths over those pairs                  */
    uint64_t r_bases;   /* sum over indexed references of Dr = 2*Lr + 3*mrd       */
    uint64_t h2d_bytes; /* bytes copied host->device inside the call              */
    uint64_t d2h_bytes; /* bytes copied device->host inside the call              */
    uint32_t launches;  /* kernels launched inside the call                       */
    uint32_t n_batches; /* reference batches the index memory budget forced       */
    double filter_ms;   /* lzani_filter_pairs: tables H2D, F1-F3, surviving pairs D2H */
    uint64_t filter_d2h_bytes; /* bytes lzani_filter_pairs copied device->host    */
    double pairlist_ms; /* lzani_build_pair_list: edges H2D, expand, sort, CSR, D2H  */
} lzani_timing_t;

enum {
    LZANI_OK = 0,
    LZANI_ERR_ARG = 1,    /* bad argument / unsupported parameter combination */
    LZANI_ERR_CUDA = 2,   /* CUDA runtime error or no device                  */
    LZANI_ERR_STATE = 3,  /* call order (e.g. run before set_sequences)       */
    LZANI_ERR_NOMEM = 4   /* host or device allocation failed                 */
};

typedef struct lzani_ctx lzani_ctx;

/* Replaces `CParser parser(params)` and the worker pool around it (parser.h:237-241,
 * lz_matcher.cpp:184-200).  device_ids/n_devices: the CUDA devices this context drives.
 * n_devices == 1 (device_ids==NULL: the current device): one context per GPU, e.g. one per
 * process under torchrun.  n_devices > 1: ONE context that owns all the listed GPUs of the
 * box -- the library runs one worker thread per device, replicates the sequences, deals
 * the reference rows of every lzani_run* call over the devices (dense lists: row r to
 * device r mod n; sparse lists: heaviest rows first) and writes the results back in the
 * caller's CSR order; lzani_filter_pairs() on such a context reads the other devices' rows
 * over NVLink peer access.  Returns LZANI_OK or an error code; on error *ctx is still a
 * valid object holding the message when allocation was possible. */
int lzani_create(lzani_ctx** ctx, const lzani_params_t* params, const int* device_ids, int n_devices);

/* Replaces the seq_view hand-off of prepare_reference/prepare_data (parser.cpp:16-50).
 * data[i] = sequence i in the reservoir's default `three_in_byte` packing
 * (seq_reservoir.cpp:39-57: byte = 36a+6b+c, symbols 0..3 = ACGT, 4/5 = N, last partial
 * byte high-aligned), len[i] = number of bases.  The bytes are copied; the caller may free
 * them after the call returns.  May be called again to replace the set. */
int lzani_set_sequences(lzani_ctx* ctx, uint32_t n, const uint8_t* const* data, const uint32_t* len);

/* Device-side ingest: the same hand-off one step earlier in the loader.  text[i] = the
 * symbols of sequence i exactly as CSeqReservoir::append receives them (seq_reservoir.cpp:18-87): one
 * byte per base, FASTA letters in either case, anything that is not ACGTacgt -- N, IUPAC codes, the
 * byte-5 contig separators load_fasta inserts (:110-127) -- is N (dna_code, seq_reservoir.h:243-247).
 * The device builds the bit-plane records and the reverse complement straight from the text; the
 * host never packs base-6 triples.  Results are identical to lzani_set_sequences on append's output. */
int lzani_set_sequences_text(lzani_ctx* ctx, uint32_t n, const uint8_t* const* text, const uint32_t* len);

/* Replaces the whole worker loop of do_matching (lz_matcher.cpp:192-269).
 * CSR pair list: for reference r in [0, n_refs) the queries are
 * q_ids[row_off[r] .. row_off[r+1]); n_refs must equal the number of sequences.
 * q_ids == NULL means all2all: every q != r in increasing q (row_off ignored, may be NULL).
 * out[k] receives the result of the k-th (r, q) in CSR order.  Blocking. */
int lzani_run(lzani_ctx* ctx, uint32_t n_refs, const uint64_t* row_off, const uint32_t* q_ids,
              lzani_result_t* out);

/* Same, for a subset of reference rows (the multi-GPU shard of one rank):
 * row j is reference ref_ids[j]; q_ids==NULL again means every q != ref_ids[j]. */
int lzani_run_rows(lzani_ctx* ctx, uint32_t n_rows, const uint32_t* ref_ids, const uint64_t* row_off,
                   const uint32_t* q_ids, lzani_result_t* out);

/* Replaces `parser.get_parsing()` (parser.h:253, calc_regions parser.cpp:786-837) as used for
 * --out-alignment (lz_matcher.cpp:226-227): the same rows as lzani_run_rows() and, besides out[],
 * the kept regions of every pair in CSR form: the regions of the k-th pair are
 * regions[region_off[k] .. region_off[k+1]), sorted the way calc_regions sorts them (length
 * descending, then seq_start).  region_off has n_pairs + 1 entries.  *n_regions receives the total;
 * when it exceeds regions_cap nothing is written to regions[] / region_off[] and LZANI_ERR_NOMEM is
 * returned: call again with a buffer of at least *n_regions entries.  Reference coordinates are
 * those of seq_ref (forward | 2*mrd N | reverse complement, parser.cpp:16-34), as in the reference. */
int lzani_run_rows_regions(lzani_ctx* ctx, uint32_t n_rows, const uint32_t* ref_ids, const uint64_t* row_off,
                           const uint32_t* q_ids, lzani_result_t* out, uint64_t* region_off,
                           lzani_region_t* regions, uint64_t regions_cap, uint64_t* n_regions);

/* Device-resident variant used to time the kernels alone: runs the same rows `repeat`
 * times back to back without copying results to the host; the last pass's results are
 * left on the device and copied to `out` only if out != NULL. */
int lzani_run_rows_resident(lzani_ctx* ctx, uint32_t n_rows, const uint32_t* ref_ids,
                            const uint64_t* row_off, const uint32_t* q_ids, int repeat,
                            lzani_result_t* out);

/* Replaces the pairing, derived-metric and --out-filter arithmetic of store_results
 * (lz_matcher.cpp:404-463) as a device-side pre-filter.  Call after
 * lzani_run_rows_resident(ctx, ..., repeat 1, out NULL) -- or any lzani_run_rows* -- whose results are
 * still on the device.  For every ordered pair (a, b), a < b, of ctx's rows the kernel finds the
 * opposite direction (row b, entry a), computes gani/ani/tani/qcov/rcov of both output lines in IEEE
 * double exactly as the reference does, and keeps the pair when at least one line passes.  Only the
 * surviving pairs are copied to pairs[] (CSR order of ctx's rows; *n_out = their number; when it
 * exceeds cap nothing is written and LZANI_ERR_NOMEM is returned).
 * group / n_group: all contexts of this process whose rows together hold both directions (one per
 * GPU, ctx included; NULL = ctx alone).  Rows owned by other GPUs are read in place over NVLink peer
 * access -- no host round trip, no collective; none of the group's contexts may start a new run until
 * every lzani_filter_pairs call of the group has returned.  eff_len[i] = length of sequence i without
 * its contig separators (`len - (no_parts-1)*mrd`, lz_matcher.cpp:437).  Query ids must ascend within
 * every CSR row; a missing opposite direction is an error (the reference asserts, :417). */
int lzani_filter_pairs(lzani_ctx* ctx, lzani_ctx* const* group, int n_group, const uint32_t* eff_len,
                       const lzani_out_filter_t* flt, lzani_pair_t* pairs, uint64_t cap, uint64_t* n_out);

/* Replaces the symmetric closure of CFilter::load_filter (filter.cpp:243-280) and CFilter::reorder_items
 * (:293-345) -- the construction of the sparse pair list from the edges of a kmer-db filter file
 * (SURVEY 8f-4).  edge_i[e], edge_j[e]: 0-based ids, in the file's own sequence order, of the e-th
 * entry that passed the threshold (row i of the file lists `j+1:value`, filter.cpp:64-84); every edge
 * is entered in both rows, as the reference does.  reorder_map[old id] = new id (the permutation
 * CSeqReservoir::reorder_items returns, seq_reservoir.cpp:215-251; NULL = identity).  Output: the CSR
 * pair list lzani_run takes -- row_off[n + 1], q_ids[*n_entries = 2 * n_edges], ids ascending within
 * each row (the reference's in-row order is thread-arrival order and unobservable).  The multiset of
 * (row, query) entries is the reference's: nothing is deduplicated or dropped.  When 2 * n_edges > cap
 * nothing is written and LZANI_ERR_NOMEM is returned. */
int lzani_build_pair_list(lzani_ctx* ctx, uint32_t n, uint64_t n_edges, const uint32_t* edge_i,
                          const uint32_t* edge_j, const uint32_t* reorder_map, uint64_t* row_off,
                          uint32_t* q_ids, uint64_t cap, uint64_t* n_entries);

/* NO COUNTERPART IN THE REFERENCE -- added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  The --out-filter pre-filter of lzani_filter_pairs() over a DENSE all2all result matrix the
 * caller already holds in device memory, allocating nothing: the kernels' scratch (a flag byte per
 * ordered pair, a count and an offset per 256 pairs) is one file-scope __device__ array of about
 * 4.2 MB, sized for LZANI_NOALLOC_MAX_SEQS sequences, that every GPU which loads the library holds
 * until the process exits.  The call makes no cudaMalloc / cudaFree / allocating library call and
 * does not synchronise: it launches three kernels on `stream` (a cudaStream_t, NULL = the default
 * stream) of `device` and returns.  (The first call per device looks the array's address up with
 * cudaGetSymbolAddress; make that first call outside a graph capture.)
 *   d_res     DEVICE: n_seqs * (n_seqs - 1) results, row r = reference r, its queries = every id != r
 *             ascending (what lzani_run / lzani_run_rows write for the dense pair list)
 *   d_eff_len DEVICE: n_seqs lengths without contig separators (as lzani_filter_pairs' eff_len)
 *   flt       HOST: the thresholds
 *   d_pairs   DEVICE: room for `cap` surviving pairs, written in CSR order like lzani_filter_pairs
 *   d_n_out   DEVICE: the number of surviving pairs -- the full count even when it exceeds cap (then
 *             only the first cap are written); 2^64-1 if the scratch address check failed on the device
 *   err       HOST, may be NULL: receives a message when the return value is not LZANI_OK
 * n_seqs > LZANI_NOALLOC_MAX_SEQS is rejected with LZANI_ERR_ARG (use lzani_filter_pairs).  All calls
 * on one device share the scratch array: calls that overlap on the GPU race -- keep them on one stream
 * or order them yourself.  The array keeps the previous call's contents; every word a call reads it
 * has written itself, so nothing needs clearing between calls. */
#define LZANI_NOALLOC_MAX_SEQS 2048
int lzani_filter_dense_noalloc(int device, uint32_t n_seqs, const lzani_result_t* d_res,
                               const uint32_t* d_eff_len, const lzani_out_filter_t* flt,
                               lzani_pair_t* d_pairs, uint64_t cap, uint64_t* d_n_out, void* stream,
                               char* err, size_t err_len);

int lzani_get_timing(const lzani_ctx* ctx, lzani_timing_t* t);

/* Bytes of device memory the index batches may use (default: 60% of free memory). */
int lzani_set_index_budget(lzani_ctx* ctx, uint64_t bytes);

const char* lzani_last_error(const lzani_ctx* ctx);
void lzani_destroy(lzani_ctx* ctx);
int lzani_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* LZANI_B200_H */
