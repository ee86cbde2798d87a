// kernels.h -- launch wrappers of kernels.cu, used by the C ABI (capi.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "layout.h"
#include "regions.h"
#include "outfilter.h"

#ifndef LZ_PARSE_THREADS
#define LZ_PARSE_THREADS 128
#endif
#ifndef LZ_PARSE_MIN_BLOCKS
#define LZ_PARSE_MIN_BLOCKS 5
#endif

#ifndef LZ_WALK_THREADS
#define LZ_WALK_THREADS 128
#endif
#ifndef LZ_WALK_MIN_BLOCKS
#define LZ_WALK_MIN_BLOCKS 5
#endif
// lane-per-pair walk (k_walk_lanes): block size, resident blocks, pairs per lane, plain events per round, lanes with
// a waiting pair that trigger the general step / the seed scan, rounds a waiting pair waits at most
#ifndef LZ_LANE_THREADS
#define LZ_LANE_THREADS 128
#endif
#ifndef LZ_LANE_MIN_BLOCKS
#define LZ_LANE_MIN_BLOCKS 4
#endif
#ifndef LZ_LANE_SLOTS
#define LZ_LANE_SLOTS 3
#endif
#ifndef LZ_LANE_RUN_STEPS
#define LZ_LANE_RUN_STEPS 8
#endif
#ifndef LZ_LANE_PARK_SLOW
#define LZ_LANE_PARK_SLOW 24
#endif
#ifndef LZ_LANE_PARK_SCAN
#define LZ_LANE_PARK_SCAN 24
#endif
#ifndef LZ_LANE_WAIT
#define LZ_LANE_WAIT 32
#endif
// a lane hands its pair to the warp-per-pair walk after BASE + (query positions >> SHIFT) parked steps (an unrelated
// 50 kb pair parks ~120 times, a related one 600-1300 times)
// ... and a pair whose hit runs number this many percent of what an unrelated pair would show goes there at once
#ifndef LZ_LANE_DENSE_PCT
#define LZ_LANE_DENSE_PCT 125
#endif
#ifndef LZ_LANE_HARD_BASE
#define LZ_LANE_HARD_BASE 64
#endif
#ifndef LZ_LANE_HARD_SHIFT
#define LZ_LANE_HARD_SHIFT 8
#endif
// the hand-over to the warp-per-pair walk pays when related pairs are the exception in a chunk (they would stall the
// warp's other lanes); when more than 1 pair in LZ_LANE_KEEP_DEN is related, every warp is busy with general steps
// anyway and the lane walk keeps them all (config 4, 20 000 genomes: parse 246 -> 177 ms)
#ifndef LZ_LANE_KEEP_DEN
#define LZ_LANE_KEEP_DEN 4
#endif
#ifndef LZ_RECORD_THREADS
#define LZ_RECORD_THREADS 256
#endif
#ifndef LZ_SPEC_THREADS
#define LZ_SPEC_THREADS 256
#endif

namespace lzani {

cudaError_t launch_pack(cudaStream_t st, const uint8_t* src, const SeqMeta* meta, const uint64_t* rec_base,
                        uint32_t n_seqs, uint64_t n_recs, int mrd, Rec* recs, bool ascii = false);

cudaError_t launch_build_index(cudaStream_t st, const Rec* recs, const SeqMeta* meta, const uint32_t* row_ref,
                               const RefIndex* row_idx, const uint32_t* blk_base, uint32_t n_rows,
                               uint32_t n_blocks, const Prm& P, uint32_t* bitmap_pool, uint64_t* table_pool);

int parse_blocks_per_sm(const Prm& P);

// one region of one pair as the regions kernel appends it (host groups by pair and sorts)
struct DevRegion { unsigned long long pair; RegionRec g; };

cudaError_t launch_parse_regions(cudaStream_t st, int grid_blocks, const Rec* recs, const SeqMeta* meta,
                                 const uint32_t* row_ref, const RefIndex* row_idx, const uint64_t* row_off,
                                 uint32_t n_rows, const uint32_t* q_ids, uint64_t pair_begin, uint64_t pair_end,
                                 const Prm& P, const uint32_t* bitmap_pool, const uint64_t* table_pool,
                                 unsigned long long* counter, void* out, DevRegion* regs,
                                 unsigned long long* reg_counter, uint64_t reg_cap);

cudaError_t launch_parse(cudaStream_t st, int grid_blocks, const Rec* recs, const SeqMeta* meta,
                         const uint32_t* row_ref, const RefIndex* row_idx, const uint64_t* row_off,
                         uint32_t n_rows, const uint32_t* q_ids, uint64_t pair_begin, uint64_t pair_end,
                         const Prm& P, const uint32_t* bitmap_pool, const uint64_t* table_pool,
                         unsigned long long* counter, void* out);

// ---- staged pipeline (kernels.cu "K3 staged"): per-chunk scratch and launches
struct PairDesc {                 // everything the stage kernels need to know about one pair of the chunk
    uint64_t q_rec, r_rec;        // first records of the query's Q-form / the reference's R-form
    uint64_t bitmap_word, table_slot;
    int Dq, Dr;
    uint32_t bbits, tbits;
};
struct StagedBufs {
    PairDesc* desc;               // [pairs]
    uint32_t* n_words;            // [pairs + 1]
    uint32_t* word_off;           // [pairs + 1]  first word of each pair
    uint32_t* hm;                 // [words + 1]  hit masks
    uint32_t* wpair;              // [words]      pair of each word
    uint32_t* rec_off;            // [words + 1]  first record of each word; [words] = records of the chunk
    HitEntry* list;               // [records]
    Rec* recs;                    // [records]
    void* tmp; size_t tmp_bytes;  // CUB scan scratch
    unsigned long long* counter;  // [4] work counters of the walk kernels; [3]: 64-bit hit-run total
    uint32_t* hard;               // [pairs] pairs the lane walk hands to the warp walk
};
// all2all form of the scan (k_bitmap_matrix + k_scan_matrix): the chunk is made of whole dense rows
struct MatrixScan {
    uint32_t* mc;                 // [2^bbits] one word per bitmap bit (direct bitmaps: per k-mer): which of the block's <= 32 references have it
    uint32_t bbits;               // log2(bits) of every bitmap of the chunk's rows (they must agree)
    uint32_t row_first;           // first row of the chunk, relative to the row arrays handed to the launch
    uint32_t n_rows_chunk;
    uint32_t n_seqs;              // queries = every sequence
    uint32_t max_words;           // longest query, in 32-position words
};
int walk_blocks_per_sm(const Prm& P);
int lane_blocks_per_sm(const Prm& P);
size_t staged_scan_bytes(uint64_t n_items);
// S0, scan, S1, scan: afterwards rec_off[n_words] holds the number of records of the chunk
cudaError_t launch_staged_scan(cudaStream_t st, const StagedBufs& B, const Rec* recs, const SeqMeta* meta,
                               const uint32_t* row_ref, const RefIndex* row_idx, const uint64_t* row_off,
                               uint32_t n_rows, const uint32_t* q_ids, uint64_t pair_begin, uint32_t n_pairs,
                               uint64_t n_words, const Prm& P, const uint32_t* bitmap_pool, int scan_blocks,
                               const MatrixScan* mx = nullptr);
// exact 64-bit number of hit runs of the chunk -> counter[3] (chunks whose 32-bit scan total could wrap)
cudaError_t launch_run_total(cudaStream_t st, const StagedBufs& B, uint64_t n_words);
// S2, S3, S4
cudaError_t launch_staged_walk(cudaStream_t st, int grid_blocks, const StagedBufs& B, const Rec* recs,
                               uint64_t pair_begin, uint32_t n_pairs, uint64_t n_words, uint64_t n_recs, const Prm& P,
                               const uint32_t* bitmap_pool, const uint64_t* table_pool, void* out,
                               int lane_grid_blocks = 0);       // > 0: the lane-per-pair walk (k_walk_lanes) on that many blocks

// ---- postproc.cu: pairing of the two directions, derived metrics, --out-filter compaction
// where the results of reference row r live (this GPU or a peer): res[i] is the result of the i-th
// query of the row, q its ascending query ids (nullptr = dense row: every id != r)
struct RowLoc { const void* res; const uint32_t* q; uint32_t cnt; uint32_t pad_; };
struct FilterArgs {
    const uint32_t* row_ref; const uint64_t* row_off; const uint32_t* q_ids; const void* res;   // this context's last run
    uint32_t n_rows; uint64_t n_pairs;
    const RowLoc* loc;            // per sequence id
    const uint32_t* eff_len;      // per sequence id: length without contig separators
    OutFilter flt;
};
struct PairOut { uint32_t a, b; int32_t ab[3], ba[3]; uint32_t lines; };   // == lzani_pair_t
uint64_t filter_blocks(uint64_t n_pairs);
cudaError_t launch_filter_count(cudaStream_t st, const FilterArgs& A, uint8_t* flags, uint32_t* blk_cnt,
                                unsigned long long* blk_off, unsigned long long* n_missing);
cudaError_t launch_filter_emit(cudaStream_t st, const FilterArgs& A, const uint8_t* flags,
                               const unsigned long long* blk_off, PairOut* out, uint64_t cap);

// ---- pairlist.cu: kmer-db edges -> symmetric, id-remapped, row-sorted CSR pair list
size_t pairlist_sort_scratch_bytes(uint64_t n_keys, int end_bit);
cudaError_t launch_pairlist(cudaStream_t st, const uint32_t* ei, const uint32_t* ej, uint64_t n_edges,
                            const uint32_t* map, uint32_t n, unsigned long long* keys_a, unsigned long long* keys_b,
                            void* scratch, size_t scratch_bytes, int end_bit, unsigned long long* row_off,
                            uint32_t* q_ids, unsigned long long* n_bad);

}  // namespace lzani
