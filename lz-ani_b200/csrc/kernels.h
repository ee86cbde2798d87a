// kernels.h -- launch wrappers of kernels.cu, used by the C ABI (capi.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "layout.h"
#include "regions.h"

#define LZ_PARSE_THREADS 128
#ifndef LZ_PARSE_MIN_BLOCKS
#define LZ_PARSE_MIN_BLOCKS 5
#endif

namespace lzani {

cudaError_t launch_pack(cudaStream_t st, const uint8_t* src, const SeqMeta* meta, const uint64_t* rec_base,
                        uint32_t n_seqs, uint64_t n_recs, int mrd, Rec* recs);

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

}  // namespace lzani
