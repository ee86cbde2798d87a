// pairlist.cu -- the sparse pair list on the device: from the edges a kmer-db filter
// file yields (reference CFilter::load_filter, src/filter.cpp:20-290: row i lists j < i with a value;
// every edge passing the threshold is entered in BOTH rows -- the symmetric closure :243-280) and the
// length-descending id permutation (CFilter::reorder_items, :293-345) to the CSR rows lzani_run takes.
//
//   P1 k_expand_edges : one thread per edge (i, j): keys (map[i] << 32 | map[j]) and (map[j] << 32 | map[i])
//   sort              : cub::DeviceRadixSort over the 2E 64-bit keys (library sort, only the bits in use)
//   P2 k_rows_from_keys: one thread per sorted key: q_ids[k] = column; row_off[] filled at row changes
//
// The multiset of (row, query) entries equals the reference's; within a row the reference's order is
// the arrival order of its loader threads (unobservable: do_matching sorts each result row by id,
// lz_matcher.cpp:253), here it is ascending, which lzani_filter_pairs needs for its binary search.
#include <cuda_runtime.h>
#include <stdint.h>

#include <cub/device/device_radix_sort.cuh>

#include "kernels.h"

namespace lzani {

namespace {

__global__ void __launch_bounds__(256)
k_expand_edges(const uint32_t* __restrict__ ei, const uint32_t* __restrict__ ej, uint64_t n_edges,
               const uint32_t* __restrict__ map, uint32_t n, unsigned long long* __restrict__ keys,
               unsigned long long* __restrict__ n_bad)
{
    const uint64_t e = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (e >= n_edges) return;
    uint32_t i = ei[e], j = ej[e];
    if (i >= n || j >= n) {                       // the reference would index out of bounds (filter.cpp:75-76)
        atomicAdd(n_bad, 1ull);
        i = j = 0;
    } else if (map) { i = map[i]; j = map[j]; }
    keys[2 * e] = ((unsigned long long)i << 32) | j;
    keys[2 * e + 1] = ((unsigned long long)j << 32) | i;
}

__global__ void __launch_bounds__(256)
k_rows_from_keys(const unsigned long long* __restrict__ keys, uint64_t n_keys, uint32_t n,
                 unsigned long long* __restrict__ row_off, uint32_t* __restrict__ q_ids)
{
    const uint64_t k = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (k > n_keys) return;
    if (k == n_keys) {                            // rows after the last key's row, and the end marker
        const uint32_t last = n_keys ? (uint32_t)(keys[n_keys - 1] >> 32) + 1u : 0u;
        for (uint32_t r = last; r <= n; ++r) row_off[r] = n_keys;
        return;
    }
    const unsigned long long key = keys[k];
    q_ids[k] = (uint32_t)key;
    const uint32_t row = (uint32_t)(key >> 32);
    const uint32_t first = k ? (uint32_t)(keys[k - 1] >> 32) + 1u : 0u;
    for (uint32_t r = first; r <= row; ++r) row_off[r] = k;   // runs only where the row changes
}

}  // namespace

size_t pairlist_sort_scratch_bytes(uint64_t n_keys, int end_bit)
{
    size_t bytes = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, bytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                   (int64_t)n_keys, 0, end_bit);
    return bytes;
}

cudaError_t launch_pairlist(cudaStream_t st, const uint32_t* ei, const uint32_t* ej, uint64_t n_edges,
                            const uint32_t* map, uint32_t n, unsigned long long* keys_a, unsigned long long* keys_b,
                            void* scratch, size_t scratch_bytes, int end_bit, unsigned long long* row_off,
                            uint32_t* q_ids, unsigned long long* n_bad)
{
    const uint64_t n_keys = 2 * n_edges;
    if (n_edges) {
        k_expand_edges<<<(unsigned)((n_edges + 255) / 256), 256, 0, st>>>(ei, ej, n_edges, map, n, keys_a, n_bad);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        e = cub::DeviceRadixSort::SortKeys(scratch, scratch_bytes, (const unsigned long long*)keys_a, keys_b, (int64_t)n_keys, 0,
                                           end_bit, st);
        if (e != cudaSuccess) return e;
    }
    k_rows_from_keys<<<(unsigned)((n_keys + 1 + 255) / 256), 256, 0, st>>>(keys_b, n_keys, n, row_off, q_ids);
    return cudaGetLastError();
}

}  // namespace lzani
