// postproc.cu -- what follows the parse on the device: pairing the two directions of
// every unordered pair, the derived metrics, the --out-filter test and the compaction of the
// surviving pairs, so that only those cross PCIe and reach the host writer.
//
//   F1 k_filter_flags : one thread per ordered pair (a, b) of this context's rows with a < b: find the
//                       result of the opposite direction (row b, entry a) -- in this GPU's memory or,
//                       through NVLink peer access, in the memory of the GPU that owns row b --
//                       evaluate out_filter_mask (outfilter.h), count survivors per block
//   F2 k_scan_blocks  : exclusive scan of the block counts (one block)
//   F3 k_filter_emit  : block-level scan, stable compaction of the surviving pairs in CSR order
//
// Replaces the filter arithmetic of CLZMatcher::store_results (reference src/lz_matcher.cpp:434-463)
// as a pre-filter: the host writer stays the reference's own and prints the same text.
#include <cuda_runtime.h>
#include <stdint.h>

#include "kernels.h"
#include "outfilter.h"

namespace lzani {

namespace {

template <class T>
__device__ __forceinline__ uint32_t seg_of(const T* __restrict__ prefix, uint32_t n, T v)
{
    uint32_t lo = 0, hi = n;   // prefix[lo] <= v < prefix[hi]
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (prefix[mid] <= v) lo = mid; else hi = mid;
    }
    return lo;
}

constexpr int FT = 256;   // threads per block of F1/F3

struct PairAt {
    uint32_t a, b;
    const int3* rev;     // result of the opposite direction, nullptr when absent
};

// ordered pair k of this context -> (a, b) and where the opposite direction's result lives
__device__ __forceinline__ bool locate(const FilterArgs& A, uint64_t k, PairAt& p)
{
    const uint32_t j = seg_of<uint64_t>(A.row_off, A.n_rows, k);
    p.a = A.row_ref[j];
    if (A.q_ids) p.b = A.q_ids[k];
    else {
        p.b = (uint32_t)(k - A.row_off[j]);
        if (p.b >= p.a) ++p.b;
    }
    p.rev = nullptr;
    if (p.b <= p.a) return false;              // the unordered pair is handled from its smaller id
    const RowLoc L = A.loc[p.b];
    if (!L.res) return true;
    if (!L.q) {                                // dense row b: every id != b in increasing order, a < b
        p.rev = reinterpret_cast<const int3*>(L.res) + p.a;
        return true;
    }
    uint32_t lo = 0, hi = L.cnt;               // sparse row b: ids ascending (checked by the host)
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (L.q[mid] < p.a) lo = mid + 1; else hi = mid;
    }
    if (lo < L.cnt && L.q[lo] == p.a) p.rev = reinterpret_cast<const int3*>(L.res) + lo;
    return true;
}

__global__ void __launch_bounds__(FT)
k_filter_flags(FilterArgs A, uint8_t* __restrict__ flags, uint32_t* __restrict__ blk_cnt,
               unsigned long long* __restrict__ n_missing)
{
    const uint64_t k = (uint64_t)blockIdx.x * FT + threadIdx.x;
    uint32_t mask = 0;
    if (k < A.n_pairs) {
        PairAt p;
        if (locate(A, k, p)) {
            if (!p.rev) atomicAdd(n_missing, 1ull);
            else {
                const int3 ab = reinterpret_cast<const int3*>(A.res)[k];
                const int3 ba = *p.rev;        // possibly a load from a peer GPU over NVLink
                mask = out_filter_mask(A.flt, DirStats{ab.x, ab.y, ab.z}, DirStats{ba.x, ba.y, ba.z},
                                       A.eff_len[p.a], A.eff_len[p.b]);
            }
        }
        flags[k] = (uint8_t)mask;
    }
    const int c = __syncthreads_count(mask != 0);
    if (threadIdx.x == 0) blk_cnt[blockIdx.x] = (uint32_t)c;
}

// exclusive scan of cnt[0..n) into off[0..n], off[n] = total; one block of 1024 threads
__global__ void __launch_bounds__(1024)
k_scan_blocks(const uint32_t* __restrict__ cnt, uint64_t n, unsigned long long* __restrict__ off)
{
    __shared__ unsigned long long part[1024];
    const uint64_t per = (n + 1023) / 1024;
    const uint64_t lo = (uint64_t)threadIdx.x * per;
    const uint64_t hi = lo + per < n ? lo + per : n;
    unsigned long long s = 0;
    for (uint64_t i = lo; i < hi; ++i) s += cnt[i];
    part[threadIdx.x] = s;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {       // Hillis-Steele inclusive scan of the 1024 partial sums
        const unsigned long long v = threadIdx.x >= d ? part[threadIdx.x - d] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned long long run = threadIdx.x ? part[threadIdx.x - 1] : 0;
    for (uint64_t i = lo; i < hi; ++i) { off[i] = run; run += cnt[i]; }
    if (threadIdx.x == 1023) off[n] = part[1023];
}

__global__ void __launch_bounds__(FT)
k_filter_emit(FilterArgs A, const uint8_t* __restrict__ flags, const unsigned long long* __restrict__ blk_off,
              PairOut* __restrict__ out, unsigned long long cap)
{
    __shared__ uint32_t warp_base[FT / 32];
    const uint64_t k = (uint64_t)blockIdx.x * FT + threadIdx.x;
    const uint32_t mask = k < A.n_pairs ? flags[k] : 0u;
    const uint32_t bal = __ballot_sync(0xFFFFFFFFu, mask != 0);
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    if (lane == 0) warp_base[wid] = (uint32_t)__popc(bal);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = 0;
        for (int i = 0; i < FT / 32; ++i) { const uint32_t c = warp_base[i]; warp_base[i] = run; run += c; }
    }
    __syncthreads();
    if (!mask) return;
    const unsigned long long dst = blk_off[blockIdx.x] + warp_base[wid] + (uint32_t)__popc(bal & ((1u << lane) - 1u));
    if (dst >= cap) return;
    PairAt p;
    locate(A, k, p);
    const int3 ab = reinterpret_cast<const int3*>(A.res)[k];
    const int3 ba = *p.rev;
    PairOut o;
    o.a = p.a; o.b = p.b;
    o.ab[0] = ab.x; o.ab[1] = ab.y; o.ab[2] = ab.z;
    o.ba[0] = ba.x; o.ba[1] = ba.y; o.ba[2] = ba.z;
    o.lines = mask;
    out[dst] = o;
}

}  // namespace

uint64_t filter_blocks(uint64_t n_pairs) { return (n_pairs + FT - 1) / FT; }

cudaError_t launch_filter_count(cudaStream_t st, const FilterArgs& A, uint8_t* flags, uint32_t* blk_cnt,
                                unsigned long long* blk_off, unsigned long long* n_missing)
{
    const uint64_t nb = filter_blocks(A.n_pairs);
    if (nb) {
        k_filter_flags<<<(unsigned)nb, FT, 0, st>>>(A, flags, blk_cnt, n_missing);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    k_scan_blocks<<<1, 1024, 0, st>>>(blk_cnt, nb, blk_off);
    return cudaGetLastError();
}

cudaError_t launch_filter_emit(cudaStream_t st, const FilterArgs& A, const uint8_t* flags,
                               const unsigned long long* blk_off, PairOut* out, uint64_t cap)
{
    const uint64_t nb = filter_blocks(A.n_pairs);
    if (!nb) return cudaSuccess;
    k_filter_emit<<<(unsigned)nb, FT, 0, st>>>(A, flags, blk_off, out, (unsigned long long)cap);
    return cudaGetLastError();
}

}  // namespace lzani
