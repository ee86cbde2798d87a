// kernels.cu -- the three sm_100a kernels of the LZ-ANI pair path.
//
//   K1 k_pack_sequences : reservoir bytes (three_in_byte) or the loader's text (one byte per base)
//                         -> Q-form / R-form bit-plane records
//   K2 k_build_index    : per reference row, anchor presence bitmap + anchor table
//   K3 k_parse_pairs    : persistent grid, one warp per ordered pair (parse_core.h)
//
// Compile: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo (see __graft_entry__.build()).
#include <cuda_runtime.h>
#include <stdint.h>

#include "kernels.h"
#include "parse_core.h"

namespace lzani {

// index of the last element of a sorted prefix array (n+1 entries) that is <= v
template <class T>
__device__ __forceinline__ uint32_t find_segment(const T* __restrict__ prefix, uint32_t n, T v)
{
    uint32_t lo = 0, hi = n;   // invariant: prefix[lo] <= v < prefix[hi]
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (prefix[mid] <= v) lo = mid; else hi = mid;
    }
    return lo;
}

// ------------------------------------------------------------------------------------------
// K1.  One thread per output record.  rec_base[s] = first record of sequence s (its Q-form,
// directly followed by its R-form).
template <class Sym>
__global__ void __launch_bounds__(256)
k_pack_sequences(const uint8_t* __restrict__ src, const SeqMeta* __restrict__ meta,
                 const uint64_t* __restrict__ rec_base, uint32_t n_seqs, uint64_t n_recs, int mrd,
                 Rec* __restrict__ recs)
{
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_recs) return;
    const uint32_t s = find_segment<uint64_t>(rec_base, n_seqs, t);
    const SeqMeta m = meta[s];
    const uint32_t local = (uint32_t)(t - m.q_rec);
    const uint32_t nq = (uint32_t)(m.r_rec - m.q_rec);
    Rec o;
    const Sym sym{src + m.src_off};
    if (local < nq) o = pack_q_record_from(sym, m.len, local);
    else o = pack_r_record_from(sym, m.len, (uint32_t)mrd, local - nq);
    reinterpret_cast<uint4*>(recs)[t] = make_uint4(o.x, o.y, o.z, o.w);
}

// ------------------------------------------------------------------------------------------
// K2.  blk_base[j] = first block of batch row j; each block covers 256 consecutive R-form
// positions of its row's reference.  The bitmap and table regions were cleared beforehand
// (bitmap = 0, table = 0xFF..).
__global__ void __launch_bounds__(256)
k_build_index(const Rec* __restrict__ recs, const SeqMeta* __restrict__ meta,
              const uint32_t* __restrict__ row_ref, const RefIndex* __restrict__ row_idx,
              const uint32_t* __restrict__ blk_base, uint32_t n_rows, Prm P,
              uint32_t* __restrict__ bitmap_pool, unsigned long long* __restrict__ table_pool)
{
    const uint32_t j = find_segment<uint32_t>(blk_base, n_rows, blockIdx.x);
    const SeqMeta m = meta[row_ref[j]];
    const RefIndex ix = row_idx[j];
    const int Dr = r_form_len((int)m.len, P.mrd);
    const int pos = (int)((blockIdx.x - blk_base[j]) * 256u + threadIdx.x);
    if (pos >= Dr) return;
    uint64_t key;
    if (!ref_kmer_at(recs + m.r_rec, Dr, pos, P.mal, key)) return;
    const uint32_t b = bitmap_index(key, P.mal, ix.bbits);
    atomicOr(bitmap_pool + ix.bitmap_word + (b >> 5), 1u << (b & 31u));
    const uint32_t tmask = (1u << ix.tbits) - 1u;
    uint32_t h = table_home(key, ix.tbits);
    const unsigned long long e = ((unsigned long long)table_tag(key, P.mal) << 32) | (uint32_t)pos;
    unsigned long long* tab = table_pool + ix.table_slot;
    for (;;) {
        const unsigned long long old = atomicCAS(tab + h, (unsigned long long)LZ_EMPTY, e);
        if (old == (unsigned long long)LZ_EMPTY) break;
        h = (h + 1u) & tmask;
    }
}

// ------------------------------------------------------------------------------------------
// K3.  Persistent warps pull pair numbers from a global counter.  Pair k (relative to the
// batch's first pair) belongs to the batch row found in row_off; its query is q_ids[k] or, for
// the dense all2all list, the (k - row_off[row])-th id different from the reference.
template <class PT>
__global__ void __launch_bounds__(LZ_PARSE_THREADS, LZ_PARSE_MIN_BLOCKS)
k_parse_pairs(const Rec* __restrict__ recs, const SeqMeta* __restrict__ meta,
              const uint32_t* __restrict__ row_ref, const RefIndex* __restrict__ row_idx,
              const uint64_t* __restrict__ row_off, uint32_t n_rows, const uint32_t* __restrict__ q_ids,
              uint64_t pair_begin, uint64_t pair_end, PT P, const uint32_t* __restrict__ bitmap_pool,
              const uint64_t* __restrict__ table_pool, unsigned long long* __restrict__ counter,
              int3* __restrict__ out)
{
    DevWarp w((int)(threadIdx.x & 31u));
    __shared__ WarpScratch scratch[LZ_PARSE_THREADS / 32];
    WarpScratch* sc = &scratch[threadIdx.x >> 5];
    for (;;) {
        unsigned long long k = 0;
        if (w.lane == 0) k = atomicAdd(counter, 1ull);
        k = __shfl_sync(0xFFFFFFFFu, k, 0) + pair_begin;
        if (k >= pair_end) return;
        const uint32_t j = find_segment<uint64_t>(row_off, n_rows, (uint64_t)k);
        const uint32_t r = row_ref[j];
        uint32_t q;
        if (q_ids) q = q_ids[k];
        else {
            q = (uint32_t)(k - row_off[j]);
            if (q >= r) ++q;
        }
        const SeqMeta mr = meta[r], mq = meta[q];
        const RefIndex ix = row_idx[j];
        PairView pv;
        pv.q = recs + mq.q_rec;
        pv.Dq = (int)mq.len + P.mrd;
        pv.r = recs + mr.r_rec;
        pv.Dr = r_form_len((int)mr.len, P.mrd);
        pv.bitmap = bitmap_pool + ix.bitmap_word;
        pv.bbits = ix.bbits;
        pv.table = table_pool + ix.table_slot;
        pv.tbits = ix.tbits;
        const PairResult res = parse_pair(w, pv, P, sc);
        if (w.lane == 0) out[k] = make_int3(res.m, res.x, res.n);
    }
}

// ------------------------------------------------------------------------------------------
// K3r.  The same parse with the region fold switched on (regions.h): besides the counts, every kept
// region of every pair is appended to `regs` (any order; the host groups and sorts them the way
// calc_regions does).  Used for --out-alignment only; no speculative records, no fast path.
struct DevRegionOut {
    DevRegion* buf;
    unsigned long long* counter;
    unsigned long long cap;
    unsigned long long pair;
    int lane;
    __device__ __forceinline__ void emit(const RegionRec& g)
    {
        if (lane == 0) {
            const unsigned long long idx = atomicAdd(counter, 1ull);
            if (idx < cap) { DevRegion d; d.pair = pair; d.g = g; buf[idx] = d; }
        }
    }
};

template <class PT>
__global__ void __launch_bounds__(LZ_PARSE_THREADS)
k_parse_regions(const Rec* __restrict__ recs, const SeqMeta* __restrict__ meta,
                const uint32_t* __restrict__ row_ref, const RefIndex* __restrict__ row_idx,
                const uint64_t* __restrict__ row_off, uint32_t n_rows, const uint32_t* __restrict__ q_ids,
                uint64_t pair_begin, uint64_t pair_end, PT P, const uint32_t* __restrict__ bitmap_pool,
                const uint64_t* __restrict__ table_pool, unsigned long long* __restrict__ counter,
                int3* __restrict__ out, DevRegion* __restrict__ regs, unsigned long long* __restrict__ reg_counter,
                unsigned long long reg_cap)
{
    DevWarp w((int)(threadIdx.x & 31u));
    __shared__ WarpScratch scratch[LZ_PARSE_THREADS / 32];
    WarpScratch* sc = &scratch[threadIdx.x >> 5];
    for (;;) {
        unsigned long long k = 0;
        if (w.lane == 0) k = atomicAdd(counter, 1ull);
        k = __shfl_sync(0xFFFFFFFFu, k, 0) + pair_begin;
        if (k >= pair_end) return;
        const uint32_t j = find_segment<uint64_t>(row_off, n_rows, (uint64_t)k);
        const uint32_t r = row_ref[j];
        uint32_t q;
        if (q_ids) q = q_ids[k];
        else {
            q = (uint32_t)(k - row_off[j]);
            if (q >= r) ++q;
        }
        const SeqMeta mr = meta[r], mq = meta[q];
        const RefIndex ix = row_idx[j];
        PairView pv;
        pv.q = recs + mq.q_rec;
        pv.Dq = (int)mq.len + P.mrd;
        pv.r = recs + mr.r_rec;
        pv.Dr = r_form_len((int)mr.len, P.mrd);
        pv.bitmap = bitmap_pool + ix.bitmap_word;
        pv.bbits = ix.bbits;
        pv.table = table_pool + ix.table_slot;
        pv.tbits = ix.tbits;
        DevRegionOut ro{regs, reg_counter, reg_cap, k, w.lane};
        RegionTracker<DevRegionOut> rt;
        rt.out = &ro;
        const PairResult res = parse_pair(w, pv, P, sc, rt);
        if (w.lane == 0) out[k] = make_int3(res.m, res.x, res.n);
    }
}

// ------------------------------------------------------------------------------------------
// launch wrappers (called by capi.cu)
cudaError_t launch_pack(cudaStream_t st, const uint8_t* src, const SeqMeta* meta, const uint64_t* rec_base,
                        uint32_t n_seqs, uint64_t n_recs, int mrd, Rec* recs, bool ascii)
{
    if (!n_recs) return cudaSuccess;
    const uint64_t blocks = (n_recs + 255) / 256;
    if (ascii) k_pack_sequences<SymAscii><<<(unsigned)blocks, 256, 0, st>>>(src, meta, rec_base, n_seqs, n_recs, mrd, recs);
    else k_pack_sequences<SymPacked3><<<(unsigned)blocks, 256, 0, st>>>(src, meta, rec_base, n_seqs, n_recs, mrd, recs);
    return cudaGetLastError();
}

cudaError_t launch_build_index(cudaStream_t st, const Rec* recs, const SeqMeta* meta, const uint32_t* row_ref,
                               const RefIndex* row_idx, const uint32_t* blk_base, uint32_t n_rows,
                               uint32_t n_blocks, const Prm& P, uint32_t* bitmap_pool, uint64_t* table_pool)
{
    if (!n_blocks) return cudaSuccess;
    k_build_index<<<n_blocks, 256, 0, st>>>(recs, meta, row_ref, row_idx, blk_base, n_rows, P, bitmap_pool,
                                            reinterpret_cast<unsigned long long*>(table_pool));
    return cudaGetLastError();
}

int parse_blocks_per_sm(const Prm& P)
{
    int nb = 0;
    const cudaError_t e = is_default_kernel_params(P)
        ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_parse_pairs<PrmDefault>, LZ_PARSE_THREADS, 0)
        : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_parse_pairs<Prm>, LZ_PARSE_THREADS, 0);
    if (e != cudaSuccess || nb < 1) nb = 1;
    return nb;
}

cudaError_t launch_parse(cudaStream_t st, int grid_blocks, const Rec* recs, const SeqMeta* meta,
                         const uint32_t* row_ref, const RefIndex* row_idx, const uint64_t* row_off,
                         uint32_t n_rows, const uint32_t* q_ids, uint64_t pair_begin, uint64_t pair_end,
                         const Prm& P, const uint32_t* bitmap_pool, const uint64_t* table_pool,
                         unsigned long long* counter, void* out)
{
    if (pair_end <= pair_begin) return cudaSuccess;
    if (is_default_kernel_params(P)) {
        const PrmDefault D{P.mrd, P.mqd, P.reg};
        k_parse_pairs<PrmDefault><<<grid_blocks, LZ_PARSE_THREADS, 0, st>>>(recs, meta, row_ref, row_idx, row_off, n_rows,
                                                                        q_ids, pair_begin, pair_end, D, bitmap_pool,
                                                                        table_pool, counter, reinterpret_cast<int3*>(out));
    } else
        k_parse_pairs<Prm><<<grid_blocks, LZ_PARSE_THREADS, 0, st>>>(recs, meta, row_ref, row_idx, row_off, n_rows, q_ids,
                                                                 pair_begin, pair_end, P, bitmap_pool, table_pool,
                                                                 counter, reinterpret_cast<int3*>(out));
    return cudaGetLastError();
}

cudaError_t launch_parse_regions(cudaStream_t st, int grid_blocks, const Rec* recs, const SeqMeta* meta,
                                 const uint32_t* row_ref, const RefIndex* row_idx, const uint64_t* row_off,
                                 uint32_t n_rows, const uint32_t* q_ids, uint64_t pair_begin, uint64_t pair_end,
                                 const Prm& P, const uint32_t* bitmap_pool, const uint64_t* table_pool,
                                 unsigned long long* counter, void* out, DevRegion* regs,
                                 unsigned long long* reg_counter, uint64_t reg_cap)
{
    if (pair_end <= pair_begin) return cudaSuccess;
    if (is_default_kernel_params(P)) {
        const PrmDefault D{P.mrd, P.mqd, P.reg};
        k_parse_regions<PrmDefault><<<grid_blocks, LZ_PARSE_THREADS, 0, st>>>(
            recs, meta, row_ref, row_idx, row_off, n_rows, q_ids, pair_begin, pair_end, D, bitmap_pool, table_pool, counter,
            reinterpret_cast<int3*>(out), regs, reg_counter, (unsigned long long)reg_cap);
    } else
        k_parse_regions<Prm><<<grid_blocks, LZ_PARSE_THREADS, 0, st>>>(
            recs, meta, row_ref, row_idx, row_off, n_rows, q_ids, pair_begin, pair_end, P, bitmap_pool, table_pool, counter,
            reinterpret_cast<int3*>(out), regs, reg_counter, (unsigned long long)reg_cap);
    return cudaGetLastError();
}

}  // namespace lzani
