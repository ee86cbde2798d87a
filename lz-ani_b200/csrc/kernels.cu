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

#include <cub/cub.cuh>
#include <thrust/iterator/transform_iterator.h>

#include "kernels.h"
#include "parse_core.h"
#include "lane_walk.h"

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
// K3 staged.  The parse of a CHUNK of consecutive pairs as a pipeline of plain kernels (parse_core.h,
// "staged pipeline"): what depends on the query position alone is computed for all pairs of the chunk
// by data-parallel kernels at full occupancy and stored; only the sequential walk stays warp-per-pair.
//
//   S0 k_pair_desc     thread per pair        pair descriptor + number of 32-position words of its loop
//      (exclusive scan -> word_off)
//   S1 k_scan_hits     thread per word        hit mask of 32 query positions (anchor presence bitmap)
//      (exclusive scan of the run counts -> rec_off)
//   S2 k_expand_hits   thread per word        hit list: (pair, position, run length) per hit run
//   S3 k_hit_probe     thread per hit run     table probe, match length, both approximate extensions
//      k_hit_spec      thread per hit run     speculative seed scan (jobs sorted by length per block)
//                                             + successor link -> one 16-byte record per hit run
//   S4 k_walk_pairs    warp per pair          the state machine of parser.cpp:501-708 over the records
template <class PT>
__global__ void __launch_bounds__(256)
k_pair_desc(const SeqMeta* __restrict__ meta, const uint32_t* __restrict__ row_ref,
            const RefIndex* __restrict__ row_idx, const uint64_t* __restrict__ row_off, uint32_t n_rows,
            const uint32_t* __restrict__ q_ids, uint64_t pair_begin, uint32_t n_pairs, PT P,
            PairDesc* __restrict__ desc, uint32_t* __restrict__ n_words)
{
    const uint32_t pl = blockIdx.x * 256u + threadIdx.x;
    if (pl > n_pairs) return;
    if (pl == n_pairs) { n_words[pl] = 0; return; }
    const uint64_t k = pair_begin + pl;
    const uint32_t j = find_segment<uint64_t>(row_off, n_rows, k);
    const uint32_t r = row_ref[j];
    uint32_t q;
    if (q_ids) q = q_ids[k];
    else {
        q = (uint32_t)(k - row_off[j]);
        if (q >= r) ++q;
    }
    const SeqMeta mr = meta[r], mq = meta[q];
    const RefIndex ix = row_idx[j];
    PairDesc d;
    d.q_rec = mq.q_rec; d.r_rec = mr.r_rec;
    d.bitmap_word = ix.bitmap_word; d.table_slot = ix.table_slot;
    d.Dq = (int)mq.len + P.mrd;
    d.Dr = r_form_len((int)mr.len, P.mrd);
    d.bbits = ix.bbits; d.tbits = ix.tbits;
    desc[pl] = d;
    const int n_loop = d.Dq - P.msl;
    n_words[pl] = n_loop > 0 ? (uint32_t)((n_loop + 31) / 32) : 0u;
}

__device__ __forceinline__ PairView view_of(const PairDesc& d, const Rec* recs, const uint32_t* bitmap_pool,
                                            const uint64_t* table_pool)
{
    PairView pv;
    pv.q = recs + d.q_rec; pv.Dq = d.Dq;
    pv.r = recs + d.r_rec; pv.Dr = d.Dr;
    pv.bitmap = bitmap_pool + d.bitmap_word; pv.bbits = d.bbits;
    pv.table = table_pool + d.table_slot; pv.tbits = d.tbits;
    return pv;
}

// S1.  Persistent blocks of LZ_SCAN_THREADS threads, one per SM, each over a contiguous range of tiles of
// LZ_SCAN_TILE words.  The anchor presence bitmap of the reference is random-accessed once per query
// position -- 32-byte L2 sectors for 1 bit -- so the block keeps a folded copy (<= 128 KiB, coarse_fold_word)
// in shared memory and only the positions that pass it touch the exact bitmap.  Consecutive pairs of a
// chunk share their reference (CSR order), so the copy is re-folded only when the reference changes.
constexpr int LZ_SCAN_THREADS = 1024;
constexpr uint32_t LZ_SCAN_TILE = 8192;
template <class PT>
__global__ void __launch_bounds__(LZ_SCAN_THREADS, 1)
k_scan_hits(const Rec* __restrict__ recs, const PairDesc* __restrict__ desc, const uint32_t* __restrict__ word_off,
            uint32_t n_pairs, PT P, const uint32_t* __restrict__ bitmap_pool, uint32_t* __restrict__ hm,
            uint32_t* __restrict__ wpair)
{
    extern __shared__ uint32_t coarse[];
    __shared__ uint32_t s_pl, s_end;
    const uint32_t n_words = word_off[n_pairs];
    const uint32_t n_tiles = (n_words + LZ_SCAN_TILE - 1) / LZ_SCAN_TILE;
    const uint32_t t0 = (uint32_t)((uint64_t)n_tiles * blockIdx.x / gridDim.x);
    const uint32_t t1 = (uint32_t)((uint64_t)n_tiles * (blockIdx.x + 1) / gridDim.x);
    unsigned long long have = ~0ull;          // bitmap whose folded copy is in shared memory
    for (uint32_t tile = t0; tile < t1; ++tile) {
        const uint32_t wb = tile * LZ_SCAN_TILE;
        const uint32_t we = wb + LZ_SCAN_TILE < n_words ? wb + LZ_SCAN_TILE : n_words;
        uint32_t w = wb;
        while (w < we) {
            // segment: the words of consecutive pairs with the same reference, inside the tile
            if (threadIdx.x == 0) {
                uint32_t pl = find_segment<uint32_t>(word_off, n_pairs, w);
                const unsigned long long bm = desc[pl].bitmap_word;
                uint32_t e = pl;
                while (e + 1 < n_pairs && word_off[e + 1] < we && desc[e + 1].bitmap_word == bm) ++e;
                s_pl = pl;
                s_end = word_off[e + 1] < we ? word_off[e + 1] : we;
            }
            __syncthreads();
            const uint32_t pl0 = s_pl, seg_end = s_end;
            const PairDesc d0 = desc[pl0];
            if (d0.bitmap_word != have) {
                const uint32_t cw = coarse_words(d0.bbits);
                for (uint32_t c = threadIdx.x; c < cw; c += LZ_SCAN_THREADS)
                    coarse[c] = coarse_fold_word(bitmap_pool + d0.bitmap_word, d0.bbits, c);
                have = d0.bitmap_word;
                __syncthreads();
            }
            uint32_t pl = pl0;
            for (uint32_t gw = w + threadIdx.x; gw < seg_end; gw += LZ_SCAN_THREADS) {
                while (gw >= word_off[pl + 1]) ++pl;
                const PairDesc d = desc[pl];
                const PairView pv = view_of(d, recs, bitmap_pool, nullptr);
                hm[gw] = scan_word_coarse(pv, P, (int)(32u * (gw - word_off[pl])), d.Dq - P.msl, coarse);
                wpair[gw] = pl;
            }
            w = seg_end;
            __syncthreads();
        }
    }
}

// S1m.  The all2all form of S1 (dense rows, direct-addressed bitmaps): instead of one probe per query
// position and REFERENCE, the presence bitmaps of up to 32 references of the chunk are transposed into one
// 32-bit word per k-mer (k_bitmap_matrix: Mc[key] bit i <=> reference i has the k-mer), a query word is
// scanned once against all of them -- one 4-byte gather per position -- and a register transpose turns the
// 32 gathered words into the 32 hit masks.  ~40x fewer instructions per (pair, word) than the probe scan.
__global__ void __launch_bounds__(256)
k_bitmap_matrix(const uint32_t* __restrict__ bitmap_pool, const RefIndex* __restrict__ row_idx, uint32_t nr,
                uint32_t n_key_words, uint32_t* __restrict__ mc)
{
    const uint32_t t = blockIdx.x * 256u + threadIdx.x;
    if (t >= n_key_words) return;
    uint32_t a[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) a[i] = (uint32_t)i < nr ? bitmap_pool[row_idx[i].bitmap_word + t] : 0u;
    transpose32(a);
    uint4* o = reinterpret_cast<uint4*>(mc + 32ull * t);
#pragma unroll
    for (int b = 0; b < 8; ++b) o[b] = make_uint4(a[4 * b], a[4 * b + 1], a[4 * b + 2], a[4 * b + 3]);
}

// grid: (blocks over the words of a query, query id).  row_ref / row_off point at the first of the nr rows
// of the block; chunk_k0 = first pair of the chunk (global pair number, as in row_off).
template <class PT>
__global__ void __launch_bounds__(128)
k_scan_matrix(const Rec* __restrict__ recs, const SeqMeta* __restrict__ meta, const uint32_t* __restrict__ row_ref,
              const uint64_t* __restrict__ row_off, uint32_t nr, uint64_t chunk_k0,
              const uint32_t* __restrict__ word_off, PT P, const uint32_t* __restrict__ mc, uint32_t key_mask, uint32_t bbits,
              uint32_t* __restrict__ hm, uint32_t* __restrict__ wpair)
{
    __shared__ uint32_t s_ref[32];
    __shared__ uint32_t s_pl0[32];
    const uint32_t q = blockIdx.y;
    if (threadIdx.x < 32) {
        const uint32_t i = threadIdx.x;
        uint32_t r = 0xFFFFFFFFu, p0 = 0;
        if (i < nr) {
            r = row_ref[i];
            p0 = (uint32_t)(row_off[i] - chunk_k0) + (q < r ? q : q - 1u);     // dense row: the (q - [q > r])-th pair
        }
        s_ref[i] = r;
        s_pl0[i] = p0;
    }
    __syncthreads();
    const SeqMeta mq = meta[q];
    const int n_loop = (int)mq.len + P.mrd - P.msl;
    const uint32_t w = blockIdx.x * 128u + threadIdx.x;
    if (n_loop <= 0 || w >= (uint32_t)((n_loop + 31) / 32)) return;
    const Rec* qr = recs + mq.q_rec;
    const Rec a = ld_rec(qr + w), b = ld_rec(qr + w + 1);
    const uint32_t km = low_mask32(P.mal);
    uint32_t v[32];
    if (2 * P.mal <= LZ_DIRECT_BITMAP_MAX_BITS) {     // direct bitmaps: the matrix is indexed by the k-mer itself
        const uint32_t h0 = a.y << P.mal, h1 = funnel_r(a.y, b.y, 32 - P.mal);
#pragma unroll
        for (int t = 0; t < 32; ++t) {
            const uint32_t lo = funnel_r(a.x, b.x, t), hs = funnel_r(h0, h1, t);
            v[t] = __ldg(mc + (((lo & km) | (hs & ~km)) & key_mask));
        }
    } else {                                          // hashed bitmaps of one size: by the bitmaps' common bit index
#pragma unroll
        for (int t = 0; t < 32; ++t) {
            const uint32_t lo = funnel_r(a.x, b.x, t), hi = funnel_r(a.y, b.y, t);
            v[t] = __ldg(mc + bitmap_index(kmer_key(lo, hi, P.mal), P.mal, bbits));
        }
    }
    transpose32(v);
    const uint32_t valid = scan_valid_mask(a, b, P.mal, (int)(32u * w), n_loop);
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        if ((uint32_t)i < nr && s_ref[i] != q) {
            const uint32_t pl = s_pl0[i];
            const uint32_t gw = word_off[pl] + w;
            hm[gw] = v[i] & valid;
            wpair[gw] = pl;
        }
    }
}

// input of the second scan: hit runs per word
struct RunCount64 {
    __host__ __device__ __forceinline__ unsigned long long operator()(uint32_t m) const { return (unsigned long long)__builtin_popcount(run_first_bits(m)); }
};
struct RunCount {
    __host__ __device__ __forceinline__ uint32_t operator()(uint32_t m) const { return (uint32_t)__builtin_popcount(run_first_bits(m)); }
};

__global__ void __launch_bounds__(256)
k_expand_hits(const uint32_t* __restrict__ hm, const uint32_t* __restrict__ wpair, const uint32_t* __restrict__ word_off,
              uint32_t n_words, const uint32_t* __restrict__ rec_off, HitEntry* __restrict__ list)
{
    const uint32_t gw = blockIdx.x * 256u + threadIdx.x;
    if (gw >= n_words) return;
    const uint32_t m = hm[gw];
    if (!m) return;
    const uint32_t pl = wpair[gw];
    const uint32_t pos0 = 32u * (gw - word_off[pl]);
    uint32_t at = rec_off[gw];
    uint32_t f = run_first_bits(m);
    while (f) {
        const int b = ctz32(f);
        f &= f - 1u;
        HitEntry e;
        e.pair = pl;
        e.pr = (pos0 + (uint32_t)b) | ((uint32_t)run_following(m, b) << HIT_ENTRY_RUN_SHIFT);
        list[at++] = e;
    }
}

struct WarpSyncFn {
    __device__ __forceinline__ void operator()() const { __syncwarp(); }
};

// S3a.  One thread per hit run: table probe, match length, both approximate extensions (latency-bound:
// random table and sequence loads; few registers, full occupancy).
template <class PT>
__global__ void __launch_bounds__(LZ_RECORD_THREADS)
k_hit_probe(const Rec* __restrict__ recs, const PairDesc* __restrict__ desc, const HitEntry* __restrict__ list,
            const uint32_t* __restrict__ rec_total, PT P, const uint32_t* __restrict__ bitmap_pool,
            const uint64_t* __restrict__ table_pool, Rec* __restrict__ out)
{
    const uint32_t total = *rec_total;
    const uint32_t t = blockIdx.x * (uint32_t)LZ_RECORD_THREADS + threadIdx.x;
    const bool live = t < total;
    PairView pv{};
    if (live) pv = view_of(desc[list[t].pair], recs, bitmap_pool, table_pool);
    const Rec o = hit_record_probe(pv, P, list, (uint64_t)t, live, WarpSyncFn{});
    if (live) reinterpret_cast<uint4*>(out)[t] = make_uint4(o.x, o.y, o.z, o.w);
}

// S3b.  The speculative seed scans (ALU-bound: ~33 instructions per row, up to mqd+7 rows).  The scan
// length K differs from record to record (it ends at the next hit run), so the jobs of a thread block --
// two records per thread -- are counting-sorted by K through shared memory and dealt out in two passes:
// thread i takes the i-th longest job, then the i-th shortest.  The lanes of a warp then run about the same
// number of rows in each pass, and every warp of the block about the same number in total.
constexpr int LZ_SPEC_JOBS = 2 * LZ_SPEC_THREADS;
// hit records are read AND written by the threads of k_hit_spec (each thread its own record; verified_link also reads a
// successor's record, of which it uses only bits no k_hit_spec thread changes -- static_assert below): coherent loads
// (ld.global.cg), never the read-only path
__device__ __forceinline__ Rec ld_rec_cg(const Rec* p)
{
    const uint4 v = __ldcg(reinterpret_cast<const uint4*>(p));
    Rec r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w;
    return r;
}
// what hit_spec_finish() may change: x[29,32) (link), z[25,32) (seed offset), w[18,32) (scan length, first seed);
// what verified_link() reads of the successor: x[0,29), y, z flags [18,22), w[12,18)
static_assert(HIT_ENTRY_RUN_SHIFT == 29 && HIT_X_SHIFT == 25 && HIT_X_SHIFT > 21, "spec fields overlap what verified_link reads of a successor");

template <class PT>
__global__ void __launch_bounds__(LZ_SPEC_THREADS)
k_hit_spec(const Rec* __restrict__ recs, const PairDesc* __restrict__ desc, const HitEntry* __restrict__ list,
           const uint32_t* __restrict__ rec_total, PT P, Rec* hit_recs)
{
    struct Item { const Rec* q; const Rec* r; int si0, sP0, K; uint32_t succ; };
    __shared__ uint32_t s_hist[64];
    __shared__ uint16_t s_order[LZ_SPEC_JOBS];
    __shared__ Item s_item[LZ_SPEC_JOBS];
    const uint32_t total = *rec_total;
    const uint32_t tid = threadIdx.x;
    const uint32_t base = blockIdx.x * (uint32_t)LZ_SPEC_JOBS;
    if (tid < 64) s_hist[tid] = 0;
    __syncthreads();
    uint32_t key[2], rank[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint32_t slot = tid + (uint32_t)h * LZ_SPEC_THREADS;
        const uint32_t t = base + slot;
        key[h] = 63; rank[h] = 0;
        if (t < total) {
            const Rec rec = ld_rec_cg(hit_recs + t);
            const PairDesc d = desc[list[t].pair];
            const SpecJob j = hit_spec_job(P, d.Dq - P.msl, list, (uint64_t)t, (uint64_t)total, rec);
            if (j.need) {
                s_item[slot] = Item{recs + d.q_rec, recs + d.r_rec, j.si0, j.sP0, j.K, j.succ};
                key[h] = j.K < 0 ? 62u : (uint32_t)(SPEC_MAX_K - j.K);        // K <= SPEC_MAX_K = 57: longest first
                rank[h] = atomicAdd(&s_hist[key[h]], 1u);
            }
        }
    }
    __syncthreads();
    if (tid < 32) {                                                 // exclusive prefix over the 64 bins
        const uint32_t a = s_hist[2 * tid], b = s_hist[2 * tid + 1];
        uint32_t v = a + b;
#pragma unroll
        for (int dlt = 1; dlt < 32; dlt <<= 1) {
            const uint32_t u = __shfl_up_sync(0xFFFFFFFFu, v, dlt);
            if ((int)tid >= dlt) v += u;
        }
        s_hist[2 * tid] = v - a - b;
        s_hist[2 * tid + 1] = v - b;
    }
    __syncthreads();
#pragma unroll
    for (int h = 0; h < 2; ++h)
        if (key[h] != 63u) s_order[s_hist[key[h]] + rank[h]] = (uint16_t)(tid + (uint32_t)h * LZ_SPEC_THREADS);
    const uint32_t n_jobs = s_hist[63];                             // bin 63 is never used: its prefix is the job count
    __syncthreads();
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        // pass 0: the tid-th longest job; pass 1: the tid-th shortest of those left
        uint32_t jx;
        if (pass == 0) { if (tid >= n_jobs) continue; jx = tid; }
        else { if (n_jobs <= (uint32_t)LZ_SPEC_THREADS || tid >= n_jobs - (uint32_t)LZ_SPEC_THREADS) continue; jx = n_jobs - 1u - tid; }
        const uint32_t o = s_order[jx];
        const Item it = s_item[o];
        PairView pv{};
        pv.q = it.q; pv.r = it.r;
        const int ks = spec_first_seed(pv, P, it.si0, it.sP0, it.K, P.mrd);
        const uint32_t to = base + o;
        Rec r2 = ld_rec_cg(hit_recs + to);
        SpecJob jj; jj.si0 = it.si0; jj.sP0 = it.sP0; jj.K = it.K; jj.succ = it.succ; jj.need = true;
        hit_spec_finish(r2, jj, ks);
        // successor link (verified_link): needs this record complete and, of the successor, only what
        // k_hit_probe wrote (position, anchor, flags, backward extension) -- bits no k_hit_spec thread changes,
        // so it does not matter whether the successor's own scan has been folded in yet
        const uint32_t d = r2.x >> HIT_ENTRY_RUN_SHIFT;
        if (d && verified_link(P, r2, ld_rec_cg(hit_recs + to + d)) != d) r2.x &= (uint32_t)(LZ_STAGED_MAX_POS - 1);
        reinterpret_cast<uint4*>(hit_recs)[to] = make_uint4(r2.x, r2.y, r2.z, r2.w);
    }
}

template <class PT>
__global__ void __launch_bounds__(LZ_WALK_THREADS, LZ_WALK_MIN_BLOCKS)
k_walk_pairs(const Rec* __restrict__ recs, const PairDesc* __restrict__ desc, const uint32_t* __restrict__ word_off,
             const uint32_t* __restrict__ rec_off, const Rec* __restrict__ hit_recs, uint32_t n_pairs,
             uint64_t pair_begin, PT P, const uint32_t* __restrict__ bitmap_pool,
             const uint64_t* __restrict__ table_pool, unsigned long long* __restrict__ counter,
             int3* __restrict__ out, const uint32_t* __restrict__ list, const unsigned long long* __restrict__ n_list)
{
    // list != nullptr: only the pairs list[0 .. *n_list) -- the ones k_walk_lanes handed over
    DevWarp w((int)(threadIdx.x & 31u));
    const unsigned long long n_todo = list ? *n_list : (unsigned long long)n_pairs;
    for (;;) {
        unsigned long long k = 0;
        if (w.lane == 0) k = atomicAdd(counter, 1ull);
        k = __shfl_sync(0xFFFFFFFFu, k, 0);
        if (k >= n_todo) return;
        const uint32_t pl = list ? list[k] : (uint32_t)k;
        const PairView pv = view_of(desc[pl], recs, bitmap_pool, table_pool);
        const uint32_t r0 = rec_off[word_off[pl]], r1 = rec_off[word_off[pl + 1]];
        const PairResult res = walk_pair(w, pv, P, hit_recs + r0, r1 - r0);
        if (w.lane == 0) out[pair_begin + pl] = make_int3(res.m, res.x, res.n);
    }
}

// S4, lane form (lane_walk.h): every LANE walks its own pairs -- LZ_LANE_SLOTS of them at a time, their contexts
// (21 words each) parked in shared memory.  Per round a lane takes up to LZ_LANE_RUN_STEPS plain events of one of
// its runnable pairs straight from the records (lane_run_step).  A pair that meets anything else waits: for the
// general step (lane_slow_step: seed and close matches, deferred anchors, long extensions, the tail) or for the
// seed scan of a window no record describes (lane_scan_step).  The warp enters such a step only once
// LZ_LANE_PARK_SLOW / LZ_LANE_PARK_SCAN lanes have a pair waiting for it (or nobody can run, or LZ_LANE_WAIT rounds
// have passed), so the expensive code runs with many active lanes -- and because a lane holds several pairs it
// usually still has a runnable one meanwhile.  Lanes pull new pairs as theirs finish (consecutive pairs of a
// chunk share their reference); a pair that lives in the general code (related genomes: hundreds of seed / close
// matches) is handed to the warp-per-pair walk, which does those steps 32 lanes wide.
constexpr int LANE_EMPTY = 4;            // slot without a pair (the chunk has run out)
constexpr int LANE_CTX_WORDS = 21;
struct LaneSlots {
    uint32_t w[LZ_LANE_SLOTS][LANE_CTX_WORDS][LZ_LANE_THREADS];
};
__device__ __forceinline__ void lane_ctx_store(LaneSlots& m, int j, const LaneCtx& L, uint32_t r0, uint32_t pl, int parks)
{
    uint32_t (*w)[LZ_LANE_THREADS] = m.w[j];
    const int t = threadIdx.x;
    w[0][t] = r0; w[1][t] = L.n; w[2][t] = L.c; w[3][t] = (uint32_t)L.n_loop;
    w[4][t] = (uint32_t)L.s.i; w[5][t] = (uint32_t)L.s.pred; w[6][t] = (uint32_t)L.s.lit;
    w[7][t] = (uint32_t)L.s.rs; w[8][t] = (uint32_t)L.s.re; w[9][t] = (uint32_t)L.s.kept_end;
    w[10][t] = (uint32_t)L.s.curM; w[11][t] = (uint32_t)L.s.curX;
    w[12][t] = (uint32_t)L.s.totM; w[13][t] = (uint32_t)L.s.totX; w[14][t] = (uint32_t)L.s.totN;
    w[15][t] = (uint32_t)L.spec_K1; w[16][t] = (uint32_t)L.spec_ks; w[17][t] = (uint32_t)L.spec_x;
    w[18][t] = (uint32_t)parks; w[19][t] = pl; w[20][t] = (uint32_t)L.chain;
}
__device__ __forceinline__ void lane_ctx_load(const LaneSlots& m, int j, LaneCtx& L, const Rec* hit_recs, uint32_t& r0, uint32_t& pl,
                                              int& parks, int mode)
{
    const uint32_t (*w)[LZ_LANE_THREADS] = m.w[j];
    const int t = threadIdx.x;
    r0 = w[0][t]; L.recs = hit_recs + r0; L.n = w[1][t]; L.c = w[2][t]; L.n_loop = (int)w[3][t];
    L.s.i = (int)w[4][t]; L.s.pred = (int)w[5][t]; L.s.lit = (int)w[6][t];
    L.s.rs = (int)w[7][t]; L.s.re = (int)w[8][t]; L.s.kept_end = (int)w[9][t];
    L.s.curM = (int)w[10][t]; L.s.curX = (int)w[11][t];
    L.s.totM = (int)w[12][t]; L.s.totX = (int)w[13][t]; L.s.totN = (int)w[14][t];
    L.spec_K1 = (int)w[15][t]; L.spec_ks = (int)w[16][t]; L.spec_x = (int)w[17][t];
    parks = (int)w[18][t]; pl = w[19][t]; L.chain = (int)w[20][t];
    L.mode = mode;
}

// Related genomes show in the hit records before any walking: an unrelated pair has about 32 * rho * (1 - rho) hit
// runs per 32 query positions, rho = the share of all mal-mers the reference holds.  A pair with LZ_LANE_DENSE_PCT %
// of that or more is a related one.
template <class PT>
__device__ __forceinline__ bool pair_is_dense(const PT& P, const PairDesc& pd, uint32_t nw, uint32_t n_rec)
{
    const float rho = fminf(1.0f, (float)pd.Dr * exp2f(-2.0f * (float)P.mal));
    return (float)n_rec * 100.0f > (float)LZ_LANE_DENSE_PCT * 32.0f * rho * (1.0f - rho) * (float)nw + 6400.0f;
}

// how many pairs of the chunk are related ones (decides whether k_walk_lanes hands them over, see LZ_LANE_KEEP_DEN)
template <class PT>
__global__ void __launch_bounds__(256)
k_count_dense(const PairDesc* __restrict__ desc, const uint32_t* __restrict__ word_off, const uint32_t* __restrict__ rec_off,
              uint32_t n_pairs, PT P, unsigned long long* __restrict__ n_dense)
{
    const uint32_t pl = blockIdx.x * 256u + threadIdx.x;
    bool dense = false;
    if (pl < n_pairs) {
        const uint32_t w0 = word_off[pl], w1 = word_off[pl + 1];
        dense = pair_is_dense(P, desc[pl], w1 - w0, rec_off[w1] - rec_off[w0]);
    }
    const int c = __syncthreads_count(dense);
    if (threadIdx.x == 0 && c) atomicAdd(n_dense, (unsigned long long)c);
}

template <class PT>
__global__ void __launch_bounds__(LZ_LANE_THREADS, LZ_LANE_MIN_BLOCKS)
k_walk_lanes(const Rec* __restrict__ recs, const PairDesc* __restrict__ desc, const uint32_t* __restrict__ word_off,
             const uint32_t* __restrict__ rec_off, const Rec* __restrict__ hit_recs, uint32_t n_pairs,
             uint64_t pair_begin, PT P, const uint64_t* __restrict__ table_pool, unsigned long long* __restrict__ counter,
             int3* __restrict__ out, uint32_t* __restrict__ hard_list, unsigned long long* __restrict__ hard_count,
             const unsigned long long* __restrict__ n_dense)
{
    __shared__ LaneSlots slots;
    // related pairs go to the warp-per-pair walk only while they are the exception in this chunk (the flag is read in
    // two rare places: shared memory, not a register of the hot loop)
    __shared__ int s_hand_over;
    if (threadIdx.x == 0) s_hand_over = *n_dense * (unsigned long long)LZ_LANE_KEEP_DEN <= (unsigned long long)n_pairs;
    __syncthreads();
    int mode[LZ_LANE_SLOTS];
#pragma unroll
    for (int j = 0; j < LZ_LANE_SLOTS; ++j) mode[j] = LANE_DONE;
    bool more = true;
    int w_slow = 0, w_scan = 0;           // rounds the parked pairs have waited
    for (;;) {
        // ---- new pairs into the slots whose pair is done
        bool live = false;
#pragma unroll
        for (int j = 0; j < LZ_LANE_SLOTS; ++j) {
            if (mode[j] == LANE_DONE) {
                mode[j] = LANE_EMPTY;
                if (more) {
                    const unsigned long long k = atomicAdd(counter, 1ull);
                    if (k < n_pairs) {
                        const uint32_t pl = (uint32_t)k;
                        const uint32_t r0 = rec_off[word_off[pl]], r1 = rec_off[word_off[pl + 1]];
                        const PairDesc pd = desc[pl];
                        // a related pair would only park its way to the hand-over: it goes to the warp-per-pair walk at once
                        if (s_hand_over && pair_is_dense(P, pd, word_off[pl + 1] - word_off[pl], r1 - r0)) {
                            hard_list[atomicAdd(hard_count, 1ull)] = pl;
                            mode[j] = LANE_DONE;        // the slot takes another pair next round
                            live = true;
                            continue;
                        }
                        LaneCtx L;
                        PairView pv{};
                        pv.Dq = pd.Dq;
                        lane_begin(L, pv, hit_recs + r0, r1 - r0, P.msl);
                        lane_ctx_store(slots, j, L, r0, pl, 0);
                        mode[j] = LANE_RUN;
                    } else more = false;
                }
            }
            live = live || mode[j] != LANE_EMPTY;
        }
        if (!__any_sync(0xFFFFFFFFu, live)) return;

        // ---- plain events of one runnable pair
        int jr = -1, js = -1, jc = -1;
#pragma unroll
        for (int j = LZ_LANE_SLOTS - 1; j >= 0; --j) if (mode[j] == LANE_RUN) jr = j;
        if (jr >= 0) {
            LaneCtx L;
            uint32_t r0, pl;
            int parks;
            lane_ctx_load(slots, jr, L, hit_recs, r0, pl, parks, LANE_RUN);
            RecAhead ahead;
            ahead.base = L.c;
#pragma unroll
            for (uint32_t k = 0; k < 4u; ++k) {
                ahead.r[k] = Rec{0x7FFFFFFFu, 0, 0, 0};
                if (L.c + k < L.n) ahead.r[k] = ld_rec(L.recs + L.c + k);
            }
#pragma unroll 1
            for (int r = 0; r < LZ_LANE_RUN_STEPS && L.mode == LANE_RUN; ++r) lane_run_step(L, P, ahead);
            lane_ctx_store(slots, jr, L, r0, pl, parks);
#pragma unroll
            for (int j = 0; j < LZ_LANE_SLOTS; ++j) if (j == jr) mode[j] = L.mode;
        }
        __syncwarp();
        bool has_run = false;
#pragma unroll
        for (int j = LZ_LANE_SLOTS - 1; j >= 0; --j) {
            has_run = has_run || mode[j] == LANE_RUN;
            if (mode[j] == LANE_SLOW) js = j;
        }
        bool any_run = __any_sync(0xFFFFFFFFu, has_run);

        // ---- the general step for the pairs waiting for it
        const int n_slow = __popc(__ballot_sync(0xFFFFFFFFu, js >= 0));
        if (n_slow && (n_slow >= LZ_LANE_PARK_SLOW || !any_run || ++w_slow >= LZ_LANE_WAIT)) {
            if (js >= 0) {
                LaneCtx L;
                uint32_t r0, pl;
                int parks;
                lane_ctx_load(slots, js, L, hit_recs, r0, pl, parks, LANE_SLOW);
                ++parks;
                if (parks > LZ_LANE_HARD_BASE + (L.n_loop >> LZ_LANE_HARD_SHIFT) && L.s.i < L.n_loop && s_hand_over) {
                    hard_list[atomicAdd(hard_count, 1ull)] = pl;       // k_walk_pairs restarts it
                    L.mode = LANE_DONE;
                } else {
                    L.pv = view_of(desc[pl], recs, nullptr, table_pool);
                    lane_slow_step(L, P);
                    if (L.mode == LANE_DONE) out[pair_begin + pl] = make_int3(L.s.totM, L.s.totX, L.s.totN);
                    else lane_ctx_store(slots, js, L, r0, pl, parks);
                }
#pragma unroll
                for (int j = 0; j < LZ_LANE_SLOTS; ++j) if (j == js) mode[j] = L.mode;
            }
            __syncwarp();
            w_slow = 0;
            has_run = false;
#pragma unroll
            for (int j = 0; j < LZ_LANE_SLOTS; ++j) has_run = has_run || mode[j] == LANE_RUN;
            any_run = __any_sync(0xFFFFFFFFu, has_run);
        }

        // ---- the seed scan for the pairs waiting for it
#pragma unroll
        for (int j = LZ_LANE_SLOTS - 1; j >= 0; --j) if (mode[j] == LANE_SCAN) jc = j;
        const int n_scan = __popc(__ballot_sync(0xFFFFFFFFu, jc >= 0));
        if (n_scan && (n_scan >= LZ_LANE_PARK_SCAN || !any_run || ++w_scan >= LZ_LANE_WAIT)) {
            if (jc >= 0) {
                LaneCtx L;
                uint32_t r0, pl;
                int parks;
                lane_ctx_load(slots, jc, L, hit_recs, r0, pl, parks, LANE_SCAN);
                ++parks;
                L.pv = view_of(desc[pl], recs, nullptr, table_pool);
                lane_scan_step(L, P);
                lane_ctx_store(slots, jc, L, r0, pl, parks);
#pragma unroll
                for (int j = 0; j < LZ_LANE_SLOTS; ++j) if (j == jc) mode[j] = L.mode;
            }
            __syncwarp();
            w_scan = 0;
        }
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


// ---- staged pipeline launches.  PT dispatch as launch_parse(): compile-time defaults or run-time values.
template <class F>
static cudaError_t with_prm(const Prm& P, F f)
{
    if (is_default_kernel_params(P)) return f(PrmDefault{P.mrd, P.mqd, P.reg});
    if (is_long_kernel_params(P)) return f(PrmLong{P.mrd, P.mqd, P.reg});
    return f(P);
}

int walk_blocks_per_sm(const Prm& P)
{
    int nb = 0;
    const cudaError_t e = is_default_kernel_params(P)
        ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_walk_pairs<PrmDefault>, LZ_WALK_THREADS, 0)
        : is_long_kernel_params(P)
        ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_walk_pairs<PrmLong>, LZ_WALK_THREADS, 0)
        : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_walk_pairs<Prm>, LZ_WALK_THREADS, 0);
    if (e != cudaSuccess || nb < 1) nb = 1;
    return nb;
}

int lane_blocks_per_sm(const Prm& P)
{
    int nb = 0;
    const cudaError_t e = is_default_kernel_params(P)
        ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_walk_lanes<PrmDefault>, LZ_LANE_THREADS, 0)
        : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_walk_lanes<Prm>, LZ_LANE_THREADS, 0);
    if (e != cudaSuccess || nb < 1) nb = 1;
    return nb;
}

size_t staged_scan_bytes(uint64_t n_items)
{
    size_t a = 0, b = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, a, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int64_t)n_items);
    auto it = thrust::make_transform_iterator((const uint32_t*)nullptr, RunCount{});
    cub::DeviceScan::ExclusiveSum(nullptr, b, it, (uint32_t*)nullptr, (int64_t)n_items);
    size_t r = 0;
    auto it64 = thrust::make_transform_iterator((const uint32_t*)nullptr, RunCount64{});
    cub::DeviceReduce::Sum(nullptr, r, it64, (unsigned long long*)nullptr, (int64_t)n_items);
    a = a > b ? a : b;
    return a > r ? a : r;
}

cudaError_t launch_staged_scan(cudaStream_t st, const StagedBufs& B, const Rec* recs, const SeqMeta* meta,
                               const uint32_t* row_ref, const RefIndex* row_idx, const uint64_t* row_off,
                               uint32_t n_rows, const uint32_t* q_ids, uint64_t pair_begin, uint32_t n_pairs,
                               uint64_t n_words, const Prm& P, const uint32_t* bitmap_pool, int scan_blocks,
                               const MatrixScan* mx)
{
    cudaError_t e = with_prm(P, [&](auto PT) {
        k_pair_desc<<<n_pairs / 256 + 1, 256, 0, st>>>(meta, row_ref, row_idx, row_off, n_rows, q_ids, pair_begin, n_pairs,
                                                         PT, B.desc, B.n_words);
        return cudaGetLastError();
    });
    if (e != cudaSuccess) return e;
    size_t tb = B.tmp_bytes;
    e = cub::DeviceScan::ExclusiveSum(B.tmp, tb, (const uint32_t*)B.n_words, B.word_off, (int64_t)n_pairs + 1, st);
    if (e != cudaSuccess) return e;
    if (n_words && mx) {
        // all2all form: blocks of <= 32 whole rows
        const uint32_t key_words = (uint32_t)((((uint64_t)1 << mx->bbits) + 31u) / 32u);
        const uint32_t key_mask = (uint32_t)(((uint64_t)1 << mx->bbits) - 1u);
        for (uint32_t i0 = 0; i0 < mx->n_rows_chunk; i0 += 32) {
            const uint32_t nr = mx->n_rows_chunk - i0 < 32u ? mx->n_rows_chunk - i0 : 32u;
            k_bitmap_matrix<<<(key_words + 255) / 256, 256, 0, st>>>(bitmap_pool, row_idx + mx->row_first + i0, nr, key_words, mx->mc);
            e = cudaGetLastError();
            if (e != cudaSuccess) return e;
            e = with_prm(P, [&](auto PT) {
                const dim3 grid((mx->max_words + 127) / 128, mx->n_seqs);
                k_scan_matrix<<<grid, 128, 0, st>>>(recs, meta, row_ref + mx->row_first + i0, row_off + mx->row_first + i0, nr,
                                                    pair_begin, B.word_off, PT, mx->mc, key_mask, mx->bbits, B.hm, B.wpair);
                return cudaGetLastError();
            });
            if (e != cudaSuccess) return e;
        }
    } else if (n_words) {
        e = with_prm(P, [&](auto PT) {
            using K = decltype(PT);
            const size_t smem = sizeof(uint32_t) << (LZ_COARSE_BITS - 5);
            // per device (a process may drive several GPUs, one context each): set on every launch
            cudaError_t ea = cudaFuncSetAttribute(k_scan_hits<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (ea != cudaSuccess) return ea;
            k_scan_hits<K><<<scan_blocks, LZ_SCAN_THREADS, smem, st>>>(recs, B.desc, B.word_off, n_pairs, PT, bitmap_pool, B.hm, B.wpair);
            return cudaGetLastError();
        });
        if (e != cudaSuccess) return e;
    }
    auto it = thrust::make_transform_iterator((const uint32_t*)B.hm, RunCount{});
    tb = B.tmp_bytes;
    return cub::DeviceScan::ExclusiveSum(B.tmp, tb, it, B.rec_off, (int64_t)n_words + 1, st);
}

cudaError_t launch_run_total(cudaStream_t st, const StagedBufs& B, uint64_t n_words)
{
    auto it = thrust::make_transform_iterator((const uint32_t*)B.hm, RunCount64{});
    size_t tb = B.tmp_bytes;
    return cub::DeviceReduce::Sum(B.tmp, tb, it, B.counter + 3, (int64_t)n_words, st);
}

cudaError_t launch_staged_walk(cudaStream_t st, int grid_blocks, const StagedBufs& B, const Rec* recs,
                               uint64_t pair_begin, uint32_t n_pairs, uint64_t n_words, uint64_t n_recs, const Prm& P,
                               const uint32_t* bitmap_pool, const uint64_t* table_pool, void* out, int lane_grid_blocks)
{
    if (n_words) {
        k_expand_hits<<<(unsigned)((n_words + 255) / 256), 256, 0, st>>>(B.hm, B.wpair, B.word_off, (uint32_t)n_words, B.rec_off, B.list);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    if (n_recs) {
        cudaError_t e = with_prm(P, [&](auto PT) {
            k_hit_probe<<<(unsigned)((n_recs + LZ_RECORD_THREADS - 1) / LZ_RECORD_THREADS), LZ_RECORD_THREADS, 0, st>>>(
                recs, B.desc, B.list, B.rec_off + n_words, PT, bitmap_pool, table_pool, B.recs);
            cudaError_t e1 = cudaGetLastError();
            if (e1 != cudaSuccess) return e1;
            if constexpr (decltype(PT)::kSpecSeeds) if (PT.mqd <= SPEC_MAX_K)
                k_hit_spec<<<(unsigned)((n_recs + LZ_SPEC_JOBS - 1) / LZ_SPEC_JOBS), LZ_SPEC_THREADS, 0, st>>>(
                    recs, B.desc, B.list, B.rec_off + n_words, PT, B.recs);
            return cudaGetLastError();
        });
        if (e != cudaSuccess) return e;
    }
    // counter[0]: pairs pulled by the first walk kernel, [1]: by the second, [2]: pairs handed over, [4]: related pairs
    // of the chunk ([3] is launch_run_total's)
    cudaError_t e = cudaMemsetAsync(B.counter, 0, 3 * sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(B.counter + 4, 0, sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    return with_prm(P, [&](auto PT) {
        bool lanes = false;
        if constexpr (decltype(PT)::kSpecSeeds) lanes = lane_grid_blocks > 0 && PT.mqd <= SPEC_MAX_K;
        if constexpr (decltype(PT)::kSpecSeeds) if (lanes) {
            // lane per pair; the few pairs that live in the general code are handed to the warp-per-pair walk
            k_count_dense<<<n_pairs / 256 + 1, 256, 0, st>>>(B.desc, B.word_off, B.rec_off, n_pairs, PT, B.counter + 4);
            const cudaError_t e0 = cudaGetLastError();
            if (e0 != cudaSuccess) return e0;
            k_walk_lanes<<<lane_grid_blocks, LZ_LANE_THREADS, 0, st>>>(recs, B.desc, B.word_off, B.rec_off, B.recs, n_pairs, pair_begin,
                                                                     PT, table_pool, B.counter, reinterpret_cast<int3*>(out), B.hard,
                                                                     B.counter + 2, B.counter + 4);
            const cudaError_t e1 = cudaGetLastError();
            if (e1 != cudaSuccess) return e1;
            k_walk_pairs<<<grid_blocks, LZ_WALK_THREADS, 0, st>>>(recs, B.desc, B.word_off, B.rec_off, B.recs, n_pairs, pair_begin,
                                                                PT, bitmap_pool, table_pool, B.counter + 1, reinterpret_cast<int3*>(out),
                                                                B.hard, B.counter + 2);
        }
        if (!lanes)
            k_walk_pairs<<<grid_blocks, LZ_WALK_THREADS, 0, st>>>(recs, B.desc, B.word_off, B.rec_off, B.recs, n_pairs, pair_begin,
                                                                PT, bitmap_pool, table_pool, B.counter, reinterpret_cast<int3*>(out),
                                                                nullptr, nullptr);
        return cudaGetLastError();
    });
}

}  // namespace lzani
