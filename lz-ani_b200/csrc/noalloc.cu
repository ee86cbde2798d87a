// noalloc.cu -- lzani_filter_dense_noalloc(): the --out-filter pre-filter (postproc.cu, F1-F3) over a DENSE
// all2all result matrix the caller already holds on the device, for callers that may not allocate: code
// under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.  No existing code path goes through this file.
//
// The three kernels need scratch (one flag byte per ordered pair, one count and one offset per block of
// 256 pairs).  lzani_filter_pairs() takes it from cudaMalloc; here it is ONE file-scope __device__ array of
// sizeof(Scratch) bytes (about 4.4 MB, sized for LZANI_NOALLOC_MAX_SEQS sequences) that every GPU which
// loads the library holds until the process exits.  The entry point calls no cudaMalloc / cudaFree and no
// allocating library routine; it only launches three kernels on the caller's stream.
//
//   * Every call on a device shares that array: calls that overlap on the GPU (two streams, two threads)
//     race.  Serialise them.
//   * The array keeps the last call's contents.  Nothing here relies on them: every word a kernel reads
//     (flags[0..n_pairs), blk_cnt[0..nb), blk_off[0..nb]) is written earlier in the SAME call, and
//     *d_n_out is stored, not accumulated -- so nothing has to be cleared.
//   * Host code never uses the array's name as an address (there it is a host-side stand-in): the device
//     address comes from cudaGetSymbolAddress, once per device, with that device current.
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <mutex>

#include "../../include/lzani_b200.h"
#include "outfilter.h"

namespace lzani {
namespace {

constexpr int NA_FT = 256;                                                     // threads (= pairs) per block
constexpr uint64_t NA_MAX_PAIRS = (uint64_t)LZANI_NOALLOC_MAX_SEQS * (LZANI_NOALLOC_MAX_SEQS - 1);
constexpr uint64_t NA_MAX_BLOCKS = (NA_MAX_PAIRS + NA_FT - 1) / NA_FT;
constexpr int NA_MAX_DEVICES = 64;

struct Scratch {
    unsigned long long blk_off[NA_MAX_BLOCKS + 1];   // exclusive scan of blk_cnt, [nb] = total
    uint32_t blk_cnt[NA_MAX_BLOCKS];                 // survivors per block
    uint8_t flags[NA_MAX_PAIRS];                     // out_filter_mask of every ordered pair (0 for a > b)
};

static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(Scratch)];

struct NaPair { uint32_t a, b; uint64_t rev; };

// ordered pair k of the dense matrix: row a = k / (n-1), the j-th id != a
__device__ __forceinline__ NaPair na_locate(uint32_t n, uint64_t k)
{
    NaPair p;
    p.a = (uint32_t)(k / (n - 1));
    const uint32_t j = (uint32_t)(k - (uint64_t)p.a * (n - 1));
    p.b = j + (j >= p.a ? 1u : 0u);
    p.rev = (uint64_t)p.b * (n - 1) + p.a;           // a < b: a is entry a of row b
    return p;
}

__global__ void __launch_bounds__(NA_FT)
k_na_flags(uint32_t n, uint64_t n_pairs, const int3* __restrict__ res, const uint32_t* __restrict__ eff_len,
           OutFilter flt, Scratch* __restrict__ S)
{
    const uint64_t k = (uint64_t)blockIdx.x * NA_FT + threadIdx.x;
    uint32_t mask = 0;
    if (k < n_pairs) {
        const NaPair p = na_locate(n, k);
        if (p.a < p.b) {
            const int3 ab = res[k], ba = res[p.rev];
            mask = out_filter_mask(flt, DirStats{ab.x, ab.y, ab.z}, DirStats{ba.x, ba.y, ba.z}, eff_len[p.a], eff_len[p.b]);
        }
        S->flags[k] = (uint8_t)mask;
    }
    const int c = __syncthreads_count(mask != 0);
    if (threadIdx.x == 0) S->blk_cnt[blockIdx.x] = (uint32_t)c;
}

// one block: exclusive scan of blk_cnt[0..nb) -> blk_off[0..nb], the total also to *n_out
__global__ void __launch_bounds__(1024)
k_na_scan(uint64_t nb, Scratch* __restrict__ S, unsigned long long* __restrict__ n_out)
{
    __shared__ unsigned long long warp_sum[32];
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    if (reinterpret_cast<const unsigned char*>(S) != g_noalloc_scratch) {   // the host passed something else than
        if (threadIdx.x == 0) *n_out = ~0ull;                               // this device's copy of the array
        return;
    }
    const uint64_t per = (nb + 1023) / 1024;
    const uint64_t lo = (uint64_t)threadIdx.x * per < nb ? (uint64_t)threadIdx.x * per : nb;
    const uint64_t hi = lo + per < nb ? lo + per : nb;
    unsigned long long s = 0;
    for (uint64_t i = lo; i < hi; ++i) s += S->blk_cnt[i];
    unsigned long long inc = s;                      // inclusive scan inside the warp
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long v = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= (uint32_t)d) inc += v;
    }
    if (lane == 31) warp_sum[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        unsigned long long w = warp_sum[lane], winc = w;
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long v = __shfl_up_sync(0xFFFFFFFFu, winc, d);
            if (lane >= (uint32_t)d) winc += v;
        }
        warp_sum[lane] = winc - w;                   // exclusive base of warp `lane`
        if (lane == 31) { S->blk_off[nb] = winc; *n_out = winc; }
    }
    __syncthreads();
    unsigned long long run = warp_sum[wid] + inc - s;
    for (uint64_t i = lo; i < hi; ++i) { S->blk_off[i] = run; run += S->blk_cnt[i]; }
}

__global__ void __launch_bounds__(NA_FT)
k_na_emit(uint32_t n, uint64_t n_pairs, const int3* __restrict__ res, const Scratch* __restrict__ S,
          lzani_pair_t* __restrict__ out, unsigned long long cap)
{
    __shared__ uint32_t warp_base[NA_FT / 32];
    const uint64_t k = (uint64_t)blockIdx.x * NA_FT + threadIdx.x;
    const uint32_t mask = k < n_pairs ? S->flags[k] : 0u;
    const uint32_t bal = __ballot_sync(0xFFFFFFFFu, mask != 0);
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    if (lane == 0) warp_base[wid] = (uint32_t)__popc(bal);
    __syncthreads();
    if (threadIdx.x < 32) {                          // exclusive scan of the 8 warp counts
        const uint32_t c = lane < NA_FT / 32 ? warp_base[lane] : 0u;
        uint32_t inc = c;
        for (int d = 1; d < NA_FT / 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, inc, d);
            if (lane >= (uint32_t)d) inc += v;
        }
        if (lane < NA_FT / 32) warp_base[lane] = inc - c;
    }
    __syncthreads();
    if (!mask) return;
    const unsigned long long dst = S->blk_off[blockIdx.x] + warp_base[wid] + (uint32_t)__popc(bal & ((1u << lane) - 1u));
    if (dst >= cap) return;                          // *d_n_out still reports the full count
    const NaPair p = na_locate(n, k);
    const int3 ab = res[k], ba = res[p.rev];
    lzani_pair_t o;
    o.a = p.a; o.b = p.b;
    o.ab.sym_in_matches = ab.x; o.ab.sym_in_literals = ab.y; o.ab.no_components = ab.z;
    o.ba.sym_in_matches = ba.x; o.ba.sym_in_literals = ba.y; o.ba.no_components = ba.z;
    o.lines = mask;
    out[dst] = o;
}

// device address of g_noalloc_scratch per device, looked up once with that device current
std::mutex g_na_mutex;
Scratch* g_na_ptr[NA_MAX_DEVICES] = {};

struct ErrOut {                                      // the caller's message buffer (may be absent)
    char* buf; size_t len;
    int fail(int code, const char* what, cudaError_t e = cudaSuccess) const
    {
        if (buf && len) {
            if (e != cudaSuccess) snprintf(buf, len, "lzani_filter_dense_noalloc: %s: %s", what, cudaGetErrorString(e));
            else snprintf(buf, len, "lzani_filter_dense_noalloc: %s", what);
        }
        return code;
    }
};

struct DeviceGuard {                                 // restores the caller's current device
    int prev = -1;
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace
}  // namespace lzani

using namespace lzani;

extern "C" int lzani_filter_dense_noalloc(int device, uint32_t n_seqs, const lzani_result_t* d_res,
                                          const uint32_t* d_eff_len, const lzani_out_filter_t* flt,
                                          lzani_pair_t* d_pairs, uint64_t cap, uint64_t* d_n_out, void* stream,
                                          char* err, size_t err_len)
{
    const ErrOut E{err, err_len};
    if (err && err_len) err[0] = 0;
    if (!flt || !d_n_out) return E.fail(LZANI_ERR_ARG, "flt and d_n_out must not be NULL");
    if (n_seqs > LZANI_NOALLOC_MAX_SEQS) {
        char msg[160];
        snprintf(msg, sizeof msg, "%u sequences exceed LZANI_NOALLOC_MAX_SEQS = %d (the static scratch array is sized for "
                 "that many; use lzani_filter_pairs)", n_seqs, (int)LZANI_NOALLOC_MAX_SEQS);
        return E.fail(LZANI_ERR_ARG, msg);
    }
    const uint64_t n_pairs = n_seqs > 1 ? (uint64_t)n_seqs * (n_seqs - 1) : 0;
    if (n_pairs && (!d_res || !d_eff_len)) return E.fail(LZANI_ERR_ARG, "d_res and d_eff_len must not be NULL");
    if (cap && !d_pairs) return E.fail(LZANI_ERR_ARG, "d_pairs is NULL but cap > 0");
    if (device < 0 || device >= NA_MAX_DEVICES) return E.fail(LZANI_ERR_ARG, "device id out of range");

    DeviceGuard guard;
    cudaError_t e = cudaGetDevice(&guard.prev);
    if (e != cudaSuccess) { guard.prev = -1; return E.fail(LZANI_ERR_CUDA, "cudaGetDevice", e); }
    if (guard.prev == device) guard.prev = -1;
    else if ((e = cudaSetDevice(device)) != cudaSuccess) return E.fail(LZANI_ERR_CUDA, "cudaSetDevice", e);

    Scratch* S;
    {
        std::lock_guard<std::mutex> lock(g_na_mutex);
        if (!g_na_ptr[device]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_noalloc_scratch);     // `device` is current here
            if (e != cudaSuccess || !p) return E.fail(LZANI_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch)", e);
            g_na_ptr[device] = static_cast<Scratch*>(p);
        }
        S = g_na_ptr[device];
    }

    const cudaStream_t st = static_cast<cudaStream_t>(stream);
    const uint64_t nb = (n_pairs + NA_FT - 1) / NA_FT;
    const OutFilter f{flt->gani, flt->ani, flt->tani, flt->qcov, flt->rcov};
    const int3* res = reinterpret_cast<const int3*>(d_res);
    if (nb) {
        k_na_flags<<<(unsigned)nb, NA_FT, 0, st>>>(n_seqs, n_pairs, res, d_eff_len, f, S);
        if ((e = cudaGetLastError()) != cudaSuccess) return E.fail(LZANI_ERR_CUDA, "k_na_flags", e);
    }
    k_na_scan<<<1, 1024, 0, st>>>(nb, S, reinterpret_cast<unsigned long long*>(d_n_out));
    if ((e = cudaGetLastError()) != cudaSuccess) return E.fail(LZANI_ERR_CUDA, "k_na_scan", e);
    if (nb && cap) {
        k_na_emit<<<(unsigned)nb, NA_FT, 0, st>>>(n_seqs, n_pairs, res, S, d_pairs, (unsigned long long)cap);
        if ((e = cudaGetLastError()) != cudaSuccess) return E.fail(LZANI_ERR_CUDA, "k_na_emit", e);
    }
    return LZANI_OK;
}
