// simt.h -- the warp abstraction the parse kernel is written against.
//
// The kernel logic (parse_core.h) is warp-synchronous: uniform control flow, per-lane data
// in `Lanes<T>` variables, and explicit collectives (ballot / redux / shuffle / scan).
// Two policies implement that vocabulary:
//
//   DevWarp   -- the product: one CUDA warp; Lanes<T> is one register per lane, LZ_EACH runs
//                its body once (for this thread's lane), collectives are __ballot_sync,
//                __reduce_*_sync (REDUX on sm_100a), __shfl_sync.
//   HostWarp  -- TEST ONLY (tests/emul): 32 lanes emulated by a loop so that the very same
//                kernel source can be checked against the oracle on a machine without a GPU.
//                It is compiled only by tests/emul/Makefile and never linked into the library.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define LZ_HD __host__ __device__ __forceinline__
#define LZ_D __device__ __forceinline__
// rarely executed warp-cooperative routines: kept out of line so that the hot loop of the
// parse kernel stays inside the instruction cache
#define LZ_COLD __host__ __device__ __noinline__
// the parse kernel is bound by instruction issue AND sits at the edge of the instruction cache:
// loops outside the innermost bit loops are kept rolled
#define LZ_ROLLED _Pragma("unroll 1")
#else
#define LZ_ROLLED
#define LZ_HD inline
#define LZ_D inline
#define LZ_COLD inline
#endif

namespace lzani {

// ---- scalar bit helpers usable on both sides ----
LZ_HD int popc32(uint32_t x)
{
#if defined(__CUDA_ARCH__)
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}
LZ_HD int ctz32(uint32_t x)  // x != 0
{
#if defined(__CUDA_ARCH__)
    return __ffs((int)x) - 1;
#else
    return __builtin_ctz(x);
#endif
}
LZ_HD int ctz64(uint64_t x)  // x != 0
{
#if defined(__CUDA_ARCH__)
    return __ffsll((long long)x) - 1;
#else
    return __builtin_ctzll(x);
#endif
}
LZ_HD int popc64(uint64_t x)
{
#if defined(__CUDA_ARCH__)
    return __popcll(x);
#else
    return __builtin_popcountll(x);
#endif
}
LZ_HD int clz32(uint32_t x)  // x != 0
{
#if defined(__CUDA_ARCH__)
    return __clz((int)x);
#else
    return __builtin_clz(x);
#endif
}
LZ_HD uint32_t brev32(uint32_t x)
{
#if defined(__CUDA_ARCH__)
    return __brev(x);
#else
    x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
    x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
    x = ((x >> 4) & 0x0F0F0F0Fu) | ((x & 0x0F0F0F0Fu) << 4);
    x = ((x >> 8) & 0x00FF00FFu) | ((x & 0x00FF00FFu) << 8);
    return (x >> 16) | (x << 16);
#endif
}
// 32 bits starting at bit `sh` (0..31) of the 64-bit value hi:lo
LZ_HD uint32_t funnel_r(uint32_t lo, uint32_t hi, int sh)
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, (uint32_t)sh);
#else
    return sh ? (lo >> sh) | (hi << (32 - sh)) : lo;
#endif
}
LZ_HD uint32_t low_mask32(int n)  // n in [0,32]
{
    return n >= 32 ? 0xFFFFFFFFu : ((1u << n) - 1u);
}
LZ_HD uint64_t low_mask64(int n)  // n in [0,64]
{
    return n >= 64 ? ~0ull : ((1ull << n) - 1ull);
}

#if defined(__CUDACC__)
// ------------------------------------------------------------------------------------------
struct DevWarp {
    static constexpr bool kHost = false;
    int lane;
    template <class T>
    struct Lanes {
        T v;
        __device__ __forceinline__ T& operator[](int) { return v; }
        __device__ __forceinline__ const T& operator[](int) const { return v; }
    };
    __device__ __forceinline__ explicit DevWarp(int l) : lane(l) {}
    __device__ __forceinline__ int first() const { return lane; }
    __device__ __forceinline__ int end() const { return lane + 1; }

    __device__ __forceinline__ uint32_t ballot(const Lanes<int>& p) const { return __ballot_sync(0xFFFFFFFFu, p.v); }
    __device__ __forceinline__ int min_i32(const Lanes<int>& x) const { return __reduce_min_sync(0xFFFFFFFFu, x.v); }
    __device__ __forceinline__ int max_i32(const Lanes<int>& x) const { return __reduce_max_sync(0xFFFFFFFFu, x.v); }
    __device__ __forceinline__ int add_i32(const Lanes<int>& x) const { return __reduce_add_sync(0xFFFFFFFFu, x.v); }
    __device__ __forceinline__ uint32_t min_u32(const Lanes<uint32_t>& x) const { return __reduce_min_sync(0xFFFFFFFFu, x.v); }
    __device__ __forceinline__ uint32_t or_u32(const Lanes<uint32_t>& x) const { return __reduce_or_sync(0xFFFFFFFFu, x.v); }
    __device__ __forceinline__ uint64_t or_u64(const Lanes<uint64_t>& x) const
    {
        const uint32_t lo = __reduce_or_sync(0xFFFFFFFFu, (uint32_t)x.v);
        const uint32_t hi = __reduce_or_sync(0xFFFFFFFFu, (uint32_t)(x.v >> 32));
        return ((uint64_t)hi << 32) | lo;
    }
    // orders this warp's shared-memory scratch writes before its later reads
    __device__ __forceinline__ void sync() const { __syncwarp(); }
    // hint: the line at p will be read soon
    __device__ __forceinline__ void prefetch(const void* p) const { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
    // value of lane `src` (uniform src)
    template <class T>
    __device__ __forceinline__ T get(const Lanes<T>& x, int src) const { return __shfl_sync(0xFFFFFFFFu, x.v, src); }
    // per-lane source lane
    template <class T>
    __device__ __forceinline__ void gather(Lanes<T>& dst, const Lanes<T>& x, const Lanes<int>& src) const
    {
        dst.v = __shfl_sync(0xFFFFFFFFu, x.v, src.v);
    }
    // dst[lane] = x[lane-1], lane 0 gets `fill`
    template <class T>
    __device__ __forceinline__ void shift_up(Lanes<T>& dst, const Lanes<T>& x, T fill) const
    {
        T t = __shfl_up_sync(0xFFFFFFFFu, x.v, 1);
        dst.v = lane == 0 ? fill : t;
    }
    // exclusive prefix sum over lanes
    __device__ __forceinline__ void exscan_add(Lanes<int>& dst, const Lanes<int>& x) const
    {
        int v = x.v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(0xFFFFFFFFu, v, d);
            if (lane >= d) v += t;
        }
        dst.v = v - x.v;
    }
};
#endif

#if !defined(__CUDA_ARCH__)
// ------------------------------------------------------------------------------------------
// TEST ONLY: sequential emulation of the 32 lanes.
struct HostWarp {
    static constexpr bool kHost = true;
    template <class T>
    struct Lanes {
        T v[32];
        T& operator[](int l) { return v[l]; }
        const T& operator[](int l) const { return v[l]; }
    };
    int first() const { return 0; }
    int end() const { return 32; }
    uint32_t ballot(const Lanes<int>& p) const
    {
        uint32_t m = 0;
        for (int l = 0; l < 32; ++l) m |= (p.v[l] ? 1u : 0u) << l;
        return m;
    }
    int min_i32(const Lanes<int>& x) const { int r = x.v[0]; for (int l = 1; l < 32; ++l) r = x.v[l] < r ? x.v[l] : r; return r; }
    int max_i32(const Lanes<int>& x) const { int r = x.v[0]; for (int l = 1; l < 32; ++l) r = x.v[l] > r ? x.v[l] : r; return r; }
    int add_i32(const Lanes<int>& x) const { int r = 0; for (int l = 0; l < 32; ++l) r += x.v[l]; return r; }
    uint32_t min_u32(const Lanes<uint32_t>& x) const { uint32_t r = x.v[0]; for (int l = 1; l < 32; ++l) r = x.v[l] < r ? x.v[l] : r; return r; }
    uint32_t or_u32(const Lanes<uint32_t>& x) const { uint32_t r = 0; for (int l = 0; l < 32; ++l) r |= x.v[l]; return r; }
    uint64_t or_u64(const Lanes<uint64_t>& x) const { uint64_t r = 0; for (int l = 0; l < 32; ++l) r |= x.v[l]; return r; }
    void sync() const {}
    void prefetch(const void*) const {}
    template <class T>
    T get(const Lanes<T>& x, int src) const { return x.v[src & 31]; }
    template <class T>
    void gather(Lanes<T>& dst, const Lanes<T>& x, const Lanes<int>& src) const
    {
        Lanes<T> t = x;
        for (int l = 0; l < 32; ++l) dst.v[l] = t.v[src.v[l] & 31];
    }
    template <class T>
    void shift_up(Lanes<T>& dst, const Lanes<T>& x, T fill) const
    {
        Lanes<T> t = x;
        dst.v[0] = fill;
        for (int l = 1; l < 32; ++l) dst.v[l] = t.v[l - 1];
    }
    void exscan_add(Lanes<int>& dst, const Lanes<int>& x) const
    {
        int acc = 0;
        for (int l = 0; l < 32; ++l) { int t = x.v[l]; dst.v[l] = acc; acc += t; }
    }
};
#endif

}  // namespace lzani

// Runs its body for "every lane this thread stands for": once on the device, 32x on the host.
#define LZ_EACH(w, ln) for (int ln = (w).first(), ln##_end = (w).end(); ln < ln##_end; ++ln)
