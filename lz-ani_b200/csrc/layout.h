// layout.h -- how sequences and reference indices live in HBM, and the per-element functions
// that build them.  Shared by the CUDA kernels (kernels.cu) and the test-only host emulation.
//
// SEQUENCE RECORDS.  Every 32 consecutive bases are one 16-byte record (one 128-bit load):
//     .x = low  bit of the 2-bit code of each base (bit j = base j of the record)
//     .y = high bit of the 2-bit code
//     .z = N mask (1 = not ACGT: an input N/IUPAC symbol or padding)
//     .w = 0
// i.e. the 2-bit code is stored as two bit planes so that a 32-base comparison is
// (x^x') | (y^y') | z | z' with no bit compaction.  A mismatch word has bit j = 1 where base j
// differs; N / padding never equals anything (reference: code_N_ref 4 != code_N_seq 5,
// defs.h:28-30, parser.cpp:20-24, 41-42).
//
// Each sequence s is stored twice:
//   Q-form: the sequence as a QUERY, seq_data of CParser::prepare_data (parser.cpp:37-50):
//           Lq bases, then N forever (covers the mrd-long N tail; Dq = Lq + mrd is a length
//           bound kept in registers).
//   R-form: the sequence as a REFERENCE, seq_ref of CParser::prepare_reference
//           (parser.cpp:16-34): fwd | 2*mrd N | reverse complement | mrd N,  Dr = 2*Lr + 3*mrd.
// Both are followed by LZ_PAD_RECS all-N records so that any 64-base window starting below the
// end can be loaded without a bounds test.
//
// REFERENCE INDEX (replaces ht_long of prepare_ht_long, parser.cpp:146-189; the seed table
// ht_short of prepare_ht_short :106-143 has no stored counterpart -- seeds are found by direct
// comparison along diagonals, see parse_core.h):
//   presence bitmap: 2^bbits bits; bit hash_b(key) is set iff some valid mal-mer of the R-form
//                    has that key (direct-addressed when 2*mal <= bbits: exact).
//   anchor table:    2^tbits 64-bit slots, open addressing / linear probing, slot =
//                    (tag << 32) | position, EMPTY = ~0.  All positions of one k-mer sit in one
//                    probe chain; their order is irrelevant because the parse takes the
//                    longest match and breaks ties by smallest position explicitly.
#pragma once
#include "simt.h"

namespace lzani {

constexpr int LZ_PAD_RECS = 3;
constexpr uint64_t LZ_EMPTY = ~0ull;

struct Rec { uint32_t x, y, z, w; };   // same layout as uint4

struct Prm {
    int mal, msl, mrd, mqd, reg, aw, am, ar;
    static constexpr bool kSpecSeeds = false;   // speculative per-hit seed scan: msl == 7 kernels only
};
// The reference's default k-mer / extension parameters (params.h:38-45) as compile-time
// constants: the parse kernel is instantiated once for them (bit loops unroll, masks fold) and
// once for arbitrary run-time values (Prm).  The kernel code only says P.mal, P.msl, ...
struct PrmDefault {
    static constexpr int mal = 11, msl = 7, aw = 15, am = 7, ar = 3;
    static constexpr bool kSpecSeeds = true;
    int mrd, mqd, reg;
};
LZ_HD bool is_default_kernel_params(const Prm& p) { return p.mal == 11 && p.msl == 7 && p.aw == 15 && p.am == 7 && p.ar == 3; }
// The long-genome parameters BASELINE configs[4] names for bacterial sets (--mal 15 --msl 9 --aw 20 --am 9 --ar 4),
// also compile-time for the staged kernels: the run tests of the seed scan (msl) and the extension windows unroll.
struct PrmLong {
    static constexpr int mal = 15, msl = 9, aw = 20, am = 9, ar = 4;
    static constexpr bool kSpecSeeds = false;
    int mrd, mqd, reg;
};
LZ_HD bool is_long_kernel_params(const Prm& p) { return p.mal == 15 && p.msl == 9 && p.aw == 20 && p.am == 9 && p.ar == 4; }

// per-sequence placement in the record pool
struct SeqMeta {
    uint64_t q_rec;    // first record of the Q-form
    uint64_t r_rec;    // first record of the R-form
    uint64_t src_off;  // byte offset of the three_in_byte source
    uint32_t len;      // bases
    uint32_t pad_;
};

// per-reference placement of its index inside the current batch
struct RefIndex {
    uint64_t bitmap_word;  // first 32-bit word of the presence bitmap
    uint64_t table_slot;   // first slot of the anchor table
    uint32_t bbits;        // log2(bitmap bits)
    uint32_t tbits;        // log2(table slots)
};

// staged parse: one element of a chunk's hit list (a run of consecutive anchor-bitmap hits of one pair)
struct HitEntry { uint32_t pair, pr; };          // pair (chunk-relative), position | run length << 29
constexpr int HIT_ENTRY_RUN_SHIFT = 29;
constexpr int LZ_STAGED_MAX_POS = 1 << HIT_ENTRY_RUN_SHIFT;

LZ_HD int q_form_recs(int len, int mrd) { return (len + mrd + 31) / 32 + LZ_PAD_RECS; }
LZ_HD int r_form_len(int len, int mrd) { return 2 * len + 3 * mrd; }
LZ_HD int r_form_recs(int len, int mrd) { return (r_form_len(len, mrd) + 31) / 32 + LZ_PAD_RECS; }

LZ_HD int ceil_log2_u64(uint64_t v)
{
    int b = 0;
    while (((uint64_t)1 << b) < v) ++b;
    return b;
}
// bitmap: exact (direct-addressed, 4^mal bits) whenever the key space is at most 2^24 bits (2 MiB;
// 512 KiB for the default mal = 11) -- a compile-time fact for the default-parameter kernel --
// else hashed at ~1/32 occupancy
constexpr int LZ_DIRECT_BITMAP_MAX_BITS = 24;
LZ_HD uint32_t index_bbits(int Dr, int mal)
{
    if (2 * mal <= LZ_DIRECT_BITMAP_MAX_BITS) return (uint32_t)(2 * mal);
    int want = ceil_log2_u64((uint64_t)(Dr > 1 ? Dr : 2)) + 5;
    if (want < 10) want = 10;
    return (uint32_t)want;
}
// table: load factor <= 0.4
LZ_HD uint32_t index_tbits(int Dr)
{
    int b = ceil_log2_u64((uint64_t)(Dr > 1 ? Dr : 2) * 5 / 2);
    return (uint32_t)(b < 6 ? 6 : b);
}

LZ_HD Rec ld_rec(const Rec* p)
{
#if defined(__CUDA_ARCH__)
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
    Rec r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w;
    return r;
#else
    return *p;
#endif
}

// the three planes of the 32 bases starting at pos (pos >= 0, pos below the padded end)
LZ_HD void ext3(const Rec* s, int pos, uint32_t& lo, uint32_t& hi, uint32_t& nm)
{
    const int w = pos >> 5, sh = pos & 31;
    const Rec a = ld_rec(s + w), b = ld_rec(s + w + 1);
    lo = funnel_r(a.x, b.x, sh);
    hi = funnel_r(a.y, b.y, sh);
    nm = funnel_r(a.z, b.z, sh);
}

// mismatch word of q[qpos .. qpos+32) against r[rpos .. rpos+32); past either end = mismatch
LZ_HD uint32_t mism32(const Rec* q, int Dq, int qpos, const Rec* r, int Dr, int rpos)
{
    if (qpos >= Dq || rpos >= Dr) return 0xFFFFFFFFu;
    uint32_t ql, qh, qn, rl, rh, rn;
    ext3(q, qpos, ql, qh, qn);
    ext3(r, rpos, rl, rh, rn);
    return (ql ^ rl) | (qh ^ rh) | qn | rn;
}

// k-mer key (k <= 32) from extracted planes: low-plane bits in [0,k), high-plane bits in [k,2k)
LZ_HD uint64_t kmer_key(uint32_t lo, uint32_t hi, int k)
{
    const uint32_t km = low_mask32(k);
    return (uint64_t)(lo & km) | ((uint64_t)(hi & km) << k);
}

LZ_HD uint64_t mix64(uint64_t key, uint64_t mul) { return key * mul; }

LZ_HD uint32_t bitmap_index(uint64_t key, int mal, uint32_t bbits)
{
    if (2 * mal <= LZ_DIRECT_BITMAP_MAX_BITS) return (uint32_t)key;     // bbits == 2*mal (index_bbits)
    return (uint32_t)(mix64(key, 0xD6E8FEB86659FD93ull) >> (64 - bbits));
}
LZ_HD uint32_t table_home(uint64_t key, uint32_t tbits)
{
    return (uint32_t)(mix64(key, 0x9E3779B97F4A7C15ull) >> (64 - tbits));
}
// 32-bit tag stored next to the position: the key itself when it fits (exact), else hash bits
LZ_HD uint32_t table_tag(uint64_t key, int mal)
{
    if (mal <= 16) return (uint32_t)key;
    return (uint32_t)(mix64(key, 0xA24BAED4963EE407ull) >> 32);
}

// ------------------------------------------------------------------------------------------
// K1 element: one output record of the Q-form or R-form of one sequence, decoded from the
// reservoir's three_in_byte bytes (seq_reservoir.cpp:39-57; byte = 36a+6b+c).
LZ_HD uint32_t sym_at(const uint8_t* src, uint32_t j)
{
    const uint32_t b = src[j / 3u];
    const uint32_t r = j % 3u;
    return r == 0 ? b / 36u : (r == 1 ? (b / 6u) % 6u : b % 6u);
}

// symbol sources: the reservoir's base-6 triples, or the loader's one-byte-per-base text as
// CSeqReservoir::append receives it (seq_reservoir.cpp:18-87; dna_code :243-247: ACGT in either case
// = 0..3, every other byte -- N, IUPAC codes, the byte-5 contig separator of load_fasta :110-127 -- = N)
struct SymPacked3 {
    const uint8_t* src;
    LZ_HD uint32_t operator()(uint32_t j) const { return sym_at(src, j); }
};
struct SymAscii {
    const uint8_t* src;
    LZ_HD uint32_t operator()(uint32_t j) const
    {
        const uint32_t ch = (uint32_t)src[j] | 0x20u;
        return ch == 'a' ? 0u : ch == 'c' ? 1u : ch == 'g' ? 2u : ch == 't' ? 3u : 5u;
    }
};

template <class Sym>
LZ_HD Rec pack_q_record_from(Sym sym, uint32_t len, uint32_t rec)
{
    Rec o; o.x = o.y = 0; o.z = 0; o.w = 0;
    const uint32_t base = rec * 32u;
    for (uint32_t j = 0; j < 32u; ++j) {
        const uint32_t p = base + j;
        uint32_t c = 5;
        if (p < len) c = sym(p);
        if (c < 4) { o.x |= (c & 1u) << j; o.y |= (c >> 1) << j; }
        else o.z |= 1u << j;
    }
    return o;
}

template <class Sym>
LZ_HD Rec pack_r_record_from(Sym sym, uint32_t len, uint32_t mrd, uint32_t rec)
{
    Rec o; o.x = o.y = 0; o.z = 0; o.w = 0;
    const uint32_t base = rec * 32u;
    const uint32_t rc0 = len + 2u * mrd;  // first base of the reverse complement
    for (uint32_t j = 0; j < 32u; ++j) {
        const uint32_t p = base + j;
        uint32_t c = 5;
        if (p < len) c = sym(p);
        else if (p >= rc0 && p < rc0 + len) {
            c = sym(len - 1u - (p - rc0));
            if (c < 4) c = 3u - c;
        }
        if (c < 4) { o.x |= (c & 1u) << j; o.y |= (c >> 1) << j; }
        else o.z |= 1u << j;
    }
    return o;
}

LZ_HD Rec pack_q_record(const uint8_t* src, uint32_t len, uint32_t rec) { return pack_q_record_from(SymPacked3{src}, len, rec); }
LZ_HD Rec pack_r_record(const uint8_t* src, uint32_t len, uint32_t mrd, uint32_t rec)
{
    return pack_r_record_from(SymPacked3{src}, len, mrd, rec);
}

// ------------------------------------------------------------------------------------------
// K2 element: key of the mal-mer starting at R-form position pos, if all mal symbols are ACGT
LZ_HD bool ref_kmer_at(const Rec* r, int Dr, int pos, int mal, uint64_t& key)
{
    if (pos + mal > Dr) return false;
    uint32_t lo, hi, nm;
    ext3(r, pos, lo, hi, nm);
    if (nm & low_mask32(mal)) return false;
    key = kmer_key(lo, hi, mal);
    return true;
}

}  // namespace lzani
