// parse_core.h -- one warp parses one ordered (reference, query) pair.
//
// This is the B200 re-design of CParser::parse + calc_stats (reference src/parser.cpp:482-716,
// 734-783).  The reference walks the query base by base on one CPU thread, building a factor
// vector and summarising it afterwards.  Here:
//
//  * 32 query positions are examined per step (one per lane): each lane probes the
//    reference's anchor presence bitmap with its mal-mer; in the mqd-window that follows a
//    match, lanes additionally compare the query against the reference along the <= mqd+mrd
//    diagonals of the prediction window (this replaces the ht_short seed table: a seed hit at
//    query position k on diagonal d is a run of >= msl equal bases starting at k).  A ballot
//    finds the first position where the sequential algorithm would have found a match;
//    positions before it are, by construction, literal steps.
//  * the match itself (longest anchor, best seed, the double-precision arbitration, the
//    two-sided gap re-alignment, approximate backward/forward extension) is evaluated
//    cooperatively on 32-base mismatch words: lanes own consecutive words, ballots / REDUX
//    find first mismatches, window overflows and best splits.
//  * no factor vector: the per-pair counts are accumulated in O(1) state (SURVEY Appendix A;
//    the oracle keeps the factor vector, so the two formulations check each other).
//
// Everything is written against the warp policy W of simt.h.
#pragma once
#include "layout.h"
#include "regions.h"

namespace lzani {

struct PairView {
    const Rec* q; int Dq;          // Q-form of the query, Dq = Lq + mrd
    const Rec* r; int Dr;          // R-form of the reference
    const uint32_t* bitmap; uint32_t bbits;
    const uint64_t* table; uint32_t tbits;
};

struct PairResult { int m, x, n; };   // sym_in_matches, sym_in_literals, no_components

LZ_HD uint32_t ld_u32(const uint32_t* p)
{
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}
LZ_HD uint64_t ld_u64(const uint64_t* p)
{
#if defined(__CUDA_ARCH__)
    return __ldg(reinterpret_cast<const unsigned long long*>(p));
#else
    return *p;
#endif
}

// ---- 1 - 4^-len, the base of the chance estimates (prob_len, parser.h:134-172).  For
// len >= 27 the difference rounds to exactly 1.0 in double, as it does in the reference.
LZ_HD double one_minus_4pow(int len)
{
    if (len >= 27) return 1.0;
    if (len <= 0) return 0.0;
    // 4^-len is an exact power of two: build it from its exponent bits
    union { uint64_t u; double d; } cv;
    cv.u = (uint64_t)(1023 - 2 * len) << 52;
    return 1.0 - cv.d;
}
// ipow<double> (parser.h:174-188): square-and-multiply, LSB first, multiplications only
LZ_HD double ipow_u32(double base, uint32_t e)
{
    double r = 1.0;
    while (e) {
#if defined(__CUDA_ARCH__)
        if (e & 1u) r = __dmul_rn(r, base);
        base = __dmul_rn(base, base);
#else
        if (e & 1u) r *= base;
        base *= base;
#endif
        e >>= 1;
    }
    return r;
}

// The anchor / seed arbitration of parser.cpp:606-610: take the anchor iff pa > pc with
// pa = (1 - 4^-a_len)^ea, pc = (1 - 4^-best_len)^ec (ea, ec as the reference's uint32).  Two outcomes need no
// multiplication and are exact: a power of exactly 1.0 (len >= 27, or exponent 0) is exactly 1.0, and a base
// below 1 raised to an exponent >= 1 stays below 1 (the product of two doubles <= 1 - 2^-53 is at most
// 1 - 2^-52 + 2^-106, which rounds to 1 - 2^-52, never up to 1.0).  Long anchors in related genomes take them.
LZ_HD bool anchor_beats_seed(int a_len, uint32_t ea, int best_len, uint32_t ec)
{
    if (best_len >= 27 || ec == 0u) return false;        // pc == 1.0 >= pa
    if (a_len >= 27 || ea == 0u) return true;            // pa == 1.0 > pc
    return ipow_u32(one_minus_4pow(a_len), ea) > ipow_u32(one_minus_4pow(best_len), ec);
}

// ---- mismatch streams.  A stream is a virtual bit string (1 = mismatch); stream word `wi`
// holds stream positions [32*wi, 32*wi+32).
struct FwdStream {   // position k <-> q[q0+k] vs r[r0+k]
    const PairView* pv; int q0, r0;
    LZ_HD uint32_t word(int wi) const { return mism32(pv->q, pv->Dq, q0 + 32 * wi, pv->r, pv->Dr, r0 + 32 * wi); }
};
struct BwdStream {   // position k <-> q[qe-1-k] vs r[re-1-k]
    const PairView* pv; int qe, re;
    LZ_HD uint32_t word(int wi) const
    {
        const int a = qe - 32 * (wi + 1), b = re - 32 * (wi + 1);
        int d = 0;
        if (-a > d) d = -a;
        if (-b > d) d = -b;
        if (d >= 32) return 0xFFFFFFFFu;
        uint32_t f = mism32(pv->q, pv->Dq, a + d, pv->r, pv->Dr, b + d);
        if (d) f = (f << d) | low_mask32(d);
        return brev32(f);
    }
};

struct AnyStream {   // direction chosen at run time: one instantiation of approx_extend()
    const PairView* pv; int a, b; bool back;
    LZ_HD uint32_t word(int wi) const
    {
        if (back) { BwdStream s{pv, a, b}; return s.word(wi); }
        FwdStream s{pv, a, b};
        return s.word(wi);
    }
};

// ---- equal_len (parser.cpp:192-207): common prefix of q[qpos..] and r[rpos..], the first
// `from` bases being known equal.  Uniform arguments, uniform result.
template <class W>
LZ_COLD int equal_len_coop(W& w, const PairView& pv, int qpos, int rpos, int from)
{
    // first word by every lane alike (one broadcast load); most matches end inside it
    const uint32_t w0 = mism32(pv.q, pv.Dq, qpos + from, pv.r, pv.Dr, rpos + from);
    if (w0) return from + ctz32(w0);
    int base = from + 32;
    for (;;) {
        typename W::template Lanes<uint32_t> mw;
        typename W::template Lanes<int> nz;
        LZ_EACH(w, ln)
        {
            mw[ln] = mism32(pv.q, pv.Dq, qpos + base + 32 * ln, pv.r, pv.Dr, rpos + base + 32 * ln);
            nz[ln] = mw[ln] != 0;
        }
        const uint32_t b = w.ballot(nz);
        if (b) {
            const int l0 = ctz32(b);
            return base + 32 * l0 + ctz32(w.get(mw, l0));
        }
        base += 1024;
    }
}

// ---- try_extend_forward / try_extend_backward (parser.cpp:377-441) on a mismatch stream of
// `limit` positions.  Walking base by base the reference keeps the number of mismatches among
// the last `aw` positions and stops after the first position where it exceeds `am`; the result
// is (end of) the last position that closes a run of >= ar matches (the run counter starts at
// ar, i.e. positions before the stream count as matches).  Here each lane owns one stream word:
// the window count at position j is a popcount over the word and its predecessor, the first
// overflow is a min-reduction, the last run end a max-reduction.  Also returns the number of
// matching positions below the returned length (what compare_ranges, parser.cpp:210-248, would
// emit as match factors).  Requires aw <= 32 and ar <= 32 (checked by the host).
template <class W, class PT>
LZ_COLD void approx_extend(W& w, const AnyStream& st, int limit, const PT& P, int& out_len, int& out_matches)
{
    out_len = 0;
    out_matches = 0;
    if (limit <= 0) return;
    uint32_t carry = 0;           // stream word preceding the chunk (0: "matches" before the start)
    int matches_before = 0;       // matches in stream positions below the chunk
    const uint64_t wmask = low_mask64(P.aw);
    LZ_ROLLED
    for (int base = 0;; base += 1024) {
        typename W::template Lanes<uint32_t> cur, prev;
        LZ_EACH(w, ln)
        {
            const int p0 = base + 32 * ln;
            cur[ln] = p0 < limit ? st.word(p0 >> 5) : 0u;
        }
        w.shift_up(prev, cur, carry);
        typename W::template Lanes<int> fb;     // first overflow position of the lane's word
        LZ_EACH(w, ln)
        {
            const int p0 = base + 32 * ln;
            const uint64_t x = ((uint64_t)cur[ln] << 32) | prev[ln];
            int f = 0x7FFFFFFF;
            // cheap bound: the window around any position of this word is inside these bits
            if (p0 < limit && popc32(cur[ln]) + popc32(prev[ln] >> (33 - P.aw > 31 ? 31 : 33 - P.aw)) > P.am) {
                uint32_t t = cur[ln];                 // the window can only overflow at a mismatch
                while (t) {
                    const int j = ctz32(t);
                    if (popc64((x >> (33 + j - P.aw)) & wmask) > P.am) { f = p0 + j; break; }
                    t &= t - 1u;
                }
            }
            fb[ln] = f;
        }
        int brk = w.min_i32(fb);
        const int chunk_end = base + 1024 < limit ? base + 1024 : limit;   // positions that exist
        int end = chunk_end;                                              // positions processed
        if (brk < end) end = brk + 1; else brk = 0x7FFFFFFF;
        // run ends: position j closes a run of >= ar matches
        typename W::template Lanes<int> hi, cnt, tot;
        LZ_EACH(w, ln)
        {
            const int p0 = base + 32 * ln;
            const uint64_t y = ~(((uint64_t)cur[ln] << 32) | prev[ln]);
            uint64_t t = y;
            for (int s = 1; s < P.ar; ++s) t &= y << s;
            uint32_t ok = (uint32_t)(t >> 32);
            const int nvalid = end - p0;                 // positions of this word below `end`
            ok &= nvalid <= 0 ? 0u : low_mask32(nvalid > 32 ? 32 : nvalid);
            hi[ln] = ok ? p0 + 32 - clz32(ok) : 0;       // (last position) + 1
        }
        const int last = w.max_i32(hi);
        LZ_EACH(w, ln)
        {
            const int p0 = base + 32 * ln;
            const uint32_t mt = ~cur[ln];
            const int nl = last - p0, ne = chunk_end - p0;
            cnt[ln] = nl <= 0 ? 0 : popc32(mt & low_mask32(nl > 32 ? 32 : nl));
            tot[ln] = ne <= 0 ? 0 : popc32(mt & low_mask32(ne > 32 ? 32 : ne));
        }
        if (last > 0) {
            out_len = last;
            out_matches = matches_before + w.add_i32(cnt);
        }
        if (brk != 0x7FFFFFFF || chunk_end >= limit) return;
        matches_before += w.add_i32(tot);
        carry = w.get(cur, 31);
    }
}

// ---- compare_ranges_both_ways (parser.cpp:251-374), counts only: the best number of matching
// columns when the `len` query bases ending at qe are aligned, up to a split point, against
// the reference continuing from r_left and, after it, against the reference ENDING at r_right
// (the caller passes the end of the new match, parser.cpp:632 -- kept).  len <= 1024.
template <class W>
LZ_COLD int both_ways_best(W& w, const PairView& pv, int qe, int r_left, int r_right, int len)
{
    if (len <= 0) return 0;
    const int q0 = qe - len;
    const int span = r_right < r_left ? len : (r_right - r_left < len ? r_right - r_left : len);
    const int rmax = span < r_right ? span : r_right;
    FwdStream L{&pv, q0, r_left};
    BwdStream R{&pv, qe, r_right};
    if (span < 32) {
        // the usual gap (a few literals before a close seed): both sides are one word, every lane loads the
        // same two words and lane s evaluates split s
        const uint32_t e_l = ~L.word(0) & low_mask32(span), e_r = ~R.word(0) & low_mask32(rmax);
        typename W::template Lanes<int> v;
        LZ_EACH(w, ln)
        {
            const int s = ln > span ? span : ln;
            const int u = span - s;
            v[ln] = popc32(e_l & low_mask32(s)) + (u > rmax ? 0 : popc32(e_r & low_mask32(u)));
        }
        return w.max_i32(v);
    }
    typename W::template Lanes<uint32_t> el, er;   // EQUAL bits of the two sides, per word
    typename W::template Lanes<int> cl, cr, pl, pr;
    LZ_EACH(w, ln)
    {
        const int p0 = 32 * ln;
        const int nl = span - p0, nr = rmax - p0;
        el[ln] = nl <= 0 ? 0u : (~L.word(ln) & low_mask32(nl > 32 ? 32 : nl));
        er[ln] = nr <= 0 ? 0u : (~R.word(ln) & low_mask32(nr > 32 ? 32 : nr));
        cl[ln] = popc32(el[ln]);
        cr[ln] = popc32(er[ln]);
    }
    w.exscan_add(pl, cl);
    w.exscan_add(pr, cr);
    int best = 0;
    LZ_ROLLED
    for (int s0 = 0; s0 <= span; s0 += 32) {
        typename W::template Lanes<int> sl, sr;
        typename W::template Lanes<uint32_t> wl, wr;
        typename W::template Lanes<int> bl, br, v;
        LZ_EACH(w, ln)
        {
            int s = s0 + ln;
            if (s > span) s = span;
            const int u = span - s;
            sl[ln] = s >> 5;
            sr[ln] = u >> 5;
        }
        w.gather(wl, el, sl);
        w.gather(bl, pl, sl);
        w.gather(wr, er, sr);
        w.gather(br, pr, sr);
        LZ_EACH(w, ln)
        {
            int s = s0 + ln;
            if (s > span) s = span;
            const int u = span - s;
            const int left = bl[ln] + popc32(wl[ln] & low_mask32(s & 31));
            const int right = u > rmax ? 0 : br[ln] + popc32(wr[ln] & low_mask32(u & 31));
            v[ln] = left + right;
        }
        const int m = w.max_i32(v);
        if (m > best) best = m;
    }
    return best;
}

// ---- regions kernel only: what compare_ranges (parser.cpp:210-248) would emit for `len` columns of a
// forward stream, condensed to what the region fold needs (regions.h).
template <class W>
LZ_COLD RangeSummary summarize_range(W& w, const FwdStream& st, int len)
{
    RangeSummary u;
    u.first = -1; u.first_run = 0; u.last_end = 0; u.matches = 0;
    bool growing = false;
    LZ_ROLLED
    for (int base = 0; base < len; base += 1024) {
        typename W::template Lanes<uint32_t> eq;
        typename W::template Lanes<int> nz, c;
        LZ_EACH(w, ln)
        {
            const int p0 = base + 32 * ln, nv = len - p0;
            eq[ln] = nv <= 0 ? 0u : (~st.word(p0 >> 5) & low_mask32(nv > 32 ? 32 : nv));
            nz[ln] = eq[ln] != 0;
            c[ln] = popc32(eq[ln]);
        }
        const uint32_t any = w.ballot(nz);
        u.matches += w.add_i32(c);
        if (any) {
            const int lh = 31 - clz32(any);
            u.last_end = base + 32 * lh + 32 - clz32(w.get(eq, lh));
            if (u.first < 0) {
                const int l0 = ctz32(any);
                u.first = base + 32 * l0 + ctz32(w.get(eq, l0));
                growing = true;
            }
        }
        if (growing) {
            // first column at or after u.first (inside this chunk, below len) that does not match
            typename W::template Lanes<int> zp;
            LZ_EACH(w, ln)
            {
                const int p0 = base + 32 * ln, nv = len - p0;
                uint32_t z = nv <= 0 ? 0u : (~eq[ln] & low_mask32(nv > 32 ? 32 : nv));
                const int rel = u.first - p0;
                if (rel >= 32) z = 0;
                else if (rel > 0) z &= ~low_mask32(rel);
                zp[ln] = z ? p0 + ctz32(z) : 0x7FFFFFFF;
            }
            const int mn = w.min_i32(zp);
            if (mn != 0x7FFFFFFF) { u.first_run = mn - u.first; growing = false; }
        }
    }
    if (growing) u.first_run = len - u.first;
    return u;
}

// ---- regions kernel only: the factors compare_ranges_both_ways (parser.cpp:251-374) emits for the gap
// of `len` query bases ending at qe, fed to the region fold column by column (uniform serial code:
// the gap is at most mqd+1 long).  Same split rule as both_ways_best() -- the LAST split with the
// maximal sum -- and the reference's position bookkeeping of the right-hand part (:353-368).
template <class W, class RS>
LZ_COLD void both_ways_regions(W& w, const PairView& pv, RS& rt, int qe, int r_left, int r_right, int len)
{
    if (len <= 0) return;
    const int q0 = qe - len;
    const int span = r_right < r_left ? len : (r_right - r_left < len ? r_right - r_left : len);
    const int rmax = span < r_right ? span : r_right;
    FwdStream L{&pv, q0, r_left};
    BwdStream R{&pv, qe, r_right};
    typename W::template Lanes<uint32_t> el, er;
    LZ_EACH(w, ln)
    {
        const int p0 = 32 * ln;
        const int nl = span - p0, nr = rmax - p0;
        el[ln] = nl <= 0 ? 0u : (~L.word(ln) & low_mask32(nl > 32 ? 32 : nl));
        er[ln] = nr <= 0 ? 0u : (~R.word(ln) & low_mask32(nr > 32 ? 32 : nr));
    }
    // best split: left[s] + right[span - s], right[u] = 0 for u > rmax (the reference zero-pads)
    int best = 0, split = 0;
    {
        int right_total = 0;
        for (int t = 0; t < rmax; ++t) right_total += (int)((w.get(er, t >> 5) >> (t & 31)) & 1u);
        int left = 0, right = right_total;      // right = equal columns among the first (span - s) of R
        int ru = span;                          // = span - s
        for (int sp = 0; sp <= span; ++sp) {
            const int rv = ru > rmax ? 0 : right;
            if (left + rv >= best) { best = left + rv; split = sp; }
            if (sp < span) {
                left += (int)((w.get(el, sp >> 5) >> (sp & 31)) & 1u);
                // moving to span - (sp+1): drop R column (ru - 1) when it lies inside rmax
                if (ru - 1 < rmax) right -= (int)((w.get(er, (ru - 1) >> 5) >> ((ru - 1) & 31)) & 1u);
                --ru;
            }
        }
    }
    // emit: a pending match factor (q, r, n) is flushed when a column of the other kind follows
    int fq = 0, fr = 0, fn = 0;          // pending match factor, fn == 0: none
    int last_flag = -1;                  // flag of the last factor of this gap: 1 match, 0 literals, -1 none
    int qp = q0;
    for (int sp = 1; sp <= split; ++sp, ++qp) {
        const int same = (int)((w.get(el, (sp - 1) >> 5) >> ((sp - 1) & 31)) & 1u);
        if (same) {
            if (last_flag == 1) ++fn;
            else { fq = qp; fr = r_left + sp - 1; fn = 1; }
        } else {
            if (fn) { rt.close(fq, fr, fn); fn = 0; }
            rt.lit(1);
        }
        last_flag = same;
    }
    if (span < len) {                    // unscanned middle
        if (fn) { rt.close(fq, fr, fn); fn = 0; }
        rt.lit(len - span);
        last_flag = 0;
    }
    if (split < span) {
        const int shift = len - span;
        const int from_right = span - split;
        int same = (int)((w.get(er, (from_right - 1) >> 5) >> ((from_right - 1) & 31)) & 1u);
        qp = q0 + split + shift;
        if (!same && (split > 0 || shift > 0) && last_flag == 0) rt.lit(1);      // merged: qp NOT advanced
        else {
            if (fn) { rt.close(fq, fr, fn); fn = 0; }
            if (same) { fq = qp; fr = r_right - from_right; fn = 1; }
            else rt.lit(1);
            last_flag = same;
            ++qp;
        }
        for (int t = from_right - 1; t > 0; --t, ++qp) {
            same = (int)((w.get(er, (t - 1) >> 5) >> ((t - 1) & 31)) & 1u);
            if (same) {
                if (last_flag == 1) ++fn;
                else { fq = qp; fr = r_right - t; fn = 1; }
            } else {
                if (fn) { rt.close(fq, fr, fn); fn = 0; }
                rt.lit(1);
            }
            last_flag = same;
        }
    }
    if (fn) rt.close(fq, fr, fn);
}

// ---- anchor lookup: the probe chain of ht_long (parser.cpp:516-530 / :587-601).  Returns the
// longest common prefix over all reference positions whose mal-mer equals the query's at
// qpos; ties -> smallest position (the reference's strict '>' over ascending positions).
// (0,0) when there is none.  32 slots of the chain are inspected per step, one per lane.
template <class W, class PT>
LZ_COLD void anchor_lookup(W& w, const PairView& pv, const PT& P, int qpos, uint64_t key, int& a_pos, int& a_len)
{
    a_pos = 0;
    a_len = 0;
    const uint32_t tmask = (1u << pv.tbits) - 1u;
    const uint32_t tag = table_tag(key, P.mal);
    uint32_t h = table_home(key, pv.tbits);
    for (;;) {
        typename W::template Lanes<uint64_t> slot;
        typename W::template Lanes<int> is_empty, is_cand;
        LZ_EACH(w, ln)
        {
            slot[ln] = ld_u64(pv.table + ((h + (uint32_t)ln) & tmask));
            is_empty[ln] = slot[ln] == LZ_EMPTY;
            is_cand[ln] = !is_empty[ln] && (uint32_t)(slot[ln] >> 32) == tag;
        }
        const uint32_t em = w.ballot(is_empty);
        uint32_t cm = w.ballot(is_cand);
        if (em) cm &= (em & (0u - em)) - 1u;      // only slots before the first empty one
        if (cm) {
            // every candidate lane measures the first 32 bases of its own candidate
            typename W::template Lanes<int> len, more;
            LZ_EACH(w, ln)
            {
                len[ln] = 0;
                more[ln] = 0;
                if ((cm >> ln) & 1u) {
                    const uint32_t m0 = mism32(pv.q, pv.Dq, qpos, pv.r, pv.Dr, (int)(uint32_t)slot[ln]);
                    if (m0) len[ln] = ctz32(m0);
                    else more[ln] = 1;
                }
            }
            uint32_t lm = w.ballot(more);
            while (lm) {                            // long matches: finish them warp-wide, one by one
                const int l = ctz32(lm);
                lm &= lm - 1u;
                const int rp = (int)(uint32_t)w.get(slot, l);
                const int full = equal_len_coop(w, pv, qpos, rp, 32);
                LZ_EACH(w, ln) if (ln == l) len[ln] = full;
            }
            typename W::template Lanes<int> v;
            LZ_EACH(w, ln) v[ln] = (((cm >> ln) & 1u) && len[ln] >= P.mal) ? len[ln] : 0;
            const int bl = w.max_i32(v);
            if (bl > 0) {
                typename W::template Lanes<uint32_t> pp;
                LZ_EACH(w, ln) pp[ln] = v[ln] == bl ? (uint32_t)slot[ln] : 0xFFFFFFFFu;
                const int bp = (int)w.min_u32(pp);
                if (bl > a_len || (bl == a_len && bp < a_pos)) { a_len = bl; a_pos = bp; }
            }
        }
        if (em) return;
        h += 32u;
    }
}

// ---- the seeds of one step.  Query positions i0+k, k in [0,nB), are in "prediction mode":
// the reference expects the match to continue near P0+lit0+k (P0 = reference position right
// after the previous match, lit0 = literals since).  A seed for position k is a reference
// position x in [P0, P0+lit0+k+mrd) with r[x..x+msl) == q[i0+k..+msl)  (parser.cpp:548-580).
// Diagonal d = x - P0 - k ranges over [-(nB-1), lit0+mrd).  For one diagonal the equality word
// E (bit k = q[i0+k]==r[P0+d+k]) is built from two 32-base comparisons; a seed at k is a run
// of msl ones starting at bit k.
struct DiagWord {
    uint64_t E;     // equality bits, bit k <-> query position i0+k  (bits below k0 are 0)
};
LZ_HD uint64_t diag_equal_bits(const PairView& pv, int i0, int P0, int d)
{
    const int k0 = d < 0 ? -d : 0;
    const uint32_t m0 = mism32(pv.q, pv.Dq, i0 + k0, pv.r, pv.Dr, P0 + d + k0);
    const uint32_t m1 = mism32(pv.q, pv.Dq, i0 + k0 + 32, pv.r, pv.Dr, P0 + d + k0 + 32);
    const uint64_t e = ~(((uint64_t)m1 << 32) | m0);
    return k0 ? e << k0 : e;      // k0 <= 31
}
LZ_HD uint64_t run_starts(uint64_t E, int n)   // bit k set <=> bits k..k+n-1 of E all set
{
    uint64_t t = E;
    int have = 1;
    while (have * 2 <= n) { t &= t >> have; have *= 2; }
    if (have < n) t &= t >> (n - have);
    return t;
}


// ------------------------------------------------------------------------------------------
// 64 bases of the three planes starting at pos (three 128-bit loads)
struct Planes64 { uint32_t lo0, lo1, hi0, hi1, nm0, nm1; };
LZ_HD Planes64 ext64(const Rec* s, int pos)
{
    const int wd = pos >> 5, sh = pos & 31;
    const Rec a = ld_rec(s + wd), b = ld_rec(s + wd + 1), c = ld_rec(s + wd + 2);
    Planes64 o;
    o.lo0 = funnel_r(a.x, b.x, sh); o.lo1 = funnel_r(b.x, c.x, sh);
    o.hi0 = funnel_r(a.y, b.y, sh); o.hi1 = funnel_r(b.y, c.y, sh);
    o.nm0 = funnel_r(a.z, b.z, sh); o.nm1 = funnel_r(b.z, c.z, sh);
    return o;
}
// equality word of two 64-base windows (bit t = bases t equal, N never equal)
LZ_HD uint64_t equal64(const Planes64& a, const Planes64& b)
{
    const uint32_t m0 = (a.lo0 ^ b.lo0) | (a.hi0 ^ b.hi0) | a.nm0 | b.nm0;
    const uint32_t m1 = (a.lo1 ^ b.lo1) | (a.hi1 ^ b.hi1) | a.nm1 | b.nm1;
    return ~(((uint64_t)m1 << 32) | m0);
}

// Equality word of one diagonal of a prediction-mode chunk: bit k <-> q[i0+k] == r[P0+d+k]
// with Q = ext64(q, i0).  Bits k < -d (reference positions before P0) are 0.
LZ_HD uint64_t diag_equal64(const PairView& pv, const Planes64& Q, int i0, int P0, int d)
{
    const int rp = P0 + d;
    if (rp >= 0) {
        if (rp >= pv.Dr) return 0;
        uint64_t e = equal64(Q, ext64(pv.r, rp));
        if (d < 0) e &= ~low_mask64(-d);
        return e;
    }
    return diag_equal_bits(pv, i0, P0, d);    // the window starts before the reference does (rare)
}

// ------------------------------------------------------------------------------------------
// Approximate extension inside ONE stream word, evaluated by a single lane (the speculative
// per-hit path).  mw: bit j = mismatch at stream position j; nv <= 32 positions exist in this
// word; `more`: the stream continues behind it.  Same rule as approx_extend() above
// (try_extend_forward/backward, parser.cpp:377-441).  `examined` = positions the sequential
// walk would have looked at; complete == false when the walk would run on into the next word.
struct ExtWord { int len, matches, examined; bool complete; };
template <class PT>
LZ_HD ExtWord extend_in_word(uint32_t mw, int nv, bool more, const PT& P)
{
    ExtWord o;
    o.len = 0; o.matches = 0; o.examined = 0; o.complete = true;
    if (nv <= 0) return o;
    const uint32_t m = mw & low_mask32(nv);
    int brk = -1;
    {
        // the window can only overflow at a mismatch, and not before the (am+1)-th one
        uint32_t t = m;
        const int na = P.am < 32 ? P.am : 32;
        for (int c = 0; c < na; ++c) t &= t - 1u;       // no early exit: lanes stay together
        const uint64_t x = (uint64_t)m << 32;          // positions before the stream: matches
        const uint64_t wmask = low_mask64(P.aw);
        while (t) {
            const int j = ctz32(t);
            if (popc64((x >> (33 + j - P.aw)) & wmask) > P.am) { brk = j; break; }
            t &= t - 1u;
        }
    }
    const int end = brk >= 0 ? brk + 1 : nv;
    o.examined = end;
    o.complete = brk >= 0 || !more;
    const uint64_t y = ~((uint64_t)m << 32);
    uint64_t t = y;
    for (int s = 1; s < P.ar; ++s) t &= y << s;
    const uint32_t ok = (uint32_t)(t >> 32) & low_mask32(end);
    if (ok) {
        o.len = 32 - clz32(ok);
        o.matches = popc32(~m & low_mask32(o.len));
    }
    return o;
}

// try_extend_forward / try_extend_backward when the walk ends inside the first stream word (nearly always:
// with the default parameters a random continuation overflows the window after ~10 positions): one
// broadcast comparison evaluated by every lane alike instead of the 1024-position cooperative chunk of
// approx_extend().  false: the walk runs on into the next word -- use approx_extend().
template <class PT>
LZ_COLD bool approx_extend_short(const AnyStream& st, int limit, const PT& P, int& out_len, int& out_matches)
{
    out_len = 0;
    out_matches = 0;
    if (limit <= 0) return true;
    const ExtWord e = extend_in_word(st.word(0), limit > 32 ? 32 : limit, limit > 32, P);
    if (!e.complete) return false;
    out_len = e.len;
    out_matches = e.matches;
    return true;
}

// ------------------------------------------------------------------------------------------
// Speculative seed scan, evaluated by ONE lane for ITS hit ahead of the walk (msl == 7 kernels).
// Should the walk take the hit's anchor as a match (with the forward extension of its record),
// it continues in prediction mode at query position i0, reference position P0, literal run 0.
// The seeds of the next positions k = 0..K (parser.cpp:548-580: reference positions
// x in [P0, P0+k+mrd) whose msl-mer equals the query's at i0+k) then depend on nothing else, so
// the first k that has one can be found now.  Row-wise bit-parallel: for query base m the mismatch
// mask Mis_m of the 96-base reference window against that base; a seed at (k, x) is a 0 bit x of
//   S_k = Mis_k | Mis_{k+1}>>1 | ... | Mis_{k+6}>>6   (built by doubling: pairs, fours, 4+3).
// Returns the first k with a seed, 64 when there is none in [0, K], -1 when not applicable (an N
// or the sequence end inside the window: left to the warp-cooperative path).  With a seed at k the
// value also carries, in bits [8,15), the window offset x of its reference position P0+x when that
// is the ONLY seed of position k (SPEC_MULTI set otherwise): best_seed() then has nothing to search.
struct V3 { uint32_t a, b, c; };
LZ_HD V3 v3_or_shr(const V3& x, const V3& y, int j)   // x | (y >> j), 0 < j < 32
{
    V3 r;
    r.a = x.a | funnel_r(y.a, y.b, j);
    r.b = x.b | funnel_r(y.b, y.c, j);
    r.c = x.c | (y.c >> j);
    return r;
}
template <int NW>
LZ_HD V3 v3_and_shr(const V3& x, const V3& y, int j)  // x & (y >> j) on the first NW words, 0 < j < 32, zeros shifted in
{
    V3 r;
    r.a = x.a & funnel_r(y.a, y.b, j);
    r.b = NW > 2 ? x.b & funnel_r(y.b, y.c, j) : x.b & (y.b >> j);
    r.c = NW > 2 ? x.c & (y.c >> j) : 0u;
    return r;
}
constexpr int SPEC_MAX_K = 57;
constexpr int SPEC_MULTI = 1 << 15;
#ifndef LZ_CHAIN
#define LZ_CHAIN 0
#endif
// rounds of the diagonal seed scan (32 diagonals each) issued together: their loads overlap
#ifndef LZ_DIAG_UNROLL
#define LZ_DIAG_UNROLL 1
#endif
#if defined(__CUDACC__)
#define LZ_PRAGMA_(x) _Pragma(#x)
#define LZ_PRAGMA(x) LZ_PRAGMA_(x)
#define LZ_DIAG_LOOP LZ_PRAGMA(unroll LZ_DIAG_UNROLL)
#else
#define LZ_DIAG_LOOP
#endif
constexpr bool kChain = LZ_CHAIN != 0;   // successor links (measured slower on B200: off)
#ifndef LZ_SCAN_UNROLL
#define LZ_SCAN_UNROLL 4
#endif
#ifndef LZ_WIN_WORDS_N
#define LZ_WIN_WORDS_N 8
#endif
constexpr int kScanUnroll = LZ_SCAN_UNROLL;
constexpr int LZ_WIN_WORDS = LZ_WIN_WORDS_N;                    // hit-mask words per lane and window (8192 positions)
constexpr int LZ_WIN_LANE = 32 * LZ_WIN_WORDS;     // query positions per lane and window
// per-warp scratch (shared memory on the device): the hit masks of the current window, one column
// per lane, and the hit list of the current batch
struct WarpScratch {
    uint32_t hm[LZ_WIN_WORDS][32];
    uint16_t list[32];
};
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
static long lz_spec_used = 0, lz_spec_mismatches = 0, lz_pred_steps = 0, lz_batches = 0, lz_hits = 0, lz_matches = 0, lz_fast = 0, lz_chain = 0, lz_why[12] = {0};   // TEST ONLY counters (tests/emul)
#endif
// four rows of the speculative seed scan (spec_first_seed below) on the first NW words of the window
constexpr int SPEC_GO_ON = -2;
template <int NW, bool kRefN, class PT>
LZ_HD int spec_rows4(const PT& P, const V3& rl, const V3& rh, const V3& rn, uint64_t ql, uint64_t qh, uint64_t qn, int mb, int K,
                     int mrdx, V3& e_prev, V3 (&a2)[4], V3 (&a4)[4])
{
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int u = 0; u < 4; ++u) {      // ring slots (m mod 4) are compile-time
        const int m = mb + u;
        // all-ones where the query base's plane bit is 0: equal <=> (r ^ ~q) on both planes
        const uint32_t nl = ((uint32_t)(ql >> u) & 1u) - 1u, nh = ((uint32_t)(qh >> u) & 1u) - 1u;
        V3 eq;
        if (kRefN) {       // an N equals nothing: reference N masked per word (still two LOP3), a query N empties its row
            const uint32_t nq = ((uint32_t)(qn >> u) & 1u) - 1u;
            eq.a = (rl.a ^ nl) & ((rh.a ^ nh) & ~rn.a) & nq;
            eq.b = (rl.b ^ nl) & ((rh.b ^ nh) & ~rn.b) & nq;
            eq.c = NW > 2 ? (rl.c ^ nl) & ((rh.c ^ nh) & ~rn.c) & nq : 0u;
        } else {
            eq.a = (rl.a ^ nl) & (rh.a ^ nh);
            eq.b = (rl.b ^ nl) & (rh.b ^ nh);
            eq.c = NW > 2 ? (rl.c ^ nl) & (rh.c ^ nh) : 0u;
        }
        const V3 p2 = v3_and_shr<NW>(e_prev, eq, 1);                // A2_{m-1} = E_{m-1} & E_m>>1
        const V3 p4 = v3_and_shr<NW>(a2[(u + 1) % 4], p2, 2);       // A4_{m-3} = A2_{m-3} & A2_{m-1}>>2
        const V3 s7 = v3_and_shr<NW>(a4[(u + 2) % 4], p4, 3);       // S_{m-6}  = A4_{m-6} & A4_{m-3}>>3
        a2[(u + 3) % 4] = p2;                                       // slot (m-1) mod 4
        a4[(u + 1) % 4] = p4;                                       // slot (m-3) mod 4
        e_prev = eq;
        if ((s7.a | s7.b | s7.c) && m >= 6) {
            const int k = m - 6;
            const int nv = k + mrdx;                                // valid reference offsets of row k: x < k + mrd (+ lit0)
            const uint32_t c0 = s7.a & low_mask32(nv > 32 ? 32 : nv);
            const uint32_t c1 = nv > 32 ? s7.b & low_mask32(nv - 32 > 32 ? 32 : nv - 32) : 0u;
            const uint32_t c2 = nv > 64 ? s7.c & low_mask32(nv - 64 > 32 ? 32 : nv - 64) : 0u;
            if (c0 | c1 | c2) {
                if (k > K) return 64;
                const int x = c0 ? ctz32(c0) : (c1 ? 32 + ctz32(c1) : 64 + ctz32(c2));
                const bool multi = popc32(c0) + popc32(c1) + popc32(c2) > 1;
                return k | (x << 8) | (multi ? SPEC_MULTI : 0);
            }
        }
    }
    return SPEC_GO_ON;
}

// mrdx = mrd + (literals already behind the window's first position): P.mrd for the window a match opens
// (k_hit_spec); the lane walk (lane_walk.h) also scans windows it has already walked lit0 positions into.
// kRefN: N inside the reference window is handled (masked out of every row) instead of giving up -- the lane walk's
// own scans; k_hit_spec uses the plain form (three registers fewer) and leaves such windows to the walk.
template <bool kRefN, class PT>
LZ_HD int spec_first_seed_t(const PairView& pv, const PT& P, int i0, int P0, int K, int mrdx)
{
    // window: query bases [i0, i0+K+7), reference bases [P0, P0+K+mrdx+6)
    if (K < 0 || K > SPEC_MAX_K || K + mrdx + 6 > 96 || P0 < 0) return -1;
    const Planes64 Q = ext64(pv.q, i0);
    uint64_t qn = (((uint64_t)Q.nm1 << 32) | Q.nm0) & low_mask64(K + 7);
    if (qn) {
        // an msl-mer over an N is no seed.  N from some base to the end of the window (every query ends like that):
        // the rows from the first N - 6 on are empty -- scan the ones before.  N in the middle: masked row by row
        // (kRefN) or left to the walk.
        if ((qn | (qn - 1ull)) == low_mask64(K + 7)) {
            K = ctz64(qn) - 7;
            if (K < 0) return 64;
            qn = 0;
        } else if (!kRefN) return -1;
    }
    V3 rl, rh, rn{0, 0, 0};
    {
        const int wd = P0 >> 5, sh = P0 & 31;
        const Rec a = ld_rec(pv.r + wd), b = ld_rec(pv.r + wd + 1), c = ld_rec(pv.r + wd + 2), d = ld_rec(pv.r + wd + 3);
        rl.a = funnel_r(a.x, b.x, sh); rl.b = funnel_r(b.x, c.x, sh); rl.c = funnel_r(c.x, d.x, sh);
        rh.a = funnel_r(a.y, b.y, sh); rh.b = funnel_r(b.y, c.y, sh); rh.c = funnel_r(c.y, d.y, sh);
        const uint32_t n0 = funnel_r(a.z, b.z, sh), n1 = funnel_r(b.z, c.z, sh), n2 = funnel_r(c.z, d.z, sh);
        const int nr = K + mrdx + 6;       // <= 96
        if (kRefN) { rn.a = n0; rn.b = n1; rn.c = n2; }
        else if ((nr >= 32 ? n0 : (n0 & low_mask32(nr))) | (nr > 32 ? (n1 & low_mask32(nr - 32 > 32 ? 32 : nr - 32)) : 0u) |
                 (nr > 64 ? (n2 & low_mask32(nr - 64)) : 0u))
            return -1;
    }
    uint64_t ql = ((uint64_t)Q.lo1 << 32) | Q.lo0, qh = ((uint64_t)Q.hi1 << 32) | Q.hi0;
    // MATCH masks E_m (bit x: r[P0+x] == q[i0+m]), AND-combined by doubling: pairs A2, fours A4, sevens S.
    // Shifted-in zeros and rows that do not exist read as "no match", so nothing has to be masked inside the
    // loop: the validity of a candidate (reference offset x < k + mrd, position k <= K) is tested only when one
    // appears (about 1 row in 100).  Rows are run in groups of four; up to three rows past the last one are
    // evaluated for nothing (they can only produce candidates at k > K, which are dropped).
    // The third window word (offsets 64..95) only matters for rows k with k + mrd + 6 > 64: the first rows
    // run on two words.
    V3 e_prev{0, 0, 0}, a2[4], a4[4];
    for (int u = 0; u < 4; ++u) a2[u] = a4[u] = V3{0, 0, 0};
    const int m_last = K + 6;
    int m_two = 58 - mrdx;                  // rows below this (rounded down to a group) need two words
    m_two = m_two > 0 ? m_two & ~3 : 0;
    int mb = 0;
    for (; mb < m_two && mb <= m_last; mb += 4) {
        const int r = spec_rows4<2, kRefN>(P, rl, rh, rn, ql, qh, qn, mb, K, mrdx, e_prev, a2, a4);
        if (r != SPEC_GO_ON) return r;
        ql >>= 4;
        qh >>= 4;
        qn >>= 4;
    }
    for (; mb <= m_last; mb += 4) {
        const int r = spec_rows4<3, kRefN>(P, rl, rh, rn, ql, qh, qn, mb, K, mrdx, e_prev, a2, a4);
        if (r != SPEC_GO_ON) return r;
        ql >>= 4;
        qh >>= 4;
        qn >>= 4;
    }
    return 64;
}

template <class PT>
LZ_HD int spec_first_seed(const PairView& pv, const PT& P, int i0, int P0, int K, int mrdx)
{
    return spec_first_seed_t<false>(pv, P, i0, P0, K, mrdx);
}

// ------------------------------------------------------------------------------------------
// One anchor hit, evaluated by ONE lane ahead of the sequential walk.  The result depends only
// on the query position p (not on the parse state): the longest reference occurrence of the
// mal-mer at p (ties -> smallest position, parser.cpp:516-530), and -- should the walk take it
// as a match -- the approximate extension behind it (parser.cpp:691-695) and in front of it
// (:659-672, valid when the literal run before p is at least `examined` long).  Anything that
// does not fit one 32-base word (long matches, long extensions, long probe chains) is left to
// the warp-cooperative routines above by setting HIT_DEFER / clearing HIT_FWD / HIT_BWD.
//
//   w2: [0,6) anchor length  [6,12) fwd ext  [12,18) fwd matches  + flags
//   w3: [0,6) bwd ext        [6,12) bwd matches  [12,18) bwd positions examined
constexpr uint32_t HIT_LIVE = 1u << 18;    // there is (or may be) an anchor at this position
constexpr uint32_t HIT_DEFER = 1u << 19;   // evaluate with anchor_lookup() when reached
constexpr uint32_t HIT_FWD = 1u << 20;     // fwd fields valid
constexpr uint32_t HIT_BWD = 1u << 21;     // bwd fields valid
constexpr int HIT_MAX_CHAIN = 6;           // matching-tag slots one lane inspects before deferring

// Stage A: table probe + first-word comparison.  Single-exit loop so that the lanes of a batch
// reconverge behind it (the caller also puts a warp sync between the stages).
template <class PT>
LZ_HD void hit_probe(const PairView& pv, const PT& P, int p, uint32_t& o_pos, uint32_t& o_w2)
{
    o_pos = 0; o_w2 = 0;
    uint32_t lo, hi, nm;
    ext3(pv.q, p, lo, hi, nm);
    const uint64_t key = kmer_key(lo, hi, P.mal);
    const uint32_t tmask = (1u << pv.tbits) - 1u;
    const uint32_t tag = table_tag(key, P.mal);
    uint32_t h = table_home(key, pv.tbits);
    int best_len = 0, best_pos = 0, seen = 0;
    bool defer = false;
    bool go = !(nm & low_mask32(P.mal));
    while (go) {
        const uint64_t slot = ld_u64(pv.table + (h & tmask));
        ++h;
        if (slot == LZ_EMPTY) go = false;
        else if ((uint32_t)(slot >> 32) == tag) {
            const int rp = (int)(uint32_t)slot;
            uint32_t rl, rh, rn;
            ext3(pv.r, rp, rl, rh, rn);
            const uint32_t m0 = (lo ^ rl) | (hi ^ rh) | nm | rn;
            if (++seen > HIT_MAX_CHAIN || !m0) { defer = true; go = false; }   // long chain / >= 32 equal bases
            else {
                const int len = ctz32(m0);
                if (len >= P.mal && (len > best_len || (len == best_len && rp < best_pos))) { best_len = len; best_pos = rp; }
            }
        }
    }
    if (defer) o_w2 = HIT_LIVE | HIT_DEFER;
    else if (best_len) { o_pos = (uint32_t)best_pos; o_w2 = HIT_LIVE | (uint32_t)best_len; }
}

// Stage B: both approximate extensions of a probed hit, inside one stream word each
template <class PT>
LZ_HD void hit_extend(const PairView& pv, const PT& P, int p, int best_pos, uint32_t& w2, uint32_t& w3)
{
    const int best_len = (int)(w2 & 63u);
    // both directions through ONE copy of extend_in_word (code size: I-cache)
    uint32_t mw[2];
    int lim[2];
    {
        const int qn = p + best_len, pn = best_pos + best_len;
        lim[0] = pv.Dq - qn;
        if (pv.Dr - pn < lim[0]) lim[0] = pv.Dr - pn;
        mw[0] = lim[0] > 0 ? mism32(pv.q, pv.Dq, qn, pv.r, pv.Dr, pn) : 0xFFFFFFFFu;
        lim[1] = p < best_pos ? p : best_pos;
        BwdStream bs{&pv, p, best_pos};
        mw[1] = lim[1] > 0 ? bs.word(0) : 0xFFFFFFFFu;
    }
    w3 = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
    for (int dir = 0; dir < 2; ++dir) {
        const uint32_t m = dir ? mw[1] : mw[0];
        const int l = dir ? lim[1] : lim[0];
        const ExtWord e = extend_in_word(m, l > 32 ? 32 : l, l > 32, P);
        if (e.complete) {
            if (dir == 0) w2 |= HIT_FWD | ((uint32_t)e.len << 6) | ((uint32_t)e.matches << 12);
            else {
                w2 |= HIT_BWD;
                w3 = (uint32_t)e.len | ((uint32_t)e.matches << 6) | ((uint32_t)e.examined << 12);
            }
        }
    }
}

constexpr uint32_t KV_KEEP = 1u << 16;
constexpr int KV_LINK_SHIFT = 17;
struct ParseState {
    int i, pred, lit;            // query position, ref_pred_pos (<0: none), cur_lit_run_len
    int rs, re, kept_end;        // prev_region_start / prev_region_end / end of last kept region
    int curM, curX;              // matches / counted literals of the region being built
    int totM, totX, totN;        // kept regions so far
};

// One match found at query position qi (parser.cpp:626-698), counts only.  (w2, w3) is the
// speculative record of hit_probe()/hit_extend() when the match IS that anchor, else w2 == 0.
template <class W, class PT, class RS>
LZ_HD void take_match(W& w, const PairView& pv, const PT& P, ParseState& s, RS& rt, int qi, int lit_k, int pred_k,
                      int best_pos, int best_len, uint32_t w2, uint32_t w3)
{
    int dist = best_pos - pred_k;
    if (dist < 0) dist = -dist;
    if (pred_k >= 0 && dist <= P.mrd) {
        // close match: the literal gap is re-aligned from both ends (:630-635)
        const int g = both_ways_best(w, pv, qi, pred_k - lit_k, best_pos + best_len, lit_k);
        s.curM += g + best_len;
        s.curX += lit_k - g;
        if (RS::kActive) {
            both_ways_regions(w, pv, rt, qi, pred_k - lit_k, best_pos + best_len, lit_k);
            rt.close(qi, best_pos, best_len);
        }
    } else {
        // distant match: judge the region built so far (:643, eval_region :446-449) ...
        if (s.rs >= 0 && s.re - s.rs >= P.reg) {
            s.totM += s.curM;
            s.totX += s.curX;
            ++s.totN;
            s.kept_end = s.re;
            if (RS::kActive) rt.end_region(P.reg);
        } else if (RS::kActive) rt.drop();
        // ... everything since the last kept region is now one literal run (:645-655), which the
        // new match may reclaim backwards (:659-672)
        const int run = qi - s.kept_end;
        int back = 0, bm = 0;
        if (run > 0) {
            int lim = run;
            if (qi < lim) lim = qi;
            if (best_pos < lim) lim = best_pos;
            if ((w2 & HIT_BWD) && lim >= (int)((w3 >> 12) & 63u)) {
                back = (int)(w3 & 63u);
                bm = (int)((w3 >> 6) & 63u);
            } else {
                AnyStream bs{&pv, qi, best_pos, true};
                if (!approx_extend_short(bs, lim, P, back, bm)) approx_extend(w, bs, lim, P, back, bm);
            }
        }
        s.curM = bm + best_len;
        s.curX = back - bm;
        s.rs = qi - back;
        if (RS::kActive) {
            if (back > 0) {
                // the reclaimed columns: first run a distant match, the rest and the match itself close (:664-672)
                const FwdStream bw{&pv, qi - back, best_pos - back};
                rt.range(qi - back, best_pos - back, back, summarize_range(w, bw, back), true);
                rt.close(qi, best_pos, best_len);
            } else rt.dist(qi, best_pos, best_len);
        }
    }
    const int qn = qi + best_len;
    const int pn = best_pos + best_len;
    int fwd = 0, fm = 0;
    if (w2 & HIT_FWD) {
        fwd = (int)((w2 >> 6) & 63u);
        fm = (int)((w2 >> 12) & 63u);
    } else {
        int lim = pv.Dq - qn;
        if (pv.Dr - pn < lim) lim = pv.Dr - pn;
        AnyStream fs{&pv, qn, pn, false};
        if (!approx_extend_short(fs, lim, P, fwd, fm)) approx_extend(w, fs, lim, P, fwd, fm);      // :691-695
    }
    s.curM += fm;
    s.curX += fwd - fm;
    if (RS::kActive && fwd > 0) {
        const FwdStream fs{&pv, qn, pn};
        rt.range(qn, pn, fwd, summarize_range(w, fs, fwd), false);
    }
    s.i = qn + fwd;
    s.pred = pn + fwd;
    s.lit = 0;
    s.re = s.i;
}

// query positions of the chunk at which some diagonal of [d_lo, d_hi) starts a run of >= msl equal
// bases (bit k <-> i0+k).  Lanes own diagonals.
template <class W, class PT>
LZ_COLD uint64_t diag_seed_starts(W& w, const PairView& pv, const PT& P, const Planes64& Q, int i0, int P0, int d_lo, int d_hi)
{
    typename W::template Lanes<uint64_t> sm;
    LZ_EACH(w, ln) sm[ln] = 0;
    LZ_DIAG_LOOP
    for (int dbase = d_lo; dbase < d_hi; dbase += 32) {
        LZ_EACH(w, ln)
        {
            const int d = dbase + ln;
            if (d < d_hi) sm[ln] |= run_starts(diag_equal64(pv, Q, i0, P0, d), P.msl);
        }
    }
    return w.or_u64(sm);
}

// best seed at chunk position k (parser.cpp:548-580): longest; ties -> closest to the
// prediction, then smallest position.  Lanes own diagonals.
template <class W, class PT>
LZ_COLD void best_seed(W& w, const PairView& pv, const PT& P, const Planes64& Q, int i0, int P0, int lit0, int k,
                     int d_lo, int d_hi, int& best_pos, int& best_len)
{
    int sl = 0;
    uint32_t skey = 0xFFFFFFFFu;
    const int qi = i0 + k;
    LZ_ROLLED
    for (int dbase = d_lo; dbase < d_hi; dbase += 32) {
        typename W::template Lanes<int> len, more;
        typename W::template Lanes<uint32_t> dk;
        LZ_EACH(w, ln)
        {
            const int d = dbase + ln;
            len[ln] = 0;
            more[ln] = 0;
            dk[ln] = 0xFFFFFFFFu;
            if (d < d_hi && d >= -k) {
                const uint64_t E = diag_equal64(pv, Q, i0, P0, d);
                if ((run_starts(E, P.msl) >> k) & 1ull) {
                    const uint64_t z = ~(E >> k);            // first 0 of the run
                    const int rl = z ? ctz64(z) : 64;
                    if (rl >= 64 - k) more[ln] = 1;          // run reaches the word's end
                    len[ln] = rl > 64 - k ? 64 - k : rl;
                    const int off = d - lit0;                // x - pred_k
                    dk[ln] = off < 0 ? (uint32_t)(-off) * 2u : (uint32_t)off * 2u + 1u;
                }
            }
        }
        uint32_t lm = w.ballot(more);
        while (lm) {
            const int l = ctz32(lm);
            lm &= lm - 1u;
            const int d = dbase + l;
            const int full = equal_len_coop(w, pv, qi, P0 + d + k, 64 - k);
            LZ_EACH(w, ln) if (ln == l) len[ln] = full;
        }
        const int bl = w.max_i32(len);
        if (bl > 0) {
            typename W::template Lanes<uint32_t> kk;
            LZ_EACH(w, ln) kk[ln] = len[ln] == bl ? dk[ln] : 0xFFFFFFFFu;
            const uint32_t bk = w.min_u32(kk);
            if (bl > sl || (bl == sl && bk < skey)) { sl = bl; skey = bk; }
        }
    }
    best_pos = 0;
    best_len = 0;
    if (sl > 0) {
        const int off = (skey & 1u) ? (int)(skey >> 1) : -(int)(skey >> 1);
        best_len = sl;
        best_pos = P0 + lit0 + k + off;
    }
}

// ---- the next anchor event at or after query position i0 (and below `lim`).
// Consecutive hit positions are almost always one longer match seen twice, and the second is then
// skipped by the walk.  So only the first position of a run of consecutive hits (a run never crosses
// a multiple of 8) gets a lane and a record; the run length (<= 7) rides in bits [22,25) of w2.
// Should the walk nevertheless come to stand inside a run -- or on one whose first hit does not
// count -- the event is "virtual": LIVE | DEFER, evaluated by anchor_lookup() when taken.
// need_pos: prediction mode, where an anchor at reference position 0 counts as none (:604).
constexpr uint32_t HIT_RUN_SHIFT = 22;
constexpr uint32_t HIT_X_SHIFT = 25;       // w2 [25,32): window offset of the first seed's reference position (w3 bit 31: not unique)
struct Event { int j, qi, a_pos; uint32_t w2, w3; };
template <class W, class LI, class LU>
LZ_HD Event next_event(W& w, const LI& r_p, const LU& r_pos, const LU& r_w2, const LU& r_w3, int i0, int lim, bool need_pos)
{
    typename W::template Lanes<int> c1, c2;
    LZ_EACH(w, ln)
    {
        const int p = r_p[ln];
        const uint32_t w2 = r_w2[ln];
        const int run = (int)((w2 >> HIT_RUN_SHIFT) & 7u);
        const bool counts = (w2 & HIT_LIVE) && (!need_pos || (w2 & HIT_DEFER) || r_pos[ln] != 0u);
        const bool first = p >= i0 && counts;
        c1[ln] = first && p < lim;
        // rare: the walk stands inside the run, or the run's first hit does not count
        c2[ln] = !first && run > 0 && p + run >= i0;
    }
    const uint32_t vm1 = w.ballot(c1), vm2 = w.ballot(c2);
    Event o;
    o.j = -1; o.qi = 0x7FFFFFFF; o.a_pos = 0; o.w2 = 0; o.w3 = 0;
    if (vm2 && (!vm1 || (vm2 & (0u - vm2)) < (vm1 & (0u - vm1)))) {
        const int j2 = ctz32(vm2);
        const int p2 = w.get(r_p, j2);
        const int e = p2 + 1 > i0 ? p2 + 1 : i0;
        if (e < lim) { o.j = j2; o.qi = e; o.w2 = HIT_LIVE | HIT_DEFER; }
        // (e >= lim: the records are in position order, so nothing else lies below lim either)
    } else if (vm1) {
        o.j = ctz32(vm1);
        o.qi = w.get(r_p, o.j);
        o.w2 = w.get(r_w2, o.j);
        o.w3 = w.get(r_w3, o.j);
        o.a_pos = (int)w.get(r_pos, o.j);
    }
    return o;
}

// ---- hit mask of the 32 query positions starting at pos0 (a multiple of 32): bit t is set iff the
// mal-mer at pos0+t has no N and its bit in the reference's anchor presence bitmap is set -- i.e. the
// probe chain of ht_long (parser.cpp:516-518) would not end at once.  Positions >= end are cleared.
template <class PT>
LZ_HD uint32_t scan_word(const PairView& pv, const PT& P, int pos0, int end)
{
    uint32_t m = 0;
    if (pos0 < end) {
        const Rec a = ld_rec(pv.q + (pos0 >> 5)), b = ld_rec(pv.q + (pos0 >> 5) + 1);
        // positions whose mal-mer covers an N (rare: one test per 32 positions)
        uint32_t bad = 0;
        if (a.z | b.z) {
            const uint64_t nz = ((uint64_t)b.z << 32) | a.z;
            uint64_t u = nz;
            for (int j = 1; j < P.mal; ++j) u |= nz >> j;
            bad = (uint32_t)u;
        }
#if defined(__CUDA_ARCH__)
#pragma unroll kScanUnroll
#endif
        for (int t = 0; t < 32; ++t) {     // small body: the kernel is I-cache sensitive
            const uint32_t lo = funnel_r(a.x, b.x, t), hi = funnel_r(a.y, b.y, t);
            const uint32_t bi = bitmap_index(kmer_key(lo, hi, P.mal), P.mal, pv.bbits);
            const uint32_t bw = ld_u32(pv.bitmap + (bi >> 5)) >> (bi & 31u);
            m = (m >> 1) | (bw << 31);      // bit of position t ends up at bit t
        }
        m &= ~bad;
        const int nv = end - pos0;
        if (nv < 32) m &= low_mask32(nv);
    }
    return m;
}
// ---- the same hit mask through a two-level bitmap: `coarse` (shared memory in k_scan_hits) is the
// reference's presence bitmap OR-folded down to at most 2^LZ_COARSE_BITS bits; only positions that pass it
// (about 4x the true hit rate for the default parameters) look at the exact bitmap in L2.
constexpr int LZ_COARSE_BITS = 20;                       // 128 KiB
LZ_HD uint32_t coarse_words(uint32_t bbits) { return bbits >= (uint32_t)LZ_COARSE_BITS ? (1u << (LZ_COARSE_BITS - 5)) : (bbits > 5 ? (1u << (bbits - 5)) : 1u); }
// word c of the folded bitmap: the OR of the exact words c, c + Cw, c + 2 Cw, ...
LZ_HD uint32_t coarse_fold_word(const uint32_t* bitmap, uint32_t bbits, uint32_t c)
{
    const uint32_t cw = coarse_words(bbits);
    const uint32_t ew = bbits > 5 ? (1u << (bbits - 5)) : 1u;
    uint32_t v = 0;
    for (uint32_t t = c; t < ew; t += cw) v |= ld_u32(bitmap + t);
    return v;
}
template <class PT>
LZ_HD uint32_t scan_word_coarse(const PairView& pv, const PT& P, int pos0, int end, const uint32_t* coarse)
{
    uint32_t m = 0;
    if (pos0 < end) {
        const Rec a = ld_rec(pv.q + (pos0 >> 5)), b = ld_rec(pv.q + (pos0 >> 5) + 1);
        uint32_t bad = 0;
        if (a.z | b.z) {
            const uint64_t nz = ((uint64_t)b.z << 32) | a.z;
            uint64_t u = nz;
            for (int j = 1; j < P.mal; ++j) u |= nz >> j;
            bad = (uint32_t)u;
        }
        const uint32_t cbits = pv.bbits < (uint32_t)LZ_COARSE_BITS ? pv.bbits : (uint32_t)LZ_COARSE_BITS;
        const uint32_t cmask = (1u << cbits) - 1u;
        if (2 * P.mal <= LZ_DIRECT_BITMAP_MAX_BITS) {
            // direct-addressed bitmap: index = key = low-plane bits | high-plane bits << mal.  With the high
            // plane shifted up by mal once per word the index is one funnel shift per plane and one select
            // (the bits above 2*mal are dropped by the fold mask); the byte offset into the folded copy is
            // taken straight from it.
            const uint32_t km = low_mask32(P.mal);
            const uint32_t h0 = a.y << P.mal, h1 = funnel_r(a.y, b.y, 32 - P.mal);
            const uint32_t omask = (cmask >> 5) << 2;
            const uint32_t bmask = cmask < 31u ? cmask : 31u;       // bitmaps below 32 bits (mal <= 2): the bit index needs masking too
#if defined(__CUDA_ARCH__)
#pragma unroll 8
#endif
            for (int t = 0; t < 32; ++t) {
                const uint32_t lo = funnel_r(a.x, b.x, t), hs = funnel_r(h0, h1, t);
                const uint32_t sel = (lo & km) | (hs & ~km);
                const uint32_t cwd = *reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(coarse) + ((sel >> 3) & omask));
                m = funnel_r(m, funnel_r(cwd, 0u, (int)(sel & bmask)), 1);
            }
        } else {
#if defined(__CUDA_ARCH__)
#pragma unroll 8
#endif
        for (int t = 0; t < 32; ++t) {
            const uint32_t lo = funnel_r(a.x, b.x, t), hi = funnel_r(a.y, b.y, t);
            const uint32_t ci = bitmap_index(kmer_key(lo, hi, P.mal), P.mal, pv.bbits) & cmask;
            const uint32_t bw = coarse[ci >> 5] >> (ci & 31u);
            m = (m >> 1) | (bw << 31);
        }
        }
        m &= ~bad;
        const int nv = end - pos0;
        if (nv < 32) m &= low_mask32(nv);
        if (pv.bbits > (uint32_t)LZ_COARSE_BITS) {
            uint32_t t = m;
            while (t) {                     // the exact bitmap decides
                const int j = ctz32(t);
                t &= t - 1u;
                const uint32_t lo = funnel_r(a.x, b.x, j), hi = funnel_r(a.y, b.y, j);
                const uint32_t bi = bitmap_index(kmer_key(lo, hi, P.mal), P.mal, pv.bbits);
                if (!((ld_u32(pv.bitmap + (bi >> 5)) >> (bi & 31u)) & 1u)) m &= ~(1u << j);
            }
        }
    }
    return m;
}
// ---- 32 x 32 bit-matrix transpose in registers: afterwards bit t of a[i] is what bit i of a[t] was.
// Used by the all2all form of the scan (k_bitmap_matrix / k_scan_matrix): with the presence bitmaps of 32
// references transposed into one 32-bit word per k-mer, ONE gather per query position answers the probe for
// all 32 references, and a second transpose turns the 32 gathered words of a query word into the 32 hit
// masks (one per reference) that scan_word() computes one probe at a time.
LZ_HD void transpose32(uint32_t (&a)[32])
{
    uint32_t m = 0x0000FFFFu;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int s = 16; s; s >>= 1) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int k = 0; k < 32; ++k) {
            if (k & s) continue;
            // swap the high-s columns of row k with the low-s columns of row k+s
            const uint32_t t = ((a[k] >> s) ^ a[k + s]) & m;
            a[k + s] ^= t;
            a[k] ^= t << s;
        }
        m ^= m << (s >> 1);          // 0x0000FFFF -> 0x00FF00FF -> 0x0F0F0F0F -> 0x33333333 -> 0x55555555
    }
}
// what scan_word() masks away, for the 32 positions starting at pos0 of a query whose loop ends at `end`:
// positions whose mal-mer covers an N, positions >= end
LZ_HD uint32_t scan_valid_mask(const Rec& a, const Rec& b, int mal, int pos0, int end)
{
    uint32_t bad = 0;
    if (a.z | b.z) {
        const uint64_t nz = ((uint64_t)b.z << 32) | a.z;
        uint64_t u = nz;
        for (int j = 1; j < mal; ++j) u |= nz >> j;
        bad = (uint32_t)u;
    }
    const int nv = end - pos0;
    return ~bad & (nv >= 32 ? 0xFFFFFFFFu : (nv <= 0 ? 0u : low_mask32(nv)));
}
// first positions of the runs of consecutive hits inside a hit mask (a run never crosses a multiple of 8)
LZ_HD uint32_t run_first_bits(uint32_t m) { return m & ~(m & (m << 1) & ~0x01010101u); }

// ---- staged pipeline (kernels.cu: k_scan_hits -> k_expand_hits -> k_hit_records -> k_walk_pairs).
// Everything a hit record holds depends on the query position alone, never on the parse state, so the
// records of ALL hits of a chunk of pairs are computed by plain data-parallel kernels (one thread per
// 32 query positions, then one thread per hit run) and stored in HBM in position order; the walk kernel
// then only loads them, 32 at a time.  A stored record is one Rec: x = query position, y = reference
// position of the best anchor, z = w2, w = w3 (the fields of hit_probe / hit_extend / spec_first_seed).
// (HitEntry, the hit-list element, is declared in layout.h.)

// length of the run of hits that follow the run-first bit b of hit mask m (0..7)
LZ_HD int run_following(uint32_t m, int b)
{
    const uint32_t sec = m & (m << 1) & ~0x01010101u;
    const uint32_t follow = b < 31 ? ~(sec >> (b + 1)) : 1u;     // run: sec bits right behind b
    return ctz32(follow);
}

// The stored record of hit-list entry t, stages A and B (k_hit_probe, one thread per entry): table probe,
// match length, both approximate extensions.  `sync` is the reconvergence point between the stages
// (lanes take different numbers of probe iterations).
template <class PT, class SyncFn>
LZ_HD Rec hit_record_probe(const PairView& pv, const PT& P, const HitEntry* list, uint64_t t, bool live, SyncFn sync)
{
    Rec o; o.x = 0; o.y = 0; o.z = 0; o.w = 0;
    HitEntry e; e.pair = 0; e.pr = 0;
    if (live) e = list[t];
    const int p = (int)(e.pr & (uint32_t)(LZ_STAGED_MAX_POS - 1));
    uint32_t pos = 0, w2 = 0, w3 = 0;
    if (live) hit_probe(pv, P, p, pos, w2);
    w2 |= (e.pr >> HIT_ENTRY_RUN_SHIFT) << HIT_RUN_SHIFT;
    sync();
    if (live && (w2 & (HIT_LIVE | HIT_DEFER)) == HIT_LIVE) hit_extend(pv, P, p, (int)pos, w2, w3);
    o.x = (uint32_t)p; o.y = pos; o.z = w2; o.w = w3;
    return o;
}

// Stage C (k_hit_spec): the speculative seed scan of the window the hit opens if it is taken.  The job is
// described first (cheap: where the window starts, how far to scan, the tentative successor link) so that the
// scans of a thread block can be dealt to its threads in order of their length (lanes of a warp then run
// about the same number of rows), then run (spec_first_seed) and folded into the record.
struct SpecJob { int si0, sP0, K; uint32_t succ; bool need; };
template <class PT>
LZ_HD SpecJob hit_spec_job(const PT& P, int n_loop, const HitEntry* list, uint64_t t, uint64_t total, const Rec& rec)
{
    SpecJob j; j.si0 = 0; j.sP0 = 0; j.K = -1; j.succ = 0; j.need = false;
    if (!(PT::kSpecSeeds && P.mqd <= SPEC_MAX_K && (rec.z & HIT_FWD))) return j;
    const int adv = (int)(rec.z & 63u) + (int)((rec.z >> 6) & 63u);
    j.si0 = (int)rec.x + adv;
    j.sP0 = (int)rec.y + adv;
    // scanned up to the next hit run of this pair at or after the window's start (inclusive), or the last
    // position of the loop
    int lim_p = n_loop - 1;
    bool inside = false;                 // the walk would come to stand inside a run on the way
    const uint32_t pair = list[t].pair;
    for (uint64_t u = t + 1; u < total && list[u].pair == pair; ++u) {
        const int pu = (int)(list[u].pr & (uint32_t)(LZ_STAGED_MAX_POS - 1));
        if (pu >= j.si0) {
            lim_p = pu;
            if (!inside && u - t <= 7u) j.succ = (uint32_t)(u - t);
            break;
        }
        if (pu + (int)(list[u].pr >> HIT_ENTRY_RUN_SHIFT) >= j.si0) inside = true;
    }
    j.K = lim_p - j.si0;
    if (j.K > P.mqd) j.K = P.mqd;
    j.need = true;
    return j;
}
// ks = spec_first_seed(pv, P, j.si0, j.sP0, j.K, P.mrd)
LZ_HD void hit_spec_finish(Rec& rec, const SpecJob& j, int ks)
{
    if (ks < 0) return;              // no speculative seed record: the walk goes through the general code (no link)
    rec.w |= ((uint32_t)(j.K + 1) << 18) | ((uint32_t)(ks & 127) << 24) | ((ks & SPEC_MULTI) ? 0x80000000u : 0u);
    rec.z |= (uint32_t)((ks >> 8) & 127) << HIT_X_SHIFT;
    rec.x |= j.succ << HIT_ENTRY_RUN_SHIFT;
}

// ---- successor links (decided at the end of k_hit_spec, one thread per record).  hit_record() left in x[29,32) the offset of
// the first record at or after the window its record opens (0: none).  If the walk takes record `a` entirely
// from its fields (the fast path of parse_pair), what happens next depends on nothing but a and that
// successor b: this decides whether it is again one of the two plain fast-path events -- b inside the window,
// no seed before or at it, far from the prediction; or nothing for mqd+1 positions (the prediction expires)
// and b next -- with b's backward record usable whatever region was kept last (the literal run before b is at
// least qi_b - si0_a long).  The offset is kept only then; the walk follows such links without re-deriving
// anything (parse_pair, "chain").
template <class PT>
LZ_HD uint32_t verified_link(const PT& P, const Rec& a, const Rec& b)
{
    const uint32_t d = a.x >> HIT_ENTRY_RUN_SHIFT;
    if (!d) return 0;
    if ((a.z & (HIT_DEFER | HIT_FWD | HIT_BWD)) != (HIT_FWD | HIT_BWD)) return 0;
    if ((b.z & (HIT_LIVE | HIT_DEFER | HIT_FWD | HIT_BWD)) != (HIT_LIVE | HIT_FWD | HIT_BWD)) return 0;
    const int adv = (int)(a.z & 63u) + (int)((a.z >> 6) & 63u);
    const int si0 = (int)(a.x & (uint32_t)(LZ_STAGED_MAX_POS - 1)) + adv, sP0 = (int)a.y + adv;
    const int K1 = (int)((a.w >> 18) & 63u), ks = (int)((a.w >> 24) & 127u);
    const int qb = (int)(b.x & (uint32_t)(LZ_STAGED_MAX_POS - 1)), pb = (int)b.y;
    const int ka = qb - si0;
    if (ka <= 0 || K1 <= 0) return 0;
    if (ka > P.mqd) {
        if (!(K1 > P.mqd && ks >= 64)) return 0;
    } else {
        int dist = pb - (sP0 + ka);
        if (dist < 0) dist = -dist;
        if (!(ka < K1 && ks > ka && pb != 0 && dist > P.mrd)) return 0;
    }
    int lim = ka < qb ? ka : qb;
    if (pb < lim) lim = pb;
    if (lim < (int)((b.w >> 12) & 63u)) return 0;
    return d;
}

// where parse_pair() gets its batches of hit records from
struct InlineBatches {                   // built by the warp itself, window by window (one-kernel path, regions kernel)
    static constexpr bool kStored = false;
    const Rec* recs = nullptr; uint32_t n = 0;
};
struct StoredBatches {                   // loaded from the records the staged kernels wrote
    static constexpr bool kStored = true;
    const Rec* recs; uint32_t n;         // this pair's records, position order
};

// ---- the whole pair.
//
// Three nested granularities:
//   window (<= 1024 query positions): every lane probes the anchor presence bitmap for its own
//          32 consecutive positions -> hit masks, one word per lane;
//   batch  (<= 32 hits of the window, in position order): lane j evaluates hit j on its own
//          (hit_probe, hit_extend, spec_first_seed) -- table probe, match length, both approximate extensions -- ahead of
//          the walk; this is where the lanes do independent work;
//   walk   : the sequential state machine of parser.cpp:501-708 over chunks of <= 65-msl
//          positions, consuming the batch records in order.  Without a prediction it jumps
//          from hit to hit; with one (the mqd positions after a match) it also looks for seeds
//          along the diagonals of the prediction window.
// `list` is 32 uint16 of per-warp scratch (shared memory on the device).
template <class W, class PT, class RS, class BS>
LZ_HD PairResult parse_pair(W& w, const PairView& pv, const PT& P, WarpScratch* sc, RS& rt, const BS& bs)
{
    uint16_t* list = BS::kStored ? nullptr : sc->list;
    if (RS::kActive) rt.clear_all();
    ParseState s;
    s.i = 0; s.pred = -1; s.lit = 0;
    s.rs = -1; s.re = 0; s.kept_end = 0;
    s.curM = s.curX = 0;
    s.totM = s.totX = s.totN = 0;
    const int Dq = pv.Dq;
    const int n_loop = Dq - P.msl;     // loop runs while i < n_loop (i + msl < Dq, :501)
    const int NC = 65 - P.msl;         // chunk size: msl-runs starting in the chunk fit 64 bits

    // speculative seed record of the match just taken; spec_K1 > 0 <=> it describes the window at s.i
    // (cleared by every literal step)
    int spec_K1 = 0, spec_ks = 64, spec_x = -1;
    int chain_nx = 32;                 // its successor link (32: none)
    int win_base = 0, win_end = 0;     // hit masks cover [win_base, win_end)
    int b_cover = 0;                   // every hit in [batch start, b_cover) has a record
    typename W::template Lanes<uint32_t> r_pos, r_w2, r_w3;
    typename W::template Lanes<int> r_p, r_nx;     // r_nx: successor links (see "successor links" below)
    typename W::template Lanes<uint32_t> r_kv;     // stored batches: chain summary of each record
    uint32_t cur = 0;                  // stored batches: first record of the current batch
    int b_loaded = 0;                  //                 and how many records it holds
    LZ_EACH(w, ln)
    {
        if (!BS::kStored) for (int q = 0; q < LZ_WIN_WORDS; ++q) sc->hm[q][ln] = 0;
        r_pos[ln] = 0; r_w2[ln] = 0; r_w3[ln] = 0; r_p[ln] = 0x7FFFFFFF; r_nx[ln] = 32; r_kv[ln] = 0;
    }

    for (;;) {
        const int i0 = s.i;
        // the match this step finds, if any (one take_match call site keeps the code small)
        bool matched = false, seed_cut = false;
        int m_qi = 0, m_lit = 0, m_pred = -1, m_pos = 0, m_len = 0;
        uint32_t m_w2 = 0, m_w3 = 0;
        int m_j = -1;                  // batch record the match was taken from (when m_w2 != 0)
        if (i0 >= b_cover) {
            if (i0 >= n_loop) break;       // b_cover <= n_loop: the end of the loop (:501) is only met here
            if (BS::kStored) {
                // ---- next batch: the first <= 32 stored records not entirely behind the walk.  The records
                // are in position order and runs do not overlap, so "behind" is a prefix property.
                cur += (uint32_t)b_loaded;             // b_cover <= i0: the whole previous batch lies behind
                for (;;) {
                    typename W::template Lanes<int> behind;
                    LZ_EACH(w, ln)
                    {
                        const uint32_t idx = cur + (uint32_t)ln;
                        Rec h; h.x = 0x7FFFFFFFu; h.y = 0; h.z = 0; h.w = 0;
                        if (idx < bs.n) h = ld_rec(bs.recs + idx);
                        const int hp = (int)(h.x & 0x7FFFFFFFu & ~(7u << HIT_ENTRY_RUN_SHIFT) ) | (idx < bs.n ? 0 : 0x7FFFFFFF);
                        r_p[ln] = hp; r_pos[ln] = h.y; r_w2[ln] = h.z; r_w3[ln] = h.w;
                        // chain summary: matches / literals of the region the record opens when taken from its
                        // fields, whether that region is kept, link offset
                        {
                            const int len = (int)(h.z & 63u), fwd = (int)((h.z >> 6) & 63u), fm = (int)((h.z >> 12) & 63u);
                            const int back = (int)(h.w & 63u), bm = (int)((h.w >> 6) & 63u);
                            r_kv[ln] = (uint32_t)(bm + len + fm) | ((uint32_t)(back - bm + fwd - fm) << 8) |
                                       (back + len + fwd >= P.reg ? KV_KEEP : 0u) | ((h.x >> HIT_ENTRY_RUN_SHIFT) << KV_LINK_SHIFT);
                        }
                        behind[ln] = idx < bs.n && hp + (int)((h.z >> HIT_RUN_SHIFT) & 7u) < i0;
                    }
                    const int nb = popc32(w.ballot(behind));
                    if (!nb) break;
                    cur += (uint32_t)nb;
                    if (nb == 32) {
                        // a long match jumped far ahead: gallop over blocks of 32 records
                        for (;;) {
                            typename W::template Lanes<int> far;
                            LZ_EACH(w, ln)
                            {
                                const uint64_t idx = (uint64_t)cur + 32u * (uint32_t)ln + 31u;
                                far[ln] = 0;
                                if (idx < bs.n) {
                                    const Rec h = ld_rec(bs.recs + idx);
                                    far[ln] = (int)(h.x & (uint32_t)(LZ_STAGED_MAX_POS - 1)) + (int)((h.z >> HIT_RUN_SHIFT) & 7u) < i0;
                                }
                            }
                            const int nf = popc32(w.ballot(far));
                            cur += 32u * (uint32_t)nf;
                            if (nf < 32) break;
                        }
                    }
                }
                const uint32_t left = bs.n - cur;
                LZ_EACH(w, ln)
                {
                    // the next batch: 32 records = four 128-byte lines, most likely still in HBM
                    if (ln < 4 && 32u + 8u * (uint32_t)ln < left) w.prefetch(bs.recs + cur + 32u + 8u * (uint32_t)ln);
                }
                b_loaded = left < 32u ? (int)left : 32;
                b_cover = left > 32u ? w.get(r_p, 31) + (int)((w.get(r_w2, 31) >> HIT_RUN_SHIFT) & 7u) + 1 : n_loop;
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
                lz_batches++; lz_hits += b_loaded;
#endif
            } else {
            if (i0 >= win_end) {
                // ---- new window: lane l probes positions win_base + LZ_WIN_LANE*l + [0, LZ_WIN_LANE)
                // (large enough that batches of 32 hits are normally full: ~2.4 % of the positions of
                // an unrelated pair hit the bitmap)
                win_base = i0 & ~31;
                win_end = win_base + 32 * LZ_WIN_LANE < n_loop ? win_base + 32 * LZ_WIN_LANE : n_loop;
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
                for (int q = 0; q < LZ_WIN_WORDS; ++q) {
                    LZ_EACH(w, ln)
                    {
                        sc->hm[q][ln] = scan_word(pv, P, win_base + LZ_WIN_LANE * ln + 32 * q, win_end);
                    }
                }
            }
            // ---- new batch: the first <= 32 hit runs at or after i0
            typename W::template Lanes<int> cnt, pre;
            LZ_EACH(w, ln)
            {
                int c = 0;
                LZ_ROLLED
                for (int q = 0; q < LZ_WIN_WORDS; ++q) {
                    const int rel = i0 - (win_base + LZ_WIN_LANE * ln + 32 * q);
                    if (rel >= 32) sc->hm[q][ln] = 0;                        // behind the walk: never needed again
                    else if (rel > 0) sc->hm[q][ln] &= ~low_mask32(rel);
                    const uint32_t m = sc->hm[q][ln];
                    c += popc32(m & ~(m & (m << 1) & ~0x01010101u));     // first positions of runs
                }
                cnt[ln] = c;
            }
            w.exscan_add(pre, cnt);
            const int total = w.add_i32(cnt);
            LZ_EACH(w, ln)
            {
                int idx = pre[ln];
                LZ_ROLLED
                for (int q = 0; q < LZ_WIN_WORDS && idx < 32; ++q) {
                    const uint32_t hmw = sc->hm[q][ln];
                    const uint32_t sec = hmw & (hmw << 1) & ~0x01010101u;
                    uint32_t m = hmw & ~sec;
                    while (m && idx < 32) {
                        const int b = ctz32(m);
                        const uint32_t follow = b < 31 ? ~(sec >> (b + 1)) : 1u;     // run: sec bits right behind b
                        list[idx++] = (uint16_t)((LZ_WIN_LANE * ln + 32 * q + b) | (ctz32(follow) << 13));
                        m &= m - 1u;
                    }
                }
            }
            w.sync();
            const int b_cnt = total < 32 ? total : 32;
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
            lz_batches++; lz_hits += b_cnt;
#endif
            typename W::template Lanes<int> r_run;
            LZ_EACH(w, ln)
            {
                r_p[ln] = ln < b_cnt ? win_base + (int)(list[ln] & 0x1FFFu) : 0x7FFFFFFF;
                r_run[ln] = ln < b_cnt ? (int)(list[ln] >> 13) : 0;
            }
            w.sync();
            b_cover = total <= 32 ? win_end : w.get(r_p, 31) + w.get(r_run, 31) + 1;
            // three stages, each entered by the whole warp (sync = reconvergence point): lanes take
            // different numbers of probe / scan iterations and must not stay split afterwards
            LZ_EACH(w, ln)
            {
                uint32_t a = 0, b = 0;
                if (ln < b_cnt) hit_probe(pv, P, r_p[ln], a, b);
                r_pos[ln] = a; r_w2[ln] = b | ((uint32_t)r_run[ln] << HIT_RUN_SHIFT); r_w3[ln] = 0;
            }
            w.sync();
            if (!RS::kActive)       // the regions kernel needs the columns, not their counts: no records
            LZ_EACH(w, ln)
            {
                if ((r_w2[ln] & (HIT_LIVE | HIT_DEFER)) == HIT_LIVE) hit_extend(pv, P, r_p[ln], (int)r_pos[ln], r_w2[ln], r_w3[ln]);
            }
            w.sync();
            if (PT::kSpecSeeds && !RS::kActive && P.mqd <= SPEC_MAX_K) {
                typename W::template Lanes<int> tc;      // the record the walk meets next if it takes this one
                LZ_EACH(w, ln)
                {
                    tc[ln] = 32;
                    if (r_w2[ln] & HIT_FWD) {
                        // the prediction window this hit opens if it is taken (see spec_first_seed):
                        // scanned up to the next hit of the batch (inclusive) or the end of what the
                        // batch covers
                        const uint32_t b = r_w2[ln];
                        const int adv = (int)(b & 63u) + (int)((b >> 6) & 63u);
                        const int si0 = r_p[ln] + adv, sP0 = (int)r_pos[ln] + adv;
                        int idx = ln + 1;
                        while (idx < b_cnt && win_base + (int)(list[idx] & 0x1FFFu) < si0) ++idx;
                        int K = idx < b_cnt ? win_base + (int)(list[idx] & 0x1FFFu) - si0 : b_cover - 1 - si0;
                        if (K > P.mqd) K = P.mqd;
                        const int ks = spec_first_seed(pv, P, si0, sP0, K, P.mrd);
                        if (ks >= 0) {
                            r_w3[ln] |= ((uint32_t)(K + 1) << 18) | ((uint32_t)(ks & 127) << 24) | ((ks & SPEC_MULTI) ? 0x80000000u : 0u);
                            r_w2[ln] |= (uint32_t)((ks >> 8) & 127) << HIT_X_SHIFT;
                            // not when the walk would come to stand inside the run of the record before idx
                            const int prev = idx - 1;
                            const bool inside = prev > ln && win_base + (int)(list[prev] & 0x1FFFu) + (int)(list[prev] >> 13) >= si0;
                            if (idx < b_cnt && !inside) tc[ln] = idx;
                        }
                    }
                }
                w.sync();
                // ---- successor links.  If the walk takes record ln entirely from its fields (fast path),
                // what happens next depends on nothing but the batch: the next record t it meets, and
                // whether that is one of the two plain events of the fast path below (a distant anchor
                // taken from ITS record, possibly after the prediction expired).  Lane ln decides that
                // now, for all records at once; the walk then follows nx[] instead of searching.
                if (kChain) {
                typename W::template Lanes<uint32_t> t_w2, t_pos;
                w.gather(t_w2, r_w2, tc);
                w.gather(t_pos, r_pos, tc);
                LZ_EACH(w, ln)
                {
                    int nx = 32;
                    if (tc[ln] < 32 && (t_w2[ln] & (HIT_DEFER | HIT_FWD)) == HIT_FWD) {
                        const uint32_t b = r_w2[ln], c = r_w3[ln];
                        const int adv = (int)(b & 63u) + (int)((b >> 6) & 63u);
                        const int si0 = r_p[ln] + adv, sP0 = (int)r_pos[ln] + adv;
                        const int K1 = (int)((c >> 18) & 63u), ks = (int)((c >> 24) & 127u);
                        const int ka = win_base + (int)(list[tc[ln]] & 0x1FFFu) - si0;
                        bool ok;
                        if (ka > P.mqd) ok = K1 > P.mqd && ks >= 64;           // the prediction expires first
                        else {
                            int dist = (int)t_pos[ln] - (sP0 + ka);
                            if (dist < 0) dist = -dist;
                            ok = ka < K1 && ks > ka && t_pos[ln] != 0u && dist > P.mrd;
                        }
                        if (ok) nx = tc[ln];
                    }
                    r_nx[ln] = nx;
                }
                }
            }
            }   // !BS::kStored
            chain_nx = 32;
        }
        // what the speculative seed record says about the positions this batch covers (a stored record's
        // scan may reach beyond the batch; an inline one never does)
        int sK1 = spec_K1;
        if (BS::kStored && sK1 > b_cover - i0) sK1 = b_cover - i0;
        const int sks = spec_ks < sK1 ? spec_ks : 64;

        // ---- FAST PATH.  The walk stands right behind an anchor match taken from its record
        // (prediction on, no literals yet, seeds of the window known from the speculative scan) and
        // the next thing that happens is one of the two plain events of an unrelated pair:
        //   * the next anchor record lies inside the window, no seed before or at it, and it is far
        //     from the prediction (> mrd): a distant match taken entirely from its record;
        //   * nothing at all for mqd+1 positions: the prediction expires.
        // Everything else (seeds, close matches, deferred or incomplete records) falls through to the
        // general code below, which re-derives the same step.
        if (PT::kSpecSeeds && !RS::kActive && (s.pred < 0 || sK1 > 0)) {
            Event ev;
            bool plain;
            int nx_next = -1;          // successor link of the event, when already fetched
            bool expired = false;      // the state moved on inside this step: i0 is stale
            const bool chained = kChain && chain_nx < 32 && s.pred >= 0;
            if (chained) {
                // the successor link of the record just taken: a plain event, decided in the batch stage
                ev.j = chain_nx;
                ev.qi = w.get(r_p, ev.j);
                ev.w2 = w.get(r_w2, ev.j);
                ev.w3 = w.get(r_w3, ev.j);
                ev.a_pos = (int)w.get(r_pos, ev.j);
                nx_next = w.get(r_nx, ev.j);
                plain = true;
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
                {   // TEST ONLY (tests/emul): the link must be what the search below finds
                    Event e2 = next_event(w, r_p, r_pos, r_w2, r_w3, i0, 0x7FFFFFFF, false);
                    bool ok = e2.j >= 0 && (e2.w2 & (HIT_DEFER | HIT_FWD)) == HIT_FWD;
                    const int ka2 = e2.qi - i0;
                    if (ok && ka2 > P.mqd) {
                        ok = spec_K1 > P.mqd && spec_ks >= 64;
                        if (ok) e2 = next_event(w, r_p, r_pos, r_w2, r_w3, i0 + P.mqd + 1, 0x7FFFFFFF, false);
                        ok = ok && e2.j >= 0 && (e2.w2 & (HIT_DEFER | HIT_FWD)) == HIT_FWD;
                    } else if (ok) {
                        int dist = e2.a_pos - (s.pred + ka2);
                        if (dist < 0) dist = -dist;
                        ok = ka2 < spec_K1 && spec_ks > ka2 && e2.a_pos != 0 && dist > P.mrd;
                    }
                    if (!ok || e2.j != ev.j || e2.qi != ev.qi) lz_spec_mismatches++;
                    lz_chain++;
                }
#endif
            } else {
            ev = next_event(w, r_p, r_pos, r_w2, r_w3, i0, 0x7FFFFFFF, false);
            const int ka = ev.qi - i0;
            plain = (ev.w2 & (HIT_DEFER | HIT_FWD)) == HIT_FWD;     // a complete record
            if (s.pred >= 0) {
                if (ka > P.mqd) {
                    if (sK1 > P.mqd && sks >= 64) {
                        // mqd+1 literal steps, then the prediction is dropped (:699-707) ...
                        s.i = i0 + P.mqd + 1;
                        s.lit = P.mqd + 1;
                        s.pred = -1;
                        spec_K1 = 0;
                        // ... and the same event, still ahead (ka > mqd), is the next one without a prediction
                        expired = true;
                        if (!plain) continue;
                    } else plain = false;
                } else {
                    int dist = ev.a_pos - (s.pred + ka);
                    if (dist < 0) dist = -dist;
                    plain = plain && ka < sK1 && sks > ka && ev.a_pos != 0 && dist > P.mrd;
                }
            }
            }
            const int qi = ev.qi;
            const uint32_t w2 = ev.w2;
            const int a_pos = ev.a_pos;
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
            if (!plain) {   // TEST ONLY: why the fast path was left
                const int ka = qi - i0;
                if (ev.j < 0) lz_why[0]++;
                else if (w2 & HIT_DEFER) lz_why[1]++;
                else if (!(w2 & HIT_FWD)) lz_why[2]++;
                else if (s.pred >= 0 && ka > P.mqd) lz_why[3]++;
                else if (s.pred >= 0 && ka >= spec_K1) lz_why[4]++;
                else if (s.pred >= 0 && spec_ks <= ka) lz_why[5]++;
                else if (s.pred >= 0 && a_pos == 0) lz_why[6]++;
                else lz_why[7]++;          // close anchor
            }
#endif
            // ---- SEED SHORT CUT.  Prediction on, the record of the match just taken names the first seed of its
            // window (position sks, the only one there: reference offset spec_x) and no anchor event lies at or
            // before it: the step the general code below would take is exactly that seed match
            // (parser.cpp:548-580, no arbitration), so measure it and go straight to take_match().
            if (BS::kStored && !plain && !expired && s.pred >= 0 && s.lit == 0 && sks < 64 && spec_x >= 0 && qi - i0 > sks) {
                const int sq = i0 + sks, sp = s.pred + spec_x;
                const uint32_t m0 = mism32(pv.q, pv.Dq, sq, pv.r, pv.Dr, sp);
                const int sl = m0 ? ctz32(m0) : equal_len_coop(w, pv, sq, sp, 32);
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
                {   // TEST ONLY (tests/emul): the general code's answer for this step
                    const Planes64 Qc = ext64(pv.q, i0);
                    int cp = 0, cl = 0;
                    uint64_t chk = 0;
                    for (int d = -sks; d < P.mrd; ++d) chk |= run_starts(diag_equal64(pv, Qc, i0, s.pred, d), P.msl);
                    best_seed(w, pv, P, Qc, i0, s.pred, 0, sks, -sks, P.mrd, cp, cl);
                    const Event e2 = next_event(w, r_p, r_pos, r_w2, r_w3, i0, i0 + sks + 1, true);
                    if ((chk & low_mask64(sks + 1)) != (1ull << sks) || cp != sp || cl != sl || e2.j >= 0 || sl < P.msl) lz_spec_mismatches++;
                    lz_why[10]++;
                }
#endif
                matched = true;
                seed_cut = true;
                m_qi = sq; m_lit = sks; m_pred = s.pred + sks; m_pos = sp; m_len = sl; m_w2 = 0; m_w3 = 0; m_j = -1;
            }
            if (plain) {
                const uint32_t w3 = ev.w3;
                const bool keep = s.rs >= 0 && s.re - s.rs >= P.reg;
                const int kept_end = keep ? s.re : s.kept_end;
                const int run = qi - kept_end;
                int lim = run < qi ? run : qi;
                if (a_pos < lim) lim = a_pos;
                if (run <= 0 || ((w2 & HIT_BWD) && lim >= (int)((w3 >> 12) & 63u))) {
                    if (keep) { s.totM += s.curM; s.totX += s.curX; ++s.totN; s.kept_end = kept_end; }
                    const int back = run > 0 ? (int)(w3 & 63u) : 0, bm = run > 0 ? (int)((w3 >> 6) & 63u) : 0;
                    const int len = (int)(w2 & 63u), fwd = (int)((w2 >> 6) & 63u), fm = (int)((w2 >> 12) & 63u);
                    s.curM = bm + len + fm;
                    s.curX = back - bm + fwd - fm;
                    s.rs = qi - back;
                    s.i = qi + len + fwd;
                    s.pred = a_pos + len + fwd;
                    s.lit = 0;
                    s.re = s.i;
                    spec_K1 = (int)((w3 >> 18) & 63u);
                    spec_ks = (int)((w3 >> 24) & 127u);
                    spec_x = (w3 >> 31) ? -1 : (int)(w2 >> HIT_X_SHIFT);
                    if (nx_next < 0) nx_next = w.get(r_nx, ev.j);
                    chain_nx = spec_K1 > 0 ? nx_next : 32;
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
                    lz_matches++; lz_fast++;
#endif
                    if (BS::kStored) {
                        // ---- chain: verified successor links (verified_link) inside the batch.  Every step is a
                        // distant match taken from its record right after one taken from its record: the region of
                        // the record left behind is judged by its length, nothing else carries over.
                        int j = ev.j;
                        uint32_t v = w.get(r_kv, j);
                        int d = (int)(v >> KV_LINK_SHIFT);
                        if (d && run > 0 && j + d < b_loaded) {      // (run > 0: the entry record's region is the one its summary describes)
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
                            ParseState chk = s;      // TEST ONLY (tests/emul): the serial fast path, step by step
                            int chk_j = j;
#endif
                            do {
                                if (v & KV_KEEP) {
                                    s.totM += (int)(v & 255u);
                                    s.totX += (int)((v >> 8) & 255u);
                                    ++s.totN;
                                    const uint32_t jw2 = w.get(r_w2, j);
                                    s.kept_end = w.get(r_p, j) + (int)(jw2 & 63u) + (int)((jw2 >> 6) & 63u);
                                }
                                j += d;
                                v = w.get(r_kv, j);
                                d = (int)(v >> KV_LINK_SHIFT);
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
                                lz_matches++; lz_fast++; lz_chain++;
#endif
                            } while (d && j + d < b_loaded);
                            const int cq = w.get(r_p, j), cp = (int)w.get(r_pos, j);
                            const uint32_t cw2 = w.get(r_w2, j), cw3 = w.get(r_w3, j);
                            const int back = (int)(cw3 & 63u), len = (int)(cw2 & 63u), fwd = (int)((cw2 >> 6) & 63u);
                            s.curM = (int)(v & 255u);
                            s.curX = (int)((v >> 8) & 255u);
                            s.rs = cq - back;
                            s.i = cq + len + fwd;
                            s.pred = cp + len + fwd;
                            s.re = s.i;
                            spec_K1 = (int)((cw3 >> 18) & 63u);
                            spec_ks = (int)((cw3 >> 24) & 127u);
                            spec_x = (cw3 >> 31) ? -1 : (int)(cw2 >> HIT_X_SHIFT);
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
                            {   // TEST ONLY: replay the same steps with the serial rules of the fast path
                                while (chk_j != j) {
                                    const int dd = (int)(w.get(r_kv, chk_j) >> KV_LINK_SHIFT);
                                    const int t = chk_j + dd;
                                    const int tq = w.get(r_p, t), tp = (int)w.get(r_pos, t);
                                    const uint32_t tw2 = w.get(r_w2, t), tw3 = w.get(r_w3, t);
                                    const uint32_t aw2 = w.get(r_w2, chk_j), aw3 = w.get(r_w3, chk_j);
                                    const int K1 = (int)((aw3 >> 18) & 63u), ks = (int)((aw3 >> 24) & 127u);
                                    const int ka = tq - chk.i;
                                    bool ok = (tw2 & (HIT_LIVE | HIT_DEFER | HIT_FWD)) == (HIT_LIVE | HIT_FWD) && K1 > 0 && (aw2 & HIT_FWD);
                                    const Event e2 = next_event(w, r_p, r_pos, r_w2, r_w3, chk.i, 0x7FFFFFFF, false);
                                    if (e2.j != t || e2.qi != tq) ok = false;
                                    if (ka > P.mqd) ok = ok && K1 > P.mqd && ks >= 64;
                                    else {
                                        int dist = tp - (chk.pred + ka);
                                        if (dist < 0) dist = -dist;
                                        ok = ok && ka < K1 && ks > ka && tp != 0 && dist > P.mrd;
                                    }
                                    const bool keep = chk.rs >= 0 && chk.re - chk.rs >= P.reg;
                                    const int kept_end = keep ? chk.re : chk.kept_end;
                                    const int run = tq - kept_end;
                                    int lim = run < tq ? run : tq;
                                    if (tp < lim) lim = tp;
                                    ok = ok && run > 0 && (tw2 & HIT_BWD) && lim >= (int)((tw3 >> 12) & 63u);
                                    if (!ok) lz_spec_mismatches++;
                                    if (keep) { chk.totM += chk.curM; chk.totX += chk.curX; ++chk.totN; chk.kept_end = kept_end; }
                                    const int tb = (int)(tw3 & 63u), tbm = (int)((tw3 >> 6) & 63u);
                                    const int tl = (int)(tw2 & 63u), tf = (int)((tw2 >> 6) & 63u), tfm = (int)((tw2 >> 12) & 63u);
                                    chk.curM = tbm + tl + tfm; chk.curX = tb - tbm + tf - tfm;
                                    chk.rs = tq - tb; chk.i = tq + tl + tf; chk.pred = tp + tl + tf; chk.lit = 0; chk.re = chk.i;
                                    chk_j = t;
                                }
                                if (chk.i != s.i || chk.pred != s.pred || chk.rs != s.rs || chk.re != s.re || chk.kept_end != s.kept_end ||
                                    chk.curM != s.curM || chk.curX != s.curX || chk.totM != s.totM || chk.totX != s.totX || chk.totN != s.totN)
                                    lz_spec_mismatches++;
                            }
#endif
                        }
                    }
                    continue;
                }
                chain_nx = 32;             // backward record not usable: the general code takes this step
                if (expired) continue;
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
                lz_why[8]++;
#endif
            }
        }
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
        else lz_why[9]++;                  // no speculative record for this state
#endif

        if (seed_cut) {
            // the match was found above
        } else if (s.pred < 0) {
            // ---- no prediction (:507-532): the next event is the next live anchor record
            const Event ev = next_event(w, r_p, r_pos, r_w2, r_w3, i0, 0x7FFFFFFF, false);
            if (ev.j < 0) {
                s.lit += b_cover - i0;
                s.i = b_cover;
                continue;
            }
            const int qi = ev.qi;
            uint32_t w2 = ev.w2, w3 = ev.w3;
            int a_pos = ev.a_pos, a_len = (int)(w2 & 63u);
            if (w2 & HIT_DEFER) {
                uint32_t lo, hi, nm;
                ext3(pv.q, qi, lo, hi, nm);
                anchor_lookup(w, pv, P, qi, kmer_key(lo, hi, P.mal), a_pos, a_len);
                w2 = 0;
            }
            if (a_len > 0 && a_len >= P.msl) {
                matched = true;
                m_qi = qi; m_lit = s.lit + (qi - i0); m_pred = -1; m_pos = a_pos; m_len = a_len; m_w2 = w2; m_w3 = w3; m_j = ev.j;
            } else {
                s.lit += qi + 1 - i0;
                s.i = qi + 1;
            }
        } else {
        // ---- prediction mode (:533-624): pred >= 0 and lit <= mqd
        const int lit0 = s.lit;
        const int P0 = s.pred - s.lit;
        int n = P.mqd + 1 - lit0;
        if (n > NC) n = NC;
        if (n > b_cover - i0) n = b_cover - i0;
        // seeds of this step already known?  (record of the anchor match that ended at i0)
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
        lz_pred_steps++;
#endif
        const bool use_spec = PT::kSpecSeeds && sK1 > 0;
        if (use_spec && n > sK1) n = sK1;
        // the first anchor record inside the chunk that counts here (position 0 == "none", :604)
        const Event eva = next_event(w, r_p, r_pos, r_w2, r_w3, i0, i0 + n, true);
        const int ja = eva.j;
        const int ka = ja >= 0 ? eva.qi - i0 : n;
        const int ns = ja >= 0 ? ka + 1 : n;          // positions examined by this step

        // seeds: runs of >= msl equal bases along the diagonals [d_lo, d_hi) of the window
        Planes64 Q{};                    // loaded only when a seed has to be looked at
        const int d_lo = -(ns - 1), d_hi = lit0 + P.mrd;
        uint64_t cand;
        if (use_spec) cand = sks < ns ? 1ull << sks : 0ull;   // only the first seed matters
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
        if (use_spec) {   // TEST ONLY (tests/emul): the record must agree with the diagonal scan
            uint64_t chk = 0;
            const Planes64 Qc = ext64(pv.q, i0);
            for (int d = d_lo; d < d_hi; ++d) chk |= run_starts(diag_equal64(pv, Qc, i0, P0, d), P.msl);
            chk &= low_mask64(ns);
            const uint64_t first = chk ? chk & (0 - chk) : 0;
            if (first != cand) { lz_spec_mismatches++; cand = first; }
            lz_spec_used++;
        }
#endif
        if (!use_spec || (cand && spec_x < 0)) Q = ext64(pv.q, i0);
        if (!use_spec) {
            cand = diag_seed_starts(w, pv, P, Q, i0, P0, d_lo, d_hi) & low_mask64(ns);
        }

        for (;;) {
            const int ks = cand ? ctz64(cand) : 64;
            const int k = ks < ka ? ks : ka;
            if (k >= ns) break;
            const int qi = i0 + k;
            int best_pos = 0, best_len = 0;
            if (ks == k) {
                if (use_spec && spec_x >= 0) {
                    // the record names the only seed of this position: measure it (equal_len, :192-207)
                    best_pos = P0 + spec_x;
                    const uint32_t m0 = mism32(pv.q, pv.Dq, qi, pv.r, pv.Dr, best_pos);
                    best_len = m0 ? ctz32(m0) : equal_len_coop(w, pv, qi, best_pos, 32);
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
                    {   // TEST ONLY (tests/emul): must agree with the diagonal scan
                        int cp = 0, cl = 0;
                        const Planes64 Qc = ext64(pv.q, i0);
                        best_seed(w, pv, P, Qc, i0, P0, lit0, k, d_lo, d_hi, cp, cl);
                        if (cp != best_pos || cl != best_len) { lz_spec_mismatches++; best_pos = cp; best_len = cl; }
                    }
#endif
                } else
                    best_seed(w, pv, P, Q, i0, P0, lit0, k, d_lo, d_hi, best_pos, best_len);
                cand &= cand - 1ull;
            }
            uint32_t w2 = 0, w3 = 0;
            if (ja >= 0 && k == ka) {
                w2 = eva.w2;
                w3 = eva.w3;
                int a_pos = eva.a_pos, a_len = (int)(w2 & 63u);
                if (w2 & HIT_DEFER) {
                    uint32_t lo, hi, nm;
                    ext3(pv.q, qi, lo, hi, nm);
                    anchor_lookup(w, pv, P, qi, kmer_key(lo, hi, P.mal), a_pos, a_len);
                    w2 = 0;
                }
                bool take_anchor = false;
                if (a_pos) {                              // :604
                    if (!best_pos) take_anchor = true;    // :606-610
                    else {
                        take_anchor = anchor_beats_seed(a_len, (uint32_t)(2 * (pv.Dr + 1 - a_len)), best_len,
                                                        (uint32_t)(lit0 + k + P.mrd + 1 - best_len));
                    }
                }
                if (take_anchor) { best_pos = a_pos; best_len = a_len; }
                else w2 = 0;
            }
            if (best_len >= P.msl) {
                matched = true;
                m_qi = qi; m_lit = lit0 + k; m_pred = P0 + lit0 + k; m_pos = best_pos; m_len = best_len; m_w2 = w2; m_w3 = w3; m_j = ja;
                break;
            }
            if (ja >= 0 && k == ka) break;       // the step ends at the anchor's position
        }
        if (!matched) {
            // ns literal steps (:699-707)
            s.i = i0 + ns;
            s.lit = lit0 + ns;
            s.pred = s.lit > P.mqd ? -1 : s.pred + ns;
            spec_K1 = 0;
        }
        }
        if (matched) {
            take_match(w, pv, P, s, rt, m_qi, m_lit, m_pred, m_pos, m_len, m_w2, m_w3);
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
            lz_matches++;
#endif
            spec_K1 = 0;
            chain_nx = 32;
            if (PT::kSpecSeeds && (m_w2 & HIT_FWD)) {    // s.i / s.pred came from the record
                spec_K1 = (int)((m_w3 >> 18) & 63u);
                spec_ks = (int)((m_w3 >> 24) & 127u);
                spec_x = (m_w3 >> 31) ? -1 : (int)(m_w2 >> HIT_X_SHIFT);
                if (spec_K1 > 0 && m_j >= 0) chain_nx = w.get(r_nx, m_j);
            }
        }
    }

    // tail (:710-713): with a live prediction the rest is compared against the reference shifted
    // by -msl; what counts are the columns up to the last matching one.
    if (s.pred >= 0) {
        const int n = s.lit + (Dq - s.i);
        FwdStream ts{&pv, s.i - s.lit, s.pred - s.lit - P.msl};
        int last = 0, mt = 0;
        LZ_ROLLED
        for (int base = 0; base < n; base += 1024) {
            typename W::template Lanes<uint32_t> eq;
            typename W::template Lanes<int> hi, c;
            LZ_EACH(w, ln)
            {
                const int p0 = base + 32 * ln, nv = n - p0;
                eq[ln] = nv <= 0 ? 0u : (~ts.word(p0 >> 5) & low_mask32(nv > 32 ? 32 : nv));
                hi[ln] = eq[ln] ? p0 + 32 - clz32(eq[ln]) : 0;
                c[ln] = popc32(eq[ln]);
            }
            const int h = w.max_i32(hi);
            const int cs = w.add_i32(c);
            if (h > 0) { last = h; }
            mt += cs;      // every match lies at or below the overall last matching column
        }
        s.curM += mt;
        s.curX += last - mt;
        if (RS::kActive) rt.range(s.i - s.lit, s.pred - s.lit - P.msl, n, summarize_range(w, ts, n), false);
    }
    if (RS::kActive) rt.end_region(P.reg);      // the last region is never dropped by the parse
    if (s.rs >= 0 && s.curM + s.curX >= P.reg) {
        s.totM += s.curM;
        s.totX += s.curX;
        ++s.totN;
    }
    PairResult r;
    r.m = s.totM; r.x = s.totX; r.n = s.totN;
    return r;
}

// one-kernel form: the warp builds its batches itself
template <class W, class PT, class RS>
LZ_HD PairResult parse_pair(W& w, const PairView& pv, const PT& P, WarpScratch* sc, RS& rt)
{
    const InlineBatches ib{};
    return parse_pair(w, pv, P, sc, rt, ib);
}
// the counting kernel: no region records
template <class W, class PT>
LZ_HD PairResult parse_pair(W& w, const PairView& pv, const PT& P, WarpScratch* sc)
{
    NoRegions none;
    return parse_pair(w, pv, P, sc, none);
}
// the walk over stored records (staged pipeline)
template <class W, class PT>
LZ_HD PairResult walk_pair(W& w, const PairView& pv, const PT& P, const Rec* recs, uint32_t n_recs)
{
    NoRegions none;
    const StoredBatches sb{recs, n_recs};
    return parse_pair(w, pv, P, (WarpScratch*)nullptr, none, sb);
}

}  // namespace lzani
