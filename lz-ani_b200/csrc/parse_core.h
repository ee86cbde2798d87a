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

// ---- equal_len (parser.cpp:192-207): common prefix of q[qpos..] and r[rpos..], the first
// `from` bases being known equal.  Uniform arguments, uniform result.
template <class W>
LZ_HD int equal_len_coop(W& w, const PairView& pv, int qpos, int rpos, int from)
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
template <class W, class S>
LZ_HD void approx_extend(W& w, const S& st, int limit, const Prm& P, int& out_len, int& out_matches)
{
    out_len = 0;
    out_matches = 0;
    if (limit <= 0) return;
    uint32_t carry = 0;           // stream word preceding the chunk (0: "matches" before the start)
    int matches_before = 0;       // matches in stream positions below the chunk
    const uint64_t wmask = low_mask64(P.aw);
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
                for (int j = 0; j < 32; ++j) {
                    const uint64_t win = (x >> (33 + j - P.aw)) & wmask;
                    const int c = popc32((uint32_t)win) + popc32((uint32_t)(win >> 32));
                    if (c > P.am) { f = p0 + j; break; }
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
LZ_HD int both_ways_best(W& w, const PairView& pv, int qe, int r_left, int r_right, int len)
{
    if (len <= 0) return 0;
    const int q0 = qe - len;
    const int span = r_right < r_left ? len : (r_right - r_left < len ? r_right - r_left : len);
    const int rmax = span < r_right ? span : r_right;
    FwdStream L{&pv, q0, r_left};
    BwdStream R{&pv, qe, r_right};
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

// ---- anchor lookup: the probe chain of ht_long (parser.cpp:516-530 / :587-601).  Returns the
// longest common prefix over all reference positions whose mal-mer equals the query's at
// qpos; ties -> smallest position (the reference's strict '>' over ascending positions).
// (0,0) when there is none.  32 slots of the chain are inspected per step, one per lane.
template <class W>
LZ_HD void anchor_lookup(W& w, const PairView& pv, const Prm& P, int qpos, uint64_t key, int& a_pos, int& a_len)
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

struct ParseState {
    int i, pred, lit;            // query position, ref_pred_pos (<0: none), cur_lit_run_len
    int rs, re, kept_end;        // prev_region_start / prev_region_end / end of last kept region
    int curM, curX;              // matches / counted literals of the region being built
    int totM, totX, totN;        // kept regions so far
};

// One match found at query position qi (parser.cpp:626-698), counts only.
template <class W>
LZ_HD void take_match(W& w, const PairView& pv, const Prm& P, ParseState& s, int qi, int lit_k, int pred_k,
                      int best_pos, int best_len)
{
    int dist = best_pos - pred_k;
    if (dist < 0) dist = -dist;
    if (pred_k >= 0 && dist <= P.mrd) {
        // close match: the literal gap is re-aligned from both ends (:630-635)
        const int g = both_ways_best(w, pv, qi, pred_k - lit_k, best_pos + best_len, lit_k);
        s.curM += g + best_len;
        s.curX += lit_k - g;
    } else {
        // distant match: judge the region built so far (:643, eval_region :446-449) ...
        if (s.rs >= 0 && s.re - s.rs >= P.reg) {
            s.totM += s.curM;
            s.totX += s.curX;
            ++s.totN;
            s.kept_end = s.re;
        }
        // ... everything since the last kept region is now one literal run (:645-655), which the
        // new match may reclaim backwards (:659-672)
        const int run = qi - s.kept_end;
        int back = 0, bm = 0;
        if (run > 0) {
            int lim = run;
            if (qi < lim) lim = qi;
            if (best_pos < lim) lim = best_pos;
            BwdStream bs{&pv, qi, best_pos};
            approx_extend(w, bs, lim, P, back, bm);
        }
        s.curM = bm + best_len;
        s.curX = back - bm;
        s.rs = qi - back;
    }
    int qn = qi + best_len;
    int pn = best_pos + best_len;
    int fwd = 0, fm = 0;
    {
        int lim = pv.Dq - qn;
        if (pv.Dr - pn < lim) lim = pv.Dr - pn;
        FwdStream fs{&pv, qn, pn};
        approx_extend(w, fs, lim, P, fwd, fm);      // :691-695
    }
    s.curM += fm;
    s.curX += fwd - fm;
    s.i = qn + fwd;
    s.pred = pn + fwd;
    s.lit = 0;
    s.re = s.i;
}

// ---- the whole pair.
template <class W>
LZ_HD PairResult parse_pair(W& w, const PairView& pv, const Prm& P)
{
    ParseState s;
    s.i = 0; s.pred = -1; s.lit = 0;
    s.rs = -1; s.re = 0; s.kept_end = 0;
    s.curM = s.curX = 0;
    s.totM = s.totX = s.totN = 0;
    const int Dq = pv.Dq;
    const int n_loop = Dq - P.msl;     // loop runs while i < n_loop (i + msl < Dq, :501)

    while (s.i < n_loop) {
        const int i0 = s.i;
        int n = n_loop - i0;
        if (n > 32) n = 32;
        // positions i0+k with k < nB are in prediction mode (pred >= 0 and lit <= mqd)
        int nB = 0;
        if (s.pred >= 0) {
            nB = P.mqd + 1 - s.lit;
            if (nB > n) nB = n;
        }
        const int lit0 = s.lit;
        const int P0 = s.pred - s.lit;   // meaningful only when nB > 0

        // (1) anchor presence: lane k tests the mal-mer at i0+k
        typename W::template Lanes<uint64_t> key;
        typename W::template Lanes<int> hit;
        LZ_EACH(w, ln)
        {
            hit[ln] = 0;
            key[ln] = 0;
            if (ln < n) {
                uint32_t lo, hi, nm;
                ext3(pv.q, i0 + ln, lo, hi, nm);
                if (!(nm & low_mask32(P.mal))) {
                    key[ln] = kmer_key(lo, hi, P.mal);
                    const uint32_t b = bitmap_index(key[ln], P.mal, pv.bbits);
                    hit[ln] = (ld_u32(pv.bitmap + (b >> 5)) >> (b & 31u)) & 1u;
                }
            }
        }
        uint32_t amask = w.ballot(hit);

        // (2) seed presence along the diagonals of the prediction window
        uint32_t smask = 0;
        const int d_lo = -(nB - 1), d_hi = lit0 + P.mrd;   // diagonals [d_lo, d_hi)
        if (nB > 0) {
            typename W::template Lanes<uint32_t> sm;
            LZ_EACH(w, ln) sm[ln] = 0;
            for (int dbase = d_lo; dbase < d_hi; dbase += 32) {
                LZ_EACH(w, ln)
                {
                    const int d = dbase + ln;
                    if (d < d_hi) sm[ln] |= (uint32_t)run_starts(diag_equal_bits(pv, i0, P0, d), P.msl);
                }
            }
            smask = w.or_u32(sm) & low_mask32(nB);
        }

        // (3) walk the candidate positions in order; the first real match ends the step
        uint32_t cand = amask | smask;
        bool matched = false;
        while (cand) {
            const int k = ctz32(cand);
            cand &= cand - 1u;
            const int qi = i0 + k;
            const bool modeB = k < nB;
            int best_pos = 0, best_len = 0;
            if (modeB && ((smask >> k) & 1u)) {
                // best seed at k: longest; ties -> closest to the prediction, then smallest position
                int sl = 0;
                uint32_t skey = 0xFFFFFFFFu;
                for (int dbase = d_lo; dbase < d_hi; dbase += 32) {
                    typename W::template Lanes<int> len, more;
                    typename W::template Lanes<uint32_t> dk;
                    LZ_EACH(w, ln)
                    {
                        const int d = dbase + ln;
                        len[ln] = 0;
                        more[ln] = 0;
                        dk[ln] = 0xFFFFFFFFu;
                        if (d < d_hi) {
                            const uint64_t E = diag_equal_bits(pv, i0, P0, d);
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
                if (sl > 0) {
                    const int off = (skey & 1u) ? (int)(skey >> 1) : -(int)(skey >> 1);
                    best_len = sl;
                    best_pos = P0 + lit0 + k + off;
                }
            }
            int a_pos = 0, a_len = 0;
            if ((amask >> k) & 1u) anchor_lookup(w, pv, P, qi, w.get(key, k), a_pos, a_len);
            if (!modeB) {
                best_pos = a_pos;                     // :507-532 (a hit at position 0 counts here)
                best_len = a_len;
            } else if (a_pos) {                       // :604 (position 0 == "none")
                if (!best_pos) { best_pos = a_pos; best_len = a_len; }
                else {
                    const double pa = ipow_u32(one_minus_4pow(a_len), (uint32_t)(2 * (pv.Dr + 1 - a_len)));
                    const double pc = ipow_u32(one_minus_4pow(best_len), (uint32_t)(lit0 + k + P.mrd + 1 - best_len));
                    if (pa > pc) { best_pos = a_pos; best_len = a_len; }
                }
            }
            if (best_len >= P.msl) {
                take_match(w, pv, P, s, qi, lit0 + k, modeB ? P0 + lit0 + k : -1, best_pos, best_len);
                matched = true;
                break;
            }
        }
        if (!matched) {
            // n literal steps (:699-707)
            s.i = i0 + n;
            s.lit = lit0 + n;
            if (s.pred >= 0) s.pred = s.lit > P.mqd ? -1 : s.pred + n;
        } else if (s.lit > P.mqd) s.pred = -1;       // (:706 after a match: lit is 0, never true)
    }

    // tail (:710-713): with a live prediction the rest is compared against the reference shifted
    // by -msl; what counts are the columns up to the last matching one.
    if (s.pred >= 0) {
        const int n = s.lit + (Dq - s.i);
        FwdStream ts{&pv, s.i - s.lit, s.pred - s.lit - P.msl};
        int last = 0, mt = 0;
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
    }
    if (s.rs >= 0 && s.curM + s.curX >= P.reg) {
        s.totM += s.curM;
        s.totX += s.curX;
        ++s.totN;
    }
    PairResult r;
    r.m = s.totM; r.x = s.totX; r.n = s.totN;
    return r;
}

}  // namespace lzani
