// lane_walk.h -- the sequential walk of the staged parse, ONE LANE PER PAIR.
//
// parse_core.h's parse_pair() gives a whole warp to one pair: its lanes cooperate on the rare wide
// operations (long matches, diagonal scans) but execute the state machine of parser.cpp:501-708 itself --
// which record comes next, is it plain, region bookkeeping -- 32 times over (ncu r01zg: ~60 % of the walk's
// instructions ran at 32 threads per instruction doing one pair's scalar work).  With the hit records of the
// staged pipeline everything the common event needs is already in its 16-byte record, so here every lane
// walks its OWN pair:
//
//   RUN   lane_run_step(): the plain events of an unrelated pair (86 % of all events) -- the next anchor record
//         is complete, lies inside the seed-free part of the window the previous record opened (or the
//         prediction expires first), is far from the prediction, and its backward extension is usable:
//         a distant match taken entirely from the record's fields.  A few dozen scalar instructions and one
//         16-byte load per event.
//   SLOW  lane_slow_step(): one general step of the reference loop, scalar: seed matches (best seed of the
//         window, parser.cpp:548-580), close matches with the two-sided gap re-alignment (:251-374), the
//         anchor-vs-seed arbitration (:604-623), deferred anchors (probe chain + long equal_len), extensions
//         that leave their first word (:377-441), the tail (:710-713).
//
//   SCAN  lane_scan_step(): the seeds of a prediction window no record describes (after a seed match, a close
//         match, a deferred anchor): the row-wise scan of k_hit_spec run by the lane itself.
//
// A lane that meets something non-plain PARKS (mode = SLOW or SCAN); the kernel (k_walk_lanes) runs a parked
// step for all lanes of a warp waiting for it once enough of them do, so the expensive code is entered with
// many active lanes instead of once per event, and lanes wanting different expensive things do not serialise.  Everything in this file is plain scalar code over loads of the
// sequence records and the reference's anchor table: it runs unchanged on the host in tests/emul.
//
// Semantics: SURVEY Appendix A (verified against the reference), the same quirks as parse_pair() -- position 0
// = "no anchor/seed", right side of the two-sided gap anchored at the match END, tail shifted by -msl, uint32
// wrap of a negative exponent, '>' / '>=' tie rules.
#pragma once
#include "parse_core.h"
#if defined(LZ_SPEC_CHECK) && !defined(__CUDACC__)
#include <stdio.h>      // TEST ONLY (tests/emul): self-check diagnostics
#endif

namespace lzani {

constexpr int LANE_RUN = 0, LANE_SLOW = 1, LANE_SCAN = 2, LANE_DONE = 3;
constexpr int LANE_NONE = 0x7FFFFFFF;

struct LaneCtx {
    PairView pv;
    const Rec* recs;             // this pair's hit records, position order
    uint32_t n, c;               // their number; first record not entirely behind s.i
    int n_loop;                  // the loop of parser.cpp:501 runs while i < n_loop = Dq - msl
    ParseState s;
    int spec_K1, spec_ks, spec_x;    // speculative seed record of the match just taken (spec_K1 > 0 <=> valid for the window at s.i)
    int chain;                   // verified successor link of the record just taken (0: none): the next event is the
                                 // record that many places on, plain, whatever the region bookkeeping says (verified_link)
    int mode;
};

LZ_HD void lane_begin(LaneCtx& L, const PairView& pv, const Rec* recs, uint32_t n_recs, int msl)
{
    L.pv = pv;
    L.recs = recs; L.n = n_recs; L.c = 0;
    L.n_loop = pv.Dq - msl;
    L.s.i = 0; L.s.pred = -1; L.s.lit = 0;
    L.s.rs = -1; L.s.re = 0; L.s.kept_end = 0;
    L.s.curM = L.s.curX = 0;
    L.s.totM = L.s.totX = L.s.totN = 0;
    L.spec_K1 = 0; L.spec_ks = 64; L.spec_x = -1;
    L.chain = 0;
    L.mode = LANE_RUN;
}

// ---- scalar forms of the warp-cooperative routines of parse_core.h

// equal_len (parser.cpp:192-207), the first `from` bases known equal
LZ_HD int equal_len_lane(const PairView& pv, int qpos, int rpos, int from)
{
    int base = from;
    for (;;) {
        const uint32_t m = mism32(pv.q, pv.Dq, qpos + base, pv.r, pv.Dr, rpos + base);   // past either end: all mismatches
        if (m) return base + ctz32(m);
        base += 32;
    }
}

// try_extend_forward / try_extend_backward (parser.cpp:377-441) on a mismatch stream of `limit` positions,
// word by word; same rule as approx_extend()
template <class PT>
LZ_HD void approx_extend_lane(const AnyStream& st, int limit, const PT& P, int& out_len, int& out_matches)
{
    out_len = 0;
    out_matches = 0;
    uint32_t prev = 0;            // positions before the stream count as matches
    int matches_before = 0;
    const uint64_t wmask = low_mask64(P.aw);
    LZ_ROLLED
    for (int p0 = 0; p0 < limit; p0 += 32) {
        const uint32_t cur = st.word(p0 >> 5);
        const int nv = limit - p0 > 32 ? 32 : limit - p0;
        const uint64_t x = ((uint64_t)cur << 32) | prev;
        int brk = -1;
        uint32_t t = cur & low_mask32(nv);      // the window can only overflow at a mismatch
        if (popc32(t) + popc32(prev >> (33 - P.aw > 31 ? 31 : 33 - P.aw)) > P.am) {
            while (t) {
                const int j = ctz32(t);
                if (popc64((x >> (33 + j - P.aw)) & wmask) > P.am) { brk = j; break; }
                t &= t - 1u;
            }
        }
        const int end = brk >= 0 ? brk + 1 : nv;
        const uint64_t y = ~x;
        uint64_t r = y;
        for (int s = 1; s < P.ar; ++s) r &= y << s;
        const uint32_t ok = (uint32_t)(r >> 32) & low_mask32(end);
        if (ok) {
            const int l = 32 - clz32(ok);
            out_len = p0 + l;
            out_matches = matches_before + popc32(~cur & low_mask32(l));
        }
        if (brk >= 0) return;
        matches_before += popc32(~cur & low_mask32(nv));
        prev = cur;
    }
}

// compare_ranges_both_ways (parser.cpp:251-374), counts only: see both_ways_best().  The split is moved one
// column at a time (len <= mqd + 1).
LZ_HD int both_ways_best_lane(const PairView& pv, int qe, int r_left, int r_right, int len)
{
    if (len <= 0) return 0;
    const int q0 = qe - len;
    const int span = r_right < r_left ? len : (r_right - r_left < len ? r_right - r_left : len);
    const int rmax = span < r_right ? span : r_right;
    const FwdStream Ls{&pv, q0, r_left};
    const BwdStream Rs{&pv, qe, r_right};
    if (span <= 64) {
        // the usual gap (at most mqd literals before a close seed): both sides in registers
        uint64_t e_l = (uint64_t)(uint32_t)~Ls.word(0), e_r = (uint64_t)(uint32_t)~Rs.word(0);
        if (span > 32) e_l |= (uint64_t)(uint32_t)~Ls.word(1) << 32;
        if (rmax > 32) e_r |= (uint64_t)(uint32_t)~Rs.word(1) << 32;
        e_l &= low_mask64(span);
        e_r &= low_mask64(rmax);
        int best = 0, left = 0, right = popc64(e_r);       // right = equal columns among the first min(u, rmax) of R, u = span - s
        LZ_ROLLED
        for (int sp = 0;; ++sp) {
            const int u = span - sp;
            const int v = left + (u > rmax ? 0 : right);
            if (v > best) best = v;
            if (sp == span) break;
            left += (int)((e_l >> sp) & 1ull);
            if (u - 1 < rmax) right -= (int)((e_r >> (u - 1)) & 1ull);
        }
        return best;
    }
    int rc = 0;                                  // equal columns among the first min(ru, rmax) of the right side
    LZ_ROLLED
    for (int w0 = 0; 32 * w0 < rmax; ++w0) {
        const int nr = rmax - 32 * w0;
        rc += popc32(~Rs.word(w0) & low_mask32(nr > 32 ? 32 : nr));
    }
    int best = 0, left = 0, ru = span;
    uint32_t lw = 0, rw = 0;
    int lwi = -1, rwi = -1;
    LZ_ROLLED
    for (int sp = 0; sp <= span; ++sp) {
        const int v = left + (ru > rmax ? 0 : rc);       // the reference zero-pads the right side beyond rmax
        if (v > best) best = v;
        if (sp < span) {
            if ((sp >> 5) != lwi) { lwi = sp >> 5; lw = ~Ls.word(lwi); }
            left += (int)((lw >> (sp & 31)) & 1u);
            const int b = ru - 1;
            if (b < rmax) {
                if ((b >> 5) != rwi) { rwi = b >> 5; rw = ~Rs.word(rwi); }
                rc -= (int)((rw >> (b & 31)) & 1u);
            }
            --ru;
        }
    }
    return best;
}

// the probe chain of ht_long (parser.cpp:516-530 / :587-601): longest occurrence, ties -> smallest position
template <class PT>
LZ_HD void anchor_lookup_lane(const PairView& pv, const PT& P, int qpos, int& a_pos, int& a_len)
{
    a_pos = 0;
    a_len = 0;
    uint32_t lo, hi, nm;
    ext3(pv.q, qpos, lo, hi, nm);
    if (nm & low_mask32(P.mal)) return;
    const uint64_t key = kmer_key(lo, hi, P.mal);
    const uint32_t tmask = (1u << pv.tbits) - 1u;
    const uint32_t tag = table_tag(key, P.mal);
    uint32_t h = table_home(key, pv.tbits);
    for (;;) {
        const uint64_t slot = ld_u64(pv.table + (h & tmask));
        ++h;
        if (slot == LZ_EMPTY) return;
        if ((uint32_t)(slot >> 32) != tag) continue;
        const int rp = (int)(uint32_t)slot;
        const int len = equal_len_lane(pv, qpos, rp, 0);
        if (len >= P.mal && (len > a_len || (len == a_len && rp < a_pos))) { a_len = len; a_pos = rp; }
    }
}

// ---- seeds of a prediction window (parser.cpp:548-580).  Window: query positions i0 + k, k in [0, n); the
// reference expects the match to continue at P0 + lit0 + k; a seed at k is a reference position P0 + x,
// x in [0, lit0 + k + mrd), whose msl-mer equals the query's.

// msl == 7: every seed of row k, as a mask over x in [0, 96) (caller masks to the valid offsets)
LZ_HD V3 seeds_at_row7(const PairView& pv, int qi, int P0)
{
    V3 rl, rh, rn;
    {
        const int wd = P0 >> 5, sh = P0 & 31;
        const Rec a = ld_rec(pv.r + wd), b = ld_rec(pv.r + wd + 1), c = ld_rec(pv.r + wd + 2), d = ld_rec(pv.r + wd + 3);
        rl.a = funnel_r(a.x, b.x, sh); rl.b = funnel_r(b.x, c.x, sh); rl.c = funnel_r(c.x, d.x, sh);
        rh.a = funnel_r(a.y, b.y, sh); rh.b = funnel_r(b.y, c.y, sh); rh.c = funnel_r(c.y, d.y, sh);
        rn.a = funnel_r(a.z, b.z, sh); rn.b = funnel_r(b.z, c.z, sh); rn.c = funnel_r(c.z, d.z, sh);
    }
    uint32_t lo, hi, nm;
    ext3(pv.q, qi, lo, hi, nm);
    if (nm & 0x7Fu) return V3{0, 0, 0};
    V3 s{0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
    LZ_ROLLED
    for (int j = 0; j < 7; ++j) {
        const uint32_t nl = ((lo >> j) & 1u) - 1u, nh = ((hi >> j) & 1u) - 1u;
        V3 eq;
        eq.a = (rl.a ^ nl) & (rh.a ^ nh) & ~rn.a;
        eq.b = (rl.b ^ nl) & (rh.b ^ nh) & ~rn.b;
        eq.c = (rl.c ^ nl) & (rh.c ^ nh) & ~rn.c;
        if (j) {
            s.a &= funnel_r(eq.a, eq.b, j);
            s.b &= funnel_r(eq.b, eq.c, j);
            s.c &= eq.c >> j;
        } else { s.a &= eq.a; s.b &= eq.b; s.c &= eq.c; }
    }
    return s;
}

// best seed among the candidates of one row (longest; ties -> closest to the prediction, then smallest position)
LZ_HD void best_seed_of_mask(const PairView& pv, V3 cand, int qi, int P0, int off0, int& best_pos, int& best_len)
{
    int bl = 0, bx = 0;
    uint32_t bk = 0xFFFFFFFFu;
    LZ_ROLLED
    for (int wi = 0; wi < 3; ++wi) {
        uint32_t m = wi == 0 ? cand.a : (wi == 1 ? cand.b : cand.c);
        while (m) {
            const int x = 32 * wi + ctz32(m);
            m &= m - 1u;
            const int len = equal_len_lane(pv, qi, P0 + x, 0);
            const int off = x - off0;
            const uint32_t dk = off < 0 ? (uint32_t)(-off) * 2u : (uint32_t)off * 2u + 1u;
            if (len > bl || (len == bl && dk < bk)) { bl = len; bk = dk; bx = x; }
        }
    }
    best_pos = bl ? P0 + bx : 0;
    best_len = bl;
}

// any parameters / N inside the window: the diagonals of the window one by one (scalar diag_seed_starts / best_seed)
template <class PT>
LZ_HD uint64_t diag_seed_starts_lane(const PairView& pv, const PT& P, const Planes64& Q, int i0, int P0, int d_lo, int d_hi)
{
    uint64_t sm = 0;
    LZ_ROLLED
    for (int d = d_lo; d < d_hi; ++d) sm |= run_starts(diag_equal64(pv, Q, i0, P0, d), P.msl);
    return sm;
}
template <class PT>
LZ_HD void best_seed_diag_lane(const PairView& pv, const PT& P, const Planes64& Q, int i0, int P0, int lit0, int k, int d_lo,
                               int d_hi, int& best_pos, int& best_len)
{
    int sl = 0;
    uint32_t skey = 0xFFFFFFFFu;
    const int qi = i0 + k;
    LZ_ROLLED
    for (int d = d_lo < -k ? -k : d_lo; d < d_hi; ++d) {
        const uint64_t E = diag_equal64(pv, Q, i0, P0, d);
        if (!((run_starts(E, P.msl) >> k) & 1ull)) continue;
        const uint64_t z = ~(E >> k);
        const int rl = z ? ctz64(z) : 64;
        const int len = rl >= 64 - k ? equal_len_lane(pv, qi, P0 + d + k, 64 - k) : rl;
        const int off = d - lit0;
        const uint32_t dk = off < 0 ? (uint32_t)(-off) * 2u : (uint32_t)off * 2u + 1u;
        if (len > sl || (len == sl && dk < skey)) { sl = len; skey = dk; }
    }
    best_pos = 0;
    best_len = 0;
    if (sl > 0) {
        const int off = (skey & 1u) ? (int)(skey >> 1) : -(int)(skey >> 1);
        best_len = sl;
        best_pos = P0 + lit0 + k + off;
    }
}

#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
static long lz_lane_why[16] = {0};      // TEST ONLY: why lanes parked / what the slow steps did
// TEST ONLY (tests/emul): what the diagonal scan says about the first `ns` positions of the window at (i0, P0, lit0);
// every use of a speculative record or of the row-wise scan is compared with it
template <class PT>
inline int lane_check_first_seed(const PairView& pv, const PT& P, int i0, int P0, int lit0, int ns)
{
    if (ns <= 0) return 64;
    const Planes64 Qc = ext64(pv.q, i0);
    const uint64_t chk = diag_seed_starts_lane(pv, P, Qc, i0, P0, -(ns - 1), lit0 + P.mrd) & low_mask64(ns);
    return chk ? ctz64(chk) : 64;
}
#endif

// ---- the next anchor event at or after query position i0 and below lim (next_event(), one lane): the first
// record that is not behind the walk and either counts itself or has a run of hits reaching i0 ("virtual":
// LIVE | DEFER, evaluated by anchor_lookup_lane when taken).  Advances the cursor over records behind i0.
struct LaneEvent { int qi, a_pos; uint32_t w2, w3, idx, link; };
LZ_HD int rec_pos(const Rec& h) { return (int)(h.x & (uint32_t)(LZ_STAGED_MAX_POS - 1)); }
LZ_HD int rec_run(const Rec& h) { return (int)((h.z >> HIT_RUN_SHIFT) & 7u); }
LZ_HD LaneEvent lane_next_event(LaneCtx& L, int i0, int lim, bool need_pos)
{
    LaneEvent o;
    o.qi = LANE_NONE; o.a_pos = 0; o.w2 = 0; o.w3 = 0; o.idx = L.n; o.link = 0;
    uint32_t c = L.c;
    bool behind_prefix = true;
    int skipped = 0;
    while (c < L.n) {
        const Rec h = ld_rec(L.recs + c);
        const int p = rec_pos(h), run = rec_run(h);
        if (p + run < i0) {
            ++c;
            if (behind_prefix) {
                L.c = c;
                if (++skipped == 8) {
                    // a long match jumped far ahead: gallop over the records behind the walk, then bisect
                    uint32_t step = 8u;
                    while (L.c + step < L.n) {
                        const Rec hn = ld_rec(L.recs + L.c + step);
                        if (rec_pos(hn) + rec_run(hn) >= i0) break;
                        L.c += step + 1u;               // position order: everything up to that record lies behind
                        step *= 2u;
                    }
                    while (step > 1u) {
                        step >>= 1;
                        if (L.c + step < L.n) {
                            const Rec hn = ld_rec(L.recs + L.c + step);
                            if (rec_pos(hn) + rec_run(hn) < i0) L.c += step + 1u;
                        }
                    }
                    c = L.c;
                }
            }
            continue;
        }
        behind_prefix = false;
        if (p >= lim) break;                    // position order: nothing else lies below lim either
        const bool counts = (h.z & HIT_LIVE) && (!need_pos || (h.z & HIT_DEFER) || h.y != 0u);
        if (p >= i0 && counts) {
            o.qi = p; o.a_pos = (int)h.y; o.w2 = h.z; o.w3 = h.w; o.idx = c; o.link = h.x >> HIT_ENTRY_RUN_SHIFT;
            break;
        }
        if (run > 0) {                          // the walk stands inside the run, or the run's first hit does not count
            const int e = p + 1 > i0 ? p + 1 : i0;
            if (e < lim) { o.qi = e; o.w2 = HIT_LIVE | HIT_DEFER; o.idx = c; }
            break;
        }
        ++c;
    }
    return o;
}

// ---- one match (parser.cpp:626-698), counts only: take_match() for one lane
template <class PT>
LZ_HD void lane_take_match(LaneCtx& L, const PT& P, int qi, int lit_k, int pred_k, int best_pos, int best_len, uint32_t w2, uint32_t w3)
{
    ParseState& s = L.s;
    const PairView& pv = L.pv;
    int dist = best_pos - pred_k;
    if (dist < 0) dist = -dist;
    if (pred_k >= 0 && dist <= P.mrd) {
        const int g = both_ways_best_lane(pv, qi, pred_k - lit_k, best_pos + best_len, lit_k);
        s.curM += g + best_len;
        s.curX += lit_k - g;
    } else {
        if (s.rs >= 0 && s.re - s.rs >= P.reg) {
            s.totM += s.curM;
            s.totX += s.curX;
            ++s.totN;
            s.kept_end = s.re;
        }
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
                const AnyStream bs{&pv, qi, best_pos, true};
                approx_extend_lane(bs, lim, P, back, bm);
            }
        }
        s.curM = bm + best_len;
        s.curX = back - bm;
        s.rs = qi - back;
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
        const AnyStream fs{&pv, qn, pn, false};
        approx_extend_lane(fs, lim, P, fwd, fm);
    }
    s.curM += fm;
    s.curX += fwd - fm;
    s.i = qn + fwd;
    s.pred = pn + fwd;
    s.lit = 0;
    s.re = s.i;
}

// the end of the parse: tail (parser.cpp:710-713, reference side shifted by -msl) and the last region
template <class PT>
LZ_HD void lane_finish(LaneCtx& L, const PT& P)
{
    ParseState& s = L.s;
    const PairView& pv = L.pv;
    if (s.pred >= 0) {
        const int n = s.lit + (pv.Dq - s.i);
        const FwdStream ts{&pv, s.i - s.lit, s.pred - s.lit - P.msl};
        int last = 0, mt = 0;
        LZ_ROLLED
        for (int base = 0; base < n; base += 32) {
            const int nv = n - base;
            const uint32_t eq = ~ts.word(base >> 5) & low_mask32(nv > 32 ? 32 : nv);
            if (eq) last = base + 32 - clz32(eq);
            mt += popc32(eq);
        }
        s.curM += mt;
        s.curX += last - mt;
    }
    if (s.rs >= 0 && s.curM + s.curX >= P.reg) {
        s.totM += s.curM;
        s.totX += s.curX;
        ++s.totN;
    }
    L.mode = LANE_DONE;
}

LZ_HD void lane_set_spec(LaneCtx& L, uint32_t w2, uint32_t w3)
{
    L.spec_K1 = (int)((w3 >> 18) & 63u);
    L.spec_ks = (int)((w3 >> 24) & 127u);
    L.spec_x = (w3 >> 31) ? -1 : (int)(w2 >> HIT_X_SHIFT);
}

// ---- RUN: one plain event, or park.  (The fast path of parse_pair() without the batch machinery.)
// Window info (spec_K1 > 0): the rows [0, spec_K1) of the prediction window at (s.i, s.lit) are known -- the
// first one with a seed is spec_ks (64: none) -- from the record of the match just taken (k_hit_spec, lit 0) or
// from this lane's own scan (lane_scan_step).
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
#define LZ_LANE_WHY(n) lz_lane_why[n]++
#else
#define LZ_LANE_WHY(n)
#endif
// the records [base, base + 4) of the pair, loaded together ahead of the steps that will want them (the kernel
// issues the four loads at once: one memory latency per round instead of one per step)
struct RecAhead {
    Rec r[4];
    uint32_t base;
};
LZ_HD Rec rec_ahead_get(const RecAhead& a, uint32_t k)      // k < 4; selects, not an indexed (local-memory) access
{
    Rec h = a.r[0];
    if (k == 1u) h = a.r[1];
    if (k == 2u) h = a.r[2];
    if (k == 3u) h = a.r[3];
    return h;
}

template <class PT>
LZ_HD void lane_run_step(LaneCtx& L, const PT& P, const RecAhead& ahead)
{
    ParseState& s = L.s;
    if (L.chain) {
        // ---- chained event (62 % of the events of an unrelated pair): the record just taken carries a verified link,
        // so the next event is that record, a distant match taken from its fields -- nothing to decide
        const uint32_t idx = L.c - 1u + (uint32_t)L.chain;
        const uint32_t ka = idx - ahead.base;
        const Rec h = ka < 4u ? rec_ahead_get(ahead, ka) : ld_rec(L.recs + idx);
        const int qi = rec_pos(h), a_pos = (int)h.y;
        const uint32_t w2 = h.z, w3 = h.w;
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
        {   // TEST ONLY (tests/emul): the general rules must find the same event, plain, with a usable backward record
            LaneCtx G = L;
            const LaneEvent e2 = lane_next_event(G, s.i, LANE_NONE, false);
            bool ok = e2.idx == idx && e2.qi == qi && (w2 & (HIT_LIVE | HIT_DEFER | HIT_FWD | HIT_BWD)) == (HIT_LIVE | HIT_FWD | HIT_BWD) &&
                      s.pred >= 0 && L.spec_K1 > 0;
            const int kq = qi - s.i;
            if (kq > P.mqd) ok = ok && L.spec_K1 > P.mqd && L.spec_ks >= 64;
            else {
                int dist = a_pos - (s.pred + kq);
                if (dist < 0) dist = -dist;
                ok = ok && kq < L.spec_K1 && L.spec_ks > kq && a_pos != 0 && dist > P.mrd;
            }
            const bool keep2 = s.rs >= 0 && s.re - s.rs >= P.reg;
            const int run2 = qi - (keep2 ? s.re : s.kept_end);
            int lim2 = run2 < qi ? run2 : qi;
            if (a_pos < lim2) lim2 = a_pos;
            ok = ok && run2 > 0 && lim2 >= (int)((w3 >> 12) & 63u);
            if (!ok) { lz_spec_mismatches++; fprintf(stderr, "DBG chain: i %d qi %d idx %u/%u\n", s.i, qi, idx, e2.idx); }
            lz_chain++; lz_fast++;
        }
#endif
        if (s.re - s.rs >= P.reg) { s.totM += s.curM; s.totX += s.curX; ++s.totN; s.kept_end = s.re; }
        const int len = (int)(w2 & 63u), fwd = (int)((w2 >> 6) & 63u), fm = (int)((w2 >> 12) & 63u);
        const int back = (int)(w3 & 63u), bm = (int)((w3 >> 6) & 63u);
        s.curM = bm + len + fm;
        s.curX = back - bm + fwd - fm;
        s.rs = qi - back;
        s.i = qi + len + fwd;
        s.pred = a_pos + len + fwd;
        s.re = s.i;
        lane_set_spec(L, w2, w3);
        L.chain = (int)(h.x >> HIT_ENTRY_RUN_SHIFT);
        L.c = idx + 1u;
        return;
    }
    const int i0 = s.i;
    if (i0 >= L.n_loop) { LZ_LANE_WHY(0); L.mode = LANE_SLOW; return; }                   // the tail
    if (s.pred >= 0 && L.spec_K1 <= 0) { LZ_LANE_WHY(1); L.mode = LANE_SCAN; return; }     // a window nobody has scanned
    // the next anchor event: nearly always the record under the cursor itself (one load, no search)
    LaneEvent ev;
    bool direct = false;
    if (L.c < L.n) {
        const uint32_t ka = L.c - ahead.base;
        const Rec h = ka < 4u ? rec_ahead_get(ahead, ka) : ld_rec(L.recs + L.c);
        const int p = rec_pos(h);
        if (p >= i0 && (h.z & HIT_LIVE)) {
            ev.qi = p; ev.a_pos = (int)h.y; ev.w2 = h.z; ev.w3 = h.w; ev.idx = L.c; ev.link = h.x >> HIT_ENTRY_RUN_SHIFT;
            direct = true;
        }
    }
    if (!direct) ev = lane_next_event(L, i0, LANE_NONE, false);
    const int qi = ev.qi;
    const uint32_t w2 = ev.w2, w3 = ev.w3;
    const int a_pos = ev.a_pos;
    const int len = (int)(w2 & 63u);
    bool plain = (w2 & (HIT_LIVE | HIT_DEFER | HIT_FWD)) == (HIT_LIVE | HIT_FWD) && len >= P.msl;     // a complete record
    if (s.pred < 0) {
        if (qi == LANE_NONE) {
            // no anchor left: literals to the end of the loop
            s.lit += L.n_loop - i0;
            s.i = L.n_loop;
            L.mode = LANE_SLOW;
            return;
        }
        if (!plain) LZ_LANE_WHY(2);
    } else {
        const int n_rem = P.mqd + 1 - s.lit;          // positions the prediction lives for (:699-707)
        const int K1 = L.spec_K1, ks = L.spec_ks;
        const int ka = qi == LANE_NONE ? LANE_NONE : qi - i0;
        if (ks < K1 && ks <= ka) { LZ_LANE_WHY(3); L.mode = LANE_SLOW; return; }          // a seed comes first (or ties with the anchor)
        if (ka < K1 && ka < n_rem) {
            // the anchor lies inside the rows known to hold no seed
            int dist = a_pos - (s.pred + ka);
            if (dist < 0) dist = -dist;
            plain = plain && a_pos != 0 && dist > P.mrd;
            if (!plain) LZ_LANE_WHY(4);
        } else {
            // neither a seed nor an anchor in the known rows
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
            if (lane_check_first_seed(L.pv, P, i0, s.pred - s.lit, s.lit, K1 < n_rem ? K1 : n_rem) != 64) { lz_spec_mismatches++; fprintf(stderr, "DBG run literal: i0 %d pred %d lit %d K1 %d ks %d\n", i0, s.pred, s.lit, K1, ks); }
            lz_spec_used++;
#endif
            if (K1 >= n_rem) {
                // literal steps until the prediction is dropped (:699-707); the same event, still ahead, is met again
                // without a prediction
                s.i = i0 + n_rem;
                s.lit = P.mqd + 1;
                s.pred = -1;
            } else {
                // the known rows end first: K1 literal steps, then the window needs a scan
                s.i = i0 + K1;
                s.lit += K1;
                s.pred += K1;
            }
            L.spec_K1 = 0;
            return;
        }
    }
    if (plain) {
        const bool keep = s.rs >= 0 && s.re - s.rs >= P.reg;
        const int kept_end = keep ? s.re : s.kept_end;
        const int run = qi - kept_end;
        int lim = run < qi ? run : qi;
        if (a_pos < lim) lim = a_pos;
        if (run <= 0 || ((w2 & HIT_BWD) && lim >= (int)((w3 >> 12) & 63u))) {
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
            if (s.pred >= 0) {      // no seed at or before the anchor's position
                if (lane_check_first_seed(L.pv, P, i0, s.pred - s.lit, s.lit, qi - i0 + 1) != 64) { lz_spec_mismatches++; fprintf(stderr, "DBG run plain: i0 %d pred %d lit %d qi %d K1 %d ks %d\n", i0, s.pred, s.lit, qi, L.spec_K1, L.spec_ks); }
                lz_spec_used++;
            }
            lz_fast++;
#endif
            if (keep) { s.totM += s.curM; s.totX += s.curX; ++s.totN; s.kept_end = kept_end; }
            const int back = run > 0 ? (int)(w3 & 63u) : 0, bm = run > 0 ? (int)((w3 >> 6) & 63u) : 0;
            const int fwd = (int)((w2 >> 6) & 63u), fm = (int)((w2 >> 12) & 63u);
            s.curM = bm + len + fm;
            s.curX = back - bm + fwd - fm;
            s.rs = qi - back;
            s.i = qi + len + fwd;
            s.pred = a_pos + len + fwd;
            s.lit = 0;
            s.re = s.i;
            lane_set_spec(L, w2, w3);
            // the record is behind the walk now -- unless the match is shorter than the record's run of hits, which takes
            // mal <= 7 (a match is at least mal long, a run at most 8): then the walk stands inside that run
            L.c = (P.mal <= 7 && qi + (int)((w2 >> HIT_RUN_SHIFT) & 7u) >= s.i) ? ev.idx : ev.idx + 1u;
            // (run > 0: the region just opened is the one the record's fields describe -- what verified_link assumed)
            if (run > 0 && L.spec_K1 > 0) L.chain = (int)ev.link;
            return;
        }
        LZ_LANE_WHY(5);
    }
    L.mode = LANE_SLOW;
}

// ---- SCAN: the seeds of the window at (s.i, s.lit), up to the next anchor event (parser.cpp:548-580 for every
// position at once): row-wise bit-parallel (spec_first_seed) when the window qualifies, else diagonal by diagonal
template <class PT>
LZ_HD void lane_scan_step(LaneCtx& L, const PT& P)
{
    ParseState& s = L.s;
    const PairView& pv = L.pv;
    const int i0 = s.i;
    L.mode = LANE_RUN;
    if (i0 >= L.n_loop || s.pred < 0 || L.spec_K1 > 0) return;
    LZ_LANE_WHY(10);
    const int lit0 = s.lit;
    const int P0 = s.pred - lit0;
    int n = P.mqd + 1 - lit0;
    if (n > 65 - P.msl) n = 65 - P.msl;          // msl-runs starting in the chunk fit 64 bits
    if (n > L.n_loop - i0) n = L.n_loop - i0;
    // rows up to the first anchor record that counts here (position 0 == "none", :604)
    const LaneEvent eva = lane_next_event(L, i0, i0 + n, true);
    const int ns = eva.qi != LANE_NONE ? eva.qi - i0 + 1 : n;
    int ks = 64, sx = -1;
    int r = -1;
    if (PT::kSpecSeeds) r = spec_first_seed_t<true>(pv, P, i0, P0, ns - 1, P.mrd + lit0);
    if (r >= 0) {
        if (r != 64) {                           // k | x << 8 | SPEC_MULTI, k <= ns - 1
            ks = r & 0xFF;
            sx = (r & SPEC_MULTI) ? -1 : ((r >> 8) & 127);
        }
    } else {
        const Planes64 Q = ext64(pv.q, i0);
        const uint64_t cand = diag_seed_starts_lane(pv, P, Q, i0, P0, -(ns - 1), lit0 + P.mrd) & low_mask64(ns);
        if (cand) ks = ctz64(cand);
        sx = -2;                                 // candidates of a row: diagonal by diagonal
    }
    L.spec_K1 = ns;
    L.spec_ks = ks;
    L.spec_x = sx;
}

// ---- SLOW: one general step of the loop (parser.cpp:501-708) or, at its end, the tail.  In prediction mode the
// window info must be there (lane_scan_step).
template <class PT>
LZ_HD void lane_slow_step(LaneCtx& L, const PT& P)
{
    ParseState& s = L.s;
    const PairView& pv = L.pv;
    const int i0 = s.i;
    L.chain = 0;
    if (i0 >= L.n_loop) { lane_finish(L, P); return; }
    if (s.pred >= 0 && L.spec_K1 <= 0) { L.mode = LANE_SCAN; return; }
    L.mode = LANE_RUN;
    LZ_LANE_WHY(9);
    bool matched = false;
    int m_qi = 0, m_lit = 0, m_pred = -1, m_pos = 0, m_len = 0;
    uint32_t m_w2 = 0, m_w3 = 0;
    if (s.pred < 0) {
        // ---- no prediction (:507-532): the next live anchor
        const LaneEvent ev = lane_next_event(L, i0, LANE_NONE, false);
        if (ev.qi == LANE_NONE) {
            s.lit += L.n_loop - i0;
            s.i = L.n_loop;
            lane_finish(L, P);
            return;
        }
        const int qi = ev.qi;
        uint32_t w2 = ev.w2;
        int a_pos = ev.a_pos, a_len = (int)(w2 & 63u);
        if (w2 & HIT_DEFER) {
            anchor_lookup_lane(pv, P, qi, a_pos, a_len);
            w2 = 0;
        }
        if (a_len > 0 && a_len >= P.msl) {
            matched = true;
            m_qi = qi; m_lit = s.lit + (qi - i0); m_pred = -1; m_pos = a_pos; m_len = a_len; m_w2 = w2; m_w3 = ev.w3;
        } else {
            s.lit += qi + 1 - i0;
            s.i = qi + 1;
            return;
        }
    } else {
        // ---- prediction mode (:533-624): pred >= 0 and lit <= mqd
        const int lit0 = s.lit;
        const int P0 = s.pred - lit0;
        int n = P.mqd + 1 - lit0;
        if (n > 65 - P.msl) n = 65 - P.msl;
        if (n > L.n_loop - i0) n = L.n_loop - i0;
        if (n > L.spec_K1) n = L.spec_K1;
        // the first anchor record inside the chunk that counts here (position 0 == "none", :604)
        const LaneEvent eva = lane_next_event(L, i0, i0 + n, true);
        const bool has_a = eva.qi != LANE_NONE;
        const int ka = has_a ? eva.qi - i0 : n;
        const int ns = has_a ? ka + 1 : n;           // positions examined by this step
        const int ks = L.spec_ks < ns ? L.spec_ks : 64;
        const int sx = L.spec_x;
        const int d_lo = -(ns - 1), d_hi = lit0 + P.mrd;
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
        if (lane_check_first_seed(pv, P, i0, P0, lit0, ns) != ks) { lz_spec_mismatches++; fprintf(stderr, "DBG slow first-seed: i0 %d P0 %d lit0 %d ns %d ks %d chk %d K1 %d sx %d Dq %d Dr %d\n", i0, P0, lit0, ns, ks, lane_check_first_seed(pv, P, i0, P0, lit0, ns), L.spec_K1, sx, pv.Dq, pv.Dr); }
        lz_spec_used++;
        lz_pred_steps++;
#endif
        const int k = ks < ka ? ks : ka;
        if (k < ns) {
            const int qi = i0 + k;
            int best_pos = 0, best_len = 0;
            if (ks == k) {
                if (sx >= 0) {
                    // the only seed of this position: measure it (equal_len, :192-207)
                    best_pos = P0 + sx;
                    best_len = equal_len_lane(pv, qi, best_pos, 0);
                } else if (sx == -1) {
                    V3 cand = seeds_at_row7(pv, qi, P0);
                    const int nv = lit0 + k + P.mrd;
                    cand.a &= low_mask32(nv > 32 ? 32 : nv);
                    cand.b &= nv > 32 ? low_mask32(nv - 32 > 32 ? 32 : nv - 32) : 0u;
                    cand.c &= nv > 64 ? low_mask32(nv - 64 > 32 ? 32 : nv - 64) : 0u;
                    best_seed_of_mask(pv, cand, qi, P0, lit0 + k, best_pos, best_len);
                } else {
                    const Planes64 Q = ext64(pv.q, i0);
                    best_seed_diag_lane(pv, P, Q, i0, P0, lit0, k, d_lo, d_hi, best_pos, best_len);
                }
#if defined(LZ_SPEC_CHECK) && !defined(__CUDA_ARCH__)
                if (sx != -2) {      // TEST ONLY: the row-wise candidates against the diagonal scan
                    int cp = 0, cl = 0;
                    const Planes64 Qc = ext64(pv.q, i0);
                    best_seed_diag_lane(pv, P, Qc, i0, P0, lit0, k, d_lo, d_hi, cp, cl);
                    if (cp != best_pos || cl != best_len) { lz_spec_mismatches++; fprintf(stderr, "DBG best seed: i0 %d P0 %d lit0 %d k %d sx %d got %d/%d want %d/%d\n", i0, P0, lit0, k, sx, best_pos, best_len, cp, cl); }
                    lz_why[10]++;
                }
#endif
            }
            uint32_t w2 = 0, w3 = 0;
            if (has_a && k == ka) {
                w2 = eva.w2;
                w3 = eva.w3;
                int a_pos = eva.a_pos, a_len = (int)(w2 & 63u);
                if (w2 & HIT_DEFER) {
                    anchor_lookup_lane(pv, P, qi, a_pos, a_len);
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
                m_qi = qi; m_lit = lit0 + k; m_pred = P0 + lit0 + k; m_pos = best_pos; m_len = best_len; m_w2 = w2; m_w3 = w3;
            }
        }
        if (!matched) {
            // ns literal steps (:699-707)
            s.i = i0 + ns;
            s.lit = lit0 + ns;
            s.pred = s.lit > P.mqd ? -1 : s.pred + ns;
            L.spec_K1 = 0;
        }
    }
    if (matched) {
        lane_take_match(L, P, m_qi, m_lit, m_pred, m_pos, m_len, m_w2, m_w3);
        L.spec_K1 = 0;
        if (PT::kSpecSeeds && (m_w2 & HIT_FWD)) lane_set_spec(L, m_w2, m_w3);    // s.i / s.pred came from the record
    }
    if (s.i >= L.n_loop) lane_finish(L, P);
    else if (s.pred >= 0 && L.spec_K1 <= 0) L.mode = LANE_SCAN;
}

// the whole pair on one lane (host emulation; the kernel interleaves the steps across a warp)
template <class PT>
LZ_HD PairResult lane_walk_pair(const PairView& pv, const PT& P, const Rec* recs, uint32_t n_recs)
{
    LaneCtx L;
    lane_begin(L, pv, recs, n_recs, P.msl);
    RecAhead none;
    none.base = 0x80000000u;     // no record index comes within 4 of it
    for (int k = 0; k < 4; ++k) none.r[k] = Rec{0, 0, 0, 0};
    while (L.mode != LANE_DONE) {
        if (L.mode == LANE_RUN) lane_run_step(L, P, none);
        else if (L.mode == LANE_SCAN) lane_scan_step(L, P);
        else lane_slow_step(L, P);
    }
    PairResult r;
    r.m = L.s.totM; r.x = L.s.totX; r.n = L.s.totN;
    return r;
}

}  // namespace lzani
