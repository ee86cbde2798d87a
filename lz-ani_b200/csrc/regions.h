// regions.h -- region records for --out-alignment: the B200 counterpart of CParser::calc_regions /
// get_parsing (reference src/parser.cpp:786-837) with region_t's update rules (src/defs.h:67-153).
//
// The reference keeps the whole factor vector and folds it into regions afterwards.  Here the fold
// runs alongside the parse with O(1) state: the parse reports what the reference would append to
// v_parsing -- a distant match, a close match, a run of literals, the run-length encoding of two
// aligned ranges (compare_ranges, parser.cpp:210-248) or of a two-sided gap
// (compare_ranges_both_ways, :251-374) -- and the tracker applies the same update rules in the same
// order, including the ones that only make sense historically:
//   * before a close match the pending literal count is ADDED to ref_end and seq_end
//     (extend_region) and only then maxed with the match's own end;
//   * the query coordinate of the right-hand columns of a two-sided gap lags by one when the first
//     of them is merged into a preceding literal run (parser.cpp:353-368).
// A region enters the output when the parse did not drop it (prev_region_end - prev_region_start
// >= reg, parser.cpp:449/643) and calc_regions keeps it (seq_end - seq_start >= reg, :800/:826).
//
// Two sinks implement the same calls: NoRegions (everything compiles away: the counting kernel) and
// RegionTracker<Out> (the regions kernel; Out::emit receives finished regions).
#pragma once
#include "layout.h"

namespace lzani {

struct RegionRec { int ref_start, ref_end, seq_start, seq_end, num_matches, num_mismatches; };

// what one run-length-encoded diagonal range looks like to the fold
struct RangeSummary {
    int first;       // first matching column, -1 if none
    int first_run;   // length of the run of matches starting there
    int last_end;    // one past the last matching column
    int matches;     // matching columns in total
};

struct NoRegions {
    static constexpr bool kActive = false;
    LZ_HD void lit(int) {}
    LZ_HD void dist(int, int, int) {}
    LZ_HD void close(int, int, int) {}
    LZ_HD void range(int, int, int, const RangeSummary&, bool) {}
    LZ_HD void drop() {}
    LZ_HD void end_region(int) {}
    LZ_HD void clear_all() {}
};

template <class Out>
struct RegionTracker {
    static constexpr bool kActive = true;
    Out* out;
    int rs, re, qs, qe, m, x, pend;     // region_t of the region being built + buf_n_lit

    LZ_HD void clear() { rs = re = qs = qe = -1; m = x = 0; pend = 0; }
    LZ_HD void clear_all() { clear(); }
    LZ_HD void cover(int q, int r, int len)        // the four update_* calls + update_matches
    {
        if (qs < 0 || q < qs) qs = q;
        if (qe < 0 || q + len > qe) qe = q + len;
        if (rs < 0 || r < rs) rs = r;
        if (re < 0 || r + len > re) re = r + len;
        m += len;
    }
    LZ_HD void lit(int n) { pend += n; }                               // run_literals (:826-829)
    LZ_HD void dist(int q, int r, int len) { clear(); cover(q, r, len); }   // match_distant (:797-809)
    LZ_HD void close(int q, int r, int len)                            // match_close (:810-821)
    {
        re += pend; qe += pend; x += pend; pend = 0;
        cover(q, r, len);
    }
    // compare_ranges over `len` columns from (q0, r0); the first matching run is a distant match
    // when first_distant (the backward extension, parser.cpp:664-668), else a close one
    LZ_HD void range(int q0, int r0, int len, const RangeSummary& u, bool first_distant)
    {
        if (u.first < 0) { pend += len; return; }
        if (first_distant) dist(q0 + u.first, r0 + u.first, u.first_run);
        else { pend += u.first; close(q0 + u.first, r0 + u.first, u.first_run); }
        const int d0 = u.first + u.first_run;          // columns done
        if (u.last_end > d0) {
            // further runs (g_i literals, m_i matches) on the same diagonal: every step is
            //   end = max(end + g_i, diagonal end after the run), so the lead of `end` over the
            // diagonal shrinks by m_i per run and never grows back
            const int more = u.matches - u.first_run;
            int lead_r = re - (r0 + d0), lead_q = qe - (q0 + d0);
            lead_r = lead_r > more ? lead_r - more : 0;
            lead_q = lead_q > more ? lead_q - more : 0;
            re = r0 + u.last_end + lead_r;
            qe = q0 + u.last_end + lead_q;
            x += (u.last_end - d0) - more;
            m += more;
        }
        pend = len - u.last_end;
    }
    LZ_HD void drop() { clear(); }          // the parse removed the region (parser.cpp:643-657)
    // a distant match or the end of the parse closes the region; calc_regions keeps it by length
    LZ_HD void end_region(int reg)
    {
        if (qe - qs >= reg) {
            RegionRec g;
            g.ref_start = rs; g.ref_end = re; g.seq_start = qs; g.seq_end = qe; g.num_matches = m; g.num_mismatches = x;
            out->emit(g);
        }
        clear();
    }
};

}  // namespace lzani
