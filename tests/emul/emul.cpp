// TEST ONLY.  Host emulation of the CUDA path's SOURCE: compiles lz-ani_b200/csrc/{layout.h,
// parse_core.h} with the HostWarp policy of simt.h (32 lanes run as a loop) so that the kernel
// logic can be checked against the oracle on a machine without a GPU (`pytest -m "not gpu"`).
// This is not a CPU fallback: it is built only by tests/emul/Makefile into
// tests/emul/liblzani_emul.so, is never linked into liblzani_b200.so and is ~100x slower than
// the reference.  The library itself fails loudly without a CUDA device.
#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "../../lz-ani_b200/csrc/parse_core.h"
#include "../../lz-ani_b200/csrc/lane_walk.h"
#include "../../lz-ani_b200/csrc/outfilter.h"

using namespace lzani;

namespace {

struct HostSeq {
    std::vector<Rec> q, r;
    int len = 0;
};
struct HostIndex {
    std::vector<uint32_t> bitmap;
    std::vector<uint64_t> table;
    uint32_t bbits = 0, tbits = 0;
};

void build_seq(HostSeq& s, const uint8_t* src, uint32_t len, const Prm& P)
{
    s.len = (int)len;
    int nq = q_form_recs((int)len, P.mrd), nr = r_form_recs((int)len, P.mrd);
    s.q.resize(nq);
    s.r.resize(nr);
    for (int i = 0; i < nq; ++i) s.q[i] = pack_q_record(src, len, (uint32_t)i);
    for (int i = 0; i < nr; ++i) s.r[i] = pack_r_record(src, len, (uint32_t)P.mrd, (uint32_t)i);
}

// same element logic as the K2 kernel (kernels.cu: k_build_index)
void build_index(HostIndex& ix, const HostSeq& s, const Prm& P)
{
    const int Dr = r_form_len(s.len, P.mrd);
    ix.bbits = index_bbits(Dr, P.mal);
    ix.tbits = index_tbits(Dr);
    ix.bitmap.assign(((size_t)1 << ix.bbits) / 32 ? ((size_t)1 << ix.bbits) / 32 : 1, 0u);
    ix.table.assign((size_t)1 << ix.tbits, LZ_EMPTY);
    const uint32_t tmask = (1u << ix.tbits) - 1u;
    for (int pos = 0; pos < Dr; ++pos) {
        uint64_t key;
        if (!ref_kmer_at(s.r.data(), Dr, pos, P.mal, key)) continue;
        const uint32_t b = bitmap_index(key, P.mal, ix.bbits);
        ix.bitmap[b >> 5] |= 1u << (b & 31u);
        uint32_t h = table_home(key, ix.tbits);
        const uint64_t e = ((uint64_t)table_tag(key, P.mal) << 32) | (uint32_t)pos;
        // insertion order is scrambled on purpose: the parse must not depend on chain order
        while (ix.table[h] != LZ_EMPTY) h = (h + 1u) & tmask;
        ix.table[h] = e;
    }
}

// the staged pipeline of kernels.cu for one pair: k_scan_hits / k_expand_hits / k_hit_records element by
// element, then the walk over the stored records (k_walk_pairs)
template <class PT>
std::vector<Rec> staged_records(const PairView& pv, const PT& P, uint32_t& n_recs);

template <class PT>
PairResult staged_pair(HostWarp& w, const PairView& pv, const PT& P, bool lane_walk)
{
    uint32_t n_recs = 0;
    const std::vector<Rec> recs = staged_records(pv, P, n_recs);
    // k_walk_lanes (one lane per pair, lane_walk.h) or k_walk_pairs (one warp per pair).  The launcher uses the lane
    // walk only for the parameter sets that have seed records (PrmDefault's constants, mqd <= SPEC_MAX_K); here every
    // parameter set goes through it when asked to, so that its general code is checked beyond what the device runs
    if (lane_walk) return lane_walk_pair(pv, P, recs.data(), n_recs);
    return walk_pair(w, pv, P, recs.data(), n_recs);
}

// the hit records of one pair as k_scan_* / k_expand_hits / k_hit_probe / k_hit_spec leave them
template <class PT>
std::vector<Rec> staged_records(const PairView& pv, const PT& P, uint32_t& n_recs)
{
    const int n_loop = pv.Dq - P.msl;
    const int n_words = n_loop > 0 ? (n_loop + 31) / 32 : 0;
    std::vector<HitEntry> list;
    std::vector<uint32_t> coarse(coarse_words(pv.bbits));
    for (uint32_t c = 0; c < coarse.size(); ++c) coarse[c] = coarse_fold_word(pv.bitmap, pv.bbits, c);
    for (int wd = 0; wd < n_words; ++wd) {
        const uint32_t m = scan_word_coarse(pv, P, 32 * wd, n_loop, coarse.data());
        if (m != scan_word(pv, P, 32 * wd, n_loop)) ++lz_spec_mismatches;     // two-level scan == exact scan
        uint32_t f = run_first_bits(m);
        while (f) {
            const int b = ctz32(f);
            f &= f - 1u;
            list.push_back(HitEntry{0u, (uint32_t)(32 * wd + b) | ((uint32_t)run_following(m, b) << HIT_ENTRY_RUN_SHIFT)});
        }
    }
    std::vector<Rec> recs(list.size() + 1);
    for (size_t t = 0; t < list.size(); ++t) recs[t] = hit_record_probe(pv, P, list.data(), t, true, [] {});      // k_hit_probe
    // k_hit_spec: scan, then the successor link.  The kernel's threads run in any order: the link must come out
    // the same whether the successor's own scan has been folded in yet (here: ascending = never; checked below
    // against descending = always)
    std::vector<uint32_t> link(list.size(), 0u);
    for (size_t t = 0; t < list.size(); ++t) {
        const SpecJob j = hit_spec_job(P, n_loop, list.data(), t, list.size(), recs[t]);
        if (j.need) hit_spec_finish(recs[t], j, spec_first_seed(pv, P, j.si0, j.sP0, j.K, P.mrd));
        const uint32_t d = recs[t].x >> HIT_ENTRY_RUN_SHIFT;
        link[t] = d ? verified_link(P, recs[t], recs[t + d]) : 0u;          // successor as k_hit_probe left it
    }
    for (size_t t = 0; t < list.size(); ++t) {
        const uint32_t d = recs[t].x >> HIT_ENTRY_RUN_SHIFT;
        const uint32_t after = d ? verified_link(P, recs[t], recs[t + d]) : 0u;   // successor with its scan (and tentative link)
        if (after != link[t]) ++lz_spec_mismatches;
    }
    for (size_t t = 0; t < list.size(); ++t)
        recs[t].x = (recs[t].x & (uint32_t)(LZ_STAGED_MAX_POS - 1)) | (link[t] << HIT_ENTRY_RUN_SHIFT);
    n_recs = (uint32_t)list.size();
    return recs;
}

}  // namespace

static int staged = 0;       // 0: one-kernel form, 1: staged + warp-per-pair walk, 2: staged + lane-per-pair walk
extern "C" void lzani_emul_set_staged(int on) { staged = on; }

extern "C" double lzani_emul_run_rows(const int32_t* p8, uint32_t n_seqs, const uint8_t* const* data,
                                      const uint32_t* len, uint32_t n_rows, const uint32_t* ref_ids,
                                      const uint64_t* row_off, const uint32_t* q_ids, int32_t* out3,
                                      int n_threads)
{
    Prm P{p8[0], p8[1], p8[2], p8[3], p8[4], p8[5], p8[6], p8[7]};
    if (n_threads <= 0) n_threads = 1;
    std::vector<HostSeq> seqs(n_seqs);
    for (uint32_t i = 0; i < n_seqs; ++i) build_seq(seqs[i], data[i], len[i], P);
    std::vector<uint64_t> dense;
    if (!q_ids) {
        dense.assign((size_t)n_rows + 1, 0);
        for (uint32_t j = 0; j < n_rows; ++j) dense[j + 1] = dense[j] + (n_seqs ? n_seqs - 1 : 0);
        row_off = dense.data();
    }
    std::atomic<uint64_t> next{0};
    auto worker = [&]() {
        HostWarp w;
        WarpScratch scratch;
        HostIndex ix;
        for (;;) {
            uint64_t j = next.fetch_add(1);
            if (j >= n_rows) break;
            uint32_t r = ref_ids ? ref_ids[j] : (uint32_t)j;
            uint64_t k = row_off[j];
            if (row_off[j + 1] == k) continue;
            build_index(ix, seqs[r], P);
            auto one = [&](uint32_t q) {
                PairView pv;
                pv.q = seqs[q].q.data(); pv.Dq = seqs[q].len + P.mrd;
                pv.r = seqs[r].r.data(); pv.Dr = r_form_len(seqs[r].len, P.mrd);
                pv.bitmap = ix.bitmap.data(); pv.bbits = ix.bbits;
                pv.table = ix.table.data(); pv.tbits = ix.tbits;
                // same dispatch as launch_parse(): compile-time defaults or run-time parameters
                PairResult res;
                if (staged) {
                    if (is_default_kernel_params(P)) { const PrmDefault D{P.mrd, P.mqd, P.reg}; res = staged_pair(w, pv, D, staged == 2); }
                    else if (is_long_kernel_params(P)) { const PrmLong D{P.mrd, P.mqd, P.reg}; res = staged_pair(w, pv, D, staged == 2); }
                    else res = staged_pair(w, pv, P, staged == 2);
                } else if (is_default_kernel_params(P)) { const PrmDefault D{P.mrd, P.mqd, P.reg}; res = parse_pair(w, pv, D, &scratch); }
                else res = parse_pair(w, pv, P, &scratch);
                out3[3 * k] = res.m; out3[3 * k + 1] = res.x; out3[3 * k + 2] = res.n;
                ++k;
            };
            if (q_ids) for (uint64_t e = row_off[j]; e < row_off[j + 1]; ++e) one(q_ids[e]);
            else for (uint32_t q = 0; q < n_seqs; ++q) if (q != r) one(q);
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < n_threads; ++t) pool.emplace_back(worker);
    worker();
    for (auto& th : pool) th.join();
    return 0.0;
}

// how often the speculative seed records were used / disagreed with the diagonal scan
extern "C" void lzani_emul_counters(long* c)
{
    c[0] = lz_pred_steps; c[1] = lz_batches; c[2] = lz_hits; c[3] = lz_matches; c[4] = lz_spec_used; c[5] = lz_fast;
    for (int i = 0; i < 10; ++i) c[6 + i] = lz_why[i];
    c[16] = lz_chain;
    c[17] = lz_why[10];      // seed short cuts
}
extern "C" void lzani_emul_lane_why(long* c) { for (int i = 0; i < 16; ++i) c[i] = lz_lane_why[i]; }
extern "C" void lzani_emul_spec_stats(long* used, long* mismatches)
{
    *used = lz_spec_used;
    *mismatches = lz_spec_mismatches;
}

namespace {
struct VecOut {
    std::vector<RegionRec> v;
    void emit(const RegionRec& g) { v.push_back(g); }
};
}  // namespace

// regions of one pair through the kernel source's region fold (regions.h), sorted the way
// calc_regions sorts them (length descending, then seq_start)
extern "C" int lzani_emul_regions(const int32_t* p8, const uint8_t* ref3, uint32_t ref_len, const uint8_t* qry3,
                                  uint32_t qry_len, int32_t* out6, uint32_t cap, int32_t* stats3)
{
    Prm P{p8[0], p8[1], p8[2], p8[3], p8[4], p8[5], p8[6], p8[7]};
    HostSeq r, q;
    build_seq(r, ref3, ref_len, P);
    build_seq(q, qry3, qry_len, P);
    HostIndex ix;
    build_index(ix, r, P);
    PairView pv;
    pv.q = q.q.data(); pv.Dq = q.len + P.mrd;
    pv.r = r.r.data(); pv.Dr = r_form_len(r.len, P.mrd);
    pv.bitmap = ix.bitmap.data(); pv.bbits = ix.bbits;
    pv.table = ix.table.data(); pv.tbits = ix.tbits;
    HostWarp w;
    WarpScratch scratch;
    VecOut out;
    RegionTracker<VecOut> rt;
    rt.out = &out;
    PairResult res;
    if (is_default_kernel_params(P)) { const PrmDefault D{P.mrd, P.mqd, P.reg}; res = parse_pair(w, pv, D, &scratch, rt); }
    else res = parse_pair(w, pv, P, &scratch, rt);
    if (stats3) { stats3[0] = res.m; stats3[1] = res.x; stats3[2] = res.n; }
    std::stable_sort(out.v.begin(), out.v.end(), [](const RegionRec& a, const RegionRec& b) {
        const int la = a.seq_end - a.seq_start, lb = b.seq_end - b.seq_start;
        if (la != lb) return la > lb;
        return a.seq_start < b.seq_start;
    });
    for (size_t i = 0; i < out.v.size() && i < cap; ++i) {
        const RegionRec& g = out.v[i];
        int32_t* o = out6 + 6 * i;
        o[0] = g.ref_start; o[1] = g.ref_end; o[2] = g.seq_start; o[3] = g.seq_end; o[4] = g.num_matches; o[5] = g.num_mismatches;
    }
    return (int)out.v.size();
}

// The --out-filter decision of postproc.cu's F1 kernel (outfilter.h) for n unordered pairs, on the host:
// thr = gani, ani, tani, qcov, rcov; ab/ba = n x 3 results; masks[i] = line bits.
extern "C" void lzani_emul_out_filter(const double* thr, uint64_t n, const int32_t* ab, const int32_t* ba,
                                      const uint32_t* len_a, const uint32_t* len_b, uint32_t* masks)
{
    const OutFilter f{thr[0], thr[1], thr[2], thr[3], thr[4]};
    for (uint64_t i = 0; i < n; ++i)
        masks[i] = out_filter_mask(f, DirStats{ab[3 * i], ab[3 * i + 1], ab[3 * i + 2]},
                                   DirStats{ba[3 * i], ba[3 * i + 1], ba[3 * i + 2]}, len_a[i], len_b[i]);
}

// K1's element functions on the host: the Q-form and R-form records of one sequence from either symbol
// source (ascii = 0: three_in_byte bytes; 1: the loader's text, one byte per base).
extern "C" void lzani_emul_pack_records(const uint8_t* src, uint32_t len, int mrd, int ascii, uint32_t* q_out, uint32_t* r_out)
{
    const int nq = q_form_recs((int)len, mrd), nr = r_form_recs((int)len, mrd);
    for (int i = 0; i < nq; ++i) {
        const Rec o = ascii ? pack_q_record_from(SymAscii{src}, len, (uint32_t)i) : pack_q_record_from(SymPacked3{src}, len, (uint32_t)i);
        memcpy(q_out + 4 * i, &o, 16);
    }
    for (int i = 0; i < nr; ++i) {
        const Rec o = ascii ? pack_r_record_from(SymAscii{src}, len, (uint32_t)mrd, (uint32_t)i)
                            : pack_r_record_from(SymPacked3{src}, len, (uint32_t)mrd, (uint32_t)i);
        memcpy(r_out + 4 * i, &o, 16);
    }
}
// the register bit-matrix transpose of the all2all scan (parse_core.h: transpose32)
extern "C" void lzani_emul_transpose32(uint32_t* a32)
{
    uint32_t a[32];
    memcpy(a, a32, sizeof(a));
    transpose32(a);
    memcpy(a32, a, sizeof(a));
}
// the matrix form of the scan for one (reference, query) pair: hit masks of every query word, computed the
// way k_bitmap_matrix + k_scan_matrix do (this reference in column `col` of a 32-reference block whose other
// columns hold arbitrary bits) -- must equal scan_word().  Returns the number of differing words.
extern "C" int lzani_emul_matrix_scan_check(const int32_t* p8, const uint8_t* ref3, uint32_t ref_len, const uint8_t* qry3,
                                            uint32_t qry_len, int col, uint32_t noise)
{
    Prm P{p8[0], p8[1], p8[2], p8[3], p8[4], p8[5], p8[6], p8[7]};
    if (P.mal > 32) return -1;
    HostSeq r, q;
    build_seq(r, ref3, ref_len, P);
    build_seq(q, qry3, qry_len, P);
    HostIndex ix;
    build_index(ix, r, P);
    PairView pv;
    pv.q = q.q.data(); pv.Dq = q.len + P.mrd;
    pv.r = r.r.data(); pv.Dr = r_form_len(r.len, P.mrd);
    pv.bitmap = ix.bitmap.data(); pv.bbits = ix.bbits;
    pv.table = ix.table.data(); pv.tbits = ix.tbits;
    // k_bitmap_matrix: 32 keys per step, column `col` = this reference, the others pseudo-random
    const uint32_t n_keys = 1u << ix.bbits;
    std::vector<uint32_t> mc(n_keys);
    uint32_t x = noise | 1u;
    const uint32_t kw = n_keys / 32 ? n_keys / 32 : 1;
    for (uint32_t t = 0; t < kw; ++t) {
        uint32_t a[32];
        for (int i = 0; i < 32; ++i) { x = x * 1664525u + 1013904223u; a[i] = i == col ? ix.bitmap[t] : x; }
        transpose32(a);
        for (int b = 0; b < 32 && 32 * t + b < n_keys; ++b) mc[32 * t + b] = a[b];
    }
    // k_scan_matrix
    const int n_loop = pv.Dq - P.msl;
    const int n_words = n_loop > 0 ? (n_loop + 31) / 32 : 0;
    const uint32_t kmask = n_keys - 1u, km = low_mask32(P.mal);
    int bad = 0;
    for (int wd = 0; wd < n_words; ++wd) {
        const Rec a = pv.q[wd], b = pv.q[wd + 1];
        uint32_t v[32];
        if (2 * P.mal <= LZ_DIRECT_BITMAP_MAX_BITS) {
            const uint32_t h0 = a.y << P.mal, h1 = funnel_r(a.y, b.y, 32 - P.mal);
            for (int t = 0; t < 32; ++t) {
                const uint32_t lo = funnel_r(a.x, b.x, t), hs = funnel_r(h0, h1, t);
                v[t] = mc[((lo & km) | (hs & ~km)) & kmask];
            }
        } else {                       // hashed bitmaps of one size
            for (int t = 0; t < 32; ++t) {
                const uint32_t lo = funnel_r(a.x, b.x, t), hi = funnel_r(a.y, b.y, t);
                v[t] = mc[bitmap_index(kmer_key(lo, hi, P.mal), P.mal, ix.bbits)];
            }
        }
        transpose32(v);
        const uint32_t m = v[col] & scan_valid_mask(a, b, P.mal, 32 * wd, n_loop);
        if (m != scan_word(pv, P, 32 * wd, n_loop)) ++bad;
    }
    return bad;
}

// ---- TEST ONLY: the segment-parallel walk of DESIGN.md section 8 ("few long pairs"), tried out on the host.
// A walker started at an arbitrary record in the CLEAN state (no prediction, no open region) is claimed to merge
// with the true walk at the first clean point both pass through, and to reproduce the true totals from there on.
// Clean point: the walk is in anchor-search mode; its key is the query position of the next anchor event.  Totals
// are taken with the pending region folded in (kept iff re - rs >= reg), as the next distant match would judge it.
namespace {
struct CleanPoint { int key; long m, x, n; int kept; };
template <class PT>
void walk_logged(LaneCtx L, const PT& P, std::vector<CleanPoint>& log, long fin[3])
{
    RecAhead none;
    none.base = 0x80000000u;
    for (int k = 0; k < 4; ++k) none.r[k] = Rec{0, 0, 0, 0};
    int last_key = -2;
    while (L.mode != LANE_DONE) {
        if (L.s.pred < 0 && L.s.i < L.n_loop) {
            LaneCtx C = L;
            const int key = lane_next_event(C, C.s.i, LANE_NONE, false).qi;
            if (key != last_key) {
                const bool keep = L.s.rs >= 0 && L.s.re - L.s.rs >= P.reg;
                log.push_back(CleanPoint{key, L.s.totM + (keep ? L.s.curM : 0), L.s.totX + (keep ? L.s.curX : 0), L.s.totN + (keep ? 1 : 0), keep ? L.s.re : L.s.kept_end});
                last_key = key;
            }
        }
        if (L.mode == LANE_RUN) lane_run_step(L, P, none);
        else if (L.mode == LANE_SCAN) lane_scan_step(L, P);
        else lane_slow_step(L, P);
    }
    fin[0] = L.s.totM; fin[1] = L.s.totX; fin[2] = L.s.totN;
}
}  // namespace

// st[0] boundaries tried, st[1] walkers that merged with the true walk, st[2] of those: totals predicted wrongly,
// st[3] sum over the merged walkers of (clean points of their own before the merge), st[4] merged walkers for which a fresh
// walker started at the merge point WITH the true end of the last kept region reproduces the true totals (must be all)
extern "C" int lzani_emul_segment_check(const int32_t* p8, const uint8_t* ref3, uint32_t ref_len, const uint8_t* qry3,
                                        uint32_t qry_len, int n_segments, long* st)
{
    Prm P{p8[0], p8[1], p8[2], p8[3], p8[4], p8[5], p8[6], p8[7]};
    HostSeq r, q;
    build_seq(r, ref3, ref_len, P);
    build_seq(q, qry3, qry_len, P);
    HostIndex ix;
    build_index(ix, r, P);
    PairView pv;
    pv.q = q.q.data(); pv.Dq = q.len + P.mrd;
    pv.r = r.r.data(); pv.Dr = r_form_len(r.len, P.mrd);
    pv.bitmap = ix.bitmap.data(); pv.bbits = ix.bbits;
    pv.table = ix.table.data(); pv.tbits = ix.tbits;
    uint32_t n = 0;
    const std::vector<Rec> recs = staged_records(pv, P, n);
    LaneCtx L0;
    lane_begin(L0, pv, recs.data(), n, P.msl);
    std::vector<CleanPoint> truth;
    long fin[3];
    walk_logged(L0, P, truth, fin);
    for (int j = 1; j < n_segments && n; ++j) {
        const uint32_t c = (uint32_t)((uint64_t)n * j / n_segments);
        if (c == 0 || c >= n) continue;
        ++st[0];
        LaneCtx L;
        lane_begin(L, pv, recs.data(), n, P.msl);
        L.c = c;
        L.s.i = rec_pos(recs[c]);
        L.s.lit = P.mqd + 1;
        std::vector<CleanPoint> mine;
        long f2[3];
        walk_logged(L, P, mine, f2);
        // first clean point of this walker that the true walk also has
        for (size_t a = 0; a < mine.size(); ++a) {
            size_t b = 0;
            while (b < truth.size() && truth[b].key != mine[a].key) ++b;
            if (b == truth.size()) continue;
            ++st[1];
            st[3] += (long)a;
            const long pm = truth[b].m + (f2[0] - mine[a].m), px = truth[b].x + (f2[1] - mine[a].x), pn = truth[b].n + (f2[2] - mine[a].n);
            if (pm != fin[0] || px != fin[1] || pn != fin[2]) {
                ++st[2];
                if (getenv("LZ_SEGMENT_DEBUG"))
                    fprintf(stderr, "segment %d/%d start rec %u key %d: predicted %ld %ld %ld, true %ld %ld %ld (true totals at the merge %ld %ld %ld, walker's %ld %ld %ld)\n",
                            j, n_segments, c, mine[a].key, pm, px, pn, fin[0], fin[1], fin[2], truth[b].m, truth[b].x, truth[b].n, mine[a].m, mine[a].x, mine[a].n);
            }
            {
                // a fresh walker in the clean state at the merge point, told the true end of the last kept region, must
                // reproduce the rest of the true walk: the clean state + kept_end is ALL the state there is
                LaneCtx L2;
                lane_begin(L2, pv, recs.data(), n, P.msl);
                uint32_t ck = 0;
                while (ck < n && rec_pos(recs[ck]) + rec_run(recs[ck]) < mine[a].key) ++ck;
                L2.c = ck;
                L2.s.i = mine[a].key;
                L2.s.lit = P.mqd + 1;
                L2.s.kept_end = truth[b].kept;
                std::vector<CleanPoint> again;
                long f3[3];
                walk_logged(L2, P, again, f3);
                if (truth[b].m + f3[0] == fin[0] && truth[b].x + f3[1] == fin[1] && truth[b].n + f3[2] == fin[2]) ++st[4];
            }
            break;
        }
    }
    return 0;
}

extern "C" int lzani_emul_q_recs(uint32_t len, int mrd) { return q_form_recs((int)len, mrd); }
extern "C" int lzani_emul_r_recs(uint32_t len, int mrd) { return r_form_recs((int)len, mrd); }
