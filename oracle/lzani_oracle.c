/*
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may build, load or call this file; the product path
 * (lz-ani_b200/csrc, the C ABI in include/lzani_b200.h) never does.
 *
 * lzani_oracle.c -- plain-C CPU restatement of the LZ-ANI per-pair hot path
 * (reference class CParser, src/parser.{h,cpp}).  Parity status: PINNED -- tests/test_oracle.py
 * checks it (a) against the committed golden vectors generated from the reference itself
 * (tests/golden/, produced by tests/golden/make_golden.py from oracle/_ref) and, where
 * oracle/_ref exists, (b) live against the unmodified reference CParser on vir61, the example
 * set, non-default parameter sets and seeded synthetic families.
 *
 * It is a restatement, not a copy: the index is a sorted (k-mer, position) array + a
 * counting-sorted seed list instead of the reference's MurmurHash open-addressing table
 * (the hash function and probe order are unobservable); the factor list,
 * the parse state machine, the approximate extensions, the two-sided gap re-alignment and
 * the region statistics follow the reference line by line in MEANING, each function citing
 * the lines it restates.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

enum { SYM_N_REF = 4, SYM_N_QRY = 5 };               /* defs.h:28-30 */
enum { F_CLOSE = 1, F_DISTANT = 2, F_LITERALS = 4 }; /* flag_t values used on the path, defs.h:32 */

typedef struct { int qpos, flag, rpos, len; } factor_t; /* defs.h:37-46 */

typedef struct {
    int mal, msl, mrd, mqd, reg, aw, am, ar;
} prm_t;

typedef struct { uint64_t kmer; int pos; } kp_t;

typedef struct {
    prm_t p;
    /* reference side */
    uint8_t* ref; int Dr, ref_cap;
    kp_t* anchors; int n_anchors, anchors_cap;   /* sorted by (kmer, pos): replaces ht_long   */
    int* seed_off; int* seed_pos; int seed_pos_cap; /* CSR over 4^msl: replaces ht_short(_desc) */
    /* query side */
    uint8_t* qry; int Dq, qry_cap;
    int64_t* q_anchor_kmer; int64_t* q_seed_kmer; int qk_cap;
    /* factor list */
    factor_t* f; int nf, f_cap;
    int* win; int win_cap;
    int* lft; int* rgt; uint8_t* lftm; uint8_t* rgtm; int side_cap;
} oracle_t;

static void* xrealloc(void* p, size_t n) { void* q = realloc(p, n ? n : 1); if (!q) abort(); return q; }

/* seq_view::unpack3 (seq_reservoir.h:145-169) for the three_in_byte packing of
 * CSeqReservoir::append (seq_reservoir.cpp:39-57). */
static void unpack3(const uint8_t* src, uint32_t len, uint8_t* dst)
{
    uint32_t full = len / 3;
    for (uint32_t i = 0; i < full; ++i) {
        unsigned b = src[i];
        dst[3 * i] = (uint8_t)(b / 36);
        dst[3 * i + 1] = (uint8_t)((b / 6) % 6);
        dst[3 * i + 2] = (uint8_t)(b % 6);
    }
    unsigned rem = len % 3;
    if (rem) {
        unsigned b = src[full];
        dst[3 * full] = (uint8_t)(b / 36);
        if (rem == 2) dst[3 * full + 1] = (uint8_t)((b / 6) % 6);
    }
}

/* k-mer value at every start position, -1 where the k symbols are not all ACGT or the
 * k-mer would run off the end (prepare_kmers, parser.cpp:53-103, store_all = true). */
static void kmers_at_every_pos(const uint8_t* s, int n, int k, int64_t* out)
{
    uint64_t mask = (~0ull) >> (64 - 2 * k);
    uint64_t v = 0; int good = 0;
    for (int i = 0; i < n; ++i) out[i] = -1;
    for (int i = 0; i < n; ++i) {
        if (s[i] >= 4) { good = 0; v = 0; }
        else { v = ((v << 2) | s[i]) & mask; ++good; }
        if (good >= k) out[i + 1 - k] = (int64_t)v;
    }
}

static int cmp_kp(const void* a, const void* b)
{
    const kp_t* x = (const kp_t*)a; const kp_t* y = (const kp_t*)b;
    if (x->kmer != y->kmer) return x->kmer < y->kmer ? -1 : 1;
    return (x->pos > y->pos) - (x->pos < y->pos);
}

/* prepare_reference (parser.cpp:16-34): seq_ref = fwd | mrd N | mrd N | revcomp | mrd N with
 * every N rewritten to 4; then the two indices (prepare_ht_long :146-189, prepare_ht_short
 * :106-143), here as sorted arrays giving the same ordered candidate sets. */
static void set_reference(oracle_t* o, const uint8_t* packed3, uint32_t len)
{
    const prm_t* p = &o->p;
    int Dr = 2 * (int)len + 3 * p->mrd;
    if (Dr + 64 > o->ref_cap) { o->ref_cap = Dr + 64; o->ref = xrealloc(o->ref, (size_t)o->ref_cap); }
    uint8_t* r = o->ref;
    unpack3(packed3, len, r);
    for (uint32_t i = 0; i < len; ++i) if (r[i] >= 4) r[i] = SYM_N_REF;
    memset(r + len, SYM_N_REF, 2 * (size_t)p->mrd);
    uint8_t* rc = r + len + 2 * p->mrd;
    for (uint32_t i = 0; i < len; ++i) { uint8_t c = r[len - 1 - i]; rc[i] = c < 4 ? (uint8_t)(3 - c) : SYM_N_REF; }
    memset(rc + len, SYM_N_REF, (size_t)p->mrd);
    o->Dr = Dr;

    int64_t* tmp = malloc(sizeof(int64_t) * (size_t)(Dr + 1));
    if (!tmp) abort();
    /* anchors */
    kmers_at_every_pos(r, Dr, p->mal, tmp);
    if (Dr > o->anchors_cap) { o->anchors_cap = Dr; o->anchors = xrealloc(o->anchors, sizeof(kp_t) * (size_t)Dr); }
    int na = 0;
    for (int i = 0; i < Dr; ++i) if (tmp[i] >= 0) { o->anchors[na].kmer = (uint64_t)tmp[i]; o->anchors[na].pos = i; ++na; }
    qsort(o->anchors, (size_t)na, sizeof(kp_t), cmp_kp);
    o->n_anchors = na;
    /* seeds: counting sort by k-mer keeps positions ascending inside a bucket */
    kmers_at_every_pos(r, Dr, p->msl, tmp);
    size_t nb = (size_t)1 << (2 * p->msl);
    o->seed_off = xrealloc(o->seed_off, sizeof(int) * (nb + 1));
    memset(o->seed_off, 0, sizeof(int) * (nb + 1));
    if (Dr > o->seed_pos_cap) { o->seed_pos_cap = Dr; o->seed_pos = xrealloc(o->seed_pos, sizeof(int) * (size_t)Dr); }
    for (int i = 0; i < Dr; ++i) if (tmp[i] >= 0) ++o->seed_off[tmp[i] + 1];
    for (size_t b = 0; b < nb; ++b) o->seed_off[b + 1] += o->seed_off[b];
    int* cur = malloc(sizeof(int) * nb);
    if (!cur) abort();
    memcpy(cur, o->seed_off, sizeof(int) * nb);
    for (int i = 0; i < Dr; ++i) if (tmp[i] >= 0) o->seed_pos[cur[tmp[i]]++] = i;
    free(cur);
    free(tmp);
}

/* prepare_data (parser.cpp:37-50): seq_data = query (N stays 5) | mrd x 5, plus its k-mers. */
static void set_query(oracle_t* o, const uint8_t* packed3, uint32_t len)
{
    const prm_t* p = &o->p;
    int Dq = (int)len + p->mrd;
    if (Dq + 64 > o->qry_cap) { o->qry_cap = Dq + 64; o->qry = xrealloc(o->qry, (size_t)o->qry_cap); }
    unpack3(packed3, len, o->qry);
    for (uint32_t i = 0; i < len; ++i) if (o->qry[i] >= 4) o->qry[i] = SYM_N_QRY;
    memset(o->qry + len, SYM_N_QRY, (size_t)p->mrd);
    o->Dq = Dq;
    if (Dq + 1 > o->qk_cap) {
        o->qk_cap = Dq + 1;
        o->q_anchor_kmer = xrealloc(o->q_anchor_kmer, sizeof(int64_t) * (size_t)o->qk_cap);
        o->q_seed_kmer = xrealloc(o->q_seed_kmer, sizeof(int64_t) * (size_t)o->qk_cap);
    }
    kmers_at_every_pos(o->qry, Dq, p->mal, o->q_anchor_kmer);
    kmers_at_every_pos(o->qry, Dq, p->msl, o->q_seed_kmer);
}

static void push(oracle_t* o, int qpos, int flag, int rpos, int len)
{
    if (o->nf == o->f_cap) { o->f_cap = o->f_cap ? 2 * o->f_cap : 1024; o->f = xrealloc(o->f, sizeof(factor_t) * (size_t)o->f_cap); }
    factor_t* x = &o->f[o->nf++];
    x->qpos = qpos; x->flag = flag; x->rpos = rpos; x->len = len;
}

/* equal_len (parser.cpp:192-207) */
static int common_prefix(const oracle_t* o, int rpos, int qpos, int from)
{
    int lim = o->Dr - rpos < o->Dq - qpos ? o->Dr - rpos : o->Dq - qpos;
    int n = from;
    while (n < lim && o->ref[rpos + n] == o->qry[qpos + n]) ++n;
    return n;
}

/* compare_ranges (parser.cpp:210-248): run-length encode the columns of two aligned ranges.
 * The first matching run carries `first_flag`, later ones are close matches. */
static void emit_columns(oracle_t* o, int q0, int r0, int len, int first_flag)
{
    int flag = first_flag;
    int j = 0;
    while (j < len) {
        int same = o->ref[r0 + j] == o->qry[q0 + j];
        int e = j + 1;
        while (e < len && (o->ref[r0 + e] == o->qry[q0 + e]) == same) ++e;
        if (same) { push(o, q0 + j, flag, r0 + j, e - j); if (e < len) flag = F_CLOSE; }
        else push(o, q0 + j, F_LITERALS, 0, e - j);
        j = e;
    }
}

static void need_sides(oracle_t* o, int n)
{
    if (n <= o->side_cap) return;
    o->side_cap = n + 64;
    o->lft = xrealloc(o->lft, sizeof(int) * (size_t)o->side_cap);
    o->rgt = xrealloc(o->rgt, sizeof(int) * (size_t)o->side_cap);
    o->lftm = xrealloc(o->lftm, (size_t)o->side_cap);
    o->rgtm = xrealloc(o->rgtm, (size_t)o->side_cap);
}

/* compare_ranges_both_ways (parser.cpp:251-374).  Gap of `len` query bases starting at q0 that
 * precedes a close match: prefix counts against the reference continuing from `r_left`,
 * suffix counts against the reference ending at `r_right_end` (which the caller passes as the
 * END of the new match -- parser.cpp:632 -- and which is kept that way), the LAST split with
 * the maximal sum wins, then the columns are emitted; the position bookkeeping of the right
 * part reproduces parser.cpp:353-368 including the case where the first right column is
 * merged into a preceding literal run without advancing the running position. */
static void emit_gap_both_ways(oracle_t* o, int q0, int r_left, int r_right_end, int len)
{
    int span = r_right_end < r_left ? len : (r_right_end - r_left < len ? r_right_end - r_left : len);
    need_sides(o, span + 2);
    int acc = 0;
    o->lft[0] = 0; o->lftm[0] = 0;
    for (int t = 0; t < span; ++t) {
        int same = o->ref[r_left + t] == o->qry[q0 + t];
        acc += same; o->lft[t + 1] = acc; o->lftm[t + 1] = (uint8_t)same;
    }
    acc = 0;
    o->rgt[0] = 0; o->rgtm[0] = 0;
    int rmax = span < r_right_end ? span : r_right_end;
    for (int t = 1; t <= span; ++t) {
        if (t <= rmax) {
            int same = o->ref[r_right_end - t] == o->qry[q0 + len - t];
            acc += same; o->rgt[t] = acc; o->rgtm[t] = (uint8_t)same;
        } else { o->rgt[t] = 0; o->rgtm[t] = 0; }
    }
    int best = 0, split = 0;
    for (int s = 0; s <= span; ++s) {
        int v = o->lft[s] + o->rgt[span - s];
        if (v >= best) { best = v; split = s; }
    }
    /* left columns */
    if (split > 0) {
        int same = o->lftm[1];
        int qp = q0;
        push(o, qp++, same ? F_CLOSE : F_LITERALS, same ? r_left : 0, 1);
        for (int s = 2; s <= split; ++s, ++qp) {
            same = o->lftm[s];
            int fl = same ? F_CLOSE : F_LITERALS;
            if (o->f[o->nf - 1].flag == fl) ++o->f[o->nf - 1].len;
            else push(o, qp, fl, same ? r_left + s - 1 : 0, 1);
        }
    }
    /* unscanned middle */
    if (span < len) {
        if (split > 0 && o->f[o->nf - 1].flag == F_LITERALS) o->f[o->nf - 1].len += len - span;
        else push(o, q0 + split, F_LITERALS, 0, len - span);
    }
    /* right columns */
    if (split < span) {
        int shift = len - span;
        int from_right = span - split;
        int same = o->rgtm[from_right];
        int qp = q0 + split + shift;
        if (!same && (split > 0 || shift > 0) && o->f[o->nf - 1].flag == F_LITERALS) ++o->f[o->nf - 1].len;
        else push(o, qp++, same ? F_CLOSE : F_LITERALS, same ? r_right_end - from_right : 0, 1);
        for (int t = from_right - 1; t > 0; --t, ++qp) {
            same = o->rgtm[t];
            int fl = same ? F_CLOSE : F_LITERALS;
            if (o->f[o->nf - 1].flag == fl) ++o->f[o->nf - 1].len;
            else push(o, qp, fl, same ? r_right_end - t : 0, 1);
        }
    }
}

/* try_extend_forward (parser.cpp:377-409) / try_extend_backward (:412-441): walk base by base
 * with a ring of the last `aw` mismatch flags; stop once more than `am` of them are set; the
 * answer is the end of the last run of >= `ar` matches (the run counter starts AT `ar`). */
static int extend(oracle_t* o, int q0, int r0, int dir, int max_len)
{
    const prm_t* p = &o->p;
    if (p->aw > o->win_cap) { o->win_cap = p->aw; o->win = xrealloc(o->win, sizeof(int) * (size_t)p->aw); }
    memset(o->win, 0, sizeof(int) * (size_t)p->aw);
    int bad = 0, last = 0, run = p->ar;
    for (int e = 0;; ++e) {
        int qi, ri;
        if (dir > 0) { qi = q0 + e; ri = r0 + e; if (!(qi < o->Dq && ri < o->Dr)) break; }
        else { if (!(q0 - e > 0 && r0 - e > 0 && e < max_len)) break; qi = q0 - e - 1; ri = r0 - e - 1; }
        int mm = o->qry[qi] != o->ref[ri];
        int slot = e % p->aw;
        bad += mm - o->win[slot];
        o->win[slot] = mm;
        if (!mm) { if (++run >= p->ar) last = e + 1; }
        else run = 0;
        if (bad > p->am) break;
    }
    return last;
}

/* prob_len (parser.h:134-172) and ipow (parser.h:174-188). */
static double four_to_minus(int len)
{
    if (len < 30) return ldexp(1.0, -2 * len); /* the table entries are exactly 4^-len */
    return pow(4, -len);
}
static double ipow_u32(double base, uint32_t e)
{
    double r = 1.0;
    while (e) { if (e & 1) r *= base; base *= base; e >>= 1; }
    return r;
}

/* all reference positions whose mal-mer equals `kmer`, ascending (the ht_long probe chain,
 * parser.cpp:516-530 / :587-601, minus foreign keys).  Returns the best (longest, then
 * smallest position because of the strict '>'). */
static void best_anchor(const oracle_t* o, int64_t kmer, int qpos, int* out_pos, int* out_len)
{
    *out_pos = 0; *out_len = 0;
    if (kmer < 0) return;
    int lo = 0, hi = o->n_anchors;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (o->anchors[mid].kmer < (uint64_t)kmer) lo = mid + 1; else hi = mid; }
    for (int j = lo; j < o->n_anchors && o->anchors[j].kmer == (uint64_t)kmer; ++j) {
        int L = common_prefix(o, o->anchors[j].pos, qpos, 0);
        if (L < o->p.mal) continue;
        if (L > *out_len) { *out_len = L; *out_pos = o->anchors[j].pos; }
    }
}

/* CParser::parse (parser.cpp:482-716). */
static void parse_pair(oracle_t* o)
{
    const prm_t* p = &o->p;
    const int Dq = o->Dq;
    int pred = -Dq;          /* ref_pred_pos */
    int lit = 0;             /* cur_lit_run_len */
    int reg_start = -1;      /* prev_region_start */
    int reg_end = 0;         /* prev_region_end */
    int i = 0;
    o->nf = 0;

    while (i + p->msl < Dq) {
        int best_pos = 0, best_len = 0;
        if (pred < 0) {
            /* no prediction: anchors only (:507-532) */
            best_anchor(o, o->q_anchor_kmer[i], i, &best_pos, &best_len);
        } else {
            /* seeds near the predicted position (:548-580) */
            int64_t sk = o->q_seed_kmer[i];
            if (sk >= 0) {
                const int* b = o->seed_pos + o->seed_off[sk];
                int nb = o->seed_off[sk + 1] - o->seed_off[sk];
                int from = pred - lit;
                int lo = 0, hi = nb;
                while (lo < hi) { int mid = (lo + hi) >> 1; if (b[mid] < from) lo = mid + 1; else hi = mid; }
                for (int j = lo; j < nb && b[j] < pred + p->mrd; ++j) {
                    int L = common_prefix(o, b[j], i, p->msl);
                    if (L > best_len) { best_len = L; best_pos = b[j]; }
                    else if (L == best_len && abs(b[j] - pred) < abs(best_pos - pred)) best_pos = b[j];
                }
            }
            /* anchors anywhere (:582-602) and the chance arbitration (:604-623) */
            int a_pos, a_len;
            best_anchor(o, o->q_anchor_kmer[i], i, &a_pos, &a_len);
            if (a_pos) {
                if (!best_pos) { best_pos = a_pos; best_len = a_len; }
                else {
                    double pa = ipow_u32(1 - four_to_minus(a_len), (uint32_t)(int)(2 * ((size_t)o->Dr + 1 - (size_t)a_len)));
                    double pc = ipow_u32(1 - four_to_minus(best_len), (uint32_t)(lit + p->mrd + 1 - best_len));
                    if (pa > pc) { best_pos = a_pos; best_len = a_len; }
                }
            }
        }

        if (best_len >= p->msl) {
            int flag = F_DISTANT;
            if (pred >= 0 && abs(best_pos - pred) <= p->mrd) {
                /* close match: re-align the literal gap from both ends (:630-635) */
                emit_gap_both_ways(o, i - lit, pred - lit, best_pos + best_len, lit);
                push(o, i, F_CLOSE, best_pos, best_len);
            } else {
                /* distant match (:636-685) */
                if (lit) push(o, i - lit, F_LITERALS, 0, lit);
                if (reg_start >= 0 && !(reg_end - reg_start >= p->reg)) {   /* eval_region :446-449 */
                    while (o->nf && o->f[o->nf - 1].qpos >= reg_start) --o->nf;
                    int run = i - reg_start;
                    while (o->nf && o->f[o->nf - 1].flag == F_LITERALS) { run += o->f[o->nf - 1].len; --o->nf; }
                    push(o, i - run, F_LITERALS, 0, run);
                    reg_start = -1;
                }
                if (o->nf && o->f[o->nf - 1].flag == F_LITERALS) {
                    int back = extend(o, i, best_pos, -1, o->f[o->nf - 1].len);
                    if (back) {
                        o->f[o->nf - 1].len -= back;
                        if (o->f[o->nf - 1].len == 0) --o->nf;
                        emit_columns(o, i - back, best_pos - back, back, F_DISTANT);
                        flag = F_CLOSE;
                        reg_start = i - back;
                    }
                }
                push(o, i, flag, best_pos, best_len);
                if (flag == F_DISTANT) reg_start = i;
                if (reg_start < 0)                               /* :678-684 */
                    for (int j = o->nf - 1; j >= 0; --j)
                        if (o->f[j].flag == F_DISTANT) { reg_start = o->f[j].qpos; break; }
            }
            i += best_len;
            pred = best_pos + best_len;
            lit = 0;
            int fwd = extend(o, i, pred, +1, 0);
            emit_columns(o, i, pred, fwd, F_CLOSE);
            i += fwd;
            pred += fwd;
            reg_end = i;
        } else {
            ++i; ++pred; ++lit;
        }
        if (lit > p->mqd) pred = -Dq;
    }
    /* tail (:710-713) */
    if (pred < 0) push(o, i - lit, F_LITERALS, 0, lit + (Dq - i));
    else {
        /* the reference reads seq_ref[] here without a bound; positions at or past Dr hold no
         * defined value there -- they are treated as "different" (only reachable with mqd > mrd). */
        int q0 = i - lit, r0 = pred - lit - p->msl, n = lit + (Dq - i);
        int safe = n;
        if (r0 + safe > o->Dr) safe = o->Dr - r0 > 0 ? o->Dr - r0 : 0;
        emit_columns(o, q0, r0, safe, F_CLOSE);
        if (safe < n) {
            if (o->nf && o->f[o->nf - 1].flag == F_LITERALS && safe > 0) o->f[o->nf - 1].len += n - safe;
            else push(o, q0 + safe, F_LITERALS, 0, n - safe);
        }
    }
}

/* calc_stats (parser.cpp:734-783).  The descending sort there does not change the sums. */
static void stats_of_parse(const oracle_t* o, int32_t* out3)
{
    int M = 0, X = 0, pend = 0;
    int tm = 0, tx = 0, tn = 0;
    for (int j = 0; j < o->nf; ++j) {
        const factor_t* x = &o->f[j];
        if (x->flag == F_DISTANT) {
            if (M && M + X >= o->p.reg) { tm += M; tx += X; ++tn; }
            M = x->len; X = 0; pend = 0;
        } else if (x->flag == F_CLOSE) { M += x->len; X += pend; pend = 0; }
        else pend += x->len;
    }
    if (M && M + X >= o->p.reg) { tm += M; tx += X; ++tn; }
    out3[0] = tm; out3[1] = tx; out3[2] = tn;
}

/* calc_regions (parser.cpp:786-837) with region_t's update rules (defs.h:67-153). */
typedef struct { int rs, re, qs, qe, m, x; } reg_t;

static void reg_clear(reg_t* g) { g->rs = g->re = g->qs = g->qe = -1; g->m = g->x = 0; }
static void reg_cover(reg_t* g, const factor_t* f)
{
    if (g->qs < 0 || f->qpos < g->qs) g->qs = f->qpos;
    if (g->qe < 0 || f->qpos + f->len > g->qe) g->qe = f->qpos + f->len;
    if (g->rs < 0 || f->rpos < g->rs) g->rs = f->rpos;
    if (g->re < 0 || f->rpos + f->len > g->re) g->re = f->rpos + f->len;
    g->m += f->len;
}
static int cmp_reg(const void* a, const void* b)
{
    const reg_t* x = (const reg_t*)a; const reg_t* y = (const reg_t*)b;
    int lx = x->qe - x->qs, ly = y->qe - y->qs;
    if (lx != ly) return lx > ly ? -1 : 1;
    return (x->qs > y->qs) - (x->qs < y->qs);
}
static int regions_of_parse(const oracle_t* o, reg_t** out)
{
    reg_t* v = NULL; int n = 0, cap = 0;
    reg_t cur; reg_clear(&cur);
    int pend = 0;
    for (int j = 0; j <= o->nf; ++j) {
        int fl = j < o->nf ? o->f[j].flag : F_DISTANT;
        if (fl == F_DISTANT) {
            if (cur.qe - cur.qs >= o->p.reg) {
                if (n == cap) { cap = cap ? 2 * cap : 64; v = xrealloc(v, sizeof(reg_t) * (size_t)cap); }
                v[n++] = cur;
            }
            if (j == o->nf) break;
            reg_clear(&cur);
            reg_cover(&cur, &o->f[j]);
            pend = 0;
        } else if (fl == F_CLOSE) {
            cur.re += pend; cur.qe += pend;      /* extend_region */
            cur.x += pend; pend = 0;
            reg_cover(&cur, &o->f[j]);
        } else pend += o->f[j].len;
    }
    /* std::sort is not stable, but (length, seq_start) is a total order over regions of one
     * parse except for exact duplicates, so any sort gives the reference's order. */
    qsort(v, (size_t)n, sizeof(reg_t), cmp_reg);
    *out = v;
    return n;
}

static oracle_t* oracle_new(const int32_t* p8)
{
    oracle_t* o = calloc(1, sizeof(oracle_t));
    if (!o) abort();
    o->p.mal = p8[0]; o->p.msl = p8[1]; o->p.mrd = p8[2]; o->p.mqd = p8[3];
    o->p.reg = p8[4]; o->p.aw = p8[5]; o->p.am = p8[6]; o->p.ar = p8[7];
    return o;
}
static void oracle_free(oracle_t* o)
{
    free(o->ref); free(o->anchors); free(o->seed_off); free(o->seed_pos); free(o->qry);
    free(o->q_anchor_kmer); free(o->q_seed_kmer); free(o->f); free(o->win);
    free(o->lft); free(o->rgt); free(o->lftm); free(o->rgtm); free(o);
}

/* ---- batch driver: same contract as lzani_run_rows (include/lzani_b200.h) and as the worker
 * loop of CLZMatcher::do_matching (lz_matcher.cpp:192-269): a shared row counter, one private
 * parser state per thread, the reference prepared once per row. ---- */
typedef struct {
    const int32_t* p8; uint32_t n_seqs; const uint8_t* const* data; const uint32_t* len;
    uint32_t n_rows; const uint32_t* ref_ids; const uint64_t* row_off; const uint32_t* q_ids;
    int32_t* out3; uint64_t next_row; pthread_mutex_t mu;
} job_t;

static void* job_worker(void* arg)
{
    job_t* J = (job_t*)arg;
    oracle_t* o = oracle_new(J->p8);
    for (;;) {
        pthread_mutex_lock(&J->mu);
        uint64_t j = J->next_row++;
        pthread_mutex_unlock(&J->mu);
        if (j >= J->n_rows) break;
        uint32_t r = J->ref_ids ? J->ref_ids[j] : (uint32_t)j;
        uint64_t k = J->row_off[j];
        if (J->row_off[j + 1] == k) continue;
        set_reference(o, J->data[r], J->len[r]);
        if (J->q_ids) {
            for (uint64_t e = J->row_off[j]; e < J->row_off[j + 1]; ++e, ++k) {
                uint32_t q = J->q_ids[e];
                set_query(o, J->data[q], J->len[q]); parse_pair(o); stats_of_parse(o, J->out3 + 3 * k);
            }
        } else {
            for (uint32_t q = 0; q < J->n_seqs; ++q) {
                if (q == r) continue;
                set_query(o, J->data[q], J->len[q]); parse_pair(o); stats_of_parse(o, J->out3 + 3 * k);
                ++k;
            }
        }
    }
    oracle_free(o);
    return NULL;
}

double lzani_oracle_run_rows(const int32_t* params8, uint32_t n_seqs, const uint8_t* const* data,
                             const uint32_t* len, uint32_t n_rows, const uint32_t* ref_ids,
                             const uint64_t* row_off, const uint32_t* q_ids, int32_t* out3, int n_threads)
{
    if (!params8 || !data || !len || !out3 || (q_ids && !row_off)) return -1.0;
    if (n_threads <= 0) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    uint64_t* dense = NULL;
    if (!q_ids) {
        dense = malloc(sizeof(uint64_t) * ((size_t)n_rows + 1));
        if (!dense) abort();
        dense[0] = 0;
        for (uint32_t j = 0; j < n_rows; ++j) dense[j + 1] = dense[j] + (n_seqs ? n_seqs - 1 : 0);
        row_off = dense;
    }
    job_t J = { params8, n_seqs, data, len, n_rows, ref_ids, row_off, q_ids, out3, 0, PTHREAD_MUTEX_INITIALIZER };
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    pthread_t th[256];
    for (int t = 1; t < n_threads; ++t) pthread_create(&th[t], NULL, job_worker, &J);
    job_worker(&J);
    for (int t = 1; t < n_threads; ++t) pthread_join(th[t], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    free(dense);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

int lzani_oracle_regions(const int32_t* params8, const uint8_t* ref3, uint32_t ref_len, const uint8_t* qry3,
                         uint32_t qry_len, int32_t* out6, uint32_t cap, int32_t* stats3)
{
    oracle_t* o = oracle_new(params8);
    set_reference(o, ref3, ref_len);
    set_query(o, qry3, qry_len);
    parse_pair(o);
    if (stats3) stats_of_parse(o, stats3);
    reg_t* v = NULL;
    int n = regions_of_parse(o, &v);
    for (int i = 0; i < n && (uint32_t)i < cap; ++i) {
        out6[6 * i] = v[i].rs; out6[6 * i + 1] = v[i].re; out6[6 * i + 2] = v[i].qs;
        out6[6 * i + 3] = v[i].qe; out6[6 * i + 4] = v[i].m; out6[6 * i + 5] = v[i].x;
    }
    free(v);
    oracle_free(o);
    return n;
}

/* Factor dump of one pair, for debugging the CUDA path: 4 ints per factor. */
int lzani_oracle_factors(const int32_t* params8, const uint8_t* ref3, uint32_t ref_len, const uint8_t* qry3,
                         uint32_t qry_len, int32_t* out4, uint32_t cap)
{
    oracle_t* o = oracle_new(params8);
    set_reference(o, ref3, ref_len);
    set_query(o, qry3, qry_len);
    parse_pair(o);
    int n = o->nf;
    for (int i = 0; i < n && (uint32_t)i < cap; ++i) {
        out4[4 * i] = o->f[i].qpos; out4[4 * i + 1] = o->f[i].flag; out4[4 * i + 2] = o->f[i].rpos; out4[4 * i + 3] = o->f[i].len;
    }
    oracle_free(o);
    return n;
}
