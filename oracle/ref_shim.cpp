// TEST INFRASTRUCTURE ONLY -- never linked into, imported by or called from the product path.
//
// ref_shim.cpp: thin extern "C" wrapper around the UNMODIFIED reference `CParser`
// (/root/reference/src/parser.{h,cpp}), compiled from the sources where they lie by
// oracle/Makefile into oracle/_ref/libref_parser.so.  It drives CParser exactly the way the
// worker lambda of CLZMatcher::do_matching does (reference src/lz_matcher.cpp:192-269):
// an atomic counter over reference rows, one private CParser per thread,
// prepare_reference once per row, then prepare_data / parse / calc_stats per query.
//
// Used by tests/ (to pin oracle/lzani_oracle.c and to check the CUDA path) and by
// bench.py's cpu_baseline / --impl reference legs (kind "reference").
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "parser.h"   // reference header, found through -I/root/reference/src

namespace {

CParams make_params(const int32_t* p)
{
    CParams cp;
    cp.min_anchor_len = p[0];
    cp.min_seed_len = p[1];
    cp.max_dist_in_ref = p[2];
    cp.max_dist_in_query = p[3];
    cp.min_region_len = p[4];
    cp.approx_window = p[5];
    cp.approx_mismatches = p[6];
    cp.approx_run_len = p[7];
    return cp;
}

}  // namespace

extern "C" {

// params8 = {mal, msl, mrd, mqd, reg, aw, am, ar}.  data[i] is three_in_byte packed (the
// reservoir's default, reference src/params.h:51), len[i] bases.  Rows as in lzani_run_rows:
// row j = reference ref_ids[j] (NULL: j), queries q_ids[row_off[j]..row_off[j+1]) or, with
// q_ids == NULL, every q != ref.  out3 receives 3 ints per pair in CSR order.
// Returns the wall-clock seconds of the matching loop (the "LZ matching" stage,
// lz_matcher.cpp:606-608), or a negative value on bad input.
double lzani_ref_run_rows(const int32_t* params8, uint32_t n_seqs, const uint8_t* const* data,
                          const uint32_t* len, uint32_t n_rows, const uint32_t* ref_ids,
                          const uint64_t* row_off, const uint32_t* q_ids, int32_t* out3,
                          int n_threads)
{
    if (!params8 || !data || !len || !out3 || (q_ids && !row_off))
        return -1.0;
    if (n_threads <= 0) {
        n_threads = (int)std::thread::hardware_concurrency();
        if (n_threads <= 0) n_threads = 1;
    }
    const CParams cp = make_params(params8);

    std::vector<uint64_t> dense_off;
    if (!q_ids) {
        dense_off.resize((size_t)n_rows + 1, 0);
        for (uint32_t j = 0; j < n_rows; ++j)
            dense_off[j + 1] = dense_off[j] + (n_seqs ? n_seqs - 1 : 0);
        row_off = dense_off.data();
    }

    std::atomic<uint64_t> next_row{0};
    auto t0 = std::chrono::high_resolution_clock::now();

    auto worker = [&]() {
        CParser parser(cp);
        while (true) {
            uint64_t j = next_row.fetch_add(1);
            if (j >= n_rows) break;
            uint32_t r = ref_ids ? ref_ids[j] : (uint32_t)j;
            uint64_t k = row_off[j];
            if (row_off[j + 1] == k) continue;
            parser.prepare_reference(seq_view(data[r], len[r], internal_packing_t::three_in_byte), 1);
            auto one = [&](uint32_t q) {
                parser.prepare_data(seq_view(data[q], len[q], internal_packing_t::three_in_byte), 1);
                parser.parse();
                results_t res = parser.calc_stats();
                out3[3 * k + 0] = res.sym_in_matches;
                out3[3 * k + 1] = res.sym_in_literals;
                out3[3 * k + 2] = res.no_components;
                ++k;
            };
            if (q_ids) {
                for (uint64_t e = row_off[j]; e < row_off[j + 1]; ++e) one(q_ids[e]);
            } else {
                for (uint32_t q = 0; q < n_seqs; ++q)
                    if (q != r) one(q);
            }
        }
    };

    std::vector<std::thread> pool;
    for (int t = 1; t < n_threads; ++t) pool.emplace_back(worker);
    worker();
    for (auto& th : pool) th.join();

    auto t1 = std::chrono::high_resolution_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

// One ordered pair: regions as CParser::get_parsing() returns them (parser.cpp:786-837).
// Writes up to cap regions of 6 ints (ref_start, ref_end, seq_start, seq_end, num_matches,
// num_mismatches); returns the full count, stats3 (optional) gets calc_stats().
int lzani_ref_regions(const int32_t* params8, const uint8_t* ref3, uint32_t ref_len, const uint8_t* qry3,
                      uint32_t qry_len, int32_t* out6, uint32_t cap, int32_t* stats3)
{
    const CParams cp = make_params(params8);
    CParser parser(cp);
    parser.prepare_reference(seq_view(ref3, ref_len, internal_packing_t::three_in_byte), 1);
    parser.prepare_data(seq_view(qry3, qry_len, internal_packing_t::three_in_byte), 1);
    parser.parse();
    std::vector<region_t> v = parser.get_parsing();
    for (size_t i = 0; i < v.size() && i < cap; ++i) {
        out6[6 * i + 0] = v[i].ref_start;
        out6[6 * i + 1] = v[i].ref_end;
        out6[6 * i + 2] = v[i].seq_start;
        out6[6 * i + 3] = v[i].seq_end;
        out6[6 * i + 4] = v[i].num_matches;
        out6[6 * i + 5] = v[i].num_mismatches;
    }
    if (stats3) {
        results_t res = parser.calc_stats();
        stats3[0] = res.sym_in_matches;
        stats3[1] = res.sym_in_literals;
        stats3[2] = res.no_components;
    }
    return (int)v.size();
}

}  // extern "C"
