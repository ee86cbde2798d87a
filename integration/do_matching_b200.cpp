// do_matching_b200.cpp -- the ONE function of the reference's host program that this repository
// replaces: CLZMatcher::do_matching (reference src/lz_matcher.cpp:172-277).
//
// The reference's worker threads each own a CParser and loop
//     prepare_reference(r); for q: prepare_data(q); parse(); calc_stats()
// (lz_matcher.cpp:192-269).  Here the same contract is a few batch submits through the C ABI of
// include/lzani_b200.h, from this one host thread: ONE context drives all the GPUs of LZANI_DEVICES
// (lzani_create with n_devices = G: the library owns the devices, its worker threads and the dealing
// of reference rows over them); the reservoir's three_in_byte bytes go to the devices
// once (lzani_set_sequences), the pair list goes as CSR rows (all2all: q_ids == NULL; --flt-kmerdb:
// the filter rows), and `results[r]` is filled with id-sorted id_results_t exactly as the reference
// leaves it for store_results (lz_matcher.cpp:252-255, :404-418).
//
// With --out-filter (two-TSV output, no --out-alignment) the results stay on the GPUs: after the parse
// every GPU pairs up the two directions of its unordered pairs -- reading the rows owned by the other
// GPUs in place over NVLink peer access -- evaluates the reference's filter arithmetic
// (lz_matcher.cpp:434-463) and ships only the surviving pairs (lzani_filter_pairs).
// `results` then holds just those pairs; the unmodified store_results re-tests and formats them, so
// the text is identical while the D2H volume and the host-side result table shrink to the survivors.
//
// Everything else of the reference (CLI, params, loaders, filter, reorder, TSV writers) is compiled
// UNMODIFIED from /root/reference/src by integration/Makefile; this file is the only host source of
// ours in the binary.  There is no CPU fallback: without a B200 the program prints the library's
// error and exits 1 (the reference's own error convention, lz_matcher.h:61-65).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <iostream>
#include <string>
#include <thread>
#include <vector>

#include "lz_matcher.h"          // reference header (-I /root/reference/src)
#include "lzani_b200.h"          // our C ABI (-I include)

bool lzani_device_ingest();      // loader_hooks_b200.cpp

namespace {

std::vector<int> devices_from_env()
{
    // LZANI_DEVICES="0,1,2,3": the GPUs to use (default: device 0)
    std::vector<int> d;
    if (const char* e = getenv("LZANI_DEVICES")) {
        std::string s(e);
        size_t p = 0;
        while (p < s.size()) {
            size_t c = s.find(',', p);
            if (c == std::string::npos) c = s.size();
            if (c > p) d.push_back(atoi(s.substr(p, c - p).c_str()));
            p = c + 1;
        }
    }
    if (d.empty()) d.push_back(0);
    return d;
}

[[noreturn]] void die(const std::string& what, lzani_ctx* c)
{
    std::cerr << "lzani-b200: " << what << ": " << (c ? lzani_last_error(c) : "no context") << std::endl;
    exit(1);
}

// f(j) for j in [0, n) on the host cores (building and sorting the result rows is the host's share of the matching stage)
template <class F>
void parallel_rows(size_t n, unsigned n_threads, F f)
{
    std::atomic<size_t> next{0};
    auto work = [&] { for (size_t j; (j = next.fetch_add(1)) < n;) f(j); };
    std::vector<std::thread> th;
    for (unsigned t = 1; t < n_threads && t < n; ++t) th.emplace_back(work);
    work();
    for (auto& t : th) t.join();
}

}  // namespace

void CLZMatcher::do_matching()
{
    if (params.verbosity_level >= 1)
        cerr << "All2all sparse (B200)" << endl;

    const uint32_t n = (uint32_t)seq_reservoir.size();
    results.clear();
    results.resize(n);

    lzani_params_t P;
    P.mal = (int32_t)params.min_anchor_len;
    P.msl = (int32_t)params.min_seed_len;
    P.mrd = (int32_t)params.max_dist_in_ref;
    P.mqd = (int32_t)params.max_dist_in_query;
    P.reg = (int32_t)params.min_region_len;
    P.aw = (int32_t)params.approx_window;
    P.am = (int32_t)params.approx_mismatches;
    P.ar = (int32_t)params.approx_run_len;

    std::vector<const uint8_t*> data(n);
    std::vector<uint32_t> len(n);
    for (uint32_t i = 0; i < n; ++i) {
        auto it = seq_reservoir.get_sequence(i);
        data[i] = it->data;
        len[i] = it->len;
    }
    if (params.internal_packing != internal_packing_t::three_in_byte) {
        cerr << "lzani-b200: only the default three_in_byte reservoir packing is supported" << endl;
        exit(1);
    }

    const bool dense = filter.is_empty();
    const std::vector<int> devs = devices_from_env();
    const uint32_t G = (uint32_t)devs.size();
    const unsigned host_threads = std::max(1u, params.no_threads ? (unsigned)params.no_threads : std::thread::hardware_concurrency());

    // where the matching stage goes (-V 2): context creation, sequence upload, device submits, building the result rows
    using clk = std::chrono::steady_clock;
    auto secs = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };
    double t_submit = 0, t_rows = 0;
    const auto t_begin = clk::now();
    lzani_ctx* ctx = nullptr;
    if (lzani_create(&ctx, &P, devs.data(), (int)G) != LZANI_OK) die("lzani_create", ctx);
    const auto t_created = clk::now();
    if (lzani_device_ingest()) {
        // LZANI_DEVICE_INGEST=1: the reservoir holds the loader's text (loader_hooks_b200.cpp); packing happens on the device
        if (lzani_set_sequences_text(ctx, n, data.data(), len.data()) != LZANI_OK) die("lzani_set_sequences_text", ctx);
    } else if (lzani_set_sequences(ctx, n, data.data(), len.data()) != LZANI_OK) die("lzani_set_sequences", ctx);
    const auto t_loaded = clk::now();
    auto report = [&] {
        if (params.verbosity_level >= 2)
            cerr << "  B200 matching stage: create " << secs(t_begin, t_created) << " s, sequences to " << G << " GPU(s) " << secs(t_created, t_loaded)
                 << " s, submits " << t_submit << " s, result rows " << t_rows << " s" << endl;
    };

    // the pair list in CSR form (sparse lists: the filter's rows, ids ascending)
    uint64_t total_pairs = 0;
    if (dense) total_pairs = (uint64_t)n * (n ? n - 1 : 0);
    else for (uint32_t r = 0; r < n; ++r) total_pairs += filter.get_row(r).size();

    // ---- device-side --out-filter (see the header comment).  The results of all rows stay in HBM (12 B per pair
    // + 1 flag byte): when that would not fit comfortably, the host filter path below takes over.
    bool device_filter = params.output_filter_mask != 0 && params.output_type == output_type_t::two_tsv &&
                         !f_alignment && !getenv("LZANI_HOST_FILTER");
    if (device_filter && total_pairs / G * 13ull > (64ull << 30)) device_filter = false;
    if (device_filter && total_pairs) {
        lzani_out_filter_t flt{};
        flt.gani = params.output_filter_vals[(uint32_t)output_component_t::gani];
        flt.ani = params.output_filter_vals[(uint32_t)output_component_t::ani];
        flt.tani = params.output_filter_vals[(uint32_t)output_component_t::tani];
        flt.qcov = params.output_filter_vals[(uint32_t)output_component_t::qcov];
        flt.rcov = params.output_filter_vals[(uint32_t)output_component_t::rcov];
        std::vector<uint32_t> eff_len(n);
        for (uint32_t i = 0; i < n; ++i) {
            auto it = seq_reservoir.get_sequence(i);
            eff_len[i] = it->len - (it->no_parts - 1) * (uint32_t)params.max_dist_in_ref;   // lz_matcher.cpp:437
        }
        std::vector<uint32_t> q_ids;
        std::vector<uint64_t> row_off(1, 0);
        if (!dense) {
            q_ids.reserve(total_pairs);
            for (uint32_t r = 0; r < n; ++r) {
                const size_t a = q_ids.size();
                for (auto id : filter.get_row(r)) q_ids.push_back((uint32_t)id);
                std::sort(q_ids.begin() + a, q_ids.end());
                row_off.push_back(q_ids.size());
                filter.clear_row(r);
            }
        }
        // one submit with every row; the results stay on the GPUs that parsed them
        const auto t_s0 = clk::now();
        if (lzani_run_rows_resident(ctx, n, nullptr, dense ? nullptr : row_off.data(), dense ? nullptr : q_ids.data(), 1, nullptr) != LZANI_OK)
            die("lzani_run_rows_resident", ctx);
        std::vector<lzani_pair_t> sv(1024);
        for (;;) {
            uint64_t need = 0;
            const int rc = lzani_filter_pairs(ctx, nullptr, 0, eff_len.data(), &flt, sv.data(), sv.size(), &need);
            if (rc == LZANI_ERR_NOMEM && need > sv.size()) { sv.resize(need); continue; }
            if (rc != LZANI_OK) die("lzani_filter_pairs", ctx);
            sv.resize(need);
            break;
        }
        const auto t_s1 = clk::now();
        t_submit = secs(t_s0, t_s1);
        // only the surviving pairs reach the host: both directions go into `results`, id-sorted, as
        // store_results expects them (lz_matcher.cpp:404-418)
        for (const lzani_pair_t& p : sv) {
            results[p.a].emplace_back(p.b, results_t(p.ab.sym_in_matches, p.ab.sym_in_literals, p.ab.no_components));
            results[p.b].emplace_back(p.a, results_t(p.ba.sym_in_matches, p.ba.sym_in_literals, p.ba.no_components));
        }
        parallel_rows(n, host_threads, [&](size_t r) { std::sort(results[r].begin(), results[r].end()); });
        t_rows = secs(t_s1, clk::now());
        lzani_destroy(ctx);
        if (params.verbosity_level >= 2) cerr << endl;
        report();
        return;
    }

    // ---- submits of consecutive rows: bounded result buffers, and a progress tick between submits.  With
    // --out-alignment every pair also returns its region records (parser.get_parsing(), lz_matcher.cpp:226-227):
    // smaller submits keep that buffer bounded.
    const uint64_t max_pairs_per_call = (f_alignment ? (1ull << 20) : (16ull << 20)) * G;
    std::vector<uint32_t> rows, q_ids;
    std::vector<uint64_t> row_off;
    std::vector<lzani_result_t> out;
    std::vector<uint64_t> region_off;
    std::vector<lzani_region_t> regions;
    uint32_t r = 0;
    while (r < n) {
        rows.clear(); q_ids.clear(); row_off.assign(1, 0);
        uint64_t pairs = 0;
        while (r < n && (rows.empty() || pairs < max_pairs_per_call)) {
            rows.push_back(r);
            if (dense) pairs += n - 1;
            else {
                const size_t a = q_ids.size();
                for (auto id : filter.get_row(r)) q_ids.push_back((uint32_t)id);
                std::sort(q_ids.begin() + a, q_ids.end());
                pairs = q_ids.size();
            }
            row_off.push_back(pairs);
            ++r;
        }
        out.resize(pairs);
        const auto t_s0 = clk::now();
        if (pairs) {        // (a set of one sequence, or a filter nothing passes: nothing to submit, header-only output)
            if (f_alignment) {
                region_off.resize(pairs + 1);
                if (regions.size() < std::max<uint64_t>(1024, 4 * pairs)) regions.resize(std::max<uint64_t>(1024, 4 * pairs));
                for (;;) {
                    uint64_t need = 0;
                    const int rc = lzani_run_rows_regions(ctx, (uint32_t)rows.size(), rows.data(), dense ? nullptr : row_off.data(),
                                                          dense ? nullptr : q_ids.data(), out.data(), region_off.data(),
                                                          regions.data(), regions.size(), &need);
                    if (rc == LZANI_ERR_NOMEM && need > regions.size()) { regions.resize(need + need / 8); continue; }
                    if (rc != LZANI_OK) die("lzani_run_rows_regions", ctx);
                    break;
                }
            } else {
                const int rc = lzani_run_rows(ctx, (uint32_t)rows.size(), rows.data(), dense ? nullptr : row_off.data(),
                                              dense ? nullptr : q_ids.data(), out.data());
                if (rc != LZANI_OK) die("lzani_run_rows", ctx);
            }
        }
        const auto t_s1 = clk::now();
        t_submit += secs(t_s0, t_s1);
        auto fill_row = [&](size_t j) {
            const uint32_t ref = rows[j];
            vec_id_results_t res_row;
            const uint64_t a = row_off[j], b = row_off[j + 1];
            res_row.reserve(b - a);
            for (uint64_t k = a; k < b; ++k) {
                uint32_t q;
                if (dense) { q = (uint32_t)(k - a); if (q >= ref) ++q; }
                else q = q_ids[k];
                res_row.emplace_back(q, results_t(out[k].sym_in_matches, out[k].sym_in_literals, out[k].no_components));
            }
            std::sort(res_row.begin(), res_row.end());
            results[ref] = std::move(res_row);       // each row is written by exactly one thread
        };
        parallel_rows(rows.size(), host_threads, fill_row);
        if (f_alignment) {
            // the reference's own writer (lz_matcher.cpp:102-169) takes a mutex per call: one thread
            std::vector<region_t> parsing;
            for (size_t j = 0; j < rows.size(); ++j)
                for (uint64_t k = row_off[j]; k < row_off[j + 1]; ++k) {
                    uint32_t q;
                    if (dense) { q = (uint32_t)(k - row_off[j]); if (q >= rows[j]) ++q; }
                    else q = q_ids[k];
                    parsing.clear();
                    for (uint64_t e = region_off[k]; e < region_off[k + 1]; ++e) {
                        const lzani_region_t& g = regions[e];
                        parsing.emplace_back(g.ref_start, g.ref_end, g.seq_start, g.seq_end, g.num_matches, g.num_mismatches);
                    }
                    store_alignment(rows[j], q, parsing);
                }
        }
        if (!dense) for (uint32_t ref : rows) filter.clear_row(ref);
        t_rows += secs(t_s1, clk::now());
        if (params.verbosity_level >= 2)
            cerr << "Sequences: " + std::to_string(100ull * r / (n ? n : 1)) + "%       \r" << std::flush;
    }
    lzani_destroy(ctx);
    if (params.verbosity_level >= 2)
        cerr << endl;
    report();
}
