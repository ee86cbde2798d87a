// loader_hooks_b200.cpp -- the two rows of SURVEY 8f that sit one step OUTSIDE do_matching, reachable from the
// drop-in binary behind environment switches (default off: the reference's own code runs):
//
//   LZANI_DEVICE_INGEST=1    CSeqReservoir::append (reference src/seq_reservoir.cpp:18-87) keeps the loader's text --
//                            one byte per base, as it receives it -- instead of packing base-6 triples on the host;
//                            do_matching_b200.cpp then hands the text to lzani_set_sequences_text and the K1 kernel
//                            applies dna_code (seq_reservoir.h:243-247), builds the bit-plane records and the reverse
//                            complement on the device (8f-3).
//   LZANI_DEVICE_PAIRLIST=1  CFilter::reorder_items (reference src/filter.cpp:293-345) hands the filter's edges and
//                            the reordering map to lzani_build_pair_list: id remap, both directions, row sort and
//                            CSR on the device (8f-4); the rows it writes back are what the reference's own loop
//                            produces, ids ascending.
//
// Both are member functions of reference classes, swapped in the way do_matching is (integration/Makefile: the
// reference's definition is made weak and stays reachable under an alias, so the switch-off path IS the reference's
// code, not a restatement of it).
#include <cstdlib>
#include <cstring>
#include <iostream>

#include "seq_reservoir.h"       // reference headers (-I /root/reference/src)
#include "filter.h"
#include "lzani_b200.h"

// the reference's own definitions (aliases added by integration/Makefile with objcopy --add-symbol)
extern "C" void lzani_ref_seq_append(CSeqReservoir*, const std::string&, const std::string&);
extern "C" void lzani_ref_filter_reorder(CFilter*, const std::vector<uint32_t>&, uint32_t, uint32_t);

bool lzani_device_ingest()
{
    static const bool on = [] { const char* e = getenv("LZANI_DEVICE_INGEST"); return e && atoi(e) != 0; }();
    return on;
}
static bool device_pairlist()
{
    const char* e = getenv("LZANI_DEVICE_PAIRLIST");
    return e && atoi(e) != 0;
}

void CSeqReservoir::append(const string& name, const string& seq)
{
    if (!lzani_device_ingest()) {
        lzani_ref_seq_append(this, name, seq);
        return;
    }
    uint8_t* text = (uint8_t*)mma_seq.allocate(seq.length() ? seq.length() : 1);
    memcpy(text, seq.data(), seq.length());
    const size_t name_len = name.find(' ') == string::npos ? name.size() : name.find(' ');     // the id: up to the first space
    char* id = (char*)mma_name.allocate(name_len + 1);
    memcpy(id, name.data(), name_len);
    id[name_len] = 0;
    if (items.size() == items.capacity()) items.reserve(items.size() * 2);
    items.emplace_back(id, text, seq.length(), 1);
}

void CFilter::reorder_items(const vector<uint32_t>& reordering_map, uint32_t no_threads, uint32_t verbosity_level)
{
    if (!device_pairlist() || filter.empty()) {
        lzani_ref_filter_reorder(this, reordering_map, no_threads, verbosity_level);
        return;
    }
    if (verbosity_level >= 1) cerr << "Reordering filter (B200 pair list)" << endl;
    const uint32_t n = (uint32_t)filter.size();
    // the filter is symmetric by now (load_filter entered every edge in both rows): one direction is the edge list
    std::vector<uint32_t> ei, ej;
    for (uint32_t i = 0; i < n; ++i)
        for (auto j : filter[i])
            if ((uint32_t)j < i) { ei.push_back(i); ej.push_back((uint32_t)j); }
    lzani_params_t P{11, 7, 40, 40, 35, 15, 7, 3};       // the pair list does not depend on the parse parameters
    lzani_ctx* ctx = nullptr;
    int dev = 0;
    if (const char* e = getenv("LZANI_DEVICES")) dev = atoi(e);
    auto die = [&](const char* what) {
        cerr << "lzani-b200: " << what << ": " << (ctx ? lzani_last_error(ctx) : "no context") << endl;
        exit(1);
    };
    if (lzani_create(&ctx, &P, &dev, 1) != LZANI_OK) die("lzani_create");
    std::vector<uint64_t> row_off((size_t)n + 1, 0);
    std::vector<uint32_t> q_ids(2 * ei.size() + 1);
    uint64_t n_ent = 0;
    if (lzani_build_pair_list(ctx, n, ei.size(), ei.data(), ej.data(), reordering_map.data(), row_off.data(), q_ids.data(),
                              q_ids.size(), &n_ent) != LZANI_OK)
        die("lzani_build_pair_list");
    lzani_destroy(ctx);
    for (uint32_t r = 0; r < n; ++r) {
        filter[r].assign(q_ids.begin() + row_off[r], q_ids.begin() + row_off[r + 1]);
        filter[r].shrink_to_fit();
    }
}
