// capi.cu -- the C ABI of include/lzani_b200.h: context, HBM residency of the sequence set,
// reference-row batching under the index memory budget, launches, timing.
//
// There is no CPU fallback here: without a CUDA device every entry point returns
// LZANI_ERR_CUDA with a message.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "../../include/lzani_b200.h"
#include "kernels.h"

using namespace lzani;

struct lzani_ctx {
    Prm P{};
    // n_devices > 1 at lzani_create: this object is the GROUP -- one full single-device context per GPU in `subs`, driven by
    // worker threads the library spawns per call (the reference's own thread pool, lz_matcher.cpp:184-200, is inside
    // do_matching too); reference rows are dealt over the devices, sequences replicated, no collective.
    std::vector<lzani_ctx*> subs;
    int device = -1;
    int n_sm = 0;
    int blocks_per_sm = 1;
    cudaStream_t stream = nullptr;
    std::string err;

    // sequence set (resident)
    uint32_t n_seqs = 0;
    std::vector<SeqMeta> meta;
    uint64_t n_recs = 0;
    uint8_t* d_src = nullptr;
    SeqMeta* d_meta = nullptr;
    uint64_t* d_rec_base = nullptr;
    Rec* d_recs = nullptr;
    uint64_t src_cap = 0, meta_cap = 0, recs_cap = 0;     // capacities: buffers are re-used by the next set_sequences
    uint8_t* h_src = nullptr; uint64_t h_src_cap = 0;     // pinned staging of the reservoir bytes
    double pack_ms = 0;          // of the last set_sequences
    uint64_t seq_h2d_bytes = 0;

    // index pools for the current batch
    uint64_t budget = 0;
    uint32_t* d_bitmap = nullptr; uint64_t bitmap_words_cap = 0;
    uint64_t* d_table = nullptr;  uint64_t table_slots_cap = 0;

    // pair list + results
    uint32_t* d_row_ref = nullptr; RefIndex* d_row_idx = nullptr; uint64_t* d_row_off = nullptr;
    uint32_t* d_blk_base = nullptr; uint64_t rows_cap = 0;
    uint32_t* d_q_ids = nullptr; uint64_t q_cap = 0;
    void* d_out = nullptr; uint64_t out_cap = 0;
    unsigned long long* d_counter = nullptr;
    DevRegion* d_regs = nullptr; uint64_t regs_cap = 0;     // regions kernel (--out-alignment)
    unsigned long long* d_reg_counter = nullptr;

    // the last run, kept for lzani_filter_pairs (results stay in d_out until the next run)
    std::vector<uint32_t> last_row_ref;
    std::vector<uint64_t> last_row_off;
    bool last_valid = false, last_sparse = false, last_sorted = true;

    // out-filter compaction (postproc.cu)
    uint8_t* d_flags = nullptr; uint64_t flags_cap = 0;
    uint32_t* d_blk_cnt = nullptr; uint64_t blk_cnt_cap = 0;
    unsigned long long* d_blk_off = nullptr; uint64_t blk_off_cap = 0;
    RowLoc* d_loc = nullptr; uint64_t loc_cap = 0;
    uint32_t* d_eff_len = nullptr; uint64_t eff_cap = 0;
    PairOut* d_pairs = nullptr; uint64_t pairs_cap = 0;

    // staged parse (kernels.cu "K3 staged"): two chunk slots, each with its own stream, so that one chunk's
    // data-parallel stages fill the tail of the other chunk's walk
    struct Slot {
        StagedBufs B{};
        uint64_t pairs_cap = 0, words_cap = 0, recs_cap = 0;
        cudaStream_t st = nullptr;
        cudaEvent_t done = nullptr;
        cudaEvent_t dbg[4] = {nullptr, nullptr, nullptr, nullptr};   // LZANI_DEBUG_CHUNKS: scan start / scan end / walk start / walk end
        uint32_t dbg_chunk = 0; bool dbg_set = false;
        uint32_t* h_total = nullptr;      // pinned: records of the chunk
        uint32_t* mc = nullptr;           // all2all scan: bitmap-bit x reference matrix (2^bbits words)
        uint64_t mc_words = 0;
    } slot[4];
    int n_slots = 2;                      // chunks in flight (LZANI_SLOTS, 1..4): slot i's walk overlaps the stages and walks of the others
    int matrix_scan = 1;                  // 0: always the probe scan (LZANI_MATRIX_SCAN=0)
    int staged = 1;                       // 0: one-kernel parse (LZANI_STAGED=0)
    uint64_t staged_words = 0;            // words per chunk (LZANI_STAGED_WORDS; 0: sized from the free HBM at the first run)
    bool debug_chunks = false;            // LZANI_DEBUG_CHUNKS=1: one line per chunk on stderr
    bool density_seen = false;            // rec_per_word is an observation, not the initial guess
    uint64_t staged_grow = 0;             // words per chunk after the one-time growth for long genomes (0: not grown)
    bool staged_words_fixed = false;      // set by LZANI_STAGED_WORDS: chunks never grow beyond it
    uint64_t staged_recs = 0;             // record capacity per chunk (LZANI_STAGED_RECS; 0: 1.25 per word)
    int walk_blocks_per_sm = 1;
    int lane_walk = 1;                    // 0: warp-per-pair walk k_walk_pairs (LZANI_LANE_WALK=0)
    int lane_blocks_per_sm = 1;
    double rec_per_word = 1.0;            // observed hit runs per 32 query positions (chunk sizing)
    uint32_t staged_chunks = 0;           // of the last run

    cudaEvent_t ev[8]{};
    lzani_timing_t timing{};
};

namespace {

int fail(lzani_ctx* c, int code, const std::string& msg)
{
    if (c) c->err = msg;
    return code;
}
#define LZ_CUDA(c, call)                                                                         \
    do {                                                                                         \
        cudaError_t e_ = (call);                                                                 \
        if (e_ != cudaSuccess)                                                                   \
            return fail((c), e_ == cudaErrorMemoryAllocation ? LZANI_ERR_NOMEM : LZANI_ERR_CUDA, \
                        std::string(#call) + ": " + cudaGetErrorString(e_));                     \
    } while (0)

template <class T>
int ensure(lzani_ctx* c, T*& p, uint64_t& cap, uint64_t need)
{
    if (need <= cap && p) return LZANI_OK;
    if (p) { cudaFree(p); p = nullptr; cap = 0; }
    if (!need) need = 1;
    LZ_CUDA(c, cudaMalloc((void**)&p, need * sizeof(T)));
    cap = need;
    return LZANI_OK;
}

void free_sequences(lzani_ctx* c)
{
    cudaFree(c->d_src); c->d_src = nullptr;
    cudaFree(c->d_meta); c->d_meta = nullptr;
    cudaFree(c->d_rec_base); c->d_rec_base = nullptr;
    cudaFree(c->d_recs); c->d_recs = nullptr;
    if (c->h_src) { cudaFreeHost(c->h_src); c->h_src = nullptr; }
    c->src_cap = c->meta_cap = c->recs_cap = c->h_src_cap = 0;
    c->n_seqs = 0; c->n_recs = 0; c->meta.clear();
    c->last_valid = false;
    c->staged_grow = 0; c->density_seen = false; c->rec_per_word = 1.0;      // chunk sizing observations belong to a set
}

void free_slot(lzani_ctx::Slot& s)
{
    cudaFree(s.B.desc); cudaFree(s.B.n_words); cudaFree(s.B.word_off); cudaFree(s.B.hm); cudaFree(s.B.rec_off);
    cudaFree(s.B.list); cudaFree(s.B.recs); cudaFree(s.B.tmp); cudaFree(s.B.wpair); cudaFree(s.B.hard);
    s.B.hard = nullptr;
    s.B.desc = nullptr; s.B.n_words = nullptr; s.B.word_off = nullptr; s.B.hm = nullptr; s.B.rec_off = nullptr; s.B.wpair = nullptr;
    s.B.list = nullptr; s.B.recs = nullptr; s.B.tmp = nullptr; s.B.tmp_bytes = 0;
    s.pairs_cap = s.words_cap = s.recs_cap = 0;
}

// chunk scratch of one slot, grown on demand (the caller has synchronised the slot's stream)
int ensure_slot(lzani_ctx* c, lzani_ctx::Slot& s, uint64_t pairs, uint64_t words, uint64_t recs)
{
    if (pairs > s.pairs_cap) {
        cudaFree(s.B.desc); cudaFree(s.B.n_words); cudaFree(s.B.word_off); cudaFree(s.B.hard);
        s.B.desc = nullptr; s.B.n_words = nullptr; s.B.word_off = nullptr; s.B.hard = nullptr; s.pairs_cap = 0;
        LZ_CUDA(c, cudaMalloc((void**)&s.B.desc, sizeof(PairDesc) * pairs));
        LZ_CUDA(c, cudaMalloc((void**)&s.B.hard, sizeof(uint32_t) * pairs));
        LZ_CUDA(c, cudaMalloc((void**)&s.B.n_words, sizeof(uint32_t) * (pairs + 1)));
        LZ_CUDA(c, cudaMalloc((void**)&s.B.word_off, sizeof(uint32_t) * (pairs + 1)));
        s.pairs_cap = pairs;
    }
    if (words > s.words_cap) {
        cudaFree(s.B.hm); cudaFree(s.B.rec_off); cudaFree(s.B.wpair);
        s.B.hm = nullptr; s.B.rec_off = nullptr; s.B.wpair = nullptr; s.words_cap = 0;
        LZ_CUDA(c, cudaMalloc((void**)&s.B.wpair, sizeof(uint32_t) * (words + 1)));
        LZ_CUDA(c, cudaMalloc((void**)&s.B.hm, sizeof(uint32_t) * (words + 1)));
        LZ_CUDA(c, cudaMalloc((void**)&s.B.rec_off, sizeof(uint32_t) * (words + 1)));
        s.words_cap = words;
    }
    if (recs > s.recs_cap) {
        cudaFree(s.B.list); cudaFree(s.B.recs);
        s.B.list = nullptr; s.B.recs = nullptr; s.recs_cap = 0;
        LZ_CUDA(c, cudaMalloc((void**)&s.B.list, 8 * recs));
        LZ_CUDA(c, cudaMalloc((void**)&s.B.recs, sizeof(Rec) * (recs + 1)));
        s.recs_cap = recs;
    }
    const size_t tb = staged_scan_bytes(std::max(s.words_cap, s.pairs_cap) + 1);
    if (tb > s.B.tmp_bytes) {
        cudaFree(s.B.tmp); s.B.tmp = nullptr; s.B.tmp_bytes = 0;
        LZ_CUDA(c, cudaMalloc(&s.B.tmp, tb));
        s.B.tmp_bytes = tb;
    }
    return LZANI_OK;
}

float ms_between(cudaEvent_t a, cudaEvent_t b)
{
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    return ms;
}

}  // namespace

extern "C" int lzani_abi_version(void) { return LZANI_ABI_VERSION; }

extern "C" int lzani_create(lzani_ctx** out, const lzani_params_t* p, const int* device_ids, int n_devices)
{
    if (!out) return LZANI_ERR_ARG;
    *out = nullptr;
    lzani_ctx* c = new (std::nothrow) lzani_ctx();
    if (!c) return LZANI_ERR_NOMEM;
    *out = c;
    if (!p) return fail(c, LZANI_ERR_ARG, "params == NULL");
    c->P = Prm{p->mal, p->msl, p->mrd, p->mqd, p->reg, p->aw, p->am, p->ar};
    if (n_devices > 1) {
        if (!device_ids) return fail(c, LZANI_ERR_ARG, "device_ids == NULL with n_devices > 1");
        // (a device may be named twice: two contexts sharing one GPU -- how the group path is tested on a one-GPU box)
        // one thread per device: CUDA context creation (~1 s per GPU) runs side by side
        c->subs.assign((size_t)n_devices, nullptr);
        std::vector<int> rcs((size_t)n_devices, LZANI_OK);
        std::vector<std::thread> th;
        for (int g = 0; g < n_devices; ++g) th.emplace_back([&, g] { rcs[g] = lzani_create(&c->subs[g], p, device_ids + g, 1); });
        for (auto& t : th) t.join();
        int bad = -1;
        for (int g = 0; g < n_devices; ++g) if (rcs[g] != LZANI_OK && bad < 0) bad = g;
        if (bad >= 0) {
            const int rc = rcs[bad];
            const std::string msg = "device " + std::to_string(device_ids[bad]) + ": " + (c->subs[bad] ? c->subs[bad]->err : std::string("allocation failed"));
            for (lzani_ctx*& sub : c->subs) { if (sub) lzani_destroy(sub); sub = nullptr; }
            c->subs.clear();
            return fail(c, rc, msg);
        }
        return LZANI_OK;
    }
    const Prm& P = c->P;
    // what the kernels support (the reference has no explicit limits except mal <= 32, parser.cpp:59)
    if (P.mal < 1 || P.mal > 32) return fail(c, LZANI_ERR_ARG, "mal must be in [1,32]");
    if (P.msl < 1 || P.msl > 32) return fail(c, LZANI_ERR_ARG, "msl must be in [1,32]");
    if (P.mrd < 1 || P.mrd > 100000) return fail(c, LZANI_ERR_ARG, "mrd must be in [1,100000]");
    if (P.mqd < 0 || P.mqd > 1000) return fail(c, LZANI_ERR_ARG, "mqd must be in [0,1000]");
    if (P.aw < 1 || P.aw > 32) return fail(c, LZANI_ERR_ARG, "aw must be in [1,32]");
    if (P.ar < 1 || P.ar > 32) return fail(c, LZANI_ERR_ARG, "ar must be in [1,32]");
    if (P.am < 0) return fail(c, LZANI_ERR_ARG, "am must be >= 0");

    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0)
        return fail(c, LZANI_ERR_CUDA, std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "count is 0") +
                                           " (this library has no CPU fallback)");
    int dev = 0;
    if (device_ids && n_devices == 1) dev = device_ids[0];
    else LZ_CUDA(c, cudaGetDevice(&dev));
    if (dev < 0 || dev >= n_dev) return fail(c, LZANI_ERR_ARG, "device id out of range");
    LZ_CUDA(c, cudaSetDevice(dev));
    c->device = dev;
    cudaDeviceProp prop;
    LZ_CUDA(c, cudaGetDeviceProperties(&prop, dev));
    if (prop.major < 10) return fail(c, LZANI_ERR_CUDA, std::string("device ") + prop.name + " is not sm_100-class; the kernels are built for sm_100a only");
    c->n_sm = prop.multiProcessorCount;
    LZ_CUDA(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    for (auto& ev : c->ev) LZ_CUDA(c, cudaEventCreate(&ev));
    LZ_CUDA(c, cudaMalloc((void**)&c->d_counter, sizeof(unsigned long long)));
    LZ_CUDA(c, cudaMalloc((void**)&c->d_reg_counter, sizeof(unsigned long long)));
    c->blocks_per_sm = parse_blocks_per_sm(c->P);
    c->walk_blocks_per_sm = walk_blocks_per_sm(c->P);
    // the lane walk is bound by dependent-instruction latency, not by occupancy: three resident blocks per SM do as well
    // as five (measured, profiles/r02_walk_ab.txt) and leave registers for the other slot's data-parallel stages to run beside it
    c->lane_blocks_per_sm = std::min(lane_blocks_per_sm(c->P), 3);
    if (const char* e = getenv("LZANI_LANE_WALK")) c->lane_walk = atoi(e);
    if (const char* e = getenv("LZANI_LANE_BLOCKS")) { const int v = atoi(e); if (v > 0 && v <= lane_blocks_per_sm(c->P)) c->lane_blocks_per_sm = v; }
    if (const char* e = getenv("LZANI_STAGED")) c->staged = atoi(e);
    if (const char* e = getenv("LZANI_STAGED_RECS")) { const long long v = atoll(e); if (v > 0) c->staged_recs = (uint64_t)v; }
    if (const char* e = getenv("LZANI_MATRIX_SCAN")) c->matrix_scan = atoi(e);
    // fewer resident walk blocks leave registers for the other slot's data-parallel stages to run beside the walk
    if (const char* e = getenv("LZANI_WALK_BLOCKS")) { const int v = atoi(e); if (v > 0 && v < c->walk_blocks_per_sm) c->walk_blocks_per_sm = v; }
    if (const char* e = getenv("LZANI_DEBUG_CHUNKS")) c->debug_chunks = atoi(e) != 0;
    if (const char* e = getenv("LZANI_SLOTS")) { const int v = atoi(e); if (v >= 1 && v <= 4) c->n_slots = v; }
    if (const char* e = getenv("LZANI_STAGED_WORDS")) { const long long v = atoll(e); if (v > 0) { c->staged_words = (uint64_t)v; c->staged_words_fixed = true; } }
    for (auto& sl : c->slot) {
        LZ_CUDA(c, cudaStreamCreateWithFlags(&sl.st, cudaStreamNonBlocking));
        LZ_CUDA(c, cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
        for (auto& e : sl.dbg) LZ_CUDA(c, cudaEventCreate(&e));
        LZ_CUDA(c, cudaMallocHost((void**)&sl.h_total, 4 * sizeof(uint32_t)));
        LZ_CUDA(c, cudaMalloc((void**)&sl.B.counter, 6 * sizeof(unsigned long long)));
    }
    return LZANI_OK;
}

extern "C" void lzani_destroy(lzani_ctx* c)
{
    if (!c) return;
    for (lzani_ctx* sub : c->subs) lzani_destroy(sub);
    c->subs.clear();
    if (c->device >= 0) {
        cudaSetDevice(c->device);
        free_sequences(c);
        cudaFree(c->d_bitmap); cudaFree(c->d_table);
        cudaFree(c->d_row_ref); cudaFree(c->d_row_idx); cudaFree(c->d_row_off); cudaFree(c->d_blk_base);
        cudaFree(c->d_q_ids); cudaFree(c->d_out); cudaFree(c->d_counter);
        cudaFree(c->d_regs); cudaFree(c->d_reg_counter);
        cudaFree(c->d_flags); cudaFree(c->d_blk_cnt); cudaFree(c->d_blk_off); cudaFree(c->d_loc);
        cudaFree(c->d_eff_len); cudaFree(c->d_pairs);
        for (auto& sl : c->slot) {
            free_slot(sl);
            cudaFree(sl.mc);
            cudaFree(sl.B.counter);
            if (sl.h_total) cudaFreeHost(sl.h_total);
            if (sl.done) cudaEventDestroy(sl.done);
            for (auto& e : sl.dbg) if (e) cudaEventDestroy(e);
            if (sl.st) cudaStreamDestroy(sl.st);
        }
        for (auto& ev : c->ev) if (ev) cudaEventDestroy(ev);
        if (c->stream) cudaStreamDestroy(c->stream);
    }
    delete c;
}

extern "C" const char* lzani_last_error(const lzani_ctx* c) { return c ? c->err.c_str() : "null context"; }

extern "C" int lzani_set_index_budget(lzani_ctx* c, uint64_t bytes)
{
    if (!c) return LZANI_ERR_ARG;
    c->budget = bytes;
    for (lzani_ctx* sub : c->subs) sub->budget = bytes;       // per device
    return LZANI_OK;
}

extern "C" int lzani_get_timing(const lzani_ctx* c, lzani_timing_t* t)
{
    if (!c || !t) return LZANI_ERR_ARG;
    *t = c->timing;
    return LZANI_OK;
}

namespace {
int set_sequences_impl(lzani_ctx* c, uint32_t n, const uint8_t* const* data, const uint32_t* len, bool ascii);

// ---- group plumbing: f(g, sub) on every device of the group, one host thread each; the first error wins
template <class F>
int group_each(lzani_ctx* c, F f)
{
    const size_t G = c->subs.size();
    std::vector<int> rc(G, LZANI_OK);
    std::vector<std::thread> th;
    for (size_t g = 1; g < G; ++g) th.emplace_back([&, g] { rc[g] = f((uint32_t)g, c->subs[g]); });
    rc[0] = f(0u, c->subs[0]);
    for (auto& t : th) t.join();
    for (size_t g = 0; g < G; ++g)
        if (rc[g] != LZANI_OK) return fail(c, rc[g], "device " + std::to_string(c->subs[g]->device) + ": " + c->subs[g]->err);
    return LZANI_OK;
}
// timings of a group call: the devices run side by side -- times are the slowest device's, counts add up
void group_timing(lzani_ctx* c)
{
    lzani_timing_t T{};
    for (const lzani_ctx* s : c->subs) {
        const lzani_timing_t& t = s->timing;
        T.h2d_ms = std::max(T.h2d_ms, t.h2d_ms); T.pack_ms = std::max(T.pack_ms, t.pack_ms);
        T.index_ms = std::max(T.index_ms, t.index_ms); T.parse_ms = std::max(T.parse_ms, t.parse_ms);
        T.d2h_ms = std::max(T.d2h_ms, t.d2h_ms); T.total_ms = std::max(T.total_ms, t.total_ms);
        T.filter_ms = std::max(T.filter_ms, t.filter_ms); T.pairlist_ms = std::max(T.pairlist_ms, t.pairlist_ms);
        T.n_pairs += t.n_pairs; T.q_bases += t.q_bases; T.r_bases += t.r_bases;
        T.h2d_bytes += t.h2d_bytes; T.d2h_bytes += t.d2h_bytes; T.filter_d2h_bytes += t.filter_d2h_bytes;
        T.launches += t.launches; T.n_batches = std::max(T.n_batches, t.n_batches);
    }
    c->timing = T;
}
int group_set_sequences(lzani_ctx* c, uint32_t n, const uint8_t* const* data, const uint32_t* len, bool ascii)
{
    const int rc = group_each(c, [&](uint32_t, lzani_ctx* s) { return set_sequences_impl(s, n, data, len, ascii); });
    c->n_seqs = rc == LZANI_OK ? n : 0;
    group_timing(c);
    return rc;
}
}  // namespace

extern "C" int lzani_set_sequences(lzani_ctx* c, uint32_t n, const uint8_t* const* data, const uint32_t* len)
{
    if (c && !c->subs.empty()) return group_set_sequences(c, n, data, len, false);
    return set_sequences_impl(c, n, data, len, false);
}

extern "C" int lzani_set_sequences_text(lzani_ctx* c, uint32_t n, const uint8_t* const* text, const uint32_t* len)
{
    if (c && !c->subs.empty()) return group_set_sequences(c, n, text, len, true);
    return set_sequences_impl(c, n, text, len, true);
}

namespace {
// bytes of one sequence in the staging buffer: base-6 triples, or one byte per base (16-byte aligned starts)
inline uint64_t src_bytes_of(uint32_t len, bool ascii) { return ascii ? (((uint64_t)len + 15) & ~15ull) : ((uint64_t)len + 2) / 3; }

int set_sequences_impl(lzani_ctx* c, uint32_t n, const uint8_t* const* data, const uint32_t* len, bool ascii)
{
    if (!c) return LZANI_ERR_ARG;
    if (c->device < 0) return fail(c, LZANI_ERR_CUDA, "context has no device (lzani_create failed)");
    if (n && (!data || !len)) return fail(c, LZANI_ERR_ARG, "data/len == NULL");
    LZ_CUDA(c, cudaSetDevice(c->device));
    // the previous set is gone from here on (its buffers are re-used when large enough)
    c->n_seqs = 0; c->n_recs = 0; c->meta.clear();
    c->last_valid = false;
    const Prm& P = c->P;
    std::vector<SeqMeta> meta(n);
    std::vector<uint64_t> rec_base((size_t)n + 1, 0);
    uint64_t src_bytes = 0, recs = 0;
    for (uint32_t i = 0; i < n; ++i) {
        // positions are 32-bit ints in the reference (combined reference 2L+3mrd < 2^31)
        if ((uint64_t)len[i] * 2 + 3ull * (uint64_t)P.mrd >= 0x7FFFFF00ull)
            return fail(c, LZANI_ERR_ARG, "sequence too long for 32-bit positions");
        if (len[i] && !data[i]) return fail(c, LZANI_ERR_ARG, "data[i] == NULL");
        meta[i].src_off = src_bytes;
        meta[i].len = len[i];
        meta[i].pad_ = 0;
        meta[i].q_rec = recs;
        meta[i].r_rec = recs + (uint64_t)q_form_recs((int)len[i], P.mrd);
        rec_base[i] = recs;
        recs = meta[i].r_rec + (uint64_t)r_form_recs((int)len[i], P.mrd);
        src_bytes += src_bytes_of(len[i], ascii);
    }
    rec_base[n] = recs;

    // stage the reservoir bytes contiguously in pinned memory, one H2D copy
    if (src_bytes > c->h_src_cap || !c->h_src) {
        if (c->h_src) { cudaFreeHost(c->h_src); c->h_src = nullptr; c->h_src_cap = 0; }
        LZ_CUDA(c, cudaMallocHost((void**)&c->h_src, src_bytes ? src_bytes : 1));
        c->h_src_cap = src_bytes;
    }
    uint8_t* h_src = c->h_src;
    for (uint32_t i = 0; i < n; ++i)
        if (len[i]) memcpy(h_src + meta[i].src_off, data[i], ascii ? (size_t)len[i] : ((size_t)len[i] + 2) / 3);

    int rc = LZANI_OK;
    do {
        cudaError_t e;
#define LZ_TRY(call) if ((e = (call)) != cudaSuccess) { rc = fail(c, e == cudaErrorMemoryAllocation ? LZANI_ERR_NOMEM : LZANI_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e)); break; }
        if (!c->d_src || src_bytes > c->src_cap) {
            cudaFree(c->d_src); c->d_src = nullptr; c->src_cap = 0;
            LZ_TRY(cudaMalloc((void**)&c->d_src, src_bytes ? src_bytes : 1));
            c->src_cap = src_bytes;
        }
        if (!c->d_meta || !c->d_rec_base || n > c->meta_cap) {
            cudaFree(c->d_meta); cudaFree(c->d_rec_base); c->d_meta = nullptr; c->d_rec_base = nullptr; c->meta_cap = 0;
            LZ_TRY(cudaMalloc((void**)&c->d_meta, sizeof(SeqMeta) * (n ? n : 1)));
            LZ_TRY(cudaMalloc((void**)&c->d_rec_base, sizeof(uint64_t) * ((size_t)n + 1)));
            c->meta_cap = n;
        }
        if (!c->d_recs || recs > c->recs_cap) {
            cudaFree(c->d_recs); c->d_recs = nullptr; c->recs_cap = 0;
            LZ_TRY(cudaMalloc((void**)&c->d_recs, sizeof(Rec) * (recs ? recs : 1)));
            c->recs_cap = recs;
        }
        LZ_TRY(cudaEventRecord(c->ev[0], c->stream));
        LZ_TRY(cudaMemcpyAsync(c->d_src, h_src, src_bytes, cudaMemcpyHostToDevice, c->stream));
        LZ_TRY(cudaMemcpyAsync(c->d_meta, meta.data(), sizeof(SeqMeta) * n, cudaMemcpyHostToDevice, c->stream));
        LZ_TRY(cudaMemcpyAsync(c->d_rec_base, rec_base.data(), sizeof(uint64_t) * ((size_t)n + 1), cudaMemcpyHostToDevice, c->stream));
        LZ_TRY(cudaEventRecord(c->ev[1], c->stream));
        LZ_TRY(launch_pack(c->stream, c->d_src, c->d_meta, c->d_rec_base, n, recs, P.mrd, c->d_recs, ascii));
        LZ_TRY(cudaEventRecord(c->ev[2], c->stream));
        LZ_TRY(cudaStreamSynchronize(c->stream));
#undef LZ_TRY
    } while (0);
    if (rc != LZANI_OK) { free_sequences(c); return rc; }
    c->n_seqs = n;
    c->n_recs = recs;
    c->meta.swap(meta);
    c->pack_ms = ms_between(c->ev[1], c->ev[2]);
    c->seq_h2d_bytes = src_bytes + sizeof(SeqMeta) * n + sizeof(uint64_t) * ((size_t)n + 1);
    c->timing = lzani_timing_t{};
    c->timing.h2d_ms = ms_between(c->ev[0], c->ev[1]);
    c->timing.pack_ms = c->pack_ms;
    c->timing.h2d_bytes = c->seq_h2d_bytes;
    c->timing.launches = recs ? 1 : 0;
    return LZANI_OK;
}
}  // namespace

namespace {

struct RowPlan {
    uint64_t bitmap_words, table_slots;   // 0 for rows without pairs
    uint32_t bbits, tbits, blocks;
};

// The whole hot path for a set of reference rows.  `repeat` > 1 re-runs index build + parse
// with everything resident (bench "value" leg).
struct RegionReq {          // lzani_run_rows_regions: where the regions go
    uint64_t* region_off;
    lzani_region_t* regions;
    uint64_t cap;
    uint64_t* n_regions;
};

int run_rows_impl(lzani_ctx* c, uint32_t n_rows, const uint32_t* ref_ids, const uint64_t* row_off_in,
                  const uint32_t* q_ids, int repeat, lzani_result_t* out, bool copy_out, const RegionReq* rq = nullptr)
{
    if (!c) return LZANI_ERR_ARG;
    if (c->device < 0) return fail(c, LZANI_ERR_CUDA, "context has no device (lzani_create failed)");
    if (!c->d_recs) return fail(c, LZANI_ERR_STATE, "lzani_set_sequences has not been called");
    if (q_ids && !row_off_in) return fail(c, LZANI_ERR_ARG, "q_ids given without row_off");
    if (repeat < 1) repeat = 1;
    LZ_CUDA(c, cudaSetDevice(c->device));
    const Prm& P = c->P;
    const uint32_t n = c->n_seqs;

    // ---- row table on the host
    std::vector<uint32_t> row_ref(n_rows);
    std::vector<uint64_t> row_off((size_t)n_rows + 1, 0);
    uint64_t q_bases = 0, r_bases = 0, total_len = 0;
    for (uint32_t i = 0; i < n; ++i) total_len += c->meta[i].len;
    for (uint32_t j = 0; j < n_rows; ++j) {
        const uint32_t r = ref_ids ? ref_ids[j] : j;
        if (r >= n) return fail(c, LZANI_ERR_ARG, "reference id out of range");
        row_ref[j] = r;
        if (q_ids) {
            if (row_off_in[j + 1] < row_off_in[j]) return fail(c, LZANI_ERR_ARG, "row_off is not non-decreasing");
            row_off[j + 1] = row_off_in[j + 1] - row_off_in[0];
        } else {
            row_off[j + 1] = row_off[j] + (n ? n - 1 : 0);
            q_bases += total_len - c->meta[r].len;
        }
    }
    const uint64_t n_pairs = row_off[n_rows];
    // (a submit without pairs -- a single sequence, a filter nothing passes, a GPU whose rows are empty -- is legal and
    // returns nothing: the reference writes header-only output then)
    if (copy_out && !out && n_pairs) return fail(c, LZANI_ERR_ARG, "out == NULL");
    const uint32_t* q_src = q_ids ? q_ids + row_off_in[0] : nullptr;
    c->last_valid = false;
    bool rows_sorted = true;
    if (q_ids) {
        for (uint64_t k = 0; k < n_pairs; ++k) {
            if (q_src[k] >= n) return fail(c, LZANI_ERR_ARG, "query id out of range");
            q_bases += c->meta[q_src[k]].len;
        }
        for (uint32_t j = 0; j < n_rows && rows_sorted; ++j)
            for (uint64_t k = row_off[j] + 1; k < row_off[j + 1]; ++k)
                if (q_src[k] <= q_src[k - 1]) { rows_sorted = false; break; }
    }
    std::vector<RowPlan> plan(n_rows);
    for (uint32_t j = 0; j < n_rows; ++j) {
        RowPlan& rp = plan[j];
        rp = RowPlan{0, 0, 0, 0, 0};
        if (row_off[j + 1] == row_off[j]) continue;
        const int Dr = r_form_len((int)c->meta[row_ref[j]].len, P.mrd);
        rp.bbits = index_bbits(Dr, P.mal);
        rp.tbits = index_tbits(Dr);
        rp.bitmap_words = std::max<uint64_t>(1, ((uint64_t)1 << rp.bbits) / 32);
        rp.table_slots = (uint64_t)1 << rp.tbits;
        rp.blocks = (uint32_t)((Dr + 255) / 256);
        r_bases += (uint64_t)Dr;
    }

    // ---- device buffers for the pair list and the results
    if (c->rows_cap < (uint64_t)n_rows + 1) {
        cudaFree(c->d_row_ref); cudaFree(c->d_row_idx); cudaFree(c->d_row_off); cudaFree(c->d_blk_base);
        c->d_row_ref = nullptr; c->d_row_idx = nullptr; c->d_row_off = nullptr; c->d_blk_base = nullptr;
        c->rows_cap = 0;
        const uint64_t need = (uint64_t)n_rows + 1;
        LZ_CUDA(c, cudaMalloc((void**)&c->d_row_ref, need * sizeof(uint32_t)));
        LZ_CUDA(c, cudaMalloc((void**)&c->d_row_idx, need * sizeof(RefIndex)));
        LZ_CUDA(c, cudaMalloc((void**)&c->d_row_off, need * sizeof(uint64_t)));
        LZ_CUDA(c, cudaMalloc((void**)&c->d_blk_base, need * sizeof(uint32_t)));
        c->rows_cap = need;
    }
    if (q_ids) if (int rc = ensure(c, c->d_q_ids, c->q_cap, n_pairs)) return rc;
    {
        uint64_t cap_bytes = c->out_cap;
        uint8_t* p = (uint8_t*)c->d_out;
        if (int rc = ensure(c, p, cap_bytes, n_pairs * sizeof(lzani_result_t))) return rc;
        c->d_out = p; c->out_cap = cap_bytes;
    }

    if (rq) {
        if (!rq->n_regions || (rq->cap && (!rq->regions || !rq->region_off))) return fail(c, LZANI_ERR_ARG, "region buffers == NULL");
        if (int rc = ensure(c, c->d_regs, c->regs_cap, rq->cap ? rq->cap : 1)) return rc;
        LZ_CUDA(c, cudaMemsetAsync(c->d_reg_counter, 0, sizeof(unsigned long long), c->stream));
    }

    // ---- staged parse: words (32 query positions) per sequence as a query, chunk scratch.  Allocated
    // before the index budget is taken from what is left.
    std::vector<uint32_t> wq;
    uint64_t words_all = 0, words_run = 0;
    uint32_t words_max = 0;
    bool staged = !rq && c->staged != 0 && n_pairs > 0;
    if (staged) {
        wq.resize(n);
        for (uint32_t i = 0; i < n; ++i) {
            const int64_t n_loop = (int64_t)c->meta[i].len + P.mrd - P.msl;
            if ((int64_t)c->meta[i].len + P.mrd >= (int64_t)LZ_STAGED_MAX_POS) staged = false;   // hit-list entries hold 29-bit positions
            wq[i] = n_loop > 0 ? (uint32_t)((n_loop + 31) / 32) : 0u;
            words_all += wq[i];
            words_max = std::max(words_max, wq[i]);
        }
    }
    if (staged && !c->staged_words) {
        // chunk size: the slots together may take a fifth of the free HBM (~40 B of scratch per query word),
        // between 96 M and 768 M words -- large chunks keep many rows together (the all2all scan handles 32
        // references per gather) and the walk's tails rare; buffers still only grow to what a run needs
        size_t fr = 0, tot = 0;
        LZ_CUDA(c, cudaMemGetInfo(&fr, &tot));
        uint64_t wds = (uint64_t)((double)fr * 0.2 / ((q_ids ? 60.0 : 40.0) * c->n_slots));     // sparse lists: twice the records (below)
        wds = std::max<uint64_t>(96ull << 20, std::min<uint64_t>(wds, 768ull << 20));
        c->staged_words = wds;
    }
    if (staged) {
        if (q_ids) for (uint64_t k = 0; k < n_pairs; ++k) words_run += wq[q_src[k]];
        else for (uint32_t j = 0; j < n_rows; ++j) words_run += words_all - wq[row_ref[j]];
        const uint64_t wcap = std::max<uint64_t>(std::min<uint64_t>(c->staged_words, words_run), std::max<uint32_t>(words_max, 1u));
        const uint64_t pcap = std::min<uint64_t>(n_pairs, (uint64_t)1 << 19);
        // 1.25 hit runs per query word hold any mix of mostly unrelated pairs (0.5 at the bench config).  A sparse pair list
        // comes from a similarity filter: its pairs are related ones (1.2-1.3 runs per word for phages at >= 70 % identity,
        // up to 4 for near-identical genomes) -- twice the room, so that the first chunk does not overflow and get re-scanned
        uint64_t rcap = (q_ids ? 2 * wcap : wcap + wcap / 4) + 1024;
        if (c->staged_recs) rcap = c->staged_recs;       // test knob: forces the chunk-halving / growth paths
        const int n_slots = words_run > wcap ? c->n_slots : 1;
        for (int i = 0; i < n_slots; ++i) {
            lzani_ctx::Slot& sl = c->slot[i];
            if (pcap > sl.pairs_cap || wcap > sl.words_cap || rcap > sl.recs_cap)
                if (int rc = ensure_slot(c, sl, std::max(pcap, sl.pairs_cap), std::max(wcap, sl.words_cap), std::max(rcap, sl.recs_cap))) return rc;
        }
    }

    // ---- batches of consecutive rows under the index budget
    uint64_t budget = c->budget;
    if (!budget) {
        size_t fr = 0, tot = 0;
        LZ_CUDA(c, cudaMemGetInfo(&fr, &tot));
        fr += c->bitmap_words_cap * 4 + c->table_slots_cap * 8;   // pools we may re-use
        budget = (uint64_t)(0.6 * (double)fr);
    }
    struct Batch { uint32_t a, b; uint64_t words, slots; };
    std::vector<Batch> batches;
    {
        Batch cur{0, 0, 0, 0};
        for (uint32_t j = 0; j < n_rows; ++j) {
            const uint64_t bytes = plan[j].bitmap_words * 4 + plan[j].table_slots * 8;
            if (bytes > budget) return fail(c, LZANI_ERR_NOMEM, "the index of one reference exceeds the index budget");
            if ((cur.words + plan[j].bitmap_words) * 4 + (cur.slots + plan[j].table_slots) * 8 > budget && cur.b > cur.a) {
                batches.push_back(cur);
                cur = Batch{j, j, 0, 0};
            }
            cur.b = j + 1;
            cur.words += plan[j].bitmap_words;
            cur.slots += plan[j].table_slots;
        }
        if (cur.b > cur.a) batches.push_back(cur);
    }
    uint64_t max_words = 1, max_slots = 1;
    for (const Batch& b : batches) { max_words = std::max(max_words, b.words); max_slots = std::max(max_slots, b.slots); }
    if (int rc = ensure(c, c->d_bitmap, c->bitmap_words_cap, max_words)) return rc;
    if (int rc = ensure(c, c->d_table, c->table_slots_cap, max_slots)) return rc;

    lzani_timing_t T{};
    T.n_batches = (uint32_t)batches.size();
    c->staged_chunks = 0;
    cudaStream_t st = c->stream;

    // ---- H2D: the pair list
    LZ_CUDA(c, cudaEventRecord(c->ev[0], st));
    LZ_CUDA(c, cudaMemcpyAsync(c->d_row_ref, row_ref.data(), sizeof(uint32_t) * n_rows, cudaMemcpyHostToDevice, st));
    LZ_CUDA(c, cudaMemcpyAsync(c->d_row_off, row_off.data(), sizeof(uint64_t) * ((size_t)n_rows + 1), cudaMemcpyHostToDevice, st));
    T.h2d_bytes += sizeof(uint32_t) * n_rows + sizeof(uint64_t) * ((size_t)n_rows + 1);
    if (q_ids && n_pairs) {
        LZ_CUDA(c, cudaMemcpyAsync(c->d_q_ids, q_src, sizeof(uint32_t) * n_pairs, cudaMemcpyHostToDevice, st));
        T.h2d_bytes += sizeof(uint32_t) * n_pairs;
    }
    LZ_CUDA(c, cudaEventRecord(c->ev[1], st));

    const int grid = c->n_sm * c->blocks_per_sm;
    std::vector<RefIndex> h_idx;
    std::vector<uint32_t> h_blk;
    for (int rep = 0; rep < repeat; ++rep) {
        for (const Batch& b : batches) {
            const uint32_t nb = b.b - b.a;
            h_idx.assign(nb, RefIndex{0, 0, 0, 0});
            h_blk.assign((size_t)nb + 1, 0);
            uint64_t w = 0, s = 0;
            for (uint32_t j = 0; j < nb; ++j) {
                const RowPlan& rp = plan[b.a + j];
                h_idx[j] = RefIndex{w, s, rp.bbits, rp.tbits};
                w += rp.bitmap_words;
                s += rp.table_slots;
                h_blk[j + 1] = h_blk[j] + rp.blocks;
            }
            LZ_CUDA(c, cudaEventRecord(c->ev[2], st));
            LZ_CUDA(c, cudaMemcpyAsync(c->d_row_idx + b.a, h_idx.data(), sizeof(RefIndex) * nb, cudaMemcpyHostToDevice, st));
            LZ_CUDA(c, cudaMemcpyAsync(c->d_blk_base, h_blk.data(), sizeof(uint32_t) * ((size_t)nb + 1), cudaMemcpyHostToDevice, st));
            T.h2d_bytes += sizeof(RefIndex) * nb + sizeof(uint32_t) * ((size_t)nb + 1);
            LZ_CUDA(c, cudaEventRecord(c->ev[3], st));
            if (b.words) LZ_CUDA(c, cudaMemsetAsync(c->d_bitmap, 0, b.words * 4, st));
            if (b.slots) LZ_CUDA(c, cudaMemsetAsync(c->d_table, 0xFF, b.slots * 8, st));
            LZ_CUDA(c, launch_build_index(st, c->d_recs, c->d_meta, c->d_row_ref + b.a, c->d_row_idx + b.a, c->d_blk_base,
                                          nb, h_blk[nb], P, c->d_bitmap, c->d_table));
            LZ_CUDA(c, cudaEventRecord(c->ev[4], st));
            LZ_CUDA(c, cudaMemsetAsync(c->d_counter, 0, sizeof(unsigned long long), st));
            if (rq)
                LZ_CUDA(c, launch_parse_regions(st, grid, c->d_recs, c->d_meta, c->d_row_ref + b.a, c->d_row_idx + b.a,
                                                c->d_row_off + b.a, nb, q_ids ? c->d_q_ids : nullptr, row_off[b.a],
                                                row_off[b.b], P, c->d_bitmap, c->d_table, c->d_counter, c->d_out,
                                                c->d_regs, c->d_reg_counter, rq->cap));
            else if (staged) {
                // ---- chunks of consecutive pairs through the staged pipeline, alternating between the two slots
                const uint64_t kb = row_off[b.a], ke = row_off[b.b];
                for (auto& sl : c->slot) LZ_CUDA(c, cudaStreamWaitEvent(sl.st, c->ev[4], 0));
                const int wgrid = c->n_sm * c->walk_blocks_per_sm;
                auto pair_q = [&](uint32_t j, uint64_t k) -> uint32_t {
                    if (q_ids) return q_src[k];
                    uint32_t q = (uint32_t)(k - row_off[j]);
                    if (q >= row_ref[j]) ++q;
                    return q;
                };
                auto range_words = [&](uint64_t k0, uint64_t k1) -> uint64_t {
                    uint64_t wsum = 0;
                    uint32_t j = (uint32_t)(std::upper_bound(row_off.begin(), row_off.end(), k0) - row_off.begin()) - 1;
                    for (uint64_t k = k0; k < k1; ++k) {
                        while (row_off[j + 1] <= k) ++j;
                        wsum += wq[pair_q(j, k)];
                    }
                    return wsum;
                };
                struct Range { uint64_t k0, k1, words; };
                std::vector<Range> redo;             // halves of a chunk whose records did not fit
                uint64_t k_next = kb;
                uint32_t j_cur = b.a;
                while (k_next < ke || !redo.empty()) {
                    Range ch{0, 0, 0};
                    lzani_ctx::Slot& sl = c->slot[c->staged_chunks % (uint32_t)c->n_slots];
                    if (!redo.empty()) { ch = redo.back(); redo.pop_back(); }
                    else {
                        // the next chunk: as many pairs as the word / record / pair capacities allow
                        uint64_t w_target = c->staged_words;
                        const uint64_t have_r = sl.recs_cap ? sl.recs_cap : c->slot[0].recs_cap;
                        const uint64_t by_recs = (uint64_t)((double)have_r / (1.15 * c->rec_per_word));
                        // long genomes: a chunk of the configured size holds few pairs (about one wave of the walk) and few
                        // rows (the all2all scan costs the same for 3 rows as for 32).  Once the hit density has been
                        // observed, such chunks grow to what the record buffers hold at that density -- only the 12 B
                        // per word of hit masks / offsets grow with them -- up to twice the configured words.
                        // Decided ONCE (growing a buffer means cudaFree, which waits for the other slot's walk).
                        if (!c->staged_grow && c->density_seen && !c->staged_words_fixed &&
                            words_run / std::max<uint64_t>(n_pairs, 1) * 32768ull > c->staged_words) {
                            c->staged_grow = std::min<uint64_t>(std::min<uint64_t>(2 * c->staged_words, 0xE0000000ull), by_recs);
                            if (c->staged_grow < c->staged_words) c->staged_grow = c->staged_words;
                        }
                        if (c->staged_grow) w_target = c->staged_grow;
                        if (by_recs < w_target) w_target = by_recs;
                        const uint64_t p_max = (uint64_t)1 << 19;
                        uint64_t k = k_next, wsum = 0;
                        while (row_off[j_cur + 1] <= k) ++j_cur;
                        const uint32_t j_first = j_cur;
                        while (k < ke && k - k_next < p_max) {
                            while (row_off[j_cur + 1] <= k) ++j_cur;
                            if (!q_ids && k == row_off[j_cur]) {          // a whole dense row at once
                                const uint64_t rw = words_all - wq[row_ref[j_cur]], np = row_off[j_cur + 1] - k;
                                if (wsum + rw <= w_target && k - k_next + np <= p_max) { wsum += rw; k += np; continue; }
                                if (k > k_next && k_next == row_off[j_first]) break;   // keep the chunk made of whole rows
                            }
                            const uint32_t w1 = wq[pair_q(j_cur, k)];
                            if (k > k_next && wsum + w1 > w_target) break;
                            wsum += w1;
                            ++k;
                        }
                        ch = Range{k_next, k, wsum};
                        k_next = k;
                    }
                    const uint64_t np = ch.k1 - ch.k0;
                    if (np > sl.pairs_cap || ch.words > sl.words_cap || !sl.recs_cap) {
                        LZ_CUDA(c, cudaStreamSynchronize(sl.st));
                        if (int rc = ensure_slot(c, sl, std::max(np, std::max(sl.pairs_cap, c->slot[0].pairs_cap)),
                                                 std::max(std::max(ch.words, c->staged_grow), std::max(sl.words_cap, c->slot[0].words_cap)),
                                                 std::max(sl.recs_cap, c->slot[0].recs_cap))) return rc;
                    }
                    auto dbg_report = [&](lzani_ctx::Slot& s_) {
                        if (!c->debug_chunks || !s_.dbg_set) return;
                        cudaEventSynchronize(s_.dbg[3]);
                        float t0 = 0, t1 = 0, t2 = 0, t3 = 0;
                        cudaEventElapsedTime(&t0, c->ev[4], s_.dbg[0]); cudaEventElapsedTime(&t1, c->ev[4], s_.dbg[1]);
                        cudaEventElapsedTime(&t2, c->ev[4], s_.dbg[2]); cudaEventElapsedTime(&t3, c->ev[4], s_.dbg[3]);
                        fprintf(stderr, "[lzani] chunk %u timeline (ms since the batch's index): scan %.1f-%.1f  records+walk %.1f-%.1f\n",
                                s_.dbg_chunk, t0, t1, t2, t3);
                        s_.dbg_set = false;
                    };
                    dbg_report(sl);
                    if (c->debug_chunks) LZ_CUDA(c, cudaEventRecord(sl.dbg[0], sl.st));
                    // all2all form of the scan when the chunk is made of whole dense rows and the bitmaps are direct
                    MatrixScan mx{};
                    bool use_mx = false;
                    if (!q_ids && c->matrix_scan && P.mal <= 32 && n <= 65535u && np) {
                        const uint32_t ja = (uint32_t)(std::upper_bound(row_off.begin(), row_off.end(), ch.k0) - row_off.begin()) - 1;
                        const uint32_t jb = (uint32_t)(std::upper_bound(row_off.begin(), row_off.end(), ch.k1 - 1) - row_off.begin());
                        // one bitmap size for every row of the chunk (direct bitmaps: always; hashed ones: references of
                        // similar length), at most 2^30 bits = a 4 GB matrix
                        uint32_t bb = plan[ja].bbits;
                        for (uint32_t j = ja; j < jb; ++j) if (plan[j].bbits != bb) bb = 0;
                        if (row_off[ja] == ch.k0 && row_off[jb] == ch.k1 && bb && bb <= 30u) {
                            // 2^bbits words, written 32 at a time (k_bitmap_matrix): never fewer than 32
                            const uint64_t mw = std::max<uint64_t>(32, (uint64_t)1 << bb);
                            if (mw > sl.mc_words) {
                                LZ_CUDA(c, cudaStreamSynchronize(sl.st));
                                cudaFree(sl.mc); sl.mc = nullptr; sl.mc_words = 0;
                                // no room for the matrix (up to 4 GB): not an error, the probe scan does the chunk
                                if (cudaMalloc((void**)&sl.mc, sizeof(uint32_t) * mw) == cudaSuccess) sl.mc_words = mw;
                                else { sl.mc = nullptr; (void)cudaGetLastError(); }
                            }
                            if (sl.mc) {
                                mx.mc = sl.mc; mx.bbits = bb; mx.row_first = ja - b.a; mx.n_rows_chunk = jb - ja; mx.n_seqs = n; mx.max_words = words_max;
                                use_mx = true;
                            }
                        }
                    }
                    LZ_CUDA(c, launch_staged_scan(sl.st, sl.B, c->d_recs, c->d_meta, c->d_row_ref + b.a, c->d_row_idx + b.a,
                                                  c->d_row_off + b.a, nb, q_ids ? c->d_q_ids : nullptr, ch.k0, (uint32_t)np,
                                                  ch.words, P, c->d_bitmap, c->n_sm, use_mx ? &mx : nullptr));
                    if (c->debug_chunks) LZ_CUDA(c, cudaEventRecord(sl.dbg[1], sl.st));
                    LZ_CUDA(c, cudaMemcpyAsync(sl.h_total, sl.B.rec_off + ch.words, sizeof(uint32_t), cudaMemcpyDeviceToHost, sl.st));
                    // rec_off is a 32-bit scan: a word holds up to 16 hit runs, so beyond 2^28 words the total could wrap --
                    // such chunks also get the exact 64-bit count (one more pass over the hit masks, ~0.5 ms)
                    const bool may_wrap = ch.words * 16ull >= (1ull << 32);
                    if (may_wrap) {
                        LZ_CUDA(c, launch_run_total(sl.st, sl.B, ch.words));
                        LZ_CUDA(c, cudaMemcpyAsync(sl.h_total + 2, sl.B.counter + 3, sizeof(unsigned long long), cudaMemcpyDeviceToHost, sl.st));
                    }
                    LZ_CUDA(c, cudaStreamSynchronize(sl.st));
                    uint64_t n_hit = *sl.h_total;
                    if (may_wrap) {
                        unsigned long long exact;
                        memcpy(&exact, sl.h_total + 2, sizeof(exact));
                        n_hit = exact;
                    }
                    T.launches += 1 + (ch.words ? (use_mx ? 2 * ((mx.n_rows_chunk + 31) / 32) : 1) : 0);     // our kernels (CUB's scans not counted)
                    if (ch.words) {
                        const double ratio = (double)n_hit / (double)ch.words;
                        c->rec_per_word = ratio > c->rec_per_word || !c->density_seen ? ratio : 0.5 * (c->rec_per_word + ratio);
                        c->density_seen = true;
                        if (c->rec_per_word < 0.25) c->rec_per_word = 0.25;
                    }
                    if (c->debug_chunks)
                        fprintf(stderr, "[lzani] chunk %u slot %u: pairs %llu words %llu hit runs %llu (capacity %llu)%s\n", c->staged_chunks,
                                c->staged_chunks % (uint32_t)c->n_slots, (unsigned long long)np, (unsigned long long)ch.words,
                                (unsigned long long)n_hit, (unsigned long long)sl.recs_cap, n_hit > sl.recs_cap ? "  -> halved" : "");
                    if (n_hit > sl.recs_cap || n_hit > 0xFFFFFFFFull) {
                        if (np > 1) {                      // too many hit runs for the record buffers: halve the chunk
                            const uint64_t mid = ch.k0 + np / 2;
                            const uint64_t w0 = range_words(ch.k0, mid);
                            redo.push_back(Range{mid, ch.k1, ch.words - w0});
                            redo.push_back(Range{ch.k0, mid, w0});
                            continue;
                        }
                        if (n_hit > 0xFFFFFFFFull) return fail(c, LZANI_ERR_NOMEM, "one pair has more than 2^32 hit runs");
                        if (int rc = ensure_slot(c, sl, sl.pairs_cap, sl.words_cap, n_hit)) return rc;
                    }
                    if (c->debug_chunks) LZ_CUDA(c, cudaEventRecord(sl.dbg[2], sl.st));
                    LZ_CUDA(c, launch_staged_walk(sl.st, wgrid, sl.B, c->d_recs, ch.k0, (uint32_t)np, ch.words, n_hit, P,
                                                  c->d_bitmap, c->d_table, c->d_out,
                                                  c->lane_walk ? c->n_sm * c->lane_blocks_per_sm : 0));
                    if (c->debug_chunks) { LZ_CUDA(c, cudaEventRecord(sl.dbg[3], sl.st)); sl.dbg_chunk = c->staged_chunks; sl.dbg_set = true; }
                    T.launches += 1 + (ch.words ? 1 : 0) + (n_hit ? (is_default_kernel_params(P) && P.mqd <= 57 ? 2 : 1) : 0) +
                                  (c->lane_walk && is_default_kernel_params(P) && P.mqd <= 57 ? 2 : 0);
                    ++c->staged_chunks;
                }
                for (auto& sl : c->slot) {
                    LZ_CUDA(c, cudaEventRecord(sl.done, sl.st));
                    LZ_CUDA(c, cudaStreamWaitEvent(st, sl.done, 0));
                }
            } else
                LZ_CUDA(c, launch_parse(st, grid, c->d_recs, c->d_meta, c->d_row_ref + b.a, c->d_row_idx + b.a,
                                        c->d_row_off + b.a, nb, q_ids ? c->d_q_ids : nullptr, row_off[b.a], row_off[b.b], P,
                                        c->d_bitmap, c->d_table, c->d_counter, c->d_out));
            LZ_CUDA(c, cudaEventRecord(c->ev[5], st));
            LZ_CUDA(c, cudaStreamSynchronize(st));
            T.h2d_ms += ms_between(c->ev[2], c->ev[3]);
            T.index_ms += ms_between(c->ev[3], c->ev[4]);
            T.parse_ms += ms_between(c->ev[4], c->ev[5]);
            T.launches += (h_blk[nb] ? 1 : 0) + (!staged && row_off[b.b] > row_off[b.a] ? 1 : 0);
        }
    }
    LZ_CUDA(c, cudaEventRecord(c->ev[6], st));
    if (out && n_pairs) {
        LZ_CUDA(c, cudaMemcpyAsync(out, c->d_out, n_pairs * sizeof(lzani_result_t), cudaMemcpyDeviceToHost, st));
        T.d2h_bytes = n_pairs * sizeof(lzani_result_t);
    }
    LZ_CUDA(c, cudaEventRecord(c->ev[7], st));
    LZ_CUDA(c, cudaStreamSynchronize(st));
    T.h2d_ms += ms_between(c->ev[0], c->ev[1]);
    T.d2h_ms = ms_between(c->ev[6], c->ev[7]);
    T.total_ms = ms_between(c->ev[0], c->ev[7]);
    T.n_pairs = n_pairs * (uint64_t)repeat;
    T.q_bases = q_bases * (uint64_t)repeat;
    T.r_bases = r_bases * (uint64_t)repeat;
    c->timing = T;
    c->last_row_ref = row_ref;
    c->last_row_off = row_off;
    c->last_sparse = q_ids != nullptr;
    c->last_sorted = rows_sorted;
    c->last_valid = true;
    if (rq) {
        // regions: bring them back, group by pair (counting sort), order each group like calc_regions
        unsigned long long total = 0;
        LZ_CUDA(c, cudaMemcpy(&total, c->d_reg_counter, sizeof(total), cudaMemcpyDeviceToHost));
        *rq->n_regions = total;
        if (total > rq->cap) return fail(c, LZANI_ERR_NOMEM, "regions buffer too small: " + std::to_string(total) + " regions");
        std::vector<DevRegion> h(total);
        if (total) LZ_CUDA(c, cudaMemcpy(h.data(), c->d_regs, sizeof(DevRegion) * total, cudaMemcpyDeviceToHost));
        c->timing.d2h_bytes += sizeof(DevRegion) * total;
        std::vector<uint64_t> off(n_pairs + 1, 0);
        for (const DevRegion& d : h) ++off[d.pair + 1];
        for (uint64_t k = 0; k < n_pairs; ++k) off[k + 1] += off[k];
        std::vector<uint64_t> cur(off.begin(), off.end() - 1);
        for (const DevRegion& d : h) {
            lzani_region_t& o = rq->regions[cur[d.pair]++];
            o.ref_start = d.g.ref_start; o.ref_end = d.g.ref_end; o.seq_start = d.g.seq_start; o.seq_end = d.g.seq_end;
            o.num_matches = d.g.num_matches; o.num_mismatches = d.g.num_mismatches;
        }
        for (uint64_t k = 0; k < n_pairs; ++k)
            std::stable_sort(rq->regions + off[k], rq->regions + off[k + 1], [](const lzani_region_t& a, const lzani_region_t& b) {
                const int la = a.seq_end - a.seq_start, lb = b.seq_end - b.seq_start;     // parser.cpp:828-833
                if (la != lb) return la > lb;
                return a.seq_start < b.seq_start;
            });
        memcpy(rq->region_off, off.data(), sizeof(uint64_t) * (n_pairs + 1));
    }
    return LZANI_OK;
}

}  // namespace

namespace {

// ---- a run on a group: the rows are dealt over the devices (dense rows: row j -> device j mod G, every device gets the
// same length mix of the length-sorted ids; sparse rows: heaviest first onto the least loaded device, by query bases),
// every device parses its rows (run_rows_impl, its own host thread), and each worker scatters its results into the
// caller's CSR order.  No collective: the only shared data are the read-only sequences, replicated at set_sequences.
int group_run_rows(lzani_ctx* c, uint32_t n_rows, const uint32_t* ref_ids, const uint64_t* row_off_in, const uint32_t* q_ids,
                   int repeat, lzani_result_t* out, bool copy_out, const RegionReq* rq = nullptr)
{
    const uint32_t G = (uint32_t)c->subs.size();
    const uint32_t n = c->n_seqs;
    if (q_ids && !row_off_in) return fail(c, LZANI_ERR_ARG, "q_ids given without row_off");
    if (copy_out && !out) {
        bool any = false;
        if (q_ids) any = row_off_in[n_rows] > row_off_in[0]; else any = n_rows && n > 1;
        if (any) return fail(c, LZANI_ERR_ARG, "out == NULL");
    }
    if (rq && (!rq->n_regions || (rq->cap && (!rq->regions || !rq->region_off)))) return fail(c, LZANI_ERR_ARG, "region buffers == NULL");
    // global CSR offsets of the rows (relative to the first)
    std::vector<uint64_t> goff((size_t)n_rows + 1, 0);
    for (uint32_t j = 0; j < n_rows; ++j) {
        if (q_ids) {
            if (row_off_in[j + 1] < row_off_in[j]) return fail(c, LZANI_ERR_ARG, "row_off is not non-decreasing");
            goff[j + 1] = row_off_in[j + 1] - row_off_in[0];
        } else goff[j + 1] = goff[j] + (n ? n - 1 : 0);
    }
    const uint32_t* q_src = q_ids ? q_ids + row_off_in[0] : nullptr;
    std::vector<uint32_t> owner(n_rows);
    if (!q_ids) for (uint32_t j = 0; j < n_rows; ++j) owner[j] = j % G;
    else {
        const lzani_ctx* s0 = c->subs[0];
        std::vector<uint64_t> cost(n_rows, 0), load(G, 0);
        for (uint32_t j = 0; j < n_rows; ++j)
            for (uint64_t k = goff[j]; k < goff[j + 1]; ++k) {
                if (q_src[k] >= n) return fail(c, LZANI_ERR_ARG, "query id out of range");
                cost[j] += s0->meta[q_src[k]].len;
            }
        std::vector<uint32_t> order(n_rows);
        for (uint32_t j = 0; j < n_rows; ++j) order[j] = j;
        std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return cost[a] > cost[b]; });
        for (uint32_t j : order) {
            const uint32_t g = (uint32_t)(std::min_element(load.begin(), load.end()) - load.begin());
            owner[j] = g;
            load[g] += cost[j] + 1;
        }
    }
    struct Part {
        std::vector<uint32_t> rows_local, refs, q;
        std::vector<uint64_t> off;
        std::vector<lzani_result_t> res;
        std::vector<uint64_t> region_off;
        std::vector<lzani_region_t> regions;
        uint64_t n_regions = 0;
    };
    std::vector<Part> part(G);
    for (uint32_t g = 0; g < G; ++g) part[g].off.assign(1, 0);
    for (uint32_t j = 0; j < n_rows; ++j) {
        Part& pt = part[owner[j]];
        pt.rows_local.push_back(j);
        pt.refs.push_back(ref_ids ? ref_ids[j] : j);
        if (q_ids) pt.q.insert(pt.q.end(), q_src + goff[j], q_src + goff[j + 1]);
        pt.off.push_back(pt.off.back() + (goff[j + 1] - goff[j]));
    }
    const int rc = group_each(c, [&](uint32_t g, lzani_ctx* s) {
        Part& pt = part[g];
        const uint64_t np = pt.off.back();
        if (copy_out) pt.res.resize(np);
        int r;
        if (rq) {
            pt.region_off.assign(np + 1, 0);
            uint64_t cap = std::max<uint64_t>(1024, 4 * np);
            for (;;) {
                pt.regions.resize(cap);
                const RegionReq sub_rq{pt.region_off.data(), pt.regions.data(), cap, &pt.n_regions};
                r = run_rows_impl(s, (uint32_t)pt.refs.size(), pt.refs.data(), q_ids ? pt.off.data() : nullptr, q_ids ? pt.q.data() : nullptr,
                                  repeat, pt.res.data(), copy_out, &sub_rq);
                if (r == LZANI_ERR_NOMEM && pt.n_regions > cap) { cap = pt.n_regions + pt.n_regions / 8; continue; }
                break;
            }
        } else
            r = run_rows_impl(s, (uint32_t)pt.refs.size(), pt.refs.data(), q_ids ? pt.off.data() : nullptr, q_ids ? pt.q.data() : nullptr,
                              repeat, copy_out ? pt.res.data() : nullptr, copy_out, nullptr);
        if (r != LZANI_OK) return r;
        if (copy_out)
            for (size_t jl = 0; jl < pt.rows_local.size(); ++jl) {
                const uint32_t j = pt.rows_local[jl];
                if (goff[j + 1] > goff[j])
                    memcpy(out + goff[j], pt.res.data() + pt.off[jl], sizeof(lzani_result_t) * (goff[j + 1] - goff[j]));
            }
        return (int)LZANI_OK;
    });
    group_timing(c);
    if (rc != LZANI_OK) return rc;
    if (rq) {
        uint64_t total = 0;
        for (const Part& pt : part) total += pt.n_regions;
        *rq->n_regions = total;
        if (total > rq->cap) return fail(c, LZANI_ERR_NOMEM, "regions buffer too small: " + std::to_string(total) + " regions");
        // stitch the per-device region lists back into the caller's CSR order
        const uint64_t np_all = goff[n_rows];
        std::vector<uint64_t> cnt(np_all + 1, 0);
        for (const Part& pt : part)
            for (size_t jl = 0; jl < pt.rows_local.size(); ++jl) {
                const uint32_t j = pt.rows_local[jl];
                for (uint64_t k = 0; k < goff[j + 1] - goff[j]; ++k)
                    cnt[goff[j] + k + 1] = pt.region_off[pt.off[jl] + k + 1] - pt.region_off[pt.off[jl] + k];
            }
        for (uint64_t k = 0; k < np_all; ++k) cnt[k + 1] += cnt[k];
        for (const Part& pt : part)
            for (size_t jl = 0; jl < pt.rows_local.size(); ++jl) {
                const uint32_t j = pt.rows_local[jl];
                const uint64_t a = pt.region_off[pt.off[jl]], b = pt.region_off[pt.off[jl + 1]];
                if (b > a) memcpy(rq->regions + cnt[goff[j]], pt.regions.data() + a, sizeof(lzani_region_t) * (b - a));
            }
        memcpy(rq->region_off, cnt.data(), sizeof(uint64_t) * (np_all + 1));
    }
    return LZANI_OK;
}

int group_filter_pairs(lzani_ctx* c, const uint32_t* eff_len, const lzani_out_filter_t* flt, lzani_pair_t* pairs, uint64_t cap,
                       uint64_t* n_out)
{
    if (!eff_len || !flt || !n_out || (cap && !pairs)) return fail(c, LZANI_ERR_ARG, "eff_len / flt / pairs / n_out == NULL");
    const size_t G = c->subs.size();
    std::vector<std::vector<lzani_pair_t>> sv(G);
    const int rc = group_each(c, [&](uint32_t g, lzani_ctx* s) {
        std::vector<lzani_pair_t>& v = sv[g];
        v.resize(1024);
        for (;;) {
            uint64_t need = 0;
            const int r = lzani_filter_pairs(s, c->subs.data(), (int)G, eff_len, flt, v.data(), v.size(), &need);
            if (r == LZANI_ERR_NOMEM && need > v.size()) { v.resize(need); continue; }
            if (r != LZANI_OK) return r;
            v.resize(need);
            return (int)LZANI_OK;
        }
    });
    group_timing(c);
    if (rc != LZANI_OK) return rc;
    uint64_t total = 0;
    for (const auto& v : sv) total += v.size();
    *n_out = total;
    if (total > cap) return fail(c, LZANI_ERR_NOMEM, "pairs buffer too small: " + std::to_string(total) + " pairs survive the filter");
    uint64_t at = 0;
    for (const auto& v : sv) {
        if (!v.empty()) memcpy(pairs + at, v.data(), sizeof(lzani_pair_t) * v.size());
        at += v.size();
    }
    return LZANI_OK;
}

}  // namespace

extern "C" int lzani_run(lzani_ctx* c, uint32_t n_refs, const uint64_t* row_off, const uint32_t* q_ids,
                         lzani_result_t* out)
{
    if (!c) return LZANI_ERR_ARG;
    if (!c->subs.empty()) {
        if (n_refs != c->n_seqs) return fail(c, LZANI_ERR_ARG, "n_refs must equal the number of sequences");
        return group_run_rows(c, n_refs, nullptr, row_off, q_ids, 1, out, true);
    }
    if (c->d_recs && n_refs != c->n_seqs) return fail(c, LZANI_ERR_ARG, "n_refs must equal the number of sequences");
    return run_rows_impl(c, n_refs, nullptr, row_off, q_ids, 1, out, true);
}

extern "C" int lzani_run_rows(lzani_ctx* c, uint32_t n_rows, const uint32_t* ref_ids, const uint64_t* row_off,
                              const uint32_t* q_ids, lzani_result_t* out)
{
    if (c && !c->subs.empty()) return group_run_rows(c, n_rows, ref_ids, row_off, q_ids, 1, out, true);
    return run_rows_impl(c, n_rows, ref_ids, row_off, q_ids, 1, out, true);
}

extern "C" int lzani_run_rows_regions(lzani_ctx* c, uint32_t n_rows, const uint32_t* ref_ids, const uint64_t* row_off,
                                      const uint32_t* q_ids, lzani_result_t* out, uint64_t* region_off,
                                      lzani_region_t* regions, uint64_t regions_cap, uint64_t* n_regions)
{
    const RegionReq rq{region_off, regions, regions_cap, n_regions};
    if (c && !c->subs.empty()) return group_run_rows(c, n_rows, ref_ids, row_off, q_ids, 1, out, true, &rq);
    return run_rows_impl(c, n_rows, ref_ids, row_off, q_ids, 1, out, true, &rq);
}

extern "C" int lzani_run_rows_resident(lzani_ctx* c, uint32_t n_rows, const uint32_t* ref_ids,
                                       const uint64_t* row_off, const uint32_t* q_ids, int repeat,
                                       lzani_result_t* out)
{
    if (c && !c->subs.empty()) return group_run_rows(c, n_rows, ref_ids, row_off, q_ids, repeat, out, out != nullptr);
    return run_rows_impl(c, n_rows, ref_ids, row_off, q_ids, repeat, out, false);
}

// ------------------------------------------------------------------------------------------
// derived metrics + --out-filter compaction on the device.
extern "C" int lzani_filter_pairs(lzani_ctx* c, lzani_ctx* const* group, int n_group, const uint32_t* eff_len,
                                  const lzani_out_filter_t* flt, lzani_pair_t* pairs, uint64_t cap, uint64_t* n_out)
{
    static_assert(sizeof(lzani_pair_t) == sizeof(PairOut), "lzani_pair_t layout");
    if (!c) return LZANI_ERR_ARG;
    if (!c->subs.empty()) return group_filter_pairs(c, eff_len, flt, pairs, cap, n_out);       // the group is its own `group`
    if (c->device < 0) return fail(c, LZANI_ERR_CUDA, "context has no device (lzani_create failed)");
    if (!eff_len || !flt || !n_out || (cap && !pairs)) return fail(c, LZANI_ERR_ARG, "eff_len / flt / pairs / n_out == NULL");
    if (!c->last_valid) return fail(c, LZANI_ERR_STATE, "no results on the device: run lzani_run_rows* first");
    lzani_ctx* self[1] = {c};
    if (!group || n_group < 1) { group = self; n_group = 1; }
    const uint32_t n = c->n_seqs;
    LZ_CUDA(c, cudaSetDevice(c->device));

    // where every reference row's results live, across the group
    std::vector<RowLoc> loc(n ? n : 1, RowLoc{nullptr, nullptr, 0, 0});
    for (int g = 0; g < n_group; ++g) {
        const lzani_ctx* m = group[g];
        if (!m || !m->last_valid) return fail(c, LZANI_ERR_STATE, "a context of the group holds no results");
        if (m->n_seqs != n) return fail(c, LZANI_ERR_ARG, "the contexts of the group hold different sequence sets");
        if (m->last_sparse && !m->last_sorted)
            return fail(c, LZANI_ERR_ARG, "query ids must ascend within every row for lzani_filter_pairs");
        if (m->device != c->device) {
            int can = 0;
            LZ_CUDA(c, cudaDeviceCanAccessPeer(&can, c->device, m->device));
            if (!can) return fail(c, LZANI_ERR_CUDA, "no peer access from device " + std::to_string(c->device) + " to " + std::to_string(m->device));
            const cudaError_t e = cudaDeviceEnablePeerAccess(m->device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) (void)cudaGetLastError();
            else if (e != cudaSuccess) return fail(c, LZANI_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
        }
        for (size_t j = 0; j < m->last_row_ref.size(); ++j) {
            const uint64_t a = m->last_row_off[j], b = m->last_row_off[j + 1];
            RowLoc& L = loc[m->last_row_ref[j]];
            L.res = (const uint8_t*)m->d_out + a * sizeof(lzani_result_t);
            L.q = m->last_sparse ? m->d_q_ids + a : nullptr;
            L.cnt = (uint32_t)(b - a);
        }
    }

    const uint64_t n_pairs = c->last_row_off.empty() ? 0 : c->last_row_off.back();
    const uint64_t nb = filter_blocks(n_pairs);
    if (int rc = ensure(c, c->d_loc, c->loc_cap, (uint64_t)loc.size())) return rc;
    if (int rc = ensure(c, c->d_eff_len, c->eff_cap, (uint64_t)(n ? n : 1))) return rc;
    if (int rc = ensure(c, c->d_flags, c->flags_cap, n_pairs)) return rc;
    if (int rc = ensure(c, c->d_blk_cnt, c->blk_cnt_cap, nb)) return rc;
    if (int rc = ensure(c, c->d_blk_off, c->blk_off_cap, nb + 1)) return rc;

    cudaStream_t st = c->stream;
    FilterArgs A;
    A.row_ref = c->d_row_ref; A.row_off = c->d_row_off; A.q_ids = c->last_sparse ? c->d_q_ids : nullptr; A.res = c->d_out;
    A.n_rows = (uint32_t)c->last_row_ref.size(); A.n_pairs = n_pairs;
    A.loc = c->d_loc; A.eff_len = c->d_eff_len;
    A.flt = OutFilter{flt->gani, flt->ani, flt->tani, flt->qcov, flt->rcov};

    LZ_CUDA(c, cudaEventRecord(c->ev[0], st));
    LZ_CUDA(c, cudaMemcpyAsync(c->d_loc, loc.data(), sizeof(RowLoc) * loc.size(), cudaMemcpyHostToDevice, st));
    if (n) LZ_CUDA(c, cudaMemcpyAsync(c->d_eff_len, eff_len, sizeof(uint32_t) * n, cudaMemcpyHostToDevice, st));
    LZ_CUDA(c, cudaMemsetAsync(c->d_reg_counter, 0, sizeof(unsigned long long), st));
    LZ_CUDA(c, launch_filter_count(st, A, c->d_flags, c->d_blk_cnt, c->d_blk_off, c->d_reg_counter));
    unsigned long long total = 0, missing = 0;
    LZ_CUDA(c, cudaMemcpyAsync(&total, c->d_blk_off + nb, sizeof(total), cudaMemcpyDeviceToHost, st));
    LZ_CUDA(c, cudaMemcpyAsync(&missing, c->d_reg_counter, sizeof(missing), cudaMemcpyDeviceToHost, st));
    LZ_CUDA(c, cudaStreamSynchronize(st));
    *n_out = total;
    if (missing)
        return fail(c, LZANI_ERR_ARG, "the pair list is not symmetric: the opposite direction of " + std::to_string(missing) +
                                          " pairs is missing (the reference asserts it, lz_matcher.cpp:417)");
    if (total > cap) return fail(c, LZANI_ERR_NOMEM, "pairs buffer too small: " + std::to_string(total) + " pairs survive the filter");
    if (int rc = ensure(c, c->d_pairs, c->pairs_cap, (uint64_t)total)) return rc;
    LZ_CUDA(c, launch_filter_emit(st, A, c->d_flags, c->d_blk_off, c->d_pairs, total));
    if (total) LZ_CUDA(c, cudaMemcpyAsync(pairs, c->d_pairs, sizeof(PairOut) * total, cudaMemcpyDeviceToHost, st));
    LZ_CUDA(c, cudaEventRecord(c->ev[1], st));
    LZ_CUDA(c, cudaStreamSynchronize(st));
    c->timing.filter_ms = ms_between(c->ev[0], c->ev[1]);
    c->timing.filter_d2h_bytes = sizeof(PairOut) * total + 16;
    c->timing.launches += nb ? 3 : 1;
    return LZANI_OK;
}

// ------------------------------------------------------------------------------------------
// the sparse pair list built on the device.
extern "C" int lzani_build_pair_list(lzani_ctx* c, uint32_t n, uint64_t n_edges, const uint32_t* edge_i,
                                     const uint32_t* edge_j, const uint32_t* reorder_map, uint64_t* row_off,
                                     uint32_t* q_ids, uint64_t cap, uint64_t* n_entries)
{
    if (!c) return LZANI_ERR_ARG;
    if (!c->subs.empty()) {          // any one device does: the list is host data again afterwards
        const int rc = lzani_build_pair_list(c->subs[0], n, n_edges, edge_i, edge_j, reorder_map, row_off, q_ids, cap, n_entries);
        if (rc != LZANI_OK) c->err = c->subs[0]->err;
        c->timing.pairlist_ms = c->subs[0]->timing.pairlist_ms;
        return rc;
    }
    if (c->device < 0) return fail(c, LZANI_ERR_CUDA, "context has no device (lzani_create failed)");
    if (!row_off || !n_entries || (n_edges && (!edge_i || !edge_j))) return fail(c, LZANI_ERR_ARG, "edge / row_off / n_entries == NULL");
    const uint64_t n_keys = 2 * n_edges;
    *n_entries = n_keys;
    if (n_keys > cap) return fail(c, LZANI_ERR_NOMEM, "q_ids buffer too small: " + std::to_string(n_keys) + " entries");
    if (n_keys && !q_ids) return fail(c, LZANI_ERR_ARG, "q_ids == NULL");
    LZ_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    int end_bit = 32 + ceil_log2_u64((uint64_t)(n > 1 ? n : 2));
    if (end_bit > 64) end_bit = 64;
    const size_t scratch_bytes = pairlist_sort_scratch_bytes(n_keys, end_bit);

    uint32_t *d_i = nullptr, *d_j = nullptr, *d_map = nullptr, *d_q = nullptr;
    unsigned long long *d_ka = nullptr, *d_kb = nullptr, *d_off = nullptr;
    void* d_scratch = nullptr;
    int rc = LZANI_OK;
    unsigned long long bad = 0;
    do {
        cudaError_t e;
#define LZ_TRY(call) if ((e = (call)) != cudaSuccess) { rc = fail(c, e == cudaErrorMemoryAllocation ? LZANI_ERR_NOMEM : LZANI_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e)); break; }
        LZ_TRY(cudaMalloc((void**)&d_i, sizeof(uint32_t) * (n_edges ? n_edges : 1)));
        LZ_TRY(cudaMalloc((void**)&d_j, sizeof(uint32_t) * (n_edges ? n_edges : 1)));
        LZ_TRY(cudaMalloc((void**)&d_map, sizeof(uint32_t) * (n ? n : 1)));
        LZ_TRY(cudaMalloc((void**)&d_q, sizeof(uint32_t) * (n_keys ? n_keys : 1)));
        LZ_TRY(cudaMalloc((void**)&d_ka, sizeof(unsigned long long) * (n_keys ? n_keys : 1)));
        LZ_TRY(cudaMalloc((void**)&d_kb, sizeof(unsigned long long) * (n_keys ? n_keys : 1)));
        LZ_TRY(cudaMalloc((void**)&d_off, sizeof(unsigned long long) * ((size_t)n + 1)));
        LZ_TRY(cudaMalloc(&d_scratch, scratch_bytes ? scratch_bytes : 1));
        LZ_TRY(cudaEventRecord(c->ev[0], st));
        if (n_edges) {
            LZ_TRY(cudaMemcpyAsync(d_i, edge_i, sizeof(uint32_t) * n_edges, cudaMemcpyHostToDevice, st));
            LZ_TRY(cudaMemcpyAsync(d_j, edge_j, sizeof(uint32_t) * n_edges, cudaMemcpyHostToDevice, st));
        }
        if (reorder_map && n) LZ_TRY(cudaMemcpyAsync(d_map, reorder_map, sizeof(uint32_t) * n, cudaMemcpyHostToDevice, st));
        LZ_TRY(cudaMemsetAsync(c->d_reg_counter, 0, sizeof(unsigned long long), st));
        LZ_TRY(launch_pairlist(st, d_i, d_j, n_edges, reorder_map ? d_map : nullptr, n, d_ka, d_kb, d_scratch, scratch_bytes,
                               end_bit, d_off, d_q, c->d_reg_counter));
        LZ_TRY(cudaMemcpyAsync(row_off, d_off, sizeof(uint64_t) * ((size_t)n + 1), cudaMemcpyDeviceToHost, st));
        if (n_keys) LZ_TRY(cudaMemcpyAsync(q_ids, d_q, sizeof(uint32_t) * n_keys, cudaMemcpyDeviceToHost, st));
        LZ_TRY(cudaMemcpyAsync(&bad, c->d_reg_counter, sizeof(bad), cudaMemcpyDeviceToHost, st));
        LZ_TRY(cudaEventRecord(c->ev[1], st));
        LZ_TRY(cudaStreamSynchronize(st));
#undef LZ_TRY
    } while (0);
    cudaFree(d_i); cudaFree(d_j); cudaFree(d_map); cudaFree(d_q); cudaFree(d_ka); cudaFree(d_kb); cudaFree(d_off); cudaFree(d_scratch);
    if (rc != LZANI_OK) return rc;
    if (bad) return fail(c, LZANI_ERR_ARG, std::to_string(bad) + " edges name a sequence id >= n");
    c->timing.pairlist_ms = ms_between(c->ev[0], c->ev[1]);
    return LZANI_OK;
}
