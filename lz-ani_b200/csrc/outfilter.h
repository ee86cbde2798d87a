// outfilter.h -- the derived metrics and the --out-filter test of one unordered pair, exactly as
// CLZMatcher::store_results evaluates them (reference src/lz_matcher.cpp:434-463), written once for
// the device kernels (postproc.cu) and for the host checker (tests/emul).
//
// The reference computes five IEEE doubles per direction from the six integers of the two parses
// and the two lengths and drops an output line when any is below its threshold.  Every operation is
// an int -> double conversion, one double division or one double comparison: all correctly rounded
// on both machines (div.rn.f64 on sm_100a), no sums of products, so there is nothing for FMA
// contraction to change and the device decisions equal the host's bit for bit.
#pragma once
#include <stdint.h>

#include "simt.h"   // LZ_HD

namespace lzani {

// thresholds in the order gani, ani, tani, qcov, rcov (0 = not set: no metric is ever below 0)
struct OutFilter { double gani, ani, tani, qcov, rcov; };

struct DirStats { int32_t mat, lit, cnt; };   // == results_t of one direction

// Bit i of the result: output line i of the unordered pair (a < b) survives the filter.
//   line 0: reference a, query b -- statistics of the parse of b against the index of a  (`q->results`)
//   line 1: reference b, query a -- statistics of the parse of a against the index of b  (`p->results`)
// len_a / len_b: the sequences' lengths without the inter-contig separators
// (`len - (no_parts-1)*mrd`, lz_matcher.cpp:437).
LZ_HD uint32_t out_filter_mask(const OutFilter& f, DirStats ab, DirStats ba, uint32_t len_a, uint32_t len_b)
{
    // index 0 = line 0: lengths are the QUERY lengths of each line (len[0] = len(b), len[1] = len(a))
    const uint32_t len[2] = {len_b, len_a};
    const int32_t mat[2] = {ab.mat, ba.mat};
    const int32_t lit[2] = {ab.lit, ba.lit};
    const double total_ani = (double)(mat[0] + mat[1]) / (double)(uint32_t)(len[0] + len[1]);
    double gani[2], ani[2], cov[2];
    for (int i = 0; i < 2; ++i) {
        gani[i] = (double)mat[i] / (double)len[i];
        ani[i] = (mat[i] + lit[i] != 0) ? (double)mat[i] / (double)(mat[i] + lit[i]) : 0.0;
        cov[i] = (double)(mat[i] + lit[i]) / (double)len[i];
    }
    uint32_t mask = 0;
    for (int i = 0; i < 2; ++i) {
        if (gani[i] < f.gani) continue;
        if (ani[i] < f.ani) continue;
        if (total_ani < f.tani) continue;
        if (cov[i] < f.qcov) continue;
        if (cov[i ^ 1] < f.rcov) continue;
        mask |= 1u << i;
    }
    return mask;
}

}  // namespace lzani
