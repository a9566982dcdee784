// batch_plan.h — how a render's samples are cut into batches and spread over lanes (host code, no CUDA: api.cu: enqueue_render
// uses it, tests/hostsim checks it on the CPU).
//
// A batch is whole images of samples. Rules, in order (measured on the B200: profiles/r02_tune16 .. r02_tune22):
//   1. at most `batch_paths` camera paths per batch, halved until `lanes` pools of that size fit the memory budget, then fewer lanes;
//   2. a job smaller than one batch per lane is split into one batch per lane as long as each keeps >= 4 Mi paths;
//   3. a job of between one and two batches per lane runs as ONE larger batch per lane when that fits the budget
//      (a second round of half-size batches costs every lane a second tail);
//   4. otherwise equal batches, a whole number of them per lane (125 samples in batches of at most 29 on 4 lanes are 8 batches of
//      16, not 4 of 29 and one of 9 with three lanes idle behind the fifth).
#pragma once
#include <algorithm>
#include <cstddef>
#include <cstdint>

namespace rt {

struct BatchPlan {
    uint64_t samples_per_batch = 0, n_batches = 0;
    size_t n_lanes = 0;
    bool ok = false; // false: not even one batch of one sample per pixel fits the budget
};

// bytes_for(paths) = device bytes ONE lane needs for batches of `paths` camera paths
template <class BytesFor>
inline BatchPlan plan_batches(uint64_t total_samples, uint64_t npix, uint64_t batch_paths, size_t want_lanes, double budget, BytesFor bytes_for) {
    BatchPlan p;
    if (total_samples == 0 || npix == 0 || want_lanes == 0) return p;
    uint64_t spb = std::max<uint64_t>(1, batch_paths / npix);
    spb = std::min<uint64_t>(spb, total_samples);
    uint64_t n_batches = 0;
    size_t n_lanes = 0;
    for (;;) { // rule 1
        n_batches = (total_samples + spb - 1) / spb;
        n_lanes = (size_t)std::min<uint64_t>(n_batches, want_lanes);
        if ((double)n_lanes * bytes_for(spb * npix) <= budget) break;
        if (spb > 1) spb = (spb + 1) / 2;   // smaller batches first,
        else if (want_lanes > 1) want_lanes--; // then fewer of them in flight
        else return p;
    }
    if (n_batches < want_lanes) { // rule 2
        const uint64_t nb = std::min<uint64_t>(std::min<uint64_t>(want_lanes, total_samples), total_samples * npix >> 22);
        if (nb > n_batches) {
            spb = (total_samples + nb - 1) / nb;
            n_batches = (total_samples + spb - 1) / spb;
            n_lanes = (size_t)std::min<uint64_t>(n_batches, want_lanes);
        }
    }
    if (n_batches > n_lanes && n_batches < 2 * n_lanes) { // rule 3
        const uint64_t one = (total_samples + n_lanes - 1) / n_lanes;
        if (one * npix < 0x7FFF0000ull && (double)n_lanes * bytes_for(one * npix) <= budget) {
            spb = one;
            n_batches = (total_samples + spb - 1) / spb;
        }
    }
    if (n_batches > n_lanes) { // rule 4
        const uint64_t nb = (n_batches + n_lanes - 1) / n_lanes * n_lanes;
        spb = (total_samples + nb - 1) / nb;
        n_batches = (total_samples + spb - 1) / spb;
    }
    p.samples_per_batch = spb; p.n_batches = n_batches; p.n_lanes = n_lanes; p.ok = true;
    return p;
}

} // namespace rt
