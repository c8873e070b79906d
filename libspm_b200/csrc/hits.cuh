// hits.cuh -- the compacted hit buffer kernels append to.
#pragma once
#include <stdint.h>
#include "../../include/spmb.h"

struct HitSink {
    spmb_hit *hits;             // device buffer
    unsigned long long *count;  // total hits produced (may exceed cap: finish() then grows and re-runs)
    unsigned long long cap;
};
