// msm_table.cuh -- the fixed-base window table descriptor shared by both MSM shapes (msm.cuh, msm_ba.cuh).
#pragma once
#include "g1.cuh"

namespace kb200 {

#define KB_N 4096
#define KB_BLOB_BYTES 131072

struct MsmTable {
    const uint4* data;   // 6 x uint4 per entry
    int c;               // window bits
    int nwin;            // windows
    int dlast;           // digit magnitudes stored for the top (unsigned) window: (r >> c*(nwin-1)) + 1
    int glv;             // 1: nwin windows cover a 128-bit GLV half (glv.cuh) and dlast is ((N >> c*(nwin-1)) + 1); every
                         //    scalar contributes 2 * nwin points, the second half through Q = (beta x, -y)
};

}  // namespace kb200
