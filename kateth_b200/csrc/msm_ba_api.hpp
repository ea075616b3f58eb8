// msm_ba_api.hpp -- host-side interface of the batch-affine MSM (msm_ba.cuh), which is compiled as its own translation unit
// (msm_ba.cu): the kernel is the one that gets tuned, and a variant of it can be built and linked without recompiling the
// rest of the engine.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace kb200 {

struct MsmTable;
struct G1;

struct BaPlan {
    int ctas_per_sm;        // resident CTAs of k_msm_ba per SM (occupancy query)
    int threads;            // threads per CTA
};
// copies the field constants into this translation unit's constant memory and queries the kernel's occupancy
int ba_init(BaPlan* plan);
// workspace of an MSM batch of n items on a persistent grid of `grid` CTAs
void ba_workspace(const MsmTable& tab, int grid, size_t n, size_t* dig_bytes, size_t* pts_bytes, size_t* pre_bytes);
// recode (fmt 0: blobs, big-endian field elements; 1: little-endian limbs) + the persistent batch-affine kernel; the caller has
// zeroed *counter on the same stream.  partials: one XYZZ sum per item.
void ba_launch(const MsmTable& tab, int fmt, const uint32_t* d_scalars, const int32_t* d_status, uint32_t* d_dig, void* d_scr_pts, void* d_scr_pre,
               G1* d_partials, unsigned int* d_counter, int n, unsigned grid, cudaStream_t stream);

}  // namespace kb200
