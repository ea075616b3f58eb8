// msm_ba.cu -- translation unit of the batch-affine fixed-base MSM (kernels in msm_ba.cuh, interface in msm_ba_api.hpp).
#include "msm_ba.cuh"
#include "msm_ba_api.hpp"
#include <string.h>

namespace kb200 {

static BaOffset ba_offset(const MsmTable& tab) {
    // signed-digit offset: adding 2^(c-1) to every window below the top one turns the unsigned window value v into v - 2^(c-1)
    BaOffset off; memset(&off, 0, sizeof off);
    for (int j = 0; j < tab.nwin - 1; j++) { int bit = tab.c * j + tab.c - 1; off.l[bit >> 5] |= 1u << (bit & 31); }
    return off;
}

int ba_init(BaPlan* plan) {
    DevConsts h;
    using namespace kb200::consts;
    memcpy(h.fp_mod, H_FP_P, 48); memcpy(h.fp_one, H_FP_ONE, 48); memcpy(h.fp_r2, H_FP_R2, 48); memcpy(h.fp_b4, H_FP_B4, 48);
    memcpy(h.fp_half, H_FP_HALF, 48); memcpy(h.fp_sqrt_exp, H_FP_SQRT_EXP, 48); memcpy(h.fp_pm2, H_FP_PM2, 48); memcpy(h.fp_beta, H_FP_BETA, 48);
    memcpy(h.fr_one, H_FR_ONE, 32); memcpy(h.fr_r2, H_FR_R2, 32); memcpy(h.fr_rm2, H_FR_RM2, 32); memcpy(h.fr_omega, H_FR_OMEGA, 32);
    memcpy(h.fr_ninv, H_FR_NINV, 32); memcpy(h.fr_mod, H_FR_R, 32);
    memcpy(h.g1_gen_x, H_G1_GEN_X, 48); memcpy(h.g1_gen_y, H_G1_GEN_Y, 48); memcpy(h.g1_neg_gen_y, H_G1_NEG_GEN_Y, 48);
    if (cudaMemcpyToSymbol(c_k, &h, sizeof h) != cudaSuccess) return -1;
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_msm_ba, BA_THREADS, BA_SMEM_BYTES) != cudaSuccess) return -1;
    if (BA_SMEM_BYTES > 48 * 1024)
        if (cudaFuncSetAttribute(k_msm_ba, cudaFuncAttributeMaxDynamicSharedMemorySize, BA_SMEM_BYTES) != cudaSuccess) return -1;
    if (BA_SMEM_BYTES > 48 * 1024)
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_msm_ba, BA_THREADS, BA_SMEM_BYTES) != cudaSuccess) return -1;
    plan->ctas_per_sm = occ < 1 ? 1 : occ;
    plan->threads = BA_THREADS;
    return 0;
}

void ba_workspace(const MsmTable& tab, int grid, size_t n, size_t* dig_bytes, size_t* pts_bytes, size_t* pre_bytes) {
    const size_t K1 = (size_t)BA_UB * tab.nwin * (tab.glv + 1) / 2;
    *dig_bytes = n * KB_N * (size_t)tab.nwin * (tab.glv + 1) * 4;
    *pts_bytes = (size_t)grid * K1 * BA_PTS_SLOTS_NUM / BA_PTS_SLOTS_DEN * 6 * BA_THREADS * sizeof(uint4);
    *pre_bytes = (size_t)grid * K1 * BA_PRE_SLOTS_NUM / BA_PRE_SLOTS_DEN * 3 * BA_THREADS * sizeof(uint4);
}

void ba_launch(const MsmTable& tab, int fmt, const uint32_t* d_scalars, const int32_t* d_status, uint32_t* d_dig, void* d_scr_pts, void* d_scr_pre,
               G1* d_partials, unsigned int* d_counter, int n, unsigned grid, cudaStream_t stream) {
    const BaOffset off = ba_offset(tab);
    const unsigned rb = (unsigned)(((size_t)n * KB_N + 255) / 256);
    if (fmt == 0) k_msm_recode<0><<<rb, 256, 0, stream>>>(tab, off, d_scalars, d_status, d_dig, n);
    else k_msm_recode<1><<<rb, 256, 0, stream>>>(tab, off, d_scalars, d_status, d_dig, n);
    BaArgs a;
    a.tab = tab; a.dig = d_dig; a.scr_pts = (uint4*)d_scr_pts; a.scr_pre = (uint4*)d_scr_pre;
    a.partials = d_partials; a.status = d_status; a.counter = d_counter; a.n = n;
    k_msm_ba<<<grid, BA_THREADS, BA_SMEM_BYTES, stream>>>(a);
}

}  // namespace kb200
