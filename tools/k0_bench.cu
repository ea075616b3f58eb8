// k0_bench.cu -- K0: integer-pipe microbenchmarks on sm_100a (roofline denominators for the field kernels).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/k0_bench tools/k0_bench.cu ; run on the GPU box.
// Each kernel runs ITERS iterations of a fully unrolled body of independent (or chained) integer ops per thread and
// reports warp-instructions per cycle per SMSP (from the SM clock read by %clock64 on one warp) and G ops/s.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../kateth_b200/csrc/mont_gen.cuh"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

// 1. plain 32-bit IMAD: 16 independent chains
__global__ void __launch_bounds__(256) k_imad32(uint32_t* out, int iters, uint32_t a, uint32_t b) {
    uint32_t acc[16];
#pragma unroll
    for (int k = 0; k < 16; k++) acc[k] = threadIdx.x + k;
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 16; k++) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(acc[k]) : "r"(a), "r"(b));
    }
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) s ^= acc[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 2. IMAD.WIDE with a 64-bit register addend: 16 independent chains
__global__ void __launch_bounds__(256) k_wide(uint64_t* out, int iters, uint32_t a, uint32_t b) {
    uint64_t acc[16];
#pragma unroll
    for (int k = 0; k < 16; k++) acc[k] = threadIdx.x + k;
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 16; k++) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"(a), "r"(b));
    }
    uint64_t s = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) s ^= acc[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 3. IMAD.WIDE.X carry chains as the Montgomery product emits them: 4 independent chains of 6 (lo.cc / hi.cc pairs)
__device__ __forceinline__ void chain6(uint64_t (&t)[6], uint32_t a, uint32_t b) {
    asm volatile(
        "{\n\t.reg .u32 l0,h0,l1,h1,l2,h2,l3,h3,l4,h4,l5,h5;\n\t"
        "mov.b64 {l0,h0}, %0;\n\tmov.b64 {l1,h1}, %1;\n\tmov.b64 {l2,h2}, %2;\n\tmov.b64 {l3,h3}, %3;\n\tmov.b64 {l4,h4}, %4;\n\tmov.b64 {l5,h5}, %5;\n\t"
        "mad.lo.cc.u32 l0, %6, %7, l0;\n\tmadc.hi.cc.u32 h0, %6, %7, h0;\n\t"
        "madc.lo.cc.u32 l1, %6, %7, l1;\n\tmadc.hi.cc.u32 h1, %6, %7, h1;\n\t"
        "madc.lo.cc.u32 l2, %6, %7, l2;\n\tmadc.hi.cc.u32 h2, %6, %7, h2;\n\t"
        "madc.lo.cc.u32 l3, %6, %7, l3;\n\tmadc.hi.cc.u32 h3, %6, %7, h3;\n\t"
        "madc.lo.cc.u32 l4, %6, %7, l4;\n\tmadc.hi.cc.u32 h4, %6, %7, h4;\n\t"
        "madc.lo.cc.u32 l5, %6, %7, l5;\n\tmadc.hi.u32 h5, %6, %7, h5;\n\t"
        "mov.b64 %0, {l0,h0};\n\tmov.b64 %1, {l1,h1};\n\tmov.b64 %2, {l2,h2};\n\tmov.b64 %3, {l3,h3};\n\tmov.b64 %4, {l4,h4};\n\tmov.b64 %5, {l5,h5};\n\t}"
        : "+l"(t[0]), "+l"(t[1]), "+l"(t[2]), "+l"(t[3]), "+l"(t[4]), "+l"(t[5]) : "r"(a), "r"(b));
}
template <int NCH>
__global__ void __launch_bounds__(256) k_wide_x(uint64_t* out, int iters, uint32_t a, uint32_t b) {
    uint64_t t[NCH][6];
#pragma unroll
    for (int c = 0; c < NCH; c++)
#pragma unroll
        for (int k = 0; k < 6; k++) t[c][k] = threadIdx.x + k + c;
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int c = 0; c < NCH; c++) chain6(t[c], a + c, b);
    }
    uint64_t s = 0;
#pragma unroll
    for (int c = 0; c < NCH; c++)
#pragma unroll
        for (int k = 0; k < 6; k++) s ^= t[c][k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 4. IMAD.WIDE (no carry) alternating with IADD3 on the ALU pipe: does the ALU work hide behind the multiplier?
__global__ void __launch_bounds__(256) k_wide_alu(uint64_t* out, int iters, uint32_t a, uint32_t b) {
    uint64_t acc[12];
    uint32_t x[12];
#pragma unroll
    for (int k = 0; k < 12; k++) { acc[k] = threadIdx.x + k; x[k] = k * 3 + threadIdx.x; }
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 12; k++) {
            asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[k]) : "r"(a), "r"(b));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(x[k]) : "r"(a));
        }
    }
    uint64_t s = 0;
#pragma unroll
    for (int k = 0; k < 12; k++) s ^= acc[k] + x[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 5. IMAD.WIDE with an immediate multiplicand (the reduction half of the Montgomery product)
__global__ void __launch_bounds__(256) k_wide_imm(uint64_t* out, int iters, uint32_t a) {
    uint64_t acc[16];
#pragma unroll
    for (int k = 0; k < 16; k++) acc[k] = threadIdx.x + k;
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 16; k++) asm volatile("mad.wide.u32 %0, %1, 0x64774b84, %0;" : "+l"(acc[k]) : "r"(a));
    }
    uint64_t s = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) s ^= acc[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 6. mul.wide (no addend) + separate 64-bit add: IMAD.WIDE with RZ addend
__global__ void __launch_bounds__(256) k_mulwide(uint64_t* out, int iters, uint32_t a, uint32_t b) {
    uint64_t acc[16];
    uint32_t v[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { acc[k] = 0; v[k] = a + k + threadIdx.x; }
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 16; k++) { asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(acc[k]) : "r"(v[k]), "r"(b)); v[k] = (uint32_t)acc[k]; }
    }
    uint64_t s = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) s ^= acc[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// 7. the real Fp Montgomery product (300 IMAD.WIDE + ~90 ALU ops): ILP independent dependent-chains per thread
template <int ILP>
__global__ void __launch_bounds__(256) k_fpmul(uint32_t* out, int iters, uint32_t seed) {
    uint32_t a[ILP][12], b[12];
#pragma unroll
    for (int k = 0; k < 12; k++) { b[k] = seed * 31 + blockIdx.x + k;
#pragma unroll
        for (int c = 0; c < ILP; c++) a[c][k] = seed + threadIdx.x + k + c; }
    b[11] &= 0x0fffffff;
#pragma unroll
    for (int c = 0; c < ILP; c++) a[c][11] &= 0x0fffffff;
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int c = 0; c < ILP; c++) kb200::fp_mul(a[c], a[c], b);
    }
    uint32_t s = 0;
#pragma unroll
    for (int c = 0; c < ILP; c++)
#pragma unroll
        for (int k = 0; k < 12; k++) s ^= a[c][k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// 8. FP64 pipe: 16 independent DFMA chains; and DFMA interleaved 1:1 with IMAD.WIDE (separate pipes?)
__global__ void __launch_bounds__(256) k_dfma(double* out, int iters, double a, double b) {
    double acc[16];
#pragma unroll
    for (int k = 0; k < 16; k++) acc[k] = threadIdx.x + k;
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 16; k++) asm volatile("fma.rz.f64 %0, %0, %1, %2;" : "+d"(acc[k]) : "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) s += acc[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NI>
__global__ void __launch_bounds__(256) k_dfma_imad(double* out, int iters, double a, double b, uint32_t x, uint32_t y) {
    double acc[8];
    uint64_t w[8];
#pragma unroll
    for (int k = 0; k < 8; k++) { acc[k] = threadIdx.x + k; w[k] = threadIdx.x * 7 + k; }
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            asm volatile("fma.rz.f64 %0, %0, %1, %2;" : "+d"(acc[k]) : "d"(a), "d"(b));
#pragma unroll
            for (int q = 0; q < NI; q++) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[(k + q) & 7]) : "r"((uint32_t)w[(k + q + 3) & 7]), "r"(y));
        }
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) s += acc[k] + (double)w[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// 8b. co-issue, corrected: eight DFMA chains and eight IMAD.WIDE chains that share NO register (round 1's version fed the
// multiplier from the accumulators of other chains, so its streams were not independent).  NI IMAD.WIDE per DFMA.
template <int NI>
__global__ void __launch_bounds__(256) k_dfma_imad_indep(double* out, int iters, double a, double b, uint32_t x, uint32_t y) {
    double acc[8];
    uint64_t w[8];
#pragma unroll
    for (int k = 0; k < 8; k++) { acc[k] = threadIdx.x + k; w[k] = threadIdx.x * 7 + k; }
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            asm volatile("fma.rz.f64 %0, %0, %1, %2;" : "+d"(acc[k]) : "d"(a), "d"(b));
#pragma unroll
            for (int q = 0; q < NI; q++) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[(k + 4 * q) & 7]) : "r"(x), "r"(y));
        }
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) s += acc[k] + (double)w[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 8c. the same eight IMAD.WIDE chains alone (the reference rate of 8b's integer stream)
__global__ void __launch_bounds__(256) k_imad_indep8(uint64_t* out, int iters, uint32_t x, uint32_t y) {
    uint64_t w[8];
#pragma unroll
    for (int k = 0; k < 8; k++) w[k] = threadIdx.x * 7 + k;
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 8; k++) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[k]) : "r"(x), "r"(y));
    }
    uint64_t s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) s += w[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 8d. IADD3 stream beside DFMA (the integer adds a DFMA-based product would need): 1 DFMA + NA IADD3, no shared registers
template <int NA>
__global__ void __launch_bounds__(256) k_dfma_iadd(double* out, int iters, double a, double b, uint32_t x) {
    double acc[8];
    uint32_t w[8];
#pragma unroll
    for (int k = 0; k < 8; k++) { acc[k] = threadIdx.x + k; w[k] = threadIdx.x * 7 + k; }
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            asm volatile("fma.rz.f64 %0, %0, %1, %2;" : "+d"(acc[k]) : "d"(a), "d"(b));
#pragma unroll
            for (int q = 0; q < NA; q++) asm volatile("add.u32 %0, %0, %1;" : "+r"(w[(k + 3 * q) & 7]) : "r"(x));
        }
    }
    double s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) s += acc[k] + (double)w[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class F>
static int run(const char* name, F launch, double ops_per_thread_iter, int iters, int grid, int block, int sms, double clock_ghz) {
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    launch(64);
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
        CK(cudaEventRecord(e0)); launch(iters); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    double ops = (double)grid * block * iters * ops_per_thread_iter;
    double tops = ops / (best * 1e-3) / 1e12;
    double per_clk_sm = ops / (best * 1e-3) / (sms * clock_ghz * 1e9);
    printf("%-28s %8.3f ms  %7.3f T thread-ops/s  %6.2f lanes/clk/SM (at %.3f GHz)  warps/SM=%d\n", name, best, tops, per_clk_sm, clock_ghz, grid * block / 32 / sms);
    return 0;
}

int main(int argc, char** argv) {
    int dev = 0; CK(cudaSetDevice(dev));
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, dev));
    int sms = p.multiProcessorCount; double ghz = p.clockRate * 1e-6;
    printf("%s: %d SMs, clockRate %.3f GHz\n", p.name, sms, ghz);
    void* buf; CK(cudaMalloc(&buf, (size_t)sms * 8 * 256 * 8));
    int iters = 20000;
    {   // ONE warp per SM sub-partition (1 CTA of 128 threads per SM): what a single warp can pull through the multiplier on its own
        int grid = sms, block = 128;
        printf("--- 1 CTA x 128 threads per SM (one warp per scheduler)\n");
        run("imad.wide.x 1 chain of 6", [&](int it) { k_wide_x<1><<<grid, block>>>((uint64_t*)buf, it, 3, 5); }, 6, iters, grid, block, sms, ghz);
        run("imad.wide.x 2 chains of 6", [&](int it) { k_wide_x<2><<<grid, block>>>((uint64_t*)buf, it, 3, 5); }, 12, iters, grid, block, sms, ghz);
        run("imad.wide.x 4 chains of 6", [&](int it) { k_wide_x<4><<<grid, block>>>((uint64_t*)buf, it, 3, 5); }, 24, iters, grid, block, sms, ghz);
        run("imad.wide 8 indep chains", [&](int it) { k_imad_indep8<<<grid, block>>>((uint64_t*)buf, it, 3, 5); }, 8, iters, grid, block, sms, ghz);
        run("fp_mul chain ILP1 (x300 MAC)", [&](int it) { k_fpmul<1><<<grid, block>>>((uint32_t*)buf, it / 16, 3); }, 300.0 / 16, iters, grid, block, sms, ghz);
        run("fp_mul chain ILP2 (x300 MAC)", [&](int it) { k_fpmul<2><<<grid, block>>>((uint32_t*)buf, it / 16, 3); }, 600.0 / 16, iters, grid, block, sms, ghz);
    }
    for (int wps = 1; wps <= 8; wps *= 2) {    // CTAs of 256 threads per SM: 8 / 16 / 32 / 64 warps per SM
        int grid = sms * wps, block = 256;
        printf("--- %d CTAs x 256 threads per SM\n", wps);
        run("imad32 (16 chains)", [&](int it) { k_imad32<<<grid, block>>>((uint32_t*)buf, it, 3, 5); }, 16, iters, grid, block, sms, ghz);
        run("imad.wide reg addend", [&](int it) { k_wide<<<grid, block>>>((uint64_t*)buf, it, 3, 5); }, 16, iters, grid, block, sms, ghz);
        run("imad.wide imm multiplicand", [&](int it) { k_wide_imm<<<grid, block>>>((uint64_t*)buf, it, 3); }, 16, iters, grid, block, sms, ghz);
        run("imad.wide.x 1 chain of 6", [&](int it) { k_wide_x<1><<<grid, block>>>((uint64_t*)buf, it, 3, 5); }, 6, iters, grid, block, sms, ghz);
        run("imad.wide.x 2 chains of 6", [&](int it) { k_wide_x<2><<<grid, block>>>((uint64_t*)buf, it, 3, 5); }, 12, iters, grid, block, sms, ghz);
        run("imad.wide.x 4 chains of 6", [&](int it) { k_wide_x<4><<<grid, block>>>((uint64_t*)buf, it, 3, 5); }, 24, iters, grid, block, sms, ghz);
        run("imad.wide + iadd3 pairs", [&](int it) { k_wide_alu<<<grid, block>>>((uint64_t*)buf, it, 3, 5); }, 12, iters, grid, block, sms, ghz);
        run("fp_mul chain ILP1 (x300 MAC)", [&](int it) { k_fpmul<1><<<grid, block>>>((uint32_t*)buf, it / 16, 3); }, 300.0 / 16, iters, grid, block, sms, ghz);
        run("fp_mul chain ILP2 (x300 MAC)", [&](int it) { k_fpmul<2><<<grid, block>>>((uint32_t*)buf, it / 16, 3); }, 600.0 / 16, iters, grid, block, sms, ghz);
        run("dfma (16 chains)", [&](int it) { k_dfma<<<grid, block>>>((double*)buf, it, 1.0000001, 0.5); }, 16, iters, grid, block, sms, ghz);
        run("dfma + 1 imad.wide (pairs)", [&](int it) { k_dfma_imad<1><<<grid, block>>>((double*)buf, it, 1.0000001, 0.5, 3, 5); }, 8, iters, grid, block, sms, ghz);
        run("dfma + 2 imad.wide (triples)", [&](int it) { k_dfma_imad<2><<<grid, block>>>((double*)buf, it, 1.0000001, 0.5, 3, 5); }, 8, iters, grid, block, sms, ghz);
        run("mul.wide dependent", [&](int it) { k_mulwide<<<grid, block>>>((uint64_t*)buf, it, 3, 5); }, 16, iters, grid, block, sms, ghz);
        // round 2: the co-issue question with independent streams (ops counted: DFMA only, so "pairs/s" compares with the dfma line)
        run("imad.wide 8 indep chains", [&](int it) { k_imad_indep8<<<grid, block>>>((uint64_t*)buf, it, 3, 5); }, 8, iters, grid, block, sms, ghz);
        run("dfma || 1 imad.wide indep", [&](int it) { k_dfma_imad_indep<1><<<grid, block>>>((double*)buf, it, 1.0000001, 0.5, 3, 5); }, 8, iters, grid, block, sms, ghz);
        run("dfma || 2 imad.wide indep", [&](int it) { k_dfma_imad_indep<2><<<grid, block>>>((double*)buf, it, 1.0000001, 0.5, 3, 5); }, 8, iters, grid, block, sms, ghz);
        run("dfma || 2 iadd indep", [&](int it) { k_dfma_iadd<2><<<grid, block>>>((double*)buf, it, 1.0000001, 0.5, 3); }, 8, iters, grid, block, sms, ghz);
        run("dfma || 4 iadd indep", [&](int it) { k_dfma_iadd<4><<<grid, block>>>((double*)buf, it, 1.0000001, 0.5, 3); }, 8, iters, grid, block, sms, ghz);
    }
    return 0;
}
