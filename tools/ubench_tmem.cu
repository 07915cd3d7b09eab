// Micro-benchmark: tcgen05.ld read rate out of tensor memory on sm_100a, per SM, with the epilogue's warp layout
// (8 warps, two per TMEM lane quarter).  Variants: 32x32b.x32 / .x64 / .x128 and .pack::16b.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_tmem tools/ubench_tmem.cu && tools/ubench_tmem
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int MODE> __device__ __forceinline__ uint32_t ld_and_fold(uint32_t taddr);

// MODE 0: 32x32b.x32   (32 columns -> 32 registers)
template <> __device__ __forceinline__ uint32_t ld_and_fold<0>(uint32_t taddr) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];\n"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    uint32_t a = 0;
#pragma unroll
    for (int i = 0; i < 32; i += 8) a ^= r[i];
    return a;
}
// MODE 1: 32x32b.x32.pack::16b   (64 columns of 16-bit data -> 32 registers)
template <> __device__ __forceinline__ uint32_t ld_and_fold<1>(uint32_t taddr) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.pack::16b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];\n"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    uint32_t a = 0;
#pragma unroll
    for (int i = 0; i < 32; i += 8) a ^= r[i];
    return a;
}
// MODE 2: 32x32b.x64 (64 columns -> 64 registers)
template <> __device__ __forceinline__ uint32_t ld_and_fold<2>(uint32_t taddr) {
    uint32_t r[64];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,"
        "%32,%33,%34,%35,%36,%37,%38,%39,%40,%41,%42,%43,%44,%45,%46,%47,"
        "%48,%49,%50,%51,%52,%53,%54,%55,%56,%57,%58,%59,%60,%61,%62,%63}, [%64];\n"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]),
          "=r"(r[32]), "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]),
          "=r"(r[40]), "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]),
          "=r"(r[48]), "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]),
          "=r"(r[56]), "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])
        : "r"(taddr) : "memory");
    uint32_t a = 0;
#pragma unroll
    for (int i = 0; i < 64; i += 8) a ^= r[i];
    return a;
}
// MODE 3: 16x256b.x4 (16 lanes x 32 columns per instruction... 4 x 256 bits per lane pair) -> 16 registers
template <> __device__ __forceinline__ uint32_t ld_and_fold<3>(uint32_t taddr) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x4.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
    uint32_t a = 0;
#pragma unroll
    for (int i = 0; i < 16; i += 4) a ^= r[i];
    return a;
}

template <int MODE>
__global__ void __launch_bounds__(384) tmem_read_kernel(int iters, unsigned long long* cycles, uint32_t* sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = slot;
    uint32_t acc = 0;
    long long t0 = 0, t1 = 0;
    if (warp >= 4) {
        const int q = warp & 3, half = (warp - 4) >> 2;
        const uint32_t lane_base = base + ((uint32_t)(q * 32) << 16);
        t0 = clock64();
        for (int i = 0; i < iters; ++i) {
            // one "tile": this warp's 128 columns of a 256-column accumulator (two accumulators alternate)
            const uint32_t col0 = (uint32_t)((i & 1) * 256 + half * 128);
            if (MODE == 0 || MODE == 3) {
#pragma unroll
                for (int c = 0; c < 4; ++c) acc ^= ld_and_fold<MODE>(lane_base + col0 + c * 32);
            } else {   // 64 columns per instruction
#pragma unroll
                for (int c = 0; c < 2; ++c) acc ^= ld_and_fold<MODE>(lane_base + col0 + c * 64);
            }
        }
        t1 = clock64();
    }
    if (threadIdx.x == 128) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
    if (acc == 0x12345678u) sink[0] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(512) : "memory");
}

template <int MODE> void run(const char* name, double cols_per_warp_iter) {
    unsigned long long* cyc; uint32_t* sink;
    cudaMalloc(&cyc, 148 * 8); cudaMalloc(&sink, 4);
    const int iters = 20000;
    tmem_read_kernel<MODE><<<148, 384>>>(iters, cyc, sink);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    tmem_read_kernel<MODE><<<148, 384>>>(iters, cyc, sink);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    unsigned long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    // 8 warps x 32 lanes x cols columns of 32-bit TMEM cells per iteration and SM
    const double cells = 8.0 * 32.0 * cols_per_warp_iter * iters;
    printf("%-28s %s  %.3f ms  %.1f TMEM cells/clk/SM (%.1f B/clk/SM of 32-bit cells)  %.1f Gcells/s/SM\n", name,
           cudaGetErrorString(err), ms, cells / (double)h[0], 4.0 * cells / (double)h[0], cells / (ms * 1e6));
    cudaFree(cyc); cudaFree(sink);
}

int main() {
    run<0>("32x32b.x32", 128);
    run<2>("32x32b.x64", 128);
    run<1>("32x32b.x32.pack::16b", 128);
    run<3>("16x256b.x4 (half the lanes)", 128);
    return 0;
}
