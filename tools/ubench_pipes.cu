// Micro-benchmark: issue throughput (warp-instructions per cycle per SM sub-partition) of the ALU /
// FMA-pipe instructions the tensor kernel's epilogue is made of.  Bring-up tool, not part of the library.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_pipes tools/ubench_pipes.cu && ./ubench_pipes
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_fp16.h>

#define CHAINS 8
#define ITERS 4096

template <int OP>
__global__ void bench(float* out, unsigned long long* cycles, float seed) {
    float v[CHAINS];
    unsigned u[CHAINS];
    __half2 h[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) { v[c] = seed + c + threadIdx.x; u[c] = __float_as_uint(v[c]); h[c] = __floats2half2_rn(v[c], v[c] + 1.f); }
    float w = seed * 1.0001f, z = seed * 0.37f;
    unsigned uw = __float_as_uint(w), uz = __float_as_uint(z);
    __half2 hw = __floats2half2_rn(w, z);
    __syncthreads();
    unsigned long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < ITERS; ++i) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            if (OP == 0) v[c] = fminf(v[c], w + c);                      // FADD + FMNMX
            if (OP == 1) v[c] = fminf(fminf(v[c], w), z);                // FMNMX3
            if (OP == 2) v[c] = v[c] + w;                                // FADD
            if (OP == 3) u[c] = (u[c] & uw) | uz;                        // LOP3
            if (OP == 4) u[c] = (unsigned)__vimin3_s32((int)u[c], (int)uw, (int)uz);   // VIMNMX3
            if (OP == 5) h[c] = __hmin2(h[c], hw);                       // HMNMX2
            if (OP == 6) u[c] = min(u[c], uw) ;                          // IMNMX / VIMNMX
            if (OP == 7) v[c] = fminf(v[c], w);                          // FMNMX
            if (OP == 8) u[c] = u[c] * uw + uz;                          // IMAD
            if (OP == 9) u[c] = __vimin3_s16x2(u[c], uw, uz);            // VIMNMX3.S16x2
            if (OP == 10) u[c] = __vminu2(u[c], uw);                     // VIMNMX.U16x2 (or emulated)
            if (OP == 11) v[c] = fmaf(v[c], w, z);                       // FFMA
            if (OP == 12) u[c] = (v[c] < w) ? u[c] : uz, v[c] += 1.f;    // FSETP + SEL + FADD
            if (OP == 13) { float2 t = __fadd2_rn(make_float2(v[c], __uint_as_float(u[c])), make_float2(w, z)); v[c] = t.x; u[c] = __float_as_uint(t.y); }   // FADD2
            if (OP == 15) h[c] = __hmin2(__hmin2(h[c], hw), __halves2half2(__low2half(hw), __high2half(h[(c + 1) % CHAINS])));   // VHMNMX (3-input half2 min)
            if (OP == 16) { h[c] = __hmin2(h[c], hw); h[c] = __hmin2_nan(h[c], __halves2half2(__low2half(hw), __high2half(h[(c + 1) % CHAINS]))); }   // HMNMX2 + HMNMX2.NAN: two instructions the assembler cannot fuse
            if (OP == 14) { float2 t = __ffma2_rn(make_float2(v[c], __uint_as_float(u[c])), make_float2(w, z), make_float2(z, w)); v[c] = t.x; u[c] = __float_as_uint(t.y); }   // FFMA2
        }
        if (OP == 0 || OP == 7 || OP == 1) { w += 1e-7f; }
        if (OP == 3 || OP == 4 || OP == 6 || OP == 9 || OP == 10) { uw += 1; }
    }
    unsigned long long t1 = clock64();
    float acc = 0.f;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) acc += v[c] + __uint_as_float(u[c] & 0x3fffffff) + __low2float(h[c]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, int warps) {
    float* out; unsigned long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    bench<OP><<<148, warps * 32>>>(out, cyc, 1.5f);
    bench<OP><<<148, warps * 32>>>(out, cyc, 1.5f);
    cudaDeviceSynchronize();
    unsigned long long h[148]; cudaMemcpy(h, cyc, 148 * 8, cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    double inst = (double)ITERS * CHAINS * warps;       // warp-instructions of the op per SM
    printf("%-28s warps/SM=%2d  cycles=%9.0f  warp-inst/clk/SM=%.3f  per-SMSP=%.3f\n", name, warps, avg, inst / avg, inst / avg / 4);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int warps : {4, 8, 16}) {
        run<2>("FADD", warps); run<11>("FFMA", warps); run<7>("FMNMX", warps); run<1>("FMNMX3", warps);
        run<0>("FADD+FMNMX", warps); run<3>("LOP3", warps); run<6>("IMNMX(u32)", warps); run<4>("VIMNMX3.s32", warps);
        run<9>("VIMNMX3.s16x2", warps); run<10>("vminu2", warps); run<5>("HMNMX2", warps); run<8>("IMAD", warps);
        run<15>("VHMNMX (min3 half2)", warps); run<16>("HMNMX2 + HMNMX2.NAN pair", warps);
        run<12>("FSETP+SEL+FADD", warps); run<13>("FADD2 (2 adds/inst)", warps); run<14>("FFMA2 (2 fmas/inst)", warps);
        printf("\n");
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
