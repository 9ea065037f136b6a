// EXPERIMENT: issue / pipe rates of the instructions the traversal loop is made of (B200, sm_100a).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/pipes tools/exp/pipes.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 2048
#define UNR 8

template <int MODE>
__global__ void __launch_bounds__(1024) k(float *out, float a0, float b0, float one, unsigned long long *cyc)
{
    float x[UNR], y[UNR];
#pragma unroll
    for (int i = 0; i < UNR; i++) { x[i] = a0 + threadIdx.x * 1e-3f + i; y[i] = b0 + i * 0.5f; }
    unsigned long long p[UNR];
#pragma unroll
    for (int i = 0; i < UNR; i++) asm("mov.b64 %0, {%1, %2};" : "=l"(p[i]) : "f"(x[i]), "f"(y[i]));
    unsigned long long cb, ob;
    asm("mov.b64 %0, {%1, %1};" : "=l"(cb) : "f"(b0));
    asm("mov.b64 %0, {%1, %1};" : "=l"(ob) : "f"(one));
    long long t0 = clock64();
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < UNR; i++) {
            if (MODE == 0) { x[i] = __fmul_rn(x[i], b0); }                          // FMUL
            if (MODE == 1) { x[i] = __fmul_rn(x[i], b0); x[i] = __fadd_rn(x[i], a0); }   // FMUL + FADD
            if (MODE == 2) { x[i] = __fmaf_rn(x[i], b0, a0); }                      // FFMA
            if (MODE == 3) { asm("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(cb), "l"(ob)); }   // FFMA2
            if (MODE == 4) { x[i] = fminf(x[i], y[i]); y[i] = fmaxf(y[i], b0); }    // 2 FMNMX
            if (MODE == 5) { asm("min.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(y[i]), "f"(b0)); }     // FMNMX3
            if (MODE == 6) { x[i] = x[i] <= y[i] ? x[i] : a0; }                     // FSETP + FSEL
            if (MODE == 7) { x[i] = __fmul_rn(x[i], b0); y[i] = fminf(y[i], x[i]); }  // FMUL + FMNMX (two pipes)
            if (MODE == 8) { asm("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(cb), "l"(ob)); y[i] = fminf(y[i], a0 + i); y[i] = fmaxf(y[i], b0); }  // FFMA2 + 2 FMNMX
            if (MODE == 9) { unsigned u = __float_as_uint(x[i]); u = (u & 0xFFFFu) ^ (u >> 3); x[i] = __uint_as_float(u | 0x3F800000u); }  // LOP3/SHF
        }
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < UNR; i++) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[i]));
        s += x[i] + y[i] + lo + hi;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = (unsigned long long)(t1 - t0);
}

template <int MODE>
void run(const char *name, int per_iter, float *out, unsigned long long *cyc)
{
    for (int warps : {4, 8, 16, 32}) {
        k<MODE><<<148, warps * 32>>>(out, 1.0001f, 0.9999f, 1.0f, cyc);
        cudaDeviceSynchronize();
        k<MODE><<<148, warps * 32>>>(out, 1.0001f, 0.9999f, 1.0f, cyc);
        cudaDeviceSynchronize();
        unsigned long long c;
        cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        double inst = (double)ITERS * UNR * per_iter * (warps / 4.0);       // warp-instructions per SMSP
        printf("%-28s warps/SM %2d: %.3f cycles per warp-instruction per SMSP (%.2f IPC/SMSP)\n", name, warps,
               c / inst, inst / c);
    }
}

int main()
{
    float *out; unsigned long long *cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
    run<0>("FMUL", 1, out, cyc);
    run<1>("FMUL+FADD", 2, out, cyc);
    run<2>("FFMA", 1, out, cyc);
    run<3>("FFMA2", 1, out, cyc);
    run<4>("FMNMX x2", 2, out, cyc);
    run<5>("FMNMX3", 1, out, cyc);
    run<6>("FSETP+FSEL", 2, out, cyc);
    run<7>("FMUL+FMNMX", 2, out, cyc);
    run<8>("FFMA2+2 FMNMX", 3, out, cyc);
    run<9>("LOP3/SHF x4", 4, out, cyc);
    printf("done %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
