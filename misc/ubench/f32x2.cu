// Issue-rate probe: does FADD2/FMUL2 (add/mul.f32x2) free issue slots next to ALU work on sm_100a?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o f32x2 f32x2.cu && ./f32x2
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float add1(float a, float b) { float r; asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ unsigned lop(unsigned a, unsigned b) { unsigned r; asm volatile("lop3.b32 %0, %1, %2, 0x80000000, 0xD2;" : "=r"(r) : "r"(a), "r"(b)); return r; }

template <int MODE> __global__ void k(float* out, int iters, float s) {
    float a[8]; u64 p[4]; unsigned q[4];
    for (int i = 0; i < 8; i++) a[i] = threadIdx.x * 0.001f + i;
    for (int i = 0; i < 4; i++) { p[i] = pk(a[2 * i], a[2 * i + 1]); q[i] = threadIdx.x + i; }
    const u64 s2 = pk(s, s);
    for (int it = 0; it < iters; it++) {
        if (MODE == 0) { for (int i = 0; i < 8; i++) a[i] = add1(a[i], s); }                       // 8 FADD
        if (MODE == 1) { for (int i = 0; i < 4; i++) p[i] = add2(p[i], s2); }                      // 4 FADD2 (same flops)
        if (MODE == 2) { for (int i = 0; i < 8; i++) a[i] = add1(a[i], s); for (int i = 0; i < 4; i++) { q[i] = lop(q[i], it); q[i] = lop(q[i], it + 1); } }   // 8 FADD + 8 LOP3
        if (MODE == 3) { for (int i = 0; i < 4; i++) p[i] = add2(p[i], s2); for (int i = 0; i < 4; i++) { q[i] = lop(q[i], it); q[i] = lop(q[i], it + 1); } }  // 4 FADD2 + 8 LOP3
        if (MODE == 4) { for (int i = 0; i < 4; i++) { q[i] = lop(q[i], it); q[i] = lop(q[i], it + 1); } }   // 8 LOP3
    }
    float r = 0; for (int i = 0; i < 8; i++) r += a[i];
    for (int i = 0; i < 4; i++) { float x, y; asm("mov.b64 {%0,%1}, %2;" : "=f"(x), "=f"(y) : "l"(p[i])); r += x + y + q[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE> void run(const char* name, float* d) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    k<MODE><<<148 * 8, 256>>>(d, 100, 1.0f);
    cudaEventRecord(e0); k<MODE><<<148 * 8, 256>>>(d, iters, 1.0f); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    // warp-iterations per SMSP: 8 CTAs x 8 warps / 4 SMSP = 16 warps per SMSP
    const double cyc = ms * 1e-3 * 1.965e9 / (16.0 * iters);
    printf("%-22s %.3f ms  -> %.2f issue-cycles per warp-iteration (at 1.965 GHz)\n", name, ms, cyc);
}
int main() {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    run<0>("8 FADD", d); run<1>("4 FADD2", d); run<2>("8 FADD + 8 LOP3", d); run<3>("4 FADD2 + 8 LOP3", d); run<4>("8 LOP3", d);
    return 0;
}
