// pipes.cu — issue-rate microbenchmark for the instruction classes of the BVH node visit on sm_100a:
// FFMA, FFMA2 (fma.rn.f32x2), FMNMX, FMNMX3, FSETP+SEL, and mixes. Prints warp-instructions per cycle per SM sub-partition.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu ; run on the GPU box (tools/gpu_ubench.sh).
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long f2pack(float a, float b)
{
    return ((unsigned long long)__float_as_uint(b) << 32) | __float_as_uint(a);
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c)
{
    unsigned long long d;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
#define N_ACC 8
#define ITERS 4096

template <int MODE> __global__ void __launch_bounds__(1024, 1) k(float* out, float s, float t, long long* cycles)
{
    float a[N_ACC];
    unsigned long long p[N_ACC];
    for (int i = 0; i < N_ACC; i++) { a[i] = threadIdx.x * 0.001f + i; p[i] = f2pack(a[i], a[i] + 1.f); }
    const unsigned long long ps = f2pack(s, s), pt = f2pack(t, t);
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < N_ACC; i++) {
            if (MODE == 0) a[i] = fmaf(a[i], s, t);                                        // FFMA
            if (MODE == 1) p[i] = fma2(p[i], ps, pt);                                      // FFMA2
            if (MODE == 2) asm volatile("max.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(s));      // FMNMX
            if (MODE == 3) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(s), "f"(t)); // FMNMX3
            if (MODE == 4) { p[i] = fma2(p[i], ps, pt); asm volatile("max.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(s)); } // 1:1 FFMA2 : FMNMX
            if (MODE == 5) { a[i] = fmaf(a[i], s, t); asm volatile("max.f32 %0, %0, %1;" : "+f"(a[(i + 4) % N_ACC]) : "f"(s)); } // FFMA : FMNMX... same regs
            if (MODE == 6) { p[i] = fma2(p[i], ps, pt); p[i] = fma2(p[i], pt, ps); asm volatile("max.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(s)); } // 2:1
            if (MODE == 7) { asm volatile("{.reg .pred q; setp.lt.f32 q, %0, %1; selp.f32 %0, %2, %0, q;}" : "+f"(a[i]) : "f"(s), "f"(t)); } // FSETP + SEL
            if (MODE == 9) { asm volatile("{.reg .pred q; setp.lt.f32 q, %0, %1; @q add.f32 %0, %0, %2;}" : "+f"(a[i]) : "f"(s), "f"(t)); } // FSETP + predicated FADD
            if (MODE == 11) { unsigned u = __float_as_uint(a[i]); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u) : "r"(__float_as_uint(s)), "r"(__float_as_uint(t))); a[i] = __uint_as_float(u); }
            if (MODE == 12) { unsigned u = __float_as_uint(a[i]); asm volatile("add.u32 %0, %0, %1;" : "+r"(u) : "r"(__float_as_uint(s))); a[i] = __uint_as_float(u); }
            if (MODE == 13) { unsigned u = __float_as_uint(a[i]); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(u) : "r"(__float_as_uint(s)), "r"(__float_as_uint(t))); a[i] = __uint_as_float(u); }
            if (MODE == 14) { unsigned u = __float_as_uint(a[i]); asm volatile("{.reg .pred q; setp.lt.s32 q, %0, %1; selp.u32 %0, %1, %0, q;}" : "+r"(u) : "r"(__float_as_uint(s))); a[i] = __uint_as_float(u); } // ISETP + SEL
            if (MODE == 15) { unsigned u = __float_as_uint(a[i]); asm volatile("max.s32 %0, %0, %1;" : "+r"(u) : "r"(__float_as_uint(s))); a[i] = __uint_as_float(u); } // IMNMX
            if (MODE == 16) { asm volatile("add.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(s)); }
            if (MODE == 17) { asm volatile("mul.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(s)); }
            if (MODE == 18) { asm volatile("{.reg .pred q; setp.lt.f32 q, %0, %1; @q mov.f32 %0, %2;}" : "+f"(a[i]) : "f"(s), "f"(t)); } // FSETP + predicated MOV
            if (MODE == 19) { unsigned u = __float_as_uint(a[i]); asm volatile("shf.l.wrap.b32 %0, %0, %1, 3;" : "+r"(u) : "r"(__float_as_uint(s))); a[i] = __uint_as_float(u); }
            if (MODE == 20) { asm volatile("{.reg .pred q; setp.lt.f32 q, %0, %1; setp.lt.and.f32 q, %0, %2, q; @q mov.f32 %0, %2;}" : "+f"(a[i]) : "f"(s), "f"(t)); } // 2 FSETP + MOV
            if (MODE == 8) { a[i] = fmaf(a[i], s, t); a[i] = fmaf(a[i], t, s); asm volatile("max.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(s)); } // FFMA 2:1 FMNMX
        }
    }
    const long long t1 = clock64();
    float acc = 0.f;
    for (int i = 0; i < N_ACC; i++) acc += a[i] + __uint_as_float((unsigned)p[i]) + __uint_as_float((unsigned)(p[i] >> 32));
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE> void run(const char* name, int per_iter)
{
    float* out;
    long long* cyc;
    cudaMalloc(&out, 148 * 1024 * sizeof(float));
    cudaMalloc(&cyc, 148 * sizeof(long long));
    k<MODE><<<148, 1024>>>(out, 1.0001f, 0.5f, cyc);
    k<MODE><<<148, 1024>>>(out, 1.0001f, 0.5f, cyc);
    cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double c = 0;
    for (int i = 0; i < 148; i++) c += (double)h[i];
    c /= 148;
    // warp-instructions per SMSP: 8 warps per SMSP x ITERS x N_ACC x per_iter
    const double wi = 8.0 * ITERS * N_ACC * per_iter;
    printf("%-28s %8.0f cycles  %.3f warp-instr/cycle/SMSP\n", name, c, wi / c);
    cudaFree(out);
    cudaFree(cyc);
}

int main()
{
    run<0>("FFMA", 1);
    run<1>("FFMA2", 1);
    run<2>("FMNMX", 1);
    run<3>("FMNMX3", 1);
    run<4>("FFMA2 + FMNMX", 2);
    run<5>("FFMA + FMNMX", 2);
    run<6>("2 FFMA2 + FMNMX", 3);
    run<7>("FSETP + SEL", 2);
    run<8>("2 FFMA + FMNMX", 3);
    run<9>("FSETP + @p FADD", 2);
    run<11>("LOP3", 1);
    run<12>("IADD", 1);
    run<13>("IMAD", 1);
    run<14>("ISETP + SEL", 2);
    run<15>("IMNMX", 1);
    run<16>("FADD", 1);
    run<17>("FMUL", 1);
    run<18>("FSETP + @p MOV", 2);
    run<19>("SHF", 1);
    run<20>("2 FSETP + @p MOV", 3);
    return 0;
}
