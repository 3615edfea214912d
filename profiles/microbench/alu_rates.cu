// Microbenchmark: issue rates of FMNMX3 / FMNMX / FSETP / FADD(|x|) on sm_100a, 16 warps per SM (4 per scheduler).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(int iters, float* out, long long* cyc, float seed)
{
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = seed * (threadIdx.x + i);
    float m0 = 0.f, m1 = 0.f, m2 = 0.f, m3 = 0.f, m4 = 0.f, m5 = 0.f, m6 = 0.f, m7 = 0.f;
    bool p = false;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {  // 3-input max with abs: 8 independent chains, 1 op each per iteration
            m0 = fmaxf(m0, fmaxf(fabsf(a[0]), fabsf(a[1])));  m1 = fmaxf(m1, fmaxf(fabsf(a[2]), fabsf(a[3])));
            m2 = fmaxf(m2, fmaxf(fabsf(a[4]), fabsf(a[5])));  m3 = fmaxf(m3, fmaxf(fabsf(a[6]), fabsf(a[7])));
            m4 = fmaxf(m4, fmaxf(fabsf(a[8]), fabsf(a[9])));  m5 = fmaxf(m5, fmaxf(fabsf(a[10]), fabsf(a[11])));
            m6 = fmaxf(m6, fmaxf(fabsf(a[12]), fabsf(a[13]))); m7 = fmaxf(m7, fmaxf(fabsf(a[14]), fabsf(a[15])));
        } else if (MODE == 1) {  // 2-input max with abs
            m0 = fmaxf(m0, fabsf(a[0])); m1 = fmaxf(m1, fabsf(a[2])); m2 = fmaxf(m2, fabsf(a[4])); m3 = fmaxf(m3, fabsf(a[6]));
            m4 = fmaxf(m4, fabsf(a[8])); m5 = fmaxf(m5, fabsf(a[10])); m6 = fmaxf(m6, fabsf(a[12])); m7 = fmaxf(m7, fabsf(a[14]));
        } else if (MODE == 2) {  // compare with predicate accumulate
            p |= fabsf(a[0]) >= m0; p |= fabsf(a[1]) >= m0; p |= fabsf(a[2]) >= m0; p |= fabsf(a[3]) >= m0;
            p |= fabsf(a[4]) >= m0; p |= fabsf(a[5]) >= m0; p |= fabsf(a[6]) >= m0; p |= fabsf(a[7]) >= m0;
        } else {  // fadd with abs (fma pipe)
            m0 += fabsf(a[0]); m1 += fabsf(a[2]); m2 += fabsf(a[4]); m3 += fabsf(a[6]);
            m4 += fabsf(a[8]); m5 += fabsf(a[10]); m6 += fabsf(a[12]); m7 += fabsf(a[14]);
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) asm volatile("" : "+f"(a[i]));  // keep the loop body from being hoisted
        asm volatile("" : "+f"(m0), "+f"(m1), "+f"(m2), "+f"(m3), "+f"(m4), "+f"(m5), "+f"(m6), "+f"(m7));
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = m0 + m1 + m2 + m3 + m4 + m5 + m6 + m7 + (p ? 1.f : 0.f);
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main()
{
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&cyc, 148 * 8);
    const int iters = 20000;
    const char* names[4] = {"FMNMX3 |a|,|b| (8 per iter)", "FMNMX |a| (8 per iter)", "FSETP.OR |a| (8 per iter)", "FADD |a| (8 per iter)"};
    for (int mode = 0; mode < 4; ++mode) {
        if (mode == 0) k<0><<<148, 512>>>(iters, out, cyc, 1.5f);
        if (mode == 1) k<1><<<148, 512>>>(iters, out, cyc, 1.5f);
        if (mode == 2) k<2><<<148, 512>>>(iters, out, cyc, 1.5f);
        if (mode == 3) k<3><<<148, 512>>>(iters, out, cyc, 1.5f);
        cudaDeviceSynchronize();
        long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        // 16 warps per SM = 4 per scheduler; ops per scheduler per iteration = 4 warps * 8
        printf("%-30s cycles/iter=%.2f  -> %.2f cycles per warp-instruction per scheduler\n", names[mode], (double)h / iters, (double)h / iters / 32.0);
    }
    return 0;
}
