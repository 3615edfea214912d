// Microbenchmark: TMEM -> register read bandwidth per SM for several tcgen05.ld shapes and warp counts.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_ld_bw tmem_ld_bw.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define LD32(shape, regs)                                                                                              \
    asm volatile("tcgen05.ld.sync.aligned." shape ".b32 "                                                               \
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                              \
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"              \
                 : "=r"(regs[0]), "=r"(regs[1]), "=r"(regs[2]), "=r"(regs[3]), "=r"(regs[4]), "=r"(regs[5]),            \
                   "=r"(regs[6]), "=r"(regs[7]), "=r"(regs[8]), "=r"(regs[9]), "=r"(regs[10]), "=r"(regs[11]),          \
                   "=r"(regs[12]), "=r"(regs[13]), "=r"(regs[14]), "=r"(regs[15]), "=r"(regs[16]), "=r"(regs[17]),      \
                   "=r"(regs[18]), "=r"(regs[19]), "=r"(regs[20]), "=r"(regs[21]), "=r"(regs[22]), "=r"(regs[23]),      \
                   "=r"(regs[24]), "=r"(regs[25]), "=r"(regs[26]), "=r"(regs[27]), "=r"(regs[28]), "=r"(regs[29]),      \
                   "=r"(regs[30]), "=r"(regs[31])                                                                      \
                 : "r"(taddr))

template <int SHAPE>
__global__ void bench(int iters, int warps_used, long long* cycles, uint32_t* sink)
{
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = slot;
    uint32_t acc = 0;
    long long t0 = 0, t1 = 0;
    __syncthreads();
    if (warp < warps_used) {
        const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
        t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            uint32_t r[32];
            const uint32_t col = (uint32_t)(((it * 32) + (warp >> 2) * 128) & 511);
            uint32_t taddr = base + lane_base + col;
            if (SHAPE == 0) { LD32("32x32b.x32", r); }
            if (SHAPE == 1) { LD32("16x256b.x8", r); }
            if (SHAPE == 2) { LD32("16x128b.x16", r); }
            if (SHAPE == 3) { LD32("16x64b.x32", r); }
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int k = 0; k < 32; ++k) acc ^= r[k];
        }
        t1 = clock64();
    }
    __syncthreads();
    if (threadIdx.x % 32 == 0 && warp < warps_used) cycles[blockIdx.x * 32 + warp] = t1 - t0;
    if (acc == 0x12345678u) sink[0] = acc;
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(512));
}

// two loads in flight before each wait
template <int SHAPE>
__global__ void bench2(int iters, int warps_used, long long* cycles, uint32_t* sink)
{
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = slot;
    uint32_t acc = 0;
    long long t0 = 0, t1 = 0;
    __syncthreads();
    if (warp < warps_used) {
        const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
        t0 = clock64();
        for (int it = 0; it < iters; it += 2) {
            uint32_t r[32], q[32];
            uint32_t taddr = base + lane_base + (uint32_t)((it * 32) & 511);
            if (SHAPE == 0) { LD32("32x32b.x32", r); }
            if (SHAPE == 1) { LD32("16x256b.x8", r); }
            taddr = base + lane_base + (uint32_t)((it * 32 + 32) & 511);
            if (SHAPE == 0) { LD32("32x32b.x32", q); }
            if (SHAPE == 1) { LD32("16x256b.x8", q); }
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int k = 0; k < 32; ++k) acc ^= r[k] ^ q[k];
        }
        t1 = clock64();
    }
    __syncthreads();
    if (threadIdx.x % 32 == 0 && warp < warps_used) cycles[blockIdx.x * 32 + warp] = t1 - t0;
    if (acc == 0x12345678u) sink[0] = acc;
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(512));
}

int main()
{
    long long* d_cycles; uint32_t* d_sink;
    cudaMalloc(&d_cycles, 148 * 32 * 8); cudaMalloc(&d_sink, 4);
    const int iters = 4096;
    const char* names[4] = {"32x32b.x32 (4096 B/warp-instr)", "16x256b.x8 (4096 B)", "16x128b.x16 (4096 B)", "16x64b.x32 (4096 B)"};
    for (int shape = 0; shape < 4; ++shape)
        for (int warps : {1, 4, 8, 16}) {
            cudaMemset(d_cycles, 0, 148 * 32 * 8);
            if (shape == 0) bench<0><<<148, 512>>>(iters, warps, d_cycles, d_sink);
            if (shape == 1) bench<1><<<148, 512>>>(iters, warps, d_cycles, d_sink);
            if (shape == 2) bench<2><<<148, 512>>>(iters, warps, d_cycles, d_sink);
            if (shape == 3) bench<3><<<148, 512>>>(iters, warps, d_cycles, d_sink);
            cudaError_t e = cudaDeviceSynchronize();
            long long h[32];
            cudaMemcpy(h, d_cycles, sizeof(h), cudaMemcpyDeviceToHost);
            long long mx = 0; for (int w = 0; w < warps; ++w) mx = h[w] > mx ? h[w] : mx;
            double bytes = (double)iters * warps * 4096.0;
            printf("%-34s warps=%2d  %s  cycles=%lld  B/clk/SM=%.1f  cyc/instr/warp=%.1f\n", names[shape], warps, cudaGetErrorString(e), mx, bytes / mx, (double)mx / iters);
        }
    for (int shape = 0; shape < 2; ++shape)
        for (int warps : {4, 8, 16}) {
            cudaMemset(d_cycles, 0, 148 * 32 * 8);
            if (shape == 0) bench2<0><<<148, 512>>>(iters, warps, d_cycles, d_sink);
            if (shape == 1) bench2<1><<<148, 512>>>(iters, warps, d_cycles, d_sink);
            cudaError_t e = cudaDeviceSynchronize();
            long long h[32];
            cudaMemcpy(h, d_cycles, sizeof(h), cudaMemcpyDeviceToHost);
            long long mx = 0; for (int w = 0; w < warps; ++w) mx = h[w] > mx ? h[w] : mx;
            double bytes = (double)iters * warps * 4096.0;
            printf("2-in-flight %-22s warps=%2d  %s  cycles=%lld  B/clk/SM=%.1f\n", names[shape], warps, cudaGetErrorString(e), mx, bytes / mx);
        }
    return 0;
}
