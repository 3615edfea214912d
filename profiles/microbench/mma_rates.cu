// Microbenchmark: issue rate of tcgen05.mma (cta_group::1, M = 128, N = 256) per SM for kind::f16 (K = 16 per
// instruction), kind::i8 (K = 32) and kind::f8f6f4 (K = 32).  Operands are whatever shared memory holds: the tensor pipe's
// timing does not depend on the values.  One CTA per SM, one thread issues `chain` instructions back to back into one
// accumulator stage, commits to an mbarrier and waits; the clock is read around `reps` such chains.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rates mma_rates.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t sbo_bytes, uint32_t layout_type)
{
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)layout_type << 61;
    return d;
}

template <int KIND>
__device__ __forceinline__ void mma(uint32_t d_tmem, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc)
{
    if (KIND == 0)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                     "l"(a), "l"(b), "r"(idesc), "r"(acc)
                     : "memory");
    if (KIND == 1)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                     "l"(a), "l"(b), "r"(idesc), "r"(acc)
                     : "memory");
    if (KIND == 2)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                     "l"(a), "l"(b), "r"(idesc), "r"(acc)
                     : "memory");
}

template <int KIND>
__global__ void __launch_bounds__(128) bench(int chain, int reps, long long* cycles)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t slot;
    __shared__ __align__(8) uint64_t bar;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 48 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = slot;
    if (threadIdx.x == 0) {
        const uint32_t sa = (uint32_t)__cvta_generic_to_shared(smem);
        // A: 128 rows x 128 B (SWIZZLE_128B, 8-row groups 1024 B apart) = 16 KB; B: 256 rows x 128 B = 32 KB
        const uint64_t adesc = make_desc(sa, 1024, 2), bdesc = make_desc(sa + 16 * 1024, 1024, 2);
        // c_format at bit 4 (F32 = 1, S32 = 2), a/b format at bits 7 / 10 (F16 = 0; UINT8 = 0; E4M3 = 0), N>>3 at 17, M>>4 at 24
        const uint32_t cfmt = KIND == 1 ? 2u : 1u;
        const uint32_t idesc = (cfmt << 4) | ((256u >> 3) << 17) | ((128u >> 4) << 24);
        const uint32_t barp = (uint32_t)__cvta_generic_to_shared(&bar);
        uint32_t parity = 0;
        const long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            for (int k = 0; k < chain; ++k) mma<KIND>(base + (uint32_t)((r & 1) * 256), adesc, bdesc, idesc, k > 0);
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(barp) : "memory");
            uint32_t ok = 0;
            const long long w0 = clock64();
            while (!ok && clock64() - w0 < 400000000ll)   // bounded: a faulting instruction must not hang the GPU
                asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
                             : "=r"(ok)
                             : "r"(barp), "r"(parity)
                             : "memory");
            parity ^= 1;
        }
        cycles[blockIdx.x] = clock64() - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(512));
}

template <int KIND>
static void run(const char* name, int k_per_instr, int sms, int clock_khz)
{
    long long* d;
    cudaMalloc(&d, sizeof(long long) * sms);
    cudaFuncSetAttribute(bench<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    const int chain = 64, reps = 200;
    bench<KIND><<<sms, 128, 64 * 1024>>>(chain, 4, d);   // warm-up
    bench<KIND><<<sms, 128, 64 * 1024>>>(chain, reps, d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    long long* h = new long long[sms];
    cudaMemcpy(h, d, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    long long worst = 0;
    for (int i = 0; i < sms; ++i) worst = h[i] > worst ? h[i] : worst;
    const double per_instr = (double)worst / ((double)chain * reps);
    const double ops = 2.0 * 128 * 256 * k_per_instr;
    printf("%-8s K=%2d per instruction: %.1f cycles per instruction, %.0f dense (fl)op/cycle/SM, %.0f T(FL)OP/s on %d SMs at %.2f GHz\n",
           name, k_per_instr, per_instr, ops / per_instr, ops / per_instr * sms * clock_khz * 1e3 / 1e12, sms, clock_khz / 1e6);
    cudaFree(d);
    delete[] h;
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    run<0>("f16", 16, p.multiProcessorCount, khz);
    run<1>("i8", 32, p.multiProcessorCount, khz);
    run<2>("f8f6f4", 32, p.multiProcessorCount, khz);
    return 0;
}
