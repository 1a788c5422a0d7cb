// Micro-benchmark: tcgen05.ld (TMEM -> registers) throughput per SM on sm_100a, by shape and warp count.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench_tmem profiles/ubench_tmem.cu
// Output: bytes per SM clock for 4 / 8 / 16 warps issuing back-to-back loads of their own lane quarter.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

#define REGS32(v)                                                                                              \
    "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),   \
            "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), \
            "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),           \
            "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])

// SHAPE 0: 32x32b.x32 (32 lanes x 32 columns, 4 KB)   SHAPE 1: 16x256b.x4 (16 lanes x 8 cols x4 ... 32 regs)
// SHAPE 2: 32x32b.x16 (2 KB)                            SHAPE 3: 16x128b.x8 (32 regs)
template<int SHAPE>
__device__ __forceinline__ void ld(uint32_t taddr, uint32_t* v) {
    if (SHAPE == 0)
        asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                : REGS32(v)
                : "r"(taddr)
                : "memory");
    else if (SHAPE == 1)
        asm volatile(
                "tcgen05.ld.sync.aligned.16x256b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                : REGS32(v)
                : "r"(taddr)
                : "memory");
    else if (SHAPE == 2)
        asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                  "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                : "r"(taddr)
                : "memory");
    else
        asm volatile(
                "tcgen05.ld.sync.aligned.16x128b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                : REGS32(v)
                : "r"(taddr)
                : "memory");
}

template<int SHAPE>
__global__ void __launch_bounds__(512, 1) k(int iters, unsigned long long* out, uint32_t* sink) {
    __shared__ uint32_t slot;
    const int           warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16) + ((warp >> 2) * 64) % 512;
    uint32_t       v[32], acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        ld<SHAPE>(base, v);
        ld<SHAPE>(base + 32, v);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        acc ^= v[0] ^ v[7] ^ v[15];
    }
    const long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0)
        out[blockIdx.x] = (unsigned long long)(t1 - t0);
    if (acc == 0x12345678u)
        sink[threadIdx.x] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(slot));
}

template<int SHAPE>
void run(const char* name, int bytesPerLd) {
    unsigned long long* d;
    uint32_t*           sink;
    cudaMalloc(&d, 148 * 8);
    cudaMalloc(&sink, 4096);
    const int iters = 4000;
    for (int warps : {4, 8, 16}) {
        k<SHAPE><<<148, warps * 32>>>(iters, d, sink);
        k<SHAPE><<<148, warps * 32>>>(iters, d, sink);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
            printf("%s: %s\n", name, cudaGetErrorString(e));
            return;
        }
        unsigned long long h[148];
        cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        double cyc = 0;
        for (int i = 0; i < 148; ++i)
            cyc += h[i];
        cyc /= 148;
        const double bytes = (double)iters * 2 * bytesPerLd * warps;
        printf("%-14s warps %2d : %.1f bytes / clk / SM  (%.1f clk per load per warp)\n", name, warps, bytes / cyc,
               cyc / (iters * 2.0));
    }
    cudaFree(d);
    cudaFree(sink);
}

int main() {
    run<0>("32x32b.x32", 4096);
    run<2>("32x32b.x16", 2048);
    run<1>("16x256b.x8", 4096);
    run<3>("16x128b.x16", 4096);
    return 0;
}
