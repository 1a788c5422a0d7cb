// Micro-benchmark: issue rate of tcgen05.mma kind::f16 (SS mode, operands in 128B-swizzled shared memory) on one SM.
// One thread issues `iters` groups of `perTile` MMAs (K = 16 each), commits to an mbarrier after every group, and
// waits for the last commit.  Reports SM clocks per MMA for several N; operand values are irrelevant (zeros).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o profiles/ubench_mma.bin profiles/ubench_mma.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ uint64_t desc128(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
            "l"(a), "l"(b), "r"(idesc), "r"(acc)
            : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
    asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
              "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
              "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
              "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(taddr)
            : "memory");
}

// mode 0: distinct k-slices like the scorer (9 per tile over two swizzle atoms); mode 1: same slice every time
// noise (what warps 4..19 do meanwhile, once per `pace` clocks): 0 nothing, 1 ALU only, 2 each warp reads 8 KB of
// TMEM, 3 each warp stores + loads 2 KB of shared memory, 4 one thread bulk-copies 32 KB global -> shared
__global__ void __launch_bounds__(640, 1) k(int iters, int perTile, int N, int M, int mode, unsigned long long* out,
                                            int noise, int pace, const uint8_t* gsrc, uint32_t* sink, int work) {
    extern __shared__ uint8_t raw[];
    uint8_t*                  smem = raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u);
    __shared__ uint32_t       slot;
    __shared__ uint64_t       bar, barLast, barBulk;
    __shared__ volatile int   stop;
    const int                 warp = threadIdx.x >> 5;
    for (uint32_t i = threadIdx.x; i < (64 + 32) * 1024 / 4; i += blockDim.x)
        reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&barLast)));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&barBulk)));
        stop = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (warp == 1 && mode == 2) {
        // lean issue loop of the scorer: warp converged, one elected lane, descriptors from warp-uniform values
        uint32_t elected;
        asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(elected));
        const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        const uint32_t bLo = (smem_u32(smem) & 0x3FFFF) >> 4, aLo = bLo + (64 * 1024 >> 4);
        const uint32_t hi = (uint32_t)(1024 >> 4) | (1u << 14) | (2u << 29);
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            const uint32_t d = slot + (it & 1) * 256;
            if (elected) {
                for (uint32_t ks = 0; ks < 3; ++ks) {
                    const uint32_t kl = 3 + ks;
                    const uint32_t ah = aLo + (ks / 4) * 1024 + (ks % 4) * 2, al = aLo + (kl / 4) * 1024 + (kl % 4) * 2;
                    const uint32_t bh = bLo + (ks / 4) * 2048 + (ks % 4) * 2, bl = bLo + (kl / 4) * 2048 + (kl % 4) * 2;
                    mma(d, ((uint64_t)hi << 32) | ah, ((uint64_t)hi << 32) | bh, idesc, ks ? 1u : 0u);
                    mma(d, ((uint64_t)hi << 32) | al, ((uint64_t)hi << 32) | bh, idesc, 1u);
                    mma(d, ((uint64_t)hi << 32) | ah, ((uint64_t)hi << 32) | bl, idesc, 1u);
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                                     smem_u32(it + 1 == iters ? &barLast : &bar))
                             : "memory");
            }
            __syncwarp();
        }
        const long long t1 = clock64();
        asm volatile(
                "{\n\t.reg .pred p;\n\tWL_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DL_%=;\n\tbra "
                "WL_%=;\n\tDL_%=:\n\t}" ::"r"(smem_u32(&barLast)),
                "r"(0u)
                : "memory");
        const long long t2 = clock64();
        if (elected) {
            out[blockIdx.x * 2]     = (unsigned long long)(t1 - t0);
            out[blockIdx.x * 2 + 1] = (unsigned long long)(t2 - t0);
        }
        __syncwarp();
        stop = 1;
    }
    else if (threadIdx.x == 32) {
        const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        const uint32_t bAddr = smem_u32(smem), aAddr = bAddr + 64 * 1024;
        const long long t0   = clock64();
        for (int it = 0; it < iters; ++it) {
            const uint32_t d = slot + (it & 1) * 256;
            for (int j = 0; j < perTile; ++j) {
                const uint32_t ks = mode ? 0 : (j % 6);
                const uint32_t ao = (ks / 4) * (128 * 128) + (ks % 4) * 32;
                const uint32_t bo = (ks / 4) * (256 * 128) + (ks % 4) * 32;
                mma(d, desc128(aAddr + ao), desc128(bAddr + bo), idesc, j ? 1u : 0u);
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                                 smem_u32(it + 1 == iters ? &barLast : &bar))
                         : "memory");
        }
        const long long t1 = clock64();
        // wait for the last commit (its own barrier, first phase)
        asm volatile(
                "{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra "
                "W_%=;\n\tD_%=:\n\t}" ::"r"(smem_u32(&barLast)),
                "r"(0u)
                : "memory");
        const long long t2 = clock64();
        out[blockIdx.x * 2]     = (unsigned long long)(t1 - t0);
        out[blockIdx.x * 2 + 1] = (unsigned long long)(t2 - t0);
        stop                    = 1;
    }
    else if (warp >= 4 && noise) {
        const int       e = warp - 4, lane = threadIdx.x & 31;
        uint8_t*        scratch = smem + 100 * 1024;  // 64 KB of noise area behind the operand tiles
        uint32_t        v[32], acc = 0, ph = 0;
        float           x = (float)threadIdx.x;
        long long       next = clock64();
        while (!stop) {
            if (noise >= 5) {  // sleep instead of spinning, then `work` dependent FMAs (4 chains)
                while (clock64() < next && !stop)
                    __nanosleep(100);
                next += pace;
                if (noise == 6 && (warp & 3) == 1)
                    continue;
                float y = x + 1.f, z = x + 2.f, w2 = x + 3.f;
                for (int i = 0; i < work; i += 4) {
                    x  = fmaf(x, 1.0001f, 0.5f);
                    y  = fmaf(y, 1.0001f, 0.5f);
                    z  = fmaf(z, 1.0001f, 0.5f);
                    w2 = fmaf(w2, 1.0001f, 0.5f);
                }
                x += y + z + w2;
                continue;
            }
            while (clock64() < next && !stop) {
            }
            next += pace;
            if (noise == 1) {
                for (int i = 0; i < 256; ++i)
                    x = fmaf(x, 1.0001f, 0.5f);
            }
            else if (noise == 2) {
                const uint32_t ta = slot + ((uint32_t)((warp & 3) * 32) << 16) + (e >> 2) * 64;
                tmem_ld32(ta, v);
                tmem_ld32(ta + 32, v);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                acc ^= v[3];
            }
            else if (noise == 3) {
                uint32_t* w = reinterpret_cast<uint32_t*>(scratch) + e * 1024 + lane;
                for (int i = 0; i < 8; ++i)
                    w[i * 32] = acc + i;
                __syncwarp();
                for (int i = 0; i < 8; ++i)
                    acc += w[i * 32 + 1 - 2 * (lane & 1)];
            }
            else if (noise == 4 && e == 0 && lane == 0) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&barBulk)), "r"(32768)
                             : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                     smem_u32(scratch)),
                             "l"(gsrc + (size_t)(ph & 63) * 32768), "r"(32768), "r"(smem_u32(&barBulk))
                             : "memory");
                asm volatile(
                        "{\n\t.reg .pred p;\n\tWB_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra "
                        "DB_%=;\n\tbra WB_%=;\n\tDB_%=:\n\t}" ::"r"(smem_u32(&barBulk)),
                        "r"(ph & 1)
                        : "memory");
                ++ph;
            }
        }
        if (acc == 0x1234567u || x == 1.2345f)
            sink[threadIdx.x] = acc;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(slot));
}

int main() {
    unsigned long long* d;
    uint8_t*            gsrc;
    uint32_t*           sink;
    cudaMalloc(&d, 148 * 16);
    cudaMalloc(&gsrc, 64 * 32768);
    cudaMemset(gsrc, 0, 64 * 32768);
    cudaMalloc(&sink, 4096);
    const size_t smem = 165 * 1024 + 1024;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int iters = 2000, perTile = 9;
    struct {
        int N, M, mode;
    } cases[] = {{256, 128, 0}, {128, 128, 0}, {64, 128, 0}, {256, 64, 0}};
    for (auto c : cases) {
        k<<<148, 640, smem>>>(iters, perTile, c.N, c.M, c.mode, d, 0, 1152, gsrc, sink, 0);
        k<<<148, 640, smem>>>(iters, perTile, c.N, c.M, c.mode, d, 0, 1152, gsrc, sink, 0);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
            printf("error: %s\n", cudaGetErrorString(e));
            return 1;
        }
        unsigned long long h[2];
        cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        printf("M %3d N %3d K 16 alone : %.1f clk/MMA (peak math = %.0f)\n", c.M, c.N,
               (double)h[1] / (iters * perTile), (double)c.M * c.N * 16 / 4096.0);
    }
    const char* names[] = {"none", "ALU on 16 warps", "TMEM reads 128 KB/period", "shared st+ld 64 KB/period",
                           "bulk copy 32 KB/period"};
    for (int noise = 0; noise <= 4; ++noise)
        for (int pace : {1152, 576}) {
            k<<<148, 640, smem>>>(iters, perTile, 256, 128, 0, d, noise, pace, gsrc, sink, 0);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) {
                printf("error: %s\n", cudaGetErrorString(e));
                return 1;
            }
            unsigned long long h[2];
            cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
            printf("M 128 N 256 with %-28s every %4d clk : %.1f clk/MMA\n", names[noise], pace,
                   (double)h[1] / (iters * perTile));
        }
    for (int mode : {0, 2})
    for (int noise : {5, 6})
        for (int work : {0, 128, 256, 512, 768}) {
            if (noise == 6 && work != 768)
                continue;
            k<<<148, 640, smem>>>(iters, perTile, 256, 128, mode, d, noise, 1152, gsrc, sink, work);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) {
                printf("error: %s\n", cudaGetErrorString(e));
                return 1;
            }
            unsigned long long h[2];
            cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
            printf("%s issue loop, %4d FMA per warp per 1152 clk on %s : %.1f clk/MMA\n",
                   mode == 2 ? "lean (converged, uniform)" : "one divergent lane", work,
                   noise == 5 ? "16 warps (4 per scheduler)" : "12 warps (none on the MMA warp's scheduler)",
                   (double)h[1] / (iters * perTile));
        }
    cudaFree(d);
    return 0;
}
