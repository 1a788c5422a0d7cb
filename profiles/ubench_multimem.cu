// Micro-benchmark: can the mixture-sharded push (gmm_gather.cu) leave the GPU ONCE and be replicated by the NVSwitch
// (multimem.st on a multicast object, CUDA driver API cuMulticast*) instead of being stored N-1 times over unicast
// peer mappings?  One process drives all N GPUs of the box:
//   * every GPU gets a block of `bytes` of physical memory (cuMemCreate), mapped on all GPUs (unicast peer mappings)
//     and bound to one multicast object that is mapped on GPU 0;
//   * GPU 0 writes a [rows x rowBytes] column range (pitch = ld bytes) of its block into every block
//       (a) with N-1 unicast 16-byte stores per element (what gather_push_kernel does today),
//       (b) with ONE multimem.st.v4.f32 per element;
//   * both are timed with CUDA events on GPU 0 and every peer's copy is checked.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench_multimem profiles/ubench_multimem.cu -lcuda
// Usage: ubench_multimem [rows] [rowFloats] [ldFloats]
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>

#define CU(x)                                                                             \
    do {                                                                                  \
        CUresult r_ = (x);                                                                \
        if (r_ != CUDA_SUCCESS) {                                                         \
            const char* s_ = nullptr;                                                     \
            cuGetErrorString(r_, &s_);                                                    \
            printf("%s failed: %d (%s)\n", #x, (int)r_, s_ ? s_ : "?");                   \
            return 1;                                                                     \
        }                                                                                 \
    } while (0)
#define RT(x)                                                                             \
    do {                                                                                  \
        cudaError_t e_ = (x);                                                             \
        if (e_ != cudaSuccess) {                                                          \
            printf("%s failed: %s\n", #x, cudaGetErrorString(e_));                        \
            return 1;                                                                     \
        }                                                                                 \
    } while (0)

struct Peers {
    float* p[8];
};

__global__ void fill(float* a, size_t n, float v) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        a[i] = v + (float)(i % 1000);
}

__global__ void __launch_bounds__(256) push_unicast(Peers peers, int world, const float* src, size_t rows, size_t w4, size_t ld) {
    const size_t total = rows * w4;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / w4, c = i - r * w4, off = r * ld + c * 4;
        const float4 v = *reinterpret_cast<const float4*>(src + off);
        for (int k = 1; k < world; ++k)
            *reinterpret_cast<float4*>(peers.p[k] + off) = v;
    }
}

__global__ void __launch_bounds__(256) push_multicast(float* mc, const float* src, size_t rows, size_t w4, size_t ld) {
    const size_t total = rows * w4;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / w4, c = i - r * w4, off = r * ld + c * 4;
        const float4 v = *reinterpret_cast<const float4*>(src + off);
        asm volatile("multimem.st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc + off), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                     : "memory");
    }
}

int main(int argc, char** argv) {
    const size_t rows = argc > 1 ? atol(argv[1]) : 8192, rowF = argc > 2 ? atol(argv[2]) : 8192;
    size_t       ld   = argc > 3 ? atol(argv[3]) : 0;
    CU(cuInit(0));
    int n = 0;
    CU(cuDeviceGetCount(&n));
    if (n > 8)
        n = 8;
    if (ld == 0)
        ld = rowF * n;  // rank 0's columns of an [rows x N * rowF] block
    printf("devices: %d; column range %zu rows x %zu floats, pitch %zu floats (%.1f MB per destination)\n", n, rows, rowF, ld,
           rows * rowF * 4 / 1e6);
    if (n < 2) {
        printf("needs two GPUs\n");
        return 0;
    }
    std::vector<CUdevice> dev(n);
    for (int i = 0; i < n; ++i) {
        CU(cuDeviceGet(&dev[i], i));
        int mc = 0;
        CU(cuDeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev[i]));
        printf("device %d: multicast supported = %d\n", i, mc);
        if (!mc)
            return 0;
        RT(cudaSetDevice(i));
        RT(cudaFree(0));
        for (int j = 0; j < n; ++j)
            if (j != i)
                cudaDeviceEnablePeerAccess(j, 0);
    }
    size_t bytes = rows * ld * 4;
    CUmulticastObjectProp mp = {};
    mp.numDevices  = n;
    mp.handleTypes = 0;
    size_t gran = 0, granMin = 0;
    mp.size = bytes;
    CU(cuMulticastGetGranularity(&granMin, &mp, CU_MULTICAST_GRANULARITY_MINIMUM));
    CU(cuMulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    bytes   = (bytes + gran - 1) / gran * gran;
    mp.size = bytes;
    printf("multicast granularity: minimum %zu, recommended %zu; block %zu bytes\n", granMin, gran, bytes);
    CUmemGenericAllocationHandle mcH;
    CU(cuMulticastCreate(&mcH, &mp));
    for (int i = 0; i < n; ++i)
        CU(cuMulticastAddDevice(mcH, dev[i]));
    std::vector<CUmemGenericAllocationHandle> mem(n);
    std::vector<CUdeviceptr>                  va(n);
    std::vector<CUmemAccessDesc>              acc(n);
    for (int i = 0; i < n; ++i) {
        acc[i].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        acc[i].location.id   = i;
        acc[i].flags         = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    }
    for (int i = 0; i < n; ++i) {
        CUmemAllocationProp ap = {};
        ap.type          = CU_MEM_ALLOCATION_TYPE_PINNED;
        ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        ap.location.id   = i;
        CU(cuMemCreate(&mem[i], bytes, &ap, 0));
        CU(cuMemAddressReserve(&va[i], bytes, gran, 0, 0));
        CU(cuMemMap(va[i], bytes, 0, mem[i], 0));
        CU(cuMemSetAccess(va[i], bytes, acc.data(), n));  // every GPU may touch every block (unicast)
        CU(cuMulticastBindMem(mcH, 0, mem[i], 0, bytes, 0));
    }
    CUdeviceptr mcVa;
    CU(cuMemAddressReserve(&mcVa, bytes, gran, 0, 0));
    CU(cuMemMap(mcVa, bytes, 0, mcH, 0));
    CU(cuMemSetAccess(mcVa, bytes, &acc[0], 1));  // the multicast mapping is used by GPU 0

    RT(cudaSetDevice(0));
    Peers peers = {};
    for (int i = 0; i < n; ++i)
        peers.p[i] = reinterpret_cast<float*>(va[i]);
    float*       src  = peers.p[0];
    const size_t w4   = rowF / 4;
    cudaEvent_t  e0, e1;
    RT(cudaEventCreate(&e0));
    RT(cudaEventCreate(&e1));
    fill<<<592, 256>>>(src, rows * ld, 1.0f);
    RT(cudaDeviceSynchronize());
    std::vector<float> host(rowF);
    auto check = [&](const char* what, float base) -> int {
        int bad = 0;
        for (int k = 1; k < n; ++k) {
            const size_t r = rows - 1;
            if (cudaMemcpy(host.data(), peers.p[k] + r * ld, rowF * 4, cudaMemcpyDeviceToHost) != cudaSuccess)
                return -1;
            for (size_t c = 0; c < rowF; ++c)
                bad += host[c] != base + (float)((r * ld + c) % 1000);
        }
        printf("  %s: last row on %d peers %s\n", what, n - 1, bad ? "WRONG" : "ok");
        return bad;
    };
    for (int ctas : {64, 128, 296, 592}) {
        for (int variant = 0; variant < 2; ++variant) {
            float best = 1e30f;
            for (int rep = 0; rep < 6; ++rep) {
                RT(cudaEventRecord(e0));
                if (variant == 0)
                    push_unicast<<<ctas, 256>>>(peers, n, src, rows, w4, ld);
                else
                    push_multicast<<<ctas, 256>>>(reinterpret_cast<float*>(mcVa), src, rows, w4, ld);
                RT(cudaEventRecord(e1));
                RT(cudaEventSynchronize(e1));
                float ms;
                RT(cudaEventElapsedTime(&ms, e0, e1));
                if (rep)
                    best = ms < best ? ms : best;
            }
            const double payload = (double)rows * rowF * 4;
            printf("%s, %3d CTAs: %.3f ms  -> %.0f GB/s delivered to the %d peers (%.0f GB/s leaving GPU 0)\n",
                   variant ? "multimem.st   " : "unicast stores", ctas, best, payload * (n - 1) / best / 1e6, n - 1,
                   payload * (variant ? 1 : n - 1) / best / 1e6);
        }
    }
    fill<<<592, 256>>>(src, rows * ld, 7.0f);
    push_unicast<<<128, 256>>>(peers, n, src, rows, w4, ld);
    RT(cudaDeviceSynchronize());
    check("unicast", 7.0f);
    fill<<<592, 256>>>(src, rows * ld, 9.0f);
    push_multicast<<<128, 256>>>(reinterpret_cast<float*>(mcVa), src, rows, w4, ld);
    RT(cudaDeviceSynchronize());
    check("multimem", 9.0f);
    return 0;
}
