// Micro-benchmark: how fast can 148 persistent CTAs write a [T x ld] f32 matrix in the access pattern of the scoring
// kernels' epilogues -- per work item one column block of `piece` bytes over a range of 128-row tiles, items dealt
// round-robin to the CTAs (adjacent column blocks are written by different CTAs at about the same time)?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench_rowpiece profiles/ubench_rowpiece.cu
// Usage: ubench_rowpiece T M   (M = columns in units of 4-byte elements; ld = M rounded up to 4)
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

template<int PIECE>  // bytes per row piece: 64, 128, 256
__global__ void __launch_bounds__(512, 1) k(float* out, size_t ld, int T, int nBlocks, int split) {
    constexpr int Q   = PIECE / 16;        // 16-byte accesses per row piece
    constexpr int RPP = 512 / Q;           // rows per pass of the CTA
    const int frameTiles = (T + 127) / 128, per = (frameTiles + split - 1) / split;
    const int items = nBlocks * split;
    const int qi = threadIdx.x % Q, r0 = threadIdx.x / Q;
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
        const int nb = item / split, chunk = item % split;
        const int f0 = chunk * per, f1 = min(frameTiles, f0 + per);
        for (int f = f0; f < f1; ++f) {
#pragma unroll
            for (int ps = 0; ps < 128 / RPP; ++ps) {
                const int row = f * 128 + r0 + ps * RPP;
                if (row < T)
                    *reinterpret_cast<float4*>(reinterpret_cast<char*>(out + (size_t)row * ld) + (size_t)nb * PIECE + qi * 16) =
                            make_float4(1.f, 2.f, 3.f, (float)item);
            }
        }
    }
}

int main(int argc, char** argv) {
    const int T = argc > 1 ? atoi(argv[1]) : 32000, M = argc > 2 ? atoi(argv[2]) : 4504;
    const size_t ld = (M + 3) / 4 * 4;
    float* d;
    cudaMalloc(&d, (size_t)T * ld * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int piece : {64, 128, 256}) {
        const int nBlocks = (int)(ld * 4 / piece);
        for (int split : {1, 4, 11}) {
            float best = 1e30f;
            for (int rep = 0; rep < 5; ++rep) {
                cudaEventRecord(e0);
                if (piece == 64) k<64><<<148, 512>>>(d, ld, T, nBlocks, split);
                else if (piece == 128) k<128><<<148, 512>>>(d, ld, T, nBlocks, split);
                else k<256><<<148, 512>>>(d, ld, T, nBlocks, split);
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                if (ms < best) best = ms;
            }
            printf("T %d ld %zu piece %3d B split %2d: %.3f ms  %.0f GB/s written\n", T, ld, piece, split, best,
                   (double)T * nBlocks * piece / best / 1e6);
        }
    }
    cudaMemset(d, 0, (size_t)T * ld * 4);
    cudaEventRecord(e0);
    cudaMemsetAsync(d, 1, (size_t)T * ld * 4);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    printf("cudaMemset of the same block: %.3f ms  %.0f GB/s\n", ms, (double)T * ld * 4 / ms / 1e6);
    return cudaGetLastError() != cudaSuccess;
}
