// Front end fused into the frame upload: the reference's lda.flow network
// (/root/reference/src/Tools/FeatureExtraction/share/lda.flow) evaluated on the device, so that the host uploads the
// narrow base features (e.g. 16 floats per frame) instead of the transformed ones (45).
//   window         signal-vector-f32-sequence-concatenation, max-size = length, right = right, margin-condition
//                  present-not-empty (one output per input frame), margin-policy copy (a missing neighbour is the
//                  closest frame of the segment); oldest frame first
//                  (src/Signal/VectorSequenceConcatenation.hh:85-100, src/Signal/Delay.hh:32-46, Delay.cc:30-31,171-174)
//   multiplication signal-matrix-multiplication-f32: out = matrix * in
//                  (src/Signal/MatrixMult.hh:70,243; src/Math/Matrix.hh:482-489; src/Math/Vector.hh:95-101:
//                  result += a[i] * b[i] in f32, left to right -- the reference build emits scalar mulss / addss for
//                  this loop, no vectorisation, no FMA, so the device repeats exactly that and is bit-identical)
//
// One CTA stages the base rows a tile of 64 frames needs (64 + length - 1 rows, clamped to the segment) in shared
// memory -- the window of frame f is then the contiguous slice starting at row f -- and keeps the matrix there as
// well; one thread per (frame, output component) runs the sequential dot product.
#include "gmm_internal.cuh"

#include <algorithm>
#include <vector>

namespace {

constexpr uint32_t kLdaTileFrames = 64;  // frames per tile; halved for very wide windows until the rows fit

struct LdaTile {
    uint32_t f0, n;              // frames [f0, f0 + n)
    uint32_t segBegin, segEnd;   // of the segment [segBegin, segEnd)
};

__global__ void __launch_bounds__(256) lda_kernel(const LdaTile* __restrict__ tiles, uint32_t nTiles,
                                                  const float* __restrict__ base, uint32_t baseDim, uint32_t length,
                                                  uint32_t right, const float* __restrict__ matrix, uint32_t D,
                                                  uint32_t matrixInSmem, uint32_t tileFrames,
                                                  float* __restrict__ out) {
    extern __shared__ float sm[];
    const uint32_t Din = baseDim * length, left = length - 1 - right, pitch = Din + 1;
    float*         win = sm;                                               // [(64 + length - 1) x baseDim]
    float*         mat = sm + (size_t)(tileFrames + length - 1) * baseDim;  // [D x pitch]
    if (matrixInSmem)
        for (uint32_t i = threadIdx.x; i < D * Din; i += blockDim.x)
            mat[(i / Din) * pitch + i % Din] = matrix[i];
    for (uint32_t ti = blockIdx.x; ti < nTiles; ti += gridDim.x) {
        const LdaTile  tl   = tiles[ti];
        const uint32_t rows = tl.n + length - 1;
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < rows * baseDim; i += blockDim.x) {
            const uint32_t r = i / baseDim, c = i % baseDim;
            long long      f = (long long)tl.f0 - left + r;  // margin policy copy: the closest frame of the segment
            f                = f < (long long)tl.segBegin ? tl.segBegin : f >= (long long)tl.segEnd ? tl.segEnd - 1 : f;
            win[i]           = base[(size_t)f * baseDim + c];
        }
        __syncthreads();
        // one thread: output component o of 4 frames -- a[k] is loaded once for 4 independent sequential sums
        const uint32_t groups = (tl.n + 3) / 4;
        for (uint32_t i = threadIdx.x; i < groups * D; i += blockDim.x) {
            const uint32_t f = (i / D) * 4, o = i % D;
            const uint32_t nf = min(4u, tl.n - f);
            const float*   a  = matrixInSmem ? mat + (size_t)o * pitch : matrix + (size_t)o * Din;
            const float*   w0 = win + (size_t)f * baseDim;
            // frames beyond the tile read rows that exist in the staging buffer only up to `rows`: clamp them
            const float* w1 = w0 + (nf > 1 ? baseDim : 0);
            const float* w2 = w0 + (nf > 2 ? 2 * baseDim : 0);
            const float* w3 = w0 + (nf > 3 ? 3 * baseDim : 0);
            float        acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
#pragma unroll 4
            for (uint32_t k = 0; k < Din; ++k) {
                const float ak = a[k];
                acc0           = __fadd_rn(acc0, __fmul_rn(ak, w0[k]));
                acc1           = __fadd_rn(acc1, __fmul_rn(ak, w1[k]));
                acc2           = __fadd_rn(acc2, __fmul_rn(ak, w2[k]));
                acc3           = __fadd_rn(acc3, __fmul_rn(ak, w3[k]));
            }
            float* dst = out + (size_t)(tl.f0 + f) * D + o;
            dst[0]     = acc0;
            if (nf > 1)
                dst[D] = acc1;
            if (nf > 2)
                dst[2 * (size_t)D] = acc2;
            if (nf > 3)
                dst[3 * (size_t)D] = acc3;
        }
    }
}

}  // namespace

// frameOffsets: host, nSeg + 1 entries relative to d_base / d_x.
int gmm_lda_launch(gmm_handle* h, uint32_t nSeg, const uint32_t* frameOffsets, const float* d_base, float* d_x) {
    const LdaModel& lm = h->lda;
    uint32_t tileFrames = kLdaTileFrames;
    while (tileFrames > 4 && (size_t)(tileFrames + lm.length - 1) * lm.baseDim * 4 > 64 * 1024)
        tileFrames /= 2;
    std::vector<LdaTile> tiles;
    for (uint32_t s = 0; s < nSeg; ++s) {
        if (frameOffsets[s + 1] < frameOffsets[s]) {
            h->setError("gmm_lda: frame offsets not ascending");
            return GMM_ERR_ARG;
        }
        for (uint32_t f = frameOffsets[s]; f < frameOffsets[s + 1]; f += tileFrames)
            tiles.push_back({f, std::min(tileFrames, frameOffsets[s + 1] - f), frameOffsets[s], frameOffsets[s + 1]});
    }
    if (tiles.empty())
        return GMM_OK;
    void* dt = nullptr;
    int   rc = h->reserve(SCR_TILES, tiles.size() * sizeof(LdaTile), &dt);
    if (rc)
        return rc;
    GMM_CUDA_TRY(h, cudaMemcpyAsync(dt, tiles.data(), tiles.size() * sizeof(LdaTile), cudaMemcpyHostToDevice, h->stream));
    const uint32_t Din      = lm.baseDim * lm.length;
    const size_t   winBytes = (size_t)(tileFrames + lm.length - 1) * lm.baseDim * 4;
    const size_t   matBytes = (size_t)lm.D * (Din + 1) * 4;
    const bool     inSmem   = winBytes + matBytes <= 160 * 1024;
    const size_t   smem     = winBytes + (inSmem ? matBytes : 0);
    if (smem > 200 * 1024) {
        h->setError("gmm_lda: window too large for shared memory");
        return GMM_ERR_UNSUPPORTED;
    }
    if (smem > 48 * 1024)
        GMM_CUDA_TRY(h, cudaFuncSetAttribute(lda_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = (int)std::min<size_t>(tiles.size(), (size_t)h->smCount * 4);
    h->timeBegin();
    lda_kernel<<<grid, 256, smem, h->stream>>>((const LdaTile*)dt, (uint32_t)tiles.size(), d_base, lm.baseDim, lm.length,
                                               lm.right, lm.matrix, lm.D, inSmem ? 1u : 0u, tileFrames, d_x);
    h->timeEnd();
    h->launches++;
    GMM_CUDA_TRY(h, cudaGetLastError());
    // the tile list is pageable host memory: the copy above has consumed it before cudaMemcpyAsync returned
    return GMM_OK;
}
