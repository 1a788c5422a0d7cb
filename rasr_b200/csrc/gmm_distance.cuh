// Distance of the reference's float scorers, operation for operation: four lane sums over d < 4*floor(D/4),
// (l0+l1)+(l2+l3), then the scalar tail, no FMA contraction
// (/root/reference/src/Mm/GaussDiagonalMaximumFeatureScorer.cc:144-218).  Shared by the direct scoring kernel
// (gmm_direct.cu) and the training E-step (gmm_accum.cu).
#pragma once
#include <cuda_runtime.h>

template<int DP, int FR>
struct Frames {
    float x[FR][DP];
};

// distance of the reference: lanes over d < eff, hadd, tail.  D, eff are warp-uniform run-time values.
template<int DP, int FR>
__device__ __forceinline__ void distance(const Frames<DP, FR>& f, const float4* __restrict__ mean4,
                                         const float4* __restrict__ isr4, int D, int eff, float* out) {
    float lane[FR][4];
#pragma unroll
    for (int r = 0; r < FR; ++r)
        lane[r][0] = lane[r][1] = lane[r][2] = lane[r][3] = 0.f;
    float4 tm = make_float4(0.f, 0.f, 0.f, 0.f), ts = tm;
#pragma unroll
    for (int q = 0; q < DP / 4; ++q) {
        if (4 * q < D) {  // uniform
            float4 m = __ldg(mean4 + q);
            float4 s = __ldg(isr4 + q);
            if (4 * q < eff) {
#pragma unroll
                for (int r = 0; r < FR; ++r) {
                    float d0 = __fmul_rn(__fsub_rn(m.x, f.x[r][4 * q + 0]), s.x);
                    float d1 = __fmul_rn(__fsub_rn(m.y, f.x[r][4 * q + 1]), s.y);
                    float d2 = __fmul_rn(__fsub_rn(m.z, f.x[r][4 * q + 2]), s.z);
                    float d3 = __fmul_rn(__fsub_rn(m.w, f.x[r][4 * q + 3]), s.w);
                    lane[r][0] = __fadd_rn(lane[r][0], __fmul_rn(d0, d0));
                    lane[r][1] = __fadd_rn(lane[r][1], __fmul_rn(d1, d1));
                    lane[r][2] = __fadd_rn(lane[r][2], __fmul_rn(d2, d2));
                    lane[r][3] = __fadd_rn(lane[r][3], __fmul_rn(d3, d3));
                }
            }
            else {
                tm = m;
                ts = s;
            }
        }
    }
    const int tail = D - eff;  // 0..3
#pragma unroll
    for (int r = 0; r < FR; ++r) {
        float b0  = __fadd_rn(lane[r][0], lane[r][1]);
        float b1  = __fadd_rn(lane[r][2], lane[r][3]);
        float res = __fadd_rn(0.f, __fadd_rn(b0, b1));
        // tail elements live in quad q = eff/4; select them statically
        float tx[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int q = 0; q < DP / 4; ++q) {
            if (4 * q == eff) {
                tx[0] = f.x[r][4 * q + 0];
                tx[1] = f.x[r][4 * q + 1];
                tx[2] = f.x[r][4 * q + 2];
            }
        }
        if (tail > 0) {
            float d = __fmul_rn(__fsub_rn(tm.x, tx[0]), ts.x);
            res     = __fadd_rn(res, __fmul_rn(d, d));
        }
        if (tail > 1) {
            float d = __fmul_rn(__fsub_rn(tm.y, tx[1]), ts.y);
            res     = __fadd_rn(res, __fmul_rn(d, d));
        }
        if (tail > 2) {
            float d = __fmul_rn(__fsub_rn(tm.z, tx[2]), ts.z);
            res     = __fadd_rn(res, __fmul_rn(d, d));
        }
        out[r] = res;
    }
}
