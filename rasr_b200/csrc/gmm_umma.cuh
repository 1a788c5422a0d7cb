// Shared device-side building blocks of the tcgen05 scorers (float: gmm_tc.cu, quantized: gmm_qtc.cu): tile
// geometry, the per-tile metadata that travels with a B tile, and thin wrappers over the PTX the kernels use
// (mbarrier, cp.async.bulk, tcgen05.mma / commit / ld, UMMA shared-memory descriptors).  sm_100a only.
#ifndef GMM_UMMA_CUH
#define GMM_UMMA_CUH
#include <cuda_runtime.h>
#include <stdint.h>

namespace {

constexpr int kTileN       = 256;  // packed columns per B tile (UMMA N)
constexpr int kTileM       = 128;  // frames per A tile (UMMA M, TMEM lanes)
constexpr int kEpiWarp0    = 2;    // warp 0: bulk-copy producer + TMEM allocator, warp 1: MMA issuer
constexpr int kMaxStagesA  = 4;
constexpr int kRing        = 4;    // tile-indexed barrier rings (>= kMaxStagesA, >= 2 accumulators + 2)
constexpr int kMaxKSteps   = 16;   // 16-element k-steps per precision part (K <= 256)
constexpr int kMetaBytes   = kTileN * 2 + 48;  // slot[256] u16, TileMeta
constexpr uint32_t kSmemLimit = 232448;                     // 227 KB

struct TileMeta {
    uint32_t segEnd;       // bit s: 8-column segment s closes a mixture
    uint32_t segEmpty;     // bit s (on a closing segment): the mixture has no live density at all
    uint32_t mixBase[4];   // first mixture of column part p (part = 128 or 64 columns, one epilogue warp each)
    uint32_t live[4];      // 4 bits per segment: live columns - 1 (pads repeat the first live column)
    uint32_t uni;          // byte p: segments per mixture if every mixture of part p has that many and none is
                           // empty (the part is full), else 0
};
static_assert(sizeof(TileMeta) == 44, "TileMeta layout");

// ---- PTX wrappers ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "WAIT_%=:\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
            "@p bra DONE_%=;\n\t"
            "bra WAIT_%=;\n\t"
            "DONE_%=:\n\t}" ::"r"(bar),
            "r"(parity)
            : "memory");
}
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {  // one probe, no spinning
    uint32_t ok;
    asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    return ok != 0;
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t dTmem, uint64_t aDesc, uint64_t bDesc, uint32_t idesc,
                                           uint32_t accumulate) {
    asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(dTmem),
            "l"(aDesc), "l"(bDesc), "r"(idesc), "r"(accumulate)
            : "memory");
}
// accumulating fp16 MMA whose accumulator INPUT is scaled: D = A.B + D * 2^-SHIFT (scale-input-d, kind::f16 / tf32 only)
template<int SHIFT>
__device__ __forceinline__ void tc_mma_f16_scaled(uint32_t dTmem, uint64_t aDesc, uint64_t bDesc, uint32_t idesc) {
    asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p, %4;\n\t}" ::"r"(dTmem),
            "l"(aDesc), "l"(bDesc), "r"(idesc), "n"(SHIFT)
            : "memory");
}
// 8-bit floating point operands (formats in the instruction descriptor), K = 32 bytes per instruction
__device__ __forceinline__ void tc_mma_f8(uint32_t dTmem, uint64_t aDesc, uint64_t bDesc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(dTmem),
            "l"(aDesc), "l"(bDesc), "r"(idesc), "r"(accumulate)
            : "memory");
}
__device__ __forceinline__ bool elect_one() {  // true in exactly one lane of the (converged) warp
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
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
__device__ __forceinline__ void tmem_ld_wait() {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void sts_f32(uint32_t addr, float v) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void sts_u32(uint32_t addr, uint32_t v) {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ float lds_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void sts_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// UMMA shared-memory descriptor, K-major, swizzled canonical layout (8-row groups of SWZ-byte rows).
template<int SWZ>
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    constexpr uint64_t layout = SWZ == 128 ? 2ull : 4ull;  // SWIZZLE_128B : SWIZZLE_64B
    constexpr uint64_t sbo    = (8 * SWZ) >> 4;            // stride between 8-row groups
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | (sbo << 32) | (1ull << 46) | (layout << 61);
}


// ---- swizzled tile image addressing (host + device) ---------------------------------------------------
// byte offset of 16-byte chunk `c16` of row `row` inside one atom block [rows][SWZ bytes]
__host__ __device__ inline uint32_t swz_offset(uint32_t row, uint32_t c16, uint32_t swz) {
    const uint32_t lin  = row * swz + c16 * 16;
    const uint32_t mask = (swz / 16) - 1;  // 7 for 128B, 3 for 64B
    return lin ^ (((lin >> 7) & mask) << 4);
}


}  // namespace
#endif  // GMM_UMMA_CUH
