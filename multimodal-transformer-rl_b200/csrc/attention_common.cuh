// Warp-level building blocks shared by the attention forward / backward kernels.
#pragma once
#include "common.cuh"

namespace dtb {

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const bf16* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_u32(p)));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], const bf16* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_u32(p)));
}
// c(16x8, fp32) += a(16x16, bf16 row) * b(16x8, bf16 col)
__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// 64 rows x D bf16 from global (row stride ld elements, rows r0.. of T) into smem with pitch D+8; rows >= T zero-filled.
// 128 threads, 16-byte loads.
template <int D>
__device__ __forceinline__ void load_tile(bf16* dst, const bf16* src, int ld, int r0, int T) {
  constexpr int P = D + 8;
  constexpr int CH = D / 8;  // 16-byte chunks per row
#pragma unroll
  for (int i = threadIdx.x; i < 64 * CH; i += 128) {
    const int r = i / CH, c = (i % CH) * 8;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (r0 + r < T) v = ldg_u4(src + (size_t)(r0 + r) * ld + c);
    *reinterpret_cast<uint4*>(dst + r * P + c) = v;
  }
}

}  // namespace dtb
