// extern "C" boundary of libdtb200.so (declared in include/dtb200.h).  No torch types, no exceptions.
#include "../../include/dtb200.h"
#include "kernels.h"

#include <string.h>

namespace dtb {
const char* last_error();
int require_sm100();
}  // namespace dtb

using namespace dtb;

extern "C" {

int dtb_abi_version(void) { return DTB_ABI_VERSION; }
const char* dtb_last_error(void) { return dtb::last_error(); }
int dtb_check_device(void) { return dtb::require_sm100(); }

int dtb_gemm_bf16(void* stream, const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major, int M, int N, int K,
                  const dtb_gemm_epilogue* e, int split_k, int bn_override) {
  DTB_REQUIRE(A && B && e, DTB_ERR_BAD_ARG, "dtb_gemm_bf16: null operand");
  DTB_TRY(require_sm100());
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  GemmEpilogue ep;
  memset(&ep, 0, sizeof(ep));
  ep.alpha = e->alpha == 0.f ? 1.f : e->alpha;
  const bool split = split_k > 1;
  if (split) {
    DTB_REQUIRE(e->partial && e->out_f32, DTB_ERR_BAD_ARG, "dtb_gemm_bf16: split-K needs partial and out_f32");
    ep.partial = e->partial;
    int rc = gemm_bf16(st, (const bf16*)A, lda, a_mn_major, (const bf16*)B, ldb, b_mn_major, M, N, K, ep, split_k, bn_override);
    if (rc < 0) return rc;
    return splitk_reduce(st, e->partial, rc, M, N, e->out_f32, e->ldo_f, e->accumulate, ep.alpha);
  }
  ep.bias = e->bias;
  ep.residual = e->residual;
  ep.ldr = e->ldr;
  ep.aux = (const bf16*)e->aux;
  ep.ldaux = e->ldaux;
  ep.act = e->act;
  ep.dact = e->dact;
  ep.out_bf16 = (bf16*)e->out_bf16;
  ep.ldo_b = e->ldo_b;
  ep.out_f32 = e->out_f32;
  ep.ldo_f = e->ldo_f;
  ep.out_preact = (bf16*)e->out_preact;
  int rc = gemm_bf16(st, (const bf16*)A, lda, a_mn_major, (const bf16*)B, ldb, b_mn_major, M, N, K, ep, 1, bn_override);
  return rc < 0 ? rc : DTB_OK;
}

}  // extern "C"
