// Internal host-side launch API shared by the .cu translation units (not part of the C ABI).
#pragma once
#include "common.cuh"

namespace dtb {

struct GemmEpilogue {
  const float* bias;      // [N] or null
  const float* residual;  // fp32 [M, ldr] or null (added after activation)
  int ldr;
  const bf16* aux;        // for dact: saved activation output (relu) or pre-activation (gelu)
  int ldaux;
  int act;                // 0 none, 1 relu, 2 gelu_new   (applied after bias)
  int dact;               // 0 none, 1 relu', 2 gelu_new' (multiplies the accumulator by act'(aux))
  bf16* out_bf16;         // optional bf16 output [M, ldo_b]
  int ldo_b;
  float* out_f32;         // optional fp32 output [M, ldo_f]
  int ldo_f;
  bf16* out_preact;       // optional: bias-added pre-activation (bf16, ld = ldo_b) saved for gelu backward
  float* partial;         // split-K: fp32 partials [split][M][N]; epilogue fields above are then ignored
  float alpha;            // scale applied to the accumulator first
  DropSite drop;          // dropout of (acc + bias [act]) before the residual add; element index = row * N + col
  // Fused LayerNorm of the finished output rows (single-CTA kernel with BN == N <= 256, N % 32 == 0: a CTA owns whole rows, TMEM lane =
  // row): ln_out[row] = (v - mean) * rstd * ln_gamma + ln_beta in bf16, v = the fp32 value written to out_f32 (bias / dropout / residual
  // applied); mean / rstd are saved for ln_bwd.  Replaces the standalone ln_fwd launch after every c_proj (HF:272, :303, :450).
  const float* ln_gamma;
  const float* ln_beta;
  bf16* ln_out;
  int ld_ln;
  float* ln_mean;
  float* ln_rstd;
  float ln_eps;
  // Split-K reduced inside the kernel: every epilogue warp takes a ticket on its (tile, row block) counter after storing its partial
  // rows; the last of the split_k arrivals sums the slices in index order (deterministic) and writes sk_out.  Counters return to zero.
  unsigned int* sk_counters;
  float* sk_out;
  int sk_ldo;
};

struct GemmParams {
  int M, N, K;
  int BN;            // N tile: 32..256, multiple of 32 (multiple of 64 when B is MN-major)
  int a_mn_major, b_mn_major;
  int split_k;       // >= 1
  int kb_per_split;  // k-blocks (of 64) per split
  int m_tiles, n_tiles;
  GemmEpilogue ep;
};


// ---- gemm_host.cu ------------------------------------------------------------------------------
// Returns the number of split-K slices written (>=1) or a negative error code.
int gemm_bf16(cudaStream_t stream, const bf16* A, int lda, int a_mn_major, const bf16* B, int ldb, int b_mn_major, int M, int N, int K,
              const GemmEpilogue& ep, int split_k, int bn_override);
int gemm_pick_split(int M, int N, int K, int BN);
int gemm_pick_bn(int M, int N, int b_mn_major);
int gemm_wgrad_split(int Mout, int Nout, int Kred, size_t partial_floats);
int gemm_wgrad(cudaStream_t stream, const bf16* A, int lda, const bf16* B, int ldb, int Mout, int Nout, int Kred, float* out, int ldo,
               float* partial, size_t partial_floats);
// `need` zero-initialised device counters for a "last arrival finishes the reduction" kernel (self-resetting; a slice of a ring, so
// kernels running concurrently on different streams never share one).  *out = nullptr when the request is too large.
int ticket_counters_take(unsigned int need, unsigned int** out);
int splitk_reduce(cudaStream_t stream, const float* partial, int splits, int M, int N, float* out, int ldo, int accumulate, float scale);

// 2-D bf16 tensor map (cached): dims {d0 contiguous, d1}, row stride in elements, box {b0, b1}, 128B swizzle.
int make_tmap(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t stride_elems, uint32_t b0, uint32_t b1);

int make_tmap3(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t stride1, uint64_t stride2, uint32_t b0,
               uint32_t b1);

// ---- elementwise.cu ----------------------------------------------------------------------------
// fp32 [R, C] -> bf16 copy [R, C] and (optional) bf16 transpose [C, R]
int cast_transpose(cudaStream_t stream, const float* src, int R, int C, bf16* dst, bf16* dst_t);
int cast_bf16(cudaStream_t stream, const float* src, size_t n, bf16* dst);
// column sums of a bf16 matrix [M, N] (ld) -> fp32 out[N] (+= if accumulate)
int colsum_bf16(cudaStream_t stream, const bf16* x, int M, int N, int ld, float* out, int accumulate, float* scratch);
size_t colsum_scratch_floats(int N);
int colsum_bf16_static(cudaStream_t stream, const bf16* x, int M, int N, int ld, float* out, int accumulate);  // scratch = static device array

// dropout plumbing: advance the device-resident step counter; export keep bits for the tests (groups = 0: no group folding)
int dropout_advance(cudaStream_t stream, void* state);
int dropout_mask(cudaStream_t stream, const DropSite& d, int groups, long long n, uint8_t* out);

int pad_left(cudaStream_t stream, int nseg, const void* const* srcs, void* const* dsts, const int* row_bytes, int t, int K, float* mask);

// ---- get_batch.cu --------------------------------------------------------------------------------
int get_batch(cudaStream_t stream, const dtb_traj_store* st, const long long* traj_idx, const long long* start, int B, int K, int max_ep_len,
              double scale, const dtb_batch_out* out);

// ---- optim.cu ------------------------------------------------------------------------------------
size_t adamw_scratch_floats();
int clip_adamw_step(cudaStream_t stream, int nseg, const dtb_adamw_segment* segs, const dtb_adamw_hparams* hp, int* step_dev, float* scratch);

// ---- layernorm.cu ------------------------------------------------------------------------------
int ln_fwd(cudaStream_t stream, const float* x, const float* gamma, const float* beta, int M, int H, float eps, bf16* y, float* mean,
           float* rstd);
// dx_out = dres + LNbwd(dy); also bf16 copy; dgamma/dbeta accumulated (+=) after internal partial reduce
int ln_bwd(cudaStream_t stream, const bf16* dy, const float* x, const float* mean, const float* rstd, const float* gamma,
           const float* dres, int M, int H, float* dx_out, bf16* dx_out_bf16, float* dgamma, float* dbeta, float* scratch,
           const DropSite& drop_bf16 = drop_none(),  // dropout mask of the consumer site applied to the bf16 copy only (idx = row * H + col)
           float* dcolsum = nullptr);                // optional [H]: column sums of the (masked) bf16-copy values = bias gradient of the Linear that produced this row
size_t ln_bwd_scratch_floats(int H);

// ---- attention.cu ------------------------------------------------------------------------------
// qkv: bf16 [B, T, 3H] (q | k | v, head h at columns h*d), mask: fp32 [B, K] (T = 3K, token j uses mask[b, j/3]) or null
int attn_fwd(cudaStream_t stream, const bf16* qkv, const float* mask, int B, int T, int nh, int d, bf16* out, float* lse,
             const DropSite& drop = drop_none());  // dropout on the softmax weights: group = b * nh + h, idx = i * T + j
int attn_bwd(cudaStream_t stream, const bf16* qkv, const float* mask, const bf16* out, const bf16* dout, const float* lse, int B, int T,
             int nh, int d, bf16* dqkv, float* delta, const DropSite& drop = drop_none());

// ---- attention_tc.cu: tcgen05 / TMEM / TMA flash attention (head_dim 64) ----------------------------
bool attn_tc_supported(int T, int d);
int attn_fwd_tc(cudaStream_t stream, const bf16* qkv, const float* mask, int B, int T, int nh, int d, bf16* out, float* lse,
                const DropSite& drop);
int attn_bwd_tc(cudaStream_t stream, const bf16* qkv, const float* mask, const bf16* out, const bf16* dout, const float* lse, int B, int T,
                int nh, int d, bf16* dqkv, float* delta, const DropSite& drop);

// ---- embed.cu ----------------------------------------------------------------------------------
int embed_fwd(cudaStream_t stream, const float* states, const float* state_proj, const float* actions, const float* rtg,
              const long long* timesteps, const float* table, const float* wR, const float* bR, const float* Ws, const float* bs,
              const float* Wa, const float* ba, const float* gamma, const float* beta, int BK, int sd, int ad, int H, int E, float eps,
              float* x, float* pre, float* mean, float* rstd, bf16* U, int P8, int u_states, const DropSite& drop = drop_none(),
              const float* ln1_gamma = nullptr, const float* ln1_beta = nullptr, bf16* h1 = nullptr, float* mean1 = nullptr,
              float* rstd1 = nullptr);
int embed_te_scatter(cudaStream_t stream, const bf16* d_pre, const long long* timesteps, const float* mask /* [BK] or null */, int BK, int H,
                     int E, float* d_table);
int embed_dstates(cudaStream_t stream, const bf16* d_pre, const float* Ws, int BK, int H, int sd, float* d_states);
int embed_unpack(cudaStream_t stream, const float* G, const float* bsum, int H, int sd, int ad, int P8, int u_states, float* dwR,
                 float* dbR, float* dWs, float* dbs, float* dWa, float* dba);

// ---- heads.cu ----------------------------------------------------------------------------------
int heads_fwd(cudaStream_t stream, const bf16* y, const float* Wa, const float* ba, const float* Ws, const float* bs, const float* Wr,
              const float* br, int BK, int H, int sd, int ad, int tanh_act, float* action_preds, float* state_preds,
              float* return_preds);
size_t mse_scratch_floats();
int mse_loss(cudaStream_t stream, const float* pred, const float* target, const float* mask, int BK, int ad, float grad_scale,
             float* loss_out, float* d_pred, float* scratch);
int heads_bwd(cudaStream_t stream, const float* d_action, const float* d_state, const float* d_return, const float* action_preds,
              const float* Wa, const float* Ws, const float* Wr, int BK, int H, int sd, int ad, int tanh_act, int ZA8, int ZS8, bf16* dy,
              bf16* dzA, bf16* dzS);
int copy2d(cudaStream_t stream, const float* src, int lds, int R, int C, float* dst, int ldd);

// ---- conv1.cu (implicit-GEMM first convolution) -------------------------------------------------
bool conv1_implicit_supported(int C, int W, int stride, int cout);
int conv1_fwd(cudaStream_t stream, const void* x, int x_u8, int N, int C, int H, int W, int stride, float scale, const bf16* w_staged,
              const float* bias, int cout, bf16* y);
size_t conv1_wgrad_partial_floats(int C, int cout);
int conv1_wgrad(cudaStream_t stream, const void* x, int x_u8, int N, int C, int H, int W, int stride, float scale, const bf16* dz, int cout,
                float* partial, float* dw, float* db /* bias gradient from the same pass (ones chunk); may be null */);

// ---- convn.cu (implicit-GEMM conv2 / conv3: forward, data gradient, weight + bias gradient) ---------------------
bool convn_supported(int cin, int cout, int k, int stride);
int convn_fwd(cudaStream_t stream, const bf16* x, int N, int hin, int win, int cin, int k, int stride, const bf16* w_staged, const float* bias,
              int cout, bf16* y);
size_t convn_dgrad_weight_elems(int cin, int cout, int k);
int convn_stage_dgrad(cudaStream_t stream, const float* w, int cin, int cout, int k, int stride, bf16* dst);
int convn_dgrad(cudaStream_t stream, const bf16* dz, int N, int hin, int win, int cin, int k, int stride, const bf16* wd_staged, int cout,
                const bf16* act, bf16* dx);
size_t convn_wgrad_partial_floats(int cin, int k);
int convn_wgrad(cudaStream_t stream, const bf16* x, int N, int hin, int win, int cin, int k, int stride, const bf16* dz, int cout,
                float* partial, float* dw, float* db);

// ---- conv.cu -----------------------------------------------------------------------------------
int im2col_nchw_k8(cudaStream_t stream, const float* x, int N, int C, int H, int W, int stride, float scale, bf16* p);
int im2col_nhwc(cudaStream_t stream, const bf16* x, int N, int H, int W, int C, int k, int stride, bf16* p);
int col2im_nhwc_relu(cudaStream_t stream, const bf16* dp, const bf16* y, int N, int H, int W, int C, int k, int stride, bf16* dz);
int permute_stage(cudaStream_t stream, const float* src, int O, int A, int B, bf16* dst);
int permute_unstage(cudaStream_t stream, const float* src, int O, int A, int B, float* dst);
int add_bf16(cudaStream_t stream, const bf16* a, const bf16* b, size_t n, bf16* out);
int relu_mask_bf16(cudaStream_t stream, const bf16* dy, int ldy, const bf16* y, int ld, int rows, int cols, bf16* dz, int ldz);

}  // namespace dtb
