// GPU-side get_batch: the step in front of train_step().  Upstream (kzl experiment.py get_batch, restated in
// oracle/get_batch.py — its source is the un-vendored submodule, so parity is pinned to that restatement) samples (trajectory, start)
// pairs on the host, then for every sample slices a window of at most K steps, LEFT-pads it to K, normalises states, divides the
// return-to-go suffix sums by `scale`, clamps timesteps and builds the mask — dozens of small numpy ops and a host->device copy of the
// whole batch per step.  Here the trajectories live in HBM once (flat per-timestep rows + trajectory offsets + the reward suffix sums,
// computed once on the host with upstream's own loop), only the 2 B sampled indices cross PCIe, and one launch emits the batch in the
// layout the model reads.  Arithmetic follows numpy's promotions exactly (padding with float64 zeros promotes the window to float64
// before `(s - mean) / std` and `rtg / scale`, the result is then cast to fp32), so every output is bit-identical to the oracle's.
#include "kernels.h"

namespace dtb {

struct GetBatchArgs {
  dtb_traj_store st;
  const long long* traj_idx;
  const long long* start;
  int B, K, max_ep_len;
  double scale;
  dtb_batch_out out;
};

// one warp per output timestep (b, k); lanes walk the row's elements.  blockIdx.y = b.
__global__ void __launch_bounds__(256) get_batch_kernel(const GetBatchArgs a) {
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.y, K = a.K;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const long long tr = a.traj_idx[b], si = a.start[b];
  const long long t0 = a.st.traj_offset[tr], tlen_traj = a.st.traj_offset[tr + 1] - t0;
  long long tl = tlen_traj - si;          // steps left in the trajectory from the sampled start
  const int tlen = (int)(tl < K ? tl : K);  // window length (>= 1: si <= len - 1)
  const int pad = K - tlen;
  const int sd = a.st.state_dim, ad = a.st.act_dim;
  for (int k = blockIdx.x * nwarp + warp; k < K + 1; k += gridDim.x * nwarp) {
    const bool valid = k >= pad && k < K;
    const long long src = t0 + si + (k - pad);  // flat timestep index of output row k (when valid)
    if (k < K) {
      const size_t row = (size_t)b * K + k;
      if (a.out.states) {
        for (int j = lane; j < sd; j += 32) {
          const double v = valid ? (double)a.st.obs[(size_t)src * sd + j] : 0.0;  // padding is normalised too, as upstream does
          a.out.states[row * sd + j] = (float)((v - a.st.state_mean[j]) / a.st.state_std[j]);
        }
      }
      if (a.out.actions)
        for (int j = lane; j < ad; j += 32) a.out.actions[row * ad + j] = valid ? a.st.act[(size_t)src * ad + j] : -10.f;
      for (int f = 0; f < a.st.n_raw; ++f) {  // raw per-timestep rows (frames, spectrograms): copied, zero padding
        const int words = a.st.raw_row_bytes[f] >> 4;
        const uint4* s4 = reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(a.st.raw_src[f]) + (size_t)(valid ? src : 0) * a.st.raw_row_bytes[f]);
        uint4* d4 = reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(a.out.raw_dst[f]) + row * a.st.raw_row_bytes[f]);
        for (int j = lane; j < words; j += 32) d4[j] = valid ? __ldg(s4 + j) : make_uint4(0, 0, 0, 0);
      }
      if (lane == 0) {
        if (a.out.rewards) a.out.rewards[row] = valid ? a.st.rew[src] : 0.f;
        if (a.out.dones) a.out.dones[row] = valid ? (long long)a.st.done[src] : 2LL;
        if (a.out.timesteps) {
          long long ts = valid ? si + (k - pad) : 0;
          if (ts >= a.max_ep_len) ts = a.max_ep_len - 1;  // "padding cutoff"
          a.out.timesteps[row] = ts;
        }
        if (a.out.mask) a.out.mask[row] = valid ? 1.f : 0.f;
      }
    }
    // returns-to-go: K + 1 entries, entry k >= pad is the suffix sum at step si + (k - pad), 0 once past the trajectory's end
    if (lane == 0 && a.out.rtg) {
      double r = 0.0;
      if (k >= pad && si + (k - pad) < tlen_traj) r = a.st.rtg_suffix[src];
      a.out.rtg[(size_t)b * (K + 1) + k] = (float)(r / a.scale);
    }
  }
}

int get_batch(cudaStream_t stream, const dtb_traj_store* st, const long long* traj_idx, const long long* start, int B, int K, int max_ep_len,
              double scale, const dtb_batch_out* out) {
  DTB_REQUIRE(st && traj_idx && start && out && B >= 0 && K >= 1 && max_ep_len >= 1 && scale != 0.0, DTB_ERR_BAD_ARG, "get_batch: bad arguments");
  DTB_REQUIRE(st->traj_offset != nullptr, DTB_ERR_BAD_ARG, "get_batch: trajectory offsets missing");
  DTB_REQUIRE(!out->states || (st->obs && st->state_mean && st->state_std && st->state_dim >= 1), DTB_ERR_BAD_ARG, "get_batch: states need obs / mean / std");
  DTB_REQUIRE(!out->actions || (st->act && st->act_dim >= 1), DTB_ERR_BAD_ARG, "get_batch: actions need the action store");
  DTB_REQUIRE(!out->rewards || st->rew, DTB_ERR_BAD_ARG, "get_batch: rewards need the reward store");
  DTB_REQUIRE(!out->dones || st->done, DTB_ERR_BAD_ARG, "get_batch: dones need the done store");
  DTB_REQUIRE(!out->rtg || st->rtg_suffix, DTB_ERR_BAD_ARG, "get_batch: rtg needs the suffix sums");
  DTB_REQUIRE(st->n_raw >= 0 && st->n_raw <= 4, DTB_ERR_BAD_ARG, "get_batch: at most 4 raw fields");
  for (int f = 0; f < st->n_raw; ++f)
    DTB_REQUIRE(st->raw_src[f] && out->raw_dst[f] && st->raw_row_bytes[f] > 0 && st->raw_row_bytes[f] % 16 == 0, DTB_ERR_BAD_ARG,
                "get_batch: raw field %d needs 16-byte-multiple rows and non-null pointers", f);
  if (B == 0) return DTB_OK;
  GetBatchArgs a;
  memset(&a, 0, sizeof(a));
  a.st = *st; a.traj_idx = traj_idx; a.start = start; a.B = B; a.K = K; a.max_ep_len = max_ep_len; a.scale = scale; a.out = *out;
  const int bx = (K + 1 + 7) / 8;
  DTB_LAUNCH(get_batch_kernel, dim3(bx, B), 256, 0, stream, a);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

}  // namespace dtb
