// Fork/join helper: weight-gradient work (GEMM + split-K reduce + bias column sums) runs on a library-owned side stream while
// the data-gradient chain continues on the caller's stream.  Every API call joins before it returns, so the caller still sees plain
// stream-ordered semantics (and a CUDA-graph capture of the call records the fork/join as parallel branches).
#pragma once
#include "common.cuh"
#include <stdlib.h>

namespace dtb {

struct SideStream {
  cudaStream_t s = nullptr;
  cudaEvent_t fork_ev = nullptr, join_ev = nullptr, ev_a = nullptr, ev_b = nullptr;
  int state = 0;  // 0 = not created, 1 = ready, -1 = disabled (DTB_NO_SIDE_STREAM=1)

  int init() {
    if (state != 0) return DTB_OK;
    const char* off = getenv("DTB_NO_SIDE_STREAM");
    if (off && off[0] == '1') { state = -1; return DTB_OK; }
    DTB_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    DTB_CUDA(cudaEventCreateWithFlags(&fork_ev, cudaEventDisableTiming));
    DTB_CUDA(cudaEventCreateWithFlags(&join_ev, cudaEventDisableTiming));
    DTB_CUDA(cudaEventCreateWithFlags(&ev_a, cudaEventDisableTiming));
    DTB_CUDA(cudaEventCreateWithFlags(&ev_b, cudaEventDisableTiming));
    state = 1;
    return DTB_OK;
  }
  bool on() const { return state == 1; }
  // stream the weight-gradient work should be enqueued on (the main stream when disabled)
  cudaStream_t pick(cudaStream_t main) const { return on() ? s : main; }
  // side waits for everything enqueued on main so far
  int fork(cudaStream_t main) {
    if (!on()) return DTB_OK;
    DTB_CUDA(cudaEventRecord(fork_ev, main));
    DTB_CUDA(cudaStreamWaitEvent(s, fork_ev, 0));
    return DTB_OK;
  }
  // record a side-progress marker / make main wait for it
  int mark(cudaEvent_t ev) {
    if (!on()) return DTB_OK;
    DTB_CUDA(cudaEventRecord(ev, s));
    return DTB_OK;
  }
  int wait(cudaStream_t main, cudaEvent_t ev) {
    if (!on()) return DTB_OK;
    DTB_CUDA(cudaStreamWaitEvent(main, ev, 0));
    return DTB_OK;
  }
  int join(cudaStream_t main) {
    if (!on()) return DTB_OK;
    DTB_CUDA(cudaEventRecord(join_ev, s));
    DTB_CUDA(cudaStreamWaitEvent(main, join_ev, 0));
    return DTB_OK;
  }
  void destroy() {
    if (state == 1) {
      cudaEventDestroy(fork_ev); cudaEventDestroy(join_ev); cudaEventDestroy(ev_a); cudaEventDestroy(ev_b);
      cudaStreamDestroy(s);
    }
    state = 0;
  }
};

}  // namespace dtb
