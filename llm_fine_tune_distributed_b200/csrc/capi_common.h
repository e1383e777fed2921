// Error plumbing shared by the C-ABI translation units.  Every entry point returns 0 on success
// and a negative code on failure; b200sft_last_error() returns the message for the calling thread.
#pragma once
#include <cuda_runtime.h>
#include <stdio.h>

#define B200SFT_OK 0
#define B200SFT_ERR_INVALID -1
#define B200SFT_ERR_CUDA -2
#define B200SFT_ERR_WORKSPACE -3
#define B200SFT_ERR_UNSUPPORTED -4

namespace b200 {

char* last_error_buf();  // thread-local, 512 bytes

template <class... Args>
inline int fail(int code, const char* fmt, Args... args) {
  snprintf(last_error_buf(), 512, fmt, args...);
  return code;
}
inline int fail(int code, const char* msg) {
  snprintf(last_error_buf(), 512, "%s", msg);
  return code;
}

#define B200_CUDA_CHECK(expr)                                                                              \
  do {                                                                                                     \
    cudaError_t _e = (expr);                                                                               \
    if (_e != cudaSuccess)                                                                                 \
      return ::b200::fail(B200SFT_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                          __LINE__);                                                                       \
  } while (0)

// elementwise.cu: stable compaction list of shifted int32 labels (the label front-end kernel, shift skipped)
int launch_compact_labels(const int* labels, int N, int ignore_index, int* row_index, int* labels_c, int* pos,
                          int* n_live, cudaStream_t stream);

}  // namespace b200
