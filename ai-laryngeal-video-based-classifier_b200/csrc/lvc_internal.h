// Internal helpers shared by the translation units of liblvc_b200.so.
#pragma once
#include "../../include/lvc_b200.h"

// Records a thread-local message (printf-style) and returns `code`.
int lvc_fail(int code, const char* fmt, ...);

#ifdef __CUDACC__
#define LVC_CUDA_TRY(expr)                                                                  \
  do {                                                                                      \
    cudaError_t e_ = (expr);                                                                \
    if (e_ != cudaSuccess)                                                                  \
      return lvc_fail(LVC_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), \
                      __FILE__, __LINE__);                                                  \
  } while (0)
#endif
