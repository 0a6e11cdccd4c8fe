// Internal helpers shared by the translation units of liblvc_b200.so.
#pragma once
#include "../../include/lvc_b200.h"

// Records a thread-local message (printf-style) and returns `code`.
int lvc_fail(int code, const char* fmt, ...);

// Profiling knobs (LVC_CLIP_*, LVC_GEMM_*, LVC_PIL_*, LVC_NO_PDL, ...): environment variables that change the launch
// geometry or pick a diagnostic kernel variant.  They are honoured ONLY when LVC_PROFILING_KNOBS=1 was in the environment
// the first time the library looked (the gate is read once per process); otherwise no launch path reads the environment at all and a
// stray variable cannot change which kernel runs.  With the gate on, a knob is re-read at every launch (sweep scripts
// change them between launches).
bool lvc_knobs_enabled();
int lvc_knob_int(const char* name, int dflt);   // value of the knob, `dflt` when unset or the gate is off
bool lvc_knob_set(const char* name);            // knob present (any value) and the gate on

// Device address (current device) of the sticky error word the kernels OR contract violations into (LVC_DEVERR_*);
// read by lvc_device_status.  NULL when no device is usable.
unsigned* lvc_dev_error_word();

// The clip kernel as the producer half of lvc_clip_swin_embed_forward (defined in lvc_clip.cu, used by lvc_embed.cu).
int lvc_clip_forward_fused(const lvc_clip_desc* d, const lvc_video* videos, const int32_t* indices, const int32_t* boxes,
                           const float* lut, void* out, unsigned* tile_ready, int tile_shift, int smem_cap, void* stream);

#ifdef __CUDACC__
#define LVC_CUDA_TRY(expr)                                                                  \
  do {                                                                                      \
    cudaError_t e_ = (expr);                                                                \
    if (e_ != cudaSuccess)                                                                  \
      return lvc_fail(LVC_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), \
                      __FILE__, __LINE__);                                                  \
  } while (0)
#endif
