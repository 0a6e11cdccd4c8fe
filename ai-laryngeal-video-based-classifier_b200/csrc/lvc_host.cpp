// Error reporting and the host-side normalisation look-up table.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "lvc_internal.h"

namespace {
thread_local char g_err[512] = "";
}

int lvc_fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
  return code;
}

bool lvc_knobs_enabled() {
  static const bool on = [] {
    const char* v = getenv("LVC_PROFILING_KNOBS");
    return v && v[0] == '1' && v[1] == 0;
  }();
  return on;
}
int lvc_knob_int(const char* name, int dflt) {
  if (!lvc_knobs_enabled()) return dflt;
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}
bool lvc_knob_set(const char* name) { return lvc_knobs_enabled() && getenv(name) != nullptr; }

extern "C" {

const char* lvc_last_error(void) { return g_err; }
int lvc_version(void) { return 100; }

// Every step is rounded to float32 exactly where the reference rounds it (compiled with
// -ffp-contract=off): resnet50-2d-lstm/inference.py:166-174 for LVC_LUT_DIV; transformers
// image_transforms.rescale (float64 multiply, cast to float32) + normalize for LVC_LUT_MUL64.
int lvc_build_lut(int kind, double scale, int offset, const float mean[3], const float std[3],
                  float* out) {
  if (!mean || !std || !out) return lvc_fail(LVC_E_INVAL, "lvc_build_lut: NULL argument");
  if (kind != LVC_LUT_DIV && kind != LVC_LUT_MUL64) return lvc_fail(LVC_E_INVAL, "lvc_build_lut: unknown kind %d", kind);
  for (int c = 0; c < 3; ++c) {
    if (std[c] == 0.0f) return lvc_fail(LVC_E_INVAL, "lvc_build_lut: std[%d] is zero", c);
    for (int v = 0; v < 256; ++v) {
      volatile float x;
      if (kind == LVC_LUT_DIV) {
        x = (float)v / (float)scale;
      } else {
        volatile double d = (double)v * scale;
        x = (float)d;
      }
      if (offset) x = x - 1.0f;
      x = x - mean[c];
      x = x / std[c];
      out[c * 256 + v] = x;
    }
  }
  return LVC_OK;
}

}  // extern "C"
