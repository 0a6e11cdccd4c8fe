// Host-side frame-index samplers with bit-exact CPython `random` and numpy-legacy RNG streams.
// 32 integers per video: the host is the right place; the table is uploaded once per batch.
// Compiled with -ffp-contract=off so float64 expressions round exactly like CPython's.
//
// Reference behaviour followed (paths relative to the reference root):
//   vivit_transformer/vivit_classifier/data_config/dataset.py:129-193   get_sampling_indices
//   visualize_sampling.py:105-178                                      pure-function copies
//   resnet50-2d-lstm/inference.py:74-130                               numpy-RNG copies
//   data_handling.py:41-56                                             sample_frame_indices
// RNG algorithms restated from their published definitions (Matsumoto & Nishimura MT19937;
// CPython Lib/random.py `_randbelow_with_getrandbits`, `sample`, `choices`; numpy legacy
// `RandomState.randint` masked rejection and `shuffle`).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

#include "lvc_internal.h"

namespace {

constexpr int kN = 624, kM = 397;

void init_genrand(lvc_rng* st, uint32_t s) {
  st->mt[0] = s;
  for (int i = 1; i < kN; ++i)
    st->mt[i] = 1812433253u * (st->mt[i - 1] ^ (st->mt[i - 1] >> 30)) + (uint32_t)i;
  st->pos = kN;
}

void init_by_array(lvc_rng* st, const uint32_t* key, int len) {
  init_genrand(st, 19650218u);
  uint32_t* mt = st->mt;
  int i = 1, j = 0;
  for (int k = (kN > len ? kN : len); k; --k) {
    mt[i] = (mt[i] ^ ((mt[i - 1] ^ (mt[i - 1] >> 30)) * 1664525u)) + key[j] + (uint32_t)j;
    if (++i >= kN) { mt[0] = mt[kN - 1]; i = 1; }
    if (++j >= len) j = 0;
  }
  for (int k = kN - 1; k; --k) {
    mt[i] = (mt[i] ^ ((mt[i - 1] ^ (mt[i - 1] >> 30)) * 1566083941u)) - (uint32_t)i;
    if (++i >= kN) { mt[0] = mt[kN - 1]; i = 1; }
  }
  mt[0] = 0x80000000u;
  st->pos = kN;
}

uint32_t next_u32(lvc_rng* st) {
  uint32_t* mt = st->mt;
  if (st->pos >= kN) {
    int kk = 0;
    for (; kk < kN - kM; ++kk) {
      uint32_t y = (mt[kk] & 0x80000000u) | (mt[kk + 1] & 0x7fffffffu);
      mt[kk] = mt[kk + kM] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    for (; kk < kN - 1; ++kk) {
      uint32_t y = (mt[kk] & 0x80000000u) | (mt[kk + 1] & 0x7fffffffu);
      mt[kk] = mt[kk + (kM - kN)] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    uint32_t y = (mt[kN - 1] & 0x80000000u) | (mt[0] & 0x7fffffffu);
    mt[kN - 1] = mt[kM - 1] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    st->pos = 0;
  }
  uint32_t y = mt[st->pos++];
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}

// ---- CPython random ----
double py_random(lvc_rng* st) {
  uint32_t a = next_u32(st) >> 5, b = next_u32(st) >> 6;
  return (a * 67108864.0 + b) * (1.0 / 9007199254740992.0);
}

int bit_length(uint64_t n) {
  int k = 0;
  while (n) { ++k; n >>= 1; }
  return k;
}

// getrandbits(k), k <= 64 (CPython fills 32-bit words least-significant first)
uint64_t py_getrandbits(lvc_rng* st, int k) {
  if (k <= 32) return next_u32(st) >> (32 - k);
  uint64_t lo = next_u32(st);
  uint64_t hi = next_u32(st) >> (64 - k);
  return lo | (hi << 32);
}

uint64_t py_randbelow(lvc_rng* st, uint64_t n) {
  int k = bit_length(n);
  uint64_t r = py_getrandbits(st, k);
  while (r >= n) r = py_getrandbits(st, k);
  return r;
}

// ---- numpy legacy ----
uint64_t np_bounded(lvc_rng* st, uint64_t rng) {  // uniform in [0, rng], masked rejection
  if (rng == 0) return 0;
  uint64_t mask = rng;
  mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4;
  mask |= mask >> 8; mask |= mask >> 16; mask |= mask >> 32;
  uint64_t v;
  if (rng <= 0xffffffffull) {
    if (rng == 0xffffffffull) return next_u32(st);
    while ((v = (next_u32(st) & mask)) > rng) {}
  } else {
    // 64-bit draws: high word first (numpy mt19937_next64)
    for (;;) {
      uint64_t hi = next_u32(st), lo = next_u32(st);
      v = ((hi << 32) | lo) & mask;
      if (v <= rng) break;
    }
  }
  return v;
}

}  // namespace

extern "C" {

int lvc_rng_seed_py(lvc_rng* st, uint64_t seed) {
  if (!st) return lvc_fail(LVC_E_INVAL, "lvc_rng_seed_py: NULL state");
  uint32_t key[2] = {(uint32_t)(seed & 0xffffffffu), (uint32_t)(seed >> 32)};
  init_by_array(st, key, key[1] ? 2 : 1);
  return LVC_OK;
}

int lvc_rng_seed_np(lvc_rng* st, uint32_t seed) {
  if (!st) return lvc_fail(LVC_E_INVAL, "lvc_rng_seed_np: NULL state");
  init_genrand(st, seed);
  return LVC_OK;
}

double lvc_rng_random(lvc_rng* st) { return py_random(st); }
uint32_t lvc_rng_u32(lvc_rng* st) { return next_u32(st); }

// ---- torch CPU generator stream (at::mt19937 seeded by torch.manual_seed) -------------------
// torch.randint(low, high, (1,)): ATen random_from_to -> uniform_int_from_to(random(), range, base)
// = uint32 draw % range + base for range < 2^32.
int lvc_torch_randint(lvc_rng* st, int64_t low, int64_t high, int64_t* out) {
  if (!st || !out) return lvc_fail(LVC_E_INVAL, "lvc_torch_randint: NULL argument");
  if (high <= low || (uint64_t)(high - low) >= (1ull << 32)) return lvc_fail(LVC_E_INVAL, "lvc_torch_randint: need 0 < high - low < 2^32");
  *out = (int64_t)((uint64_t)next_u32(st) % (uint64_t)(high - low)) + low;
  return LVC_OK;
}
// torch.rand(1) (float32): uniform_real(random(), 0, 1) = (uint32 draw & (2^24 - 1)) * 2^-24.
float lvc_torch_rand(lvc_rng* st) { return st ? (float)(next_u32(st) & ((1u << 24) - 1u)) * (1.0f / 16777216.0f) : -1.0f; }

// One clip's draws of the train chain in the reference's order (videoswintransformer/.../dataset.py:156-161):
// RandomShortSideScale: size = randint(min, max+1); RandomCrop.get_params on the scaled (nh, nw): if it is
// already 224x224 no draw, else top = randint(0, nh-224+1), left = randint(0, nw-224+1);
// RandomHorizontalFlip: flip = rand(1) < p.   out = {size, top, left, flip}.
int lvc_train_aug_draw(lvc_rng* st, int32_t height, int32_t width, int32_t min_size, int32_t max_size,
                       double flip_p, int32_t* out) {
  if (!st || !out) return lvc_fail(LVC_E_INVAL, "lvc_train_aug_draw: NULL argument");
  if (height <= 0 || width <= 0 || min_size < 224 || max_size < min_size) return lvc_fail(LVC_E_INVAL, "lvc_train_aug_draw: need positive extents and 224 <= min_size <= max_size");
  int64_t size = 0, top = 0, left = 0;
  lvc_torch_randint(st, min_size, (int64_t)max_size + 1, &size);
  int64_t nh, nw;
  if (width < height) { nw = size; nh = (int64_t)std::floor(((double)height / (double)width) * (double)size); }
  else { nh = size; nw = (int64_t)std::floor(((double)width / (double)height) * (double)size); }
  if (!(nh == 224 && nw == 224)) {
    lvc_torch_randint(st, 0, nh - 224 + 1, &top);
    lvc_torch_randint(st, 0, nw - 224 + 1, &left);
  }
  const float u = lvc_torch_rand(st);
  out[0] = (int32_t)size; out[1] = (int32_t)top; out[2] = (int32_t)left; out[3] = (u < (float)flip_p) ? 1 : 0;
  return LVC_OK;
}

double lvc_dynamic_fps(int64_t total_frames, int32_t num_frames, double original_fps) {
  if (total_frames >= num_frames) return std::numeric_limits<double>::quiet_NaN();
  if (original_fps <= 0) original_fps = 30.0;
  double duration = (double)total_frames / original_fps;
  return (double)num_frames / duration;
}

int lvc_sample_indices(int method, int variant, int64_t total_frames, int32_t num_frames,
                       lvc_rng* st, int64_t* out, int32_t* n_out) {
  if (!out || !n_out) return lvc_fail(LVC_E_INVAL, "lvc_sample_indices: NULL output");
  if (total_frames <= 0 || num_frames <= 0)
    return lvc_fail(LVC_E_INVAL, "lvc_sample_indices: total_frames and num_frames must be positive");
  if (method < LVC_SAMPLE_UNIFORM || method > LVC_SAMPLE_RANDOM)
    return lvc_fail(LVC_E_INVAL, "lvc_sample_indices: unknown sampling method");
  if (variant < LVC_VARIANT_DATASET || variant > LVC_VARIANT_NUMPY_INFER)
    return lvc_fail(LVC_E_INVAL, "lvc_sample_indices: unknown variant");
  if (method != LVC_SAMPLE_UNIFORM && !st)
    return lvc_fail(LVC_E_INVAL, "lvc_sample_indices: random methods need an RNG state");
  const int64_t N = total_frames;
  int64_t T = num_frames;
  if (variant != LVC_VARIANT_DATASET && T > N) T = N;
  const bool np_stream = variant == LVC_VARIANT_NUMPY_INFER;
  *n_out = (int32_t)T;

  if (N >= T) {
    if (method == LVC_SAMPLE_RANDOM) {
      if (np_stream) {
        // np.random.choice(N, T, replace=False) = permutation(N)[:T]; legacy shuffle
        std::vector<int64_t> p((size_t)N);
        for (int64_t i = 0; i < N; ++i) p[(size_t)i] = i;
        for (int64_t i = N - 1; i >= 1; --i) {
          int64_t j = (int64_t)np_bounded(st, (uint64_t)i);
          std::swap(p[(size_t)i], p[(size_t)j]);
        }
        for (int64_t i = 0; i < T; ++i) out[i] = p[(size_t)i];
      } else {
        // random.sample(range(N), T): pool for small populations, rejection set otherwise
        int64_t setsize = 21;
        if (T > 5) setsize += (int64_t)std::pow(4.0, std::ceil(std::log(3.0 * (double)T) / std::log(4.0)));
        if (N <= setsize) {
          std::vector<int64_t> pool((size_t)N);
          for (int64_t i = 0; i < N; ++i) pool[(size_t)i] = i;
          for (int64_t i = 0; i < T; ++i) {
            int64_t j = (int64_t)py_randbelow(st, (uint64_t)(N - i));
            out[i] = pool[(size_t)j];
            pool[(size_t)j] = pool[(size_t)(N - i - 1)];
          }
        } else {
          for (int64_t i = 0; i < T; ++i) {
            int64_t j;
            bool dup;
            do {
              j = (int64_t)py_randbelow(st, (uint64_t)N);
              dup = false;
              for (int64_t q = 0; q < i; ++q) dup |= (out[q] == j);
            } while (dup);
            out[i] = j;
          }
        }
      }
      std::sort(out, out + T);
    } else if (method == LVC_SAMPLE_RANDOM_WINDOW) {
      double w = (double)N / (double)T;
      for (int64_t i = 0; i < T; ++i) {
        int64_t start = (int64_t)((double)i * w);
        int64_t end = std::min((int64_t)((double)(i + 1) * w), N);
        end = std::max(end, start + 1);
        uint64_t width = (uint64_t)(end - start);
        out[i] = start + (int64_t)(np_stream ? np_bounded(st, width - 1) : py_randbelow(st, width));
      }
    } else {
      if (T == 1) {
        out[0] = N / 2;
      } else {
        double step = (double)(N - 1) / (double)(T - 1);
        for (int64_t i = 0; i < T; ++i) out[i] = std::min((int64_t)((double)i * step), N - 1);
      }
    }
    return LVC_OK;
  }
  // short videos (N < T): only the DATASET variant gets here
  if (method == LVC_SAMPLE_RANDOM) {
    double n = (double)N + 0.0;
    for (int64_t i = 0; i < T; ++i) out[i] = (int64_t)std::floor(py_random(st) * n);
    std::sort(out, out + T);
  } else if (method == LVC_SAMPLE_RANDOM_WINDOW) {
    double w = (double)N / (double)T;
    for (int64_t i = 0; i < T; ++i) {
      double vs = (double)i * w;
      double ve = (double)(i + 1) * w;
      double r = py_random(st);
      double d = (ve - vs) * r;
      out[i] = std::min((int64_t)std::floor(vs + d), N - 1);
    }
  } else {
    if (T == 1) {
      out[0] = N / 2;
    } else {
      double step = (double)N / (double)T;
      for (int64_t i = 0; i < T; ++i) out[i] = std::min((int64_t)((double)i * step), N - 1);
    }
  }
  return LVC_OK;
}

int lvc_sample_frame_indices(int32_t clip_len, double frame_sample_rate, int64_t seg_len,
                             lvc_rng* st, int64_t* out) {
  if (!st || !out || clip_len <= 0) return lvc_fail(LVC_E_INVAL, "lvc_sample_frame_indices: bad argument");
  int64_t conv = (int64_t)((double)clip_len * frame_sample_rate);
  if (seg_len <= conv)
    return lvc_fail(LVC_E_INVAL, "lvc_sample_frame_indices: low >= high (numpy raises ValueError)");
  int64_t end = conv + (int64_t)np_bounded(st, (uint64_t)(seg_len - 1 - conv));
  int64_t start = end - conv;
  // np.linspace(start, end, num): arange(num) * step + start, last element = stop
  double step = clip_len > 1 ? ((double)end - (double)start) / (double)(clip_len - 1) : 0.0;
  for (int32_t i = 0; i < clip_len; ++i) {
    double v = (double)i * step;
    v = v + (double)start;
    if (clip_len > 1 && i == clip_len - 1) v = (double)end;
    v = std::min(std::max(v, (double)start), (double)(end - 1));
    out[i] = (int64_t)v;
  }
  return LVC_OK;
}

int lvc_uniform_temporal_subsample(int64_t t, int32_t num_samples, int64_t* out) {
  if (!out || t <= 0 || num_samples <= 0)
    return lvc_fail(LVC_E_INVAL, "lvc_uniform_temporal_subsample: bad argument");
  // torch.linspace(0, t-1, n) in float32: symmetric evaluation from both ends
  float start = 0.0f, end = (float)(t - 1);
  if (num_samples == 1) { out[0] = 0; return LVC_OK; }
  float step = (end - start) / (float)(num_samples - 1);
  int64_t half = num_samples / 2;
  for (int64_t i = 0; i < num_samples; ++i) {
    float v = i < half ? start + step * (float)i : end - step * (float)(num_samples - i - 1);
    v = std::min(std::max(v, 0.0f), end);
    out[i] = (int64_t)v;
  }
  return LVC_OK;
}

}  // extern "C"
