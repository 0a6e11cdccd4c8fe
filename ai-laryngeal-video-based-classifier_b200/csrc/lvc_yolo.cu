// Detector pre- and post-processing around the (out-of-scope) YOLOv5 network, SURVEY.md 8f.4:
//   letterbox 640 + BGR->RGB + /255 + CHW      src/yolo_detector.py:52-61
//   scale_boxes(...).round() + int()           src/yolo_detector.py:76-81
//   highest-confidence box per frame           src/yolo_detector.py:161 (the reference gates on max confidence)
// so that ROI boxes are produced and consumed on the device: the int32 rows these kernels write are rows of
// lvc_clip_forward's `boxes` table.  yolov5's letterbox / scale_boxes are restated from the published
// ultralytics/yolov5 definitions (utils/augmentations.py, utils/general.py): the reference's yolov5/ submodule
// is empty and unpinned.  Byte / integer work, HBM-bound: no tensor cores.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <type_traits>

#include "lvc_internal.h"

namespace {

constexpr int kLbThreads = 256;
#ifndef LVC_LB_ROWS
#define LVC_LB_ROWS 8
#endif
constexpr int kLbRows = LVC_LB_ROWS;        // output rows per work item (the grid is persistent: blocks walk the items round-robin)
constexpr int kLbMaxW = 4096;     // widest resize target the tap table holds
constexpr int kLbUnroll = 3;      // pixels per thread and pass (768 columns per pass)

struct XTap { int off; int a01; };  // byte offset of the left pixel; a0 | a1 << 16; bit 31 of off: right tap clamps

// OpenCV INTER_LINEAR table set-up (cv::resize, 8-bit): fx = (float)((d+0.5)*scale - 0.5) in double without
// contraction, floor / fraction in float, 11-bit coefficients.
__device__ __forceinline__ void cv_tap(int d, double scale, int& s, float& fr) {
  const float f = (float)__dadd_rn(__dmul_rn((double)d + 0.5, scale), -0.5);
  const float fl = floorf(f);
  s = (int)fl;
  fr = __fsub_rn(f, fl);
}
__device__ __forceinline__ double cv_scale(int src, int dst) { return __ddiv_rn(1.0, __ddiv_rn((double)dst, (double)src)); }
__device__ __forceinline__ int coef(float w) { return __float2int_rn(__fmul_rn(w, 2048.0f)); }

struct LbParams {
  lvc_letterbox_geom g;
  const uint8_t* frames;
  void* out;
  int n, out_bf16;
  int row_blocks, items;   // items = n * row_blocks, item = frame * row_blocks + row block
  int rows, pitch;         // staged kernel: output rows per item, bytes per source-row slot
};

__global__ void __launch_bounds__(kLbThreads) lvc_letterbox_direct_kernel(const __grid_constant__ LbParams p) {
  __shared__ XTap xt[kLbMaxW];
  const lvc_letterbox_geom& g = p.g;
  const bool area = g.src_w == 2 * g.new_w && g.src_h == 2 * g.new_h;   // cv::resize switches to INTER_AREA
  const double sx_scale = cv_scale(g.src_w, g.new_w), sy_scale = cv_scale(g.src_h, g.new_h);
  for (int x = threadIdx.x; x < g.new_w; x += kLbThreads) {
    XTap t;
    if (area) { t.off = 6 * x; t.a01 = 1 | (1 << 16); }
    else {
      int sx; float fr;
      cv_tap(x, sx_scale, sx, fr);
      if (sx < 0) { sx = 0; fr = 0.f; }
      if (sx >= g.src_w - 1) { sx = g.src_w - 1; fr = 0.f; }
      t.off = 3 * sx | ((sx + 1 <= g.src_w - 1) ? 0 : (1 << 31));
      t.a01 = coef(__fsub_rn(1.0f, fr)) | (coef(fr) << 16);
    }
    xt[x] = t;
  }
  // x / 255 for every uint8 value (IEEE division = what torch's CPU `im /= 255` computes), once per block
  __shared__ float s_div[256];
  __shared__ int4 s_row[2][kLbRows];   // y0, y1, b0, b1 of the item's rows (y0 < 0: border row), double-buffered over items
  for (int i = threadIdx.x; i < 256; i += kLbThreads) s_div[i] = __fdiv_rn((float)i, 255.0f);
  const uint8_t* frames_end = p.frames + (size_t)p.n * g.src_h * g.src_w * 3;
  const size_t plane = (size_t)g.out_h * g.out_w;
  // Persistent grid: the tables above are built once per block (they depend on the geometry only), the blocks then walk
  // the (frame, row block) items round-robin -- one wave of resident blocks instead of 1.5 waves of 16-row blocks.
  int buf = 0;
  for (int item = blockIdx.x; item < p.items; item += gridDim.x, buf ^= 1) {
  const int frame = item / p.row_blocks, rblk = item - frame * p.row_blocks;
  if (threadIdx.x < kLbRows) {
    const int iy = rblk * kLbRows + threadIdx.x - g.top;
    int4 rt = make_int4(-1, 0, 0, 0);
    if (iy >= 0 && iy < g.new_h) {
      if (area) rt = make_int4(2 * iy, 2 * iy + 1, 0, 0);
      else {
        int sy; float fr;
        cv_tap(iy, sy_scale, sy, fr);
        // cv::resize does not reset the fraction of clamped ROWS
        rt = make_int4(min(max(sy, 0), g.src_h - 1), min(max(sy + 1, 0), g.src_h - 1), coef(__fsub_rn(1.0f, fr)), coef(fr));
      }
    }
    s_row[buf][threadIdx.x] = rt;
  }
  __syncthreads();   // also publishes xt / s_div on the first trip; the other buffer is free (everyone passed the previous sync)
  const uint8_t* src = p.frames + (size_t)frame * g.src_h * g.src_w * 3;
  const float pad = s_div[114];
  for (int r = 0; r < kLbRows; ++r) {
    const int oy = rblk * kLbRows + r;
    if (oy >= g.out_h) break;
    const int4 rt = s_row[buf][r];
    const bool row_in = rt.x >= 0;
    const int y0 = max(rt.x, 0), y1 = rt.y, b0 = rt.z, b1 = rt.w;
    const uint8_t* r0 = src + (size_t)y0 * g.src_w * 3;
    const uint8_t* r1 = src + (size_t)y1 * g.src_w * 3;
    // kLbUnroll pixels per thread and pass: all source words are requested before the first one is used, so
    // a thread keeps up to 18 loads in flight (the kernel is latency-bound otherwise)
    for (int ox0 = threadIdx.x; ox0 < g.out_w; ox0 += kLbThreads * kLbUnroll) {
      uint32_t lo0[kLbUnroll], hi0[kLbUnroll], lo1[kLbUnroll], hi1[kLbUnroll], a01[kLbUnroll];
      bool in[kLbUnroll];
#pragma unroll
      for (int u = 0; u < kLbUnroll; ++u) {
        const int ox = ox0 + u * kLbThreads, ix = ox - g.left;
        in[u] = row_in && ox < g.out_w && ix >= 0 && ix < g.new_w;
        lo0[u] = hi0[u] = lo1[u] = hi1[u] = a01[u] = 0u;
        if (!in[u]) continue;
        const XTap t = xt[ix];
        const int off = t.off & 0x7fffffff, d1 = (t.off < 0) ? 0 : 3;
        a01[u] = (uint32_t)t.a01;
        const uint8_t* q0 = r0 + off;
        const uint8_t* q1 = r1 + off;
        if (d1 == 3 && q0 + 12 <= frames_end && q1 + 12 <= frames_end) {
          // 6 bytes (2 pixels x 3 channels) per row from three aligned 32-bit loads and a funnel shift
          const uint32_t sh0 = ((uint32_t)(uintptr_t)q0 & 3u) * 8u, sh1 = ((uint32_t)(uintptr_t)q1 & 3u) * 8u;
          const uint32_t* w0 = reinterpret_cast<const uint32_t*>((uintptr_t)q0 & ~(uintptr_t)3);
          const uint32_t* w1 = reinterpret_cast<const uint32_t*>((uintptr_t)q1 & ~(uintptr_t)3);
          const uint32_t a = __ldg(w0), bq = __ldg(w0 + 1), cq = __ldg(w0 + 2), d = __ldg(w1), e = __ldg(w1 + 1), fq = __ldg(w1 + 2);
          lo0[u] = __funnelshift_r(a, bq, sh0); hi0[u] = __funnelshift_r(bq, cq, sh0);
          lo1[u] = __funnelshift_r(d, e, sh1); hi1[u] = __funnelshift_r(e, fq, sh1);
        } else {   // clamped right edge or the last bytes of the buffer: byte loads into the same layout
          lo0[u] = __ldg(q0) | (__ldg(q0 + 1) << 8) | (__ldg(q0 + 2) << 16) | (__ldg(q0 + d1) << 24);
          hi0[u] = __ldg(q0 + d1 + 1) | (__ldg(q0 + d1 + 2) << 8);
          lo1[u] = __ldg(q1) | (__ldg(q1 + 1) << 8) | (__ldg(q1 + 2) << 16) | (__ldg(q1 + d1) << 24);
          hi1[u] = __ldg(q1 + d1 + 1) | (__ldg(q1 + d1 + 2) << 8);
        }
      }
#pragma unroll
      for (int u = 0; u < kLbUnroll; ++u) {
        const int ox = ox0 + u * kLbThreads;
        if (ox >= g.out_w) continue;
        float v[3] = {pad, pad, pad};
        if (in[u]) {
          const uint32_t sel[3] = {0x4430u, 0x4441u, 0x4452u};   // (c of pixel 0, c of pixel 1) of the 8 bytes
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const int s0 = (int)__dp2a_lo(a01[u], __byte_perm(lo0[u], hi0[u], sel[c]), 0u);
            const int s1 = (int)__dp2a_lo(a01[u], __byte_perm(lo1[u], hi1[u], sel[c]), 0u);
            const int uu = area ? ((s0 + s1 + 2) >> 2) : ((((b0 * (s0 >> 4)) >> 16) + ((b1 * (s1 >> 4)) >> 16) + 2) >> 2);
            v[c] = s_div[uu];
          }
        }
        // transpose(2,0,1)[::-1]: output channel k is input channel 2 - k
        const size_t e = (size_t)frame * 3 * plane + (size_t)oy * g.out_w + ox;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          if (p.out_bf16) reinterpret_cast<__nv_bfloat16*>(p.out)[e + k * plane] = __float2bfloat16_rn(v[2 - k]);
          else __stcs(reinterpret_cast<float*>(p.out) + e + k * plane, v[2 - k]);
        }
      }
    }
  }
  }   // items
}

// ---- staged variant: the item's tapped source rows travel HBM -> shared memory as whole-row bulk copies (cp.async.bulk,
// the TMA engine; SASS UBLKCP) on an mbarrier, double-buffered over items, and the taps are shared-memory word loads.
// The direct kernel above issues 6 scattered 4-byte global loads per output pixel and is latency-bound at half the
// copy rate; whole rows keep the memory system streaming (the clip kernel's ingest, csrc/lvc_clip.cu).
__device__ __forceinline__ uint32_t lb_s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void lb_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(lb_s32(bar)), "r"(count));
}
__device__ __forceinline__ void lb_mbar_expect(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(lb_s32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void lb_mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(lb_s32(bar)), "r"(parity), "r"(0x989680u) : "memory");
  } while (!done);
}
__device__ __forceinline__ void lb_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(lb_s32(dst)), "l"(src), "r"(bytes), "r"(lb_s32(bar)) : "memory");
}
constexpr int kLbMaxRows = 8;   // output rows per item of the staged kernel (2 * rows source-row slots <= 16 lanes)

constexpr int kLbFixed = 16 + 4 * 1024 + 2 * kLbMaxRows * 16;   // barriers, value table, row tables (bytes before xt[])
__host__ __device__ inline size_t lb_slots_offset(int new_w) { return ((size_t)kLbFixed + (size_t)new_w * 8 + 127) & ~(size_t)127; }

// One output pixel of the staged kernel from its two tapped rows in shared memory: three aligned words per row + funnel
// shift, PRMT into (c0 p0, c0 p1, c1 p0, c1 p1) / (c2 p0, c2 p1), cv2's exact 11-bit horizontal pass as DP2A, its vertical
// pass ((b * (S >> 4)) >> 16) as one multiply-high per term (the row coefficients arrive as b << 16).  The exact-2x
// INTER_AREA case is the same code with a0 = a1 = 32, b = 2^31: (p00 + p01 + p10 + p11 + 2) >> 2.  Returns the UNROUNDED
// vertical sums t[c] (<= 1020): the value table is indexed by them and folds cv2's (+ 2) >> 2.
__device__ __forceinline__ void lb_pixel(uint32_t q0, uint32_t q1, uint32_t a01, uint32_t b0s, uint32_t b1s, unsigned (&t)[3]) {
  uint32_t a, bq, cq, d, e, fq;
  asm volatile("ld.shared.u32 %0, [%3];\n\tld.shared.u32 %1, [%3+4];\n\tld.shared.u32 %2, [%3+8];"
               : "=r"(a), "=r"(bq), "=r"(cq) : "r"(q0 & ~3u));
  asm volatile("ld.shared.u32 %0, [%3];\n\tld.shared.u32 %1, [%3+4];\n\tld.shared.u32 %2, [%3+8];"
               : "=r"(d), "=r"(e), "=r"(fq) : "r"(q1 & ~3u));
  const uint32_t sh0 = q0 << 3, sh1 = q1 << 3;   // funnel shifts wrap at 32: (q & 3) * 8
  const uint32_t lo0 = __funnelshift_r(a, bq, sh0), hi0 = __funnelshift_r(bq, cq, sh0);
  const uint32_t lo1 = __funnelshift_r(d, e, sh1), hi1 = __funnelshift_r(e, fq, sh1);
  const uint32_t wa0 = __byte_perm(lo0, hi0, 0x4130u), wb0 = __byte_perm(lo0, hi0, 0x4452u);
  const uint32_t wa1 = __byte_perm(lo1, hi1, 0x4130u), wb1 = __byte_perm(lo1, hi1, 0x4452u);
  const unsigned s00 = __dp2a_lo(a01, wa0, 0u), s01 = __dp2a_hi(a01, wa0, 0u), s02 = __dp2a_lo(a01, wb0, 0u);
  const unsigned s10 = __dp2a_lo(a01, wa1, 0u), s11 = __dp2a_hi(a01, wa1, 0u), s12 = __dp2a_lo(a01, wb1, 0u);
  t[0] = __umulhi(b0s, s00 >> 4) + __umulhi(b1s, s10 >> 4);
  t[1] = __umulhi(b0s, s01 >> 4) + __umulhi(b1s, s11 >> 4);
  t[2] = __umulhi(b0s, s02 >> 4) + __umulhi(b1s, s12 >> 4);
}

// shared memory: [0,16) two mbarriers | s_val[1024] (x / 255 of (t + 2) >> 2 for the unrounded vertical sum t) |
// s_row[2][kLbMaxRows] (off0, off1, b0 << 16, b1 << 16; off0 < 0: border row) | xt[new_w] | row slots [2][2 * rows][pitch]
#ifndef LVC_LB_MINBLOCKS
#define LVC_LB_MINBLOCKS 3   // 80 registers, three blocks per SM: 0.088 ms for 64 x 1080p; capped at 64 registers (four blocks) it spills: 0.100 ms
#endif
__global__ void __launch_bounds__(kLbThreads, LVC_LB_MINBLOCKS) lvc_letterbox_kernel(const __grid_constant__ LbParams p) {
  extern __shared__ __align__(128) uint8_t lb_smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(lb_smem);
  float* s_val = reinterpret_cast<float*>(lb_smem + 16);
  int4* s_row = reinterpret_cast<int4*>(lb_smem + 16 + 4 * 1024);
  XTap* xt = reinterpret_cast<XTap*>(lb_smem + kLbFixed);
  const lvc_letterbox_geom& g = p.g;
  uint8_t* slots = lb_smem + lb_slots_offset(g.new_w);
  const int R = p.rows, nslot = 2 * R;
  const bool area = g.src_w == 2 * g.new_w && g.src_h == 2 * g.new_h;   // cv::resize switches to INTER_AREA
  const double sx_scale = cv_scale(g.src_w, g.new_w), sy_scale = cv_scale(g.src_h, g.new_h);
  for (int x = threadIdx.x; x < g.new_w; x += kLbThreads) {
    XTap t;
    if (area) { t.off = 6 * x; t.a01 = 32 | (32 << 16); }
    else {
      int sx; float fr;
      cv_tap(x, sx_scale, sx, fr);
      if (sx < 0) { sx = 0; fr = 0.f; }
      if (sx >= g.src_w - 1) { sx = g.src_w - 1; fr = 0.f; }   // clamped edge: a1 == 0, the slack bytes read at +3 weigh nothing
      t.off = 3 * sx;
      t.a01 = coef(__fsub_rn(1.0f, fr)) | (coef(fr) << 16);
    }
    xt[x] = t;
  }
  // x / 255 (IEEE division = what torch's CPU `im /= 255` computes) of the rounded value, indexed by the unrounded sum
  for (int i = threadIdx.x; i < 1024; i += kLbThreads) s_val[i] = __fdiv_rn((float)min((i + 2) >> 2, 255), 255.0f);
  if (threadIdx.x == 0) {
    lb_mbar_init(&full[0], 1);
    lb_mbar_init(&full[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const uint64_t row_bytes = (uint64_t)g.src_w * 3u;
  const uint64_t buf_end = reinterpret_cast<uint64_t>(p.frames) + (uint64_t)p.n * g.src_h * row_bytes;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  // warp 0: lane j < 2R owns tap (j & 1) of output row j >> 1 of the item; duplicates inside the item are fetched once
  auto issue = [&](int item, int b) {
    const int frame = item / p.row_blocks, rblk = item - frame * p.row_blocks;
    const unsigned lanes = nslot >= 32 ? 0xffffffffu : ((1u << nslot) - 1u);
    uint32_t bulk_bytes = 0;
    if (lane < nslot) {
      const int r = lane >> 1, iy = rblk * R + r - g.top;
      const bool row_in = iy >= 0 && iy < g.new_h && rblk * R + r < g.out_h;
      int ysrc = -1 - lane, cb = 0;   // border rows: a distinct key per lane, nothing fetched
      if (row_in) {
        if (area) { ysrc = 2 * iy + (lane & 1); cb = (int)0x80000000u; }
        else {
          int sy; float fr;
          cv_tap(iy, sy_scale, sy, fr);
          // cv::resize does not reset the fraction of clamped ROWS
          ysrc = min(max(sy + (lane & 1), 0), g.src_h - 1);
          cb = ((lane & 1) ? coef(fr) : coef(__fsub_rn(1.0f, fr))) << 16;   // pre-shifted: the blend is a multiply-high
        }
      }
      const unsigned same = __match_any_sync(lanes, ysrc);
      const int lead = __ffs(same) - 1;
      const uint64_t addr = reinterpret_cast<uint64_t>(p.frames) + ((uint64_t)frame * g.src_h + (uint64_t)max(ysrc, 0)) * row_bytes;
      const uint64_t a = addr & ~(uint64_t)15;
      const uint32_t shift = (uint32_t)(addr - a);
      const uint32_t my_bytes = (shift + (uint32_t)row_bytes + 15u) & ~15u;   // <= pitch - 16 by construction
      const int my_off = row_in ? (b * nslot + lead) * p.pitch + (int)shift : -1;
      uint8_t* dst = slots + (size_t)(b * nslot + lane) * p.pitch;
      if (row_in && lead == lane) {
        bulk_bytes = my_bytes;
        if (a + my_bytes > buf_end) {   // last bytes of the buffer: not a whole 16-byte granule, copied by hand
          bulk_bytes = (uint32_t)((buf_end - a) & ~(uint64_t)15);
          for (uint64_t q = a + bulk_bytes; q < buf_end; ++q) dst[q - a] = *reinterpret_cast<const uint8_t*>(q);
        }
      }
      const int off_next = __shfl_down_sync(lanes, my_off, 1);
      const int cb_next = __shfl_down_sync(lanes, cb, 1);
      if (!(lane & 1)) s_row[b * kLbMaxRows + r] = make_int4(my_off, off_next, cb, cb_next);
      uint32_t total = bulk_bytes;
      // xor-reduction over the 2R active lanes (R is 1, 2, 4 or 8: 2R is a power of two, partners stay inside `lanes`)
      for (int o = nslot >> 1; o; o >>= 1) total += __shfl_xor_sync(lanes, total, o);
      __syncwarp(lanes);   // the row table and any hand-copied tail bytes are written before the barrier is armed
      if (lane == 0) lb_mbar_expect(&full[b], total);
      __syncwarp(lanes);
      if (bulk_bytes) lb_bulk_g2s(dst, reinterpret_cast<const void*>(a), bulk_bytes, &full[b]);
    }
  };

  const size_t plane = (size_t)g.out_h * g.out_w;
  const float pad = s_val[114 * 4];
  const uint32_t slots_s = lb_s32(slots);
  // The usual network input is at most 768 columns wide: a thread then owns the SAME (up to three) columns in every row
  // of every item, and their taps live in registers for the whole kernel (no table look-up, no bounds logic per pixel).
  const bool narrow = g.out_w <= kLbThreads * kLbUnroll;
  uint32_t c_off[kLbUnroll], c_a01[kLbUnroll];
  bool c_in[kLbUnroll], c_val[kLbUnroll];
#pragma unroll
  for (int u = 0; u < kLbUnroll; ++u) {
    const int ox = (int)threadIdx.x + u * kLbThreads, ix = ox - g.left;
    c_val[u] = ox < g.out_w;
    c_in[u] = c_val[u] && ix >= 0 && ix < g.new_w;
    const XTap t = c_in[u] ? xt[ix] : XTap{0, 0};
    c_off[u] = (uint32_t)t.off; c_a01[u] = (uint32_t)t.a01;
  }
  // store of one pixel: transpose(2,0,1)[::-1] -- output channel k is input channel 2 - k
  auto store3 = [&](auto BF, size_t e, const float (&v)[3]) {
    if (decltype(BF)::value) {
      __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + e;
      o[0] = __float2bfloat16_rn(v[2]); o[plane] = __float2bfloat16_rn(v[1]); o[2 * plane] = __float2bfloat16_rn(v[0]);
    } else {
      float* o = reinterpret_cast<float*>(p.out) + e;
      __stcs(o, v[2]); __stcs(o + plane, v[1]); __stcs(o + 2 * plane, v[0]);
    }
  };
  auto rows = [&](auto BF, int frame, int rblk, int b) {
    for (int r = 0; r < R; ++r) {
      const int oy = rblk * R + r;
      if (oy >= g.out_h) break;
      const int4 rt = s_row[b * kLbMaxRows + r];
      const bool row_in = rt.x >= 0;
      const uint32_t base0 = slots_s + (uint32_t)max(rt.x, 0), base1 = slots_s + (uint32_t)max(rt.y, 0);
      const uint32_t b0s = (uint32_t)rt.z, b1s = (uint32_t)rt.w;
      const size_t erow = (size_t)frame * 3 * plane + (size_t)oy * g.out_w;
      if (narrow) {
        unsigned t[kLbUnroll][3];
#pragma unroll
        for (int u = 0; u < kLbUnroll; ++u)   // all shared loads of the thread's pixels are requested before the first is used
          if (row_in && c_in[u]) lb_pixel(base0 + c_off[u], base1 + c_off[u], c_a01[u], b0s, b1s, t[u]);
#pragma unroll
        for (int u = 0; u < kLbUnroll; ++u) {
          if (!c_val[u]) continue;
          float v[3] = {pad, pad, pad};
          if (row_in && c_in[u]) { v[0] = s_val[t[u][0]]; v[1] = s_val[t[u][1]]; v[2] = s_val[t[u][2]]; }
          store3(BF, erow + threadIdx.x + u * kLbThreads, v);
        }
        continue;
      }
      for (int ox0 = threadIdx.x; ox0 < g.out_w; ox0 += kLbThreads * kLbUnroll) {
        unsigned t[kLbUnroll][3];
        bool in[kLbUnroll];
#pragma unroll
        for (int u = 0; u < kLbUnroll; ++u) {
          const int ox = ox0 + u * kLbThreads, ix = ox - g.left;
          in[u] = row_in && ox < g.out_w && ix >= 0 && ix < g.new_w;
          if (!in[u]) continue;
          const XTap tp = xt[ix];
          lb_pixel(base0 + (uint32_t)tp.off, base1 + (uint32_t)tp.off, (uint32_t)tp.a01, b0s, b1s, t[u]);
        }
#pragma unroll
        for (int u = 0; u < kLbUnroll; ++u) {
          const int ox = ox0 + u * kLbThreads;
          if (ox >= g.out_w) continue;
          float v[3] = {pad, pad, pad};
          if (in[u]) { v[0] = s_val[t[u][0]]; v[1] = s_val[t[u][1]]; v[2] = s_val[t[u][2]]; }
          store3(BF, erow + ox, v);
        }
      }
    }
  };
  if (warp == 0 && (int)blockIdx.x < p.items) issue(blockIdx.x, 0);
  uint32_t it = 0;
  for (int item = blockIdx.x; item < p.items; item += gridDim.x, ++it) {
    const int b = it & 1;
    // the other buffer was read by the previous trip, which every thread has left (the barrier at the end of the trip)
    if (warp == 0 && item + (int)gridDim.x < p.items) issue(item + gridDim.x, b ^ 1);
    lb_mbar_wait(&full[b], (it >> 1) & 1u);
    const int frame = item / p.row_blocks, rblk = item - frame * p.row_blocks;
    if (p.out_bf16) rows(std::true_type{}, frame, rblk, b);
    else rows(std::false_type{}, frame, rblk, b);
    __syncthreads();   // buffer b (rows and s_row) may be refilled by the next trip's issue
  }
}

// yolov5 scale_boxes: gain and pad in double (python floats), applied to float32 box coordinates; clip; round
// half to even; int().
struct BoxScale { float gain, pad_x, pad_y; int w, h; };
BoxScale box_scale(const lvc_letterbox_geom& g) {
  const double gain = std::fmin((double)g.out_h / (double)g.src_h, (double)g.out_w / (double)g.src_w);
  BoxScale b;
  b.gain = (float)gain;
  b.pad_x = (float)(((double)g.out_w - (double)g.src_w * gain) / 2.0);
  b.pad_y = (float)(((double)g.out_h - (double)g.src_h * gain) / 2.0);
  b.w = g.src_w; b.h = g.src_h;
  return b;
}
__device__ __forceinline__ int4 scale_one(const float* d, const BoxScale& s) {
  const float x1 = __fdiv_rn(__fsub_rn(d[0], s.pad_x), s.gain), y1 = __fdiv_rn(__fsub_rn(d[1], s.pad_y), s.gain);
  const float x2 = __fdiv_rn(__fsub_rn(d[2], s.pad_x), s.gain), y2 = __fdiv_rn(__fsub_rn(d[3], s.pad_y), s.gain);
  return make_int4((int)rintf(fminf(fmaxf(x1, 0.f), (float)s.w)), (int)rintf(fminf(fmaxf(y1, 0.f), (float)s.h)),
                   (int)rintf(fminf(fmaxf(x2, 0.f), (float)s.w)), (int)rintf(fminf(fmaxf(y2, 0.f), (float)s.h)));
}

__global__ void lvc_scale_boxes_kernel(const float* __restrict__ det, int n, int row_stride, BoxScale s, int4* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = scale_one(det + (size_t)i * row_stride, s);
}

// One block per frame: arg-max of the confidence over the frame's detections (ties: lowest row index).
__global__ void __launch_bounds__(128) lvc_best_boxes_kernel(const float* __restrict__ det, const int32_t* __restrict__ frame_of,
                                                            int n_det, float conf_thres, BoxScale s, int4* __restrict__ out) {
  __shared__ float s_conf[128];
  __shared__ int s_idx[128];
  const int frame = blockIdx.x;
  float best = -1.f;
  int bi = 0x7fffffff;
  for (int i = threadIdx.x; i < n_det; i += blockDim.x) {
    if (frame_of[i] != frame) continue;
    const float c = det[(size_t)i * 6 + 4];
    if (c >= conf_thres && (c > best || (c == best && i < bi))) { best = c; bi = i; }
  }
  s_conf[threadIdx.x] = best; s_idx[threadIdx.x] = bi;
  __syncthreads();
  for (int o = 64; o; o >>= 1) {
    if (threadIdx.x < o) {
      const float c = s_conf[threadIdx.x + o];
      const int j = s_idx[threadIdx.x + o];
      if (c > s_conf[threadIdx.x] || (c == s_conf[threadIdx.x] && j < s_idx[threadIdx.x])) { s_conf[threadIdx.x] = c; s_idx[threadIdx.x] = j; }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0)
    out[frame] = s_conf[0] >= 0.f ? scale_one(det + (size_t)s_idx[0] * 6, s) : make_int4(-1, -1, -1, -1);
}

int check_geom(const lvc_letterbox_geom* g, const char* who) {
  if (!g) return lvc_fail(LVC_E_INVAL, "%s: NULL geometry", who);
  if (g->src_h <= 0 || g->src_w <= 0 || g->new_h <= 0 || g->new_w <= 0 || g->top < 0 || g->left < 0 ||
      g->out_h < g->top + g->new_h || g->out_w < g->left + g->new_w)
    return lvc_fail(LVC_E_INVAL, "%s: inconsistent letterbox geometry (use lvc_letterbox_geometry)", who);
  return LVC_OK;
}

}  // namespace

extern "C" int lvc_letterbox_geometry(int32_t src_h, int32_t src_w, int32_t shape_h, int32_t shape_w, int32_t stride,
                                      int32_t auto_pad, int32_t scaleup, lvc_letterbox_geom* out) {
  if (!out) return lvc_fail(LVC_E_INVAL, "letterbox_geometry: NULL output");
  if (src_h <= 0 || src_w <= 0 || shape_h <= 0 || shape_w <= 0 || stride <= 0)
    return lvc_fail(LVC_E_INVAL, "letterbox_geometry: extents and stride must be positive");
  // python: r = min(new_shape[0] / shape[0], new_shape[1] / shape[1]); new_unpad = int(round(shape[1] * r)), ...
  double r = std::fmin((double)shape_h / (double)src_h, (double)shape_w / (double)src_w);
  if (!scaleup) r = std::fmin(r, 1.0);
  const int new_w = (int)std::nearbyint((double)src_w * r), new_h = (int)std::nearbyint((double)src_h * r);
  if (new_w <= 0 || new_h <= 0) return lvc_fail(LVC_E_SHAPE, "letterbox_geometry: frame collapses to zero size");
  int dw = shape_w - new_w, dh = shape_h - new_h;
  if (auto_pad) { dw %= stride; dh %= stride; }   // np.mod: minimum rectangle
  const double hw = dw / 2.0, hh = dh / 2.0;
  const int top = (int)std::nearbyint(hh - 0.1), bottom = (int)std::nearbyint(hh + 0.1);
  const int left = (int)std::nearbyint(hw - 0.1), right = (int)std::nearbyint(hw + 0.1);
  out->src_h = src_h; out->src_w = src_w; out->new_h = new_h; out->new_w = new_w;
  out->top = top; out->left = left; out->out_h = new_h + top + bottom; out->out_w = new_w + left + right;
  return LVC_OK;
}

extern "C" int lvc_letterbox_forward(const lvc_letterbox_geom* g, const uint8_t* frames, int32_t n, int32_t out_dtype,
                                     void* out, void* stream) {
  int rc = check_geom(g, "letterbox_forward");
  if (rc != LVC_OK) return rc;
  if (!frames || !out) return lvc_fail(LVC_E_INVAL, "letterbox_forward: NULL pointer");
  if (n <= 0 || n > 65535) return lvc_fail(LVC_E_INVAL, "letterbox_forward: 1..65535 frames per call");
  if (out_dtype != LVC_DT_F32 && out_dtype != LVC_DT_BF16) return lvc_fail(LVC_E_INVAL, "letterbox_forward: out_dtype must be f32 or bf16");
  if (g->new_w > kLbMaxW) return lvc_fail(LVC_E_SHAPE, "letterbox_forward: resize target wider than %d", kLbMaxW);
  LbParams p;
  p.g = *g; p.frames = frames; p.out = out; p.n = n; p.out_bf16 = out_dtype == LVC_DT_BF16;
  int dev = 0, sms = 0, per_sm = 0, max_smem = 0;
  LVC_CUDA_TRY(cudaGetDevice(&dev));
  LVC_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  LVC_CUDA_TRY(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  // staged kernel: the most rows per item (4, 2, 1) whose double-buffered row slots leave room for four blocks per SM
  // (measured on B200, 64 frames: 1080p 2 rows 0.094 ms vs 4 rows 0.112, 8 rows 0.167; 480p 4 rows 0.108 vs 2 rows 0.116),
  // else for two, else for one; frames too wide even for that take the direct-global kernel
  p.pitch = ((g->src_w * 3 + 15 + 15) & ~15) + 16;
  const size_t fixed = lb_slots_offset(g->new_w);
  p.rows = 0;
  const int forced_rows = lvc_knob_int("LVC_LB_STAGE_ROWS", -1);   // profiling knob (gated): 0 = direct kernel
  for (int pass = 0; pass < 3 && !p.rows && forced_rows != 0; ++pass)
    for (int r = forced_rows > 0 ? kLbMaxRows : 4; r >= 1; r >>= 1) {
      if (forced_rows > 0 && r != forced_rows) continue;
      const size_t need = fixed + (size_t)2 * 2 * r * p.pitch;
      const size_t budget = pass == 0 ? (size_t)(max_smem / 4 - 1024) : (pass == 1 ? (size_t)(max_smem / 2 - 1024) : (size_t)max_smem);
      if (need <= budget && (pass == 2 || r >= 2 || forced_rows > 0)) { p.rows = r; break; }
    }
  if (p.rows) {
    const size_t smem = fixed + (size_t)2 * 2 * p.rows * p.pitch;
    p.row_blocks = (g->out_h + p.rows - 1) / p.rows;
    p.items = n * p.row_blocks;
    LVC_CUDA_TRY(cudaFuncSetAttribute(lvc_letterbox_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    LVC_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lvc_letterbox_kernel, kLbThreads, smem));
    const int grid = std::min(p.items, sms * std::max(per_sm, 1));
    lvc_letterbox_kernel<<<grid, kLbThreads, smem, static_cast<cudaStream_t>(stream)>>>(p);
    LVC_CUDA_TRY(cudaGetLastError());
    return LVC_OK;
  }
  p.row_blocks = (g->out_h + kLbRows - 1) / kLbRows;
  p.items = n * p.row_blocks;
  LVC_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lvc_letterbox_direct_kernel, kLbThreads, 0));
  const int grid = std::min(p.items, sms * std::max(per_sm, 1));
  lvc_letterbox_direct_kernel<<<grid, kLbThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
  LVC_CUDA_TRY(cudaGetLastError());
  return LVC_OK;
}

extern "C" int lvc_scale_boxes(const lvc_letterbox_geom* g, const float* det, int32_t n, int32_t row_stride,
                               int32_t* boxes_out, void* stream) {
  int rc = check_geom(g, "scale_boxes");
  if (rc != LVC_OK) return rc;
  if (n == 0) return LVC_OK;
  if (!det || !boxes_out || n < 0 || row_stride < 4) return lvc_fail(LVC_E_INVAL, "scale_boxes: bad argument");
  if (reinterpret_cast<uintptr_t>(boxes_out) & 15) return lvc_fail(LVC_E_ALIGN, "scale_boxes: boxes_out must be 16-byte aligned");
  lvc_scale_boxes_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(det, n, row_stride, box_scale(*g),
                                                                                      reinterpret_cast<int4*>(boxes_out));
  LVC_CUDA_TRY(cudaGetLastError());
  return LVC_OK;
}

extern "C" int lvc_best_boxes(const lvc_letterbox_geom* g, const float* det, const int32_t* frame_of, int32_t n_det,
                              int32_t n_frames, float conf_thres, int32_t* boxes_out, void* stream) {
  int rc = check_geom(g, "best_boxes");
  if (rc != LVC_OK) return rc;
  if (n_frames <= 0 || n_det < 0 || !boxes_out || (n_det > 0 && (!det || !frame_of))) return lvc_fail(LVC_E_INVAL, "best_boxes: bad argument");
  if (reinterpret_cast<uintptr_t>(boxes_out) & 15) return lvc_fail(LVC_E_ALIGN, "best_boxes: boxes_out must be 16-byte aligned");
  lvc_best_boxes_kernel<<<n_frames, 128, 0, static_cast<cudaStream_t>(stream)>>>(det, frame_of, n_det, conf_thres, box_scale(*g),
                                                                                reinterpret_cast<int4*>(boxes_out));
  LVC_CUDA_TRY(cudaGetLastError());
  return LVC_OK;
}
