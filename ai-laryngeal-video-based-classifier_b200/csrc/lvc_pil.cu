// HF image processors on RAW frames (any size), fused: Pillow BILINEAR resize of the shortest edge to E
// (antialiased: the triangle filter widens with the down-scale), centre crop 224, rescale + normalise, layout.
//   VivitImageProcessor      E = 256   data_preprocessing.py:16, vivit_transformer/inference.py:187
//   VideoMAE-class processor E = 224   data_preprocessing_timesformer.py:16
// (transformers image_transforms.resize -> PIL.Image.resize(BILINEAR); Pillow's ImagingResample: horizontal
// pass -> uint8 -> vertical pass -> uint8, 22-bit fixed-point coefficients; restated and pinned in
// oracle/resize.py).  The 224x224 inputs of the trainer path keep the faster lvc_clip_kernel presets; this kernel
// closes the gap for callers that hand the processor decoded frames directly.
// Work item = (frame, band of R cropped output rows): the band's source rows go through the horizontal pass
// into a shared-memory intermediate (one word per pixel), then the vertical pass + LUT + store.  Integer / byte
// work, no tensor cores.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>

#include "lvc_internal.h"

namespace {

constexpr int S = 224;        // crop / output size
constexpr int KT = 24;        // most taps per output sample: ceil(scale)*2+1 <= 24, scale <= 11.5
constexpr int kMaxThreads = 512;
constexpr int kCH = 32;       // source rows per staged block: one row per lane in the horizontal pass
constexpr int IP = S + 1;     // words per row of the intermediate ring (odd: lane = row writes hit 32 banks)
#ifndef LVC_PIL_WORD_LOADS_FROM
#define LVC_PIL_WORD_LOADS_FROM 6
#endif
constexpr int kPilWordLoadsFrom = LVC_PIL_WORD_LOADS_FROM;   // tap count from which the horizontal pass reads words instead of bytes

struct PilParams {
  const lvc_video* videos;
  const int32_t* indices;
  const float* lut;
  void* out;
  int T, layout, out_dtype, bgr, edge;
  int bands, R, total_items;    // item = (frame, band of R output rows)
  int nseg;                     // column segments a row block is staged in (S % nseg == 0)
  int kt;                       // taps the tables hold per sample: 2*ceil(max scale) + 1
  int pitch_words;              // words per staged source-row segment (odd)
  int ir;                       // rows in the intermediate ring: kCH + kt + 1
  int kt_tok, kh, kw;           // LVC_LAYOUT_TOKENS: patch size (output = GEMM operand A[M][K], bf16)
  int bulk;                     // rows staged with 16-byte-granular bulk copies (frames in pinned HOST memory, read over PCIe)
  int max_w, max_h;             // the frame size the tables and buffers were sized for (desc.max_width, max_height)
  unsigned* err;                // device error word (lvc_device_status)
};

struct Taps { int xmin, n; };

// Pillow precompute_coeffs + normalize_coeffs_8bpc for output sample `xx` of an in_size -> out_size resize with
// the triangle (BILINEAR) filter; all arithmetic in double, no contraction, C truncation.  kk[i * stride].
__device__ Taps pil_taps(int xx, int in_size, int out_size, int* kk, int stride, int kt) {
  const double scale = __ddiv_rn((double)in_size, (double)out_size);
  const double filterscale = scale < 1.0 ? 1.0 : scale;
  const double support = filterscale;   // bilinear support 1.0 * filterscale
  const double center = __dmul_rn((double)xx + 0.5, scale);
  const double ss = __ddiv_rn(1.0, filterscale);
  int xmin = (int)__dadd_rn(__dadd_rn(center, -support), 0.5);
  if (xmin < 0) xmin = 0;
  int xmax = (int)__dadd_rn(__dadd_rn(center, support), 0.5);
  if (xmax > in_size) xmax = in_size;
  // n <= kt always: the kernel clamps frames larger than the descriptor's max_width / max_height before it gets here.
  // (Stated for the compiler: its register allocation of the unrolled passes depends on knowing these bounds.)
  const int n = xmax - xmin;
  __builtin_assume(n <= kt);
  double w[KT];
  double ww = 0.0;
  for (int x = 0; x < n; ++x) {
    double t = __dmul_rn(__dadd_rn(__dadd_rn((double)(x + xmin), -center), 0.5), ss);
    t = t < 0.0 ? -t : t;
    w[x] = t < 1.0 ? __dadd_rn(1.0, -t) : 0.0;
    ww = __dadd_rn(ww, w[x]);
  }
  for (int x = 0; x < n; ++x) {
    const double k = ww != 0.0 ? __ddiv_rn(w[x], ww) : w[x];
    kk[x * stride] = (int)(k < 0.0 ? __dadd_rn(-0.5, __dmul_rn(k, 4194304.0)) : __dadd_rn(0.5, __dmul_rn(k, 4194304.0)));
  }
  for (int x = n; x < kt; ++x) kk[x * stride] = 0;   // the passes run a fixed tap count: padding taps weigh nothing
  Taps t; t.xmin = xmin; t.n = n;
  return t;
}

// HF get_resize_output_image_size(shortest_edge = E, default_to_square = False): python int(E * long / short)
__device__ __forceinline__ void hf_geom(int H, int W, int E, int& nh, int& nw) {
  if (E == 0) { nh = S; nw = S; }   // size = {"height": 224, "width": 224}: both axes to 224, aspect not kept
  else if (H <= W) { nh = E; nw = (int)__ddiv_rn((double)(E * (long long)W), (double)H); }
  else { nw = E; nh = (int)__ddiv_rn((double)(E * (long long)H), (double)W); }
}

__device__ __forceinline__ void cp_async4(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4s(uint32_t dst_shared, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst_shared), "l"(src) : "memory");
}
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(s32(bar)), "r"(parity), "r"(0x989680u) : "memory");   // suspend-time hint: poll less
  } while (!done);
}
// 1-D bulk async copy global (or pinned host) -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(s32(dst)), "l"(src), "r"(bytes), "r"(s32(bar)) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait0() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Two CTAs per SM (64 registers) for the tap counts of down-scales up to ~4.5x (1080p -> 256).  ptxas lands there on its
// own only by a hair (unrelated edits moved NT = 8, 9 to 95 registers and cost 15 % at 1080p), so it is stated.
template <int NT>   // taps per sample held in the tables (>= 2*ceil(scale) + 1), fully unrolled in both passes
__global__ void __launch_bounds__(kMaxThreads, (NT <= 9 ? 2 : 1)) lvc_pil_kernel(const __grid_constant__ PilParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  constexpr int kt = NT;
  int* hk = reinterpret_cast<int*>(smem);                 // [kt][S]  horizontal weights, tap-major
  int* vk = hk + kt * S;                                  // [kt][S]  vertical weights of the 224 cropped rows
  int* hmin = vk + kt * S;                                // [S] first source column / tap count per output column
  int* hn = hmin + S;
  int* vmin = hn + S;                                     // [S] first source row / tap count per output row
  int* vn = vmin + S;
  int* rsh = vn + S;                                      // [2][kCH] byte shift of each staged row
  uint64_t* full = reinterpret_cast<uint64_t*>(rsh + 2 * kCH);   // [2] bulk mode: bytes of a slot have landed
  float* lut = reinterpret_cast<float*>(full + 2);          // [3*256]
  uint32_t* inter = reinterpret_cast<uint32_t*>(lut + 768);            // [ir][IP] ring of horizontally resized rows
  // [2][kCH][pitch_words], 16-byte aligned (bulk copies land here)
  const size_t rows_off = ((size_t)(reinterpret_cast<uint8_t*>(inter + (size_t)p.ir * IP) - smem) + 15) & ~(size_t)15;
  uint32_t* rows = reinterpret_cast<uint32_t*>(smem + rows_off);   // offset arithmetic keeps the shared address space
  // a kernel launched behind this one with programmatic stream serialization (the patch-embedding GEMM) may be
  // scheduled as SMs free up; it waits for this grid to complete before it touches global data
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kThreads = blockDim.x, kWarps = kThreads >> 5;
  if (p.lut)
    for (int i = tid; i < 768; i += kThreads) lut[i] = p.lut[i];
  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  uint32_t gq = 0;         // chunks staged so far by this CTA: slot = gq & 1, barrier phase = (gq >> 1) & 1
  int cw = -1, ch_ = -1;   // frame size the tables were built for
  const size_t plane = (size_t)S * S;
  const int xseg = S / p.nseg;

  for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
    const int f = item / p.bands, band = item - f * p.bands;
    const int b = f / p.T, t = f - b * p.T;
    const lvc_video v = p.videos[b];
    const int idx = min(max(p.indices[f], 0), v.n_frames - 1);
    // A video larger than the descriptor's max_width x max_height (which sized the tap tables and the staging buffers) is
    // a contract violation: it is processed AS IF it had the maximum size -- every access stays inside its buffer and
    // inside shared memory, its output frames are meaningless -- and the device error word is raised.  (Skipping the
    // item instead costs the kernel its two-CTAs-per-SM register budget: 63 -> 95 registers.)
    const int H = min(v.height, p.max_h), W = min(v.width, p.max_w);
    if ((v.width > p.max_w || v.height > p.max_h) && tid == 0) atomicOr(p.err, (unsigned)LVC_DEVERR_FRAME_TOO_LARGE);
    __syncthreads();   // previous item is done with the tables and buffers
    if (W != cw || H != ch_) {
      int nh, nw;
      hf_geom(H, W, p.edge, nh, nw);
      const int top = (nh - S) / 2, left = (nw - S) / 2;
      if (tid < S) {
        Taps tp = pil_taps(left + tid, W, nw, hk + tid, S, kt);
        hmin[tid] = tp.xmin; hn[tid] = tp.n;
        tp = pil_taps(top + tid, H, nh, vk + tid, S, kt);
        vmin[tid] = tp.xmin; vn[tid] = tp.n;
      }
      cw = W; ch_ = H;
      __syncthreads();
    }
    const int oy0 = band * p.R, oy1 = oy0 + p.R;               // output rows [oy0, oy1)
    const int r0 = vmin[oy0], r1 = vmin[oy1 - 1] + vn[oy1 - 1];   // source rows [r0, r1)
    const uint64_t buf_end = v.base + (uint64_t)v.nbytes;
    const uint64_t frame_base = v.base + (uint64_t)idx * (uint64_t)H * (uint64_t)W * 3u;
    const int nblk = (r1 - r0 + kCH - 1) / kCH;
    const int nchunk = nblk * p.nseg;

    // stage chunk (row block, column segment) into slot `sl`: 4-byte async copies from HBM (any alignment, odd word
    // pitch), or -- frames in pinned host memory -- one 16-byte-granular bulk copy per row (PCIe wants large requests;
    // the 4-way bank conflicts of the 16-byte pitch hide behind the bus)
    auto issue = [&](int sl, int rb, int seg) {
      const int xa = seg * xseg, xb = xa + xseg;
      const int c0 = hmin[xa], c1 = hmin[xb - 1] + hn[xb - 1];
      const int span = (c1 - c0) * 3;
      uint32_t* slot = rows + (size_t)sl * kCH * p.pitch_words;
      if (p.bulk) {
        if (warp == 0) {
          const int r = r0 + rb * kCH + lane;
          uint32_t bulk_bytes = 0;
          uint64_t a16 = 0;
          uint8_t* dst = reinterpret_cast<uint8_t*>(slot + (size_t)lane * p.pitch_words);
          if (r < r1) {
            const uint64_t a = frame_base + ((uint64_t)r * (uint64_t)W + (uint64_t)c0) * 3u;
            a16 = a & ~(uint64_t)15;
            const uint32_t sh = (uint32_t)(a - a16);
            const uint32_t bytes = (sh + (uint32_t)span + 15u) & ~15u;
            __builtin_assume(bytes <= (uint32_t)p.pitch_words * 4u);
            rsh[sl * kCH + lane] = (int)sh;
            bulk_bytes = bytes;
            if (a16 + bytes > buf_end) {   // last bytes of the buffer: not a whole 16-byte granule, copied by hand
              bulk_bytes = (uint32_t)((buf_end - a16) & ~(uint64_t)15);
              for (uint64_t qq = a16 + bulk_bytes; qq < buf_end; ++qq) dst[qq - a16] = *reinterpret_cast<const uint8_t*>(qq);
            }
          }
          uint32_t total = bulk_bytes;
#pragma unroll
          for (int o = 16; o; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
          if (lane == 0) mbar_arrive_expect_tx(&full[sl], total);
          __syncwarp();
          if (bulk_bytes) bulk_g2s(dst, reinterpret_cast<const void*>(a16), bulk_bytes, &full[sl]);
        }
        return;
      }
      for (int j = warp; j < kCH; j += kWarps) {
        const int r = r0 + rb * kCH + j;
        if (r >= r1) break;
        const uint64_t a = frame_base + ((uint64_t)r * (uint64_t)W + (uint64_t)c0) * 3u;
        const uint64_t a4 = a & ~(uint64_t)3;
        const int sh = (int)(a - a4);
        const int nwords = (sh + span + 3) >> 2;
        __builtin_assume(nwords <= p.pitch_words);
        if (lane == 0) rsh[sl * kCH + j] = sh;
        uint32_t* dst = slot + (size_t)j * p.pitch_words;
        if (a4 + 4u * (uint64_t)nwords <= buf_end) {   // every row but the buffer's last: no per-word bound check
          const uint8_t* g = reinterpret_cast<const uint8_t*>(a4) + 4 * lane;
          uint32_t sa = (uint32_t)__cvta_generic_to_shared(dst) + 4u * lane;
          int left = nwords - lane;
          for (; left > 96; left -= 128, g += 512, sa += 512) {
            cp_async4s(sa, g); cp_async4s(sa + 128, g + 128); cp_async4s(sa + 256, g + 256); cp_async4s(sa + 384, g + 384);
          }
          for (; left > 0; left -= 32, g += 128, sa += 128) cp_async4s(sa, g);
        } else {
          for (int i = lane; i < nwords; i += 32) {
            const uint64_t wa = a4 + 4u * (uint64_t)i;
            if (wa + 4 <= buf_end) cp_async4(dst + i, reinterpret_cast<const void*>(wa));
            else {   // last bytes of the buffer
              uint32_t w = 0;
              for (int qq = 0; qq < 4; ++qq)
                if (wa + qq < buf_end) w |= (uint32_t)__ldg(reinterpret_cast<const uint8_t*>(wa + qq)) << (8 * qq);
              dst[i] = w;
            }
          }
        }
      }
    };

    issue((int)(gq & 1u), 0, 0);
    cp_async_commit();
    int next_emit = oy0;
    int rb = 0, seg = 0;               // chunk q = (row block rb, column segment seg), tracked without divisions
    int ring0 = r0 % p.ir;             // ring slot of the row block's first row
    for (int q = 0; q < nchunk; ++q) {
      const int sl = (int)(gq & 1u);
      if (p.bulk) mbar_wait(&full[sl], (gq >> 1) & 1u);
      else cp_async_wait0();
      ++gq;
      __syncthreads();   // chunk q landed (every thread's copies), rsh[] is visible; everyone is past chunk q-1:
                         // its slot may be refilled and the ring rows its vertical pass read may be overwritten
      if (q + 1 < nchunk) {               // overlaps the passes below
        const bool wrap = seg + 1 == p.nseg;
        issue(sl ^ 1, wrap ? rb + 1 : rb, wrap ? 0 : seg + 1);
      }
      cp_async_commit();
      const int xa = seg * xseg;
      const int c0 = hmin[xa];
      const int r = r0 + rb * kCH + lane;   // this lane's source row
      // ---- horizontal pass: lane = source row, warp = output column (weights are warp-uniform broadcasts)
      if (r < r1) {
        const uint8_t* rowp = reinterpret_cast<const uint8_t*>(rows + ((size_t)sl * kCH + lane) * p.pitch_words) +
                              rsh[sl * kCH + lane];
        int rslot = ring0 + lane;
        if (rslot >= p.ir) rslot -= p.ir;   // kCH <= ir: one wrap at most
        uint32_t* irow = inter + (size_t)rslot * IP;
        // Filters of six taps and more (down-scales from 2.5x: 720p, 1080p): the window's 3 * NT bytes are read as aligned 32-bit
        // words + funnel shift and split with PRMT -- the kernel is bound by the LSU / MIO queue (ncu at 1080p: mio_throttle 1.7
        // stall cycles per issue, LSU pipe 62 % with one LDS.U8 per sample) while the ALU pipe idles at 28 %.  Measured, 8 x 32
        // frames: 1080p -> 256 (NT = 9) 0.606 -> 0.564 ms, 720p -> 256 (NT = 6) 0.352 -> 0.341, 720p -> 224 (NT = 7) 0.418 -> 0.400;
        // narrower filters measure the same either way and keep the byte loads.
        const uint32_t rowp_s = s32(rowp);
        if constexpr (NT >= kPilWordLoadsFrom)
        for (int x = xa + warp; x < xa + xseg; x += kWarps) {
          const uint32_t boff = rowp_s + (uint32_t)((hmin[x] - c0) * 3);
          const uint32_t wa = boff & ~3u, sh = boff << 3;   // the funnel shift wraps at 32: (boff & 3) * 8
          const int* k = hk + x;
          int a[3] = {1 << 21, 1 << 21, 1 << 21};
          uint32_t cur, nxt;
          asm volatile("ld.shared.u32 %0, [%1];" : "=r"(cur) : "r"(wa));
#pragma unroll
          for (int j = 0; j < (3 * NT + 3) / 4; ++j) {
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(nxt) : "r"(wa + 4u * (uint32_t)(j + 1)));
            const uint32_t sw = __funnelshift_r(cur, nxt, sh);   // bytes 4j .. 4j + 3 of the window
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const int b = 4 * j + q;   // byte b of the window = channel b % 3 of tap b / 3
              if (b < 3 * NT) a[b % 3] += (int)__byte_perm(sw, 0u, 0x4440u + q) * k[(b / 3) * S];
            }
            cur = nxt;
          }
          // non-negative weights summing to 2^22 (+- n/2): the result is already inside [0, 255]
          irow[x] = (uint32_t)(a[0] >> 22) | ((uint32_t)(a[1] >> 22) << 8) | ((uint32_t)(a[2] >> 22) << 16);
        }
        else
        for (int x = xa + warp; x < xa + xseg; x += kWarps) {
          const uint8_t* src = rowp + (hmin[x] - c0) * 3;
          const int* k = hk + x;
          int a0 = 1 << 21, a1 = 1 << 21, a2 = 1 << 21;
#pragma unroll
          for (int i = 0; i < NT; ++i) {
            const int kv = k[i * S];
            a0 += src[3 * i] * kv; a1 += src[3 * i + 1] * kv; a2 += src[3 * i + 2] * kv;
          }
          // non-negative weights summing to 2^22 (+- n/2): the result is already inside [0, 255]
          irow[x] = (uint32_t)(a0 >> 22) | ((uint32_t)(a1 >> 22) << 8) | ((uint32_t)(a2 >> 22) << 16);
        }
      }
      if (seg == p.nseg - 1) {
        __syncthreads();   // the block's rows are in the ring
        // ---- vertical pass for the output rows whose window is complete: warp = output row, lane = column
        const int done = min(r0 + (rb + 1) * kCH, r1);
        int ne = next_emit;
        while (ne < oy1 && vmin[ne] + vn[ne] <= done) ++ne;
        for (int oy = next_emit + warp; oy < ne; oy += kWarps) {
          const int* k = vk + oy;
          const int s0 = vmin[oy] % p.ir;
          for (int x = lane; x < S; x += 32) {
            int a0 = 1 << 21, a1 = 1 << 21, a2 = 1 << 21;
            int sl = s0;
#pragma unroll
            for (int i = 0; i < NT; ++i) {
              const uint32_t w = inter[(size_t)sl * IP + x];
              const int kv = k[i * S];
              a0 += (int)(w & 0xffu) * kv; a1 += (int)((w >> 8) & 0xffu) * kv; a2 += (int)((w >> 16) & 0xffu) * kv;
              sl = sl + 1 == p.ir ? 0 : sl + 1;
            }
            int u[3] = {a0 >> 22, a1 >> 22, a2 >> 22};
            if (p.bgr) { const int tmp = u[0]; u[0] = u[2]; u[2] = tmp; }
            size_t e, cstride;
            if (p.layout == LVC_LAYOUT_BTCHW) { cstride = plane; e = (size_t)f * 3 * plane + (size_t)oy * S + x; }
            else if (p.layout == LVC_LAYOUT_BCTHW) { cstride = (size_t)p.T * plane; e = ((size_t)b * 3 * p.T + t) * plane + (size_t)oy * S + x; }
            else if (p.layout == LVC_LAYOUT_BTHWC) { cstride = 1; e = ((size_t)f * plane + (size_t)oy * S + x) * 3; }
            else {   // tokens: m = ((b*To + t/kt)*Ho + y/kh)*Wo + x/kw, k = ((c*kt + t%kt)*kh + y%kh)*kw + x%kw
              const unsigned K = 3u * p.kt_tok * p.kh * p.kw, Wo = S / p.kw, Ho = S / p.kh, To = p.T / p.kt_tok;
              const unsigned to = (unsigned)t / p.kt_tok, dt = (unsigned)t - to * p.kt_tok;
              cstride = (size_t)p.kt_tok * p.kh * p.kw;
              const size_t m = (((size_t)b * To + to) * Ho + oy / p.kh) * Wo + x / p.kw;
              e = m * K + ((size_t)dt * p.kh + oy % p.kh) * p.kw + x % p.kw;
            }
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              if (p.out_dtype == LVC_DT_F32) __stcs(reinterpret_cast<float*>(p.out) + e + c * cstride, lut[c * 256 + u[c]]);
              else if (p.out_dtype == LVC_DT_BF16) reinterpret_cast<__nv_bfloat16*>(p.out)[e + c * cstride] = __float2bfloat16_rn(lut[c * 256 + u[c]]);
              else reinterpret_cast<uint8_t*>(p.out)[e + c * cstride] = (uint8_t)u[c];
            }
          }
        }
        next_emit = ne;
        ring0 += kCH;
        if (ring0 >= p.ir) ring0 -= p.ir;
        ++rb;
        seg = 0;
      } else {
        ++seg;
      }
    }
    cp_async_wait0();
  }
}

template <int NT>
int pil_launch(const PilParams& p, int grid, int threads, size_t smem, void* stream) {
  LVC_CUDA_TRY(cudaFuncSetAttribute(lvc_pil_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  lvc_pil_kernel<NT><<<grid, threads, smem, static_cast<cudaStream_t>(stream)>>>(p);
  LVC_CUDA_TRY(cudaGetLastError());
  return LVC_OK;
}

int pil_device(int* sms, int* max_smem) {
  int dev = -1;
  LVC_CUDA_TRY(cudaGetDevice(&dev));
  LVC_CUDA_TRY(cudaDeviceGetAttribute(sms, cudaDevAttrMultiProcessorCount, dev));
  LVC_CUDA_TRY(cudaDeviceGetAttribute(max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  return LVC_OK;
}

}  // namespace

extern "C" int lvc_hf_clip_forward_ex(const lvc_clip_desc* d, int32_t shortest_edge, int32_t max_height, int32_t flags,
                                      const lvc_video* videos, const int32_t* indices, const float* lut, void* out,
                                      void* stream) {
  if (!d || !videos || !indices || !out) return lvc_fail(LVC_E_INVAL, "lvc_hf_clip_forward: NULL argument");
  if (d->out_size != S) return lvc_fail(LVC_E_SHAPE, "lvc_hf_clip_forward: out_size %d unsupported (224 only)", d->out_size);
  if (shortest_edge != 0 && shortest_edge < S) return lvc_fail(LVC_E_SHAPE, "lvc_hf_clip_forward: shortest_edge %d < crop %d (the processor would pad)", shortest_edge, S);
  if (d->batch <= 0 || d->num_frames <= 0 || d->max_width <= 0 || max_height <= 0)
    return lvc_fail(LVC_E_INVAL, "lvc_hf_clip_forward: batch, num_frames, max_width and max_height must be positive");
  if (d->layout < LVC_LAYOUT_BTCHW || d->layout > LVC_LAYOUT_TOKENS || d->out_dtype < LVC_DT_F32 || d->out_dtype > LVC_DT_U8)
    return lvc_fail(LVC_E_INVAL, "lvc_hf_clip_forward: unknown layout or dtype");
  if (d->layout == LVC_LAYOUT_TOKENS) {
    if (d->out_dtype != LVC_DT_BF16) return lvc_fail(LVC_E_NOTSUP, "lvc_hf_clip_forward: the token layout is the bf16 GEMM operand");
    if (d->tok_kt <= 0 || d->tok_kh <= 0 || d->tok_kw <= 0 || d->num_frames % d->tok_kt || S % d->tok_kh || S % d->tok_kw)
      return lvc_fail(LVC_E_SHAPE, "lvc_hf_clip_forward: patch (%d,%d,%d) must divide (T,224,224)", d->tok_kt, d->tok_kh, d->tok_kw);
  }
  if (d->layout == LVC_LAYOUT_BTHWC && d->out_dtype != LVC_DT_U8)
    return lvc_fail(LVC_E_NOTSUP, "lvc_hf_clip_forward: BTHWC layout is the uint8 clip only");
  if (d->out_dtype != LVC_DT_U8 && !lut) return lvc_fail(LVC_E_INVAL, "lvc_hf_clip_forward: float output needs a LUT (lvc_build_lut)");
  if (d->src_float) return lvc_fail(LVC_E_NOTSUP, "lvc_hf_clip_forward: Pillow resizes uint8 frames");
  if ((int64_t)d->batch * d->num_frames > (1 << 22)) return lvc_fail(LVC_E_SHAPE, "lvc_hf_clip_forward: too many frames in one call");
  int sms = 0, max_smem = 0;
  int rc = pil_device(&sms, &max_smem);
  if (rc != LVC_OK) return rc;

  // Down-scale bound: short side <= min(max_w, max_h); the long side's scale W / int(E*W/H) exceeds short/E by at
  // most a factor (1 + 1/E).  The tap tables hold 2*ceil(scale) + 1 <= KT entries per sample.
  const double scale = shortest_edge == 0
      ? std::max(1.0, (double)std::max(d->max_width, max_height) / S)   // exact 224 x 224: each axis has its own scale
      : std::max(1.0, (double)std::min(d->max_width, max_height) / (double)shortest_edge * (1.0 + 1.0 / shortest_edge));
  const int cs = (int)std::ceil(scale);
  if (std::ceil(2.0 * scale) > KT - 1) return lvc_fail(LVC_E_SHAPE, "lvc_hf_clip_forward: down-scale %.2f needs more than %d taps", scale, KT);
  PilParams p{};
  p.videos = videos; p.indices = indices; p.lut = lut; p.out = out;
  p.T = d->num_frames; p.layout = d->layout; p.out_dtype = d->out_dtype; p.bgr = d->src_order == LVC_SRC_BGR;
  p.bulk = (flags & LVC_HF_SRC_HOST) != 0;
  p.max_w = d->max_width; p.max_h = max_height;
  p.err = lvc_dev_error_word();
  if (!p.err) return lvc_fail(LVC_E_CUDA, "lvc_hf_clip_forward: cannot resolve the device error word");
  p.edge = shortest_edge; p.kt_tok = d->tok_kt; p.kh = d->tok_kh; p.kw = d->tok_kw;
  // a sample's window [int(c - s + .5), int(c + s + .5)) holds fewer than 2s + 1, i.e. at most ceil(2s), samples
  static const int kTapSteps[] = {2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 15, 17, 19, 21, 23};
  const int need_taps = (int)std::ceil(2.0 * scale);
  p.kt = 23;
  for (int i = 16; i >= 0; --i) if (kTapSteps[i] >= need_taps) p.kt = kTapSteps[i];
  p.ir = kCH + p.kt + 1;
  // Column segments: the fewest that keep a CTA under 128 KB -- tuned on B200 over 480p / 720p / 1080p at both
  // edges: bigger stages gain nothing, smaller ones add barriers and column overlap (LVC_PIL_NSEG overrides).
  const size_t fixed = (size_t)(2 * p.kt * S + 4 * S + 2 * kCH + 4 + 768) * 4 + (size_t)p.ir * IP * 4;
  static const int kSegs[] = {1, 2, 4, 7, 14, 28};
  size_t smem = 0;
  int pick = -1;
  const int force_seg = lvc_knob_int("LVC_PIL_NSEG", 0);   // tuning override (gated: lvc_internal.h)
  for (int pass = 0; pass < 2 && pick < 0; ++pass)
    for (int i = 0; i < 6; ++i) {
      if (force_seg > 0 && (kSegs[i] != force_seg || pass == 0)) continue;
      if (p.bulk && pass == 0) continue;   // over PCIe: the fewest, largest row copies that fit at all
      const int xs = S / kSegs[i];
      const int span_px = std::min((int)(xs * scale) + 2 * cs + 3, d->max_width);
      int pw = (span_px * 3 + 6) / 4 + 1;
      pw |= 1;
      if (p.bulk) pw = (((span_px * 3 + 15 + 15) & ~15) + 16) / 4;   // 16-byte granules for the bulk copies
      const size_t need = fixed + (size_t)2 * kCH * pw * 4 + 144;   // slack: padding taps read past the last staged row
      if (need <= (pass == 0 ? (size_t)128 * 1024 : (size_t)max_smem)) {
        pick = i; p.nseg = kSegs[i]; p.pitch_words = pw; smem = need;
        break;
      }
    }
  if (pick < 0) return lvc_fail(LVC_E_SHAPE, "lvc_hf_clip_forward: frames too large for shared memory");
  int threads = 512;
  if (const int env_t = lvc_knob_int("LVC_PIL_THREADS", 0)) threads = std::max(256, std::min(kMaxThreads, env_t / 32 * 32));
  const int per_sm = std::max(1, std::min(std::min(2, 2048 / threads), (int)((size_t)max_smem / (smem + 1024))));
  const int slots = sms * per_sm;
  // Bands per frame: minimise rounds x source rows per item (each band re-reads a halo of ~kt rows)
  const int frames = d->batch * d->num_frames;
  static const int kBands[] = {1, 2, 4, 7, 8, 14};
  double best = 1e30;
  for (int i = 0; i < 6; ++i) {
    const int nb = kBands[i], R = S / nb;
    const double cost = std::ceil((double)frames * nb / slots) * (R * scale + p.kt + 16);
    if (cost < best * 0.97) { best = cost; p.bands = nb; p.R = R; }
  }
  if (const int nb = lvc_knob_int("LVC_PIL_BANDS", 0)) {   // tuning override
    if (nb > 0 && S % nb == 0) { p.bands = nb; p.R = S / nb; }
  }
  p.total_items = frames * p.bands;
  const int grid = std::min(p.total_items, slots);
  switch (p.kt) {
#define LVC_PIL_CASE(N) case N: return pil_launch<N>(p, grid, threads, smem, stream);
    LVC_PIL_CASE(2) LVC_PIL_CASE(3) LVC_PIL_CASE(4) LVC_PIL_CASE(5) LVC_PIL_CASE(6) LVC_PIL_CASE(7) LVC_PIL_CASE(8) LVC_PIL_CASE(9)
    LVC_PIL_CASE(10) LVC_PIL_CASE(11) LVC_PIL_CASE(12) LVC_PIL_CASE(13) LVC_PIL_CASE(15) LVC_PIL_CASE(17)
    LVC_PIL_CASE(19) LVC_PIL_CASE(21)
#undef LVC_PIL_CASE
    default: return pil_launch<23>(p, grid, threads, smem, stream);
  }
}

extern "C" int lvc_hf_clip_forward(const lvc_clip_desc* d, int32_t shortest_edge, int32_t max_height,
                                   const lvc_video* videos, const int32_t* indices, const float* lut, void* out,
                                   void* stream) {
  return lvc_hf_clip_forward_ex(d, shortest_edge, max_height, 0, videos, indices, lut, out, stream);
}
