// Fused clip-construction kernel for sm_100a: frame gather by index -> optional ROI window ->
// cv2-exact fixed-point bilinear resize to 224x224 -> LUT normalisation -> clip layout, one pass
// over HBM.  Replaces (paths relative to the reference root)
//   vivit_transformer/vivit_classifier/data_config/dataset.py:246-284   gather + cv2.resize loop
//   resnet50-2d-lstm/inference.py:158-178                               resize + /255 + mean/std + permute
//   vivit_transformer/vivit_classifier/trainers/trainer.py:62-95        image processor + torch.cat
//
// Design (DESIGN.md "clip kernel"): a persistent CTA per SM; four producer warps (stage n -> warp n mod 4) stream
// the TAPPED source rows (2 of every ~scale rows; duplicates inside a stage are fetched once) into a
// shared-memory ring with 1-D bulk async copies (cp.async.bulk, the TMA engine) completing on
// mbarriers, and publish each stage's header (tap coefficients, ring offsets, output addressing);
// 14 consumer warps (2 groups x 224 threads -- 4 groups for the ViViT preset -- one thread per output column,
// x taps in registers for the whole frame) do the horizontal + vertical fixed-point passes out of shared
// memory, look the result up in the LUT and write coalesced rows.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <type_traits>

#include "lvc_internal.h"

namespace {

constexpr int S = 224;                    // output width/height
enum { MODE_CV2 = 0, MODE_VIVIT = 1, MODE_PTV = 2 };
// Consumer groups work on different rows of a stage (224 threads each, thread = output column).  Two groups for the
// cv2 / torch chains; the ViViT preset runs two Pillow passes on top of the resize and is bound by consumer issue,
// so it gets four (measured with 2+ producer warps: 155 -> 135 us on config 2; the other modes lose with four).
#ifndef LVC_CLIP_MINBLOCKS
#define LVC_CLIP_MINBLOCKS 1   // 2: cap the cv2 / torch kernels at 56 registers so that two CTAs share an SM (experiment)
#endif
#ifndef LVC_CLIP_GROUPS_CV2
#define LVC_CLIP_GROUPS_CV2 2
#endif
__host__ __device__ constexpr int groups_of(int mode) { return mode == MODE_VIVIT ? 4 : LVC_CLIP_GROUPS_CV2; }
// Producer warps: stage n of a CTA's stream belongs to warp n % kProducers.  One warp needs 0.5 us + 0.165 us per
// output row to set up and issue a stage (serial tap / address / header chain, then UBLKCP lane by lane: it is a
// uniform-datapath instruction), which bounded the small-source configs; four warps work on four stages concurrently.
#ifndef LVC_CLIP_PRODUCERS
#define LVC_CLIP_PRODUCERS 4
#endif
constexpr int kProducers = LVC_CLIP_PRODUCERS;
__host__ __device__ constexpr int threads_of(int mode, int producers = kProducers) { return 32 * producers + groups_of(mode) * S; }
// The variant that shares its SM with the patch-embedding kernel (lvc_clip_swin_embed_forward) runs two producer warps:
// on ROI batches two are as good as four (cfg 4: 91 vs 90 us) and 64 fewer threads are 5.6 K registers for the neighbour.
constexpr int kProducersFused = 2;
constexpr int kMaxRPS = 16;               // output rows per stage (2*RPS source-row slots <= 32 lanes)
constexpr int kMaxStages = 16;

struct RowDesc {  // one output row of a stage
  int off0, off1; // byte offsets in the ring of source pixel (x_origin, y0) / (x_origin, y1)
  int b0, b1;     // 11-bit vertical coefficients
};
struct alignas(16) StageHdr {   // (the lean consumers read the leading words as int4: keep their order)
  int frame;      // output frame id b*T + t
  int sw;         // source window width (pixels); PTV: full frame width
  int oy0;        // first output row of this stage
  int flags;      // bit 0: exact 2x2 down-scale (cv::resize switches to INTER_AREA); bit 1: hflip (PTV train chain);
                  // bit 2: the item's rows do not fit the ring (contract violation): nothing was fetched, skip the rows
  int ox;         // PTV: first source column held in the row slots
  int nw;         // PTV: width of the short-side-scaled image
  int left;       // PTV: centre-crop offset in the scaled image
  int rlo;        // VIVIT bands: first resized row held in the intermediate buffer (nw = first row the NEXT band owns)
  // output addressing, worked out once per item by the producer (no divisions in the consumers)
  unsigned long long obase;    // element offset of (frame, channel 0, row 0, column 0) -- tokens: of patch row 0
  unsigned long long cstride;  // channel stride in elements
  unsigned rowmul;             // planar layouts: elements per output row; tokens: Wo*K (per patch row)
  int clip;                    // b = frame / T (stats slot)
  double xscale;               // cv::resize column scale of the item's window (two double divisions, done once by the producer)
  unsigned roff[kMaxRPS];      // element offset of each output row of the stage inside the frame
  RowDesc rows[kMaxRPS];
};
static_assert(sizeof(StageHdr) % 16 == 0 && offsetof(StageHdr, flags) == 12 && offsetof(StageHdr, ox) == 16 && offsetof(StageHdr, rlo) == 28,
              "lean consumers load the header's leading words as two int4");
struct ClipParams {
  const lvc_video* videos;
  const int32_t* indices;
  const int32_t* boxes;
  const int32_t* aug;        // PTV: per-clip {short side, crop top, crop left, flip} or NULL (eval chain)
  const float* lut;
  void* out;
  unsigned long long* stats;
  int T;
  int layout, out_dtype;
  int bands, rows_per_band;  // bands per frame, output rows per band
  int v_stages;              // VIVIT: stages per band (resized rows incl. the Pillow halo, rounded up to rps)
  int ring_off;              // byte offset of the source-row ring in shared memory
  int total_items;           // B*T*bands
  int rps, stages, pitch;    // rows per stage, barrier slots in use (= kMaxStages), bytes per row slot of the WIDEST video
  int ring_bytes;            // size of the byte-granular source-row ring
  int lut4_off;              // lean cv2 consumer: byte offset of the 3 x 1024 table indexed by the unrounded vertical sum
  int no_fma;                // profiling knob LVC_CLIP_NO_FMA: keep the bf16 look-up even when the affine form verifies
  int no_tok44;              // profiling knob LVC_CLIP_NO_TOK44: 4 x 4 token blocks as four 2-byte stores per channel
  unsigned* tile_ready;      // fused patch embedding: counters of token-row slices written per tile of A, else NULL
  int tile_shift;            // log2(tokens per tile)
  unsigned long long* tile_stamps;   // fused patch embedding: CTA 0 stamps its start / end here (timeline of the pair)
  unsigned* err;             // device error word (LVC_DEVERR_*), never NULL
  unsigned long long* fetched;  // device counter: bytes the producers requested from the source buffers (bulk copies)
  unsigned long long* trace;    // profiling (LVC_CLIP_TRACE=1 under the knob gate): CTA 0 logs 4 timestamps per stage, else NULL
  int producers;             // active producer warps (<= kProducers, <= stages)
  int merge_rows;            // fetch the (y, y+1) rows of a full-width tap pair with one copy
  float mean[3], std[3];     // PTV normalisation
  int kt, kh, kw;            // LVC_LAYOUT_TOKENS: patch size; output is the GEMM operand A[M][K] (bf16)
  int kh_shift;              // log2(kh) when kh is a power of two, else -1
  int f32src;                // cv2.resize as the reference runs it on FLOAT frames, result truncated to uint8
  int bgr;                   // source frames are BGR (cv2.VideoCapture order): swap channels in the gather
};

// ---- PTX wrappers -------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"   // suspend-time hint: the thread still wakes
        "selp.u32 %0, 1, 0, p;\n\t}"                                        // when the phase completes, it just polls less
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity), "r"(0x989680u)
        : "memory");
  } while (!done);
}
// 1-D bulk async copy global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP).
// Source rows are read once: evict-first L2 policy.
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar,
                                         uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
      ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}
// The per-stage timeline (scripts/clip_trace.py) is a profiling aid: its checks sit in the stage loops of every consumer and
// producer, so it is compiled in only with -DLVC_CLIP_TRACE_BUILD=1 (`make trace` builds liblvc_b200_trace.so beside the
// shipped library; the shipped kernels carry none of it).
#ifndef LVC_CLIP_TRACE_BUILD
#define LVC_CLIP_TRACE_BUILD 0
#endif
constexpr bool kTraceBuild = LVC_CLIP_TRACE_BUILD != 0;
constexpr int kTraceStages = 4096;
constexpr int kTraceWords = kTraceStages * 4 + 2 * 1024;   // per-stage stamps of CTA 0, then (start, end) of every CTA
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}

// OpenCV INTER_LINEAR tap for destination coordinate d: fx = (float)((d+0.5)*scale - 0.5) in
// double without contraction, then floor / fraction in float (cv::resize table set-up).
__device__ __forceinline__ void cv_tap(int d, double scale, int& s, float& fr) {
  float f = (float)__dadd_rn(__dmul_rn((double)d + 0.5, scale), -0.5);
  float fl = floorf(f);
  s = (int)fl;
  fr = __fsub_rn(f, fl);
}
// Float-frame variant (cv2.resize on float32 input, vivit .../dataset.py:244-284): the fraction keeps
// double precision until the final cast, which reproduces the installed OpenCV to 1 float32 ulp.
__device__ __forceinline__ void cv_tap64(int d, double scale, int& s, float& fr) {
  const double f = __dadd_rn(__dmul_rn((double)d + 0.5, scale), -0.5);
  const double fl = floor(f);
  s = (int)fl;
  fr = (float)__dsub_rn(f, fl);
}
__device__ __forceinline__ double cv_scale(int src) {
  double inv = __ddiv_rn((double)S, (double)src);
  return __ddiv_rn(1.0, inv);
}
__device__ __forceinline__ int coef(float w) { return __float2int_rn(__fmul_rn(w, 2048.0f)); }

}  // namespace

namespace {

// Shared memory map: [0,256) full/empty mbarriers | StageHdr[kMaxStages] | LUT 3x256 f32 | LUT 3x256 bf16 |
// (MODE_VIVIT: Pillow tap table, 2x2 row buffers, intermediate rows of one band x 672 B) | ring.
constexpr int kHdrOff = 256;
constexpr int kLutOff = kHdrOff + kMaxStages * (int)sizeof(StageHdr);
constexpr int kLut16Off = kLutOff + 3 * 256 * 4;    // the same table rounded to bf16 (bf16 outputs skip the per-sample convert)
constexpr int kAuxOff = (kLut16Off + 3 * 256 * 2 + 127) / 128 * 128;
constexpr int kRowBytes = S * 4;                    // one resized row, a 32-bit word (R | G<<8 | B<<16) per pixel
constexpr int kVivitTapOff = kAuxOff;               // int4[224]: ymin, k0, k1, k2
constexpr int kVivitRowPartOff = kVivitTapOff + S * 16;  // unsigned[224]: row term of the output offset (row_part) of every output row
constexpr int kVivitRowBufOff = kVivitRowPartOff + S * 4;    // [group][parity][224 words]
constexpr int kVivitInterOff = kVivitRowBufOff + groups_of(MODE_VIVIT) * 2 * kRowBytes;

// Pillow BILINEAR 224 -> 256 taps for cropped output coordinate d (source coordinate d + 16 of the
// 256-wide image): precompute_coeffs + normalize_coeffs_8bpc in double, support 1 (up-scale).
__device__ __forceinline__ int4 pil_tap(int d) {
  const double scale = 224.0 / 256.0;
  const double center = __dmul_rn((double)(d + 16) + 0.5, scale);
  int xmin = (int)(__dadd_rn(__dadd_rn(center, -1.0), 0.5));
  if (xmin < 0) xmin = 0;
  int xmax = (int)(__dadd_rn(__dadd_rn(center, 1.0), 0.5));
  if (xmax > S) xmax = S;
  const int n = xmax - xmin;
  double w[3] = {0.0, 0.0, 0.0}, ww = 0.0;
  for (int i = 0; i < 3; ++i)
    if (i < n) {
      double t = __dadd_rn(__dadd_rn((double)(i + xmin), -center), 0.5);
      t = t < 0 ? -t : t;
      w[i] = t < 1.0 ? __dadd_rn(1.0, -t) : 0.0;
      ww = __dadd_rn(ww, w[i]);
    }
  int k[3];
  for (int i = 0; i < 3; ++i) {
    double v = (ww != 0.0) ? __ddiv_rn(w[i], ww) : w[i];
    k[i] = (int)(__dadd_rn(0.5, __dmul_rn(v, 4194304.0)));
  }
  return make_int4(xmin, k[0], k[1], k[2]);
}
// Channel c of a packed pixel word is taken with one PRMT (__byte_perm(w, 0, 0x4440 + c)) instead of shift + mask.
// Pillow's clip8 is the identity here: bilinear weights are non-negative and their integer sum is within
// 2^22 +- 2 (<= 3 taps, each rounded to nearest), so 2^21 + sum(p*k) >> 22 stays in [0, 255] for p in [0, 255].
__device__ __forceinline__ int clip8(int v) { return v; }

// torch bilinear (align_corners=False) source coordinate, exactly as ATen's CPU kernel evaluates it:
// src = max(fma(scale, dst + 0.5, -0.5), 0); i0 = min(floor(src), in-1); lambda = clamp(src - i0, 0, 1).
__device__ __forceinline__ void torch_tap(int dst, float scale, int in, int& i0, int& i1, float& l0, float& l1) {
  float src = fmaxf(__fmaf_rn(scale, (float)dst + 0.5f, -0.5f), 0.0f);
  i0 = min((int)floorf(src), in - 1);
  l1 = fminf(fmaxf(__fsub_rn(src, (float)i0), 0.0f), 1.0f);
  l0 = __fsub_rn(1.0f, l1);
  i1 = i0 + (i0 < in - 1 ? 1 : 0);
}
// pytorchvideo ShortSideScale(size) geometry for an HxW frame, then the crop origin: torchvision
// CenterCrop(224) when `aug` is absent (eval chain), else the drawn RandomCrop origin (train chain).
__device__ __forceinline__ void ptv_geom(int H, int W, int size, int& nh, int& nw) {
  // python: int(math.floor((float(h) / w) * size)), in double
  if (W < H) { nw = size; nh = (int)floor(__dmul_rn(__ddiv_rn((double)H, (double)W), (double)size)); }
  else { nh = size; nw = (int)floor(__dmul_rn(__ddiv_rn((double)W, (double)H), (double)size)); }
}
__device__ __forceinline__ int center_crop_origin(int n) { return (int)rint((double)(n - S) / 2.0); }
__device__ __forceinline__ void ring_next(int& slot, uint32_t& phase, int stages) {
  if (++slot == stages) { slot = 0; phase ^= 1u; }
}
__device__ __forceinline__ void group_bar(int g) {
  asm volatile("bar.sync %0, %1;" ::"r"(2 + g), "r"(S) : "memory");
}
__device__ __forceinline__ void consumers_bar(int consumers) {
  asm volatile("bar.sync 1, %0;" ::"r"(consumers) : "memory");
}

}  // namespace

namespace {

// Element offset of output sample (frame f = b*T + t, channel 0, row y, column 0) and the channel
// stride, for every layout; the per-thread column term comes from col_offset().  LVC_LAYOUT_TOKENS
// writes the patch-embedding operand A[m][k] directly: m = ((b*To + t/kt)*Ho + y/kh)*Wo + x/kw,
// k = ((c*kt + t%kt)*kh + y%kh)*kw + x%kw (token order of the reference modules).
struct FrameOut {   // per-frame part of the output addressing, computed once per stage
  size_t base;      // element offset of (channel 0, row 0, column 0) -- tokens: of patch row 0
  size_t cstride;   // channel stride in elements
  unsigned rowmul;  // planar layouts: elements per output row; tokens: Wo*K (per patch row)
};
__device__ __forceinline__ FrameOut frame_out(const ClipParams& p, int f, int b, int t) {
  const size_t plane = (size_t)S * S;
  FrameOut o;
  switch (p.layout) {
    case LVC_LAYOUT_BTCHW: o.cstride = plane; o.base = (size_t)f * 3 * plane; o.rowmul = S; break;
    case LVC_LAYOUT_BCTHW: o.cstride = (size_t)p.T * plane; o.base = ((size_t)b * 3 * p.T + t) * plane; o.rowmul = S; break;
    case LVC_LAYOUT_BTHWC: o.cstride = 1; o.base = (size_t)f * plane * 3; o.rowmul = S * 3; break;
    default: {
      const unsigned To = p.T / p.kt, Ho = S / p.kh, Wo = S / p.kw, K = 3 * p.kt * p.kh * p.kw;
      const unsigned to = (unsigned)t / (unsigned)p.kt, dt = (unsigned)t - to * p.kt;
      o.cstride = (size_t)p.kt * p.kh * p.kw;
      o.rowmul = Wo * K;
      o.base = ((size_t)b * To + to) * Ho * o.rowmul + (size_t)dt * p.kh * p.kw;
    }
  }
  return o;
}
// row term of the output offset (< 2^32 elements inside one frame)
__device__ __forceinline__ unsigned row_part(const ClipParams& p, unsigned rowmul, int y) {
  if (p.layout != LVC_LAYOUT_TOKENS) return (unsigned)y * rowmul;
  const unsigned ho = p.kh_shift >= 0 ? ((unsigned)y >> p.kh_shift) : ((unsigned)y / (unsigned)p.kh);
  const unsigned dy = (unsigned)y - ho * p.kh;
  return ho * rowmul + dy * (unsigned)p.kw;
}
__device__ __forceinline__ unsigned col_offset(const ClipParams& p, int x) {
  if (p.layout == LVC_LAYOUT_BTHWC) return (unsigned)x * 3u;
  if (p.layout == LVC_LAYOUT_TOKENS) return (unsigned)(x / p.kw) * (unsigned)(3 * p.kt * p.kh * p.kw) + (unsigned)(x % p.kw);
  return (unsigned)x;
}


// ---- lean consumer of the cv2 chain with a compile-time output type -----------------------------------------------
// Same arithmetic as the general consumer below (it stays for float frames, the ViViT and the torch chains), built for
// instruction count: the consumers of the small-source configs are issue-bound (ncu: 114 warp instructions per output
// row per warp, 0.5 IPC per scheduler, data always waiting in the ring), so every instruction removed from the row body
// is time.  Against the general path: output type / stats / source type are template parameters (no per-stage dispatch);
// shared addresses are 32-bit window addresses with the ring base folded into the per-frame column offset (add, and,
// shift per source row instead of five instructions); one PRMT serves two channels (DP2A on the low and on the high
// byte pair); float outputs without a checksum index a 1024-entry table by the unrounded vertical sum t0 + t1, which
// folds cv2's (+2) >> 2 into the table; the layout's channel stride is a per-launch constant, not a header field.
__device__ __forceinline__ void lds3(uint32_t addr, uint32_t& a, uint32_t& b, uint32_t& c) {
  asm volatile("ld.shared.u32 %0, [%3];\n\tld.shared.u32 %1, [%3+4];\n\tld.shared.u32 %2, [%3+8];"
               : "=r"(a), "=r"(b), "=r"(c) : "r"(addr));
}
__device__ __forceinline__ uint32_t shf_r_wrap(uint32_t lo, uint32_t hi, uint32_t s) {
  uint32_t d;
  asm("shf.r.wrap.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(lo), "r"(hi), "r"(s));
  return d;
}

// Token layout of a 4 x 4 patch (Video Swin): a token's 16 samples of one (channel, tubelet frame) are 32 contiguous bytes
// = one sector, but a thread (= column) holds one sample of each of the patch's four rows.  Stored directly that is four
// 2-byte stores per channel whose warp-wide footprint is 8-byte fragments of eight different sectors (9.5 M written
// sectors for 77 MB; the stage took 86 us against 68 us for planar bf16).  Instead the four lanes of a token transpose
// their 4 x 4 block with two shuffles: P0 = (row 0, row 1), P1 = (row 2, row 3) of this lane's column, packed bf16;
// afterwards lane dx holds ROW dx of the token, columns 0..3 = 8 contiguous bytes, and four lanes write one whole sector.
__device__ __forceinline__ void tok44_transpose(uint32_t P0, uint32_t P1, int dx, uint32_t& w0, uint32_t& w1) {
  const bool hi2 = (dx & 2) != 0, odd = (dx & 1) != 0;
  const uint32_t keep1 = hi2 ? P1 : P0;
  const uint32_t recv1 = __shfl_xor_sync(0xffffffffu, hi2 ? P0 : P1, 2);   // rows (rb, rb + 1), rb = dx & 2, of column dx ^ 2
  const uint32_t clo = hi2 ? recv1 : keep1, chi = hi2 ? keep1 : recv1;     // ... ordered by column: dx & 1, (dx & 1) + 2
  const uint32_t rowA = __byte_perm(clo, chi, 0x5410), rowB = __byte_perm(clo, chi, 0x7632);   // row rb / rb + 1 of both
  const uint32_t keep2 = odd ? rowB : rowA;
  const uint32_t recv2 = __shfl_xor_sync(0xffffffffu, odd ? rowA : rowB, 1);   // the same row of the neighbour's two columns
  const uint32_t ev = odd ? recv2 : keep2, od = odd ? keep2 : recv2;            // columns (0, 2) / (1, 3) of row dx
  w0 = __byte_perm(ev, od, 0x5410);   // columns 0, 1
  w1 = __byte_perm(ev, od, 0x7632);   // columns 2, 3
}

template <int ODT, bool STATS>
__device__ __forceinline__ void lean_cv2_consumer(const ClipParams& p, uint8_t* smem, uint64_t* full, uint64_t* empty,
                                                  const StageHdr* hdrs, const int ct, const int lane, const bool use_fma,
                                                  const float (&fa)[3], const float (&fb)[3]) {
  constexpr int kGroups = groups_of(MODE_CV2), kConsumers = kGroups * S;
  constexpr int kEsz = ODT == LVC_DT_F32 ? 4 : (ODT == LVC_DT_BF16 ? 2 : 1);
  constexpr bool kT4 = !STATS && ODT != LVC_DT_U8;   // table indexed by the unrounded vertical sum
  const int g = ct / S, x = ct - g * S;
  const uint32_t ring_s = smem_u32(smem) + (uint32_t)p.ring_off;
  const unsigned xoff = col_offset(p, x);
  // channel stride of the layout in bytes (a per-launch constant)
  unsigned cs;
  switch (p.layout) {
    case LVC_LAYOUT_BTCHW: cs = S * S; break;
    case LVC_LAYOUT_BCTHW: cs = (unsigned)p.T * S * S; break;
    case LVC_LAYOUT_BTHWC: cs = 1; break;
    default: cs = (unsigned)(p.kt * p.kh * p.kw);
  }
  cs *= kEsz;
  const float* lutf = reinterpret_cast<const float*>(smem + (kT4 ? p.lut4_off : kLutOff));
  const __nv_bfloat16* luth = reinterpret_cast<const __nv_bfloat16*>(smem + (kT4 ? p.lut4_off : kLut16Off));
  constexpr int kLutStride = kT4 ? 1024 : 256;
  const uint32_t selA = p.bgr ? 0x4152u : 0x4130u;   // (c0 of pixel 0, c0 of pixel 1, c1 of pixel 0, c1 of pixel 1)
  const uint32_t selB = p.bgr ? 0x4430u : 0x4452u;   // (c2 of pixel 0, c2 of pixel 1, -, -); BGR sources swap c0 / c2
  int cached_sw = -1, cached_flags = -1;
  uint32_t xo_s = 0, a01 = 0;
  unsigned long long acc = 0;
  int acc_b = -1;
  int slot = 0, cn = 0;
  uint32_t phase = 0;
  const int stages_per_item = p.rows_per_band / p.rps;
  const bool pair_rows = (p.rps & (2 * kGroups - 1)) == 0;
  // 4 x 4 patches with one patch row per group and stage: whole-sector token stores (tok44_transpose)
  const bool tok44 = ODT == LVC_DT_BF16 && p.layout == LVC_LAYOUT_TOKENS && p.kh == 4 && p.kw == 4 && p.rps == 4 * kGroups && !p.no_tok44;
  const unsigned xoff44 = (unsigned)(x >> 2) * (unsigned)(3 * p.kt * 16) + (unsigned)(x & 3) * 4u;   // token, then ROW x & 3 of its block

  for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
    for (int st = 0; st < stages_per_item; ++st, ring_next(slot, phase, kMaxStages), ++cn) {
      const bool tr = kTraceBuild && p.trace && blockIdx.x == 0 && ct == 0 && cn < kTraceStages;
      if (tr) p.trace[cn * 4 + 1] = globaltimer_ns();
      mbar_wait(&full[slot], phase);
      if (tr) p.trace[cn * 4 + 2] = globaltimer_ns();
      const StageHdr* h = &hdrs[slot];
      const int4 h0 = reinterpret_cast<const int4*>(h)[0];   // (frame, sw, oy0, flags) in one 16-byte load
      const int sw = h0.y, flags = h0.w;
      if (sw != cached_sw || flags != cached_flags) {
        cached_sw = sw; cached_flags = flags;
        // exact 2x2 (INTER_AREA): a0 = a1 = 32 and b0 = b1 = 2^31 turn the fixed-point blend into
        // (p00 + p01 + p10 + p11 + 2) >> 2: s >> 4 = 2 (p0 + p1), umulhi(2^31, 2 t) = t
        int xo, a0, a1;
        if (flags & 1) { xo = 6 * x; a0 = 32; a1 = 32; }
        else {
          int sx; float fr;
          cv_tap(x, h->xscale, sx, fr);
          if (sx < 0) { sx = 0; fr = 0.f; }
          if (sx >= sw - 1) { sx = sw - 1; fr = 0.f; }   // clamped edge: a1 == 0, the bytes read at +3 do not matter
          a0 = coef(__fsub_rn(1.0f, fr));
          a1 = coef(fr);
          xo = 3 * sx;
        }
        xo_s = ring_s + (uint32_t)xo;
        a01 = (uint32_t)a0 | ((uint32_t)a1 << 16);
      }
      if (STATS) {
        const int b = h->clip;
        if (b != acc_b) {
          if (acc_b >= 0 && p.stats) {
#pragma unroll
            for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (lane == 0 && acc) atomicAdd(&p.stats[acc_b], acc);
          }
          acc = 0; acc_b = b;
        }
      }
      char* const obase = reinterpret_cast<char*>(p.out) + (size_t)h->obase * kEsz;
      unsigned acc32 = 0;
      // one output row: 2 x 3 aligned shared words + funnel shift, exact 11-bit horizontal pass (PRMT + DP2A), cv2's
      // vertical pass as two multiply-highs, table, three stores
      auto vert = [&](const RowDesc& rd, uint32_t a, uint32_t b, uint32_t c, uint32_t o, uint32_t d, uint32_t e, uint32_t f,
                      uint32_t o1, unsigned (&t)[3]) {
        const uint32_t sh0 = o << 3, sh1 = o1 << 3;   // shf wraps at 32: (o & 3) * 8
        const uint32_t lo0 = shf_r_wrap(a, b, sh0), hi0 = shf_r_wrap(b, c, sh0);
        const uint32_t lo1 = shf_r_wrap(d, e, sh1), hi1 = shf_r_wrap(e, f, sh1);
        const uint32_t wa0 = __byte_perm(lo0, hi0, selA), wb0 = __byte_perm(lo0, hi0, selB);
        const uint32_t wa1 = __byte_perm(lo1, hi1, selA), wb1 = __byte_perm(lo1, hi1, selB);
        const unsigned s00 = __dp2a_lo(a01, wa0, 0u), s01 = __dp2a_hi(a01, wa0, 0u), s02 = __dp2a_lo(a01, wb0, 0u);
        const unsigned s10 = __dp2a_lo(a01, wa1, 0u), s11 = __dp2a_hi(a01, wa1, 0u), s12 = __dp2a_lo(a01, wb1, 0u);
        // ((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16): the coefficients arrive as b << 16, each term is one multiply-high
        t[0] = __umulhi((unsigned)rd.b0, s00 >> 4) + __umulhi((unsigned)rd.b1, s10 >> 4);
        t[1] = __umulhi((unsigned)rd.b0, s01 >> 4) + __umulhi((unsigned)rd.b1, s11 >> 4);
        t[2] = __umulhi((unsigned)rd.b0, s02 >> 4) + __umulhi((unsigned)rd.b1, s12 >> 4);
      };
      auto emit = [&](auto FMA, const unsigned eo, unsigned (&t)[3]) {
        constexpr bool kFma = decltype(FMA)::value;
        char* q = obase + (size_t)eo * kEsz;
        if (!kT4 || kFma) {
#pragma unroll
          for (int c = 0; c < 3; ++c) t[c] = (t[c] + 2u) >> 2;
          if (STATS) acc32 += t[0] + t[1] + t[2];
        }
        if (kFma) {   // bf16 only: verified equal to the table for every 8-bit value (kernel prologue)
#pragma unroll
          for (int c = 0; c < 3; ++c)
            *reinterpret_cast<__nv_bfloat16*>(q + (size_t)c * cs) = __float2bfloat16_rn(__fmaf_rn((float)t[c], fa[c], fb[c]));
          return;
        }
        if (ODT == LVC_DT_F32) {
#pragma unroll
          for (int c = 0; c < 3; ++c) __stcs(reinterpret_cast<float*>(q + (size_t)c * cs), lutf[c * kLutStride + t[c]]);
        } else if (ODT == LVC_DT_BF16) {
#pragma unroll
          for (int c = 0; c < 3; ++c) *reinterpret_cast<__nv_bfloat16*>(q + (size_t)c * cs) = luth[c * kLutStride + t[c]];
        } else {
#pragma unroll
          for (int c = 0; c < 3; ++c) *reinterpret_cast<uint8_t*>(q + (size_t)c * cs) = (uint8_t)t[c];
        }
      };
      auto rows = [&](auto FMA) {
        constexpr bool kFma = decltype(FMA)::value;
        if (ODT == LVC_DT_BF16 && tok44) {
          const int r0 = 4 * g;   // rows [4g, 4g + 4) of the stage = one patch row
          unsigned t[4][3];
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            const RowDesc ra = h->rows[r0 + 2 * hh], rb = h->rows[r0 + 2 * hh + 1];
            const uint32_t oa0 = (uint32_t)ra.off0 + xo_s, oa1 = (uint32_t)ra.off1 + xo_s;
            const uint32_t ob0 = (uint32_t)rb.off0 + xo_s, ob1 = (uint32_t)rb.off1 + xo_s;
            uint32_t w[12];
            lds3(oa0 & ~3u, w[0], w[1], w[2]);
            lds3(oa1 & ~3u, w[3], w[4], w[5]);
            lds3(ob0 & ~3u, w[6], w[7], w[8]);
            lds3(ob1 & ~3u, w[9], w[10], w[11]);
            vert(ra, w[0], w[1], w[2], oa0, w[3], w[4], w[5], oa1, t[2 * hh]);
            vert(rb, w[6], w[7], w[8], ob0, w[9], w[10], w[11], ob1, t[2 * hh + 1]);
          }
          if (!kT4 || kFma) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
              for (int c = 0; c < 3; ++c) { t[i][c] = (t[i][c] + 2u) >> 2; if (STATS) acc32 += t[i][c]; }
          }
          char* const q = obase + (size_t)(h->roff[r0] + xoff44) * 2;
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            uint32_t P0, P1;
            if (kFma) {
              const __nv_bfloat162 a2 = __floats2bfloat162_rn(__fmaf_rn((float)t[0][c], fa[c], fb[c]), __fmaf_rn((float)t[1][c], fa[c], fb[c]));
              const __nv_bfloat162 b2 = __floats2bfloat162_rn(__fmaf_rn((float)t[2][c], fa[c], fb[c]), __fmaf_rn((float)t[3][c], fa[c], fb[c]));
              P0 = *reinterpret_cast<const uint32_t*>(&a2);
              P1 = *reinterpret_cast<const uint32_t*>(&b2);
            } else {
              const unsigned short* lb = reinterpret_cast<const unsigned short*>(luth) + c * kLutStride;
              P0 = (uint32_t)lb[t[0][c]] | ((uint32_t)lb[t[1][c]] << 16);
              P1 = (uint32_t)lb[t[2][c]] | ((uint32_t)lb[t[3][c]] << 16);
            }
            uint32_t w0, w1;
            tok44_transpose(P0, P1, x & 3, w0, w1);
            *reinterpret_cast<uint2*>(q + (size_t)c * cs) = make_uint2(w0, w1);
          }
          return;
        }
        if (pair_rows) {
          for (int r = g; r < p.rps; r += 2 * kGroups) {
            const RowDesc ra = h->rows[r], rb = h->rows[r + kGroups];
            const unsigned ea = h->roff[r] + xoff, eb = h->roff[r + kGroups] + xoff;
            const uint32_t oa0 = (uint32_t)ra.off0 + xo_s, oa1 = (uint32_t)ra.off1 + xo_s;
            const uint32_t ob0 = (uint32_t)rb.off0 + xo_s, ob1 = (uint32_t)rb.off1 + xo_s;
            uint32_t w[12];
            lds3(oa0 & ~3u, w[0], w[1], w[2]);
            lds3(oa1 & ~3u, w[3], w[4], w[5]);
            lds3(ob0 & ~3u, w[6], w[7], w[8]);
            lds3(ob1 & ~3u, w[9], w[10], w[11]);
            unsigned ta[3], tb[3];
            vert(ra, w[0], w[1], w[2], oa0, w[3], w[4], w[5], oa1, ta);
            vert(rb, w[6], w[7], w[8], ob0, w[9], w[10], w[11], ob1, tb);
            emit(FMA, ea, ta);
            emit(FMA, eb, tb);
          }
        } else {
          for (int r = g; r < p.rps; r += kGroups) {
            const RowDesc ra = h->rows[r];
            const unsigned ea = h->roff[r] + xoff;
            const uint32_t oa0 = (uint32_t)ra.off0 + xo_s, oa1 = (uint32_t)ra.off1 + xo_s;
            uint32_t w[6];
            lds3(oa0 & ~3u, w[0], w[1], w[2]);
            lds3(oa1 & ~3u, w[3], w[4], w[5]);
            unsigned ta[3];
            vert(ra, w[0], w[1], w[2], oa0, w[3], w[4], w[5], oa1, ta);
            emit(FMA, ea, ta);
          }
        }
      };
      if (!(flags & 4)) {
        if (ODT == LVC_DT_BF16 && use_fma) rows(std::true_type{});
        else rows(std::false_type{});
      }
      if (STATS) acc += acc32;
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);
      if (tr) p.trace[cn * 4 + 3] = globaltimer_ns();
    }
    if (ODT == LVC_DT_BF16 && p.tile_ready) {
      // Fused patch embedding: tell the GEMM kernel running beside this one which tokens gained their rows.  The item's
      // band of R output rows of frame f (R % kh == 0) holds, for the (R / kh) * Wo consecutive tokens starting at
      // (f / kt * Ho + y0 / kh) * Wo, the kh row slices of tubelet frame f % kt: every tile the range overlaps gets
      // overlap * kh.  Once per item, not per stage: the release needs a gpu-scope fence behind ALL of the CTA's stores,
      // which is worth a microsecond.  Order: stores -> consumer barrier (CTA scope) -> fence (gpu scope, cumulative) -> red.
      consumers_bar(kConsumers);
      if (ct == 0) {
        const unsigned Wo = (unsigned)(S / p.kw), Ho = (unsigned)(S / p.kh), tile = 1u << p.tile_shift;
        const unsigned f = (unsigned)item / (unsigned)p.bands, band = (unsigned)item - f * (unsigned)p.bands;
        const unsigned y0 = band * (unsigned)p.rows_per_band;
        unsigned tok = ((f / (unsigned)p.kt) * Ho + y0 / (unsigned)p.kh) * Wo;
        unsigned left = ((unsigned)p.rows_per_band / (unsigned)p.kh) * Wo;
        __threadfence();
        while (left) {
          const unsigned n = min(left, tile - (tok & (tile - 1u)));
          asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(p.tile_ready + (tok >> p.tile_shift)), "r"(n * (unsigned)p.kh) : "memory");
          tok += n; left -= n;
        }
      }
    }
  }
  if (STATS && p.stats && acc_b >= 0) {
#pragma unroll
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0 && acc) atomicAdd(&p.stats[acc_b], acc);
  }
  if (kTraceBuild && p.trace && ct == 0) p.trace[kTraceStages * 4 + blockIdx.x * 2 + 1] = globaltimer_ns();   // consumer warp 0 done
  if (p.tile_stamps && ct == 0 && blockIdx.x == 0) p.tile_stamps[1] = globaltimer_ns();
}

// ---- lean consumer of the ViViT preset ---------------------------------------------------------------------------------
// cv2 resize row (lean arithmetic as above) -> Pillow horizontal 224 -> 256 pass (cropped columns) into the band's
// intermediate buffer; after the band's last stage the Pillow vertical pass, normalisation and store.  This preset runs
// 28 consumer warps and IS issue-bound (ncu: issue slots 76 % busy), so the instruction count is the time: against the
// general consumer the per-stage bookkeeping shrinks to what this mode reads, the resize uses the lean row body, the output
// type is resolved once per band instead of once per sample, and bf16 outputs take the verified affine form instead of
// three bank-conflicted table look-ups per sample.
// Pillow's BILINEAR 224 -> 256 pass on one packed pixel pair (R | G << 8 | B << 16).  Pillow computes, per channel,
// (2^21 + p0*k0 + p1*k1 [+ p2*k2]) >> 22 with 22-bit weights; for THIS geometry (scale 7/8, sample centres on multiples of
// 1/16) every sample has exactly two taps and the weights are k = m * 2^18 with m0 + m1 = 16 (pil_tap; the kernel prologue
// checks the table), so the expression is 2^18 * (8 + m0*p0 + m1*p1) >> 22 = (m0*p0 + m1*p1 + 8) >> 4 -- exactly.  With
// 4-bit weights the sums stay below 2^12: R and B go through one 32-bit multiply-add chain in 16-bit fields, G through a
// second (13 instructions per pixel instead of 9 PRMT + 9 IMAD + packing).
__device__ __forceinline__ uint32_t pil16(const uint32_t w0, const uint32_t w1, const uint32_t m0, const uint32_t m1) {
  const uint32_t rb = (w0 & 0x00ff00ffu) * m0 + ((w1 & 0x00ff00ffu) * m1 + 0x00080008u);
  const uint32_t gg = (w0 & 0x0000ff00u) * m0 + ((w1 & 0x0000ff00u) * m1 + 0x00000800u);
  return ((rb >> 4) & 0x00ff00ffu) | ((gg >> 4) & 0x0000ff00u);
}

__device__ __forceinline__ void cv2_row_lean(const RowDesc& rd, const uint32_t xo_s, const uint32_t a01, const uint32_t selA,
                                             const uint32_t selB, unsigned (&v)[3]) {
  const uint32_t o0 = (uint32_t)rd.off0 + xo_s, o1 = (uint32_t)rd.off1 + xo_s;
  uint32_t a, b, c, d, e, f;
  lds3(o0 & ~3u, a, b, c);
  lds3(o1 & ~3u, d, e, f);
  const uint32_t sh0 = o0 << 3, sh1 = o1 << 3;   // shf wraps at 32: (o & 3) * 8
  const uint32_t lo0 = shf_r_wrap(a, b, sh0), hi0 = shf_r_wrap(b, c, sh0);
  const uint32_t lo1 = shf_r_wrap(d, e, sh1), hi1 = shf_r_wrap(e, f, sh1);
  const uint32_t wa0 = __byte_perm(lo0, hi0, selA), wb0 = __byte_perm(lo0, hi0, selB);
  const uint32_t wa1 = __byte_perm(lo1, hi1, selA), wb1 = __byte_perm(lo1, hi1, selB);
  const unsigned s00 = __dp2a_lo(a01, wa0, 0u), s01 = __dp2a_hi(a01, wa0, 0u), s02 = __dp2a_lo(a01, wb0, 0u);
  const unsigned s10 = __dp2a_lo(a01, wa1, 0u), s11 = __dp2a_hi(a01, wa1, 0u), s12 = __dp2a_lo(a01, wb1, 0u);
  v[0] = (__umulhi((unsigned)rd.b0, s00 >> 4) + __umulhi((unsigned)rd.b1, s10 >> 4) + 2u) >> 2;
  v[1] = (__umulhi((unsigned)rd.b0, s01 >> 4) + __umulhi((unsigned)rd.b1, s11 >> 4) + 2u) >> 2;
  v[2] = (__umulhi((unsigned)rd.b0, s02 >> 4) + __umulhi((unsigned)rd.b1, s12 >> 4) + 2u) >> 2;
}

template <bool STATS>
__device__ __forceinline__ void lean_vivit_consumer(const ClipParams& p, uint8_t* smem, uint64_t* full, uint64_t* empty,
                                                    const StageHdr* hdrs, const int ct, const int lane, const bool use_fma,
                                                    const float (&fa)[3], const float (&fb)[3]) {
  constexpr int kGroups = groups_of(MODE_VIVIT), kConsumers = kGroups * S;
  const int g = ct / S, x = ct - g * S;
  const uint32_t ring_s = smem_u32(smem) + (uint32_t)p.ring_off;
  const float* lut = reinterpret_cast<const float*>(smem + kLutOff);
  const __nv_bfloat16* lut16 = reinterpret_cast<const __nv_bfloat16*>(smem + kLut16Off);
  const int4* vt = reinterpret_cast<const int4*>(smem + kVivitTapOff);
  const unsigned* rpart = reinterpret_cast<const unsigned*>(smem + kVivitRowPartOff);
  const int4 hx = vt[x];   // Pillow taps of this thread's (cropped) column: first source column, 22-bit weights
  const uint32_t hm0 = (uint32_t)hx.y >> 18, hm1 = (uint32_t)hx.z >> 18;   // ... which are m * 2^18 here: see pil16
  uint32_t* rbuf = reinterpret_cast<uint32_t*>(smem + kVivitRowBufOff) + g * 2 * S;   // [parity][S] per group
  uint32_t* inter = reinterpret_cast<uint32_t*>(smem + kVivitInterOff);
  const uint32_t selA = p.bgr ? 0x4152u : 0x4130u, selB = p.bgr ? 0x4430u : 0x4452u;
  const unsigned xoff = col_offset(p, x);
  // output addressing of the layout: channel stride (elements) and the multiplier of row_part()
  unsigned cs;   // (the row term comes from the kernel prologue's table, rpart[])
  switch (p.layout) {
    case LVC_LAYOUT_BTCHW: cs = S * S; break;
    case LVC_LAYOUT_BCTHW: cs = (unsigned)p.T * S * S; break;
    case LVC_LAYOUT_BTHWC: cs = 1; break;
    default: cs = (unsigned)(p.kt * p.kh * p.kw);
  }
  int cached_sw = -1, cached_flags = -1;
  uint32_t xo_s = 0, a01 = 0;
  unsigned long long acc = 0;
  int acc_b = -1, parity = 0;
  int slot = 0, cn = 0;
  uint32_t phase = 0;

  for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
    bool item_bad = false;
    unsigned long long obase = 0;
    for (int st = 0; st < p.v_stages; ++st, ring_next(slot, phase, kMaxStages), ++cn) {
      const bool tr = kTraceBuild && p.trace && blockIdx.x == 0 && ct == 0 && cn < kTraceStages;
      if (tr) p.trace[cn * 4 + 1] = globaltimer_ns();
      mbar_wait(&full[slot], phase);
      if (tr) p.trace[cn * 4 + 2] = globaltimer_ns();
      const StageHdr* h = &hdrs[slot];
      // the eight leading header words as two 16-byte loads: (frame, sw, oy0, flags), (ox, own_hi, left, rlo)
      const int4 h0 = reinterpret_cast<const int4*>(h)[0], h1 = reinterpret_cast<const int4*>(h)[1];
      const int sw = h0.y, flags = h0.w;
      if (sw != cached_sw || flags != cached_flags) {
        cached_sw = sw; cached_flags = flags;
        int xo, a0, a1;
        if (flags & 1) { xo = 6 * x; a0 = 32; a1 = 32; }   // exact 2x2: INTER_AREA as the same fixed-point blend
        else {
          int sx; float fr;
          cv_tap(x, h->xscale, sx, fr);
          if (sx < 0) { sx = 0; fr = 0.f; }
          if (sx >= sw - 1) { sx = sw - 1; fr = 0.f; }
          a0 = coef(__fsub_rn(1.0f, fr));
          a1 = coef(fr);
          xo = 3 * sx;
        }
        xo_s = ring_s + (uint32_t)xo;
        a01 = (uint32_t)a0 | ((uint32_t)a1 << 16);
      }
      if (STATS) {
        const int b = h->clip;
        if (b != acc_b) {
          if (acc_b >= 0 && p.stats) {
#pragma unroll
            for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (lane == 0 && acc) atomicAdd(&p.stats[acc_b], acc);
          }
          acc = 0; acc_b = b;
        }
      }
      if (st == 0) obase = h->obase;
      if (flags & 4) item_bad = true;   // nothing was fetched for this item (contract violation, flagged by the producer)
      else {
        const int oy0 = h0.z, rlo = h1.w, own_hi = h1.y;
        unsigned acc32 = 0;
        for (int r = g; r < p.rps; r += kGroups) {
          const RowDesc rd = h->rows[r];
          const int oyr = oy0 + r;
          unsigned v[3];
          cv2_row_lean(rd, xo_s, a01, selA, selB, v);
          // checksum of the resized rows this band OWNS (halo rows belong to the neighbour band)
          if (STATS && oyr < S && oyr < own_hi) acc32 += v[0] + v[1] + v[2];
          // Pillow horizontal pass of the row (224 -> cropped 224 of 256), rounded to uint8; one word per pixel
          uint32_t* rb = rbuf + parity * S;
          parity ^= 1;
          rb[x] = v[0] | (v[1] << 8) | (v[2] << 16);
          group_bar(g);
          inter[(oyr - rlo) * S + x] = pil16(rb[hx.x], rb[hx.x + 1], hm0, hm1);
        }
        if (STATS) acc += acc32;
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);
      if (tr) p.trace[cn * 4 + 3] = globaltimer_ns();
    }
    // the band's resized rows went through the horizontal pass: Pillow vertical pass + normalisation + store
    consumers_bar(kConsumers);
    if (!item_bad) {
      const int band = item % p.bands, ylo = band * p.rows_per_band, rlo = vt[ylo].x;
      auto vpass = [&](auto ODTV) {
        constexpr int kOdt = decltype(ODTV)::value;   // 0 f32, 1 bf16 table, 2 u8, 3 bf16 affine
#pragma unroll 2
        for (int y = ylo + g; y < ylo + p.rows_per_band; y += kGroups) {
          const size_t eoff = (size_t)obase + (size_t)(rpart[y] + xoff);
          const int4 ty = vt[y];
          const uint32_t* tp = inter + (ty.x - rlo) * S + x;
          const uint32_t pk = pil16(tp[0], tp[S], (uint32_t)ty.y >> 18, (uint32_t)ty.z >> 18);
          const int v[3] = {(int)(pk & 0xffu), (int)((pk >> 8) & 0xffu), (int)(pk >> 16)};
          if (kOdt == 0) {
            float* o = reinterpret_cast<float*>(p.out) + eoff;
#pragma unroll
            for (int c = 0; c < 3; ++c) __stcs(o + (size_t)c * cs, lut[c * 256 + v[c]]);
          } else if (kOdt == 2) {
            uint8_t* o = reinterpret_cast<uint8_t*>(p.out) + eoff;
#pragma unroll
            for (int c = 0; c < 3; ++c) o[(size_t)c * cs] = (uint8_t)v[c];
          } else {
            __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + eoff;
#pragma unroll
            for (int c = 0; c < 3; ++c)
              o[(size_t)c * cs] = kOdt == 3 ? __float2bfloat16_rn(__fmaf_rn((float)v[c], fa[c], fb[c])) : lut16[c * 256 + v[c]];
          }
        }
      };
      using std::integral_constant;
      if (p.out_dtype == LVC_DT_F32) vpass(integral_constant<int, 0>{});
      else if (p.out_dtype == LVC_DT_U8) vpass(integral_constant<int, 2>{});
      else if (use_fma) vpass(integral_constant<int, 3>{});
      else vpass(integral_constant<int, 1>{});
    }
    consumers_bar(kConsumers);   // the next band's horizontal pass overwrites the intermediate buffer
  }
  if (STATS && p.stats && acc_b >= 0) {
#pragma unroll
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0 && acc) atomicAdd(&p.stats[acc_b], acc);
  }
  if (kTraceBuild && p.trace && ct == 0) p.trace[kTraceStages * 4 + blockIdx.x * 2 + 1] = globaltimer_ns();
}

// ODT: compile-time output type of the lean cv2 consumer (LVC_DT_*), or -1 = the general consumer (type from the descriptor)
template <int MODE, int ODT, bool STATS, int NPW = kProducers>   // NPW: producer warps the CTA is built with
// (The variant that shares its SM declares a bound of 736 threads -- it is launched with 512 -- to cap it at 88 registers:
// registers are allocated per FOUR warps, so the neighbour's 6 warps cost 8 x 32 x 80 = 20 480 and 16 x 32 x 88 = 45 056
// is what is left of the 65 536.)
__global__ void __launch_bounds__(NPW == kProducers ? threads_of(MODE, NPW) : 736, (MODE == MODE_VIVIT ? 1 : LVC_CLIP_MINBLOCKS)) lvc_clip_kernel(const __grid_constant__ ClipParams p) {
  constexpr int kGroups = groups_of(MODE), kConsumers = kGroups * S, kThreads = threads_of(MODE, NPW);
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty = full + kMaxStages;
  StageHdr* hdrs = reinterpret_cast<StageHdr*>(smem + kHdrOff);
  float* lut = reinterpret_cast<float*>(smem + kLutOff);
  __nv_bfloat16* lut16 = reinterpret_cast<__nv_bfloat16*>(smem + kLut16Off);
  uint8_t* ring = smem + p.ring_off;

  // a kernel launched behind this one with programmatic stream serialization (the patch-embedding GEMM) may be
  // scheduled as SMs free up; it waits for this grid to complete before it touches global data
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int tid = threadIdx.x;
  if (kTraceBuild && p.trace && tid == 0) p.trace[kTraceStages * 4 + blockIdx.x * 2] = globaltimer_ns();   // CTA start
  if (p.tile_stamps && tid == 0 && blockIdx.x == 0) p.tile_stamps[0] = globaltimer_ns();
  const int warp = tid >> 5, lane = tid & 31;
  const int stages_per_item = MODE == MODE_VIVIT ? p.v_stages : p.rows_per_band / p.rps;

  if (tid == 0) {
    for (int s = 0; s < kMaxStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kConsumers / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (MODE == MODE_VIVIT)   // (the producers read this table too: item geometry of the bands)
    for (int i = tid; i < S; i += kThreads) {
      const int4 t = pil_tap(i);
      reinterpret_cast<int4*>(smem + kVivitTapOff)[i] = t;
      // the lean consumer's 4-bit form of the Pillow pass (pil16) needs two taps, multiples of 2^18, summing to 2^22
      if (((t.y | t.z) & 0x3ffff) || t.w || t.y + t.z != (1 << 22)) atomicOr(p.err, (unsigned)LVC_DEVERR_INTERNAL);
      // row term of the output offset, once per kernel instead of once per thread and output row (warp-uniform arithmetic)
      const unsigned rowmul = p.layout == LVC_LAYOUT_TOKENS ? (unsigned)(S / p.kw) * (unsigned)(3 * p.kt * p.kh * p.kw)
                                                             : (p.layout == LVC_LAYOUT_BTHWC ? S * 3 : S);
      reinterpret_cast<unsigned*>(smem + kVivitRowPartOff)[i] = row_part(p, rowmul, i);
    }
  __syncthreads();   // barriers (and the ViViT tap table) are ready: the producer warps start fetching NOW
  // The value tables are the consumers' business: they build them while the first stages are in flight (the set-up used to
  // sit in front of the producers' first copy: ~1.5 us of every launch) and meet on the consumers' own barrier.
  const bool is_consumer = warp >= NPW;
  const int ct0 = tid - 32 * NPW;
  if (is_consumer) {
    if (p.lut)
      for (int i = ct0; i < 3 * 256; i += kConsumers) {
        const float w = p.lut[i];
        lut[i] = w;
        lut16[i] = __float2bfloat16_rn(w);
      }
    if (MODE == MODE_CV2 && ODT >= 0 && ODT != LVC_DT_U8 && !STATS && p.lut) {
      // table of the lean consumer: entry t = LUT[(t + 2) >> 2], t = unrounded sum of the two vertical terms (<= 1020)
      for (int i = ct0; i < 3 * 1024; i += kConsumers) {
        const float w = p.lut[(i >> 10) * 256 + min(((i & 1023) + 2) >> 2, 255)];
        if (ODT == LVC_DT_F32) reinterpret_cast<float*>(smem + p.lut4_off)[i] = w;
        else reinterpret_cast<__nv_bfloat16*>(smem + p.lut4_off)[i] = __float2bfloat16_rn(w);
      }
    }
    consumers_bar(kConsumers);
  }
  // bf16 outputs: the look-up costs three bank-conflicted shared loads per sample (random 8-bit indices: ~3.5 wavefronts
  // each -- ncu: 41 % of this kernel's shared-memory wavefronts on the small-source configs, whose consumers are bound by
  // the LSU data pipe).  The reference's normalisation is affine in the pixel value, so y = fma(v, a, b) rounded to bf16
  // usually equals the rounded table entry for all 3 x 256 values; every consumer WARP verifies that here (all 768 values,
  // 24 per lane: the same answer in every warp, no vote across warps needed) and only then replaces the look-up by two
  // FMA-pipe instructions (bit-identical by construction; any table that fails keeps the look-up).
  float fa[3] = {0.f, 0.f, 0.f}, fb[3] = {0.f, 0.f, 0.f};
  bool use_fma = false;
  if (is_consumer && MODE != MODE_PTV && p.out_dtype == LVC_DT_BF16 && p.lut && !p.no_fma) {
    bool ok = true;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      fb[c] = lut[c * 256];
      fa[c] = __fdiv_rn(__fsub_rn(lut[c * 256 + 255], lut[c * 256]), 255.0f);
    }
    for (int i = lane; i < 3 * 256; i += 32) {
      const int c = i >> 8;
      const float a = c == 0 ? fa[0] : (c == 1 ? fa[1] : fa[2]), b = c == 0 ? fb[0] : (c == 1 ? fb[1] : fb[2]);
      const __nv_bfloat16 y = __float2bfloat16_rn(__fmaf_rn((float)(i & 255), a, b));
      ok = ok && (__bfloat16_as_ushort(y) == __bfloat16_as_ushort(lut16[i]));
    }
    use_fma = __all_sync(0xffffffffu, ok) != 0;
  }

  if (warp < NPW) {
    // ================= producer warps: lane j fetches source row (j&1 ? y1 : y0) of output row j>>1
    // No more active producers than ring slots: a warp that finished stage n has seen stage n - np - stages consumed;
    // its parity wait on the slot of stage n is unambiguous only if stage n - 2*stages is consumed too (np <= stages).
    const int np = p.producers;
    const bool env_merge = p.merge_rows != 0;
    if (warp >= np) return;
    int turn = 0;          // stage counter modulo np: this warp acts on the stages with turn == warp
    const uint64_t policy = policy_evict_first();
    const int nrows = 2 * p.rps;
    const unsigned lanes = nrows >= 32 ? 0xffffffffu : ((1u << nrows) - 1u);
    // Byte-granular ring: a stage takes 2*rps row slots of THIS item's row size (an ROI row is a third of a frame
    // row: the same shared memory then holds 2-3x as many stages in flight), placed by a bump allocator that wraps
    // when the stage does not fit before the end.  Every producer warp runs the allocator for every stage (sizes
    // depend only on the item), so all warps agree on the offsets without talking to each other.  Stage n uses
    // barrier slot n % 16; before filling it the owner waits for the YOUNGEST earlier stage whose bytes overlap the
    // new range, or stage n - 16 (the slot's previous user) -- consumers release stages in order, so that one wait
    // covers every older stage.  Lane l < 16 remembers the byte range of the latest stage with n % 16 == l.
    int n = 0, head = 0;   // stage sequence number, allocator position
    int hs = 0, he = 0;    // lane's remembered byte range
    unsigned long long fetched = 0;   // bytes this warp asked the copy engine for (lane 0)
    for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
      const int f = item / p.bands, band = item - f * p.bands;
      const int b = f / p.T;
      const FrameOut fo = frame_out(p, f, b, f - b * p.T);
      const lvc_video v = p.videos[b];
      int idx = p.indices[f];
      idx = min(max(idx, 0), v.n_frames - 1);
      int ox = 0, oy = 0, sw = v.width, sh = v.height;  // source window: origin and size
      int nh = 0, nw = 0, top = 0, left = 0;
      float scale_h = 0.f;
      int flip = 0;
      if (MODE == MODE_PTV) {
        if (p.aug) {   // train chain: RandomShortSideScale size, RandomCrop origin, RandomHorizontalFlip (per clip)
          const int4 a = reinterpret_cast<const int4*>(p.aug)[b];
          ptv_geom(v.height, v.width, max(a.x, S), nh, nw);
          top = min(max(a.y, 0), nh - S);
          left = min(max(a.z, 0), nw - S);
          flip = a.w != 0;
        } else {
          ptv_geom(v.height, v.width, 256, nh, nw);
          top = center_crop_origin(nh);
          left = center_crop_origin(nw);
        }
        scale_h = __fdiv_rn((float)v.height, (float)nh);
        const float scale_w = __fdiv_rn((float)v.width, (float)nw);
        int xa, xb, xc, xd; float l0, l1;
        torch_tap(left, scale_w, v.width, xa, xb, l0, l1);          // first and last tapped columns
        torch_tap(left + S - 1, scale_w, v.width, xc, xd, l0, l1);
        ox = xa; sw = xd - xa + 1;
      } else if (p.boxes) {
        const int4 bx = reinterpret_cast<const int4*>(p.boxes)[f];
        if (bx.x >= 0) {  // numpy half-open slice of the box clipped to the frame, at least 1x1
          ox = min(max(bx.x, 0), v.width - 1);
          oy = min(max(bx.y, 0), v.height - 1);
          sw = min(max(bx.z, ox + 1), v.width) - ox;
          sh = min(max(bx.w, oy + 1), v.height) - oy;
        }
      }
      // VIVIT bands: output rows [band*R, band*R+R) of the cropped Pillow image need the resized rows
      // [rlo, rhi] (Pillow's vertical window, 2-3 taps); the band OWNS (checksums) rows [rlo, next band's rlo)
      int rlo = 0, own_hi = S;
      if (MODE == MODE_VIVIT) {
        const int4* vt = reinterpret_cast<const int4*>(smem + kVivitTapOff);
        rlo = vt[band * p.rows_per_band].x;
        if (band + 1 < p.bands) own_hi = vt[(band + 1) * p.rows_per_band].x;
        else own_hi = vt[S - 1].x + (vt[S - 1].w ? 3 : 2);   // one past the last resized row the crop reads
      }
      const bool area = MODE != MODE_PTV && (sw == 2 * S) && (sh == 2 * S);
      const double scale_y = cv_scale(sh);
      const double scale_x = MODE == MODE_PTV ? 0.0 : cv_scale(sw);   // consumers' column taps (header)
      const uint64_t row_bytes = (uint64_t)v.width * 3u;
      const bool full_width = MODE != MODE_PTV && ox == 0 && sw == v.width && env_merge;
      const uint64_t frame_base = v.base + (uint64_t)idx * (uint64_t)v.height * row_bytes;
      const uint64_t buf_end = v.base + (uint64_t)v.nbytes;
      // row slot of this item: the window's bytes + alignment slack, rounded to 16 (+16 to stagger banks)
      const int item_pitch = ((sw * 3 + 15 + 15) & ~15) + 16;
      const int stage_bytes = nrows * item_pitch;
      // a window that does not fit the ring even alone (a video much wider than desc.max_width, which sized the ring):
      // fetch nothing, tell the consumers to skip the rows and raise the device error word (lvc_device_status)
      const bool bad = stage_bytes > p.ring_bytes;
      if (bad && warp == 0 && lane == 0) atomicOr(p.err, (unsigned)LVC_DEVERR_FRAME_TOO_WIDE);
      for (int st = 0; st < stages_per_item; ++st, ++n) {
        const int slot = n & (kMaxStages - 1);
        const int sbytes = bad ? 0 : stage_bytes;
        int start = head;
        if (start + sbytes > p.ring_bytes) start = 0;
        head = start + sbytes;
        const bool mine = (turn == warp);
        turn = (turn + 1 == np) ? 0 : turn + 1;
        if (mine) {
          int m = n - ((n - lane) & (kMaxStages - 1));   // latest stage < n with m % 16 == lane ...
          if (m == n) m -= kMaxStages;                   // ... which for lane == slot is the slot's previous user
          const bool dep_l = lane < kMaxStages && m >= 0 && (lane == slot || (hs < start + sbytes && start < he));
          const int dep = __reduce_max_sync(0xffffffffu, dep_l ? m : -1);
          if (dep >= 0 && lane == 0) mbar_wait(&empty[dep & (kMaxStages - 1)], (uint32_t)(dep >> 4) & 1u);
          __syncwarp();
        }
        if (lane == slot) { hs = start; he = start + sbytes; }
        if (!mine) continue;
        const int oy_first = MODE == MODE_VIVIT ? rlo + st * p.rps : band * p.rows_per_band + st * p.rps;
        uint32_t my_bytes = 0, bulk_bytes = 0;
        int my_off = 0, cb = 0;
        unsigned roff = 0;
        uint64_t a = 0;
        bool leader = false;
        uint8_t* stage_base = ring + start;
        if (lane < nrows) {
          const int r = lane >> 1, oyr = min(oy_first + r, S - 1);   // VIVIT halo stages may run past the last row
          if (MODE != MODE_VIVIT) roff = row_part(p, fo.rowmul, oyr);
          int ysrc;
          if (MODE == MODE_PTV) {
            int y0, y1; float l0, l1;
            torch_tap(oyr + top, scale_h, v.height, y0, y1, l0, l1);
            ysrc = (lane & 1) ? y1 : y0;
            cb = __float_as_int((lane & 1) ? l1 : l0);
          } else if (area) {
            ysrc = 2 * oyr + (lane & 1);
            cb = (int)0x80000000u;   // with a0 = a1 = 32 the fixed-point blend below IS the 2x2 average (see the consumer)
          } else {
            int sy; float fr;
            if (p.f32src) cv_tap64(oyr, scale_y, sy, fr); else cv_tap(oyr, scale_y, sy, fr);
            ysrc = min(max(sy + (lane & 1), 0), sh - 1);
            if (p.f32src) cb = __float_as_int((lane & 1) ? fr : __fsub_rn(1.0f, fr));
            else cb = ((lane & 1) ? coef(fr) : coef(__fsub_rn(1.0f, fr))) << 16;   // pre-shifted: the blend is a multiply-high
          }
          const uint64_t addr = frame_base + (uint64_t)(oy + ysrc) * row_bytes + (uint64_t)ox * 3u;
          a = addr & ~(uint64_t)15;
          const uint32_t shift = (uint32_t)(addr - a);
          my_bytes = (shift + (uint32_t)sw * 3u + 15u) & ~15u;   // <= item_pitch - 16 by construction
          const unsigned same = __match_any_sync(lanes, ysrc);
          const int lead = __ffs(same) - 1;
          leader = (lead == lane) && !bad;
          my_off = start + lead * item_pitch + (int)shift;
          bulk_bytes = bad ? 0u : my_bytes;
          if (a + my_bytes > buf_end) bulk_bytes = (uint32_t)((buf_end - a) & ~(uint64_t)15);
          if (!leader) bulk_bytes = 0;
          // Full-width rows y and y+1 are contiguous in memory: when the two taps of an output row are such a pair
          // (and neither row serves another lane) the even lane fetches both with ONE copy into its two slots --
          // half as many UBLKCP issues per stage, twice the request size on the bus.
          const unsigned peer_same = __shfl_xor_sync(lanes, same, 1);
          const int y_peer = __shfl_xor_sync(lanes, ysrc, 1);
          const uint32_t merged = (shift + 2u * (uint32_t)row_bytes + 15u) & ~15u;
          int pair = !(lane & 1) && full_width && !bad && same == (1u << lane) && peer_same == (2u << lane) && y_peer == ysrc + 1 &&
                     a + merged <= buf_end;
          pair |= __shfl_xor_sync(lanes, pair, 1);
          const int off_even = __shfl_xor_sync(lanes, my_off, 1);
          if (pair) {
            if (lane & 1) { leader = false; bulk_bytes = 0; my_off = off_even + (int)row_bytes; }
            else bulk_bytes = merged;
          }
        }
        uint32_t total = bulk_bytes;
#pragma unroll
        for (int o = 16; o; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
        const int off_next = __shfl_down_sync(0xffffffffu, my_off, 1);
        const int cb_next = __shfl_down_sync(0xffffffffu, cb, 1);
        StageHdr* h = &hdrs[slot];
        if (lane < nrows && !(lane & 1)) {
          RowDesc rd;
          rd.off0 = my_off; rd.off1 = off_next; rd.b0 = cb; rd.b1 = cb_next;
          h->rows[lane >> 1] = rd;
          h->roff[lane >> 1] = roff;
        }
        if (lane == 0) {
          h->frame = f; h->sw = (MODE == MODE_PTV) ? v.width : sw; h->oy0 = oy_first;
          h->flags = (area ? 1 : 0) | (flip << 1) | (bad ? 4 : 0);
          h->ox = ox; h->nw = (MODE == MODE_VIVIT) ? own_hi : nw; h->left = left; h->rlo = rlo;
          h->obase = fo.base; h->cstride = fo.cstride; h->rowmul = fo.rowmul; h->clip = b; h->xscale = scale_x;
        }
        if (leader && bulk_bytes < my_bytes) {
          // last bytes of the buffer: not a whole 16-byte granule, copy them by hand
          uint8_t* slot_ptr = stage_base + (size_t)lane * item_pitch;
          for (uint64_t q = a + bulk_bytes; q < buf_end && q < a + my_bytes; ++q)
            slot_ptr[q - a] = *reinterpret_cast<const uint8_t*>(q);
        }
        __syncwarp();
        if (lane == 0) { mbar_arrive_expect_tx(&full[slot], total); fetched += total; }
        if (kTraceBuild && p.trace && blockIdx.x == 0 && lane == 0 && n < kTraceStages) p.trace[n * 4 + 0] = globaltimer_ns();
        __syncwarp();
        if (leader && bulk_bytes)
          bulk_g2s(stage_base + (size_t)lane * item_pitch, reinterpret_cast<const void*>(a), bulk_bytes,
                   &full[slot], policy);
      }
    }
    if (lane == 0 && fetched) atomicAdd(p.fetched, fetched);
    return;
  }

  // ================= consumer warps: thread -> output column x, group -> rows g, g+G, ...
  const int ct = tid - 32 * NPW;
  if constexpr (MODE == MODE_CV2 && ODT >= 0) {
    lean_cv2_consumer<ODT, STATS>(p, smem, full, empty, hdrs, ct, lane, use_fma, fa, fb);
    return;
  }
  if constexpr (MODE == MODE_VIVIT) {
    if (!p.f32src) {   // float frames keep the general consumer below
      lean_vivit_consumer<STATS>(p, smem, full, empty, hdrs, ct, lane, use_fma, fa, fb);
      return;
    }
  }
  const int g = ct / S;
  const int x = ct - g * S;
  int cached_sw = -1, cached_flags = -1, cached_ox = -1;
  unsigned long long obase = 0, cstride = 0;   // output addressing of the current item (from the stage header)
  unsigned rowmul = 0;
  int xo = 0, d1 = 0, a0 = 0, a1 = 0;
  float lx0 = 0.f, lx1 = 0.f;
  unsigned long long acc = 0;  // checksum of resized bytes for the clip being processed
  const bool want_stats = p.stats != nullptr;
  int acc_b = -1;
  int slot = 0, cn = 0;
  uint32_t phase = 0;
  const unsigned xoff = col_offset(p, x);
  // MODE_VIVIT state: Pillow horizontal taps of this thread's column, row-buffer parity
  int4 hx = make_int4(0, 0, 0, 0);
  int parity = 0;
  uint8_t* inter = smem + kVivitInterOff;
  if (MODE == MODE_VIVIT) hx = reinterpret_cast<const int4*>(smem + kVivitTapOff)[x];

  // LUT + store of one output sample per channel (element offset eoff of channel 0, channel stride cstride)
  auto store_px = [&](auto ODT_, const size_t eoff, const int (&v)[3]) {
    constexpr int kOdt = decltype(ODT_)::value;   // compile-time output dtype, or -1 = look at the descriptor
    const int odt = kOdt >= 0 ? kOdt : p.out_dtype;
    if (odt == LVC_DT_F32) {
      float* o = reinterpret_cast<float*>(p.out) + eoff;
#pragma unroll
      for (int c = 0; c < 3; ++c) __stcs(o + c * cstride, lut[c * 256 + v[c]]);
    } else if (odt == LVC_DT_BF16) {
      __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + eoff;
      if (use_fma) {
#pragma unroll
        for (int c = 0; c < 3; ++c) o[c * cstride] = __float2bfloat16_rn(__fmaf_rn((float)v[c], fa[c], fb[c]));
      } else {
#pragma unroll
        for (int c = 0; c < 3; ++c) o[c * cstride] = lut16[c * 256 + v[c]];
      }
    } else {
      uint8_t* o = reinterpret_cast<uint8_t*>(p.out) + eoff;
#pragma unroll
      for (int c = 0; c < 3; ++c) o[c * cstride] = (uint8_t)v[c];
    }
  };

  for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
    bool item_bad = false;
    for (int st = 0; st < stages_per_item; ++st, ring_next(slot, phase, kMaxStages), ++cn) {
      const bool tr = kTraceBuild && p.trace && blockIdx.x == 0 && ct == 0 && cn < kTraceStages;
      if (tr) p.trace[cn * 4 + 1] = globaltimer_ns();   // consumer starts waiting
      mbar_wait(&full[slot], phase);
      if (tr) p.trace[cn * 4 + 2] = globaltimer_ns();   // stage landed (or had landed)
      const StageHdr* h = &hdrs[slot];
      const int sw = h->sw, flags = h->flags, oy0 = h->oy0;
      obase = h->obase; cstride = h->cstride; rowmul = h->rowmul;
      if (MODE == MODE_PTV) {
        const int hox = h->ox;
        const int geo = h->nw | (h->left << 12) | ((flags & 2) << 28);   // nw, left < 4096
        if (sw != cached_sw || hox != cached_ox || geo != cached_flags) {
          cached_sw = sw; cached_ox = hox; cached_flags = geo;
          int x0, x1;
          const int xs = (flags & 2) ? S - 1 - x : x;   // hflip: output column x shows crop column 223 - x
          torch_tap(xs + h->left, __fdiv_rn((float)sw, (float)h->nw), sw, x0, x1, lx0, lx1);
          xo = 3 * (x0 - hox);
          d1 = 3 * (x1 - x0);
        }
      } else if (sw != cached_sw || flags != cached_flags) {
        cached_sw = sw; cached_flags = flags;
        // exact 2x2 (INTER_AREA): a0 = a1 = 32 and b0 = b1 = 2^31 turn the fixed-point blend into
        // (p00 + p01 + p10 + p11 + 2) >> 2: s >> 4 = 2 (p0 + p1), umulhi(2^31, 2 t) = t
        if (flags & 1) { xo = 6 * x; d1 = 3; a0 = 32; a1 = 32; }
        else {
          int sx; float fr;
          if (p.f32src) cv_tap64(x, h->xscale, sx, fr); else cv_tap(x, h->xscale, sx, fr);
          if (sx < 0) { sx = 0; fr = 0.f; }
          if (sx >= sw - 1) { sx = sw - 1; fr = 0.f; }
          a0 = coef(__fsub_rn(1.0f, fr));
          a1 = coef(fr);
          lx0 = __fsub_rn(1.0f, fr); lx1 = fr;
          xo = 3 * sx;
          d1 = (sx + 1 <= sw - 1) ? 3 : 0;
        }
      }
      const int b = h->clip;
      if (MODE != MODE_PTV && want_stats && b != acc_b) {
        if (acc_b >= 0) {
#pragma unroll
          for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
          if (lane == 0 && acc) atomicAdd(&p.stats[acc_b], acc);
        }
        acc = 0; acc_b = b;
      }
      unsigned acc32 = 0;   // this stage's share of the checksum (<= 4 rows x 765)
      // F32SRC / ODT are compile-time copies of p.f32src / p.out_dtype: the row body has no branches, so the two
      // rows of a trip are scheduled as one block (their shared-memory loads overlap)
      auto do_row = [&](auto F32SRC, auto ODT_, const int r) {
        constexpr bool kF32Src = decltype(F32SRC)::value;
        const RowDesc rd = h->rows[r];
        const uint8_t* p0 = ring + rd.off0 + xo;
        const uint8_t* p1 = ring + rd.off1 + xo;
        const int oyr = oy0 + r;
        const size_t eoff = (size_t)obase + (size_t)(h->roff[r] + xoff);  // element offset of channel 0
        if (MODE == MODE_PTV) {
          // torch CPU blend order: r = fma(p_x0, lx0, p_x1*lx1) per row, out = fma(r_y0, ly0, r_y1*ly1)
          const float ly0 = __int_as_float(rd.b0), ly1 = __int_as_float(rd.b1);
          float* o = reinterpret_cast<float*>(p.out) + eoff;
          __nv_bfloat16* ob = reinterpret_cast<__nv_bfloat16*>(p.out) + eoff;
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const int sc = p.bgr ? 2 - c : c;
            const float r0 = __fmaf_rn((float)p0[sc], lx0, __fmul_rn((float)p0[sc + d1], lx1));
            const float r1 = __fmaf_rn((float)p1[sc], lx0, __fmul_rn((float)p1[sc + d1], lx1));
            const float val = __fmaf_rn(r0, ly0, __fmul_rn(r1, ly1));
            const float y = __fdiv_rn(__fsub_rn(val, p.mean[c]), p.std[c]);
            if (p.out_dtype == LVC_DT_F32) __stcs(o + c * cstride, y);
            else ob[c * cstride] = __float2bfloat16_rn(y);
          }
          return;
        }
        int v[3];
        {
          // 6 source bytes (2 pixels x RGB) per row come from three aligned 32-bit shared loads and a
          // funnel shift instead of six byte loads (halves the bank-conflicted wavefronts); the
          // horizontal pass a0*p0 + a1*p1 is one PRMT + one DP2A per channel.  At a clamped right
          // edge a1 == 0, so the bytes read at +3 (still inside the slot's slack) do not matter.
          const uint32_t o0 = (uint32_t)(rd.off0 + xo), o1 = (uint32_t)(rd.off1 + xo);
          const uint32_t* w0 = reinterpret_cast<const uint32_t*>(ring + (o0 & ~3u));
          const uint32_t* w1 = reinterpret_cast<const uint32_t*>(ring + (o1 & ~3u));
          const uint32_t a = w0[0], bq = w0[1], cq = w0[2], d = w1[0], e = w1[1], fq = w1[2];
          const uint32_t sh0 = (o0 & 3u) * 8u, sh1 = (o1 & 3u) * 8u;
          const uint32_t lo0 = __funnelshift_r(a, bq, sh0), hi0 = __funnelshift_r(bq, cq, sh0);
          const uint32_t lo1 = __funnelshift_r(d, e, sh1), hi1 = __funnelshift_r(e, fq, sh1);
          const uint32_t a01 = (uint32_t)a0 | ((uint32_t)a1 << 16);
          // byte pairs (c of pixel 0, c of pixel 1) of the 8 source bytes; a BGR source swaps ends
          const uint32_t sel[3] = {p.bgr ? 0x4452u : 0x4430u, 0x4441u, p.bgr ? 0x4430u : 0x4452u};
          if (kF32Src) {
            // float frames: two float products and an add per pass (no FMA contraction), truncation to uint8;
            // the exact 2x2 case is INTER_AREA = sum * 0.25f, truncated
            const float by0 = __int_as_float(rd.b0), by1 = __int_as_float(rd.b1);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              const uint32_t w0c = __byte_perm(lo0, hi0, sel[c]), w1c = __byte_perm(lo1, hi1, sel[c]);
              if (flags & 1) {
                v[c] = (int)((w0c & 0xffu) + ((w0c >> 8) & 0xffu) + (w1c & 0xffu) + ((w1c >> 8) & 0xffu)) >> 2;
              } else {
                const float r0 = __fadd_rn(__fmul_rn((float)(w0c & 0xffu), lx0), __fmul_rn((float)((w0c >> 8) & 0xffu), lx1));
                const float r1 = __fadd_rn(__fmul_rn((float)(w1c & 0xffu), lx0), __fmul_rn((float)((w1c >> 8) & 0xffu), lx1));
                v[c] = (int)__fadd_rn(__fmul_rn(r0, by0), __fmul_rn(r1, by1));
              }
            }
          } else {
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              const int s0 = (int)__dp2a_lo(a01, __byte_perm(lo0, hi0, sel[c]), 0u);
              const int s1 = (int)__dp2a_lo(a01, __byte_perm(lo1, hi1, sel[c]), 0u);
              // cv2's vertical pass ((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16): the coefficients arrive as b << 16,
              // so each term is one multiply-high (all operands are non-negative)
              v[c] = (int)((__umulhi((unsigned)rd.b0, (unsigned)s0 >> 4) + __umulhi((unsigned)rd.b1, (unsigned)s1 >> 4) + 2u) >> 2);
            }
          }
        }
        // checksum share (dropped after the stage loop when no stats buffer was passed); bands: owned rows only
        if (MODE != MODE_VIVIT) acc32 += (unsigned)(v[0] + v[1] + v[2]);   // branch-free body: always counted
        else if (want_stats && oyr < S && oyr < h->nw) acc32 += (unsigned)(v[0] + v[1] + v[2]);
        if (MODE == MODE_VIVIT) {
          // Pillow horizontal pass of this row (224 -> cropped 224 of 256), rounded to uint8
          // (one word per pixel: a store and up to three loads per thread instead of 3 + 9 byte accesses)
          uint32_t* rb = reinterpret_cast<uint32_t*>(smem + kVivitRowBufOff + (g * 2 + parity) * kRowBytes);
          parity ^= 1;
          rb[x] = (uint32_t)v[0] | ((uint32_t)v[1] << 8) | ((uint32_t)v[2] << 16);
          group_bar(g);
          const uint32_t w0 = rb[hx.x], w1 = rb[hx.x + 1], w2 = hx.w ? rb[hx.x + 2] : 0u;
          uint32_t packed = 0;
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const int ss = (1 << 21) + (int)__byte_perm(w0, 0u, 0x4440u + c) * hx.y + (int)__byte_perm(w1, 0u, 0x4440u + c) * hx.z +
                           (int)__byte_perm(w2, 0u, 0x4440u + c) * hx.w;
            packed |= (uint32_t)clip8(ss >> 22) << (8 * c);
          }
          reinterpret_cast<uint32_t*>(inter)[(oyr - h->rlo) * S + x] = packed;
          return;
        }
        store_px(ODT_, eoff, v);
      };
      // two rows per trip (their shared-memory loads overlap) whenever the group has an even number of rows
      auto do_rows = [&](auto F32SRC, auto ODT_) {
        if (MODE != MODE_VIVIT && (p.rps & (2 * kGroups - 1)) == 0) {
          for (int r = g; r < p.rps; r += 2 * kGroups) { do_row(F32SRC, ODT_, r); do_row(F32SRC, ODT_, r + kGroups); }
        } else {
          for (int r = g; r < p.rps; r += kGroups) do_row(F32SRC, ODT_, r);
        }
      };
      using std::integral_constant;
      if (flags & 4) item_bad = true;                                                          // nothing was fetched
      else if (MODE == MODE_PTV) do_rows(std::false_type{}, integral_constant<int, -1>{});        // stores by itself
      else if (p.f32src) do_rows(std::true_type{}, integral_constant<int, -1>{});             // float frames: rare
      else if (MODE == MODE_VIVIT) do_rows(std::false_type{}, integral_constant<int, -1>{});  // rows go to the intermediate buffer
      else if (p.out_dtype == LVC_DT_BF16) do_rows(std::false_type{}, integral_constant<int, LVC_DT_BF16>{});
      else if (p.out_dtype == LVC_DT_F32) do_rows(std::false_type{}, integral_constant<int, LVC_DT_F32>{});
      else do_rows(std::false_type{}, integral_constant<int, LVC_DT_U8>{});
      if (want_stats) acc += acc32;
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);
      if (tr) p.trace[cn * 4 + 3] = globaltimer_ns();   // this warp is done with the stage
    }
    if (MODE == MODE_VIVIT) {
      // the whole frame went through the horizontal pass: Pillow vertical pass + LUT + store
      consumers_bar(kConsumers);
      const int4* vt = reinterpret_cast<const int4*>(smem + kVivitTapOff);
      const int band = item % p.bands, ylo = band * p.rows_per_band, rlo = vt[ylo].x;
      for (int y = ylo + g; y < ylo + p.rows_per_band && !item_bad; y += kGroups) {
        const size_t eoff = (size_t)obase + (size_t)(row_part(p, rowmul, y) + xoff);
        const int4 ty = vt[y];
        const uint32_t* tp = reinterpret_cast<const uint32_t*>(inter) + (ty.x - rlo) * S + x;
        const uint32_t w0 = tp[0], w1 = tp[S], w2 = ty.w ? tp[2 * S] : 0u;
        int v[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const int ss = (1 << 21) + (int)__byte_perm(w0, 0u, 0x4440u + c) * ty.y + (int)__byte_perm(w1, 0u, 0x4440u + c) * ty.z +
                         (int)__byte_perm(w2, 0u, 0x4440u + c) * ty.w;
          v[c] = clip8(ss >> 22);
        }
        store_px(std::integral_constant<int, -1>{}, eoff, v);
      }
      consumers_bar(kConsumers);  // the next frame's horizontal pass overwrites the intermediate
    }
  }
  if (MODE != MODE_PTV && p.stats && acc_b >= 0) {
#pragma unroll
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0 && acc) atomicAdd(&p.stats[acc_b], acc);
  }
}

}  // namespace

// ------------------------------------------------------------------------------------------
namespace {

int env_int(const char* name, int dflt) { return lvc_knob_int(name, dflt); }   // gated profiling knobs (lvc_internal.h)

struct DeviceInfo {
  int sms = 0, max_smem = 0;
  cudaError_t err = cudaSuccess;
};
const DeviceInfo& device_info() {
  static thread_local DeviceInfo di;
  static thread_local int cached_dev = -1;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e == cudaSuccess && dev != cached_dev) {
    e = cudaDeviceGetAttribute(&di.sms, cudaDevAttrMultiProcessorCount, dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&di.max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (e == cudaSuccess) cached_dev = dev;
  }
  di.err = e;
  return di;
}

// Contract violations found on the device (sticky until read): lvc_device_status()
__device__ unsigned g_dev_error;
// Bytes the clip kernels' producers fetched from the source buffers since the last read: lvc_device_bytes_fetched()
__device__ unsigned long long g_bytes_fetched;

template <int MODE, int ODT = -1, bool STATS = true, int NPW = kProducers>
int launch(const ClipParams& p, int grid, size_t smem, cudaStream_t st) {
  LVC_CUDA_TRY(cudaFuncSetAttribute(lvc_clip_kernel<MODE, ODT, STATS, NPW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (NPW != kProducers)   // the variant that shares its SM: the L1 / shared split must leave room for the neighbour's CTA
    LVC_CUDA_TRY(cudaFuncSetAttribute(lvc_clip_kernel<MODE, ODT, STATS, NPW>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
  lvc_clip_kernel<MODE, ODT, STATS, NPW><<<grid, threads_of(MODE, NPW), smem, st>>>(p);
  LVC_CUDA_TRY(cudaGetLastError());
  return LVC_OK;
}

int check_desc(const lvc_clip_desc* d, bool boxes, bool aug, bool lut) {
  if (d->out_size != S) return lvc_fail(LVC_E_SHAPE, "lvc_clip_forward: out_size %d unsupported (224 only)", d->out_size);
  if (d->batch <= 0 || d->num_frames <= 0 || d->max_width <= 0)
    return lvc_fail(LVC_E_INVAL, "lvc_clip_forward: batch, num_frames and max_width must be positive");
  if (d->layout < LVC_LAYOUT_BTCHW || d->layout > LVC_LAYOUT_TOKENS || d->out_dtype < LVC_DT_F32 || d->out_dtype > LVC_DT_U8)
    return lvc_fail(LVC_E_INVAL, "lvc_clip_forward: unknown layout or dtype");
  if (d->layout == LVC_LAYOUT_TOKENS) {
    if (d->out_dtype != LVC_DT_BF16) return lvc_fail(LVC_E_NOTSUP, "lvc_clip_forward: the token layout is the bf16 GEMM operand");
    if (d->tok_kt <= 0 || d->tok_kh <= 0 || d->tok_kw <= 0 || d->num_frames % d->tok_kt || S % d->tok_kh || S % d->tok_kw)
      return lvc_fail(LVC_E_SHAPE, "lvc_clip_forward: patch (%d,%d,%d) must divide (T,224,224)", d->tok_kt, d->tok_kh, d->tok_kw);
  }
  if (d->resample < LVC_RESAMPLE_CV2 || d->resample > LVC_RESAMPLE_PTV_EVAL)
    return lvc_fail(LVC_E_INVAL, "lvc_clip_forward: unknown resample mode %d", d->resample);
  if (d->layout == LVC_LAYOUT_BTHWC && d->out_dtype != LVC_DT_U8)
    return lvc_fail(LVC_E_NOTSUP, "lvc_clip_forward: BTHWC layout is the uint8 Dataset clip only");
  const bool ptv = d->resample == LVC_RESAMPLE_PTV_EVAL;
  if (ptv) {
    if (d->out_dtype == LVC_DT_U8) return lvc_fail(LVC_E_NOTSUP, "lvc_clip_forward: the pytorchvideo chain is a float pipeline (no uint8 output)");
    if (boxes) return lvc_fail(LVC_E_NOTSUP, "lvc_clip_forward: ROI boxes are not defined for the pytorchvideo chain");
    for (int c = 0; c < 3; ++c)
      if (d->std[c] == 0.0f) return lvc_fail(LVC_E_INVAL, "lvc_clip_forward: std[%d] is zero", c);
  } else if (d->out_dtype != LVC_DT_U8 && !lut) {
    return lvc_fail(LVC_E_INVAL, "lvc_clip_forward: float output needs a LUT (lvc_build_lut)");
  }
  if (aug && !ptv) return lvc_fail(LVC_E_NOTSUP, "lvc_clip_forward: the augmentation table belongs to the pytorchvideo chain (LVC_RESAMPLE_PTV_EVAL)");
  if ((int64_t)d->batch * d->num_frames > (1 << 24)) return lvc_fail(LVC_E_SHAPE, "lvc_clip_forward: too many frames in one call");
  return LVC_OK;
}

// Launch geometry for a descriptor on a device with `sms` SMs and `max_smem` bytes of opt-in shared memory per CTA.
// Pure host arithmetic (lvc_clip_plan_query runs it without a device).
int make_plan(const lvc_clip_desc* d, bool boxes, bool stats, int sms, int max_smem, ClipParams& p, size_t& smem, int& grid,
              int smem_cap = 0, int built_producers = kProducers) {
  const bool ptv = d->resample == LVC_RESAMPLE_PTV_EVAL;
  const int mode = d->resample == LVC_RESAMPLE_CV2 ? MODE_CV2 : (ptv ? MODE_PTV : MODE_VIVIT);
  // row slot of the widest video: window bytes + alignment slack, rounded to 16 (+16 to stagger banks)
  p.pitch = ((d->max_width * 3 + 15 + 15) & ~15) + 16;
  // the whole opt-in shared memory of the SM (227 KB on B200): the kernel keeps next to nothing in L1, and the ViViT preset --
  // whose Pillow intermediate takes 50-60 KB -- gains a fifth ring stage (cfg 2: 133 -> 121 us); LVC_CLIP_SMEM_KB narrows it
  int budget_all = std::min(max_smem, std::max(env_int("LVC_CLIP_SMEM_KB", max_smem / 1024 + 1) * 1024, 96 * 1024));
  if (smem_cap > 0) budget_all = std::min(budget_all, smem_cap);   // the SM is shared with another kernel (fused patch embedding)
  // MODE_VIVIT: the Pillow stage needs its vertical window on chip.  A band of R output rows spans at most
  // floor((R-1)*7/8) + 4 resized rows (halo included); pick the band count that minimises rounds x rows per
  // item (whole frames when the batch is large enough that quantisation does not matter) AMONG the candidates whose
  // intermediate buffer leaves the source-row ring three 4-row stages of the widest video (a large batch of 1080p
  // frames would otherwise pick whole-frame items and starve -- or not fit -- the ring); when none does, the
  // smallest buffer.
  int v_nb = 1, v_res = S;
  auto vivit_ring_base = [](int res) { return (kVivitInterOff + (res + 7) / 8 * 8 * kRowBytes + 127) / 128 * 128; };
  if (mode == MODE_VIVIT) {
    static const int kVB[] = {1, 2, 4, 7, 8, 14};
    const long long nframes = (long long)d->batch * d->num_frames;
    const int forced_nb = env_int("LVC_CLIP_VBANDS", 0);
    long long best_cost = -1;
    bool best_ok = false;
    for (int nb : kVB) {
      const int R = S / nb;
      const int res = ((R - 1) * 7) / 8 + 4;   // nb == 1: rows 13..210, the crop never reads the rest
      const long long rounds = (nframes * nb + sms - 1) / sms;
      const long long cost = rounds * ((res + 7) / 8 * 8 + R / 2 + 6);   // resize rows + Pillow vertical rows + per-item fixed
      const bool ok = budget_all - vivit_ring_base(res) >= 3 * 8 * p.pitch;
      bool take;
      if (forced_nb) take = nb == forced_nb;
      else take = best_cost < 0 || (ok && !best_ok) || (ok == best_ok && (ok ? cost < best_cost : true));
      if (take) { best_cost = cost; best_ok = ok; v_nb = nb; v_res = res; }
    }
  }
  int fixed = mode == MODE_VIVIT ? vivit_ring_base(v_res) : kAuxOff;
  p.lut4_off = 0;
  if (mode == MODE_CV2 && !stats && !d->src_float && d->out_dtype != LVC_DT_U8) {   // the lean consumer's 3 x 1024 table
    p.lut4_off = fixed;
    fixed += 3 * 1024 * (d->out_dtype == LVC_DT_F32 ? 4 : 2);
  }
  p.ring_off = fixed;
  const int budget = budget_all - fixed;
  p.rps = env_int("LVC_CLIP_RPS", 0);
  if (p.rps == 0) {
    // full-frame sources keep the bus busy only with a finely sliced ring (a stage being consumed cannot be refilled:
    // 3 stages leave a third of the ring idle, 6 a sixth; cfg 3: 60 -> 53 us with 4 producers); ROI rows are much
    // shorter than a frame row (the ring is byte-granular: their stages take what they need) and the ViViT preset is
    // consumer-bound: those prefer fewer, larger stages
    // (with boxes the slots of max_width rows are only the worst case: two such stages must fit, the usual ROI stage is
    // a fraction of that and the byte-granular ring keeps several in flight)
    const int want_stages = boxes ? 2 : (mode == MODE_VIVIT ? 3 : 6);
    p.rps = (budget / (16 * p.pitch) >= want_stages) ? 8 : 4;
  }
  if (p.rps < 2 || p.rps > kMaxRPS || (S % p.rps)) return lvc_fail(LVC_E_INVAL, "LVC_CLIP_RPS=%d invalid", p.rps);
  int stage_bytes = 2 * p.rps * p.pitch;
  while (budget / stage_bytes < 2 && p.rps > 2) {   // very wide sources: fewer rows per stage (groups beyond rps idle)
    do --p.rps; while (S % p.rps);
    stage_bytes = 2 * p.rps * p.pitch;
  }
  if (budget / stage_bytes < 2) return lvc_fail(LVC_E_SHAPE, "lvc_clip_forward: max_width %d does not fit the shared-memory ring", d->max_width);
  p.stages = kMaxStages;
  // the ring never needs more than 16 stages of the widest rows (small sources then leave room for two CTAs per SM)
  p.ring_bytes = (int)std::min<long long>((long long)budget, (long long)kMaxStages * stage_bytes) & ~127;
  smem = (size_t)fixed + (size_t)p.ring_bytes;
  p.merge_rows = env_int("LVC_CLIP_MERGE", 1);
  p.no_fma = env_int("LVC_CLIP_NO_FMA", 0);
  p.no_tok44 = env_int("LVC_CLIP_NO_TOK44", 0);
  p.producers = std::max(1, std::min(built_producers, env_int("LVC_CLIP_NP", built_producers)));

  const int ctas_per_sm = smem_cap > 0 ? 1 : (int)std::max<size_t>(1, std::min<size_t>(2, (size_t)max_smem / (smem + 1024)));
  const int grid_max = sms * ctas_per_sm;
  const int frames = d->batch * d->num_frames;
  // band = output rows per work item.  A CTA walks its items round-robin, and every item boundary costs about
  // four rows' worth of time (descriptor fetch, tap set-up, a partly drained ring; measured on cfg 3 / cfg 4 with
  // scripts/clip_knobs.py): pick the divisor of 224 that minimises rounds x (rows + 4)
  static const int kBands[] = {224, 112, 56, 32, 28, 16, 8, 4, 2};
  p.rows_per_band = p.rps;
  const int forced = env_int("LVC_CLIP_BAND", 0);
  long long band_cost = -1;
  for (int rb : kBands) {
    if (rb % p.rps) continue;
    if (smem_cap > 0 && d->tok_kh > 0 && rb % d->tok_kh) continue;   // fused embedding: an item signals whole patch rows
    if (forced) { if (rb == forced) p.rows_per_band = rb; continue; }
    const long long items = (long long)frames * (S / rb);
    const long long cost = ((items + grid_max - 1) / grid_max) * (rb + 4);
    if (band_cost < 0 || cost < band_cost) { band_cost = cost; p.rows_per_band = rb; }
  }
  p.bands = S / p.rows_per_band;
  if (mode == MODE_VIVIT) { p.bands = v_nb; p.rows_per_band = S / v_nb; p.v_stages = (v_res + p.rps - 1) / p.rps; }
  p.total_items = frames * p.bands;
  grid = std::min(p.total_items, grid_max);
  return LVC_OK;
}

}  // namespace

namespace { unsigned long long* g_trace_buf = nullptr; }
// Profiling aid (not part of the public header): copies the last traced launch's timeline, 4 timestamps (ns) per stage of
// CTA 0 -- producer issued, consumer warp 0 began waiting, stage landed, consumer warp 0 done.
extern "C" int lvc_clip_trace_read(unsigned long long* out, int stages) {
  if (!g_trace_buf || !out) return lvc_fail(LVC_E_INVAL, "lvc_clip_trace_read: no trace (LVC_PROFILING_KNOBS=1 LVC_CLIP_TRACE=1)");
  (void)stages;   // the whole buffer: kTraceStages x 4 stage stamps, then 1024 x (CTA start, CTA end)
  LVC_CUDA_TRY(cudaMemcpy(out, g_trace_buf, (size_t)kTraceWords * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  return LVC_OK;
}

extern "C" int lvc_clip_launch_count(const lvc_clip_desc* d) { return d ? 1 : 0; }

extern "C" int lvc_clip_plan_query(const lvc_clip_desc* d, int32_t flags, int32_t sm_count, int32_t max_smem,
                                   lvc_clip_plan* out) {
  if (!d || !out) return lvc_fail(LVC_E_INVAL, "lvc_clip_plan_query: NULL argument");
  if (int rc = check_desc(d, (flags & LVC_PLAN_BOXES) != 0, false, true)) return rc;
  if (sm_count <= 0 || max_smem <= 0) {
    const DeviceInfo& di = device_info();
    if (di.err != cudaSuccess) return lvc_fail(LVC_E_CUDA, "lvc_clip_plan_query: no usable CUDA device: %s", cudaGetErrorString(di.err));
    sm_count = di.sms; max_smem = di.max_smem;
  }
  ClipParams p{};
  size_t smem = 0;
  int grid = 0;
  if (int rc = make_plan(d, (flags & LVC_PLAN_BOXES) != 0, (flags & LVC_PLAN_STATS) != 0, sm_count, max_smem, p, smem, grid)) return rc;
  const int mode = d->resample == LVC_RESAMPLE_CV2 ? MODE_CV2 : (d->resample == LVC_RESAMPLE_PTV_EVAL ? MODE_PTV : MODE_VIVIT);
  *out = lvc_clip_plan{};
  out->rows_per_stage = p.rps; out->ring_bytes = p.ring_bytes; out->ring_offset = p.ring_off; out->slot_pitch = p.pitch;
  out->stages_of_widest = p.ring_bytes / (2 * p.rps * p.pitch);
  out->bands = p.bands; out->rows_per_band = p.rows_per_band; out->producers = p.producers; out->grid = grid;
  out->threads = mode == MODE_VIVIT ? threads_of(MODE_VIVIT) : threads_of(MODE_CV2);
  out->smem_bytes = (int32_t)smem;
  return LVC_OK;
}

extern "C" int lvc_clip_validate(const lvc_clip_desc* d, int32_t max_height, const lvc_video* host_videos) {
  if (!d || !host_videos) return lvc_fail(LVC_E_INVAL, "lvc_clip_validate: NULL argument");
  for (int b = 0; b < d->batch; ++b) {
    const lvc_video& v = host_videos[b];
    if (max_height > 0 && v.height > max_height)
      return lvc_fail(LVC_E_SHAPE, "video %d is %d pixels high, higher than max_height = %d (which sizes the shared-memory buffers)", b, v.height, max_height);
    if (v.n_frames <= 0 || v.height <= 0 || v.width <= 0)
      return lvc_fail(LVC_E_INVAL, "video %d: n_frames, height and width must be positive (%d, %d, %d)", b, v.n_frames, v.height, v.width);
    if (v.width > d->max_width)
      return lvc_fail(LVC_E_SHAPE, "video %d is %d pixels wide, wider than desc.max_width = %d (which sizes the shared-memory ring)", b, v.width, d->max_width);
    if (v.base == 0 || (v.base & 15)) return lvc_fail(LVC_E_ALIGN, "video %d: base must be a 16-byte aligned device-readable address", b);
    if (v.nbytes < (int64_t)v.n_frames * v.height * v.width * 3)
      return lvc_fail(LVC_E_INVAL, "video %d: nbytes %lld is smaller than n_frames*height*width*3", b, (long long)v.nbytes);
  }
  return LVC_OK;
}

extern "C" int lvc_device_status(void* stream, uint32_t* flags, int32_t clear) {
  if (!flags) return lvc_fail(LVC_E_INVAL, "lvc_device_status: NULL argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unsigned* dptr = nullptr;
  LVC_CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void**>(&dptr), g_dev_error));
  unsigned h = 0;
  LVC_CUDA_TRY(cudaMemcpyAsync(&h, dptr, sizeof h, cudaMemcpyDeviceToHost, st));
  if (clear) LVC_CUDA_TRY(cudaMemsetAsync(dptr, 0, sizeof h, st));
  LVC_CUDA_TRY(cudaStreamSynchronize(st));
  *flags = h;
  return LVC_OK;
}

extern "C" int lvc_device_bytes_fetched(void* stream, uint64_t* bytes, int32_t clear) {
  if (!bytes) return lvc_fail(LVC_E_INVAL, "lvc_device_bytes_fetched: NULL argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unsigned long long* dptr = nullptr;
  LVC_CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void**>(&dptr), g_bytes_fetched));
  unsigned long long h = 0;
  LVC_CUDA_TRY(cudaMemcpyAsync(&h, dptr, sizeof h, cudaMemcpyDeviceToHost, st));
  if (clear) LVC_CUDA_TRY(cudaMemsetAsync(dptr, 0, sizeof h, st));
  LVC_CUDA_TRY(cudaStreamSynchronize(st));
  *bytes = h;
  return LVC_OK;
}

extern "C" int lvc_host_alloc(int64_t nbytes, int32_t flags, void** out) {
  if (!out || nbytes <= 0) return lvc_fail(LVC_E_INVAL, "lvc_host_alloc: bad argument");
  unsigned f = cudaHostAllocPortable | cudaHostAllocMapped;
  if (flags & LVC_HOST_WRITE_COMBINED) f |= cudaHostAllocWriteCombined;
  LVC_CUDA_TRY(cudaHostAlloc(out, (size_t)nbytes, f));
  return LVC_OK;
}
extern "C" int lvc_host_free(void* p) {
  if (p) LVC_CUDA_TRY(cudaFreeHost(p));
  return LVC_OK;
}

// device address of the error word for the other translation units' kernels
unsigned* lvc_dev_error_word() {
  static thread_local unsigned* ptr = nullptr;
  static thread_local int dev_of = -1;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
  if (dev != dev_of) {
    if (cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_dev_error) != cudaSuccess) return nullptr;
    dev_of = dev;
  }
  return ptr;
}

extern "C" int lvc_clip_forward(const lvc_clip_desc* d, const lvc_video* videos, const int32_t* indices,
                                const int32_t* boxes, const float* lut, void* out, uint64_t* stats,
                                void* stream) {
  return lvc_clip_forward_aug(d, videos, indices, boxes, nullptr, lut, out, stats, stream);
}

extern "C" int lvc_clip_forward_aug(const lvc_clip_desc* d, const lvc_video* videos, const int32_t* indices,
                                    const int32_t* boxes, const int32_t* aug, const float* lut, void* out,
                                    uint64_t* stats, void* stream) {
  if (!d || !videos || !indices || !out) return lvc_fail(LVC_E_INVAL, "lvc_clip_forward: NULL argument");
  if (int rc = check_desc(d, boxes != nullptr, aug != nullptr, lut != nullptr)) return rc;
  if (boxes && (reinterpret_cast<uintptr_t>(boxes) & 15)) return lvc_fail(LVC_E_ALIGN, "lvc_clip_forward: boxes must be 16-byte aligned");
  if (aug && (reinterpret_cast<uintptr_t>(aug) & 15)) return lvc_fail(LVC_E_ALIGN, "lvc_clip_forward: aug must be 16-byte aligned");

  const DeviceInfo& di = device_info();
  if (di.err != cudaSuccess) return lvc_fail(LVC_E_CUDA, "lvc_clip_forward: no usable CUDA device: %s", cudaGetErrorString(di.err));

  const bool ptv = d->resample == LVC_RESAMPLE_PTV_EVAL;
  const int mode = d->resample == LVC_RESAMPLE_CV2 ? MODE_CV2 : (ptv ? MODE_PTV : MODE_VIVIT);
  ClipParams p{};
  p.videos = videos; p.indices = indices; p.boxes = boxes; p.aug = aug; p.lut = lut; p.out = out;
  p.stats = reinterpret_cast<unsigned long long*>(stats);
  p.T = d->num_frames; p.layout = d->layout; p.out_dtype = d->out_dtype;
  for (int c = 0; c < 3; ++c) { p.mean[c] = d->mean[c]; p.std[c] = d->std[c]; }
  p.kt = d->tok_kt; p.kh = d->tok_kh; p.kw = d->tok_kw;
  p.bgr = d->src_order == LVC_SRC_BGR;
  p.f32src = d->src_float != 0 && !ptv;
  p.kh_shift = -1;
  for (int sft = 0; sft < 8; ++sft) if (p.kh == (1 << sft)) p.kh_shift = sft;
  p.err = lvc_dev_error_word();
  if (!p.err) return lvc_fail(LVC_E_CUDA, "lvc_clip_forward: cannot resolve the device error word");
  {
    static thread_local unsigned long long* fptr = nullptr;
    static thread_local int fdev = -1;
    int dev = -1;
    LVC_CUDA_TRY(cudaGetDevice(&dev));
    if (dev != fdev) {
      LVC_CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void**>(&fptr), g_bytes_fetched));
      fdev = dev;
    }
    p.fetched = fptr;
  }
  size_t smem = 0;
  int grid = 0;
  if (int rc = make_plan(d, boxes != nullptr, stats != nullptr, di.sms, di.max_smem, p, smem, grid)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (env_int("LVC_CLIP_TRACE", 0)) {   // profiling only: per-stage timeline of CTA 0 (lvc_clip_trace_read)
    if (!kTraceBuild) return lvc_fail(LVC_E_NOTSUP, "LVC_CLIP_TRACE needs the trace build of the library (make -C csrc trace; LVC_B200_LIB=.../liblvc_b200_trace.so)");
    if (!g_trace_buf) LVC_CUDA_TRY(cudaMalloc(&g_trace_buf, kTraceWords * sizeof(unsigned long long)));
    LVC_CUDA_TRY(cudaMemsetAsync(g_trace_buf, 0, kTraceWords * sizeof(unsigned long long), st));
    p.trace = g_trace_buf;
  }
  if (mode == MODE_CV2) {
    if (p.f32src) return launch<MODE_CV2>(p, grid, smem, st);   // float frames: the general consumer
    const bool st8 = stats != nullptr;
    switch (d->out_dtype) {
      case LVC_DT_F32: return st8 ? launch<MODE_CV2, LVC_DT_F32, true>(p, grid, smem, st) : launch<MODE_CV2, LVC_DT_F32, false>(p, grid, smem, st);
      case LVC_DT_BF16: return st8 ? launch<MODE_CV2, LVC_DT_BF16, true>(p, grid, smem, st) : launch<MODE_CV2, LVC_DT_BF16, false>(p, grid, smem, st);
      default: return st8 ? launch<MODE_CV2, LVC_DT_U8, true>(p, grid, smem, st) : launch<MODE_CV2, LVC_DT_U8, false>(p, grid, smem, st);
    }
  }
  if (mode == MODE_VIVIT)
    return stats ? launch<MODE_VIVIT, -1, true>(p, grid, smem, st) : launch<MODE_VIVIT, -1, false>(p, grid, smem, st);
  return launch<MODE_PTV, -1, false>(p, grid, smem, st);
}

// Clip kernel for lvc_clip_swin_embed_forward (csrc/lvc_embed.cu): bf16 token layout, no checksum, two producer warps,
// shared memory capped at `smem_cap` so that the patch-embedding kernel's CTA fits on the same SM, and per-tile ready
// counters for that kernel (`tile_ready[t]` += token-row slices written into tile t of A; tiles of 1 << tile_shift tokens).
int lvc_clip_forward_fused(const lvc_clip_desc* d, const lvc_video* videos, const int32_t* indices, const int32_t* boxes,
                           const float* lut, void* out, unsigned* tile_ready, int tile_shift, int smem_cap, void* stream) {
  if (!d || !videos || !indices || !out || !tile_ready || !lut) return lvc_fail(LVC_E_INVAL, "clip_swin_embed: NULL argument");
  if (int rc = check_desc(d, boxes != nullptr, false, true)) return rc;
  if (d->resample != LVC_RESAMPLE_CV2 || d->layout != LVC_LAYOUT_TOKENS || d->out_dtype != LVC_DT_BF16 || d->src_float)
    return lvc_fail(LVC_E_NOTSUP, "clip_swin_embed: the fused path is the cv2 chain on uint8 frames with the bf16 token layout");
  if (boxes && (reinterpret_cast<uintptr_t>(boxes) & 15)) return lvc_fail(LVC_E_ALIGN, "clip_swin_embed: boxes must be 16-byte aligned");
  const DeviceInfo& di = device_info();
  if (di.err != cudaSuccess) return lvc_fail(LVC_E_CUDA, "clip_swin_embed: no usable CUDA device: %s", cudaGetErrorString(di.err));
  ClipParams p{};
  p.videos = videos; p.indices = indices; p.boxes = boxes; p.lut = lut; p.out = out;
  p.T = d->num_frames; p.layout = d->layout; p.out_dtype = d->out_dtype;
  p.kt = d->tok_kt; p.kh = d->tok_kh; p.kw = d->tok_kw;
  p.bgr = d->src_order == LVC_SRC_BGR;
  p.kh_shift = -1;
  for (int sft = 0; sft < 8; ++sft) if (p.kh == (1 << sft)) p.kh_shift = sft;
  p.tile_ready = tile_ready; p.tile_shift = tile_shift;
  {
    const long long tiles = ((long long)d->batch * (d->num_frames / d->tok_kt) * (S / d->tok_kh) * (S / d->tok_kw) + (1 << tile_shift) - 1) >> tile_shift;
    p.tile_stamps = reinterpret_cast<unsigned long long*>(tile_ready + (tiles + 63) / 64 * 64);
  }
  p.err = lvc_dev_error_word();
  if (!p.err) return lvc_fail(LVC_E_CUDA, "clip_swin_embed: cannot resolve the device error word");
  {
    unsigned long long* fptr = nullptr;
    LVC_CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void**>(&fptr), g_bytes_fetched));
    p.fetched = fptr;
  }
  size_t smem = 0;
  int grid = 0;
  if (int rc = make_plan(d, boxes != nullptr, false, di.sms, std::min(di.max_smem, smem_cap), p, smem, grid, smem_cap, kProducersFused)) return rc;
  return launch<MODE_CV2, LVC_DT_BF16, false, kProducersFused>(p, grid, smem, static_cast<cudaStream_t>(stream));
}
