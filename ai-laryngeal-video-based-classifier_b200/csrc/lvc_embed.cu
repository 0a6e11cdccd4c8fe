// Patch / tubelet embedding for sm_100a: Conv3d/Conv2d with kernel == stride as a bf16 GEMM on the
// 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM, operands staged by TMA).
// Replaces the first layer of the three transformer models the reference fine-tunes:
//   VivitTubeletEmbeddings       (vivit_transformer/vivit_classifier/models/vivit_model.py:27-41)
//   TimesformerPatchEmbeddings   (timesformer/timesformer_classifier/models/timesformer_model.py:27-42)
//   torchvision PatchEmbed3d     (videoswintransformer/swin_video_classifier/models/swin3d.py:22-44)
//
// Pipeline per call:  im2col (x -> A[M][K] bf16, token-major, K = (c,kt,kh,kw))  ->  persistent
// warp-specialised GEMM  D[M][N] = A . W^T + bias (+ LayerNorm over N for Swin):
//   warp 0    TMA producer   cp.async.bulk.tensor.2d, 128B/64B swizzle, mbarrier ring
//   warp 1    MMA issuer     one elected lane: tcgen05.mma.cta_group::1.kind::f16, M=128, N=BN, K=16
//   warps 2-5 epilogue       tcgen05.ld 32x32b -> bias/LayerNorm -> bf16/fp32 rows to global
// Two TMEM accumulator buffers overlap tile i's epilogue with tile i+1's MMAs.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <utility>

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdlib>

#include "lvc_internal.h"

namespace {

constexpr int BM = 128;          // rows (tokens) per tile = UMMA M
constexpr int kGemmThreads = 192;

struct EmbedGeom {
  int B, T, H, W;      // input extents (channels = 3)
  int kt, kh, kw;
  int To, Ho, Wo;      // output token grid (ceil: PatchEmbed3d pads)
  int M, K, N;
};

int make_geom(const lvc_embed_desc* d, EmbedGeom* g) {
  if (!d) return lvc_fail(LVC_E_INVAL, "patch_embed: NULL descriptor");
  if (d->kind < LVC_EMBED_VIVIT || d->kind > LVC_EMBED_SWIN) return lvc_fail(LVC_E_INVAL, "patch_embed: unknown kind %d", d->kind);
  if (d->batch <= 0 || d->frames <= 0 || d->height <= 0 || d->width <= 0 || d->kt <= 0 || d->kh <= 0 || d->kw <= 0 || d->embed_dim <= 0)
    return lvc_fail(LVC_E_INVAL, "patch_embed: extents must be positive");
  g->B = d->batch; g->T = d->frames; g->H = d->height; g->W = d->width;
  g->kt = d->kt; g->kh = d->kh; g->kw = d->kw;
  if (d->kind == LVC_EMBED_SWIN) {  // pads up to a multiple of the patch (swin_transformer.py:353-358)
    g->To = (g->T + g->kt - 1) / g->kt; g->Ho = (g->H + g->kh - 1) / g->kh; g->Wo = (g->W + g->kw - 1) / g->kw;
  } else {                          // Conv with stride == kernel and no padding floors
    if (d->kind == LVC_EMBED_TIMESFORMER && d->kt != 1) return lvc_fail(LVC_E_INVAL, "patch_embed: TimeSformer patches are 2-D (kt must be 1)");
    g->To = g->T / g->kt; g->Ho = g->H / g->kh; g->Wo = g->W / g->kw;
  }
  const long long M = 1LL * g->B * g->To * g->Ho * g->Wo;
  if (M <= 0 || M > 0x7fffffffLL) return lvc_fail(LVC_E_SHAPE, "patch_embed: token count %lld out of range", M);
  g->M = (int)M; g->K = 3 * g->kt * g->kh * g->kw; g->N = d->embed_dim;
  if (g->K % 32) return lvc_fail(LVC_E_SHAPE, "patch_embed: K = 3*kt*kh*kw = %d must be a multiple of 32", g->K);
  if (g->N % 16 || g->N < 16) return lvc_fail(LVC_E_SHAPE, "patch_embed: embed_dim %d must be a multiple of 16", g->N);
  if (d->kind == LVC_EMBED_SWIN && g->N > 256) return lvc_fail(LVC_E_SHAPE, "patch_embed: fused LayerNorm needs embed_dim <= 256");
  if (g->N > 256 && g->N % 128) return lvc_fail(LVC_E_SHAPE, "patch_embed: embed_dim %d > 256 must be a multiple of 128", g->N);
  return LVC_OK;
}

// ---------------------------------------------------------------------------- im2col ----
// One thread per (token, c, dt, dy) run of kw contiguous pixels.  Reads are contiguous along x
// (kw elements) and across the w' tokens of an image row; writes are kw*2-byte runs of A.
template <typename TIn>
__global__ void lvc_im2col_kernel(const TIn* __restrict__ x, __nv_bfloat16* __restrict__ a, EmbedGeom g, int bcthw) {
  const long long runs = 1LL * g.M * 3 * g.kt * g.kh;
  const long long stride = 1LL * gridDim.x * blockDim.x;
  for (long long r = 1LL * blockIdx.x * blockDim.x + threadIdx.x; r < runs; r += stride) {
    // order: x-token fastest, then dy, c/dt, so a warp reads one image row contiguously
    long long q = r;
    const int wo = (int)(q % g.Wo); q /= g.Wo;
    const int dy = (int)(q % g.kh); q /= g.kh;
    const int ho = (int)(q % g.Ho); q /= g.Ho;
    const int dt = (int)(q % g.kt); q /= g.kt;
    const int c = (int)(q % 3); q /= 3;
    const int to = (int)(q % g.To); q /= g.To;
    const int b = (int)q;
    const int t = to * g.kt + dt, y = ho * g.kh + dy, x0 = wo * g.kw;
    const long long m = ((1LL * b * g.To + to) * g.Ho + ho) * g.Wo + wo;
    __nv_bfloat16* dst = a + m * g.K + ((c * g.kt + dt) * g.kh + dy) * g.kw;
    const bool in = (t < g.T) && (y < g.H);
    const long long src = bcthw ? (((1LL * b * 3 + c) * g.T + t) * g.H + y) * g.W + x0
                                : (((1LL * b * g.T + t) * 3 + c) * g.H + y) * g.W + x0;
    for (int i = 0; i < g.kw; ++i) {
      float v = 0.f;
      if (in && x0 + i < g.W) v = (float)x[src + i];
      dst[i] = __float2bfloat16_rn(v);
    }
  }
}

// Vectorised variant for the shipped patch shapes (kw = 16 or 4, no padding, 16-byte aligned rows):
// a thread converts ROWS x KW pixels with 16-byte loads and writes one contiguous bf16 run of A.
template <typename TIn, int KW, int ROWS>
__global__ void lvc_im2col_vec_kernel(const TIn* __restrict__ x, __nv_bfloat16* __restrict__ a, EmbedGeom g, int bcthw) {
  const unsigned dys = g.kh / ROWS;
  const unsigned runs = (unsigned)g.M * 3u * g.kt * dys;
  for (unsigned r = blockIdx.x * blockDim.x + threadIdx.x; r < runs; r += gridDim.x * blockDim.x) {
    unsigned q = r;
    const unsigned wo = q % g.Wo; q /= g.Wo;
    const unsigned dyb = q % dys; q /= dys;
    const unsigned ho = q % g.Ho; q /= g.Ho;
    const unsigned dt = q % g.kt; q /= g.kt;
    const unsigned c = q % 3u; q /= 3u;
    const unsigned to = q % g.To; q /= g.To;
    const unsigned b = q;
    const unsigned t = to * g.kt + dt, y = ho * g.kh + dyb * ROWS, x0 = wo * KW;
    const size_t m = (((size_t)b * g.To + to) * g.Ho + ho) * g.Wo + wo;
    __nv_bfloat16* dst = a + m * g.K + ((c * g.kt + dt) * g.kh + dyb * ROWS) * KW;
    const size_t src = bcthw ? ((((size_t)b * 3 + c) * g.T + t) * g.H + y) * g.W + x0
                             : ((((size_t)b * g.T + t) * 3 + c) * g.H + y) * g.W + x0;
    uint32_t packed[ROWS * KW / 2];
#pragma unroll
    for (int rr = 0; rr < ROWS; ++rr) {
      const TIn* sp = x + src + (size_t)rr * g.W;
      if constexpr (sizeof(TIn) == 4) {
#pragma unroll
        for (int i = 0; i < KW / 4; ++i) {
          const float4 f = __ldcs(reinterpret_cast<const float4*>(sp) + i);
          __nv_bfloat162 lo = __floats2bfloat162_rn(f.x, f.y), hi = __floats2bfloat162_rn(f.z, f.w);
          packed[rr * KW / 2 + 2 * i] = *reinterpret_cast<uint32_t*>(&lo);
          packed[rr * KW / 2 + 2 * i + 1] = *reinterpret_cast<uint32_t*>(&hi);
        }
      } else {
#pragma unroll
        for (int i = 0; i < KW / 4; ++i) {
          const uint2 u = __ldcs(reinterpret_cast<const uint2*>(sp) + i);
          packed[rr * KW / 2 + 2 * i] = u.x;
          packed[rr * KW / 2 + 2 * i + 1] = u.y;
        }
      }
    }
    uint4* d4 = reinterpret_cast<uint4*>(dst);
#pragma unroll
    for (int i = 0; i < ROWS * KW / 8; ++i)
      d4[i] = make_uint4(packed[4 * i], packed[4 * i + 1], packed[4 * i + 2], packed[4 * i + 3]);
  }
}

// conv weight (N, 3, kt, kh, kw) fp32 is already [N][K] row-major: packing = bf16 rounding.
__global__ void lvc_pack_w_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ o, long long n) {
  for (long long i = 1LL * blockIdx.x * blockDim.x + threadIdx.x; i < n; i += 1LL * gridDim.x * blockDim.x)
    o[i] = __float2bfloat16_rn(w[i]);
}

}  // namespace

namespace {

// ------------------------------------------------------------------------ PTX helpers ----
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mb_init(uint64_t* b, uint32_t n) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(n));
}
__device__ __forceinline__ void mb_expect_tx(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mb_arrive(uint64_t* b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(b)) : "memory");
}
__device__ __forceinline__ void mb_wait(uint64_t* b, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(s32(b)), "r"(parity) : "memory");
  } while (!done);
}
// wait with back-off: for warps that idle for long stretches next to busy warps of the same SM
// sub-partition (a bare try_wait spin competes with them for issue slots)
__device__ __forceinline__ void mb_wait_backoff(uint64_t* b, uint32_t parity) {
  uint32_t done;
  for (;;) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(s32(b)), "r"(parity), "r"(2000u) : "memory");
    if (done) break;
    __nanosleep(100);
  }
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
// 2-D tiled TMA load (SASS: UTMALDG): coordinates {c0 = element along K, c1 = row}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(s32(dst)), "l"(map), "r"(s32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
// tcgen05
__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(slot)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// D[tmem] (+)= A[smem] . B[smem]^T, bf16 inputs, fp32 accumulate (SASS: UTCHMMA)
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// mbarrier arrives once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(bar)) : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread i of the warp gets TMEM lane (base lane + i)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}


// Asynchronous variant: the 32 destination registers are valid only after tmem_wait_dep(v), whose
// "+r" operands give the compiler a true dependency (nothing that reads v can move above the wait).
__device__ __forceinline__ void tmem_ld32_async(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_wait_dep(float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                 "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                 "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
                 "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :: "memory");
}
// 2-D tiled TMA store (SASS: UTMASTG) of a swizzled shared-memory tile; rows/columns beyond the
// tensor's extents are clipped by the hardware.
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(map), "r"(s32(src)), "r"(c0), "r"(c1) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Packed fp32x2 arithmetic (SASS FFMA2 / FADD2): two lanes per instruction, halves the issue slots of
// the LayerNorm epilogue.
__device__ __forceinline__ uint64_t pack2(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}

// Shared-memory matrix descriptor for a K-major bf16 tile whose rows are one swizzle span wide
// (BK*2 = 128 B -> SWIZZLE_128B, 64 B -> SWIZZLE_64B): 8-row groups are `span*8` bytes apart (SBO);
// LBO is unused for swizzled K-major operands; bits 46-47 = descriptor version 1 (sm_100).
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, int span_bytes) {
  const uint64_t layout = span_bytes == 128 ? 2ull : 4ull;  // SWIZZLE_128B = 2, SWIZZLE_64B = 4
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3ffff) >> 4);
  d |= (uint64_t)(((uint32_t)span_bytes * 8u) >> 4) << 32;
  d |= 1ull << 46;
  d |= layout << 61;
  return d;
}
// kind::f16 instruction descriptor: D = F32 (bits 4-5 = 1), A = B = BF16 (bits 7-9, 10-12 = 1),
// both K-major, N>>3 at bits 17-22, M>>4 at bits 24-28.
__device__ __forceinline__ uint32_t make_idesc(int m, int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

}  // namespace

namespace {

struct GemmParams {
  int M, N, K;
  int BN, BK;          // tile columns, K elements per stage (64 -> 128B swizzle, 32 -> 64B swizzle)
  int stages, num_k_blocks, m_tiles, n_tiles;
  uint32_t tmem_cols;  // power of two >= 2*BN
  const float* bias;
  const float* ln_w;
  const float* ln_b;
  float ln_eps;
  int do_ln, out_f32;
  // fused embeddings epilogue: emb_mode 1 = VivitEmbeddings, 2 = TimesformerEmbeddings
  const float* cls;    // [N]
  const float* pos;    // [(1 + P)][N] position_embeddings
  const float* temb;   // [T][N] time_embeddings (mode 2, may be NULL = space_only)
  int emb_mode, emb_P, emb_T;
  // fused clip -> embed (lvc_clip_swin_embed_forward): ready[tile] counts the token-row slices the clip kernel has
  // written into tile `tile` of A; the TMA producer loads a tile once ready[tile] >= tokens in the tile x ready_need.  NULL = A is complete.
  const unsigned* ready;
  unsigned ready_need;
  const void* diag_a;  // A base pointer for the linear-copy experiment
  int diag;            // LVC_GEMM_DIAG (profiling only): 1 = no output stores, 2 = no MMAs, 4 = no TMA loads
  void* out;
};

// Coalesced epilogue store: every thread owns one output row (TMEM lane); its `row_bytes` (64 or 128)
// of packed results go through a warp-private, XOR-swizzled shared-memory tile so that the global
// stores cover whole 128-byte lines (8 lanes x 16 B per row) instead of 32 scattered 16-byte pieces.
__device__ __forceinline__ void stage_store_rows(uint8_t* stg, const uint32_t* pk, int lane, uint8_t* gbase,
                                                 size_t row_stride, int rows_valid, int row_bytes) {
  const int cpr = row_bytes >> 4;  // 16-byte chunks per row: 4 or 8
  __syncwarp();
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (j < cpr)
      *reinterpret_cast<uint4*>(stg + lane * row_bytes + ((j ^ (lane & (cpr - 1))) << 4)) =
          make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
  __syncwarp();
  const int rpi = 32 / cpr;  // rows covered by one warp-wide store
  const int c = lane % cpr, r0 = lane / cpr;
#pragma unroll
  for (int it = 0; it < 8; ++it)
    if (it < cpr) {
      const int row = it * rpi + r0;
      const uint4 val = *reinterpret_cast<const uint4*>(stg + row * row_bytes + ((c ^ (row & (cpr - 1))) << 4));
      if (row < rows_valid) *reinterpret_cast<uint4*>(gbase + (size_t)row * row_stride + (c << 4)) = val;
    }
}
// Packs 32 fp32 accumulator columns into the thread's staging words: bf16 -> 16 words at pk[half*16],
// fp32 -> 32 words.
template <int HALF>
__device__ __forceinline__ void pack_bf16(const float* v, uint32_t* pk) {
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
    pk[HALF * 16 + j] = *reinterpret_cast<uint32_t*>(&h);
  }
}
// Emits chunk `ch` (32 columns starting at n0) of this thread's row; flushes a full 128-byte staging
// row (two bf16 chunks or one fp32 chunk) or the trailing 64-byte half row.
__device__ __forceinline__ void emit_chunk(const float* v, uint32_t* pk, const GemmParams& p, uint8_t* stg, int lane,
                                           long long m_warp0, int n0, int ch, int nchunks) {
  const int rows_valid = (int)max(0LL, min(32LL, (long long)p.M - m_warp0));
  const size_t esz = p.out_f32 ? 4 : 2;
  const size_t stride = (size_t)p.N * esz;
  uint8_t* rowbase = reinterpret_cast<uint8_t*>(p.out) + (size_t)m_warp0 * stride;
  if (p.out_f32) {
#pragma unroll
    for (int j = 0; j < 32; ++j) pk[j] = __float_as_uint(v[j]);
    stage_store_rows(stg, pk, lane, rowbase + (size_t)n0 * 4, stride, rows_valid, 128);
  } else if (ch & 1) {
    pack_bf16<1>(v, pk);
    stage_store_rows(stg, pk, lane, rowbase + (size_t)(n0 - 32) * 2, stride, rows_valid, 128);
  } else {
    pack_bf16<0>(v, pk);
    if (ch == nchunks - 1) stage_store_rows(stg, pk, lane, rowbase + (size_t)n0 * 2, stride, rows_valid, 64);
  }
}

__global__ void __launch_bounds__(kGemmThreads, 2)
lvc_embed_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                      const __grid_constant__ GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (s32(smem_raw) & 1023u)) & 1023u);  // 1024-aligned, stays in the shared window
  const int span = p.BK * 2;                       // bytes per tile row = swizzle span
  const uint32_t a_bytes = BM * span, b_bytes = p.BN * span;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + (size_t)p.stages * a_bytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_b + (size_t)p.stages * b_bytes);
  uint64_t* empty = full + p.stages;
  uint64_t* tmem_full = empty + p.stages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  float* s_bias = reinterpret_cast<float*>(tmem_slot + 4);  // N floats, then LayerNorm weight / bias (BN each)
  float* s_lnw = s_bias + p.N;
  float* s_lnb = s_lnw + p.BN;
  uint8_t* stg = reinterpret_cast<uint8_t*>(s_lnb + p.BN) + (size_t)(threadIdx.x >> 5 & 3) * 4096;  // per epilogue warp
  for (int i = threadIdx.x; i < p.N; i += kGemmThreads) s_bias[i] = p.bias[i];
  if (p.do_ln)
    for (int i = threadIdx.x; i < p.BN; i += kGemmThreads) { s_lnw[i] = p.ln_w[i]; s_lnb[i] = p.ln_b[i]; }

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_tiles = p.m_tiles * p.n_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    for (int s = 0; s < p.stages; ++s) { mb_init(&full[s], 1); mb_init(&empty[s], 1); }
    for (int a = 0; a < 2; ++a) { mb_init(&tmem_full[a], 1); mb_init(&tmem_empty[a], 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m_blk = tile / p.n_tiles, n_blk = tile - m_blk * p.n_tiles;
        for (int kb = 0; kb < p.num_k_blocks; ++kb, ++it) {
          const int s = it % p.stages;
          mb_wait(&empty[s], ((it / p.stages) & 1u) ^ 1u);
          mb_expect_tx(&full[s], a_bytes + b_bytes);
          tma_load_2d(smem_a + (size_t)s * a_bytes, &map_a, kb * p.BK, m_blk * BM, &full[s]);
          tma_load_2d(smem_b + (size_t)s * b_bytes, &map_b, kb * p.BK, n_blk * p.BN, &full[s]);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      const uint32_t idesc = make_idesc(BM, p.BN);
      uint32_t it = 0, local = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
        const uint32_t acc = local & 1u, acc_phase = (local >> 1) & 1u;
        mb_wait(&tmem_empty[acc], acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * (uint32_t)p.BN;
        for (int kb = 0; kb < p.num_k_blocks; ++kb, ++it) {
          const int s = it % p.stages;
          mb_wait(&full[s], (it / p.stages) & 1u);
          tc_fence_after();
          const uint64_t adesc = make_smem_desc(s32(smem_a + (size_t)s * a_bytes), span);
          const uint64_t bdesc = make_smem_desc(s32(smem_b + (size_t)s * b_bytes), span);
          for (int k = 0; k < p.BK / 16; ++k)  // +32 bytes (2 x 16 B units) per K=16 step inside the swizzle span
            umma_bf16(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (uint32_t)((kb | k) != 0));
          umma_commit(&empty[s]);    // smem stage free once these MMAs have read it
        }
        umma_commit(&tmem_full[acc]);  // accumulator complete
      }
    }
  } else {
    // ===================== epilogue: 4 warps, warp%4 selects its TMEM lane quarter =====================
    const int q = warp & 3;
    uint32_t local = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      const int m_blk = tile / p.n_tiles, n_blk = tile - m_blk * p.n_tiles;
      const uint32_t acc = local & 1u, acc_phase = (local >> 1) & 1u;
      mb_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * (uint32_t)p.BN;
      const long long m_warp0 = (long long)m_blk * BM + q * 32;
      const int nchunks = p.BN / 32;
      const int NCH = nchunks;
      uint32_t pk[32];
      float mean = 0.f, rstd = 1.f;
      float v[32];
      const float* sb = s_bias + n_blk * p.BN;
      if (p.do_ln) {  // LayerNorm over the N = BN columns this thread owns (one token row): one stats pass
        float sum = 0.f, sq = 0.f;
        for (int ch = 0; ch < nchunks; ++ch) {
          tmem_ld32(taddr + ch * 32, v);
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
            const float4 b4 = reinterpret_cast<const float4*>(sb + ch * 32)[j4];
            const float z0 = v[4 * j4] + b4.x, z1 = v[4 * j4 + 1] + b4.y, z2 = v[4 * j4 + 2] + b4.z, z3 = v[4 * j4 + 3] + b4.w;
            sum += (z0 + z1) + (z2 + z3);
            sq = fmaf(z0, z0, fmaf(z1, z1, fmaf(z2, z2, fmaf(z3, z3, sq))));
          }
        }
        mean = sum / (float)p.N;
        rstd = rsqrtf(fmaxf(sq / (float)p.N - mean * mean, 0.f) + p.ln_eps);
      }
      for (int ch = 0; ch < nchunks; ++ch) {
        tmem_ld32(taddr + ch * 32, v);
        const int n0 = n_blk * p.BN + ch * 32;
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 b4 = reinterpret_cast<const float4*>(sb + ch * 32)[j4];
          v[4 * j4] += b4.x; v[4 * j4 + 1] += b4.y; v[4 * j4 + 2] += b4.z; v[4 * j4 + 3] += b4.w;
          if (p.do_ln) {
            const float4 g4 = reinterpret_cast<const float4*>(s_lnw + ch * 32)[j4];
            const float4 o4 = reinterpret_cast<const float4*>(s_lnb + ch * 32)[j4];
            float tt;
            tt = rstd * g4.x; v[4 * j4] = fmaf(v[4 * j4], tt, fmaf(-mean, tt, o4.x));
            tt = rstd * g4.y; v[4 * j4 + 1] = fmaf(v[4 * j4 + 1], tt, fmaf(-mean, tt, o4.y));
            tt = rstd * g4.z; v[4 * j4 + 2] = fmaf(v[4 * j4 + 2], tt, fmaf(-mean, tt, o4.z));
            tt = rstd * g4.w; v[4 * j4 + 3] = fmaf(v[4 * j4 + 3], tt, fmaf(-mean, tt, o4.w));
          }
        }
        emit_chunk(v, pk, p, stg, lane, m_warp0, n0, ch, NCH);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mb_arrive(&tmem_empty[acc]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

}  // namespace

// ---------------------------------------------------------------- 2-CTA (cta_group::2) GEMM ----
// CTA pairs (cluster 2x1x1) cooperate on a 256 x BN tile: each CTA stages its own 128 rows of A and
// HALF of the B tile, the leader CTA issues tcgen05.mma.cta_group::2 (M = 256) which reads both
// halves, and each CTA's TMEM receives its 128 accumulator rows.  Per k-block a CTA pulls
// (128 + BN/2) x 64 bf16 = 32 KB from L2 instead of 48 KB: the 1-CTA kernel is L2->SM feed-bound
// (MMA issuer stalls on the `full` barriers), this cuts the feed by a third and deepens the ring.
namespace {

// Programmatic dependent launch: the kernel may start while its predecessor in the stream is still running (its
// prologue overlaps the predecessor's tail); nothing the predecessor wrote may be read, and nothing it may still read
// may be overwritten, before this returns.  A no-op when the launch carries no programmatic-serialization attribute.
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `p` (a local shared address) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
__device__ __forceinline__ void mb_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_2d_2sm(void* dst, const CUtensorMap* map, int c0, int c1, uint32_t bar_cluster_addr) {
  asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(s32(dst)), "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t* slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(slot)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_bf16_2sm(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive (once the MMAs issued so far completed) on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(s32(bar)), "h"((uint16_t)3) : "memory");
}

constexpr int kGemm2Threads = 320;  // producer warp, MMA warp, 8 epilogue warps (two per TMEM lane quarter)

template <bool EMB>  // EMB: fused cls / position / time embeddings epilogue (separate instantiation keeps the plain one lean)
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemm2Threads, 1)
lvc_embed_gemm2_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                       const __grid_constant__ CUtensorMap map_o, const __grid_constant__ GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  if (p.diag & 8) return;
  uint8_t* smem = smem_raw + ((1024u - (s32(smem_raw) & 1023u)) & 1023u);  // 1024-aligned, stays in the shared window
  const int span = p.BK * 2;
  const int half_n = p.BN / 2;
  const uint32_t a_bytes = BM * span, b_bytes = half_n * span;
  // staging tiles of the TMA stores first (128B-swizzled boxes need 1024-byte alignment): 8 warps x 4 KB
  uint8_t* stg = smem + (size_t)((threadIdx.x >> 5) - 2 & 7) * 4096;
  uint8_t* smem_a = smem + 8 * 4096;
  uint8_t* smem_b = smem_a + (size_t)p.stages * a_bytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_b + (size_t)p.stages * b_bytes);
  uint64_t* empty = full + p.stages;
  uint64_t* tmem_full = empty + p.stages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_rank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int num_units = p.m_tiles * p.n_tiles;  // m_tiles counts 256-row pair tiles here

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    for (int s = 0; s < p.stages; ++s) { mb_init(&full[s], 1); mb_init(&empty[s], 1); }
    tma_prefetch_desc(&map_o);
    for (int a = 0; a < 2; ++a) { mb_init(&tmem_full[a], 1); mb_init(&tmem_empty[a], 16); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc_2sm(tmem_slot, p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // peer barriers are initialised before any remote signal
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_wait();       // everything above touched no global data

  if (EMB && warp == 0) {
    // row 0 of every batch entry: cls_token + position_embeddings[0] (time embeddings do not touch it)
    const int rows_per_b = 1 + p.emb_P * (p.emb_mode == 2 ? p.emb_T : 1);
    const int nb = p.M / (p.emb_P * (p.emb_mode == 2 ? p.emb_T : 1));
    for (int bb = blockIdx.x; bb < nb; bb += gridDim.x)
      for (int n = lane; n < p.N; n += 32) {
        const float v = p.cls[n] + p.pos[n];
        const size_t eo = (size_t)bb * rows_per_b * p.N + n;
        if (p.out_f32) reinterpret_cast<float*>(p.out)[eo] = v;
        else reinterpret_cast<__nv_bfloat16*>(p.out)[eo] = __float2bfloat16_rn(v);
      }
  }
  if (warp == 0) {
    // ===================== TMA producer (both CTAs; bytes are counted on the LEADER's full barrier) =====
    if (elect_one()) {
      uint32_t it = 0;
      for (int u = pair; u < num_units; u += num_pairs) {
        const int m_pair = u / p.n_tiles, n_blk = u - m_pair * p.n_tiles;
        const int row_a = m_pair * 2 * BM + (int)rank * BM;
        const int row_b = n_blk * p.BN + (int)rank * half_n;
        for (int kb = 0; kb < p.num_k_blocks; ++kb, ++it) {
          const int s = it % p.stages;
          mb_wait(&empty[s], ((it / p.stages) & 1u) ^ 1u);
          if (p.diag & 4) { if (leader) mb_arrive(&full[s]); continue; }
          if (leader) mb_expect_tx(&full[s], 2u * (a_bytes + b_bytes));
          const uint32_t bar = mapa(s32(&full[s]), 0);
          tma_load_2d_2sm(smem_a + (size_t)s * a_bytes, &map_a, kb * p.BK, row_a, bar);
          tma_load_2d_2sm(smem_b + (size_t)s * b_bytes, &map_b, kb * p.BK, row_b, bar);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && elect_one()) {
      const uint32_t idesc = make_idesc(2 * BM, p.BN);
      uint32_t it = 0, local = 0;
      for (int u = pair; u < num_units; u += num_pairs, ++local) {
        const uint32_t acc = local & 1u, acc_phase = (local >> 1) & 1u;
        mb_wait(&tmem_empty[acc], acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * (uint32_t)p.BN;
        for (int kb = 0; kb < p.num_k_blocks; ++kb, ++it) {
          const int s = it % p.stages;
          mb_wait(&full[s], (it / p.stages) & 1u);
          tc_fence_after();
          const uint64_t adesc = make_smem_desc(s32(smem_a + (size_t)s * a_bytes), span);
          const uint64_t bdesc = make_smem_desc(s32(smem_b + (size_t)s * b_bytes), span);
          for (int k = 0; k < p.BK / 16 && !(p.diag & 2); ++k)
            umma_bf16_2sm(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (uint32_t)((kb | k) != 0));
          umma_commit_2sm(&empty[s]);
        }
        umma_commit_2sm(&tmem_full[acc]);
      }
    }
  } else {
    // ===================== epilogue (both CTAs, own 128 rows; 8 warps) =====================
    // warp%4 selects the TMEM lane quarter (hardware rule), (warp-2)/4 the column half of the tile.
    // Per 128-byte fill: two TMEM loads are in flight while the previous fill is packed, written to
    // the warp's swizzled staging tile and handed to the TMA store engine.
    const int q = warp & 3, half = (warp - 2) >> 2;
    const uint32_t leader_tmem_empty0 = mapa(s32(&tmem_empty[0]), 0);
    const bool f32_fill = p.out_f32 || EMB;   // the fused-embeddings epilogue stages fp32
    const int cpf = f32_fill ? 32 : 64;      // columns per fill (128 bytes of staging)
    const int nfill = half_n / cpf;
    uint32_t local = 0;
    for (int u = pair; u < num_units; u += num_pairs, ++local) {
      const int m_pair = u / p.n_tiles, n_blk = u - m_pair * p.n_tiles;
      const uint32_t acc = local & 1u, acc_phase = (local >> 1) & 1u;
      mb_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * (uint32_t)p.BN + (uint32_t)(half * half_n);
      const int m_warp0 = m_pair * 2 * BM + (int)rank * BM + q * 32;
      const int n_base = n_blk * p.BN + half * half_n;
      float va[32], vb[32];
      uint32_t pk[32];
      if (p.diag & 16) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mb_arrive_cluster(leader_tmem_empty0 + acc * 8u);
        continue;
      }
      // fused embeddings: where this lane's token row goes and which table rows it adds
      //   ViViT        m = b*P + tok            -> out row b*(1+P) + 1 + tok,        + pos[1 + tok]
      //   TimeSformer  m = (b*T + t)*P + patch  -> out row b*(1+P*T) + 1 + patch*T + t, + pos[1 + patch] + time[t]
      int e_out = -1, e_pos = 0, e_t = 0;
      if (EMB) {
        const int m = m_warp0 + lane;
        if (m < p.M) {
          const int fr = m / p.emb_P, tok = m - fr * p.emb_P;
          e_pos = 1 + tok;
          if (p.emb_mode == 1) e_out = fr * (1 + p.emb_P) + 1 + tok;
          else { const int bb = fr / p.emb_T; e_t = fr - bb * p.emb_T; e_out = bb * (1 + p.emb_P * p.emb_T) + 1 + tok * p.emb_T + e_t; }
        }
      }
      // coalesced pass geometry: 8 lanes cover the 128 bytes of a staged row, 4 rows per step, 8 steps
      const int c4 = lane & 7;
      int orow_[8], prow_[8], trow_[8];
      if (EMB) {
#pragma unroll
        for (int itr = 0; itr < 8; ++itr) {
          const int r = itr * 4 + (lane >> 3);
          orow_[itr] = __shfl_sync(0xffffffffu, e_out, r);
          prow_[itr] = __shfl_sync(0xffffffffu, e_pos, r);
          trow_[itr] = __shfl_sync(0xffffffffu, e_t, r);
        }
      }
      tmem_ld32_async(taddr, va);
      if (!f32_fill) tmem_ld32_async(taddr + 32, vb);
      for (int f = 0; f < nfill; ++f) {
        const int n0 = n_base + f * cpf;
        tmem_wait_dep(va);
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + n0) + j4);
          va[4 * j4] += b4.x; va[4 * j4 + 1] += b4.y; va[4 * j4 + 2] += b4.z; va[4 * j4 + 3] += b4.w;
        }
        if (f32_fill) {
#pragma unroll
          for (int j = 0; j < 32; ++j) pk[j] = __float_as_uint(va[j]);
        } else {
          tmem_wait_dep(vb);
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
            const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + n0 + 32) + j4);
            vb[4 * j4] += b4.x; vb[4 * j4 + 1] += b4.y; vb[4 * j4 + 2] += b4.z; vb[4 * j4 + 3] += b4.w;
          }
          pack_bf16<0>(va, pk);
          pack_bf16<1>(vb, pk);
        }
        if (f + 1 < nfill) {   // next fill's accumulator columns travel while this one is stored
          tmem_ld32_async(taddr + (uint32_t)((f + 1) * cpf), va);
          if (!f32_fill) tmem_ld32_async(taddr + (uint32_t)((f + 1) * cpf + 32), vb);
        } else {               // accumulator fully read: hand the TMEM buffer back to the MMA issuer
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mb_arrive_cluster(leader_tmem_empty0 + acc * 8u);
        }
        if (!EMB && lane == 0) tma_store_wait_read();   // the previous store has finished reading the tile
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(stg + lane * 128 + ((j ^ (lane & 7)) << 4)) =
              make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
        if (!EMB) {
          fence_async_smem();
          __syncwarp();
          if (lane == 0 && !(p.diag & 1)) tma_store_2d(&map_o, stg, n0, m_warp0);
        } else {
        // fused embeddings: coalesced pass over the staged fp32 tile; position (+ time) rows are added in fp32,
        // rows scatter to their output slots.  All table loads of the pass are issued before the first use.
        float4 pe_[8], te_[8];
#pragma unroll
        for (int itr = 0; itr < 8; ++itr) {
          pe_[itr] = make_float4(0.f, 0.f, 0.f, 0.f);
          te_[itr] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (orow_[itr] >= 0) {
            pe_[itr] = __ldg(reinterpret_cast<const float4*>(p.pos + (size_t)prow_[itr] * p.N + n0) + c4);
            if (p.temb) te_[itr] = __ldg(reinterpret_cast<const float4*>(p.temb + (size_t)trow_[itr] * p.N + n0) + c4);
          }
        }
        __syncwarp();
#pragma unroll
        for (int itr = 0; itr < 8; ++itr) {
          const int r = itr * 4 + (lane >> 3);
          if (orow_[itr] < 0) continue;
          float4 z = *reinterpret_cast<const float4*>(stg + r * 128 + ((c4 ^ (r & 7)) << 4));
          z.x += pe_[itr].x + te_[itr].x; z.y += pe_[itr].y + te_[itr].y;
          z.z += pe_[itr].z + te_[itr].z; z.w += pe_[itr].w + te_[itr].w;
          const size_t eo = (size_t)orow_[itr] * p.N + n0 + c4 * 4;
          if (p.out_f32) {
            *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + eo) = z;
          } else {
            __nv_bfloat162 lo = __floats2bfloat162_rn(z.x, z.y), hi = __floats2bfloat162_rn(z.z, z.w);
            *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.out) + eo) =
                make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
          }
        }
        __syncwarp();   // the tile is rewritten by the next fill
        }
      }
    }
    if (lane == 0) tma_store_wait_all();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // nobody exits while the peer may still signal its barriers
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_2sm(tmem_base, p.tmem_cols);
  }
}

}  // namespace

// ------------------------------------------------- short-K GEMM + LayerNorm (Video Swin) ----
// PatchEmbed3d is HBM-bound (K = 96, N = 128: 55 FLOP/B), so this kernel is organised around the
// byte streams, not the MMAs: the whole weight matrix B[N][K] is loaded ONCE per CTA and stays in
// shared memory; A tiles (128 tokens x K) stream through a TMA ring, one mbarrier per tile; two
// epilogue warpgroups own one TMEM accumulator each and alternate tiles.  A thread holds its token's
// full row (N fp32 values) in registers: bias, two-pass LayerNorm statistics and the affine are done
// without re-reading TMEM, and rows leave through 128-byte-swizzled staging tiles and TMA stores.
namespace {

constexpr int kSwinGroups = 4;                           // epilogue warpgroups = TMEM accumulators (stand-alone kernel)
constexpr int kSwinThreads = 64 + kSwinGroups * 128;     // producer warp, MMA warp, 16 epilogue warps
__host__ __device__ inline int num_tiles_of(const GemmParams& p) { return p.m_tiles; }
constexpr int kSwinGroupsFused = 1;                      // the variant that shares its SM with the clip kernel: 6 warps
__host__ __device__ constexpr int swin_threads(int ng) { return 64 + ng * 128; }

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// NG = epilogue warpgroups = TMEM accumulators.  NG = 4: the stand-alone kernel (576 threads, the SM to itself).
// NG = 1: 192 threads, 96 registers, ~110 KB of shared memory -- small enough to be CO-RESIDENT with the clip kernel
// (lvc_clip_swin_embed_forward): it is launched behind the clip kernel with programmatic stream serialization, starts
// while the clip kernel runs, and its producer loads a tile of A when the clip kernel's ready counter says the tile
// is complete (p.ready).  It must then NOT execute griddepcontrol.wait (that would wait for the whole clip grid).
template <int NCH, bool OUT_F32, int NG = kSwinGroups>  // N = 32 * NCH columns, NG * N <= 512 TMEM columns
__global__ void __launch_bounds__(swin_threads(NG), (NG == kSwinGroupsFused ? 3 : 1))   // fused variant: <= 112 registers (it takes 96: 8 x 32 x 96 + the clip variant's 16 x 32 x 80 = 65 536)
lvc_embed_swin_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                      const __grid_constant__ CUtensorMap map_o, const __grid_constant__ GemmParams p) {
  constexpr int N = 32 * NCH;
  constexpr int kSwinThreads = swin_threads(NG);
  constexpr int NA = NG == 1 ? 2 : NG;   // TMEM accumulators: one per epilogue group; the single-group variant ping-pongs two
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (s32(smem_raw) & 1023u)) & 1023u);
  const int span = p.BK * 2;
  const uint32_t a_sub = BM * span, b_sub = N * span;          // one k-block of A / of B
  const uint32_t a_bytes = a_sub * p.num_k_blocks, b_bytes = b_sub * p.num_k_blocks;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* stg = smem + (size_t)((warp - 2) & (4 * NG - 1)) * 4096;   // one 4 KB store tile per epilogue warp
  uint8_t* smem_b = smem + 4 * NG * 4096;
  // bias rides in the MMA as one extra K=16 step: D += ONES[128 x 16] . BIASPARTS[N x 16]^T, where the three
  // leading K entries of row n hold a bf16 hi/mid/lo split of bias[n] (their fp32 sum is bias[n] to ~2^-24)
  uint8_t* smem_ones = smem_b + b_bytes;                 // 128 rows x 64 B, every element bf16 1.0
  uint8_t* smem_bias = smem_ones + BM * 64;              // N rows x 64 B, 64B-swizzled
  uint8_t* smem_a = smem_bias + N * 64;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_a + (size_t)p.stages * a_bytes);
  uint64_t* empty = full + p.stages;
  uint64_t* tmem_full = empty + p.stages;
  uint64_t* tmem_empty = tmem_full + NA;
  uint64_t* b_full = tmem_empty + NA;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(b_full + 1);
  // parameters after the barriers, 16-byte aligned for float4 reads (barrier block is < 512 bytes)
  float* s_bias = reinterpret_cast<float*>(smem_a + (size_t)p.stages * a_bytes + 512);
  float* s_lnw = s_bias + N;
  float* s_lnb = s_lnw + N;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    tma_prefetch_desc(&map_o);
    for (int s = 0; s < p.stages; ++s) { mb_init(&full[s], 1); mb_init(&empty[s], 1); }
    for (int a = 0; a < NA; ++a) { mb_init(&tmem_full[a], 1); mb_init(&tmem_empty[a], 4); }
    mb_init(b_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, p.tmem_cols);
  // barrier set-up and the TMEM allocation overlap the previous kernel's tail; parameters do not.  Gated on ready
  // counters the kernel runs BESIDE its producer kernel: the counters, not grid completion, order the reads of A
  // (bias / LayerNorm parameters and the weights were written before the producer kernel was launched).
  if (!p.ready) griddep_wait();
  unsigned long long* stamps = p.ready ? reinterpret_cast<unsigned long long*>(const_cast<unsigned*>(p.ready) + ((num_tiles_of(p) + 63) / 64) * 64) : nullptr;
  if (stamps && blockIdx.x == 0 && threadIdx.x == 0) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); stamps[2] = t; }
  for (int i = threadIdx.x; i < N; i += kSwinThreads) {
    s_lnw[i] = p.do_ln ? p.ln_w[i] : 1.f;
    s_lnb[i] = p.do_ln ? p.ln_b[i] : 0.f;
    const float b = p.bias[i];
    const __nv_bfloat16 hi = __float2bfloat16_rn(b);
    const float r1 = b - __bfloat162float(hi);
    const __nv_bfloat16 mid = __float2bfloat16_rn(r1);
    const __nv_bfloat16 lo = __float2bfloat16_rn(r1 - __bfloat162float(mid));
    uint4* row = reinterpret_cast<uint4*>(smem_bias + i * 64);
    const int c0 = (i >> 1) & 3;   // SWIZZLE_64B: 16-byte chunk c of row i lives at chunk c ^ ((i >> 1) & 3)
#pragma unroll
    for (int c = 0; c < 4; ++c) row[c] = make_uint4(0u, 0u, 0u, 0u);
    row[c0] = make_uint4((uint32_t)__bfloat16_as_ushort(hi) | ((uint32_t)__bfloat16_as_ushort(mid) << 16),
                         (uint32_t)__bfloat16_as_ushort(lo), 0u, 0u);
  }
  for (int i = threadIdx.x; i < BM * 64 / 16; i += kSwinThreads)
    reinterpret_cast<uint4*>(smem_ones)[i] = make_uint4(0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u);
  fence_async_smem();   // generic-proxy writes above are read by the tensor core (async proxy)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int num_tiles = p.m_tiles;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      mb_expect_tx(b_full, b_bytes);
      for (int kb = 0; kb < p.num_k_blocks; ++kb) tma_load_2d(smem_b + (size_t)kb * b_sub, &map_b, kb * p.BK, 0, b_full);
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int s = it % p.stages;
        if (p.ready) mb_wait_backoff(&empty[s], ((it / p.stages) & 1u) ^ 1u); else mb_wait(&empty[s], ((it / p.stages) & 1u) ^ 1u);
        if (p.diag & 4) { mb_arrive(&full[s]); continue; }
        if (p.ready) {   // the clip kernel is still writing A: wait for this tile's token rows, then order the TMA read
          const unsigned need = (unsigned)min(BM, p.M - tile * BM) * p.ready_need;   // the last tile may be partial
          while (ld_acquire_gpu(p.ready + tile) < need) __nanosleep(200);
          asm volatile("fence.proxy.async;" ::: "memory");   // generic-proxy writes (other SMs) -> async-proxy read
        }
        mb_expect_tx(&full[s], a_bytes);
        if (p.diag & 128) {  // timing experiment: the tile as ONE linear bulk copy (layout is wrong for the MMA)
          const uint8_t* src = reinterpret_cast<const uint8_t*>(p.diag_a) + (size_t)tile * a_bytes;
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                       ::"r"(s32(smem_a + (size_t)s * a_bytes)), "l"(src), "r"(a_bytes), "r"(s32(&full[s])) : "memory");
          continue;
        }
        for (int kb = 0; kb < p.num_k_blocks; ++kb)
          tma_load_2d(smem_a + (size_t)s * a_bytes + (size_t)kb * a_sub, &map_a, kb * p.BK, tile * BM, &full[s]);
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      const uint32_t idesc = make_idesc(BM, N);
      mb_wait(b_full, 0);
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const uint32_t acc = it % NA, acc_phase = (it / NA) & 1u;
        const int s = it % p.stages;
        if (p.ready) {   // beside the clip kernel: tiles arrive at ITS pace, and a bare spin would take issue slots from its warps
          mb_wait_backoff(&tmem_empty[acc], acc_phase ^ 1u);
          mb_wait_backoff(&full[s], (it / p.stages) & 1u);
        } else {
          mb_wait(&tmem_empty[acc], acc_phase ^ 1u);
          mb_wait(&full[s], (it / p.stages) & 1u);
        }
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * (uint32_t)N;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          const uint64_t adesc = make_smem_desc(s32(smem_a + (size_t)s * a_bytes + (size_t)kb * a_sub), span);
          const uint64_t bdesc = make_smem_desc(s32(smem_b + (size_t)kb * b_sub), span);
          for (int k = 0; k < p.BK / 16 && !(p.diag & 2); ++k)
            umma_bf16(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (uint32_t)((kb | k) != 0));
        }
        umma_commit(&empty[s]);
        umma_bf16(d_tmem, make_smem_desc(s32(smem_ones), 64), make_smem_desc(s32(smem_bias), 64), idesc, 1u);  // + bias
        umma_commit(&tmem_full[acc]);
      }
    }
  } else {
    // ===================== epilogue: warpgroup g = (warp-2)/4 owns accumulator g =====================
    // A thread owns one token row.  Pass 1 over TMEM: bias + sum / sum of squares (four independent
    // chains each); pass 2: normalise, affine, pack, 128-byte fills through the warp's swizzled store
    // tile.  The next chunk's tcgen05.ld is always in flight while the current one is processed.
    const int q = warp & 3, grp = (warp - 2) >> 2;
    const float inv_n = 1.f / (float)N;
    constexpr int CPF = OUT_F32 ? 1 : 2;  // 32-column chunks per 128-byte fill
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      if ((int)(it % NG) != grp) continue;
      const int acc = (int)(it % NA);   // == grp when every group owns an accumulator
      if (p.diag & 256) mb_wait(&tmem_full[acc], (it / NA) & 1u); else mb_wait_backoff(&tmem_full[acc], (it / NA) & 1u);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * N);
      if (p.diag & 16) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mb_arrive(&tmem_empty[acc]);
        continue;
      }
      float v[2][32];
      float mean = 0.f, rstd = 1.f;
      if (p.do_ln && !(p.diag & 32)) {
        // pass 1: sum and sum of squares, two packed chains each (4 scalar chains)
        uint64_t s2[2] = {0ull, 0ull}, q2[2] = {0ull, 0ull};
        tmem_ld32_async(taddr, v[0]);
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          tmem_wait_dep(v[c & 1]);
          if (c + 1 < NCH) tmem_ld32_async(taddr + (c + 1) * 32, v[(c + 1) & 1]);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const uint64_t z = pack2(v[c & 1][2 * j], v[c & 1][2 * j + 1]);
            s2[j & 1] = add2(s2[j & 1], z);
            q2[j & 1] = fma2(z, z, q2[j & 1]);
          }
        }
        float sa, sb, qa, qb;
        unpack2(add2(s2[0], s2[1]), sa, sb);
        unpack2(add2(q2[0], q2[1]), qa, qb);
        mean = (sa + sb) * inv_n;
        rstd = rsqrtf(fmaxf((qa + qb) * inv_n - mean * mean, 0.f) + p.ln_eps);
      }
      const int m0 = tile * BM + q * 32;
      const uint64_t rstd2 = pack2(rstd, rstd), nmr2 = pack2(-mean * rstd, -mean * rstd);
      uint32_t pk[32];
      tmem_ld32_async(taddr, v[0]);
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        tmem_wait_dep(v[c & 1]);
        if (c + 1 < NCH) {
          tmem_ld32_async(taddr + (c + 1) * 32, v[(c + 1) & 1]);
        } else {   // accumulator fully read: the MMA issuer may overwrite it
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mb_arrive(&tmem_empty[acc]);
        }
        const int cc = c % CPF;
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 g4 = reinterpret_cast<const float4*>(s_lnw + c * 32)[j4];
          const float4 o4 = reinterpret_cast<const float4*>(s_lnb + c * 32)[j4];
          // (z - mean) * rstd * gamma + beta  ==  fma(fma(z, rstd, -mean*rstd), gamma, beta), z = acc (bias included)
          const uint64_t ya = fma2(fma2(pack2(v[c & 1][4 * j4], v[c & 1][4 * j4 + 1]), rstd2, nmr2), pack2(g4.x, g4.y), pack2(o4.x, o4.y));
          const uint64_t yb = fma2(fma2(pack2(v[c & 1][4 * j4 + 2], v[c & 1][4 * j4 + 3]), rstd2, nmr2), pack2(g4.z, g4.w), pack2(o4.z, o4.w));
          float y0, y1, y2, y3;
          unpack2(ya, y0, y1);
          unpack2(yb, y2, y3);
          if (OUT_F32) {
            pk[4 * j4] = __float_as_uint(y0); pk[4 * j4 + 1] = __float_as_uint(y1);
            pk[4 * j4 + 2] = __float_as_uint(y2); pk[4 * j4 + 3] = __float_as_uint(y3);
          } else {
            __nv_bfloat162 lo = __floats2bfloat162_rn(y0, y1), hi = __floats2bfloat162_rn(y2, y3);
            pk[cc * 16 + 2 * j4] = *reinterpret_cast<uint32_t*>(&lo);
            pk[cc * 16 + 2 * j4 + 1] = *reinterpret_cast<uint32_t*>(&hi);
          }
        }
        if (cc == CPF - 1) {   // a 128-byte fill is complete
          if (lane == 0) tma_store_wait_read();   // the tile's previous store has been read out
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<uint4*>(stg + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
          fence_async_smem();
          __syncwarp();
          if (lane == 0 && !(p.diag & 1)) tma_store_2d(&map_o, stg, (c + 1 - CPF) * 32, m0);
        }
      }
    }
    if (lane == 0) tma_store_wait_all();
  }
  tc_fence_before();
  __syncthreads();
  if (stamps && blockIdx.x == 0 && threadIdx.x == 0) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); stamps[3] = t; }
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

}  // namespace

// ------------------------------------------------------------------------------ host ----
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// [rows][K] bf16 row-major, box = BK x box_rows, swizzle = BK*2 bytes
int make_map(CUtensorMap* map, const void* base, uint64_t rows, uint64_t K, uint32_t BK, uint32_t box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return lvc_fail(LVC_E_CUDA, "patch_embed: cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t dims[2] = {K, rows};
  cuuint64_t strides[1] = {K * 2};
  cuuint32_t box[2] = {BK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, BK == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return lvc_fail(LVC_E_CUDA, "patch_embed: cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return LVC_OK;
}

// out[M][N] (bf16 or fp32) row-major, box = 128 bytes x 32 rows, SWIZZLE_128B: the epilogue's store tile
int make_out_map(CUtensorMap* map, void* base, uint64_t rows, uint64_t N, bool f32) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return lvc_fail(LVC_E_CUDA, "patch_embed: cuTensorMapEncodeTiled is not available from the driver");
  const uint64_t esz = f32 ? 4 : 2;
  cuuint64_t dims[2] = {N, rows};
  cuuint64_t strides[1] = {N * esz};
  cuuint32_t box[2] = {(cuuint32_t)(128 / esz), 32};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return lvc_fail(LVC_E_CUDA, "patch_embed: cuTensorMapEncodeTiled(out) failed with CUresult %d", (int)r);
  return LVC_OK;
}

int env_i(const char* name, int dflt) { return lvc_knob_int(name, dflt); }   // gated: lvc_internal.h

// Launch with programmatic stream serialization: the grid may be scheduled while the previous kernel in the stream
// (the clip kernel of a pipeline, the im2col pass) drains; the kernels call griddep_wait() before touching global data.
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kern)(KArgs...), int grid, int block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = env_i("LVC_NO_PDL", 0) ? 0 : 1;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

}  // namespace

extern "C" int64_t lvc_patch_embed_workspace(const lvc_embed_desc* d) {
  EmbedGeom g;
  if (make_geom(d, &g) != LVC_OK) return -1;
  return ((int64_t)g.M * g.K * 2 + 255) / 256 * 256;
}

extern "C" int lvc_patch_embed_launch_count(const lvc_embed_desc* d) { return d ? (d->x_layout == LVC_LAYOUT_TOKENS ? 1 : 2) : 0; }

extern "C" int lvc_patch_embed_pack_weights(const lvc_embed_desc* d, const float* w_f32, void* w_packed, void* stream) {
  EmbedGeom g;
  int rc = make_geom(d, &g);
  if (rc != LVC_OK) return rc;
  if (!w_f32 || !w_packed) return lvc_fail(LVC_E_INVAL, "patch_embed_pack_weights: NULL pointer");
  const long long n = 1LL * g.N * g.K;
  lvc_pack_w_kernel<<<(int)std::min<long long>((n + 255) / 256, 1184), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      w_f32, static_cast<__nv_bfloat16*>(w_packed), n);
  LVC_CUDA_TRY(cudaGetLastError());
  return LVC_OK;
}

namespace {
struct EmbTables { const float* cls; const float* pos; const float* temb; int mode; };
int patch_embed_impl(const lvc_embed_desc* d, const void* x, const void* w_packed, const float* bias,
                     const float* ln_w, const float* ln_b, const EmbTables& et, void* out, void* workspace, void* stream);
}  // namespace

extern "C" int lvc_patch_embed(const lvc_embed_desc* d, const void* x, const void* w_packed, const float* bias,
                               const float* ln_w, const float* ln_b, void* out, void* workspace, void* stream) {
  return patch_embed_impl(d, x, w_packed, bias, ln_w, ln_b, EmbTables{nullptr, nullptr, nullptr, 0}, out, workspace, stream);
}

extern "C" int lvc_embeddings_forward(const lvc_embed_desc* d, const void* x, const void* w_packed, const float* bias,
                                      const float* cls_token, const float* position_embeddings,
                                      const float* time_embeddings, void* out, void* workspace, void* stream) {
  if (!d) return lvc_fail(LVC_E_INVAL, "embeddings_forward: NULL descriptor");
  if (d->kind != LVC_EMBED_VIVIT && d->kind != LVC_EMBED_TIMESFORMER)
    return lvc_fail(LVC_E_NOTSUP, "embeddings_forward: ViViT and TimeSformer only (Video Swin has no cls / position tables)");
  if (!cls_token || !position_embeddings) return lvc_fail(LVC_E_INVAL, "embeddings_forward: NULL cls_token / position_embeddings");
  if (d->kind == LVC_EMBED_VIVIT && time_embeddings) return lvc_fail(LVC_E_INVAL, "embeddings_forward: ViViT has no time embeddings");
  return patch_embed_impl(d, x, w_packed, bias, nullptr, nullptr,
                          EmbTables{cls_token, position_embeddings, time_embeddings, d->kind == LVC_EMBED_VIVIT ? 1 : 2},
                          out, workspace, stream);
}

namespace {
int patch_embed_impl(const lvc_embed_desc* d, const void* x, const void* w_packed, const float* bias,
                     const float* ln_w, const float* ln_b, const EmbTables& et, void* out, void* workspace, void* stream) {
  EmbedGeom g;
  int rc = make_geom(d, &g);
  if (rc != LVC_OK) return rc;
  if (!x || !w_packed || !bias || !out || (!workspace && d->x_layout != LVC_LAYOUT_TOKENS)) return lvc_fail(LVC_E_INVAL, "patch_embed: NULL pointer");
  if (d->kind == LVC_EMBED_SWIN && (!ln_w || !ln_b)) return lvc_fail(LVC_E_INVAL, "patch_embed: Swin needs LayerNorm weight and bias");
  if (d->x_dtype != LVC_DT_F32 && d->x_dtype != LVC_DT_BF16) return lvc_fail(LVC_E_INVAL, "patch_embed: x_dtype must be f32 or bf16");
  if (d->out_dtype != LVC_DT_F32 && d->out_dtype != LVC_DT_BF16) return lvc_fail(LVC_E_INVAL, "patch_embed: out_dtype must be f32 or bf16");
  const bool tokens_in = d->x_layout == LVC_LAYOUT_TOKENS;
  if (d->x_layout != LVC_LAYOUT_BTCHW && d->x_layout != LVC_LAYOUT_BCTHW && !tokens_in) return lvc_fail(LVC_E_INVAL, "patch_embed: x_layout must be BTCHW, BCTHW or TOKENS");
  if (tokens_in) {
    if (d->x_dtype != LVC_DT_BF16) return lvc_fail(LVC_E_INVAL, "patch_embed: token input must be bf16");
    if (reinterpret_cast<uintptr_t>(x) & 127) return lvc_fail(LVC_E_ALIGN, "patch_embed: token input needs 128-byte alignment");
    if (g.T % g.kt || g.H % g.kh || g.W % g.kw) return lvc_fail(LVC_E_SHAPE, "patch_embed: token input cannot carry padding");
    workspace = const_cast<void*>(x);  // the GEMM reads A straight from the clip kernel's output
  }
  if ((reinterpret_cast<uintptr_t>(workspace) & 127) || (reinterpret_cast<uintptr_t>(w_packed) & 127) || (reinterpret_cast<uintptr_t>(out) & 15))
    return lvc_fail(LVC_E_ALIGN, "patch_embed: workspace / packed weights need 128-byte, out 16-byte alignment");
  if (g.N % 32) return lvc_fail(LVC_E_SHAPE, "patch_embed: embed_dim %d must be a multiple of 32", g.N);
  cudaStream_t st = static_cast<cudaStream_t>(stream);

  int dev = 0, sms = 0, max_smem = 0;
  LVC_CUDA_TRY(cudaGetDevice(&dev));
  LVC_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  LVC_CUDA_TRY(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));

  // 1) im2col: x -> A[M][K] bf16 (skipped when the clip kernel already wrote the tokens)
  if (!tokens_in) {
    const long long runs = 1LL * g.M * 3 * g.kt * g.kh;
    const int blocks = (int)std::min<long long>((runs + 255) / 256, (long long)sms * 16);
    const int bcthw = d->x_layout == LVC_LAYOUT_BCTHW;
    const size_t esz = d->x_dtype == LVC_DT_F32 ? 4 : 2;
    const bool vec_ok = g.T % g.kt == 0 && g.H % g.kh == 0 && g.W % g.kw == 0 && (g.W * esz) % 16 == 0 &&
                        (reinterpret_cast<uintptr_t>(x) & 15) == 0 && runs < (1LL << 31) && !lvc_knob_set("LVC_IM2COL_SCALAR");
    __nv_bfloat16* A = static_cast<__nv_bfloat16*>(workspace);
    if (vec_ok && g.kw == 16) {
      const int nb = (int)((runs + 255) / 256);
      if (d->x_dtype == LVC_DT_F32) lvc_im2col_vec_kernel<float, 16, 1><<<nb, 256, 0, st>>>(static_cast<const float*>(x), A, g, bcthw);
      else lvc_im2col_vec_kernel<__nv_bfloat16, 16, 1><<<nb, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), A, g, bcthw);
    } else if (vec_ok && g.kw == 4 && g.kh == 4) {
      const int nb = (int)((runs / 4 + 255) / 256);
      if (d->x_dtype == LVC_DT_F32) lvc_im2col_vec_kernel<float, 4, 4><<<nb, 256, 0, st>>>(static_cast<const float*>(x), A, g, bcthw);
      else lvc_im2col_vec_kernel<__nv_bfloat16, 4, 4><<<nb, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), A, g, bcthw);
    } else if (d->x_dtype == LVC_DT_F32)
      lvc_im2col_kernel<float><<<blocks, 256, 0, st>>>(static_cast<const float*>(x), static_cast<__nv_bfloat16*>(workspace), g, bcthw);
    else
      lvc_im2col_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(workspace), g, bcthw);
    LVC_CUDA_TRY(cudaGetLastError());
  }
  // 2) GEMM
  GemmParams p{};
  p.M = g.M; p.N = g.N; p.K = g.K;
  p.BN = g.N <= 256 ? g.N : (g.N % 256 == 0 ? 256 : 128);
  p.BK = (g.K % 64 == 0) ? 64 : 32;
  p.num_k_blocks = g.K / p.BK;
  p.m_tiles = (g.M + BM - 1) / BM;
  p.n_tiles = g.N / p.BN;
  p.tmem_cols = 32;
  while (p.tmem_cols < 2u * p.BN) p.tmem_cols <<= 1;
  p.bias = bias; p.ln_w = ln_w; p.ln_b = ln_b; p.ln_eps = d->ln_eps;
  p.do_ln = d->kind == LVC_EMBED_SWIN;
  p.out_f32 = d->out_dtype == LVC_DT_F32;
  p.out = out;
  // Directed-experiment knobs (scripts/gemm_probe.py): results are WRONG with any bit set and some combinations
  // (2 | 4 with the embeddings epilogue) never terminate, so they need a second, explicit opt-in.
  p.diag = env_i("LVC_GEMM_DIAG", 0);
  p.diag_a = workspace;
  p.cls = et.cls; p.pos = et.pos; p.temb = et.temb; p.emb_mode = et.mode;
  p.emb_P = g.Ho * g.Wo * (et.mode == 1 ? g.To : 1);   // tokens per batch entry (ViViT) / per frame (TimeSformer)
  p.emb_T = g.To;
  if (et.mode && ((reinterpret_cast<uintptr_t>(et.pos) & 15) || (et.temb && (reinterpret_cast<uintptr_t>(et.temb) & 15))))
    return lvc_fail(LVC_E_ALIGN, "embeddings_forward: position / time tables need 16-byte alignment");
  const int stage_bytes = (BM + p.BN) * p.BK * 2;
  const int fixed_bytes = 1024 + 2048 + (g.N + 2 * p.BN) * 4 + 16384;  // alignment, barriers, params, store staging
  // two CTAs per SM when TMEM (2 x <=256 columns) and shared memory allow: doubles the epilogue warps,
  // which is what bounds the HBM-bound Swin shape
  int per_sm = (p.tmem_cols <= 256 && !lvc_knob_set("LVC_GEMM_ONE_PER_SM")) ? 2 : 1;
  int sm_smem = 0;
  LVC_CUDA_TRY(cudaDeviceGetAttribute(&sm_smem, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev));
  for (;; per_sm = 1) {
    const int budget = per_sm == 2 ? std::min(max_smem, sm_smem / 2 - 1024) : max_smem;
    p.stages = std::min({env_i("LVC_GEMM_STAGES", 8), (budget - fixed_bytes) / stage_bytes, std::max(2, p.num_k_blocks * 2)});
    if (per_sm == 1 || p.stages >= std::min(3, p.num_k_blocks + 1)) break;
  }
  if (p.stages < 2) return lvc_fail(LVC_E_SHAPE, "patch_embed: tile does not fit shared memory");
  const size_t smem = (size_t)fixed_bytes + (size_t)p.stages * stage_bytes;

  // tensor-bound shapes (no LayerNorm, 128B-swizzle K blocks, even tile split): CTA-pair kernel
  const bool two_cta = !p.do_ln && p.BK == 64 && p.BN % 128 == 0 && !lvc_knob_set("LVC_GEMM_1CTA") && sms >= 2;
  if (et.mode && !two_cta) return lvc_fail(LVC_E_NOTSUP, "embeddings_forward: needs K %% 64 == 0 and embed_dim %% 128 == 0");
  if (two_cta) {
    const int stage2 = (BM + p.BN / 2) * p.BK * 2;
    const int fixed2 = 1024 + 8 * 4096 + 512;  // alignment slack, 8 store tiles, barriers + TMEM slot
    p.m_tiles = (g.M + 2 * BM - 1) / (2 * BM);
    p.stages = std::min({env_i("LVC_GEMM_STAGES", 8), (max_smem - fixed2) / stage2, std::max(2, p.num_k_blocks * 2)});
    if (p.stages < 2) return lvc_fail(LVC_E_SHAPE, "patch_embed: tile does not fit shared memory");
    const size_t smem2 = (size_t)fixed2 + (size_t)p.stages * stage2;
    CUtensorMap ma, mb, mo;
    rc = make_map(&ma, workspace, (uint64_t)g.M, (uint64_t)g.K, (uint32_t)p.BK, BM);
    if (rc != LVC_OK) return rc;
    rc = make_map(&mb, w_packed, (uint64_t)g.N, (uint64_t)g.K, (uint32_t)p.BK, (uint32_t)(p.BN / 2));
    if (rc != LVC_OK) return rc;
    rc = make_out_map(&mo, out, (uint64_t)g.M, (uint64_t)g.N, p.out_f32 != 0);
    if (rc != LVC_OK) return rc;
    const int pairs = std::min(sms / 2, p.m_tiles * p.n_tiles);
    if (et.mode) {
      LVC_CUDA_TRY(cudaFuncSetAttribute(lvc_embed_gemm2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
      LVC_CUDA_TRY(launch_pdl(lvc_embed_gemm2_kernel<true>, 2 * pairs, kGemm2Threads, smem2, st, ma, mb, mo, p));
    } else {
      LVC_CUDA_TRY(cudaFuncSetAttribute(lvc_embed_gemm2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
      LVC_CUDA_TRY(launch_pdl(lvc_embed_gemm2_kernel<false>, 2 * pairs, kGemm2Threads, smem2, st, ma, mb, mo, p));
    }
    LVC_CUDA_TRY(cudaGetLastError());
    return LVC_OK;
  }
  // short-K shapes with N = 128 (Video Swin-B): weights resident in shared memory, row-in-registers LayerNorm
  if (g.N == 128 && g.K <= 256 && !lvc_knob_set("LVC_GEMM_GENERIC")) {
    const int a_tile = BM * g.K * 2, b_all = g.N * g.K * 2;
    const int fixed3 = 1024 + 4 * kSwinGroups * 4096 + b_all + (BM + g.N) * 64 + 512 + 3 * g.N * 4 + 64;
    p.stages = std::min(env_i("LVC_GEMM_STAGES", 8), (max_smem - fixed3) / a_tile);
    if (p.stages >= 2) {
      const size_t smem3 = (size_t)fixed3 + (size_t)p.stages * a_tile;
      CUtensorMap ma, mb, mo;
      rc = make_map(&ma, workspace, (uint64_t)g.M, (uint64_t)g.K, (uint32_t)p.BK, BM);
      if (rc != LVC_OK) return rc;
      rc = make_map(&mb, w_packed, (uint64_t)g.N, (uint64_t)g.K, (uint32_t)p.BK, (uint32_t)g.N);
      if (rc != LVC_OK) return rc;
      rc = make_out_map(&mo, out, (uint64_t)g.M, (uint64_t)g.N, p.out_f32 != 0);
      if (rc != LVC_OK) return rc;
      const int grid3 = std::min(sms, p.m_tiles);
      if (lvc_knob_set("LVC_SWIN_NG1")) {   // profiling: the 192-thread variant of the fused path, stand-alone and ungated
        const int fixed1 = 1024 + 4 * kSwinGroupsFused * 4096 + b_all + (BM + g.N) * 64 + 512 + 3 * g.N * 4 + 64;
        p.stages = env_i("LVC_FUSED_EMBED_STAGES", 2);
        p.tmem_cols = 256;
        const size_t smem1 = (size_t)fixed1 + (size_t)p.stages * a_tile;
        LVC_CUDA_TRY(cudaFuncSetAttribute(lvc_embed_swin_kernel<4, false, kSwinGroupsFused>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
        LVC_CUDA_TRY(launch_pdl(lvc_embed_swin_kernel<4, false, kSwinGroupsFused>, grid3, swin_threads(kSwinGroupsFused), smem1, st, ma, mb, mo, p));
        return LVC_OK;
      }
      p.tmem_cols = 512;  // kSwinGroups accumulators of N columns
      if (p.out_f32) {
        LVC_CUDA_TRY(cudaFuncSetAttribute(lvc_embed_swin_kernel<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem3));
        LVC_CUDA_TRY(launch_pdl(lvc_embed_swin_kernel<4, true>, grid3, kSwinThreads, smem3, st, ma, mb, mo, p));
      } else {
        LVC_CUDA_TRY(cudaFuncSetAttribute(lvc_embed_swin_kernel<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem3));
        LVC_CUDA_TRY(launch_pdl(lvc_embed_swin_kernel<4, false>, grid3, kSwinThreads, smem3, st, ma, mb, mo, p));
      }
      LVC_CUDA_TRY(cudaGetLastError());
      return LVC_OK;
    }
  }
  CUtensorMap map_a, map_b;
  rc = make_map(&map_a, workspace, (uint64_t)g.M, (uint64_t)g.K, (uint32_t)p.BK, BM);
  if (rc != LVC_OK) return rc;
  rc = make_map(&map_b, w_packed, (uint64_t)g.N, (uint64_t)g.K, (uint32_t)p.BK, (uint32_t)p.BN);
  if (rc != LVC_OK) return rc;
  LVC_CUDA_TRY(cudaFuncSetAttribute(lvc_embed_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int grid = std::min(sms * per_sm, p.m_tiles * p.n_tiles);
  lvc_embed_gemm_kernel<<<grid, kGemmThreads, smem, st>>>(map_a, map_b, p);
  LVC_CUDA_TRY(cudaGetLastError());
  return LVC_OK;
}
}  // namespace

// ---------------------------------------------------------------------------------------------------------------
// lvc_patch_embed_static: lvc_patch_embed for callers that cannot allocate (CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).  The patch matrix A[M][K] lives in a file-scope device array instead of a
// caller workspace; every GPU that loads the library holds LVC_STATIC_SCRATCH_BYTES of it until the process exits.
// The array's name is a host-side stand-in in host code, so the device address is taken once per device with
// cudaGetSymbolAddress (that device current) and cached.  No cudaMalloc / cudaFree / allocating routine is called.
__device__ __align__(256) static unsigned char g_static_scratch[LVC_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxStaticDevices = 64;
std::atomic<void*> g_static_scratch_ptr[kMaxStaticDevices];

int static_scratch_address(void** out) {
  int dev = -1;
  LVC_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxStaticDevices) return lvc_fail(LVC_E_NOTSUP, "patch_embed_static: device ordinal %d out of range", dev);
  void* p = g_static_scratch_ptr[dev].load(std::memory_order_acquire);
  if (!p) {
    LVC_CUDA_TRY(cudaGetSymbolAddress(&p, g_static_scratch));   // address on the CURRENT device (= dev)
    if (!p || (reinterpret_cast<uintptr_t>(p) & 255))
      return lvc_fail(LVC_E_CUDA, "patch_embed_static: cudaGetSymbolAddress returned an unusable address %p", p);
    g_static_scratch_ptr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return LVC_OK;
}
}  // namespace

extern "C" int64_t lvc_patch_embed_static_capacity(void) { return (int64_t)LVC_STATIC_SCRATCH_BYTES; }

extern "C" int lvc_patch_embed_static(const lvc_embed_desc* d, const void* x, const void* w_packed, const float* bias,
                                      const float* ln_w, const float* ln_b, void* out, void* stream) {
  if (!d) return lvc_fail(LVC_E_INVAL, "patch_embed_static: NULL descriptor");
  const int64_t need = lvc_patch_embed_workspace(d);
  if (need < 0) return LVC_E_INVAL;   // message already recorded by the geometry check
  if (d->x_layout != LVC_LAYOUT_TOKENS && need > (int64_t)LVC_STATIC_SCRATCH_BYTES)
    return lvc_fail(LVC_E_SHAPE, "patch_embed_static: the patch matrix needs %lld bytes of scratch, the built-in array holds "
                    "%lld (LVC_STATIC_SCRATCH_BYTES); use lvc_patch_embed with a caller workspace",
                    (long long)need, (long long)LVC_STATIC_SCRATCH_BYTES);
  void* scratch = nullptr;
  int rc = static_scratch_address(&scratch);
  if (rc != LVC_OK) return rc;
  // Nothing has to start at zero: the im2col pass overwrites all M*K elements of A before the GEMM reads them, and the
  // tensor map is built on exactly M rows (rows past M are zero-filled by TMA, never read from the array).
  return patch_embed_impl(d, x, w_packed, bias, ln_w, ln_b, EmbTables{nullptr, nullptr, nullptr, 0}, out, scratch, stream);
}

// ---------------------------------------------------------------------------------------------------------------
// lvc_clip_swin_embed_forward: the Video Swin pipeline (clip kernel -> PatchEmbed3d GEMM + LayerNorm) with the two
// kernels CO-RESIDENT on every SM.  The clip stage of ROI batches is bound by its consumers' latency (half the issue
// slots idle, 2 of 6.5 TB/s of DRAM) and the embedding is bound by HBM, so run in sequence they waste each other's
// resource; side by side the embedding hides behind the clip stage.  Mechanics: (1) the clip kernel runs in a variant
// with two producer warps, 80 registers and its shared memory capped so that a 192-thread / 96-register / 110 KB embedding
// CTA fits on the same SM (registers are allocated per four warps: 16 x 32 x 80 + 8 x 32 x 96 = 65 536); (2) it counts, per 128-token tile of A, the
// token-row slices it has stored (red.release.gpu); (3) the embedding kernel is launched behind it with programmatic
// stream serialization -- the clip kernel executes griddepcontrol.launch_dependents on entry, so the embedding grid
// starts as soon as every clip CTA is resident -- and its TMA producer loads a tile when the tile's counter is complete
// (ld.acquire.gpu + fence.proxy.async).  The clip kernel never waits for the embedding kernel, so if the two do not
// fit together (another kernel holds the SM) the embedding simply runs afterwards: correctness does not depend on
// co-residency.  A stays in L2 (77 MB of 126 MB) between the kernels' SMs.  Results are bit-identical to
// lvc_clip_forward + lvc_patch_embed (same arithmetic in both kernels).
extern "C" int lvc_clip_swin_embed_workspace(const lvc_embed_desc* ed, int64_t* tokens_bytes, int64_t* ready_bytes) {
  EmbedGeom g;
  if (!ed || !tokens_bytes || !ready_bytes) return lvc_fail(LVC_E_INVAL, "clip_swin_embed_workspace: NULL argument");
  int rc = make_geom(ed, &g);
  if (rc != LVC_OK) return rc;
  *tokens_bytes = ((int64_t)g.M * g.K * 2 + 255) / 256 * 256;
  // per-tile counters, then 256 bytes of timestamps (CTA 0 of either kernel: start / end, globaltimer ns) for scripts/perf_swin_fused.py
  *ready_bytes = (((int64_t)g.M + BM - 1) / BM * 4 + 255) / 256 * 256 + 256;
  return LVC_OK;
}

extern "C" int lvc_clip_swin_embed_forward(const lvc_clip_desc* cd, const lvc_video* videos, const int32_t* indices,
                                           const int32_t* boxes, const float* lut, const lvc_embed_desc* ed,
                                           const void* w_packed, const float* bias, const float* ln_w, const float* ln_b,
                                           void* tokens, void* ready, void* out, void* stream) {
  if (!cd || !ed || !videos || !indices || !lut || !w_packed || !bias || !ln_w || !ln_b || !tokens || !ready || !out)
    return lvc_fail(LVC_E_INVAL, "clip_swin_embed: NULL argument");
  EmbedGeom g;
  int rc = make_geom(ed, &g);
  if (rc != LVC_OK) return rc;
  if (ed->kind != LVC_EMBED_SWIN || g.N != 128 || g.K > 256 || g.K % 32)
    return lvc_fail(LVC_E_NOTSUP, "clip_swin_embed: Video Swin PatchEmbed3d shapes only (embed_dim 128, K <= 256)");
  if (cd->layout != LVC_LAYOUT_TOKENS || cd->tok_kt != g.kt || cd->tok_kh != g.kh || cd->tok_kw != g.kw ||
      cd->batch != ed->batch || cd->num_frames != ed->frames || ed->height != 224 || ed->width != 224)
    return lvc_fail(LVC_E_SHAPE, "clip_swin_embed: the clip descriptor's token layout must be the embedding's patch (%d,%d,%d) on (B,T,224,224)", g.kt, g.kh, g.kw);
  if (g.T % g.kt) return lvc_fail(LVC_E_SHAPE, "clip_swin_embed: T must divide by the tubelet");
  if (ed->out_dtype != LVC_DT_F32 && ed->out_dtype != LVC_DT_BF16) return lvc_fail(LVC_E_INVAL, "clip_swin_embed: out_dtype must be f32 or bf16");
  if ((reinterpret_cast<uintptr_t>(tokens) & 127) || (reinterpret_cast<uintptr_t>(w_packed) & 127) || (reinterpret_cast<uintptr_t>(out) & 15) ||
      (reinterpret_cast<uintptr_t>(ready) & 3))
    return lvc_fail(LVC_E_ALIGN, "clip_swin_embed: tokens / packed weights need 128-byte, out 16-byte alignment");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int dev = 0, sms = 0, max_smem = 0;
  LVC_CUDA_TRY(cudaGetDevice(&dev));
  LVC_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  LVC_CUDA_TRY(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));

  GemmParams p{};
  p.M = g.M; p.N = g.N; p.K = g.K;
  p.BN = g.N;
  p.BK = (g.K % 64 == 0) ? 64 : 32;
  p.num_k_blocks = g.K / p.BK;
  p.m_tiles = (g.M + BM - 1) / BM;
  p.n_tiles = 1;
  p.tmem_cols = 256;   // two accumulators of N columns (the one epilogue group ping-pongs)
  p.bias = bias; p.ln_w = ln_w; p.ln_b = ln_b; p.ln_eps = ed->ln_eps;
  p.do_ln = 1;
  p.out_f32 = ed->out_dtype == LVC_DT_F32;
  p.out = out;
  p.ready = static_cast<const unsigned*>(ready);
  p.ready_need = (unsigned)(g.kt * g.kh);   // per token: all kt * kh row slices (3 channels x kw columns each) are stored
  const int a_tile = BM * g.K * 2, b_all = g.N * g.K * 2;
  const int fixed3 = 1024 + 4 * kSwinGroupsFused * 4096 + b_all + (BM + g.N) * 64 + 512 + 3 * g.N * 4 + 64;
  // ONE stage of A: the embedding has slack, the clip kernel does not -- every 24 KB given back to its source-row ring
  // counts (with two stages the ring no longer holds two 8-row stages of a full 720p frame, the kernel falls back to
  // 7-row stages and loses the whole-sector token stores: pair 138 us instead of 109)
  p.stages = env_i("LVC_FUSED_EMBED_STAGES", 1);
  const size_t smem3 = (size_t)fixed3 + (size_t)p.stages * a_tile;
  // what is left of the SM for the clip kernel: the opt-in maximum minus the embedding CTA and the per-CTA reserve
  const int clip_cap = max_smem - (int)smem3 - 2048;
  if (clip_cap < 64 * 1024) return lvc_fail(LVC_E_SHAPE, "clip_swin_embed: K = %d leaves the clip kernel %d bytes of shared memory", g.K, clip_cap);

  const int tiles = p.m_tiles;
  int tile_shift = 0;
  while ((1 << tile_shift) < BM) ++tile_shift;
  LVC_CUDA_TRY(cudaMemsetAsync(ready, 0, (size_t)tiles * 4, st));
  rc = lvc_clip_forward_fused(cd, videos, indices, boxes, lut, tokens, static_cast<unsigned*>(ready), tile_shift, clip_cap, stream);
  if (rc != LVC_OK) return rc;

  if (lvc_knob_set("LVC_FUSED_NO_EMBED")) return LVC_OK;   // profiling: the clip half alone
  CUtensorMap ma, mb, mo;
  rc = make_map(&ma, tokens, (uint64_t)g.M, (uint64_t)g.K, (uint32_t)p.BK, BM);
  if (rc != LVC_OK) return rc;
  rc = make_map(&mb, w_packed, (uint64_t)g.N, (uint64_t)g.K, (uint32_t)p.BK, (uint32_t)g.N);
  if (rc != LVC_OK) return rc;
  rc = make_out_map(&mo, out, (uint64_t)g.M, (uint64_t)g.N, p.out_f32 != 0);
  if (rc != LVC_OK) return rc;
  const int grid3 = std::min(sms, tiles);
  LVC_CUDA_TRY(cudaFuncSetAttribute(lvc_embed_swin_kernel<4, true, kSwinGroupsFused>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
  LVC_CUDA_TRY(cudaFuncSetAttribute(lvc_embed_swin_kernel<4, false, kSwinGroupsFused>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
  if (p.out_f32) {
    LVC_CUDA_TRY(cudaFuncSetAttribute(lvc_embed_swin_kernel<4, true, kSwinGroupsFused>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem3));
    LVC_CUDA_TRY(launch_pdl(lvc_embed_swin_kernel<4, true, kSwinGroupsFused>, grid3, swin_threads(kSwinGroupsFused), smem3, st, ma, mb, mo, p));
  } else {
    LVC_CUDA_TRY(cudaFuncSetAttribute(lvc_embed_swin_kernel<4, false, kSwinGroupsFused>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem3));
    LVC_CUDA_TRY(launch_pdl(lvc_embed_swin_kernel<4, false, kSwinGroupsFused>, grid3, swin_threads(kSwinGroupsFused), smem3, st, ma, mb, mo, p));
  }
  LVC_CUDA_TRY(cudaGetLastError());
  return LVC_OK;
}
