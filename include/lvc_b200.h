/* lvc_b200 -- C ABI of the B200-native clip-construction path.
 *
 * The reference (shaunliewsmu/ai-laryngeal-video-based-classifier) is pure Python and has no
 * FFI; its seams for this path are Python call signatures.  Each entry point
 * below names the reference interface it replaces (paths relative to the reference root).
 * INTEGRATION.md shows the ctypes binding a maintainer adds on the reference side.
 *
 * Conventions: every function returns 0 (LVC_OK) or a negative lvc_status; lvc_last_error()
 * returns a thread-local message for the last failure.  No ownership transfer: all buffers
 * are caller-allocated.  Device work is enqueued on the caller's stream (a cudaStream_t passed
 * as void*) and never synchronises.  There is no CPU fallback: without a CUDA device the
 * device entry points return LVC_E_CUDA.
 */
#ifndef LVC_B200_H
#define LVC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  LVC_OK = 0,
  LVC_E_INVAL = -1,  /* bad argument (NULL, negative size, unknown enum) */
  LVC_E_SHAPE = -2,  /* shape the kernels do not support (e.g. output size != 224) */
  LVC_E_CUDA = -3,   /* CUDA runtime / driver error, message in lvc_last_error() */
  LVC_E_ALIGN = -4,  /* pointer alignment requirement violated */
  LVC_E_NOTSUP = -5
} lvc_status;

const char* lvc_last_error(void);
int lvc_version(void);

/* ------------------------------------------------------------------ samplers (host) ---- */

/* MT19937 state shared by the CPython-`random` and numpy-legacy emulations. */
typedef struct {
  uint32_t mt[624];
  int32_t pos;
  int32_t reserved;
} lvc_rng;

/* random.seed(int)  -- e.g. vivit_transformer/vivit_classifier/data_config/dataset.py:39,
 * resnet50-2d-lstm/src/data_config/dataset.py:97 (init_by_array on |seed|). */
int lvc_rng_seed_py(lvc_rng* st, uint64_t seed);
/* np.random.seed(int) -- resnet50-2d-lstm/inference.py:77,114 (init_genrand). */
int lvc_rng_seed_np(lvc_rng* st, uint32_t seed);
/* random.random() / next 32-bit word, exposed for tests of the stream itself. */
double lvc_rng_random(lvc_rng* st);
uint32_t lvc_rng_u32(lvc_rng* st);

enum { LVC_SAMPLE_UNIFORM = 0, LVC_SAMPLE_RANDOM_WINDOW = 1, LVC_SAMPLE_RANDOM = 2 };
enum {
  /* VideoDataset.get_sampling_indices: vivit .../data_config/dataset.py:129-193 and its copies
   * (timesformer :137-207, swin :64-149, resnet50-3d-video :79-169, resnet50-2d-lstm :85-163).
   * CPython stream; short videos (N < T) use the duplicate-allowing branch. */
  LVC_VARIANT_DATASET = 0,
  /* visualize_sampling.py:105-178: T clamped to N, CPython stream. */
  LVC_VARIANT_CLAMPED = 1,
  /* resnet50-2d-lstm/inference.py:74-130: T clamped to N, numpy legacy stream
   * (np.random.choice without replacement / np.random.randint). */
  LVC_VARIANT_NUMPY_INFER = 2
};

/* Writes *n_out (<= num_frames) indices to out[num_frames].  `st` is consumed exactly like the
 * reference consumes its global stream; may be NULL for LVC_SAMPLE_UNIFORM. */
int lvc_sample_indices(int method, int variant, int64_t total_frames, int32_t num_frames,
                       lvc_rng* st, int64_t* out, int32_t* n_out);
/* data_handling.py:41-56 sample_frame_indices(clip_len, frame_sample_rate, seg_len); numpy stream. */
int lvc_sample_frame_indices(int32_t clip_len, double frame_sample_rate, int64_t seg_len,
                             lvc_rng* st, int64_t* out);
/* pytorchvideo UniformTemporalSubsample(T) index rule (videoswintransformer/.../dataset.py:157):
 * clamp(linspace_f32(0, t-1, T), 0, t-1).long(). */
int lvc_uniform_temporal_subsample(int64_t t, int32_t num_samples, int64_t* out);
/* Swin copy's second return value (videoswintransformer/.../dataset.py:104-121); NaN = None. */
double lvc_dynamic_fps(int64_t total_frames, int32_t num_frames, double original_fps);

/* torch CPU generator stream: after lvc_rng_seed_np(st, seed) the state equals torch.manual_seed(seed)'s
 * at::mt19937.  lvc_torch_randint = torch.randint(low, high, (1,)).item(); lvc_torch_rand = torch.rand(1).item(). */
int lvc_torch_randint(lvc_rng* st, int64_t low, int64_t high, int64_t* out);
float lvc_torch_rand(lvc_rng* st);
/* The train chain's per-clip draws in the reference's order -- RandomShortSideScale(min,max), RandomCrop(224),
 * RandomHorizontalFlip(p) (videoswintransformer/.../dataset.py:156-161) -- for a (height, width) video.
 * out = {short-side size, crop top, crop left, flip}: one row of lvc_clip_forward_aug's `aug` table. */
int lvc_train_aug_draw(lvc_rng* st, int32_t height, int32_t width, int32_t min_size, int32_t max_size,
                       double flip_p, int32_t* out);

/* -------------------------------------------------- normalisation look-up table (host) ---- */

enum {
  /* f32(x) / f32(scale)      -- resnet50-2d-lstm/inference.py:166-167 (`astype(float32) / 255.0`) */
  LVC_LUT_DIV = 0,
  /* f32(f64(x) * scale)      -- transformers image_transforms.rescale as called by
   *                             VivitImageProcessor (scale 1/127.5, offset) and the VideoMAE-class
   *                             processor (scale 1/255): vivit .../trainers/trainer.py:77,
   *                             timesformer .../trainers/trainer.py:91-98 */
  LVC_LUT_MUL64 = 1
};
/* out[c*256 + v] = ((rescale(v) [- 1 if offset]) - mean[c]) / std[c], every step rounded to
 * float32 in the reference's order, so a uint8 input gives the reference's float32 bit for bit. */
int lvc_build_lut(int kind, double scale, int offset, const float mean[3], const float std[3],
                  float* out /* [3*256] host */);

/* --------------------------------------------------------------- fused clip kernel ---- */

/* One decoded video resident in device memory: uint8 RGB, (n_frames, height, width, 3) contiguous.
 * `base` must be 16-byte aligned; `nbytes` is the readable size of the buffer. */
typedef struct {
  uint64_t base;
  int64_t nbytes;
  int32_t n_frames;
  int32_t height;
  int32_t width;
  int32_t reserved;
} lvc_video;

enum {
  /* cv2.resize(frame, (224,224)) INTER_LINEAR uint8, bit-exact (vivit .../dataset.py:275,
   * resnet50-2d-lstm/inference.py:162), then the LUT. */
  LVC_RESAMPLE_CV2 = 0,
  /* LVC_RESAMPLE_CV2 followed on-chip by the VivitImageProcessor geometry: Pillow BILINEAR
   * 224->256 and centre crop 224 (vivit .../trainers/trainer.py:77), then the LUT. */
  LVC_RESAMPLE_CV2_PIL256 = 1,
  /* pytorchvideo eval chain on float frames: ShortSideScale(256) via torch bilinear
   * (align_corners=False), CenterCrop(224), (x-mean)/std without /255
   * (videoswintransformer/.../dataset.py:164-172).  Uses mean/std, not the LUT. */
  LVC_RESAMPLE_PTV_EVAL = 2,
  /* HF image processor on RAW frames of any size: Pillow BILINEAR (antialiased) resize of the shortest edge,
   * centre crop 224, LUT.  Only through lvc_hf_clip_forward (lvc_clip_forward rejects it). */
  LVC_RESAMPLE_PIL_SHORT = 3
};
enum {
  LVC_LAYOUT_BTCHW = 0,
  LVC_LAYOUT_BCTHW = 1,
  LVC_LAYOUT_BTHWC = 2,
  /* the clip kernel writes the patch-embedding GEMM operand A[M][K] (bf16, token-major, K ordered
   * (c,kt,kh,kw)) directly, so lvc_patch_embed can skip its im2col pass */
  LVC_LAYOUT_TOKENS = 3
};
enum { LVC_DT_F32 = 0, LVC_DT_BF16 = 1, LVC_DT_U8 = 2 };
enum { LVC_SRC_RGB = 0, LVC_SRC_BGR = 1 };

typedef struct {
  int32_t resample;   /* LVC_RESAMPLE_* */
  int32_t layout;     /* LVC_LAYOUT_*; BTHWC only with LVC_DT_U8 (the Dataset's uint8 clip,
                         vivit .../dataset.py:287-292) */
  int32_t out_dtype;  /* LVC_DT_* ; U8 = resized bytes before normalisation */
  int32_t out_size;   /* 224 */
  int32_t batch;      /* B clips, one per video entry */
  int32_t num_frames; /* T */
  int32_t max_width;  /* largest frame width (pixels) among the B videos: sizes the shared-memory
                         ring.  A wider video in `videos` is a contract violation: lvc_clip_validate
                         reports it on the host before the launch (LVC_E_SHAPE); a launch that still
                         carries one never faults -- frames whose rows do not fit the ring are left
                         unwritten and LVC_DEVERR_FRAME_TOO_WIDE is raised in lvc_device_status. */
  int32_t tok_kt, tok_kh, tok_kw; /* LVC_LAYOUT_TOKENS: patch size, e.g. (2,16,16); else ignored */
  float mean[3];      /* PTV_EVAL only */
  float std[3];
  int32_t src_order;  /* LVC_SRC_RGB, or LVC_SRC_BGR: frames as cv2.VideoCapture.read() returns them; the
                         kernel's gather swaps the channels, replacing cv2.cvtColor(frame, COLOR_BGR2RGB)
                         (vivit .../dataset.py:336, resnet50-2d-lstm/inference.py:66) */
  int32_t src_float;  /* CV2 / CV2_PIL256 only.  1: resize like the reference's PRIMARY Dataset branch, where
                         pytorchvideo hands over float32 frames (vivit .../dataset.py:244, :271-284): cv2.resize
                         computes in float and `.astype(np.uint8)` truncates.  Differs from the uint8 path by
                         0/+1 LSB on ~37 % of pixels.  Frames are still stored as uint8 (they hold integers). */
} lvc_clip_desc;

/* Replaces, for a batch of B videos: the gather (vivit .../dataset.py:246-268), the per-frame
 * cv2.resize loop (:271-277), the image-processor call in _process_batch
 * (vivit .../trainers/trainer.py:62-95) or the manual normalise + permute
 * (resnet50-2d-lstm/inference.py:166-178), and the collate (model_configuration.py:11-31).
 *   videos   device  lvc_video[B]
 *   indices  device  int32[B*T]   frame index per output frame (clamped to the video)
 *   boxes    device  int32[B*T*4] xyxy ROI per output frame (src/yolo_detector.py:76-85), x1 < 0 =
 *                    no detection = full frame; NULL = no ROI
 *   lut      device  float[3*256] from lvc_build_lut (ignored for PTV_EVAL / U8 output)
 *   out      device  B*T*3*224*224 elements of out_dtype in `layout`
 *   stats    device  uint64[B] or NULL: += sum of the clip's resized uint8 samples (checksum); for
 *                    LVC_RESAMPLE_CV2_PIL256 the samples of the resized rows the 224-crop reads (13..210)
 */
int lvc_clip_forward(const lvc_clip_desc* desc, const lvc_video* videos, const int32_t* indices,
                     const int32_t* boxes, const float* lut, void* out, uint64_t* stats,
                     void* stream);

/* Train-time chain of the pytorchvideo datasets (videoswintransformer/.../dataset.py:153-162,
 * resnet50-3d-video/.../dataset.py:173-182, resnet50-2d-lstm/.../dataset.py:63-72): RandomShortSideScale(256,320)
 * -> RandomCrop(224) -> RandomHorizontalFlip(0.5) -> Normalize, drawn ONCE per clip like the reference's
 * clip-level transforms.  Same as lvc_clip_forward with resample = LVC_RESAMPLE_PTV_EVAL plus
 *   aug      device  int32[B*4]  per clip {short-side size, crop top, crop left, flip}; the draws themselves are
 *                    made on the host on a torch-compatible MT19937 stream (lvc_torch_randint / lvc_torch_rand)
 *                    so that a seeded run reproduces the reference's torch.manual_seed stream.  NULL = eval chain. */
int lvc_clip_forward_aug(const lvc_clip_desc* desc, const lvc_video* videos, const int32_t* indices,
                         const int32_t* boxes, const int32_t* aug, const float* lut, void* out,
                         uint64_t* stats, void* stream);

/* The HF image processors applied to RAW decoded frames, as the reference's stand-alone preprocessing scripts
 * call them (vivit_transformer/data_preprocessing.py:16 `image_processor(list(video), return_tensors="pt")`,
 * timesformer_transformer/data_preprocessing_timesformer.py:16, vivit_transformer/inference.py:187):
 * resize(shortest_edge) with PIL BILINEAR -- Pillow widens the triangle filter with the down-scale and rounds to
 * uint8 after each of its two passes -- centre crop 224, rescale + normalise (the LUT), layout.
 *   shortest_edge  256 for VivitImageProcessor, 224 for the VideoMAE-class processor; >= 224.  0 = the call form
 *                  `size={"height": 224, "width": 224}` (timesformer .../trainers/trainer.py:91-98): both axes are
 *                  resized to 224 independently (aspect not kept), the crop is the identity
 *   max_height     largest frame height among the B videos (with desc->max_width it sizes the shared-memory
 *                  buffers; a larger video is a contract violation: lvc_clip_validate reports it on the host,
 *                  the kernel treats it as max-sized -- every access stays inside its buffer, its output frames
 *                  are meaningless -- and raises LVC_DEVERR_FRAME_TOO_LARGE).  Down-scales up to 11.5x.
 * desc->resample is ignored (LVC_RESAMPLE_PIL_SHORT implied), desc->src_float must be 0; no ROI boxes, no stats.
 * Output size is (int(E*W/H)|E) x (E|int(E*H/W)) as get_resize_output_image_size computes it, cropped at
 * ((nh-224)//2, (nw-224)//2).  Bit-exact against the processor (tests/test_hfraw_gpu.py). */
int lvc_hf_clip_forward(const lvc_clip_desc* desc, int32_t shortest_edge, int32_t max_height,
                        const lvc_video* videos, const int32_t* indices, const float* lut, void* out,
                        void* stream);
/* Same, with flags.  LVC_HF_SRC_HOST: the frames live in PINNED HOST memory (what the reference's scripts hold: decoded
 * numpy frames) and the kernel reads them directly over PCIe -- one 16-byte-granular bulk copy per source row, and only
 * the rows and columns the centre crop needs (43 % of a 1080p frame at shortest_edge 256), no staging copy. */
enum { LVC_HF_SRC_HOST = 1 };
int lvc_hf_clip_forward_ex(const lvc_clip_desc* desc, int32_t shortest_edge, int32_t max_height, int32_t flags,
                           const lvc_video* videos, const int32_t* indices, const float* lut, void* out,
                           void* stream);

/* Number of kernel launches lvc_clip_forward enqueues for this descriptor (for accounting). */
int lvc_clip_launch_count(const lvc_clip_desc* desc);

/* Host-side check of the `videos` table against the descriptor BEFORE a launch, from a host copy of the table (the
 * device entry points only see a device pointer and never synchronise): every video has frames, a 16-byte aligned
 * base, nbytes >= n_frames*height*width*3, width <= desc->max_width and -- max_height > 0, the lvc_hf_clip_forward
 * contract -- height <= max_height.  LVC_E_SHAPE / LVC_E_INVAL / LVC_E_ALIGN with the offending entry in
 * lvc_last_error().  The reference's seam raises ValueError for unusable videos (vivit .../dataset.py:217). */
int lvc_clip_validate(const lvc_clip_desc* desc, int32_t max_height, const lvc_video* host_videos);

/* Contract violations a kernel found on the device (it skips the offending frames instead of faulting): a sticky
 * bit mask per device, OR-ed by the kernels.  Synchronises `stream`, returns the mask and (clear != 0) resets it. */
enum {
  LVC_DEVERR_FRAME_TOO_WIDE = 1,  /* lvc_clip_forward: a source window does not fit the ring sized by max_width */
  LVC_DEVERR_FRAME_TOO_LARGE = 2, /* lvc_hf_clip_forward: a video exceeds max_width x max_height */
  LVC_DEVERR_INTERNAL = 4         /* a kernel's own consistency check failed (never expected; report it) */
};
int lvc_device_status(void* stream, uint32_t* flags, int32_t clear);

/* Bytes the clip kernels' bulk-copy producers requested from the source buffers (HBM, or pinned host memory read over
 * PCIe) since the counter was last cleared, on the current device: a device-side counter, one atomic per producer
 * warp per launch.  Synchronises `stream`. */
int lvc_device_bytes_fetched(void* stream, uint64_t* bytes, int32_t clear);

/* Page-locked host memory for decoded frames the kernels read in place (zero-copy ingest): portable + mapped, and with
 * LVC_HOST_WRITE_COMBINED not snooped by the CPU caches on device reads (fast to fill sequentially from a decoder,
 * slow to READ from the CPU).  The reference holds decoded frames in numpy arrays (vivit .../dataset.py:336-341). */
enum { LVC_HOST_WRITE_COMBINED = 1 };
int lvc_host_alloc(int64_t nbytes, int32_t flags, void** out);
int lvc_host_free(void* p);

/* Launch geometry lvc_clip_forward derives from a descriptor (pure host arithmetic; for tests and tuning reports).
 * sm_count / max_smem: the device's SM count and opt-in shared memory per CTA; <= 0 = query the current device
 * (B200: 148, 232448). */
typedef struct {
  int32_t rows_per_stage;   /* output rows per ring stage (2 source rows each) */
  int32_t ring_bytes;       /* byte-granular source-row ring */
  int32_t ring_offset;      /* where the ring starts in shared memory (tables, LUTs, ViViT intermediate before it) */
  int32_t slot_pitch;       /* bytes per row slot of the widest video */
  int32_t stages_of_widest; /* stages of max_width rows the ring holds (narrower windows: proportionally more, <= 16) */
  int32_t bands;            /* work items per frame */
  int32_t rows_per_band;
  int32_t producers;        /* bulk-copy producer warps */
  int32_t grid, threads, smem_bytes;
  int32_t reserved[5];
} lvc_clip_plan;
enum { LVC_PLAN_BOXES = 1 /* an ROI table will be passed */, LVC_PLAN_STATS = 2 /* a stats buffer will be passed */ };
int lvc_clip_plan_query(const lvc_clip_desc* desc, int32_t flags, int32_t sm_count, int32_t max_smem,
                        lvc_clip_plan* out);

/* ------------------------------------------------------------ patch / tubelet embedding ---- */

enum {
  LVC_EMBED_VIVIT = 0,       /* VivitTubeletEmbeddings: Conv3d(3,E,k=s=(2,16,16)); vivit_model.py:27-41 */
  LVC_EMBED_TIMESFORMER = 1, /* TimesformerPatchEmbeddings: Conv2d(3,E,k=s=16); timesformer_model.py:27-42 */
  LVC_EMBED_SWIN = 2         /* PatchEmbed3d: Conv3d(3,E,k=s=(2,4,4)) + LayerNorm(E); swin3d.py:22-44 */
};

typedef struct {
  int32_t kind;      /* LVC_EMBED_* */
  int32_t x_dtype;   /* LVC_DT_F32 or LVC_DT_BF16 */
  int32_t x_layout;  /* LVC_LAYOUT_BTCHW (HF models), LVC_LAYOUT_BCTHW (torchvision swin), or
                        LVC_LAYOUT_TOKENS: x already is A[M][K] bf16 from lvc_clip_forward */
  int32_t out_dtype; /* LVC_DT_BF16 or LVC_DT_F32 */
  int32_t batch, frames, height, width; /* B, T, H, W of x; channels are 3 */
  int32_t kt, kh, kw; /* patch size: (2,16,16) / (1,16,16) / (2,4,4) */
  int32_t embed_dim;  /* N: 768 / 768 / 128 */
  float ln_eps;       /* SWIN: LayerNorm eps (1e-5); ignored otherwise */
  int32_t reserved;
} lvc_embed_desc;

/* Bytes of device scratch lvc_patch_embed needs (the bf16 token-major patch matrix A[M][K]). */
int64_t lvc_patch_embed_workspace(const lvc_embed_desc* d);
/* Packs conv weights (E, 3, kt, kh, kw) fp32 (device) into the bf16 [N][K] K-major operand the
 * tcgen05 GEMM reads; call once per weight update. */
int lvc_patch_embed_pack_weights(const lvc_embed_desc* d, const float* w_f32, void* w_packed_bf16,
                                 void* stream);
/* out[M][N] = patches(x)[M][K] . W[K][N] + bias (+ LayerNorm for SWIN), token order as the
 * reference modules produce it: VIVIT (B, T/2*H/16*W/16, E); TIMESFORMER (B*T, H/16*W/16, E);
 * SWIN (B, T/2, H/4, W/4, E).  x, w_packed, bias, ln_w, ln_b, out, workspace: device pointers. */
int lvc_patch_embed(const lvc_embed_desc* d, const void* x, const void* w_packed_bf16,
                    const float* bias, const float* ln_w, const float* ln_b, void* out,
                    void* workspace, void* stream);
int lvc_patch_embed_launch_count(const lvc_embed_desc* d);

/* lvc_patch_embed without a caller workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  The bf16 patch matrix A[M][K] is written to a file-scope
 * `static __device__` array of LVC_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library holds it
 * until the process exits); the entry point calls no cudaMalloc / cudaFree / allocating library routine.
 * Shapes whose lvc_patch_embed_workspace() exceeds LVC_STATIC_SCRATCH_BYTES return LVC_E_SHAPE (16 MiB holds one
 * clip of every configured model: ViViT 32x224x224 = 9.6 MB, TimeSformer 8x224x224 = 2.4 MB, Swin-B 32x224x224 =
 * 9.6 MB).  All calls on a device share the one array: it keeps the previous call's contents (nothing needs clearing,
 * the im2col pass overwrites every element the GEMM reads), and two calls that overlap on the GPU (different streams,
 * or graphs replayed concurrently) RACE -- serialise them on one stream.  Other arguments and results are exactly
 * lvc_patch_embed's (same kernels).  The first call on a device resolves the array's address (cudaGetSymbolAddress) and loads
 * the kernels: make it once outside a stream capture; later calls only enqueue launches. */
#define LVC_STATIC_SCRATCH_BYTES (16 * 1024 * 1024)
int64_t lvc_patch_embed_static_capacity(void);
int lvc_patch_embed_static(const lvc_embed_desc* d, const void* x, const void* w_packed_bf16,
                           const float* bias, const float* ln_w, const float* ln_b, void* out, void* stream);

/* The step after the patch embedding, fused into the GEMM epilogue:
 *   VIVIT        VivitEmbeddings.forward (transformers modeling_vivit.py:141-155): cat(cls_token, patches) +
 *                position_embeddings -> out (B, 1 + T/2*14*14, E)
 *   TIMESFORMER  TimesformerEmbeddings.forward (modeling_timesformer.py:92-144): cat(cls, patches) +
 *                position_embeddings, then + time_embeddings per frame and the (B, patch*T + t) token order of
 *                divided space-time attention -> out (B, 1 + 196*T, E); time_embeddings = NULL is "space_only"
 *                ((B*T, 197, E) is NOT produced: the cls row of each batch entry is written once).
 * Tables are fp32 device pointers: cls_token [E], position_embeddings [(1 + tokens)][E], time_embeddings [T][E].
 * Dropout layers are identity (inference / p = 0, the configs' default).  Other arguments as lvc_patch_embed. */
int lvc_embeddings_forward(const lvc_embed_desc* d, const void* x, const void* w_packed_bf16, const float* bias,
                           const float* cls_token, const float* position_embeddings,
                           const float* time_embeddings, void* out, void* workspace, void* stream);

/* The Video Swin front end in one call: lvc_clip_forward (cv2 chain, bf16 token layout of the (2,4,4) patch) and
 * lvc_patch_embed (PatchEmbed3d + LayerNorm) as two CO-RESIDENT kernels -- replaces, in one pass,
 * videoswintransformer/swin_video_classifier/data_config/dataset.py:151-173 (+ src/yolo_detector.py:76-85 boxes) and
 * models/swin3d.py:22-44 (torchvision swin_transformer.py:353-363).  The embedding kernel is launched behind the clip
 * kernel with programmatic stream serialization and loads each 128-token tile of the patch matrix as soon as the clip
 * kernel's per-tile counter says it is complete, so the HBM-bound embedding hides behind the latency-bound clip stage.
 * `tokens` (bf16 A[M][K]) and `ready` (uint32 per tile) are device workspaces of the sizes
 * lvc_clip_swin_embed_workspace reports (ready is cleared by the call).  clip_desc: LVC_RESAMPLE_CV2, LVC_LAYOUT_TOKENS,
 * LVC_DT_BF16, tok_* = the embedding's patch; embed_desc: LVC_EMBED_SWIN, x_layout LVC_LAYOUT_TOKENS, 224 x 224.
 * Results are bit-identical to the two separate calls.  Other arguments as lvc_clip_forward / lvc_patch_embed.
 * EXPERIMENTAL, NOT THE DEFAULT PATH: on BASELINE config 4 the pair measures 109 us against 100 us for the two launches
 * (the clip stage runs on a smaller ring, 80 instead of 68 us, and the embedding ends 16 us after it);
 * pipeline.ClipEmbedPipeline and bench.py keep the two launches.  Correctness never depends on co-residency: the clip
 * kernel does not wait for the embedding kernel, which simply runs afterwards when the two do not fit on an SM together. */
int lvc_clip_swin_embed_workspace(const lvc_embed_desc* embed_desc, int64_t* tokens_bytes, int64_t* ready_bytes);
int lvc_clip_swin_embed_forward(const lvc_clip_desc* clip_desc, const lvc_video* videos, const int32_t* indices,
                                const int32_t* boxes, const float* lut, const lvc_embed_desc* embed_desc,
                                const void* w_packed_bf16, const float* bias, const float* ln_w, const float* ln_b,
                                void* tokens, void* ready, void* out, void* stream);

/* --------------------------------------------- detector pre / post-processing ---- */

/* Geometry of yolov5 `letterbox(im, new_shape, stride=stride)` with its defaults auto=True, scaleFill=False,
 * scaleup=True (src/yolo_detector.py:52; ultralytics/yolov5 utils/augmentations.py -- an un-vendored, unpinned
 * submodule of the reference: restated from its published definition). */
typedef struct {
  int32_t src_h, src_w;   /* decoded frame */
  int32_t new_h, new_w;   /* cv2.resize target: round(src * r), r = min(shape_h / src_h, shape_w / src_w) */
  int32_t top, left;      /* border before the image: round(d/2 - 0.1) */
  int32_t out_h, out_w;   /* network input: new + top + bottom, new + left + right (multiples of stride when auto) */
} lvc_letterbox_geom;
int lvc_letterbox_geometry(int32_t src_h, int32_t src_w, int32_t shape_h, int32_t shape_w, int32_t stride,
                           int32_t auto_pad, int32_t scaleup, lvc_letterbox_geom* out);
/* src/yolo_detector.py:52-61 for n frames in one launch: cv2.resize(INTER_LINEAR, bit-exact uint8) to
 * (new_w, new_h), constant border 114, HWC->CHW with the channel order REVERSED (`transpose(2,0,1)[::-1]`: the
 * reference feeds BGR frames, the network sees RGB), float(x) / 255.
 *   frames  device  uint8 (n, src_h, src_w, 3);  out  device  (n, 3, out_h, out_w) of out_dtype (F32 or BF16) */
int lvc_letterbox_forward(const lvc_letterbox_geom* g, const uint8_t* frames, int32_t n, int32_t out_dtype,
                          void* out, void* stream);
/* src/yolo_detector.py:76: `scale_boxes(im.shape[2:], det[:, :4], im0.shape).round()` (yolov5 utils/general.py:
 * subtract the letterbox pad, divide by the gain, clip to the frame) followed by the int() of :81.
 *   det  device  float rows of `row_stride` floats starting with xyxy (yolov5 NMS output rows are xyxy,conf,cls);
 *   boxes_out  device  int32 [n][4] -- directly usable as rows of lvc_clip_forward's `boxes` table. */
int lvc_scale_boxes(const lvc_letterbox_geom* g, const float* det, int32_t n, int32_t row_stride,
                    int32_t* boxes_out, void* stream);
/* Per frame, the highest-confidence detection becomes the frame's ROI (the reference gates on max confidence,
 * src/yolo_detector.py:161); frames without a detection >= conf_thres get x1 = -1 (= full frame).
 *   det  device  float [n_det][6] xyxy,conf,cls in letterboxed coordinates; frame_of  device int32 [n_det];
 *   boxes_out  device int32 [n_frames][4], written completely. */
int lvc_best_boxes(const lvc_letterbox_geom* g, const float* det, const int32_t* frame_of, int32_t n_det,
                   int32_t n_frames, float conf_thres, int32_t* boxes_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LVC_B200_H */
