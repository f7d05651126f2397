/* libmab200 -- C ABI of the B200-native MapAnything inference hot path.
 *
 * The reference has no native boundary on this path (its only native entry point is cuRoPE's
 * pybind `rope_2d`, uniception/models/libs/croco/curope/curope.cpp:49-69, which the default model
 * never calls).  These entry points are what a ctypes/cffi binding inside the reference's
 * `MapAnything.forward` (mapanything/models/mapanything/model.py:1657-2152) and
 * `postprocess_model_outputs_for_inference` (mapanything/utils/inference.py:314-506) would bind;
 * see INTEGRATION.md for the stub.
 *
 * Conventions: plain pointers and sizes, no torch types.  Every pointer is a DEVICE pointer unless
 * the name ends in _host.  Every call is asynchronous on the caller's `stream` (a cudaStream_t
 * passed as void*), never synchronises the device and returns 0 on success, a negative value for
 * an argument error, or a positive cudaError_t.  mab_last_error() returns a static description.
 * A handle may be used by one host thread at a time.
 */
#ifndef MAB200_H
#define MAB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mab_handle mab_handle;

enum { MAB_F32 = 0, MAB_BF16 = 1 };
enum { MAB_ACT_NONE = 0, MAB_ACT_GELU = 1, MAB_ACT_RELU = 2 };

/* Architecture of the production model (configs/model/mapanything.yaml and SURVEY.md section 0). */
typedef struct {
  int patch;          /* 14 */
  int enc_dim;        /* 1024 */
  int enc_depth;      /* 24 */
  int enc_heads;      /* 16 */
  int enc_pos_grid;   /* 37 (pos_embed is 1 + 37*37 rows) */
  int aat_dim;        /* 768 (1024 in the 48-layer configs) */
  int aat_depth;      /* 24 */
  int aat_heads;      /* 12 = aat_dim / 64 */
  int aat_tap0;       /* 11 */
  int aat_tap1;       /* 17 */
  int head_chunk;     /* views per dense-head pass (workspace knob; outputs do not depend on it) */
  /* info_sharing module_args variants (configs/model/info_sharing/aat_ifr_24_layers_*.yaml) */
  int aat_ref_view_pe;      /* 1: distinguish_ref_and_non_ref_views (production); 0: *_no_ref_view.yaml */
  int aat_entropy_scaling;  /* 1: use_entropy_scaling (*_escaling.yaml): q *= sqrt(1.4 log N / log 444) */
  /* aat_ifr_48_layers*.yaml: aat_dim 1024 / aat_heads 16 / aat_depth 48 (no proj_embed at that width,
   * alternating_attention_transformer.py:121-124) and a third intermediate feature, indices [11,23,35]:
   * the dense head then takes [tap0, tap1, tap2, final] instead of [encoder, tap0, tap1, final]
   * (model.py:324-327,1724-1768).  <= 0: two intermediate features. */
  int aat_tap2;
  /* 1: use_pe_for_non_reference_views (alternating_attention_transformer.py:30,628-645): every view adds
   * a row of info_sharing.view_pos_table [max_num_views_for_pe, aat_dim] to its tokens; the rows come
   * from mab_set_view_pe_rows before every forward (the reference draws them with torch.randint). */
  int aat_nonref_view_pe;
  /* 1: model_type "global_attention" (gat_ifr_24_layers*.yaml, global_attention_transformer.py): every
   * block attends over all tokens of the scene (the alternating trunk: even blocks only); that model
   * always encodes every view (aat_ref_view_pe = aat_nonref_view_pe = 1, table of max_num_views rows). */
  int aat_all_global;
  /* Dense head (DPT feature + regressor) arithmetic.  1: fp32 activations and TF32 tensor-core
   * contractions -- the reference's own precision for this region (autocast off at model.py:1774, TF32
   * allowed at model.py:89-93; uniception/models/prediction_heads/dpt.py:180-232,285-311) -- the Python
   * mirror's default.  0: bf16 operands (about 2 ms faster per 8 views; the mask-logit channel is then
   * 1.6e-2 from fp32 on identical inputs instead of < 5e-3). */
  int head_tf32;
  /* Views per encoder pass (SURVEY.md 8(f) many-view streaming; the reference chunks only its heads,
   * model.py:1440-1477,1533-1618).  The DINOv2 encoder and the geometric-input encoders are per-view
   * independent, so they run enc_chunk views at a time through ONE pass-sized set of activations
   * (38 KB per token); only the fused tokens, the info-sharing stream / activations and its taps stay
   * resident per view (about 35 MB per view at 518x392 instead of 60 MB).  Outputs do not depend on it.
   * <= 0 or >= the views held: one pass. */
  int enc_chunk;
  /* Encoder lanes: 2 = the passes of the encoder alternate between TWO streams (the caller's and one the
   * handle owns), each with its own pass-sized activations, so that the idle SMs of one lane's last wave
   * and draining tiles could run the other lane's kernels; 0 / 1 = one stream (the Python mirror's
   * default: measured on one box, 8 views, 21.76 / 21.86 ms with one lane against 21.82 / 21.92 ms with
   * two -- the halves' own wave quantisation eats what the overlap buys).  Outputs do not depend on it. */
  int enc_lanes;
} mab_config;

/* One scene, `views` views of H x W pixels held by this rank. */
typedef struct {
  int views;             /* views held by this rank */
  int H, W;              /* multiples of 14 */
  const float* images;   /* [views,3,H,W] fp32, already normalised (model.py:2229-2232) */
  /* optional geometric inputs (multi-modal, model.py:1377-1436); NULL = absent for every view */
  const float* ray_dirs;        /* [views,H,W,3] unit rays in the camera frame */
  const float* depth_along_ray; /* [views,H,W]   0 = missing */
  const float* cam_quats;       /* [views,4] xyzw cam2world, already relative to view 0 */
  const float* cam_trans;       /* [views,3] already relative to view 0 */
  const unsigned char* ray_mask;    /* [views] per-view availability flags (0/1) */
  const unsigned char* depth_mask;  /* [views] */
  const unsigned char* cam_mask;    /* [views] */
  const unsigned char* depth_metric_mask; /* [views] is_metric_scale & use_depth_scale */
  const unsigned char* pose_metric_mask;  /* [views] is_metric_scale & use_pose_scale */
  float pose_trans_norm;  /* mean non-zero |t| over ALL views of the scene (geometry.py:1629-1666) */
} mab_inputs;

/* Raw forward outputs (model.py:1907-1923, 2117-2150), all fp32 unless noted. */
typedef struct {
  float* pts3d;            /* [views,H,W,3] */
  float* pts3d_cam;        /* [views,H,W,3] */
  float* ray_directions;   /* [views,H,W,3] */
  float* depth_along_ray;  /* [views,H,W]   */
  float* conf;             /* [views,H,W]   */
  float* mask_logits;      /* [views,H,W]   */
  unsigned char* non_ambiguous_mask; /* [views,H,W] 0/1 */
  float* pose;             /* [views,8]: unscaled t(3), unit q(4), pad */
  float* metric_scale;     /* [1] */
  float* logits6;          /* optional [views,H,W,6] raw regressor output, may be NULL */
} mab_outputs;

/* Post-processed extras (inference.py:314-506). */
typedef struct {
  float* img_no_norm;      /* [views,H,W,3] */
  float* depth_z;          /* [views,H,W] */
  float* intrinsics;       /* [views,3,3] */
  float* camera_poses;     /* [views,4,4] */
  float* cam_trans;        /* [views,3] (x metric scale) */
  float* cam_quats;        /* [views,4] */
  unsigned char* mask;     /* [views,H,W] 0/1 */
} mab_post_outputs;

typedef struct {
  int apply_mask, mask_edges;
  float edge_normal_threshold_deg, edge_depth_threshold;
  int apply_confidence_mask;
  float confidence_percentile;
} mab_post_options;

const char* mab_last_error(void);
int mab_version(void);

int mab_create(const mab_config* cfg, int device, mab_handle** out);
int mab_destroy(mab_handle* h);
/* Upload one tensor of the reference state dict (reference key names, fp32, host or device). */
int mab_set_weight(mab_handle* h, const char* ref_key, const void* ptr, int dtype, int ndim,
                   const int64_t* shape);
/* view_pos_table row of every view THIS RANK holds (host array of n ints, copied; row 0 for the scene's
 * reference view, -1 = none), consumed by the next mab_forward / mab_stage_encode of a handle created
 * with aat_nonref_view_pe.  n must equal that call's in->views. */
int mab_set_view_pe_rows(mab_handle* h, const int* rows, int n);
/* Pack bf16 GEMM/conv operands, merge aliases, drop the fp32 staging copies that are not needed. */
int mab_finalize_weights(mab_handle* h, void* stream);

size_t mab_workspace_bytes(mab_handle* h, int views_local, int views_total, int H, int W);
/* Whole forward on one GPU (views_total == views). */
int mab_forward(mab_handle* h, const mab_inputs* in, mab_outputs* out, void* workspace,
                size_t workspace_bytes, void* stream);
/* inference.py:314-506 for all views of this rank.  workspace: views*H*W*18 + views*8 + 512 bytes. */
int mab_postprocess(mab_handle* h, const mab_inputs* in, mab_outputs* raw, mab_post_outputs* post,
                    const mab_post_options* opt, void* workspace, size_t workspace_bytes,
                    void* stream);

/* ---- pre-processing of the optional geometric inputs (mapanything/utils/inference.py:222-311) ----
 * Device pointers, fp32.  K [views,3,3] -> unit rays [views,H,W,3] (geometry.py:186-241);
 * rays / (|rays| + 1e-8) (inference.py:251-256); z-depth -> depth along the ray (inference.py:263-274). */
int mab_rays_from_intrinsics(const void* K, int views, int H, int W, void* rays, void* stream);
int mab_normalize_rays(const void* rays_in, long long pixels, void* rays_out, void* stream);
int mab_depth_z_to_along_ray(const void* depth_z, const void* rays, long long pixels, void* out,
                             void* stream);

/* ---- staged forward (view-sharded multi-GPU: the caller exchanges K/V between the stages) ----
 * Rank r holds a contiguous block of the scene's views; the rank that holds view 0 also owns the
 * scale token.  has_scale_token: bit 0 = this rank owns the scale token, bit 1 = this rank holds
 * view 0 (gets the reference-view positional encoding, alternating_attention_transformer.py:618-626).
 * views_total > in->views reserves the ring-attention carry buffers in the workspace. */
int mab_stage_encode(mab_handle* h, const mab_inputs* in, int views_total, void* workspace,
                     size_t workspace_bytes, int has_scale_token, void* stream);
/* LN1 + qkv projection of info-sharing block `layer`; returns the device pointer of this rank's
 * packed [rows, 3*aat_dim] bf16 qkv buffer (inside the workspace) and its row count. */
int mab_stage_aat_qkv(mab_handle* h, int layer, void** qkv_out, int* rows_out, void* stream);
/* Attention of block `layer` against ONE K/V block: kv = [kv_rows, 3*aat_dim] bf16 qkv rows of the
 * rank whose block is currently held (NULL = this rank's own).  A global (even) layer of a
 * view-sharded scene takes one call per ring step, flags marking the first / last step; the
 * online-softmax state is carried between the calls.  Everything else: one call, both flags. */
enum { MAB_RING_FIRST = 1, MAB_RING_LAST = 2 };
int mab_stage_aat_attend(mab_handle* h, int layer, const void* kv, int kv_rows, int flags,
                         void* stream);
/* Gather variant of the exchange step: export this rank's K|V rows ([rows, 2*aat_dim] bf16,
 * contiguous) for the other ranks, and attend against a gathered K|V block of any rank(s).  Same
 * flags / carried state as mab_stage_aat_attend; blocks may be presented in any order. */
int mab_stage_aat_export_kv(mab_handle* h, void* dst, void* stream);
int mab_stage_aat_attend_kv(mab_handle* h, int layer, const void* kv, int kv_rows, int flags,
                            void* stream);
/* proj + MLP of block `layer` (+ intermediate taps / final norm) */
int mab_stage_aat_finish(mab_handle* h, int layer, void* stream);
/* metric scale from the final scale token (rank owning it); broadcast the float to the others */
int mab_stage_scale_head(mab_handle* h, float* metric_scale_out, void* stream);
/* pose head + dense head + fused adaptor/pointmap.  metric_scale_in: device float broadcast from
 * the scale-token rank, or NULL to compute it here (single GPU). */
int mab_stage_heads(mab_handle* h, const mab_inputs* in, mab_outputs* out,
                    const float* metric_scale_in, void* stream);
/* The same for the views [view_lo, view_hi) of this rank only; out->metric_scale must already hold
 * the scene's metric scale.  Lets the caller post-process and copy out finished views while the
 * heads of later views compute (MapAnything.infer(host_outputs=True)). */
int mab_stage_heads_range(mab_handle* h, const mab_inputs* in, mab_outputs* out, int view_lo,
                          int view_hi, void* stream);

/* The view-sharded forward as ONE call (SURVEY.md 8(b) `mab_set_comm`, 8(e)): no reference counterpart
 * (the reference has no multi-GPU inference; its only NCCL use is training, train_tools.py:362-403).
 * mab_set_comm: `nccl_comm` is the caller's ncclComm_t over the `world` ranks that share the scene (rank r
 * holds the r-th block of parallel.shard_table), `nccl_library` the path of the libnccl.so.2 that created it
 * (resolved with dlopen: the library has no link-time NCCL dependency).  The communicator stays the
 * caller's.  mab_forward_sharded = mab_stage_encode, then per block qkv / attention / finish with the K|V
 * rows of the global layers exchanged by grouped ncclSend / ncclRecv on a stream the handle owns (every
 * rank attends to its own block while the blocks travel, carried online softmax), the metric scale
 * broadcast from rank 0, and (with_heads != 0) the heads.  kv_all: caller-owned scratch of
 * mab_kv_gather_bytes(views_total, H, W) bytes (the scene's K|V rows, bf16). */
int mab_set_comm(mab_handle* h, void* nccl_comm, int rank, int world, const char* nccl_library);
size_t mab_kv_gather_bytes(mab_handle* h, int views_total, int H, int W);
int mab_forward_sharded(mab_handle* h, const mab_inputs* in, int views_total, mab_outputs* out, void* workspace,
                        size_t workspace_bytes, void* kv_all, size_t kv_all_bytes, int with_heads, void* stream);

/* Launch accounting for bench.py: kernels launched by this library since load, and an optional
 * per-launch CUDA-event profiler (events recorded on the launch stream around every kernel). */
enum { MAB_CAT_GEMM = 0, MAB_CAT_CONV3 = 1, MAB_CAT_CONV3_REGRESS = 2, MAB_CAT_ATTENTION = 3,
       MAB_CAT_LAYERNORM = 4, MAB_CAT_ELEMENTWISE = 5, MAB_CAT_POSTPROCESS = 6,
       MAB_CAT_GEMM_TF32 = 7,  /* kind::tf32 linear layers (fp32 dense head 1x1 convs, geometric encoders) */
       MAB_CAT_CONV3_TF32 = 8, /* kind::tf32 3x3 convolutions; MAB_CAT_CONV3_REGRESS is bf16 or tf32 by head_tf32 */
       MAB_CAT_COUNT = 9 };
/* K/V length up to which mab_attention uses the one-Q-tile kernel (two CTAs per SM) instead of the
 * two-Q-tile kernel; < 0 restores the default (6000).  Test hook: both kernels at small sizes. */
int mab_set_attention_short_max(int kv_len);
long long mab_launch_count(void);
int mab_prof_enable(int on); /* also clears the records */
/* synchronises the device; sums device time, algorithmic FLOPs and bytes of one category */
int mab_prof_query(int cat, long long* launches, double* ms, double* flops, double* bytes);
/* the same records one by one, in launch order (arrays of `max` entries, any may be NULL); returns the number
 * of records held, -1 on a CUDA error.  profiles/step_breakdown.py groups them by operator shape. */
long long mab_prof_records(int* cat, double* ms, double* flops, double* bytes, long long max);

/* Diagnostic hook for the parity tests: device pointer, element count and dtype (MAB_F32 /
 * MAB_BF16) of a named intermediate of the last forward ("enc_tokens", "fused", "aat_stream",
 * "tap0", "tap1", "tap2", "final", "scale_token", "pose_raw", "pose_pool", "feat8"). */
int mab_debug_buffer(mab_handle* h, const char* name, void** ptr, long long* numel, int* dtype);

/* ---- per-kernel entry points (parity tests call these through the same ABI) ---- */
int mab_layernorm(const float* x, int rows, int C, const float* w, const float* b, float eps,
                  void* out_bf16, float* out_f32, void* stream);
/* C = act(A[M,K] * B[N,K]^T + bias) * gamma + res ; A,B bf16 row-major; outputs optional */
int mab_gemm(const void* A, const void* B, int M, int N, int K, const float* bias,
             const float* gamma, const float* res_f32, int act, float* out_f32, void* out_bf16,
             void* stream);
/* qkv: [rows, 3*heads*64] bf16 packed q|k|v ; out: [rows, heads*64] bf16 */
int mab_attention(const void* q_src, int q_rows_total, const void* kv_src, int kv_rows_total,
                  int heads, int num_seqs, int q_len, int kv_len, int q_seq_stride,
                  int kv_seq_stride, float scale, void* out, float* lse, void* stream);
/* one ring step of attention with carried online-softmax state (carry_o [rows, heads*64] fp32,
 * carry_ml [rows, heads, 2] fp32); carry_out = 0 writes the normalised bf16 `out` */
int mab_attention_carry(const void* q_src, int q_rows_total, const void* kv_src, int kv_rows_total,
                        int heads, int num_seqs, int q_len, int kv_len, int q_seq_stride,
                        int kv_seq_stride, float scale, void* out, float* carry_o, float* carry_ml,
                        int carry_in, int carry_out, void* stream);
/* 3x3/s1/p1 conv over NHWC bf16; weight packed [Cout, 9, Cpad] bf16 (mab_pack_conv3_weight) */
int mab_conv3x3(const void* x, int batch, int H, int W, int C, const void* w_packed, int Cout,
                const float* bias, const void* res1, const void* res2, int act, void* out_bf16,
                void* out_relu, void* stream);
int mab_pack_conv3_weight(const float* w_oihw, int Cout, int Cin, void* out_packed, void* stream);
/* TF32 variants for the reference's fp32 regions (model.py:1377): fp32 row-major / NHWC operands,
 * fp32 output, out = act(A B^T + bias) + res (act_post = 1: act(A B^T + bias + res)).
 * Conv weight packed [Cout, 9, Cpad] fp32 with Cpad = the activation's (padded) channel count. */
int mab_gemm_tf32(const float* A, const float* B, int M, int N, int K, const float* bias,
                  const float* res_f32, int act, int act_post, float* out_f32, void* stream);
int mab_pack_conv3_weight_f32(const float* w_oihw, int Cout, int Cin, int Cpad, float* out_packed,
                              void* stream);
int mab_conv3x3_tf32(const float* x, int batch, int H, int W, int C, const float* w_packed, int Cout,
                     const float* bias, const float* res_f32, int act, int act_post, float* out_f32,
                     void* stream);
int mab_resize_bilinear(const void* x, int batch, int Hi, int Wi, int C, int Hv, int Wv, int Ho,
                        int Wo, void* out, void* out_relu, void* stream);
int mab_posembed_resize(const float* pos, int grid, int C, int h, int w, float* out, void* stream);
/* RoPE2D in place -- the sm_100a equivalent of the reference tree's one native entry point,
 * `rope_2d(tokens, positions, base, fwd)` (uniception/models/libs/croco/curope/curope.cpp:49-69,
 * kernels.cu:17-108).  Token i, head h starts at tokens + i*token_pitch + h*head_dim (elements of
 * `dtype`), so the q and k thirds of a packed [rows, 3*heads*head_dim] qkv buffer are rotated where
 * they lie; positions [n_tokens, 2] int64 (y, x); fwd = -1 is the inverse rotation.
 * head_dim % 16 == 0 (curope: % 4), token_pitch % 4 == 0, tokens 16-byte aligned. */
int mab_rope2d(void* tokens, int dtype, long long n_tokens, int heads, int head_dim,
               long long token_pitch, const long long* positions, float base, float fwd, void* stream);
/* Image input path: what `load_images` (mapanything/utils/image.py:134-332)
 * does per decoded image -- PIL LANCZOS/BICUBIC resize so that the image contains the target
 * (mapanything/utils/cropping.py:230-244), centred crop (:441-447), ToTensor + Normalize
 * (image.py:297-300) -- bit-exact with Pillow's 8-bit fixed-point resampling.  rgb: DEVICE u8
 * [in_h,in_w,3]; mean/std: HOST float[3]; out: [3,out_h,out_w] fp32; out_u8 optional [out_h,out_w,3]
 * (the resized + cropped 8-bit image).  File decoding stays on the host. */
typedef struct {
  int in_w, in_h;          /* decoded image */
  int resize_w, resize_h;  /* floor(in * (max(target / in) + 1e-8)) */
  int filter;              /* 0 LANCZOS (shrinking), 1 BICUBIC */
  int crop_left, crop_top; /* (resize - target) // 2 */
  int out_w, out_h;        /* target */
} mab_image_plan;
int mab_image_plan_make(int in_w, int in_h, int target_w, int target_h, mab_image_plan* plan);
size_t mab_preprocess_workspace_bytes(const mab_image_plan* plan);
/* The resampling coefficient tables of a plan are computed at its first use and kept in pinned host memory
 * (one cudaHostAlloc per distinct plan), so every call is one asynchronous table upload + two kernels. */
int mab_preprocess_image(const unsigned char* rgb, const mab_image_plan* plan, const float* mean_host,
                         const float* std_host, float* out, unsigned char* out_u8, void* workspace,
                         size_t workspace_bytes, void* stream);

/* `recover_pinhole_intrinsics_from_ray_directions` (mapanything/utils/geometry.py:304-447, least-squares
 * branch): rays [views,H,W,3] -> K [views,3,3]; used by preprocess_inputs for views that carry
 * ray_directions instead of intrinsics (image.py:572-590). */
int mab_recover_intrinsics(const float* rays, int views, int H, int W, float* K_out, void* stream);

/* Depth maps of `preprocess_inputs` (image.py:335-675): cv2.INTER_NEAREST resize to
 * (resize_w, resize_h) (cropping.py:246-253) + crop (cropping.py:345-349), one fp32 channel. */
int mab_resize_nearest_crop(const float* src, int in_h, int in_w, int resize_w, int resize_h,
                            int crop_left, int crop_top, int out_w, int out_h, float* out, void* stream);

/* Dense export.  depth_z [views,H,W], intrinsics [views,3,3], camera_poses
 * [views,4,4] or NULL (camera frame) -> world_points [views,H,W,3] (optional; geometry.py:24-114
 * depthmap_to_world_frame) and the mask-filtered cloud the demos build on the CPU
 * (scripts/demo_local_weight.py:122-140, viz.py:233-239,275-279): pixels with mask != 0 (NULL = all)
 * and depth > 0, in view-major / row-major order (numpy boolean-mask order), out_xyz [capacity,3]
 * fp32 and out_rgb [capacity,3] u8 = (img_no_norm * 255).astype(uint8).  counts [views+1] int64 on
 * the device: kept pixels per view, then the total (points beyond `capacity` are counted, not stored). */
size_t mab_export_workspace_bytes(int views, int H, int W);
int mab_export_points(const float* depth_z, const float* intrinsics, const float* camera_poses,
                      const unsigned char* mask, const float* img_no_norm, int views, int H, int W,
                      float* world_points, float* out_xyz, unsigned char* out_rgb, long long capacity,
                      long long* counts, void* workspace, size_t workspace_bytes, void* stream);
/* Allocation-free variant of mab_export_points.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments and results as mab_export_points without the
 * workspace: the per-tile counts / offsets (12 bytes per tile of 1024 pixels) live in a file-scope
 * `static __device__` array of MAB_EXPORT_STATIC_MAX_TILES tiles (768 KiB, held by every GPU that loads
 * the library until exit).  views * ceil(H*W / 1024) above that limit is an argument error (-1, text in
 * mab_last_error()); 65536 tiles = 329 views of 392 x 518.  The call allocates nothing (no cudaMalloc /
 * cudaFree, no allocating library routine; the launch profiler does not bracket it).  Every call on a
 * device shares the ONE array, which keeps the previous call's contents (each call overwrites what it
 * reads, nothing needs clearing): two calls that overlap on the GPU -- different streams, or graphs
 * replayed concurrently -- race; calls ordered on one stream are safe. */
#define MAB_EXPORT_STATIC_MAX_TILES 65536
int mab_export_points_static(const float* depth_z, const float* intrinsics, const float* camera_poses,
                             const unsigned char* mask, const float* img_no_norm, int views, int H, int W,
                             float* world_points, float* out_xyz, unsigned char* out_rgb,
                             long long capacity, long long* counts, void* stream);
int mab_edge_mask(const float* pts3d, float* pts3d_cam, const unsigned char* fmask, int views, int H,
                  int W, float normal_tol_deg, float depth_rtol, void* workspace,
                  unsigned char* mask_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif
