/*
 * ggd.h — C ABI of the B200-native (sm_100a) decode → rasterise hot path.
 *
 * Every entry point takes plain pointers and sizes; device pointers are raw
 * CUDA device addresses, `stream` is a cudaStream_t passed as void*.  No torch
 * type crosses this boundary.  All functions return 0 on success and a
 * non-zero GGD_E_* code otherwise; ggd_last_error() returns a thread-local,
 * human-readable message for the last failure.  Nothing here allocates device
 * memory: the caller owns every buffer (the Python host side uses torch uint8
 * tensors so ownership matches the reference, whose three state buffers are
 * torch tensors returned to Python, DGR/rasterize_points.cu:68-78).
 *
 * Citations are into /root/reference; DGR abbreviates
 * gaussian_splatting/submodules/diff-gaussian-rasterization.
 */
#ifndef GGD_H_INCLUDED
#define GGD_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define GGD_API __attribute__((visibility("default")))
#else
#define GGD_API
#endif

#define GGD_OK 0
#define GGD_E_INVALID 1   /* bad argument / unsupported combination            */
#define GGD_E_CUDA 2      /* a CUDA runtime call or kernel launch failed        */
#define GGD_E_CAPACITY 3  /* a caller-provided buffer is too small              */

/* Tile size of the rasteriser. Part of the bit-exact contract: tile keys and
 * ranges depend on it (DGR/cuda_rasterizer/config.h:16-17). */
#define GGD_TILE_X 16
#define GGD_TILE_Y 16

GGD_API const char* ggd_last_error(void);
/* ABI version of this header; bumped on any signature change. */
GGD_API int ggd_abi_version(void);
/* Number of CUDA kernels this library has launched so far in this process (every entry point, device and host
 * thread; memsets and copies are not kernels).  Benchmarks read it before and after a timed region to report
 * how many kernels really ran.  No reference counterpart. */
GGD_API uint64_t ggd_launch_count(void);

/* ------------------------------------------------------------------------ */
/* Rasteriser forward                                                        */
/* ------------------------------------------------------------------------ */

/* Bytes required for the three state buffers.  Replaces
 * CudaRasterizer::required<GeometryState|BinningState|ImageState>
 * (DGR/cuda_rasterizer/rasterizer_impl.h:66-72) and the fromChunk layouts
 * (rasterizer_impl.cu:155-194).  `bin_capacity` is the number of
 * (Gaussian, tile) instances the binning buffer can hold; unlike the
 * reference the binning buffer is sized BEFORE the frame runs, which is what
 * removes the reference's mid-frame device→host sync (rasterizer_impl.cu:281).
 * sizes_out = { geom_bytes, binning_bytes, img_bytes }. */
GGD_API int ggd_raster_workspace_sizes(int P, int W, int H, uint64_t bin_capacity, size_t sizes_out[3]);

/* Stage 1: per-Gaussian projection, covariance, SH→RGB, culling, tile rects,
 * then the inclusive prefix sum of tiles_touched.  Replaces
 * FORWARD::preprocess → preprocessCUDA (DGR/cuda_rasterizer/forward.cu:155-256,
 * :402-455) and cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:277).
 * Pointer arguments follow CudaRasterizer::Rasterizer::forward
 * (DGR/cuda_rasterizer/rasterizer.h:31-54): exactly one of shs /
 * colors_precomp and exactly one of (scales, rotations) / cov3D_precomp is
 * non-NULL.  `radii` may be NULL (an internal array is used).
 * `num_rendered_host`, if non-NULL, receives the instance count R through an
 * asynchronous device→host copy on `stream` (valid after the stream is
 * synchronised; use pinned memory to keep the copy asynchronous).
 * `binning`: how stage 2 will order this frame's instances (GGD_BINNING_* below);
 * in modes 1 and 2 this stage also counts the instances per tile and marks the
 * geometry state accordingly, so ggd_raster_render can never mistake a state
 * that was prepared without counts for an empty frame. */
GGD_API int ggd_raster_preprocess(void* stream, int P, int D, int M, int W, int H,
                          const float* means3D, const float* shs, const float* colors_precomp,
                          const float* opacities, const float* scales, float scale_modifier,
                          const float* rotations, const float* cov3D_precomp,
                          const float* viewmatrix, const float* projmatrix, const float* campos,
                          float tan_fovx, float tan_fovy, int prefiltered,
                          void* geom_ws, int* radii, uint32_t* num_rendered_host, int binning);

/* Stage 2: key duplication, radix sort, tile ranges, blending.  Replaces
 * duplicateWithKeys (rasterizer_impl.cu:70-111), getHigherMsb (:35-50),
 * cub::DeviceRadixSort::SortPairs (:303-308), the ranges memset +
 * identifyTileRanges (:116-138, :310-318) and FORWARD::render → renderCUDA
 * (forward.cu:261-400).  R is read from geom_ws ON THE DEVICE; if R exceeds
 * bin_capacity nothing past the capacity is written, the overflow flag in
 * geom_ws is raised (see ggd_raster_status) and the image is undefined — the
 * caller re-runs this stage with a larger binning buffer.  `binning` should be the
 * mode stage 1 ran under; mode 1 on a state prepared in mode 0 falls back to the
 * radix sort on the device, mode 2 on such a state raises status flag 2. */
GGD_API int ggd_raster_render(void* stream, int P, int W, int H, const float* background,
                      const float* colors_precomp, const int* radii,
                      void* geom_ws, void* bin_ws, uint64_t bin_capacity, void* img_ws,
                      float* out_color, int binning);

/* Stages 1+2 back to back on `stream`, no host synchronisation.  Replaces
 * CudaRasterizer::Rasterizer::forward (rasterizer_impl.cu:198-337). */
GGD_API int ggd_raster_forward(void* stream, int P, int D, int M, const float* background, int W, int H,
                       const float* means3D, const float* shs, const float* colors_precomp,
                       const float* opacities, const float* scales, float scale_modifier,
                       const float* rotations, const float* cov3D_precomp,
                       const float* viewmatrix, const float* projmatrix, const float* campos,
                       float tan_fovx, float tan_fovy, int prefiltered,
                       void* geom_ws, void* bin_ws, uint64_t bin_capacity, void* img_ws,
                       float* out_color, int* radii, uint32_t* num_rendered_host, int binning);

/* Backward of ggd_raster_forward.  Replaces CudaRasterizer::Rasterizer::backward
 * (DGR/cuda_rasterizer/rasterizer_impl.cu:340-435): BACKWARD::render -> renderCUDA
 * (backward.cu:399-557), computeCov2DCUDA (:144-274), preprocessCUDA (:346-396).
 * Inputs are the forward's inputs plus its three state buffers (same bin_capacity)
 * and dL/dpixel [3][H][W].  Every output must be ZERO-FILLED by the caller, as the
 * reference's binding does (DGR/rasterize_points.cu:151-159); shapes follow it:
 * dL_dmean2D [P][3] (only .xy written), dL_dconic [P][2][2] (slots 0,1,3),
 * dL_dopacity [P], dL_dcolor [P][3], dL_dmean3D [P][3], dL_dcov3D [P][6],
 * dL_dsh [P][M][3], dL_dscale [P][3], dL_drot [P][4] (no quaternion-norm
 * Jacobian, like backward.cu:340).  Sums are floating point: results agree with
 * the reference to rounding (its own atomics are order-nondeterministic). */
GGD_API int ggd_raster_backward(void* stream, int P, int D, int M, const float* background, int W, int H,
                                const float* means3D, const float* shs, const float* colors_precomp,
                                const float* scales, float scale_modifier, const float* rotations,
                                const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                                const float* campos, float tan_fovx, float tan_fovy, const int* radii,
                                const void* geom_ws, const void* bin_ws, uint64_t bin_capacity, const void* img_ws,
                                const float* dL_dpix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                                float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                                float* dL_dscale, float* dL_drot);

/* Blocking read-back of { num_rendered, flags } from geom_ws (synchronises `stream`).  flags: 1 = more
 * instances than bin_capacity, 2 = mode 2 frame that needs the radix kernels, 4 = the per-tile counts in
 * geom_ws do not add up to num_rendered (state buffers of different frames were mixed), 8 = a tile list longer
 * than 24576 has more than 24576 instances inside 1/128 of its depth interval (the per-tile sort does not take
 * that: render the frame again with GGD_BINNING_RADIX); with any flag set the image is undefined.  The reference returns num_rendered from forward
 * (rasterizer_impl.cu:336); here it is a separate, optional query. */
GGD_API int ggd_raster_status(void* stream, const void* geom_ws, int P, uint32_t out[2]);

/* How stage 2 orders the instances — the `binning` argument of the three entry points above.
 * GGD_BINNING_AUTO (1): frames whose longest tile list is at most 98304 are binned per tile and each
 * list is sorted in shared memory (csrc/tile_binning.cu); other frames take the global radix sort;
 * the per-frame choice is made on the device.  GGD_BINNING_RADIX (0): always the radix sort.  Both
 * produce the bits of the reference's duplicateWithKeys + stable SortPairs + identifyTileRanges
 * (rasterizer_impl.cu:70-138, :303-318).  GGD_BINNING_TILES (2): like 1 but the radix kernels are not
 * even enqueued (saves their launches); a frame that turns out to need them raises flag 2 in
 * ggd_raster_status's second word, its image is undefined and the caller runs it again in mode 1 --
 * for frame loops whose scenes are known to qualify.  GGD_BINNING_DEFAULT (-1): the process default,
 * which starts from env GGD_BINNING=radix|auto|tiles (auto if unset).
 * The mode is an argument of each call (it used to be process-wide state: two callers could change it
 * under each other between stage 1 and stage 2).  No reference counterpart. */
#define GGD_BINNING_DEFAULT (-1)
#define GGD_BINNING_RADIX 0
#define GGD_BINNING_AUTO 1
#define GGD_BINNING_TILES 2
/* Or-ed into a mode: a frame binned per tile does not write the reference-order arrays "point_list_keys_unsorted" /
 * "point_list_unsorted" (duplicateWithKeys' output, 12 bytes per instance, read by nothing but ggd_raster_field
 * users).  On images of more than 4096 tiles it also sorts a long tile list only as far as the blend reads it (the
 * nearest depth chunk; a tile whose pixels are not saturated by then is sorted completely and blended again): image,
 * final T, contributor counts and the list prefix the backward replays are unchanged, "point_list" /
 * "point_list_keys" beyond that prefix are not in order.  For frame loops. */
#define GGD_BINNING_LEAN 4
/* Sets the process DEFAULT used by calls that pass GGD_BINNING_DEFAULT; pass -1 to query.  Returns the
 * previous default. */
GGD_API int ggd_raster_binning(int mode);

/* Blocking read-back of { binned_per_tile, longest_tile_list } of the last frame in geom_ws
 * (longest_tile_list is only counted in modes 1 and 2).  No reference counterpart. */
GGD_API int ggd_raster_binning_info(void* stream, const void* geom_ws, int P, uint32_t out[2]);

/* Frustum test only.  Replaces Rasterizer::markVisible → checkFrustum
 * (rasterizer_impl.cu:54-66, :141-152).  `present` is P bytes (bool). */
GGD_API int ggd_mark_visible(void* stream, int P, const float* means3D, const float* viewmatrix,
                     const float* projmatrix, uint8_t* present);

/* Field locator for tests/debugging: byte offset and element count of a named
 * array inside a state buffer.  The reference exposes its layout only through
 * fromChunk (rasterizer_impl.cu:155-194).  which: 0 geom, 1 binning, 2 img.
 * geom: "depths" f32[P], "radii" i32[P], "means2D" f32[2P], "cov3D" f32[6P],
 *       "conic_opacity" f32[4P], "rgb" f32[3P], "tiles_touched" u32[P],
 *       "point_offsets" u32[P], "clamped" u8[3P], "num_rendered" u32[2]
 * binning (needs R = num_rendered ≤ capacity): "point_list" u32[R],
 *       "point_list_unsorted" u32[R], "point_list_keys" u64[R],
 *       "point_list_keys_unsorted" u64[R]
 * img:  "accum_alpha" f32[WH], "n_contrib" u32[WH], "ranges" u32[2*tiles] */
GGD_API int ggd_raster_field(int which, const char* name, int P, int W, int H, uint64_t bin_capacity,
                     size_t* byte_offset, size_t* count);

/* Per-stage device timing for benchmarks (the reference has only the debug
 * sync-after-each-stage CHECK_CUDA, DGR/cuda_rasterizer/auxiliary.h:166-173).
 * When enabled, every entry point records CUDA events around each stage on the
 * caller's stream; ggd_stage_times waits for them and returns milliseconds of
 * the LAST recorded occurrence, -1 for a stage that did not run.  Order:
 * planes_channels_last, decode, preprocess, scan, duplicate, sort, ranges, blend. */
GGD_API int ggd_stage_timing(int enable);
GGD_API int ggd_stage_times(float ms_out[8]);

/* ------------------------------------------------------------------------ */
/* Stand-alone stable LSD radix sort of (u64 key, u32 value) pairs           */
/* ------------------------------------------------------------------------ */

/* Replaces the reference's call into cub::DeviceRadixSort::SortPairs
 * (CUB of CUDA 12.x, not vendored; call site rasterizer_impl.cu:303-308):
 * stable sort on key bits [begin_bit, end_bit).  `n_dev` points to the element
 * count in DEVICE memory (so it can be produced by an earlier kernel without a
 * host sync); `capacity` bounds it.  keys_in/vals_in are preserved; the sorted
 * result is written to keys_out/vals_out.  `scratch` must hold
 * ggd_sort_scratch_bytes(capacity) bytes. */
GGD_API size_t ggd_sort_scratch_bytes(uint64_t capacity);
GGD_API int ggd_sort_pairs(void* stream, const uint64_t* keys_in, uint64_t* keys_out,
                   const uint32_t* vals_in, uint32_t* vals_out,
                   const uint32_t* n_dev, uint64_t capacity, int begin_bit, int end_bit,
                   uint64_t* keys_tmp, uint32_t* vals_tmp, void* scratch, size_t scratch_bytes);


/* Same sort for callers that cannot allocate device memory.  NO COUNTERPART IN
 * THE REFERENCE (its only sort is the CUB call above, whose temp storage the
 * reference takes from a torch tensor it grows on demand).
 *
 * Who needs this: a host that owns the whole device-memory budget itself — a
 * viewer or serving process that splits the GPU into fixed arenas at start-up, or
 * an engine embedding this library as a plug-in — can pass keys, values and the
 * two ping-pong buffers out of its own arena, but has no way to obtain the
 * sort's bookkeeping memory, and a cudaMalloc in the middle of its frame loop
 * would stall every stream it runs.  Here that bookkeeping (digit histograms,
 * tile tickets, look-back words, about 4 MB) is a fixed array inside the library
 * image; each GPU that loads the library holds it until the process exits.  The
 * call makes no cudaMalloc/cudaFree and calls no allocating library routine.
 *
 *  - capacity may not exceed GGD_SORT_STATIC_MAX_PAIRS; larger requests fail
 *    with GGD_E_CAPACITY and a message naming the limit.
 *  - The first call on each device looks up the array's device address (one-time
 *    set-up, with that device current).  Make that first call during start-up,
 *    before any stream capture.
 *  - All calls on a device share the one array, and it keeps whatever the
 *    previous call left in it; each call zeroes what it needs first, on `stream`.
 *    Two calls whose GPU work can overlap (different streams, or different host
 *    threads on the same device) therefore race on it: order them yourself.
 *  Arguments and results are exactly those of ggd_sort_pairs. */
#define GGD_SORT_STATIC_MAX_PAIRS (1u << 21)
GGD_API int ggd_sort_pairs_static(void* stream, const uint64_t* keys_in, uint64_t* keys_out,
                                  const uint32_t* vals_in, uint32_t* vals_out,
                                  const uint32_t* n_dev, uint64_t capacity, int begin_bit, int end_bit,
                                  uint64_t* keys_tmp, uint32_t* vals_tmp);

/* ------------------------------------------------------------------------ */
/* Fused tri-plane sampling + five-head decoder MLP                          */
/* ------------------------------------------------------------------------ */

/* Static description of a decoder.  Mirrors the constructor / forward
 * arguments of the reference decoder classes
 * (main/decoder_models/parallel_decoder.py:13-36, sequential_decoder.py:13-36,
 * sequential_decoder_reverse.py:13-36) and the generator's rendering_kwargs
 * (box_warp, triplane_depth). */
typedef struct ggd_decoder_config {
    int arch;          /* 0 = EG3D tri-planes (bilinear), 1 = PanoHead tri-grids (trilinear)  */
    int variant;       /* 0 = parallel, 1 = sequential, 2 = sequential_reverse                */
    int plane_depth;   /* triplane_depth D (1 for EG3D)                                       */
    int plane_h, plane_w;
    int pos_freqs;     /* 0 = raw xyz, 10 = Embedder(num_freqs=10) -> 63 dims                 */
    int hidden;        /* must be 128, the width the kernels are specialised for; the host classes run a
                          narrower decoder by zero-padding its weights into this layout              */
    float box_warp;
    float scale_bias;  /* activate_scale offset: -2 (parallel, sequential), -2.5 (reverse)    */
    int precision;     /* 0 = fp32 CUDA-core path (max-abs 1e-3 contract),
                          1 = fp16 tcgen05 tensor-core path, fp32 accumulation (relative 1e-2 contract) */
} ggd_decoder_config;

/* Number of floats of (a) the flat parameter buffer in nn.Linear layout, heads
 * in get_params_custom() order xyz, scale, rotation, opacity, color, each
 * W0,b0,..,W3,b3 with W[out][in] (main/decoder_models/base_decoder.py:11-19),
 * and (b) the packed buffer ggd_decode_pack_weights produces. */
GGD_API int ggd_decode_param_counts(const ggd_decoder_config* cfg, size_t* flat_floats, size_t* packed_floats);
/* Offset (in floats) of one tensor inside the flat buffer: head 0..4, layer 0..3,
 * is_bias 0/1; also returns its fan-in and fan-out. */
GGD_API int ggd_decode_param_offset(const ggd_decoder_config* cfg, int head, int layer, int is_bias,
                                    size_t* offset, int* fan_in, int* fan_out);
/* Re-lays the flat parameters out for the kernels (transposed, zero padded).
 * No counterpart call in the reference (ATen reads nn.Linear weights in place). */
GGD_API int ggd_decode_pack_weights(void* stream, const ggd_decoder_config* cfg,
                                    const float* flat_params, float* packed);
/* planes: the generator's feature_planes [3][32*D][H][W] fp32 (batch 1;
 * eg3d/training/triplane.py:93, PanoHead/training/triplane.py:224) ->
 * channels-last [3][D][H][W][32], fp32 (precision 0: one texel = one 128-byte
 * line) or fp16 (precision 1: 64 bytes per texel). */
GGD_API int ggd_planes_channels_last(void* stream, const ggd_decoder_config* cfg,
                                     const float* planes_nchw, void* planes_cl);
/* sample_from_planes + (Embedder) + 5 x Decoder + scale/xyz post-math, and
 * optionally the GaussianModel activations, for P points.  Replaces
 * {Parallel,Sequential,SequentialReverse}Decoder.forward after G.synthesis
 * (parallel_decoder.py:41-80) and gaussian_model.py:99-121.  Outputs are the
 * reference's result fields: xyz[P][3] scale[P][3] rotation[P][4]
 * opacity[P][1] color[P][3]; act_* may be NULL. */
GGD_API int ggd_decode_forward(void* stream, const ggd_decoder_config* cfg, const void* planes_cl,
                               const float* xyz, int P, const float* packed,
                               float* out_xyz, float* out_scale, float* out_rotation,
                               float* out_opacity, float* out_color,
                               float* act_scale, float* act_rotation, float* act_opacity);

/* Diagnostics: cycle accounting of the tcgen05 decoder's roles for CTA 0 of the last ggd_decode_forward launched with
 * env GGD_DECODE_PROF=1 (an instrumented instantiation; the default build of the kernel carries no counters).
 * out = { kernel, sampling, waiting for an accumulator, hidden-layer epilogues, last-layer epilogues (one epilogue
 * warp); MMA thread: waiting for an operand, waiting for weights, issuing, of which layer-0 steps, last-layer steps;
 * two spare words }.  Synchronises the device.
 * No reference counterpart. */
GGD_API int ggd_debug_decode_profile(uint64_t out[12]);

/* One decoder head on caller-supplied features: mean over the three plane
 * samples, concat with gaussian_features, Linear-GELU x3, Linear.  Replaces
 * Decoder.forward(triplane_features[3][P][32], gaussian_features[P][k]) ->
 * [P][out] (main/decoder_models/base_decoder.py:21-27) when that module is called
 * on its own; weights are the nn.Linear tensors themselves (W[out][in]). */
GGD_API int ggd_mlp_forward(void* stream, const float* tri_features, const float* gaussian_features, int k, int P,
                            const float* W0, const float* b0, const float* W1, const float* b1, const float* W2,
                            const float* b2, const float* W3, const float* b3, int hidden, int out_features, float* out);

/* Backward of ggd_decode_forward: gradients of the decoder parameters (flat
 * nn.Linear layout, same offsets as ggd_decode_param_offset) and, optionally, of
 * the feature planes in the generator's own layout [3][32*D][H][W] (what
 * grid_sample backward scatters into; needed when use_gen_finetune trains G).
 * Replaces PyTorch autograd through the reference decoder classes
 * (main/decoder_models/base_decoder.py:21-27 x5, parallel_decoder.py:58-80 and
 * the sequential variants) and grid_sample backward.  No gradient flows to
 * init_position, as in the reference (it is an input).
 *   planes_cl_f32 : fp32 channels-last planes (ggd_planes_channels_last, precision 0); may be NULL when planes_cl_f16 is
 *                   given and the tensor-core chain runs (precision 1, pos_freqs 0): it reads only the fp16 planes
 *   planes_cl_f16 : optional (may be NULL): the fp16 channels-last planes the fp16 forward sampled (precision 1).  The
 *                   tensor-core chain then recomputes the forward from exactly what the forward read, at half the bytes
 *                   (measured: its plane sampling drops from 29 % of the kernel to 5 %); ignored by the fp32 chain
 *   packed_f32    : ggd_decode_pack_weights output for precision 0
 *   out_*         : the forward results (the chained variants feed them to later heads)
 *   dL_d*         : gradients of the five results; any may be NULL (= zero)
 *   scratch       : ggd_decode_backward_scratch_bytes(cfg, P) bytes of staging
 *   dflat         : [flat_floats] overwritten;  dplanes: overwritten if non-NULL
 * cfg->precision selects the flavour: 0 = exact fp32 CUDA-core chain; 1 = tcgen05 chain
 * (fp16 operands, fp32 accumulation; decoders without the positional embedding — with it
 * the fp32 chain is used).  Either way the forward is recomputed inside, the planes are the
 * fp32 channels-last ones and the weight gradients are accumulated on tensor cores. */
GGD_API size_t ggd_decode_backward_scratch_bytes(const ggd_decoder_config* cfg, int P);
GGD_API int ggd_decode_backward(void* stream, const ggd_decoder_config* cfg, const float* planes_cl_f32,
                                const void* planes_cl_f16, const float* xyz, int P, const float* flat_params, const float* packed_f32,
                                const float* out_xyz, const float* out_scale, const float* out_rotation,
                                const float* out_opacity, const float* out_color,
                                const float* dL_dxyz, const float* dL_dscale, const float* dL_drotation,
                                const float* dL_dopacity, const float* dL_dcolor,
                                void* scratch, size_t scratch_bytes, float* dflat, float* dplanes);

/* Backward of the activations the rasteriser's inputs went through: scaling = exp(scale), rotation =
 * F.normalize(rotation) (eps 1e-12), opacity = sigmoid(opacity) — the GaussianModel getters
 * (gaussian_splatting/scene/gaussian_model.py:99-121), three autograd nodes in the reference's training loop
 * (main/train_pano2gaussian_decoder.py:223-230).  In place: on entry g_* hold dL/d(activated value) as
 * ggd_raster_backward wrote them (dL_dscale, dL_drot, dL_dopacity); on return they hold dL/d(raw decoder output),
 * ready for ggd_decode_backward.  `rotation` is the RAW decoder output, act_* the activated values. */
GGD_API int ggd_activation_backward(void* stream, int P, const float* act_scale, const float* rotation,
                                    const float* act_opacity, float* g_scale, float* g_rotation, float* g_opacity);

/* ---- point-set producer -------------------------------------------------
 * Iso-surface points of a density grid, on the GPU.  Replaces the CPU detour of
 * main/decoder_utils/target_dataloader.py:96-118 + :185-190 (skimage.measure.marching_cubes on the
 * 128^3 sigma grid, trimesh, one random barycentric point per face repeated until n_points exist,
 * truncation, radial scale clip(randn * thickness + 1, 0, 1)).  sigma is [res][res][res] fp32 in the
 * reference's index order; a mesh vertex at index coordinates v maps to v / res - 0.5.  The cells are
 * tessellated with six tetrahedra each (no case tables), point k lies on triangle k mod n_triangles,
 * random numbers are counter-based from `seed`: results are reproducible and match
 * oracle/surface_oracle.py; against the reference the agreement is statistical (different
 * triangulation and random stream).  scratch: ggd_surface_points_scratch_bytes(res), 256-byte
 * aligned.  num_triangles_host (optional, pinned) receives the triangle count asynchronously;
 * with no triangle at all the points are zeros.  points may be null when n_points is 0. */
GGD_API size_t ggd_surface_points_scratch_bytes(int res);
GGD_API int ggd_surface_points(void* stream, const float* sigma, int res, float level, uint64_t seed, int n_points,
                               float thickness, void* scratch, size_t scratch_bytes, float* points,
                               uint32_t* num_triangles_host);

#ifdef __cplusplus
}
#endif
#endif /* GGD_H_INCLUDED */
