/*
 * gsb200.h -- C ABI of the B200-native Gaussian-splat hot path (libgsb200.so).
 *
 * Plain pointers and sizes only: no torch, no C++ types.  Every pointer is a DEVICE pointer unless
 * the parameter name ends in `_host`.  Every entry point takes the CUDA stream it must run on as a
 * `void*` (a cudaStream_t) and returns 0 on success or a negative GSB_E* code; nothing here
 * synchronises the device except where stated.
 *
 * Each entry point cites the reference interface it replaces.  "DGR/" is
 * /root/reference/diff-gaussian-rasterization-w-depth/ ; "cr/" is DGR/cuda_rasterizer/.
 */
#ifndef GSB200_H
#define GSB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSB_OK 0
#define GSB_EINVAL (-1)   /* bad argument (null pointer, non-positive size, unsupported channel count) */
#define GSB_ECUDA (-2)    /* a CUDA runtime call or launch failed; see gsb_last_error() */
#define GSB_ENOMEM (-3)   /* a buffer callback returned NULL / capacity exceeded */
#define GSB_EPREFILTER (-4) /* prefiltered=1 was promised but a Gaussian was culled (cr/auxiliary.h:154-161) */

#define GSB_TILE 16       /* cr/config.h:16-17 BLOCK_X = BLOCK_Y = 16 */
#define GSB_MAX_CHANNELS 3 /* cr/config.h:15 NUM_CHANNELS */

/* Growable byte buffer callback: replaces std::function<char*(size_t)> of
 * Rasterizer::forward (cr/rasterizer.h:32-34) / resizeFunctional (DGR/rasterize_points.cu:27-33).
 * Must return a device pointer to at least `bytes` bytes (256-B aligned), valid until backward. */
typedef char* (*gsb_alloc_fn)(void* user, size_t bytes);

/* One camera (or a batch of n_cams cameras that share W, H, bg).  Replaces the argument list of
 * Rasterizer::forward (cr/rasterizer.h:31-54).  Matrices are the reference's transposed (column-major
 * flat) 4x4s, `n_cams` of them back to back. */
typedef struct gsb_raster_args {
  int32_t P;               /* Gaussians */
  int32_t n_cams;          /* 1 for the drop-in path */
  int32_t width, height;
  int32_t prefiltered;
  int32_t color_cam_stride; /* 0: one colour table [P][3] shared by all cameras; else floats between tables */
  float scale_modifier;
  const float* tanfovx_host; /* HOST [n_cams] */
  const float* tanfovy_host; /* HOST [n_cams] */
  const float* background;   /* [3] */
  const float* means3D;      /* [P][3] */
  const float* colors;       /* [P][3] colors_precomp (SH path is dead code in the reference repo: shared.py:32,121) */
  const float* opacities;    /* [P] */
  const float* scales;       /* [P][3] or NULL when cov3D_precomp is given */
  const float* rotations;    /* [P][4] (r,x,y,z) or NULL */
  const float* cov3D_precomp;/* [P][6] or NULL */
  const float* viewmatrix;   /* [n_cams][16] */
  const float* projmatrix;   /* [n_cams][16] */
  /* Tile-band sharding (new; NULL = every entry is a whole camera).  DEVICE int32 [n_cams][4] per batch entry:
   * (tile_row0, tile_row1, split_slot, count_visibility).  The entry renders only the tile rows [tile_row0, tile_row1) of
   * its camera: instance lists exist for those tiles only, all other tiles are empty (background).  Per-Gaussian
   * state (radii, means2D, conics) is the whole camera's, so a Gaussian's gradients are the sums over the rendered
   * tiles.  split_slot >= 0 marks a camera whose rows are shared with another rank: gsb_raster_backward_flat then
   * writes the entry's screen-space gradient into the extra columns 20 + 2 slot, 21 + 2 slot of the flat record instead
   * of adding its norm to col 14 (the norm must be taken of the camera's SUMMED gradient, after the exchange), and
   * only the entry with count_visibility != 0 counts the camera in cols 15 / 16. */
  const int32_t* entries;
  /* Two colour tables in one walk (new; NULL = one render).  colors2 [P][3]: a second colour table of the SAME
   * Gaussians -- the segmentation colours that densify.py:136-150 renders right after the colours, through the same
   * camera -- blended by the same forward pass into out_color2 [n_cams][3][H][W] (the two renders share every alpha,
   * transmittance and termination decision, so each image equals its own render).  gsb_raster_backward_flat then takes
   * the second image's gradient as dL_dpix2 and back-propagates both in one walk. */
  const float* colors2;
  float* out_color2;
} gsb_raster_args;

/* Saved state of one forward call: replaces the (geomBuffer, binningBuffer, imgBuffer, num_rendered)
 * tuple the reference hands back to Python (DGR/rasterize_points.cu:114) and re-decodes with
 * fromChunk (cr/rasterizer_impl.cu:155-194, 373-375).  POD; pointers point into the three buffers.
 *
 * Binning layout.  The reference sorts R instances on 64-bit (tile << 32 | depth) keys, 6 radix passes
 * over 12 B records.  A stable LSD sort is digit-separable and the low 32 key bits are the same for all
 * instances of a Gaussian, so the depth digits are sorted BEFORE duplication, on the n_cams*P Gaussians
 * (one segment per camera, 32-bit keys), instances are emitted in that order with their camera-local
 * tile id as a 16-bit key, and one segmented pass with a 10-bit digit (two 8-bit passes above 1024
 * tiles per camera) puts them in tile order.  point_list and ranges are bit-identical to the
 * reference's; in the one-pass case the sorted tile ids are never materialised (the per-camera digit
 * histogram is the per-tile count, so ranges are its exclusive scan). */
typedef struct gsb_raster_ctx {
  int64_t num_rendered;      /* R: total (Gaussian, tile) instances over all cameras */
  int32_t P, n_cams, width, height, tiles_x, tiles_y;
  int32_t sort_bits;         /* 32 + getHigherMsb(tiles): what the reference passes to SortPairs */
  int32_t tile_bits;         /* getHigherMsb(tiles per camera): bits sorted at instance granularity */
  /* geometry state, [n_cams][P] each (cr/rasterizer_impl.h:31-45); cov3D is [P][6], one table for all cameras */
  float* depths; int32_t* radii; float* means2D; float* cov3D; float* conic_opacity;
  uint32_t* tiles_touched;
  uint32_t* gauss_order;     /* Gaussian index per sorted slot: camera-major, depth-ascending, stable;
                                culled Gaussians (depth key 0xFFFFFFFF) sort behind every visible one */
  uint32_t* point_offsets;   /* inclusive scan of tiles_touched taken in gauss_order */
  uint32_t* seg_start;       /* [n_cams + 1]: camera c owns point_list[seg_start[c] .. seg_start[c+1]) */
  /* image state (cr/rasterizer_impl.h:47-54) */
  float* final_T; uint32_t* n_contrib; uint32_t* ranges; /* ranges: [n_cams*tiles][2] */
  /* binning state (cr/rasterizer_impl.h:56-66) */
  uint32_t* point_list;               /* sorted Gaussian indices == the reference's point_list */
  uint32_t* point_list_unsorted;      /* emission order (camera-major, depth order, row-major rect) */
  uint16_t* tile_keys_unsorted;       /* camera-local tile id per emitted instance */
  uint16_t* tile_keys_sorted;         /* sorted camera-local tile ids; NULL when one pass suffices */
  void* sort_temp;
  /* capacity-guarded path (gsb_raster_forward_async): the binning buffer holds `capacity` instances; the instance
   * count is only known on the device.  status (DEVICE int32[4]): [0] error code of the call (0, GSB_EPREFILTER,
   * GSB_ENOMEM when the instances did not fit), [1] num_rendered, [2] instances that would have been needed.
   * On the blocking path capacity == num_rendered. */
  int64_t capacity;
  int32_t* status;
} gsb_raster_ctx;

/* Library / build identification. */
const char* gsb_version(void);
const char* gsb_last_error(void);            /* text of the last CUDA error seen by this thread */

/* Byte sizes of the three state buffers: replaces required<GeometryState/ImageState/BinningState>
 * (cr/rasterizer_impl.h:68-73). */
size_t gsb_geom_bytes(int32_t P, int32_t n_cams);
size_t gsb_image_bytes(int32_t width, int32_t height, int32_t n_cams);
size_t gsb_binning_bytes(int64_t num_rendered, int32_t tiles_per_cam, int32_t n_cams);

/* getHigherMsb (cr/rasterizer_impl.cu:35-50). */
uint32_t gsb_higher_msb(uint32_t n);

/* Rasterizer::forward (cr/rasterizer.h:31-54, cr/rasterizer_impl.cu:198-339).
 * out_color [n_cams][3][H][W], out_depth [n_cams][1][H][W], radii [n_cams][P] (may be NULL).
 * Blocks the host once (4-byte D2H of num_rendered, like cr/rasterizer_impl.cu:281-282).
 * Returns num_rendered in ctx->num_rendered. */
int gsb_raster_forward(const gsb_raster_args* args, gsb_alloc_fn geom_alloc, gsb_alloc_fn binning_alloc,
                       gsb_alloc_fn image_alloc, void* user, float* out_color, float* out_depth,
                       int32_t* radii, gsb_raster_ctx* ctx, void* stream);

/* Rasterizer::forward without the host in the loop (the reference blocks on a 4-byte
 * D2H of num_rendered per call, cr/rasterizer_impl.cu:281-282).  The three state buffers are caller-owned arenas
 * (>= gsb_geom_bytes / gsb_image_bytes / gsb_binning_bytes(capacity, ...) bytes, 256-B aligned, reusable from call
 * to call); `capacity` bounds the (Gaussian, tile) instances.  Nothing synchronises and nothing is copied from
 * pageable or pinned host memory, so the call can be captured into a CUDA graph: the per-camera scalars come from
 * tanfov_dev (DEVICE float [n_cams][2] = tanfovx, tanfovy; args->tanfov*_host are ignored).  If the instances do not
 * fit, ctx->status[0] becomes GSB_ENOMEM, ctx->status[2] the count that was needed, every tile list is empty (the
 * images are pure background, all gradients zero) and nothing is written out of bounds: the caller reads the
 * status words whenever it next synchronises and repeats the step with a larger arena.  ctx->num_rendered is -1
 * (unknown on the host); ctx->status[1] holds it on the device. */
int gsb_raster_forward_async(const gsb_raster_args* args, const float* tanfov_dev, char* geom, size_t geom_bytes,
                             char* image, size_t image_bytes, char* binning, size_t binning_bytes, int64_t capacity,
                             float* out_color, float* out_depth, int32_t* radii, gsb_raster_ctx* ctx, void* stream);

/* Work counters of a forward state, per camera: counts (DEVICE uint64 [n_cams][6]) = (sum over pixels of n_contrib,
 * pairs with alpha >= 1/255 that were blended -- B and B_alpha of the metric definition --, pairs the inner loop
 * evaluated for live pixels, lane slots it issued (32 per warp iteration), list entries staged after tile culling,
 * list entries); a measurement entry point (re-runs the forward walk without writing images). */
int gsb_raster_count_pairs(const gsb_raster_args* args, const gsb_raster_ctx* ctx, uint64_t* counts, void* stream);

/* Re-blend an existing forward state with different colours: everything a forward call computes before the
 * blend (preprocess, sorts, instance lists, tile ranges) depends only on the geometry, opacities and cameras, so
 * a second render of the SAME Gaussians through the SAME cameras -- the segmentation render that follows every
 * colour render in densify.py:116-150 -- only needs the blend kernel: no sorts and no host synchronisation.
 * `a` = the forward call's arguments with `colors` replaced; `ctx` = the state that call left behind (its
 * final_T / n_contrib are rewritten with identical values).  gsb_raster_backward then takes the same `ctx`
 * together with these arguments. */
int gsb_raster_forward_colors(const gsb_raster_args* a, const gsb_raster_ctx* ctx, float* out_color, float* out_depth,
                              void* stream);

/* Rasterizer::backward (cr/rasterizer.h:56-83, cr/rasterizer_impl.cu:343-436).
 * dL_dpix [n_cams][3][H][W].  Outputs (all written in full, no pre-zeroing required):
 *   dL_dmean2D [n_cams][P][3] (z = 0), dL_dconic [n_cams][P][4] (scratch: on return it holds the raw second
 *   moments sum_pixels dL/dG G (dx dx, dx dy, -, dy dy); dL/dconic = -0.5 x these, consumed internally),
 *   dL_dopacity [P], dL_dcolor [P][3] (or [n_cams][P][3] if color_cam_stride != 0),
 *   dL_dmean3D [P][3], dL_dcov3D [P][6] (n_cams==1 only, else may be NULL), dL_dscale [P][3], dL_drot [P][4]
 * summed over the cameras of the batch. */
int gsb_raster_backward(const gsb_raster_args* args, const gsb_raster_ctx* ctx, const float* dL_dpix,
                        float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                        float* dL_dmean3D, float* dL_dcov3D, float* dL_dscale, float* dL_drot, void* stream);

/* The backward of a camera-batched training step, written for the exchange: instead of separate gradient tables the
 * step's whole per-Gaussian record goes into ONE row-major buffer flat [P][GSB_FLAT_WIDTH] that the multi-GPU
 * exchange reduces in place and the optimiser consumes:
 *   cols 0-2 dL/dmeans3D, 3-5 dL/dcolors, 6-9 dL/drotations, 10 dL/dopacity, 11-13 dL/dscales   (summed over cameras)
 *   col 14  sum over the batch's cameras with radius > 0 of ||dL/dmean2D.xy||   (accumulate_mean_2d_gradients,
 *           external.py:113-124: the norm is taken per camera, BEFORE any reduction)
 *   col 15  number of the batch's cameras with radius > 0                       (visibility_count increment)
 *   col 16  max over the batch's cameras of the radius                          (max_2d_radii, densify.py:154-162)
 *   cols 17-19 dL/d(second colour table) -- the segmentation render of densify.py:136-150
 * mode 0: first render of the iteration, every column written (17-19 zero).  mode 1: a second render of the same
 * Gaussians with another colour table: cols 0-2, 6-13 are added to, 17-19 written, 3-5 and 14-16 left alone.
 * mode 2: another camera group of the SAME step (a caller that renders its cameras as several batches on concurrent
 * streams, so that one batch's long tiles overlap another's bulk): every column is added to, col 16 maximised.
 * Adding GSB_FLAT_BLEND_ONLY runs only the per-tile blend backward (into scratch), GSB_FLAT_FINISH_ONLY only the
 * per-Gaussian pass from scratch into flat -- the two halves of one call, so that a caller can run the blends of
 * several groups concurrently and order only the short passes that touch flat.
 * flat_stride: floats per row, a multiple of 4, >= GSB_FLAT_WIDTH; columns from 20 on are the split-camera slots of
 * tile-band sharding (gsb_raster_args.entries), zero where this batch has nothing to say.
 * scratch: gsb_backward_scratch_bytes(P, n_cams) bytes.  dL_dmean2D_out (optional, [n_cams][P][3]) receives the
 * per-camera screen-space gradients like gsb_raster_backward. */
#define GSB_FLAT_WIDTH 20
#define GSB_FLAT_COL_GRADNORM 14
#define GSB_FLAT_COL_VISCOUNT 15
#define GSB_FLAT_COL_MAXRADIUS 16
#define GSB_FLAT_BLEND_ONLY 0x100
#define GSB_FLAT_FINISH_ONLY 0x200
size_t gsb_backward_scratch_bytes(int32_t P, int32_t n_cams);
int gsb_raster_backward_flat(const gsb_raster_args* args, const gsb_raster_ctx* ctx, const float* dL_dpix,
                             void* scratch, float* flat, int32_t flat_stride, int32_t mode, float* dL_dmean2D_out,
                             void* stream);

/* The backward of a two-table forward (gsb_raster_args.colors2): dL_dpix of the first image and dL_dpix2 of the second
 * are back-propagated in ONE walk of the lists (they share alpha, T and every decision; dL/dalpha is the sum over both
 * images' channels).  Writes every column of flat like mode 0, cols 17-19 = the second table's colour gradient: the
 * result equals gsb_raster_backward_flat mode 0 on the first render followed by mode 1 on a second render. */
int gsb_raster_backward_flat2(const gsb_raster_args* args, const gsb_raster_ctx* ctx, const float* dL_dpix,
                              const float* dL_dpix2, void* scratch, float* flat, int32_t flat_stride,
                              float* dL_dmean2D_out, void* stream);

/* DensificationVariables update from a (reduced) flat buffer: grad_accum += col 14, vis_count += col 15,
 * max_2d_radii = max(max_2d_radii, col 16) (shared.py:20-26, densify.py:154-162, external.py:113-124); every
 * split-camera slot (cols 20 + 2 j, 21 + 2 j < flat_stride) adds the norm of its summed gradient to grad_accum. */
int gsb_flat_apply_stats(int32_t P, const float* flat, int32_t flat_stride, float* grad_accum, float* vis_count,
                         float* max_2d_radii, void* stream);

/* Multi-GPU exchange over NVLink peer memory (new; the reference has no distributed code).  One process per GPU;
 * every rank allocates its exchange block with gsb_comm_alloc, exports it (64-byte CUDA IPC handle), the hosts swap
 * handles (any side channel) and open the peers' blocks.  gsb_allreduce_flat then reduces `n_float4` 16-byte words
 * of the block in place across `world` ranks in ONE kernel: every rank pushes its part of slice q into rank q's
 * staging rows (P2P stores), rank r reduces slice r in rank order -- SUM, except component `max_comp` of every
 * `row_float4`-th word counted from word `max_word`, which takes MAX (col 16 of a flat row) -- and pushes the result
 * into every peer's block (P2P stores); ranks meet through flag words in the blocks themselves (release/acquire at
 * system scope), no host involvement, graph-capturable.  All ranks end with bit-identical contents.  peers_host: HOST array of `world`
 * device pointers (own block at index rank); every block is gsb_comm_bytes(payload_bytes) long, payload first. */
#define GSB_COMM_MAX_RANKS 16
size_t gsb_comm_bytes(size_t payload_bytes);
int gsb_comm_alloc(size_t bytes, void** block, unsigned char* handle64);   /* cudaMalloc + zero + cudaIpcGetMemHandle */
int gsb_comm_open(const unsigned char* handle64, void** peer_block);       /* cudaIpcOpenMemHandle (peer access) */
int gsb_comm_close(void* peer_block);
int gsb_comm_free(void* block);
int gsb_allreduce_flat(int32_t rank, int32_t world, void* const* peers_host, size_t payload_bytes, int64_t n_float4,
                       int32_t row_float4, int32_t max_word, int32_t max_comp, void* stream);

/* Rasterizer::markVisible (cr/rasterizer.h:24-29, cr/rasterizer_impl.cu:141-153). present: uint8 [P]. */
int gsb_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream);

/* Fused L1 + SSIM loss (replaces external.calc_ssim, external.py:49-110, torch l1_loss and the
 * 0.8/0.2 mix of densify.py:127-129 / train.py:362-364).  img, target: [n][3][H][W].
 * loss_out: float [n][3] = (l1 mean, ssim mean, w_l1*l1 + w_ssim*(1-ssim)) per image.
 * If dL_dimg != NULL it receives grad_scale * d loss/d img.  scratch: gsb_ssim_scratch_bytes(). */
size_t gsb_ssim_scratch_bytes(int32_t n, int32_t height, int32_t width);
int gsb_ssim_l1(int32_t n, int32_t height, int32_t width, const float* img, const float* target,
                float w_l1, float w_ssim, float grad_scale, float* loss_out, float* dL_dimg,
                void* scratch, void* stream);
/* The same for image BANDS (tile-band sharding): rows = DEVICE int32 [n][4] per image (loss_y0, loss_y1, write_y0,
 * write_y1).  Only the pixel rows [loss_y0, loss_y1) enter the two means (still normalised by the whole image's
 * 3 H W, so the bands of an image add up to its loss; the third number is w_l1 l1 + w_ssim ((loss_y1 - loss_y0) / H - ssim));
 * the image must be valid on [loss_y0 - 5, loss_y1 + 5), and dL_dimg is written on [write_y0, write_y1) (zero
 * beyond loss +- 5) -- the rows of the tiles the band's rasterizer backward reads. */
int gsb_ssim_l1_rows(int32_t n, int32_t height, int32_t width, const float* img, const float* target, const int32_t* rows,
                     float w_l1, float w_ssim, float grad_scale, float* loss_out, float* dL_dimg, void* scratch,
                     void* stream);

/* Densification statistics: update_max_2d_radii_and_visibility_mask (densify.py:154-162) fused with
 * accumulate_mean_2d_gradients (external.py:113-124).  radii [P], grad_means2D [P][3]. */
int gsb_densify_accumulate(int32_t P, const int32_t* radii, const float* grad_means2D, float* max_2d_radii,
                           float* grad_accum, float* vis_count, uint8_t* visible, void* stream);

/* The same for the n_cams cameras of a batch in one launch: radii [n_cams][P], grad_means2D [n_cams][P][3];
 * equals n_cams successive gsb_densify_accumulate calls (same accumulation order). */
int gsb_densify_accumulate_batch(int32_t P, int32_t n_cams, const int32_t* radii, const float* grad_means2D,
                                 float* max_2d_radii, float* grad_accum, float* vis_count, uint8_t* visible,
                                 void* stream);

/* accumulate_mean_2d_gradients (external.py:113-124) taking the visibility mask (uint8 [P]) like the reference. */
int gsb_densify_accumulate_masked(int32_t P, const uint8_t* visible, const float* grad_means2D, float* grad_accum,
                                  float* vis_count, void* stream);

/* Split step of densify_gaussians (external.py:251-271).  sel uint8 [P], rank = gsb_mask_scan(sel), S selected rows,
 * samples [2S][3] (the torch.normal draw of external.py:261).  Row rank r yields output rows r and S + r:
 * out_means = mean + R(normalize(q)) * sample,  out_log_scales = log(exp(log_scale) / 1.6). */
int gsb_split_rows(int32_t P, int32_t S, const uint8_t* sel, const int32_t* rank, const float* means,
                   const float* quats, const float* log_scales, const float* samples, float* out_means,
                   float* out_log_scales, void* stream);

/* Stable stream compaction of row tables: replaces the boolean-mask indexing of
 * remove_points / cat_params_to_optimizer (external.py:127-204).
 *   gsb_mask_scan: keep[P] (uint8) -> dst_index[P] (exclusive rank of kept rows), count_out (device int32[1]).
 *   gsb_gather_rows: dst[dst_index[i]] = src[i] for kept rows, `row_floats` floats per row. */
size_t gsb_scan_scratch_bytes(int32_t n);
int gsb_mask_scan(int32_t P, const uint8_t* keep, int32_t* dst_index, int32_t* count_out, void* scratch, void* stream);
int gsb_gather_rows(int32_t P, const uint8_t* keep, const int32_t* dst_index, const float* src, float* dst,
                    int32_t row_floats, void* stream);

/* One densify event -- clone, split (n = 2), removal of the split parents, prune (external.py:211-314 with
 * cat_params_to_optimizer / remove_points, external.py:127-204) -- as ONE pass over the Gaussians instead of four
 * re-allocations of every table.  gsb_densify_plan decides, per source row, which of its up to four output rows
 * survive (original / clone / two split children) and ranks them (order = the reference's: originals, clones,
 * first children, second children, each in source order); totals_dev (DEVICE int32 [4]) receives the numbers of
 * surviving originals, clones, child pairs and of ALL split rows (the split draw of external.py:261 has 2 x that
 * many rows).  Thresholds: mean accumulated gradient >= grad_threshold; clone when max scale <= clone_max_scale else
 * split; prune sigmoid(opacity) < prune_opacity or (prune_big and max scale > big_scale).  The caller reads the four
 * totals (its one synchronisation), allocates the new tables (K0 + K1 + 2 K2 rows) and calls gsb_densify_apply with
 * src18 / dst18 = HOST arrays of the 18 table pointers: means [.,3], colors [.,3], segmentation colours [.,3],
 * quaternions [.,4], opacity logits [.,1], log scales [.,3], then their exp_avg and their exp_avg_sq in the same
 * order (new rows get zero moments), and samples [2 S][3].  scratch: gsb_densify_plan_scratch_bytes(P), shared by
 * both calls. */
size_t gsb_densify_plan_scratch_bytes(int32_t P);
int gsb_densify_plan(int32_t P, const float* grad_accum, const float* vis_count, const float* opacity_logits,
                     const float* log_scales, float grad_threshold, float clone_max_scale, float prune_opacity,
                     int32_t prune_big, float big_scale, void* scratch, int32_t* totals_dev, void* stream);
int gsb_densify_apply(int32_t P, const void* scratch, const int32_t* totals_host, const float* const* src18,
                      float* const* dst18, const float* samples, void* stream);

/* create_render_arguments (shared.py:29-42), the three activations in one launch:
 * rotations = normalize(quats) [P][4] (torch.nn.functional.normalize: q / max(||q||, 1e-12)),
 * opacities = sigmoid(opacity_logits) [P], scales = exp(log_scales) [P][3].  New entry point:
 * the reference runs three torch kernels + autograd nodes per render. */
int gsb_activate(int32_t P, const float* quats, const float* opacity_logits, const float* log_scales,
                 float* rotations, float* opacities, float* scales, void* stream);
/* Its adjoint: gradients w.r.t. the activated tables -> gradients w.r.t. the raw parameter tables.  Any d_* input
 * may be NULL (treated as zeros); any output may be NULL (skipped). */
int gsb_activate_backward(int32_t P, const float* quats, const float* opacities, const float* scales,
                          const float* d_rotations, const float* d_opacities, const float* d_scales,
                          float* d_quats, float* d_opacity_logits, float* d_log_scales, void* stream);

/* torch.optim.Adam as create_optimizer builds it (densify.py:68-86; no weight decay, no amsgrad) over up to 16
 * parameter tables in ONE launch.  `groups_host` is a HOST array; step is the 1-based step count of the group
 * AFTER this update (torch increments before use).  Groups with grad == NULL are skipped like torch skips
 * parameters without a gradient. */
typedef struct gsb_adam_group {
  float* param; const float* grad; float* exp_avg; float* exp_avg_sq;
  int64_t numel; int64_t step; double lr;
} gsb_adam_group;
int gsb_adam_step(int32_t n_groups, const gsb_adam_group* groups_host, double beta1, double beta2, double eps,
                  void* stream);

/* One optimiser step of the first-frame loop straight from the (exchanged) flat step record of
 * gsb_raster_backward_flat: the adjoint of create_render_arguments (shared.py:29-42: normalize / sigmoid / exp,
 * evaluated on the raw parameters) + torch.optim.Adam's update (densify.py:68-86) over the six per-Gaussian tables
 * in ONE launch.  Table order: means [P][3] (cols 0-2), colors [P][3] (3-5), segmentation colours [P][3] (17-19),
 * rotation_quaternions [P][4] (6-9 through the normalisation), opacity_logits [P] (10 through the sigmoid),
 * log_scales [P][3] (11-13 through the exponential).  *step_dev (DEVICE int64) is the number of updates applied so
 * far; the call increments it on the device and uses the new value for the bias corrections, so it can be replayed
 * from a CUDA graph. */
typedef struct gsb_flat_adam_args {
  int32_t P;
  float* param[6]; float* exp_avg[6]; float* exp_avg_sq[6];
  double lr[6];
} gsb_flat_adam_args;
int gsb_adam_flat(const gsb_flat_adam_args* args, const float* flat, int32_t flat_stride, int64_t* step_dev, double beta1,
                  double beta2, double eps, void* stream);

/* Rigidity loss of the per-timestep step (train.py:317-351 with quat_mult :301-308, weighted_l2_loss_v2 :311-315,
 * build_rotation external.py:27-47): loss = mean_ij sqrt(w_ij |R_i^T (mean_nbr(i,j) - mean_i) - prev_offset_ij|^2
 * + 1e-20), R_i from normalize(quats_i) (x) prev_inv_rot_i, over F foreground Gaussians x K neighbours.
 * fg_index[F] maps a foreground Gaussian to its row of the [P,*] tables (NULL = identity); nbr holds neighbour
 * ids in foreground space.  The per-pair tables are addressed as base[i * stride_i + j * stride_j (+ c *
 * stride_c)] (strides in elements), so reference-layout [F,K] / [F,K,3] tensors and transposed copies both
 * work.  *loss (a DEVICE double) receives the loss.  With d_means / d_quats non-NULL the gradients
 * dloss/dmeans [P,3] and dloss/dquats [P,4] (raw, through the normalisation) are ADDED to them in the same
 * launch (float atomics for the neighbour shares); pass both or neither. */
int gsb_rigidity_loss(int32_t F, int32_t K, const int32_t* fg_index, const float* means, const float* quats,
                      const float* prev_inv_rot, const int32_t* nbr, int64_t nbr_stride_i, int64_t nbr_stride_j,
                      const float* weights, int64_t w_stride_i, int64_t w_stride_j, const float* prev_offsets,
                      int64_t prev_stride_i, int64_t prev_stride_j, int64_t prev_stride_c, double* loss,
                      float* d_means, float* d_quats, void* stream);

/* create_foreground_info (train.py:226-247): inv_rot[F,4] = conjugate of normalize(quats[fg]); offsets
 * (strided like prev_offsets above) = means[fg][nbr] - means[fg]. */
int gsb_foreground_info(int32_t F, int32_t K, const int32_t* fg_index, const float* means, const float* quats,
                        const int32_t* nbr, int64_t nbr_stride_i, int64_t nbr_stride_j, float* inv_rot,
                        float* offsets, int64_t off_stride_i, int64_t off_stride_j, int64_t off_stride_c,
                        void* stream);

/* Dense layer of the DeformationNetwork (train.py:57-110: nn.Linear over all Gaussians) on the tcgen05 tensor
 * cores: Y[rows, N] = X[rows, K] . W[N, K]^T (+ bias[N] when non-NULL), all FP32 in memory, row pitches ldx / ldw /
 * ldy in floats.  TF32 multiply, FP32 accumulate (what torch computes with allow_tf32).  Constraints: N <= 256;
 * X and W 16-byte aligned with ldx, ldw multiples of 4 (TMA); rows < 2^31.
 * col_stats (optional, DEVICE double[2][N], zeroed by the caller): the epilogue adds the column sums and sums of
 * squares of Y over the rows -- the BatchNorm1d batch statistics of the layer that follows (train.py:57-78) --
 * so no extra pass over Y is needed; requires a 16-byte aligned Y with ldy and N multiples of 4. */
int gsb_linear_tf32(int64_t rows, int32_t K, int32_t N, const float* X, int64_t ldx, const float* W, int64_t ldw,
                    const float* bias, float* Y, int64_t ldy, double* col_stats, void* stream);
/* Same layer with the 3xTF32 operand split (hi/lo halves of X in shared memory, of W in `scratch`): FP32-level
 * accuracy (~2^-22 relative per product) at three tensor-core instructions per tile.  scratch: device buffer of
 * gsb_linear_tf32x3_scratch_bytes(K, N) bytes, 16-byte aligned; W must be dense (ldw == K). */
size_t gsb_linear_tf32x3_scratch_bytes(int32_t K, int32_t N);
int gsb_linear_tf32x3(int64_t rows, int32_t K, int32_t N, const float* X, int64_t ldx, const float* W, int64_t ldw,
                      const float* bias, float* Y, int64_t ldy, double* col_stats, void* scratch, void* stream);

/* Y = X . W^T + residual[rows, N] (no bias): the dense layer with a second tensor added in the epilogue -- the input
 * gradient of a residual block, dX = dZ1 . W1 + (gradient of the skip branch), without a separate pass over dX
 * (train.py:57-78 under autograd).  precise != 0: the 3xTF32 form (scratch as for gsb_linear_tf32x3, ldw == K), else TF32
 * (scratch ignored).  Y and residual 16-byte aligned, ldy, ldr and N multiples of 4; Y may alias residual. */
int gsb_linear_residual(int64_t rows, int32_t K, int32_t N, const float* X, int64_t ldx, const float* W, int64_t ldw,
                        const float* residual, int64_t ldr, float* Y, int64_t ldy, int32_t precise, void* scratch,
                        void* stream);

/* Weight gradient of the dense layer: dW[N, K] (dense, row pitch K) = G[rows, N]^T . X[rows, K] (+= when accumulate),
 * TF32 multiply / FP32 accumulate on tcgen05 with both operands MN-major (the reduction runs over the rows).
 * N, K multiples of 32 and <= 256; G, X 16-byte aligned with pitches multiples of 4; scratch: device buffer of
 * gsb_linear_wgrad_scratch_bytes(N, K) bytes (per-CTA partial sums).  The input gradient dX = G . W is
 * gsb_linear_tf32 with the transposed weight. */
size_t gsb_linear_wgrad_scratch_bytes(int32_t N, int32_t K);
int gsb_linear_wgrad_tf32(int64_t rows, int32_t N, int32_t K, const float* G, int64_t ldg, const float* X, int64_t ldx,
                          float* dW, int32_t accumulate, void* scratch, void* stream);

/* The same product with the 3xTF32 operand split: both streamed operands are split where they land in shared memory
 * (hi = upper 19 bits, lo = x - hi) and Ghi Xhi + Ghi Xlo + Glo Xhi is accumulated -- FP32-level accuracy, what the
 * reference's FP32 autograd matmuls compute (train.py:57-110).  Same constraints and scratch. */
int gsb_linear_wgrad_tf32x3(int64_t rows, int32_t N, int32_t K, const float* G, int64_t ldg, const float* X, int64_t ldx,
                            float* dW, int32_t accumulate, void* scratch, void* stream);

/* The two backward products of a THIN dense layer -- the network's 7-column output layer (train.py:80-110); autograd
 * runs two cuBLAS SGEMMs in the reference -- as FP32 streaming kernels (7 is not a tensor-core tile), M <= 8:
 *   gsb_linear_thin_dx:    dX[rows, H] = G[rows, M] . W[M, H]   (H % 4 == 0, H <= 1024, dX 16-byte aligned, ldx % 4 == 0)
 *   gsb_linear_thin_wgrad: dW[M, H] (dense) = G^T . X[rows, H] and db[M] = column sums of G (optional);
 *                          8 <= H <= 1024, H % 4 == 0, X 16-byte aligned, ldx % 4 == 0;
 *                          scratch: gsb_linear_thin_wgrad_scratch_bytes(M, H) bytes (per-block partial sums). */
int gsb_linear_thin_dx(int64_t rows, int32_t M, int32_t H, const float* G, int64_t ldg, const float* W, int64_t ldw,
                       float* dX, int64_t ldx, void* stream);
size_t gsb_linear_thin_wgrad_scratch_bytes(int32_t M, int32_t H);
int gsb_linear_thin_wgrad(int64_t rows, int32_t M, int32_t H, const float* G, int64_t ldg, const float* X, int64_t ldx,
                          float* dW, float* db, void* scratch, void* stream);

/* BatchNorm1d in training mode (batch statistics; the reference never switches the network to eval) + optional
 * residual + exact (erf) GELU in one pass over Z [rows, N]: Y = act(gamma (Z - mean) / sqrt(var + eps) + beta
 * [+ residual]), mean / biased var from col_stats = the [2][N] sums gsb_linear_tf32 accumulated (train.py:57-78).
 * running_mean / running_var (both or neither) receive torch's momentum update with the unbiased variance.
 * N <= 256, N and all pitches multiples of 4.  batch_rows: the number of rows col_stats was summed over -- 0 or rows
 * for a whole batch; when the batch is sharded by rows over several GPUs (gsb200.deform row sharding) every rank passes
 * its own rows, the column sums added over the ranks, and the batch's total row count here. */
int gsb_bn_gelu(int64_t rows, int64_t batch_rows, int32_t N, const float* Z, int64_t ldz, const double* col_stats,
                const float* gamma, const float* beta, float eps, const float* residual, int64_t ldr, float* Y, int64_t ldy,
                float* running_mean, float* running_var, float momentum, int32_t apply_gelu, void* stream);

/* Adjoint of gsb_bn_gelu (same arguments for the forward quantities): given dY it forms G = dY * act'(pre-activation)
 * (also the gradient of the residual input), the per-column sums (`sums`, DEVICE double[2][N], zeroed here) that are
 * d_beta = sum G and d_gamma = sum G xhat, and dZ = gamma inv_std (G - d_beta / rows - xhat d_gamma / rows).
 * dZ may alias G (in place) or be a separate buffer (G is then kept).  d_gamma / d_beta [N] optional.
 * batch_rows as in gsb_bn_gelu.  phase 0: the whole adjoint; phase 1: only the first pass (G and `sums` over this
 * rank's rows); phase 2: only the second pass with `sums` as given -- a row-sharded batch adds the ranks' sums between
 * the two. */
int gsb_bn_gelu_backward(int64_t rows, int64_t batch_rows, int32_t phase, int32_t N, const float* dY, int64_t ldd,
                         const float* Z, int64_t ldz, const double* col_stats, const float* gamma, const float* beta,
                         float eps, const float* residual, int64_t ldr, int32_t apply_gelu, float* G, int64_t ldg,
                         float* dZ, int64_t lddz, double* sums, float* d_gamma, float* d_beta, void* stream);

/* PositionalEncoding (train.py:113-127) of x [rows, D] with F frequencies 2^k pi into out[:, col0 + j D + d],
 * j in [0, 2F): even j -> sin(f x), odd j -> cos(sin(f x)) (the reference's in-place quirk).  With lo / span [D]
 * the input is first mapped to 2 (x - lo) / span - 1 (normalize_and_encode_means_and_rotations, train.py:186-205). */
int gsb_positional_encoding(int64_t rows, int32_t D, int32_t F, const float* x, int64_t ldx, const float* lo,
                            const float* span, float* out, int64_t ldo, int32_t col0, void* stream);

/* Debug / test accessors: the individual stages, so that tests can compare every intermediate
 * bit-for-bit. */
int gsb_inclusive_scan_u32(int32_t n, const uint32_t* in, uint32_t* out, void* scratch, void* stream);
/* Allocation-free variant of gsb_inclusive_scan_u32.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same kernel and same result, but the scan's scratch (ticket + tile status words) is a
 * file-scope `static __device__` array inside the library -- gsb_scan_scratch_bytes(GSB_SCAN_STATIC_MAX_ITEMS) =
 * 65 792 bytes, held by every GPU that loads the library until exit -- whose device address is taken once per device
 * with cudaGetSymbolAddress.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine);
 * the caller still passes `in` and `out`.  n > GSB_SCAN_STATIC_MAX_ITEMS returns GSB_EINVAL (gsb_last_error() says
 * why).  Every call on a device shares the one array, which keeps the previous call's contents: the call clears the
 * words it uses first (a cudaMemsetAsync on `stream`), and two calls that OVERLAP on the same GPU (different streams,
 * or graphs replayed concurrently) race on it -- serialise them. */
#define GSB_SCAN_STATIC_MAX_ITEMS (1 << 24)
int gsb_inclusive_scan_u32_static(int32_t n, const uint32_t* in, uint32_t* out, void* stream);
size_t gsb_sort_temp_bytes(int64_t n);
int gsb_sort_pairs(int64_t n, const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in,
                   uint32_t* vals_out, int32_t end_bit, void* temp, void* stream);

/* The segmented wide-digit onesweep of the product path (csrc/segsort.cu), exposed for the stability /
 * edge-size tests: n elements in n_seg segments, seg_start DEVICE uint32 [n_seg + 1]; each segment is
 * sorted stably and independently on key bits [0, end_bit) with digits of `digit_bits` bits (u16 keys:
 * 5, 8 or 10; u32 keys: 8 or 11).  vals_in NULL: the payload is the element's index in its segment.
 * keys_out may be NULL when only one pass is needed (end_bit <= digit_bits). */
size_t gsb_seg_sort_temp_bytes(int64_t n, int32_t n_seg, int32_t key_bytes, int32_t end_bit, int32_t digit_bits);
int gsb_seg_sort_u16(int64_t n, int32_t n_seg, const uint32_t* seg_start, const uint16_t* keys_in, uint16_t* keys_out,
                     const uint32_t* vals_in, uint32_t* vals_out, int32_t end_bit, int32_t digit_bits, void* temp,
                     void* stream);
int gsb_seg_sort_u32(int64_t n, int32_t n_seg, const uint32_t* seg_start, const uint32_t* keys_in, uint32_t* keys_out,
                     const uint32_t* vals_in, uint32_t* vals_out, int32_t end_bit, int32_t digit_bits, void* temp,
                     void* stream);

/* Per-stage CUDA-event timing of the last forward/backward on this thread's library instance (bench
 * instrumentation; off by default).  Stage order: fwd_begin, preprocess, depth_sort, scan, d2h_sync, duplicate_keys,
 * sort, tile_ranges, render_fwd, bwd_begin, bwd_zero, render_bwd, preprocess_bwd (GSB_PROFILE_STAGES). */
#define GSB_PROFILE_STAGES 13
int gsb_profile_enable(int on);
int gsb_profile_read(float* ms_out_host, int32_t n);

#ifdef __cplusplus
}
#endif
#endif /* GSB200_H */
