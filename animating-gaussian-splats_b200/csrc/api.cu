// api.cu -- the extern "C" surface of libgsb200.so (see include/gsb200.h) and the host orchestration
// that replaces CudaRasterizer::Rasterizer::forward/backward (cr/rasterizer_impl.cu:198-436).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <nvtx3/nvToolsExt.h>
#include "common.cuh"

namespace gsb {

static thread_local char g_err[512] = "";
void set_last_error(const char* what, cudaError_t e) {
  snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
}

// Small per-thread pinned staging block: per-camera scalars go up, num_rendered + error flag come down.
struct HostStage {
  void* pinned = nullptr;      // [0..63]: downloads; [64..]: CamParams upload area
  size_t bytes = 0;
  cudaEvent_t ev = nullptr;    // guards re-use of the upload area by a later call
  bool ev_pending = false;
};
static thread_local HostStage g_stage;

static int stage_reserve(size_t bytes) {
  if (g_stage.bytes >= bytes) return GSB_OK;
  if (g_stage.ev_pending) { GSB_CUDA_CHECK(cudaEventSynchronize(g_stage.ev)); g_stage.ev_pending = false; }
  if (g_stage.pinned) GSB_CUDA_CHECK(cudaFreeHost(g_stage.pinned));
  g_stage.pinned = nullptr;
  g_stage.bytes = 0;
  GSB_CUDA_CHECK(cudaMallocHost(&g_stage.pinned, bytes));
  g_stage.bytes = bytes;
  if (!g_stage.ev) GSB_CUDA_CHECK(cudaEventCreateWithFlags(&g_stage.ev, cudaEventDisableTiming));
  return GSB_OK;
}

// Optional per-stage CUDA-event timing (gsb_profile_enable / gsb_profile_read), used by bench.py to time
// each kernel live on the launching stream.
enum { PF_FWD_BEGIN = 0, PF_PREPROCESS, PF_DEPTH_SORT, PF_SCAN, PF_D2H, PF_DUPKEYS, PF_SORT, PF_RANGES, PF_RENDER_FWD,
       PF_BWD_BEGIN, PF_BWD_ZERO, PF_RENDER_BWD, PF_PREPROCESS_BWD, PF_COUNT };
static thread_local bool g_prof_on = false;   // per thread, like g_stage: two threads profiling two streams do not race
static thread_local cudaEvent_t g_prof_ev[PF_COUNT];
static thread_local bool g_prof_seen[PF_COUNT];
static const char* const kStageName[PF_COUNT] = {"gsb.fwd", "gsb.preprocess", "gsb.depth_sort", "gsb.scan", "gsb.d2h_sync",
                                                 "gsb.duplicate_keys", "gsb.sort", "gsb.tile_ranges", "gsb.render_fwd",
                                                 "gsb.bwd", "gsb.bwd_zero", "gsb.render_bwd", "gsb.preprocess_bwd"};
// NVTX ranges per stage: GSB_NVTX=1 brackets the host-side issue of each stage, so a
// timeline tool shows which launches belong to which stage.  Off by default (one getenv at first use).
static bool nvtx_on() {
  static const bool on = [] { const char* e = getenv("GSB_NVTX"); return e && atoi(e) != 0; }();
  return on;
}
static thread_local int g_nvtx_depth = 0;
static inline void prof_mark(int i, cudaStream_t st) {
  if (nvtx_on()) {
    if (i != PF_FWD_BEGIN && i != PF_BWD_BEGIN && g_nvtx_depth > 0) { nvtxRangePop(); g_nvtx_depth--; }
    if (i != PF_RENDER_FWD && i != PF_PREPROCESS_BWD && i + 1 < PF_COUNT) { nvtxRangePushA(kStageName[i + 1]); g_nvtx_depth++; }
  }
  if (!g_prof_on) return;
  cudaEventRecord(g_prof_ev[i], st);
  g_prof_seen[i] = true;
}

// Bump-pointer carve-up of a byte buffer, 256-B aligned (the reference aligns to 128 B,
// cr/rasterizer_impl.h:21-29).
struct Carver {
  char* p;
  size_t off = 0;
  explicit Carver(char* base) : p(base) {}
  template <typename T>
  T* take(size_t count) {
    off = align_up(off, 256);
    T* r = reinterpret_cast<T*>(p ? p + off : nullptr);
    off += count * sizeof(T);
    return r;
  }
};

// Digit widths of the two segmented sorts (overridable for experiments: GSB_DEPTH_DIGIT = 8 | 11,
// GSB_TILE_DIGIT = 5 | 8 | 10).
static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
}
static int depth_digit() {
  static int d = env_int("GSB_DEPTH_DIGIT", 8);
  return d == 11 ? 11 : 8;
}
struct TileSortPlan { int log_nb, passes; };
static TileSortPlan tile_sort_plan(int tile_bits) {
  static int forced = env_int("GSB_TILE_DIGIT", 0);
  TileSortPlan p;
  if (forced == 5 || forced == 8 || forced == 10) p.log_nb = forced;
  else p.log_nb = tile_bits <= 5 ? 5 : tile_bits <= 8 ? 8 : tile_bits <= 10 ? 10 : 8;
  p.passes = (std::max(tile_bits, 1) + p.log_nb - 1) / p.log_nb;
  return p;
}

struct GeomExtras {
  CamParams* cams = nullptr;
  int* err = nullptr;
  void* scan_scratch = nullptr;
  uint32_t* gkeys_in = nullptr;    // unsorted depth keys [n_cams][P], written by preprocess
  uint32_t* gkeys_b = nullptr;     // ping-pong key buffers of the depth sort
  uint32_t* gkeys_c = nullptr;
  uint32_t* gorder_alt = nullptr;  // ping-pong payload buffer of the depth sort
  void* gsort_temp = nullptr;
  uint32_t* dseg_start = nullptr;  // [n_cams + 1] fixed-length segments of the depth sort
  uint32_t* dseg_tile0 = nullptr;
  uint32_t* seg_tile0 = nullptr;   // [n_cams + 1] sort-tile prefix of the instance sort
};

struct GeomLayout {
  static size_t carve(char* base, int64_t P, int n_cams, gsb_raster_ctx* c, GeomExtras* x) {
    Carver k(base);
    const size_t n = (size_t)P * n_cams;
    float* depths = k.take<float>(n);
    int32_t* radii = k.take<int32_t>(n);
    float* means2D = k.take<float>(2 * n);
    float* cov3D = k.take<float>(6 * (size_t)P);   // shared by the cameras of the batch
    float* conic = k.take<float>(4 * n);
    uint32_t* tt = k.take<uint32_t>(n);
    uint32_t* go = k.take<uint32_t>(n);
    uint32_t* po = k.take<uint32_t>(n);
    uint32_t* ss = k.take<uint32_t>((size_t)n_cams + 1);
    CamParams* cp = k.take<CamParams>(n_cams);
    int* e = k.take<int>(4);
    void* sc = k.take<char>(scan_scratch_bytes((int64_t)n));
    uint32_t* gki = k.take<uint32_t>(n);
    uint32_t* gkb = k.take<uint32_t>(n);
    uint32_t* gkc = k.take<uint32_t>(n);
    uint32_t* goa = k.take<uint32_t>(n);
    uint32_t* ds = k.take<uint32_t>((size_t)n_cams + 1);
    uint32_t* dt = k.take<uint32_t>(2 * ((size_t)n_cams + 1));   // tile prefix | chunk prefix
    uint32_t* st0 = k.take<uint32_t>(2 * ((size_t)n_cams + 1));
    const int dd = depth_digit();
    void* gt = k.take<char>(seg_sort_temp_bytes((int64_t)n, n_cams, 4, dd));
    if (c) {
      c->depths = depths; c->radii = radii; c->means2D = means2D; c->cov3D = cov3D; c->conic_opacity = conic;
      c->tiles_touched = tt; c->gauss_order = go; c->point_offsets = po; c->seg_start = ss;
    }
    if (x) {
      x->cams = cp; x->err = e; x->scan_scratch = sc; x->gkeys_in = gki; x->gkeys_b = gkb; x->gkeys_c = gkc;
      x->gorder_alt = goa; x->gsort_temp = gt; x->dseg_start = ds; x->dseg_tile0 = dt; x->seg_tile0 = st0;
    }
    return align_up(k.off, 256);
  }
};

static size_t carve_image(char* base, int W, int H, int n_cams, int tiles_total, gsb_raster_ctx* c) {
  Carver k(base);
  const size_t n = (size_t)W * H * n_cams;
  float* ft = k.take<float>(n);
  uint32_t* nc = k.take<uint32_t>(n);
  uint32_t* rg = k.take<uint32_t>(3 * (size_t)tiles_total);   // ranges (uint2 per tile) + heaviest-first tile order
  if (c) { c->final_T = ft; c->n_contrib = nc; c->ranges = rg; }
  return align_up(k.off, 256);
}

struct BinExtras {
  uint32_t* vals_alt = nullptr;
  uint16_t* keys_alt = nullptr;
};

static size_t carve_binning(char* base, int64_t R, int tile_bits, int n_cams, gsb_raster_ctx* c, BinExtras* x) {
  Carver k(base);
  const size_t n = (size_t)(R > 0 ? R : 1);
  const TileSortPlan plan = tile_sort_plan(tile_bits);
  uint32_t* pl = k.take<uint32_t>(n);
  uint32_t* plu = k.take<uint32_t>(n);
  uint16_t* tku = k.take<uint16_t>(n);
  uint32_t* va = nullptr;
  uint16_t* tks = nullptr;
  uint16_t* tka = nullptr;
  if (plan.passes > 1) {
    va = k.take<uint32_t>(n);
    tks = k.take<uint16_t>(n);
    tka = k.take<uint16_t>(n);
  }
  void* tmp = k.take<char>(seg_sort_temp_bytes(R, n_cams, 2, plan.log_nb));
  if (c) {
    c->point_list = pl; c->point_list_unsorted = plu; c->tile_keys_unsorted = tku; c->tile_keys_sorted = tks;
    c->sort_temp = tmp;
  }
  if (x) { x->vals_alt = va; x->keys_alt = tka; }
  return align_up(k.off, 256);
}

static uint32_t higher_msb(uint32_t n) {  // getHigherMsb, cr/rasterizer_impl.cu:35-50
  uint32_t msb = sizeof(n) * 4, step = msb;
  while (step > 1) {
    step /= 2;
    if (n >> msb) msb += step; else msb -= step;
  }
  if (n >> msb) msb++;
  return msb;
}

}  // namespace gsb

using namespace gsb;

extern "C" {

const char* gsb_version(void) { return "gsb200 0.1 (sm_100a)"; }
const char* gsb_last_error(void) { return g_err; }

size_t gsb_geom_bytes(int32_t P, int32_t n_cams) {
  return GeomLayout::carve(nullptr, P, n_cams, nullptr, nullptr);
}
size_t gsb_image_bytes(int32_t width, int32_t height, int32_t n_cams) {
  const int tiles = ((width + GSB_TILE - 1) / GSB_TILE) * ((height + GSB_TILE - 1) / GSB_TILE) * n_cams;
  return carve_image(nullptr, width, height, n_cams, tiles, nullptr);
}
size_t gsb_binning_bytes(int64_t num_rendered, int32_t tiles_per_cam, int32_t n_cams) {
  return carve_binning(nullptr, num_rendered, (int)higher_msb((uint32_t)std::max(tiles_per_cam, 1)), n_cams, nullptr,
                       nullptr);
}
uint32_t gsb_higher_msb(uint32_t n) { return higher_msb(n); }

}  // extern "C"

namespace gsb {
// Rasterizer::forward host orchestration (cr/rasterizer_impl.cu:198-339), shared by the blocking entry point
// (callbacks, one D2H of num_rendered that sizes the binning buffer) and the asynchronous one (caller-owned arenas,
// capacity-guarded instance count that never leaves the device).
struct FwdBuffers {
  bool async = false;
  char* geom = nullptr; char* img = nullptr; char* bin = nullptr;
  int64_t capacity = 0;
  const float* tanfov_dev = nullptr;
  gsb_alloc_fn geom_alloc = nullptr, binning_alloc = nullptr, image_alloc = nullptr;
  void* user = nullptr;
};

static int forward_impl(const gsb_raster_args* a, FwdBuffers& fb, float* out_color, float* out_depth, int32_t* radii,
                        gsb_raster_ctx* ctx, cudaStream_t st) {
  if (a->P < 0 || a->n_cams <= 0 || a->width <= 0 || a->height <= 0) return GSB_EINVAL;
  memset(ctx, 0, sizeof(*ctx));
  ctx->P = a->P; ctx->n_cams = a->n_cams; ctx->width = a->width; ctx->height = a->height;
  ctx->tiles_x = (a->width + GSB_TILE - 1) / GSB_TILE;
  ctx->tiles_y = (a->height + GSB_TILE - 1) / GSB_TILE;
  const int tiles_per_cam = ctx->tiles_x * ctx->tiles_y;
  const int tiles_total = tiles_per_cam * a->n_cams;
  // 16-bit camera-local tile ids; the backward keeps one 128-byte matrix pair per camera in shared memory
  if (tiles_per_cam > 65535 || a->n_cams > GSB_MAX_CAMS) return GSB_EINVAL;
  if (tiles_total >= (1 << 24)) return GSB_EINVAL;                    // blend kernels: float-assisted tile coordinates
  ctx->tile_bits = (int)higher_msb((uint32_t)tiles_per_cam);
  ctx->sort_bits = 32 + (int)higher_msb((uint32_t)tiles_total);
  if (a->P == 0) return GSB_OK;  // the glue returns zero images, like DGR/rasterize_points.cu:81
  if (!a->means3D || !a->colors || !a->opacities || !a->viewmatrix || !a->projmatrix || !a->background ||
      !out_color || !out_depth)
    return GSB_EINVAL;
  if (!a->cov3D_precomp && (!a->scales || !a->rotations)) return GSB_EINVAL;

  // geometry + image state
  char* geom = fb.async ? fb.geom : fb.geom_alloc(fb.user, gsb_geom_bytes(a->P, a->n_cams));
  char* img = fb.async ? fb.img : fb.image_alloc(fb.user, gsb_image_bytes(a->width, a->height, a->n_cams));
  if (!geom || !img) return GSB_ENOMEM;
  GeomExtras gx;
  GeomLayout::carve(geom, a->P, a->n_cams, ctx, &gx);
  CamParams* cams_dev = gx.cams; int* err_dev = gx.err; void* scan_scratch = gx.scan_scratch;
  ctx->status = err_dev;
  carve_image(img, a->width, a->height, a->n_cams, tiles_total, ctx);

  int rc;
  if (fb.async) {
    // per-camera scalars computed on the device from the caller's table: nothing host-resident enters the stream
    if ((rc = launch_cam_params(a->n_cams, a->width, a->height, fb.tanfov_dev, a->entries, cams_dev, st))) return rc;
  } else {
    // per-camera scalars: focal = dim / (2 tanfov) (cr/rasterizer_impl.cu:222-223)
    if ((rc = stage_reserve(64 + sizeof(CamParams) * (size_t)a->n_cams))) return rc;
    if (g_stage.ev_pending) { GSB_CUDA_CHECK(cudaEventSynchronize(g_stage.ev)); g_stage.ev_pending = false; }
    CamParams* cams_host = reinterpret_cast<CamParams*>(reinterpret_cast<char*>(g_stage.pinned) + 64);
    for (int i = 0; i < a->n_cams; i++) {
      cams_host[i].tanfovx = a->tanfovx_host[i];
      cams_host[i].tanfovy = a->tanfovy_host[i];
      cams_host[i].focal_y = (float)a->height / (2.0f * a->tanfovy_host[i]);
      cams_host[i].focal_x = (float)a->width / (2.0f * a->tanfovx_host[i]);
      cams_host[i].row0 = 0; cams_host[i].row1 = ctx->tiles_y; cams_host[i].split_slot = -1; cams_host[i].count_vis = 1;
    }
    GSB_CUDA_CHECK(cudaMemcpyAsync(cams_dev, cams_host, sizeof(CamParams) * (size_t)a->n_cams, cudaMemcpyHostToDevice, st));
    if (a->entries)   // tile-band sharding: the entry fields come from the caller's device table
      if ((rc = launch_cam_params(a->n_cams, a->width, a->height, nullptr, a->entries, cams_dev, st))) return rc;
  }
  GSB_CUDA_CHECK(cudaMemsetAsync(err_dev, 0, 16, st));

  prof_mark(PF_FWD_BEGIN, st);
  if ((rc = launch_preprocess(*a, cams_dev, *ctx, radii, gx.gkeys_in, err_dev, st))) return rc;
  prof_mark(PF_PREPROCESS, st);
  const int64_t n_slots = (int64_t)a->P * a->n_cams;
  // the depth digits are sorted here, on Gaussians, one segment per camera, before duplication
  if ((rc = launch_seg_setup(a->n_cams, 1, a->P, 4, nullptr, gx.dseg_start, gx.dseg_tile0, st))) return rc;
  if ((rc = launch_seg_sort_u32(n_slots, a->n_cams, gx.dseg_start, gx.dseg_tile0, gx.gkeys_in, gx.gkeys_b, gx.gkeys_c,
                                nullptr, ctx->gauss_order, gx.gorder_alt, 32, depth_digit(), false, gx.gsort_temp, st,
                                nullptr)))
    return rc;
  prof_mark(PF_DEPTH_SORT, st);
  if ((rc = launch_inclusive_scan_u32(n_slots, ctx->tiles_touched, ctx->point_offsets, scan_scratch, st,
                                      ctx->gauss_order, a->P)))
    return rc;
  // instance segments per camera; writes num_rendered to status[1] and applies the capacity guard
  if ((rc = launch_seg_setup(a->n_cams, 2, a->P, 2, ctx->point_offsets, ctx->seg_start, gx.seg_tile0, st,
                             fb.async ? fb.capacity : (int64_t)-1, err_dev)))
    return rc;
  prof_mark(PF_SCAN, st);

  char* bin;
  if (fb.async) {
    ctx->num_rendered = -1;   // known on the device only (status[1])
    ctx->capacity = fb.capacity;
    bin = fb.bin;
  } else {
    // num_rendered: one blocking 8-byte D2H (cr/rasterizer_impl.cu:281-282)
    uint32_t* down = reinterpret_cast<uint32_t*>(g_stage.pinned);
    GSB_CUDA_CHECK(cudaMemcpyAsync(down, ctx->point_offsets + n_slots - 1, 4, cudaMemcpyDeviceToHost, st));
    GSB_CUDA_CHECK(cudaMemcpyAsync(down + 1, err_dev, 4, cudaMemcpyDeviceToHost, st));
    GSB_CUDA_CHECK(cudaStreamSynchronize(st));
    if ((int)down[1] != 0) return (int)down[1];
    ctx->num_rendered = (int64_t)down[0];
    if (ctx->num_rendered >= (1ll << 30)) return GSB_EINVAL;
    ctx->capacity = ctx->num_rendered;
    bin = fb.binning_alloc(fb.user, gsb_binning_bytes(ctx->num_rendered, tiles_per_cam, a->n_cams));
  }
  if (!bin) return GSB_ENOMEM;
  BinExtras bx;
  carve_binning(bin, ctx->capacity, ctx->tile_bits, a->n_cams, ctx, &bx);

  prof_mark(PF_D2H, st);
  if ((rc = launch_duplicate_with_keys(*ctx, cams_dev, st))) return rc;
  prof_mark(PF_DUPKEYS, st);
  const TileSortPlan plan = tile_sort_plan(ctx->tile_bits);
  const uint32_t* tile_hist = nullptr;
  if (ctx->capacity > 0) {
    if ((rc = launch_seg_sort_u16(ctx->capacity, a->n_cams, ctx->seg_start, gx.seg_tile0, ctx->tile_keys_unsorted,
                                  ctx->tile_keys_sorted, bx.keys_alt, ctx->point_list_unsorted, ctx->point_list,
                                  bx.vals_alt, std::max(ctx->tile_bits, 1), plan.log_nb, plan.passes > 1, ctx->sort_temp,
                                  st, &tile_hist, (int)higher_msb((uint32_t)a->P))))   // payloads are Gaussian ids < P
      return rc;
  }
  prof_mark(PF_SORT, st);
  if (plan.passes == 1 && ctx->capacity > 0) {
    if ((rc = launch_ranges_from_hist(a->n_cams, tiles_per_cam, plan.log_nb, ctx->seg_start, tile_hist, ctx->ranges, st)))
      return rc;
  } else {
    if ((rc = launch_tile_ranges_seg(ctx->capacity, a->n_cams, tiles_per_cam, ctx->seg_start, ctx->tile_keys_sorted,
                                     ctx->ranges, st)))
      return rc;
  }
  if ((rc = launch_tile_order(tiles_total, ctx->ranges, ctx->ranges + 2 * (size_t)tiles_total, st))) return rc;
  prof_mark(PF_RANGES, st);
  if ((rc = launch_render_fwd(*a, *ctx, out_color, out_depth, st))) return rc;
  prof_mark(PF_RENDER_FWD, st);
  return GSB_OK;
}
}  // namespace gsb

extern "C" {

int gsb_raster_forward(const gsb_raster_args* a, gsb_alloc_fn geom_alloc, gsb_alloc_fn binning_alloc,
                       gsb_alloc_fn image_alloc, void* user, float* out_color, float* out_depth, int32_t* radii,
                       gsb_raster_ctx* ctx, void* stream) {
  if (!a || !ctx || !geom_alloc || !binning_alloc || !image_alloc) return GSB_EINVAL;
  if (!a->tanfovx_host || !a->tanfovy_host) return GSB_EINVAL;
  FwdBuffers fb;
  fb.geom_alloc = geom_alloc; fb.binning_alloc = binning_alloc; fb.image_alloc = image_alloc; fb.user = user;
  return forward_impl(a, fb, out_color, out_depth, radii, ctx, reinterpret_cast<cudaStream_t>(stream));
}

int gsb_raster_forward_async(const gsb_raster_args* a, const float* tanfov_dev, char* geom, size_t geom_bytes,
                             char* image, size_t image_bytes, char* binning, size_t binning_bytes, int64_t capacity,
                             float* out_color, float* out_depth, int32_t* radii, gsb_raster_ctx* ctx, void* stream) {
  if (!a || !ctx || !tanfov_dev || !geom || !image || !binning) return GSB_EINVAL;
  if (capacity < 0 || capacity >= (1ll << 30) || a->n_cams <= 0 || a->width <= 0 || a->height <= 0 || a->P < 0)
    return GSB_EINVAL;
  const int tiles_per_cam = ((a->width + GSB_TILE - 1) / GSB_TILE) * ((a->height + GSB_TILE - 1) / GSB_TILE);
  if (geom_bytes < gsb_geom_bytes(a->P, a->n_cams) || image_bytes < gsb_image_bytes(a->width, a->height, a->n_cams) ||
      binning_bytes < gsb_binning_bytes(capacity, tiles_per_cam, a->n_cams))
    return GSB_ENOMEM;
  if (((uintptr_t)geom | (uintptr_t)image | (uintptr_t)binning) & 255u) return GSB_EINVAL;
  FwdBuffers fb;
  fb.async = true; fb.geom = geom; fb.img = image; fb.bin = binning; fb.capacity = capacity; fb.tanfov_dev = tanfov_dev;
  return forward_impl(a, fb, out_color, out_depth, radii, ctx, reinterpret_cast<cudaStream_t>(stream));
}

int gsb_raster_count_pairs(const gsb_raster_args* a, const gsb_raster_ctx* ctx, uint64_t* counts, void* stream) {
  if (!a || !ctx || !counts) return GSB_EINVAL;
  if (a->P != ctx->P || a->n_cams != ctx->n_cams || a->width != ctx->width || a->height != ctx->height || !ctx->ranges)
    return GSB_EINVAL;
  if (!a->colors || !a->background) return GSB_EINVAL;
  return launch_count_pairs(*a, *ctx, reinterpret_cast<unsigned long long*>(counts), reinterpret_cast<cudaStream_t>(stream));
}

int gsb_raster_forward_colors(const gsb_raster_args* a, const gsb_raster_ctx* ctx, float* out_color, float* out_depth,
                              void* stream) {
  if (!a || !ctx) return GSB_EINVAL;
  if (a->P == 0) return GSB_OK;
  if (!a->colors || !a->background || !out_color || !out_depth) return GSB_EINVAL;
  if (a->P != ctx->P || a->n_cams != ctx->n_cams || a->width != ctx->width || a->height != ctx->height || !ctx->ranges)
    return GSB_EINVAL;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  prof_mark(PF_FWD_BEGIN, st);
  int rc;
  if ((rc = launch_render_fwd(*a, *ctx, out_color, out_depth, st))) return rc;
  if (nvtx_on() && g_nvtx_depth > 0) { nvtxRangePop(); g_nvtx_depth--; }
  return GSB_OK;
}

int gsb_raster_backward(const gsb_raster_args* a, const gsb_raster_ctx* ctx, const float* dL_dpix,
                        float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                        float* dL_dmean3D, float* dL_dcov3D, float* dL_dscale, float* dL_drot, void* stream) {
  if (!a || !ctx) return GSB_EINVAL;
  if (a->P == 0) return GSB_OK;
  if (!dL_dpix || !dL_dmean2D || !dL_dconic || !dL_dopacity || !dL_dcolor || !dL_dmean3D) return GSB_EINVAL;
  if (a->P != ctx->P || a->n_cams != ctx->n_cams || a->width != ctx->width || a->height != ctx->height)
    return GSB_EINVAL;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const size_t n = (size_t)a->P * a->n_cams;
  prof_mark(PF_BWD_BEGIN, st);
  GSB_CUDA_CHECK(cudaMemsetAsync(dL_dmean2D, 0, n * 3 * sizeof(float), st));
  GSB_CUDA_CHECK(cudaMemsetAsync(dL_dconic, 0, n * 4 * sizeof(float), st));
  GSB_CUDA_CHECK(cudaMemsetAsync(dL_dopacity, 0, (size_t)a->P * sizeof(float), st));
  const size_t ncol = a->color_cam_stride ? ((size_t)(a->n_cams - 1) * a->color_cam_stride + (size_t)a->P * 3)
                                          : (size_t)a->P * 3;
  GSB_CUDA_CHECK(cudaMemsetAsync(dL_dcolor, 0, ncol * sizeof(float), st));
  int rc;
  prof_mark(PF_BWD_ZERO, st);
  if (ctx->num_rendered != 0)   // -1: known on the device only (asynchronous forward)
    if ((rc = launch_render_bwd(*a, *ctx, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, st))) return rc;
  prof_mark(PF_RENDER_BWD, st);
  // the per-camera scalar table still sits in the geometry buffer next to the state arrays
  GeomExtras gx;
  GeomLayout::carve(reinterpret_cast<char*>(ctx->depths), a->P, a->n_cams, nullptr, &gx);
  CamParams* cams_dev = gx.cams;
  if ((rc = launch_preprocess_bwd(*a, cams_dev, *ctx, dL_dmean2D, dL_dconic, dL_dmean3D, dL_dcov3D, dL_dscale,
                                  dL_drot, st)))
    return rc;
  prof_mark(PF_PREPROCESS_BWD, st);
  return GSB_OK;
}

size_t gsb_backward_scratch_bytes(int32_t P, int32_t n_cams) {
  // raw second moments [n_cams][P][4] (read as float4: first, so they stay 16-byte aligned for any P) | raw first
  // moments [n_cams][P][3] | dL/dopacity [P] | dL/dcolor [P][3] | two-table backward: dL/dcolor of the second table
  // [P][3] | first moments of the first image alone [n_cams][P][2]
  return align_up(((size_t)P * n_cams * 9 + (size_t)P * 7) * sizeof(float), 256);
}

}  // extern "C"

static int backward_flat_impl(const gsb_raster_args* a, const gsb_raster_ctx* ctx, const float* dL_dpix,
                              const float* dL_dpix2, void* scratch, float* flat, int32_t flat_stride, int32_t mode,
                              float* dL_dmean2D_out, void* stream) {
  if (!a || !ctx) return GSB_EINVAL;
  if (a->P == 0) return GSB_OK;
  const int part = mode & ~0xFF;
  mode &= 0xFF;
  if (!scratch || !flat || mode < 0 || mode > 2 || a->color_cam_stride != 0) return GSB_EINVAL;
  if (flat_stride < GSB_FLAT_WIDTH || (flat_stride & 3)) return GSB_EINVAL;
  if (part != 0 && part != GSB_FLAT_BLEND_ONLY && part != GSB_FLAT_FINISH_ONLY) return GSB_EINVAL;
  if (!dL_dpix && part != GSB_FLAT_FINISH_ONLY) return GSB_EINVAL;
  if (a->P != ctx->P || a->n_cams != ctx->n_cams || a->width != ctx->width || a->height != ctx->height)
    return GSB_EINVAL;
  if (((uintptr_t)flat & 15u) || ((uintptr_t)scratch & 15u)) return GSB_EINVAL;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const size_t n = (size_t)a->P * a->n_cams;
  float* m2 = reinterpret_cast<float*>(scratch);
  float* m1 = m2 + 4 * n;
  float* dop = m1 + 3 * n;
  float* dcol = dop + a->P;
  float* dcol2 = dL_dpix2 ? dcol + 3 * (size_t)a->P : nullptr;
  float* m1f = dL_dpix2 ? dcol2 + 3 * (size_t)a->P : nullptr;
  if (dL_dpix2 && (mode != 0 || part != 0 || !a->colors2)) return GSB_EINVAL;
  int rc;
  if (part != GSB_FLAT_FINISH_ONLY) {
    prof_mark(PF_BWD_BEGIN, st);
    GSB_CUDA_CHECK(cudaMemsetAsync(scratch, 0, (n * (dL_dpix2 ? 9 : 7) + (size_t)a->P * (dL_dpix2 ? 7 : 4)) * sizeof(float), st));   // one fill
    prof_mark(PF_BWD_ZERO, st);
    if (ctx->num_rendered != 0)
      if ((rc = launch_render_bwd(*a, *ctx, dL_dpix, m1, m2, dop, dcol, st, dL_dpix2, dcol2, m1f))) return rc;
    prof_mark(PF_RENDER_BWD, st);
    if (part == GSB_FLAT_BLEND_ONLY) return GSB_OK;
  }
  GeomExtras gx;
  GeomLayout::carve(reinterpret_cast<char*>(ctx->depths), a->P, a->n_cams, nullptr, &gx);
  if ((rc = launch_preprocess_bwd(*a, gx.cams, *ctx, m1, m2, nullptr, nullptr, nullptr, nullptr, st, flat, mode, dop, dcol,
                                  dL_dmean2D_out, flat_stride, dcol2, m1f)))
    return rc;
  prof_mark(PF_PREPROCESS_BWD, st);
  return GSB_OK;
}

extern "C" {

int gsb_raster_backward_flat(const gsb_raster_args* a, const gsb_raster_ctx* ctx, const float* dL_dpix, void* scratch,
                             float* flat, int32_t flat_stride, int32_t mode, float* dL_dmean2D_out, void* stream) {
  return backward_flat_impl(a, ctx, dL_dpix, nullptr, scratch, flat, flat_stride, mode, dL_dmean2D_out, stream);
}

int gsb_raster_backward_flat2(const gsb_raster_args* a, const gsb_raster_ctx* ctx, const float* dL_dpix,
                              const float* dL_dpix2, void* scratch, float* flat, int32_t flat_stride,
                              float* dL_dmean2D_out, void* stream) {
  if (!dL_dpix2) return GSB_EINVAL;
  return backward_flat_impl(a, ctx, dL_dpix, dL_dpix2, scratch, flat, flat_stride, 0, dL_dmean2D_out, stream);
}

int gsb_flat_apply_stats(int32_t P, const float* flat, int32_t flat_stride, float* grad_accum, float* vis_count,
                         float* max_2d_radii, void* stream) {
  if (P < 0 || (P > 0 && (!flat || !grad_accum || !vis_count))) return GSB_EINVAL;
  if (flat_stride < GSB_FLAT_WIDTH || (flat_stride & 3)) return GSB_EINVAL;
  return launch_flat_apply_stats(P, flat, flat_stride, grad_accum, vis_count, max_2d_radii,
                                 reinterpret_cast<cudaStream_t>(stream));
}

size_t gsb_comm_bytes(size_t payload_bytes) { return comm_bytes(payload_bytes); }

int gsb_comm_alloc(size_t bytes, void** block, unsigned char* handle64) {
  if (!block || !handle64 || bytes == 0) return GSB_EINVAL;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  GSB_CUDA_CHECK(cudaMalloc(&p, bytes));
  GSB_CUDA_CHECK(cudaMemset(p, 0, bytes));
  cudaIpcMemHandle_t h;
  GSB_CUDA_CHECK(cudaIpcGetMemHandle(&h, p));
  memcpy(handle64, &h, 64);
  *block = p;
  return GSB_OK;
}

int gsb_comm_open(const unsigned char* handle64, void** peer_block) {
  if (!handle64 || !peer_block) return GSB_EINVAL;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  GSB_CUDA_CHECK(cudaIpcOpenMemHandle(peer_block, h, cudaIpcMemLazyEnablePeerAccess));
  return GSB_OK;
}

int gsb_comm_close(void* peer_block) {
  if (peer_block) GSB_CUDA_CHECK(cudaIpcCloseMemHandle(peer_block));
  return GSB_OK;
}

int gsb_comm_free(void* block) {
  if (block) GSB_CUDA_CHECK(cudaFree(block));
  return GSB_OK;
}

int gsb_allreduce_flat(int32_t rank, int32_t world, void* const* peers_host, size_t payload_bytes, int64_t n_float4,
                       int32_t row_float4, int32_t max_word, int32_t max_comp, void* stream) {
  if (world < 1 || world > GSB_COMM_MAX_RANKS || rank < 0 || rank >= world || !peers_host || n_float4 < 0) return GSB_EINVAL;
  if ((size_t)n_float4 * 16 > payload_bytes || row_float4 <= 0 || max_comp < -1 || max_comp > 3) return GSB_EINVAL;
  for (int i = 0; i < world; i++)
    if (!peers_host[i] || ((uintptr_t)peers_host[i] & 15u)) return GSB_EINVAL;
  return launch_allreduce_flat(rank, world, peers_host, payload_bytes, n_float4, row_float4, max_word, max_comp,
                               reinterpret_cast<cudaStream_t>(stream));
}

int gsb_profile_enable(int on) {
  if (on && !g_prof_ev[0]) {
    for (int i = 0; i < PF_COUNT; i++) GSB_CUDA_CHECK(cudaEventCreate(&g_prof_ev[i]));
  }
  for (int i = 0; i < PF_COUNT; i++) g_prof_seen[i] = false;
  g_prof_on = on != 0;
  return GSB_OK;
}

int gsb_profile_read(float* ms_out, int32_t n) {
  // ms_out[i] = elapsed between stage event i-1 and i of the LAST forward / backward (0 for the two begins)
  if (!ms_out || n < PF_COUNT) return GSB_EINVAL;
  for (int i = 0; i < PF_COUNT; i++) {
    ms_out[i] = 0.f;
    if (i == PF_FWD_BEGIN || i == PF_BWD_BEGIN || !g_prof_seen[i] || !g_prof_seen[i - 1]) continue;
    GSB_CUDA_CHECK(cudaEventSynchronize(g_prof_ev[i]));
    GSB_CUDA_CHECK(cudaEventElapsedTime(&ms_out[i], g_prof_ev[i - 1], g_prof_ev[i]));
  }
  return GSB_OK;
}

int gsb_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream) {
  (void)projmatrix;
  if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) return GSB_EINVAL;
  return launch_mark_visible(P, means3D, viewmatrix, present, reinterpret_cast<cudaStream_t>(stream));
}

size_t gsb_scan_scratch_bytes(int32_t n) { return scan_scratch_bytes(n); }
int gsb_inclusive_scan_u32(int32_t n, const uint32_t* in, uint32_t* out, void* scratch, void* stream) {
  if (n < 0 || (n > 0 && (!in || !out || !scratch))) return GSB_EINVAL;
  return launch_inclusive_scan_u32(n, in, out, scratch, reinterpret_cast<cudaStream_t>(stream));
}
size_t gsb_sort_temp_bytes(int64_t n) { return sort_temp_bytes(n); }
int gsb_sort_pairs(int64_t n, const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in,
                   uint32_t* vals_out, int32_t end_bit, void* temp, void* stream) {
  if (n < 0 || (n > 0 && (!keys_in || !keys_out || !vals_in || !vals_out || !temp))) return GSB_EINVAL;
  return launch_sort_pairs(n, keys_in, keys_out, vals_in, vals_out, end_bit, temp,
                           reinterpret_cast<cudaStream_t>(stream));
}
}  // extern "C"
static int seg_passes(int end_bit, int digit_bits) { return (end_bit + digit_bits - 1) / digit_bits; }
extern "C" size_t gsb_seg_sort_temp_bytes(int64_t n, int32_t n_seg, int32_t key_bytes, int32_t end_bit, int32_t digit_bits) {
  if (n < 0 || n_seg <= 0 || digit_bits <= 0 || end_bit <= 0) return 0;
  const size_t nn = (size_t)(n > 0 ? n : 1);
  // sort temp + seg_tile0 + alternate key / payload buffers
  return seg_sort_temp_bytes(n, n_seg, key_bytes, digit_bits) + align_up(((size_t)n_seg + 1) * 8, 256) +
         align_up(nn * (size_t)key_bytes, 256) + align_up(nn * 4, 256);
}
template <typename KeyT, typename Fn>
static int seg_sort_debug(int64_t n, int32_t n_seg, const uint32_t* seg_start, const KeyT* kin, KeyT* kout,
                          const uint32_t* vin, uint32_t* vout, int32_t end_bit, int32_t digit_bits, void* temp,
                          cudaStream_t st, Fn launch) {
  if (n < 0 || n_seg <= 0 || n_seg > 1024 || !seg_start || !temp || end_bit <= 0) return GSB_EINVAL;
  if (n == 0) return GSB_OK;
  if (!kin || !vout) return GSB_EINVAL;
  const int passes = seg_passes(end_bit, digit_bits);
  char* t = reinterpret_cast<char*>(temp);
  const size_t s0 = seg_sort_temp_bytes(n, n_seg, (int)sizeof(KeyT), digit_bits);
  uint32_t* tile0 = reinterpret_cast<uint32_t*>(t + s0);
  KeyT* k_alt = reinterpret_cast<KeyT*>(t + s0 + align_up(((size_t)n_seg + 1) * 8, 256));
  uint32_t* v_alt = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(k_alt) + align_up((size_t)n * sizeof(KeyT), 256));
  int rc = launch_seg_setup(n_seg, 0, 0, (int)sizeof(KeyT), nullptr, const_cast<uint32_t*>(seg_start), tile0, st);
  if (rc) return rc;
  return launch(n, n_seg, seg_start, tile0, kin, kout, k_alt, vin, vout, v_alt, end_bit, digit_bits, kout != nullptr, temp, st,
                nullptr);
}
extern "C" {
int gsb_seg_sort_u16(int64_t n, int32_t n_seg, const uint32_t* seg_start, const uint16_t* keys_in, uint16_t* keys_out,
                     const uint32_t* vals_in, uint32_t* vals_out, int32_t end_bit, int32_t digit_bits, void* temp,
                     void* stream) {
  return seg_sort_debug<uint16_t>(n, n_seg, seg_start, keys_in, keys_out, vals_in, vals_out, end_bit, digit_bits, temp,
                                  reinterpret_cast<cudaStream_t>(stream),
                                  [](int64_t n_, int ns, const uint32_t* ss, const uint32_t* t0, const uint16_t* ki, uint16_t* ko,
                                     uint16_t* ka, const uint32_t* vi, uint32_t* vo, uint32_t* va, int eb, int db, bool wk,
                                     void* tp, cudaStream_t s, const uint32_t** h) {
                                    return launch_seg_sort_u16(n_, ns, ss, t0, ki, ko, ka, vi, vo, va, eb, db, wk, tp, s, h, 32);
                                  });
}
int gsb_seg_sort_u32(int64_t n, int32_t n_seg, const uint32_t* seg_start, const uint32_t* keys_in, uint32_t* keys_out,
                     const uint32_t* vals_in, uint32_t* vals_out, int32_t end_bit, int32_t digit_bits, void* temp,
                     void* stream) {
  return seg_sort_debug<uint32_t>(n, n_seg, seg_start, keys_in, keys_out, vals_in, vals_out, end_bit, digit_bits, temp,
                                  reinterpret_cast<cudaStream_t>(stream), launch_seg_sort_u32);
}

}  // extern "C"
