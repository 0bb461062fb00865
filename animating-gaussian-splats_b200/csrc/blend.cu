// blend.cu -- per-tile alpha blending, forward and backward.
//
// Forward replaces renderCUDA (cr/forward.cu:261-393): a 16x16 block per tile walks the tile's depth-
// sorted instance list in batches staged into shared memory (position, conic+opacity and colour+depth
// packed as float2/float4/float4, so the inner loop never touches global memory -- the reference
// gathers the colour from global per blended pair), front to back, with per-pixel early termination.
// The power/alpha/T arithmetic is pinned with round-to-nearest intrinsics in the reference's rounding
// order so every threshold decision (power > 0, alpha < 1/255, T < 1e-4, median depth) and therefore
// n_contrib is identical.
//
// Backward replaces renderCUDA (cr/backward.cu:399-557): same lists back to front.  The nine
// per-Gaussian partial gradients are first summed over the 32 pixels of the warp with a butterfly
// multi-value shuffle reduction (14 SHFL for 9 values) and only the per-warp sums go to global
// atomics in one RED instruction (the reference issues 9 global atomics per pixel-Gaussian pair).
// Warps skip list entries behind their deepest contributor, and whole batches behind the tile's.
#include "common.cuh"

namespace gsb {

// power = -0.5 (A dx^2 + C dy^2) - B dx dy with the reference's contraction (SASS of renderCUDA):
//   fma(fma(dx, dx*A, dy*(dy*C)), -0.5, -(dy*(dx*B)))
__device__ __forceinline__ float blend_power(float dx, float dy, float A, float B, float C) {
  const float q = FMA(dx, MUL(dx, A), MUL(dy, MUL(dy, C)));
  return FMA(q, -0.5f, -MUL(dy, MUL(dx, B)));
}

// ------------------------------------------------------------------------------------------------
// Exact-conservative tile culling, done by the thread that stages a list entry.
//
// The reference's lists hold every Gaussian whose 3-sigma bounding SQUARE touches the tile; most of
// them cannot reach alpha >= 1/255 at any pixel of the tile.  An entry may be dropped from the staged
// batch iff for every pixel of the tile the reference would `continue` (cr/forward.cu:342-351), i.e.
//   opacity < 1/255                      (alpha = min(.99, o*exp(power)) <= o, power <= 0), or
//   max over the tile of power < -ln(255 o)   with a slack that covers every rounding difference.
// max power = -0.5 * min q, q = A dx^2 + 2 B dx dy + C dy^2 (convex): 0 if the centre lies in the tile,
// else attained on one of the four edges (a clamped 1-D quadratic each).  Dropping such entries leaves
// colour, depth, final_T and n_contrib (= original list position of the last blended entry) unchanged.
// `thr` (returned) is the per-entry power below which a pixel certainly has alpha < 1/255, so the inner
// loop can skip it before paying for expf.
// ------------------------------------------------------------------------------------------------
// Packs the 8-bit patch mask into the low mantissa bits of the threshold (a 3e-5 relative change of a
// value that carries a 2e-3 slack).
__device__ __forceinline__ float pack_thr(float thr, unsigned mask) {
  return __uint_as_float((__float_as_uint(thr) & ~0xFFu) | (mask & 0xFFu));
}

template <bool WITH_MASK>
__device__ __forceinline__ bool tile_cull_keep(float cx, float cy, float4 co, float tx0, float ty0, float& thr) {
  const float A = co.x, B = co.y, C = co.z, o = co.w;
  thr = pack_thr(-3.0e38f, 0xFFu);
  if (o < 1.0f / 255.0f) return false;
  if (!(o >= 1.0f / 255.0f)) return true;                 // NaN opacity: keep, no early threshold
  const float L = logf(255.0f * o);                        // >= 0
  thr = pack_thr(-(L + 2.0e-3f + 1.0e-3f * L), 0xFFu);
  const float det = A * C - B * B;
  if (!(A > 0.f && C > 0.f && det > 0.f)) return true;      // not positive definite: no claim
  const float dxlo = cx - (tx0 + 15.0f), dxhi = cx - tx0;  // d = centre - pixel over the tile
  const float dylo = cy - (ty0 + 15.0f), dyhi = cy - ty0;
  const float DX = fmaxf(fabsf(dxlo), fabsf(dxhi)), DY = fmaxf(fabsf(dylo), fabsf(dyhi));
  const float Mmax = A * DX * DX + 2.f * fabsf(B) * DX * DY + C * DY * DY;  // bounds every rounding error
  const float lim = L + 2.0e-3f + 1.0e-3f * L + 1.0e-5f * Mmax;              // a pixel contributes only if q/2 <= lim
  if (WITH_MASK) {
    // Patch mask: the axis-aligned bounding box of the ellipse q <= 2 lim has half-widths
    // sqrt(2 lim C / det), sqrt(2 lim A / det); warp w owns the 8x4 patch (column w & 1, row w >> 1).
    const float idet = __fdividef(1.0f, det);
    const float hx = sqrtf(2.f * lim * C * idet) * 1.0001f + 1.0e-3f;
    const float hy = sqrtf(2.f * lim * A * idet) * 1.0001f + 1.0e-3f;
    unsigned cols = 0, rows = 0;
#pragma unroll
    for (int j = 0; j < 2; j++)
      if (cx + hx >= tx0 + 8.f * j && cx - hx <= tx0 + 8.f * j + 7.f) cols |= 1u << j;
#pragma unroll
    for (int r = 0; r < 4; r++)
      if (cy + hy >= ty0 + 4.f * r && cy - hy <= ty0 + 4.f * r + 3.f) rows |= 1u << r;
    unsigned mask = 0;
#pragma unroll
    for (int r = 0; r < 4; r++)
      if (rows & (1u << r)) mask |= cols << (2 * r);
    if (!(hx == hx && hy == hy)) mask = 0xFFu;             // NaN: no claim
    thr = pack_thr(thr, mask);
    if (mask == 0) return false;
  }
  if (dxlo <= 0.f && dxhi >= 0.f && dylo <= 0.f && dyhi >= 0.f) return true;
  const float iA = __fdividef(1.0f, A), iC = __fdividef(1.0f, C);
  float qmin = 3.0e38f;
#pragma unroll
  for (int k = 0; k < 2; k++) {
    const float e = k ? dxhi : dxlo;                       // vertical edges: dx = e
    const float dy = fminf(fmaxf(-B * e * iC, dylo), dyhi);
    qmin = fminf(qmin, A * e * e + 2.f * B * e * dy + C * dy * dy);
    const float f = k ? dyhi : dylo;                       // horizontal edges: dy = f
    const float dx = fminf(fmaxf(-B * f * iA, dxlo), dxhi);
    qmin = fminf(qmin, A * dx * dx + 2.f * B * dx * f + C * f * f);
  }
  if (!(qmin < 3.0e38f)) return true;                      // NaN geometry: no claim
  return !(0.5f * qmin > lim);
}

// Block-wide stable compaction slot for `keep`; returns the number of kept entries of the batch.
__device__ __forceinline__ int compact_slot(bool keep, unsigned tid, int* s_wcnt, int& slot) {
  const unsigned bal = __ballot_sync(0xffffffffu, keep);
  if ((tid & 31u) == 0) s_wcnt[tid >> 5] = __popc(bal);
  __syncthreads();
  int before = 0, total = 0;
#pragma unroll
  for (int w = 0; w < GSB_BLOCK_SIZE / 32; w++) {
    const int c = s_wcnt[w];
    if (w < (int)(tid >> 5)) before += c;
    total += c;
  }
  slot = before + __popc(bal & lanemask_lt());
  return total;
}

// ------------------------------------------------------------------------------------------------
// Heaviest tiles first.  Blocks are dispatched in blockIdx order; a tile's cost is roughly its list length, and
// the lengths span three orders of magnitude, so with tiles in raster order the last wave of a launch is a few
// long tiles running alone (at C2's 24 840 blocks that tail is small, with 3-4 cameras per GPU it is a fifth of
// the kernel).  For short launches (use_tile_order) one block buckets the tiles by floor(log2(length)) (a counting sort, 33 buckets, longest
// first) into `order`; the blend kernels map blockIdx.x through it.  Order within a bucket is arbitrary, which
// is harmless: no result depends on block order.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
tile_order_kernel(int tiles_total, const uint2* __restrict__ ranges, uint32_t* __restrict__ order) {
  __shared__ unsigned s_cnt[33], s_base[33];
  if (threadIdx.x < 33) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  for (int t = threadIdx.x; t < tiles_total; t += 1024) {
    const uint2 r = ranges[t];
    const unsigned len = r.y - r.x;
    atomicAdd(&s_cnt[len ? 32 - __clz(len) : 0], 1u);   // bucket 0: empty, bucket b: 2^(b-1) <= len < 2^b
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned run = 0;
    for (int b = 32; b >= 0; b--) { s_base[b] = run; run += s_cnt[b]; }
  }
  __syncthreads();
  for (int t = threadIdx.x; t < tiles_total; t += 1024) {
    const uint2 r = ranges[t];
    const unsigned len = r.y - r.x;
    order[atomicAdd(&s_base[len ? 32 - __clz(len) : 0], 1u)] = (uint32_t)t;
  }
}

// Reordering costs locality (neighbouring tiles share Gaussians, and in raster order they run together and share
// L1/L2 lines): measured at C2, 27 cameras (28 waves of blocks) forward +4%; 4 cameras (4 waves) forward -28%,
// backward -22%.  So the order is used only when the launch is short enough for its tail to matter.
bool use_tile_order(int tiles_total) { return tiles_total < 12 * 148 * 6; }

int launch_tile_order(int tiles_total, const uint32_t* ranges, uint32_t* order, cudaStream_t st) {
  if (tiles_total <= 0 || !use_tile_order(tiles_total)) return GSB_OK;
  tile_order_kernel<<<1, 1024, 0, st>>>(tiles_total, reinterpret_cast<const uint2*>(ranges), order);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

// blockIdx.x -> (camera, tile column, tile row) through the heaviest-first order
struct TileCoord { int cam; uint32_t tile, bx, by; };
// n / d for n < 2^24 without the ~40-instruction integer division: float estimate + one correction either way
__device__ __forceinline__ uint32_t small_div(uint32_t n, uint32_t d, uint32_t& rem) {
  uint32_t q = (uint32_t)(__uint2float_rz(n) * __frcp_rn(__uint2float_rz(d)));
  int r = (int)n - (int)(q * d);
  if (r < 0) { q--; r += (int)d; }
  if (r >= (int)d) { q++; r -= (int)d; }
  rem = (uint32_t)r;
  return q;
}
__device__ __forceinline__ TileCoord tile_coord(const uint32_t* __restrict__ order, int tiles_x, int tiles_per_cam) {
  const uint32_t t = order ? order[blockIdx.x] : blockIdx.x;   // nullptr: raster order
  TileCoord c;
  c.cam = (int)small_div(t, (uint32_t)tiles_per_cam, c.tile);
  c.by = small_div(c.tile, (uint32_t)tiles_x, c.bx);
  return c;
}

#ifndef GSB_FWD_MINB
#define GSB_FWD_MINB 6
#endif
// COUNT: the measurement variant behind gsb_raster_count_pairs -- the same walk, no image outputs; per camera it
// adds up the walked pairs (sum of n_contrib) and the pairs actually composited (alpha >= 1/255 and not the
// terminating entry): the B and B_alpha of SURVEY.md section 8(d).
// (The patch mask of the backward -- a warp skips entries whose alpha >= 1/255 ellipse cannot reach its 8x4 patch --
// was measured SLOWER here: 1.40 vs 1.21 ms at C2, 3.10 vs 2.64 ms at C4.  The forward's per-entry cost for a warp
// that rejects is already only the 9-instruction power evaluation, and the mask's square roots and divisions in the
// staging phase plus one more branch per entry cost more than they save.)
// TWO: a second colour table (the segmentation colours of densify.py:136-150) is blended in the same walk into a
// second image: the two renders of a first-frame iteration share every alpha, T and termination decision, so the
// second one costs three more FMAs per composited pair instead of a second pass over the lists.
template <bool COUNT, bool TWO>
__global__ void __launch_bounds__(GSB_BLOCK_SIZE, GSB_FWD_MINB)
render_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ order, int tiles_x, int tiles_pc,
                  const uint32_t* __restrict__ point_list, int W, int H, int P,
                  const float2* __restrict__ means2D, const float* __restrict__ colors, int color_cam_stride,
                  const float4* __restrict__ conic_opacity, const float* __restrict__ depths,
                  const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                  float* __restrict__ out_color, float* __restrict__ out_depth,
                  unsigned long long* __restrict__ pair_counts, const float* __restrict__ colors2,
                  float* __restrict__ out_color2) {
  // one array, so that the inner loop addresses all of an entry's records from ONE register with immediate offsets
  // (as separate arrays every access recomputed its shared-window base: ~8 of the loop's 58 instructions)
  __shared__ float4 s_stage[TWO ? 4 : 3][GSB_BLOCK_SIZE];
  float4* const s_a = s_stage[0];    // x, y, early-reject power threshold, bits(list position + 1)
  float4* const s_co = s_stage[1];   // conic a, b, c, opacity
  float4* const s_cd = s_stage[2];   // r, g, b, depth
  float4* const s_c2 = s_stage[TWO ? 3 : 2];   // TWO: r, g, b of the second table
  __shared__ int s_wcnt[GSB_BLOCK_SIZE / 32];

  const TileCoord tc = tile_coord(order, tiles_x, tiles_pc);
  const int cam = tc.cam;
  const uint32_t tiles_per_cam = (uint32_t)tiles_pc;
  const uint32_t tile = tc.tile;
  const unsigned tid = threadIdx.y * GSB_BLOCK_X + threadIdx.x;
  // a warp owns an 8x4 pixel patch (not the reference's 16x2 strip): a splat's footprint then meets fewer
  // warps, and the warps it meets have more live lanes.  Per-pixel results do not depend on the mapping.
  const uint32_t px = tc.bx * GSB_BLOCK_X + ((tid >> 5) & 1u) * 8u + (tid & 7u);
  const uint32_t py = tc.by * GSB_BLOCK_Y + (tid >> 6) * 4u + ((tid >> 3) & 3u);
  const bool inside = px < (uint32_t)W && py < (uint32_t)H;
  const float pfx = (float)px, pfy = (float)py;
  const float tx0 = (float)(tc.bx * GSB_BLOCK_X), ty0 = (float)(tc.by * GSB_BLOCK_Y);
  bool done = !inside;

  const uint2 range = ranges[(size_t)cam * tiles_per_cam + tile];
  const int total = (int)(range.y - range.x);
  const int rounds = (total + GSB_BLOCK_SIZE - 1) / GSB_BLOCK_SIZE;
  const size_t cam_off = (size_t)cam * P;
  const float* col_tab = colors + (size_t)cam * color_cam_stride;

  float T = 1.0f;
  uint32_t last_contributor = 0;
  float C0 = 0.f, C1 = 0.f, C2 = 0.f;
  float E0 = 0.f, E1 = 0.f, E2 = 0.f;   // TWO: the second image
  float D = 15.0f;  // median depth default (cr/forward.cu:308)
  uint32_t blended = 0, evaluated = 0, staged = 0;   // COUNT only
  float slots = 0.f;

  for (int i = 0; i < rounds; i++) {
    if (__syncthreads_count(done) == GSB_BLOCK_SIZE) break;
    const int progress = i * GSB_BLOCK_SIZE + tid;
    bool keep = false;
    uint32_t id = 0;
    float2 xy = make_float2(0.f, 0.f);
    float4 co = make_float4(0.f, 0.f, 0.f, 0.f);
    float thr = 0.f;
    if (progress < total) {
      id = point_list[range.x + progress];
      xy = means2D[cam_off + id];
      co = conic_opacity[cam_off + id];
      keep = tile_cull_keep<false>(xy.x, xy.y, co, tx0, ty0, thr);
    }
    int slot;
    const int n = compact_slot(keep, tid, s_wcnt, slot);
    if (keep) {
      s_a[slot] = make_float4(xy.x, xy.y, thr, __uint_as_float((uint32_t)progress + 1u));
      s_co[slot] = co;
      s_cd[slot] = make_float4(col_tab[3 * (size_t)id], col_tab[3 * (size_t)id + 1], col_tab[3 * (size_t)id + 2],
                               depths[cam_off + id]);
      if (TWO) s_c2[slot] = make_float4(colors2[3 * (size_t)id], colors2[3 * (size_t)id + 1], colors2[3 * (size_t)id + 2], 0.f);
    }
    __syncthreads();
    if (COUNT && tid == 0) staged += (uint32_t)n;
    for (int j = 0; !done && j < n; j++) {
      if (COUNT) { evaluated++; slots += 32.0f / (float)__popc(__activemask()); }   // 32 lane slots per warp iteration
      const float4 a = s_a[j];
      const float4 co2 = s_co[j];
      const float dx = SUB(a.x, pfx), dy = SUB(a.y, pfy);
      const float power = blend_power(dx, dy, co2.x, co2.y, co2.z);
      if (power > 0.0f || power < a.z) continue;   // below a.z: certainly alpha < 1/255
      const float alpha = fminf(0.99f, MUL(co2.w, expf(power)));
      if (alpha < 1.0f / 255.0f) continue;
      const float test_T = MUL(T, SUB(1.0f, alpha));
      if (test_T < 0.0001f) { done = true; break; }
      const float4 cd = s_cd[j];
      C0 = FMA(T, MUL(alpha, cd.x), C0);
      C1 = FMA(T, MUL(alpha, cd.y), C1);
      C2 = FMA(T, MUL(alpha, cd.z), C2);
      if (TWO) {
        const float4 c2 = s_c2[j];
        E0 = FMA(T, MUL(alpha, c2.x), E0);
        E1 = FMA(T, MUL(alpha, c2.y), E1);
        E2 = FMA(T, MUL(alpha, c2.z), E2);
      }
      if (T > 0.5f && test_T < 0.5f) D = cd.w;
      T = test_T;
      last_contributor = __float_as_uint(a.w);
      if (COUNT) blended++;
    }
  }

  if (COUNT) {
    // per camera: walked pairs (sum of n_contrib), composited pairs, evaluated pairs (inner-loop iterations of live
    // pixels), lane slots issued (32 per warp iteration: evaluated / slots = lane utilisation), entries staged after
    // tile culling, entries listed
    unsigned long long v[4] = {last_contributor, blended, evaluated, (unsigned long long)(slots + 0.5f)};
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1)
#pragma unroll
      for (int k = 0; k < 4; k++) v[k] += __shfl_xor_sync(0xffffffffu, v[k], s);
    if ((tid & 31u) == 0)
      for (int k = 0; k < 4; k++) atomicAdd(pair_counts + 6 * cam + k, v[k]);
    if (tid == 0) {
      atomicAdd(pair_counts + 6 * cam + 4, (unsigned long long)staged);
      atomicAdd(pair_counts + 6 * cam + 5, (unsigned long long)total);
    }
    return;
  }

  // the tile coordinates are recomputed here rather than kept in registers across the loop (40-register budget)
  const TileCoord te = tile_coord(order, tiles_x, tiles_pc);
  const uint32_t ex = te.bx * GSB_BLOCK_X + ((tid >> 5) & 1u) * 8u + (tid & 7u);
  const uint32_t ey = te.by * GSB_BLOCK_Y + (tid >> 6) * 4u + ((tid >> 3) & 3u);
  if (ex < (uint32_t)W && ey < (uint32_t)H) {
    const size_t HW = (size_t)H * W;
    const size_t pid = (size_t)ey * W + ex;
    final_T[te.cam * HW + pid] = T;
    n_contrib[te.cam * HW + pid] = last_contributor;
    float* oc = out_color + (size_t)te.cam * 3 * HW;
    oc[pid] = FMA(bg[0], T, C0);
    oc[HW + pid] = FMA(bg[1], T, C1);
    oc[2 * HW + pid] = FMA(bg[2], T, C2);
    if (TWO) {
      float* o2 = out_color2 + (size_t)te.cam * 3 * HW;
      o2[pid] = FMA(bg[0], T, E0);
      o2[HW + pid] = FMA(bg[1], T, E1);
      o2[2 * HW + pid] = FMA(bg[2], T, E2);
    }
    out_depth[te.cam * HW + pid] = D;
  }
}

// A two-pixel-per-thread forward (128 threads, the layout of render_bwd2_kernel) was measured SLOWER (1.28 vs
// 1.21 ms at C2): the forward has no per-visit reduction to amortise, and a thread now runs until both of its
// pixels have terminated.

int launch_render_fwd(const gsb_raster_args& a, const gsb_raster_ctx& c, float* out_color, float* out_depth,
                      cudaStream_t st) {
  const int tiles_pc = c.tiles_x * c.tiles_y, tiles_total = tiles_pc * a.n_cams;
  const uint32_t* order = use_tile_order(tiles_total) ? c.ranges + 2 * (size_t)tiles_total : nullptr;
  dim3 block(GSB_BLOCK_X, GSB_BLOCK_Y);
  if (a.colors2 && (!a.out_color2 || a.color_cam_stride != 0)) return GSB_EINVAL;
  auto kern = a.colors2 ? render_fwd_kernel<false, true> : render_fwd_kernel<false, false>;
  kern<<<tiles_total, block, 0, st>>>(
      reinterpret_cast<const uint2*>(c.ranges), order, c.tiles_x, tiles_pc, c.point_list, a.width, a.height, a.P,
      reinterpret_cast<const float2*>(c.means2D), a.colors, a.color_cam_stride,
      reinterpret_cast<const float4*>(c.conic_opacity), c.depths, a.background, c.final_T, c.n_contrib, out_color,
      out_depth, nullptr, a.colors2, a.out_color2);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

int launch_count_pairs(const gsb_raster_args& a, const gsb_raster_ctx& c, unsigned long long* counts, cudaStream_t st) {
  const int tiles_pc = c.tiles_x * c.tiles_y, tiles_total = tiles_pc * a.n_cams;
  GSB_CUDA_CHECK(cudaMemsetAsync(counts, 0, sizeof(unsigned long long) * 6 * (size_t)a.n_cams, st));
  dim3 block(GSB_BLOCK_X, GSB_BLOCK_Y);
  render_fwd_kernel<true, false><<<tiles_total, block, 0, st>>>(
      reinterpret_cast<const uint2*>(c.ranges), nullptr, c.tiles_x, tiles_pc, c.point_list, a.width, a.height, a.P,
      reinterpret_cast<const float2*>(c.means2D), a.colors, a.color_cam_stride,
      reinterpret_cast<const float4*>(c.conic_opacity), c.depths, a.background, nullptr, nullptr, nullptr, nullptr,
      counts, nullptr, nullptr);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

// ------------------------------------------------------------------------------------------------
// Backward
// ------------------------------------------------------------------------------------------------

// Sum 8 values over the 32 lanes with 9 shuffles.  Afterwards lane L (L % 4 == 0) holds in v[0] the
// full warp sum of component L / 4.
__device__ __forceinline__ void warp_reduce8(float (&v)[8], unsigned lane) {
  const unsigned full = 0xffffffffu;
  {
    const bool up = lane & 16u;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const float send = up ? v[i] : v[i + 4];
      const float keep = up ? v[i + 4] : v[i];
      v[i] = keep + __shfl_xor_sync(full, send, 16);
    }
  }
  {
    const bool up = lane & 8u;
#pragma unroll
    for (int i = 0; i < 2; i++) {
      const float send = up ? v[i] : v[i + 2];
      const float keep = up ? v[i + 2] : v[i];
      v[i] = keep + __shfl_xor_sync(full, send, 8);
    }
  }
  {
    const bool up = lane & 4u;
    const float send = up ? v[0] : v[1];
    const float keep = up ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(full, send, 4);
  }
  v[0] += __shfl_xor_sync(full, v[0], 2);
  v[0] += __shfl_xor_sync(full, v[0], 1);
}

// ------------------------------------------------------------------------------------------------
// Backward, two pixels per thread: 128 threads per tile, a warp owns an 8x8 block = two of the 8x4 patches
// above (rows y and y + 4 of the same column), so one staged entry that touches both patches costs ONE
// loop iteration, ONE 9-value warp reduction and ONE atomic instruction instead of two.  Per-pixel
// arithmetic is the reference's; only the order in which partial sums meet changes.  (A one-pixel-per-thread kernel,
// 256 threads per tile, preceded this one: 2.64 vs 2.21 ms at C2; removed.)
// ------------------------------------------------------------------------------------------------
struct BwdPixel {
  float T, ar0, ar1, ar2, dpx0, dpx1, dpx2, nTf_bg, pfy;
  int last;
  float br0, br1, br2, dq0, dq1, dq2, nTf_bgi;   // TWO: accum_rec / dL/dpixel of the second image, background term of the first
};

// TWO: the pixel belongs to two images blended from two colour tables in one walk (render_fwd_kernel<.., TWO>): alpha, T
// and the decisions are shared, dL/dalpha is the sum over both images' channels, the second table's colour gradient
// goes to w[1..3] (w[0] carries the opacity term g_op), and w[4..5] are the first moments of the FIRST image's share
// alone: the densification statistic is the norm of the colour render's screen-space gradient, not of the sum
// (densify.py:147-150 reads means_2d.grad of the colour render's carrier).
template <bool TWO>
__device__ __forceinline__ void bwd_pixel(BwdPixel& s, bool valid, float alpha, float G, float dx, float dy, float4 ci,
                                          float4 ci2, float opac, float (&v)[8], float (&w)[8]) {
  float& g_op = w[0];
  if (!valid) return;
  const float om = 1.f - alpha;
  float rcp;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rcp) : "f"(om));
  rcp = fmaf(rcp, fmaf(-om, rcp, 1.f), rcp);
  s.T = s.T * rcp;
  const float dchannel_dcolor = alpha * s.T;
  const float d0 = ci.x - s.ar0, d1 = ci.y - s.ar1, d2 = ci.z - s.ar2;
  float dL_dalpha = d0 * s.dpx0;
  dL_dalpha = fmaf(d1, s.dpx1, dL_dalpha);
  dL_dalpha = fmaf(d2, s.dpx2, dL_dalpha);
  s.ar0 = fmaf(alpha, d0, s.ar0);
  s.ar1 = fmaf(alpha, d1, s.ar1);
  s.ar2 = fmaf(alpha, d2, s.ar2);
  v[0] = fmaf(dchannel_dcolor, s.dpx0, v[0]);
  v[1] = fmaf(dchannel_dcolor, s.dpx1, v[1]);
  v[2] = fmaf(dchannel_dcolor, s.dpx2, v[2]);
  if (TWO) {
    const float wgi = opac * (G * fmaf(dL_dalpha, s.T, s.nTf_bgi * rcp));   // the first image's share of opac * dL/dG * G
    w[4] = fmaf(wgi, dx, w[4]);
    w[5] = fmaf(wgi, dy, w[5]);
    const float e0 = ci2.x - s.br0, e1 = ci2.y - s.br1, e2 = ci2.z - s.br2;
    dL_dalpha = fmaf(e0, s.dq0, dL_dalpha);
    dL_dalpha = fmaf(e1, s.dq1, dL_dalpha);
    dL_dalpha = fmaf(e2, s.dq2, dL_dalpha);
    s.br0 = fmaf(alpha, e0, s.br0);
    s.br1 = fmaf(alpha, e1, s.br1);
    s.br2 = fmaf(alpha, e2, s.br2);
    w[1] = fmaf(dchannel_dcolor, s.dq0, w[1]);
    w[2] = fmaf(dchannel_dcolor, s.dq1, w[2]);
    w[3] = fmaf(dchannel_dcolor, s.dq2, w[3]);
  }
  dL_dalpha = fmaf(dL_dalpha, s.T, s.nTf_bg * rcp);
  const float g = G * dL_dalpha;
  g_op += g;
  const float wg = opac * g;
  const float s1 = wg * dx, s2 = wg * dy;
  v[3] += s1;
  v[4] += s2;
  v[5] = fmaf(s1, dx, v[5]);
  v[6] = fmaf(s1, dy, v[6]);
  v[7] = fmaf(s2, dy, v[7]);
}

#ifndef GSB_BWD2_MINB
#define GSB_BWD2_MINB 6
#endif

template <bool TWO>
__global__ void __launch_bounds__(128, TWO ? 5 : GSB_BWD2_MINB)
render_bwd2_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ order, int tiles_x, int tiles_pc,
                   const uint32_t* __restrict__ point_list, int W, int H, int P,
                   const float* __restrict__ bg, const float2* __restrict__ means2D,
                   const float4* __restrict__ conic_opacity, const float* __restrict__ colors, int color_cam_stride,
                   const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                   const float* __restrict__ dL_dpixels, float* __restrict__ dL_dmean2D,
                   float* __restrict__ dL_dconic, float* __restrict__ dL_dopacity, float* __restrict__ dL_dcolor,
                   const float* __restrict__ colors2, const float* __restrict__ dL_dpixels2,
                   float* __restrict__ dL_dcolor2, float* __restrict__ dL_dmean2D_first /* [n_cams][P][2] */) {
  // one array: an entry's records are addressed from one register with immediate offsets (see render_fwd_kernel)
  __shared__ float4 s_stage[TWO ? 4 : 3][GSB_BLOCK_SIZE];
  float4* const s_a = s_stage[0];    // x, y, early-reject power threshold | patch mask, bits(list position)
  float4* const s_co = s_stage[1];   // conic a, b, c, opacity
  float4* const s_ci = s_stage[2];   // r, g, b, bits(id)
  float4* const s_c2 = s_stage[TWO ? 3 : 2];   // TWO: r, g, b of the second colour table
  __shared__ int s_wcnt[8];
  __shared__ int s_block_last;
  __shared__ int s_patch_last[8];

  const TileCoord tc = tile_coord(order, tiles_x, tiles_pc);
  const int cam = tc.cam;
  const uint32_t tiles_per_cam = (uint32_t)tiles_pc;
  const uint32_t tile = tc.tile;
  const unsigned tid = threadIdx.x;
  const unsigned lane = tid & 31u, warp = tid >> 5;
  const unsigned pt = (warp & 1u) + 4u * (warp >> 1), pb = pt + 2u;   // the two 8x4 patches of this warp
  const uint32_t px = tc.bx * GSB_BLOCK_X + (warp & 1u) * 8u + (lane & 7u);
  const uint32_t pyA = tc.by * GSB_BLOCK_Y + (warp >> 1) * 8u + (lane >> 3);
  const uint32_t pyB = pyA + 4u;
  const float pfx = (float)px;
  const float tx0 = (float)(tc.bx * GSB_BLOCK_X), ty0 = (float)(tc.by * GSB_BLOCK_Y);
  const size_t HW = (size_t)H * W;

  const uint2 range = ranges[(size_t)cam * tiles_per_cam + tile];
  const size_t cam_off = (size_t)cam * P;
  const float* col_tab = colors + (size_t)cam * color_cam_stride;
  float* dcol_tab = dL_dcolor + (size_t)cam * color_cam_stride;
  const float b0 = bg[0], b1 = bg[1], b2 = bg[2];

  BwdPixel A, B;
  auto init = [&](BwdPixel& s, uint32_t py) {
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const size_t pid = (size_t)py * W + px;
    const float T_final = inside ? final_Ts[cam * HW + pid] : 0.f;
    s.T = T_final;
    s.last = inside ? (int)n_contrib[cam * HW + pid] : 0;
    s.dpx0 = s.dpx1 = s.dpx2 = 0.f;
    if (inside) {
      const float* dp = dL_dpixels + (size_t)cam * 3 * HW;
      s.dpx0 = dp[pid]; s.dpx1 = dp[HW + pid]; s.dpx2 = dp[2 * HW + pid];
    }
    float bg_dot = 0.f;
    bg_dot += b0 * s.dpx0; bg_dot += b1 * s.dpx1; bg_dot += b2 * s.dpx2;
    s.ar0 = s.ar1 = s.ar2 = 0.f;
    s.br0 = s.br1 = s.br2 = 0.f;
    s.dq0 = s.dq1 = s.dq2 = 0.f;
    s.nTf_bgi = -T_final * bg_dot;
    if (TWO) {
      if (inside) {
        const float* dq = dL_dpixels2 + (size_t)cam * 3 * HW;
        s.dq0 = dq[pid]; s.dq1 = dq[HW + pid]; s.dq2 = dq[2 * HW + pid];
      }
      bg_dot += b0 * s.dq0; bg_dot += b1 * s.dq1; bg_dot += b2 * s.dq2;   // both images sit on the same background
    }
    s.nTf_bg = -T_final * bg_dot;
    s.pfy = (float)py;
  };
  init(A, pyA);
  init(B, pyB);

  // writer lanes: after warp_reduce8 the four lanes 4 k .. 4 k + 3 hold component k.  Lane 4 k writes component k of
  // the first group; lane 4 k + 1 component k of the second group (k = 0: the opacity term; TWO: k = 1..3 the second
  // table's colour gradient, k = 4, 5 the first image's first moments)
  float* lane_base;
  unsigned lane_stride;
  {
    const unsigned comp = lane >> 2;
    if (lane == 1) { lane_base = dL_dopacity; lane_stride = 1; }
    else if (TWO && (lane & 3u) == 1u && comp < 4) { lane_base = dL_dcolor2 + (comp - 1u); lane_stride = 3; }
    else if (TWO && (lane & 3u) == 1u) { lane_base = dL_dmean2D_first + 2 * cam_off + (comp & 1u); lane_stride = 2; }
    else if (comp < 3) { lane_base = dcol_tab + comp; lane_stride = 3; }
    else if (comp < 5) { lane_base = dL_dmean2D + 3 * cam_off + (comp - 3); lane_stride = 3; }
    else { lane_base = dL_dconic + 4 * cam_off + (comp == 7 ? 3 : comp - 5); lane_stride = 4; }
  }
  const bool lane_writes = ((lane & 3u) == 0) || lane == 1 || (TWO && (lane & 3u) == 1u && lane < 24u);

  int lastA = A.last, lastB = B.last;
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    lastA = max(lastA, __shfl_xor_sync(0xffffffffu, lastA, s));
    lastB = max(lastB, __shfl_xor_sync(0xffffffffu, lastB, s));
  }
  if (tid == 0) s_block_last = 0;
  __syncthreads();
  if (lane == 0) {
    s_patch_last[pt] = lastA;
    s_patch_last[pb] = lastB;
    const int m = max(lastA, lastB);
    if (m > 0) atomicMax(&s_block_last, m);
  }
  __syncthreads();
  const int block_last = s_block_last;
  if (block_last == 0) return;

  const int rounds = (block_last + GSB_BLOCK_SIZE - 1) / GSB_BLOCK_SIZE;
  for (int i = 0; i < rounds; i++) {
    if (i > 0) __syncthreads();
    // each thread stages two list positions of the 256-entry batch: k = tid and k = tid + 128
    bool keep[2];
    uint32_t id[2];
    float2 xy[2];
    float4 co[2];
    float thr[2];
    int pos[2];
#pragma unroll
    for (int h = 0; h < 2; h++) {
      pos[h] = block_last - 1 - i * GSB_BLOCK_SIZE - (int)(tid + 128u * h);
      keep[h] = false; id[h] = 0; xy[h] = make_float2(0.f, 0.f); co[h] = make_float4(0.f, 0.f, 0.f, 0.f); thr[h] = 0.f;
      if (pos[h] >= 0) {
        id[h] = point_list[range.x + (uint32_t)pos[h]];
        xy[h] = means2D[cam_off + id[h]];
        co[h] = conic_opacity[cam_off + id[h]];
      }
    }
#pragma unroll
    for (int h = 0; h < 2; h++) {
      if (pos[h] >= 0) {
        keep[h] = tile_cull_keep<true>(xy[h].x, xy[h].y, co[h], tx0, ty0, thr[h]);
        if (keep[h]) {
          unsigned m = __float_as_uint(thr[h]) & 0xFFu;
#pragma unroll
          for (int w = 0; w < 8; w++)
            if (pos[h] >= s_patch_last[w]) m &= ~(1u << w);
          thr[h] = pack_thr(thr[h], m);
          keep[h] = m != 0;
        }
      }
    }
    const unsigned bal0 = __ballot_sync(0xffffffffu, keep[0]), bal1 = __ballot_sync(0xffffffffu, keep[1]);
    if (lane == 0) { s_wcnt[warp] = __popc(bal0); s_wcnt[4 + warp] = __popc(bal1); }
    __syncthreads();
    int before0 = 0, before1 = 0, n = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
      const int c = s_wcnt[w];
      if (w < (int)warp) before0 += c;
      if (w < (int)warp + 4) before1 += c;
      n += c;
    }
    const int slot[2] = {before0 + __popc(bal0 & lanemask_lt()), before1 + __popc(bal1 & lanemask_lt())};
#pragma unroll
    for (int h = 0; h < 2; h++)
      if (keep[h]) {
        s_a[slot[h]] = make_float4(xy[h].x, xy[h].y, thr[h], __int_as_float(pos[h]));
        s_co[slot[h]] = co[h];
        s_ci[slot[h]] = make_float4(col_tab[3 * (size_t)id[h]], col_tab[3 * (size_t)id[h] + 1],
                                    col_tab[3 * (size_t)id[h] + 2], __uint_as_float(id[h]));
        if (TWO)
          s_c2[slot[h]] = make_float4(colors2[3 * (size_t)id[h]], colors2[3 * (size_t)id[h] + 1],
                                      colors2[3 * (size_t)id[h] + 2], 0.f);
      }
    __syncthreads();
    for (int j = 0; j < n; j++) {
      const float4 a = s_a[j];
      const unsigned m = __float_as_uint(a.z);
      const bool mt = (m >> pt) & 1u, mb = (m >> pb) & 1u;   // warp-uniform
      if (!(mt || mb)) continue;
      const int pos_j = __float_as_int(a.w);
      const float4 co2 = s_co[j];
      const float dx = SUB(a.x, pfx);
      const float dyA = SUB(a.y, A.pfy), dyB = SUB(a.y, B.pfy);
      bool preA = false, preB = false;
      float powA = 0.f, powB = 0.f;
      if (mt) {
        powA = blend_power(dx, dyA, co2.x, co2.y, co2.z);
        preA = (pos_j < A.last) && !(powA > 0.0f) && !(powA < a.z);
      }
      if (mb) {
        powB = blend_power(dx, dyB, co2.x, co2.y, co2.z);
        preB = (pos_j < B.last) && !(powB > 0.0f) && !(powB < a.z);
      }
      if (!__any_sync(0xffffffffu, preA || preB)) continue;
      const float GA = expf(powA), GB = expf(powB);
      const float alphaA = fminf(0.99f, MUL(co2.w, GA)), alphaB = fminf(0.99f, MUL(co2.w, GB));
      const bool validA = preA && !(alphaA < 1.0f / 255.0f), validB = preB && !(alphaB < 1.0f / 255.0f);
      if (!__any_sync(0xffffffffu, validA || validB)) continue;

      float v[8], w[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int k = 0; k < 8; k++) v[k] = 0.f;
      const float4 ci = s_ci[j];
      const float4 ci2 = TWO ? s_c2[j] : make_float4(0.f, 0.f, 0.f, 0.f);
      bwd_pixel<TWO>(A, validA, alphaA, GA, dx, dyA, ci, ci2, co2.w, v, w);
      bwd_pixel<TWO>(B, validB, alphaB, GB, dx, dyB, ci, ci2, co2.w, v, w);
      warp_reduce8(v, lane);
      float extra;
      if (TWO) {
        warp_reduce8(w, lane);   // lanes 4 k .. 4 k + 3 hold component k
        extra = w[0];
      } else {
        extra = w[0];
#pragma unroll
        for (int s = 16; s >= 1; s >>= 1) extra += __shfl_xor_sync(0xffffffffu, extra, s);
      }
      const uint32_t id2 = __float_as_uint(ci.w);
      const float val = ((lane & 3u) == 1u) ? extra : v[0];
      if (lane_writes && val != 0.f) atomicAdd(lane_base + id2 * lane_stride, val);
    }
  }
}

int launch_render_bwd(const gsb_raster_args& a, const gsb_raster_ctx& c, const float* dL_dpix, float* dL_dmean2D,
                      float* dL_dconic, float* dL_dopacity, float* dL_dcolor, cudaStream_t st, const float* dL_dpix2,
                      float* dL_dcolor2, float* dL_dmean2D_first) {
  const int tiles_pc = c.tiles_x * c.tiles_y, tiles_total = tiles_pc * a.n_cams;
  const uint32_t* order = use_tile_order(tiles_total) ? c.ranges + 2 * (size_t)tiles_total : nullptr;
  if (dL_dpix2) {   // both images of a two-table forward in one walk
    if (!a.colors2 || !dL_dcolor2 || !dL_dmean2D_first || a.color_cam_stride != 0) return GSB_EINVAL;
    render_bwd2_kernel<true><<<tiles_total, 128, 0, st>>>(
        reinterpret_cast<const uint2*>(c.ranges), order, c.tiles_x, tiles_pc, c.point_list, a.width, a.height, a.P,
        a.background,
        reinterpret_cast<const float2*>(c.means2D), reinterpret_cast<const float4*>(c.conic_opacity), a.colors,
        a.color_cam_stride, c.final_T, c.n_contrib, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, a.colors2,
        dL_dpix2, dL_dcolor2, dL_dmean2D_first);
    GSB_LAUNCH_CHECK();
    return GSB_OK;
  }
  render_bwd2_kernel<false><<<tiles_total, 128, 0, st>>>(
      reinterpret_cast<const uint2*>(c.ranges), order, c.tiles_x, tiles_pc, c.point_list, a.width, a.height, a.P,
      a.background,
      reinterpret_cast<const float2*>(c.means2D), reinterpret_cast<const float4*>(c.conic_opacity), a.colors,
      a.color_cam_stride, c.final_T, c.n_contrib, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, nullptr,
      nullptr, nullptr, nullptr);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

}  // namespace gsb
