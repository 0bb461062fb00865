// binning.cu -- inclusive scan, (tile|depth) key emission, hand-written onesweep radix sort, tile ranges.
//
// Replaces cub::DeviceScan::InclusiveSum (cr/rasterizer_impl.cu:277-278), duplicateWithKeys (:70-111),
// cub::DeviceRadixSort::SortPairs (:304-309) and identifyTileRanges (:116-138).  No CUB/Thrust.
// All integer work: results are bit-identical to the reference by construction (inclusive prefix
// sum; *stable* ascending LSD sort on bits [0, end_bit)).
#include <algorithm>
#include <mutex>
#include "common.cuh"

namespace gsb {

// ------------------------------------------------------------------------------------------------
// Single-pass inclusive scan with decoupled look-back (one read + one write of the data).
// Tile = 256 threads x 8 items.  Tile status word: (flag << 32) | value, flag 1 = aggregate,
// 2 = inclusive prefix.  Tiles take their index from an atomic ticket, so a tile can only wait on
// tiles that already started (forward-progress safe).
// ------------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

size_t scan_scratch_bytes(int64_t n) {
  const int64_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  return align_up(16 + (size_t)(tiles > 0 ? tiles : 1) * sizeof(unsigned long long), 256);
}

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(SCAN_THREADS)
scan_u32_kernel(long long n, const uint32_t* in, uint32_t* out,
                unsigned int* __restrict__ ticket, unsigned long long* __restrict__ status,
                const uint32_t* __restrict__ perm, int perm_block) {
  __shared__ uint32_t s_warp[SCAN_THREADS / 32];
  __shared__ uint32_t s_tile, s_prefix;
  if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
  __syncthreads();
  const uint32_t tile = s_tile;
  const long long base = (long long)tile * SCAN_TILE + (long long)threadIdx.x * SCAN_ITEMS;

  uint32_t v[SCAN_ITEMS];
  if (perm) {
    // element i is in[(i / perm_block) * perm_block + perm[i]]: the scan runs in a permuted order.  One division
    // per thread; the block index advances by at most one per element.
    uint32_t pidx[SCAN_ITEMS];
    if (base + SCAN_ITEMS <= n && ((reinterpret_cast<uintptr_t>(perm) & 15u) == 0)) {
      const uint4 a = *reinterpret_cast<const uint4*>(perm + base);
      const uint4 b = *reinterpret_cast<const uint4*>(perm + base + 4);
      pidx[0] = a.x; pidx[1] = a.y; pidx[2] = a.z; pidx[3] = a.w; pidx[4] = b.x; pidx[5] = b.y; pidx[6] = b.z; pidx[7] = b.w;
    } else {
#pragma unroll
      for (int k = 0; k < SCAN_ITEMS; k++) pidx[k] = (base + k < n) ? perm[base + k] : 0u;
    }
    long long blk0 = (base / perm_block) * (long long)perm_block;
    int rem = (int)(base - blk0);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
      if (rem == perm_block) { rem = 0; blk0 += perm_block; }
      v[k] = (base + k < n) ? in[blk0 + pidx[k]] : 0u;
      rem++;
    }
  } else if (base + SCAN_ITEMS <= n && ((reinterpret_cast<uintptr_t>(in) & 15u) == 0)) {
    const uint4 a = *reinterpret_cast<const uint4*>(in + base);
    const uint4 b = *reinterpret_cast<const uint4*>(in + base + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) v[k] = (base + k < n) ? in[base + k] : 0u;
  }
  uint32_t run = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { run += v[k]; v[k] = run; }

  // block exclusive scan of the per-thread totals
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  uint32_t inc = run;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += t;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  uint32_t warp_off = 0, block_total = 0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; w++) {
    const uint32_t t = s_warp[w];
    if (w < warp) warp_off += t;
    block_total += t;
  }
  const uint32_t thread_excl = warp_off + inc - run;

  // decoupled look-back by warp 0: 32 predecessors per round trip (a first version walked them one at a time from
  // thread 0 -- ~17 dependent L2 round trips per tile, which made the kernel latency-bound at 0.17 of HBM speed)
  if (warp == 0) {
    uint32_t prefix = 0;
    if (tile == 0) {
      if (lane == 0) st_relaxed_u64(status, (2ull << 32) | block_total);
    } else {
      if (lane == 0) st_relaxed_u64(status + tile, (1ull << 32) | block_total);
      long long p = (long long)tile - 1;   // nearest predecessor not yet accounted for
      while (true) {
        const long long idx = p - (long long)lane;
        unsigned long long sv = (2ull << 32);   // in front of tile 0: an inclusive prefix of 0
        if (idx >= 0) {
          sv = ld_relaxed_u64(status + idx);
          while ((sv >> 32) == 0) sv = ld_relaxed_u64(status + idx);
        }
        const unsigned incl = __ballot_sync(0xffffffffu, (sv >> 32) == 2);
        const int first = incl ? __ffs(incl) - 1 : 31;          // lanes 0..first are needed
        uint32_t v = ((int)lane <= first) ? (uint32_t)sv : 0u;
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
        prefix += v;
        if (incl) break;
        p -= 32;
      }
      if (lane == 0) st_relaxed_u64(status + tile, (2ull << 32) | (uint32_t)(prefix + block_total));
    }
    if (lane == 0) s_prefix = prefix;
  }
  __syncthreads();
  const uint32_t add = s_prefix + thread_excl;
  if (base + SCAN_ITEMS <= n && ((reinterpret_cast<uintptr_t>(out) & 15u) == 0)) {
    *reinterpret_cast<uint4*>(out + base) = make_uint4(v[0] + add, v[1] + add, v[2] + add, v[3] + add);
    *reinterpret_cast<uint4*>(out + base + 4) = make_uint4(v[4] + add, v[5] + add, v[6] + add, v[7] + add);
  } else {
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++)
      if (base + k < n) out[base + k] = v[k] + add;
  }
}

int launch_inclusive_scan_u32(int64_t n, const uint32_t* in, uint32_t* out, void* scratch, cudaStream_t st,
                              const uint32_t* perm, int perm_block) {
  if (n <= 0) return GSB_OK;
  const int64_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  GSB_CUDA_CHECK(cudaMemsetAsync(scratch, 0, scan_scratch_bytes(n), st));
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  unsigned long long* status = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(scratch) + 16);
  scan_u32_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, st>>>(n, in, out, ticket, status, perm, perm_block);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free variant of the scan (gsb_inclusive_scan_u32_static): the ticket + tile status words live in a
// file-scope device array instead of a caller buffer.  Every GPU that loads the library holds the array until exit.
// In host code `g_scan_static` is only a stand-in for the symbol, NOT a device address: the address is taken once per
// device with cudaGetSymbolAddress (that device current) and that pointer is what the kernel receives.
// ------------------------------------------------------------------------------------------------
constexpr int64_t SCAN_STATIC_MAX_ITEMS = (int64_t)GSB_SCAN_STATIC_MAX_ITEMS;
constexpr size_t SCAN_STATIC_BYTES =
    (16 + (size_t)((SCAN_STATIC_MAX_ITEMS + SCAN_TILE - 1) / SCAN_TILE) * sizeof(unsigned long long) + 255) / 256 * 256;
static __device__ __align__(256) unsigned char g_scan_static[SCAN_STATIC_BYTES];

constexpr int SCAN_STATIC_MAX_DEVICES = 64;
static void* g_scan_static_ptr[SCAN_STATIC_MAX_DEVICES];   // device address of g_scan_static, per device ordinal
static std::mutex g_scan_static_mu;

static int scan_static_address(void** out) {
  int dev = -1;
  GSB_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= SCAN_STATIC_MAX_DEVICES) return GSB_EINVAL;
  std::lock_guard<std::mutex> lock(g_scan_static_mu);
  if (!g_scan_static_ptr[dev]) {
    void* p = nullptr;
    GSB_CUDA_CHECK(cudaGetSymbolAddress(&p, g_scan_static));
    if (!p) return GSB_ECUDA;
    g_scan_static_ptr[dev] = p;
  }
  *out = g_scan_static_ptr[dev];
  return GSB_OK;
}

static int launch_inclusive_scan_u32_static(int64_t n, const uint32_t* in, uint32_t* out, cudaStream_t st) {
  if (n <= 0) return GSB_OK;
  if (n > SCAN_STATIC_MAX_ITEMS) {
    set_last_error("gsb_inclusive_scan_u32_static: n exceeds GSB_SCAN_STATIC_MAX_ITEMS (use gsb_inclusive_scan_u32 "
                   "with a caller scratch buffer)", cudaErrorInvalidValue);
    return GSB_EINVAL;
  }
  void* scratch = nullptr;
  const int rc = scan_static_address(&scratch);
  if (rc) return rc;
  // the array keeps the previous call's ticket and status words: clear what this call will use
  return launch_inclusive_scan_u32(n, in, out, scratch, st, nullptr, 0);
}
}  // namespace gsb

extern "C" int gsb_inclusive_scan_u32_static(int32_t n, const uint32_t* in, uint32_t* out, void* stream) {
  if (n < 0 || (n > 0 && (!in || !out))) return GSB_EINVAL;
  return gsb::launch_inclusive_scan_u32_static(n, in, out, reinterpret_cast<cudaStream_t>(stream));
}

namespace gsb {

// ------------------------------------------------------------------------------------------------
// duplicateWithKeys (cr/rasterizer_impl.cu:70-111), warp-cooperative: a warp owns 32 consecutive
// (camera, Gaussian) slots, and its lanes walk the warp's contiguous instance range together, so the
// key/value stores are coalesced (the reference runs a divergent serial loop per thread).
// The slots arrive camera-major and depth-sorted (gauss_order), so the depth digits of the reference's key
// are already in order; what is emitted per instance is the camera-local tile id y*grid.x + x (16 bits)
// and the Gaussian index in its camera.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void tile_rect_dup(float px, float py, int rad, int gx, int gy, uint32_t& x0, uint32_t& y0,
                                              uint32_t& x1, uint32_t& y1) {
  const float rf = (float)rad;
  const int ix0 = __float2int_rz(MUL(SUB(px, rf), 0.0625f));
  const int iy0 = __float2int_rz(MUL(SUB(py, rf), 0.0625f));
  const int ix1 = __float2int_rz(MUL(ADD(ADD(ADD(px, rf), 16.0f), -1.0f), 0.0625f));
  const int iy1 = __float2int_rz(MUL(ADD(ADD(ADD(py, rf), 16.0f), -1.0f), 0.0625f));
  x0 = min((uint32_t)max(ix0, 0), (uint32_t)gx);
  y0 = min((uint32_t)max(iy0, 0), (uint32_t)gy);
  x1 = min((uint32_t)max(ix1, 0), (uint32_t)gx);
  y1 = min((uint32_t)max(iy1, 0), (uint32_t)gy);
}

__global__ void __launch_bounds__(256)
duplicate_with_keys_kernel(long long total, int P, int gx, int gy, const uint32_t* __restrict__ gauss_order,
                           const float2* __restrict__ means2D, const int32_t* __restrict__ radii,
                           const uint32_t* __restrict__ offsets, uint16_t* __restrict__ tile_keys,
                           uint32_t* __restrict__ vals, const uint32_t* __restrict__ n_emit,
                           const CamParams* __restrict__ cams) {
  const long long slot = (long long)blockIdx.x * 256 + threadIdx.x;
  const unsigned lane = threadIdx.x & 31u;
  uint32_t end = 0, start = 0, x0 = 0, y0 = 0, w = 1, gidx = 0;
  const long long warp_first = slot - lane;
  if (warp_first >= total) return;
  if (slot < total) {
    end = offsets[slot];
    const uint32_t cam = (uint32_t)(slot / P);
    gidx = gauss_order[slot];
    const size_t o = (size_t)cam * P + gidx;
    const int rad = radii[o];
    start = end;
    if (rad > 0) {
      const float2 xy = means2D[o];
      uint32_t x1, y1;
      tile_rect_dup(xy.x, xy.y, rad, gx, gy, x0, y0, x1, y1);
      const uint32_t r0 = (uint32_t)cams[cam].row0, r1 = (uint32_t)cams[cam].row1;   // tile-band sharding
      y0 = min(max(y0, r0), r1);
      y1 = min(max(y1, r0), r1);
      w = x1 - x0;
      start = end - w * (y1 - y0);
      if (w == 0) w = 1;
    }
  }
  // slots past `total` inherit the last valid end so that the search below stays monotone
  uint32_t last_end = __shfl_sync(0xffffffffu, end, (unsigned)min(31ll, total - 1 - warp_first));
  if (slot >= total) { end = last_end; start = last_end; }
  const uint32_t warp_start = __shfl_sync(0xffffffffu, start, 0);
  // *n_emit = seg_start[n_cams]: the instance total, or 0 when it did not fit the arena (capacity guard)
  const uint32_t warp_end = min(last_end, __ldg(n_emit));
  // owner of output position i = number of lanes whose `end` <= i (ends are non-decreasing).  When every lane
  // of the warp owns at least one instance the ends are distinct, so that count is a population count over
  // the bit set of ends that fall into the current 32-wide window (one warp OR-reduction); warps holding empty
  // slots (culled Gaussians at the tail of a camera) take a 5-step shuffle search instead.
  const bool has_empty = __any_sync(0xffffffffu, end == start);
  const uint32_t xy0 = x0 | (y0 << 16);
  const float inv_w = __frcp_rn((float)w);
  for (uint32_t b = warp_start; b < warp_end; b += 32) {
    const uint32_t i = b + lane;
    unsigned j;
    if (!has_empty) {
      const uint32_t rel = end - b;                      // wraps for ends before the window
      const unsigned before = __popc(__ballot_sync(0xffffffffu, end <= b));
      const unsigned bits = __reduce_or_sync(0xffffffffu, (end > b && rel <= 32u) ? (1u << (rel - 1u)) : 0u);
      // bit t set: some lane ends at b + t + 1, i.e. it owns position b + t but not b + t + 1
      j = before + __popc(bits & ((1u << lane) - 1u));
    } else {
      j = 0;
#pragma unroll
      for (int step = 16; step >= 1; step >>= 1) {
        const uint32_t e = __shfl_sync(0xffffffffu, end, (j + step - 1) & 31u);
        if (e <= i) j += step;
      }
    }
    j &= 31u;
    const uint32_t s = __shfl_sync(0xffffffffu, start, j);
    const uint32_t oxy = __shfl_sync(0xffffffffu, xy0, j);
    const uint32_t ow = __shfl_sync(0xffffffffu, w, j);
    const float oinv = __shfl_sync(0xffffffffu, inv_w, j);
    const uint32_t og = __shfl_sync(0xffffffffu, gidx, j);
    if (i < warp_end) {
      const uint32_t k = i - s;
      // k / ow by reciprocal multiplication with an exact fix-up (k < 65536 tiles)
      uint32_t ry = __float2uint_rz(__uint2float_rn(k) * oinv);
      int rx = (int)k - (int)(ry * ow);
      if (rx < 0) { ry--; rx += (int)ow; }
      else if (rx >= (int)ow) { ry++; rx -= (int)ow; }
      tile_keys[i] = (uint16_t)(((oxy >> 16) + ry) * (uint32_t)gx + ((oxy & 0xFFFFu) + (uint32_t)rx));
      vals[i] = og;
    }
  }
}

int launch_duplicate_with_keys(const gsb_raster_ctx& c, const CamParams* cams_dev, cudaStream_t st) {
  const long long total = (long long)c.n_cams * c.P;
  if (total <= 0 || c.capacity <= 0) return GSB_OK;
  duplicate_with_keys_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(
      total, c.P, c.tiles_x, c.tiles_y, c.gauss_order, reinterpret_cast<const float2*>(c.means2D), c.radii,
      c.point_offsets, c.tile_keys_unsorted, c.point_list_unsorted, c.seg_start + c.n_cams, cams_dev);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

// ------------------------------------------------------------------------------------------------
// General-purpose onesweep LSD radix sort, 8-bit digits, 64-bit keys + 32-bit payloads, stable
// (gsb_sort_pairs; the product path uses the segmented wide-digit sort of segsort.cu).
//   kernel 1: all per-pass digit histograms in one read of the keys;
//   kernel 2 (per pass): each tile ranks its keys (warp match), publishes its digit counts and
//     resolves its global offsets with a decoupled look-back over earlier tiles, then scatters
//     through shared memory so that global stores are in runs.
// Temp layout: hist[8][256] u32 | ticket[8] u32 (padded) | status[passes][tiles][256] u32.
// Status word: top 2 bits flag (1 = tile aggregate, 2 = inclusive prefix), low 30 bits value.
// ------------------------------------------------------------------------------------------------
constexpr int RS_BITS = 8;
constexpr int RS_RADIX = 1 << RS_BITS;
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;  // 4096 keys
constexpr int RS_MAX_PASSES = 8;
constexpr uint32_t RS_FLAG_AGG = 1u << 30, RS_FLAG_INC = 2u << 30, RS_VAL_MASK = (1u << 30) - 1;
constexpr size_t RS_HEADER = (size_t)RS_MAX_PASSES * RS_RADIX * 4 + 256;

size_t sort_temp_bytes(int64_t n) {
  const int64_t tiles = (n + RS_TILE - 1) / RS_TILE;
  const size_t t = (size_t)(tiles > 0 ? tiles : 1);
  const size_t nn = (size_t)(n > 0 ? n : 1);  // sized for 64-bit keys; 32-bit keys use less
  return RS_HEADER + align_up((size_t)RS_MAX_PASSES * t * RS_RADIX * 4, 256) + align_up(nn * 8, 256) + align_up(nn * 4, 256);
}

template <typename KeyT>
__global__ void __launch_bounds__(256)
rs_histogram_kernel(long long n, const KeyT* __restrict__ keys, int end_bit, int passes,
                    uint32_t* __restrict__ ghist) {
  __shared__ uint32_t s_hist[RS_MAX_PASSES * RS_RADIX];
  for (int i = threadIdx.x; i < passes * RS_RADIX; i += 256) s_hist[i] = 0;
  __syncthreads();
  const long long stride = (long long)gridDim.x * 256;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
    const KeyT k = keys[i];
    for (int p = 0; p < passes; p++) {
      const int shift = p * RS_BITS;
      const int nb = min(RS_BITS, end_bit - shift);
      const uint32_t d = (uint32_t)(k >> shift) & ((1u << nb) - 1u);
      atomicAdd(&s_hist[p * RS_RADIX + d], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * RS_RADIX; i += 256) {
    const uint32_t c = s_hist[i];
    if (c) atomicAdd(&ghist[i], c);
  }
}

template <typename KeyT>
struct RsSmem {
  KeyT keys[RS_TILE];
  uint32_t vals[RS_TILE];
  uint32_t warp_cnt[RS_WARPS][RS_RADIX];
  uint32_t digit_start[RS_RADIX];   // exclusive prefix of this tile's digit counts
  uint32_t digit_gbase[RS_RADIX];   // global position of the tile's first key of digit d, minus digit_start[d]
  uint32_t warp_tot[RS_WARPS];
  uint32_t block_cnt[RS_RADIX];     // early digit counts of the tile (published before ranking)
  uint32_t tile;
};

template <typename KeyT>
__global__ void __launch_bounds__(RS_THREADS, 3)
rs_onesweep_kernel(long long n, const KeyT* __restrict__ kin, KeyT* __restrict__ kout,
                   const uint32_t* __restrict__ vin, uint32_t* __restrict__ vout, int shift, int nbits,
                   const uint32_t* __restrict__ ghist, uint32_t* __restrict__ ticket,
                   uint32_t* __restrict__ status) {
  extern __shared__ __align__(16) unsigned char rs_smem_raw[];
  RsSmem<KeyT>& S = *reinterpret_cast<RsSmem<KeyT>*>(rs_smem_raw);
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  if (tid == 0) S.tile = atomicAdd(ticket, 1u);
#pragma unroll
  for (int w = 0; w < RS_WARPS; w++) S.warp_cnt[w][tid] = 0;
  S.block_cnt[tid] = 0;
  __syncthreads();
  const uint32_t tile = S.tile;
  const long long tile_base = (long long)tile * RS_TILE;
  const int n_valid = (int)min((long long)RS_TILE, n - tile_base);
  const uint32_t dmask = (1u << nbits) - 1u;

  // load (warp-striped: item i of lane l is element warp*512 + i*32 + l of the tile)
  KeyT key[RS_ITEMS];
  uint32_t val[RS_ITEMS];
  uint32_t rank[RS_ITEMS];
  const int wbase = warp * (32 * RS_ITEMS);
#pragma unroll
  for (int i = 0; i < RS_ITEMS; i++) {
    const int e = wbase + i * 32 + lane;
    if (e < n_valid) {
      key[i] = kin[tile_base + e];
      val[i] = vin[tile_base + e];
    } else {
      key[i] = (KeyT)~(KeyT)0;
      val[i] = 0;
    }
  }
  // early counts: the tile's digit histogram via shared atomics, published as the tile's aggregate BEFORE
  // the (long) ranking phase, so that by the time this tile looks back its predecessors have published too
#pragma unroll
  for (int i = 0; i < RS_ITEMS; i++) {
    const int e = wbase + i * 32 + lane;
    const uint32_t d = (e < n_valid) ? ((uint32_t)(key[i] >> shift) & dmask) : (RS_RADIX - 1);
    atomicAdd(&S.block_cnt[d], 1u);
  }
  __syncthreads();
  {
    const uint32_t c = S.block_cnt[tid];
    uint32_t* my_status = status + (size_t)tile * RS_RADIX + tid;
    st_relaxed_u32(my_status, (tile == 0 ? RS_FLAG_INC : RS_FLAG_AGG) | c);
  }
  // rank within the warp, in (item, lane) order == memory order => stable.  Peers (lanes holding the same
  // digit) come from 8 ballots, one per digit bit: independent across items, unlike match.any whose
  // latency serialised the loop.
  const unsigned lt = lanemask_lt();
#pragma unroll
  for (int i = 0; i < RS_ITEMS; i++) {
    const int e = wbase + i * 32 + lane;
    const uint32_t d = (e < n_valid) ? ((uint32_t)(key[i] >> shift) & dmask) : (RS_RADIX - 1);
    unsigned peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < RS_BITS; b++) {
      const bool bit = (d >> b) & 1u;
      const unsigned bal = __ballot_sync(0xffffffffu, bit);
      peers &= bit ? bal : ~bal;
    }
    const int leader = __ffs(peers) - 1;
    uint32_t before = 0;
    if ((int)lane == leader) {
      before = S.warp_cnt[warp][d];
      S.warp_cnt[warp][d] = before + __popc(peers);
    }
    before = __shfl_sync(0xffffffffu, before, leader);
    rank[i] = before + __popc(peers & lt);
    __syncwarp();
  }
  __syncthreads();

  // thread d: per-warp exclusive offsets of digit d, tile count, look-back
  {
    const uint32_t d = tid;
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; w++) {
      const uint32_t c = S.warp_cnt[w][d];
      S.warp_cnt[w][d] = run;
      run += c;
    }
    const uint32_t cnt = run;  // tile's count of digit d (the last tile's digit 255 also counts its padding)
    uint32_t* my_status = status + (size_t)tile * RS_RADIX + d;

    // exclusive scans over the 256 digits: tile counts and global histogram
    const uint32_t g = ghist[d];
    uint32_t inc_c = cnt, inc_g = g;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const uint32_t tc = __shfl_up_sync(0xffffffffu, inc_c, s);
      const uint32_t tg = __shfl_up_sync(0xffffffffu, inc_g, s);
      if (lane >= (unsigned)s) { inc_c += tc; inc_g += tg; }
    }
    if (lane == 31) { S.warp_tot[warp] = inc_c; S.digit_start[warp] = inc_g; }  // digit_start reused as scratch
    __syncthreads();
    uint32_t off_c = 0, off_g = 0;
#pragma unroll
    for (int w = 0; w < RS_WARPS; w++)
      if (w < (int)warp) { off_c += S.warp_tot[w]; off_g += S.digit_start[w]; }
    __syncthreads();
    const uint32_t excl_c = off_c + inc_c - cnt;
    const uint32_t excl_g = off_g + inc_g - g;

    uint32_t prefix = 0;
    if (tile > 0) {
      long long p = (long long)tile - 1;
      while (true) {
        const uint32_t* ps = status + (size_t)p * RS_RADIX + d;
        uint32_t s = ld_relaxed_u32(ps);
        while ((s & ~RS_VAL_MASK) == 0) s = ld_relaxed_u32(ps);
        prefix += s & RS_VAL_MASK;
        if (s & RS_FLAG_INC) break;
        p--;
      }
      st_relaxed_u32(my_status, RS_FLAG_INC | ((prefix + cnt) & RS_VAL_MASK));
    }
    S.digit_start[d] = excl_c;
    S.digit_gbase[d] = excl_g + prefix - excl_c;
  }
  __syncthreads();

  // scatter into tile-sorted order in shared memory
#pragma unroll
  for (int i = 0; i < RS_ITEMS; i++) {
    const int e = wbase + i * 32 + lane;
    const uint32_t d = (e < n_valid) ? ((uint32_t)(key[i] >> shift) & dmask) : (RS_RADIX - 1);
    const uint32_t pos = S.digit_start[d] + S.warp_cnt[warp][d] + rank[i];
    S.keys[pos] = key[i];
    S.vals[pos] = val[i];
  }
  __syncthreads();
  // write out in runs; padding sits at the very end of the tile-sorted order
#pragma unroll
  for (int i = 0; i < RS_ITEMS; i++) {
    const int p = i * RS_THREADS + tid;
    if (p < n_valid) {
      const KeyT k = S.keys[p];
      const uint32_t d = (uint32_t)(k >> shift) & dmask;
      const uint32_t dst = S.digit_gbase[d] + (uint32_t)p;
      kout[dst] = k;
      vout[dst] = S.vals[p];
    }
  }
}

template <typename KeyT>
static int sort_pairs_impl(int64_t n, const KeyT* kin, KeyT* kout, const uint32_t* vin, uint32_t* vout,
                           int end_bit, void* temp, cudaStream_t st) {
  if (n <= 0) return GSB_OK;
  if (n >= (1ll << 30) || end_bit <= 0 || end_bit > (int)sizeof(KeyT) * 8) return GSB_EINVAL;
  const int passes = (end_bit + RS_BITS - 1) / RS_BITS;
  const int64_t tiles = (n + RS_TILE - 1) / RS_TILE;
  static bool attr_set = false;
  if (!attr_set) {
    GSB_CUDA_CHECK(cudaFuncSetAttribute(rs_onesweep_kernel<KeyT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)sizeof(RsSmem<KeyT>)));
    GSB_CUDA_CHECK(cudaFuncSetAttribute(rs_onesweep_kernel<KeyT>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                        cudaSharedmemCarveoutMaxShared));
    attr_set = true;
  }
  GSB_CUDA_CHECK(cudaMemsetAsync(temp, 0, RS_HEADER + (size_t)passes * tiles * RS_RADIX * 4, st));
  uint32_t* ghist = reinterpret_cast<uint32_t*>(temp);
  uint32_t* tickets = ghist + RS_MAX_PASSES * RS_RADIX;
  uint32_t* status = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(temp) + RS_HEADER);
  const int hist_blocks = (int)std::min<int64_t>((n + 256 * 16 - 1) / (256 * 16), 148 * 8);
  rs_histogram_kernel<KeyT><<<hist_blocks, 256, 0, st>>>(n, kin, end_bit, passes, ghist);
  GSB_LAUNCH_CHECK();
  // ping-pong between the outputs and an alternate pair inside `temp`, arranged so that the last pass
  // lands in (kout, vout) and the caller's input is never written (like CUB's non-DoubleBuffer API).
  const size_t status_bytes = align_up((size_t)RS_MAX_PASSES * tiles * RS_RADIX * 4, 256);
  KeyT* alt_k = reinterpret_cast<KeyT*>(reinterpret_cast<char*>(temp) + RS_HEADER + status_bytes);
  uint32_t* alt_v = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(alt_k) + align_up((size_t)n * sizeof(KeyT), 256));
  const KeyT* a = kin;
  const uint32_t* va = vin;
  for (int p = 0; p < passes; p++) {
    const int shift = p * RS_BITS;
    const int nb = (end_bit - shift < RS_BITS) ? end_bit - shift : RS_BITS;
    const bool to_out = ((passes - 1 - p) % 2 == 0);
    KeyT* b = to_out ? kout : alt_k;
    uint32_t* vb = to_out ? vout : alt_v;
    rs_onesweep_kernel<KeyT><<<(unsigned)tiles, RS_THREADS, sizeof(RsSmem<KeyT>), st>>>(
        n, a, b, va, vb, shift, nb, ghist + p * RS_RADIX, tickets + p, status + (size_t)p * tiles * RS_RADIX);
    GSB_LAUNCH_CHECK();
    a = b;
    va = vb;
  }
  return GSB_OK;
}

int launch_sort_pairs(int64_t n, const uint64_t* kin, uint64_t* kout, const uint32_t* vin, uint32_t* vout,
                      int end_bit, void* temp, cudaStream_t st) {
  return sort_pairs_impl<unsigned long long>(n, reinterpret_cast<const unsigned long long*>(kin),
                                             reinterpret_cast<unsigned long long*>(kout), vin, vout, end_bit, temp, st);
}
}  // namespace gsb
