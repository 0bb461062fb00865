// segsort.cu -- segmented onesweep LSD radix sort with wide digits, and tile ranges from its histogram.
//
// Replaces cub::DeviceRadixSort::SortPairs (cr/rasterizer_impl.cu:304-309) on the product path.  The
// reference sorts R instances on 64-bit (tile << 32 | depth) keys in 6 eight-bit passes.  A stable LSD
// sort is digit-separable, so here
//   * the depth digits are sorted on the n_cams*P GAUSSIANS, one segment per camera (32-bit keys),
//   * instances are emitted in that order (duplicate_with_keys) with their camera-local tile id as a
//     16-bit key, and
//   * ONE pass with a 10-bit digit (<= 1024 tiles per camera; two passes beyond) puts them in tile order,
//     again one segment per camera.  The sorted tile ids are never written in the one-pass case: the
//     per-camera digit histogram IS the per-tile instance count, so the tile ranges
//     (identifyTileRanges, cr/rasterizer_impl.cu:116-138) are its exclusive scan.
// Results are integer and bit-identical to the reference's point_list / ranges by construction.
//
// Kernel structure (per pass): a tile of THREADS x ITEMS keys ranks its keys per warp (ballot-based digit
// matching, rank order = memory order => stable), publishes its digit counts EARLY (before ranking),
// resolves its global offsets by decoupled look-back over the earlier tiles of its segment (tile ids
// come from an atomic ticket, so a tile only waits on tiles that already started), scatters into
// tile-sorted order in shared memory and writes out in runs.
#include <algorithm>
#include <cstdlib>
#include <type_traits>
#include "common.cuh"

namespace gsb {

constexpr uint32_t SS_FLAG_AGG = 1u << 30, SS_FLAG_INC = 2u << 30, SS_VAL_MASK = (1u << 30) - 1;
constexpr int SS_THREADS = 256;
constexpr int SS_ITEMS = 16;
constexpr int SS_TILE = SS_THREADS * SS_ITEMS;
constexpr int SS_MAX_PASSES = 4;
constexpr size_t SS_HEADER = 256;  // tickets

__device__ __forceinline__ uint32_t ss_ld_relaxed(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ss_ld_relaxed_v4(const uint32_t* p) {
  uint4 v;
  asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void ss_st_relaxed(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ------------------------------------------------------------------------------------------------
// Segment tables.  seg_start[n_seg + 1]: element range of each segment; seg_tile0[n_seg + 1]: exclusive
// prefix of the segments' sort-tile counts.  mode 0: seg_start is given; 1: fixed-length segments;
// 2: from the inclusive scan of tiles_touched taken camera-major (segment c ends at offsets[(c+1)*len-1]).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
seg_setup_kernel(int n_seg, int mode, int len, const uint32_t* __restrict__ offsets, int tile,
                 uint32_t* __restrict__ seg_start, uint32_t* __restrict__ seg_tile0) {
  __shared__ uint32_t s_tiles[1024];
  for (int s = threadIdx.x; s <= n_seg; s += blockDim.x) {
    if (mode == 1) seg_start[s] = (uint32_t)s * (uint32_t)len;
    else if (mode == 2) seg_start[s] = (s == 0) ? 0u : offsets[(size_t)s * len - 1];
  }
  __syncthreads();
  for (int s = threadIdx.x; s < n_seg; s += blockDim.x)
    s_tiles[s] = (seg_start[s + 1] - seg_start[s] + (uint32_t)tile - 1u) / (uint32_t)tile;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    for (int s = 0; s < n_seg; s++) { seg_tile0[s] = run; run += s_tiles[s]; }
    seg_tile0[n_seg] = run;
  }
}

// ------------------------------------------------------------------------------------------------
// Per-segment digit histograms of every pass in one read of the keys.  hist[pass][seg][NB].
// ------------------------------------------------------------------------------------------------
template <typename KeyT, int LOG_NB>
__global__ void __launch_bounds__(256)
seg_hist_kernel(const uint32_t* __restrict__ seg_start, int n_seg, const KeyT* __restrict__ keys, int passes,
                int end_bit, uint32_t* __restrict__ hist) {
  constexpr int NB = 1 << LOG_NB;
  constexpr int G = 16 / (int)sizeof(KeyT);
  extern __shared__ uint32_t s_hist[];  // [passes][NB]
  const int seg = blockIdx.y;
  const uint32_t b = seg_start[seg], e = seg_start[seg + 1];
  for (int i = threadIdx.x; i < passes * NB; i += 256) s_hist[i] = 0;
  __syncthreads();
  const bool aligned = (reinterpret_cast<uintptr_t>(keys) & 15u) == 0;
  const uint32_t g0 = b / G, g1 = (e + G - 1) / G;
  for (uint32_t g = g0 + blockIdx.x * 256u + threadIdx.x; g < g1; g += gridDim.x * 256u) {
    const uint32_t i0 = g * G;
    KeyT k[G];
    if (aligned && i0 >= b && i0 + G <= e) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(keys) + g);
      memcpy(k, &v, 16);
    } else {
#pragma unroll
      for (int j = 0; j < G; j++) k[j] = (i0 + j >= b && i0 + j < e) ? keys[i0 + j] : (KeyT)0;
    }
#pragma unroll
    for (int j = 0; j < G; j++) {
      if (i0 + j < b || i0 + j >= e) continue;
      for (int p = 0; p < passes; p++) {
        const int shift = p * LOG_NB;
        const int nb = min(LOG_NB, end_bit - shift);
        const uint32_t d = (uint32_t)(k[j] >> shift) & ((1u << nb) - 1u);
        atomicAdd(&s_hist[p * NB + d], 1u);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * NB; i += 256) {
    const uint32_t c = s_hist[i];
    const int p = i / NB, d = i - p * NB;
    if (c) atomicAdd(&hist[((size_t)p * n_seg + seg) * NB + d], c);
  }
}

// ------------------------------------------------------------------------------------------------
// One onesweep pass.  Two ranking schemes:
//   OPT = false  ballot ranking: peers (lanes with the same digit) from one ballot per digit bit; the
//                tile's digit counts are taken EARLY with shared atomics and published as its aggregate
//                before the (long) ranking phase.  Right for digits with many equal keys per warp (depth
//                bits).
//   OPT = true   optimistic ranking: every lane takes its slot with a returning shared atomic on the warp's
//                digit counter; a warp in which two lanes met on one counter (detected by re-reading it)
//                redoes that item with the ballot scheme, so the result is still in lane order (stable).
//                Right for wide digits with few equal keys per warp (camera-local tile ids: 32
//                consecutive instances rarely share a tile).  Ranking is short, so the aggregate is
//                published after it.
// ------------------------------------------------------------------------------------------------
template <typename KeyT, int LOG_NB, bool OPT>
struct SsLayout {
  static constexpr int NB = 1 << LOG_NB;
  static constexpr int WARPS = SS_THREADS / 32;
  static constexpr int WCNT_BYTES = OPT ? 4 : 2;
  static constexpr size_t vals = 0;                                        // u32 [TILE]
  static constexpr size_t keys = vals + 4 * (size_t)SS_TILE;               // KeyT [TILE]
  static constexpr size_t wcnt = keys + sizeof(KeyT) * (size_t)SS_TILE;    // u16|u32 [WARPS][NB]
  static constexpr size_t bcnt = wcnt + WCNT_BYTES * (size_t)WARPS * NB;   // u32 [NB]: tile counts, then digit starts
  static constexpr size_t gbase = bcnt + 4 * (size_t)NB;                   // u32 [NB]
  static constexpr size_t misc = gbase + 4 * (size_t)NB;                   // u32 [4 + 2 WARPS]
  static constexpr size_t bytes = misc + 4 * (size_t)(4 + 2 * WARPS);
};

template <int LOG_NB>
__device__ __forceinline__ unsigned ss_peers(uint32_t d) {
  unsigned peers = 0xffffffffu;
#pragma unroll
  for (int b = 0; b < LOG_NB; b++) {
    const bool bit = (d >> b) & 1u;
    const unsigned bal = __ballot_sync(0xffffffffu, bit);
    peers &= bit ? bal : ~bal;
  }
  return peers;
}

// Body of one tile.  FULL: all SS_TILE slots hold keys (no bounds checks in the unrolled loops).
// For 16-bit keys the key and its rank share one register (key low, rank high).
#ifdef SS_PROFILE   // keep the body out of line so that -lineinfo attributes its instructions to its own lines
#define SS_BODY_ATTR __noinline__
#else
#define SS_BODY_ATTR __forceinline__
#endif
template <typename KeyT, int LOG_NB, bool OPT, bool FULL>
__device__ SS_BODY_ATTR void ss_tile_body(unsigned char* ss_smem, const uint32_t tile, const int seg,
                                             const uint32_t local_tile, const uint32_t sbeg, const uint32_t tbase,
                                             const int n_valid, const KeyT* __restrict__ kin, KeyT* __restrict__ kout,
                                             const uint32_t* __restrict__ vin, uint32_t* __restrict__ vout,
                                             const int shift, const uint32_t dmask,
                                             const uint32_t* __restrict__ seg_hist, uint32_t* __restrict__ status) {
  using Lay = SsLayout<KeyT, LOG_NB, OPT>;
  using CntT = typename std::conditional<OPT, uint32_t, uint16_t>::type;
  constexpr int NB = Lay::NB, WARPS = Lay::WARPS;
  constexpr int D = NB >= SS_THREADS ? NB / SS_THREADS : 1;  // digits per thread in the digit phase
  constexpr bool PACK = sizeof(KeyT) == 2;
  uint32_t* s_vals = reinterpret_cast<uint32_t*>(ss_smem + Lay::vals);
  KeyT* s_keys = reinterpret_cast<KeyT*>(ss_smem + Lay::keys);
  CntT* s_wcnt = reinterpret_cast<CntT*>(ss_smem + Lay::wcnt);
  uint32_t* s_bcnt = reinterpret_cast<uint32_t*>(ss_smem + Lay::bcnt);
  uint32_t* s_gbase = reinterpret_cast<uint32_t*>(ss_smem + Lay::gbase);
  uint32_t* s_misc = reinterpret_cast<uint32_t*>(ss_smem + Lay::misc);
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const unsigned full = 0xffffffffu;
  uint32_t* my_status = status + (size_t)tile * NB;

  // load (warp-striped: item i of lane l is element warp*512 + i*32 + l of the tile)
  uint32_t key[SS_ITEMS];   // zero-extended key (16-bit keys: rank in the high half later)
  uint32_t val[SS_ITEMS];
  uint32_t rank[PACK ? 1 : SS_ITEMS];
  const int e0 = warp * (32 * SS_ITEMS) + (int)lane;
  {
    const KeyT* kp = kin + tbase + e0;
#pragma unroll
    for (int i = 0; i < SS_ITEMS; i++)
      key[i] = (FULL || e0 + i * 32 < n_valid) ? (uint32_t)kp[i * 32] : (uint32_t)(KeyT)~(KeyT)0;
    if (vin) {
      const uint32_t* vp = vin + tbase + e0;
#pragma unroll
      for (int i = 0; i < SS_ITEMS; i++) val[i] = (FULL || e0 + i * 32 < n_valid) ? vp[i * 32] : 0u;
    } else {  // no payload given: the element's index in its segment
#pragma unroll
      for (int i = 0; i < SS_ITEMS; i++) val[i] = local_tile * SS_TILE + (uint32_t)(e0 + i * 32);
    }
  }
  const unsigned lt = lanemask_lt();
  CntT* my_wcnt = s_wcnt + warp * NB;

  if (!OPT) {
    // early counts, published as the tile's aggregate BEFORE the (long) ranking phase, so that by the time
    // this tile looks back its predecessors have published too
#pragma unroll
    for (int i = 0; i < SS_ITEMS; i++)
      if (FULL || e0 + i * 32 < n_valid) atomicAdd(&s_bcnt[(key[i] >> shift) & dmask], 1u);
    __syncthreads();
    for (int d = tid; d < NB; d += SS_THREADS)
      ss_st_relaxed(my_status + d, (local_tile == 0 ? SS_FLAG_INC : SS_FLAG_AGG) | s_bcnt[d]);
    // rank within the warp, in (item, lane) order == memory order => stable.  Padding takes the last digit
    // and, being last in memory order, ranks behind every valid key of that digit.
#pragma unroll
    for (int i = 0; i < SS_ITEMS; i++) {
      const uint32_t d = (FULL || e0 + i * 32 < n_valid) ? ((key[i] >> shift) & dmask) : (uint32_t)(NB - 1);
      const unsigned peers = ss_peers<LOG_NB>(d);
      const int leader = __ffs(peers) - 1;
      uint32_t before = 0;
      if ((int)lane == leader) {
        before = my_wcnt[d];
        my_wcnt[d] = (CntT)(before + __popc(peers));
      }
      before = __shfl_sync(full, before, leader);
      const uint32_t r = before + __popc(peers & lt);
      if (PACK) key[i] |= r << 16; else rank[PACK ? 0 : i] = r;
      __syncwarp();
    }
  } else {
#pragma unroll
    for (int i = 0; i < SS_ITEMS; i++) {
      const uint32_t d = (FULL || e0 + i * 32 < n_valid) ? ((key[i] >> shift) & dmask) : (uint32_t)(NB - 1);
      uint32_t r = atomicAdd(reinterpret_cast<uint32_t*>(&my_wcnt[d]), 1u);
      __syncwarp();
      const uint32_t now = my_wcnt[d];
      if (__any_sync(full, now != r + 1u)) {   // two lanes met on one counter: redo this item in lane order
        const unsigned peers = ss_peers<LOG_NB>(d);
        r = now - __popc(peers) + __popc(peers & lt);
      }
      __syncwarp();
      if (PACK) key[i] |= r << 16; else rank[PACK ? 0 : i] = r;
    }
  }
  __syncthreads();

  // digit phase: thread t owns digits [t*D, t*D + D)
  {
    const bool act = (int)(tid * D) < NB;
    uint32_t gh[D], excl_c[D], excl_g[D], bc[D];
    uint32_t tc = 0, tg = 0;
#pragma unroll
    for (int k = 0; k < D; k++) {
      gh[k] = 0; excl_c[k] = 0; excl_g[k] = 0; bc[k] = 0;
      if (act) {
        const int d = tid * D + k;
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < WARPS; w++) {
          const uint32_t c = s_wcnt[w * NB + d];
          s_wcnt[w * NB + d] = (CntT)run;
          run += c;
        }
        gh[k] = seg_hist[(size_t)seg * NB + d];
        bc[k] = OPT ? run : s_bcnt[d];   // OPT: the last tile of a segment counts its padding in the last digit;
                                          // nothing looks back at it (the next segment starts a new chain)
        excl_c[k] = tc; excl_g[k] = tg;
        tc += run; tg += gh[k];
      }
    }
    if (OPT && act) {
#pragma unroll
      for (int k = 0; k < D; k++)
        ss_st_relaxed(my_status + tid * D + k, (local_tile == 0 ? SS_FLAG_INC : SS_FLAG_AGG) | bc[k]);
    }
    uint32_t inc_c = tc, inc_g = tg;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const uint32_t a = __shfl_up_sync(full, inc_c, s);
      const uint32_t b = __shfl_up_sync(full, inc_g, s);
      if (lane >= (unsigned)s) { inc_c += a; inc_g += b; }
    }
    if (lane == 31) { s_misc[4 + warp] = inc_c; s_misc[4 + WARPS + warp] = inc_g; }
    __syncthreads();
    uint32_t off_c = inc_c - tc, off_g = inc_g - tg;
#pragma unroll
    for (int w = 0; w < WARPS; w++)
      if (w < (int)warp) { off_c += s_misc[4 + w]; off_g += s_misc[4 + WARPS + w]; }

    uint32_t prefix[D];
#pragma unroll
    for (int k = 0; k < D; k++) prefix[k] = 0;
    if (act && local_tile > 0) {
      // decoupled look-back, the thread's D digit chains in lock step over the same predecessor tile
      uint32_t p = tile - 1;
      unsigned pend = (1u << D) - 1u;
      while (pend) {
        const uint32_t* ps = status + (size_t)p * NB + tid * D;
        uint32_t s[D];
        if (D % 4 == 0) {
#pragma unroll
          for (int q = 0; q < D / 4; q++) {
            const uint4 v = ss_ld_relaxed_v4(ps + 4 * q);
            s[4 * q] = v.x; s[4 * q + 1] = v.y; s[4 * q + 2] = v.z; s[4 * q + 3] = v.w;
          }
        } else {
#pragma unroll
          for (int k = 0; k < D; k++) s[k] = ss_ld_relaxed(ps + k);
        }
        bool ready = true;
#pragma unroll
        for (int k = 0; k < D; k++)
          if (((pend >> k) & 1u) && (s[k] & ~SS_VAL_MASK) == 0) ready = false;
        if (!ready) continue;
#pragma unroll
        for (int k = 0; k < D; k++) {
          if ((pend >> k) & 1u) {
            prefix[k] += s[k] & SS_VAL_MASK;
            if (s[k] & SS_FLAG_INC) pend &= ~(1u << k);
          }
        }
        p--;
      }
#pragma unroll
      for (int k = 0; k < D; k++)
        ss_st_relaxed(my_status + tid * D + k, SS_FLAG_INC | ((prefix[k] + bc[k]) & SS_VAL_MASK));
    }
    if (act) {
#pragma unroll
      for (int k = 0; k < D; k++) {
        const int d = tid * D + k;
        const uint32_t dstart = off_c + excl_c[k];
        s_bcnt[d] = dstart;                                             // start of digit d in tile-sorted order
        s_gbase[d] = sbeg + off_g + excl_g[k] + prefix[k] - dstart;     // + position in the tile = global slot
      }
    }
  }
  __syncthreads();

  // scatter into tile-sorted order in shared memory
#pragma unroll
  for (int i = 0; i < SS_ITEMS; i++) {
    const uint32_t k = PACK ? (key[i] & 0xFFFFu) : key[i];
    const uint32_t r = PACK ? (key[i] >> 16) : rank[PACK ? 0 : i];
    const uint32_t d = (FULL || e0 + i * 32 < n_valid) ? ((k >> shift) & dmask) : (uint32_t)(NB - 1);
    const uint32_t pos = s_bcnt[d] + my_wcnt[d] + r;
    s_keys[pos] = (KeyT)k;
    s_vals[pos] = val[i];
  }
  __syncthreads();
  // write out in runs; padding sits at the very end of the tile-sorted order
#pragma unroll
  for (int i = 0; i < SS_ITEMS; i++) {
    const int p = i * SS_THREADS + (int)tid;
    if (FULL || p < n_valid) {
      const KeyT k = s_keys[p];
      const uint32_t d = ((uint32_t)k >> shift) & dmask;
      const uint32_t dst = s_gbase[d] + (uint32_t)p;
      if (kout) kout[dst] = k;
      vout[dst] = s_vals[p];
    }
  }
}

#ifndef SS_MINB16
#define SS_MINB16 3
#endif
#ifndef SS_MINB32
#define SS_MINB32 3
#endif
template <typename KeyT, int LOG_NB, bool OPT>
__global__ void __launch_bounds__(SS_THREADS, sizeof(KeyT) > 2 ? SS_MINB32 : SS_MINB16)
seg_onesweep_kernel(const uint32_t* __restrict__ seg_start, const uint32_t* __restrict__ seg_tile0, int n_seg,
                    const KeyT* __restrict__ kin, KeyT* __restrict__ kout, const uint32_t* __restrict__ vin,
                    uint32_t* __restrict__ vout, int shift, uint32_t dmask, const uint32_t* __restrict__ seg_hist,
                    uint32_t* __restrict__ ticket, uint32_t* __restrict__ status) {
  using Lay = SsLayout<KeyT, LOG_NB, OPT>;
  constexpr int NB = Lay::NB, WARPS = Lay::WARPS;
  extern __shared__ __align__(16) unsigned char ss_smem[];
  uint32_t* s_bcnt = reinterpret_cast<uint32_t*>(ss_smem + Lay::bcnt);
  uint32_t* s_misc = reinterpret_cast<uint32_t*>(ss_smem + Lay::misc);
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const unsigned full = 0xffffffffu;

  if (warp == 0) {
    uint32_t t = 0;
    if (lane == 0) t = atomicAdd(ticket, 1u);
    t = __shfl_sync(full, t, 0);
    // segment of sort tile t = #{ e in [1, n_seg] : seg_tile0[e] <= t }  (n_seg: past the last tile)
    int seg = 0;
    for (int e0 = 1; e0 <= n_seg; e0 += 32) {
      const int e = e0 + (int)lane;
      const bool le = (e <= n_seg) && (seg_tile0[e] <= t);
      seg += __popc(__ballot_sync(full, le));
    }
    if (lane == 0) { s_misc[0] = t; s_misc[1] = (uint32_t)seg; }
  }
  {
    uint4* z = reinterpret_cast<uint4*>(ss_smem + Lay::wcnt);
    for (int i = tid; i < (Lay::WCNT_BYTES * WARPS * NB) / 16; i += SS_THREADS) z[i] = make_uint4(0, 0, 0, 0);
    if (!OPT)
      for (int i = tid; i < NB; i += SS_THREADS) s_bcnt[i] = 0;
  }
  __syncthreads();
  const uint32_t tile = s_misc[0];
  const int seg = (int)s_misc[1];
  if (seg >= n_seg) return;  // the grid is an upper bound on the tile count
  const uint32_t local_tile = tile - seg_tile0[seg];
  const uint32_t sbeg = seg_start[seg], send = seg_start[seg + 1];
  const uint32_t tbase = sbeg + local_tile * SS_TILE;
  const int n_valid = (int)min((uint32_t)SS_TILE, send - tbase);
  if (n_valid == SS_TILE)
    ss_tile_body<KeyT, LOG_NB, OPT, true>(ss_smem, tile, seg, local_tile, sbeg, tbase, n_valid, kin, kout, vin, vout,
                                          shift, dmask, seg_hist, status);
  else
    ss_tile_body<KeyT, LOG_NB, OPT, false>(ss_smem, tile, seg, local_tile, sbeg, tbase, n_valid, kin, kout, vin, vout,
                                           shift, dmask, seg_hist, status);
}

// ------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------
static inline int64_t ss_max_tiles(int64_t n_upper, int n_seg) { return (n_upper + SS_TILE - 1) / SS_TILE + n_seg; }

size_t seg_sort_temp_bytes(int64_t n_upper, int n_seg, int passes, int log_nb) {
  const size_t nb = (size_t)1 << log_nb;
  return SS_HEADER + align_up((size_t)passes * n_seg * nb * 4, 256) +
         align_up((size_t)passes * (size_t)ss_max_tiles(n_upper, n_seg) * nb * 4, 256);
}

int launch_seg_setup(int n_seg, int mode, int len, const uint32_t* offsets, uint32_t* seg_start, uint32_t* seg_tile0,
                     cudaStream_t st) {
  if (n_seg <= 0 || n_seg > 1024) return GSB_EINVAL;
  seg_setup_kernel<<<1, 256, 0, st>>>(n_seg, mode, len, offsets, SS_TILE, seg_start, seg_tile0);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

template <typename KeyT, int LOG_NB, bool OPT>
static int seg_sort_impl(int64_t n_upper, int n_seg, const uint32_t* seg_start, const uint32_t* seg_tile0,
                         const KeyT* kin, KeyT* k_out, KeyT* k_alt, const uint32_t* vin, uint32_t* v_out,
                         uint32_t* v_alt, int end_bit, bool want_keys, void* temp, cudaStream_t st,
                         const uint32_t** hist_out) {
  using Lay = SsLayout<KeyT, LOG_NB, OPT>;
  constexpr int NB = Lay::NB;
  const int passes = (end_bit + LOG_NB - 1) / LOG_NB;
  if (passes < 1 || passes > SS_MAX_PASSES || n_upper >= (1ll << 30)) return GSB_EINVAL;
  if (passes > 1 && (!v_alt || !k_alt || !k_out)) return GSB_EINVAL;
  if (want_keys && !k_out) return GSB_EINVAL;
  static bool attr_set = false;
  if (!attr_set) {
    GSB_CUDA_CHECK(cudaFuncSetAttribute(seg_onesweep_kernel<KeyT, LOG_NB, OPT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)Lay::bytes));
    GSB_CUDA_CHECK(cudaFuncSetAttribute(seg_onesweep_kernel<KeyT, LOG_NB, OPT>,
                                        cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    attr_set = true;
  }
  const int64_t max_tiles = ss_max_tiles(n_upper, n_seg);
  const size_t hist_bytes = align_up((size_t)passes * n_seg * NB * 4, 256);
  GSB_CUDA_CHECK(cudaMemsetAsync(temp, 0, SS_HEADER + hist_bytes + (size_t)passes * max_tiles * NB * 4, st));
  uint32_t* tickets = reinterpret_cast<uint32_t*>(temp);
  uint32_t* hist = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(temp) + SS_HEADER);
  uint32_t* status = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(temp) + SS_HEADER + hist_bytes);
  if (hist_out) *hist_out = hist;
  {
    // enough blocks per segment to fill the GPU, few enough that the flush atomics stay cheap
    const int per_seg = std::max(1, std::min<int>((int)((n_upper / std::max(n_seg, 1) + 16383) / 16384), (148 * 8 + n_seg - 1) / n_seg));
    dim3 grid(per_seg, n_seg);
    seg_hist_kernel<KeyT, LOG_NB><<<grid, 256, (size_t)passes * NB * 4, st>>>(seg_start, n_seg, kin, passes, end_bit, hist);
    GSB_LAUNCH_CHECK();
  }
  const KeyT* a = kin;
  const uint32_t* va = vin;
  for (int p = 0; p < passes; p++) {
    const int shift = p * LOG_NB;
    const int nb = std::min(LOG_NB, end_bit - shift);
    const bool to_out = ((passes - 1 - p) % 2 == 0);
    const bool last = (p == passes - 1);
    KeyT* b = (last && !want_keys) ? nullptr : (to_out ? k_out : k_alt);
    uint32_t* vb = to_out ? v_out : v_alt;
    seg_onesweep_kernel<KeyT, LOG_NB, OPT><<<(unsigned)max_tiles, SS_THREADS, Lay::bytes, st>>>(
        seg_start, seg_tile0, n_seg, a, b, va, vb, shift, (1u << nb) - 1u, hist + (size_t)p * n_seg * NB, tickets + p,
        status + (size_t)p * max_tiles * NB);
    GSB_LAUNCH_CHECK();
    a = b;
    va = vb;
  }
  return GSB_OK;
}

int launch_seg_sort_u16(int64_t n_upper, int n_seg, const uint32_t* seg_start, const uint32_t* seg_tile0,
                        const uint16_t* kin, uint16_t* k_out, uint16_t* k_alt, const uint32_t* vin, uint32_t* v_out,
                        uint32_t* v_alt, int end_bit, int log_nb, bool want_keys, void* temp, cudaStream_t st,
                        const uint32_t** hist_out) {
#define GSB_SS_CASE(B)                                                                                              \
  case B:                                                                                                           \
    return opt ? seg_sort_impl<uint16_t, B, true>(n_upper, n_seg, seg_start, seg_tile0, kin, k_out, k_alt, vin, v_out,  \
                                                  v_alt, end_bit, want_keys, temp, st, hist_out)                    \
               : seg_sort_impl<uint16_t, B, false>(n_upper, n_seg, seg_start, seg_tile0, kin, k_out, k_alt, vin, v_out, \
                                                   v_alt, end_bit, want_keys, temp, st, hist_out)
  static const bool opt = getenv("GSB_SS_OPT") ? atoi(getenv("GSB_SS_OPT")) != 0 : true;
  switch (log_nb) {
    GSB_SS_CASE(5);
    GSB_SS_CASE(8);
    GSB_SS_CASE(10);
    default: return GSB_EINVAL;
  }
#undef GSB_SS_CASE
}

int launch_seg_sort_u32(int64_t n_upper, int n_seg, const uint32_t* seg_start, const uint32_t* seg_tile0,
                        const uint32_t* kin, uint32_t* k_out, uint32_t* k_alt, const uint32_t* vin, uint32_t* v_out,
                        uint32_t* v_alt, int end_bit, int log_nb, bool want_keys, void* temp, cudaStream_t st,
                        const uint32_t** hist_out) {
#define GSB_SS_CASE(B)                                                                                              \
  case B:                                                                                                           \
    return seg_sort_impl<uint32_t, B, false>(n_upper, n_seg, seg_start, seg_tile0, kin, k_out, k_alt, vin, v_out, v_alt,   \
                                      end_bit, want_keys, temp, st, hist_out)
  switch (log_nb) {
    GSB_SS_CASE(8);
    GSB_SS_CASE(11);
    default: return GSB_EINVAL;
  }
#undef GSB_SS_CASE
}

// ------------------------------------------------------------------------------------------------
// Tile ranges from the one-pass histogram: hist[cam][tile] is the instance count of the tile, the
// instances of camera `cam` occupy [seg_start[cam], seg_start[cam+1]) in tile order.  Untouched tiles
// keep (0, 0) like the reference's memset (cr/rasterizer_impl.cu:311).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
ranges_from_hist_kernel(int tiles_per_cam, int nb, const uint32_t* __restrict__ seg_start,
                        const uint32_t* __restrict__ hist, uint2* __restrict__ ranges) {
  __shared__ uint32_t s_warp[8];
  const int cam = blockIdx.x;
  const int per = (nb + 255) / 256;
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  const uint32_t* h = hist + (size_t)cam * nb;
  uint32_t tot = 0;
  for (int k = 0; k < per; k++) {
    const int t = threadIdx.x * per + k;
    if (t < tiles_per_cam) tot += h[t];
  }
  uint32_t inc = tot;
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) {
    const uint32_t a = __shfl_up_sync(0xffffffffu, inc, s);
    if (lane >= (unsigned)s) inc += a;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  uint32_t run = seg_start[cam] + inc - tot;
  for (int w = 0; w < 8; w++)
    if (w < (int)warp) run += s_warp[w];
  for (int k = 0; k < per; k++) {
    const int t = threadIdx.x * per + k;
    if (t < tiles_per_cam) {
      const uint32_t c = h[t];
      ranges[(size_t)cam * tiles_per_cam + t] = c ? make_uint2(run, run + c) : make_uint2(0u, 0u);
      run += c;
    }
  }
}

int launch_ranges_from_hist(int n_cams, int tiles_per_cam, int log_nb, const uint32_t* seg_start, const uint32_t* hist,
                            uint32_t* ranges, cudaStream_t st) {
  ranges_from_hist_kernel<<<n_cams, 256, 0, st>>>(tiles_per_cam, 1 << log_nb, seg_start, hist,
                                                  reinterpret_cast<uint2*>(ranges));
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

// identifyTileRanges on sorted camera-local tile ids (multi-pass sorts: > 1024 tiles per camera).
// ranges pre-zeroed by the caller.
__global__ void __launch_bounds__(256)
tile_ranges_seg_kernel(uint32_t R_upper, int n_cams, int tiles_per_cam, const uint32_t* __restrict__ seg_start,
                       const uint16_t* __restrict__ keys, uint2* __restrict__ ranges) {
  const uint32_t i = blockIdx.x * 256u + threadIdx.x;
  const uint32_t R = seg_start[n_cams];
  if (i >= R || i >= R_upper) return;
  int cam = 0;  // camera of instance i: last c with seg_start[c] <= i
  for (int lo = 0, hi = n_cams; lo < hi;) {
    const int mid = (lo + hi + 1) >> 1;
    if (seg_start[mid] <= i) { lo = mid; cam = mid; } else hi = mid - 1;
  }
  const uint32_t cur = keys[i];
  uint2* rg = ranges + (size_t)cam * tiles_per_cam;
  const bool first = (i == seg_start[cam]);
  if (first) rg[cur].x = i;
  else {
    const uint32_t prev = keys[i - 1];
    if (prev != cur) { rg[prev].y = i; rg[cur].x = i; }
  }
  if (i + 1 == seg_start[cam + 1]) rg[cur].y = i + 1;
}

int launch_tile_ranges_seg(int64_t R, int n_cams, int tiles_per_cam, const uint32_t* seg_start, const uint16_t* keys,
                           uint32_t* ranges, cudaStream_t st) {
  GSB_CUDA_CHECK(cudaMemsetAsync(ranges, 0, (size_t)n_cams * tiles_per_cam * 8, st));
  if (R <= 0) return GSB_OK;
  tile_ranges_seg_kernel<<<(unsigned)((R + 255) / 256), 256, 0, st>>>((uint32_t)R, n_cams, tiles_per_cam, seg_start, keys,
                                                                      reinterpret_cast<uint2*>(ranges));
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

}  // namespace gsb
