// segsort.cu -- segmented LSD radix sort with wide digits and two-level offsets; tile ranges from its histogram.
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
// Pass structure.  A first version was a onesweep (chained scan with decoupled look-back).  With ~450
// sort tiles resident on 148 SMs and a 1024-digit status row per tile, the serial look-back (one L2 round
// trip per predecessor, ~17 predecessors on average) was a third of the kernel's instructions and most of
// its stall samples.  A pass is therefore three kernels with no inter-block waiting:
//   K1 seg_tile_hist   per sort tile (4096 keys) a u16 row: the digit counts of the EARLIER tiles of its chunk
//                      (the block walks its chunk's tiles in order, so the row is its running total); per chunk
//                      (16 tiles; 4 in the depth sort) the digit totals;
//   K2 seg_chunk_scan  per segment: exclusive scan of the chunk sums over the chunks and of the digit totals
//                      over the digits (+ the segment's start) -> chunk_base, digit_base, totals;
//   K3 seg_scatter     per sort tile: rank the keys per warp; global slot of digit d = digit_base[d] +
//                      chunk_base[chunk][d] + tile row [d] (three independent loads, one latency) + rank;
//                      scatter through shared memory, write out in runs.
#include <algorithm>
#include <cstdlib>
#include <type_traits>
#include "common.cuh"

namespace gsb {

constexpr int SS_THREADS = 256;
constexpr int SS_ITEMS = 16;
constexpr int SS_TILE = SS_THREADS * SS_ITEMS;   // keys per sort tile
// Sort tiles per chunk (= per block of the histogram kernel).  16 for the instance sort (20 k tiles: plenty of
// blocks, few chunk rows to scan); 4 for the depth sort, whose ~2 k tiles would otherwise be histogrammed by
// ~150 blocks of 8 warps, one per SM, each walking 16 tiles serially (12% of the warp slots, 1.4 TB/s).
template <typename KeyT> struct SsChunk { static constexpr int value = sizeof(KeyT) > 2 ? 4 : 16; };
constexpr int SS_MAX_PASSES = 4;

// ------------------------------------------------------------------------------------------------
// Segment tables.  seg_start[n_seg + 1]: element range of each segment; seg_tile0[n_seg + 1]: exclusive
// prefix of the segments' sort-tile counts, followed by seg_chunk0[n_seg + 1], the same for chunks.
// mode 0: seg_start is given; 1: fixed-length segments; 2: from the inclusive scan of tiles_touched
// taken camera-major (segment c ends at offsets[(c+1)*len-1]).
// Mode 2 also publishes the element total in status[1] and applies the capacity guard of the asynchronous
// forward: a total above `capacity` (>= 0) sets status[0] = GSB_ENOMEM (unless an earlier error is recorded),
// status[2] = the total, and empties every segment, so that nothing downstream touches memory past the arena.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
seg_setup_kernel(int n_seg, int mode, int len, int chunk_tiles, const uint32_t* __restrict__ offsets,
                 uint32_t* __restrict__ seg_start, uint32_t* __restrict__ seg_tile0, long long capacity,
                 int* __restrict__ status) {
  __shared__ uint32_t s_tiles[1024];
  uint32_t* seg_chunk0 = seg_tile0 + n_seg + 1;
  bool overflow = false;
  if (mode == 2) {
    const uint32_t total = offsets[(size_t)n_seg * len - 1];
    overflow = capacity >= 0 && (long long)total > capacity;
    if (threadIdx.x == 0 && status) {
      status[1] = (int)total;
      if (overflow) { if (status[0] == 0) status[0] = GSB_ENOMEM; status[2] = (int)total; }
    }
  }
  for (int s = threadIdx.x; s <= n_seg; s += blockDim.x) {
    if (mode == 1) seg_start[s] = (uint32_t)s * (uint32_t)len;
    else if (mode == 2) seg_start[s] = (s == 0 || overflow) ? 0u : offsets[(size_t)s * len - 1];
  }
  __syncthreads();
  for (int s = threadIdx.x; s < n_seg; s += blockDim.x)
    s_tiles[s] = (seg_start[s + 1] - seg_start[s] + (uint32_t)SS_TILE - 1u) / (uint32_t)SS_TILE;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0, crun = 0;
    for (int s = 0; s < n_seg; s++) {
      seg_tile0[s] = run; seg_chunk0[s] = crun;
      run += s_tiles[s]; crun += (s_tiles[s] + chunk_tiles - 1) / chunk_tiles;
    }
    seg_tile0[n_seg] = run; seg_chunk0[n_seg] = crun;
  }
}

// index of the segment that holds unit t (sort tile or chunk), given the exclusive prefix of the segments'
// unit counts: #{ e in [1, n_seg] : prefix[e] <= t }; n_seg when t is past the last unit.  Whole warp.
__device__ __forceinline__ int ss_find_segment(const uint32_t* __restrict__ prefix, int n_seg, uint32_t t, unsigned lane) {
  int seg = 0;
  for (int e0 = 1; e0 <= n_seg; e0 += 32) {
    const int e = e0 + (int)lane;
    const bool le = (e <= n_seg) && (prefix[e] <= t);
    seg += __popc(__ballot_sync(0xffffffffu, le));
  }
  return seg;
}

// ------------------------------------------------------------------------------------------------
// K1: digit counts per sort tile (u16 rows) and per chunk (u32 rows).  One block per chunk.
// ------------------------------------------------------------------------------------------------
template <typename KeyT, int LOG_NB>
__global__ void __launch_bounds__(SS_THREADS)
seg_tile_hist_kernel(const uint32_t* __restrict__ seg_start, const uint32_t* __restrict__ seg_tile0, int n_seg,
                     const KeyT* __restrict__ keys, int shift, uint32_t dmask, uint16_t* __restrict__ tile_cnt,
                     uint32_t* __restrict__ chunk_tot) {
  constexpr int NB = 1 << LOG_NB;
  constexpr int PER = (NB + SS_THREADS - 1) / SS_THREADS;
  constexpr int SS_CHUNK = SsChunk<KeyT>::value;
  __shared__ uint32_t s_hist[NB];
  __shared__ int s_seg;
  const uint32_t* seg_chunk0 = seg_tile0 + n_seg + 1;
  const unsigned tid = threadIdx.x, lane = tid & 31u;
  const uint32_t chunk = blockIdx.x;
  if (tid < 32) {
    const int sg = ss_find_segment(seg_chunk0, n_seg, chunk, lane);
    if (lane == 0) s_seg = sg;
  }
  __syncthreads();
  const int seg = s_seg;
  if (seg >= n_seg) return;  // the grid is an upper bound on the chunk count
  const uint32_t sbeg = seg_start[seg], send = seg_start[seg + 1];
  const uint32_t first_local = (chunk - seg_chunk0[seg]) * SS_CHUNK;
  const uint32_t seg_tiles = seg_tile0[seg + 1] - seg_tile0[seg];
  const int n_tiles = (int)min((uint32_t)SS_CHUNK, seg_tiles - first_local);
  uint32_t tot[PER];
#pragma unroll
  for (int k = 0; k < PER; k++) tot[k] = 0;
  for (int t = 0; t < n_tiles; t++) {
#pragma unroll
    for (int k = 0; k < PER; k++)
      if (k * SS_THREADS + (int)tid < NB) s_hist[k * SS_THREADS + tid] = 0;
    __syncthreads();
    const uint32_t tbase = sbeg + (first_local + t) * SS_TILE;
    const int n_valid = (int)min((uint32_t)SS_TILE, send - tbase);
    const KeyT* kp = keys + tbase + tid;
    if (n_valid == SS_TILE) {
      uint32_t k[SS_ITEMS];
#pragma unroll
      for (int i = 0; i < SS_ITEMS; i++) k[i] = (uint32_t)kp[i * SS_THREADS];
#pragma unroll
      for (int i = 0; i < SS_ITEMS; i++) atomicAdd(&s_hist[(k[i] >> shift) & dmask], 1u);
    } else {
      for (int i = 0; i < SS_ITEMS; i++)
        if (i * SS_THREADS + (int)tid < n_valid)
          atomicAdd(&s_hist[((uint32_t)kp[i * SS_THREADS] >> shift) & dmask], 1u);
    }
    __syncthreads();
    uint16_t* row = tile_cnt + (size_t)(seg_tile0[seg] + first_local + t) * NB;
#pragma unroll
    for (int k = 0; k < PER; k++) {
      const int d = k * SS_THREADS + (int)tid;
      if (d < NB) {
        const uint32_t c = s_hist[d];
        row[d] = (uint16_t)tot[k];   // keys of digit d in the EARLIER tiles of this chunk (<= 15 * 4096)
        tot[k] += c;
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int k = 0; k < PER; k++) {
    const int d = k * SS_THREADS + (int)tid;
    if (d < NB) chunk_tot[(size_t)chunk * NB + d] = tot[k];
  }
}

// ------------------------------------------------------------------------------------------------
// K2: one block per segment.  chunk_tot -> chunk_base (exclusive over the segment's chunks);
// digit_base[seg][d] = seg_start[seg] + exclusive scan over the digits of the segment's totals;
// totals[seg][d] = the segment's count of digit d.
// ------------------------------------------------------------------------------------------------
template <int LOG_NB>
__global__ void __launch_bounds__(LOG_NB >= 10 ? 1024 : (1 << LOG_NB) < 32 ? 32 : (1 << LOG_NB))
seg_chunk_scan_kernel(const uint32_t* __restrict__ seg_start, const uint32_t* __restrict__ seg_tile0, int n_seg,
                      const uint32_t* __restrict__ chunk_tot, uint32_t* __restrict__ chunk_base,
                      uint32_t* __restrict__ digit_base, uint32_t* __restrict__ totals) {
  constexpr int NB = 1 << LOG_NB;
  constexpr int THREADS = LOG_NB >= 10 ? 1024 : NB < 32 ? 32 : NB;
  constexpr int D = NB / THREADS > 0 ? NB / THREADS : 1;
  __shared__ uint32_t s_warp[32];
  const uint32_t* seg_chunk0 = seg_tile0 + n_seg + 1;
  const int seg = blockIdx.x;
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const bool act = (int)(tid * D) < NB;
  const uint32_t c0 = seg_chunk0[seg], c1 = seg_chunk0[seg + 1];
  uint32_t run[D];
#pragma unroll
  for (int k = 0; k < D; k++) run[k] = 0;
  if (act) {
    // reads and writes go to different arrays so that the loads of the next chunks are in flight while the
    // running sums of this one are stored (an in-place scan serialised one memory round trip per chunk)
    uint32_t c = c0;
    for (; c + 8 <= c1; c += 8) {
      uint32_t v[8][D];
#pragma unroll
      for (int u = 0; u < 8; u++)
#pragma unroll
        for (int k = 0; k < D; k++) v[u][k] = __ldg(chunk_tot + (size_t)(c + u) * NB + tid * D + k);
#pragma unroll
      for (int u = 0; u < 8; u++)
#pragma unroll
        for (int k = 0; k < D; k++) {
          chunk_base[(size_t)(c + u) * NB + tid * D + k] = run[k];
          run[k] += v[u][k];
        }
    }
    for (; c < c1; c++) {
#pragma unroll
      for (int k = 0; k < D; k++) {
        const uint32_t v = __ldg(chunk_tot + (size_t)c * NB + tid * D + k);
        chunk_base[(size_t)c * NB + tid * D + k] = run[k];
        run[k] += v;
      }
    }
  }
  uint32_t tsum = 0, excl[D];
#pragma unroll
  for (int k = 0; k < D; k++) { excl[k] = tsum; tsum += run[k]; }
  uint32_t inc = tsum;
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) {
    const uint32_t a = __shfl_up_sync(0xffffffffu, inc, s);
    if (lane >= (unsigned)s) inc += a;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  uint32_t off = seg_start[seg] + inc - tsum;
  for (int w = 0; w < (int)warp; w++) off += s_warp[w];
  if (act) {
#pragma unroll
    for (int k = 0; k < D; k++) {
      digit_base[(size_t)seg * NB + tid * D + k] = off + excl[k];
      totals[(size_t)seg * NB + tid * D + k] = run[k];
    }
  }
}

// ------------------------------------------------------------------------------------------------
// K3: rank + scatter of one sort tile.  Two ranking schemes:
//   OPT = false  ballot ranking: peers (lanes with the same digit) from one ballot per digit bit.  Right for
//                digits with many equal keys per warp (depth bits).
//   OPT = true   optimistic ranking: every lane takes its slot with a returning shared atomic on the warp's
//                digit counter (16 bits, two per word); a warp in which two lanes met on one counter (detected by re-reading it)
//                redoes that item with the ballot scheme, so the result is still in lane order (stable).
//                Right for wide digits with few equal keys per warp (camera-local tile ids: 32
//                consecutive instances rarely share a tile).
// ------------------------------------------------------------------------------------------------
template <typename KeyT, int LOG_NB, bool OPT>
struct SsLayout {
  static constexpr int NB = 1 << LOG_NB;
  static constexpr int WARPS = SS_THREADS / 32;
  static constexpr int WCNT_BYTES = 2;
  static constexpr size_t vals = 0;                                        // u32 [TILE]
  static constexpr size_t keys = vals + 4 * (size_t)SS_TILE;               // KeyT [TILE]
  static constexpr size_t wcnt = keys + sizeof(KeyT) * (size_t)SS_TILE;    // u16|u32 [WARPS][NB]
  static constexpr size_t dstart = wcnt + WCNT_BYTES * (size_t)WARPS * NB; // u32 [NB]
  static constexpr size_t gbase = dstart + 4 * (size_t)NB;                 // u32 [NB]
  static constexpr size_t misc = gbase + 4 * (size_t)NB;                   // u32 [4 + WARPS]
  static constexpr size_t bytes = misc + 4 * (size_t)(4 + WARPS);
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
// PACKV (one-pass sorts whose payload fits beside the digit in 32 bits, e.g. Gaussian ids < 2^22 with a 10-bit
// digit): digit and payload are staged as ONE shared-memory word, so the scatter into tile order is one random
// store instead of two and the write-out one load instead of two (the kernel is bound by shared-memory wavefronts).
template <typename KeyT, int LOG_NB, bool OPT, bool FULL, bool PACKV>
__device__ SS_BODY_ATTR void ss_tile_body(unsigned char* ss_smem, const uint32_t tile, const int seg,
                                          const uint32_t local_tile, const uint32_t chunk, const uint32_t tbase,
                                          const int n_valid, const KeyT* __restrict__ kin, KeyT* __restrict__ kout,
                                          const uint32_t* __restrict__ vin, uint32_t* __restrict__ vout,
                                          const int shift, const uint32_t dmask,
                                          const uint16_t* __restrict__ tile_cnt,
                                          const uint32_t* __restrict__ chunk_base,
                                          const uint32_t* __restrict__ digit_base) {
  using Lay = SsLayout<KeyT, LOG_NB, OPT>;
  using CntT = uint16_t;
  constexpr int NB = Lay::NB, WARPS = Lay::WARPS;
  constexpr int D = NB >= SS_THREADS ? NB / SS_THREADS : 1;  // digits per thread in the digit phase
  constexpr bool PACK = sizeof(KeyT) == 2;
  uint32_t* s_vals = reinterpret_cast<uint32_t*>(ss_smem + Lay::vals);
  KeyT* s_keys = reinterpret_cast<KeyT*>(ss_smem + Lay::keys);
  CntT* s_wcnt = reinterpret_cast<CntT*>(ss_smem + Lay::wcnt);
  uint32_t* s_dstart = reinterpret_cast<uint32_t*>(ss_smem + Lay::dstart);
  uint32_t* s_gbase = reinterpret_cast<uint32_t*>(ss_smem + Lay::gbase);
  uint32_t* s_misc = reinterpret_cast<uint32_t*>(ss_smem + Lay::misc);
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const unsigned full = 0xffffffffu;

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
  // global slot of the tile's first key of each digit: independent loads, issued before the ranking
  // phase so that their latency hides behind it.  Thread t owns digits [t*D, t*D + D).
  const bool act = (int)(tid * D) < NB;
  uint32_t gb[D];
#pragma unroll
  for (int k = 0; k < D; k++) gb[k] = 0;
  if (act) {
    const int d0 = tid * D;
    if (D == 4) {
      const uint4 a = *reinterpret_cast<const uint4*>(chunk_base + (size_t)chunk * NB + d0);
      const uint4 b = *reinterpret_cast<const uint4*>(digit_base + (size_t)seg * NB + d0);
      const uint2 c = *reinterpret_cast<const uint2*>(tile_cnt + (size_t)tile * NB + d0);
      gb[0] = a.x + b.x + (c.x & 0xFFFFu); gb[1 % D] = a.y + b.y + (c.x >> 16);
      gb[2 % D] = a.z + b.z + (c.y & 0xFFFFu); gb[3 % D] = a.w + b.w + (c.y >> 16);
    } else {
#pragma unroll
      for (int k = 0; k < D; k++)
        gb[k] = chunk_base[(size_t)chunk * NB + d0 + k] + digit_base[(size_t)seg * NB + d0 + k] +
                tile_cnt[(size_t)tile * NB + d0 + k];
    }
  }

  const unsigned lt = lanemask_lt();
  CntT* my_wcnt = s_wcnt + warp * NB;
  if (!OPT) {
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
      // two 16-bit counters per word (a warp holds 512 keys): the atomic adds into the digit's half
      const uint32_t sh = (d & 1u) * 16u;
      uint32_t r = (atomicAdd(reinterpret_cast<uint32_t*>(my_wcnt) + (d >> 1), 1u << sh) >> sh) & 0xFFFFu;
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

  // digit phase: per-warp exclusive offsets of each digit, start of each digit in tile-sorted order
  {
    uint32_t excl_c[D];
    uint32_t tc = 0;
    if (D == 4) {
      // a thread's four 16-bit counters of one warp row are one 8-byte word
      uint32_t run[4] = {0, 0, 0, 0};
      if (act) {
#pragma unroll
        for (int w = 0; w < WARPS; w++) {
          uint2* p = reinterpret_cast<uint2*>(s_wcnt + w * NB + tid * 4);
          const uint2 v = *p;
          *p = make_uint2(run[0] | (run[1] << 16), run[2] | (run[3] << 16));
          run[0] += v.x & 0xFFFFu; run[1] += v.x >> 16; run[2] += v.y & 0xFFFFu; run[3] += v.y >> 16;
        }
      }
#pragma unroll
      for (int k = 0; k < D; k++) { excl_c[k] = tc; tc += run[k % 4]; }
    } else {
#pragma unroll
      for (int k = 0; k < D; k++) {
        excl_c[k] = 0;
        if (act) {
          const int d = tid * D + k;
          uint32_t run = 0;
#pragma unroll
          for (int w = 0; w < WARPS; w++) {
            const uint32_t c = s_wcnt[w * NB + d];
            s_wcnt[w * NB + d] = (CntT)run;
            run += c;
          }
          excl_c[k] = tc;
          tc += run;
        }
      }
    }
    uint32_t inc_c = tc;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const uint32_t a = __shfl_up_sync(full, inc_c, s);
      if (lane >= (unsigned)s) inc_c += a;
    }
    if (lane == 31) s_misc[4 + warp] = inc_c;
    __syncthreads();
    uint32_t off_c = inc_c - tc;
#pragma unroll
    for (int w = 0; w < WARPS; w++)
      if (w < (int)warp) off_c += s_misc[4 + w];
    if (act) {
#pragma unroll
      for (int k = 0; k < D; k++) {
        const int d = tid * D + k;
        const uint32_t ds = off_c + excl_c[k];
        s_dstart[d] = ds;            // start of digit d in tile-sorted order
        s_gbase[d] = gb[k] - ds;     // + position in the tile = global slot
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
    const uint32_t pos = s_dstart[d] + my_wcnt[d] + r;
    if (PACKV) s_vals[pos] = (d << (32 - LOG_NB)) | val[i];
    else { s_keys[pos] = (KeyT)k; s_vals[pos] = val[i]; }
  }
  __syncthreads();
  // write out in runs; padding sits at the very end of the tile-sorted order
#pragma unroll
  for (int i = 0; i < SS_ITEMS; i++) {
    const int p = i * SS_THREADS + (int)tid;
    if (FULL || p < n_valid) {
      if (PACKV) {
        const uint32_t w = s_vals[p];
        vout[s_gbase[w >> (32 - LOG_NB)] + (uint32_t)p] = w & ((1u << (32 - LOG_NB)) - 1u);
      } else {
        const KeyT k = s_keys[p];
        const uint32_t d = ((uint32_t)k >> shift) & dmask;
        const uint32_t dst = s_gbase[d] + (uint32_t)p;
        if (kout) kout[dst] = k;
        vout[dst] = s_vals[p];
      }
    }
  }
}

#ifndef SS_MINB16
#define SS_MINB16 4
#endif
#ifndef SS_MINB32
#define SS_MINB32 3
#endif
template <typename KeyT, int LOG_NB, bool OPT, bool PACKV>
__global__ void __launch_bounds__(SS_THREADS, sizeof(KeyT) > 2 ? SS_MINB32 : SS_MINB16)
seg_scatter_kernel(const uint32_t* __restrict__ seg_start, const uint32_t* __restrict__ seg_tile0, int n_seg,
                   const KeyT* __restrict__ kin, KeyT* __restrict__ kout, const uint32_t* __restrict__ vin,
                   uint32_t* __restrict__ vout, int shift, uint32_t dmask, const uint16_t* __restrict__ tile_cnt,
                   const uint32_t* __restrict__ chunk_base, const uint32_t* __restrict__ digit_base) {
  using Lay = SsLayout<KeyT, LOG_NB, OPT>;
  constexpr int NB = Lay::NB, WARPS = Lay::WARPS;
  extern __shared__ __align__(16) unsigned char ss_smem[];
  uint32_t* s_misc = reinterpret_cast<uint32_t*>(ss_smem + Lay::misc);
  const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const uint32_t tile = blockIdx.x;
  if (warp == 0) {
    const int sg = ss_find_segment(seg_tile0, n_seg, tile, lane);
    if (lane == 0) s_misc[1] = (uint32_t)sg;
  }
  {
    uint4* z = reinterpret_cast<uint4*>(ss_smem + Lay::wcnt);
    for (int i = tid; i < (Lay::WCNT_BYTES * WARPS * NB) / 16; i += SS_THREADS) z[i] = make_uint4(0, 0, 0, 0);
  }
  __syncthreads();
  const int seg = (int)s_misc[1];
  if (seg >= n_seg) return;  // the grid is an upper bound on the tile count
  const uint32_t* seg_chunk0 = seg_tile0 + n_seg + 1;
  const uint32_t local_tile = tile - seg_tile0[seg];
  const uint32_t chunk = seg_chunk0[seg] + local_tile / SsChunk<KeyT>::value;
  const uint32_t sbeg = seg_start[seg], send = seg_start[seg + 1];
  const uint32_t tbase = sbeg + local_tile * SS_TILE;
  const int n_valid = (int)min((uint32_t)SS_TILE, send - tbase);
  if (n_valid == SS_TILE)
    ss_tile_body<KeyT, LOG_NB, OPT, true, PACKV>(ss_smem, tile, seg, local_tile, chunk, tbase, n_valid, kin, kout, vin, vout,
                                          shift, dmask, tile_cnt, chunk_base, digit_base);
  else
    ss_tile_body<KeyT, LOG_NB, OPT, false, PACKV>(ss_smem, tile, seg, local_tile, chunk, tbase, n_valid, kin, kout, vin, vout,
                                           shift, dmask, tile_cnt, chunk_base, digit_base);
}

// ------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------
static inline int64_t ss_max_tiles(int64_t n_upper, int n_seg) { return (n_upper + SS_TILE - 1) / SS_TILE + n_seg; }
static inline int64_t ss_max_chunks(int64_t n_upper, int n_seg, int chunk_tiles) {
  return (n_upper + (int64_t)SS_TILE * chunk_tiles - 1) / ((int64_t)SS_TILE * chunk_tiles) + n_seg;
}

// temp: tile_cnt u16 [max_tiles][NB] | chunk_tot u32 [max_chunks][NB] | chunk_base u32 [max_chunks][NB] |
//       digit_base u32 [n_seg][NB] | totals u32 [n_seg][NB]; reused by every pass.  Nothing needs zeroing.
size_t seg_sort_temp_bytes(int64_t n_upper, int n_seg, int key_bytes, int log_nb) {
  const int chunk_tiles = key_bytes > 2 ? SsChunk<uint32_t>::value : SsChunk<uint16_t>::value;
  const size_t nb = (size_t)1 << log_nb;
  return align_up((size_t)ss_max_tiles(n_upper, n_seg) * nb * 2, 256) +
         2 * align_up((size_t)ss_max_chunks(n_upper, n_seg, chunk_tiles) * nb * 4, 256) + 2 * align_up((size_t)n_seg * nb * 4, 256);
}

int launch_seg_setup(int n_seg, int mode, int len, int key_bytes, const uint32_t* offsets, uint32_t* seg_start,
                     uint32_t* seg_tile0, cudaStream_t st, int64_t capacity, int* status) {
  if (n_seg <= 0 || n_seg > 1024) return GSB_EINVAL;
  const int chunk_tiles = key_bytes > 2 ? SsChunk<uint32_t>::value : SsChunk<uint16_t>::value;
  seg_setup_kernel<<<1, 256, 0, st>>>(n_seg, mode, len, chunk_tiles, offsets, seg_start, seg_tile0, (long long)capacity,
                                      status);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

template <typename KeyT, int LOG_NB, bool OPT>
static int seg_sort_impl(int64_t n_upper, int n_seg, const uint32_t* seg_start, const uint32_t* seg_tile0,
                         const KeyT* kin, KeyT* k_out, KeyT* k_alt, const uint32_t* vin, uint32_t* v_out,
                         uint32_t* v_alt, int end_bit, bool want_keys, void* temp, cudaStream_t st,
                         const uint32_t** hist_out, int val_bits = 32) {
  using Lay = SsLayout<KeyT, LOG_NB, OPT>;
  constexpr int NB = Lay::NB;
  const int passes = (end_bit + LOG_NB - 1) / LOG_NB;
  if (passes < 1 || passes > SS_MAX_PASSES || n_upper >= (1ll << 30)) return GSB_EINVAL;
  if (passes > 1 && (!v_alt || !k_alt || !k_out)) return GSB_EINVAL;
  if (want_keys && !k_out) return GSB_EINVAL;
  constexpr bool CAN_PACK = sizeof(KeyT) == 2 && OPT;
  const bool packv = CAN_PACK && passes == 1 && !want_keys && vin && val_bits + LOG_NB <= 32;
  static bool attr_set = false;
  if (!attr_set) {
    GSB_CUDA_CHECK(cudaFuncSetAttribute(seg_scatter_kernel<KeyT, LOG_NB, OPT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)Lay::bytes));
    GSB_CUDA_CHECK(cudaFuncSetAttribute(seg_scatter_kernel<KeyT, LOG_NB, OPT, false>,
                                        cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    if (CAN_PACK) {
      GSB_CUDA_CHECK(cudaFuncSetAttribute(seg_scatter_kernel<KeyT, LOG_NB, OPT, CAN_PACK>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Lay::bytes));
      GSB_CUDA_CHECK(cudaFuncSetAttribute(seg_scatter_kernel<KeyT, LOG_NB, OPT, CAN_PACK>,
                                          cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    }
    attr_set = true;
  }
  const int64_t max_tiles = ss_max_tiles(n_upper, n_seg);
  const int64_t max_chunks = ss_max_chunks(n_upper, n_seg, SsChunk<KeyT>::value);
  char* t = reinterpret_cast<char*>(temp);
  uint16_t* tile_cnt = reinterpret_cast<uint16_t*>(t);
  t += align_up((size_t)max_tiles * NB * 2, 256);
  uint32_t* chunk_tot = reinterpret_cast<uint32_t*>(t);
  t += align_up((size_t)max_chunks * NB * 4, 256);
  uint32_t* chunk_base = reinterpret_cast<uint32_t*>(t);
  t += align_up((size_t)max_chunks * NB * 4, 256);
  uint32_t* digit_base = reinterpret_cast<uint32_t*>(t);
  t += align_up((size_t)n_seg * NB * 4, 256);
  uint32_t* totals = reinterpret_cast<uint32_t*>(t);
  if (hist_out) *hist_out = totals;   // digit counts per segment of the LAST pass
  const KeyT* a = kin;
  const uint32_t* va = vin;
  for (int p = 0; p < passes; p++) {
    const int shift = p * LOG_NB;
    const int nb = std::min(LOG_NB, end_bit - shift);
    const uint32_t dmask = (1u << nb) - 1u;
    const bool to_out = ((passes - 1 - p) % 2 == 0);
    const bool last = (p == passes - 1);
    KeyT* b = (last && !want_keys) ? nullptr : (to_out ? k_out : k_alt);
    uint32_t* vb = to_out ? v_out : v_alt;
    seg_tile_hist_kernel<KeyT, LOG_NB><<<(unsigned)max_chunks, SS_THREADS, 0, st>>>(seg_start, seg_tile0, n_seg, a, shift,
                                                                                   dmask, tile_cnt, chunk_tot);
    GSB_LAUNCH_CHECK();
    seg_chunk_scan_kernel<LOG_NB><<<n_seg, LOG_NB >= 10 ? 1024 : std::max(NB, 32), 0, st>>>(seg_start, seg_tile0, n_seg, chunk_tot, chunk_base, digit_base,
                                                                                  totals);
    GSB_LAUNCH_CHECK();
    if (packv)
      seg_scatter_kernel<KeyT, LOG_NB, OPT, CAN_PACK><<<(unsigned)max_tiles, SS_THREADS, Lay::bytes, st>>>(
          seg_start, seg_tile0, n_seg, a, b, va, vb, shift, dmask, tile_cnt, chunk_base, digit_base);
    else
      seg_scatter_kernel<KeyT, LOG_NB, OPT, false><<<(unsigned)max_tiles, SS_THREADS, Lay::bytes, st>>>(
          seg_start, seg_tile0, n_seg, a, b, va, vb, shift, dmask, tile_cnt, chunk_base, digit_base);
    GSB_LAUNCH_CHECK();
    a = b;
    va = vb;
  }
  return GSB_OK;
}

int launch_seg_sort_u16(int64_t n_upper, int n_seg, const uint32_t* seg_start, const uint32_t* seg_tile0,
                        const uint16_t* kin, uint16_t* k_out, uint16_t* k_alt, const uint32_t* vin, uint32_t* v_out,
                        uint32_t* v_alt, int end_bit, int log_nb, bool want_keys, void* temp, cudaStream_t st,
                        const uint32_t** hist_out, int val_bits) {
#define GSB_SS_CASE(B)                                                                                              \
  case B:                                                                                                           \
    return opt ? seg_sort_impl<uint16_t, B, true>(n_upper, n_seg, seg_start, seg_tile0, kin, k_out, k_alt, vin, v_out,  \
                                                  v_alt, end_bit, want_keys, temp, st, hist_out, val_bits)          \
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
// ranges pre-zeroed by the caller.  A thread owns 8 consecutive instances (one 16-byte load of their ids + the id in
// front of them); the segment table sits in shared memory, the camera of the thread's first instance comes from one
// binary search there and advances with the instances (the first version searched global memory once per instance
// and read every id twice: 1.05 ms for C4's 249 M instances).
constexpr int TR_ITEMS = 8;
__global__ void __launch_bounds__(256)
tile_ranges_seg_kernel(uint32_t R_upper, int n_cams, int tiles_per_cam, const uint32_t* __restrict__ seg_start,
                       const uint16_t* __restrict__ keys, uint2* __restrict__ ranges) {
  __shared__ uint32_t s_seg[GSB_MAX_CAMS + 2];
  for (int c = threadIdx.x; c <= n_cams; c += 256) s_seg[c] = seg_start[c];
  __syncthreads();
  const uint32_t R = min(s_seg[n_cams], R_upper);
  const uint32_t base = (blockIdx.x * 256u + threadIdx.x) * TR_ITEMS;
  if (base >= R) return;
  uint32_t k[TR_ITEMS];
  if (base + TR_ITEMS <= R) {
    const uint4 v = *reinterpret_cast<const uint4*>(keys + base);
    k[0] = v.x & 0xFFFFu; k[1] = v.x >> 16; k[2] = v.y & 0xFFFFu; k[3] = v.y >> 16;
    k[4] = v.z & 0xFFFFu; k[5] = v.z >> 16; k[6] = v.w & 0xFFFFu; k[7] = v.w >> 16;
  } else {
#pragma unroll
    for (int j = 0; j < TR_ITEMS; j++) k[j] = base + j < R ? keys[base + j] : 0u;
  }
  uint32_t prev = base > 0 ? keys[base - 1] : 0u;
  int cam = 0;  // camera of instance `base`: last c with seg_start[c] <= base
  for (int lo = 0, hi = n_cams; lo < hi;) {
    const int mid = (lo + hi + 1) >> 1;
    if (s_seg[mid] <= base) { lo = mid; cam = mid; } else hi = mid - 1;
  }
#pragma unroll
  for (int j = 0; j < TR_ITEMS; j++) {
    const uint32_t i = base + j;
    if (i >= R) break;
    while (i >= s_seg[cam + 1]) cam++;   // also steps over cameras without instances
    const uint32_t cur = k[j];
    uint2* rg = ranges + (size_t)cam * tiles_per_cam;
    if (i == s_seg[cam]) rg[cur].x = i;
    else if (prev != cur) { rg[prev].y = i; rg[cur].x = i; }
    if (i + 1 == s_seg[cam + 1]) rg[cur].y = i + 1;
    prev = cur;
  }
}

int launch_tile_ranges_seg(int64_t R, int n_cams, int tiles_per_cam, const uint32_t* seg_start, const uint16_t* keys,
                           uint32_t* ranges, cudaStream_t st) {
  GSB_CUDA_CHECK(cudaMemsetAsync(ranges, 0, (size_t)n_cams * tiles_per_cam * 8, st));
  if (R <= 0) return GSB_OK;
  if (n_cams > GSB_MAX_CAMS) return GSB_EINVAL;
  const int64_t per_block = 256 * TR_ITEMS;
  tile_ranges_seg_kernel<<<(unsigned)((R + per_block - 1) / per_block), 256, 0, st>>>((uint32_t)R, n_cams, tiles_per_cam,
                                                                                   seg_start, keys,
                                                                                   reinterpret_cast<uint2*>(ranges));
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

}  // namespace gsb
