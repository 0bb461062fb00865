// gemm.cu -- the dense layers of the DeformationNetwork (train.py:57-110: Linear 192 -> H,
// residual blocks of Linear H -> H, Linear H -> 7 over ALL Gaussians, N ~ 3e5 rows) on the 5th-generation tensor
// cores:  Y[rows, N] = X[rows, K] . W[N, K]^T (+ bias), FP32 in HBM, TF32 multiply / FP32 accumulate.
//
// This is the only GEMM-shaped work in the product.  With K, N <= 256 the layer is HBM-bound (rows x (K + N) x 4 B
// against 2 rows K N flop: 128 flop/B at K = N = 256, below the ~170 flop/B TF32 ridge), so the design goal is to
// stream X once at full bandwidth and keep everything else out of the way:
//   * persistent kernel, one CTA per SM, each looping over 128-row tiles of X;
//   * warp 0 / one thread: TMA producer.  X and W tiles ([128 x 32] and [N x 32] floats, 128-byte rows, hardware
//     128B swizzle) land in a multi-stage shared-memory ring; FP32 data is fed to the tensor core as it is
//     (kind::tf32 reads the upper 19 bits), so there is no conversion pass over the activations;
//   * warp 1 / one thread: tcgen05.mma issuer, 128 x N x 8 instructions, accumulator in TMEM, two accumulator
//     stages (2 x N <= 512 columns) so the epilogue of tile i overlaps the loads and MMAs of tile i + 1;
//   * warps 2-5: epilogue.  tcgen05.ld (lane = row) 32 columns at a time, bias add, swizzled staging in the warp's
//     own shared-memory buffer, coalesced 16-byte global stores (8 lanes per 128-byte row segment).  A TMA store
//     was measured slower here: it queues behind the producer's prefetched loads in the SM's TMA unit.
// W (<= 256 KB) is re-read per tile from L2.  mbarrier pipelines: full/empty per smem stage (TMA <-> MMA),
// full/empty per accumulator stage (MMA <-> epilogue).
//
// gsb_linear_tf32x3 (the same kernel with SPLIT = true) recovers FP32-level accuracy: four extra warps split every
// landed tile in place into hi = x & ~0x1fff and lo = x - hi (lo in a second buffer with the same swizzled
// offsets) and the issuer accumulates Xhi Whi + Xhi Wlo + Xlo Whi.
#include <cuda.h>

#include "common.cuh"

namespace gsb {

constexpr int GM_BLOCK_M = 128;
constexpr int GM_BLOCK_K = 32;                      // floats: one 128-byte swizzle row
constexpr int GM_UMMA_K = 8;                        // tf32: 32 bytes per instruction
constexpr int GM_A_BYTES = GM_BLOCK_M * GM_BLOCK_K * 4;   // 16 KB
constexpr int GM_THREADS = 192;                     // warp 0 TMA, warp 1 MMA, warps 2-5 epilogue
constexpr int GM_SPLIT_WARPS = 4;                   // SPLIT only: warps 6-9 split tiles into hi / lo
constexpr int GM_MAX_STAGES = 8;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, "
      "[%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tc_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand tile with 128-byte rows and the 128B swizzle (what TMA SWIZZLE_128B writes): 8-row groups are
// 1024 bytes apart (SBO), LBO unused, descriptor version 1 (sm_100), layout type 2 = SWIZZLE_128B.
__device__ __forceinline__ uint64_t umma_desc_k_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;                  // leading byte offset (ignored for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;        // stride byte offset
  d |= (uint64_t)1 << 46;                  // version
  d |= (uint64_t)2 << 61;                  // SWIZZLE_128B
  return d;
}

struct GemmParams {
  long long rows;
  int K, N, umma_n, stages, b_bytes;
  const float* bias;
  const float* residual;   // optional [rows][N] tensor added to the output in the epilogue (row pitch ldr floats)
  long long ldr;
  float* Y;
  long long ldy;
  double* col_stats;   // [2][N] column sums and sums of squares of Y over the rows (BatchNorm statistics), or NULL
  int tma_store;   // Y goes out through shared-memory staging as coalesced 16-byte stores (aligned Y, N % 4 == 0)
};
constexpr int GM_EPI_BYTES = 4 * 32 * 128;   // per epilogue warp: one [32 rows x 32 floats] swizzled staging buffer

template <bool SPLIT>
__global__ void __launch_bounds__(GM_THREADS + (SPLIT ? GM_SPLIT_WARPS * 32 : 0), 1)
linear_tf32_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                   const __grid_constant__ CUtensorMap map_wlo, const __grid_constant__ CUtensorMap map_y,
                   GemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[GM_MAX_STAGES], empty_bar[GM_MAX_STAGES], split_bar[GM_MAX_STAGES];
  __shared__ __align__(8) uint64_t acc_full[2], acc_empty[2];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float bias_s[256];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // 1024-byte aligned stage ring: [A | B] (+ [A_lo | B_lo] for SPLIT) per stage
  uint8_t* epi = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* smem = epi + GM_EPI_BYTES;
  const int stage_bytes = (GM_A_BYTES + p.b_bytes) * (SPLIT ? 2 : 1);
  const int k_blocks = (p.K + GM_BLOCK_K - 1) / GM_BLOCK_K;
  const long long m_tiles = (p.rows + GM_BLOCK_M - 1) / GM_BLOCK_M;

  for (int i = threadIdx.x; i < 256; i += blockDim.x) bias_s[i] = (p.bias && i < p.N) ? __ldg(p.bias + i) : 0.f;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; s++) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
      mbar_init(&split_bar[s], GM_SPLIT_WARPS * 32);
    }
    for (int s = 0; s < 2; s++) {
      mbar_init(&acc_full[s], 1);
      mbar_init(&acc_empty[s], 4 * 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {   // TMEM: 2 accumulator stages of umma_n columns (power of two >= 32)
    uint32_t cols = 32;
    while (cols < 2u * p.umma_n) cols <<= 1;
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)),
                 "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (long long t = blockIdx.x; t < m_tiles; t += gridDim.x) {
        for (int kb = 0; kb < k_blocks; kb++) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* a = smem + (size_t)stage * stage_bytes;
          uint8_t* b = a + GM_A_BYTES;
          mbar_expect_tx(&full_bar[stage], GM_A_BYTES + p.b_bytes + (SPLIT ? p.b_bytes : 0));
          tma_load_2d(&map_x, &full_bar[stage], a, kb * GM_BLOCK_K, (int)(t * GM_BLOCK_M));
          tma_load_2d(&map_w, &full_bar[stage], b, kb * GM_BLOCK_K, 0);
          if (SPLIT) tma_load_2d(&map_wlo, &full_bar[stage], b + GM_A_BYTES + p.b_bytes, kb * GM_BLOCK_K, 0);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      // instruction descriptor: D = F32, A = B = TF32, both K-major, N >> 3, M >> 4
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(p.umma_n >> 3) << 17) |
                             ((uint32_t)(GM_BLOCK_M >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (long long t = blockIdx.x; t < m_tiles; t += gridDim.x) {
        mbar_wait(&acc_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d = tmem_base + (uint32_t)(acc * p.umma_n);
        for (int kb = 0; kb < k_blocks; kb++) {
          mbar_wait(SPLIT ? &split_bar[stage] : &full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a = smem_u32(smem + (size_t)stage * stage_bytes);
          const uint32_t b = a + GM_A_BYTES;
          const uint64_t ad = umma_desc_k_sw128(a), bd = umma_desc_k_sw128(b);
#pragma unroll
          for (int k = 0; k < GM_BLOCK_K / GM_UMMA_K; k++) {
            const uint64_t step = (uint64_t)((k * GM_UMMA_K * 4) >> 4);
            tc_mma_tf32(d, ad + step, bd + step, idesc, (kb | k) != 0);
            if (SPLIT) {
              const uint64_t alo = umma_desc_k_sw128(a + GM_A_BYTES + p.b_bytes) + step;
              const uint64_t blo = umma_desc_k_sw128(b + GM_A_BYTES + p.b_bytes) + step;
              tc_mma_tf32(d, ad + step, blo, idesc, 1);
              tc_mma_tf32(d, alo, bd + step, idesc, 1);
            }
          }
          tc_commit(&empty_bar[stage]);   // frees the smem stage when the MMAs above have read it
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        tc_commit(&acc_full[acc]);        // accumulator complete
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp < 6) {
    // ===== epilogue: TMEM -> registers -> (+ bias) -> global =====
    const int q = warp & 3;               // TMEM lane quarter this warp may read
    double st_sum[8], st_sq[8];           // column statistics: lane = column within the 32-column chunk
#pragma unroll
    for (int i = 0; i < 8; i++) st_sum[i] = st_sq[i] = 0.0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (long long t = blockIdx.x; t < m_tiles; t += gridDim.x) {
      mbar_wait(&acc_full[acc], acc_phase);
      tc_fence_after();
      const long long row = t * GM_BLOCK_M + q * 32 + lane;
      if (p.tma_store) {
        // 32 columns at a time: registers -> this warp's swizzled [32 x 128 B] buffer -> one TMA store per chunk
#pragma unroll
        for (int ci = 0; ci < 8; ci++) {
          const int c = ci * 32;
          if (c >= p.umma_n) break;
          uint32_t v[32];
          tc_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * p.umma_n + c), *reinterpret_cast<uint32_t(*)[16]>(v));
          if (c + 16 < p.umma_n)
            tc_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * p.umma_n + c + 16),
                    *reinterpret_cast<uint32_t(*)[16]>(v + 16));
          tc_ld_wait();
          uint8_t* buf = epi + q * 4096;
          __syncwarp();   // the previous chunk's reads of this buffer are done
#pragma unroll
          for (int j = 0; j < 8; j++) {
            float4 o = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]),
                                   __uint_as_float(v[4 * j + 3]));
            const float4 bb = *reinterpret_cast<const float4*>(bias_s + c + 4 * j);   // zeros without a bias
            o.x += bb.x; o.y += bb.y; o.z += bb.z; o.w += bb.w;
            *reinterpret_cast<float4*>(buf + lane * 128 + ((j ^ (lane & 7)) << 4)) = o;
          }
          __syncwarp();
          // coalesced write-out: 8 lanes cover one 128-byte row segment, 4 rows per instruction
          const int j = lane & 7;
          float4 rq[8];
          if (p.residual) {   // all eight loads in flight before the first store (Y may alias the residual element-wise)
#pragma unroll
            for (int i = 0; i < 8; i++) {
              const long long grow = t * GM_BLOCK_M + q * 32 + i * 4 + (lane >> 3);
              rq[i] = (grow < p.rows && c + 4 * j < p.N)
                          ? __ldcs(reinterpret_cast<const float4*>(p.residual + grow * p.ldr + c + 4 * j))
                          : make_float4(0.f, 0.f, 0.f, 0.f);
            }
          }
#pragma unroll
          for (int i = 0; i < 8; i++) {
            const int r = i * 4 + (lane >> 3);
            const long long grow = t * GM_BLOCK_M + q * 32 + r;
            float4 o = *reinterpret_cast<const float4*>(buf + r * 128 + ((j ^ (r & 7)) << 4));
            if (p.residual) { o.x += rq[i].x; o.y += rq[i].y; o.z += rq[i].z; o.w += rq[i].w; }
            if (grow < p.rows && c + 4 * j < p.N) *reinterpret_cast<float4*>(p.Y + grow * p.ldy + c + 4 * j) = o;
          }
          if (p.col_stats) {   // column `lane` of the chunk over this warp's valid rows (conflict-free: 32 consecutive floats)
            const long long left = p.rows - (t * GM_BLOCK_M + q * 32);
            const int nr = left < 32 ? (int)left : 32;
            float s1 = 0.f, s2 = 0.f;
            for (int r = 0; r < nr; r++) {
              const float x = *reinterpret_cast<const float*>(buf + r * 128 + (((lane >> 2) ^ (r & 7)) << 4) + ((lane & 3) << 2));
              s1 += x;
              s2 = fmaf(x, x, s2);
            }
            st_sum[ci] += (double)s1;
            st_sq[ci] += (double)s2;
          }
        }
      } else {
      float* yrow = p.Y + row * p.ldy;
      for (int c = 0; c < p.umma_n; c += 16) {
        uint32_t v[16];
        tc_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * p.umma_n + c), v);
        tc_ld_wait();
        if (row < p.rows) {
#pragma unroll
          for (int i = 0; i < 16; i++)
            if (c + i < p.N) yrow[c + i] = __uint_as_float(v[i]) + bias_s[c + i];
        }
      }
      }
      tc_fence_before();
      mbar_arrive(&acc_empty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (p.col_stats && p.tma_store) {
#pragma unroll
      for (int ci = 0; ci < 8; ci++)
        if (ci * 32 + lane < p.N) {
          atomicAdd(p.col_stats + ci * 32 + lane, st_sum[ci]);
          atomicAdd(p.col_stats + p.N + ci * 32 + lane, st_sq[ci]);
        }
    }
  } else if (SPLIT) {
    // ===== splitter: X tile -> hi (in place) + lo; the W halves arrive pre-split through TMA =====
    const int tid = threadIdx.x - GM_THREADS;
    int stage = 0;
    uint32_t phase = 0;
    for (long long t = blockIdx.x; t < m_tiles; t += gridDim.x) {
      for (int kb = 0; kb < k_blocks; kb++) {
        mbar_wait(&full_bar[stage], phase);
        float4* a = reinterpret_cast<float4*>(smem + (size_t)stage * stage_bytes);
        float4* alo = reinterpret_cast<float4*>(smem + (size_t)stage * stage_bytes + GM_A_BYTES + p.b_bytes);
#pragma unroll
        for (int i = 0; i < GM_A_BYTES / 16 / (GM_SPLIT_WARPS * 32); i++) {
          const int o = tid + i * GM_SPLIT_WARPS * 32;
          const float4 x = a[o];
          float4 h;
          h.x = __uint_as_float(__float_as_uint(x.x) & 0xFFFFE000u);
          h.y = __uint_as_float(__float_as_uint(x.y) & 0xFFFFE000u);
          h.z = __uint_as_float(__float_as_uint(x.z) & 0xFFFFE000u);
          h.w = __uint_as_float(__float_as_uint(x.w) & 0xFFFFE000u);
          a[o] = h;
          alo[o] = make_float4(x.x - h.x, x.y - h.y, x.z - h.z, x.w - h.w);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the MMA
        mbar_arrive(&split_bar[stage]);
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    uint32_t cols = 32;
    while (cols < 2u * p.umma_n) cols <<= 1;
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(cols) : "memory");
  }
}

// ------------------------------------------------------------------------------------------------
// Weight gradient: dW[N, K] = G[rows, N]^T . X[rows, K], the reduction runs over the ROWS (3e5), so both operands are
// "MN-major" for the tensor core: A = G^T is contiguous along M (= n), B = X along N (= k).  For 32-bit MN-major operands
// the only shared-memory layout is the 128-byte swizzle with 32-byte atoms (UMMA layout type SWIZZLE_128B_BASE32B:
// 128-byte rows, the four 32-byte chunks of a row XOR-permuted by row % 4), which is what a TMA box of
// [32 rows x 32 floats] with CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B writes: blocks of 32 floats along M / N are LBO = 4096 B
// apart, groups of 4 reduction rows SBO = 512 B apart, and one kind::tf32 instruction (K = 8) consumes two groups.  Each CTA owns a contiguous range of 32-row blocks, accumulates its
// whole range in TMEM (up to 2 x [128 x 256] FP32 = all 512 columns) and writes ONE partial [N, K]; a second kernel sums
// the partials.  N, K multiples of 32, <= 256.
// ------------------------------------------------------------------------------------------------

template <int BK>
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((BK * 128) >> 4) << 16;      // leading byte offset: next block of 32 floats along M / N
  d |= (uint64_t)(512 >> 4) << 32;             // stride byte offset: next group of 4 reduction rows
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)1 << 61;                      // SWIZZLE_128B_BASE32B: the only MN-major layout for 32-bit operands
  return d;
}

struct WgradParams {
  long long rows;
  int N, K, npad, stages;
  float* partial;   // [gridDim.x][N][K]
};

// SPLIT (gsb_linear_wgrad_tf32x3): the 3xTF32 form.  Both operands stream, so both are split where they land: four
// extra warps turn every landed stage in place into hi = x & ~0x1fff and write lo = x - hi into a second copy of the
// stage (same swizzled offsets), and the issuer accumulates Ghi Xhi + Ghi Xlo + Glo Xhi.  BK = reduction rows per stage
// (16 with SPLIT so that three double-size stages fit in shared memory at N = K = 256).
template <bool SPLIT, int BK>
__global__ void __launch_bounds__(GM_THREADS + (SPLIT ? GM_SPLIT_WARPS * 32 : 0), 1)
wgrad_tf32_kernel(const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_x, WgradParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[GM_MAX_STAGES], empty_bar[GM_MAX_STAGES], split_bar[GM_MAX_STAGES], acc_full;
  __shared__ uint32_t tmem_base_s;
  constexpr int WG_BK = BK;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int a_bytes = p.npad * (WG_BK * 4), b_bytes = p.K * (WG_BK * 4);
  const int half_bytes = a_bytes + b_bytes;            // [A | B]; SPLIT: followed by [A_lo | B_lo]
  const int stage_bytes = half_bytes * (SPLIT ? 2 : 1);
  const long long blocks = (p.rows + WG_BK - 1) / WG_BK;
  const long long b0 = blocks * blockIdx.x / gridDim.x, b1 = blocks * (blockIdx.x + 1) / gridDim.x;
  const int halves = p.npad / 128;
  const uint32_t cols = halves * p.K <= 32 ? 32 : halves * p.K <= 64 ? 64 : halves * p.K <= 128 ? 128
                        : halves * p.K <= 256 ? 256 : 512;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; s++) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
      mbar_init(&split_bar[s], GM_SPLIT_WARPS * 32);
    }
    mbar_init(&acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)),
                 "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (long long b = b0; b < b1; b++) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* a = smem + (size_t)stage * stage_bytes;
        uint8_t* bt = a + a_bytes;
        mbar_expect_tx(&full_bar[stage], (uint32_t)((p.N + p.K) * (WG_BK * 4)));
        for (int nb = 0; nb < p.N / 32; nb++) tma_load_2d(&map_g, &full_bar[stage], a + nb * (WG_BK * 128), nb * 32, (int)(b * WG_BK));
        for (int kb = 0; kb < p.K / 32; kb++) tma_load_2d(&map_x, &full_bar[stage], bt + kb * (WG_BK * 128), kb * 32, (int)(b * WG_BK));
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // D = F32, A = B = TF32, both MN-major (bits 15, 16), N = K_dim, M = 128
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) |
                             ((uint32_t)(p.K >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      for (long long b = b0; b < b1; b++) {
        mbar_wait(SPLIT ? &split_bar[stage] : &full_bar[stage], phase);
        tc_fence_after();
        const uint32_t a = smem_u32(smem + (size_t)stage * stage_bytes);
        const uint32_t bt = a + a_bytes;
#pragma unroll
        for (int s = 0; s < WG_BK / GM_UMMA_K; s++) {
          const uint64_t bd = umma_desc_mn_sw128<BK>(bt + s * 1024);
          const uint64_t bdlo = umma_desc_mn_sw128<BK>(bt + half_bytes + s * 1024);
          for (int h = 0; h < halves; h++) {
            const uint32_t ah = a + h * (4 * WG_BK * 128) + s * 1024;
            const uint64_t ad = umma_desc_mn_sw128<BK>(ah);
            const uint32_t d = tmem_base + (uint32_t)(h * p.K);
            tc_mma_tf32(d, ad, bd, idesc, (b != b0 || s != 0) ? 1u : 0u);
            if (SPLIT) {
              tc_mma_tf32(d, ad, bdlo, idesc, 1u);
              tc_mma_tf32(d, umma_desc_mn_sw128<BK>(ah + half_bytes), bd, idesc, 1u);
            }
          }
        }
        tc_commit(&empty_bar[stage]);
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
      tc_commit(&acc_full);
    }
  } else if (warp >= 6) {
    // ===== splitter (SPLIT only): G and X blocks of the landed stage -> hi (in place) + lo =====
    const int tid = threadIdx.x - GM_THREADS;
    const int a_f4 = p.N * (WG_BK * 4) / 16, b_f4 = p.K * (WG_BK * 4) / 16;   // the loaded parts (padding rows stay as they are)
    int stage = 0;
    uint32_t phase = 0;
    for (long long b = b0; b < b1; b++) {
      mbar_wait(&full_bar[stage], phase);
      uint8_t* base = smem + (size_t)stage * stage_bytes;
      for (int part = 0; part < 2; part++) {
        float4* hi = reinterpret_cast<float4*>(base + (part ? a_bytes : 0));
        float4* lo = reinterpret_cast<float4*>(base + (part ? a_bytes : 0) + half_bytes);
        const int n4 = part ? b_f4 : a_f4;
        for (int o = tid; o < n4; o += GM_SPLIT_WARPS * 32) {
          const float4 x = hi[o];
          float4 h;
          h.x = __uint_as_float(__float_as_uint(x.x) & 0xFFFFE000u);
          h.y = __uint_as_float(__float_as_uint(x.y) & 0xFFFFE000u);
          h.z = __uint_as_float(__float_as_uint(x.z) & 0xFFFFE000u);
          h.w = __uint_as_float(__float_as_uint(x.w) & 0xFFFFE000u);
          hi[o] = h;
          lo[o] = make_float4(x.x - h.x, x.y - h.y, x.z - h.z, x.w - h.w);
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the MMA
      mbar_arrive(&split_bar[stage]);
      if (++stage == p.stages) { stage = 0; phase ^= 1; }
    }
  } else {
    const int q = warp & 3;
    float* out = p.partial + (size_t)blockIdx.x * p.N * p.K;
    if (b1 > b0) {
      mbar_wait(&acc_full, 0);
      tc_fence_after();
    }
    for (int h = 0; h < halves; h++) {
      const int n = h * 128 + q * 32 + lane;
      for (int c = 0; c < p.K; c += 16) {
        uint32_t v[16];
        if (b1 > b0) {
          tc_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * p.K + c), v);
          tc_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 16; i++) v[i] = 0u;
        }
        if (n < p.N) {
#pragma unroll
          for (int i = 0; i < 16; i += 4)
            *reinterpret_cast<float4*>(out + (size_t)n * p.K + c + i) =
                make_float4(__uint_as_float(v[i]), __uint_as_float(v[i + 1]), __uint_as_float(v[i + 2]), __uint_as_float(v[i + 3]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(cols) : "memory");
}

__global__ void wgrad_reduce_kernel(int parts, long long n, const float* __restrict__ partial, float* __restrict__ dW,
                                    int accumulate) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float s = 0.f;
  for (int c = 0; c < parts; c++) s += partial[(size_t)c * n + i];
  dW[i] = accumulate ? dW[i] + s : s;
}

}  // namespace gsb

// ------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------
namespace gsb {

__global__ void split_tf32_kernel(long long n, const float* __restrict__ w, float* __restrict__ hi,
                                  float* __restrict__ lo) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float x = w[i];
  const float h = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
  hi[i] = h;
  lo[i] = x - h;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 2-D FP32 tensor [rows][cols] with row pitch ld (floats), box [box_rows][32 floats], 128B swizzle, zero fill.
static bool make_map(CUtensorMap* m, const float* base, long long rows, int cols, long long ld, int box_rows,
                     CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return false;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
  cuuint32_t box[2] = {(cuuint32_t)GM_BLOCK_K, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static int sm_count() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  }
  return n;
}

template <bool SPLIT>
static int launch_linear(long long rows, int K, int N, const float* X, long long ldx, const float* W, long long ldw,
                         const float* Wlo, const float* bias, float* Y, long long ldy, double* col_stats,
                         cudaStream_t st, const float* residual = nullptr, long long ldr = 0) {
  GemmParams p;
  p.residual = residual; p.ldr = ldr;
  p.col_stats = col_stats;
  p.rows = rows; p.K = K; p.N = N;
  p.umma_n = (N + 15) / 16 * 16;
  p.b_bytes = p.umma_n * GM_BLOCK_K * 4;
  p.bias = bias; p.Y = Y; p.ldy = ldy;
  const int stage_bytes = (GM_A_BYTES + p.b_bytes) * (SPLIT ? 2 : 1);
  p.tma_store = ((reinterpret_cast<uintptr_t>(Y) & 15u) == 0 && (ldy & 3) == 0 && (N & 3) == 0) ? 1 : 0;
  if (col_stats && !p.tma_store) return GSB_EINVAL;   // statistics come from the staged (vector) epilogue
  if (residual && (!p.tma_store || (ldr & 3) || (reinterpret_cast<uintptr_t>(residual) & 15u))) return GSB_EINVAL;
  const int budget = 223 * 1024 - GM_EPI_BYTES;   // 227 KB per block minus static shared memory and alignment slack
  p.stages = budget / stage_bytes;
  if (p.stages > GM_MAX_STAGES) p.stages = GM_MAX_STAGES;
  if (p.stages < 2) return GSB_EINVAL;
  CUtensorMap mx, mw, mwlo, my;
  if (!make_map(&mx, X, rows, K, ldx, GM_BLOCK_M) || !make_map(&mw, W, N, K, ldw, p.umma_n) ||
      !make_map(&mwlo, SPLIT ? Wlo : W, N, K, ldw, p.umma_n) ||
      !make_map(&my, p.tma_store ? Y : X, rows, p.tma_store ? N : K, p.tma_store ? ldy : ldx, 32)) {
    set_last_error("cuTensorMapEncodeTiled", cudaErrorInvalidValue);
    return GSB_ECUDA;
  }
  const size_t smem = (size_t)p.stages * stage_bytes + GM_EPI_BYTES + 1024;
  auto kern = linear_tf32_kernel<SPLIT>;
  static bool attr_set = false;
  if (!attr_set) {
    GSB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    attr_set = true;
  }
  const long long m_tiles = (rows + GM_BLOCK_M - 1) / GM_BLOCK_M;
  const int grid = (int)(m_tiles < sm_count() ? m_tiles : sm_count());
  kern<<<grid, GM_THREADS + (SPLIT ? GM_SPLIT_WARPS * 32 : 0), smem, st>>>(mx, mw, mwlo, my, p);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

}  // namespace gsb

using namespace gsb;

extern "C" {

static int linear_args_ok(int64_t rows, int32_t K, int32_t N, const float* X, int64_t ldx, const float* W, int64_t ldw,
                          const float* Y, int64_t ldy) {
  if (rows < 0 || K <= 0 || N <= 0 || N > 256 || rows > 0x7fffffffLL) return 0;
  if (!X || !W || !Y || ldx < K || ldw < K || ldy < N) return 0;
  // TMA: 16-byte aligned bases and row pitches
  if ((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(W)) & 15u) return 0;
  if ((ldx & 3) || (ldw & 3)) return 0;
  return 1;
}

int gsb_linear_tf32(int64_t rows, int32_t K, int32_t N, const float* X, int64_t ldx, const float* W, int64_t ldw,
                    const float* bias, float* Y, int64_t ldy, double* col_stats, void* stream) {
  if (rows == 0) return GSB_OK;
  if (!linear_args_ok(rows, K, N, X, ldx, W, ldw, Y, ldy)) return GSB_EINVAL;
  return launch_linear<false>(rows, K, N, X, ldx, W, ldw, nullptr, bias, Y, ldy, col_stats,
                              reinterpret_cast<cudaStream_t>(stream));
}

size_t gsb_linear_tf32x3_scratch_bytes(int32_t K, int32_t N) { return (size_t)2 * N * K * sizeof(float); }

int gsb_linear_tf32x3(int64_t rows, int32_t K, int32_t N, const float* X, int64_t ldx, const float* W, int64_t ldw,
                      const float* bias, float* Y, int64_t ldy, double* col_stats, void* scratch, void* stream) {
  if (rows == 0) return GSB_OK;
  if (!linear_args_ok(rows, K, N, X, ldx, W, ldw, Y, ldy) || !scratch || ldw != K) return GSB_EINVAL;
  if (reinterpret_cast<uintptr_t>(scratch) & 15u) return GSB_EINVAL;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  float* hi = reinterpret_cast<float*>(scratch);
  float* lo = hi + (size_t)N * K;
  const long long n = (long long)N * K;
  split_tf32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, W, hi, lo);
  GSB_LAUNCH_CHECK();
  return launch_linear<true>(rows, K, N, X, ldx, hi, K, lo, bias, Y, ldy, col_stats, st);
}

int gsb_linear_residual(int64_t rows, int32_t K, int32_t N, const float* X, int64_t ldx, const float* W, int64_t ldw,
                        const float* residual, int64_t ldr, float* Y, int64_t ldy, int32_t precise, void* scratch,
                        void* stream) {
  if (rows == 0) return GSB_OK;
  if (!linear_args_ok(rows, K, N, X, ldx, W, ldw, Y, ldy) || !residual || ldr < N) return GSB_EINVAL;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (!precise)
    return launch_linear<false>(rows, K, N, X, ldx, W, ldw, nullptr, nullptr, Y, ldy, nullptr, st, residual, ldr);
  if (!scratch || ldw != K || (reinterpret_cast<uintptr_t>(scratch) & 15u)) return GSB_EINVAL;
  float* hi = reinterpret_cast<float*>(scratch);
  float* lo = hi + (size_t)N * K;
  const long long n = (long long)N * K;
  split_tf32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, W, hi, lo);
  GSB_LAUNCH_CHECK();
  return launch_linear<true>(rows, K, N, X, ldx, hi, K, lo, nullptr, Y, ldy, nullptr, st, residual, ldr);
}

size_t gsb_linear_wgrad_scratch_bytes(int32_t N, int32_t K) { return (size_t)148 * 2 * N * K * sizeof(float); }

}  // extern "C"

template <bool SPLIT, int BK>
static int launch_wgrad(int64_t rows, int32_t N, int32_t K, const float* G, int64_t ldg, const float* X, int64_t ldx,
                        float* dW, int32_t accumulate, void* scratch, void* stream) {
  if (rows < 0 || N <= 0 || K <= 0 || N > 256 || K > 256 || (N & 31) || (K & 31) || rows > 0x7fffffffLL) return GSB_EINVAL;
  if (!dW || !scratch) return GSB_EINVAL;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (rows == 0) {
    if (!accumulate) GSB_CUDA_CHECK(cudaMemsetAsync(dW, 0, (size_t)N * K * sizeof(float), st));
    return GSB_OK;
  }
  if (!G || !X || ldg < N || ldx < K || (ldg & 3) || (ldx & 3)) return GSB_EINVAL;
  if ((reinterpret_cast<uintptr_t>(G) | reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(scratch)) & 15u)
    return GSB_EINVAL;
  WgradParams p;
  p.rows = rows; p.N = N; p.K = K; p.npad = (N + 127) / 128 * 128;
  p.partial = reinterpret_cast<float*>(scratch);
  const int stage_bytes = (p.npad + K) * (BK * 4) * (SPLIT ? 2 : 1);
  p.stages = (220 * 1024) / stage_bytes;
  if (p.stages > GM_MAX_STAGES) p.stages = GM_MAX_STAGES;
  if (p.stages < 2) return GSB_EINVAL;
  CUtensorMap mg, mx;
  if (!make_map(&mg, G, rows, N, ldg, BK, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B) ||
      !make_map(&mx, X, rows, K, ldx, BK, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)) {
    set_last_error("cuTensorMapEncodeTiled", cudaErrorInvalidValue);
    return GSB_ECUDA;
  }
  auto kern = wgrad_tf32_kernel<SPLIT, BK>;
  static bool attr_set = false;
  if (!attr_set) {
    GSB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    attr_set = true;
  }
  const long long blocks = (rows + BK - 1) / BK;
  int grid = sm_count() < 296 ? sm_count() : 296;   // gsb_linear_wgrad_scratch_bytes holds 296 partials
  if (grid > blocks) grid = (int)blocks;
  const size_t smem = (size_t)p.stages * stage_bytes + 1024;
  kern<<<grid, GM_THREADS + (SPLIT ? GM_SPLIT_WARPS * 32 : 0), smem, st>>>(mg, mx, p);
  GSB_LAUNCH_CHECK();
  const long long n = (long long)N * K;
  wgrad_reduce_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(grid, n, p.partial, dW, accumulate);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

extern "C" {
int gsb_linear_wgrad_tf32(int64_t rows, int32_t N, int32_t K, const float* G, int64_t ldg, const float* X, int64_t ldx,
                          float* dW, int32_t accumulate, void* scratch, void* stream) {
  return launch_wgrad<false, 32>(rows, N, K, G, ldg, X, ldx, dW, accumulate, scratch, stream);
}
int gsb_linear_wgrad_tf32x3(int64_t rows, int32_t N, int32_t K, const float* G, int64_t ldg, const float* X, int64_t ldx,
                            float* dW, int32_t accumulate, void* scratch, void* stream) {
  return launch_wgrad<true, 16>(rows, N, K, G, ldg, X, ldx, dW, accumulate, scratch, stream);
}

}  // extern "C"
