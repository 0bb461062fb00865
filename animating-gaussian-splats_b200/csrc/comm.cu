// comm.cu -- the multi-GPU exchange of the camera-sharded step as ONE kernel over NVLink peer memory.
//
// The reference has no distributed code; SURVEY.md section 8(e) defines the exchange: after its cameras' backward
// every rank holds a flat [P][GSB_FLAT_WIDTH] buffer (gradients, statistic increments, this batch's largest radius)
// that must be summed (the radius column: maximised) over the ranks.  Every rank's buffer lives in a block that all
// peers have mapped (CUDA IPC, gsb_comm_alloc / gsb_comm_open), and the reduction is a two-shot all-reduce written
// directly against those mappings:
//
//   scatter   rank r stores its part of every other rank's slice into that rank's staging rows (P2P stores);
//   meet      per block, release / acquire on flag words in the blocks themselves (st.release.sys / ld.acquire.sys);
//   reduce    rank r owns slice r: it sums its own part and the staged rows in rank order (local loads) and stores
//             the result into every peer's buffer (P2P stores) -- the reduce-scatter and the all-gather of a ring /
//             tree library collapse into one launch, the NVSwitch carries (world - 1) / world of the buffer in each
//             direction per GPU and phase, every byte as a posted store;
//   leave     "my slice is everywhere" to every peer, wait for all of them: when the kernel ends the local buffer
//             holds the complete result and no peer reads or writes it any more, so the next step may overwrite it.
//
// Only rank r ever computes slice r, so all replicas end bit-identical.  No host involvement, no NCCL call, no
// staging copy: the producer (preprocess_bwd_kernel<FLAT>) writes straight into the mapped block.  Flag words carry a
// per-block epoch that the kernel itself advances, so launches can be replayed from a CUDA graph.
#include <algorithm>
#include <cstdlib>
#include "common.cuh"

namespace gsb {

constexpr int COMM_MAX_BLOCKS = 512;   // flag slots per block of the exchange
constexpr int COMM_SLOT_WORDS = 64;    // ready[16] | done[16] | epoch | error | pad
constexpr int COMM_THREADS = 512;
constexpr unsigned long long COMM_TIMEOUT_NS = 10ull * 1000 * 1000 * 1000;

// block layout: payload | staging (one row of ceil(n / world) words per rank: <= payload + 16 world bytes) | flags
static size_t comm_stage_bytes(size_t payload_bytes) { return align_up(payload_bytes + 16 * GSB_COMM_MAX_RANKS, 256); }
size_t comm_bytes(size_t payload_bytes) {
  return align_up(payload_bytes, 256) + comm_stage_bytes(payload_bytes) +
         (size_t)COMM_MAX_BLOCKS * COMM_SLOT_WORDS * sizeof(uint32_t);
}

struct CommPeers { char* p[GSB_COMM_MAX_RANKS]; };

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ float4 ld_peer(const float4* p) {   // L2 only: peer lines must never be served from a stale L1
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}

// wait until *flag has reached `epoch`; false on timeout (a peer died): the caller records it and leaves
__device__ __forceinline__ bool comm_wait(const uint32_t* flag, uint32_t epoch) {
  if ((int)(ld_acquire_sys(flag) - epoch) >= 0) return true;
  const unsigned long long t0 = global_ns();
  while ((int)(ld_acquire_sys(flag) - epoch) < 0) {
    if (global_ns() - t0 > COMM_TIMEOUT_NS) return false;
    __nanosleep(40);
  }
  return true;
}

__device__ __forceinline__ float4 comm_combine(float4 a, float4 b, int max_comp) {
  float4 r = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
  if (max_comp == 0) r.x = fmaxf(a.x, b.x);
  else if (max_comp == 1) r.y = fmaxf(a.y, b.y);
  else if (max_comp == 2) r.z = fmaxf(a.z, b.z);
  else if (max_comp == 3) r.w = fmaxf(a.w, b.w);
  return r;
}

// WORLD > 0: compile-time rank count (loops over the peers fully unrolled); WORLD == 0: run-time count (<= 16).
//
// Push form.  A first version PULLED slice r from every peer with P2P loads; at 8 GPUs it moved 24 MB in 91 us
// (~230 GB/s per direction): remote loads pay a round trip each and the links' read tracking limits how many are
// outstanding.  Here every byte crosses NVLink as a posted STORE:
//   scatter   rank r copies its part of slice q into peer q's staging area (row r), for every q != r;
//   meet      release / acquire on the flag words, per block: block b of every rank handles the same words of every
//             slice, so block b of rank q only needs the blocks b of its peers;
//   reduce    rank r sums its own part and the world - 1 staged rows of slice r in rank order (local loads, MAX for
//             the radius component) and stores the result into every peer's buffer;
//   leave     second release / acquire: when the kernel ends the local buffer is complete and nobody writes it.
template <int WORLD>
__global__ void __launch_bounds__(COMM_THREADS, (WORLD >= 8 || WORLD == 0) ? 1 : 2)
allreduce_flat_kernel(CommPeers peers, int rank, int world_rt, size_t stage_off, size_t flags_off, long long n4,
                      long long slice_cap, int row4, int max_word, int max_comp) {
  const int world = WORLD > 0 ? WORLD : world_rt;
  const int b = blockIdx.x, G = gridDim.x;
  const unsigned tid = threadIdx.x;
  uint32_t* my_slot = reinterpret_cast<uint32_t*>(peers.p[rank] + flags_off) + (size_t)b * COMM_SLOT_WORDS;
  __shared__ uint32_t s_epoch;
  __shared__ int s_fail;
  if (tid == 0) { s_epoch = my_slot[32] + 1u; s_fail = 0; }
  __syncthreads();
  const uint32_t epoch = s_epoch;
  const float4* mine = reinterpret_cast<const float4*>(peers.p[rank]);
  const long long stride = (long long)G * COMM_THREADS;
  const long long first = (long long)b * COMM_THREADS + tid;

  // scatter: my contribution to every other rank's slice goes into that rank's staging row `rank`
  for (long long o = first; o < slice_cap; o += stride) {
#pragma unroll
    for (int q = 0; q < (WORLD > 0 ? WORLD : GSB_COMM_MAX_RANKS); q++) {
      if (q >= world || q == rank) continue;
      const long long lo = n4 * q / world, hi = n4 * (q + 1) / world;
      if (lo + o < hi)
        (reinterpret_cast<float4*>(peers.p[q] + stage_off) + (size_t)rank * slice_cap)[o] = mine[lo + o];
    }
  }
  __syncthreads();
  if (tid < (unsigned)world) {
    uint32_t* peer_slot = reinterpret_cast<uint32_t*>(peers.p[tid] + flags_off) + (size_t)b * COMM_SLOT_WORDS;
    st_release_sys(peer_slot + rank, epoch);   // release: ordered after this block's stores (bar.sync above)
    if (!comm_wait(my_slot + tid, epoch)) s_fail = 1;
  }
  __syncthreads();

  if (!s_fail) {
    const long long lo = n4 * rank / world, hi = n4 * (rank + 1) / world;
    const float4* stage = reinterpret_cast<const float4*>(peers.p[rank] + stage_off);
    long long i = lo + first;
    int phase = (int)(((i - max_word) % row4 + row4) % row4);   // 0: word i holds the MAX component
    const int step = (int)(stride % row4);
    for (; i < hi; i += stride) {
      const int mc = (max_comp >= 0 && i >= max_word && phase == 0) ? max_comp : -1;
      float4 v[WORLD > 0 ? WORLD : GSB_COMM_MAX_RANKS];
#pragma unroll
      for (int q = 0; q < (WORLD > 0 ? WORLD : GSB_COMM_MAX_RANKS); q++)
        if (q < world) v[q] = (q == rank) ? mine[i] : ld_peer(stage + (size_t)q * slice_cap + (i - lo));
      float4 acc = v[0];
#pragma unroll
      for (int q = 1; q < (WORLD > 0 ? WORLD : GSB_COMM_MAX_RANKS); q++)
        if (q < world) acc = comm_combine(acc, v[q], mc);
#pragma unroll
      for (int q = 0; q < (WORLD > 0 ? WORLD : GSB_COMM_MAX_RANKS); q++)
        if (q < world) reinterpret_cast<float4*>(peers.p[q])[i] = acc;
      phase += step;
      if (phase >= row4) phase -= row4;
    }
  }
  __syncthreads();

  // leave: my slice is in every peer's buffer; wait until every peer's slice is in mine
  if (tid < (unsigned)world) {
    uint32_t* peer_slot = reinterpret_cast<uint32_t*>(peers.p[tid] + flags_off) + (size_t)b * COMM_SLOT_WORDS;
    st_release_sys(peer_slot + 16 + rank, epoch);
    if (!comm_wait(my_slot + 16 + tid, epoch)) s_fail = 1;
  }
  __syncthreads();
  if (tid == 0) {
    my_slot[32] = epoch;
    if (s_fail) my_slot[33] = 1u;   // a peer never arrived: the buffer contents are undefined
  }
}

int launch_allreduce_flat(int rank, int world, void* const* peers, size_t payload_bytes, int64_t n_float4, int row_float4,
                          int max_word, int max_comp, cudaStream_t st) {
  if (world == 1 || n_float4 == 0) return GSB_OK;
  CommPeers cp;
  for (int i = 0; i < GSB_COMM_MAX_RANKS; i++) cp.p[i] = i < world ? reinterpret_cast<char*>(peers[i]) : nullptr;
  const size_t stage_off = align_up(payload_bytes, 256);
  const size_t flags_off = stage_off + comm_stage_bytes(payload_bytes);
  const long long slice_cap = (n_float4 + world - 1) / world;
  if ((size_t)slice_cap * world * 16 > comm_stage_bytes(payload_bytes)) return GSB_EINVAL;
  int dev = 0, sms = 148;
  GSB_CUDA_CHECK(cudaGetDevice(&dev));
  GSB_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const long long per_rank = (n_float4 + world - 1) / world;
  static const int per_sm = [] { const char* e = getenv("GSB_COMM_BLOCKS_PER_SM"); return e ? std::max(1, atoi(e)) : 0; }();
  const int bps = per_sm ? per_sm : (world == 2 || world == 4 ? 2 : 1);   // what the launch bounds of the instance allow
  int grid = (int)std::min<long long>(std::min(bps * sms, COMM_MAX_BLOCKS), (per_rank + 4 * COMM_THREADS - 1) / (4 * COMM_THREADS));   // >= 4 pipelined words per thread
  if (grid < 1) grid = 1;
#define GSB_AR_CASE(W)                                                                                             \
  case W:                                                                                                          \
    allreduce_flat_kernel<W><<<grid, COMM_THREADS, 0, st>>>(cp, rank, world, stage_off, flags_off, (long long)n_float4, \
                                                             slice_cap, row_float4, max_word, max_comp);           \
    break
  switch (world) {
    GSB_AR_CASE(2);
    GSB_AR_CASE(4);
    GSB_AR_CASE(8);
    default:
      allreduce_flat_kernel<0><<<grid, COMM_THREADS, 0, st>>>(cp, rank, world, stage_off, flags_off, (long long)n_float4,
                                                              slice_cap, row_float4, max_word, max_comp);
  }
#undef GSB_AR_CASE
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

}  // namespace gsb
