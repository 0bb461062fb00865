// comm.cu -- the multi-GPU exchange of the camera-sharded step as ONE kernel over NVLink peer memory.
//
// The reference has no distributed code; SURVEY.md section 8(e) defines the exchange: after its cameras' backward
// every rank holds a flat [P][GSB_FLAT_WIDTH] buffer (gradients, statistic increments, this batch's largest radius)
// that must be summed (the radius column: maximised) over the ranks.  Every rank's buffer lives in a block that all
// peers have mapped (CUDA IPC, gsb_comm_alloc / gsb_comm_open), and the reduction is a two-shot all-reduce written
// directly against those mappings:
//
//   meet      every block b of every rank tells block b of every peer "my buffer is complete" (st.release.sys on a
//             flag word in the peer's block) and waits for the same from all of them (ld.acquire.sys);
//   reduce    rank r owns slice r: it loads that slice from every peer's buffer (P2P loads, all in flight before the
//             first add), reduces in rank order and stores the result into every peer's buffer (P2P stores) -- the
//             reduce-scatter and the all-gather of a ring/tree library collapse into one pass, the NVSwitch carries
//             (world - 1) / world of the buffer in each direction per GPU;
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

size_t comm_bytes(size_t payload_bytes) {
  return align_up(payload_bytes, 256) + (size_t)COMM_MAX_BLOCKS * COMM_SLOT_WORDS * sizeof(uint32_t);
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

// WORLD > 0: compile-time rank count (loads fully unrolled, all in flight); WORLD == 0: run-time count (<= 16)
template <int WORLD>
__global__ void __launch_bounds__(COMM_THREADS, (WORLD >= 8 || WORLD == 0) ? 1 : 2)
allreduce_flat_kernel(CommPeers peers, int rank, int world_rt, size_t flags_off, long long n4, int row4, int max_word,
                      int max_comp) {
  const int world = WORLD > 0 ? WORLD : world_rt;
  const int b = blockIdx.x, G = gridDim.x;
  const unsigned tid = threadIdx.x;
  uint32_t* my_slot = reinterpret_cast<uint32_t*>(peers.p[rank] + flags_off) + (size_t)b * COMM_SLOT_WORDS;
  __shared__ uint32_t s_epoch;
  __shared__ int s_fail;
  if (tid == 0) { s_epoch = my_slot[32] + 1u; s_fail = 0; }
  __syncthreads();
  const uint32_t epoch = s_epoch;

  // meet: my buffer is complete (this kernel follows its producer on the stream); wait for every peer's
  if (tid < (unsigned)world) {
    uint32_t* peer_slot = reinterpret_cast<uint32_t*>(peers.p[tid] + flags_off) + (size_t)b * COMM_SLOT_WORDS;
    st_release_sys(peer_slot + rank, epoch);
    if (!comm_wait(my_slot + tid, epoch)) s_fail = 1;
  }
  __syncthreads();

  if (!s_fail) {
    const long long lo = n4 * rank / world, hi = n4 * (rank + 1) / world;
    const long long stride = (long long)G * COMM_THREADS;
    long long i = lo + (long long)b * COMM_THREADS + tid;
    int phase = (int)(((i - max_word) % row4 + row4) % row4);   // 0: word i holds the MAX component
    const int step = (int)(stride % row4);
    if (WORLD > 0) {
      // software pipeline: the peer loads of word k + 1 are issued before the reduction and the peer stores of word
      // k, so the inbound (pull) and outbound (push) directions of the links are busy at the same time (a first
      // version loaded everything, then stored everything: each direction idle half the time)
      float4 v[WORLD > 0 ? WORLD : 1];
      if (i < hi) {
#pragma unroll
        for (int q = 0; q < WORLD; q++) v[q] = ld_peer(reinterpret_cast<const float4*>(peers.p[q]) + i);
      }
      while (i < hi) {
        const long long nx = i + stride;
        const bool has = nx < hi;
        float4 w[WORLD > 0 ? WORLD : 1];
        if (has) {
#pragma unroll
          for (int q = 0; q < WORLD; q++) w[q] = ld_peer(reinterpret_cast<const float4*>(peers.p[q]) + nx);
        }
        const int mc = (max_comp >= 0 && i >= max_word && phase == 0) ? max_comp : -1;
        float4 acc = v[0];
#pragma unroll
        for (int q = 1; q < WORLD; q++) acc = comm_combine(acc, v[q], mc);
#pragma unroll
        for (int q = 0; q < WORLD; q++) reinterpret_cast<float4*>(peers.p[q])[i] = acc;
        if (!has) break;
#pragma unroll
        for (int q = 0; q < WORLD; q++) v[q] = w[q];
        i = nx;
        phase += step;
        if (phase >= row4) phase -= row4;
      }
    } else {
      for (; i < hi; i += stride) {
        const int mc = (max_comp >= 0 && i >= max_word && phase == 0) ? max_comp : -1;
        float4 acc = ld_peer(reinterpret_cast<const float4*>(peers.p[0]) + i);
        for (int q = 1; q < world; q++) acc = comm_combine(acc, ld_peer(reinterpret_cast<const float4*>(peers.p[q]) + i), mc);
        for (int q = 0; q < world; q++) reinterpret_cast<float4*>(peers.p[q])[i] = acc;
        phase += step;
        if (phase >= row4) phase -= row4;
      }
    }
  }
  __syncthreads();

  // leave: my slice is in every peer's buffer; wait until every peer's slice is in mine
  if (tid < (unsigned)world) {
    uint32_t* peer_slot = reinterpret_cast<uint32_t*>(peers.p[tid] + flags_off) + (size_t)b * COMM_SLOT_WORDS;
    st_release_sys(peer_slot + 16 + rank, epoch);   // release: ordered after this block's stores (bar.sync above)
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
  const size_t flags_off = align_up(payload_bytes, 256);
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
    allreduce_flat_kernel<W><<<grid, COMM_THREADS, 0, st>>>(cp, rank, world, flags_off, (long long)n_float4, row_float4, \
                                                             max_word, max_comp);                                  \
    break
  switch (world) {
    GSB_AR_CASE(2);
    GSB_AR_CASE(4);
    GSB_AR_CASE(8);
    default:
      allreduce_flat_kernel<0><<<grid, COMM_THREADS, 0, st>>>(cp, rank, world, flags_off, (long long)n_float4, row_float4,
                                                              max_word, max_comp);
  }
#undef GSB_AR_CASE
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

}  // namespace gsb
