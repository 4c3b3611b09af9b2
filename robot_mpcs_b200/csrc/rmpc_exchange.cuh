// Result exchange between the ranks of one node over peer memory (NVLink / NVSwitch), SURVEY.md 8e.
//
// The scenario batch is sharded over one process per GPU and the solve needs no collective; what a control loop wants back
// is every scenario's (action, exitflag, iterations) on every rank.  Round 1 used one NCCL all_gather per step for that; its
// kernel (one 512-thread CTA per channel) is resident from the end of the rank's own solve until the slowest peer is done and
// takes SM room from the one-warp solve blocks.  Here every rank owns one device region that its peers
// map through CUDA IPC, and
//   publish : a small kernel behind the rank's solve stores its rows straight into the region of every rank (peer stores are
//             posted: nothing waits for the link), then one flag per (step, rank);
//   collect : one resident warp waits for the world's flags of that step, a second small kernel unpacks the rows and
//             acknowledges the step to every rank so that its ring slot can be written again.
// No rank ever reads remote memory, nothing of this sits on an SM while the solve runs except the one waiting warp.
//
// Protocol.  Step numbers (`seq`) are counted by the order of the publish calls, which has to be the same on every rank (as for
// any collective); a step lives in ring slot seq % depth.  Region of rank r (written by the peers, read by r only):
//   status[4] | flags[depth][world] | acks[world] | counters[2][depth] | rows[depth][world][rows_per_rank][width]
//   flags[slot][w] = seq + 1 once rank w's rows of step seq are in the slot;  acks[w] = steps rank w has collected (in order).
// A publish of step seq >= depth first waits until every rank has collected step seq - depth.  Every wait goes to work that
// was enqueued EARLIER in the canonical per-rank order publish(s), collect(s), publish(s+1), ... (on this or another rank), so
// the protocol cannot deadlock even where two CUDA streams share a hardware queue; each wait also gives up after
// kTimeoutNs and raises the region's status word instead of hanging the GPU.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace rmpc_x {

constexpr int kMaxWorld = 16;
constexpr unsigned long long kTimeoutNs = 20ull * 1000ull * 1000ull * 1000ull;

struct Layout {
  int world, rows, width, depth;
  __host__ __device__ size_t off_flags() const { return 4 * sizeof(uint64_t); }
  __host__ __device__ size_t off_acks() const { return off_flags() + size_t(depth) * world * sizeof(uint64_t); }
  __host__ __device__ size_t off_counters() const { return off_acks() + size_t(world) * sizeof(uint64_t); }
  __host__ __device__ size_t off_rows() const { return (off_counters() + 2 * size_t(depth) * sizeof(uint32_t) + 255) / 256 * 256; }
  __host__ __device__ size_t bytes() const { return off_rows() + size_t(depth) * world * rows * width * sizeof(double); }
};

struct View {   // passed to the kernels by value
  Layout l;
  int rank;
  char* region[kMaxWorld];   // region[w] = rank w's region as mapped into this process
  __device__ uint64_t* status(int w) const { return reinterpret_cast<uint64_t*>(region[w]); }
  __device__ uint64_t* flags(int w, int slot) const { return reinterpret_cast<uint64_t*>(region[w] + l.off_flags()) + size_t(slot) * l.world; }
  __device__ uint64_t* acks(int w) const { return reinterpret_cast<uint64_t*>(region[w] + l.off_acks()); }
  __device__ uint32_t* counter(int kind, int slot) const { return reinterpret_cast<uint32_t*>(region[rank] + l.off_counters()) + kind * l.depth + slot; }
  __device__ double* rows(int w, int slot, int from) const {
    return reinterpret_cast<double*>(region[w] + l.off_rows()) + (size_t(slot) * l.world + from) * l.rows * l.width;
  }
};

__device__ __forceinline__ unsigned long long now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// one warp: lane w waits until word[w] satisfies (exact ? == want : >= want)
__global__ void __launch_bounds__(32) wait_kernel(const __grid_constant__ View v, int what /* 0 acks, 1 flags */, int slot, unsigned long long want) {
  const int w = threadIdx.x;
  if (w < v.l.world) {
    const volatile uint64_t* p = (what == 0 ? v.acks(v.rank) : v.flags(v.rank, slot)) + w;
    const unsigned long long t0 = now_ns();
    for (;;) {
      const uint64_t x = *p;
      if (what == 0 ? x >= want : x == want) break;
      if (now_ns() - t0 > kTimeoutNs) { atomicMax(reinterpret_cast<unsigned long long*>(v.status(v.rank)), (unsigned long long)(what + 1)); break; }
      __nanosleep(200);
    }
  }
  __threadfence_system();
}

// thread per row of this rank's shard: the row goes to the slot of every rank (its own included)
__global__ void __launch_bounds__(256) publish_kernel(const __grid_constant__ View v, int slot, unsigned long long tag, int B, int nu,
                                                      const double* __restrict__ action, const int* __restrict__ exitflag,
                                                      const int* __restrict__ iters) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) {
    double row[8];
    for (int j = 0; j < nu; ++j) row[j] = action[size_t(b) * nu + j];
    const int2 packed = make_int2(exitflag ? exitflag[b] : 0, iters ? iters[b] : 0);
    row[nu] = *reinterpret_cast<const double*>(&packed);
    for (int p = 0; p < v.l.world; ++p) {
      const int w = (v.rank + p) % v.l.world;   // ranks start with different peers
      double* dst = v.rows(w, slot, v.rank) + size_t(b) * v.l.width;
      for (int j = 0; j <= nu; ++j) dst[j] = row[j];
    }
  }
  __threadfence_system();
  __syncthreads();
  __shared__ bool last;
  if (threadIdx.x == 0) {
    uint32_t* c = v.counter(0, slot);
    last = atomicAdd(c, 1u) == gridDim.x - 1;
    if (last) *c = 0;
  }
  __syncthreads();
  if (last && threadIdx.x < v.l.world) {
    __threadfence_system();
    *reinterpret_cast<volatile uint64_t*>(v.flags(threadIdx.x, slot) + v.rank) = tag;
  }
}

// thread per row of the whole batch: unpack the slot, then acknowledge the step to every rank
__global__ void __launch_bounds__(256) unpack_kernel(const __grid_constant__ View v, int slot, unsigned long long collected, int nu,
                                                     double* __restrict__ action, int* __restrict__ exitflag, int* __restrict__ iters) {
  const int total = v.l.world * v.l.rows;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < total) {
    const double* src = v.rows(v.rank, slot, 0) + size_t(i) * v.l.width;
    for (int j = 0; j < nu; ++j) action[size_t(i) * nu + j] = __ldcg(src + j);
    const double pk = __ldcg(src + nu);
    const int2 packed = *reinterpret_cast<const int2*>(&pk);
    if (exitflag) exitflag[i] = packed.x;
    if (iters) iters[i] = packed.y;
  }
  __syncthreads();
  __shared__ bool last;
  if (threadIdx.x == 0) {
    __threadfence();
    uint32_t* c = v.counter(1, slot);
    last = atomicAdd(c, 1u) == gridDim.x - 1;
    if (last) *c = 0;
  }
  __syncthreads();
  if (last && threadIdx.x < v.l.world) {
    __threadfence_system();
    *reinterpret_cast<volatile uint64_t*>(v.acks(threadIdx.x) + v.rank) = collected;
  }
}

}  // namespace rmpc_x
