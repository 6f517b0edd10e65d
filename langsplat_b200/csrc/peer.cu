// peer.cu — the cross-GPU synchronisation of a view-sharded step.  One process per GPU; every rank owns
// a buffer allocated with lsr_peer_alloc whose first LSR_PEER_HEADER_BYTES hold the words below, and has the other
// ranks' buffers mapped (lsr_peer_open, CUDA IPC over NVLink / NVSwitch).  The barrier is the only ordering the
// exchange needs: after it, every rank's factors of this step are complete and visible, so the SH-gradient kernel
// (preprocess_bwd.cu: sh_grad_views_kernel) can read them straight out of peer memory.
#include <algorithm>

#include "common.cuh"

namespace lsr {
namespace {

struct PeerHeader {
  uint32_t arrived[LSR_MAX_PEERS];  // arrived[r] = last epoch rank r signalled to the owner of this header
  uint32_t timed_out;               // set by the owner's barrier kernel when a peer did not show up in time
  uint32_t auto_epoch;              // the owner's own barrier count when the epochs are kept on the device (mode bit 2)
};
static_assert(sizeof(PeerHeader) <= LSR_PEER_HEADER_BYTES, "header does not fit");

struct PeerHeaders { PeerHeader* h[LSR_MAX_PEERS]; };

__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// One warp.  Lane r < world: publish this rank's epoch in rank r's header (release at system scope, after a system
// fence that orders every earlier write of this stream — the factors — before it), then wait until rank r's epoch
// shows up in our own header (acquire).  Epochs only grow, compared as a signed difference so that wrap-around is
// harmless.  A peer that never arrives must not hang the GPU: after the time-out (lsr_peer_set_timeout_ms, default
// 120 s — long enough for a rank that checkpoints, evaluates or loads data before its step) the wait gives up and
// raises the sticky timed_out word in the own header.  Nothing downstream may then pass for a result: the consumers'
// outputs are overwritten with NaN (peer_poison_kernel, enqueued behind them by the exchange), lsr_peer_status
// reports it, and the Python exchange raises at its next begin_step().

// mode bit 0: signal, bit 1: wait.  Splitting the two lets a rank publish "my remote writes are done", run local work
// that does not depend on the peers, and only then wait for them.
// mode bit 2: the epoch lives in the own header instead of the kernel's arguments — a signal advances it, a wait alone
// uses the value its signal left — so that the launch is the same every step and can be replayed from a CUDA graph
// (every rank runs the same sequence of barriers, so the counts agree without the host).
__global__ void __launch_bounds__(32) peer_barrier_kernel(PeerHeaders hs, int rank, int world, uint32_t epoch,
                                                          int mode, uint64_t timeout_ns) {
  const int r = threadIdx.x;
  if (mode & 4) {
    uint32_t e = 0;
    if (r == 0) {
      e = hs.h[rank]->auto_epoch + ((mode & 1) ? 1u : 0u);
      if (mode & 1) hs.h[rank]->auto_epoch = e;
    }
    epoch = __shfl_sync(0xffffffffu, e, 0);
  }
  __threadfence_system();
  if (r < world && (mode & 1)) {
    uint32_t* theirs = &hs.h[r]->arrived[rank];
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(epoch) : "memory");
  }
  if (r < world && (mode & 2)) {
    const uint32_t* mine = &hs.h[rank]->arrived[r];
    const uint64_t t0 = globaltimer_ns();
    for (;;) {
      uint32_t seen;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(mine) : "memory");
      if ((int32_t)(seen - epoch) >= 0) break;
      if (globaltimer_ns() - t0 > timeout_ns) {
        hs.h[rank]->timed_out = 1u;
        break;
      }
      __nanosleep(64);
    }
  }
  __syncwarp();
  __threadfence_system();
}

// After a time-out the step's sums are incomplete: overwrite `data` with NaN so that no consumer mistakes them for
// gradients.  One word read per thread in the normal case.
__global__ void __launch_bounds__(256) peer_poison_kernel(const PeerHeader* own, float* __restrict__ data, int64_t n) {
  if (own->timed_out == 0u) return;
  const float nan = __int_as_float(0x7fc00000);
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) data[i] = nan;
}

// Two-shot sum over the ranks through peer memory.  Rank r owns the r-th chunk of the vector: it loads that chunk from
// every rank's input buffer (its own included, fixed rank order, so the sum is the same bits whoever computes it),
// adds, and stores the result into EVERY rank's output buffer — the reduce-scatter reads and the all-gather writes of
// an all-reduce, fused into one pass over NVLink with 16-byte accesses.  Callers order it with the barrier: inputs
// complete on all ranks before (signal+wait), outputs complete on all ranks after (signal, ..., wait).
struct PeerBufs {
  const float4* in[LSR_MAX_PEERS];
  float4* out[LSR_MAX_PEERS];
};

constexpr int PAR_THREADS = 256;

template <int WORLD>
__global__ void __launch_bounds__(PAR_THREADS)
peer_allreduce_kernel(PeerBufs b, int world, int64_t first, int64_t end) {
  const int64_t stride = (int64_t)gridDim.x * PAR_THREADS;
  for (int64_t i = first + (int64_t)blockIdx.x * PAR_THREADS + threadIdx.x; i < end; i += stride) {
    const int n = WORLD > 0 ? WORLD : world;
    float4 v[WORLD > 0 ? WORLD : LSR_MAX_PEERS];
#pragma unroll
    for (int r = 0; r < (WORLD > 0 ? WORLD : LSR_MAX_PEERS); r++)
      if (r < n) v[r] = b.in[r][i];  // all loads in flight before the first add
    float4 acc = v[0];
#pragma unroll
    for (int r = 1; r < (WORLD > 0 ? WORLD : LSR_MAX_PEERS); r++)
      if (r < n) { acc.x += v[r].x; acc.y += v[r].y; acc.z += v[r].z; acc.w += v[r].w; }
#pragma unroll
    for (int r = 0; r < (WORLD > 0 ? WORLD : LSR_MAX_PEERS); r++)
      if (r < n) b.out[r][i] = acc;
  }
}

std::atomic<uint64_t> g_timeout_ms{120000};

}  // namespace

void peer_set_timeout_ms(uint64_t ms) { g_timeout_ms.store(ms ? ms : 1, std::memory_order_relaxed); }
uint64_t peer_timeout_ms() { return g_timeout_ms.load(std::memory_order_relaxed); }

void launch_peer_barrier(int rank, int world, void* const* headers, uint32_t epoch, int mode, cudaStream_t s) {
  PeerHeaders hs;
  for (int i = 0; i < LSR_MAX_PEERS; i++) hs.h[i] = i < world ? reinterpret_cast<PeerHeader*>(headers[i]) : nullptr;
  peer_barrier_kernel<<<1, 32, 0, s>>>(hs, rank, world, epoch, mode, peer_timeout_ms() * 1000000ull);
  count_launch();
}

void launch_peer_poison(const void* own_buffer, float* data, int64_t n, cudaStream_t s) {
  if (n <= 0) return;
  const int blocks = (int)std::min<int64_t>((n + 255) / 256, 148 * 8);
  peer_poison_kernel<<<blocks, 256, 0, s>>>(reinterpret_cast<const PeerHeader*>(own_buffer), data, n);
  count_launch();
}

void launch_peer_allreduce(int rank, int world, const float* const* in, float* const* out, int64_t n_floats,
                           cudaStream_t s) {
  const int64_t n4 = n_floats / 4;
  const int64_t chunk = (n4 + world - 1) / world;
  const int64_t first = std::min(n4, (int64_t)rank * chunk), end = std::min(n4, first + chunk);
  if (end <= first) return;
  PeerBufs b;
  for (int i = 0; i < LSR_MAX_PEERS; i++) {
    b.in[i] = i < world ? reinterpret_cast<const float4*>(in[i]) : nullptr;
    b.out[i] = i < world ? reinterpret_cast<float4*>(out[i]) : nullptr;
  }
  const int blocks = (int)std::min<int64_t>((end - first + PAR_THREADS - 1) / PAR_THREADS, 148 * 8);
  switch (world) {
    case 2: peer_allreduce_kernel<2><<<blocks, PAR_THREADS, 0, s>>>(b, world, first, end); break;
    case 4: peer_allreduce_kernel<4><<<blocks, PAR_THREADS, 0, s>>>(b, world, first, end); break;
    case 8: peer_allreduce_kernel<8><<<blocks, PAR_THREADS, 0, s>>>(b, world, first, end); break;
    default: peer_allreduce_kernel<0><<<blocks, PAR_THREADS, 0, s>>>(b, world, first, end); break;
  }
  count_launch();
}

}  // namespace lsr
