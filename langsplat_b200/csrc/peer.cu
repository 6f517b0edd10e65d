// peer.cu — the cross-GPU synchronisation of a view-sharded step.  One process per GPU; every rank owns
// a buffer allocated with lsr_peer_alloc whose first LSR_PEER_HEADER_BYTES hold the words below, and has the other
// ranks' buffers mapped (lsr_peer_open, CUDA IPC over NVLink / NVSwitch).  The barrier is the only ordering the
// exchange needs: after it, every rank's factors of this step are complete and visible, so the SH-gradient kernel
// (preprocess_bwd.cu: sh_grad_views_kernel) can read them straight out of peer memory.
#include "common.cuh"

namespace lsr {
namespace {

struct PeerHeader {
  uint32_t arrived[LSR_MAX_PEERS];  // arrived[r] = last epoch rank r signalled to the owner of this header
  uint32_t timed_out;               // set by the owner's barrier kernel when a peer did not show up in time
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
// harmless.  A peer that never arrives must not hang the GPU: after LSR_PEER_TIMEOUT_NS the wait gives up and raises
// timed_out in the own header (lsr_peer_status), the gradients of that step are then garbage and the caller aborts.
constexpr uint64_t LSR_PEER_TIMEOUT_NS = 4000000000ull;

__global__ void __launch_bounds__(32) peer_barrier_kernel(PeerHeaders hs, int rank, int world, uint32_t epoch) {
  const int r = threadIdx.x;
  __threadfence_system();
  if (r < world) {
    uint32_t* theirs = &hs.h[r]->arrived[rank];
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(epoch) : "memory");
    const uint32_t* mine = &hs.h[rank]->arrived[r];
    const uint64_t t0 = globaltimer_ns();
    for (;;) {
      uint32_t seen;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(mine) : "memory");
      if ((int32_t)(seen - epoch) >= 0) break;
      if (globaltimer_ns() - t0 > LSR_PEER_TIMEOUT_NS) {
        hs.h[rank]->timed_out = 1u;
        break;
      }
      __nanosleep(64);
    }
  }
  __syncwarp();
  __threadfence_system();
}

}  // namespace

void launch_peer_barrier(int rank, int world, void* const* headers, uint32_t epoch, cudaStream_t s) {
  PeerHeaders hs;
  for (int i = 0; i < LSR_MAX_PEERS; i++) hs.h[i] = i < world ? reinterpret_cast<PeerHeader*>(headers[i]) : nullptr;
  peer_barrier_kernel<<<1, 32, 0, s>>>(hs, rank, world, epoch);
  count_launch();
}

}  // namespace lsr
