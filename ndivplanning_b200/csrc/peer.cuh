// Peer window: every rank of one node owns a cudaMalloc'ed window that every other rank maps through CUDA IPC, so
// kernels exchange results with plain loads / stores over NVLink and a flag word per (slot, source rank) instead of a
// library collective (NCCL's all-gather costs 15-20 us of latency for the 16 KB .. 0.5 MB these paths exchange).
//
//   window = [ header 4 KB : flags[kWinSlots][kMaxPeers] u32 | error word ] [ payload ]
//
// Protocol ("epochs"): every collective entry point bumps its sequence number ("epoch") on the host (all ranks make the
// same calls in the same order, like any collective).  A producer finishes its stores, fences at system scope and
// stores the epoch into flags[slot][its rank] of EVERY rank's window; a consumer spins on its OWN window's flag
// (local memory) until it reaches the epoch, then reads the producer's payload through the mapped pointer.  Spins are
// bounded (kSpinLimit cycles, ~30 s): on expiry the consumer records an error word the host reports at the next call, and
// carries on, so a missing peer cannot hang the GPU.  Payloads are double-buffered by epoch parity by the callers.
#pragma once
#include <stdint.h>
#include "common.cuh"

namespace ndp {

constexpr int kMaxPeers = 8;
constexpr int kWinSlots = 16;
constexpr size_t kWinHeader = 4096;
constexpr size_t kWinErrOff = 2048;                 // u32 error word inside the header
constexpr long long kSpinLimit = 60000000000ll;     // ~30 s of SM clocks: a hang guard (a dead peer), not a latency bound —
                                                    // ranks may legitimately be seconds apart (start-up, data loading)

struct PeerWin {
  char* peer[kMaxPeers];                            // base of every rank's window as mapped in THIS process
  int rank, world;
};

#ifdef __CUDACC__
__device__ __forceinline__ unsigned* win_flag(char* base, int slot, int src) {
  return reinterpret_cast<unsigned*>(base) + slot * kMaxPeers + src;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// payload reads through a peer mapping: system-scope relaxed loads (never served from a stale L1 line)
__device__ __forceinline__ float ld_peer(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer4(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_peer_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
// one WARP, after the block's (or the grid's last block's) results are written: publish `epoch` to every rank
__device__ __forceinline__ void peer_signal_all(const PeerWin& w, int slot, unsigned epoch) {
  __threadfence_system();
  const int lane = threadIdx.x & 31;
  if (lane < w.world) st_release_sys(win_flag(w.peer[lane], slot, w.rank), epoch);
}
// one THREAD: wait until rank `src` has published `epoch` (or later) in `slot`; false on timeout
__device__ __forceinline__ bool peer_wait(const PeerWin& w, int slot, int src, unsigned epoch) {
  const unsigned* f = win_flag(w.peer[w.rank], slot, src);
  const long long t0 = clock64();
  while ((int)(ld_acquire_sys(f) - epoch) < 0) {
    if (clock64() - t0 > kSpinLimit) {
      atomicExch(reinterpret_cast<unsigned*>(w.peer[w.rank] + kWinErrOff), 1u + (unsigned)src);
      return false;
    }
    __nanosleep(64);
  }
  return true;
}
#endif  // __CUDACC__

// host side (comm.cu)
struct ncclComm;
}  // namespace ndp

struct ndp_comm {
  ndp::ncclComm* comm;
  int rank, world;
  char* win;                       // this rank's window (cudaMalloc), nullptr until first use
  char* peer[ndp::kMaxPeers];      // mapped windows (peer[rank] == win)
  size_t win_bytes;                // header + payload
  unsigned epoch;                  // sequence number of the windowed ndiv calls (payload parity = epoch & 1)
  unsigned epoch_rec;              // sequence number of the windowed plan-record exchanges (slot = epoch % 8)
  int win_state;                   // 0 = untried, 1 = mapped, -1 = peers not reachable (callers use NCCL)
};

namespace ndp {
// Collective: make sure every rank's window holds `payload` bytes (re-allocates and re-maps when it grows).
// Returns NDP_OK with *w filled when the window path is usable, NDP_OK with w->world == 0 when the ranks are not
// peer-reachable (callers fall back to the NCCL collectives), or an error.
int comm_window(ndp_comm* c, size_t payload, cudaStream_t s, PeerWin* w);
// error word of the local window (non-zero after a spin timed out); clears it
int comm_window_check(ndp_comm* c, cudaStream_t s);
}  // namespace ndp
