// Peer exchange of the per-shard results over NVLink: the "all-gather" of the sharded step done
// by this library's own kernels on memory the ranks map into each other with CUDA IPC.
//
// Every rank owns one allocation [2 parities][world][slot_bytes] plus a flag word per
// (parity, source rank).  k_xchg_push stores this rank's record into slot[rank] of EVERY rank's
// buffer (plain 16-byte stores through the NVSwitch; a store needs no answer, so the 1.25 MiB of
// config 4 leave at link speed instead of paying a round trip per load), and the last CTA to finish
// raises flag[parity][rank] = epoch on every peer after a system-scope fence.  k_xchg_wait then
// spins, one thread per source rank, until the local flags show the epoch; the merge kernel that
// follows on the stream reads the local buffer only.  Two parities make the buffer of step e free
// again before any peer can push step e+2: a peer pushes e+2 only after its own wait for e+1, which
// needs this rank's push of e+1, which this rank's stream orders after its merge of e.
//
// No multicast, no TMA: plain st.global / ld.acquire.sys on peer-mapped memory.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#define DSM_XCHG_MAX_WORLD 32

struct XchgPeers {
  char* data[DSM_XCHG_MAX_WORLD];      // every rank's buffer base (own included), as mapped here
  uint32_t* flags[DSM_XCHG_MAX_WORLD]; // every rank's flag array [2][world]
};

struct XchgPushParams {
  XchgPeers peers;
  const uint4* src;
  uint32_t n_vec;       // 16-byte units to send
  uint32_t world, rank, parity, epoch;
  uint64_t slot_bytes;
  uint32_t* ticket;     // local: CTAs that finished this push
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// grid = (chunks, world): CTA (c, d) copies chunk c of the record into rank d's slot[rank].
__global__ void __launch_bounds__(256) k_xchg_push(XchgPushParams p) {
  const uint32_t d = blockIdx.y;
  uint4* dst = reinterpret_cast<uint4*>(p.peers.data[d] + ((uint64_t)p.parity * p.world + p.rank) * p.slot_bytes);
  // the source is read once per destination; after the first pass it comes from L2
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < p.n_vec; i += gridDim.x * blockDim.x)
    dst[i] = __ldg(p.src + i);
  __threadfence_system();
  __syncthreads();
  __shared__ bool last;
  if (threadIdx.x == 0) last = atomicAdd(p.ticket, 1u) == gridDim.x * gridDim.y - 1;
  __syncthreads();
  if (!last) return;
  if (threadIdx.x == 0) *p.ticket = 0;  // ready for the next push (stream-ordered)
  __threadfence_system();
  if (threadIdx.x < p.world)
    st_release_sys(p.peers.flags[threadIdx.x] + p.parity * p.world + p.rank, p.epoch);
}

// One CTA, one thread per source rank.  Gives up after `timeout_ns` (a peer that died must not
// leave this GPU spinning) and records the rank it was waiting for in *err (0 = fine).
__global__ void k_xchg_wait(const uint32_t* flags, uint32_t world, uint32_t parity, uint32_t epoch,
                            unsigned long long timeout_ns, uint32_t* err) {
  const uint32_t r = threadIdx.x;
  if (r >= world) return;
  const uint32_t* f = flags + parity * world + r;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  // a flag only ever moves forward (e, e+2, ...); "not yet" = still behind this epoch
  while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (t - t0 > timeout_ns) {
      atomicMax(err, r + 1);
      return;
    }
    __nanosleep(200);
  }
}
