// peer.cuh -- exchange steps of the hot path done by the compute kernels themselves over NVLink peer memory.
//
// Every rank owns one "box" in device memory that its peers map through CUDA IPC (dist.cu).  Two exchange steps use it:
//
//  * sum-allreduce of <= PEER_ARMAX doubles, fused into the reduction kernels (gmres.cu): the block that finishes the
//    local reduction stores its result into slot [rank] of every peer's box, raises a flag there, waits for the flags
//    of all peers in its own box and adds the slots in rank order.  Same summation order on every rank, so all ranks
//    hold bit-identical results; one kernel instead of kernel + ncclAllReduce (Epetra_Comm::SumAll in the reference).
//  * halo exchange of a stored vector: a push kernel gathers the owned cells each neighbour needs and stores them
//    straight into the neighbour's receive buffer, the last block raises the flags; the pull kernel waits for the
//    neighbours' flags and scatters the received cells into the halo ring (Epetra_Import in the reference).
//
// Flags carry a monotonically increasing epoch; slots and buffers are double-buffered by epoch parity, which is enough
// because a rank can only complete exchange e+1 after every peer has finished reading exchange e (stream order).
// A wait that lasts longer than the time-out (box word pad[1], nanoseconds; IEMIC_PEER_TIMEOUT_S, default 120 s) gives up: it
// raises the error word pad[0] of the rank's own box and lets the kernel finish with meaningless numbers; the host turns the
// word into IEMIC_B200_ENCCL at its next synchronisation (dist_peer_error).  No __trap(): the CUDA context survives, so a caller
// like Continuation::reset() keeps working; the handle itself is out of step with its peers from then on and must be recreated.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace thcm {

constexpr int PEER_MAX = 8;          // ranks of one NVSwitch domain handled here
constexpr int PEER_ARMAX = 512;      // doubles per fused allreduce (= MAXV of gmres.cu)
enum { PEER_PAD_ERROR = 0, PEER_PAD_TIMEOUT_NS = 1 }; // words of PeerBox::pad

struct PeerBox {
  unsigned long long ar_flag[2][PEER_MAX];
  unsigned long long halo_flag[2][PEER_MAX];
  unsigned long long pad[32];
  double ar_val[2][PEER_MAX][PEER_ARMAX];
  // the halo receive buffers (2 x 6*nrecv doubles) follow at byte offset sizeof(PeerBox)
};

// argument block of a fused allreduce; nranks == 0 switches it off
struct PeerRed {
  PeerBox* box[PEER_MAX]; // box[rank] is the local one
  int rank, nranks;
  unsigned long long epoch;
};

// argument block of the halo push / pull kernels
struct PeerHalo {
  double* rbuf[PEER_MAX];                 // per neighbour q: where my cells go in q's receive buffer (this epoch's half)
  unsigned long long* flag[PEER_MAX];     // per neighbour q: q's halo_flag[parity][my rank]
  int send_end[PEER_MAX];                 // prefix ends (cells) of the per-neighbour segments of the send list
  const unsigned long long* wait[PEER_MAX]; // my own halo_flag[parity][q's rank]
  PeerBox* me;                            // my own box (error word, time-out)
  int npeers;
  unsigned long long epoch;
};

#ifdef __CUDACC__
__device__ __forceinline__ void peer_st_release(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long peer_ld_acquire(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double peer_ld_f64(const double* p) { // never served from a stale L1 line
  double v;
  asm volatile("ld.volatile.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long peer_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void peer_wait(const unsigned long long* flag, unsigned long long epoch, PeerBox* me) {
  if (peer_ld_acquire(flag) >= epoch) return;
  const unsigned long long t0 = peer_now(), limit = me->pad[PEER_PAD_TIMEOUT_NS];
  while (peer_ld_acquire(flag) < epoch)
    if (peer_now() - t0 > limit) { atomicExch(&me->pad[PEER_PAD_ERROR], 1ull); __threadfence_system(); return; }
}

// Called by every thread of ONE block after out[0..nv) holds this rank's sums; on return out[] holds the global sums.
__device__ __forceinline__ void peer_allreduce_block(const PeerRed& p, double* out, int nv) {
  const int tid = threadIdx.x, nt = blockDim.x, b = (int)(p.epoch & 1ull);
  __syncthreads();
  for (int k = tid; k < nv; k += nt) {
    const double v = out[k];
    for (int r = 0; r < p.nranks; r++) p.box[r]->ar_val[b][p.rank][k] = v;
  }
  __threadfence_system();
  __syncthreads();
  if (tid < p.nranks) {
    peer_st_release(&p.box[tid]->ar_flag[b][p.rank], p.epoch);
    peer_wait(&p.box[p.rank]->ar_flag[b][tid], p.epoch, p.box[p.rank]);
  }
  __syncthreads();
  const PeerBox* me = p.box[p.rank];
  for (int k = tid; k < nv; k += nt) {
    double s = 0.0;
    for (int r = 0; r < p.nranks; r++) s += peer_ld_f64(&me->ar_val[b][r][k]);
    out[k] = s;
  }
}
#endif

} // namespace thcm
