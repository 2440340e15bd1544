// dist.h -- multi-GPU plumbing of the hot path: the reference's 2-D lon x lat domain decomposition
// (TRIOS::Domain, src/trios/TRIOS_Domain.C:81-195, 517-664) with NCCL over NVLink in place of Epetra Import/Export
// and Epetra_Comm::SumAll.  NCCL is bound at run time (dlopen of the libnccl the host process already uses).
#pragma once
#include <cuda_runtime.h>
#include "thcm_host.h"
#include "peer.cuh"

namespace thcm {
struct Dist;
Dist* dist_create(const Host& H, int rank, int nranks, const void* nccl_id);
void dist_destroy(Dist*);
int dist_unique_id(void* id128);
int dist_allreduce_sum(Dist*, double* d_buf, int count, cudaStream_t st);
int dist_halo_exchange(Dist*, double* d_x, cudaStream_t st);
int dist_halo_launches(Dist*);
// split form for overlap: begin() packs on `st` and runs the transfers on the communicator's own stream; end() makes `st`
// wait for the unpacked halo.  Work enqueued on `st` in between must not touch the halo part of d_x.
int dist_halo_begin(Dist*, double* d_x, cudaStream_t st);
int dist_halo_end(Dist*, cudaStream_t st);
// peer-memory path (peer.cuh): true when every rank mapped every other rank's box; dist_next_red() then hands out the
// argument block that makes a reducing kernel finish the sum across ranks itself (nranks == 0: not available, use
// dist_allreduce_sum after the kernel).  Every rank must request them in the same order.
bool dist_p2p(const Dist*);
PeerRed dist_next_red(Dist*);
int dist_peer_error(Dist*); // 0, or -1 when a peer wait timed out since the handle was created (message in dist_last_error)
const char* dist_last_error();
} // namespace thcm
