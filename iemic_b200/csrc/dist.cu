// dist.cu -- see dist.h.  NCCL over NVLink/NVSwitch for the two exchange steps of the hot path:
//   * halo exchange of a stored vector (state before an assembly, x before an SpMV): one gather kernel packs the
//     owned cells every peer needs (6 unknowns per cell, all levels) into one send buffer, a single NCCL group issues
//     the <= 8 point-to-point transfers, one scatter kernel writes the received cells into the halo ring.
//     This replaces Epetra_Import (TRIOS_Domain.C:527-560, Epetra_CrsMatrix::Apply's column-map import).
//   * sum-allreduce of a few doubles (dot products, norms): replaces Epetra_Comm::SumAll / MPI_Allreduce.
// When all ranks sit on distinct GPUs of one NVSwitch domain and CUDA IPC is available, both steps bypass NCCL and go
// through peer memory instead (peer.cuh): push/pull halo kernels and reductions finished inside the reducing kernel.
// NCCL then only bootstraps (handle exchange) and stays as the path for IEMIC_P2P=0 or boxes without peer access.
// NCCL is resolved with dlopen so that a single-GPU build has no link-time dependency on it.
#include "dist.h"
#include <dlfcn.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace thcm {
namespace {
std::string g_derr;

// the stable part of NCCL's C ABI that is used here
typedef void* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { NCCL_INT8 = 0, NCCL_FLOAT64 = 8, NCCL_SUM = 0 };
struct Nccl {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
} N;

bool load_nccl() {
  if (N.lib) return true;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    N.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (N.lib) break;
  }
  if (!N.lib) { g_derr = std::string("cannot load libnccl.so.2: ") + dlerror(); return false; }
#define SYM(field, name)                                                      \
  *(void**)(&N.field) = dlsym(N.lib, name);                                   \
  if (!N.field) { g_derr = std::string("NCCL symbol missing: ") + name; N.lib = nullptr; return false; }
  SYM(GetUniqueId, "ncclGetUniqueId") SYM(CommInitRank, "ncclCommInitRank") SYM(CommDestroy, "ncclCommDestroy")
  SYM(AllReduce, "ncclAllReduce") SYM(AllGather, "ncclAllGather") SYM(Send, "ncclSend") SYM(Recv, "ncclRecv") SYM(GroupStart, "ncclGroupStart")
  SYM(GroupEnd, "ncclGroupEnd") SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
  return true;
}
bool nccl_ok(ncclResult_t r, const char* what) {
  if (r == 0) return true;
  g_derr = std::string(what) + ": " + (N.GetErrorString ? N.GetErrorString(r) : "NCCL error");
  return false;
}

// one thread per (cell, unknown): coalesced on the buffer side, 48-byte runs on the vector side
__global__ void __launch_bounds__(256) halo_pack_kernel(const double* __restrict__ x, const int* __restrict__ cells, int ncells, double* __restrict__ buf) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < 6 * ncells) buf[t] = x[(size_t)6 * cells[t / 6] + (t % 6)];
}
__global__ void __launch_bounds__(256) halo_unpack_kernel(double* __restrict__ x, const int* __restrict__ cells, int ncells, const double* __restrict__ buf) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < 6 * ncells) x[(size_t)6 * cells[t / 6] + (t % 6)] = buf[t];
}

// ---- peer-memory halo exchange (peer.cuh) ----
__global__ void __launch_bounds__(256) halo_push_kernel(const double* __restrict__ x, const int* __restrict__ cells, int ncells, const PeerHalo p,
                                                        unsigned int* ticket) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < 6 * ncells) {
    const int c = t / 6, v = t - 6 * c;
    int q = 0, beg = 0;
    while (q < p.npeers - 1 && c >= p.send_end[q]) { beg = p.send_end[q]; q++; }
    p.rbuf[q][(size_t)6 * (c - beg) + v] = x[(size_t)6 * cells[c] + v];
  }
  // the block that finishes last raises my flag in every neighbour's box
  __shared__ bool is_last;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int k = atomicAdd(ticket, 1u);
    is_last = (k == gridDim.x - 1);
    if (is_last) *ticket = 0u;
  }
  __syncthreads();
  if (is_last && threadIdx.x < p.npeers) {
    __threadfence_system();
    peer_st_release(p.flag[threadIdx.x], p.epoch);
  }
}
__global__ void __launch_bounds__(256) halo_pull_kernel(double* __restrict__ x, const int* __restrict__ cells, int ncells, const double* rbuf,
                                                        const PeerHalo p) {
  if (threadIdx.x < p.npeers) peer_wait(p.wait[threadIdx.x], p.epoch, p.me);
  __syncthreads();
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < 6 * ncells) x[(size_t)6 * cells[t / 6] + (t % 6)] = peer_ld_f64(rbuf + t);
}

struct Boot { // what every rank tells the others once
  cudaIpcMemHandle_t handle;
  long long roff[PEER_MAX], nr[PEER_MAX]; // my receive segment (cells) per source rank
  long long nrecv, ok;
  char busid[32];
};
} // namespace

struct Dist {
  ncclComm_t comm = nullptr;
  int rank = 0, nranks = 1;
  struct Peer { int rank; size_t soff, ns, roff, nr; };
  std::vector<Peer> peers;
  size_t nsend = 0, nrecv = 0;     // cells
  int *d_send = nullptr, *d_recv = nullptr;
  double *d_sbuf = nullptr, *d_rbuf = nullptr;
  // cells exchanged with myself (periodic images when a rank is its own neighbour)
  size_t nself = 0; int *d_self_src = nullptr, *d_self_dst = nullptr;
  cudaStream_t cs = nullptr; cudaEvent_t ev_pack = nullptr, ev_done = nullptr; // overlap path
  // peer-memory path
  bool p2p = false;
  void* box_local = nullptr; void* box_remote[PEER_MAX] = {};
  long long peer_nrecv[PEER_MAX] = {}, peer_roff_me[PEER_MAX] = {}; // per rank: its receive buffer size, my segment in it
  unsigned long long halo_epoch = 0, ar_epoch = 0;
  unsigned int* d_ticket = nullptr;
  PeerHalo pending; double* pending_x = nullptr;
};

namespace {
// Map every rank's box; on any failure all ranks agree to stay on NCCL.
void p2p_setup(Dist* d) {
  const char* env = getenv("IEMIC_P2P");
  long long ok = !(env && env[0] == '0') && d->nranks <= PEER_MAX && d->peers.size() <= (size_t)PEER_MAX;
  const size_t box_bytes = sizeof(PeerBox) + 2 * 6 * d->nrecv * sizeof(double) + 64;
  Boot mine; std::memset(&mine, 0, sizeof(mine));
  for (int r = 0; r < PEER_MAX; r++) mine.roff[r] = -1;
  if (ok && (cudaMalloc(&d->box_local, box_bytes) != cudaSuccess || cudaMemset(d->box_local, 0, box_bytes) != cudaSuccess ||
             cudaMalloc((void**)&d->d_ticket, sizeof(unsigned int)) != cudaSuccess || cudaMemset(d->d_ticket, 0, sizeof(unsigned int)) != cudaSuccess ||
             cudaIpcGetMemHandle(&mine.handle, d->box_local) != cudaSuccess)) ok = 0;
  if (ok) for (const auto& p : d->peers) { mine.roff[p.rank] = (long long)p.roff; mine.nr[p.rank] = (long long)p.nr; }
  mine.nrecv = (long long)d->nrecv;
  int dev = 0; cudaGetDevice(&dev);
  if (cudaDeviceGetPCIBusId(mine.busid, sizeof(mine.busid), dev) != cudaSuccess) ok = 0;
  mine.ok = ok;
  cudaGetLastError();
  // exchange (NCCL bootstrap)
  std::vector<Boot> all(d->nranks);
  Boot *d_mine = nullptr, *d_all = nullptr;
  bool xok = cudaMalloc((void**)&d_mine, sizeof(Boot)) == cudaSuccess && cudaMalloc((void**)&d_all, sizeof(Boot) * d->nranks) == cudaSuccess &&
             cudaMemcpy(d_mine, &mine, sizeof(Boot), cudaMemcpyHostToDevice) == cudaSuccess && cudaDeviceSynchronize() == cudaSuccess &&
             N.AllGather(d_mine, d_all, sizeof(Boot), NCCL_INT8, d->comm, nullptr) == 0 && cudaStreamSynchronize(nullptr) == cudaSuccess &&
             cudaMemcpy(all.data(), d_all, sizeof(Boot) * d->nranks, cudaMemcpyDeviceToHost) == cudaSuccess;
  if (!xok) { ok = 0; for (auto& b : all) b.ok = 0; }
  for (int r = 0; r < d->nranks && ok; r++) {
    if (!all[r].ok) ok = 0;
    for (int q = 0; q < r && ok; q++) if (std::strncmp(all[r].busid, all[q].busid, sizeof(mine.busid)) == 0) ok = 0; // two ranks on one GPU
  }
  if (ok) {
    for (int r = 0; r < d->nranks && ok; r++) {
      d->peer_nrecv[r] = all[r].nrecv; d->peer_roff_me[r] = all[r].roff[d->rank];
      if (r == d->rank) { d->box_remote[r] = d->box_local; continue; }
      if (cudaIpcOpenMemHandle(&d->box_remote[r], all[r].handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { d->box_remote[r] = nullptr; ok = 0; }
    }
    for (const auto& p : d->peers) if (ok && (all[p.rank].nr[d->rank] != (long long)p.ns || all[p.rank].roff[d->rank] < 0) && p.ns) ok = 0;
    cudaGetLastError();
  }
  // agreement: everybody or nobody
  double agree = ok ? 1.0 : 0.0, *d_agree = (double*)d_mine;
  if (xok && cudaMemcpy(d_agree, &agree, sizeof(double), cudaMemcpyHostToDevice) == cudaSuccess &&
      N.AllReduce(d_agree, d_agree, 1, NCCL_FLOAT64, NCCL_SUM, d->comm, nullptr) == 0 && cudaStreamSynchronize(nullptr) == cudaSuccess &&
      cudaMemcpy(&agree, d_agree, sizeof(double), cudaMemcpyDeviceToHost) == cudaSuccess)
    d->p2p = ok && agree == (double)d->nranks;
  if (d->p2p) { // time-out of a wait on a peer's flag (nanoseconds), kept in the rank's own box
    const char* ts = getenv("IEMIC_PEER_TIMEOUT_S");
    const double sec = ts && atof(ts) > 0.0 ? atof(ts) : 120.0;
    const unsigned long long ns = (unsigned long long)(sec * 1.0e9);
    cudaMemcpy(&((PeerBox*)d->box_local)->pad[PEER_PAD_TIMEOUT_NS], &ns, sizeof(ns), cudaMemcpyHostToDevice);
  }
  if (d_mine) cudaFree(d_mine);
  if (d_all) cudaFree(d_all);
  cudaGetLastError();
  if (getenv("IEMIC_DEBUG")) fprintf(stderr, "[iemic_b200] rank %d: peer-memory exchange %s\n", d->rank, d->p2p ? "on" : "off (NCCL)");
}
void p2p_teardown(Dist* d) {
  cudaDeviceSynchronize();
  for (int r = 0; r < PEER_MAX; r++) if (d->box_remote[r] && r != d->rank) cudaIpcCloseMemHandle(d->box_remote[r]);
  if (d->box_local) cudaFree(d->box_local);
  if (d->d_ticket) cudaFree(d->d_ticket);
  cudaGetLastError();
}
PeerHalo peer_halo_args(Dist* d) {
  PeerHalo a; std::memset(&a, 0, sizeof(a));
  a.epoch = ++d->halo_epoch;
  const int b = (int)(a.epoch & 1ull);
  PeerBox* me = (PeerBox*)d->box_local;
  a.me = me;
  int i = 0;
  for (const auto& p : d->peers) {
    PeerBox* rb = (PeerBox*)d->box_remote[p.rank];
    double* rbuf = (double*)((char*)rb + sizeof(PeerBox));
    a.rbuf[i] = rbuf + (size_t)b * 6 * d->peer_nrecv[p.rank] + (size_t)6 * (d->peer_roff_me[p.rank] < 0 ? 0 : d->peer_roff_me[p.rank]);
    a.flag[i] = &rb->halo_flag[b][d->rank];
    a.wait[i] = &me->halo_flag[b][p.rank];
    a.send_end[i] = (int)(p.soff + p.ns);
    i++;
  }
  a.npeers = i;
  return a;
}
} // namespace

Dist* dist_create(const Host& H, int rank, int nranks, const void* nccl_id) {
  if (!nccl_id) { g_derr = "nccl_id is required when nranks > 1"; return nullptr; }
  if (!load_nccl()) return nullptr;
  Dist* d = new Dist();
  d->rank = rank; d->nranks = nranks;
  ncclUniqueId id;
  std::memcpy(&id, nccl_id, sizeof(id));
  if (!nccl_ok(N.CommInitRank(&d->comm, nranks, id, rank), "ncclCommInitRank")) { delete d; return nullptr; }
  std::vector<int> send, recv, self_src, self_dst;
  for (const auto& p : H.halo_plan()) {
    if (p.rank == rank) {
      self_src.insert(self_src.end(), p.send_cells.begin(), p.send_cells.end());
      self_dst.insert(self_dst.end(), p.recv_cells.begin(), p.recv_cells.end());
      continue;
    }
    Dist::Peer q{p.rank, send.size(), p.send_cells.size(), recv.size(), p.recv_cells.size()};
    send.insert(send.end(), p.send_cells.begin(), p.send_cells.end());
    recv.insert(recv.end(), p.recv_cells.begin(), p.recv_cells.end());
    d->peers.push_back(q);
  }
  d->nsend = send.size(); d->nrecv = recv.size(); d->nself = self_src.size();
  auto up = [](int** dp, const std::vector<int>& v) {
    if (v.empty()) return true;
    if (cudaMalloc((void**)dp, v.size() * sizeof(int)) != cudaSuccess) return false;
    return cudaMemcpy(*dp, v.data(), v.size() * sizeof(int), cudaMemcpyHostToDevice) == cudaSuccess;
  };
  bool ok = up(&d->d_send, send) && up(&d->d_recv, recv) && up(&d->d_self_src, self_src) && up(&d->d_self_dst, self_dst);
  ok = ok && cudaMalloc((void**)&d->d_sbuf, (6 * std::max<size_t>(d->nsend, d->nself) + 8) * sizeof(double)) == cudaSuccess;
  ok = ok && cudaMalloc((void**)&d->d_rbuf, (6 * d->nrecv + 8) * sizeof(double)) == cudaSuccess;
  ok = ok && cudaStreamCreateWithFlags(&d->cs, cudaStreamNonBlocking) == cudaSuccess;
  ok = ok && cudaEventCreateWithFlags(&d->ev_pack, cudaEventDisableTiming) == cudaSuccess;
  ok = ok && cudaEventCreateWithFlags(&d->ev_done, cudaEventDisableTiming) == cudaSuccess;
  if (!ok) { g_derr = "cudaMalloc failed in dist_create"; dist_destroy(d); return nullptr; }
  p2p_setup(d);
  return d;
}

void dist_destroy(Dist* d) {
  if (!d) return;
  p2p_teardown(d);
  if (d->comm && N.CommDestroy) N.CommDestroy(d->comm);
  void* ptrs[] = {d->d_send, d->d_recv, d->d_sbuf, d->d_rbuf, d->d_self_src, d->d_self_dst};
  for (void* p : ptrs) if (p) cudaFree(p);
  if (d->cs) cudaStreamDestroy(d->cs);
  if (d->ev_pack) cudaEventDestroy(d->ev_pack);
  if (d->ev_done) cudaEventDestroy(d->ev_done);
  delete d;
}

int dist_unique_id(void* id128) {
  if (!load_nccl()) return -1;
  ncclUniqueId id;
  if (!nccl_ok(N.GetUniqueId(&id), "ncclGetUniqueId")) return -1;
  std::memcpy(id128, &id, sizeof(id));
  return 0;
}

int dist_allreduce_sum(Dist* d, double* buf, int count, cudaStream_t st) {
  return nccl_ok(N.AllReduce(buf, buf, (size_t)count, NCCL_FLOAT64, NCCL_SUM, d->comm, st), "ncclAllReduce") ? 0 : -1;
}

namespace {
int halo_transfers(Dist* d, double* x, cudaStream_t st) { // NCCL group + unpack, all on `st`
  if (!nccl_ok(N.GroupStart(), "ncclGroupStart")) return -1;
  for (const auto& p : d->peers) {
    if (p.ns && !nccl_ok(N.Send(d->d_sbuf + 6 * p.soff, 6 * p.ns, NCCL_FLOAT64, p.rank, d->comm, st), "ncclSend")) return -1;
    if (p.nr && !nccl_ok(N.Recv(d->d_rbuf + 6 * p.roff, 6 * p.nr, NCCL_FLOAT64, p.rank, d->comm, st), "ncclRecv")) return -1;
  }
  if (!nccl_ok(N.GroupEnd(), "ncclGroupEnd")) return -1;
  if (d->nrecv) halo_unpack_kernel<<<(int)((6 * d->nrecv + 255) / 256), 256, 0, st>>>(x, d->d_recv, (int)d->nrecv, d->d_rbuf);
  return 0;
}
void halo_self(Dist* d, double* x, cudaStream_t st) { // my own periodic images
  if (!d->nself) return;
  halo_pack_kernel<<<(int)((6 * d->nself + 255) / 256), 256, 0, st>>>(x, d->d_self_src, (int)d->nself, d->d_sbuf);
  halo_unpack_kernel<<<(int)((6 * d->nself + 255) / 256), 256, 0, st>>>(x, d->d_self_dst, (int)d->nself, d->d_sbuf);
}
int halo_pack(Dist* d, double* x, cudaStream_t st) {
  halo_self(d, x, st);
  if (d->nsend) halo_pack_kernel<<<(int)((6 * d->nsend + 255) / 256), 256, 0, st>>>(x, d->d_send, (int)d->nsend, d->d_sbuf);
  return 0;
}
void peer_push(Dist* d, const double* x, const PeerHalo& a, cudaStream_t st) {
  const int nb = (int)((6 * d->nsend + 255) / 256);
  halo_push_kernel<<<nb < 1 ? 1 : nb, 256, 0, st>>>(x, d->d_send, (int)d->nsend, a, d->d_ticket);
}
void peer_pull(Dist* d, double* x, const PeerHalo& a, cudaStream_t st) {
  const int nb = (int)((6 * d->nrecv + 255) / 256);
  const double* rbuf = (const double*)((const char*)d->box_local + sizeof(PeerBox)) + (size_t)(a.epoch & 1ull) * 6 * d->nrecv;
  halo_pull_kernel<<<nb < 1 ? 1 : nb, 256, 0, st>>>(x, d->d_recv, (int)d->nrecv, rbuf, a);
}
int halo_check(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { g_derr = std::string(what) + ": " + cudaGetErrorString(e); return -1; }
  return 0;
}
} // namespace

int dist_halo_exchange(Dist* d, double* x, cudaStream_t st) {
  if (d->p2p) {
    halo_self(d, x, st);
    const PeerHalo a = peer_halo_args(d);
    peer_push(d, x, a, st);
    peer_pull(d, x, a, st);
    return halo_check("halo exchange (peer memory)");
  }
  halo_pack(d, x, st);
  if (halo_transfers(d, x, st)) return -1;
  return halo_check("halo exchange");
}
int dist_halo_begin(Dist* d, double* x, cudaStream_t st) {
  if (d->p2p) { // push now, pull in end(): kernels enqueued in between overlap the flight time of the neighbours' data
    halo_self(d, x, st);
    d->pending = peer_halo_args(d); d->pending_x = x;
    peer_push(d, x, d->pending, st);
    return halo_check("halo begin (peer memory)");
  }
  halo_pack(d, x, st);
  if (cudaEventRecord(d->ev_pack, st) != cudaSuccess || cudaStreamWaitEvent(d->cs, d->ev_pack, 0) != cudaSuccess) return halo_check("halo begin");
  if (halo_transfers(d, x, d->cs)) return -1;
  if (cudaEventRecord(d->ev_done, d->cs) != cudaSuccess) return halo_check("halo begin");
  return halo_check("halo begin");
}
int dist_halo_end(Dist* d, cudaStream_t st) {
  if (d->p2p) {
    peer_pull(d, d->pending_x, d->pending, st);
    return halo_check("halo end (peer memory)");
  }
  if (cudaStreamWaitEvent(st, d->ev_done, 0) != cudaSuccess) return halo_check("halo end");
  return 0;
}
int dist_halo_launches(Dist* d) { return d->p2p ? 2 + (d->nself ? 2 : 0) : (d->nsend ? 1 : 0) + (d->nrecv ? 1 : 0) + 1 + (d->nself ? 2 : 0); }
bool dist_p2p(const Dist* d) { return d && d->p2p; }
PeerRed dist_next_red(Dist* d) {
  PeerRed r; std::memset(&r, 0, sizeof(r));
  if (!d || !d->p2p) return r;
  for (int q = 0; q < d->nranks; q++) r.box[q] = (PeerBox*)d->box_remote[q];
  r.rank = d->rank; r.nranks = d->nranks; r.epoch = ++d->ar_epoch;
  return r;
}
// Did a wait on a peer's flag give up (peer.cuh)?  Call after a synchronisation of the stream the kernels ran on.
int dist_peer_error(Dist* d) {
  if (!d || !d->p2p) return 0;
  unsigned long long w = 0;
  if (cudaMemcpy(&w, &((PeerBox*)d->box_local)->pad[PEER_PAD_ERROR], sizeof(w), cudaMemcpyDeviceToHost) != cudaSuccess) { g_derr = "reading the peer error word failed"; return -1; }
  if (w) { g_derr = "a wait on a peer GPU's flag timed out (IEMIC_PEER_TIMEOUT_S): a rank died or fell more than the time-out behind; this handle is out of step with its peers, recreate it"; return -1; }
  return 0;
}
const char* dist_last_error() { return g_derr.c_str(); }

} // namespace thcm
