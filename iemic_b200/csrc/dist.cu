// dist.cu -- see dist.h.  NCCL over NVLink/NVSwitch for the two exchange steps of the hot path:
//   * halo exchange of a stored vector (state before an assembly, x before an SpMV): one gather kernel packs the
//     owned cells every peer needs (6 unknowns per cell, all levels) into one send buffer, a single NCCL group issues
//     the <= 8 point-to-point transfers, one scatter kernel writes the received cells into the halo ring.
//     This replaces Epetra_Import (TRIOS_Domain.C:527-560, Epetra_CrsMatrix::Apply's column-map import).
//   * sum-allreduce of a few doubles (dot products, norms): replaces Epetra_Comm::SumAll / MPI_Allreduce.
// NCCL is resolved with dlopen so that a single-GPU build has no link-time dependency on it.
#include "dist.h"
#include <dlfcn.h>
#include <cstdio>
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
enum { NCCL_FLOAT64 = 8, NCCL_SUM = 0 };
struct Nccl {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
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
  SYM(AllReduce, "ncclAllReduce") SYM(Send, "ncclSend") SYM(Recv, "ncclRecv") SYM(GroupStart, "ncclGroupStart")
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
};

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
  return d;
}

void dist_destroy(Dist* d) {
  if (!d) return;
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
int halo_pack(Dist* d, double* x, cudaStream_t st) {
  if (d->nself) { // my own periodic images
    halo_pack_kernel<<<(int)((6 * d->nself + 255) / 256), 256, 0, st>>>(x, d->d_self_src, (int)d->nself, d->d_sbuf);
    halo_unpack_kernel<<<(int)((6 * d->nself + 255) / 256), 256, 0, st>>>(x, d->d_self_dst, (int)d->nself, d->d_sbuf);
  }
  if (d->nsend) halo_pack_kernel<<<(int)((6 * d->nsend + 255) / 256), 256, 0, st>>>(x, d->d_send, (int)d->nsend, d->d_sbuf);
  return 0;
}
int halo_check(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { g_derr = std::string(what) + ": " + cudaGetErrorString(e); return -1; }
  return 0;
}
} // namespace

int dist_halo_exchange(Dist* d, double* x, cudaStream_t st) {
  halo_pack(d, x, st);
  if (halo_transfers(d, x, st)) return -1;
  return halo_check("halo exchange");
}
int dist_halo_begin(Dist* d, double* x, cudaStream_t st) {
  halo_pack(d, x, st);
  if (cudaEventRecord(d->ev_pack, st) != cudaSuccess || cudaStreamWaitEvent(d->cs, d->ev_pack, 0) != cudaSuccess) return halo_check("halo begin");
  if (halo_transfers(d, x, d->cs)) return -1;
  if (cudaEventRecord(d->ev_done, d->cs) != cudaSuccess) return halo_check("halo begin");
  return halo_check("halo begin");
}
int dist_halo_end(Dist* d, cudaStream_t st) {
  if (cudaStreamWaitEvent(st, d->ev_done, 0) != cudaSuccess) return halo_check("halo end");
  return 0;
}
int dist_halo_launches(Dist* d) { return (d->nsend ? 1 : 0) + (d->nrecv ? 1 : 0) + 1 + (d->nself ? 2 : 0); }
const char* dist_last_error() { return g_derr.c_str(); }

} // namespace thcm
