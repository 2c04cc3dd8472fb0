// d4b_comm.cu -- halo exchange between ranks (one process per GPU).
//
// Replaces d4est_ghost_data_exchange (src/Mesh/d4est_ghost_data.c:142-255: per
// neighbour rank an MPI_Isend of the mirror elements' nodal values and an
// MPI_Irecv into the ghost array) and the sc_allreduce calls of the solvers
// (src/Solver/d4est_solver_cg_eigs.c:181-245) by grouped NCCL point-to-point
// and a scalar all-reduce over NVLink.  NCCL is bound at run time with dlopen
// ("libnccl.so.2": the copy already loaded by the host process, e.g. PyTorch's,
// or the system one) so that single-rank users need no NCCL at all; a
// multi-rank mesh without an initialised communicator aborts.
#include <dlfcn.h>
#include <nccl.h>
#include <cstring>
#include "d4b_internal.h"

namespace {

struct NcclApi {
  void *handle = nullptr;
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRank) CommInitRank = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclSend) Send = nullptr;
  decltype(&ncclRecv) Recv = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
  decltype(&ncclAllReduce) AllReduce = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
};

NcclApi g_nccl;
ncclComm_t g_comm = nullptr;
int g_rank = 0, g_size = 1;

#define D4B_NCCL(call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) \
    D4B_ABORT("NCCL error %s: %s", #call, g_nccl.GetErrorString ? g_nccl.GetErrorString(r_) : "?"); } while (0)

void load_nccl() {
  if (g_nccl.handle) return;
  const char *names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char *n : names) {
    g_nccl.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (g_nccl.handle) break;
  }
  D4B_CHECK(g_nccl.handle, "multi-rank run requested but libnccl.so.2 cannot be loaded: %s", dlerror());
#define D4B_SYM(field, name) \
  g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(g_nccl.handle, name)); \
  D4B_CHECK(g_nccl.field, "libnccl: symbol %s missing", name)
  D4B_SYM(GetUniqueId, "ncclGetUniqueId");
  D4B_SYM(CommInitRank, "ncclCommInitRank");
  D4B_SYM(CommDestroy, "ncclCommDestroy");
  D4B_SYM(Send, "ncclSend");
  D4B_SYM(Recv, "ncclRecv");
  D4B_SYM(GroupStart, "ncclGroupStart");
  D4B_SYM(GroupEnd, "ncclGroupEnd");
  D4B_SYM(AllReduce, "ncclAllReduce");
  D4B_SYM(GetErrorString, "ncclGetErrorString");
#undef D4B_SYM
}

// one block per (peer, mirror) entry: copy the element's nodal values into the send buffer
__global__ void pack_kernel(const int64_t *__restrict__ src, const int64_t *__restrict__ dst,
                            const int32_t *__restrict__ cnt, const double *__restrict__ u,
                            double *__restrict__ sendbuf) {
  const int64_t s = src[blockIdx.x], d = dst[blockIdx.x];
  const int n = cnt[blockIdx.x];
  for (int i = threadIdx.x; i < n; i += blockDim.x) sendbuf[d + i] = u[s + i];
}

}  // namespace

extern "C" void d4b_comm_get_unique_id(void *id_out) {
  load_nccl();
  static_assert(sizeof(ncclUniqueId) == D4B_COMM_ID_BYTES, "ncclUniqueId size");
  ncclUniqueId id;
  D4B_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(id_out, &id, sizeof(id));
}

extern "C" void d4b_comm_init(const void *id_in, int rank, int size) {
  D4B_CHECK(size >= 1 && rank >= 0 && rank < size, "d4b_comm_init: bad rank %d / size %d", rank, size);
  if (g_comm) d4b_comm_destroy();
  g_rank = rank; g_size = size;
  if (size == 1) return;
  load_nccl();
  ncclUniqueId id;
  memcpy(&id, id_in, sizeof(id));
  D4B_NCCL(g_nccl.CommInitRank(&g_comm, size, id, rank));
}

extern "C" void d4b_comm_destroy(void) {
  if (g_comm) { g_nccl.CommDestroy(g_comm); g_comm = nullptr; }
  g_rank = 0; g_size = 1;
}

extern "C" int d4b_comm_size(void) { return g_size; }
extern "C" int d4b_comm_rank(void) { return g_rank; }

// ---- exchange plan ------------------------------------------------------------
void d4b_comm_setup(d4b_mesh *m, const d4b_mesh_desc_t *d) {
  m->mpirank = d->mpirank; m->mpisize = d->mpisize > 0 ? d->mpisize : 1;
  if (m->mpisize == 1) {
    D4B_CHECK(d->n_ghost == 0, "single-rank mesh with %d ghost elements", d->n_ghost);
    return;
  }
  D4B_CHECK(d->mirror_proc_offsets && d->ghost_proc_offsets && (d->n_mirrors == 0 || d->mirror_elem),
            "multi-rank mesh without an exchange plan");
  std::vector<int64_t> src, dst;
  std::vector<int32_t> cnt;
  int64_t send_total = 0;
  for (int q = 0; q < m->mpisize; ++q) {
    D4bPeer p;
    p.rank = q;
    p.send_off = send_total;
    for (int i = d->mirror_proc_offsets[q]; i < d->mirror_proc_offsets[q + 1]; ++i) {
      const int e = d->mirror_elem[d->mirror_proc_mirrors[i]];
      D4B_CHECK(e >= 0 && e < d->n_local, "mirror element out of range");
      const int n = (d->deg[e] + 1) * (d->deg[e] + 1) * (d->deg[e] + 1);
      src.push_back(d->nodal_stride[e]); dst.push_back(send_total); cnt.push_back(n);
      send_total += n;
    }
    p.send_cnt = send_total - p.send_off;
    const int g0 = d->ghost_proc_offsets[q], g1 = d->ghost_proc_offsets[q + 1];
    p.recv_off = (g0 < d->n_ghost) ? d->ghost_nodal_stride[g0] : d->ghost_nodes;
    p.recv_cnt = 0;
    for (int g = g0; g < g1; ++g) {
      const int n = (d->ghost_deg[g] + 1) * (d->ghost_deg[g] + 1) * (d->ghost_deg[g] + 1);
      // the receive lands in place: ghosts of one rank must be contiguous in ghost-id order
      D4B_CHECK(d->ghost_nodal_stride[g] == p.recv_off + p.recv_cnt, "ghost buffer is not contiguous per rank");
      p.recv_cnt += n;
    }
    D4B_CHECK(q != m->mpirank || (p.send_cnt == 0 && p.recv_cnt == 0), "rank exchanges with itself");
    if (p.send_cnt > 0 || p.recv_cnt > 0) m->peers.push_back(p);
  }
  m->n_send_elems = (int32_t)src.size();
  m->d_send_src = d4b_dev_upload<int64_t>(m, src.data(), src.size());
  m->d_send_dst = d4b_dev_upload<int64_t>(m, dst.data(), dst.size());
  m->d_send_cnt = d4b_dev_upload<int32_t>(m, cnt.data(), cnt.size());
  m->d_sendbuf = d4b_dev_alloc<double>(m, (size_t)send_total);
  // highest priority: the halo and the few elements that wait for it are scheduled into
  // the SM slots the interior launch frees, instead of queueing behind it
  int prio_lo = 0, prio_hi = 0;
  D4B_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
  D4B_CUDA(cudaStreamCreateWithPriority(&m->comm_stream, cudaStreamNonBlocking, prio_hi));
  D4B_CUDA(cudaEventCreateWithFlags(&m->ev_interp, cudaEventDisableTiming));
  D4B_CUDA(cudaEventCreateWithFlags(&m->ev_boundary, cudaEventDisableTiming));
  D4B_CUDA(cudaEventCreateWithFlags(&m->ev_packed, cudaEventDisableTiming));
  D4B_CUDA(cudaEventCreateWithFlags(&m->ev_recv, cudaEventDisableTiming));
}

void d4b_comm_teardown(d4b_mesh *m) {
  if (m->comm_stream) cudaStreamDestroy(m->comm_stream);
  if (m->ev_packed) cudaEventDestroy(m->ev_packed);
  if (m->ev_recv) cudaEventDestroy(m->ev_recv);
  if (m->ev_interp) cudaEventDestroy(m->ev_interp);
  if (m->ev_boundary) cudaEventDestroy(m->ev_boundary);
}

// pack on the compute stream, exchange on the communication stream; the caller
// makes the compute stream wait for ev_recv before it touches ghost data
void d4b_comm_exchange_begin(d4b_mesh *m, const double *d_u) {
  if (m->mpisize == 1) return;
  D4B_CHECK(g_comm && g_size == m->mpisize && g_rank == m->mpirank,
            "mesh of rank %d/%d but d4b_comm_init was %s", m->mpirank, m->mpisize,
            g_comm ? "called with another rank/size" : "not called");
  if (m->n_send_elems > 0) {
    pack_kernel<<<m->n_send_elems, 128, 0, m->stream>>>(m->d_send_src, m->d_send_dst, m->d_send_cnt, d_u, m->d_sendbuf);
    D4B_CUDA(cudaGetLastError());
  }
  D4B_CUDA(cudaEventRecord(m->ev_packed, m->stream));
  D4B_CUDA(cudaStreamWaitEvent(m->comm_stream, m->ev_packed, 0));
  D4B_NCCL(g_nccl.GroupStart());
  for (const D4bPeer &p : m->peers) {
    if (p.send_cnt > 0)
      D4B_NCCL(g_nccl.Send(m->d_sendbuf + p.send_off, (size_t)p.send_cnt, ncclDouble, p.rank, g_comm, m->comm_stream));
    if (p.recv_cnt > 0)
      D4B_NCCL(g_nccl.Recv(m->d_ughost + p.recv_off, (size_t)p.recv_cnt, ncclDouble, p.rank, g_comm, m->comm_stream));
  }
  D4B_NCCL(g_nccl.GroupEnd());
  D4B_CUDA(cudaEventRecord(m->ev_recv, m->comm_stream));
}

void d4b_comm_exchange_wait(d4b_mesh *m) {
  if (m->mpisize == 1) return;
  D4B_CUDA(cudaStreamWaitEvent(m->stream, m->ev_recv, 0));
}

// sum n doubles (n <= 4) over the ranks, in place on the device, on the mesh's stream
void d4b_comm_allreduce_sum(d4b_mesh *m, double *d_x, int n) {
  if (m->mpisize == 1) return;
  D4B_CHECK(g_comm, "d4b_comm_init was not called");
  D4B_NCCL(g_nccl.AllReduce(d_x, d_x, (size_t)n, ncclDouble, ncclSum, g_comm, m->stream));
}
