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
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
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
  decltype(&ncclAllGather) AllGather = nullptr;
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
  D4B_SYM(AllGather, "ncclAllGather");
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

static void d4b_comm_release_allreduce();
extern "C" void d4b_comm_destroy(void) {
  d4b_comm_release_allreduce();
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
    p.send_elem_off = (int32_t)src.size();
    for (int i = d->mirror_proc_offsets[q]; i < d->mirror_proc_offsets[q + 1]; ++i) {
      const int e = d->mirror_elem[d->mirror_proc_mirrors[i]];
      D4B_CHECK(e >= 0 && e < d->n_local, "mirror element out of range");
      const int n = (d->deg[e] + 1) * (d->deg[e] + 1) * (d->deg[e] + 1);
      src.push_back(d->nodal_stride[e]); dst.push_back(send_total); cnt.push_back(n);
      send_total += n;
    }
    p.send_cnt = send_total - p.send_off;
    p.send_elem_cnt = (int32_t)src.size() - p.send_elem_off;
    const int g0 = d->ghost_proc_offsets[q], g1 = d->ghost_proc_offsets[q + 1];
    p.recv_off = (g0 < d->n_ghost) ? d->ghost_nodal_stride[g0] : d->ghost_nodes;
    p.recv_cnt = 0;
    p.ghost_first = g0; p.ghost_cnt = g1 - g0;
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

namespace {
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

}  // namespace

// ---- all-reduce of a few doubles over peer memory ------------------------------------------
// The solvers' dot products end in a sum of ONE double over the ranks (sc_allreduce in the reference,
// d4est_solver_cg_eigs.c:181-245), two per CG iteration on every multigrid level: pure latency.  On one node every
// rank maps a small slot array of every other rank (CUDA IPC) and one kernel does the whole exchange: it writes this
// rank's values into slot [rank] of every rank -- each 8-byte store carries 4 bytes of data and the 4-byte epoch
// of the call, so a value is complete when both of its words show the epoch (no flag, no fence) -- then polls its
// own R slots and adds them in rank order: every rank gets the same bits.  Two slot sets alternate: a rank can only
// be two calls ahead of a peer that has not yet read.  Falls back to ncclAllReduce when the mappings cannot be made.
namespace {

constexpr int kArMax = 4;   // doubles per call

struct PeerAllreduce {
  int state = 0;                         // 0 not tried, 1 active, -1 unavailable
  unsigned long long *slots = nullptr;   // [2][R][kArMax][2]
  unsigned long long **d_peer = nullptr; // [R] slot arrays of all ranks (own included)
  std::vector<void *> mapped;
  unsigned *d_epoch = nullptr;           // call counter, advanced by the kernel itself (a replayed CUDA graph of a
                                         // V-cycle must not carry the epochs of the cycle it was captured from)
  cudaStream_t last_stream = nullptr;    // calls must execute in issue order: a call on another stream waits for the last
  cudaEvent_t ev = nullptr;
} g_ar;

__global__ void peer_allreduce_k(double *__restrict__ x, int n, unsigned long long *__restrict__ mine,
                                 unsigned long long *const *__restrict__ peer, int rank, int R,
                                 unsigned *__restrict__ epoch_ctr) {
  __shared__ double got[64 * kArMax];
  const int t = threadIdx.x;
  unsigned epoch = *epoch_ctr + 1;
  if (epoch == 0) epoch = 1;   // 0 marks an empty slot
  const size_t set = (size_t)(epoch & 1) * R * kArMax * 2;
  if (t < R * n) {
    const int q = t / n, j = t % n;
    const unsigned long long bits = (unsigned long long)__double_as_longlong(x[j]);
    unsigned long long *dst = peer[q] + set + ((size_t)rank * kArMax + j) * 2;
    const unsigned long long w0 = (bits & 0xffffffffull) | ((unsigned long long)epoch << 32);
    const unsigned long long w1 = (bits >> 32) | ((unsigned long long)epoch << 32);
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(dst), "l"(w0) : "memory");
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(dst + 1), "l"(w1) : "memory");
    const unsigned long long *src = mine + set + ((size_t)q * kArMax + j) * 2;
    const unsigned long long t0 = global_ns();
    unsigned long long a, b;
    for (;;) {
      asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(a) : "l"(src) : "memory");
      asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(b) : "l"(src + 1) : "memory");
      if ((unsigned)(a >> 32) == epoch && (unsigned)(b >> 32) == epoch) break;
      if (global_ns() - t0 > 30000000000ull) {
        printf("[d4b] rank %d: all-reduce %u timed out waiting for rank %d\n", rank, epoch, q);
        __trap();
      }
    }
    got[q * kArMax + j] = __longlong_as_double((long long)((a & 0xffffffffull) | (b << 32)));
  }
  __syncthreads();
  if (t == 0) *epoch_ctr = epoch;
  if (t < n) {
    double s = 0.;
    for (int q = 0; q < R; ++q) s += got[q * kArMax + t];
    x[t] = s;
  }
}

void peer_allreduce_release() {
  for (void *p : g_ar.mapped) cudaIpcCloseMemHandle(p);
  g_ar.mapped.clear();
  if (g_ar.slots) { cudaFree(g_ar.slots); g_ar.slots = nullptr; }
  if (g_ar.d_peer) { cudaFree(g_ar.d_peer); g_ar.d_peer = nullptr; }
  if (g_ar.d_epoch) { cudaFree(g_ar.d_epoch); g_ar.d_epoch = nullptr; }
  if (g_ar.ev) { cudaEventDestroy(g_ar.ev); g_ar.ev = nullptr; }
  g_ar.last_stream = nullptr;
  cudaGetLastError();
}

// collective; every rank reaches it in its first all-reduce
bool peer_allreduce_ready(cudaStream_t st) {
  if (g_ar.state) return g_ar.state > 0;
  g_ar.state = -1;
  const char *env = getenv("D4B_PEER_ALLREDUCE");
  const int R = g_size;
  int ok = !(env && atoi(env) == 0) && R <= 64;
  struct X { cudaIpcMemHandle_t h; int64_t ok; } mine;
  memset(&mine, 0, sizeof(mine));
  const size_t bytes = (size_t)2 * R * kArMax * 2 * sizeof(unsigned long long);
  if (ok) ok = cudaMalloc(&g_ar.slots, bytes) == cudaSuccess;
  if (ok) {
    D4B_CUDA(cudaMemsetAsync(g_ar.slots, 0, bytes, st));   // epoch 0 is never used
    ok = cudaIpcGetMemHandle(&mine.h, g_ar.slots) == cudaSuccess;
  }
  cudaGetLastError();
  mine.ok = ok;
  char *d_x = nullptr;
  D4B_CUDA(cudaMalloc(&d_x, (size_t)(R + 1) * sizeof(X)));
  D4B_CUDA(cudaMemcpyAsync(d_x + (size_t)R * sizeof(X), &mine, sizeof(mine), cudaMemcpyHostToDevice, st));
  D4B_NCCL(g_nccl.AllGather(d_x + (size_t)R * sizeof(X), d_x, sizeof(X), ncclChar, g_comm, st));
  std::vector<X> all(R);
  D4B_CUDA(cudaMemcpyAsync(all.data(), d_x, (size_t)R * sizeof(X), cudaMemcpyDeviceToHost, st));
  D4B_CUDA(cudaStreamSynchronize(st));
  for (int q = 0; q < R; ++q) ok = ok && all[q].ok;
  std::vector<unsigned long long *> ptr(R, nullptr);
  for (int q = 0; q < R && ok; ++q) {
    if (q == g_rank) { ptr[q] = g_ar.slots; continue; }
    void *a = nullptr;
    ok = cudaIpcOpenMemHandle(&a, all[q].h, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
    if (ok) { g_ar.mapped.push_back(a); ptr[q] = static_cast<unsigned long long *>(a); }
  }
  cudaGetLastError();
  // agree (over NCCL, which always works here)
  int64_t v = ok, *d_v = reinterpret_cast<int64_t *>(d_x);
  D4B_CUDA(cudaMemcpyAsync(d_v, &v, sizeof(v), cudaMemcpyHostToDevice, st));
  D4B_NCCL(g_nccl.AllReduce(d_v, d_v, 1, ncclInt64, ncclMin, g_comm, st));
  D4B_CUDA(cudaMemcpyAsync(&v, d_v, sizeof(v), cudaMemcpyDeviceToHost, st));
  D4B_CUDA(cudaStreamSynchronize(st));
  cudaFree(d_x);
  if (!v) { peer_allreduce_release(); return false; }
  D4B_CUDA(cudaMalloc(&g_ar.d_peer, (size_t)R * sizeof(void *)));
  D4B_CUDA(cudaMemcpy(g_ar.d_peer, ptr.data(), (size_t)R * sizeof(void *), cudaMemcpyHostToDevice));
  D4B_CUDA(cudaEventCreateWithFlags(&g_ar.ev, cudaEventDisableTiming));
  D4B_CUDA(cudaMalloc(&g_ar.d_epoch, sizeof(unsigned)));
  D4B_CUDA(cudaMemset(g_ar.d_epoch, 0, sizeof(unsigned)));
  g_ar.last_stream = nullptr;
  g_ar.state = 1;
  return true;
}

}  // namespace

// sum n doubles (n <= 4) over the ranks, in place on the device, on the mesh's stream
void d4b_comm_allreduce_sum(d4b_mesh *m, double *d_x, int n) {
  if (m->mpisize == 1) return;
  D4B_CHECK(g_comm, "d4b_comm_init was not called");
  D4B_CHECK(n >= 1 && n <= kArMax, "all-reduce of %d doubles", n);
  if (peer_allreduce_ready(m->stream)) {
    if (g_ar.last_stream && g_ar.last_stream != m->stream) D4B_CUDA(cudaStreamWaitEvent(m->stream, g_ar.ev, 0));
    peer_allreduce_k<<<1, 64 * kArMax, 0, m->stream>>>(d_x, n, g_ar.slots, g_ar.d_peer, g_rank, g_size, g_ar.d_epoch);
    D4B_CUDA(cudaGetLastError());
    D4B_CUDA(cudaEventRecord(g_ar.ev, m->stream));
    g_ar.last_stream = m->stream;
    return;
  }
  D4B_NCCL(g_nccl.AllReduce(d_x, d_x, (size_t)n, ncclDouble, ncclSum, g_comm, m->stream));
}

// ---- halo over peer memory -------------------------------------------------------
// No counterpart in the reference (its exchange is MPI point-to-point of nodal values,
// d4est_ghost_data.c:142-255).  On one node every rank maps the Gauss-collocated vector uhat of
// its neighbour ranks (cudaIpc*, NVLink) and the kernels of its boundary elements read the
// neighbours' lines where they lie: pack kernel, send/recv and the interpolation of the ghost
// elements leave the critical path of an apply.  Synchronisation per apply k:
//   every rank   : uhat of its own elements -> copy (k & 1) of its buffer; flags[me] on every neighbour := k
//   before the boundary elements: wait until flags[q] >= k for every neighbour q
// Two copies suffice: a rank can only reach apply k + 2 (and overwrite copy k & 1) after its
// neighbours signalled k + 1, which they do after their boundary elements of apply k are done.
namespace {

// the epoch is a device counter advanced here, once per apply (or smoother sweep): waits read it back.  Kernel
// arguments would freeze the epochs of the V-cycle a CUDA graph was captured from into its replays.
__global__ void peer_signal_k(unsigned long long *const *__restrict__ dst, int n, unsigned long long *__restrict__ ctr) {
  const unsigned long long epoch = *ctr + 1;
  __syncthreads();
  if (threadIdx.x == 0) *ctr = epoch;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(dst[i]), "l"(epoch) : "memory");
  }
}

__global__ void peer_wait_k(const unsigned long long *__restrict__ flags, const int32_t *__restrict__ ranks, int n,
                            const unsigned long long *__restrict__ ctr, int rank) {
  const unsigned long long epoch = *ctr;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const unsigned long long *p = flags + ranks[i];
    const unsigned long long t0 = global_ns();
    for (;;) {
      unsigned long long v;
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
      if (v >= epoch) break;
      if (global_ns() - t0 > 30000000000ull) {   // 30 s: a neighbour never produced this apply
        printf("[d4b] rank %d: peer halo timed out waiting for rank %d (have %llu, need %llu)\n", rank, ranks[i], v, epoch);
        __trap();
      }
      __nanosleep(64);
    }
  }
}

struct PeerXchg {
  cudaIpcMemHandle_t uh, flags;
  int64_t ok;
};

// min over the ranks of a local 0/1
int all_agree(d4b_mesh *m, int64_t *d_scratch, int local_ok) {
  int64_t v = local_ok;
  D4B_CUDA(cudaMemcpyAsync(d_scratch, &v, sizeof(v), cudaMemcpyHostToDevice, m->stream));
  D4B_NCCL(g_nccl.AllReduce(d_scratch, d_scratch, 1, ncclInt64, ncclMin, g_comm, m->stream));
  D4B_CUDA(cudaMemcpyAsync(&v, d_scratch, sizeof(v), cudaMemcpyDeviceToHost, m->stream));
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  return (int)v;
}

void peer_release(d4b_mesh *m) {
  D4bPeerHalo &P = m->peer;
  for (void *p : P.mapped) cudaIpcCloseMemHandle(p);
  P.mapped.clear();
  if (P.uh2) { if (m->d_uhat >= P.uh2 && m->d_uhat < P.uh2 + 2 * P.H) m->d_uhat = nullptr; cudaFree(P.uh2); P.uh2 = nullptr; }
  if (P.flags) { cudaFree(P.flags); P.flags = nullptr; }
  if (P.d_signal) { cudaFree(P.d_signal); P.d_signal = nullptr; }
  if (P.d_wait) { cudaFree(P.d_wait); P.d_wait = nullptr; }
  if (P.d_ghost_uh_off) { cudaFree(P.d_ghost_uh_off); P.d_ghost_uh_off = nullptr; }
  if (P.d_epoch) { cudaFree(P.d_epoch); P.d_epoch = nullptr; }
  cudaGetLastError();
}

}  // namespace

// Collective over the communicator (every rank reaches it in its first apply on this mesh).
// D4B_PEER_HALO=0 keeps the NCCL exchange.
bool d4b_comm_peer_ready(d4b_mesh *m) {
  D4bPeerHalo &P = m->peer;
  if (P.state) return P.state > 0;
  P.state = -1;
  if (m->mpisize == 1) return false;
  D4B_CHECK(g_comm && g_size == m->mpisize && g_rank == m->mpirank,
            "mesh of rank %d/%d but d4b_comm_init was %s", m->mpirank, m->mpisize,
            g_comm ? "called with another rank/size" : "not called");
  const char *env = getenv("D4B_PEER_HALO");   // 0: keep the NCCL exchange
  // the general kernel reads nodal ghost values: such meshes keep the nodal exchange
  int ok = !(env && atoi(env) == 0) && m->n_generic == 0 && !m->fast.empty() && m->d_uhat != nullptr;
  const int R = m->mpisize;
  int64_t *d_sc = nullptr;
  D4B_CUDA(cudaMalloc(&d_sc, 4 * sizeof(int64_t)));
  int64_t H = (m->local_nodes + 31) / 32 * 32;
  D4B_CUDA(cudaMemcpyAsync(d_sc + 1, &H, sizeof(H), cudaMemcpyHostToDevice, m->stream));
  D4B_NCCL(g_nccl.AllReduce(d_sc + 1, d_sc + 1, 1, ncclInt64, ncclMax, g_comm, m->stream));
  D4B_CUDA(cudaMemcpyAsync(&H, d_sc + 1, sizeof(H), cudaMemcpyDeviceToHost, m->stream));
  if (!all_agree(m, d_sc, ok)) { cudaFree(d_sc); return false; }

  // buffers and their IPC handles
  P.H = H;
  PeerXchg mine;
  memset(&mine, 0, sizeof(mine));
  ok = cudaMalloc(&P.uh2, (size_t)2 * H * sizeof(double)) == cudaSuccess &&
       cudaMalloc(&P.flags, (size_t)R * sizeof(unsigned long long)) == cudaSuccess;
  if (ok) {
    D4B_CUDA(cudaMemsetAsync(P.uh2, 0, (size_t)2 * H * sizeof(double), m->stream));
    D4B_CUDA(cudaMemsetAsync(P.flags, 0, (size_t)R * sizeof(unsigned long long), m->stream));
    ok = cudaIpcGetMemHandle(&mine.uh, P.uh2) == cudaSuccess && cudaIpcGetMemHandle(&mine.flags, P.flags) == cudaSuccess;
  }
  cudaGetLastError();
  mine.ok = ok;
  char *d_x = nullptr;
  D4B_CUDA(cudaMalloc(&d_x, (size_t)(R + 1) * sizeof(PeerXchg)));
  D4B_CUDA(cudaMemcpyAsync(d_x + (size_t)R * sizeof(PeerXchg), &mine, sizeof(mine), cudaMemcpyHostToDevice, m->stream));
  D4B_NCCL(g_nccl.AllGather(d_x + (size_t)R * sizeof(PeerXchg), d_x, sizeof(PeerXchg), ncclChar, g_comm, m->stream));
  std::vector<PeerXchg> all(R);
  D4B_CUDA(cudaMemcpyAsync(all.data(), d_x, (size_t)R * sizeof(PeerXchg), cudaMemcpyDeviceToHost, m->stream));
  D4B_CUDA(cudaStreamSynchronize(m->stream));   // also: every rank's flags are zero before anybody can signal
  cudaFree(d_x);
  for (int q = 0; q < R; ++q) ok = ok && all[q].ok;

  // map the neighbours, learn where their mirrors' lines are (the nodal offset of each of my ghosts on its owner)
  std::vector<double *> puh(m->peers.size(), nullptr);
  std::vector<unsigned long long *> pfl(m->peers.size(), nullptr);
  for (size_t i = 0; i < m->peers.size() && ok; ++i) {
    const int q = m->peers[i].rank;
    void *a = nullptr, *b = nullptr;
    ok = cudaIpcOpenMemHandle(&a, all[q].uh, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
    if (ok) P.mapped.push_back(a);
    ok = ok && cudaIpcOpenMemHandle(&b, all[q].flags, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
    if (ok) P.mapped.push_back(b);
    puh[i] = static_cast<double *>(a); pfl[i] = static_cast<unsigned long long *>(b);
  }
  cudaGetLastError();
  int64_t *d_remote = nullptr;
  D4B_CUDA(cudaMalloc(&d_remote, (size_t)(m->n_ghost > 0 ? m->n_ghost : 1) * sizeof(int64_t)));
  D4B_NCCL(g_nccl.GroupStart());
  for (const D4bPeer &p : m->peers) {
    if (p.send_elem_cnt > 0)
      D4B_NCCL(g_nccl.Send(m->d_send_src + p.send_elem_off, (size_t)p.send_elem_cnt, ncclInt64, p.rank, g_comm, m->stream));
    if (p.ghost_cnt > 0)
      D4B_NCCL(g_nccl.Recv(d_remote + p.ghost_first, (size_t)p.ghost_cnt, ncclInt64, p.rank, g_comm, m->stream));
  }
  D4B_NCCL(g_nccl.GroupEnd());
  std::vector<int64_t> off((size_t)(m->n_ghost > 0 ? m->n_ghost : 1), 0);
  D4B_CUDA(cudaMemcpyAsync(off.data(), d_remote, (size_t)m->n_ghost * sizeof(int64_t), cudaMemcpyDeviceToHost, m->stream));
  const int agreed = all_agree(m, d_sc, ok);
  cudaFree(d_sc);
  if (!agreed) { cudaFree(d_remote); peer_release(m); return false; }

  for (size_t i = 0; i < m->peers.size(); ++i) {
    const D4bPeer &p = m->peers[i];
    const intptr_t delta = reinterpret_cast<intptr_t>(puh[i]) - reinterpret_cast<intptr_t>(P.uh2);
    D4B_CHECK(delta % (intptr_t)sizeof(double) == 0, "peer mapping is not 8-byte aligned");
    for (int g = p.ghost_first; g < p.ghost_first + p.ghost_cnt; ++g) {
      D4B_CHECK(off[g] >= 0 && off[g] < H, "peer halo: remote offset of ghost %d out of range", g);
      off[g] += delta / (intptr_t)sizeof(double);
    }
  }
  D4B_CUDA(cudaMemcpyAsync(d_remote, off.data(), (size_t)m->n_ghost * sizeof(int64_t), cudaMemcpyHostToDevice, m->stream));
  d4b_fast_patch_ghost_offsets(m, d_remote);          // conforming classes: neighbour offsets in their face descriptors
  for (int64_t &o : off) o -= m->local_nodes;         // mortar kernel: uhat + local_nodes + d_ghost_uh_off[g]
  D4B_CUDA(cudaMalloc(&P.d_ghost_uh_off, off.size() * sizeof(int64_t)));
  D4B_CUDA(cudaMemcpyAsync(P.d_ghost_uh_off, off.data(), off.size() * sizeof(int64_t), cudaMemcpyHostToDevice, m->stream));
  D4B_CUDA(cudaMalloc(&P.d_epoch, sizeof(unsigned long long)));
  D4B_CUDA(cudaMemsetAsync(P.d_epoch, 0, sizeof(unsigned long long), m->stream));
  P.n_nbr = (int)m->peers.size();
  std::vector<unsigned long long *> sig(P.n_nbr > 0 ? P.n_nbr : 1, nullptr);
  std::vector<int32_t> wr(P.n_nbr > 0 ? P.n_nbr : 1, 0);
  for (int i = 0; i < P.n_nbr; ++i) { sig[i] = pfl[i] + m->mpirank; wr[i] = m->peers[i].rank; }
  D4B_CUDA(cudaMalloc(&P.d_signal, sig.size() * sizeof(void *)));
  D4B_CUDA(cudaMalloc(&P.d_wait, wr.size() * sizeof(int32_t)));
  D4B_CUDA(cudaMemcpyAsync(P.d_signal, sig.data(), sig.size() * sizeof(void *), cudaMemcpyHostToDevice, m->stream));
  D4B_CUDA(cudaMemcpyAsync(P.d_wait, wr.data(), wr.size() * sizeof(int32_t), cudaMemcpyHostToDevice, m->stream));
  D4B_CUDA(cudaStreamSynchronize(m->stream));
  cudaFree(d_remote);
  // the one-copy buffer of d4b_fast_setup is no longer used
  cudaFree(m->d_uhat);
  m->device_bytes += (int64_t)(2 * H - (m->local_nodes + m->ghost_nodes)) * (int64_t)sizeof(double);
  m->d_uhat = P.uh2;
  P.state = 1;
  return true;
}

void d4b_comm_peer_begin(d4b_mesh *m) {
  D4bPeerHalo &P = m->peer;
  ++P.epoch;
  m->d_uhat = P.uh2 + (P.epoch & 1) * P.H;
}

// the copy of uhat the NEXT epoch will use (the Gauss-resident smoother writes its new iterate there)
double *d4b_comm_peer_next_copy(d4b_mesh *m) {
  const D4bPeerHalo &P = m->peer;
  return P.uh2 + ((P.epoch + 1) & 1) * P.H;
}

void d4b_comm_peer_signal(d4b_mesh *m, cudaStream_t st) {
  const D4bPeerHalo &P = m->peer;
  if (P.n_nbr == 0) return;
  peer_signal_k<<<1, 32, 0, st>>>(P.d_signal, P.n_nbr, P.d_epoch);
  D4B_CUDA(cudaGetLastError());
}

void d4b_comm_peer_wait(d4b_mesh *m, cudaStream_t st) {
  const D4bPeerHalo &P = m->peer;
  if (P.n_nbr == 0) return;
  peer_wait_k<<<1, 32, 0, st>>>(P.flags, P.d_wait, P.n_nbr, P.d_epoch, m->mpirank);
  D4B_CUDA(cudaGetLastError());
}

// Neighbours may still be reading this rank's uhat: meshes with an active peer halo are destroyed
// collectively (every rank drains its own device, then all meet before anything is unmapped or freed).
void d4b_comm_peer_teardown(d4b_mesh *m) {
  if (m->peer.state != 1) return;
  cudaDeviceSynchronize();
  if (g_comm && g_size == m->mpisize) {
    int64_t *d_sc = nullptr;
    if (cudaMalloc(&d_sc, sizeof(int64_t)) == cudaSuccess) {
      all_agree(m, d_sc, 1);
      cudaFree(d_sc);
    }
  }
  peer_release(m);
  m->peer.state = -1;
}

// d4b_comm_destroy is collective like d4b_comm_init: every rank drains its device, all meet, then unmap
static void d4b_comm_release_allreduce() {
  if (g_ar.state == 1) {
    cudaDeviceSynchronize();
    if (g_comm) {
      int64_t *d_v = nullptr;
      if (cudaMalloc(&d_v, sizeof(int64_t)) == cudaSuccess) {
        cudaMemset(d_v, 0, sizeof(int64_t));
        g_nccl.AllReduce(d_v, d_v, 1, ncclInt64, ncclMin, g_comm, (cudaStream_t)0);
        cudaStreamSynchronize((cudaStream_t)0);
        cudaFree(d_v);
      }
    }
    peer_allreduce_release();
  }
  g_ar.state = 0;
}

// ---- CUDA-graph capture of solver sequences on multi-rank meshes (d4b_mg.cu) --------------------
// A mesh's applies can be captured when nothing in them calls NCCL and nothing is a kernel argument that must change
// from replay to replay: peer halo and peer all-reduce active (their epochs are device counters), no phase timing.
bool d4b_comm_graph_safe(d4b_mesh *m) {
  if (m->mpisize == 1) return true;
  static const bool timing = getenv("D4B_HALO_TIMING") && atoi(getenv("D4B_HALO_TIMING")) != 0;
  return m->peer.state == 1 && g_ar.state == 1 && !timing;
}
// the all-reduce orders itself across streams with an event: a capture must neither wait for an event recorded
// outside it nor leave one behind that is waited for outside (call before BeginCapture and after EndCapture,
// with the streams involved idle or ordered otherwise)
void d4b_comm_allreduce_reset_order() { g_ar.last_stream = nullptr; }
// which copy of uhat the mesh's next apply writes (0 / 1): part of the key of a captured sequence, whose kernels carry
// the pointers of the copies they were captured with
int d4b_comm_peer_parity(const d4b_mesh *m) { return m->peer.state == 1 ? (int)((m->peer.epoch + 1) & 1) : 0; }
// a replayed sequence has advanced the mesh by `epochs` applies
void d4b_comm_peer_advance(d4b_mesh *m, unsigned long long epochs) {
  if (m->peer.state != 1 || epochs == 0) return;
  m->peer.epoch += epochs;
  m->d_uhat = m->peer.uh2 + (m->peer.epoch & 1) * m->peer.H;
}
unsigned long long d4b_comm_peer_epoch(const d4b_mesh *m) { return m->peer.state == 1 ? m->peer.epoch : 0; }
