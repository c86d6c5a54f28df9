// Multi-GPU plumbing: one process per GPU, NCCL over NVLink 5 / NVSwitch.
//
// replaces: the MPI traffic that is implicit in the reference -- PyOP2 halo
//           exchanges behind every assemble/interpolate/assign and the MPI_Allreduce inside
//           every PETSc KSP dot product / norm and adapt_common's function_data_{max,min,sum}.
//
// The path shards by vertex ownership (owner computes): the only data-path collectives are
//   * forward halo exchange of a vertex field (owned -> ghost), grouped ncclSend/ncclRecv,
//     receiving straight into the contiguous ghost segment of the vector;
//   * ncclAllReduce of a handful of device-resident fp64 scalars (sum / max / min).
// NCCL is dlopen'ed so that the single-GPU library has no link-time dependency on it.
#include <dlfcn.h>
#include <nccl.h>

#include "mv_internal.cuh"

namespace mv {

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi g_nccl;
static ncclComm_t g_comm = nullptr;
static int g_rank = 0, g_nranks = 1;

#define MV_NCCL(call)                                                                   \
  do {                                                                                  \
    ncclResult_t r__ = (call);                                                          \
    if (r__ != ncclSuccess) {                                                           \
      set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call,                           \
                g_nccl.GetErrorString ? g_nccl.GetErrorString(r__) : "nccl error");     \
      return MV_ERR_COMM;                                                               \
    }                                                                                   \
  } while (0)

static int load_nccl(const char* path) {
  if (g_nccl.lib) return MV_OK;
  void* lib = dlopen(path && path[0] ? path : "libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) {
    set_error("cannot dlopen NCCL (%s): %s", path ? path : "libnccl.so.2", dlerror());
    return MV_ERR_COMM;
  }
#define SYM(field, name)                                         \
  *(void**)(&g_nccl.field) = dlsym(lib, name);                   \
  if (!g_nccl.field) {                                           \
    set_error("NCCL symbol %s missing", name);                   \
    return MV_ERR_COMM;                                          \
  }
  SYM(GetUniqueId, "ncclGetUniqueId");
  SYM(CommInitRank, "ncclCommInitRank");
  SYM(CommDestroy, "ncclCommDestroy");
  SYM(AllReduce, "ncclAllReduce");
  SYM(Broadcast, "ncclBroadcast");
  SYM(AllGather, "ncclAllGather");
  SYM(Send, "ncclSend");
  SYM(Recv, "ncclRecv");
  SYM(GroupStart, "ncclGroupStart");
  SYM(GroupEnd, "ncclGroupEnd");
  SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
  g_nccl.lib = lib;
  return MV_OK;
}

bool comm_active() { return g_comm != nullptr && g_nranks > 1; }
int comm_size() { return g_nranks; }

int allreduce_dev(double* d, int n, int op) {
  if (!comm_active()) return MV_OK;
  if (p2p_allreduce_fits(n)) return p2p_allreduce(d, n, op, nullptr);   // one kernel over the peer windows (mv_p2p.cu)
  const ncclRedOp_t rop = op == 0 ? ncclSum : (op == 1 ? ncclMax : ncclMin);
  MV_NCCL(g_nccl.AllReduce(d, d, (size_t)n, ncclDouble, rop, g_comm, g_stream));
  return MV_OK;
}
int allreduce_sum_dev(double* d, int n) { return allreduce_dev(d, n, 0); }
// In-place allgather of C component rows: rank r holds vec[c*ld + r*chunk .. + chunk) and receives the others'
// chunks; ld >= chunk * ranks (work-split replicated multigrid levels, mv_amg_cycle.cu).
int allgather_rows_dev(double* vec, int64_t ld, int C, int64_t chunk) {
  if (!comm_active()) return MV_OK;
  if (C > 1) MV_NCCL(g_nccl.GroupStart());
  for (int c = 0; c < C; ++c)
    MV_NCCL(g_nccl.AllGather(vec + c * ld + (int64_t)g_rank * chunk, vec + c * ld, (size_t)chunk, ncclDouble, g_comm, g_stream));
  if (C > 1) MV_NCCL(g_nccl.GroupEnd());
  return MV_OK;
}
int allreduce_volstats_dev(double* dout) {
  MV_CHECK(allreduce_dev(dout + 0, 1, 0));
  MV_CHECK(allreduce_dev(dout + 1, 1, 1));
  MV_CHECK(allreduce_dev(dout + 2, 1, 2));
  return MV_OK;
}

template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_halo_pack(int64_t n, const int* __restrict__ idx, const double* __restrict__ vec, double* __restrict__ buf) {
  const int64_t tot = n * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t k = i / C;
    const int c = (int)(i - k * C);
    buf[i] = vec[(int64_t)__ldg(&idx[k]) * C + c];
  }
}

int halo_exchange(mv_mesh* m, int C, double* vec) {
  if (!comm_active() || m->halo.nn == 0) return MV_OK;
  Halo& h = m->halo;
  if (C <= h.p2p_maxC) return p2p_halo_exchange(h, m->Vown, C, vec, m->V, false, nullptr);
  const int64_t nsend = h.send_ptr[h.nn];
  const int grid = grid_for(nsend * C);
  switch (C) {
    case 1: MV_L k_halo_pack<1><<<grid, MV_BLOCK, 0, g_stream>>>(nsend, h.send_idx, vec, h.send_buf); break;
    case 2: MV_L k_halo_pack<2><<<grid, MV_BLOCK, 0, g_stream>>>(nsend, h.send_idx, vec, h.send_buf); break;
    case 3: MV_L k_halo_pack<3><<<grid, MV_BLOCK, 0, g_stream>>>(nsend, h.send_idx, vec, h.send_buf); break;
    case 4: MV_L k_halo_pack<4><<<grid, MV_BLOCK, 0, g_stream>>>(nsend, h.send_idx, vec, h.send_buf); break;
    case 9: MV_L k_halo_pack<9><<<grid, MV_BLOCK, 0, g_stream>>>(nsend, h.send_idx, vec, h.send_buf); break;
    default: set_error("halo_exchange: unsupported ncomp %d", C); return MV_ERR_UNSUPPORTED;
  }
  MV_KERNEL_CHECK();
  MV_NCCL(g_nccl.GroupStart());
  for (int k = 0; k < h.nn; ++k) {
    const int64_t ns = h.send_ptr[k + 1] - h.send_ptr[k], nr = h.recv_ptr[k + 1] - h.recv_ptr[k];
    if (ns > 0) MV_NCCL(g_nccl.Send(h.send_buf + h.send_ptr[k] * C, (size_t)(ns * C), ncclDouble, h.ranks[k], g_comm, g_stream));
    if (nr > 0)
      MV_NCCL(g_nccl.Recv(vec + (m->Vown + h.recv_ptr[k]) * C, (size_t)(nr * C), ncclDouble, h.ranks[k], g_comm, g_stream));
  }
  MV_NCCL(g_nccl.GroupEnd());
  return MV_OK;
}

int comm_rank() { return g_rank; }

// Host-visible small collectives for set-up code (sizes, offsets, flags): through a device scratch.
int allreduce_host(double* h, int n, int op) {
  if (!comm_active()) return MV_OK;
  double* d = nullptr;
  MV_CHECK(dev_alloc(&d, n));
  MV_CUDA(cudaMemcpyAsync(d, h, n * sizeof(double), cudaMemcpyHostToDevice, g_stream));
  int st = allreduce_dev(d, n, op);
  if (st == MV_OK) {
    MV_CUDA(cudaMemcpyAsync(h, d, n * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
  }
  dev_free(d);
  return st;
}
// all[r] = value of rank r (exact for |value| < 2^53): a zero-padded sum
int allgather_i64(int64_t mine, std::vector<int64_t>& all) {
  all.assign(g_nranks, 0);
  all[g_rank] = mine;
  if (!comm_active()) return MV_OK;
  std::vector<double> h(g_nranks, 0.0);
  h[g_rank] = (double)mine;
  MV_CHECK(allreduce_host(h.data(), g_nranks, 0));
  for (int r = 0; r < g_nranks; ++r) all[r] = (int64_t)(h[r] + (h[r] >= 0 ? 0.5 : -0.5));
  return MV_OK;
}
// In-place broadcast of raw device bytes from `root`.
int bcast_bytes(void* buf, size_t bytes, int root) {
  if (!comm_active() || bytes == 0) return MV_OK;
  MV_NCCL(g_nccl.Broadcast(buf, buf, bytes, ncclChar, root, g_comm, g_stream));
  return MV_OK;
}
// Neighbour exchange of raw device bytes: neighbour k gets sbuf[k] (sbytes[k]) and fills rbuf[k] (rbytes[k]).
int neighbor_exchange_bytes(int nn, const int* ranks, const void* const* sbuf, const size_t* sbytes, void* const* rbuf,
                            const size_t* rbytes) {
  if (!comm_active() || nn == 0) return MV_OK;
  MV_NCCL(g_nccl.GroupStart());
  for (int k = 0; k < nn; ++k) {
    if (sbytes[k] > 0) MV_NCCL(g_nccl.Send(sbuf[k], sbytes[k], ncclChar, ranks[k], g_comm, g_stream));
    if (rbytes[k] > 0) MV_NCCL(g_nccl.Recv(rbuf[k], rbytes[k], ncclChar, ranks[k], g_comm, g_stream));
  }
  MV_NCCL(g_nccl.GroupEnd());
  return MV_OK;
}

// component-major vectors ([C][ld]): one contiguous segment per component and neighbour, received in place
__global__ void k_halo_pack_soa(int64_t nsend, int C, const int* __restrict__ idx, const double* __restrict__ vec, int64_t ld,
                                double* __restrict__ buf) {
  const int64_t n = nsend * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t c = i / nsend, t = i - c * nsend;
    buf[i] = vec[c * ld + idx[t]];
  }
}

// Packed variant for C > 1: ONE message per neighbour ([C][ns_k] contiguous in the send buffer, received into a
// staging buffer and scattered by a kernel) instead of C messages per neighbour.  At 8 GPUs the C = 9 exchange of
// the Hessian-recovery iteration cost ~0.27 ms as 36 point-to-point operations against 0.22 ms for the kernel it feeds.
__global__ void k_halo_pack_nb(int64_t ns, int C, const int* __restrict__ idx, const double* __restrict__ vec, int64_t ld,
                               double* __restrict__ buf) {
  const int64_t n = ns * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t c = i / ns, t = i - c * ns;
    buf[i] = vec[c * ld + idx[t]];
  }
}
__global__ void k_halo_unpack_nb(int64_t nr, int C, const double* __restrict__ buf, double* __restrict__ ghost, int64_t ld) {
  const int64_t n = nr * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t c = i / nr, t = i - c * nr;
    ghost[c * ld + t] = buf[i];
  }
}
static double* g_halo_rbuf = nullptr;
static size_t g_halo_rbuf_cap = 0;

static int halo_exchange_soa_packed(const Halo& h, int64_t nown, int C, double* vec, int64_t ld) {
  const size_t need = (size_t)h.recv_ptr[h.nn] * C;
  if (need > g_halo_rbuf_cap) {
    dev_free(g_halo_rbuf);
    MV_CHECK(dev_alloc(&g_halo_rbuf, need));
    g_halo_rbuf_cap = need;
  }
  for (int k = 0; k < h.nn; ++k) {
    const int64_t ns = h.send_ptr[k + 1] - h.send_ptr[k];
    if (ns > 0)
      MV_L k_halo_pack_nb<<<grid_for(ns * C), MV_BLOCK, 0, g_stream>>>(ns, C, h.send_idx + h.send_ptr[k], vec, ld,
                                                                  h.send_buf + (size_t)C * h.send_ptr[k]);
  }
  MV_KERNEL_CHECK();
  MV_NCCL(g_nccl.GroupStart());
  for (int k = 0; k < h.nn; ++k) {
    const int64_t ns = h.send_ptr[k + 1] - h.send_ptr[k], nr = h.recv_ptr[k + 1] - h.recv_ptr[k];
    if (ns > 0) MV_NCCL(g_nccl.Send(h.send_buf + (size_t)C * h.send_ptr[k], (size_t)(ns * C), ncclDouble, h.ranks[k], g_comm, g_stream));
    if (nr > 0) MV_NCCL(g_nccl.Recv(g_halo_rbuf + (size_t)C * h.recv_ptr[k], (size_t)(nr * C), ncclDouble, h.ranks[k], g_comm, g_stream));
  }
  MV_NCCL(g_nccl.GroupEnd());
  for (int k = 0; k < h.nn; ++k) {
    const int64_t nr = h.recv_ptr[k + 1] - h.recv_ptr[k];
    if (nr > 0)
      MV_L k_halo_unpack_nb<<<grid_for(nr * C), MV_BLOCK, 0, g_stream>>>(nr, C, g_halo_rbuf + (size_t)C * h.recv_ptr[k],
                                                                    vec + nown + h.recv_ptr[k], ld);
  }
  MV_KERNEL_CHECK();
  return MV_OK;
}

// Collective (every rank, once per halo plan): the peer-memory path carries an exchange of C components only if the
// largest message of ANY rank fits a window slot -- a per-rank decision could pair a peer-memory sender with an NCCL receiver.
int halo_finalize(Halo& h) {
  h.p2p_maxC = 0;
  if (!comm_active() || !p2p_active()) return MV_OK;
  double m = 0.0;
  for (int k = 0; k < h.nn; ++k) {
    const double ns = (double)(h.send_ptr[k + 1] - h.send_ptr[k]), nr = (double)(h.recv_ptr[k + 1] - h.recv_ptr[k]);
    m = fmax(m, fmax(ns, nr));
  }
  double nn = (double)h.nn;
  MV_CHECK(allreduce_host(&m, 1, 1));
  MV_CHECK(allreduce_host(&nn, 1, 1));
  if (nn > 16.0) return MV_OK;
  const double fit = m > 0.0 ? floor((double)p2p_slot_bytes() / (sizeof(double) * m)) : 64.0;
  h.p2p_maxC = fit > 64.0 ? 64 : (int)fit;
  return MV_OK;
}

int halo_exchange_soa(const Halo& h, int64_t nown, int C, double* vec, int64_t ld) {
  if (!comm_active() || h.nn == 0) return MV_OK;
  if (C <= h.p2p_maxC) return p2p_halo_exchange(h, nown, C, vec, ld, true, nullptr);
  static int packed = -1;
  if (packed < 0) { const char* e = getenv("MV_HALO_PACKED"); packed = e ? atoi(e) : 0; }
  if (packed && C > 1) return halo_exchange_soa_packed(h, nown, C, vec, ld);
  const int64_t nsend = h.send_ptr[h.nn];
  if (nsend > 0) MV_L k_halo_pack_soa<<<grid_for(nsend * C), MV_BLOCK, 0, g_stream>>>(nsend, C, h.send_idx, vec, ld, h.send_buf);
  MV_KERNEL_CHECK();
  MV_NCCL(g_nccl.GroupStart());
  for (int k = 0; k < h.nn; ++k) {
    const int64_t ns = h.send_ptr[k + 1] - h.send_ptr[k], nr = h.recv_ptr[k + 1] - h.recv_ptr[k];
    for (int c = 0; c < C; ++c) {
      if (ns > 0) MV_NCCL(g_nccl.Send(h.send_buf + c * nsend + h.send_ptr[k], (size_t)ns, ncclDouble, h.ranks[k], g_comm, g_stream));
      if (nr > 0) MV_NCCL(g_nccl.Recv(vec + c * ld + nown + h.recv_ptr[k], (size_t)nr, ncclDouble, h.ranks[k], g_comm, g_stream));
    }
  }
  MV_NCCL(g_nccl.GroupEnd());
  return MV_OK;
}

// single-precision vectors (the correction iterates of the mixed-precision Chebyshev sweeps); the send buffer of the
// plan is sized for MV_MAX_COMP doubles per entry, so the float payload always fits
__global__ void k_halo_pack_soa_f32(int64_t nsend, int C, const int* __restrict__ idx, const float* __restrict__ vec, int64_t ld,
                                    float* __restrict__ buf) {
  const int64_t n = nsend * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t c = i / nsend, t = i - c * nsend;
    buf[i] = vec[c * ld + idx[t]];
  }
}
int halo_exchange_soa_f32(const Halo& h, int64_t nown, int C, float* vec, int64_t ld) {
  if (!comm_active() || h.nn == 0) return MV_OK;
  if (C <= h.p2p_maxC) return p2p_halo_exchange_f32(h, nown, C, vec, ld, nullptr);
  const int64_t nsend = h.send_ptr[h.nn];
  float* sbuf = reinterpret_cast<float*>(h.send_buf);
  if (nsend > 0) MV_L k_halo_pack_soa_f32<<<grid_for(nsend * C), MV_BLOCK, 0, g_stream>>>(nsend, C, h.send_idx, vec, ld, sbuf);
  MV_KERNEL_CHECK();
  MV_NCCL(g_nccl.GroupStart());
  for (int k = 0; k < h.nn; ++k) {
    const int64_t ns = h.send_ptr[k + 1] - h.send_ptr[k], nr = h.recv_ptr[k + 1] - h.recv_ptr[k];
    for (int c = 0; c < C; ++c) {
      if (ns > 0) MV_NCCL(g_nccl.Send(sbuf + c * nsend + h.send_ptr[k], (size_t)ns, ncclFloat, h.ranks[k], g_comm, g_stream));
      if (nr > 0) MV_NCCL(g_nccl.Recv(vec + c * ld + nown + h.recv_ptr[k], (size_t)nr, ncclFloat, h.ranks[k], g_comm, g_stream));
    }
  }
  MV_NCCL(g_nccl.GroupEnd());
  return MV_OK;
}

}  // namespace mv

using namespace mv;

extern "C" int mv_comm_unique_id(const char* nccl_lib_path, void* unique_id_128) {
  if (!unique_id_128) return MV_ERR_BADARG;
  MV_CHECK(load_nccl(nccl_lib_path));
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  MV_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(unique_id_128, &id, sizeof(id));
  return MV_OK;
}

// host bytes of every rank, concatenated in rank order (set-up only): staged through the device for ncclAllGather
static int allgather_host_bytes(const void* sendbuf, void* recvbuf, size_t bytes) {
  unsigned char *ds = nullptr, *dr = nullptr;
  MV_CUDA(cudaMalloc((void**)&ds, bytes));
  MV_CUDA(cudaMalloc((void**)&dr, bytes * g_nranks));
  MV_CUDA(cudaMemcpyAsync(ds, sendbuf, bytes, cudaMemcpyHostToDevice, g_stream));
  MV_NCCL(g_nccl.AllGather(ds, dr, bytes, ncclChar, g_comm, g_stream));
  MV_CUDA(cudaMemcpyAsync(recvbuf, dr, bytes * g_nranks, cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  cudaFree(ds);
  cudaFree(dr);
  return MV_OK;
}

extern "C" int mv_comm_init(const char* nccl_lib_path, const void* unique_id_128, int rank, int nranks) {
  if (!unique_id_128 || rank < 0 || rank >= nranks) return MV_ERR_BADARG;
  MV_CHECK(load_nccl(nccl_lib_path));
  ncclUniqueId id;
  memcpy(&id, unique_id_128, sizeof(id));
  MV_NCCL(g_nccl.CommInitRank(&g_comm, nranks, id, rank));
  g_rank = rank;
  g_nranks = nranks;
  // peer-memory windows for the latency-critical collectives (mv_p2p.cu); NCCL stays for set-up and large messages
  MV_CHECK(p2p_init(rank, nranks, allgather_host_bytes));
  return MV_OK;
}

// info[0] = 1 if the peer-memory path is active, info[1] / info[2] = halo exchanges / allreduces it has carried
extern "C" int mv_comm_info(int64_t* info) {
  if (!info) return MV_ERR_BADARG;
  long long h = 0, r = 0;
  p2p_counts(&h, &r);
  info[0] = p2p_active() ? 1 : 0;
  info[1] = h;
  info[2] = r;
  info[3] = g_nranks;
  return MV_OK;
}

extern "C" int mv_comm_finalize(void) {
  cudaStreamSynchronize(g_stream);
  pcg_graph_release_all();   // graphs with NCCL nodes first: ncclCommDestroy waits for them
  p2p_finalize();
  if (g_comm) {
    cudaStreamSynchronize(g_stream);
    g_nccl.CommDestroy(g_comm);
    g_comm = nullptr;
  }
  g_nranks = 1;
  g_rank = 0;
  return MV_OK;
}

extern "C" int mv_mesh_set_halo(mv_mesh* m, int nn, const int32_t* ranks, const int64_t* send_ptr, const int32_t* send_idx,
                                const int64_t* recv_ptr) {
  if (!m || nn < 0 || (nn > 0 && (!ranks || !send_ptr || !send_idx || !recv_ptr))) return MV_ERR_BADARG;
  Halo& h = m->halo;
  h.nn = nn;
  h.ranks.assign(ranks, ranks + nn);
  h.send_ptr.assign(send_ptr, send_ptr + nn + 1);
  h.recv_ptr.assign(recv_ptr, recv_ptr + nn + 1);
  dev_free(h.send_idx);
  dev_free(h.send_buf);
  const int64_t nsend = nn ? send_ptr[nn] : 0;
  if (m->Vown + (nn ? recv_ptr[nn] : 0) != m->V) {
    set_error("mv_mesh_set_halo: ghost count %lld does not match V - n_owned = %lld", (long long)(nn ? recv_ptr[nn] : 0),
              (long long)(m->V - m->Vown));
    return MV_ERR_BADARG;
  }
  MV_CHECK(dev_alloc(&h.send_idx, nsend));
  MV_CHECK(dev_alloc(&h.send_buf, nsend * (size_t)(m->dim * m->dim)));
  if (nsend) MV_CUDA(cudaMemcpyAsync(h.send_idx, send_idx, nsend * sizeof(int), cudaMemcpyDefault, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  return halo_finalize(h);
}

extern "C" int mv_halo_exchange(mv_mesh* m, int ncomp, double* vec) {
  if (!m || !vec) return MV_ERR_BADARG;
  MV_CHECK(halo_exchange(m, ncomp, vec));
  // the public entry hands the ghosts to the host side: a peer-memory wait that timed out must not pass as data
  if (comm_active() && p2p_active()) MV_CHECK(p2p_check_status());
  return MV_OK;
}

extern "C" int mv_allreduce_sum(double* device_scalars, int n) { return allreduce_sum_dev(device_scalars, n); }
