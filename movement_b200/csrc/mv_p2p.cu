// Peer-memory collectives over NVLink 5 / NVSwitch: the halo exchange and the scalar allreduce of the inner solvers as
// ONE kernel each, writing straight into the neighbour's memory.
//
// replaces: the MPI halo exchange / MPI_Allreduce implicit in PyOP2 and PETSc on the latency-critical
//           inner-solver path, where round 1 used grouped ncclSend/ncclRecv (36 point-to-point operations for the
//           9-component exchange of an inner rank: ~0.27 ms in front of a 0.22 ms kernel at 8 GPUs) and ncclAllReduce of
//           <= 9 doubles (a full launch + rendezvous each, two per PCG iteration).
//
// Every rank owns a WINDOW (cudaMalloc, exported with cudaIpcGetMemHandle and mapped by every peer at mv_comm_init):
//   flags   halo_flag[src]  : sequence number of the last halo message of rank `src` that is complete in this window
//           red_flag[src]   : same for the allreduce mailboxes
//   red     [2 parities][nranks][P2P_RED_MAX] doubles
//   slots   [nranks sources][2 parities][slot_bytes] halo payloads
// Halo exchange (k_p2p_halo): every CTA gathers its share of the send list and stores it directly into the neighbour's
// slot (peer stores over NVLink); __threadfence_system + a ticket: the last CTA to finish releases the flag of every
// neighbour (st.release.sys); then every CTA acquires the flags of the incoming messages (ld.acquire.sys on LOCAL
// memory) and copies the payload from its own window into the ghost segment of the vector.  No signal ever waits for
// a receipt, so ranks cannot deadlock; two parities per (source, destination) pair are enough because a rank sends
// message e + 2 to a peer only after its kernel e + 1 has seen that peer's message e + 1, which the peer sent after
// its kernel e (and the unpack inside it) had completed.  Sequence numbers live in device memory and are advanced by the
// kernels themselves, so the exchange needs no host bookkeeping between launches.
// Allreduce of n <= P2P_RED_MAX doubles (k_p2p_allreduce): one CTA stores its values into every peer's mailbox,
// releases, acquires all peers' flags and combines the contributions in rank order -- bitwise identical on every rank.
// Spin loops are bounded by wall time (%globaltimer): MV_P2P_TIMEOUT_S seconds, default 600, 0 = wait for ever like NCCL
// does.  A wait that times out (a neighbour stalled in host code for that long: hierarchy rebuild, a slow monitor
// callback, a debugger) raises a device status that the next host poll reports ONCE as MV_ERR_COMM and clears: the
// numbers of that solve are void, but the sequence counters advance at send time, so once the neighbour is back the
// following exchanges pair up again and the communicator stays usable.
#include "mv_internal.cuh"

namespace mv {

#define P2P_MAX_RANKS 16
#define P2P_RED_MAX 64
#define P2P_TIMEOUT_DEFAULT_S 600.0

struct P2PWindowLayout {
  size_t off_halo_flag, off_red_flag, off_red, off_slots, slot_bytes, total;
  unsigned long long timeout_ns;   // 0: no time-out (rides along because every kernel receives the layout by value)
};

struct P2P {
  bool on = false;
  int rank = 0, nranks = 1;
  unsigned char* win = nullptr;                 // own window
  unsigned char* peer[P2P_MAX_RANKS] = {nullptr};   // mapped windows (peer[rank] == win)
  P2PWindowLayout lay;
  unsigned long long* seq = nullptr;            // device: [0..nranks) halo sequence per peer, [nranks] allreduce sequence
  unsigned* ticket = nullptr;                   // device ticket of the halo kernel
  int* status = nullptr;                        // device sticky status (0 ok, 1 timeout)
  int* h_status = nullptr;                      // pinned mirror
  long long n_halo = 0, n_red = 0;              // launches (for the bench line)
};
static P2P g_p2p;

bool p2p_active() { return g_p2p.on; }
void p2p_counts(long long* halo, long long* red) { *halo = g_p2p.n_halo; *red = g_p2p.n_red; }

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// wait until *p >= want; false on timeout
__device__ __forceinline__ bool spin_until(const unsigned long long* p, unsigned long long want, unsigned long long timeout_ns) {
  if (ld_acquire_sys(p) >= want) return true;
  const unsigned long long t0 = global_ns();
  for (unsigned it = 1;; ++it) {
    if (ld_acquire_sys(p) >= want) return true;
    __nanosleep(20);
    if (timeout_ns && (it & 1023u) == 0 && global_ns() - t0 > timeout_ns) return false;
  }
}

struct HaloArgs {
  int nn;
  int ranks[P2P_MAX_RANKS];
  long long send_ptr[P2P_MAX_RANKS + 1], recv_ptr[P2P_MAX_RANKS + 1];
  unsigned char* peer_win[P2P_MAX_RANKS];   // window of neighbour k
};

// vec: component-major [C][ld] (SOA) or interleaved [ld][C]; owned rows [0, nown), ghosts after.
// LAYOUT 2: packed groups of four, [C/4][ld][4] (the float4 iterates of the mixed-precision Chebyshev sweeps, mv_cheb.cu)
template <int LAYOUT, typename T>
__global__ void __launch_bounds__(MV_BLOCK)
k_p2p_halo(HaloArgs h, int me, int C, const int* __restrict__ send_idx, T* __restrict__ vec, int64_t ld, int64_t nown,
           unsigned char* __restrict__ mywin, P2PWindowLayout lay, unsigned long long* __restrict__ seq,
           unsigned* __restrict__ ticket, int* __restrict__ status, const int* __restrict__ done) {
  if (done && *done) return;
  __shared__ bool s_last;
  __shared__ unsigned long long s_seq[P2P_MAX_RANKS];
  if (threadIdx.x < h.nn) s_seq[threadIdx.x] = seq[h.ranks[threadIdx.x]] + 1ull;
  __syncthreads();
  const int64_t tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
  // 1. pack + push: payload of neighbour k is [C][ns_k] (the receiver's ghost order), written into ITS window
  for (int k = 0; k < h.nn; ++k) {
    const int64_t ns = h.send_ptr[k + 1] - h.send_ptr[k];
    const int* idx = send_idx + h.send_ptr[k];
    T* dst = reinterpret_cast<T*>(h.peer_win[k] + lay.off_slots + ((size_t)me * 2 + (s_seq[k] & 1ull)) * lay.slot_bytes);
    const int64_t tot = ns * C;
    for (int64_t i = tid; i < tot; i += nth) {
      const int64_t c = i / ns, t = i - c * ns;
      const int64_t v = __ldg(&idx[t]);
      dst[i] = LAYOUT == 1 ? vec[c * ld + v] : (LAYOUT == 0 ? vec[v * C + c] : vec[((c >> 2) * ld + v) * 4 + (c & 3)]);
    }
  }
  // 2. the last CTA to finish its stores releases the flags
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (s_last) {
    __threadfence_system();
    if (threadIdx.x < h.nn) {
      unsigned long long* flag = reinterpret_cast<unsigned long long*>(h.peer_win[threadIdx.x] + lay.off_halo_flag) + me;
      st_release_sys(flag, s_seq[threadIdx.x]);
      seq[h.ranks[threadIdx.x]] = s_seq[threadIdx.x];
    }
    if (threadIdx.x == 0) *ticket = 0u;
  }
  // 3. acquire the incoming messages and unpack them into the ghost segment
  for (int k = 0; k < h.nn; ++k) {
    const int64_t nr = h.recv_ptr[k + 1] - h.recv_ptr[k];
    const unsigned long long* flag = reinterpret_cast<const unsigned long long*>(mywin + lay.off_halo_flag) + h.ranks[k];
    __shared__ bool s_ok;
    if (threadIdx.x == 0) s_ok = spin_until(flag, s_seq[k], lay.timeout_ns);
    __syncthreads();
    if (!s_ok) {
      if (threadIdx.x == 0) *status = 1;
      return;
    }
    const T* src = reinterpret_cast<const T*>(mywin + lay.off_slots + ((size_t)h.ranks[k] * 2 + (s_seq[k] & 1ull)) * lay.slot_bytes);
    const int64_t tot = nr * C, g0 = nown + h.recv_ptr[k];
    for (int64_t i = tid; i < tot; i += nth) {
      const int64_t c = i / nr, t = i - c * nr;
      const T val = __ldcv(&src[i]);      // written by the peer: never from a stale cache line
      if (LAYOUT == 1) vec[c * ld + g0 + t] = val;
      else if (LAYOUT == 0) vec[(g0 + t) * C + c] = val;
      else vec[((c >> 2) * ld + g0 + t) * 4 + (c & 3)] = val;
    }
    __syncthreads();
  }
}

struct RedArgs {
  unsigned char* peer_win[P2P_MAX_RANKS];
};

// in-place allreduce of n <= P2P_RED_MAX doubles; op: 0 sum, 1 max, 2 min.  One CTA.
__global__ void __launch_bounds__(128)
k_p2p_allreduce(RedArgs a, int me, int nranks, int n, int op, double* __restrict__ data, unsigned char* __restrict__ mywin,
                P2PWindowLayout lay, unsigned long long* __restrict__ seq, int* __restrict__ status, const int* __restrict__ done) {
  if (done && *done) return;
  const unsigned long long s = seq[nranks] + 1ull;
  const size_t par = (size_t)(s & 1ull);
  __shared__ bool s_ok;
  const int t = threadIdx.x;
  // 1. my contribution into every mailbox (my own included)
  for (int i = t; i < n * nranks; i += blockDim.x) {
    const int r = i / n, j = i - r * n;
    double* box = reinterpret_cast<double*>(a.peer_win[r] + lay.off_red) + (par * nranks + me) * P2P_RED_MAX;
    box[j] = data[j];
  }
  __threadfence_system();
  __syncthreads();
  if (t < nranks) st_release_sys(reinterpret_cast<unsigned long long*>(a.peer_win[t] + lay.off_red_flag) + me, s);
  // 2. wait for everybody's contribution
  if (t == 0) s_ok = true;
  __syncthreads();
  if (t < nranks) {
    if (!spin_until(reinterpret_cast<const unsigned long long*>(mywin + lay.off_red_flag) + t, s, lay.timeout_ns)) s_ok = false;
  }
  __syncthreads();
  if (!s_ok) {
    if (t == 0) *status = 1;
    return;
  }
  // 3. combine in rank order (identical bits on every rank)
  const double* boxes = reinterpret_cast<const double*>(mywin + lay.off_red) + par * nranks * P2P_RED_MAX;
  for (int j = t; j < n; j += blockDim.x) {
    double acc = __ldcv(&boxes[j]);
    for (int r = 1; r < nranks; ++r) {
      const double v = __ldcv(&boxes[(size_t)r * P2P_RED_MAX + j]);
      acc = op == 0 ? acc + v : (op == 1 ? fmax(acc, v) : fmin(acc, v));
    }
    data[j] = acc;
  }
  if (t == 0) seq[nranks] = s;
}

static P2PWindowLayout make_layout(int nranks, size_t window_bytes) {
  P2PWindowLayout l;
  size_t off = 0;
  l.off_halo_flag = off; off += P2P_MAX_RANKS * sizeof(unsigned long long);
  l.off_red_flag = off; off += P2P_MAX_RANKS * sizeof(unsigned long long);
  off = (off + 255) & ~(size_t)255;
  l.off_red = off; off += (size_t)2 * nranks * P2P_RED_MAX * sizeof(double);
  off = (off + 255) & ~(size_t)255;
  l.off_slots = off;
  l.slot_bytes = ((window_bytes - off) / ((size_t)2 * nranks)) & ~(size_t)255;
  l.total = off + l.slot_bytes * 2 * nranks;
  const char* e = getenv("MV_P2P_TIMEOUT_S");
  const double sec = e ? atof(e) : P2P_TIMEOUT_DEFAULT_S;
  l.timeout_ns = sec > 0.0 ? (unsigned long long)(sec * 1e9) : 0ull;
  return l;
}

// Called from mv_comm_init once NCCL is up: allocate the window, exchange the IPC handles (through the NCCL
// communicator), map the peers.  Any failure leaves the NCCL path in charge (g_p2p.on stays false).
int p2p_init(int rank, int nranks, int (*allgather_bytes)(const void* sendbuf, void* recvbuf, size_t bytes)) {
  P2P& p = g_p2p;
  p.on = false;
  const char* e = getenv("MV_P2P");
  if (e && atoi(e) == 0) return MV_OK;
  if (nranks < 2 || nranks > P2P_MAX_RANKS) return MV_OK;
  const char* wmb = getenv("MV_P2P_WINDOW_MB");
  const size_t window_bytes = (size_t)(wmb ? atoi(wmb) : 512) << 20;
  p.rank = rank;
  p.nranks = nranks;
  p.lay = make_layout(nranks, window_bytes);
  bool local_ok = cudaMalloc((void**)&p.win, p.lay.total) == cudaSuccess;
  if (!local_ok) { cudaGetLastError(); p.win = nullptr; }
  if (local_ok) local_ok = cudaMemset(p.win, 0, p.lay.off_slots) == cudaSuccess;
  MV_CUDA(cudaMalloc((void**)&p.seq, (P2P_MAX_RANKS + 1) * sizeof(unsigned long long)));
  MV_CUDA(cudaMemset(p.seq, 0, (P2P_MAX_RANKS + 1) * sizeof(unsigned long long)));
  MV_CUDA(cudaMalloc((void**)&p.ticket, sizeof(unsigned)));
  MV_CUDA(cudaMemset(p.ticket, 0, sizeof(unsigned)));
  MV_CUDA(cudaMalloc((void**)&p.status, sizeof(int)));
  MV_CUDA(cudaMemset(p.status, 0, sizeof(int)));
  MV_CUDA(cudaMallocHost((void**)&p.h_status, sizeof(int)));
  *p.h_status = 0;
  MV_CUDA(cudaDeviceSynchronize());
  // handle exchange: [handle (64 B) | ok flag]; every rank takes part in both exchanges whatever happened locally
  struct Msg { cudaIpcMemHandle_t h; int ok; int pad[15]; };
  static_assert(sizeof(Msg) == 128, "message size");
  Msg mine;
  memset(&mine, 0, sizeof(mine));
  mine.ok = (local_ok && cudaIpcGetMemHandle(&mine.h, p.win) == cudaSuccess) ? 1 : 0;
  if (!mine.ok) cudaGetLastError();
  std::vector<Msg> all(nranks);
  MV_CHECK(allgather_bytes(&mine, all.data(), sizeof(Msg)));
  bool ok = true;
  for (int r = 0; r < nranks; ++r) ok = ok && all[r].ok;
  int mapped = ok ? 1 : 0;
  if (ok) {
    for (int r = 0; r < nranks && mapped; ++r) {
      if (r == rank) { p.peer[r] = p.win; continue; }
      void* ptr = nullptr;
      if (cudaIpcOpenMemHandle(&ptr, all[r].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        cudaGetLastError();
        mapped = 0;
      }
      p.peer[r] = (unsigned char*)ptr;
    }
  }
  // everybody must agree, otherwise nobody uses the windows
  Msg vote;
  memset(&vote, 0, sizeof(vote));
  vote.ok = mapped;
  MV_CHECK(allgather_bytes(&vote, all.data(), sizeof(Msg)));
  for (int r = 0; r < nranks; ++r) mapped = mapped && all[r].ok;
  p.on = mapped != 0;
  if (getenv("MV_P2P_VERBOSE") && rank == 0)
    fprintf(stderr, "[mv p2p] %s: %d ranks, window %zu MiB, halo slot %zu KiB\n", p.on ? "peer windows mapped" : "unavailable, NCCL path",
            nranks, p.lay.total >> 20, p.lay.slot_bytes >> 10);
  return MV_OK;
}

void p2p_finalize() {
  P2P& p = g_p2p;
  if (!p.win) return;
  cudaDeviceSynchronize();
  for (int r = 0; r < p.nranks; ++r)
    if (r != p.rank && p.peer[r]) cudaIpcCloseMemHandle(p.peer[r]);
  cudaFree(p.win); cudaFree(p.seq); cudaFree(p.ticket); cudaFree(p.status);
  if (p.h_status) cudaFreeHost(p.h_status);
  p = P2P();
}

size_t p2p_slot_bytes() { return g_p2p.on ? g_p2p.lay.slot_bytes : 0; }

template <typename T>
static int p2p_halo_exchange_t(const Halo& h, int64_t nown, int C, T* vec, int64_t ld, int layout, const int* done) {
  P2P& p = g_p2p;
  HaloArgs a;
  a.nn = h.nn;
  int64_t work = 0;
  for (int k = 0; k < h.nn; ++k) {
    a.ranks[k] = h.ranks[k];
    a.peer_win[k] = p.peer[h.ranks[k]];
    const int64_t ns = h.send_ptr[k + 1] - h.send_ptr[k], nr = h.recv_ptr[k + 1] - h.recv_ptr[k];
    work += (ns > nr ? ns : nr) * C;
  }
  for (int k = 0; k <= h.nn; ++k) { a.send_ptr[k] = h.send_ptr[k]; a.recv_ptr[k] = h.recv_ptr[k]; }
  // every CTA must be able to become resident while earlier ones spin: stay within one wave, as the occupancy
  // calculator sees it for this instantiation (and never more than two CTAs per SM: room for a co-resident kernel)
  int grid = (int)((work + MV_BLOCK * 4 - 1) / (MV_BLOCK * 4));
  static int per_sm[3][2] = {{0, 0}, {0, 0}, {0, 0}};
  int& occ = per_sm[layout][sizeof(T) == 8 ? 1 : 0];
  if (occ == 0) {
    int nb = 0;
    cudaError_t qe = layout == 1 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_p2p_halo<1, T>, MV_BLOCK, 0)
                   : layout == 0 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_p2p_halo<0, T>, MV_BLOCK, 0)
                                 : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_p2p_halo<2, T>, MV_BLOCK, 0);
    if (qe != cudaSuccess) { cudaGetLastError(); nb = 1; }
    occ = nb < 1 ? 1 : (nb > 2 ? 2 : nb);
  }
  const int cap = (g_nsm > 0 ? g_nsm : 64) * occ;
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  if (layout == 1)
    MV_L k_p2p_halo<1, T><<<grid, MV_BLOCK, 0, g_stream>>>(a, p.rank, C, h.send_idx, vec, ld, nown, p.win, p.lay, p.seq, p.ticket, p.status, done);
  else if (layout == 0)
    MV_L k_p2p_halo<0, T><<<grid, MV_BLOCK, 0, g_stream>>>(a, p.rank, C, h.send_idx, vec, ld, nown, p.win, p.lay, p.seq, p.ticket, p.status, done);
  else
    MV_L k_p2p_halo<2, T><<<grid, MV_BLOCK, 0, g_stream>>>(a, p.rank, C, h.send_idx, vec, ld, nown, p.win, p.lay, p.seq, p.ticket, p.status, done);
  MV_KERNEL_CHECK();
  ++p.n_halo;
  return MV_OK;
}

int p2p_halo_exchange(const Halo& h, int64_t nown, int C, double* vec, int64_t ld, bool soa, const int* done) {
  return p2p_halo_exchange_t<double>(h, nown, C, vec, ld, soa ? 1 : 0, done);
}
int p2p_halo_exchange_f32(const Halo& h, int64_t nown, int C, float* vec, int64_t ld, const int* done) {
  return p2p_halo_exchange_t<float>(h, nown, C, vec, ld, 1, done);
}
// packed float4 groups: C = 4 * groups values per vertex, [groups][ld][4]
int p2p_halo_exchange_f32_packed4(const Halo& h, int64_t nown, int groups, float* vec, int64_t ld, const int* done) {
  return p2p_halo_exchange_t<float>(h, nown, 4 * groups, vec, ld, 2, done);
}

int p2p_allreduce(double* d, int n, int op, const int* done) {
  P2P& p = g_p2p;
  RedArgs a;
  for (int r = 0; r < p.nranks; ++r) a.peer_win[r] = p.peer[r];
  MV_L k_p2p_allreduce<<<1, 128, 0, g_stream>>>(a, p.rank, p.nranks, n, op, d, p.win, p.lay, p.seq, p.status, done);
  MV_KERNEL_CHECK();
  ++p.n_red;
  return MV_OK;
}

bool p2p_allreduce_fits(int n) { return g_p2p.on && n <= P2P_RED_MAX; }

// host poll (call after a stream synchronisation): a timed-out spin is a communication failure
int p2p_check_status() {
  P2P& p = g_p2p;
  if (!p.on) return MV_OK;
  MV_CUDA(cudaMemcpyAsync(p.h_status, p.status, sizeof(int), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  if (*p.h_status) {
    set_error("peer-memory collective timed out waiting for a neighbour (rank %d; MV_P2P_TIMEOUT_S = %.0f s); the results of "
              "this call are void, the communicator stays usable", p.rank, (double)p.lay.timeout_ns * 1e-9);
    *p.h_status = 0;
    MV_CUDA(cudaMemsetAsync(p.status, 0, sizeof(int), g_stream));   // reported once, not for ever
    return MV_ERR_COMM;
  }
  return MV_OK;
}

}  // namespace mv
