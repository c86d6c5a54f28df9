// Multi-GPU part of the multigrid set-up: every level is partitioned like the fine level (owned rows
// first, ghost columns after, grouped by owner rank), aggregates never straddle ranks and the
// prolongator is rank-local, so the only set-up communication per level is ONE forward exchange of
// sparse rows: the prolongator rows of the vertices on each rank's send list (the neighbours' ghosts).
// From them follow, without further messages, the coarse ghost set (distinct owner-local aggregate
// ids received, ascending), the coarse send lists (the same set, computed by the sender from what it
// sent) and the rows of A P for the owned vertices -- hence the owned rows of A_c = P^T (A P).
//
// replaces: the parallel PtAP / PCGAMG set-up PETSc performs over MPI for
//           movement/solver_parameters.py:68-71 (MPI is implicit in the reference).
#include <cub/device/device_scan.cuh>

#include "mv_internal.cuh"

namespace mv {

#define GS(i, n) for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < (n); i += (int64_t)gridDim.x * blockDim.x)

__global__ void k_send_rowlen(int64_t nsend, const int* __restrict__ idx, const int64_t* __restrict__ ptr, int64_t* __restrict__ cnt) {
  GS(t, nsend + 1) cnt[t] = t < nsend ? ptr[idx[t] + 1] - ptr[idx[t]] : 0;
}
__global__ void k_send_rows(int64_t nsend, const int* __restrict__ idx, const int64_t* __restrict__ ptr, const int* __restrict__ col,
                            const double* __restrict__ val, const int64_t* __restrict__ soff, int* __restrict__ scol,
                            double* __restrict__ sval) {
  GS(t, nsend) {
    const int64_t b = ptr[idx[t]], e = ptr[idx[t] + 1], o = soff[t];
    for (int64_t k = b; k < e; ++k) { scol[o + k - b] = col[k]; sval[o + k - b] = val[k]; }
  }
}
__global__ void k_flag_cols(int64_t n, const int* __restrict__ col, int64_t* __restrict__ flag) {
  GS(t, n) flag[col[t]] = 1;   // benign race: every writer stores 1
}
// distinct flagged ids, ascending: list[rank[J]] = J ; optional global ids
__global__ void k_compact_flags(int64_t n, const int64_t* __restrict__ flag, const int64_t* __restrict__ rank, int* __restrict__ list,
                                int64_t* __restrict__ gid, int64_t goff) {
  GS(j, n) if (flag[j]) {
    if (list) list[rank[j]] = (int)j;
    if (gid) gid[rank[j]] = goff + j;
  }
}
__global__ void k_map_cols(int64_t n, int* __restrict__ col, const int64_t* __restrict__ rank, int64_t base) {
  GS(t, n) col[t] = (int)(base + rank[col[t]]);
}
__global__ void k_shift_ptr(int64_t n, const int64_t* __restrict__ roff, int64_t base, int64_t* __restrict__ ptr) {
  GS(t, n) ptr[t] = base + roff[t];
}
__global__ void k_iota_gid(int64_t n, int64_t goff, int64_t* __restrict__ gid) {
  GS(i, n) gid[i] = goff + i;
}

static int scan64(const int64_t* in, int64_t* out, int64_t n) { return exclusive_scan_i64(in, out, n); }

// In: level halo H (of the fine level), local prolongator P [n x nc_own] (CSR), nc of every rank.
// Out: Pext [ntot x ntot_c] = P rows of owned AND ghost vertices in local coarse numbering, the coarse
// level's halo, column count and global ids.
int amg_exchange_prolongator(const Halo& H, int64_t n, int64_t ntot, const Csr& P, const std::vector<int64_t>& nc_all,
                             const std::vector<int64_t>& coff_all, Csr& Pext, Halo& Hc, int64_t* ntot_c, int64_t** gid_c) {
  const int nn = H.nn;
  const int me = comm_rank();
  const int64_t nc_own = nc_all[me];
  const int64_t nsend = nn ? H.send_ptr[nn] : 0, nrecv = ntot - n;
  // (a) row lengths
  int64_t *scnt = nullptr, *rcnt = nullptr, *soff = nullptr, *roff = nullptr;
  MV_CHECK(dev_alloc(&scnt, nsend + 1));
  MV_CHECK(dev_alloc(&rcnt, nrecv + 1));
  MV_CHECK(dev_alloc(&soff, nsend + 1));
  MV_CHECK(dev_alloc(&roff, nrecv + 1));
  MV_L k_send_rowlen<<<grid_for(nsend + 1), MV_BLOCK, 0, g_stream>>>(nsend, H.send_idx, P.ptr, scnt);
  MV_KERNEL_CHECK();
  MV_CUDA(cudaMemsetAsync(rcnt, 0, (nrecv + 1) * sizeof(int64_t), g_stream));
  std::vector<const void*> sb(nn);
  std::vector<void*> rb(nn);
  std::vector<size_t> sby(nn), rby(nn);
  for (int k = 0; k < nn; ++k) {
    sb[k] = scnt + H.send_ptr[k]; sby[k] = (size_t)(H.send_ptr[k + 1] - H.send_ptr[k]) * sizeof(int64_t);
    rb[k] = rcnt + H.recv_ptr[k]; rby[k] = (size_t)(H.recv_ptr[k + 1] - H.recv_ptr[k]) * sizeof(int64_t);
  }
  MV_CHECK(neighbor_exchange_bytes(nn, H.ranks.data(), sb.data(), sby.data(), rb.data(), rby.data()));
  MV_CHECK(scan64(scnt, soff, nsend + 1));
  MV_CHECK(scan64(rcnt, roff, nrecv + 1));
  std::vector<int64_t> hso(nsend + 1), hro(nrecv + 1);
  MV_CUDA(cudaMemcpyAsync(hso.data(), soff, (nsend + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaMemcpyAsync(hro.data(), roff, (nrecv + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  const int64_t stot = hso[nsend], rtot = hro[nrecv];
  // (b) the rows themselves
  int *scol = nullptr, *gcol = nullptr;
  double *sval = nullptr, *gval = nullptr;
  MV_CHECK(dev_alloc(&scol, stot));
  MV_CHECK(dev_alloc(&sval, stot));
  MV_CHECK(dev_alloc(&gcol, rtot));
  MV_CHECK(dev_alloc(&gval, rtot));
  if (nsend) MV_L k_send_rows<<<grid_for(nsend), MV_BLOCK, 0, g_stream>>>(nsend, H.send_idx, P.ptr, P.col, P.val, soff, scol, sval);
  MV_KERNEL_CHECK();
  for (int pass = 0; pass < 2; ++pass) {
    const size_t es = pass == 0 ? sizeof(int) : sizeof(double);
    for (int k = 0; k < nn; ++k) {
      const int64_t s0 = hso[H.send_ptr[k]], s1 = hso[H.send_ptr[k + 1]], r0 = hro[H.recv_ptr[k]], r1 = hro[H.recv_ptr[k + 1]];
      sb[k] = pass == 0 ? (const void*)(scol + s0) : (const void*)(sval + s0);
      rb[k] = pass == 0 ? (void*)(gcol + r0) : (void*)(gval + r0);
      sby[k] = (size_t)(s1 - s0) * es;
      rby[k] = (size_t)(r1 - r0) * es;
    }
    MV_CHECK(neighbor_exchange_bytes(nn, H.ranks.data(), sb.data(), sby.data(), rb.data(), rby.data()));
  }
  // (c) coarse ghosts (receiver) and coarse send lists (sender): distinct ids per neighbour, ascending
  Hc.nn = nn;
  Hc.ranks = H.ranks;
  Hc.send_ptr.assign(nn + 1, 0);
  Hc.recv_ptr.assign(nn + 1, 0);
  std::vector<int*> send_lists(nn, nullptr);
  std::vector<int64_t*> ghost_gids(nn, nullptr);
  for (int k = 0; k < nn; ++k) {
    const int r = H.ranks[k];
    // receiver side
    {
      const int64_t nck = nc_all[r];
      const int64_t r0 = hro[H.recv_ptr[k]], r1 = hro[H.recv_ptr[k + 1]];
      int64_t *flag = nullptr, *rank = nullptr;
      MV_CHECK(dev_alloc(&flag, nck + 1));
      MV_CHECK(dev_alloc(&rank, nck + 1));
      MV_CUDA(cudaMemsetAsync(flag, 0, (nck + 1) * sizeof(int64_t), g_stream));
      if (r1 > r0) MV_L k_flag_cols<<<grid_for(r1 - r0), MV_BLOCK, 0, g_stream>>>(r1 - r0, gcol + r0, flag);
      MV_CHECK(scan64(flag, rank, nck + 1));
      int64_t ng = 0;
      MV_CUDA(cudaMemcpyAsync(&ng, rank + nck, sizeof(int64_t), cudaMemcpyDeviceToHost, g_stream));
      MV_CUDA(cudaStreamSynchronize(g_stream));
      Hc.recv_ptr[k + 1] = Hc.recv_ptr[k] + ng;
      MV_CHECK(dev_alloc(&ghost_gids[k], ng));
      if (nck) MV_L k_compact_flags<<<grid_for(nck), MV_BLOCK, 0, g_stream>>>(nck, flag, rank, nullptr, ghost_gids[k], coff_all[r]);
      if (r1 > r0) MV_L k_map_cols<<<grid_for(r1 - r0), MV_BLOCK, 0, g_stream>>>(r1 - r0, gcol + r0, rank, nc_own + Hc.recv_ptr[k]);
      MV_KERNEL_CHECK();
      MV_CUDA(cudaStreamSynchronize(g_stream));
      dev_free(flag); dev_free(rank);
    }
    // sender side
    {
      const int64_t s0 = hso[H.send_ptr[k]], s1 = hso[H.send_ptr[k + 1]];
      int64_t *flag = nullptr, *rank = nullptr;
      MV_CHECK(dev_alloc(&flag, nc_own + 1));
      MV_CHECK(dev_alloc(&rank, nc_own + 1));
      MV_CUDA(cudaMemsetAsync(flag, 0, (nc_own + 1) * sizeof(int64_t), g_stream));
      if (s1 > s0) MV_L k_flag_cols<<<grid_for(s1 - s0), MV_BLOCK, 0, g_stream>>>(s1 - s0, scol + s0, flag);
      MV_CHECK(scan64(flag, rank, nc_own + 1));
      int64_t ns = 0;
      MV_CUDA(cudaMemcpyAsync(&ns, rank + nc_own, sizeof(int64_t), cudaMemcpyDeviceToHost, g_stream));
      MV_CUDA(cudaStreamSynchronize(g_stream));
      Hc.send_ptr[k + 1] = Hc.send_ptr[k] + ns;
      MV_CHECK(dev_alloc(&send_lists[k], ns));
      if (nc_own) MV_L k_compact_flags<<<grid_for(nc_own), MV_BLOCK, 0, g_stream>>>(nc_own, flag, rank, send_lists[k], nullptr, 0);
      MV_KERNEL_CHECK();
      MV_CUDA(cudaStreamSynchronize(g_stream));
      dev_free(flag); dev_free(rank);
    }
  }
  const int64_t nsend_c = Hc.send_ptr[nn], nghost_c = Hc.recv_ptr[nn];
  MV_CHECK(dev_alloc(&Hc.send_idx, nsend_c));
  MV_CHECK(dev_alloc(&Hc.send_buf, nsend_c * 3));
  int64_t* gid = nullptr;
  MV_CHECK(dev_alloc(&gid, nc_own + nghost_c));
  if (nc_own) MV_L k_iota_gid<<<grid_for(nc_own), MV_BLOCK, 0, g_stream>>>(nc_own, coff_all[me], gid);
  for (int k = 0; k < nn; ++k) {
    const int64_t ns = Hc.send_ptr[k + 1] - Hc.send_ptr[k], ng = Hc.recv_ptr[k + 1] - Hc.recv_ptr[k];
    if (ns) MV_CUDA(cudaMemcpyAsync(Hc.send_idx + Hc.send_ptr[k], send_lists[k], ns * sizeof(int), cudaMemcpyDeviceToDevice, g_stream));
    if (ng) MV_CUDA(cudaMemcpyAsync(gid + nc_own + Hc.recv_ptr[k], ghost_gids[k], ng * sizeof(int64_t), cudaMemcpyDeviceToDevice, g_stream));
  }
  MV_CUDA(cudaStreamSynchronize(g_stream));
  for (int k = 0; k < nn; ++k) { dev_free(send_lists[k]); dev_free(ghost_gids[k]); }
  // (d) extended prolongator: owned rows, then the received ghost rows
  Pext.nrows = ntot;
  Pext.ncols = nc_own + nghost_c;
  Pext.nnz = P.nnz + rtot;
  MV_CHECK(dev_alloc(&Pext.ptr, ntot + 1));
  MV_CHECK(dev_alloc(&Pext.col, Pext.nnz));
  MV_CHECK(dev_alloc(&Pext.val, Pext.nnz));
  MV_CUDA(cudaMemcpyAsync(Pext.ptr, P.ptr, (n + 1) * sizeof(int64_t), cudaMemcpyDeviceToDevice, g_stream));
  MV_L k_shift_ptr<<<grid_for(nrecv + 1), MV_BLOCK, 0, g_stream>>>(nrecv + 1, roff, P.nnz, Pext.ptr + n);
  MV_CUDA(cudaMemcpyAsync(Pext.col, P.col, P.nnz * sizeof(int), cudaMemcpyDeviceToDevice, g_stream));
  MV_CUDA(cudaMemcpyAsync(Pext.val, P.val, P.nnz * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
  if (rtot) {
    MV_CUDA(cudaMemcpyAsync(Pext.col + P.nnz, gcol, rtot * sizeof(int), cudaMemcpyDeviceToDevice, g_stream));
    MV_CUDA(cudaMemcpyAsync(Pext.val + P.nnz, gval, rtot * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
  }
  MV_KERNEL_CHECK();
  MV_CUDA(cudaStreamSynchronize(g_stream));
  dev_free(scnt); dev_free(rcnt); dev_free(soff); dev_free(roff);
  dev_free(scol); dev_free(sval); dev_free(gcol); dev_free(gval);
  *ntot_c = nc_own + nghost_c;
  *gid_c = gid;
  return halo_finalize(Hc);
}

// ------------------------------------------------------------------ coarsest level, gathered on every rank
__global__ void k_dense_fill_global(int64_t n, int64_t N, int64_t goff, const int64_t* __restrict__ gid,
                                    const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
                                    const double* __restrict__ val, const int* __restrict__ row_nnz, double* __restrict__ a) {
  GS(i, n) {
    const int64_t base = slice_ptr[i >> 5] + (i & 31);
    const int nz = row_nnz[i];
    for (int k = 0; k < nz; ++k) {
      const int j = cols[base + (int64_t)k * MV_SLICE];
      a[(goff + i) * N + gid[j]] += val[base + (int64_t)k * MV_SLICE];
    }
  }
}
__global__ void k_add_shift(int64_t NN, double shift, double* __restrict__ a) { GS(i, NN) a[i] += shift; }

int amg_dense_gather(const AmgLevel& L, double shift, double* a /* [N*N], zeroed here */) {
  const int64_t N = L.nglobal;
  MV_CUDA(cudaMemsetAsync(a, 0, (size_t)N * N * sizeof(double), g_stream));
  if (L.n) MV_L k_dense_fill_global<<<grid_for(L.n), MV_BLOCK, 0, g_stream>>>(L.n, N, L.goff, L.gid, L.pat.slice_ptr, L.pat.cols, L.val,
                                                                        L.pat.row_nnz, a);
  MV_KERNEL_CHECK();
  // every entry is non-zero on exactly one rank, so the sum is exact whatever the reduction order
  for (int64_t o = 0; o < N * N; o += (1 << 20)) {
    const int64_t len = N * N - o < (1 << 20) ? N * N - o : (1 << 20);
    MV_CHECK(allreduce_sum_dev(a + o, (int)len));
  }
  if (shift != 0.0) MV_L k_add_shift<<<grid_for(N * N), MV_BLOCK, 0, g_stream>>>(N * N, shift, a);
  MV_KERNEL_CHECK();
  return MV_OK;
}

}  // namespace mv
