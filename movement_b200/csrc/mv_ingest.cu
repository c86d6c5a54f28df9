// Ingest: mesh topology -> device-resident SELL-32 operators.
//
// replaces: movement/mover.py:49-102 (mesh copy, P0/P1 spaces, CellVolume),
//           movement/monge_ampere.py:176-184 (total volume) and the matrix assembly that
//           Firedrake performs for the bilinear forms at monge_ampere.py:327 (P1 mass),
//           :437 (P1 stiffness), :476 (tensor mass), laplacian_smoothing.py:68.
//
// Pipeline (all on the device, atomic-free where values are produced):
//   1. vertex -> (cell,loc) CSR       count (int atomics) / scan / fill / per-row sort
//   2. row pattern                    per-vertex sorted unique neighbour list
//   3. SELL-32 layout                 slice width = max row length in 32 rows
//   4. values of M, K, m_l            "owner computes": one thread per row walks its
//                                     incident cells in ascending cell order, so every
//                                     sum has a fixed order (bit-reproducible, no atomics)
//   5. boundary facets                adjacent cell, outward normal, measure
#include <cub/device/device_scan.cuh>

#include "mv_internal.cuh"

namespace mv {

// ------------------------------------------------------------------ generic CSR inverse map
__global__ void k_count_keys(const int* __restrict__ keys, int64_t n, int64_t nown, int* __restrict__ cnt) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int k = keys[i];
    if (k < nown) atomicAdd(&cnt[k], 1);
  }
}
__global__ void k_fill_keys(const int* __restrict__ keys, int64_t n, int64_t nown, const int* __restrict__ ptr,
                            int* __restrict__ cursor, int* __restrict__ list) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int k = keys[i];
    if (k < nown) {
      int pos = atomicAdd(&cursor[k], 1);
      list[ptr[k] + pos] = (int)i;
    }
  }
}
__global__ void k_sort_rows(const int* __restrict__ ptr, int64_t nrows, int* __restrict__ list) {
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    int b = ptr[v], e = ptr[v + 1];
    for (int i = b + 1; i < e; ++i) {  // insertion sort, rows are short
      int key = list[i];
      int j = i - 1;
      while (j >= b && list[j] > key) {
        list[j + 1] = list[j];
        --j;
      }
      list[j + 1] = key;
    }
  }
}

int exclusive_scan_i32(const int* in, int* out, int64_t n) {
  void* tmp = nullptr;
  size_t bytes = 0;
  MV_CUDA(cub::DeviceScan::ExclusiveSum(tmp, bytes, in, out, (int)n, g_stream));
  MV_CUDA(cudaMallocAsync(&tmp, bytes, g_stream));
  MV_CUDA(cub::DeviceScan::ExclusiveSum(tmp, bytes, in, out, (int)n, g_stream));
  MV_CUDA(cudaFreeAsync(tmp, g_stream));
  return MV_OK;
}
int exclusive_scan_i64(const int64_t* in, int64_t* out, int64_t n) {
  void* tmp = nullptr;
  size_t bytes = 0;
  MV_CUDA(cub::DeviceScan::ExclusiveSum(tmp, bytes, in, out, (int)n, g_stream));
  MV_CUDA(cudaMallocAsync(&tmp, bytes, g_stream));
  MV_CUDA(cub::DeviceScan::ExclusiveSum(tmp, bytes, in, out, (int)n, g_stream));
  MV_CUDA(cudaFreeAsync(tmp, g_stream));
  return MV_OK;
}

// keys[n] in [0, nkeys) -> ptr[nown+1], list[...] (indices i ascending within a key)
static int build_inverse_map(const int* keys, int64_t n, int64_t nown, int** ptr_out, int** list_out) {
  int *cnt = nullptr, *ptr = nullptr, *list = nullptr;
  MV_CHECK(dev_alloc(&cnt, nown + 1));
  MV_CHECK(dev_alloc(&ptr, nown + 1));
  MV_CUDA(cudaMemsetAsync(cnt, 0, (nown + 1) * sizeof(int), g_stream));
  int grid = grid_for(n);
  if (n > 0) MV_L k_count_keys<<<grid, MV_BLOCK, 0, g_stream>>>(keys, n, nown, cnt);
  MV_KERNEL_CHECK();
  MV_CHECK(exclusive_scan_i32(cnt, ptr, nown + 1));
  int total = 0;
  MV_CUDA(cudaMemcpyAsync(&total, ptr + nown, sizeof(int), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  MV_CHECK(dev_alloc(&list, (size_t)total));
  MV_CUDA(cudaMemsetAsync(cnt, 0, (nown + 1) * sizeof(int), g_stream));
  if (n > 0) MV_L k_fill_keys<<<grid, MV_BLOCK, 0, g_stream>>>(keys, n, nown, ptr, cnt, list);
  MV_KERNEL_CHECK();
  if (nown > 0) MV_L k_sort_rows<<<grid_for(nown), MV_BLOCK, 0, g_stream>>>(ptr, nown, list);
  MV_KERNEL_CHECK();
  dev_free(cnt);
  *ptr_out = ptr;
  *list_out = list;
  return MV_OK;
}

// ------------------------------------------------------------------ row pattern
// Sorted unique neighbour list (including the vertex itself) of vertex v.
template <int D>
__device__ __forceinline__ int row_neighbours(const int* __restrict__ cells, const int* __restrict__ v2c_ptr,
                                              const int* __restrict__ v2c, int64_t v, int* nb, bool& overflow) {
  int n = 0;
  const int b = v2c_ptr[v], e = v2c_ptr[v + 1];
  // always contain self (isolated vertices still get a diagonal)
  nb[n++] = (int)v;
  for (int t = b; t < e; ++t) {
    const int64_t c = v2c[t] / (D + 1);
#pragma unroll
    for (int j = 0; j <= D; ++j) {
      const int u = __ldg(&cells[c * (D + 1) + j]);
      // sorted insert
      int lo = 0, hi = n;
      while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (nb[mid] < u) lo = mid + 1; else hi = mid;
      }
      if (lo < n && nb[lo] == u) continue;
      if (n >= MV_MAX_ROW) { overflow = true; continue; }
      for (int k = n; k > lo; --k) nb[k] = nb[k - 1];
      nb[lo] = u;
      ++n;
    }
  }
  return n;
}

template <int D>
__global__ void k_row_count(const int* __restrict__ cells, const int* __restrict__ v2c_ptr, const int* __restrict__ v2c,
                            int64_t nrows, int* __restrict__ row_nnz, int* __restrict__ err) {
  int nb[MV_MAX_ROW];
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    bool ovf = false;
    row_nnz[v] = row_neighbours<D>(cells, v2c_ptr, v2c, v, nb, ovf);
    if (ovf) atomicExch(err, 1);
  }
}

__global__ void k_slice_width(const int* __restrict__ row_nnz, int64_t nrows, int64_t nslices, int64_t* __restrict__ slice_len) {
  const int lane = threadIdx.x & 31;
  int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t s = warp; s < nslices; s += nwarps) {
    int64_t row = s * MV_SLICE + lane;
    int w = row < nrows ? row_nnz[row] : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) w = max(w, __shfl_xor_sync(0xffffffffu, w, o));
    if (lane == 0) slice_len[s] = (int64_t)w * MV_SLICE;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) slice_len[nslices] = 0;
}

template <int D>
__global__ void k_fill_cols(const int* __restrict__ cells, const int* __restrict__ v2c_ptr, const int* __restrict__ v2c,
                            int64_t nrows, int64_t nslices, const int64_t* __restrict__ slice_ptr, int* __restrict__ cols) {
  int nb[MV_MAX_ROW];
  const int64_t total = nslices * MV_SLICE;
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < total; v += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = v / MV_SLICE;
    const int lane = (int)(v % MV_SLICE);
    const int64_t base = slice_ptr[s];
    const int w = (int)((slice_ptr[s + 1] - base) / MV_SLICE);
    int n = 0;
    int self = 0;
    if (v < nrows) {
      bool ovf = false;
      n = row_neighbours<D>(cells, v2c_ptr, v2c, v, nb, ovf);
      self = (int)v;
    }
    for (int k = 0; k < w; ++k) cols[base + (int64_t)k * MV_SLICE + lane] = k < n ? nb[k] : self;
  }
}

// ------------------------------------------------------------------ values
template <int D>
__global__ void k_cell_volume(const double* __restrict__ x, const int* __restrict__ cells, int dg, int64_t C, double* __restrict__ vol) {
  for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < C; c += (int64_t)gridDim.x * blockDim.x) {
    int cv[D + 1];
    double xc[D + 1][D];
    CellGeom<D> g;
    load_cell<D>(cells, c, cv);
    load_cell_coords<D>(x, cv, xc, dg, c);
    cell_geom(xc, g);
    vol[c] = g.vol;
  }
}

template <int D>
__global__ void k_assemble_MK(const double* __restrict__ x, const int* __restrict__ cells, int dg, const int* __restrict__ v2c_ptr,
                              const int* __restrict__ v2c, int64_t nrows, const int64_t* __restrict__ slice_ptr,
                              const int* __restrict__ cols, double* __restrict__ Mval, double* __restrict__ Kval,
                              double* __restrict__ ml, double* __restrict__ dinvM, double* __restrict__ dinvK) {
  const double mfac = 1.0 / ((D + 1) * (D + 2));
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = v / MV_SLICE;
    const int lane = (int)(v % MV_SLICE);
    const int64_t base = slice_ptr[s] + lane;
    const int w = (int)((slice_ptr[s + 1] - slice_ptr[s]) / MV_SLICE);
    for (int k = 0; k < w; ++k) {
      Mval[base + (int64_t)k * MV_SLICE] = 0.0;
      Kval[base + (int64_t)k * MV_SLICE] = 0.0;
    }
    double mlv = 0.0, dM = 0.0, dK = 0.0;
    const int b = v2c_ptr[v], e = v2c_ptr[v + 1];
    for (int t = b; t < e; ++t) {
      const int entry = v2c[t];
      const int64_t c = entry / (D + 1);
      const int loc = entry % (D + 1);
      int cv[D + 1];
      double xc[D + 1][D];
      CellGeom<D> g;
      load_cell<D>(cells, c, cv);
      load_cell_coords<D>(x, cv, xc, dg, c);
      cell_geom(xc, g);
      mlv += g.vol / (D + 1);
      double gl[D];
#pragma unroll
      for (int a = 0; a < D; ++a) {
        gl[a] = 0.0;
#pragma unroll
        for (int jj = 0; jj <= D; ++jj)
          if (jj == loc) gl[a] = g.g[jj][a];
      }
#pragma unroll
      for (int j = 0; j <= D; ++j) {
        const int u = cv[j];
        double kij = 0.0;
#pragma unroll
        for (int a = 0; a < D; ++a) kij += gl[a] * g.g[j][a];
        kij *= g.vol;
        const double mij = g.vol * mfac * (j == loc ? 2.0 : 1.0);
        // columns are sorted: find u
        int pos = 0;
        for (; pos < w; ++pos)
          if (cols[base + (int64_t)pos * MV_SLICE] == u) break;
        Mval[base + (int64_t)pos * MV_SLICE] += mij;
        Kval[base + (int64_t)pos * MV_SLICE] += kij;
        if (j == loc) { dM += mij; dK += kij; }
      }
    }
    ml[v] = mlv;
    dinvM[v] = dM > 0.0 ? 1.0 / dM : 0.0;
    dinvK[v] = dK > 0.0 ? 1.0 / dK : 0.0;
  }
}

// Sum of `n` values (and of their squares) in a fixed order: per-block partials + last-block ticket.
__global__ void __launch_bounds__(MV_BLOCK)
k_sum_fixed(const double* __restrict__ v, int64_t n, double* __restrict__ out, double* __restrict__ out_sq,
            double* __restrict__ partials, unsigned* __restrict__ counter) {
  double acc[2] = {0.0, 0.0};
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double t = v[i];
    acc[0] += t;
    acc[1] = fma(t, t, acc[1]);
  }
  grid_reduce_sum<2>(acc, partials, counter, [&](double(&t)[2]) { out[0] = t[0]; if (out_sq) out_sq[0] = t[1]; });
}

// ------------------------------------------------------------------ boundary facets
template <int D>
__global__ void k_facet_geometry(const double* __restrict__ x, const int* __restrict__ cells, const int* __restrict__ v2c_ptr,
                                 const int* __restrict__ v2c, int64_t F, int64_t Vown, const int* __restrict__ bf_verts,
                                 int* __restrict__ bf_cell, double* __restrict__ bf_normal, double* __restrict__ bf_area) {
  for (int64_t f = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; f < F; f += (int64_t)gridDim.x * blockDim.x) {
    int fv[D];
#pragma unroll
    for (int a = 0; a < D; ++a) fv[a] = bf_verts[f * D + a];
    // adjacent cell: search the cells around an owned facet vertex
    int cell = -1, oppv = -1;
    for (int a = 0; a < D && cell < 0; ++a) {
      if (fv[a] >= Vown) continue;
      for (int t = v2c_ptr[fv[a]]; t < v2c_ptr[fv[a] + 1] && cell < 0; ++t) {
        const int64_t c = v2c[t] / (D + 1);
        int hit = 0, opp = -1;
        for (int j = 0; j <= D; ++j) {
          const int u = cells[c * (D + 1) + j];
          bool in = false;
          for (int q = 0; q < D; ++q) in |= (u == fv[q]);
          if (in) ++hit; else opp = u;
        }
        if (hit == D) { cell = (int)c; oppv = opp; }
      }
    }
    bf_cell[f] = cell;
    double n[D], area;
    if constexpr (D == 1) {
      // a boundary "facet" of an interval mesh is a point: measure 1, normal pointing away from the cell
      area = 1.0;
      n[0] = 1.0;
    } else if constexpr (D == 2) {
      const double ex = x[(size_t)fv[1] * D] - x[(size_t)fv[0] * D], ey = x[(size_t)fv[1] * D + 1] - x[(size_t)fv[0] * D + 1];
      area = sqrt(ex * ex + ey * ey);
      n[0] = ey / area;
      n[1] = -ex / area;
    } else {
      double e1[3], e2[3];
      for (int a = 0; a < 3; ++a) {
        e1[a] = x[(size_t)fv[1] * D + a] - x[(size_t)fv[0] * D + a];
        e2[a] = x[(size_t)fv[D - 1] * D + a] - x[(size_t)fv[0] * D + a];
      }
      double c[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
      const double nn = sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
      area = 0.5 * nn;
      for (int a = 0; a < D; ++a) n[a] = c[a] / nn;
    }
    if (oppv >= 0) {
      double dot = 0.0;
      for (int a = 0; a < D; ++a) dot += n[a] * (x[(size_t)fv[0] * D + a] - x[(size_t)oppv * D + a]);
      if (dot < 0.0)
        for (int a = 0; a < D; ++a) n[a] = -n[a];
    }
    for (int a = 0; a < D; ++a) bf_normal[f * D + a] = n[a];
    bf_area[f] = area;
  }
}

int grid_for(int64_t n, int blocks_per_sm) {
  int64_t need = (n + MV_BLOCK - 1) / MV_BLOCK;
  int64_t cap = (int64_t)g_nsm * blocks_per_sm;
  if (cap > MV_MAX_BLOCKS) cap = MV_MAX_BLOCKS;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

template <int D>
static int ingest_impl(mv_mesh* m) {
  const int64_t V = m->V, C = m->C, F = m->F, Vown = m->Vown;
  // 1. vertex -> cell map
  MV_CHECK(build_inverse_map(m->cells, C * (D + 1), Vown, &m->v2c_ptr, &m->v2c));
  // 2./3. pattern
  Sell& p = m->pat;
  p.nrows = Vown;
  p.nslices = (Vown + MV_SLICE - 1) / MV_SLICE;
  MV_CHECK(dev_alloc(&p.row_nnz, Vown));
  int* d_err = nullptr;
  MV_CHECK(dev_alloc(&d_err, 1));
  MV_CUDA(cudaMemsetAsync(d_err, 0, sizeof(int), g_stream));
  MV_L k_row_count<D><<<grid_for(Vown), MV_BLOCK, 0, g_stream>>>(m->cells, m->v2c_ptr, m->v2c, Vown, p.row_nnz, d_err);
  MV_KERNEL_CHECK();
  int64_t* slice_len = nullptr;
  MV_CHECK(dev_alloc(&slice_len, p.nslices + 1));
  MV_CHECK(dev_alloc(&p.slice_ptr, p.nslices + 1));
  MV_L k_slice_width<<<grid_for(p.nslices * 32), MV_BLOCK, 0, g_stream>>>(p.row_nnz, Vown, p.nslices, slice_len);
  MV_KERNEL_CHECK();
  MV_CHECK(exclusive_scan_i64(slice_len, p.slice_ptr, p.nslices + 1));
  int h_err = 0;
  MV_CUDA(cudaMemcpyAsync(&p.len, p.slice_ptr + p.nslices, sizeof(int64_t), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaMemcpyAsync(&h_err, d_err, sizeof(int), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  dev_free(slice_len);
  dev_free(d_err);
  if (h_err) {
    set_error("a vertex has more than %d neighbours", MV_MAX_ROW);
    return MV_ERR_VALENCE;
  }
  MV_CHECK(dev_alloc(&p.cols, p.len));
  MV_L k_fill_cols<D><<<grid_for(p.nslices * MV_SLICE), MV_BLOCK, 0, g_stream>>>(m->cells, m->v2c_ptr, m->v2c, Vown, p.nslices,
                                                                         p.slice_ptr, p.cols);
  MV_KERNEL_CHECK();
  // 4. values
  MV_CHECK(dev_alloc(&m->Mval, p.len));
  MV_CHECK(dev_alloc(&m->Kval, p.len));
  MV_CHECK(dev_alloc(&m->ml, Vown));
  MV_CHECK(dev_alloc(&m->dinvM, Vown));
  MV_CHECK(dev_alloc(&m->dinvK, Vown));
  MV_CHECK(dev_alloc(&m->vol, C));
  MV_CUDA(cudaMemsetAsync(m->Mval, 0, p.len * sizeof(double), g_stream));
  MV_CUDA(cudaMemsetAsync(m->Kval, 0, p.len * sizeof(double), g_stream));
  if (C > 0) MV_L k_cell_volume<D><<<grid_for(C), MV_BLOCK, 0, g_stream>>>(m->xi, m->cells, m->dg, C, m->vol);
  MV_KERNEL_CHECK();
  MV_L k_assemble_MK<D><<<grid_for(Vown), MV_BLOCK, 0, g_stream>>>(m->xi, m->cells, m->dg, m->v2c_ptr, m->v2c, Vown, p.slice_ptr, p.cols,
                                                            m->Mval, m->Kval, m->ml, m->dinvM, m->dinvK);
  MV_KERNEL_CHECK();
  // scalars (fixed-order sums)
  double* d_s = nullptr;
  MV_CHECK(dev_alloc(&d_s, 4));
  // work buffers of the deterministic grid reductions
  MV_CHECK(dev_alloc(&m->red.partials, (size_t)MV_MAX_BLOCKS * MV_RED_SLOTS));
  MV_CHECK(dev_alloc(&m->red.counter, 4));
  MV_CUDA(cudaMemsetAsync(m->red.counter, 0, 4 * sizeof(unsigned), g_stream));
  MV_L k_sum_fixed<<<grid_for(m->Cown), MV_BLOCK, 0, g_stream>>>(m->vol, m->Cown, d_s, nullptr, m->red.partials, m->red.counter);
  MV_L k_sum_fixed<<<grid_for(Vown), MV_BLOCK, 0, g_stream>>>(m->ml, Vown, d_s + 1, d_s + 2, m->red.partials, m->red.counter);
  MV_KERNEL_CHECK();
  double h_s[3];
  int h_nnz_dummy = 0;
  (void)h_nnz_dummy;
  MV_CUDA(cudaMemcpyAsync(h_s, d_s, 3 * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
  // nnz = sum row_nnz: reuse scan
  int* nnz_scan = nullptr;
  MV_CHECK(dev_alloc(&nnz_scan, Vown + 1));
  MV_CHECK(exclusive_scan_i32(p.row_nnz, nnz_scan, Vown));
  int h_last[2] = {0, 0};
  if (Vown > 0) {
    MV_CUDA(cudaMemcpyAsync(&h_last[0], nnz_scan + Vown - 1, sizeof(int), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaMemcpyAsync(&h_last[1], p.row_nnz + Vown - 1, sizeof(int), cudaMemcpyDeviceToHost, g_stream));
  }
  MV_CUDA(cudaStreamSynchronize(g_stream));
  dev_free(nnz_scan);
  dev_free(d_s);
  p.nnz = (int64_t)h_last[0] + h_last[1];
  p.maxw = 0;
  m->total_volume = h_s[0];
  m->ml_sum = h_s[1];
  m->ml_sumsq = h_s[2];
  // 5. boundary facets
  if (F > 0) {
    MV_CHECK(dev_alloc(&m->bf_cell, F));
    MV_CHECK(dev_alloc(&m->bf_normal, F * D));
    MV_CHECK(dev_alloc(&m->bf_area, F));
    MV_L k_facet_geometry<D><<<grid_for(F), MV_BLOCK, 0, g_stream>>>(m->xi, m->cells, m->v2c_ptr, m->v2c, F, Vown, m->bf_verts,
                                                              m->bf_cell, m->bf_normal, m->bf_area);
    MV_KERNEL_CHECK();
  }
  MV_CHECK(build_inverse_map(m->bf_verts, F * D, Vown, &m->v2f_ptr, &m->v2f));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  (void)V;
  return MV_OK;
}

}  // namespace mv

using namespace mv;

// P1 vertex -> its coordinate node in the last cell that contains it (periodic meshes)
__global__ void k_dg_rep(int64_t nnodes, const int* __restrict__ cells_flat, int* __restrict__ rep) {
  for (int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; n < nnodes; n += (int64_t)gridDim.x * blockDim.x)
    atomicMax(&rep[cells_flat[n]], (int)n);
}

extern "C" int mv_mesh_create(int dim, int64_t V, int64_t C, const double* coords, const int32_t* cells, int64_t F,
                              const int32_t* bfacet_verts, const int32_t* bfacet_tags, int64_t n_owned_vertices,
                              int64_t n_owned_cells, mv_mesh** out) {
  return mv_mesh_create_ex(dim, V, C, coords, cells, F, bfacet_verts, bfacet_tags, n_owned_vertices, n_owned_cells, 0, out);
}

extern "C" int mv_mesh_create_ex(int dim, int64_t V, int64_t C, const double* coords, const int32_t* cells, int64_t F,
                                 const int32_t* bfacet_verts, const int32_t* bfacet_tags, int64_t n_owned_vertices,
                                 int64_t n_owned_cells, int coord_dg, mv_mesh** out) {
  if (!out) return MV_ERR_BADARG;
  *out = nullptr;
  if (coord_dg && (F != 0 || n_owned_vertices != V || n_owned_cells != C)) {
    set_error("mv_mesh_create_ex: a DG coordinate field (periodic mesh) has no boundary facets and is not partitioned");
    return MV_ERR_UNSUPPORTED;
  }
  if (dim < 1 || dim > 3) {
    set_error("mv_mesh_create: dim must be 1, 2 or 3 (got %d)", dim);
    return MV_ERR_UNSUPPORTED;
  }
  if (V <= 0 || C < 0 || F < 0 || !coords || (C > 0 && !cells) || n_owned_vertices < 0 || n_owned_vertices > V ||
      n_owned_cells < 0 || n_owned_cells > C) {
    set_error("mv_mesh_create: bad sizes / null arrays");
    return MV_ERR_BADARG;
  }
  if (C * (int64_t)(dim + 1) >= (int64_t)INT32_MAX || V * (int64_t)MV_MAX_COMP >= (int64_t)4 * INT32_MAX) {
    set_error("mv_mesh_create: mesh too large for 32-bit connectivity offsets");
    return MV_ERR_UNSUPPORTED;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    set_error("mv_mesh_create: no CUDA device (there is no CPU fallback)");
    return MV_ERR_CUDA;
  }
  if (g_nsm == 0) {
    int dev = 0;
    MV_CUDA(cudaGetDevice(&dev));
    MV_CUDA(cudaDeviceGetAttribute(&g_nsm, cudaDevAttrMultiProcessorCount, dev));
  }
  mv_mesh* m = new mv_mesh();
  m->dim = dim; m->V = V; m->C = C; m->F = F; m->Vown = n_owned_vertices; m->Cown = n_owned_cells;
  m->dg = coord_dg ? 1 : 0;
  m->Vc = coord_dg ? C * (int64_t)(dim + 1) : V;
  int st = MV_OK;
  auto fail = [&](int s) { mv_mesh_destroy(m); return s; };
  if ((st = dev_alloc(&m->xi, m->Vc * dim)) != MV_OK) return fail(st);
  if ((st = dev_alloc(&m->cells, C * (dim + 1))) != MV_OK) return fail(st);
  if ((st = dev_alloc(&m->bf_verts, F * dim)) != MV_OK) return fail(st);
  if ((st = dev_alloc(&m->bf_tags, F)) != MV_OK) return fail(st);
  cudaError_t e = cudaMemcpyAsync(m->xi, coords, m->Vc * dim * sizeof(double), cudaMemcpyDefault, g_stream);
  if (e == cudaSuccess && C) e = cudaMemcpyAsync(m->cells, cells, C * (dim + 1) * sizeof(int), cudaMemcpyDefault, g_stream);
  if (e == cudaSuccess && F) e = cudaMemcpyAsync(m->bf_verts, bfacet_verts, F * dim * sizeof(int), cudaMemcpyDefault, g_stream);
  if (e == cudaSuccess && F) e = cudaMemcpyAsync(m->bf_tags, bfacet_tags, F * sizeof(int), cudaMemcpyDefault, g_stream);
  if (e != cudaSuccess) {
    set_error("mv_mesh_create: copy failed: %s", cudaGetErrorString(e));
    return fail(MV_ERR_CUDA);
  }
  m->h_bf_tags.resize(F);
  if (F) {
    e = cudaMemcpyAsync(m->h_bf_tags.data(), m->bf_tags, F * sizeof(int), cudaMemcpyDeviceToHost, g_stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(g_stream);
    if (e != cudaSuccess) { set_error("mv_mesh_create: %s", cudaGetErrorString(e)); return fail(MV_ERR_CUDA); }
  }
  if (m->dg) {
    if ((st = dev_alloc(&m->rep, V)) != MV_OK) return fail(st);
    if (cudaMemsetAsync(m->rep, 0, V * sizeof(int), g_stream) != cudaSuccess) return fail(MV_ERR_CUDA);
    MV_L k_dg_rep<<<grid_for(m->Vc), MV_BLOCK, 0, g_stream>>>(m->Vc, m->cells, m->rep);
  }
  st = dim == 1 ? ingest_impl<1>(m) : (dim == 2 ? ingest_impl<2>(m) : ingest_impl<3>(m));
  if (st != MV_OK) return fail(st);
  *out = m;
  return MV_OK;
}

extern "C" void mv_mesh_destroy(mv_mesh* m) {
  if (!m) return;
  cudaStreamSynchronize(g_stream);
  dev_free(m->xi); dev_free(m->cells); dev_free(m->v2c_ptr); dev_free(m->v2c);
  dev_free(m->bf_verts); dev_free(m->bf_tags); dev_free(m->bf_cell); dev_free(m->bf_normal); dev_free(m->bf_area);
  dev_free(m->v2f_ptr); dev_free(m->v2f);
  dev_free(m->pat.slice_ptr); dev_free(m->pat.cols); dev_free(m->pat.row_nnz); dev_free(m->pat.sched);
  dev_free(m->Mval); dev_free(m->Kval); dev_free(m->MKval); dev_free(m->ml); dev_free(m->vol);
  dev_free(m->dinvM); dev_free(m->dinvK); dev_free(m->dinvMK); dev_free(m->edge_opp); dev_free(m->cell_gid); dev_free(m->rep);
  dev_free(m->red.partials); dev_free(m->red.counter);
  dev_free(m->pcg.r); dev_free(m->pcg.p); dev_free(m->pcg.q); dev_free(m->pcg.xs); dev_free(m->pcg.scal); dev_free(m->pcg.flags);
  for (int i = 0; i < 3; ++i) dev_free(m->pcg.f32[i]);
  dev_free(m->pcg.valf); dev_free(m->pcg.dinvf);
  mv::pcg_graph_release(m->pcg);
  if (m->pcg.h_flags) cudaFreeHost(m->pcg.h_flags);
  if (m->pcg.h_scal) cudaFreeHost(m->pcg.h_scal);
  dev_free(m->halo.send_idx); dev_free(m->halo.send_buf);
  dev_free(m->tmp_vec);
  dev_free(m->cell_tmp);
  mv::amg_release(m);
  mv::spring_pc_release(m);
  delete m;
}

extern "C" int mv_mesh_info(const mv_mesh* m, int64_t* info) {
  if (!m || !info) return MV_ERR_BADARG;
  info[0] = m->V; info[1] = m->C; info[2] = m->F; info[3] = m->pat.nnz; info[4] = m->pat.len;
  info[5] = m->pat.nslices; info[6] = m->pat.maxw; info[7] = m->dim; info[8] = m->Vown; info[9] = m->Cown;
  info[10] = m->pat.sched ? (m->pat.sched_tiled ? 2 : 1) : 0; info[11] = m->pat.sched_len;
  return MV_OK;
}

extern "C" int mv_mesh_scalars(const mv_mesh* m, double* s) {
  if (!m || !s) return MV_ERR_BADARG;
  s[0] = m->total_volume; s[1] = m->ml_sumsq; s[2] = m->ml_sum;
  return MV_OK;
}

extern "C" int mv_mesh_get(const mv_mesh* m, int what, void* dst) {
  if (!m || !dst) return MV_ERR_BADARG;
  const void* src = nullptr;
  size_t bytes = 0;
  const int d = m->dim;
  switch (what) {
    case MV_GET_ROW_NNZ: src = m->pat.row_nnz; bytes = m->Vown * sizeof(int); break;
    case MV_GET_SLICE_PTR: src = m->pat.slice_ptr; bytes = (m->pat.nslices + 1) * sizeof(int64_t); break;
    case MV_GET_COLS: src = m->pat.cols; bytes = m->pat.len * sizeof(int); break;
    case MV_GET_MASS: src = m->Mval; bytes = m->pat.len * sizeof(double); break;
    case MV_GET_STIFF: src = m->Kval; bytes = m->pat.len * sizeof(double); break;
    case MV_GET_LUMPED: src = m->ml; bytes = m->Vown * sizeof(double); break;
    case MV_GET_CELL_VOL: src = m->vol; bytes = m->C * sizeof(double); break;
    case MV_GET_BF_NORMAL: src = m->bf_normal; bytes = m->F * d * sizeof(double); break;
    case MV_GET_BF_AREA: src = m->bf_area; bytes = m->F * sizeof(double); break;
    case MV_GET_BF_CELL: src = m->bf_cell; bytes = m->F * sizeof(int); break;
    default: set_error("mv_mesh_get: unknown item %d", what); return MV_ERR_BADARG;
  }
  if (bytes == 0) return MV_OK;
  MV_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  return MV_OK;
}
