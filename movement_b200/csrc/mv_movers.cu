// Laplacian smoothing and lineal springs.
#include "mv_internal.cuh"

namespace mv {

// ------------------------------------------------------------------ Laplacian smoothing
// replaces: movement/laplacian_smoothing.py:64-77 (K v = 0 with Dirichlet data) and
//           :116-117 (x += v dt).  The constrained dofs of v hold the boundary values; the
//           free dofs are zeroed so that the initial residual is exactly -K_fb g.
template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_zero_free(int64_t nrows, const uint16_t* __restrict__ mask, double* __restrict__ v) {
  const int64_t n = nrows * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t vert = i / C;
    const int c = (int)(i - vert * C);
    if (!((mask[vert] >> c) & 1)) v[i] = 0.0;
  }
}

// x += s*v ; disp = s*v (optional) ; sum (s*v)^2
__global__ void __launch_bounds__(MV_BLOCK)
k_axpy_norm(int64_t n, double s, const double* __restrict__ v, double* __restrict__ x, double* __restrict__ disp,
            double* __restrict__ partials, unsigned* __restrict__ counter, double* __restrict__ out) {
  double acc[1] = {0.0};
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double d = s * v[i];
    x[i] += d;
    if (disp) disp[i] = d;
    acc[0] = fma(d, d, acc[0]);
  }
  grid_reduce_sum<1>(acc, partials, counter, [&](double(&t)[1]) { out[0] = t[0]; });
}

// ------------------------------------------------------------------ lineal springs
// Edge (i,j) geometry exactly as the reference obtains it from HDiv-Trace solves
// (movement/spring.py:70-146): length L, unit tangent t = perp(n+) = (-n_y, n_x) with n+ the
// outward normal of the edge's '+' cell (the lowest-numbered incident cell; on the boundary
// the only cell), angle = arccos(clamp(t_x)).  The 2x2 block is
// (1/L) [[c^2, sc], [sc, s^2]] with c = cos(angle), s = sin(angle) >= 0 (spring.py:160-180).
struct EdgeBlock {
  double c2, sc, s2, L, tx, ty, angle;
};

__device__ __forceinline__ EdgeBlock edge_block(const double* __restrict__ X, int i, int j, int opp) {
  // endpoints ordered by vertex number so that (i,j) and (j,i) give identical bits
  const int a = i < j ? i : j, b = i < j ? j : i;
  const double ax = __ldg(&X[2 * (int64_t)a]), ay = __ldg(&X[2 * (int64_t)a + 1]);
  const double bx = __ldg(&X[2 * (int64_t)b]), by = __ldg(&X[2 * (int64_t)b + 1]);
  const double ox = __ldg(&X[2 * (int64_t)opp]), oy = __ldg(&X[2 * (int64_t)opp + 1]);
  const double ex = bx - ax, ey = by - ay;
  EdgeBlock e;
  e.L = sqrt(ex * ex + ey * ey);
  double nx = ey / e.L, ny = -ex / e.L;
  if (nx * (ax - ox) + ny * (ay - oy) < 0.0) {
    nx = -nx;
    ny = -ny;
  }
  e.tx = -ny;
  e.ty = nx;
  e.angle = acos(fmin(fmax(e.tx, -1.0), 1.0));
  const double c = cos(e.angle), s = sin(e.angle);
  e.c2 = c * c / e.L;
  e.sc = s * c / e.L;
  e.s2 = s * s / e.L;
  return e;
}

// opposite vertex of the '+' cell for every stored (row, col) pair, col != row
__global__ void __launch_bounds__(MV_BLOCK)
k_edge_opp(const int* __restrict__ cells, const int* __restrict__ v2c_ptr, const int* __restrict__ v2c, int64_t nrows,
           const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols, const int* __restrict__ row_nnz,
           const int64_t* __restrict__ cell_gid, int* __restrict__ opp) {
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = v / MV_SLICE;
    const int64_t base = slice_ptr[s] + (v % MV_SLICE);
    const int w = (int)((slice_ptr[s + 1] - slice_ptr[s]) / MV_SLICE);
    const int nn = row_nnz[v];
    for (int k = 0; k < w; ++k) {
      const int64_t pos = base + (int64_t)k * MV_SLICE;
      const int u = cols[pos];
      int o = -1;
      if (k < nn && u != v) {
        // the lowest-numbered cell containing both v and u.  One GPU: cells are listed ascending per vertex, the first hit
        // wins.  Partitioned: the local cell order (owned cells first) is not the global one, so the hit with the smallest
        // GLOBAL cell number wins (cell_gid, mv_mesh_set_cell_order) -- every rank then picks the same '+' cell.
        int64_t best = -1;
        for (int t = v2c_ptr[v]; t < v2c_ptr[v + 1] && (o < 0 || cell_gid); ++t) {
          const int64_t c = v2c[t] / 3;
          const int c0 = cells[c * 3], c1 = cells[c * 3 + 1], c2 = cells[c * 3 + 2];
          if (c0 == u || c1 == u || c2 == u) {
            const int64_t g = cell_gid ? cell_gid[c] : c;
            if (o < 0 || g < best) {
              best = g;
              o = (c0 != u && c0 != v) ? c0 : ((c1 != u && c1 != v) ? c1 : c2);
            }
          }
        }
      }
      opp[pos] = o;
    }
  }
}

// y = K(X) x with Dirichlet rows replaced by identity, plus (optionally) the partial x.y
template <bool DOT>
__global__ void __launch_bounds__(MV_BLOCK)
k_spring_apply(int64_t nslices, int64_t nrows, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
               const int* __restrict__ opp, const double* __restrict__ X, const uint16_t* __restrict__ mask,
               const double* __restrict__ x, double* __restrict__ y, const int* __restrict__ done,
               double* __restrict__ partials, unsigned* __restrict__ counter, double* __restrict__ scal) {
  if (DOT && done != nullptr && *done) return;
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  double dot[2] = {0.0, 0.0};
  for (int64_t s = warp0; s < nslices; s += nwarps) {
    const int64_t base = __ldg(&slice_ptr[s]);
    const int w = (int)((__ldg(&slice_ptr[s + 1]) - base) >> 5);
    const int64_t row = s * MV_SLICE + lane;
    if (row >= nrows) continue;
    const double xi0 = __ldg(&x[2 * row]), xi1 = __ldg(&x[2 * row + 1]);
    double a0 = 0.0, a1 = 0.0;
    for (int k = 0; k < w; ++k) {
      const int64_t pos = base + lane + (int64_t)k * MV_SLICE;
      const int o = __ldcs(&opp[pos]);
      if (o < 0) continue;
      const int col = __ldcs(&cols[pos]);
      const EdgeBlock e = edge_block(X, (int)row, col, o);
      const double d0 = xi0 - __ldg(&x[2 * (int64_t)col]), d1 = xi1 - __ldg(&x[2 * (int64_t)col + 1]);
      a0 += e.c2 * d0 + e.sc * d1;
      a1 += e.sc * d0 + e.s2 * d1;
    }
    const unsigned mk = mask ? __ldg(&mask[row]) : 0u;
    if (mk & 1u) a0 = DOT ? 0.0 : xi0;
    if (mk & 2u) a1 = DOT ? 0.0 : xi1;
    y[2 * row] = a0;
    y[2 * row + 1] = a1;
    if (DOT) {
      dot[0] = fma(a0, xi0, dot[0]);
      dot[1] = fma(a1, xi1, dot[1]);
    }
  }
  if (DOT) {
    grid_reduce_sum<2>(dot, partials, counter, [&](double(&t)[2]) {
      const double tot = t[0] + t[1];
      scal[SC(S_PQ, 0)] = tot;
      scal[SC(S_PQ, 1)] = tot;
    });
  }
}

// inverse diagonal per dof
__global__ void __launch_bounds__(MV_BLOCK)
k_spring_diag(int64_t nrows, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
              const int* __restrict__ opp, const double* __restrict__ X, double* __restrict__ dinv) {
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = v / MV_SLICE;
    const int64_t base = slice_ptr[s] + (v % MV_SLICE);
    const int w = (int)((slice_ptr[s + 1] - slice_ptr[s]) / MV_SLICE);
    double d0 = 0.0, d1 = 0.0;
    for (int k = 0; k < w; ++k) {
      const int64_t pos = base + (int64_t)k * MV_SLICE;
      const int o = opp[pos];
      if (o < 0) continue;
      const EdgeBlock e = edge_block(X, (int)v, cols[pos], o);
      d0 += e.c2;
      d1 += e.s2;
    }
    dinv[2 * v] = d0 > 0.0 ? 1.0 / d0 : 1.0;
    dinv[2 * v + 1] = d1 > 0.0 ? 1.0 / d1 : 1.0;
  }
}

// edge list export: edges with col > row, in (row, col) order
__global__ void k_edge_count(int64_t nrows, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
                             const int* __restrict__ opp, int* __restrict__ cnt) {
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = v / MV_SLICE;
    const int64_t base = slice_ptr[s] + (v % MV_SLICE);
    const int w = (int)((slice_ptr[s + 1] - slice_ptr[s]) / MV_SLICE);
    int n = 0;
    for (int k = 0; k < w; ++k) {
      const int64_t pos = base + (int64_t)k * MV_SLICE;
      if (opp[pos] >= 0 && cols[pos] > v) ++n;
    }
    cnt[v] = n;
  }
}
__global__ void k_edge_fill(int64_t nrows, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
                            const int* __restrict__ opp, const int* __restrict__ ptr, const double* __restrict__ X,
                            int* __restrict__ edges, double* __restrict__ length, double* __restrict__ tangent,
                            double* __restrict__ angle) {
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = v / MV_SLICE;
    const int64_t base = slice_ptr[s] + (v % MV_SLICE);
    const int w = (int)((slice_ptr[s + 1] - slice_ptr[s]) / MV_SLICE);
    int e = ptr[v];
    for (int k = 0; k < w; ++k) {
      const int64_t pos = base + (int64_t)k * MV_SLICE;
      const int col = cols[pos];
      if (opp[pos] < 0 || col <= v) continue;
      const EdgeBlock b = edge_block(X, (int)v, col, opp[pos]);
      if (edges) { edges[2 * e] = (int)v; edges[2 * e + 1] = col; }
      if (length) length[e] = b.L;
      if (tangent) { tangent[2 * e] = b.tx; tangent[2 * e + 1] = b.ty; }
      if (angle) angle[e] = b.angle;
      ++e;
    }
  }
}

static int ensure_edge_opp(mv_mesh* m) {
  if (m->dim != 2) {
    set_error("Spring-based Movers only currently support triangular meshes.");
    return MV_ERR_UNSUPPORTED;
  }
  if (m->edge_opp) return MV_OK;
  MV_CHECK(dev_alloc(&m->edge_opp, m->pat.len));
  MV_L k_edge_opp<<<grid_for(m->Vown), MV_BLOCK, 0, g_stream>>>(m->cells, m->v2c_ptr, m->v2c, m->Vown, m->pat.slice_ptr, m->pat.cols,
                                                         m->pat.row_nnz, m->cell_gid, m->edge_opp);
  MV_KERNEL_CHECK();
  return MV_OK;
}

int spring_apply_dev(mv_mesh* m, const double* coords, const uint16_t* mask, const double* x, double* y, bool dot) {
  const Sell& p = m->pat;
  const int grid = grid_for(p.nslices * MV_SLICE, 8);
  if (dot)
    MV_L k_spring_apply<true><<<grid, MV_BLOCK, 0, g_stream>>>(p.nslices, p.nrows, p.slice_ptr, p.cols, m->edge_opp, coords, mask, x, y,
                                                          m->pcg.flags, m->red.partials, m->red.counter, m->pcg.scal);
  else
    MV_L k_spring_apply<false><<<grid, MV_BLOCK, 0, g_stream>>>(p.nslices, p.nrows, p.slice_ptr, p.cols, m->edge_opp, coords, mask, x,
                                                           y, nullptr, nullptr, nullptr, nullptr);
  MV_KERNEL_CHECK();
  return MV_OK;
}

}  // namespace mv

using namespace mv;

extern "C" int mv_laplace_move(mv_mesh* m, double* v, const uint16_t* mask, double dt, double* x, double rtol, double atol,
                               int maxit, int* iters, double* displacement_norm) {
  if (!m || !v || !mask || !x) return MV_ERR_BADARG;
  if (!is_device_ptr(v) || !is_device_ptr(x)) {
    set_error("mv_laplace_move: v and x must be device pointers");
    return MV_ERR_BADARG;
  }
  Staged sm;
  MV_CHECK(sm.in(mask, m->Vown * sizeof(uint16_t)));
  const int d = m->dim;
  const int64_t n = m->Vown;
  if (d == 2) MV_L k_zero_free<2><<<grid_for(n * d), MV_BLOCK, 0, g_stream>>>(n, sm.as<uint16_t>(), v);
  else MV_L k_zero_free<3><<<grid_for(n * d), MV_BLOCK, 0, g_stream>>>(n, sm.as<uint16_t>(), v);
  MV_KERNEL_CHECK();
  int st = pcg_solve(m, MV_OP_STIFFNESS, d, nullptr, v, sm.as<uint16_t>(), rtol, atol, maxit, 0, iters, nullptr);
  if (st != MV_OK) return st;
  double* dout = m->pcg.scal + SC(S_SUM, 8);
  MV_L k_axpy_norm<<<grid_for(n * d), MV_BLOCK, 0, g_stream>>>(n * d, dt, v, x, nullptr, m->red.partials, m->red.counter, dout);
  MV_KERNEL_CHECK();
  if (comm_active()) {
    MV_CHECK(allreduce_sum_dev(dout, 1));
    MV_CHECK(halo_exchange(m, d, x));
  }
  double h = 0.0;
  MV_CUDA(cudaMemcpyAsync(&h, dout, sizeof(double), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  if (comm_active()) MV_CHECK(p2p_check_status());
  if (displacement_norm) *displacement_norm = sqrt(h);
  return MV_OK;
}

// global cell numbers of the local cells (partitioned meshes): the '+' cell of an edge is its lowest GLOBAL cell
extern "C" int mv_mesh_set_cell_order(mv_mesh* m, const int64_t* global_cell) {
  if (!m || !global_cell) return MV_ERR_BADARG;
  dev_free(m->cell_gid);
  dev_free(m->edge_opp);
  MV_CHECK(dev_alloc(&m->cell_gid, (size_t)m->C));
  MV_CUDA(cudaMemcpyAsync(m->cell_gid, global_cell, m->C * sizeof(int64_t), cudaMemcpyDefault, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  return MV_OK;
}

extern "C" int mv_spring_edge_count(mv_mesh* m, int64_t* n_edges) {
  if (!m || !n_edges) return MV_ERR_BADARG;
  *n_edges = (m->pat.nnz - m->Vown) / 2;
  return MV_OK;
}

extern "C" int mv_spring_edge_geometry(mv_mesh* m, const double* coords, int32_t* edges, double* length, double* tangent,
                                       double* angle) {
  if (!m || !coords) return MV_ERR_BADARG;
  MV_CHECK(ensure_edge_opp(m));
  const int64_t E = (m->pat.nnz - m->Vown) / 2;
  Staged sx, se, sl, st, sa;
  MV_CHECK(sx.in(coords, m->V * 2 * sizeof(double)));
  MV_CHECK(se.out_begin(edges, E * 2 * sizeof(int)));
  MV_CHECK(sl.out_begin(length, E * sizeof(double)));
  MV_CHECK(st.out_begin(tangent, E * 2 * sizeof(double)));
  MV_CHECK(sa.out_begin(angle, E * sizeof(double)));
  int *cnt = nullptr, *ptr = nullptr;
  MV_CHECK(dev_alloc(&cnt, m->Vown + 1));
  MV_CHECK(dev_alloc(&ptr, m->Vown + 1));
  MV_CUDA(cudaMemsetAsync(cnt, 0, (m->Vown + 1) * sizeof(int), g_stream));
  const int grid = grid_for(m->Vown);
  MV_L k_edge_count<<<grid, MV_BLOCK, 0, g_stream>>>(m->Vown, m->pat.slice_ptr, m->pat.cols, m->edge_opp, cnt);
  MV_CHECK(exclusive_scan_i32(cnt, ptr, m->Vown + 1));
  MV_L k_edge_fill<<<grid, MV_BLOCK, 0, g_stream>>>(m->Vown, m->pat.slice_ptr, m->pat.cols, m->edge_opp, ptr, sx.as<double>(),
                                              se.as<int>(), sl.as<double>(), st.as<double>(), sa.as<double>());
  MV_KERNEL_CHECK();
  MV_CHECK(se.out_end(edges, E * 2 * sizeof(int)));
  MV_CHECK(sl.out_end(length, E * sizeof(double)));
  MV_CHECK(st.out_end(tangent, E * 2 * sizeof(double)));
  MV_CHECK(sa.out_end(angle, E * sizeof(double)));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  dev_free(cnt);
  dev_free(ptr);
  return MV_OK;
}

extern "C" int mv_spring_apply(mv_mesh* m, const double* coords, const uint16_t* mask, const double* x, double* y) {
  if (!m || !coords || !x || !y) return MV_ERR_BADARG;
  MV_CHECK(ensure_edge_opp(m));
  Staged sc, sm, sx, sy;
  MV_CHECK(sc.in(coords, m->V * 2 * sizeof(double)));
  MV_CHECK(sm.in(mask, m->Vown * sizeof(uint16_t)));
  MV_CHECK(sx.in(x, m->V * 2 * sizeof(double)));
  MV_CHECK(sy.out_begin(y, m->Vown * 2 * sizeof(double)));
  MV_CHECK(spring_apply_dev(m, sc.as<double>(), sm.as<uint16_t>(), sx.as<double>(), sy.as<double>(), false));
  return sy.out_end(y, m->Vown * 2 * sizeof(double));
}

extern "C" int mv_spring_move(mv_mesh* m, double* coords, const double* forcing, const uint16_t* mask, double dt,
                              double* displacement, double rtol, double atol, int maxit, int* iters, double* displacement_norm) {
  if (!m || !coords || !forcing || !mask || !displacement) return MV_ERR_BADARG;
  if (!is_device_ptr(coords) || !is_device_ptr(displacement)) {
    set_error("mv_spring_move: coords and displacement must be device pointers");
    return MV_ERR_BADARG;
  }
  MV_CHECK(ensure_edge_opp(m));
  MV_CHECK(ensure_pcg_workspace(m));
  const int64_t n = m->Vown;
  Staged sf, sm;
  MV_CHECK(sf.in(forcing, m->V * 2 * sizeof(double)));
  MV_CHECK(sm.in(mask, n * sizeof(uint16_t)));
  // u: Dirichlet dofs hold the forcing (= boundary displacement), free dofs start at zero
  if (!m->tmp_vec) MV_CHECK(dev_alloc(&m->tmp_vec, m->V * (size_t)(m->dim * m->dim)));
  double* u = m->tmp_vec;
  double* dinv = m->tmp_vec + m->V * 2;
  MV_CUDA(cudaMemcpyAsync(u, sf.dev, m->V * 2 * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
  MV_L k_zero_free<2><<<grid_for(n * 2), MV_BLOCK, 0, g_stream>>>(n, sm.as<uint16_t>(), u);
  MV_L k_spring_diag<<<grid_for(n), MV_BLOCK, 0, g_stream>>>(n, m->pat.slice_ptr, m->pat.cols, m->edge_opp, coords, dinv);
  MV_KERNEL_CHECK();
  // right-hand side = the forcing vector: its rows at Dirichlet dofs are replaced (u = forcing there), its rows at free
  // dofs are whatever earlier calls left in _forcing (spring.py:229-230 never clears it), normally zero
  int st = pcg_solve_spring(m, coords, dinv, sf.as<double>(), u, sm.as<uint16_t>(), rtol, atol, maxit, iters);
  if (st != MV_OK) return st;
  double* dout = m->pcg.scal + SC(S_SUM, 8);
  MV_L k_axpy_norm<<<grid_for(n * 2), MV_BLOCK, 0, g_stream>>>(n * 2, dt, u, coords, displacement, m->red.partials, m->red.counter,
                                                         dout);
  MV_KERNEL_CHECK();
  if (comm_active()) {
    MV_CHECK(allreduce_sum_dev(dout, 1));
    MV_CHECK(halo_exchange(m, 2, coords));   // the next call's stiffness reads the ghost coordinates
  }
  double h = 0.0;
  MV_CUDA(cudaMemcpyAsync(&h, dout, sizeof(double), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  if (comm_active()) MV_CHECK(p2p_check_status());
  if (displacement_norm) *displacement_norm = sqrt(h);
  return MV_OK;
}
