// Gradient / Hessian recovery for the solution-based monitor builders.
//
// replaces: movement/monitor.py:232-240 (recover_gradient_l2) and :286-293 (recover_hessian_clement), both from the
//           third-party `animate.recovery` (not vendored, not pinned by the reference: parity of this file is anchored
//           on the reference's own known-answer tests, test/test_monitor.py:64-152, and on the test oracle's
//           restatement of the published algorithms):
//             recover_gradient_l2(f)     = project(grad f, P1_vec): consistent mass matrix, cellwise-constant gradient
//             clement_interpolant(P0 g)  = volume-weighted mean over the cells around a vertex
//             recover_hessian_clement(f) = g = clement(grad f), H = clement(grad g)         (f in P1)
// Both act on the mesh the operators were ingested on (mv_mesh::xi): the host mirror ingests the mesh the monitor is
// asked for.  The L2 projection reuses the A1 right-hand side kernel and the mass solver of the hot path.
#include "mv_internal.cuh"

namespace mv {

// out[v, c, a] = sum_{K ni v} |K| d_a u_c|_K / sum_{K ni v} |K|   (u interleaved [V, NC])
template <int D, int NC>
__global__ void __launch_bounds__(MV_BLOCK)
k_clement_gradient(const double* __restrict__ xi, const int* __restrict__ cells, int dg, const int* __restrict__ v2c_ptr,
                   const int* __restrict__ v2c, int64_t nrows, const double* __restrict__ u, double* __restrict__ out) {
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    double acc[NC][D], den = 0.0;
#pragma unroll
    for (int c = 0; c < NC; ++c)
#pragma unroll
      for (int a = 0; a < D; ++a) acc[c][a] = 0.0;
    const int tb = __ldg(&v2c_ptr[v]), te = __ldg(&v2c_ptr[v + 1]);
    for (int t = tb; t < te; ++t) {
      const int64_t cell = __ldg(&v2c[t]) / (D + 1);
      int cv[D + 1];
      double xc[D + 1][D];
      CellGeom<D> g;
      load_cell<D>(cells, cell, cv);
      load_cell_coords<D>(xi, cv, xc, dg, cell);
      cell_geom(xc, g);
      den += g.vol;
#pragma unroll
      for (int c = 0; c < NC; ++c)
#pragma unroll
        for (int a = 0; a < D; ++a) {
          double gk = 0.0;
#pragma unroll
          for (int j = 0; j <= D; ++j) gk = fma(__ldg(&u[(size_t)cv[j] * NC + c]), g.g[j][a], gk);
          acc[c][a] = fma(g.vol, gk, acc[c][a]);
        }
    }
#pragma unroll
    for (int c = 0; c < NC; ++c)
#pragma unroll
      for (int a = 0; a < D; ++a) out[((size_t)v * NC + c) * D + a] = acc[c][a] / den;
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Projection of a P1 field between two meshes of the same domain (replaces: movement/monitor.py:191-201,
// `Function(fs).project(self.solution)`, Firedrake's supermesh projection -- third-party, unpinned).  Restated as the
// Galerkin L2 projection  M_t u_t = b,  b_i = int psi_i^t u_s dx,  with the integral taken by the (d+1)-point rule of
// degree 2 on every TARGET cell (points with one barycentric coordinate A_D and the others B_D, equal weights): exact
// when u_s is linear on the target cell -- so projecting onto the same mesh is the identity -- and second-order
// accurate where a target cell straddles source cells.  u_s(x_q) needs the source cell that contains x_q: a straight
// walk through face neighbours (found through the vertex->cell lists) that starts in the source cell with the target
// cell's number -- for a mover's moved copy of the solution's mesh that is a few steps.
// ---------------------------------------------------------------------------------------------------------------
template <int D> struct QuadRule;
template <> struct QuadRule<1> { static constexpr double A = 0.7886751345948129, B = 0.2113248654051871; };
template <> struct QuadRule<2> { static constexpr double A = 2.0 / 3.0, B = 1.0 / 6.0; };
template <> struct QuadRule<3> { static constexpr double A = 0.5854101966249685, B = 0.1381966011250105; };

#define MV_WALK_MAX_STEPS (1 << 20)
#define MV_WALK_TOL 1.0e-12

// neighbour of cell c across the face opposite local vertex j (-1: boundary face)
template <int D>
__device__ __forceinline__ int64_t face_neighbour(const int* __restrict__ cells, const int* __restrict__ v2c_ptr,
                                                  const int* __restrict__ v2c, int64_t c, const int (&cv)[D + 1], int j) {
  const int f0 = cv[(j + 1) % (D + 1)];
  for (int t = __ldg(&v2c_ptr[f0]), te = __ldg(&v2c_ptr[f0 + 1]); t < te; ++t) {
    const int64_t c2 = __ldg(&v2c[t]) / (D + 1);
    if (c2 == c) continue;
    int cw[D + 1];
    load_cell<D>(cells, c2, cw);
    bool all = true;
#pragma unroll
    for (int k = 0; k <= D; ++k) {
      if (k == j) continue;
      bool found = false;
#pragma unroll
      for (int l = 0; l <= D; ++l) found = found || (cw[l] == cv[k]);
      all = all && found;
    }
    if (all) return c2;
  }
  return -1;
}

// one thread per quadrature point of the target mesh: uq[c*(D+1)+q] = u_s(x_q); *outside counts the points the walk
// could not place inside a source cell (left the domain through a boundary face, or ran out of steps): those take the
// linear extension of the last cell visited
template <int D>
__global__ void __launch_bounds__(MV_BLOCK)
k_project_sample(const double* __restrict__ xt, const int* __restrict__ cells_t, int64_t Ct, const double* __restrict__ xs,
                 const int* __restrict__ cells_s, int64_t Cs, const int* __restrict__ v2c_ptr, const int* __restrict__ v2c,
                 const double* __restrict__ us, double* __restrict__ uq, unsigned long long* __restrict__ outside) {
  const int64_t n = Ct * (D + 1);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t ct = i / (D + 1);
    const int q = (int)(i - ct * (D + 1));
    int cv[D + 1];
    double xc[D + 1][D], p[D];
    load_cell<D>(cells_t, ct, cv);
    load_cell_coords<D>(xt, cv, xc);
#pragma unroll
    for (int a = 0; a < D; ++a) {
      double t = 0.0;
#pragma unroll
      for (int j = 0; j <= D; ++j) t += (j == q ? QuadRule<D>::A : QuadRule<D>::B) * xc[j][a];
      p[a] = t;
    }
    int64_t c = ct < Cs ? ct : 0;
    double lam[D + 1];
    bool inside = false;
    for (int step = 0; step < MV_WALK_MAX_STEPS; ++step) {
      CellGeom<D> g;
      load_cell<D>(cells_s, c, cv);
      load_cell_coords<D>(xs, cv, xc);
      cell_geom(xc, g);
      lam[0] = 1.0;
#pragma unroll
      for (int j = 1; j <= D; ++j) {
        double l = 0.0;
#pragma unroll
        for (int a = 0; a < D; ++a) l += g.g[j][a] * (p[a] - xc[0][a]);
        lam[j] = l;
        lam[0] -= l;
      }
      // leave through the face with the most negative coordinate that has a neighbour behind it
      bool tried[D + 1];
#pragma unroll
      for (int j = 0; j <= D; ++j) tried[j] = false;
      int64_t next = -1;
      bool any_negative = false;
      for (int round = 0; round <= D && next < 0; ++round) {
        int jm = -1;
        double lm = -MV_WALK_TOL;
#pragma unroll
        for (int j = 0; j <= D; ++j)
          if (!tried[j] && lam[j] < lm) { lm = lam[j]; jm = j; }
        if (jm < 0) break;
        any_negative = true;
        tried[jm] = true;
        next = face_neighbour<D>(cells_s, v2c_ptr, v2c, c, cv, jm);
      }
      if (!any_negative) { inside = true; break; }
      if (next < 0) break;   // only boundary faces in the way
      c = next;
    }
    if (!inside) atomicAdd(outside, 1ull);
    double val = 0.0;
#pragma unroll
    for (int j = 0; j <= D; ++j) val += lam[j] * __ldg(&us[cv[j]]);
    uq[i] = val;
  }
}

// b_i = sum_{K ni i} |K| / (d+1) sum_q lambda_i(x_q) u_q  -- gathered per target vertex (no atomics, fixed order)
template <int D>
__global__ void __launch_bounds__(MV_BLOCK)
k_project_rhs(const double* __restrict__ xt, const int* __restrict__ cells_t, const int* __restrict__ v2c_ptr,
              const int* __restrict__ v2c, int64_t nrows, const double* __restrict__ uq, double* __restrict__ b) {
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    double acc = 0.0;
    for (int t = __ldg(&v2c_ptr[v]), te = __ldg(&v2c_ptr[v + 1]); t < te; ++t) {
      const int e = __ldg(&v2c[t]);
      const int64_t cell = e / (D + 1);
      const int loc = e - (int)cell * (D + 1);
      int cv[D + 1];
      double xc[D + 1][D];
      CellGeom<D> g;
      load_cell<D>(cells_t, cell, cv);
      load_cell_coords<D>(xt, cv, xc);
      cell_geom(xc, g);
      double sq = 0.0;
#pragma unroll
      for (int q = 0; q <= D; ++q) sq += (q == loc ? QuadRule<D>::A : QuadRule<D>::B) * __ldg(&uq[cell * (D + 1) + q]);
      acc += g.vol * (1.0 / (D + 1)) * sq;
    }
    b[v] = acc;
  }
}

template <int D>
static int project_t(mv_mesh* src, const double* us, mv_mesh* tgt, double* uq, double* b, unsigned long long* d_out) {
  MV_L k_project_sample<D><<<grid_for(tgt->C * (D + 1)), MV_BLOCK, 0, g_stream>>>(tgt->xi, tgt->cells, tgt->C, src->xi, src->cells, src->C,
                                                                            src->v2c_ptr, src->v2c, us, uq, d_out);
  MV_KERNEL_CHECK();
  MV_L k_project_rhs<D><<<grid_for(tgt->V), MV_BLOCK, 0, g_stream>>>(tgt->xi, tgt->cells, tgt->v2c_ptr, tgt->v2c, tgt->V, uq, b);
  MV_KERNEL_CHECK();
  return MV_OK;
}

template <int D>
static int clement_t(mv_mesh* m, int ncomp, const double* u, double* out) {
  const int grid = grid_for(m->Vown);
  if (ncomp == 1)
    MV_L k_clement_gradient<D, 1><<<grid, MV_BLOCK, 0, g_stream>>>(m->xi, m->cells, m->dg, m->v2c_ptr, m->v2c, m->Vown, u, out);
  else if (ncomp == D)
    MV_L k_clement_gradient<D, D><<<grid, MV_BLOCK, 0, g_stream>>>(m->xi, m->cells, m->dg, m->v2c_ptr, m->v2c, m->Vown, u, out);
  else {
    set_error("mv_clement_gradient: ncomp must be 1 or the mesh dimension");
    return MV_ERR_BADARG;
  }
  MV_KERNEL_CHECK();
  return MV_OK;
}

}  // namespace mv

using namespace mv;

extern "C" int mv_clement_gradient(mv_mesh* m, int ncomp, double* u, double* out) {
  if (!m || !u || !out) return MV_ERR_BADARG;
  if (!is_device_ptr(u) || !is_device_ptr(out)) {
    set_error("mv_clement_gradient: device pointers required");
    return MV_ERR_BADARG;
  }
  if (comm_active()) MV_CHECK(halo_exchange(m, ncomp, u));
  MV_DISPATCH_DIM(m->dim, MV_CHECK(clement_t<DIM>(m, ncomp, u, out)));
  if (comm_active()) MV_CHECK(halo_exchange(m, ncomp * m->dim, out));
  return MV_OK;
}

extern "C" int mv_recover_gradient_l2(mv_mesh* m, double* u, double* g, double rtol, int maxit, int* iters) {
  if (!m || !u || !g) return MV_ERR_BADARG;
  if (!is_device_ptr(u) || !is_device_ptr(g)) {
    set_error("mv_recover_gradient_l2: device pointers required");
    return MV_ERR_BADARG;
  }
  if (!m->tmp_vec) MV_CHECK(dev_alloc(&m->tmp_vec, m->V * (size_t)(m->dim * m->dim)));
  if (comm_active()) MV_CHECK(halo_exchange(m, 1, u));
  MV_CHECK(l2_rhs_dev(m, u, m->tmp_vec));
  MV_CUDA(cudaMemsetAsync(g, 0, (size_t)m->V * m->dim * sizeof(double), g_stream));
  MV_CHECK(pcg_solve(m, MV_OP_MASS, m->dim, m->tmp_vec, g, nullptr, rtol, 1e-300, maxit, 0, iters, nullptr));
  if (comm_active()) MV_CHECK(halo_exchange(m, m->dim, g));
  return MV_OK;
}

extern "C" int mv_project_p1(mv_mesh* src, const double* u_src, mv_mesh* tgt, double* u_tgt, double rtol, int maxit, int* iters,
                             int64_t* n_outside) {
  if (!src || !tgt || !u_src || !u_tgt) return MV_ERR_BADARG;
  if (!is_device_ptr(u_src) || !is_device_ptr(u_tgt)) {
    set_error("mv_project_p1: device pointers required");
    return MV_ERR_BADARG;
  }
  if (src->dim != tgt->dim || src->dg || tgt->dg) {
    set_error("mv_project_p1: both meshes must have the same dimension and P1 (not periodic DG) coordinates");
    return MV_ERR_UNSUPPORTED;
  }
  if (comm_active()) {
    set_error("mv_project_p1: a quadrature point may lie in another rank's part of the source mesh; not supported on a partitioned mesh");
    return MV_ERR_UNSUPPORTED;
  }
  double *uq = nullptr, *b = nullptr;
  unsigned long long* d_out = nullptr;
  MV_CHECK(dev_alloc(&uq, (size_t)tgt->C * (tgt->dim + 1)));
  MV_CHECK(dev_alloc(&b, (size_t)tgt->V));
  MV_CHECK(dev_alloc(&d_out, 1));
  MV_CUDA(cudaMemsetAsync(d_out, 0, sizeof(unsigned long long), g_stream));
  int st = MV_OK;
  MV_DISPATCH_DIM(tgt->dim, st = project_t<DIM>(src, u_src, tgt, uq, b, d_out));
  if (st == MV_OK) {
    cudaMemsetAsync(u_tgt, 0, (size_t)tgt->V * sizeof(double), g_stream);
    st = pcg_solve(tgt, MV_OP_MASS, 1, b, u_tgt, nullptr, rtol, 1e-300, maxit, 0, iters, nullptr);
  }
  unsigned long long h = 0;
  cudaMemcpyAsync(&h, d_out, sizeof(h), cudaMemcpyDeviceToHost, g_stream);
  cudaStreamSynchronize(g_stream);
  if (n_outside) *n_outside = (int64_t)h;
  dev_free(uq);
  dev_free(b);
  dev_free(d_out);
  return st;
}
