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
