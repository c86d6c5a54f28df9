// Non-axis-aligned free boundary segments of the L2 projector: the EquationBC branch.
//
// replaces: movement/monge_ampere.py:274-309 -- for a straight / planar segment that is neither fixed nor axis-aligned
//           the reference imposes two EquationBCs on the segment's rows of the projection M g = int v . grad(phi_old):
//             bc1:  int (v.n)(u.n) ds = 0                                   (no movement normal to the boundary)
//             bc2:  int v_t . u_t ds = int v_t . grad(phi_old)_t ds         (tangential movement), with the nodes the
//                   segment shares with any other segment pinned to zero (nested DirichletBC `ffacet_bc`)
//           Firedrake adds the forms of successive EquationBCs into the same rows, and n n^T + (I - n n^T) = I, so the
//           rows of the segment read  int v . u ds = int v . P_t grad(phi_old) ds  -- the L2(Gamma) projection of the
//           tangential gradient with the boundary mass matrix.  Its unique solution is zero at the shared ("corner")
//           nodes and has u.n = 0 on the whole segment; the rows of all other vertices see it as Dirichlet data.
//
// Here: set-up (host, once; boundary-sized) numbers the nodes of all such segments, assembles the boundary mass
// matrices (facet matrix area / (d (d+1)) (1 + delta_ab)) over the non-corner nodes in CSR, and the node -> facet
// lists of the right-hand side.  Every projection then runs, on the device,
//   k_eq_rhs      r_i = sum_{F ni i} area(F)/d  P_t(F) grad(phi_old)|_cell(F)          (gather, no atomics)
//   k_eq_cheb     Chebyshev iteration on D^-1 M_Gamma over Wathen's interval [1/2, 1 + (d-1)/2] for P1 facets
//   k_eq_check    measured ||r - M u|| / ||r||  (a fixed number of iterations is run; the bound 2 rho^k is 1e-15)
//   k_eq_scatter  grad_phi at the segment's nodes <- u (corners: 0)
// and the mass solve that follows keeps those rows (all components masked).  One GPU only: a partitioned segment
// would need its own halo plan; mv_ma_set_boundary_ex reports MV_ERR_UNSUPPORTED on every rank in that case.
#include <algorithm>
#include <math.h>

#include <vector>

#include "mv_internal.cuh"

namespace mv {

struct EqBC {
  int64_t nb = 0, nu = 0;        // nodes of all EquationBC segments (unknowns first), unknown (non-corner) nodes
  int* node = nullptr;           // [nb] vertex number
  int64_t* mptr = nullptr;       // boundary mass matrix over the unknowns, CSR
  int* mcol = nullptr;
  double* mval = nullptr;
  double* dinv = nullptr;        // [nu]
  int64_t* fptr = nullptr;       // unknown -> facets of its segment
  int* flist = nullptr;
  double *rhs = nullptr, *u[2] = {nullptr, nullptr};   // [nu*d]
  double* norms = nullptr;       // device [2]: ||r - M u||^2, ||r||^2
  double* h_norms = nullptr;     // pinned
  int iters = 0;
};

void eqbc_free(mv_ma* ma) {
  EqBC* e = ma->eqbc;
  if (!e) return;
  dev_free(e->node); dev_free(e->mptr); dev_free(e->mcol); dev_free(e->mval); dev_free(e->dinv); dev_free(e->fptr);
  dev_free(e->flist); dev_free(e->rhs); dev_free(e->u[0]); dev_free(e->u[1]); dev_free(e->norms);
  if (e->h_norms) cudaFreeHost(e->h_norms);
  delete e;
  ma->eqbc = nullptr;
}

template <typename T>
static int upload(const std::vector<T>& h, T** d) {
  MV_CHECK(dev_alloc(d, h.size() ? h.size() : 1));
  if (h.size()) MV_CUDA(cudaMemcpyAsync(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, g_stream));
  return MV_OK;
}

// tagidx[f]: index of the facet's tag; is_eq[k]: tag k takes the EquationBC branch.  Host arrays of the F facets.
int eqbc_build(mv_ma* ma, const std::vector<int>& tagidx, const std::vector<char>& is_eq, const std::vector<int>& verts,
               const std::vector<double>& area) {
  eqbc_free(ma);
  mv_mesh* m = ma->mesh;
  const int d = m->dim;
  const int64_t F = m->F;
  // (vertex, tag) incidences of the whole boundary
  std::vector<std::pair<int, int>> vt;
  vt.reserve((size_t)F * d);
  for (int64_t f = 0; f < F; ++f)
    for (int a = 0; a < d; ++a) vt.emplace_back(verts[f * d + a], tagidx[f]);
  std::sort(vt.begin(), vt.end());
  vt.erase(std::unique(vt.begin(), vt.end()), vt.end());
  // nodes of EquationBC segments: unknown if the vertex lies on exactly one segment, corner otherwise
  std::vector<int> unk, cor;
  for (size_t i = 0; i < vt.size();) {
    size_t j = i;
    bool eq = false;
    while (j < vt.size() && vt[j].first == vt[i].first) { eq |= is_eq[vt[j].second] != 0; ++j; }
    if (eq) (j - i == 1 ? unk : cor).push_back(vt[i].first);
    i = j;
  }
  EqBC* e = new EqBC();
  ma->eqbc = e;
  e->nu = (int64_t)unk.size();
  e->nb = e->nu + (int64_t)cor.size();
  std::vector<int> node(unk);
  node.insert(node.end(), cor.begin(), cor.end());
  auto loc = [&](int v) -> int64_t {   // unknown number of vertex v, -1 if none (unk is sorted)
    auto it = std::lower_bound(unk.begin(), unk.end(), v);
    return (it != unk.end() && *it == v) ? (int64_t)(it - unk.begin()) : -1;
  };
  // boundary mass matrix and facet lists: count, then fill (facet order = ascending f: deterministic sums)
  std::vector<int64_t> mcnt(e->nu + 1, 0), fcnt(e->nu + 1, 0);
  std::vector<int64_t> lv((size_t)F * d, -1);
  for (int64_t f = 0; f < F; ++f) {
    if (!is_eq[tagidx[f]]) continue;
    for (int a = 0; a < d; ++a) lv[f * d + a] = loc(verts[f * d + a]);
    for (int a = 0; a < d; ++a) {
      if (lv[f * d + a] < 0) continue;
      ++fcnt[lv[f * d + a] + 1];
      for (int b = 0; b < d; ++b)
        if (lv[f * d + b] >= 0) ++mcnt[lv[f * d + a] + 1];
    }
  }
  for (int64_t i = 0; i < e->nu; ++i) { mcnt[i + 1] += mcnt[i]; fcnt[i + 1] += fcnt[i]; }
  // raw (row, col, val) triplets per row, then merge duplicates of a column
  std::vector<int> rcol(mcnt[e->nu]), flist(fcnt[e->nu]);
  std::vector<double> rval(mcnt[e->nu]);
  {
    std::vector<int64_t> mpos(mcnt.begin(), mcnt.end() - 1), fpos(fcnt.begin(), fcnt.end() - 1);
    for (int64_t f = 0; f < F; ++f) {
      if (!is_eq[tagidx[f]]) continue;
      const double w = area[f] / (double)(d * (d + 1));
      for (int a = 0; a < d; ++a) {
        const int64_t i = lv[f * d + a];
        if (i < 0) continue;
        flist[fpos[i]++] = (int)f;
        for (int b = 0; b < d; ++b) {
          const int64_t j = lv[f * d + b];
          if (j < 0) continue;
          rcol[mpos[i]] = (int)j;
          rval[mpos[i]++] = a == b ? 2.0 * w : w;
        }
      }
    }
  }
  std::vector<int64_t> mptr(e->nu + 1, 0);
  std::vector<int> mcol;
  std::vector<double> mval, dinv(e->nu, 0.0);
  mcol.reserve(rcol.size());
  mval.reserve(rcol.size());
  std::vector<std::pair<int, double>> row;
  for (int64_t i = 0; i < e->nu; ++i) {
    row.clear();
    for (int64_t t = mcnt[i]; t < mcnt[i + 1]; ++t) row.emplace_back(rcol[t], rval[t]);
    std::stable_sort(row.begin(), row.end(), [](const std::pair<int, double>& x, const std::pair<int, double>& y) { return x.first < y.first; });
    for (size_t t = 0; t < row.size();) {
      double s = 0.0;
      size_t q = t;
      while (q < row.size() && row[q].first == row[t].first) s += row[q++].second;
      mcol.push_back(row[t].first);
      mval.push_back(s);
      if (row[t].first == (int)i) dinv[i] = 1.0 / s;
      t = q;
    }
    mptr[i + 1] = (int64_t)mcol.size();
  }
  MV_CHECK(upload(node, &e->node));
  MV_CHECK(upload(mptr, &e->mptr));
  MV_CHECK(upload(mcol, &e->mcol));
  MV_CHECK(upload(mval, &e->mval));
  MV_CHECK(upload(dinv, &e->dinv));
  MV_CHECK(upload(fcnt, &e->fptr));
  MV_CHECK(upload(flist, &e->flist));
  const size_t nv = (size_t)(e->nu ? e->nu : 1) * d;
  MV_CHECK(dev_alloc(&e->rhs, nv));
  MV_CHECK(dev_alloc(&e->u[0], nv));
  MV_CHECK(dev_alloc(&e->u[1], nv));
  MV_CHECK(dev_alloc(&e->norms, 2));
  MV_CUDA(cudaMallocHost((void**)&e->h_norms, 2 * sizeof(double)));
  // iterations for 2 rho^k <= 1e-15 on [1/2, 1 + (d-1)/2]
  const double kappa = 2.0 + (d - 1);
  const double rho = (sqrt(kappa) - 1.0) / (sqrt(kappa) + 1.0);
  e->iters = (int)ceil(log(2e15) / -log(rho));
  MV_CUDA(cudaStreamSynchronize(g_stream));   // the host vectors go out of scope
  return MV_OK;
}

// right-hand side of the boundary projection at unknown node i
template <int D>
__global__ void __launch_bounds__(MV_BLOCK)
k_eq_rhs(int64_t nu, const int64_t* __restrict__ fptr, const int* __restrict__ flist, const int* __restrict__ bf_cell,
         const double* __restrict__ bf_normal, const double* __restrict__ bf_area, const double* __restrict__ xi,
         const int* __restrict__ cells, int dg, const double* __restrict__ phi, double* __restrict__ rhs) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nu; i += (int64_t)gridDim.x * blockDim.x) {
    double acc[D];
#pragma unroll
    for (int a = 0; a < D; ++a) acc[a] = 0.0;
    for (int64_t t = fptr[i]; t < fptr[i + 1]; ++t) {
      const int f = flist[t];
      const int64_t c = bf_cell[f];
      int cv[D + 1];
      double xc[D + 1][D];
      CellGeom<D> g;
      load_cell<D>(cells, c, cv);
      load_cell_coords<D>(xi, cv, xc, dg, c);
      cell_geom(xc, g);
      double gk[D], gn = 0.0;
#pragma unroll
      for (int a = 0; a < D; ++a) {
        gk[a] = 0.0;
#pragma unroll
        for (int j = 0; j <= D; ++j) gk[a] = fma(__ldg(&phi[cv[j]]), g.g[j][a], gk[a]);
        gn = fma(gk[a], bf_normal[(size_t)f * D + a], gn);
      }
      const double w = bf_area[f] / D;
#pragma unroll
      for (int a = 0; a < D; ++a) acc[a] = fma(w, gk[a] - gn * bf_normal[(size_t)f * D + a], acc[a]);
    }
#pragma unroll
    for (int a = 0; a < D; ++a) rhs[i * D + a] = acc[a];
  }
}

// one Chebyshev iteration: xout holds x_{k-1} on entry, x_{k+1} on exit (first: ak = 0 and xin = 0 are implied)
template <int D, bool FIRST>
__global__ void __launch_bounds__(MV_BLOCK)
k_eq_cheb(int64_t nu, const int64_t* __restrict__ mptr, const int* __restrict__ mcol, const double* __restrict__ mval,
          const double* __restrict__ dinv, const double* __restrict__ rhs, const double* __restrict__ xin,
          double* __restrict__ xout, double ak, double ck) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nu; i += (int64_t)gridDim.x * blockDim.x) {
    double acc[D];
#pragma unroll
    for (int a = 0; a < D; ++a) acc[a] = 0.0;
    if (!FIRST) {
      for (int64_t t = mptr[i]; t < mptr[i + 1]; ++t) {
        const double w = mval[t];
        const double* xp = xin + (size_t)mcol[t] * D;
#pragma unroll
        for (int a = 0; a < D; ++a) acc[a] = fma(w, xp[a], acc[a]);
      }
    }
    const double di = dinv[i];
#pragma unroll
    for (int a = 0; a < D; ++a) {
      const double xk = FIRST ? 0.0 : xin[i * D + a];
      const double z = di * (rhs[i * D + a] - acc[a]);
      xout[i * D + a] = FIRST ? ck * z : xk + ak * (xk - xout[i * D + a]) + ck * z;
    }
  }
}

// ||rhs - M x||^2 and ||rhs||^2, one CTA, fixed order
template <int D>
__global__ void __launch_bounds__(1024)
k_eq_check(int64_t nu, const int64_t* __restrict__ mptr, const int* __restrict__ mcol, const double* __restrict__ mval,
           const double* __restrict__ rhs, const double* __restrict__ x, double* __restrict__ norms) {
  __shared__ double s[2][32];
  double r2 = 0.0, b2 = 0.0;
  for (int64_t i = threadIdx.x; i < nu; i += blockDim.x) {
    double acc[D];
#pragma unroll
    for (int a = 0; a < D; ++a) acc[a] = 0.0;
    for (int64_t t = mptr[i]; t < mptr[i + 1]; ++t) {
      const double w = mval[t];
#pragma unroll
      for (int a = 0; a < D; ++a) acc[a] = fma(w, x[(size_t)mcol[t] * D + a], acc[a]);
    }
#pragma unroll
    for (int a = 0; a < D; ++a) {
      const double b = rhs[i * D + a], r = b - acc[a];
      r2 = fma(r, r, r2);
      b2 = fma(b, b, b2);
    }
  }
  r2 = warp_sum(r2);
  b2 = warp_sum(b2);
  if ((threadIdx.x & 31) == 0) { s[0][threadIdx.x >> 5] = r2; s[1][threadIdx.x >> 5] = b2; }
  __syncthreads();
  if (threadIdx.x < 32) {
    r2 = warp_sum(s[0][threadIdx.x]);
    b2 = warp_sum(s[1][threadIdx.x]);
    if (threadIdx.x == 0) { norms[0] = r2; norms[1] = b2; }
  }
}

template <int D>
__global__ void __launch_bounds__(MV_BLOCK)
k_eq_scatter(int64_t nb, int64_t nu, const int* __restrict__ node, const double* __restrict__ u, double* __restrict__ g) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nb; i += (int64_t)gridDim.x * blockDim.x) {
    const size_t v = (size_t)node[i];
#pragma unroll
    for (int a = 0; a < D; ++a) g[v * D + a] = i < nu ? u[i * D + a] : 0.0;
  }
}

template <int D>
static int eqbc_apply_t(mv_ma* ma, const double* phi, double* grad_phi) {
  EqBC* e = ma->eqbc;
  mv_mesh* m = ma->mesh;
  const int gu = grid_for(e->nu > 0 ? e->nu : 1), gb = grid_for(e->nb);
  double* x[2] = {e->u[0], e->u[1]};
  if (e->nu > 0) {
    MV_L k_eq_rhs<D><<<gu, MV_BLOCK, 0, g_stream>>>(e->nu, e->fptr, e->flist, m->bf_cell, m->bf_normal, m->bf_area, m->xi, m->cells, m->dg, phi,
                                               e->rhs);
    const double lmin = 0.5, lmax = 1.0 + 0.5 * (D - 1);
    const double theta = 0.5 * (lmax + lmin), delta = 0.5 * (lmax - lmin);
    if (delta == 0.0) {   // d = 1 never gets here (every 1D tag is a plain DirichletBC)
      set_error("equation bc: degenerate interval");
      return MV_ERR_BADARG;
    }
    const double sigma1 = theta / delta;
    double rho = 1.0 / sigma1;
    MV_L k_eq_cheb<D, true><<<gu, MV_BLOCK, 0, g_stream>>>(e->nu, e->mptr, e->mcol, e->mval, e->dinv, e->rhs, nullptr, x[1], 0.0, 1.0 / theta);
    int cur = 1;   // x[cur] = x_k, x[cur ^ 1] = x_{k-1} (x_0 = 0: set below for k = 1)
    MV_CUDA(cudaMemsetAsync(x[0], 0, (size_t)e->nu * D * sizeof(double), g_stream));
    for (int k = 1; k < e->iters; ++k) {
      const double rk = 1.0 / (2.0 * sigma1 - rho);
      MV_L k_eq_cheb<D, false><<<gu, MV_BLOCK, 0, g_stream>>>(e->nu, e->mptr, e->mcol, e->mval, e->dinv, e->rhs, x[cur], x[cur ^ 1], rk * rho,
                                                         2.0 * rk / delta);
      rho = rk;
      cur ^= 1;
    }
    MV_L k_eq_check<D><<<1, 1024, 0, g_stream>>>(e->nu, e->mptr, e->mcol, e->mval, e->rhs, x[cur], e->norms);
    MV_L k_eq_scatter<D><<<gb, MV_BLOCK, 0, g_stream>>>(e->nb, e->nu, e->node, x[cur], grad_phi);
    MV_KERNEL_CHECK();
    MV_CUDA(cudaMemcpyAsync(e->h_norms, e->norms, 2 * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    const double r2 = e->h_norms[0], b2 = e->h_norms[1];
    if (!(r2 <= 1e-24 * b2) && !(r2 <= 1e-300)) {
      set_error("equation bc: boundary projection not converged after %d Chebyshev iterations (relative residual %.3e)", e->iters,
                sqrt(r2 / b2));
      return MV_ERR_NOT_CONVERGED;
    }
  } else {
    MV_L k_eq_scatter<D><<<gb, MV_BLOCK, 0, g_stream>>>(e->nb, 0, e->node, x[0], grad_phi);
    MV_KERNEL_CHECK();
  }
  return MV_OK;
}

// grad_phi at the nodes of the EquationBC segments <- boundary projection of the tangential gradient of phi
int eqbc_apply(mv_ma* ma, const double* phi, double* grad_phi) {
  if (!ma->eqbc || ma->eqbc->nb == 0) return MV_OK;
  if (ma->mesh->dim == 2) return eqbc_apply_t<2>(ma, phi, grad_phi);
  if (ma->mesh->dim == 3) return eqbc_apply_t<3>(ma, phi, grad_phi);
  set_error("equation bc: dimension %d", ma->mesh->dim);
  return MV_ERR_BADARG;
}

int64_t eqbc_counts(const mv_ma* ma, int64_t* nu) {
  if (nu) *nu = ma->eqbc ? ma->eqbc->nu : 0;
  return ma->eqbc ? ma->eqbc->nb : 0;
}

}  // namespace mv
