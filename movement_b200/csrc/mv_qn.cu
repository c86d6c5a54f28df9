// Quasi-Newton pieces: nonlinear residual F, Jacobian action, block
// preconditioner and a flexible GMRES for the Newton correction.
//
// replaces: the PETSc SNES/KSP stack behind movement/monge_ampere.py:640-645 (F), its UFL-derived
//           Jacobian, Jp :670-675 and movement/solver_parameters.py:1-47
//           (gmres + multiplicative fieldsplit: block 0 ~ (M+K)^-1, block 1 ~ M^-1, off-diagonal
//           blocks from the true Jacobian, constant nullspace on phi :682-684).
//
// Unknown u = (phi[V], H[V,d,d]); Krylov vectors hold the owned part, laid out [phi | H].
//   F_psi = -(a - theta m_l)                         a_i = int psi_i m det(I+H)
//   F_tau = M H_ab - S_ab phi                        S phi = the A8 right-hand side (hess_rhs)
//   J (dphi, dH) = [ -A dH ; M dH - S dphi ]         A dH = int psi m cof(I+H):dH   (m, theta frozen)
// The Newton/line-search control flow lives in movement_b200/monge_ampere.py.
#include <math.h>

#include <vector>

#include "mv_internal.cuh"

namespace mv {

#define QN_NB 8  // basis vectors handled per multi-dot / multi-axpy launch

struct VecPtrs {
  const double* v[QN_NB];
};

// dots[k] = w . V_k  for k < nb (deterministic grid reduction)
__global__ void __launch_bounds__(MV_BLOCK)
k_multi_dot(int64_t n, const double* __restrict__ w, VecPtrs vp, int nb, double* __restrict__ partials,
            unsigned* __restrict__ counter, double* __restrict__ out) {
  double acc[QN_NB];
#pragma unroll
  for (int k = 0; k < QN_NB; ++k) acc[k] = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double wi = w[i];
#pragma unroll
    for (int k = 0; k < QN_NB; ++k)
      if (k < nb) acc[k] = fma(wi, __ldg(&vp.v[k][i]), acc[k]);
  }
  grid_reduce_sum<QN_NB>(acc, partials, counter, [&](double(&t)[QN_NB]) {
#pragma unroll
    for (int k = 0; k < QN_NB; ++k) out[k] = t[k];
  });
}

// w += sum_k coef[k] * V_k   (coefficients read from device memory, scaled by `scale`)
__global__ void __launch_bounds__(MV_BLOCK)
k_multi_axpy(int64_t n, double* __restrict__ w, VecPtrs vp, int nb, const double* __restrict__ coef, double scale) {
  double c[QN_NB];
#pragma unroll
  for (int k = 0; k < QN_NB; ++k) c[k] = k < nb ? scale * coef[k] : 0.0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    double wi = w[i];
#pragma unroll
    for (int k = 0; k < QN_NB; ++k)
      if (k < nb) wi = fma(c[k], __ldg(&vp.v[k][i]), wi);
    w[i] = wi;
  }
}

__global__ void __launch_bounds__(MV_BLOCK)
k_scale_copy(int64_t n, double s, const double* __restrict__ a, double* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = s * a[i];
}
__global__ void __launch_bounds__(MV_BLOCK)
k_sub(int64_t n, const double* __restrict__ a, const double* __restrict__ b, double* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = a[i] - b[i];
}
__global__ void __launch_bounds__(MV_BLOCK)
k_add_inplace(int64_t n, const double* __restrict__ a, double* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] += a[i];
}
// out = base - lambda * y
__global__ void __launch_bounds__(MV_BLOCK)
k_axpy_from(int64_t n, const double* __restrict__ base, double lambda, const double* __restrict__ y, double* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = fma(-lambda, y[i], base[i]);
}
__global__ void __launch_bounds__(MV_BLOCK)
k_sum1(int64_t n, const double* __restrict__ a, double* __restrict__ partials, unsigned* __restrict__ counter, double* __restrict__ out) {
  double acc[1] = {0.0};
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) acc[0] += a[i];
  grid_reduce_sum<1>(acc, partials, counter, [&](double(&t)[1]) { out[0] = t[0]; });
}
__global__ void __launch_bounds__(MV_BLOCK)
k_shift(int64_t n, double* __restrict__ a, const double* __restrict__ sum, double inv_n) {
  const double s = sum[0] * inv_n;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) a[i] -= s;
}

struct Qn {
  mv_ma* ma;
  mv_mesh* m;
  int d, dd;
  int64_t nv, N;       // owned vertices, owned unknowns nv*(1+dd)
  double nglobal;      // global vertex count (mean projection)
};

static int qn_setup(mv_ma* ma, Qn& q) {
  q.ma = ma;
  q.m = ma->mesh;
  q.d = q.m->dim;
  q.dd = q.d * q.d;
  q.nv = q.m->Vown;
  q.N = q.nv * (1 + q.dd);
  mv_mesh* m = q.m;
  for (int fld : {MV_F_PHI, MV_F_H, MV_F_MONITOR, MV_F_RESIDUAL}) {
    if (!ma->f[fld]) {
      set_error("mv_qn: field %d not bound", fld);
      return MV_ERR_BADARG;
    }
  }
  if (!ma->qn_w) {
    MV_CHECK(dev_alloc(&ma->qn_w, (size_t)q.N));
    MV_CHECK(dev_alloc(&ma->qn_gphi, (size_t)m->V));
    MV_CHECK(dev_alloc(&ma->qn_gH, (size_t)m->V * q.dd));
    MV_CHECK(dev_alloc(&ma->qn_t1, (size_t)m->V * q.dd));
    MV_CHECK(dev_alloc(&ma->qn_t2, (size_t)m->V * q.dd));
    MV_CHECK(dev_alloc(&ma->qn_base, (size_t)q.N));
    MV_CHECK(dev_alloc(&ma->qn_dots, 64));
    MV_CUDA(cudaMallocHost((void**)&ma->qn_hdots, 64 * sizeof(double)));
    MV_CUDA(cudaMemsetAsync(ma->qn_gphi, 0, m->V * sizeof(double), g_stream));
    MV_CUDA(cudaMemsetAsync(ma->qn_gH, 0, m->V * q.dd * sizeof(double), g_stream));
  }
  double hn = (double)q.nv;
  if (comm_active()) {
    MV_CUDA(cudaMemcpyAsync(ma->qn_dots, &hn, sizeof(double), cudaMemcpyHostToDevice, g_stream));
    MV_CHECK(allreduce_sum_dev(ma->qn_dots, 1));
    MV_CUDA(cudaMemcpyAsync(&hn, ma->qn_dots, sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
  }
  q.nglobal = hn;
  return MV_OK;
}

// remove the mean of the phi part (constant nullspace / transpose nullspace, monge_ampere.py:682-684)
static int project_phi(Qn& q, double* vec) {
  mv_mesh* m = q.m;
  double* s = q.ma->qn_dots + 48;
  MV_L k_sum1<<<grid_for(q.nv), MV_BLOCK, 0, g_stream>>>(q.nv, vec, m->red.partials, m->red.counter, s);
  if (comm_active()) MV_CHECK(allreduce_sum_dev(s, 1));
  MV_L k_shift<<<grid_for(q.nv), MV_BLOCK, 0, g_stream>>>(q.nv, vec, s, 1.0 / q.nglobal);
  MV_KERNEL_CHECK();
  return MV_OK;
}

// dots of w against nb basis vectors -> host (global over ranks)
static int dots_host(Qn& q, const double* w, const double* const* vecs, int nb, double* out) {
  mv_mesh* m = q.m;
  double* dd = q.ma->qn_dots;
  for (int k0 = 0; k0 < nb; k0 += QN_NB) {
    VecPtrs vp;
    const int n = nb - k0 < QN_NB ? nb - k0 : QN_NB;
    for (int k = 0; k < QN_NB; ++k) vp.v[k] = vecs[k0 + (k < n ? k : 0)];
    MV_L k_multi_dot<<<grid_for(q.N), MV_BLOCK, 0, g_stream>>>(q.N, w, vp, n, m->red.partials, m->red.counter, dd + k0);
  }
  MV_KERNEL_CHECK();
  const int padded = ((nb + QN_NB - 1) / QN_NB) * QN_NB;
  if (comm_active()) MV_CHECK(allreduce_sum_dev(dd, padded));
  MV_CUDA(cudaMemcpyAsync(q.ma->qn_hdots, dd, padded * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  for (int k = 0; k < nb; ++k) out[k] = q.ma->qn_hdots[k];
  return MV_OK;
}

// w += scale * sum_k coef[k] vecs[k], coefficients given on the host
static int axpys(Qn& q, double* w, const double* const* vecs, int nb, const double* coef, double scale) {
  double* dd = q.ma->qn_dots;
  MV_CUDA(cudaMemcpyAsync(dd, coef, nb * sizeof(double), cudaMemcpyHostToDevice, g_stream));
  for (int k0 = 0; k0 < nb; k0 += QN_NB) {
    VecPtrs vp;
    const int n = nb - k0 < QN_NB ? nb - k0 : QN_NB;
    for (int k = 0; k < QN_NB; ++k) vp.v[k] = vecs[k0 + (k < n ? k : 0)];
    MV_L k_multi_axpy<<<grid_for(q.N), MV_BLOCK, 0, g_stream>>>(q.N, w, vp, n, dd + k0, scale);
  }
  MV_KERNEL_CHECK();
  MV_CUDA(cudaStreamSynchronize(g_stream));  // coef is a host buffer the caller may reuse
  return MV_OK;
}

// ghosted copies of an owned [phi | H] vector (halo-exchanged when distributed)
static int scatter_ghosted(Qn& q, const double* vec, double* gphi, double* gH) {
  mv_mesh* m = q.m;
  MV_CUDA(cudaMemcpyAsync(gphi, vec, q.nv * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
  MV_CUDA(cudaMemcpyAsync(gH, vec + q.nv, q.nv * q.dd * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
  if (comm_active()) {
    MV_CHECK(halo_exchange(m, 1, gphi));
    MV_CHECK(halo_exchange(m, q.dd, gH));
  }
  return MV_OK;
}

// out = J z :  out_phi = -A(m,H) zH ;  out_H = M zH - S zphi      (Q2)
static int jac_apply(Qn& q, const double* z, double* out) {
  mv_ma* ma = q.ma;
  mv_mesh* m = q.m;
  MV_CHECK(scatter_ghosted(q, z, ma->qn_gphi, ma->qn_gH));
  MV_CHECK(jac_A_dev(m, ma->f[MV_F_MONITOR], ma->f[MV_F_H], ma->qn_gH, -1.0, out));
  MV_CHECK(spmv(m, MV_OP_MASS, q.dd, ma->qn_gH, out + q.nv));
  MV_CHECK(hess_rhs_dev(m, ma->qn_gphi, ma->qn_t1));
  const int64_t n = q.nv * q.dd;
  MV_L k_sub<<<grid_for(n), MV_BLOCK, 0, g_stream>>>(n, out + q.nv, ma->qn_t1, out + q.nv);
  MV_KERNEL_CHECK();
  return project_phi(q, out);
}

// z = P^-1 r, multiplicative fieldsplit with the true off-diagonal block (Q3):
//   z_phi = (M+K)^-1 r_phi (inexact PCG) ; z_H = M^-1 (r_H + S z_phi) (inexact PCG)
static int pc_apply(Qn& q, const double* r, double* z, double pc_rtol, int pc_maxit) {
  mv_ma* ma = q.ma;
  mv_mesh* m = q.m;
  double* yphi = ma->qn_gphi;   // ghosted work (solution vectors of the inner solves)
  double* yH = ma->qn_gH;
  MV_CUDA(cudaMemsetAsync(yphi, 0, m->V * sizeof(double), g_stream));
  int st = pcg_solve(m, MV_OP_MASS_PLUS_STIFFNESS, 1, r, yphi, nullptr, pc_rtol, 0.0, pc_maxit, 0, nullptr, nullptr);
  if (st != MV_OK && st != MV_ERR_NOT_CONVERGED) return st;
  if (comm_active()) MV_CHECK(halo_exchange(m, 1, yphi));
  MV_CHECK(hess_rhs_dev(m, yphi, ma->qn_t1));
  const int64_t n = q.nv * q.dd;
  MV_L k_add_inplace<<<grid_for(n), MV_BLOCK, 0, g_stream>>>(n, r + q.nv, ma->qn_t1);
  MV_KERNEL_CHECK();
  MV_CUDA(cudaMemsetAsync(yH, 0, m->V * q.dd * sizeof(double), g_stream));
  st = pcg_solve(m, MV_OP_MASS, q.dd, ma->qn_t1, yH, nullptr, pc_rtol, 0.0, pc_maxit, 0, nullptr, nullptr);
  if (st != MV_OK && st != MV_ERR_NOT_CONVERGED) return st;
  MV_CUDA(cudaMemcpyAsync(z, yphi, q.nv * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
  MV_CUDA(cudaMemcpyAsync(z + q.nv, yH, n * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
  return project_phi(q, z);
}

}  // namespace mv

using namespace mv;

// F at the current iterate (phi, H) with the monitor and H_old already updated by the caller
// (update_monitor, monge_ampere.py:648-662).  out[0] = ||F||_2^2, out[1] = theta,
// out[2] = relative l2 residual (monge_ampere.py:212-223).
extern "C" int mv_qn_residual(mv_ma* ma, double* F, double* out) {
  if (!ma || !F || !out) return MV_ERR_BADARG;
  Qn q;
  MV_CHECK(qn_setup(ma, q));
  mv_mesh* m = q.m;
  const bool dist = comm_active();
  double gtot = m->total_volume, gml2 = m->ml_sumsq;
  if (dist) {
    double h[2] = {gtot, gml2};
    MV_CUDA(cudaMemcpyAsync(ma->qn_dots + 32, h, sizeof(h), cudaMemcpyHostToDevice, g_stream));
    MV_CHECK(allreduce_sum_dev(ma->qn_dots + 32, 2));
    MV_CUDA(cudaMemcpyAsync(h, ma->qn_dots + 32, sizeof(h), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    gtot = h[0]; gml2 = h[1];
    MV_CHECK(halo_exchange(m, q.dd, ma->f[MV_F_H]));
    MV_CHECK(halo_exchange(m, 1, ma->f[MV_F_PHI]));
  }
  // F_psi = -(a - theta m_l)
  MV_CHECK(assemble_a_dev(m, ma->f[MV_F_MONITOR], ma->f[MV_F_H], ma->a, ma->dscal));
  if (dist) MV_CHECK(allreduce_sum_dev(ma->dscal + D_SUMA, 1));
  double suma = 0.0;
  MV_CUDA(cudaMemcpyAsync(&suma, ma->dscal + D_SUMA, sizeof(double), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  const double theta = suma / gtot;
  // residual vector r = a - theta*ml into F_psi with a minus sign: reuse k_axpy_from (out = base - lambda*y)
  MV_L k_axpy_from<<<grid_for(q.nv), MV_BLOCK, 0, g_stream>>>(q.nv, ma->a, theta, m->ml, ma->f[MV_F_RESIDUAL]);
  MV_L k_scale_copy<<<grid_for(q.nv), MV_BLOCK, 0, g_stream>>>(q.nv, -1.0, ma->f[MV_F_RESIDUAL], F);
  // F_tau = M H - S phi
  MV_CHECK(spmv(m, MV_OP_MASS, q.dd, ma->f[MV_F_H], F + q.nv));
  MV_CHECK(hess_rhs_dev(m, ma->f[MV_F_PHI], ma->qn_t1));
  const int64_t n = q.nv * q.dd;
  MV_L k_sub<<<grid_for(n), MV_BLOCK, 0, g_stream>>>(n, F + q.nv, ma->qn_t1, F + q.nv);
  MV_KERNEL_CHECK();
  // norms: ||F||^2 and ||F_psi||^2 (= ||r||^2)
  const double* vecs[1] = {F};
  double f2 = 0.0;
  MV_CHECK(dots_host(q, F, vecs, 1, &f2));
  // ||r||^2 over the phi part only: dot of F against itself restricted -> reuse the multi-dot on nv entries
  {
    VecPtrs vp;
    for (int k = 0; k < QN_NB; ++k) vp.v[k] = F;
    MV_L k_multi_dot<<<grid_for(q.nv), MV_BLOCK, 0, g_stream>>>(q.nv, F, vp, 1, m->red.partials, m->red.counter, ma->qn_dots);
    MV_KERNEL_CHECK();
    if (dist) MV_CHECK(allreduce_sum_dev(ma->qn_dots, 1));
    MV_CUDA(cudaMemcpyAsync(ma->qn_hdots, ma->qn_dots, sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
  }
  out[0] = f2;
  out[1] = theta;
  out[2] = sqrt(ma->qn_hdots[0]) / (fabs(theta) * sqrt(gml2));
  return MV_OK;
}

// remember the current iterate u0 = (phi, H); set (phi, H) = u0 - lambda * Y (line-search trial points)
extern "C" int mv_qn_save(mv_ma* ma) {
  if (!ma) return MV_ERR_BADARG;
  Qn q;
  MV_CHECK(qn_setup(ma, q));
  MV_CUDA(cudaMemcpyAsync(ma->qn_base, ma->f[MV_F_PHI], q.nv * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
  MV_CUDA(cudaMemcpyAsync(ma->qn_base + q.nv, ma->f[MV_F_H], q.nv * q.dd * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
  return MV_OK;
}
extern "C" int mv_qn_set_iterate(mv_ma* ma, double lambda, const double* Y) {
  if (!ma || !Y) return MV_ERR_BADARG;
  Qn q;
  MV_CHECK(qn_setup(ma, q));
  const int64_t n = q.nv * q.dd;
  MV_L k_axpy_from<<<grid_for(q.nv), MV_BLOCK, 0, g_stream>>>(q.nv, ma->qn_base, lambda, Y, ma->f[MV_F_PHI]);
  MV_L k_axpy_from<<<grid_for(n), MV_BLOCK, 0, g_stream>>>(n, ma->qn_base + q.nv, lambda, Y + q.nv, ma->f[MV_F_H]);
  MV_KERNEL_CHECK();
  if (comm_active()) {
    MV_CHECK(halo_exchange(q.m, 1, ma->f[MV_F_PHI]));
    MV_CHECK(halo_exchange(q.m, q.dd, ma->f[MV_F_H]));
  }
  return MV_OK;
}

// y = J z at the current (monitor, H) -- exposed for the parity tests
extern "C" int mv_qn_jacobian_apply(mv_ma* ma, const double* z, double* y) {
  if (!ma || !z || !y) return MV_ERR_BADARG;
  Qn q;
  MV_CHECK(qn_setup(ma, q));
  return jac_apply(q, z, y);
}

// Solve J Y = F by right-preconditioned flexible GMRES(restart) from Y = 0.
// Stops on ||F - J Y|| <= ksp_rtol ||F|| (PETSc default 1e-5) or ksp_maxit (solver_parameters.py:11-13).
// info[0] = iterations, info[1] = relative residual estimate.
extern "C" int mv_qn_solve(mv_ma* ma, const double* F, double* Y, double ksp_rtol, int ksp_maxit, int restart,
                           double pc_rtol, int pc_maxit, double* info) {
  if (!ma || !F || !Y) return MV_ERR_BADARG;
  Qn q;
  MV_CHECK(qn_setup(ma, q));
  // ksp_gmres_restart (movement/solver_parameters.py:13 sets 200 = ksp_max_it, i.e. no restart in the reference).  The
  // Krylov basis of flexible GMRES costs 2 * restart + 1 vectors of V (1 + d^2) doubles, so the value is honoured up to what
  // fits in 60 % of the free device memory (never below 5); info[2] reports the value used.  A solve that needs fewer
  // iterations than that -- the normal case, 10-30 with this preconditioner -- is identical to the unrestarted one.
  if (restart < 1) restart = 200;
  if (restart > ksp_maxit && ksp_maxit >= 1) restart = ksp_maxit;
  {
    size_t free_b = 0, total_b = 0;
    MV_CUDA(cudaMemGetInfo(&free_b, &total_b));
    if (ma->qn_basis) free_b += (size_t)q.N * (2 * ma->qn_restart + 1) * sizeof(double);
    const double cap = (0.6 * (double)free_b / ((double)q.N * sizeof(double)) - 1.0) / 2.0;
    const int icap = cap < 5.0 ? 5 : (cap > 1e6 ? 1000000 : (int)cap);
    if (restart > icap) restart = icap;
  }
  if (ma->qn_restart != restart) {
    dev_free(ma->qn_basis);
    MV_CHECK(dev_alloc(&ma->qn_basis, (size_t)q.N * (2 * restart + 1)));
    ma->qn_restart = restart;
  }
  const int m = restart;
  double* Vb = ma->qn_basis;                      // V_0 .. V_m
  double* Zb = ma->qn_basis + (size_t)q.N * (m + 1);  // Z_0 .. Z_{m-1}
  double* w = ma->qn_w;
  const int gN = grid_for(q.N);
  MV_CUDA(cudaMemsetAsync(Y, 0, q.N * sizeof(double), g_stream));
  std::vector<double> Hh((size_t)(m + 1) * m, 0.0), cs(m), sn(m), g(m + 1), y(m), h(m + 2), h2(m + 2);
  std::vector<const double*> vptr(m + 1), zptr(m);
  for (int k = 0; k <= m; ++k) vptr[k] = Vb + (size_t)q.N * k;
  for (int k = 0; k < m; ++k) zptr[k] = Zb + (size_t)q.N * k;
  double bnorm = -1.0, rel = 1.0;
  int total = 0;
  while (true) {
    // r = F - J Y   (Y = 0 on the first cycle)
    if (total == 0) {
      MV_CUDA(cudaMemcpyAsync(w, F, q.N * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
    } else {
      MV_CHECK(jac_apply(q, Y, w));
      MV_L k_sub<<<gN, MV_BLOCK, 0, g_stream>>>(q.N, F, w, w);
      MV_KERNEL_CHECK();
    }
    MV_CHECK(project_phi(q, w));
    const double* wp[1] = {w};
    double r2 = 0.0;
    MV_CHECK(dots_host(q, w, wp, 1, &r2));
    const double beta = sqrt(r2);
    if (bnorm < 0.0) bnorm = beta;
    rel = bnorm > 0.0 ? beta / bnorm : 0.0;
    if (!(beta > 0.0) || rel <= ksp_rtol || total >= ksp_maxit) break;
    MV_L k_scale_copy<<<gN, MV_BLOCK, 0, g_stream>>>(q.N, 1.0 / beta, w, Vb);
    MV_KERNEL_CHECK();
    std::fill(g.begin(), g.end(), 0.0);
    g[0] = beta;
    int j = 0;
    for (; j < m && total < ksp_maxit; ++j, ++total) {
      double* Zj = Zb + (size_t)q.N * j;
      MV_CHECK(pc_apply(q, vptr[j], Zj, pc_rtol, pc_maxit));
      MV_CHECK(jac_apply(q, Zj, w));
      // classical Gram-Schmidt, twice
      MV_CHECK(dots_host(q, w, vptr.data(), j + 1, h.data()));
      MV_CHECK(axpys(q, w, vptr.data(), j + 1, h.data(), -1.0));
      MV_CHECK(dots_host(q, w, vptr.data(), j + 1, h2.data()));
      MV_CHECK(axpys(q, w, vptr.data(), j + 1, h2.data(), -1.0));
      for (int i = 0; i <= j; ++i) h[i] += h2[i];
      double n2 = 0.0;
      MV_CHECK(dots_host(q, w, wp, 1, &n2));
      const double hn = sqrt(n2);
      h[j + 1] = hn;
      if (hn > 0.0) {
        MV_L k_scale_copy<<<gN, MV_BLOCK, 0, g_stream>>>(q.N, 1.0 / hn, w, Vb + (size_t)q.N * (j + 1));
        MV_KERNEL_CHECK();
      }
      // Givens rotations on the new Hessenberg column
      for (int i = 0; i < j; ++i) {
        const double t = cs[i] * h[i] + sn[i] * h[i + 1];
        h[i + 1] = -sn[i] * h[i] + cs[i] * h[i + 1];
        h[i] = t;
      }
      const double den = hypot(h[j], h[j + 1]);
      cs[j] = den > 0.0 ? h[j] / den : 1.0;
      sn[j] = den > 0.0 ? h[j + 1] / den : 0.0;
      h[j] = cs[j] * h[j] + sn[j] * h[j + 1];
      h[j + 1] = 0.0;
      g[j + 1] = -sn[j] * g[j];
      g[j] = cs[j] * g[j];
      for (int i = 0; i <= j; ++i) Hh[(size_t)i * m + j] = h[i];
      rel = fabs(g[j + 1]) / bnorm;
      if (rel <= ksp_rtol || hn == 0.0) { ++j; ++total; break; }
    }
    // back substitution, Y += Z y
    for (int i = j - 1; i >= 0; --i) {
      double s = g[i];
      for (int k = i + 1; k < j; ++k) s -= Hh[(size_t)i * m + k] * y[k];
      y[i] = s / Hh[(size_t)i * m + i];
    }
    MV_CHECK(axpys(q, Y, zptr.data(), j, y.data(), 1.0));
    if (rel <= ksp_rtol || total >= ksp_maxit) {
      if (rel > ksp_rtol) continue;  // one more true-residual evaluation decides
      break;
    }
  }
  if (info) { info[0] = (double)total; info[1] = rel; info[2] = (double)restart; }
  if (rel > ksp_rtol) {
    // PETSc: KSP_DIVERGED_ITS -> SNES_DIVERGED_LINEAR_SOLVE -> firedrake ConvergenceError (monge_ampere.py:744-745)
    set_error("gmres: not converged in %d iterations (relative residual %.3e > %.3e)", total, rel, ksp_rtol);
    return MV_ERR_NOT_CONVERGED;
  }
  return MV_OK;
}
