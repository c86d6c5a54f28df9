// Smoothed-aggregation multigrid: the V-cycle (solve phase), used as the PCG preconditioner.
//
// replaces: the PCApply of PETSc `gamg` (movement/solver_parameters.py:68-71) inside the KSP at
//           movement/monge_ampere.py:441-448, and of the ILU/gamg sub-solves at
//           movement/laplacian_smoothing.py:75 and solver_parameters.py:26-33.
//
// One cycle = per level: Chebyshev(k) pre-smoothing from a zero guess, residual, restriction;
// dense solve on the coarsest level; prolongation + Chebyshev(k) post-smoothing on the way up.
// Every kernel is a streaming SELL-32 pass (one warp per 32-row slice, matrix read once with
// evict-first loads, vectors gathered through the read-only path) with the vector updates of the
// smoother fused into the SpMV epilogue, so a Chebyshev step costs one matrix stream:
//   k_smooth_first   r = b - A x ; d = D^-1 r / theta ; x' = x + d           12 nnz + 4 + 40c B/row
//   k_smooth_next    r -= A d ; d' = c1 d + c2 D^-1 r ; x += d'               12 nnz + 4 + 56c B/row
//   k_residual       r = b - A x                                              12 nnz + 4 + 24c B/row
//   k_sell_apply     y (+)= S x  for the transfer operators P and R = P^T
// The smoother is a fixed polynomial in D^-1 A and pre/post smoothing are equal, so the cycle is a
// symmetric positive definite operator, as PCG needs.  Kernels turn into no-ops once the solve's
// `done` flag is set (the host polls it once per batch of iterations).
#include "mv_internal.cuh"

namespace mv {

template <int C, typename TV = double>
__device__ __forceinline__ void sell_dot(const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
                                         const TV* __restrict__ val, const double* __restrict__ x, int64_t ldx, int64_t s,
                                         int lane, double (&acc)[C]) {
  const int64_t base = __ldg(&slice_ptr[s]);
  const int w = (int)((__ldg(&slice_ptr[s + 1]) - base) >> 5);
  const int* cp = cols + base + lane;
  const TV* vp = val + base + lane;
#pragma unroll
  for (int c = 0; c < C; ++c) acc[c] = 0.0;
  // column indices of a chunk first, then the gathers (breaks the index -> gather latency chain)
  constexpr int KB = 8;
  for (int k0 = 0; k0 < w; k0 += KB) {
    int col[KB];
    double a[KB];
#pragma unroll
    for (int j = 0; j < KB; ++j) {
      const bool in = k0 + j < w;
      col[j] = in ? __ldcs(cp + (k0 + j) * MV_SLICE) : 0;
      a[j] = in ? (double)__ldcs(vp + (k0 + j) * MV_SLICE) : 0.0;
    }
#pragma unroll
    for (int j = 0; j < KB; ++j) {
      const double* xp = x + col[j];
#pragma unroll
      for (int c = 0; c < C; ++c) acc[c] = fma(a[j], __ldg(xp + c * ldx), acc[c]);
    }
  }
}

#define SLICE_LOOP                                                                         \
  const int lane = threadIdx.x & 31;                                                       \
  const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;             \
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;                           \
  for (int64_t s = s0 + warp0; s < nslices; s += nwarps)

// x = d = D^-1 b / theta (zero initial guess)
template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_smooth_zero(int64_t n, int64_t ld, const double* __restrict__ b, const double* __restrict__ dinv, double sc, double* __restrict__ x,
              double* __restrict__ d, const int* __restrict__ done) {
  if (done && *done) return;
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < n; v += (int64_t)gridDim.x * blockDim.x) {
    const double di = sc * __ldg(&dinv[v]);
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const double t = di * b[c * ld + v];
      x[c * ld + v] = t;
      if (d) d[c * ld + v] = t;
    }
  }
}

template <int C, typename TV>
__global__ void __launch_bounds__(MV_BLOCK)
k_smooth_first(int64_t s0, int64_t nslices, int64_t nrows, int64_t ld, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
               const TV* __restrict__ val, const double* __restrict__ b, const double* __restrict__ xin,
               const double* __restrict__ dinv, double sc, double* __restrict__ xout, double* __restrict__ d,
               double* __restrict__ r, const int* __restrict__ done) {
  if (done && *done) return;
  SLICE_LOOP {
    double acc[C];
    sell_dot<C, TV>(slice_ptr, cols, val, xin, ld, s, lane, acc);
    const int64_t row = s * MV_SLICE + lane;
    if (row < nrows) {
      const double di = __ldg(&dinv[row]);
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const int64_t i = c * ld + row;
        const double t = di != 0.0 ? b[i] - acc[c] : 0.0;
        const double dd = sc * di * t;
        xout[i] = xin[i] + dd;
        if (d) d[i] = dd;
        if (r) r[i] = t;
      }
    }
  }
}

template <int C, typename TV>
__global__ void __launch_bounds__(MV_BLOCK)
k_smooth_next(int64_t s0, int64_t nslices, int64_t nrows, int64_t ld, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
              const TV* __restrict__ val, const double* __restrict__ din, double* __restrict__ dout,
              const double* __restrict__ rin, double* __restrict__ rout, const double* __restrict__ dinv, double c1, double c2,
              double* __restrict__ x, const int* __restrict__ done) {
  if (done && *done) return;
  SLICE_LOOP {
    double acc[C];
    sell_dot<C, TV>(slice_ptr, cols, val, din, ld, s, lane, acc);
    const int64_t row = s * MV_SLICE + lane;
    if (row < nrows) {
      const double di = __ldg(&dinv[row]);
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const int64_t i = c * ld + row;
        const double t = di != 0.0 ? rin[i] - acc[c] : 0.0;
        const double dd = c1 * din[i] + c2 * di * t;
        dout[i] = dd;
        x[i] += dd;
        if (rout) rout[i] = t;
      }
    }
  }
}

template <int C, typename TV>
__global__ void __launch_bounds__(MV_BLOCK)
k_residual(int64_t s0, int64_t nslices, int64_t nrows, int64_t ld, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
           const TV* __restrict__ val, const double* __restrict__ b, const double* __restrict__ x,
           const double* __restrict__ dinv, double* __restrict__ r, const int* __restrict__ done) {
  if (done && *done) return;
  SLICE_LOOP {
    double acc[C];
    sell_dot<C, TV>(slice_ptr, cols, val, x, ld, s, lane, acc);
    const int64_t row = s * MV_SLICE + lane;
    if (row < nrows) {
      const bool free_row = __ldg(&dinv[row]) != 0.0;
#pragma unroll
      for (int c = 0; c < C; ++c) r[c * ld + row] = free_row ? b[c * ld + row] - acc[c] : 0.0;
    }
  }
}

// y = S x (ADD = false) or y += S x (ADD = true) for the rectangular transfer operators
template <int C, bool ADD>
__global__ void __launch_bounds__(MV_BLOCK)
k_sell_apply(int64_t nslices, int64_t nrows, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
             const double* __restrict__ val, const double* __restrict__ x, int64_t ldx, double* __restrict__ y, int64_t ldy,
             const int* __restrict__ done) {
  if (done && *done) return;
  const int64_t s0 = 0;
  SLICE_LOOP {
    double acc[C];
    sell_dot<C>(slice_ptr, cols, val, x, ldx, s, lane, acc);
    const int64_t row = s * MV_SLICE + lane;
    if (row < nrows) {
#pragma unroll
      for (int c = 0; c < C; ++c) y[c * ldy + row] = ADD ? y[c * ldy + row] + acc[c] : acc[c];
    }
  }
}

// y = S x with S in CSR, one warp per row (long rows: restriction on the coarse levels); the lanes' partial
// sums are combined by a fixed-order shuffle tree, so the result is deterministic
template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_csr_apply_warp(int64_t nrows, const int64_t* __restrict__ ptr, const int* __restrict__ col, const double* __restrict__ val,
                 const double* __restrict__ x, int64_t ldx, double* __restrict__ y, int64_t ldy, const int* __restrict__ done) {
  if (done && *done) return;
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t row = warp0; row < nrows; row += nwarps) {
    double acc[C];
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = 0.0;
    const int64_t b = ptr[row], e = ptr[row + 1];
    for (int64_t t = b + lane; t < e; t += 32) {
      const double a = val[t];
      const double* xp = x + col[t];
#pragma unroll
      for (int c = 0; c < C; ++c) acc[c] = fma(a, __ldg(xp + c * ldx), acc[c]);
    }
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const double t = warp_sum(acc[c]);
      if (lane == 0) y[c * ldy + row] = t;
    }
  }
}

// The smoother passes of a small coarse level in CSR, one warp per row (rows of ~80 non-zeros: the SELL kernels above walk
// them serially, one thread per row).  Same arithmetic per row up to the order of the row sum (fixed shuffle tree).
template <int C>
__device__ __forceinline__ void csr_row_dot(const int64_t* __restrict__ ptr, const int* __restrict__ col, const double* __restrict__ val,
                                            const double* __restrict__ x, int64_t ld, int64_t row, int lane, double (&acc)[C]) {
#pragma unroll
  for (int c = 0; c < C; ++c) acc[c] = 0.0;
  const int64_t b = ptr[row], e = ptr[row + 1];
  for (int64_t t = b + lane; t < e; t += 32) {
    const double a = val[t];
    const double* xp = x + col[t];
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = fma(a, __ldg(xp + c * ld), acc[c]);
  }
#pragma unroll
  for (int c = 0; c < C; ++c) acc[c] = warp_sum(acc[c]);
}
#define ROW_LOOP                                                                           \
  const int lane = threadIdx.x & 31;                                                       \
  const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;             \
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;                           \
  for (int64_t row = warp0; row < nrows; row += nwarps)

// MODE 0: k_smooth_first, 1: k_smooth_next, 2: k_residual (arguments as there; unused ones nullptr)
template <int C, int MODE>
__global__ void __launch_bounds__(MV_BLOCK)
k_smooth_csr(int64_t nrows, int64_t ld, const int64_t* __restrict__ ptr, const int* __restrict__ col, const double* __restrict__ val,
             const double* __restrict__ b, const double* __restrict__ xin, const double* __restrict__ dinv, double c1, double c2,
             double* __restrict__ xout, double* __restrict__ dout, double* __restrict__ rout, const int* __restrict__ done) {
  if (done && *done) return;
  ROW_LOOP {
    double acc[C];
    csr_row_dot<C>(ptr, col, val, xin, ld, row, lane, acc);
    if (lane < C) {
      double a = acc[0];
#pragma unroll
      for (int c = 1; c < C; ++c) a = lane == c ? acc[c] : a;
      const int64_t i = lane * ld + row;
      const double di = dinv[row];
      const double t = di != 0.0 ? b[i] - a : 0.0;
      if (MODE == 0) {            // b = right-hand side, xin = iterate: x' = x + sc D^-1 r
        const double dd = c2 * di * t;
        xout[i] = xin[i] + dd;
        if (dout) dout[i] = dd;
        if (rout) rout[i] = t;
      } else if (MODE == 1) {     // b = r_in, xin = d_in: d' = c1 d + c2 D^-1 (r - A d), x += d'
        const double dd = c1 * xin[i] + c2 * di * t;
        dout[i] = dd;
        xout[i] += dd;
        if (rout) rout[i] = t;
      } else {
        rout[i] = t;
      }
    }
  }
}

// x = Ainv b on the coarsest level: one warp per owned row (global row goff + row), lanes stride the N
// columns, fixed-order shuffle sum.  Multi-GPU: b is the gathered right-hand side of all ranks.
template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_dense_apply(int n, int N, int64_t goff, const double* __restrict__ inv, const double* __restrict__ b, int64_t ldb,
              double* __restrict__ x, int64_t ldx, const int* __restrict__ done) {
  if (done && *done) return;
  const int lane = threadIdx.x & 31;
  const int row = (int)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5);
  if (row >= n) return;
  double acc[C];
#pragma unroll
  for (int c = 0; c < C; ++c) acc[c] = 0.0;
  for (int j = lane; j < N; j += 32) {
    const double a = inv[(size_t)(goff + row) * N + j];
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = fma(a, b[c * ldb + j], acc[c]);
  }
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const double t = warp_sum(acc[c]);
    if (lane == 0) x[c * ldx + row] = t;
  }
}
// own segment of the coarse right-hand side into the zeroed global vector [C][N]
template <int C>
__global__ void k_coarse_scatter(int n, int N, int64_t goff, const double* __restrict__ b, int64_t ld, double* __restrict__ bg) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
#pragma unroll
    for (int c = 0; c < C; ++c) bg[(size_t)c * N + goff + i] = b[c * ld + i];
}

int amg_ensure_work(Amg* a, int C) {
  for (size_t l = 0; l < a->lv.size(); ++l) {
    AmgLevel& L = a->lv[l];
    if (L.work_C >= C) continue;
    if (l > 0) { dev_free(L.x0); dev_free(L.b); }
    else dev_free(L.x0);
    dev_free(L.x1); dev_free(L.r); dev_free(L.d0); dev_free(L.d1);
    const size_t nt = (size_t)L.ntot * C;
    MV_CHECK(dev_alloc(&L.x0, nt));
    MV_CUDA(cudaMemsetAsync(L.x0, 0, nt * sizeof(double), g_stream));
    if (l > 0) {
      MV_CHECK(dev_alloc(&L.b, nt));
      MV_CUDA(cudaMemsetAsync(L.b, 0, nt * sizeof(double), g_stream));
      MV_CHECK(dev_alloc(&L.x1, nt));
      MV_CUDA(cudaMemsetAsync(L.x1, 0, nt * sizeof(double), g_stream));
    }
    MV_CHECK(dev_alloc(&L.r, nt));
    MV_CUDA(cudaMemsetAsync(L.r, 0, nt * sizeof(double), g_stream));
    if (L.degree > 1) {
      MV_CHECK(dev_alloc(&L.d0, nt));
      MV_CHECK(dev_alloc(&L.d1, nt));
      MV_CUDA(cudaMemsetAsync(L.d0, 0, nt * sizeof(double), g_stream));
      MV_CUDA(cudaMemsetAsync(L.d1, 0, nt * sizeof(double), g_stream));
    }
    L.work_C = C;
  }
  if (a->lv.back().dist && a->coarse_inv && a->z_C < C) {
    dev_free(a->cglob);
    MV_CHECK(dev_alloc(&a->cglob, (size_t)a->nc * C));
  }
  if (a->z_C < C) {
    dev_free(a->z);
    MV_CHECK(dev_alloc(&a->z, (size_t)a->lv[0].ntot * C));
    MV_CUDA(cudaMemsetAsync(a->z, 0, (size_t)a->lv[0].ntot * C * sizeof(double), g_stream));
    a->z_C = C;
  }
  return MV_OK;
}

static inline int slice_grid(const Sell& p) { return grid_for(p.nslices * MV_SLICE, 8); }

// Chebyshev(L.degree) on D^-1 A over [rho/ratio, rho].  xin == nullptr: zero initial guess.  Result in xout
// (xout != xin).
template <int C>
static int smooth(AmgLevel& L, const double* b, const double* xin, double* xout, const int* done) {
  const int deg = L.degree;
  const double ratio = deg <= 1 ? 4.0 : 8.0;
  const double lmax = L.rho, lmin = L.rho / ratio;
  const double theta = 0.5 * (lmax + lmin), delta = 0.5 * (lmax - lmin), sigma = theta / delta;
  const Sell& p = L.pat;
  // work-split replicated level: this rank's slices only, vectors allgathered after each matrix pass
  const bool split = L.split_chunk > 0;
  const int64_t s0 = split ? L.split_lo / MV_SLICE : 0;
  const int64_t s1 = split ? (L.split_hi > L.split_lo ? (L.split_hi + MV_SLICE - 1) / MV_SLICE : s0) : p.nslices;
  const int grid = split ? grid_for((s1 - s0 > 0 ? s1 - s0 : 1) * MV_SLICE, 8) : slice_grid(p);
  bool x_partial = false;      // xout is current on this rank's rows only
  double* dcur = deg > 1 ? L.d0 : nullptr;
  double* dnext = L.d1;
  const double* rin = b;
  if (xin && L.dist) MV_CHECK(halo_exchange_soa(L.halo, L.n, C, const_cast<double*>(xin), L.ntot));
  if (!xin) {
    MV_L k_smooth_zero<C><<<grid_for(L.n), MV_BLOCK, 0, g_stream>>>(L.n, L.ntot, b, L.dinv, 1.0 / theta, xout, dcur, done);
  } else {
    x_partial = split;
    if (L.Ac.ptr)
      MV_L k_smooth_csr<C, 0><<<grid_for(L.n * 32), MV_BLOCK, 0, g_stream>>>(L.n, L.ntot, L.Ac.ptr, L.Ac.col, L.Ac.val, b, xin, L.dinv, 0.0, 1.0 / theta,
                                                                        xout, dcur, deg > 1 ? L.r : nullptr, done);
    else if (L.valf)
      MV_L k_smooth_first<C, float><<<grid, MV_BLOCK, 0, g_stream>>>(s0, s1, L.n, L.ntot, p.slice_ptr, p.cols, L.valf, b, xin, L.dinv, 1.0 / theta,
                                                                xout, dcur, deg > 1 ? L.r : nullptr, done);
    else
      MV_L k_smooth_first<C, double><<<grid, MV_BLOCK, 0, g_stream>>>(s0, s1, L.n, L.ntot, p.slice_ptr, p.cols, L.val, b, xin, L.dinv, 1.0 / theta,
                                                                 xout, dcur, deg > 1 ? L.r : nullptr, done);
    rin = L.r;
  }
  double rho_k = 1.0 / sigma;
  for (int k = 1; k < deg; ++k) {
    const double rho_n = 1.0 / (2.0 * sigma - rho_k);
    const bool last = k == deg - 1;
    if (L.dist) MV_CHECK(halo_exchange_soa(L.halo, L.n, C, dcur, L.ntot));
    if (split && x_partial) MV_CHECK(allgather_rows_dev(dcur, L.ntot, C, L.split_chunk));   // (after k_smooth_zero d is complete everywhere)
    x_partial = split;
    if (L.Ac.ptr)
      MV_L k_smooth_csr<C, 1><<<grid_for(L.n * 32), MV_BLOCK, 0, g_stream>>>(L.n, L.ntot, L.Ac.ptr, L.Ac.col, L.Ac.val, rin, dcur, L.dinv, rho_n * rho_k,
                                                                        2.0 * rho_n / delta, xout, dnext, last ? nullptr : L.r, done);
    else if (L.valf)
      MV_L k_smooth_next<C, float><<<grid, MV_BLOCK, 0, g_stream>>>(s0, s1, L.n, L.ntot, p.slice_ptr, p.cols, L.valf, dcur, dnext, rin,
                                                               last ? nullptr : L.r, L.dinv, rho_n * rho_k, 2.0 * rho_n / delta, xout, done);
    else
      MV_L k_smooth_next<C, double><<<grid, MV_BLOCK, 0, g_stream>>>(s0, s1, L.n, L.ntot, p.slice_ptr, p.cols, L.val, dcur, dnext, rin,
                                                                last ? nullptr : L.r, L.dinv, rho_n * rho_k, 2.0 * rho_n / delta, xout, done);
    rin = L.r;
    rho_k = rho_n;
    double* t = dcur; dcur = dnext; dnext = t;
  }
  MV_KERNEL_CHECK();
  if (x_partial) MV_CHECK(allgather_rows_dev(xout, L.ntot, C, L.split_chunk));
  return MV_OK;
}

template <int C>
static int vcycle(mv_mesh* m, Amg* a, const double* rfine, double* z, const int* done) {
  const int nl = (int)a->lv.size();
  const double* b = rfine;
  for (int l = 0; l + 1 < nl; ++l) {
    AmgLevel& L = a->lv[l];
    AmgLevel& N = a->lv[l + 1];
    MV_CHECK(smooth<C>(L, b, nullptr, L.x0, done));
    if (L.dist) MV_CHECK(halo_exchange_soa(L.halo, L.n, C, L.x0, L.ntot));
    {
      const bool split = L.split_chunk > 0;
      const int64_t s0 = split ? L.split_lo / MV_SLICE : 0;
      const int64_t s1 = split ? (L.split_hi > L.split_lo ? (L.split_hi + MV_SLICE - 1) / MV_SLICE : s0) : L.pat.nslices;
      const int rg = split ? grid_for((s1 - s0 > 0 ? s1 - s0 : 1) * MV_SLICE, 8) : slice_grid(L.pat);
      if (L.Ac.ptr)
        MV_L k_smooth_csr<C, 2><<<grid_for(L.n * 32), MV_BLOCK, 0, g_stream>>>(L.n, L.ntot, L.Ac.ptr, L.Ac.col, L.Ac.val, b, L.x0, L.dinv, 0.0, 0.0, nullptr,
                                                                          nullptr, L.r, done);
      else if (L.valf)
        MV_L k_residual<C, float><<<rg, MV_BLOCK, 0, g_stream>>>(s0, s1, L.n, L.ntot, L.pat.slice_ptr, L.pat.cols, L.valf, b, L.x0, L.dinv, L.r, done);
      else
        MV_L k_residual<C, double><<<rg, MV_BLOCK, 0, g_stream>>>(s0, s1, L.n, L.ntot, L.pat.slice_ptr, L.pat.cols, L.val, b, L.x0, L.dinv, L.r, done);
      if (split) MV_CHECK(allgather_rows_dev(L.r, L.ntot, C, L.split_chunk));   // the restriction below gathers all of r
    }
    if (L.Rc.ptr)
      MV_L k_csr_apply_warp<C><<<grid_for(L.Rc.nrows * 32), MV_BLOCK, 0, g_stream>>>(L.Rc.nrows, L.Rc.ptr, L.Rc.col, L.Rc.val, L.r, L.ntot, N.b,
                                                                              N.ntot, done);
    else
      MV_L k_sell_apply<C, false><<<slice_grid(L.R), MV_BLOCK, 0, g_stream>>>(L.R.nslices, N.n, L.R.slice_ptr, L.R.cols, L.Rval, L.r, L.ntot, N.b, N.ntot, done);
    if (L.reduce_next) MV_CHECK(allreduce_sum_dev(N.b, (int)(N.ntot * C)));   // replicated coarse level: sum of the ranks' partial restrictions
    b = N.b;
  }
  AmgLevel& Lc = a->lv[nl - 1];
  double* xc = nl == 1 ? z : Lc.x0;
  if (a->coarse_inv) {
    const int64_t threads = (int64_t)Lc.n * 32;
    const unsigned blocks = (unsigned)((threads + MV_BLOCK - 1) / MV_BLOCK);
    if (Lc.dist) {
      // gather the coarse right-hand side of all ranks (each entry is non-zero on one rank: exact sum)
      const int N = (int)a->nc;
      MV_CUDA(cudaMemsetAsync(a->cglob, 0, (size_t)N * C * sizeof(double), g_stream));
      if (Lc.n) MV_L k_coarse_scatter<C><<<grid_for(Lc.n), MV_BLOCK, 0, g_stream>>>((int)Lc.n, N, Lc.goff, b, Lc.ntot, a->cglob);
      MV_CHECK(allreduce_sum_dev(a->cglob, N * C));
      if (Lc.n) MV_L k_dense_apply<C><<<blocks, MV_BLOCK, 0, g_stream>>>((int)Lc.n, N, Lc.goff, a->coarse_inv, a->cglob, N, xc, Lc.ntot, done);
    } else {
      MV_L k_dense_apply<C><<<blocks, MV_BLOCK, 0, g_stream>>>((int)Lc.n, (int)Lc.n, 0, a->coarse_inv, b, Lc.ntot, xc, Lc.ntot, done);
    }
  } else {
    MV_CHECK(smooth<C>(Lc, b, nullptr, xc, done));
  }
  MV_KERNEL_CHECK();
  const double* xcoarse = xc;
  for (int l = nl - 2; l >= 0; --l) {
    AmgLevel& L = a->lv[l];
    MV_L k_sell_apply<C, true><<<slice_grid(L.P), MV_BLOCK, 0, g_stream>>>(L.P.nslices, L.n, L.P.slice_ptr, L.P.cols, L.Pval, xcoarse, a->lv[l + 1].ntot, L.x0, L.ntot, done);
    double* out = l == 0 ? z : L.x1;
    MV_CHECK(smooth<C>(L, l == 0 ? rfine : L.b, L.x0, out, done));   // (exchanges the ghosts of L.x0 first)
    xcoarse = out;
  }
  (void)m;
  return MV_OK;
}

int amg_apply(mv_mesh* m, Amg* a, int C, const double* r, double* z, const int* done) {
  MV_CHECK(amg_ensure_work(a, C));
  switch (C) {
    case 1: return vcycle<1>(m, a, r, z, done);
    case 2: return vcycle<2>(m, a, r, z, done);
    case 3: return vcycle<3>(m, a, r, z, done);
  }
  set_error("amg: unsupported number of right-hand sides %d", C);
  return MV_ERR_UNSUPPORTED;
}

}  // namespace mv
