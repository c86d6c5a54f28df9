// SELL-32 SpMV and Jacobi-preconditioned CG with C interleaved right-hand sides.
//
// replaces: PETSc KSP `cg` + PC `bjacobi`/`ilu` (movement/solver_parameters.py:62-66) used
//           for the consistent-mass solves at movement/monge_ampere.py:340 (d components)
//           and :483 (d*d components), the vector Laplacian at
//           movement/laplacian_smoothing.py:75, and KSP `cg` + `gamg` with a constant
//           nullspace (solver_parameters.py:68-71) at monge_ampere.py:441-448.
//
// One PCG iteration is three kernels (no host involvement, scalars stay on the device):
//   k_pcg_spmv    q = A p                      + partial p.q        12*nnz+4 +16c B/vertex
//   k_pcg_update  x += a p ; r -= a q ; z=D^-1 r + partial r.z          56c B/vertex
//   k_pcg_dir     p = D^-1 r + b p                                      32c B/vertex
// which is the B_cg(c) = 12*nnz + 4 + 104*c bytes/vertex/iteration model of SURVEY 8d.
// All dot products use the deterministic last-block reduction (mv_internal.cuh), so a
// solve is bit-reproducible run to run.  After convergence the `done` flag turns the
// remaining launches of a batch into no-ops; the host polls the flag once per batch.
#include "mv_internal.cuh"

namespace mv {

template <int C>
__device__ __forceinline__ bool masked(const uint16_t* __restrict__ mask, int64_t v, int c) {
  return mask != nullptr && ((__ldg(&mask[v]) >> c) & 1);
}

// coupled systems (springs): the C interleaved components form ONE vector, so every inner
// product is the sum over components, broadcast back to all C slots
template <int C, int N>
__device__ __forceinline__ void couple(double (&t)[N]) {
  double s = 0.0;
#pragma unroll
  for (int c = 0; c < C; ++c) s += t[c];
#pragma unroll
  for (int c = 0; c < C; ++c) t[c] = s;
}

// ------------------------------------------------------------------ SpMV
// One thread per row, one warp per 32-row slice; matrix streams (cols, vals) are read
// once, fully coalesced, with streaming (evict-first) loads; x is gathered through the
// read-only path so neighbouring rows share lines in L1/L2.
template <int C, bool DOT>
__global__ void __launch_bounds__(MV_BLOCK)
k_spmv(int64_t nslices, int64_t nrows, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
       const double* __restrict__ val, const double* __restrict__ x, double* __restrict__ y,
       const int* __restrict__ done, double* __restrict__ partials, unsigned* __restrict__ counter,
       double* __restrict__ scal, int dot_slot) {
  if (DOT && done != nullptr && *done) return;
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  double dot[C];
#pragma unroll
  for (int c = 0; c < C; ++c) dot[c] = 0.0;
  for (int64_t s = warp0; s < nslices; s += nwarps) {
    const int64_t base = __ldg(&slice_ptr[s]);
    const int w = (int)((__ldg(&slice_ptr[s + 1]) - base) >> 5);
    const int64_t row = s * MV_SLICE + lane;
    const int* cp = cols + base + lane;
    const double* vp = val + base + lane;
    double acc[C];
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = 0.0;
#pragma unroll 4
    for (int k = 0; k < w; ++k) {
      const int col = __ldcs(cp + k * MV_SLICE);
      const double a = __ldcs(vp + k * MV_SLICE);
      const double* xp = x + (int64_t)col * C;
#pragma unroll
      for (int c = 0; c < C; ++c) acc[c] = fma(a, __ldg(xp + c), acc[c]);
    }
    if (row < nrows) {
#pragma unroll
      for (int c = 0; c < C; ++c) {
        y[row * C + c] = acc[c];
        if (DOT) dot[c] = fma(acc[c], __ldg(&x[row * C + c]), dot[c]);
      }
    }
  }
  if (DOT) {
    grid_reduce_sum<C>(dot, partials, counter, [&](double(&t)[C]) {
#pragma unroll
      for (int c = 0; c < C; ++c) scal[SC(dot_slot, c)] = t[c];
    });
  }
}

// ------------------------------------------------------------------ PCG pieces
// r = P (b - mean - A x), p = z = D^-1 r ; sums r.z and b.D^-1.b (free dofs)
template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_pcg_init(int64_t nrows, const double* __restrict__ b, const double* __restrict__ q /* = A x */,
           const double* __restrict__ dinv, const uint16_t* __restrict__ mask, double* __restrict__ r,
           double* __restrict__ p, double* __restrict__ partials, unsigned* __restrict__ counter,
           double* __restrict__ scal, double inv_nglobal, int project_mean, int dpd, int coupled) {
  double acc[2 * C];
#pragma unroll
  for (int c = 0; c < 2 * C; ++c) acc[c] = 0.0;
  double mean[C];
#pragma unroll
  for (int c = 0; c < C; ++c) mean[c] = project_mean ? scal[SC(S_SUM, c)] * inv_nglobal : 0.0;
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int64_t i = v * C + c;
      const double di = __ldg(&dinv[dpd ? i : v]);
      const bool fix = masked<C>(mask, v, c);
      const double bb = b ? (b[i] - mean[c]) : 0.0;
      const double rr = fix ? 0.0 : bb - q[i];
      const double zz = di * rr;
      r[i] = rr;
      p[i] = zz;
      acc[c] = fma(rr, zz, acc[c]);
      acc[C + c] = fix ? acc[C + c] : fma(bb * di, bb, acc[C + c]);
    }
  }
  grid_reduce_sum<2 * C>(acc, partials, counter, [&](double(&t)[2 * C]) {
    if (coupled) couple<C>(t), couple<C>(*(double(*)[C])(&t[C]));
#pragma unroll
    for (int c = 0; c < C; ++c) {
      scal[SC(S_RZ, c)] = t[c];
      scal[SC(S_BB, c)] = t[C + c];
    }
  });
}

// thresholds + initial convergence test (one thread)
__global__ void k_pcg_setup(int C, double rtol, double atol, double* __restrict__ scal, int* __restrict__ flags) {
  bool all = true;
  for (int c = 0; c < C; ++c) {
    const double rz = scal[SC(S_RZ, c)], bb = scal[SC(S_BB, c)];
    const double ref = fmax(rz, bb);
    scal[SC(S_REF, c)] = ref;
    const double thr = fmax(rtol * rtol * ref, atol * atol);
    scal[SC(S_THR, c)] = thr;
    all = all && (rz <= thr);
  }
  flags[0] = all ? 1 : 0;
  flags[1] = 0;
  flags[2] = 0;
}

// after r.z (new) is known: beta, convergence, iteration count (one thread)
__device__ __forceinline__ void pcg_post_update(int C, double* __restrict__ scal, int* __restrict__ flags) {
  bool all = true;
  for (int c = 0; c < C; ++c) {
    const double rzn = scal[SC(S_RZNEW, c)], rz = scal[SC(S_RZ, c)];
    scal[SC(S_BETA, c)] = rz > 0.0 ? rzn / rz : 0.0;
    scal[SC(S_RZ, c)] = rzn;
    all = all && (rzn <= scal[SC(S_THR, c)]);
    if (!(rzn == rzn)) flags[2] = 1;  // NaN: breakdown
  }
  flags[1] += 1;
  if (all || flags[2]) flags[0] = 1;
}
__global__ void k_pcg_post_update(int C, double* __restrict__ scal, int* __restrict__ flags) {
  if (flags[0]) return;
  pcg_post_update(C, scal, flags);
}

// x += alpha p ; r -= alpha q ; partial r.D^-1.r
template <int C, bool LOCAL_POST>
__global__ void __launch_bounds__(MV_BLOCK)
k_pcg_update(int64_t nrows, const double* __restrict__ p, const double* __restrict__ q, const double* __restrict__ dinv,
             const uint16_t* __restrict__ mask, double* __restrict__ x, double* __restrict__ r,
             int* __restrict__ flags, double* __restrict__ partials, unsigned* __restrict__ counter,
             double* __restrict__ scal, int dpd, int coupled) {
  if (flags[0]) return;
  __shared__ double s_alpha[C];
  if (threadIdx.x < C) {
    const double pq = scal[SC(S_PQ, threadIdx.x)], rz = scal[SC(S_RZ, threadIdx.x)];
    s_alpha[threadIdx.x] = pq > 0.0 ? rz / pq : 0.0;
  }
  __syncthreads();
  double acc[C];
#pragma unroll
  for (int c = 0; c < C; ++c) acc[c] = 0.0;
  const int64_t n = nrows * C;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * C;
  // thread t handles elements t*C .. t*C+C-1 of each stride block?  No: keep it flat and
  // coalesced -- element i belongs to vertex i/C, component i%C.
  for (int64_t i0 = (blockIdx.x * (int64_t)blockDim.x) * C + threadIdx.x; i0 < n; i0 += stride) {
#pragma unroll
    for (int k = 0; k < C; ++k) {
      const int64_t i = i0 + (int64_t)k * blockDim.x;
      if (i < n) {
        const int64_t v = i / C;
        const int c = (int)(i - v * C);
        const double al = s_alpha[c];
        const double pp = p[i], qq = __ldcs(&q[i]);
        const double xx = x[i] + al * pp;
        const double rr = r[i] - al * qq;
        x[i] = xx;
        r[i] = rr;
        const bool fix = masked<C>(mask, v, c);
        const double zz = fix ? 0.0 : __ldg(&dinv[dpd ? i : v]) * rr;
        // accumulate into the slot of this element's component
#pragma unroll
        for (int cc = 0; cc < C; ++cc) acc[cc] = (cc == c) ? fma(rr, zz, acc[cc]) : acc[cc];
      }
    }
  }
  grid_reduce_sum<C>(acc, partials, counter, [&](double(&t)[C]) {
    if (coupled) couple<C>(t);
#pragma unroll
    for (int c = 0; c < C; ++c) scal[SC(S_RZNEW, c)] = t[c];
    if (LOCAL_POST) pcg_post_update(C, scal, flags);
  });
}

// p = D^-1 r + beta p
template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_pcg_dir(int64_t nrows, const double* __restrict__ r, const double* __restrict__ dinv, const uint16_t* __restrict__ mask,
          double* __restrict__ p, const int* __restrict__ flags, const double* __restrict__ scal, int dpd) {
  if (flags[0]) return;
  __shared__ double s_beta[C];
  if (threadIdx.x < C) s_beta[threadIdx.x] = scal[SC(S_BETA, threadIdx.x)];
  __syncthreads();
  const int64_t n = nrows * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t v = i / C;
    const int c = (int)(i - v * C);
    const bool fix = masked<C>(mask, v, c);
    const double zz = fix ? 0.0 : __ldg(&dinv[dpd ? i : v]) * __ldcs(&r[i]);
    p[i] = fma(s_beta[c], p[i], zz);
  }
}

// sum of each component (for the mean projection)
template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_comp_sum(int64_t nrows, const double* __restrict__ x, double* __restrict__ partials, unsigned* __restrict__ counter,
           double* __restrict__ scal) {
  double acc[C];
#pragma unroll
  for (int c = 0; c < C; ++c) acc[c] = 0.0;
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x)
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] += x[v * C + c];
  grid_reduce_sum<C>(acc, partials, counter, [&](double(&t)[C]) {
#pragma unroll
    for (int c = 0; c < C; ++c) scal[SC(S_SUM, c)] = t[c];
  });
}
template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_sub_mean(int64_t nrows, double* __restrict__ x, const double* __restrict__ scal, double inv_nglobal) {
  const int64_t n = nrows * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    x[i] -= scal[SC(S_SUM, (int)(i % C))] * inv_nglobal;
}

__global__ void k_add_vals(int64_t n, const double* __restrict__ a, const double* __restrict__ b, double* __restrict__ c) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) c[i] = a[i] + b[i];
}
__global__ void k_dinv_sum(int64_t n, const double* __restrict__ a, const double* __restrict__ b, double* __restrict__ c) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double d = (a[i] != 0.0 ? 1.0 / a[i] : 0.0) + (b[i] != 0.0 ? 1.0 / b[i] : 0.0);
    c[i] = d > 0.0 ? 1.0 / d : 0.0;
  }
}

int op_arrays(mv_mesh* m, int op, const double** val, const double** dinv) {
  if (!m->Mval) {
    set_error("no operators on a %dD mesh (interval meshes carry geometry for the tangling check only)", m->dim);
    return MV_ERR_UNSUPPORTED;
  }
  switch (op) {
    case MV_OP_MASS: *val = m->Mval; *dinv = m->dinvM; return MV_OK;
    case MV_OP_STIFFNESS: *val = m->Kval; *dinv = m->dinvK; return MV_OK;
    case MV_OP_MASS_PLUS_STIFFNESS:
      if (!m->MKval) {
        MV_CHECK(dev_alloc(&m->MKval, m->pat.len));
        MV_CHECK(dev_alloc(&m->dinvMK, m->Vown));
        MV_L k_add_vals<<<grid_for(m->pat.len), MV_BLOCK, 0, g_stream>>>(m->pat.len, m->Mval, m->Kval, m->MKval);
        MV_L k_dinv_sum<<<grid_for(m->Vown), MV_BLOCK, 0, g_stream>>>(m->Vown, m->dinvM, m->dinvK, m->dinvMK);
        MV_KERNEL_CHECK();
      }
      *val = m->MKval; *dinv = m->dinvMK; return MV_OK;
  }
  set_error("unknown operator %d", op);
  return MV_ERR_BADARG;
}

static int spmv_grid(const mv_mesh* m) {
  // one warp per slice, persistent grid: 8 CTAs of 256 threads per SM
  return grid_for(m->pat.nslices * MV_SLICE, 8);
}

template <int C>
static int spmv_launch(mv_mesh* m, const double* val, const double* x, double* y, bool dot, int slot) {
  const Sell& p = m->pat;
  const int grid = spmv_grid(m);
  if (dot)
    MV_L k_spmv<C, true><<<grid, MV_BLOCK, 0, g_stream>>>(p.nslices, p.nrows, p.slice_ptr, p.cols, val, x, y, m->pcg.flags,
                                                     m->red.partials, m->red.counter, m->pcg.scal, slot);
  else
    MV_L k_spmv<C, false><<<grid, MV_BLOCK, 0, g_stream>>>(p.nslices, p.nrows, p.slice_ptr, p.cols, val, x, y, nullptr, nullptr,
                                                      nullptr, nullptr, 0);
  MV_KERNEL_CHECK();
  return MV_OK;
}

static int ensure_pcg_work(mv_mesh* m) {
  PcgWork& w = m->pcg;
  if (w.r) return MV_OK;
  const size_t maxc = (size_t)m->dim * m->dim;
  MV_CHECK(dev_alloc(&w.r, m->V * maxc));
  MV_CHECK(dev_alloc(&w.q, m->V * maxc));
  MV_CHECK(dev_alloc(&w.p, m->V * maxc));
  MV_CHECK(dev_alloc(&w.xs, m->V * maxc));
  MV_CUDA(cudaMemsetAsync(w.xs, 0, m->V * maxc * sizeof(double), g_stream));
  MV_CHECK(dev_alloc(&w.scal, S_NSLOTS * 16));
  MV_CHECK(dev_alloc(&w.flags, 4));
  MV_CUDA(cudaMemsetAsync(w.scal, 0, S_NSLOTS * 16 * sizeof(double), g_stream));
  MV_CUDA(cudaMemsetAsync(w.flags, 0, 4 * sizeof(int), g_stream));
  MV_CUDA(cudaMemsetAsync(w.p, 0, m->V * maxc * sizeof(double), g_stream));
  MV_CUDA(cudaMallocHost((void**)&w.h_flags, 4 * sizeof(int)));
  MV_CUDA(cudaMallocHost((void**)&w.h_scal, S_NSLOTS * 16 * sizeof(double)));
  return MV_OK;
}

template <int C>
static int apply_op(mv_mesh* m, const LinOp& op, const uint16_t* mask, const double* x, double* y, bool dot) {
  if (op.val) return spmv_launch<C>(m, op.val, x, y, dot, S_PQ);
  return spring_apply_dev(m, op.spring_coords, mask, x, y, dot);
}

template <int C>
static int pcg_impl(mv_mesh* m, const LinOp& op, const double* rhs, double* x, const uint16_t* mask,
                    double rtol, double atol, int maxit, int project_mean, int* iters, double* relres) {
  MV_CHECK(ensure_pcg_work(m));
  const double* dinv = op.dinv;
  const int dpd = op.dpd, coupled = op.coupled;
  PcgWork& w = m->pcg;
  const int64_t n = m->Vown;
  const bool dist = comm_active();
  int64_t nglobal = n;
  if (dist) {
    // global row count for the mean projection
    double hn = (double)n;
    MV_CUDA(cudaMemcpyAsync(w.scal + SC(S_SUM, 15), &hn, sizeof(double), cudaMemcpyHostToDevice, g_stream));
    MV_CHECK(allreduce_sum_dev(w.scal + SC(S_SUM, 15), 1));
    MV_CUDA(cudaMemcpyAsync(&hn, w.scal + SC(S_SUM, 15), sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    nglobal = (int64_t)(hn + 0.5);
  }
  const double inv_ng = 1.0 / (double)nglobal;
  const int gv = grid_for(n);
  const int ge = grid_for(n * C);
  if (project_mean && rhs) {
    MV_L k_comp_sum<C><<<gv, MV_BLOCK, 0, g_stream>>>(n, rhs, m->red.partials, m->red.counter, w.scal);
    MV_KERNEL_CHECK();
    if (dist) MV_CHECK(allreduce_sum_dev(w.scal + SC(S_SUM, 0), C));
  }
  // initial residual
  if (dist) MV_CHECK(halo_exchange(m, C, x));
  MV_CHECK(apply_op<C>(m, op, mask, x, w.q, false));
  MV_L k_pcg_init<C><<<gv, MV_BLOCK, 0, g_stream>>>(n, rhs, w.q, dinv, mask, w.r, w.p, m->red.partials, m->red.counter, w.scal,
                                               inv_ng, project_mean && rhs ? 1 : 0, dpd, coupled);
  MV_KERNEL_CHECK();
  if (dist) {
    MV_CHECK(allreduce_sum_dev(w.scal + SC(S_RZ, 0), C));
    MV_CHECK(allreduce_sum_dev(w.scal + SC(S_BB, 0), C));
  }
  MV_L k_pcg_setup<<<1, 1, 0, g_stream>>>(C, rtol, atol, w.scal, w.flags);
  MV_KERNEL_CHECK();

  int it = 0, batch = 8;
  w.h_flags[0] = 0;
  while (true) {
    MV_CUDA(cudaMemcpyAsync(w.h_flags, w.flags, 4 * sizeof(int), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    if (g_prof.used) prof_drain();
    if (w.h_flags[0] || it >= maxit) break;
    const int nb = batch < maxit - it ? batch : maxit - it;
    for (int k = 0; k < nb; ++k) {
      if (dist) MV_CHECK(halo_exchange(m, C, w.p));
      const int pslot = prof_begin(C);
      MV_CHECK(apply_op<C>(m, op, mask, w.p, w.q, true));
      prof_end(pslot);
      if (dist) {
        MV_CHECK(allreduce_sum_dev(w.scal + SC(S_PQ, 0), C));
        MV_L k_pcg_update<C, false><<<ge, MV_BLOCK, 0, g_stream>>>(n, w.p, w.q, dinv, mask, x, w.r, w.flags, m->red.partials,
                                                              m->red.counter, w.scal, dpd, coupled);
        MV_CHECK(allreduce_sum_dev(w.scal + SC(S_RZNEW, 0), C));
        MV_L k_pcg_post_update<<<1, 1, 0, g_stream>>>(C, w.scal, w.flags);
      } else {
        MV_L k_pcg_update<C, true><<<ge, MV_BLOCK, 0, g_stream>>>(n, w.p, w.q, dinv, mask, x, w.r, w.flags, m->red.partials,
                                                             m->red.counter, w.scal, dpd, coupled);
      }
      MV_L k_pcg_dir<C><<<ge, MV_BLOCK, 0, g_stream>>>(n, w.r, dinv, mask, w.p, w.flags, w.scal, dpd);
    }
    MV_KERNEL_CHECK();
    it += nb;
    if (batch < 64) batch *= 2;
  }
  const int done = w.h_flags[0], niter = w.h_flags[1], breakdown = w.h_flags[2];
  g_prof.pcg_iters[C & 15] += niter;
  if (project_mean) {
    MV_L k_comp_sum<C><<<gv, MV_BLOCK, 0, g_stream>>>(n, x, m->red.partials, m->red.counter, w.scal);
    if (dist) MV_CHECK(allreduce_sum_dev(w.scal + SC(S_SUM, 0), C));
    MV_L k_sub_mean<C><<<ge, MV_BLOCK, 0, g_stream>>>(n, x, w.scal, inv_ng);
    MV_KERNEL_CHECK();
  }
  if (iters) *iters = niter;
  if (relres) {
    MV_CUDA(cudaMemcpyAsync(w.h_scal, w.scal, S_NSLOTS * 16 * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    for (int c = 0; c < C; ++c) {
      const double ref = w.h_scal[SC(S_REF, c)];
      relres[c] = ref > 0.0 ? sqrt(w.h_scal[SC(S_RZ, c)] / ref) : 0.0;
    }
  }
  if (breakdown) {
    set_error("pcg: breakdown (NaN) after %d iterations", niter);
    return MV_ERR_NOT_CONVERGED;
  }
  if (!done) {
    set_error("pcg: not converged in %d iterations", maxit);
    return MV_ERR_NOT_CONVERGED;
  }
  return MV_OK;
}

int pcg_solve(mv_mesh* m, int op, int ncomp, const double* rhs, double* x, const uint16_t* mask, double rtol, double atol,
              int maxit, int project_mean, int* iters, double* relres) {
  LinOp lop;
  MV_CHECK(op_arrays(m, op, &lop.val, &lop.dinv));
  if (ncomp > m->dim * m->dim) {
    set_error("pcg: ncomp %d exceeds dim^2", ncomp);
    return MV_ERR_BADARG;
  }
  if (op == MV_OP_MASS && g_mass_solver == MV_MASS_CHEBYSHEV && !project_mean)
    return cheb_mass_solve(m, ncomp, lop.val, lop.dinv, rhs, x, mask, rtol, atol, maxit, iters, relres);
  Amg* pc = nullptr;
  if (ncomp <= 3) MV_CHECK(amg_get(m, op, mask, ncomp, project_mean, &pc));
  return pcg_solve_soa(m, ncomp, lop.val, lop.dinv, rhs, x, mask, rtol, atol, maxit, project_mean, iters, relres, pc);
}

int pcg_solve_spring(mv_mesh* m, const double* coords, const double* dinv_dof, const double* rhs, double* x, const uint16_t* mask,
                     double rtol, double atol, int maxit, int* iters) {
  LinOp lop;
  lop.val = nullptr;
  lop.dinv = dinv_dof;
  lop.dpd = 1;
  lop.coupled = 1;
  lop.spring_coords = coords;
  return pcg_impl<2>(m, lop, rhs, x, mask, rtol, atol, maxit, 0, iters, nullptr);
}

int ensure_pcg_workspace(mv_mesh* m) { return ensure_pcg_work(m); }

int spmv(mv_mesh* m, int op, int ncomp, const double* x, double* y) {
  const double *val, *dinv;
  MV_CHECK(op_arrays(m, op, &val, &dinv));
  switch (ncomp) {
    case 1: return spmv_launch<1>(m, val, x, y, false, 0);
    case 2: return spmv_launch<2>(m, val, x, y, false, 0);
    case 3: return spmv_launch<3>(m, val, x, y, false, 0);
    case 4: return spmv_launch<4>(m, val, x, y, false, 0);
    case 9: return spmv_launch<9>(m, val, x, y, false, 0);
  }
  set_error("spmv: unsupported ncomp %d (1,2,3,4,9)", ncomp);
  return MV_ERR_UNSUPPORTED;
}

}  // namespace mv

using namespace mv;

extern "C" int mv_spmv(mv_mesh* m, int op, int ncomp, const double* x, double* y) {
  if (!m || !x || !y) return MV_ERR_BADARG;
  Staged sx, sy;
  MV_CHECK(sx.in(x, m->V * ncomp * sizeof(double)));
  MV_CHECK(sy.out_begin(y, m->Vown * ncomp * sizeof(double)));
  MV_CHECK(spmv(m, op, ncomp, sx.as<double>(), sy.as<double>()));
  return sy.out_end(y, m->Vown * ncomp * sizeof(double));
}

extern "C" int mv_pcg_solve(mv_mesh* m, int op, int ncomp, const double* rhs, double* x, const uint16_t* mask, double rtol,
                            double atol, int maxit, int project_mean, int* iters, double* relres) {
  if (!m || !x) return MV_ERR_BADARG;
  Staged sb, sx, sm;
  const size_t nb = m->V * ncomp * sizeof(double);
  MV_CHECK(sb.in(rhs, m->Vown * ncomp * sizeof(double)));
  MV_CHECK(sx.in(x, nb));  // initial guess in, solution out
  MV_CHECK(sm.in(mask, m->Vown * sizeof(uint16_t)));
  int st = pcg_solve(m, op, ncomp, sb.as<double>(), sx.as<double>(), sm.as<uint16_t>(), rtol, atol, maxit, project_mean,
                     iters, relres);
  if (sx.owned) {
    MV_CUDA(cudaMemcpyAsync(x, sx.dev, nb, cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
  }
  return st;
}
