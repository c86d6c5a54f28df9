// Preconditioned CG for the assembled (SELL-32) operators with C right-hand sides sharing one
// matrix stream, working vectors held COMPONENT-MAJOR ([C][ld], "SoA").
//
// replaces: PETSc KSP `cg` (movement/solver_parameters.py:62-71) at movement/monge_ampere.py:340
//           (d mass solves), :441-448 (Poisson, gamg), :483 (d*d mass solves) and
//           movement/laplacian_smoothing.py:75 (d stiffness solves).
//
// Why component-major inside the solve: the user-facing fields are Firedrake's interleaved
// [V,C] arrays, but a warp that gathers x[col*C + c] for 32 consecutive rows touches 32 sectors per
// load instruction (stride 8C bytes) and the C = 9 SpMV became LSU-bound at 2.1 TB/s (ncu, round 1).
// With [C][ld] every gather/stream instruction of a warp covers 256 contiguous bytes, and the
// per-component dot products need no index arithmetic.  The right-hand side and the initial guess
// are converted on entry, the solution on exit (two passes per solve, < 1 % of its traffic).
//
// One iteration: k_spmv_soa (q = A p, fused p.q) -> k_update_soa (x += a p, r -= a q, fused
// r.D^-1 r, beta, convergence flag) -> k_dir_soa (p = D^-1 r + b p); with the multigrid
// preconditioner the update carries no reduction and z = V-cycle(r) follows (mv_amg_cycle.cu).
// Scalars stay on the device; reductions are deterministic (mv_internal.cuh).
#include <string.h>

#include "mv_internal.cuh"

namespace mv {

// Graphs that hold NCCL nodes must be gone before the communicator is destroyed (ncclCommDestroy waits for them: a mover
// still alive at mv_comm_finalize hung the 8-GPU converged run at exit, round 2), so every captured graph is registered.
static std::vector<PcgWork*> g_graph_owners;
void pcg_graph_release(PcgWork& w) {
  if (w.graph_exec) {
    cudaGraphExecDestroy((cudaGraphExec_t)w.graph_exec);
    w.graph_exec = nullptr;
  }
  for (size_t i = 0; i < g_graph_owners.size(); ++i)
    if (g_graph_owners[i] == &w) { g_graph_owners.erase(g_graph_owners.begin() + i); break; }
}
void pcg_graph_release_all() {
  for (PcgWork* w : g_graph_owners)
    if (w->graph_exec) {
      cudaGraphExecDestroy((cudaGraphExec_t)w->graph_exec);
      w->graph_exec = nullptr;
    }
  g_graph_owners.clear();
}

#define VERTEX_LOOP(v, n) for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < (n); v += (int64_t)gridDim.x * blockDim.x)

template <int C>
__device__ __forceinline__ bool fixed_dof(const uint16_t* __restrict__ mask, int64_t v, int c) {
  return mask != nullptr && ((__ldg(&mask[v]) >> c) & 1);
}

// ------------------------------------------------------------------ layout conversion
template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_aos_to_soa(int64_t n, const double* __restrict__ src, double* __restrict__ dst, int64_t ld) {
  __shared__ double tile[MV_BLOCK * C];
  for (int64_t v0 = blockIdx.x * (int64_t)MV_BLOCK; v0 < n; v0 += (int64_t)gridDim.x * MV_BLOCK) {
    const int nv = (int)(n - v0 < MV_BLOCK ? n - v0 : MV_BLOCK);
    for (int e = threadIdx.x; e < nv * C; e += MV_BLOCK) tile[e] = src[v0 * C + e];
    __syncthreads();
    if ((int)threadIdx.x < nv)
#pragma unroll
      for (int c = 0; c < C; ++c) dst[c * ld + v0 + threadIdx.x] = tile[threadIdx.x * C + c];
    __syncthreads();
  }
}
template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_soa_to_aos(int64_t n, const double* __restrict__ src, int64_t ld, double* __restrict__ dst) {
  __shared__ double tile[MV_BLOCK * C];
  for (int64_t v0 = blockIdx.x * (int64_t)MV_BLOCK; v0 < n; v0 += (int64_t)gridDim.x * MV_BLOCK) {
    const int nv = (int)(n - v0 < MV_BLOCK ? n - v0 : MV_BLOCK);
    if ((int)threadIdx.x < nv)
#pragma unroll
      for (int c = 0; c < C; ++c) tile[threadIdx.x * C + c] = src[c * ld + v0 + threadIdx.x];
    __syncthreads();
    for (int e = threadIdx.x; e < nv * C; e += MV_BLOCK) dst[v0 * C + e] = tile[e];
    __syncthreads();
  }
}

// ------------------------------------------------------------------ SpMV, component-major
template <int C, bool DOT, int VAR>
__global__ void __launch_bounds__(MV_BLOCK)
k_spmv_soa(int64_t nslices, int64_t nrows, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
           const double* __restrict__ val, const double* __restrict__ x, int64_t ld, double* __restrict__ y,
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
    // the column indices of a chunk of KB entries are fetched first, so KB*C gathers are in flight
    // per thread instead of one dependent (index -> gather) chain per entry
    constexpr int KB = C >= 9 ? 5 : (C >= 3 ? 8 : 8);
    for (int k0 = 0; k0 < w; k0 += KB) {
      int col[KB];
      double a[KB];
#pragma unroll
      for (int j = 0; j < KB; ++j) {
        const bool in = k0 + j < w;
        col[j] = in ? __ldcs(cp + (k0 + j) * MV_SLICE) : 0;
        a[j] = in ? __ldcs(vp + (k0 + j) * MV_SLICE) : 0.0;
      }
      if (VAR == 0) {
#pragma unroll
        for (int j = 0; j < KB; ++j) {
          const double* xp = x + col[j];
#pragma unroll
          for (int c = 0; c < C; ++c) acc[c] = fma(a[j], __ldg(xp + c * ld), acc[c]);
        }
      } else {   // component-outer: neighbouring columns of one component are gathered back to back (L1 reuse)
#pragma unroll
        for (int c = 0; c < C; ++c) {
          const double* xc = x + c * ld;
#pragma unroll
          for (int j = 0; j < KB; ++j) acc[c] = fma(a[j], __ldg(xc + col[j]), acc[c]);
        }
      }
    }
    if (row < nrows) {
#pragma unroll
      for (int c = 0; c < C; ++c) {
        y[c * ld + row] = acc[c];
        if (DOT) dot[c] = fma(acc[c], __ldg(&x[c * ld + row]), dot[c]);
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
// r = P (b - mean - A x), p = z = D^-1 r ; sums r.z and b.D^-1.b over the free dofs.  b is interleaved [n,C].
// EUCL (multigrid-preconditioned solves, C <= 3): also the Euclidean r.r and b.b, kept at component index 8 + c of the
// same scalar rows -- the solve then stops only when BOTH sqrt(r.M^-1 r) and ||r||_2 have dropped by rtol, so that
// "rtol" holds in the unpreconditioned 2-norm the true-residual audit measures (the two norms differ by up to
// sqrt(cond) on fine meshes: 2.2e-11 instead of 1e-12 at 256^2)
#define EU 8
template <int C, bool EUCL>
__global__ void __launch_bounds__(MV_BLOCK)
k_init_soa(int64_t n, int64_t ld, const double* __restrict__ b, const double* __restrict__ q, const double* __restrict__ dinv,
           const uint16_t* __restrict__ mask, double* __restrict__ r, double* __restrict__ p, double* __restrict__ partials,
           unsigned* __restrict__ counter, double* __restrict__ scal, double inv_nglobal, int project_mean) {
  constexpr int NA = EUCL ? 4 * C : 2 * C;
  double acc[NA];
#pragma unroll
  for (int c = 0; c < NA; ++c) acc[c] = 0.0;
  double mean[C];
#pragma unroll
  for (int c = 0; c < C; ++c) mean[c] = project_mean ? scal[SC(S_SUM, c)] * inv_nglobal : 0.0;
  VERTEX_LOOP(v, n) {
    const double di = __ldg(&dinv[v]);
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const bool fix = fixed_dof<C>(mask, v, c);
      const double bb = b ? (b[v * C + c] - mean[c]) : 0.0;
      const double rr = fix ? 0.0 : bb - q[c * ld + v];
      const double zz = di * rr;
      r[c * ld + v] = rr;
      p[c * ld + v] = zz;
      acc[c] = fma(rr, zz, acc[c]);
      acc[C + c] = fix ? acc[C + c] : fma(bb * di, bb, acc[C + c]);
      if (EUCL) {
        acc[2 * C + c] = fma(rr, rr, acc[2 * C + c]);
        acc[3 * C + c] = fix ? acc[3 * C + c] : fma(bb, bb, acc[3 * C + c]);
      }
    }
  }
  grid_reduce_sum<NA>(acc, partials, counter, [&](double(&t)[NA]) {
#pragma unroll
    for (int c = 0; c < C; ++c) {
      scal[SC(S_RZ, c)] = t[c];
      scal[SC(S_BB, c)] = t[C + c];
      if (EUCL) {
        scal[SC(S_RZ, EU + c)] = t[2 * C + c];
        scal[SC(S_BB, EU + c)] = t[3 * C + c];
      }
    }
  });
}

// Multigrid-preconditioned solves: the stopping rule is relative to max(b, r0) in the preconditioned norm, but only
// r0.M^-1 r0 is available without an extra V-cycle.  The ratio of the two in the Jacobi norm (both known from
// k_init_soa) carries it over: ref = r0.z0 * max(1, b.D^-1 b / r0.D^-1 r0).  With a zero initial guess r0 = b and
// nothing changes; with a good guess the target stays tied to b instead of tightening with the guess.
__global__ void k_ref_ratio(int C, int phase, double* __restrict__ scal) {
  for (int c = 0; c < C; ++c) {
    if (phase == 0) {
      const double rz = scal[SC(S_RZ, c)], bb = scal[SC(S_BB, c)];
      scal[SC(S_SUM, c)] = (rz > 0.0 && bb > rz) ? bb / rz : 1.0;
    } else {
      scal[SC(S_BB, c)] *= scal[SC(S_SUM, c)];
    }
  }
}

__global__ void k_setup_soa(int C, int eucl, double rtol, double atol, double* __restrict__ scal, int* __restrict__ flags) {
  bool all = true;
  for (int c = 0; c < C; ++c) {
    const double rz = scal[SC(S_RZ, c)], bb = scal[SC(S_BB, c)];
    const double ref = fmax(rz, bb);
    scal[SC(S_REF, c)] = ref;
    const double thr = fmax(rtol * rtol * ref, atol * atol);
    scal[SC(S_THR, c)] = thr;
    all = all && (rz <= thr);
    if (C <= EU) {
      // Euclidean twin of the test (eucl == 0: switched off, threshold +inf and a zero in the r.r slot)
      const double rr = eucl ? scal[SC(S_RZ, EU + c)] : 0.0, b2 = eucl ? scal[SC(S_BB, EU + c)] : 0.0;
      const double thr2 = eucl ? fmax(rtol * rtol * fmax(rr, b2), atol * atol) : 1.7976931348623157e308;
      scal[SC(S_THR, EU + c)] = thr2;
      scal[SC(S_RZNEW, EU + c)] = rr;
      all = all && (rr <= thr2);
    }
  }
  flags[0] = all ? 1 : 0;
  flags[1] = 0;
  flags[2] = 0;
  flags[3] = 0;
}

// flags[3]: iteration at which the preconditioned-norm test was first met (0: not yet).  The Euclidean twin may need a
// few more iterations; it is given max(8, 25 %) extra ones and then the iterate is accepted on the preconditioned test
// alone (the two norms differ by up to sqrt(cond); on the finest 2D meshes the recurrence residual's 2-norm can idle
// just above the threshold, and an unbounded wait would spin to maxit).
__device__ __forceinline__ void post_update_soa(int C, double* __restrict__ scal, int* __restrict__ flags) {
  bool pre = true, euc = true;
  for (int c = 0; c < C; ++c) {
    const double rzn = scal[SC(S_RZNEW, c)], rz = scal[SC(S_RZ, c)];
    scal[SC(S_BETA, c)] = rz > 0.0 ? rzn / rz : 0.0;
    scal[SC(S_RZ, c)] = rzn;
    pre = pre && (rzn <= scal[SC(S_THR, c)]);
    if (C <= EU) euc = euc && (scal[SC(S_RZNEW, EU + c)] <= scal[SC(S_THR, EU + c)]);
    if (!(rzn == rzn)) flags[2] = 1;  // NaN: breakdown
  }
  flags[1] += 1;
  if (pre && flags[3] == 0) flags[3] = flags[1];
  const int waited = flags[3] > 0 ? flags[1] - flags[3] : 0;
  const int patience = flags[3] / 4 > 8 ? flags[3] / 4 : 8;
  if ((pre && (euc || waited >= patience)) || flags[2]) flags[0] = 1;
}
__global__ void k_post_update_soa(int C, double* __restrict__ scal, int* __restrict__ flags) {
  if (flags[0]) return;
  post_update_soa(C, scal, flags);
}

// x += alpha p ; r -= alpha q ; JACOBI: partial r.D^-1.r (and, single GPU, beta / convergence)
template <int C, bool JACOBI, bool LOCAL_POST>
__global__ void __launch_bounds__(MV_BLOCK)
k_update_soa(int64_t n, int64_t ld, const double* __restrict__ p, const double* __restrict__ q, const double* __restrict__ dinv,
             const uint16_t* __restrict__ mask, double* __restrict__ x, double* __restrict__ r, int* __restrict__ flags,
             double* __restrict__ partials, unsigned* __restrict__ counter, double* __restrict__ scal) {
  if (flags[0]) return;
  double alpha[C], acc[C];
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const double pq = scal[SC(S_PQ, c)], rz = scal[SC(S_RZ, c)];
    alpha[c] = pq > 0.0 ? rz / pq : 0.0;
    acc[c] = 0.0;
  }
  VERTEX_LOOP(v, n) {
    const double di = JACOBI ? __ldg(&dinv[v]) : 0.0;
    const unsigned mk = mask ? __ldg(&mask[v]) : 0u;
    // all loads first (the compiler cannot move a load of x/r above an earlier store to x/r, which
    // serialised the C components: 2.7 TB/s at C = 9), then the arithmetic, then all stores
    double xv[C], rv[C], pv[C], qv[C];
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int64_t i = c * ld + v;
      pv[c] = p[i];
      qv[c] = __ldcs(&q[i]);
      xv[c] = x[i];
      rv[c] = r[i];
    }
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const bool fix = (mk >> c) & 1u;
      xv[c] += alpha[c] * pv[c];
      rv[c] = fix ? 0.0 : rv[c] - alpha[c] * qv[c];
      if (JACOBI) acc[c] = fma(rv[c] * di, rv[c], acc[c]);
    }
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int64_t i = c * ld + v;
      x[i] = xv[c];
      r[i] = rv[c];
    }
  }
  if (JACOBI) {
    grid_reduce_sum<C>(acc, partials, counter, [&](double(&t)[C]) {
#pragma unroll
      for (int c = 0; c < C; ++c) scal[SC(S_RZNEW, c)] = t[c];
      if (LOCAL_POST) post_update_soa(C, scal, flags);
    });
  }
}

// p = D^-1 r + beta p
template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_dir_soa(int64_t n, int64_t ld, const double* __restrict__ r, const double* __restrict__ dinv, double* __restrict__ p,
          const int* __restrict__ flags, const double* __restrict__ scal) {
  if (flags[0]) return;
  double beta[C];
#pragma unroll
  for (int c = 0; c < C; ++c) beta[c] = scal[SC(S_BETA, c)];
  VERTEX_LOOP(v, n) {
    const double di = __ldg(&dinv[v]);
    double pv[C], rv[C];
#pragma unroll
    for (int c = 0; c < C; ++c) { pv[c] = p[c * ld + v]; rv[c] = __ldcs(&r[c * ld + v]); }
#pragma unroll
    for (int c = 0; c < C; ++c) p[c * ld + v] = fma(beta[c], pv[c], di * rv[c]);   // r is zero at constrained dofs
  }
}

// multigrid variant: r.z -> slot (POST: beta / convergence / iteration count)
template <int C, int POST>
__global__ void __launch_bounds__(MV_BLOCK)
k_dot_rz_soa(int64_t n, int64_t ld, const double* __restrict__ r, const double* __restrict__ z, int* __restrict__ flags,
             double* __restrict__ partials, unsigned* __restrict__ counter, double* __restrict__ scal, int slot, int slot2) {
  if (POST && flags[0]) return;
  constexpr bool EUCL = POST != 0 && C <= 3;   // the iteration's r.z carries the Euclidean r.r along (same pass over r)
  constexpr int NA = EUCL ? 2 * C : C;
  double acc[NA];
#pragma unroll
  for (int c = 0; c < NA; ++c) acc[c] = 0.0;
  VERTEX_LOOP(v, n)
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const double rv = r[c * ld + v];
      acc[c] = fma(rv, z[c * ld + v], acc[c]);
      if (EUCL) acc[C + c] = fma(rv, rv, acc[C + c]);
    }
  grid_reduce_sum<NA>(acc, partials, counter, [&](double(&t)[NA]) {
#pragma unroll
    for (int c = 0; c < C; ++c) {
      scal[SC(slot, c)] = t[c];
      if (slot2 >= 0) scal[SC(slot2, c)] = t[c];
      if (EUCL) scal[SC(slot, EU + c)] = t[C + c];
    }
    if (POST == 1) post_update_soa(C, scal, flags);
  });
}
// p = z + beta p  (first: p = z)
template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_dir_z_soa(int64_t n, int64_t ld, const double* __restrict__ z, double* __restrict__ p, const int* __restrict__ flags,
            const double* __restrict__ scal, int first) {
  if (!first && flags[0]) return;
  double beta[C];
#pragma unroll
  for (int c = 0; c < C; ++c) beta[c] = first ? 0.0 : scal[SC(S_BETA, c)];
  VERTEX_LOOP(v, n) {
    double pv[C], zv[C];
#pragma unroll
    for (int c = 0; c < C; ++c) { pv[c] = first ? 0.0 : p[c * ld + v]; zv[c] = z[c * ld + v]; }
#pragma unroll
    for (int c = 0; c < C; ++c) p[c * ld + v] = fma(beta[c], pv[c], zv[c]);
  }
}

// audit: sums |P(b - mean - q)|^2 and |P(b - mean)|^2 per component, q = A x component-major, b interleaved
template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_true_res_soa(int64_t n, int64_t ld, const double* __restrict__ b, const double* __restrict__ q, const uint16_t* __restrict__ mask,
               double* __restrict__ partials, unsigned* __restrict__ counter, double* __restrict__ scal, double inv_nglobal,
               int project_mean) {
  double acc[2 * C];
#pragma unroll
  for (int c = 0; c < 2 * C; ++c) acc[c] = 0.0;
  double mean[C];
#pragma unroll
  for (int c = 0; c < C; ++c) mean[c] = project_mean ? scal[SC(S_SUM, c)] * inv_nglobal : 0.0;
  VERTEX_LOOP(v, n) {
#pragma unroll
    for (int c = 0; c < C; ++c) {
      if (fixed_dof<C>(mask, v, c)) continue;
      const double bb = b ? (b[v * C + c] - mean[c]) : 0.0;
      const double rr = bb - q[c * ld + v];
      acc[c] = fma(rr, rr, acc[c]);
      acc[C + c] = fma(bb, bb, acc[C + c]);
    }
  }
  grid_reduce_sum<2 * C>(acc, partials, counter, [&](double(&t)[2 * C]) {
#pragma unroll
    for (int c = 0; c < C; ++c) {
      scal[SC(S_RZNEW, c)] = t[c];
      scal[SC(S_PQ, c)] = t[C + c];
    }
  });
}

template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_comp_sum_aos(int64_t n, const double* __restrict__ x, double* __restrict__ partials, unsigned* __restrict__ counter,
               double* __restrict__ scal) {
  double acc[C];
#pragma unroll
  for (int c = 0; c < C; ++c) acc[c] = 0.0;
  VERTEX_LOOP(v, n)
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] += x[v * C + c];
  grid_reduce_sum<C>(acc, partials, counter, [&](double(&t)[C]) {
#pragma unroll
    for (int c = 0; c < C; ++c) scal[SC(S_SUM, c)] = t[c];
  });
}
template <int C>
__global__ void __launch_bounds__(MV_BLOCK)
k_sub_mean_aos(int64_t n, double* __restrict__ x, const double* __restrict__ scal, double inv_nglobal) {
  const int64_t ne = n * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < ne; i += (int64_t)gridDim.x * blockDim.x)
    x[i] -= scal[SC(S_SUM, (int)(i % C))] * inv_nglobal;
}

}  // namespace mv

namespace mv {

// layout conversion for other translation units (mv_cheb.cu)
int aos_to_soa(int C, int64_t n, const double* src, double* dst, int64_t ld) {
  const int g = grid_for(n);
  switch (C) {
    case 1: MV_CUDA(cudaMemcpyAsync(dst, src, n * sizeof(double), cudaMemcpyDeviceToDevice, g_stream)); break;
    case 2: MV_L k_aos_to_soa<2><<<g, MV_BLOCK, 0, g_stream>>>(n, src, dst, ld); break;
    case 3: MV_L k_aos_to_soa<3><<<g, MV_BLOCK, 0, g_stream>>>(n, src, dst, ld); break;
    case 4: MV_L k_aos_to_soa<4><<<g, MV_BLOCK, 0, g_stream>>>(n, src, dst, ld); break;
    case 9: MV_L k_aos_to_soa<9><<<g, MV_BLOCK, 0, g_stream>>>(n, src, dst, ld); break;
    default: set_error("aos_to_soa: unsupported ncomp %d", C); return MV_ERR_UNSUPPORTED;
  }
  MV_KERNEL_CHECK();
  return MV_OK;
}
int soa_to_aos(int C, int64_t n, const double* src, int64_t ld, double* dst) {
  const int g = grid_for(n);
  switch (C) {
    case 1: MV_CUDA(cudaMemcpyAsync(dst, src, n * sizeof(double), cudaMemcpyDeviceToDevice, g_stream)); break;
    case 2: MV_L k_soa_to_aos<2><<<g, MV_BLOCK, 0, g_stream>>>(n, src, ld, dst); break;
    case 3: MV_L k_soa_to_aos<3><<<g, MV_BLOCK, 0, g_stream>>>(n, src, ld, dst); break;
    case 4: MV_L k_soa_to_aos<4><<<g, MV_BLOCK, 0, g_stream>>>(n, src, ld, dst); break;
    case 9: MV_L k_soa_to_aos<9><<<g, MV_BLOCK, 0, g_stream>>>(n, src, ld, dst); break;
    default: set_error("soa_to_aos: unsupported ncomp %d", C); return MV_ERR_UNSUPPORTED;
  }
  MV_KERNEL_CHECK();
  return MV_OK;
}

template <int C>
static int spmv_soa_launch(mv_mesh* m, const double* val, const double* x, double* y, int64_t ld, bool dot) {
  const Sell& p = m->pat;
  const int grid = grid_for(p.nslices * MV_SLICE, 8);
  static int variant = -1;
  if (variant < 0) { const char* e = getenv("MV_SPMV_VARIANT"); variant = e ? atoi(e) : 0; }
#define SPMV_LAUNCH(D, V)                                                                                                   \
  MV_L k_spmv_soa<C, D, V><<<grid, MV_BLOCK, 0, g_stream>>>(p.nslices, p.nrows, p.slice_ptr, p.cols, val, x, ld, y,          \
                                                       D ? m->pcg.flags : nullptr, D ? m->red.partials : nullptr,        \
                                                       D ? m->red.counter : nullptr, D ? m->pcg.scal : nullptr, S_PQ)
  if (dot) { if (variant == 1) SPMV_LAUNCH(true, 1); else SPMV_LAUNCH(true, 0); }
  else { if (variant == 1) SPMV_LAUNCH(false, 1); else SPMV_LAUNCH(false, 0); }
#undef SPMV_LAUNCH
  MV_KERNEL_CHECK();
  return MV_OK;
}

template <int C>
int pcg_soa(mv_mesh* m, const double* val, const double* dinv, const double* rhs, double* x, const uint16_t* mask, double rtol,
            double atol, int maxit, int project_mean, int* iters, double* relres, Amg* pc) {
  MV_CHECK(ensure_pcg_workspace(m));
  PcgWork& w = m->pcg;
  const int64_t n = m->Vown, ld = m->V;
  const bool dist = comm_active();
  int64_t nglobal = n;
  if (dist) {
    double hn = (double)n;
    MV_CUDA(cudaMemcpyAsync(w.scal + SC(S_SUM, 15), &hn, sizeof(double), cudaMemcpyHostToDevice, g_stream));
    MV_CHECK(allreduce_sum_dev(w.scal + SC(S_SUM, 15), 1));
    MV_CUDA(cudaMemcpyAsync(&hn, w.scal + SC(S_SUM, 15), sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    nglobal = (int64_t)(hn + 0.5);
  }
  const double inv_ng = 1.0 / (double)nglobal;
  const int gv = grid_for(n);
  if (project_mean && rhs) {
    MV_L k_comp_sum_aos<C><<<gv, MV_BLOCK, 0, g_stream>>>(n, rhs, m->red.partials, m->red.counter, w.scal);
    MV_KERNEL_CHECK();
    if (dist) MV_CHECK(allreduce_sum_dev(w.scal + SC(S_SUM, 0), C));
  }
  // initial guess -> component-major, initial residual
  double* xs = C == 1 ? x : w.xs;   // one component: the two layouts coincide
  if (C > 1) MV_L k_aos_to_soa<C><<<grid_for(n), MV_BLOCK, 0, g_stream>>>(n, x, xs, ld);
  if (dist) MV_CHECK(halo_exchange_soa(m->halo, n, C, xs, ld));
  MV_CHECK(spmv_soa_launch<C>(m, val, xs, w.q, ld, false));
  constexpr bool CAN_EUCL = C <= 3;
  const bool eucl = CAN_EUCL && pc != nullptr;
  const int nred = eucl ? EU + C : C;   // r.z in [0, C) and r.r in [EU, EU + C) of one scalar row: one allreduce
  if (eucl) {
    MV_CUDA(cudaMemsetAsync(w.scal + SC(S_RZ, 0), 0, 16 * sizeof(double), g_stream));
    MV_CUDA(cudaMemsetAsync(w.scal + SC(S_BB, 0), 0, 16 * sizeof(double), g_stream));
    MV_CUDA(cudaMemsetAsync(w.scal + SC(S_RZNEW, 0), 0, 16 * sizeof(double), g_stream));
    MV_L k_init_soa<C, CAN_EUCL><<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, rhs, w.q, dinv, mask, w.r, w.p, m->red.partials, m->red.counter,
                                                           w.scal, inv_ng, project_mean && rhs ? 1 : 0);
  } else {
    MV_L k_init_soa<C, false><<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, rhs, w.q, dinv, mask, w.r, w.p, m->red.partials, m->red.counter, w.scal,
                                                        inv_ng, project_mean && rhs ? 1 : 0);
  }
  MV_KERNEL_CHECK();
  if (dist) {
    MV_CHECK(allreduce_sum_dev(w.scal + SC(S_RZ, 0), nred));
    MV_CHECK(allreduce_sum_dev(w.scal + SC(S_BB, 0), nred));
  }
  if (pc) {
    // z0 = M^-1 r0 by one V-cycle; reference norm r0.z0
    MV_CHECK(amg_ensure_work(pc, C));
    MV_L k_ref_ratio<<<1, 1, 0, g_stream>>>(C, 0, w.scal);
    MV_CHECK(amg_apply(m, pc, C, w.r, pc->z, nullptr));
    MV_L k_dot_rz_soa<C, 0><<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, w.r, pc->z, w.flags, m->red.partials, m->red.counter, w.scal, S_RZ, S_BB);
    if (dist) {
      MV_CHECK(allreduce_sum_dev(w.scal + SC(S_RZ, 0), C));
      MV_CHECK(allreduce_sum_dev(w.scal + SC(S_BB, 0), C));
    }
    MV_L k_ref_ratio<<<1, 1, 0, g_stream>>>(C, 1, w.scal);
    MV_L k_dir_z_soa<C><<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, pc->z, w.p, w.flags, w.scal, 1);
    MV_KERNEL_CHECK();
  }
  MV_L k_setup_soa<<<1, 1, 0, g_stream>>>(C, eucl ? 1 : 0, rtol, atol, w.scal, w.flags);
  MV_KERNEL_CHECK();

  int it = 0, batch = pc ? 4 : 8;
  w.h_flags[0] = 0;
  // One iteration as a sequence of launches; `timed` brackets the SpMV with the profiling events (not inside a graph).
  auto iteration = [&](bool timed) -> int {
      if (dist) MV_CHECK(halo_exchange_soa(m->halo, n, C, w.p, ld));
      const int pslot = timed ? prof_begin(C) : -1;
      MV_CHECK(spmv_soa_launch<C>(m, val, w.p, w.q, ld, true));
      prof_end(pslot);
      if (dist) MV_CHECK(allreduce_sum_dev(w.scal + SC(S_PQ, 0), C));
      if (pc) {
        MV_L k_update_soa<C, false, false><<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, w.p, w.q, dinv, mask, xs, w.r, w.flags, m->red.partials,
                                                                     m->red.counter, w.scal);
        MV_CHECK(amg_apply(m, pc, C, w.r, pc->z, w.flags));
        if (dist) {
          MV_L k_dot_rz_soa<C, 2><<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, w.r, pc->z, w.flags, m->red.partials, m->red.counter, w.scal,
                                                            S_RZNEW, -1);
          MV_CHECK(allreduce_sum_dev(w.scal + SC(S_RZNEW, 0), nred));
          MV_L k_post_update_soa<<<1, 1, 0, g_stream>>>(C, w.scal, w.flags);
        } else {
          MV_L k_dot_rz_soa<C, 1><<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, w.r, pc->z, w.flags, m->red.partials, m->red.counter, w.scal,
                                                            S_RZNEW, -1);
        }
        MV_L k_dir_z_soa<C><<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, pc->z, w.p, w.flags, w.scal, 0);
      } else {
        if (dist) {
          MV_L k_update_soa<C, true, false><<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, w.p, w.q, dinv, mask, xs, w.r, w.flags, m->red.partials,
                                                                      m->red.counter, w.scal);
          MV_CHECK(allreduce_sum_dev(w.scal + SC(S_RZNEW, 0), C));
          MV_L k_post_update_soa<<<1, 1, 0, g_stream>>>(C, w.scal, w.flags);
        } else {
          MV_L k_update_soa<C, true, true><<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, w.p, w.q, dinv, mask, xs, w.r, w.flags, m->red.partials,
                                                                     m->red.counter, w.scal);
        }
        MV_L k_dir_soa<C><<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, w.r, dinv, w.p, w.flags, w.scal);
      }
      return MV_OK;
  };
  // CUDA graph of the multigrid-preconditioned iteration: ~25 launches at one GPU, ~45 with the peer-memory exchanges on
  // several, most of them a few microseconds long on the coarse levels -- replayed as one graph launch.  All arguments
  // are device pointers / device-resident scalars that do not change between iterations or solves, so the graph is
  // captured once per (operator, hierarchy, buffers).  The first batch of a solve stays un-graphed while the profiling
  // events are on (bench.py's per-launch SpMV timing).  MV_PCG_GRAPH=0 switches it off.
  static int graph_env = -1;
  if (graph_env < 0) { const char* e = getenv("MV_PCG_GRAPH"); graph_env = e ? atoi(e) : 1; }
  const bool want_graph = pc != nullptr && graph_env != 0 && !w.graph_failed;
  unsigned long long key[12] = {(unsigned long long)C, pc ? pc->uid : 0ull, (unsigned long long)(uintptr_t)val, (unsigned long long)(uintptr_t)dinv,
                                (unsigned long long)(uintptr_t)mask, (unsigned long long)(uintptr_t)xs, (unsigned long long)n,
                                (unsigned long long)ld, dist ? 1ull : 0ull, (unsigned long long)(uintptr_t)w.r,
                                (unsigned long long)(uintptr_t)(pc ? pc->z : nullptr), (unsigned long long)(uintptr_t)g_stream};
  auto ensure_graph = [&]() -> bool {
    if (w.graph_exec && memcmp(key, w.graph_key, sizeof(key)) == 0) return true;
    pcg_graph_release(w);
    const long long l0 = g_launches;
    // the library's stream is normally torch's current stream, i.e. the legacy default stream, which cannot be captured:
    // capture on a private stream (every launch site uses the global g_stream), replay on the caller's
    static cudaStream_t cap_stream = nullptr;
    if (!cap_stream && cudaStreamCreateWithFlags(&cap_stream, cudaStreamNonBlocking) != cudaSuccess) {
      cudaGetLastError(); w.graph_failed = true; return false;
    }
    cudaStream_t user_stream = g_stream;
    g_stream = cap_stream;
    cudaError_t cb = cudaStreamBeginCapture(g_stream, cudaStreamCaptureModeThreadLocal);
    if (cb != cudaSuccess) {
      g_stream = user_stream;
      if (getenv("MV_PCG_VERBOSE")) fprintf(stderr, "[mv pcg] graph capture not started: %s\n", cudaGetErrorString(cb));
      cudaGetLastError(); w.graph_failed = true; return false;
    }
    const int st = iteration(false);
    cudaGraph_t graph = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(g_stream, &graph);
    g_stream = user_stream;
    w.graph_launches = g_launches - l0;
    g_launches = l0;                       // nothing ran yet
    cudaGraphExec_t exec = nullptr;
    cudaError_t ci = cudaSuccess;
    if (st != MV_OK || ce != cudaSuccess || !graph || (ci = cudaGraphInstantiate(&exec, graph, 0)) != cudaSuccess) {
      if (getenv("MV_PCG_VERBOSE"))
        fprintf(stderr, "[mv pcg] graph capture failed: iteration status %d, end capture '%s', instantiate '%s'\n", st, cudaGetErrorString(ce),
                cudaGetErrorString(ci));
      cudaGetLastError();
      if (graph) cudaGraphDestroy(graph);
      w.graph_failed = true;
      return false;
    }
    if (getenv("MV_PCG_VERBOSE")) fprintf(stderr, "[mv pcg] iteration graph captured: %lld kernel launches\n", w.graph_launches);
    cudaGraphDestroy(graph);
    w.graph_exec = exec;
    g_graph_owners.push_back(&w);
    memcpy(w.graph_key, key, sizeof(key));
    return true;
  };
  while (true) {
    MV_CUDA(cudaMemcpyAsync(w.h_flags, w.flags, 4 * sizeof(int), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    if (g_prof.used) prof_drain();
    if (w.h_flags[0] || it >= maxit) break;
    const int nb = batch < maxit - it ? batch : maxit - it;
    const bool graphed = want_graph && !(g_prof.on && it == 0) && ensure_graph();
    for (int k = 0; k < nb; ++k) {
      if (graphed) {
        MV_CUDA(cudaGraphLaunch((cudaGraphExec_t)w.graph_exec, g_stream));
        g_launches += w.graph_launches;
      } else {
        MV_CHECK(iteration(true));
      }
    }
    MV_KERNEL_CHECK();
    it += nb;
    if (batch < (pc ? 16 : 64)) batch *= 2;
  }
  const int done = w.h_flags[0], niter = w.h_flags[1], breakdown = w.h_flags[2];
  g_prof.pcg_iters[C & 15] += niter;
  static int verbose = -1;
  if (verbose < 0) { const char* e = getenv("MV_PCG_VERBOSE"); verbose = e ? atoi(e) : 0; }
  if (verbose) {
    MV_CUDA(cudaMemcpyAsync(w.h_scal, w.scal, S_NSLOTS * 16 * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    fprintf(stderr, "[mv pcg] C %d n %lld %s: %d iterations (preconditioned test met at %d), done %d;", C, (long long)n, pc ? "amg" : "jacobi", niter,
            w.h_flags[3], done);
    for (int c = 0; c < C; ++c)
      fprintf(stderr, "  rz/thr %.3e rr/thr2 %.3e", w.h_scal[SC(S_RZ, c)] / w.h_scal[SC(S_THR, c)],
              eucl ? w.h_scal[SC(S_RZNEW, EU + c)] / w.h_scal[SC(S_THR, EU + c)] : 0.0);
    fprintf(stderr, "\n");
  }
  if (g_trueres.on) {
    // audit of the returned iterate: q = A x, then ||P(b - q)|| / ||P b|| (Euclidean) per right-hand side
    if (project_mean && rhs) {
      MV_L k_comp_sum_aos<C><<<gv, MV_BLOCK, 0, g_stream>>>(n, rhs, m->red.partials, m->red.counter, w.scal);
      if (dist) MV_CHECK(allreduce_sum_dev(w.scal + SC(S_SUM, 0), C));
    }
    if (dist) MV_CHECK(halo_exchange_soa(m->halo, n, C, xs, ld));
    MV_CHECK(spmv_soa_launch<C>(m, val, xs, w.q, ld, false));
    MV_L k_true_res_soa<C><<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, rhs, w.q, mask, m->red.partials, m->red.counter, w.scal, inv_ng,
                                                     project_mean && rhs ? 1 : 0);
    MV_KERNEL_CHECK();
    if (dist) {
      MV_CHECK(allreduce_sum_dev(w.scal + SC(S_RZNEW, 0), C));
      MV_CHECK(allreduce_sum_dev(w.scal + SC(S_PQ, 0), C));
    }
    MV_CUDA(cudaMemcpyAsync(w.h_scal, w.scal, S_NSLOTS * 16 * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    double worst = 0.0;
    for (int c = 0; c < C; ++c) {
      const double num = w.h_scal[SC(S_RZNEW, c)], den = w.h_scal[SC(S_PQ, c)];
      const double rel = den > 0.0 ? sqrt(num / den) : 0.0;
      if (rel > worst) worst = rel;
    }
    TrueRes& t = g_trueres;
    t.last_rel[C & 15] = worst;
    if (worst > t.max_rel[C & 15]) t.max_rel[C & 15] = worst;
    t.count[C & 15] += 1;
  }
  if (C > 1) MV_L k_soa_to_aos<C><<<grid_for(n), MV_BLOCK, 0, g_stream>>>(n, xs, ld, x);
  if (project_mean) {
    MV_L k_comp_sum_aos<C><<<gv, MV_BLOCK, 0, g_stream>>>(n, x, m->red.partials, m->red.counter, w.scal);
    if (dist) MV_CHECK(allreduce_sum_dev(w.scal + SC(S_SUM, 0), C));
    MV_L k_sub_mean_aos<C><<<grid_for(n * C), MV_BLOCK, 0, g_stream>>>(n, x, w.scal, inv_ng);
  }
  MV_KERNEL_CHECK();
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

int pcg_solve_soa(mv_mesh* m, int ncomp, const double* val, const double* dinv, const double* rhs, double* x, const uint16_t* mask,
                  double rtol, double atol, int maxit, int project_mean, int* iters, double* relres, Amg* pc) {
  switch (ncomp) {
    case 1: return pcg_soa<1>(m, val, dinv, rhs, x, mask, rtol, atol, maxit, project_mean, iters, relres, pc);
    case 2: return pcg_soa<2>(m, val, dinv, rhs, x, mask, rtol, atol, maxit, project_mean, iters, relres, pc);
    case 3: return pcg_soa<3>(m, val, dinv, rhs, x, mask, rtol, atol, maxit, project_mean, iters, relres, pc);
    case 4: return pcg_soa<4>(m, val, dinv, rhs, x, mask, rtol, atol, maxit, project_mean, iters, relres, pc);
    case 9: return pcg_soa<9>(m, val, dinv, rhs, x, mask, rtol, atol, maxit, project_mean, iters, relres, pc);
  }
  set_error("pcg: unsupported ncomp %d (1,2,3,4,9)", ncomp);
  return MV_ERR_UNSUPPORTED;
}

}  // namespace mv
