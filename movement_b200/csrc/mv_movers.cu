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


// ------------------------------------------------------------------ assembled spring operator + multigrid preconditioner
// The matrix-free operator above recomputes every edge block (sqrt, acos, cos, sin) on each application and its
// Jacobi-preconditioned CG needs O(n) iterations (4096^2: thousands per solve).  For large meshes the three distinct
// blocks are assembled once per call on the mesh pattern,
//     K = [ Kxx  Kxy ]      Kxx_ij = -c^2/L, Kxy_ij = -sc/L, Kyy_ij = -s^2/L  (i != j),  diagonal = -(row sum),
//         [ Kxy  Kyy ]
// and the solve is preconditioned by diag(V-cycle(Kxx), V-cycle(Kyy)): each diagonal block is a weighted graph Laplacian
// (an M-matrix, what smoothed aggregation is built for) and K <= 2 diag(Kxx, Kyy) edge by edge (v v^T <= 2 diag(v^2)), so
// the preconditioned operator has a mesh-independent condition number (exact block solves: 25 PCG iterations at
// 32^2, 64^2, 128^2 against 226 / 450 / 880 with Jacobi).  Dirichlet rows are row-replaced as in
// movement/spring.py:225-234: constrained dofs keep their value, the masks of the two hierarchies are the two bits.
__global__ void __launch_bounds__(MV_BLOCK)
k_spring_assemble(int64_t nrows, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols, const int* __restrict__ opp,
                  const int* __restrict__ row_nnz, const double* __restrict__ X, double* __restrict__ vxx, double* __restrict__ vxy,
                  double* __restrict__ vyy) {
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = v / MV_SLICE;
    const int64_t base = slice_ptr[s] + (v % MV_SLICE);
    const int w = (int)((slice_ptr[s + 1] - slice_ptr[s]) / MV_SLICE);
    const int nn = row_nnz[v];
    double d0 = 0.0, d1 = 0.0, d2 = 0.0;
    int64_t dpos = -1;
    for (int k = 0; k < w; ++k) {
      const int64_t pos = base + (int64_t)k * MV_SLICE;
      const int o = opp[pos];
      double a0 = 0.0, a1 = 0.0, a2 = 0.0;
      if (o >= 0) {
        const EdgeBlock e = edge_block(X, (int)v, cols[pos], o);
        a0 = -e.c2; a1 = -e.sc; a2 = -e.s2;
        d0 += e.c2; d1 += e.sc; d2 += e.s2;
      } else if (k < nn && cols[pos] == v && dpos < 0) {
        dpos = pos;
      }
      vxx[pos] = a0; vxy[pos] = a1; vyy[pos] = a2;
    }
    if (dpos >= 0) { vxx[dpos] = d0; vxy[dpos] = d1; vyy[dpos] = d2; }
  }
}

// q = K p on component-major vectors [2][ld], rows of constrained dofs give 0 (DOT: inside PCG, p is zero there) or the
// dof itself (row replacement); DOT: p.q summed over both components into slot S_PQ
template <bool DOT>
__global__ void __launch_bounds__(MV_BLOCK)
k_spring_spmv(int64_t nslices, int64_t nrows, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
              const double* __restrict__ vxx, const double* __restrict__ vxy, const double* __restrict__ vyy,
              const uint16_t* __restrict__ mask, const double* __restrict__ x, int64_t ld, double* __restrict__ y,
              const int* __restrict__ done, double* __restrict__ partials, unsigned* __restrict__ counter, double* __restrict__ scal) {
  if (DOT && done != nullptr && *done) return;
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  double dot[1] = {0.0};
  for (int64_t s = warp0; s < nslices; s += nwarps) {
    const int64_t base = __ldg(&slice_ptr[s]);
    const int w = (int)((__ldg(&slice_ptr[s + 1]) - base) >> 5);
    const int64_t row = s * MV_SLICE + lane;
    double a0 = 0.0, a1 = 0.0;
    constexpr int KB = 4;
    for (int k0 = 0; k0 < w; k0 += KB) {
      int col[KB];
      double axx[KB], axy[KB], ayy[KB];
#pragma unroll
      for (int j = 0; j < KB; ++j) {
        const bool in = k0 + j < w;
        const int64_t pos = base + lane + (int64_t)(k0 + j) * MV_SLICE;
        col[j] = in ? __ldcs(&cols[pos]) : 0;
        axx[j] = in ? __ldcs(&vxx[pos]) : 0.0;
        axy[j] = in ? __ldcs(&vxy[pos]) : 0.0;
        ayy[j] = in ? __ldcs(&vyy[pos]) : 0.0;
      }
#pragma unroll
      for (int j = 0; j < KB; ++j) {
        const double px = __ldg(&x[col[j]]), py = __ldg(&x[ld + col[j]]);
        a0 = fma(axx[j], px, fma(axy[j], py, a0));
        a1 = fma(axy[j], px, fma(ayy[j], py, a1));
      }
    }
    if (row < nrows) {
      const unsigned mk = mask ? __ldg(&mask[row]) : 0u;
      const double x0 = __ldg(&x[row]), x1 = __ldg(&x[ld + row]);
      if (mk & 1u) a0 = DOT ? 0.0 : x0;
      if (mk & 2u) a1 = DOT ? 0.0 : x1;
      y[row] = a0;
      y[ld + row] = a1;
      if (DOT) dot[0] = fma(a0, x0, fma(a1, x1, dot[0]));
    }
  }
  if (DOT) grid_reduce_sum<1>(dot, partials, counter, [&](double(&t)[1]) { scal[SC(S_PQ, 0)] = t[0]; });
}

// level-0 masks of the two hierarchies + number of entries that differ from the stored ones
__global__ void k_spring_masks(int64_t n, const uint16_t* __restrict__ mask, uint16_t* __restrict__ m0, uint16_t* __restrict__ m1,
                               int init, int* __restrict__ ndiff_ncon) {
  int diff = 0, c0 = 0, c1 = 0;
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < n; v += (int64_t)gridDim.x * blockDim.x) {
    const uint16_t b0 = mask[v] & 1u, b1 = (mask[v] >> 1) & 1u;
    if (!init && (m0[v] != b0 || m1[v] != b1)) ++diff;
    m0[v] = b0; m1[v] = b1;
    c0 += b0; c1 += b1;
  }
  if (diff) atomicAdd(&ndiff_ncon[0], diff);
  if (c0) atomicAdd(&ndiff_ncon[1], c0);
  if (c1) atomicAdd(&ndiff_ncon[2], c1);
}

// r = P(b - q) (b interleaved [n,2], q, r component-major), Jacobi and Euclidean norms of r and of the free part of b
__global__ void __launch_bounds__(MV_BLOCK)
k_sp_init(int64_t n, int64_t ld, const double* __restrict__ b, const double* __restrict__ q, const double* __restrict__ dinv_dof,
          const uint16_t* __restrict__ mask, double* __restrict__ r, double* __restrict__ partials, unsigned* __restrict__ counter,
          double* __restrict__ scal) {
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < n; v += (int64_t)gridDim.x * blockDim.x) {
    const unsigned mk = mask ? mask[v] : 0u;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const bool fix = (mk >> c) & 1u;
      const double bb = b[2 * v + c], di = dinv_dof[2 * v + c];
      const double rr = fix ? 0.0 : bb - q[c * ld + v];
      r[c * ld + v] = rr;
      acc[0] = fma(rr * di, rr, acc[0]);
      acc[2] = fma(rr, rr, acc[2]);
      if (!fix) { acc[1] = fma(bb * di, bb, acc[1]); acc[3] = fma(bb, bb, acc[3]); }
    }
  }
  grid_reduce_sum<4>(acc, partials, counter, [&](double(&t)[4]) {
    scal[SC(S_SUM, 0)] = t[0]; scal[SC(S_SUM, 1)] = t[1]; scal[SC(S_SUM, 2)] = t[2]; scal[SC(S_SUM, 3)] = t[3];
  });
}
// r.z and r.r over both components -> scal[slot][0], scal[slot][1]
__global__ void __launch_bounds__(MV_BLOCK)
k_sp_dot(int64_t n, int64_t ld, const double* __restrict__ r, const double* __restrict__ zx, const double* __restrict__ zy,
         const int* __restrict__ flags, int check_done, double* __restrict__ partials, unsigned* __restrict__ counter,
         double* __restrict__ scal, int slot) {
  if (check_done && flags[0]) return;
  double acc[2] = {0.0, 0.0};
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < n; v += (int64_t)gridDim.x * blockDim.x) {
    const double r0 = r[v], r1 = r[ld + v];
    acc[0] = fma(r0, zx[v], fma(r1, zy[v], acc[0]));
    acc[1] = fma(r0, r0, fma(r1, r1, acc[1]));
  }
  grid_reduce_sum<2>(acc, partials, counter, [&](double(&t)[2]) { scal[SC(slot, 0)] = t[0]; scal[SC(slot, 1)] = t[1]; });
}
// thresholds from the initial residual: stopping reference tied to max(b, r0) through the Jacobi-norm ratio (as the
// multigrid PCG of mv_pcg_soa.cu), Euclidean twin of the test
__global__ void k_sp_setup(double rtol, double atol, double* __restrict__ scal, int* __restrict__ flags) {
  const double rz = scal[SC(S_RZ, 0)], rr = scal[SC(S_RZ, 1)];
  const double rj = scal[SC(S_SUM, 0)], bj = scal[SC(S_SUM, 1)], b2 = scal[SC(S_SUM, 3)];
  const double ratio = (rj > 0.0 && bj > rj) ? bj / rj : 1.0;
  const double ref = rz * ratio;
  scal[SC(S_REF, 0)] = ref;
  scal[SC(S_THR, 0)] = fmax(rtol * rtol * ref, atol * atol);
  scal[SC(S_THR, 1)] = fmax(rtol * rtol * fmax(rr, b2), atol * atol);
  flags[0] = (rz <= scal[SC(S_THR, 0)] && rr <= scal[SC(S_THR, 1)]) ? 1 : 0;
  flags[1] = 0; flags[2] = 0; flags[3] = 0;
}
// after the new r.z / r.r are known (slot S_RZNEW): beta, convergence (preconditioned test, then the Euclidean twin with
// bounded patience, as post_update_soa), iteration count
__global__ void k_sp_post(double* __restrict__ scal, int* __restrict__ flags) {
  if (flags[0]) return;
  const double rzn = scal[SC(S_RZNEW, 0)], rrn = scal[SC(S_RZNEW, 1)], rz = scal[SC(S_RZ, 0)];
  scal[SC(S_BETA, 0)] = rz > 0.0 ? rzn / rz : 0.0;
  scal[SC(S_RZ, 0)] = rzn;
  scal[SC(S_RZ, 1)] = rrn;
  if (!(rzn == rzn) || rzn < 0.0) flags[2] = 1;   // NaN, or a preconditioner that is not positive definite
  flags[1] += 1;
  const bool pre = rzn <= scal[SC(S_THR, 0)], euc = rrn <= scal[SC(S_THR, 1)];
  if (pre && flags[3] == 0) flags[3] = flags[1];
  const int waited = flags[3] > 0 ? flags[1] - flags[3] : 0;
  const int patience = flags[3] / 4 > 8 ? flags[3] / 4 : 8;
  if ((pre && (euc || waited >= patience)) || flags[2]) flags[0] = 1;
}
// x += alpha p ; r -= alpha q (constrained dofs stay zero)
__global__ void __launch_bounds__(MV_BLOCK)
k_sp_update(int64_t n, int64_t ld, const double* __restrict__ p, const double* __restrict__ q, const uint16_t* __restrict__ mask,
            double* __restrict__ x, double* __restrict__ r, int* __restrict__ flags, const double* __restrict__ scal) {
  if (flags[0]) return;
  const double pq = scal[SC(S_PQ, 0)], rz = scal[SC(S_RZ, 0)];
  const double alpha = pq > 0.0 ? rz / pq : 0.0;
  // p.K p <= 0 or NaN with r != 0: the operator is not positive definite (degenerate edge, NaN coordinates) -- a breakdown,
  // not a reason to iterate to maxit with alpha = 0
  if (!(pq > 0.0) && rz > 0.0 && blockIdx.x == 0 && threadIdx.x == 0) flags[2] = 1;
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < n; v += (int64_t)gridDim.x * blockDim.x) {
    const unsigned mk = mask ? mask[v] : 0u;
    const double p0 = p[v], p1 = p[ld + v], q0 = __ldcs(&q[v]), q1 = __ldcs(&q[ld + v]);
    const double x0 = x[v], x1 = x[ld + v], r0 = r[v], r1 = r[ld + v];
    x[v] = fma(alpha, p0, x0);
    x[ld + v] = fma(alpha, p1, x1);
    r[v] = (mk & 1u) ? 0.0 : fma(-alpha, q0, r0);
    r[ld + v] = (mk & 2u) ? 0.0 : fma(-alpha, q1, r1);
  }
}
// p = z + beta p
__global__ void __launch_bounds__(MV_BLOCK)
k_sp_dir(int64_t n, int64_t ld, const double* __restrict__ zx, const double* __restrict__ zy, double* __restrict__ p,
         const int* __restrict__ flags, const double* __restrict__ scal, int first) {
  if (!first && flags[0]) return;
  const double beta = first ? 0.0 : scal[SC(S_BETA, 0)];
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < n; v += (int64_t)gridDim.x * blockDim.x) {
    p[v] = first ? zx[v] : fma(beta, p[v], zx[v]);
    p[ld + v] = first ? zy[v] : fma(beta, p[ld + v], zy[v]);
  }
}

void spring_pc_release(mv_mesh* m) {
  for (int c = 0; c < 2; ++c) {
    if (m->spring_pc[c]) { delete m->spring_pc[c]; m->spring_pc[c] = nullptr; }
    dev_free(m->spring_mask[c]);
  }
  for (int c = 0; c < 3; ++c) dev_free(m->spring_val[c]);
  m->spring_pc_its = m->spring_pc_age = 0;
}

static int g_spring_min_rows = -1;   // MV_SPRING_AMG_MIN: global vertex count from which the multigrid path is used
static int g_spring_last_setup = 0;  // 1: the last solve rebuilt the hierarchies

// Solve K(coords) u = rhs with row-replaced Dirichlet dofs (u holds their values on entry), interleaved [V,2] vectors.
static int spring_solve_amg(mv_mesh* m, const double* coords, const double* dinv_dof, const double* rhs, double* u, const uint16_t* mask,
                            double rtol, double atol, int maxit, int* iters) {
  MV_CHECK(ensure_pcg_workspace(m));
  PcgWork& w = m->pcg;
  const Sell& pt = m->pat;
  const int64_t n = m->Vown, ld = m->V;
  const bool dist = comm_active();
  const int gv = grid_for(n);
  const int gs = grid_for(pt.nslices * MV_SLICE, 8);
  // operator of this call
  const bool first = m->spring_val[0] == nullptr;
  if (first)
    for (int c = 0; c < 3; ++c) MV_CHECK(dev_alloc(&m->spring_val[c], pt.len));
  MV_L k_spring_assemble<<<gv, MV_BLOCK, 0, g_stream>>>(n, pt.slice_ptr, pt.cols, m->edge_opp, pt.row_nnz, coords, m->spring_val[0],
                                                   m->spring_val[1], m->spring_val[2]);
  MV_KERNEL_CHECK();
  // constrained sets: a change rebuilds the hierarchies, and so does a preconditioner that has aged (the mesh has moved
  // away from the configuration it was built on: 30 % more iterations than right after its set-up)
  int* cnt = nullptr;
  MV_CHECK(dev_alloc(&cnt, 4));
  MV_CUDA(cudaMemsetAsync(cnt, 0, 4 * sizeof(int), g_stream));
  const bool init = m->spring_mask[0] == nullptr;
  if (init)
    for (int c = 0; c < 2; ++c) MV_CHECK(dev_alloc(&m->spring_mask[c], n));
  MV_L k_spring_masks<<<gv, MV_BLOCK, 0, g_stream>>>(n, mask, m->spring_mask[0], m->spring_mask[1], init ? 1 : 0, cnt);
  MV_KERNEL_CHECK();
  int hc[4] = {0, 0, 0, 0};
  MV_CUDA(cudaMemcpyAsync(hc, cnt, 4 * sizeof(int), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  dev_free(cnt);
  double gl[3] = {(double)hc[0], (double)hc[1], (double)hc[2]};
  if (dist) MV_CHECK(allreduce_host(gl, 3, 0));
  int reuse = 1;
  { const char* e = getenv("MV_SPRING_PC_REUSE"); if (e) reuse = atoi(e); }
  const bool aged = m->spring_pc_age > 0 && m->spring_pc_its > 0 && m->spring_pc_age_its * 10 > m->spring_pc_its * 13 + 20;
  bool rebuild = !m->spring_pc[0] || !m->spring_pc[1] || gl[0] > 0 || aged || !reuse;
  g_spring_last_setup = rebuild ? 1 : 0;
  if (rebuild) {
    for (int c = 0; c < 2; ++c) {
      if (m->spring_pc[c]) { delete m->spring_pc[c]; m->spring_pc[c] = nullptr; }
      MV_CHECK(amg_build_values(m, 100 + c, m->spring_val[2 * c], m->spring_mask[c], gl[1 + c] > 0 ? 0 : 1, &m->spring_pc[c]));
    }
    m->spring_pc_age = 0;
    m->spring_pc_its = 0;
  }
  Amg *px = m->spring_pc[0], *py = m->spring_pc[1];
  MV_CHECK(amg_ensure_work(px, 1));
  MV_CHECK(amg_ensure_work(py, 1));
  // component-major iterate, initial residual
  double* xs = w.xs;
  MV_CHECK(aos_to_soa(2, n, u, xs, ld));
  if (dist) MV_CHECK(halo_exchange_soa(m->halo, n, 2, xs, ld));
#define SPRING_SPMV(D, X, Y)                                                                                                         \
  MV_L k_spring_spmv<D><<<gs, MV_BLOCK, 0, g_stream>>>(pt.nslices, pt.nrows, pt.slice_ptr, pt.cols, m->spring_val[0], m->spring_val[1],  \
                                                  m->spring_val[2], mask, X, ld, Y, D ? w.flags : nullptr, m->red.partials,          \
                                                  m->red.counter, w.scal)
  SPRING_SPMV(false, xs, w.q);
  MV_L k_sp_init<<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, rhs, w.q, dinv_dof, mask, w.r, m->red.partials, m->red.counter, w.scal);
  MV_KERNEL_CHECK();
  if (dist) MV_CHECK(allreduce_sum_dev(w.scal + SC(S_SUM, 0), 4));
  MV_CHECK(amg_apply(m, px, 1, w.r, px->z, nullptr));
  MV_CHECK(amg_apply(m, py, 1, w.r + ld, py->z, nullptr));
  MV_L k_sp_dot<<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, w.r, px->z, py->z, w.flags, 0, m->red.partials, m->red.counter, w.scal, S_RZ);
  if (dist) MV_CHECK(allreduce_sum_dev(w.scal + SC(S_RZ, 0), 2));
  MV_L k_sp_setup<<<1, 1, 0, g_stream>>>(rtol, atol, w.scal, w.flags);
  MV_L k_sp_dir<<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, px->z, py->z, w.p, w.flags, w.scal, 1);
  MV_KERNEL_CHECK();
  int it = 0, batch = 4;
  w.h_flags[0] = 0;
  if (maxit > 5000) maxit = 5000;   // mesh-independent convergence: thousands of multigrid-preconditioned iterations mean a broken operator
  while (true) {
    MV_CUDA(cudaMemcpyAsync(w.h_flags, w.flags, 4 * sizeof(int), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    if (g_prof.used) prof_drain();
    if (w.h_flags[0] || it >= maxit) break;
    const int nb = batch < maxit - it ? batch : maxit - it;
    for (int k = 0; k < nb; ++k) {
      if (dist) MV_CHECK(halo_exchange_soa(m->halo, n, 2, w.p, ld));
      const int pslot = prof_begin(2);
      SPRING_SPMV(true, w.p, w.q);
      prof_end(pslot);
      if (dist) MV_CHECK(allreduce_sum_dev(w.scal + SC(S_PQ, 0), 1));
      MV_L k_sp_update<<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, w.p, w.q, mask, xs, w.r, w.flags, w.scal);
      MV_CHECK(amg_apply(m, px, 1, w.r, px->z, w.flags));
      MV_CHECK(amg_apply(m, py, 1, w.r + ld, py->z, w.flags));
      MV_L k_sp_dot<<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, w.r, px->z, py->z, w.flags, 1, m->red.partials, m->red.counter, w.scal, S_RZNEW);
      if (dist) MV_CHECK(allreduce_sum_dev(w.scal + SC(S_RZNEW, 0), 2));
      MV_L k_sp_post<<<1, 1, 0, g_stream>>>(w.scal, w.flags);
      MV_L k_sp_dir<<<gv, MV_BLOCK, 0, g_stream>>>(n, ld, px->z, py->z, w.p, w.flags, w.scal, 0);
    }
    MV_KERNEL_CHECK();
    it += nb;
    if (batch < 16) batch *= 2;
  }
#undef SPRING_SPMV
  const int done = w.h_flags[0], niter = w.h_flags[1], breakdown = w.h_flags[2];
  g_prof.pcg_iters[2] += niter;
  if (getenv("MV_PCG_VERBOSE")) {
    MV_CUDA(cudaMemcpyAsync(w.h_scal, w.scal, S_NSLOTS * 16 * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    fprintf(stderr, "[mv spring] n %lld: %d iterations (preconditioned test met at %d), done %d, breakdown %d, hierarchies %s (age %d, levels %zu/%zu); "
            "rz/thr %.3e rr/thr2 %.3e ref %.3e pq %.3e\n", (long long)n, niter, w.h_flags[3], done, breakdown, rebuild ? "rebuilt" : "reused",
            m->spring_pc_age, px->lv.size(), py->lv.size(), w.h_scal[SC(S_RZ, 0)] / w.h_scal[SC(S_THR, 0)],
            w.h_scal[SC(S_RZ, 1)] / w.h_scal[SC(S_THR, 1)], w.h_scal[SC(S_REF, 0)], w.h_scal[SC(S_PQ, 0)]);
  }
  MV_CHECK(soa_to_aos(2, n, xs, ld, u));
  if (m->spring_pc_age == 0) m->spring_pc_its = niter;
  m->spring_pc_age += 1;
  m->spring_pc_age_its = niter;
  if (iters) *iters = niter;
  if (breakdown) { set_error("spring pcg: breakdown (NaN) after %d iterations", niter); return MV_ERR_NOT_CONVERGED; }
  if (!done) { set_error("spring pcg: not converged in %d iterations", maxit); return MV_ERR_NOT_CONVERGED; }
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
  // large meshes: assembled blocks + block-diagonal multigrid; small ones: matrix-free Jacobi-PCG
  { const char* e = getenv("MV_SPRING_AMG_MIN"); g_spring_min_rows = e ? atoi(e) : 4096; }   // read per call: tests switch it
  double vg = (double)n;
  if (comm_active()) MV_CHECK(allreduce_host(&vg, 1, 0));
  int st;
  if (m->pc_kind == MV_PC_AMG && vg >= (double)g_spring_min_rows) {
    st = spring_solve_amg(m, coords, dinv, sf.as<double>(), u, sm.as<uint16_t>(), rtol, atol, maxit, iters);
    if (st == MV_ERR_NOT_CONVERGED && !g_spring_last_setup) {
      // a REUSED preconditioner failed (breakdown or the iteration cap): the mesh may have moved too far from the
      // configuration it was built on -- rebuild the hierarchies on the current operator and solve once more from the
      // same initial guess (the status is the same on every rank: it derives from allreduced scalars)
      for (int c = 0; c < 2; ++c)
        if (m->spring_pc[c]) { delete m->spring_pc[c]; m->spring_pc[c] = nullptr; }
      MV_CUDA(cudaMemcpyAsync(u, sf.dev, m->V * 2 * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
      MV_L k_zero_free<2><<<grid_for(n * 2), MV_BLOCK, 0, g_stream>>>(n, sm.as<uint16_t>(), u);
      MV_KERNEL_CHECK();
      st = spring_solve_amg(m, coords, dinv, sf.as<double>(), u, sm.as<uint16_t>(), rtol, atol, maxit, iters);
      if (st == MV_OK) set_error("%s", "");   // the first attempt's message must not outlive a successful retry
    }
  } else {
    st = pcg_solve_spring(m, coords, dinv, sf.as<double>(), u, sm.as<uint16_t>(), rtol, atol, maxit, iters);
  }
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

extern "C" int mv_spring_solver_info(const mv_mesh* m, int64_t* info) {
  if (!m || !info) return MV_ERR_BADARG;
  info[0] = m->spring_pc[0] ? (int64_t)m->spring_pc[0]->lv.size() : 0;
  info[1] = m->spring_pc[1] ? (int64_t)m->spring_pc[1]->lv.size() : 0;
  info[2] = g_spring_last_setup;
  info[3] = m->spring_pc_its;
  info[4] = m->spring_pc_age;
  info[5] = m->spring_pc_age_its;
  return MV_OK;
}
