// Monge-Ampere relaxation kernels.
//
// All assembly kernels are "owner computes": one thread per (owned) vertex walks the
// vertex->cell list in ascending cell order, recomputes the P1 geometry of each incident
// simplex from the xi coordinates and accumulates its own row.  No atomics, fixed
// summation order, bit-reproducible.
#include <algorithm>

#include "mv_internal.cuh"

namespace mv {


// ------------------------------------------------------------------ quadrature tables
// 2D: Dunavant 6-point rule, exact to degree 4 = d+2.  3D: 14-point rule exact to degree
// 5 = d+2 (both verified against the barycentric monomial formula in tests/).
__constant__ double c_q2_lam[6][3] = {
    {0.10810301816807022736334149223390, 0.44594849091596488631832925388305, 0.44594849091596488631832925388305},
    {0.44594849091596488631832925388305, 0.10810301816807022736334149223390, 0.44594849091596488631832925388305},
    {0.44594849091596488631832925388305, 0.44594849091596488631832925388305, 0.10810301816807022736334149223390},
    {0.81684757298045851308085707319560, 0.09157621350977074345957146340220, 0.09157621350977074345957146340220},
    {0.09157621350977074345957146340220, 0.81684757298045851308085707319560, 0.09157621350977074345957146340220},
    {0.09157621350977074345957146340220, 0.09157621350977074345957146340220, 0.81684757298045851308085707319560}};
__constant__ double c_q2_w[6] = {0.22338158967801146569500700843312, 0.22338158967801146569500700843312,
                                 0.22338158967801146569500700843312, 0.10995174365532186763832632490021,
                                 0.10995174365532186763832632490021, 0.10995174365532186763832632490021};
#define Q3A1 0.31088591926330060980
#define Q3C1 (1.0 - 3.0 * Q3A1)
#define Q3A2 0.09273525031089122640
#define Q3C2 (1.0 - 3.0 * Q3A2)
#define Q3B 0.04550370412564964949
#define Q3D (0.5 - Q3B)
__constant__ double c_q3_lam[14][4] = {
    {Q3C1, Q3A1, Q3A1, Q3A1}, {Q3A1, Q3C1, Q3A1, Q3A1}, {Q3A1, Q3A1, Q3C1, Q3A1}, {Q3A1, Q3A1, Q3A1, Q3C1},
    {Q3C2, Q3A2, Q3A2, Q3A2}, {Q3A2, Q3C2, Q3A2, Q3A2}, {Q3A2, Q3A2, Q3C2, Q3A2}, {Q3A2, Q3A2, Q3A2, Q3C2},
    {Q3B, Q3B, Q3D, Q3D},     {Q3B, Q3D, Q3B, Q3D},     {Q3B, Q3D, Q3D, Q3B},     {Q3D, Q3B, Q3B, Q3D},
    {Q3D, Q3B, Q3D, Q3B},     {Q3D, Q3D, Q3B, Q3B}};
__constant__ double c_q3_w[14] = {0.11268792571801585080, 0.11268792571801585080, 0.11268792571801585080, 0.11268792571801585080,
                                  0.07349304311636194954, 0.07349304311636194954, 0.07349304311636194954, 0.07349304311636194954,
                                  0.04254602077708146644, 0.04254602077708146644, 0.04254602077708146644, 0.04254602077708146644,
                                  0.04254602077708146644, 0.04254602077708146644};

// 1D: two-point Gauss rule, exact to degree 3 = d+2
__constant__ double c_q1_lam[2][2] = {{0.78867513459481288225, 0.21132486540518711775}, {0.21132486540518711775, 0.78867513459481288225}};
template <int D> struct Quad;
template <> struct Quad<1> {
  static constexpr int Q = 2;
  __device__ static double lam(int q, int j) { return c_q1_lam[q][j]; }
  __device__ static double w(int q) { return 0.5; }
};
template <> struct Quad<2> {
  static constexpr int Q = 6;
  __device__ static double lam(int q, int j) { return c_q2_lam[q][j]; }
  __device__ static double w(int q) { return c_q2_w[q]; }
};
template <> struct Quad<3> {
  static constexpr int Q = 14;
  __device__ static double lam(int q, int j) { return c_q3_lam[q][j]; }
  __device__ static double w(int q) { return c_q3_w[q]; }
};

__device__ __forceinline__ double det_I_plus(const double (&h)[1][1]) { return 1.0 + h[0][0]; }
__device__ __forceinline__ double det_I_plus(const double (&h)[2][2]) {
  return (1.0 + h[0][0]) * (1.0 + h[1][1]) - h[0][1] * h[1][0];
}
__device__ __forceinline__ double det_I_plus(const double (&h)[3][3]) {
  const double a = 1.0 + h[0][0], e = 1.0 + h[1][1], i = 1.0 + h[2][2];
  return a * (e * i - h[1][2] * h[2][1]) - h[0][1] * (h[1][0] * i - h[1][2] * h[2][0]) +
         h[0][2] * (h[1][0] * h[2][1] - e * h[2][0]);
}

// ------------------------------------------------------------------ A1 right-hand side
// b_a,i = sum_{K ni i} vol(K)/(d+1) * (grad phi|_K)_a       (monge_ampere.py:328)
template <int D>
__global__ void __launch_bounds__(MV_BLOCK)
k_l2_rhs(const double* __restrict__ xi, const int* __restrict__ cells, int dg, const int* __restrict__ v2c_ptr,
         const int* __restrict__ v2c, int64_t nrows, const double* __restrict__ phi, double* __restrict__ b) {
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    double acc[D];
#pragma unroll
    for (int a = 0; a < D; ++a) acc[a] = 0.0;
    const int tb = __ldg(&v2c_ptr[v]), te = __ldg(&v2c_ptr[v + 1]);
    for (int t = tb; t < te; ++t) {
      const int64_t c = __ldg(&v2c[t]) / (D + 1);
      int cv[D + 1];
      double xc[D + 1][D];
      CellGeom<D> g;
      load_cell<D>(cells, c, cv);
      load_cell_coords<D>(xi, cv, xc, dg, c);
      cell_geom(xc, g);
      const double wgt = g.vol / (D + 1);
#pragma unroll
      for (int a = 0; a < D; ++a) {
        double gk = 0.0;
#pragma unroll
        for (int j = 0; j <= D; ++j) gk = fma(__ldg(&phi[cv[j]]), g.g[j][a], gk);
        acc[a] = fma(wgt, gk, acc[a]);
      }
    }
#pragma unroll
    for (int a = 0; a < D; ++a) b[v * D + a] = acc[a];
  }
}

// ------------------------------------------------------------------ A8 right-hand side
// c_ab,i = - sum_K vol(K) d_b psi_i (grad phi|_K)_a + sum_{F ni i} area(F)/d t_a n_b,
// t = g - (g.n) n with g = grad phi on the cell behind F      (monge_ampere.py:477-480)
template <int D>
__global__ void __launch_bounds__(MV_BLOCK)
k_hess_rhs(const double* __restrict__ xi, const int* __restrict__ cells, int dg, const int* __restrict__ v2c_ptr,
           const int* __restrict__ v2c, const int* __restrict__ v2f_ptr, const int* __restrict__ v2f,
           const int* __restrict__ bf_cell, const double* __restrict__ bf_normal, const double* __restrict__ bf_area,
           int64_t nrows, const double* __restrict__ phi, double* __restrict__ out, double* __restrict__ l2_out) {
  // l2_out != nullptr: also the right-hand side of the NEXT L2 projection of grad(phi) (k_l2_rhs, same phi, same cell
  // geometry and cell gradients -- the same operations in the same order, so the same bits)
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    double acc[D][D], acc1[D];
#pragma unroll
    for (int a = 0; a < D; ++a) {
      acc1[a] = 0.0;
#pragma unroll
      for (int b = 0; b < D; ++b) acc[a][b] = 0.0;
    }
    const int tb = __ldg(&v2c_ptr[v]), te = __ldg(&v2c_ptr[v + 1]);
    for (int t = tb; t < te; ++t) {
      const int entry = __ldg(&v2c[t]);
      const int64_t c = entry / (D + 1);
      const int loc = entry - (int)c * (D + 1);
      int cv[D + 1];
      double xc[D + 1][D];
      CellGeom<D> g;
      load_cell<D>(cells, c, cv);
      load_cell_coords<D>(xi, cv, xc, dg, c);
      cell_geom(xc, g);
      double gk[D], gl[D];
#pragma unroll
      for (int a = 0; a < D; ++a) {
        gk[a] = 0.0;
        gl[a] = 0.0;
#pragma unroll
        for (int j = 0; j <= D; ++j) {
          gk[a] = fma(__ldg(&phi[cv[j]]), g.g[j][a], gk[a]);
          if (j == loc) gl[a] = g.g[j][a];
        }
      }
#pragma unroll
      for (int a = 0; a < D; ++a)
#pragma unroll
        for (int b = 0; b < D; ++b) acc[a][b] -= g.vol * gl[b] * gk[a];
      const double wgt = g.vol / (D + 1);
#pragma unroll
      for (int a = 0; a < D; ++a) acc1[a] = fma(wgt, gk[a], acc1[a]);
    }
    const int fb = __ldg(&v2f_ptr[v]), fe = __ldg(&v2f_ptr[v + 1]);
    for (int t = fb; t < fe; ++t) {
      const int f = __ldg(&v2f[t]) / D;
      const int64_t c = __ldg(&bf_cell[f]);
      int cv[D + 1];
      double xc[D + 1][D];
      CellGeom<D> g;
      load_cell<D>(cells, c, cv);
      load_cell_coords<D>(xi, cv, xc);
      cell_geom(xc, g);
      double gk[D], n[D], gn = 0.0;
#pragma unroll
      for (int a = 0; a < D; ++a) {
        gk[a] = 0.0;
#pragma unroll
        for (int j = 0; j <= D; ++j) gk[a] = fma(__ldg(&phi[cv[j]]), g.g[j][a], gk[a]);
        n[a] = __ldg(&bf_normal[(int64_t)f * D + a]);
        gn += gk[a] * n[a];
      }
      const double wgt = __ldg(&bf_area[f]) / D;
#pragma unroll
      for (int a = 0; a < D; ++a) {
        const double ta = gk[a] - gn * n[a];
#pragma unroll
        for (int b = 0; b < D; ++b) acc[a][b] += wgt * ta * n[b];
      }
    }
#pragma unroll
    for (int a = 0; a < D; ++a)
#pragma unroll
      for (int b = 0; b < D; ++b) out[v * D * D + a * D + b] = acc[a][b];
    if (l2_out) {
#pragma unroll
      for (int a = 0; a < D; ++a) l2_out[v * D + a] = acc1[a];
    }
  }
}

// ------------------------------------------------------------------ A5 nonlinear pass
// a_i = int psi_i m det(I + H) dxi  (degree d+2 exact quadrature), fused with sum_i a_i.
// replaces: monge_ampere.py:406 (theta_form), :409 (_residual_l2_form)
template <int D>
__global__ void __launch_bounds__(MV_BLOCK)
k_assemble_a(const double* __restrict__ xi, const int* __restrict__ cells, int dg, const int* __restrict__ v2c_ptr,
             const int* __restrict__ v2c, int64_t nrows, const double* __restrict__ mon, const double* __restrict__ H,
             double* __restrict__ a_out, double* __restrict__ partials, unsigned* __restrict__ counter,
             double* __restrict__ dscal) {
  double tot[1] = {0.0};
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    double acc = 0.0;
    const int tb = __ldg(&v2c_ptr[v]), te = __ldg(&v2c_ptr[v + 1]);
    for (int t = tb; t < te; ++t) {
      const int entry = __ldg(&v2c[t]);
      const int64_t c = entry / (D + 1);
      const int loc = entry - (int)c * (D + 1);
      int cv[D + 1];
      double xc[D + 1][D];
      CellGeom<D> g;
      load_cell<D>(cells, c, cv);
      load_cell_coords<D>(xi, cv, xc, dg, c);
      cell_geom(xc, g);
      double mj[D + 1], Hj[D + 1][D][D];
#pragma unroll
      for (int j = 0; j <= D; ++j) {
        mj[j] = __ldg(&mon[cv[j]]);
#pragma unroll
        for (int a = 0; a < D; ++a)
#pragma unroll
          for (int b = 0; b < D; ++b) Hj[j][a][b] = __ldg(&H[(int64_t)cv[j] * D * D + a * D + b]);
      }
      double cell = 0.0;
#pragma unroll 2
      for (int q = 0; q < Quad<D>::Q; ++q) {
        double mq = 0.0, hq[D][D];
#pragma unroll
        for (int a = 0; a < D; ++a)
#pragma unroll
          for (int b = 0; b < D; ++b) hq[a][b] = 0.0;
        double lam_loc = 0.0;
#pragma unroll
        for (int j = 0; j <= D; ++j) {
          const double l = Quad<D>::lam(q, j);
          if (j == loc) lam_loc = l;
          mq = fma(l, mj[j], mq);
#pragma unroll
          for (int a = 0; a < D; ++a)
#pragma unroll
            for (int b = 0; b < D; ++b) hq[a][b] = fma(l, Hj[j][a][b], hq[a][b]);
        }
        cell = fma(Quad<D>::w(q) * lam_loc, mq * det_I_plus(hq), cell);
      }
      acc = fma(g.vol, cell, acc);
    }
    a_out[v] = acc;
    tot[0] += acc;
  }
  grid_reduce_sum<1>(tot, partials, counter, [&](double(&t)[1]) { dscal[D_SUMA] = t[0]; });
}


// Cell-major variant of the same assembly: one thread per cell evaluates the quadrature ONCE and stores the
// d+1 vertex contributions; a vertex pass then adds them in ascending cell order (the order of v2c), so the
// result is deterministic and every (m, H) value is gathered once per cell instead of once per (vertex, cell)
// pair: 4x (3D) less gather traffic than k_assemble_a, which ncu showed to be bound by it (24 ms at 256^3).
template <int D>
__global__ void __launch_bounds__(MV_BLOCK)
k_assemble_a_cells(const int* __restrict__ cells, const double* __restrict__ vol, int64_t ncells, const double* __restrict__ mon,
                   const double* __restrict__ H, double* __restrict__ contrib) {
  for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < ncells; c += (int64_t)gridDim.x * blockDim.x) {
    int cv[D + 1];
    load_cell<D>(cells, c, cv);
    double mj[D + 1], Hj[D + 1][D][D];
#pragma unroll
    for (int j = 0; j <= D; ++j) {
      mj[j] = __ldg(&mon[cv[j]]);
#pragma unroll
      for (int a = 0; a < D; ++a)
#pragma unroll
        for (int b = 0; b < D; ++b) Hj[j][a][b] = __ldg(&H[(int64_t)cv[j] * D * D + a * D + b]);
    }
    double cell[D + 1];
#pragma unroll
    for (int j = 0; j <= D; ++j) cell[j] = 0.0;
#pragma unroll 2
    for (int q = 0; q < Quad<D>::Q; ++q) {
      double mq = 0.0, hq[D][D];
#pragma unroll
      for (int a = 0; a < D; ++a)
#pragma unroll
        for (int b = 0; b < D; ++b) hq[a][b] = 0.0;
      double lam[D + 1];
#pragma unroll
      for (int j = 0; j <= D; ++j) {
        const double l = Quad<D>::lam(q, j);
        lam[j] = l;
        mq = fma(l, mj[j], mq);
#pragma unroll
        for (int a = 0; a < D; ++a)
#pragma unroll
          for (int b = 0; b < D; ++b) hq[a][b] = fma(l, Hj[j][a][b], hq[a][b]);
      }
      const double f = mq * det_I_plus(hq);
      const double wq = Quad<D>::w(q);
#pragma unroll
      for (int j = 0; j <= D; ++j) cell[j] = fma(wq * lam[j], f, cell[j]);
    }
    const double vk = __ldg(&vol[c]);
#pragma unroll
    for (int j = 0; j <= D; ++j) contrib[c * (D + 1) + j] = vk * cell[j];
  }
}
__global__ void __launch_bounds__(MV_BLOCK)
k_assemble_a_gather(const int* __restrict__ v2c_ptr, const int* __restrict__ v2c, int64_t nrows, const double* __restrict__ contrib,
                    double* __restrict__ a_out, double* __restrict__ partials, unsigned* __restrict__ counter,
                    double* __restrict__ dscal) {
  double tot[1] = {0.0};
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    double acc = 0.0;
    const int tb = __ldg(&v2c_ptr[v]), te = __ldg(&v2c_ptr[v + 1]);
    for (int t = tb; t < te; ++t) acc += __ldg(&contrib[__ldg(&v2c[t])]);
    a_out[v] = acc;
    tot[0] += acc;
  }
  grid_reduce_sum<1>(tot, partials, counter, [&](double(&t)[1]) { dscal[D_SUMA] = t[0]; });
}

// cof(I + h) : dh  (directional derivative of det(I + h) in direction dh)
__device__ __forceinline__ double cof_contract(const double (&)[1][1], const double (&dh)[1][1]) { return dh[0][0]; }
__device__ __forceinline__ double cof_contract(const double (&h)[2][2], const double (&dh)[2][2]) {
  return (1.0 + h[1][1]) * dh[0][0] - h[1][0] * dh[0][1] - h[0][1] * dh[1][0] + (1.0 + h[0][0]) * dh[1][1];
}
__device__ __forceinline__ double cof_contract(const double (&h)[3][3], const double (&dh)[3][3]) {
  double B[3][3];
#pragma unroll
  for (int a = 0; a < 3; ++a)
#pragma unroll
    for (int b = 0; b < 3; ++b) B[a][b] = h[a][b] + (a == b ? 1.0 : 0.0);
  double s = 0.0;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const int a1 = (a + 1) % 3, a2 = (a + 2) % 3;
#pragma unroll
    for (int b = 0; b < 3; ++b) {
      const int b1 = (b + 1) % 3, b2 = (b + 2) % 3;
      // cofactor with the cyclic-index identity (sign absorbed by the cyclic order)
      s = fma(B[a1][b1] * B[a2][b2] - B[a1][b2] * B[a2][b1], dh[a][b], s);
    }
  }
  return s;
}

// Q2: (A dH)_i = int psi_i m cof(I + H) : dH dxi   (m, theta frozen; same quadrature as A5)
// replaces: the UFL-derived Jacobian of movement/monge_ampere.py:640-645 (psi-row, H-column block)
template <int D>
__global__ void __launch_bounds__(MV_BLOCK)
k_jac_A(const double* __restrict__ xi, const int* __restrict__ cells, int dg, const int* __restrict__ v2c_ptr,
        const int* __restrict__ v2c, int64_t nrows, const double* __restrict__ mon, const double* __restrict__ H,
        const double* __restrict__ dH, double sign, double* __restrict__ out) {
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    double acc = 0.0;
    const int tb = __ldg(&v2c_ptr[v]), te = __ldg(&v2c_ptr[v + 1]);
    for (int t = tb; t < te; ++t) {
      const int entry = __ldg(&v2c[t]);
      const int64_t c = entry / (D + 1);
      const int loc = entry - (int)c * (D + 1);
      int cv[D + 1];
      double xc[D + 1][D];
      CellGeom<D> g;
      load_cell<D>(cells, c, cv);
      load_cell_coords<D>(xi, cv, xc, dg, c);
      cell_geom(xc, g);
      double cell = 0.0;
#pragma unroll 1
      for (int q = 0; q < Quad<D>::Q; ++q) {
        double mq = 0.0, hq[D][D], dq[D][D];
#pragma unroll
        for (int a = 0; a < D; ++a)
#pragma unroll
          for (int b = 0; b < D; ++b) { hq[a][b] = 0.0; dq[a][b] = 0.0; }
        double lam_loc = 0.0;
#pragma unroll
        for (int j = 0; j <= D; ++j) {
          const double l = Quad<D>::lam(q, j);
          if (j == loc) lam_loc = l;
          mq = fma(l, __ldg(&mon[cv[j]]), mq);
#pragma unroll
          for (int a = 0; a < D; ++a)
#pragma unroll
            for (int b = 0; b < D; ++b) {
              hq[a][b] = fma(l, __ldg(&H[(int64_t)cv[j] * D * D + a * D + b]), hq[a][b]);
              dq[a][b] = fma(l, __ldg(&dH[(int64_t)cv[j] * D * D + a * D + b]), dq[a][b]);
            }
        }
        cell = fma(Quad<D>::w(q) * lam_loc, mq * cof_contract(hq, dq), cell);
      }
      acc = fma(g.vol, cell, acc);
    }
    out[v] = sign * acc;
  }
}

// residual r = a - theta m_l ; sum r^2       (monge_ampere.py:212-223, :409, :514)
__global__ void __launch_bounds__(MV_BLOCK)
k_residual(int64_t nrows, const double* __restrict__ a, const double* __restrict__ ml, double inv_total_volume,
           double* __restrict__ res, double* __restrict__ partials, unsigned* __restrict__ counter,
           double* __restrict__ dscal) {
  const double theta = dscal[D_SUMA] * inv_total_volume;
  double s[1] = {0.0};
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nrows; v += (int64_t)gridDim.x * blockDim.x) {
    const double r = a[v] - theta * __ldg(&ml[v]);
    res[v] = r;
    s[0] = fma(r, r, s[0]);
  }
  grid_reduce_sum<1>(s, partials, counter, [&](double(&t)[1]) {
    dscal[D_RES2] = t[0];
    dscal[D_THETA] = theta;
  });
}

// ------------------------------------------------------------------ A2 + A3
struct MonParams {
  int kind;
  double p[8];
};

template <int D>
__device__ __forceinline__ double monitor_raw(const MonParams& mp, const double (&x)[D]) {
  switch (mp.kind) {
    case MV_MON_BALL:
    case MV_MON_BALL_RAW: {
      double dist = 0.0;
#pragma unroll
      for (int a = 0; a < D; ++a) {
        const double dd = x[a] - mp.p[a];
        dist += dd * dd;
      }
      const double rr = mp.kind == MV_MON_BALL ? mp.p[3] * mp.p[3] : mp.p[3];
      const double ch = cosh(mp.p[5] * (dist - rr));
      return 1.0 + mp.p[4] / (ch * ch);
    }
    case MV_MON_GAUSS_HESS: {
      // u = exp(-((x-x0)^2+(y-y0)^2)/w^2); |Hess u|_F
      const double w2 = mp.p[2] * mp.p[2];
      const double dx = x[0] - mp.p[0], dy = x[1 % D] - mp.p[1];
      const double u = exp(-(dx * dx + dy * dy) / w2);
      const double uxx = u * (4.0 * dx * dx / (w2 * w2) - 2.0 / w2);
      const double uyy = u * (4.0 * dy * dy / (w2 * w2) - 2.0 / w2);
      const double uxy = u * (4.0 * dx * dy / (w2 * w2));
      return sqrt(uxx * uxx + uyy * uyy + 2.0 * uxy * uxy);
    }
    case MV_MON_SINATAN3: {
      // u = 0.1 sin(50xz) + atan2(0.1, sin(5y) - 2xz); H:H (no sqrt)
      const double X = x[0], Y = x[1 % D], Z = x[2 % D];
      double sn, cs;
      sincos(50.0 * X * Z, &sn, &cs);
      double h[3][3] = {{-250.0 * Z * Z * sn, 0.0, 5.0 * cs - 250.0 * X * Z * sn},
                        {0.0, 0.0, 0.0},
                        {5.0 * cs - 250.0 * X * Z * sn, 0.0, -250.0 * X * X * sn}};
      double s5, c5;
      sincos(5.0 * Y, &s5, &c5);
      const double q = s5 - 2.0 * X * Z;
      const double Dn = 0.01 + q * q;
      const double g = -0.1 / Dn;
      const double gp = 0.2 * q / (Dn * Dn);
      const double gq[3] = {-2.0 * Z, 5.0 * c5, -2.0 * X};
      double hh = 0.0;
#pragma unroll
      for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) {
          double v = h[a][b] + gp * gq[a] * gq[b];
          if (a == 1 && b == 1) v += g * (-25.0 * s5);
          if ((a == 0 && b == 2) || (a == 2 && b == 0)) v += g * (-2.0);
          hh = fma(v, v, hh);
        }
      return hh;
    }
    default:
      return 1.0;
  }
}

__host__ __device__ inline bool monitor_normalised(int kind) { return kind == MV_MON_GAUSS_HESS || kind == MV_MON_SINATAN3; }

// x = xi + g (optional) ; raw monitor at x ; running max for the normalised kinds
// rep != nullptr (periodic mesh, DG coordinate field): the loop runs over the P1 vertices, vertex v is evaluated at its
// representative coordinate node rep[v]; the DG field x itself is written by k_update_x_dg
template <int D, bool UPDATE_X>
__global__ void __launch_bounds__(MV_BLOCK)
k_update_x_monitor(int64_t n, const double* __restrict__ xi, const double* __restrict__ g, double* __restrict__ x,
                   MonParams mp, double* __restrict__ mon, double* __restrict__ dscal, const int* __restrict__ rep) {
  double vmax = 0.0;
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < n; v += (int64_t)gridDim.x * blockDim.x) {
    double xv[D];
    const int64_t node = rep ? (int64_t)rep[v] : v;
#pragma unroll
    for (int a = 0; a < D; ++a) {
      if (UPDATE_X) {
        xv[a] = xi[node * D + a] + g[v * D + a];
        if (!rep) x[v * D + a] = xv[a];
      } else {
        xv[a] = x[node * D + a];
      }
    }
    if (mp.kind != MV_MON_USER) {
      const double r = monitor_raw<D>(mp, xv);
      mon[v] = r;
      vmax = fmax(vmax, r);
    }
  }
  if (monitor_normalised(mp.kind)) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) vmax = fmax(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    if ((threadIdx.x & 31) == 0) atomic_max_pos(&dscal[D_MONMAX], vmax);
  }
}

__global__ void __launch_bounds__(MV_BLOCK)
k_monitor_normalise(int64_t n, double alpha, double* __restrict__ mon, const double* __restrict__ dscal) {
  const double s = alpha / dscal[D_MONMAX];
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < n; v += (int64_t)gridDim.x * blockDim.x)
    mon[v] = fma(s, mon[v], 1.0);
}

// periodic meshes: x = xi + grad_phi interpolated from the periodic P1 space into the DG coordinate space
// (movement/monge_ampere.py:236-239), one thread per coordinate node
template <int D>
__global__ void __launch_bounds__(MV_BLOCK)
k_update_x_dg(int64_t nnodes, const int* __restrict__ cells_flat, const double* __restrict__ xi, const double* __restrict__ g,
              double* __restrict__ x) {
  for (int64_t nd = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; nd < nnodes; nd += (int64_t)gridDim.x * blockDim.x) {
    const int64_t v = cells_flat[nd];
#pragma unroll
    for (int a = 0; a < D; ++a) x[nd * D + a] = xi[nd * D + a] + g[v * D + a];
  }
}

__global__ void k_set_scalar(double* p, double v) { *p = v; }

__global__ void __launch_bounds__(MV_BLOCK)
k_scale(int64_t n, double s, const double* __restrict__ a, double* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = s * a[i];
}
__global__ void __launch_bounds__(MV_BLOCK)
k_add(int64_t n, const double* __restrict__ a, const double* __restrict__ b, double* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = a[i] + b[i];
}

// boundary mask: for every boundary facet with tag index k, OR tagbits[k] into its vertices
__global__ void k_bc_mask(int64_t F, int D, const int* __restrict__ bf_verts, const int* __restrict__ bf_tagidx,
                          const unsigned* __restrict__ tagbits, int64_t nown, unsigned* __restrict__ mask32) {
  for (int64_t f = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; f < F; f += (int64_t)gridDim.x * blockDim.x) {
    const unsigned bits = tagbits[bf_tagidx[f]];
    if (!bits) continue;
    for (int a = 0; a < D; ++a) {
      const int v = bf_verts[f * D + a];
      if (v < nown) atomicOr(&mask32[v >> 1], bits << ((v & 1) * 16));
    }
  }
}

template <int D>
static int eval_monitor(mv_ma* ma, bool update_x, const double* xin, double* xout, double* mon) {
  mv_mesh* m = ma->mesh;
  MonParams mp;
  mp.kind = ma->mon_kind;
  for (int i = 0; i < 8; ++i) mp.p[i] = ma->mon_params[i];
  const int64_t n = m->V;  // ghosts too: every rank evaluates the monitor at all its local vertices
  const int grid = grid_for(n);
  if (monitor_normalised(mp.kind)) MV_L k_set_scalar<<<1, 1, 0, g_stream>>>(ma->dscal + D_MONMAX, 0.0);
  if (update_x) {
    MV_L k_update_x_monitor<D, true><<<grid, MV_BLOCK, 0, g_stream>>>(n, xin, ma->f[MV_F_GRAD_PHI], xout, mp, mon, ma->dscal, m->rep);
    if (m->dg) MV_L k_update_x_dg<D><<<grid_for(m->Vc), MV_BLOCK, 0, g_stream>>>(m->Vc, m->cells, xin, ma->f[MV_F_GRAD_PHI], xout);
  } else {
    MV_L k_update_x_monitor<D, false><<<grid, MV_BLOCK, 0, g_stream>>>(n, nullptr, nullptr, xout, mp, mon, ma->dscal, m->rep);
  }
  MV_KERNEL_CHECK();
  if (monitor_normalised(mp.kind)) {
    // NB the reference demos take a rank-local max (demos/monge_ampere_3d.py:67); a global
    // max is the only definition that is partition independent.
    if (comm_active()) MV_CHECK(allreduce_dev(ma->dscal + D_MONMAX, 1, 1));
    const double alpha = mp.kind == MV_MON_GAUSS_HESS ? mp.p[3] : mp.p[0];
    MV_L k_monitor_normalise<<<grid, MV_BLOCK, 0, g_stream>>>(n, alpha, mon, ma->dscal);
    MV_KERNEL_CHECK();
  }
  return MV_OK;
}

}  // namespace mv

using namespace mv;

#define MA_REQUIRE(ma, field)                                         \
  if (!(ma)->f[field]) {                                              \
    set_error("mv_ma: field %d not bound", (int)(field));             \
    return MV_ERR_BADARG;                                             \
  }

extern "C" int mv_ma_create(mv_mesh* m, mv_ma** out) {
  if (!m || !out) return MV_ERR_BADARG;
  mv_ma* ma = new mv_ma();
  ma->mesh = m;
  const size_t maxc = (size_t)m->dim * m->dim;
  int st;
  if ((st = dev_alloc(&ma->rhs, m->V * maxc)) != MV_OK || (st = dev_alloc(&ma->a, m->V)) != MV_OK ||
      (st = dev_alloc(&ma->dphi, m->V)) != MV_OK || (st = dev_alloc(&ma->dscal, D_N)) != MV_OK ||
      (st = dev_alloc(&ma->bc_mask, m->V + 2)) != MV_OK) {
    mv_ma_destroy(ma);
    return st;
  }
  MV_CUDA(cudaMemsetAsync(ma->dscal, 0, D_N * sizeof(double), g_stream));
  MV_CUDA(cudaMemsetAsync(ma->bc_mask, 0, (m->V + 2) * sizeof(uint16_t), g_stream));
  MV_CUDA(cudaMemsetAsync(ma->dphi, 0, m->V * sizeof(double), g_stream));
  MV_CUDA(cudaMallocHost((void**)&ma->h_scal, D_N * sizeof(double)));
  *out = ma;
  return MV_OK;
}

extern "C" void mv_ma_destroy(mv_ma* ma) {
  if (!ma) return;
  cudaStreamSynchronize(g_stream);
  dev_free(ma->qn_basis); dev_free(ma->qn_w); dev_free(ma->qn_gphi); dev_free(ma->qn_gH); dev_free(ma->qn_t1);
  dev_free(ma->qn_t2); dev_free(ma->qn_base); dev_free(ma->qn_dots);
  if (ma->qn_hdots) cudaFreeHost(ma->qn_hdots);
  dev_free(ma->rhs); dev_free(ma->a); dev_free(ma->dphi); dev_free(ma->dscal); dev_free(ma->bc_mask);
  for (int i = 0; i < 8; ++i) { dev_free(ma->hist_x[i]); dev_free(ma->hist_b[i]); }
  dev_free(ma->hist_G);
  dev_free(ma->l2_rhs);
  if (ma->h_scal) cudaFreeHost(ma->h_scal);
  eqbc_free(ma);
  delete ma;
}

extern "C" int mv_ma_bind(mv_ma* ma, int field, double* p) {
  if (!ma || field < 0 || field >= MV_F_COUNT) return MV_ERR_BADARG;
  if (p && !is_device_ptr(p)) {
    set_error("mv_ma_bind: field %d must be a device pointer", field);
    return MV_ERR_BADARG;
  }
  ma->f[field] = p;
  return MV_OK;
}

extern "C" int mv_ma_set_solver(mv_ma* ma, double rtol, double atol, int maxit_mass, int maxit_poisson) {
  if (!ma) return MV_ERR_BADARG;
  ma->rtol = rtol; ma->atol = atol; ma->maxit_mass = maxit_mass; ma->maxit_poisson = maxit_poisson;
  return MV_OK;
}

extern "C" int mv_ma_set_monitor(mv_ma* ma, int kind, int nparams, const double* params) {
  if (!ma || nparams > 8 || nparams < 0) return MV_ERR_BADARG;
  const int d = ma->mesh->dim;
  if ((kind == MV_MON_GAUSS_HESS && d != 2) || (kind == MV_MON_SINATAN3 && d != 3)) {
    set_error("mv_ma_set_monitor: monitor kind %d is not defined in %dD", kind, d);
    return MV_ERR_BADARG;
  }
  ma->mon_kind = kind;
  for (int i = 0; i < 8; ++i) ma->mon_params[i] = i < nparams ? params[i] : 0.0;
  return MV_OK;
}

extern "C" int mv_ma_set_boundary(mv_ma* ma, int n_fixed, const int32_t* fixed_tags) {
  return mv_ma_set_boundary_ex(ma, n_fixed, fixed_tags, 0, nullptr, nullptr);
}

extern "C" int mv_ma_set_boundary_ex(mv_ma* ma, int n_fixed, const int32_t* fixed_tags, int n_all, const int32_t* all_tags,
                                     int32_t* kinds) {
  if (!ma || (n_all > 0 && !all_tags)) return MV_ERR_BADARG;
  mv_mesh* m = ma->mesh;
  const int d = m->dim;
  const int64_t F = m->F;
  MV_CUDA(cudaMemsetAsync(ma->bc_mask, 0, (m->V + 2) * sizeof(uint16_t), g_stream));
  ma->have_bc = false;
  eqbc_free(ma);
  if (F == 0 && n_all == 0) return MV_OK;
  std::vector<double> nrm(F * d), area(F);
  if (F > 0) {
    MV_CUDA(cudaMemcpyAsync(nrm.data(), m->bf_normal, F * d * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaMemcpyAsync(area.data(), m->bf_area, F * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
  }
  // the tags the decision is taken over: the caller's list (every tag of the whole mesh, the same on every rank), else
  // the tags of this rank's facets
  std::vector<int> tags;
  if (n_all > 0) tags.assign(all_tags, all_tags + n_all);
  else tags = m->h_bf_tags;
  std::sort(tags.begin(), tags.end());
  tags.erase(std::unique(tags.begin(), tags.end()), tags.end());
  for (int i = 0; i < n_fixed; ++i)
    if (!std::binary_search(tags.begin(), tags.end(), fixed_tags[i])) {
      set_error("Provided boundary_tag '%d' is invalid.", fixed_tags[i]);
      return MV_ERR_BADARG;
    }
  std::vector<int> tagidx(F);
  std::vector<double> absn(tags.size() * d, 0.0);
  for (int64_t f = 0; f < F; ++f) {
    const int k = (int)(std::lower_bound(tags.begin(), tags.end(), m->h_bf_tags[f]) - tags.begin());
    if (k >= (int)tags.size() || tags[k] != m->h_bf_tags[f]) {
      set_error("mv_ma_set_boundary_ex: facet tag %d is not in the list of all tags", m->h_bf_tags[f]);
      return MV_ERR_BADARG;
    }
    tagidx[f] = k;
    for (int a = 0; a < d; ++a) absn[k * d + a] += fabs(nrm[f * d + a]) * area[f];
  }
  // assemble(abs(n_i) * ds(tag)) is a global integral: sum the ranks' parts, so that every rank takes the same branch
  if (comm_active() && n_all > 0) MV_CHECK(allreduce_host(absn.data(), (int)absn.size(), 0));
  std::vector<unsigned> bits(tags.size(), 0u);
  std::vector<char> is_eq(tags.size(), 0);
  const unsigned all = (1u << d) - 1u;
  bool any_eq = false;
  for (size_t k = 0; k < tags.size(); ++k) {
    bool fixed = false;
    for (int i = 0; i < n_fixed; ++i) fixed |= (fixed_tags[i] == tags[k]);
    if (kinds) kinds[k] = 0;
    if (fixed || d == 1) { bits[k] = all; continue; }   // any tag in 1D: the whole vector is fixed (monge_ampere.py:258-260)
    int nzero = 0, idx = -1;
    for (int a = 0; a < d; ++a) {
      if (fabs(absn[k * d + a]) <= 1e-8) ++nzero; else idx = a;   // np.isclose(., 0.0); NaN (degenerate facets) is not close
    }
    if (nzero == d - 1) {
      bits[k] = 1u << idx;
      if (kinds) kinds[k] = 1 + idx;
    } else {
      // EquationBC branch (monge_ampere.py:274-309): every component at the segment's nodes is prescribed by the boundary solve
      bits[k] = all;
      is_eq[k] = 1;
      any_eq = true;
      if (kinds) kinds[k] = -1;
    }
  }
  if (any_eq && comm_active()) {   // the same verdict on every rank (absn was summed over the ranks)
    set_error("non-axis-aligned free boundary segments (EquationBC branch) are not partitioned over several GPUs yet");
    return MV_ERR_UNSUPPORTED;
  }
  if (F == 0) return MV_OK;
  if (any_eq) {
    std::vector<int> verts(F * d);
    MV_CUDA(cudaMemcpyAsync(verts.data(), m->bf_verts, F * d * sizeof(int), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    MV_CHECK(eqbc_build(ma, tagidx, is_eq, verts, area));
  }
  int* d_tagidx = nullptr;
  unsigned* d_bits = nullptr;
  MV_CUDA(cudaMallocAsync((void**)&d_tagidx, F * sizeof(int), g_stream));
  MV_CUDA(cudaMallocAsync((void**)&d_bits, bits.size() * sizeof(unsigned), g_stream));
  MV_CUDA(cudaMemcpyAsync(d_tagidx, tagidx.data(), F * sizeof(int), cudaMemcpyHostToDevice, g_stream));
  MV_CUDA(cudaMemcpyAsync(d_bits, bits.data(), bits.size() * sizeof(unsigned), cudaMemcpyHostToDevice, g_stream));
  MV_L k_bc_mask<<<grid_for(F), MV_BLOCK, 0, g_stream>>>(F, d, m->bf_verts, d_tagidx, d_bits, m->Vown, (unsigned*)ma->bc_mask);
  MV_KERNEL_CHECK();
  MV_CUDA(cudaFreeAsync(d_tagidx, g_stream));
  MV_CUDA(cudaFreeAsync(d_bits, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  ma->have_bc = true;
  return MV_OK;
}

extern "C" int mv_ma_get_bcmask(mv_ma* ma, uint16_t* dst) {
  if (!ma || !dst) return MV_ERR_BADARG;
  MV_CUDA(cudaMemcpyAsync(dst, ma->bc_mask, ma->mesh->Vown * sizeof(uint16_t), cudaMemcpyDefault, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  return MV_OK;
}

namespace mv {
int l2_rhs_dev(mv_mesh* m, const double* phi, double* b) {
  const int grid = grid_for(m->Vown);
  MV_DISPATCH_DIM(m->dim, MV_L k_l2_rhs<DIM><<<grid, MV_BLOCK, 0, g_stream>>>(m->xi, m->cells, m->dg, m->v2c_ptr, m->v2c, m->Vown, phi, b));
  MV_KERNEL_CHECK();
  return MV_OK;
}
int hess_rhs_dev(mv_mesh* m, const double* phi, double* c, double* l2_out) {
  const int grid = grid_for(m->Vown);
  MV_DISPATCH_DIM(m->dim, MV_L k_hess_rhs<DIM><<<grid, MV_BLOCK, 0, g_stream>>>(m->xi, m->cells, m->dg, m->v2c_ptr, m->v2c, m->v2f_ptr, m->v2f,
                                                                           m->bf_cell, m->bf_normal, m->bf_area, m->Vown, phi, c, l2_out));
  MV_KERNEL_CHECK();
  return MV_OK;
}
int assemble_a_dev(mv_mesh* m, const double* mon, const double* H, double* a, double* dscal) {
  if (!m->cell_tmp) MV_CHECK(dev_alloc(&m->cell_tmp, (size_t)m->C * (m->dim + 1)));
  const int gc = grid_for(m->C), gv = grid_for(m->Vown);
  MV_DISPATCH_DIM(m->dim, MV_L k_assemble_a_cells<DIM><<<gc, MV_BLOCK, 0, g_stream>>>(m->cells, m->vol, m->C, mon, H, m->cell_tmp));
  MV_L k_assemble_a_gather<<<gv, MV_BLOCK, 0, g_stream>>>(m->v2c_ptr, m->v2c, m->Vown, m->cell_tmp, a, m->red.partials, m->red.counter,
                                                     dscal);
  MV_KERNEL_CHECK();
  return MV_OK;
}

int jac_A_dev(mv_mesh* m, const double* mon, const double* H, const double* dH, double sign, double* out) {
  const int grid = grid_for(m->Vown);
  MV_DISPATCH_DIM(m->dim, MV_L k_jac_A<DIM><<<grid, MV_BLOCK, 0, g_stream>>>(m->xi, m->cells, m->dg, m->v2c_ptr, m->v2c, m->Vown, mon, H, dH, sign, out));
  MV_KERNEL_CHECK();
  return MV_OK;
}
}  // namespace mv

extern "C" int mv_ma_l2_rhs(mv_mesh* m, const double* phi, double* b) {
  if (!m || !phi || !b) return MV_ERR_BADARG;
  Staged sp, sb;
  MV_CHECK(sp.in(phi, m->V * sizeof(double)));
  MV_CHECK(sb.out_begin(b, m->Vown * m->dim * sizeof(double)));
  MV_CHECK(l2_rhs_dev(m, sp.as<double>(), sb.as<double>()));
  return sb.out_end(b, m->Vown * m->dim * sizeof(double));
}
extern "C" int mv_ma_hess_rhs(mv_mesh* m, const double* phi, double* c) {
  if (!m || !phi || !c) return MV_ERR_BADARG;
  Staged sp, sc;
  const size_t nb = m->Vown * m->dim * m->dim * sizeof(double);
  MV_CHECK(sp.in(phi, m->V * sizeof(double)));
  MV_CHECK(sc.out_begin(c, nb));
  MV_CHECK(hess_rhs_dev(m, sp.as<double>(), sc.as<double>()));
  return sc.out_end(c, nb);
}
extern "C" int mv_ma_assemble_a(mv_mesh* m, const double* mon, const double* H, double* a) {
  if (!m || !mon || !H || !a) return MV_ERR_BADARG;
  Staged sm, sh, sa;
  MV_CHECK(sm.in(mon, m->V * sizeof(double)));
  MV_CHECK(sh.in(H, m->V * m->dim * m->dim * sizeof(double)));
  MV_CHECK(sa.out_begin(a, m->Vown * sizeof(double)));
  double* dscal = nullptr;
  MV_CUDA(cudaMallocAsync((void**)&dscal, D_N * sizeof(double), g_stream));
  int st = assemble_a_dev(m, sm.as<double>(), sh.as<double>(), sa.as<double>(), dscal);
  MV_CUDA(cudaFreeAsync(dscal, g_stream));
  MV_CHECK(st);
  return sa.out_end(a, m->Vown * sizeof(double));
}

extern "C" int mv_ma_l2_project(mv_ma* ma, int* iters) {
  if (!ma) return MV_ERR_BADARG;
  MA_REQUIRE(ma, MV_F_PHI_OLD);
  MA_REQUIRE(ma, MV_F_GRAD_PHI);
  mv_mesh* m = ma->mesh;
  if (comm_active()) MV_CHECK(halo_exchange(m, 1, ma->f[MV_F_PHI_OLD]));
  // inside mv_ma_relax_move the right-hand side came with the previous iteration's Hessian right-hand side (same phi)
  const double* rhs = ma->rhs;
  if (ma->l2_rhs_valid && ma->l2_rhs_phi == ma->f[MV_F_PHI_OLD]) rhs = ma->l2_rhs;
  else MV_CHECK(l2_rhs_dev(m, ma->f[MV_F_PHI_OLD], ma->rhs));
  ma->l2_rhs_valid = false;
  MV_CHECK(eqbc_apply(ma, ma->f[MV_F_PHI_OLD], ma->f[MV_F_GRAD_PHI]));   // EquationBC segments: boundary solve first, its values are data
  MV_CHECK(pcg_solve(m, MV_OP_MASS, m->dim, rhs, ma->f[MV_F_GRAD_PHI], ma->have_bc ? ma->bc_mask : nullptr, ma->rtol,
                     ma->atol, ma->maxit_mass, 0, iters, nullptr));
  if (comm_active()) MV_CHECK(halo_exchange(m, m->dim, ma->f[MV_F_GRAD_PHI]));
  return MV_OK;
}

extern "C" int mv_ma_update_coordinates(mv_ma* ma) {
  if (!ma) return MV_ERR_BADARG;
  MA_REQUIRE(ma, MV_F_XI);
  MA_REQUIRE(ma, MV_F_X);
  MA_REQUIRE(ma, MV_F_GRAD_PHI);
  if (ma->mon_kind != MV_MON_USER) MA_REQUIRE(ma, MV_F_MONITOR);
  MV_DISPATCH_DIM(ma->mesh->dim, return eval_monitor<DIM>(ma, true, ma->f[MV_F_XI], ma->f[MV_F_X], ma->f[MV_F_MONITOR]));
  return MV_OK;
}

extern "C" int mv_ma_eval_monitor(mv_ma* ma, const double* coords, double* m_out) {
  if (!ma || !coords || !m_out) return MV_ERR_BADARG;
  if (ma->mon_kind == MV_MON_USER) {
    set_error("mv_ma_eval_monitor: no built-in monitor set");
    return MV_ERR_BADARG;
  }
  if (!is_device_ptr(coords) || !is_device_ptr(m_out)) {
    set_error("mv_ma_eval_monitor: device pointers required");
    return MV_ERR_BADARG;
  }
  MV_DISPATCH_DIM(ma->mesh->dim, return eval_monitor<DIM>(ma, false, nullptr, (double*)coords, m_out));
  return MV_OK;
}

extern "C" int mv_ma_diagnostics(mv_ma* ma, int use_H_old, double* out) {
  if (!ma || !out) return MV_ERR_BADARG;
  MA_REQUIRE(ma, MV_F_X);
  MA_REQUIRE(ma, MV_F_MONITOR);
  MA_REQUIRE(ma, MV_F_VOLUME);
  MA_REQUIRE(ma, MV_F_ORIG_VOLUME);
  MA_REQUIRE(ma, MV_F_RESIDUAL);
  const int hf = use_H_old ? MV_F_H_OLD : MV_F_H;
  MA_REQUIRE(ma, hf);
  mv_mesh* m = ma->mesh;
  const bool dist = comm_active();
  double gcells = (double)m->Cown, gtotvol = m->total_volume, gml2 = m->ml_sumsq;
  if (dist) {
    // three constants of the xi-mesh: summed over the ranks once per mover, not once per relaxation iteration (the
    // H2D + allreduce + D2H + stream synchronisation cost a visible share of a 33 ms step at 8 GPUs)
    if (!ma->gconst_ok) {
      double h[3] = {gcells, gtotvol, gml2};
      MV_CUDA(cudaMemcpyAsync(ma->dscal + 8, h, sizeof(h), cudaMemcpyHostToDevice, g_stream));
      MV_CHECK(allreduce_sum_dev(ma->dscal + 8, 3));
      MV_CUDA(cudaMemcpyAsync(h, ma->dscal + 8, sizeof(h), cudaMemcpyDeviceToHost, g_stream));
      MV_CUDA(cudaStreamSynchronize(g_stream));
      ma->gconst[0] = h[0]; ma->gconst[1] = h[1]; ma->gconst[2] = h[2];
      ma->gconst_ok = true;
    }
    gcells = ma->gconst[0]; gtotvol = ma->gconst[1]; gml2 = ma->gconst[2];
    MV_CHECK(halo_exchange(m, m->dim * m->dim, ma->f[hf]));
  }
  // A4 volumes (monge_ampere.py:507-510)
  MV_CHECK(volume_stats_device(m, ma->f[MV_F_X], ma->f[MV_F_MONITOR], ma->f[MV_F_ORIG_VOLUME], ma->f[MV_F_VOLUME],
                               ma->dscal + D_VSUM, (int64_t)(gcells + 0.5)));
  if (dist) MV_CHECK(allreduce_sum_dev(ma->dscal + D_VSSD, 1));
  // A5 theta + residual
  MV_CHECK(assemble_a_dev(m, ma->f[MV_F_MONITOR], ma->f[hf], ma->a, ma->dscal));
  if (dist) MV_CHECK(allreduce_sum_dev(ma->dscal + D_SUMA, 1));
  MV_L k_residual<<<grid_for(m->Vown), MV_BLOCK, 0, g_stream>>>(m->Vown, ma->a, m->ml, 1.0 / gtotvol, ma->f[MV_F_RESIDUAL],
                                                          m->red.partials, m->red.counter, ma->dscal);
  MV_KERNEL_CHECK();
  if (dist) MV_CHECK(allreduce_sum_dev(ma->dscal + D_RES2, 1));
  MV_CUDA(cudaMemcpyAsync(ma->h_scal, ma->dscal, D_N * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  if (comm_active()) MV_CHECK(p2p_check_status());   // once per relaxation iteration: a timed-out peer-memory wait is an error, not a wrong number
  const double* h = ma->h_scal;
  const double theta = h[D_THETA];
  const double mean = h[D_VSUM] / gcells;
  out[0] = theta;
  out[1] = sqrt(h[D_RES2]) / (fabs(theta) * sqrt(gml2));
  out[2] = h[D_VMAX];
  out[3] = h[D_VMIN];
  out[4] = mean;
  out[5] = sqrt(h[D_VSSD] / gcells) / mean;
  return MV_OK;
}

// Initial guess of the Poisson solve K x = b from the previous solves of the same mover: the Galerkin (energy-optimal)
// combination x0 = sum_i a_i x_i over the last m solutions, G a = f with G_ij = x_i.K x_j = x_i.b_j (the right-hand sides
// b_j are kept, so no matrix pass is needed) and f_i = x_i.b.  G is kept on the device and only its newest row is computed
// when a pair is added (m fused dot products); a solve costs m more for f and one combination pass.  The small system is
// solved by Cholesky in the order newest -> oldest and truncated at the first pivot that is (numerically) zero, i.e. when
// an older solution adds nothing to the span.  The right-hand side dt * (a - theta m_l) changes smoothly from one relaxation
// iteration to the next, so x0 is typically two to three digits accurate; the solve itself and its stopping rule (rtol
// relative to b) are unchanged.  Depth: MV_POISSON_HISTORY (0 off, 1 = default depth, k >= 2 = k).
#define MV_HIST_MAX 8
#define MV_HIST_DEFAULT 3
enum { H_C = 8, H_S0 = 8 + MV_HIST_MAX, H_END = 9 + MV_HIST_MAX };
static_assert((int)H_END <= (int)D_N, "history scalars must fit the device scalar block");
struct HistPtrs {
  const double* x[MV_HIST_MAX];
  const double* b[MV_HIST_MAX];
};
// c_i = X_i . y for the m stored vectors (X = x or b side), s0 = z . y (optional)
template <int M>
__global__ void __launch_bounds__(MV_BLOCK)
k_hist_dots(int64_t n, HistPtrs h, int side_b, const double* __restrict__ y, const double* __restrict__ z,
            double* __restrict__ partials, unsigned* __restrict__ counter, double* __restrict__ dscal) {
  double acc[M + 1];
#pragma unroll
  for (int i = 0; i <= M; ++i) acc[i] = 0.0;
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < n; v += (int64_t)gridDim.x * blockDim.x) {
    const double yy = y[v];
#pragma unroll
    for (int i = 0; i < M; ++i) acc[i] = fma((side_b ? h.b[i] : h.x[i])[v], yy, acc[i]);
    if (z) acc[M] = fma(z[v], yy, acc[M]);
  }
  grid_reduce_sum<M + 1>(acc, partials, counter, [&](double(&t)[M + 1]) {
#pragma unroll
    for (int i = 0; i < M; ++i) dscal[H_C + i] = t[i];
    dscal[H_S0] = t[M];
  });
}
// G a = f (m x m, newest first) by Cholesky, truncated at the first negligible pivot; a -> dscal[H_C..]
__global__ void k_hist_solve(int m, const double* __restrict__ G, double* __restrict__ dscal) {
  double Lm[MV_HIST_MAX][MV_HIST_MAX], y[MV_HIST_MAX], a[MV_HIST_MAX];
  int mm = 0;
  for (int j = 0; j < m; ++j) {
    double d = G[j * MV_HIST_MAX + j];
    for (int k = 0; k < j; ++k) d -= Lm[j][k] * Lm[j][k];
    if (!(d > 1e-10 * G[j * MV_HIST_MAX + j]) || !(G[j * MV_HIST_MAX + j] > 0.0)) break;
    Lm[j][j] = sqrt(d);
    for (int i = j + 1; i < m; ++i) {
      double t = G[i * MV_HIST_MAX + j];
      for (int k = 0; k < j; ++k) t -= Lm[i][k] * Lm[j][k];
      Lm[i][j] = t / Lm[j][j];
    }
    mm = j + 1;
  }
  for (int i = 0; i < mm; ++i) {
    double t = dscal[H_C + i];
    for (int k = 0; k < i; ++k) t -= Lm[i][k] * y[k];
    y[i] = t / Lm[i][i];
  }
  for (int i = mm - 1; i >= 0; --i) {
    double t = y[i];
    for (int k = i + 1; k < mm; ++k) t -= Lm[k][i] * a[k];
    a[i] = t / Lm[i][i];
  }
  for (int i = 0; i < MV_HIST_MAX; ++i) {
    const double v = i < mm ? a[i] : 0.0;
    dscal[H_C + i] = v == v ? v : 0.0;
  }
}
// x0 = sum_i a_i x_i
template <int M>
__global__ void __launch_bounds__(MV_BLOCK)
k_hist_guess(int64_t n, HistPtrs h, const double* __restrict__ dscal, double* __restrict__ x0) {
  double c[M > 0 ? M : 1];
#pragma unroll
  for (int i = 0; i < M; ++i) c[i] = dscal[H_C + i];
  for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < n; v += (int64_t)gridDim.x * blockDim.x) {
    double a = 0.0;
#pragma unroll
    for (int i = 0; i < M; ++i) a = fma(c[i], h.x[i][v], a);
    x0[v] = a;
  }
}
// the pair that was just stored becomes index 0: shift G by one and insert its row (dscal[H_S0] = x0.b0, dscal[H_C + i] = x0.b_{i+1})
__global__ void k_hist_push(int keep, double* __restrict__ G, const double* __restrict__ dscal) {
  for (int i = keep; i >= 1; --i)
    for (int j = keep; j >= 1; --j) G[i * MV_HIST_MAX + j] = G[(i - 1) * MV_HIST_MAX + (j - 1)];
  G[0] = dscal[H_S0];
  for (int i = 1; i <= keep; ++i) G[i] = G[i * MV_HIST_MAX] = dscal[H_C + i - 1];
}

#define HIST_DISPATCH(m, ...)                                              \
  switch (m) {                                                             \
    case 0: { constexpr int M = 0; __VA_ARGS__; } break;                   \
    case 1: { constexpr int M = 1; __VA_ARGS__; } break;                   \
    case 2: { constexpr int M = 2; __VA_ARGS__; } break;                   \
    case 3: { constexpr int M = 3; __VA_ARGS__; } break;                   \
    case 4: { constexpr int M = 4; __VA_ARGS__; } break;                   \
    case 5: { constexpr int M = 5; __VA_ARGS__; } break;                   \
    case 6: { constexpr int M = 6; __VA_ARGS__; } break;                   \
    case 7: { constexpr int M = 7; __VA_ARGS__; } break;                   \
    default: { constexpr int M = 8; __VA_ARGS__; } break;                  \
  }

static int g_poisson_history = -1;   // -1: read MV_POISSON_HISTORY (default on); otherwise the depth (0 = off)
extern "C" int mv_set_poisson_history(int on) {
  g_poisson_history = on <= 0 ? 0 : (on == 1 ? MV_HIST_DEFAULT : (on > MV_HIST_MAX ? MV_HIST_MAX : on));
  return MV_OK;
}

extern "C" int mv_ma_pseudotimestep(mv_ma* ma, double dt, int* iters) {
  if (!ma) return MV_ERR_BADARG;
  MA_REQUIRE(ma, MV_F_PHI);
  MA_REQUIRE(ma, MV_F_PHI_OLD);
  MA_REQUIRE(ma, MV_F_RESIDUAL);
  mv_mesh* m = ma->mesh;
  const int64_t n = m->Vown;
  const int gv = grid_for(n);
  MV_L k_scale<<<gv, MV_BLOCK, 0, g_stream>>>(n, dt, ma->f[MV_F_RESIDUAL], ma->rhs);
  if (g_poisson_history < 0) { const char* e = getenv("MV_POISSON_HISTORY"); mv_set_poisson_history(e ? atoi(e) : 1); }
  const int depth = g_poisson_history;
  HistPtrs hp;
  const int mh = ma->hist_n < depth ? ma->hist_n : depth;   // stored pairs in use (the newest mh)
  for (int i = 0; i < MV_HIST_MAX; ++i) { hp.x[i] = i < mh ? ma->hist_x[i] : nullptr; hp.b[i] = i < mh ? ma->hist_b[i] : nullptr; }
  const bool dist = comm_active();
  if (depth > 0 && !ma->hist_G) {
    MV_CHECK(dev_alloc(&ma->hist_G, MV_HIST_MAX * MV_HIST_MAX));
    MV_CUDA(cudaMemsetAsync(ma->hist_G, 0, MV_HIST_MAX * MV_HIST_MAX * sizeof(double), g_stream));
  }
  if (depth > 0 && mh > 0) {
    HIST_DISPATCH(mh, MV_L k_hist_dots<M><<<gv, MV_BLOCK, 0, g_stream>>>(n, hp, 0, ma->rhs, nullptr, m->red.partials, m->red.counter, ma->dscal));
    if (dist) MV_CHECK(allreduce_sum_dev(ma->dscal + H_C, mh));
    MV_L k_hist_solve<<<1, 1, 0, g_stream>>>(mh, ma->hist_G, ma->dscal);
    HIST_DISPATCH(mh, MV_L k_hist_guess<M><<<gv, MV_BLOCK, 0, g_stream>>>(n, hp, ma->dscal, ma->dphi));
    MV_KERNEL_CHECK();
  } else {
    MV_CUDA(cudaMemsetAsync(ma->dphi, 0, m->V * sizeof(double), g_stream));
  }
  MV_CHECK(pcg_solve(m, MV_OP_STIFFNESS, 1, ma->rhs, ma->dphi, nullptr, ma->rtol, ma->atol, ma->maxit_poisson, 1, iters,
                     nullptr));
  MV_L k_add<<<gv, MV_BLOCK, 0, g_stream>>>(n, ma->f[MV_F_PHI_OLD], ma->dphi, ma->f[MV_F_PHI]);
  MV_KERNEL_CHECK();
  if (depth > 0) {
    // the new pair takes the place of the oldest one (slot depth-1 -> slot 0); its row of G from the stored right-hand sides
    for (int i = 0; i < depth; ++i)
      if (!ma->hist_x[i]) {
        MV_CHECK(dev_alloc(&ma->hist_x[i], n));
        MV_CHECK(dev_alloc(&ma->hist_b[i], n));
      }
    const int keep = mh < depth - 1 ? mh : depth - 1;   // older pairs that stay
    for (int i = 0; i < MV_HIST_MAX; ++i) { if (i >= keep) { hp.x[i] = nullptr; hp.b[i] = nullptr; } }
    // x_new . b_i (i kept) and x_new . b_new; x_new has zero mean, so the mean of the right-hand sides drops out
    HIST_DISPATCH(keep, MV_L k_hist_dots<M><<<gv, MV_BLOCK, 0, g_stream>>>(n, hp, 1, ma->dphi, ma->rhs, m->red.partials, m->red.counter, ma->dscal));
    if (dist) MV_CHECK(allreduce_sum_dev(ma->dscal + H_C, MV_HIST_MAX + 1));
    MV_L k_hist_push<<<1, 1, 0, g_stream>>>(keep, ma->hist_G, ma->dscal);
    double* xn = ma->hist_x[depth - 1];
    double* bn = ma->hist_b[depth - 1];
    for (int i = depth - 1; i > 0; --i) { ma->hist_x[i] = ma->hist_x[i - 1]; ma->hist_b[i] = ma->hist_b[i - 1]; }
    ma->hist_x[0] = xn;
    ma->hist_b[0] = bn;
    MV_CUDA(cudaMemcpyAsync(xn, ma->dphi, n * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
    MV_CUDA(cudaMemcpyAsync(bn, ma->rhs, n * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
    MV_KERNEL_CHECK();
    if (ma->hist_n < depth) ++ma->hist_n;
  }
  return MV_OK;
}

extern "C" int mv_ma_equidistribute(mv_ma* ma, int* iters) {
  if (!ma) return MV_ERR_BADARG;
  MA_REQUIRE(ma, MV_F_PHI);
  MA_REQUIRE(ma, MV_F_H);
  mv_mesh* m = ma->mesh;
  if (comm_active()) MV_CHECK(halo_exchange(m, 1, ma->f[MV_F_PHI]));
  double* l2_out = nullptr;
  if (ma->fuse_l2_rhs) {
    if (!ma->l2_rhs) MV_CHECK(dev_alloc(&ma->l2_rhs, (size_t)m->V * m->dim));
    l2_out = ma->l2_rhs;
  }
  MV_CHECK(hess_rhs_dev(m, ma->f[MV_F_PHI], ma->rhs, l2_out));
  ma->l2_rhs_phi = l2_out ? ma->f[MV_F_PHI] : nullptr;
  MV_CHECK(pcg_solve(m, MV_OP_MASS, m->dim * m->dim, ma->rhs, ma->f[MV_F_H], nullptr, ma->rtol, ma->atol, ma->maxit_mass, 0,
                     iters, nullptr));
  return MV_OK;
}

extern "C" int mv_ma_advance(mv_ma* ma) {
  if (!ma) return MV_ERR_BADARG;
  MA_REQUIRE(ma, MV_F_PHI);
  MA_REQUIRE(ma, MV_F_PHI_OLD);
  MA_REQUIRE(ma, MV_F_H);
  MA_REQUIRE(ma, MV_F_H_OLD);
  mv_mesh* m = ma->mesh;
  const int dd = m->dim * m->dim;
  MV_CUDA(cudaMemcpyAsync(ma->f[MV_F_PHI_OLD], ma->f[MV_F_PHI], m->V * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
  MV_CUDA(cudaMemcpyAsync(ma->f[MV_F_H_OLD], ma->f[MV_F_H], m->V * dd * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
  return MV_OK;
}

extern "C" int mv_ma_relax_iterate(mv_ma* ma, double dt, int niter, double* trace, int* cg_iters) {
  if (!ma) return MV_ERR_BADARG;
  if (ma->mon_kind == MV_MON_USER) {
    set_error("mv_ma_relax_iterate needs a built-in monitor");
    return MV_ERR_BADARG;
  }
  for (int i = 0; i < niter; ++i) {
    int it[3] = {0, 0, 0};
    double out[6];
    MV_CHECK(mv_ma_l2_project(ma, &it[0]));
    MV_CHECK(mv_ma_update_coordinates(ma));
    MV_CHECK(mv_ma_diagnostics(ma, 1, out));
    MV_CHECK(mv_ma_pseudotimestep(ma, dt, &it[1]));
    MV_CHECK(mv_ma_equidistribute(ma, &it[2]));
    MV_CHECK(mv_ma_advance(ma));
    if (trace) for (int k = 0; k < 6; ++k) trace[i * 6 + k] = out[k];
    if (cg_iters) for (int k = 0; k < 3; ++k) cg_iters[i * 3 + k] = it[k];
  }
  return MV_OK;
}

// The whole relaxation loop of movement/monge_ampere.py:488-540 behind one call: iterations i0, i0+1, ... until the
// residual test passes, the divergence test fires, `maxiter` is reached or `nmax` iterations were done in this call
// (the caller prints the trace lines of a chunk and calls again).  phi/phi_old and H/H_old are advanced by swapping
// the bound pointers instead of copying 8(1+d^2) B per vertex; on return all four user buffers hold what the
// reference's fields hold at that point (phi == phi_old, H == H_old after an advance).
//   res0: in/out initial residual (set at i == 0);  info[0] = last iteration index i, info[1] = status
//   (0 converged, 1 diverged, 2 maxiter reached, 3 chunk exhausted -- call again with i0 = info[0] + 1).
extern "C" int mv_ma_relax_move(mv_ma* ma, double dt, int i0, int nmax, int maxiter, double rtol, double dtol, double* res0,
                                double* trace, int* cg_iters, int* info) {
  if (!ma || !res0 || !info || nmax < 1) return MV_ERR_BADARG;
  if (ma->mon_kind == MV_MON_USER) {
    set_error("mv_ma_relax_move needs a built-in monitor");
    return MV_ERR_BADARG;
  }
  MA_REQUIRE(ma, MV_F_PHI); MA_REQUIRE(ma, MV_F_PHI_OLD); MA_REQUIRE(ma, MV_F_H); MA_REQUIRE(ma, MV_F_H_OLD);
  mv_mesh* m = ma->mesh;
  const int dd = m->dim * m->dim;
  double* const user_phi = ma->f[MV_F_PHI];
  double* const user_phi_old = ma->f[MV_F_PHI_OLD];
  double* const user_H = ma->f[MV_F_H];
  double* const user_H_old = ma->f[MV_F_H_OLD];
  int status = 3, i = i0, st = MV_OK, done = 0;
  bool advanced = false;
  static int fuse_env = -1;
  if (fuse_env < 0) { const char* e = getenv("MV_FUSE_L2_RHS"); fuse_env = e ? atoi(e) : 1; }
  ma->fuse_l2_rhs = fuse_env != 0;
  ma->l2_rhs_valid = false;
  for (int k = 0; k < nmax && i < maxiter; ++k, ++i) {
    int it[3] = {0, 0, 0};
    double out[6];
    if ((st = mv_ma_l2_project(ma, &it[0])) != MV_OK) break;
    if ((st = mv_ma_update_coordinates(ma)) != MV_OK) break;
    if ((st = mv_ma_diagnostics(ma, 1, out)) != MV_OK) break;
    if (trace) for (int q = 0; q < 6; ++q) trace[k * 6 + q] = out[q];
    ++done;
    const double residual = out[1];
    if (i == 0) *res0 = residual;
    if (residual < rtol) { status = 0; }
    else if (residual > dtol * *res0) { status = 1; }
    else if (i == maxiter - 1) { status = 2; }
    if (status != 3) {
      if (cg_iters) { cg_iters[k * 3] = it[0]; cg_iters[k * 3 + 1] = 0; cg_iters[k * 3 + 2] = 0; }
      break;
    }
    if ((st = mv_ma_pseudotimestep(ma, dt, &it[1])) != MV_OK) break;
    if ((st = mv_ma_equidistribute(ma, &it[2])) != MV_OK) break;
    // advance (monge_ampere.py:537-538) by exchanging the roles of the buffers
    double* t = ma->f[MV_F_PHI]; ma->f[MV_F_PHI] = ma->f[MV_F_PHI_OLD]; ma->f[MV_F_PHI_OLD] = t;
    t = ma->f[MV_F_H]; ma->f[MV_F_H] = ma->f[MV_F_H_OLD]; ma->f[MV_F_H_OLD] = t;
    advanced = true;
    ma->l2_rhs_valid = ma->fuse_l2_rhs && ma->l2_rhs_phi == ma->f[MV_F_PHI_OLD];   // the buffer the Hessian pass read is phi_old now
    if (cg_iters) for (int q = 0; q < 3; ++q) cg_iters[k * 3 + q] = it[q];
  }
  ma->fuse_l2_rhs = false;      // between calls the caller may change phi: the next call starts from k_l2_rhs again
  ma->l2_rhs_valid = false;
  if (status == 3 && st == MV_OK) --i;   // chunk exhausted: i is the last completed iteration
  // hand the fields back: the newest iterate sits in f[*_OLD] after an advance; both user buffers get it
  const double* newest_phi = ma->f[MV_F_PHI_OLD];
  const double* newest_H = ma->f[MV_F_H_OLD];
  ma->f[MV_F_PHI] = user_phi; ma->f[MV_F_PHI_OLD] = user_phi_old; ma->f[MV_F_H] = user_H; ma->f[MV_F_H_OLD] = user_H_old;
  if (advanced) {
    if (newest_phi != user_phi_old) MV_CUDA(cudaMemcpyAsync(user_phi_old, newest_phi, m->V * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
    else MV_CUDA(cudaMemcpyAsync(user_phi, newest_phi, m->V * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
    if (newest_H != user_H_old) MV_CUDA(cudaMemcpyAsync(user_H_old, newest_H, m->V * dd * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
    else MV_CUDA(cudaMemcpyAsync(user_H, newest_H, m->V * dd * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
  }
  info[0] = i;
  info[1] = status;
  info[2] = done;
  return st;
}
