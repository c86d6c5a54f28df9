// Smoothed-aggregation algebraic multigrid: hierarchy SETUP, entirely on the device.
//
// replaces: PETSc `pc_type gamg` on the Poisson pseudo-timestep
//           (movement/solver_parameters.py:68-71, movement/monge_ampere.py:441-448), the
//           ILU-preconditioned KSP of movement/laplacian_smoothing.py:55-77 and the gamg
//           sub-preconditioner of the quasi-Newton fieldsplit (solver_parameters.py:26-33).
//
// Every level keeps the layout of the fine level (SELL-32 operator, owned rows first) so the
// solve phase re-uses the same streaming kernels.  Set-up pipeline per level, all data parallel
// and deterministic (integer atomics only; every floating-point sum has a fixed order):
//   1. strength graph   |a_ij| > theta sqrt(a_ii a_jj), both ends unconstrained and owned
//   2. distance-2 maximal independent set, Luby rounds with a hashed priority     -> roots
//   3. aggregates       root + strong neighbours; the rest join their strongest aggregated neighbour
//   4. prolongator      P = (I - omega D'^-1 A_loc) T, omega = 4/(3 rho); couplings to ghost
//                       vertices are lumped into D' so P stays rank-local and P 1 = 1
//   5. Galerkin         A_c = P^T (A P): two row-wise sparse products with per-thread sorted
//                       accumulators (symbolic count -> scan -> numeric fill)
//   6. rho(D^-1 A)      power iteration, 12 steps, x1.1
// The coarsest operator (<= MV_AMG_DENSE rows) is inverted densely by Gauss-Jordan in one CTA.
// oracle/amg_prototype.py is the numpy statement of the same algorithm.
#include <algorithm>
#include <cub/device/device_radix_sort.cuh>

#include "mv_internal.cuh"

#define MV_AMG_DENSE 1024       // coarsest level is inverted densely when it has at most this many rows
#define MV_AMG_MAX_LEVELS 12
#define MV_AMG_ACC 768          // per-thread accumulator capacity (distinct columns of one product row)

namespace mv {

__global__ void k_to_float(int64_t n, const double* __restrict__ a, float* __restrict__ b) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) b[i] = (float)a[i];
}

void Csr::release() {
  dev_free(ptr);
  dev_free(col);
  dev_free(val);
  nrows = ncols = nnz = 0;
}

static void sell_free(Sell& s) {
  dev_free(s.slice_ptr);
  dev_free(s.cols);
  dev_free(s.row_nnz);
}

Amg::~Amg() {
  for (size_t l = 0; l < lv.size(); ++l) {
    AmgLevel& L = lv[l];
    if (L.owns_A) {
      sell_free(L.pat);
      double* v = const_cast<double*>(L.val);
      dev_free(v);
    }
    dev_free(L.dinv);
    dev_free(L.valf);
    sell_free(L.P);
    sell_free(L.R);
    dev_free(L.Pval);
    dev_free(L.Rval);
    L.Rc.release();
    L.Ac.release();
    dev_free(L.x0);
    if (l > 0) dev_free(L.b);
    dev_free(L.x1); dev_free(L.r); dev_free(L.d0); dev_free(L.d1);
    if (L.owns_halo) { dev_free(L.halo.send_idx); dev_free(L.halo.send_buf); }
    dev_free(L.gid);
  }
  dev_free(cglob);
  dev_free(mask_copy);
  dev_free(coarse_inv);
  dev_free(z);
}

void amg_release(mv_mesh* m) {
  for (Amg* a : m->amg) delete a;
  m->amg.clear();
}

// ------------------------------------------------------------------ small helpers
__device__ __forceinline__ unsigned hash32(unsigned x) {
  x += 0x9E3779B9u;
  x ^= x >> 16; x *= 0x85EBCA6Bu;
  x ^= x >> 16; x *= 0xC2B2AE35u;
  x ^= x >> 16;
  return x;
}
__device__ __forceinline__ unsigned long long mis_priority(int64_t i) {
  return ((unsigned long long)hash32((unsigned)i) << 32) | (unsigned long long)(i + 1);
}

// Walk the stored entries of SELL row `row`: f(col, value, k).
template <typename F>
__device__ __forceinline__ void sell_row(const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
                                         const double* __restrict__ val, const int* __restrict__ row_nnz, int64_t row, F f) {
  const int64_t base = slice_ptr[row >> 5] + (row & 31);
  const int n = row_nnz[row];
  for (int k = 0; k < n; ++k) f(cols[base + (int64_t)k * MV_SLICE], val ? val[base + (int64_t)k * MV_SLICE] : 0.0, k);
}

#define GRID_STRIDE(i, n) for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < (n); i += (int64_t)gridDim.x * blockDim.x)

__global__ void k_amg_diag(int64_t n, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
                           const double* __restrict__ val, const int* __restrict__ row_nnz, const uint16_t* __restrict__ mask,
                           double* __restrict__ diag, double* __restrict__ dinv) {
  GRID_STRIDE(i, n) {
    double d = 0.0;
    sell_row(slice_ptr, cols, val, row_nnz, i, [&](int c, double a, int) { if (c == i) d += a; });
    diag[i] = d;
    const bool fix = mask && (mask[i] & 1);
    dinv[i] = (!fix && d != 0.0) ? 1.0 / d : 0.0;
  }
}

// ------------------------------------------------------------------ MIS-2 (Luby) on the strength graph
struct Graph {
  int64_t n;
  const int64_t* slice_ptr;
  const int* cols;
  const double* val;
  const int* row_nnz;
  const double* diag;
  const uint16_t* mask;
  double theta2;   // theta^2
};
__device__ __forceinline__ bool g_active(const Graph& g, int64_t i) { return !(g.mask && (g.mask[i] & 1)); }
__device__ __forceinline__ bool g_strong(const Graph& g, int64_t i, int j, double a) {
  if (j == i || j >= g.n || !g_active(g, j)) return false;
  return a * a > g.theta2 * fabs(g.diag[i] * g.diag[j]);
}

enum { MIS_UND = 0, MIS_IN = 1, MIS_OUT = 2 };

__global__ void k_mis_init(Graph g, int* __restrict__ state) {
  GRID_STRIDE(i, g.n) state[i] = g_active(g, i) ? MIS_UND : MIS_OUT;
}
// t1[i] = max priority of undecided vertices in the closed neighbourhood of i
__global__ void k_mis_max1(Graph g, const int* __restrict__ state, unsigned long long* __restrict__ t1) {
  GRID_STRIDE(i, g.n) {
    unsigned long long m = 0;
    if (g_active(g, i)) {
      if (state[i] == MIS_UND) m = mis_priority(i);
      sell_row(g.slice_ptr, g.cols, g.val, g.row_nnz, i, [&](int j, double a, int) {
        if (g_strong(g, i, j, a) && state[j] == MIS_UND) {
          const unsigned long long p = mis_priority(j);
          m = p > m ? p : m;
        }
      });
    }
    t1[i] = m;
  }
}
__global__ void k_mis_max2(Graph g, int* __restrict__ state, const unsigned long long* __restrict__ t1) {
  GRID_STRIDE(i, g.n) {
    if (state[i] != MIS_UND) continue;
    unsigned long long m = t1[i];
    sell_row(g.slice_ptr, g.cols, g.val, g.row_nnz, i, [&](int j, double a, int) {
      if (g_strong(g, i, j, a)) m = t1[j] > m ? t1[j] : m;
    });
    if (m == mis_priority(i)) state[i] = MIS_IN;   // no other thread reads state[] in this kernel
  }
}
__global__ void k_mis_near1(Graph g, const int* __restrict__ state, int* __restrict__ f1) {
  GRID_STRIDE(i, g.n) {
    int f = state[i] == MIS_IN;
    if (!f && g_active(g, i))
      sell_row(g.slice_ptr, g.cols, g.val, g.row_nnz, i, [&](int j, double a, int) {
        if (g_strong(g, i, j, a) && state[j] == MIS_IN) f = 1;
      });
    f1[i] = f;
  }
}
__global__ void k_mis_near2(Graph g, int* __restrict__ state, const int* __restrict__ f1, int* __restrict__ remaining) {
  int left = 0;
  GRID_STRIDE(i, g.n) {
    if (state[i] != MIS_UND) continue;
    int f = f1[i];
    if (!f)
      sell_row(g.slice_ptr, g.cols, g.val, g.row_nnz, i, [&](int j, double a, int) {
        if (g_strong(g, i, j, a) && f1[j]) f = 1;
      });
    if (f) state[i] = MIS_OUT; else ++left;
  }
  if (left) atomicAdd(remaining, left);
}

// ------------------------------------------------------------------ aggregates
__global__ void k_root_flag(int64_t n, const int* __restrict__ state, int64_t* __restrict__ flag) {
  GRID_STRIDE(i, n + 1) flag[i] = (i < n && state[i] == MIS_IN) ? 1 : 0;
}
__global__ void k_agg_roots(int64_t n, const int* __restrict__ state, const int64_t* __restrict__ rank, int* __restrict__ agg) {
  GRID_STRIDE(i, n) agg[i] = state[i] == MIS_IN ? (int)rank[i] : -1;
}
// neighbours of a root (at most one root is adjacent to a vertex)
__global__ void k_agg_pass1(Graph g, const int* __restrict__ state, const int* __restrict__ agg0, int* __restrict__ agg1) {
  GRID_STRIDE(i, g.n) {
    int a0 = agg0[i];
    if (a0 < 0 && g_active(g, i))
      sell_row(g.slice_ptr, g.cols, g.val, g.row_nnz, i, [&](int j, double a, int) {
        if (g_strong(g, i, j, a) && state[j] == MIS_IN) a0 = max(a0, agg0[j]);
      });
    agg1[i] = a0;
  }
}
// the rest: aggregate of the strongest neighbour aggregated in pass 1 (ties -> lowest index)
__global__ void k_agg_pass2(Graph g, const int* __restrict__ agg1, int* __restrict__ agg, int* __restrict__ orphans) {
  int lost = 0;
  GRID_STRIDE(i, g.n) {
    int a1 = agg1[i];
    if (a1 < 0 && g_active(g, i)) {
      double best = -1.0;
      sell_row(g.slice_ptr, g.cols, g.val, g.row_nnz, i, [&](int j, double a, int) {
        if (g_strong(g, i, j, a) && agg1[j] >= 0 && fabs(a) > best) { best = fabs(a); a1 = agg1[j]; }
      });
      if (a1 < 0) ++lost;
    }
    agg[i] = a1;
  }
  if (lost) atomicAdd(orphans, lost);
}


// ------------------------------------------------------------------ per-thread sorted accumulator
template <int CAP>
struct RowAcc {
  int key[CAP];
  double val[CAP];
  int n = 0;
  bool overflow = false;
  __device__ __forceinline__ void add(int k, double v) {
    int lo = 0, hi = n;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (key[mid] < k) lo = mid + 1; else hi = mid;
    }
    if (lo < n && key[lo] == k) { val[lo] += v; return; }
    if (n >= CAP) { overflow = true; return; }
    for (int t = n; t > lo; --t) { key[t] = key[t - 1]; val[t] = val[t - 1]; }
    key[lo] = k;
    val[lo] = v;
    ++n;
  }
  template <bool FILL>
  __device__ __forceinline__ void emit(int64_t row, int64_t* __restrict__ ptr_or_cnt, int* __restrict__ col, double* __restrict__ v,
                                       int* __restrict__ err) {
    if (overflow) atomicExch(err, 1);
    if (!FILL) { ptr_or_cnt[row] = n; return; }
    const int64_t b = ptr_or_cnt[row];
    for (int t = 0; t < n; ++t) { col[b + t] = key[t]; v[b + t] = val[t]; }
  }
};

// ------------------------------------------------------------------ prolongator P = (I - omega D'^-1 A_f) T
template <bool FILL>
__global__ void k_prolongator(Graph g, const int* __restrict__ agg, double omega, int64_t* __restrict__ ptr_or_cnt,
                              int* __restrict__ col, double* __restrict__ val, int* __restrict__ err) {
  GRID_STRIDE(i, g.n) {
    RowAcc<64> acc;
    const int ai = agg[i];
    if (ai >= 0) {
      double dl = g.diag[i];   // diagonal + lumped weak / off-rank couplings
      sell_row(g.slice_ptr, g.cols, g.val, g.row_nnz, i, [&](int j, double a, int) {
        if (j == i) return;
        if (j >= g.n) { dl += a; return; }            // ghost column: lumped, keeps P rank-local with P 1 = 1
        if (!g_active(g, j)) return;                   // constrained neighbour: belongs to the eliminated block
        if (g_strong(g, i, j, a) && agg[j] >= 0) acc.add(agg[j], a); else dl += a;
      });
      acc.add(ai, dl);
      const double s = dl != 0.0 ? omega / dl : 0.0;
      for (int t = 0; t < acc.n; ++t) acc.val[t] = (acc.key[t] == ai ? 1.0 : 0.0) - s * acc.val[t];
    }
    acc.emit<FILL>(i, ptr_or_cnt, col, val, err);
  }
}

// ------------------------------------------------------------------ transpose (CSR -> CSR)
__global__ void k_tr_count(int64_t nnz, const int* __restrict__ col, unsigned long long* __restrict__ cnt) {
  GRID_STRIDE(t, nnz) atomicAdd(&cnt[col[t]], 1ull);
}
__global__ void k_tr_fill(int64_t nrows, const int64_t* __restrict__ ptr, const int* __restrict__ col, const double* __restrict__ val,
                          const int64_t* __restrict__ tptr, unsigned long long* __restrict__ cursor, int* __restrict__ tcol,
                          double* __restrict__ tval) {
  GRID_STRIDE(i, nrows) {
    for (int64_t t = ptr[i]; t < ptr[i + 1]; ++t) {
      const int c = col[t];
      const int64_t pos = tptr[c] + (int64_t)atomicAdd(&cursor[c], 1ull);
      tcol[pos] = (int)i;
      tval[pos] = val[t];
    }
  }
}
__global__ void k_csr_sort_rows(int64_t nrows, const int64_t* __restrict__ ptr, int* __restrict__ col, double* __restrict__ val) {
  GRID_STRIDE(r, nrows) {
    const int64_t b = ptr[r], e = ptr[r + 1];
    for (int64_t i = b + 1; i < e; ++i) {
      const int kc = col[i];
      const double kv = val[i];
      int64_t j = i - 1;
      while (j >= b && col[j] > kc) { col[j + 1] = col[j]; val[j + 1] = val[j]; --j; }
      col[j + 1] = kc;
      val[j + 1] = kv;
    }
  }
}

// ------------------------------------------------------------------ row-wise sparse products
// C = A B with A in SELL (rows with mask bit 0 set are skipped) and B in CSR
template <bool FILL>
__global__ void k_spgemm_sell_csr(int64_t n, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
                                  const double* __restrict__ aval, const int* __restrict__ row_nnz, const uint16_t* __restrict__ mask,
                                  const int64_t* __restrict__ bptr, const int* __restrict__ bcol, const double* __restrict__ bval,
                                  int64_t brows, int64_t* __restrict__ ptr_or_cnt, int* __restrict__ col, double* __restrict__ val,
                                  int* __restrict__ err) {
  GRID_STRIDE(i, n) {
    RowAcc<MV_AMG_ACC> acc;
    if (!(mask && (mask[i] & 1)))
      sell_row(slice_ptr, cols, aval, row_nnz, i, [&](int j, double a, int) {
        if (a == 0.0 || j >= brows) return;
        for (int64_t t = bptr[j]; t < bptr[j + 1]; ++t) acc.add(bcol[t], a * bval[t]);
      });
    acc.emit<FILL>(i, ptr_or_cnt, col, val, err);
  }
}
// C = A B, both CSR
template <bool FILL>
__global__ void k_spgemm_csr_csr(int64_t n, const int64_t* __restrict__ aptr, const int* __restrict__ acol,
                                 const double* __restrict__ aval, const int64_t* __restrict__ bptr, const int* __restrict__ bcol,
                                 const double* __restrict__ bval, int64_t* __restrict__ ptr_or_cnt, int* __restrict__ col,
                                 double* __restrict__ val, int* __restrict__ err) {
  GRID_STRIDE(i, n) {
    RowAcc<MV_AMG_ACC> acc;
    for (int64_t s = aptr[i]; s < aptr[i + 1]; ++s) {
      const int j = acol[s];
      const double a = aval[s];
      for (int64_t t = bptr[j]; t < bptr[j + 1]; ++t) acc.add(bcol[t], a * bval[t]);
    }
    acc.emit<FILL>(i, ptr_or_cnt, col, val, err);
  }
}

// ------------------------------------------------------------------ CSR -> SELL-32
__global__ void k_csr_row_nnz(int64_t nrows, const int64_t* __restrict__ ptr, int* __restrict__ row_nnz) {
  GRID_STRIDE(i, nrows) row_nnz[i] = (int)(ptr[i + 1] - ptr[i]);
}
__global__ void k_amg_slice_width(const int* __restrict__ row_nnz, int64_t nrows, int64_t nslices, int64_t* __restrict__ slice_len,
                                  int* __restrict__ maxw) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t s = warp; s < nslices; s += nwarps) {
    const int64_t row = s * MV_SLICE + lane;
    int w = row < nrows ? row_nnz[row] : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) w = max(w, __shfl_xor_sync(0xffffffffu, w, o));
    if (lane == 0) { slice_len[s] = (int64_t)w * MV_SLICE; atomicMax(maxw, w); }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) slice_len[nslices] = 0;
}
// padding: value 0, column = the row's own index when that is a valid column, else its first column (or 0)
__global__ void k_csr_to_sell(int64_t nrows, int64_t ncols, int64_t nslices, const int64_t* __restrict__ ptr, const int* __restrict__ col,
                              const double* __restrict__ val, const int64_t* __restrict__ slice_ptr, int* __restrict__ scols,
                              double* __restrict__ svals) {
  GRID_STRIDE(v, nslices * MV_SLICE) {
    const int64_t s = v >> 5;
    const int lane = (int)(v & 31);
    const int64_t base = slice_ptr[s];
    const int w = (int)((slice_ptr[s + 1] - base) >> 5);
    int64_t b = 0, e = 0;
    if (v < nrows) { b = ptr[v]; e = ptr[v + 1]; }
    int pad = 0;
    if (v < nrows && v < ncols) pad = (int)v; else if (e > b) pad = col[b];
    for (int k = 0; k < w; ++k) {
      const bool in = b + k < e;
      scols[base + (int64_t)k * MV_SLICE + lane] = in ? col[b + k] : pad;
      svals[base + (int64_t)k * MV_SLICE + lane] = in ? val[b + k] : 0.0;
    }
  }
}

static int csr_to_sell(const Csr& c, Sell& s, double** sval) {
  s.nrows = c.nrows;
  s.nslices = (c.nrows + MV_SLICE - 1) / MV_SLICE;
  s.nnz = c.nnz;
  MV_CHECK(dev_alloc(&s.row_nnz, c.nrows));
  MV_CHECK(dev_alloc(&s.slice_ptr, s.nslices + 1));
  int64_t* slice_len = nullptr;
  int* maxw = nullptr;
  MV_CHECK(dev_alloc(&slice_len, s.nslices + 1));
  MV_CHECK(dev_alloc(&maxw, 1));
  MV_CUDA(cudaMemsetAsync(maxw, 0, sizeof(int), g_stream));
  MV_L k_csr_row_nnz<<<grid_for(c.nrows), MV_BLOCK, 0, g_stream>>>(c.nrows, c.ptr, s.row_nnz);
  MV_L k_amg_slice_width<<<grid_for(s.nslices * 32), MV_BLOCK, 0, g_stream>>>(s.row_nnz, c.nrows, s.nslices, slice_len, maxw);
  MV_KERNEL_CHECK();
  MV_CHECK(exclusive_scan_i64(slice_len, s.slice_ptr, s.nslices + 1));
  MV_CUDA(cudaMemcpyAsync(&s.len, s.slice_ptr + s.nslices, sizeof(int64_t), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaMemcpyAsync(&s.maxw, maxw, sizeof(int), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  dev_free(slice_len);
  dev_free(maxw);
  MV_CHECK(dev_alloc(&s.cols, s.len));
  MV_CHECK(dev_alloc(sval, s.len));
  MV_L k_csr_to_sell<<<grid_for(s.nslices * 32), MV_BLOCK, 0, g_stream>>>(c.nrows, c.ncols, s.nslices, c.ptr, c.col, c.val, s.slice_ptr,
                                                                     s.cols, *sval);
  MV_KERNEL_CHECK();
  return MV_OK;
}

// counts[nrows] (+1 slot) -> ptr by exclusive scan; allocates col/val
static int csr_finish_symbolic(Csr& c, int64_t* counts) {
  MV_CUDA(cudaMemsetAsync(counts + c.nrows, 0, sizeof(int64_t), g_stream));
  MV_CHECK(dev_alloc(&c.ptr, c.nrows + 1));
  MV_CHECK(exclusive_scan_i64(counts, c.ptr, c.nrows + 1));
  MV_CUDA(cudaMemcpyAsync(&c.nnz, c.ptr + c.nrows, sizeof(int64_t), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  MV_CHECK(dev_alloc(&c.col, c.nnz));
  MV_CHECK(dev_alloc(&c.val, c.nnz));
  return MV_OK;
}

// (column << 32 | row) keys of every stored entry, in storage order
__global__ void k_tr_keys(int64_t nrows, const int64_t* __restrict__ ptr, const int* __restrict__ col,
                          unsigned long long* __restrict__ key) {
  GRID_STRIDE(i, nrows)
    for (int64_t t = ptr[i]; t < ptr[i + 1]; ++t) key[t] = ((unsigned long long)(unsigned)col[t] << 32) | (unsigned long long)(unsigned)i;
}
__global__ void k_tr_unpack(int64_t nnz, const unsigned long long* __restrict__ key, int* __restrict__ tcol) {
  GRID_STRIDE(t, nnz) tcol[t] = (int)(key[t] & 0xffffffffull);
}

// Transpose by one radix sort of (column, row) keys: rows of the result come out sorted, no atomics on values.
static int csr_transpose(const Csr& a, Csr& t) {
  t.nrows = a.ncols;
  t.ncols = a.nrows;
  unsigned long long* cnt = nullptr;
  MV_CHECK(dev_alloc(&cnt, t.nrows + 1));
  MV_CUDA(cudaMemsetAsync(cnt, 0, (t.nrows + 1) * sizeof(unsigned long long), g_stream));
  if (a.nnz) MV_L k_tr_count<<<grid_for(a.nnz), MV_BLOCK, 0, g_stream>>>(a.nnz, a.col, cnt);
  MV_KERNEL_CHECK();
  MV_CHECK(csr_finish_symbolic(t, (int64_t*)cnt));
  dev_free(cnt);
  if (a.nnz == 0) return MV_OK;
  unsigned long long *k0 = nullptr, *k1 = nullptr;
  MV_CHECK(dev_alloc(&k0, a.nnz));
  MV_CHECK(dev_alloc(&k1, a.nnz));
  MV_L k_tr_keys<<<grid_for(a.nrows), MV_BLOCK, 0, g_stream>>>(a.nrows, a.ptr, a.col, k0);
  MV_KERNEL_CHECK();
  int bits_row = 1, bits_col = 1;
  while ((1ll << bits_row) < a.nrows) ++bits_row;
  while ((1ll << bits_col) < a.ncols) ++bits_col;
  void* tmp = nullptr;
  size_t bytes = 0;
  MV_CUDA(cub::DeviceRadixSort::SortPairs(tmp, bytes, k0, k1, a.val, t.val, (int)a.nnz, 0, 32 + bits_col, g_stream));
  MV_CUDA(cudaMallocAsync(&tmp, bytes, g_stream));
  MV_CUDA(cub::DeviceRadixSort::SortPairs(tmp, bytes, k0, k1, a.val, t.val, (int)a.nnz, 0, 32 + bits_col, g_stream));
  MV_L k_tr_unpack<<<grid_for(a.nnz), MV_BLOCK, 0, g_stream>>>(a.nnz, k1, t.col);
  MV_KERNEL_CHECK();
  MV_CUDA(cudaStreamSynchronize(g_stream));
  cudaFreeAsync(tmp, g_stream);
  dev_free(k0); dev_free(k1);
  (void)bits_row;
  return MV_OK;
}

// ------------------------------------------------------------------ rho(D^-1 A) by power iteration
__global__ void k_pow_init(int64_t n, const double* __restrict__ dinv, double* __restrict__ x) {
  GRID_STRIDE(i, n) x[i] = dinv[i] != 0.0 ? (double)hash32((unsigned)i) * (1.0 / 4294967296.0) - 0.5 : 0.0;
}
__global__ void __launch_bounds__(MV_BLOCK)
k_pow_step(int64_t n, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols, const double* __restrict__ val,
           const int* __restrict__ row_nnz, const double* __restrict__ dinv, const double* __restrict__ x, double* __restrict__ y,
           double* __restrict__ partials, unsigned* __restrict__ counter, double* __restrict__ out) {
  double acc[2] = {0.0, 0.0};
  GRID_STRIDE(i, n) {
    double t = 0.0;
    sell_row(slice_ptr, cols, val, row_nnz, i, [&](int j, double a, int) { t = fma(a, x[j], t); });
    t *= dinv[i];
    y[i] = t;
    acc[0] = fma(t, t, acc[0]);
    acc[1] = fma(x[i], x[i], acc[1]);
  }
  grid_reduce_sum<2>(acc, partials, counter, [&](double(&t)[2]) { out[0] = t[0]; out[1] = t[1]; });
}
__global__ void k_pow_scale(int64_t n, const double* __restrict__ y, const double* __restrict__ out, double* __restrict__ x) {
  const double s = out[0] > 0.0 ? rsqrt(out[0]) : 0.0;
  GRID_STRIDE(i, n) x[i] = y[i] * s;
}

int amg_exchange_prolongator(const Halo& H, int64_t n, int64_t ntot, const Csr& P, const std::vector<int64_t>& nc_all,
                             const std::vector<int64_t>& coff_all, Csr& Pext, Halo& Hc, int64_t* ntot_c, int64_t** gid_c);
int amg_dense_gather(const AmgLevel& L, double shift, double* a);

// While the replicated tail of a multi-GPU hierarchy is being built every rank works on the same
// gathered operator, so the set-up must not communicate.
static bool g_serial = false;
static bool amg_dist() { return comm_active() && !g_serial; }

// set-up decisions must be identical on every rank: reduce the local status / flags
static int global_max_int(int v, int* out) {
  double h = (double)v;
  MV_CHECK(allreduce_host(&h, 1, 1));
  *out = (int)h;
  return MV_OK;
}
static int global_sum_i64(int64_t v, int64_t* out) {
  double h = (double)v;
  MV_CHECK(allreduce_host(&h, 1, 0));
  *out = (int64_t)(h + 0.5);
  return MV_OK;
}
// st: local status of a set-up stage; returns the first non-OK status of any rank (as UNSUPPORTED if remote)
static int agree(int st) {
  if (!amg_dist()) return st;
  int g = 0;
  if (global_max_int(st == MV_OK ? 0 : 1, &g) != MV_OK) return MV_ERR_COMM;
  if (g && st == MV_OK) { set_error("amg: set-up stage failed on another rank"); return MV_ERR_UNSUPPORTED; }
  return st;
}

static int estimate_rho(mv_mesh* m, const AmgLevel& L, double* rho) {
  double *x = nullptr, *y = nullptr, *out = nullptr;
  MV_CHECK(dev_alloc(&x, L.ntot));
  MV_CHECK(dev_alloc(&y, L.ntot));
  MV_CHECK(dev_alloc(&out, 2));
  MV_CUDA(cudaMemsetAsync(x, 0, L.ntot * sizeof(double), g_stream));
  const int grid = grid_for(L.n);
  MV_L k_pow_init<<<grid, MV_BLOCK, 0, g_stream>>>(L.n, L.dinv, x);
  const int iters = 15;
  for (int k = 0; k < iters; ++k) {
    if (amg_dist()) MV_CHECK(halo_exchange_soa(L.halo, L.n, 1, x, L.ntot));
    MV_L k_pow_step<<<grid, MV_BLOCK, 0, g_stream>>>(L.n, L.pat.slice_ptr, L.pat.cols, L.val, L.pat.row_nnz, L.dinv, x, y,
                                                m->red.partials, m->red.counter, out);
    if (amg_dist()) MV_CHECK(allreduce_sum_dev(out, 2));
    if (k + 1 < iters) MV_L k_pow_scale<<<grid, MV_BLOCK, 0, g_stream>>>(L.n, y, out, x);
  }
  MV_KERNEL_CHECK();
  double h[2] = {0, 0};
  MV_CUDA(cudaMemcpyAsync(h, out, 2 * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  dev_free(x); dev_free(y); dev_free(out);
  const double lam = h[1] > 0.0 ? sqrt(h[0] / h[1]) : 1.0;
  *rho = 1.1 * (lam > 0.0 && lam == lam ? lam : 1.0);
  return MV_OK;
}

// ------------------------------------------------------------------ dense inverse of the coarsest operator
__global__ void k_dense_fill(int64_t n, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
                             const double* __restrict__ val, const int* __restrict__ row_nnz, double shift, double* __restrict__ a) {
  GRID_STRIDE(i, n) {
    for (int64_t j = 0; j < n; ++j) a[i * n + j] = shift;
    sell_row(slice_ptr, cols, val, row_nnz, i, [&](int j, double v, int) { if (j < n) a[i * n + j] += v; });
  }
}
// In-place Gauss-Jordan without pivoting (the matrix is SPD), one CTA.
__global__ void __launch_bounds__(1024) k_gauss_jordan(int n, double* __restrict__ a, int* __restrict__ err) {
  __shared__ double prow[MV_AMG_DENSE];
  __shared__ double fcol[MV_AMG_DENSE];
  for (int k = 0; k < n; ++k) {
    const double piv = a[(size_t)k * n + k];
    if (!(fabs(piv) > 0.0)) { if (threadIdx.x == 0) *err = 1; return; }
    const double ip = 1.0 / piv;
    for (int j = threadIdx.x; j < n; j += blockDim.x) {
      prow[j] = (j == k ? 1.0 : a[(size_t)k * n + j]) * ip;
      fcol[j] = a[(size_t)j * n + k];
    }
    __syncthreads();
    for (int64_t e = threadIdx.x; e < (int64_t)n * n; e += blockDim.x) {
      const int i = (int)(e / n), j = (int)(e - (int64_t)i * n);
      if (i == k) a[e] = prow[j];
      else a[e] = (j == k ? 0.0 : a[e]) - fcol[i] * prow[j];
    }
    __syncthreads();
  }
}

__global__ void k_mean_diag(int64_t n, const double* __restrict__ diag, double* __restrict__ out) {
  // tiny (n <= MV_AMG_DENSE): one thread, fixed order
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    double s = 0.0;
    for (int64_t i = 0; i < n; ++i) s += fabs(diag[i]);
    *out = s;
  }
}

// ------------------------------------------------------------------ constrained-set signature
__global__ void k_mask_sig(int64_t n, const uint16_t* __restrict__ mask, int allbits, unsigned long long* __restrict__ sig) {
  unsigned long long h = 0, cnt = 0, bad = 0;
  GRID_STRIDE(i, n) {
    const int b = mask[i] & allbits;
    if (b == allbits) { h += (unsigned long long)hash32((unsigned)i) * 0x9E3779B97F4A7C15ull + 1ull; ++cnt; }
    else if (b != 0) ++bad;
  }
  if (h) atomicAdd(&sig[0], h);
  if (cnt) atomicAdd(&sig[1], cnt);
  if (bad) atomicAdd(&sig[2], bad);
}
__global__ void k_mask_copy(int64_t n, const uint16_t* __restrict__ mask, int allbits, uint16_t* __restrict__ out) {
  GRID_STRIDE(i, n) out[i] = (mask[i] & allbits) == allbits ? 1 : 0;
}


// ------------------------------------------------------------------ host driver
static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e && *e ? atoi(e) : dflt;
}
static double env_double(const char* name, double dflt) {
  const char* e = getenv(name);
  return e && *e ? atof(e) : dflt;
}

// aggregates of level L -> agg[n] (device), *nc
static int aggregate(mv_mesh* m, const Graph& g, int** agg_out, int64_t* nc) {
  const int64_t n = g.n;
  const int grid = grid_for(n);
  int *state = nullptr, *f1 = nullptr, *agg0 = nullptr, *agg1 = nullptr, *agg = nullptr, *counter = nullptr;
  unsigned long long* t1 = nullptr;
  int64_t *flag = nullptr, *rank = nullptr;
  MV_CHECK(dev_alloc(&state, n));
  MV_CHECK(dev_alloc(&f1, n));
  MV_CHECK(dev_alloc(&t1, n));
  MV_CHECK(dev_alloc(&counter, 2));
  MV_L k_mis_init<<<grid, MV_BLOCK, 0, g_stream>>>(g, state);
  int remaining = 1, rounds = 0;
  while (remaining > 0) {
    if (++rounds > 200) { set_error("amg: MIS-2 did not terminate"); return MV_ERR_NOT_CONVERGED; }
    MV_CUDA(cudaMemsetAsync(counter, 0, 2 * sizeof(int), g_stream));
    MV_L k_mis_max1<<<grid, MV_BLOCK, 0, g_stream>>>(g, state, t1);
    MV_L k_mis_max2<<<grid, MV_BLOCK, 0, g_stream>>>(g, state, t1);
    MV_L k_mis_near1<<<grid, MV_BLOCK, 0, g_stream>>>(g, state, f1);
    MV_L k_mis_near2<<<grid, MV_BLOCK, 0, g_stream>>>(g, state, f1, counter);
    MV_KERNEL_CHECK();
    MV_CUDA(cudaMemcpyAsync(&remaining, counter, sizeof(int), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
  }
  MV_CHECK(dev_alloc(&flag, n + 1));
  MV_CHECK(dev_alloc(&rank, n + 1));
  MV_L k_root_flag<<<grid_for(n + 1), MV_BLOCK, 0, g_stream>>>(n, state, flag);
  MV_CHECK(exclusive_scan_i64(flag, rank, n + 1));
  MV_CUDA(cudaMemcpyAsync(nc, rank + n, sizeof(int64_t), cudaMemcpyDeviceToHost, g_stream));
  MV_CHECK(dev_alloc(&agg0, n));
  MV_CHECK(dev_alloc(&agg1, n));
  MV_CHECK(dev_alloc(&agg, n));
  MV_CUDA(cudaMemsetAsync(counter, 0, 2 * sizeof(int), g_stream));
  MV_L k_agg_roots<<<grid, MV_BLOCK, 0, g_stream>>>(n, state, rank, agg0);
  MV_L k_agg_pass1<<<grid, MV_BLOCK, 0, g_stream>>>(g, state, agg0, agg1);
  MV_L k_agg_pass2<<<grid, MV_BLOCK, 0, g_stream>>>(g, agg1, agg, counter);
  MV_KERNEL_CHECK();
  int orphans = 0;
  MV_CUDA(cudaMemcpyAsync(&orphans, counter, sizeof(int), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  dev_free(state); dev_free(f1); dev_free(t1); dev_free(counter); dev_free(flag); dev_free(rank); dev_free(agg0); dev_free(agg1);
  if (orphans) {   // cannot happen for a maximal set; guard against a broken (non-symmetric) graph
    dev_free(agg);
    set_error("amg: %d vertices left without an aggregate (non-symmetric operator?)", orphans);
    return MV_ERR_UNSUPPORTED;
  }
  *agg_out = agg;
  return MV_OK;
}

template <typename Count, typename Fill>
static int two_phase(Csr& c, Count count, Fill fill, int* err_flag) {
  int64_t* cnt = nullptr;
  MV_CHECK(dev_alloc(&cnt, c.nrows + 1));
  count(cnt);
  MV_KERNEL_CHECK();
  MV_CHECK(csr_finish_symbolic(c, cnt));
  dev_free(cnt);
  fill();
  MV_KERNEL_CHECK();
  int h = 0;
  MV_CUDA(cudaMemcpyAsync(&h, err_flag, sizeof(int), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  if (h) { set_error("amg: product row wider than %d entries", MV_AMG_ACC); return MV_ERR_UNSUPPORTED; }
  return MV_OK;
}

// Build level l+1 from level l.  Returns MV_ERR_UNSUPPORTED when coarsening should stop here (same answer on every rank).
static int coarsen(mv_mesh* m, Amg* a, double theta, double* diag) {
  AmgLevel& L = a->lv.back();
  const bool dist = amg_dist();
  Graph g{L.n, L.pat.slice_ptr, L.pat.cols, L.val, L.pat.row_nnz, diag, L.mask, theta * theta};
  int* agg = nullptr;
  int64_t nc = 0;
  int st = agree(aggregate(m, g, &agg, &nc));
  if (st != MV_OK) { dev_free(agg); return st; }
  std::vector<int64_t> nc_all, coff_all;
  if (dist) MV_CHECK(allgather_i64(nc, nc_all));
  else nc_all.assign(1, nc);
  coff_all.assign(nc_all.size() + 1, 0);
  for (size_t r = 0; r < nc_all.size(); ++r) coff_all[r + 1] = coff_all[r] + nc_all[r];
  const int64_t ncg = coff_all.back();
  if (ncg <= 0 || ncg * 10 > L.nglobal * 8) {   // stalled
    dev_free(agg);
    return MV_ERR_UNSUPPORTED;
  }
  int* err = nullptr;
  MV_CHECK(dev_alloc(&err, 1));
  MV_CUDA(cudaMemsetAsync(err, 0, sizeof(int), g_stream));
  const int grid = grid_for(L.n);
  const double omega = 4.0 / (3.0 * L.rho);
  Csr P, R, Pext, AP, Ac;
  AmgLevel N;
  P.nrows = L.n; P.ncols = nc;
  st = two_phase(P, [&](int64_t* cnt) { MV_L k_prolongator<false><<<grid, MV_BLOCK, 0, g_stream>>>(g, agg, omega, cnt, nullptr, nullptr, err); },
                 [&]() { MV_L k_prolongator<true><<<grid, MV_BLOCK, 0, g_stream>>>(g, agg, omega, P.ptr, P.col, P.val, err); }, err);
  dev_free(agg);
  if (st == MV_OK) st = csr_transpose(P, R);
  st = agree(st);
  int64_t ntot_c = nc;
  if (st == MV_OK && dist) {
    st = amg_exchange_prolongator(L.halo, L.n, L.ntot, P, nc_all, coff_all, Pext, N.halo, &ntot_c, &N.gid);
    N.owns_halo = true;
    st = agree(st);
  }
  const Csr& B = dist ? Pext : P;
  if (st == MV_OK) {
    AP.nrows = L.n; AP.ncols = ntot_c;
    st = two_phase(AP,
                   [&](int64_t* cnt) {
                     MV_L k_spgemm_sell_csr<false><<<grid, MV_BLOCK, 0, g_stream>>>(L.n, L.pat.slice_ptr, L.pat.cols, L.val, L.pat.row_nnz, L.mask,
                                                                             B.ptr, B.col, B.val, B.nrows, cnt, nullptr, nullptr, err);
                   },
                   [&]() {
                     MV_L k_spgemm_sell_csr<true><<<grid, MV_BLOCK, 0, g_stream>>>(L.n, L.pat.slice_ptr, L.pat.cols, L.val, L.pat.row_nnz, L.mask,
                                                                            B.ptr, B.col, B.val, B.nrows, AP.ptr, AP.col, AP.val, err);
                   },
                   err);
  }
  if (st == MV_OK) {
    Ac.nrows = nc; Ac.ncols = ntot_c;
    const int gc = grid_for(nc);
    st = two_phase(Ac,
                   [&](int64_t* cnt) {
                     MV_L k_spgemm_csr_csr<false><<<gc, MV_BLOCK, 0, g_stream>>>(nc, R.ptr, R.col, R.val, AP.ptr, AP.col, AP.val, cnt, nullptr,
                                                                          nullptr, err);
                   },
                   [&]() {
                     MV_L k_spgemm_csr_csr<true><<<gc, MV_BLOCK, 0, g_stream>>>(nc, R.ptr, R.col, R.val, AP.ptr, AP.col, AP.val, Ac.ptr, Ac.col,
                                                                         Ac.val, err);
                   },
                   err);
  }
  AP.release();
  Pext.release();
  dev_free(err);
  if (st == MV_OK) st = csr_to_sell(P, L.P, &L.Pval);
  if (st == MV_OK) st = csr_to_sell(R, L.R, &L.Rval);
  if (st == MV_OK) {
    double* v = nullptr;
    st = csr_to_sell(Ac, N.pat, &v);
    N.val = v;
    N.owns_A = true;
  }
  if (st == MV_OK && R.nrows > 0 && R.nnz > 48 * R.nrows) { L.Rc = R; R = Csr(); }   // long rows: keep CSR for the warp-per-row kernel
  P.release(); R.release(); Ac.release();
  st = agree(st);
  if (st != MV_OK) {
    sell_free(L.P); sell_free(L.R); dev_free(L.Pval); dev_free(L.Rval);
    L.P = Sell(); L.R = Sell();
    if (N.owns_A) { sell_free(N.pat); double* v = const_cast<double*>(N.val); dev_free(v); }
    if (N.owns_halo) { dev_free(N.halo.send_idx); dev_free(N.halo.send_buf); }
    dev_free(N.gid);
    return st;
  }
  N.n = nc;
  N.ntot = ntot_c;
  N.nglobal = ncg;
  N.goff = dist ? coff_all[comm_rank()] : 0;
  N.dist = dist;
  a->lv.push_back(N);
  return MV_OK;
}

// SELL rows [row0, row0+nrows) -> CSR; gid (optional) maps the local column numbers
__global__ void k_sell_rows_count(int64_t nrows, int64_t row0, const int* __restrict__ row_nnz, int64_t* __restrict__ cnt) {
  GRID_STRIDE(i, nrows + 1) cnt[i] = i < nrows ? row_nnz[row0 + i] : 0;
}
__global__ void k_sell_rows_fill(int64_t nrows, int64_t row0, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
                                 const double* __restrict__ val, const int* __restrict__ row_nnz, const int64_t* __restrict__ gid,
                                 const int64_t* __restrict__ ptr, int* __restrict__ col, double* __restrict__ v) {
  GRID_STRIDE(i, nrows) {
    const int64_t o = ptr[i];
    sell_row(slice_ptr, cols, val, row_nnz, row0 + i, [&](int c, double a, int k) {
      col[o + k] = gid ? (int)gid[c] : c;
      v[o + k] = a;
    });
  }
}
static int sell_rows_to_csr(const Sell& p, const double* val, int64_t row0, int64_t nrows, int64_t ncols, const int64_t* gid, Csr& c) {
  c.nrows = nrows;
  c.ncols = ncols;
  int64_t* cnt = nullptr;
  MV_CHECK(dev_alloc(&cnt, nrows + 1));
  MV_L k_sell_rows_count<<<grid_for(nrows + 1), MV_BLOCK, 0, g_stream>>>(nrows, row0, p.row_nnz, cnt);
  MV_KERNEL_CHECK();
  MV_CHECK(csr_finish_symbolic(c, cnt));
  dev_free(cnt);
  MV_L k_sell_rows_fill<<<grid_for(nrows), MV_BLOCK, 0, g_stream>>>(nrows, row0, p.slice_ptr, p.cols, val, p.row_nnz, gid, c.ptr, c.col, c.val);
  MV_KERNEL_CHECK();
  return MV_OK;
}
__global__ void k_add_i64(int64_t n, int64_t* __restrict__ p, int64_t off) { GRID_STRIDE(i, n) p[i] += off; }

// The rows of a distributed level, gathered on every rank in global numbering (CSR).
static int gather_level(const AmgLevel& L, Csr& G) {
  Csr loc;
  MV_CHECK(sell_rows_to_csr(L.pat, L.val, 0, L.n, L.nglobal, L.gid, loc));
  std::vector<int64_t> n_all, z_all;
  MV_CHECK(allgather_i64(L.n, n_all));
  MV_CHECK(allgather_i64(loc.nnz, z_all));
  const int R = (int)n_all.size(), me = comm_rank();
  std::vector<int64_t> roff(R + 1, 0), zoff(R + 1, 0);
  for (int r = 0; r < R; ++r) { roff[r + 1] = roff[r] + n_all[r]; zoff[r + 1] = zoff[r] + z_all[r]; }
  G.nrows = G.ncols = roff[R];
  G.nnz = zoff[R];
  MV_CHECK(dev_alloc(&G.ptr, G.nrows + 1));
  MV_CHECK(dev_alloc(&G.col, G.nnz));
  MV_CHECK(dev_alloc(&G.val, G.nnz));
  if (L.n) {
    MV_L k_add_i64<<<grid_for(L.n), MV_BLOCK, 0, g_stream>>>(L.n, loc.ptr, zoff[me]);
    MV_CUDA(cudaMemcpyAsync(G.ptr + roff[me], loc.ptr, L.n * sizeof(int64_t), cudaMemcpyDeviceToDevice, g_stream));
    MV_CUDA(cudaMemcpyAsync(G.col + zoff[me], loc.col, loc.nnz * sizeof(int), cudaMemcpyDeviceToDevice, g_stream));
    MV_CUDA(cudaMemcpyAsync(G.val + zoff[me], loc.val, loc.nnz * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
  }
  MV_CUDA(cudaMemcpyAsync(G.ptr + G.nrows, &G.nnz, sizeof(int64_t), cudaMemcpyHostToDevice, g_stream));
  for (int r = 0; r < R; ++r) {
    MV_CHECK(bcast_bytes(G.ptr + roff[r], (size_t)n_all[r] * sizeof(int64_t), r));
    MV_CHECK(bcast_bytes(G.col + zoff[r], (size_t)z_all[r] * sizeof(int), r));
    MV_CHECK(bcast_bytes(G.val + zoff[r], (size_t)z_all[r] * sizeof(double), r));
  }
  MV_CUDA(cudaStreamSynchronize(g_stream));
  loc.release();
  return MV_OK;
}

struct BuildParams {
  double theta;
  int deg0, degc;
  int64_t dense_max, repl_max;
};
static int build_levels(mv_mesh* m, Amg* a, const BuildParams& bp);

// Multi-GPU: from this (distributed) level down, the hierarchy is built from the gathered operator and
// kept REPLICATED on every rank: coarse levels are tiny, and a rank-local prolongator on them (where the
// partition interface is a large share of the rows) costs iterations (oracle/amg_prototype.py: 21 -> 30+).
static int replicate_tail(mv_mesh* m, Amg* a, const BuildParams& bp) {
  Csr G;
  MV_CHECK(gather_level(a->lv.back(), G));
  Amg tail;
  tail.op = a->op;
  tail.singular = a->singular;
  {
    AmgLevel T0;
    T0.n = T0.ntot = T0.nglobal = G.nrows;
    double* v = nullptr;
    MV_CHECK(csr_to_sell(G, T0.pat, &v));
    T0.val = v;
    T0.owns_A = true;
    tail.lv.push_back(T0);
  }
  G.release();
  g_serial = true;
  BuildParams tp = bp;
  tp.deg0 = bp.degc;
  int st = build_levels(m, &tail, tp);
  g_serial = false;
  if (st != MV_OK) return st;
  AmgLevel& L = a->lv.back();
  if (tail.lv.size() == 1) {
    // could not coarsen further: this level is the coarsest; keep the (gathered) dense inverse if there is one
    a->coarse_inv = tail.coarse_inv;
    tail.coarse_inv = nullptr;
    a->nc = L.nglobal;
    if (!a->coarse_inv) L.degree = 8;
    return MV_OK;
  }
  // own rows of the global prolongator; the restriction is its transpose (partial sums, allreduced in the cycle)
  const AmgLevel& T0 = tail.lv[0];
  const int64_t nc = tail.lv[1].n;
  Csr Ploc, Rpart;
  MV_CHECK(sell_rows_to_csr(T0.P, T0.Pval, L.goff, L.n, nc, nullptr, Ploc));
  MV_CHECK(csr_transpose(Ploc, Rpart));
  MV_CHECK(csr_to_sell(Ploc, L.P, &L.Pval));
  MV_CHECK(csr_to_sell(Rpart, L.R, &L.Rval));
  if (Rpart.nrows > 0 && Rpart.nnz > 48 * Rpart.nrows) { L.Rc = Rpart; Rpart = Csr(); }
  Ploc.release(); Rpart.release();
  L.reduce_next = true;
  a->coarse_inv = tail.coarse_inv;
  tail.coarse_inv = nullptr;
  a->nc = tail.nc;
  std::vector<AmgLevel> moved(tail.lv.begin() + 1, tail.lv.end());
  tail.lv.resize(1);   // the tail's destructor frees the gathered operator and the global P, R
  for (AmgLevel& N : moved) a->lv.push_back(N);
  return MV_OK;
}

// Coarsen from the last level of a->lv until the coarsest level is reached; sets a->nc / a->coarse_inv.
static int build_levels(mv_mesh* m, Amg* a, const BuildParams& bp) {
  int st = MV_OK;
  while (true) {
    AmgLevel& L = a->lv.back();
    const bool dist = amg_dist();
    L.dist = dist;
    L.degree = a->lv.size() == 1 ? bp.deg0 : bp.degc;
    double* diag = nullptr;
    if ((st = dev_alloc(&diag, L.n)) != MV_OK) break;
    if ((st = dev_alloc(&L.dinv, L.n)) != MV_OK) break;
    MV_L k_amg_diag<<<grid_for(L.n), MV_BLOCK, 0, g_stream>>>(L.n, L.pat.slice_ptr, L.pat.cols, L.val, L.pat.row_nnz, L.mask, diag, L.dinv);
    if ((st = estimate_rho(m, L, &L.rho)) != MV_OK) { dev_free(diag); break; }
    const bool last = L.nglobal <= bp.dense_max || (int)a->lv.size() >= MV_AMG_MAX_LEVELS;
    if (!last && dist && L.gid != nullptr && L.nglobal <= bp.repl_max) {
      dev_free(diag);
      return replicate_tail(m, a, bp);
    }
    int cs = last ? MV_ERR_UNSUPPORTED : coarsen(m, a, bp.theta, diag);
    if (cs == MV_OK) { dev_free(diag); continue; }
    if (cs != MV_ERR_UNSUPPORTED) { st = cs; dev_free(diag); break; }
    // this level is the coarsest
    AmgLevel& Lc = a->lv.back();
    const int64_t N = Lc.nglobal;
    a->nc = N;
    // gathered dense solve needs the global column ids, which exist from level 1 on
    if (N <= MV_AMG_DENSE && Lc.mask == nullptr && (!dist || Lc.gid != nullptr)) {
      double* shift = nullptr;
      int* err = nullptr;
      if ((st = dev_alloc(&a->coarse_inv, (size_t)N * N)) != MV_OK) break;
      if ((st = dev_alloc(&shift, 1)) != MV_OK) break;
      if ((st = dev_alloc(&err, 1)) != MV_OK) break;
      cudaMemsetAsync(err, 0, sizeof(int), g_stream);
      MV_L k_mean_diag<<<1, 32, 0, g_stream>>>(Lc.n, diag, shift);
      double hs = 0.0;
      cudaMemcpyAsync(&hs, shift, sizeof(double), cudaMemcpyDeviceToHost, g_stream);
      cudaStreamSynchronize(g_stream);
      if (dist && (st = allreduce_host(&hs, 1, 0)) != MV_OK) break;
      hs = a->singular ? hs / ((double)N * (double)N) : 0.0;   // A + s 1 1^T: exact pseudo-inverse action on vectors orthogonal to 1
      if (dist) {
        if ((st = amg_dense_gather(Lc, hs, a->coarse_inv)) != MV_OK) break;
      } else {
        MV_L k_dense_fill<<<grid_for(Lc.n), MV_BLOCK, 0, g_stream>>>(Lc.n, Lc.pat.slice_ptr, Lc.pat.cols, Lc.val, Lc.pat.row_nnz, hs, a->coarse_inv);
      }
      MV_L k_gauss_jordan<<<1, 1024, 0, g_stream>>>((int)N, a->coarse_inv, err);
      int herr = 0;
      cudaMemcpyAsync(&herr, err, sizeof(int), cudaMemcpyDeviceToHost, g_stream);
      cudaStreamSynchronize(g_stream);
      dev_free(shift); dev_free(err);
      if (herr) { dev_free(a->coarse_inv); }   // singular pivot: fall back to smoothing on the coarsest level
    }
    if (!a->coarse_inv) Lc.degree = 8;
    dev_free(diag);
    break;
  }
  return st;
}

static int amg_build(mv_mesh* m, int op, const uint16_t* mask_bit0, int singular, Amg** out) {
  const double *val, *dinv_unused;
  MV_CHECK(op_arrays(m, op, &val, &dinv_unused));
  return amg_build_values(m, op, val, mask_bit0, singular, out);
}

// hierarchy of ANY symmetric M-matrix-like operator stored on the mesh pattern (borrowed values; the caller keeps them
// alive): the assembled operators above, the diagonal blocks of the spring stiffness (mv_movers.cu)
int amg_build_values(mv_mesh* m, int op, const double* val, const uint16_t* mask_bit0, int singular, Amg** out) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0, g_stream);
  Amg* a = new Amg();
  static unsigned long long next_uid = 1;
  a->uid = next_uid++;
  a->op = op;
  a->singular = singular;
  BuildParams bp;
  bp.theta = env_double("MV_AMG_THETA", 1e-8);
  bp.deg0 = env_int("MV_AMG_DEG0", 1);
  bp.degc = env_int("MV_AMG_DEGC", 2);   // measured at 256^3: degree 2 = 24.8 PCG iterations, 192.6 ms per step (N = 4: 67.4); degree 3 = 23.5, 196.3 (70.8)
  bp.dense_max = env_int("MV_AMG_DENSE_MAX", 640);
  bp.repl_max = (int64_t)env_int("MV_AMG_REPLICATE_MAX", 4000000);
  AmgLevel L0;
  L0.n = m->Vown; L0.ntot = m->V;
  L0.pat = m->pat; L0.val = val; L0.owns_A = false;
  L0.mask = mask_bit0;
  L0.halo = m->halo;          // borrowed (device lists belong to the mesh)
  L0.owns_halo = false;
  L0.nglobal = L0.n;
  if (comm_active()) {
    MV_CHECK(global_sum_i64(L0.n, &L0.nglobal));
    std::vector<int64_t> all;
    MV_CHECK(allgather_i64(L0.n, all));
    for (int r = 0; r < comm_rank(); ++r) L0.goff += all[r];
  }
  a->lv.push_back(L0);
  int st = build_levels(m, a, bp);
  // The V-cycle streams the level operators several times per PCG iteration; it reads single-precision copies of the
  // values (8 instead of 12 B per non-zero).  The cycle is still one fixed symmetric operator (fl(K_ij) = fl(K_ji)),
  // vectors, diagonal and the outer PCG stay in double precision, so the solve reaches the same 1e-12.
  if (st == MV_OK && env_int("MV_AMG_FP32", 1) && a->lv.size() > 1) {
    const size_t nsm = a->lv.size() - (a->coarse_inv ? 1 : 0);   // the smoothed levels
    for (size_t l = 0; l < nsm && st == MV_OK; ++l) {
      AmgLevel& F = a->lv[l];
      if (!F.val || F.pat.len == 0) continue;
      st = dev_alloc(&F.valf, F.pat.len);
      if (st == MV_OK) MV_L k_to_float<<<grid_for(F.pat.len), MV_BLOCK, 0, g_stream>>>(F.pat.len, F.val, F.valf);
    }
  }
  // Coarse levels of a few 10^4 rows with ~80 non-zeros per row: one thread per row walks the row serially (36 us per
  // pass at 21 k rows, ncu round 2: 9 % of any unit -- latency); kept in CSR as well and smoothed one warp per row.
  if (st == MV_OK && env_int("MV_AMG_CSR_COARSE", 1)) {
    const size_t nsm = a->lv.size() - (a->coarse_inv ? 1 : 0);
    for (size_t l = 1; l < nsm && st == MV_OK; ++l) {
      AmgLevel& S = a->lv[l];
      if (S.dist || !S.val || S.n == 0 || S.n > 400000 || S.pat.nnz < 24 * S.n) continue;
      st = sell_rows_to_csr(S.pat, S.val, 0, S.n, S.ntot, nullptr, S.Ac);
    }
  }
  // Work split of large replicated levels: the tail of a multi-GPU hierarchy is replicated, so its cost does not
  // shrink with the number of ranks (256^3: level 1 has 1.5 M rows, ~25 ms per relaxation step on every rank,
  // half of the step at 8 GPUs).  Each rank then smooths only its chunk of slices and the vectors are allgathered.
  if (st == MV_OK && comm_active() && env_int("MV_AMG_SPLIT", 1)) {
    const int64_t split_min = (int64_t)env_int("MV_AMG_SPLIT_MIN", 200000);
    const size_t nsm = a->lv.size() - (a->coarse_inv ? 1 : 0);
    for (size_t l = 1; l < nsm; ++l) {
      AmgLevel& S = a->lv[l];
      if (S.dist || S.n < split_min || S.ntot != S.n || S.Ac.ptr) continue;
      const int64_t per = (S.pat.nslices + comm_size() - 1) / comm_size();
      S.split_chunk = per * MV_SLICE;
      S.split_lo = std::min<int64_t>(S.n, (int64_t)comm_rank() * S.split_chunk);
      S.split_hi = std::min<int64_t>(S.n, (int64_t)(comm_rank() + 1) * S.split_chunk);
      S.ntot = S.split_chunk * comm_size();   // leading dimension of the level's vectors (>= n)
    }
  }
  cudaEventRecord(e1, g_stream);
  cudaEventSynchronize(e1);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  a->setup_ms = ms;
  if (st == MV_OK && cudaGetLastError() != cudaSuccess) { set_error("amg: CUDA error during set-up"); st = MV_ERR_CUDA; }
  if (st != MV_OK) { delete a; return st; }
  if (getenv("MV_AMG_VERBOSE")) {
    fprintf(stderr, "[mv amg] rank %d op %d: %zu levels, set-up %.1f ms, singular %d, dense coarse %d\n", comm_rank(), op, a->lv.size(), ms,
            singular, a->coarse_inv ? 1 : 0);
    for (size_t l = 0; l < a->lv.size(); ++l)
      fprintf(stderr, "   level %zu: n %lld (+%lld ghosts, global %lld, %s) nnz %lld maxw %d rho %.3f degree %d\n", l, (long long)a->lv[l].n,
              (long long)(a->lv[l].ntot - a->lv[l].n), (long long)a->lv[l].nglobal, a->lv[l].dist ? "distributed" : "replicated/serial",
              (long long)a->lv[l].pat.nnz, a->lv[l].pat.maxw, a->lv[l].rho, a->lv[l].degree);
  }
  *out = a;
  return MV_OK;
}

int amg_get(mv_mesh* m, int op, const uint16_t* mask, int ncomp, int singular_hint, Amg** out) {
  *out = nullptr;
  if (m->pc_kind != MV_PC_AMG || op == MV_OP_MASS) return MV_OK;
  const bool dist = comm_active();
  int64_t vglobal = m->Vown;
  if (dist) MV_CHECK(global_sum_i64(m->Vown, &vglobal));
  if (vglobal < 64) return MV_OK;
  unsigned long long hsig[3] = {0, 0, 0};
  const int allbits = (1 << ncomp) - 1;
  if (mask) {
    unsigned long long* sig = nullptr;
    MV_CHECK(dev_alloc(&sig, 3));
    MV_CUDA(cudaMemsetAsync(sig, 0, 3 * sizeof(unsigned long long), g_stream));
    MV_L k_mask_sig<<<grid_for(m->Vown), MV_BLOCK, 0, g_stream>>>(m->Vown, mask, allbits, sig);
    MV_KERNEL_CHECK();
    MV_CUDA(cudaMemcpyAsync(hsig, sig, sizeof(hsig), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    dev_free(sig);
  }
  double gl[2] = {(double)hsig[1], (double)hsig[2]};
  if (dist) MV_CHECK(allreduce_host(gl, 2, 0));
  if (gl[1] > 0) return MV_OK;   // components constrained on different vertex sets: one hierarchy cannot serve them
  const bool constrained = gl[0] > 0;
  const int singular = (op == MV_OP_STIFFNESS && !constrained) ? 1 : 0;
  (void)singular_hint;
  Amg* hit = nullptr;
  for (Amg* a : m->amg)
    if (a->op == op && a->mask_hash == hsig[0] && a->singular == singular) hit = a;
  int miss = hit ? 0 : 1;
  if (dist) MV_CHECK(global_max_int(miss, &miss));   // a change of the constrained set on any rank rebuilds everywhere
  if (!miss) { *out = hit; return MV_OK; }
  if (hit) {
    for (size_t k = 0; k < m->amg.size(); ++k)
      if (m->amg[k] == hit) { delete hit; m->amg.erase(m->amg.begin() + k); break; }
  }
  uint16_t* mcopy = nullptr;
  if (constrained && hsig[1] > 0) {
    MV_CHECK(dev_alloc(&mcopy, m->Vown));
    MV_L k_mask_copy<<<grid_for(m->Vown), MV_BLOCK, 0, g_stream>>>(m->Vown, mask, allbits, mcopy);
    MV_KERNEL_CHECK();
  }
  Amg* a = nullptr;
  int st = amg_build(m, op, mcopy, singular, &a);
  if (st != MV_OK) { dev_free(mcopy); return st; }
  a->mask_hash = hsig[0];
  a->mask_copy = mcopy;
  if (m->amg.size() >= 4) { delete m->amg.front(); m->amg.erase(m->amg.begin()); }
  m->amg.push_back(a);
  *out = a;
  return MV_OK;
}

}  // namespace mv

using namespace mv;

extern "C" int mv_set_preconditioner(mv_mesh* m, int kind) {
  if (!m || (kind != MV_PC_JACOBI && kind != MV_PC_AMG)) return MV_ERR_BADARG;
  m->pc_kind = kind;
  return MV_OK;
}

// info[0] = levels, info[1] = coarsest rows, info[2] = dense coarse solve (0/1), then per level (rows, nnz)
extern "C" int mv_amg_info(mv_mesh* m, int op, int64_t* info, int max_levels, double* rho, double* setup_ms) {
  if (!m || !info) return MV_ERR_BADARG;
  for (Amg* a : m->amg)
    if (a->op == op) {
      info[0] = (int64_t)a->lv.size();
      info[1] = a->nc;
      info[2] = a->coarse_inv ? 1 : 0;
      for (int l = 0; l < (int)a->lv.size() && l < max_levels; ++l) {
        info[3 + 2 * l] = a->lv[l].n;
        info[4 + 2 * l] = a->lv[l].pat.nnz;
        if (rho) rho[l] = a->lv[l].rho;
      }
      if (setup_ms) *setup_ms = a->setup_ms;
      return MV_OK;
    }
  info[0] = 0;
  return MV_OK;
}

// which: 0 = operator A_l, 1 = prolongator P_l, 2 = restriction R_l.  size[0..3] = rows, slices, padded length, nnz.
// Any of the destination pointers (host or device) may be NULL.
extern "C" int mv_amg_level_get(mv_mesh* m, int op, int level, int which, int64_t* size, int64_t* slice_ptr, int32_t* cols,
                                double* vals, int32_t* row_nnz) {
  if (!m || !size) return MV_ERR_BADARG;
  for (Amg* a : m->amg) {
    if (a->op != op) continue;
    if (level < 0 || level >= (int)a->lv.size()) break;
    const AmgLevel& L = a->lv[level];
    const Sell& p = which == 0 ? L.pat : which == 1 ? L.P : L.R;
    const double* v = which == 0 ? L.val : which == 1 ? L.Pval : L.Rval;
    if (!p.slice_ptr) break;
    size[0] = p.nrows; size[1] = p.nslices; size[2] = p.len; size[3] = p.nnz;
    cudaMemcpyKind k = cudaMemcpyDefault;
    if (slice_ptr) MV_CUDA(cudaMemcpyAsync(slice_ptr, p.slice_ptr, (p.nslices + 1) * sizeof(int64_t), k, g_stream));
    if (cols) MV_CUDA(cudaMemcpyAsync(cols, p.cols, p.len * sizeof(int), k, g_stream));
    if (vals) MV_CUDA(cudaMemcpyAsync(vals, v, p.len * sizeof(double), k, g_stream));
    if (row_nnz) MV_CUDA(cudaMemcpyAsync(row_nnz, p.row_nnz, p.nrows * sizeof(int), k, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    return MV_OK;
  }
  set_error("mv_amg_level_get: no such level / operator");
  return MV_ERR_BADARG;
}
