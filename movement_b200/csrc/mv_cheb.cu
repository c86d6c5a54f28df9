// Consistent-mass solves by a FUSED Chebyshev iteration: one kernel per iteration, no inner products.
//
// replaces: PETSc KSP `cg` + `bjacobi/ilu` (movement/solver_parameters.py:62-66) at
//           movement/monge_ampere.py:340 (d mass solves, L2 projection of grad(phi)) and :483
//           (d*d mass solves, Hessian recovery).
//
// Why not CG here: for P1 simplices the Jacobi-scaled consistent mass matrix has its spectrum inside
// [1/2, 1 + d/2] on ANY mesh (Wathen 1987, element-by-element bound; a principal submatrix -- Dirichlet
// rows removed -- interlaces, so the bound survives the masks of monge_ampere.py:258-272).  With the
// interval known a priori the optimal polynomial is Chebyshev's, whose coefficients are host constants:
//     x_{k+1} = x_k + a_k (x_k - x_{k-1}) + c_k D^-1 (b - M x_k)
// Every term is local to the row except the gather of x_k, so the whole iteration is ONE kernel that
// streams the matrix once and touches 32 B per component per vertex (b, x_{k-1} in, x_{k+1} out over
// x_{k-1}, own x_k) against 104 B for a CG iteration (SpMV + update + direction): per vertex and
// iteration  B_cheb(c) = 12 nnz + 4 + 8 + 32 c  (3D: 480 B at c = 9; CG: 1120 B).  Chebyshev needs
// ~25 % more iterations than CG for the same reduction (it cannot exploit the discrete spectrum), so a
// solve moves about half the bytes, launches a third of the kernels, and on several GPUs needs no
// allreduce except at the sampled convergence checks.
//
// Convergence is MEASURED, not assumed: checked launches also reduce r_k . D^-1 r_k per component
// (the residual of their input iterate, deterministic grid reduction) and apply the stopping rule of
// mv_pcg_solve, sqrt(r.D^-1 r) <= max(rtol * ref, atol), ref = max(b.D^-1 b, r_0.D^-1 r_0).  The iterate
// returned is the one whose residual passed the test.
#include <math.h>

#include <type_traits>

#include "mv_internal.cuh"

namespace mv {

// Solver selection for MV_OP_MASS (mv_set_mass_solver): 1 = fused Chebyshev (default), 0 = Jacobi-PCG.
int g_mass_solver = 1;
int g_mass_mixed = -1;   // -1: MV_CHEB_MIXED / default (on); 0 / 1 forced through mv_set_mass_solver

enum { CHEB_PLAIN = 0, CHEB_CHECK = 1, CHEB_FIRST = 2 };

// flags: [0] done, [1] index k of the verified iterate, [2] breakdown (NaN)
template <int C>
__device__ __forceinline__ void cheb_decide(int k, double* __restrict__ scal, int* __restrict__ flags) {
  bool all = true;
  for (int c = 0; c < C; ++c) {
    const double rz = scal[SC(S_RZ, c)];
    all = all && (rz <= scal[SC(S_THR, c)]);
    if (!(rz == rz)) flags[2] = 1;
  }
  if (all || flags[2]) { flags[0] = 1; flags[1] = k; }
}
template <int C>
__device__ __forceinline__ void cheb_setup(double rtol, double atol, double* __restrict__ scal) {
  for (int c = 0; c < C; ++c) {
    const double ref = fmax(scal[SC(S_RZ, c)], scal[SC(S_BB, c)]);
    scal[SC(S_REF, c)] = ref;
    scal[SC(S_THR, c)] = fmax(rtol * rtol * ref, atol * atol);
  }
}
// multi-GPU: the sums were allreduced, now decide
template <int C>
__global__ void k_cheb_post(int k, int first, double rtol, double atol, double* __restrict__ scal, int* __restrict__ flags) {
  if (flags[0]) return;
  if (first) cheb_setup<C>(rtol, atol, scal);
  cheb_decide<C>(k, scal, flags);
}

// Deterministic grid reduction of the per-thread sums of a checked launch (same scheme as grid_reduce_sum, with
// the component of a slot depending on the warp): totals [kind][component] -> scal, then the decision.
template <int C, int CG, int SPB, int MODE, bool LOCAL, int NS>
__device__ __forceinline__ void cheb_grid_reduce(double (&sum)[NS], int k, double rtol, double atol, int* __restrict__ flags,
                                                 double* __restrict__ partials, unsigned* __restrict__ counter,
                                                 double* __restrict__ scal) {
  constexpr int NG = C / CG, NW = NG * SPB;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  constexpr int NT = MODE == CHEB_FIRST ? 2 * C : C;   // totals: [kind][component]
  __shared__ double s_w[NW][NS];
  __shared__ double s_tot[MV_RED_SLOTS];
  __shared__ bool s_last;
#pragma unroll
  for (int n = 0; n < NS; ++n) {
    const double t = warp_sum(sum[n]);
    if (lane == 0) s_w[wid][n] = t;
  }
  __syncthreads();
  if (threadIdx.x < NT) {
    const int kind = threadIdx.x / C, c = threadIdx.x % C, gg = c / CG, j = c % CG;
    double t = 0.0;
    for (int q = 0; q < SPB; ++q) t += s_w[q * NG + gg][kind * CG + j];
    partials[(size_t)blockIdx.x * MV_RED_SLOTS + threadIdx.x] = t;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // last block: warp n sums slot n over the blocks in a fixed order
  for (int n = wid; n < NT; n += NW) {
    double t = 0.0;
    for (unsigned q = lane; q < gridDim.x; q += 32) t += __ldcg(&partials[(size_t)q * MV_RED_SLOTS + n]);
    t = warp_sum(t);
    if (lane == 0) s_tot[n] = t;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    *counter = 0u;
    for (int c = 0; c < C; ++c) {
      scal[SC(S_RZ, c)] = s_tot[c];
      if (MODE == CHEB_FIRST) scal[SC(S_BB, c)] = s_tot[C + c];
    }
    if (LOCAL) {
      if (MODE == CHEB_FIRST) cheb_setup<C>(rtol, atol, scal);
      cheb_decide<C>(k, scal, flags);
    }
  }
}

// One Chebyshev iteration.  xin = x_k ([C][ld], ghosts valid), xout holds x_{k-1} on entry and x_{k+1} on exit
// (owned rows), b is component-major [C][ld].  MODE: CHEB_CHECK also sums r_k.D^-1 r_k, CHEB_FIRST (k = 0,
// a = 0, xout not read) sums b.D^-1 b as well and fixes the thresholds.  LOCAL: single GPU, decide in-kernel.
//
// Work split: a CTA of 32*NG*SPB threads owns SPB consecutive slices; the NG = C/CG warps of one slice each
// carry CG components (ncu, round 1: with all 9 components in one thread the kernel needs 64 registers,
// 4 CTAs/SM, and keeps too few gathers in flight -- 35 % of any unit, long-scoreboard bound).  The warps of a
// slice read the same column/value stream; the first one brings it from HBM, the others hit L1.
template <int C, int CG, int SPB, int MINB, int MODE, bool LOCAL>
__global__ void __launch_bounds__(32 * (C / CG) * SPB, MINB)
k_cheb_soa(const int* __restrict__ sched, int64_t ngroups, int64_t nrows, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
           const double* __restrict__ val, const double* __restrict__ dinv, const uint16_t* __restrict__ mask,
           const double* __restrict__ b, const double* __restrict__ xin, double* __restrict__ xout, int64_t ld,
           double ak, double ck, int k, double rtol, double atol, int* __restrict__ flags,
           double* __restrict__ partials, unsigned* __restrict__ counter, double* __restrict__ scal) {
  constexpr int NG = C / CG;
  static_assert(C % CG == 0, "component groups must tile C");
  if (MODE != CHEB_FIRST && *flags) return;
  constexpr int NS = MODE == CHEB_FIRST ? 2 * CG : (MODE == CHEB_CHECK ? CG : 1);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int g = wid % NG, sl = wid / NG;
  const int64_t cb = (int64_t)g * CG * ld;   // offset of this warp's first component
  b += cb; xin += cb; xout += cb;
  double sum[NS];
#pragma unroll
  for (int c = 0; c < NS; ++c) sum[c] = 0.0;
  for (int64_t sg = blockIdx.x; sg < ngroups; sg += gridDim.x) {
    const int64_t s = __ldg(&sched[sg * SPB + sl]);   // the SPB slices of a group are neighbours in the mesh (mv_sched.cu)
    if (s < 0) continue;
    const int64_t base = __ldg(&slice_ptr[s]);
    const int w = (int)((__ldg(&slice_ptr[s + 1]) - base) >> 5);
    const int64_t row = s * MV_SLICE + lane;
    const bool live = row < nrows;
    const int64_t rr = live ? row : nrows - 1;
    const int* cp = cols + base + lane;
    const double* vp = val + base + lane;
    // the streamed operands of this row are requested before the gathers so that they are in flight together
    double bv[CG], xo[CG];
#pragma unroll
    for (int c = 0; c < CG; ++c) {
      bv[c] = __ldcs(&b[c * ld + rr]);
      xo[c] = MODE == CHEB_FIRST ? 0.0 : __ldcs(&xout[c * ld + rr]);
    }
    const double di = __ldg(&dinv[rr]);
    const unsigned mk = mask ? (unsigned)__ldg(&mask[rr]) >> (g * CG) : 0u;
    double acc[CG];
#pragma unroll
    for (int c = 0; c < CG; ++c) acc[c] = 0.0;
    constexpr int KB = CG >= 3 ? 5 : 8;
    for (int k0 = 0; k0 < w; k0 += KB) {
      int col[KB];
      double a[KB];
#pragma unroll
      for (int j = 0; j < KB; ++j) {
        const bool in = k0 + j < w;
        // NG > 1: default caching, the sibling warps of this slice re-read the stream from L1
        col[j] = in ? (NG > 1 ? __ldg(cp + (k0 + j) * MV_SLICE) : __ldcs(cp + (k0 + j) * MV_SLICE)) : 0;
        a[j] = in ? (NG > 1 ? __ldg(vp + (k0 + j) * MV_SLICE) : __ldcs(vp + (k0 + j) * MV_SLICE)) : 0.0;
      }
#pragma unroll
      for (int j = 0; j < KB; ++j) {
        const double* xp = xin + col[j];
#pragma unroll
        for (int c = 0; c < CG; ++c) acc[c] = fma(a[j], __ldg(xp + c * ld), acc[c]);
      }
    }
    if (live) {
#pragma unroll
      for (int c = 0; c < CG; ++c) {
        const bool fix = (mk >> c) & 1u;
        const double xk = __ldg(&xin[c * ld + row]);
        const double r = fix ? 0.0 : bv[c] - acc[c];
        const double z = di * r;
        // x_{k+1} = x_k + a_k (x_k - x_{k-1}) + c_k z_k ; a constrained dof keeps its value
        const double xn = fix ? xk : xk + (MODE == CHEB_FIRST ? 0.0 : ak * (xk - xo[c])) + ck * z;
        xout[c * ld + row] = xn;
        if (MODE != CHEB_PLAIN) sum[c] = fma(r, z, sum[c]);
        if (MODE == CHEB_FIRST) sum[CG + c] = fix ? sum[CG + c] : fma(bv[c] * di, bv[c], sum[CG + c]);
      }
    }
  }
  if (MODE == CHEB_PLAIN) return;
  cheb_grid_reduce<C, CG, SPB, MODE, LOCAL, NS>(sum, k, rtol, atol, flags, partials, counter, scal);
}

// ------------------------------------------------------------------------------------------------------
// TMA-staged variant: the matrix stream of a slice (column indices + values, two contiguous runs of
// w*128 B and w*256 B) is brought into shared memory by bulk asynchronous copies (cp.async.bulk, completion
// on an mbarrier) ONE ITERATION AHEAD of its use.  ncu of k_cheb_soa<9> showed no unit above 50 % and every
// warp stalled on the scoreboard: each chunk of a row waits for its indices from HBM and only then can
// issue the gathers that depend on them (two memory latencies in series, three chunks per row, few
// registers to keep loads in flight).  With the indices already in shared memory the gathers of a whole row
// are issued back to back, and the HBM stream needs no registers at all.
// ------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// MIX selects the role in the mixed-precision solver further down (0: the all-double iteration described above):
//   MIX_RES  TV = TX = double: residual pass -- r = b - M xin and the sums of MODE, r -> f_r (single), first correction
//            iterate ck D^-1 r -> f_d (single); xout is not touched
//   MIX_IT   TV = TX = float: one iteration on the correction, b = the single-precision residual, xin / xout the
//            correction iterates (MODE == CHEB_FIRST: the previous iterate is zero; no sums in any mode);
//            x64 != nullptr: the last iteration of a sweep, x64 += the new iterate
//   MIX_RES4 / MIX_IT4: the same two roles with the single-precision iterates PACKED per component group: CG = 3 components of
//            a vertex + one pad in one float4 ([C/3][ld] float4), so that a neighbour's three values arrive with ONE 128-bit
//            gather instead of three 32-bit ones.  ncu (round 2, 256^3) showed the SoA fp32 iteration issue-bound, not
//            bandwidth-bound: 2.2 warp instructions per cycle and SM, 1750 per slice, 135 of them gathers -- 1.41 ms at
//            2.8 TB/s of DRAM traffic where the fp64 kernel it replaced took 1.45 ms.
enum { MIX_NONE = 0, MIX_RES = 1, MIX_IT = 2, MIX_RES4 = 3, MIX_IT4 = 4 };

template <int C, int CG, int SPB, int MINB, int NSTAGE, bool VS, int MODE, bool LOCAL, typename TV = double, typename TX = double, int MIX = MIX_NONE>
__global__ void __launch_bounds__(32 * (C / CG) * SPB, MINB)
k_cheb_tma(const int* __restrict__ sched, int64_t ngroups, int64_t nrows, const int64_t* __restrict__ slice_ptr,
           const int* __restrict__ cols, const TV* __restrict__ val, const TV* __restrict__ dinv,
           const uint16_t* __restrict__ mask, const TX* __restrict__ b, const TX* __restrict__ xin,
           TX* __restrict__ xout, int64_t ld, double ak, double ck, int k, double rtol, double atol,
           int* __restrict__ flags, double* __restrict__ partials, unsigned* __restrict__ counter,
           double* __restrict__ scal, int maxw, float* __restrict__ f_r = nullptr, float* __restrict__ f_d = nullptr,
           double* __restrict__ x64 = nullptr) {
  constexpr int NG = C / CG;
  constexpr bool IS_RES = MIX == MIX_RES || MIX == MIX_RES4, IS_IT = MIX == MIX_IT || MIX == MIX_IT4;
  constexpr bool PK = MIX == MIX_RES4 || MIX == MIX_IT4;
  static_assert(!PK || CG == 3, "packed iterates carry three components per float4");
  static_assert(MIX != MIX_IT4 || sizeof(TX) == 4, "packed iterates are single precision");
  if ((MODE != CHEB_FIRST || IS_IT) && *flags) return;
  constexpr int NS = IS_IT ? 1 : (MODE == CHEB_FIRST ? 2 * CG : (MODE == CHEB_CHECK ? CG : 1));
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int slice_elems = maxw * MV_SLICE;
  // VS: the values are staged too; otherwise only the indices (the gathers depend on them, the values do not) and
  // the shared-memory carve-out stays small, which leaves the L1 to the gathered vector
  TV* s_val = reinterpret_cast<TV*>(smem_raw);                                        // [NSTAGE][SPB][maxw*32] if VS
  int* s_col = reinterpret_cast<int*>(smem_raw + (VS ? (size_t)NSTAGE * SPB * slice_elems * sizeof(TV) : 0));   // [NSTAGE][SPB][maxw*32]
  __shared__ int64_t s_base[NSTAGE][SPB];
  __shared__ __align__(8) uint64_t s_bar[NSTAGE];
  __shared__ int s_width[NSTAGE][SPB];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int g = wid % NG, sl = wid / NG;
  const int64_t cb = (int64_t)g * CG * ld;
  const int64_t cb4 = (int64_t)g * 4 * ld;   // packed iterates: group g starts at float 4 g ld
  b += cb;
  xin += MIX == MIX_IT4 ? cb4 : cb;
  if (!IS_RES) xout += MIX == MIX_IT4 ? cb4 : cb;
  if (IS_RES) { f_r += cb; f_d += PK ? cb4 : cb; }
  if (IS_IT && x64) x64 += cb;
  const TX akx = (TX)ak, ckx = (TX)ck;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int q = 0; q < NSTAGE; ++q) mbar_init(&s_bar[q], SPB);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // producer: lane q < SPB of warp 0 fetches the stream of slot q of group `sg` into stage `st`
  auto produce = [&](int64_t sg, int st) {
    if (wid == 0 && lane < SPB) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic reads of this stage precede the async writes
      const int64_t s = sg < ngroups ? (int64_t)__ldg(&sched[sg * SPB + lane]) : -1;
      if (s >= 0) {
        const int64_t base = __ldg(&slice_ptr[s]);
        const int w = (int)((__ldg(&slice_ptr[s + 1]) - base) >> 5);
        s_width[st][lane] = w;
        s_base[st][lane] = base;
        const uint32_t nb_col = (uint32_t)w * MV_SLICE * sizeof(int), nb_val = VS ? (uint32_t)w * MV_SLICE * sizeof(TV) : 0u;
        mbar_arrive_expect_tx(&s_bar[st], nb_col + nb_val);
        if (w > 0) {
          if (VS) bulk_g2s(s_val + ((size_t)st * SPB + lane) * slice_elems, val + base, nb_val, &s_bar[st]);
          bulk_g2s(s_col + ((size_t)st * SPB + lane) * slice_elems, cols + base, nb_col, &s_bar[st]);
        }
      } else {
        s_width[st][lane] = 0;
        mbar_arrive_expect_tx(&s_bar[st], 0u);
      }
    }
  };
  double sum[NS];
#pragma unroll
  for (int c = 0; c < NS; ++c) sum[c] = 0.0;
#pragma unroll
  for (int q = 0; q < NSTAGE - 1; ++q) produce(blockIdx.x + (int64_t)q * gridDim.x, q);
  int64_t s_cur = blockIdx.x < ngroups ? (int64_t)__ldg(&sched[(int64_t)blockIdx.x * SPB + sl]) : -1;
  int it = 0;
  for (int64_t sg = blockIdx.x; sg < ngroups; sg += gridDim.x, ++it) {
    const int st = it % NSTAGE;
    // every warp is done with the stage read in the previous iteration: refill it for iteration it + NSTAGE - 1
    __syncthreads();
    produce(sg + (int64_t)(NSTAGE - 1) * gridDim.x, (it + NSTAGE - 1) % NSTAGE);
    const int64_t s = s_cur;
    const int64_t sgn = sg + gridDim.x;
    s_cur = sgn < ngroups ? (int64_t)__ldg(&sched[sgn * SPB + sl]) : -1;
    const int64_t row = (s < 0 ? 0 : s) * MV_SLICE + lane;
    const bool live = s >= 0 && row < nrows;
    const int64_t rr = live ? row : 0;
    // streamed operands first: they are in flight while the warp waits for the matrix tile
    TX bv[CG], xo[CG];
#pragma unroll
    for (int c = 0; c < CG; ++c) bv[c] = __ldcs(&b[c * ld + rr]);
    if (MIX == MIX_IT4) {
      const float4 t = MODE == CHEB_FIRST ? make_float4(0.f, 0.f, 0.f, 0.f) : __ldcs(reinterpret_cast<const float4*>(xout) + rr);
      xo[0] = (TX)t.x; xo[1 % CG] = (TX)t.y; xo[2 % CG] = (TX)t.z;
    } else {
#pragma unroll
      for (int c = 0; c < CG; ++c) xo[c] = (MODE == CHEB_FIRST || IS_RES) ? (TX)0 : __ldcs(&xout[c * ld + rr]);
    }
    const TV di = __ldg(&dinv[rr]);
    const unsigned mk = mask ? (unsigned)__ldg(&mask[rr]) >> (g * CG) : 0u;
    mbar_wait(&s_bar[st], (uint32_t)((it / NSTAGE) & 1));
    const int w = s >= 0 ? s_width[st][sl] : 0;
    const int* cp = s_col + ((size_t)st * SPB + sl) * slice_elems + lane;
    const TV* vp = VS ? s_val + ((size_t)st * SPB + sl) * slice_elems + lane : val + (s >= 0 ? s_base[st][sl] : 0) + lane;
    TX acc[CG];
#pragma unroll
    for (int c = 0; c < CG; ++c) acc[c] = (TX)0;
    if (MIX == MIX_IT4) {
      // one 128-bit gather per neighbour
      constexpr int KB = MINB >= 4 ? 5 : 8;   // fewer registers per thread at higher occupancy targets
      const float4* x4 = reinterpret_cast<const float4*>(xin);
      for (int k0 = 0; k0 < w; k0 += KB) {
        int col[KB];
#pragma unroll
        for (int j = 0; j < KB; ++j) col[j] = k0 + j < w ? cp[(k0 + j) * MV_SLICE] : (int)rr;
        float4 xv[KB];
#pragma unroll
        for (int j = 0; j < KB; ++j) xv[j] = __ldg(x4 + col[j]);
#pragma unroll
        for (int j = 0; j < KB; ++j) {
          const TV a = k0 + j < w ? (VS ? vp[(k0 + j) * MV_SLICE] : __ldg(vp + (k0 + j) * MV_SLICE)) : (TV)0;
          acc[0] = fma(a, (TX)xv[j].x, acc[0]);
          acc[1 % CG] = fma(a, (TX)xv[j].y, acc[1 % CG]);
          acc[2 % CG] = fma(a, (TX)xv[j].z, acc[2 % CG]);
        }
      }
    } else {
      constexpr int KB = CG >= 9 ? (sizeof(TX) == 4 ? 3 : 2) : (CG >= 3 ? (sizeof(TX) == 4 ? 8 : 5) : 8);
      for (int k0 = 0; k0 < w; k0 += KB) {
        int col[KB];
#pragma unroll
        for (int j = 0; j < KB; ++j) col[j] = k0 + j < w ? cp[(k0 + j) * MV_SLICE] : (int)rr;
        TX xv[KB][CG];
#pragma unroll
        for (int j = 0; j < KB; ++j) {
          const TX* xp = xin + col[j];
#pragma unroll
          for (int c = 0; c < CG; ++c) xv[j][c] = __ldg(xp + c * ld);
        }
#pragma unroll
        for (int j = 0; j < KB; ++j) {
          const TV a = k0 + j < w ? (VS ? vp[(k0 + j) * MV_SLICE] : (NG > 1 ? __ldg(vp + (k0 + j) * MV_SLICE) : __ldcs(vp + (k0 + j) * MV_SLICE))) : (TV)0;
#pragma unroll
          for (int c = 0; c < CG; ++c) acc[c] = fma(a, xv[j][c], acc[c]);
        }
      }
    }
    if (live) {
      TX xkv[CG], outv[CG];
      if (MIX == MIX_IT4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(xin) + row);
        xkv[0] = (TX)t.x; xkv[1 % CG] = (TX)t.y; xkv[2 % CG] = (TX)t.z;
      } else if (!IS_RES) {
#pragma unroll
        for (int c = 0; c < CG; ++c) xkv[c] = __ldg(&xin[c * ld + row]);
      }
#pragma unroll
      for (int c = 0; c < CG; ++c) {
        const bool fix = (mk >> c) & 1u;
        const TX r = fix ? (TX)0 : bv[c] - acc[c];
        const TX z = di * r;
        if (IS_RES) {
          f_r[c * ld + row] = (float)r;
          outv[c] = ckx * z;
          if (!PK) f_d[c * ld + row] = (float)outv[c];
        } else if (IS_IT) {
          const TX xk = xkv[c];
          const TX xn = fix ? (TX)0 : xk + akx * (xk - xo[c]) + ckx * z;
          outv[c] = xn;
          if (!PK) xout[c * ld + row] = xn;
          if (x64) x64[c * ld + row] += (double)xn;
        } else {
          const TX xk = xkv[c];
          const TX xn = fix ? xk : xk + (MODE == CHEB_FIRST ? (TX)0 : akx * (xk - xo[c])) + ckx * z;
          xout[c * ld + row] = xn;
        }
        if (!IS_IT) {
          if (MODE != CHEB_PLAIN) sum[c] = fma((double)r, (double)z, sum[c]);
          if (MODE == CHEB_FIRST) sum[CG + c] = fix ? sum[CG + c] : fma((double)(bv[c] * di), (double)bv[c], sum[CG + c]);
        }
      }
      if (PK) {
        const float4 o = make_float4((float)outv[0], (float)outv[1 % CG], (float)outv[2 % CG], 0.f);
        if (IS_RES) reinterpret_cast<float4*>(f_d)[row] = o;
        else reinterpret_cast<float4*>(xout)[row] = o;
      }
    }
  }
  if (MODE == CHEB_PLAIN || IS_IT) return;
  cheb_grid_reduce<C, CG, SPB, MODE, LOCAL, NS>(sum, k, rtol, atol, flags, partials, counter, scal);
}

// Wathen's interval for diag(M)^-1 M, P1 simplices in d dimensions.
static inline void mass_spectrum(int dim, double* lmin, double* lmax) {
  *lmin = 0.5;
  *lmax = 1.0 + 0.5 * dim;
}

template <int C, int CG, int SPB, int MINB, int NSTAGE, bool VS>
struct ChebLaunch {
  static int grid(int64_t ngroups) {
    static int per_sm = 0;
    if (per_sm == 0) {
      int nb = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_cheb_soa<C, CG, SPB, MINB, CHEB_CHECK, true>, 32 * (C / CG) * SPB, 0) !=
              cudaSuccess || nb < 1)
        nb = 1;
      per_sm = nb;
    }
    const int64_t need = ngroups;
    int64_t g = (int64_t)g_nsm * per_sm;
    if (g > MV_MAX_BLOCKS) g = MV_MAX_BLOCKS;
    if (g > need) g = need;
    return (int)(g < 1 ? 1 : g);
  }
  static int run(mv_mesh* m, const double* val, const double* dinv, const uint16_t* mask, const double* b, const double* xin,
                 double* xout, int64_t ld, double ak, double ck, int k, int mode, bool local, double rtol, double atol) {
    const Sell& p = m->pat;
    static_assert(MV_SCHED_TILE % SPB == 0, "a CTA works on an aligned part of a schedule tile");
    const int64_t ngroups = p.sched_len / SPB;
    constexpr int NT = 32 * (C / CG) * SPB;
    PcgWork& w = m->pcg;
    static int use_tma = -1;
    if (use_tma < 0) { const char* e = getenv("MV_CHEB_TMA"); use_tma = e ? atoi(e) : 1; }
    constexpr int NST = NSTAGE > 0 ? NSTAGE : 2;
    const size_t smem = (size_t)NST * SPB * p.maxw * MV_SLICE * ((VS ? sizeof(double) : 0) + sizeof(int));
    if (NSTAGE > 0 && use_tma && smem <= 200 * 1024) {
      static size_t smem_set = 0;
      static int per_sm = 0;
      if (smem != smem_set) {
        if (smem > 48 * 1024) {
#define CHEB_ATTR(MODE, LOC) MV_CUDA(cudaFuncSetAttribute(k_cheb_tma<C, CG, SPB, MINB, NST, VS, MODE, LOC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem))
          CHEB_ATTR(CHEB_PLAIN, true); CHEB_ATTR(CHEB_CHECK, true); CHEB_ATTR(CHEB_CHECK, false); CHEB_ATTR(CHEB_FIRST, true); CHEB_ATTR(CHEB_FIRST, false);
#undef CHEB_ATTR
        }
        int nb = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_cheb_tma<C, CG, SPB, MINB, NST, VS, CHEB_CHECK, true>, NT, smem) != cudaSuccess || nb < 1) nb = 1;
        per_sm = nb;
        smem_set = smem;
      }
      int64_t g = (int64_t)g_nsm * per_sm;
      if (g > MV_MAX_BLOCKS) g = MV_MAX_BLOCKS;
      if (g > ngroups) g = ngroups;
      const int gr = (int)(g < 1 ? 1 : g);
#define CHEB_GO(MODE, LOC)                                                                                                  \
  MV_L k_cheb_tma<C, CG, SPB, MINB, NST, VS, MODE, LOC><<<gr, NT, smem, g_stream>>>(p.sched, ngroups, p.nrows, p.slice_ptr, p.cols, val, dinv, mask, \
                                                                     b, xin, xout, ld, ak, ck, k, rtol, atol, w.flags,       \
                                                                     m->red.partials, m->red.counter, w.scal, p.maxw)
      if (mode == CHEB_PLAIN) CHEB_GO(CHEB_PLAIN, true);
      else if (mode == CHEB_CHECK) { if (local) CHEB_GO(CHEB_CHECK, true); else CHEB_GO(CHEB_CHECK, false); }
      else { if (local) CHEB_GO(CHEB_FIRST, true); else CHEB_GO(CHEB_FIRST, false); }
#undef CHEB_GO
      MV_KERNEL_CHECK();
      return MV_OK;
    }
    const int gr = grid(ngroups);
#define CHEB_GO(MODE, LOC)                                                                                                  \
  MV_L k_cheb_soa<C, CG, SPB, MINB, MODE, LOC><<<gr, NT, 0, g_stream>>>(p.sched, ngroups, p.nrows, p.slice_ptr, p.cols, val, dinv, mask, b, \
                                                                  xin, xout, ld, ak, ck, k, rtol, atol, w.flags,            \
                                                                  m->red.partials, m->red.counter, w.scal)
    if (mode == CHEB_PLAIN) CHEB_GO(CHEB_PLAIN, true);
    else if (mode == CHEB_CHECK) { if (local) CHEB_GO(CHEB_CHECK, true); else CHEB_GO(CHEB_CHECK, false); }
    else { if (local) CHEB_GO(CHEB_FIRST, true); else CHEB_GO(CHEB_FIRST, false); }
#undef CHEB_GO
    MV_KERNEL_CHECK();
    return MV_OK;
  }
};

template <int C, int CG, int SPB, int MINB, int NSTAGE, bool VS = false>
static int cheb_soa(mv_mesh* m, const double* val, const double* dinv, const double* rhs, double* x, const uint16_t* mask,
                    double rtol, double atol, int maxit, int* iters, double* relres) {
  MV_CHECK(ensure_pcg_workspace(m));
  MV_CHECK(ensure_slice_schedule(m));
  PcgWork& w = m->pcg;
  const int64_t n = m->Vown, ld = m->V;
  const bool dist = comm_active();
  double lmin, lmax;
  mass_spectrum(m->dim, &lmin, &lmax);
  const double theta = 0.5 * (lmax + lmin), delta = 0.5 * (lmax - lmin), sigma1 = theta / delta;

  double* X[2] = {w.xs, w.p};   // both span the ghosts
  double* bs = w.r;
  MV_CHECK(aos_to_soa(C, n, x, X[0], ld));
  if (rhs) MV_CHECK(aos_to_soa(C, n, rhs, bs, ld));
  else MV_CUDA(cudaMemsetAsync(bs, 0, (size_t)ld * C * sizeof(double), g_stream));
  MV_CUDA(cudaMemsetAsync(w.flags, 0, 4 * sizeof(int), g_stream));

  const int kmin = 6;           // first sampled convergence check (after the k = 0 launch, which sets the thresholds)
  static int kdense_env = -1;
  if (kdense_env < 0) { const char* e = getenv("MV_CHEB_KDENSE"); kdense_env = e ? atoi(e) : 18; }
  const int kdense = kdense_env;        // from here on every launch checks (the bound 2 rho^k reaches 1e-12 at k = 26..30)
  int k = 0;
  double rho = 1.0 / sigma1;    // rho_{k-1}
  w.h_flags[0] = 0;
  int batch = 20;
  while (true) {
    const int kend = k + batch < maxit ? k + batch : maxit;
    for (; k < kend; ++k) {
      double ak = 0.0, ck = 1.0 / theta;
      if (k > 0) {
        const double rk = 1.0 / (2.0 * sigma1 - rho);
        ak = rk * rho;
        ck = 2.0 * rk / delta;
        rho = rk;
      }
      const int mode = k == 0 ? CHEB_FIRST : ((k >= kdense || (k >= kmin && ((k - kmin) & 1) == 0)) ? CHEB_CHECK : CHEB_PLAIN);
      if (dist) MV_CHECK(halo_exchange_soa(m->halo, n, C, X[k & 1], ld));
      const int pslot = prof_begin(C);
      MV_CHECK((ChebLaunch<C, CG, SPB, MINB, NSTAGE, VS>::run(m, val, dinv, mask, bs, X[k & 1], X[(k + 1) & 1], ld, ak, ck, k, mode, !dist, rtol, atol)));
      prof_end(pslot);
      if (dist && mode != CHEB_PLAIN) {
        MV_CHECK(allreduce_sum_dev(w.scal + SC(S_RZ, 0), C));
        if (mode == CHEB_FIRST) MV_CHECK(allreduce_sum_dev(w.scal + SC(S_BB, 0), C));
        MV_L k_cheb_post<C><<<1, 1, 0, g_stream>>>(k, mode == CHEB_FIRST, rtol, atol, w.scal, w.flags);
        MV_KERNEL_CHECK();
      }
    }
    MV_CUDA(cudaMemcpyAsync(w.h_flags, w.flags, 4 * sizeof(int), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    if (g_prof.used) prof_drain();
    if (w.h_flags[0] || k >= maxit) break;
    batch = 8;
  }
  const int done = w.h_flags[0], kv = done ? w.h_flags[1] : k, breakdown = w.h_flags[2];
  g_prof.pcg_iters[C & 15] += kv;
  // the verified iterate x_kv was the input of launch kv
  MV_CHECK(soa_to_aos(C, n, X[kv & 1], ld, x));
  if (iters) *iters = kv;
  if (relres) {
    MV_CUDA(cudaMemcpyAsync(w.h_scal, w.scal, S_NSLOTS * 16 * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    for (int c = 0; c < C; ++c) {
      const double ref = w.h_scal[SC(S_REF, c)];
      relres[c] = ref > 0.0 ? sqrt(w.h_scal[SC(S_RZ, c)] / ref) : 0.0;
    }
  }
  if (breakdown) {
    set_error("chebyshev: breakdown (NaN) after %d iterations", kv);
    return MV_ERR_NOT_CONVERGED;
  }
  if (!done) {
    set_error("chebyshev: not converged in %d iterations", maxit);
    return MV_ERR_NOT_CONVERGED;
  }
  return MV_OK;
}


// ======================================================================================================
// MIXED-PRECISION variant (default): fp64 results at fp32 bandwidth.
//
// Iterative refinement around the same Chebyshev polynomial: a RESIDUAL pass in double (r = b - M x with the fp64
// matrix, the measured stopping test r.D^-1 r <= thr exactly as above) hands the residual to a SWEEP of k Chebyshev
// iterations carried out entirely in single precision -- fp32 copy of the matrix values, fp32 iterates d_j for the
// CORRECTION M d = r from d_0 = 0 -- whose last iteration adds d_k to the fp64 solution.  A sweep reduces the residual
// by 2 rho^k down to the fp32 floor (~1e-6), so 1e-12 takes two sweeps and three residual passes, and the returned iterate is
// still the one whose fp64 residual passed the test.  Per vertex an fp32 iteration moves 8 nnz + 8 + 16 c bytes
// (3D, c = 9: 272 B against 480 B) and gathers half the bytes through L1/L2, which is what bounds the fp64 kernel
// (ncu, round 1: L1/TEX 68-76 %, L2->L1 2.2 x the DRAM bytes).  31 fp32 iterations + 3 fp64 passes replace 30 fp64
// iterations from a cold start (oracle prototype, 14^3): 10.0 kB instead of 14.4 kB per vertex.
// fp32 range: residuals are stored unscaled; |r_i| ~ 1e-12 h^d |b| stays above 1e-30 for every mesh the 32-bit
// connectivity can hold, far from the 1.2e-38 denormal threshold.
// ======================================================================================================
enum { MK_RES_FIRST = 0, MK_RES_CHECK = 1, MK_IT = 2, MK_IT_FIRST = 3 };

template <int C, int CG, int SPB, int MINB, int KIND, bool LAST, bool LOCAL, bool PK = false>
__global__ void __launch_bounds__(32 * (C / CG) * SPB, MINB)
k_mcheb(const int* __restrict__ sched, int64_t ngroups, int64_t nrows, const int64_t* __restrict__ slice_ptr, const int* __restrict__ cols,
        const double* __restrict__ val64, const float* __restrict__ val32, const double* __restrict__ dinv64,
        const float* __restrict__ dinv32, const uint16_t* __restrict__ mask, const double* __restrict__ b, double* __restrict__ x64,
        float* __restrict__ r32, const float* __restrict__ din, float* __restrict__ dout, int64_t ld, double ak, double ck, int k,
        double rtol, double atol, int* __restrict__ flags, double* __restrict__ partials, unsigned* __restrict__ counter,
        double* __restrict__ scal) {
  constexpr int NG = C / CG;
  constexpr bool RES = KIND == MK_RES_FIRST || KIND == MK_RES_CHECK;
  static_assert(C % CG == 0, "component groups must tile C");
  if (KIND != MK_RES_FIRST && *flags) return;
  constexpr int NS = KIND == MK_RES_FIRST ? 2 * CG : (KIND == MK_RES_CHECK ? CG : 1);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int g = wid % NG, sl = wid / NG;
  const int64_t cb = (int64_t)g * CG * ld;
  const int64_t cb4 = (int64_t)g * 4 * ld;   // PK: the single-precision iterates are [C/3][ld] float4 (three components + pad)
  static_assert(!PK || CG == 3, "packed iterates carry three components per float4");
  double sum[NS];
#pragma unroll
  for (int c = 0; c < NS; ++c) sum[c] = 0.0;
  const float akf = (float)ak, ckf = (float)ck;
  for (int64_t sg = blockIdx.x; sg < ngroups; sg += gridDim.x) {
    const int64_t s = __ldg(&sched[sg * SPB + sl]);
    if (s < 0) continue;
    const int64_t base = __ldg(&slice_ptr[s]);
    const int w = (int)((__ldg(&slice_ptr[s + 1]) - base) >> 5);
    const int64_t row = s * MV_SLICE + lane;
    const bool live = row < nrows;
    const int64_t rr = live ? row : nrows - 1;
    const int* cp = cols + base + lane;
    const unsigned mk = mask ? (unsigned)__ldg(&mask[rr]) >> (g * CG) : 0u;
    if (RES) {
      // ---- fp64 residual pass: r = b - M x, r.D^-1 r, r -> fp32, d_1 = D^-1 r / theta
      const double* vp = val64 + base + lane;
      const double* xg = x64 + cb;
      double bv[CG];
#pragma unroll
      for (int c = 0; c < CG; ++c) bv[c] = __ldcs(&b[cb + c * ld + rr]);
      const double di = __ldg(&dinv64[rr]);
      double acc[CG];
#pragma unroll
      for (int c = 0; c < CG; ++c) acc[c] = 0.0;
      constexpr int KB = CG >= 3 ? 5 : 8;
      for (int k0 = 0; k0 < w; k0 += KB) {
        int col[KB];
        double a[KB];
#pragma unroll
        for (int j = 0; j < KB; ++j) {
          const bool in = k0 + j < w;
          col[j] = in ? (NG > 1 ? __ldg(cp + (k0 + j) * MV_SLICE) : __ldcs(cp + (k0 + j) * MV_SLICE)) : 0;
          a[j] = in ? (NG > 1 ? __ldg(vp + (k0 + j) * MV_SLICE) : __ldcs(vp + (k0 + j) * MV_SLICE)) : 0.0;
        }
#pragma unroll
        for (int j = 0; j < KB; ++j) {
          const double* xp = xg + col[j];
#pragma unroll
          for (int c = 0; c < CG; ++c) acc[c] = fma(a[j], __ldg(xp + c * ld), acc[c]);
        }
      }
      if (live) {
        float d1[CG];
#pragma unroll
        for (int c = 0; c < CG; ++c) {
          const bool fix = (mk >> c) & 1u;
          const double r = fix ? 0.0 : bv[c] - acc[c];
          const double z = di * r;
          r32[cb + c * ld + row] = (float)r;
          d1[c] = (float)(ck * z);
          if (!PK) dout[cb + c * ld + row] = d1[c];
          sum[c] = fma(r, z, sum[c]);
          if (KIND == MK_RES_FIRST) sum[CG + c] = fix ? sum[CG + c] : fma(bv[c] * di, bv[c], sum[CG + c]);
        }
        if (PK) reinterpret_cast<float4*>(dout + cb4)[row] = make_float4(d1[0], d1[1 % CG], d1[2 % CG], 0.f);
      }
    } else {
      // ---- fp32 Chebyshev iteration on the correction: d_{j+1} = d_j + a (d_j - d_{j-1}) + c D^-1 (r - M d_j)
      const float* vp = val32 + base + lane;
      const float* dg = din + (PK ? cb4 : cb);
      const float4* dg4 = reinterpret_cast<const float4*>(dg);
      float rv[CG], dp[CG];
#pragma unroll
      for (int c = 0; c < CG; ++c) rv[c] = __ldcs(&r32[cb + c * ld + rr]);
      if (PK) {
        const float4 t = KIND == MK_IT_FIRST ? make_float4(0.f, 0.f, 0.f, 0.f) : __ldcs(reinterpret_cast<const float4*>(dout + cb4) + rr);
        dp[0] = t.x; dp[1 % CG] = t.y; dp[2 % CG] = t.z;
      } else {
#pragma unroll
        for (int c = 0; c < CG; ++c) dp[c] = KIND == MK_IT_FIRST ? 0.0f : __ldcs(&dout[cb + c * ld + rr]);
      }
      const float di = __ldg(&dinv32[rr]);
      float acc[CG];
#pragma unroll
      for (int c = 0; c < CG; ++c) acc[c] = 0.0f;
      constexpr int KB = CG >= 9 ? 4 : 8;
      for (int k0 = 0; k0 < w; k0 += KB) {
        int col[KB];
        float a[KB];
#pragma unroll
        for (int j = 0; j < KB; ++j) {
          const bool in = k0 + j < w;
          col[j] = in ? (NG > 1 ? __ldg(cp + (k0 + j) * MV_SLICE) : __ldcs(cp + (k0 + j) * MV_SLICE)) : 0;
          a[j] = in ? (NG > 1 ? __ldg(vp + (k0 + j) * MV_SLICE) : __ldcs(vp + (k0 + j) * MV_SLICE)) : 0.0f;
        }
        if (PK) {
          float4 xv[KB];
#pragma unroll
          for (int j = 0; j < KB; ++j) xv[j] = __ldg(dg4 + col[j]);   // one 128-bit gather per neighbour
#pragma unroll
          for (int j = 0; j < KB; ++j) {
            acc[0] = fmaf(a[j], xv[j].x, acc[0]);
            acc[1 % CG] = fmaf(a[j], xv[j].y, acc[1 % CG]);
            acc[2 % CG] = fmaf(a[j], xv[j].z, acc[2 % CG]);
          }
        } else {
#pragma unroll
          for (int j = 0; j < KB; ++j) {
            const float* xp = dg + col[j];
#pragma unroll
            for (int c = 0; c < CG; ++c) acc[c] = fmaf(a[j], __ldg(xp + c * ld), acc[c]);
          }
        }
      }
      if (live) {
        float dkv[CG], dnv[CG];
        if (PK) {
          const float4 t = __ldg(dg4 + row);
          dkv[0] = t.x; dkv[1 % CG] = t.y; dkv[2 % CG] = t.z;
        } else {
#pragma unroll
          for (int c = 0; c < CG; ++c) dkv[c] = __ldg(&dg[c * ld + row]);
        }
#pragma unroll
        for (int c = 0; c < CG; ++c) {
          const bool fix = (mk >> c) & 1u;
          const float dk = dkv[c];
          const float res = fix ? 0.0f : rv[c] - acc[c];
          const float dn = fix ? 0.0f : dk + akf * (dk - dp[c]) + ckf * (di * res);
          dnv[c] = dn;
          if (!PK) dout[cb + c * ld + row] = dn;
          if (LAST) x64[cb + c * ld + row] += (double)dn;
        }
        if (PK) reinterpret_cast<float4*>(dout + cb4)[row] = make_float4(dnv[0], dnv[1 % CG], dnv[2 % CG], 0.f);
      }
    }
  }
  if (!RES) return;
  cheb_grid_reduce<C, CG, SPB, KIND == MK_RES_FIRST ? CHEB_FIRST : CHEB_CHECK, LOCAL, NS>(sum, k, rtol, atol, flags, partials, counter, scal);
}

__global__ void k_cheb_to_float(int64_t n, const double* __restrict__ a, float* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = (float)a[i];
}
// x += d (fp32 correction of a one-iteration sweep)
__global__ void k_add_float(int64_t n, const float* __restrict__ d, double* __restrict__ x) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) x[i] += (double)d[i];
}

static int ensure_mixed_buffers(mv_mesh* m, const double* val, const double* dinv) {
  PcgWork& w = m->pcg;
  const size_t maxc = (size_t)m->dim * m->dim;
  if (!w.f32[0]) {
    for (int i = 0; i < 3; ++i) {
      // room for the packed layout of the iterates: ceil(maxc / 3) float4 per vertex
      const size_t per_vertex = ((maxc + 2) / 3) * 4 > maxc ? ((maxc + 2) / 3) * 4 : maxc;
      MV_CHECK(dev_alloc(&w.f32[i], (size_t)m->V * per_vertex));
      MV_CUDA(cudaMemsetAsync(w.f32[i], 0, (size_t)m->V * per_vertex * sizeof(float), g_stream));
    }
  }
  if (w.f32_src != val) {     // single-precision copy of the operator the sweeps stream (one per mesh: the mass matrix)
    if (!w.valf) MV_CHECK(dev_alloc(&w.valf, (size_t)m->pat.len));
    if (!w.dinvf) MV_CHECK(dev_alloc(&w.dinvf, (size_t)m->Vown));
    MV_L k_cheb_to_float<<<grid_for(m->pat.len), MV_BLOCK, 0, g_stream>>>(m->pat.len, val, w.valf);
    MV_L k_cheb_to_float<<<grid_for(m->Vown), MV_BLOCK, 0, g_stream>>>(m->Vown, dinv, w.dinvf);
    MV_KERNEL_CHECK();
    w.f32_src = val;
  }
  return MV_OK;
}

template <int C, int CG, int SPB, int MINB, bool PK = false>
struct MChebLaunch {
  static int run(mv_mesh* m, int kind, bool last, bool local, const double* val, const double* dinv, const uint16_t* mask, const double* b,
                 double* x64, float* r32, const float* din, float* dout, int64_t ld, double ak, double ck, int k, double rtol, double atol) {
    const Sell& p = m->pat;
    static_assert(MV_SCHED_TILE % SPB == 0, "a CTA works on an aligned part of a schedule tile");
    const int64_t ngroups = p.sched_len / SPB;
    constexpr int NT = 32 * (C / CG) * SPB;
    PcgWork& w = m->pcg;
    static int per_sm = 0;
    if (per_sm == 0) {
      int nb = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_mcheb<C, CG, SPB, MINB, MK_IT, false, true, PK>, NT, 0) != cudaSuccess || nb < 1) nb = 1;
      per_sm = nb;
    }
    int64_t g = (int64_t)g_nsm * per_sm;
    if (g > MV_MAX_BLOCKS) g = MV_MAX_BLOCKS;
    if (g > ngroups) g = ngroups;
    const int gr = (int)(g < 1 ? 1 : g);
#define MCHEB_GO(KIND, LASTF, LOC)                                                                                                       \
  MV_L k_mcheb<C, CG, SPB, MINB, KIND, LASTF, LOC, PK><<<gr, NT, 0, g_stream>>>(p.sched, ngroups, p.nrows, p.slice_ptr, p.cols, val, w.valf, dinv,  \
                                                                       w.dinvf, mask, b, x64, r32, din, dout, ld, ak, ck, k, rtol, atol,  \
                                                                       w.flags, m->red.partials, m->red.counter, w.scal)
    if (kind == MK_RES_FIRST) { if (local) MCHEB_GO(MK_RES_FIRST, false, true); else MCHEB_GO(MK_RES_FIRST, false, false); }
    else if (kind == MK_RES_CHECK) { if (local) MCHEB_GO(MK_RES_CHECK, false, true); else MCHEB_GO(MK_RES_CHECK, false, false); }
    else if (kind == MK_IT_FIRST) { if (last) MCHEB_GO(MK_IT_FIRST, true, true); else MCHEB_GO(MK_IT_FIRST, false, true); }
    else { if (last) MCHEB_GO(MK_IT, true, true); else MCHEB_GO(MK_IT, false, true); }
#undef MCHEB_GO
    MV_KERNEL_CHECK();
    return MV_OK;
  }
};

// the TMA-staged kernel in its two mixed-precision roles
template <int C, int CG, int SPB, int MINB, int NST, int MIX>
struct MixTmaLaunch {
  static constexpr bool IT = MIX == MIX_IT || MIX == MIX_IT4;
  using TV = typename std::conditional<IT, float, double>::type;
  static bool fits(const mv_mesh* m) { return (size_t)NST * SPB * m->pat.maxw * MV_SLICE * (sizeof(TV) + sizeof(int)) <= 200 * 1024; }
  // MIX_RES: b, xin double, mode CHEB_FIRST / CHEB_CHECK;  MIX_IT: b = r32, xin / xout the float iterates, first: previous iterate zero
  static int run(mv_mesh* m, int mode, bool local, const uint16_t* mask, const void* b, const void* xin, void* xout, float* f_r, float* f_d,
                 double* x64, const double* val, const double* dinv, int64_t ld, double ak, double ck, int k, double rtol, double atol) {
    const Sell& p = m->pat;
    static_assert(MV_SCHED_TILE % SPB == 0, "a CTA works on an aligned part of a schedule tile");
    const int64_t ngroups = p.sched_len / SPB;
    constexpr int NT = 32 * (C / CG) * SPB;
    PcgWork& w = m->pcg;
    const size_t smem = (size_t)NST * SPB * p.maxw * MV_SLICE * (sizeof(TV) + sizeof(int));
    static size_t smem_set = 0;
    static int per_sm = 0;
#define MIX_K(MODE, LOC) k_cheb_tma<C, CG, SPB, MINB, NST, true, MODE, LOC, TV, TV, MIX>
    if (smem != smem_set) {
      if (smem > 48 * 1024) {
        MV_CUDA(cudaFuncSetAttribute(MIX_K(CHEB_PLAIN, true), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        MV_CUDA(cudaFuncSetAttribute(MIX_K(CHEB_FIRST, true), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (!IT) {
          MV_CUDA(cudaFuncSetAttribute(MIX_K(CHEB_CHECK, true), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
          MV_CUDA(cudaFuncSetAttribute(MIX_K(CHEB_CHECK, false), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
          MV_CUDA(cudaFuncSetAttribute(MIX_K(CHEB_FIRST, false), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        }
      }
      int nb = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, MIX_K(CHEB_FIRST, true), NT, smem) != cudaSuccess || nb < 1) nb = 1;
      per_sm = nb;
      smem_set = smem;
    }
    int64_t g = (int64_t)g_nsm * per_sm;
    if (g > MV_MAX_BLOCKS) g = MV_MAX_BLOCKS;
    if (g > ngroups) g = ngroups;
    const int gr = (int)(g < 1 ? 1 : g);
    const TV* valt = IT ? reinterpret_cast<const TV*>(w.valf) : reinterpret_cast<const TV*>(val);
    const TV* dinvt = IT ? reinterpret_cast<const TV*>(w.dinvf) : reinterpret_cast<const TV*>(dinv);
#define MIX_GO(MODE, LOC)                                                                                                                \
  MV_L MIX_K(MODE, LOC)<<<gr, NT, smem, g_stream>>>(p.sched, ngroups, p.nrows, p.slice_ptr, p.cols, valt, dinvt, mask, static_cast<const TV*>(b), \
                                                    static_cast<const TV*>(xin), static_cast<TV*>(xout), ld, ak, ck, k, rtol, atol, w.flags, \
                                                    m->red.partials, m->red.counter, w.scal, p.maxw, f_r, f_d, x64)
    if (IT) { if (mode == CHEB_FIRST) MIX_GO(CHEB_FIRST, true); else MIX_GO(CHEB_PLAIN, true); }
    else if (mode == CHEB_FIRST) { if (local) MIX_GO(CHEB_FIRST, true); else MIX_GO(CHEB_FIRST, false); }
    else { if (local) MIX_GO(CHEB_CHECK, true); else MIX_GO(CHEB_CHECK, false); }
#undef MIX_GO
#undef MIX_K
    MV_KERNEL_CHECK();
    return MV_OK;
  }
};

// R.. = launch shape of the fp64 residual passes, I.. = of the fp32 iterations; ..NST = stages of the TMA pipeline (0: plain loads)
template <int C, int RCG, int RSPB, int RMINB, int RNST, int ICG, int ISPB, int IMINB, int INST>
static int cheb_mixed(mv_mesh* m, const double* val, const double* dinv, const double* rhs, double* x, const uint16_t* mask,
                      double rtol, double atol, int maxit, int* iters, double* relres) {
  MV_CHECK(ensure_pcg_workspace(m));
  MV_CHECK(ensure_slice_schedule(m));
  MV_CHECK(ensure_mixed_buffers(m, val, dinv));
  PcgWork& w = m->pcg;
  const int64_t n = m->Vown, ld = m->V;
  const bool dist = comm_active();
  double lmin, lmax;
  mass_spectrum(m->dim, &lmin, &lmax);
  const double theta = 0.5 * (lmax + lmin), delta = 0.5 * (lmax - lmin), sigma1 = theta / delta;
  const double fac = (sqrt(lmax / lmin) - 1.0) / (sqrt(lmax / lmin) + 1.0);   // asymptotic reduction per iteration
  static int kcap_env = -1;
  if (kcap_env < 0) { const char* e = getenv("MV_CHEB_SWEEP"); kcap_env = e ? atoi(e) : 16; }
  const int kcap = kcap_env < 2 ? 2 : kcap_env;   // 2 rho^16 = 4e-7: the single-precision floor of a sweep

  double* X = w.xs;     // fp64 solution, component-major, ghosts included
  double* bs = w.r;
  float* r32 = w.f32[0];
  float* D[2] = {w.f32[1], w.f32[2]};
  MV_CHECK(aos_to_soa(C, n, x, X, ld));
  if (rhs) MV_CHECK(aos_to_soa(C, n, rhs, bs, ld));
  else MV_CUDA(cudaMemsetAsync(bs, 0, (size_t)ld * C * sizeof(double), g_stream));
  MV_CUDA(cudaMemsetAsync(w.flags, 0, 4 * sizeof(int), g_stream));

  // packed single-precision iterates (three components + pad per float4, one 128-bit gather per neighbour): whenever the
  // launch shapes carry three components per warp; on several GPUs only if the peer-memory halo exchange can carry them
  constexpr bool CAN_PK = C % 3 == 0 && RCG == 3 && ICG == 3;
  constexpr int NGRP = C / 3 > 0 ? C / 3 : 1;
  // measured at 256^3 (round 2): C = 9 iteration 1.50 -> 1.28 ms packed; C = 3 0.58 -> 0.59 ms (already at 0.8 of the HBM
  // peak, the pad only adds bytes), so three-component solves stay component-major unless MV_CHEB_PACK=2
  int pack_env = C >= 9 ? 1 : 0;
  { const char* e = getenv("MV_CHEB_PACK"); if (e) pack_env = atoi(e) == 2 ? 1 : (atoi(e) == 0 ? 0 : pack_env); }
  const bool pk = CAN_PK && pack_env && (!dist || (m->halo.nn > 0 && 2 * NGRP <= m->halo.p2p_maxC) || m->halo.nn == 0);
  constexpr int RES_MIX = CAN_PK ? MIX_RES4 : MIX_RES, IT_MIX = CAN_PK ? MIX_IT4 : MIX_IT;

  int cycle = 0, its32 = 0, passes = 0;
  bool done = false, breakdown = false;
  while (true) {
    // ---- residual pass (double): test of the current iterate, residual and first correction iterate in single
    if (dist) MV_CHECK(halo_exchange_soa(m->halo, n, C, X, ld));
    {
      const int pslot = prof_begin(C + 4);
      if (RNST > 0 && MixTmaLaunch<C, RCG, RSPB, RMINB, RNST ? RNST : 2, MIX_RES>::fits(m)) {
        if (pk)
          MV_CHECK((MixTmaLaunch<C, RCG, RSPB, RMINB, RNST ? RNST : 2, RES_MIX>::run(m, cycle == 0 ? CHEB_FIRST : CHEB_CHECK, !dist, mask, bs, X, nullptr, r32, D[1],
                                                                          nullptr, val, dinv, ld, 0.0, 1.0 / theta, cycle, rtol, atol)));
        else
          MV_CHECK((MixTmaLaunch<C, RCG, RSPB, RMINB, RNST ? RNST : 2, MIX_RES>::run(m, cycle == 0 ? CHEB_FIRST : CHEB_CHECK, !dist, mask, bs, X, nullptr, r32, D[1],
                                                                          nullptr, val, dinv, ld, 0.0, 1.0 / theta, cycle, rtol, atol)));
      } else if (pk) {
        MV_CHECK((MChebLaunch<C, RCG, RSPB, RMINB, CAN_PK>::run(m, cycle == 0 ? MK_RES_FIRST : MK_RES_CHECK, false, !dist, val, dinv, mask, bs, X, r32,
                                                                nullptr, D[1], ld, 0.0, 1.0 / theta, cycle, rtol, atol)));
      } else {
        MV_CHECK((MChebLaunch<C, RCG, RSPB, RMINB>::run(m, cycle == 0 ? MK_RES_FIRST : MK_RES_CHECK, false, !dist, val, dinv, mask, bs, X, r32,
                                                        nullptr, D[1], ld, 0.0, 1.0 / theta, cycle, rtol, atol)));
      }
      prof_end(pslot);
    }
    ++passes;
    if (dist) {
      MV_CHECK(allreduce_sum_dev(w.scal + SC(S_RZ, 0), C));
      if (cycle == 0) MV_CHECK(allreduce_sum_dev(w.scal + SC(S_BB, 0), C));
      MV_L k_cheb_post<C><<<1, 1, 0, g_stream>>>(cycle, cycle == 0, rtol, atol, w.scal, w.flags);
      MV_KERNEL_CHECK();
    }
    MV_CUDA(cudaMemcpyAsync(w.h_flags, w.flags, 4 * sizeof(int), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaMemcpyAsync(w.h_scal, w.scal, S_NSLOTS * 16 * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    if (g_prof.used) prof_drain();
    breakdown = w.h_flags[2] != 0;
    done = w.h_flags[0] != 0 && !breakdown;
    if (done || breakdown || its32 + passes - 1 >= maxit) break;
    // ---- how many single-precision iterations the remaining reduction needs (bound 2 fac^k), capped by the fp32 floor
    double need = 1.0;
    for (int c = 0; c < C; ++c) {
      const double rz = w.h_scal[SC(S_RZ, c)], thr = w.h_scal[SC(S_THR, c)];
      if (rz > thr && rz > 0.0) need = fmin(need, sqrt(thr / rz));
    }
    int k = (int)ceil(log(0.25 * need) / log(fac));
    if (k < 2) k = 2;
    if (k > kcap) k = kcap;
    // ---- the sweep: d_1 came with the residual pass; iterations 1 .. k-1 in single precision, the last one adds d to x
    double rho = 1.0 / sigma1;
    int cur = 1;   // D[cur] = d_j, D[cur ^ 1] = d_{j-1} -> d_{j+1}
    for (int j = 1; j < k; ++j) {
      const double rk = 1.0 / (2.0 * sigma1 - rho);
      const double ak = rk * rho, ck = 2.0 * rk / delta;
      rho = rk;
      if (dist) {
        if (pk) { if (m->halo.nn > 0) MV_CHECK(p2p_halo_exchange_f32_packed4(m->halo, n, NGRP, D[cur], ld, nullptr)); }
        else MV_CHECK(halo_exchange_soa_f32(m->halo, n, C, D[cur], ld));
      }
      const int pslot = prof_begin(C);
      if (INST > 0 && MixTmaLaunch<C, ICG, ISPB, IMINB, INST ? INST : 2, MIX_IT>::fits(m)) {
        if (pk)
          MV_CHECK((MixTmaLaunch<C, ICG, ISPB, IMINB, INST ? INST : 2, IT_MIX>::run(m, j == 1 ? CHEB_FIRST : CHEB_PLAIN, true, mask, r32, D[cur], D[cur ^ 1], nullptr,
                                                                         nullptr, j == k - 1 ? X : nullptr, val, dinv, ld, ak, ck, cycle, rtol, atol)));
        else
          MV_CHECK((MixTmaLaunch<C, ICG, ISPB, IMINB, INST ? INST : 2, MIX_IT>::run(m, j == 1 ? CHEB_FIRST : CHEB_PLAIN, true, mask, r32, D[cur], D[cur ^ 1], nullptr,
                                                                         nullptr, j == k - 1 ? X : nullptr, val, dinv, ld, ak, ck, cycle, rtol, atol)));
      } else if (pk) {
        MV_CHECK((MChebLaunch<C, ICG, ISPB, IMINB, CAN_PK>::run(m, j == 1 ? MK_IT_FIRST : MK_IT, j == k - 1, true, val, dinv, mask, bs, X, r32, D[cur],
                                                                D[cur ^ 1], ld, ak, ck, cycle, rtol, atol)));
      } else {
        MV_CHECK((MChebLaunch<C, ICG, ISPB, IMINB>::run(m, j == 1 ? MK_IT_FIRST : MK_IT, j == k - 1, true, val, dinv, mask, bs, X, r32, D[cur],
                                                        D[cur ^ 1], ld, ak, ck, cycle, rtol, atol)));
      }
      prof_end(pslot);
      cur ^= 1;
    }
    its32 += k;
    ++cycle;
  }
  const int total = its32 + passes - 1;   // matrix passes after the test of the initial guess
  g_prof.pcg_iters[C & 15] += its32;
  g_prof.pcg_iters[(C + 4) & 15] += passes;
  MV_CHECK(soa_to_aos(C, n, X, ld, x));
  if (iters) *iters = total;
  if (relres) {
    for (int c = 0; c < C; ++c) {
      const double ref = w.h_scal[SC(S_REF, c)];
      relres[c] = ref > 0.0 ? sqrt(w.h_scal[SC(S_RZ, c)] / ref) : 0.0;
    }
  }
  if (breakdown) {
    set_error("chebyshev: breakdown (NaN) after %d matrix passes", total);
    return MV_ERR_NOT_CONVERGED;
  }
  if (!done) {
    set_error("chebyshev: not converged in %d iterations", maxit);
    return MV_ERR_NOT_CONVERGED;
  }
  return MV_OK;
}

int cheb_mass_solve(mv_mesh* m, int ncomp, const double* val, const double* dinv, const double* rhs, double* x,
                    const uint16_t* mask, double rtol, double atol, int maxit, int* iters, double* relres) {
  static int var = -1;   // experiment knob: occupancy target (min CTAs/SM -> register cap)
  if (var < 0) { const char* e = getenv("MV_CHEB_VAR"); var = e ? atoi(e) : 0; }
#define CHEB_ARGS m, val, dinv, rhs, x, mask, rtol, atol, maxit, iters, relres
  // defaults = the fastest of the variants measured at 256^3 (B200, round 1): C = 1: 64 registers, 32 warps/SM,
  // 0.64 ms; C = 3: 40 registers, 48 warps/SM, 0.97 ms; C = 9: three 3-component warps per slice, 40 registers, 2.9 ms
  // last template argument: stages of the TMA pipeline of the matrix stream (0 = plain loads; measured faster for C <= 3)
  static int mixed = -1;   // MV_CHEB_MIXED=0: every iteration in double (round-1 solver)
  if (mixed < 0) { const char* e = getenv("MV_CHEB_MIXED"); mixed = e ? atoi(e) : 1; }
  if (g_mass_mixed >= 0) mixed = g_mass_mixed;
  if (mixed) {
    switch (ncomp) {
      case 1: return cheb_mixed<1, 1, 8, 4, 0, 1, 8, 4, 0>(CHEB_ARGS);
      case 2: return cheb_mixed<2, 2, 8, 4, 0, 2, 8, 4, 0>(CHEB_ARGS);
      case 3:
        if (var == 21) return cheb_mixed<3, 3, 8, 4, 0, 3, 8, 4, 0>(CHEB_ARGS);
        if (var == 22) return cheb_mixed<3, 3, 8, 6, 0, 3, 8, 3, 2>(CHEB_ARGS);
        if (var == 23) return cheb_mixed<3, 3, 8, 6, 0, 1, 4, 4, 2>(CHEB_ARGS);
        if (var == 24) return cheb_mixed<3, 3, 8, 6, 0, 1, 4, 4, 0>(CHEB_ARGS);
        if (var == 25) return cheb_mixed<3, 3, 8, 6, 0, 3, 8, 6, 0>(CHEB_ARGS);
        // measured at 256^3, cold start to 1e-12 (33 matrix passes): iteration kernel <3,8,4> (64 registers, no spills)
        // 19.6 ms per solve, <3,8,6> (40 registers) 23.9, TMA-staged <3,8,3> 21.2; all-double solver 30.1
        return cheb_mixed<3, 3, 8, 6, 0, 3, 8, 4, 0>(CHEB_ARGS);
      case 4: return cheb_mixed<4, 4, 8, 4, 0, 4, 8, 4, 0>(CHEB_ARGS);
      case 9:
        if (var == 11) return cheb_mixed<9, 3, 4, 3, 2, 9, 8, 2, 0>(CHEB_ARGS);
        if (var == 12) return cheb_mixed<9, 3, 4, 3, 2, 3, 2, 8, 0>(CHEB_ARGS);
        if (var == 13) return cheb_mixed<9, 3, 4, 3, 2, 3, 4, 4, 2>(CHEB_ARGS);
        if (var == 14) return cheb_mixed<9, 3, 4, 3, 2, 9, 8, 2, 2>(CHEB_ARGS);
        if (var == 15) return cheb_mixed<9, 3, 4, 3, 2, 3, 8, 2, 2>(CHEB_ARGS);
        if (var == 16) return cheb_mixed<9, 3, 4, 3, 2, 1, 2, 3, 2>(CHEB_ARGS);
        if (var == 31) return cheb_mixed<9, 3, 4, 3, 2, 3, 4, 4, 2>(CHEB_ARGS);
        if (var == 32) return cheb_mixed<9, 3, 4, 3, 2, 3, 8, 2, 2>(CHEB_ARGS);
        if (var == 33) return cheb_mixed<9, 3, 4, 3, 2, 3, 2, 6, 2>(CHEB_ARGS);
        if (var == 34) return cheb_mixed<9, 3, 4, 3, 2, 3, 4, 3, 0>(CHEB_ARGS);
        if (var == 35) return cheb_mixed<9, 3, 4, 3, 2, 3, 2, 8, 2>(CHEB_ARGS);
        if (var == 36) return cheb_mixed<9, 3, 4, 3, 2, 3, 4, 3, 3>(CHEB_ARGS);
        if (var == 37) return cheb_mixed<9, 3, 4, 3, 2, 3, 8, 2, 0>(CHEB_ARGS);
        if (var == 38) return cheb_mixed<9, 3, 4, 3, 2, 3, 4, 4, 0>(CHEB_ARGS);
        if (var == 39) return cheb_mixed<9, 3, 4, 3, 2, 3, 1, 12, 2>(CHEB_ARGS);
        if (var == 40) return cheb_mixed<9, 3, 4, 3, 2, 3, 2, 7, 2>(CHEB_ARGS);
        if (var == 41) return cheb_mixed<9, 3, 4, 3, 2, 3, 2, 6, 3>(CHEB_ARGS);
        if (var == 42) return cheb_mixed<9, 3, 4, 3, 2, 3, 1, 10, 2>(CHEB_ARGS);
        if (var == 43) return cheb_mixed<9, 3, 4, 3, 2, 3, 2, 5, 2>(CHEB_ARGS);
        if (var == 44) return cheb_mixed<9, 3, 2, 6, 2, 3, 2, 6, 2>(CHEB_ARGS);
        if (var == 45) return cheb_mixed<9, 3, 2, 5, 2, 3, 2, 6, 2>(CHEB_ARGS);
        if (var == 46) return cheb_mixed<9, 3, 1, 10, 2, 3, 1, 12, 2>(CHEB_ARGS);
        // measured at 256^3, cold start to 1e-12 (33 matrix passes), packed iterates: iteration shape <3,4,3,2> (12 warps per CTA,
        // 3 CTAs/SM) 43.5 ms per solve, barrier stall 3.9 per issue (ncu); <3,2,6,2> (6 warps per CTA, 6 CTAs/SM) 39.1;
        // <3,1,10..12,2> 39.0; plain loads instead of the TMA stage 54-70 (profiles/r2_cheb9_variants.txt)
        if (var == 30) return cheb_mixed<9, 3, 4, 3, 2, 3, 4, 3, 2>(CHEB_ARGS);
        return cheb_mixed<9, 3, 4, 3, 2, 3, 2, 6, 2>(CHEB_ARGS);
    }
  }
  switch (ncomp) {
    case 1: return cheb_soa<1, 1, 8, 4, 0>(CHEB_ARGS);
    case 2: return cheb_soa<2, 2, 8, 4, 0>(CHEB_ARGS);
    case 3: return cheb_soa<3, 3, 8, 6, 0>(CHEB_ARGS);
    case 4: return cheb_soa<4, 4, 8, 4, 0>(CHEB_ARGS);
    case 9:
      // measured at 256^3, ms per iteration: plain loads <9,3,2,8> 2.8-3.0; TMA-staged indices + values, 2 stages,
      // 4 slices x 3 component groups per CTA, 56 registers: 2.0; indices only: 2.9; 3 stages: 2.9 (the larger
      // carve-out takes the L1 away from the gathered vector)
      if (var == 1) return cheb_soa<9, 3, 4, 3, 2, false>(CHEB_ARGS);
      if (var == 2) return cheb_soa<9, 3, 2, 6, 2, true>(CHEB_ARGS);
      if (var == 6) return cheb_soa<9, 3, 2, 8, 0>(CHEB_ARGS);
      return cheb_soa<9, 3, 4, 3, 2, true>(CHEB_ARGS);
  }
#undef CHEB_ARGS
  set_error("chebyshev: unsupported ncomp %d (1,2,3,4,9)", ncomp);
  return MV_ERR_UNSUPPORTED;
}

}  // namespace mv

extern "C" int mv_set_mass_solver(int kind) {
  if (kind != MV_MASS_PCG && kind != MV_MASS_CHEBYSHEV && kind != MV_MASS_CHEBYSHEV_FP64) return MV_ERR_BADARG;
  mv::g_mass_solver = kind == MV_MASS_PCG ? MV_MASS_PCG : MV_MASS_CHEBYSHEV;
  mv::g_mass_mixed = kind == MV_MASS_CHEBYSHEV_FP64 ? 0 : (kind == MV_MASS_CHEBYSHEV ? 1 : mv::g_mass_mixed);
  return MV_OK;
}
