// Internal declarations shared by the translation units of libmovement_b200.so.
// Nothing here is part of the ABI (see include/movement_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "movement_b200.h"

#define MV_MAX_ROW 96        // max neighbours (incl. self) of one vertex
#define MV_SLICE 32          // SELL slice height == warp size
#define MV_BLOCK 256
#define MV_MAX_COMP 9
#define MV_RED_SLOTS 24      // doubles per block in the partials buffer
#define MV_MAX_BLOCKS 2368   // 148 SMs x 16: upper bound on persistent grids
#define MV_SCHED_TILE 16     // slices per tile of the slice schedule (mv_sched.cu)

namespace mv {

void set_error(const char* fmt, ...);
extern cudaStream_t g_stream;
extern int g_nsm;
extern long long g_launches;   // kernels launched by this library (every launch site is prefixed MV_L)
#define MV_L (void)++mv::g_launches,

// Optional live timing of the SpMV launches inside PCG (bench.py's roofline figure): CUDA events
// on the launching stream around each sampled launch, drained whenever the host polls the solve.
struct Prof {
  bool on = false;
  static const int NPAIR = 128;
  cudaEvent_t ev[2 * NPAIR];
  int cls[NPAIR];
  int used = 0;
  bool created = false;
  double ms[16] = {0};       // by number of right-hand sides
  long long count[16] = {0};
  long long pcg_iters[16] = {0};
};
extern Prof g_prof;
// True-residual audit of the preconditioned CG (mv_true_residual_*): at the exit of a solve the Euclidean
// ||P(b - A x)|| / ||P b|| of the RETURNED iterate is recomputed with one extra SpMV (the recurrence r -= alpha q
// that the stopping rule watches can drift from it), indexed by the number of right-hand sides like Prof.
struct TrueRes {
  bool on = false;
  double max_rel[16] = {0};
  double last_rel[16] = {0};
  long long count[16] = {0};
};
extern TrueRes g_trueres;
int prof_begin(int ncomp);   // returns slot or -1
void prof_end(int slot);
void prof_drain();           // stream must be synchronised

#define MV_CUDA(call)                                                                  \
  do {                                                                                 \
    cudaError_t e__ = (call);                                                          \
    if (e__ != cudaSuccess) {                                                          \
      mv::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
      return MV_ERR_CUDA;                                                              \
    }                                                                                  \
  } while (0)

#define MV_CHECK(call)              \
  do {                              \
    int s__ = (call);               \
    if (s__ != MV_OK) return s__;   \
  } while (0)

#define MV_KERNEL_CHECK() MV_CUDA(cudaGetLastError())

// ---------------------------------------------------------------------------------
// device memory helpers
// ---------------------------------------------------------------------------------
// Device memory comes from the CUDA stream-ordered pool (cudaMallocAsync / cudaFreeAsync on g_stream): the set-up
// phases (ingest, multigrid hierarchy) allocate and drop gigabytes of temporaries, and with cudaMalloc/cudaFree
// every free synchronised the device and every second mover paid the driver for the same pages again (multigrid
// set-up 260 ms for the first mover of a process, 680 ms for the next one, round 1).  The pool keeps up to
// MV_POOL_KEEP_GB (default: half of the device memory) of freed memory for reuse.
int pool_init();
template <typename T>
int dev_alloc(T** p, size_t n) {
  *p = nullptr;
  if (n == 0) n = 1;
  MV_CHECK(pool_init());
  MV_CUDA(cudaMallocAsync((void**)p, n * sizeof(T), g_stream));
  return MV_OK;
}
template <typename T>
void dev_free(T*& p) {
  if (p) cudaFreeAsync(p, g_stream);
  p = nullptr;
}
bool is_device_ptr(const void* p);

// Stage an input that may live on the host: returns a device pointer valid until the
// Staged object dies (async copy on g_stream).
struct Staged {
  void* dev = nullptr;
  bool owned = false;
  ~Staged() {
    if (owned && dev) cudaFreeAsync(dev, g_stream);
  }
  int in(const void* src, size_t bytes);              // host or device -> device (read)
  int out_begin(void* dst, size_t bytes);             // prepare device buffer for an output
  int out_end(void* dst, size_t bytes);               // copy back if dst is host
  template <typename T>
  T* as() { return (T*)dev; }
};

// ---------------------------------------------------------------------------------
// structures
// ---------------------------------------------------------------------------------
struct Sell {
  int64_t nrows = 0;       // owned rows
  int64_t nslices = 0;
  int64_t len = 0;         // padded length
  int64_t nnz = 0;
  int maxw = 0;
  int64_t* slice_ptr = nullptr;  // [nslices+1]
  int* cols = nullptr;           // [len]
  int* row_nnz = nullptr;        // [nrows]
  // slice schedule (mv_sched.cu, mesh pattern only): tiles of MV_SCHED_TILE slice indices, -1 = empty slot
  int* sched = nullptr;          // [sched_len]
  int64_t sched_len = 0;
  int sched_tiled = 0;           // 1: lattice tiling, 0: identity
};

struct Halo {
  int nn = 0;
  std::vector<int> ranks;
  std::vector<int64_t> send_ptr, recv_ptr;  // host copies
  int* send_idx = nullptr;                  // device [send_ptr[nn]]
  double* send_buf = nullptr;               // device [send_ptr[nn]*MV_MAX_COMP]
  int p2p_maxC = 0;                         // components per exchange that fit the peer-memory slots on EVERY rank (0: NCCL path)
};

struct Reduce {
  double* partials = nullptr;   // [MV_MAX_BLOCKS*MV_RED_SLOTS]
  unsigned* counter = nullptr;  // ticket for the last-block pattern
};

struct PcgWork {
  double *r = nullptr, *p = nullptr, *q = nullptr;  // [V*MV_MAX_COMP] (p spans ghosts)
  double* xs = nullptr;     // component-major copy of the iterate (mv_pcg_soa.cu)
  double* scal = nullptr;   // device scalars, see pcg.cu
  int* flags = nullptr;     // device: [0]=done, [1]=iterations, [2]=breakdown
  int* h_flags = nullptr;   // pinned host mirror
  double* h_scal = nullptr; // pinned host mirror
  // mixed-precision Chebyshev (mv_cheb.cu): residual and two correction iterates in single, [V*MV_MAX_COMP] each,
  // single-precision copy of the operator values / inverse diagonal the sweeps stream
  // CUDA graph of ONE multigrid-preconditioned CG iteration (mv_pcg_soa.cu): captured once per (operator, hierarchy,
  // buffers) and replayed, so the ~40 small launches of an iteration cost one
  void* graph_exec = nullptr;          // cudaGraphExec_t
  unsigned long long graph_key[12] = {0};
  long long graph_launches = 0;        // kernels inside the graph (launch accounting)
  bool graph_failed = false;           // capture / instantiation failed once: plain launches from then on
  float* f32[3] = {nullptr, nullptr, nullptr};
  float* valf = nullptr;
  float* dinvf = nullptr;
  const double* f32_src = nullptr;
};

// ---------------------------------------------------------------------------------
// smoothed-aggregation multigrid hierarchy (mv_amg.cu builds it, mv_amg_cycle.cu applies it)
// ---------------------------------------------------------------------------------
struct Csr {
  int64_t nrows = 0, ncols = 0, nnz = 0;
  int64_t* ptr = nullptr;   // [nrows+1]
  int* col = nullptr;       // sorted ascending inside a row
  double* val = nullptr;
  void release();
};

struct AmgLevel {
  int64_t n = 0;            // owned rows
  int64_t ntot = 0;         // owned + ghost columns of A
  Sell pat;                 // operator of this level (level 0 borrows the mesh pattern)
  const double* val = nullptr;
  float* valf = nullptr;    // single-precision copy of val for the V-cycle's matrix streams (smoothed levels; mv_amg_cycle.cu)
  bool owns_A = false;
  double* dinv = nullptr;   // [n]; 0 at constrained vertices
  const uint16_t* mask = nullptr;  // level 0 only: bit 0 set = constrained vertex
  double rho = 0.0;         // upper bound of spec(D^-1 A)
  int degree = 1;           // Chebyshev smoother degree
  Sell P;  double* Pval = nullptr;   // [n x n_coarse]
  Sell R;  double* Rval = nullptr;   // [n_coarse x n] = P^T
  Csr Rc;                   // the same in CSR when its rows are long (coarse levels): applied one warp per row
  Csr Ac;                   // small coarse levels with long rows: the operator in CSR, smoothed one warp per row
  double *x0 = nullptr, *x1 = nullptr, *b = nullptr, *r = nullptr, *d0 = nullptr, *d1 = nullptr;  // [ntot*C]
  int work_C = 0;
  // multi-GPU: forward halo of this level (level 0 borrows the mesh's), global numbering
  Halo halo;
  bool owns_halo = false;
  bool dist = false;        // operator rows are spread over the ranks (halo exchange before every gather of a vector)
  // replicated level whose matrix passes are split over the ranks (rows [split_lo, split_hi) here, vectors
  // allgathered in place after each pass; ntot is padded to split_chunk * ranks); 0 = every rank does every row
  int64_t split_chunk = 0, split_lo = 0, split_hi = 0;
  bool reduce_next = false; // next level is replicated on every rank: its right-hand side is the allreduce of the partial restrictions
  int64_t nglobal = 0;      // rows over all ranks
  int64_t goff = 0;         // global index of owned row 0
  int64_t* gid = nullptr;   // [ntot] global index of every local column (levels >= 1)
};

struct Amg {
  std::vector<AmgLevel> lv;
  int op = 0;
  int singular = 0;
  uint64_t mask_hash = 0;
  uint16_t* mask_copy = nullptr;  // level-0 constrained set (bit 0), owned by the hierarchy
  int64_t nc = 0;                 // coarsest size
  double* coarse_inv = nullptr;   // dense [nc*nc] inverse (shifted if singular), or nullptr -> smoother only
  double* z = nullptr;            // [n0*C] preconditioned residual for PCG
  int z_C = 0;
  double* cglob = nullptr;        // multi-GPU coarse solve: [2][C][nc] gathered right-hand side / scratch
  double setup_ms = 0.0;
  unsigned long long uid = 0;     // unique per built hierarchy (a rebuilt one may reuse the address)
  ~Amg();
};

}  // namespace mv

struct mv_mesh {
  int dim = 0;
  int64_t V = 0, C = 0, F = 0, Vown = 0, Cown = 0;
  // periodic meshes (movement/mover.py:70-75, monge_ampere.py:236-239): `cells` is the periodic P1 map, the coordinate
  // field is DG1 -- node c*(dim+1)+j belongs to corner j of cell c, Vc = C*(dim+1) coordinate nodes (dg = 1);
  // rep[v] = the coordinate node of P1 vertex v in the last cell that contains it
  int dg = 0;
  int64_t Vc = 0;
  int* rep = nullptr;       // [V] (dg only)
  double* xi = nullptr;     // [Vc*dim] coordinates at ingest (computational mesh)
  int* cells = nullptr;     // [C*(dim+1)]
  int* v2c_ptr = nullptr;   // [Vown+1]
  int* v2c = nullptr;       // entries cell*(dim+1)+loc, ascending per vertex
  // boundary facets
  int* bf_verts = nullptr;  // [F*dim]
  int* bf_tags = nullptr;   // [F]
  int* bf_cell = nullptr;   // [F]
  double* bf_normal = nullptr;  // [F*dim]
  double* bf_area = nullptr;    // [F]
  int* v2f_ptr = nullptr;   // [Vown+1]
  int* v2f = nullptr;       // entries f*dim+loc
  std::vector<int> h_bf_tags;   // host copy (small)
  // operators
  mv::Sell pat;
  double* Mval = nullptr;
  double* Kval = nullptr;
  double* MKval = nullptr;   // lazily built M+K
  double* ml = nullptr;      // [Vown] lumped mass
  double* vol = nullptr;     // [C] cell volumes on xi
  double* dinvM = nullptr;   // [Vown]
  double* dinvK = nullptr;
  double* dinvMK = nullptr;
  double total_volume = 0, ml_sumsq = 0, ml_sum = 0;
  // spring data (lazy)
  int* edge_opp = nullptr;   // [sell_len] opposite vertex of the '+' cell for (row,col)
  // assembled spring stiffness (xx, xy, yy blocks on the mesh pattern) and the block-diagonal multigrid preconditioner
  // of its solves (mv_movers.cu): hierarchies of the xx and yy blocks, kept across calls while they still precondition well
  double* spring_val[3] = {nullptr, nullptr, nullptr};
  mv::Amg* spring_pc[2] = {nullptr, nullptr};
  uint16_t* spring_mask[2] = {nullptr, nullptr};   // bit 0 = component constrained (the hierarchies' level-0 masks)
  int spring_pc_its = 0;     // PCG iterations of the solve that followed the last set-up
  int spring_pc_age = 0;     // solves since
  int spring_pc_age_its = 0; // PCG iterations of the latest solve
  int64_t* cell_gid = nullptr;   // [C] global cell numbers (partitioned meshes only; mv_mesh_set_cell_order)
  // work
  mv::Reduce red;
  mv::PcgWork pcg;
  mv::Halo halo;
  double* tmp_vec = nullptr;   // [V*MV_MAX_COMP]
  double* cell_tmp = nullptr;  // [C*(dim+1)] per-cell vertex contributions of the cell-major assemblies
  // preconditioner of the stiffness-type operators: MV_PC_JACOBI or MV_PC_AMG
  int pc_kind = 1;
  std::vector<mv::Amg*> amg;   // cached hierarchies (one per operator / constrained set)
};

namespace mv { struct EqBC; }
struct mv_ma {
  mv_mesh* mesh = nullptr;
  mv::EqBC* eqbc = nullptr;     // non-axis-aligned free boundary segments (mv_eqbc.cu)
  double* f[MV_F_COUNT] = {nullptr};
  uint16_t* bc_mask = nullptr;  // [V]
  bool have_bc = false;
  double* rhs = nullptr;        // [V*MV_MAX_COMP] assembled right-hand sides
  double* a = nullptr;          // [V] assembled int psi m det(I+H)
  double* dphi = nullptr;       // [V]
  // the two previous Poisson solves (update, right-hand side): Galerkin initial guess of the next one
  double* hist_x[8] = {nullptr};   // [V] previous updates, newest first
  double* hist_b[8] = {nullptr};   // [V] their right-hand sides (= K x_i up to a constant)
  double* hist_G = nullptr;        // [8*8] Gram matrix x_i . b_j
  int hist_n = 0;
  // right-hand side of the next L2 projection, produced by the Hessian right-hand-side kernel inside mv_ma_relax_move
  double* l2_rhs = nullptr;     // [V*d]
  const double* l2_rhs_phi = nullptr;   // the phi buffer it was computed from
  bool l2_rhs_valid = false, fuse_l2_rhs = false;
  double* mon_raw = nullptr;    // [V]
  double* dscal = nullptr;      // device scalars
  double* h_scal = nullptr;     // pinned
  double gconst[3] = {0, 0, 0};  // allreduced constants of the mesh (cells, total volume, sum m_l^2); valid when gconst_ok
  bool gconst_ok = false;
  int mon_kind = MV_MON_USER;
  double mon_params[8] = {0};
  double rtol = 1e-12, atol = 1e-50;
  int maxit_mass = 500, maxit_poisson = 100000;
  // quasi-Newton workspace (lazy, mv_qn.cu)
  double* qn_basis = nullptr;   // V_0..V_m then Z_0..Z_{m-1}, each Vown*(1+d*d)
  int qn_restart = 0;
  double *qn_w = nullptr, *qn_gphi = nullptr, *qn_gH = nullptr, *qn_t1 = nullptr, *qn_t2 = nullptr;
  double* qn_base = nullptr;    // iterate saved by mv_qn_save
  double* qn_dots = nullptr;    // device [64]
  double* qn_hdots = nullptr;   // pinned [64]
};

namespace mv {

// device scalar slots of mv_ma::dscal (mv_ma.cu, mv_qn.cu)
enum { D_MONMAX = 0, D_SUMA = 1, D_THETA = 2, D_RES2 = 3, D_VSUM = 4, D_VMAX = 5, D_VMIN = 6, D_VSSD = 7, D_N = 32 };

// EquationBC branch of the L2 projector (mv_eqbc.cu)
int eqbc_build(mv_ma* ma, const std::vector<int>& tagidx, const std::vector<char>& is_eq, const std::vector<int>& verts,
               const std::vector<double>& area);
int eqbc_apply(mv_ma* ma, const double* phi, double* grad_phi);
void eqbc_free(mv_ma* ma);
int64_t eqbc_counts(const mv_ma* ma, int64_t* nu);

// PCG device scalar slots: scal[slot*16 + comp]
enum { S_RZ = 0, S_PQ = 1, S_RZNEW = 2, S_BETA = 3, S_THR = 4, S_REF = 5, S_BB = 6, S_SUM = 7, S_NSLOTS = 8 };
#define SC(slot, c) ((slot) * 16 + (c))

struct LinOp {
  const double* val = nullptr;            // SELL values; nullptr selects the matrix-free spring operator
  const double* dinv = nullptr;           // inverse diagonal, per vertex (dpd=0) or per dof (dpd=1)
  int dpd = 0;
  int coupled = 0;                        // components form one vector (one alpha/beta)
  const double* spring_coords = nullptr;
};
int spring_apply_dev(mv_mesh* m, const double* coords, const uint16_t* mask, const double* x, double* y, bool dot);
int pcg_solve_spring(mv_mesh* m, const double* coords, const double* dinv_dof, const double* rhs, double* x, const uint16_t* mask,
                     double rtol, double atol, int maxit, int* iters);
int ensure_pcg_workspace(mv_mesh* m);
void pcg_graph_release(PcgWork& w);   // captured iteration graph of a mesh (mv_pcg_soa.cu)
void pcg_graph_release_all();         // before the communicator goes away
int exclusive_scan_i32(const int* in, int* out, int64_t n);

// host-side entry points shared across translation units
int pcg_solve(mv_mesh* m, int op, int ncomp, const double* rhs, double* x, const uint16_t* mask,
              double rtol, double atol, int maxit, int project_mean, int* iters, double* relres);
int spmv(mv_mesh* m, int op, int ncomp, const double* x, double* y);
int pcg_solve_soa(mv_mesh* m, int ncomp, const double* val, const double* dinv, const double* rhs, double* x, const uint16_t* mask,
                  double rtol, double atol, int maxit, int project_mean, int* iters, double* relres, Amg* pc);
int aos_to_soa(int C, int64_t n, const double* src, double* dst, int64_t ld);
int soa_to_aos(int C, int64_t n, const double* src, int64_t ld, double* dst);
int ensure_slice_schedule(mv_mesh* m);
// fused Chebyshev iteration for the consistent mass matrix (mv_cheb.cu)
extern int g_mass_solver;
extern int g_mass_mixed;
int cheb_mass_solve(mv_mesh* m, int ncomp, const double* val, const double* dinv, const double* rhs, double* x,
                    const uint16_t* mask, double rtol, double atol, int maxit, int* iters, double* relres);
int halo_exchange_soa(const Halo& h, int64_t nown, int C, double* vec, int64_t ld);
int halo_finalize(Halo& h);   // collective: decides whether the peer-memory path can carry this halo
// peer-memory collectives (mv_p2p.cu)
bool p2p_active();
int p2p_init(int rank, int nranks, int (*allgather_bytes)(const void* sendbuf, void* recvbuf, size_t bytes));
void p2p_finalize();
size_t p2p_slot_bytes();
int p2p_halo_exchange(const Halo& h, int64_t nown, int C, double* vec, int64_t ld, bool soa, const int* done);
int p2p_halo_exchange_f32(const Halo& h, int64_t nown, int C, float* vec, int64_t ld, const int* done);
int halo_exchange_soa_f32(const Halo& h, int64_t nown, int C, float* vec, int64_t ld);
int p2p_halo_exchange_f32_packed4(const Halo& h, int64_t nown, int groups, float* vec, int64_t ld, const int* done);
bool p2p_allreduce_fits(int n);
int p2p_allreduce(double* d, int n, int op, const int* done);
int p2p_check_status();
void p2p_counts(long long* halo, long long* red);
int op_arrays(mv_mesh* m, int op, const double** val, const double** dinv);
int halo_exchange(mv_mesh* m, int ncomp, double* vec);
int allreduce_sum_dev(double* dscal, int n);
int allgather_rows_dev(double* vec, int64_t ld, int C, int64_t chunk);
int allreduce_dev(double* dscal, int n, int op);  // op: 0 sum, 1 max, 2 min
int allreduce_volstats_dev(double* dout);  // [0] sum, [1] max, [2] min
int volume_stats_device(mv_mesh* m, const double* x, const double* mon, const double* vol0, double* volume,
                        double* dout, int64_t global_cells);
int l2_rhs_dev(mv_mesh* m, const double* phi, double* b);
int hess_rhs_dev(mv_mesh* m, const double* phi, double* c, double* l2_out = nullptr);
int assemble_a_dev(mv_mesh* m, const double* mon, const double* H, double* a, double* dscal);
int jac_A_dev(mv_mesh* m, const double* mon, const double* H, const double* dH, double sign, double* out);
bool comm_active();
int comm_size();
int comm_rank();
int allreduce_host(double* h, int n, int op);
int allgather_i64(int64_t mine, std::vector<int64_t>& all);
int bcast_bytes(void* buf, size_t bytes, int root);
int neighbor_exchange_bytes(int nn, const int* ranks, const void* const* sbuf, const size_t* sbytes, void* const* rbuf,
                            const size_t* rbytes);
int grid_for(int64_t nthreads_needed, int blocks_per_sm = 8);
// run a statement with DIM = the mesh dimension as a compile-time constant
#define MV_DISPATCH_DIM(dim, ...)                                   \
  do {                                                              \
    if ((dim) == 1) { constexpr int DIM = 1; __VA_ARGS__; }         \
    else if ((dim) == 2) { constexpr int DIM = 2; __VA_ARGS__; }    \
    else { constexpr int DIM = 3; __VA_ARGS__; }                    \
  } while (0)
int exclusive_scan_i64(const int64_t* in, int64_t* out, int64_t n);
// multigrid (mv_amg.cu / mv_amg_cycle.cu)
int amg_get(mv_mesh* m, int op, const uint16_t* mask, int ncomp, int singular, Amg** out);  // *out = nullptr -> use Jacobi
int amg_apply(mv_mesh* m, Amg* a, int ncomp, const double* r, double* z, const int* done);
int amg_ensure_work(Amg* a, int ncomp);
void amg_release(mv_mesh* m);
int amg_build_values(mv_mesh* m, int op, const double* val, const uint16_t* mask_bit0, int singular, Amg** out);
void spring_pc_release(mv_mesh* m);

// ---------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------
#ifdef __CUDACC__

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Deterministic grid-wide sum of N doubles per thread.
// Every thread of every block must call this (no early exits).  The block that takes
// the last ticket adds the per-block partials in a fixed order and hands the totals to
// `fin(total)` (executed by thread 0 of that block).  Results are independent of block
// scheduling order, so repeated runs give identical bits.
template <int N, typename Fin>
__device__ __forceinline__ void grid_reduce_sum(double (&v)[N], double* partials,
                                                unsigned* counter, Fin fin) {
  static_assert(N <= MV_RED_SLOTS, "too many reduction slots");
  __shared__ double s_red[MV_BLOCK / 32][MV_RED_SLOTS];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
  for (int n = 0; n < N; ++n) {
    double t = warp_sum(v[n]);
    if (lane == 0) s_red[wid][n] = t;
  }
  __syncthreads();
  if (threadIdx.x < N) {
    double t = 0.0;
    for (int w = 0; w < nw; ++w) t += s_red[w][threadIdx.x];
    partials[(size_t)blockIdx.x * MV_RED_SLOTS + threadIdx.x] = t;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned ticket = atomicAdd(counter, 1u);
    s_last = (ticket == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  double tot[N];
#pragma unroll
  for (int n = 0; n < N; ++n) {
    double t = 0.0;
    for (unsigned b = threadIdx.x; b < gridDim.x; b += blockDim.x)
      t += __ldcg(&partials[(size_t)b * MV_RED_SLOTS + n]);
    t = warp_sum(t);
    if (lane == 0) s_red[wid][n] = t;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int n = 0; n < N; ++n) {
      double t = 0.0;
      for (int w = 0; w < nw; ++w) t += s_red[w][n];
      tot[n] = t;
    }
    *counter = 0u;
    fin(tot);
  }
}

// atomic max / min for non-negative doubles through their order-preserving bit pattern
__device__ __forceinline__ void atomic_max_pos(double* addr, double v) {
  atomicMax((unsigned long long*)addr, (unsigned long long)__double_as_longlong(v));
}
__device__ __forceinline__ void atomic_min_pos(double* addr, double v) {
  atomicMin((unsigned long long*)addr, (unsigned long long)__double_as_longlong(v));
}

// P1 geometry of one simplex: volume and basis-function gradients.
template <int D>
struct CellGeom {
  double vol;
  double det;
  double g[D + 1][D];
};

// coordinates of the corners of cell c: through the P1 map (cv), or, on a periodic mesh (dg), the cell's own DG nodes
template <int D>
__device__ __forceinline__ void load_cell_coords(const double* __restrict__ x, const int* cv,
                                                 double (&xc)[D + 1][D], int dg = 0, int64_t c = 0) {
#pragma unroll
  for (int j = 0; j <= D; ++j) {
    const size_t node = dg ? (size_t)c * (D + 1) + j : (size_t)cv[j];
#pragma unroll
    for (int a = 0; a < D; ++a) xc[j][a] = __ldg(&x[node * D + a]);
  }
}

__device__ __forceinline__ void cell_geom(const double (&xc)[2][1], CellGeom<1>& o) {
  const double det = xc[1][0] - xc[0][0];
  o.det = det;
  o.vol = fabs(det);
  o.g[1][0] = 1.0 / det;
  o.g[0][0] = -o.g[1][0];
}

__device__ __forceinline__ void cell_geom(const double (&xc)[3][2], CellGeom<2>& o) {
  const double a = xc[1][0] - xc[0][0], b = xc[2][0] - xc[0][0];
  const double c = xc[1][1] - xc[0][1], d = xc[2][1] - xc[0][1];
  const double det = a * d - b * c;
  const double inv = 1.0 / det;
  o.det = det;
  o.vol = 0.5 * fabs(det);
  o.g[1][0] = d * inv;
  o.g[1][1] = -b * inv;
  o.g[2][0] = -c * inv;
  o.g[2][1] = a * inv;
  o.g[0][0] = -(o.g[1][0] + o.g[2][0]);
  o.g[0][1] = -(o.g[1][1] + o.g[2][1]);
}

__device__ __forceinline__ void cell_geom(const double (&xc)[4][3], CellGeom<3>& o) {
  double e1[3], e2[3], e3[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    e1[a] = xc[1][a] - xc[0][a];
    e2[a] = xc[2][a] - xc[0][a];
    e3[a] = xc[3][a] - xc[0][a];
  }
  // rows of J^-1 are (e2 x e3, e3 x e1, e1 x e2)/det
  double c1[3] = {e2[1] * e3[2] - e2[2] * e3[1], e2[2] * e3[0] - e2[0] * e3[2], e2[0] * e3[1] - e2[1] * e3[0]};
  double c2[3] = {e3[1] * e1[2] - e3[2] * e1[1], e3[2] * e1[0] - e3[0] * e1[2], e3[0] * e1[1] - e3[1] * e1[0]};
  double c3[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
  const double det = e1[0] * c1[0] + e1[1] * c1[1] + e1[2] * c1[2];
  const double inv = 1.0 / det;
  o.det = det;
  o.vol = fabs(det) * (1.0 / 6.0);
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    o.g[1][a] = c1[a] * inv;
    o.g[2][a] = c2[a] * inv;
    o.g[3][a] = c3[a] * inv;
    o.g[0][a] = -(o.g[1][a] + o.g[2][a] + o.g[3][a]);
  }
}

template <int D>
__device__ __forceinline__ void load_cell(const int* __restrict__ cells, int64_t c, int (&cv)[D + 1]) {
#pragma unroll
  for (int j = 0; j <= D; ++j) cv[j] = __ldg(&cells[c * (D + 1) + j]);
}

#endif  // __CUDACC__

}  // namespace mv
