// Slice schedule: the ORDER in which the SELL-32 slices of the mesh pattern are handed to the CTAs of the
// gather-heavy kernels (mv_cheb.cu).  Any permutation is correct; a good one makes the 16 slices a CTA works
// on at the same time share their neighbour columns, so that the gathered vector is fetched from L2 once per
// CTA instead of once per slice.
//
// Why: with c = 9 right-hand sides a row gathers 135 values.  In the natural (lexicographic) numbering of
// UnitSquareMesh / UnitCubeMesh a slice of 32 consecutive rows reads 3 (2D) / 7 (3D) windows of
// the vector, one per neighbouring mesh line; consecutive slices lie on the same line and share nothing but
// the x+-1 overlap, and ncu (round 1) shows the kernel bound by L2->L1 traffic (L1 hit rate 37 %, 15 GB over
// the crossbar per launch for 6.5 GB of DRAM traffic).  A tile of 4 x 4 lines (y x z) at the same x-range
// touches 34 windows instead of 16 x 7 = 112.
//
// The tiling needs the line length L and the plane size P of the numbering.  They are READ OFF THE PATTERN
// (column offsets of a row in the middle of the owned range), not assumed: a mesh whose pattern does not
// look lexicographic (offsets {1, L, L+1} in 2D, {1, L, L+1, P, ..} with L | P in 3D) gets the identity
// schedule, i.e. consecutive slices per CTA as before.  The table is verified on the device (every slice
// exactly once) before it is used.
#include <algorithm>
#include <vector>

#include "mv_internal.cuh"

namespace mv {

// tile t = ((kt * NJT) + jt) * A + a ; slot q: bits (dj0, dk0, dj1, dk1) in 3D, dj in 2D
__global__ void k_build_sched(int dim, int64_t L, int64_t NY, int64_t NZ, int64_t A, int64_t NJT, int64_t ntiles, int64_t nslices,
                              int64_t nrows, int* __restrict__ sched, int* __restrict__ seen) {
  const int64_t n = ntiles * MV_SCHED_TILE;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t t = e / MV_SCHED_TILE;
    const int q = (int)(e % MV_SCHED_TILE);
    const int64_t a = t % A, jt = (t / A) % NJT, kt = t / (A * NJT);
    int64_t j, k;
    if (dim == 3) {
      const int dj = (q & 1) | ((q >> 1) & 2), dk = ((q >> 1) & 1) | ((q >> 2) & 2);
      j = 4 * jt + dj;
      k = 4 * kt + dk;
    } else {
      j = MV_SCHED_TILE * jt + q;
      k = 0;
    }
    int s = -1;
    if (j < NY && k < NZ) {
      const int64_t line = (k * NY + j) * L;
      const int64_t lo = line + 32 * a;
      const int64_t hi = lo + 32 < line + L ? lo + 32 : line + L;
      const int64_t c = (lo + 31) / 32;
      if (32 * c < hi && c < nslices && 32 * c < nrows) s = (int)c;
    }
    sched[e] = s;
    if (s >= 0) atomicAdd(&seen[s], 1);
  }
}
__global__ void k_sched_identity(int64_t nslices, int64_t len, int* __restrict__ sched) {
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < len; e += (int64_t)gridDim.x * blockDim.x)
    sched[e] = e < nslices ? (int)e : -1;
}
__global__ void k_sched_verify(int64_t nslices, const int* __restrict__ seen, int* __restrict__ bad) {
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < nslices; e += (int64_t)gridDim.x * blockDim.x)
    if (seen[e] != 1) atomicAdd(bad, 1);
}

__global__ void k_max_slice_width(int64_t nslices, const int64_t* __restrict__ slice_ptr, int* __restrict__ out) {
  int w = 0;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < nslices; e += (int64_t)gridDim.x * blockDim.x) {
    const int we = (int)((slice_ptr[e + 1] - slice_ptr[e]) / MV_SLICE);
    w = we > w ? we : w;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const int t = __shfl_xor_sync(0xffffffffu, w, o);
    w = t > w ? t : w;
  }
  if ((threadIdx.x & 31) == 0 && w > 0) atomicMax(out, w);
}

// Line length and plane size of a lexicographic numbering, from the column offsets of one interior row.
static bool detect_lattice(mv_mesh* m, int64_t row, int64_t* L, int64_t* P) {
  const Sell& p = m->pat;
  const int64_t s = row / MV_SLICE;
  const int lane = (int)(row % MV_SLICE);
  int64_t h_ptr[2];
  if (cudaMemcpyAsync(h_ptr, p.slice_ptr + s, 2 * sizeof(int64_t), cudaMemcpyDeviceToHost, g_stream) != cudaSuccess) return false;
  if (cudaStreamSynchronize(g_stream) != cudaSuccess) return false;
  const int64_t len = h_ptr[1] - h_ptr[0];
  if (len <= 0 || len > (int64_t)MV_SLICE * MV_MAX_ROW) return false;
  std::vector<int> cols((size_t)len);
  if (cudaMemcpyAsync(cols.data(), p.cols + h_ptr[0], len * sizeof(int), cudaMemcpyDeviceToHost, g_stream) != cudaSuccess) return false;
  if (cudaStreamSynchronize(g_stream) != cudaSuccess) return false;
  std::vector<int64_t> off;
  for (int64_t k = 0; k < len / MV_SLICE; ++k) {
    const int64_t o = (int64_t)cols[(size_t)(k * MV_SLICE + lane)] - row;
    if (o > 0) off.push_back(o);
  }
  std::sort(off.begin(), off.end());
  off.erase(std::unique(off.begin(), off.end()), off.end());
  if (m->dim == 2) {
    // one diagonal per quadrilateral: {1, L, L+1} or {1, L-1, L}; whole lines are owned, so L divides the row count
    if (off.size() != 3 || off[0] != 1 || off[2] != off[1] + 1 || off[1] < 2 * MV_SLICE) return false;
    *L = (p.nrows % off[2] == 0 && p.nrows % off[1] != 0) ? off[2] : off[1];
    *P = 0;
    return true;
  }
  // six Kuhn tetrahedra per hexahedron: {1, L, L+1, P, P+1, P+L+1} (+ P+L for other splittings)
  if (off.size() < 6 || off[0] != 1 || off[2] != off[1] + 1 || off[1] < 2 * MV_SLICE) return false;
  *L = off[1];
  *P = off[3];
  if (*P < 2 * *L || *P % *L != 0) return false;
  for (size_t i = 3; i < off.size(); ++i) {
    const int64_t r = off[i] - *P;
    if (r != 0 && r != 1 && r != *L && r != *L + 1) return false;
  }
  return true;
}

int ensure_slice_schedule(mv_mesh* m) {
  Sell& p = m->pat;
  if (p.sched) return MV_OK;
  {  // widest slice (shared-memory tile of the TMA-staged kernels)
    int* d_w = nullptr;
    MV_CHECK(dev_alloc(&d_w, 1));
    MV_CUDA(cudaMemsetAsync(d_w, 0, sizeof(int), g_stream));
    MV_L k_max_slice_width<<<grid_for(p.nslices), MV_BLOCK, 0, g_stream>>>(p.nslices, p.slice_ptr, d_w);
    MV_KERNEL_CHECK();
    MV_CUDA(cudaMemcpyAsync(&p.maxw, d_w, sizeof(int), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    dev_free(d_w);
  }
  // measured at 256^3 (round 1): the lattice tiles do NOT pay -- c = 9: 3.2 ms per iteration tiled against 2.8 ms in
  // natural order (the kernel is bound by load latency, not by L2->L1 bytes) -- so the identity is the default and
  // MV_SCHED=1 (read when a mesh builds its schedule) selects the tiles for experiments and tests
  const char* e = getenv("MV_SCHED");
  const bool disable = !(e && atoi(e) == 1);
  int64_t L = 0, P = 0;
  bool tiled = !disable && p.nslices >= 4 * MV_SCHED_TILE && detect_lattice(m, p.nrows / 2, &L, &P);
  if (tiled) {
    const int64_t NY = m->dim == 3 ? P / L : (p.nrows + L - 1) / L;
    const int64_t NZ = m->dim == 3 ? (p.nrows + P - 1) / P : 1;
    const int64_t A = (L + 31) / 32;
    const int64_t NJT = m->dim == 3 ? (NY + 3) / 4 : (NY + MV_SCHED_TILE - 1) / MV_SCHED_TILE;
    const int64_t NKT = m->dim == 3 ? (NZ + 3) / 4 : 1;
    const int64_t ntiles = NKT * NJT * A;
    int *seen = nullptr, *bad = nullptr;
    MV_CHECK(dev_alloc(&p.sched, ntiles * MV_SCHED_TILE));
    MV_CHECK(dev_alloc(&seen, p.nslices + 1));
    MV_CUDA(cudaMemsetAsync(seen, 0, (p.nslices + 1) * sizeof(int), g_stream));
    bad = seen + p.nslices;
    MV_L k_build_sched<<<grid_for(ntiles * MV_SCHED_TILE), MV_BLOCK, 0, g_stream>>>(m->dim, L, NY, NZ, A, NJT, ntiles, p.nslices, p.nrows,
                                                                              p.sched, seen);
    MV_L k_sched_verify<<<grid_for(p.nslices), MV_BLOCK, 0, g_stream>>>(p.nslices, seen, bad);
    MV_KERNEL_CHECK();
    int h_bad = 1;
    MV_CUDA(cudaMemcpyAsync(&h_bad, bad, sizeof(int), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    dev_free(seen);
    if (h_bad == 0) {
      p.sched_len = ntiles * MV_SCHED_TILE;
      p.sched_tiled = 1;
      return MV_OK;
    }
    dev_free(p.sched);   // not a permutation: fall through to the identity
  }
  const int64_t len = (p.nslices + MV_SCHED_TILE - 1) / MV_SCHED_TILE * MV_SCHED_TILE;
  MV_CHECK(dev_alloc(&p.sched, len));
  MV_L k_sched_identity<<<grid_for(len), MV_BLOCK, 0, g_stream>>>(p.nslices, len, p.sched);
  MV_KERNEL_CHECK();
  p.sched_len = len;
  p.sched_tiled = 0;
  return MV_OK;
}

}  // namespace mv
