// mv_cell_volumes_static: the cell-volume diagnostic (mv_cell_volumes, mv_core.cu) for callers that must not allocate.
//
// Who cannot allocate, in this library's world:
//  * a rank of a multi-GPU run while a neighbour's peer-memory kernel (mv_p2p.cu) spins on a flag this rank has yet to
//    raise: cudaMalloc / cudaFree synchronise the whole device, so an allocation made at the wrong moment waits for a
//    kernel that is itself waiting for this rank;
//  * a process whose HBM already belongs to somebody else -- torch's caching allocator, or a 256^3 mover that holds
//    34 GB of operators and multigrid levels: a diagnostic must not be the call that fails with out-of-memory in the
//    middle of a run, nor make the pool hand pages back and forth every timestep;
//  * a host thread that is not allowed to block on the driver's allocator lock (per-timestep monitoring beside a
//    running mover).
// mv_cell_volumes needs a mesh handle (mv_mesh_create allocates all its operators) and takes its result slots from
// cudaMallocAsync on every call.  The variant below needs neither: connectivity and coordinates come straight from the
// caller's device arrays and all scratch (per-block partial sums / maxima / minima, four result slots, one ticket) lives
// in ONE file-scope __device__ array of fixed size.  The price: an upper bound on the number of cells, one shared
// scratch per GPU (calls that overlap on the device race), and ~768 KB of HBM held by every GPU that loads the library.
//
// Has no counterpart in the reference; the arithmetic is that of mv_cell_volumes (mover.py:199-217, spring.py:274,
// laplacian_smoothing.py:119 in the reference checkout).
#include "mv_internal.cuh"

namespace mv {

#define MV_STATIC_CHUNK 4096                    // cells per CTA: 256 threads x 16 cells, independent of the SM count
#define MV_STATIC_MAX_BLOCKS 32768              // -> at most 2^27 = 134 217 728 cells (8192^2 triangles, 256^3 tets fit)
#define MV_STATIC_MAX_CELLS ((int64_t)MV_STATIC_CHUNK * MV_STATIC_MAX_BLOCKS)
// scratch layout in doubles: [0,B) block sums (pass 1: volumes, pass 2: squared deviations), [B,2B) block maxima,
// [2B,3B) block minima, [3B..3B+3] sum / max / min / sum of squared deviations, [3B+4] ticket (unsigned, 8-byte slot)
#define MV_STATIC_RES (3 * MV_STATIC_MAX_BLOCKS)
#define MV_STATIC_TICKET (MV_STATIC_RES + 4)
#define MV_STATIC_DOUBLES (MV_STATIC_RES + 8)

// The scratch itself.  In HOST code this name is only a handle for cudaGetSymbolAddress -- it is not a device address
// and must never be passed to a kernel; scratch_address() below turns it into one, once per device.
static __device__ __align__(256) double g_static_scratch[MV_STATIC_DOUBLES];

#define MV_STATIC_MAX_DEVICES 64
static double* g_scratch_dev[MV_STATIC_MAX_DEVICES] = {nullptr};

// Device address of g_static_scratch on the CURRENT device; looked up on the first call per device.  The look-up makes
// the driver load this module on that device (which is when the array gets its memory), hence "first call before any
// stream capture" in the header.
static int scratch_address(double** out) {
  int dev = -1;
  MV_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MV_STATIC_MAX_DEVICES) {
    set_error("mv_cell_volumes_static: device ordinal %d outside 0..%d", dev, MV_STATIC_MAX_DEVICES - 1);
    return MV_ERR_UNSUPPORTED;
  }
  if (!g_scratch_dev[dev]) {
    void* p = nullptr;
    MV_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    if (!p) {
      set_error("mv_cell_volumes_static: cudaGetSymbolAddress returned a null address on device %d", dev);
      return MV_ERR_CUDA;
    }
    g_scratch_dev[dev] = (double*)p;
  }
  *out = g_scratch_dev[dev];
  return MV_OK;
}

// Sum / max / min of one value per thread over the CTA, in a fixed order; valid in thread 0.
__device__ __forceinline__ void block_reduce3(double& s, double& mx, double& mn) {
  __shared__ double sh[3][MV_BLOCK / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  __syncthreads();   // sh may still be read by thread 0 from an earlier call in the same kernel
  if (lane == 0) { sh[0][wid] = s; sh[1][wid] = mx; sh[2][wid] = mn; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < MV_BLOCK / 32; ++w) {
      s += sh[0][w];
      mx = fmax(mx, sh[1][w]);
      mn = fmin(mn, sh[2][w]);
    }
  }
}

// true in every thread of the CTA that took the last ticket, after all CTAs' partials are visible to it
__device__ __forceinline__ bool last_block(unsigned* ticket) {
  __shared__ bool s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (s_last) __threadfence();
  return s_last;
}

// Pass 1: volume of every cell (P1 simplex, mv_internal.cuh cell_geom), per-CTA sum / max / min into the scratch; the
// last CTA folds the partials in block order and leaves sum, max, min in the result slots and the ticket at zero.
template <int D>
__global__ void __launch_bounds__(MV_BLOCK)
k_static_volumes(const double* __restrict__ x, const int* __restrict__ cells, int64_t C, double* __restrict__ volume,
                 double* scratch) {
  double s = 0.0, mx = 0.0, mn = 1.7976931348623157e308;
  const int64_t base = (int64_t)blockIdx.x * MV_STATIC_CHUNK + threadIdx.x;
#pragma unroll 4
  for (int k = 0; k < MV_STATIC_CHUNK / MV_BLOCK; ++k) {
    const int64_t c = base + (int64_t)k * MV_BLOCK;
    if (c < C) {
      int cv[D + 1];
      double xc[D + 1][D];
      CellGeom<D> g;
      load_cell<D>(cells, c, cv);
      load_cell_coords<D>(x, cv, xc);
      cell_geom(xc, g);
      volume[c] = g.vol;
      s += g.vol;
      mx = fmax(mx, g.vol);
      mn = fmin(mn, g.vol);
    }
  }
  block_reduce3(s, mx, mn);
  if (threadIdx.x == 0) {
    scratch[blockIdx.x] = s;
    scratch[MV_STATIC_MAX_BLOCKS + blockIdx.x] = mx;
    scratch[2 * MV_STATIC_MAX_BLOCKS + blockIdx.x] = mn;
  }
  unsigned* ticket = (unsigned*)(scratch + MV_STATIC_TICKET);
  if (!last_block(ticket)) return;
  s = 0.0; mx = 0.0; mn = 1.7976931348623157e308;
  for (unsigned b = threadIdx.x; b < gridDim.x; b += blockDim.x) {
    s += __ldcg(&scratch[b]);
    mx = fmax(mx, __ldcg(&scratch[MV_STATIC_MAX_BLOCKS + b]));
    mn = fmin(mn, __ldcg(&scratch[2 * MV_STATIC_MAX_BLOCKS + b]));
  }
  block_reduce3(s, mx, mn);
  if (threadIdx.x == 0) {
    scratch[MV_STATIC_RES + 0] = s;
    scratch[MV_STATIC_RES + 1] = mx;
    scratch[MV_STATIC_RES + 2] = mn;
    *ticket = 0u;
  }
}

// Pass 2 (as the reference does it, mover.py:199-217: mean first, then the spread around it): sum over the cells of
// (volume - mean)^2, same chunks, same fold.
__global__ void __launch_bounds__(MV_BLOCK)
k_static_volume_var(const double* __restrict__ volume, int64_t C, double inv_count, double* scratch) {
  const double mean = scratch[MV_STATIC_RES + 0] * inv_count;
  double s = 0.0, mx = 0.0, mn = 0.0;
  const int64_t base = (int64_t)blockIdx.x * MV_STATIC_CHUNK + threadIdx.x;
#pragma unroll 4
  for (int k = 0; k < MV_STATIC_CHUNK / MV_BLOCK; ++k) {
    const int64_t c = base + (int64_t)k * MV_BLOCK;
    if (c < C) {
      const double d = volume[c] - mean;
      s += d * d;
    }
  }
  block_reduce3(s, mx, mn);
  if (threadIdx.x == 0) scratch[blockIdx.x] = s;
  unsigned* ticket = (unsigned*)(scratch + MV_STATIC_TICKET);
  if (!last_block(ticket)) return;
  s = 0.0;
  for (unsigned b = threadIdx.x; b < gridDim.x; b += blockDim.x) s += __ldcg(&scratch[b]);
  block_reduce3(s, mx, mn);
  if (threadIdx.x == 0) {
    scratch[MV_STATIC_RES + 3] = s;
    *ticket = 0u;
  }
}

}  // namespace mv

using namespace mv;

// See include/movement_b200.h.  Calls no cudaMalloc / cudaFree / cudaMallocAsync and nothing that does.
extern "C" int mv_cell_volumes_static(int dim, int64_t C, const double* coords, const int32_t* cells, double* volume,
                                      double* stats) {
  if (dim < 1 || dim > 3 || C < 1 || !coords || !cells || !volume || !stats) {
    set_error("mv_cell_volumes_static: dim in 1..3, at least one cell and non-null coords, cells, volume, stats are required");
    return MV_ERR_BADARG;
  }
  if (C > MV_STATIC_MAX_CELLS) {
    set_error("mv_cell_volumes_static: %lld cells exceed the fixed scratch (at most %lld cells = %d CTAs x %d); "
              "use mv_cell_volumes, which allocates what it needs",
              (long long)C, (long long)MV_STATIC_MAX_CELLS, MV_STATIC_MAX_BLOCKS, MV_STATIC_CHUNK);
    return MV_ERR_UNSUPPORTED;
  }
  // host arrays would have to be staged through a device buffer, i.e. allocated for: refused here
  if (!is_device_ptr(coords) || !is_device_ptr(cells) || !is_device_ptr(volume)) {
    set_error("mv_cell_volumes_static: coords, cells and volume must be device pointers (staging a host array would allocate)");
    return MV_ERR_BADARG;
  }
  double* scratch = nullptr;
  MV_CHECK(scratch_address(&scratch));
  // the only word that has to start at zero: the ticket.  Both kernels leave it at zero, but a call that died half way
  // (or a kernel aborted by a fault in the caller's arrays) would not have.
  MV_CUDA(cudaMemsetAsync(scratch + MV_STATIC_TICKET, 0, sizeof(double), g_stream));
  const int grid = (int)((C + MV_STATIC_CHUNK - 1) / MV_STATIC_CHUNK);
  MV_DISPATCH_DIM(dim, MV_L k_static_volumes<DIM><<<grid, MV_BLOCK, 0, g_stream>>>(coords, cells, C, volume, scratch));
  MV_KERNEL_CHECK();
  MV_L k_static_volume_var<<<grid, MV_BLOCK, 0, g_stream>>>(volume, C, 1.0 / (double)C, scratch);
  MV_KERNEL_CHECK();
  double h[4];
  MV_CUDA(cudaMemcpyAsync(h, scratch + MV_STATIC_RES, sizeof(h), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  const double mean = h[0] / (double)C;
  stats[0] = h[1];
  stats[1] = h[2];
  stats[2] = mean;
  stats[3] = sqrt(h[3] / (double)C) / mean;
  return MV_OK;
}
