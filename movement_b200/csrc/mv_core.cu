// Library plumbing + the per-cell kernels (tangling A9, cell volumes / diagnostics A4).
#include <stdarg.h>

#include "mv_internal.cuh"

namespace mv {

static thread_local char g_err[1024] = "";
cudaStream_t g_stream = 0;

int pool_init() {
  static int done_for = -1;
  int dev = 0;
  MV_CUDA(cudaGetDevice(&dev));
  if (done_for == dev) return MV_OK;
  cudaMemPool_t pool;
  MV_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
  // freed memory kept for reuse: half of the device by default.  A 256^3 mover holds ~34 GB; with the former 32 GB
  // threshold the pool was trimmed at every synchronisation and the next mover paid the driver for its pages again
  // (end-to-end 86.7 M -> 58.8 M vertex-iters/s, round 2).
  size_t free_b = 0, total_b = 0;
  MV_CUDA(cudaMemGetInfo(&free_b, &total_b));
  const char* e = getenv("MV_POOL_KEEP_GB");
  unsigned long long keep = e ? (unsigned long long)(atof(e) * (double)(1ull << 30)) : (unsigned long long)(total_b / 2);
  MV_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
  done_for = dev;
  return MV_OK;
}
int g_nsm = 0;
long long g_launches = 0;
Prof g_prof;
TrueRes g_trueres;

int prof_begin(int ncomp) {
  Prof& p = g_prof;
  if (!p.on || p.used >= Prof::NPAIR) return -1;
  if (!p.created) {
    for (int i = 0; i < 2 * Prof::NPAIR; ++i) cudaEventCreate(&p.ev[i]);
    p.created = true;
  }
  const int s = p.used++;
  p.cls[s] = ncomp & 15;
  cudaEventRecord(p.ev[2 * s], g_stream);
  return s;
}
void prof_end(int slot) {
  if (slot >= 0) cudaEventRecord(g_prof.ev[2 * slot + 1], g_stream);
}
void prof_drain() {
  Prof& p = g_prof;
  for (int s = 0; s < p.used; ++s) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, p.ev[2 * s], p.ev[2 * s + 1]) == cudaSuccess) {
      p.ms[p.cls[s]] += ms;
      p.count[p.cls[s]] += 1;
    }
  }
  p.used = 0;
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

bool is_device_ptr(const void* p) {
  cudaPointerAttributes at;
  cudaError_t e = cudaPointerGetAttributes(&at, p);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

int Staged::in(const void* src, size_t bytes) {
  if (!src) { dev = nullptr; return MV_OK; }
  if (is_device_ptr(src)) { dev = (void*)src; owned = false; return MV_OK; }
  MV_CUDA(cudaMallocAsync(&dev, bytes ? bytes : 1, g_stream));
  owned = true;
  MV_CUDA(cudaMemcpyAsync(dev, src, bytes, cudaMemcpyHostToDevice, g_stream));
  return MV_OK;
}
int Staged::out_begin(void* dst, size_t bytes) {
  if (!dst) { dev = nullptr; return MV_OK; }
  if (is_device_ptr(dst)) { dev = dst; owned = false; return MV_OK; }
  MV_CUDA(cudaMallocAsync(&dev, bytes ? bytes : 1, g_stream));
  owned = true;
  return MV_OK;
}
int Staged::out_end(void* dst, size_t bytes) {
  if (!dst || !owned) return MV_OK;
  MV_CUDA(cudaMemcpyAsync(dst, dev, bytes, cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  return MV_OK;
}

// ---------------------------------------------------------------------------------
// A9 tangling: signed Jacobian determinant, evaluated with explicit round-to-nearest
// multiplies/subtracts (never contracted into FMAs) in the order the oracle uses, so that
// the sign test `ratio <= 0` is bit-identical on CPU and GPU.
// replaces: movement/tangling.py:26-29 (detJ), :37 (ratio), :45-46 (count)
// ---------------------------------------------------------------------------------
template <int D>
__device__ __forceinline__ double signed_detJ(const double (&xc)[D + 1][D]);

template <>
__device__ __forceinline__ double signed_detJ<1>(const double (&xc)[2][1]) {
  return __dsub_rn(xc[1][0], xc[0][0]);   // interval: x1 - x0 (movement/tangling.py:26-29 in 1D)
}
template <>
__device__ __forceinline__ double signed_detJ<2>(const double (&xc)[3][2]) {
  const double e00 = __dsub_rn(xc[1][0], xc[0][0]), e01 = __dsub_rn(xc[1][1], xc[0][1]);
  const double e10 = __dsub_rn(xc[2][0], xc[0][0]), e11 = __dsub_rn(xc[2][1], xc[0][1]);
  return __dsub_rn(__dmul_rn(e00, e11), __dmul_rn(e01, e10));
}
template <>
__device__ __forceinline__ double signed_detJ<3>(const double (&xc)[4][3]) {
  double a[3], b[3], c[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    a[k] = __dsub_rn(xc[1][k], xc[0][k]);
    b[k] = __dsub_rn(xc[2][k], xc[0][k]);
    c[k] = __dsub_rn(xc[3][k], xc[0][k]);
  }
  const double m0 = __dsub_rn(__dmul_rn(b[1], c[2]), __dmul_rn(b[2], c[1]));
  const double m1 = __dsub_rn(__dmul_rn(b[0], c[2]), __dmul_rn(b[2], c[0]));
  const double m2 = __dsub_rn(__dmul_rn(b[0], c[1]), __dmul_rn(b[1], c[0]));
  return __dadd_rn(__dsub_rn(__dmul_rn(a[0], m0), __dmul_rn(a[1], m1)), __dmul_rn(a[2], m2));
}

template <int D, bool INIT>
__global__ void __launch_bounds__(MV_BLOCK)
k_tangle(const double* __restrict__ x, const int* __restrict__ cells, int dg, int64_t C, int64_t Ccount,
         double* __restrict__ detJ0, double* __restrict__ ratio, unsigned long long* __restrict__ count) {
  unsigned local = 0;
  for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < C; c += (int64_t)gridDim.x * blockDim.x) {
    int cv[D + 1];
    double xc[D + 1][D];
    load_cell<D>(cells, c, cv);
    load_cell_coords<D>(x, cv, xc, dg, c);
    const double det = signed_detJ<D>(xc);
    if (INIT) {
      detJ0[c] = det;
    } else {
      const double r = __ddiv_rn(det, __ldg(&detJ0[c]));
      if (ratio) ratio[c] = r;
      local += (c < Ccount && r <= 0.0) ? 1u : 0u;
    }
  }
  if (!INIT) {
    // integer atomics: order-independent, hence deterministic
    local = __reduce_add_sync(0xffffffffu, local);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(count, (unsigned long long)local);
  }
}

// ---------------------------------------------------------------------------------
// A4 / mover diagnostics: per-cell volume (optionally weighted by the cell-mean monitor and
// divided by the original volume), then max / min / mean / coefficient of variation.
// replaces: movement/monge_ampere.py:507-510, laplacian_smoothing.py:119, spring.py:274,
//           movement/mover.py:199-217
// Two passes, as the reference does: pass 1 volumes + sum/max/min, pass 2 sum (v-mean)^2.
// ---------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(MV_BLOCK)
k_volumes(const double* __restrict__ x, const int* __restrict__ cells, int dg, int64_t C, const double* __restrict__ mon,
          const double* __restrict__ vol0, double* __restrict__ volume, double* __restrict__ partials,
          unsigned* __restrict__ counter, double* __restrict__ out /* [0]=sum [1]=max [2]=min */) {
  double s[1] = {0.0};
  double vmax = 0.0, vmin = 1.7976931348623157e308;
  for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < C; c += (int64_t)gridDim.x * blockDim.x) {
    int cv[D + 1];
    double xc[D + 1][D];
    CellGeom<D> g;
    load_cell<D>(cells, c, cv);
    load_cell_coords<D>(x, cv, xc, dg, c);
    cell_geom(xc, g);
    double v = g.vol;
    if (mon) {
      double mm = 0.0;
#pragma unroll
      for (int j = 0; j <= D; ++j) mm += __ldg(&mon[cv[j]]);
      v = v * (mm / (D + 1));
    }
    if (vol0) v = v / __ldg(&vol0[c]);
    volume[c] = v;
    s[0] += v;
    vmax = fmax(vmax, v);
    vmin = fmin(vmin, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    vmax = fmax(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    vmin = fmin(vmin, __shfl_xor_sync(0xffffffffu, vmin, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomic_max_pos(&out[1], fmax(vmax, 0.0));
    atomic_min_pos(&out[2], fmax(vmin, 0.0));
  }
  grid_reduce_sum<1>(s, partials, counter, [&](double(&t)[1]) { out[0] = t[0]; });
}

__global__ void __launch_bounds__(MV_BLOCK)
k_volume_var(const double* __restrict__ volume, int64_t C, double inv_count, double* __restrict__ partials,
             unsigned* __restrict__ counter, double* __restrict__ out /* in: [0]=global sum; out: [3]=sum sq dev */) {
  const double mean = out[0] * inv_count;
  double s[1] = {0.0};
  for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < C; c += (int64_t)gridDim.x * blockDim.x) {
    const double d = volume[c] - mean;
    s[0] += d * d;
  }
  grid_reduce_sum<1>(s, partials, counter, [&](double(&t)[1]) { out[3] = t[0]; });
}

__global__ void k_init_volstats(double* out) {
  out[0] = 0.0;
  out[1] = 0.0;
  out[2] = 1.7976931348623157e308;
  out[3] = 0.0;
}

// device part shared by mv_cell_volumes and mv_ma_diagnostics: leaves
// dout[0..3] = sum, max, min, sum of squared deviations (global over ranks)
int volume_stats_device(mv_mesh* m, const double* x, const double* mon, const double* vol0, double* volume,
                        double* dout, int64_t global_cells) {
  const int64_t C = m->Cown;
  MV_L k_init_volstats<<<1, 1, 0, g_stream>>>(dout);
  const int grid = grid_for(C);
  MV_DISPATCH_DIM(m->dim, MV_L k_volumes<DIM><<<grid, MV_BLOCK, 0, g_stream>>>(x, m->cells, m->dg, C, mon, vol0, volume, m->red.partials,
                                                                          m->red.counter, dout));
  MV_KERNEL_CHECK();
  if (comm_active()) MV_CHECK(allreduce_volstats_dev(dout));
  MV_L k_volume_var<<<grid, MV_BLOCK, 0, g_stream>>>(volume, C, 1.0 / (double)global_cells, m->red.partials, m->red.counter, dout);
  MV_KERNEL_CHECK();
  return MV_OK;
}

}  // namespace mv

using namespace mv;

extern "C" int mv_version(void) { return 100; }
extern "C" const char* mv_last_error(void) { return g_err; }
extern "C" int mv_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}
extern "C" int mv_set_device(int device) {
  MV_CUDA(cudaSetDevice(device));
  MV_CUDA(cudaDeviceGetAttribute(&g_nsm, cudaDevAttrMultiProcessorCount, device));
  return MV_OK;
}
extern "C" int mv_set_stream(void* s) {
  g_stream = (cudaStream_t)s;
  return MV_OK;
}
extern "C" int mv_prof_enable(int on) {
  if (g_prof.used) { cudaStreamSynchronize(g_stream); prof_drain(); }
  g_prof.on = on != 0;
  return MV_OK;
}
extern "C" int mv_prof_reset(void) {
  if (g_prof.used) { cudaStreamSynchronize(g_stream); prof_drain(); }
  for (int i = 0; i < 16; ++i) { g_prof.ms[i] = 0; g_prof.count[i] = 0; g_prof.pcg_iters[i] = 0; }
  g_launches = 0;
  return MV_OK;
}
extern "C" int mv_prof_get(double* spmv_ms16, int64_t* spmv_count16, int64_t* pcg_iters16, int64_t* launches) {
  if (g_prof.used) { cudaStreamSynchronize(g_stream); prof_drain(); }
  for (int i = 0; i < 16; ++i) {
    if (spmv_ms16) spmv_ms16[i] = g_prof.ms[i];
    if (spmv_count16) spmv_count16[i] = g_prof.count[i];
    if (pcg_iters16) pcg_iters16[i] = g_prof.pcg_iters[i];
  }
  if (launches) *launches = g_launches;
  return MV_OK;
}
extern "C" int mv_true_residual_enable(int on) {
  g_trueres.on = on != 0;
  for (int i = 0; i < 16; ++i) { g_trueres.max_rel[i] = 0; g_trueres.last_rel[i] = 0; g_trueres.count[i] = 0; }
  return MV_OK;
}
extern "C" int mv_true_residual_get(double* max_rel16, double* last_rel16, int64_t* count16) {
  for (int i = 0; i < 16; ++i) {
    if (max_rel16) max_rel16[i] = g_trueres.max_rel[i];
    if (last_rel16) last_rel16[i] = g_trueres.last_rel[i];
    if (count16) count16[i] = g_trueres.count[i];
  }
  return MV_OK;
}
extern "C" int mv_sync(void) {
  MV_CUDA(cudaStreamSynchronize(g_stream));
  return MV_OK;
}

extern "C" int mv_tangle_init(mv_mesh* m, const double* coords, double* detJ0) {
  if (!m || !coords || !detJ0) return MV_ERR_BADARG;
  Staged sx, sd;
  MV_CHECK(sx.in(coords, m->Vc * m->dim * sizeof(double)));
  MV_CHECK(sd.out_begin(detJ0, m->C * sizeof(double)));
  const int grid = grid_for(m->C);
  MV_DISPATCH_DIM(m->dim, MV_L k_tangle<DIM, true><<<grid, MV_BLOCK, 0, g_stream>>>(sx.as<double>(), m->cells, m->dg, m->C, m->C, sd.as<double>(),
                                                                              nullptr, nullptr));
  MV_KERNEL_CHECK();
  return sd.out_end(detJ0, m->C * sizeof(double));
}

extern "C" int mv_tangle_check(mv_mesh* m, const double* coords, const double* detJ0, double* ratio, int64_t* count) {
  if (!m || !coords || !detJ0 || !count) return MV_ERR_BADARG;
  Staged sx, sd, sr;
  MV_CHECK(sx.in(coords, m->Vc * m->dim * sizeof(double)));
  MV_CHECK(sd.in(detJ0, m->C * sizeof(double)));
  MV_CHECK(sr.out_begin(ratio, m->C * sizeof(double)));
  unsigned long long* d_count = (unsigned long long*)(m->red.counter + 2);  // 8-byte aligned slot
  MV_CUDA(cudaMemsetAsync(d_count, 0, sizeof(unsigned long long), g_stream));
  // the ratio field is written for all local cells, only owned cells are counted
  const int64_t C = ratio ? m->C : m->Cown;
  const int grid = grid_for(C);
  MV_DISPATCH_DIM(m->dim, MV_L k_tangle<DIM, false><<<grid, MV_BLOCK, 0, g_stream>>>(sx.as<double>(), m->cells, m->dg, C, m->Cown, sd.as<double>(),
                                                                               sr.as<double>(), d_count));
  MV_KERNEL_CHECK();
  unsigned long long h = 0;
  MV_CUDA(cudaMemcpyAsync(&h, d_count, sizeof(h), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  *count = (int64_t)h;
  return sr.out_end(ratio, m->C * sizeof(double));
}

extern "C" int mv_cell_volumes(mv_mesh* m, const double* coords, double* volume, double* stats) {
  if (!m || !coords || !volume) return MV_ERR_BADARG;
  Staged sx, sv;
  MV_CHECK(sx.in(coords, m->Vc * m->dim * sizeof(double)));
  MV_CHECK(sv.out_begin(volume, m->C * sizeof(double)));
  double* dout = nullptr;
  MV_CUDA(cudaMallocAsync((void**)&dout, 4 * sizeof(double), g_stream));
  int64_t gcells = m->Cown;
  if (comm_active()) {
    double hc = (double)m->Cown;
    MV_CUDA(cudaMemcpyAsync(dout, &hc, sizeof(double), cudaMemcpyHostToDevice, g_stream));
    MV_CHECK(allreduce_sum_dev(dout, 1));
    MV_CUDA(cudaMemcpyAsync(&hc, dout, sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
    gcells = (int64_t)(hc + 0.5);
  }
  MV_CHECK(volume_stats_device(m, sx.as<double>(), nullptr, nullptr, sv.as<double>(), dout, gcells));
  double h[4];
  MV_CUDA(cudaMemcpyAsync(h, dout, sizeof(h), cudaMemcpyDeviceToHost, g_stream));
  MV_CUDA(cudaFreeAsync(dout, g_stream));
  MV_CUDA(cudaStreamSynchronize(g_stream));
  if (comm_active()) {   // sum of squared deviations is per rank
    double* d2 = nullptr;
    MV_CUDA(cudaMallocAsync((void**)&d2, sizeof(double), g_stream));
    MV_CUDA(cudaMemcpyAsync(d2, &h[3], sizeof(double), cudaMemcpyHostToDevice, g_stream));
    MV_CHECK(allreduce_sum_dev(d2, 1));
    MV_CUDA(cudaMemcpyAsync(&h[3], d2, sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    MV_CUDA(cudaFreeAsync(d2, g_stream));
    MV_CUDA(cudaStreamSynchronize(g_stream));
  }
  if (stats) {
    const double mean = h[0] / (double)gcells;
    stats[0] = h[1];
    stats[1] = h[2];
    stats[2] = mean;
    stats[3] = sqrt(h[3] / (double)gcells) / mean;
  }
  return sv.out_end(volume, m->C * sizeof(double));
}
