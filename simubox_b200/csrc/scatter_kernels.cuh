// scatter_kernels.cuh -- fused |F|^2 / |q|-shell kernels (sm_100a).
//
// Replaces the hstack + argsort + merge loop of Scatter.sacttering_peak
// (SimuBox/calculator/scatter.py:86-118).  The spectrum is the cuFFT D2Z half spectrum
// (last axis stored 0..N/2); every stored element is binned for itself and, when it has one,
// for its Hermitian mirror, each with the k^2 the reference pairs with that full-spectrum
// position (quirk Q1).  All k^2 arithmetic is two table look-ups and
// two adds, performed with __dadd_rn so that host table and device agree bit for bit.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "peaks.cuh"
#include "scatter_types.h"

namespace sbx {

// largest s with rep[s] <= v (v >= rep[0] = smallest k^2 of the grid)
__device__ __forceinline__ int shell_search(const double* rep, int n, double v) {
  int lo = 0, hi = n;  // invariant: rep[lo] <= v < rep[hi]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (rep[mid] <= v) lo = mid; else hi = mid;
  }
  return lo;
}

// shell id for k^2 value v at full-spectrum index f, or -1 if the reference would not bin it
__device__ __forceinline__ int shell_of(const ShellDev& sh, const double* rep_s, double v, long long f) {
  const int s = shell_search(rep_s, sh.n_shells, v);
  if (fabs(__dsub_rn(rep_s[s], v)) > sh.threshold) return -1;  // beyond the table (scatter.py:108)
  if (sh.truncated && s == sh.n_shells - 1 && f != sh.designated) return -1;  // scatter.py:106-107
  return s;
}

__device__ __forceinline__ double intensity(const double2 z, double factor) {
  // factor * |F^2| (scatter.py:92); |F^2| = re^2 + im^2 up to 1 ulp
  return factor * (z.x * z.x + z.y * z.y);
}

// ------------------------------------------------------------------------------------------
// Dense path: full scan of the stored spectrum, shared-memory shell histogram per block,
// one global RED per touched bin per block.  grid = (blocks_per_field, batch).
// The last block to finish a field optionally runs the peak search on the finished row.
// 3D grids walk the spectrum ROW-wise: a warp takes 32 rows (a, b) at a time, every lane tests one row's pair of
// plane k^2 values against the largest listed k^2 (one coalesced 16-byte load per row), and only rows that can hold
// a listed mode are read at all -- by the whole warp, 32 consecutive c per step.  With cutoff = 30 that is ~1 % of a
// 128^3 spectrum, so the scan costs one test per row instead of a division and a table look-up per element.
// ------------------------------------------------------------------------------------------
template <int NDIM>
__global__ void __launch_bounds__(256)
shell_histogram_kernel(const double2* __restrict__ spec, GridDev g, ShellDev sh,
                       double* __restrict__ shell_I, unsigned int* __restrict__ tickets,
                       int do_peaks, const double* __restrict__ q, PeakParams pp,
                       double* __restrict__ plot_y, int32_t* __restrict__ peak_idx,
                       int32_t* __restrict__ n_peaks) {
  extern __shared__ double smem[];
  double* rep_s = smem;                 // n_shells
  double* bins = smem + sh.n_shells;    // n_shells
  const int field = blockIdx.y;
  const int tid = threadIdx.x;
  for (int s = tid; s < sh.n_shells; s += blockDim.x) {
    rep_s[s] = sh.rep[s];
    bins[s] = 0.0;
  }
  __syncthreads();

  const double2* fs = spec + (long long)field * g.nspec;
  const long long per_block = (g.nspec + gridDim.x - 1) / gridDim.x;
  const long long e0 = (long long)blockIdx.x * per_block;
  const long long e1 = min(e0 + per_block, g.nspec);
  const int halfLast = g.H - 1;  // N_last / 2
  bool any = false;

  if (NDIM == 3) {
    const int lane = tid & 31;
    const long long nrows = g.nspec / g.Hs;  // local rows (slab layout: N0 * n1loc), Hs stored columns each
    const long long wglobal = (long long)blockIdx.x * (blockDim.x >> 5) + (tid >> 5), wtotal = (long long)gridDim.x * (blockDim.x >> 5);
    for (long long r0 = wglobal * 32; r0 < nrows; r0 += wtotal * 32) {
      const long long lrow = r0 + lane;
      long long row = lrow;
      double2 p2 = make_double2(INFINITY, INFINITY);
      if (lrow < nrows) {
        if (g.n1loc != g.N1) {  // slab layout [x][j_local][c] (slab.cu): global row = x*N1 + joff + j_local
          const long long x = lrow / g.n1loc;
          row = x * g.N1 + g.joff + (lrow - x * g.n1loc);
        }
        p2 = g.plane2[row];
      }
      unsigned todo = __ballot_sync(0xffffffffu, fmin(p2.x, p2.y) <= sh.vcut);  // other rows are never loaded
      while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const long long wl = r0 + src, wrow = __shfl_sync(0xffffffffu, row, src);
        const double px = __shfl_sync(0xffffffffu, p2.x, src), py = __shfl_sync(0xffffffffu, p2.y, src);
        for (int c = lane; c < g.Hs; c += 32) {
          const double kc = g.kvlast[c];
          const double v1 = __dadd_rn(px, kc), v2 = __dadd_rn(py, kc);
          const bool has_mirror = (c != 0) && (c != halfLast);
          const bool h1 = v1 <= sh.vcut, h2 = has_mirror && v2 <= sh.vcut;
          if (!(h1 || h2)) continue;
          const double inten = intensity(fs[wl * g.Hs + c], g.factor);
          if (h1) {
            const int s = shell_of(sh, rep_s, v1, wrow * g.N2 + c);
            if (s >= 0) { atomicAdd(&bins[s], inten); any = true; }
          }
          if (h2) {
            long long f2 = -1;
            if (sh.truncated) {  // only needed for the designated-member test
              const int a = (int)(wrow / g.N1), b = (int)(wrow - (long long)a * g.N1);
              f2 = ((long long)(a ? g.N0 - a : 0) * g.N1 + (b ? g.N1 - b : 0)) * g.N2 + (g.N2 - c);
            }
            const int s = shell_of(sh, rep_s, v2, f2);
            if (s >= 0) { atomicAdd(&bins[s], inten); any = true; }
          }
        }
      }
    }
  } else
  for (long long e = e0 + tid; e < e1; e += blockDim.x) {
    double v1, v2;
    long long f1, f2;
    bool has_mirror;
    if (NDIM == 2) {
      const int a = (int)(e / g.H), b = (int)(e - (long long)a * g.H);
      f1 = (long long)a * g.N1 + b;
      has_mirror = (b != 0) && (b != halfLast);
      f2 = (long long)(a ? g.N0 - a : 0) * g.N1 + (g.N1 - b);
      v1 = g.plane[f1];
      v2 = has_mirror ? g.plane[f2] : 0.0;
    } else {
      f1 = e;
      has_mirror = (e != 0) && (e != halfLast);
      f2 = g.N0 - e;
      v1 = g.plane[f1];
      v2 = has_mirror ? g.plane[f2] : 0.0;
    }
    const bool h1 = v1 <= sh.vcut;
    const bool h2 = has_mirror && v2 <= sh.vcut;
    if (!(h1 || h2)) continue;
    const double inten = intensity(fs[e], g.factor);
    if (h1) {
      const int s = shell_of(sh, rep_s, v1, f1);
      if (s >= 0) { atomicAdd(&bins[s], inten); any = true; }
    }
    if (h2) {
      const int s = shell_of(sh, rep_s, v2, f2);
      if (s >= 0) { atomicAdd(&bins[s], inten); any = true; }
    }
  }
  const int block_any = __syncthreads_or(any ? 1 : 0);
  if (block_any) {
    double* out = shell_I + (long long)field * sh.n_shells;
    for (int s = tid; s < sh.n_shells; s += blockDim.x) {
      const double v = bins[s];
      if (v != 0.0) atomicAdd(&out[s], v);
    }
  }
  if (!tickets) return;

  // last-arriving block of this field: peak search on the finished shell row
  __shared__ int s_last;
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const unsigned t = atomicAdd(&tickets[field], 1u);
    s_last = (t == gridDim.x - 1);
    if (s_last) tickets[field] = 0;  // re-arm for the next launch
  }
  __syncthreads();
  if (!s_last || !do_peaks) return;
  __threadfence();
  double* row = smem;  // reuse: n_shells doubles for I, then the peaks scratch
  for (int s = tid; s < sh.n_shells; s += blockDim.x)
    row[s] = __ldcg(shell_I + (long long)field * sh.n_shells + s);
  __syncthreads();
  peaks_block(q, row, sh.n_shells, pp, plot_y ? plot_y + (long long)field * pp.step : nullptr,
              peak_idx ? peak_idx + (long long)field * pp.max_peaks : nullptr,
              n_peaks ? n_peaks + field : nullptr, smem + sh.n_shells);
}

// ------------------------------------------------------------------------------------------
// Gather path: the shell members are known on the host (field independent), so only those
// spectrum elements are read.  One block per field, one warp per shell at a time; every lane
// sums its members in list order and the warp combines with a fixed shuffle tree, so the
// result is bitwise reproducible.  The same block then runs the peak search.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
shell_gather_kernel(const double2* __restrict__ spec, long long nspec, double factor,
                    const long long* __restrict__ elem, const int* __restrict__ shell_off,
                    int n_shells, double* __restrict__ shell_I, int do_peaks,
                    const double* __restrict__ q, PeakParams pp, double* __restrict__ plot_y,
                    int32_t* __restrict__ peak_idx, int32_t* __restrict__ n_peaks) {
  extern __shared__ double smem[];
  const int field = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const double2* fs = spec + (long long)field * nspec;
  for (int s = warp; s < n_shells; s += nwarp) {
    double acc = 0.0;
    for (int m = shell_off[s] + lane; m < shell_off[s + 1]; m += 32) acc += intensity(__ldg(&fs[elem[m]]), factor);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) {
      smem[s] = acc;
      shell_I[(long long)field * n_shells + s] = acc;
    }
  }
  if (!do_peaks) return;
  __syncthreads();
  peaks_block(q, smem, n_shells, pp, plot_y ? plot_y + (long long)field * pp.step : nullptr,
              peak_idx ? peak_idx + (long long)field * pp.max_peaks : nullptr,
              n_peaks ? n_peaks + field : nullptr, smem + n_shells);
}

// ------------------------------------------------------------------------------------------
// Pruned-DFT finish: stage 3 (x transform) of pruned_dft.cuh for the listed modes only, then
// |F|^2/N, shell sums in fixed order (bitwise reproducible) and the peak search.  One block
// per field; one thread per mode, so no cross-lane reduction is needed.
// planes: [field][x][16][16] doubles written by pruned_planes_kernel.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
pruned_finish_kernel(const double* __restrict__ planes, int N0, double factor, const int4* __restrict__ umodes,
                     int n_umodes, const double2* __restrict__ tw0, const long long* __restrict__ memb_u,
                     const int* __restrict__ shell_off, int n_shells, double* __restrict__ shell_I, int do_peaks,
                     const double* __restrict__ q, PeakParams pp, double* __restrict__ plot_y,
                     int32_t* __restrict__ peak_idx, int32_t* __restrict__ n_peaks) {
  extern __shared__ double smem[];
  double* row = smem;                     // n_shells
  double* Iu = smem + n_shells;           // n_umodes   (region reused by the peak search afterwards)
  double2* tw = reinterpret_cast<double2*>(Iu + n_umodes + ((n_shells + n_umodes) & 1));  // N0, 16-byte aligned
  const int field = blockIdx.x, tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
  for (int j = tid; j < N0; j += blockDim.x) tw[j] = tw0[j];
  __syncthreads();
  const double* fp = planes + (long long)field * N0 * 256;
  for (int u = tid; u < n_umodes; u += blockDim.x) {
    const int4 md = umodes[u];  // x = a index (0..N0-1), y = signed b, z = c
    const int ab = md.y < 0 ? -md.y : md.y;
    const int o_rc = md.z * 16 + ab, o_rs = o_rc + 8, o_sc = (8 + md.z) * 16 + ab, o_ss = o_sc + 8;
    double fr = 0.0, fi = 0.0;
    int j = 0;
#pragma unroll 4
    for (int x = 0; x < N0; ++x) {
      const double* hp = fp + (long long)x * 256;
      const double rc = hp[o_rc], rs = hp[o_rs], sc = hp[o_sc], ss = hp[o_ss];
      const double hr = md.y >= 0 ? rc - ss : rc + ss;
      const double hi = md.y >= 0 ? -(rs + sc) : rs - sc;
      const double2 w = tw[j];  // (cos, sin)(2 pi a x / N0)
      fr += hr * w.x + hi * w.y;
      fi += hi * w.x - hr * w.y;
      j += md.x;
      if (j >= N0) j -= N0;
    }
    Iu[u] = factor * (fr * fr + fi * fi);
  }
  __syncthreads();
  for (int s = warp; s < n_shells; s += nwarp) {
    double acc = 0.0;
    for (int m = shell_off[s] + lane; m < shell_off[s + 1]; m += 32) acc += Iu[memb_u[m]];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) {
      row[s] = acc;
      shell_I[(long long)field * n_shells + s] = acc;
    }
  }
  if (!do_peaks) return;
  __syncthreads();
  peaks_block(q, row, n_shells, pp, plot_y ? plot_y + (long long)field * pp.step : nullptr,
              peak_idx ? peak_idx + (long long)field * pp.max_peaks : nullptr,
              n_peaks ? n_peaks + field : nullptr, smem + n_shells);
}

// stand-alone peak search (sbx_peaks_execute): one block per field
__global__ void __launch_bounds__(256)
peaks_kernel(const double* __restrict__ q, const double* __restrict__ I, int n, PeakParams pp,
             double* __restrict__ plot_y, int32_t* __restrict__ peak_idx, int32_t* __restrict__ n_peaks) {
  extern __shared__ double smem[];
  const int field = blockIdx.x;
  peaks_block(q, I + (long long)field * n, n, pp, plot_y ? plot_y + (long long)field * pp.step : nullptr,
              peak_idx ? peak_idx + (long long)field * pp.max_peaks : nullptr,
              n_peaks ? n_peaks + field : nullptr, smem);
}

}  // namespace sbx
