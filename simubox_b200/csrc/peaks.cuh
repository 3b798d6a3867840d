// peaks.cuh -- block-level Gaussian broadening + strict-local-maxima peak search.
//
// Device restatement of the numeric part of Scatter.show_peak
// (SimuBox/calculator/scatter.py:137-149), peak.gaussian_expansion (calculator/peak.py:11-17)
// and scipy.signal.find_peaks(y, height=h) (plateau midpoint, edges excluded).
// One thread block handles one field.  All fp64 arithmetic uses explicit round-to-nearest
// intrinsics so that nvcc cannot contract a*b+c into an FMA the numpy path does not have.
#pragma once
#include <cstdint>

namespace sbx {

struct PeakParams {
  double width;
  double q_cutoff;
  double min_height;
  int step;
  int max_peaks;
};

// shared-memory doubles needed by peaks_block for n rows and `step` samples
__host__ __device__ inline size_t peaks_smem_doubles(int n, int step) { return (size_t)step + 2 * (size_t)n; }

// q: n shell positions; I: n intensities (row 0 treated as 0); sm: peaks_smem_doubles() doubles.
// plot_y (step doubles) / peak_idx (max_peaks) / n_peaks (1) may each be NULL.
// n_peaks = -1 flags "no row with q <= q_cutoff" (the reference raises IndexError there).
__device__ inline void peaks_block(const double* __restrict__ q, const double* __restrict__ I, int n,
                                   const PeakParams& pp, double* __restrict__ plot_y,
                                   int32_t* __restrict__ peak_idx, int32_t* __restrict__ n_peaks,
                                   double* sm) {
  double* y = sm;
  double* ctr = sm + pp.step;
  double* amp = ctr + n;
  __shared__ int s_rows;
  __shared__ int s_warp[32];
  __shared__ int s_total;
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int lane = tid & 31, warp = tid >> 5, nwarp = (nthr + 31) >> 5;

  // stage q and I in shared memory with parallel loads (a serial walk over global memory costs one L2
  // round trip per row), then compact in order:  q_Intensity[q_Intensity[:,0] <= cutoff]  (scatter.py:141)
  for (int r = tid; r < n; r += nthr) {
    amp[r] = q[r];
    ctr[r] = r == 0 ? 0.0 : I[r];  // scatter.py:138
  }
  __syncthreads();
  if (tid == 0) {
    int m = 0;
    for (int r = 0; r < n; ++r) {  // m <= r: slot m is rewritten only after row r has been read
      const double qr = amp[r], ir = ctr[r];
      if (qr <= pp.q_cutoff) {
        ctr[m] = qr;
        amp[m] = ir;
        ++m;
      }
    }
    s_rows = m;
    s_total = 0;
  }
  __syncthreads();
  const int rows = s_rows;
  if (rows == 0) {
    if (tid == 0 && n_peaks) *n_peaks = -1;
    return;
  }
  // x = np.linspace(q[0], q[-1], step): i*dx + start, last sample = stop exactly
  const double start = ctr[0], stop = ctr[rows - 1];
  const int div = pp.step - 1;
  const double delta = __dsub_rn(stop, start);
  const double dx = div > 0 ? __ddiv_rn(delta, (double)div) : 0.0;
  // width = 2^k (the default 0.5 is): 1 / width is exact, so RN(d * (1 / width)) == RN(d / width) bit for bit and
  // the ~20-instruction division leaves the step x rows inner loop
  const long long wbits = __double_as_longlong(pp.width);
  const int wexp = (int)((wbits >> 52) & 0x7ff);
  const bool w_pow2 = (wbits & 0x000fffffffffffffLL) == 0 && wexp > 0x3ff - 1000 && wexp < 0x3ff + 1000;
  const double w_inv = w_pow2 ? __ddiv_rn(1.0, pp.width) : 0.0;
  for (int i = tid; i < pp.step; i += nthr) {
    double x;
    if (div > 0 && dx != 0.0) x = __dadd_rn(__dmul_rn((double)i, dx), start);
    else if (div > 0) x = __dadd_rn(__dmul_rn(__ddiv_rn((double)i, (double)div), delta), start);
    else x = __dadd_rn(__dmul_rn((double)i, delta), start);
    if (div > 0 && i == div) x = stop;
    double acc = 0.0;
    for (int r = 0; r < rows; ++r) {
      const double d = __dsub_rn(x, ctr[r]);
      const double t = w_pow2 ? __dmul_rn(d, w_inv) : __ddiv_rn(d, pp.width);
      const double g = __dmul_rn(amp[r], exp(-__dmul_rn(t, t)));
      acc = __dadd_rn(acc, g);
    }
    y[i] = acc;
    if (plot_y) plot_y[i] = acc;
  }
  __syncthreads();

  // strict local maxima with plateau midpoint; ordered compaction
  const int i_max = pp.step - 1;
  for (int base = 0; base < pp.step; base += nthr) {
    const int i = base + tid;
    int pk = -1;
    if (i >= 1 && i < i_max && y[i - 1] < y[i]) {
      int j = i + 1;
      while (j < i_max && y[j] == y[i]) ++j;
      if (y[j] < y[i]) {
        const int mid = (i + j - 1) >> 1;
        if (y[mid] >= pp.min_height) pk = mid;
      }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, pk >= 0);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    int before = s_total;
    for (int w = 0; w < warp; ++w) before += s_warp[w];
    if (pk >= 0) {
      const int slot = before + __popc(bal & ((1u << lane) - 1u));
      if (peak_idx && slot < pp.max_peaks) peak_idx[slot] = pk;
    }
    __syncthreads();
    if (tid == 0) {
      int add = 0;
      for (int w = 0; w < nwarp; ++w) add += s_warp[w];
      s_total += add;
    }
    __syncthreads();
  }
  if (tid == 0 && n_peaks) *n_peaks = s_total;
}

}  // namespace sbx
