// pruned_fused.cuh -- single-kernel pruned DFT: folded z stage, y stage, x accumulation, per-field
// finish (|F|^2/N, shell sums, peaks) by the last CTA that touches the field (sm_100a).
//
// Same mathematics as pruned_dft.cuh (the low-|q| modes the shell table of
// SimuBox/calculator/scatter.py:99-115 keeps, out of the full FFT of :85), with two changes:
//
//  * radix-2 fold along z: with e[z] = rho[z] + rho[z + N2/2], o[z] = rho[z] - rho[z + N2/2]
//    (z < N2/2), the even frequencies c = 0,2,4,6 are sums over e and the odd ones over o, because
//    exp(-2 pi i c (z + N2/2) / N2) = (-1)^c exp(-2 pi i c z / N2).  Two 8-column DMMA tiles over
//    K = N2/2 replace one 16-column product over K = N2: half the tensor instructions per byte
//    loaded, and both halves of a row are read with forward, 16-byte aligned loads.
//  * the x transform is accumulated plane by plane in registers (one thread per listed mode), so the
//    per-plane 16x16 results never go to HBM.  A work unit = `xc` consecutive planes of one field;
//    units are handed out by an atomic counter; a unit ends by writing its partial sums; the CTA that
//    completes a field (ticket) adds the partials in fixed order -- the result does not depend on the
//    schedule -- and runs the shell sums and the peak search.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "peaks.cuh"
#include "pruned_dft.cuh"

namespace sbx {

constexpr int kFusedModesPerThread = 2;  // listed modes <= 2 * 256

struct FusedArgs {
  const double* fields;
  int n_fields;
  int N0, N1, N2;
  int ks1;            // stage-2 k-steps: N1 / 4
  int gstride;        // N1 + 4
  int ds;             // fold steps: N2 / 16
  const double* twf;  // [ds][4][32]  B fragments: {even freq, odd freq} x {even t, odd t}
  const double* tw1;  // [2][ks1][32] B fragments of stage 2
  const double2* tw0; // N0: (cos, sin)(2 pi j / N0)
  const int4* umodes; // x = a index (0..N0-1), y = signed b, z = c
  int n_umodes;
  const long long* memb_u;
  const int* shell_off;
  int n_shells;
  double factor;
  int xc, chunks;     // planes per unit, plane chunks per field
  int parts;          // units per field: chunks (x yblocks in the blocked kernel)
  // blocked ("giant grid" / slab) kernel only:
  int planes;         // x-planes of each field held in `fields` (N0, or this rank's slab)
  int x_base;         // global x index of the first of them
  int yb_rows, yblocks;  // a unit covers yb_rows rows of its planes; N1 = yb_rows * yblocks
  int finish;         // 0: leave the partial sums (they are reduced across ranks first)
  double2* partial;   // [n_fields][chunks][n_umodes]
  unsigned* ctrl;     // [0] unit counter, [1 + field] tickets (zeroed before the launch)
  int work_doubles;   // size of the per-CTA scratch region
  int ring_slots;     // TMA-ring variant: slots of 8 rows (power of two)
  int slots_shift, groups_shift, xc_shift;  // log2 of ring_slots, N1 / 8, xc
  double* shell_I;
  int do_peaks;
  const double* q;
  PeakParams pp;
  double* plot_y;
  int32_t* peak_idx;
  int32_t* n_peaks;
};

__host__ __device__ inline size_t fused_work_doubles(int N1, int n_shells, int n_umodes, int step, bool peaks) {
  const size_t a = (size_t)16 * (N1 + 4) + 512;
  const size_t b = (size_t)n_shells + n_umodes + 2 + (peaks ? peaks_smem_doubles(n_shells, step) : 0);
  return a > b ? a : b;
}
__host__ __device__ inline size_t fused_smem_bytes(int N0, int N1, int N2, size_t work_doubles) {
  return ((size_t)(N2 / 16) * 128 + 2 * (size_t)(N1 / 4) * 32 + 2 * (size_t)N0 + work_doubles) * sizeof(double);
}

// shell sums + peak search of one finished field; `work` = per-CTA scratch (fused_work_doubles)
__device__ inline void fused_finish(const FusedArgs& A, int field, double* work) {
  double* row = work;              // n_shells
  double* Iu = work + A.n_shells;  // n_umodes (region reused by the peak search afterwards)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
  const double2* part = A.partial + (long long)field * A.parts * A.n_umodes;
  if (A.parts <= 64) {
    for (int u = tid; u < A.n_umodes; u += blockDim.x) {
      double sr = 0.0, si = 0.0;
      for (int ch = 0; ch < A.parts; ++ch) {  // fixed order: independent of which CTA produced which unit
        const double2 p = __ldcg(part + (long long)ch * A.n_umodes + u);
        sr += p.x;
        si += p.y;
      }
      Iu[u] = A.factor * (sr * sr + si * si);
    }
  } else {
    // many units per field (giant grids): a warp per mode, lane l adds units l, l + 32, ... in order and a fixed
    // shuffle tree adds the lanes -- still independent of the schedule, with 32 loads in flight instead of one
    for (int u = warp; u < A.n_umodes; u += nwarp) {
      double sr = 0.0, si = 0.0;
      for (int ch = lane; ch < A.parts; ch += 32) {
        const double2 p = __ldcg(part + (long long)ch * A.n_umodes + u);
        sr += p.x;
        si += p.y;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        sr += __shfl_xor_sync(0xffffffffu, sr, o);
        si += __shfl_xor_sync(0xffffffffu, si, o);
      }
      if (lane == 0) Iu[u] = A.factor * (sr * sr + si * si);
    }
  }
  __syncthreads();
  for (int s = warp; s < A.n_shells; s += nwarp) {
    double acc = 0.0;
    for (int m = A.shell_off[s] + lane; m < A.shell_off[s + 1]; m += 32) acc += Iu[A.memb_u[m]];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) {
      row[s] = acc;
      A.shell_I[(long long)field * A.n_shells + s] = acc;
    }
  }
  if (!A.do_peaks) return;
  __syncthreads();
  peaks_block(A.q, row, A.n_shells, A.pp, A.plot_y ? A.plot_y + (long long)field * A.pp.step : nullptr,
              A.peak_idx ? A.peak_idx + (long long)field * A.pp.max_peaks : nullptr,
              A.n_peaks ? A.n_peaks + field : nullptr, work + A.n_shells);
}

__global__ void __launch_bounds__(256, 3)
pruned_fused_kernel(const FusedArgs A) {
  extern __shared__ double sm[];
  double* twf = sm;                                   // ds * 128
  double* tw1 = twf + (size_t)A.ds * 128;             // 2 * ks1 * 32
  double2* tw0 = reinterpret_cast<double2*>(tw1 + (size_t)2 * A.ks1 * 32);  // N0
  double* work = reinterpret_cast<double*>(tw0 + A.N0);
  double* gT = work;                                  // 16 * gstride
  double* c2 = gT + (size_t)16 * A.gstride;           // 256
  double* hp = c2 + 256;                              // 256
  __shared__ int s_unit, s_last;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < A.ds * 128; i += 256) twf[i] = A.twf[i];
  for (int i = tid; i < 2 * A.ks1 * 32; i += 256) tw1[i] = A.tw1[i];
  for (int i = tid; i < A.N0; i += 256) tw0[i] = A.tw0[i];

  // this thread's listed modes
  int4 md[kFusedModesPerThread];
  int orc[kFusedModesPerThread];
#pragma unroll
  for (int k = 0; k < kFusedModesPerThread; ++k) {
    const int u = tid + 256 * k;
    md[k] = u < A.n_umodes ? A.umodes[u] : make_int4(0, 0, 0, 0);
    orc[k] = md[k].z * 16 + (md[k].y < 0 ? -md[k].y : md[k].y);
  }
  const int m = lane >> 2, kk = lane & 3;
  const int groups = A.N1 >> 3;
  const int n_units = A.n_fields * A.chunks;
  const long long plane_doubles = (long long)A.N1 * A.N2;
  // gT rows of this lane's accumulator columns n = 2 kk, 2 kk + 1 (even-frequency tile; odd tile = +1)
  const int n0 = 2 * kk, n1 = 2 * kk + 1;
  const int re0 = n0 < 4 ? 2 * n0 : 8 + 2 * (n0 - 4), re1 = n1 < 4 ? 2 * n1 : 8 + 2 * (n1 - 4);
  __syncthreads();

  for (;;) {
    if (tid == 0) s_unit = (int)atomicAdd(&A.ctrl[0], 1u);
    __syncthreads();
    const int unit = s_unit;
    if (unit >= n_units) break;
    const int field = unit / A.chunks, chunk = unit - field * A.chunks;
    const int x0 = chunk * A.xc;
    double fr[kFusedModesPerThread], fi[kFusedModesPerThread];
    int jj[kFusedModesPerThread];
#pragma unroll
    for (int k = 0; k < kFusedModesPerThread; ++k) {
      fr[k] = 0.0; fi[k] = 0.0;
      jj[k] = (int)(((long long)md[k].x * x0) % A.N0);
    }
    for (int x = x0; x < x0 + A.xc; ++x) {
      const double* base = A.fields + ((long long)field * A.N0 + x) * plane_doubles;
      // ---- stage 1 (z, folded): each warp transforms 8 rows at a time
      for (int grp = warp; grp < groups; grp += 8) {
        {
          // The warp's next row group (same plane, or its first group of the unit's next plane) on its way into L2
          // while this one is transformed: 8 N2 contiguous doubles, one 128-byte line per lane and step.  No registers
          // held across the transform; measured 0.886 -> 0.829 ms per 256 x 128^3 with the same register allocation.
          const long long nd = grp + 8 < groups ? (long long)(grp + 8) * 8 * A.N2
                                                : (x + 1 < x0 + A.xc ? plane_doubles + (long long)warp * 8 * A.N2 : -1);
          if (nd >= 0)
            for (int off = lane * 128; off < A.N2 * 64; off += 4096)
              asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(base + nd) + off));
        }
        const double2* row = reinterpret_cast<const double2*>(base + (long long)(grp * 8 + m) * A.N2) + kk;
        const double2* row2 = row + (A.N2 >> 2);
        double e0 = 0.0, e1 = 0.0, o0 = 0.0, o1 = 0.0;
#pragma unroll 4
        for (int s = 0; s < A.ds; ++s) {
          const double2 a = __ldg(row + 4 * s), b = __ldg(row2 + 4 * s);  // z = 8 s + 2 kk (+1), and + N2/2
          const double* tw = twf + s * 128 + lane;
          dmma_m8n8k4(e0, e1, a.x + b.x, tw[0]);
          dmma_m8n8k4(e0, e1, a.y + b.y, tw[32]);
          dmma_m8n8k4(o0, o1, a.x - b.x, tw[64]);
          dmma_m8n8k4(o0, o1, a.y - b.y, tw[96]);
        }
        const int y = grp * 8 + m;
        gT[(size_t)re0 * A.gstride + y] = e0;
        gT[(size_t)re1 * A.gstride + y] = e1;
        gT[(size_t)(re0 + 1) * A.gstride + y] = o0;
        gT[(size_t)(re1 + 1) * A.gstride + y] = o1;
      }
      __syncthreads();
      // ---- stage 2 (y): 16 x 16 x N1 product, 4 output tiles x 2 K halves over the 8 warps
      {
        const int mt = warp & 1, nt = (warp >> 1) & 1, kh = warp >> 2;
        const int s0 = kh * (A.ks1 >> 1), s1 = s0 + (A.ks1 >> 1);
        const double* arow = gT + (size_t)(8 * mt + m) * A.gstride;
        const double* brow = tw1 + (size_t)nt * A.ks1 * 32;
        double d0 = 0.0, d1 = 0.0;
#pragma unroll 4
        for (int s = s0; s < s1; ++s) dmma_m8n8k4(d0, d1, arow[4 * s + kk], brow[s * 32 + lane]);
        const int r = 8 * mt + m, c = 8 * nt + 2 * kk;
        if (kh == 1) { c2[r * 16 + c] = d0; c2[r * 16 + c + 1] = d1; }
        __syncthreads();
        if (kh == 0) { hp[r * 16 + c] = d0 + c2[r * 16 + c]; hp[r * 16 + c + 1] = d1 + c2[r * 16 + c + 1]; }
      }
      __syncthreads();
      // ---- stage 3 (x): one thread per listed mode accumulates H[x,b,c] e^{-2 pi i a x / N0}
#pragma unroll
      for (int k = 0; k < kFusedModesPerThread; ++k) {
        if (tid + 256 * k < A.n_umodes) {
          const double rc = hp[orc[k]], rs = hp[orc[k] + 8], sc = hp[orc[k] + 128], ss = hp[orc[k] + 136];
          const double hr = md[k].y >= 0 ? rc - ss : rc + ss;
          const double hi = md[k].y >= 0 ? -(rs + sc) : rs - sc;
          const double2 w = tw0[jj[k]];
          fr[k] += hr * w.x + hi * w.y;
          fi[k] += hi * w.x - hr * w.y;
          jj[k] += md[k].x;
          if (jj[k] >= A.N0) jj[k] -= A.N0;
        }
      }
      // hp is rewritten only after the next plane's stage-1 barrier; gT after this plane's stage 2
    }
    double2* part = A.partial + (long long)unit * A.n_umodes;
#pragma unroll
    for (int k = 0; k < kFusedModesPerThread; ++k)
      if (tid + 256 * k < A.n_umodes) part[tid + 256 * k] = make_double2(fr[k], fi[k]);
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(&A.ctrl[1 + field], 1u) == (unsigned)(A.chunks - 1);
    __syncthreads();
    if (s_last) {
      __threadfence();
      fused_finish(A, field, work);
    }
    __syncthreads();  // work / s_unit are reused by the next unit
  }
}

// ------------------------------------------------------------------------------------------
// Blocked variant for grids whose planes do not fit the scheme above (512^3, 1024^3) and for slabs of
// one field spread over several GPUs.  Same stages; differences:
//  * a unit = xc planes x ONE block of yb_rows rows, so the stage-1 buffer is 16 x (yb_rows + 4) whatever N1
//    is; the y transform of a block uses the block's slice of the twiddle table and its 16 x 16 result goes
//    straight through the (linear) x stage -- partial sums simply have more terms;
//  * z and y twiddles are read through L1 (__ldg) instead of being copied to shared memory (64 KB +
//    128 KB at 1024^3), leaving room for 3 CTAs per SM;
//  * `planes` / `x_base`: the field pointer holds only this rank's planes, the x twiddle uses the global
//    plane index; `finish = 0` stops after the partial sums (summed over ranks before the shell sums).
__host__ __device__ inline size_t blocked_work_doubles(int yb_rows, int n_shells, int n_umodes, int step, bool peaks) {
  const size_t a = (size_t)16 * (yb_rows + 4) + 512;
  const size_t b = (size_t)n_shells + n_umodes + 2 + (peaks ? (size_t)step + 2 * (size_t)n_shells : 0);
  return a > b ? a : b;
}
// TWF_SMEM: the z twiddles (N2 / 16 KB) are copied to shared memory when they are at most 32 KB (N2 <= 512) --
// reading them through L1 costs 10 % at 128^3
__host__ __device__ inline bool blocked_twf_in_smem(int N2) { return (size_t)(N2 / 16) * 1024 <= 32 * 1024; }
__host__ __device__ inline size_t blocked_smem_bytes(int N0, int N2, size_t work_doubles) {
  return (2 * (size_t)N0 + work_doubles + (blocked_twf_in_smem(N2) ? (size_t)(N2 / 16) * 128 : 0)) * sizeof(double);
}

template <bool TWF_SMEM>
__global__ void __launch_bounds__(256, 3)
pruned_blocked_kernel(const FusedArgs A) {
  extern __shared__ double sm[];
  double2* tw0 = reinterpret_cast<double2*>(sm);      // N0
  double* twf_s = reinterpret_cast<double*>(tw0 + A.N0);  // ds * 128 when TWF_SMEM
  double* work = twf_s + (TWF_SMEM ? (size_t)A.ds * 128 : 0);
  const int gstride = A.yb_rows + 4;
  double* gT = work;                                  // 16 * gstride
  double* c2 = gT + (size_t)16 * gstride;             // 256
  double* hp = c2 + 256;                              // 256
  __shared__ int s_unit, s_last;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < A.N0; i += 256) tw0[i] = A.tw0[i];
  if (TWF_SMEM)
    for (int i = tid; i < A.ds * 128; i += 256) twf_s[i] = A.twf[i];
  int4 md[kFusedModesPerThread];
  int orc[kFusedModesPerThread];
#pragma unroll
  for (int k = 0; k < kFusedModesPerThread; ++k) {
    const int u = tid + 256 * k;
    md[k] = u < A.n_umodes ? A.umodes[u] : make_int4(0, 0, 0, 0);
    orc[k] = md[k].z * 16 + (md[k].y < 0 ? -md[k].y : md[k].y);
  }
  const int m = lane >> 2, kk = lane & 3;
  const int groups = A.yb_rows >> 3, ksb = A.yb_rows >> 2;
  const int n_units = A.n_fields * A.parts;
  const long long plane_doubles = (long long)A.N1 * A.N2;
  const int n0 = 2 * kk, n1 = 2 * kk + 1;
  const int re0 = n0 < 4 ? 2 * n0 : 8 + 2 * (n0 - 4), re1 = n1 < 4 ? 2 * n1 : 8 + 2 * (n1 - 4);
  __syncthreads();

  for (;;) {
    if (tid == 0) s_unit = (int)atomicAdd(&A.ctrl[0], 1u);
    __syncthreads();
    const int unit = s_unit;
    if (unit >= n_units) break;
    const int field = unit / A.parts, part = unit - field * A.parts;
    const int chunk = part / A.yblocks, yb = part - chunk * A.yblocks;
    const int x0 = chunk * A.xc;
    double fr[kFusedModesPerThread], fi[kFusedModesPerThread];
    int jj[kFusedModesPerThread];
#pragma unroll
    for (int k = 0; k < kFusedModesPerThread; ++k) {
      fr[k] = 0.0; fi[k] = 0.0;
      jj[k] = (int)(((long long)md[k].x * (A.x_base + x0)) % A.N0);
    }
    for (int x = x0; x < x0 + A.xc; ++x) {
      const double* base = A.fields + ((long long)field * A.planes + x) * plane_doubles + (long long)yb * A.yb_rows * A.N2;
      // ---- stage 1 (z, folded): each warp transforms 8 rows at a time
      for (int grp = warp; grp < groups; grp += 8) {
        const double2* row = reinterpret_cast<const double2*>(base + (long long)(grp * 8 + m) * A.N2) + kk;
        const double2* row2 = row + (A.N2 >> 2);
        double e0 = 0.0, e1 = 0.0, o0 = 0.0, o1 = 0.0;
#pragma unroll 4
        for (int s = 0; s < A.ds; ++s) {
          const double2 a = __ldg(row + 4 * s), b = __ldg(row2 + 4 * s);
          const double* tw = (TWF_SMEM ? twf_s : A.twf) + s * 128 + lane;
          dmma_m8n8k4(e0, e1, a.x + b.x, TWF_SMEM ? tw[0] : __ldg(tw));
          dmma_m8n8k4(e0, e1, a.y + b.y, TWF_SMEM ? tw[32] : __ldg(tw + 32));
          dmma_m8n8k4(o0, o1, a.x - b.x, TWF_SMEM ? tw[64] : __ldg(tw + 64));
          dmma_m8n8k4(o0, o1, a.y - b.y, TWF_SMEM ? tw[96] : __ldg(tw + 96));
        }
        const int y = grp * 8 + m;
        gT[(size_t)re0 * gstride + y] = e0;
        gT[(size_t)re1 * gstride + y] = e1;
        gT[(size_t)(re0 + 1) * gstride + y] = o0;
        gT[(size_t)(re1 + 1) * gstride + y] = o1;
      }
      __syncthreads();
      // ---- stage 2 (y): 16 x 16 x yb_rows product with this block's slice of the twiddles
      {
        const int mt = warp & 1, nt = (warp >> 1) & 1, kh = warp >> 2;
        const int s0 = kh * (ksb >> 1), s1 = s0 + (ksb >> 1);
        const double* arow = gT + (size_t)(8 * mt + m) * gstride;
        const double* brow = A.tw1 + ((size_t)nt * A.ks1 + (size_t)yb * ksb) * 32;
        double d0 = 0.0, d1 = 0.0;
#pragma unroll 4
        for (int s = s0; s < s1; ++s) dmma_m8n8k4(d0, d1, arow[4 * s + kk], __ldg(brow + s * 32 + lane));
        const int r = 8 * mt + m, c = 8 * nt + 2 * kk;
        if (kh == 1) { c2[r * 16 + c] = d0; c2[r * 16 + c + 1] = d1; }
        __syncthreads();
        if (kh == 0) { hp[r * 16 + c] = d0 + c2[r * 16 + c]; hp[r * 16 + c + 1] = d1 + c2[r * 16 + c + 1]; }
      }
      __syncthreads();
      // ---- stage 3 (x)
#pragma unroll
      for (int k = 0; k < kFusedModesPerThread; ++k) {
        if (tid + 256 * k < A.n_umodes) {
          const double rc = hp[orc[k]], rs = hp[orc[k] + 8], sc = hp[orc[k] + 128], ss = hp[orc[k] + 136];
          const double hr = md[k].y >= 0 ? rc - ss : rc + ss;
          const double hi = md[k].y >= 0 ? -(rs + sc) : rs - sc;
          const double2 w = tw0[jj[k]];
          fr[k] += hr * w.x + hi * w.y;
          fi[k] += hi * w.x - hr * w.y;
          jj[k] += md[k].x;
          if (jj[k] >= A.N0) jj[k] -= A.N0;
        }
      }
    }
    double2* part_out = A.partial + (long long)unit * A.n_umodes;
#pragma unroll
    for (int k = 0; k < kFusedModesPerThread; ++k)
      if (tid + 256 * k < A.n_umodes) part_out[tid + 256 * k] = make_double2(fr[k], fi[k]);
    if (A.finish) {
      __threadfence();
      __syncthreads();
      if (tid == 0) s_last = atomicAdd(&A.ctrl[1 + field], 1u) == (unsigned)(A.parts - 1);
      __syncthreads();
      if (s_last) {
        __threadfence();
        fused_finish(A, field, work);
      }
    }
    __syncthreads();
  }
}

// slab path: sums[mode] = sum over this rank's units in fixed order (one warp per mode: lane l adds units l, l + 32, ...
// in order, then a fixed shuffle tree)
__global__ void fused_reduce_parts_kernel(const double2* __restrict__ partial, int parts, int n_umodes, double2* __restrict__ sums) {
  const int u = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (u >= n_umodes) return;
  double sr = 0.0, si = 0.0;
  for (int p = lane; p < parts; p += 32) {
    const double2 v = partial[(long long)p * n_umodes + u];
    sr += v.x;
    si += v.y;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sr += __shfl_xor_sync(0xffffffffu, sr, o);
    si += __shfl_xor_sync(0xffffffffu, si, o);
  }
  if (lane == 0) sums[u] = make_double2(sr, si);
}

// slab path: shell sums (+ peaks) from mode sums that are already complete (A.partial = sums, A.parts = 1)
__global__ void __launch_bounds__(256) fused_finish_kernel(const FusedArgs A) {
  extern __shared__ double sm[];
  fused_finish(A, 0, sm);
}

// ------------------------------------------------------------------------------------------
// TMA-ring variant (the default when it applies).  The LDG kernel above only has loads in flight while
// a warp sits in its stage-1 loop; across the plane barriers and the y / x stages HBM idles.  Here one
// persistent CTA of 16 warps per SM owns a ring of shared-memory slots, each holding a whole 8-row
// group (8 N2 doubles, contiguous in the field) that arrives by ONE TMA bulk copy (cp.async.bulk +
// mbarrier complete_tx; SASS UBLKCP).  Groups are numbered q = 0, 1, 2, ... through the CTA's whole
// work (all planes of all its units); group q lives in slot q % S.  Warps take the next group number
// from a shared counter, wait for its slot, transform it, and the same warp immediately refills the
// slot with group q + S -- which may belong to the next plane or the next unit (units come from an
// atomic counter, fetched two ahead) -- so ~180 KB per SM are in flight whatever the CTA is doing
// (plane barriers, y / x stages, a field's finish).  Stage-2 twiddles are read through L1 to leave the
// shared memory to the ring.  Requires N2 % 16 == 0, N1 % 8 == 0, at most 512 listed modes.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pf_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void pf_mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(pf_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void pf_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(pf_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void pf_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   pf_smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(pf_smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void pf_mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "PF_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra PF_DONE;\n"
      "bra PF_WAIT;\n"
      "PF_DONE:\n"
      "}\n" ::"r"(pf_smem_u32(bar)),
      "r"(parity)
      : "memory");
}


constexpr int kRingThreads = 512;   // 16 warps, one CTA per SM
constexpr int kRingMaxSlots = 64;

__host__ __device__ inline size_t ring_slot_bytes(int N2) { return (size_t)8 * N2 * 8; }
__host__ __device__ inline size_t ring_work_doubles(int N1, int n_shells, int n_umodes, int step, bool peaks) {
  const size_t a = (size_t)16 * (N1 + 4) + 1024;  // gT | c2 (3 x 256) | hp (256)
  const size_t b = (size_t)n_shells + n_umodes + 2 + (peaks ? peaks_smem_doubles(n_shells, step) : 0);
  return a > b ? a : b;
}
__host__ __device__ inline size_t ring_fixed_bytes(int N0, int N2, size_t work_doubles) {
  return ((size_t)(N2 / 16) * 128 + 2 * (size_t)N0 + work_doubles) * sizeof(double) + kRingMaxSlots * 8;
}

// NW warps per CTA: 16 (one CTA per SM) or 8 (two CTAs per SM, so that one CTA's plane barriers, y / x stages and
// field finish overlap the other's z stage).
template <int NW>
__global__ void __launch_bounds__(32 * NW, 16 / NW)
pruned_ring_kernel(const FusedArgs A) {
  constexpr int NT = 32 * NW;          // threads
  constexpr int KQ = NW / 4;           // K splits of the y-stage product
  constexpr int MPT = kRingThreads / NT;  // listed modes per thread (n_umodes <= kRingThreads)
  extern __shared__ __align__(128) unsigned char smraw[];
  const int S = A.ring_slots;
  const size_t slot_bytes = ring_slot_bytes(A.N2);
  unsigned char* ring = smraw;
  uint64_t* full = reinterpret_cast<uint64_t*>(ring + (size_t)S * slot_bytes);
  double* twf = reinterpret_cast<double*>(full + kRingMaxSlots);          // ds * 128
  double2* tw0 = reinterpret_cast<double2*>(twf + (size_t)A.ds * 128);    // N0
  double* work = reinterpret_cast<double*>(tw0 + A.N0);
  double* gT = work;                                  // 16 * gstride
  double* c2 = gT + (size_t)16 * A.gstride;           // 3 * 256
  double* hp = c2 + 768;                              // 256
  __shared__ int s_units[4];
  __shared__ const double* s_ubase[4];
  __shared__ int s_q, s_last;
  __shared__ volatile int s_issued[kRingMaxSlots];  // how many times each slot has been requested
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < kRingMaxSlots) s_issued[tid] = 0;
  for (int i = tid; i < A.ds * 128; i += NT) twf[i] = A.twf[i];
  for (int i = tid; i < A.N0; i += NT) tw0[i] = A.tw0[i];
  if (tid < S) pf_mbar_init(&full[tid], 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  // first plane of unit u (nullptr past the end), so that a refill needs no division
  auto unit_base = [&](int u) -> const double* {
    if (u >= A.n_fields * A.chunks) return nullptr;
    const int f = u / A.chunks, ch = u - f * A.chunks;
    return A.fields + ((long long)f * A.N0 + (long long)ch * A.xc) * ((long long)A.N1 * A.N2);
  };
  if (tid == 0) {
    s_units[0] = (int)atomicAdd(&A.ctrl[0], 1u);
    s_units[1] = (int)atomicAdd(&A.ctrl[0], 1u);
    s_ubase[0] = unit_base(s_units[0]);
    s_ubase[1] = unit_base(s_units[1]);
    s_q = NW;  // group numbers 0..NW-1 start out with warps 0..NW-1
  }
  // this thread's listed mode
  int4 md[MPT];
  int orc[MPT];
#pragma unroll
  for (int j = 0; j < MPT; ++j) {
    const int u = tid + NT * j;
    md[j] = u < A.n_umodes ? A.umodes[u] : make_int4(0, 0, 0, 0);
    orc[j] = md[j].z * 16 + (md[j].y < 0 ? -md[j].y : md[j].y);
  }
  const int m = lane >> 2, kk = lane & 3;
  const int groups = A.N1 >> 3;
  const int n_units = A.n_fields * A.chunks;
  const long long plane_doubles = (long long)A.N1 * A.N2;
  const int n0 = 2 * kk, n1 = 2 * kk + 1;
  const int re0 = n0 < 4 ? 2 * n0 : 8 + 2 * (n0 - 4), re1 = n1 < 4 ? 2 * n1 : 8 + 2 * (n1 - 4);
  __syncthreads();

  // lane 0: request group q (CTA-local numbering) into its slot; nothing to do past the last unit.
  // groups per plane, planes per unit and ring slots are powers of two: shifts, no integer division.
  const int g_sh = A.groups_shift, x_sh = A.xc_shift, s_mask = S - 1, s_sh = A.slots_shift;
  auto issue = [&](int q) {
    if (lane == 0) {
      const int pn = q >> g_sh, g = q & (groups - 1);
      const int kn = pn >> x_sh, xi = pn & (A.xc - 1);
      const double* ubase = s_ubase[kn & 3];
      if (ubase) {
        const double* rows = ubase + (long long)xi * plane_doubles + (long long)g * 8 * A.N2;
        const int sl = q & s_mask;
        pf_mbar_expect_tx(&full[sl], (uint32_t)slot_bytes);
        pf_bulk_g2s(ring + (size_t)sl * slot_bytes, rows, (uint32_t)slot_bytes, &full[sl]);
        __threadfence_block();
        s_issued[sl] = (q >> s_sh) + 1;
      }
    }
  };
  for (int q = warp; q < S; q += NW) issue(q);
  int myq = warp;

  for (int k = 0;; ++k) {
    const int unit = s_units[k & 3];
    if (unit >= n_units) break;
    if (tid == 0) {  // read (after barriers) from unit k+1 on
      const int u2 = (int)atomicAdd(&A.ctrl[0], 1u);
      s_units[(k + 2) & 3] = u2;
      s_ubase[(k + 2) & 3] = unit_base(u2);
    }
    const int field = unit / A.chunks, chunk = unit - field * A.chunks;
    const int x0 = chunk * A.xc;
    double fr[MPT], fi[MPT];
    int jj[MPT];
#pragma unroll
    for (int j = 0; j < MPT; ++j) {
      fr[j] = 0.0; fi[j] = 0.0;
      jj[j] = (int)(((long long)md[j].x * x0) % A.N0);
    }
    for (int xi = 0; xi < A.xc; ++xi) {
      const int plane_start = (k * A.xc + xi) * groups, plane_end = plane_start + groups;
      // ---- stage 1 (z, folded): warps take group numbers until the plane is used up
      while (myq < plane_end) {
        const int sl = myq & s_mask, use = myq >> s_sh;
        // a parity wait is only meaningful once the slot's previous use is over: wait until this use has been
        // requested (which happens after the previous group was consumed), then for its bytes
        if (lane == 0) while (s_issued[sl] <= use) {}
        __syncwarp();
        pf_mbar_wait(&full[sl], (uint32_t)use & 1u);
        const unsigned char* slot = ring + (size_t)sl * slot_bytes;
        const double2* ra = reinterpret_cast<const double2*>(slot + (size_t)m * A.N2 * 8) + kk;
        const double2* rb = ra + (A.N2 >> 2);
        double e0 = 0.0, e1 = 0.0, o0 = 0.0, o1 = 0.0;
#pragma unroll 4
        for (int s = 0; s < A.ds; ++s) {
          const double2 a = ra[4 * s], b = rb[4 * s];
          const double* tw = twf + s * 128 + lane;
          dmma_m8n8k4(e0, e1, a.x + b.x, tw[0]);
          dmma_m8n8k4(e0, e1, a.y + b.y, tw[32]);
          dmma_m8n8k4(o0, o1, a.x - b.x, tw[64]);
          dmma_m8n8k4(o0, o1, a.y - b.y, tw[96]);
        }
        __syncwarp();
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        issue(myq + S);
        const int y = (myq - plane_start) * 8 + m;
        gT[(size_t)re0 * A.gstride + y] = e0;
        gT[(size_t)re1 * A.gstride + y] = e1;
        gT[(size_t)(re0 + 1) * A.gstride + y] = o0;
        gT[(size_t)(re1 + 1) * A.gstride + y] = o1;
        int nq = 0;
        if (lane == 0) nq = atomicAdd(&s_q, 1);
        myq = __shfl_sync(0xffffffffu, nq, 0);
      }
      __syncthreads();
      // ---- stage 2 (y): 16 x 16 x N1 product, 4 output tiles x KQ K-parts over the warps
      {
        const int mt = warp & 1, nt = (warp >> 1) & 1, kh = warp >> 2;
        const int s0 = kh * (A.ks1 / KQ), s1 = s0 + (A.ks1 / KQ);
        const double* arow = gT + (size_t)(8 * mt + m) * A.gstride;
        const double* brow = A.tw1 + (size_t)nt * A.ks1 * 32;
        double d0 = 0.0, d1 = 0.0;
#pragma unroll 4
        for (int s = s0; s < s1; ++s) dmma_m8n8k4(d0, d1, arow[4 * s + kk], __ldg(brow + s * 32 + lane));
        const int r = 8 * mt + m, c = 8 * nt + 2 * kk;
        if (kh > 0) { c2[(kh - 1) * 256 + r * 16 + c] = d0; c2[(kh - 1) * 256 + r * 16 + c + 1] = d1; }
        __syncthreads();
        if (kh == 0) {
#pragma unroll
          for (int j = 0; j < KQ - 1; ++j) {
            d0 += c2[j * 256 + r * 16 + c];
            d1 += c2[j * 256 + r * 16 + c + 1];
          }
          hp[r * 16 + c] = d0;
          hp[r * 16 + c + 1] = d1;
        }
      }
      __syncthreads();
      // ---- stage 3 (x): one thread per listed mode
#pragma unroll
      for (int j = 0; j < MPT; ++j) {
        if (tid + NT * j < A.n_umodes) {
          const double rc = hp[orc[j]], rs = hp[orc[j] + 8], sc = hp[orc[j] + 128], ss = hp[orc[j] + 136];
          const double hr = md[j].y >= 0 ? rc - ss : rc + ss;
          const double hi = md[j].y >= 0 ? -(rs + sc) : rs - sc;
          const double2 w = tw0[jj[j]];
          fr[j] += hr * w.x + hi * w.y;
          fi[j] += hi * w.x - hr * w.y;
          jj[j] += md[j].x;
          if (jj[j] >= A.N0) jj[j] -= A.N0;
        }
      }
    }
#pragma unroll
    for (int j = 0; j < MPT; ++j)
      if (tid + NT * j < A.n_umodes) A.partial[(long long)unit * A.n_umodes + tid + NT * j] = make_double2(fr[j], fi[j]);
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(&A.ctrl[1 + field], 1u) == (unsigned)(A.chunks - 1);
    __syncthreads();
    if (s_last) {
      __threadfence();
      fused_finish(A, field, work);
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------
// Bulk-staged blocked variant (giant grids: 512^3, 1024^3, and the slabs of one field spread over GPUs).
// The LDG blocked kernel reads a row group as 16 streams of 64-byte pieces (8 rows x 2 halves, rows 4-8 KB
// apart): at N2 = 1024 that reaches ~52 % of the HBM peak.  Here every warp owns a private two-stage pipeline
// in shared memory.  One pipeline item = one segment of 64 samples of both halves of the warp's 8 rows:
// 16 pieces of 512 contiguous bytes, each fetched by one TMA bulk copy (cp.async.bulk + mbarrier complete_tx;
// SASS UBLKCP) issued by lanes 0..15.  The warp transforms the stage that has landed while the other one is
// in flight; the item sequence of a warp runs through all segments, row groups and planes of its unit, so
// copies stay in flight across the plane barriers and the y / x stages.  Accumulators live in registers across
// the segments of a row group.  Stage layout: pieces of even rows | 64 B | pieces of odd rows, so that the two
// rows a quarter-warp reads with one 16-byte load per lane fall into disjoint banks.
// One CTA of 8 warps per SM: 2 x 8 x 8.06 KB of stages + the z twiddles (N2 x 64 B) + the unit's work area.
// Requires N2 % 128 == 0, yb_rows % 64 == 0 and at least two items per unit.  Same unit / partial-sum / finish
// scheme, hence (for the same planes per unit) the same bits, as pruned_blocked_kernel.  A CTA knows its next unit
// one unit ahead, and the item sequence of a warp runs on into it: no pipeline bubble between units.
// ------------------------------------------------------------------------------------------
constexpr int kBulkSeg = 64;                                  // doubles per piece
constexpr int kBulkPiece = kBulkSeg * 8;                      // 512 bytes
constexpr int kBulkStageBytes = 16 * kBulkPiece + 64;
constexpr int kBulkStages = 2;
__host__ __device__ inline size_t bulk_smem_bytes(int N2, size_t work_doubles) {
  return (size_t)8 * kBulkStages * kBulkStageBytes + (size_t)8 * kBulkStages * 8 + ((size_t)(N2 / 16) * 128 + work_doubles) * sizeof(double);
}
__device__ __forceinline__ int bulk_piece_offset(int row, int half) {
  return ((row >> 1) * 2 + half) * kBulkPiece + (row & 1) * (8 * kBulkPiece + 64);
}

__global__ void __launch_bounds__(256, 1)
pruned_bulk_kernel(const FusedArgs A) {
  extern __shared__ __align__(128) unsigned char smraw[];
  unsigned char* stages = smraw;                                            // [8 warps][kBulkStages][kBulkStageBytes]
  uint64_t* bars = reinterpret_cast<uint64_t*>(stages + (size_t)8 * kBulkStages * kBulkStageBytes);
  double* twf = reinterpret_cast<double*>(bars + 8 * kBulkStages);         // ds * 128
  double* work = twf + (size_t)A.ds * 128;
  const int gstride = A.yb_rows + 4;
  double* gT = work;                                  // 16 * gstride
  double* c2 = gT + (size_t)16 * gstride;             // 256
  double* hp = c2 + 256;                              // 256
  __shared__ int s_units[2], s_last;  // units k and k + 1 of this CTA (k & 1): the pipeline runs one unit ahead
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < A.ds * 128; i += 256) twf[i] = A.twf[i];
  if (tid < 8 * kBulkStages) pf_mbar_init(&bars[tid], 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  int4 md[kFusedModesPerThread];
  int orc[kFusedModesPerThread];
#pragma unroll
  for (int k = 0; k < kFusedModesPerThread; ++k) {
    const int u = tid + 256 * k;
    md[k] = u < A.n_umodes ? A.umodes[u] : make_int4(0, 0, 0, 0);
    orc[k] = md[k].z * 16 + (md[k].y < 0 ? -md[k].y : md[k].y);
  }
  const int m = lane >> 2, kk = lane & 3;
  const int gpw = A.yb_rows >> 6;          // row groups per warp and plane (groups / 8)
  const int nseg = A.N2 >> 7;              // segments per row: (N2 / 2) / kBulkSeg
  const int ksb = A.yb_rows >> 2;
  const int n_units = A.n_fields * A.parts;
  const int n_items = A.xc * gpw * nseg;   // pipeline items of one warp in one unit
  const long long plane_doubles = (long long)A.N1 * A.N2;
  const int n0 = 2 * kk, n1 = 2 * kk + 1;
  const int re0 = n0 < 4 ? 2 * n0 : 8 + 2 * (n0 - 4), re1 = n1 < 4 ? 2 * n1 : 8 + 2 * (n1 - 4);
  unsigned char* my_stages = stages + (size_t)warp * kBulkStages * kBulkStageBytes;
  uint64_t* my_bars = bars + warp * kBulkStages;
  // this lane's two pieces of a stage (A fragment: row m, samples 2 kk, 2 kk + 1 of every 8)
  const int off_a = bulk_piece_offset(m, 0) + kk * 16;
  // lanes 0..15 fetch piece (row = lane & 7, half = lane >> 3)
  const int put_off = bulk_piece_offset(lane & 7, (lane >> 3) & 1);
  const long long put_src = (long long)(lane & 7) * A.N2 + (long long)((lane >> 3) & 1) * (A.N2 >> 1);
  unsigned it = 0;  // items this warp has consumed so far (all units): stage = it % 2, parity = (it / 2) % 2
  if (tid == 0) {
    s_units[0] = (int)atomicAdd(&A.ctrl[0], 1u);
    s_units[1] = (int)atomicAdd(&A.ctrl[0], 1u);
  }
  __syncthreads();
  // first row of unit u (nullptr past the end)
  auto unit_rows = [&](int u) -> const double* {
    if (u >= n_units) return nullptr;
    const int f = u / A.parts, pt = u - f * A.parts;
    const int ch = pt / A.yblocks, b = pt - ch * A.yblocks;
    return A.fields + ((long long)f * A.planes + (long long)ch * A.xc) * plane_doubles + (long long)b * A.yb_rows * A.N2;
  };

  for (int k = 0;; ++k) {
    const int unit = s_units[k & 1];
    if (unit >= n_units) break;
    const int field = unit / A.parts, part = unit - field * A.parts;
    const int chunk = part / A.yblocks, yb = part - chunk * A.yblocks;
    const int x0 = chunk * A.xc;
    const double* ubase = unit_rows(unit);
    const double* nbase = unit_rows(s_units[(k + 1) & 1]);  // the CTA's next unit: its first two items are requested
                                                             // while this unit finishes (n_items >= 2)
    // request item i into stage st (whole warp); items n_items, n_items + 1 are items 0, 1 of the next unit
    auto issue = [&](int i, int st) {
      const double* ub = ubase;
      if (i >= n_items) { ub = nbase; i -= n_items; }
      if (!ub) return;
      const int sg = i % nseg, t2 = i / nseg;
      const int gi = t2 % gpw, xi = t2 / gpw;
      const double* src = ub + (long long)xi * plane_doubles + (long long)((warp + 8 * gi) * 8) * A.N2 + sg * kBulkSeg;
      if (lane == 0) pf_mbar_expect_tx(&my_bars[st], 16 * kBulkPiece);
      __syncwarp();
      if (lane < 16) pf_bulk_g2s(my_stages + (size_t)st * kBulkStageBytes + put_off, src + put_src, kBulkPiece, &my_bars[st]);
    };
    if (k == 0) {  // later units find their first two items already requested
      issue(0, it & 1u);
      issue(1, (it + 1u) & 1u);
    }
    double fr[kFusedModesPerThread], fi[kFusedModesPerThread];
    int jj[kFusedModesPerThread];
#pragma unroll
    for (int k = 0; k < kFusedModesPerThread; ++k) {
      fr[k] = 0.0; fi[k] = 0.0;
      jj[k] = (int)(((long long)md[k].x * (A.x_base + x0)) % A.N0);
    }
    int item = 0;
    for (int xi = 0; xi < A.xc; ++xi) {
      // ---- stage 1 (z, folded): this warp's row groups, segment by segment out of its pipeline
      for (int gi = 0; gi < gpw; ++gi) {
        double e0 = 0.0, e1 = 0.0, o0 = 0.0, o1 = 0.0;
        for (int sg = 0; sg < nseg; ++sg, ++item, ++it) {
          const int st = it & 1u;
          pf_mbar_wait(&my_bars[st], (it >> 1) & 1u);
          const double2* ra = reinterpret_cast<const double2*>(my_stages + (size_t)st * kBulkStageBytes + off_a);
          const double2* rb = ra + kBulkPiece / 16;
          const double* tws = twf + (size_t)sg * (kBulkSeg / 8) * 128 + lane;
#pragma unroll
          for (int s = 0; s < kBulkSeg / 8; ++s) {
            const double2 a = ra[4 * s], b = rb[4 * s];
            const double* tw = tws + s * 128;
            dmma_m8n8k4(e0, e1, a.x + b.x, tw[0]);
            dmma_m8n8k4(e0, e1, a.y + b.y, tw[32]);
            dmma_m8n8k4(o0, o1, a.x - b.x, tw[64]);
            dmma_m8n8k4(o0, o1, a.y - b.y, tw[96]);
          }
          __syncwarp();
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          issue(item + 2, st);
        }
        const int y = (warp + 8 * gi) * 8 + m;
        gT[(size_t)re0 * gstride + y] = e0;
        gT[(size_t)re1 * gstride + y] = e1;
        gT[(size_t)(re0 + 1) * gstride + y] = o0;
        gT[(size_t)(re1 + 1) * gstride + y] = o1;
      }
      __syncthreads();
      // ---- stage 2 (y): 16 x 16 x yb_rows product with this block's slice of the twiddles
      {
        const int mt = warp & 1, nt = (warp >> 1) & 1, kh = warp >> 2;
        const int s0 = kh * (ksb >> 1), s1 = s0 + (ksb >> 1);
        const double* arow = gT + (size_t)(8 * mt + m) * gstride;
        const double* brow = A.tw1 + ((size_t)nt * A.ks1 + (size_t)yb * ksb) * 32;
        double d0 = 0.0, d1 = 0.0;
#pragma unroll 4
        for (int s = s0; s < s1; ++s) dmma_m8n8k4(d0, d1, arow[4 * s + kk], __ldg(brow + s * 32 + lane));
        const int r = 8 * mt + m, c = 8 * nt + 2 * kk;
        if (kh == 1) { c2[r * 16 + c] = d0; c2[r * 16 + c + 1] = d1; }
        __syncthreads();
        if (kh == 0) { hp[r * 16 + c] = d0 + c2[r * 16 + c]; hp[r * 16 + c + 1] = d1 + c2[r * 16 + c + 1]; }
      }
      __syncthreads();
      // ---- stage 3 (x)
#pragma unroll
      for (int k = 0; k < kFusedModesPerThread; ++k) {
        if (tid + 256 * k < A.n_umodes) {
          const double rc = hp[orc[k]], rs = hp[orc[k] + 8], sc = hp[orc[k] + 128], ss = hp[orc[k] + 136];
          const double hr = md[k].y >= 0 ? rc - ss : rc + ss;
          const double hi = md[k].y >= 0 ? -(rs + sc) : rs - sc;
          const double2 w = __ldg(A.tw0 + jj[k]);
          fr[k] += hr * w.x + hi * w.y;
          fi[k] += hi * w.x - hr * w.y;
          jj[k] += md[k].x;
          if (jj[k] >= A.N0) jj[k] -= A.N0;
        }
      }
    }
    double2* part_out = A.partial + (long long)unit * A.n_umodes;
#pragma unroll
    for (int k = 0; k < kFusedModesPerThread; ++k)
      if (tid + 256 * k < A.n_umodes) part_out[tid + 256 * k] = make_double2(fr[k], fi[k]);
    if (A.finish) {
      __threadfence();
      __syncthreads();
      if (tid == 0) s_last = atomicAdd(&A.ctrl[1 + field], 1u) == (unsigned)(A.parts - 1);
      __syncthreads();
      if (s_last) {
        __threadfence();
        fused_finish(A, field, work);
      }
    }
    if (tid == 0) s_units[k & 1] = (int)atomicAdd(&A.ctrl[0], 1u);  // unit k + 2 (everyone read unit k's slot long ago)
    __syncthreads();
  }
}

}  // namespace sbx
