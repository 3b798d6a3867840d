// pruned_dft.cuh -- direct (pruned) DFT of the few low-|q| modes the shell table needs (sm_100a).
//
// The reference takes a full N^3 FFT (SimuBox/calculator/scatter.py:85) and then keeps the
// cutoff+1 lowest shells (:99-115), i.e. a few hundred of the N^3 coefficients.  This path
// computes only those coefficients, reading every field exactly once from HBM:
//
//   stage 1 (z)  G[x,y,c]  = sum_z rho[x,y,z] e^{-2 pi i c z/N2},  c = 0..7
//                real input: pair z with N2-z, so  Gr = sum (rho[z]+rho[N2-z]) cos,
//                                                  Gs = sum (rho[z]-rho[N2-z]) sin   (G = Gr - i Gs)
//   stage 2 (y)  the four real sums  {Gr,Gs} x {cos,sin}(2 pi b y/N1),  b = 0..7, which give
//                H[x,+-b,c] = (RC -+ SS) -+ i (RS +- SC)
//   stage 3 (x)  F[a,b,c] = sum_x H[x,b,c] e^{-2 pi i a x/N0} for the listed modes only
//                (pruned_finish in scatter_kernels.cuh), then |F|^2/N, shell sums, peaks.
//
// Stages 1 and 2 are fp64 matrix products (rows x z-samples x 8 frequencies) and run on the
// tensor-core DMMA path (mma.sync m8n8k4 f64): the MMA performs the cross-lane reduction
// that a scalar formulation would pay for in shuffles.  One CTA handles whole x-planes; the
// stage-1 result of a plane never leaves shared memory.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace sbx {

struct PrunedGeo {
  int N0, N1, N2;
  int ks2;       // stage-1 k-steps: ceil((N2/2 + 1) / 4)
  int ks1;       // stage-2 k-steps: N1 / 4
  int gstride;   // row stride (doubles) of the transposed stage-1 plane in shared memory: N1 + 4
  const double* tw2;  // [ks2][2][32]  B fragments of stage 1: cos | sin of 2 pi c t / N2
  const double* tw1;  // [2][ks1][32]  B fragments of stage 2: cos | sin of 2 pi b y / N1
};

__host__ __device__ inline size_t pruned_smem_bytes(int N1, int N2) {
  const size_t ks2 = ((size_t)N2 / 2 + 1 + 3) / 4, ks1 = (size_t)N1 / 4;
  return (ks2 * 64 + 2 * ks1 * 32 + 16 * ((size_t)N1 + 4) + 256) * sizeof(double);
}

__device__ __forceinline__ void dmma_m8n8k4(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

// grid: persistent CTAs striding over (field, x) planes; 256 threads.
// out: [field][x][16][16] doubles:  rows 0-7 = Gr(c), 8-15 = Gs(c);  cols 0-7 = cos(b), 8-15 = sin(b)
__global__ void __launch_bounds__(256, 3)
pruned_planes_kernel(const double* __restrict__ fields, PrunedGeo g, int n_fields, double* __restrict__ out) {
  extern __shared__ double sm[];
  double* tw2 = sm;                                   // ks2 * 64
  double* tw1 = tw2 + (size_t)g.ks2 * 64;             // 2 * ks1 * 32
  double* gT = tw1 + (size_t)2 * g.ks1 * 32;          // 16 * gstride
  double* c2 = gT + (size_t)16 * g.gstride;           // 256
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < g.ks2 * 64; i += 256) tw2[i] = g.tw2[i];
  for (int i = tid; i < 2 * g.ks1 * 32; i += 256) tw1[i] = g.tw1[i];
  __syncthreads();

  const int m = lane >> 2, kk = lane & 3;  // fragment coordinates: row m, k index kk
  const int half2 = g.N2 >> 1;
  const int groups = g.N1 >> 3;            // 8-row groups per plane
  const long long planes = (long long)n_fields * g.N0;
  for (long long plane = blockIdx.x; plane < planes; plane += gridDim.x) {
    const double* base = fields + plane * ((long long)g.N1 * g.N2);
    // ---- stage 1: each warp transforms 8 rows at a time along z
    for (int grp = warp; grp < groups; grp += 8) {
      const double* row = base + (long long)(grp * 8 + m) * g.N2;
      double cr0 = 0.0, cr1 = 0.0, cs0 = 0.0, cs1 = 0.0;
#pragma unroll 4
      for (int s = 0; s < g.ks2; ++s) {
        const int t = 4 * s + kk;
        const double a1 = (t <= half2) ? __ldg(row + t) : 0.0;
        const double a2 = (t > 0 && t < half2) ? __ldg(row + (g.N2 - t)) : 0.0;
        dmma_m8n8k4(cr0, cr1, a1 + a2, tw2[s * 64 + lane]);
        dmma_m8n8k4(cs0, cs1, a1 - a2, tw2[s * 64 + 32 + lane]);
      }
      const int y = grp * 8 + m, n = 2 * kk;
      gT[(size_t)n * g.gstride + y] = cr0;
      gT[(size_t)(n + 1) * g.gstride + y] = cr1;
      gT[(size_t)(8 + n) * g.gstride + y] = cs0;
      gT[(size_t)(9 + n) * g.gstride + y] = cs1;
    }
    __syncthreads();
    // ---- stage 2: 16 x 16 x N1 product, 4 output tiles x 2 K halves over the 8 warps
    {
      const int mt = warp & 1, nt = (warp >> 1) & 1, kh = warp >> 2;
      const int s0 = kh * (g.ks1 >> 1), s1 = s0 + (g.ks1 >> 1);
      const double* arow = gT + (size_t)(8 * mt + m) * g.gstride;
      const double* brow = tw1 + (size_t)nt * g.ks1 * 32;
      double d0 = 0.0, d1 = 0.0;
#pragma unroll 4
      for (int s = s0; s < s1; ++s) dmma_m8n8k4(d0, d1, arow[4 * s + kk], brow[s * 32 + lane]);
      const int r = 8 * mt + m, c = 8 * nt + 2 * kk;
      if (kh == 1) { c2[r * 16 + c] = d0; c2[r * 16 + c + 1] = d1; }
      __syncthreads();
      if (kh == 0) {
        double2 v = make_double2(d0 + c2[r * 16 + c], d1 + c2[r * 16 + c + 1]);
        *reinterpret_cast<double2*>(out + plane * 256 + r * 16 + c) = v;
      }
    }
    __syncthreads();  // gT / c2 are rewritten by the next plane
  }
}

// Variant B: no z / N2-z pairing.  Every lane issues one aligned 16-byte load per 8 z-samples
// (half the L1 wavefronts of the paired form) and pays for it with twice the DMMA count.
// tw2f: [N2/8][4][32] B fragments: {cos, sin} x {even t, odd t} of the 8 samples of a double step.
__global__ void __launch_bounds__(256, 3)
pruned_planes_fullk_kernel(const double* __restrict__ fields, PrunedGeo g, const double* __restrict__ tw2f,
                           int n_fields, double* __restrict__ out) {
  extern __shared__ double sm[];
  const int ds = g.N2 >> 3;                           // double steps
  double* tw2 = sm;                                   // ds * 128
  double* tw1 = tw2 + (size_t)ds * 128;               // 2 * ks1 * 32
  double* gT = tw1 + (size_t)2 * g.ks1 * 32;          // 16 * gstride
  double* c2 = gT + (size_t)16 * g.gstride;           // 256
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < ds * 128; i += 256) tw2[i] = tw2f[i];
  for (int i = tid; i < 2 * g.ks1 * 32; i += 256) tw1[i] = g.tw1[i];
  __syncthreads();
  const int m = lane >> 2, kk = lane & 3;
  const int groups = g.N1 >> 3;
  const long long planes = (long long)n_fields * g.N0;
  for (long long plane = blockIdx.x; plane < planes; plane += gridDim.x) {
    const double* base = fields + plane * ((long long)g.N1 * g.N2);
    for (int grp = warp; grp < groups; grp += 8) {
      const double2* row = reinterpret_cast<const double2*>(base + (long long)(grp * 8 + m) * g.N2) + kk;
      double cr0 = 0.0, cr1 = 0.0, cs0 = 0.0, cs1 = 0.0;
#pragma unroll 8
      for (int s = 0; s < ds; ++s) {
        const double2 a = __ldg(row + 4 * s);  // samples t = 8 s + 2 kk, + 1
        const double* tw = tw2 + s * 128 + lane;
        dmma_m8n8k4(cr0, cr1, a.x, tw[0]);
        dmma_m8n8k4(cs0, cs1, a.x, tw[32]);
        dmma_m8n8k4(cr0, cr1, a.y, tw[64]);
        dmma_m8n8k4(cs0, cs1, a.y, tw[96]);
      }
      const int y = grp * 8 + m, n = 2 * kk;
      gT[(size_t)n * g.gstride + y] = cr0;
      gT[(size_t)(n + 1) * g.gstride + y] = cr1;
      gT[(size_t)(8 + n) * g.gstride + y] = cs0;
      gT[(size_t)(9 + n) * g.gstride + y] = cs1;
    }
    __syncthreads();
    {
      const int mt = warp & 1, nt = (warp >> 1) & 1, kh = warp >> 2;
      const int s0 = kh * (g.ks1 >> 1), s1 = s0 + (g.ks1 >> 1);
      const double* arow = gT + (size_t)(8 * mt + m) * g.gstride;
      const double* brow = tw1 + (size_t)nt * g.ks1 * 32;
      double d0 = 0.0, d1 = 0.0;
#pragma unroll 4
      for (int s = s0; s < s1; ++s) dmma_m8n8k4(d0, d1, arow[4 * s + kk], brow[s * 32 + lane]);
      const int r = 8 * mt + m, c = 8 * nt + 2 * kk;
      if (kh == 1) { c2[r * 16 + c] = d0; c2[r * 16 + c + 1] = d1; }
      __syncthreads();
      if (kh == 0) {
        double2 v = make_double2(d0 + c2[r * 16 + c], d1 + c2[r * 16 + c + 1]);
        *reinterpret_cast<double2*>(out + plane * 256 + r * 16 + c) = v;
      }
    }
    __syncthreads();
  }
}

__host__ __device__ inline size_t pruned_fullk_smem_bytes(int N1, int N2) {
  return ((size_t)(N2 / 8) * 128 + 2 * (size_t)(N1 / 4) * 32 + 16 * ((size_t)N1 + 4) + 256) * sizeof(double);
}

}  // namespace sbx
