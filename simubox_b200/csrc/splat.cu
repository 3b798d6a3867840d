// splat.cu -- particle -> grid density splat.
//
// Device restatement of XmlTransformer.transform (SimuBox/toolkit/xmltrans.py:117-153): every particle adds 1
// to each grid point of its type's density that lies within r_cut of it.  The reference's arithmetic is kept
// operation for operation, quirks included:
//   stencil   idx[a] = int(r_cut * 2 / h[a] + 2) points per axis                                  (:130)
//   base      int((floor((x - r_cut) / h) * h + L / 2) / h), truncation toward zero               (:136-143)
//   point     g = (offset + base) mod N (non-negative), coordinate g * h - L / 2                  (:146-147)
//   test      sqrt(((cx - x)^2 + (cy - y)^2) + (cz - z)^2) <= r_cut                               (:148-149)
// The coordinate is that of the WRAPPED grid point, not of its periodic image next to the particle, so a
// particle close to a box face does not reach across it (reference behaviour).  Counts are integers, so the
// fp64 atomic adds are exact and their order does not matter: results are bit-identical to the reference.
//
// One warp per particle: lanes stride over the stencil, (offset -> ix, iy, iz) by 32-bit division.
#include <algorithm>
#include <cstdint>

#include "common.h"

namespace sbx {
namespace {

struct SplatGeo {
  int N[3];
  int idx[3];       // stencil points per axis
  double h[3], L[3], half[3];  // spacing, box, L / 2.0
  double r_cut;
  long long P;
  int n_types;
};

__global__ void __launch_bounds__(256)
splat_kernel(const SplatGeo g, const double* __restrict__ xyz, const int32_t* __restrict__ type, double* __restrict__ phi) {
  const int lane = threadIdx.x & 31;
  const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
  const int n_st = g.idx[0] * g.idx[1] * g.idx[2];
  const long long nxyz = (long long)g.N[0] * g.N[1] * g.N[2];
  for (long long p = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; p < g.P; p += warps) {
    const int t = type[p];
    if (t < 0 || t >= g.n_types) continue;
    double x[3];
    long long base[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      x[a] = xyz[p * 3 + a];
      const double fl = floor(__ddiv_rn(__dsub_rn(x[a], g.r_cut), g.h[a]));
      const double v = __ddiv_rn(__dadd_rn(__dmul_rn(fl, g.h[a]), g.half[a]), g.h[a]);
      base[a] = (long long)v;  // np.array(..., dtype=int): toward zero
    }
    double* out = phi + (long long)t * nxyz;
    for (int s = lane; s < n_st; s += 32) {
      const int ix = s / (g.idx[1] * g.idx[2]);
      const int rem = s - ix * (g.idx[1] * g.idx[2]);
      const int iy = rem / g.idx[2], iz = rem - iy * g.idx[2];
      const int o[3] = {ix, iy, iz};
      int gi[3];
      double acc = 0.0;
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        long long m = (o[a] + base[a]) % g.N[a];
        if (m < 0) m += g.N[a];
        gi[a] = (int)m;
        const double c = __dsub_rn(__dmul_rn((double)gi[a], g.h[a]), g.half[a]);
        const double d = __dsub_rn(c, x[a]);
        acc = a == 0 ? __dmul_rn(d, d) : __dadd_rn(acc, __dmul_rn(d, d));
      }
      if (__dsqrt_rn(acc) <= g.r_cut) atomicAdd(out + ((long long)gi[0] * g.N[1] + gi[1]) * g.N[2] + gi[2], 1.0);
    }
  }
}

}  // namespace
}  // namespace sbx

using namespace sbx;

extern "C" int sbx_particle_splat(int device, const double* xyz, const int32_t* type, int64_t n_particles, int n_types,
                                  const int64_t* N, const double* L, double r_cut, double* phi, void* stream) {
  if (!xyz || !type || !N || !L || !phi) return fail(SBX_EINVAL, "NULL argument");
  if (n_particles < 0 || n_types < 1) return fail(SBX_EINVAL, "bad particle / type count");
  if (!(r_cut > 0.0)) return fail(SBX_EINVAL, "r_cut must be positive");
  int rc = require_sm100(device);
  if (rc) return rc;
  SplatGeo g{};
  long long st = 1;
  for (int a = 0; a < 3; ++a) {
    if (N[a] < 1 || N[a] > 65536 || !(L[a] > 0.0)) return fail(SBX_EINVAL, "bad grid or box on axis %d", a);
    g.N[a] = (int)N[a];
    g.L[a] = L[a];
    g.h[a] = L[a] / (double)N[a];        // lxlylz / NxNyNz (xmltrans.py:102)
    g.half[a] = L[a] / 2.0;
    g.idx[a] = (int)(r_cut * 2 / g.h[a] + 2);  // np.array(r_cut * 2 / grid_spacing + 2, dtype=int)
    st *= g.idx[a];
  }
  if (st > (1 << 24)) return fail(SBX_EINVAL, "stencil of %lld points per particle is too large", st);
  g.r_cut = r_cut;
  g.P = n_particles;
  g.n_types = n_types;
  DeviceGuard guard(device);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  SBX_CUDA(cudaMemsetAsync(phi, 0, (size_t)n_types * N[0] * N[1] * N[2] * sizeof(double), s));
  if (n_particles == 0) return SBX_OK;
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  const long long want = (n_particles + 7) / 8;  // 8 warps per CTA, one particle per warp per round
  const int grid = (int)std::min<long long>(want, (long long)sms * 8);
  splat_kernel<<<grid, 256, 0, s>>>(g, xyz, type, phi);
  SBX_CUDA(cudaGetLastError());
  return SBX_OK;
}
