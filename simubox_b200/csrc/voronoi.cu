// voronoi.cu -- grid Voronoi labelling of domain centres (sm_100a).
//
// Replaces the per-pixel Python loop of VoronoiCell.weighted_voronoi_diagrams
// (SimuBox/calculator/voronoi.py:284-292) with VoronoiCell.additive (:244-246) or
// VoronoiCell.power (:249-255) as the distance.  Two kernels:
//   1. tile_candidates_kernel: per voxel tile, the seeds that can win anywhere in the tile
//      (exact interval bounds with slack; never prunes a possible winner), ascending index.
//   2. label_tiles_kernel: stages the tile's candidate records in shared memory with a TMA
//      bulk copy (cp.async.bulk + mbarrier) and evaluates the reference expression in fp64
//      with round-to-nearest intrinsics (no FMA contraction), first minimum wins (np.argmin).
// Periodic boxes use the minimum image per axis, which equals the minimum over the 3^d
// replicated images of the reference expression because every fp64 step is monotone.
#include <cstdint>
#include <cuda_runtime.h>

#include "common.h"

namespace sbx {

constexpr int kCap = 64;        // candidate records staged per tile
constexpr int kTileVox = 512;   // voxels (= threads) per tile

struct VorGeo {
  int ndim;
  int N[3];       // grid (ndim entries used)
  int T[3];       // tile extent per axis
  int nt[3];      // tiles per axis
  int periodic;
  int mode;       // 0 additive, 1 power
  double box[3];
  long long S;
};

struct alignas(16) SeedRec { double x, y, z, w; };

__device__ __forceinline__ void tile_origin(const VorGeo& g, int tile, int lo[3]) {
  int t = tile;
  for (int a = g.ndim - 1; a >= 0; --a) {
    lo[a] = (t % g.nt[a]) * g.T[a];
    t /= g.nt[a];
  }
}

// conservative per-axis distance interval between coordinate p and voxel range [lo, hi]
__device__ __forceinline__ void axis_bounds(double p, double lo, double hi, double& dmin, double& dmax) {
  dmin = fmax(0.0, fmax(lo - p, p - hi));
  dmax = fmax(fabs(p - lo), fabs(p - hi));
}

template <int ND>
__device__ __forceinline__ void seed_bounds(const VorGeo& g, const double* __restrict__ seeds,
                                            const double* __restrict__ weights, long long s, const int lo[3],
                                            const int hi[3], double& LB, double& UB) {
  double smin = 0.0, smax = 0.0;
#pragma unroll
  for (int a = 0; a < ND; ++a) {
    double p = seeds[s * ND + a];
    if (g.mode == 1) p += (double)s;  // voronoi.py:251 adds the seed index to every coordinate
    double dmin, dmax;
    axis_bounds(p, (double)lo[a], (double)hi[a], dmin, dmax);
    if (g.periodic) {
      double d0, d1;
      axis_bounds(p - g.box[a], (double)lo[a], (double)hi[a], d0, d1);
      dmin = fmin(dmin, d0); dmax = fmin(dmax, d1);
      axis_bounds(p + g.box[a], (double)lo[a], (double)hi[a], d0, d1);
      dmin = fmin(dmin, d0); dmax = fmin(dmax, d1);
    }
    smin += dmin * dmin;
    smax += dmax * dmax;
  }
  const double w = weights ? weights[s] : 0.0;
  if (g.mode == 0) { LB = sqrt(smin) - w; UB = sqrt(smax) - w; }
  else { LB = smin - w; UB = smax - w; }
}

template <int ND>
__global__ void __launch_bounds__(256)
tile_candidates_kernel(VorGeo g, const double* __restrict__ seeds, const double* __restrict__ weights,
                       SeedRec* __restrict__ rec, int* __restrict__ ridx, int* __restrict__ rcount) {
  const int tile = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
  tile_origin(g, tile, lo);
  for (int a = 0; a < ND; ++a) hi[a] = min(lo[a] + g.T[a], g.N[a]) - 1;

  __shared__ double s_red[8];
  __shared__ double s_U;
  __shared__ int s_warp[8];
  __shared__ int s_total;
  double best = INFINITY;
  for (long long s = tid; s < g.S; s += blockDim.x) {
    double LB, UB;
    seed_bounds<ND>(g, seeds, weights, s, lo, hi, LB, UB);
    best = fmin(best, UB);
  }
  for (int o = 16; o > 0; o >>= 1) best = fmin(best, __shfl_xor_sync(0xffffffffu, best, o));
  if (lane == 0) s_red[warp] = best;
  if (tid == 0) s_total = 0;
  __syncthreads();
  if (tid == 0) {
    double u = s_red[0];
    for (int w = 1; w < 8; ++w) u = fmin(u, s_red[w]);
    s_U = u;
  }
  __syncthreads();
  const double U = s_U;

  SeedRec* trec = rec + (long long)tile * kCap;
  int* tidx = ridx + (long long)tile * kCap;
  for (long long base = 0; base < g.S; base += blockDim.x) {
    const long long s = base + tid;
    bool keep = false;
    if (s < g.S) {
      double LB, UB;
      seed_bounds<ND>(g, seeds, weights, s, lo, hi, LB, UB);
      const double slack = 1e-10 * (fabs(U) + fabs(LB) + 1.0);
      keep = (LB - slack <= U + slack);  // never drops a seed that could tie or win
    }
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    int before = s_total;
    for (int w = 0; w < warp; ++w) before += s_warp[w];
    if (keep) {
      const int slot = before + __popc(bal & ((1u << lane) - 1u));
      if (slot < kCap) {
        SeedRec r;
        r.x = seeds[s * ND + 0];
        r.y = seeds[s * ND + 1];
        r.z = ND == 3 ? seeds[s * ND + 2] : 0.0;
        r.w = weights ? weights[s] : 0.0;
        trec[slot] = r;
        tidx[slot] = (int)s;
      }
    }
    __syncthreads();
    if (tid == 0) {
      int add = 0;
      for (int w = 0; w < 8; ++w) add += s_warp[w];
      s_total += add;
    }
    __syncthreads();
  }
  if (tid == 0) rcount[tile] = s_total;
}

// ---- mbarrier / TMA bulk copy helpers (PTX; SASS: SYNCS.* / UBLKCP) --------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// the reference expression for one seed at one voxel (voronoi.py:245 / :250-254)
template <int ND>
__device__ __forceinline__ double seed_value(const VorGeo& g, const double p[3], double w, double shift,
                                             const double x[3]) {
  double ss = 0.0;
#pragma unroll
  for (int a = 0; a < ND; ++a) {
    double d = __dsub_rn(p[a], x[a]);
    if (g.mode == 1) d = __dadd_rn(d, shift);
    if (g.periodic) {
      double dm = __dsub_rn(__dsub_rn(p[a], g.box[a]), x[a]);
      double dp = __dsub_rn(__dadd_rn(p[a], g.box[a]), x[a]);
      if (g.mode == 1) { dm = __dadd_rn(dm, shift); dp = __dadd_rn(dp, shift); }
      d = fmin(fabs(d), fmin(fabs(dm), fabs(dp)));
    }
    const double sq = __dmul_rn(d, d);
    ss = a == 0 ? sq : __dadd_rn(ss, sq);  // np.linalg.norm: sequential sum x -> y -> z
  }
  const double r = __dsqrt_rn(ss);
  return __dsub_rn(g.mode == 1 ? __dmul_rn(r, r) : r, w);
}

template <int ND>
__global__ void __launch_bounds__(kTileVox)
label_tiles_kernel(VorGeo g, const double* __restrict__ seeds, const double* __restrict__ weights,
                   const SeedRec* __restrict__ rec, const int* __restrict__ ridx,
                   const int* __restrict__ rcount, const uint8_t* __restrict__ mask,
                   int32_t* __restrict__ labels, unsigned long long* __restrict__ stats) {
  __shared__ alignas(128) SeedRec s_rec[kCap];
  __shared__ alignas(16) int s_idx[kCap];
  __shared__ alignas(8) uint64_t s_bar;
  const int tile = blockIdx.x, tid = threadIdx.x;
  const int count = rcount[tile];
  const bool staged = count <= kCap;
  if (tid == 0) mbar_init(&s_bar, 1);
  __syncthreads();
  if (staged && tid == 0) {
    const uint32_t b_rec = (uint32_t)count * (uint32_t)sizeof(SeedRec);
    const uint32_t b_idx = (uint32_t)((count + 3) & ~3) * 4u;
    mbar_expect_tx(&s_bar, b_rec + b_idx);
    tma_bulk_g2s(s_rec, rec + (long long)tile * kCap, b_rec, &s_bar);
    tma_bulk_g2s(s_idx, ridx + (long long)tile * kCap, b_idx, &s_bar);
  }

  int lo[3] = {0, 0, 0};
  tile_origin(g, tile, lo);
  int v[3] = {0, 0, 0};
  {
    int t = tid;
    for (int a = ND - 1; a >= 0; --a) {
      v[a] = lo[a] + t % g.T[a];
      t /= g.T[a];
    }
  }
  bool inside = true;
  long long lin = 0;
  for (int a = 0; a < ND; ++a) {
    inside = inside && v[a] < g.N[a];
    lin = lin * g.N[a] + v[a];
  }
  const bool live = inside && (!mask || mask[lin] != 0);
  const double x[3] = {(double)v[0], (double)v[1], (double)v[2]};

  if (staged) mbar_wait(&s_bar, 0);
  int best_idx = -1;
  if (live) {
    double best = INFINITY;
    if (staged) {
      for (int c = 0; c < count; ++c) {
        const SeedRec r = s_rec[c];
        const int si = s_idx[c];
        const double p[3] = {r.x, r.y, r.z};
        const double val = seed_value<ND>(g, p, r.w, (double)si, x);
        if (val < best || best_idx < 0) { best = val; best_idx = si; }
      }
    } else {  // more possible winners than the staging buffer holds: scan every seed
      for (long long s = 0; s < g.S; ++s) {
        const double p[3] = {seeds[s * ND], seeds[s * ND + 1], ND == 3 ? seeds[s * ND + 2] : 0.0};
        const double val = seed_value<ND>(g, p, weights ? weights[s] : 0.0, (double)s, x);
        if (val < best || best_idx < 0) { best = val; best_idx = (int)s; }
      }
    }
  }
  if (inside) labels[lin] = best_idx;
  if (stats) {
    const unsigned long long evals = live ? (unsigned long long)(staged ? count : g.S) : 0ull;
    unsigned long long tot = evals;
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    if ((tid & 31) == 0 && tot) atomicAdd(&stats[0], tot);
    if (tid == 0) {
      atomicAdd(&stats[1], 1ull);
      if (!staged) atomicAdd(&stats[2], 1ull);
    }
  }
}

}  // namespace sbx

using namespace sbx;

extern "C" int sbx_voronoi_label(int device, int ndim, const int64_t* N, const double* seeds, int64_t S,
                                 const double* weights, int mode, const double* box_host, const uint8_t* mask,
                                 int32_t* labels, void* stream, int64_t* stats_host) {
  if (!N || !seeds || !labels) return fail(SBX_EINVAL, "NULL argument");
  if (ndim != 2 && ndim != 3) return fail(SBX_EINVAL, "ndim must be 2 or 3, got %d", ndim);
  if (S < 1 || S > 0x7fffffff) return fail(SBX_EINVAL, "seed count %lld out of range", (long long)S);
  if (mode != 0 && mode != 1) return fail(SBX_EINVAL, "mode must be 0 (additive) or 1 (power)");
  int rc = require_sm100(device);
  if (rc) return rc;
  DeviceGuard guard(device);
  cudaStream_t st = (cudaStream_t)stream;

  VorGeo g{};
  g.ndim = ndim;
  g.mode = mode;
  g.S = S;
  g.periodic = box_host ? 1 : 0;
  const int tile3[3] = {8, 8, 8}, tile2[3] = {16, 32, 1};
  long long tiles = 1, voxels = 1;
  for (int a = 0; a < 3; ++a) {
    g.N[a] = 1; g.T[a] = 1; g.nt[a] = 1; g.box[a] = 0.0;
  }
  for (int a = 0; a < ndim; ++a) {
    if (N[a] < 1 || N[a] > (1 << 20)) return fail(SBX_EINVAL, "N[%d]=%lld out of range", a, (long long)N[a]);
    g.N[a] = (int)N[a];
    g.T[a] = ndim == 3 ? tile3[a] : tile2[a];
    g.nt[a] = (g.N[a] + g.T[a] - 1) / g.T[a];
    tiles *= g.nt[a];
    voxels *= g.N[a];
    if (box_host) {
      if (!(box_host[a] > 0.0)) return fail(SBX_EINVAL, "box[%d] must be > 0", a);
      g.box[a] = box_host[a];
    }
  }
  if (tiles > 0x7fffffff) return fail(SBX_EINVAL, "grid too large (%lld tiles)", tiles);

  SeedRec* rec = nullptr;
  int* ridx = nullptr;
  int* rcount = nullptr;
  unsigned long long* stats = nullptr;
  SBX_CUDA(cudaMallocAsync((void**)&rec, (size_t)tiles * kCap * sizeof(SeedRec), st));
  SBX_CUDA(cudaMallocAsync((void**)&ridx, (size_t)tiles * kCap * sizeof(int), st));
  SBX_CUDA(cudaMallocAsync((void**)&rcount, (size_t)tiles * sizeof(int), st));
  if (stats_host) {
    SBX_CUDA(cudaMallocAsync((void**)&stats, 3 * sizeof(unsigned long long), st));
    SBX_CUDA(cudaMemsetAsync(stats, 0, 3 * sizeof(unsigned long long), st));
  }
  if (ndim == 3) {
    tile_candidates_kernel<3><<<(unsigned)tiles, 256, 0, st>>>(g, seeds, weights, rec, ridx, rcount);
    label_tiles_kernel<3><<<(unsigned)tiles, kTileVox, 0, st>>>(g, seeds, weights, rec, ridx, rcount, mask, labels, stats);
  } else {
    tile_candidates_kernel<2><<<(unsigned)tiles, 256, 0, st>>>(g, seeds, weights, rec, ridx, rcount);
    label_tiles_kernel<2><<<(unsigned)tiles, kTileVox, 0, st>>>(g, seeds, weights, rec, ridx, rcount, mask, labels, stats);
  }
  SBX_CUDA(cudaGetLastError());
  if (stats_host) {
    unsigned long long h[3];
    SBX_CUDA(cudaMemcpyAsync(h, stats, sizeof(h), cudaMemcpyDeviceToHost, st));
    SBX_CUDA(cudaStreamSynchronize(st));
    for (int i = 0; i < 3; ++i) stats_host[i] = (int64_t)h[i];
    SBX_CUDA(cudaFreeAsync(stats, st));
  }
  SBX_CUDA(cudaFreeAsync(rec, st));
  SBX_CUDA(cudaFreeAsync(ridx, st));
  SBX_CUDA(cudaFreeAsync(rcount, st));
  return SBX_OK;
}
