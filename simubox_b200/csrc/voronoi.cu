// voronoi.cu -- grid Voronoi labelling of domain centres (sm_100a).
//
// Replaces the per-pixel Python loop of VoronoiCell.weighted_voronoi_diagrams
// (SimuBox/calculator/voronoi.py:284-292) with VoronoiCell.additive (:244-246) or
// VoronoiCell.power (:249-255) as the distance.  Three kernels:
//   1. candidates_kernel<ND,0>: per COARSE tile (4 fine tiles per axis) the seeds that can win
//      anywhere inside it, out of all S seeds;
//   2. candidates_kernel<ND,1>: per FINE tile (8x8x8 voxels, 16x32 pixels) the possible winners
//      out of the parent's list -- exact interval bounds with slack, never prunes a possible
//      winner, ascending seed index -- written as staged records;
//   3. label_tiles_kernel: stages the tile's records in shared memory with a TMA bulk copy
//      (cp.async.bulk + mbarrier) and evaluates the reference expression per voxel in fp64 with
//      round-to-nearest intrinsics (no FMA contraction); first minimum wins (np.argmin).
// Periodic boxes use the minimum image per axis, which equals the minimum over the 3^d
// replicated images of the reference expression because every fp64 step is monotone.  When
// one image is nearest for the whole tile (the usual case) the record stores fl(p + m L) and the
// per-voxel work is three subtractions and a sum of squares.
#include <cstdint>
#include <cuda_runtime.h>

#include "common.h"

namespace sbx {

constexpr int kCap = 64;          // candidate records staged per fine tile
constexpr int kCoarseCap = 1024;  // candidate indices kept per coarse tile
constexpr int kCF = 4;            // fine tiles per coarse tile and axis
constexpr int kTileVox = 512;     // voxels (= threads) per fine tile
constexpr int kFlagGeneric = 1 << 30;  // fine tile needs the generic 3-image evaluation

struct VorGeo {
  int ndim;
  int N[3];       // grid (ndim entries used)
  int T[3];       // fine tile extent per axis
  int nt[3];      // fine tiles per axis
  int ct[3];      // coarse tiles per axis
  int periodic;
  int mode;       // 0 additive, 1 power
  double box[3];
  long long S;
};

struct alignas(16) SeedRec { double x, y, z, w; };

__device__ __forceinline__ void tile_coords(const int n[3], int ndim, int tile, int t[3]) {
  for (int a = ndim - 1; a >= 0; --a) {
    t[a] = tile % n[a];
    tile /= n[a];
  }
}

// conservative per-axis distance interval between coordinate p and voxel range [lo, hi]
__device__ __forceinline__ void axis_bounds(double p, double lo, double hi, double& dmin, double& dmax) {
  dmin = fmax(0.0, fmax(lo - p, p - hi));
  dmax = fmax(fabs(p - lo), fabs(p - hi));
}

// bounds of the seed's value over the voxel box [lo, hi]; for periodic additive tiles also
// reports per axis which image is nearest everywhere in the box (img[a] in -1..1), or sets
// `ambiguous` when no single image is.
template <int ND>
__device__ __forceinline__ void seed_bounds(const VorGeo& g, const double* __restrict__ seeds,
                                            const double* __restrict__ weights, long long s, const int lo[3],
                                            const int hi[3], double& LB, double& UB, int img[3], bool& ambiguous) {
  double smin = 0.0, smax = 0.0;
  ambiguous = false;
#pragma unroll
  for (int a = 0; a < ND; ++a) {
    double p = seeds[s * ND + a];
    if (g.mode == 1) p += (double)s;  // voronoi.py:251 adds the seed index to every coordinate
    double dmin, dmax;
    axis_bounds(p, (double)lo[a], (double)hi[a], dmin, dmax);
    img[a] = 0;
    if (g.periodic) {
      double n0, n1, p0, p1;
      axis_bounds(p - g.box[a], (double)lo[a], (double)hi[a], n0, n1);
      axis_bounds(p + g.box[a], (double)lo[a], (double)hi[a], p0, p1);
      // nearest image = smallest worst-case distance; unambiguous if the others are farther everywhere
      double bmax = dmax, o1 = n0, o2 = p0;
      if (n1 < bmax) { bmax = n1; img[a] = -1; o1 = dmin; o2 = p0; }
      if (p1 < bmax) { bmax = p1; img[a] = 1; o1 = dmin; o2 = n0; }
      const double margin = 1e-9 * (bmax + 1.0);
      if (!(fmin(o1, o2) > bmax + margin)) ambiguous = true;
      dmin = fmin(dmin, fmin(n0, p0));
      dmax = bmax;
    }
    smin += dmin * dmin;
    smax += dmax * dmax;
  }
  const double w = weights ? weights[s] : 0.0;
  if (g.mode == 0) { LB = sqrt(smin) - w; UB = sqrt(smax) - w; }
  else { LB = smin - w; UB = smax - w; }
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

// ------------------------------------------------------------------------------------------
// Candidate selection.  LEVEL 0: one block per coarse tile, source = all seeds, output = seed
// indices.  LEVEL 1: one block per fine tile, source = the parent's list (or all seeds if the
// parent overflowed), output = staged records + indices.  A seed is kept iff its lower bound
// over the tile does not exceed the smallest upper bound of any seed (with slack), so a seed
// that wins or ties anywhere in the tile is never dropped.  Output order = ascending seed index.
// ------------------------------------------------------------------------------------------
template <int ND, int LEVEL>
__global__ void __launch_bounds__(LEVEL == 0 ? 256 : 128)
candidates_kernel(VorGeo g, const double* __restrict__ seeds, const double* __restrict__ weights,
                  int* __restrict__ clist, int* __restrict__ ccount,           // coarse lists (out for LEVEL 0, in for 1)
                  SeedRec* __restrict__ rec, int* __restrict__ ridx, int* __restrict__ rcount) {
  constexpr int NW = (LEVEL == 0 ? 256 : 128) / 32;
  const int tile = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0}, t[3] = {0, 0, 0};
  const int* src = nullptr;
  long long n_src = g.S;
  if (LEVEL == 0) {
    tile_coords(g.ct, ND, tile, t);
    for (int a = 0; a < ND; ++a) {
      lo[a] = t[a] * g.T[a] * kCF;
      hi[a] = min(lo[a] + g.T[a] * kCF, g.N[a]) - 1;
    }
  } else {
    tile_coords(g.nt, ND, tile, t);
    int parent = 0;
    for (int a = 0; a < ND; ++a) {
      lo[a] = t[a] * g.T[a];
      hi[a] = min(lo[a] + g.T[a], g.N[a]) - 1;
      parent = parent * g.ct[a] + t[a] / kCF;
    }
    const int pc = ccount[parent];
    if (pc <= kCoarseCap) {
      src = clist + (long long)parent * kCoarseCap;
      n_src = pc;
    }
  }

  __shared__ double s_red[NW];
  __shared__ double s_U;
  __shared__ int s_warp[NW];
  __shared__ int s_total;
  __shared__ int s_flag;
  int img[3];
  bool amb;
  double best = INFINITY;
  for (long long i = tid; i < n_src; i += blockDim.x) {
    const long long s = src ? src[i] : i;
    double LB, UB;
    seed_bounds<ND>(g, seeds, weights, s, lo, hi, LB, UB, img, amb);
    best = fmin(best, UB);
  }
  for (int o = 16; o > 0; o >>= 1) best = fmin(best, __shfl_xor_sync(0xffffffffu, best, o));
  if (lane == 0) s_red[warp] = best;
  if (tid == 0) { s_total = 0; s_flag = 0; }
  __syncthreads();
  if (tid == 0) {
    double u = s_red[0];
    for (int w = 1; w < NW; ++w) u = fmin(u, s_red[w]);
    s_U = u;
  }
  __syncthreads();
  const double U = s_U;
  const int cap = LEVEL == 0 ? kCoarseCap : kCap;

  for (long long base = 0; base < n_src; base += blockDim.x) {
    const long long i = base + tid;
    bool keep = false;
    long long s = 0;
    if (i < n_src) {
      s = src ? src[i] : i;
      double LB, UB;
      seed_bounds<ND>(g, seeds, weights, s, lo, hi, LB, UB, img, amb);
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
      if (slot < cap) {
        if (LEVEL == 0) {
          clist[(long long)tile * kCoarseCap + slot] = (int)s;
        } else {
          SeedRec r;
          double p[3] = {seeds[s * ND], seeds[s * ND + 1], ND == 3 ? seeds[s * ND + 2] : 0.0};
          if (g.periodic && g.mode == 0) {
            if (amb) s_flag = 1;  // benign race: every writer stores 1
#pragma unroll
            for (int a = 0; a < ND; ++a)  // fl(p + m L): the image coordinate the reference evaluates
              p[a] = img[a] < 0 ? __dsub_rn(p[a], g.box[a]) : (img[a] > 0 ? __dadd_rn(p[a], g.box[a]) : p[a]);
          }
          r.x = p[0]; r.y = p[1]; r.z = p[2];
          r.w = weights ? weights[s] : 0.0;
          rec[(long long)tile * kCap + slot] = r;
          ridx[(long long)tile * kCap + slot] = (int)s;
        }
      }
    }
    __syncthreads();
    if (tid == 0) {
      int add = 0;
      for (int w = 0; w < NW; ++w) add += s_warp[w];
      s_total += add;
    }
    __syncthreads();
  }
  if (tid == 0) {
    if (LEVEL == 0) ccount[tile] = s_total;
    else rcount[tile] = s_total | ((s_flag || g.mode == 1) ? kFlagGeneric : 0);
  }
}

// the reference expression for one seed at one voxel, all three images (voronoi.py:245 / :250-254)
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

// sum of squares of (p - x), same operation order as the reference, image already resolved
template <int ND>
__device__ __forceinline__ double sumsq(const SeedRec& r, const double x[3]) {
  const double dx = __dsub_rn(r.x, x[0]), dy = __dsub_rn(r.y, x[1]);
  double ss = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
  if (ND == 3) {
    const double dz = __dsub_rn(r.z, x[2]);
    ss = __dadd_rn(ss, __dmul_rn(dz, dz));
  }
  return ss;
}

// MODE 0: generic evaluation only (power mode).  MODE 1: additive, no weights -- compare sums of
// squares and take the square root only inside a 2^-50 relative window, where rounding of the
// root could turn "smaller" into "equal" (then the lower index keeps the voxel, like argmin).
// MODE 2: additive with weights -- sqrt(ss) - w per candidate.
template <int ND, int MODE>
__global__ void __launch_bounds__(kTileVox)
label_tiles_kernel(VorGeo g, const double* __restrict__ seeds, const double* __restrict__ weights,
                   const SeedRec* __restrict__ rec, const int* __restrict__ ridx,
                   const int* __restrict__ rcount, const uint8_t* __restrict__ mask,
                   int32_t* __restrict__ labels, unsigned long long* __restrict__ stats) {
  __shared__ alignas(128) SeedRec s_rec[kCap];
  __shared__ alignas(16) int s_idx[kCap];
  __shared__ alignas(8) uint64_t s_bar;
  const int tile = blockIdx.x, tid = threadIdx.x;
  const int raw = rcount[tile];
  const int count = raw & (kFlagGeneric - 1);
  const bool generic = MODE == 0 || (raw & kFlagGeneric) != 0;
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

  int t[3] = {0, 0, 0}, v[3] = {0, 0, 0};
  tile_coords(g.nt, ND, tile, t);
  {
    int r = tid;
    for (int a = ND - 1; a >= 0; --a) {
      v[a] = t[a] * g.T[a] + r % g.T[a];
      r /= g.T[a];
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
    if (!staged) {  // more possible winners than the staging buffer holds: scan every seed
      double best = INFINITY;
      for (long long s = 0; s < g.S; ++s) {
        const double p[3] = {seeds[s * ND], seeds[s * ND + 1], ND == 3 ? seeds[s * ND + 2] : 0.0};
        const double val = seed_value<ND>(g, p, weights ? weights[s] : 0.0, (double)s, x);
        if (val < best || best_idx < 0) { best = val; best_idx = (int)s; }
      }
    } else if (generic) {  // records hold the original coordinates: evaluate all three images
      double best = INFINITY;
      for (int c = 0; c < count; ++c) {
        const int si = s_idx[c];
        const double p[3] = {seeds[(long long)si * ND], seeds[(long long)si * ND + 1], ND == 3 ? seeds[(long long)si * ND + 2] : 0.0};
        const double val = seed_value<ND>(g, p, s_rec[c].w, (double)si, x);
        if (val < best || best_idx < 0) { best = val; best_idx = si; }
      }
    } else if (MODE == 1) {
      double best_ss = INFINITY, best_lo = INFINITY;
#pragma unroll 2
      for (int c = 0; c < count; ++c) {
        const double ss = sumsq<ND>(s_rec[c], x);
        bool take = ss < best_lo;
        if (!take && ss < best_ss) take = __dsqrt_rn(ss) < __dsqrt_rn(best_ss);  // rare: roots may round together
        if (take || best_idx < 0) {
          best_ss = ss;
          best_lo = __dmul_rn(ss, 1.0 - 0x1p-50);
          best_idx = s_idx[c];
        }
      }
    } else {
      double best = INFINITY;
#pragma unroll 2
      for (int c = 0; c < count; ++c) {
        const SeedRec r = s_rec[c];
        const double val = __dsub_rn(__dsqrt_rn(sumsq<ND>(r, x)), r.w);
        if (val < best || best_idx < 0) { best = val; best_idx = s_idx[c]; }
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

template <int ND>
static void launch_voronoi(const VorGeo& g, long long ctiles, long long tiles, const double* seeds, const double* weights,
                           int* clist, int* ccount, SeedRec* rec, int* ridx, int* rcount, const uint8_t* mask,
                           int32_t* labels, unsigned long long* stats, cudaStream_t st) {
  candidates_kernel<ND, 0><<<(unsigned)ctiles, 256, 0, st>>>(g, seeds, weights, clist, ccount, nullptr, nullptr, nullptr);
  candidates_kernel<ND, 1><<<(unsigned)tiles, 128, 0, st>>>(g, seeds, weights, clist, ccount, rec, ridx, rcount);
  if (g.mode == 1)
    label_tiles_kernel<ND, 0><<<(unsigned)tiles, kTileVox, 0, st>>>(g, seeds, weights, rec, ridx, rcount, mask, labels, stats);
  else if (!weights)
    label_tiles_kernel<ND, 1><<<(unsigned)tiles, kTileVox, 0, st>>>(g, seeds, weights, rec, ridx, rcount, mask, labels, stats);
  else
    label_tiles_kernel<ND, 2><<<(unsigned)tiles, kTileVox, 0, st>>>(g, seeds, weights, rec, ridx, rcount, mask, labels, stats);
}

extern "C" int sbx_voronoi_label(int device, int ndim, const int64_t* N, const double* seeds, int64_t S,
                                 const double* weights, int mode, const double* box_host, const uint8_t* mask,
                                 int32_t* labels, void* stream, int64_t* stats_host) {
  if (!N || !seeds || !labels) return fail(SBX_EINVAL, "NULL argument");
  if (ndim != 2 && ndim != 3) return fail(SBX_EINVAL, "ndim must be 2 or 3, got %d", ndim);
  if (S < 1 || S > 0x3fffffff) return fail(SBX_EINVAL, "seed count %lld out of range", (long long)S);
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
  long long tiles = 1, ctiles = 1;
  for (int a = 0; a < 3; ++a) {
    g.N[a] = 1; g.T[a] = 1; g.nt[a] = 1; g.ct[a] = 1; g.box[a] = 0.0;
  }
  for (int a = 0; a < ndim; ++a) {
    if (N[a] < 1 || N[a] > (1 << 20)) return fail(SBX_EINVAL, "N[%d]=%lld out of range", a, (long long)N[a]);
    g.N[a] = (int)N[a];
    g.T[a] = ndim == 3 ? tile3[a] : tile2[a];
    g.nt[a] = (g.N[a] + g.T[a] - 1) / g.T[a];
    g.ct[a] = (g.nt[a] + kCF - 1) / kCF;
    tiles *= g.nt[a];
    ctiles *= g.ct[a];
    if (box_host) {
      if (!(box_host[a] > 0.0)) return fail(SBX_EINVAL, "box[%d] must be > 0", a);
      g.box[a] = box_host[a];
    }
  }
  if (tiles > 0x7fffffff) return fail(SBX_EINVAL, "grid too large (%lld tiles)", tiles);

  SeedRec* rec = nullptr;
  int *ridx = nullptr, *rcount = nullptr, *clist = nullptr, *ccount = nullptr;
  unsigned long long* stats = nullptr;
  SBX_CUDA(cudaMallocAsync((void**)&rec, (size_t)tiles * kCap * sizeof(SeedRec), st));
  SBX_CUDA(cudaMallocAsync((void**)&ridx, (size_t)tiles * kCap * sizeof(int), st));
  SBX_CUDA(cudaMallocAsync((void**)&rcount, (size_t)tiles * sizeof(int), st));
  SBX_CUDA(cudaMallocAsync((void**)&clist, (size_t)ctiles * kCoarseCap * sizeof(int), st));
  SBX_CUDA(cudaMallocAsync((void**)&ccount, (size_t)ctiles * sizeof(int), st));
  if (stats_host) {
    SBX_CUDA(cudaMallocAsync((void**)&stats, 3 * sizeof(unsigned long long), st));
    SBX_CUDA(cudaMemsetAsync(stats, 0, 3 * sizeof(unsigned long long), st));
  }
  if (ndim == 3) launch_voronoi<3>(g, ctiles, tiles, seeds, weights, clist, ccount, rec, ridx, rcount, mask, labels, stats, st);
  else launch_voronoi<2>(g, ctiles, tiles, seeds, weights, clist, ccount, rec, ridx, rcount, mask, labels, stats, st);
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
  SBX_CUDA(cudaFreeAsync(clist, st));
  SBX_CUDA(cudaFreeAsync(ccount, st));
  return SBX_OK;
}
