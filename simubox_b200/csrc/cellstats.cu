// cellstats.cu -- per-cell statistics on a Voronoi label grid.
//
// What the reference derives downstream of its Voronoi / Delaunay constructions
// (SimuBox/calculator/voronoi.py:150-213: coord_dict = neighbours of every centre; demos/scripts.ipynb
// cells 1-2: in-cell density ratios with a Python point-in-polygon loop), computed directly from the
// label array the labelling kernel writes:
//   volume[s]    = number of voxels with label s                         (exact, int64)
//   mass[s]      = sum of field over those voxels                        (fp64, atomics: order not fixed)
//   adjacency    = symmetric S x S bit matrix: s ~ t iff two face-adjacent voxels (periodic wrap
//                  optional) carry the labels s and t;  degree[s] = number of neighbours of s
// One pass over the labels (4 B/voxel) and the field (8 B/voxel); label < 0 (masked voxel) is skipped.
//
// Each thread owns a run of 8 consecutive voxels of one z-row; a warp covers 16 rows x 2 runs, a compact
// 16 x 16 patch that meets only a handful of cells.  Labels are spatially coherent, so a
// thread folds its run into a head and a tail partial (anything in between goes straight to global
// atomics), and the warp then reduces by key (match.any + shuffle tree per distinct label), leaving
// ~2 global atomics per 256 voxels.  Adjacency bits are tested with a plain load before the atomicOr,
// so after the first few hits per pair the traffic is read-only and L2-resident (S = 4096: 2 MB).
#include <algorithm>
#include <cstdint>

#include "common.h"

namespace sbx {
namespace {

constexpr int kRun = 8;
constexpr unsigned kFull = 0xffffffffu;

struct StatGeo {
  int N0, N1, N2;      // 2D grids arrive as 1 x X x Y
  int runs_per_row;    // ceil(N2 / 8)
  int periodic;
  int tiles_y, tiles_z;  // a warp covers 16 rows x 2 runs (16 x 16 voxels): few distinct labels per warp
  long long n_runs;    // padded: N0 * tiles_y * tiles_z * 32
  int words;           // adjacency row length in 32-bit words
  long long n_cells;
  int vec;             // 1: N2 % 8 == 0 and 16-byte aligned bases -> vector loads
};

__device__ __forceinline__ void load_labels(const int32_t* __restrict__ p, long long off, int n, int vec, int (&out)[kRun]) {
  if (vec) {
    const int4 a = __ldg(reinterpret_cast<const int4*>(p + off));
    const int4 b = __ldg(reinterpret_cast<const int4*>(p + off) + 1);
    out[0] = a.x; out[1] = a.y; out[2] = a.z; out[3] = a.w;
    out[4] = b.x; out[5] = b.y; out[6] = b.z; out[7] = b.w;
  } else {
#pragma unroll
    for (int k = 0; k < kRun; ++k) out[k] = k < n ? __ldg(p + off + k) : -1;
  }
}

__device__ __forceinline__ void set_pair(uint32_t* adj, int words, int a, int b) {
  uint32_t* w = adj + (size_t)a * words + (b >> 5);
  const uint32_t bit = 1u << (b & 31);
  // bits only ever go 0 -> 1: a stale L1 line can at worst cause a redundant atomicOr
  if (!(*w & bit)) atomicOr(w, bit);
}

// (pa, pb): the pair this thread linked last -- neighbouring voxels mostly repeat it
__device__ __forceinline__ void link(uint32_t* adj, const StatGeo& g, int a, int b, int& pa, int& pb) {
  if (a != b && a >= 0 && b >= 0 && a < g.n_cells && b < g.n_cells && !(a == pa && b == pb)) {
    set_pair(adj, g.words, a, b);
    set_pair(adj, g.words, b, a);
    pa = a;
    pb = b;
  }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(kFull, v, off);
  return v;
}

// every lane holds (key, count, sum); one lane per distinct key >= 0 adds the warp's total to global memory
__device__ __forceinline__ void warp_reduce_by_key(int key, int cnt, double sum, const StatGeo& g, unsigned long long* volume,
                                                   double* mass, int lane) {
  const unsigned grp = __match_any_sync(kFull, key);
  unsigned todo = kFull;
  while (todo) {  // uniform loop: one round per distinct key in the warp
    const int leader = __ffs(todo) - 1;
    const unsigned gm = __shfl_sync(kFull, grp, leader);
    const bool mine = grp == gm;
    const int c = __reduce_add_sync(kFull, mine ? cnt : 0);
    const double s = mass ? warp_sum(mine ? sum : 0.0) : 0.0;
    if (lane == leader && key >= 0 && key < g.n_cells && c > 0) {
      atomicAdd(volume + key, (unsigned long long)c);
      if (mass) atomicAdd(mass + key, s);
    }
    todo &= ~gm;
  }
}

__global__ void __launch_bounds__(256)
cell_stats_kernel(const StatGeo g, const int32_t* __restrict__ labels, const double* __restrict__ field,
                  unsigned long long* __restrict__ volume, double* __restrict__ mass, uint32_t* adj) {
  const int lane = threadIdx.x & 31;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long plane = (long long)g.N1 * g.N2;
  // whole warps iterate together (the reduce-by-key needs all 32 lanes), idle lanes carry key -1
  const long long n_iter = (g.n_runs + stride - 1) / stride;
  for (long long it = 0; it < n_iter; ++it) {
    const long long r = it * stride + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int hk = -1, hc = 0, tk = -1, tc = 0;
    double hs = 0.0, ts = 0.0;
    // warp-tile w = (x, ty, tz); lane = (ly, lz): row y = 16 ty + ly, run 2 tz + lz
    const unsigned w = (unsigned)(r >> 5);  // < 2^31 warp tiles (checked on the host): 32-bit divisions
    const unsigned wy = w / (unsigned)g.tiles_z;
    const int tz = (int)(w - wy * (unsigned)g.tiles_z);
    const unsigned xq = wy / (unsigned)g.tiles_y;
    const int ty = (int)(wy - xq * (unsigned)g.tiles_y);
    const int yq = ty * 16 + (lane >> 1), run = tz * 2 + (lane & 1);
    if (r < g.n_runs && yq < g.N1 && run < g.runs_per_row) {
      const long long row = (long long)xq * g.N1 + yq;
      const int z0 = run * kRun;
      const int n = min(kRun, g.N2 - z0);
      const long long off = row * g.N2 + z0;
      int l[kRun];
      load_labels(labels, off, n, g.vec, l);
      double f[kRun];
      if (field) {
        if (g.vec) {
#pragma unroll
          for (int k = 0; k < kRun; k += 2) {
            const double2 v = __ldg(reinterpret_cast<const double2*>(field + off + k));
            f[k] = v.x; f[k + 1] = v.y;
          }
        } else {
#pragma unroll
          for (int k = 0; k < kRun; ++k) f[k] = k < n ? __ldg(field + off + k) : 0.0;
        }
      } else {
#pragma unroll
        for (int k = 0; k < kRun; ++k) f[k] = 0.0;
      }
      // ---- fold the run: head partial, tail partial, anything in between straight to global memory
      int ck = l[0], cc = 0, seg = 0;
      double cs = 0.0;
#pragma unroll
      for (int k = 0; k < kRun; ++k) {
        if (k < n) {
          if (l[k] != ck) {
            if (seg == 0) { hk = ck; hc = cc; hs = cs; }
            else if (ck >= 0 && ck < g.n_cells) {
              atomicAdd(volume + ck, (unsigned long long)cc);
              if (mass) atomicAdd(mass + ck, cs);
            }
            ++seg;
            ck = l[k]; cc = 0; cs = 0.0;
          }
          ++cc;
          cs += f[k];
        }
      }
      if (seg == 0) { hk = ck; hc = cc; hs = cs; }
      else { tk = ck; tc = cc; ts = cs; }
      // ---- face neighbours: +z inside the run and across its end, +y, +x
      if (adj) {
        const int x = (int)xq, y = yq;
        int pa = -1, pb = -1;
#pragma unroll
        for (int k = 0; k + 1 < kRun; ++k)
          if (k + 1 < n) link(adj, g, l[k], l[k + 1], pa, pb);
        {
          const int zn = z0 + n;  // first voxel after the run
          if (zn < g.N2) link(adj, g, ck, __ldg(labels + row * g.N2 + zn), pa, pb);  // ck = l[n - 1] after the fold
          else if (g.periodic && g.N2 > 1) link(adj, g, ck, __ldg(labels + row * g.N2), pa, pb);
        }
        if (g.N1 > 1 && (y + 1 < g.N1 || g.periodic)) {
          const long long row2 = y + 1 < g.N1 ? row + 1 : row - (g.N1 - 1);
          int m[kRun];
          load_labels(labels, row2 * g.N2 + z0, n, g.vec, m);
#pragma unroll
          for (int k = 0; k < kRun; ++k)
            if (k < n) link(adj, g, l[k], m[k], pa, pb);
        }
        if (g.N0 > 1 && (x + 1 < g.N0 || g.periodic)) {
          const long long off2 = x + 1 < g.N0 ? off + plane : off - (long long)(g.N0 - 1) * plane;
          int m[kRun];
          load_labels(labels, off2, n, g.vec, m);
#pragma unroll
          for (int k = 0; k < kRun; ++k)
            if (k < n) link(adj, g, l[k], m[k], pa, pb);
        }
      }
    }
    warp_reduce_by_key(hk, hc, hs, g, volume, mass, lane);
    if (__any_sync(kFull, tk >= 0)) warp_reduce_by_key(tk, tc, ts, g, volume, mass, lane);
  }
}

__global__ void degree_kernel(const uint32_t* __restrict__ adj, int words, long long n_cells, int32_t* __restrict__ degree) {
  const long long s = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (s >= n_cells) return;
  const int lane = threadIdx.x & 31;
  int c = 0;
  for (int w = lane; w < words; w += 32) c += __popc(adj[(size_t)s * words + w]);
  c = __reduce_add_sync(kFull, c);
  if (lane == 0) degree[s] = c;
}

}  // namespace
}  // namespace sbx

using namespace sbx;

extern "C" int sbx_cell_stats(int device, int ndim, const int64_t* N, const int32_t* labels, const double* field,
                              int64_t n_cells, int periodic, int64_t* volume, double* mass, uint32_t* adjacency,
                              int32_t* degree, void* stream) {
  if (!N || !labels || !volume) return fail(SBX_EINVAL, "NULL argument");
  if (ndim != 2 && ndim != 3) return fail(SBX_EINVAL, "ndim must be 2 or 3, got %d", ndim);
  if (n_cells < 1 || n_cells > (1 << 24)) return fail(SBX_EINVAL, "n_cells must be in 1..2^24, got %lld", (long long)n_cells);
  if (degree && !adjacency) return fail(SBX_EINVAL, "degree needs the adjacency buffer");
  if (mass && !field) return fail(SBX_EINVAL, "mass needs a field");
  int rc = require_sm100(device);
  if (rc) return rc;
  StatGeo g{};
  g.N0 = ndim == 3 ? (int)N[0] : 1;
  g.N1 = (int)N[ndim - 2];
  g.N2 = (int)N[ndim - 1];
  for (int a = 0; a < ndim; ++a)
    if (N[a] < 1 || N[a] > 65536) return fail(SBX_EINVAL, "grid extent %lld out of range", (long long)N[a]);
  g.runs_per_row = (g.N2 + kRun - 1) / kRun;
  g.tiles_y = (g.N1 + 15) / 16;
  g.tiles_z = (g.runs_per_row + 1) / 2;
  g.n_runs = (long long)g.N0 * g.tiles_y * g.tiles_z * 32;
  if (g.n_runs >> 5 >= (1LL << 31)) return fail(SBX_EINVAL, "grid too large for the cell-statistics kernel");
  g.periodic = periodic ? 1 : 0;
  g.words = (int)((n_cells + 31) / 32);
  g.n_cells = n_cells;
  g.vec = g.N2 % kRun == 0 && (reinterpret_cast<uintptr_t>(labels) & 15) == 0 && (!field || (reinterpret_cast<uintptr_t>(field) & 15) == 0);
  DeviceGuard guard(device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  SBX_CUDA(cudaMemsetAsync(volume, 0, (size_t)n_cells * sizeof(int64_t), st));
  if (mass) SBX_CUDA(cudaMemsetAsync(mass, 0, (size_t)n_cells * sizeof(double), st));
  if (adjacency) SBX_CUDA(cudaMemsetAsync(adjacency, 0, (size_t)n_cells * g.words * sizeof(uint32_t), st));
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  const long long want = (g.n_runs + 255) / 256;
  const int grid = (int)std::min<long long>(want, (long long)sms * 8);
  cell_stats_kernel<<<grid, 256, 0, st>>>(g, labels, mass ? field : nullptr, reinterpret_cast<unsigned long long*>(volume), mass, adjacency);
  SBX_CUDA(cudaGetLastError());
  if (degree) {
    degree_kernel<<<(unsigned)((n_cells + 7) / 8), 256, 0, st>>>(adjacency, g.words, n_cells, degree);
    SBX_CUDA(cudaGetLastError());
  }
  return SBX_OK;
}
