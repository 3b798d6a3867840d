// seeds.cu -- periodic 2D/3D domain-centre finder.
//
// The reference finds domain centres only in 2D: it tiles the field 3 x 3, thresholds it to a uint8
// image and takes contour moments with OpenCV (SimuBox/calculator/voronoi.py:31-68), and refuses 3D
// (web/pages/3_Voronoi.py:156-157).  Here: threshold -> connected components of the foreground under
// face connectivity WITH wrap-around (no tiling) -> centroid of every component, 2D or 3D.
//
//   foreground   field > threshold                                   (strict, like cv2.THRESH_BINARY)
//   components   union-find over the +x/+y/+z face neighbours; unions keep the SMALLER voxel index as
//                root, so a component's root is its minimum linear index and components come out
//                ordered by that index -- independent of thread scheduling
//   centroid     root + (sum of minimum-image integer displacements voxel - root) / count, wrapped into
//                [0, N): integer sums, hence bit-reproducible
//
// Kernels: init (threshold + own parent), merge (lock-free union with atomicMin), flatten (parent =
// root), rank (order roots: per-CTA count, one-CTA scan, scatter), accumulate (warp reduce-by-key on the
// component id, then one atomic per distinct component per warp).
#include <algorithm>
#include <cstdint>

#include "common.h"

namespace sbx {
namespace {

constexpr unsigned kFull = 0xffffffffu;

struct SeedGeo {
  int N0, N1, N2;  // 2D grids arrive as 1 x X x Y
  int periodic;
  long long n;     // voxels (< 2^31)
};

// parent links only ever decrease; reads go to L2 (ld.cg) so that a retry sees the other threads' unions
__device__ __forceinline__ int find_root(const int* parent, int i) {
  int p = __ldcg(parent + i);
  while (p != i) {
    i = p;
    p = __ldcg(parent + i);
  }
  return i;
}

// union keeping the smaller index as root (Playne-Hawick / Komura label equivalence)
__device__ __forceinline__ void unite(int* parent, int a, int b) {
  for (;;) {
    a = find_root(parent, a);
    b = find_root(parent, b);
    if (a == b) return;
    if (a > b) {
      const int t = a;
      a = b;
      b = t;
    }
    const int old = atomicMin(parent + b, a);  // b was a root: hang it under a
    if (old == b) return;
    b = old;  // somebody re-parented b meanwhile: retry from there
  }
}

// parent = first voxel of the z-run the voxel belongs to, as far as the warp's 32 consecutive voxels show it (runs
// are stitched across warp segments and around the box by the merge kernel): most +z unions are done right here
__global__ void seeds_init_kernel(SeedGeo g, const double* __restrict__ field, double threshold, int* __restrict__ parent) {
  const int lane = threadIdx.x & 31;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long n_iter = (g.n + stride - 1) / stride;
  for (long long it = 0; it < n_iter; ++it) {
    const long long i = it * stride + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool in = i < g.n;
    const bool fg = in && field[i] > threshold;
    const bool row_start = in && ((unsigned)i % (unsigned)g.N2) == 0;
    const unsigned f = __ballot_sync(kFull, fg), brk = __ballot_sync(kFull, row_start);
    if (in) {
      int p = -1;
      if (fg) {
        const unsigned below = (1u << lane) - 1u;
        const unsigned gaps = ~f & below;                       // background voxels before me
        const int after_gap = gaps ? 32 - __clz(gaps) : 0;      // first lane after the last gap
        const unsigned rows = brk & (below | (1u << lane));     // row starts at or before me
        const int at_row = rows ? 31 - __clz(rows) : 0;
        p = (int)(i - lane + max(after_gap, at_row));
      }
      parent[i] = p;
    }
  }
}

__global__ void seeds_merge_kernel(SeedGeo g, int* parent) {
  const long long plane = (long long)g.N1 * g.N2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < g.n; i += (long long)gridDim.x * blockDim.x) {
    if (parent[i] < 0) continue;
    const unsigned ui = (unsigned)i, urow = ui / (unsigned)g.N2;  // n < 2^31: 32-bit divisions
    const int z = (int)(ui - urow * (unsigned)g.N2);
    const int x = (int)(urow / (unsigned)g.N1), y = (int)(urow - (unsigned)x * (unsigned)g.N1);
    // +z: inside a warp segment the init kernel already joined the run; stitch segment ends and the wrap face
    if (g.N2 > 1) {
      if (z + 1 < g.N2) {
        if ((ui & 31u) == 31u && parent[i + 1] >= 0) unite(parent, (int)i, (int)(i + 1));
      } else if (g.periodic && parent[i - (g.N2 - 1)] >= 0) {
        unite(parent, (int)i, (int)(i - (g.N2 - 1)));
      }
    }
    // +y / +x: the face (i, j) is redundant when the previous voxels of both runs are foreground too
    // (i ~ i-1, j ~ j-1 along z, and the face (i-1, j-1) is united by the thread of i-1)
    if (g.N1 > 1 && (y + 1 < g.N1 || g.periodic)) {
      const long long j = y + 1 < g.N1 ? i + g.N2 : i - (long long)(g.N1 - 1) * g.N2;
      if (parent[j] >= 0 && !(z > 0 && parent[i - 1] >= 0 && parent[j - 1] >= 0)) unite(parent, (int)i, (int)j);
    }
    if (g.N0 > 1 && (x + 1 < g.N0 || g.periodic)) {
      const long long j = x + 1 < g.N0 ? i + plane : i - (long long)(g.N0 - 1) * plane;
      if (parent[j] >= 0 && !(z > 0 && parent[i - 1] >= 0 && parent[j - 1] >= 0)) unite(parent, (int)i, (int)j);
    }
  }
}

__global__ void seeds_flatten_kernel(SeedGeo g, int* parent) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < g.n; i += (long long)gridDim.x * blockDim.x)
    if (parent[i] >= 0) parent[i] = find_root(parent, (int)i);
}

// ---- order the roots: chunk c = voxels [c * kChunk, (c + 1) * kChunk)
constexpr int kChunk = 4096;

__global__ void __launch_bounds__(256) seeds_count_kernel(SeedGeo g, const int* __restrict__ parent, int* __restrict__ chunk_count) {
  __shared__ int s_n;
  if (threadIdx.x == 0) s_n = 0;
  __syncthreads();
  const long long base = (long long)blockIdx.x * kChunk;
  int c = 0;
  for (int k = threadIdx.x; k < kChunk; k += 256) {
    const long long i = base + k;
    if (i < g.n && parent[i] == (int)i) ++c;
  }
  c = __reduce_add_sync(kFull, c);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_n, c);
  __syncthreads();
  if (threadIdx.x == 0) chunk_count[blockIdx.x] = s_n;
}

// exclusive scan of chunk_count in place (one CTA), total -> *n_components
__global__ void __launch_bounds__(1024) seeds_scan_kernel(int* chunk_count, int n_chunks, int* n_components) {
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int base = 0; base < n_chunks; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = i < n_chunks ? chunk_count[i] : 0;
    int incl = v;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int t = __shfl_up_sync(kFull, incl, off);
      if (lane >= off) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int w = s_warp[lane];
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const int t = __shfl_up_sync(kFull, w, off);
        if (lane >= off) w += t;
      }
      s_warp[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const int before = s_carry + (warp ? s_warp[warp - 1] : 0) + incl - v;
    if (i < n_chunks) chunk_count[i] = before;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) *n_components = s_carry;
}

// roots of a chunk get consecutive ids in voxel order: comp[root] = id, roots[id] = root
__global__ void __launch_bounds__(256) seeds_rank_kernel(SeedGeo g, const int* __restrict__ parent, const int* __restrict__ chunk_first,
                                                         int* __restrict__ comp, int* __restrict__ roots, long long max_components) {
  __shared__ int s_warp[8];
  __shared__ int s_base;
  if (threadIdx.x == 0) s_base = chunk_first[blockIdx.x];
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long base = (long long)blockIdx.x * kChunk;
  for (int k0 = 0; k0 < kChunk; k0 += 256) {
    const long long i = base + k0 + threadIdx.x;
    const bool is_root = i < g.n && parent[i] == (int)i;
    const unsigned bal = __ballot_sync(kFull, is_root);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    int before = s_base;
    for (int w = 0; w < warp; ++w) before += s_warp[w];
    if (is_root) {
      const int id = before + __popc(bal & ((1u << lane) - 1u));
      comp[i] = id;
      if (id < max_components) roots[id] = (int)i;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int add = 0;
      for (int w = 0; w < 8; ++w) add += s_warp[w];
      s_base += add;
    }
    __syncthreads();
  }
}

// comp[i] = id of i's root (-1 on the background); root entries were written by the rank kernel
__global__ void seeds_relabel_kernel(SeedGeo g, const int* __restrict__ parent, int* comp) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < g.n; i += (long long)gridDim.x * blockDim.x) {
    const int p = parent[i];
    if (p < 0) comp[i] = -1;
    else if (p != (int)i) comp[i] = comp[p];
  }
}

__device__ __forceinline__ int min_image(int d, int N, int periodic) {
  if (!periodic) return d;
  const int h = N >> 1;
  int m = (d + h) % N;
  if (m < 0) m += N;
  return m - h;  // in [-N/2, N - 1 - N/2]
}

struct Part {
  int key, cnt, sx, sy, sz;
};

__device__ __forceinline__ void flush_part(const Part& p, long long* acc, long long max_components) {
  if (p.key >= 0 && p.key < max_components && p.cnt > 0) {
    unsigned long long* a = reinterpret_cast<unsigned long long*>(acc) + (size_t)p.key * 4;
    atomicAdd(a + 0, (unsigned long long)(long long)p.cnt);
    atomicAdd(a + 1, (unsigned long long)(long long)p.sx);
    atomicAdd(a + 2, (unsigned long long)(long long)p.sy);
    atomicAdd(a + 3, (unsigned long long)(long long)p.sz);
  }
}

__device__ __forceinline__ void warp_flush_by_key(const Part& p, long long* acc, long long max_components, int lane) {
  const unsigned grp = __match_any_sync(kFull, p.key);
  unsigned todo = kFull;
  while (todo) {
    const int leader = __ffs(todo) - 1;
    const unsigned gm = __shfl_sync(kFull, grp, leader);
    const bool mine = grp == gm;
    Part t;
    t.key = p.key;
    t.cnt = __reduce_add_sync(kFull, mine ? p.cnt : 0);
    t.sx = __reduce_add_sync(kFull, mine ? p.sx : 0);
    t.sy = __reduce_add_sync(kFull, mine ? p.sy : 0);
    t.sz = __reduce_add_sync(kFull, mine ? p.sz : 0);
    if (lane == leader) flush_part(t, acc, max_components);
    todo &= ~gm;
  }
}

// acc[id] += (count, sum dx, sum dy, sum dz) with d = minimum-image displacement from the component's root
__global__ void __launch_bounds__(256)
seeds_accumulate_kernel(SeedGeo g, const int* __restrict__ parent, const int* __restrict__ comp, long long* acc, long long max_components,
                        int tiles_y, int tiles_z, int runs_per_row, long long n_runs) {
  const int lane = threadIdx.x & 31;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long n_iter = (n_runs + stride - 1) / stride;
  const long long plane = (long long)g.N1 * g.N2;
  for (long long it = 0; it < n_iter; ++it) {
    const long long r = it * stride + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned w = (unsigned)(r >> 5);
    const unsigned wy = w / (unsigned)tiles_z;
    const int tz = (int)(w - wy * (unsigned)tiles_z);
    const unsigned xq = wy / (unsigned)tiles_y;
    const int ty = (int)(wy - xq * (unsigned)tiles_y);
    const int yq = ty * 16 + (lane >> 1), run = tz * 2 + (lane & 1);
    Part head{-1, 0, 0, 0, 0}, tail{-1, 0, 0, 0, 0};
    if (r < n_runs && yq < g.N1 && run < runs_per_row) {
      const int z0 = run * 8;
      const int n = min(8, g.N2 - z0);
      const long long off = ((long long)xq * g.N1 + yq) * g.N2 + z0;
      Part cur{-2, 0, 0, 0, 0};
      int seg = 0, rx = 0, ry = 0, rz = 0;
      for (int k = 0; k < n; ++k) {
        const int id = comp[off + k];
        if (id != cur.key) {
          if (cur.key != -2) {
            if (seg == 0) head = cur;
            else flush_part(cur, acc, max_components);
            ++seg;
          }
          cur = Part{id, 0, 0, 0, 0};
          if (id >= 0) {
            const unsigned root = (unsigned)parent[off + k];
            rx = (int)(root / (unsigned)plane);
            const int rem = (int)(root - (unsigned)rx * (unsigned)plane);
            ry = rem / g.N2;
            rz = rem - ry * g.N2;
          }
        }
        if (id >= 0) {
          ++cur.cnt;
          cur.sx += min_image((int)xq - rx, g.N0, g.periodic);
          cur.sy += min_image(yq - ry, g.N1, g.periodic);
          cur.sz += min_image(z0 + k - rz, g.N2, g.periodic);
        }
      }
      if (seg == 0) head = cur;
      else tail = cur;
      if (head.key == -2) head.key = -1;
    }
    warp_flush_by_key(head, acc, max_components, lane);
    if (__any_sync(kFull, tail.key >= 0)) warp_flush_by_key(tail, acc, max_components, lane);
  }
}

__global__ void seeds_finalize_kernel(SeedGeo g, int ndim, const int* __restrict__ roots, const long long* __restrict__ acc,
                                      long long n, double* __restrict__ centres, long long* __restrict__ volumes) {
  const long long id = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= n) return;
  const long long plane = (long long)g.N1 * g.N2;
  const int root = roots[id];
  const int rx = (int)(root / plane);
  const int rem = (int)(root - (long long)rx * plane);
  const int r[3] = {rx, rem / g.N2, rem % g.N2};
  const int N[3] = {g.N0, g.N1, g.N2};
  const long long cnt = acc[id * 4];
  if (volumes) volumes[id] = cnt;
  for (int a = 3 - ndim; a < 3; ++a) {
    double c = __dadd_rn((double)r[a], __ddiv_rn((double)acc[id * 4 + 1 + a], (double)cnt));
    if (g.periodic) {
      if (c < 0.0) c = __dadd_rn(c, (double)N[a]);
      else if (c >= (double)N[a]) c = __dsub_rn(c, (double)N[a]);
    }
    centres[id * ndim + (a - (3 - ndim))] = c;
  }
}

}  // namespace
}  // namespace sbx

using namespace sbx;

extern "C" int sbx_domain_centres(int device, int ndim, const int64_t* N, const double* field, double threshold,
                                  int periodic, int32_t* comp_labels, int64_t max_components, double* centres,
                                  int64_t* volumes, int64_t* n_components_host, void* stream) {
  if (!N || !field || !centres || !n_components_host) return fail(SBX_EINVAL, "NULL argument");
  if (ndim != 2 && ndim != 3) return fail(SBX_EINVAL, "ndim must be 2 or 3, got %d", ndim);
  if (max_components < 1) return fail(SBX_EINVAL, "max_components must be >= 1");
  int rc = require_sm100(device);
  if (rc) return rc;
  SeedGeo g{};
  g.N0 = ndim == 3 ? (int)N[0] : 1;
  g.N1 = (int)N[ndim - 2];
  g.N2 = (int)N[ndim - 1];
  g.periodic = periodic ? 1 : 0;
  g.n = 1;
  for (int a = 0; a < ndim; ++a) {
    if (N[a] < 1 || N[a] > 65536) return fail(SBX_EINVAL, "grid extent %lld out of range", (long long)N[a]);
    g.n *= N[a];
  }
  if (g.n >= (1LL << 31)) return fail(SBX_EINVAL, "grids of 2^31 voxels or more are not supported by the seed finder");
  DeviceGuard guard(device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int n_chunks = (int)((g.n + kChunk - 1) / kChunk);
  // stream-ordered scratch, returned to the pool on every exit path
  struct Scratch {
    cudaStream_t st;
    void* ptrs[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    int n = 0;
    explicit Scratch(cudaStream_t s) : st(s) {}
    ~Scratch() { for (int i = 0; i < n; ++i) cudaFreeAsync(ptrs[i], st); }
    cudaError_t get(void** p, size_t bytes) {
      const cudaError_t e = cudaMallocAsync(p, bytes, st);
      if (e == cudaSuccess) ptrs[n++] = *p;
      return e;
    }
  } scratch(st);
  int *parent = nullptr, *chunk = nullptr, *roots = nullptr, *own_comp = nullptr;
  long long* acc = nullptr;
  SBX_CUDA(scratch.get((void**)&parent, (size_t)g.n * sizeof(int)));
  SBX_CUDA(scratch.get((void**)&chunk, ((size_t)n_chunks + 1) * sizeof(int)));
  SBX_CUDA(scratch.get((void**)&roots, (size_t)max_components * sizeof(int)));
  SBX_CUDA(scratch.get((void**)&acc, (size_t)max_components * 4 * sizeof(long long)));
  if (!comp_labels) {
    SBX_CUDA(scratch.get((void**)&own_comp, (size_t)g.n * sizeof(int)));
    comp_labels = own_comp;
  }
  SBX_CUDA(cudaMemsetAsync(acc, 0, (size_t)max_components * 4 * sizeof(long long), st));
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  const int grid = (int)std::min<long long>((g.n + 255) / 256, (long long)sms * 8);
  seeds_init_kernel<<<grid, 256, 0, st>>>(g, field, threshold, parent);
  seeds_merge_kernel<<<grid, 256, 0, st>>>(g, parent);
  seeds_flatten_kernel<<<grid, 256, 0, st>>>(g, parent);
  seeds_count_kernel<<<n_chunks, 256, 0, st>>>(g, parent, chunk);
  seeds_scan_kernel<<<1, 1024, 0, st>>>(chunk, n_chunks, chunk + n_chunks);
  seeds_rank_kernel<<<n_chunks, 256, 0, st>>>(g, parent, chunk, comp_labels, roots, max_components);
  seeds_relabel_kernel<<<grid, 256, 0, st>>>(g, parent, comp_labels);
  const int runs_per_row = (g.N2 + 7) / 8, tiles_y = (g.N1 + 15) / 16, tiles_z = (runs_per_row + 1) / 2;
  const long long n_runs = (long long)g.N0 * tiles_y * tiles_z * 32;
  const int grid_acc = (int)std::min<long long>((n_runs + 255) / 256, (long long)sms * 8);
  seeds_accumulate_kernel<<<grid_acc, 256, 0, st>>>(g, parent, comp_labels, acc, max_components, tiles_y, tiles_z, runs_per_row, n_runs);
  SBX_CUDA(cudaGetLastError());
  int n_comp = 0;
  SBX_CUDA(cudaMemcpyAsync(&n_comp, chunk + n_chunks, sizeof(int), cudaMemcpyDeviceToHost, st));
  SBX_CUDA(cudaStreamSynchronize(st));
  *n_components_host = n_comp;
  const long long filled = std::min<long long>(n_comp, max_components);
  if (filled > 0) {
    seeds_finalize_kernel<<<(unsigned)((filled + 127) / 128), 128, 0, st>>>(g, ndim, roots, acc, filled, centres, reinterpret_cast<long long*>(volumes));
    SBX_CUDA(cudaGetLastError());
  }
  return SBX_OK;
}
