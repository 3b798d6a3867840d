// voronoi_cells.cuh -- cell-list labelling path (additive distance, no weights) for the per-pixel loop of
// VoronoiCell.weighted_voronoi_diagrams (SimuBox/calculator/voronoi.py:284-292 with VoronoiCell.additive, :244-246,
// weights == 0).  Included by voronoi.cu, which chooses between this path and the tile path (coarse / mid / label
// kernels): plane ranges of a grid sharded across GPUs, dense or many seeds, and the allocation-free entry point run
// here; whole grids with sparse, few seeds run the tile path (measurements: DESIGN.md section 4a).
//
// Periodicity in the reference is "replicate every seed into its 3^d images p + m L and take the argmin"
// (SURVEY.md 8c).  Here the images are made explicit once: every (seed, image) whose image can be the nearest
// one of that seed for some voxel of the grid becomes an IMAGE POINT q = fl(p + m L) carrying the seed index,
// and the labelling is a plain nearest-point query over image points -- no minimum-image logic, no ambiguous
// tiles.  Three launches, chained with programmatic dependent launch:
//   1. cell_count_kernel   one thread per (seed, image): cell of the image point (uniform cells of 2^sh voxels
//                          on a grid extended by E cells on either side, clamped: outermost cells stand for
//                          everything beyond); atomic count per cell and per cell row.
//   2. cell_scatter_kernel every CTA scans the (few hundred) row totals in shared memory; position of a point =
//                          row start + counts of the cells before it in its row + a per-cell cursor.  The CTAs
//                          also write the per-cell `start` table the label kernel reads.
//   3. label_cells_kernel  one WARP per gather tile (8 x 8 x 16 voxels / 32 x 32 pixels), no CTA-level barrier;
//                          on their way out its threads zero the counters for the next labelling.
//                          Cells are C-ordered, so all points of the cells [cl2, ch2] of one (c0, c1) row are
//                          one contiguous range: the warp reads the ranges of the rows that intersect
//                          tile +- R and stages the fp64 records of the points within R of the tile's box (cells
//                          overhang the box; a point farther than R on some axis is farther than R from every
//                          voxel).  A = the point nearest to the tile centre; if R does not cover the farthest
//                          voxel's distance to A the search is repeated with that radius.
//                          Pruning and evaluation use g_c(x) = d_c^2(x) - d_A^2(x), which is AFFINE in x: its four
//                          coefficients are formed in fp64, its minimum over a box is a closed form, so "can c beat
//                          A anywhere in the box" is exact up to a 1e-9 margin (tile level, then once more per half
//                          tile of 4 x 8 x 8 / 32 x 8 voxels against the half's own nearest point).
//                          Evaluation: a lane owns a column of 8 voxels; per candidate one fp32 FMA per voxel
//                          gives g_c, and the smallest and second smallest are tracked as integer keys with the
//                          candidate number in the low mantissa bits (FFMA, LOP3, 3 x VIMNMX per voxel and
//                          candidate).  A voxel whose two smallest differ by more than the certified fp32 error
//                          has the reference's label; a column holding any other voxel (near ties: ~1e-4 on random
//                          seeds, every cell face on a lattice) is redone with the reference's fp64 expression,
//                          rounded roots compared, equal roots to the lower seed index (np.argmin).
// A tile whose search box holds more than kStage points within R is labelled by a direct scan of the box's cells
// (exact, slow; counted in stats[2]).
#pragma once

namespace sbx {

constexpr int kStage = 160;  // image points staged per warp (fp64 records, 5 KB)

struct CellGeo {
  int nd, periodic;
  int N[3];      // grid, internal axes
  int xlo, xhi;  // labelled range on the first real axis (internal axis 3 - nd)
  int sh[3];     // log2 of the cell edge
  int E[3];      // extension cells on either side of the grid
  int n[3];      // cells per axis (extension included)
  int nt[3];     // gather tiles per axis (over the labelled range)
  int ncell, nrow;  // nrow = n[0] * n[1]
  long long S;
  double box[3];
  double R0;     // first search radius
};

// The image point of (seed s, combination `combo`), or false when that image can never be the nearest image of
// its seed for a voxel of the grid.  Per axis the three images q- = p - L < q0 = p < q+ = p + L split the line at
// p -+ L/2: q- is nearest below p - L/2, q+ above p + L/2.  An image is dropped only if every voxel coordinate
// 0..N-1 is on the far side of its bisector by a margin (1e-9 relative) that dwarfs the rounding of the
// reference's own differences, so the minimum over the kept images equals the minimum over all 3^d.
template <int ND, bool PER>
__device__ __forceinline__ bool image_point(const CellGeo& g, const double* __restrict__ seeds, unsigned s, unsigned combo,
                                            double q[3], int& row, int& c2) {
  int cell = 0;
  q[0] = 0.0;
#pragma unroll
  for (int a = 3 - ND; a < 3; ++a) {
    const double p = seed_coord<ND>(seeds, s, a);
    double qa = p;
    if (PER) {
      const unsigned m = combo % 3u;  // 0: p, 1: p - L, 2: p + L
      combo /= 3u;
      const double L = g.box[a], hL = 0.5 * L, top = (double)(g.N[a] - 1);
      const double mg = 1e-9 * (fabs(p) + L + top + 1.0);
      bool ok;
      if (m == 0) ok = (p - hL < top + mg) && (p + hL > -mg);
      else if (m == 1) { ok = p - hL > -mg; qa = __dsub_rn(p, L); }
      else { ok = p + hL < top + mg; qa = __dadd_rn(p, L); }
      if (!ok) return false;
    }
    q[a] = qa;
    int ci = (__double2int_rd(qa) >> g.sh[a]) + g.E[a];  // saturating conversion, then clamp: monotone in qa
    ci = min(max(ci, 0), g.n[a] - 1);
    if (a == 2) { row = cell; c2 = ci; }
    else cell = cell * g.n[a] + ci;
  }
  return true;
}

template <int ND, bool PER>
__global__ void __launch_bounds__(256)
cell_count_kernel(CellGeo g, const double* __restrict__ seeds, int* __restrict__ count, int* __restrict__ rowcount) {
  constexpr unsigned K = PER ? (ND == 3 ? 27u : 9u) : 1u;
  pdl_trigger();
  const unsigned t = blockIdx.x * 256u + threadIdx.x;
  if (t < (unsigned)g.S * K) {
    const unsigned s = t / K;
    double q[3];
    int row, c2;
    if (image_point<ND, PER>(g, seeds, s, t - s * K, q, row, c2)) {
      atomicAdd(&count[row * g.n[2] + c2], 1);
      atomicAdd(&rowcount[row], 1);
    }
  }
}

// Every CTA scans the row totals (nrow <= kRowMax, shared memory); a point's position = start of its row + the
// counts of the cells before it in the row + the next value of the cell's cursor.  CTA b also writes the `start`
// entries of the rows b, b + gridDim.x, ...
constexpr int kRowMax = 8192;
template <int ND, bool PER>
__global__ void __launch_bounds__(256)
cell_scatter_kernel(CellGeo g, const double* __restrict__ seeds, const int* __restrict__ count, const int* __restrict__ rowcount,
                    int* __restrict__ cursor, int* __restrict__ start, SeedRec* __restrict__ pts) {
  constexpr unsigned K = PER ? (ND == 3 ? 27u : 9u) : 1u;
  __shared__ int s_row[kRowMax + 1];
  __shared__ int s_warp[8];
  pdl_trigger();
  const unsigned t = blockIdx.x * 256u + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  bool ok = false;
  double q[3];
  int row = 0, c2 = 0;
  unsigned s = 0;
  if (t < (unsigned)g.S * K) {
    s = t / K;
    ok = image_point<ND, PER>(g, seeds, s, t - s * K, q, row, c2);
  }
  pdl_wait();  // the counts
  // exclusive scan of the row totals: thread t owns `per` consecutive rows
  const int per = (g.nrow + 255) >> 8;
  const int r_lo = min((int)threadIdx.x * per, g.nrow), r_hi = min(r_lo + per, g.nrow);
  int sum = 0;
  for (int r = r_lo; r < r_hi; ++r) {
    const int v = __ldcg(rowcount + r);
    s_row[r] = v;
    sum += v;
  }
  int inc = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int up = __shfl_up_sync(kFull, inc, o);
    if (lane >= o) inc += up;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  int run = inc - sum;
  for (int w = 0; w < warp; ++w) run += s_warp[w];
  for (int r = r_lo; r < r_hi; ++r) {
    const int v = s_row[r];
    s_row[r] = run;
    run += v;
  }
  if (threadIdx.x == 255) s_row[g.nrow] = run;
  __syncthreads();
  // per-cell starts of this CTA's rows (one warp per row, 32 cells a step)
  const int n2 = g.n[2];
  for (int r = blockIdx.x * 8 + warp; r < g.nrow; r += gridDim.x * 8) {
    int base = s_row[r];
    for (int c0 = 0; c0 < n2; c0 += 32) {
      const int c = c0 + lane;
      const int v = c < n2 ? __ldcg(count + r * n2 + c) : 0;
      int w = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(kFull, w, o);
        if (lane >= o) w += up;
      }
      if (c < n2) start[r * n2 + c] = base + w - v;
      base += __shfl_sync(kFull, w, 31);
    }
    if (r == g.nrow - 1 && lane == 0) start[g.ncell] = base;
  }
  if (ok) {
    int pos = s_row[row];
    const int* crow = count + row * n2;
    for (int c = 0; c < c2; ++c) pos += __ldcg(crow + c);
    pos += atomicAdd(&cursor[row * n2 + c2], 1);
    SeedRec rec;
    rec.x = q[0]; rec.y = q[1]; rec.z = q[2];
    rec.w = __longlong_as_double((long long)s);  // seed index in the low word
    pts[pos] = rec;
  }
}

__device__ __forceinline__ int rec_idx(const SeedRec& r) { return __double2loint(r.w); }
__device__ __forceinline__ SeedRec ld_rec(const SeedRec* __restrict__ p) {
  const double2 a = __ldg(reinterpret_cast<const double2*>(p)), b = __ldg(reinterpret_cast<const double2*>(p) + 1);
  SeedRec r;
  r.x = a.x; r.y = a.y; r.z = b.x; r.w = b.y;
  return r;
}

// the reference expression for one column of kCol voxels, candidates in ANY order: rounded roots compared, equal
// roots go to the lower seed index (= np.argmin over the seed-major image list)
template <int ND>
__device__ __forceinline__ void column_update_unordered(const SeedRec& R, double X, double Y, int z0, double bv[kCol], Col8& out) {
  const int id = rec_idx(R);
  const double dy = __dsub_rn(R.y, Y);
  double sxy = __dmul_rn(dy, dy);
  if (ND == 3) {
    const double dx = __dsub_rn(R.x, X);
    sxy = __dadd_rn(__dmul_rn(dx, dx), sxy);
  }
#pragma unroll
  for (int k = 0; k < kCol; ++k) {
    const double dz = __dsub_rn(R.z, (double)(z0 + k));
    const double val = __dsqrt_rn(__dadd_rn(sxy, __dmul_rn(dz, dz)));
    if (out.v[k] < 0 || val < bv[k] || (val == bv[k] && id < out.v[k])) { bv[k] = val; out.v[k] = id; }
  }
}

struct CellBox { int cl[3], ch[3]; };

// overflow path: every image point of the search box, straight from global memory
template <int ND>
__device__ __noinline__ Col8 column_scan_box(const SeedRec* __restrict__ pts, const int* __restrict__ start, CellBox b, int n1,
                                             int n2, double X, double Y, int z0) {
  Col8 out;
  double bv[kCol];
#pragma unroll
  for (int k = 0; k < kCol; ++k) { bv[k] = INFINITY; out.v[k] = -1; }
  for (int c0 = b.cl[0]; c0 <= b.ch[0]; ++c0)
    for (int c1 = b.cl[1]; c1 <= b.ch[1]; ++c1) {
      const int base = (c0 * n1 + c1) * n2;
      const int e = start[base + b.ch[2] + 1];
      for (int i = start[base + b.cl[2]]; i < e; ++i) column_update_unordered<ND>(ld_rec(pts + i), X, Y, z0, bv, out);
    }
  return out;
}

// Pruning and evaluation work on g_c(x) = d_c^2(x) - d_A^2(x) = sum_a e_a (s_a - 2 x_a), e = c - A, s = c + A, which is
// AFFINE in the voxel x: for a box with centre x0 and half extents h its minimum is g_c(x0) - 2 sum_a |e_a| h_a, and
// argmin_c g_c(x) = argmin_c d_c^2(x).  The four coefficients are formed in fp64 from the fp64 records (exact to
// 1e-16), so dropping a point is safe with a 1e-9 relative margin, and only their ROUNDING to fp32 and three fp32
// FMAs enter the per-voxel values: an absolute error below 4 * 2^-24 * (|g0| + 2 sum |e_a| h_a), ~1e-4 in d^2 units.
struct GForm { double g0, ex, ey, ez; };
template <int ND>
__device__ __forceinline__ GForm g_form(const SeedRec& A, const SeedRec& C, const double x0[3]) {
  GForm f;
  f.ex = ND == 3 ? C.x - A.x : 0.0;
  f.ey = C.y - A.y;
  f.ez = C.z - A.z;
  f.g0 = (ND == 3 ? f.ex * ((C.x + A.x) - 2.0 * x0[0]) : 0.0) + f.ey * ((C.y + A.y) - 2.0 * x0[1]) + f.ez * ((C.z + A.z) - 2.0 * x0[2]);
  return f;
}
// can C beat or tie A somewhere in the box (x0, h)?  `spread` = 2 sum |e_a| h_a.  NaN keeps the point.
__device__ __forceinline__ bool g_keep(const GForm& f, const double h[3], double& spread) {
  spread = 2.0 * (fabs(f.ex) * h[0] + fabs(f.ey) * h[1] + fabs(f.ez) * h[2]);
  return !(f.g0 - spread > 1e-9 * (fabs(f.g0) + spread + 1.0));
}

// bit k of the result = byte k of m is non-zero
__device__ __forceinline__ unsigned live_nibble(unsigned m) {
  const unsigned t = (((m & 0x7f7f7f7fu) + 0x7f7f7f7fu) | m) & 0x80808080u;
  return ((t >> 7) * 0x10204080u) >> 28;
}

// Gather tile of one warp and its four half tiles.  3D: 8 x 8 x 16, half h = (h & 1, h >> 1) of 4 x 8 x 8 voxels; a
// lane owns the column (lane >> 3, lane & 7, 0..7) of each.  2D (1 x X x Y): 32 x 32, halves 32 x 8 along the last
// axis; lane = row.
template <int ND> struct GTile;
template <> struct GTile<3> {
  static constexpr int G0 = 8, G1 = 8, G2 = 16, H = 4, W0 = 4, W1 = 8;
  __device__ static int off0(int h) { return 4 * (h & 1); }
  __device__ static int off2(int h) { return 8 * (h >> 1); }
  __device__ static void lane_xy(int lane, int& lx, int& ly) { lx = lane >> 3; ly = lane & 7; }
};
template <> struct GTile<2> {
  static constexpr int G0 = 1, G1 = 32, G2 = 32, H = 4, W0 = 1, W1 = 32;
  __device__ static int off0(int) { return 0; }
  __device__ static int off2(int h) { return 8 * h; }
  __device__ static void lane_xy(int lane, int& lx, int& ly) { lx = 0; ly = lane; }
};

__device__ __noinline__ unsigned load_live_generic(const uint8_t* __restrict__ mask, long long lin, int nz) {
  unsigned l = 0;
  for (int k = 0; k < nz; ++k)
    if (mask[lin + k]) l |= 1u << k;
  return l;
}

__device__ __noinline__ void store_column_generic(int32_t* __restrict__ labels, long long lin, int nz, unsigned live, Col8 b) {
  for (int k = 0; k < nz; ++k) labels[lin + k] = (live >> k & 1u) ? b.v[k] : -1;
}

// exact evaluation of one column against the staged records rec[list[0..n)] (list == nullptr: rec[0..n))
template <int ND>
__device__ __noinline__ Col8 column_listed_cells(const SeedRec* rec, const unsigned char* list, int n, double X, double Y, int z0) {
  Col8 out;
  double bv[kCol];
#pragma unroll
  for (int k = 0; k < kCol; ++k) { bv[k] = INFINITY; out.v[k] = -1; }
  for (int j = 0; j < n; ++j) column_update_unordered<ND>(rec[list ? list[j] : j], X, Y, z0, bv, out);
  return out;
}

// one gather tile (warp-level; all lanes of the warp call it together)
template <int ND>
__device__ __forceinline__ void label_tile(const CellGeo& g, const int t[3], const SeedRec* __restrict__ pts, const int* __restrict__ start,
                                           const uint8_t* __restrict__ mask, int32_t* __restrict__ labels,
                                           unsigned long long* __restrict__ stats, SeedRec* s_stage, float4* s_G, int* s_kidx,
                                           unsigned char* s_klist, int lane) {
  using GT = GTile<ND>;
  constexpr int H = GT::H, AX = 3 - ND;
  const unsigned lt = (1u << lane) - 1u;
  int lx, ly;
  GT::lane_xy(lane, lx, ly);
  const long long plane = ND == 3 ? (long long)g.N[1] * g.N[2] : (long long)g.N[2];  // stride of the first real axis
  const int n1 = g.n[1], n2 = g.n[2];

  int lo[3], hi[3];
  bool full = true;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const int G = a == 0 ? GT::G0 : (a == 1 ? GT::G1 : GT::G2);
    const int first = a == AX ? g.xlo : 0, end = a == AX ? g.xhi : g.N[a];
    lo[a] = first + t[a] * G;
    hi[a] = min(lo[a] + G, end) - 1;
    full = full && hi[a] - lo[a] + 1 == G;
  }
  // vector path: whole tile inside the grid and every column 8-byte (mask) / 32-byte (labels) aligned
  const bool vec = full && (g.N[2] & 7) == 0 && (reinterpret_cast<uintptr_t>(labels) & 31) == 0 &&
                   (!mask || (reinterpret_cast<uintptr_t>(mask) & 7) == 0);
  // lin(h) = tile_lin + GT::off0(h) * N1 * N2 + GT::off2(h)   (first voxel of this lane's column in half h)
  const long long tile_lin = (((long long)(lo[0] + lx) * g.N[1] + lo[1] + ly) * g.N[2] + lo[2]) - (long long)g.xlo * plane;
  const long long h_stride = (long long)g.N[1] * g.N[2];
  // ---- liveness of the half tiles, 8 bits each (the mask bytes are the longest latency: issued first)
  unsigned live_all = 0, any_h = 0;
  if (vec) {
    if (mask) {
      uint2 m[H];
#pragma unroll
      for (int h = 0; h < H; ++h) m[h] = __ldg(reinterpret_cast<const uint2*>(mask + tile_lin + GT::off0(h) * h_stride + GT::off2(h)));
#pragma unroll
      for (int h = 0; h < H; ++h) live_all |= (live_nibble(m[h].x) | live_nibble(m[h].y) << 4) << (8 * h);
    } else {
      live_all = 0xffffffffu;
    }
  } else {
#pragma unroll 1
    for (int h = 0; h < H; ++h) {
      const int vx = lo[0] + GT::off0(h) + lx, vy = lo[1] + ly, z0 = lo[2] + GT::off2(h);
      const int nz = min(kCol, hi[2] - z0 + 1);
      if (vx <= hi[0] && vy <= hi[1] && nz > 0) {
        const unsigned l = mask ? load_live_generic(mask, tile_lin + GT::off0(h) * h_stride + GT::off2(h), nz) : (1u << nz) - 1u;
        live_all |= l << (8 * h);
      }
    }
  }
#pragma unroll
  for (int h = 0; h < H; ++h) any_h |= __any_sync(kFull, (live_all >> (8 * h)) & 0xffu) ? 1u << h : 0u;

  int n = 0;
  bool over = false;
  CellBox cb;
  if (any_h) {
    pdl_wait();  // the sorted image points
    // ---- gather: points of the cell rows that intersect tile +- R; A = nearest to the tile centre
    const double c[3] = {0.5 * (double)(lo[0] + hi[0]), 0.5 * (double)(lo[1] + hi[1]), 0.5 * (double)(lo[2] + hi[2])};
    const double hh[3] = {(double)hi[0] - c[0], (double)hi[1] - c[1], (double)hi[2] - c[2]};  // half extents
    double R = g.R0;
    int slotA = 0;
#pragma unroll 1
    for (;;) {
      bool all = true;
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        cb.cl[a] = min(max((__double2int_rd((double)lo[a] - R) >> g.sh[a]) + g.E[a], 0), g.n[a] - 1);
        cb.ch[a] = min(max((__double2int_rd((double)hi[a] + R) >> g.sh[a]) + g.E[a], 0), g.n[a] - 1);
        all = all && cb.cl[a] == 0 && cb.ch[a] == g.n[a] - 1;
      }
      const int ny = cb.ch[1] - cb.cl[1] + 1, nrows = (cb.ch[0] - cb.cl[0] + 1) * ny;
      const float inv_ny = 1.0f / (float)ny;
      unsigned bkey = 0xffffffffu;  // (bits of the fp32 squared distance to the centre, low bits dropped)
      float bneed = 0.f;
      n = 0;
#pragma unroll 1
      for (int r0 = 0; r0 < nrows; r0 += 32) {
        const int r = r0 + lane;
        int s = 0, len = 0;
        if (r < nrows) {
          int rx = (int)(((float)r + 0.5f) * inv_ny);
          int ry = r - rx * ny;
          if (ry < 0) { --rx; ry += ny; }
          else if (ry >= ny) { ++rx; ry -= ny; }
          const int base = ((cb.cl[0] + rx) * n1 + cb.cl[1] + ry) * n2;
          s = __ldg(start + base + cb.cl[2]);
          len = __ldg(start + base + cb.ch[2] + 1) - s;
        }
        int inc = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int up = __shfl_up_sync(kFull, inc, o);
          if (lane >= o) inc += up;
        }
        const int tot = __shfl_sync(kFull, inc, 31), off = inc - len;
#pragma unroll 1
        for (int j0 = 0; j0 < tot; j0 += 32) {
          const int j = j0 + lane;
          int pos = 0;  // largest lane whose exclusive offset is <= j (empty rows share an offset with their successor)
#pragma unroll
          for (int step = 16; step > 0; step >>= 1) {
            const int tv = __shfl_sync(kFull, off, pos + step);
            if (tv <= j) pos += step;
          }
          const int srow = __shfl_sync(kFull, s, pos), orow = __shfl_sync(kFull, off, pos);
          unsigned key = 0xffffffffu;
          float need = 0.f;
          bool inside = false;
          SeedRec v;
          if (j < tot) {
            v = ld_rec(pts + srow + (j - orow));
            const double ax = ND == 3 ? fabs(v.x - c[0]) : 0.0, ay = fabs(v.y - c[1]), az = fabs(v.z - c[2]);
            key = (__float_as_uint((float)(ax * ax + ay * ay + az * az)) & ~31u) | (unsigned)lane;
            // farthest voxel of the tile from this point, rounded up: what the search box has to cover if it becomes A
            const float fx = (float)(ax + hh[0]), fy = (float)(ay + hh[1]), fz = (float)(az + hh[2]);
            need = sqrtf(fx * fx + fy * fy + fz * fz) * 1.0001f + 1e-3f;
            // the cells of the search box overhang it: a point farther than R from the tile's box on some axis is
            // farther than R from every voxel of the tile and is not staged (when the box is the whole cell grid
            // there is no R to trust and everything is kept)
            inside = all || (ax <= hh[0] + R && ay <= hh[1] + R && az <= hh[2] + R);
          }
          const unsigned bal = __ballot_sync(kFull, inside);
          const int slot = n + __popc(bal & lt);
          if (inside && slot < kStage) s_stage[slot] = v;
          const unsigned kmin = __reduce_min_sync(kFull, key);
          if (kmin < bkey) {  // uniform
            bkey = kmin;
            const int src = kmin & 31u;
            bneed = __shfl_sync(kFull, need, src);
            slotA = __shfl_sync(kFull, inside ? slot : -1, src);  // -1: outside the box, then bneed > R and the search repeats
          }
          n += __popc(bal);
        }
      }
      if (bkey == 0xffffffffu) {  // no point in the box yet
        if (all) break;  // cannot happen: S >= 1 and every seed keeps an image
        R *= 2.0;
        continue;
      }
      if ((double)bneed <= R || all) break;
      R = (double)bneed * 1.001;
    }
    over = n > kStage;
    __syncwarp();
    if (!over) {
      // ---- tile-level pruning against A, compacted in place
      const SeedRec A = s_stage[slotA];
      int m = 0;
#pragma unroll 1
      for (int c0 = 0; c0 < n; c0 += 32) {
        const int ci = c0 + lane;
        SeedRec B;
        bool keep = false;
        if (ci < n) {
          B = s_stage[ci];
          double spread;
          keep = g_keep(g_form<ND>(A, B, c), hh, spread);
        }
        const unsigned bal = __ballot_sync(kFull, keep);
        __syncwarp();  // every read of this round is done; the targets (<= ci) were read this round or earlier
        if (keep) s_stage[m + __popc(bal & lt)] = B;
        m += __popc(bal);
      }
      __syncwarp();
      n = m;
    }
  }
  unsigned long long evals = 0;

#pragma unroll 1
  for (int h = 0; h < H; ++h) {
    int bi[kCol];
#pragma unroll
    for (int k = 0; k < kCol; ++k) bi[k] = -1;
    const int vx = lo[0] + GT::off0(h) + lx, vy = lo[1] + ly, z0 = lo[2] + GT::off2(h);
    const int nz = min(kCol, hi[2] - z0 + 1);
    const unsigned live = (live_all >> (8 * h)) & 0xffu;
    if (any_h >> h & 1u) {
      if (over) {
        if (live) {
          const Col8 ex = column_scan_box<ND>(pts, start, cb, n1, n2, (double)vx, (double)vy, z0);
#pragma unroll
          for (int k = 0; k < kCol; ++k) bi[k] = ex.v[k];
        }
        evals += (unsigned long long)__popc(live) * (unsigned long long)n;
      } else {
        // ---- half box: centre x0, half extents hb; A2 = staged point nearest to x0
        const int ux = min(lo[0] + GT::off0(h) + GT::W0 - 1, hi[0]), uy = min(lo[1] + GT::W1 - 1, hi[1]), uz = min(z0 + kCol - 1, hi[2]);
        const double x0[3] = {0.5 * (double)(lo[0] + GT::off0(h) + ux), 0.5 * (double)(lo[1] + uy), 0.5 * (double)(z0 + uz)};
        const double hb[3] = {(double)ux - x0[0], (double)uy - x0[1], (double)uz - x0[2]};
        unsigned best = 0xffffffffu;
        for (int ci = lane; ci < n; ci += 32) {
          const SeedRec P = s_stage[ci];
          const double ex = ND == 3 ? P.x - x0[0] : 0.0, ey = P.y - x0[1], ez = P.z - x0[2];
          best = min(best, (__float_as_uint((float)(ex * ex + ey * ey + ez * ez)) & ~255u) | (unsigned)ci);
        }
        best = __reduce_min_sync(kFull, best);
        const SeedRec A2 = s_stage[best & 255u];
        // d_A2^2 is at most ua2 on the half, so g_c(x) >= -ua2 for every point c
        float ua2;
        {
          const double bx = ND == 3 ? fabs(A2.x - x0[0]) + hb[0] : 0.0, by = fabs(A2.y - x0[1]) + hb[1], bz = fabs(A2.z - x0[2]) + hb[2];
          ua2 = (float)(bx * bx + by * by + bz * bz);
        }
        int kept = 0;
        float mmax = 0.f;  // largest |g0| + spread of a kept point
#pragma unroll 1
        for (int c0 = 0; c0 < n; c0 += 32) {
          const int ci = c0 + lane;
          GForm f;
          double spread = 0.0;
          bool keep = false;
          int idx = 0;
          if (ci < n) {
            const SeedRec B = s_stage[ci];
            idx = rec_idx(B);
            f = g_form<ND>(A2, B, x0);
            keep = g_keep(f, hb, spread);
          }
          const unsigned bal = __ballot_sync(kFull, keep);
          const int d = kept + __popc(bal & lt);
          if (keep && d < 32) {
            s_G[d] = make_float4((float)f.g0, (float)(-2.0 * f.ex), (float)(-2.0 * f.ey), (float)(-2.0 * f.ez));
            s_kidx[d] = idx;
            s_klist[d] = (unsigned char)ci;
            mmax = fmaxf(mmax, (float)(fabs(f.g0) + spread));
          }
          kept += __popc(bal);
        }
        __syncwarp();
        const double X = (double)vx, Y = (double)vy;
        if (kept == 1) {
          const int id = s_kidx[0];
#pragma unroll
          for (int k = 0; k < kCol; ++k) bi[k] = id;
        } else if (kept <= 32) {
          // ---- fp32 evaluation of the affine forms: smallest and second smallest per voxel, candidate number in the
          // low mantissa bits (3 bits when kept <= 8)
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) mmax = fmaxf(mmax, __shfl_xor_sync(kFull, mmax, o));
          // the bias keeps every computed value positive (its bits then order like the value): -g <= ua2, and the fp32
          // evaluation is off by less than 5 * 2^-24 * (bias + |g0| + spread)
          const float bias = ua2 * 1.001f + 1e-6f * mmax + 0.01f;
          const unsigned tmask = kept <= 8 ? ~7u : ~31u;
          const float Xr = (float)(X - x0[0]), Yr = (float)(Y - x0[1]), Zr = (float)((double)z0 - x0[2]);
          unsigned m1[kCol], m2[kCol];
#pragma unroll
          for (int k = 0; k < kCol; ++k) m1[k] = m2[k] = 0x7f800000u;  // +inf
#pragma unroll 2
          for (int j = 0; j < kept; ++j) {
            const float4 G = s_G[j];
            const float gxy = fmaf(G.y, Xr, fmaf(G.z, Yr, fmaf(G.w, Zr, G.x + bias)));
#pragma unroll
            for (int k = 0; k < kCol; ++k) {
              const unsigned key = (__float_as_uint(fmaf(G.w, (float)k, gxy)) & tmask) | (unsigned)j;
              m2[k] = min(m2[k], max(m1[k], key));
              m1[k] = min(m1[k], key);
            }
          }
          // certified where the two smallest differ by more than the error of both: rounding of the coefficients and
          // four FMAs, < 5 * 2^-24 * (bias + |g0| + spread) each, plus the dropped mantissa bits, < 2^-20 (2^-18) relative
          const float cert = 2e-6f * (bias + mmax) + 1e-7f, rel = kept <= 8 ? 0.999996f : 0.999984f;
          float worst = INFINITY;
#pragma unroll
          for (int k = 0; k < kCol; ++k) {
            const float a1 = __uint_as_float(m1[k] & tmask), a2 = __uint_as_float(m2[k] & tmask);
            const float slack = fmaf(a2, rel, -a1);
            worst = (live >> k & 1u) ? fminf(worst, slack) : worst;
            bi[k] = s_kidx[m1[k] & ~tmask];
          }
          if (!(worst > cert)) {  // this lane's column again, the reference's way
            const Col8 ex = column_listed_cells<ND>(s_stage, s_klist, kept, X, Y, z0);
#pragma unroll
            for (int k = 0; k < kCol; ++k) bi[k] = ex.v[k];
          }
        } else {
          // more survivors than the compact list holds (clustered / duplicate seeds): all staged points, exactly
          const Col8 ex = column_listed_cells<ND>(s_stage, nullptr, n, X, Y, z0);
#pragma unroll
          for (int k = 0; k < kCol; ++k) bi[k] = ex.v[k];
        }
        evals += (unsigned long long)kept * (unsigned long long)__popc(live);
        __syncwarp();
      }
    }
    if (vx <= hi[0] && vy <= hi[1] && nz > 0) {
      const long long lin = tile_lin + GT::off0(h) * h_stride + GT::off2(h);
      if (vec) {
        int4 a, b;
        a.x = live & 1u ? bi[0] : -1; a.y = live & 2u ? bi[1] : -1; a.z = live & 4u ? bi[2] : -1; a.w = live & 8u ? bi[3] : -1;
        b.x = live & 16u ? bi[4] : -1; b.y = live & 32u ? bi[5] : -1; b.z = live & 64u ? bi[6] : -1; b.w = live & 128u ? bi[7] : -1;
        int4* dst = reinterpret_cast<int4*>(labels + lin);
        dst[0] = a;
        dst[1] = b;
      } else {
        Col8 cb8;
#pragma unroll
        for (int k = 0; k < kCol; ++k) cb8.v[k] = bi[k];
        store_column_generic(labels, lin, nz, live, cb8);
      }
    }
  }
  if (stats) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) evals += __shfl_xor_sync(kFull, evals, off);
    if (lane == 0) {
      if (evals) atomicAdd(&stats[0], evals);
      if (any_h) atomicAdd(&stats[1], 1ull);
      if (over) atomicAdd(&stats[2], 1ull);
    }
  }
}

// grid = (ceil(nt[2] / 4), nt[1], nt[0]): warp w of a CTA owns tile (blockIdx.z, blockIdx.y, 4 blockIdx.x + w).
// On their way out the threads put the counters of the binning kernels back to zero for the next labelling.
template <int ND, int MINB>
__global__ void __launch_bounds__(128, MINB)
label_cells_kernel(CellGeo g, const SeedRec* __restrict__ pts, const int* __restrict__ start, int* __restrict__ zero_me, int n_zero,
                   const uint8_t* __restrict__ mask, int32_t* __restrict__ labels, unsigned long long* __restrict__ stats) {
  constexpr int W = 4;
  __shared__ alignas(16) SeedRec s_stage_all[W][kStage];
  __shared__ alignas(16) float4 s_G_all[W][32];
  __shared__ int s_kidx_all[W][32];
  __shared__ unsigned char s_klist_all[W][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int t[3] = {(int)blockIdx.z, (int)blockIdx.y, (int)blockIdx.x * W + warp};
  if (t[2] < g.nt[2])
    label_tile<ND>(g, t, pts, start, mask, labels, stats, s_stage_all[warp], s_G_all[warp], s_kidx_all[warp], s_klist_all[warp], lane);
  const int cta = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
  const int nthreads = gridDim.x * gridDim.y * gridDim.z * 128;
  int i = cta * 128 + threadIdx.x;
  if (i < n_zero) {
    pdl_wait();  // the scatter kernel has finished with the counters
    for (; i < n_zero; i += nthreads) zero_me[i] = 0;
  }
}

}  // namespace sbx
