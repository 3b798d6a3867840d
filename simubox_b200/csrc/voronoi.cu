// voronoi.cu -- grid Voronoi labelling of domain centres (sm_100a).
//
// Replaces the per-pixel Python loop of VoronoiCell.weighted_voronoi_diagrams
// (SimuBox/calculator/voronoi.py:284-292) with VoronoiCell.additive (:244-246) or
// VoronoiCell.power (:249-255) as the distance.  Two paths behind one plan (the host code at the end of this file
// chooses: sbx_voronoi_plan_execute_range): the TILE path below (any mode, weights; whole grids with sparse, few
// seeds) and the CELL-LIST path of voronoi_cells.cuh (additive, no weights: plane ranges, dense or many seeds,
// sbx_voronoi_label_noalloc).  The grid is always handled as three internal
// axes (a 2D image is 1 x X x Y); the tile path is three kernels:
//   1. coarse_kernel: per COARSE tile (32^3 voxels / 256^2 pixels) the seeds that can win anywhere
//      inside it, out of all S seeds (interval bounds with slack, ascending seed index);
//   2. mid_kernel: one warp per MID tile (16^3 / 64^2) filters its parent's list and writes staged
//      records: fl(p + m L) of the single periodic image that is nearest for the whole tile;
//   3. label_kernel: one CTA per half mid tile.  The records arrive in shared memory by TMA bulk copy
//      (cp.async.bulk + mbarrier).  Every warp owns a 4x8x8 (1x32x8) sub-tile, prunes the list once
//      more -- for unweighted seeds with the exact bisector test (d_B^2 - d_A^2 is affine in x, so its
//      minimum over a box sits on a corner) -- and keeps the survivors in a per-warp list.  Every
//      thread owns a column of 8 voxels along the last axis: (dx^2 + dy^2) is formed once per
//      candidate and dz^2 (computed once per warp) added per voxel, in the reference's summation order, with round-to-nearest
//      intrinsics (no FMA contraction).  First minimum wins (np.argmin).
// Periodic boxes use the minimum image per axis, which equals the minimum over the 3^d replicated
// images of the reference expression because every fp64 step is monotone.  Tiles where no single
// image is nearest everywhere (tiny boxes), power mode, and tiles whose candidate list overflows
// fall back to a generic per-voxel evaluation of all three images per axis.
#include <cstdint>
#include <cstdlib>
#include <type_traits>
#include <cstdlib>
#include <cuda_runtime.h>
#include <new>

#include "common.h"

namespace sbx {

constexpr int kCap = 128;          // candidate records staged per mid tile
constexpr int kCoarseCap = 1024;   // candidate indices kept per coarse tile
constexpr int kCol = 8;            // voxels per thread along the last axis
constexpr int kFlagGeneric = 1 << 30;  // mid tile needs the generic 3-image evaluation
constexpr unsigned kFull = 0xffffffffu;

struct VorGeo {
  int nd;         // real dimensions (2 or 3); internal axis 0 is a dummy of extent 1 when nd == 2
  int N[3];       // grid, internal axes
  int M[3];       // mid tile extent per axis
  int C[3];       // coarse tile extent per axis
  int nm[3];      // mid tiles per axis
  int nc[3];      // coarse tiles per axis
  int periodic;
  int mode;       // 0 additive, 1 power
  double box[3];  // internal axes
  long long S;
};

struct alignas(16) SeedRec { double x, y, z, w; };

__device__ __forceinline__ void tile_coords(const int n[3], int tile, int t[3]) {
  for (int a = 2; a >= 0; --a) {
    t[a] = tile % n[a];
    tile /= n[a];
  }
}

// seed coordinate on internal axis a (a >= 3 - ND)
template <int ND>
__device__ __forceinline__ double seed_coord(const double* __restrict__ seeds, long long s, int a) {
  return seeds[s * ND + (a - (3 - ND))];
}

// Conservative bounds of the seed's value over the voxel box with centre c and half extent h
// (internal axes).  Periodic boxes use the allowed image (p, p - L or p + L) nearest to the centre:
// for every voxel x of the box and every allowed image q', |q' - x| >= |q' - c| - h >= |q - c| - h,
// and the true value is at most |q - x| <= |q - c| + h, so both bounds hold whichever image is
// nearest to x.  `amb` is set when another image may be nearer for some voxel of the box (the
// evaluation then has to look at all three images).  Keys: additive without weights compares
// squared distances (no square root); additive with weights sqrt(.) - w; power (.) - w.
template <int ND>
__device__ __forceinline__ void tile_bounds(const VorGeo& g, const double* __restrict__ seeds,
                                            const double* __restrict__ weights, long long s, const double c[3],
                                            const double h[3], double& LB, double& UB, bool& amb) {
  double smin = 0.0, smax = 0.0;
  amb = false;
#pragma unroll
  for (int a = 3 - ND; a < 3; ++a) {
    double p = seed_coord<ND>(seeds, s, a);
    if (g.mode == 1) p += (double)s;  // voronoi.py:251 adds the seed index to every coordinate
    double d = p - c[a];
    if (g.periodic) {
      const double L = g.box[a], hL = 0.5 * L;
      if (d > hL) d -= L;
      else if (d < -hL) d += L;
      amb |= !(fabs(d) + h[a] < hL - 1e-9 * (L + 1.0));
    }
    const double t = fabs(d);
    const double dmax = t + h[a], dmin = fmax(t - h[a], 0.0);
    smin += dmin * dmin;
    smax += dmax * dmax;
  }
  if (g.mode == 1) {
    const double w = weights ? weights[s] : 0.0;
    LB = smin - w; UB = smax - w;
  } else if (weights) {
    const double w = weights[s];
    LB = sqrt(smin) - w; UB = sqrt(smax) - w;
  } else {
    LB = smin; UB = smax;
  }
}

// fl(p + m L) of the image tile_bounds picked: the number the reference evaluates for that image
template <int ND>
__device__ __forceinline__ void resolve_image(const VorGeo& g, const double* __restrict__ seeds, long long s,
                                              const double c[3], double pr[3]) {
  pr[0] = 0.0;
#pragma unroll
  for (int a = 3 - ND; a < 3; ++a) {
    const double p0 = seed_coord<ND>(seeds, s, a);
    pr[a] = p0;
    if (g.periodic) {
      const double d = (g.mode == 1 ? p0 + (double)s : p0) - c[a];
      const double L = g.box[a], hL = 0.5 * L;
      if (d > hL) pr[a] = __dsub_rn(p0, L);
      else if (d < -hL) pr[a] = __dadd_rn(p0, L);
    }
  }
}

// tile_bounds that also returns which periodic image it used, as a code: bits 2a..2a+1 for internal axis a
// (0 = the seed itself, 1 = p - L, 2 = p + L), bit 6 = `amb`.  The mid kernel parks this byte next to the lower
// bound, so a seed that survives is not evaluated a second time: its record is fl(p -+ L) per the code.
template <int ND>
__device__ __forceinline__ void tile_bounds_code(const VorGeo& g, const double* __restrict__ seeds,
                                                 const double* __restrict__ weights, long long s, const double c[3],
                                                 const double h[3], double& LB, double& UB, unsigned& code) {
  double smin = 0.0, smax = 0.0;
  code = 0;
#pragma unroll
  for (int a = 3 - ND; a < 3; ++a) {
    const double p0 = seed_coord<ND>(seeds, s, a);
    double d = (g.mode == 1 ? p0 + (double)s : p0) - c[a];
    if (g.periodic) {
      const double L = g.box[a], hL = 0.5 * L;
      if (d > hL) { d -= L; code |= 1u << (2 * a); }
      else if (d < -hL) { d += L; code |= 2u << (2 * a); }
      if (!(fabs(d) + h[a] < hL - 1e-9 * (L + 1.0))) code |= 64u;
    }
    const double t = fabs(d);
    const double dmax = t + h[a], dmin = fmax(t - h[a], 0.0);
    smin += dmin * dmin;
    smax += dmax * dmax;
  }
  if (g.mode == 1) {
    const double w = weights ? weights[s] : 0.0;
    LB = smin - w; UB = smax - w;
  } else if (weights) {
    const double w = weights[s];
    LB = sqrt(smin) - w; UB = sqrt(smax) - w;
  } else {
    LB = smin; UB = smax;
  }
}

__device__ __forceinline__ bool may_win(double LB, double U) {
  const double slack = 1e-10 * (fabs(U) + fabs(LB) + 1.0);
  return LB - slack <= U + slack;  // never drops a seed that could tie or win
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

// Programmatic dependent launch (PDL): the three kernels of one labelling are chained with
// cudaLaunchAttributeProgrammaticStreamSerialization, so a kernel's CTAs are scheduled (and run their
// prologue) while the previous kernel drains; pdl_wait() returns once the previous grid has completed and its
// writes are visible.  Both are no-ops in a kernel launched without the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(kFull, v, o));
  return v;
}

// Bisector test (unweighted seeds): can B beat or tie A anywhere in the voxel box [lo, hi]?
// d_B^2 - d_A^2 = sum_a e_a (s_a - 2 x_a) with e = B - A, s = B + A is affine in x, so its minimum over
// the box is attained per axis at lo or hi.  UA = max over the box of d_A^2.  The 1e-11 relative
// margin covers the rounding of this bound and of the reference's own evaluation, so a seed is
// dropped only if it loses to A at every voxel of the box by far more than an ulp.
template <int ND>
__device__ __forceinline__ bool bisector_keep(const SeedRec& A, const SeedRec& B, const double lo[3], const double hi[3], double UA) {
  double gmin = 0.0, scale = 0.0;
  if (ND == 3) {
    const double e = B.x - A.x, s = B.x + A.x;
    gmin += fmin(e * (s - 2.0 * lo[0]), e * (s - 2.0 * hi[0]));
    scale += fabs(e) * (fabs(A.x) + fabs(B.x) + 2.0 * hi[0] + 1.0);
  }
  {
    const double e = B.y - A.y, s = B.y + A.y;
    gmin += fmin(e * (s - 2.0 * lo[1]), e * (s - 2.0 * hi[1]));
    scale += fabs(e) * (fabs(A.y) + fabs(B.y) + 2.0 * hi[1] + 1.0);
  }
  {
    const double e = B.z - A.z, s = B.z + A.z;
    gmin += fmin(e * (s - 2.0 * lo[2]), e * (s - 2.0 * hi[2]));
    scale += fabs(e) * (fabs(A.z) + fabs(B.z) + 2.0 * hi[2] + 1.0);
  }
  return !(gmin > 1e-11 * (scale + UA + 1.0));  // also keeps NaN
}

template <int ND>
__device__ __forceinline__ double box_max_d2(const SeedRec& A, const double lo[3], const double hi[3]) {
  const double ax0 = A.x - lo[0], ax1 = A.x - hi[0], ay0 = A.y - lo[1], ay1 = A.y - hi[1], az0 = A.z - lo[2], az1 = A.z - hi[2];
  return (ND == 3 ? fmax(ax0 * ax0, ax1 * ax1) : 0.0) + fmax(ay0 * ay0, ay1 * ay1) + fmax(az0 * az0, az1 * az1);
}

// warp argmin of squared distance to the point (cx, cy, cz) over the records rec[0..count)
template <int ND>
__device__ __forceinline__ int nearest_record(const SeedRec* rec, int count, double cx, double cy, double cz, int lane) {
  double bd = INFINITY;
  int bc = 0;
  for (int c = lane; c < count; c += 32) {
    const SeedRec R = rec[c];
    const double dx = R.x - cx, dy = R.y - cy, dz = R.z - cz;
    const double d2 = (ND == 3 ? dx * dx : 0.0) + dy * dy + dz * dz;
    if (d2 < bd) { bd = d2; bc = c; }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const double od = __shfl_xor_sync(kFull, bd, off);
    const int oc = __shfl_xor_sync(kFull, bc, off);
    if (od < bd || (od == bd && oc < bc)) { bd = od; bc = oc; }
  }
  return bc;
}

// ------------------------------------------------------------------------------------------
// Coarse level: one CTA per coarse tile scans ALL seeds.  Seeds are taken interleaved (thread t,
// round k -> seed 4096 c + 256 k + t: coalesced loads); lower bounds of a 4096-seed chunk are
// cached in shared memory, the keep flags go to a shared bitmap whose prefix popcount gives every
// kept seed its slot, so the output is in ascending seed index without a barrier per round.
// ------------------------------------------------------------------------------------------
template <int ND>
__global__ void __launch_bounds__(256, 4)
coarse_kernel(VorGeo g, const double* __restrict__ seeds, const double* __restrict__ weights,
              int* __restrict__ clist, int* __restrict__ ccount) {
  constexpr int kRounds = 16, kChunk = 256 * kRounds;
  const int tile = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  pdl_trigger();
  int t[3];
  tile_coords(g.nc, tile, t);
  double c[3], h[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const double lo = (double)(t[a] * g.C[a]);
    const double hi = (double)(min(t[a] * g.C[a] + g.C[a], g.N[a]) - 1);
    c[a] = 0.5 * (lo + hi);
    h[a] = 0.5 * (hi - lo);
  }
  __shared__ double s_lb[kChunk];
  __shared__ double s_red[8];
  __shared__ unsigned s_bits[kRounds * 8];
  __shared__ int s_pref[kRounds * 8 + 1];
  const long long n = g.S;
  const bool cache = n <= kChunk;
  double best = INFINITY;
  bool amb;
  for (long long c0 = 0; c0 < n; c0 += kChunk) {
#pragma unroll 2
    for (int k = 0; k < kRounds; ++k) {
      const long long s = c0 + k * 256 + tid;
      double LB = INFINITY, UB = INFINITY;
      if (s < n) tile_bounds<ND>(g, seeds, weights, s, c, h, LB, UB, amb);
      best = fmin(best, UB);
      if (cache) s_lb[k * 256 + tid] = LB;
    }
  }
  best = warp_min(best);
  if (lane == 0) s_red[warp] = best;
  __syncthreads();
  double U = s_red[0];
#pragma unroll
  for (int w = 1; w < 8; ++w) U = fmin(U, s_red[w]);

  int base = 0;
  for (long long c0 = 0; c0 < n; c0 += kChunk) {
#pragma unroll 2
    for (int k = 0; k < kRounds; ++k) {
      const long long s = c0 + k * 256 + tid;
      double LB = INFINITY, UB;
      if (cache) LB = s_lb[k * 256 + tid];
      else if (s < n) tile_bounds<ND>(g, seeds, weights, s, c, h, LB, UB, amb);
      const unsigned bal = __ballot_sync(kFull, s < n && may_win(LB, U));
      if (lane == 0) s_bits[k * 8 + warp] = bal;
    }
    __syncthreads();
    if (warp == 0) {  // exclusive prefix popcount over the 128 bitmap words
      int v[4], sum = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) { v[j] = __popc(s_bits[lane * 4 + j]); sum += v[j]; }
      int inc = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(kFull, inc, o);
        if (lane >= o) inc += up;
      }
      int run = inc - sum;
#pragma unroll
      for (int j = 0; j < 4; ++j) { s_pref[lane * 4 + j] = run; run += v[j]; }
      if (lane == 31) s_pref[kRounds * 8] = inc;
    }
    __syncthreads();
#pragma unroll 4
    for (int k = 0; k < kRounds; ++k) {
      const unsigned word = s_bits[k * 8 + warp];
      if (word >> lane & 1u) {
        const int slot = base + s_pref[k * 8 + warp] + __popc(word & ((1u << lane) - 1u));
        if (slot < kCoarseCap) clist[(long long)tile * kCoarseCap + slot] = (int)(c0 + k * 256 + tid);
      }
    }
    base += s_pref[kRounds * 8];
    __syncthreads();
  }
  if (tid == 0) ccount[tile] = base;
}

// ------------------------------------------------------------------------------------------
// Mid level: one warp per mid tile filters its parent's list (all seeds if the parent overflowed)
// and writes staged records + indices in ascending seed index.  rcount = count | generic flag.
// ------------------------------------------------------------------------------------------
template <int ND>
__global__ void __launch_bounds__(128, 6)
mid_kernel(VorGeo g, int n_mid, const double* __restrict__ seeds, const double* __restrict__ weights,
           const int* __restrict__ clist, const int* __restrict__ ccount, SeedRec* __restrict__ rec,
           int* __restrict__ ridx, int* __restrict__ rcount) {
  constexpr int kRounds = 8, kChunk = 32 * kRounds;
  __shared__ double s_lb[4][kChunk];
  __shared__ unsigned char s_code[4][kChunk];   // image code of every candidate (tile_bounds_code)
  __shared__ alignas(16) SeedRec s_kr[4][kCap];  // records that passed the interval test, per warp
  __shared__ int s_ki[4][kCap];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int tile = blockIdx.x * 4 + warp;
  pdl_trigger();
  if (tile >= n_mid) return;
  int t[3];
  tile_coords(g.nm, tile, t);
  double c[3], h[3];
  int parent = 0;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const double lo = (double)(t[a] * g.M[a]);
    const double hi = (double)(min(t[a] * g.M[a] + g.M[a], g.N[a]) - 1);
    c[a] = 0.5 * (lo + hi);
    h[a] = 0.5 * (hi - lo);
    parent = parent * g.nc[a] + (t[a] * g.M[a]) / g.C[a];
  }
  pdl_wait();  // the coarse lists
  const int pc = ccount[parent];
  const int* src = pc <= kCoarseCap ? clist + (long long)parent * kCoarseCap : nullptr;
  const long long n = src ? pc : g.S;
  const bool cache = n <= kChunk;
  double best = INFINITY;
  bool amb;
  int base = 0;
  bool any_amb = false;
  if (cache) {
    // The usual case (a coarse list of <= 256 seeds): one evaluation per candidate.  All seed indices are fetched
    // first (one dependent load less per round), the lower bound and the image code of every candidate are parked
    // in shared memory, and the second pass only compares, shifts the coordinates of the survivors and stores.
    int sid[kRounds];
#pragma unroll
    for (int k = 0; k < kRounds; ++k) {
      const int i = k * 32 + lane;
      sid[k] = i < n ? (src ? src[i] : i) : -1;
    }
#pragma unroll
    for (int k = 0; k < kRounds; ++k) {
      double LB = INFINITY, UB = INFINITY;
      unsigned code = 0;
      if (sid[k] >= 0) tile_bounds_code<ND>(g, seeds, weights, sid[k], c, h, LB, UB, code);
      best = fmin(best, UB);
      s_lb[warp][k * 32 + lane] = LB;
      s_code[warp][k * 32 + lane] = (unsigned char)code;
    }
    const double U = warp_min(best);
#pragma unroll
    for (int k = 0; k < kRounds; ++k) {
      if (k * 32 >= n) break;
      const bool keep = sid[k] >= 0 && may_win(s_lb[warp][k * 32 + lane], U);
      const unsigned bal = __ballot_sync(kFull, keep);
      if (keep) {
        const int slot = base + __popc(bal & ((1u << lane) - 1u));
        const unsigned code = s_code[warp][k * 32 + lane];
        any_amb |= (code & 64u) != 0;
        if (slot < kCap) {
          double pr[3] = {0.0, 0.0, 0.0};
#pragma unroll
          for (int a = 3 - ND; a < 3; ++a) {
            const double p0 = seed_coord<ND>(seeds, sid[k], a);
            const unsigned sh = (code >> (2 * a)) & 3u;
            pr[a] = sh == 1u ? __dsub_rn(p0, g.box[a]) : (sh == 2u ? __dadd_rn(p0, g.box[a]) : p0);
          }
          SeedRec r;
          r.x = pr[0]; r.y = pr[1]; r.z = pr[2];
          r.w = weights ? weights[sid[k]] : 0.0;
          s_kr[warp][slot] = r;
          s_ki[warp][slot] = sid[k];
        }
      }
      base += __popc(bal);
    }
  } else {
  for (long long c0 = 0; c0 < n; c0 += kChunk) {
#pragma unroll 2
    for (int k = 0; k < kRounds; ++k) {
      const long long i = c0 + k * 32 + lane;
      double LB = INFINITY, UB = INFINITY;
      if (i < n) tile_bounds<ND>(g, seeds, weights, src ? src[i] : i, c, h, LB, UB, amb);
      best = fmin(best, UB);
    }
  }
  const double U = warp_min(best);
  for (long long c0 = 0; c0 < n; c0 += kChunk) {
#pragma unroll 2
    for (int k = 0; k < kRounds; ++k) {
      if (c0 + k * 32 >= n) break;
      const long long i = c0 + k * 32 + lane;
      const long long s = i < n ? (src ? src[i] : i) : 0;
      double LB = INFINITY, UB;
      if (i < n) tile_bounds<ND>(g, seeds, weights, s, c, h, LB, UB, amb);
      const bool keep = i < n && may_win(LB, U);
      const unsigned bal = __ballot_sync(kFull, keep);
      if (keep) {
        const int slot = base + __popc(bal & ((1u << lane) - 1u));
        any_amb |= amb;
        if (slot < kCap) {
          double pr[3];
          resolve_image<ND>(g, seeds, s, c, pr);
          SeedRec r;
          r.x = pr[0]; r.y = pr[1]; r.z = pr[2];
          r.w = weights ? weights[s] : 0.0;
          s_kr[warp][slot] = r;
          s_ki[warp][slot] = (int)s;
        }
      }
      base += __popc(bal);
    }
  }
  }
  const bool generic = __any_sync(kFull, any_amb) || g.mode == 1;
  __syncwarp();
  SeedRec* out_r = rec + (long long)tile * kCap;
  int* out_i = ridx + (long long)tile * kCap;
  for (int j = lane; j < min(base, kCap); j += 32) {
    out_r[j] = s_kr[warp][j];
    out_i[j] = s_ki[warp][j];
  }
  if (lane == 0) rcount[tile] = base | (generic ? kFlagGeneric : 0);
}

// the reference expression for one seed at one voxel, all three images (voronoi.py:245 / :250-254)
template <int ND>
__device__ __forceinline__ double seed_value(const VorGeo& g, const double* __restrict__ seeds, long long s, double w,
                                             const double x[3]) {
  double ss = 0.0;
  const double shift = (double)s;
#pragma unroll
  for (int a = 3 - ND; a < 3; ++a) {
    const double p = seed_coord<ND>(seeds, s, a);
    double d = __dsub_rn(p, x[a]);
    if (g.mode == 1) d = __dadd_rn(d, shift);
    if (g.periodic) {
      double dm = __dsub_rn(__dsub_rn(p, g.box[a]), x[a]);
      double dp = __dsub_rn(__dadd_rn(p, g.box[a]), x[a]);
      if (g.mode == 1) { dm = __dadd_rn(dm, shift); dp = __dadd_rn(dp, shift); }
      d = fmin(fabs(d), fmin(fabs(dm), fabs(dp)));
    }
    const double sq = __dmul_rn(d, d);
    ss = a == 3 - ND ? sq : __dadd_rn(ss, sq);  // np.linalg.norm: sequential sum x -> y -> z
  }
  const double r = __dsqrt_rn(ss);
  return __dsub_rn(g.mode == 1 ? __dmul_rn(r, r) : r, w);
}

// Rare path of the unweighted compare, kept out of line so that the inner loop stays small: one column of kCol
// voxels against the warp's surviving records, in the reference's order.  Sums of squares are compared; when a
// strictly smaller sum is within 8 ulps of the running minimum the rounded roots decide, as in the reference
// (equal roots leave the voxel with the earlier, lower index: np.argmin).
struct Col8 { int v[kCol]; };
template <int ND>
__device__ __noinline__ Col8 column_exact(const SeedRec* rec, const int* idx, const unsigned char* list, int kept, double X,
                                          double Y, int z0) {
  Col8 out;
  double bs[kCol];
#pragma unroll
  for (int k = 0; k < kCol; ++k) { bs[k] = INFINITY; out.v[k] = -1; }
  for (int j = 0; j < kept; ++j) {
    const int c = list[j];
    const SeedRec R = rec[c];
    const int id = idx[c];
    const double dy = __dsub_rn(R.y, Y);
    double sxy = __dmul_rn(dy, dy);
    if (ND == 3) {
      const double dx = __dsub_rn(R.x, X);
      sxy = __dadd_rn(__dmul_rn(dx, dx), sxy);
    }
#pragma unroll
    for (int k = 0; k < kCol; ++k) {
      const double dz = __dsub_rn(R.z, (double)(z0 + k));
      const double ss = __dadd_rn(sxy, __dmul_rn(dz, dz));
      bool take = ss < bs[k];
      if (take && __double_as_longlong(bs[k]) - __double_as_longlong(ss) <= 8) take = __dsqrt_rn(ss) < __dsqrt_rn(bs[k]);
      if (take) { bs[k] = ss; out.v[k] = id; }
    }
  }
  return out;
}

// Warp / lane geometry of the label kernel.  A CTA of 8 warps labels HALF a mid tile (split along the
// last axis).  3D: mid tile 16^3, half = 16 x 16 x 8 = 4 x 2 warps of 4 x 8 x 8 voxels;
// 2D (1 x X x Y): mid tile 64 x 64, half = 64 x 32 = 2 x 4 warps of 32 x 8 pixels.  A lane owns one
// column of 8 voxels along the last axis.
template <int ND> struct WarpTile;
template <> struct WarpTile<3> {
  static constexpr int M0 = 16, M1 = 16, M2 = 16, W0 = 4, W1 = 8;
  // gw = 0..15: the warp's number within the mid tile
  __device__ static void origin(int gw, int o[3]) { o[0] = (gw & 3) * 4; o[1] = ((gw >> 2) & 1) * 8; o[2] = (gw >> 3) * 8; }
  __device__ static void lane_xy(int lane, int& lx, int& ly) { lx = lane >> 3; ly = lane & 7; }
};
template <> struct WarpTile<2> {
  static constexpr int M0 = 1, M1 = 64, M2 = 64, W0 = 1, W1 = 32;
  __device__ static void origin(int gw, int o[3]) { o[0] = 0; o[1] = (gw & 1) * 32; o[2] = (gw >> 1) * 8; }
  __device__ static void lane_xy(int lane, int& lx, int& ly) { lx = 0; ly = lane; }
};

// MODE 0: generic evaluation only (power mode).  MODE 1: additive, no weights -- bisector pruning,
// compare sums of squares and take the square root only when the two sums are within 8 ulps, where
// rounding of the root could turn "smaller" into "equal" (then the lower index keeps the voxel, like
// argmin).  MODE 2: additive with weights -- interval pruning, sqrt(ss) - w per candidate.
// W warps per CTA (a mid tile = 16 warp tiles = 16 / W CTAs): warps finish at very different times (a tile inside
// one cell keeps a single candidate), and a CTA's slot is only recycled when its slowest warp is done.
// Each warp works through R warp tiles of the same mid tile one after the other (no barrier in between: the staged
// records are read-only), so the staging, the barrier and the TMA wait are paid once per R * W warp tiles.
template <int ND, int MODE, int W, int R>
__global__ void __launch_bounds__(32 * W, 24 / W)
label_kernel(VorGeo g, const double* __restrict__ seeds, const double* __restrict__ weights,
             const SeedRec* __restrict__ rec, const int* __restrict__ ridx, const int* __restrict__ rcount,
             const int* __restrict__ clist, const int* __restrict__ ccount, const uint8_t* __restrict__ mask,
             int32_t* __restrict__ labels, unsigned long long* __restrict__ stats) {
  using WT = WarpTile<ND>;
  __shared__ alignas(128) SeedRec s_rec[kCap];
  __shared__ alignas(16) int s_idx[kCap];
  __shared__ alignas(16) double s_dz[W][32];      // per warp: (z_c - z_k)^2 of a batch of 4 candidates x 8 voxels
  __shared__ unsigned char s_list[W][kCap];       // per warp: surviving record slots, ascending
  __shared__ alignas(8) uint64_t s_bar;
  // grid = (kParts * nm[2], nm[1], nm[0]): no integer division to find the tile
  constexpr int kParts = 16 / (W * R);
  const int t[3] = {(int)blockIdx.z, (int)blockIdx.y, (int)(blockIdx.x / kParts)};
  const int part = blockIdx.x % kParts;
  const int tile = (t[0] * g.nm[1] + t[1]) * g.nm[2] + t[2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int lx, ly;
  WT::lane_xy(lane, lx, ly);
  pdl_wait();  // the staged records of the mid kernel
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
  unsigned long long evals = 0;
  bool waited = false;
#pragma unroll 1
  for (int round = 0; round < R; ++round) {
  int o[3];
  WT::origin((part * R + round) * W + warp, o);
  const int x0 = t[0] * WT::M0 + o[0], y0 = t[1] * WT::M1 + o[1], z0 = t[2] * WT::M2 + o[2];
  const bool w_in = x0 < g.N[0] && y0 < g.N[1] && z0 < g.N[2];
  const int vx = x0 + lx, vy = y0 + ly;
  const bool col_in = w_in && vx < g.N[0] && vy < g.N[1];
  const int nz = min(kCol, g.N[2] - z0);
  const long long lin = ((long long)vx * g.N[1] + vy) * g.N[2] + z0;
  const bool mfast = mask && col_in && nz == kCol && (reinterpret_cast<uintptr_t>(mask + lin) & 7) == 0;
  uint2 mraw = make_uint2(0u, 0u);
  if (mfast) mraw = __ldg(reinterpret_cast<const uint2*>(mask + lin));  // issued first: the longest latency of the round
  if (!w_in) continue;  // whole warp outside the grid (no CTA barrier follows)
  unsigned live = col_in ? (1u << nz) - 1u : 0u;
  if (mask && col_in) {
    live = 0;
    if (mfast) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if ((mraw.x >> (8 * k)) & 0xffu) live |= 1u << k;
        if ((mraw.y >> (8 * k)) & 0xffu) live |= 1u << (k + 4);
      }
    } else {
      for (int k = 0; k < nz; ++k)
        if (mask[lin + k]) live |= 1u << k;
    }
  }
  const bool any_live = __any_sync(kFull, live != 0);
  int bi[kCol];
#pragma unroll
  for (int k = 0; k < kCol; ++k) bi[k] = -1;

  if (staged && !waited) {
    mbar_wait(&s_bar, 0);
    waited = true;
  }
  if (any_live) {
    if (!staged || generic) {
      // generic path: every thread scans the candidate set with the three-image expression
      const int* src = nullptr;
      long long n = g.S;
      if (staged) n = count;
      else {
        int parent = 0;
#pragma unroll
        for (int a = 0; a < 3; ++a) parent = parent * g.nc[a] + (t[a] * g.M[a]) / g.C[a];
        const int pc = ccount[parent];
        if (pc <= kCoarseCap) { src = clist + (long long)parent * kCoarseCap; n = pc; }
      }
      for (int k = 0; k < nz; ++k) {
        if (!(live >> k & 1u)) continue;
        const double x[3] = {(double)vx, (double)vy, (double)(z0 + k)};
        double best = INFINITY;
        int b = -1;
        for (long long cidx = 0; cidx < n; ++cidx) {
          const long long s = staged ? s_idx[cidx] : (src ? src[cidx] : cidx);
          const double val = seed_value<ND>(g, seeds, s, weights ? weights[s] : 0.0, x);
          if (val < best || b < 0) { best = val; b = (int)s; }
        }
        bi[k] = b;
        evals += (unsigned long long)n;
      }
    } else {
      // ---- warp-level pruning of the staged list; survivors -> s_list[warp] (ascending seed index)
      const double lo[3] = {(double)x0, (double)y0, (double)z0};
      const double hi[3] = {(double)(min(x0 + WT::W0, g.N[0]) - 1), (double)(min(y0 + WT::W1, g.N[1]) - 1),
                            (double)(z0 + nz - 1)};
      constexpr int kR = kCap / 32;
      const int rounds = (count + 31) >> 5;
      int kept = 0;
      if (MODE == 1) {
        // reference seed A: nearest to the sub-tile centre
        const int bc = nearest_record<ND>(s_rec, count, 0.5 * (lo[0] + hi[0]), 0.5 * (lo[1] + hi[1]), 0.5 * (lo[2] + hi[2]), lane);
        const SeedRec A = s_rec[bc];
        const double UA = box_max_d2<ND>(A, lo, hi);
        for (int r = 0; r < rounds; ++r) {
          const int c = 32 * r + lane;
          const bool keep = c < count && bisector_keep<ND>(A, s_rec[c], lo, hi, UA);
          const unsigned bal = __ballot_sync(kFull, keep);
          if (keep) s_list[warp][kept + __popc(bal & ((1u << lane) - 1u))] = (unsigned char)c;
          kept += __popc(bal);
        }
      } else {
        double lb[kR], ub = INFINITY;
#pragma unroll
        for (int r = 0; r < kR; ++r) {
          const int c = 32 * r + lane;
          lb[r] = INFINITY;
          if (c < count) {
            const SeedRec R = s_rec[c];
            double smin = 0.0, smax = 0.0;
            const double q[3] = {R.x, R.y, R.z};
#pragma unroll
            for (int a = 3 - ND; a < 3; ++a) {
              const double u = lo[a] - q[a], v = q[a] - hi[a];
              const double dmin = fmax(0.0, fmax(u, v)), dmax = fmax(-u, -v);
              smin += dmin * dmin;
              smax += dmax * dmax;
            }
            lb[r] = sqrt(smin) - R.w;
            ub = fmin(ub, sqrt(smax) - R.w);
          }
        }
        const double U = warp_min(ub);
#pragma unroll
        for (int r = 0; r < kR; ++r) {
          const int c = 32 * r + lane;
          const bool keep = c < count && may_win(lb[r], U);
          const unsigned bal = __ballot_sync(kFull, keep);
          if (keep) s_list[warp][kept + __popc(bal & ((1u << lane) - 1u))] = (unsigned char)c;
          kept += __popc(bal);
        }
      }
      __syncwarp();

      // ---- evaluation: column of kCol voxels per thread; (z_c - z_k)^2 shared by the warp, 4 candidates a batch
      const double X = (double)vx, Y = (double)vy;
      double bs[kCol];
      unsigned gap = 0xffffffffu;
#pragma unroll
      for (int k = 0; k < kCol; ++k) bs[k] = INFINITY;
      for (int b0 = 0; b0 < kept; b0 += 4) {
        {
          const int j = b0 + (lane >> 3);
          if (j < kept) {
            const double dz = __dsub_rn(s_rec[s_list[warp][j]].z, (double)(z0 + (lane & 7)));
            s_dz[warp][lane] = __dmul_rn(dz, dz);
          }
        }
        __syncwarp();
        const int nb = min(4, kept - b0);
        for (int j = 0; j < nb; ++j) {
          const int c = s_list[warp][b0 + j];
          const double2 Rxy = *reinterpret_cast<const double2*>(&s_rec[c]);
          const int idx = s_idx[c];
          const double dy = __dsub_rn(Rxy.y, Y);
          double sxy = __dmul_rn(dy, dy);
          if (ND == 3) {
            const double dx = __dsub_rn(Rxy.x, X);
            sxy = __dadd_rn(__dmul_rn(dx, dx), sxy);
          }
          double w = 0.0;
          if (MODE == 2) w = s_rec[c].w;
          const double2* dzp = reinterpret_cast<const double2*>(&s_dz[warp][8 * j]);
#pragma unroll
          for (int k2 = 0; k2 < kCol / 2; ++k2) {
            const double2 d2 = dzp[k2];
#pragma unroll
            for (int q = 0; q < 2; ++q) {
              const int k = 2 * k2 + q;
              const double ss = __dadd_rn(sxy, q ? d2.y : d2.x);
              if (MODE == 1) {
                // branch-free: the 8 voxels of the column are independent chains.  A strictly smaller sum within
                // 8 ulps of the running minimum is the only case where the reference's rounded roots could tie and
                // leave the voxel with the lower index.  Here only the smallest gap between the high words of a
                // replaced minimum and its replacement is kept (<= 1 for anything within 2^-20 relative, a superset
                // of the 8-ulp case); such a column is redone exactly below.
                const bool take = ss < bs[k];
                const unsigned dh = (unsigned)__double2hiint(bs[k]) - (unsigned)__double2hiint(ss);
                gap = take ? min(gap, dh) : gap;
                bs[k] = take ? ss : bs[k];
                bi[k] = take ? idx : bi[k];
              } else {
                const double val = __dsub_rn(__dsqrt_rn(ss), w);
                if (val < bs[k] || bi[k] < 0) { bs[k] = val; bi[k] = idx; }
              }
            }
          }
        }
        __syncwarp();
      }
      if (MODE == 1 && gap <= 1u) {  // rare (near ties): this thread's column again, roots compared where needed
        const Col8 ex = column_exact<ND>(s_rec, s_idx, s_list[warp], kept, X, Y, z0);
#pragma unroll
        for (int k = 0; k < kCol; ++k) bi[k] = ex.v[k];
      }
      evals += (unsigned long long)kept * (unsigned long long)__popc(live);
    }
  }
  if (col_in) {
    if (nz == kCol && (reinterpret_cast<uintptr_t>(labels + lin) & 15) == 0) {
      int4 a, b;
      a.x = live & 1u ? bi[0] : -1; a.y = live & 2u ? bi[1] : -1; a.z = live & 4u ? bi[2] : -1; a.w = live & 8u ? bi[3] : -1;
      b.x = live & 16u ? bi[4] : -1; b.y = live & 32u ? bi[5] : -1; b.z = live & 64u ? bi[6] : -1; b.w = live & 128u ? bi[7] : -1;
      int4* dst = reinterpret_cast<int4*>(labels + lin);
      dst[0] = a;
      dst[1] = b;
    } else {
      for (int k = 0; k < nz; ++k) labels[lin + k] = (live >> k & 1u) ? bi[k] : -1;
    }
  }
  }  // rounds
  if (stats) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) evals += __shfl_xor_sync(kFull, evals, off);
    if (lane == 0 && evals) atomicAdd(&stats[0], evals);
    if (tid == 0 && part == 0) {
      atomicAdd(&stats[1], 1ull);
      if (!staged) atomicAdd(&stats[2], 1ull);
    }
  }
}

}  // namespace sbx

#include "voronoi_cells.cuh"

using namespace sbx;

#include <cmath>

static cudaLaunchConfig_t pdl_config(dim3 grid, unsigned threads, cudaStream_t st, cudaLaunchAttribute* attr, bool pdl) {
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cfg;
}

static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
}

template <int ND>
static void launch_voronoi(const VorGeo& g, long long ctiles, long long tiles, const double* seeds, const double* weights,
                           int* clist, int* ccount, SeedRec* rec, int* ridx, int* rcount, const uint8_t* mask,
                           int32_t* labels, unsigned long long* stats, cudaStream_t st) {
  coarse_kernel<ND><<<(unsigned)ctiles, 256, 0, st>>>(g, seeds, weights, clist, ccount);
  static const int pdl = env_int("SBX_VOR_PDL", 1);
  cudaLaunchAttribute pdl_attr[1];
  auto config = [&](dim3 grid, unsigned threads) { return pdl_config(grid, threads, st, pdl_attr, pdl != 0); };
  {
    const cudaLaunchConfig_t cfg = config(dim3((unsigned)((tiles + 3) / 4)), 128);
    cudaLaunchKernelEx(&cfg, mid_kernel<ND>, g, (int)tiles, seeds, weights, (const int*)clist, (const int*)ccount, rec, ridx, rcount);
  }
  // (warps per CTA, warp tiles per warp).  Measured at 256^3 / 4096 seeds (whole labelling, r01 v6): (8,1) 0.097 ms,
  // (4,1) 0.089 ms, (4,2) 0.101, (2,8) 0.105 (v5: 0.124 / 0.110 / 0.118 / 0.126): small CTAs recycle their SM slot
  // sooner; more rounds per warp cost registers.
  auto launch_label = [&](auto mode_tag) {
    constexpr int MODE = decltype(mode_tag)::value;
    const dim3 lgrid((unsigned)(4 * g.nm[2]), (unsigned)g.nm[1], (unsigned)g.nm[0]);
    const cudaLaunchConfig_t cfg = config(lgrid, 128);
    cudaLaunchKernelEx(&cfg, label_kernel<ND, MODE, 4, 1>, g, seeds, weights, (const SeedRec*)rec, (const int*)ridx,
                       (const int*)rcount, (const int*)clist, (const int*)ccount, mask, labels, stats);
  };
  if (g.mode == 1) launch_label(std::integral_constant<int, 0>{});
  else if (!weights) launch_label(std::integral_constant<int, 1>{});
  else launch_label(std::integral_constant<int, 2>{});
}

// A plan owns the per-grid workspace, so that repeated labellings of one grid size cost three kernel launches and
// nothing else on the host.  Two workspaces, each allocated on first use:
//   cell path (additive, no weights -- the default of the reference's entry point): cell counters / starts and the
//   sorted image points (voronoi_cells.cuh);
//   tile path (weights, power mode): candidate lists of the coarse and mid tiles.
struct sbx_voronoi_plan {
  int device = 0, ndim = 3;
  VorGeo g{};
  long long tiles = 1, ctiles = 1;
  SeedRec* rec = nullptr;
  int *ridx = nullptr, *rcount = nullptr, *clist = nullptr, *ccount = nullptr;
  unsigned long long* stats = nullptr;
  // cell path
  int *cell_count = nullptr, *cell_start = nullptr;
  SeedRec* pts = nullptr;
  long long cell_cap = 0, pts_cap = 0;
  int cell_ncell = -1;
  int last_path = -1;  // 0 tile path (coarse + mid + label), 1 cell-list path
  // tile path on a plane range: full-grid scratch
  int32_t* tmp_labels = nullptr;
  uint8_t* tmp_mask = nullptr;
};

extern "C" int sbx_voronoi_plan_destroy(sbx_voronoi_plan* p) {
  if (!p) return SBX_OK;
  DeviceGuard guard(p->device);
  cudaFree(p->rec); cudaFree(p->ridx); cudaFree(p->rcount); cudaFree(p->clist); cudaFree(p->ccount); cudaFree(p->stats);
  cudaFree(p->cell_count); cudaFree(p->cell_start); cudaFree(p->pts);
  cudaFree(p->tmp_labels); cudaFree(p->tmp_mask);
  delete p;
  return SBX_OK;
}

extern "C" int sbx_voronoi_plan_create(sbx_voronoi_plan** out, int device, int ndim, const int64_t* N) {
  if (!out || !N) return fail(SBX_EINVAL, "NULL argument");
  *out = nullptr;
  if (ndim != 2 && ndim != 3) return fail(SBX_EINVAL, "ndim must be 2 or 3, got %d", ndim);
  int rc = require_sm100(device);
  if (rc) return rc;
  DeviceGuard guard(device);
  sbx_voronoi_plan* p = new (std::nothrow) sbx_voronoi_plan();
  if (!p) return fail(SBX_ENOMEM, "out of host memory");
  p->device = device;
  p->ndim = ndim;
  VorGeo& g = p->g;
  g.nd = ndim;
  const int mid3[3] = {16, 16, 16}, mid2[3] = {1, 64, 64};
  const int coarse3[3] = {32, 32, 32}, coarse2[3] = {1, 256, 256};
  const int off = 3 - ndim;  // internal axis of real axis 0
  for (int a = 0; a < 3; ++a) {
    g.N[a] = 1;
    g.box[a] = 1.0;
    g.M[a] = ndim == 3 ? mid3[a] : mid2[a];
    g.C[a] = ndim == 3 ? coarse3[a] : coarse2[a];
  }
  for (int a = 0; a < ndim; ++a) {
    if (N[a] < 1 || N[a] > (1 << 20)) { delete p; return fail(SBX_EINVAL, "N[%d]=%lld out of range", a, (long long)N[a]); }
    g.N[a + off] = (int)N[a];
  }
  for (int a = 0; a < 3; ++a) {
    g.nm[a] = (g.N[a] + g.M[a] - 1) / g.M[a];
    g.nc[a] = (g.N[a] + g.C[a] - 1) / g.C[a];
    p->tiles *= g.nm[a];
    p->ctiles *= g.nc[a];
  }
  if (p->tiles > 0x3fffffff || g.nm[0] > 65535 || g.nm[1] > 65535) {
    const long long t = p->tiles;
    delete p;
    return fail(SBX_EINVAL, "grid too large (%lld tiles)", t);
  }
  cudaError_t e = cudaMalloc((void**)&p->stats, 3 * sizeof(unsigned long long));
  if (e != cudaSuccess) {
    sbx_voronoi_plan_destroy(p);
    return fail(e == cudaErrorMemoryAllocation ? SBX_ENOMEM : SBX_ECUDA, "voronoi plan allocation: %s", cudaGetErrorString(e));
  }
  *out = p;
  return SBX_OK;
}

static int alloc_fail(cudaError_t e, const char* what) {
  return fail(e == cudaErrorMemoryAllocation ? SBX_ENOMEM : SBX_ECUDA, "voronoi %s allocation: %s", what, cudaGetErrorString(e));
}

static int ensure_tile_workspace(sbx_voronoi_plan* p) {
  if (p->rec) return SBX_OK;
  cudaError_t e = cudaMalloc((void**)&p->rec, (size_t)p->tiles * kCap * sizeof(SeedRec));
  if (e == cudaSuccess) e = cudaMalloc((void**)&p->ridx, (size_t)p->tiles * kCap * sizeof(int));
  if (e == cudaSuccess) e = cudaMalloc((void**)&p->rcount, (size_t)p->tiles * sizeof(int));
  if (e == cudaSuccess) e = cudaMalloc((void**)&p->clist, (size_t)p->ctiles * kCoarseCap * sizeof(int));
  if (e == cudaSuccess) e = cudaMalloc((void**)&p->ccount, (size_t)p->ctiles * sizeof(int));
  if (e != cudaSuccess) {
    cudaFree(p->rec); cudaFree(p->ridx); cudaFree(p->rcount); cudaFree(p->clist); cudaFree(p->ccount);
    p->rec = nullptr; p->ridx = p->rcount = p->clist = p->ccount = nullptr;
    return alloc_fail(e, "tile workspace");
  }
  return SBX_OK;
}

// Cell geometry of one labelling (host).  Mean seed spacing h = (voxels / S)^(1/d); cell edge = the power of two
// nearest to h (>= 8 voxels, the gather tile); first search radius R0 = 1.25 h; E cells of extension so that
// an ordinary search box never reaches the outermost (catch-all) layer.
static void cell_geometry(const sbx_voronoi_plan* p, const VorGeo& vg, long long S, int xlo, int xhi, CellGeo& c) {
  const int nd = p->ndim, off = 3 - nd;
  c.nd = nd;
  c.periodic = vg.periodic;
  c.S = S;
  c.xlo = xlo;
  c.xhi = xhi;
  double V = 1.0;
  for (int a = 0; a < 3; ++a) {
    c.N[a] = vg.N[a];
    c.box[a] = vg.box[a];
    if (a >= off) V *= (double)vg.N[a];
  }
  const double h = pow(V / (double)S, 1.0 / nd);
  static const int force_sh = env_int("SBX_VOR_CELL", 0);       // log2 of the cell edge, 0 = automatic
  static const int r0_pct = env_int("SBX_VOR_R0", 125);         // first search radius, percent of h
  int sh = 3;
  while ((double)(1 << sh) * 1.4142135623730951 < h && sh < 16) ++sh;
  if (force_sh >= 3 && force_sh <= 16) sh = force_sh;
  c.R0 = fmax(0.01 * r0_pct * h, 4.0);
  for (;;) {
    const double cs = (double)(1 << sh);
    const int E = vg.periodic ? (int)fmin(ceil(1.5 * c.R0 / cs) + 1.0, 6.0) : 1;
    long long ncell = 1;
    for (int a = 0; a < 3; ++a) {
      if (a < off) { c.sh[a] = 0; c.E[a] = 0; c.n[a] = 1; continue; }
      c.sh[a] = sh;
      c.E[a] = E;
      c.n[a] = ((c.N[a] + (1 << sh) - 1) >> sh) + 2 * E;
      ncell *= c.n[a];
    }
    c.nrow = c.n[0] * c.n[1];
    if (ncell <= (1ll << 26) && c.nrow <= kRowMax) { c.ncell = (int)ncell; break; }
    ++sh;
  }
  const int G[3] = {nd == 3 ? GTile<3>::G0 : GTile<2>::G0, nd == 3 ? GTile<3>::G1 : GTile<2>::G1, nd == 3 ? GTile<3>::G2 : GTile<2>::G2};
  for (int a = 0; a < 3; ++a) {
    const int extent = a == off ? xhi - xlo : c.N[a];
    c.nt[a] = (extent + G[a] - 1) / G[a];
  }
}

static int ensure_cell_workspace(sbx_voronoi_plan* p, const CellGeo& c, cudaStream_t st) {
  // counters = count[ncell] | cursor[ncell] | rowcount[nrow], zero between labellings (the label kernel resets them);
  // the layout depends on ncell, so a change of geometry re-zeroes the block
  const long long need_cnt = 2ll * c.ncell + c.nrow;
  if (need_cnt > p->cell_cap || c.ncell != p->cell_ncell) {
    if (need_cnt > p->cell_cap) {
      cudaFree(p->cell_count); cudaFree(p->cell_start);  // implicit device synchronisation: no kernel still reads them
      p->cell_count = p->cell_start = nullptr;
      p->cell_cap = 0;
      cudaError_t e = cudaMalloc((void**)&p->cell_count, (size_t)need_cnt * sizeof(int));
      if (e == cudaSuccess) e = cudaMalloc((void**)&p->cell_start, (size_t)(need_cnt + 1) * sizeof(int));
      if (e != cudaSuccess) return alloc_fail(e, "cell table");
      p->cell_cap = need_cnt;
    }
    SBX_CUDA(cudaMemsetAsync(p->cell_count, 0, (size_t)p->cell_cap * sizeof(int), st));
    p->cell_ncell = c.ncell;
  }
  const long long need = c.S * (c.periodic ? (c.nd == 3 ? 27 : 9) : 1);
  if (need > p->pts_cap) {
    cudaFree(p->pts);
    p->pts = nullptr;
    p->pts_cap = 0;
    cudaError_t e = cudaMalloc((void**)&p->pts, (size_t)need * sizeof(SeedRec));
    if (e != cudaSuccess) return alloc_fail(e, "image point");
    p->pts_cap = need;
  }
  return SBX_OK;
}

template <int ND, bool PER>
static void launch_cells(const CellGeo& c, const double* seeds, int* counters, int* start, SeedRec* pts,
                         const uint8_t* mask, int32_t* labels, unsigned long long* stats, cudaStream_t st) {
  static const int pdl = env_int("SBX_VOR_PDL", 1);
  const long long threads = c.S * (PER ? (ND == 3 ? 27 : 9) : 1);
  int *count = counters, *cursor = counters + c.ncell, *rowcount = counters + 2 * (size_t)c.ncell;  // one block: zeroed together
  const unsigned ctas = (unsigned)((threads + 255) / 256);
  cell_count_kernel<ND, PER><<<ctas, 256, 0, st>>>(c, seeds, count, rowcount);
  cudaLaunchAttribute attr[1];
  {
    const cudaLaunchConfig_t cfg = pdl_config(dim3(ctas), 256, st, attr, pdl != 0);
    cudaLaunchKernelEx(&cfg, cell_scatter_kernel<ND, PER>, c, seeds, (const int*)count, (const int*)rowcount, cursor, start, pts);
  }
  const cudaLaunchConfig_t cfg = pdl_config(dim3((unsigned)((c.nt[2] + 3) / 4), (unsigned)c.nt[1], (unsigned)c.nt[0]), 128, st, attr, pdl != 0);
  static const int minb = env_int("SBX_VOR_MINB", 6);
  if (minb == 4)
    cudaLaunchKernelEx(&cfg, label_cells_kernel<ND, 4>, c, (const SeedRec*)pts, (const int*)start, counters, 2 * c.ncell + c.nrow, mask, labels, stats);
  else if (minb == 5)
    cudaLaunchKernelEx(&cfg, label_cells_kernel<ND, 5>, c, (const SeedRec*)pts, (const int*)start, counters, 2 * c.ncell + c.nrow, mask, labels, stats);
  else if (minb == 8)
    cudaLaunchKernelEx(&cfg, label_cells_kernel<ND, 8>, c, (const SeedRec*)pts, (const int*)start, counters, 2 * c.ncell + c.nrow, mask, labels, stats);
  else
    cudaLaunchKernelEx(&cfg, label_cells_kernel<ND, 6>, c, (const SeedRec*)pts, (const int*)start, counters, 2 * c.ncell + c.nrow, mask, labels, stats);
}

// Labels of the planes [x_lo, x_hi) of the first grid axis.  `labels` (and `mask`) point to the first plane of that
// range: a rank that owns a slab of the grid passes its local buffers.  Seeds are the complete set on every rank.
extern "C" int sbx_voronoi_plan_execute_range(sbx_voronoi_plan* p, const double* seeds, int64_t S, const double* weights,
                                              int mode, const double* box_host, const uint8_t* mask, int32_t* labels,
                                              int64_t x_lo, int64_t x_hi, void* stream, int64_t* stats_host) {
  if (!p || !seeds || !labels) return fail(SBX_EINVAL, "NULL argument");
  if (S < 1 || S > 0x3fffffff) return fail(SBX_EINVAL, "seed count %lld out of range", (long long)S);
  if (mode != 0 && mode != 1) return fail(SBX_EINVAL, "mode must be 0 (additive) or 1 (power)");
  const int off = 3 - p->ndim;
  const int64_t N0 = p->g.N[off];
  if (x_lo < 0 || x_hi > N0 || x_lo >= x_hi) return fail(SBX_EINVAL, "plane range [%lld, %lld) outside [0, %lld)", (long long)x_lo, (long long)x_hi, (long long)N0);
  DeviceGuard guard(p->device);
  cudaStream_t st = (cudaStream_t)stream;
  VorGeo g = p->g;
  g.mode = mode;
  g.S = S;
  g.periodic = box_host ? 1 : 0;
  for (int a = 0; a < p->ndim && box_host; ++a) {
    if (!(box_host[a] > 0.0)) return fail(SBX_EINVAL, "box[%d] must be > 0", a);
    g.box[a + off] = box_host[a];
  }
  unsigned long long* stats = stats_host ? p->stats : nullptr;
  if (stats) SBX_CUDA(cudaMemsetAsync(stats, 0, 3 * sizeof(unsigned long long), st));
  // Path choice, from the measured sweep profiles/r02/voronoi_paths.jsonl (whole grids, B200; tiles vs cells):
  //   spacing h = (voxels / S)^(1/d)   256^3: h 32  75 vs  97 us | h 16  91 vs 122 us | h 10.7  187 vs 182 us | h 8  1.8 s vs 0.88 ms
  //                                    512^3: h 32 0.52 vs 0.58 ms | h 16 (32768 seeds) 1.58 vs 0.76 ms;  128^3 h 8: 39.9 vs 0.069 ms
  // The tile path stages at most kCap = 128 candidates per 16^3 mid tile and falls back to a full scan beyond (dense
  // seeds: catastrophic), and its coarse kernel scans ALL S seeds once per 32^3 tile (S / 32768 tests per voxel);
  // the cell path has neither limit but a higher fixed cost.  So: whole grids run the tile path when the seeds are
  // sparse (h >= 11 voxels) and few (S < 12000), the cell path otherwise; a plane range always runs the cell path,
  // which labels only the tiles of the range.  SBX_VOR_PATH=cells|tiles forces one (read once).
  static const int force = [] { const char* e = getenv("SBX_VOR_PATH"); return !e ? 0 : (e[0] == 'c' ? 1 : (e[0] == 't' ? 2 : 0)); }();
  const bool whole = x_lo == 0 && x_hi == N0;
  double vox = 1.0;
  for (int a = off; a < 3; ++a) vox *= (double)g.N[a];
  const double spacing = pow(vox / (double)S, 1.0 / p->ndim);
  const bool cells = mode == 0 && !weights && (long long)S * (g.periodic ? 27 : 1) <= (1ll << 27) &&
                     (force == 1 || (force == 0 && (!whole || S >= 12000 || spacing < 11.0)));
  if (cells) {
    CellGeo c{};
    cell_geometry(p, g, S, (int)x_lo, (int)x_hi, c);
    if (c.nt[0] > 65535 || c.nt[1] > 65535) return fail(SBX_EINVAL, "grid too large for the cell path");
    p->last_path = 1;
    int rc = ensure_cell_workspace(p, c, st);
    if (rc) return rc;
    if (p->ndim == 3) {
      if (g.periodic) launch_cells<3, true>(c, seeds, p->cell_count, p->cell_start, p->pts, mask, labels, stats, st);
      else launch_cells<3, false>(c, seeds, p->cell_count, p->cell_start, p->pts, mask, labels, stats, st);
    } else {
      if (g.periodic) launch_cells<2, true>(c, seeds, p->cell_count, p->cell_start, p->pts, mask, labels, stats, st);
      else launch_cells<2, false>(c, seeds, p->cell_count, p->cell_start, p->pts, mask, labels, stats, st);
    }
  } else {
    p->last_path = 0;
    int rc = ensure_tile_workspace(p);
    if (rc) return rc;
    const size_t plane = (size_t)(p->ndim == 3 ? (long long)g.N[1] * g.N[2] : (long long)g.N[2]);
    const bool full = whole;
    int32_t* lab = labels;
    const uint8_t* msk = mask;
    if (!full) {  // the tile path labels whole grids: run it into scratch, copy the planes out
      const size_t vox = plane * (size_t)N0;
      if (!p->tmp_labels) {
        cudaError_t e = cudaMalloc((void**)&p->tmp_labels, vox * sizeof(int32_t));
        if (e != cudaSuccess) return alloc_fail(e, "range scratch");
      }
      lab = p->tmp_labels;
      if (mask) {
        if (!p->tmp_mask) {
          cudaError_t e = cudaMalloc((void**)&p->tmp_mask, vox);
          if (e != cudaSuccess) return alloc_fail(e, "range scratch");
        }
        SBX_CUDA(cudaMemsetAsync(p->tmp_mask, 0, vox, st));
        SBX_CUDA(cudaMemcpyAsync(p->tmp_mask + plane * (size_t)x_lo, mask, plane * (size_t)(x_hi - x_lo), cudaMemcpyDeviceToDevice, st));
        msk = p->tmp_mask;
      }
    }
    if (p->ndim == 3)
      launch_voronoi<3>(g, p->ctiles, p->tiles, seeds, weights, p->clist, p->ccount, p->rec, p->ridx, p->rcount, msk, lab, stats, st);
    else
      launch_voronoi<2>(g, p->ctiles, p->tiles, seeds, weights, p->clist, p->ccount, p->rec, p->ridx, p->rcount, msk, lab, stats, st);
    if (!full)
      SBX_CUDA(cudaMemcpyAsync(labels, lab + plane * (size_t)x_lo, plane * (size_t)(x_hi - x_lo) * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
  }
  SBX_CUDA(cudaGetLastError());
  if (stats_host) {
    unsigned long long h[3];
    SBX_CUDA(cudaMemcpyAsync(h, stats, sizeof(h), cudaMemcpyDeviceToHost, st));
    SBX_CUDA(cudaStreamSynchronize(st));
    for (int i = 0; i < 3; ++i) stats_host[i] = (int64_t)h[i];
  }
  return SBX_OK;
}

extern "C" int sbx_voronoi_plan_last_path(const sbx_voronoi_plan* p) { return p ? p->last_path : SBX_EINVAL; }

extern "C" int sbx_voronoi_plan_execute(sbx_voronoi_plan* p, const double* seeds, int64_t S, const double* weights,
                                        int mode, const double* box_host, const uint8_t* mask, int32_t* labels,
                                        void* stream, int64_t* stats_host) {
  if (!p) return fail(SBX_EINVAL, "NULL argument");
  return sbx_voronoi_plan_execute_range(p, seeds, S, weights, mode, box_host, mask, labels, 0, p->g.N[3 - p->ndim], stream, stats_host);
}

// one-shot convenience: plan + execute + destroy (synchronises `stream` before the workspace is freed)
extern "C" int sbx_voronoi_label(int device, int ndim, const int64_t* N, const double* seeds, int64_t S,
                                 const double* weights, int mode, const double* box_host, const uint8_t* mask,
                                 int32_t* labels, void* stream, int64_t* stats_host) {
  sbx_voronoi_plan* p = nullptr;
  int rc = sbx_voronoi_plan_create(&p, device, ndim, N);
  if (rc) return rc;
  rc = sbx_voronoi_plan_execute(p, seeds, S, weights, mode, box_host, mask, labels, stream, stats_host);
  {
    DeviceGuard guard(device);
    cudaStreamSynchronize((cudaStream_t)stream);
  }
  sbx_voronoi_plan_destroy(p);
  return rc;
}

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free labelling (no counterpart in the reference; added for callers
// that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it).
// The cell-list path above needs three pieces of scratch: the sorted image points, the cell counters and the cell
// start table.  Here they live in ONE file-scope device array of fixed size, which every GPU that loads this library
// holds until the process exits:
//     image points  kNaMaxPoints x 32 B = 8 MiB   (seeds x enumerated periodic images: S x 3^d, S when non-periodic)
//     counters      3 x kNaMaxCells ints = 3 MiB   (count | cursor | rowcount, 2 ncell + nrow <= 3 ncell)
//     start table   (kNaMaxCells + 64) ints = 1 MiB
// In HOST code the array's name is only a stand-in for the symbol, not a device address: its address is taken once
// per device, with that device current, by cudaGetSymbolAddress, and that pointer is what the kernels get.
// All calls share the array and it keeps the last call's contents: the counters are cleared at the start of every
// call (cudaMemsetAsync on the caller's stream: no allocation), everything else is written before it is read.  Calls
// that overlap on one GPU (two streams, two threads) race on it -- issue them in stream order.
// ---------------------------------------------------------------------------------------------------------------
namespace {
constexpr long long kNaMaxPoints = 1ll << 18;
constexpr long long kNaMaxCells = 1ll << 18;
constexpr size_t kNaPtsBytes = (size_t)kNaMaxPoints * sizeof(SeedRec);
constexpr size_t kNaCntBytes = (size_t)3 * kNaMaxCells * sizeof(int);
constexpr size_t kNaStartBytes = (size_t)(kNaMaxCells + 64) * sizeof(int);
constexpr size_t kNaBytes = kNaPtsBytes + kNaCntBytes + kNaStartBytes;
constexpr int kNaMaxDevices = 64;
}  // namespace

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaBytes];

#include <mutex>

// device address of g_noalloc_scratch on `device` (which must be current); resolved once per device
static int noalloc_scratch(int device, unsigned char** out) {
  static std::mutex mu;
  static unsigned char* ptrs[kNaMaxDevices] = {};
  if (device < 0 || device >= kNaMaxDevices) return fail(SBX_EINVAL, "device %d outside the %d the static scratch table covers", device, kNaMaxDevices);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[device]) {
    unsigned char* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) return fail(SBX_ECUDA, "cudaGetSymbolAddress(static Voronoi scratch) on device %d: %s", device, cudaGetErrorString(e));
    ptrs[device] = p;
  }
  *out = ptrs[device];
  return SBX_OK;
}

extern "C" int sbx_voronoi_label_noalloc(int device, int ndim, const int64_t* N, const double* seeds, int64_t S,
                                         const double* box_host, const uint8_t* mask, int32_t* labels, int64_t x_lo,
                                         int64_t x_hi, void* stream) {
  if (!N || !seeds || !labels) return fail(SBX_EINVAL, "NULL argument");
  if (ndim != 2 && ndim != 3) return fail(SBX_EINVAL, "ndim must be 2 or 3, got %d", ndim);
  if (S < 1) return fail(SBX_EINVAL, "seed count %lld out of range", (long long)S);
  const long long images = box_host ? (ndim == 3 ? 27 : 9) : 1;
  if ((long long)S > kNaMaxPoints / images)
    return fail(SBX_ERANGE, "%lld seeds x %lld periodic images exceed the %lld image points of the static scratch; use sbx_voronoi_plan_execute_range",
                (long long)S, images, kNaMaxPoints);
  int rc = require_sm100(device);
  if (rc) return rc;
  // geometry exactly as sbx_voronoi_plan_create + sbx_voronoi_plan_execute_range set it up, on the stack
  sbx_voronoi_plan tmp;
  tmp.device = device;
  tmp.ndim = ndim;
  VorGeo& g = tmp.g;
  g = VorGeo{};
  g.nd = ndim;
  const int off = 3 - ndim;
  for (int a = 0; a < 3; ++a) { g.N[a] = 1; g.box[a] = 1.0; }
  for (int a = 0; a < ndim; ++a) {
    if (N[a] < 1 || N[a] > (1 << 20)) return fail(SBX_EINVAL, "N[%d]=%lld out of range", a, (long long)N[a]);
    g.N[a + off] = (int)N[a];
  }
  g.mode = 0;
  g.S = S;
  g.periodic = box_host ? 1 : 0;
  for (int a = 0; a < ndim && box_host; ++a) {
    if (!(box_host[a] > 0.0)) return fail(SBX_EINVAL, "box[%d] must be > 0", a);
    g.box[a + off] = box_host[a];
  }
  const int64_t N0 = g.N[off];
  if (x_lo < 0 || x_hi > N0 || x_lo >= x_hi) return fail(SBX_EINVAL, "plane range [%lld, %lld) outside [0, %lld)", (long long)x_lo, (long long)x_hi, (long long)N0);
  CellGeo c{};
  cell_geometry(&tmp, g, S, (int)x_lo, (int)x_hi, c);
  if (c.ncell > kNaMaxCells)
    return fail(SBX_ERANGE, "%d cells exceed the %lld cells of the static scratch; use sbx_voronoi_plan_execute_range", c.ncell, kNaMaxCells);
  if (c.nt[0] > 65535 || c.nt[1] > 65535) return fail(SBX_EINVAL, "grid too large for the cell path");
  DeviceGuard guard(device);
  unsigned char* base = nullptr;
  if ((rc = noalloc_scratch(device, &base))) return rc;
  SeedRec* pts = reinterpret_cast<SeedRec*>(base);
  int* counters = reinterpret_cast<int*>(base + kNaPtsBytes);
  int* start = reinterpret_cast<int*>(base + kNaPtsBytes + kNaCntBytes);
  cudaStream_t st = (cudaStream_t)stream;
  SBX_CUDA(cudaMemsetAsync(counters, 0, (size_t)(2ll * c.ncell + c.nrow) * sizeof(int), st));  // whatever the last call left
  if (ndim == 3) {
    if (g.periodic) launch_cells<3, true>(c, seeds, counters, start, pts, mask, labels, nullptr, st);
    else launch_cells<3, false>(c, seeds, counters, start, pts, mask, labels, nullptr, st);
  } else {
    if (g.periodic) launch_cells<2, true>(c, seeds, counters, start, pts, mask, labels, nullptr, st);
    else launch_cells<2, false>(c, seeds, counters, start, pts, mask, labels, nullptr, st);
  }
  SBX_CUDA(cudaGetLastError());
  return SBX_OK;
}
