// scatter_types.h -- device-side geometry structs shared by scatter.cu and slab.cu
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace sbx {

struct ShellDev {
  const double* rep;  // n_shells ascending k^2 of each shell's first member
  int n_shells;
  int truncated;       // last shell keeps only `designated`
  long long designated;
  double threshold;
  double vcut;  // conservative upper bound: k^2 > vcut can never be binned
};

struct GridDev {
  int ndim;
  int N0, N1, N2;      // unused trailing dims = 1
  int H;               // half-spectrum length of the last axis = N_last/2 + 1 (c = H - 1 is the Nyquist column)
  int Hs;              // 3D: columns STORED per row, c < Hs <= H (slab FFT mode exchanges only the columns that can hold a listed mode)
  long long nspec;     // stored elements per field
  long long nreal;     // real samples per field
  const double2* plane2;  // 3D: per (a,b) row {plane k^2 of F[a,b,.], plane k^2 of its mirror row}
  const double* plane;    // 1D/2D: k^2 per full-spectrum position (N0 or N0*N1 entries)
  const double* kvlast;   // 3D: kv2[c], c < H
  double factor;          // 1 / prod(N)  (scatter.py:84)
  int n1loc;              // 3D: rows of axis 1 held locally (= N1 except in the slab decomposition)
  int joff;               // 3D: first axis-1 index held locally (slab decomposition)
};

// Launch the fused |F|^2 / shell-histogram kernel (3D, no peak tail) on a half spectrum laid out as
// GridDev describes; shell_I (n_shells doubles) must be zeroed by the caller.  Defined in scatter.cu.
int launch_shell_histogram3(const double2* spec, const GridDev& g, const ShellDev& sh, double* shell_I,
                            int blocks, cudaStream_t st);

// ---- pruned-DFT slab helpers (defined in scatter.cu; used by slab.cu) --------------------------
// A plan created with SBX_SCATTER_PRUNED_FUSED can transform `planes` consecutive x-planes of ONE field that start
// at global plane index x_base: modesum (n_umodes complex sums over those planes, fixed summation order) is what
// the ranks add up (all-reduce) before scatter_slab_finish turns the totals into the n_shells intensities.
}  // namespace sbx
struct sbx_scatter_plan;
namespace sbx {
bool scatter_slab_supported(const sbx_scatter_plan* p, int planes);
int scatter_slab_n_umodes(const sbx_scatter_plan* p);
int scatter_slab_partial(sbx_scatter_plan* p, const double* slab, int planes, int x_base, double2* modesum, cudaStream_t st);
int scatter_slab_finish(sbx_scatter_plan* p, const double2* modesum, double* shell_I, cudaStream_t st);

}  // namespace sbx
