// shell_table.h -- field-independent |q|-shell geometry (host side).
//
// Reproduces, without sorting N^3 rows, what the reference derives per call by
// argsort + merge loop (SimuBox/calculator/scatter.py:97-115) from the k^2 grid of
// Scatter.k_space (scatter.py:50-65).  See DESIGN.md "Shell table".
#pragma once
#include <cstdint>
#include <vector>

namespace sbx {

struct ShellTable {
  int ndim = 0;
  int64_t N[3] = {1, 1, 1};  // first ndim entries used
  double L[3] = {1, 1, 1};
  int cutoff = 30;
  double threshold = 1e-12;

  std::vector<double> kv[3];     // kv[d][t] = (2*pi*kvec(N_d)[t]/L_d)^2, t = 0..N_d-1  (scatter.py:60)
  std::vector<double> rep;       // k^2 of the first (smallest) member of each returned shell
  std::vector<int64_t> count;    // members the reference accumulates (1 for a truncated last row)
  bool truncated = false;        // last row is the one-member row of scatter.py:106-107
  int64_t designated = -1;       // full-spectrum index kept for a truncated last row
  // members[s]: ascending full-spectrum C-order indices f = (a*N1+b)*N2+c of the modes F[a,b,c]
  // binned into shell s (for a truncated last row: only `designated`).
  std::vector<std::vector<int64_t>> members;

  int64_t total() const { return N[0] * N[1] * N[2]; }
  int n_shells() const { return (int)rep.size(); }
  int64_t n_members() const {
    int64_t n = 0;
    for (auto& m : members) n += (int64_t)m.size();
    return n;
  }
  // k^2 the reference pairs with spectrum element f (quirk Q1: meshgrid 'xy' pairing)
  double k2_of(int64_t f) const;
  // F[a,b,c] -> index of its Hermitian mirror F[-a,-b,-c]
  int64_t mirror_of(int64_t f) const;
};

// Returns 0 or an SBX_E* code (message through sbx::fail).
int build_shell_table(ShellTable& t, int ndim, const int64_t* N, const double* L, int cutoff,
                      double threshold);

}  // namespace sbx
