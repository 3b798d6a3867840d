// shell_table.cpp -- best-first enumeration of the lowest |q| shells (host, fp64, no FMA).
#include "shell_table.h"

#include <algorithm>
#include <cmath>
#include <queue>

#include "common.h"

namespace sbx {

namespace {

// Scatter.k_vector (scatter.py:40-47): signed frequency of table slot t.
inline double signed_freq(int64_t t, int64_t n) { return t <= n / 2 ? (double)t : (double)(t - n); }

struct Node {
  double v;
  int32_t i, j, k;  // non-negative frequency indices 0..N_d/2 in "k^2 space"
  bool operator>(const Node& o) const { return v > o.v; }
};

}  // namespace

double ShellTable::k2_of(int64_t f) const {
  if (ndim == 1) return kv[0][f];
  const int64_t n2 = ndim == 3 ? N[2] : 1;
  const int64_t m = f / n2, c = f % n2;
  const double plane = kv[0][m % N[0]] + kv[1][m / N[0]];
  return ndim == 3 ? plane + kv[2][c] : plane;
}

int64_t ShellTable::mirror_of(int64_t f) const {
  int64_t idx[3] = {0, 0, 0};
  int64_t r = f;
  for (int d = ndim - 1; d >= 0; --d) {
    idx[d] = r % N[d];
    r /= N[d];
  }
  int64_t g = 0;
  for (int d = 0; d < ndim; ++d) g = g * N[d] + (idx[d] ? N[d] - idx[d] : 0);
  return g;
}

int build_shell_table(ShellTable& t, int ndim, const int64_t* N, const double* L, int cutoff,
                      double threshold) {
  if (ndim < 1 || ndim > 3) return fail(SBX_EINVAL, "ndim must be 1..3, got %d", ndim);
  if (cutoff < 0) return fail(SBX_EINVAL, "cutoff must be >= 0, got %d", cutoff);
  if (!(threshold >= 0.0)) return fail(SBX_EINVAL, "threshold must be >= 0");
  t = ShellTable();
  t.ndim = ndim;
  t.cutoff = cutoff;
  t.threshold = threshold;
  for (int d = 0; d < ndim; ++d) {
    if (N[d] < 2 || (N[d] & 1))  // scatter.py:57 (size-1 axes are dropped before this call)
      return fail(SBX_EINVAL, "grid sizes must be even and >= 2, got N[%d]=%lld", d, (long long)N[d]);
    if (!(L[d] > 0.0)) return fail(SBX_EINVAL, "box length L[%d] must be > 0", d);
    t.N[d] = N[d];
    t.L[d] = L[d];
    t.kv[d].resize(N[d]);
    for (int64_t s = 0; s < N[d]; ++s) {
      // (2 * np.pi * k_vec / l) ** 2, evaluated left to right in fp64
      volatile double x = 6.283185307179586 * signed_freq(s, N[d]);
      volatile double y = x / L[d];
      volatile double z = y * y;
      t.kv[d][s] = z;
    }
  }
  for (int d = ndim; d < 3; ++d) t.kv[d].assign(1, 0.0);

  const int64_t half[3] = {t.N[0] / 2, ndim > 1 ? t.N[1] / 2 : 0, ndim > 2 ? t.N[2] / 2 : 0};
  auto value = [&](int32_t i, int32_t j, int32_t k) {
    volatile double p = t.kv[0][i] + t.kv[1][j];
    volatile double q = p + t.kv[2][k];
    return (double)q;
  };

  std::priority_queue<Node, std::vector<Node>, std::greater<Node>> heap;
  heap.push({value(0, 0, 0), 0, 0, 0});
  std::vector<std::vector<int64_t>> mem;  // per shell, unsorted
  bool closing_last = false;              // collecting the rest of a truncated last shell
  std::vector<int64_t> tied_first;        // members of that shell whose k^2 equals its first value

  auto emit = [&](const Node& nd, std::vector<int64_t>& dst) {
    // expand +-i, +-j, +-k and map k^2-space (i,j,k) to the spectrum element the reference
    // pairs it with: meshgrid plane index m = j*N0 + i  ->  F[m / N1, m % N1, k]
    int64_t is[2] = {nd.i, t.N[0] - nd.i}, js[2] = {nd.j, t.N[1] - nd.j}, ks[2] = {nd.k, t.N[2] - nd.k};
    const int ni = (nd.i == 0 || nd.i == half[0]) ? 1 : 2;
    const int nj = (ndim < 2 || nd.j == 0 || nd.j == half[1]) ? 1 : 2;
    const int nk = (ndim < 3 || nd.k == 0 || nd.k == half[2]) ? 1 : 2;
    for (int a = 0; a < ni; ++a)
      for (int b = 0; b < nj; ++b)
        for (int c = 0; c < nk; ++c) {
          int64_t f;
          if (ndim == 1) f = is[a];
          else {
            const int64_t m = js[b] * t.N[0] + is[a];
            f = ndim == 3 ? m * t.N[2] + ks[c] : m;
          }
          dst.push_back(f);
        }
  };

  while (!heap.empty()) {
    Node nd = heap.top();
    heap.pop();
    bool opens = t.rep.empty() || std::fabs(t.rep.back() - std::fabs(nd.v)) > threshold;
    if (opens) {
      if (closing_last) break;  // first value beyond the truncated shell: done
      t.rep.push_back(nd.v);
      mem.emplace_back();
      if ((int)t.rep.size() > cutoff) closing_last = true;  // scatter.py:106-107
    }
    emit(nd, mem.back());
    if (closing_last && nd.v == t.rep.back()) emit(nd, tied_first);
    if (nd.i < half[0]) heap.push({value(nd.i + 1, nd.j, nd.k), nd.i + 1, nd.j, nd.k});
    if (nd.i == 0 && nd.j < half[1]) heap.push({value(0, nd.j + 1, nd.k), 0, nd.j + 1, nd.k});
    if (nd.i == 0 && nd.j == 0 && nd.k < half[2]) heap.push({value(0, 0, nd.k + 1), 0, 0, nd.k + 1});
  }

  const int ns = (int)t.rep.size();
  // Once row `cutoff` has been opened the merge loop breaks on its next iteration, so that
  // row keeps exactly one member whatever follows it.
  t.truncated = (ns == cutoff + 1);
  t.members.resize(ns);
  t.count.resize(ns);
  for (int s = 0; s < ns; ++s) {
    std::sort(mem[s].begin(), mem[s].end());
    if (t.truncated && s == ns - 1) {
      // the merge loop keeps the first sorted row: smallest k^2, and among exact ties the
      // lowest spectrum index (= what a stable sort keeps; the reference's sort is unstable, Q3)
      t.designated = *std::min_element(tied_first.begin(), tied_first.end());
      t.members[s].assign(1, t.designated);
    } else {
      t.members[s] = std::move(mem[s]);
    }
    t.count[s] = (int64_t)t.members[s].size();
  }
  return SBX_OK;
}

}  // namespace sbx
