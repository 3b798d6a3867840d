// scatter.cu -- scatter plan + C ABI (see include/sbx.h).  cuFFT D2Z is a stated library call;
// everything after the transform is the hand-written kernels of scatter_kernels.cuh.
#include <cufft.h>

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <new>
#include <vector>

#include "common.h"
#include "pruned_dft.cuh"
#include "scatter_kernels.cuh"
#include "pruned_fused.cuh"
#include "shell_table.h"

namespace sbx {

thread_local char g_err[512] = "";
char* last_error_buf() { return g_err; }

int require_sm100(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    cudaGetLastError();
    return fail(SBX_ENODEVICE, "no CUDA device visible (this library has no CPU fallback)");
  }
  if (device < 0 || device >= n) return fail(SBX_ENODEVICE, "device %d out of range (%d visible)", device, n);
  int major = 0;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
  if (major != 10) return fail(SBX_ENODEVICE, "device %d has compute capability %d.x; sm_100 required", device, major);
  return SBX_OK;
}

#define SBX_CUFFT(call)                                                                       \
  do {                                                                                        \
    cufftResult r_ = (call);                                                                  \
    if (r_ != CUFFT_SUCCESS) return ::sbx::fail(SBX_ECUFFT, "%s:%d %s -> cufftResult %d", __FILE__, __LINE__, #call, (int)r_); \
  } while (0)

template <class T>
static int upload(T** dptr, const std::vector<T>& h) {
  *dptr = nullptr;
  if (h.empty()) return SBX_OK;
  SBX_CUDA(cudaMalloc((void**)dptr, h.size() * sizeof(T)));
  SBX_CUDA(cudaMemcpy(*dptr, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
  return SBX_OK;
}

int launch_shell_histogram3(const double2* spec, const GridDev& g, const ShellDev& sh, double* shell_I, int blocks,
                            cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(shell_histogram_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    attr_set = true;
  }
  PeakParams pp{};
  shell_histogram_kernel<3><<<dim3(blocks, 1), 256, (size_t)2 * sh.n_shells * sizeof(double), st>>>(
      spec, g, sh, shell_I, nullptr, 0, nullptr, pp, nullptr, nullptr, nullptr);
  SBX_CUDA(cudaGetLastError());
  return SBX_OK;
}

}  // namespace sbx

using namespace sbx;

struct sbx_scatter_plan {
  int device = 0;
  int path = SBX_SCATTER_FFT_DENSE;
  int max_batch = 1;
  int chunk = 1;  // fields per cuFFT launch (spectrum workspace = chunk fields)
  ShellTable tab;
  GridDev grid{};
  ShellDev shell{};
  std::vector<double> q_host;
  // device tables
  double* d_rep = nullptr;
  double* d_q = nullptr;
  double2* d_plane2 = nullptr;
  double* d_plane = nullptr;
  double* d_kvlast = nullptr;
  long long* d_elem = nullptr;
  int* d_shell_off = nullptr;
  unsigned int* d_tickets = nullptr;
  double2* d_spec = nullptr;
  // pruned-DFT path
  PrunedGeo pruned{};
  double* d_tw2 = nullptr;
  double* d_tw2f = nullptr;   // full-K stage-1 fragments (variant B), only when N2 % 8 == 0
  int pruned_variant = 0;     // 0 = paired z / N2-z (variant A), 1 = full K with 16-byte loads (variant B)
  int pruned_ctas_per_sm = 3;
  double* d_tw1 = nullptr;
  double2* d_tw0 = nullptr;
  int4* d_umodes = nullptr;
  long long* d_memb_u = nullptr;
  double* d_planes = nullptr;
  int n_umodes = 0;
  bool pruned_ok = false;
  // fused single-kernel variant (pruned_fused.cuh)
  // debugging switches, read ONCE at plan creation (SBX_BULK=0: LDG row-blocked kernel instead of the bulk-staged one;
  // SBX_FUSED_BLOCKED=1: row-blocked kernels on grids that would run the whole-plane kernels)
  bool env_bulk_off = false, env_force_blocked = false;
  bool fused_ok = false;    // whole planes fit the shared-memory scheme of pruned_fused_kernel / pruned_ring_kernel
  bool modes_ok = false;    // every listed mode lies in the 8 x 8 (k_y, k_z) tile: tables below are built
  bool blocked_ok = false;  // pruned_blocked_kernel applies (any N1 = yb_rows * yblocks)
  int yb_rows = 0, yblocks = 1;
  double2* d_modesum = nullptr;  // slab path: this rank's / the reduced mode sums
  const char* last_kernel = "";  // dominant hand-written kernel of the last execute (bench roofline)
  double* d_twf = nullptr;
  double2* d_partial = nullptr;
  unsigned* d_ctrl = nullptr;
  int fused_xc = 1;
  size_t partial_units = 0;  // capacity of d_partial in work units (all fields of one launch)
  cudaStream_t s_aux = nullptr;
  cudaEvent_t ev_planes[2] = {nullptr, nullptr}, ev_finish[2] = {nullptr, nullptr};
  std::map<int, cufftHandle> fft;  // batch -> plan
  int blocks_per_field = 1;
  int sm_count = 148;
  // host-streaming state (sbx_scatter_execute_host)
  cudaStream_t s_copy = nullptr, s_comp = nullptr;
  double* d_stage[2] = {nullptr, nullptr};
  double* d_out[2] = {nullptr, nullptr};
  int32_t* d_pidx[2] = {nullptr, nullptr};
  int32_t* d_npk[2] = {nullptr, nullptr};
  int stage_max_peaks = 0;
  cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr}, ev_out[2] = {nullptr, nullptr};
  // live profiling (sbx_scatter_plan_set_profiling): event triples around transform / fused kernel
  bool profiling = false;
  std::vector<cudaEvent_t> prof_events;  // 3 per chunk launch: start, after transform, after fused kernel
  size_t prof_used = 0;
  long long own_launches = 0, fft_calls = 0;
};

static cudaEvent_t prof_event(sbx_scatter_plan* p, cudaStream_t st) {
  if (p->prof_used == p->prof_events.size()) {
    cudaEvent_t e;
    cudaEventCreate(&e);
    p->prof_events.push_back(e);
  }
  cudaEvent_t e = p->prof_events[p->prof_used++];
  cudaEventRecord(e, st);
  return e;
}

static int get_fft(sbx_scatter_plan* p, int batch, cufftHandle* out) {
  auto it = p->fft.find(batch);
  if (it != p->fft.end()) { *out = it->second; return SBX_OK; }
  cufftHandle h;
  int n[3];
  for (int d = 0; d < p->tab.ndim; ++d) n[d] = (int)p->tab.N[d];
  SBX_CUFFT(cufftCreate(&h));
  size_t ws = 0;
  SBX_CUFFT(cufftMakePlanMany(h, p->tab.ndim, n, nullptr, 1, 0, nullptr, 1, 0, CUFFT_D2Z, batch, &ws));
  p->fft[batch] = h;
  *out = h;
  return SBX_OK;
}

static void fill_q(sbx_scatter_plan* p) {
  const ShellTable& t = p->tab;
  p->q_host.resize(t.n_shells());
  for (int s = 0; s < t.n_shells(); ++s) p->q_host[s] = std::sqrt(t.rep[s]);  // scatter.py:118
}

static int build_device_tables(sbx_scatter_plan* p) {
  const ShellTable& t = p->tab;
  GridDev& g = p->grid;
  g.ndim = t.ndim;
  g.N0 = (int)t.N[0]; g.N1 = (int)t.N[1]; g.N2 = (int)t.N[2];
  const int64_t nlast = t.N[t.ndim - 1];
  g.H = (int)(nlast / 2 + 1);
  g.Hs = g.H;
  g.nreal = t.total();
  g.nspec = g.nreal / nlast * g.H;
  g.factor = 1.0 / (double)g.nreal;
  g.n1loc = g.N1;
  g.joff = 0;
  int rc;
  if (t.ndim == 3) {
    const int64_t rows = t.N[0] * t.N[1];
    std::vector<double2> p2(rows);
    for (int64_t row = 0; row < rows; ++row) {
      const int64_t a = row / t.N[1], b = row % t.N[1];
      const int64_t mrow = (a ? t.N[0] - a : 0) * t.N[1] + (b ? t.N[1] - b : 0);
      volatile double x = t.kv[0][row % t.N[0]] + t.kv[1][row / t.N[0]];
      volatile double y = t.kv[0][mrow % t.N[0]] + t.kv[1][mrow / t.N[0]];
      p2[row] = make_double2(x, y);
    }
    if ((rc = upload(&p->d_plane2, p2))) return rc;
    std::vector<double> kl(t.kv[2].begin(), t.kv[2].begin() + g.H);
    if ((rc = upload(&p->d_kvlast, kl))) return rc;
  } else {
    std::vector<double> pl(t.total());
    for (int64_t f = 0; f < t.total(); ++f) pl[f] = t.k2_of(f);
    if ((rc = upload(&p->d_plane, pl))) return rc;
  }
  g.plane2 = p->d_plane2; g.plane = p->d_plane; g.kvlast = p->d_kvlast;

  if ((rc = upload(&p->d_rep, t.rep))) return rc;
  if ((rc = upload(&p->d_q, p->q_host))) return rc;
  ShellDev& sh = p->shell;
  sh.rep = p->d_rep;
  sh.n_shells = t.n_shells();
  sh.truncated = t.truncated ? 1 : 0;
  sh.designated = t.designated;
  sh.threshold = t.threshold;
  // Anything above vcut fails |rep_last - v| <= threshold with margin (pre-filter only; the exact
  // reference comparison is repeated in shell_of()).
  const double top = t.rep.back() + t.threshold;
  sh.vcut = top + 8.0 * std::fabs(top) * 2.220446049250313e-16 + 1e-300;

  // gather list: members -> stored element (direct, or the Hermitian mirror if c > N_last/2)
  std::vector<long long> elem;
  std::vector<int> off(1, 0);
  elem.reserve((size_t)t.n_members());
  for (int s = 0; s < t.n_shells(); ++s) {
    for (int64_t f : t.members[s]) {
      int64_t c = f % nlast;
      int64_t src = f;
      if (c > nlast / 2) { src = t.mirror_of(f); c = src % nlast; }
      elem.push_back((src / nlast) * g.H + c);
    }
    off.push_back((int)elem.size());
  }
  if ((rc = upload(&p->d_elem, elem))) return rc;
  if ((rc = upload(&p->d_shell_off, off))) return rc;
  return SBX_OK;
}

// ---- pruned-DFT tables: which modes, and the twiddle B-fragments of pruned_dft.cuh --------
static double cos2pi(long long j, long long n) { return (double)cosl(6.283185307179586476925286766559L * (long double)(j % n) / (long double)n); }
static double sin2pi(long long j, long long n) { return (double)sinl(6.283185307179586476925286766559L * (long double)(j % n) / (long double)n); }

static int build_pruned_tables(sbx_scatter_plan* p) {
  const ShellTable& t = p->tab;
  p->pruned_ok = false;
  if (t.ndim != 3) return SBX_OK;
  const int64_t N0 = t.N[0], N1 = t.N[1], N2 = t.N[2];
  if (N1 % 8 || N1 < 16 || N2 < 16) return SBX_OK;
  const bool planes_ok = pruned_smem_bytes((int)N1, (int)N2) <= 100 * 1024;  // two-kernel path only
  if (t.n_members() > (int64_t)1 << 16) return SBX_OK;
  std::map<std::array<int64_t, 3>, int> index;
  std::vector<int4> umodes;
  std::vector<long long> memb;
  for (int s = 0; s < t.n_shells(); ++s)
    for (int64_t f : t.members[s]) {
      int64_t c = f % N2, b = (f / N2) % N1, a = f / (N2 * N1);
      int64_t fa = a <= N0 / 2 ? a : a - N0, fb = b <= N1 / 2 ? b : b - N1, fc = c <= N2 / 2 ? c : c - N2;
      if (fc < 0 || (fc == 0 && (fb < 0 || (fb == 0 && fa < 0)))) { fa = -fa; fb = -fb; fc = -fc; }  // Hermitian mirror
      if (fc > 7 || fc >= N2 / 2 || fb > 7 || fb < -7 || 2 * (fb < 0 ? -fb : fb) >= N1) return SBX_OK;  // outside the 8x8 tile
      std::array<int64_t, 3> key{fa, fb, fc};
      auto it = index.find(key);
      if (it == index.end()) {
        it = index.emplace(key, (int)umodes.size()).first;
        umodes.push_back(make_int4((int)(((fa % N0) + N0) % N0), (int)fb, (int)fc, 0));
      }
      memb.push_back(it->second);
    }
  if (umodes.size() > 8192) return SBX_OK;
  PrunedGeo& g = p->pruned;
  g.N0 = (int)N0; g.N1 = (int)N1; g.N2 = (int)N2;
  g.ks2 = (int)((N2 / 2 + 1 + 3) / 4);
  g.ks1 = (int)(N1 / 4);
  g.gstride = (int)N1 + 4;
  std::vector<double> tw2((size_t)g.ks2 * 64, 0.0), tw1((size_t)2 * g.ks1 * 32, 0.0);
  for (int s = 0; s < g.ks2; ++s)
    for (int lane = 0; lane < 32; ++lane) {
      const int64_t tt = 4 * s + (lane & 3), n = lane >> 2;
      if (tt > N2 / 2) continue;
      tw2[(size_t)s * 64 + lane] = cos2pi(n * tt, N2);
      tw2[(size_t)s * 64 + 32 + lane] = sin2pi(n * tt, N2);
    }
  for (int s = 0; s < g.ks1; ++s)
    for (int lane = 0; lane < 32; ++lane) {
      const int64_t y = 4 * s + (lane & 3), b = lane >> 2;
      tw1[(size_t)s * 32 + lane] = cos2pi(b * y, N1);
      tw1[((size_t)g.ks1 + s) * 32 + lane] = sin2pi(b * y, N1);
    }
  std::vector<double2> tw0(N0);
  for (int64_t j = 0; j < N0; ++j) tw0[j] = make_double2(cos2pi(j, N0), sin2pi(j, N0));
  int rc;
  if ((rc = upload(&p->d_tw2, tw2))) return rc;
  if ((rc = upload(&p->d_tw1, tw1))) return rc;
  if (N2 % 8 == 0) {
    std::vector<double> tw2f((size_t)(N2 / 8) * 128);
    for (int64_t s = 0; s < N2 / 8; ++s)
      for (int lane = 0; lane < 32; ++lane) {
        const int64_t n = lane >> 2, te = 8 * s + 2 * (lane & 3), to = te + 1;
        tw2f[(size_t)s * 128 + lane] = cos2pi(n * te, N2);
        tw2f[(size_t)s * 128 + 32 + lane] = sin2pi(n * te, N2);
        tw2f[(size_t)s * 128 + 64 + lane] = cos2pi(n * to, N2);
        tw2f[(size_t)s * 128 + 96 + lane] = sin2pi(n * to, N2);
      }
    if ((rc = upload(&p->d_tw2f, tw2f))) return rc;
    p->pruned_variant = 1;
  }
  if (const char* e = getenv("SBX_PRUNED_VARIANT")) p->pruned_variant = (atoi(e) == 1 && p->d_tw2f) ? 1 : 0;
  if (const char* e = getenv("SBX_PRUNED_CTAS")) p->pruned_ctas_per_sm = std::max(1, std::min(3, atoi(e)));
  if (N2 % 16 == 0 && N1 % 8 == 0 && umodes.size() <= (size_t)256 * kFusedModesPerThread) {
    // fold fragments (pruned_fused.cuh): step s covers z = 8 s .. 8 s + 7 of both halves of a row
    const int64_t ds = N2 / 16;
    std::vector<double> twf((size_t)ds * 128);
    for (int64_t s = 0; s < ds; ++s)
      for (int lane = 0; lane < 32; ++lane) {
        const int64_t k = lane & 3, n = lane >> 2, te = 8 * s + 2 * k, to = te + 1;
        const int64_t ce = n < 4 ? 2 * n : 2 * (n - 4), co = ce + 1;  // even / odd frequency of column n
        const bool is_cos = n < 4;
        twf[(size_t)s * 128 + lane] = is_cos ? cos2pi(ce * te, N2) : sin2pi(ce * te, N2);
        twf[(size_t)s * 128 + 32 + lane] = is_cos ? cos2pi(ce * to, N2) : sin2pi(ce * to, N2);
        twf[(size_t)s * 128 + 64 + lane] = is_cos ? cos2pi(co * te, N2) : sin2pi(co * te, N2);
        twf[(size_t)s * 128 + 96 + lane] = is_cos ? cos2pi(co * to, N2) : sin2pi(co * to, N2);
      }
    if ((rc = upload(&p->d_twf, twf))) return rc;
    p->fused_xc = N0 % 4 == 0 ? 4 : (N0 % 2 == 0 ? 2 : 1);
    if (const char* e = getenv("SBX_FUSED_XC")) {
      const int v = atoi(e);
      if (v >= 1 && N0 % v == 0) p->fused_xc = v;
    }
    const size_t work = fused_work_doubles((int)N1, t.n_shells(), (int)umodes.size(), 2000, true);
    p->fused_ok = fused_smem_bytes((int)N0, (int)N1, (int)N2, work) <= 200 * 1024;
    // blocked kernel: blocks of 128 rows (or the whole plane when N1 is small)
    p->yb_rows = N1 % 128 == 0 ? 128 : (N1 <= 256 ? (int)N1 : 0);
    if (p->yb_rows) {
      p->yblocks = (int)(N1 / p->yb_rows);
      p->blocked_ok = blocked_smem_bytes((int)N0, (int)N2, blocked_work_doubles(p->yb_rows, t.n_shells(), (int)umodes.size(), 2000, true)) <= 200 * 1024;
    }
  }
  if ((rc = upload(&p->d_tw0, tw0))) return rc;
  if ((rc = upload(&p->d_umodes, umodes))) return rc;
  if ((rc = upload(&p->d_memb_u, memb))) return rc;
  g.tw2 = p->d_tw2;
  g.tw1 = p->d_tw1;
  p->n_umodes = (int)umodes.size();
  p->modes_ok = true;
  p->pruned_ok = planes_ok;
  return SBX_OK;
}

extern "C" int sbx_version(void) { return SBX_VERSION; }
extern "C" const char* sbx_last_error(void) { return g_err; }

extern "C" int sbx_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  int ok = 0;
  for (int d = 0; d < n; ++d) {
    int major = 0;
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, d);
    ok += (major == 10);
  }
  return ok;
}

// ------------------------------------------------------------------------------ plan life cycle
static void free_plan(sbx_scatter_plan* p) {
  if (!p) return;
  if (p->device == SBX_DEVICE_NONE) { delete p; return; }
  DeviceGuard guard(p->device);
  for (auto& kv : p->fft) cufftDestroy(kv.second);
  for (cudaEvent_t e : p->prof_events) cudaEventDestroy(e);
  cudaFree(p->d_rep); cudaFree(p->d_q); cudaFree(p->d_plane2); cudaFree(p->d_plane);
  cudaFree(p->d_kvlast); cudaFree(p->d_elem); cudaFree(p->d_shell_off); cudaFree(p->d_tickets);
  cudaFree(p->d_spec);
  cudaFree(p->d_tw2); cudaFree(p->d_tw2f); cudaFree(p->d_tw1); cudaFree(p->d_tw0); cudaFree(p->d_umodes); cudaFree(p->d_memb_u);
  cudaFree(p->d_planes);
  cudaFree(p->d_twf); cudaFree(p->d_partial); cudaFree(p->d_ctrl); cudaFree(p->d_modesum);
  for (int i = 0; i < 2; ++i) {
    cudaFree(p->d_stage[i]); cudaFree(p->d_out[i]); cudaFree(p->d_pidx[i]); cudaFree(p->d_npk[i]);
    if (p->ev_in[i]) cudaEventDestroy(p->ev_in[i]);
    if (p->ev_done[i]) cudaEventDestroy(p->ev_done[i]);
    if (p->ev_out[i]) cudaEventDestroy(p->ev_out[i]);
  }
  for (int i = 0; i < 2; ++i) {
    if (p->ev_planes[i]) cudaEventDestroy(p->ev_planes[i]);
    if (p->ev_finish[i]) cudaEventDestroy(p->ev_finish[i]);
  }
  if (p->s_aux) cudaStreamDestroy(p->s_aux);
  if (p->s_copy) cudaStreamDestroy(p->s_copy);
  if (p->s_comp) cudaStreamDestroy(p->s_comp);
  delete p;
}

static const size_t kMaxDynSmem = 200 * 1024;
static const size_t kMaxTmaSmem = 226 * 1024;

extern "C" int sbx_scatter_plan_create_ex(sbx_scatter_plan** out, int device, int ndim, const int64_t* N,
                                          const double* L, int cutoff, double threshold, int max_batch,
                                          int path) {
  if (!out || !N || !L) return fail(SBX_EINVAL, "NULL argument");
  *out = nullptr;
  if (max_batch < 1) return fail(SBX_EINVAL, "max_batch must be >= 1");
  int rc;
  if (device == SBX_DEVICE_NONE) {
    // geometry-only plan: shell table on the host, no CUDA call; *_execute refuses to run
    sbx_scatter_plan* hp = new (std::nothrow) sbx_scatter_plan();
    if (!hp) return fail(SBX_ENOMEM, "out of host memory");
    hp->device = SBX_DEVICE_NONE;
    if ((rc = build_shell_table(hp->tab, ndim, N, L, cutoff, threshold))) { delete hp; return rc; }
    fill_q(hp);
    hp->path = SBX_SCATTER_AUTO;
    *out = hp;
    return SBX_OK;
  }
  if ((rc = require_sm100(device))) return rc;
  DeviceGuard guard(device);
  sbx_scatter_plan* p = new (std::nothrow) sbx_scatter_plan();
  if (!p) return fail(SBX_ENOMEM, "out of host memory");
  p->device = device;
  p->max_batch = max_batch;
  { const char* e = getenv("SBX_BULK"); p->env_bulk_off = e && atoi(e) == 0; }
  { const char* e = getenv("SBX_FUSED_BLOCKED"); p->env_force_blocked = e && atoi(e) != 0; }
  if ((rc = build_shell_table(p->tab, ndim, N, L, cutoff, threshold))) { free_plan(p); return rc; }
  fill_q(p);
  if ((rc = build_device_tables(p))) { free_plan(p); return rc; }
  cudaDeviceGetAttribute(&p->sm_count, cudaDevAttrMultiProcessorCount, device);

  const GridDev& g = p->grid;
  if ((rc = build_pruned_tables(p))) { free_plan(p); return rc; }
  if (path == SBX_SCATTER_AUTO) {
    if (p->modes_ok && (p->fused_ok || p->blocked_ok) && !getenv("SBX_NO_PRUNED") && !getenv("SBX_NO_FUSED")) path = SBX_SCATTER_PRUNED_FUSED;
    else if (p->pruned_ok && !getenv("SBX_NO_PRUNED")) path = SBX_SCATTER_PRUNED;
    else path = (p->tab.n_members() * 4 <= g.nspec) ? SBX_SCATTER_FFT_GATHER : SBX_SCATTER_FFT_DENSE;
  }
  if (path == SBX_SCATTER_PRUNED && !p->pruned_ok) {
    free_plan(p);
    return fail(SBX_EINVAL, "pruned-DFT path needs a 3D grid (N1 %% 8 == 0, N1,N2 >= 16) whose shell modes have |k_y|,k_z <= 7");
  }
  if (path == SBX_SCATTER_PRUNED_FUSED && !(p->modes_ok && (p->fused_ok || p->blocked_ok))) {
    free_plan(p);
    return fail(SBX_EINVAL, "fused pruned-DFT path needs the pruned path's conditions plus N2 %% 16 == 0 and at most %d listed modes", 256 * kFusedModesPerThread);
  }
  if (path != SBX_SCATTER_FFT_DENSE && path != SBX_SCATTER_FFT_GATHER && path != SBX_SCATTER_PRUNED && path != SBX_SCATTER_PRUNED_FUSED) {
    free_plan(p);
    return fail(SBX_EINVAL, "unknown scatter path %d", path);
  }
  p->path = path;
  if ((size_t)p->tab.n_shells() * 3 * sizeof(double) > kMaxDynSmem) {
    free_plan(p);
    return fail(SBX_EINVAL, "%d shells exceed the shared-memory shell table (max %zu)", p->tab.n_shells(),
                kMaxDynSmem / 24);
  }

  // fields per cuFFT launch and per histogram launch: a batched plan over as many fields as 1 GiB of half spectra holds
  // (at most 64).  r01 sized this to the L2 (one 128^3 field per cuFFT exec, one histogram launch per field: 512 launches
  // per 256 fields, each ending in the serial peak search of its last CTA); the row-wise histogram no longer reads
  // the rows beyond the shell table, so L2 residency of the spectrum buys nothing and batching buys the launches back.
  size_t chunk = ((size_t)1 << 30) / ((size_t)g.nspec * 16);
  chunk = std::min<size_t>(std::max<size_t>(1, chunk), 64);
  if (path == SBX_SCATTER_FFT_GATHER) chunk = (size_t)max_batch;  // one CTA per field: launch many fields at once
  if (const char* e = getenv("SBX_FFT_CHUNK")) chunk = std::max(1, atoi(e));
  p->chunk = (int)std::min<size_t>(chunk, (size_t)max_batch);
  if (path == SBX_SCATTER_PRUNED || path == SBX_SCATTER_PRUNED_FUSED) p->chunk = max_batch;
  const long long want = (4LL * p->sm_count + p->chunk - 1) / p->chunk;
  long long bpf = std::max<long long>((g.nspec + 16383) / 16384, want);
  bpf = std::min<long long>(bpf, std::max<long long>(1, g.nspec / 256));
  p->blocks_per_field = (int)std::min<long long>(bpf, 65535);

  cudaError_t e = cudaSuccess;
  if (path == SBX_SCATTER_PRUNED) {
    e = cudaMalloc((void**)&p->d_planes, (size_t)2 * p->chunk * g.N0 * 256 * sizeof(double));
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&p->s_aux, cudaStreamNonBlocking);
    for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
      e = cudaEventCreateWithFlags(&p->ev_planes[i], cudaEventDisableTiming);
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_finish[i], cudaEventDisableTiming);
    }
  }
  else if (path == SBX_SCATTER_PRUNED_FUSED) {
    // giant planes (>= 512^2: the row-blocked kernels) get room for one-plane units, so that a single field or a slab of
    // it still splits into several units per SM (blocked_unit_planes)
    const bool giant = p->blocked_ok && (long long)g.N1 * g.N2 >= 512LL * 512LL;
    const size_t parts = (size_t)g.N0 / (giant ? 1 : p->fused_xc) * (p->blocked_ok ? p->yblocks : 1);
    p->partial_units = (size_t)p->chunk * parts;
    e = cudaMalloc((void**)&p->d_partial, (size_t)p->chunk * parts * std::max(p->n_umodes, 1) * sizeof(double2));
    if (e == cudaSuccess) e = cudaMalloc((void**)&p->d_modesum, (size_t)std::max(p->n_umodes, 1) * sizeof(double2));
    if (e == cudaSuccess) e = cudaMalloc((void**)&p->d_ctrl, ((size_t)p->chunk + 1) * sizeof(unsigned));
  }
  else e = cudaMalloc((void**)&p->d_spec, (size_t)p->chunk * g.nspec * sizeof(double2));
  if (e == cudaSuccess) e = cudaMalloc((void**)&p->d_tickets, (size_t)max_batch * sizeof(unsigned));
  if (e == cudaSuccess) e = cudaMemset(p->d_tickets, 0, (size_t)max_batch * sizeof(unsigned));
  if (e != cudaSuccess) {
    free_plan(p);
    return fail(e == cudaErrorMemoryAllocation ? SBX_ENOMEM : SBX_ECUDA, "plan allocation: %s", cudaGetErrorString(e));
  }
  cufftHandle h;
  if (path != SBX_SCATTER_PRUNED && path != SBX_SCATTER_PRUNED_FUSED && (rc = get_fft(p, p->chunk, &h))) { free_plan(p); return rc; }
  cudaFuncSetAttribute(pruned_planes_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
  cudaFuncSetAttribute(pruned_planes_fullk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
  cudaFuncSetAttribute(pruned_finish_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
  cudaFuncSetAttribute(pruned_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
  cudaFuncSetAttribute(pruned_blocked_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
  cudaFuncSetAttribute(pruned_blocked_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
  cudaFuncSetAttribute(fused_finish_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
  cudaFuncSetAttribute(pruned_ring_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxTmaSmem);
  cudaFuncSetAttribute(pruned_ring_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxTmaSmem);
  cudaFuncSetAttribute(pruned_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxTmaSmem);
  cudaFuncSetAttribute(shell_histogram_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
  cudaFuncSetAttribute(shell_histogram_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
  cudaFuncSetAttribute(shell_histogram_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
  cudaFuncSetAttribute(shell_gather_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
  cudaFuncSetAttribute(peaks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
  *out = p;
  return SBX_OK;
}

extern "C" int sbx_scatter_plan_create(sbx_scatter_plan** out, int device, int ndim, const int64_t* N,
                                       const double* L, int cutoff, double threshold, int max_batch) {
  return sbx_scatter_plan_create_ex(out, device, ndim, N, L, cutoff, threshold, max_batch, SBX_SCATTER_AUTO);
}

extern "C" int sbx_scatter_plan_destroy(sbx_scatter_plan* plan) {
  free_plan(plan);
  return SBX_OK;
}

extern "C" int sbx_scatter_plan_path(const sbx_scatter_plan* plan) { return plan ? plan->path : SBX_EINVAL; }
extern "C" const char* sbx_scatter_plan_kernel(const sbx_scatter_plan* plan) { return plan ? plan->last_kernel : ""; }

extern "C" int sbx_scatter_plan_shells(const sbx_scatter_plan* plan, int* n_shells, double* q_host,
                                       int64_t* count_host) {
  if (!plan || !n_shells) return fail(SBX_EINVAL, "NULL argument");
  *n_shells = plan->tab.n_shells();
  for (int s = 0; s < *n_shells; ++s) {
    if (q_host) q_host[s] = plan->q_host[s];
    if (count_host) count_host[s] = plan->tab.count[s];
  }
  return SBX_OK;
}

extern "C" int sbx_scatter_plan_members(const sbx_scatter_plan* plan, int shell, int64_t* n, double* k2xyz_host,
                                        int64_t capacity) {
  if (!plan || !n) return fail(SBX_EINVAL, "NULL argument");
  const ShellTable& t = plan->tab;
  if (shell < 0 || shell >= t.n_shells()) return fail(SBX_EINVAL, "shell %d out of range", shell);
  const auto& mem = t.members[shell];
  *n = (int64_t)mem.size();
  if (!k2xyz_host) return SBX_OK;
  if (capacity < *n) return fail(SBX_ERANGE, "member buffer holds %lld, need %lld", (long long)capacity, (long long)*n);
  for (size_t j = 0; j < mem.size(); ++j) {  // txyz row of scatter.py:82: (kv0[a], kv1[b], kv2[c])
    int64_t r = mem[j];
    for (int d = t.ndim - 1; d >= 0; --d) {
      k2xyz_host[j * t.ndim + d] = t.kv[d][r % t.N[d]];
      r /= t.N[d];
    }
  }
  return SBX_OK;
}

// ------------------------------------------------------------------------------ execution
static int check_peaks_args(int n, int step, int max_peaks) {
  if (step < 1 || step > 16384) return fail(SBX_EINVAL, "step must be in 1..16384, got %d", step);
  if (max_peaks < 0) return fail(SBX_EINVAL, "max_peaks must be >= 0");
  if ((peaks_smem_doubles(n, step) + (size_t)n) * sizeof(double) > kMaxDynSmem)
    return fail(SBX_EINVAL, "n=%d rows with step=%d exceed the peak-search shared memory", n, step);
  return SBX_OK;
}

namespace sbx { static int blocked_unit_planes(const sbx_scatter_plan* p, bool bulk, int planes); }

// The bulk-staged form of the blocked kernel (pruned_bulk_kernel) applies when rows split into 512-byte pieces per
// half (N2 % 128 == 0), a block of rows gives every warp whole row groups (yb_rows % 64 == 0) and the stages, the z
// twiddles and the work area fit one CTA's shared memory.  Measured on B200 (profiles/r01): see DESIGN.md section 3.
// SBX_BULK=0 at plan creation keeps the LDG blocked kernel.
static bool bulk_applies(const sbx_scatter_plan* p, size_t bwork) {
  if (p->env_bulk_off) return false;
  const bool two_items = (p->yb_rows / 64) * (p->grid.N2 / 128) >= 2 || p->fused_xc >= 2;  // per warp and unit
  return p->grid.N2 % 128 == 0 && p->yb_rows % 64 == 0 && two_items && bulk_smem_bytes(p->grid.N2, bwork) <= kMaxTmaSmem;
}

// scatter (+ optional fused peaks) on `batch` device-resident fields
static int run_scatter(sbx_scatter_plan* p, const double* fields, int batch, double* shell_I, bool do_peaks,
                       const PeakParams& pp, double* plot_y, int32_t* peak_idx, int32_t* n_peaks,
                       cudaStream_t st) {
  if (!p || !fields || !shell_I) return fail(SBX_EINVAL, "NULL argument");
  if (p->device == SBX_DEVICE_NONE) return fail(SBX_ENODEVICE, "geometry-only plan: no device to execute on");
  if (batch < 1) return fail(SBX_EINVAL, "batch must be >= 1, got %d", batch);
  DeviceGuard guard(p->device);
  const GridDev& g = p->grid;
  const ShellDev& sh = p->shell;
  const int ns = sh.n_shells;
  int rc;
  if (do_peaks && (rc = check_peaks_args(ns, pp.step, pp.max_peaks))) return rc;
  const size_t smem_peaks = do_peaks ? (peaks_smem_doubles(ns, pp.step) + ns) * sizeof(double) : 0;
  if (p->path == SBX_SCATTER_PRUNED_FUSED) {
    FusedArgs A{};
    A.N0 = g.N0; A.N1 = g.N1; A.N2 = g.N2;
    A.ks1 = g.N1 / 4; A.gstride = g.N1 + 4; A.ds = g.N2 / 16;
    A.twf = p->d_twf; A.tw1 = p->d_tw1; A.tw0 = p->d_tw0;
    A.umodes = p->d_umodes; A.n_umodes = p->n_umodes; A.memb_u = p->d_memb_u; A.shell_off = p->d_shell_off;
    A.n_shells = ns; A.factor = g.factor;
    A.xc = p->fused_xc; A.chunks = g.N0 / p->fused_xc;
    A.partial = p->d_partial; A.ctrl = p->d_ctrl;
    A.do_peaks = do_peaks ? 1 : 0; A.q = p->d_q; A.pp = pp;
    A.parts = A.chunks; A.planes = g.N0; A.x_base = 0; A.yb_rows = g.N1; A.yblocks = 1; A.finish = 1;
    const int force_blocked = p->env_force_blocked ? 1 : 0;
    // whole-plane kernels need >= 2 CTAs per SM to hide their barriers; beyond that (512^3, 1024^3): row-blocked kernel
    const bool whole_plane_thin = fused_smem_bytes(g.N0, g.N1, g.N2, fused_work_doubles(g.N1, ns, p->n_umodes, do_peaks ? pp.step : 0, do_peaks)) > 113 * 1024;
    if ((!p->fused_ok || force_blocked || whole_plane_thin) && p->blocked_ok) {
      A.yb_rows = p->yb_rows; A.yblocks = p->yblocks;
      const size_t bwork = blocked_work_doubles(p->yb_rows, ns, p->n_umodes, do_peaks ? pp.step : 0, do_peaks);
      const bool bulk = bulk_applies(p, bwork);
      A.xc = blocked_unit_planes(p, bulk, g.N0);
      A.chunks = g.N0 / A.xc; A.parts = A.chunks * A.yblocks;
      A.work_doubles = (int)bwork;
      const size_t bsmem = blocked_smem_bytes(g.N0, g.N2, bwork);
      if (bsmem > kMaxDynSmem) return fail(SBX_EINVAL, "blocked pruned kernel needs %zu bytes of shared memory", bsmem);
      p->last_kernel = bulk ? "pruned_bulk_kernel" : "pruned_blocked_kernel";
      for (int c0 = 0; c0 < batch; c0 += p->chunk) {
        const int nb = std::min(p->chunk, batch - c0);
        A.fields = fields + (size_t)c0 * g.nreal;
        A.n_fields = nb;
        A.shell_I = shell_I + (size_t)c0 * ns;
        A.plot_y = plot_y ? plot_y + (size_t)c0 * pp.step : nullptr;
        A.peak_idx = peak_idx ? peak_idx + (size_t)c0 * pp.max_peaks : nullptr;
        A.n_peaks = n_peaks ? n_peaks + c0 : nullptr;
        SBX_CUDA(cudaMemsetAsync(p->d_ctrl, 0, ((size_t)nb + 1) * sizeof(unsigned), st));
        const long long units = (long long)nb * A.parts;
        const int grid = (int)std::min<long long>(units, (long long)p->pruned_ctas_per_sm * p->sm_count);
        if (p->profiling) prof_event(p, st);
        if (bulk) pruned_bulk_kernel<<<(int)std::min<long long>(units, p->sm_count), 256, bulk_smem_bytes(g.N2, bwork), st>>>(A);
        else if (blocked_twf_in_smem(g.N2)) pruned_blocked_kernel<true><<<grid, 256, bsmem, st>>>(A);
        else pruned_blocked_kernel<false><<<grid, 256, bsmem, st>>>(A);
        SBX_CUDA(cudaGetLastError());
        if (p->profiling) { prof_event(p, st); prof_event(p, st); }
        p->own_launches += 1;
      }
      return SBX_OK;
    }
    const size_t work = fused_work_doubles(g.N1, ns, p->n_umodes, do_peaks ? pp.step : 0, do_peaks);
    A.work_doubles = (int)work;
    const size_t smem = fused_smem_bytes(g.N0, g.N1, g.N2, work);
    if (smem > kMaxDynSmem) return fail(SBX_EINVAL, "fused pruned kernel needs %zu bytes of shared memory", smem);
    // TMA-ring variant when the grid allows it and a useful ring fits in shared memory
    static const int tma_on = [] { const char* e = getenv("SBX_FUSED_TMA"); return e ? atoi(e) : 1; }();
    const size_t rwork = ring_work_doubles(g.N1, ns, p->n_umodes, do_peaks ? pp.step : 0, do_peaks);
    const size_t rfixed = ring_fixed_bytes(g.N0, g.N2, rwork);
    auto ilog2 = [](int v) { int l = 0; while ((1 << (l + 1)) <= v) ++l; return l; };
    auto pow2 = [](int v) { return v > 0 && (v & (v - 1)) == 0; };
    // ring size for a CTA that may use `budget` bytes: a power of two (slot = q & (S - 1)), at most one unit ahead
    auto ring_slots_for = [&](size_t budget) {
      int sl = rfixed < budget ? (int)((budget - rfixed) / ring_slot_bytes(g.N2)) : 0;
      sl = std::min(sl, std::min(kRingMaxSlots, A.xc * (g.N1 / 8)));
      if (const char* e = getenv("SBX_RING_SLOTS")) sl = std::min(sl, std::max(1, atoi(e)));
      return sl > 0 ? 1 << ilog2(sl) : 0;
    };
    // two 8-warp CTAs per SM when each still gets a ring of >= 8 slots (one CTA's barriers, y / x stages and field
    // finish then overlap the other's z stage), else one 16-warp CTA per SM
    static const int ring_warps = [] { const char* e = getenv("SBX_RING_WARPS"); return e ? atoi(e) : 0; }();
    const int slots8 = ring_slots_for(kMaxTmaSmem / 2 - 1024), slots16 = ring_slots_for(kMaxTmaSmem);
    const int nw = ring_warps == 8 || ring_warps == 16 ? ring_warps : (slots8 >= 8 ? 8 : 16);
    int slots = nw == 8 ? slots8 : slots16;
    // Measured on B200 (profiles/r01): the ring wins from 256^2 planes on (0.81 vs 0.98 ms per 32 x 256^3), the
    // 3-CTA/SM LDG kernel on smaller planes (0.94 vs 1.06 ms per 256 x 128^3); SBX_FUSED_TMA=2 forces the ring.
    const bool ring_fits = g.N1 % 16 == 0 && pow2(g.N1 / 8) && pow2(A.xc) && p->n_umodes <= kRingThreads && slots >= 4;
    const bool use_tma = ring_fits && (tma_on >= 2 || (tma_on == 1 && (long long)g.N1 * g.N2 >= 256LL * 256LL));
    p->last_kernel = use_tma ? "pruned_ring_kernel" : "pruned_fused_kernel";
    A.slots_shift = slots > 0 ? ilog2(slots) : 0;
    A.groups_shift = ilog2(g.N1 / 8);
    A.xc_shift = ilog2(A.xc);
    const size_t smem_tma = rfixed + (size_t)slots * ring_slot_bytes(g.N2);
    A.ring_slots = slots;
    if (use_tma) A.work_doubles = (int)rwork;
    for (int c0 = 0; c0 < batch; c0 += p->chunk) {
      const int nb = std::min(p->chunk, batch - c0);
      A.fields = fields + (size_t)c0 * g.nreal;
      A.n_fields = nb;
      A.shell_I = shell_I + (size_t)c0 * ns;
      A.plot_y = plot_y ? plot_y + (size_t)c0 * pp.step : nullptr;
      A.peak_idx = peak_idx ? peak_idx + (size_t)c0 * pp.max_peaks : nullptr;
      A.n_peaks = n_peaks ? n_peaks + c0 : nullptr;
      SBX_CUDA(cudaMemsetAsync(p->d_ctrl, 0, ((size_t)nb + 1) * sizeof(unsigned), st));
      const long long units = (long long)nb * A.chunks;
      const int grid1 = (int)std::min<long long>(units, (long long)p->pruned_ctas_per_sm * p->sm_count);
      if (p->profiling) prof_event(p, st);
      if (use_tma) {
        const int grid2 = (int)std::min<long long>(units, (long long)p->sm_count * (16 / nw));
        if (nw == 8) pruned_ring_kernel<8><<<grid2, 256, smem_tma, st>>>(A);
        else pruned_ring_kernel<16><<<grid2, 512, smem_tma, st>>>(A);
      } else
      pruned_fused_kernel<<<grid1, 256, smem, st>>>(A);
      SBX_CUDA(cudaGetLastError());
      if (p->profiling) { prof_event(p, st); prof_event(p, st); }
      p->own_launches += 1;
    }
    return SBX_OK;
  }
  if (p->path == SBX_SCATTER_PRUNED) {
    const size_t sm1 = p->pruned_variant == 1 ? pruned_fullk_smem_bytes(g.N1, g.N2) : pruned_smem_bytes(g.N1, g.N2);
    const size_t need2 = ((size_t)p->n_umodes + 2 + 2 * (size_t)g.N0) * sizeof(double);
    const size_t sm2 = (size_t)ns * sizeof(double) + std::max(need2, do_peaks ? peaks_smem_doubles(ns, pp.step) * sizeof(double) : 0);
    if (sm2 > kMaxDynSmem) return fail(SBX_EINVAL, "pruned finish kernel needs %zu bytes of shared memory", sm2);
    // The finish kernel (one CTA per field, latency bound) runs on an auxiliary stream so that it
    // overlaps the bandwidth-bound planes kernel of the next chunk; plane buffers are double buffered.
    // With profiling on, everything is serialised on `st` to time each kernel alone.
    const bool overlap = !p->profiling && batch > p->chunk;
    const size_t plane_doubles = (size_t)p->chunk * g.N0 * 256;
    int i = 0;
    for (int c0 = 0; c0 < batch; c0 += p->chunk, ++i) {
      const int nb = std::min(p->chunk, batch - c0);
      const int slot = overlap ? (i & 1) : 0;
      double* planes_buf = p->d_planes + (size_t)slot * plane_doubles;
      const long long planes = (long long)nb * g.N0;
      const int grid1 = (int)std::min<long long>(planes, (long long)p->pruned_ctas_per_sm * p->sm_count);
      if (overlap) SBX_CUDA(cudaStreamWaitEvent(st, p->ev_finish[slot], 0));
      if (p->profiling) prof_event(p, st);
      if (p->pruned_variant == 1)
        pruned_planes_fullk_kernel<<<grid1, 256, sm1, st>>>(fields + (size_t)c0 * g.nreal, p->pruned, p->d_tw2f, nb, planes_buf);
      else
        pruned_planes_kernel<<<grid1, 256, sm1, st>>>(fields + (size_t)c0 * g.nreal, p->pruned, nb, planes_buf);
      if (p->profiling) prof_event(p, st);
      cudaStream_t fs = st;
      if (overlap) {
        SBX_CUDA(cudaEventRecord(p->ev_planes[slot], st));
        SBX_CUDA(cudaStreamWaitEvent(p->s_aux, p->ev_planes[slot], 0));
        fs = p->s_aux;
      }
      pruned_finish_kernel<<<nb, 1024, sm2, fs>>>(planes_buf, g.N0, g.factor, p->d_umodes, p->n_umodes, p->d_tw0,
                                                 p->d_memb_u, p->d_shell_off, ns, shell_I + (size_t)c0 * ns,
                                                 do_peaks ? 1 : 0, p->d_q, pp, plot_y ? plot_y + (size_t)c0 * pp.step : nullptr,
                                                 peak_idx ? peak_idx + (size_t)c0 * pp.max_peaks : nullptr,
                                                 n_peaks ? n_peaks + c0 : nullptr);
      SBX_CUDA(cudaGetLastError());
      if (overlap) SBX_CUDA(cudaEventRecord(p->ev_finish[slot], p->s_aux));
      if (p->profiling) prof_event(p, st);
      p->own_launches += 2;
    }
    if (overlap) {
      SBX_CUDA(cudaStreamWaitEvent(st, p->ev_finish[0], 0));
      SBX_CUDA(cudaStreamWaitEvent(st, p->ev_finish[1], 0));
    }
    return SBX_OK;
  }
  for (int c0 = 0; c0 < batch; c0 += p->chunk) {
    const int nb = std::min(p->chunk, batch - c0);
    cufftHandle h;
    if ((rc = get_fft(p, nb, &h))) return rc;
    SBX_CUFFT(cufftSetStream(h, st));
    if (p->profiling) prof_event(p, st);
    SBX_CUFFT(cufftExecD2Z(h, const_cast<double*>(fields) + (size_t)c0 * g.nreal, (cufftDoubleComplex*)p->d_spec));
    if (p->profiling) prof_event(p, st);
    p->fft_calls += 1;
    p->own_launches += 1;
    double* outI = shell_I + (size_t)c0 * ns;
    double* py = plot_y ? plot_y + (size_t)c0 * pp.step : nullptr;
    int32_t* pi = peak_idx ? peak_idx + (size_t)c0 * pp.max_peaks : nullptr;
    int32_t* np_ = n_peaks ? n_peaks + c0 : nullptr;
    if (p->path == SBX_SCATTER_FFT_GATHER) {
      const size_t smem = std::max(smem_peaks, (size_t)ns * sizeof(double));
      shell_gather_kernel<<<nb, 256, smem, st>>>(p->d_spec, g.nspec, g.factor, p->d_elem, p->d_shell_off, ns, outI,
                                                 do_peaks ? 1 : 0, p->d_q, pp, py, pi, np_);
    } else {
      SBX_CUDA(cudaMemsetAsync(outI, 0, (size_t)nb * ns * sizeof(double), st));
      const size_t smem = std::max(smem_peaks, (size_t)2 * ns * sizeof(double));
      dim3 grid(p->blocks_per_field, nb);
      unsigned* tk = do_peaks ? p->d_tickets : nullptr;
      if (g.ndim == 3)
        shell_histogram_kernel<3><<<grid, 256, smem, st>>>(p->d_spec, g, sh, outI, tk, do_peaks, p->d_q, pp, py, pi, np_);
      else if (g.ndim == 2)
        shell_histogram_kernel<2><<<grid, 256, smem, st>>>(p->d_spec, g, sh, outI, tk, do_peaks, p->d_q, pp, py, pi, np_);
      else
        shell_histogram_kernel<1><<<grid, 256, smem, st>>>(p->d_spec, g, sh, outI, tk, do_peaks, p->d_q, pp, py, pi, np_);
    }
    SBX_CUDA(cudaGetLastError());
    if (p->profiling) prof_event(p, st);
  }
  return SBX_OK;
}

extern "C" int sbx_scatter_plan_set_profiling(sbx_scatter_plan* plan, int enable) {
  if (!plan) return fail(SBX_EINVAL, "NULL argument");
  plan->profiling = enable != 0;
  return SBX_OK;
}

extern "C" int sbx_scatter_plan_get_profile(sbx_scatter_plan* plan, double* transform_ms, double* fused_ms,
                                            int64_t* own_launches, int64_t* transform_calls, int reset) {
  if (!plan) return fail(SBX_EINVAL, "NULL argument");
  double a = 0.0, b = 0.0;
  if (plan->device != SBX_DEVICE_NONE && plan->prof_used) {
    DeviceGuard guard(plan->device);
    SBX_CUDA(cudaEventSynchronize(plan->prof_events[plan->prof_used - 1]));
    for (size_t i = 0; i + 3 <= plan->prof_used; i += 3) {
      float t0 = 0.f, t1 = 0.f;
      cudaEventElapsedTime(&t0, plan->prof_events[i], plan->prof_events[i + 1]);
      cudaEventElapsedTime(&t1, plan->prof_events[i + 1], plan->prof_events[i + 2]);
      a += t0;
      b += t1;
    }
  }
  if (transform_ms) *transform_ms = a;
  if (fused_ms) *fused_ms = b;
  if (own_launches) *own_launches = plan->own_launches;
  if (transform_calls) *transform_calls = plan->fft_calls;
  if (reset) {
    plan->prof_used = 0;
    plan->own_launches = 0;
    plan->fft_calls = 0;
  }
  return SBX_OK;
}

// ---- pruned-DFT slab helpers (scatter_types.h) --------------------------------------------------
namespace sbx {

// Planes per unit of the row-blocked kernels.  Fixed per plan and plane count: the summation order, hence the bits of
// the result, must not depend on how many fields a launch carries.  The LDG kernel keeps fused_xc (halving it for
// single-field launches was measured slower: its per-unit epilogue outweighs the tail).  The bulk-staged kernel runs
// one CTA per SM and pipelines across units, so small units cost nothing: planes per unit are halved until a field
// (or a rank's slab of it) gives every SM about six units -- as far as the partial-sum buffer of the plan allows
// (giant planes get room for one-plane units) and every warp keeps two pipeline items per unit.
static int blocked_unit_planes(const sbx_scatter_plan* p, bool bulk, int planes) {
  int xc = p->fused_xc;
  if (!bulk) return xc;
  const long long per_plane_items = (long long)(p->yb_rows / 64) * (p->grid.N2 / 128);
  while (xc % 2 == 0 && (long long)(planes / xc) * p->yblocks < 6LL * p->sm_count &&
         (size_t)p->chunk * (size_t)(p->grid.N0 / (xc / 2)) * p->yblocks <= p->partial_units && (xc / 2) * per_plane_items >= 2)
    xc /= 2;
  return xc;
}

static void slab_args(const sbx_scatter_plan* p, FusedArgs& A) {
  const GridDev& g = p->grid;
  A = FusedArgs{};
  A.N0 = g.N0; A.N1 = g.N1; A.N2 = g.N2;
  A.ks1 = g.N1 / 4; A.gstride = g.N1 + 4; A.ds = g.N2 / 16;
  A.twf = p->d_twf; A.tw1 = p->d_tw1; A.tw0 = p->d_tw0;
  A.umodes = p->d_umodes; A.n_umodes = p->n_umodes; A.memb_u = p->d_memb_u; A.shell_off = p->d_shell_off;
  A.n_shells = p->shell.n_shells; A.factor = g.factor;
  A.xc = p->fused_xc;
  A.yb_rows = p->yb_rows; A.yblocks = p->yblocks;
  A.partial = p->d_partial; A.ctrl = p->d_ctrl;
  A.q = p->d_q;
}

bool scatter_slab_supported(const sbx_scatter_plan* p, int planes) {
  return p && p->path == SBX_SCATTER_PRUNED_FUSED && p->blocked_ok && planes >= 1 && planes <= p->grid.N0 && planes % p->fused_xc == 0;
}

int scatter_slab_n_umodes(const sbx_scatter_plan* p) { return p ? p->n_umodes : 0; }

int scatter_slab_partial(sbx_scatter_plan* p, const double* slab, int planes, int x_base, double2* modesum, cudaStream_t st) {
  if (!scatter_slab_supported(p, planes)) return fail(SBX_EINVAL, "plan cannot run the pruned slab path on %d planes", planes);
  if (!slab || !modesum) return fail(SBX_EINVAL, "NULL argument");
  DeviceGuard guard(p->device);
  FusedArgs A;
  slab_args(p, A);
  A.fields = slab; A.n_fields = 1; A.planes = planes; A.x_base = x_base; A.finish = 0;
  const size_t bwork = blocked_work_doubles(p->yb_rows, A.n_shells, p->n_umodes, 0, false);
  const bool bulk = bulk_applies(p, bwork);
  A.xc = blocked_unit_planes(p, bulk, planes);
  A.chunks = planes / A.xc; A.parts = A.chunks * A.yblocks;
  A.work_doubles = (int)bwork;
  const size_t bsmem = blocked_smem_bytes(A.N0, A.N2, bwork);
  if (bsmem > kMaxDynSmem) return fail(SBX_EINVAL, "blocked pruned kernel needs %zu bytes of shared memory", bsmem);
  SBX_CUDA(cudaMemsetAsync(p->d_ctrl, 0, 2 * sizeof(unsigned), st));
  const int grid = (int)std::min<long long>(A.parts, (long long)p->pruned_ctas_per_sm * p->sm_count);
  if (bulk) pruned_bulk_kernel<<<(int)std::min<long long>(A.parts, p->sm_count), 256, bulk_smem_bytes(A.N2, bwork), st>>>(A);
  else if (blocked_twf_in_smem(A.N2)) pruned_blocked_kernel<true><<<grid, 256, bsmem, st>>>(A);
  else pruned_blocked_kernel<false><<<grid, 256, bsmem, st>>>(A);
  SBX_CUDA(cudaGetLastError());
  fused_reduce_parts_kernel<<<(p->n_umodes + 3) / 4, 128, 0, st>>>(p->d_partial, A.parts, p->n_umodes, modesum);
  SBX_CUDA(cudaGetLastError());
  p->own_launches += 2;
  p->last_kernel = bulk ? "pruned_bulk_kernel" : "pruned_blocked_kernel";
  return SBX_OK;
}

int scatter_slab_finish(sbx_scatter_plan* p, const double2* modesum, double* shell_I, cudaStream_t st) {
  if (!p || !modesum || !shell_I) return fail(SBX_EINVAL, "NULL argument");
  DeviceGuard guard(p->device);
  FusedArgs A;
  slab_args(p, A);
  A.n_fields = 1; A.parts = 1; A.chunks = 1;
  A.partial = const_cast<double2*>(modesum);
  A.shell_I = shell_I;
  const size_t smem = ((size_t)A.n_shells + p->n_umodes + 2) * sizeof(double);
  fused_finish_kernel<<<1, 256, smem, st>>>(A);
  SBX_CUDA(cudaGetLastError());
  p->own_launches += 1;
  return SBX_OK;
}

}  // namespace sbx

extern "C" int sbx_scatter_execute(sbx_scatter_plan* plan, const double* fields, int batch, double* shell_I,
                                   void* stream) {
  PeakParams pp{};
  return run_scatter(plan, fields, batch, shell_I, false, pp, nullptr, nullptr, nullptr, (cudaStream_t)stream);
}

extern "C" int sbx_scatter_peaks_execute(sbx_scatter_plan* plan, const double* fields, int batch, double* shell_I,
                                         double width, int step, double q_cutoff, double min_height,
                                         double* plot_y, int32_t* peak_idx, int32_t* n_peaks, int max_peaks,
                                         void* stream) {
  PeakParams pp{width, q_cutoff, min_height, step, max_peaks};
  return run_scatter(plan, fields, batch, shell_I, true, pp, plot_y, peak_idx, n_peaks, (cudaStream_t)stream);
}

extern "C" int sbx_peaks_execute(const double* q, const double* I, int n, int batch, double width, int step,
                                 double q_cutoff, double min_height, double* plot_y, int32_t* peak_idx,
                                 int32_t* n_peaks, int max_peaks, void* stream) {
  if (!q || !I) return fail(SBX_EINVAL, "NULL argument");
  if (n < 1 || batch < 1) return fail(SBX_EINVAL, "n and batch must be >= 1");
  int rc = check_peaks_args(n, step, max_peaks);
  if (rc) return rc;
  static thread_local bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(peaks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
    attr_set = true;
  }
  PeakParams pp{width, q_cutoff, min_height, step, max_peaks};
  const size_t smem = peaks_smem_doubles(n, step) * sizeof(double);
  peaks_kernel<<<batch, 256, smem, (cudaStream_t)stream>>>(q, I, n, pp, plot_y, peak_idx, n_peaks);
  SBX_CUDA(cudaGetLastError());
  return SBX_OK;
}

// host-buffer streaming: H2D of chunk i+1 overlaps the transform + binning of chunk i
static int ensure_stage(sbx_scatter_plan* p, int max_peaks) {
  const GridDev& g = p->grid;
  if (!p->s_copy) {
    SBX_CUDA(cudaStreamCreateWithFlags(&p->s_copy, cudaStreamNonBlocking));
    SBX_CUDA(cudaStreamCreateWithFlags(&p->s_comp, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
      SBX_CUDA(cudaMalloc((void**)&p->d_stage[i], (size_t)p->max_batch * g.nreal * sizeof(double)));
      SBX_CUDA(cudaMalloc((void**)&p->d_out[i], (size_t)p->max_batch * p->shell.n_shells * sizeof(double)));
      SBX_CUDA(cudaMalloc((void**)&p->d_npk[i], (size_t)p->max_batch * sizeof(int32_t)));
      SBX_CUDA(cudaEventCreateWithFlags(&p->ev_in[i], cudaEventDisableTiming));
      SBX_CUDA(cudaEventCreateWithFlags(&p->ev_done[i], cudaEventDisableTiming));
      SBX_CUDA(cudaEventCreateWithFlags(&p->ev_out[i], cudaEventDisableTiming));
    }
  }
  if (max_peaks > p->stage_max_peaks) {
    for (int i = 0; i < 2; ++i) {
      cudaFree(p->d_pidx[i]);
      SBX_CUDA(cudaMalloc((void**)&p->d_pidx[i], (size_t)p->max_batch * max_peaks * sizeof(int32_t)));
    }
    p->stage_max_peaks = max_peaks;
  }
  return SBX_OK;
}

extern "C" int sbx_scatter_execute_host(sbx_scatter_plan* plan, const double* fields_host, int64_t total,
                                        double* shell_I_host, double width, int step, double q_cutoff,
                                        double min_height, int32_t* peak_idx_host, int32_t* n_peaks_host,
                                        int max_peaks) {
  sbx_scatter_plan* p = plan;
  if (!p || !fields_host || !shell_I_host) return fail(SBX_EINVAL, "NULL argument");
  if (p->device == SBX_DEVICE_NONE) return fail(SBX_ENODEVICE, "geometry-only plan: no device to execute on");
  if (total < 1) return fail(SBX_EINVAL, "total must be >= 1");
  const bool do_peaks = n_peaks_host != nullptr;
  if (do_peaks && !peak_idx_host) max_peaks = 0;
  DeviceGuard guard(p->device);
  int rc = ensure_stage(p, do_peaks ? std::max(max_peaks, 1) : 0);
  if (rc) return rc;
  const GridDev& g = p->grid;
  const int ns = p->shell.n_shells;
  PeakParams pp{width, q_cutoff, min_height, step, max_peaks};
  int64_t done = 0;
  for (int64_t i = 0; done < total; ++i) {
    const int slot = (int)(i & 1);
    const int nb = (int)std::min<int64_t>(p->max_batch, total - done);
    // the stage buffer is free once the compute that read it (two chunks ago) has finished,
    // the output buffers once their D2H (two chunks ago) has been issued and completed
    SBX_CUDA(cudaStreamWaitEvent(p->s_copy, p->ev_done[slot], 0));
    SBX_CUDA(cudaMemcpyAsync(p->d_stage[slot], fields_host + (size_t)done * g.nreal,
                             (size_t)nb * g.nreal * sizeof(double), cudaMemcpyHostToDevice, p->s_copy));
    SBX_CUDA(cudaEventRecord(p->ev_in[slot], p->s_copy));
    SBX_CUDA(cudaStreamWaitEvent(p->s_comp, p->ev_in[slot], 0));
    SBX_CUDA(cudaStreamWaitEvent(p->s_comp, p->ev_out[slot], 0));
    rc = run_scatter(p, p->d_stage[slot], nb, p->d_out[slot], do_peaks, pp, nullptr,
                     do_peaks && max_peaks ? p->d_pidx[slot] : nullptr, do_peaks ? p->d_npk[slot] : nullptr,
                     p->s_comp);
    if (rc) return rc;
    // results ride the compute stream so that they never queue behind the next big H2D
    SBX_CUDA(cudaMemcpyAsync(shell_I_host + (size_t)done * ns, p->d_out[slot], (size_t)nb * ns * sizeof(double),
                             cudaMemcpyDeviceToHost, p->s_comp));
    if (do_peaks) {
      SBX_CUDA(cudaMemcpyAsync(n_peaks_host + done, p->d_npk[slot], (size_t)nb * sizeof(int32_t),
                               cudaMemcpyDeviceToHost, p->s_comp));
      if (max_peaks)
        SBX_CUDA(cudaMemcpyAsync(peak_idx_host + (size_t)done * max_peaks, p->d_pidx[slot],
                                 (size_t)nb * max_peaks * sizeof(int32_t), cudaMemcpyDeviceToHost, p->s_comp));
    }
    SBX_CUDA(cudaEventRecord(p->ev_done[slot], p->s_comp));
    SBX_CUDA(cudaEventRecord(p->ev_out[slot], p->s_comp));
    done += nb;
  }
  SBX_CUDA(cudaStreamSynchronize(p->s_comp));
  SBX_CUDA(cudaStreamSynchronize(p->s_copy));
  return SBX_OK;
}
