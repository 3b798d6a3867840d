// slab.cu -- one giant field across GPUs: slab-decomposed D2Z FFT + NCCL all-to-all (sbx.h).
//
// The reference cannot run this case at all (a 1024^3 field needs ~320 GB of temporaries in
// SimuBox/calculator/scatter.py:82-97); the only distributed FFT in the repository is the
// FFTW-MPI slab decomposition of the detached SCFT solver
// (work/.../ABC_star/abc_star_short.c:270-286).  B200 layout, one process per GPU:
//
//   rank r owns planes x in [r*n0, (r+1)*n0) of the N0 x N1 x N2 field           (n0 = N0/P)
//   A  cuFFT 2D D2Z over (y,z), K chunks of n0/K planes     F[x][j][c]     c < H = N2/2+1
//      pack by destination rank, columns c < Hc only        Y[d][x][jl][c] j = d*n1 + jl  (n1 = N1/P)
//   X  all-to-all (ncclSend/ncclRecv group per chunk)       X[s][x][jl][c] = spectrum of plane s*n0 + x
//   B  cuFFT 1D Z2Z along x (stride n1*Hc)      full spectrum for columns j in [r*n1,(r+1)*n1), c < Hc
//      fused |F|^2 / shell histogram (scatter_kernels.cuh, row remap) -> partial shell sums
//   R  ncclAllReduce(sum) of the n_shells partial sums
// Hc = number of columns c whose kv2[c] alone does not exceed the largest listed k^2: every k^2 of column c is
// kv2[c] plus non-negative terms (whatever quirk Q1 pairs with it), so the other columns can never be binned and
// are neither exchanged nor transformed along x -- exact, and it shrinks the exchange by Hc / H.
// Overlap: chunk k's exchange runs on a second stream while the 2D FFT + pack of chunk k+1 run on the caller's
// stream (the x chunks are independent until the 1D transform, which needs every plane and therefore follows the
// last chunk).  With profiling on, the stages are serialised on the caller's stream so that each can be timed.
//
// Fused pack + exchange (r02, sbx_scatter_slab_ipc_export / _attach): once the ranks have mapped each other's receive
// buffers with CUDA IPC, the pack kernel stores every destination's piece STRAIGHT INTO THAT PEER'S d_x over NVLink
// (slab_push_kernel: one kernel packs and exchanges, tile by tile, no send buffer, no ncclSend/ncclRecv); chunk k is
// pushed on the second stream while chunk k + 1 is transformed (two chunk buffers).  What is left for NCCL is a
// 4-byte all-reduce as the "every piece has landed" barrier before the 1D transform.  No barrier is needed before
// the pushes: every execute ends with the all-reduce of the shell sums, which no rank passes before all ranks have
// finished reading their d_x.
//
// NCCL is loaded with dlopen (the library the process already has -- torch's bundled
// libnccl.so.2 -- or $SBX_NCCL_LIB), so libsimubox_b200.so has no link-time NCCL dependency.
// sbx_scatter_slab_emulate runs the same stages for all P ranks on ONE device with device
// memcpys as the exchange: the single-GPU check of the decomposition.
#include <cufft.h>
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "common.h"
#include "scatter_types.h"
#include "shell_table.h"

using namespace sbx;

#define SBX_CUFFT(call)                                                                       \
  do {                                                                                        \
    cufftResult r_ = (call);                                                                  \
    if (r_ != CUFFT_SUCCESS) return ::sbx::fail(SBX_ECUFFT, "%s:%d %s -> cufftResult %d", __FILE__, __LINE__, #call, (int)r_); \
  } while (0)

// ------------------------------------------------------------------------------------ NCCL via dlopen
namespace {

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
};
NcclApi g_nccl;

int nccl_load() {
  if (g_nccl.handle) return SBX_OK;
  void* h = nullptr;
  const char* names[] = {getenv("SBX_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);  // the copy this process already uses, if any
  for (int i = 0; i < 3 && !h; ++i)
    if (names[i] && *names[i]) h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
  if (!h) return fail(SBX_ENCCL, "cannot load NCCL (set SBX_NCCL_LIB): %s", dlerror());
#define SBX_SYM(field, name)                                                          \
  *(void**)(&g_nccl.field) = dlsym(h, name);                                          \
  if (!g_nccl.field) return fail(SBX_ENCCL, "NCCL library lacks symbol %s", name)
  SBX_SYM(GetUniqueId, "ncclGetUniqueId");
  SBX_SYM(CommInitRank, "ncclCommInitRank");
  SBX_SYM(CommDestroy, "ncclCommDestroy");
  SBX_SYM(GroupStart, "ncclGroupStart");
  SBX_SYM(GroupEnd, "ncclGroupEnd");
  SBX_SYM(Send, "ncclSend");
  SBX_SYM(Recv, "ncclRecv");
  SBX_SYM(AllReduce, "ncclAllReduce");
  SBX_SYM(GetErrorString, "ncclGetErrorString");
  SBX_SYM(GetVersion, "ncclGetVersion");
#undef SBX_SYM
  g_nccl.handle = h;
  return SBX_OK;
}

#define SBX_NCCL(call)                                                                          \
  do {                                                                                          \
    ncclResult_t r_ = (call);                                                                   \
    if (r_ != ncclSuccess)                                                                      \
      return ::sbx::fail(SBX_ENCCL, "%s:%d %s -> %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r_)); \
  } while (0)

// F[xl][j][c < H] (one chunk of planes, xl local to the chunk) -> Y[d][x0 + xl][jl][c < Hc], j = d*n1 + jl:
// every (xl, d) pair moves n1 rows of Hc elements.
__global__ void __launch_bounds__(256)
slab_pack_kernel(const double2* __restrict__ F, double2* __restrict__ Y, int n0, int x0, int P, int n1, int H, int Hc) {
  const int xl = blockIdx.y / P, d = blockIdx.y - xl * P;
  const long long run = (long long)n1 * Hc;
  const double2* src = F + ((long long)xl * P + d) * n1 * H;
  double2* dst = Y + ((long long)d * n0 + x0 + xl) * run;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < run; i += (long long)gridDim.x * blockDim.x) {
    const long long jl = i / Hc;
    dst[i] = src[jl * H + (i - jl * Hc)];
  }
}

// Fused pack + exchange: like slab_pack_kernel, but piece d goes straight into rank d's receive buffer
// X_d[(sender*n0 + x0 + xl)][jl][c < Hc] (peer memory mapped through CUDA IPC; the own piece is local).
struct PeerPtrs { double2* x[16]; };
// PERSISTENT grid of a few CTAs (enough to saturate NVLink, SBX_SLAB_PUSH_CTAS, default 32): the kernel shares the GPU
// with the cuFFT passes of the next chunk, and a grid that filled every SM would simply run before or after them.
// Work item = (plane xl of the chunk, destination d, 4096-element segment of the n1 x Hc run), remote destinations
// first in rotated order so that the ranks do not all write to the same peer at once.
__global__ void __launch_bounds__(512)
slab_push_kernel(const double2* __restrict__ F, PeerPtrs peers, int sender, int n0, int x0, int nc0, int P, int n1, int H, int Hc) {
  const long long run = (long long)n1 * Hc;
  constexpr int kSeg = 4096;  // = blockDim.x * kU below
  const int segs = (int)((run + kSeg - 1) / kSeg);
  const long long items = (long long)nc0 * P * segs;
  for (long long it = blockIdx.x; it < items; it += gridDim.x) {
    const int seg = (int)(it % segs);
    const long long pd = it / segs;
    const int xl = (int)(pd % nc0);
    const int d = (sender + 1 + (int)(pd / nc0)) % P;  // own piece last
    const double2* src = F + ((long long)xl * P + d) * n1 * H;
    double2* dst = peers.x[d] + ((long long)sender * n0 + x0 + xl) * run;
    const long long i0 = (long long)seg * kSeg, i1 = min(run, i0 + kSeg);
    // kU independent 16-byte loads in flight per thread before the stores: a few CTAs have to cover ~2 us of NVLink
    // latency (one load -> store pair per thread gave 13 GB/s per CTA)
    constexpr int kU = 8;  // 512 threads x 8 = one 4096-element segment (64 KB) per pass
    double2 v[kU];
    long long idx[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      idx[u] = i0 + threadIdx.x + (long long)u * blockDim.x;
      if (idx[u] < i1) {
        const long long jl = idx[u] / Hc;
        v[u] = __ldcs(src + jl * H + (idx[u] - jl * Hc));
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u)
      if (idx[u] < i1) dst[idx[u]] = v[u];
  }
  __threadfence_system();  // the stores to peer memory are performed before the kernel counts as complete
}

template <class T>
int upload(T** dptr, const std::vector<T>& h) {
  *dptr = nullptr;
  if (h.empty()) return SBX_OK;
  SBX_CUDA(cudaMalloc((void**)dptr, h.size() * sizeof(T)));
  SBX_CUDA(cudaMemcpy(*dptr, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
  return SBX_OK;
}

}  // namespace

struct sbx_slab_plan {
  int rank = 0, world = 1, device = 0;
  ncclComm_t comm = nullptr;
  ShellTable tab;
  GridDev grid{};
  ShellDev shell{};
  std::vector<double> q_host;
  int n0 = 0, n1 = 0, H = 0;
  int Hc = 0;         // columns c < Hc are exchanged and transformed along x
  int K = 1, nc0 = 0; // chunks of the local planes and planes per chunk (n0 = K * nc0)
  long long blk = 0;  // elements exchanged per rank pair: n0*n1*Hc
  cufftHandle fft2d = 0, fft1d = 0;
  bool have2d = false, have1d = false;
  double2 *d_f = nullptr;                  // 2D spectrum of one chunk of planes: nc0 x N1 x H
  double2 *d_x = nullptr, *d_y = nullptr;  // receive / send buffers: world x blk
  cudaStream_t s_comm = nullptr;           // exchange stream
  cudaEvent_t ev_chunk[16] = {}, ev_comm = nullptr;
  // fused pack + exchange over peer memory (after sbx_scatter_slab_ipc_attach)
  bool p2p = false;
  double2* d_f2 = nullptr;                 // second chunk buffer (chunk k + 1 is transformed while chunk k is pushed)
  double2* peer_x[16] = {};                // receive buffers of all ranks (own: d_x), IPC-mapped
  bool peer_open[16] = {};
  cudaEvent_t ev_push[16] = {};
  int* d_flag = nullptr;                   // operand of the barrier all-reduce
  int push_ctas = 48;                      // persistent grid of slab_push_kernel
  double *d_rep = nullptr, *d_kvlast = nullptr, *d_partial = nullptr;
  double2* d_plane2 = nullptr;
  int blocks = 1;
  cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};  // stage timing (sbx_scatter_slab_profile)
  bool timed = false;
  // pruned-DFT mode: no transform of the whole field, no all-to-all -- every rank reduces its planes to the few
  // hundred low-|q| mode sums the shells need (pruned_blocked_kernel) and the ranks all-reduce those 5 kB
  bool pruned = false;
  sbx_scatter_plan* inner = nullptr;
  double2* d_modesum = nullptr;
};

static void slab_free(sbx_slab_plan* p) {
  if (!p) return;
  DeviceGuard guard(p->device);
  if (p->have2d) cufftDestroy(p->fft2d);
  if (p->have1d) cufftDestroy(p->fft1d);
  cudaFree(p->d_f); cudaFree(p->d_x); cudaFree(p->d_y); cudaFree(p->d_rep); cudaFree(p->d_kvlast); cudaFree(p->d_partial);
  for (int r = 0; r < 16; ++r) if (p->peer_open[r]) cudaIpcCloseMemHandle(p->peer_x[r]);
  cudaFree(p->d_f2); cudaFree(p->d_flag);
  for (cudaEvent_t e : p->ev_push) if (e) cudaEventDestroy(e);
  if (p->s_comm) cudaStreamDestroy(p->s_comm);
  for (cudaEvent_t e : p->ev_chunk) if (e) cudaEventDestroy(e);
  if (p->ev_comm) cudaEventDestroy(p->ev_comm);
  cudaFree(p->d_plane2);
  cudaFree(p->d_modesum);
  if (p->inner) sbx_scatter_plan_destroy(p->inner);
  for (cudaEvent_t e : p->ev) if (e) cudaEventDestroy(e);
  delete p;
}

static int slab_build(sbx_slab_plan** out, ncclComm_t comm, int rank, int world, int device, const int64_t* N,
                      const double* L, int cutoff, double threshold) {
  if (!out || !N || !L) return fail(SBX_EINVAL, "NULL argument");
  *out = nullptr;
  if (world < 1 || rank < 0 || rank >= world) return fail(SBX_EINVAL, "rank %d / world %d out of range", rank, world);
  if (N[0] % world)
    return fail(SBX_EINVAL, "N0=%lld must be divisible by the number of ranks (%d)", (long long)N[0], world);
  int rc = require_sm100(device);
  if (rc) return rc;
  DeviceGuard guard(device);
  sbx_slab_plan* p = new (std::nothrow) sbx_slab_plan();
  if (!p) return fail(SBX_ENOMEM, "out of host memory");
  p->rank = rank; p->world = world; p->device = device; p->comm = comm;
  if ((rc = build_shell_table(p->tab, 3, N, L, cutoff, threshold))) { slab_free(p); return rc; }
  const ShellTable& t = p->tab;
  p->q_host.resize(t.n_shells());
  for (int s = 0; s < t.n_shells(); ++s) p->q_host[s] = std::sqrt(t.rep[s]);  // scatter.py:118
  p->n0 = (int)(N[0] / world);
  p->n1 = (int)(N[1] / world);
  p->H = (int)(N[2] / 2 + 1);

  // ---- mode: pruned DFT when every shell mode is a low one (the default cutoff = 30 case), else slab FFT.
  // SBX_SLAB_PATH=fft / pruned forces one of them.
  {
    const char* sp = getenv("SBX_SLAB_PATH");
    const bool want_fft = sp && !strcmp(sp, "fft"), want_pruned = sp && !strcmp(sp, "pruned");
    if (!want_fft) {
      // The mode must come out the same on every rank, or their collectives no longer match: it may depend on the
      // geometry only (shell modes, grid, planes per rank -- SBX_EINVAL from the inner plan, identical everywhere).
      // Any other failure of the inner plan (allocation, CUDA) is an error of this rank, not a reason to change mode.
      sbx_scatter_plan* inner = nullptr;
      const int irc = sbx_scatter_plan_create_ex(&inner, device, 3, N, L, cutoff, threshold, 1, SBX_SCATTER_PRUNED_FUSED);
      if (irc == SBX_OK) {
        if (scatter_slab_supported(inner, p->n0)) {
          p->inner = inner;
          p->pruned = true;
        } else {
          sbx_scatter_plan_destroy(inner);
        }
      } else if (irc != SBX_EINVAL) {
        slab_free(p);
        return irc;
      }
      if (want_pruned && !p->pruned) {
        slab_free(p);
        return fail(SBX_EINVAL, "SBX_SLAB_PATH=pruned, but the pruned slab path does not apply to this grid / cutoff / rank count");
      }
    }
    if (p->pruned) {
      p->shell.n_shells = t.n_shells();
      cudaError_t e = cudaMalloc((void**)&p->d_modesum, (size_t)std::max(scatter_slab_n_umodes(p->inner), 1) * sizeof(double2));
      for (int i = 0; i < 5 && e == cudaSuccess; ++i) e = cudaEventCreate(&p->ev[i]);
      if (e != cudaSuccess) {
        slab_free(p);
        return fail(SBX_ECUDA, "slab plan allocation: %s", cudaGetErrorString(e));
      }
      *out = p;
      return SBX_OK;
    }
  }

  if (N[1] % world) {  // only the transpose of the slab FFT splits axis 1
    slab_free(p);
    return fail(SBX_EINVAL, "slab FFT mode: N1=%lld must be divisible by the number of ranks (%d)", (long long)N[1], world);
  }
  // k^2 tables: same construction as the single-GPU plan (scatter.cu build_device_tables)
  GridDev& g = p->grid;
  g.ndim = 3;
  g.N0 = (int)N[0]; g.N1 = (int)N[1]; g.N2 = (int)N[2];
  g.H = p->H;
  g.nreal = t.total();
  g.factor = 1.0 / (double)g.nreal;
  g.n1loc = p->n1;
  g.joff = rank * p->n1;
  {
    const int64_t rows = N[0] * N[1];
    std::vector<double2> p2(rows);
    for (int64_t row = 0; row < rows; ++row) {
      const int64_t a = row / N[1], b = row % N[1];
      const int64_t mrow = (a ? N[0] - a : 0) * N[1] + (b ? N[1] - b : 0);
      volatile double x = t.kv[0][row % N[0]] + t.kv[1][row / N[0]];
      volatile double y = t.kv[0][mrow % N[0]] + t.kv[1][mrow / N[0]];
      p2[row] = make_double2(x, y);
    }
    if ((rc = upload(&p->d_plane2, p2))) { slab_free(p); return rc; }
    std::vector<double> kl(t.kv[2].begin(), t.kv[2].begin() + g.H);
    if ((rc = upload(&p->d_kvlast, kl))) { slab_free(p); return rc; }
    if ((rc = upload(&p->d_rep, t.rep))) { slab_free(p); return rc; }
  }
  g.plane2 = p->d_plane2; g.plane = nullptr; g.kvlast = p->d_kvlast;
  ShellDev& sh = p->shell;
  sh.rep = p->d_rep;
  sh.n_shells = t.n_shells();
  sh.truncated = t.truncated ? 1 : 0;
  sh.designated = t.designated;
  sh.threshold = t.threshold;
  const double top = t.rep.back() + t.threshold;
  sh.vcut = top + 8.0 * std::fabs(top) * 2.220446049250313e-16 + 1e-300;
  if ((size_t)sh.n_shells * 3 * sizeof(double) > 200 * 1024) {
    slab_free(p);
    return fail(SBX_EINVAL, "%d shells exceed the shared-memory shell table", sh.n_shells);
  }
  // columns worth exchanging: kv2 grows with c on 0..N2/2, and a column whose kv2[c] alone is beyond the table
  // cannot hold a listed mode (SBX_SLAB_ALLCOLS=1 keeps all of them: the exchange-bound case for measurements)
  p->Hc = 1;
  while (p->Hc < p->H && t.kv[2][p->Hc] <= sh.vcut) ++p->Hc;
  { const char* ac = getenv("SBX_SLAB_ALLCOLS"); if (ac && atoi(ac) == 1) p->Hc = p->H; }
  p->blk = (long long)p->n0 * p->n1 * p->Hc;
  g.Hs = p->Hc;
  g.nspec = (long long)N[0] * p->n1 * p->Hc;  // elements this rank bins
  // chunks of the local planes: the largest count <= SBX_SLAB_CHUNKS (default 4, at most 16) that divides n0
  {
    const char* ce = getenv("SBX_SLAB_CHUNKS");
    int want = ce ? atoi(ce) : 4;
    want = std::max(1, std::min(want, 16));
    p->K = 1;
    for (int k = want; k >= 1; --k) if (p->n0 % k == 0) { p->K = k; break; }
    p->nc0 = p->n0 / p->K;
  }

  const size_t spec_bytes = (size_t)world * p->blk * sizeof(double2);
  cudaError_t e = cudaMalloc((void**)&p->d_f, (size_t)p->nc0 * N[1] * p->H * sizeof(double2));
  if (e == cudaSuccess && world > 1) e = cudaMalloc((void**)&p->d_x, spec_bytes);
  if (e == cudaSuccess) e = cudaMalloc((void**)&p->d_y, spec_bytes);
  if (e == cudaSuccess) e = cudaMalloc((void**)&p->d_partial, (size_t)sh.n_shells * sizeof(double));
  for (int i = 0; i < 5 && e == cudaSuccess; ++i) e = cudaEventCreate(&p->ev[i]);
  if (e == cudaSuccess) {  // exchange stream at the highest priority: its few CTAs are placed as soon as an SM has room
    int lo_pri = 0, hi_pri = 0;
    cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri);
    e = cudaStreamCreateWithPriority(&p->s_comm, cudaStreamNonBlocking, hi_pri);
  }
  { const char* pc = getenv("SBX_SLAB_PUSH_CTAS"); if (pc && atoi(pc) > 0) p->push_ctas = std::min(atoi(pc), 4096); }
  for (int i = 0; i < p->K && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&p->ev_chunk[i], cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_comm, cudaEventDisableTiming);
  if (e == cudaSuccess && world > 1 && world <= 16) {
    e = cudaMalloc((void**)&p->d_f2, (size_t)p->nc0 * N[1] * p->H * sizeof(double2));
    if (e == cudaSuccess) e = cudaMalloc((void**)&p->d_flag, sizeof(int));
    if (e == cudaSuccess) e = cudaMemset(p->d_flag, 0, sizeof(int));
    for (int i = 0; i < p->K && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&p->ev_push[i], cudaEventDisableTiming);
  }
  if (e != cudaSuccess) {
    slab_free(p);
    return fail(e == cudaErrorMemoryAllocation ? SBX_ENOMEM : SBX_ECUDA, "slab plan allocation: %s", cudaGetErrorString(e));
  }
  {
    int n2[2] = {(int)N[1], (int)N[2]};
    size_t ws = 0;
    cufftResult r = cufftCreate(&p->fft2d);
    if (r == CUFFT_SUCCESS) { p->have2d = true; r = cufftMakePlanMany(p->fft2d, 2, n2, nullptr, 1, 0, nullptr, 1, 0, CUFFT_D2Z, p->nc0, &ws); }
    if (r == CUFFT_SUCCESS) r = cufftCreate(&p->fft1d);
    if (r == CUFFT_SUCCESS) {
      p->have1d = true;
      int n1d[1] = {(int)N[0]};
      int embed[1] = {(int)N[0]};
      const long long batch = (long long)p->n1 * p->Hc;
      if (batch > 0x7fffffffLL) r = CUFFT_INVALID_SIZE;
      else r = cufftMakePlanMany(p->fft1d, 1, n1d, embed, (int)batch, 1, embed, (int)batch, 1, CUFFT_Z2Z, (int)batch, &ws);
    }
    if (r != CUFFT_SUCCESS) { slab_free(p); return fail(SBX_ECUFFT, "slab cuFFT plan creation failed: cufftResult %d", (int)r); }
  }
  int sm = 148;
  cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device);
  long long bpf = std::max<long long>((g.nspec + 16383) / 16384, 4LL * sm);
  bpf = std::min<long long>(bpf, std::max<long long>(1, g.nspec / 256));
  p->blocks = (int)std::min<long long>(bpf, 8LL * sm * 8);
  *out = p;
  return SBX_OK;
}

// stage A of chunk k: planes [k*nc0, (k+1)*nc0) of the local slab -> their part of the packed send buffer d_y
static int slab_stage_a(sbx_slab_plan* p, const double* local_slab, int k, cudaStream_t st) {
  const GridDev& g = p->grid;
  SBX_CUFFT(cufftSetStream(p->fft2d, st));
  SBX_CUFFT(cufftExecD2Z(p->fft2d, const_cast<double*>(local_slab) + (size_t)k * p->nc0 * g.N1 * g.N2, (cufftDoubleComplex*)p->d_f));
  const long long run = (long long)p->n1 * p->Hc;
  dim3 grid((unsigned)std::max<long long>(1, std::min<long long>((run + 255) / 256, 64)), (unsigned)(p->nc0 * p->world));
  slab_pack_kernel<<<grid, 256, 0, st>>>(p->d_f, p->d_y, p->n0, k * p->nc0, p->world, p->n1, p->H, p->Hc);
  SBX_CUDA(cudaGetLastError());
  return SBX_OK;
}

// stage A of chunk k in the fused form: transform into chunk buffer `buf`; slab_push then packs it into the peers
static int slab_fft_chunk(sbx_slab_plan* p, const double* local_slab, int k, double2* buf, cudaStream_t st) {
  const GridDev& g = p->grid;
  SBX_CUFFT(cufftSetStream(p->fft2d, st));
  SBX_CUFFT(cufftExecD2Z(p->fft2d, const_cast<double*>(local_slab) + (size_t)k * p->nc0 * g.N1 * g.N2, (cufftDoubleComplex*)buf));
  return SBX_OK;
}
static int slab_push(sbx_slab_plan* p, int k, const double2* buf, cudaStream_t st) {
  PeerPtrs pp;
  for (int r = 0; r < 16; ++r) pp.x[r] = r < p->world ? p->peer_x[r] : nullptr;
  slab_push_kernel<<<p->push_ctas, 512, 0, st>>>(buf, pp, p->rank, p->n0, k * p->nc0, p->nc0, p->world, p->n1, p->H, p->Hc);
  SBX_CUDA(cudaGetLastError());
  return SBX_OK;
}

// the pencils stage B works on: the receive buffer, or the packed buffer itself when there is nothing to exchange
static double2* slab_pencils(sbx_slab_plan* p) { return p->world > 1 ? p->d_x : p->d_y; }

// exchange of chunk k (planes [k*nc0, (k+1)*nc0) of every slab): piece d of d_y goes to rank d, piece s of d_x
// comes from rank s; the own piece is a device copy
static int slab_exchange(sbx_slab_plan* p, int k, cudaStream_t st) {
  const size_t piece = (size_t)p->nc0 * p->n1 * p->Hc, off = (size_t)k * piece;
  SBX_NCCL(g_nccl.GroupStart());
  for (int i = 1; i < p->world; ++i) {
    const int to = (p->rank + i) % p->world, from = (p->rank - i + p->world) % p->world;
    SBX_NCCL(g_nccl.Send(p->d_y + (size_t)to * p->blk + off, piece * 2, ncclDouble, to, p->comm, st));
    SBX_NCCL(g_nccl.Recv(p->d_x + (size_t)from * p->blk + off, piece * 2, ncclDouble, from, p->comm, st));
  }
  SBX_NCCL(g_nccl.GroupEnd());
  SBX_CUDA(cudaMemcpyAsync(p->d_x + (size_t)p->rank * p->blk + off, p->d_y + (size_t)p->rank * p->blk + off,
                           piece * sizeof(double2), cudaMemcpyDeviceToDevice, st));
  return SBX_OK;
}

// stage B: pencils -> partial shell sums d_partial
static int slab_stage_b(sbx_slab_plan* p, cudaStream_t st) {
  double2* pen = slab_pencils(p);
  SBX_CUFFT(cufftSetStream(p->fft1d, st));
  SBX_CUFFT(cufftExecZ2Z(p->fft1d, (cufftDoubleComplex*)pen, (cufftDoubleComplex*)pen, CUFFT_FORWARD));
  if (p->timed) cudaEventRecord(p->ev[3], st);
  const int ns = p->shell.n_shells;
  SBX_CUDA(cudaMemsetAsync(p->d_partial, 0, (size_t)ns * sizeof(double), st));
  return launch_shell_histogram3(pen, p->grid, p->shell, p->d_partial, p->blocks, st);
}

// ------------------------------------------------------------------------------------ C ABI
extern "C" int sbx_nccl_unique_id(void* id_host) {
  if (!id_host) return fail(SBX_EINVAL, "NULL argument");
  int rc = nccl_load();
  if (rc) return rc;
  ncclUniqueId id;
  SBX_NCCL(g_nccl.GetUniqueId(&id));
  static_assert(sizeof(ncclUniqueId) == SBX_NCCL_ID_BYTES, "ncclUniqueId size");
  memcpy(id_host, &id, sizeof(id));
  return SBX_OK;
}

extern "C" int sbx_nccl_comm_create(void** comm, const void* id_host, int rank, int world, int device) {
  if (!comm || !id_host) return fail(SBX_EINVAL, "NULL argument");
  *comm = nullptr;
  int rc = nccl_load();
  if (rc) return rc;
  if ((rc = require_sm100(device))) return rc;
  SBX_CUDA(cudaSetDevice(device));
  ncclUniqueId id;
  memcpy(&id, id_host, sizeof(id));
  ncclComm_t c = nullptr;
  SBX_NCCL(g_nccl.CommInitRank(&c, world, id, rank));
  *comm = (void*)c;
  return SBX_OK;
}

extern "C" int sbx_nccl_comm_destroy(void* comm) {
  if (!comm) return SBX_OK;
  int rc = nccl_load();
  if (rc) return rc;
  SBX_NCCL(g_nccl.CommDestroy((ncclComm_t)comm));
  return SBX_OK;
}

extern "C" int sbx_scatter_slab_plan_create(sbx_slab_plan** out, void* nccl_comm, int rank, int world, int device,
                                            const int64_t* N, const double* L, int cutoff, double threshold) {
  if (world > 1 && !nccl_comm) {
    if (out) *out = nullptr;
    return fail(SBX_ENCCL, "world=%d needs an NCCL communicator (sbx_nccl_comm_create)", world);
  }
  if (world > 1) {
    int rc = nccl_load();
    if (rc) return rc;
  }
  return slab_build(out, (ncclComm_t)nccl_comm, rank, world, device, N, L, cutoff, threshold);
}

extern "C" int sbx_scatter_slab_shells(const sbx_slab_plan* plan, int* n_shells, double* q_host, int64_t* count_host) {
  if (!plan || !n_shells) return fail(SBX_EINVAL, "NULL argument");
  *n_shells = plan->tab.n_shells();
  for (int s = 0; s < *n_shells; ++s) {
    if (q_host) q_host[s] = plan->q_host[s];
    if (count_host) count_host[s] = plan->tab.count[s];
  }
  return SBX_OK;
}

extern "C" int sbx_scatter_slab_execute(sbx_slab_plan* p, const double* local_slab, double* shell_I, void* stream) {
  if (!p || !local_slab || !shell_I) return fail(SBX_EINVAL, "NULL argument");
  DeviceGuard guard(p->device);
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  if (p->timed) cudaEventRecord(p->ev[0], st);
  if (p->pruned) {
    if ((rc = scatter_slab_partial(p->inner, local_slab, p->n0, p->rank * p->n0, p->d_modesum, st))) return rc;
    if (p->timed) cudaEventRecord(p->ev[1], st);
    if (p->world > 1)
      SBX_NCCL(g_nccl.AllReduce(p->d_modesum, p->d_modesum, (size_t)2 * scatter_slab_n_umodes(p->inner), ncclDouble, ncclSum, p->comm, st));
    if (p->timed) { cudaEventRecord(p->ev[2], st); cudaEventRecord(p->ev[3], st); }
    if ((rc = scatter_slab_finish(p->inner, p->d_modesum, shell_I, st))) return rc;
    if (p->timed) cudaEventRecord(p->ev[4], st);
    return SBX_OK;
  }
  if (p->p2p && p->world > 1) {
    if (p->timed) {
      // serial form for stage timing: [transform + push of every chunk] [barrier] ...
      for (int k = 0; k < p->K; ++k) {
        if ((rc = slab_fft_chunk(p, local_slab, k, p->d_f, st))) return rc;
        if ((rc = slab_push(p, k, p->d_f, st))) return rc;
      }
      cudaEventRecord(p->ev[1], st);
    } else {
      // chunk k is pushed into the peers on s_comm while chunk k + 1 is transformed on the caller's stream
      for (int k = 0; k < p->K; ++k) {
        double2* buf = (k & 1) ? p->d_f2 : p->d_f;
        if (k >= 2) SBX_CUDA(cudaStreamWaitEvent(st, p->ev_push[k - 2], 0));  // the buffer's previous push has read it
        if ((rc = slab_fft_chunk(p, local_slab, k, buf, st))) return rc;
        SBX_CUDA(cudaEventRecord(p->ev_chunk[k], st));
        SBX_CUDA(cudaStreamWaitEvent(p->s_comm, p->ev_chunk[k], 0));
        if ((rc = slab_push(p, k, buf, p->s_comm))) return rc;
        SBX_CUDA(cudaEventRecord(p->ev_push[k], p->s_comm));
      }
      for (int k = std::max(0, p->K - 2); k < p->K; ++k) SBX_CUDA(cudaStreamWaitEvent(st, p->ev_push[k], 0));
    }
    // every rank's pieces have landed in every d_x once all ranks have passed this point
    SBX_NCCL(g_nccl.AllReduce(p->d_flag, p->d_flag, 1, ncclInt, ncclSum, p->comm, st));
  } else if (p->timed || p->world == 1) {
    // serial form (stage timing; or nothing to exchange): all chunks through stage A, then all exchanges
    for (int k = 0; k < p->K; ++k)
      if ((rc = slab_stage_a(p, local_slab, k, st))) return rc;
    if (p->timed) cudaEventRecord(p->ev[1], st);
    for (int k = 0; k < p->K && p->world > 1; ++k)
      if ((rc = slab_exchange(p, k, st))) return rc;
  } else {
    // pipelined form: the exchange of chunk k on s_comm overlaps stage A of chunk k + 1 on the caller's stream
    for (int k = 0; k < p->K; ++k) {
      if ((rc = slab_stage_a(p, local_slab, k, st))) return rc;
      SBX_CUDA(cudaEventRecord(p->ev_chunk[k], st));
      SBX_CUDA(cudaStreamWaitEvent(p->s_comm, p->ev_chunk[k], 0));
      if ((rc = slab_exchange(p, k, p->s_comm))) return rc;
    }
    SBX_CUDA(cudaEventRecord(p->ev_comm, p->s_comm));
    SBX_CUDA(cudaStreamWaitEvent(st, p->ev_comm, 0));
  }
  if (p->timed) cudaEventRecord(p->ev[2], st);
  if ((rc = slab_stage_b(p, st))) return rc;
  const int ns = p->shell.n_shells;
  if (p->world > 1) SBX_NCCL(g_nccl.AllReduce(p->d_partial, shell_I, (size_t)ns, ncclDouble, ncclSum, p->comm, st));
  else SBX_CUDA(cudaMemcpyAsync(shell_I, p->d_partial, (size_t)ns * sizeof(double), cudaMemcpyDeviceToDevice, st));
  if (p->timed) cudaEventRecord(p->ev[4], st);
  return SBX_OK;
}

extern "C" int sbx_scatter_slab_ipc_export(sbx_slab_plan* p, void* handle_host) {
  if (!p || !handle_host) return fail(SBX_EINVAL, "NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == SBX_IPC_HANDLE_BYTES, "cudaIpcMemHandle_t size");
  if (p->pruned || p->world < 2 || !p->d_x) { memset(handle_host, 0, SBX_IPC_HANDLE_BYTES); return SBX_OK; }  // nothing to map
  DeviceGuard guard(p->device);
  cudaIpcMemHandle_t h;
  SBX_CUDA(cudaIpcGetMemHandle(&h, p->d_x));
  memcpy(handle_host, &h, sizeof(h));
  return SBX_OK;
}

extern "C" int sbx_scatter_slab_ipc_attach(sbx_slab_plan* p, const void* handles_host) {
  if (!p || !handles_host) return fail(SBX_EINVAL, "NULL argument");
  if (p->pruned || p->world < 2) return SBX_OK;  // pruned mode has no exchange
  if (p->world > 16) return fail(SBX_EINVAL, "the fused exchange supports up to 16 ranks, got %d", p->world);
  if (!p->d_f2) return fail(SBX_EINVAL, "plan has no second chunk buffer");
  DeviceGuard guard(p->device);
  for (int r = 0; r < p->world; ++r) {
    if (r == p->rank) { p->peer_x[r] = p->d_x; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)handles_host + (size_t)r * SBX_IPC_HANDLE_BYTES, sizeof(h));
    void* ptr = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      for (int q = 0; q < r; ++q) if (p->peer_open[q]) { cudaIpcCloseMemHandle(p->peer_x[q]); p->peer_open[q] = false; }
      return fail(SBX_ECUDA, "cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(e));
    }
    p->peer_x[r] = (double2*)ptr;
    p->peer_open[r] = true;
  }
  p->p2p = true;
  return SBX_OK;
}

extern "C" int sbx_scatter_slab_ipc_detach(sbx_slab_plan* p) {
  if (!p) return SBX_OK;
  DeviceGuard guard(p->device);
  cudaDeviceSynchronize();
  for (int r = 0; r < 16; ++r)
    if (p->peer_open[r]) { cudaIpcCloseMemHandle(p->peer_x[r]); p->peer_open[r] = false; p->peer_x[r] = nullptr; }
  p->p2p = false;
  return SBX_OK;
}

extern "C" int sbx_scatter_slab_exchange(const sbx_slab_plan* p) { return !p ? SBX_EINVAL : (p->pruned ? 0 : (p->p2p ? 2 : 1)); }

extern "C" const char* sbx_scatter_slab_kernel(const sbx_slab_plan* p) {
  return p && p->inner ? sbx_scatter_plan_kernel(p->inner) : "";
}

extern "C" int sbx_scatter_slab_path(const sbx_slab_plan* p) {
  if (!p) return SBX_EINVAL;
  return p->pruned ? SBX_SCATTER_PRUNED_FUSED : SBX_SCATTER_FFT_DENSE;
}

extern "C" int sbx_scatter_slab_set_profiling(sbx_slab_plan* p, int enable) {
  if (!p) return fail(SBX_EINVAL, "NULL argument");
  p->timed = enable != 0;
  return SBX_OK;
}

// ms of the last timed execute: [2D FFT + pack, all-to-all, 1D FFT, histogram + all-reduce]
extern "C" int sbx_scatter_slab_profile(sbx_slab_plan* p, double* ms4_host) {
  if (!p || !ms4_host) return fail(SBX_EINVAL, "NULL argument");
  DeviceGuard guard(p->device);
  SBX_CUDA(cudaEventSynchronize(p->ev[4]));
  for (int i = 0; i < 4; ++i) {
    float t = 0.f;
    SBX_CUDA(cudaEventElapsedTime(&t, p->ev[i], p->ev[i + 1]));
    ms4_host[i] = t;
  }
  return SBX_OK;
}

extern "C" int sbx_scatter_slab_plan_destroy(sbx_slab_plan* plan) {
  slab_free(plan);
  return SBX_OK;
}

// Single-device emulation of the `world`-rank decomposition: the same stage A / stage B code per
// rank, device memcpys in place of ncclSend/ncclRecv, a host loop in place of the all-reduce.
extern "C" int sbx_scatter_slab_emulate(int device, int world, const int64_t* N, const double* L, int cutoff,
                                        double threshold, const double* field, double* shell_I, void* stream) {
  if (!N || !L || !field || !shell_I) return fail(SBX_EINVAL, "NULL argument");
  if (world < 1 || world > 64) return fail(SBX_EINVAL, "world must be in 1..64");
  cudaStream_t st = (cudaStream_t)stream;
  std::vector<sbx_slab_plan*> plans(world, nullptr);
  int rc = SBX_OK;
  for (int r = 0; r < world && !rc; ++r) rc = slab_build(&plans[r], nullptr, r, world, device, N, L, cutoff, threshold);
  DeviceGuard guard(device);
  if (!rc && plans[0]->pruned) {
    // pruned mode: per-rank mode sums, added on the host in rank order, then one finish
    const long long slab = (long long)plans[0]->n0 * N[1] * N[2];
    const int nu = scatter_slab_n_umodes(plans[0]->inner);
    std::vector<double> acc((size_t)2 * nu, 0.0), part((size_t)2 * nu);
    for (int r = 0; r < world && !rc; ++r) {
      rc = scatter_slab_partial(plans[r]->inner, field + (size_t)r * slab, plans[r]->n0, r * plans[r]->n0, plans[r]->d_modesum, st);
      if (!rc && cudaMemcpyAsync(part.data(), plans[r]->d_modesum, part.size() * sizeof(double), cudaMemcpyDeviceToHost, st) != cudaSuccess)
        rc = fail(SBX_ECUDA, "mode-sum copy failed");
      if (!rc && cudaStreamSynchronize(st) != cudaSuccess) rc = fail(SBX_ECUDA, "stream sync failed: %s", cudaGetErrorString(cudaGetLastError()));
      for (size_t k = 0; k < acc.size() && !rc; ++k) acc[k] += part[k];
    }
    if (!rc && cudaMemcpyAsync(plans[0]->d_modesum, acc.data(), acc.size() * sizeof(double), cudaMemcpyHostToDevice, st) != cudaSuccess)
      rc = fail(SBX_ECUDA, "mode-sum upload failed");
    if (!rc) rc = scatter_slab_finish(plans[0]->inner, plans[0]->d_modesum, shell_I, st);
    if (!rc) cudaStreamSynchronize(st);
  } else if (!rc) {
    const long long slab = (long long)plans[0]->n0 * N[1] * N[2];
    for (int r = 0; r < world && !rc; ++r)
      for (int k = 0; k < plans[r]->K && !rc; ++k) rc = slab_stage_a(plans[r], field + (size_t)r * slab, k, st);
    if (!rc && world > 1)
      for (int r = 0; r < world; ++r)      // receiver
        for (int s = 0; s < world; ++s) {  // sender
          const size_t blk = (size_t)plans[r]->blk;
          if (cudaMemcpyAsync(plans[r]->d_x + s * blk, plans[s]->d_y + r * blk, blk * sizeof(double2), cudaMemcpyDeviceToDevice, st) != cudaSuccess)
            rc = fail(SBX_ECUDA, "emulated exchange failed");
        }
    const int ns = plans[0]->shell.n_shells;
    std::vector<double> acc(ns, 0.0), part(ns);
    for (int r = 0; r < world && !rc; ++r) {
      rc = slab_stage_b(plans[r], st);
      if (!rc && cudaMemcpyAsync(part.data(), plans[r]->d_partial, ns * sizeof(double), cudaMemcpyDeviceToHost, st) != cudaSuccess)
        rc = fail(SBX_ECUDA, "partial copy failed");
      if (!rc && cudaStreamSynchronize(st) != cudaSuccess) rc = fail(SBX_ECUDA, "stream sync failed: %s", cudaGetErrorString(cudaGetLastError()));
      for (int s = 0; s < ns && !rc; ++s) acc[s] += part[s];
    }
    if (!rc && cudaMemcpyAsync(shell_I, acc.data(), ns * sizeof(double), cudaMemcpyHostToDevice, st) != cudaSuccess)
      rc = fail(SBX_ECUDA, "result copy failed");
    if (!rc) cudaStreamSynchronize(st);
  }
  for (sbx_slab_plan* p : plans) slab_free(p);
  return rc;
}
