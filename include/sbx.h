/* sbx.h -- C ABI of libsimubox_b200.so: B200 (sm_100a) density-field analysis hot path.
 *
 * The reference (KangruiYuan/SimuBox) has no FFI for this path -- it is plain Python
 * (SURVEY.md 8(b)).  Each entry point below names the reference code it replaces; the
 * Python layer (simubox_b200/scatter.py, voronoi.py) binds them with ctypes and keeps the
 * reference's call signatures.  INTEGRATION.md shows the stub a SimuBox maintainer adds.
 *
 * Conventions
 *   - return 0 (SBX_OK) or a negative SBX_E* code; text via sbx_last_error() (thread-local).
 *   - pointers are CUDA DEVICE pointers unless the name ends in _host / the comment says host.
 *   - the library never allocates caller-visible buffers; a plan owns its cuFFT workspace,
 *     tables and scratch.  A plan is bound to one device and is not thread-safe.
 *   - *_execute calls are asynchronous on `stream` (a cudaStream_t passed as void*; NULL =
 *     legacy default stream) and never synchronise, except the *_host variants which
 *     return after the results are in host memory.
 *   - no CPU fallback: without an sm_100 device plan creation fails with SBX_ENODEVICE.
 */
#ifndef SBX_H
#define SBX_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SBX_VERSION 100 /* 0.1.0 */

enum {
  SBX_OK = 0,
  SBX_EINVAL = -1,    /* bad argument (odd grid size, ndim, NULL pointer ...)          */
  SBX_ENODEVICE = -2, /* no CUDA device / not compute capability 10.x                   */
  SBX_ECUDA = -3,     /* CUDA runtime error                                             */
  SBX_ECUFFT = -4,    /* cuFFT error                                                    */
  SBX_ENOMEM = -5,    /* host or device allocation failed                               */
  SBX_ENCCL = -6,     /* NCCL error / NCCL library not loadable                         */
  SBX_ERANGE = -7     /* output capacity too small (max_peaks, member buffer ...)        */
};

/* device = SBX_DEVICE_NONE builds a geometry-only scatter plan on the host (shell table,
 * q, counts, members); no CUDA call is made and the *_execute entry points return
 * SBX_ENODEVICE.  Used to inspect the shell geometry and by the CPU-side tests. */
#define SBX_DEVICE_NONE (-1)

/* scatter-plan flags (sbx_scatter_plan_create_ex) */
enum {
  SBX_SCATTER_AUTO = 0,       /* choose the fastest valid path                                */
  SBX_SCATTER_FFT_DENSE = 1,  /* cuFFT D2Z + fused full-spectrum |F|^2/shell-histogram kernel */
  SBX_SCATTER_FFT_GATHER = 2, /* cuFFT D2Z + fused gather of the shell members only           */
  SBX_SCATTER_PRUNED = 3,     /* hand-written pruned DFT of the member modes (no spectrum),
                                 two kernels: planes + finish                                  */
  SBX_SCATTER_PRUNED_FUSED = 4 /* the same in ONE kernel: folded z stage, x accumulation in
                                 registers, finish by the last CTA of each field               */
};

typedef struct sbx_scatter_plan sbx_scatter_plan;
typedef struct sbx_slab_plan sbx_slab_plan;

int sbx_version(void);
const char* sbx_last_error(void);
/* number of visible CUDA devices with compute capability 10.x; <0 on error */
int sbx_device_count(void);

/* ---- scattering: replaces Scatter.k_space + Scatter.sacttering_peak -----------------
 * (reference SimuBox/calculator/scatter.py:40-65 and :68-122).
 * ndim in 1..3; N[d] even; L[d] box lengths; cutoff/threshold as in sacttering_peak.
 * Builds: k^2 component tables (scatter.py:60), the field-independent shell table that the
 * reference obtains by argsort + merge loop (scatter.py:97-115, incl. the meshgrid 'xy'
 * pairing of :62-63 and the truncated last shell of :106-107), cuFFT D2Z plan(s) for up to
 * max_batch fields per launch. */
int sbx_scatter_plan_create(sbx_scatter_plan** out, int device, int ndim, const int64_t* N,
                            const double* L, int cutoff, double threshold, int max_batch);
int sbx_scatter_plan_create_ex(sbx_scatter_plan** out, int device, int ndim, const int64_t* N,
                               const double* L, int cutoff, double threshold, int max_batch,
                               int path /* SBX_SCATTER_* */);
int sbx_scatter_plan_destroy(sbx_scatter_plan* plan);

/* Geometry of the result rows = q_Intensity[:,0] and member counts (scatter.py:117-118).
 * q_host / count_host may be NULL to query n_shells only.  Host pointers. */
int sbx_scatter_plan_shells(const sbx_scatter_plan* plan, int* n_shells, double* q_host,
                            int64_t* count_host);
/* q_idx[shell] (scatter.py:103,110-115): n member vectors of ndim squared k components each,
 * in ascending spectrum-index order.  Pass k2xyz_host = NULL to query n.  Host pointers. */
int sbx_scatter_plan_members(const sbx_scatter_plan* plan, int shell, int64_t* n,
                             double* k2xyz_host, int64_t capacity);
/* which SBX_SCATTER_* path the plan runs */
int sbx_scatter_plan_path(const sbx_scatter_plan* plan);
/* name of the dominant hand-written kernel the last execute launched ("" before the first one) */
const char* sbx_scatter_plan_kernel(const sbx_scatter_plan* plan);

/* Live profiling for bench.py: when enabled every chunk launch is bracketed by CUDA events on
 * the launching stream.  get_profile synchronises on the last event and returns the summed
 * device time of the transform stage and of the hand-written fused kernel, the number of this
 * library's own kernel launches and of transform calls since the last reset. */
int sbx_scatter_plan_set_profiling(sbx_scatter_plan* plan, int enable);
int sbx_scatter_plan_get_profile(sbx_scatter_plan* plan, double* transform_ms, double* fused_ms,
                                 int64_t* own_launches, int64_t* transform_calls, int reset);

/* fields: batch x prod(N) fp64, C order (= DensityParseResult.mat of each field);
 * shell_I: batch x n_shells fp64 = q_Intensity[:,1] per field.  Any batch >= 1: the plan works
 * through it in chunks of max_batch (its workspace size). */
int sbx_scatter_execute(sbx_scatter_plan* plan, const double* fields, int batch, double* shell_I,
                        void* stream);

/* ---- peaks: replaces the numeric part of Scatter.show_peak (scatter.py:137-149) and
 * peak.gaussian_expansion (calculator/peak.py:11-17) + scipy.signal.find_peaks(height=).
 * q: n shell positions (shared by the batch); I: batch x n intensities (row 0 is zeroed
 * internally, scatter.py:138); rows with q <= q_cutoff are kept (:141).
 * plot_y: batch x step; peak_idx: batch x max_peaks ascending sample indices;
 * n_peaks: batch counts (may exceed max_peaks: then only the first max_peaks are stored). */
int sbx_peaks_execute(const double* q, const double* I, int n, int batch, double width, int step,
                      double q_cutoff, double min_height, double* plot_y, int32_t* peak_idx,
                      int32_t* n_peaks, int max_peaks, void* stream);

/* scatter + peaks in one call (the fused tail runs in the last-arriving block of each field) */
int sbx_scatter_peaks_execute(sbx_scatter_plan* plan, const double* fields, int batch,
                              double* shell_I, double width, int step, double q_cutoff,
                              double min_height, double* plot_y, int32_t* peak_idx,
                              int32_t* n_peaks, int max_peaks, void* stream);

/* End-to-end variant with HOST buffers (pinned memory recommended): total fields are
 * streamed through the plan in chunks of max_batch with copy/compute overlap; returns after
 * shell_I_host (total x n_shells) -- and, if non-NULL, peak_idx_host / n_peaks_host -- are
 * complete.  This is the call Scatter.sacttering_peak_batch makes. */
int sbx_scatter_execute_host(sbx_scatter_plan* plan, const double* fields_host, int64_t total,
                             double* shell_I_host, double width, int step, double q_cutoff,
                             double min_height, int32_t* peak_idx_host, int32_t* n_peaks_host,
                             int max_peaks);

/* ---- Voronoi labelling: replaces the per-pixel loop of
 * VoronoiCell.weighted_voronoi_diagrams (calculator/voronoi.py:284-292) with
 * VoronoiCell.additive (:244-246, mode 0) or VoronoiCell.power (:249-255, mode 1).
 * ndim 2 or 3; N grid size; seeds S x ndim fp64 (voxel units); weights S fp64 or NULL (= 0);
 * box_host: ndim periodic lengths on the HOST or NULL for the reference's non-periodic
 * behaviour; mask: prod(N) uint8 or NULL (voxels with mask == 0 get label -1);
 * labels: prod(N) int32, C order (x-major like product(range(imgx), range(imgy))).
 * Ties go to the lowest seed index (np.argmin).  stats_host (may be NULL) receives
 * [distance evaluations, tiles, overflowed tiles] after a stream sync when non-NULL.
 * One-shot form: allocates and frees its workspace and synchronises `stream` before returning. */
int sbx_voronoi_label(int device, int ndim, const int64_t* N, const double* seeds, int64_t S,
                      const double* weights, int mode, const double* box_host,
                      const uint8_t* mask, int32_t* labels, void* stream, int64_t* stats_host);
/* Plan form of the same call for repeated labellings of one grid size (sweeps): the plan owns the
 * workspace (cell table + sorted image points, or the tile candidate lists when weights / power mode are used),
 * execute is three kernel launches, asynchronous on `stream` unless
 * stats_host is non-NULL.  A plan is bound to one device; executes on one plan must be issued in
 * stream order (the workspace is reused). */
typedef struct sbx_voronoi_plan sbx_voronoi_plan;
int sbx_voronoi_plan_create(sbx_voronoi_plan** out, int device, int ndim, const int64_t* N);
int sbx_voronoi_plan_execute(sbx_voronoi_plan* plan, const double* seeds, int64_t S,
                             const double* weights, int mode, const double* box_host,
                             const uint8_t* mask, int32_t* labels, void* stream,
                             int64_t* stats_host);
/* The same labelling restricted to the planes [x_lo, x_hi) of the first grid axis -- the voxel-slab
 * decomposition of one grid across GPUs (seeds replicated, planes [r N0/P, (r+1) N0/P) on
 * rank r, no collective; replaces the same loop, voronoi.py:284-292, over that rank's pixels).
 * `labels` and `mask` point to the first plane of the range ((x_hi - x_lo) x N1 [x N2] elements); the
 * result equals the corresponding planes of a whole-grid labelling bit for bit. */
int sbx_voronoi_plan_execute_range(sbx_voronoi_plan* plan, const double* seeds, int64_t S,
                                   const double* weights, int mode, const double* box_host,
                                   const uint8_t* mask, int32_t* labels, int64_t x_lo, int64_t x_hi,
                                   void* stream, int64_t* stats_host);
/* which kernels the last execute of this plan ran (introspection for tests and bench records; no reference
 * counterpart): 0 = tile path (coarse + mid + label kernels), 1 = cell-list path, -1 = no execute yet */
int sbx_voronoi_plan_last_path(const sbx_voronoi_plan* plan);
int sbx_voronoi_plan_destroy(sbx_voronoi_plan* plan);
/* Allocation-free labelling.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * sbx_voronoi_plan_execute_range with weights = NULL, mode = 0 (additive), bit for bit; it calls no cudaMalloc /
 * cudaFree / allocating library routine.  Its scratch (sorted image points, cell counters, cell start table) is one
 * file-scope device array of 12 MiB that every GPU loading this library holds until exit, which sets the limits:
 *   S x (27 | 9 | 1 images for periodic 3D | periodic 2D | non-periodic) <= 262144 image points
 *   (4096 seeds in a periodic 3D box use 110592) and at most 262144 cells of the internal cell list;
 * beyond either the call returns SBX_ERANGE and names the plan entry point to use instead.  Asynchronous on `stream`.
 * All calls share the array: calls that overlap on one GPU (other streams or threads) race on it -- issue them in
 * stream order.  The first call on a device resolves the array's address (cudaGetSymbolAddress): make it outside
 * graph capture. */
int sbx_voronoi_label_noalloc(int device, int ndim, const int64_t* N, const double* seeds, int64_t S,
                              const double* box_host, const uint8_t* mask, int32_t* labels, int64_t x_lo,
                              int64_t x_hi, void* stream);

/* ---- one giant field across GPUs: slab-decomposed FFT + NCCL all-to-all ---------------
 * The reference cannot run a 1024^3 field (scatter.py:82-97 materialises ~300 B/voxel); its only
 * distributed FFT is the FFTW-MPI slab code of the detached solver
 * (work/.../ABC_star/abc_star_short.c:270-286).  One process per GPU:
 *   sbx_nccl_unique_id on rank 0 -> broadcast the SBX_NCCL_ID_BYTES bytes by any means
 *   (torch.distributed, MPI, a file) -> sbx_nccl_comm_create on every rank.
 * NCCL is loaded at run time with dlopen (the copy already in the process, else $SBX_NCCL_LIB,
 * else libnccl.so.2).  N0 must be divisible by `world`; the slab FFT mode
 * (not the pruned mode) needs N1 divisible by `world` as well.
 * local_slab = planes [rank*N0/world, (rank+1)*N0/world) of the field (fp64, C order);
 * shell_I: n_shells fp64, identical on all ranks after the call completes on `stream`. */
#define SBX_NCCL_ID_BYTES 128
int sbx_nccl_unique_id(void* id_host /* SBX_NCCL_ID_BYTES */);
int sbx_nccl_comm_create(void** comm, const void* id_host, int rank, int world, int device);
int sbx_nccl_comm_destroy(void* comm);
int sbx_scatter_slab_plan_create(sbx_slab_plan** out, void* nccl_comm, int rank, int world,
                                 int device, const int64_t* N, const double* L, int cutoff,
                                 double threshold);
int sbx_scatter_slab_shells(const sbx_slab_plan* plan, int* n_shells, double* q_host,
                            int64_t* count_host);
int sbx_scatter_slab_execute(sbx_slab_plan* plan, const double* local_slab, double* shell_I,
                             void* stream);
/* Fused pack + exchange over peer memory for the slab FFT mode (no counterpart in the reference: it has no
 * multi-GPU path; the exchange it stands for is the FFTW-MPI transpose of work/.../abc_star_short.c:270-286).
 * After plan creation every rank exports the CUDA IPC handle of its receive buffer, the SBX_IPC_HANDLE_BYTES-byte
 * handles of all ranks are gathered by any means (rank order) and attached on every rank; from then on
 * sbx_scatter_slab_execute packs each destination's piece straight into that rank's buffer over NVLink (one kernel,
 * no ncclSend/ncclRecv; a 4-byte all-reduce is the arrival barrier).  Ranks must be processes on one node with
 * peer access (NVLink / NVSwitch); up to 16 ranks.  A plan in pruned mode ignores both calls.  Before destroying the
 * plans: detach on every rank, then a barrier of the caller's, then destroy (the exporter must outlive the mappings).
 * sbx_scatter_slab_exchange: 0 = none (pruned mode), 1 = NCCL send/recv, 2 = fused peer stores. */
#define SBX_IPC_HANDLE_BYTES 64
int sbx_scatter_slab_ipc_export(sbx_slab_plan* plan, void* handle_host /* SBX_IPC_HANDLE_BYTES */);
int sbx_scatter_slab_ipc_attach(sbx_slab_plan* plan, const void* handles_host /* world x SBX_IPC_HANDLE_BYTES */);
int sbx_scatter_slab_ipc_detach(sbx_slab_plan* plan);
int sbx_scatter_slab_exchange(const sbx_slab_plan* plan);
/* stage timing of the last execute after set_profiling(1): ms4_host = [2D FFT + pack, all-to-all,
 * 1D FFT, histogram + all-reduce]; synchronises on the last event */
/* SBX_SCATTER_PRUNED_FUSED when the plan runs the pruned DFT (every shell mode low: each rank reduces its planes
 * to the mode sums, one small all-reduce, no all-to-all), SBX_SCATTER_FFT_DENSE for the slab FFT.  The
 * environment variable SBX_SLAB_PATH=fft|pruned forces one at plan creation. */
int sbx_scatter_slab_path(const sbx_slab_plan* plan);
/* dominant hand-written kernel of the last execute in pruned mode ("pruned_bulk_kernel" / "pruned_blocked_kernel"),
 * "" in slab FFT mode or before the first execute */
const char* sbx_scatter_slab_kernel(const sbx_slab_plan* plan);
int sbx_scatter_slab_set_profiling(sbx_slab_plan* plan, int enable);
int sbx_scatter_slab_profile(sbx_slab_plan* plan, double* ms4_host);
int sbx_scatter_slab_plan_destroy(sbx_slab_plan* plan);
/* The same decomposition for all `world` ranks on ONE device (device memcpys as the exchange):
 * field = the whole N0 x N1 x N2 field.  Validation of the slab path without several GPUs;
 * synchronises `stream`. */
int sbx_scatter_slab_emulate(int device, int world, const int64_t* N, const double* L, int cutoff,
                             double threshold, const double* field, double* shell_I, void* stream);

/* ---- per-cell statistics on a label grid (device pointers) -------------------------------------
 * What the reference derives downstream of its Voronoi / Delaunay constructions
 * (SimuBox/calculator/voronoi.py:150-213 coord_dict; demos/scripts.ipynb cells 1-2 in-cell densities),
 * computed from the int32 label array sbx_voronoi_label writes (label < 0 = masked voxel, skipped):
 *   volume[s]  voxels labelled s;  mass[s]  sum of field over them (field/mass may be NULL);
 *   adjacency  n_cells x ceil(n_cells/32) uint32 bit matrix, bit (s,t) set iff two face-adjacent
 *              voxels (wrap-around when periodic != 0) carry labels s and t (NULL = skip);
 *   degree[s]  number of set bits of row s (needs adjacency).
 * All outputs are overwritten.  Asynchronous on `stream`. */
int sbx_cell_stats(int device, int ndim, const int64_t* N, const int32_t* labels, const double* field,
                   int64_t n_cells, int periodic, int64_t* volume, double* mass, uint32_t* adjacency,
                   int32_t* degree, void* stream);

/* ---- periodic domain-centre finder (device pointers; n_components_host is a host pointer) -------
 * 3D-capable replacement for the seed finding of VoronoiCell.OpenCV (SimuBox/calculator/voronoi.py:31-68:
 * 3 x 3 tiling + cv2 threshold + contour moments, 2D only): foreground = field > threshold, connected
 * components under face connectivity with wrap-around when periodic != 0, centroid of every component
 * = root + mean minimum-image displacement (integer sums: bit-reproducible), wrapped into [0, N).
 * Components are ordered by their smallest voxel index.
 *   comp_labels  Nxyz int32 out, component id per voxel, -1 background (NULL = not wanted)
 *   centres      max_components x ndim fp64 (voxel units), volumes max_components int64 (NULL ok)
 *   *n_components_host = number of components found; only the first max_components are written.
 * Synchronises on `stream` (it has to return the count). */
int sbx_domain_centres(int device, int ndim, const int64_t* N, const double* field, double threshold,
                       int periodic, int32_t* comp_labels, int64_t max_components, double* centres,
                       int64_t* volumes, int64_t* n_components_host, void* stream);

/* ---- particle -> grid density splat (device pointers; N, L host) --------------------------------
 * XmlTransformer.transform (SimuBox/toolkit/xmltrans.py:117-153): every particle adds 1 to each grid point of
 * its type's density within r_cut, with the reference's stencil, base-index and wrapped-coordinate
 * arithmetic (bit-identical counts).  xyz: P x 3 fp64 in [-L/2, L/2); type: P int32 in [0, n_types)
 * (others are skipped); phi: n_types x N0 x N1 x N2 fp64, overwritten.  Asynchronous on `stream`. */
int sbx_particle_splat(int device, const double* xyz, const int32_t* type, int64_t n_particles, int n_types,
                       const int64_t* N, const double* L, double r_cut, double* phi, void* stream);

/* ---- phin / phout text reader (host code, no device needed) ---------------------------------
 * Replaces pd.read_csv(path, skiprows=skip, header=None, sep="[ ]+", engine="python")
 * (SimuBox/toolkit/reader.py:170-172).  buf = the whole file in host memory.  `skip` header lines are
 * skipped, blank lines ignored, tokens are blank-separated and converted with the arithmetic of
 * pandas' precise_xstrtod (what that read_csv call uses: <= 17 digits accumulated in a double, then one multiply or
 * divide by a power of ten -- NOT correctly rounded beyond 15 digits, and deliberately so: the doubles are the
 * reference's bit for bit).  threads <= 0: one per core.
 * shape: rows = non-blank lines after the header, cols = tokens of the first of them.
 * parse: out = rows x cols fp64, C order; SBX_EINVAL when a row has another width or a token is not a
 * number (the Python layer raises ValueError, which drives the reference's header-guess retry,
 * reader.py:173-186).  is_ascii: 1 when every byte < 0x80 (then the file certainly decodes as UTF-8). */
int sbx_text_is_ascii(const char* buf, int64_t len);
int sbx_text_table_shape(const char* buf, int64_t len, int skip, int threads, int64_t* rows, int32_t* cols);
int sbx_text_table_parse(const char* buf, int64_t len, int skip, int64_t rows, int32_t cols, double* out,
                         int threads);

#ifdef __cplusplus
}
#endif
#endif /* SBX_H */
