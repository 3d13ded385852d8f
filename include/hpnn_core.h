/*
 * hpnn_core.h -- C ABI of the hpnn-b200 native core (libhpnn_core.so).
 *
 * The core is libtorch-free: it owns the file formats, the CPU/OpenMP oracle
 * and every sm_100a kernel.  Three users sit on top of it:
 *   - the PyTorch framework layer (hpnn_b200/, via ctypes: tensors are passed
 *     as raw device pointers, launches go to the caller's CUDA stream),
 *   - the libhpnn.h compatibility shim (csrc/capi/libhpnn.cpp),
 *   - plain C/C++ host programs.
 *
 * Reference parity notes cite /root/reference paths as file:line.
 */
#ifndef HPNN_CORE_H
#define HPNN_CORE_H

#include <stddef.h>
#include <stdint.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HPNN_MAX_LAYERS 16 /* weight layers (hidden layers + output layer)   */
#define HPNN_MAX_WORLD 8   /* ranks (GPUs) in one NVLink domain              */
#define HPNN_TINY 1E-14    /* include/libhpnn/common.h:79                    */

/* network type / trainer: same numeric values as include/libhpnn.h:51-66 */
enum { HPNN_TYPE_ANN = 0, HPNN_TYPE_LNN = 1, HPNN_TYPE_SNN = 2, HPNN_TYPE_UKN = -1 };
enum { HPNN_TRAIN_BP = 0, HPNN_TRAIN_BPM = 1, HPNN_TRAIN_CG = 2, HPNN_TRAIN_SPLX = 3, HPNN_TRAIN_UKN = -1 };

/* ------------------------------------------------------------------------ */
/* host-side network ("kernel" in libhpnn parlance): topology + fp64 weights */
/* ------------------------------------------------------------------------ */
typedef struct hpnn_kernel hpnn_kernel;

/* dims = {n_in, h_1, ..., h_K, n_out}; n_dims = K+2; weights zeroed. */
hpnn_kernel *hpnn_kernel_create(int n_dims, const int *dims);
void hpnn_kernel_destroy(hpnn_kernel *k);
int hpnn_kernel_n_layers(const hpnn_kernel *k);      /* weight layers = n_dims-1 */
int hpnn_kernel_dim(const hpnn_kernel *k, int i);    /* i in [0, n_layers]       */
double *hpnn_kernel_weights(hpnn_kernel *k, int l);  /* row-major [dim(l+1) x dim(l)] */
const char *hpnn_kernel_name(const hpnn_kernel *k);  /* NULL if unnamed          */
void hpnn_kernel_set_name(hpnn_kernel *k, const char *name);
/* glibc srandom/random init, 2(r-0.5)/sqrt(fan_in), layer by layer, row-major
 * (src/ann.c:632-703).  *seed==0 -> time(NULL), written back. */
hpnn_kernel *hpnn_kernel_generate(unsigned *seed, int n_dims, const int *dims);
hpnn_kernel *hpnn_kernel_load(const char *path);            /* src/ann.c:206-631 */
int hpnn_kernel_dump(const hpnn_kernel *k, FILE *fp);       /* src/ann.c:770-857 */
int hpnn_kernel_dump_path(const hpnn_kernel *k, const char *path);
/* exact fp64 sidecar (the text format loses bits) */
int hpnn_kernel_dump_binary(const hpnn_kernel *k, const char *path);
hpnn_kernel *hpnn_kernel_load_binary(const char *path);

/* ------------------------------------------------------------------------ */
/* nn.conf (src/libhpnn.c:658-937)                                           */
/* ------------------------------------------------------------------------ */
typedef struct {
  char *name;
  int type;       /* HPNN_TYPE_*  */
  int need_init;  /* [init] generate */
  unsigned seed;
  char *f_kernel; /* [init] <file>  */
  int train;      /* HPNN_TRAIN_* */
  char *samples;
  char *tests;
  int n_in, n_out, n_hid;
  int *hid;
} hpnn_conf;

/* returns 0 on success; strings/arrays are malloc'd, release with _free */
int hpnn_conf_load(const char *path, hpnn_conf *out);
void hpnn_conf_free(hpnn_conf *c);
/* writer of src/libhpnn.c:885-937; dims may be NULL (sizes printed as 0) */
int hpnn_conf_dump(const hpnn_conf *c, int n_dims, const int *dims, FILE *fp);

/* ------------------------------------------------------------------------ */
/* samples and directory walks (src/libhpnn.c:1070-1145, 1184-1229)          */
/* ------------------------------------------------------------------------ */
/* *in / *out are malloc'd (free with hpnn_free); returns 0 on success */
int hpnn_sample_read(const char *path, double **in, int *n_in, double **out, int *n_out);
int hpnn_sample_write(const char *path, const double *in, int n_in, const double *out, int n_out,
                      const char *comment);
/* parallel loader: file i -> row i of X [n x n_in] and T [n x n_out]; ok[i]=1 if it matched */
int hpnn_sample_read_many(const char *const *paths, int n, int n_in, int n_out, double *X, double *T,
                          unsigned char *ok);
/* readdir() order, dot-files skipped; names is a malloc'd array of malloc'd strings */
int hpnn_dir_list(const char *dir, char ***names, int *n);
void hpnn_dir_list_free(char **names, int n);
/* the reference's srandom(seed) + rejection draw without replacement */
void hpnn_shuffle_order(unsigned seed, int n, int *order);
void hpnn_free(void *p);

/* ------------------------------------------------------------------------ */
/* CPU / OpenMP engine: fallback and test oracle (src/ann.c, src/snn.c)      */
/* ------------------------------------------------------------------------ */
typedef struct {
  double init_err;  /* error of the first forward ("init=")                 */
  double final_de;  /* last E_before - E_after ("final=")                   */
  double last_err;  /* error after the last iteration                       */
  int n_iter;       /* N_ITER                                               */
  int first_ok;     /* argmax matched the target after iteration 1 (OK/NO)  */
  int success;      /* loop-exit is_ok (SUCCESS!/FAIL!)                      */
  int status;       /* device engine: 0 = ran to completion, 1 = abandoned (a wait for another CTA or
                     * rank exceeded the spin time-out: weights and numbers above are meaningless)     */
} hpnn_train_result;

typedef struct {
  int type;        /* HPNN_TYPE_ANN | HPNN_TYPE_SNN                          */
  int train;       /* HPNN_TRAIN_BP | HPNN_TRAIN_BPM                         */
  double lr;       /* <=0 -> backend default                                 */
  double alpha;    /* momentum factor (BPM), reference passes 0.2            */
  double delta;    /* <=0 -> 1e-6                                            */
  int min_iter;    /* <0 -> 31 (BP) / 15 (BPM)                               */
  int max_iter;    /* <0 -> 102399                                           */
  int stable_softmax; /* 0 = reference formula exp(z-1) (parity), 1 = max-shifted */
} hpnn_train_opts;

void hpnn_train_opts_default(hpnn_train_opts *o, int type, int train, int cuda_defaults);
void hpnn_set_omp_threads(int n);
/* out must hold dim(n_layers) doubles */
int hpnn_cpu_forward(const hpnn_kernel *k, int type, int stable_softmax, const double *in, double *out);
/* single-step API (ann_kernel_run / ann_kernel_train / ann_kernel_train_momentum):
 * act(0) is the input buffer, act(i+1) the output of layer i; momentum persists across
 * hpnn_cpu_train_iteration calls until zeroed. */
double *hpnn_cpu_act(hpnn_kernel *k, int i);
double *hpnn_cpu_momentum(hpnn_kernel *k, int l);
void hpnn_cpu_momentum_init(hpnn_kernel *k);
void hpnn_cpu_momentum_zero(hpnn_kernel *k);
void hpnn_cpu_momentum_free(hpnn_kernel *k);
int hpnn_cpu_forward_inplace(hpnn_kernel *k, int type, int stable_softmax);
double hpnn_cpu_error(hpnn_kernel *k, int type, const double *target);
double hpnn_cpu_train_iteration(hpnn_kernel *k, const hpnn_train_opts *o, const double *target);
/* whole per-sample convergence loop */
int hpnn_cpu_train_sample(hpnn_kernel *k, const hpnn_train_opts *o, const double *in,
                          const double *target, hpnn_train_result *res, double *out);

/* ------------------------------------------------------------------------ */
/* device engine (sm_100a)                                                   */
/* ------------------------------------------------------------------------ */
/* All pointers are device pointers owned by the caller.  W rows are stored with
 * a padded leading dimension ld[l] (multiple of 8 doubles, pad = 0) so that
 * every row is 64-byte aligned.  In a row-split (tensor-parallel) job each
 * rank holds rows [row0[l], row0[l]+rows[l]) of every layer.  The "arena" is a
 * per-rank buffer with the SAME layout on every rank; arena[p] is rank p's
 * arena as mapped into this process (peer memory), arena[rank] the local one. */
typedef struct {
  int n_layers;
  int dims[HPNN_MAX_LAYERS + 1];
  int ld[HPNN_MAX_LAYERS];      /* leading dim of W[l], >= dims[l]          */
  int row0[HPNN_MAX_LAYERS];
  int rows[HPNN_MAX_LAYERS];
  double *W[HPNN_MAX_LAYERS];   /* [rows[l] x ld[l]]                        */
  double *dW[HPNN_MAX_LAYERS];  /* momentum, same shape, or NULL            */
  int rank, world;
  double *arena[HPNN_MAX_WORLD];
  /* NVLS multicast address of the arenas (hpnn_symm_mc_ptr), or NULL: a multimem.st to
   * arena_mc + off lands at arena[q] + off on EVERY rank q (all-gather in one store)     */
  double *arena_mc;
  unsigned *local_sync;         /* 256 zeroed u32 words, private to the rank */
  /* arena layout (offsets in doubles), filled by hpnn_dev_arena_layout      */
  long long off_act[HPNN_MAX_LAYERS + 1]; /* act[0]=input, act[l+1]=out of l */
  long long off_part[HPNN_MAX_LAYERS];    /* per-rank partial W^T.delta      */
  long long off_target;
  long long off_out;                      /* final output vector             */
  long long off_flags;                    /* cross-rank barrier flags        */
  long long arena_doubles;                /* total size                      */
  int part_stride[HPNN_MAX_LAYERS];       /* max owned rows of layer l       */
  /* single-rank only: per-CTA partial W^T.delta slots of the shared-memory-resident trainer,
   * cta_slots x ld[l] {value, tag} pairs for layer l >= 1 (0 slots = variant not available) */
  long long off_cta_part[HPNN_MAX_LAYERS];
  int cta_slots;
  /* deferred rank-1 updates of the wide-layer BP trainer (train_lazy.cu): for layer l >= 1 the
   * last hist_depth (<= 8) update factors, u = lr*delta per owned row and v = input per column,
   * stored as 8-double chunks per row / column (0 = variant not available for this shape) */
  long long off_hist_u[HPNN_MAX_LAYERS];
  long long off_hist_v[HPNN_MAX_LAYERS];
  /* per-(rank, CTA) partial dot products v_j . x of layer l >= 1, world x 160 chunks of 8 */
  long long off_hist_dot[HPNN_MAX_LAYERS];
  int hist_depth;
  /* single rank: per-CTA partial W^T.delta vectors of the row-panel variant (train_lazy1.cu),
   * 2 buffers x 160 CTAs x panel_stride doubles (panel_stride = widest ld + 8; 0 = none) */
  long long off_panel;
  int panel_stride;
  /* row-panel trainer, output layers of at most 32 neurons: the output layer is folded into the
   * phases of the layer below it.  off_wlast / off_mlast: working copy of the WHOLE output layer
   * (dims[n_layers] x ld[n_layers-1]) and of its momentum; off_zpart: per-(rank, CTA) partial
   * output sums, world x 160 slots of 32 doubles.  0 = not reserved.                          */
  long long off_wlast, off_mlast, off_zpart;
} hpnn_dev_net;

/* Deferred-update BP trainer for layers much larger than L2 (train_lazy.cu).  depth: write-back
 * period, 2..8; 0 disables the variant; -1 restores the default (HPNN_LAZY / HPNN_LAZY_DEPTH).
 * min_bytes: smallest local weight volume of the layers >= 1 for which hpnn_dev_arena_layout
 * reserves the factor history (-1 = default 64 MiB); set it BEFORE laying out a network. */
void hpnn_dev_set_lazy(int depth, long long min_bytes);
/* one rank only: row-panel organisation of the same trainer (train_lazy1.cu); 1 = on, 0 = off
 * (column tasks, train_lazy.cu), -1 = default (on unless HPNN_LAZY_PANEL=0) */
void hpnn_dev_set_lazy_panel(int on);
/* several ranks, panels of at most 8 rows: forward passes wait per producer rank inside the pass instead of behind
 * a full cross-rank barrier; 1 = on, 0 = off, -1 = default (off unless HPNN_SLICEWISE=1: measured slower) */
void hpnn_dev_set_slicewise(int on);

/* fills ld/row0/rows/offsets from dims/rank/world; returns arena size in doubles */
long long hpnn_dev_arena_layout(hpnn_dev_net *net);

/* plain device memory helpers (cudaMalloc/cudaFree, zero-filled) + CUDA IPC */
int hpnn_cuda_device_count(void);
int hpnn_cuda_set_device(int dev);
void *hpnn_dev_alloc(size_t bytes);
void hpnn_dev_free(void *p);
int hpnn_dev_memset(void *p, int v, size_t bytes, void *stream);
int hpnn_dev_h2d(void *dst, const void *src, size_t bytes, void *stream);
int hpnn_dev_d2h(void *dst, const void *src, size_t bytes, void *stream);
int hpnn_dev_sync(void *stream);
void *hpnn_host_alloc_pinned(size_t bytes);
void hpnn_host_free_pinned(void *p);
int hpnn_ipc_export(void *dev_ptr, unsigned char handle[64]);
void *hpnn_ipc_open(const unsigned char handle[64]);
int hpnn_ipc_close(void *p);
int hpnn_enable_peer_access(int dev, int peer);

/* Symmetric memory of one NVLink / NVSwitch domain (csrc/core/symm.cu): the same-sized allocation on
 * every rank, each rank's copy mapped into every participant, plus -- where the fabric has NVLS -- one
 * multicast mapping through which a single multimem.st / multimem.red reaches all copies and a
 * multimem.ld_reduce returns their sum.  Replaces the reference's gather-to-GPU-0 + broadcast copies
 * (src/cuda_ann.cu:610-666) and its MPI_Allgather calls (src/ann.c:925).
 * local:  one process drives all devices.  shared: one process per device; the allocation of every rank
 * and the multicast object travel between the processes as POSIX file descriptors (the caller moves
 * them: SCM_RIGHTS over a unix socket, hpnn_b200/parallel/symm.py).  Order for the multicast object:
 * rank 0 mc_create -> others mc_import -> all mc_add_device -> barrier -> all mc_bind; any failure:
 * every rank calls mc_drop and the peer mappings are used alone.  NULL / -1 on failure
 * (hpnn_last_error()); callers then fall back to cudaMalloc + CUDA IPC. */
typedef struct hpnn_symm hpnn_symm;
int hpnn_symm_multicast_supported(int dev);
hpnn_symm *hpnn_symm_create_local(const int *devs, int n, size_t bytes, int want_multicast);
hpnn_symm *hpnn_symm_create_shared(int dev, int rank, int world, size_t bytes, int want_multicast);
int hpnn_symm_export_fd(hpnn_symm *s);
int hpnn_symm_import_peer(hpnn_symm *s, int peer, int fd);
int hpnn_symm_mc_create(hpnn_symm *s);
int hpnn_symm_mc_import(hpnn_symm *s, int fd);
int hpnn_symm_mc_add_device(hpnn_symm *s);
int hpnn_symm_mc_bind(hpnn_symm *s);
void hpnn_symm_mc_drop(hpnn_symm *s);
void *hpnn_symm_ptr(const hpnn_symm *s, int rank); /* rank's copy as mapped here (zero-filled)        */
void *hpnn_symm_mc_ptr(const hpnn_symm *s);        /* multicast address of the set, or NULL           */
size_t hpnn_symm_size(const hpnn_symm *s);
void hpnn_symm_destroy(hpnn_symm *s);
const char *hpnn_last_error(void);

/* persistent per-sample kernel (K-P).  Input must already be in act[0] and the
 * target in off_target of the LOCAL arena (every rank loads the same sample).
 * result is a device pointer.  mode 0: train to convergence, 1: forward only
 * (the softmaxed / activated output lands at off_out).  Returns 0 on success. */
int hpnn_dev_train_sample(const hpnn_dev_net *net, const hpnn_train_opts *o, int mode,
                          hpnn_train_result *result_dev, void *stream);
/* Virtual ranks: the same training step for ALL `world` ranks of a row-split job in ONE cooperative
 * launch on the current device (each rank gets SMs / world CTAs).  nets[r] is rank r's descriptor:
 * own weight slice, own arena, own sync words, arena[q] = rank q's arena -- all on this device.
 * For verifying the multi-rank protocols where one GPU is available (the reference's counterpart
 * is the DEBUG build that fakes three GPUs on device 0, include/libhpnn/common.h:511-572);
 * results_dev[r] may be NULL.  Not a performance path. */
int hpnn_dev_train_sample_vranks(const hpnn_dev_net *nets, int world, const hpnn_train_opts *o, int mode,
                                 hpnn_train_result *const *results_dev, void *stream);
/* Longest time any single wait inside the persistent kernels (grid barrier, cross-rank flag, tagged
 * pair) may last before the whole job is abandoned: every rank leaves its kernel and reports
 * hpnn_train_result.status = 1 instead of hanging (a rank whose launch failed, a dead peer process).
 * ms < 0: default (HPNN_SPIN_TIMEOUT_MS, else 10 000); 0: never give up.  The reference has no such
 * wait: it synchronises through the host and exits on any CUDA error (include/libhpnn/common.h:337-353). */
void hpnn_dev_set_spin_timeout_ms(long long ms);
/* Fault injection for tests: kind 1 = the kernels launched for rank `rank` return at once, as if that
 * rank's launch had failed; kind 0 = off. */
void hpnn_dev_set_fault(int kind, int rank);
/* Schedule perturbation for tests: ns > 0 makes every barrier arrival, cross-rank signal and tagged store of the
 * trainers wait a pseudo-random time of up to ns nanoseconds first.  All summation orders are fixed, so the
 * results must not change by a single bit (tests/test_gpu_vranks.py); 0 = off. */
void hpnn_dev_set_jitter_ns(int ns);
/* number of co-resident CTAs the persistent kernel will use on this device */
int hpnn_dev_persistent_grid(void);
/* set grid (CTAs) used by K-P: 0 = one per SM (default) */
void hpnn_dev_set_persistent_grid(int ctas);
/* observability: when set, CTA 0 of K-P appends (tag, clock64) pairs at every phase boundary
 * (tags: 1 iteration start, 100+l column phase l done, 200+l its barrier passed, 300+l row
 * phase l done, 400+l its barrier passed, 2 evaluation done).  NULL switches it off. */
void hpnn_dev_set_phase_trace(long long *dev_buf, int capacity_pairs);

/* stand-alone phase kernels (K-A..K-E), one launch each, single rank or one
 * rank's slice.  Used by unit tests, by the multi-launch trainer and by the
 * NCCL-collective baseline of the row split. */
/* y[r] = f(W[r,:].x): act_kind 0 = 2/(1+e^-x)-1, 1 = identity               */
int hpnn_dev_forward_layer(const double *W, int ld, int rows, int cols, const double *x, double *y,
                           int act_kind, void *stream);
/* output statistics: type ANN: o given; SNN: z given and o (softmax) written.
 * delta_out[j] (j<n), stats[0]=E, stats[1]=argmax, stats[2]=target index     */
int hpnn_dev_output_stats(int type, int stable_softmax, const double *z, const double *target, int n,
                          double *o, double *delta_out, double *stats, void *stream);
/* part[c] = sum_r W[r,c]*delta[r] (old W), then optionally the BP/BPM update
 * W[r,c] += lr*delta[r]*h[c] in the same pass.  update: 0 none, 1 BP, 2 BPM   */
int hpnn_dev_backward_update(double *W, double *dW, int ld, int rows, int cols, const double *delta,
                             const double *h, double *part, int update, double lr, double alpha,
                             void *stream);
/* delta[r] = dact(h[r]) * sum_g part[g*stride + r]                            */
int hpnn_dev_delta_finalize(const double *part, int n_parts, int stride, const double *h, int n,
                            double *delta, void *stream);
/* first layer: W[r,:] += lr*delta[r]*x (or BPM triple), then y[r] = f(W[r,:].x) */
int hpnn_dev_update_forward(double *W, double *dW, int ld, int rows, int cols, const double *delta,
                            const double *x, double *y, int update, double lr, double alpha,
                            int act_kind, void *stream);

/* batched inference (extension): Y[B x N] = f(X[B x M] . W^T), fp64 */
int hpnn_dev_forward_batch_f64(const double *W, int ld, int rows, int cols, const double *X, int ldx,
                               int batch, double *Y, int ldy, int act_kind, void *stream);
/* row-split layer of a batched forward: GEMM whose epilogue stores the result into EVERY rank's
 * activation matrix at column col0 (all-gather over NVLink peer memory fused into the GEMM),
 * and the one-CTA cross-rank barrier that separates two such layers.  Y_all[p] / flags_all[p]
 * are rank p's buffers as mapped into this process; epoch must grow by one per call. */
int hpnn_dev_forward_batch_f64_tp(const double *W, int ld, int rows, int cols, const double *X, int ldx, int batch,
                                  double *const *Y_all, int n_ranks, int col0, int ldy, int act_kind, void *stream);
/* same with the all-gather through the NVLS multicast mapping of the activation matrices (Y_mc from
 * hpnn_symm_mc_ptr, NULL = peer stores): one 16-byte multimem.st per pair of results */
int hpnn_dev_forward_batch_f64_tp_mc(const double *W, int ld, int rows, int cols, const double *X, int ldx, int batch,
                                     double *const *Y_all, double *Y_mc, int n_ranks, int col0, int ldy, int act_kind,
                                     void *stream);
int hpnn_dev_world_barrier(void *const *flags_all, int rank, int world, unsigned long long epoch, void *stream);
/* opt-in reduced precision (K-T): bf16 operands, fp32 accumulation on the tcgen05 tensor cores
 * (TMA -> shared memory -> tcgen05.mma -> TMEM -> epilogue), FP64 result.  Never used for parity.
 * hpnn_dev_cast_bf16 converts an fp64 matrix (zero-filling up to ld_dst); Wb/Xb must be 16-byte
 * aligned with leading dimensions that are multiples of 8. */
int hpnn_dev_cast_bf16(const double *src, int ld_src, int rows, int cols, void *dst_bf16, int ld_dst, void *stream);
int hpnn_dev_gemm_bf16(const void *Wb, int ldw, int rows, int cols, const void *Xb, int ldx, int batch, double *Y,
                       int ldy, int act_kind, void *stream);
/* same GEMM with a bf16 result [batch x ldyb] that feeds the next layer's GEMM directly (hidden layers) */
int hpnn_dev_gemm_bf16_bf16(const void *Wb, int ldw, int rows, int cols, const void *Xb, int ldx, int batch, void *Yb,
                            int ldyb, int act_kind, void *stream);
/* tile mapping of K-T: 2 pairs of CTAs on one 256 x 256 tile (tcgen05.mma.cta_group::2, default), 1 pairs of
 * CTAs that share each block of W through TMA multicast within a cluster, 0 one CTA per tile, -1 environment
 * HPNN_KT_CLUSTER.  The pair mappings are used for operands in this GPU's own memory and two or more batch tiles. */
void hpnn_dev_set_kt_cluster(int mode);
/* diagnostic: CTA 0 of every following K-T launch writes clock64 stamps into buf (>= 160 int64 on the device); NULL: off */
void hpnn_dev_set_kt_trace(void *buf);
/* row-wise softmax exp(z-1)/(TINY+sum) on a [batch x n] matrix (in place) */
int hpnn_dev_softmax_rows(double *Z, int ldz, int batch, int n, int stable, void *stream);

/* data-parallel gradient merge (extension): W += sum_ranks G_rank */
int hpnn_dev_grad_merge(double *W, long long n, double *const *grads, int n_ranks, double scale,
                        void *stream);

/* Factored form of the same merge (csrc/kernels/dp_merge.cu): a BP update is lr*delta (x) h, so the
 * ranks exchange the FACTORS (O(N+M) per layer) instead of gradient matrices (O(N*M)).
 * dp_publish copies `n` doubles into this rank's slot of every rank's factor buffer (dst_all[q], or one
 * multimem.st per element through dst_mc when non-NULL), then fence.sys and sets this rank's step flag
 * on every rank.  dp_apply waits on the device until flags[0..world) >= step (timeout_ms > 0: gives up,
 * leaves W untouched and sets *status_dev = 1) and adds sum_r U[r*u_stride + i] * V[r*v_stride + j] to
 * W[i][j], ranks in index order.  grad_merge_nvls is the matrix form through NVLS: rank r reduces slice
 * r of all gradient buffers with multimem.ld_reduce and writes the new weights into every replica with
 * multimem.st (W_mc / G_mc: multicast addresses from hpnn_symm_mc_ptr).  No reference counterpart. */
int hpnn_dev_dp_publish(const double *src, long long n, double *const *dst_all, double *dst_mc,
                        unsigned long long *const *flag_all, unsigned long long *flag_mc, int world,
                        unsigned long long step, void *stream);
int hpnn_dev_dp_apply(double *W, int ld, int rows, int cols, const double *U, long long u_stride, const double *V,
                      long long v_stride, int world, const unsigned long long *flags, unsigned long long step,
                      long long timeout_ms, int *status_dev, void *stream);
int hpnn_dev_grad_merge_nvls(const double *W_local, double *W_mc, const double *G_mc, long long n, int rank, int world,
                             double scale, void *stream);

/* ------------------------------------------------------------------------ */
/* hpnn_net: device-resident network driven from one host process            */
/* ------------------------------------------------------------------------ */
/* Owns its device memory.  n_gpus <= 0 means "all visible GPUs": the layers are then
 * row-split over them exactly as in the multi-process PyTorch path (one persistent kernel
 * per GPU per sample, collectives issued from inside the kernels over peer memory). */
typedef struct hpnn_net hpnn_net;
hpnn_net *hpnn_net_create(const hpnn_kernel *k, int type, int momentum, int n_gpus);
/* same, on devices first_dev .. first_dev + n_gpus - 1 (one process per GPU under torchrun / mpirun:
 * first_dev = the local rank, n_gpus = 1) */
hpnn_net *hpnn_net_create_on(const hpnn_kernel *k, int type, int momentum, int n_gpus, int first_dev);
void hpnn_net_destroy(hpnn_net *n);
int hpnn_net_world(const hpnn_net *n);
void hpnn_net_set_type(hpnn_net *n, int type);
long long hpnn_net_device_bytes(const hpnn_net *n);
int hpnn_net_upload(hpnn_net *n, const hpnn_kernel *k);   /* host weights -> device */
int hpnn_net_download(hpnn_net *n, hpnn_kernel *k);       /* device weights -> host */
/* per-sample convergent training (K-P) / single forward; `out` may be NULL */
int hpnn_net_train_sample(hpnn_net *n, const hpnn_train_opts *o, const double *in, const double *target,
                          hpnn_train_result *res, double *out);
int hpnn_net_run(hpnn_net *n, const double *in, double *out, int stable_softmax);
/* batched forward, host matrices X [batch x n_in] -> Y [batch x n_out] (single-GPU nets; FP64 DMMA path) */
int hpnn_net_forward_batch(hpnn_net *n, const double *X, int batch, double *Y, int stable_softmax);

/* Forward pass for callers that must not allocate device memory.
 * (No counterpart in the reference: its CUDA path calls cudaMalloc per sample and per
 * iteration, src/ann.c:2305-2310, src/cuda_ann.cu:2114-2117.)
 * Hosts that gave the whole GPU to their own pool, or that cannot afford the device-wide
 * synchronisation cudaMalloc/cudaFree imply, pass the padded weight matrices, the input and
 * the output buffer (all device memory they already own); the persistent kernel's work space
 * is a fixed array inside the library instead: HPNN_NOALLOC_SCRATCH_BYTES on every GPU the
 * entry point has been called on, held until the process exits.
 *   type / stable_softmax  as in hpnn_train_opts
 *   dims[0..n_layers]      layer widths; W[l] is [dims[l+1] x pad8(dims[l])], pad columns zero
 *   x (dims[0] doubles)    input;  out (dims[n_layers] doubles) receives the activated /
 *                          soft-maxed output; the call is asynchronous on `stream`
 * Returns 0, -1 on bad arguments or a CUDA error, -2 when the network's arena layout needs
 * more than HPNN_NOALLOC_MAX_DOUBLES doubles (nothing is launched; see hpnn_last_error()):
 *   need = pad8(dims[0]) + 12 * sum_{i>=1} pad8(dims[i]) + 2 * pad8(dims[n_layers]) + 16,
 * i.e. the hidden and output widths may add up to about 10 900.
 * It calls no cudaMalloc/cudaFree.  All calls on a device share the one array: calls whose
 * GPU work overlaps race, so keep them on one stream.  The first call on each device resolves
 * the array's address (cudaGetSymbolAddress) and, because CUDA loads modules lazily, is also
 * the moment the driver reserves the array and the kernel image: make that call at start-up,
 * before any stream capture begins and before the host hands the rest of the GPU to its own
 * pool (or run with CUDA_MODULE_LOADING=EAGER).  The cached address does not survive a
 * cudaDeviceReset by the host. */
#define HPNN_NOALLOC_MAX_DOUBLES (128 * 1024)
#define HPNN_NOALLOC_SCRATCH_BYTES (HPNN_NOALLOC_MAX_DOUBLES * 8 + 1024)
int hpnn_dev_forward_noalloc(int type, int stable_softmax, int n_layers, const int *dims,
                             const double *const *W, const double *x, double *out, void *stream);

/* launch counter: number of kernels this library launched since the last reset */
long long hpnn_dev_launch_count(int reset);
/* sizeof of the ABI structs as this library was compiled, for bindings that mirror them
 * (0 hpnn_train_result, 1 hpnn_train_opts, 2 hpnn_conf, 3 hpnn_dev_net; -1 = unknown index) */
long long hpnn_abi_sizeof(int which);

#ifdef __cplusplus
}
#endif
#endif /* HPNN_CORE_H */
