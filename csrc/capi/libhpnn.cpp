// libhpnn.so: the libhpnn v0.2 C API (include/libhpnn.h, libhpnn/ann.h, libhpnn/snn.h)
// implemented on the hpnn-b200 core.  Behavioural reference: src/libhpnn.c (runtime, knobs,
// configuration, drivers), src/ann.c / src/snn.c (model layer).  No code is shared with it:
// every kernel object is a view over an hpnn_kernel (host weights + CPU work space) plus, when
// a GPU is present, an hpnn_net that keeps the live weights on the device(s).
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "hpnn_core.h"
#include "libhpnn.h"
#include "libhpnn/ann.h"
#include "libhpnn/snn.h"
#include "libhpnn/cuda_ann.h"
#include "libhpnn/cuda_snn.h"

extern "C" {
nn_runtime lib_runtime;  // one runtime per process, as in the reference (src/libhpnn.c:60)
}

namespace {

// The object behind nn_def.kernel / kernel_ann*.  Kernels made by ann_generate / ann_load /
// nn_load_conf hand out &own; a kernel_ann the CALLER allocated and passed to ann_kernel_allocate
// (the reference's contract, include/libhpnn/ann.h:59, src/ann.c:113-202) is tracked by its address
// and refreshed in place.  Nothing is ever read behind a kernel_ann* that is not in the registry.
struct Shim {
  kernel_ann own;
  kernel_ann *pub = nullptr;  // &own, or the caller's struct
  hpnn_kernel *host = nullptr;
  hpnn_net *dev = nullptr;
  bool dev_newer = false;   // device weights are ahead of the host mirror
  bool dev_stale = false;   // host weights changed after the upload (CPU-side training)
  bool dev_momentum = false;
  int dev_gpus = 0;
  std::vector<layer_ann> hidden_views;
  std::vector<double *> dw_views;
  std::string name;
};

std::mutex g_reg_mu;
std::unordered_map<const kernel_ann *, Shim *> g_registry;

// the Shim behind a public kernel pointer, or nullptr for pointers this library does not know
Shim *as_shim(const kernel_ann *k) {
  if (!k) return nullptr;
  std::lock_guard<std::mutex> g(g_reg_mu);
  auto it = g_registry.find(k);
  return it == g_registry.end() ? nullptr : it->second;
}
void register_shim(Shim *s) {
  std::lock_guard<std::mutex> g(g_reg_mu);
  g_registry[s->pub] = s;
}
void unregister_shim(Shim *s) {
  std::lock_guard<std::mutex> g(g_reg_mu);
  g_registry.erase(s->pub);
}

int g_gpu_cap = 0;          // nn_set_n_gpu: 0 = all
int g_first_dev = 0;        // one process per GPU (torchrun / mpirun): this process's device
bool g_cuda_ready = false;  // nn_init_CUDA succeeded

bool use_device() {
  return g_cuda_ready && (lib_runtime.capability & NN_CAP_CUDA) && lib_runtime.cudas.n_gpu > 0;
}

void refresh_views(Shim *s) {
  hpnn_kernel *h = s->host;
  const int NL = hpnn_kernel_n_layers(h);
  kernel_ann &k = *s->pub;
  k.n_inputs = (UINT)hpnn_kernel_dim(h, 0);
  k.n_hiddens = (UINT)(NL - 1);
  k.n_outputs = (UINT)hpnn_kernel_dim(h, NL);
  k.in = hpnn_cpu_act(h, 0);
  s->hidden_views.resize(NL - 1);
  UINT mx = k.n_inputs > k.n_outputs ? k.n_inputs : k.n_outputs;
  for (int l = 0; l < NL - 1; ++l) {
    layer_ann &v = s->hidden_views[l];
    v.n_neurons = (UINT)hpnn_kernel_dim(h, l + 1);
    v.n_inputs = (UINT)hpnn_kernel_dim(h, l);
    v.weights = hpnn_kernel_weights(h, l);
    v.vec = hpnn_cpu_act(h, l + 1);
    if (v.n_neurons > mx) mx = v.n_neurons;
  }
  k.hiddens = s->hidden_views.data();
  k.output.n_neurons = k.n_outputs;
  k.output.n_inputs = (UINT)hpnn_kernel_dim(h, NL - 1);
  k.output.weights = hpnn_kernel_weights(h, NL - 1);
  k.output.vec = hpnn_cpu_act(h, NL);
  k.max_index = mx;
  const char *nm = hpnn_kernel_name(h);
  if (nm) {
    s->name = nm;
    k.name = &s->name[0];
  } else {
    k.name = nullptr;
  }
  s->dw_views.assign(NL, nullptr);
  bool have_dw = hpnn_cpu_momentum(h, 0) != nullptr;
  for (int l = 0; l < NL; ++l) s->dw_views[l] = hpnn_cpu_momentum(h, l);
  k.dw = have_dw ? s->dw_views.data() : nullptr;
  k.tmp_cpu = k.output.vec;
  k.tmp_gpu = nullptr;
  k.kerns = nullptr;
}

Shim *wrap(hpnn_kernel *h) {
  if (!h) return nullptr;
  Shim *s = new Shim();
  memset(&s->own, 0, sizeof(s->own));
  s->pub = &s->own;
  s->host = h;
  register_shim(s);
  refresh_views(s);
  const int NL = hpnn_kernel_n_layers(h);
  unsigned long cpu_bytes = 0;
  for (int l = 0; l < NL; ++l)
    cpu_bytes += sizeof(double) * ((unsigned long)hpnn_kernel_dim(h, l) + 1) * hpnn_kernel_dim(h, l + 1);
  NN_OUT(stdout, "[CPU] ANN total allocation: %lu (bytes)\n", cpu_bytes);
  if (use_device()) {
    // reported here, like the reference does at allocation time (src/ann.c:197-200); the device
    // copy itself is created on first use: padded rows + activation / partial buffers
    unsigned long gpu_bytes = 0;
    for (int l = 0; l < NL; ++l)
      gpu_bytes += sizeof(double) * (unsigned long)((hpnn_kernel_dim(h, l) + 7) & ~7) * hpnn_kernel_dim(h, l + 1);
    NN_OUT(stdout, "[GPU] ANN total allocation: %lu (bytes)\n", gpu_bytes);
  }
  return s;
}

void drop_device(Shim *s) {
  if (s->dev) hpnn_net_destroy(s->dev);
  s->dev = nullptr;
  s->dev_newer = s->dev_stale = false;
}

// make the host mirror current
void sync_to_host(Shim *s) {
  if (s->dev && s->dev_newer) {
    hpnn_net_download(s->dev, s->host);
    s->dev_newer = false;
  }
}

// make a device net current (creating / re-uploading as needed); nullptr -> use the CPU engine
hpnn_net *device_for(Shim *s, int type, bool momentum) {
  if (!use_device()) return nullptr;
  const int want_gpus = g_gpu_cap > 0 ? g_gpu_cap : (int)lib_runtime.cudas.n_gpu;
  if (s->dev && ((momentum && !s->dev_momentum) || s->dev_gpus != want_gpus)) {
    sync_to_host(s);
    drop_device(s);
  }
  if (!s->dev) {
    s->dev = hpnn_net_create_on(s->host, type, momentum, want_gpus, g_first_dev);
    if (!s->dev) {
      NN_WARN(stderr, "device engine unavailable (%s): using the CPU engine\n", hpnn_last_error());
      return nullptr;
    }
    s->dev_momentum = momentum;
    s->dev_gpus = want_gpus;
  } else if (s->dev_stale) {
    hpnn_net_upload(s->dev, s->host);
  }
  s->dev_stale = false;
  hpnn_net_set_type(s->dev, type);
  return s->dev;
}

void fill_opts(hpnn_train_opts *o, int type, int train, double alpha, double delta, bool on_device) {
  hpnn_train_opts_default(o, type, train, on_device ? 1 : 0);
  o->alpha = alpha;
  if (delta > 0.) o->delta = delta;
  const char *lr = getenv("HPNN_LR");  // the reference's rates differ per back-end
  if (lr && atof(lr) > 0.) o->lr = atof(lr);
}

// per-sample convergence loop + the reference's log line (src/ann.c:2322-2366)
double train_sample(kernel_ann *kernel, int type, int train, DOUBLE *in, DOUBLE *out, DOUBLE alpha, DOUBLE delta) {
  Shim *s = as_shim(kernel);
  if (!s || !s->host) {
    NN_ERROR(stderr, "training step failed: not a kernel of this library\n");
    return 0.;
  }
  hpnn_train_result r;
  memset(&r, 0, sizeof(r));
  hpnn_train_opts o;
  hpnn_net *dev = device_for(s, type, train == HPNN_TRAIN_BPM);
  fill_opts(&o, type, train, alpha, delta, dev != nullptr);
  int rc;
  if (dev) {
    rc = hpnn_net_train_sample(dev, &o, in, out, &r, kernel->output.vec);
    s->dev_newer = true;
    memcpy(kernel->in, in, sizeof(double) * kernel->n_inputs);
  } else {
    sync_to_host(s);  // the device may hold newer weights (CUDA switched off after device training)
    rc = hpnn_cpu_train_sample(s->host, &o, in, out, &r, nullptr);
    if (s->dev) s->dev_stale = true;
    refresh_views(s);
  }
  if (rc != 0) {
    NN_ERROR(stderr, "training step failed: %s\n", hpnn_last_error());
    return 0.;
  }
  NN_COUT(stdout, " init=%15.10f", r.init_err);
  NN_COUT(stdout, r.first_ok ? " OK" : " NO");
  NN_COUT(stdout, " N_ITER=%8i", r.n_iter);
  NN_COUT(stdout, " final=%15.10f", r.final_de);
  if (type == HPNN_TYPE_SNN && train == HPNN_TRAIN_BP) {
    NN_COUT(stdout, "\n");  // snn_train_BP prints no verdict (src/snn.c:1496-1497)
  } else {
    NN_COUT(stdout, r.success ? " SUCCESS!\n" : " FAIL!\n");
  }
  fflush(stdout);
  return r.final_de;
}

void run_once(kernel_ann *kernel, int type) {
  Shim *s = as_shim(kernel);
  if (!s || !s->host) return;
  hpnn_net *dev = device_for(s, type, s->dev_momentum);
  if (dev) {
    // only the output vector comes back; the hidden layers' `vec` keep their previous contents
    // on the device path (the reference's ann_kernel_run fills every vec, src/ann.c:892-1242)
    hpnn_net_run(dev, kernel->in, kernel->output.vec, 0);
  } else {
    sync_to_host(s);
    hpnn_cpu_forward_inplace(s->host, type, 0);
  }
}

double single_iteration(kernel_ann *kernel, int type, int train, const DOUBLE *target, DOUBLE alpha) {
  // the single-iteration entry points work on the host mirror (momentum must persist between
  // calls, which the per-sample device kernel deliberately does not do)
  Shim *s = as_shim(kernel);
  if (!s || !s->host) return 0.;
  sync_to_host(s);
  hpnn_train_opts o;
  fill_opts(&o, type, train, alpha, -1., use_device());
  const double d = hpnn_cpu_train_iteration(s->host, &o, target);
  if (s->dev) s->dev_stale = true;
  refresh_views(s);
  return d;
}

char *dup_or_null(const char *p) {
  if (!p || !*p) return nullptr;
  char *d = (char *)calloc(strlen(p) + 1, 1);
  strcpy(d, p);
  return d;
}

std::vector<std::string> listed_files(const char *dir) {
  std::vector<std::string> v;
  char **names = nullptr;
  int n = 0;
  if (hpnn_dir_list(dir, &names, &n) != 0) return v;
  for (int i = 0; i < n; ++i) v.push_back(names[i]);
  hpnn_dir_list_free(names, n);
  return v;
}

// ---- several processes started by torchrun / mpirun / srun (the reference's MPI build, src/libhpnn.c:
// 182-200, src/ann.c:242-245, 549-614): rank-0-only output and file writing, ONE GPU per process, and
// every rank ends up with the same kernel.  The reference broadcasts the kernel rank 0 read or
// generated; here every rank reads the same file / draws the same glibc random() stream, so only the
// seed has to be agreed on when the configuration says "0 = pick one" (rank 0 picks, the others read
// it from a file named after the launcher's process id, which all local ranks share).
int env_int(const char *const *keys, int n_keys, int fallback) {
  for (int i = 0; i < n_keys; ++i) {
    const char *v = getenv(keys[i]);
    if (v && *v) return atoi(v);
  }
  return fallback;
}
int launcher_world() {
  static const char *k[] = {"WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", "SLURM_NTASKS"};
  return env_int(k, 4, 1);
}
int launcher_rank() {
  static const char *k[] = {"RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK", "SLURM_PROCID"};
  return env_int(k, 4, 0);
}
int launcher_local_rank() {
  static const char *k[] = {"LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "MPI_LOCALRANKID", "SLURM_LOCALID"};
  return env_int(k, 4, launcher_rank());
}
unsigned agreed_seed(unsigned seed) {
  if (seed != 0 || launcher_world() < 2) return seed ? seed : (unsigned)time(NULL);
  static unsigned serial = 0;  // several "pick one" requests per run: one file each
  char path[256];
  snprintf(path, sizeof(path), "/tmp/.hpnn_seed_%ld_%u", (long)getppid(), serial++);
  if (launcher_rank() == 0) {
    const unsigned s = (unsigned)time(NULL);
    char tmp[280];
    snprintf(tmp, sizeof(tmp), "%s.tmp", path);
    FILE *fp = fopen(tmp, "w");
    if (fp) {
      fprintf(fp, "%u\n", s);
      fclose(fp);
      rename(tmp, path);
    }
    return s;
  }
  for (int tries = 0; tries < 3000; ++tries) {  // up to 30 s
    FILE *fp = fopen(path, "r");
    unsigned s = 0;
    if (fp) {
      const int got = fscanf(fp, "%u", &s);
      fclose(fp);
      if (got == 1 && s != 0) return s;
    }
    usleep(10000);
  }
  NN_WARN(stderr, "rank %i: no seed from rank 0, using the clock\n", launcher_rank());
  return (unsigned)time(NULL);
}

}  // namespace

extern "C" {

int hpnn_compat_is_root(void) {
  static int cached = -1;
  if (cached < 0) {
    cached = 1;
    const char *keys[] = {"RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK", "SLURM_PROCID"};
    for (const char *k : keys) {
      const char *v = getenv(k);
      if (v && *v) {
        cached = atoi(v) == 0;
        break;
      }
    }
  }
  return cached;
}

/* ------------------------------- runtime -------------------------------- */
void nn_inc_verbose() {
  if (lib_runtime.nn_verbose > 2) return;
  lib_runtime.nn_verbose++;
  NN_DBG(stdout, "verbosity set to %i.\n", lib_runtime.nn_verbose);
}
void nn_dec_verbose() {
  if (lib_runtime.nn_verbose < 1) return;
  lib_runtime.nn_verbose--;
}
void nn_set_verbose(SHORT v) { lib_runtime.nn_verbose = v; }
void nn_get_verbose(SHORT *v) { *v = lib_runtime.nn_verbose; }
SHORT nn_return_verbose() { return lib_runtime.nn_verbose; }
void nn_toggle_dry() { lib_runtime.nn_dry = !lib_runtime.nn_dry; }  // the reference's toggle is a no-op (quirk 1); the flag is never read

nn_cap nn_return_capabilities() {
  unsigned res = NN_CAP_OMP;
  if (launcher_world() > 1) res |= NN_CAP_MPI;
  if (hpnn_cuda_device_count() > 0) res |= NN_CAP_CUDA;
  return (nn_cap)res;
}
void nn_get_capabilities(nn_cap *c) { *c = nn_return_capabilities(); }
void nn_unset_capability(nn_cap c) { lib_runtime.capability = (nn_cap)(lib_runtime.capability & ~(unsigned)c); }

void nn_init_runtime() {
  lib_runtime.capability = nn_return_capabilities();
  lib_runtime.nn_verbose = 0;
  lib_runtime.nn_dry = FALSE;
  lib_runtime.nn_num_threads = 1;
  lib_runtime.nn_num_blas = 1;
  lib_runtime.nn_num_tasks = 1;
  lib_runtime.cudas.n_gpu = 1;
  lib_runtime.cudas.cuda_handle = NULL;
  lib_runtime.cudas.cuda_n_streams = 1;
  lib_runtime.cudas.cuda_streams = NULL;
  lib_runtime.cudas.mem_model = CUDA_MEM_NONE;
}
BOOL nn_init_OMP() {
  NN_OUT(stdout, "NN: OMP init done.\n");
  return TRUE;
}
BOOL nn_init_MPI() {
  if (launcher_world() < 2) {
    NN_WARN(stdout, "failed to init MPI (no capability).\n");
    return FALSE;
  }
  lib_runtime.nn_num_tasks = (UINT)launcher_world();
  NN_OUT(stdout, "MPI started %i tasks.\n", lib_runtime.nn_num_tasks);
  return TRUE;
}
BOOL nn_init_CUDA() {
  const int n = hpnn_cuda_device_count();
  if (n < 1) {
    NN_WARN(stderr, "CUDA error: no CUDA-capable device reported.\n");
    g_cuda_ready = false;
    return FALSE;
  }
  if (launcher_world() > 1) {
    // one process per GPU: this rank drives device (local rank mod devices) and nothing else; the
    // ranks do not split a net between them (each holds the whole net on its GPU, all train the
    // same samples), which is what the reference's MPI + CUDA combination amounts to as well
    // (its device code bypasses MPI, src/ann.c:893-896)
    g_first_dev = launcher_local_rank() % n;
    hpnn_cuda_set_device(g_first_dev);
    lib_runtime.cudas.n_gpu = 1;
    NN_WARN(stdout, "CUDA started, found %i GPU(s).\n", n);
    NN_OUT(stdout, "task %i of %i uses GPU[%i].\n", launcher_rank(), launcher_world(), g_first_dev);
    lib_runtime.cudas.mem_model = CUDA_MEM_NONE;
    g_cuda_ready = true;
    return TRUE;
  }
  lib_runtime.cudas.n_gpu = (UINT)n;
  NN_WARN(stdout, "CUDA started, found %i GPU(s).\n", n);
  lib_runtime.cudas.mem_model = n > 1 ? CUDA_MEM_P2P : CUDA_MEM_NONE;
  if (n > 1) NN_DBG(stdout, "multi-GPU: weights are row-partitioned, peers exchange vectors only\n");
  g_cuda_ready = true;
  return TRUE;
}
BOOL nn_init_BLAS() {
  NN_WARN(stdout, "NOT USING BLAS.\n");
  return FALSE;
}
int nn_init_all(UINT init_verbose) {
  BOOL ok = FALSE;
  nn_init_runtime();
  lib_runtime.nn_verbose = (SHORT)init_verbose;
  const unsigned cap = lib_runtime.capability;
  if (cap & NN_CAP_MPI) ok |= nn_init_MPI();
  if (cap & NN_CAP_OMP) ok |= nn_init_OMP();
  if (cap & NN_CAP_CUDA) ok |= nn_init_CUDA();
  lib_runtime.nn_verbose = 0;
  return ok ? 0 : -1;
}
BOOL nn_deinit_OMP() { return TRUE; }
BOOL nn_deinit_MPI() { return (lib_runtime.capability & NN_CAP_MPI) ? TRUE : FALSE; }
BOOL nn_deinit_CUDA() {
  const BOOL was = g_cuda_ready ? TRUE : FALSE;
  g_cuda_ready = false;
  return was;
}
BOOL nn_deinit_BLAS() { return FALSE; }
int nn_deinit_all() {
  BOOL ok = nn_deinit_OMP();
  ok |= nn_deinit_MPI();
  ok |= nn_deinit_CUDA();
  return ok ? 0 : -1;
}

/* --------------------------------- knobs --------------------------------- */
BOOL nn_set_omp_threads(UINT n) {
  lib_runtime.nn_num_threads = n;
  hpnn_set_omp_threads((int)n);
  return TRUE;
}
BOOL nn_get_omp_threads(UINT *n) {
  *n = lib_runtime.nn_num_threads;
  return TRUE;
}
int nn_return_omp_threads() { return (int)lib_runtime.nn_num_threads; }
BOOL nn_set_mpi_tasks(UINT) {
  NN_WARN(stdout, "Changing MPI num_tasks is not implemented yet.\n");
  return (lib_runtime.capability & NN_CAP_MPI) ? TRUE : FALSE;
}
BOOL nn_get_mpi_tasks(UINT *n) {
  *n = lib_runtime.nn_num_tasks ? lib_runtime.nn_num_tasks : 1;
  return (lib_runtime.capability & NN_CAP_MPI) ? TRUE : FALSE;
}
BOOL nn_get_curr_mpi_task(UINT *task) {
  if (!(lib_runtime.capability & NN_CAP_MPI)) return FALSE;
  *task = (UINT)launcher_rank();
  return TRUE;
}
BOOL nn_set_n_gpu(UINT n) {
  if (!g_cuda_ready || n < 1 || n > lib_runtime.cudas.n_gpu) return FALSE;
  g_gpu_cap = (int)n;
  return TRUE;
}
BOOL nn_get_n_gpu(UINT *n) {
  if (!g_cuda_ready) {
    *n = 0;
    return FALSE;
  }
  *n = g_gpu_cap > 0 ? (UINT)g_gpu_cap : lib_runtime.cudas.n_gpu;
  return TRUE;
}
BOOL nn_set_cuda_streams(UINT n) {
  if (!g_cuda_ready) return FALSE;
  // The reference slices every layer over n_gpu x n_streams streams to overlap its small
  // kernels; the persistent kernel already owns the whole GPU, so the value is only recorded.
  lib_runtime.cudas.cuda_n_streams = n < 2 ? 1 : n;
  return TRUE;
}
BOOL nn_get_cuda_streams(UINT *n) {
  if (!g_cuda_ready) return FALSE;
  *n = lib_runtime.cudas.cuda_n_streams;
  return TRUE;
}
BOOL nn_set_omp_blas(UINT) { return FALSE; }
BOOL nn_get_omp_blas(UINT *) { return FALSE; }
cudastreams *nn_return_cudas() { return &lib_runtime.cudas; }

/* ---------------------------- configuration object ----------------------- */
void nn_init_conf(nn_def *c) {
  c->rr = &lib_runtime;
  c->name = NULL;
  c->type = NN_TYPE_UKN;
  c->need_init = FALSE;
  c->seed = 0;
  c->kernel = NULL;
  c->f_kernel = NULL;
  c->train = NN_TRAIN_UKN;
  c->samples = NULL;
  c->tests = NULL;
}
void nn_deinit_conf(nn_def *c) {
  if (!c) return;
  if (c->kernel) nn_free_kernel(c);
  c->kernel = NULL;
  c->rr = NULL;
  FREE(c->name);
  c->type = NN_TYPE_UKN;
  c->need_init = FALSE;
  c->seed = 0;
  FREE(c->f_kernel);
  c->train = NN_TRAIN_UKN;
  FREE(c->samples);
  FREE(c->tests);
}
#define STR_FIELD(fn, field, ctype)                          \
  void nn_set_##fn(nn_def *c, ctype v) {                      \
    FREE(c->field);                                           \
    c->field = dup_or_null(v);                                \
  }                                                           \
  void nn_get_##fn(nn_def *c, CHAR **v) { *v = dup_or_null(c->field); } \
  char *nn_return_##fn(nn_def *c) { return c->field; }
STR_FIELD(name, name, const CHAR *)
STR_FIELD(kernel_filename, f_kernel, CHAR *)
STR_FIELD(samples_directory, samples, CHAR *)
STR_FIELD(tests_directory, tests, CHAR *)
#undef STR_FIELD
#define VAL_FIELD(fn, field, ctype)                            \
  void nn_set_##fn(nn_def *c, ctype v) { c->field = v; }        \
  void nn_get_##fn(nn_def *c, ctype *v) { *v = c->field; }      \
  ctype nn_return_##fn(nn_def *c) { return c->field; }
VAL_FIELD(type, type, nn_type)
VAL_FIELD(need_init, need_init, BOOL)
VAL_FIELD(seed, seed, UINT)
VAL_FIELD(train, train, nn_train)
#undef VAL_FIELD

nn_def *nn_load_conf(const CHAR *filename) {
  hpnn_conf pc;
  if (hpnn_conf_load(filename, &pc) != 0) return NULL;
  nn_def *c = (nn_def *)calloc(1, sizeof(nn_def));
  nn_init_conf(c);
  c->name = dup_or_null(pc.name);
  c->type = (nn_type)pc.type;
  c->need_init = pc.need_init ? TRUE : FALSE;
  c->seed = pc.seed;
  c->f_kernel = dup_or_null(pc.f_kernel);
  c->train = (nn_train)pc.train;
  c->samples = dup_or_null(pc.samples);
  c->tests = dup_or_null(pc.tests);
  BOOL ok;
  if (c->need_init) {
    NN_OUT(stdout, "generating kernel!\n");
    std::vector<UINT> hid(pc.hid, pc.hid + pc.n_hid);
    ok = nn_generate_kernel(c, (UINT)pc.n_in, (UINT)pc.n_hid, (UINT)pc.n_out, hid.data());
    if (!ok) {
      NN_ERROR(stderr, "FAILED to generate NN kernel!\n");
      NN_ERROR(stderr, "[type] unsupported...\n");
    }
  } else {
    NN_OUT(stdout, "loading kernel!\n");
    ok = nn_load_kernel(c);
    if (!ok) NN_ERROR(stderr, "FAILED to load the NN kernel!\n");
  }
  hpnn_conf_free(&pc);
  if (!ok || !c->kernel) {
    if (ok) NN_ERROR(stderr, "Initialization or load of NN kernel FAILED!\n");
    nn_deinit_conf(c);
    free(c);
    return NULL;
  }
  NN_OUT(stdout, "NN definition allocation: %lu (bytes)\n", (unsigned long)sizeof(nn_def));
  return c;
}

void nn_dump_conf(nn_def *c, FILE *fp) {
  if (!c || !fp) return;
  hpnn_conf pc;
  memset(&pc, 0, sizeof(pc));
  pc.name = c->name;
  pc.type = c->type;
  pc.need_init = c->need_init;
  pc.seed = c->seed;
  pc.f_kernel = c->f_kernel;
  pc.train = c->train;
  pc.samples = c->samples;
  pc.tests = c->tests;
  std::vector<int> dims;
  if (c->kernel) {  // sizes are printed for generated kernels too (the reference prints 0, quirk 3)
    Shim *ks = as_shim((kernel_ann *)c->kernel);
    hpnn_kernel *h = ks ? ks->host : nullptr;
    for (int i = 0; h && i <= hpnn_kernel_n_layers(h); ++i) dims.push_back(hpnn_kernel_dim(h, i));
  }
  hpnn_conf_dump(&pc, (int)dims.size(), dims.empty() ? nullptr : dims.data(), fp);
}

/* ------------------------------ kernel management ------------------------ */
static bool ann_like(nn_def *c) { return c && (c->type == NN_TYPE_ANN || c->type == NN_TYPE_SNN); }

void nn_free_kernel(nn_def *c) {
  if (!ann_like(c) || !c->kernel) return;
  ann_kernel_free((kernel_ann *)c->kernel);
  if (Shim *ks = as_shim((kernel_ann *)c->kernel)) {  // still registered: a library-made kernel
    unregister_shim(ks);
    delete ks;
  }
  c->kernel = NULL;
}
BOOL nn_generate_kernel(nn_def *c, ...) {
  if (!ann_like(c)) return FALSE;
  va_list ap;
  va_start(ap, c);
  const UINT n_in = va_arg(ap, UINT), n_hid = va_arg(ap, UINT), n_out = va_arg(ap, UINT);
  UINT *hid = va_arg(ap, UINT *);
  va_end(ap);
  c->kernel = (void *)ann_generate(&c->seed, n_in, n_hid, n_out, hid);
  return c->kernel ? TRUE : FALSE;
}
BOOL nn_load_kernel(nn_def *c) {
  if (!ann_like(c) || !c->f_kernel) return FALSE;
  c->kernel = (void *)ann_load(c->f_kernel);
  return c->kernel ? TRUE : FALSE;
}
void nn_dump_kernel(nn_def *c, FILE *out) {
  if (!ann_like(c) || !c->kernel) return;
  ann_dump((kernel_ann *)c->kernel, out);
}

/* -------------------------------- introspection -------------------------- */
UINT nn_get_n_inputs(nn_def *c) { return (ann_like(c) && c->kernel) ? ((kernel_ann *)c->kernel)->n_inputs : 0; }
UINT nn_get_n_hiddens(nn_def *c) { return (ann_like(c) && c->kernel) ? ((kernel_ann *)c->kernel)->n_hiddens : 0; }
UINT nn_get_n_outputs(nn_def *c) { return (ann_like(c) && c->kernel) ? ((kernel_ann *)c->kernel)->n_outputs : 0; }
UINT nn_get_h_neurons(nn_def *c, UINT layer) {
  if (!ann_like(c) || !c->kernel) return 0;
  kernel_ann *k = (kernel_ann *)c->kernel;
  return layer < k->n_hiddens ? k->hiddens[layer].n_neurons : 0;
}

/* ---------------------------------- samples ------------------------------ */
BOOL nn_read_sample(CHAR *filename, DOUBLE **in, DOUBLE **out) {
  int ni = 0, no = 0;
  return hpnn_sample_read(filename, in, &ni, out, &no) == 0 ? TRUE : FALSE;
}

/* ---------------------------------- drivers ------------------------------ */
BOOL nn_train_kernel(nn_def *c) {
  if (!c || !c->kernel || !c->samples || c->type == NN_TYPE_UKN) return FALSE;
  if (!ann_like(c)) {
    NN_ERROR(stdout, "unimplemented NN type!\n");
    return FALSE;
  }
  kernel_ann *k = (kernel_ann *)c->kernel;
  // host momentum arrays are only needed by the CPU engine (the device kernel owns its own)
  const bool host_momentum = c->train == NN_TRAIN_BPM && !use_device();
  if (host_momentum) ann_momentum_init(k);
  else if (c->train == NN_TRAIN_BPM) {
    unsigned long bytes = 0;
    for (UINT l = 0; l < k->n_hiddens; ++l) bytes += sizeof(double) * (unsigned long)((k->hiddens[l].n_inputs + 7) & ~7u) * k->hiddens[l].n_neurons;
    bytes += sizeof(double) * (unsigned long)((k->output.n_inputs + 7) & ~7u) * k->output.n_neurons;
    NN_OUT(stdout, "[GPU] MOMENTUM ALLOC: %lu (bytes)\n", bytes);
  }
  std::vector<std::string> files = listed_files(c->samples);
  DIR *probe = opendir(c->samples);
  if (!probe) {
    NN_ERROR(stderr, "can't open sample directory: %s\n", c->samples);
    return FALSE;
  }
  closedir(probe);
  if (c->seed == 0) c->seed = (UINT)agreed_seed(0);
  std::vector<int> order(files.size());
  hpnn_shuffle_order(c->seed, (int)files.size(), order.data());
  for (int idx : order) {
    const std::string &name = files[idx];
    NN_OUT(stdout, "TRAINING FILE: %16.16s\t", name.c_str());
    std::string path = std::string(c->samples) + "/" + name;
    DOUBLE *in = NULL, *out = NULL;
    nn_read_sample(&path[0], &in, &out);
    if (!in || !out) {
      FREE(in);
      FREE(out);
      continue;
    }
    DOUBLE res = 0.;
    const bool snn = c->type == NN_TYPE_SNN;
    if (c->train == NN_TRAIN_BPM) res = snn ? snn_train_BPM(k, in, out, .2, -1.) : ann_train_BPM(k, in, out, .2, -1.);
    else if (c->train == NN_TRAIN_BP) res = snn ? snn_train_BP(k, in, out, -1.) : ann_train_BP(k, in, out, -1.);
    if (res > 0.1) NN_DBG(stdout, "bad optimization!\n");
    FREE(in);
    FREE(out);
  }
  if (Shim *ks = as_shim(k)) sync_to_host(ks);
  if (host_momentum) ann_momentum_free(k);
  return TRUE;
}

// verdict lines of nn_run_kernel for one sample (src/libhpnn.c:1443-1514)
static void report_test(nn_def *c, const kernel_ann *k, const DOUBLE *o, const DOUBLE *t) {
  if (c->type == NN_TYPE_SNN) {
    DOUBLE res = 0.;
    UINT guess = 0, want = 0;
    NN_DBG(stdout, " CLASS | PROBABILITY (%%)\n");
    NN_DBG(stdout, "-------|----------------\n");
    for (UINT i = 0; i < k->n_outputs; ++i) {
      NN_DBG(stdout, " %5i | %15.10f\n", i + 1, o[i] * 100.);
      if (o[i] > res) {
        res = o[i];
        guess = i;
      }
      if (t[i] > 0.1) want = i;
    }
    NN_DBG(stdout, "-------|----------------\n");
    NN_COUT(stdout, " BEST CLASS idx=%i P=%15.10f", guess + 1, res * 100);
    if (guess == want) NN_COUT(stdout, " [PASS]\n");
    else NN_COUT(stdout, " [FAIL idx=%i]\n", want + 1);
  } else {
    DOUBLE res = -1.;
    UINT guess = k->n_outputs, want = TRUE;  // `is_ok=TRUE` start value of the reference
    for (UINT i = 0; i < k->n_outputs; ++i) {
      if (res < o[i]) {
        guess = i;
        res = o[i];
      }
      if (t[i] > 0.5) want = i;
    }
    if (guess == want) NN_COUT(stdout, " [PASS]\n");
    else NN_COUT(stdout, " [FAIL idx=%i]\n", want + 1);
  }
  fflush(stdout);
}

void nn_run_kernel(nn_def *c) {
  if (!c || !c->kernel || !c->tests || !ann_like(c)) return;
  kernel_ann *k = (kernel_ann *)c->kernel;
  DIR *probe = opendir(c->tests);
  if (!probe) {
    NN_ERROR(stderr, "can't open test directory: %s\n", c->tests);
    return;
  }
  closedir(probe);
  std::vector<std::string> files = listed_files(c->tests);
  if (c->seed == 0) c->seed = (UINT)agreed_seed(0);
  std::vector<int> order(files.size());
  hpnn_shuffle_order(c->seed, (int)files.size(), order.data());
  const int type = c->type == NN_TYPE_SNN ? HPNN_TYPE_SNN : HPNN_TYPE_ANN;
  Shim *s = as_shim(k);
  if (!s || !s->host) {
    NN_ERROR(stderr, "nn_run_kernel: the configuration's kernel was not created by this library\n");
    return;
  }
  // Fast path (same verdicts, far fewer host round trips): parse every file in parallel, push the
  // whole directory through the batched FP64 forward.  HPNN_RUN_BATCH=0 restores one launch per file.
  const char *rb = getenv("HPNN_RUN_BATCH");
  const int batch = rb ? atoi(rb) : 1024;
  hpnn_net *dev = (batch > 0 && !files.empty()) ? device_for(s, type, s->dev_momentum) : nullptr;
  if (dev && hpnn_net_world(dev) == 1) {
    const int n_in = (int)k->n_inputs, n_out = (int)k->n_outputs, nf = (int)files.size();
    std::vector<std::string> paths(nf);
    std::vector<const char *> cpaths(nf);
    for (int j = 0; j < nf; ++j) {
      paths[j] = std::string(c->tests) + "/" + files[order[j]];
      cpaths[j] = paths[j].c_str();
    }
    std::vector<double> X((size_t)nf * n_in), T((size_t)nf * n_out), Y((size_t)nf * n_out);
    std::vector<unsigned char> ok(nf);
    hpnn_sample_read_many(cpaths.data(), nf, n_in, n_out, X.data(), T.data(), ok.data());
    bool fine = true;
    for (int b0 = 0; b0 < nf && fine; b0 += batch) {
      const int nb = nf - b0 < batch ? nf - b0 : batch;
      fine = hpnn_net_forward_batch(dev, X.data() + (size_t)b0 * n_in, nb, Y.data() + (size_t)b0 * n_out, 0) == 0;
    }
    if (fine) {
      for (int j = 0; j < nf; ++j) {
        NN_OUT(stdout, "TESTING FILE: %16.16s\t", files[order[j]].c_str());
        if (!ok[j]) continue;  // unreadable sample: skipped silently, as in the reference
        report_test(c, k, Y.data() + (size_t)j * n_out, T.data() + (size_t)j * n_out);
      }
      return;
    }
    NN_WARN(stderr, "batched run failed (%s): falling back to one sample per launch\n", hpnn_last_error());
  }
  for (int idx : order) {
    const std::string &name = files[idx];
    NN_OUT(stdout, "TESTING FILE: %16.16s\t", name.c_str());
    std::string path = std::string(c->tests) + "/" + name;
    DOUBLE *in = NULL, *t = NULL;
    nn_read_sample(&path[0], &in, &t);
    if (!in || !t) {
      FREE(in);
      FREE(t);
      continue;
    }
    memcpy(k->in, in, sizeof(double) * k->n_inputs);
    if (c->type == NN_TYPE_SNN) snn_kernel_run(k);
    else ann_kernel_run(k);
    report_test(c, k, k->output.vec, t);
    FREE(in);
    FREE(t);
  }
}

/* ============================ model layer: ann.h ========================== */
// Releases the storage behind a kernel.  A kernel of ann_generate / ann_load stays registered
// (its struct is zeroed, nn_free_kernel deletes the object); a caller-owned struct that went
// through ann_kernel_allocate is forgotten altogether, the struct itself remains the caller's.
BOOL ann_kernel_free(kernel_ann *kernel) {
  Shim *s = as_shim(kernel);
  if (!s) return FALSE;  // not ours (or already released): nothing behind it may be touched
  drop_device(s);
  if (s->host) hpnn_kernel_destroy(s->host);
  s->host = nullptr;
  memset(s->pub, 0, sizeof(*s->pub));
  if (s->pub != &s->own) {
    unregister_shim(s);
    delete s;
  }
  return TRUE;
}

// Reference contract (src/ann.c:113-202): the caller ALLOCs a kernel_ann and this call fills
// it.  The weights and work vectors it points to afterwards are owned by the library and live
// until ann_kernel_free(kernel).  Called on a kernel of ann_generate / ann_load it re-shapes that
// kernel (weights zeroed).
BOOL ann_kernel_allocate(kernel_ann *kernel, UINT n_inputs, UINT n_hiddens, UINT *h_neurons, UINT n_outputs) {
  if (!kernel || !h_neurons || n_hiddens < 1) return FALSE;
  std::vector<int> dims;
  dims.push_back((int)n_inputs);
  for (UINT i = 0; i < n_hiddens; ++i) dims.push_back((int)h_neurons[i]);
  dims.push_back((int)n_outputs);
  hpnn_kernel *h = hpnn_kernel_create((int)dims.size(), dims.data());
  if (!h) return FALSE;
  Shim *s = as_shim(kernel);
  if (!s) {  // a struct of the caller's: track it by address, never read behind it
    s = new Shim();
    memset(&s->own, 0, sizeof(s->own));
    s->pub = kernel;
    register_shim(s);
  } else {
    drop_device(s);
    if (s->host) hpnn_kernel_destroy(s->host);
  }
  s->host = h;
  refresh_views(s);
  return TRUE;
}

kernel_ann *ann_load(CHAR *f_kernel) {
  Shim *s = wrap(hpnn_kernel_load(f_kernel));
  return s ? s->pub : NULL;
}

kernel_ann *ann_generate(UINT *seed, UINT n_inputs, UINT n_hiddens, UINT n_outputs, UINT *hiddens) {
  if (!hiddens || n_hiddens < 1) return NULL;
  std::vector<int> dims;
  dims.push_back((int)n_inputs);
  for (UINT i = 0; i < n_hiddens; ++i) dims.push_back((int)hiddens[i]);
  dims.push_back((int)n_outputs);
  unsigned s0 = agreed_seed(seed ? *seed : 0);  // several processes: rank 0's choice when the seed is 0
  Shim *s = wrap(hpnn_kernel_generate(&s0, (int)dims.size(), dims.data()));
  if (seed) *seed = s0;
  return s ? s->pub : NULL;
}

void ann_dump(kernel_ann *kernel, FILE *out) {
  if (!kernel) {
    NN_ERROR(stderr, "CAN'T SAVE KERNEL! kernel=NULL\n");
    return;
  }
  Shim *s = as_shim(kernel);
  if (!s || !s->host) {
    NN_ERROR(stderr, "CAN'T SAVE KERNEL! not a kernel of this library\n");
    return;
  }
  sync_to_host(s);
  if (hpnn_compat_is_root()) hpnn_kernel_dump(s->host, out);
}

BOOL ann_validate_kernel(kernel_ann *k) {
  if (!k || k->n_inputs < 1 || k->n_outputs < 1 || k->n_hiddens < 1 || !k->in) return FALSE;
  for (UINT i = 0; i < k->n_hiddens; ++i)
    if (k->hiddens[i].n_neurons < 1 || k->hiddens[i].n_inputs < 1 || !k->hiddens[i].weights || !k->hiddens[i].vec)
      return FALSE;
  return (k->output.n_neurons >= 1 && k->output.n_inputs >= 1 && k->output.weights && k->output.vec) ? TRUE : FALSE;
}

DOUBLE ann_act(DOUBLE x) { return 2.0 / (1.0 + exp(-1.0 * x)) - 1.0; }
DOUBLE ann_dact(DOUBLE y) { return -0.5 * (y * y - 1.0); }

void ann_kernel_run(kernel_ann *kernel) {
  if (kernel) run_once(kernel, HPNN_TYPE_ANN);
}
DOUBLE ann_kernel_train(kernel_ann *kernel, const DOUBLE *train) {
  return kernel ? single_iteration(kernel, HPNN_TYPE_ANN, HPNN_TRAIN_BP, train, 0.) : 0.;
}
void ann_momentum_init(kernel_ann *kernel) {
  Shim *s = as_shim(kernel);
  if (!s || !s->host) return;
  hpnn_cpu_momentum_init(s->host);
  refresh_views(s);
  unsigned long bytes = 0;
  for (int l = 0; l < hpnn_kernel_n_layers(s->host); ++l)
    bytes += sizeof(double) * (unsigned long)hpnn_kernel_dim(s->host, l) * hpnn_kernel_dim(s->host, l + 1);
  NN_OUT(stdout, "[CPU] MOMENTUM ALLOC: %lu (bytes)\n", bytes);  // src/ann.c:1904-1907
}
void ann_raz_momentum(kernel_ann *kernel) {
  Shim *s = as_shim(kernel);
  if (s && s->host) hpnn_cpu_momentum_zero(s->host);
}
void ann_momentum_free(kernel_ann *kernel) {
  Shim *s = as_shim(kernel);
  if (!s || !s->host) return;
  hpnn_cpu_momentum_free(s->host);
  refresh_views(s);
}
DOUBLE ann_kernel_train_momentum(kernel_ann *kernel, const DOUBLE *train, DOUBLE alpha) {
  return kernel ? single_iteration(kernel, HPNN_TYPE_ANN, HPNN_TRAIN_BPM, train, alpha) : 0.;
}
DOUBLE ann_train_BP(kernel_ann *kernel, DOUBLE *in, DOUBLE *out, DOUBLE delta) {
  return kernel ? train_sample(kernel, HPNN_TYPE_ANN, HPNN_TRAIN_BP, in, out, 0., delta) : 0.;
}
DOUBLE ann_train_BPM(kernel_ann *kernel, DOUBLE *in, DOUBLE *out, DOUBLE alpha, DOUBLE delta) {
  return kernel ? train_sample(kernel, HPNN_TYPE_ANN, HPNN_TRAIN_BPM, in, out, alpha, delta) : 0.;
}

/* ============================ model layer: snn.h ========================== */
void snn_kernel_run(kernel_ann *kernel) {
  if (kernel) run_once(kernel, HPNN_TYPE_SNN);
}
DOUBLE snn_kernel_train(kernel_ann *kernel, const DOUBLE *train) {
  return kernel ? single_iteration(kernel, HPNN_TYPE_SNN, HPNN_TRAIN_BP, train, 0.) : 0.;
}
DOUBLE snn_kernel_train_momentum(kernel_ann *kernel, const DOUBLE *train, DOUBLE alpha) {
  return kernel ? single_iteration(kernel, HPNN_TYPE_SNN, HPNN_TRAIN_BPM, train, alpha) : 0.;
}
DOUBLE snn_train_BP(kernel_ann *kernel, DOUBLE *in, DOUBLE *out, DOUBLE delta) {
  return kernel ? train_sample(kernel, HPNN_TYPE_SNN, HPNN_TRAIN_BP, in, out, 0., delta) : 0.;
}
DOUBLE snn_train_BPM(kernel_ann *kernel, DOUBLE *in, DOUBLE *out, DOUBLE alpha, DOUBLE delta) {
  return kernel ? train_sample(kernel, HPNN_TYPE_SNN, HPNN_TRAIN_BPM, in, out, alpha, delta) : 0.;
}

/* ==================== device layer: cuda_ann.h / cuda_snn.h ==================== */
// The reference's scuda_* functions (src/cuda_ann.cu:192-4103, src/cuda_snn.cu) operate on device
// pointers kept inside kernel_ann.  Here the device copy lives behind the kernel (Shim::dev), so these
// entry points are the same operations phrased on the host mirror; `cudas` is accepted and ignored.
static layer_ann *layer_at(kernel_ann *k, int index) {
  if (!k || index < 0) return nullptr;
  return (UINT)index < k->n_hiddens ? &k->hiddens[index] : &k->output;
}
int64_t scuda_ann_allocate(kernel_ann *kernel, cudastreams *) {
  Shim *s = as_shim(kernel);
  if (!s || !s->host) return 0;
  hpnn_net *dev = device_for(s, HPNN_TYPE_ANN, s->dev_momentum);
  return dev ? (int64_t)hpnn_net_device_bytes(dev) : 0;
}
void scuda_ann_deallocate(kernel_ann *kernel, cudastreams *) {
  Shim *s = as_shim(kernel);
  if (!s) return;
  sync_to_host(s);
  drop_device(s);
}
int64_t scuda_ann_allocate_momentum(kernel_ann *kernel, cudastreams *) {
  Shim *s = as_shim(kernel);
  if (!s || !s->host) return 0;
  ann_momentum_init(kernel);
  int64_t bytes = 0;
  for (int l = 0; l < hpnn_kernel_n_layers(s->host); ++l)
    bytes += (int64_t)sizeof(double) * hpnn_kernel_dim(s->host, l) * hpnn_kernel_dim(s->host, l + 1);
  return bytes;
}
BOOL scuda_ann_free_momentum(kernel_ann *kernel, cudastreams *) {
  if (!as_shim(kernel)) return FALSE;
  ann_momentum_free(kernel);
  return TRUE;
}
void scuda_ann_weight_transfer_C2G(kernel_ann *kernel, int index, DOUBLE *weight, cudastreams *) {
  Shim *s = as_shim(kernel);
  layer_ann *L = layer_at(kernel, index);
  if (!s || !s->host || !L || !weight) return;
  sync_to_host(s);
  memcpy(L->weights, weight, sizeof(double) * (size_t)L->n_neurons * L->n_inputs);
  if (s->dev) s->dev_stale = true;  // re-uploaded before the next device step
}
void scuda_ann_weight_transfer_G2C(kernel_ann *kernel, int index, DOUBLE **weight, cudastreams *) {
  Shim *s = as_shim(kernel);
  layer_ann *L = layer_at(kernel, index);
  if (!s || !s->host || !L || !weight || !*weight) return;
  sync_to_host(s);
  memcpy(*weight, L->weights, sizeof(double) * (size_t)L->n_neurons * L->n_inputs);
}
void scuda_ann_forward(kernel_ann *kernel, cudastreams *) { ann_kernel_run(kernel); }
void scuda_snn_forward(kernel_ann *kernel, cudastreams *) { snn_kernel_run(kernel); }
static double device_layer_error(kernel_ann *kernel, int type, const double *train) {
  Shim *s = as_shim(kernel);
  if (!s || !s->host || !train) return 0.;
  return hpnn_cpu_error(s->host, type, train);  // from the output vector the last forward left in the mirror
}
double scuda_ann_error(kernel_ann *kernel, double *train, cudastreams *) { return device_layer_error(kernel, HPNN_TYPE_ANN, train); }
double scuda_snn_error(kernel_ann *kernel, double *train, cudastreams *) { return device_layer_error(kernel, HPNN_TYPE_SNN, train); }
double scuda_ann_train(kernel_ann *kernel, double *train, cudastreams *) { return ann_kernel_train(kernel, train); }
double scuda_snn_train(kernel_ann *kernel, double *train, cudastreams *) { return snn_kernel_train(kernel, train); }
void scuda_ann_raz_momentum(kernel_ann *kernel, cudastreams *) { ann_raz_momentum(kernel); }
double scuda_ann_train_momentum(kernel_ann *kernel, double *train, double moment, cudastreams *) {
  return ann_kernel_train_momentum(kernel, train, moment);
}
double scuda_snn_train_momentum(kernel_ann *kernel, double *train, double moment, cudastreams *) {
  return snn_kernel_train_momentum(kernel, train, moment);
}

}  // extern "C"
