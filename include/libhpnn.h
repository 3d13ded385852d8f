/*
 * libhpnn.h -- public C API of hpnn-b200, source compatible with libhpnn v0.2.
 *
 * Every nn_* entry point of the reference header (include/libhpnn.h:126-215) exists with
 * the same signature, so host programs -- including the reference's own tests/train_nn.c and
 * tests/run_nn.c -- compile unchanged and link against libhpnn.so of this project.
 * What runs underneath is new: file formats and the CPU engine in csrc/core + csrc/cpu, the
 * persistent sm_100a trainer in csrc/kernels (see include/hpnn_core.h).
 */
#ifndef LIBHPNN_H
#define LIBHPNN_H
#include <libhpnn/common.h>

/* ---- capabilities (bit values are part of the ABI) ---- */
typedef enum {
  NN_CAP_NONE = 0,
  NN_CAP_OMP = (1 << 0),
  NN_CAP_MPI = (1 << 1), /* here: multi-rank launch detected (torchrun / mpirun environment) */
  NN_CAP_CUDA = (1 << 2),
  NN_CAP_CUBLAS = (1 << 3), /* never set: no library GEMV on the hot path */
  /* (1<<4) reserved by the reference for OpenCL */
  NN_CAP_PBLAS = (1 << 5), /* never set */
  NN_CAP_SBLAS = (1 << 6), /* never set */
} nn_cap;

typedef struct {
  nn_cap capability;
  SHORT nn_verbose;
  BOOL nn_dry;
  UINT nn_num_threads;
  UINT nn_num_blas;
  UINT nn_num_tasks;
  cudastreams cudas;
} nn_runtime;

typedef enum {
  NN_TYPE_ANN = 0,  /* feed-forward, sigmoid-type activation everywhere */
  NN_TYPE_LNN = 1,  /* declared by the reference, implemented by neither */
  NN_TYPE_SNN = 2,  /* ANN with a softmax output layer */
  NN_TYPE_UKN = -1,
} nn_type;

typedef enum {
  NN_TRAIN_BP = 0,
  NN_TRAIN_BPM = 1,
  NN_TRAIN_CG = 2,   /* enum value only, as in the reference */
  NN_TRAIN_SPLX = 3, /* enum value only, as in the reference */
  NN_TRAIN_UKN = -1,
} nn_train;

#define BP_LEARN_RATE 0.001
#define MIN_BP_ITER 31
#define MAX_BP_ITER 102399
#define DELTA_BP 1E-6
#define BPM_LEARN_RATE 0.0005
#define MIN_BPM_ITER 15
#define MAX_BPM_ITER 102399
#define DELTA_BPM 1E-6

typedef struct {
  nn_runtime *rr;
  CHAR *name;
  nn_type type;
  BOOL need_init;
  UINT seed;
  void *kernel; /* kernel_ann* (include/libhpnn/ann.h) */
  CHAR *f_kernel;
  nn_train train;
  CHAR *samples;
  CHAR *tests;
} nn_def;

#define _NN(a, b) nn_##a##_##b

/* verbosity-gated output: WARN > 0, OUT/COUT > 1, DBG > 2 */
#define NN_DBG(_file, ...)                  \
  do {                                      \
    if ((_NN(return, verbose)()) > 2) {     \
      _OUT((_file), "NN(DBG): ");           \
      _OUT((_file), __VA_ARGS__);           \
    }                                       \
  } while (0)
#define NN_OUT(_file, ...)                  \
  do {                                      \
    if ((_NN(return, verbose)()) > 1) {     \
      _OUT((_file), "NN: ");                \
      _OUT((_file), __VA_ARGS__);           \
    }                                       \
  } while (0)
#define NN_COUT(_file, ...)                 \
  do {                                      \
    if ((_NN(return, verbose)()) > 1) {     \
      _OUT((_file), __VA_ARGS__);           \
    }                                       \
  } while (0)
#define NN_WARN(_file, ...)                 \
  do {                                      \
    if ((_NN(return, verbose)()) > 0) {     \
      _OUT((_file), "NN(WARN): ");          \
      _OUT((_file), __VA_ARGS__);           \
    }                                       \
  } while (0)
#define NN_ERROR(_file, ...)                \
  do {                                      \
    _OUT((_file), "NN(ERR): ");             \
    _OUT((_file), __VA_ARGS__);             \
  } while (0)
#define NN_WRITE _OUT

#ifdef __cplusplus
extern "C" {
#endif
/* ---- runtime ---- */
void _NN(inc, verbose)();
void _NN(dec, verbose)();
void _NN(set, verbose)(SHORT verbosity);
void _NN(get, verbose)(SHORT *verbosity);
SHORT _NN(return, verbose)();
void _NN(toggle, dry)();
void _NN(get, capabilities)(nn_cap *capabilities);
void _NN(unset, capability)(nn_cap capability);
nn_cap _NN(return, capabilities)();
BOOL _NN(init, OMP)();
BOOL _NN(init, MPI)();
BOOL _NN(init, CUDA)();
BOOL _NN(init, BLAS)();
int _NN(init, all)(UINT init_verbose);
BOOL _NN(deinit, OMP)();
BOOL _NN(deinit, MPI)();
BOOL _NN(deinit, CUDA)();
BOOL _NN(deinit, BLAS)();
int _NN(deinit, all)();
/* ---- knobs ---- */
BOOL _NN(set, omp_threads)(UINT n_threads);
BOOL _NN(get, omp_threads)(UINT *n_threads);
int _NN(return, omp_threads)();
BOOL _NN(set, mpi_tasks)(UINT n_tasks);
BOOL _NN(get, mpi_tasks)(UINT *n_tasks);
BOOL _NN(get, curr_mpi_task)(UINT *task);
BOOL _NN(set, n_gpu)(UINT n_gpu); /* implemented here: caps the GPUs used by later kernels */
BOOL _NN(get, n_gpu)(UINT *n_gpu);
BOOL _NN(set, cuda_streams)(UINT n_streams);
BOOL _NN(get, cuda_streams)(UINT *n_streams);
BOOL _NN(set, omp_blas)(UINT n_blas);
BOOL _NN(get, omp_blas)(UINT *n_blas);
cudastreams *_NN(return, cudas)();
/* ---- configuration object ---- */
void _NN(init, conf)(nn_def *conf);
void _NN(deinit, conf)(nn_def *conf);
void _NN(set, name)(nn_def *conf, const CHAR *name);
void _NN(get, name)(nn_def *conf, CHAR **name);
char *_NN(return, name)(nn_def *conf);
void _NN(set, type)(nn_def *conf, nn_type type);
void _NN(get, type)(nn_def *conf, nn_type *type);
nn_type _NN(return, type)(nn_def *conf);
void _NN(set, need_init)(nn_def *conf, BOOL need_init);
void _NN(get, need_init)(nn_def *conf, BOOL *need_init);
BOOL _NN(return, need_init)(nn_def *conf);
void _NN(set, seed)(nn_def *conf, UINT seed);
void _NN(get, seed)(nn_def *conf, UINT *seed);
UINT _NN(return, seed)(nn_def *conf);
void _NN(set, kernel_filename)(nn_def *conf, CHAR *f_kernel);
void _NN(get, kernel_filename)(nn_def *conf, CHAR **f_kernel);
char *_NN(return, kernel_filename)(nn_def *conf);
void _NN(set, train)(nn_def *conf, nn_train train);
void _NN(get, train)(nn_def *conf, nn_train *train);
nn_train _NN(return, train)(nn_def *conf);
void _NN(set, samples_directory)(nn_def *conf, CHAR *samples);
void _NN(get, samples_directory)(nn_def *conf, CHAR **samples);
char *_NN(return, samples_directory)(nn_def *conf);
void _NN(set, tests_directory)(nn_def *conf, CHAR *tests);
void _NN(get, tests_directory)(nn_def *conf, CHAR **tests);
char *_NN(return, tests_directory)(nn_def *conf);
nn_def *_NN(load, conf)(const CHAR *filename);
void _NN(dump, conf)(nn_def *conf, FILE *fp);
/* ---- kernel management ---- */
void _NN(free, kernel)(nn_def *conf);
BOOL _NN(generate, kernel)(nn_def *conf, ...); /* UINT n_in, UINT n_hid, UINT n_out, UINT *hiddens */
BOOL _NN(load, kernel)(nn_def *conf);
void _NN(dump, kernel)(nn_def *conf, FILE *output);
/* ---- introspection ---- */
UINT _NN(get, n_inputs)(nn_def *conf);
UINT _NN(get, n_hiddens)(nn_def *conf);
UINT _NN(get, n_outputs)(nn_def *conf);
UINT _NN(get, h_neurons)(nn_def *conf, UINT layer);
/* ---- sample I/O ---- */
BOOL _NN(read, sample)(CHAR *filename, DOUBLE **in, DOUBLE **out);
/* ---- execution ---- */
BOOL _NN(train, kernel)(nn_def *conf);
void _NN(run, kernel)(nn_def *conf);
#ifdef __cplusplus
}
#endif
#endif /* LIBHPNN_H */
