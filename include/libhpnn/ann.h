/*
 * libhpnn/ann.h -- ANN model layer of the compatibility API.
 *
 * Same structs and entry points as the reference (include/libhpnn/ann.h:35-78).  The arrays
 * reachable from a kernel_ann are HOST mirrors: the live copy of the weights is on the GPU(s)
 * while a network trains; the mirrors are refreshed whenever a call that may have changed
 * them returns (ann_kernel_train*, ann_train_BP/BPM, nn_train_kernel) and before ann_dump.
 */
#ifndef ANN_H
#define ANN_H
#include <libhpnn/common.h>

#define _2D_IDX(len, j, i) (len) * (j) + i

#ifndef ANN_UNROLL
#define ANN_UNROLL 4
#endif

typedef struct {
  UINT n_neurons;  /* rows of the weight matrix      */
  UINT n_inputs;   /* columns                        */
  DOUBLE *weights; /* row-major [n_neurons x n_inputs] */
  DOUBLE *vec;     /* layer output                   */
} layer_ann;

typedef struct kann {
  CHAR *name;
  UINT n_inputs;
  DOUBLE *in;
  UINT n_hiddens;
  layer_ann *hiddens;
  UINT n_outputs;
  layer_ann output;
  DOUBLE **dw;         /* momentum, one array per layer (NULL until ann_momentum_init) */
  UINT max_index;      /* widest vector */
  DOUBLE *tmp_cpu;
  DOUBLE *tmp_gpu;     /* unused: scratch lives inside the device engine */
  struct kann **kerns; /* unused: there are no per-GPU replicas */
} kernel_ann;

#ifdef __cplusplus
extern "C" {
#endif
BOOL ann_kernel_free(kernel_ann *kernel);
BOOL ann_kernel_allocate(kernel_ann *kernel, UINT n_inputs, UINT n_hiddens, UINT *h_neurons, UINT n_outputs);
kernel_ann *ann_load(CHAR *f_kernel);
kernel_ann *ann_generate(UINT *seed, UINT n_inputs, UINT n_hiddens, UINT n_outputs, UINT *hiddens);
void ann_dump(kernel_ann *kernel, FILE *out);
BOOL ann_validate_kernel(kernel_ann *kernel);
DOUBLE ann_act(DOUBLE x);
DOUBLE ann_dact(DOUBLE y);
void ann_kernel_run(kernel_ann *kernel);
DOUBLE ann_kernel_train(kernel_ann *kernel, const DOUBLE *train);
void ann_momentum_init(kernel_ann *kernel);
void ann_raz_momentum(kernel_ann *kernel);
void ann_momentum_free(kernel_ann *kernel);
DOUBLE ann_kernel_train_momentum(kernel_ann *kernel, const DOUBLE *train, DOUBLE alpha);
DOUBLE ann_train_BP(kernel_ann *kernel, DOUBLE *train_in, DOUBLE *train_out, DOUBLE delta);
DOUBLE ann_train_BPM(kernel_ann *kernel, DOUBLE *train_in, DOUBLE *train_out, DOUBLE alpha, DOUBLE delta);
#ifdef __cplusplus
}
#endif
#endif /* ANN_H */
