/*
 * libhpnn/cuda_ann.h -- device-layer entry points of the compatibility API, ANN flavour
 * (reference: include/libhpnn/cuda_ann.h:40-52; its __global__ kernels are not part of the API).
 *
 * In the reference these functions work on DEVICE pointers stored inside kernel_ann.  Here a
 * kernel_ann is a host mirror and the device copy (an hpnn_net, row-split over the GPUs when the
 * net is large enough) is managed by the library, so: `train` / `weight` are HOST arrays, the
 * cudastreams argument is accepted and ignored, the *_allocate calls create the device copy and
 * report its size, and the single-step calls keep the reference's return values (error, dEp).
 */
#ifndef CUDA_ANN_H
#define CUDA_ANN_H
#include <stdint.h>
#include <libhpnn/ann.h>

#ifdef __cplusplus
extern "C" {
#endif
void scuda_ann_deallocate(kernel_ann *kernel, cudastreams *cudas);
int64_t scuda_ann_allocate(kernel_ann *kernel, cudastreams *cudas);
BOOL scuda_ann_free_momentum(kernel_ann *kernel, cudastreams *cudas);
int64_t scuda_ann_allocate_momentum(kernel_ann *kernel, cudastreams *cudas);
/* index < n_hiddens: hidden layer `index`, otherwise the output layer; row-major [neurons x inputs] */
void scuda_ann_weight_transfer_C2G(kernel_ann *kernel, int index, DOUBLE *weight, cudastreams *cudas);
void scuda_ann_weight_transfer_G2C(kernel_ann *kernel, int index, DOUBLE **weight, cudastreams *cudas);
void scuda_ann_forward(kernel_ann *kernel, cudastreams *cudas);
double scuda_ann_error(kernel_ann *kernel, double *train, cudastreams *cudas);
double scuda_ann_train(kernel_ann *kernel, double *train, cudastreams *cudas);
void scuda_ann_raz_momentum(kernel_ann *kernel, cudastreams *cudas);
double scuda_ann_train_momentum(kernel_ann *kernel, double *train, double moment, cudastreams *cudas);
#ifdef __cplusplus
}
#endif
#endif /* CUDA_ANN_H */
