/*
 * libhpnn/cuda_snn.h -- device-layer entry points of the compatibility API, SNN flavour
 * (reference: include/libhpnn/cuda_snn.h:33-39).  Same conventions as libhpnn/cuda_ann.h.
 */
#ifndef CUDA_SNN_H
#define CUDA_SNN_H
#include <libhpnn/cuda_ann.h>

#ifdef __cplusplus
extern "C" {
#endif
void scuda_snn_forward(kernel_ann *kernel, cudastreams *cudas);
double scuda_snn_error(kernel_ann *kernel, double *train, cudastreams *cudas);
double scuda_snn_train(kernel_ann *kernel, double *train, cudastreams *cudas);
double scuda_snn_train_momentum(kernel_ann *kernel, double *train, double moment, cudastreams *cudas);
#ifdef __cplusplus
}
#endif
#endif /* CUDA_SNN_H */
