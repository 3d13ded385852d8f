/*
 * libhpnn/snn.h -- SNN (softmax output) model layer of the compatibility API
 * (reference: include/libhpnn/snn.h:25-36).  An SNN shares kernel_ann with the ANN.
 */
#ifndef SNN_H
#define SNN_H
#include <libhpnn/ann.h>

#ifdef __cplusplus
extern "C" {
#endif
void snn_kernel_run(kernel_ann *kernel);
DOUBLE snn_kernel_train(kernel_ann *kernel, const DOUBLE *train);
DOUBLE snn_kernel_train_momentum(kernel_ann *kernel, const DOUBLE *train, DOUBLE alpha);
DOUBLE snn_train_BP(kernel_ann *kernel, DOUBLE *train_in, DOUBLE *train_out, DOUBLE delta);
DOUBLE snn_train_BPM(kernel_ann *kernel, DOUBLE *train_in, DOUBLE *train_out, DOUBLE alpha, DOUBLE delta);
#ifdef __cplusplus
}
#endif
#endif /* SNN_H */
