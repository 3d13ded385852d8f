// Host-side network object shared by the I/O code, the CPU engine and the shim.
#pragma once
#include <string>
#include <vector>

struct hpnn_kernel {
  bool has_name = false;
  std::string name;
  std::vector<int> dims;                 // n_in, hidden..., n_out
  std::vector<std::vector<double>> W;    // W[l] row-major [dims[l+1] x dims[l]]
  // CPU-engine work space (lazily sized)
  std::vector<std::vector<double>> act;  // act[0]=input, act[l+1]=output of layer l
  std::vector<std::vector<double>> delta;
  std::vector<std::vector<double>> hold;  // activations of the previous sweep (CPU engine, fused update+forward)
  std::vector<std::vector<double>> dW;   // momentum
  // CPU engine, per-sample training loop: weights / momenta with 8 rows interleaved per column
  std::vector<std::vector<double>> packW, packM;
  int n_layers() const { return (int)dims.size() - 1; }
};
