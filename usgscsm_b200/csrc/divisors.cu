// divisors.cu -- host helper: evaluate the reciprocals of a model's constant divisors on the
// device (device_math.cuh refinedReciprocal: the value depends on the GPU's MUFU.RCP64H
// table, so it cannot be computed on the host).
#include <vector>

#include "model.h"

namespace ub200 {

void fillDivisors(const std::vector<Divisor *> &divisors) {
  const int n = (int)divisors.size();
  if (n == 0) return;
  std::vector<double> b(n), r(n);
  for (int i = 0; i < n; i++) b[i] = divisors[i]->b;
  double *d = nullptr;
  UB_CUDA(cudaMalloc(&d, 2 * n * sizeof(double)));
  cudaError_t e = cudaMemcpy(d, b.data(), n * sizeof(double), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = launchReciprocals(n, d, d + n, nullptr);
  if (e == cudaSuccess) e = cudaMemcpy(r.data(), d + n, n * sizeof(double), cudaMemcpyDeviceToHost);
  cudaFree(d);
  UB_CUDA(e);
  for (int i = 0; i < n; i++) divisors[i]->r = r[i];
}

void appendLagrangeDivisors(LagrangeDivisors &L, std::vector<Divisor *> &out) {
  L.d5040.b = 5040.0;
  L.d720.b = 720.0;
  L.d240.b = 240.0;
  L.d144.b = 144.0;
  L.d120.b = 120.0;
  L.d24.b = 24.0;
  L.d12.b = 12.0;
  L.d6.b = 6.0;
  for (Divisor *d : {&L.d5040, &L.d720, &L.d240, &L.d144, &L.d120, &L.d24, &L.d12, &L.d6})
    out.push_back(d);
}

// A UniformTable keeps its step as a plain double plus the reciprocal; route it through a
// temporary Divisor.
void fillTableSteps(const std::vector<UniformTable *> &tables) {
  std::vector<Divisor> tmp(tables.size());
  std::vector<Divisor *> ptrs;
  for (size_t i = 0; i < tables.size(); i++) {
    tmp[i].b = tables[i]->dt;
    ptrs.push_back(&tmp[i]);
  }
  fillDivisors(ptrs);
  for (size_t i = 0; i < tables.size(); i++) tables[i]->rdt = tmp[i].r;
}

}  // namespace ub200
