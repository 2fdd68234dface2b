// conv_halo_kernel instantiations, group 3 (see conv_halo_kernel.cuh)
#include "conv_halo_kernel.cuh"

namespace pdac {

HaloKernel halo_kernel_k3(int mode, int ki, int kj) {
  static const HaloKernel t5[4][5] = HALO_MODE_D(5);
  if (mode == 5) return t5[ki][kj];
  return nullptr;
}

}  // namespace pdac
