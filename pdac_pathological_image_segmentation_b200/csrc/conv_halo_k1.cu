// conv_halo_kernel instantiations, group 1 (see conv_halo_kernel.cuh)
#include "conv_halo_kernel.cuh"

namespace pdac {

HaloKernel halo_kernel_k1(int mode, int ki, int kj) {
  static const HaloKernel t2[4][5] = HALO_MODE(2);
  static const HaloKernel t7[4][5] = HALO_MODE(7);
  if (mode == 2) return t2[ki][kj];
  if (mode == 7) return t7[ki][kj];
  return nullptr;
}

}  // namespace pdac
