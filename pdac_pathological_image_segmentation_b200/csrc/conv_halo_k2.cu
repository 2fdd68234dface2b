// conv_halo_kernel instantiations, group 2 (see conv_halo_kernel.cuh)
#include "conv_halo_kernel.cuh"

namespace pdac {

HaloKernel halo_kernel_k2(int mode, int ki, int kj) {
  static const HaloKernel t3[4][5] = HALO_MODE(3);
  static const HaloKernel t4[4][5] = HALO_MODE(4);
  if (mode == 3) return t3[ki][kj];
  if (mode == 4) return t4[ki][kj];
  return nullptr;
}

}  // namespace pdac
