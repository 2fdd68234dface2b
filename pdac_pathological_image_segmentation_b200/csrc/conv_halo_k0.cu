// conv_halo_kernel instantiations, group 0 (see conv_halo_kernel.cuh)
#include "conv_halo_kernel.cuh"

namespace pdac {

HaloKernel halo_kernel_k0(int mode, int ki, int kj) {
  static const HaloKernel t0[4][5] = HALO_MODE(0);
  static const HaloKernel t1[4][5] = HALO_MODE(1);
  if (mode == 0) return t0[ki][kj];
  if (mode == 1) return t1[ki][kj];
  return nullptr;
}

}  // namespace pdac
