"""pdac_pathological_image_segmentation_b200 — B200-native (sm_100a) hot path of
moksu27/PDAC_pathological_image_segmentation: drop-in ResNet18-encoder segmentation modules, Dice loss / masks,
DDP gradient reduction and tile stitching on hand-written CUDA kernels behind a C ABI (include/pdac_b200.h)."""
__version__ = "0.1.0"
