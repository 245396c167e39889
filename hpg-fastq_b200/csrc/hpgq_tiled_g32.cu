// hpgq_tiled_g32.cu -- instantiations of the tiled QC kernel for read groups of 32 lanes (see hpgq_tiled.cuh)
#include "hpgq_tiled_body.inc"
namespace hqt {
HQT_INSTANTIATE(3, 32)
HQT_INSTANTIATE(4, 32)
HQT_INSTANTIATE(5, 32)
}  // namespace hqt
