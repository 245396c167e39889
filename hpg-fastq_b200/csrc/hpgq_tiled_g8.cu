// hpgq_tiled_g8.cu -- instantiations of the tiled QC kernel for read groups of 8 lanes (see hpgq_tiled.cuh)
#include "hpgq_tiled_body.inc"
namespace hqt {
HQT_INSTANTIATE(3, 8)
HQT_INSTANTIATE(4, 8)
HQT_INSTANTIATE(5, 8)
}  // namespace hqt
