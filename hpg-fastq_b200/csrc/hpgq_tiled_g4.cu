// hpgq_tiled_g4.cu -- instantiations of the tiled QC kernel for read groups of 4 lanes (see hpgq_tiled.cuh)
#include "hpgq_tiled_body.inc"
namespace hqt {
HQT_INSTANTIATE(2, 4)
HQT_INSTANTIATE(3, 4)
HQT_INSTANTIATE(4, 4)
HQT_INSTANTIATE(5, 4)
HQT_INSTANTIATE(7, 4)
}  // namespace hqt
