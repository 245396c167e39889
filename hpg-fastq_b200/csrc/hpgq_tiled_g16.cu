// hpgq_tiled_g16.cu -- instantiations of the tiled QC kernel for read groups of 16 lanes (see hpgq_tiled.cuh)
#include "hpgq_tiled_body.inc"
namespace hqt {
HQT_INSTANTIATE(3, 16)
HQT_INSTANTIATE(4, 16)
HQT_INSTANTIATE(5, 16)
}  // namespace hqt
