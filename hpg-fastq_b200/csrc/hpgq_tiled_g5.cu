// hpgq_tiled_g5.cu -- instantiations of the tiled QC kernel for read groups of 5 lanes (see hpgq_tiled.cuh)
#include "hpgq_tiled_body.inc"
namespace hqt {
HQT_INSTANTIATE(5, 5)
}  // namespace hqt
