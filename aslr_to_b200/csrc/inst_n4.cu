// explicit instantiation of every kernel for the 4-DOF models (nx = 16; SEA nu = 4, VSA nu = 8)
#include "kernels.cuh"
namespace aslr {
Launchers launchers_n4(bool vsa) { return vsa ? Launch<4, true>::table() : Launch<4, false>::table(); }
}  // namespace aslr
