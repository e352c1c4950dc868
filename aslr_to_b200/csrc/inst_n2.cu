// explicit instantiation of every kernel for the 2-DOF models (nx = 8; SEA nu = 2, VSA nu = 4)
#include "kernels.cuh"
namespace aslr {
Launchers launchers_n2(bool vsa) { return vsa ? Launch<2, true>::table() : Launch<2, false>::table(); }
}  // namespace aslr
