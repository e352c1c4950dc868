// explicit instantiation of every kernel for the 3-DOF models (nx = 12; SEA nu = 3, VSA nu = 6)
#include "kernels.cuh"
namespace aslr {
Launchers launchers_n3(bool vsa) { return vsa ? Launch<3, true>::table() : Launch<3, false>::table(); }
}  // namespace aslr
