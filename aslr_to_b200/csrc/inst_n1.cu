// explicit instantiation of every kernel for the 1-DOF models (nx = 4; SEA nu = 1, VSA nu = 2)
#include "kernels.cuh"
namespace aslr {
Launchers launchers_n1(bool vsa) { return vsa ? Launch<1, true>::table() : Launch<1, false>::table(); }
}  // namespace aslr
