// explicit instantiation of every kernel for the 7-DOF models (nx = 28; SEA nu = 7, VSA nu = 14)
#include "kernels.cuh"
namespace aslr {
Launchers launchers_n7(bool vsa) { return vsa ? Launch<7, true>::table() : Launch<7, false>::table(); }
}  // namespace aslr
