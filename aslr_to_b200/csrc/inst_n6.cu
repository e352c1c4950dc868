// explicit instantiation of every kernel for the 6-DOF models (nx = 24; SEA nu = 6, VSA nu = 12)
#include "kernels.cuh"
namespace aslr {
Launchers launchers_n6(bool vsa) { return vsa ? Launch<6, true>::table() : Launch<6, false>::table(); }
}  // namespace aslr
