// explicit instantiation of every kernel for the 5-DOF models (nx = 20; SEA nu = 5, VSA nu = 10)
#include "kernels.cuh"
namespace aslr {
Launchers launchers_n5(bool vsa) { return vsa ? Launch<5, true>::table() : Launch<5, false>::table(); }
}  // namespace aslr
