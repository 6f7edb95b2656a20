// Kernel (b) instantiations for cascades of 3 biquad(s).
#include "iir_kernel.cuh"
namespace cbl {
iir_launch_fn iir_launch_k3(int D, int in_kind, bool par) { return iir_pick_launch<3>(D, in_kind, par); }
iir_occ_fn iir_occ_k3(int D, int in_kind, bool par) { return iir_pick_occ<3>(D, in_kind, par); }
}
