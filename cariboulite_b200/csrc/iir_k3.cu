// Kernel (b) instantiations for cascades of 3 biquad(s), double-staged.  The float-staged ones: iir_k3y.cu.
#include "iir_kernel.cuh"
namespace cbl {
iir_launch_fn iir_launch_k3y(int D, int in_kind, bool par);
iir_launch_fn iir_launch_k3(int D, int in_kind, bool par, bool y32)
{
    return y32 ? iir_launch_k3y(D, in_kind, par) : iir_pick_launch<3, false>(D, in_kind, par);
}
}
