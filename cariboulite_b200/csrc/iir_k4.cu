// Kernel (b) instantiations for cascades of 4 biquad(s), double-staged.  The float-staged ones: iir_k4y.cu.
#include "iir_kernel.cuh"
namespace cbl {
iir_launch_fn iir_launch_k4y(int D, int in_kind, bool par);
iir_launch_fn iir_launch_k4(int D, int in_kind, bool par, bool y32)
{
    return y32 ? iir_launch_k4y(D, in_kind, par) : iir_pick_launch<4, false>(D, in_kind, par);
}
}
