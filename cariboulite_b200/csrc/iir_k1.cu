// Kernel (b) instantiations for cascades of 1 biquad(s), double-staged.  The float-staged ones: iir_k1y.cu.
#include "iir_kernel.cuh"
namespace cbl {
iir_launch_fn iir_launch_k1y(int D, int in_kind, bool par);
iir_launch_fn iir_launch_k1(int D, int in_kind, bool par, bool y32)
{
    return y32 ? iir_launch_k1y(D, in_kind, par) : iir_pick_launch<1, false>(D, in_kind, par);
}
}
