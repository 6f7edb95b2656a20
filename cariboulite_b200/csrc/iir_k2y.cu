// Kernel (b) instantiations for cascades of 2 biquad(s), float-staged (CF32 output without int16 truncation).
#include "iir_kernel.cuh"
namespace cbl {
iir_launch_fn iir_launch_k2y(int D, int in_kind, bool par) { return iir_pick_launch<2, true>(D, in_kind, par); }
}
