// Kernel (b), rerun form (iir_rerun.cuh), instantiations for cascades of 4 biquad(s).
#include "iir_rerun.cuh"
namespace cbl {
iir_launch_fn iir_rerun_k4(int D, int in_kind) { return rerun_pick_launch<4>(D, in_kind); }
}
