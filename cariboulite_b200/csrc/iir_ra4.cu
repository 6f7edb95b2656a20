// Kernel (b), rerun form with the AP phase 3 (iir_rerun.cuh), instantiations for cascades of 4 biquad(s).
#include "iir_rerun.cuh"
namespace cbl {
iir_launch_fn iir_rerun_ap_k4(int D, int in_kind) { return rerun_ap_pick_launch<4>(D, in_kind); }
}
