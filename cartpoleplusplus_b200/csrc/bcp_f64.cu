// FP64 parity mode: compiled with -fmad=false so every operation rounds once, in the
// order written - the same IEEE sequence the CPU oracle executes.
#include "bcp_launch_impl.cuh"
namespace bcp {
template struct Launch<double>;
}  // namespace bcp
