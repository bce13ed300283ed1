// osc_fused.cuh -- single-launch tick (updateModel + updateControllerTaskModels + computeControlTorques).
// Placeholder until the fused kernel lands: osc_tick falls back to the API-faithful kernel sequence.
#pragma once
#include "osc_host.h"
#include "osc_kernels.cuh"

namespace osc {
template <class T, int N> bool fused_tick_supported(Batch&) { return false; }
template <class T, int N> int fused_tick_launch(Batch&, const ChainC<T, N>&, int) { return OSC_ERR_UNSUPPORTED; }
}  // namespace osc
