// Host-side access to the tensor-parallel "arm" state (tp_comm.cu): a launcher of the GEMM family calls
// tp_take_armed() exactly once per launch and copies the result into its kernel arguments. The returned struct has
// wait == signal == 0 unless hds_tp_arm() was called on this thread since the previous launch.
#pragma once
#include "hds_ptx.cuh"

namespace hds {
TpSync tp_take_armed();
bool tp_get_context(int ctx, TpSync* out);
}  // namespace hds
