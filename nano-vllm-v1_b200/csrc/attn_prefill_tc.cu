// (2, variant 2) tcgen05 / TMEM / TMA prefill -- placeholder until the kernel lands.
#include "launch.h"
namespace b200 {
bool prefill_tc_available() { return false; }
int launch_attn_prefill_tc(const PrefillArgs&, cudaStream_t) { return -6; }
}  // namespace b200
