// Fused cluster path (placeholder until the DSMEM kernel lands): reports "not supported" so
// sfpe_project_batch takes the generic path.
#include "sfpe_internal.h"

namespace sfpe {

bool fused_supported(const DevGrid&, const PointSrc&, const DevOut&, int) { return false; }

cudaError_t launch_fused(const DevGrid&, const PointSrc&, const FrameSet&, long long, const Scratch&, const DevOut&,
                         int, cudaStream_t, long long*) {
  return cudaErrorNotSupported;
}

}  // namespace sfpe
