// builder.cu -- index construction on the device.  Placeholder until
// the query path is parity-green on hardware.
#include "hb_internal.h"

namespace hb {
Index* build_hnsw(const uint64_t*, uint64_t, const hb_build_params&, int, bool) {
  throw Error(HB_ERR_UNSUPPORTED, "hb_build_hnsw is not implemented yet");
}
}  // namespace hb
