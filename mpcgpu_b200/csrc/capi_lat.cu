// capi_lat.cu -- the fp32 cluster "latency tilings" (launch.cuh: MPCB200_LAT_CFGS), instantiated here so that
// they compile in parallel with the throughput tilings of capi.cu.
#include "launch.cuh"

namespace mpcb200 {
MPCB200_LAT_CFGS(MPCB200_LAT_DEF)
}  // namespace mpcb200
