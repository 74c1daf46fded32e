// Library-internal hooks between translation units of libaps.so (hidden visibility: not part of the ABI).
#pragma once
#include "../../include/aps.h"

namespace aps {
// The configuration a live store was opened with (defaults resolved).  APS_ECLOSED on a closed handle.
int store_config(aps_store* s, aps_config* out);
}  // namespace aps
