// Internal declarations shared by ti_host.cpp and ti_api.cu (not part of the C ABI).
#ifndef TI_INTERNAL_H
#define TI_INTERNAL_H

#include <stdarg.h>

#include "../../include/traceingest.h"

namespace ti {
void set_error(const char *fmt, ...);
}

// host finalisation of the counter block for an explicit configuration
ti_status ti_counters_finalize_cfg(const ti_config *cfg, const ti_counters *c, ti_power_counters *out);

#endif
