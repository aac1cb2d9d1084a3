// TEST INFRASTRUCTURE: stands in for gpu-simulator/trace-driven/trace_driven.h (see loop_engine.h).
#include "trace-parser/trace_parser.h"
#include "loop_engine.h"
