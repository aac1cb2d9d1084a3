// TEST INFRASTRUCTURE: stands in for the simulator header of the same name (see loop_engine.h)
