// TEST INFRASTRUCTURE: stands in for ISA_Def/trace_opcode.h (the command loop needs nothing of it)
