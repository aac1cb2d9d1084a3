// TEST INFRASTRUCTURE: the trace parser the loop is built against -- the reference's own header or the drop-in's,
// chosen by the recipe in oracle/Makefile (-DTI_LOOP_TRACE_PARSER_H="...").
#include TI_LOOP_TRACE_PARSER_H
