// Helpers shared by the translation units of librpde_b200.so (defined in rpde.cu); not part of the C ABI.
#pragma once
int rpde_internal_fail(int code, const char* msg);   // records msg for rpde_last_error() (thread-local), returns code
void rpde_internal_count_launch(int tensor_core);    // launch statistics (rpde_launch_count / rpde_tc_launch_count)
int rpde_internal_precision(void);                   // the calling thread's engine selection (RPDE_PREC_*)
