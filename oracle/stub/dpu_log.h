/* TEST INFRASTRUCTURE ONLY -- empty stand-in for the UPMEM <dpu_log.h>; see dpu.h in this directory. */
