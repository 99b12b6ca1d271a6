/* TEST INFRASTRUCTURE ONLY -- empty stand-in for the UPMEM <dpu_memory.h>; see dpu.h in this directory. */
