// Library identification for libb200sde.so.
#include "common.cuh"
#include "../../include/b200_sde.h"

extern "C" int b200_abi_version(void) { return 1; }
extern "C" const char* b200_build_info(void) {
  return "libb200sde sm_100a (compute_100a) built " __DATE__ " " __TIME__;
}
