// Error plumbing and version for the C ABI.
#include "../../include/vistaf.h"
#include "common.cuh"

char g_vistaf_err[512] = "";

int vistaf_set_err(const char* what, cudaError_t e, const char* file, int line) {
  snprintf(g_vistaf_err, sizeof(g_vistaf_err), "%s: %s (%s:%d)", what, cudaGetErrorString(e), file, line);
  return 1;
}

extern "C" const char* vistaf_last_error(void) { return g_vistaf_err; }
extern "C" int vistaf_version(void) { return 100; }
