// Error reporting and version string of the C ABI (include/origami_b200.h).
#include <cstdarg>
#include <cstdio>

#include "og_common.cuh"

static thread_local char g_err[512] = "";

int og_set_error(int code, const char *fmt, ...) {
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_err, sizeof(g_err), fmt, ap);
	va_end(ap);
	return code;
}

int og_check_launch(const char *what) {
	cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess)
		return og_set_error((int)e, "%s: %s", what, cudaGetErrorString(e));
	return OG_OK;
}

extern "C" const char *og_last_error(void) { return g_err; }

extern "C" const char *og_version(void) { return "origami_b200 0.1 sm_100a"; }
