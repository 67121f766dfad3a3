// Error reporting and version string of the C ABI (include/origami_b200.h).
#include <cstdarg>
#include <cstdio>

#include "og_common.cuh"
#include "og_internal.cuh"

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

#define OG_MAX_DEVICES 64

int og_device_ordinal() {
	int dev = -1;
	return cudaGetDevice(&dev) == cudaSuccess ? dev : -1;
}

// Persistent grids are sized in multiples of the SM count of the device the call runs on.
int og_num_sms() {
	static int cached[OG_MAX_DEVICES] = {0};
	const int dev = og_device_ordinal();
	if (dev < 0 || dev >= OG_MAX_DEVICES) return 148;
	if (cached[dev] == 0) {
		int v = 0;
		if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
		cached[dev] = v;  // benign race: every thread writes the same value
	}
	return cached[dev];
}

bool og_first_use(int slot) {
	static unsigned char seen[OG_MAX_DEVICES][64] = {{0}};
	const int dev = og_device_ordinal();
	if (dev < 0 || dev >= OG_MAX_DEVICES || slot < 0 || slot >= 64) return true;
	if (seen[dev][slot]) return false;
	seen[dev][slot] = 1;
	return true;
}

extern "C" const char *og_last_error(void) { return g_err; }

extern "C" const char *og_version(void) { return "origami_b200 0.1 sm_100a"; }
