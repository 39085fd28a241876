// e3_exclusive_scan_noalloc: the instance-offset prefix sum for callers that cannot provide
// (or obtain) scratch memory.
//
// Who needs this. The regular entry point e3_exclusive_scan() wants a caller-owned scratch buffer.
// Some hosts of this library have no way to produce one at the moment they call us:
//   * an application that grabbed its whole device-memory budget at start-up and runs us from a
//     frame/step loop in which cudaMalloc is banned (it would stall the device: cudaMalloc and
//     cudaFree synchronise, and a real-time viewer or a serving loop cannot afford that);
//   * a plugin host that hands us input and output arrays only and gives us no allocator hook;
//   * code that already sits inside a region where allocation calls are not permitted.
// For them the scratch lives in this translation unit, in a fixed-size device array that every GPU
// loading the library carries from load to exit (32.8 KB), and the entry point calls no cudaMalloc,
// cudaFree or any other allocating routine.
//
// Limits and rules for the caller:
//   * n <= E3_NOALLOC_SCAN_MAX (16,777,216 values); larger inputs are refused with
//     E3_ERR_TOO_LARGE and nothing is launched;
//   * the FIRST call on each device looks up the device address of the array (one
//     cudaGetSymbolAddress, with that device current). Make that first call during start-up,
//     before any stream capture begins;
//   * all calls on one device share the one array. It keeps whatever the previous call left in
//     it; that is harmless here because each call writes every scratch word before it reads it
//     (block sums, then their scan), so nothing has to start at zero and nothing is cleared.
//     But two calls that overlap on the GPU (different streams, or different host threads) would
//     scribble over each other's block sums: serialise them on one stream or with events.
#include "common.cuh"
#include "../../include/event3dgs.h"

#include <mutex>

namespace {

constexpr size_t kScratchWords = E3_NOALLOC_SCAN_MAX / E3_SCAN_TILE + 8;

// In host code this name is only a handle for the runtime, not a device address.
static __device__ __align__(256) uint32_t g_noalloc_scan_scratch[kScratchWords];

constexpr int kMaxDevices = 64;
uint32_t* g_scratch_addr[kMaxDevices] = {};
std::mutex g_scratch_mu;

// Device address of the array on the CURRENT device (looked up once per device).
cudaError_t scratch_for_current_device(uint32_t** out)
{
	int dev = 0;
	cudaError_t e = cudaGetDevice(&dev);
	if (e != cudaSuccess) return e;
	if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
	std::lock_guard<std::mutex> lock(g_scratch_mu);
	if (!g_scratch_addr[dev]) {
		uint32_t* ptr = nullptr;
		e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scan_scratch);
		if (e != cudaSuccess) return e;
		g_scratch_addr[dev] = ptr;
	}
	*out = g_scratch_addr[dev];
	return cudaSuccess;
}

}  // namespace

extern "C" {

size_t e3_exclusive_scan_noalloc_max(void) { return (size_t)E3_NOALLOC_SCAN_MAX; }

int e3_exclusive_scan_noalloc(const uint32_t* in, uint32_t* out, size_t n, uint32_t* total, void* stream)
{
	cudaStream_t st = (cudaStream_t)stream;
	if (n > (size_t)E3_NOALLOC_SCAN_MAX) return E3_ERR_TOO_LARGE;
	if (n == 0) {
		if (total) { cudaError_t e = cudaMemsetAsync(total, 0, sizeof(uint32_t), st); if (e != cudaSuccess) return (int)e; }
		return 0;
	}
	if (!in || !out) return E3_ERR_BAD_ARGUMENT;
	static_assert(kScratchWords >= (E3_NOALLOC_SCAN_MAX + E3_SCAN_TILE - 1) / E3_SCAN_TILE + 2, "scratch too small");
	uint32_t* scratch = nullptr;
	cudaError_t e = scratch_for_current_device(&scratch);
	if (e != cudaSuccess) return (int)e;
	e = e3_exclusive_scan_u32(in, nullptr, out, n, scratch, total, st);
	return (int)e;
}

}  // extern "C"
