// noalloc.cu -- the one entry point of the library that allocates nothing (sb_pairs_checksum_noalloc).
//
// For callers that may not call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid
// it. It is sb_pairs_checksum (refine.cu k_pairs_checksum, whose two accumulators live in the context's scratch arena)
// restated so that its scratch -- one (count, sum) partial per 4096-pair tile -- is a file-scope __device__ array of a
// fixed size. No context, no arena, no pool: the call is two kernel launches on the caller's stream.
//
// The array is resident on every GPU that loads the library (256 KiB) until the process exits. ALL calls on a device
// share it: two calls that overlap on the GPU (different streams, or two host threads) race on the partials and give
// wrong sums; calls on one stream are ordered and safe. Nothing in it has to start at zero: pass 1 overwrites the
// partial of every tile pass 2 reads, so the contents a previous call left behind are never observed.
#include <mutex>

#include "common.cuh"

#define NA_TILE 4096u                        // pairs per block of pass 1 (256 threads x 16)
#define NA_MAX_TILES (SB_NOALLOC_MAX_PAIRS / NA_TILE)
#define NA_MAX_DEVICES 64

// the scratch: [tile][0] = pairs of the tile inside [lo, hi), [tile][1] = wrapping sum of their mixes
static __device__ __align__(256) unsigned long long g_na_partials[2 * NA_MAX_TILES];

// same mix as refine.cu sb_pair_mix (and oracle/ref_harness.cpp, tests)
__device__ __forceinline__ unsigned long long na_pair_mix(uint32_t p, uint32_t b) {
	unsigned long long k = (unsigned long long)p * 2654435761ull + (unsigned long long)b * 2246822519ull;
	k ^= k >> 29;
	k *= 0xBF58476D1CE4E5B9ull;
	k ^= k >> 32;
	return k;
}

__global__ void __launch_bounds__(256) k_na_partials(const uint32_t *__restrict__ pp, const uint32_t *__restrict__ pb,
                                                     uint64_t n, uint32_t lo, uint32_t hi,
                                                     unsigned long long *__restrict__ partials) {
	__shared__ unsigned long long s_cnt[8], s_sum[8];
	const uint64_t base = (uint64_t)blockIdx.x * NA_TILE;
	unsigned long long cnt = 0, sum = 0;
#pragma unroll 4
	for (uint32_t k = 0; k < NA_TILE / 256; k++) {
		const uint64_t i = base + k * 256u + threadIdx.x;
		if (i < n) {
			const uint32_t p = pp[i];
			if (p >= lo && p < hi) {
				cnt++;
				sum += na_pair_mix(p, pb[i]);
			}
		}
	}
#pragma unroll
	for (int s = 16; s > 0; s >>= 1) {
		cnt += __shfl_xor_sync(0xFFFFFFFFu, cnt, s);
		sum += __shfl_xor_sync(0xFFFFFFFFu, sum, s);
	}
	if ((threadIdx.x & 31) == 0) {
		s_cnt[threadIdx.x >> 5] = cnt;
		s_sum[threadIdx.x >> 5] = sum;
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		for (int w = 1; w < 8; w++) {
			cnt += s_cnt[w];
			sum += s_sum[w];
		}
		partials[2 * (uint64_t)blockIdx.x] = cnt; // every tile is written, so stale contents never reach pass 2
		partials[2 * (uint64_t)blockIdx.x + 1] = sum;
	}
}

__global__ void __launch_bounds__(1024) k_na_total(const unsigned long long *__restrict__ partials, uint32_t tiles,
                                                   unsigned long long *__restrict__ out2) {
	__shared__ unsigned long long s_cnt[32], s_sum[32];
	unsigned long long cnt = 0, sum = 0;
	for (uint32_t t = threadIdx.x; t < tiles; t += 1024) {
		cnt += partials[2 * (uint64_t)t];
		sum += partials[2 * (uint64_t)t + 1];
	}
#pragma unroll
	for (int s = 16; s > 0; s >>= 1) {
		cnt += __shfl_xor_sync(0xFFFFFFFFu, cnt, s);
		sum += __shfl_xor_sync(0xFFFFFFFFu, sum, s);
	}
	if ((threadIdx.x & 31) == 0) {
		s_cnt[threadIdx.x >> 5] = cnt;
		s_sum[threadIdx.x >> 5] = sum;
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		for (int w = 1; w < 32; w++) {
			cnt += s_cnt[w];
			sum += s_sum[w];
		}
		out2[0] = cnt;
		out2[1] = sum;
	}
}

// In host code `g_na_partials` is only the host-side stand-in of the symbol, not a device address: the address the
// kernels get is looked up once per device, with that device current.
static std::mutex g_na_mu;
static unsigned long long *g_na_dev_ptr[NA_MAX_DEVICES];

static int na_scratch_for_current_device(int device, unsigned long long **out) {
	std::lock_guard<std::mutex> lock(g_na_mu);
	if (!g_na_dev_ptr[device]) {
		unsigned long long *ptr = nullptr;
		SB_CUDA(cudaGetSymbolAddress((void **)&ptr, g_na_partials));
		g_na_dev_ptr[device] = ptr;
	}
	*out = g_na_dev_ptr[device];
	return SB_OK;
}

namespace {
struct device_guard {
	int prev = -1;
	bool switched = false;
	~device_guard() {
		if (switched) {
			cudaSetDevice(prev);
		}
	}
};
} // namespace

extern "C" int sb_pairs_checksum_noalloc(int device, void *cuda_stream, const uint32_t *pair_probe,
                                         const uint32_t *pair_build, uint64_t count, uint32_t probe_lo,
                                         uint32_t probe_hi, uint64_t *out_count_hash) {
	SB_ARG(out_count_hash != nullptr, "out_count_hash (device pointer to two 64-bit words)");
	if (count > (uint64_t)SB_NOALLOC_MAX_PAIRS) {
		sb_set_error("sb_pairs_checksum_noalloc: count %llu exceeds SB_NOALLOC_MAX_PAIRS (%llu): the static scratch holds "
		             "one partial per %u-pair tile; use sb_pairs_checksum, or split the list and add the results",
		             (unsigned long long)count, (unsigned long long)SB_NOALLOC_MAX_PAIRS, NA_TILE);
		return SB_ERR_ARG;
	}
	SB_ARG(count == 0 || (pair_probe != nullptr && pair_build != nullptr), "pairs");
	device_guard guard;
	SB_CUDA(cudaGetDevice(&guard.prev));
	if (device < 0) {
		device = guard.prev;
	}
	SB_ARG(device < NA_MAX_DEVICES, "device");
	if (device != guard.prev) {
		SB_CUDA(cudaSetDevice(device));
		guard.switched = true;
	}
	unsigned long long *partials = nullptr;
	SB_TRY(na_scratch_for_current_device(device, &partials));
	cudaStream_t stream = (cudaStream_t)cuda_stream;
	const uint32_t tiles = (uint32_t)((count + NA_TILE - 1) / NA_TILE);
	if (tiles) {
		k_na_partials<<<tiles, 256, 0, stream>>>(pair_probe, pair_build, count, probe_lo, probe_hi, partials);
		SB_CUDA(cudaGetLastError());
	}
	k_na_total<<<1, 1024, 0, stream>>>(partials, tiles, (unsigned long long *)out_count_hash);
	SB_CUDA(cudaGetLastError());
	return SB_OK;
}
