// common.cuh -- shared device helpers and the context object of libspatial_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <float.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <map>
#include <string>
#include <vector>

#include "../../include/spatial_b200.h"

#define SB_NUM_SMS 148 // B200: 2 dies x 74 SMs; grids are sized in multiples of this

// ---- error plumbing ---------------------------------------------------------------------------------------
void sb_set_error(const char *fmt, ...);

#define SB_CUDA(expr)                                                                                                \
	do {                                                                                                             \
		cudaError_t _e = (expr);                                                                                     \
		if (_e != cudaSuccess) {                                                                                     \
			sb_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__);                \
			return SB_ERR_CUDA;                                                                                      \
		}                                                                                                            \
	} while (0)

#define SB_TRY(expr)                                                                                                 \
	do {                                                                                                             \
		int _s = (expr);                                                                                             \
		if (_s != SB_OK) {                                                                                           \
			return _s;                                                                                               \
		}                                                                                                            \
	} while (0)

#define SB_ARG(cond, msg)                                                                                            \
	do {                                                                                                             \
		if (!(cond)) {                                                                                               \
			sb_set_error("invalid argument: %s", msg);                                                               \
			return SB_ERR_ARG;                                                                                       \
		}                                                                                                            \
	} while (0)

// ---- context ----------------------------------------------------------------------------------------------
// Scratch arena: bump allocator over device blocks. Reset at the start of every top-level call; blocks are only
// recycled in stream order (all work of a context runs on one stream), and coalesced into one block lazily.
struct sb_arena {
	struct block {
		char *ptr;
		size_t size;
	};
	std::vector<block> blocks;
	size_t cur_block = 0, cur_off = 0, high_water = 0, used = 0;
	cudaStream_t *stream = nullptr; // blocks come from the context's private stream-ordered pool, ordered on *stream
	cudaMemPool_t pool = nullptr;
	int alloc(size_t bytes, void **out);
	// mark / rewind: give back everything allocated after the mark (same stream => safe to reuse in stream order)
	struct mark_t {
		size_t cur_block, cur_off, used;
	};
	mark_t mark() const {
		return {cur_block, cur_off, used};
	}
	void rewind(const mark_t &m) {
		cur_block = m.cur_block, cur_off = m.cur_off, used = m.used;
	}
	void reset();
	void release();
};

struct sb_prof_entry {
	double ms = 0;
	uint64_t launches = 0;
	std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending;
};

struct sb_context {
	int device = 0;
	cudaStream_t stream = nullptr;
	cudaStream_t own_stream = nullptr;
	// private stream-ordered pool (release threshold = max): nothing this library caches shows up in, or changes the
	// behaviour of, the host process's default pool; destroyed (memory returned) with the context
	cudaMemPool_t pool = nullptr;
	sb_arena arena;
	uint64_t launches = 0;
	bool profiling = false;
	std::map<std::string, sb_prof_entry> prof;
	std::vector<cudaEvent_t> event_pool;
	// sb_context_set_option("probe_path"): 0 = auto (bulk point probes take the bucketed kernel), 1 = always the
	// per-probe tree walk, 2 = always bucketed
	int probe_path = 0;
	// sb_context_set_option("pip_path"): 0 = auto, 1 = FlatRTree candidates + per-pair refinement, 2 = polygon-centric
	int pip_path = 0;
	// sb_context_set_option("hist_sample"): 1 = size the bucket regions of the cell partition from a 1/16 sample of the
	// probe rows for large inputs (exact counts are recovered afterwards, overflow repeats with the exact histogram),
	// 0 = always the exact histogram pass
	int hist_sample = 1;
	// sb_context_set_option("str_path"): 0 = auto (STR leaf level: one-pass record partition when the layer has <= 512
	// slices, leaf nodes by a second select + partition per slice), 1 = always the bin-id sort + gather form used for
	// larger layers, 2 = record partition + segmented y sort, 3 = node partition of a slice by separate count / scan /
	// scatter kernels instead of inside the select kernel (what slices above 65535 entries take) (tests, A/B).
	int str_path = 0;
	// sb_context_set_option("trace", 1): synchronise at SB_TRACE points and print the wall time of each phase
	int trace = 0;
	double trace_t0 = 0;
	// second and third stream of the host-pointer join pipeline (H2D of the next chunk / D2H of the previous one)
	cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;
	uint64_t pipe_chunk = 8u << 20;  // probe rows per chunk of that pipeline (option "pipe_chunk")
	// pinned staging for small host<->device transfers (counters)
	uint64_t *pinned_u64 = nullptr, *pinned_u64_dev = nullptr;

	int scratch(size_t bytes, void **out) {
		return arena.alloc(bytes, out);
	}
	template <class T>
	int scratch_t(size_t count, T **out) {
		return arena.alloc(count * sizeof(T), reinterpret_cast<void **>(out));
	}
};

// RAII profiling scope: records two events around the launches of one kernel class when profiling is on.
struct sb_prof_scope {
	sb_context *ctx;
	cudaEvent_t a = nullptr, b = nullptr;
	std::string tag;
	sb_prof_scope(sb_context *c, const char *t);
	~sb_prof_scope();
};

// long-lived device buffers (trees): from the context's pool, on its stream
static inline cudaError_t sb_pool_alloc(sb_context *ctx, void **out, size_t bytes) {
	return cudaMallocFromPoolAsync(out, bytes, ctx->pool, ctx->stream);
}
// frees a tree buffer in stream order on its building context's stream when that context is still alive (registry in
// context.cu); otherwise after a device synchronise
void sb_owner_free(sb_context *owner, int device, void *ptr);

void sb_trace_point(sb_context *ctx, const char *what);
#define SB_TRACE(ctx, what)                                                                                          \
	do {                                                                                                             \
		if ((ctx)->trace & 1) {                                                                                      \
			sb_trace_point(ctx, what);                                                                               \
		}                                                                                                            \
	} while (0)

// stage helpers for SB_HOST calls
template <class T>
static inline int sb_to_device(sb_context *ctx, sb_mem mem, const T *src, size_t count, const T **dev) {
	if (mem == SB_DEVICE || src == nullptr || count == 0) {
		*dev = src;
		return SB_OK;
	}
	T *d = nullptr;
	SB_TRY(ctx->scratch_t<T>(count, &d));
	SB_CUDA(cudaMemcpyAsync(d, src, count * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
	*dev = d;
	return SB_OK;
}

template <class T>
static inline int sb_out_buffer(sb_context *ctx, sb_mem mem, T *dst, size_t count, T **dev) {
	if (mem == SB_DEVICE || dst == nullptr || count == 0) {
		*dev = dst;
		return SB_OK;
	}
	return ctx->scratch_t<T>(count, dev);
}

template <class T>
static inline int sb_to_host(sb_context *ctx, sb_mem mem, T *dst, const T *dev, size_t count) {
	if (mem == SB_DEVICE || dst == nullptr || count == 0) {
		return SB_OK;
	}
	SB_CUDA(cudaMemcpyAsync(dst, dev, count * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
	return SB_OK;
}

static inline unsigned sb_grid(uint64_t work_items, unsigned per_block, unsigned max_waves = 64) {
	// enough blocks to cover the work, capped at a multiple of the SM count (persistent grid-stride beyond that)
	uint64_t b = (work_items + per_block - 1) / per_block;
	const uint64_t cap = (uint64_t)SB_NUM_SMS * max_waves;
	if (b > cap) {
		b = cap;
	}
	return b == 0 ? 1u : (unsigned)b;
}

#define SB_LAUNCH_CHECK(ctx)                                                                                         \
	do {                                                                                                             \
		(ctx)->launches++;                                                                                           \
		SB_CUDA(cudaGetLastError());                                                                                 \
	} while (0)

// Opt a kernel into more than 48 KB of dynamic shared memory. The attribute is per device and the C-ABI allows contexts
// on several devices (and threads) of one process, so it is set once per (call site, device), race-free.
#define SB_SMEM_OPTIN(ctx, func, bytes)                                                                              \
	do {                                                                                                             \
		static std::atomic<uint64_t> _done{0};                                                                       \
		const uint64_t _bit = 1ull << ((ctx)->device & 63);                                                          \
		if (!(_done.load(std::memory_order_acquire) & _bit)) {                                                       \
			SB_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)));          \
			_done.fetch_or(_bit, std::memory_order_release);                                                         \
		}                                                                                                            \
	} while (0)

// ---- device helpers ---------------------------------------------------------------------------------------
#ifdef __CUDACC__

// MathUtil::DoubleToFloatDown / Up (src/spatial/util/math.hpp:19-47): clamp to +-FLT_MAX, else directed rounding.
__device__ __forceinline__ float sb_f64_to_f32_down(double d) {
	if (d > (double)FLT_MAX) {
		return FLT_MAX;
	}
	if (d <= (double)-FLT_MAX) {
		return -FLT_MAX;
	}
	return __double2float_rd(d);
}
__device__ __forceinline__ float sb_f64_to_f32_up(double d) {
	if (d >= (double)FLT_MAX) {
		return FLT_MAX;
	}
	if (d < (double)-FLT_MAX) {
		return -FLT_MAX;
	}
	return __double2float_ru(d);
}

// static_cast<uint32_t>(fp) as the x86-64 reference build executes it: cvtts[sd]2si to a 64-bit register, low
// 32 bits kept; out of range / NaN = "integer indefinite" 0x8000000000000000 -> 0.
__device__ __forceinline__ uint32_t sb_cast_f64_u32(double v) {
	if (!(v > -9223372036854775808.0 && v < 9223372036854775808.0)) {
		return 0u;
	}
	return (uint32_t)(unsigned long long)__double2ll_rz(v);
}
__device__ __forceinline__ uint32_t sb_cast_f32_u32(float v) {
	if (!(v > -9223372036854775808.0f && v < 9223372036854775808.0f)) {
		return 0u;
	}
	return (uint32_t)(unsigned long long)__float2ll_rz(v);
}

__device__ __forceinline__ uint32_t sb_interleave16(uint32_t x) {
	x = (x | (x << 8)) & 0x00FF00FFu;
	x = (x | (x << 4)) & 0x0F0F0F0Fu;
	x = (x | (x << 2)) & 0x33333333u;
	x = (x | (x << 1)) & 0x55555555u;
	return x;
}

// sgl::util::hilbert_encode(16, x, y), src/sgl/sgl.hpp:1799-1848, for n = 16 (no pre-shift). x, y may carry
// bits above 16 (ST_Hilbert(geom) feeds 32-bit values); they flow through exactly as on the CPU.
__device__ __forceinline__ uint32_t sb_hilbert16(uint32_t x, uint32_t y) {
	uint32_t A, B, C, D;
	{
		const uint32_t a = x ^ y;
		const uint32_t b = 0xFFFFu ^ a;
		const uint32_t c = 0xFFFFu ^ (x | y);
		const uint32_t d = x & (y ^ 0xFFFFu);
		A = a | (b >> 1);
		B = (a >> 1) ^ a;
		C = ((c >> 1) ^ (b & (d >> 1))) ^ c;
		D = ((a & (c >> 1)) ^ (d >> 1)) ^ d;
	}
#pragma unroll
	for (int s = 2; s <= 4; s <<= 1) {
		const uint32_t a = A, b = B, c = C, d = D;
		A = (a & (a >> s)) ^ (b & (b >> s));
		B = (a & (b >> s)) ^ (b & ((a ^ b) >> s));
		C ^= (a & (c >> s)) ^ (b & (d >> s));
		D ^= (b & (c >> s)) ^ ((a ^ b) & (d >> s));
	}
	{
		const uint32_t a = A, b = B, c = C, d = D;
		C ^= (a & (c >> 8)) ^ (b & (d >> 8));
		D ^= (b & (c >> 8)) ^ ((a ^ b) & (d >> 8));
	}
	const uint32_t a = C ^ (C >> 1);
	const uint32_t b = D ^ (D >> 1);
	const uint32_t i0 = x ^ y;
	const uint32_t i1 = b | (0xFFFFu ^ (i0 | a));
	return (sb_interleave16(i1) << 1) | sb_interleave16(i0);
}

// sgl::util::hilbert_f32_to_u32, sgl.hpp:1850-1862
__device__ __forceinline__ uint32_t sb_f32_order_u32(float f) {
	if (isnan(f)) {
		return 0xFFFFFFFFu;
	}
	const uint32_t r = __float_as_uint(f);
	return (r & 0x80000000u) ? ~r : (r | 0x80000000u);
}

// order-preserving float -> uint32 for radix sorting, with -0.0f folded onto +0.0f (a comparison sort treats
// them as equal) so ties break by input position exactly like the oracle's comparator.
__device__ __forceinline__ uint32_t sb_f32_sort_key(float f) {
	f = f + 0.0f; // -0 -> +0
	const uint32_t r = __float_as_uint(f);
	return (r & 0x80000000u) ? ~r : (r | 0x80000000u);
}

// Box2D<float>::Intersects, src/spatial/geometry/bbox.hpp:27-29 (closed intervals)
__device__ __forceinline__ bool sb_box_intersects(const float4 a, const float4 b) {
	return !(a.x > b.z || a.z < b.x || a.y > b.w || a.w < b.y);
}

__device__ __forceinline__ unsigned sb_lane() {
	return threadIdx.x & 31u;
}
__device__ __forceinline__ unsigned sb_lanemask_lt() {
	unsigned m;
	asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
	return m;
}

#endif // __CUDACC__
