// context.cu -- context, scratch arena, error string, kernel-class profiling.
#include "common.cuh"

#include <stdarg.h>

#include <mutex>
#include <set>

static thread_local char g_err[512] = "";

static std::mutex g_live_mu;
static std::set<sb_context *> g_live; // contexts that exist: trees free on their owner's stream only while it lives

void sb_set_error(const char *fmt, ...) {
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_err, sizeof(g_err), fmt, ap);
	va_end(ap);
}

extern "C" const char *sb_last_error(void) {
	return g_err;
}

extern "C" const char *sb_version(void) {
	return "spatial_b200 0.1 (sm_100a)";
}

#include <chrono>
void sb_trace_point(sb_context *ctx, const char *what) {
	cudaStreamSynchronize(ctx->stream);
	const double t = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
	fprintf(stderr, "[sb trace] %-28s +%9.3f ms\n", what, ctx->trace_t0 ? t - ctx->trace_t0 : 0.0);
	ctx->trace_t0 = t;
}

// ---- small transfers that stay off the copy engines ----------------------------------------------------------
// A counter read-back or a small memset issued as cudaMemcpyAsync / cudaMemsetAsync runs on a copy engine; inside the
// host-pointer join pipeline that engine is busy with ~20 ms bulk copies of the previous chunk's pairs and the small
// transfer queues behind them, which serialised probe and copy-out (measured: 282 -> see DESIGN.md section 5). So
// counters are stored to mapped pinned memory by a one-block kernel and small clears are kernels too.
__global__ void k_read_words(const uint64_t *__restrict__ src, uint64_t *__restrict__ dst, int n) {
	if ((int)threadIdx.x < n) {
		dst[threadIdx.x] = src[threadIdx.x];
	}
	__threadfence_system();
}

int sbi_read_words(sb_context *ctx, const void *dev, int nwords, uint64_t *host) {
	SB_ARG(nwords >= 1 && nwords <= 64, "sbi_read_words: 1..64 words");
	k_read_words<<<1, 64, 0, ctx->stream>>>(static_cast<const uint64_t *>(dev), ctx->pinned_u64_dev, nwords);
	SB_CUDA(cudaGetLastError());
	SB_CUDA(cudaStreamSynchronize(ctx->stream));
	memcpy(host, ctx->pinned_u64, (size_t)nwords * sizeof(uint64_t));
	return SB_OK;
}

__global__ void __launch_bounds__(256) k_zero_words(uint32_t *__restrict__ p, uint64_t nwords) {
	for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < nwords; i += (uint64_t)gridDim.x * blockDim.x) {
		p[i] = 0;
	}
}

int sbi_zero(sb_context *ctx, void *dev, size_t bytes) {
	SB_ARG((bytes & 3) == 0 && (reinterpret_cast<uintptr_t>(dev) & 3) == 0, "sbi_zero: 4-byte granules");
	if (bytes == 0) {
		return SB_OK;
	}
	const uint64_t nwords = bytes / 4;
	uint64_t blocks = (nwords + 255) / 256;
	blocks = blocks > 1184 ? 1184 : blocks;
	k_zero_words<<<(unsigned)blocks, 256, 0, ctx->stream>>>(static_cast<uint32_t *>(dev), nwords);
	SB_CUDA(cudaGetLastError());
	return SB_OK;
}

// ---- arena --------------------------------------------------------------------------------------------------
int sb_arena::alloc(size_t bytes, void **out) {
	bytes = (bytes + 255) & ~(size_t)255;
	if (bytes == 0) {
		bytes = 256;
	}
	while (cur_block < blocks.size()) {
		block &b = blocks[cur_block];
		if (cur_off + bytes <= b.size) {
			*out = b.ptr + cur_off;
			cur_off += bytes;
			used += bytes;
			if (used > high_water) {
				high_water = used;
			}
			return SB_OK;
		}
		cur_block++;
		cur_off = 0;
	}
	size_t sz = bytes > ((size_t)64 << 20) ? bytes : ((size_t)64 << 20);
	block nb;
	cudaError_t e = cudaMallocFromPoolAsync((void **)&nb.ptr, sz, pool, *stream);
	if (e != cudaSuccess) {
		sb_set_error("cudaMalloc(%zu) failed: %s", sz, cudaGetErrorString(e));
		return SB_ERR_CUDA;
	}
	nb.size = sz;
	blocks.push_back(nb);
	cur_block = blocks.size() - 1;
	*out = nb.ptr;
	cur_off = bytes;
	used += bytes;
	if (used > high_water) {
		high_water = used;
	}
	return SB_OK;
}

void sb_arena::reset() {
	// coalesce: if the last call spilled over several blocks, replace them by one block of the high-water size
	if (blocks.size() > 1) {
		size_t want = high_water + (high_water >> 3);
		release(); // stream-ordered free: work already enqueued on the stream still sees the old blocks
		block nb;
		if (cudaMallocFromPoolAsync((void **)&nb.ptr, want, pool, *stream) == cudaSuccess) {
			nb.size = want;
			blocks.push_back(nb);
		} else {
			cudaGetLastError();
		}
	}
	cur_block = 0;
	cur_off = 0;
	used = 0;
}

void sb_arena::release() {
	for (auto &b : blocks) {
		cudaFreeAsync(b.ptr, *stream);
	}
	blocks.clear();
	cur_block = 0;
	cur_off = 0;
	used = 0;
}

// ---- context ------------------------------------------------------------------------------------------------
extern "C" int sb_context_create(int device, sb_context **out) {
	SB_ARG(out != nullptr, "out");
	int count = 0;
	cudaError_t e = cudaGetDeviceCount(&count);
	if (e != cudaSuccess || count == 0) {
		sb_set_error("no CUDA device: %s (libspatial_b200 has no CPU fallback)", cudaGetErrorString(e));
		cudaGetLastError();
		return SB_ERR_CUDA;
	}
	SB_ARG(device >= 0 && device < count, "device index");
	cudaDeviceProp prop;
	SB_CUDA(cudaGetDeviceProperties(&prop, device));
	if (prop.major != 10) {
		sb_set_error("device %d is sm_%d%d; this library only contains sm_100a code", device, prop.major, prop.minor);
		return SB_ERR_CUDA;
	}
	SB_CUDA(cudaSetDevice(device));
	sb_context *ctx = new sb_context();
	ctx->device = device;
	SB_CUDA(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
	ctx->stream = ctx->own_stream;
	SB_CUDA(cudaStreamCreateWithFlags(&ctx->h2d_stream, cudaStreamNonBlocking));
	SB_CUDA(cudaStreamCreateWithFlags(&ctx->d2h_stream, cudaStreamNonBlocking));
	ctx->arena.stream = &ctx->stream;
	// Device buffers (scratch arena, trees) come from a stream-ordered pool that keeps what is freed:
	// cudaMalloc/cudaFree of a few hundred MB per call cost milliseconds (measured: 3.5 -> 17 ms per index build
	// depending on the allocator state), pool allocations microseconds. The pool is PRIVATE to the context, so the
	// release threshold does not leak into the host process's default pool (DuckDB, torch, ...).
	{
		cudaMemPoolProps props = {};
		props.allocType = cudaMemAllocationTypePinned;
		props.handleTypes = cudaMemHandleTypeNone;
		props.location.type = cudaMemLocationTypeDevice;
		props.location.id = device;
		SB_CUDA(cudaMemPoolCreate(&ctx->pool, &props));
		uint64_t keep = UINT64_MAX;
		SB_CUDA(cudaMemPoolSetAttribute(ctx->pool, cudaMemPoolAttrReleaseThreshold, &keep));
		ctx->arena.pool = ctx->pool;
	}
	// mapped: the counters are written by a kernel (sbi_read_words), not by a copy engine
	SB_CUDA(cudaHostAlloc((void **)&ctx->pinned_u64, 64 * sizeof(uint64_t), cudaHostAllocMapped));
	SB_CUDA(cudaHostGetDevicePointer((void **)&ctx->pinned_u64_dev, ctx->pinned_u64, 0));
	{
		std::lock_guard<std::mutex> g(g_live_mu);
		g_live.insert(ctx);
	}
	*out = ctx;
	return SB_OK;
}

void sb_owner_free(sb_context *owner, int device, void *ptr) {
	if (!ptr) {
		return;
	}
	cudaSetDevice(device);
	std::lock_guard<std::mutex> g(g_live_mu);
	if (owner && g_live.count(owner)) {
		// ordered after everything the owning context has enqueued so far (build, probes of that context)
		cudaFreeAsync(ptr, owner->stream);
	} else {
		cudaDeviceSynchronize();
		cudaFree(ptr);
	}
}

extern "C" int sb_context_trim(sb_context *ctx) {
	SB_ARG(ctx != nullptr, "ctx");
	SB_CUDA(cudaSetDevice(ctx->device));
	SB_CUDA(cudaStreamSynchronize(ctx->stream));
	ctx->arena.release();
	SB_CUDA(cudaStreamSynchronize(ctx->stream));
	SB_CUDA(cudaMemPoolTrimTo(ctx->pool, 0));
	return SB_OK;
}

extern "C" void sb_context_destroy(sb_context *ctx) {
	if (!ctx) {
		return;
	}
	{
		std::lock_guard<std::mutex> g(g_live_mu);
		g_live.erase(ctx);
	}
	cudaSetDevice(ctx->device);
	cudaStreamSynchronize(ctx->stream);
	for (auto &kv : ctx->prof) {
		for (auto &p : kv.second.pending) {
			cudaEventDestroy(p.first);
			cudaEventDestroy(p.second);
		}
	}
	for (auto ev : ctx->event_pool) {
		cudaEventDestroy(ev);
	}
	ctx->arena.release();
	cudaStreamSynchronize(ctx->stream);
	if (ctx->pool) {
		cudaMemPoolDestroy(ctx->pool); // memory goes back to the driver (outstanding tree buffers keep theirs until freed)
	}
	cudaFreeHost(ctx->pinned_u64);
	cudaStreamDestroy(ctx->h2d_stream);
	cudaStreamDestroy(ctx->d2h_stream);
	cudaStreamDestroy(ctx->own_stream);
	delete ctx;
}

extern "C" int sb_context_set_stream(sb_context *ctx, void *cuda_stream) {
	SB_ARG(ctx != nullptr, "ctx");
	SB_CUDA(cudaStreamSynchronize(ctx->stream));
	ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
	return SB_OK;
}

extern "C" int sb_context_set_option(sb_context *ctx, const char *key, int64_t value) {
	SB_ARG(ctx != nullptr && key != nullptr, "ctx/key");
	if (strcmp(key, "probe_path") == 0) {
		SB_ARG(value >= 0 && value <= 3, "probe_path must be 0 (auto), 1 (tree walk), 2 (bucketed v2) or 3 (cell-sorted v3)");
		ctx->probe_path = (int)value;
		return SB_OK;
	}
	if (strcmp(key, "pip_path") == 0) {
		SB_ARG(value >= 0 && value <= 2, "pip_path must be 0 (auto), 1 (candidate list + refine) or 2 (polygon-centric)");
		ctx->pip_path = (int)value;
		return SB_OK;
	}
	if (strcmp(key, "str_path") == 0) {
		SB_ARG(value >= 0 && value <= 3, "str_path must be 0 (auto), 1 (bin-id sort + gather), 2 (record partition + y sort) or 3 "
		                                 "(node partition by separate kernels, the form of slices above 65535 entries)");
		ctx->str_path = (int)value;
		return SB_OK;
	}
	if (strcmp(key, "hist_sample") == 0) {
		SB_ARG(value >= 0 && value <= 1, "hist_sample must be 0 (exact bucket histogram) or 1 (sampled for large inputs)");
		ctx->hist_sample = (int)value;
		return SB_OK;
	}
	if (strcmp(key, "pipe_chunk") == 0) {
		SB_ARG(value >= 1024, "pipe_chunk must be >= 1024 rows");
		ctx->pipe_chunk = (uint64_t)value;
		return SB_OK;
	}
	if (strcmp(key, "trace") == 0) {
		ctx->trace = (int)value;
		return SB_OK;
	}
	sb_set_error("unknown option '%s'", key);
	return SB_ERR_ARG;
}

extern "C" int sb_context_sync(sb_context *ctx) {
	SB_ARG(ctx != nullptr, "ctx");
	SB_CUDA(cudaStreamSynchronize(ctx->stream));
	return SB_OK;
}

extern "C" uint64_t sb_context_launches(const sb_context *ctx) {
	return ctx ? ctx->launches : 0;
}

extern "C" int sb_context_profile(sb_context *ctx, int enable) {
	SB_ARG(ctx != nullptr, "ctx");
	SB_CUDA(cudaStreamSynchronize(ctx->stream));
	ctx->profiling = enable != 0;
	for (auto &kv : ctx->prof) {
		for (auto &p : kv.second.pending) {
			ctx->event_pool.push_back(p.first);
			ctx->event_pool.push_back(p.second);
		}
	}
	ctx->prof.clear();
	// event creation is slow on a busy device (measured: >100 ms per step of host stalls when the pool was filled
	// lazily inside the timed region), so the pool is filled here, outside any timed region
	while (enable && ctx->event_pool.size() < 2048) {
		cudaEvent_t e = nullptr;
		SB_CUDA(cudaEventCreate(&e));
		ctx->event_pool.push_back(e);
	}
	return SB_OK;
}

extern "C" int sb_context_profile_read(sb_context *ctx, const char *tag, double *ms, uint64_t *launches) {
	SB_ARG(ctx != nullptr && tag != nullptr, "ctx/tag");
	SB_CUDA(cudaStreamSynchronize(ctx->stream));
	auto it = ctx->prof.find(tag);
	double total = 0;
	uint64_t n = 0;
	if (it != ctx->prof.end()) {
		sb_prof_entry &e = it->second;
		for (auto &p : e.pending) {
			float t = 0;
			SB_CUDA(cudaEventElapsedTime(&t, p.first, p.second));
			e.ms += t;
			ctx->event_pool.push_back(p.first);
			ctx->event_pool.push_back(p.second);
		}
		e.pending.clear();
		total = e.ms;
		n = e.launches;
	}
	if (ms) {
		*ms = total;
	}
	if (launches) {
		*launches = n;
	}
	return SB_OK;
}

static cudaEvent_t prof_event(sb_context *ctx) {
	if (!ctx->event_pool.empty()) {
		cudaEvent_t e = ctx->event_pool.back();
		ctx->event_pool.pop_back();
		return e;
	}
	cudaEvent_t e = nullptr;
	cudaEventCreate(&e);
	return e;
}

sb_prof_scope::sb_prof_scope(sb_context *c, const char *t) : ctx(c) {
	if (ctx->profiling) {
		tag = t;
		a = prof_event(ctx);
		b = prof_event(ctx);
		cudaEventRecord(a, ctx->stream);
	}
}

sb_prof_scope::~sb_prof_scope() {
	if (a) {
		cudaEventRecord(b, ctx->stream);
		sb_prof_entry &e = ctx->prof[tag];
		e.launches++;
		e.pending.push_back({a, b});
	}
}
