// generate.cu -- synthetic benchmark input: ST_GeneratePoints on the device.
// Restates ST_GeneratePoints (src/spatial/modules/main/spatial_functions_table.cpp:82-99) over DuckDB's RandomEngine
// (duckdb/src/common/random_engine.cpp:50-63; pcg32 = setseq_xsh_rr_64_32, duckdb/third_party/pcg/pcg_random.hpp)
// with LCG skip-ahead so that point i of the sequential stream is produced independently:
//   x_i = min_x + ldexp(u32_{2i}, -32) * (max_x - min_x),  y_i = min_y + ldexp(u32_{2i+1}, -32) * (max_y - min_y).
// Harness utility: it produces the inputs, it is not part of the timed path.
#include "internal.cuh"

#define PCG_MULT 6364136223846793005ull
#define PCG_INC 1442695040888963407ull
#define GEN_RUN 8 // consecutive points per thread (one skip-ahead each)

__device__ __forceinline__ uint64_t pcg_advance(uint64_t state, uint64_t delta) {
	uint64_t cur_mult = PCG_MULT, cur_plus = PCG_INC, acc_mult = 1, acc_plus = 0;
	while (delta > 0) {
		if (delta & 1) {
			acc_mult *= cur_mult;
			acc_plus = acc_plus * cur_mult + cur_plus;
		}
		cur_plus = (cur_mult + 1) * cur_plus;
		cur_mult *= cur_mult;
		delta >>= 1;
	}
	return acc_mult * state + acc_plus;
}

__device__ __forceinline__ uint32_t pcg_output(uint64_t old) {
	const uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
	const uint32_t rot = (uint32_t)(old >> 59u);
	return (xs >> rot) | (xs << ((32u - rot) & 31u));
}

__global__ void __launch_bounds__(256) k_generate_points(uint64_t seed, uint64_t skip, uint64_t n, double min_x,
                                                         double min_y, double w, double h, double *__restrict__ x,
                                                         double *__restrict__ y) {
	const uint64_t runs = (n + GEN_RUN - 1) / GEN_RUN;
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	const uint64_t s0 = (seed + PCG_INC) * PCG_MULT + PCG_INC; // pcg32::seed(seed)
	for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < runs; r += stride) {
		const uint64_t first = r * GEN_RUN;
		uint64_t st = pcg_advance(s0, 2 * (skip + first));
		const uint64_t last = min(first + GEN_RUN, n);
		for (uint64_t i = first; i < last; i++) {
			const uint32_t a = pcg_output(st);
			st = st * PCG_MULT + PCG_INC;
			const uint32_t b = pcg_output(st);
			st = st * PCG_MULT + PCG_INC;
			// ldexp(u32, -32) is exact; min + u * (max - min) as two rounded operations (-fmad=false)
			x[i] = __dadd_rn(min_x, __dmul_rn((double)a * 0x1p-32, w));
			y[i] = __dadd_rn(min_y, __dmul_rn((double)b * 0x1p-32, h));
		}
	}
}

extern "C" int sb_generate_points(sb_context *ctx, sb_mem mem, uint64_t seed, uint64_t skip, uint64_t n,
                                  const double *bounds4, double *x, double *y) {
	SB_ARG(ctx != nullptr && bounds4 != nullptr, "ctx/bounds");
	SB_CUDA(cudaSetDevice(ctx->device));
	ctx->arena.reset();
	if (n == 0) {
		return SB_OK;
	}
	double *dx, *dy;
	SB_TRY(sb_out_buffer(ctx, mem, x, n, &dx));
	SB_TRY(sb_out_buffer(ctx, mem, y, n, &dy));
	k_generate_points<<<sb_grid((n + GEN_RUN - 1) / GEN_RUN, 256, 16), 256, 0, ctx->stream>>>(
	    seed, skip, n, bounds4[0], bounds4[1], bounds4[2] - bounds4[0], bounds4[3] - bounds4[1], dx, dy);
	SB_LAUNCH_CHECK(ctx);
	SB_TRY(sb_to_host(ctx, mem, x, dx, n));
	SB_TRY(sb_to_host(ctx, mem, y, dy, n));
	if (mem == SB_HOST) {
		SB_CUDA(cudaStreamSynchronize(ctx->stream));
	}
	return SB_OK;
}
