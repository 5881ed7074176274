// refine.cu -- K6: exact fp64 refinement of candidate pairs, compacting survivors in place.
// Compiled with -fmad=false: every product/difference is individually rounded like the reference's x86-64 build.
#include "internal.cuh"
#include "pip.cuh"
#include "exact.cuh"

int sbi_read_u64(sb_context *ctx, const uint64_t *dev, uint64_t *host) {
	return sbi_read_words(ctx, dev, 1, host);
}

__global__ void __launch_bounds__(256) k_pack_box2d(const double *__restrict__ x0, const double *__restrict__ y0,
                                                    const double *__restrict__ x1, const double *__restrict__ y1,
                                                    uint64_t n, double4 *__restrict__ out) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		out[i] = make_double4(x0[i], y0[i], x1[i], y1[i]);
	}
}

int sbi_pack_box2d(sb_context *ctx, const double *x0, const double *y0, const double *x1, const double *y1, uint64_t n,
                   double *out4) {
	if (n == 0) {
		return SB_OK;
	}
	k_pack_box2d<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(x0, y0, x1, y1, n, reinterpret_cast<double4 *>(out4));
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

// Order-preserving in-place compaction of a pair list is done in two kernels: flags + per-block counts, then a
// decoupled look-back is overkill here -- survivors are appended through a warp-aggregated atomic cursor into a
// scratch copy and copied back (pair order is not part of the contract; pair SETS are).
template <class PRED>
__global__ void __launch_bounds__(256) k_refine(PRED pred, const uint32_t *__restrict__ pa,
                                                const uint32_t *__restrict__ pb, uint64_t n, uint32_t *__restrict__ oa,
                                                uint32_t *__restrict__ ob, double *__restrict__ od,
                                                unsigned long long *__restrict__ counter) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	const uint64_t n_round = (n + 31) & ~(uint64_t)31;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
		bool keep = false;
		uint32_t a = 0, b = 0;
		double dist = 0.0;
		if (i < n) {
			a = pa[i];
			b = pb[i];
			keep = pred(a, b, dist);
		}
		const unsigned m = __ballot_sync(0xFFFFFFFFu, keep);
		if (m) {
			unsigned long long base = 0;
			const int leader = __ffs(m) - 1;
			if ((int)sb_lane() == leader) {
				base = atomicAdd(counter, (unsigned long long)__popc(m));
			}
			base = __shfl_sync(0xFFFFFFFFu, base, leader);
			if (keep) {
				const unsigned long long o = base + __popc(m & sb_lanemask_lt());
				oa[o] = a;
				ob[o] = b;
				if (od) {
					od[o] = dist;
				}
			}
		}
	}
}

struct pred_point_box {
	const double *px, *py, *x0, *y0, *x1, *y1;
	int strict;
	__device__ bool operator()(uint32_t a, uint32_t b, double &) const {
		const double x = px[a], y = py[a];
		const double bx0 = x0[b], by0 = y0[b], bx1 = x1[b], by1 = y1[b];
		return strict ? (bx0 < x && x < bx1 && by0 < y && y < by1) : (bx0 <= x && x <= bx1 && by0 <= y && y <= by1);
	}
};

// ST_Contains::Operation, spatial_functions_scalar.cpp:1938-2051 -- literal control flow, plain fp64 (no FMA):
// an on-edge hit only leaves the coordinate loop; the ring is then judged by the winding number so far.
struct pred_point_polygon {
	const double *px, *py, *vx, *vy;
	const uint64_t *ring_off, *poly_off;
	__device__ bool operator()(uint32_t a, uint32_t p, double &) const {
		return sb_pip_native(px[a], py[a], vx, vy, ring_off, poly_off[p], poly_off[p + 1]);
	}
};

// GEOMETRY-argument predicates (the GEOS route, exact.cuh): location of the point relative to the polygon with exact
// orientation signs; intersects <=> not exterior, within / contains <=> interior.
struct pred_point_polygon_exact {
	const double *px, *py, *vx, *vy;
	const uint64_t *ring_off, *poly_off;
	int pred;
	__device__ bool operator()(uint32_t a, uint32_t p, double &) const {
		const int loc = sb_point_polygon_locate(px[a], py[a], vx, vy, ring_off, poly_off[p], poly_off[p + 1]);
		return pred == SB_PRED_INTERSECTS ? loc != 0 : loc == 2;
	}
};

// ST_Distance(POINT_2D, POINT_2D) :2189-2214: sqrt(pow(dx,2) + pow(dy,2)); pow(v,2) == v*v exactly (glibc),
// sqrt is IEEE correctly rounded on both sides.
__device__ __forceinline__ double dist_pp(double ax, double ay, double bx, double by) {
	const double dx = __dsub_rn(ax, bx), dy = __dsub_rn(ay, by);
	return __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
}

struct pred_dwithin {
	const double *ax, *ay, *bx, *by;
	double r;
	__device__ bool operator()(uint32_t a, uint32_t b, double &d) const {
		d = dist_pp(ax[a], ay[a], bx[b], by[b]);
		return d <= r;
	}
};

template <class PRED>
static int run_refine(sb_context *ctx, const char *tag, PRED pred, uint32_t *pa, uint32_t *pb, double *od, uint64_t n,
                      uint64_t *count) {
	*count = 0;
	if (n == 0) {
		return SB_OK;
	}
	uint32_t *oa = nullptr, *ob = nullptr;
	double *tmp_d = nullptr;
	unsigned long long *d_counter = nullptr;
	SB_TRY(ctx->scratch_t(n, &oa));
	SB_TRY(ctx->scratch_t(n, &ob));
	if (od) {
		SB_TRY(ctx->scratch_t(n, &tmp_d));
	}
	SB_TRY(ctx->scratch_t(1, &d_counter));
	SB_TRY(sbi_zero(ctx, d_counter, sizeof(*d_counter)));
	{
		sb_prof_scope ps(ctx, tag);
		k_refine<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(pred, pa, pb, n, oa, ob, tmp_d, d_counter);
		SB_LAUNCH_CHECK(ctx);
	}
	SB_TRY(sbi_read_u64(ctx, reinterpret_cast<const uint64_t *>(d_counter), count));
	if (*count) {
		SB_CUDA(cudaMemcpyAsync(pa, oa, *count * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
		SB_CUDA(cudaMemcpyAsync(pb, ob, *count * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
		if (od) {
			SB_CUDA(cudaMemcpyAsync(od, tmp_d, *count * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
		}
	}
	return SB_OK;
}

int sbi_refine_point_box(sb_context *ctx, int pred, const double *px, const double *py, const double *x0,
                         const double *y0, const double *x1, const double *y1, uint32_t *pp, uint32_t *pb, uint64_t n,
                         uint64_t *count) {
	pred_point_box p = {px, py, x0, y0, x1, y1, pred == SB_PRED_WITHIN};
	return run_refine(ctx, "refine_box", p, pp, pb, nullptr, n, count);
}

int sbi_refine_point_polygon(sb_context *ctx, const double *px, const double *py, const double *vx, const double *vy,
                             const uint64_t *ring_off, const uint64_t *poly_off, uint32_t *pp, uint32_t *pb, uint64_t n,
                             uint64_t *count) {
	pred_point_polygon p = {px, py, vx, vy, ring_off, poly_off};
	return run_refine(ctx, "refine_pip", p, pp, pb, nullptr, n, count);
}

int sbi_refine_point_polygon_exact(sb_context *ctx, int pred, const double *px, const double *py, const double *vx,
                                   const double *vy, const uint64_t *ring_off, const uint64_t *poly_off, uint32_t *pp,
                                   uint32_t *pb, uint64_t n, uint64_t *count) {
	pred_point_polygon_exact p = {px, py, vx, vy, ring_off, poly_off, pred};
	return run_refine(ctx, "refine_pip_exact", p, pp, pb, nullptr, n, count);
}

int sbi_refine_dwithin(sb_context *ctx, const double *ax, const double *ay, const double *bx, const double *by,
                       double r, uint32_t *pa, uint32_t *pb, double *out_d, uint64_t n, uint64_t *count) {
	pred_dwithin p = {ax, ay, bx, by, r};
	return run_refine(ctx, "refine_dist", p, pa, pb, out_d, n, count);
}

__global__ void __launch_bounds__(256) k_distance(const double *__restrict__ ax, const double *__restrict__ ay,
                                                  const double *__restrict__ bx, const double *__restrict__ by,
                                                  uint64_t n, double *__restrict__ out) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		out[i] = dist_pp(ax[i], ay[i], bx[i], by[i]);
	}
}

int sbi_distance_points(sb_context *ctx, const double *ax, const double *ay, const double *bx, const double *by,
                        uint64_t n, double *out) {
	if (n == 0) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "refine_dist");
	k_distance<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(ax, ay, bx, by, n, out);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

// conservative float boxes of points grown by r (radius-join build side): any b with dist(a,b) <= r as computed
// by dist_pp has |ax-bx| <= r(1+2^-52) and likewise in y, so a's float box intersects this one.
__global__ void __launch_bounds__(256) k_expand_points(const double *__restrict__ x, const double *__restrict__ y,
                                                       uint64_t n, double r, float4 *__restrict__ out) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		const double X = x[i], Y = y[i];
		const double ex = (fabs(X) + r) * 0x1p-48, ey = (fabs(Y) + r) * 0x1p-48;
		out[i] = make_float4(sb_f64_to_f32_down(X - r - ex), sb_f64_to_f32_down(Y - r - ey),
		                     sb_f64_to_f32_up(X + r + ex), sb_f64_to_f32_up(Y + r + ey));
	}
}

int sbi_expand_points(sb_context *ctx, const double *x, const double *y, uint64_t n, double r, float *out_box4) {
	if (n == 0) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "bbox");
	k_expand_points<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(x, y, n, r, reinterpret_cast<float4 *>(out_box4));
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

// ---- A11: LEFT / RIGHT / FULL join bookkeeping --------------------------------------------------------------------
// The reference marks matched probe rows in left_outer_marker[] (spatial_join_physical.cpp:867-876) and matched build
// rows in a bool column of the build tuple (:878-891); unmatched probe rows are emitted NULL-padded per input chunk
// (:917-938), unmatched build rows by the source scan after all probes (:1022-1068). Here: flag arrays over whole
// columns + an order-preserving selection of the rows whose flag is 0.
__global__ void __launch_bounds__(256) k_mark_matched(const uint32_t *__restrict__ pp, const uint32_t *__restrict__ pb,
                                                      uint64_t n, uint32_t probe_base, uint8_t *__restrict__ pm,
                                                      uint64_t n_probe, uint8_t *__restrict__ bm, uint64_t n_build) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		if (pm) {
			const uint64_t p = (uint64_t)pp[i] - probe_base;
			if (p < n_probe) {
				pm[p] = 1;
			}
		}
		if (bm) {
			const uint64_t b = pb[i];
			if (b < n_build) {
				bm[b] = 1;
			}
		}
	}
}

int sbi_mark_matched(sb_context *ctx, const uint32_t *pp, const uint32_t *pb, uint64_t n, uint32_t probe_base,
                     uint8_t *pm, uint64_t n_probe, uint8_t *bm, uint64_t n_build) {
	if (n == 0 || (!pm && !bm)) {
		return SB_OK;
	}
	k_mark_matched<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(pp, pb, n, probe_base, pm, n_probe, bm, n_build);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

// Order-independent checksum of a pair list restricted to probe ids in [lo, hi): count and the wrapping 64-bit sum of
// mix(probe, build). Used to compare a join result of any size with a reference pair set without moving or sorting
// it (bench.py's in-run parity check, tests). mix() is restated in oracle/ref_harness.cpp and tests.
__device__ __forceinline__ unsigned long long sb_pair_mix(uint32_t p, uint32_t b) {
	unsigned long long k = (unsigned long long)p * 2654435761ull + (unsigned long long)b * 2246822519ull;
	k ^= k >> 29;
	k *= 0xBF58476D1CE4E5B9ull;
	k ^= k >> 32;
	return k;
}
__global__ void __launch_bounds__(256) k_pairs_checksum(const uint32_t *__restrict__ pp, const uint32_t *__restrict__ pb,
                                                        uint64_t n, uint32_t lo, uint32_t hi,
                                                        unsigned long long *__restrict__ out2) {
	unsigned long long cnt = 0, sum = 0;
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		const uint32_t p = pp[i];
		if (p >= lo && p < hi) {
			cnt++;
			sum += sb_pair_mix(p, pb[i]);
		}
	}
#pragma unroll
	for (int s = 16; s > 0; s >>= 1) {
		cnt += __shfl_xor_sync(0xFFFFFFFFu, cnt, s);
		sum += __shfl_xor_sync(0xFFFFFFFFu, sum, s);
	}
	if (sb_lane() == 0 && cnt) {
		atomicAdd(out2, cnt);
		atomicAdd(out2 + 1, sum);
	}
}

int sbi_pairs_checksum(sb_context *ctx, const uint32_t *pp, const uint32_t *pb, uint64_t n, uint32_t lo, uint32_t hi,
                       unsigned long long *d_out2) {
	SB_TRY(sbi_zero(ctx, d_out2, 2 * sizeof(unsigned long long)));
	if (n) {
		k_pairs_checksum<<<sb_grid(n, 256 * 8, 8), 256, 0, ctx->stream>>>(pp, pb, n, lo, hi, d_out2);
		SB_LAUNCH_CHECK(ctx);
	}
	return SB_OK;
}

#define UM_TILE 4096
__global__ void __launch_bounds__(256) k_unmatched_count(const uint8_t *__restrict__ flags, uint64_t n,
                                                         uint32_t *__restrict__ tile_cnt) {
	__shared__ uint32_t s_cnt;
	for (uint64_t tile = blockIdx.x; tile * UM_TILE < n; tile += gridDim.x) {
		if (threadIdx.x == 0) {
			s_cnt = 0;
		}
		__syncthreads();
		uint32_t c = 0;
		for (uint32_t k = threadIdx.x; k < UM_TILE; k += 256) {
			const uint64_t i = tile * UM_TILE + k;
			c += (i < n && flags[i] == 0) ? 1u : 0u;
		}
		c = __reduce_add_sync(0xFFFFFFFFu, c);
		if (sb_lane() == 0 && c) {
			atomicAdd(&s_cnt, c);
		}
		__syncthreads();
		if (threadIdx.x == 0) {
			tile_cnt[tile] = s_cnt;
		}
		__syncthreads();
	}
}

// exclusive scan of the tile counts by one CTA (tiles <= n / 4096: a few 10^4 at the BASELINE sizes)
__global__ void __launch_bounds__(1024) k_unmatched_scan(uint32_t *__restrict__ tile_cnt, uint64_t tiles,
                                                         unsigned long long *__restrict__ total) {
	__shared__ uint32_t s_warp[32];
	__shared__ unsigned long long s_run;
	if (threadIdx.x == 0) {
		s_run = 0;
	}
	__syncthreads();
	for (uint64_t t0 = 0; t0 < tiles; t0 += 1024) {
		const uint64_t t = t0 + threadIdx.x;
		const uint32_t c = t < tiles ? tile_cnt[t] : 0u;
		const unsigned lane = sb_lane(), warp = threadIdx.x >> 5;
		uint32_t inc = c;
#pragma unroll
		for (int s = 1; s < 32; s <<= 1) {
			const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, s);
			if (lane >= (unsigned)s) {
				inc += o;
			}
		}
		if (lane == 31) {
			s_warp[warp] = inc;
		}
		__syncthreads();
		uint32_t wb = 0;
		for (unsigned w = 0; w < warp; w++) {
			wb += s_warp[w];
		}
		const unsigned long long run = s_run;
		if (t < tiles) {
			// offsets fit 32 bits: at most 2^32 rows per column (row ids are uint32)
			tile_cnt[t] = (uint32_t)(run + wb + inc - c);
		}
		__syncthreads();
		if (threadIdx.x == 1023) {
			s_run = run + wb + inc;
		}
		__syncthreads();
	}
	if (threadIdx.x == 0) {
		*total = s_run;
	}
}

__global__ void __launch_bounds__(256) k_unmatched_write(const uint8_t *__restrict__ flags, uint64_t n,
                                                         const uint32_t *__restrict__ tile_off, uint32_t row_base,
                                                         uint32_t *__restrict__ out, uint64_t capacity) {
	__shared__ uint32_t s_warp[8];
	for (uint64_t tile = blockIdx.x; tile * UM_TILE < n; tile += gridDim.x) {
		uint32_t run = tile_off[tile];
		// 16 rounds of 256 consecutive rows keep the output ascending
		for (uint32_t k0 = 0; k0 < UM_TILE; k0 += 256) {
			const uint64_t i = tile * UM_TILE + k0 + threadIdx.x;
			const bool un = i < n && flags[i] == 0;
			const uint32_t m = __ballot_sync(0xFFFFFFFFu, un);
			const unsigned lane = sb_lane(), warp = threadIdx.x >> 5;
			if (lane == 0) {
				s_warp[warp] = __popc(m);
			}
			__syncthreads();
			uint32_t wb = 0, tot = 0;
#pragma unroll
			for (unsigned w = 0; w < 8; w++) {
				const uint32_t c = s_warp[w];
				wb += w < warp ? c : 0u;
				tot += c;
			}
			if (un) {
				const uint64_t o = (uint64_t)run + wb + __popc(m & sb_lanemask_lt());
				if (o < capacity) {
					out[o] = (uint32_t)i + row_base;
				}
			}
			run += tot;
			__syncthreads();
		}
	}
}

int sbi_unmatched_rows(sb_context *ctx, const uint8_t *flags, uint64_t n, uint32_t row_base, uint32_t *out,
                       uint64_t capacity, uint64_t *count) {
	*count = 0;
	if (n == 0) {
		return SB_OK;
	}
	const uint64_t tiles = (n + UM_TILE - 1) / UM_TILE;
	uint32_t *d_tiles = nullptr;
	unsigned long long *d_total = nullptr;
	SB_TRY(ctx->scratch_t(tiles, &d_tiles));
	SB_TRY(ctx->scratch_t(1, &d_total));
	k_unmatched_count<<<sb_grid(tiles, 1, 16), 256, 0, ctx->stream>>>(flags, n, d_tiles);
	SB_LAUNCH_CHECK(ctx);
	k_unmatched_scan<<<1, 1024, 0, ctx->stream>>>(d_tiles, tiles, d_total);
	SB_LAUNCH_CHECK(ctx);
	k_unmatched_write<<<sb_grid(tiles, 1, 16), 256, 0, ctx->stream>>>(flags, n, d_tiles, row_base, out, capacity);
	SB_LAUNCH_CHECK(ctx);
	SB_TRY(sbi_read_u64(ctx, reinterpret_cast<const uint64_t *>(d_total), count));
	if (*count > capacity) {
		sb_set_error("unmatched rows: %llu, capacity %llu", (unsigned long long)*count, (unsigned long long)capacity);
		return SB_ERR_CAPACITY;
	}
	return SB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Native-type scalar predicates next to the join path
// ---------------------------------------------------------------------------------------------------------------
// ST_Intersects(BOX_2D, BOX_2D), spatial_functions_scalar.cpp:5407-5416 (same test as sgl::box_xy::intersects,
// sgl.hpp:93-95): NOT (a.min_x > b.max_x OR a.max_x < b.min_x OR a.min_y > b.max_y OR a.max_y < b.min_y); NaNs make
// every comparison false, i.e. "intersects", exactly as on the CPU. Rows whose validity byte is 0 give false
// (ST_Intersects_Extent's "no extent" rows, :5456-5481).
__global__ void __launch_bounds__(256)
    k_box2d_intersects(const double *__restrict__ a0, const double *__restrict__ a1, const double *__restrict__ a2,
                       const double *__restrict__ a3, const double *__restrict__ b0, const double *__restrict__ b1,
                       const double *__restrict__ b2, const double *__restrict__ b3, uint32_t stride,
                       const uint8_t *__restrict__ va, const uint8_t *__restrict__ vb, uint64_t n,
                       uint8_t *__restrict__ out) {
	const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += step) {
		const uint64_t k = i * stride;
		bool r = !(a0[k] > b2[k] || a2[k] < b0[k] || a1[k] > b3[k] || a3[k] < b1[k]);
		if ((va && !va[i]) || (vb && !vb[i])) {
			r = false;
		}
		out[i] = r ? 1 : 0;
	}
}

int sbi_box2d_intersects(sb_context *ctx, const double *const a[4], const double *const b[4], uint32_t stride,
                         const uint8_t *va, const uint8_t *vb, uint64_t n, uint8_t *out) {
	if (n == 0) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "box_intersects");
	k_box2d_intersects<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(a[0], a[1], a[2], a[3], b[0], b[1], b[2], b[3],
	                                                                 stride, va, vb, n, out);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

// ST_Distance(POINT_2D, LINESTRING_2D), spatial_functions_scalar.cpp:2219-2258, with ClosestPointOnSegment :2158-2176
// (ApproxEqualTo: |dx| < 1e-6 AND |dy| < 1e-6, vertex.hpp:50-52) and DistanceToSegmentSquared :2178-2184. Same operation
// order, no FMA contraction (the library is built with -fmad=false). The reference keeps the minimum of the squared
// segment distances with `<` (NaN never replaces it) and stops at 0: a minimum over non-negative values, so the
// lanes of a warp can take every 32nd segment and combine with the same `<` -- the result is the same double.
// One warp per row.
__device__ __forceinline__ double seg_dist2(double px, double py, double x1, double y1, double x2, double y2) {
	double cx, cy;
	if (fabs(x1 - x2) < 1e-6 && fabs(y1 - y2) < 1e-6) {
		cx = x1, cy = y1;
	} else {
		const double n1 = (px - x1) * (x2 - x1) + (py - y1) * (y2 - y1);
		const double n2 = (x2 - x1) * (x2 - x1) + (y2 - y1) * (y2 - y1);
		const double r = n1 / n2;
		if (r <= 0) {
			cx = x1, cy = y1;
		} else if (r >= 1) {
			cx = x2, cy = y2;
		} else {
			cx = x1 + r * (x2 - x1), cy = y1 + r * (y2 - y1);
		}
	}
	const double dx = px - cx, dy = py - cy;
	return dx * dx + dy * dy;
}

__global__ void __launch_bounds__(256)
    k_dist_point_line(const double *__restrict__ px, const double *__restrict__ py, const uint64_t *__restrict__ lists,
                      const double *__restrict__ vx, const double *__restrict__ vy, uint64_t n,
                      double *__restrict__ out) {
	const unsigned lane = sb_lane();
	const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
	for (uint64_t i = (((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5); i < n; i += warps) {
		const uint64_t off = lists[2 * i], len = lists[2 * i + 1];
		const double x = px[i], y = py[i];
		double best = DBL_MAX;
		// the reference's `length - 1` wraps for an empty list (it then reads past the list); here an empty or
		// one-vertex line has no segment and yields sqrt(DBL_MAX), the value the loop starts from
		for (uint64_t j = lane; j + 1 < len; j += 32) {
			const double d = seg_dist2(x, y, vx[off + j], vy[off + j], vx[off + j + 1], vy[off + j + 1]);
			if (d < best) {
				best = d;
			}
		}
#pragma unroll
		for (int s = 16; s > 0; s >>= 1) {
			const double o = __shfl_xor_sync(0xFFFFFFFFu, best, s);
			if (o < best) {
				best = o;
			}
		}
		if (lane == 0) {
			out[i] = sqrt(best);
		}
	}
}

int sbi_distance_point_linestring(sb_context *ctx, const double *px, const double *py, const uint64_t *lists,
                                  const double *vx, const double *vy, uint64_t n, double *out) {
	if (n == 0) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "refine_dist");
	k_dist_point_line<<<sb_grid(n * 32, 256, 16), 256, 0, ctx->stream>>>(px, py, lists, vx, vy, n, out);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}
