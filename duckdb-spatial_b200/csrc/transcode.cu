// transcode.cu -- serialized GEOMETRY blobs -> the coordinate columns the join kernels take.
//
//   sbi_points_from_blobs    GEOMETRY -> POINT_2D   (PointCasts::FromGeometryCast, spatial_functions_cast.cpp:247-270)
//   sbi_polygons_from_blobs  GEOMETRY -> POLYGON_2D (PolygonCasts::FromGeometryCast, spatial_functions_cast.cpp:473-537)
//
// Both are what a DuckDB operator needs to hand a GEOMETRY vector to sb_join_points_boxes / sb_join_points_polygons
// without deserialising on the host. The reference throws on the first bad row (wrong type, empty point, malformed
// blob); a kernel cannot throw, so every row gets a status, bad rows produce NaN / zero rings, and the call returns
// SB_ERR_GEOMETRY / SB_ERR_EMPTY after the whole column is done.
// HBM-streaming: every blob byte is read once, every coordinate written once (the polygon path reads the 8 + 4r header
// bytes of a blob twice: count, then fill).
#include "internal.cuh"
#include "blob.cuh"

// header checks shared by both casts. returns 0 or SB_ROW_*; on success `pos` is the offset of the first part header,
// `vsize` the vertex size in bytes
__device__ __forceinline__ int tc_header(const uint8_t *b, uint64_t len, uint32_t want_type, uint64_t &pos,
                                         uint32_t &vsize, uint32_t &count) {
	if (len < 8) {
		return SB_ROW_TRUNCATED;
	}
	const uint8_t type = b[0], flags = b[1];
	if (flags & 0xC0) {
		return SB_ROW_BAD_VERSION;
	}
	if (type != want_type) {
		return type <= 6 ? SB_ROW_WRONG_TYPE : SB_ROW_TRUNCATED;
	}
	const uint32_t dims = 2 + ((flags & 1) ? 1 : 0) + ((flags & 2) ? 1 : 0);
	vsize = 8 * dims;
	pos = 8 + ((flags & 0x04) ? 8 * dims : 0);
	if (pos + 8 > len) {
		return SB_ROW_TRUNCATED;
	}
	count = ld_u32(b + pos + 4);
	pos += 8;
	return 0;
}

// ---- points: one thread per row (32-byte blobs: consecutive rows share sectors) -----------------------------------
__global__ void __launch_bounds__(256) k_tc_points(const uint8_t *__restrict__ data, const uint64_t *__restrict__ off,
                                                   uint64_t n, double *__restrict__ ox, double *__restrict__ oy,
                                                   int32_t *__restrict__ status, int *__restrict__ any_bad) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		const uint64_t o = off[i], len = off[i + 1] - o;
		const uint8_t *b = data + o;
		uint64_t pos;
		uint32_t vsize, count;
		int r = tc_header(b, len, 0, pos, vsize, count);
		double x = __longlong_as_double(0x7FF8000000000000ll), y = x;
		if (r == 0) {
			if (count == 0) {
				r = SB_ROW_EMPTY;
			} else if (pos + vsize > len) {
				r = SB_ROW_TRUNCATED;
			} else {
				x = ld_f64(b + pos);
				y = ld_f64(b + pos + 8);
			}
		}
		ox[i] = x;
		oy[i] = y;
		if (status) {
			status[i] = r;
		}
		if (r != 0) {
			atomicMax(any_bad, r == SB_ROW_EMPTY ? 1 : 2);
		}
	}
}

int sbi_points_from_blobs(sb_context *ctx, const uint8_t *data, const uint64_t *off, uint64_t n, double *ox, double *oy,
                          int32_t *status, int *any_bad) {
	if (n == 0) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "transcode");
	k_tc_points<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(data, off, n, ox, oy, status, any_bad);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

// ---- polygons ---------------------------------------------------------------------------------------------------
// pass 1: rings and vertices per row (one thread per row; the ring-size array is the only part of the payload read),
// tile sums. pass 2: one CTA scans the tile sums. pass 3: per-row offsets. pass 4: one warp per row copies the rings.
#define TC_TILE 1024u

// rings / vertices of one POLYGON blob, 0 / 0 for a bad row
__device__ __forceinline__ int tc_poly_counts(const uint8_t *b, uint64_t len, uint32_t &rings, uint64_t &verts,
                                              uint64_t &vpos, uint32_t &vsize) {
	uint64_t pos;
	uint32_t count;
	rings = 0;
	verts = 0;
	int r = tc_header(b, len, 2, pos, vsize, count);
	if (r != 0) {
		return r;
	}
	vpos = pos + (uint64_t)count * 4 + ((count & 1) ? 4 : 0);
	if (vpos > len) {
		return SB_ROW_TRUNCATED;
	}
	uint64_t v = 0;
	for (uint32_t k = 0; k < count; k++) {
		v += ld_u32(b + pos + 4ull * k);
	}
	if (vpos + v * vsize > len) {
		return SB_ROW_TRUNCATED;
	}
	rings = count;
	verts = v;
	return 0;
}

__global__ void __launch_bounds__(TC_TILE) k_tc_poly_count(const uint8_t *__restrict__ data,
                                                           const uint64_t *__restrict__ off, uint64_t n,
                                                           uint32_t *__restrict__ row_rings,
                                                           unsigned long long *__restrict__ row_verts,
                                                           unsigned long long *__restrict__ tile_sums,
                                                           int32_t *__restrict__ status, int *__restrict__ any_bad) {
	__shared__ unsigned long long s_r[32], s_v[32];
	const uint64_t i = (uint64_t)blockIdx.x * TC_TILE + threadIdx.x;
	uint32_t rings = 0;
	uint64_t verts = 0;
	if (i < n) {
		uint64_t vpos;
		uint32_t vsize;
		const uint64_t o = off[i];
		const int r = tc_poly_counts(data + o, off[i + 1] - o, rings, verts, vpos, vsize);
		row_rings[i] = rings;
		row_verts[i] = verts;
		if (status) {
			status[i] = r;
		}
		if (r != 0) {
			atomicMax(any_bad, 2);
		}
	}
	unsigned long long r64 = rings, v64 = verts;
#pragma unroll
	for (int s = 16; s > 0; s >>= 1) {
		r64 += __shfl_xor_sync(0xFFFFFFFFu, r64, s);
		v64 += __shfl_xor_sync(0xFFFFFFFFu, v64, s);
	}
	if (sb_lane() == 0) {
		s_r[threadIdx.x >> 5] = r64;
		s_v[threadIdx.x >> 5] = v64;
	}
	__syncthreads();
	if (threadIdx.x < 32) {
		r64 = s_r[threadIdx.x];
		v64 = s_v[threadIdx.x];
#pragma unroll
		for (int s = 16; s > 0; s >>= 1) {
			r64 += __shfl_xor_sync(0xFFFFFFFFu, r64, s);
			v64 += __shfl_xor_sync(0xFFFFFFFFu, v64, s);
		}
		if (threadIdx.x == 0) {
			tile_sums[2ull * blockIdx.x] = r64;
			tile_sums[2ull * blockIdx.x + 1] = v64;
		}
	}
}

// exclusive scan of the (rings, vertices) tile sums by one CTA; totals[0..1] = grand totals
__global__ void __launch_bounds__(1024) k_tc_scan_tiles(unsigned long long *__restrict__ tile_sums, uint64_t tiles,
                                                        unsigned long long *__restrict__ totals) {
	__shared__ unsigned long long s_w[2][32];
	__shared__ unsigned long long s_run[2];
	if (threadIdx.x < 2) {
		s_run[threadIdx.x] = 0;
	}
	__syncthreads();
	const unsigned lane = sb_lane(), warp = threadIdx.x >> 5;
	for (uint64_t t0 = 0; t0 < tiles; t0 += 1024) {
		const uint64_t t = t0 + threadIdx.x;
		unsigned long long c[2], inc[2];
#pragma unroll
		for (int q = 0; q < 2; q++) {
			c[q] = t < tiles ? tile_sums[2 * t + q] : 0ull;
			inc[q] = c[q];
#pragma unroll
			for (int s = 1; s < 32; s <<= 1) {
				const unsigned long long o = __shfl_up_sync(0xFFFFFFFFu, inc[q], s);
				if (lane >= (unsigned)s) {
					inc[q] += o;
				}
			}
			if (lane == 31) {
				s_w[q][warp] = inc[q];
			}
		}
		__syncthreads();
#pragma unroll
		for (int q = 0; q < 2; q++) {
			unsigned long long wb = 0;
			for (unsigned w = 0; w < warp; w++) {
				wb += s_w[q][w];
			}
			if (t < tiles) {
				tile_sums[2 * t + q] = s_run[q] + wb + inc[q] - c[q];
			}
			inc[q] += wb; // inclusive over the whole chunk for the last thread
		}
		__syncthreads();
		if (threadIdx.x == 1023) {
			s_run[0] += inc[0];
			s_run[1] += inc[1];
		}
		__syncthreads();
	}
	if (threadIdx.x < 2) {
		totals[threadIdx.x] = s_run[threadIdx.x];
	}
}

// per-row exclusive offsets inside a tile + the tile base: poly_ring_off[i] (output), row_verts[i] -> first vertex
__global__ void __launch_bounds__(TC_TILE) k_tc_poly_offsets(const uint32_t *__restrict__ row_rings,
                                                             unsigned long long *__restrict__ row_verts, uint64_t n,
                                                             const unsigned long long *__restrict__ tile_sums,
                                                             const unsigned long long *__restrict__ totals,
                                                             uint64_t *__restrict__ poly_ring_off) {
	__shared__ unsigned long long s_r[32], s_v[32];
	const uint64_t i = (uint64_t)blockIdx.x * TC_TILE + threadIdx.x;
	const unsigned lane = sb_lane(), warp = threadIdx.x >> 5;
	const unsigned long long cr = i < n ? row_rings[i] : 0ull, cv = i < n ? row_verts[i] : 0ull;
	unsigned long long ir = cr, iv = cv;
#pragma unroll
	for (int s = 1; s < 32; s <<= 1) {
		const unsigned long long a = __shfl_up_sync(0xFFFFFFFFu, ir, s), b = __shfl_up_sync(0xFFFFFFFFu, iv, s);
		if (lane >= (unsigned)s) {
			ir += a;
			iv += b;
		}
	}
	if (lane == 31) {
		s_r[warp] = ir;
		s_v[warp] = iv;
	}
	__syncthreads();
	unsigned long long br = tile_sums[2ull * blockIdx.x], bv = tile_sums[2ull * blockIdx.x + 1];
	for (unsigned w = 0; w < warp; w++) {
		br += s_r[w];
		bv += s_v[w];
	}
	if (i < n) {
		poly_ring_off[i] = br + ir - cr;
		row_verts[i] = bv + iv - cv;
	}
	if (i == 0) {
		poly_ring_off[n] = totals[0];
	}
}

// one warp per row: ring offsets (warp scan over the ring sizes, 32 rings at a time), then the vertices, lanes strided
__global__ void __launch_bounds__(256) k_tc_poly_fill(const uint8_t *__restrict__ data, const uint64_t *__restrict__ off,
                                                      uint64_t n, const uint32_t *__restrict__ row_rings,
                                                      const unsigned long long *__restrict__ row_vbase,
                                                      const uint64_t *__restrict__ poly_ring_off,
                                                      const unsigned long long *__restrict__ totals,
                                                      uint64_t *__restrict__ ring_vert_off, double *__restrict__ vx,
                                                      double *__restrict__ vy) {
	const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
	const unsigned lane = sb_lane();
	if (blockIdx.x == 0 && threadIdx.x == 0) {
		ring_vert_off[totals[0]] = totals[1];
	}
	for (uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
		const uint32_t rings = row_rings[i];
		if (rings == 0) {
			continue; // empty polygon or bad row
		}
		const uint64_t o = off[i];
		const uint8_t *b = data + o;
		uint64_t pos, vpos;
		uint32_t vsize, count;
		tc_header(b, off[i + 1] - o, 2, pos, vsize, count); // validated in pass 1
		vpos = pos + (uint64_t)count * 4 + ((count & 1) ? 4 : 0);
		const uint64_t rbase = poly_ring_off[i];
		uint64_t vbase = row_vbase[i];
		const uint8_t *src = b + vpos;
		for (uint32_t r0 = 0; r0 < rings; r0 += 32) {
			const uint32_t r = r0 + lane;
			const unsigned long long c = r < rings ? ld_u32(b + pos + 4ull * r) : 0u;
			unsigned long long inc = c;
#pragma unroll
			for (int s = 1; s < 32; s <<= 1) {
				const unsigned long long a = __shfl_up_sync(0xFFFFFFFFu, inc, s);
				if (lane >= (unsigned)s) {
					inc += a;
				}
			}
			if (r < rings) {
				ring_vert_off[rbase + r] = vbase + inc - c;
			}
			const unsigned long long chunk = __shfl_sync(0xFFFFFFFFu, inc, 31); // vertices of these <= 32 rings
			for (unsigned long long k = lane; k < chunk; k += 32) {
				vx[vbase + k] = ld_f64(src + k * vsize);
				vy[vbase + k] = ld_f64(src + k * vsize + 8);
			}
			vbase += chunk;
			src += chunk * vsize;
		}
	}
}

int sbi_polygons_count(sb_context *ctx, const uint8_t *data, const uint64_t *off, uint64_t n, sb_poly_transcode *T,
                       int32_t *status, int *any_bad, uint64_t *poly_ring_off, uint64_t *n_rings, uint64_t *n_verts) {
	sb_prof_scope ps(ctx, "transcode");
	const uint64_t tiles = (n + TC_TILE - 1) / TC_TILE;
	SB_ARG(tiles < 0x7FFFFFFFull, "too many rows");
	SB_TRY(ctx->scratch_t(n, &T->row_rings));
	SB_TRY(ctx->scratch_t(n, &T->row_verts));
	SB_TRY(ctx->scratch_t(2 * tiles, &T->tile_sums));
	SB_TRY(ctx->scratch_t(2, &T->totals));
	k_tc_poly_count<<<(unsigned)tiles, TC_TILE, 0, ctx->stream>>>(data, off, n, T->row_rings, T->row_verts, T->tile_sums,
	                                                             status, any_bad);
	SB_LAUNCH_CHECK(ctx);
	k_tc_scan_tiles<<<1, 1024, 0, ctx->stream>>>(T->tile_sums, tiles, T->totals);
	SB_LAUNCH_CHECK(ctx);
	k_tc_poly_offsets<<<(unsigned)tiles, TC_TILE, 0, ctx->stream>>>(T->row_rings, T->row_verts, n, T->tile_sums, T->totals,
	                                                               poly_ring_off);
	SB_LAUNCH_CHECK(ctx);
	uint64_t tot[2];
	SB_TRY(sbi_read_words(ctx, T->totals, 2, tot));
	*n_rings = tot[0];
	*n_verts = tot[1];
	return SB_OK;
}

int sbi_polygons_fill(sb_context *ctx, const uint8_t *data, const uint64_t *off, uint64_t n, const sb_poly_transcode *T,
                      const uint64_t *poly_ring_off, uint64_t *ring_vert_off, double *vx, double *vy) {
	sb_prof_scope ps(ctx, "transcode");
	k_tc_poly_fill<<<sb_grid(n * 32, 256, 16), 256, 0, ctx->stream>>>(data, off, n, T->row_rings, T->row_verts,
	                                                                  poly_ring_off, T->totals, ring_vert_off, vx, vy);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

// ---- ST_Extent(WKB_BLOB) ------------------------------------------------------------------------------------------
// ST_Extent::ExecuteWKB (spatial_functions_scalar.cpp:2576-2629) = sgl::ops::wkb_reader_try_parse_stats
// (src/sgl/sgl.cpp:903-1058, nan_as_empty, stack of 128): one warp per blob, the part headers are walked by all lanes
// in lockstep (uniform loads), vertex runs are strided over the lanes, then a warp min / max reduction. Both byte
// orders, ISO (type + 1000 / 2000 / 3000) and EWKB (0x80000000 / 0x40000000 / 0x20000000 + SRID) flags, every ring
// of a polygon counts. min / max keep the accumulator when the vertex is NaN (std::min / std::max argument order).
#define WKB_STACK 128
#define WKB_WARPS 8

__device__ __forceinline__ uint32_t wkb_u32(const uint8_t *p, bool le) {
	const uint32_t v = ld_u32(p);
	return le ? v : __byte_perm(v, 0, 0x0123);
}
__device__ __forceinline__ double wkb_f64(const uint8_t *p, bool le) {
	if (le) {
		return ld_f64(p);
	}
	const unsigned long long hi = __byte_perm(ld_u32(p), 0, 0x0123), lo = __byte_perm(ld_u32(p + 4), 0, 0x0123);
	return __longlong_as_double((long long)(lo | (hi << 32)));
}

struct wkb_acc {
	double x0, y0, x1, y1;
};
__device__ __forceinline__ void wkb_add(wkb_acc &a, double X, double Y) {
	a.x0 = X < a.x0 ? X : a.x0;
	a.y0 = Y < a.y0 ? Y : a.y0;
	a.x1 = a.x1 < X ? X : a.x1;
	a.y1 = a.y1 < Y ? Y : a.y1;
}

// returns 1 / 0 / SB_ROW_* (all lanes agree); box valid in all lanes
__device__ int wkb_extent_warp(const uint8_t *b, uint64_t len, uint32_t *stack, double4 &box) {
	const unsigned lane = sb_lane();
	wkb_acc a = {DBL_MAX, DBL_MAX, -DBL_MAX, -DBL_MAX};
	uint64_t pos = 0;
	uint32_t depth = 0, vc = 0; // the reference counts vertices in uint32_t
	while (true) {
		if (pos + 5 > len) {
			return SB_ROW_TRUNCATED;
		}
		const bool le = b[pos] != 0;
		const uint32_t id = wkb_u32(b + pos + 1, le);
		pos += 5;
		// static_cast<geometry_type>(uint8_t): the low byte of the remainder
		const uint32_t type = ((id & 0xffffu) % 1000u) & 0xffu, fl = (id & 0xffffu) / 1000u;
		const bool has_z = fl == 1 || fl == 3 || (id & 0x80000000u), has_m = fl == 2 || fl == 3 || (id & 0x40000000u);
		if (id & 0x20000000u) {
			if (pos + 4 > len) {
				return SB_ROW_TRUNCATED;
			}
			pos += 4;
		}
		const uint32_t vsize = 16 + (has_z ? 8 : 0) + (has_m ? 8 : 0);
		bool descend = false;
		if (type == 1) {
			if (pos + vsize > len) {
				return SB_ROW_TRUNCATED;
			}
			const double x = wkb_f64(b + pos, le), y = wkb_f64(b + pos + 8, le);
			bool all_nan = x != x && y != y;
			for (uint32_t k = 16; k < vsize; k += 8) {
				const double w = wkb_f64(b + pos + k, le);
				all_nan = all_nan && w != w;
			}
			pos += vsize;
			if (!all_nan) {
				if (lane == 0) {
					wkb_add(a, x, y);
				}
				vc++;
			}
		} else if (type == 2 || type == 3) {
			uint32_t rings = 1;
			if (type == 3) {
				if (pos + 4 > len) {
					return SB_ROW_TRUNCATED;
				}
				rings = wkb_u32(b + pos, le);
				pos += 4;
			}
			for (uint32_t r = 0; r < rings; r++) {
				if (pos + 4 > len) {
					return SB_ROW_TRUNCATED;
				}
				const uint32_t np = wkb_u32(b + pos, le);
				pos += 4;
				if (pos + (uint64_t)np * vsize > len) {
					return SB_ROW_TRUNCATED;
				}
				for (uint32_t i = lane; i < np; i += 32) {
					const uint8_t *v = b + pos + (uint64_t)i * vsize;
					wkb_add(a, wkb_f64(v, le), wkb_f64(v + 8, le));
				}
				pos += (uint64_t)np * vsize;
				vc += np;
			}
		} else if (type >= 4 && type <= 7) {
			if (depth >= WKB_STACK) {
				return SB_ROW_RECURSION;
			}
			if (pos + 4 > len) {
				return SB_ROW_TRUNCATED;
			}
			const uint32_t count = wkb_u32(b + pos, le);
			pos += 4;
			if (count != 0) {
				__syncwarp();
				if (lane == 0) {
					stack[depth] = count;
				}
				__syncwarp();
				depth++;
				descend = true;
			}
		} else {
			return SB_ROW_WRONG_TYPE;
		}
		if (descend) {
			continue;
		}
		bool done = false;
		while (true) {
			if (depth == 0) {
				done = true;
				break;
			}
			__syncwarp();
			const uint32_t left = stack[depth - 1] - 1;
			__syncwarp();
			if (lane == 0) {
				stack[depth - 1] = left;
			}
			__syncwarp();
			if (left > 0) {
				break;
			}
			depth--;
		}
		if (done) {
			break;
		}
	}
#pragma unroll
	for (int s = 16; s > 0; s >>= 1) {
		const double ox0 = __shfl_xor_sync(0xFFFFFFFFu, a.x0, s), oy0 = __shfl_xor_sync(0xFFFFFFFFu, a.y0, s);
		const double ox1 = __shfl_xor_sync(0xFFFFFFFFu, a.x1, s), oy1 = __shfl_xor_sync(0xFFFFFFFFu, a.y1, s);
		a.x0 = ox0 < a.x0 ? ox0 : a.x0;
		a.y0 = oy0 < a.y0 ? oy0 : a.y0;
		a.x1 = a.x1 < ox1 ? ox1 : a.x1;
		a.y1 = a.y1 < oy1 ? oy1 : a.y1;
	}
	box = make_double4(a.x0, a.y0, a.x1, a.y1);
	return vc ? 1 : 0;
}

__global__ void __launch_bounds__(32 * WKB_WARPS) k_wkb_extent(const uint8_t *__restrict__ data,
                                                               const uint64_t *__restrict__ off, uint64_t n,
                                                               double4 *__restrict__ out_box, uint8_t *__restrict__ valid,
                                                               int32_t *__restrict__ status, int *__restrict__ any_bad) {
	__shared__ uint32_t s_stack[WKB_WARPS][WKB_STACK];
	const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
	for (uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
		const uint64_t o = off[i];
		double4 box = make_double4(0, 0, 0, 0);
		const int r = wkb_extent_warp(data + o, off[i + 1] - o, s_stack[threadIdx.x >> 5], box);
		if (sb_lane() == 0) {
			out_box[i] = r == 1 ? box : make_double4(0, 0, 0, 0);
			valid[i] = r == 1;
			if (status) {
				status[i] = r < 0 ? r : 0;
			}
			if (r < 0) {
				atomicMax(any_bad, 2);
			}
		}
	}
}

int sbi_wkb_extent(sb_context *ctx, const uint8_t *data, const uint64_t *off, uint64_t n, double *out_box, uint8_t *valid,
                   int32_t *status, int *any_bad) {
	if (n == 0) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "extent");
	k_wkb_extent<<<sb_grid(n * 32, 32 * WKB_WARPS, 16), 32 * WKB_WARPS, 0, ctx->stream>>>(
	    data, off, n, reinterpret_cast<double4 *>(out_box), valid, status, any_bad);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}
