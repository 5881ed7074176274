// bbox.cu -- K1 (bounding boxes from POINT_2D / BOX_2D / serialized GEOMETRY) and K2 (Hilbert keys).
// All kernels are HBM-streaming: one coalesced read of the coordinates, one coalesced write of the boxes / keys.
#include "internal.cuh"
#include "blob.cuh"

// ---------------------------------------------------------------------------------------------------------------
// POINT_2D -> BOX_2DF. 16 B in, 16 B out per point. Two points per thread: double2 loads, two float4 stores.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_bbox_points(const double *__restrict__ x, const double *__restrict__ y,
                                                     uint64_t n, float4 *__restrict__ out) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	const uint64_t npair = n >> 1;
	const bool aligned = ((((uintptr_t)x) | ((uintptr_t)y)) & 15) == 0;
	if (aligned) {
		const double2 *x2 = reinterpret_cast<const double2 *>(x);
		const double2 *y2 = reinterpret_cast<const double2 *>(y);
		for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < npair; i += stride) {
			const double2 xv = __ldg(x2 + i), yv = __ldg(y2 + i);
			out[2 * i] = make_float4(sb_f64_to_f32_down(xv.x), sb_f64_to_f32_down(yv.x), sb_f64_to_f32_up(xv.x),
			                         sb_f64_to_f32_up(yv.x));
			out[2 * i + 1] = make_float4(sb_f64_to_f32_down(xv.y), sb_f64_to_f32_down(yv.y), sb_f64_to_f32_up(xv.y),
			                             sb_f64_to_f32_up(yv.y));
		}
		if (blockIdx.x == 0 && threadIdx.x == 0 && (n & 1)) {
			const double xv = x[n - 1], yv = y[n - 1];
			out[n - 1] = make_float4(sb_f64_to_f32_down(xv), sb_f64_to_f32_down(yv), sb_f64_to_f32_up(xv),
			                         sb_f64_to_f32_up(yv));
		}
	} else {
		for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
			const double xv = x[i], yv = y[i];
			out[i] = make_float4(sb_f64_to_f32_down(xv), sb_f64_to_f32_down(yv), sb_f64_to_f32_up(xv),
			                     sb_f64_to_f32_up(yv));
		}
	}
}

__global__ void __launch_bounds__(256) k_bbox_box2d(const double *__restrict__ x0, const double *__restrict__ y0,
                                                    const double *__restrict__ x1, const double *__restrict__ y1,
                                                    uint64_t n, float4 *__restrict__ out) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		out[i] = make_float4(sb_f64_to_f32_down(x0[i]), sb_f64_to_f32_down(y0[i]), sb_f64_to_f32_up(x1[i]),
		                     sb_f64_to_f32_up(y1[i]));
	}
}

int sbi_bbox_points(sb_context *ctx, const double *x, const double *y, uint64_t n, float *out) {
	if (n == 0) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "bbox");
	k_bbox_points<<<sb_grid((n + 1) / 2, 256, 16), 256, 0, ctx->stream>>>(x, y, n, reinterpret_cast<float4 *>(out));
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

int sbi_bbox_box2d(sb_context *ctx, const double *x0, const double *y0, const double *x1, const double *y1, uint64_t n,
                   float *out) {
	if (n == 0) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "bbox");
	k_bbox_box2d<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(x0, y0, x1, y1, n, reinterpret_cast<float4 *>(out));
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Serialized GEOMETRY blobs. Byte-exact loads (blobs packed by the host may be only 1-byte aligned).
// ---------------------------------------------------------------------------------------------------------------
// geometry_t::TryGetCachedBounds, geometry_type.hpp:102-148. returns 1 / 0 / SB_ROW_* (<0)
__device__ __forceinline__ int cached_bounds(const uint8_t *b, uint64_t len, float4 &box) {
	if (len < 4) {
		return SB_ROW_TRUNCATED;
	}
	const uint8_t type = b[0], flags = b[1];
	if (flags & 0xC0) {
		return SB_ROW_BAD_VERSION;
	}
	if (flags & 0x04) {
		if (len < 24) {
			return SB_ROW_TRUNCATED;
		}
		box = make_float4(ld_f32(b + 8), ld_f32(b + 12), ld_f32(b + 16), ld_f32(b + 20));
		return 1;
	}
	if (type == 0) {
		if (len < 16) {
			return SB_ROW_TRUNCATED;
		}
		if (ld_u32(b + 12) == 0) {
			return 0;
		}
		if (len < 32) {
			return SB_ROW_TRUNCATED;
		}
		const double x = ld_f64(b + 16), y = ld_f64(b + 24);
		box = make_float4(sb_f64_to_f32_down(x), sb_f64_to_f32_down(y), sb_f64_to_f32_up(x), sb_f64_to_f32_up(y));
		return 1;
	}
	return 0;
}

// mode 0: ST_Extent_Approx (box + valid); mode 1: ST_Hilbert(geom) (key + valid)
template <int MODE>
__global__ void __launch_bounds__(256) k_blob_bounds(const uint8_t *__restrict__ data, const uint64_t *__restrict__ off,
                                                     uint64_t n, float4 *__restrict__ out_box,
                                                     uint32_t *__restrict__ out_key, uint8_t *__restrict__ valid,
                                                     int32_t *__restrict__ status, int *__restrict__ any_bad) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		const uint64_t o = off[i], e = off[i + 1];
		float4 box = make_float4(0.f, 0.f, 0.f, 0.f);
		const int r = cached_bounds(data + o, e - o, box);
		if (MODE == 0) {
			out_box[i] = box;
		} else {
			uint32_t h = 0;
			if (r == 1) {
				// ST_Hilbert::ExecuteGeometry, spatial_functions_scalar.cpp:5281-5286 (fp32)
				const float dx = box.x + (box.z - box.x) / 2.0f;
				const float dy = box.y + (box.w - box.y) / 2.0f;
				h = sb_hilbert16(sb_f32_order_u32(dx), sb_f32_order_u32(dy));
			}
			out_key[i] = h;
		}
		valid[i] = r == 1;
		if (status) {
			status[i] = r < 0 ? r : 0;
		}
		if (r < 0) {
			*any_bad = 1;
		}
	}
}

int sbi_blob_bounds(sb_context *ctx, int mode, const uint8_t *data, const uint64_t *off, uint64_t n, float *out_box,
                    uint32_t *out_key, uint8_t *valid, int32_t *status, int *any_bad) {
	if (n == 0) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "bbox");
	const unsigned g = sb_grid(n, 256, 16);
	if (mode == 0) {
		k_blob_bounds<0><<<g, 256, 0, ctx->stream>>>(data, off, n, reinterpret_cast<float4 *>(out_box), nullptr, valid,
		                                             status, any_bad);
	} else {
		k_blob_bounds<1><<<g, 256, 0, ctx->stream>>>(data, off, n, nullptr, out_key, valid, status, any_bad);
	}
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Exact extent (ST_Extent): one warp per blob. All lanes walk the part headers in lockstep (uniform loads), the
// vertex loops are strided over the lanes, then a warp min/max reduction. Polygons: shell ring only
// (sgl.hpp:1436-1449). min/max keep the accumulator when the vertex is NaN (std::min/std::max argument order).
// ---------------------------------------------------------------------------------------------------------------
#define SB_MAX_NEST 16

struct ext_acc {
	double x0, y0, x1, y1;
	int any;
};

// G lanes share one blob (G = 32, 16, 8 or 4: small blobs would leave most of a warp idle -- a 5-vertex envelope
// polygon kept 5 of 32 lanes busy and the kernel at 3.8 % of HBM, profiles r1j)
template <int G>
__device__ __forceinline__ void ext_vertices(const uint8_t *p, uint32_t count, uint32_t vsize, ext_acc &a) {
	for (uint32_t i = sb_lane() & (G - 1); i < count; i += G) {
		const double X = ld_f64(p + (uint64_t)i * vsize), Y = ld_f64(p + (uint64_t)i * vsize + 8);
		a.x0 = X < a.x0 ? X : a.x0;
		a.y0 = Y < a.y0 ? Y : a.y0;
		a.x1 = a.x1 < X ? X : a.x1;
		a.y1 = a.y1 < Y ? Y : a.y1;
	}
	if (count > 0) {
		a.any = 1;
	}
}

// returns 1 / 0 / SB_ROW_*; box valid in all lanes of the group after the reduction. `live` = the group has a blob
// (all 32 lanes run the final shuffles together, so no lane may leave early).
template <int G>
__device__ int extent_exact_group(const uint8_t *b, uint64_t len, bool live, double4 &box) {
	ext_acc a = {DBL_MAX, DBL_MAX, -DBL_MAX, -DBL_MAX, 0};
	int err = 0;
	if (live) {
		do {
			if (len < 8) {
				err = SB_ROW_TRUNCATED;
				break;
			}
			const uint8_t flags = b[1];
			if (flags & 0xC0) {
				err = SB_ROW_BAD_VERSION;
				break;
			}
			const uint32_t dims = 2 + ((flags & 1) ? 1 : 0) + ((flags & 2) ? 1 : 0);
			const uint32_t vsize = 8 * dims;
			uint64_t pos = 8;
			if (flags & 0x04) {
				pos += 8 * dims;
			}
			// explicit stack of "remaining parts" per open collection
			uint32_t remaining[SB_MAX_NEST];
			int depth = 0;
			bool top = true;
			uint32_t type = b[0];
			while (true) {
				// read [u32 type][u32 count] (the top-level type word repeats the header type)
				if (pos + 8 > len) {
					err = SB_ROW_TRUNCATED;
					break;
				}
				const uint32_t t = ld_u32(b + pos);
				const uint32_t count = ld_u32(b + pos + 4);
				pos += 8;
				if (!top) {
					type = t;
				}
				top = false;
				if (type <= 1) {
					if (pos + (uint64_t)count * vsize > len) {
						err = SB_ROW_TRUNCATED;
						break;
					}
					ext_vertices<G>(b + pos, count, vsize, a);
					pos += (uint64_t)count * vsize;
				} else if (type == 2) {
					const uint64_t rc = pos;
					pos += (uint64_t)count * 4 + ((count & 1) ? 4 : 0);
					if (pos > len) {
						err = SB_ROW_TRUNCATED;
						break;
					}
					for (uint32_t r = 0; r < count && !err; r++) {
						const uint32_t rcount = ld_u32(b + rc + 4ull * r);
						if (pos + (uint64_t)rcount * vsize > len) {
							err = SB_ROW_TRUNCATED;
							break;
						}
						if (r == 0) {
							ext_vertices<G>(b + pos, rcount, vsize, a);
						}
						pos += (uint64_t)rcount * vsize;
					}
					if (err) {
						break;
					}
				} else if (type <= 6) {
					if (count > 0) {
						if (depth == SB_MAX_NEST) {
							err = SB_ROW_TRUNCATED;
							break;
						}
						remaining[depth++] = count;
						continue;
					}
				} else {
					err = SB_ROW_TRUNCATED;
					break;
				}
				// part finished: pop finished collections
				while (depth > 0 && --remaining[depth - 1] == 0) {
					depth--;
				}
				if (depth == 0) {
					break;
				}
			}
		} while (false);
	}
	__syncwarp();
#pragma unroll
	for (int s = G / 2; s > 0; s >>= 1) {
		const double ox0 = __shfl_xor_sync(0xFFFFFFFFu, a.x0, s), oy0 = __shfl_xor_sync(0xFFFFFFFFu, a.y0, s);
		const double ox1 = __shfl_xor_sync(0xFFFFFFFFu, a.x1, s), oy1 = __shfl_xor_sync(0xFFFFFFFFu, a.y1, s);
		a.x0 = ox0 < a.x0 ? ox0 : a.x0;
		a.y0 = oy0 < a.y0 ? oy0 : a.y0;
		a.x1 = a.x1 < ox1 ? ox1 : a.x1;
		a.y1 = a.y1 < oy1 ? oy1 : a.y1;
	}
	box = make_double4(a.x0, a.y0, a.x1, a.y1);
	if (err) {
		return err;
	}
	return a.any ? 1 : 0;
}

// mode 0: ST_Extent; mode 1: ST_Hilbert(geom, bounds)
template <int MODE, int G>
__global__ void __launch_bounds__(256) k_blob_extent(const uint8_t *__restrict__ data, const uint64_t *__restrict__ off,
                                                     uint64_t n, double4 *__restrict__ out_box,
                                                     uint8_t *__restrict__ valid, int32_t *__restrict__ status,
                                                     double4 bounds, uint32_t *__restrict__ out_key,
                                                     int *__restrict__ any_bad) {
	constexpr int LOG_G = G == 32 ? 5 : (G == 16 ? 4 : (G == 8 ? 3 : 2));
	const uint64_t groups = ((uint64_t)gridDim.x * blockDim.x) >> LOG_G;
	const uint64_t n_round = (n + (32 / G) - 1) / (32 / G) * (32 / G); // whole warps iterate together
	for (uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> LOG_G; i < n_round; i += groups) {
		const bool live = i < n;
		const uint64_t o = live ? off[i] : 0, e = live ? off[i + 1] : 0;
		double4 box = make_double4(0, 0, 0, 0);
		const int r = extent_exact_group<G>(data + o, e - o, live, box);
		if (live && (sb_lane() & (G - 1)) == 0) {
			if (MODE == 0) {
				out_box[i] = r == 1 ? box : make_double4(0, 0, 0, 0);
				valid[i] = r == 1;
				if (status) {
					status[i] = r < 0 ? r : 0;
				}
				if (r < 0) {
					*any_bad = 1;
				}
			} else {
				uint32_t h = 0;
				if (r == 1) {
					// ExecuteGeometryWithBounds, spatial_functions_scalar.cpp:5312-5322 (fp64)
					const double dx = box.x + (box.z - box.x) / 2.0;
					const double dy = box.y + (box.w - box.y) / 2.0;
					const double hw = 65535.0 / (bounds.z - bounds.x);
					const double hh = 65535.0 / (bounds.w - bounds.y);
					h = sb_hilbert16(sb_cast_f64_u32((dx - bounds.x) * hw), sb_cast_f64_u32((dy - bounds.y) * hh));
				}
				out_key[i] = h;
				status[i] = r == 1 ? 0 : (r == 0 ? SB_ROW_EMPTY : r);
				if (r != 1) {
					*any_bad = r == 0 ? 2 : 1;
				}
			}
		}
	}
}

int sbi_blob_extent(sb_context *ctx, int mode, const uint8_t *data, const uint64_t *off, uint64_t n, double *out_box,
                    uint8_t *valid, int32_t *status, const double *bounds, uint32_t *out_key, int *any_bad,
                    uint64_t avg_bytes) {
	if (n == 0) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "extent");
	// lanes per blob from the mean blob size (avg_bytes = 0: unknown -> a whole warp): ~1 vertex (16 B) per lane and pass
	const int G = avg_bytes == 0 ? 32 : (avg_bytes <= 192 ? 4 : (avg_bytes <= 512 ? 8 : (avg_bytes <= 1536 ? 16 : 32)));
	const unsigned g = sb_grid(n * (uint64_t)G, 256, 32);
	const double4 bz = make_double4(0, 0, 0, 0);
	const double4 bb = mode == 0 ? bz : make_double4(bounds[0], bounds[1], bounds[2], bounds[3]);
	double4 *ob = reinterpret_cast<double4 *>(out_box);
#define SB_EXT_LAUNCH(GG)                                                                                            \
	if (mode == 0) {                                                                                                 \
		k_blob_extent<0, GG><<<g, 256, 0, ctx->stream>>>(data, off, n, ob, valid, status, bz, nullptr, any_bad);      \
	} else {                                                                                                         \
		k_blob_extent<1, GG><<<g, 256, 0, ctx->stream>>>(data, off, n, nullptr, nullptr, status, bb, out_key, any_bad); \
	}
	if (G == 4) {
		SB_EXT_LAUNCH(4)
	} else if (G == 8) {
		SB_EXT_LAUNCH(8)
	} else if (G == 16) {
		SB_EXT_LAUNCH(16)
	} else {
		SB_EXT_LAUNCH(32)
	}
#undef SB_EXT_LAUNCH
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// ST_Hilbert elementwise variants
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_hilbert_xy(const double *__restrict__ x, const double *__restrict__ y,
                                                    uint64_t n, double4 b, uint32_t *__restrict__ out) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	// ExecuteLonlat :5258-5265: width/height recomputed per row in the reference; same value every row
	const double hw = 65535.0 / (b.z - b.x);
	const double hh = 65535.0 / (b.w - b.y);
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		out[i] = sb_hilbert16(sb_cast_f64_u32((x[i] - b.x) * hw), sb_cast_f64_u32((y[i] - b.y) * hh));
	}
}

__global__ void __launch_bounds__(256) k_hilbert_box2d(const double4 *__restrict__ box, uint64_t n, double4 b,
                                                       uint32_t *__restrict__ out) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	const double hw = 65535.0 / (b.z - b.x);
	const double hh = 65535.0 / (b.w - b.y);
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		const double4 v = box[i];
		const double x = v.x + (v.z - v.x) / 2.0;
		const double y = v.y + (v.w - v.y) / 2.0;
		out[i] = sb_hilbert16(sb_cast_f64_u32((x - b.x) * hw), sb_cast_f64_u32((y - b.y) * hh));
	}
}

__global__ void __launch_bounds__(256) k_hilbert_box2df(const float4 *__restrict__ box, uint64_t n, float4 b,
                                                        uint32_t *__restrict__ out) {
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	const float hw = 65535.0f / (b.z - b.x);
	const float hh = 65535.0f / (b.w - b.y);
	for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		const float4 v = box[i];
		const float x = v.x + (v.z - v.x) / 2.0f;
		const float y = v.y + (v.w - v.y) / 2.0f;
		out[i] = sb_hilbert16(sb_cast_f32_u32((x - b.x) * hw), sb_cast_f32_u32((y - b.y) * hh));
	}
}

int sbi_hilbert_xy(sb_context *ctx, const double *x, const double *y, uint64_t n, const double *b, uint32_t *out) {
	if (n == 0) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "hilbert");
	k_hilbert_xy<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(x, y, n, make_double4(b[0], b[1], b[2], b[3]), out);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

int sbi_hilbert_box2d(sb_context *ctx, const double *box4, uint64_t n, const double *b, uint32_t *out) {
	if (n == 0) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "hilbert");
	k_hilbert_box2d<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(reinterpret_cast<const double4 *>(box4), n,
	                                                             make_double4(b[0], b[1], b[2], b[3]), out);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}

int sbi_hilbert_box2df(sb_context *ctx, const float *box4, uint64_t n, const float *b, uint32_t *out) {
	if (n == 0) {
		return SB_OK;
	}
	sb_prof_scope ps(ctx, "hilbert");
	k_hilbert_box2df<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(reinterpret_cast<const float4 *>(box4), n,
	                                                              make_float4(b[0], b[1], b[2], b[3]), out);
	SB_LAUNCH_CHECK(ctx);
	return SB_OK;
}
