// capi.cu -- the extern "C" boundary (include/spatial_b200.h). Host-pointer calls stage through the context's
// device arena on the context's stream and return after the results are back in host memory; device-pointer
// calls only enqueue work (except where a count has to be returned).
#include "internal.cuh"

#include <math.h>

#include <atomic>
#include <string>
#include <thread>
#include <vector>

#define SB_BEGIN(ctx)                                                                                                \
	SB_ARG((ctx) != nullptr, "ctx");                                                                                 \
	SB_CUDA(cudaSetDevice((ctx)->device));                                                                           \
	(ctx)->arena.reset()

static int finish(sb_context *ctx, sb_mem mem) {
	if (mem == SB_HOST) {
		SB_CUDA(cudaStreamSynchronize(ctx->stream));
	}
	return SB_OK;
}

// ---- K1 ---------------------------------------------------------------------------------------------------------
extern "C" int sb_bbox_from_points(sb_context *ctx, sb_mem mem, const double *x, const double *y, uint64_t n,
                                   float *out_box4) {
	SB_BEGIN(ctx);
	const double *dx, *dy;
	float *dout;
	SB_TRY(sb_to_device(ctx, mem, x, n, &dx));
	SB_TRY(sb_to_device(ctx, mem, y, n, &dy));
	SB_TRY(sb_out_buffer(ctx, mem, out_box4, 4 * n, &dout));
	SB_TRY(sbi_bbox_points(ctx, dx, dy, n, dout));
	SB_TRY(sb_to_host(ctx, mem, out_box4, dout, 4 * n));
	return finish(ctx, mem);
}

extern "C" int sb_bbox_from_box2d(sb_context *ctx, sb_mem mem, const double *min_x, const double *min_y,
                                  const double *max_x, const double *max_y, uint64_t n, float *out_box4) {
	SB_BEGIN(ctx);
	const double *a, *b, *c, *d;
	float *dout;
	SB_TRY(sb_to_device(ctx, mem, min_x, n, &a));
	SB_TRY(sb_to_device(ctx, mem, min_y, n, &b));
	SB_TRY(sb_to_device(ctx, mem, max_x, n, &c));
	SB_TRY(sb_to_device(ctx, mem, max_y, n, &d));
	SB_TRY(sb_out_buffer(ctx, mem, out_box4, 4 * n, &dout));
	SB_TRY(sbi_bbox_box2d(ctx, a, b, c, d, n, dout));
	SB_TRY(sb_to_host(ctx, mem, out_box4, dout, 4 * n));
	return finish(ctx, mem);
}

// shared by the four blob entry points. which: 0 approx box, 1 hilbert(geom), 2 exact extent, 3 hilbert(geom,bounds)
static int blob_call(sb_context *ctx, sb_mem mem, int which, const uint8_t *data, const uint64_t *offsets, uint64_t n,
                     float *out_f, double *out_d, uint32_t *out_key, uint8_t *out_valid, int32_t *row_status,
                     const double *bounds) {
	SB_BEGIN(ctx);
	if (n == 0) {
		return SB_OK;
	}
	SB_ARG(data != nullptr && offsets != nullptr, "data/offsets");
	uint64_t total = 0;
	if (mem == SB_HOST) {
		total = offsets[n];
	} else {
		SB_CUDA(cudaMemcpyAsync(&total, offsets + n, sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
		SB_CUDA(cudaStreamSynchronize(ctx->stream));
	}
	const uint8_t *d_data;
	const uint64_t *d_off;
	SB_TRY(sb_to_device(ctx, mem, data, total, &d_data));
	SB_TRY(sb_to_device(ctx, mem, offsets, n + 1, &d_off));
	float *d_f = nullptr;
	double *d_d = nullptr;
	uint32_t *d_key = nullptr;
	uint8_t *d_valid = nullptr;
	int32_t *d_status = nullptr;
	int *d_bad = nullptr;
	SB_TRY(ctx->scratch_t(1, &d_bad));
	SB_CUDA(cudaMemsetAsync(d_bad, 0, sizeof(int), ctx->stream));
	SB_TRY(sb_out_buffer(ctx, mem, out_f, 4 * n, &d_f));
	SB_TRY(sb_out_buffer(ctx, mem, out_d, 4 * n, &d_d));
	SB_TRY(sb_out_buffer(ctx, mem, out_key, n, &d_key));
	SB_TRY(sb_out_buffer(ctx, mem, out_valid, n, &d_valid));
	if (row_status || which == 3) {
		if (row_status) {
			SB_TRY(sb_out_buffer(ctx, mem, row_status, n, &d_status));
		} else {
			SB_TRY(ctx->scratch_t(n, &d_status));
		}
	}
	if (which == 0) {
		SB_TRY(sbi_blob_bounds(ctx, 0, d_data, d_off, n, d_f, nullptr, d_valid, d_status, d_bad));
	} else if (which == 1) {
		SB_TRY(sbi_blob_bounds(ctx, 1, d_data, d_off, n, nullptr, d_key, d_valid, d_status, d_bad));
	} else if (which == 2) {
		SB_TRY(sbi_blob_extent(ctx, 0, d_data, d_off, n, d_d, d_valid, d_status, nullptr, nullptr, d_bad, total / n));
	} else {
		SB_TRY(sbi_blob_extent(ctx, 1, d_data, d_off, n, nullptr, nullptr, d_status, bounds, d_key, d_bad, total / n));
	}
	SB_TRY(sb_to_host(ctx, mem, out_f, d_f, 4 * n));
	SB_TRY(sb_to_host(ctx, mem, out_d, d_d, 4 * n));
	SB_TRY(sb_to_host(ctx, mem, out_key, d_key, n));
	SB_TRY(sb_to_host(ctx, mem, out_valid, d_valid, n));
	SB_TRY(sb_to_host(ctx, mem, row_status, d_status, n));
	int bad = 0;
	SB_CUDA(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
	SB_CUDA(cudaStreamSynchronize(ctx->stream));
	if (bad == 1) {
		sb_set_error("malformed or unsupported-version geometry blob(s); see row_status");
		return SB_ERR_GEOMETRY;
	}
	if (bad == 2) {
		sb_set_error("ST_Hilbert(geom, bounds) does not support empty geometries");
		return SB_ERR_EMPTY;
	}
	return SB_OK;
}

extern "C" int sb_bbox_from_blobs(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets, uint64_t n,
                                  float *out_box4, uint8_t *out_valid, int32_t *row_status) {
	return blob_call(ctx, mem, 0, data, offsets, n, out_box4, nullptr, nullptr, out_valid, row_status, nullptr);
}
extern "C" int sb_extent_from_blobs(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets,
                                    uint64_t n, double *out_box4, uint8_t *out_valid, int32_t *row_status) {
	return blob_call(ctx, mem, 2, data, offsets, n, nullptr, out_box4, nullptr, out_valid, row_status, nullptr);
}
extern "C" int sb_hilbert_geom(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets, uint64_t n,
                               uint32_t *out, uint8_t *out_valid) {
	return blob_call(ctx, mem, 1, data, offsets, n, nullptr, nullptr, out, out_valid, nullptr, nullptr);
}
extern "C" int sb_hilbert_geom_bounds(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets,
                                      uint64_t n, const double *bounds4, uint32_t *out, int32_t *row_status) {
	SB_ARG(bounds4 != nullptr, "bounds4");
	return blob_call(ctx, mem, 3, data, offsets, n, nullptr, nullptr, out, nullptr, row_status, bounds4);
}

// ---- GEOMETRY -> POINT_2D / POLYGON_2D columns -------------------------------------------------------------------
static int blob_inputs(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets, uint64_t n,
                       const uint8_t **d_data, const uint64_t **d_off) {
	SB_ARG(data != nullptr && offsets != nullptr, "data/offsets");
	uint64_t total = 0;
	if (mem == SB_HOST) {
		total = offsets[n];
	} else {
		SB_TRY(sbi_read_words(ctx, offsets + n, 1, &total));
	}
	SB_TRY(sb_to_device(ctx, mem, data, total, d_data));
	SB_TRY(sb_to_device(ctx, mem, offsets, n + 1, d_off));
	return SB_OK;
}

static int blob_verdict(sb_context *ctx, const int *d_bad) {
	uint64_t bad = 0;
	SB_TRY(sbi_read_words(ctx, d_bad, 1, &bad)); // d_bad is an 8-byte scratch cell, low word = flag
	if ((uint32_t)bad == 2) {
		sb_set_error("GEOMETRY cast: row(s) of another geometry type, or malformed blob(s); see row_status");
		return SB_ERR_GEOMETRY;
	}
	if ((uint32_t)bad == 1) {
		sb_set_error("Cannot cast empty point GEOMETRY to POINT_2D; see row_status");
		return SB_ERR_EMPTY;
	}
	return SB_OK;
}

extern "C" int sb_points_from_blobs(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets, uint64_t n,
                                    double *out_x, double *out_y, int32_t *row_status) {
	SB_BEGIN(ctx);
	if (n == 0) {
		return SB_OK;
	}
	SB_ARG(out_x != nullptr && out_y != nullptr, "out_x/out_y");
	const uint8_t *d_data;
	const uint64_t *d_off;
	SB_TRY(blob_inputs(ctx, mem, data, offsets, n, &d_data, &d_off));
	double *dx = nullptr, *dy = nullptr;
	int32_t *d_status = nullptr;
	uint64_t *d_bad = nullptr;
	SB_TRY(ctx->scratch_t(1, &d_bad));
	SB_TRY(sbi_zero(ctx, d_bad, sizeof(uint64_t)));
	SB_TRY(sb_out_buffer(ctx, mem, out_x, n, &dx));
	SB_TRY(sb_out_buffer(ctx, mem, out_y, n, &dy));
	SB_TRY(sb_out_buffer(ctx, mem, row_status, n, &d_status));
	SB_TRY(sbi_points_from_blobs(ctx, d_data, d_off, n, dx, dy, d_status, reinterpret_cast<int *>(d_bad)));
	SB_TRY(sb_to_host(ctx, mem, out_x, dx, n));
	SB_TRY(sb_to_host(ctx, mem, out_y, dy, n));
	SB_TRY(sb_to_host(ctx, mem, row_status, d_status, n));
	const int verdict = blob_verdict(ctx, reinterpret_cast<const int *>(d_bad));
	const int s = finish(ctx, mem);
	return s != SB_OK ? s : verdict;
}

extern "C" int sb_extent_from_wkb(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets, uint64_t n,
                                  double *out_box4, uint8_t *out_valid, int32_t *row_status) {
	SB_BEGIN(ctx);
	if (n == 0) {
		return SB_OK;
	}
	SB_ARG(out_box4 != nullptr && out_valid != nullptr, "out_box4/out_valid");
	const uint8_t *d_data;
	const uint64_t *d_off;
	SB_TRY(blob_inputs(ctx, mem, data, offsets, n, &d_data, &d_off));
	double *d_box = nullptr;
	uint8_t *d_valid = nullptr;
	int32_t *d_status = nullptr;
	uint64_t *d_bad = nullptr;
	SB_TRY(ctx->scratch_t(1, &d_bad));
	SB_TRY(sbi_zero(ctx, d_bad, sizeof(uint64_t)));
	SB_TRY(sb_out_buffer(ctx, mem, out_box4, 4 * n, &d_box));
	SB_TRY(sb_out_buffer(ctx, mem, out_valid, n, &d_valid));
	SB_TRY(sb_out_buffer(ctx, mem, row_status, n, &d_status));
	SB_TRY(sbi_wkb_extent(ctx, d_data, d_off, n, d_box, d_valid, d_status, reinterpret_cast<int *>(d_bad)));
	SB_TRY(sb_to_host(ctx, mem, out_box4, d_box, 4 * n));
	SB_TRY(sb_to_host(ctx, mem, out_valid, d_valid, n));
	SB_TRY(sb_to_host(ctx, mem, row_status, d_status, n));
	uint64_t bad = 0;
	SB_TRY(sbi_read_words(ctx, d_bad, 1, &bad));
	if ((uint32_t)bad) {
		sb_set_error("Failed to parse WKB: see row_status");
		return SB_ERR_GEOMETRY;
	}
	return finish(ctx, mem);
}

extern "C" int sb_polygons_from_blobs(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets,
                                      uint64_t n, uint64_t *poly_ring_off, uint64_t *ring_vert_off, double *vx, double *vy,
                                      uint64_t ring_capacity, uint64_t vertex_capacity, uint64_t *n_rings,
                                      uint64_t *n_verts, int32_t *row_status) {
	SB_BEGIN(ctx);
	SB_ARG(n_rings != nullptr && n_verts != nullptr, "n_rings/n_verts");
	SB_ARG(poly_ring_off != nullptr, "poly_ring_off");
	*n_rings = 0;
	*n_verts = 0;
	if (n == 0) {
		if (mem == SB_HOST) {
			poly_ring_off[0] = 0;
			if (ring_vert_off) {
				ring_vert_off[0] = 0;
			}
		} else {
			SB_TRY(sbi_zero(ctx, poly_ring_off, sizeof(uint64_t)));
			if (ring_vert_off) {
				SB_TRY(sbi_zero(ctx, ring_vert_off, sizeof(uint64_t)));
			}
			SB_CUDA(cudaStreamSynchronize(ctx->stream));
		}
		return SB_OK;
	}
	const uint8_t *d_data;
	const uint64_t *d_off;
	SB_TRY(blob_inputs(ctx, mem, data, offsets, n, &d_data, &d_off));
	uint64_t *d_pro = nullptr, *d_rvo = nullptr;
	double *dvx = nullptr, *dvy = nullptr;
	int32_t *d_status = nullptr;
	uint64_t *d_bad = nullptr;
	SB_TRY(ctx->scratch_t(1, &d_bad));
	SB_TRY(sbi_zero(ctx, d_bad, sizeof(uint64_t)));
	SB_TRY(sb_out_buffer(ctx, mem, poly_ring_off, n + 1, &d_pro));
	SB_TRY(sb_out_buffer(ctx, mem, row_status, n, &d_status));
	sb_poly_transcode T;
	SB_TRY(sbi_polygons_count(ctx, d_data, d_off, n, &T, d_status, reinterpret_cast<int *>(d_bad), d_pro, n_rings, n_verts));
	SB_TRY(sb_to_host(ctx, mem, poly_ring_off, d_pro, n + 1));
	SB_TRY(sb_to_host(ctx, mem, row_status, d_status, n));
	const bool fits = *n_rings <= ring_capacity && *n_verts <= vertex_capacity && ring_vert_off != nullptr &&
	                  (*n_verts == 0 || (vx != nullptr && vy != nullptr));
	if (fits) {
		SB_TRY(sb_out_buffer(ctx, mem, ring_vert_off, *n_rings + 1, &d_rvo));
		SB_TRY(sb_out_buffer(ctx, mem, vx, *n_verts, &dvx));
		SB_TRY(sb_out_buffer(ctx, mem, vy, *n_verts, &dvy));
		SB_TRY(sbi_polygons_fill(ctx, d_data, d_off, n, &T, d_pro, d_rvo, dvx, dvy));
		SB_TRY(sb_to_host(ctx, mem, ring_vert_off, d_rvo, *n_rings + 1));
		SB_TRY(sb_to_host(ctx, mem, vx, dvx, *n_verts));
		SB_TRY(sb_to_host(ctx, mem, vy, dvy, *n_verts));
	}
	const int verdict = blob_verdict(ctx, reinterpret_cast<const int *>(d_bad));
	const int s = finish(ctx, mem);
	if (s != SB_OK) {
		return s;
	}
	if (!fits) {
		sb_set_error("polygons_from_blobs: %llu rings / %llu vertices, capacity %llu / %llu",
		             (unsigned long long)*n_rings, (unsigned long long)*n_verts, (unsigned long long)ring_capacity,
		             (unsigned long long)vertex_capacity);
		return SB_ERR_CAPACITY;
	}
	return verdict;
}

// ---- K2 ---------------------------------------------------------------------------------------------------------
extern "C" int sb_hilbert_xy(sb_context *ctx, sb_mem mem, const double *x, const double *y, uint64_t n,
                             const double *bounds4, uint32_t *out) {
	SB_BEGIN(ctx);
	SB_ARG(bounds4 != nullptr, "bounds4");
	const double *dx, *dy;
	uint32_t *dout;
	SB_TRY(sb_to_device(ctx, mem, x, n, &dx));
	SB_TRY(sb_to_device(ctx, mem, y, n, &dy));
	SB_TRY(sb_out_buffer(ctx, mem, out, n, &dout));
	SB_TRY(sbi_hilbert_xy(ctx, dx, dy, n, bounds4, dout));
	SB_TRY(sb_to_host(ctx, mem, out, dout, n));
	return finish(ctx, mem);
}

extern "C" int sb_hilbert_box2d(sb_context *ctx, sb_mem mem, const double *box4, uint64_t n, const double *bounds4,
                                uint32_t *out) {
	SB_BEGIN(ctx);
	SB_ARG(bounds4 != nullptr, "bounds4");
	const double *db;
	uint32_t *dout;
	SB_TRY(sb_to_device(ctx, mem, box4, 4 * n, &db));
	SB_TRY(sb_out_buffer(ctx, mem, out, n, &dout));
	SB_TRY(sbi_hilbert_box2d(ctx, db, n, bounds4, dout));
	SB_TRY(sb_to_host(ctx, mem, out, dout, n));
	return finish(ctx, mem);
}

extern "C" int sb_hilbert_box2df(sb_context *ctx, sb_mem mem, const float *box4, uint64_t n, const float *bounds4,
                                 uint32_t *out) {
	SB_BEGIN(ctx);
	SB_ARG(bounds4 != nullptr, "bounds4");
	const float *db;
	uint32_t *dout;
	SB_TRY(sb_to_device(ctx, mem, box4, 4 * n, &db));
	SB_TRY(sb_out_buffer(ctx, mem, out, n, &dout));
	SB_TRY(sbi_hilbert_box2df(ctx, db, n, bounds4, dout));
	SB_TRY(sb_to_host(ctx, mem, out, dout, n));
	return finish(ctx, mem);
}

// ---- K3 ---------------------------------------------------------------------------------------------------------
template <typename KeyT>
static int sort_pairs(sb_context *ctx, sb_mem mem, KeyT *keys, uint32_t *values, uint64_t n) {
	SB_BEGIN(ctx);
	if (n == 0) {
		return SB_OK;
	}
	KeyT *k = nullptr, *kb = nullptr;
	uint32_t *v = nullptr, *vb = nullptr;
	if (mem == SB_HOST) {
		SB_TRY(ctx->scratch_t(n, &k));
		SB_TRY(ctx->scratch_t(n, &v));
		SB_CUDA(cudaMemcpyAsync(k, keys, n * sizeof(KeyT), cudaMemcpyHostToDevice, ctx->stream));
		SB_CUDA(cudaMemcpyAsync(v, values, n * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
	} else {
		k = keys;
		v = values;
	}
	SB_TRY(ctx->scratch_t(n, &kb));
	SB_TRY(ctx->scratch_t(n, &vb));
	int alt = 0;
	if (sizeof(KeyT) == 4) {
		SB_TRY(sbi_radix_sort_u32(ctx, (uint32_t *)k, (uint32_t *)kb, v, vb, n, 0, 32, false, &alt));
	} else {
		SB_TRY(sbi_radix_sort_u64(ctx, (uint64_t *)k, (uint64_t *)kb, v, vb, n, 0, 64, false, &alt));
	}
	const KeyT *rk = alt ? kb : k;
	const uint32_t *rv = alt ? vb : v;
	const cudaMemcpyKind kind = mem == SB_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
	if (mem == SB_HOST || alt) {
		SB_CUDA(cudaMemcpyAsync(keys, rk, n * sizeof(KeyT), kind, ctx->stream));
		SB_CUDA(cudaMemcpyAsync(values, rv, n * sizeof(uint32_t), kind, ctx->stream));
	}
	return finish(ctx, mem);
}

extern "C" int sb_sort_pairs_u32(sb_context *ctx, sb_mem mem, uint32_t *keys, uint32_t *values, uint64_t n) {
	return sort_pairs<uint32_t>(ctx, mem, keys, values, n);
}
extern "C" int sb_sort_pairs_u64(sb_context *ctx, sb_mem mem, uint64_t *keys, uint32_t *values, uint64_t n) {
	return sort_pairs<uint64_t>(ctx, mem, keys, values, n);
}

// ---- K4a --------------------------------------------------------------------------------------------------------
extern "C" int sb_rtree_build_str(sb_context *ctx, sb_mem mem, const float *box4, const int64_t *row_ids, uint64_t n,
                                  uint32_t max_node_capacity, sb_rtree **out) {
	SB_BEGIN(ctx);
	SB_ARG(out != nullptr, "out");
	*out = nullptr;
	const float *db;
	const int64_t *dr;
	SB_TRY(sb_to_device(ctx, mem, box4, 4 * n, &db));
	SB_TRY(sb_to_device(ctx, mem, row_ids, n, &dr));
	sb_rtree *t = nullptr;
	int s = sbi_rtree_build_str(ctx, db, dr, n, max_node_capacity, &t);
	if (s != SB_OK) {
		if (t) {
			sb_rtree_destroy(t);
		}
		return s;
	}
	*out = t;
	return finish(ctx, mem);
}

extern "C" int sb_rtree_build_from_points(sb_context *ctx, sb_mem mem, const double *x, const double *y,
                                          const int64_t *row_ids, uint64_t n, uint32_t max_node_capacity, sb_rtree **out) {
	SB_BEGIN(ctx);
	SB_ARG(out != nullptr, "out");
	*out = nullptr;
	SB_ARG(n == 0 || (x && y && row_ids), "arrays");
	const double *dx, *dy;
	const int64_t *dr;
	SB_TRY(sb_to_device(ctx, mem, x, n, &dx));
	SB_TRY(sb_to_device(ctx, mem, y, n, &dy));
	SB_TRY(sb_to_device(ctx, mem, row_ids, n, &dr));
	sb_rtree *t = nullptr;
	int s = sbi_rtree_build_str(ctx, nullptr, dr, n, max_node_capacity, &t, dx, dy);
	if (s != SB_OK) {
		if (t) {
			sb_rtree_destroy(t);
		}
		return s;
	}
	*out = t;
	return finish(ctx, mem);
}

extern "C" uint64_t sb_rtree_slice_for(uint64_t n, uint32_t max_node_capacity) {
	if (n == 0 || max_node_capacity == 0) {
		return 0;
	}
	const uint64_t cap = max_node_capacity;
	return (uint64_t)ceil(sqrt((double)((n + cap - 1) / cap))) * cap;
}

extern "C" int sb_rtree_build_leaf_shard(sb_context *ctx, sb_mem mem, const float *box4, const int64_t *row_ids,
                                         uint64_t n_shard, uint32_t max_node_capacity, uint64_t slice_size,
                                         uint64_t node_base, int64_t x_key_before, int64_t x_key_after,
                                         void *out_leaf_entries, void *out_node_entries) {
	SB_BEGIN(ctx);
	SB_ARG(n_shard == 0 || (box4 && row_ids && out_leaf_entries && out_node_entries), "arrays");
	const uint64_t nodes = (n_shard + max_node_capacity - 1) / (max_node_capacity ? max_node_capacity : 1);
	const float *db;
	const int64_t *dr;
	sb_entry *d0, *d1;
	SB_TRY(sb_to_device(ctx, mem, box4, 4 * n_shard, &db));
	SB_TRY(sb_to_device(ctx, mem, row_ids, n_shard, &dr));
	SB_TRY(sb_out_buffer(ctx, mem, static_cast<sb_entry *>(out_leaf_entries), n_shard, &d0));
	SB_TRY(sb_out_buffer(ctx, mem, static_cast<sb_entry *>(out_node_entries), nodes, &d1));
	SB_TRY(sbi_rtree_leaf_shard(ctx, db, dr, n_shard, max_node_capacity, slice_size, node_base, x_key_before, x_key_after,
	                            d0, d1));
	SB_TRY(sb_to_host(ctx, mem, static_cast<sb_entry *>(out_leaf_entries), d0, n_shard));
	SB_TRY(sb_to_host(ctx, mem, static_cast<sb_entry *>(out_node_entries), d1, nodes));
	return finish(ctx, mem);
}

extern "C" int sb_rtree_from_leaf_level(sb_context *ctx, sb_mem mem, const void *leaf_entries, uint64_t n,
                                        const void *node_entries, uint32_t max_node_capacity, sb_rtree **out) {
	SB_BEGIN(ctx);
	SB_ARG(out != nullptr && leaf_entries != nullptr && node_entries != nullptr, "out/entries");
	*out = nullptr;
	const uint64_t nodes = (n + max_node_capacity - 1) / (max_node_capacity ? max_node_capacity : 1);
	const sb_entry *d0, *d1;
	SB_TRY(sb_to_device(ctx, mem, static_cast<const sb_entry *>(leaf_entries), n, &d0));
	SB_TRY(sb_to_device(ctx, mem, static_cast<const sb_entry *>(node_entries), nodes, &d1));
	sb_rtree *t = nullptr;
	int s = sbi_rtree_from_leaf_level(ctx, d0, n, d1, max_node_capacity, &t);
	if (s != SB_OK) {
		if (t) {
			sb_rtree_destroy(t);
		}
		return s;
	}
	*out = t;
	return finish(ctx, mem);
}

extern "C" int sb_rtree_build_from_blobs(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets,
                                         uint64_t n, const int64_t *row_ids, uint32_t max_node_capacity, sb_rtree **out,
                                         uint64_t *indexed_rows, int32_t *row_status) {
	SB_BEGIN(ctx);
	SB_ARG(out != nullptr, "out");
	*out = nullptr;
	if (indexed_rows) {
		*indexed_rows = 0;
	}
	SB_ARG(n == 0 || (data != nullptr && offsets != nullptr), "data/offsets");
	uint64_t total = 0;
	if (n) {
		if (mem == SB_HOST) {
			total = offsets[n];
		} else {
			SB_CUDA(cudaMemcpyAsync(&total, offsets + n, sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
			SB_CUDA(cudaStreamSynchronize(ctx->stream));
		}
	}
	const uint8_t *d_data;
	const uint64_t *d_off;
	const int64_t *d_rows;
	SB_TRY(sb_to_device(ctx, mem, data, total, &d_data));
	SB_TRY(sb_to_device(ctx, mem, offsets, n + 1, &d_off));
	SB_TRY(sb_to_device(ctx, mem, row_ids, n, &d_rows));
	float *d_box = nullptr, *d_sbox = nullptr;
	int64_t *d_srows = nullptr;
	uint8_t *d_valid = nullptr;
	int32_t *d_status = nullptr;
	int *d_bad = nullptr;
	const uint64_t m = n ? n : 1;
	SB_TRY(ctx->scratch_t(4 * m, &d_box));
	SB_TRY(ctx->scratch_t(4 * m, &d_sbox));
	SB_TRY(ctx->scratch_t(m, &d_srows));
	SB_TRY(ctx->scratch_t(m, &d_valid));
	SB_TRY(ctx->scratch_t(1, &d_bad));
	SB_TRY(sbi_zero(ctx, d_bad, sizeof(int)));
	SB_TRY(sb_out_buffer(ctx, mem, row_status, n, &d_status));
	uint64_t kept = 0;
	if (n) {
		// cached bbox per row (geometry_type.hpp:102-148); rows without one are filtered out like the planner's
		// bounds filter does (rtree_index_create_logical.cpp:42-72), malformed rows are reported in row_status
		SB_TRY(sbi_blob_bounds(ctx, 0, d_data, d_off, n, d_box, nullptr, d_valid, d_status, d_bad));
		SB_TRY(sbi_select_valid(ctx, d_valid, n, d_box, d_rows, d_sbox, d_srows, &kept));
	}
	SB_TRY(sb_to_host(ctx, mem, row_status, d_status, n));
	sb_rtree *t = nullptr;
	int s = sbi_rtree_build_str(ctx, d_sbox, d_srows, kept, max_node_capacity, &t);
	if (s != SB_OK) {
		if (t) {
			sb_rtree_destroy(t);
		}
		return s;
	}
	*out = t;
	if (indexed_rows) {
		*indexed_rows = kept;
	}
	return finish(ctx, mem);
}

extern "C" void sb_rtree_destroy(sb_rtree *tree) {
	if (!tree) {
		return;
	}
	// stream-ordered on the building context's stream: ordered after the build and after every scan / export that
	// context enqueued. Work on OTHER contexts' streams must have completed before the tree is destroyed.
	sb_owner_free(tree->owner, tree->device, tree->entries);
	delete tree;
}
extern "C" int sb_rtree_num_levels(const sb_rtree *tree) {
	return tree ? tree->nlevels : 0;
}
extern "C" uint64_t sb_rtree_level_count(const sb_rtree *tree, int level) {
	return (tree && level >= 0 && level < tree->nlevels) ? tree->counts[level] : 0;
}
extern "C" uint64_t sb_rtree_slice_size(const sb_rtree *tree) {
	return tree ? tree->slice : 0;
}

extern "C" int sb_rtree_export_level(sb_context *ctx, const sb_rtree *tree, int level, sb_mem mem, void *out_entries) {
	SB_ARG(ctx != nullptr && tree != nullptr && out_entries != nullptr, "ctx/tree/out");
	SB_ARG(level >= 0 && level < tree->nlevels, "level");
	SB_CUDA(cudaSetDevice(ctx->device));
	SB_CUDA(cudaMemcpyAsync(out_entries, tree->entries + tree->offsets[level], tree->counts[level] * sizeof(sb_entry),
	                        mem == SB_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, ctx->stream));
	return finish(ctx, mem);
}

extern "C" int sb_rtree_dump(sb_context *ctx, const sb_rtree *tree, sb_mem mem, int32_t *out_level, float *out_box4,
                             int64_t *out_row_id, uint64_t capacity, uint64_t *count) {
	SB_BEGIN(ctx);
	SB_ARG(tree != nullptr && count != nullptr, "tree/count");
	uint64_t total = 0;
	for (int l = 0; l + 1 < tree->nlevels; l++) {
		total += tree->counts[l];
	}
	*count = total;
	if (total > capacity || out_level == nullptr) {
		if (out_level == nullptr) {
			return SB_OK; // size query
		}
		sb_set_error("rtree dump: %llu rows, capacity %llu", (unsigned long long)total, (unsigned long long)capacity);
		return SB_ERR_CAPACITY;
	}
	int32_t *dl;
	float *db;
	int64_t *dr;
	SB_TRY(sb_out_buffer(ctx, mem, out_level, total, &dl));
	SB_TRY(sb_out_buffer(ctx, mem, out_box4, 4 * total, &db));
	SB_TRY(sb_out_buffer(ctx, mem, out_row_id, total, &dr));
	SB_TRY(sbi_rtree_dump(ctx, tree, dl, db, dr, total, count));
	SB_TRY(sb_to_host(ctx, mem, out_level, dl, total));
	SB_TRY(sb_to_host(ctx, mem, out_box4, db, 4 * total));
	SB_TRY(sb_to_host(ctx, mem, out_row_id, dr, total));
	return finish(ctx, mem);
}

extern "C" int sb_rtree_pages_layout(const sb_rtree *tree, uint64_t block_size, sb_rtree_pages_info *info) {
	SB_ARG(tree != nullptr && info != nullptr, "tree/info");
	return sbi_rtree_pages_info(tree, block_size, info);
}

extern "C" int sb_rtree_export_pages(sb_context *ctx, const sb_rtree *tree, uint64_t block_size, sb_mem mem,
                                     void *leaf_images, void *branch_images, void *root_entry) {
	SB_BEGIN(ctx);
	SB_ARG(tree != nullptr && root_entry != nullptr, "tree/root_entry");
	sb_rtree_pages_info info;
	SB_TRY(sbi_rtree_pages_info(tree, block_size, &info));
	uint8_t *dl, *db;
	const uint64_t lbytes = info.leaf_buffers * block_size, bbytes = info.branch_buffers * block_size;
	SB_ARG((lbytes == 0 || leaf_images != nullptr) && (bbytes == 0 || branch_images != nullptr), "image buffers");
	SB_TRY(sb_out_buffer(ctx, mem, static_cast<uint8_t *>(leaf_images), lbytes, &dl));
	SB_TRY(sb_out_buffer(ctx, mem, static_cast<uint8_t *>(branch_images), bbytes, &db));
	SB_TRY(sbi_rtree_export_pages(ctx, tree, block_size, dl, db, root_entry));
	SB_TRY(sb_to_host(ctx, mem, static_cast<uint8_t *>(leaf_images), dl, lbytes));
	SB_TRY(sb_to_host(ctx, mem, static_cast<uint8_t *>(branch_images), db, bbytes));
	return finish(ctx, mem);
}

extern "C" int sb_rtree_scan(sb_context *ctx, const sb_rtree *tree, sb_mem mem, const float *query_box4, uint32_t nq,
                             uint32_t *out_query, int64_t *out_row_id, uint64_t capacity, uint64_t *count) {
	SB_BEGIN(ctx);
	SB_ARG(tree != nullptr && count != nullptr, "tree/count");
	const float *dq;
	uint32_t *doq;
	int64_t *dor;
	SB_TRY(sb_to_device(ctx, mem, query_box4, 4 * (uint64_t)nq, &dq));
	SB_TRY(sb_out_buffer(ctx, mem, out_query, capacity, &doq));
	SB_TRY(sb_out_buffer(ctx, mem, out_row_id, capacity, &dor));
	SB_TRY(sbi_rtree_scan(ctx, tree, dq, nq, doq, dor, capacity, count));
	SB_TRY(sb_to_host(ctx, mem, out_query, doq, *count));
	SB_TRY(sb_to_host(ctx, mem, out_row_id, dor, *count));
	return finish(ctx, mem);
}

// ---- K4b / K5 -----------------------------------------------------------------------------------------------------
extern "C" int sb_flat_build(sb_context *ctx, sb_mem mem, const float *box4, uint32_t n, uint32_t node_size,
                             sb_flat **out) {
	SB_BEGIN(ctx);
	SB_ARG(out != nullptr, "out");
	*out = nullptr;
	const float *db;
	SB_TRY(sb_to_device(ctx, mem, box4, 4 * (uint64_t)n, &db));
	sb_flat *t = nullptr;
	int s = sbi_flat_build(ctx, db, n, node_size, &t);
	if (s != SB_OK) {
		if (t) {
			sb_flat_destroy(t);
		}
		return s;
	}
	*out = t;
	return finish(ctx, mem);
}

extern "C" void sb_flat_destroy(sb_flat *tree) {
	if (!tree) {
		return;
	}
	if (tree->owns && tree->buffer) {
		sb_owner_free(tree->owner, tree->device, tree->buffer); // see sb_rtree_destroy
	}
	delete tree;
}
extern "C" uint32_t sb_flat_item_count(const sb_flat *tree) {
	return tree ? tree->h.item_count : 0;
}
extern "C" uint32_t sb_flat_node_count(const sb_flat *tree) {
	return tree ? tree->h.nodes : 0;
}
extern "C" uint32_t sb_flat_num_layers(const sb_flat *tree) {
	return tree ? tree->h.nlayers : 0;
}
extern "C" const uint32_t *sb_flat_layer_bounds(const sb_flat *tree) {
	return tree ? tree->h.layer_bounds : nullptr;
}
extern "C" int sb_flat_device_buffer(const sb_flat *tree, void **dev_ptr, uint64_t *bytes) {
	SB_ARG(tree != nullptr && dev_ptr != nullptr && bytes != nullptr, "tree/out");
	*dev_ptr = tree->buffer;
	*bytes = tree->bytes;
	return SB_OK;
}

extern "C" int sb_flat_attach(sb_context *ctx, void *dev_ptr, uint64_t bytes, sb_flat **out) {
	SB_ARG(ctx != nullptr && dev_ptr != nullptr && out != nullptr, "ctx/dev_ptr/out");
	SB_ARG(bytes >= sizeof(sb_flat_header), "buffer too small");
	SB_CUDA(cudaSetDevice(ctx->device));
	sb_flat *t = new sb_flat();
	SB_CUDA(cudaMemcpyAsync(&t->h, dev_ptr, sizeof(t->h), cudaMemcpyDeviceToHost, ctx->stream));
	SB_CUDA(cudaStreamSynchronize(ctx->stream));
	// the invariants sbi_flat_build establishes (the probe kernels index layer_bounds[16], a 320-entry shared stack
	// and 32-lane node loads with them): a stale or foreign buffer must fail here, not fault in a kernel
	bool ok = t->h.magic == SB_FLAT_MAGIC && t->h.node_size >= 2 && t->h.node_size <= 32 && t->h.nlayers >= 1 &&
	          t->h.nlayers <= 16 && t->h.nodes < (1u << 27) &&
	          sizeof(sb_flat_header) + (uint64_t)t->h.nodes * (sizeof(float4) + sizeof(uint32_t)) <= bytes;
	if (ok) {
		ok = t->h.layer_bounds[0] == t->h.item_count && t->h.layer_bounds[t->h.nlayers - 1] == t->h.nodes;
		for (uint32_t l = 1; ok && l < t->h.nlayers; l++) {
			ok = t->h.layer_bounds[l] > t->h.layer_bounds[l - 1];
		}
	}
	if (!ok) {
		delete t;
		sb_set_error("sb_flat_attach: not a FlatRTree buffer (magic / node size / layer bounds)");
		return SB_ERR_ARG;
	}
	t->device = ctx->device;
	t->buffer = dev_ptr;
	t->bytes = bytes;
	t->owns = false;
	*out = t;
	return SB_OK;
}

extern "C" int sb_flat_export(sb_context *ctx, const sb_flat *tree, sb_mem mem, float *out_box4, uint32_t *out_idx) {
	SB_ARG(ctx != nullptr && tree != nullptr, "ctx/tree");
	SB_CUDA(cudaSetDevice(ctx->device));
	const cudaMemcpyKind kind = mem == SB_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
	if (tree->h.nodes) {
		SB_CUDA(cudaMemcpyAsync(out_box4, tree->boxes(), sizeof(float4) * (size_t)tree->h.nodes, kind, ctx->stream));
		SB_CUDA(cudaMemcpyAsync(out_idx, tree->idx(), sizeof(uint32_t) * (size_t)tree->h.nodes, kind, ctx->stream));
	}
	return finish(ctx, mem);
}

extern "C" int sb_flat_probe(sb_context *ctx, const sb_flat *tree, sb_mem mem, const float *probe_box4, uint64_t m,
                             uint32_t *out_probe, uint32_t *out_build, uint64_t capacity, uint64_t *count) {
	SB_BEGIN(ctx);
	SB_ARG(tree != nullptr && count != nullptr, "tree/count");
	const float *dq;
	uint32_t *dop, *dob;
	SB_TRY(sb_to_device(ctx, mem, probe_box4, 4 * m, &dq));
	SB_TRY(sb_out_buffer(ctx, mem, out_probe, capacity, &dop));
	SB_TRY(sb_out_buffer(ctx, mem, out_build, capacity, &dob));
	sb_probe_args a;
	a.probe_box = reinterpret_cast<const float4 *>(dq);
	a.m = m;
	SB_TRY(sbi_flat_probe(ctx, tree, a, dop, dob, capacity, count));
	SB_TRY(sb_to_host(ctx, mem, out_probe, dop, *count));
	SB_TRY(sb_to_host(ctx, mem, out_build, dob, *count));
	return finish(ctx, mem);
}

extern "C" int sb_flat_probe_points(sb_context *ctx, const sb_flat *tree, sb_mem mem, sb_predicate pred,
                                    const double *px, const double *py, uint64_t n_points, uint32_t probe_index_base,
                                    const double *bbox2d4, uint32_t *out_probe, uint32_t *out_build, uint64_t capacity,
                                    uint64_t *count) {
	SB_BEGIN(ctx);
	SB_ARG(tree != nullptr && count != nullptr, "tree/count");
	const double *dx, *dy, *db = nullptr;
	uint32_t *dop, *dob;
	SB_TRY(sb_to_device(ctx, mem, px, n_points, &dx));
	SB_TRY(sb_to_device(ctx, mem, py, n_points, &dy));
	if (bbox2d4) {
		SB_TRY(sb_to_device(ctx, mem, bbox2d4, 4 * (uint64_t)tree->h.item_count, &db));
	}
	SB_TRY(sb_out_buffer(ctx, mem, out_probe, capacity, &dop));
	SB_TRY(sb_out_buffer(ctx, mem, out_build, capacity, &dob));
	sb_probe_args a;
	a.px = dx;
	a.py = dy;
	a.m = n_points;
	a.index_base = probe_index_base;
	a.build_box2d = reinterpret_cast<const double4 *>(db);
	a.pred = db ? (int)pred : -1;
	SB_TRY(sbi_flat_probe(ctx, tree, a, dop, dob, capacity, count));
	SB_TRY(sb_to_host(ctx, mem, out_probe, dop, *count));
	SB_TRY(sb_to_host(ctx, mem, out_build, dob, *count));
	return finish(ctx, mem);
}

// ---- K6 ---------------------------------------------------------------------------------------------------------
extern "C" int sb_refine_point_box(sb_context *ctx, sb_mem mem, sb_predicate pred, const double *px, const double *py,
                                   const double *bmin_x, const double *bmin_y, const double *bmax_x,
                                   const double *bmax_y, uint32_t *pair_probe, uint32_t *pair_build, uint64_t *count) {
	SB_BEGIN(ctx);
	SB_ARG(count != nullptr, "count");
	SB_ARG(mem == SB_DEVICE, "refine kernels take device pointers (sizes of the coordinate columns are not known here)");
	return sbi_refine_point_box(ctx, pred, px, py, bmin_x, bmin_y, bmax_x, bmax_y, pair_probe, pair_build, *count,
	                            count);
}

// exact < 0: the native winding routine; otherwise the GEOS-route predicate `exact` (sb_predicate)
static int refine_point_polygon(sb_context *ctx, sb_mem mem, int exact, const double *px, const double *py,
                                const double *vx, const double *vy, const uint64_t *ring_vert_off,
                                const uint64_t *poly_ring_off, uint64_t n_polys, uint64_t n_rings, uint64_t n_verts,
                                uint64_t n_points, uint32_t *pair_probe, uint32_t *pair_build, uint64_t *count) {
	SB_BEGIN(ctx);
	SB_ARG(count != nullptr, "count");
	const uint64_t n = *count;
	const double *dpx, *dpy, *dvx, *dvy;
	const uint64_t *dro, *dpo;
	SB_TRY(sb_to_device(ctx, mem, px, n_points, &dpx));
	SB_TRY(sb_to_device(ctx, mem, py, n_points, &dpy));
	SB_TRY(sb_to_device(ctx, mem, vx, n_verts, &dvx));
	SB_TRY(sb_to_device(ctx, mem, vy, n_verts, &dvy));
	SB_TRY(sb_to_device(ctx, mem, ring_vert_off, n_rings + 1, &dro));
	SB_TRY(sb_to_device(ctx, mem, poly_ring_off, n_polys + 1, &dpo));
	uint32_t *dpp = pair_probe, *dpb = pair_build;
	if (mem == SB_HOST && n) {
		SB_TRY(ctx->scratch_t(n, &dpp));
		SB_TRY(ctx->scratch_t(n, &dpb));
		SB_CUDA(cudaMemcpyAsync(dpp, pair_probe, n * 4, cudaMemcpyHostToDevice, ctx->stream));
		SB_CUDA(cudaMemcpyAsync(dpb, pair_build, n * 4, cudaMemcpyHostToDevice, ctx->stream));
	}
	if (exact < 0) {
		SB_TRY(sbi_refine_point_polygon(ctx, dpx, dpy, dvx, dvy, dro, dpo, dpp, dpb, n, count));
	} else {
		SB_TRY(sbi_refine_point_polygon_exact(ctx, exact, dpx, dpy, dvx, dvy, dro, dpo, dpp, dpb, n, count));
	}
	SB_TRY(sb_to_host(ctx, mem, pair_probe, dpp, *count));
	SB_TRY(sb_to_host(ctx, mem, pair_build, dpb, *count));
	return finish(ctx, mem);
}

extern "C" int sb_refine_point_polygon(sb_context *ctx, sb_mem mem, const double *px, const double *py,
                                       const double *vx, const double *vy, const uint64_t *ring_vert_off,
                                       const uint64_t *poly_ring_off, uint64_t n_polys, uint64_t n_rings,
                                       uint64_t n_verts, uint64_t n_points, uint32_t *pair_probe, uint32_t *pair_build,
                                       uint64_t *count) {
	return refine_point_polygon(ctx, mem, -1, px, py, vx, vy, ring_vert_off, poly_ring_off, n_polys, n_rings, n_verts,
	                            n_points, pair_probe, pair_build, count);
}

extern "C" int sb_refine_point_polygon_exact(sb_context *ctx, sb_mem mem, sb_predicate pred, const double *px,
                                             const double *py, const double *vx, const double *vy,
                                             const uint64_t *ring_vert_off, const uint64_t *poly_ring_off,
                                             uint64_t n_polys, uint64_t n_rings, uint64_t n_verts, uint64_t n_points,
                                             uint32_t *pair_probe, uint32_t *pair_build, uint64_t *count) {
	SB_ARG(pred == SB_PRED_INTERSECTS || pred == SB_PRED_WITHIN, "pred");
	return refine_point_polygon(ctx, mem, (int)pred, px, py, vx, vy, ring_vert_off, poly_ring_off, n_polys, n_rings,
	                            n_verts, n_points, pair_probe, pair_build, count);
}

extern "C" int sb_refine_dwithin(sb_context *ctx, sb_mem mem, const double *ax, const double *ay, const double *bx,
                                 const double *by, double radius, uint32_t *pair_a, uint32_t *pair_b, double *out_dist,
                                 uint64_t *count) {
	SB_BEGIN(ctx);
	SB_ARG(count != nullptr, "count");
	SB_ARG(mem == SB_DEVICE, "refine kernels take device pointers (sizes of the coordinate columns are not known here)");
	return sbi_refine_dwithin(ctx, ax, ay, bx, by, radius, pair_a, pair_b, out_dist, *count, count);
}

extern "C" int sb_distance_points(sb_context *ctx, sb_mem mem, const double *ax, const double *ay, const double *bx,
                                  const double *by, uint64_t n, double *out) {
	SB_BEGIN(ctx);
	const double *a, *b, *c, *d;
	double *dout;
	SB_TRY(sb_to_device(ctx, mem, ax, n, &a));
	SB_TRY(sb_to_device(ctx, mem, ay, n, &b));
	SB_TRY(sb_to_device(ctx, mem, bx, n, &c));
	SB_TRY(sb_to_device(ctx, mem, by, n, &d));
	SB_TRY(sb_out_buffer(ctx, mem, out, n, &dout));
	SB_TRY(sbi_distance_points(ctx, a, b, c, d, n, dout));
	SB_TRY(sb_to_host(ctx, mem, out, dout, n));
	return finish(ctx, mem);
}

extern "C" int sb_intersects_box2d(sb_context *ctx, sb_mem mem, const double *a_min_x, const double *a_min_y,
                                   const double *a_max_x, const double *a_max_y, const double *b_min_x,
                                   const double *b_min_y, const double *b_max_x, const double *b_max_y,
                                   uint32_t stride, const uint8_t *valid_a, const uint8_t *valid_b, uint64_t n,
                                   uint8_t *out) {
	SB_BEGIN(ctx);
	SB_ARG(stride == 1 || stride == 4, "stride must be 1 (SoA columns) or 4 (AoS rows)");
	SB_ARG(out != nullptr || n == 0, "out");
	const double *src[8] = {a_min_x, a_min_y, a_max_x, a_max_y, b_min_x, b_min_y, b_max_x, b_max_y};
	const double *dev[8];
	if (stride == 4 && mem == SB_HOST) {
		// AoS rows: stage each box array once (columns are the four interleaved offsets of one buffer)
		for (int s = 0; s < 2; s++) {
			const double *base = src[4 * s];
			for (int c = 1; c < 4; c++) {
				SB_ARG(src[4 * s + c] == base + c, "stride 4 expects box4 + 0..3");
			}
			const double *d;
			SB_TRY(sb_to_device(ctx, mem, base, 4 * n, &d));
			for (int c = 0; c < 4; c++) {
				dev[4 * s + c] = d + c;
			}
		}
	} else {
		for (int c = 0; c < 8; c++) {
			SB_TRY(sb_to_device(ctx, mem, src[c], stride == 1 ? n : 4 * n - (c & 3), &dev[c]));
		}
	}
	const uint8_t *va = nullptr, *vb = nullptr;
	if (valid_a) {
		SB_TRY(sb_to_device(ctx, mem, valid_a, n, &va));
	}
	if (valid_b) {
		SB_TRY(sb_to_device(ctx, mem, valid_b, n, &vb));
	}
	uint8_t *dout;
	SB_TRY(sb_out_buffer(ctx, mem, out, n, &dout));
	SB_TRY(sbi_box2d_intersects(ctx, dev, dev + 4, stride, va, vb, n, dout));
	SB_TRY(sb_to_host(ctx, mem, out, dout, n));
	return finish(ctx, mem);
}

extern "C" int sb_distance_point_linestring(sb_context *ctx, sb_mem mem, const double *px, const double *py,
                                            const uint64_t *list_entries, uint64_t n, const double *vx,
                                            const double *vy, uint64_t n_vertices, double *out) {
	SB_BEGIN(ctx);
	const double *dx, *dy, *dvx, *dvy;
	const uint64_t *dl;
	double *dout;
	if (mem == SB_HOST) {
		for (uint64_t i = 0; i < n; i++) {
			const uint64_t off = list_entries[2 * i], len = list_entries[2 * i + 1];
			SB_ARG(off <= n_vertices && len <= n_vertices - off, "list entry outside the vertex columns");
		}
	}
	SB_TRY(sb_to_device(ctx, mem, px, n, &dx));
	SB_TRY(sb_to_device(ctx, mem, py, n, &dy));
	SB_TRY(sb_to_device(ctx, mem, list_entries, 2 * n, &dl));
	SB_TRY(sb_to_device(ctx, mem, vx, n_vertices, &dvx));
	SB_TRY(sb_to_device(ctx, mem, vy, n_vertices, &dvy));
	SB_TRY(sb_out_buffer(ctx, mem, out, n, &dout));
	SB_TRY(sbi_distance_point_linestring(ctx, dx, dy, dl, dvx, dvy, n, dout));
	SB_TRY(sb_to_host(ctx, mem, out, dout, n));
	return finish(ctx, mem);
}

// Host-pointer join of a large probe column: the probe column is streamed through the device in chunks so that the
// H2D copy of chunk k+1, the probe of chunk k and the D2H copy of chunk k-1's pairs overlap (three streams, PCIe is
// full duplex). The call is bound by the larger of the two copy directions instead of their sum plus the kernels.
// shared_total != nullptr (sb_join_points_boxes_multi): several devices write one pair list; a chunk's place in the
// caller's arrays is reserved from the shared counter once its pair count is known, row ids start at index_base.
static int join_points_boxes_pipelined(sb_context *ctx, sb_predicate pred, const double *px, const double *py,
                                       uint64_t n_points, const sb_flat *tree, const double4 *box2d, uint32_t n_boxes,
                                       uint32_t *out_probe, uint32_t *out_build, uint64_t capacity, uint64_t *count,
                                       uint64_t index_base = 0, std::atomic<uint64_t> *shared_total = nullptr,
                                       uint64_t chunk_rows = 0) {
	const uint64_t chunk = chunk_rows ? chunk_rows : ctx->pipe_chunk;
	// a chunk's pairs must fit one output slot; a probe row pairs with at most n_boxes build rows. When a chunk
	// overflows its slot it is split (below), so any positive slot size >= n_boxes terminates.
	uint64_t slot_cap = 12 * chunk;
	if (slot_cap < n_boxes) {
		slot_cap = n_boxes;
	}
	// Input slots: as many chunks as fit a staging budget (all of them at the BASELINE sizes), so that the H2D copies run
	// ahead at full rate and are over early: the D2H of the pair lists (4 x the input bytes at C3) then has the link to
	// itself. (With two input slots the H2D traffic was spread over the whole call and the duplex interference cost
	// ~15 % of the D2H rate: scripts/diag_pcie.py.) Output slots: two, the pairs leave as fast as the link allows.
	const uint64_t nchunks = (n_points + chunk - 1) / chunk;
	size_t free_b = 0, total_b = 0;
	SB_CUDA(cudaMemGetInfo(&free_b, &total_b));
	uint64_t in_slots = (uint64_t)(free_b / 4) / (chunk * 2 * sizeof(double));
	in_slots = in_slots < 2 ? 2 : in_slots;
	in_slots = in_slots > nchunks ? (nchunks ? nchunks : 1) : in_slots;
	std::vector<double *> dx(in_slots), dy(in_slots);
	std::vector<cudaEvent_t> h2d_done(in_slots), in_free(in_slots);
	uint32_t *dp[2], *db[2];
	cudaEvent_t d2h_done[2], probe_done[2];
	for (uint64_t k = 0; k < in_slots; k++) {
		SB_TRY(ctx->scratch_t(chunk, &dx[k]));
		SB_TRY(ctx->scratch_t(chunk, &dy[k]));
		SB_CUDA(cudaEventCreateWithFlags(&h2d_done[k], cudaEventDisableTiming));
		SB_CUDA(cudaEventCreateWithFlags(&in_free[k], cudaEventDisableTiming));
	}
	// A bulk D2H copy whose HOST address is not 4 KiB-aligned runs ~6 % slower (52.2 vs 55.7 GB/s, scripts/diag_e2e.py).
	// The pairs of a chunk land at an arbitrary offset of the caller's arrays, so each slot keeps PAGE_WORDS of slack:
	// the probe writes its pairs at the slot offset that is congruent to the host address, the partial first page goes
	// out as a small copy and the rest starts on a page boundary on both sides.
	const uint64_t PAGE_WORDS = 1024;
	auto page_align = [&](uint32_t *p) {
		return reinterpret_cast<uint32_t *>((reinterpret_cast<uintptr_t>(p) + 4095) & ~(uintptr_t)4095);
	};
	for (int k = 0; k < 2; k++) {
		SB_TRY(ctx->scratch_t(slot_cap + 2 * PAGE_WORDS, &dp[k]));
		SB_TRY(ctx->scratch_t(slot_cap + 2 * PAGE_WORDS, &db[k]));
		dp[k] = page_align(dp[k]);
		db[k] = page_align(db[k]);
		SB_CUDA(cudaEventCreateWithFlags(&d2h_done[k], cudaEventDisableTiming));
		SB_CUDA(cudaEventCreateWithFlags(&probe_done[k], cudaEventDisableTiming));
	}
	// the slots were just carved from the arena on ctx->stream: the side streams may touch them after this point
	SB_CUDA(cudaEventRecord(probe_done[0], ctx->stream));
	SB_CUDA(cudaStreamWaitEvent(ctx->h2d_stream, probe_done[0], 0));
	SB_CUDA(cudaStreamWaitEvent(ctx->d2h_stream, probe_done[0], 0));
	for (uint64_t k = 0; k < in_slots; k++) {
		SB_CUDA(cudaEventRecord(in_free[k], ctx->stream));
	}
	const sb_arena::mark_t mark = ctx->arena.mark();
	int status = SB_OK;
	uint64_t total = 0;
	// option "trace" bit 1: a timeline of the three streams (timing events, printed when the call ends)
	struct tl_event {
		cudaEvent_t ev;
		char what;
		uint64_t chunk;
	};
	std::vector<tl_event> tl;
	auto tl_mark = [&](cudaStream_t st, char what, uint64_t c) {
		if (ctx->trace & 2) {
			cudaEvent_t e;
			if (cudaEventCreate(&e) == cudaSuccess) {
				cudaEventRecord(e, st);
				tl.push_back({e, what, c});
			}
		}
	};
	tl_mark(ctx->stream, '0', 0);
	auto issue_h2d = [&](uint64_t c) -> int {
		const uint64_t islot = c % in_slots;
		const uint64_t beg = c * chunk, len = (n_points - beg < chunk) ? n_points - beg : chunk;
		// the slot's previous occupant (chunk c - in_slots) must have been probed
		SB_CUDA(cudaStreamWaitEvent(ctx->h2d_stream, in_free[islot], 0));
		tl_mark(ctx->h2d_stream, 'h', c);
		SB_CUDA(cudaMemcpyAsync(dx[islot], px + beg, len * sizeof(double), cudaMemcpyHostToDevice, ctx->h2d_stream));
		SB_CUDA(cudaMemcpyAsync(dy[islot], py + beg, len * sizeof(double), cudaMemcpyHostToDevice, ctx->h2d_stream));
		SB_CUDA(cudaEventRecord(h2d_done[islot], ctx->h2d_stream));
		tl_mark(ctx->h2d_stream, 'H', c);
		return SB_OK;
	};
	uint64_t next_h2d = 0;
	// The chunk loop runs inside a lambda: SB_CUDA / SB_TRY leave the LAMBDA on an error, and every exit then passes
	// through the three-stream drain and the event destruction below (no copy may still read the caller's columns or
	// write its pair arrays when the call returns, and the next call resets the arena slots those copies use).
	auto run_chunks = [&]() -> int {
	for (uint64_t c = 0; c < nchunks && status == SB_OK; c++) {
		const int slot = (int)(c & 1);
		const uint64_t islot = c % in_slots;
		const uint64_t beg = c * chunk, len = (n_points - beg < chunk) ? n_points - beg : chunk;
		while (next_h2d < nchunks && next_h2d < c + in_slots && status == SB_OK) {
			status = issue_h2d(next_h2d++);
		}
		if (status != SB_OK) {
			break;
		}
		SB_CUDA(cudaStreamWaitEvent(ctx->stream, h2d_done[islot], 0));
		SB_CUDA(cudaStreamWaitEvent(ctx->stream, d2h_done[slot], 0)); // pairs of chunk c-2 have left the slot
		// probe the chunk; split it when its pairs do not fit the slot
		uint64_t done = 0;
		uint64_t piece = len;
		while (done < len && status == SB_OK) {
			if (piece > len - done) {
				piece = len - done;
			}
			sb_probe_args a;
			a.px = dx[islot] + done;
			a.py = dy[islot] + done;
			a.m = piece;
			a.index_base = (uint32_t)(index_base + beg + done);
			a.build_box2d = box2d;
			a.pred = (int)pred;
			uint64_t got = 0;
			ctx->arena.rewind(mark);
			tl_mark(ctx->stream, 'p', c);
			// (shared list: the destination is only known after the probe; the slot is then written at its aligned start
			// and only the HOST side of the copy is split at a page boundary)
			uint64_t lead_p = shared_total ? 0 : (reinterpret_cast<uintptr_t>(out_probe + total) >> 2) % PAGE_WORDS;
			uint64_t lead_b = shared_total ? 0 : (reinterpret_cast<uintptr_t>(out_build + total) >> 2) % PAGE_WORDS;
			uint32_t *sp = dp[slot] + lead_p, *sb = db[slot] + lead_b;
			int s = sbi_flat_probe(ctx, tree, a, sp, sb, slot_cap, &got);
			tl_mark(ctx->stream, 'P', c);
			if (s == SB_ERR_CAPACITY && piece > 1) {
				piece = (piece + 1) / 2;
				continue;
			}
			if (s != SB_OK) {
				status = s;
				break;
			}
			// sbi_flat_probe returned => the pairs are complete in the slot (it synchronises to read the count)
			uint64_t dst = total;
			if (shared_total) {
				dst = shared_total->fetch_add(got);
				lead_p = (reinterpret_cast<uintptr_t>(out_probe + dst) >> 2) % PAGE_WORDS;
				lead_b = (reinterpret_cast<uintptr_t>(out_build + dst) >> 2) % PAGE_WORDS;
			}
			if (dst + got <= capacity) {
				tl_mark(ctx->d2h_stream, 'd', c);
				auto copy_out = [&](uint32_t *host, const uint32_t *dev, uint64_t lead) -> int {
					uint64_t head = (PAGE_WORDS - lead) % PAGE_WORDS;
					head = head < got ? head : got;
					if (head) {
						SB_CUDA(cudaMemcpyAsync(host, dev, head * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->d2h_stream));
					}
					if (got > head) {
						SB_CUDA(cudaMemcpyAsync(host + head, dev + head, (got - head) * sizeof(uint32_t), cudaMemcpyDeviceToHost,
						                        ctx->d2h_stream));
					}
					return SB_OK;
				};
				SB_TRY(copy_out(out_probe + dst, sp, lead_p));
				SB_TRY(copy_out(out_build + dst, sb, lead_b));
				tl_mark(ctx->d2h_stream, 'D', c);
			}
			total += got;
			done += piece;
			if (done < len) {
				// the same slot is written again: wait for this piece's copy-out
				SB_CUDA(cudaEventRecord(d2h_done[slot], ctx->d2h_stream));
				SB_CUDA(cudaStreamWaitEvent(ctx->stream, d2h_done[slot], 0));
			}
		}
		SB_CUDA(cudaEventRecord(d2h_done[slot], ctx->d2h_stream));
		SB_CUDA(cudaEventRecord(probe_done[slot], ctx->stream));
		SB_CUDA(cudaEventRecord(in_free[islot], ctx->stream));
	}
	return SB_OK;
	};
	{
		const int rs = run_chunks();
		if (status == SB_OK) {
			status = rs;
		}
	}
	cudaStreamSynchronize(ctx->h2d_stream);
	cudaStreamSynchronize(ctx->d2h_stream);
	cudaStreamSynchronize(ctx->stream);
	for (size_t k = 0; k < tl.size(); k++) {
		float ms = 0;
		cudaEventElapsedTime(&ms, tl[0].ev, tl[k].ev);
		fprintf(stderr, "[sb pipeline] %c %3llu %9.3f ms\n", tl[k].what, (unsigned long long)tl[k].chunk, ms);
	}
	for (size_t k = 0; k < tl.size(); k++) {
		cudaEventDestroy(tl[k].ev);
	}
	for (uint64_t k = 0; k < in_slots; k++) {
		cudaEventDestroy(h2d_done[k]);
		cudaEventDestroy(in_free[k]);
	}
	for (int k = 0; k < 2; k++) {
		cudaEventDestroy(d2h_done[k]);
		cudaEventDestroy(probe_done[k]);
	}
	*count = total;
	if (status == SB_OK && !shared_total && total > capacity) {
		sb_set_error("join: %llu pairs, capacity %llu", (unsigned long long)total, (unsigned long long)capacity);
		return SB_ERR_CAPACITY;
	}
	return status;
}

// ---- whole joins -------------------------------------------------------------------------------------------------
extern "C" int sb_join_points_boxes(sb_context *ctx, sb_mem mem, sb_predicate pred, const double *px, const double *py,
                                    uint64_t n_points, const double *bmin_x, const double *bmin_y,
                                    const double *bmax_x, const double *bmax_y, uint32_t n_boxes, uint32_t *out_probe,
                                    uint32_t *out_build, uint64_t capacity, uint64_t *count) {
	SB_BEGIN(ctx);
	SB_ARG(count != nullptr, "count");
	*count = 0;
	if (n_boxes == 0 || n_points == 0) {
		return SB_OK;
	}
	const bool pipelined = mem == SB_HOST && n_points >= 2ull * ctx->pipe_chunk;
	const double *dx = nullptr, *dy = nullptr, *b0, *b1, *b2, *b3;
	SB_TRY(sb_to_device(ctx, mem, bmin_x, (uint64_t)n_boxes, &b0));
	SB_TRY(sb_to_device(ctx, mem, bmin_y, (uint64_t)n_boxes, &b1));
	SB_TRY(sb_to_device(ctx, mem, bmax_x, (uint64_t)n_boxes, &b2));
	SB_TRY(sb_to_device(ctx, mem, bmax_y, (uint64_t)n_boxes, &b3));
	uint32_t *dop = nullptr, *dob = nullptr;
	if (!pipelined) {
		SB_TRY(sb_to_device(ctx, mem, px, n_points, &dx));
		SB_TRY(sb_to_device(ctx, mem, py, n_points, &dy));
		SB_TRY(sb_out_buffer(ctx, mem, out_probe, capacity, &dop));
		SB_TRY(sb_out_buffer(ctx, mem, out_build, capacity, &dob));
	}
	SB_TRACE(ctx, "join: begin (inputs staged)");
	// build side: BOX_2D -> BOX_2DF (cached bbox of the envelope polygons) -> FlatRTree(node_size 32, :553)
	float *fbox = nullptr;
	double *box2d = nullptr;
	SB_TRY(ctx->scratch_t(4 * (uint64_t)n_boxes, &fbox));
	SB_TRY(ctx->scratch_t(4 * (uint64_t)n_boxes, &box2d));
	SB_TRY(sbi_bbox_box2d(ctx, b0, b1, b2, b3, n_boxes, fbox));
	SB_TRY(sbi_pack_box2d(ctx, b0, b1, b2, b3, n_boxes, box2d));
	sb_flat *tree = nullptr;
	SB_TRACE(ctx, "join: bbox");
	int s = sbi_flat_build(ctx, fbox, n_boxes, 32, &tree);
	SB_TRACE(ctx, "join: tree build");
	if (s == SB_OK && pipelined) {
		s = join_points_boxes_pipelined(ctx, pred, px, py, n_points, tree, reinterpret_cast<const double4 *>(box2d),
		                                n_boxes, out_probe, out_build, capacity, count);
	} else if (s == SB_OK) {
		sb_probe_args a;
		a.px = dx;
		a.py = dy;
		a.m = n_points;
		a.build_box2d = reinterpret_cast<const double4 *>(box2d);
		a.pred = (int)pred;
		s = sbi_flat_probe(ctx, tree, a, dop, dob, capacity, count);
		if (s == SB_OK) {
			s = sb_to_host(ctx, mem, out_probe, dop, *count);
		}
		if (s == SB_OK) {
			s = sb_to_host(ctx, mem, out_build, dob, *count);
		}
		if (s == SB_OK) {
			s = finish(ctx, mem);
		} else {
			cudaStreamSynchronize(ctx->stream);
		}
	}
	SB_TRACE(ctx, "join: probe + copy out");
	if (tree) {
		cudaStreamSynchronize(ctx->stream);
		sb_flat_destroy(tree);
	}
	SB_TRACE(ctx, "join: tree free");
	return s;
}

extern "C" int sb_join_dwithin(sb_context *ctx, sb_mem mem, const double *ax, const double *ay, uint64_t na,
                               const double *bx, const double *by, uint32_t nb, double radius, uint32_t *out_a,
                               uint32_t *out_b, double *out_dist, uint64_t capacity, uint64_t *count) {
	SB_BEGIN(ctx);
	SB_ARG(count != nullptr, "count");
	SB_ARG(radius >= 0, "radius");
	*count = 0;
	const double *dax, *day, *dbx, *dby;
	SB_TRY(sb_to_device(ctx, mem, ax, na, &dax));
	SB_TRY(sb_to_device(ctx, mem, ay, na, &day));
	SB_TRY(sb_to_device(ctx, mem, bx, (uint64_t)nb, &dbx));
	SB_TRY(sb_to_device(ctx, mem, by, (uint64_t)nb, &dby));
	uint32_t *doa, *dob;
	double *dod;
	SB_TRY(sb_out_buffer(ctx, mem, out_a, capacity, &doa));
	SB_TRY(sb_out_buffer(ctx, mem, out_b, capacity, &dob));
	SB_TRY(sb_out_buffer(ctx, mem, out_dist, capacity, &dod));
	if (na == 0 || nb == 0) {
		return finish(ctx, mem);
	}
	float *fbox = nullptr;
	SB_TRY(ctx->scratch_t(4 * (uint64_t)nb, &fbox));
	SB_TRY(sbi_expand_points(ctx, dbx, dby, nb, radius, fbox));
	sb_flat *tree = nullptr;
	int s = sbi_flat_build(ctx, fbox, nb, 32, &tree);
	uint64_t cand = 0;
	if (s == SB_OK) {
		sb_probe_args a;
		a.px = dax;
		a.py = day;
		a.m = na;
		s = sbi_flat_probe(ctx, tree, a, doa, dob, capacity, &cand);
		*count = cand;
	}
	if (s == SB_OK) {
		s = sbi_refine_dwithin(ctx, dax, day, dbx, dby, radius, doa, dob, dod, cand, count);
	}
	if (s == SB_OK) {
		s = sb_to_host(ctx, mem, out_a, doa, *count);
	}
	if (s == SB_OK) {
		s = sb_to_host(ctx, mem, out_b, dob, *count);
	}
	if (s == SB_OK && out_dist) {
		s = sb_to_host(ctx, mem, out_dist, dod, *count);
	}
	cudaStreamSynchronize(ctx->stream);
	if (tree) {
		sb_flat_destroy(tree);
	}
	return s;
}

// ---- A11: LEFT / RIGHT / FULL bookkeeping --------------------------------------------------------------------------
extern "C" int sb_join_mark_matched(sb_context *ctx, sb_mem mem, const uint32_t *pair_probe, const uint32_t *pair_build,
                                    uint64_t count, uint32_t probe_index_base, uint8_t *probe_matched, uint64_t n_probe,
                                    uint8_t *build_matched, uint64_t n_build) {
	SB_BEGIN(ctx);
	SB_ARG(count == 0 || (pair_probe != nullptr && pair_build != nullptr), "pairs");
	const uint32_t *dp, *db;
	SB_TRY(sb_to_device(ctx, mem, pair_probe, count, &dp));
	SB_TRY(sb_to_device(ctx, mem, pair_build, count, &db));
	// flags accumulate: host flag arrays go to the device, are OR-ed there and come back
	const uint8_t *pm_in = nullptr, *bm_in = nullptr;
	SB_TRY(sb_to_device(ctx, mem, (const uint8_t *)probe_matched, n_probe, &pm_in));
	SB_TRY(sb_to_device(ctx, mem, (const uint8_t *)build_matched, n_build, &bm_in));
	uint8_t *pm = const_cast<uint8_t *>(pm_in), *bm = const_cast<uint8_t *>(bm_in);
	SB_TRY(sbi_mark_matched(ctx, dp, db, count, probe_index_base, pm, n_probe, bm, n_build));
	SB_TRY(sb_to_host(ctx, mem, probe_matched, pm, n_probe));
	SB_TRY(sb_to_host(ctx, mem, build_matched, bm, n_build));
	return finish(ctx, mem);
}

extern "C" int sb_pairs_checksum(sb_context *ctx, sb_mem mem, const uint32_t *pair_probe, const uint32_t *pair_build,
                                 uint64_t count, uint32_t probe_lo, uint32_t probe_hi, uint64_t *out_count,
                                 uint64_t *out_hash) {
	SB_BEGIN(ctx);
	SB_ARG(out_count != nullptr && out_hash != nullptr, "out");
	SB_ARG(count == 0 || (pair_probe != nullptr && pair_build != nullptr), "pairs");
	const uint32_t *dp, *db;
	SB_TRY(sb_to_device(ctx, mem, pair_probe, count, &dp));
	SB_TRY(sb_to_device(ctx, mem, pair_build, count, &db));
	unsigned long long *d2 = nullptr;
	SB_TRY(ctx->scratch_t(2, &d2));
	SB_TRY(sbi_pairs_checksum(ctx, dp, db, count, probe_lo, probe_hi, d2));
	uint64_t w[2];
	SB_TRY(sbi_read_words(ctx, d2, 2, w));
	*out_count = w[0];
	*out_hash = w[1];
	return SB_OK;
}

extern "C" int sb_join_unmatched_rows(sb_context *ctx, sb_mem mem, const uint8_t *matched, uint64_t n, uint32_t row_base,
                                      uint32_t *out_rows, uint64_t capacity, uint64_t *count) {
	SB_BEGIN(ctx);
	SB_ARG(count != nullptr, "count");
	SB_ARG(n == 0 || matched != nullptr, "matched");
	SB_ARG((uint64_t)row_base + n <= 0x100000000ull, "row ids must fit 32 bits");
	const uint8_t *dm;
	uint32_t *dout;
	SB_TRY(sb_to_device(ctx, mem, matched, n, &dm));
	SB_TRY(sb_out_buffer(ctx, mem, out_rows, capacity, &dout));
	const int s = sbi_unmatched_rows(ctx, dm, n, row_base, dout, capacity, count);
	if (s != SB_OK) {
		return s;
	}
	SB_TRY(sb_to_host(ctx, mem, out_rows, dout, *count));
	return finish(ctx, mem);
}

// ---- ST_Within(point, polygon) join ---------------------------------------------------------------------------------
int sbi_pip_boxes(sb_context *ctx, const double *vx, const double *vy, const uint64_t *ring_off, const uint64_t *poly_off,
                  uint64_t n_polys, float *out_box4);

static int join_points_polygons(sb_context *ctx, sb_mem mem, int exact, const double *px, const double *py,
                                uint64_t n_points, const double *vx, const double *vy, const uint64_t *ring_vert_off,
                                const uint64_t *poly_ring_off, uint64_t n_polys, uint64_t n_rings, uint64_t n_verts,
                                uint32_t *out_point, uint32_t *out_poly, uint64_t capacity, uint64_t *count) {
	SB_BEGIN(ctx);
	SB_ARG(count != nullptr, "count");
	*count = 0;
	const double *dpx, *dpy, *dvx, *dvy;
	const uint64_t *dro, *dpo;
	SB_TRY(sb_to_device(ctx, mem, px, n_points, &dpx));
	SB_TRY(sb_to_device(ctx, mem, py, n_points, &dpy));
	SB_TRY(sb_to_device(ctx, mem, vx, n_verts, &dvx));
	SB_TRY(sb_to_device(ctx, mem, vy, n_verts, &dvy));
	SB_TRY(sb_to_device(ctx, mem, ring_vert_off, n_rings + 1, &dro));
	SB_TRY(sb_to_device(ctx, mem, poly_ring_off, n_polys + 1, &dpo));
	uint32_t *dop, *dob;
	SB_TRY(sb_out_buffer(ctx, mem, out_point, capacity, &dop));
	SB_TRY(sb_out_buffer(ctx, mem, out_poly, capacity, &dob));
	if (n_points == 0 || n_polys == 0) {
		return finish(ctx, mem);
	}
	int s;
	// polygon-centric kernel when every polygon sees many probes (option "pip_path": 1 = candidate list, 2 = polygon-centric)
	const bool centric = ctx->pip_path == 2 || (ctx->pip_path == 0 && n_points >= (1u << 17) && n_points / n_polys >= 32);
	if (centric) {
		s = sbi_join_points_polygons(ctx, dpx, dpy, n_points, dvx, dvy, dro, dpo, n_polys, dop, dob, capacity, count,
		                             exact >= 0, exact >= 0 ? exact : (int)SB_PRED_WITHIN);
	} else {
		float *fbox = nullptr;
		SB_TRY(ctx->scratch_t(4 * n_polys, &fbox));
		SB_TRY(sbi_pip_boxes(ctx, dvx, dvy, dro, dpo, n_polys, fbox));
		sb_flat *tree = nullptr;
		s = sbi_flat_build(ctx, fbox, (uint32_t)n_polys, 32, &tree);
		uint64_t cand = 0;
		if (s == SB_OK) {
			sb_probe_args a;
			a.px = dpx;
			a.py = dpy;
			a.m = n_points;
			s = sbi_flat_probe(ctx, tree, a, dop, dob, capacity, &cand);
			*count = cand;
		}
		if (s == SB_OK && exact < 0) {
			s = sbi_refine_point_polygon(ctx, dpx, dpy, dvx, dvy, dro, dpo, dop, dob, cand, count);
		} else if (s == SB_OK) {
			s = sbi_refine_point_polygon_exact(ctx, exact, dpx, dpy, dvx, dvy, dro, dpo, dop, dob, cand, count);
		}
		if (tree) {
			sb_flat_destroy(tree);
		}
	}
	if (s != SB_OK) {
		return s;
	}
	SB_TRY(sb_to_host(ctx, mem, out_point, dop, *count));
	SB_TRY(sb_to_host(ctx, mem, out_poly, dob, *count));
	return finish(ctx, mem);
}

extern "C" int sb_join_points_polygons(sb_context *ctx, sb_mem mem, const double *px, const double *py, uint64_t n_points,
                                       const double *vx, const double *vy, const uint64_t *ring_vert_off,
                                       const uint64_t *poly_ring_off, uint64_t n_polys, uint64_t n_rings, uint64_t n_verts,
                                       uint32_t *out_point, uint32_t *out_poly, uint64_t capacity, uint64_t *count) {
	return join_points_polygons(ctx, mem, -1, px, py, n_points, vx, vy, ring_vert_off, poly_ring_off, n_polys, n_rings,
	                            n_verts, out_point, out_poly, capacity, count);
}

extern "C" int sb_join_points_polygons_exact(sb_context *ctx, sb_mem mem, sb_predicate pred, const double *px,
                                             const double *py, uint64_t n_points, const double *vx, const double *vy,
                                             const uint64_t *ring_vert_off, const uint64_t *poly_ring_off,
                                             uint64_t n_polys, uint64_t n_rings, uint64_t n_verts, uint32_t *out_point,
                                             uint32_t *out_poly, uint64_t capacity, uint64_t *count) {
	SB_ARG(pred == SB_PRED_INTERSECTS || pred == SB_PRED_WITHIN, "pred");
	return join_points_polygons(ctx, mem, (int)pred, px, py, n_points, vx, vy, ring_vert_off, poly_ring_off, n_polys,
	                            n_rings, n_verts, out_point, out_poly, capacity, count);
}

// One join over several devices of ONE process (the reference operator is one process,
// spatial_join_physical.cpp:695-945): device 0 builds the FlatRTree, the packed tree and the build side's exact boxes
// go to the other devices as peer copies (NVLink), every device streams a contiguous slice of the host probe column
// through its own three-stream pipeline (one host thread per device) and the pair lists land in ONE host pair list --
// each chunk reserves its range from a shared counter when its pair count is known. Row ids are global.
extern "C" int sb_join_points_boxes_multi(sb_context *const *ctxs, int n_ctx, sb_predicate pred, const double *px,
                                          const double *py, uint64_t n_points, const double *bmin_x, const double *bmin_y,
                                          const double *bmax_x, const double *bmax_y, uint32_t n_boxes,
                                          uint32_t *out_probe, uint32_t *out_build, uint64_t capacity, uint64_t *count) {
	SB_ARG(ctxs != nullptr && n_ctx >= 1 && n_ctx <= 64 && count != nullptr, "ctxs/count");
	for (int d = 0; d < n_ctx; d++) {
		SB_ARG(ctxs[d] != nullptr, "ctxs[d]");
		for (int e = 0; e < d; e++) {
			SB_ARG(ctxs[e] != ctxs[d], "contexts must be distinct");
		}
	}
	*count = 0;
	if (n_boxes == 0 || n_points == 0) {
		return SB_OK;
	}
	SB_ARG(n_points <= 0xFFFFFFFFull, "probe indices must fit 32 bits");
	sb_context *c0 = ctxs[0];
	{
		SB_BEGIN(c0);
	}
	const double *b0, *b1, *b2, *b3;
	SB_TRY(sb_to_device(c0, SB_HOST, bmin_x, (uint64_t)n_boxes, &b0));
	SB_TRY(sb_to_device(c0, SB_HOST, bmin_y, (uint64_t)n_boxes, &b1));
	SB_TRY(sb_to_device(c0, SB_HOST, bmax_x, (uint64_t)n_boxes, &b2));
	SB_TRY(sb_to_device(c0, SB_HOST, bmax_y, (uint64_t)n_boxes, &b3));
	float *fbox = nullptr;
	double *box2d = nullptr;
	SB_TRY(c0->scratch_t(4 * (uint64_t)n_boxes, &fbox));
	SB_TRY(c0->scratch_t(4 * (uint64_t)n_boxes, &box2d));
	SB_TRY(sbi_bbox_box2d(c0, b0, b1, b2, b3, n_boxes, fbox));
	SB_TRY(sbi_pack_box2d(c0, b0, b1, b2, b3, n_boxes, box2d));
	sb_flat *tree = nullptr;
	int s = sbi_flat_build(c0, fbox, n_boxes, 32, &tree);
	if (s != SB_OK) {
		if (tree) {
			sb_flat_destroy(tree);
		}
		return s;
	}
	// replicate: tree buffer + exact boxes to every other device, ordered after the build on device 0's stream
	cudaEvent_t built = nullptr;
	std::vector<sb_flat> views((size_t)n_ctx);
	std::vector<const double4 *> boxes((size_t)n_ctx, nullptr);
	views[0] = *tree;
	views[0].owns = false;
	boxes[0] = reinterpret_cast<const double4 *>(box2d);
	auto fail = [&](int code) {
		if (built) {
			cudaEventDestroy(built);
		}
		cudaSetDevice(c0->device);
		cudaStreamSynchronize(c0->stream);
		sb_flat_destroy(tree);
		return code;
	};
	if (cudaSetDevice(c0->device) != cudaSuccess || cudaEventCreateWithFlags(&built, cudaEventDisableTiming) != cudaSuccess ||
	    cudaEventRecord(built, c0->stream) != cudaSuccess) {
		sb_set_error("multi join: event setup failed");
		return fail(SB_ERR_CUDA);
	}
	for (int d = 1; d < n_ctx; d++) {
		sb_context *cd = ctxs[d];
		if (cudaSetDevice(cd->device) != cudaSuccess) {
			sb_set_error("multi join: cudaSetDevice(%d) failed", cd->device);
			return fail(SB_ERR_CUDA);
		}
		cd->arena.reset();
		void *tb = nullptr, *bb = nullptr;
		if (cd->scratch(tree->bytes, &tb) != SB_OK || cd->scratch(sizeof(double4) * (size_t)n_boxes, &bb) != SB_OK) {
			return fail(SB_ERR_CUDA);
		}
		cudaError_t e = cudaStreamWaitEvent(cd->stream, built, 0);
		if (e == cudaSuccess) {
			e = cudaMemcpyPeerAsync(tb, cd->device, tree->buffer, c0->device, tree->bytes, cd->stream);
		}
		if (e == cudaSuccess) {
			e = cudaMemcpyPeerAsync(bb, cd->device, box2d, c0->device, sizeof(double4) * (size_t)n_boxes, cd->stream);
		}
		if (e != cudaSuccess) {
			sb_set_error("multi join: peer copy to device %d failed: %s", cd->device, cudaGetErrorString(e));
			return fail(SB_ERR_CUDA);
		}
		views[d] = *tree;
		views[d].device = cd->device;
		views[d].owner = nullptr;
		views[d].buffer = tb;
		views[d].owns = false;
		boxes[d] = static_cast<const double4 *>(bb);
	}
	// contiguous probe slices, one host thread per device
	std::atomic<uint64_t> total{0};
	std::vector<int> status((size_t)n_ctx, SB_OK);
	std::vector<std::string> errors((size_t)n_ctx);
	std::vector<std::thread> workers;
	const uint64_t base = n_points / (uint64_t)n_ctx, rem = n_points % (uint64_t)n_ctx;
	auto work = [&](int d) {
		sb_context *cd = ctxs[d];
		const uint64_t beg = (uint64_t)d * base + ((uint64_t)d < rem ? (uint64_t)d : rem);
		const uint64_t len = base + ((uint64_t)d < rem ? 1 : 0);
		if (cudaSetDevice(cd->device) != cudaSuccess) {
			status[d] = SB_ERR_CUDA;
			errors[d] = "cudaSetDevice failed";
			return;
		}
		if (len == 0) {
			return;
		}
		// at least ~6 chunks per device so that copy-in, probe and copy-out overlap inside the slice
		uint64_t chunk = cd->pipe_chunk;
		while (chunk > (1u << 20) && len < 6 * chunk) {
			chunk >>= 1;
		}
		uint64_t mine = 0;
		status[d] = join_points_boxes_pipelined(cd, pred, px + beg, py + beg, len, &views[d], boxes[d], n_boxes, out_probe,
		                                        out_build, capacity, &mine, beg, &total, chunk);
		if (status[d] != SB_OK) {
			errors[d] = sb_last_error();
		}
		cudaStreamSynchronize(cd->stream);
	};
	for (int d = 1; d < n_ctx; d++) {
		workers.emplace_back(work, d);
	}
	work(0);
	for (auto &w : workers) {
		w.join();
	}
	cudaSetDevice(c0->device);
	cudaEventDestroy(built);
	built = nullptr;
	cudaStreamSynchronize(c0->stream);
	sb_flat_destroy(tree);
	*count = total.load();
	for (int d = 0; d < n_ctx; d++) {
		if (status[d] != SB_OK) {
			sb_set_error("multi join, device %d: %s", ctxs[d]->device, errors[d].c_str());
			return status[d];
		}
	}
	if (*count > capacity) {
		sb_set_error("join: %llu pairs, capacity %llu", (unsigned long long)*count, (unsigned long long)capacity);
		return SB_ERR_CAPACITY;
	}
	return SB_OK;
}

