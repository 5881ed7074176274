// internal.cuh -- device-pointer entry points shared between the .cu files (everything here takes DEVICE
// pointers and enqueues on ctx->stream; the extern "C" layer in capi.cu does the host staging).
#pragma once
#include "common.cuh"

// bbox.cu
int sbi_bbox_points(sb_context *ctx, const double *x, const double *y, uint64_t n, float *out);
int sbi_bbox_box2d(sb_context *ctx, const double *x0, const double *y0, const double *x1, const double *y1, uint64_t n,
                   float *out);
int sbi_blob_bounds(sb_context *ctx, int mode, const uint8_t *data, const uint64_t *off, uint64_t n, float *out_box,
                    uint32_t *out_key, uint8_t *valid, int32_t *status, int *any_bad);
int sbi_blob_extent(sb_context *ctx, int mode, const uint8_t *data, const uint64_t *off, uint64_t n, double *out_box,
                    uint8_t *valid, int32_t *status, const double *bounds, uint32_t *out_key, int *any_bad,
                    uint64_t avg_bytes = 0);
int sbi_hilbert_xy(sb_context *ctx, const double *x, const double *y, uint64_t n, const double *b, uint32_t *out);
int sbi_hilbert_box2d(sb_context *ctx, const double *box4, uint64_t n, const double *b, uint32_t *out);
int sbi_hilbert_box2df(sb_context *ctx, const float *box4, uint64_t n, const float *b, uint32_t *out);

// radix_sort.cu : sorts bits [begin_bit, end_bit) of the keys, stable. Ping-pongs between (keys, vals) and
// (keys_alt, vals_alt); *result_in_alt tells where the sorted data ended up. vals_iota: values start as 0..n-1
// (vals need not be initialised).
int sbi_radix_sort_u32(sb_context *ctx, uint32_t *keys, uint32_t *keys_alt, uint32_t *vals, uint32_t *vals_alt,
                       uint64_t n, int begin_bit, int end_bit, bool vals_iota, int *result_in_alt);
int sbi_radix_sort_u32_segmented(sb_context *ctx, uint32_t *keys, uint32_t *keys_alt, uint32_t *vals, uint32_t *vals_alt,
                                 uint64_t n, uint64_t seg_size, int begin_bit, int end_bit, bool vals_iota,
                                 int *result_in_alt);
int sbi_radix_sort_u64(sb_context *ctx, uint64_t *keys, uint64_t *keys_alt, uint32_t *vals, uint32_t *vals_alt,
                       uint64_t n, int begin_bit, int end_bit, bool vals_iota, int *result_in_alt);
int sbi_iota_u32(sb_context *ctx, uint32_t *out, uint64_t n);

// 24-byte RTreeEntry (rtree_node.hpp:66-75)
struct sb_entry {
	uint64_t ptr;
	float box[4];
};
static_assert(sizeof(sb_entry) == 24, "RTreeEntry must be 24 bytes");

struct sb_rtree {
	int device = 0;
	sb_context *owner = nullptr; // building context: the entries are freed in stream order on its stream
	uint32_t cap = 0;
	uint64_t slice = 0;
	int nlevels = 0;
	uint64_t counts[16] = {0};
	uint64_t offsets[16] = {0}; // entry offset of each level inside `entries`
	sb_entry *entries = nullptr; // one device allocation holding all levels
	uint64_t total = 0;
};

// FlatRTree in one device buffer: [sb_flat_header (256 B)][box_array: nodes x float4][idx_array: nodes x u32]
struct sb_flat_header {
	uint32_t magic;
	uint32_t item_count, node_size, nodes, nlayers;
	uint32_t layer_bounds[16];
	float tree_box[4];
	uint32_t pad[64 - 25];
};
static_assert(sizeof(sb_flat_header) == 256, "flat header is 256 bytes");
#define SB_FLAT_MAGIC 0x53423246u

struct sb_flat {
	int device = 0;
	sb_context *owner = nullptr; // building context: the buffer is freed in stream order on its stream
	sb_flat_header h;
	void *buffer = nullptr; // device
	uint64_t bytes = 0;
	bool owns = true;
	const float4 *boxes() const {
		return reinterpret_cast<const float4 *>(static_cast<const char *>(buffer) + sizeof(sb_flat_header));
	}
	const uint32_t *idx() const {
		return reinterpret_cast<const uint32_t *>(static_cast<const char *>(buffer) + sizeof(sb_flat_header) +
		                                          sizeof(float4) * (size_t)h.nodes);
	}
};

// where the STR build reads an entry's BOX_2DF from: the box column, or a POINT_2D column rounded outward on the fly
// (geometry_t::TryGetCachedBounds point branch + MathUtil::DoubleToFloatDown / Up, as k_bbox_points does)
struct sb_boxsrc {
	const float4 *box;
	const double *x, *y;
	template <bool POINTS>
	__device__ __forceinline__ float4 load(uint64_t i) const {
		if (!POINTS) {
			return box[i];
		}
		const double xv = x[i], yv = y[i];
		return make_float4(sb_f64_to_f32_down(xv), sb_f64_to_f32_down(yv), sb_f64_to_f32_up(xv), sb_f64_to_f32_up(yv));
	}
};

// str_partition.cu: the x order of the STR bulk load as a select + stable partition (no sort)
struct __align__(32) sb_rec32 { // one entry on its way through the leaf level: a whole 32-byte sector
	float4 box;
	int64_t row;
	uint32_t ykey;
	uint32_t pad;
};
static_assert(sizeof(sb_rec32) == 32, "partition record must be one 32-byte sector");
#define SP_MAXBINS 1025 // bins of the one-pass record partition: 2 * 512 cut keys + 1
// key_lo / key_hi: cut keys of the enclosing global order just outside a shard (-1 = none)
struct sb_str_select {
	uint32_t nsp, ncut;
	uint64_t slice;
	int64_t key_lo, key_hi;
	uint32_t *cuts;  // [ncut] sorted cut keys (device), valid after finish
	uint32_t *hist1; // [2048] histogram of the x keys' top 11 bits, to be filled between begin and finish (NULL: not needed)
	uint32_t *temp;
};
int sbi_str_select_begin(sb_context *ctx, uint64_t n, uint64_t slice, int64_t key_lo, int64_t key_hi, sb_str_select *sel);
int sbi_str_select_finish(sb_context *ctx, const uint32_t *xkeys, uint64_t n, sb_str_select *sel);
int sbi_str_bin_ids(sb_context *ctx, const uint32_t *xkeys, uint64_t n, const uint32_t *cuts, uint32_t ncut,
                    uint32_t *bins);
int sbi_str_node_partition(sb_context *ctx, const uint32_t *ykeys, uint64_t n, uint64_t slice, uint32_t cap, uint32_t *perm);
int sbi_str_partition_rec(sb_context *ctx, sb_boxsrc box, const int64_t *rows, const uint32_t *xkeys,
                          const uint32_t *ykeys, uint64_t n, const uint32_t *cuts, uint32_t ncut, sb_rec32 *rec_out,
                          uint32_t *ykey_out);

// rtree_str.cu
// px / py (optional, with box4 == NULL): build straight from a POINT_2D column
int sbi_rtree_build_str(sb_context *ctx, const float *box4, const int64_t *row_ids, uint64_t n, uint32_t cap,
                        sb_rtree **out, const double *px = nullptr, const double *py = nullptr);
int sbi_rtree_leaf_shard(sb_context *ctx, const float *box4, const int64_t *row_ids, uint64_t n, uint32_t cap,
                         uint64_t slice, uint64_t node_base, int64_t x_key_before, int64_t x_key_after, sb_entry *level0,
                         sb_entry *parents);
int sbi_rtree_from_leaf_level(sb_context *ctx, const sb_entry *level0, uint64_t n, const sb_entry *nodes_in, uint32_t cap,
                              sb_rtree **out);
int sbi_rtree_scan(sb_context *ctx, const sb_rtree *tree, const float *qbox4, uint32_t nq, uint32_t *out_query,
                   int64_t *out_row, uint64_t capacity, uint64_t *count);

int sbi_select_valid(sb_context *ctx, const uint8_t *valid, uint64_t n, const float *box4, const int64_t *rows,
                     float *obox4, int64_t *orows, uint64_t *count);
int sbi_rtree_dump(sb_context *ctx, const sb_rtree *tree, int32_t *out_level, float *out_box4, int64_t *out_row,
                   uint64_t capacity, uint64_t *count);
int sbi_fsa_layout(uint64_t segment_size, uint64_t block_size, uint64_t *segments_per_buffer, uint64_t *bitmask_count);
int sbi_rtree_pages_info(const sb_rtree *tree, uint64_t block_size, sb_rtree_pages_info *info);
int sbi_rtree_export_pages(sb_context *ctx, const sb_rtree *tree, uint64_t block_size, void *leaf_img,
                           void *branch_img, void *root_entry);

// refine.cu (native-type scalar functions)
int sbi_box2d_intersects(sb_context *ctx, const double *const a[4], const double *const b[4], uint32_t stride,
                         const uint8_t *va, const uint8_t *vb, uint64_t n, uint8_t *out);
int sbi_distance_point_linestring(sb_context *ctx, const double *px, const double *py, const uint64_t *lists,
                                  const double *vx, const double *vy, uint64_t n, double *out);

// flat_rtree.cu
int sbi_flat_build(sb_context *ctx, const float *box4, uint32_t n, uint32_t node_size, sb_flat **out);
// probe modes: boxes only (candidates), or points with a fused exact predicate against the build's BOX_2D
struct sb_probe_args {
	const float4 *probe_box = nullptr; // mode A: float probe boxes
	const double *px = nullptr, *py = nullptr; // mode B: points (float box derived in-kernel)
	const double4 *build_box2d = nullptr;      // mode B with pred >= 0: exact refinement
	int pred = -1;                             // -1 = none
	uint64_t m = 0;
	uint32_t index_base = 0;
	const uint32_t *ids = nullptr; // optional: emitted probe id = ids[i] instead of index_base + i
	bool force_walk = false;       // always take the per-probe tree walk (v1)
};
int sbi_flat_probe(sb_context *ctx, const sb_flat *tree, const sb_probe_args &args, uint32_t *out_probe,
                   uint32_t *out_build, uint64_t capacity, uint64_t *count);

// bucket_probe.cu : bulk point probe (v2), same result set as sbi_flat_probe's tree walk
int sbi_flat_probe_bucketed(sb_context *ctx, const sb_flat *tree, const sb_probe_args &args, uint32_t *out_probe,
                            uint32_t *out_build, uint64_t capacity, uint64_t *count);

// bucket_probe.cu : cell-sorted bulk point probe (v3): probes sorted by (bucket, sub-cell), warp-uniform entry lists
int sbi_flat_probe_cells(sb_context *ctx, const sb_flat *tree, const sb_probe_args &args, uint32_t *out_probe,
                         uint32_t *out_build, uint64_t capacity, uint64_t *count);

// Result of the v3 front end: probe points as 8-byte records {(ly << 16) | lx in-bucket quantised coordinate, probe
// id} grouped by cell = bucket * 1024 + sub-cell (32 x 32 buckets of 32 x 32 sub-cells over the tree box);
// cell_off[cell] .. cell_off[cell + 1] is the record range of a cell. Everything lives in the context's arena.
struct sb_cell_partition {
	const void *grid = nullptr;   // bp_grid (quantiser of the tree box)
	void *tables = nullptr;       // bp_tables
	const void *units = nullptr;  // bp_unit[n_units]: (bucket, record range <= 32768) work units
	const uint32_t *cand_pos = nullptr; // leaf positions of the tree binned by bucket (tables->cand_off)
	const void *rec_bucketed = nullptr; // records grouped by bucket only (NaN probes at the end)
	const void *rec = nullptr;          // records grouped by cell
	const uint32_t *cell_off = nullptr; // 1024 * 1024 + 1 offsets
	// want_xy: exact coordinates in bucket order (aligned with rec_bucketed) and, per cell-ordered record, its position
	// in bucket order -- the coordinates then move with two coalesced / window-local passes instead of random reads
	const void *xy_bucketed = nullptr; // double2[n]
	const uint32_t *src = nullptr;
	uint64_t n_kept = 0, n_cand = 0, n_units = 0, n_nan = 0;
};
int sbi_cell_partition(sb_context *ctx, const sb_flat *tree, const double *px, const double *py, uint64_t n,
                       uint32_t index_base, sb_cell_partition *P, bool want_xy = false);

// pip_join.cu : ST_Within(point, polygon) join, polygon-centric over the cell-sorted points
int sbi_join_points_polygons(sb_context *ctx, const double *px, const double *py, uint64_t n_points, const double *vx,
                             const double *vy, const uint64_t *ring_off, const uint64_t *poly_off, uint64_t n_polys,
                             uint32_t *out_point, uint32_t *out_poly, uint64_t capacity, uint64_t *count, int exact = 0,
                             int pred = 1);
int sbi_pip_boxes(sb_context *ctx, const double *vx, const double *vy, const uint64_t *ring_off, const uint64_t *poly_off,
                  uint64_t n_polys, float *out_box4);

// refine.cu
int sbi_pack_box2d(sb_context *ctx, const double *x0, const double *y0, const double *x1, const double *y1, uint64_t n,
                   double *out4);
int sbi_refine_point_box(sb_context *ctx, int pred, const double *px, const double *py, const double *x0,
                         const double *y0, const double *x1, const double *y1, uint32_t *pp, uint32_t *pb, uint64_t n,
                         uint64_t *d_count);
int sbi_refine_point_polygon(sb_context *ctx, const double *px, const double *py, const double *vx, const double *vy,
                             const uint64_t *ring_off, const uint64_t *poly_off, uint32_t *pp, uint32_t *pb, uint64_t n,
                             uint64_t *d_count);
int sbi_refine_point_polygon_exact(sb_context *ctx, int pred, const double *px, const double *py, const double *vx,
                                   const double *vy, const uint64_t *ring_off, const uint64_t *poly_off, uint32_t *pp,
                                   uint32_t *pb, uint64_t n, uint64_t *d_count);
int sbi_refine_dwithin(sb_context *ctx, const double *ax, const double *ay, const double *bx, const double *by,
                       double r, uint32_t *pa, uint32_t *pb, double *out_d, uint64_t n, uint64_t *d_count);
int sbi_distance_points(sb_context *ctx, const double *ax, const double *ay, const double *bx, const double *by,
                        uint64_t n, double *out);
int sbi_expand_points(sb_context *ctx, const double *x, const double *y, uint64_t n, double r, float *out_box4);

// LEFT / RIGHT / FULL join bookkeeping (refine.cu)
int sbi_mark_matched(sb_context *ctx, const uint32_t *pp, const uint32_t *pb, uint64_t n, uint32_t probe_base,
                     uint8_t *pm, uint64_t n_probe, uint8_t *bm, uint64_t n_build);
int sbi_unmatched_rows(sb_context *ctx, const uint8_t *flags, uint64_t n, uint32_t row_base, uint32_t *out,
                       uint64_t capacity, uint64_t *count);

int sbi_pairs_checksum(sb_context *ctx, const uint32_t *pp, const uint32_t *pb, uint64_t n, uint32_t lo, uint32_t hi,
                       unsigned long long *d_out2);

// read a device u64 counter back (synchronises the stream)
// transcode.cu: GEOMETRY blobs -> POINT_2D / POLYGON_2D columns
struct sb_poly_transcode {
	uint32_t *row_rings = nullptr;            // rings per row (0 for bad rows)
	unsigned long long *row_verts = nullptr;  // vertices per row, then the row's first vertex
	unsigned long long *tile_sums = nullptr, *totals = nullptr;
};
int sbi_points_from_blobs(sb_context *ctx, const uint8_t *data, const uint64_t *off, uint64_t n, double *ox, double *oy,
                          int32_t *status, int *any_bad);
int sbi_wkb_extent(sb_context *ctx, const uint8_t *data, const uint64_t *off, uint64_t n, double *out_box, uint8_t *valid,
                   int32_t *status, int *any_bad);
int sbi_polygons_count(sb_context *ctx, const uint8_t *data, const uint64_t *off, uint64_t n, sb_poly_transcode *T,
                       int32_t *status, int *any_bad, uint64_t *poly_ring_off, uint64_t *n_rings, uint64_t *n_verts);
int sbi_polygons_fill(sb_context *ctx, const uint8_t *data, const uint64_t *off, uint64_t n, const sb_poly_transcode *T,
                      const uint64_t *poly_ring_off, uint64_t *ring_vert_off, double *vx, double *vy);
int sbi_read_u64(sb_context *ctx, const uint64_t *dev, uint64_t *host);
// context.cu: counter read-back / small clears as kernels (kept off the copy engines, see there)
int sbi_read_words(sb_context *ctx, const void *dev, int nwords, uint64_t *host);
int sbi_zero(sb_context *ctx, void *dev, size_t bytes);
