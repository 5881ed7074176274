/*
 * spatial_b200.h -- C-ABI of libspatial_b200.so: the B200 (sm_100a) implementation of duckdb-spatial's
 * data-parallel hot path. This is the drop-in boundary: the DuckDB extension keeps its own ScalarFunction /
 * index / operator registrations and their bodies call these entry points on whole columns.
 * Reference paths are relative to the yuiseki/duckdb-spatial tree.
 *
 * Conventions
 *  - every function returns 0 on success, a negative sb_status otherwise; sb_last_error() gives the message of
 *    the calling thread's last failure. No C++ types, no exceptions cross this boundary.
 *  - `mem` says where ALL data pointers of that call live: SB_HOST (pageable or pinned host memory; the library
 *    stages through device memory and copies results back before returning) or SB_DEVICE (device pointers on
 *    the context's device; the call is asynchronous on the context's stream unless it returns a count).
 *  - boxes are 4 floats/doubles in the order min_x, min_y, max_x, max_y (BOX_2DF / BOX_2D child order,
 *    src/spatial/spatial_types.cpp:40-71).
 *  - blobs: a GEOMETRY column is passed as one byte buffer + n+1 byte offsets (DuckDB string_t payloads packed
 *    back to back by the host glue).
 *  - There is no CPU fallback: every entry point fails with SB_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef SPATIAL_B200_H
#define SPATIAL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SB_API __attribute__((visibility("default")))

typedef enum { SB_HOST = 0, SB_DEVICE = 1 } sb_mem;

typedef enum {
	SB_OK = 0,
	SB_ERR_CUDA = -1,     /* CUDA runtime error (message has the CUDA string) */
	SB_ERR_ARG = -2,      /* invalid argument */
	SB_ERR_CAPACITY = -3, /* caller-provided output too small; *count holds the needed size, the output arrays'
	                         contents are unspecified */
	SB_ERR_GEOMETRY = -4, /* a blob failed validation (per-row status array has details) */
	SB_ERR_EMPTY = -5     /* operation not defined on an empty geometry (ST_Hilbert(geom, bounds)) */
} sb_status;

/* per-row blob status written by the blob kernels (never a trap in-kernel) */
#define SB_ROW_OK 0
#define SB_ROW_BAD_VERSION (-1) /* geometry_properties.hpp:28-36 CheckVersion throws NotImplementedException */
#define SB_ROW_TRUNCATED (-2)   /* Cursor bounds check, util/cursor.hpp throws SerializationException */
#define SB_ROW_EMPTY 1          /* ST_Hilbert(geom,bounds) on empty input: InvalidInputException */
#define SB_ROW_WRONG_TYPE (-3)  /* GEOMETRY -> POINT_2D / POLYGON_2D cast of another geometry type: ConversionException;
                                   WKB: unsupported geometry type id (SGL_WKB_READER_UNSUPPORTED_TYPE) */
#define SB_ROW_RECURSION (-4)   /* WKB nested deeper than 128 collections (SGL_WKB_READER_RECURSION_LIMIT) */

typedef struct sb_context sb_context;
typedef struct sb_rtree sb_rtree; /* persistent RTREE index, STR bulk-loaded (CREATE INDEX ... USING RTREE) */
typedef struct sb_flat sb_flat;   /* FlatRTree of the spatial join build side */

SB_API const char *sb_last_error(void);
SB_API const char *sb_version(void);

/* One context per calling thread / GPU: owns a stream and a scratch arena. */
SB_API int sb_context_create(int device, sb_context **out);
SB_API void sb_context_destroy(sb_context *ctx);
/* adopt an external stream (e.g. torch.cuda.current_stream().cuda_stream); NULL restores the private stream, so the
 * legacy default stream has to be named as cudaStreamLegacy ((cudaStream_t)0x1) */
SB_API int sb_context_set_stream(sb_context *ctx, void *cuda_stream);
SB_API int sb_context_sync(sb_context *ctx);
/* give the context's cached device memory (scratch arena, freed trees) back to the driver. The context keeps a PRIVATE
 * stream-ordered memory pool, so nothing it caches changes the host process's default pool; sb_context_destroy trims too. */
SB_API int sb_context_trim(sb_context *ctx);
/* tuning / test knobs. "probe_path": 0 = auto, 1 = per-probe tree walk only, 2 = bucketed bulk probe only
 * (both produce the same pair set; the parity tests run every join through both). "pipe_chunk": probe rows per
 * chunk of the host-pointer join pipeline (default 8 Mi; calls with >= 2 chunks overlap H2D / probe / D2H).
 * "trace": bit 0 = print per-phase wall times to stderr (synchronises at every phase), bit 1 = print the per-chunk
 * timeline of that pipeline's three streams when the call ends (no extra synchronisation). "probe_path" 3 = the cell-sorted
 * bulk probe (v3, what auto picks), 2 = the bucketed v2 kernel. "pip_path": route of sb_join_points_polygons, 0 = auto,
 * 1 = FlatRTree candidates + sb_refine_point_polygon, 2 = polygon-centric kernel. "hist_sample": 1 (default) = the bucket
 * regions of the cell partition are sized from a 1/16 sample for inputs >= 32 Mi points (exact histogram on overflow),
 * 0 = always the exact histogram. "str_path": form of the STR leaf level, 0 = auto (select + record partition for the x
 * order, select + position partition per slice for the leaf nodes; layers above 512 slices sort bin ids and gather),
 * 1 = always the bin-id sort + gather form, 2 = record partition + segmented y sort, 3 = node partition by separate
 * kernels; all forms build the same tree (tests run every form). */
SB_API int sb_context_set_option(sb_context *ctx, const char *key, int64_t value);
/* number of kernels this context has launched so far (bench.py's gpu_launches) */
SB_API uint64_t sb_context_launches(const sb_context *ctx);
/* milliseconds the dominant kernel class spent on the GPU since the last reset, measured with CUDA events on
 * the context's stream (profiling must be enabled; adds two event records per tagged launch) */
SB_API int sb_context_profile(sb_context *ctx, int enable);
SB_API int sb_context_profile_read(sb_context *ctx, const char *tag, double *ms, uint64_t *launches);

/* ---- K1: bounding boxes --------------------------------------------------------------------------------
 * POINT_2D columns (x[], y[]) -> BOX_2DF rounded outward.
 * replaces: geometry_t::TryGetCachedBounds point branch, src/spatial/geometry/geometry_type.hpp:126-145 +
 * MathUtil::DoubleToFloatDown/Up, src/spatial/util/math.hpp:19-47 */
SB_API int sb_bbox_from_points(sb_context *ctx, sb_mem mem, const double *x, const double *y, uint64_t n,
                               float *out_box4);
/* BOX_2D columns -> BOX_2DF rounded outward (same rounding, Serde bbox write geometry_serialization.cpp:226-230) */
SB_API int sb_bbox_from_box2d(sb_context *ctx, sb_mem mem, const double *min_x, const double *min_y,
                              const double *max_x, const double *max_y, uint64_t n, float *out_box4);
/* ST_Extent_Approx(GEOMETRY): replaces ST_Extent_Approx::Execute, spatial_functions_scalar.cpp:2684-2724
 * (TryGetCachedBounds per row). valid[i]=0 -> SQL NULL. row_status may be NULL. */
SB_API int sb_bbox_from_blobs(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets, uint64_t n,
                              float *out_box4, uint8_t *out_valid, int32_t *row_status);
/* ST_Extent(GEOMETRY): replaces ST_Extent::Execute, spatial_functions_scalar.cpp:2527-2571 =
 * Serde::Deserialize (geometry_serialization.cpp:290-329) + sgl::ops::try_get_extent_xy (sgl.hpp:1408-1486) */
SB_API int sb_extent_from_blobs(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets, uint64_t n,
                                double *out_box4, uint8_t *out_valid, int32_t *row_status);

/* ST_Extent(WKB_BLOB): replaces ST_Extent::ExecuteWKB, spatial_functions_scalar.cpp:2576-2629 =
 * sgl::ops::wkb_reader_try_parse_stats (src/sgl/sgl.cpp:903-1058; nan_as_empty, 128 nesting levels). Both byte orders,
 * ISO and EWKB Z / M / SRID flags; all rings of a polygon count. valid[i] = 0 -> SQL NULL (no vertices). Rows the
 * reference fails to parse (InvalidInputException "Failed to parse WKB") get a negative row_status and the call
 * returns SB_ERR_GEOMETRY. */
SB_API int sb_extent_from_wkb(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets, uint64_t n,
                              double *out_box4, uint8_t *out_valid, int32_t *row_status);

/* ---- GEOMETRY column -> coordinate columns (what the join entry points take) ------------------------------
 * GEOMETRY -> POINT_2D: replaces PointCasts::FromGeometryCast, spatial_functions_cast.cpp:247-270 (Serde::Deserialize
 * + get_vertex_xy(0)). The reference throws on the first non-POINT / empty POINT row; here every row gets a status
 * (row_status may be NULL), bad rows yield NaN, NaN and the call returns SB_ERR_GEOMETRY (wrong type, malformed) or
 * SB_ERR_EMPTY (only empty points) after the whole column is converted. Z / M ordinates are dropped. */
SB_API int sb_points_from_blobs(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets, uint64_t n,
                                double *out_x, double *out_y, int32_t *row_status);
/* GEOMETRY -> POLYGON_2D in DuckDB list layout (the layout of sb_refine_point_polygon / sb_join_points_polygons):
 * replaces PolygonCasts::FromGeometryCast, spatial_functions_cast.cpp:473-537. poly_ring_off has n + 1 entries,
 * ring_vert_off ring_capacity + 1, vx / vy vertex_capacity. *n_rings / *n_verts always return the sizes needed;
 * SB_ERR_CAPACITY when a capacity is too small (poly_ring_off is complete even then: call once with capacities 0 to
 * size the outputs). Non-POLYGON rows: status SB_ROW_WRONG_TYPE, zero rings, SB_ERR_GEOMETRY at the end (the
 * reference throws ConversionException). POLYGON EMPTY is a valid row with zero rings. */
SB_API int sb_polygons_from_blobs(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets, uint64_t n,
                                  uint64_t *poly_ring_off, uint64_t *ring_vert_off, double *vx, double *vy,
                                  uint64_t ring_capacity, uint64_t vertex_capacity, uint64_t *n_rings,
                                  uint64_t *n_verts, int32_t *row_status);

/* ---- K2: Hilbert keys (ST_Hilbert, spatial_functions_scalar.cpp:5210-5396) -------------------------------- */
/* ST_Hilbert(x DOUBLE, y DOUBLE, bounds BOX_2D): ExecuteLonlat :5250-5268. bounds = 4 doubles (host memory). */
SB_API int sb_hilbert_xy(sb_context *ctx, sb_mem mem, const double *x, const double *y, uint64_t n,
                         const double *bounds4, uint32_t *out);
/* ST_Hilbert(box BOX_2D, bounds BOX_2D): ExecuteBox<double> :5216-5245. box4 = n x 4 doubles (AoS) */
SB_API int sb_hilbert_box2d(sb_context *ctx, sb_mem mem, const double *box4, uint64_t n, const double *bounds4,
                            uint32_t *out);
/* ST_Hilbert(box BOX_2DF, bounds BOX_2DF): ExecuteBox<float> */
SB_API int sb_hilbert_box2df(sb_context *ctx, sb_mem mem, const float *box4, uint64_t n, const float *bounds4,
                             uint32_t *out);
/* ST_Hilbert(geom GEOMETRY): ExecuteGeometry :5273-5293 (cached float box centre, f32->u32 order map) */
SB_API int sb_hilbert_geom(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets, uint64_t n,
                           uint32_t *out, uint8_t *out_valid);
/* ST_Hilbert(geom GEOMETRY, bounds BOX_2D): ExecuteGeometryWithBounds :5295-5325. Returns SB_ERR_EMPTY when a
 * row is empty (the reference throws); row_status marks the rows. */
SB_API int sb_hilbert_geom_bounds(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets,
                                  uint64_t n, const double *bounds4, uint32_t *out, int32_t *row_status);

/* ---- K3: radix sort (stable, LSD onesweep) -------------------------------------------------------------
 * replaces: DuckDB PhysicalOrder planned at rtree_index_create_logical.cpp:101-133 and FlatRTree::Sort,
 * spatial_join_physical.cpp:131-163. Sorts (key, value) pairs in place; ties keep input order. */
SB_API int sb_sort_pairs_u32(sb_context *ctx, sb_mem mem, uint32_t *keys, uint32_t *values, uint64_t n);
SB_API int sb_sort_pairs_u64(sb_context *ctx, sb_mem mem, uint64_t *keys, uint32_t *values, uint64_t n);

/* ---- K4a: persistent RTREE, STR bulk load ---------------------------------------------------------------
 * replaces: ORDER BY x-centre (rtree_index_create_logical.cpp:101-133, key = ST_Centroid::ExecuteBox<float>
 * spatial_functions_scalar.cpp:1490-1494), PhysicalCreateRTreeIndex::Sink/Finalize and BuildRTreeBottomUp,
 * src/spatial/index/rtree/rtree_index_create_physical.cpp:95-131, 136-256, 334-363.
 * Input rows must already be filtered to non-NULL, non-empty geometries (valid bbox), as the reference's plan
 * does (rtree_index_create_logical.cpp:42-72). max_node_capacity in [4, 255] (rtree_index.cpp:30-63). */
SB_API int sb_rtree_build_str(sb_context *ctx, sb_mem mem, const float *box4, const int64_t *row_ids, uint64_t n,
                              uint32_t max_node_capacity, sb_rtree **out);
/* CREATE INDEX over a POINT_2D key column: sb_bbox_from_points fused into the build (the outward-rounded BOX_2DF of a
 * point, geometry_type.hpp:126-145 + math.hpp:19-47, is computed where the build reads it and never stored: 32 bytes of
 * traffic per key less). Same tree as sb_rtree_build_str over sb_bbox_from_points(x, y), entry for entry. */
SB_API int sb_rtree_build_from_points(sb_context *ctx, sb_mem mem, const double *x, const double *y, const int64_t *row_ids,
                                      uint64_t n, uint32_t max_node_capacity, sb_rtree **out);
/* ---- multi-GPU CREATE INDEX (one process per GPU; the host side is duckdb-spatial_b200/distributed.py) ----------------
 * The reference builds the index in one process (PhysicalCreateRTreeIndex, rtree_index_create_physical.cpp:95-256); past
 * one GPU's memory the ORDER BY becomes a distributed sample sort: every rank ends up with the rows of a contiguous,
 * slice-aligned range of the global x order (ties by global input position) and builds ITS part of the leaf level; the
 * nodes' entries (one per leaf page) are gathered and the upper levels built from them. The result is the tree
 * sb_rtree_build_str builds over the whole input, node for node.
 *   sb_rtree_slice_for        the global slice size, Finalize :349-351: ceil(sqrt(ceil(n / cap))) * cap
 *   sb_rtree_build_leaf_shard leaf entries (24 B x n_shard, shard order) + one node entry per leaf page, pointer =
 *                             (LEAF_PAGE << 56) | (node_base + page); the shard must start at a multiple of slice_size
 *                             and list its rows in ascending GLOBAL input position. x_key_before / x_key_after: the
 *                             32-bit x sort key (order-preserving bits of float(min_x + max_x)) of the global order at
 *                             the shard's first rank and at the first rank after the shard, -1 at the layer's two ends
 *                             (they keep equal-y ties of the boundary slices in the single-device order)
 *   sb_rtree_from_leaf_level  the whole tree from all leaf entries (global order) and all node entries (page order) */
SB_API uint64_t sb_rtree_slice_for(uint64_t n, uint32_t max_node_capacity);
SB_API int sb_rtree_build_leaf_shard(sb_context *ctx, sb_mem mem, const float *box4, const int64_t *row_ids, uint64_t n_shard,
                                     uint32_t max_node_capacity, uint64_t slice_size, uint64_t node_base,
                                     int64_t x_key_before, int64_t x_key_after, void *out_leaf_entries,
                                     void *out_node_entries);
SB_API int sb_rtree_from_leaf_level(sb_context *ctx, sb_mem mem, const void *leaf_entries, uint64_t n,
                                    const void *node_entries, uint32_t max_node_capacity, sb_rtree **out);
/* CREATE INDEX straight from the GEOMETRY key column: TryGetCachedBounds per row (geometry_type.hpp:102-148), rows
 * without bounds dropped as the planner's filter does (rtree_index_create_logical.cpp:42-72), then the same STR bulk
 * load -- the 64.2 B/key variant of the build (32 B blob + 8 B row id in, 24 B entry out). row_ids NULL = ordinals.
 * *indexed_rows = rows that reached the index; row_status (optional) as in sb_bbox_from_blobs. */
SB_API int sb_rtree_build_from_blobs(sb_context *ctx, sb_mem mem, const uint8_t *data, const uint64_t *offsets, uint64_t n,
                                     const int64_t *row_ids, uint32_t max_node_capacity, sb_rtree **out,
                                     uint64_t *indexed_rows, int32_t *row_status);
SB_API void sb_rtree_destroy(sb_rtree *tree);
/* level 0 = leaf entries (row ids); the last level has one entry = the root entry. */
SB_API int sb_rtree_num_levels(const sb_rtree *tree);
SB_API uint64_t sb_rtree_level_count(const sb_rtree *tree, int level);
SB_API uint64_t sb_rtree_slice_size(const sb_rtree *tree);
/* Copies one level out as 24-byte RTreeEntry records {u64 pointer, float min_x, min_y, max_x, max_y}
 * (rtree_node.hpp:66-75). Node j of the level = entries [j*cap, min((j+1)*cap, count)); the host glue copies each
 * run into a FixedSizeAllocator page (rtree.cpp:28-35) and rewrites `pointer` of upper levels
 * ((type<<56) | child node index, type LEAF_PAGE=1 / BRANCH_PAGE=2) into IndexPointers. Leaf-level pointers are
 * row ids with type ROW_ID=0 (rtree.cpp:37-42). */
SB_API int sb_rtree_export_level(sb_context *ctx, const sb_rtree *tree, int level, sb_mem mem, void *out_entries);
/* rtree_index_dump('idx'): replaces the traversal of RTreeIndexDumpExecute (rtree_index_pragmas.cpp:137-227 over
 * RTreeScanner::Scan, rtree_scanner.hpp:38-83): one row (level, bounds, row_id) per entry below the root entry, in the
 * pragma's DFS order -- the root node's entries are level 0 and each entry is followed by its subtree; row_id is -1
 * where the pragma yields NULL (branch entries). out_level == NULL: only *count (= rows) is returned. */
SB_API int sb_rtree_dump(sb_context *ctx, const sb_rtree *tree, sb_mem mem, int32_t *out_level, float *out_box4,
                         int64_t *out_row_id, uint64_t capacity, uint64_t *count);
/* Persistent-index export (GetStorageInfo / checkpoint path, rtree_index.cpp:234-259): the whole tree as the block
 * images of the reference's two FixedSizeAllocators (rtree.hpp:31-36; leaf pages and branch pages), with real
 * IndexPointers (index_pointer.hpp:19-24: metadata << 56 | offset << 32 | buffer id), exactly as
 * BuildRTreeBottomUp's tree.MakePage sequence (rtree_index_create_physical.cpp:181-184, 241-246) leaves them in
 * fresh allocators: page k of an allocator = buffer k / segments_per_buffer, offset k % segments_per_buffer
 * (fixed_size_allocator.cpp:43-80, fixed_size_buffer.cpp:154-165). block_size = BlockManager::GetBlockSize()
 * (262136 for the default 256 KiB blocks). Each image is block_size bytes: bitmask_count validity words (bit set =
 * segment free), then segments of segment_size bytes = RTreeNode {u32 count, 20 B unused} + max_node_capacity
 * RTreeEntry (rtree_node.hpp:66-75, 183-189); unused entries, free segments and the slack are zero.
 * The host glue copies image b into buffers[b] of the allocator, sets its segment_count
 * (min(segments_per_buffer, pages - b * segments_per_buffer)) and calls tree.SetRoot(root_entry). */
typedef struct sb_rtree_pages_info {
	uint64_t segment_size, segments_per_buffer, bitmask_count, bitmask_offset;
	uint64_t leaf_pages, leaf_buffers, branch_pages, branch_buffers;
} sb_rtree_pages_info;
SB_API int sb_rtree_pages_layout(const sb_rtree *tree, uint64_t block_size, sb_rtree_pages_info *info);
/* leaf_images: leaf_buffers * block_size bytes, branch_images: branch_buffers * block_size bytes (`mem` says where);
 * root_entry: 24 bytes of HOST memory (one RTreeEntry; all zero for an empty tree). */
SB_API int sb_rtree_export_pages(sb_context *ctx, const sb_rtree *tree, uint64_t block_size, sb_mem mem,
                                 void *leaf_images, void *branch_images, void *root_entry);
/* Batched index scan: replaces RTreeIndex::InitializeScan/Scan (rtree_index.cpp:106-139, rtree_scanner.hpp:38-83)
 * for nq query boxes at once. Emits (query index, row id) for every leaf entry whose float box intersects the
 * query box (closed intervals, bbox.hpp:27-29). Order: grouped by nothing in particular (sort to compare).
 * capacity = number of result slots; *count = results found; SB_ERR_CAPACITY when count > capacity. */
SB_API int sb_rtree_scan(sb_context *ctx, const sb_rtree *tree, sb_mem mem, const float *query_box4, uint32_t nq,
                         uint32_t *out_query, int64_t *out_row_id, uint64_t capacity, uint64_t *count);

/* ---- K4b: FlatRTree (join build side) -------------------------------------------------------------------
 * replaces FlatRTree ctor/Push/Sort/Build, src/spatial/operators/spatial_join_physical.cpp:76-220 (Hilbert key
 * in fp32 :173-187, sort, bottom-up pack of node_size). Tie order of equal Hilbert keys = insertion order.
 * The packed tree lives in ONE device buffer (header | boxes | idx) so it can be broadcast to other GPUs. */
SB_API int sb_flat_build(sb_context *ctx, sb_mem mem, const float *box4, uint32_t n, uint32_t node_size,
                         sb_flat **out);
SB_API void sb_flat_destroy(sb_flat *tree);
SB_API uint32_t sb_flat_item_count(const sb_flat *tree);
SB_API uint32_t sb_flat_node_count(const sb_flat *tree);
SB_API uint32_t sb_flat_num_layers(const sb_flat *tree);
/* layer_bounds (host copy, num_layers entries), spatial_join_physical.cpp:86-96 */
SB_API const uint32_t *sb_flat_layer_bounds(const sb_flat *tree);
/* device buffer holding the whole tree + its size, for NCCL broadcast */
SB_API int sb_flat_device_buffer(const sb_flat *tree, void **dev_ptr, uint64_t *bytes);
/* build a tree view over a (received) device buffer produced by sb_flat_device_buffer on another GPU. The
 * buffer stays owned by the caller and must outlive the tree. */
SB_API int sb_flat_attach(sb_context *ctx, void *dev_ptr, uint64_t bytes, sb_flat **out);
/* copy box_array / idx_array out (node_count entries each) */
SB_API int sb_flat_export(sb_context *ctx, const sb_flat *tree, sb_mem mem, float *out_box4, uint32_t *out_idx);

/* ---- K5: batched probe ----------------------------------------------------------------------------------
 * replaces FlatRTree::InitScan/Scan/Lookup (:230-302) driven per probe row by
 * PhysicalSpatialJoin::ExecuteInternal (:748-849). Emits candidate pairs (probe index, build insertion index)
 * for every leaf box intersecting the probe box. */
SB_API int sb_flat_probe(sb_context *ctx, const sb_flat *tree, sb_mem mem, const float *probe_box4, uint64_t m,
                         uint32_t *out_probe, uint32_t *out_build, uint64_t capacity, uint64_t *count);

/* ---- K6: refinement -------------------------------------------------------------------------------------
 * Predicates over candidate pairs; survivors are compacted in place (pairs arrays) and *count updated. */
typedef enum {
	SB_PRED_INTERSECTS = 0, /* ST_Intersects(point, envelope polygon): closed, GEOS semantics geos_module.cpp:1128-1153 */
	SB_PRED_WITHIN = 1      /* ST_Within(point, envelope polygon) / ST_Contains(envelope, point): strict, :386-415,1885-1908 */
} sb_predicate;

SB_API int sb_refine_point_box(sb_context *ctx, sb_mem mem, sb_predicate pred, const double *px, const double *py,
                               const double *bmin_x, const double *bmin_y, const double *bmax_x, const double *bmax_y,
                               uint32_t *pair_probe, uint32_t *pair_build, uint64_t *count);
/* ST_Within(POINT_2D, POLYGON_2D): winding number with the reference's operation order,
 * ST_Contains::Operation spatial_functions_scalar.cpp:1938-2051 (reached via ST_Within :7775-7781).
 * Polygons in DuckDB list layout: polygon p = rings [poly_ring_off[p], poly_ring_off[p+1]),
 * ring r = vertices [ring_vert_off[r], ring_vert_off[r+1]) of vx[], vy[]. */
SB_API int sb_refine_point_polygon(sb_context *ctx, sb_mem mem, const double *px, const double *py, const double *vx,
                                   const double *vy, const uint64_t *ring_vert_off, const uint64_t *poly_ring_off,
                                   uint64_t n_polys, uint64_t n_rings, uint64_t n_verts, uint64_t n_points,
                                   uint32_t *pair_probe, uint32_t *pair_build, uint64_t *count);
/* The GEOS route of the same refinement, for GEOMETRY arguments: replaces GEOSPreparedIntersects_r / GEOSIntersects_r
 * (ST_Intersects, geos_module.cpp:1128-1153), GEOSContains_r (ST_Contains, :386-415) and GEOSWithin_r (ST_Within,
 * :1885-1908) for point x polygon pairs. GEOS is not part of the reference tree; implemented is its point-in-area
 * algorithm (RayCrossingCounter + locatePointInPolygon: shell, then holes) with an exact orientation sign (fp64 filter
 * + expansion arithmetic), i.e. the true OGC answer: a point on the boundary (shell or hole) intersects but is not
 * within (test/sql/geos/predicates.test:15-46, index/rtree_filter_prune.test:14-20). pred = SB_PRED_INTERSECTS or
 * SB_PRED_WITHIN (= ST_Contains(polygon, point)). Same arguments and in-place compaction as sb_refine_point_polygon. */
SB_API int sb_refine_point_polygon_exact(sb_context *ctx, sb_mem mem, sb_predicate pred, const double *px,
                                         const double *py, const double *vx, const double *vy,
                                         const uint64_t *ring_vert_off, const uint64_t *poly_ring_off, uint64_t n_polys,
                                         uint64_t n_rings, uint64_t n_verts, uint64_t n_points, uint32_t *pair_probe,
                                         uint32_t *pair_build, uint64_t *count);
/* ST_Distance(POINT_2D, POINT_2D) (:2189-2214) for each pair, and ST_DWithin filtering (dist <= r). out_dist may
 * be NULL. Survivors compacted in place. */
SB_API int sb_refine_dwithin(sb_context *ctx, sb_mem mem, const double *ax, const double *ay, const double *bx,
                             const double *by, double radius, uint32_t *pair_a, uint32_t *pair_b, double *out_dist,
                             uint64_t *count);
/* elementwise ST_Distance(POINT_2D, POINT_2D) */
SB_API int sb_distance_points(sb_context *ctx, sb_mem mem, const double *ax, const double *ay, const double *bx,
                              const double *by, uint64_t n, double *out);
/* ST_Intersects(BOX_2D, BOX_2D): replaces ST_Intersects::ExecuteBox, spatial_functions_scalar.cpp:5407-5416 (closed
 * intervals; the test of sgl::box_xy::intersects, sgl.hpp:93-95). Each box is four double columns read at index
 * i * stride: stride 1 = DuckDB's struct child vectors, stride 4 = AoS rows such as sb_extent_from_blobs writes (pass
 * box4 + 0..3). valid_a / valid_b (may be NULL): rows with a 0 byte give 0. With the extents and validity bytes of two
 * sb_extent_from_blobs calls this is ST_Intersects_Extent(GEOMETRY, GEOMETRY), :5456-5481. out: one byte per row. */
SB_API int sb_intersects_box2d(sb_context *ctx, sb_mem mem, const double *a_min_x, const double *a_min_y,
                               const double *a_max_x, const double *a_max_y, const double *b_min_x,
                               const double *b_min_y, const double *b_max_x, const double *b_max_y, uint32_t stride,
                               const uint8_t *valid_a, const uint8_t *valid_b, uint64_t n, uint8_t *out);
/* ST_Distance(POINT_2D, LINESTRING_2D): replaces ST_Distance::PointLineStringOperation, spatial_functions_scalar.cpp:
 * 2219-2258 (ClosestPointOnSegment :2158-2176 incl. its 1e-6 ApproxEqualTo vertex test, vertex.hpp:50-52; squared
 * distances, `<` minimum, one sqrt). list_entries = DuckDB list_entry_t rows {u64 offset, u64 length} into the
 * vertex columns vx / vy (n_vertices long). Lines with fewer than two vertices give sqrt(DBL_MAX) (the reference's
 * loop bound underflows for an empty list). Bit-identical doubles. */
SB_API int sb_distance_point_linestring(sb_context *ctx, sb_mem mem, const double *px, const double *py,
                                        const uint64_t *list_entries, uint64_t n, const double *vx, const double *vy,
                                        uint64_t n_vertices, double *out);

/* ---- whole join (the call PhysicalSpatialJoin::Finalize + ExecuteInternal would make) ----------------------
 * Points (probe side, LHS) x boxes (build side, RHS, given as BOX_2D columns of the envelope polygons):
 * bbox -> FlatRTree build -> probe -> exact refinement, pairs out. */
SB_API int sb_join_points_boxes(sb_context *ctx, sb_mem mem, sb_predicate pred, const double *px, const double *py,
                                uint64_t n_points, const double *bmin_x, const double *bmin_y, const double *bmax_x,
                                const double *bmax_y, uint32_t n_boxes, uint32_t *out_probe, uint32_t *out_build,
                                uint64_t capacity, uint64_t *count);
/* The same join driven over SEVERAL devices from one process -- what the single-process DuckDB operator
 * (PhysicalSpatialJoin, spatial_join_physical.cpp:695-945) would call on a multi-GPU host. ctxs[0..n_ctx) are contexts
 * on distinct devices. Host pointers only. Device ctxs[0] builds the FlatRTree; the packed tree and the build side's
 * exact boxes are replicated to the other devices with peer copies (NVLink); every device streams a contiguous slice
 * of the probe column through its own copy-in / probe / copy-out pipeline (one host thread per device, one PCIe link
 * each) and all pair lists land in the caller's ONE pair list (unordered; probe row ids are global). */
SB_API int sb_join_points_boxes_multi(sb_context *const *ctxs, int n_ctx, sb_predicate pred, const double *px,
                                      const double *py, uint64_t n_points, const double *bmin_x, const double *bmin_y,
                                      const double *bmax_x, const double *bmax_y, uint32_t n_boxes, uint32_t *out_probe,
                                      uint32_t *out_build, uint64_t capacity, uint64_t *count);
/* probe an existing tree with points and refine in the same kernel (multi-GPU path: tree was broadcast).
 * point indices emitted = probe_index_base + local index. */
SB_API int sb_flat_probe_points(sb_context *ctx, const sb_flat *tree, sb_mem mem, sb_predicate pred, const double *px,
                                const double *py, uint64_t n_points, uint32_t probe_index_base, const double *bbox2d4,
                                uint32_t *out_probe, uint32_t *out_build, uint64_t capacity, uint64_t *count);
/* ST_Within(point, polygon) / ST_Contains(polygon, point) join: every (point row, polygon row) whose cached float
 * boxes intersect (the FlatRTree candidate rule, spatial_join_physical.cpp:764-849) and for which the winding-number
 * routine ST_Contains::Operation (spatial_functions_scalar.cpp:1938-2051) holds. Polygons in DuckDB list layout (see
 * sb_refine_point_polygon). Large probe sides run polygon-centric over cell-sorted points with a per-polygon edge
 * index in shared memory (pip_join.cu); small ones take FlatRTree probe + sb_refine_point_polygon. Same pair set. */
SB_API int sb_join_points_polygons(sb_context *ctx, sb_mem mem, const double *px, const double *py, uint64_t n_points,
                                   const double *vx, const double *vy, const uint64_t *ring_vert_off,
                                   const uint64_t *poly_ring_off, uint64_t n_polys, uint64_t n_rings, uint64_t n_verts,
                                   uint32_t *out_point, uint32_t *out_poly, uint64_t capacity, uint64_t *count);
/* the same join under the GEOS-route predicates (ST_Intersects / ST_Within / ST_Contains on GEOMETRY arguments, see
 * sb_refine_point_polygon_exact): float-box candidates, then the exact point-in-area location. Large probe sides run
 * polygon-centric with the same shared-memory edge index (pip_join.cu, exact mode). */
SB_API int sb_join_points_polygons_exact(sb_context *ctx, sb_mem mem, sb_predicate pred, const double *px,
                                         const double *py, uint64_t n_points, const double *vx, const double *vy,
                                         const uint64_t *ring_vert_off, const uint64_t *poly_ring_off, uint64_t n_polys,
                                         uint64_t n_rings, uint64_t n_verts, uint32_t *out_point, uint32_t *out_poly,
                                         uint64_t capacity, uint64_t *count);
/* radius join (new capability behind the same operator): all (a, b) with distance <= radius.
 * Builds a FlatRTree over b's points expanded by radius, probes with a, refines with ST_Distance. */
SB_API int sb_join_dwithin(sb_context *ctx, sb_mem mem, const double *ax, const double *ay, uint64_t na,
                           const double *bx, const double *by, uint32_t nb, double radius, uint32_t *out_a,
                           uint32_t *out_b, double *out_dist, uint64_t capacity, uint64_t *count);

/* ---- LEFT / RIGHT / FULL join bookkeeping ----------------------------------------------------------------------
 * replaces the left_outer_marker[] stores (spatial_join_physical.cpp:867-876) and the matched-flag store into the
 * build tuple (:878-891) for a whole pair list: probe_matched[pair_probe[i] - probe_index_base] = 1 and
 * build_matched[pair_build[i]] = 1. Flags are only ever set, never cleared, so the build-side array accumulates over
 * probe batches exactly like the tuple's match column. Either flag array may be NULL. */
SB_API int sb_join_mark_matched(sb_context *ctx, sb_mem mem, const uint32_t *pair_probe, const uint32_t *pair_build,
                                uint64_t count, uint32_t probe_index_base, uint8_t *probe_matched, uint64_t n_probe,
                                uint8_t *build_matched, uint64_t n_build);
/* rows whose flag is 0, ascending, as row_base + position: the EMIT_LHS selection of unmatched probe rows
 * (:917-938, emitted NULL-padded on the build side) and the RIGHT / FULL source scan of unmatched build rows
 * (PhysicalSpatialJoin::GetData :1022-1068, NULL-padded on the probe side). SB_ERR_CAPACITY when *count > capacity. */
SB_API int sb_join_unmatched_rows(sb_context *ctx, sb_mem mem, const uint8_t *matched, uint64_t n, uint32_t row_base,
                                  uint32_t *out_rows, uint64_t capacity, uint64_t *count);
/* Order-independent checksum of a pair list: *out_count = pairs whose probe id lies in [probe_lo, probe_hi),
 * *out_hash = wrapping 64-bit sum over those pairs of mix(probe, build), mix(p, b): k = p * 2654435761 + b * 2246822519;
 * k ^= k >> 29; k *= 0xBF58476D1CE4E5B9; k ^= k >> 32. Compares a join result of any size with the pair set of the
 * reference operator (spatial_join_physical.cpp:781-912 emits the same (probe row, build row) pairs) without
 * sorting or moving it; harness utility, no reference counterpart. */
SB_API int sb_pairs_checksum(sb_context *ctx, sb_mem mem, const uint32_t *pair_probe, const uint32_t *pair_build,
                             uint64_t count, uint32_t probe_lo, uint32_t probe_hi, uint64_t *out_count,
                             uint64_t *out_hash);
/* sb_pairs_checksum for callers that may not allocate (CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc). No counterpart in the reference. It takes no context and
 * calls no cudaMalloc / cudaFree / pool routine: its scratch (one partial per 4096-pair tile) is a static 256 KiB
 * __device__ array every GPU that loads the library holds until exit. DEVICE pointers only: pair_probe / pair_build
 * [count], out_count_hash[2] = {count, hash} as sb_pairs_checksum defines them, written on `cuda_stream` (a
 * cudaStream_t, NULL = default stream) of `device` (< 0 = the current device); the call does not synchronise.
 * count > SB_NOALLOC_MAX_PAIRS fails with SB_ERR_ARG. All calls on a device share the scratch: calls that overlap on
 * the GPU (two streams, two threads) race; calls on one stream are safe. The first call per device looks the scratch
 * address up (cudaGetSymbolAddress) -- make it outside a capture. */
#define SB_NOALLOC_MAX_PAIRS 67108864ull /* 2^26 pairs */
SB_API int sb_pairs_checksum_noalloc(int device, void *cuda_stream, const uint32_t *pair_probe,
                                     const uint32_t *pair_build, uint64_t count, uint32_t probe_lo, uint32_t probe_hi,
                                     uint64_t *out_count_hash);

/* ---- probe-side operator glue: PhysicalSpatialJoin's life cycle for a STREAMING probe side ------------------
 * DuckDB hands the operator one <= 2048-row probe chunk per ExecuteInternal call (spatial_join_physical.cpp:695-945);
 * a GPU probe wants batches of >= 10^5 rows. The op buffers pushed chunks in pinned memory (CachingOperatorState-style)
 * and probes them as one device batch. Host pointers throughout (DuckDB vectors). Probe row ids count pushed rows from
 * 0 across batches; build row ids index the build columns (the row_array order of :117-129).
 *   create : Sink / Finalize :431-601 -- BOX_2D build keys -> BOX_2DF -> FlatRTree(32), kept in HBM
 *   push   : append a chunk (SB_ERR_CAPACITY when it does not fit the batch: flush first)
 *   flush  : probe + exact predicate of the pending rows -> INNER pairs; LEFT / FULL also return the pending rows
 *            without a match, ascending (EMIT_LHS :917-938: the caller emits them NULL-padded). SB_ERR_CAPACITY leaves
 *            the batch pending with *n_pairs = the capacity to come back with.
 *   finish : GetData :1022-1068 -- RIGHT / FULL: build rows no probe row matched, ascending. */
typedef struct sb_join_op sb_join_op;
typedef enum { SB_JOIN_INNER = 0, SB_JOIN_LEFT = 1, SB_JOIN_RIGHT = 2, SB_JOIN_FULL = 3 } sb_join_type;
SB_API int sb_join_op_create(sb_context *ctx, sb_predicate pred, sb_join_type join_type, const double *bmin_x,
                             const double *bmin_y, const double *bmax_x, const double *bmax_y, uint32_t n_build,
                             uint64_t batch_rows, sb_join_op **out);
SB_API int sb_join_op_push(sb_join_op *op, const double *px, const double *py, uint64_t n_rows, uint64_t *pending);
SB_API int sb_join_op_flush(sb_join_op *op, uint32_t *out_probe, uint32_t *out_build, uint64_t pair_capacity,
                            uint64_t *n_pairs, uint32_t *out_lhs_unmatched, uint64_t lhs_capacity, uint64_t *n_lhs);
SB_API int sb_join_op_finish(sb_join_op *op, uint32_t *out_rhs_unmatched, uint64_t capacity, uint64_t *n_rhs);
SB_API uint64_t sb_join_op_rows_pushed(const sb_join_op *op);
SB_API void sb_join_op_destroy(sb_join_op *op);

/* ---- harness utility: ST_GeneratePoints on the device ---------------------------------------------------------
 * restates ST_GeneratePoints (spatial_functions_table.cpp:82-99) on DuckDB's pcg32 RandomEngine with skip-ahead:
 * produces points [skip, skip+n) of the sequential stream of `seed`, bit-identical to the CPU generator. */
SB_API int sb_generate_points(sb_context *ctx, sb_mem mem, uint64_t seed, uint64_t skip, uint64_t n,
                              const double *bounds4, double *x, double *y);

#ifdef __cplusplus
}
#endif
#endif
