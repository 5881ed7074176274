// ref_duckdb.cpp -- TEST INFRASTRUCTURE ONLY (builds oracle/_ref/libref_duckdb.so).
//
// The reference's own DuckDB-linked code for the hot path, run as the reference runs it: an in-memory DuckDB
// (core built from /root/reference/duckdb by oracle/build_duckdb_ref.sh) with the UNMODIFIED reference translation
// units src/sgl/sgl.cpp and src/spatial/{geometry,util,index/rtree,operators,modules/main}/*.cpp + spatial_types.cpp
// registered through the reference's own Register* functions (spatial_extension.cpp:24-50 minus the GDAL / PROJ /
// OSM / shapefile / GEOS modules, whose third-party sources are not in the tree). Nothing is copied.
//
// It gives the tests and bench.py, through SQL:
//   * CREATE INDEX ... USING RTREE (PhysicalCreateRTreeIndex / BuildRTreeBottomUp) and rtree_index_dump (the structure
//     oracle of SURVEY 8c), index scans, persistence (FixedSizeAllocator reload) and insert / delete on a built index;
//   * ST_Within(POINT_2D, POLYGON_2D) (the native winding routine), ST_Hilbert, ST_Extent, ST_Distance, casts;
//   * the real SPATIAL_JOIN operator. GEOS being absent, the GEOMETRY x GEOMETRY predicates the join optimizer looks
//     for (ST_Intersects / ST_Within / ST_Contains) are registered here as STAND-INS for point x polygon arguments that
//     return the exact OGC answer (oracle.c orc_point_polygon_locate: GEOS' ray-crossing algorithm with an exact
//     orientation sign). Everything around the predicate -- optimizer rule, sink, FlatRTree build, probe, gather,
//     emit -- is the reference's.
#include "duckdb.hpp"
#include "duckdb/main/appender.hpp"
#include "duckdb/main/extension_util.hpp"
#include "duckdb/common/types/data_chunk.hpp"
#include "duckdb/function/scalar_function.hpp"

#include "spatial/index/rtree/rtree_module.hpp"
#include "spatial/modules/main/spatial_functions.hpp"
#include "spatial/operators/spatial_join_optimizer.hpp"
#include "spatial/operators/spatial_operator_extension.hpp"
#include "spatial/spatial_types.hpp"

#include <cstring>
#include <string>
#include <vector>

extern "C" int orc_point_polygon_locate(double x, double y, const double *vx, const double *vy, const uint64_t *ring_off,
                                        uint64_t r0, uint64_t r1);

namespace {

using namespace duckdb;

thread_local std::string g_err;

// ---- stand-in predicates (point x polygon blobs) --------------------------------------------------------------
struct BlobView {
	uint8_t type = 255;
	uint32_t count = 0;
	const uint8_t *payload = nullptr;
};
static BlobView ViewBlob(const string_t &s) {
	BlobView v;
	const auto *p = const_data_ptr_cast(s.GetData());
	const idx_t len = s.GetSize();
	if (len < 16) {
		throw InvalidInputException("stand-in predicate: blob too short");
	}
	v.type = p[0];
	const uint8_t flags = p[1];
	if (flags & 3) {
		throw NotImplementedException("stand-in predicate: Z/M geometries");
	}
	idx_t pos = 8 + ((flags & 4) ? 16 : 0);
	uint32_t t;
	memcpy(&t, p + pos, 4);
	memcpy(&v.count, p + pos + 4, 4);
	v.payload = p + pos + 8;
	(void)t;
	return v;
}
// 0 exterior, 1 boundary, 2 interior
static int LocatePointInPolygon(const BlobView &pt, const BlobView &poly, std::vector<double> &vx, std::vector<double> &vy,
                                std::vector<uint64_t> &ro) {
	if (pt.count == 0 || poly.count == 0) {
		return 0;
	}
	double x, y;
	memcpy(&x, pt.payload, 8);
	memcpy(&y, pt.payload + 8, 8);
	const uint32_t nr = poly.count;
	const uint8_t *verts = poly.payload + 4 * (idx_t)(nr + (nr & 1));
	ro.resize(nr + 1);
	uint64_t tot = 0;
	for (uint32_t r = 0; r < nr; r++) {
		uint32_t c;
		memcpy(&c, poly.payload + 4 * (idx_t)r, 4);
		ro[r] = tot;
		tot += c;
	}
	ro[nr] = tot;
	vx.resize(tot);
	vy.resize(tot);
	for (uint64_t i = 0; i < tot; i++) {
		memcpy(&vx[i], verts + 16 * i, 8);
		memcpy(&vy[i], verts + 16 * i + 8, 8);
	}
	return orc_point_polygon_locate(x, y, vx.data(), vy.data(), ro.data(), 0, nr);
}

// MODE 0: intersects (symmetric), 1: within(a, b) = a point inside polygon b, 2: contains(a, b) = polygon a contains point b
template <int MODE>
static void StandInPredicate(DataChunk &args, ExpressionState &, Vector &result) {
	std::vector<double> vx, vy;
	std::vector<uint64_t> ro;
	BinaryExecutor::Execute<string_t, string_t, bool>(args.data[0], args.data[1], result, args.size(),
	                                                  [&](const string_t &a, const string_t &b) {
		                                                  BlobView va = ViewBlob(a), vb = ViewBlob(b);
		                                                  const bool a_pt = va.type == 0, b_pt = vb.type == 0;
		                                                  const bool a_pg = va.type == 2, b_pg = vb.type == 2;
		                                                  int loc;
		                                                  if (a_pt && b_pg && MODE != 2) {
			                                                  loc = LocatePointInPolygon(va, vb, vx, vy, ro);
		                                                  } else if (a_pg && b_pt && MODE != 1) {
			                                                  loc = LocatePointInPolygon(vb, va, vx, vy, ro);
		                                                  } else {
			                                                  throw NotImplementedException(
			                                                      "stand-in predicate: only point x polygon arguments");
		                                                  }
		                                                  return MODE == 0 ? loc != 0 : loc == 2;
	                                                  });
}

static void LoadReducedSpatial(DatabaseInstance &db) {
	GeoTypes::Register(db);
	RegisterSpatialCastFunctions(db);
	RegisterSpatialScalarFunctions(db);
	RegisterSpatialAggregateFunctions(db);
	RegisterSpatialTableFunctions(db);
	SpatialJoinOptimizer::Register(db);
	RTreeModule::RegisterIndex(db);
	RTreeModule::RegisterIndexPragmas(db);
	RTreeModule::RegisterIndexScan(db);
	RTreeModule::RegisterIndexPlanScan(db);
	RegisterSpatialOperatorExtension(db);
	const auto G = GeoTypes::GEOMETRY();
	ExtensionUtil::AddFunctionOverload(db, ScalarFunction("ST_Intersects", {G, G}, LogicalType::BOOLEAN, StandInPredicate<0>));
	ExtensionUtil::AddFunctionOverload(db, ScalarFunction("ST_Within", {G, G}, LogicalType::BOOLEAN, StandInPredicate<1>));
	ExtensionUtil::AddFunctionOverload(db, ScalarFunction("ST_Contains", {G, G}, LogicalType::BOOLEAN, StandInPredicate<2>));
}

struct RdDb {
	unique_ptr<DuckDB> db;
	unique_ptr<Connection> con;
};
struct RdResult {
	unique_ptr<MaterializedQueryResult> res;
};

} // namespace

#define RD_API extern "C" __attribute__((visibility("default")))

RD_API const char *rd_last_error() {
	return g_err.c_str();
}

// path: NULL / "" = in-memory; threads <= 0 = DuckDB default (all cores)
RD_API void *rd_open(const char *path, int threads) {
	try {
		DBConfig config;
		if (threads > 0) {
			config.options.maximum_threads = (idx_t)threads;
		}
		auto h = new RdDb();
		h->db = make_uniq<DuckDB>((path && *path) ? path : nullptr, &config);
		LoadReducedSpatial(*h->db->instance);
		h->con = make_uniq<Connection>(*h->db);
		return h;
	} catch (std::exception &e) {
		g_err = e.what();
		return nullptr;
	}
}

RD_API void rd_close(void *hp) {
	auto h = static_cast<RdDb *>(hp);
	if (h) {
		h->con.reset();
		h->db.reset();
		delete h;
	}
}

RD_API int rd_exec(void *hp, const char *sql) {
	auto h = static_cast<RdDb *>(hp);
	try {
		auto r = h->con->Query(sql);
		if (r->HasError()) {
			g_err = r->GetError();
			return 1;
		}
		return 0;
	} catch (std::exception &e) {
		g_err = e.what();
		return 1;
	}
}

RD_API void *rd_query(void *hp, const char *sql) {
	auto h = static_cast<RdDb *>(hp);
	try {
		auto r = h->con->Query(sql);
		if (r->HasError()) {
			g_err = r->GetError();
			return nullptr;
		}
		auto out = new RdResult();
		out->res = std::move(r);
		return out;
	} catch (std::exception &e) {
		g_err = e.what();
		return nullptr;
	}
}

RD_API int64_t rd_rows(void *rp) {
	return (int64_t) static_cast<RdResult *>(rp)->res->RowCount();
}
RD_API int rd_cols(void *rp) {
	return (int)static_cast<RdResult *>(rp)->res->ColumnCount();
}
RD_API void rd_free_result(void *rp) {
	delete static_cast<RdResult *>(rp);
}

template <class OUT>
static int CopyColumn(RdResult *r, int col, OUT *out, uint8_t *valid) {
	try {
		auto &coll = r->res->Collection();
		idx_t row = 0;
		const LogicalType want = std::is_same<OUT, double>::value ? LogicalType::DOUBLE : LogicalType::BIGINT;
		for (auto &chunk : coll.Chunks()) {
			Vector casted(want);
			VectorOperations::Cast(*r->res->client_properties.client_context.get(), chunk.data[col], casted, chunk.size());
			UnifiedVectorFormat f;
			casted.ToUnifiedFormat(chunk.size(), f);
			auto data = UnifiedVectorFormat::GetData<OUT>(f);
			for (idx_t i = 0; i < chunk.size(); i++) {
				const idx_t k = f.sel->get_index(i);
				const bool ok = f.validity.RowIsValid(k);
				out[row] = ok ? data[k] : OUT(0);
				if (valid) {
					valid[row] = ok;
				}
				row++;
			}
		}
		return 0;
	} catch (std::exception &e) {
		g_err = e.what();
		return 1;
	}
}

// whole column `col` cast to BIGINT / DOUBLE; valid (optional) receives 0 for NULL
RD_API int rd_col_i64(void *rp, int col, int64_t *out, uint8_t *valid) {
	return CopyColumn<int64_t>(static_cast<RdResult *>(rp), col, out, valid);
}
RD_API int rd_col_f64(void *rp, int col, double *out, uint8_t *valid) {
	return CopyColumn<double>(static_cast<RdResult *>(rp), col, out, valid);
}

// CREATE TABLE name (col DOUBLE|BIGINT, ...) filled from host columns; types[c]: 0 = DOUBLE, 1 = BIGINT
RD_API int rd_create_table(void *hp, const char *name, int ncols, const char **colnames, const int *types,
                           const void **cols, int64_t n) {
	auto h = static_cast<RdDb *>(hp);
	try {
		std::string ddl = std::string("CREATE OR REPLACE TABLE ") + name + " (";
		vector<LogicalType> lt;
		for (int c = 0; c < ncols; c++) {
			ddl += std::string(c ? ", " : "") + colnames[c] + (types[c] ? " BIGINT" : " DOUBLE");
			lt.push_back(types[c] ? LogicalType::BIGINT : LogicalType::DOUBLE);
		}
		ddl += ")";
		auto r = h->con->Query(ddl);
		if (r->HasError()) {
			g_err = r->GetError();
			return 1;
		}
		Appender app(*h->con, name);
		DataChunk chunk;
		chunk.Initialize(Allocator::DefaultAllocator(), lt);
		for (int64_t base = 0; base < n; base += STANDARD_VECTOR_SIZE) {
			const idx_t m = (idx_t)std::min<int64_t>(STANDARD_VECTOR_SIZE, n - base);
			chunk.Reset();
			for (int c = 0; c < ncols; c++) {
				memcpy(chunk.data[c].GetData(), static_cast<const char *>(cols[c]) + 8 * base, 8 * m);
			}
			chunk.SetCardinality(m);
			app.AppendDataChunk(chunk);
		}
		app.Close();
		return 0;
	} catch (std::exception &e) {
		g_err = e.what();
		return 1;
	}
}

// ---------------------------------------------------------------------------------------------------------------
// N2: a bulk-loaded tree handed over as FixedSizeAllocator block images (sb_rtree_export_pages / the oracle's
// orc_rtree_export_pages) is loaded into the REFERENCE's RTree through the reference's own allocators, then the
// reference's own scan / insert / delete code (rtree.cpp:73-592, rtree_scanner.hpp:38-83) runs on it.
//   rd_rtree_load : allocates leaf_pages / branch_pages segments with FixedSizeAllocator::New() (the MakePage order of
//   a fresh index), checks that every returned IndexPointer is the one the image assigned to that page, copies the page
//   bodies in, compares the allocator's whole buffers (validity bitmask + segments) with the images byte for byte,
//   and sets the root entry. *mismatch: 0 = identical, 1 = pointer sequence differs, 2 = buffer bytes differ.
// ---------------------------------------------------------------------------------------------------------------
#include "duckdb/storage/in_memory_block_manager.hpp"
#include "duckdb/storage/buffer_manager.hpp"
#include "spatial/index/rtree/rtree.hpp"
#include "spatial/index/rtree/rtree_scanner.hpp"

namespace {
struct RdTree {
	unique_ptr<InMemoryBlockManager> bm;
	unique_ptr<RTree> tree;
};
} // namespace

RD_API void *rd_rtree_load(void *hp, uint32_t cap, uint32_t min_cap, const uint8_t *leaf_img, uint64_t leaf_pages,
                           const uint8_t *branch_img, uint64_t branch_pages, const void *root24, uint64_t block_size,
                           int *mismatch) {
	auto h = static_cast<RdDb *>(hp);
	try {
		*mismatch = 0;
		auto t = new RdTree();
		auto &buffer_manager = BufferManager::GetBufferManager(*h->db->instance);
		t->bm = make_uniq<InMemoryBlockManager>(buffer_manager, optional_idx(block_size + 8), optional_idx(8));
		RTreeConfig cfg;
		cfg.max_node_capacity = cap;
		cfg.min_node_capacity = min_cap;
		t->tree = make_uniq<RTree>(*t->bm, cfg);
		if (t->bm->GetBlockSize() != block_size) {
			g_err = "block size " + std::to_string(t->bm->GetBlockSize()) + " != " + std::to_string(block_size);
			delete t;
			return nullptr;
		}
		const idx_t seg = cfg.GetNodeByteSize();
		struct Side {
			FixedSizeAllocator *alloc;
			const uint8_t *img;
			uint64_t pages;
		} sides[2] = {{&t->tree->GetLeafAllocator(), leaf_img, leaf_pages}, {&t->tree->GetNodeAllocator(), branch_img, branch_pages}};
		for (auto &s : sides) {
			if (s.pages == 0) {
				continue;
			}
			// the allocator's geometry (private fields; this TU is compiled with -fno-access-control)
			const idx_t per_buffer = s.alloc->available_segments_per_buffer, bitmask_offset = s.alloc->bitmask_offset;
			std::vector<IndexPointer> ptrs;
			for (uint64_t k = 0; k < s.pages; k++) {
				ptrs.push_back(s.alloc->New());
			}
			for (uint64_t k = 0; k < s.pages; k++) {
				const idx_t want_buf = k / per_buffer, want_off = k % per_buffer;
				if (ptrs[k].GetBufferId() != want_buf || ptrs[k].GetOffset() != want_off) {
					*mismatch = 1; // the image numbered its pages differently from FixedSizeAllocator::New()
				}
				memcpy(s.alloc->Get(ptrs[k], true), s.img + want_buf * block_size + bitmask_offset + want_off * seg, seg);
			}
			// whole buffers: validity bitmask + segments, byte for byte
			const uint64_t nbuf = (s.pages + per_buffer - 1) / per_buffer;
			for (uint64_t b = 0; b < nbuf && *mismatch == 0; b++) {
				const data_ptr_t base = s.alloc->Get(ptrs[b * per_buffer], false) - bitmask_offset;
				const uint64_t in_buf = std::min<uint64_t>(per_buffer, s.pages - b * per_buffer);
				if (memcmp(base, s.img + b * block_size, bitmask_offset + in_buf * seg) != 0) {
					*mismatch = 2;
				}
			}
		}
		RTreeEntry root;
		memcpy(&root, root24, sizeof(RTreeEntry));
		t->tree->SetRoot(root);
		return t;
	} catch (std::exception &e) {
		g_err = e.what();
		return nullptr;
	}
}

RD_API void rd_rtree_free(void *tp) {
	delete static_cast<RdTree *>(tp);
}

// DFS dump through the reference's RTreeScanner (what rtree_index_dump does); returns rows (arrays may be NULL)
RD_API uint64_t rd_rtree_dump(void *tp, int32_t *lvl, float *box4, int64_t *rid, uint64_t cap) {
	auto t = static_cast<RdTree *>(tp);
	uint64_t n = 0;
	if (!t->tree->GetRoot().pointer.IsSet()) {
		return 0;
	}
	RTreeScanner sc;
	sc.Init(t->tree->GetRoot());
	sc.Scan(*t->tree, [&](const RTreeEntry &e, const idx_t &level) {
		if (lvl && n < cap) {
			lvl[n] = (int32_t)level;
			memcpy(box4 + 4 * n, &e.bounds, 16);
			rid[n] = e.pointer.IsRowId() ? e.pointer.GetRowId() : -1;
		}
		n++;
		return RTreeScanResult::CONTINUE;
	});
	return n;
}

// box scan like RTreeIndex::Scan (rtree_index.cpp:106-139): row ids of leaf entries intersecting q
RD_API uint64_t rd_rtree_scan(void *tp, const float *q, int64_t *out, uint64_t cap) {
	auto t = static_cast<RdTree *>(tp);
	RTreeBounds qb;
	memcpy(&qb, q, 16);
	uint64_t n = 0;
	auto &root = t->tree->GetRoot();
	if (!root.pointer.IsSet() || !qb.Intersects(root.bounds)) {
		return 0;
	}
	RTreeScanner sc;
	sc.Init(root);
	sc.Scan(*t->tree, [&](const RTreeEntry &e, const idx_t &) {
		if (!e.bounds.Intersects(qb)) {
			return RTreeScanResult::SKIP;
		}
		if (e.pointer.IsRowId()) {
			if (n < cap) {
				out[n] = e.pointer.GetRowId();
			}
			n++;
		}
		return RTreeScanResult::CONTINUE;
	});
	return n;
}

RD_API int rd_rtree_insert(void *tp, const float *box4, const int64_t *rid, uint64_t n) {
	auto t = static_cast<RdTree *>(tp);
	try {
		for (uint64_t i = 0; i < n; i++) {
			RTreeEntry e;
			e.pointer = RTree::MakeRowId(rid[i]);
			memcpy(&e.bounds, box4 + 4 * i, 16);
			t->tree->Insert(e);
		}
		return 0;
	} catch (std::exception &ex) {
		g_err = ex.what();
		return 1;
	}
}

RD_API int rd_rtree_delete(void *tp, const float *box4, const int64_t *rid, uint64_t n) {
	auto t = static_cast<RdTree *>(tp);
	try {
		for (uint64_t i = 0; i < n; i++) {
			RTreeEntry e;
			e.pointer = RTree::MakeRowId(rid[i]);
			memcpy(&e.bounds, box4 + 4 * i, 16);
			t->tree->Delete(e);
		}
		return 0;
	} catch (std::exception &ex) {
		g_err = ex.what();
		return 1;
	}
}
