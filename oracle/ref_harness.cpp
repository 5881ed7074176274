// ref_harness.cpp -- TEST INFRASTRUCTURE ONLY (builds oracle/_ref/libref.so).
//
// Thin extern "C" wrappers around the REFERENCE'S OWN code, compiled from where it lies under
// /root/reference (nothing is copied): sgl.hpp (Hilbert), util/math.hpp (outward rounding),
// geometry/bbox.hpp (Box2D::Intersects), geometry_type.hpp (TryGetCachedBounds), pcg_random.hpp
// (DuckDB's RandomEngine generator) and the FlatRTree class of operators/spatial_join_physical.cpp,
// reached by #including that translation unit.  Used to pin oracle.c (tests/test_oracle_vs_ref.py)
// and as the "reference" CPU baseline for the join probe.
//
// The persistent STR packer (BuildRTreeBottomUp) and the POLYGON_2D winding routine need a live DuckDB
// (BufferManager / Vector) and cannot be built without running DuckDB's cmake build: they are pinned by
// golden vectors instead ("oracle").
//
// The handful of DuckDB symbols the FlatRTree path needs (Allocator / AllocatedData / exception ctor)
// are defined below as malloc-backed shims; every other undefined DuckDB symbol belongs to operator
// code that is never reached from these entry points.

#include "/root/reference/src/spatial/operators/spatial_join_physical.cpp"

#include "spatial/geometry/geometry_type.hpp"
#include "sgl/sgl.hpp"
#include "pcg_random.hpp"
// the reference's WKB reader (wkb_reader_try_parse_stats, the body of ST_Extent(WKB_BLOB)), compiled where it lies
#include "/root/reference/src/sgl/sgl.cpp"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <new>

// ---- shims -----------------------------------------------------------------------------------------
namespace duckdb {

AllocatedData::AllocatedData() : allocator(nullptr), pointer(nullptr), allocated_size(0) {
}
AllocatedData::AllocatedData(Allocator &allocator_p, data_ptr_t pointer_p, idx_t size_p)
    : allocator(&allocator_p), pointer(pointer_p), allocated_size(size_p) {
}
AllocatedData::~AllocatedData() {
	std::free(pointer);
	pointer = nullptr;
}
AllocatedData &AllocatedData::operator=(AllocatedData &&other) noexcept {
	std::swap(allocator, other.allocator);
	std::swap(pointer, other.pointer);
	std::swap(allocated_size, other.allocated_size);
	return *this;
}
data_ptr_t Allocator::AllocateData(idx_t size) {
	return static_cast<data_ptr_t>(std::malloc(size ? size : 1));
}
static char g_alloc_storage[sizeof(Allocator)];
Allocator &Allocator::DefaultAllocator() {
	// never dereferenced beyond AllocateData above (which ignores `this`)
	return *reinterpret_cast<Allocator *>(g_alloc_storage);
}

struct RefFlat {
	FlatRTree *tree;
};

} // namespace duckdb

using namespace duckdb;

extern "C" {

// ---- scalar helpers -------------------------------------------------------------------------------
__attribute__((visibility("default"))) float ref_double_to_float_down(double d) {
	return MathUtil::DoubleToFloatDown(d);
}
__attribute__((visibility("default"))) float ref_double_to_float_up(double d) {
	return MathUtil::DoubleToFloatUp(d);
}
__attribute__((visibility("default"))) uint32_t ref_hilbert_encode(uint32_t n, uint32_t x, uint32_t y) {
	return sgl::util::hilbert_encode(n, x, y);
}
__attribute__((visibility("default"))) uint32_t ref_hilbert_f32_to_u32(float f) {
	return sgl::util::hilbert_f32_to_u32(f);
}
// what `static_cast<uint32_t>(fp)` compiles to with the reference's flags (x86-64, no -march)
__attribute__((visibility("default"))) uint32_t ref_cast_f32_u32(float v) {
	volatile float t = v;
	return static_cast<uint32_t>(t);
}
__attribute__((visibility("default"))) uint32_t ref_cast_f64_u32(double v) {
	volatile double t = v;
	return static_cast<uint32_t>(t);
}
__attribute__((visibility("default"))) int ref_box_intersects(const float *a, const float *b) {
	Box2D<float> x, y;
	x.min.x = a[0], x.min.y = a[1], x.max.x = a[2], x.max.y = a[3];
	y.min.x = b[0], y.min.y = b[1], y.max.x = b[2], y.max.y = b[3];
	return x.Intersects(y);
}

// RandomEngine(seed).NextRandom32(min,max) pairs, as ST_GeneratePoints draws them
// (spatial_functions_table.cpp:82-99, random_engine.cpp:50-63) on the reference's pcg32.
__attribute__((visibility("default"))) void ref_generate_points(uint64_t seed, uint64_t n, double min_x, double min_y,
                                                                 double max_x, double max_y, double *x, double *y) {
	pcg32 rng;
	rng.seed(seed);
	for (uint64_t i = 0; i < n; i++) {
		x[i] = min_x + std::ldexp(rng(), -32) * (max_x - min_x);
		y[i] = min_y + std::ldexp(rng(), -32) * (max_y - min_y);
	}
}

// geometry_t::TryGetCachedBounds on a non-inlined blob (len > 12). 1 / 0, -1 if it threw.
__attribute__((visibility("default"))) int ref_cached_bounds(const uint8_t *blob, uint32_t len, float *box4) {
	try {
		string_t s(reinterpret_cast<const char *>(blob), len);
		geometry_t g(s);
		Box2D<float> b;
		if (!g.TryGetCachedBounds(b)) {
			return 0;
		}
		box4[0] = b.min.x, box4[1] = b.min.y, box4[2] = b.max.x, box4[3] = b.max.y;
		return 1;
	} catch (...) {
		return -1;
	}
}

// ---- FlatRTree --------------------------------------------------------------------------------------
__attribute__((visibility("default"))) void *ref_flat_build(const float *box4, uint32_t n, uint32_t node_size) {
	auto t = new FlatRTree(Allocator::DefaultAllocator(), n, node_size);
	for (uint32_t i = 0; i < n; i++) {
		Box2D<float> b;
		b.min.x = box4[4 * (size_t)i + 0], b.min.y = box4[4 * (size_t)i + 1];
		b.max.x = box4[4 * (size_t)i + 2], b.max.y = box4[4 * (size_t)i + 3];
		// row "pointer" = insertion index + 1 (non-null)
		t->Push(b, reinterpret_cast<data_ptr_t>(static_cast<uintptr_t>(i) + 1));
	}
	if (n > 0) {
		t->Build();
	}
	return t;
}
__attribute__((visibility("default"))) void ref_flat_free(void *h) {
	delete static_cast<FlatRTree *>(h);
}
__attribute__((visibility("default"))) uint32_t ref_flat_nodes(void *h) {
	return (uint32_t) static_cast<FlatRTree *>(h)->box_array.size();
}
__attribute__((visibility("default"))) uint32_t ref_flat_nlayers(void *h) {
	return (uint32_t) static_cast<FlatRTree *>(h)->layer_bounds.size();
}
__attribute__((visibility("default"))) void ref_flat_export(void *h, float *box4, uint32_t *idx, uint32_t *layer_bounds) {
	auto t = static_cast<FlatRTree *>(h);
	const size_t nodes = t->box_array.size();
	for (size_t i = 0; i < nodes; i++) {
		const auto &b = t->box_array[i];
		box4[4 * i + 0] = b.min.x, box4[4 * i + 1] = b.min.y, box4[4 * i + 2] = b.max.x, box4[4 * i + 3] = b.max.y;
		idx[i] = t->idx_array[i];
	}
	for (size_t i = 0; i < t->layer_bounds.size(); i++) {
		layer_bounds[i] = t->layer_bounds[i];
	}
}

// FlatRTree::InitScan + Lookup on a scan state whose DuckDB `Vector matches` member is never constructed
// (Lookup does not touch it). Emits insertion indices. Returns the match count (writes at most cap).
struct RawScan {
	alignas(FlatRTreeScanState) unsigned char mem[sizeof(FlatRTreeScanState)];
};

static uint64_t ref_lookup_one(const FlatRTree *t, FlatRTreeScanState &st, const float *q, uint32_t *out, uint64_t cap) {
	if (t->Count() == 0) {
		return 0;
	}
	Box2D<float> b;
	b.min.x = q[0], b.min.y = q[1], b.max.x = q[2], b.max.y = q[3];
	t->InitScan(st, b);
	uint64_t cnt = 0;
	t->Lookup(st, [&](const data_ptr_t &row) {
		const auto v = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(row) - 1);
		if (out && cnt < cap) {
			out[cnt] = v;
		}
		cnt++;
		return false;
	});
	return cnt;
}

__attribute__((visibility("default"))) uint64_t ref_flat_lookup(void *h, const float *qbox, uint32_t *out, uint64_t cap) {
	RawScan raw;
	auto &st = *reinterpret_cast<FlatRTreeScanState *>(raw.mem);
	new (&st.search_queue) std::queue<size_t>();
	const auto r = ref_lookup_one(static_cast<FlatRTree *>(h), st, qbox, out, cap);
	st.search_queue.~queue();
	return r;
}

// Whole probe side of the join on the reference FlatRTree, OpenMP over 2048-row chunks (the reference probes
// chunk-parallel, spatial_join_physical.hpp:39-41), with the exact stand-in predicate for a POINT against an
// ST_MakeEnvelope polygon (GEOS is an un-vendored dependency): pred 0 = intersects (closed), 1 = within (strict).
// Counts only (the timed CPU baseline); returns refined pairs, *cand = float-box candidates.
// ref_flat_join_hash additionally returns the order-independent checksum of the pair set that the product's
// sb_pairs_checksum computes (wrapping sum of mix(probe_base + i, build)); see include/spatial_b200.h.
static inline uint64_t pair_mix(uint32_t p, uint32_t b) {
	uint64_t k = (uint64_t)p * 2654435761ull + (uint64_t)b * 2246822519ull;
	k ^= k >> 29;
	k *= 0xBF58476D1CE4E5B9ull;
	k ^= k >> 32;
	return k;
}
static uint64_t flat_join_impl(void *h, const double *px, const double *py, uint64_t np, const double *bminx,
                               const double *bminy, const double *bmaxx, const double *bmaxy, int pred, int threads,
                               uint64_t *cand_out, uint32_t probe_base, uint64_t *hash_out) {
	auto t = static_cast<FlatRTree *>(h);
	uint64_t total = 0, cand = 0, hsum = 0;
	const uint64_t chunk = STANDARD_VECTOR_SIZE;
	const int64_t nchunks = (int64_t)((np + chunk - 1) / chunk);
#pragma omp parallel num_threads(threads) reduction(+ : total, cand, hsum)
	{
		RawScan raw;
		auto &st = *reinterpret_cast<FlatRTreeScanState *>(raw.mem);
		new (&st.search_queue) std::queue<size_t>();
		std::vector<uint32_t> hits(t->Count() ? t->Count() : 1);
#pragma omp for schedule(dynamic, 16)
		for (int64_t c = 0; c < nchunks; c++) {
			const uint64_t e = std::min<uint64_t>((c + 1) * chunk, np);
			for (uint64_t i = c * chunk; i < e; i++) {
				const float q[4] = {MathUtil::DoubleToFloatDown(px[i]), MathUtil::DoubleToFloatDown(py[i]),
				                    MathUtil::DoubleToFloatUp(px[i]), MathUtil::DoubleToFloatUp(py[i])};
				const uint64_t n = ref_lookup_one(t, st, q, hits.data(), hits.size());
				cand += n;
				for (uint64_t k = 0; k < n; k++) {
					const uint32_t b = hits[k];
					const bool ok = pred == 0 ? (bminx[b] <= px[i] && px[i] <= bmaxx[b] && bminy[b] <= py[i] && py[i] <= bmaxy[b])
					                          : (bminx[b] < px[i] && px[i] < bmaxx[b] && bminy[b] < py[i] && py[i] < bmaxy[b]);
					total += ok;
					if (ok && hash_out) {
						hsum += pair_mix(probe_base + (uint32_t)i, b);
					}
				}
			}
		}
		st.search_queue.~queue();
	}
	if (cand_out) {
		*cand_out = cand;
	}
	if (hash_out) {
		*hash_out = hsum;
	}
	return total;
}
__attribute__((visibility("default"))) uint64_t ref_flat_join_count(void *h, const double *px, const double *py,
                                                                     uint64_t np, const double *bminx,
                                                                     const double *bminy, const double *bmaxx,
                                                                     const double *bmaxy, int pred, int threads,
                                                                     uint64_t *cand_out) {
	return flat_join_impl(h, px, py, np, bminx, bminy, bmaxx, bmaxy, pred, threads, cand_out, 0, nullptr);
}
__attribute__((visibility("default"))) uint64_t ref_flat_join_hash(void *h, const double *px, const double *py,
                                                                    uint64_t np, const double *bminx,
                                                                    const double *bminy, const double *bmaxx,
                                                                    const double *bmaxy, int pred, int threads,
                                                                    uint32_t probe_base, uint64_t *hash_out) {
	return flat_join_impl(h, px, py, np, bminx, bminy, bmaxx, bmaxy, pred, threads, nullptr, probe_base, hash_out);
}

// ST_Extent(WKB_BLOB): ST_Extent::ExecuteWKB, spatial_functions_scalar.cpp:2576-2629 (reader flags as set there).
// returns 1 = box written, 0 = NULL (no vertices), negative = -(sgl error code) where the reference throws.
__attribute__((visibility("default"))) int ref_wkb_extent(const uint8_t *buf, uint64_t len, double *box4, uint64_t *vertex_count) {
	uint32_t stack[128] = {};
	sgl::ops::wkb_reader reader = {};
	reader.allow_mixed_zm = true;
	reader.nan_as_empty = true;
	reader.stack_buf = stack;
	reader.stack_cap = 128;
	reader.buf = reinterpret_cast<const char *>(buf);
	reader.end = reader.buf + len;
	sgl::box_xy bbox = {};
	size_t vc = 0;
	if (!sgl::ops::wkb_reader_try_parse_stats(&reader, &bbox, &vc)) {
		return -(int)reader.error;
	}
	*vertex_count = vc;
	if (vc == 0) {
		return 0;
	}
	box4[0] = bbox.min.x;
	box4[1] = bbox.min.y;
	box4[2] = bbox.max.x;
	box4[3] = bbox.max.y;
	return 1;
}

} // extern "C"
