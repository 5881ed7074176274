// ref_scalar.cpp -- TEST INFRASTRUCTURE ONLY (builds oracle/_ref/libref_scalar.so).
//
// The reference's scalar-function translation unit, compiled from where it lies under /root/reference (nothing is
// copied), to reach the pure-arithmetic helpers inside it that need no live DuckDB:
//   ST_Distance::DistanceToSegmentSquared / ClosestPointOnSegment  (spatial_functions_scalar.cpp:2158-2184),
//   the per-segment body of ST_Distance(POINT_2D, LINESTRING_2D).
// Everything else in that file (function registration, executors over DuckDB vectors) is unreachable from the entry
// point below and is dropped by --gc-sections. Used by tests/test_oracle_vs_ref.py to pin
// oracle.c:orc_distance_point_linestring.
#include "/root/reference/src/spatial/modules/main/spatial_functions_scalar.cpp"

extern "C" __attribute__((visibility("default"))) double ref_dist_to_segment_squared(double px, double py, double ax,
                                                                                     double ay, double bx, double by) {
	using namespace duckdb;
	return ST_Distance::DistanceToSegmentSquared<double>(PointXY<double>(px, py), PointXY<double>(ax, ay),
	                                                     PointXY<double>(bx, by));
}
