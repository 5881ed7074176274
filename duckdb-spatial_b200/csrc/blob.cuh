// blob.cuh -- byte-exact loads from serialized GEOMETRY blobs (blobs packed by the host may be only 1-byte aligned).
// Layout (Serde::Serialize, src/spatial/geometry/geometry_serialization.cpp:132-242):
//   [u8 type][u8 flags][u16 0][u32 0] [4 x f32 bbox if flags & 0x04 (8 bytes per dimension)] then parts
//   [u32 type][u32 count] + POINT / LINESTRING: count vertices; POLYGON: count x u32 ring sizes (padded to 8 bytes),
//   then the rings' vertices; MULTI* / COLLECTION: count nested parts. flags bit 0 = Z, bit 1 = M (vertex = 8 x dims
//   bytes, x and y first); flags & 0xC0 = version (must be 0, geometry_properties.hpp:28-36).
#pragma once
#include <stdint.h>

__device__ __forceinline__ uint32_t ld_u32(const uint8_t *p) {
	if (((uintptr_t)p & 3) == 0) {
		return *reinterpret_cast<const uint32_t *>(p);
	}
	return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}
__device__ __forceinline__ float ld_f32(const uint8_t *p) {
	return __uint_as_float(ld_u32(p));
}
__device__ __forceinline__ double ld_f64(const uint8_t *p) {
	if (((uintptr_t)p & 7) == 0) {
		return *reinterpret_cast<const double *>(p);
	}
	const unsigned long long lo = ld_u32(p), hi = ld_u32(p + 4);
	return __longlong_as_double((long long)(lo | (hi << 32)));
}
