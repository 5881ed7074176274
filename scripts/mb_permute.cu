// Microbenchmark (diagnostic, not product): cost of the one random permute of 24-byte payloads the STR leaf pack needs,
// at C1 size (10 M records), in the forms considered in DESIGN.md: two-array gather (today), 32-byte aligned record
// gather, 32-byte aligned record scatter through the inverse permutation, and the 4-byte inverse scatter itself.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <algorithm>
#include <numeric>
#include <random>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

struct __align__(32) rec32 { float4 box; int64_t row; uint32_t y; uint32_t pad; };

__global__ void k_gather2(const float4 *box, const int64_t *row, const uint32_t *perm, uint32_t n, float4 *ob, int64_t *orow) {
	for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
		const uint32_t s = perm[p];
		ob[p] = box[s];
		orow[p] = row[s];
	}
}
__global__ void k_gather32(const rec32 *rec, const uint32_t *perm, uint32_t n, float4 *ob, int64_t *orow) {
	for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
		const uint32_t s = perm[p];
		const float4 a = *reinterpret_cast<const float4 *>(&rec[s]);
		const float4 b = *(reinterpret_cast<const float4 *>(&rec[s]) + 1);
		ob[p] = a;
		orow[p] = *reinterpret_cast<const int64_t *>(&b);
	}
}
__global__ void k_pack32(const float4 *box, const int64_t *row, uint32_t n, rec32 *rec) {
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
		rec32 r;
		r.box = box[i];
		r.row = row[i];
		r.y = i;
		r.pad = 0;
		float4 *o = reinterpret_cast<float4 *>(&rec[i]);
		o[0] = r.box;
		o[1] = *reinterpret_cast<float4 *>(&r.row);
	}
}
__global__ void k_scatter32(const float4 *box, const int64_t *row, const uint32_t *inv, uint32_t n, rec32 *out) {
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
		rec32 r;
		r.box = box[i];
		r.row = row[i];
		r.y = i;
		r.pad = 0;
		float4 *o = reinterpret_cast<float4 *>(&out[inv[i]]);
		o[0] = r.box;
		o[1] = *reinterpret_cast<float4 *>(&r.row);
	}
}
__global__ void k_inv(const uint32_t *perm, uint32_t n, uint32_t *inv) {
	for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
		inv[perm[p]] = p;
	}
}
__global__ void k_read32(const rec32 *rec, uint32_t n, float4 *ob, int64_t *orow) {
	for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
		const float4 a = *reinterpret_cast<const float4 *>(&rec[p]);
		const float4 b = *(reinterpret_cast<const float4 *>(&rec[p]) + 1);
		ob[p] = a;
		orow[p] = *reinterpret_cast<const int64_t *>(&b);
	}
}

int main(int argc, char **argv) {
	const uint32_t n = argc > 1 ? (uint32_t)atoll(argv[1]) : 10000000u;
	// window: the permutation is random inside consecutive windows of this many records (0 = fully random)
	const uint32_t window = argc > 2 ? (uint32_t)atoll(argv[2]) : 0u;
	std::vector<uint32_t> perm(n);
	std::iota(perm.begin(), perm.end(), 0u);
	std::mt19937_64 rng(7);
	if (window == 0) {
		std::shuffle(perm.begin(), perm.end(), rng);
	} else {
		for (uint32_t b = 0; b < n; b += window) {
			std::shuffle(perm.begin() + b, perm.begin() + std::min<uint64_t>(n, (uint64_t)b + window), rng);
		}
	}
	float4 *box, *ob;
	int64_t *row, *orow;
	uint32_t *dperm, *inv;
	rec32 *rec, *rec2;
	CK(cudaMalloc(&box, 16ull * n));
	CK(cudaMalloc(&ob, 16ull * n));
	CK(cudaMalloc(&row, 8ull * n));
	CK(cudaMalloc(&orow, 8ull * n));
	CK(cudaMalloc(&dperm, 4ull * n));
	CK(cudaMalloc(&inv, 4ull * n));
	CK(cudaMalloc(&rec, 32ull * n));
	CK(cudaMalloc(&rec2, 32ull * n));
	CK(cudaMemset(box, 0, 16ull * n));
	CK(cudaMemset(row, 0, 8ull * n));
	CK(cudaMemcpy(dperm, perm.data(), 4ull * n, cudaMemcpyHostToDevice));
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	const int grid = 148 * 16, block = 256;
	auto run = [&](const char *name, auto f) {
		float best = 1e9f, sum = 0;
		for (int r = 0; r < 6; r++) {
			cudaEventRecord(e0);
			f();
			cudaEventRecord(e1);
			cudaEventSynchronize(e1);
			float ms;
			cudaEventElapsedTime(&ms, e0, e1);
			if (r) { best = std::min(best, ms); sum += ms; }
		}
		printf("%-34s best %8.1f us  mean %8.1f us\n", name, best * 1e3f, sum / 5 * 1e3f);
	};
	printf("n = %u, window = %u\n", n, window);
	run("gather box16 + row8 (today)", [&] { k_gather2<<<grid, block>>>(box, row, dperm, n, ob, orow); });
	run("pack rec32 (streaming)", [&] { k_pack32<<<grid, block>>>(box, row, n, rec); });
	run("gather rec32", [&] { k_gather32<<<grid, block>>>(rec, dperm, n, ob, orow); });
	run("inverse perm scatter (4 B)", [&] { k_inv<<<grid, block>>>(dperm, n, inv); });
	run("scatter rec32 via inverse", [&] { k_scatter32<<<grid, block>>>(box, row, inv, n, rec2); });
	run("read rec32 sequential + write 24", [&] { k_read32<<<grid, block>>>(rec2, n, ob, orow); });
	CK(cudaDeviceSynchronize());
	CK(cudaGetLastError());
	return 0;
}
