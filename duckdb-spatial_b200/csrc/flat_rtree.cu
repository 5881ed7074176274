// flat_rtree.cu -- K4b (Hilbert-packed FlatRTree build) and K5 (batched probe).
// Reference: FlatRTree, src/spatial/operators/spatial_join_physical.cpp:72-320.
#include "internal.cuh"

// ---------------------------------------------------------------------------------------------------------------
// Build
// ---------------------------------------------------------------------------------------------------------------
// order-preserving float <-> uint32 (total order, -0 < +0) for atomicMin/atomicMax on floats
__device__ __forceinline__ uint32_t f32_enc(float f) {
	const uint32_t r = __float_as_uint(f);
	return (r & 0x80000000u) ? ~r : (r | 0x80000000u);
}
__device__ __forceinline__ float f32_dec(uint32_t u) {
	return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u);
}

// tree_box = union of all item boxes (FlatRTree::Push, :117-129). enc[4] must start as {~0,~0,0,0}.
__global__ void __launch_bounds__(256) k_flat_bounds(const float4 *__restrict__ box, uint32_t n,
                                                     uint32_t *__restrict__ enc) {
	float u0 = FLT_MAX, u1 = FLT_MAX, u2 = -FLT_MAX, u3 = -FLT_MAX;
	const uint32_t stride = gridDim.x * blockDim.x;
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		const float4 b = box[i];
		u0 = fminf(u0, b.x), u1 = fminf(u1, b.y), u2 = fmaxf(u2, b.z), u3 = fmaxf(u3, b.w);
	}
#pragma unroll
	for (int s = 16; s > 0; s >>= 1) {
		u0 = fminf(u0, __shfl_xor_sync(0xFFFFFFFFu, u0, s));
		u1 = fminf(u1, __shfl_xor_sync(0xFFFFFFFFu, u1, s));
		u2 = fmaxf(u2, __shfl_xor_sync(0xFFFFFFFFu, u2, s));
		u3 = fmaxf(u3, __shfl_xor_sync(0xFFFFFFFFu, u3, s));
	}
	if (sb_lane() == 0) {
		atomicMin(enc + 0, f32_enc(u0));
		atomicMin(enc + 1, f32_enc(u1));
		atomicMax(enc + 2, f32_enc(u2));
		atomicMax(enc + 3, f32_enc(u3));
	}
}

__global__ void k_flat_bounds_finish(const uint32_t *__restrict__ enc, sb_flat_header *__restrict__ hdr) {
	if (threadIdx.x < 4) {
		hdr->tree_box[threadIdx.x] = f32_dec(enc[threadIdx.x]);
	}
}

// Hilbert key, FlatRTree::Build :173-187, all fp32 (max_hilbert is a uint16_t constant promoted to float)
__global__ void __launch_bounds__(256) k_flat_curve(const float4 *__restrict__ box, uint32_t n,
                                                    const sb_flat_header *__restrict__ hdr,
                                                    uint32_t *__restrict__ curve) {
	const float tx = hdr->tree_box[0], ty = hdr->tree_box[1];
	const float hw = __fdiv_rn(65535.0f, __fsub_rn(hdr->tree_box[2], tx));
	const float hh = __fdiv_rn(65535.0f, __fsub_rn(hdr->tree_box[3], ty));
	const uint32_t stride = gridDim.x * blockDim.x;
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
		const float4 b = box[i];
		const float cx = __fdiv_rn(__fadd_rn(b.x, b.z), 2.0f);
		const float cy = __fdiv_rn(__fadd_rn(b.y, b.w), 2.0f);
		const uint32_t hx = sb_cast_f32_u32(__fmul_rn(hw, __fsub_rn(cx, tx)));
		const uint32_t hy = sb_cast_f32_u32(__fmul_rn(hh, __fsub_rn(cy, ty)));
		curve[i] = sb_hilbert16(hx, hy);
	}
}

// leaf level of the flat arrays: box_array[j] = box[perm[j]], idx_array[j] = perm[j] (perm == nullptr: identity)
__global__ void __launch_bounds__(256) k_flat_gather(const float4 *__restrict__ box, const uint32_t *__restrict__ perm,
                                                     uint32_t n, float4 *__restrict__ out_box,
                                                     uint32_t *__restrict__ out_idx) {
	const uint32_t stride = gridDim.x * blockDim.x;
	for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride) {
		const uint32_t s = perm ? perm[j] : j;
		out_box[j] = box[s];
		out_idx[j] = s;
	}
}

// one parent per group of node_size consecutive entries of the layer [beg, end): one warp per parent
__global__ void __launch_bounds__(256) k_flat_pack(float4 *__restrict__ boxes, uint32_t *__restrict__ idx,
                                                   uint32_t beg, uint32_t end, uint32_t node_size, uint32_t out_pos) {
	const uint32_t groups = (end - beg + node_size - 1) / node_size;
	const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
	for (uint32_t g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; g < groups; g += warps) {
		const uint32_t first = beg + g * node_size;
		const uint32_t last = min(first + node_size, end);
		float u0 = FLT_MAX, u1 = FLT_MAX, u2 = -FLT_MAX, u3 = -FLT_MAX;
		for (uint32_t p = first + sb_lane(); p < last; p += 32) {
			const float4 b = boxes[p];
			u0 = fminf(u0, b.x), u1 = fminf(u1, b.y), u2 = fmaxf(u2, b.z), u3 = fmaxf(u3, b.w);
		}
#pragma unroll
		for (int s = 16; s > 0; s >>= 1) {
			u0 = fminf(u0, __shfl_xor_sync(0xFFFFFFFFu, u0, s));
			u1 = fminf(u1, __shfl_xor_sync(0xFFFFFFFFu, u1, s));
			u2 = fmaxf(u2, __shfl_xor_sync(0xFFFFFFFFu, u2, s));
			u3 = fmaxf(u3, __shfl_xor_sync(0xFFFFFFFFu, u3, s));
		}
		if (sb_lane() == 0) {
			boxes[out_pos + g] = make_float4(u0, u1, u2, u3);
			idx[out_pos + g] = first;
		}
	}
}

int sbi_flat_build(sb_context *ctx, const float *box4, uint32_t n, uint32_t node_size, sb_flat **out) {
	SB_ARG(node_size >= 2 && node_size <= 32, "node_size must be in [2, 32]");
	sb_flat *t = new sb_flat();
	*out = t;
	t->device = ctx->device;
	t->owner = ctx;
	memset(&t->h, 0, sizeof(t->h));
	t->h.magic = SB_FLAT_MAGIC;
	t->h.item_count = n;
	t->h.node_size = node_size;
	// layer_bounds, FlatRTree ctor :79-96
	uint64_t count = n, nodes = n;
	t->h.layer_bounds[t->h.nlayers++] = (uint32_t)nodes;
	if (n > 0) {
		do {
			count = (count + node_size - 1) / node_size;
			nodes += count;
			SB_ARG(t->h.nlayers < 16, "too many layers");
			t->h.layer_bounds[t->h.nlayers++] = (uint32_t)nodes;
		} while (count > 1);
	}
	SB_ARG(nodes < ((uint64_t)1 << 27), "FlatRTree: more than 2^27 nodes");
	t->h.nodes = n > 0 ? (uint32_t)nodes : 0;
	t->bytes = sizeof(sb_flat_header) + (size_t)t->h.nodes * (sizeof(float4) + sizeof(uint32_t));
	t->bytes = (t->bytes + 255) & ~(uint64_t)255;
	SB_CUDA(sb_pool_alloc(ctx, &t->buffer, t->bytes));
	t->owns = true;
	SB_CUDA(cudaMemcpyAsync(t->buffer, &t->h, sizeof(t->h), cudaMemcpyHostToDevice, ctx->stream));
	if (n == 0) {
		return SB_OK;
	}
	sb_flat_header *d_hdr = static_cast<sb_flat_header *>(t->buffer);
	float4 *d_boxes = const_cast<float4 *>(t->boxes());
	uint32_t *d_idx = const_cast<uint32_t *>(t->idx());
	const float4 *box = reinterpret_cast<const float4 *>(box4);

	uint32_t *enc = nullptr;
	SB_TRY(ctx->scratch_t(4, &enc));
	static const uint32_t enc_init[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u};
	SB_CUDA(cudaMemcpyAsync(enc, enc_init, sizeof(enc_init), cudaMemcpyHostToDevice, ctx->stream));
	{
		sb_prof_scope ps(ctx, "flat_keys");
		k_flat_bounds<<<sb_grid(n, 256 * 8, 4), 256, 0, ctx->stream>>>(box, n, enc);
		SB_LAUNCH_CHECK(ctx);
		k_flat_bounds_finish<<<1, 32, 0, ctx->stream>>>(enc, d_hdr);
		SB_LAUNCH_CHECK(ctx);
	}
	if (n <= node_size) {
		// :168-171: no sort, the single parent is the tree box (idx stays 0)
		sb_prof_scope ps(ctx, "flat_pack");
		k_flat_gather<<<1, 256, 0, ctx->stream>>>(box, nullptr, n, d_boxes, d_idx);
		SB_LAUNCH_CHECK(ctx);
		k_flat_pack<<<1, 32, 0, ctx->stream>>>(d_boxes, d_idx, 0, n, node_size, n);
		SB_LAUNCH_CHECK(ctx);
		return SB_OK;
	}
	uint32_t *k = nullptr, *kb = nullptr, *v = nullptr, *vb = nullptr;
	SB_TRY(ctx->scratch_t(n, &k));
	SB_TRY(ctx->scratch_t(n, &kb));
	SB_TRY(ctx->scratch_t(n, &v));
	SB_TRY(ctx->scratch_t(n, &vb));
	{
		sb_prof_scope ps(ctx, "flat_keys");
		k_flat_curve<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(box, n, d_hdr, k);
		SB_LAUNCH_CHECK(ctx);
	}
	int alt = 0;
	SB_TRY(sbi_radix_sort_u32(ctx, k, kb, v, vb, n, 0, 32, true, &alt));
	{
		sb_prof_scope ps(ctx, "flat_pack");
		k_flat_gather<<<sb_grid(n, 256, 16), 256, 0, ctx->stream>>>(box, alt ? vb : v, n, d_boxes, d_idx);
		SB_LAUNCH_CHECK(ctx);
		uint32_t beg = 0;
		for (uint32_t layer = 0; layer + 1 < t->h.nlayers; layer++) {
			const uint32_t end = t->h.layer_bounds[layer];
			const uint32_t groups = (end - beg + node_size - 1) / node_size;
			k_flat_pack<<<sb_grid((uint64_t)groups * 32, 256, 8), 256, 0, ctx->stream>>>(d_boxes, d_idx, beg, end,
			                                                                          node_size, end);
			SB_LAUNCH_CHECK(ctx);
			beg = end;
		}
	}
	return SB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Probe (v1): one warp per chunk of 32 probes; (probe lane, node) work items on a per-warp LIFO in shared memory;
// each pop tests the node's <= 32 boxes with one coalesced 512-byte load, ballot + prefix-popcount compaction
// pushes children / stages candidate pairs; staged pairs are flushed with one global atomic per ~200 pairs.
// Traversal starts at the layer just below the root (a child hit implies all its ancestors hit, so the pair set
// is the one FlatRTree::Lookup produces, :262-302).
// ---------------------------------------------------------------------------------------------------------------
#define PR_WARPS 8
#define PR_STACK 320
#define PR_OUT 256

struct flat_meta {
	uint32_t item_count, node_size, nodes, nlayers;
	uint32_t layer_bounds[16];
};

__device__ __forceinline__ uint32_t flat_layer_end(const flat_meta &M, uint32_t pos) {
	// FlatRTree::UpperBound :222-228 -- smallest layer bound > pos. Fully unrolled so that the kernel-parameter
	// array is indexed statically (constant bank, no local-memory copy).
	uint32_t e = 0xFFFFFFFFu;
#pragma unroll
	for (int i = 15; i >= 0; i--) {
		if (i < (int)M.nlayers && M.layer_bounds[i] > pos) {
			e = M.layer_bounds[i];
		}
	}
	return e;
}

template <int MODE> // 0 = float probe boxes, 1 = points (box derived), 2 = points + exact point-in-box refinement
__global__ void __launch_bounds__(PR_WARPS * 32)
    k_flat_probe(const float4 *__restrict__ boxes, const uint32_t *__restrict__ idx, const flat_meta M,
                 const float4 *__restrict__ probe_box, const double *__restrict__ px, const double *__restrict__ py,
                 const double4 *__restrict__ build_box2d, int pred, uint64_t m, uint32_t index_base,
                 const uint32_t *__restrict__ ids, uint32_t *__restrict__ out_probe, uint32_t *__restrict__ out_build, uint64_t capacity,
                 unsigned long long *__restrict__ counter) {
	__shared__ uint32_t s_stack[PR_WARPS][PR_STACK];
	__shared__ uint2 s_out[PR_WARPS][PR_OUT];
	const unsigned warp = threadIdx.x >> 5, lane = sb_lane();
	const unsigned lt = sb_lanemask_lt();
	uint32_t *stack = s_stack[warp];
	uint2 *obuf = s_out[warp];
	uint32_t ocnt = 0;

	uint32_t start = 0; // first entry of the layer just below the root
#pragma unroll
	for (int i = 0; i < 14; i++) {
		if (i + 3 == (int)M.nlayers) {
			start = M.layer_bounds[i];
		}
	}
	const uint64_t chunks = (m + 31) / 32;
	const uint64_t nwarps = (uint64_t)gridDim.x * PR_WARPS;
	for (uint64_t chunk = (uint64_t)blockIdx.x * PR_WARPS + warp; chunk < chunks; chunk += nwarps) {
		const uint64_t pi = chunk * 32 + lane;
		const bool valid = pi < m;
		float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
		double x = 0.0, y = 0.0;
		if (valid) {
			if (MODE == 0) {
				q = probe_box[pi];
			} else {
				x = px[pi];
				y = py[pi];
				q = make_float4(sb_f64_to_f32_down(x), sb_f64_to_f32_down(y), sb_f64_to_f32_up(x), sb_f64_to_f32_up(y));
			}
		}
		const unsigned vm = __ballot_sync(0xFFFFFFFFu, valid);
		int sp = __popc(vm);
		if (valid) {
			stack[lane] = (lane << 27) | start;
		}
		__syncwarp();
		while (sp > 0) {
			const uint32_t item = stack[sp - 1];
			sp--;
			__syncwarp();
			const unsigned pl = item >> 27;
			const uint32_t beg = item & 0x07FFFFFFu;
			const float4 qq = make_float4(__shfl_sync(0xFFFFFFFFu, q.x, pl), __shfl_sync(0xFFFFFFFFu, q.y, pl),
			                              __shfl_sync(0xFFFFFFFFu, q.z, pl), __shfl_sync(0xFFFFFFFFu, q.w, pl));
			const uint32_t end = min(beg + M.node_size, flat_layer_end(M, beg));
			const uint32_t pos = beg + lane;
			bool hit = false;
			uint32_t payload = 0;
			if (pos < end) {
				hit = sb_box_intersects(qq, __ldg(boxes + pos));
				if (hit) {
					payload = __ldg(idx + pos);
				}
			}
			if (beg >= M.item_count) {
				// internal node: push the children (idx = first entry of the child node)
				const unsigned hm = __ballot_sync(0xFFFFFFFFu, hit);
				if (hit) {
					stack[sp + __popc(hm & lt)] = (pl << 27) | payload;
				}
				sp += __popc(hm);
				__syncwarp();
			} else {
				if (MODE == 2) {
					const double xx = __shfl_sync(0xFFFFFFFFu, x, pl), yy = __shfl_sync(0xFFFFFFFFu, y, pl);
					if (hit) {
						const double4 b = build_box2d[payload];
						hit = pred == SB_PRED_INTERSECTS ? (b.x <= xx && xx <= b.z && b.y <= yy && yy <= b.w)
						                                 : (b.x < xx && xx < b.z && b.y < yy && yy < b.w);
					}
				}
				const unsigned hm = __ballot_sync(0xFFFFFFFFu, hit);
				if (hit) {
					obuf[ocnt + __popc(hm & lt)] =
					    make_uint2(ids ? ids[chunk * 32 + pl] : (uint32_t)(chunk * 32 + pl) + index_base, payload);
				}
				ocnt += __popc(hm);
				__syncwarp();
				if (ocnt > PR_OUT - 32) {
					unsigned long long base = 0;
					if (lane == 0) {
						base = atomicAdd(counter, (unsigned long long)ocnt);
					}
					base = __shfl_sync(0xFFFFFFFFu, base, 0);
					for (uint32_t i = lane; i < ocnt; i += 32) {
						if (base + i < capacity) {
							const uint2 p = obuf[i];
							out_probe[base + i] = p.x;
							out_build[base + i] = p.y;
						}
					}
					ocnt = 0;
					__syncwarp();
				}
			}
		}
	}
	if (ocnt > 0) {
		unsigned long long base = 0;
		if (lane == 0) {
			base = atomicAdd(counter, (unsigned long long)ocnt);
		}
		base = __shfl_sync(0xFFFFFFFFu, base, 0);
		for (uint32_t i = lane; i < ocnt; i += 32) {
			if (base + i < capacity) {
				const uint2 p = obuf[i];
				out_probe[base + i] = p.x;
				out_build[base + i] = p.y;
			}
		}
	}
}

int sbi_flat_probe(sb_context *ctx, const sb_flat *tree, const sb_probe_args &a, uint32_t *out_probe,
                   uint32_t *out_build, uint64_t capacity, uint64_t *count) {
	*count = 0;
	if (a.m == 0 || tree->h.item_count == 0) {
		return SB_OK;
	}
	SB_ARG(a.m + a.index_base <= 0xFFFFFFFFull, "probe indices must fit 32 bits");
	// bulk point probes: bucketed kernel (same pair set); small batches and box probes: per-probe tree walk
	if (!a.force_walk && !a.probe_box && ctx->probe_path != 1 && (ctx->probe_path >= 2 || a.m >= (1u << 17))) {
		// v3 (cell-sorted) needs dense sub-cells to fill its warps: ~1 M sub-cells over the tree box, so below a few
		// 10^7 probes per call (e.g. the 16 Mi-row chunks of the host-pointer pipeline) the bucketed v2 kernel is used
		if (ctx->probe_path == 2 || (ctx->probe_path == 0 && a.m < (40ull << 20))) {
			return sbi_flat_probe_bucketed(ctx, tree, a, out_probe, out_build, capacity, count);
		}
		return sbi_flat_probe_cells(ctx, tree, a, out_probe, out_build, capacity, count);
	}
	flat_meta M;
	M.item_count = tree->h.item_count;
	M.node_size = tree->h.node_size;
	M.nodes = tree->h.nodes;
	M.nlayers = tree->h.nlayers;
	memcpy(M.layer_bounds, tree->h.layer_bounds, sizeof(M.layer_bounds));
	unsigned long long *d_counter = nullptr;
	SB_TRY(ctx->scratch_t(1, &d_counter));
	SB_TRY(sbi_zero(ctx, d_counter, sizeof(*d_counter)));
	const unsigned grid = sb_grid((a.m + 31) / 32, PR_WARPS, 8);
	{
		sb_prof_scope ps(ctx, "flat_probe");
		if (a.probe_box) {
			k_flat_probe<0><<<grid, PR_WARPS * 32, 0, ctx->stream>>>(tree->boxes(), tree->idx(), M, a.probe_box, nullptr,
			                                                        nullptr, nullptr, -1, a.m, a.index_base, a.ids, out_probe,
			                                                        out_build, capacity, d_counter);
		} else if (a.pred < 0) {
			k_flat_probe<1><<<grid, PR_WARPS * 32, 0, ctx->stream>>>(tree->boxes(), tree->idx(), M, nullptr, a.px, a.py,
			                                                        nullptr, -1, a.m, a.index_base, a.ids, out_probe, out_build,
			                                                        capacity, d_counter);
		} else {
			k_flat_probe<2><<<grid, PR_WARPS * 32, 0, ctx->stream>>>(tree->boxes(), tree->idx(), M, nullptr, a.px, a.py,
			                                                        a.build_box2d, a.pred, a.m, a.index_base, a.ids, out_probe,
			                                                        out_build, capacity, d_counter);
		}
		SB_LAUNCH_CHECK(ctx);
	}
	SB_TRY(sbi_read_u64(ctx, reinterpret_cast<const uint64_t *>(d_counter), count));
	if (*count > capacity) {
		sb_set_error("probe: %llu pairs, capacity %llu", (unsigned long long)*count, (unsigned long long)capacity);
		return SB_ERR_CAPACITY;
	}
	return SB_OK;
}
