// join_op.cu -- the probe-side operator glue: PhysicalSpatialJoin's life cycle behind the C-ABI.
//
//   reference (spatial_join_physical.cpp)                       here
//   Sink / Combine / Finalize :431-601 (build side -> FlatRTree)  sb_join_op_create   (BOX_2D build keys -> tree in HBM)
//   ExecuteInternal :695-945, one 2048-row probe chunk per call   sb_join_op_push     (chunks accumulate in pinned memory)
//                                                                 sb_join_op_flush    (one device batch: probe + exact
//                                                                                      predicate, INNER pairs, and for
//                                                                                      LEFT / FULL the unmatched probe rows
//                                                                                      of the batch, :867-938)
//   GetData :1022-1068 (unmatched build rows of RIGHT / FULL)     sb_join_op_finish
//
// The reference probes chunk by chunk because its operator is a streaming one; a GPU probe of 2048 rows is all launch
// latency, so the glue turns the stream into batches of `batch_rows` (CachingOperatorState-style buffering). Probe rows
// are numbered in push order across batches, which is what the caller needs to gather payload columns (:817-829).
#include "internal.cuh"

struct sb_join_op {
	sb_context *ctx = nullptr;
	int pred = 0, join_type = 0;
	uint32_t n_build = 0;
	uint64_t batch_rows = 0;
	sb_flat *tree = nullptr;
	double *d_box2d = nullptr;        // build keys, BOX_2D AoS (exact predicate)
	uint8_t *d_build_matched = nullptr;
	double *h_x = nullptr, *h_y = nullptr; // pinned staging of the pending probe rows
	double *d_x = nullptr, *d_y = nullptr;
	uint8_t *d_probe_matched = nullptr;
	uint64_t pending = 0;
	uint64_t row0 = 0; // probe row id of the first pending row
};

static bool op_left(const sb_join_op *op) {
	return op->join_type == SB_JOIN_LEFT || op->join_type == SB_JOIN_FULL;
}
static bool op_right(const sb_join_op *op) {
	return op->join_type == SB_JOIN_RIGHT || op->join_type == SB_JOIN_FULL;
}

extern "C" void sb_join_op_destroy(sb_join_op *op) {
	if (!op) {
		return;
	}
	if (op->ctx) {
		cudaSetDevice(op->ctx->device);
		cudaStreamSynchronize(op->ctx->stream);
	}
	if (op->tree) {
		sb_flat_destroy(op->tree);
	}
	cudaFree(op->d_box2d);
	cudaFree(op->d_build_matched);
	cudaFree(op->d_x);
	cudaFree(op->d_y);
	cudaFree(op->d_probe_matched);
	cudaFreeHost(op->h_x);
	cudaFreeHost(op->h_y);
	delete op;
}

extern "C" int sb_join_op_create(sb_context *ctx, sb_predicate pred, sb_join_type join_type, const double *bmin_x,
                                 const double *bmin_y, const double *bmax_x, const double *bmax_y, uint32_t n_build,
                                 uint64_t batch_rows, sb_join_op **out) {
	SB_ARG(ctx != nullptr && out != nullptr, "ctx/out");
	SB_CUDA(cudaSetDevice(ctx->device));
	ctx->arena.reset();
	*out = nullptr;
	SB_ARG(join_type >= SB_JOIN_INNER && join_type <= SB_JOIN_FULL, "join_type");
	SB_ARG(pred == SB_PRED_INTERSECTS || pred == SB_PRED_WITHIN, "predicate");
	SB_ARG(batch_rows >= 1 && batch_rows <= 0xFFFFFFFFull, "batch_rows");
	SB_ARG(n_build == 0 || (bmin_x && bmin_y && bmax_x && bmax_y), "build columns");
	sb_join_op *op = new sb_join_op();
	op->ctx = ctx;
	op->pred = (int)pred;
	op->join_type = (int)join_type;
	op->n_build = n_build;
	op->batch_rows = batch_rows;
	int s = SB_OK;
	auto fail = [&](int code) {
		sb_join_op_destroy(op);
		return code;
	};
#define OP_CUDA(call)                                                                                                \
	do {                                                                                                             \
		cudaError_t e_ = (call);                                                                                     \
		if (e_ != cudaSuccess) {                                                                                     \
			sb_set_error("%s failed: %s", #call, cudaGetErrorString(e_));                                            \
			return fail(SB_ERR_CUDA);                                                                                \
		}                                                                                                            \
	} while (0)
	OP_CUDA(cudaMallocHost((void **)&op->h_x, batch_rows * sizeof(double)));
	OP_CUDA(cudaMallocHost((void **)&op->h_y, batch_rows * sizeof(double)));
	OP_CUDA(cudaMalloc((void **)&op->d_x, batch_rows * sizeof(double)));
	OP_CUDA(cudaMalloc((void **)&op->d_y, batch_rows * sizeof(double)));
	if (op_left(op)) {
		OP_CUDA(cudaMalloc((void **)&op->d_probe_matched, batch_rows + 4)); // cleared in 4-byte words
	}
	if (n_build) {
		OP_CUDA(cudaMalloc((void **)&op->d_box2d, 4ull * n_build * sizeof(double)));
		if (op_right(op)) {
			OP_CUDA(cudaMalloc((void **)&op->d_build_matched, n_build));
			OP_CUDA(cudaMemsetAsync(op->d_build_matched, 0, n_build, ctx->stream));
		}
		// Finalize: BOX_2D -> BOX_2DF -> FlatRTree(node_size 32), spatial_join_physical.cpp:544-601
		const double *b0, *b1, *b2, *b3;
		float *fbox = nullptr;
		if ((s = sb_to_device(ctx, SB_HOST, bmin_x, (uint64_t)n_build, &b0)) != SB_OK ||
		    (s = sb_to_device(ctx, SB_HOST, bmin_y, (uint64_t)n_build, &b1)) != SB_OK ||
		    (s = sb_to_device(ctx, SB_HOST, bmax_x, (uint64_t)n_build, &b2)) != SB_OK ||
		    (s = sb_to_device(ctx, SB_HOST, bmax_y, (uint64_t)n_build, &b3)) != SB_OK ||
		    (s = ctx->scratch_t(4 * (uint64_t)n_build, &fbox)) != SB_OK ||
		    (s = sbi_bbox_box2d(ctx, b0, b1, b2, b3, n_build, fbox)) != SB_OK ||
		    (s = sbi_pack_box2d(ctx, b0, b1, b2, b3, n_build, op->d_box2d)) != SB_OK ||
		    (s = sbi_flat_build(ctx, fbox, n_build, 32, &op->tree)) != SB_OK) {
			return fail(s);
		}
	}
	OP_CUDA(cudaStreamSynchronize(ctx->stream));
#undef OP_CUDA
	*out = op;
	return SB_OK;
}

extern "C" int sb_join_op_push(sb_join_op *op, const double *px, const double *py, uint64_t n_rows, uint64_t *pending) {
	SB_ARG(op != nullptr, "op");
	SB_ARG(n_rows == 0 || (px != nullptr && py != nullptr), "px/py");
	if (op->pending + n_rows > op->batch_rows) {
		sb_set_error("join op: %llu rows pending + %llu pushed exceed the batch of %llu rows: flush first",
		             (unsigned long long)op->pending, (unsigned long long)n_rows, (unsigned long long)op->batch_rows);
		return SB_ERR_CAPACITY;
	}
	SB_ARG(op->row0 + op->pending + n_rows <= 0xFFFFFFFFull, "probe row ids must fit 32 bits");
	memcpy(op->h_x + op->pending, px, n_rows * sizeof(double));
	memcpy(op->h_y + op->pending, py, n_rows * sizeof(double));
	op->pending += n_rows;
	if (pending) {
		*pending = op->pending;
	}
	return SB_OK;
}

extern "C" int sb_join_op_flush(sb_join_op *op, uint32_t *out_probe, uint32_t *out_build, uint64_t pair_capacity,
                                uint64_t *n_pairs, uint32_t *out_lhs_unmatched, uint64_t lhs_capacity, uint64_t *n_lhs) {
	SB_ARG(op != nullptr && n_pairs != nullptr, "op/n_pairs");
	sb_context *ctx = op->ctx;
	SB_CUDA(cudaSetDevice(ctx->device));
	ctx->arena.reset();
	*n_pairs = 0;
	if (n_lhs) {
		*n_lhs = 0;
	}
	const uint64_t n = op->pending;
	if (n == 0) {
		return SB_OK;
	}
	SB_ARG(!op_left(op) || (n_lhs != nullptr && (out_lhs_unmatched != nullptr || lhs_capacity == 0)),
	       "LEFT / FULL joins return the unmatched probe rows: out_lhs_unmatched / n_lhs");
	if (op_left(op) && lhs_capacity < n) {
		// every pending row may be unmatched
		*n_lhs = n;
		sb_set_error("join op: lhs_capacity %llu < %llu pending rows", (unsigned long long)lhs_capacity, (unsigned long long)n);
		return SB_ERR_CAPACITY;
	}
	uint64_t got = 0;
	uint32_t *dop = nullptr, *dob = nullptr;
	if (op->n_build) {
		SB_CUDA(cudaMemcpyAsync(op->d_x, op->h_x, n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
		SB_CUDA(cudaMemcpyAsync(op->d_y, op->h_y, n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
		SB_TRY(ctx->scratch_t(pair_capacity ? pair_capacity : 1, &dop));
		SB_TRY(ctx->scratch_t(pair_capacity ? pair_capacity : 1, &dob));
		sb_probe_args a;
		a.px = op->d_x;
		a.py = op->d_y;
		a.m = n;
		a.index_base = (uint32_t)op->row0;
		a.build_box2d = reinterpret_cast<const double4 *>(op->d_box2d);
		a.pred = op->pred;
		const int s = sbi_flat_probe(ctx, op->tree, a, dop, dob, pair_capacity, &got);
		*n_pairs = got;
		if (s != SB_OK) {
			return s; // SB_ERR_CAPACITY: the batch stays pending, *n_pairs is the size to come back with
		}
		SB_ARG(got == 0 || (out_probe != nullptr && out_build != nullptr), "out_probe/out_build");
		SB_TRY(sb_to_host(ctx, SB_HOST, out_probe, dop, got));
		SB_TRY(sb_to_host(ctx, SB_HOST, out_build, dob, got));
	}
	if (op_left(op)) {
		// left_outer_marker of the batch (:867-876) -> EMIT_LHS (:917-938)
		SB_TRY(sbi_zero(ctx, op->d_probe_matched, (n + 3) & ~(uint64_t)3));
	}
	if (got && (op_left(op) || op_right(op))) {
		SB_TRY(sbi_mark_matched(ctx, dop, dob, got, (uint32_t)op->row0, op_left(op) ? op->d_probe_matched : nullptr, n,
		                        op_right(op) ? op->d_build_matched : nullptr, op->n_build));
	}
	if (op_left(op)) {
		uint32_t *d_lhs = nullptr;
		uint64_t c = 0;
		SB_TRY(ctx->scratch_t(n, &d_lhs));
		SB_TRY(sbi_unmatched_rows(ctx, op->d_probe_matched, n, (uint32_t)op->row0, d_lhs, n, &c));
		*n_lhs = c;
		SB_TRY(sb_to_host(ctx, SB_HOST, out_lhs_unmatched, d_lhs, c));
	}
	SB_CUDA(cudaStreamSynchronize(ctx->stream));
	op->row0 += n;
	op->pending = 0;
	return SB_OK;
}

extern "C" int sb_join_op_finish(sb_join_op *op, uint32_t *out_rhs_unmatched, uint64_t capacity, uint64_t *n_rhs) {
	SB_ARG(op != nullptr && n_rhs != nullptr, "op/n_rhs");
	sb_context *ctx = op->ctx;
	SB_CUDA(cudaSetDevice(ctx->device));
	ctx->arena.reset();
	*n_rhs = 0;
	if (op->pending) {
		sb_set_error("join op: %llu probe rows still pending: flush before finish", (unsigned long long)op->pending);
		return SB_ERR_ARG;
	}
	if (!op_right(op) || op->n_build == 0) {
		return SB_OK;
	}
	uint32_t *d_rhs = nullptr;
	uint64_t c = 0;
	SB_TRY(ctx->scratch_t(op->n_build, &d_rhs));
	const int s = sbi_unmatched_rows(ctx, op->d_build_matched, op->n_build, 0, d_rhs, capacity < op->n_build ? capacity : op->n_build, &c);
	*n_rhs = c;
	if (s != SB_OK) {
		return s;
	}
	SB_ARG(c == 0 || out_rhs_unmatched != nullptr, "out_rhs_unmatched");
	SB_TRY(sb_to_host(ctx, SB_HOST, out_rhs_unmatched, d_rhs, c));
	SB_CUDA(cudaStreamSynchronize(ctx->stream));
	return SB_OK;
}

extern "C" uint64_t sb_join_op_rows_pushed(const sb_join_op *op) {
	return op ? op->row0 + op->pending : 0;
}
