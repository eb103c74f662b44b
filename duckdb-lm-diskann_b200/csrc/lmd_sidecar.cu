// Routing sidecars of the device node store: the compressed neighbour vectors an expansion reads, re-laid so that one
// expansion is ONE contiguous, fully coalesced read.
//
// The reference's node block holds, per node, the ids of its R neighbours and compressed copies of THEIR vectors
// (core/index_config.cpp:104-148: ternary planes, SoA [R][words] u64 twice). That is the LM-DiskANN idea - an expansion
// touches one block - but lane j of a warp reading neighbour j's planes from that layout issues 8-byte loads 8 * words
// bytes apart (23 sectors per request at D = 768). The sidecar keeps the same information per node as
//
//     [u32 slot[R]]  [int8 codec only: u32 csq[R], f32 err[R]]  pad -> 16  [chunk-major data: 16 B per (chunk, neighbour)]
//
// ternary: chunk w = word w of the planes, entry = {pos_w, neg_w}          -> data[(w * R + j) * 16]
// int8:    chunk w = dimensions [16 w, 16 w + 16), entry = 16 int8 codes   -> data[(w * R + j) * 16]
//
// so the warp's load of chunk w is 512 contiguous bytes. The mirrored blocks stay byte-identical to the reference layout;
// the sidecars are derived views like d_vecs / d_nbrs and are rebuilt (lazily) whenever those change.
#include "lmd_common.cuh"

#include <algorithm>

using namespace lmd;

namespace {

__host__ __device__ inline uint32_t align16(uint32_t v) {
	return (v + 15u) & ~15u;
}
inline uint32_t align128(uint32_t v) {
	return (v + 127u) & ~127u;
}

// One warp per node: slots from the neighbour-slot view, planes transposed out of the node's own block.
__global__ void build_plane_side_kernel(const uint8_t *__restrict__ blocks, uint64_t block_bytes, uint64_t pos_off,
                                        uint64_t neg_off, uint32_t r, uint32_t words, uint64_t n,
                                        const uint32_t *__restrict__ nbrs, uint8_t *__restrict__ side,
                                        uint32_t side_bytes, uint32_t data_off) {
	const uint64_t node = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
	const uint32_t lane = threadIdx.x & 31;
	if (node >= n) {
		return;
	}
	const uint8_t *blk = blocks + node * block_bytes;
	uint8_t *dst = side + node * (uint64_t)side_bytes;
	for (uint32_t j = lane; j < r; j += 32) {
		reinterpret_cast<uint32_t *>(dst)[j] = nbrs[node * r + j];
	}
	const uint64_t *pp = reinterpret_cast<const uint64_t *>(blk + pos_off);
	const uint64_t *pn = reinterpret_cast<const uint64_t *>(blk + neg_off);
	ulonglong2 *out = reinterpret_cast<ulonglong2 *>(dst + data_off);
	for (uint32_t t = lane; t < r * words; t += 32) {
		const uint32_t w = t / r, j = t - w * r;
		const bool live = nbrs[node * r + j] != LMD_INVALID_SLOT;
		ulonglong2 v;
		v.x = live ? pp[(uint64_t)j * words + w] : 0ull;
		v.y = live ? pn[(uint64_t)j * words + w] : 0ull;
		out[t] = v;
	}
}

// Per-dimension sum and sum of squares over all node vectors (doubles; one atomic per (block, dimension)).
__global__ void code_stats_kernel(const float *__restrict__ vecs, uint32_t row_floats, uint32_t dims, uint64_t n,
                                  uint64_t rows_per_block, double *__restrict__ sum, double *__restrict__ sumsq) {
	const uint64_t r0 = blockIdx.x * rows_per_block;
	const uint64_t r1 = r0 + rows_per_block < n ? r0 + rows_per_block : n;
	for (uint32_t d = threadIdx.x; d < dims; d += blockDim.x) {
		double a = 0.0, b = 0.0;
		for (uint64_t row = r0; row < r1; row++) {
			const double v = (double)vecs[row * row_floats + d];
			a += v;
			b += v * v;
		}
		atomicAdd(&sum[d], a);
		atomicAdd(&sumsq[d], b);
	}
}

// COSINE stores put the UNIT rows x / || x || on the grid (the metric does not see the length): per-dimension sum of
// squares of the unit rows. One warp per row, per-block partial sums in shared memory.
__global__ void code_stats_unit_kernel(const float *__restrict__ vecs, uint32_t row_floats, uint32_t dims, uint64_t n,
                                       uint64_t rows_per_block, double *__restrict__ sumsq) {
	extern __shared__ float s_acc[];
	for (uint32_t d = threadIdx.x; d < row_floats; d += blockDim.x) {
		s_acc[d] = 0.0f;
	}
	__syncthreads();
	const uint64_t r0 = blockIdx.x * rows_per_block;
	const uint64_t r1 = r0 + rows_per_block < n ? r0 + rows_per_block : n;
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
	for (uint64_t row = r0 + warp; row < r1; row += warps) {
		const float *v = vecs + row * row_floats;
		float xx = 0.0f;
		for (uint32_t d = lane; d < dims; d += 32) {
			xx = fmaf(v[d], v[d], xx);
		}
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) {
			xx += __shfl_xor_sync(0xFFFFFFFFu, xx, o);
		}
		if (xx > 0.0f) {
			const float inv = 1.0f / xx;
			for (uint32_t d = lane; d < dims; d += 32) {
				atomicAdd(&s_acc[d], v[d] * v[d] * inv);
			}
		}
	}
	__syncthreads();
	for (uint32_t d = threadIdx.x; d < dims; d += blockDim.x) {
		atomicAdd(&sumsq[d], (double)s_acc[d]);
	}
}

// off[d] = mean of dimension d; off[row_floats] = scale = sigmas * max_d sigma_d / 127 (one grid for all dimensions, so
// that distances on the grid are plain integer arithmetic). Values beyond +-127 steps are clamped; every neighbour's
// own reconstruction error is stored next to its codes, so clamping costs bound tightness, never correctness.
// `centred` = 0 (COSINE stores): the grid has no offsets - x ~ scale * c - because a dot product, unlike a difference,
// does not cancel them; sigma_d is then the root of the second moment about zero.
__global__ void code_params_kernel(const double *__restrict__ sum, const double *__restrict__ sumsq, uint64_t n,
                                   uint32_t dims, uint32_t row_floats, float sigmas, int centred, float *__restrict__ off) {
	__shared__ float s_max[256];
	__shared__ double s_tot[256];
	float mx = 0.0f;
	double tot = 0.0; // sum over d of E[x_d^2] = mean squared row length
	for (uint32_t d = threadIdx.x; d < row_floats; d += blockDim.x) {
		float o = 0.0f;
		if (d < dims) {
			tot += sumsq[d] / (double)n;
			const double m = centred ? sum[d] / (double)n : 0.0;
			double var = sumsq[d] / (double)n - m * m;
			var = var > 0.0 ? var : 0.0;
			o = (float)m;
			mx = fmaxf(mx, (float)sqrt(var));
		}
		off[d] = o;
	}
	s_max[threadIdx.x] = mx;
	s_tot[threadIdx.x] = tot;
	__syncthreads();
	for (uint32_t st = blockDim.x / 2; st > 0; st >>= 1) {
		if (threadIdx.x < st) {
			s_max[threadIdx.x] = fmaxf(s_max[threadIdx.x], s_max[threadIdx.x + st]);
			s_tot[threadIdx.x] += s_tot[threadIdx.x + st];
		}
		__syncthreads();
	}
	if (threadIdx.x == 0) {
		const float sc = sigmas * s_max[0] / 127.0f;
		off[row_floats] = sc > 0.0f ? sc : 1.0f;
		off[row_floats + 1] = (float)sqrt(s_tot[0]); // rho: the RMS row length (IP stores scale the query to it)
	}
}

// One warp per node: for each neighbour j its vector is brought onto the grid, c = clamp(rint((x - off) / scale)),
// csq = sum c^2 (integer), err = || (x - off) - scale * c || (fp32, inflated by 1e-4 so that it is an upper bound).
// `unit` = 1 (COSINE stores): the UNIT row x / || x || goes onto the (zero-offset) grid, err = || x / || x || - scale * c ||;
// a zero row gets codes 0 and err = +inf, so that no bound ever excludes it (the reference gives it distance 1.0).
__global__ void build_code_side_kernel(const float *__restrict__ vecs, uint32_t row_floats, uint32_t dims, uint32_t r,
                                       uint64_t n, const uint32_t *__restrict__ nbrs, const float *__restrict__ off,
                                       uint8_t *__restrict__ side, uint32_t side_bytes, uint32_t data_off, int unit) {
	const uint64_t node = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
	const uint32_t lane = threadIdx.x & 31;
	if (node >= n) {
		return;
	}
	const float scale = off[row_floats], inv = 1.0f / scale;
	uint8_t *dst = side + node * (uint64_t)side_bytes;
	uint32_t *head = reinterpret_cast<uint32_t *>(dst);
	for (uint32_t j = 0; j < r; j++) {
		const uint32_t slot = nbrs[node * r + j];
		int csq = 0;
		float e = 0.0f, xmul = 1.0f;
		if (unit) {
			float xx = 0.0f;
			if (slot != LMD_INVALID_SLOT) {
				for (uint32_t f = lane; f < dims; f += 32) {
					const float v = vecs[(uint64_t)slot * row_floats + f];
					xx = fmaf(v, v, xx);
				}
			}
#pragma unroll
			for (int o = 16; o > 0; o >>= 1) {
				xx += __shfl_xor_sync(0xFFFFFFFFu, xx, o);
			}
			xmul = xx > 0.0f ? 1.0f / sqrtf(xx) : 0.0f;
			if (!(xx > 0.0f) || !isfinite(xmul)) {
				xmul = 0.0f;
				e = __int_as_float(0x7f800000); // + inf: never excluded
			}
		}
		for (uint32_t f = lane * 4; f < row_floats; f += 128) { // row_floats is a multiple of 32 floats
			int c[4] = {0, 0, 0, 0};
			if (slot != LMD_INVALID_SLOT) {
				const float4 x = *reinterpret_cast<const float4 *>(vecs + (uint64_t)slot * row_floats + f);
				const float xv[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
				for (int t = 0; t < 4; t++) {
					if (f + t < dims) {
						const float v = xv[t] * xmul - off[f + t];
						int q = __float2int_rn(v * inv);
						q = q < -127 ? -127 : (q > 127 ? 127 : q);
						const float rr = v - scale * (float)q;
						e = fmaf(rr, rr, e);
						csq += q * q;
						c[t] = q;
					}
				}
			}
			const uint32_t packed = (uint32_t)(uint8_t)(int8_t)c[0] | ((uint32_t)(uint8_t)(int8_t)c[1] << 8) |
			                        ((uint32_t)(uint8_t)(int8_t)c[2] << 16) | ((uint32_t)(uint8_t)(int8_t)c[3] << 24);
			// dimensions [f, f + 4) live in chunk f / 16, at byte f % 16 of neighbour j's 16-byte entry
			*reinterpret_cast<uint32_t *>(dst + data_off + ((uint64_t)(f / 16) * r + j) * 16 + (f % 16)) = packed;
		}
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) {
			csq += __shfl_xor_sync(0xFFFFFFFFu, csq, o);
			e += __shfl_xor_sync(0xFFFFFFFFu, e, o);
		}
		if (lane == 0) {
			head[j] = slot;
			head[r + j] = (uint32_t)csq;
			head[2 * r + j] = __float_as_uint(sqrtf(e) * 1.0001f);
		}
	}
}

lmd_status alloc_side(lmd_store *s, uint8_t *&ptr, uint64_t bytes, const char *what) {
	if (ptr) {
		return LMD_OK;
	}
	cudaError_t e = cudaMalloc((void **)&ptr, bytes);
	if (e != cudaSuccess) {
		ptr = nullptr;
		cudaGetLastError();
		set_error("cudaMalloc(%llu bytes) for %s failed: %s", (unsigned long long)bytes, what, cudaGetErrorString(e));
		return LMD_ERR_OOM;
	}
	s->device_bytes += bytes;
	return LMD_OK;
}

lmd_status build_plane_side(lmd_store *s) {
	if (s->blocks_released) {
		set_error("the ternary sidecar is stale and the block mirror was released: it cannot be rebuilt");
		return LMD_ERR_STATE;
	}
	s->side_t_off = align16(4 * s->cfg.r);
	s->side_t_bytes = align128(s->side_t_off + s->cfg.r * s->words * 16);
	lmd_status rc = alloc_side(s, s->d_side_t, s->capacity * (uint64_t)s->side_t_bytes, "ternary sidecar");
	if (rc != LMD_OK) {
		return rc;
	}
	build_plane_side_kernel<<<(unsigned)((s->n * 32 + 255) / 256), 256, 0, s->stream>>>(
	    s->d_blocks, s->cfg.block_bytes, s->lay.pos_off, s->lay.neg_off, s->cfg.r, s->words, s->n, s->d_nbrs, s->d_side_t,
	    s->side_t_bytes, s->side_t_off);
	LMD_CUDA_OK(cudaGetLastError());
	LMD_CUDA_OK(cudaStreamSynchronize(s->stream));
	s->side_t_gen = s->views_gen;
	return LMD_OK;
}

lmd_status build_code_side(lmd_store *s) {
	const uint32_t rf = s->row_floats, r = s->cfg.r;
	s->code_data_off = align16(12 * r);
	s->code_block_bytes = align128(s->code_data_off + r * rf);
	lmd_status rc = alloc_side(s, s->d_codes, s->capacity * (uint64_t)s->code_block_bytes, "int8 code sidecar");
	if (rc != LMD_OK) {
		return rc;
	}
	if (!s->d_code_off) {
		LMD_CUDA_OK(cudaMalloc((void **)&s->d_code_off, (rf + 2) * sizeof(float)));
	}
	double *d_stats = nullptr;
	LMD_CUDA_OK(cudaMalloc((void **)&d_stats, 2 * (size_t)rf * sizeof(double)));
	cudaMemsetAsync(d_stats, 0, 2 * (size_t)rf * sizeof(double), s->stream);
	const uint64_t rows_per_block = std::max<uint64_t>(64, (s->n + 8 * 148 - 1) / (8 * 148));
	const unsigned nb = (unsigned)((s->n + rows_per_block - 1) / rows_per_block);
	if (s->cfg.metric == LMD_METRIC_COSINE) {
		code_stats_unit_kernel<<<nb, 256, rf * sizeof(float), s->stream>>>(s->d_vecs, rf, s->cfg.dims, s->n, rows_per_block,
		                                                                   d_stats + rf);
	} else {
		code_stats_kernel<<<nb, 256, 0, s->stream>>>(s->d_vecs, rf, s->cfg.dims, s->n, rows_per_block, d_stats, d_stats + rf);
	}
	const float sigmas = s->tune.code_sigmas > 0 ? (float)s->tune.code_sigmas * 0.1f : 4.0f;
	const int cosine = s->cfg.metric == LMD_METRIC_COSINE ? 1 : 0;
	// L2: per-dimension means as offsets (a difference cancels them); COSINE / IP: a dot product does not, zero offsets
	code_params_kernel<<<1, 256, 0, s->stream>>>(d_stats, d_stats + rf, s->n, s->cfg.dims, rf, sigmas,
	                                             s->cfg.metric == LMD_METRIC_L2 ? 1 : 0, s->d_code_off);
	build_code_side_kernel<<<(unsigned)((s->n * 32 + 255) / 256), 256, 0, s->stream>>>(
	    s->d_vecs, rf, s->cfg.dims, r, s->n, s->d_nbrs, s->d_code_off, s->d_codes, s->code_block_bytes, s->code_data_off,
	    cosine);
	float h_par[2] = {1.0f, 0.0f}; // scale, rho
	cudaError_t e = cudaGetLastError();
	if (e == cudaSuccess) {
		e = cudaMemcpyAsync(h_par, s->d_code_off + rf, 2 * sizeof(float), cudaMemcpyDeviceToHost, s->stream);
	}
	if (e == cudaSuccess) {
		e = cudaStreamSynchronize(s->stream);
	}
	if (e == cudaSuccess) {
		s->code_scale = h_par[0];
		s->code_rho = h_par[1];
	}
	cudaFree(d_stats);
	if (e != cudaSuccess) {
		set_error("building the int8 code sidecar failed: %s", cudaGetErrorString(e));
		return LMD_ERR_CUDA;
	}
	s->codes_gen = s->views_gen;
	return LMD_OK;
}

} // namespace

namespace lmd {

bool route_ready(const lmd_store *s, uint32_t route) {
	if (s->dirty) {
		return false;
	}
	if ((route & LMD_PREP_NORMS) && (!s->d_norms || s->norms_gen != s->views_gen)) {
		return false;
	}
	route &= ~LMD_PREP_NORMS;
	if (route == LMD_ROUTE_TERNARY && s->side_t_gen != s->views_gen) {
		return false;
	}
	if ((route == LMD_ROUTE_INT8 || s->codec == LMD_CODEC_INT8) && s->codes_gen != s->views_gen) {
		return false;
	}
	return true;
}

lmd_status store_prepare(lmd_store *s, uint32_t route) {
	std::unique_lock<std::shared_mutex> g(s->mu);
	LMD_CUDA_OK(cudaSetDevice(s->device));
	lmd_status rc = finalize_locked(s);
	if (rc != LMD_OK || s->n == 0) {
		return rc;
	}
	if ((route & LMD_PREP_NORMS) && (!s->d_norms || s->norms_gen != s->views_gen) && (rc = build_row_norms(s)) != LMD_OK) {
		return rc;
	}
	route &= ~LMD_PREP_NORMS;
	if (route == LMD_ROUTE_TERNARY && s->side_t_gen != s->views_gen && (rc = build_plane_side(s)) != LMD_OK) {
		return rc;
	}
	if (route == LMD_ROUTE_INT8) {
		s->codec = LMD_CODEC_INT8;
	}
	if (s->codec == LMD_CODEC_INT8 && s->codes_gen != s->views_gen && (rc = build_code_side(s)) != LMD_OK) {
		return rc;
	}
	return LMD_OK;
}

} // namespace lmd

extern "C" {

lmd_status lmd_store_enable_codec(lmd_store *s, uint32_t codec) {
	if (!s) {
		set_error("lmd_store_enable_codec: null store");
		return LMD_ERR_INVALID_ARG;
	}
	if (codec != LMD_CODEC_NONE && codec != LMD_CODEC_INT8) {
		set_error("lmd_store_enable_codec: unknown codec %u", codec);
		return LMD_ERR_INVALID_ARG;
	}
	// L2: the grid distance || q^ - x^ || bounds and approximates the distance (pre-filter and LMD_ROUTE_INT8);
	// COSINE / IP: the grid dot product bounds it (pre-filter of the exact route only)
	{
		std::unique_lock<std::shared_mutex> g(s->mu);
		s->codec = codec;
		if (codec == LMD_CODEC_NONE) {
			if (s->d_codes) {
				LMD_CUDA_OK(cudaSetDevice(s->device));
				cudaStreamSynchronize(s->stream);
				cudaFree(s->d_codes);
				s->d_codes = nullptr;
				s->device_bytes -= s->capacity * (uint64_t)s->code_block_bytes;
				s->codes_gen = 0;
			}
			return LMD_OK;
		}
	}
	return store_prepare(s, LMD_ROUTE_EXACT);
}

lmd_status lmd_store_get_codec(lmd_store *s, float *out_offsets, float *out_scale) {
	if (!s || !out_offsets || !out_scale) {
		set_error("lmd_store_get_codec: null argument");
		return LMD_ERR_INVALID_ARG;
	}
	std::shared_lock<std::shared_mutex> g(s->mu);
	if (!s->d_codes || s->codes_gen != s->views_gen) {
		set_error("lmd_store_get_codec: no up-to-date int8 codec on this store (lmd_store_enable_codec)");
		return LMD_ERR_STATE;
	}
	LMD_CUDA_OK(cudaSetDevice(s->device));
	LMD_CUDA_OK(cudaMemcpy(out_offsets, s->d_code_off, s->cfg.dims * sizeof(float), cudaMemcpyDeviceToHost));
	*out_scale = s->code_scale;
	return LMD_OK;
}

lmd_status lmd_store_release_blocks(lmd_store *s) {
	if (!s) {
		set_error("lmd_store_release_blocks: null store");
		return LMD_ERR_INVALID_ARG;
	}
	// every view that is derived from the blocks is built first (the ternary sidecar only where the metric allows it)
	lmd_status rc = store_prepare(s, s->cfg.metric == LMD_METRIC_L2 ? LMD_ROUTE_EXACT : LMD_ROUTE_TERNARY);
	if (rc != LMD_OK) {
		return rc;
	}
	std::unique_lock<std::shared_mutex> g(s->mu);
	if (s->blocks_released) {
		return LMD_OK;
	}
	LMD_CUDA_OK(cudaSetDevice(s->device));
	LMD_CUDA_OK(cudaStreamSynchronize(s->stream));
	cudaFree(s->d_blocks);
	s->d_blocks = nullptr;
	s->device_bytes -= s->capacity * s->cfg.block_bytes;
	s->blocks_released = true;
	return LMD_OK;
}

} // extern "C"
