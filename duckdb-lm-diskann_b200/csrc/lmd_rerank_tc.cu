// Rerank as a tensor-core contraction (north_star: "3xTF32 / split-fp32, staged through bulk
// async copies"). Replaces core/Searcher.cpp:189-245 like rerank_kernel does; selected only by LMD_SEARCH_TENSOR_RERANK -
// the FMA gather kernel already runs at the HBM roofline (profiles/r2_03_rerank768_ncu_details.csv), so this variant is
// the one SURVEY.md section 7 hard part 7 asks to be built and MEASURED next to it, not the default.
//
// One CTA per SM, persistent over (query, tile of 128 candidates). Per tile, K is walked in stages of 64 dimensions:
//
//   producer warp    one cp.async.bulk per candidate row and stage (256 contiguous bytes of d_vecs) into a 2-slot staging
//                    ring, completion counted by an mbarrier (complete_tx)
//   4 worker warps   read the staged fp32 values, split each into hi = the top 19 bits (a TF32 value) and lo = a - hi,
//                    and store both into K-major 128-byte-swizzled operand tiles A_hi / A_lo [128 candidates x 64]
//   one thread       tcgen05.mma kind::tf32, M = 128, N = 16, K = 8:  D[128 x 16] += A_hi * B^T,  D += A_lo * B^T  with
//                    B row 0 = q_hi, row 1 = q_lo (rows 2..15 zero), so that D[:,0] + D[:,1] = (a_hi + a_lo)(q_hi + q_lo):
//                    the split-fp32 product with all four terms from two MMAs per K step; accumulators live in TMEM.
//                    The tensor core adds into its fp32 accumulator by TRUNCATION: one accumulator for all 192 MMAs of a
//                    768-d row came out 2^-17 low (measured, first version). So every K block of 32 dimensions
//                    accumulates into its own pair of columns (8 MMAs deep) at a pitch of 2 columns - the 14 columns an
//                    MMA writes behind its pair only ever receive exact zeros (B rows 2..15) before the next block's
//                    first MMA overwrites them - and the blocks are added in fp32 by the epilogue.
//   epilogue         tcgen05.ld of the column pairs, their fp32 sum, metric tail with the row's squared norm (per-node view
//                    d_norms) and the query's, keys into shared memory, k rounds of min extraction by warp 0
//
// L2 uses |q|^2 + |x|^2 - 2 q.x, which cancels for near-duplicate rows (SURVEY.md section 7 hard part 4): distances agree
// with rerank_kernel to 1e-5 relative on the test distributions, not for d -> 0; tests/test_gpu_round2.py states the bound.
#include "lmd_kernels.cuh"

namespace lmd {

namespace {

constexpr uint32_t TC_M = 128;                      // candidates per tile (TMEM lanes)
constexpr uint32_t TC_KS = 64;                      // dimensions per stage
constexpr uint32_t TC_ROW_BYTES = TC_KS * 4;        // 256: one bulk copy
constexpr uint32_t TC_KB_BYTES = TC_M * 128;        // one K block (32 dims) of an A tile: 128 rows x 128 bytes
constexpr uint32_t TC_A_BYTES = 2 * TC_KB_BYTES;    // 32 KB per A tile (64 dims)
constexpr uint32_t TC_STAGE_BYTES = TC_M * TC_ROW_BYTES; // 32 KB per staging slot
constexpr uint32_t TC_B_KB_BYTES = 16 * 128;        // one K block of B: 16 rows x 128 bytes
constexpr uint32_t TC_WORKERS = 128;
constexpr uint32_t TC_THREADS = TC_WORKERS + 32;
constexpr uint32_t TC_TMEM_COLS = 256;                // accumulator groups: 2 columns per K block of 32 dims (+14)
constexpr uint32_t TC_MAX_KB = (TC_TMEM_COLS - 16) / 2; // 120 K blocks = 3 840 dimensions
constexpr uint32_t TC_MAX_L = 512;

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
	return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
	uint32_t done;
	do {
		asm volatile("{\n\t"
		             ".reg .pred p;\n\t"
		             "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
		             "selp.u32 %0, 1, 0, p;\n\t"
		             "}"
		             : "=r"(done)
		             : "r"(bar), "r"(parity)
		             : "memory");
	} while (!done);
}
__device__ __forceinline__ void bulk_copy(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
	             "l"(src), "r"(bytes), "r"(bar)
	             : "memory");
}
__device__ __forceinline__ void workers_sync() { // the 128 worker threads only (barrier 1); the producer never joins
	asm volatile("bar.sync 1, 128;" ::: "memory");
}

// K-major, 128-byte swizzle (cute UMMA::SmemDescriptor: start [0,14), LBO [16,30), SBO [32,46), version [46,48) = 1,
// layout [61,64) = 2): rows 128 bytes apart, groups of 8 rows 1024 bytes apart, 16-byte chunk c of row r stored at c ^ (r & 7)
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr) {
	return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
	       ((uint64_t)2 << 61);
}
// cute UMMA::InstrDescriptor: c_format [4,6) = 1 (F32), a/b_format [7,10) / [10,13) = 2 (TF32), a/b major K (0),
// n_dim [17,23) = N >> 3, m_dim [24,29) = M >> 4
constexpr uint32_t TC_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((16u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t a, uint64_t b, uint32_t accumulate) {
	asm volatile("{\n\t"
	             ".reg .pred p;\n\t"
	             "setp.ne.b32 p, %4, 0;\n\t"
	             "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
	             "}" ::"r"(tmem_d),
	             "l"(a), "l"(b), "r"(TC_IDESC), "r"(accumulate)
	             : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
	asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void split_tf32(float a, float &hi, float &lo) {
	hi = __uint_as_float(__float_as_uint(a) & 0xFFFFE000u); // 1 + 8 + 10 bits kept: exactly representable in TF32
	lo = a - hi;                                            // exact in fp32; the MMA keeps its top 10 mantissa bits
}

struct TcParams {
	const float *__restrict__ vecs;
	const float *__restrict__ norms; // [n] squared norm of every row (fp32 sum)
	const int64_t *__restrict__ row_ids;
	const float *__restrict__ queries;
	const uint64_t *__restrict__ cand;
	int64_t *out_ids;
	float *out_dist;
	uint32_t *out_count;
	unsigned long long *counters;
	uint32_t nq, dims, row_floats, q_stride, L, k, metric;
};

template <int METRIC>
__global__ void __launch_bounds__(TC_THREADS, 1) rerank_tc_kernel(const TcParams p) {
	extern __shared__ __align__(1024) uint8_t smem[];
	const uint32_t nkb = p.row_floats / 32;          // K blocks of the query
	const uint32_t nst = p.row_floats / TC_KS;       // stages per tile
	uint8_t *a_hi = smem;
	uint8_t *a_lo = smem + TC_A_BYTES;
	uint8_t *b_op = smem + 2 * TC_A_BYTES;
	uint8_t *stage = b_op + nkb * TC_B_KB_BYTES;
	uint64_t *s_keys = reinterpret_cast<uint64_t *>(stage + 2 * TC_STAGE_BYTES); // [TC_MAX_L]
	uint32_t *s_slot = reinterpret_cast<uint32_t *>(s_keys + TC_MAX_L);          // [128] slots of the tile
	float *s_red = reinterpret_cast<float *>(s_slot + TC_M);                     // [4]
	uint32_t *s_cnt = reinterpret_cast<uint32_t *>(s_red + 4);                   // [4]
	uint32_t *s_tmem = s_cnt + 4;                                                // [1] TMEM base address
	uint64_t *s_bar = reinterpret_cast<uint64_t *>(s_tmem + 2);                  // full[2], empty[2], tiles_free, accum_full
	const uint32_t bar_full = smem_u32(s_bar), bar_empty = smem_u32(s_bar + 2), bar_tiles = smem_u32(s_bar + 4),
	               bar_accum = smem_u32(s_bar + 5);
	const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
	const uint32_t ntiles = (p.L + TC_M - 1) / TC_M;

	if (tid == 0) {
		mbar_init(bar_full, 1);
		mbar_init(bar_full + 8, 1);
		mbar_init(bar_empty, TC_WORKERS);
		mbar_init(bar_empty + 8, TC_WORKERS);
		mbar_init(bar_tiles, 1);
		mbar_init(bar_accum, 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	if (warp == 0) { // one warp allocates the accumulator columns (and frees them at the end)
		asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "n"(TC_TMEM_COLS)
		             : "memory");
		asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
	}
	// rows 2..15 of every B block stay zero for the whole kernel
	for (uint32_t i = tid; i < nkb * TC_B_KB_BYTES / 16; i += TC_THREADS) {
		reinterpret_cast<uint4 *>(b_op)[i] = make_uint4(0u, 0u, 0u, 0u);
	}
	asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
	__syncthreads();
	asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
	const uint32_t tmem_d = *s_tmem;

	if (warp == 4) {
		// ================= producer: bulk copies of the candidate rows, one stage ahead of the workers =================
		uint32_t g = 0;
		for (uint32_t q = blockIdx.x; q < p.nq; q += gridDim.x) {
			for (uint32_t tile = 0; tile < ntiles; tile++) {
				uint32_t slots[4];
				uint32_t m = 0;
#pragma unroll
				for (int j = 0; j < 4; j++) {
					const uint32_t i = tile * TC_M + j * 32 + lane;
					const uint64_t key = i < p.L ? p.cand[(uint64_t)q * p.L + i] : LMD_KEY_NONE;
					slots[j] = key_slot(key);
					m += __popc(__ballot_sync(FULL, key != LMD_KEY_NONE));
				}
				if (m == 0) {
					break; // the list is dense from the front: no later tile has candidates either
				}
				for (uint32_t st = 0; st < nst; st++, g++) {
					const uint32_t slot = g & 1;
					mbar_wait(bar_empty + slot * 8, ((g >> 1) & 1) ^ 1);
					if (lane == 0) {
						mbar_expect_tx(bar_full + slot * 8, m * TC_ROW_BYTES);
					}
					__syncwarp();
#pragma unroll
					for (int j = 0; j < 4; j++) {
						const uint32_t r = j * 32 + lane;
						if (r < m) {
							bulk_copy(smem_u32(stage + slot * TC_STAGE_BYTES + r * TC_ROW_BYTES),
							          p.vecs + (uint64_t)slots[j] * p.row_floats + st * TC_KS, TC_ROW_BYTES, bar_full + slot * 8);
						}
					}
				}
			}
		}
	} else {
		// ================= workers: operand split, MMA issue (thread 0), epilogue =================
		uint32_t g = 0, n_accum = 0;
		for (uint32_t q = blockIdx.x; q < p.nq; q += gridDim.x) {
			// ---- B operand: row 0 = q_hi, row 1 = q_lo of every K block; |q|^2 ----
			const float *qv = p.queries + (uint64_t)q * p.q_stride;
			float q2 = 0.0f;
			for (uint32_t j = tid; j < p.row_floats / 4; j += TC_WORKERS) {
				float v[4], hi[4], lo[4];
#pragma unroll
				for (int e = 0; e < 4; e++) {
					const uint32_t f = j * 4 + e;
					v[e] = f < p.dims ? __ldg(qv + f) : 0.0f;
					q2 = fmaf(v[e], v[e], q2);
					split_tf32(v[e], hi[e], lo[e]);
				}
				const uint32_t kb = j >> 3, c = j & 7;
				uint8_t *blk = b_op + kb * TC_B_KB_BYTES;
				*reinterpret_cast<float4 *>(blk + (c << 4)) = make_float4(hi[0], hi[1], hi[2], hi[3]);             // row 0
				*reinterpret_cast<float4 *>(blk + 128 + ((c ^ 1) << 4)) = make_float4(lo[0], lo[1], lo[2], lo[3]); // row 1
			}
			q2 = team_sum<32>(q2);
			if (lane == 0) {
				s_red[warp] = q2;
			}
			asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
			workers_sync();
			q2 = s_red[0] + s_red[1] + s_red[2] + s_red[3];
			uint32_t total = 0;

			for (uint32_t tile = 0; tile < ntiles; tile++) {
				const uint32_t i = tile * TC_M + tid;
				const uint64_t ckey = i < p.L ? p.cand[(uint64_t)q * p.L + i] : LMD_KEY_NONE;
				const bool valid = ckey != LMD_KEY_NONE;
				const uint32_t wcnt = __popc(__ballot_sync(FULL, valid));
				workers_sync(); // the previous tile's readers of s_cnt / s_slot are done
				if (lane == 0) {
					s_cnt[warp] = wcnt;
				}
				s_slot[tid] = key_slot(ckey);
				workers_sync();
				const uint32_t m = s_cnt[0] + s_cnt[1] + s_cnt[2] + s_cnt[3];
				if (m == 0) {
					break;
				}
				for (uint32_t st = 0; st < nst; st++, g++) {
					const uint32_t slot = g & 1;
					mbar_wait(bar_full + slot * 8, (g >> 1) & 1);
					if (g > 0) {
						mbar_wait(bar_tiles, (g - 1) & 1); // the MMAs that read the operand tiles have completed
					}
					const uint8_t *src = stage + slot * TC_STAGE_BYTES;
#pragma unroll 4
					for (uint32_t it = 0; it < 16; it++) {
						const uint32_t gi = it * TC_WORKERS + tid, row = gi >> 4, c = gi & 15;
						float4 v = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
						if (row < m) {
							v = *reinterpret_cast<const float4 *>(src + row * TC_ROW_BYTES + (c << 4));
						}
						float4 h, l;
						split_tf32(v.x, h.x, l.x);
						split_tf32(v.y, h.y, l.y);
						split_tf32(v.z, h.z, l.z);
						split_tf32(v.w, h.w, l.w);
						const uint32_t off = (c >> 3) * TC_KB_BYTES + (row >> 3) * 1024 + (row & 7) * 128 + (((c & 7) ^ (row & 7)) << 4);
						*reinterpret_cast<float4 *>(a_hi + off) = h;
						*reinterpret_cast<float4 *>(a_lo + off) = l;
					}
					asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // tile writes -> visible to the tensor core
					mbar_arrive(bar_empty + slot * 8);                           // this thread is done with the staging slot
					workers_sync();
					if (tid == 0) {
						asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
						for (uint32_t kb = 0; kb < 2; kb++) {
							const uint32_t b_addr = smem_u32(b_op + (st * 2 + kb) * TC_B_KB_BYTES);
#pragma unroll
							for (uint32_t ks = 0; ks < 4; ks++) { // K = 8 TF32 values = 32 bytes per MMA
								const uint64_t bd = umma_desc(b_addr + ks * 32);
								const uint32_t d_cols = tmem_d + (st * 2 + kb) * 2; // this K block's accumulator pair
								umma_tf32(d_cols, umma_desc(smem_u32(a_hi + kb * TC_KB_BYTES) + ks * 32), bd, ks != 0);
								umma_tf32(d_cols, umma_desc(smem_u32(a_lo + kb * TC_KB_BYTES) + ks * 32), bd, 1u);
							}
						}
						umma_commit(bar_tiles);
						if (st + 1 == nst) {
							umma_commit(bar_accum);
						}
					}
				}
				// ---- epilogue of the tile: D[row][0] + D[row][1] = q . x_row ----
				mbar_wait(bar_accum, n_accum & 1);
				n_accum++;
				asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
				float dot = 0.0f;
				for (uint32_t c0 = 0; c0 < 2 * nkb; c0 += 16) { // columns behind the last pair hold zeros
					uint32_t v[16];
					asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, "
					             "%14, %15}, [%16];"
					             : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
					               "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
					             : "r"(tmem_d + ((warp * 32u) << 16) + c0));
					asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
					float part = 0.0f;
#pragma unroll
					for (int e = 0; e < 16; e++) {
						part += __uint_as_float(v[e]);
					}
					dot += part;
				}
				asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
				uint64_t key = LMD_KEY_NONE;
				if (tid < m) {
					const uint32_t sl = s_slot[tid];
					const float x2 = __ldg(p.norms + sl);
					float dd;
					if (METRIC == LMD_METRIC_L2) {
						dd = finish<LMD_METRIC_L2>(q2 + x2 - 2.0f * dot, 0.0f, 0.0f);
					} else {
						dd = finish<METRIC>(dot, x2, q2);
					}
					key = make_key(dd, sl);
				}
				s_keys[tile * TC_M + tid] = key;
				total += m;
			}
			workers_sync();
			// ---- the k smallest keys of the query (warp 0; as rerank_kernel) ----
			if (warp == 0) {
				uint64_t keys[TC_MAX_L / 32];
				const uint32_t filled = ((total + TC_M - 1) / TC_M) * TC_M;
#pragma unroll
				for (uint32_t c = 0; c < TC_MAX_L / 32; c++) {
					const uint32_t i = c * 32 + lane;
					keys[c] = i < filled ? s_keys[i] : LMD_KEY_NONE;
				}
				const uint32_t kk = total < p.k ? total : p.k;
				for (uint32_t out = 0; out < p.k; out++) {
					uint64_t best = LMD_KEY_NONE;
#pragma unroll
					for (uint32_t c = 0; c < TC_MAX_L / 32; c++) {
						best = keys[c] < best ? keys[c] : best;
					}
					uint64_t wmin = best;
#pragma unroll
					for (int o = 16; o > 0; o >>= 1) {
						const uint64_t other = __shfl_xor_sync(FULL, wmin, o);
						wmin = other < wmin ? other : wmin;
					}
#pragma unroll
					for (uint32_t c = 0; c < TC_MAX_L / 32; c++) {
						if (keys[c] == wmin) {
							keys[c] = LMD_KEY_NONE;
						}
					}
					if (lane == 0) {
						const bool ok = out < kk && wmin != LMD_KEY_NONE;
						p.out_ids[(uint64_t)q * p.k + out] = ok ? p.row_ids[key_slot(wmin)] : -1;
						if (p.out_dist) {
							p.out_dist[(uint64_t)q * p.k + out] = ok ? key_dist(wmin) : 0.0f;
						}
					}
				}
				if (lane == 0) {
					if (p.out_count) {
						p.out_count[q] = kk;
					}
					atomicAdd(&p.counters[4], (unsigned long long)total);
				}
			}
			workers_sync(); // s_keys / B are rewritten by the next query
		}
	}
	asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
	__syncthreads();
	if (warp == 0) {
		asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "n"(TC_TMEM_COLS) : "memory");
	}
}

__global__ void row_norms_kernel(const float *__restrict__ vecs, uint32_t row_floats, uint64_t n, float *__restrict__ norms) {
	const uint64_t row = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int lane = threadIdx.x & 31;
	if (row >= n) {
		return;
	}
	const float *v = vecs + row * row_floats;
	float s = 0.0f;
	for (uint32_t c = lane; c < row_floats / 4; c += 32) {
		const float4 x = ldg4(v + c * 4);
		s = fmaf(x.x, x.x, fmaf(x.y, x.y, fmaf(x.z, x.z, fmaf(x.w, x.w, s))));
	}
	s = team_sum<32>(s);
	if (lane == 0) {
		norms[row] = s;
	}
}

} // namespace

// Squared norms of all rows (the view the tensor rerank's epilogue reads); the caller holds s->mu exclusively.
lmd_status build_row_norms(lmd_store *s) {
	if (!s->d_norms) {
		LMD_CUDA_OK(cudaMalloc((void **)&s->d_norms, s->capacity * sizeof(float)));
		s->device_bytes += s->capacity * sizeof(float);
	}
	if (s->n) {
		row_norms_kernel<<<(unsigned)((s->n * 32 + 255) / 256), 256, 0, s->stream>>>(s->d_vecs, s->row_floats, s->n, s->d_norms);
		LMD_CUDA_OK(cudaGetLastError());
		LMD_CUDA_OK(cudaStreamSynchronize(s->stream));
	}
	s->norms_gen = s->views_gen;
	return LMD_OK;
}

lmd_status launch_rerank_tensor(const RerankArgs &a, cudaStream_t st) {
	const lmd_store *s = a.s;
	if (s->row_floats % TC_KS != 0 || a.L > TC_MAX_L || s->row_floats / 32 > TC_MAX_KB) {
		set_error("LMD_SEARCH_TENSOR_RERANK needs rows padded to a multiple of %u dimensions (this store: %u) and L <= %u",
		          TC_KS, s->row_floats, TC_MAX_L);
		return LMD_ERR_UNSUPPORTED;
	}
	if (a.strict || s->tune.strict_fp) {
		set_error("LMD_SEARCH_TENSOR_RERANK and LMD_SEARCH_STRICT_FP exclude each other");
		return LMD_ERR_INVALID_ARG;
	}
	if (!s->d_norms || s->norms_gen != s->views_gen) {
		set_error("tensor rerank: the row-norm view is not built");
		return LMD_ERR_STATE;
	}
	TcParams p;
	p.vecs = s->d_vecs;
	p.norms = s->d_norms;
	p.row_ids = s->d_row_ids;
	p.queries = a.d_queries;
	p.cand = a.d_cand;
	p.out_ids = a.d_out_ids;
	p.out_dist = a.d_out_dist;
	p.out_count = a.d_out_count;
	p.counters = a.d_counters ? a.d_counters : s->d_counters;
	p.nq = a.nq;
	p.dims = s->cfg.dims;
	p.row_floats = s->row_floats;
	p.q_stride = s->cfg.dims;
	p.L = a.L;
	p.k = a.k;
	p.metric = s->cfg.metric;
	const size_t smem = 2 * (size_t)TC_A_BYTES + (size_t)(s->row_floats / 32) * TC_B_KB_BYTES + 2 * (size_t)TC_STAGE_BYTES +
	                    TC_MAX_L * 8 + TC_M * 4 + 256;
	if (smem > (size_t)227 * 1024) {
		set_error("tensor rerank: %zu bytes of shared memory needed for %u-d rows", smem, s->row_floats);
		return LMD_ERR_UNSUPPORTED;
	}
	const unsigned grid = (unsigned)std::min<uint32_t>(a.nq, (uint32_t)s->sm_count);
	void (*kernel)(const TcParams) = s->cfg.metric == LMD_METRIC_L2       ? rerank_tc_kernel<LMD_METRIC_L2>
	                                 : s->cfg.metric == LMD_METRIC_COSINE ? rerank_tc_kernel<LMD_METRIC_COSINE>
	                                                                      : rerank_tc_kernel<LMD_METRIC_IP>;
	LMD_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
	kernel<<<grid, TC_THREADS, smem, st>>>(p);
	LMD_CUDA_OK(cudaGetLastError());
	return LMD_OK;
}

} // namespace lmd
