// Launch planning of the search kernels as plain host code (no CUDA types): list-size classes, visited-filter sizing,
// and which kernel family serves a batch. lmd_search_kernels.cu calls these; tests/test_plan.py compiles this header
// with g++ and pins the plans the measurements in DESIGN.md section 6 were taken with.
#pragma once
#include <cstddef>
#include <cstdint>

namespace lmd {

inline int lpl_for(uint32_t L) {
	return L <= 32 ? 1 : L <= 64 ? 2 : L <= 96 ? 3 : L <= 128 ? 4 : L <= 192 ? 6 : L <= 256 ? 8 : L <= 384 ? 12 : L <= 512 ? 16 : 0;
}

// The small-batch, mid-size-batch and filtered kernels exist for lists of 64 / 128 / 256 / 512 entries only.
inline int lpl_coarse(int lpl) {
	return lpl <= 2 ? 2 : lpl <= 4 ? 4 : lpl <= 8 ? 8 : 16;
}

// Visited-filter sizing. `lg` = log2(entries). Mode 1 (two-choice buckets of 8 x 16-bit entries, 2 bytes per entry) is
// used whenever the store is small enough for its exact (bucket, tag) identification, i.e. nodes < 2^(lg - 3 + 15);
// mode 0 (u32 open addressing, 4 bytes per entry) otherwise. The table is sized for the nodes a search is expected to
// discover (about 0.8 R new neighbours per expansion, ~1.05 L expansions) at a load of 0.7 (mode 1) / 0.45 (mode 0), then
// shrunk until at least `min_warps` warps fit on an SM: the gathers need warps in flight more than a perfect filter.
struct FilterPlan {
	uint32_t mode, lg, bytes;
	uint32_t entries = 0; // mode 3 only: entries of the modulo-mapped table (lg is not used by the kernel then)
};

// Mode 3 (lmd_kernels.cuh modulo_test_and_set) identifies a node exactly by (id mod S, id div S) while the quotient fits
// the 16-bit tag and the multiply-shift division is exact: nodes < min(65535 S, 2^40 / S), S >= 512 (32-bit magic).
inline bool modulo_filter_fits(uint64_t nodes, uint32_t entries) {
	return entries >= 512 && entries % 8 == 0 && nodes < 65535ull * entries && nodes < (1ull << 40) / entries;
}
// `nodes`, `r`, `words`: the store's node count, graph degree and ternary plane words; `ternary`: the route;
// `tune_mode` / `tune_log2`: the "filter_mode" / "filter_log2" tuning keys (negative / 0 = not set).
inline FilterPlan plan_filter(uint64_t nodes, uint32_t r, uint32_t words, uint32_t L, size_t fixed_bytes, bool ternary,
                              uint32_t min_warps, int tune_mode, int tune_log2, int tune_entries = 0) {
	FilterPlan f;
	const double expect = 0.8 * (double)r * 1.05 * (double)L;
	int mode = tune_mode;
	if (mode == 3) {
		// "filter_entries" key: direct-mapped table of any size; a store too large for it takes the power-of-two table
		const uint32_t e = tune_entries >= 512 ? (uint32_t)tune_entries / 8 * 8 : 3504;
		if (modulo_filter_fits(nodes, e)) {
			f.mode = 3;
			f.lg = 12;
			f.entries = e;
			f.bytes = 2 * e;
			return f;
		}
		mode = 2;
	}
	if (mode != 0 && mode != 1 && mode != 2) {
		mode = 1;
	}
	if (mode == 2) {
		// direct-mapped 16-bit entries (lmd_kernels.cuh direct_test_and_set): 2 bytes per entry, exact below 2^(lg + 15)
		// nodes; sized like mode 1 at twice the slack (a collision forgets at once), shrunk to the warp budget
		uint32_t lg2 = 10;
		if (tune_log2 >= 6 && tune_log2 <= 16) {
			lg2 = (uint32_t)tune_log2;
		} else {
			while (lg2 < 15 && (double)(1u << lg2) * 0.35 < expect) {
				lg2++;
			}
			const size_t budget2 = ((size_t)227 * 1024 - (size_t)min_warps / 4 * 1024) / min_warps;
			while (lg2 > 9 && fixed_bytes + ((size_t)2 << lg2) > budget2) {
				lg2--;
			}
		}
		while (lg2 < 16 && nodes >= (1ull << (lg2 + 15))) {
			lg2++;
		}
		f.mode = 2;
		f.lg = lg2;
		f.bytes = 2u << lg2;
		return f;
	}
	auto fits_exact = [&](uint32_t lg) { return lg >= 4 && nodes < (1ull << (lg - 3 + 15)); };
	uint32_t lg;
	if (tune_log2 >= 6 && tune_log2 <= 16) {
		lg = (uint32_t)tune_log2;
	} else {
		const double load = mode == 1 ? 0.7 : 0.45;
		lg = 9;
		while (lg < 15 && (double)(1u << lg) * load < expect) {
			lg++;
		}
		// 1 KB of every CTA's shared memory belongs to the system: CTAs of 4 warps, min_warps / 4 of them
		const size_t budget = ((size_t)227 * 1024 - (size_t)min_warps / 4 * 1024) / min_warps;
		while (lg > 9 && fixed_bytes + ((size_t)(mode == 1 ? 2 : 4) << lg) > budget) {
			lg--;
		}
		// Ternary routing over planes of 3+ words (D > 128): lane j reads neighbour j's planes with 8-byte loads that sit
		// 8 * words bytes apart across the warp, so every 32-byte sector is fetched by four load instructions and the
		// kernel lives on L1 hits - and L1 is whatever the CTAs' shared memory leaves of the SM's 256 KB. A 2 KB filter
		// (4 KB per warp, 130 KB per SM) leaves L1 ~120 KB; the re-evaluations it causes cost 192 B each. Measured on
		// 1 M x 768 cosine, L = 80: 2.59 ms per 10 k queries with the 8 KB filter, 2.78 with 4 KB, 1.78 with 2 KB
		// (profiles/r23_tune_ternary.log). A warp-wide coalesced read of the plane region (every byte requested once)
		// was tried and took the L1 dependence away, but at 2.1 ms it lost to this (profiles/r24_tune_ternary.log).
		if (ternary && words > 2 && lg > 10) {
			lg = 10;
		}
	}
	if (mode == 1 && !fits_exact(lg)) {
		// Stores beyond 2^(lg + 12) nodes (8.4 M at lg = 11) need more buckets for the exact (bucket, tag) identification:
		// grow the table while the shared-memory budget allows it, and only then fall back to the 4-byte table - sized
		// for ITS load factor, not for the 2-byte one (measured on 10 M x 96, L = 48: the undersized fallback cost 25 %).
		const size_t budget = (size_t)227 * 1024 / 12;
		uint32_t lg1 = lg;
		while (lg1 < 15 && !fits_exact(lg1)) {
			lg1++;
		}
		if (!(tune_log2 >= 6) && fits_exact(lg1) && fixed_bytes + ((size_t)2 << lg1) <= budget) {
			lg = lg1;
		} else {
			mode = 0;
			if (!(tune_log2 >= 6)) {
				lg = 9;
				while (lg < 15 && (double)(1u << lg) * 0.45 < expect) {
					lg++;
				}
				while (lg > 9 && fixed_bytes + ((size_t)4 << lg) > (size_t)227 * 1024 / 16) {
					lg--;
				}
			}
		}
	}
	f.mode = (uint32_t)mode;
	f.lg = lg;
	f.bytes = (uint32_t)((mode == 1 ? 2u : 4u) << lg);
	return f;
}

// Which visited filter a launch gets when the "filter_mode" key is not set: the kernels that read a routing sidecar (ternary,
// int8 route, exact route behind the int8 pre-filter) are bound by instruction issue and re-meet forgotten nodes cheaply,
// so they take the direct-mapped filter (mode 2: a quarter of the instructions per test); the plain exact kernels are
// bound by the bytes every re-evaluated row costs and keep the two-choice buckets (mode 1). Measured on 1 M rows, 10 k
// queries (profiles/r2_19_*_fm.json): int8 route L = 56 0.66 -> 0.57 ms, ternary 768-d L = 80 1.42 -> 1.25 ms,
// pre-filtered exact L = 56 1.03 -> 0.99 ms; same ids and distances (tests: results do not depend on the filter).
inline int plan_filter_mode(int tune_mode, bool reads_sidecar, bool filtered) {
	if (tune_mode >= 0 && tune_mode <= 3) {
		return tune_mode;
	}
	return reads_sidecar && !filtered ? 2 : 1;
}

// The 128-d exact kernels (72 registers: 7 CTAs of 4 warps by the register file) held 6 CTAs per SM because the 8 KB
// power-of-two filter is what shared memory cut off; the modulo-mapped filter takes what `warps` resident warps leave
// of `smem_kb`, and no more than 3 072 entries (6 KB). smem_kb = 227 is all an SM can give. The plain kernels pass 196:
// the shared-memory carve-out comes in steps (... 164, 196, 228 KB), whatever it leaves of the SM's 256 KB is L1, and
// they measurably want the 60 KB a 196 KB carve-out leaves rather than 28 (3 000 entries beat 3 072 and 3 568 at every
// list size). 1 M x 128, 10 k queries, ms per batch alone / with two batches in flight
// (profiles/r2_31_32_modulo_filter.json, r2_34_35_modulo_plain.json):
//   pre-filtered exact L = 56   0.99 / 0.893 -> 1.00 / 0.818      96-d L = 64   1.05 / 0.965 -> 1.04 / 0.851
//   plain exact  L = 32  0.786 / 0.626 -> 0.746 / 0.588    L = 56  1.235 / 1.143 -> 1.185 / 1.074
//                L = 64  1.343 / 1.275 -> 1.300 / 1.158    L = 100 (128-entry list) no gain: not applied
// `max_load` > 0 (plain kernels: a forgotten node costs a 512-byte row) keeps the rule to searches whose expected
// discoveries (0.8 R * 1.05 L) stay below that fraction of the table. Returns entries = 0 when the rule does not apply
// (long lists; stores beyond the exact range of the 16-bit tag) and the caller keeps plan_filter().
inline FilterPlan plan_filter_modulo(uint64_t nodes, size_t fixed_bytes, uint32_t warps, uint32_t smem_kb = 227,
                                     uint32_t r = 0, uint32_t L = 0, double max_load = 0.0) {
	FilterPlan f;
	f.mode = 0;
	f.lg = 0;
	f.bytes = 0;
	const size_t budget = ((size_t)smem_kb * 1024 - (size_t)warps / 4 * 1024) / warps / 16 * 16;
	if (budget <= fixed_bytes + 4096) {
		return f;
	}
	uint32_t e = (uint32_t)((budget - fixed_bytes) / 2 / 8 * 8);
	e = e > 3072 ? 3072 : e;
	if (!modulo_filter_fits(nodes, e)) {
		return f;
	}
	if (max_load > 0.0 && 0.8 * (double)r * 1.05 * (double)L > max_load * (double)e) {
		return f;
	}
	f.mode = 3;
	f.lg = 12;
	f.entries = e;
	f.bytes = 2 * e;
	return f;
}

// Which traversal family serves an exact, unfiltered search over 128-d / 96-d rows: 0 = throughput kernels (2 rows in
// flight per team), 1 = small-batch kernels (G = 8: all gathers of an expansion at once; 168 registers, 12 warps per SM),
// 2 = mid-size-batch kernels (G = 4; 128 registers, 16 warps per SM). `tune` is the "latency_mode" key (-1 = by batch).
inline int plan_family(uint32_t nq, uint32_t sm_count, int tune) {
	if (tune == 1 || (tune < 0 && nq <= sm_count * 12)) {
		return 1;
	}
	if (tune == 2 || (tune < 0 && nq <= sm_count * 16)) {
		return 2;
	}
	return 0;
}

// Warps the chosen kernels want resident on an SM; the visited filter is shrunk until that many fit (DESIGN.md 6,
// "Long lists"). `specialised` = 128-d or 96-d rows; `lpl` = list entries / 32 after rounding to the family's sizes.
//
// The routes that read a sidecar block (ternary planes, int8 codes) are bound by instruction issue, not by bandwidth
// (ncu: issue slots 48 % busy, no eligible warp 50 % of the cycles at 22-27 resident warps, profiles/r2_04_*): they want
// 32 warps on an SM more than a filter that never forgets. Measured on 1 M rows, 10 k queries (profiles/r2_05_*_occ.json):
// int8 route L = 56 0.80 -> 0.65 ms, ternary 768-d L = 80 1.91 -> 1.48 ms with the 4 KB filter and the short tie region
// below. Longer lists keep the old rule (their filters would forget too much at 4 KB).
inline uint32_t plan_min_warps(int family, bool exact, bool specialised, bool filtered, int lpl, uint32_t L = 0) {
	if (family == 1) {
		return 12;
	}
	if (family == 2) {
		return 16;
	}
	if (!exact && !filtered && L > 0 && L <= 96) {
		return 32;
	}
	return exact && specialised && (!filtered || lpl <= 4) ? 24u : 16u;
}

// Room behind the top L for distance ties on the integer-valued routes (ternary scores, int8 code distances tie all the
// time; see list_insert): at least 8 entries, then whatever the list class (a multiple of 32 entries) leaves - a region
// of 64 put L = 56 into the 128-entry class and doubled the cost of every list operation for 0 overflows fewer.
inline uint32_t plan_tie_slots(uint32_t L) {
	const int lpl = lpl_for(L + 8);
	return lpl ? (uint32_t)lpl * 32u - L : 0u;
}

// 96-d rows (384 bytes): the G = 4 kernels also win on full batches when shared memory holds the G = 2 kernel at 24 warps
// or fewer anyway, i.e. with the 8 KB filter; they exist for lists of 64 and 128 entries.
inline bool plan_96d_prefers_g4(uint32_t row_floats, bool exact, bool filtered, int family, int tune, int lpl,
                                uint32_t filter_bytes) {
	return row_floats == 96 && exact && !filtered && family == 0 && tune < 0 && (lpl == 2 || lpl == 4) &&
	       filter_bytes >= 8192;
}

} // namespace lmd
