// tests/test_plan.py: C entry points over csrc/lmd_plan.hpp (plain host code), compiled with g++ at test time.
#include "lmd_plan.hpp"

extern "C" {
int plan_lpl(uint32_t L) {
	return lmd::lpl_for(L);
}
int plan_lpl_coarse(int lpl) {
	return lmd::lpl_coarse(lpl);
}
int plan_family(uint32_t nq, uint32_t sm_count, int tune) {
	return lmd::plan_family(nq, sm_count, tune);
}
uint32_t plan_min_warps(int family, int exact, int specialised, int filtered, int lpl) {
	return lmd::plan_min_warps(family, exact != 0, specialised != 0, filtered != 0, lpl);
}
uint32_t plan_min_warps_l(int family, int exact, int specialised, int filtered, int lpl, uint32_t L) {
	return lmd::plan_min_warps(family, exact != 0, specialised != 0, filtered != 0, lpl, L);
}
int plan_filter_mode(int tune_mode, int reads_sidecar, int filtered) {
	return lmd::plan_filter_mode(tune_mode, reads_sidecar != 0, filtered != 0);
}
uint32_t plan_tie_slots(uint32_t L) {
	return lmd::plan_tie_slots(L);
}
int plan_96d_prefers_g4(uint32_t row_floats, int exact, int filtered, int family, int tune, int lpl, uint32_t filter_bytes) {
	return lmd::plan_96d_prefers_g4(row_floats, exact != 0, filtered != 0, family, tune, lpl, filter_bytes) ? 1 : 0;
}
// out[0..2] = mode, log2(entries), bytes
void plan_filter(uint64_t nodes, uint32_t r, uint32_t words, uint32_t L, uint64_t fixed_bytes, int ternary,
                 uint32_t min_warps, int tune_mode, int tune_log2, uint32_t *out) {
	const lmd::FilterPlan f = lmd::plan_filter(nodes, r, words, L, (size_t)fixed_bytes, ternary != 0, min_warps, tune_mode, tune_log2);
	out[0] = f.mode;
	out[1] = f.lg;
	out[2] = f.bytes;
}
// out[0..2] = mode, entries, bytes of the modulo-mapped filter (entries 0 = the rule does not apply)
void plan_filter_modulo(uint64_t nodes, uint64_t fixed_bytes, uint32_t warps, uint32_t smem_kb, uint32_t r, uint32_t L,
                        double max_load, uint32_t *out) {
	const lmd::FilterPlan f = lmd::plan_filter_modulo(nodes, (size_t)fixed_bytes, warps, smem_kb, r, L, max_load);
	out[0] = f.mode;
	out[1] = f.entries;
	out[2] = f.bytes;
}
int modulo_filter_fits(uint64_t nodes, uint32_t entries) {
	return lmd::modulo_filter_fits(nodes, entries) ? 1 : 0;
}
}
