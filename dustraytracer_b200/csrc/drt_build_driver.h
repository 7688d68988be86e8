// drt_build_driver.h -- host-side level loop of the GPU BVH builder (see drt_build.cuh).
// The loop is written against a small executor interface so that the CUDA library (kernel launches,
// drt_api.cu) and the CPU logic-test harness (serial loops, tests/hostemu) drive the SAME passes.
#pragma once
#include "drt_build.cuh"

namespace drt {

enum BuildPass {
	PASS_PRIM_BOUNDS = 0, // n = primitives
	PASS_INIT_BINS,       // n = bin words of the level
	PASS_BIN,             // n = primitives
	PASS_SPLIT,           // n = nodes of the level
	PASS_PARTITION,       // n = primitives
	PASS_COLLAPSE         // n = wide nodes of the wide level; a = wide level start, b = end
};

struct BuildExec {
	virtual ~BuildExec() {}
	// Runs `n` threads of `pass`. idx_final/a/b only matter for PASS_COLLAPSE / PASS_INIT_BINS (a = words).
	virtual int launch(BuildPass pass, const BuildArrays& A, const int* idx_final, int a, int b, int n) = 0;
	virtual int read_state(const BuildArrays& A, BuildState& out) = 0; // waits for the passes issued so far
	virtual int write_state(const BuildArrays& A, const BuildState& in) = 0;
	virtual int write_root(const BuildArrays& A, const BuildNode& root, const BuildSplit& split) = 0;
};

struct BuildResult {
	int bvh2_nodes = 0, wide_nodes = 0, levels = 0, wide_levels = 0, tris_emitted = 0;
	const int* idx_final = nullptr;
};

inline size_t build_bin_slots_capacity(int n) { return (size_t)n / (DRT_SMALL_NODE + 1) + 2; }

// All arrays of A must be allocated by the caller: nodes/splits 2n, wide n (at least 1), bins
// build_bin_slots_capacity(n) * 3*DRT_BINS*DRT_BIN_WORDS ints, idx/pnode n each, out_tris/out_shade 3n.
inline int run_build(BuildExec& ex, BuildArrays A, BuildResult& res)
{
	const int n = A.n;
	if (n <= 0) { res = BuildResult(); return 0; }
	int rc;
	BuildNode root;
	init_node(&root, 0, n);
	BuildSplit rs;
	rs.axis = -1; rs.split_bin = 0; rs.n_left = 0; rs.bin_slot = n > DRT_SMALL_NODE ? 0 : -1;
	if ((rc = ex.write_root(A, root, rs))) return rc;
	BuildState st;
	st.node_count = 1; st.wide_count = 1; st.tri_cursor = 0; st.bin_slots = n > DRT_SMALL_NODE ? 1 : 0;
	if ((rc = ex.write_state(A, st))) return rc;
	if ((rc = ex.launch(PASS_PRIM_BOUNDS, A, nullptr, 0, 0, n))) return rc;

	int level_start = 0, level_end = 1, level = 0;
	int slot_base = 0, slot_end = st.bin_slots;
	while (level_end > level_start) {
		A.level_start = level_start; A.level_end = level_end; A.level = level; A.bin_slot_base = slot_base;
		const int bin_words = (slot_end - slot_base) * 3 * DRT_BINS * DRT_BIN_WORDS;
		if (bin_words > 0) {
			if ((rc = ex.launch(PASS_INIT_BINS, A, nullptr, bin_words, 0, bin_words))) return rc;
			if ((rc = ex.launch(PASS_BIN, A, nullptr, 0, 0, n))) return rc;
		}
		if ((rc = ex.launch(PASS_SPLIT, A, nullptr, 0, 0, level_end - level_start))) return rc;
		if ((rc = ex.launch(PASS_PARTITION, A, nullptr, 0, 0, n))) return rc;
		if ((rc = ex.read_state(A, st))) return rc;
		// ping-pong
		int* t = A.idx_in; A.idx_in = A.idx_out; A.idx_out = t;
		t = A.pnode_in; A.pnode_in = A.pnode_out; A.pnode_out = t;
		level_start = level_end; level_end = st.node_count;
		slot_base = slot_end; slot_end = st.bin_slots;
		level++;
	}
	res.bvh2_nodes = st.node_count;
	res.levels = level;
	res.idx_final = A.idx_in;

	// BVH2 -> wide nodes, breadth first. wide_src[0] = BVH2 root (written by write_root).
	int ws = 0, we = 1, wl = 0;
	while (we > ws) {
		if ((rc = ex.launch(PASS_COLLAPSE, A, res.idx_final, ws, we, we - ws))) return rc;
		if ((rc = ex.read_state(A, st))) return rc;
		ws = we; we = st.wide_count; wl++;
	}
	res.wide_nodes = st.wide_count;
	res.wide_levels = wl;
	res.tris_emitted = st.tri_cursor;
	return 0;
}

} // namespace drt
