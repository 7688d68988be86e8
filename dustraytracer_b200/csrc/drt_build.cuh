// drt_build.cuh -- GPU binned-SAH BVH builder, per-thread bodies (level-synchronous, atomics only).
//
// Replaces the reference's CPU builder (DustRayTracer/src/Core/BVH/BVHBuilder.cu:100-346: recursive
// top-down, 7 candidate planes x 3 axes over the vertex AABB, int-truncated cost, 20-triangle leaves)
// with a breadth-first build that runs entirely on the device:
//   level loop:  bin_prim (1 thread / primitive, large nodes)  -> global per-node bins via atomics
//                split_node (1 thread / node of the level)     -> SAH plane, child allocation;
//                                                                 small nodes are split in place here
//                partition_prim (1 thread / primitive)         -> scatter into the children's ranges
//   then:        collapse_node (1 thread / wide node, per wide level): BVH2 -> 8-wide compressed nodes
//                (drt_types.h WideNode) + triangles written in leaf order.
// Every body is a plain function of its thread index using only atomics for communication, so the
// same code runs under the CPU test harness (tests/hostemu) as a serial loop.
#pragma once
#include "drt_types.h"

namespace drt {

#ifndef DRT_BINS
#define DRT_BINS 16          // SAH bins per axis for large nodes
#endif
#define DRT_SMALL_NODE 32    // nodes with <= this many primitives are split by their own thread
#ifndef DRT_SMALL_BINS
#define DRT_SMALL_BINS 16    // (<= DRT_BINS) bins of a small node's own sweep; 8 -> 16 in round 2: +1 ... +2.7 % on the scenes of > 6000 triangles (profiles/r02c_tree_quality_ab.txt)
#endif
#ifndef DRT_LEAF_MAX
#define DRT_LEAF_MAX 3       // triangles per leaf slot (unary-coded in 3 bits of WideNode::meta: at most 3)
#endif
#define DRT_MEDIAN_LEVEL 40  // deeper than this: forced median splits (bounds the tree depth)
#define DRT_BIN_WORDS 7      // count, lo xyz, hi xyz

// ---- atomics (device: hardware; host emulation: the harness is serial) ----
DRT_HD int atomic_add_i(int* p, int v)
{
#if defined(__CUDA_ARCH__)
	return atomicAdd(p, v);
#else
	int o = *p; *p = o + v; return o;
#endif
}
DRT_HD void atomic_min_i(int* p, int v)
{
#if defined(__CUDA_ARCH__)
	atomicMin(p, v);
#else
	if (v < *p) *p = v;
#endif
}
DRT_HD void atomic_max_i(int* p, int v)
{
#if defined(__CUDA_ARCH__)
	atomicMax(p, v);
#else
	if (v > *p) *p = v;
#endif
}

// order-preserving float <-> int map so integer atomicMin/Max order floats correctly
DRT_HD int enc_f(float f) { int i = (int)as_uint(f); return i >= 0 ? i : i ^ 0x7fffffff; }
DRT_HD float dec_f(int i) { return as_float((uint32_t)(i >= 0 ? i : i ^ 0x7fffffff)); }
#define DRT_ENC_POS_INF 0x7f800000
#define DRT_ENC_NEG_INF ((int)0x807fffff) /* enc_f(-inf) = 0xff800000 ^ 0x7fffffff */

struct BuildNode {       // 64 bytes; bounds are enc_f()-encoded while the build runs
	int lo[3]; int start;
	int hi[3]; int count;
	int clo[3]; int left;    // child pair index (right = left + 1); -1: leaf
	int chi[3]; int fill;    // scatter cursor used by partition_prim
	// split decision of this node (valid when left >= 0 and count > DRT_SMALL_NODE)
};
struct BuildSplit {      // parallel array to BuildNode
	int axis;            // 0..2: SAH plane on that axis; -1: median split by position
	int split_bin;       // primitives with bin < split_bin go left
	int n_left;
	int bin_slot;        // set by the PARENT's split: slot of this node's bins if it is a large node
};

struct BuildState {
	int node_count;      // BVH2 nodes allocated
	int wide_count;      // wide nodes allocated
	int tri_cursor;      // triangles emitted in leaf order
	int bin_slots;       // cumulative count of bin slots handed to large nodes (see BuildSplit::bin_slot)
};

struct BuildArrays {
	// inputs (load order)
	const float* pos;        // 9 floats / triangle
	const float* uv;         // 6 floats / triangle
	const float* fnormal;    // 3 floats / triangle
	const int32_t* mat;      // material index (bit 31: alpha-tested), already validated
	int n;
	// per-primitive bounds
	drt_f4* tri_lo; drt_f4* tri_hi;
	// ping-pong index arrays
	int* idx_in; int* idx_out; int* pnode_in; int* pnode_out;
	BuildNode* nodes; BuildSplit* splits;
	int* bins;               // [bin_slot - bin_slot_base][axis][bin][DRT_BIN_WORDS]
	int bin_slot_base;       // first bin slot of the current level
	BuildState* state;
	int level_start, level_end, level;
	// outputs
	int* wide_src;           // BVH2 node of each wide node
	WideNode* wide;
	drt_f4* out_tris; drt_f4* out_shade;
};

// (separately rounded on the device too: the serial sweeps below, the warp-cooperative split of the single-launch builder and the
// host emulation must take the same decisions, whatever multiply-adds the compiler would contract in each code shape)
DRT_HD float box_area(float lx, float ly, float lz, float hx, float hy, float hz)
{
	float dx = hx - lx, dy = hy - ly, dz = hz - lz;
	if (!(dx >= 0.f) || !(dy >= 0.f) || !(dz >= 0.f)) return 0.f; // empty box
#if defined(__CUDA_ARCH__)
	return __fmul_rn(2.f, __fadd_rn(__fadd_rn(__fmul_rn(dx, dy), __fmul_rn(dy, dz)), __fmul_rn(dz, dx)));
#else
	return 2.f * (dx * dy + dy * dz + dz * dx);
#endif
}

// SAH cost of a candidate plane: SA(L) n_L + SA(R) n_R
DRT_HD float sah_cost(float area_l, int n_l, float area_r, int n_r)
{
#if defined(__CUDA_ARCH__)
	return __fadd_rn(__fmul_rn(area_l, (float)n_l), __fmul_rn(area_r, (float)n_r));
#else
	return area_l * (float)n_l + area_r * (float)n_r;
#endif
}

DRT_HD int bin_of(float c, float cmin, float cmax, int nbins)
{
	float ext = cmax - cmin;
	if (!(ext > 0.f)) return 0;
	int b = (int)((c - cmin) * ((float)nbins / ext));
	return b < 0 ? 0 : (b >= nbins ? nbins - 1 : b);
}

// ---- pass 0: primitive bounds + root node ----
DRT_HD void prim_bounds(const BuildArrays& A, int i)
{
	if (i >= A.n) return;
	const float* p = A.pos + 9 * (size_t)i;
	float lx = fmin3(p[0], p[3], p[6]), ly = fmin3(p[1], p[4], p[7]), lz = fmin3(p[2], p[5], p[8]);
	float hx = fmax3(p[0], p[3], p[6]), hy = fmax3(p[1], p[4], p[7]), hz = fmax3(p[2], p[5], p[8]);
	drt_f4 lo, hi;
	lo.x = lx; lo.y = ly; lo.z = lz; lo.w = 0.f;
	hi.x = hx; hi.y = hy; hi.z = hz; hi.w = 0.f;
	A.tri_lo[i] = lo; A.tri_hi[i] = hi;
	A.idx_in[i] = i;
	A.pnode_in[i] = 0;
	BuildNode* r = A.nodes;
	atomic_min_i(&r->lo[0], enc_f(lx)); atomic_min_i(&r->lo[1], enc_f(ly)); atomic_min_i(&r->lo[2], enc_f(lz));
	atomic_max_i(&r->hi[0], enc_f(hx)); atomic_max_i(&r->hi[1], enc_f(hy)); atomic_max_i(&r->hi[2], enc_f(hz));
	float cx = 0.5f * (lx + hx), cy = 0.5f * (ly + hy), cz = 0.5f * (lz + hz);
	atomic_min_i(&r->clo[0], enc_f(cx)); atomic_min_i(&r->clo[1], enc_f(cy)); atomic_min_i(&r->clo[2], enc_f(cz));
	atomic_max_i(&r->chi[0], enc_f(cx)); atomic_max_i(&r->chi[1], enc_f(cy)); atomic_max_i(&r->chi[2], enc_f(cz));
}

DRT_HD void init_node(BuildNode* nd, int start, int count)
{
	for (int a = 0; a < 3; a++) {
		nd->lo[a] = DRT_ENC_POS_INF; nd->hi[a] = DRT_ENC_NEG_INF;
		nd->clo[a] = DRT_ENC_POS_INF; nd->chi[a] = DRT_ENC_NEG_INF;
	}
	nd->start = start; nd->count = count; nd->left = -1; nd->fill = 0;
}

// ---- pass 1: bin the primitives of the large nodes of this level ----
DRT_HD void bin_prim(const BuildArrays& A, int i)
{
	if (i >= A.n) return;
	int node = A.pnode_in[i];
	if (node < A.level_start || node >= A.level_end) return;
	const BuildNode& nd = A.nodes[node];
	if (nd.count <= DRT_SMALL_NODE) return;
	int prim = A.idx_in[i];
	drt_f4 lo = A.tri_lo[prim], hi = A.tri_hi[prim];
	float c[3] = { 0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z) };
	int elo[3] = { enc_f(lo.x), enc_f(lo.y), enc_f(lo.z) }, ehi[3] = { enc_f(hi.x), enc_f(hi.y), enc_f(hi.z) };
	int* nb = A.bins + (size_t)(A.splits[node].bin_slot - A.bin_slot_base) * 3 * DRT_BINS * DRT_BIN_WORDS;
	for (int a = 0; a < 3; a++) {
		int b = bin_of(c[a], dec_f(nd.clo[a]), dec_f(nd.chi[a]), DRT_BINS);
		int* w = nb + (a * DRT_BINS + b) * DRT_BIN_WORDS;
		atomic_add_i(w, 1);
		atomic_min_i(w + 1, elo[0]); atomic_min_i(w + 2, elo[1]); atomic_min_i(w + 3, elo[2]);
		atomic_max_i(w + 4, ehi[0]); atomic_max_i(w + 5, ehi[1]); atomic_max_i(w + 6, ehi[2]);
	}
}

DRT_HD void init_bin_word(const BuildArrays& A, int j, int total_words)
{
	if (j >= total_words) return;
	int k = j % DRT_BIN_WORDS;
	A.bins[j] = k == 0 ? 0 : (k <= 3 ? DRT_ENC_POS_INF : DRT_ENC_NEG_INF);
}

struct BinAcc {
	int n;
	float lo[3], hi[3];
};
DRT_HD void acc_reset(BinAcc& b)
{
	b.n = 0;
	for (int a = 0; a < 3; a++) { b.lo[a] = INFINITY; b.hi[a] = -INFINITY; }
}
DRT_HD void acc_merge(BinAcc& d, const BinAcc& s)
{
	d.n += s.n;
	for (int a = 0; a < 3; a++) { d.lo[a] = fminf(d.lo[a], s.lo[a]); d.hi[a] = fmaxf(d.hi[a], s.hi[a]); }
}
DRT_HD float acc_area(const BinAcc& b) { return b.n ? box_area(b.lo[0], b.lo[1], b.lo[2], b.hi[0], b.hi[1], b.hi[2]) : 0.f; }

// Best plane over `nbins` bins of one axis: minimises SA(L)*nL + SA(R)*nR. Returns false if every
// plane leaves one side empty.
DRT_HD bool best_plane(const BinAcc* bins, int nbins, float& best_cost, int& best_split, BinAcc& bestL, BinAcc& bestR)
{
	bool found = false;
	BinAcc right_acc[DRT_BINS];
	BinAcc r;
	acc_reset(r);
	for (int b = nbins - 1; b >= 1; b--) { acc_merge(r, bins[b]); right_acc[b] = r; }
	BinAcc l;
	acc_reset(l);
	for (int s = 1; s < nbins; s++) {
		acc_merge(l, bins[s - 1]);
		const BinAcc& rr = right_acc[s];
		if (l.n == 0 || rr.n == 0) continue;
		float cost = sah_cost(acc_area(l), l.n, acc_area(rr), rr.n);
		if (!found || cost < best_cost) { found = true; best_cost = cost; best_split = s; bestL = l; bestR = rr; }
	}
	return found;
}

DRT_HD void set_child_box(BuildNode* c, const BinAcc& b)
{
	for (int a = 0; a < 3; a++) { c->lo[a] = enc_f(b.lo[a]); c->hi[a] = enc_f(b.hi[a]); }
}

// ---- pass 2: one thread per node of the level ----
DRT_HD void split_node(const BuildArrays& A, int j)
{
	int node = A.level_start + j;
	if (node >= A.level_end) return;
	BuildNode* nd = A.nodes + node;
	const int count = nd->count, start = nd->start;
	if (count <= DRT_LEAF_MAX) { nd->left = -1; return; }
	const bool small = count <= DRT_SMALL_NODE;
	const bool force_median = A.level >= DRT_MEDIAN_LEVEL;
	const float clo[3] = { dec_f(nd->clo[0]), dec_f(nd->clo[1]), dec_f(nd->clo[2]) };
	const float chi[3] = { dec_f(nd->chi[0]), dec_f(nd->chi[1]), dec_f(nd->chi[2]) };
	const int nbins = small ? DRT_SMALL_BINS : DRT_BINS;

	int best_axis = -1, best_split = 0;
	float best_cost = 0.f;
	BinAcc bestL, bestR;
	acc_reset(bestL); acc_reset(bestR);
	if (!force_median) {
		for (int a = 0; a < 3; a++) {
			BinAcc bins[DRT_BINS];
			for (int b = 0; b < nbins; b++) acc_reset(bins[b]);
			if (small) {
				for (int k = 0; k < count; k++) {
					int prim = A.idx_in[start + k];
					drt_f4 lo = A.tri_lo[prim], hi = A.tri_hi[prim];
					float c = a == 0 ? 0.5f * (lo.x + hi.x) : (a == 1 ? 0.5f * (lo.y + hi.y) : 0.5f * (lo.z + hi.z));
					BinAcc& bb = bins[bin_of(c, clo[a], chi[a], nbins)];
					bb.n++;
					bb.lo[0] = fminf(bb.lo[0], lo.x); bb.lo[1] = fminf(bb.lo[1], lo.y); bb.lo[2] = fminf(bb.lo[2], lo.z);
					bb.hi[0] = fmaxf(bb.hi[0], hi.x); bb.hi[1] = fmaxf(bb.hi[1], hi.y); bb.hi[2] = fmaxf(bb.hi[2], hi.z);
				}
			}
			else {
				const int* nb = A.bins + ((size_t)(A.splits[node].bin_slot - A.bin_slot_base) * 3 + a) * DRT_BINS * DRT_BIN_WORDS;
				for (int b = 0; b < nbins; b++) {
					const int* w = nb + b * DRT_BIN_WORDS;
					bins[b].n = w[0];
					for (int k = 0; k < 3; k++) { bins[b].lo[k] = dec_f(w[1 + k]); bins[b].hi[k] = dec_f(w[4 + k]); }
				}
			}
			float cost; int split; BinAcc l, r;
			if (best_plane(bins, nbins, cost, split, l, r) && (best_axis < 0 || cost < best_cost)) {
				best_axis = a; best_cost = cost; best_split = split; bestL = l; bestR = r;
			}
		}
	}
	const int n_left = best_axis >= 0 ? bestL.n : count / 2;
	const int c0 = atomic_add_i(&A.state->node_count, 2);
	BuildNode* L = A.nodes + c0;
	BuildNode* R = L + 1;
	init_node(L, start, n_left);
	init_node(R, start + n_left, count - n_left);
	if (best_axis >= 0) { set_child_box(L, bestL); set_child_box(R, bestR); }
	BuildSplit sp = A.splits[node];
	sp.axis = best_axis; sp.split_bin = best_split; sp.n_left = n_left;
	A.splits[node] = sp;
	nd->left = c0;
	for (int k = 0; k < 2; k++) {
		BuildSplit cs;
		cs.axis = -1; cs.split_bin = 0; cs.n_left = 0;
		cs.bin_slot = A.nodes[c0 + k].count > DRT_SMALL_NODE ? atomic_add_i(&A.state->bin_slots, 1) : -1;
		A.splits[c0 + k] = cs;
	}

	if (small) {
		// in-place partition of this node's own index range + exact child bounds
		int tmp[DRT_SMALL_NODE];
		int nl = 0, nr = 0;
		BinAcc boxes[2], cents[2];
		acc_reset(boxes[0]); acc_reset(boxes[1]); acc_reset(cents[0]); acc_reset(cents[1]);
		for (int k = 0; k < count; k++) {
			int prim = A.idx_in[start + k];
			drt_f4 lo = A.tri_lo[prim], hi = A.tri_hi[prim];
			float c[3] = { 0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z) };
			bool left = best_axis >= 0 ? (bin_of(c[best_axis], clo[best_axis], chi[best_axis], nbins) < best_split) : (k < n_left);
			int s = left ? 0 : 1;
			if (left) tmp[nl++] = prim; else tmp[DRT_SMALL_NODE - 1 - nr++] = prim;
			boxes[s].n++;
			boxes[s].lo[0] = fminf(boxes[s].lo[0], lo.x); boxes[s].lo[1] = fminf(boxes[s].lo[1], lo.y); boxes[s].lo[2] = fminf(boxes[s].lo[2], lo.z);
			boxes[s].hi[0] = fmaxf(boxes[s].hi[0], hi.x); boxes[s].hi[1] = fmaxf(boxes[s].hi[1], hi.y); boxes[s].hi[2] = fmaxf(boxes[s].hi[2], hi.z);
			for (int a = 0; a < 3; a++) { cents[s].lo[a] = fminf(cents[s].lo[a], c[a]); cents[s].hi[a] = fmaxf(cents[s].hi[a], c[a]); }
		}
		for (int k = 0; k < nl; k++) { A.idx_in[start + k] = tmp[k]; A.pnode_in[start + k] = c0; }
		for (int k = 0; k < nr; k++) { A.idx_in[start + nl + k] = tmp[DRT_SMALL_NODE - 1 - k]; A.pnode_in[start + nl + k] = c0 + 1; }
		set_child_box(L, boxes[0]); set_child_box(R, boxes[1]);
		for (int a = 0; a < 3; a++) {
			L->clo[a] = enc_f(cents[0].lo[a]); L->chi[a] = enc_f(cents[0].hi[a]);
			R->clo[a] = enc_f(cents[1].lo[a]); R->chi[a] = enc_f(cents[1].hi[a]);
		}
		L->fill = nl; R->fill = nr;
	}
}

// ---- pass 3: scatter the primitives of the large nodes that were split; copy the rest through ----
DRT_HD void partition_prim(const BuildArrays& A, int i)
{
	if (i >= A.n) return;
	int node = A.pnode_in[i];
	int prim = A.idx_in[i];
	bool moved = false;
	if (node >= A.level_start && node < A.level_end) {
		const BuildNode& nd = A.nodes[node];
		if (nd.left >= 0 && nd.count > DRT_SMALL_NODE) {
			const BuildSplit sp = A.splits[node];
			drt_f4 lo = A.tri_lo[prim], hi = A.tri_hi[prim];
			float c[3] = { 0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z) };
			bool left = sp.axis >= 0
				? (bin_of(c[sp.axis], dec_f(nd.clo[sp.axis]), dec_f(nd.chi[sp.axis]), DRT_BINS) < sp.split_bin)
				: (i - nd.start < sp.n_left);
			int child = nd.left + (left ? 0 : 1);
			BuildNode* cn = A.nodes + child;
			int off = atomic_add_i(&cn->fill, 1);
			int np = cn->start + off;
			A.idx_out[np] = prim;
			A.pnode_out[np] = child;
			atomic_min_i(&cn->clo[0], enc_f(c[0])); atomic_min_i(&cn->clo[1], enc_f(c[1])); atomic_min_i(&cn->clo[2], enc_f(c[2]));
			atomic_max_i(&cn->chi[0], enc_f(c[0])); atomic_max_i(&cn->chi[1], enc_f(c[1])); atomic_max_i(&cn->chi[2], enc_f(c[2]));
			if (sp.axis < 0) {
				atomic_min_i(&cn->lo[0], enc_f(lo.x)); atomic_min_i(&cn->lo[1], enc_f(lo.y)); atomic_min_i(&cn->lo[2], enc_f(lo.z));
				atomic_max_i(&cn->hi[0], enc_f(hi.x)); atomic_max_i(&cn->hi[1], enc_f(hi.y)); atomic_max_i(&cn->hi[2], enc_f(hi.z));
			}
			moved = true;
		}
	}
	if (!moved) { A.idx_out[i] = prim; A.pnode_out[i] = node; }
}

// ---- pass 4: collapse BVH2 -> compressed 8-wide nodes, one thread per wide node of the level ----
DRT_HD float node_area(const BuildNode& n)
{
	return box_area(dec_f(n.lo[0]), dec_f(n.lo[1]), dec_f(n.lo[2]), dec_f(n.hi[0]), dec_f(n.hi[1]), dec_f(n.hi[2]));
}

// smallest biased exponent byte e with 2^(e-127) * 255 >= extent
DRT_HD uint32_t quant_exponent(float extent)
{
	float s = extent / 255.f;
	uint32_t bits = as_uint(s);
	uint32_t e = (bits + 0x007fffffu) >> 23; // round the scale up to a power of two
	if (!(extent > 0.f)) e = 1u;
	return e < 1u ? 1u : (e > 254u ? 254u : e);
}

DRT_HD void write_triangle(const BuildArrays& A, int slot, int prim)
{
	const float* p = A.pos + 9 * (size_t)prim;
	const float* t = A.uv + 6 * (size_t)prim;
	const float* fn = A.fnormal + 3 * (size_t)prim;
	drt_f4 r0, r1, r2;
	// e1 = v1 - v0, e2 = v2 - v0: the same separately-rounded subtractions Intersection.cu:8-9 performs per test
	r0.x = p[0]; r0.y = p[1]; r0.z = p[2]; r0.w = as_float((uint32_t)prim);
	r1.x = p[3] - p[0]; r1.y = p[4] - p[1]; r1.z = p[5] - p[2]; r1.w = as_float((uint32_t)A.mat[prim]);
	r2.x = p[6] - p[0]; r2.y = p[7] - p[1]; r2.z = p[8] - p[2]; r2.w = 0.f;
	A.out_tris[3 * (size_t)slot] = r0; A.out_tris[3 * (size_t)slot + 1] = r1; A.out_tris[3 * (size_t)slot + 2] = r2;
	drt_f4 s0, s1, s2;
	s0.x = fn[0]; s0.y = fn[1]; s0.z = fn[2]; s0.w = t[0];
	s1.x = t[1]; s1.y = t[2]; s1.z = t[3]; s1.w = t[4];
	s2.x = t[5]; s2.y = 0.f; s2.z = 0.f; s2.w = 0.f;
	A.out_shade[3 * (size_t)slot] = s0; A.out_shade[3 * (size_t)slot + 1] = s1; A.out_shade[3 * (size_t)slot + 2] = s2;
}

// ---- collapse of one wide node, in three steps (the single-launch builder runs the middle one on a whole warp) ----
// 1. the children: open the inner child with the largest surface area until there are 8 (or only leaves); their costs per slot:
//    slot bit set (x=4,y=2,z=1) <=> child lies on the + side of that axis, so a ray travelling towards + visits slot^7 in
//    descending order = near to far (drt_trace.cuh)
DRT_HD int collapse_children(const BuildArrays& A, int w, int* child, float* cost /* [8][8] */)
{
	const int src = A.wide_src[w];
	const BuildNode& parent = A.nodes[src];
	int n = 0;
	if (parent.left < 0) child[n++] = src; // the whole scene is one leaf
	else { child[n++] = parent.left; child[n++] = parent.left + 1; }
	while (n < 8) { // open the inner child with the largest surface area
		int best = -1; float best_area = -1.f;
		for (int k = 0; k < n; k++) {
			const BuildNode& c = A.nodes[child[k]];
			if (c.left < 0) continue;
			float a = node_area(c);
			if (a > best_area) { best_area = a; best = k; }
		}
		if (best < 0) break;
		int l = A.nodes[child[best]].left;
		child[best] = l;
		child[n++] = l + 1;
	}
	const float pc[3] = { 0.5f * (dec_f(parent.lo[0]) + dec_f(parent.hi[0])), 0.5f * (dec_f(parent.lo[1]) + dec_f(parent.hi[1])), 0.5f * (dec_f(parent.lo[2]) + dec_f(parent.hi[2])) };
	for (int k = 0; k < n; k++) {
		const BuildNode& c = A.nodes[child[k]];
		float d[3];
		for (int a = 0; a < 3; a++) d[a] = 0.5f * (dec_f(c.lo[a]) + dec_f(c.hi[a])) - pc[a];
		for (int s = 0; s < 8; s++)
			cost[k * 8 + s] = ((s & 4) ? d[0] : -d[0]) + ((s & 2) ? d[1] : -d[1]) + ((s & 1) ? d[2] : -d[2]);
	}
	return n;
}

// 2. greedy slot assignment: the free (child, slot) pair of largest cost, ties to the first pair in (child, slot) order, n times
DRT_HD void collapse_match(int n, const float* cost, int* slot_of, int* child_in_slot)
{
	for (int k = 0; k < 8; k++) { slot_of[k] = -1; child_in_slot[k] = -1; }
	for (int it = 0; it < n; it++) {
		int bk = -1, bs = -1; float bc = -INFINITY;
		for (int k = 0; k < n; k++) {
			if (slot_of[k] >= 0) continue;
			for (int s = 0; s < 8; s++) {
				if (child_in_slot[s] >= 0) continue;
				if (bk < 0 || cost[k * 8 + s] > bc) { bc = cost[k * 8 + s]; bk = k; bs = s; }
			}
		}
		slot_of[bk] = bs; child_in_slot[bs] = bk;
	}
}

// quantisation grid of a wide node: per axis the smallest power-of-two step with which 255 steps from the box's low corner reach its high corner
DRT_HD void quant_grid(const float* plo, const float* phi, uint32_t* e, float* scale)
{
	for (int a = 0; a < 3; a++) {
		e[a] = quant_exponent(phi[a] - plo[a]);
		// make sure hi is reachable with q <= 255 after the outward rounding below
		for (;;) {
			scale[a] = as_float(e[a] << 23);
			if (plo[a] + 255.f * scale[a] >= phi[a] || e[a] >= 254u) break;
			e[a]++;
		}
	}
}
// a child's box on that grid, rounded outward
DRT_HD void quant_child(const BuildNode& c, const float* plo, const float* scale, uint32_t* ql, uint32_t* qh)
{
	for (int a = 0; a < 3; a++) {
		float lo = dec_f(c.lo[a]), hi = dec_f(c.hi[a]);
		float base = plo[a];
		int q0 = (int)floorf((lo - base) / scale[a]);
		int q1 = (int)ceilf((hi - base) / scale[a]);
		q0 = q0 < 0 ? 0 : (q0 > 255 ? 255 : q0);
		q1 = q1 < 0 ? 0 : (q1 > 255 ? 255 : q1);
		while (q0 > 0 && base + (float)q0 * scale[a] > lo) q0--;    // outward rounding, verified in fp32
		while (q1 < 255 && base + (float)q1 * scale[a] < hi) q1++;
		ql[a] = (uint32_t)q0; qh[a] = (uint32_t)q1;
	}
}

// 3. the node: child / triangle ranges, quantised child boxes, triangles in leaf order. `idx` is the final primitive order (leaf
// ranges of the BVH2 index into it). `defer`, when given, receives the node's triangle copies instead of having them done here:
// defer[0] = their number (<= 24), then (output slot, primitive) pairs -- the single-launch builder lets the lanes of a warp do
// them side by side (a copy is 18 dependent loads + 6 stores).
DRT_HD void collapse_emit(const BuildArrays& A, const int* idx, int w, const int* child, int n, const int* child_in_slot, int* defer)
{
	const BuildNode& parent = A.nodes[A.wide_src[w]];
	const float plo[3] = { dec_f(parent.lo[0]), dec_f(parent.lo[1]), dec_f(parent.lo[2]) };
	const float phi[3] = { dec_f(parent.hi[0]), dec_f(parent.hi[1]), dec_f(parent.hi[2]) };

	int n_inner = 0, n_tris = 0;
	for (int k = 0; k < n; k++) {
		const BuildNode& c = A.nodes[child[k]];
		if (c.left >= 0) n_inner++; else n_tris += c.count;
	}
	const int child_base = n_inner ? atomic_add_i(&A.state->wide_count, n_inner) : 0;
	const int tri_base = n_tris ? atomic_add_i(&A.state->tri_cursor, n_tris) : 0;

	WideNode out;
	out.px = plo[0]; out.py = plo[1]; out.pz = plo[2];
	uint32_t e[3];
	float scale[3];
	quant_grid(plo, phi, e, scale);
	out.ex = (uint8_t)e[0]; out.ey = (uint8_t)e[1]; out.ez = (uint8_t)e[2];
	out.imask = 0; out.child_base = (uint32_t)child_base; out.tri_base = (uint32_t)tri_base;
	int inner_rank = 0, tri_off = 0;
	for (int s = 0; s < 8; s++) {
		out.meta[s] = 0;
		out.qlox[s] = out.qloy[s] = out.qloz[s] = 0;
		out.qhix[s] = out.qhiy[s] = out.qhiz[s] = 0;
		int k = child_in_slot[s];
		if (k < 0) continue;
		const BuildNode& c = A.nodes[child[k]];
		uint32_t ql[3], qh[3];
		quant_child(c, plo, scale, ql, qh);
		out.qlox[s] = (uint8_t)ql[0]; out.qloy[s] = (uint8_t)ql[1]; out.qloz[s] = (uint8_t)ql[2];
		out.qhix[s] = (uint8_t)qh[0]; out.qhiy[s] = (uint8_t)qh[1]; out.qhiz[s] = (uint8_t)qh[2];
		if (c.left >= 0) {
			out.imask |= (uint8_t)(1u << s);
			out.meta[s] = (uint8_t)(0x20u | (24u + (uint32_t)s));
			A.wide_src[child_base + inner_rank] = child[k];
			inner_rank++;
		}
		else {
			const uint32_t unary = c.count == 1 ? 1u : (c.count == 2 ? 3u : 7u);
			out.meta[s] = (uint8_t)((unary << 5) | (uint32_t)tri_off);
			for (int t = 0; t < c.count; t++) {
				if (defer) { const int k2 = defer[0]++; defer[1 + 2 * k2] = tri_base + tri_off + t; defer[2 + 2 * k2] = idx[c.start + t]; }
				else write_triangle(A, tri_base + tri_off + t, idx[c.start + t]);
			}
			tri_off += c.count;
		}
	}
	A.wide[w] = out;
}

DRT_HD void collapse_node(const BuildArrays& A, const int* idx, int wide_level_start, int wide_level_end, int j, int* defer = nullptr)
{
	int w = wide_level_start + j;
	if (w >= wide_level_end) return;
	int child[8], slot_of[8], child_in_slot[8];
	float cost[64];
	const int n = collapse_children(A, w, child, cost);
	collapse_match(n, cost, slot_of, child_in_slot);
	collapse_emit(A, idx, w, child, n, child_in_slot, defer);
}

} // namespace drt

