// tests/hostemu/hostemu.cpp -- TEST INFRASTRUCTURE ONLY. Never linked into libdrt.so and never on a
// product path: libdrt.so has no CPU fallback and fails loudly without a CUDA device.
//
// The build container has no GPU, so the per-thread bodies of the device code (drt_*.cuh, written as
// __host__ __device__ functions) are compiled here as plain C++ and driven by serial / OpenMP loops.
// This lets the `-m "not gpu"` tests check the LOGIC of the kernels (builder passes, wide-node
// traversal, shading loop, RNG stream) against the reference oracle before any GPU time is spent.
// The real kernels (drt_kernels.cu) call the very same functions.
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../../dustraytracer_b200/csrc/drt_host.h"
#include "../../dustraytracer_b200/csrc/drt_params.h"
#include "../../dustraytracer_b200/csrc/drt_build_driver.h"
#include "../../dustraytracer_b200/csrc/drt_phased.cuh"

using namespace drt;

namespace drt {
static thread_local std::string g_err;
void set_error(const std::string& m) { g_err = m; }
}

namespace {

struct EmuExec : BuildExec {
	int launch(BuildPass pass, const BuildArrays& A, const int* idx_final, int a, int b, int n) override
	{
		for (int i = 0; i < n; i++) {
			switch (pass) {
			case PASS_PRIM_BOUNDS: prim_bounds(A, i); break;
			case PASS_INIT_BINS: init_bin_word(A, i, a); break;
			case PASS_BIN: bin_prim(A, i); break;
			case PASS_SPLIT: split_node(A, i); break;
			case PASS_PARTITION: partition_prim(A, i); break;
			case PASS_COLLAPSE: collapse_node(A, idx_final, a, b, i); break;
			}
		}
		return 0;
	}
	int read_state(const BuildArrays& A, BuildState& out) override { out = *A.state; return 0; }
	int write_state(const BuildArrays& A, const BuildState& in) override { *A.state = in; return 0; }
	int write_root(const BuildArrays& A, const BuildNode& root, const BuildSplit& split) override
	{
		A.nodes[0] = root; A.splits[0] = split; A.wide_src[0] = 0;
		return 0;
	}
};

struct EmuScene {
	HostScene host;
	std::vector<int32_t> mat_flags;
	std::vector<drt_f4> tri_lo, tri_hi, out_tris, out_shade, materials, nodes_f32;
	std::vector<float> materials_ext;
	std::vector<int> idx0, idx1, pn0, pn1, bins, wide_src;
	std::vector<BuildNode> nodes;
	std::vector<BuildSplit> splits;
	std::vector<WideNode> wide;
	std::vector<TexDesc> texdesc;
	std::vector<uint32_t> texels;
	BuildState state;
	BuildResult result;
	SceneView view;
};

void upload_tables(EmuScene* s)
{
	const HostScene& h = s->host;
	// textures: RGBA8 pool (same packing as the device upload in drt_api.cu)
	for (const HostTexture& t : h.textures) {
		TexDesc d;
		d.offset = (uint32_t)s->texels.size(); d.width = t.width; d.height = t.height; d.comps = t.comps;
		size_t np = (size_t)t.width * t.height;
		for (size_t p = 0; p < np; p++) {
			const uint8_t* px = t.pixels.data() + p * t.comps;
			uint32_t r = px[0], g = t.comps >= 3 ? px[1] : px[0], b = t.comps >= 3 ? px[2] : px[0];
			uint32_t a = t.comps == 4 ? px[3] : (t.comps == 2 ? px[1] : 255u);
			s->texels.push_back(r | (g << 8) | (b << 16) | (a << 24));
		}
		s->texdesc.push_back(d);
	}
	size_t nm = h.materials.size();
	s->materials.resize(nm + 1);
	for (size_t i = 0; i <= nm; i++) {
		drt_f4 m;
		if (i < nm) {
			int tex = h.materials[i].albedo_texture_index;
			if (tex >= (int)h.textures.size()) tex = -1;
			m.x = h.materials[i].albedo[0]; m.y = h.materials[i].albedo[1]; m.z = h.materials[i].albedo[2];
			m.w = as_float((uint32_t)tex);
		}
		else { m.x = m.y = m.z = 1.f; m.w = as_float((uint32_t)-1); } // default material for index -1
		s->materials[i] = m;
	}
	fill_materials_ext(h.materials, s->materials_ext);
	size_t n = h.tri_count();
	s->mat_flags.resize(n);
	for (size_t i = 0; i < n; i++) {
		int mi = h.mat[i];
		if (mi < 0 || (size_t)mi >= nm) mi = (int)nm;
		uint32_t flag = 0;
		if ((size_t)mi < nm) {
			int tex = h.materials[mi].albedo_texture_index;
			if (tex >= 0 && tex < (int)h.textures.size() && h.textures[tex].comps == 4) flag = 0x80000000u;
		}
		s->mat_flags[i] = (int32_t)((uint32_t)mi | flag);
	}
}

int build(EmuScene* s)
{
	upload_tables(s);
	const int n = (int)s->host.tri_count();
	s->tri_lo.resize(n); s->tri_hi.resize(n);
	s->out_tris.resize(3 * (size_t)n + 3); s->out_shade.resize(3 * (size_t)n + 3);
	s->idx0.resize(n); s->idx1.resize(n); s->pn0.resize(n); s->pn1.resize(n);
	s->nodes.resize(2 * (size_t)n + 2); s->splits.resize(2 * (size_t)n + 2);
	s->wide.resize((size_t)n + 1); s->wide_src.resize((size_t)n + 1);
	s->bins.resize(build_bin_slots_capacity(n) * 3 * DRT_BINS * DRT_BIN_WORDS);
	BuildArrays A;
	memset(&A, 0, sizeof(A));
	A.pos = s->host.pos.data(); A.uv = s->host.uv.data(); A.fnormal = s->host.fnormal.data(); A.mat = s->mat_flags.data();
	A.n = n;
	A.tri_lo = s->tri_lo.data(); A.tri_hi = s->tri_hi.data();
	A.idx_in = s->idx0.data(); A.idx_out = s->idx1.data(); A.pnode_in = s->pn0.data(); A.pnode_out = s->pn1.data();
	A.nodes = s->nodes.data(); A.splits = s->splits.data(); A.bins = s->bins.data(); A.state = &s->state;
	A.wide_src = s->wide_src.data(); A.wide = s->wide.data(); A.out_tris = s->out_tris.data(); A.out_shade = s->out_shade.data();
	EmuExec ex;
	int rc = run_build(ex, A, s->result);
	SceneView& v = s->view;
	v.nodes = (const drt_u4*)s->wide.data();
	v.tris = s->out_tris.data(); v.shade = s->out_shade.data(); v.materials = s->materials.data(); v.materials_ext = (const drt_f4*)s->materials_ext.data();
	v.textures = s->texdesc.data(); v.texels = s->texels.data();
	s->nodes_f32.resize((size_t)DRT_F32NODE_F4 * (s->result.wide_nodes > 0 ? s->result.wide_nodes : 1));
	for (int i = 0; i < s->result.wide_nodes; i++) expand_wide_node(s->wide[i], s->nodes_f32.data() + (size_t)DRT_F32NODE_F4 * i);
	v.nodes_f32 = s->nodes_f32.data();
	v.num_tris = n; v.num_nodes = s->result.wide_nodes; v.num_materials = (int)s->materials.size(); v.num_textures = (int)s->texdesc.size();
	return rc;
}

} // namespace

extern "C" {

const char* emu_last_error() { return g_err.c_str(); }

void* emu_scene_load(const char* path)
{
	EmuScene* s = new EmuScene();
	std::string err;
	if (load_gltf(path, s->host, err) != 0) { g_err = err; delete s; return nullptr; }
	if (build(s) != 0) { g_err = "build failed"; delete s; return nullptr; }
	return s;
}

void* emu_scene_from_triangles(const float* pos, const float* uv, const int32_t* mat, int n, const drt_material* mats, int nm)
{
	EmuScene* s = new EmuScene();
	HostScene& h = s->host;
	h.pos.assign(pos, pos + 9 * (size_t)n);
	if (uv) h.uv.assign(uv, uv + 6 * (size_t)n); else h.uv.assign(6 * (size_t)n, 0.f);
	h.vnormal.assign(9 * (size_t)n, 0.f);
	h.fnormal.resize(3 * (size_t)n);
	for (int i = 0; i < n; i++) {
		float zero9[9] = { 0 };
		compute_face_normal(pos + 9 * (size_t)i, zero9, h.fnormal.data() + 3 * (size_t)i);
		h.mat.push_back(mat ? mat[i] : 0);
	}
	for (int i = 0; i < nm; i++) h.materials.push_back(mats[i]);
	if (build(s) != 0) { delete s; return nullptr; }
	return s;
}

void emu_scene_free(void* h) { delete (EmuScene*)h; }

// info[0..7] = tris, materials, textures, wide nodes, bvh2 nodes, levels, wide levels, tris emitted
void emu_scene_info(void* h, long long* info)
{
	EmuScene* s = (EmuScene*)h;
	info[0] = (long long)s->host.tri_count(); info[1] = (long long)s->host.materials.size(); info[2] = (long long)s->host.textures.size();
	info[3] = s->result.wide_nodes; info[4] = s->result.bvh2_nodes; info[5] = s->result.levels; info[6] = s->result.wide_levels;
	info[7] = s->result.tris_emitted;
}

// Triangle soup in load order in the reference's 128-byte layout (padding word zeroed).
void emu_scene_get_triangles(void* h, float* out)
{
	const HostScene& s = ((EmuScene*)h)->host;
	for (size_t i = 0; i < s.tri_count(); i++) {
		float* o = out + 32 * i;
		const float* p = &s.pos[9 * i]; const float* n = &s.vnormal[9 * i]; const float* t = &s.uv[6 * i];
		o[0] = (p[0] + p[3] + p[6]) / 3; o[1] = (p[1] + p[4] + p[7]) / 3; o[2] = (p[2] + p[5] + p[8]) / 3; o[3] = 0;
		for (int k = 0; k < 3; k++) {
			float* v = o + 4 + 8 * k;
			v[0] = p[3 * k]; v[1] = p[3 * k + 1]; v[2] = p[3 * k + 2];
			v[3] = n[3 * k]; v[4] = n[3 * k + 1]; v[5] = n[3 * k + 2];
			v[6] = t[2 * k]; v[7] = t[2 * k + 1];
		}
		o[28] = s.fnormal[3 * i]; o[29] = s.fnormal[3 * i + 1]; o[30] = s.fnormal[3 * i + 2];
		int32_t m = s.mat[i];
		memcpy(o + 31, &m, 4);
	}
}

void emu_scene_get_materials(void* h, drt_material* out)
{
	const HostScene& s = ((EmuScene*)h)->host;
	memcpy(out, s.materials.data(), s.materials.size() * sizeof(drt_material));
}

void emu_scene_get_texture_info(void* h, int i, int* w, int* hh, int* c)
{
	const HostTexture& t = ((EmuScene*)h)->host.textures[i];
	*w = t.width; *hh = t.height; *c = t.comps;
}
void emu_scene_get_texture_pixels(void* h, int i, uint8_t* out)
{
	const HostTexture& t = ((EmuScene*)h)->host.textures[i];
	memcpy(out, t.pixels.data(), t.pixels.size());
}
void emu_scene_get_wide_nodes(void* h, void* out) { EmuScene* s = (EmuScene*)h; memcpy(out, s->wide.data(), (size_t)s->result.wide_nodes * sizeof(WideNode)); }
void emu_scene_get_leaf_triangles(void* h, void* out) { EmuScene* s = (EmuScene*)h; memcpy(out, s->out_tris.data(), (size_t)s->host.tri_count() * 48); }

static RenderParams make_params(EmuScene* s, const drt_camera* cam, const drt_settings* st, int W, int H)
{
	RenderParams P;
	memset(&P, 0, sizeof(P));
	P.scene = s->view;
	P.cam = make_camera_params(*cam, (uint32_t)W, (uint32_t)H);
	drt_settings dflt;
	if (!st) { default_settings(&dflt); st = &dflt; }
	P.shade = make_shade_params(*st);
	P.width = W; P.height = H;
	return P;
}

// stats[0..5] = node visits, tri tests, ties, alpha tests, max stack, texel clamps
void emu_trace_primary(void* h, const drt_camera* cam, int W, int H, unsigned frameidx, int strict, int rtl,
	drt_hit* out, float* ray_od6, unsigned long long* stats)
{
	EmuScene* s = (EmuScene*)h;
	RenderParams P = make_params(s, cam, nullptr, W, H);
	unsigned long long tot[6] = { 0, 0, 0, 0, 0, 0 };
#pragma omp parallel for schedule(dynamic, 8)
	for (int y = 0; y < H; y++) {
		TraceStats ts;
		memset(&ts, 0, sizeof(ts));
		for (int x = 0; x < W; x++) {
			uint32_t seed = (uint32_t)x + (uint32_t)y * (uint32_t)W;
			seed *= frameidx;
			RayOD r = strict ? camera_get_ray<true>(P.cam, x, y, W, H, seed, rtl != 0) : camera_get_ray<false>(P.cam, x, y, W, H, seed, rtl != 0);
			Hit hit;
			if (strict) trace_ray<true, false, true>(P.scene, r.o, r.d, FLT_MAX, hit, &ts);
			else trace_ray<false, false, true>(P.scene, r.o, r.d, FLT_MAX, hit, &ts);
			size_t p = (size_t)x + (size_t)y * W;
			out[p].prim_id = hit.prim >= 0 ? (int32_t)as_uint(P.scene.tris[3 * (size_t)hit.prim].w) : -1;
			out[p].t = hit.prim >= 0 ? hit.t : -1.f; out[p].u = hit.u; out[p].v = hit.v;
			if (ray_od6) { float* o = ray_od6 + 6 * p; o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z; }
		}
#pragma omp critical
		{
			tot[0] += ts.nodes; tot[1] += ts.tris; tot[2] += ts.ties; tot[3] += ts.alpha; tot[5] += ts.clamps;
			if (ts.max_stack > tot[4]) tot[4] = ts.max_stack;
		}
	}
	if (stats) memcpy(stats, tot, sizeof(tot));
}

// Accumulates frames first, first+stride, ... (n of them) into acc (W*H*3 floats), sequentially per pixel.
// counters[0..7] = closest rays, shadow rays, node visits, tri tests, ties, alpha tests, max stack, clamps
void emu_render(void* h, const drt_camera* cam, const drt_settings* st, int W, int H, unsigned first, unsigned stride,
	unsigned n, int strict, int rtl, float* acc, unsigned long long* counters)
{
	EmuScene* s = (EmuScene*)h;
	RenderParams P = make_params(s, cam, st, W, H);
	unsigned long long tot[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
#pragma omp parallel for schedule(dynamic, 4)
	for (int y = 0; y < H; y++) {
		TraceStats ts;
		memset(&ts, 0, sizeof(ts));
		PathCounters pc;
		pc.closest = pc.shadow = 0;
		unsigned long long c0 = 0, c1 = 0;
		for (int x = 0; x < W; x++) {
			float* a = acc + 3 * ((size_t)x + (size_t)y * W);
			for (unsigned k = 0; k < n; k++) {
				unsigned f = first + k * stride;
				V3 c = strict ? path_radiance<true, true>(P, x, y, f, rtl != 0, pc, &ts) : path_radiance<false, true>(P, x, y, f, rtl != 0, pc, &ts);
				a[0] += c.x; a[1] += c.y; a[2] += c.z;
			}
			c0 += pc.closest; c1 += pc.shadow; pc.closest = pc.shadow = 0;
		}
#pragma omp critical
		{
			tot[0] += c0; tot[1] += c1; tot[2] += ts.nodes; tot[3] += ts.tris; tot[4] += ts.ties; tot[5] += ts.alpha; tot[7] += ts.clamps;
			if (ts.max_stack > tot[6]) tot[6] = ts.max_stack;
		}
	}
	if (counters) memcpy(counters, tot, sizeof(tot));
}

// Same contract as emu_render, through the flat state machine the production kernel uses.
void emu_render_flat(void* h, const drt_camera* cam, const drt_settings* st, int W, int H, unsigned first, unsigned stride,
	unsigned n, int strict, int rtl, float* acc, unsigned long long* counters)
{
	EmuScene* s = (EmuScene*)h;
	RenderParams P = make_params(s, cam, st, W, H);
	P.frame_first = first; P.frame_stride = stride; P.n_samples = n;
	unsigned long long c0 = 0, c1 = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : c0, c1)
	for (int y = 0; y < H; y++) {
		PathCounters pc;
		pc.closest = pc.shadow = 0;
		for (int x = 0; x < W; x++) {
			float* a = acc + 3 * ((size_t)x + (size_t)y * W);
			V3 in = mk3(a[0], a[1], a[2]);
			V3 r = strict ? render_pixel_flat<true, false>(P, x, y, rtl != 0, in, pc, nullptr) : render_pixel_flat<false, false>(P, x, y, rtl != 0, in, pc, nullptr);
			a[0] = r.x; a[1] = r.y; a[2] = r.z;
		}
		c0 += pc.closest; c1 += pc.shadow;
	}
	if (counters) { counters[0] = c0; counters[1] = c1; }
}

// The reference-shaped sample loop with the extended material branch (DRT_FLAG_EXTENDED_MATERIALS), production arithmetic.
void emu_render_ext(void* h, const drt_camera* cam, const drt_settings* st, int W, int H, unsigned first, unsigned stride,
	unsigned n, int strict, int rtl, float* acc, unsigned long long* counters)
{
	EmuScene* s = (EmuScene*)h;
	RenderParams P = make_params(s, cam, st, W, H);
	unsigned long long c0 = 0, c1 = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : c0, c1)
	for (int y = 0; y < H; y++) {
		PathCounters pc;
		pc.closest = pc.shadow = 0;
		TraceStats ts;
		memset(&ts, 0, sizeof(ts));
		for (int x = 0; x < W; x++) {
			float* a = acc + 3 * ((size_t)x + (size_t)y * W);
			for (unsigned k = 0; k < n; k++) {
				const unsigned f = first + k * stride;
				V3 c = strict ? path_radiance<true, true, true>(P, x, y, f, rtl != 0, pc, &ts) : path_radiance<false, true, true>(P, x, y, f, rtl != 0, pc, &ts);
				if (strict) { a[0] = M<true>::add(a[0], c.x); a[1] = M<true>::add(a[1], c.y); a[2] = M<true>::add(a[2], c.z); }
				else { a[0] += c.x; a[1] += c.y; a[2] += c.z; }
			}
		}
		c0 += pc.closest; c1 += pc.shadow;
	}
	if (counters) { counters[0] = c0; counters[1] = c1; }
}

// Same contract as emu_render_flat, through the NODE / TRI / SHADE steps of the phased kernel (serial schedule).
// fmt: 0 = compressed nodes, 1 = expanded fp32 nodes. counters[2..3] = node visits, tri tests.
void emu_render_steps(void* h, const drt_camera* cam, const drt_settings* st, int W, int H, unsigned first, unsigned stride,
	unsigned n, int strict, int rtl, int fmt, float* acc, unsigned long long* counters)
{
	EmuScene* s = (EmuScene*)h;
	RenderParams P = make_params(s, cam, st, W, H);
	P.frame_first = first; P.frame_stride = stride; P.n_samples = n;
	unsigned long long c0 = 0, c1 = 0, c2 = 0, c3 = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : c0, c1, c2, c3)
	for (int y = 0; y < H; y++) {
		PathCounters pc;
		pc.closest = pc.shadow = 0;
		TraceStats ts;
		memset(&ts, 0, sizeof(ts));
		for (int x = 0; x < W; x++) {
			float* a = acc + 3 * ((size_t)x + (size_t)y * W);
			V3 in = mk3(a[0], a[1], a[2]), r;
			const bool rt = rtl != 0;
			// fmt: 0 = compressed nodes, 1 = expanded nodes, 2 = expanded nodes + the two-triangle TRI step (tri_step2)
			if (fmt == 2) r = strict ? render_pixel_steps<true, true, NODE_FMT_F32, true>(P, x, y, rt, in, pc, &ts) : render_pixel_steps<false, true, NODE_FMT_F32, true>(P, x, y, rt, in, pc, &ts);
			else if (strict) r = fmt ? render_pixel_steps<true, true, NODE_FMT_F32>(P, x, y, rt, in, pc, &ts) : render_pixel_steps<true, true, NODE_FMT_Q8>(P, x, y, rt, in, pc, &ts);
			else r = fmt ? render_pixel_steps<false, true, NODE_FMT_F32>(P, x, y, rt, in, pc, &ts) : render_pixel_steps<false, true, NODE_FMT_Q8>(P, x, y, rt, in, pc, &ts);
			a[0] = r.x; a[1] = r.y; a[2] = r.z;
		}
		c0 += pc.closest; c1 += pc.shadow; c2 += ts.nodes; c3 += ts.tris;
	}
	if (counters) { counters[0] = c0; counters[1] = c1; counters[2] = c2; counters[3] = c3; }
}

// Design tool (tools/sched_sim.py): the sequence of phase steps each lane of one 8x4 tile executes, as a
// string over N (node step), T (triangle step), H (closest-hit processing), U (one random-unit-vector
// attempt: sun shadow direction or one try of the scatter rejection loop), E (sample end: sky on a miss,
// post-process, next camera ray). The sequence of a lane does not depend on how a warp schedules it.
int emu_phase_trace(void* h, const drt_camera* cam, const drt_settings* st, int W, int H, int tile_x, int tile_y,
	unsigned first, unsigned n, char* out, int cap_per_lane)
{
	EmuScene* s = (EmuScene*)h;
	RenderParams P = make_params(s, cam, st, W, H);
	P.frame_first = first; P.frame_stride = 1; P.n_samples = n;
	const bool rtl = false;
	int longest = 0;
	for (int lane = 0; lane < 32; lane++) {
		char* o = out + (size_t)lane * cap_per_lane;
		int len = 0;
		const int x = tile_x * 8 + (lane & 7), y = tile_y * 4 + (lane >> 3);
		o[0] = 0;
		if (x >= W || y >= H) continue;
		LanePath S; LaneTrav L; LaneStack K; RayBox R; PathCounters pc; pc.closest = pc.shadow = 0;
		S.acc = mk3(0, 0, 0); S.k = 0; S.fn = mk3(0, 0, 0);
		path_begin_sample<false>(P, x, y, rtl, S);
		ray_box_setup(S.o, S.d, R);
		trav_begin(L, FLT_MAX);
		int phase = PH_NODE;
		while (phase != PH_DONE && len < cap_per_lane - 8) {
			if (phase == PH_NODE) { o[len++] = 'N'; node_step_q8<false>(P.scene.nodes, L, K, R, nullptr); phase = trav_advance(L, K); }
			else if (phase == PH_TRI) {
				o[len++] = 'T';
				if (tri_step<false, false>(P.scene, P.scene.tris, L, S.o, S.d, S.shadow, nullptr)) phase = PH_SHADE; else phase = trav_advance(L, K);
			}
			else {
				const bool was_shadow = S.shadow, found = L.hit.prim >= 0;
				const unsigned k0 = S.k;
				uint32_t seed = S.seed;
				const int bounce0 = S.bounce;
				int attempts = 0;
				if (was_shadow || (found && !P.shade.enable_sunlight)) { // count the tries of the scatter rejection loop
					if (!was_shadow) seed += (uint32_t)bounce0;
					for (;;) { attempts++; V3 p = random_unit_vec3<false>(seed, rtl); float l = M<false>::length(p); if (l * l < 1.f) break; }
				}
				const bool alive = shade_step<false, false>(P, x, y, rtl, L.hit, S, pc, nullptr);
				if (!was_shadow && found) o[len++] = 'H';
				if (!was_shadow && found && P.shade.enable_sunlight) o[len++] = 'U';
				for (int a = 0; a < attempts; a++) o[len++] = 'U';
				if (S.k != k0 || !alive) o[len++] = 'E';
				if (!alive) phase = PH_DONE;
				else { ray_box_setup(S.o, S.d, R); trav_begin(L, FLT_MAX); phase = PH_NODE; }
			}
		}
		o[len] = 0;
		if (len > longest) longest = len;
	}
	return longest;
}

// Five-step decomposition (drt_phased.cuh), serial schedule; nodes_first exercises triangle postponement.
void emu_render_steps5(void* h, const drt_camera* cam, const drt_settings* st, int W, int H, unsigned first, unsigned stride,
	unsigned n, int strict, int rtl, int fmt, int nodes_first, float* acc, unsigned long long* counters)
{
	EmuScene* s = (EmuScene*)h;
	RenderParams P = make_params(s, cam, st, W, H);
	P.frame_first = first; P.frame_stride = stride; P.n_samples = n;
	unsigned long long c0 = 0, c1 = 0, c2 = 0, c3 = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : c0, c1, c2, c3)
	for (int y = 0; y < H; y++) {
		PathCounters pc;
		pc.closest = pc.shadow = 0;
		TraceStats ts;
		memset(&ts, 0, sizeof(ts));
		for (int x = 0; x < W; x++) {
			float* a = acc + 3 * ((size_t)x + (size_t)y * W);
			V3 in = mk3(a[0], a[1], a[2]), r;
			const bool rt = rtl != 0, nf = nodes_first != 0;
			if (strict) r = fmt ? render_pixel_steps5<true, true, NODE_FMT_F32>(P, x, y, rt, in, pc, &ts, nf) : render_pixel_steps5<true, true, NODE_FMT_Q8>(P, x, y, rt, in, pc, &ts, nf);
			else r = fmt ? render_pixel_steps5<false, true, NODE_FMT_F32>(P, x, y, rt, in, pc, &ts, nf) : render_pixel_steps5<false, true, NODE_FMT_Q8>(P, x, y, rt, in, pc, &ts, nf);
			a[0] = r.x; a[1] = r.y; a[2] = r.z;
		}
		c0 += pc.closest; c1 += pc.shadow; c2 += ts.nodes; c3 += ts.tris;
	}
	if (counters) { counters[0] = c0; counters[1] = c1; counters[2] = c2; counters[3] = c3; }
}

// Design tool (tools/sched_sim.py): emulates ONE WARP of the five-step kernel over an 8x4 tile with the vote
// score_s = lanes wanting s * weight[s] (+ sticky bonus for the step that ran last), executes the winner for
// every lane that wants it, and reports per step: executions, lane-steps. overlap = 0 makes a lane with
// pending triangles ineligible for NODE (the exclusive schedule of the three-phase kernel).
// out[0..4] = executions, out[5..9] = lane-steps, out[10] = rays, out[11] = node visits
void emu_warp_sim(void* h, const drt_camera* cam, const drt_settings* st, int W, int H, int tile_x, int tile_y, unsigned first,
	unsigned n, const float* weight, float sticky, int overlap, int pool, unsigned long long* out)
{
	if (pool < 1) pool = 1;
	const int NL = 32 * pool; // contexts owned by the warp: `pool` horizontally adjacent tiles
	EmuScene* s = (EmuScene*)h;
	RenderParams P = make_params(s, cam, st, W, H);
	P.frame_first = first; P.frame_stride = 1; P.n_samples = n;
	std::vector<Lane5> lanes(NL);
	PathCounters pc; pc.closest = pc.shadow = 0;
	TraceStats ts; memset(&ts, 0, sizeof(ts));
	auto px = [&](int l) { return (tile_x + l / 32) * 8 + (l & 7); };
	auto py = [&](int l) { return tile_y * 4 + ((l & 31) >> 3); };
	for (int l = 0; l < NL; l++) lane5_init<false>(P, px(l), py(l), px(l) < W && py(l) < H, false, mk3(0, 0, 0), lanes[l]);
	int last = -1;
	for (;;) {
		int cnt[ST_COUNT] = { 0, 0, 0, 0, 0 };
		for (int l = 0; l < NL; l++)
			for (int k = 0; k < ST_COUNT; k++) {
				bool w = lane5_wants(lanes[l], k);
				if (k == ST_NODE && !overlap && lanes[l].L.tri_bits != 0u) w = false;
				if (w) cnt[k]++;
			}
		int best = -1; float bs = 0.f;
		for (int k = 0; k < ST_COUNT; k++) {
			if (!cnt[k]) continue;
			float sc = (cnt[k] > 32 ? 32 : cnt[k]) * weight[k] + (k == last ? sticky : 0.f);
			if (best < 0 || sc > bs) { best = k; bs = sc; }
		}
		if (best < 0) break;
		last = best;
		out[best]++;
		int served = 0;
		for (int l = 0; l < NL && served < 32; l++) {
			bool w = lane5_wants(lanes[l], best);
			if (best == ST_NODE && !overlap && lanes[l].L.tri_bits != 0u) w = false;
			if (!w) continue;
			out[5 + best]++;
			served++;
			lane5_run<false, true, NODE_FMT_Q8>(P, px(l), py(l), false, best, lanes[l], pc, &ts);
		}
	}
	out[10] += pc.closest + pc.shadow;
	out[11] += ts.nodes;
}

// Design tool (tools/sched_sim3.py): emulates ONE WARP of the three-step kernel actually shipped (render_warp_phased:
// NODE / TRI / SHADE, full vote + sticky re-issue) over an 8x4 tile, with the policy as parameters:
//   pol[0..2] = weight of the NODE / TRI / SHADE lane count in the arg-max
//   pol[3..5] = eager threshold: if > 0 and at least this many lanes want TRI (pol[4]) / SHADE (pol[5]) / NODE (pol[3]),
//               that step wins outright (checked in the order TRI, SHADE, NODE)
//   pol[6..8] = sticky threshold per step (re-issue while at least this many lanes want it again; 33 = never)
// out[0..2] = issues per step, out[3..5] = lane-steps per step, out[6] = full votes, out[7] = sticky ballots,
// out[8] = rays, out[9] = sum over SHADE issues of the largest per-lane rejection-loop draw count (proxy of its cost)
void emu_warp_sim3(void* h, const drt_camera* cam, const drt_settings* st, int W, int H, int tile_x, int tile_y, unsigned first,
	unsigned n, const float* pol, unsigned long long* out)
{
	EmuScene* s = (EmuScene*)h;
	RenderParams P = make_params(s, cam, st, W, H);
	P.frame_first = first; P.frame_stride = 1; P.n_samples = n;
	struct Lane { LanePath S; LaneTrav L; LaneStack K; RayBox R; int phase; uint32_t x, y; bool more; };
	const bool split_end = pol[10] > 0; // four-step variant: NODE / TRI / SHADE / END (post-process + next camera ray)
	bool ended_in_issue = false;
	std::vector<Lane> lanes(32);
	PathCounters pc; pc.closest = pc.shadow = 0;
	TraceStats ts; memset(&ts, 0, sizeof(ts));
	const int first_phase = P.scene.num_nodes > 0 ? PH_NODE : PH_SHADE;
	for (int l = 0; l < 32; l++) {
		Lane& A = lanes[l];
		A.x = tile_x * 8 + (l & 7); A.y = tile_y * 4 + (l >> 3);
		A.S.acc = mk3(0, 0, 0); A.S.k = 0; A.S.fn = mk3(0, 0, 0);
		A.phase = PH_DONE;
		trav_begin(A.L, FLT_MAX);
		if ((int)A.x < W && (int)A.y < H && n > 0) {
			path_begin_sample<false>(P, A.x, A.y, false, A.S);
			ray_box_setup(A.S.o, A.S.d, A.R);
			A.phase = first_phase;
		}
	}
	auto run_step = [&](int pick) -> unsigned long long {
		unsigned long long worst = 0;
		for (int l = 0; l < 32; l++) {
			Lane& A = lanes[l];
			if (A.phase != pick) continue;
			out[3 + pick]++;
			if (pick == PH_NODE) {
				node_step_f32<true>(P.scene.nodes_f32, A.L, A.K, A.R, &ts);
				A.phase = trav_advance(A.L, A.K);
			}
			else if (pick == PH_TRI) {
				if (tri_step<false, true>(P.scene, P.scene.tris, A.L, A.S.o, A.S.d, A.S.shadow, &ts)) A.phase = PH_SHADE;
				else A.phase = trav_advance(A.L, A.K);
			}
			else {
				const uint32_t seed0 = A.S.seed, k0 = A.S.k;
				const bool more = shade_step<false, true>(P, A.x, A.y, false, A.L.hit, A.S, pc, &ts);
				unsigned long long draws = 0; uint32_t q = seed0;
				while (q != A.S.seed && draws < 4096) { q = pcg_hash(q); draws++; }
				if (draws < 4096 && draws > worst) worst = draws;
				const bool ended = !more || A.S.k != k0;
				if (ended) ended_in_issue = true;
				if (ended && split_end) { A.phase = 4; A.more = more; } // PH_END: the sample end becomes a step of its own
				else if (!more) A.phase = PH_DONE;
				else { ray_box_setup(A.S.o, A.S.d, A.R); trav_begin(A.L, FLT_MAX); A.phase = first_phase; }
			}
		}
		return worst;
	};
	// pol[9] > 0: no vote at all -- every iteration runs NODE, then TRI, then SHADE for whichever lanes want them, a step
	// being skipped while fewer than pol[3] / pol[4] / pol[5] lanes want it unless nothing else can run (out[6] = iterations)
	while (pol[9] > 0) {
		int c[3] = { 0, 0, 0 };
		for (int l = 0; l < 32; l++) if (lanes[l].phase < 3) c[lanes[l].phase]++;
		if (c[0] + c[1] + c[2] == 0) { out[8] += pc.closest + pc.shadow; return; }
		out[6]++;
		bool ran = false;
		for (int k = 0; k < 3; k++) {
			int cnt = 0;
			for (int l = 0; l < 32; l++) if (lanes[l].phase == k) cnt++;
			if (cnt == 0 || cnt < pol[3 + k]) continue;
			out[k]++;
			const unsigned long long w = run_step(k);
			if (k == PH_SHADE) out[9] += w;
			ran = true;
		}
		if (!ran) { // everything below its threshold: run the fullest
			int best = 0;
			for (int k = 1; k < 3; k++) if (c[k] > c[best]) best = k;
			out[best]++;
			const unsigned long long w = run_step(best);
			if (best == PH_SHADE) out[9] += w;
		}
	}
	for (;;) {
		int c[5] = { 0, 0, 0, 0, 0 };
		for (int l = 0; l < 32; l++) if (lanes[l].phase != PH_DONE) c[lanes[l].phase]++;
		if (c[0] + c[1] + c[2] + c[4] == 0) break;
		out[6]++;
		int pick = -1;
		if (pol[4] > 0 && c[PH_TRI] >= pol[4]) pick = PH_TRI;
		else if (pol[5] > 0 && c[PH_SHADE] >= pol[5]) pick = PH_SHADE;
		else if (pol[3] > 0 && c[PH_NODE] >= pol[3]) pick = PH_NODE;
		else {
			float best = -1.f;
			for (int k = 0; k < 3; k++) if (c[k] > 0 && c[k] * pol[k] > best) { best = c[k] * pol[k]; pick = k; }
			if (c[4] > 0 && c[4] * pol[11] > best) pick = 4;
		}
		if (pick == 4) { // END step: lanes move on to their next camera ray (the work itself was done inside shade_step)
			out[11]++;
			for (int l = 0; l < 32; l++) {
				Lane& A = lanes[l];
				if (A.phase != 4) continue;
				out[12]++;
				if (!A.more) A.phase = PH_DONE;
				else { ray_box_setup(A.S.o, A.S.d, A.R); trav_begin(A.L, FLT_MAX); A.phase = first_phase; }
			}
			continue;
		}
		for (;;) {
			out[pick]++;
			ended_in_issue = false;
			const unsigned long long worst = run_step(pick);
			if (pick == PH_SHADE && ended_in_issue) out[10]++;
			if (pick == PH_SHADE) out[9] += worst;
			int again = 0;
			for (int l = 0; l < 32; l++) if (lanes[l].phase == pick) again++;
			out[7]++;
			if (again < pol[6 + pick] || again == 0) break;
		}
	}
	out[8] += pc.closest + pc.shadow;
}

// Design tool for the multi-context kernel (render_warp_dual, drt_dual.cuh): one warp whose lanes each own J pixels (contexts;
// context j of lane l = pixel l of tile (tile_x + j, tile_y)). Per iteration the warp votes for the step the most LANES can
// run (a lane can run a step if any of its contexts waits for it), each such lane advances ONE context, sticky re-issue as
// in the shipped kernel. pol[0..2] weights, pol[3] = minimum lanes for SHADE unless nothing else can run, pol[6..8] sticky.
// out as in emu_warp_sim3 (issues 0..2, lane-steps 3..5, votes 6, sticky ballots 7, rays 8, worst draws 9).
void emu_warp_sim_multi(void* h, const drt_camera* cam, const drt_settings* st, int W, int H, int tile_x, int tile_y, unsigned first,
	unsigned n, int J, const float* pol, unsigned long long* out)
{
	EmuScene* s = (EmuScene*)h;
	RenderParams P = make_params(s, cam, st, W, H);
	P.frame_first = first; P.frame_stride = 1; P.n_samples = n;
	struct Ctx { LanePath S; LaneTrav L; LaneStack K; RayBox R; int phase; uint32_t x, y; };
	std::vector<Ctx> ctx(32 * J);
	PathCounters pc; pc.closest = pc.shadow = 0;
	TraceStats ts; memset(&ts, 0, sizeof(ts));
	const int first_phase = P.scene.num_nodes > 0 ? PH_NODE : PH_SHADE;
	for (int j = 0; j < J; j++)
		for (int l = 0; l < 32; l++) {
			Ctx& A = ctx[j * 32 + l];
			A.x = (tile_x + j) * 8 + (l & 7); A.y = tile_y * 4 + (l >> 3);
			A.S.acc = mk3(0, 0, 0); A.S.k = 0; A.S.fn = mk3(0, 0, 0);
			A.phase = PH_DONE;
			trav_begin(A.L, FLT_MAX);
			if ((int)A.x < W && (int)A.y < H && n > 0) {
				path_begin_sample<false>(P, A.x, A.y, false, A.S);
				ray_box_setup(A.S.o, A.S.d, A.R);
				A.phase = first_phase;
			}
		}
	auto lanes_wanting = [&](int k) { int c = 0; for (int l = 0; l < 32; l++) { bool w = false; for (int j = 0; j < J; j++) w |= ctx[j * 32 + l].phase == k; c += w; } return c; };
	auto run_step = [&](int pick) -> unsigned long long {
		unsigned long long worst = 0;
		for (int l = 0; l < 32; l++) {
			int sel = -1;
			for (int j = 0; j < J && sel < 0; j++) if (ctx[j * 32 + l].phase == pick) sel = j;
			if (sel < 0) continue;
			Ctx& A = ctx[sel * 32 + l];
			out[3 + pick]++;
			if (pick == PH_NODE) { node_step_f32<true>(P.scene.nodes_f32, A.L, A.K, A.R, &ts); A.phase = trav_advance(A.L, A.K); }
			else if (pick == PH_TRI) {
				if (tri_step<false, true>(P.scene, P.scene.tris, A.L, A.S.o, A.S.d, A.S.shadow, &ts)) A.phase = PH_SHADE;
				else A.phase = trav_advance(A.L, A.K);
			}
			else {
				const uint32_t seed0 = A.S.seed;
				const bool more = shade_step<false, true>(P, A.x, A.y, false, A.L.hit, A.S, pc, &ts);
				unsigned long long draws = 0; uint32_t q = seed0;
				while (q != A.S.seed && draws < 4096) { q = pcg_hash(q); draws++; }
				if (draws < 4096 && draws > worst) worst = draws;
				if (!more) A.phase = PH_DONE;
				else { ray_box_setup(A.S.o, A.S.d, A.R); trav_begin(A.L, FLT_MAX); A.phase = first_phase; }
			}
		}
		return worst;
	};
	for (;;) {
		int c[3];
		for (int k = 0; k < 3; k++) c[k] = lanes_wanting(k);
		if (c[0] + c[1] + c[2] == 0) break;
		out[6]++;
		int pick = -1;
		float best = -1.f;
		for (int k = 0; k < 3; k++) {
			if (c[k] == 0) continue;
			if (k == PH_SHADE && c[k] < pol[3] && c[0] + c[1] > 0) continue;
			if (c[k] * pol[k] > best) { best = c[k] * pol[k]; pick = k; }
		}
		for (;;) {
			out[pick]++;
			const unsigned long long worst = run_step(pick);
			if (pick == PH_SHADE) out[9] += worst;
			const int again = lanes_wanting(pick);
			out[7]++;
			if (again < pol[6 + pick] || again == 0) break;
		}
	}
	out[8] += pc.closest + pc.shadow;
}

void emu_camera_params(const drt_camera* cam, int W, int H, float* out20)
{
	CameraParams c = make_camera_params(*cam, W, H);
	memcpy(out20, &c, sizeof(float) * 20);
}

void emu_default_camera(drt_camera* c) { default_camera(c); }
void emu_default_settings(drt_settings* s) { default_settings(s); }

} // extern "C"
