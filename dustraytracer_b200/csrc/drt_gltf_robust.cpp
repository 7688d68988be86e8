// drt_gltf_robust.cpp -- spec-conforming glTF 2.0 ingest, selected by DRT_LOAD_ROBUST.
//
// The reference loader (DustRayTracer/src/Core/Scene/Scene.cu:120-178,192-207) only loads files shaped like its
// own assets; drt_gltf.cpp reproduces exactly that for parity. This file is the other mode and handles what the
// reference drops or misreads:
//   * the default scene's node HIERARCHY with matrix / TRS transforms (reference: every node, no transforms; Scene.cu:192-207)
//   * accessor.byteOffset, bufferView.byteStride, accessors living in any buffer (reference: index buffer + view offset; Scene.cu:150-158)
//   * u8 / u16 / u32 indices and non-indexed primitives; TRIANGLES, TRIANGLE_STRIP and TRIANGLE_FAN (reference: u16 lists; Scene.cu:161)
//   * a missing NORMAL / TEXCOORD_0 is geometric normal / zero UV (reference: silently accessor 0; Scene.cu:129-131)
//   * normalised integer and KHR_mesh_quantization style attributes
//   * baseColorTexture.index -> textures[].source -> image (reference: uses the texture index as the image index; Scene.cu:79)
//   * alphaMode: OPAQUE ignores texture alpha, MASK thresholds at alphaCutoff, BLEND keeps the reference's
//     cut-out rule "alpha < 1 is a hole" (AnyHit.cuh:17-25). Implemented as derived texture copies so the
//     device code and the 44-byte material record stay those of the reference.
// The emitted HostScene has the same shape as the reference-compatible one: a triangle soup in emission order.
#include "drt_host.h"
#include "drt_json.h"

#include <cmath>
#include <cstring>
#include <map>

namespace drt {

namespace {

struct RView { int buffer; size_t offset, length, stride; };

struct Accessor {
	const uint8_t* base = nullptr; // nullptr: accessor without bufferView = all zeros (glTF 2.0 section 3.6.2.3)
	size_t stride = 0, count = 0;
	int comp = 5126, ncomp = 1;
	bool normalized = false;
};

int comp_size(int c)
{
	switch (c) {
	case 5120: case 5121: return 1;
	case 5122: case 5123: return 2;
	case 5125: case 5126: return 4;
	}
	return 0;
}

int type_comps(const std::string& t)
{
	if (t == "SCALAR") return 1;
	if (t == "VEC2") return 2;
	if (t == "VEC3") return 3;
	if (t == "VEC4") return 4;
	if (t == "MAT2") return 4;
	if (t == "MAT3") return 9;
	if (t == "MAT4") return 16;
	return 0;
}

struct Doc {
	const JsonValue* accessors = nullptr;
	std::vector<RView> views;
	const std::vector<const uint8_t*>* bdata = nullptr;
	const std::vector<size_t>* blen = nullptr;
};

bool resolve_accessor(const Doc& D, int a, Accessor& A, std::string& err)
{
	if (!D.accessors || a < 0 || (size_t)a >= D.accessors->size()) { err = "accessor index " + std::to_string(a) + " out of range"; return false; }
	const JsonValue& ja = D.accessors->arr[a];
	if (ja.has("sparse")) { err = "sparse accessors are not supported"; return false; }
	A.comp = ja.get_int("componentType", 0);
	A.ncomp = type_comps(ja.get_str("type"));
	if (!ja.get_size("count", 0, &A.count) || A.count > ((size_t)1 << 40)) { err = "accessor " + std::to_string(a) + ": bad count"; return false; }
	const JsonValue* nv = ja.get("normalized");
	A.normalized = nv && nv->type == JsonValue::Bool && nv->b;
	const int cs = comp_size(A.comp);
	if (!cs || !A.ncomp) { err = "accessor " + std::to_string(a) + ": bad componentType / type"; return false; }
	const size_t elem = (size_t)cs * A.ncomp;
	const int bv = ja.get_int("bufferView", -1);
	if (bv < 0) { A.base = nullptr; A.stride = elem; return true; }
	if ((size_t)bv >= D.views.size()) { err = "accessor " + std::to_string(a) + ": bufferView out of range"; return false; }
	const RView& v = D.views[bv];
	size_t off = 0;
	if (!ja.get_size("byteOffset", 0, &off)) { err = "accessor " + std::to_string(a) + ": bad byteOffset"; return false; }
	A.stride = v.stride ? v.stride : elem;
	// off + stride * (count - 1) + elem <= length, written so that nothing can wrap (stride <= 65536, count <= 2^40)
	if (A.count && (off > v.length || elem > v.length - off || (A.count - 1) > (v.length - off - elem) / A.stride)) {
		err = "accessor " + std::to_string(a) + " runs past its bufferView"; return false;
	}
	A.base = (*D.bdata)[v.buffer] + v.offset + off;
	return true;
}

double read_comp(const Accessor& A, size_t i, int k)
{
	if (!A.base) return 0.0;
	const uint8_t* p = A.base + A.stride * i + (size_t)comp_size(A.comp) * k;
	switch (A.comp) {
	case 5126: { float f; memcpy(&f, p, 4); return f; }
	case 5125: { uint32_t u; memcpy(&u, p, 4); return (double)u; }
	case 5123: { uint16_t u; memcpy(&u, p, 2); return A.normalized ? u / 65535.0 : (double)u; }
	case 5122: { int16_t s; memcpy(&s, p, 2); return A.normalized ? std::fmax(s / 32767.0, -1.0) : (double)s; }
	case 5121: { uint8_t u = *p; return A.normalized ? u / 255.0 : (double)u; }
	case 5120: { int8_t s = (int8_t)*p; return A.normalized ? std::fmax(s / 127.0, -1.0) : (double)s; }
	}
	return 0.0;
}

uint32_t read_index(const Accessor& A, size_t i)
{
	if (!A.base) return 0;
	const uint8_t* p = A.base + A.stride * i;
	switch (A.comp) {
	case 5121: return *p;
	case 5123: { uint16_t u; memcpy(&u, p, 2); return u; }
	case 5125: { uint32_t u; memcpy(&u, p, 4); return u; }
	}
	return 0;
}

// column-major 4x4, as glTF stores it
struct Mat4 { double m[16]; };

Mat4 identity()
{
	Mat4 r;
	for (int i = 0; i < 16; i++) r.m[i] = (i % 5 == 0) ? 1.0 : 0.0;
	return r;
}

Mat4 mul(const Mat4& a, const Mat4& b)
{
	Mat4 r;
	for (int c = 0; c < 4; c++)
		for (int rw = 0; rw < 4; rw++) {
			double s = 0.0;
			for (int k = 0; k < 4; k++) s += a.m[k * 4 + rw] * b.m[c * 4 + k];
			r.m[c * 4 + rw] = s;
		}
	return r;
}

Mat4 node_local(const JsonValue& n)
{
	if (const JsonValue* jm = n.get("matrix")) {
		if (jm->size() == 16) {
			Mat4 r;
			for (int i = 0; i < 16; i++) r.m[i] = jm->arr[i].num;
			return r;
		}
	}
	double t[3] = { 0, 0, 0 }, q[4] = { 0, 0, 0, 1 }, s[3] = { 1, 1, 1 };
	if (const JsonValue* v = n.get("translation")) for (size_t k = 0; k < 3 && k < v->size(); k++) t[k] = v->arr[k].num;
	if (const JsonValue* v = n.get("rotation")) for (size_t k = 0; k < 4 && k < v->size(); k++) q[k] = v->arr[k].num;
	if (const JsonValue* v = n.get("scale")) for (size_t k = 0; k < 3 && k < v->size(); k++) s[k] = v->arr[k].num;
	const double x = q[0], y = q[1], z = q[2], w = q[3];
	Mat4 r = identity();
	r.m[0] = (1 - 2 * (y * y + z * z)) * s[0]; r.m[1] = (2 * (x * y + z * w)) * s[0]; r.m[2] = (2 * (x * z - y * w)) * s[0];
	r.m[4] = (2 * (x * y - z * w)) * s[1]; r.m[5] = (1 - 2 * (x * x + z * z)) * s[1]; r.m[6] = (2 * (y * z + x * w)) * s[1];
	r.m[8] = (2 * (x * z + y * w)) * s[2]; r.m[9] = (2 * (y * z - x * w)) * s[2]; r.m[10] = (1 - 2 * (x * x + y * y)) * s[2];
	r.m[12] = t[0]; r.m[13] = t[1]; r.m[14] = t[2];
	return r;
}

bool is_identity(const Mat4& M)
{
	for (int i = 0; i < 16; i++) if (M.m[i] != ((i % 5 == 0) ? 1.0 : 0.0)) return false;
	return true;
}

void xform_point(const Mat4& M, const double* p, float* out)
{
	for (int r = 0; r < 3; r++) out[r] = (float)(M.m[r] * p[0] + M.m[4 + r] * p[1] + M.m[8 + r] * p[2] + M.m[12 + r]);
}

// normals go through the cofactor matrix of the upper 3x3 (= det * inverse transpose: right for mirrored nodes too)
void xform_normal(const Mat4& M, const double* n, float* out)
{
	const double a = M.m[0], b = M.m[4], c = M.m[8], d = M.m[1], e = M.m[5], f = M.m[9], g = M.m[2], h = M.m[6], i = M.m[10];
	const double C[9] = { e * i - f * h, -(d * i - f * g), d * h - e * g, -(b * i - c * h), a * i - c * g, -(a * h - b * g), b * f - c * e, -(a * f - c * d), a * e - b * d };
	double r[3];
	for (int k = 0; k < 3; k++) r[k] = C[3 * k] * n[0] + C[3 * k + 1] * n[1] + C[3 * k + 2] * n[2];
	const double len = std::sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
	for (int k = 0; k < 3; k++) out[k] = len > 0 ? (float)(r[k] / len) : 0.f;
}

struct Walk {
	const JsonValue* nodes;
	const JsonValue* meshes;
	Doc doc;
	HostScene* out;
	std::vector<char> on_path;
	std::string err;
};

bool emit_primitive(Walk& W, const JsonValue& prim, const Mat4& M, size_t mesh_index, size_t prim_index)
{
	const std::string where = "mesh " + std::to_string(mesh_index) + " primitive " + std::to_string(prim_index);
	const int mode = prim.get_int("mode", 4);
	if (mode != 4 && mode != 5 && mode != 6) {
		W.out->warnings.push_back(where + ": mode " + std::to_string(mode) + " is not a triangle topology; skipped");
		return true;
	}
	const JsonValue* attrs = prim.get("attributes");
	const int a_pos = attrs ? attrs->get_int("POSITION", -1) : -1;
	if (a_pos < 0) { W.out->warnings.push_back(where + ": no POSITION; skipped"); return true; }
	const int a_nrm = attrs->get_int("NORMAL", -1), a_uv = attrs->get_int("TEXCOORD_0", -1), a_idx = prim.get_int("indices", -1);
	Accessor P, N, T, I;
	if (!resolve_accessor(W.doc, a_pos, P, W.err)) return false;
	if (P.ncomp != 3) { W.err = where + ": POSITION is not VEC3"; return false; }
	if (a_nrm >= 0 && (!resolve_accessor(W.doc, a_nrm, N, W.err) || N.ncomp != 3 || N.count < P.count)) { if (W.err.empty()) W.err = where + ": bad NORMAL accessor"; return false; }
	if (a_uv >= 0 && (!resolve_accessor(W.doc, a_uv, T, W.err) || T.ncomp != 2 || T.count < P.count)) { if (W.err.empty()) W.err = where + ": bad TEXCOORD_0 accessor"; return false; }
	if (a_idx >= 0) {
		if (!resolve_accessor(W.doc, a_idx, I, W.err)) return false;
		if (I.ncomp != 1 || (I.comp != 5121 && I.comp != 5123 && I.comp != 5125)) { W.err = where + ": indices must be unsigned SCALARs"; return false; }
	}
	const size_t n_idx = a_idx >= 0 ? I.count : P.count;
	const size_t n_tri = mode == 4 ? n_idx / 3 : (n_idx >= 3 ? n_idx - 2 : 0);
	const int material = prim.get_int("material", -1);
	auto index_at = [&](size_t k) -> uint32_t { return a_idx >= 0 ? read_index(I, k) : (uint32_t)k; };
	const bool ident = is_identity(M); // untransformed meshes keep their stored bits (signed zeros, unnormalised normals)
	for (size_t t = 0; t < n_tri; t++) {
		size_t k[3];
		if (mode == 4) { k[0] = 3 * t; k[1] = 3 * t + 1; k[2] = 3 * t + 2; }
		else if (mode == 5) { k[0] = t; k[1] = t + 1 + (t & 1); k[2] = t + 2 - (t & 1); } // strips keep their winding
		else { k[0] = 0; k[1] = t + 1; k[2] = t + 2; }
		float p9[9], n9[9], t6[6];
		for (int c = 0; c < 3; c++) {
			const uint32_t vi = index_at(k[c]);
			if (vi >= P.count) { W.err = where + ": vertex index " + std::to_string(vi) + " out of range"; return false; }
			const double p[3] = { read_comp(P, vi, 0), read_comp(P, vi, 1), read_comp(P, vi, 2) };
			if (ident) for (int q = 0; q < 3; q++) p9[3 * c + q] = (float)p[q];
			else xform_point(M, p, p9 + 3 * c);
			if (a_nrm >= 0) {
				const double n[3] = { read_comp(N, vi, 0), read_comp(N, vi, 1), read_comp(N, vi, 2) };
				if (ident) for (int q = 0; q < 3; q++) n9[3 * c + q] = (float)n[q];
				else xform_normal(M, n, n9 + 3 * c);
			}
			else n9[3 * c] = n9[3 * c + 1] = n9[3 * c + 2] = 0.f; // face normal = the winding's geometric normal
			t6[2 * c] = a_uv >= 0 ? (float)read_comp(T, vi, 0) : 0.f;
			t6[2 * c + 1] = a_uv >= 0 ? (float)read_comp(T, vi, 1) : 0.f;
		}
		float fn[3];
		compute_face_normal(p9, n9, fn);
		if (!(fn[0] == fn[0]) || !(fn[1] == fn[1]) || !(fn[2] == fn[2])) continue; // zero-area triangle: nothing can hit it
		W.out->pos.insert(W.out->pos.end(), p9, p9 + 9);
		W.out->vnormal.insert(W.out->vnormal.end(), n9, n9 + 9);
		W.out->uv.insert(W.out->uv.end(), t6, t6 + 6);
		W.out->fnormal.insert(W.out->fnormal.end(), fn, fn + 3);
		W.out->mat.push_back(material);
	}
	return true;
}

bool walk_node(Walk& W, int ni, const Mat4& parent, int depth)
{
	if (ni < 0 || (size_t)ni >= W.nodes->size()) { W.err = "node index " + std::to_string(ni) + " out of range"; return false; }
	if (W.on_path[ni] || depth > 256) { W.err = "node hierarchy has a cycle at node " + std::to_string(ni); return false; }
	W.on_path[ni] = 1;
	const JsonValue& n = W.nodes->arr[ni];
	const Mat4 world = mul(parent, node_local(n));
	const int mi = n.get_int("mesh", -1);
	if (mi >= 0) {
		if ((size_t)mi >= W.meshes->size()) { W.err = "mesh index " + std::to_string(mi) + " out of range"; return false; }
		W.out->meshes++;
		if (const JsonValue* prims = W.meshes->arr[mi].get("primitives"))
			for (size_t pi = 0; pi < prims->size(); pi++)
				if (!emit_primitive(W, prims->arr[pi], world, (size_t)mi, pi)) return false;
	}
	if (const JsonValue* ch = n.get("children"))
		for (size_t c = 0; c < ch->size(); c++)
			if (!walk_node(W, (int)ch->arr[c].num, world, depth + 1)) return false;
	W.on_path[ni] = 0;
	return true;
}

} // namespace

int emit_robust_geometry(const JsonValue& root, const std::vector<const uint8_t*>& bdata, const std::vector<size_t>& blen,
	HostScene& out, std::string& err)
{
	static const JsonValue empty_array = [] { JsonValue v; v.type = JsonValue::Array; return v; }();
	auto arr = [&](const char* key) -> const JsonValue& { const JsonValue* v = root.get(key); return (v && v->type == JsonValue::Array) ? *v : empty_array; };
	if (const JsonValue* req = root.get("extensionsRequired"))
		for (size_t i = 0; i < req->size(); i++) {
			const std::string& e = req->arr[i].str;
			if (e == "KHR_draco_mesh_compression" || e == "EXT_meshopt_compression" || e == "KHR_texture_basisu") {
				err = "required extension " + e + " is not supported"; return DRT_ERR_FORMAT;
			}
		}
	Walk W;
	W.nodes = &arr("nodes");
	W.meshes = &arr("meshes");
	W.out = &out;
	W.on_path.assign(W.nodes->size(), 0);
	W.doc.accessors = &arr("accessors");
	W.doc.bdata = &bdata;
	W.doc.blen = &blen;
	const JsonValue& jviews = arr("bufferViews");
	W.doc.views.resize(jviews.size());
	for (size_t i = 0; i < jviews.size(); i++) { // ranges were validated by load_gltf
		const JsonValue& v = jviews.arr[i];
		RView rv{ v.get_int("buffer", 0), 0, 0, 0 };
		v.get_size("byteOffset", 0, &rv.offset); v.get_size("byteLength", 0, &rv.length); v.get_size("byteStride", 0, &rv.stride); // (validated by load_gltf)
		W.doc.views[i] = rv;
	}
	// roots: the default scene's, else scene 0's, else every node that is nobody's child
	std::vector<int> roots;
	const JsonValue& scenes = arr("scenes");
	if (scenes.size()) {
		int si = root.get_int("scene", 0);
		if (si < 0 || (size_t)si >= scenes.size()) si = 0;
		if (const JsonValue* rn = scenes.arr[si].get("nodes"))
			for (size_t i = 0; i < rn->size(); i++) roots.push_back((int)rn->arr[i].num);
	}
	else {
		std::vector<char> is_child(W.nodes->size(), 0);
		for (size_t i = 0; i < W.nodes->size(); i++)
			if (const JsonValue* ch = W.nodes->arr[i].get("children"))
				for (size_t c = 0; c < ch->size(); c++) {
					const int k = (int)ch->arr[c].num;
					if (k >= 0 && (size_t)k < is_child.size()) is_child[k] = 1;
				}
		for (size_t i = 0; i < W.nodes->size(); i++) if (!is_child[i]) roots.push_back((int)i);
	}
	const Mat4 I = identity();
	for (int r : roots)
		if (!walk_node(W, r, I, 0)) { err = W.err; return DRT_ERR_FORMAT; }
	return DRT_OK;
}

// Texture indirection and alpha modes. `out.textures` holds one entry per IMAGE on entry (image order); derived
// copies (alpha dropped or thresholded) are appended so that the device's rule "a 4-component texture with alpha < 255
// is a hole" implements OPAQUE / MASK / BLEND without a new material field.
int remap_robust_materials(const JsonValue& root, const std::vector<bool>& image_ok, HostScene& out, std::string& err)
{
	(void)err;
	const JsonValue* jmats = root.get("materials");
	const JsonValue* jtex = root.get("textures");
	if (!jmats) return DRT_OK;
	std::map<std::pair<int, int>, int> derived; // (image, -1 = opaque | threshold byte 0..255) -> texture index
	for (size_t i = 0; i < jmats->size() && i < out.materials.size(); i++) {
		const JsonValue& jm = jmats->arr[i];
		drt_material& m = out.materials[i];
		// fields of the extended material model (DRT_FLAG_EXTENDED_MATERIALS): metallic as a threshold on the factor
		// (the reference stores the factor in a bool, so any non-zero factor is "metallic" there), KHR extensions
		if (const JsonValue* pbr = jm.get("pbrMetallicRoughness")) m.metallic = pbr->get_num("metallicFactor", 1.0) >= 0.5 ? 1 : 0;
		if (const JsonValue* ext = jm.get("extensions")) {
			if (const JsonValue* t = ext->get("KHR_materials_transmission")) m.transmission = t->get_num("transmissionFactor", 0.0) > 0.0 ? 1 : 0;
			if (const JsonValue* t = ext->get("KHR_materials_ior")) m.refractive_index = (float)t->get_num("ior", 1.5);
			if (const JsonValue* t = ext->get("KHR_materials_emissive_strength")) {
				const float k = (float)t->get_num("emissiveStrength", 1.0);
				for (int c = 0; c < 3; c++) m.emissive_factor[c] *= k;
			}
		}
		int image = -1;
		if (const JsonValue* pbr = jm.get("pbrMetallicRoughness"))
			if (const JsonValue* t = pbr->get("baseColorTexture")) {
				const int ti = t->get_int("index", -1);
				if (jtex && ti >= 0 && (size_t)ti < jtex->size()) image = jtex->arr[ti].get_int("source", -1);
				if (t->get_int("texCoord", 0) != 0) out.warnings.push_back("material " + std::to_string(i) + ": baseColorTexture uses TEXCOORD_" + std::to_string(t->get_int("texCoord", 0)) + "; TEXCOORD_0 is sampled");
			}
		if (image < 0 || (size_t)image >= out.textures.size() || (size_t)image >= image_ok.size() || !image_ok[image]) { m.albedo_texture_index = -1; continue; }
		m.albedo_texture_index = image;
		const HostTexture& src = out.textures[image];
		if (src.comps != 4) continue; // no alpha channel: every mode behaves the same
		const std::string mode = jm.get_str("alphaMode", "OPAQUE");
		if (mode == "BLEND") continue; // the reference's cut-out rule as is
		int key = -1;
		if (mode == "MASK") {
			const double cutoff = jm.get_num("alphaCutoff", 0.5);
			key = (int)std::ceil(cutoff * 255.0 - 1e-9); // alpha/255 >= cutoff  <=>  alpha >= ceil(cutoff*255)
			key = key < 0 ? 0 : (key > 256 ? 256 : key);
		}
		auto it = derived.find(std::make_pair(image, key));
		if (it == derived.end()) {
			HostTexture t;
			t.width = src.width; t.height = src.height;
			const size_t n = (size_t)src.width * src.height;
			if (key < 0) { // OPAQUE: drop the alpha channel (also skips the alpha test on the device)
				t.comps = 3;
				t.pixels.resize(3 * n);
				for (size_t p = 0; p < n; p++) memcpy(&t.pixels[3 * p], &src.pixels[4 * p], 3);
			}
			else {
				t.comps = 4;
				t.pixels = src.pixels;
				for (size_t p = 0; p < n; p++) t.pixels[4 * p + 3] = (int)t.pixels[4 * p + 3] >= key ? 255 : 0;
			}
			out.textures.push_back(std::move(t));
			it = derived.emplace(std::make_pair(image, key), (int)out.textures.size() - 1).first;
		}
		m.albedo_texture_index = it->second;
	}
	return DRT_OK;
}

} // namespace drt
