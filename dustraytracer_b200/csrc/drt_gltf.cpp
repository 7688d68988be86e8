// drt_gltf.cpp -- glTF 2.0 / GLB scene load to a triangle soup, with the reference loader's semantics.
//
// Replaces Scene::loadGLTFmodel and helpers (DustRayTracer/src/Core/Scene/Scene.cu:22-317), which sit on
// tinygltf v2.8.10; this is an independent reader of the container (GLB chunks, JSON, buffers, PNG images).
// What must be reproduced for "identical scenes" (SURVEY.md appendix A.6):
//   * every entry of `nodes` is visited in order and its mesh emitted (no hierarchy, no transforms,
//     a mesh referenced twice is emitted twice)                                   Scene.cu:192-207
//   * per primitive: POSITION / NORMAL / TEXCOORD_0 accessors, a missing one silently becomes
//     accessor 0 (std::map::operator[])                                           Scene.cu:129-131
//   * data pointers = the INDEX buffer's bytes + each attribute's bufferView.byteOffset (accessor
//     byteOffset / byteStride ignored); indices are u16 and run over the index bufferView's byte range
//                                                                                 Scene.cu:150-171
//   * one material index per triangle = primitive.material                        Scene.cu:172-175
//   * Material.Albedo = baseColorFactor.rgb, AlbedoTextureIndex = baseColorTexture.index used as an
//     IMAGE index                                                                 Scene.cu:59-86,93-114
//   * face normal oriented to the mean vertex normal, then normalised             Scene.cu:275-300
#include "drt_host.h"
#include "drt_json.h"
#include "drt_math.cuh"

#include <cstdio>
#include <cstring>

namespace drt {

void compute_face_normal(const float* p, const float* n, float* out3)
{
	typedef M<true> m; // host build: separately rounded fp32, the reference's operation order
	V3 p0 = mk3(p[0], p[1], p[2]), p1 = mk3(p[3], p[4], p[5]), p2 = mk3(p[6], p[7], p[8]);
	V3 e0 = m::sub3(p1, p0), e1 = m::sub3(p2, p0);
	V3 fn = m::cross(e0, e1);
	V3 avg = m::divs(m::add3(m::add3(mk3(n[0], n[1], n[2]), mk3(n[3], n[4], n[5])), mk3(n[6], n[7], n[8])), 3.f);
	float ndot = m::dot(fn, avg);
	V3 sn = (ndot < 0.0f) ? mk3(-fn.x, -fn.y, -fn.z) : fn;
	V3 r = m::normalize(sn);
	out3[0] = r.x; out3[1] = r.y; out3[2] = r.z;
}

namespace {

bool read_file(const std::string& path, std::vector<uint8_t>& out)
{
	FILE* f = fopen(path.c_str(), "rb");
	if (!f) return false;
	fseek(f, 0, SEEK_END);
	long n = ftell(f);
	fseek(f, 0, SEEK_SET);
	if (n < 0) { fclose(f); return false; }
	out.resize((size_t)n);
	size_t got = n ? fread(out.data(), 1, (size_t)n, f) : 0;
	fclose(f);
	return got == (size_t)n;
}

bool base64_decode(const std::string& s, size_t start, std::vector<uint8_t>& out)
{
	out.clear();
	uint32_t acc = 0; int bits = 0;
	for (size_t i = start; i < s.size(); i++) {
		char c = s[i];
		int v;
		if (c >= 'A' && c <= 'Z') v = c - 'A';
		else if (c >= 'a' && c <= 'z') v = c - 'a' + 26;
		else if (c >= '0' && c <= '9') v = c - '0' + 52;
		else if (c == '+' || c == '-') v = 62;
		else if (c == '/' || c == '_') v = 63;
		else if (c == '=' || c == '\n' || c == '\r') continue;
		else return false;
		acc = (acc << 6) | (uint32_t)v; bits += 6;
		if (bits >= 8) { bits -= 8; out.push_back((uint8_t)((acc >> bits) & 0xff)); }
	}
	return true;
}

bool load_uri(const std::string& base_dir, const std::string& uri, std::vector<uint8_t>& out)
{
	if (uri.compare(0, 5, "data:") == 0) {
		size_t comma = uri.find(',');
		if (comma == std::string::npos) return false;
		return base64_decode(uri, comma + 1, out);
	}
	std::string decoded; // percent-decoding of plain file URIs
	for (size_t i = 0; i < uri.size(); i++) {
		if (uri[i] == '%' && i + 2 < uri.size()) { decoded += (char)strtol(uri.substr(i + 1, 2).c_str(), nullptr, 16); i += 2; }
		else decoded += uri[i];
	}
	return read_file(base_dir + decoded, out);
}

struct View { int buffer; size_t offset, length, stride; };

} // namespace

// drt_gltf_robust.cpp: spec-conforming geometry walk (scene graph, transforms, accessors); SURVEY.md 8(f) rank 1
int emit_robust_geometry(const JsonValue& root, const std::vector<const uint8_t*>& bdata, const std::vector<size_t>& blen,
	HostScene& out, std::string& err);
int remap_robust_materials(const JsonValue& root, const std::vector<bool>& image_ok, HostScene& out, std::string& err);

int load_gltf(const char* path, HostScene& out, std::string& err, unsigned flags)
{
	const bool robust = (flags & DRT_LOAD_ROBUST) != 0;
	std::vector<uint8_t> file;
	if (!read_file(path, file)) { err = std::string("cannot read ") + path; return DRT_ERR_IO; }
	std::string spath(path);
	size_t slash = spath.find_last_of("/\\");
	const std::string base_dir = slash == std::string::npos ? std::string() : spath.substr(0, slash + 1);
	// the reference decides by extension (Scene.cu:27-41); the GLB magic is checked as well
	const bool is_glb = file.size() >= 12 && memcmp(file.data(), "glTF", 4) == 0;

	const char* json_ptr = (const char*)file.data();
	size_t json_len = file.size();
	const uint8_t* glb_bin = nullptr; size_t glb_bin_len = 0;
	if (is_glb) {
		uint32_t version, total;
		memcpy(&version, file.data() + 4, 4); memcpy(&total, file.data() + 8, 4);
		if (version != 2) { err = "unsupported GLB version"; return DRT_ERR_FORMAT; }
		size_t p = 12; bool have_json = false;
		while (p + 8 <= file.size()) {
			uint32_t clen, ctype;
			memcpy(&clen, file.data() + p, 4); memcpy(&ctype, file.data() + p + 4, 4);
			if (p + 8 + (size_t)clen > file.size()) { err = "truncated GLB chunk"; return DRT_ERR_FORMAT; }
			if (ctype == 0x4E4F534Au && !have_json) { json_ptr = (const char*)file.data() + p + 8; json_len = clen; have_json = true; }
			else if (ctype == 0x004E4942u && !glb_bin) { glb_bin = file.data() + p + 8; glb_bin_len = clen; }
			p += 8 + (size_t)clen; p = (p + 3) & ~(size_t)3;
		}
		if (!have_json) { err = "GLB without JSON chunk"; return DRT_ERR_FORMAT; }
	}
	JsonValue root;
	{
		JsonParser jp(json_ptr, json_len);
		if (!jp.parse(root, err)) return DRT_ERR_FORMAT;
	}
	if (root.type != JsonValue::Object) { err = "glTF root is not an object"; return DRT_ERR_FORMAT; }
	static const JsonValue empty_array = [] { JsonValue v; v.type = JsonValue::Array; return v; }();
	auto arr = [&](const char* key) -> const JsonValue& { const JsonValue* v = root.get(key); return (v && v->type == JsonValue::Array) ? *v : empty_array; };

	// ---- buffers ----
	const JsonValue& jbuffers = arr("buffers");
	std::vector<std::vector<uint8_t>> owned(jbuffers.size());
	std::vector<const uint8_t*> bdata(jbuffers.size(), nullptr);
	std::vector<size_t> blen(jbuffers.size(), 0);
	for (size_t i = 0; i < jbuffers.size(); i++) {
		const JsonValue& b = jbuffers.arr[i];
		if (b.has("uri")) {
			if (!load_uri(base_dir, b.get_str("uri"), owned[i])) { err = "cannot load buffer " + b.get_str("uri"); return DRT_ERR_IO; }
			bdata[i] = owned[i].data(); blen[i] = owned[i].size();
		}
		else if (is_glb && glb_bin) { bdata[i] = glb_bin; blen[i] = glb_bin_len; }
		else { err = "buffer without data"; return DRT_ERR_FORMAT; }
	}
	const JsonValue& jviews = arr("bufferViews");
	std::vector<View> views(jviews.size());
	for (size_t i = 0; i < jviews.size(); i++) {
		const JsonValue& v = jviews.arr[i];
		views[i].buffer = v.get_int("buffer", 0);
		if (!v.get_size("byteOffset", 0, &views[i].offset) || !v.get_size("byteLength", 0, &views[i].length) || !v.get_size("byteStride", 0, &views[i].stride)) {
			err = "bufferView with a negative, fractional or huge byte range"; return DRT_ERR_FORMAT;
		}
		// (offset <= size first, then length against what is left: the sum could wrap)
		if (views[i].buffer < 0 || (size_t)views[i].buffer >= bdata.size() || views[i].offset > blen[views[i].buffer]
			|| views[i].length > blen[views[i].buffer] - views[i].offset || views[i].stride > 65536) {
			err = "bufferView out of range"; return DRT_ERR_FORMAT;
		}
	}
	const JsonValue& jacc = arr("accessors");
	auto accessor_view = [&](int a) -> int { return (a >= 0 && (size_t)a < jacc.size()) ? jacc.arr[a].get_int("bufferView", -1) : -1; };

	// ---- images -> textures (Scene.cu:88-117: one Texture per IMAGE, in image order) ----
	const JsonValue& jimages = arr("images");
	std::vector<bool> image_ok(jimages.size(), true);
	for (size_t i = 0; i < jimages.size(); i++) {
		const JsonValue& img = jimages.arr[i];
		HostTexture tex;
		std::string perr;
		int bv = img.get_int("bufferView", -1);
		bool ok = false;
		if (bv >= 0 && (size_t)bv < views.size()) ok = decode_image(bdata[views[bv].buffer] + views[bv].offset, views[bv].length, tex, perr);
		else if (img.has("uri")) {
			std::vector<uint8_t> bytes;
			if (load_uri(base_dir, img.get_str("uri"), bytes)) ok = decode_image(bytes.data(), bytes.size(), tex, perr);
			else perr = "cannot load image " + img.get_str("uri");
		}
		else perr = "image without bufferView/uri";
		if (!ok && robust) { // undecodable image: keep the slot so image indices stay stable, detach the materials that use it
			out.warnings.push_back("image " + std::to_string(i) + " not decoded (" + perr + "); materials using it fall back to their factor");
			image_ok[i] = false;
			tex = HostTexture();
			ok = true;
		}
		if (!ok) { err = "image " + std::to_string(i) + ": " + perr; return DRT_ERR_FORMAT; }
		out.textures.push_back(std::move(tex));
	}

	// ---- materials (Scene.cu:59-86) ----
	const JsonValue& jmats = arr("materials");
	for (size_t i = 0; i < jmats.size(); i++) {
		const JsonValue& jm = jmats.arr[i];
		drt_material m;
		memset(&m, 0, sizeof(m));
		double base[4] = { 1, 1, 1, 1 }, emis[3] = { 0, 0, 0 }, metallic = 1.0, rough = 1.0; // tinygltf defaults
		int tex = -1;
		if (const JsonValue* pbr = jm.get("pbrMetallicRoughness")) {
			if (const JsonValue* f = pbr->get("baseColorFactor"))
				for (size_t k = 0; k < 4 && k < f->size(); k++) base[k] = f->arr[k].num;
			metallic = pbr->get_num("metallicFactor", 1.0);
			rough = pbr->get_num("roughnessFactor", 1.0);
			if (const JsonValue* t = pbr->get("baseColorTexture")) tex = t->get_int("index", -1);
		}
		if (const JsonValue* e = jm.get("emissiveFactor"))
			for (size_t k = 0; k < 3 && k < e->size(); k++) emis[k] = e->arr[k].num;
		for (int k = 0; k < 3; k++) { m.albedo[k] = (float)base[k]; m.emissive_factor[k] = (float)emis[k]; }
		m.albedo_texture_index = tex;
		m.roughness = (float)rough;
		m.transmission = 0;
		m.refractive_index = 1.45f; // Material.cuh:21
		m.metallic = metallic > 0 ? 1 : 0;
		out.materials.push_back(m);
	}

	if (robust) {
		int rc = remap_robust_materials(root, image_ok, out, err);
		if (rc != DRT_OK) return rc;
		return emit_robust_geometry(root, bdata, blen, out, err);
	}

	// ---- nodes -> meshes -> primitives (Scene.cu:120-178,192-314) ----
	const JsonValue& jnodes = arr("nodes");
	const JsonValue& jmeshes = arr("meshes");
	for (size_t ni = 0; ni < jnodes.size(); ni++) {
		int mi = jnodes.arr[ni].get_int("mesh", -1);
		if (mi < 0 || (size_t)mi >= jmeshes.size()) {
			out.warnings.push_back("node " + std::to_string(ni) + " has no mesh (the reference reads out of bounds here); skipped");
			continue;
		}
		const JsonValue* prims = jmeshes.arr[mi].get("primitives");
		out.meshes++;
		if (!prims) continue;
		for (size_t pi = 0; pi < prims->size(); pi++) {
			const JsonValue& prim = prims->arr[pi];
			const JsonValue* attrs = prim.get("attributes");
			int a_pos = attrs ? attrs->get_int("POSITION", 0) : 0; // missing attribute -> accessor 0
			int a_nrm = attrs ? attrs->get_int("NORMAL", 0) : 0;
			int a_uv = attrs ? attrs->get_int("TEXCOORD_0", 0) : 0;
			int a_idx = prim.get_int("indices", -1);
			int material = prim.get_int("material", -1);
			int v_pos = accessor_view(a_pos), v_nrm = accessor_view(a_nrm), v_uv = accessor_view(a_uv), v_idx = accessor_view(a_idx);
			if (v_idx < 0 || v_pos < 0 || v_nrm < 0 || v_uv < 0 || (size_t)v_idx >= views.size() || (size_t)v_pos >= views.size()
				|| (size_t)v_nrm >= views.size() || (size_t)v_uv >= views.size()) {
				err = "primitive without u16 indices / attribute bufferViews (unsupported by the reference loader)";
				return DRT_ERR_FORMAT;
			}
			const int comp = jacc.arr[a_idx].get_int("componentType", 5123);
			if (comp != 5123) { err = "index componentType " + std::to_string(comp) + ": the reference loader reads unsigned short only (Scene.cu:161)"; return DRT_ERR_FORMAT; }
			size_t idx_count_accessor = 0;
			if (!jacc.arr[a_idx].get_size("count", 0, &idx_count_accessor)) { err = "index accessor with a bad count"; return DRT_ERR_FORMAT; }
			const View& iv = views[v_idx];
			const uint8_t* buf = bdata[iv.buffer];
			const size_t buf_len = blen[iv.buffer];
			const size_t i_begin = iv.offset / 2, i_end = (iv.length + iv.offset) / 2;
			const size_t o_pos = views[v_pos].offset, o_nrm = views[v_nrm].offset, o_uv = views[v_uv].offset;
			// the reference reads every attribute from the INDEX accessor's buffer (Scene.cu:150-170); a view of another, longer
			// buffer can start beyond this one's end, and `buf_len - o_*` below would wrap
			if (o_pos > buf_len || o_nrm > buf_len || o_uv > buf_len) { err = "attribute bufferView starts beyond the index accessor's buffer"; return DRT_ERR_FORMAT; }
			const size_t n_idx = i_end - i_begin;
			size_t n_tri = n_idx / 3;
			if (idx_count_accessor / 3 < n_tri) n_tri = idx_count_accessor / 3; // keep material/triangle arrays consistent
			for (size_t t = 0; t < n_tri; t++) {
				float p9[9], n9[9], t6[6];
				for (int k = 0; k < 3; k++) {
					uint16_t vi;
					memcpy(&vi, buf + 2 * (i_begin + 3 * t + k), 2);
					// (o_* <= buf_len was checked with the bufferViews; vi < 65536, so the products cannot wrap)
					if (12 * ((size_t)vi + 1) > buf_len - o_pos || 12 * ((size_t)vi + 1) > buf_len - o_nrm || 8 * ((size_t)vi + 1) > buf_len - o_uv) {
						err = "vertex index out of buffer range"; return DRT_ERR_FORMAT;
					}
					memcpy(p9 + 3 * k, buf + o_pos + 12 * (size_t)vi, 12);
					memcpy(n9 + 3 * k, buf + o_nrm + 12 * (size_t)vi, 12);
					memcpy(t6 + 2 * k, buf + o_uv + 8 * (size_t)vi, 8);
				}
				float fn[3];
				compute_face_normal(p9, n9, fn);
				out.pos.insert(out.pos.end(), p9, p9 + 9);
				out.vnormal.insert(out.vnormal.end(), n9, n9 + 9);
				out.uv.insert(out.uv.end(), t6, t6 + 6);
				out.fnormal.insert(out.fnormal.end(), fn, fn + 3);
				out.mat.push_back(material);
			}
		}
	}
	return DRT_OK;
}

} // namespace drt
