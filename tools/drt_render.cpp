// drt_render -- headless command-line front end over include/drt.hpp (the step after the path: replaces the editor's
// "save png", DustRayTracer/src/Editor/EditorLayer.cpp:23-31,85-97, and its hard-coded scene path, :50).
//   drt_render scene.glb [--width W] [--height H] [--spp N] [--bounces B] [--sun] [--batch K] [--device D]
//              [--pos x y z] [--dir x y z] [--robust] [--out image.ppm|image.pfm|image.png]
// --robust loads the file as the glTF specification says (DRT_LOAD_ROBUST) instead of as the reference loader does.
// Prints one JSON line with build time, render time, rays and Mrays/s. Rows are written top-down (the frame
// buffer's row 0 is the BOTTOM row, RenderKernel.cu:22-25 + EditorLayer.cpp:294-295 flip).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "drt.hpp"

// Minimal PNG writer (8-bit RGB, zlib "stored" blocks: no compression library needed); replaces stbi_write_png (EditorLayer.cpp:93-95).
static void png_chunk(FILE* f, const char* type, const std::vector<uint8_t>& data)
{
	static uint32_t table[256];
	if (!table[1])
		for (uint32_t n = 0; n < 256; n++) {
			uint32_t c = n;
			for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
			table[n] = c;
		}
	auto be32 = [&](uint32_t v) { const uint8_t b[4] = { (uint8_t)(v >> 24), (uint8_t)(v >> 16), (uint8_t)(v >> 8), (uint8_t)v }; fwrite(b, 1, 4, f); };
	uint32_t crc = 0xffffffffu;
	for (int i = 0; i < 4; i++) crc = table[(crc ^ (uint8_t)type[i]) & 0xff] ^ (crc >> 8);
	for (uint8_t d : data) crc = table[(crc ^ d) & 0xff] ^ (crc >> 8);
	be32((uint32_t)data.size());
	fwrite(type, 1, 4, f);
	if (!data.empty()) fwrite(data.data(), 1, data.size(), f);
	be32(crc ^ 0xffffffffu);
}

// rows: top-down RGB, 3 * w bytes each
static void write_png(FILE* f, uint32_t w, uint32_t h, const std::vector<uint8_t>& rgb)
{
	const uint8_t sig[8] = { 0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a };
	fwrite(sig, 1, 8, f);
	std::vector<uint8_t> ihdr = { (uint8_t)(w >> 24), (uint8_t)(w >> 16), (uint8_t)(w >> 8), (uint8_t)w, (uint8_t)(h >> 24), (uint8_t)(h >> 16), (uint8_t)(h >> 8), (uint8_t)h, 8, 2, 0, 0, 0 };
	png_chunk(f, "IHDR", ihdr);
	std::vector<uint8_t> raw;
	raw.reserve((size_t)h * (3 * w + 1));
	for (uint32_t y = 0; y < h; y++) { raw.push_back(0); raw.insert(raw.end(), rgb.begin() + (size_t)y * 3 * w, rgb.begin() + (size_t)(y + 1) * 3 * w); }
	std::vector<uint8_t> z = { 0x78, 0x01 };
	uint32_t a = 1, b = 0;
	for (size_t p = 0; p < raw.size() || p == 0;) {
		const size_t n = raw.size() - p < 65535 ? raw.size() - p : 65535;
		const bool last = p + n >= raw.size();
		z.push_back(last ? 1 : 0);
		z.push_back((uint8_t)n); z.push_back((uint8_t)(n >> 8)); z.push_back((uint8_t)~n); z.push_back((uint8_t)(~n >> 8));
		for (size_t i = 0; i < n; i++) { a = (a + raw[p + i]) % 65521u; b = (b + a) % 65521u; }
		z.insert(z.end(), raw.begin() + p, raw.begin() + p + n);
		p += n;
		if (last) break;
	}
	const uint32_t adler = (b << 16) | a;
	z.push_back((uint8_t)(adler >> 24)); z.push_back((uint8_t)(adler >> 16)); z.push_back((uint8_t)(adler >> 8)); z.push_back((uint8_t)adler);
	png_chunk(f, "IDAT", z);
	png_chunk(f, "IEND", {});
}

static int usage()
{
	fprintf(stderr, "usage: drt_render scene.glb [--width W] [--height H] [--spp N] [--bounces B] [--sun] [--batch K] [--device D] [--gpus N [--tiles]]\n"
		"                  [--pos x y z] [--dir x y z] [--robust] [--out image.ppm|image.pfm|image.png]\n");
	return 2;
}

int main(int argc, char** argv)
{
	if (argc < 2) return usage();
	std::string scene_path = argv[1], out;
	uint32_t W = 1280, H = 720, spp = 64, batch = 64;
	int device = 0, gpus = 1;
	bool tiles = false;
	uint32_t load_flags = DRT_LOAD_REFERENCE_COMPAT;
	drt::Camera cam;
	drt::RendererSettings settings;
	settings.max_samples = 1 << 30;
	for (int i = 2; i < argc; i++) {
		std::string a = argv[i];
		auto need = [&](int n) { if (i + n >= argc) { usage(); exit(2); } };
		if (a == "--width") { need(1); W = (uint32_t)atoi(argv[++i]); }
		else if (a == "--height") { need(1); H = (uint32_t)atoi(argv[++i]); }
		else if (a == "--spp") { need(1); spp = (uint32_t)atoi(argv[++i]); }
		else if (a == "--batch") { need(1); batch = (uint32_t)atoi(argv[++i]); }
		else if (a == "--bounces") { need(1); settings.ray_bounce_limit = atoi(argv[++i]); }
		else if (a == "--device") { need(1); device = atoi(argv[++i]); }
		else if (a == "--gpus") { need(1); gpus = atoi(argv[++i]); }
		else if (a == "--tiles") tiles = true;
		else if (a == "--sun") settings.enable_sunlight = 1;
		else if (a == "--robust") load_flags = DRT_LOAD_ROBUST;
		else if (a == "--out") { need(1); out = argv[++i]; }
		else if (a == "--pos") { need(3); for (int k = 0; k < 3; k++) cam.position[k] = (float)atof(argv[++i]); }
		else if (a == "--dir") { need(3); for (int k = 0; k < 3; k++) cam.forward_dir[k] = (float)atof(argv[++i]); }
		else return usage();
	}
	try {
		drt::Scene scene;
		if (!scene.loadGLTFmodel(scene_path.c_str(), load_flags)) throw drt::Error(DRT_ERR_IO, drt_last_error());
		double ms_total = 0.0;
		drt_counters c;
		uint32_t samples = 0;
		std::vector<float> px32;
		std::vector<uint8_t> px8;
		const bool pfm = out.size() > 4 && out.substr(out.size() - 4) == ".pfm";
		if (gpus > 1) {
			// N GPUs of this node: drt_multi (replicated scene, one NCCL reduce of the accumulation buffers per batch)
			std::vector<int32_t> devs;
			for (int g = 0; g < gpus; g++) devs.push_back(device + g);
			drt::MultiRenderer renderer(devs);
			renderer.m_RendererSettings = settings;
			renderer.setScene(scene);
			if (tiles) renderer.setSharding(true);
			renderer.ResizeBuffer(W, H);
			for (uint32_t done = 0; done < spp;) {
				const uint32_t n = spp - done < batch ? spp - done : batch;
				float ms = 0.f;
				renderer.Render(&cam, &ms, n);
				ms_total += ms;
				done += n;
			}
			c = renderer.counters();
			samples = renderer.getSampleCount() - 1;
			if (!out.empty()) { if (pfm) px32 = renderer.readbackRGBA32F(); else px8 = renderer.readbackRGBA8(); }
			drt::BVHBuilder builder; // (stats of one replica for the report)
			builder.device = device;
			builder.buildIterative(scene);
		}
		else {
			drt::BVHBuilder builder;
			builder.device = device;
			builder.buildIterative(scene);
			drt::Renderer renderer(device);
			renderer.m_RendererSettings = settings;
			renderer.ResizeBuffer(W, H);
			for (uint32_t done = 0; done < spp;) {
				const uint32_t n = spp - done < batch ? spp - done : batch;
				float ms = 0.f;
				renderer.Render(&cam, scene, &ms, n);
				ms_total += ms;
				done += n;
			}
			c = renderer.counters();
			samples = renderer.getSampleCount() - 1;
			if (!out.empty()) { if (pfm) px32 = renderer.readbackRGBA32F(); else px8 = renderer.readbackRGBA8(); }
		}
		const drt_scene_stats st = scene.stats();
		const double rays = (double)(c.closest_rays + c.shadow_rays);
		if (!out.empty()) {
			FILE* f = fopen(out.c_str(), "wb");
			if (!f) throw drt::Error(DRT_ERR_IO, "cannot open the output file");
			if (pfm) {
				const std::vector<float>& px = px32;
				fprintf(f, "PF\n%u %u\n-1.0\n", W, H); // PFM stores rows bottom-up, like the frame buffer
				for (size_t p = 0; p < (size_t)W * H; p++) fwrite(&px[4 * p], sizeof(float), 3, f);
			}
			else {
				const std::vector<uint8_t>& px = px8;
				std::vector<uint8_t> rgb((size_t)W * H * 3);
				for (uint32_t y = 0; y < H; y++)
					for (uint32_t x = 0; x < W; x++) memcpy(&rgb[3 * ((size_t)y * W + x)], &px[4 * ((size_t)(H - 1 - y) * W + x)], 3);
				if (out.size() > 4 && out.substr(out.size() - 4) == ".png") write_png(f, W, H, rgb);
				else {
					fprintf(f, "P6\n%u %u\n255\n", W, H);
					fwrite(rgb.data(), 1, rgb.size(), f);
				}
			}
			fclose(f);
		}
		printf("{\"scene\": \"%s\", \"triangles\": %llu, \"wide_nodes\": %llu, \"build_ms\": %.3f, \"width\": %u, \"height\": %u, \"spp\": %u, "
			"\"render_ms\": %.3f, \"rays\": %.0f, \"Mrays_s\": %.1f, \"samples\": %u, \"gpus\": %d}\n",
			scene_path.c_str(), (unsigned long long)st.triangles, (unsigned long long)st.wide_nodes, st.build_ms, W, H, spp, ms_total, rays,
			ms_total > 0 ? rays / ms_total / 1e3 : 0.0, samples, gpus);
	}
	catch (const drt::Error& e) {
		fprintf(stderr, "drt_render: %s (status %d)\n", e.what(), e.status);
		return 1;
	}
	return 0;
}
