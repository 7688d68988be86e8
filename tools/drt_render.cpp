// drt_render -- headless command-line front end over include/drt.hpp (the step after the path: replaces the editor's
// "save png", DustRayTracer/src/Editor/EditorLayer.cpp:23-31,85-97, and its hard-coded scene path, :50).
//   drt_render scene.glb [--width W] [--height H] [--spp N] [--bounces B] [--sun] [--batch K] [--device D]
//              [--pos x y z] [--dir x y z] [--out image.ppm|image.pfm]
// Prints one JSON line with build time, render time, rays and Mrays/s. Rows are written top-down (the frame
// buffer's row 0 is the BOTTOM row, RenderKernel.cu:22-25 + EditorLayer.cpp:294-295 flip).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "drt.hpp"

static int usage()
{
	fprintf(stderr, "usage: drt_render scene.glb [--width W] [--height H] [--spp N] [--bounces B] [--sun] [--batch K] [--device D]\n"
		"                  [--pos x y z] [--dir x y z] [--out image.ppm|image.pfm]\n");
	return 2;
}

int main(int argc, char** argv)
{
	if (argc < 2) return usage();
	std::string scene_path = argv[1], out;
	uint32_t W = 1280, H = 720, spp = 64, batch = 64;
	int device = 0;
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
		else if (a == "--sun") settings.enable_sunlight = 1;
		else if (a == "--out") { need(1); out = argv[++i]; }
		else if (a == "--pos") { need(3); for (int k = 0; k < 3; k++) cam.position[k] = (float)atof(argv[++i]); }
		else if (a == "--dir") { need(3); for (int k = 0; k < 3; k++) cam.forward_dir[k] = (float)atof(argv[++i]); }
		else return usage();
	}
	try {
		drt::Scene scene;
		if (!scene.loadGLTFmodel(scene_path.c_str())) throw drt::Error(DRT_ERR_IO, drt_last_error());
		drt::BVHBuilder builder;
		builder.device = device;
		builder.buildIterative(scene);
		drt::Renderer renderer(device);
		renderer.m_RendererSettings = settings;
		renderer.ResizeBuffer(W, H);
		double ms_total = 0.0;
		for (uint32_t done = 0; done < spp;) {
			const uint32_t n = spp - done < batch ? spp - done : batch;
			float ms = 0.f;
			renderer.Render(&cam, scene, &ms, n);
			ms_total += ms;
			done += n;
		}
		const drt_counters c = renderer.counters();
		const drt_scene_stats st = scene.stats();
		const double rays = (double)(c.closest_rays + c.shadow_rays);
		if (!out.empty()) {
			FILE* f = fopen(out.c_str(), "wb");
			if (!f) throw drt::Error(DRT_ERR_IO, "cannot open the output file");
			const bool pfm = out.size() > 4 && out.substr(out.size() - 4) == ".pfm";
			if (pfm) {
				std::vector<float> px = renderer.readbackRGBA32F();
				fprintf(f, "PF\n%u %u\n-1.0\n", W, H); // PFM stores rows bottom-up, like the frame buffer
				for (size_t p = 0; p < (size_t)W * H; p++) fwrite(&px[4 * p], sizeof(float), 3, f);
			}
			else {
				std::vector<uint8_t> px = renderer.readbackRGBA8();
				fprintf(f, "P6\n%u %u\n255\n", W, H);
				for (uint32_t y = 0; y < H; y++)
					for (uint32_t x = 0; x < W; x++) fwrite(&px[4 * ((size_t)(H - 1 - y) * W + x)], 1, 3, f);
			}
			fclose(f);
		}
		printf("{\"scene\": \"%s\", \"triangles\": %llu, \"wide_nodes\": %llu, \"build_ms\": %.3f, \"width\": %u, \"height\": %u, \"spp\": %u, "
			"\"render_ms\": %.3f, \"rays\": %.0f, \"Mrays_s\": %.1f, \"samples\": %u}\n",
			scene_path.c_str(), (unsigned long long)st.triangles, (unsigned long long)st.wide_nodes, st.build_ms, W, H, spp, ms_total, rays,
			ms_total > 0 ? rays / ms_total / 1e3 : 0.0, renderer.getSampleCount() - 1);
	}
	catch (const drt::Error& e) {
		fprintf(stderr, "drt_render: %s (status %d)\n", e.what(), e.status);
		return 1;
	}
	return 0;
}
