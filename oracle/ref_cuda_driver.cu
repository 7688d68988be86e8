// oracle/ref_cuda_driver.cu -- MEASUREMENT INFRASTRUCTURE ONLY (never linked into the product library).
//
// The reference's own CUDA path on the GPU box: its UNMODIFIED sources are compiled where they lie under
// /root/reference with nvcc for sm_100a (oracle/Makefile, target `refcuda`, relocatable device code like
// premake5.lua:52) and driven headless by this file:
//   Scene::loadGLTFmodel              DustRayTracer/src/Core/Scene/Scene.cu:181-317
//   BVHBuilder::build                 DustRayTracer/src/Core/BVH/BVHBuilder.cu:100-145
//   RayGen (device)                   DustRayTracer/src/Core/Kernel/Shaders/RayGen.cuh:63-172
// The reference's __global__ `kernel` (RenderKernel.cu:20-35) and InvokeRenderKernel (:37-58) include
// GL interop headers that do not exist on the box, so the 10 lines of that kernel are restated below with
// the surf2Dwrite replaced by a store to a linear float4 buffer; everything it calls is the reference's code.
// Launch shape (8x8 blocks, W/8+1 x H/8+1 grid, one launch + event pair per sample, host sync per sample)
// follows Renderer.hpp:43-44 and Renderer.cu:38-39,97-108.
#include "Core/Scene/Scene.cuh"
#include "Core/Scene/Camera.cuh"
#include "Core/BVH/BVHNode.cuh"
#include "Core/BVH/BVHBuilder.cuh"
#include "Core/Ray.cuh"
#include "Core/HitPayload.cuh"
#include "Core/CudaMath/helper_math.cuh"
#include "Core/CudaMath/Random.cuh"
#include "Core/Kernel/Shaders/RayGen.cuh"

#include <cuda_runtime.h>
#include <thrust/device_vector.h>
#include <cstring>
#include <vector>

__global__ void refcuda_kernel(float4* display, int max_x, int max_y, Camera* cam, uint32_t frameidx, float3* accumulation_buffer,
	const SceneData scenedata)
{
	int i = threadIdx.x + blockIdx.x * blockDim.x;
	int j = threadIdx.y + blockIdx.y * blockDim.y;
	if ((i >= max_x) || (j >= max_y)) return;
	float3 fcolor = RayGen(i, j, max_x, max_y, cam, frameidx, scenedata);
	accumulation_buffer[i + j * max_x] += fcolor;
	float3 accol = accumulation_buffer[i + j * max_x] / frameidx;
	display[i + j * max_x] = make_float4(accol.x, accol.y, accol.z, 1);
}

namespace {
struct RefCuda {
	Scene* scene = nullptr;
	Camera* cam = nullptr;
	thrust::device_vector<float3>* acc = nullptr;
	float4* display = nullptr;
	int W = 0, H = 0;
};
}

extern "C" {

void* refcuda_load(const char* path, int bin_count, int leaf_target)
{
	RefCuda* r = new RefCuda();
	r->scene = new Scene();
	r->scene->loadGLTFmodel(path);
	BVHBuilder builder;
	builder.m_BinCount = bin_count;
	builder.m_TargetLeafPrimitivesCount = leaf_target;
	r->scene->d_BVHTreeRoot = builder.build(r->scene->m_PrimitivesBuffer, r->scene->m_BVHNodes);
	r->cam = new Camera(); // Managed::operator new -> cudaMallocManaged (Editor/Common/Managed.cuh:6-11)
	return r;
}

int refcuda_node_count(void* h) { return (int)((RefCuda*)h)->scene->m_BVHNodes.size(); }
int refcuda_tri_count(void* h) { return (int)((RefCuda*)h)->scene->m_PrimitivesBuffer.size(); }

// Renders frames frame_begin .. frame_begin+frames-1 into a zeroed accumulation buffer, one launch per frame.
// Returns the sum of the per-launch event times in ms (what Renderer::Render reports per frame), or < 0.
// acc_out (optional): W*H*3 floats.
double refcuda_render(void* h, const void* cam_pod, const void* settings_pod, int W, int H, unsigned frame_begin, int frames,
	float* acc_out)
{
	RefCuda* r = (RefCuda*)h;
	static_assert(sizeof(Camera) == 88, "Camera POD size");
	memcpy((void*)r->cam, cam_pod, sizeof(Camera)); // managed memory: host write, like EditorLayer's camera controls
	RendererSettings settings;
	memcpy(&settings, settings_pod, sizeof(RendererSettings));
	if (W != r->W || H != r->H) {
		delete r->acc;
		if (r->display) cudaFree(r->display);
		r->acc = new thrust::device_vector<float3>((size_t)W * H);
		if (cudaMalloc(&r->display, sizeof(float4) * (size_t)W * H) != cudaSuccess) return -1.0;
		r->W = W; r->H = H;
	}
	thrust::fill(r->acc->begin(), r->acc->end(), float3()); // resetAccumulationBuffer, Renderer.cu:132-136
	const Scene& scene = *r->scene;
	SceneData sd; // InvokeRenderKernel's packing, RenderKernel.cu:41-51
	sd.DeviceBVHNodesBuffer = thrust::raw_pointer_cast(scene.m_BVHNodes.data());
	sd.DeviceTextureBufferPtr = thrust::raw_pointer_cast(scene.m_Textures.data());
	sd.DeviceMeshBufferPtr = thrust::raw_pointer_cast(scene.m_Meshes.data());
	sd.DeviceMaterialBufferPtr = thrust::raw_pointer_cast(scene.m_Material.data());
	sd.DevicePrimitivesBuffer = thrust::raw_pointer_cast(scene.m_PrimitivesBuffer.data());
	sd.DeviceMeshBufferSize = scene.m_Meshes.size();
	sd.DevicePrimitivesBufferSize = scene.m_PrimitivesBuffer.size();
	sd.DeviceBVHNodesBufferSize = scene.m_BVHNodes.size();
	sd.DeviceBVHTreeRootPtr = scene.d_BVHTreeRoot;
	sd.RenderSettings = settings;
	const dim3 threads(8, 8), blocks(W / 8 + 1, H / 8 + 1); // Renderer.hpp:43-44, Renderer.cu:38-39
	cudaEvent_t start, stop;
	cudaEventCreate(&start); cudaEventCreate(&stop);
	double total = 0.0;
	for (int f = 0; f < frames; f++) {
		cudaEventRecord(start);
		refcuda_kernel<<<blocks, threads>>>(r->display, W, H, r->cam, frame_begin + (unsigned)f, thrust::raw_pointer_cast(r->acc->data()), sd);
		cudaEventRecord(stop);
		if (cudaEventSynchronize(stop) != cudaSuccess) { total = -1.0; break; }
		float ms = 0.f;
		cudaEventElapsedTime(&ms, start, stop);
		total += ms;
	}
	cudaEventDestroy(start); cudaEventDestroy(stop);
	if (cudaGetLastError() != cudaSuccess) total = -1.0;
	if (acc_out && total >= 0.0)
		cudaMemcpy(acc_out, thrust::raw_pointer_cast(r->acc->data()), sizeof(float3) * (size_t)W * H, cudaMemcpyDeviceToHost);
	return total;
}

void refcuda_free(void* h)
{
	RefCuda* r = (RefCuda*)h;
	delete r->acc;
	if (r->display) cudaFree(r->display);
	delete r->scene;
	delete r;
}

}
