// Headless stand-in for the reference's App (src/App.cpp needs an SDL window: unavailable on this box).
// Builds the scene of App.cpp:101-122 through the reference's own Scene API and drives TWO backends through the
// unchanged render::PathTracer interface with the five calls App makes (create :98, set_settings :129,
// set_scene :130, render() :231, get_render_result() :232): the reference's CPUPathTracer (against the Embree
// stand-in) and CUDAPathTracer -- BOTH created through the registered factory
// render::PathTracer::create_path_tracer (PathTracer.cpp:11-21 with the GPU_CUDA case of
// host/reference_glue/register_gpu_cuda.patch), which is the one line App.cpp:98 changes.  Compares the RGBA8 images
// they return.  Test scaffolding (links oracle/_ref).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <memory>
#include <stdexcept>

#include "cudapt.h"
#include "engines/pathtracer/backends/cuda/CUDAPathTracer.h"  // only for the adapter's Options (static_cast below)
#include "render/Log.h"
#include "render/PathTracer.h"
#include "render/Scene.h"
#include "render/Types.h"

static std::shared_ptr<render::Scene> make_app_scene()
{
	auto scene = std::make_shared<render::Scene>();
	{
		auto s = scene->CreateNode<render::SphereObject>("123");
		s->SetRadius(1.0f);
		s->SetPosition(glm::vec3(0.0f, -1.0f, 5.0f));
	}
	{
		auto s = scene->CreateNode<render::SphereObject>("123");
		s->SetRadius(100.0f);
		s->SetPosition(glm::vec3(0.0f, -102.0f, 5.0f));
	}
	const int dims = 5;
	for (int x = -dims; x <= dims; x += 2)
		for (int y = -dims; y <= dims; y += 2)
		{
			auto s = scene->CreateNode<render::SphereObject>("sphere");
			s->SetRadius(0.5f);
			s->SetPosition(glm::vec3((float)x, (float)y, 10.0f));
		}
	return scene;
}

static double drive(render::PathTracer &pt, uint32_t w, uint32_t h, int frames, std::vector<uint32_t> &out)
{
	auto settings = std::make_shared<render::RenderSettings>();
	settings->setResolution(w, h);
	settings->setSamplesPerPixel(64);
	settings->setMaxBounces(8);
	pt.set_settings(settings);
	pt.set_scene(make_app_scene());
	const auto t0 = std::chrono::steady_clock::now();
	for (int f = 0; f < frames; ++f)
	{
		pt.render();
		const auto &res = pt.get_render_result();
		if (f == frames - 1)
			out = res.image_buffer;
	}
	return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

int main(int argc, char **argv)
{
	const uint32_t w = argc > 1 ? (uint32_t)std::atoi(argv[1]) : 640;
	const uint32_t h = argc > 2 ? (uint32_t)std::atoi(argv[2]) : 360;
	const int frames = argc > 3 ? std::atoi(argv[3]) : 16;
	render::Log::set_callback([](render::LogLevel, std::string_view) {});
	std::streambuf *old = std::cout.rdbuf(nullptr);  // the reference prints per node (Scene.h:171, CPUPathTracer.cpp:365-396)
	std::vector<uint32_t> cpu_img, gpu_img;
	auto cpu = render::PathTracer::create_path_tracer(render::PathTracer::BackendType::CPU_EMBREE);
	const double cpu_s = drive(*cpu, w, h, frames, cpu_img);
	auto gpu = render::PathTracer::create_path_tracer(render::PathTracer::BackendType::GPU_CUDA);
	const double gpu_s = drive(*gpu, w, h, frames, gpu_img);
	// resize + restart through the interface: change the resolution, render again
	std::vector<uint32_t> cpu2, gpu2;
	drive(*cpu, w / 2, h / 2, 2, cpu2);
	drive(*gpu, w / 2, h / 2, 2, gpu2);
	// adapter options: honour RenderSettings (8 bounces, roulette depth 3, exposure 2) and 4 samples per render() call;
	// the float accumulation is fingerprinted so that tests/test_gpu_adapter.py can compare it with the same render
	// made through the C ABI directly
	unsigned long long fnv = 1469598103934665603ull;
	uint32_t opt_frames = 0, opt_calls = 0;
	std::vector<uint32_t> opt_img;
	{
		auto gpu2p = render::PathTracer::create_path_tracer(render::PathTracer::BackendType::GPU_CUDA);
		auto *cu = static_cast<render::CUDAPathTracer *>(gpu2p.get());
		render::CUDAPathTracer::Options o;
		o.honour_render_settings = true;
		o.frames_per_render = 4;
		cu->set_options(o);
		auto settings = std::make_shared<render::RenderSettings>();
		settings->setResolution(w / 2, h / 2);
		settings->setMaxBounces(8);
		settings->setExposure(2.0f);
		cu->set_settings(settings);
		cu->set_scene(make_app_scene());
		for (int f = 0; f < 3; ++f, ++opt_calls) cu->render();
		opt_img = cu->get_render_result().image_buffer;
		opt_frames = cu->get_frame_count();
		std::vector<float> acc((size_t)(w / 2) * (h / 2) * 4);
		uint32_t fc = 0;
		cudapt_read_accum(cu->native_handle(), acc.data(), &fc);
		const unsigned char *b = reinterpret_cast<const unsigned char *>(acc.data());
		for (size_t i = 0; i < acc.size() * sizeof(float); ++i) fnv = (fnv ^ b[i]) * 1099511628211ull;
		unsigned long long fimg = 1469598103934665603ull;
		const unsigned char *bi = reinterpret_cast<const unsigned char *>(opt_img.data());
		for (size_t i = 0; i < opt_img.size() * 4; ++i) fimg = (fimg ^ bi[i]) * 1099511628211ull;
		std::cout.rdbuf(old);
		std::cout.clear();
		std::printf("options: %ux%u, %u render() calls -> %u frames (device %u), accum fnv %016llx, image fnv %016llx\n", w / 2, h / 2, opt_calls,
		            opt_frames, fc, fnv, fimg);
		old = std::cout.rdbuf(nullptr);
	}
	std::cout.rdbuf(old);
	std::cout.clear();
	size_t same = 0, same2 = 0;
	for (size_t i = 0; i < cpu_img.size(); ++i) same += cpu_img[i] == gpu_img[i];
	for (size_t i = 0; i < cpu2.size(); ++i) same2 += cpu2[i] == gpu2[i];
	std::printf("backend cpu: %s | gpu: %s\n", cpu->get_backend_name().c_str(), gpu->get_backend_name().c_str());
	bool unknown_throws = false;  // the factory's default branch is the reference's (PathTracer.cpp:19-20)
	try
	{
		render::PathTracer::create_path_tracer(render::PathTracer::BackendType::GPU_OPTIX);
	}
	catch (const std::runtime_error &)
	{
		unknown_throws = true;
	}
	std::printf("factory: create_path_tracer(CPU_EMBREE) -> type %d, create_path_tracer(GPU_CUDA) -> type %d, GPU_OPTIX throws: %s\n",
	            (int)cpu->get_backend_type(), (int)gpu->get_backend_type(), unknown_throws ? "yes" : "no");
	std::printf("%ux%u x %d frames: identical RGBA8 pixels %zu/%zu; after resize %zu/%zu\n", w, h, frames, same, cpu_img.size(), same2,
	            cpu2.size());
	std::printf("wall: cpu %.3f s (%.2f Msamples/s), gpu %.4f s (%.1f Msamples/s) incl. per-frame get_render_result()\n", cpu_s,
	            (double)w * h * frames / cpu_s / 1e6, gpu_s, (double)w * h * frames / gpu_s / 1e6);
	const bool ok = cpu_img.size() == (size_t)w * h && unknown_throws &&
	                cpu->get_backend_type() == render::PathTracer::BackendType::CPU_EMBREE &&
	                gpu->get_backend_type() == render::PathTracer::BackendType::GPU_CUDA && same >= cpu_img.size() - cpu_img.size() / 10000 &&
	                same2 >= cpu2.size() - cpu2.size() / 10000 - 1;
	std::printf(ok ? "OK\n" : "MISMATCH\n");
	return ok ? 0 : 1;
}
