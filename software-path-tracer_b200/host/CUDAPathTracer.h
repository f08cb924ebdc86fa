// render::PathTracer backend that forwards to the C ABI of libcudapt.so (include/cudapt.h).
//
// Drop-in sibling of the reference's CPUPathTracer
// (libs/render/src/engines/pathtracer/backends/cpu/CPUPathTracer.h:21-75): same eight overrides, same
// invalidation rules (CPUPathTracer.cpp:119-161), same fail-fast error convention (render::verify).  It is
// compiled inside the reference tree (it includes the reference's own public headers) at
// libs/render/src/engines/pathtracer/backends/cuda/, after the factory registration patch
// (host/reference_glue/register_gpu_cuda.patch) has been applied.
#pragma once

#include <memory>
#include <string>

#include "render/PathTracer.h"

struct cudapt_ctx;

namespace render
{
	class CUDAPathTracer : public PathTracer
	{
	public:
		CUDAPathTracer();
		~CUDAPathTracer() override;

		void render() override;

		void set_scene(std::shared_ptr<Scene> scene) override { m_scene = scene; }
		void set_settings(std::shared_ptr<RenderSettings> settings) override { m_renderSettings = settings; }

		std::shared_ptr<Scene> get_scene() const override { return m_scene; }
		std::shared_ptr<RenderSettings> get_settings() const override { return m_renderSettings; }

		std::string get_backend_name() const override { return "CUDA Path Tracer (sm_100a)"; }
		// PathTracer::BackendType has no CUDA enumerator upstream (PathTracer.h:16-21): the registration patch
		// (host/reference_glue/register_gpu_cuda.patch) appends GPU_CUDA and the factory case.
		BackendType get_backend_type() const override { return BackendType::GPU_CUDA; }

		const PathTracer::RenderResult &get_render_result() override;

		// Extensions (not part of render::PathTracer): float accumulation read-back and counters.
		cudapt_ctx *native_handle() const { return m_ctx; }

		// Opt-in behaviour beyond the reference's.  The defaults reproduce CPUPathTracer exactly: it renders ONE sample
		// per render() call (CPUPathTracer.cpp:57-84) and ignores every RenderSettings field except the resolution --
		// max_bounces is the literal 4 (:199), roulette starts after 2 bounces (:264), the exposure lines are
		// commented out (:101-104).
		struct Options
		{
			// honour RenderSettings::getMaxBounces / getRussianRouletteDepth (integrator) and getExposure /
			// getAutoExposure / getTargetLuminance (resolve stage); re-read whenever the settings are dirty
			bool honour_render_settings = false;
			// samples per pixel enqueued by one render() call (one UI frame); get_frame_count advances by this
			uint32_t frames_per_render = 1;
		};
		void set_options(const Options &options);
		const Options &get_options() const { return m_options; }
		uint32_t get_frame_count() const { return m_frameCount; }

	private:
		void invalidate();
		void upload_scene();
		void apply_render_settings();

		cudapt_ctx *m_ctx = nullptr;
		std::shared_ptr<Scene> m_scene;
		std::shared_ptr<RenderSettings> m_renderSettings;
		PathTracer::RenderResult m_render_result;
		uint32_t m_frameCount = 0;
		Options m_options;
		bool m_options_dirty = false;
		bool m_result_pinned = false;  // m_render_result.image_buffer is registered with cudapt_host_register
	};

}
