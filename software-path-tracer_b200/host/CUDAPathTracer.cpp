#include "CUDAPathTracer.h"

#include <algorithm>
#include <vector>

#include "cudapt.h"
#include "render/Log.h"
#include "render/Scene.h"
#include "render/Types.h"
#include "render_assert.h"

namespace render
{
	namespace
	{
		// Reference convention: no error codes on the render path, verify() prints and aborts
		// (render_assert.h:15-25, CPUPathTracer.cpp:45-46).  CUDA errors are never swallowed.
		void check_status(cudapt_ctx *ctx, int status, const char *what)
		{
			if (status != CUDAPT_OK)
				verify(false, std::string(what) + ": " + cudapt_last_error(ctx));
		}
	}

	CUDAPathTracer::CUDAPathTracer()
	{
		render::Log::info("Initializing CUDA Path Tracer (sm_100a) backend...");
		m_renderSettings = std::make_shared<RenderSettings>();  // as CPUPathTracer.cpp:33
		cudapt_config cfg;
		cudapt_default_config(&cfg);  // the reference's literals: PCG, 4 bounces, RR after 2, albedo 0.7
		const int status = cudapt_create(&cfg, &m_ctx);
		if (status != CUDAPT_OK)
			verify(false, std::string("cudapt_create: ") + cudapt_last_error(nullptr));
	}

	CUDAPathTracer::~CUDAPathTracer()
	{
		if (m_result_pinned)
			cudapt_host_unregister(m_ctx, m_render_result.image_buffer.data());
		cudapt_destroy(m_ctx);
	}

	void CUDAPathTracer::render()
	{
		verify(m_ctx != nullptr, "CUDA backend not initialized");
		verify(m_scene != nullptr, "Scene not set before rendering");  // CPUPathTracer.cpp:46
		invalidate();
		const uint32_t n = m_options.frames_per_render ? m_options.frames_per_render : 1u;
		check_status(m_ctx, cudapt_render(m_ctx, n), "cudapt_render");  // one sample per pixel per call by default (:57-84)
		m_frameCount += n;
	}

	const PathTracer::RenderResult &CUDAPathTracer::get_render_result()
	{
		verify(m_frameCount > 0, "No frames rendered yet");  // CPUPathTracer.cpp:89
		check_status(m_ctx, cudapt_resolve_rgba8(m_ctx, m_render_result.image_buffer.data()), "cudapt_resolve_rgba8");
		return m_render_result;
	}

	void CUDAPathTracer::set_options(const Options &options)
	{
		m_options = options;
		m_options_dirty = true;  // picked up by the next render(), which restarts the accumulation
	}

	// The RenderSettings fields the reference declares but never reads (Types.h:65-69).  Off: the reference's literals.
	void CUDAPathTracer::apply_render_settings()
	{
		cudapt_config cfg;
		cudapt_default_config(&cfg);
		cudapt_postprocess post;
		cudapt_default_postprocess(&post);
		if (m_options.honour_render_settings)
		{
			cfg.max_bounces = (int32_t)m_renderSettings->getMaxBounces();
			cfg.rr_start = (int32_t)m_renderSettings->getRussianRouletteDepth();
			post.exposure = m_renderSettings->getExposure();
			post.auto_exposure = m_renderSettings->getAutoExposure() ? 1 : 0;
			post.target_luminance = m_renderSettings->getTargetLuminance();
		}
		check_status(m_ctx, cudapt_set_config(m_ctx, &cfg), "cudapt_set_config");
		check_status(m_ctx, cudapt_set_postprocess(m_ctx, &post), "cudapt_set_postprocess");
	}

	// CPUPathTracer.cpp:119-161, rule for rule.
	void CUDAPathTracer::invalidate()
	{
		bool needs_rebuild = false;
		bool restart = false;
		if (m_scene->hasChanges())
		{
			restart = true;
			needs_rebuild = true;
		}
		if (m_renderSettings->isDirty() || m_options_dirty)
		{
			restart = true;
			apply_render_settings();
			m_renderSettings->clearDirty();
			m_options_dirty = false;
		}
		if (m_render_result.width != m_renderSettings->getWidth() || m_render_result.height != m_renderSettings->getHeight())
		{
			m_render_result.width = m_renderSettings->getWidth();
			m_render_result.height = m_renderSettings->getHeight();
			// the result buffer is page-locked, so that get_render_result()'s device-to-host copy lands in it directly
			if (m_result_pinned)
				cudapt_host_unregister(m_ctx, m_render_result.image_buffer.data());
			m_render_result.image_buffer.resize((size_t)m_render_result.width * m_render_result.height);
			m_result_pinned = !m_render_result.image_buffer.empty() &&
				cudapt_host_register(m_ctx, m_render_result.image_buffer.data(), m_render_result.image_buffer.size() * sizeof(uint32_t)) == CUDAPT_OK;
			check_status(m_ctx, cudapt_set_resolution(m_ctx, m_render_result.width, m_render_result.height), "cudapt_set_resolution");
			m_frameCount = 0;
			restart = false;  // set_resolution zeroes the accumulation buffer
		}
		if (needs_rebuild)
		{
			upload_scene();  // replaces the device scene and restarts accumulation
			m_scene->markChangesProcessed();
			m_frameCount = 0;
			restart = false;
		}
		if (restart)
		{
			check_status(m_ctx, cudapt_reset(m_ctx), "cudapt_reset");
			m_frameCount = 0;
		}
	}

	// CPUPathTracer.cpp:362-403: every SPHERE_OBJECT node becomes {position.xyz, radius}.  The reference walks an
	// unordered_map (unspecified order decides Embree's geomID); here nodes are sorted by NodeID so the primitive
	// order -- and therefore tie-breaking -- is deterministic.  "Rebuild = replace".
	void CUDAPathTracer::upload_scene()
	{
		render::Log::info("Uploading application scene to the CUDA backend...");
		std::vector<const SphereObject *> spheres;
		for (const auto &[id, node] : m_scene->GetAllNodes())
		{
			if (node->GetType() == NodeType::SPHERE_OBJECT)
				spheres.push_back(static_cast<const SphereObject *>(node));
		}
		std::sort(spheres.begin(), spheres.end(), [](const SphereObject *a, const SphereObject *b) { return a->GetID() < b->GetID(); });
		std::vector<float> packed;
		std::vector<uint32_t> ids;
		packed.reserve(4 * spheres.size());
		for (const SphereObject *s : spheres)
		{
			const glm::vec3 pos = s->GetPosition();
			packed.insert(packed.end(), {pos.x, pos.y, pos.z, s->GetRadius()});
			ids.push_back(s->GetID());
		}
		check_status(m_ctx, cudapt_upload_spheres(m_ctx, packed.data(), ids.data(), (uint32_t)spheres.size()), "cudapt_upload_spheres");
	}
}
