// drt_params.h -- host-side evaluation of the per-launch constants the reference recomputes per pixel.
// Same operations, same order, same libm entry points as the host-compiled reference:
//   Camera::GetRay prologue   DustRayTracer/src/Core/Scene/Camera.cu:84-103
//   sunpos / suncol           DustRayTracer/src/Core/Kernel/Shaders/RayGen.cuh:68-72
//   tonemap white scale       DustRayTracer/src/Core/Kernel/Shaders/RayGen.cuh:39-40
#pragma once
#include <math.h>
#include "drt_types.h"
#include "drt_shade.cuh"
#include "../../include/drt.h"

namespace drt {

inline float deg2rad_ref(float degree)
{
	const float PI = 3.14159265359f; // Camera.cu:125-129
	return degree * (PI / 180.f);
}

inline CameraParams make_camera_params(const drt_camera& c, uint32_t width, uint32_t height)
{
	typedef M<true> m;
	CameraParams o;
	float theta = c.vfov_rad / 2;
	float fov_factor = tanf(theta / 2.0f);
	float aspect_ratio = (float)width / (float)height;
	float plane_height = 2.0f * fov_factor * c.focus_dist;
	float plane_width = plane_height * aspect_ratio;
	V3 forward = m::normalize(mk3(c.forward_dir[0], c.forward_dir[1], c.forward_dir[2]));
	V3 right = m::normalize(m::cross(forward, mk3(0.f, 1.f, 0.f)));
	V3 up = m::cross(right, forward);
	o.plane_h = m::scale(right, plane_width);
	o.plane_v = m::scale(up, plane_height);
	float defocus_radius = c.focus_dist * tanf(deg2rad_ref(c.defocus_angle) / 2.0f);
	o.defocus_u = m::scale(right, defocus_radius);
	o.defocus_v = m::scale(up, defocus_radius);
	o.position = mk3(c.position[0], c.position[1], c.position[2]);
	o.forward_focus = m::scale(forward, c.focus_dist);
	o.defocus_angle = c.defocus_angle;
	o.exposure = c.exposure;
	return o;
}

inline ShadeParams make_shade_params(const drt_settings& s)
{
	typedef M<true> m;
	ShadeParams o;
	float sa = sinf(s.sunlight_dir[0]), ca = cosf(s.sunlight_dir[0]), sb = sinf(s.sunlight_dir[1]);
	o.sunpos = m::scale(mk3(sa * (1 - sb), sb, ca * (1 - sb)), 100.f);
	o.suncol = m::scale(mk3(s.sunlight_color[0], s.sunlight_color[1], s.sunlight_color[2]), s.sunlight_intensity);
	o.sky_color = mk3(s.sky_color[0], s.sky_color[1], s.sky_color[2]);
	o.sky_intensity = s.sky_intensity;
	o.white_scale = 1.0f / tonemap_partial<true>(11.2f);
	o.bounce_limit = s.ray_bounce_limit;
	o.enable_sunlight = s.enable_sunlight ? 1 : 0;
	o.tone_mapping = s.tone_mapping ? 1 : 0;
	o.gamma_correction = s.gamma_correction ? 1 : 0;
	o.render_mode = s.render_mode;
	o.debug_mode = s.debug_mode;
	return o;
}

inline void default_camera(drt_camera* c)
{
	// Camera.cuh:32-46 + EditorLayer.cpp:35-40
	c->exposure = 1.f;
	c->vfov_rad = deg2rad_ref(60.f);
	c->zfar = 0.f; c->znear = 0.f; c->aspect_ratio = 0.f;
	c->defocus_angle = 0.f;
	c->focus_dist = 10.f;
	c->movement_speed = 10.f;
	c->position[0] = 0.f; c->position[1] = 2.f; c->position[2] = 5.f;
	c->forward_dir[0] = 0.f; c->forward_dir[1] = 0.f; c->forward_dir[2] = -1.f;
	c->up_dir[0] = 0.f; c->up_dir[1] = 1.f; c->up_dir[2] = 0.f;
	// m_Right_dir = cross(forward, up) (Camera.cuh:17)
	c->right_dir[0] = 1.f; c->right_dir[1] = -0.f; c->right_dir[2] = 0.f;
	c->viewheight = 0; c->viewwidth = 0;
}

inline void default_settings(drt_settings* s)
{
	// RendererSettings.h:22-34
	s->gamma_correction = 1; s->tone_mapping = 1; s->enable_sunlight = 0; s->_pad0 = 0;
	s->max_samples = 500; s->ray_bounce_limit = 2; s->render_mode = 0; s->debug_mode = 0; s->_pad1 = 0;
	s->sunlight_dir[0] = -0.803f; s->sunlight_dir[1] = 0.681f;
	s->sunlight_color[0] = 1.000f; s->sunlight_color[1] = 0.944f; s->sunlight_color[2] = 0.917f;
	s->sunlight_intensity = 30.f;
	s->sky_color[0] = 0.25f; s->sky_color[1] = 0.498f; s->sky_color[2] = 0.80f;
	s->sky_intensity = 20.f;
}

} // namespace drt
