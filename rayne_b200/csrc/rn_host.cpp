// librn_cuda - host-side pieces of the path that stay on the CPU:
//   * rn_camera_build_frustum       Camera::PostUpdate / UpdateProjection / UpdateFrustum
//   * rn_scenebasic_render_order    SceneBasic::AddNode / AddRenderNode list order
// Compiled with -ffp-contract=off: the 28 frustum floats feed a bit-exact visibility test, so every operation
// below is rounded like the reference's scalar fp32 code.  No CPU fallback of any kernel lives here.
#include "rn_internal.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <vector>

namespace {

struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };
struct M4 { float m[16]; };

M4 identity()
{
	M4 r;
	for(int i = 0; i < 16; i++) r.m[i] = 0.0f;
	r.m[0] = r.m[5] = r.m[10] = r.m[15] = 1.0f;
	return r;
}

// Matrix::operator* (reference: Source/Math/RNMatrix.h:90-142): column-major, four products summed left to right.
M4 mul(const M4 &a, const M4 &b)
{
	M4 r;
	for(int c = 0; c < 4; c++)
		for(int row = 0; row < 4; row++)
			r.m[c * 4 + row] = a.m[row] * b.m[c * 4] + a.m[4 + row] * b.m[c * 4 + 1] + a.m[8 + row] * b.m[c * 4 + 2] + a.m[12 + row] * b.m[c * 4 + 3];
	return r;
}

// Matrix::operator*(Vector4) (reference: Source/Math/RNMatrix.h:226-232)
V4 mul(const M4 &a, const V4 &v)
{
	V4 r;
	r.x = a.m[0] * v.x + a.m[4] * v.y + a.m[8] * v.z + a.m[12] * v.w;
	r.y = a.m[1] * v.x + a.m[5] * v.y + a.m[9] * v.z + a.m[13] * v.w;
	r.z = a.m[2] * v.x + a.m[6] * v.y + a.m[10] * v.z + a.m[14] * v.w;
	r.w = a.m[3] * v.x + a.m[7] * v.y + a.m[11] * v.z + a.m[15] * v.w;
	return r;
}

// Quaternion::GetRotationMatrix (reference: Source/Math/RNQuaternion.h:468-493)
M4 rotation_matrix(const float q[4])
{
	const float x = q[0], y = q[1], z = q[2], w = q[3];
	const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, xz = x * z, xw = x * w, yz = y * z, yw = y * w, zw = z * w;
	M4 r = identity();
	r.m[0] = 1.0f - 2.0f * (yy + zz);
	r.m[4] = 2.0f * (xy - zw);
	r.m[8] = 2.0f * (xz + yw);
	r.m[1] = 2.0f * (xy + zw);
	r.m[5] = 1.0f - 2.0f * (xx + zz);
	r.m[9] = 2.0f * (yz - xw);
	r.m[2] = 2.0f * (xz - yw);
	r.m[6] = 2.0f * (yz + xw);
	r.m[10] = 1.0f - 2.0f * (xx + yy);
	return r;
}

// Matrix::GetDeterminant / GetSubmatrixDeterminant / GetInverse (reference: Source/Math/RNMatrix.h:245-257,438-465)
float determinant(const M4 &a)
{
	const float *m = a.m;
	float det = m[0] * (m[5] * m[10] * m[15] + m[6] * m[11] * m[13] + m[7] * m[9] * m[14] - m[7] * m[10] * m[13] - m[6] * m[9] * m[15] - m[5] * m[11] * m[14]);
	det -= m[1] * (m[4] * m[10] * m[15] + m[6] * m[11] * m[12] + m[7] * m[8] * m[14] - m[7] * m[10] * m[12] - m[6] * m[8] * m[15] - m[4] * m[11] * m[14]);
	det += m[2] * (m[4] * m[9] * m[15] + m[5] * m[11] * m[12] + m[7] * m[8] * m[13] - m[7] * m[9] * m[12] - m[5] * m[8] * m[15] - m[4] * m[11] * m[13]);
	det -= m[3] * (m[4] * m[9] * m[14] + m[5] * m[10] * m[12] + m[6] * m[8] * m[13] - m[6] * m[9] * m[12] - m[5] * m[8] * m[14] - m[4] * m[10] * m[13]);
	return det;
}

float cofactor(const M4 &a, int k)
{
	const float *m = a.m;
	const int i = k % 4, j = k / 4;
	int i1 = (i == 0);
	int i2 = i1 + 1 + (i == 1);
	int i3 = i2 + 1 + (i == 2);
	const int j1 = (j == 0);
	const int j2 = j1 + 1 + (j == 1);
	const int j3 = j2 + 1 + (j == 2);
	i1 *= 4; i2 *= 4; i3 *= 4;
	float det = m[i1 + j1] * m[i2 + j2] * m[i3 + j3] + m[i1 + j2] * m[i2 + j3] * m[i3 + j1] + m[i1 + j3] * m[i2 + j1] * m[i3 + j2]
	          - m[i1 + j3] * m[i2 + j2] * m[i3 + j1] - m[i1 + j1] * m[i2 + j3] * m[i3 + j2] - m[i1 + j2] * m[i2 + j1] * m[i3 + j3];
	det *= 1 - 2 * ((i + j) % 2);
	return det;
}

M4 inverse(const M4 &a)
{
	M4 r;
	const float det = determinant(a);
	for(int i = 0; i < 16; i++) r.m[i] = cofactor(a, i) / det;
	return r;
}

V3 sub(const V3 &a, const V3 &b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
V3 add(const V3 &a, const V3 &b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
V3 scale(const V3 &a, float n) { return {a.x * n, a.y * n, a.z * n}; }
float dot(const V3 &a, const V3 &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
V3 cross(const V3 &a, const V3 &b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }

// Vector3::Normalize (reference: Source/Math/RNVector.h:519-530; Math::InverseSqrt Source/Math/RNMath.cpp:21-24:
// `1.0 / sqrtf(x)` is a double division rounded to float on return)
V3 normalize(V3 v)
{
	const float len2 = v.x * v.x + v.y * v.y + v.z * v.z;
	if(len2 > 1.1920928955078125e-7f)
	{
		const float inv = 1.0f * static_cast<float>(1.0 / sqrtf(len2));
		v.x *= inv; v.y *= inv; v.z *= inv;
	}
	return v;
}

// Quaternion::GetRotatedVector (reference: Source/Math/RNQuaternion.h:454-459 with operator*= :64-74)
void quat_mul(const float a[4], const float b[4], float r[4])
{
	r[3] = -a[0] * b[0] - a[1] * b[1] - a[2] * b[2] + a[3] * b[3];
	r[0] = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
	r[1] = -a[0] * b[2] + a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
	r[2] = a[0] * b[1] - a[1] * b[0] + a[2] * b[3] + a[3] * b[2];
}
V3 rotate(const float q[4], const V3 &v)
{
	const float vq[4] = {v.x, v.y, v.z, 0.0f};
	const float conj[4] = {-q[0], -q[1], -q[2], q[3]};
	float t[4], r[4];
	quat_mul(q, vq, t);
	quat_mul(t, conj, r);
	return {r[0], r[1], r[2]};
}

struct Plane { V3 n; float d; };

// Plane::WithTriangle / WithPositionNormal / CalculateD (reference: Source/Math/RNPlane.h:56-107)
Plane plane_triangle(const V3 &p1, const V3 &p2, const V3 &p3, float dirfac, float offset)
{
	const V3 d1 = sub(p2, p1), d2 = sub(p3, p1);
	Plane pl;
	pl.n = normalize(scale(cross(d1, d2), dirfac));
	pl.d = dot(pl.n, p1) + offset;
	return pl;
}
Plane plane_position_normal(const V3 &p, const V3 &n)
{
	Plane pl;
	pl.n = normalize(n);
	pl.d = dot(pl.n, p) + 0.0f;
	return pl;
}

} // namespace

// Camera::PostUpdate (RNCamera.cpp:300-305) and UpdateProjection (:314-340) / SetProjectionMatrix (:221-228)
static void camera_matrices(const rn_camera_desc *cam, M4 &invView, M4 &proj)
{
	// Camera::PostUpdate (RNCamera.cpp:300-305): inverse view = T(worldPos) * R(worldRot) * S(local scale)
	invView = identity();
	invView.m[12] = cam->world_position[0];
	invView.m[13] = cam->world_position[1];
	invView.m[14] = cam->world_position[2];
	invView = mul(invView, rotation_matrix(cam->world_rotation));
	M4 sc = identity();
	sc.m[0] = cam->scale[0]; sc.m[5] = cam->scale[1]; sc.m[10] = cam->scale[2];
	invView = mul(invView, sc);

	// Camera::UpdateProjection (RNCamera.cpp:314-340) / SetProjectionMatrix (:221-228)
	proj = identity();
	if(cam->has_projection)
	{
		memcpy(proj.m, cam->projection, sizeof(proj.m));
	}
	else if(cam->flags & RN_CAMERA_ORTHOGONAL)
	{
		// Matrix::WithProjectionOrthogonal (RNMatrix.h:339-364), reverse-Z
		const float left = cam->ortho[0], right = cam->ortho[1], bottom = cam->ortho[2], top = cam->ortho[3];
		const float rl = right - left, tb = top - bottom, fn = cam->clip_far - cam->clip_near;
		const float tx = -(right + left) / (right - left);
		const float ty = -(top + bottom) / (top - bottom);
		const float tz = -(cam->clip_near) / (cam->clip_far - cam->clip_near);
		proj.m[0] = 2.0f / rl;
		proj.m[5] = 2.0f / tb;
		proj.m[10] = 1.0f / fn;
		proj.m[12] = tx;
		proj.m[13] = ty;
		proj.m[14] = static_cast<float>(-tz + 1.0);
		proj.m[15] = 1.0f;
	}
	else
	{
		// Matrix::WithProjectionPerspective(fov, aspect, near, far) (RNMatrix.h:366-386)
		const float pi = 3.14159265358979323846264338327950288f;
		const float rad = cam->fov * pi / 180.0f;
		const float h = std::cos(0.5f * rad) / std::sin(0.5f * rad);
		const float w = h / cam->aspect;
		proj.m[0] = w;
		proj.m[5] = h;
		proj.m[10] = static_cast<float>(cam->clip_far / (cam->clip_far - cam->clip_near) - 1.0);
		proj.m[11] = -1.0f;
		proj.m[14] = (cam->clip_far * cam->clip_near) / (cam->clip_far - cam->clip_near);
		proj.m[15] = 0.0f;
	}
}

extern "C" int rn_camera_build_frustum(const rn_camera_desc *cam, rn_camera_frustum *out)
{
	RN_REQUIRE(cam && out, RN_ERR_INVALID, "NULL argument");
	memset(out, 0, sizeof(*out));
	out->flags = cam->flags;
	out->render_group = cam->render_group;

	// Camera::UpdateFrustum, UseSimpleCulling branch (RNCamera.cpp:347-352): centre is camera-local (reference quirk)
	if(cam->flags & RN_CAMERA_USE_SIMPLE_CULLING)
	{
		out->center[0] = 0.0f;
		out->center[1] = 0.0f;
		out->center[2] = cam->clip_far * 0.5f;
		out->radius = static_cast<float>(cam->clip_far * 1.5);
		// the six planes keep their default-constructed value (Plane::Plane, RNPlane.h:49-54); they are never tested
		for(int k = 0; k < 6; k++) out->planes[k][1] = 1.0f;
		return RN_OK;
	}

	M4 invView, proj;
	camera_matrices(cam, invView, proj);
	const M4 invProj = inverse(proj);

	// Camera::__ToWorld (RNCamera.cpp:386-396): (invView * invProj) * ndc, then divide by w
	const M4 toWorld = mul(invView, invProj);
	auto to_world = [&](float x, float y, float z) -> V3 {
		V4 v = mul(toWorld, V4{x, y, z, 1.0f});
		const float w = v.w;
		return V3{v.x / w, v.y / w, v.z / w};
	};
	// reverse-Z: z = 1 near, z = 0 far (RNCamera.cpp:354-360)
	const V3 pos1 = to_world(-1.0f, 1.0f, 1.0f);
	const V3 pos2 = to_world(-1.0f, 1.0f, 0.0f);
	const V3 pos3 = to_world(-1.0f, -1.0f, 0.0f);
	const V3 pos4 = to_world(1.0f, -1.0f, 1.0f);
	const V3 pos5 = to_world(1.0f, 1.0f, 0.0f);
	const V3 pos6 = to_world(1.0f, -1.0f, 0.0f);

	const V3 position = {cam->world_position[0], cam->world_position[1], cam->world_position[2]};
	const V3 direction = rotate(cam->world_rotation, V3{0.0f, 0.0f, -1.0f});

	auto max6 = [](float a1, float a4, float a2, float a3, float a5, float a6) {
		return std::max(a1, std::max(a4, std::max(a2, std::max(a3, std::max(a5, a6)))));
	};
	auto min6 = [](float a1, float a4, float a2, float a3, float a5, float a6) {
		return std::min(a1, std::min(a4, std::min(a2, std::min(a3, std::min(a5, a6)))));
	};
	const V3 vmax = {max6(pos1.x, pos4.x, pos2.x, pos3.x, pos5.x, pos6.x), max6(pos1.y, pos4.y, pos2.y, pos3.y, pos5.y, pos6.y),
	                 max6(pos1.z, pos4.z, pos2.z, pos3.z, pos5.z, pos6.z)};
	const V3 vmin = {min6(pos1.x, pos4.x, pos2.x, pos3.x, pos5.x, pos6.x), min6(pos1.y, pos4.y, pos2.y, pos3.y, pos5.y, pos6.y),
	                 min6(pos1.z, pos4.z, pos2.z, pos3.z, pos5.z, pos6.z)};
	const V3 center = scale(add(vmax, vmin), 0.5f);
	const V3 diff = sub(center, vmax);
	out->center[0] = center.x; out->center[1] = center.y; out->center[2] = center.z;
	out->radius = sqrtf(diff.x * diff.x + diff.y * diff.y + diff.z * diff.z);

	// planes: left, right, top, bottom, near, far (RNCamera.cpp:378-383); offsets = top, bottom, left, right
	const float *offs = cam->plane_offsets;
	const Plane planes[6] = {
	    plane_triangle(pos1, pos2, pos3, 1.0f, -offs[2]),
	    plane_triangle(pos4, pos5, pos6, -1.0f, offs[3]),
	    plane_triangle(pos1, pos2, pos5, -1.0f, offs[0]),
	    plane_triangle(pos4, pos3, pos6, 1.0f, -offs[1]),
	    plane_position_normal(add(position, scale(direction, std::min(cam->clip_near, cam->clip_far))), V3{-direction.x, -direction.y, -direction.z}),
	    plane_position_normal(add(position, scale(direction, std::max(cam->clip_near, cam->clip_far))), direction),
	};
	for(int k = 0; k < 6; k++)
	{
		out->planes[k][0] = planes[k].n.x;
		out->planes[k][1] = planes[k].n.y;
		out->planes[k][2] = planes[k].n.z;
		out->planes[k][3] = planes[k].d;
	}
	return RN_OK;
}

// Camera::GetViewMatrix / GetProjectionMatrix after PostUpdate (RNCamera.cpp:296-311): view = GetInverse of the inverse
// view matrix; proj_view = projection * view, the product the renderers form per camera
// (Modules/Vulkan/Sources/RNVulkanRenderer.cpp:1134-1168).
extern "C" int rn_camera_build_matrices(const rn_camera_desc *cam, float *view, float *projection, float *proj_view)
{
	RN_REQUIRE(cam, RN_ERR_INVALID, "NULL argument");
	M4 invView, proj;
	camera_matrices(cam, invView, proj);
	const M4 v = inverse(invView);
	if(view) memcpy(view, v.m, sizeof(v.m));
	if(projection) memcpy(projection, proj.m, sizeof(proj.m));
	if(proj_view)
	{
		const M4 pv = mul(proj, v);
		memcpy(proj_view, pv.m, sizeof(pv.m));
	}
	return RN_OK;
}

// Light::UpdateShadows (Source/Scene/RNLight.cpp:475-529, directional branch) + Camera::MakeShadowSplit
// (Source/Scene/RNCamera.cpp:240-287): the orthographic cascade cameras of a directional light for a target camera.
extern "C" int rn_light_build_cascades(const rn_camera_desc *target, const rn_shadow_desc *sh, rn_camera_desc *out_cameras,
                                       float *out_proj_view)
{
	RN_REQUIRE(target && sh && out_cameras, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(sh->n_splits >= 1 && sh->n_splits <= RN_MAX_CAMERAS, RN_ERR_INVALID, "n_splits must be 1..%d, got %u", RN_MAX_CAMERAS, sh->n_splits);
	M4 invView, proj;
	camera_matrices(target, invView, proj);
	const M4 invProj = inverse(proj);
	// Camera::ToWorld (RNCamera.cpp:399-415): un-project (x, y) at ndc z = 0, normalise to |view z| = 1, scale to the
	// view distance dir.z, then to world space
	auto to_world = [&](float x, float y, float dist) -> V3 {
		V4 v = mul(invProj, V4{x, y, 0.0f, 1.0f});
		const float w = v.w;
		v.x /= w; v.y /= w; v.z /= w; v.w /= w;
		const float sign = v.z < 0.0f ? -1.0f : 1.0f;
		const float dz = v.z * sign;
		v.x /= dz; v.y /= dz; v.z /= dz; v.w /= dz;
		v.x *= dist; v.y *= dist; v.z *= dist;
		v.w = 1.0f;
		v = mul(invView, v);
		return V3{v.x, v.y, v.z};
	};
	const float *lq = sh->light_world_rotation;
	const V3 forward = rotate(lq, V3{0.0f, 0.0f, -1.0f});          // SceneNode::GetForward
	const M4 rot = rotation_matrix(lq);
	const M4 rotInv = inverse(rot);
	auto mul3 = [](const M4 &a, const V3 &v) -> V3 {                 // Matrix::operator*(Vector3), RNMatrix.h:145-151
		V3 r;
		r.x = a.m[0] * v.x + a.m[4] * v.y + a.m[8] * v.z + a.m[12];
		r.y = a.m[1] * v.x + a.m[5] * v.y + a.m[9] * v.z + a.m[13];
		r.z = a.m[2] * v.x + a.m[6] * v.y + a.m[10] * v.z + a.m[14];
		return r;
	};
	float near = target->clip_near;
	float far = 0.01f;
	const float n = float(sh->n_splits);
	for(uint32_t i = 0; i < sh->n_splits; i++)
	{
		// RNLight.cpp:505-514: the split distance, a blend of a linear and a logarithmic subdivision of the target's depth range
		if(sh->split_max_distance[i] < far)
		{
			const float linear = target->clip_near + (sh->max_shadow_dist * target->clip_far - target->clip_near) * (i + 1.0f) / n;
			const float log = target->clip_near * powf(sh->max_shadow_dist * target->clip_far / target->clip_near, (i + 1.0f) / n);
			far = linear * sh->distance_blend_factor + log * (1.0f - sh->distance_blend_factor);
		}
		else
			far = sh->split_max_distance[i];

		// Camera::MakeShadowSplit (RNCamera.cpp:240-287)
		const V3 nearcenter = to_world(0.0f, 0.0f, near);
		const V3 farcorner1 = to_world(1.0f, 1.0f, far);
		const V3 farcorner2 = to_world(-1.0f, -1.0f, far);
		const V3 farcenter = scale(add(farcorner1, farcorner2), 0.5f);
		const V3 center = scale(add(nearcenter, farcenter), 0.5f);
		const V3 dv = sub(center, farcorner1);
		const float dist = sqrtf(dv.x * dv.x + dv.y * dv.y + dv.z * dv.z);
		const V3 pixelsize = {dist * 2.0f / sh->resolution, dist * 2.0f / sh->resolution, 1.0f / 1.0f};
		V3 pos = sub(center, scale(forward, sh->directional_shadow_distance));
		pos = mul3(rotInv, pos);
		pos.x /= pixelsize.x; pos.y /= pixelsize.y; pos.z /= pixelsize.z;      // snap to the texel grid in light space
		pos.x = floorf(pos.x); pos.y = floorf(pos.y); pos.z = floorf(pos.z);
		pos.x *= pixelsize.x; pos.y *= pixelsize.y; pos.z *= pixelsize.z;
		pos = mul3(rot, pos);

		rn_camera_desc &c = out_cameras[i];
		memset(&c, 0, sizeof(c));
		c.flags = RN_CAMERA_ORTHOGONAL | RN_CAMERA_RENDER_EARLY;     // RNLight.cpp:318
		c.render_group = 1;                                          // Camera::Camera default (_renderGroup = 1)
		c.world_position[0] = pos.x; c.world_position[1] = pos.y; c.world_position[2] = pos.z;
		memcpy(c.world_rotation, lq, sizeof(float) * 4);
		c.scale[0] = c.scale[1] = c.scale[2] = 1.0f;
		c.fov = 70.0f; c.aspect = 1.0f;                              // unused by an orthogonal camera
		c.clip_near = sh->clip_near; c.clip_far = sh->clip_far;
		c.ortho[0] = -dist; c.ortho[1] = dist; c.ortho[2] = -dist; c.ortho[3] = dist;
		if(out_proj_view)
		{
			// _projectionMatrix * GetWorldTransform().GetInverse(); the node's world transform is T * R * S(1) (RNSceneNode.cpp:530-532)
			M4 cInvView, cProj;
			camera_matrices(&c, cInvView, cProj);
			const M4 pv = mul(cProj, inverse(cInvView));
			memcpy(out_proj_view + (size_t)i * 16, pv.m, sizeof(pv.m));
		}
		near = far;
	}
	return RN_OK;
}

// SceneBasic::AddNode (RNSceneBasic.cpp:683-736): UpdateSceneInfo adds the children first, in child order
// (RNSceneNode.cpp:286-291), then the node itself goes through AddRenderNode (:652-676):
//   (1) list empty or head.priority >= p  -> PushFront
//   (2) tail.priority <= p                -> PushBack
//   (3) otherwise walk back from the tail while priority > p, insert after.
extern "C" int rn_scenebasic_render_order(uint32_t n, const int32_t *parent, const int32_t *prio, uint32_t *order)
{
	RN_REQUIRE(parent && order, RN_ERR_INVALID, "NULL argument");
	if(n == 0) return RN_OK;
	// children in creation order (== AddChild order)
	std::vector<uint32_t> ofs(n + 1, 0), child;
	for(uint32_t i = 0; i < n; i++)
	{
		RN_REQUIRE(parent[i] < (int32_t)i, RN_ERR_INVALID, "parent[%u] = %d must precede its child", i, parent[i]);
		if(parent[i] >= 0) ofs[parent[i] + 1]++;
	}
	for(uint32_t i = 0; i < n; i++) ofs[i + 1] += ofs[i];
	child.resize(ofs[n]);
	{
		std::vector<uint32_t> fill(ofs.begin(), ofs.end() - 1);
		for(uint32_t i = 0; i < n; i++)
			if(parent[i] >= 0) child[fill[parent[i]]++] = i;
	}
	// AddRenderNode keeps the list sorted by priority.  Inside one priority a node lands BEHIND its equals (cases 2
	// and 3) unless its priority is <= the head's, i.e. <= the smallest priority seen so far (case 1: PushFront, in
	// front of its equals).  So the list is, per priority ascending, a deque: reverse(front pushes) ++ back pushes -
	// the same order as the reference's linked-list walk without its O(list) step per insert.
	struct Group { std::vector<uint32_t> front, back; };
	std::map<int32_t, Group> groups;
	bool empty = true;
	int32_t minp = 0;
	auto priority = [&](uint32_t i) { return prio ? prio[i] : 0; };
	auto insert = [&](uint32_t node) {
		const int32_t p = priority(node);
		if(empty || minp >= p)
		{
			groups[p].front.push_back(node);
			minp = p;
			empty = false;
		}
		else
			groups[p].back.push_back(node);
	};
	// iterative post-order from every root, roots in creation order
	std::vector<uint32_t> stack_node, stack_pos;
	for(uint32_t root = 0; root < n; root++)
	{
		if(parent[root] >= 0) continue;
		stack_node.push_back(root);
		stack_pos.push_back(ofs[root]);
		while(!stack_node.empty())
		{
			const uint32_t u = stack_node.back();
			uint32_t &k = stack_pos.back();
			if(k < ofs[u + 1])
			{
				const uint32_t c = child[k++];
				stack_node.push_back(c);
				stack_pos.push_back(ofs[c]);
			}
			else
			{
				insert(u);
				stack_node.pop_back();
				stack_pos.pop_back();
			}
		}
	}
	uint32_t r = 0;
	for(const auto &kv : groups)
	{
		const Group &g = kv.second;
		for(size_t q = g.front.size(); q-- > 0;) order[r++] = g.front[q];
		for(uint32_t id : g.back) order[r++] = id;
	}
	RN_REQUIRE(r == n, RN_ERR_INVALID, "internal error: %u of %u nodes ordered", r, n);
	return RN_OK;
}

// SceneWithVisibilityLists::AddNode (RNSceneWithVisibilityLists.cpp:210-236): UpdateSceneInfo adds the children first
// (RNSceneNode.cpp:286-291), then the node is push_back-ed into the default volume: plain post-order, no priorities.
extern "C" int rn_scenevis_default_volume_order(uint32_t n, const int32_t *parent, uint32_t *order)
{
	RN_REQUIRE(parent && order, RN_ERR_INVALID, "NULL argument");
	if(n == 0) return RN_OK;
	std::vector<uint32_t> ofs(n + 1, 0), child;
	for(uint32_t i = 0; i < n; i++)
	{
		RN_REQUIRE(parent[i] < (int32_t)i, RN_ERR_INVALID, "parent[%u] = %d must precede its child", i, parent[i]);
		if(parent[i] >= 0) ofs[parent[i] + 1]++;
	}
	for(uint32_t i = 0; i < n; i++) ofs[i + 1] += ofs[i];
	child.resize(ofs[n]);
	{
		std::vector<uint32_t> fill(ofs.begin(), ofs.end() - 1);
		for(uint32_t i = 0; i < n; i++)
			if(parent[i] >= 0) child[fill[parent[i]]++] = i;
	}
	std::vector<uint32_t> stack_node, stack_pos;
	uint32_t r = 0;
	for(uint32_t root = 0; root < n; root++)
	{
		if(parent[root] >= 0) continue;
		stack_node.push_back(root);
		stack_pos.push_back(ofs[root]);
		while(!stack_node.empty())
		{
			const uint32_t u = stack_node.back();
			uint32_t &k = stack_pos.back();
			if(k < ofs[u + 1])
			{
				const uint32_t c = child[k++];
				stack_node.push_back(c);
				stack_pos.push_back(ofs[c]);
			}
			else
			{
				order[r++] = u;
				stack_node.pop_back();
				stack_pos.pop_back();
			}
		}
	}
	RN_REQUIRE(r == n, RN_ERR_INVALID, "internal error: %u of %u nodes ordered", r, n);
	return RN_OK;
}

// SceneWithVisibilityLists::Render picks, per camera, the FIRST volume whose ContainsPosition is true
// (RNSceneWithVisibilityLists.cpp:157-166); AxisAlignedBoxVolume::ContainsPosition (:34-37) is the closed box test.
extern "C" int rn_scenevis_choose_volume(const float *bounds, uint32_t n_volumes, const float *camera_world_position, uint8_t *out_volume)
{
	RN_REQUIRE((bounds || n_volumes == 0) && camera_world_position && out_volume, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(n_volumes <= 254, RN_ERR_INVALID, "%u volumes, at most 254", n_volumes);
	const float x = camera_world_position[0], y = camera_world_position[1], z = camera_world_position[2];
	*out_volume = 0;
	for(uint32_t k = 0; k < n_volumes; k++)
	{
		const float *b = bounds + (size_t)k * 6;
		if(z >= b[2] && y >= b[1] && x >= b[0] && x <= b[3] && y <= b[4] && z <= b[5])
		{
			*out_volume = (uint8_t)(k + 1);
			break;
		}
	}
	return RN_OK;
}
