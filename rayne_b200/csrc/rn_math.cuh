// Device-side fp32 arithmetic of the scene hot path, in the reference's operand order.
//
// This translation unit is compiled with -fmad=false -prec-div=true -prec-sqrt=true -ftz=false: every
// `a * b + c` below is a separately rounded multiply and add, exactly as the strict-FP build of the reference
// evaluates it.  Terms the reference multiplies by a literal 0 are dropped (they only add
// +-0 and cannot change a non-zero value or a comparison).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rn {

struct Quat { float x, y, z, w; };
struct Vec3 { float x, y, z; };

// Quaternion::operator*= (reference: Source/Math/RNQuaternion.h:64-74), a = this, b = other.
__device__ __forceinline__ Quat quat_mul(const Quat &a, const Quat &b)
{
	Quat r;
	r.w = -a.x * b.x - a.y * b.y - a.z * b.z + a.w * b.w;
	r.x =  a.x * b.w + a.y * b.z - a.z * b.y + a.w * b.x;
	r.y = -a.x * b.z + a.y * b.w + a.z * b.x + a.w * b.y;
	r.z =  a.x * b.y - a.y * b.x + a.z * b.w + a.w * b.z;
	return r;
}

// Quaternion::GetRotatedVector (reference: Source/Math/RNQuaternion.h:454-459):
// (q * (v, 0)) * conj(q), two Hamilton products left to right; the `* 0` terms of the first product and the
// unused w of the second are dropped, the negations of conj(q) are folded into the signs (exact).
__device__ __forceinline__ Vec3 quat_rotate(const Quat &q, const Vec3 &v)
{
	float tw = -q.x * v.x - q.y * v.y - q.z * v.z;
	float tx =  q.y * v.z - q.z * v.y + q.w * v.x;
	float ty = -q.x * v.z + q.z * v.x + q.w * v.y;
	float tz =  q.x * v.y - q.y * v.x + q.w * v.z;
	Vec3 r;
	r.x = tx * q.w - ty * q.z + tz * q.y - tw * q.x;
	r.y = tx * q.z + ty * q.w - tz * q.x - tw * q.y;
	r.z = -(tx * q.y) + ty * q.x + tz * q.w - tw * q.z;
	return r;
}

// Affine 3x4 matrix, column-major like RN::Matrix (m[c*4 + r]); the 4th row is (0,0,0,1) for every matrix on the
// path (WithTranslation * rotation * scaling and products of those).
struct Affine
{
	float m[12]; // column c, row r at m[c*3 + r]
};

// Matrix::WithTranslation(pos); Rotate(rot); Scale(scale)  (reference: Source/Scene/RNSceneNode.cpp:530-532,
// Source/Math/RNMatrix.h:32-87,544-566; Quaternion::GetRotationMatrix Source/Math/RNQuaternion.h:468-493).
// The two full 4x4 products reduce exactly to: column c of the rotation matrix times scale_c, translation = pos.
// FMA = false: the reference's operations one by one (bit-exact with its strict-FP build).
// FMA = true : the same formula with fused multiply-adds (explicit __fmaf_rn; the TU is built with -fmad=false), one
//              rounding less per term - within the 1e-5 the north star grants the matrices, like the reference's own
//              release build (-funsafe-math-optimizations), never used for anything the cull reads.
template <bool FMA>
__device__ __forceinline__ Affine local_matrix(const Vec3 &p, const Quat &q, const Vec3 &s)
{
	Affine a;
	if(FMA)
	{
		const float sx2 = s.x + s.x, sy2 = s.y + s.y, sz2 = s.z + s.z;
		const float xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z;
		const float zw = q.z * q.w, yw = q.y * q.w, xw = q.x * q.w;
		a.m[0] = __fmaf_rn(-sx2, yy + zz, s.x);
		a.m[1] = sx2 * __fmaf_rn(q.x, q.y, zw);
		a.m[2] = sx2 * __fmaf_rn(q.x, q.z, -yw);
		a.m[3] = sy2 * __fmaf_rn(q.x, q.y, -zw);
		a.m[4] = __fmaf_rn(-sy2, xx + zz, s.y);
		a.m[5] = sy2 * __fmaf_rn(q.y, q.z, xw);
		a.m[6] = sz2 * __fmaf_rn(q.x, q.z, yw);
		a.m[7] = sz2 * __fmaf_rn(q.y, q.z, -xw);
		a.m[8] = __fmaf_rn(-sz2, xx + yy, s.z);
	}
	else
	{
		float xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z;
		float xy = q.x * q.y, xz = q.x * q.z, xw = q.x * q.w;
		float yz = q.y * q.z, yw = q.y * q.w, zw = q.z * q.w;
		float r0 = 1.0f - 2.0f * (yy + zz), r4 = 2.0f * (xy - zw), r8 = 2.0f * (xz + yw);
		float r1 = 2.0f * (xy + zw), r5 = 1.0f - 2.0f * (xx + zz), r9 = 2.0f * (yz - xw);
		float r2 = 2.0f * (xz - yw), r6 = 2.0f * (yz + xw), r10 = 1.0f - 2.0f * (xx + yy);
		a.m[0] = r0 * s.x; a.m[1] = r1 * s.x; a.m[2] = r2 * s.x;
		a.m[3] = r4 * s.y; a.m[4] = r5 * s.y; a.m[5] = r6 * s.y;
		a.m[6] = r8 * s.z; a.m[7] = r9 * s.z; a.m[8] = r10 * s.z;
	}
	a.m[9] = p.x; a.m[10] = p.y; a.m[11] = p.z;
	return a;
}
__device__ __forceinline__ Affine local_matrix(const Vec3 &p, const Quat &q, const Vec3 &s) { return local_matrix<false>(p, q, s); }

// Matrix::operator* (reference: Source/Math/RNMatrix.h:90-142) restricted to affine operands:
// r[c][r] = P[0][r]*L[c][0] + P[1][r]*L[c][1] + P[2][r]*L[c][2] (+ P[3][r] for the translation column),
// summed left to right like the reference (FMA = true: the same sums as chains of fused multiply-adds).
template <bool FMA>
__device__ __forceinline__ Affine affine_mul(const Affine &P, const Affine &L)
{
	Affine o;
#pragma unroll
	for(int c = 0; c < 3; c++)
	{
#pragma unroll
		for(int r = 0; r < 3; r++)
		{
			if(FMA)
				o.m[c * 3 + r] = __fmaf_rn(P.m[6 + r], L.m[c * 3 + 2], __fmaf_rn(P.m[3 + r], L.m[c * 3 + 1], P.m[r] * L.m[c * 3]));
			else
				o.m[c * 3 + r] = P.m[r] * L.m[c * 3] + P.m[3 + r] * L.m[c * 3 + 1] + P.m[6 + r] * L.m[c * 3 + 2];
		}
	}
#pragma unroll
	for(int r = 0; r < 3; r++)
	{
		if(FMA)
			o.m[9 + r] = __fmaf_rn(P.m[6 + r], L.m[11], __fmaf_rn(P.m[3 + r], L.m[10], __fmaf_rn(P.m[r], L.m[9], P.m[9 + r])));
		else
			o.m[9 + r] = P.m[r] * L.m[9] + P.m[3 + r] * L.m[10] + P.m[6 + r] * L.m[11] + P.m[9 + r];
	}
	return o;
}
__device__ __forceinline__ Affine affine_mul(const Affine &P, const Affine &L) { return affine_mul<false>(P, L); }

// ---- A3: inverse transforms -----------------------------------------------------------------------------------
// SceneNode::UpdateInternalInverseTransformData (reference: Source/Scene/RNSceneNode.cpp:552-554):
//   WithScaling(scale != 0 ? 1 / scale : 0); Rotate(conj(rot)); Translate(pos * -1)
// `scale != 0.0f` is the epsilon compare of all three components (Source/Math/RNVector.h:403-409).  The rotation
// matrix of the conjugate is, float for float, the transpose of the rotation matrix of rot (x*y == (-x)*(-y),
// a - (-b) == a + b), S * Rc scales ROW r by 1/scale_r, Matrix::Translate (Source/Math/RNMatrix.h:514-527) adds
// the dot products of the rows with -pos to a zero translation.
__device__ __forceinline__ Affine inverse_local_matrix(const Vec3 &p, const Quat &q, const Vec3 &s)
{
	const float eps = 1.1920928955078125e-7f; // k::EpsilonFloat = FLT_EPSILON
	Vec3 is = {0.0f, 0.0f, 0.0f};
	if(!(fabsf(s.x) <= eps && fabsf(s.y) <= eps && fabsf(s.z) <= eps))
	{
		is.x = 1.0f / s.x; is.y = 1.0f / s.y; is.z = 1.0f / s.z;
	}
	float xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z;
	float xy = q.x * q.y, xz = q.x * q.z, xw = q.x * q.w;
	float yz = q.y * q.z, yw = q.y * q.w, zw = q.z * q.w;
	// R(q), row-major names: rRC
	float r00 = 1.0f - 2.0f * (yy + zz), r01 = 2.0f * (xy - zw), r02 = 2.0f * (xz + yw);
	float r10 = 2.0f * (xy + zw), r11 = 1.0f - 2.0f * (xx + zz), r12 = 2.0f * (yz - xw);
	float r20 = 2.0f * (xz - yw), r21 = 2.0f * (yz + xw), r22 = 1.0f - 2.0f * (xx + yy);
	Affine a;   // a.m[c*3 + r] = is_r * Rc[r][c] = is_r * R[c][r]
	a.m[0] = is.x * r00; a.m[1] = is.y * r01; a.m[2] = is.z * r02;
	a.m[3] = is.x * r10; a.m[4] = is.y * r11; a.m[5] = is.z * r12;
	a.m[6] = is.x * r20; a.m[7] = is.y * r21; a.m[8] = is.z * r22;
	const float tx = p.x * -1.0f, ty = p.y * -1.0f, tz = p.z * -1.0f;
#pragma unroll
	for(int r = 0; r < 3; r++)
		a.m[9 + r] = a.m[r] * tx + a.m[3 + r] * ty + a.m[6 + r] * tz;
	return a;
}

// ---- A4: transformed AABB -------------------------------------------------------------------------------------
// AABB::operator*=(Vector3) then AABB::SetRotation (reference: Source/Math/RNAABB.h:108-114,152-177;
// Matrix::operator*(Vector3) Source/Math/RNMatrix.h:145-151 with a zero translation).
struct Box { Vec3 mn, mx; };

__device__ __forceinline__ Vec3 rot_apply(const float (&m)[9], float x, float y, float z)
{
	Vec3 r;   // m[c*3 + r]
	r.x = m[0] * x + m[3] * y + m[6] * z;
	r.y = m[1] * x + m[4] * y + m[7] * z;
	r.z = m[2] * x + m[5] * y + m[8] * z;
	return r;
}

__device__ __forceinline__ Box world_aabb(const Box &local, const Vec3 &ws, const Quat &q)
{
	Vec3 mn = {local.mn.x * ws.x, local.mn.y * ws.y, local.mn.z * ws.z};
	Vec3 mx = {local.mx.x * ws.x, local.mx.y * ws.y, local.mx.z * ws.z};
	float xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z;
	float xy = q.x * q.y, xz = q.x * q.z, xw = q.x * q.w;
	float yz = q.y * q.z, yw = q.y * q.w, zw = q.z * q.w;
	float m[9];
	m[0] = 1.0f - 2.0f * (yy + zz); m[3] = 2.0f * (xy - zw); m[6] = 2.0f * (xz + yw);
	m[1] = 2.0f * (xy + zw); m[4] = 1.0f - 2.0f * (xx + zz); m[7] = 2.0f * (yz - xw);
	m[2] = 2.0f * (xz - yw); m[5] = 2.0f * (yz + xw); m[8] = 1.0f - 2.0f * (xx + yy);
	Vec3 c[4];
	c[0] = rot_apply(m, mn.x, mn.y, mx.z);
	c[1] = rot_apply(m, mn.x, mx.y, mx.z);
	c[2] = rot_apply(m, mx.x, mx.y, mn.z);
	c[3] = rot_apply(m, mx.x, mn.y, mn.z);
	Box o;
	o.mn = rot_apply(m, mn.x, mn.y, mn.z);
	o.mx = rot_apply(m, mx.x, mx.y, mx.z);
#pragma unroll
	for(int i = 0; i < 4; i++)
	{
		o.mn.x = (c[i].x < o.mn.x) ? c[i].x : o.mn.x;
		o.mn.y = (c[i].y < o.mn.y) ? c[i].y : o.mn.y;
		o.mn.z = (c[i].z < o.mn.z) ? c[i].z : o.mn.z;
		o.mx.x = (c[i].x > o.mx.x) ? c[i].x : o.mx.x;
		o.mx.y = (c[i].y > o.mx.y) ? c[i].y : o.mx.y;
		o.mx.z = (c[i].z > o.mx.z) ? c[i].z : o.mx.z;
	}
	return o;
}

} // namespace rn
