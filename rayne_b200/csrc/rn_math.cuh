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
__device__ __forceinline__ Affine local_matrix(const Vec3 &p, const Quat &q, const Vec3 &s)
{
	float xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z;
	float xy = q.x * q.y, xz = q.x * q.z, xw = q.x * q.w;
	float yz = q.y * q.z, yw = q.y * q.w, zw = q.z * q.w;
	float r0 = 1.0f - 2.0f * (yy + zz), r4 = 2.0f * (xy - zw), r8 = 2.0f * (xz + yw);
	float r1 = 2.0f * (xy + zw), r5 = 1.0f - 2.0f * (xx + zz), r9 = 2.0f * (yz - xw);
	float r2 = 2.0f * (xz - yw), r6 = 2.0f * (yz + xw), r10 = 1.0f - 2.0f * (xx + yy);
	Affine a;
	a.m[0] = r0 * s.x; a.m[1] = r1 * s.x; a.m[2] = r2 * s.x;
	a.m[3] = r4 * s.y; a.m[4] = r5 * s.y; a.m[5] = r6 * s.y;
	a.m[6] = r8 * s.z; a.m[7] = r9 * s.z; a.m[8] = r10 * s.z;
	a.m[9] = p.x; a.m[10] = p.y; a.m[11] = p.z;
	return a;
}

// Matrix::operator* (reference: Source/Math/RNMatrix.h:90-142) restricted to affine operands:
// r[c][r] = P[0][r]*L[c][0] + P[1][r]*L[c][1] + P[2][r]*L[c][2] (+ P[3][r] for the translation column),
// summed left to right like the reference.
__device__ __forceinline__ Affine affine_mul(const Affine &P, const Affine &L)
{
	Affine o;
#pragma unroll
	for(int c = 0; c < 3; c++)
	{
#pragma unroll
		for(int r = 0; r < 3; r++)
			o.m[c * 3 + r] = P.m[r] * L.m[c * 3] + P.m[3 + r] * L.m[c * 3 + 1] + P.m[6 + r] * L.m[c * 3 + 2];
	}
#pragma unroll
	for(int r = 0; r < 3; r++)
		o.m[9 + r] = P.m[r] * L.m[9] + P.m[3 + r] * L.m[10] + P.m[6 + r] * L.m[11] + P.m[9 + r];
	return o;
}

} // namespace rn
