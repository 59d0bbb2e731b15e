/*
 * TEST INFRASTRUCTURE - NOT PART OF THE PRODUCT.
 *
 * rn_oracle.c: a plain-C, scalar, CPU restatement of the reference's scene hot path (uberpixel/Rayne @ 451a96e5),
 * written function by function from the reference sources cited below.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline leg may load it; librn_cuda never links or calls it.
 *
 * Parity pin: the reference has NO tests, golden vectors or fixtures for this path (Tests/ covers only
 * Objects/String/Number/KVO and is disabled, Tests/CMakeLists.txt:6-7).  This restatement is pinned instead
 * against the reference ITSELF: oracle/_ref/ref_harness_strict links the unmodified RN::SceneNode / RN::Camera /
 * RN::SceneBasic / RN::GenericParticleEmitter built from /root/reference (oracle/build_ref.sh) and
 * tests/test_oracle_vs_reference.py + tests/golden/ compare the two bit for bit.
 *
 * Build: gcc -O2 -ffp-contract=off -fPIC -shared (see oracle/build_oracle.py).  No fast-math: every expression is
 * evaluated in the reference's order with one rounding per operation.
 */
#include <math.h>
#include <float.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define RN_NO_PARENT 0xffffffffu
#define NODE_HIDDEN (1u << 1)
#define NODE_NOCULLING (1u << 2)
#define CAM_ORTHOGONAL (1u << 21)
#define CAM_SIMPLE (1u << 26)

typedef struct { float x, y, z; } vec3;
typedef struct { float x, y, z, w; } quat;
typedef struct { float m[16]; } mat4;

/* ---- Source/Math -------------------------------------------------------------------------------------------- */

/* Quaternion::operator*= , Source/Math/RNQuaternion.h:64-74 */
static quat quat_mul(quat a, quat o)
{
	quat r;
	r.w = -a.x * o.x - a.y * o.y - a.z * o.z + a.w * o.w;
	r.x = a.x * o.w + a.y * o.z - a.z * o.y + a.w * o.x;
	r.y = -a.x * o.z + a.y * o.w + a.z * o.x + a.w * o.y;
	r.z = a.x * o.y - a.y * o.x + a.z * o.w + a.w * o.z;
	return r;
}

/* Quaternion::GetRotatedVector, Source/Math/RNQuaternion.h:454-459 (conjugate :433-436) */
static vec3 quat_rotate(quat q, vec3 v)
{
	quat vq = {v.x, v.y, v.z, 0.0f};
	quat conj = {-q.x, -q.y, -q.z, q.w};
	quat r = quat_mul(quat_mul(q, vq), conj);
	vec3 out = {r.x, r.y, r.z};
	return out;
}

static mat4 mat_identity(void)
{
	mat4 r;
	memset(&r, 0, sizeof(r));
	r.m[0] = r.m[5] = r.m[10] = r.m[15] = 1.0f;
	return r;
}

/* Matrix::operator* / operator*=, Source/Math/RNMatrix.h:32-142 */
static mat4 mat_mul(const mat4 *a, const mat4 *o)
{
	mat4 r;
	for(int c = 0; c < 4; c++)
		for(int i = 0; i < 4; i++)
			r.m[c * 4 + i] = a->m[i] * o->m[c * 4] + a->m[4 + i] * o->m[c * 4 + 1] + a->m[8 + i] * o->m[c * 4 + 2] + a->m[12 + i] * o->m[c * 4 + 3];
	return r;
}

/* Quaternion::GetRotationMatrix, Source/Math/RNQuaternion.h:468-493 */
static mat4 quat_matrix(quat q)
{
	mat4 r = mat_identity();
	float xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z, xy = q.x * q.y, xz = q.x * q.z, xw = q.x * q.w, yz = q.y * q.z,
	      yw = q.y * q.w, zw = q.z * q.w;
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

/* Matrix::WithTranslation; Rotate; Scale - SceneNode::UpdateInternalTransformData, Source/Scene/RNSceneNode.cpp:530-532 */
static mat4 local_matrix(vec3 p, quat q, vec3 s)
{
	mat4 t = mat_identity();
	t.m[12] = p.x; t.m[13] = p.y; t.m[14] = p.z;
	mat4 r = quat_matrix(q);
	t = mat_mul(&t, &r);
	mat4 sc = mat_identity();
	sc.m[0] = s.x; sc.m[5] = s.y; sc.m[10] = s.z;
	return mat_mul(&t, &sc);
}

/* Matrix::GetDeterminant / GetSubmatrixDeterminant / GetInverse, Source/Math/RNMatrix.h:245-257,438-465 */
static float mat_det(const float *m)
{
	float det = m[0] * (m[5] * m[10] * m[15] + m[6] * m[11] * m[13] + m[7] * m[9] * m[14] - m[7] * m[10] * m[13] - m[6] * m[9] * m[15] - m[5] * m[11] * m[14]);
	det -= m[1] * (m[4] * m[10] * m[15] + m[6] * m[11] * m[12] + m[7] * m[8] * m[14] - m[7] * m[10] * m[12] - m[6] * m[8] * m[15] - m[4] * m[11] * m[14]);
	det += m[2] * (m[4] * m[9] * m[15] + m[5] * m[11] * m[12] + m[7] * m[8] * m[13] - m[7] * m[9] * m[12] - m[5] * m[8] * m[15] - m[4] * m[11] * m[13]);
	det -= m[3] * (m[4] * m[9] * m[14] + m[5] * m[10] * m[12] + m[6] * m[8] * m[13] - m[6] * m[9] * m[12] - m[5] * m[8] * m[14] - m[4] * m[10] * m[13]);
	return det;
}
static float mat_subdet(const float *m, int k)
{
	int i = k % 4, j = k / 4;
	int i1 = (i == 0), i2 = i1 + 1 + (i == 1), i3 = i2 + 1 + (i == 2);
	int j1 = (j == 0), j2 = j1 + 1 + (j == 1), j3 = j2 + 1 + (j == 2);
	i1 *= 4; i2 *= 4; i3 *= 4;
	float det = m[i1 + j1] * m[i2 + j2] * m[i3 + j3] + m[i1 + j2] * m[i2 + j3] * m[i3 + j1] + m[i1 + j3] * m[i2 + j1] * m[i3 + j2]
	          - m[i1 + j3] * m[i2 + j2] * m[i3 + j1] - m[i1 + j1] * m[i2 + j3] * m[i3 + j2] - m[i1 + j2] * m[i2 + j1] * m[i3 + j3];
	det *= 1 - 2 * ((i + j) % 2);
	return det;
}
static mat4 mat_inverse(const mat4 *a)
{
	mat4 r;
	float det = mat_det(a->m);
	for(int i = 0; i < 16; i++) r.m[i] = mat_subdet(a->m, i) / det;
	return r;
}

/* Vector3::Normalize, Source/Math/RNVector.h:519-530; Math::InverseSqrt, Source/Math/RNMath.cpp:21-24 */
static vec3 vec_normalize(vec3 v)
{
	if(v.x * v.x + v.y * v.y + v.z * v.z > FLT_EPSILON)
	{
		float inv = 1.0f * (float)(1.0 / sqrtf(v.x * v.x + v.y * v.y + v.z * v.z));
		v.x *= inv; v.y *= inv; v.z *= inv;
	}
	return v;
}
static float vec_dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; } /* RNVector.h:489-492 */
static vec3 vec_cross(vec3 a, vec3 b)                                               /* RNVector.h:494-503 */
{
	vec3 r = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
	return r;
}
static float fmax_std(float a, float b) { return (a < b) ? b : a; } /* std::max */
static float fmin_std(float a, float b) { return (b < a) ? b : a; } /* std::min */

/* ---- ordering: SceneBasic::AddNode / AddRenderNode ---------------------------------------------------------- */

/* Source/Scene/RNSceneBasic.cpp:652-676,683-736; Source/Scene/RNSceneNode.cpp:270-294.
 * parent[] in creation order (-1 = root, parent index < child index), roots are added in creation order after
 * their trees are complete.  order[rank] = creation index.  Recursive like the reference (depth-bounded inputs). */
typedef struct
{
	const int32_t *prio;
	uint32_t *next, *prev, head, tail;
	const uint32_t *child_ofs, *child;
} order_ctx;

static void add_render_node(order_ctx *c, uint32_t node)
{
	const uint32_t NIL = 0xffffffffu;
	int32_t p = c->prio ? c->prio[node] : 0;
	if(c->head == NIL || (c->prio ? c->prio[c->head] : 0) >= p)
	{
		c->next[node] = c->head; c->prev[node] = NIL;
		if(c->head != NIL) c->prev[c->head] = node; else c->tail = node;
		c->head = node;
	}
	else if((c->prio ? c->prio[c->tail] : 0) <= p)
	{
		c->prev[node] = c->tail; c->next[node] = NIL;
		c->next[c->tail] = node;
		c->tail = node;
	}
	else
	{
		uint32_t m = c->tail;
		while(c->prio[m] > p) m = c->prev[m];
		uint32_t after = c->next[m];
		c->next[m] = node; c->prev[node] = m; c->next[node] = after;
		if(after != NIL) c->prev[after] = node; else c->tail = node;
	}
}
static void add_node(order_ctx *c, uint32_t node)
{
	for(uint32_t k = c->child_ofs[node]; k < c->child_ofs[node + 1]; k++) add_node(c, c->child[k]); /* UpdateSceneInfo */
	add_render_node(c, node);
}

int rn_oracle_render_order(uint32_t n, const int32_t *parent, const int32_t *prio, uint32_t *order)
{
	uint32_t *ofs = calloc(n + 2, 4), *child = malloc(4 * (n + 1)), *fill = malloc(4 * (n + 1));
	order_ctx c;
	c.prio = prio;
	c.next = malloc(4 * (n + 1)); c.prev = malloc(4 * (n + 1));
	c.head = c.tail = 0xffffffffu;
	for(uint32_t i = 0; i < n; i++) if(parent[i] >= 0) ofs[parent[i] + 1]++;
	for(uint32_t i = 0; i < n; i++) ofs[i + 1] += ofs[i];
	memcpy(fill, ofs, 4 * (n + 1));
	for(uint32_t i = 0; i < n; i++) if(parent[i] >= 0) child[fill[parent[i]]++] = i;
	c.child_ofs = ofs; c.child = child;
	for(uint32_t i = 0; i < n; i++) if(parent[i] < 0) add_node(&c, i);
	uint32_t r = 0;
	for(uint32_t i = c.head; i != 0xffffffffu; i = c.next[i]) order[r++] = i;
	free(ofs); free(child); free(fill); free(c.next); free(c.prev);
	return r == n ? 0 : -1;
}

/* The same order without walking the list (the walk of rule 3 makes rn_oracle_render_order quadratic when priorities are
 * scattered: 11 minutes for 1 M nodes).  Derived from the three rules above, not from other code:
 *   - every rule keeps the list sorted by priority (rule 1 inserts a value <= the head's at the head, rule 2 a value >= the
 *     tail's at the tail, rule 3 behind the last element whose priority is <= the new one), so head = smallest priority seen;
 *   - rule 1 (list empty, or smallest priority so far >= p) puts the node in FRONT of the nodes of its own priority,
 *     rules 2 and 3 put it BEHIND them.
 * Hence: nodes ascending by priority; inside one priority first the rule-1 nodes in reverse call order, then the others in
 * call order.  Call order = AddNode's: children (in child order) before their parent, roots in creation order.
 * tests/test_cpu_oracle_golden.py checks it against the literal walk on random forests; oracle.py uses it for large inputs. */
typedef struct { int32_t prio; uint32_t side; int64_t seq; uint32_t node; } order_key;
static int order_key_cmp(const void *a, const void *b)
{
	const order_key *x = a, *y = b;
	if(x->prio != y->prio) return x->prio < y->prio ? -1 : 1;
	if(x->side != y->side) return x->side < y->side ? -1 : 1;
	if(x->seq != y->seq) return x->seq < y->seq ? -1 : 1;
	return 0;
}
int rn_oracle_render_order_fast(uint32_t n, const int32_t *parent, const int32_t *prio, uint32_t *order)
{
	uint32_t *ofs = calloc(n + 2, 4), *child = malloc(4 * (n + 1)), *fill = malloc(4 * (n + 1));
	uint32_t *stack = malloc(4 * (n + 1)), *next_child = malloc(4 * (n + 1));
	order_key *keys = malloc(sizeof(order_key) * (n + 1));
	for(uint32_t i = 0; i < n; i++) if(parent[i] >= 0) ofs[parent[i] + 1]++;
	for(uint32_t i = 0; i < n; i++) ofs[i + 1] += ofs[i];
	memcpy(fill, ofs, 4 * (n + 1));
	for(uint32_t i = 0; i < n; i++) if(parent[i] >= 0) child[fill[parent[i]]++] = i;
	memcpy(next_child, ofs, 4 * (n + 1));
	int64_t calls = 0;
	int have = 0;
	int32_t smallest = 0;
	for(uint32_t root = 0; root < n; root++)
	{
		if(parent[root] >= 0) continue;
		uint32_t sp = 0;
		stack[sp++] = root;
		while(sp)
		{
			const uint32_t u = stack[sp - 1];
			if(next_child[u] < ofs[u + 1]) { stack[sp++] = child[next_child[u]++]; continue; }   /* UpdateSceneInfo: children first */
			sp--;
			const int32_t p = prio ? prio[u] : 0;
			const int front = !have || smallest >= p;                                               /* rule 1 */
			keys[calls].prio = p;
			keys[calls].side = front ? 0u : 1u;
			keys[calls].seq = front ? -calls : calls;
			keys[calls].node = u;
			if(!have || p < smallest) smallest = p;
			have = 1;
			calls++;
		}
	}
	int rc = calls == (int64_t)n ? 0 : -1;
	if(rc == 0)
	{
		qsort(keys, n, sizeof(order_key), order_key_cmp);
		for(uint32_t r = 0; r < n; r++) order[r] = keys[r].node;
	}
	free(ofs); free(child); free(fill); free(stack); free(next_child); free(keys);
	return rc;
}

/* ---- SceneNode world transforms ----------------------------------------------------------------------------- */

/* SceneNode::UpdateInternalData (Source/Scene/RNSceneNode.cpp:501-524) and UpdateInternalTransformData (:526-546).
 * Arrays are indexed by any ordering in which parent[i] (an index into the same arrays, RN_NO_PARENT for roots) is
 * resolved recursively like the reference's lazy evaluation; here: iterate until all ancestors are done. */
int rn_oracle_world(uint32_t n, const uint32_t *parent, const float *pos, const float *scale, const float *rot,
                    float *wpos, float *wrot, float *wscale, float *wmat)
{
	uint8_t *done = calloc(n ? n : 1, 1);
	uint32_t *stack = malloc(4 * (n + 1));
	for(uint32_t start = 0; start < n; start++)
	{
		uint32_t sp = 0;
		uint32_t u = start;
		while(!done[u]) { stack[sp++] = u; if(parent[u] == RN_NO_PARENT) break; u = parent[u]; }
		while(sp)
		{
			u = stack[--sp];
			vec3 p = {pos[u * 3], pos[u * 3 + 1], pos[u * 3 + 2]};
			vec3 s = {scale[u * 3], scale[u * 3 + 1], scale[u * 3 + 2]};
			quat q = {rot[u * 4], rot[u * 4 + 1], rot[u * 4 + 2], rot[u * 4 + 3]};
			mat4 local = local_matrix(p, q, s);
			if(parent[u] != RN_NO_PARENT)
			{
				uint32_t P = parent[u];
				vec3 pwp = {wpos[P * 3], wpos[P * 3 + 1], wpos[P * 3 + 2]};
				vec3 pws = {wscale[P * 3], wscale[P * 3 + 1], wscale[P * 3 + 2]};
				quat pwr = {wrot[P * 4], wrot[P * 4 + 1], wrot[P * 4 + 2], wrot[P * 4 + 3]};
				vec3 rv = quat_rotate(pwr, p);
				/* _parent->GetWorldPosition() + _parent->GetWorldScale() * rotated   (:509) */
				wpos[u * 3] = pwp.x + pws.x * rv.x; wpos[u * 3 + 1] = pwp.y + pws.y * rv.y; wpos[u * 3 + 2] = pwp.z + pws.z * rv.z;
				quat wr = quat_mul(pwr, q); /* :510, never renormalised */
				wrot[u * 4] = wr.x; wrot[u * 4 + 1] = wr.y; wrot[u * 4 + 2] = wr.z; wrot[u * 4 + 3] = wr.w;
				wscale[u * 3] = pws.x * s.x; wscale[u * 3 + 1] = pws.y * s.y; wscale[u * 3 + 2] = pws.z * s.z; /* :511 */
				mat4 pm;
				memcpy(pm.m, wmat + (size_t)P * 16, 64);
				mat4 w = mat_mul(&pm, &local); /* :537 */
				memcpy(wmat + (size_t)u * 16, w.m, 64);
			}
			else
			{
				memcpy(wpos + u * 3, pos + u * 3, 12);
				memcpy(wrot + u * 4, rot + u * 4, 16);
				memcpy(wscale + u * 3, scale + u * 3, 12);
				memcpy(wmat + (size_t)u * 16, local.m, 64);
			}
			done[u] = 1;
		}
	}
	free(done); free(stack);
	return 0;
}

/* ---- Camera -------------------------------------------------------------------------------------------------- */

typedef struct
{
	uint32_t flags, render_group;
	float pos[3], rot[4], scale[3];
	float fov, aspect, clip_near, clip_far;
	float ortho[4];         /* left right bottom top */
	float plane_offsets[4]; /* top bottom left right */
	uint32_t has_projection;
	float projection[16];
} oracle_camera;

typedef struct
{
	float center[3], radius, planes[6][4];
	uint32_t flags, render_group;
} oracle_frustum;

static vec3 to_world(const mat4 *m, float x, float y, float z)
{
	/* Camera::__ToWorld, Source/Scene/RNCamera.cpp:386-396; Matrix * Vector4 Source/Math/RNMatrix.h:226-232 */
	float vx = m->m[0] * x + m->m[4] * y + m->m[8] * z + m->m[12] * 1.0f;
	float vy = m->m[1] * x + m->m[5] * y + m->m[9] * z + m->m[13] * 1.0f;
	float vz = m->m[2] * x + m->m[6] * y + m->m[10] * z + m->m[14] * 1.0f;
	float vw = m->m[3] * x + m->m[7] * y + m->m[11] * z + m->m[15] * 1.0f;
	vec3 r = {vx / vw, vy / vw, vz / vw};
	return r;
}

static void plane_triangle(vec3 p1, vec3 p2, vec3 p3, float dirfac, float offset, float out[4])
{
	/* Plane::WithTriangle + CalculateD, Source/Math/RNPlane.h:68-107 */
	vec3 d1 = {p2.x - p1.x, p2.y - p1.y, p2.z - p1.z}, d2 = {p3.x - p1.x, p3.y - p1.y, p3.z - p1.z};
	vec3 n = vec_cross(d1, d2);
	n.x *= dirfac; n.y *= dirfac; n.z *= dirfac;
	n = vec_normalize(n);
	out[0] = n.x; out[1] = n.y; out[2] = n.z;
	out[3] = vec_dot(n, p1) + offset;
}
static void plane_posnormal(vec3 p, vec3 n, float out[4])
{
	/* Plane::WithPositionNormal, Source/Math/RNPlane.h:56-66 */
	n = vec_normalize(n);
	out[0] = n.x; out[1] = n.y; out[2] = n.z;
	out[3] = vec_dot(n, p) + 0.0f;
}

/* Camera::PostUpdate (inverse view matrix) and UpdateProjection, Source/Scene/RNCamera.cpp:296-340 */
static void camera_matrices(const oracle_camera *cam, mat4 *inv_view_out, mat4 *proj_out)
{
	vec3 wp = {cam->pos[0], cam->pos[1], cam->pos[2]}, sc = {cam->scale[0], cam->scale[1], cam->scale[2]};
	quat wr = {cam->rot[0], cam->rot[1], cam->rot[2], cam->rot[3]};
	mat4 inv_view = local_matrix(wp, wr, sc); /* :303-305 */
	mat4 proj = mat_identity();
	if(cam->has_projection)
		memcpy(proj.m, cam->projection, 64);
	else if(cam->flags & CAM_ORTHOGONAL)
	{
		/* Matrix::WithProjectionOrthogonal, Source/Math/RNMatrix.h:339-364 */
		float left = cam->ortho[0], right = cam->ortho[1], bottom = cam->ortho[2], top = cam->ortho[3];
		float rl = right - left, tb = top - bottom, fn = cam->clip_far - cam->clip_near;
		float tx = -(right + left) / (right - left), ty = -(top + bottom) / (top - bottom);
		float tz = -(cam->clip_near) / (cam->clip_far - cam->clip_near);
		proj.m[0] = 2.0f / rl; proj.m[5] = 2.0f / tb; proj.m[10] = 1.0f / fn;
		proj.m[12] = tx; proj.m[13] = ty; proj.m[14] = (float)(-tz + 1.0); proj.m[15] = 1.0f;
	}
	else
	{
		/* Matrix::WithProjectionPerspective, Source/Math/RNMatrix.h:366-386 */
		float rad = cam->fov * 3.14159265358979323846264338327950288f / 180.0f;
		float h = cosf(0.5f * rad) / sinf(0.5f * rad);
		float w = h / cam->aspect;
		proj.m[0] = w; proj.m[5] = h;
		proj.m[10] = (float)(cam->clip_far / (cam->clip_far - cam->clip_near) - 1.0);
		proj.m[11] = -1.0f;
		proj.m[14] = (cam->clip_far * cam->clip_near) / (cam->clip_far - cam->clip_near);
		proj.m[15] = 0.0f;
	}
	*inv_view_out = inv_view;
	*proj_out = proj;
}

/* Camera::GetViewMatrix (= GetInverse of the inverse view matrix, RNCamera.cpp:307), GetProjectionMatrix and
 * projection * view (Modules/Vulkan/Sources/RNVulkanRenderer.cpp:1134-1168) */
void rn_oracle_camera_matrices(const oracle_camera *cam, float *view, float *projection, float *proj_view)
{
	mat4 inv_view, proj;
	camera_matrices(cam, &inv_view, &proj);
	mat4 v = mat_inverse(&inv_view), pv = mat_mul(&proj, &v);
	memcpy(view, v.m, 64); memcpy(projection, proj.m, 64); memcpy(proj_view, pv.m, 64);
}

/* Camera::PostUpdate / UpdateProjection / UpdateFrustum, Source/Scene/RNCamera.cpp:296-384 */
int rn_oracle_frustum(const oracle_camera *cam, oracle_frustum *out)
{
	memset(out, 0, sizeof(*out));
	out->flags = cam->flags;
	out->render_group = cam->render_group;
	if(cam->flags & CAM_SIMPLE)
	{
		out->center[2] = cam->clip_far * 0.5f;
		out->radius = (float)(cam->clip_far * 1.5);
		for(int k = 0; k < 6; k++) out->planes[k][1] = 1.0f; /* default Plane(), RNPlane.h:49-54 */
		return 0;
	}
	quat wr = {cam->rot[0], cam->rot[1], cam->rot[2], cam->rot[3]};
	vec3 wp = {cam->pos[0], cam->pos[1], cam->pos[2]};
	mat4 inv_view, proj;
	camera_matrices(cam, &inv_view, &proj);
	mat4 inv_proj = mat_inverse(&proj);
	mat4 tw = mat_mul(&inv_view, &inv_proj);
	vec3 pos1 = to_world(&tw, -1.0f, 1.0f, 1.0f), pos2 = to_world(&tw, -1.0f, 1.0f, 0.0f), pos3 = to_world(&tw, -1.0f, -1.0f, 0.0f);
	vec3 pos4 = to_world(&tw, 1.0f, -1.0f, 1.0f), pos5 = to_world(&tw, 1.0f, 1.0f, 0.0f), pos6 = to_world(&tw, 1.0f, -1.0f, 0.0f);
	vec3 fwd = {0.0f, 0.0f, -1.0f};
	vec3 dir = quat_rotate(wr, fwd);
#define MAX6(f) fmax_std(pos1.f, fmax_std(pos4.f, fmax_std(pos2.f, fmax_std(pos3.f, fmax_std(pos5.f, pos6.f)))))
#define MIN6(f) fmin_std(pos1.f, fmin_std(pos4.f, fmin_std(pos2.f, fmin_std(pos3.f, fmin_std(pos5.f, pos6.f)))))
	vec3 vmax = {MAX6(x), MAX6(y), MAX6(z)}, vmin = {MIN6(x), MIN6(y), MIN6(z)};
	vec3 c = {(vmax.x + vmin.x) * 0.5f, (vmax.y + vmin.y) * 0.5f, (vmax.z + vmin.z) * 0.5f};
	out->center[0] = c.x; out->center[1] = c.y; out->center[2] = c.z;
	float dx = c.x - vmax.x, dy = c.y - vmax.y, dz = c.z - vmax.z;
	out->radius = sqrtf(dx * dx + dy * dy + dz * dz);
	const float *o = cam->plane_offsets;
	plane_triangle(pos1, pos2, pos3, 1.0f, -o[2], out->planes[0]);
	plane_triangle(pos4, pos5, pos6, -1.0f, o[3], out->planes[1]);
	plane_triangle(pos1, pos2, pos5, -1.0f, o[0], out->planes[2]);
	plane_triangle(pos4, pos3, pos6, 1.0f, -o[1], out->planes[3]);
	float dn = fmin_std(cam->clip_near, cam->clip_far), df = fmax_std(cam->clip_near, cam->clip_far);
	vec3 pn = {wp.x + dir.x * dn, wp.y + dir.y * dn, wp.z + dir.z * dn}, pf = {wp.x + dir.x * df, wp.y + dir.y * df, wp.z + dir.z * df};
	vec3 ndir = {-dir.x, -dir.y, -dir.z};
	plane_posnormal(pn, ndir, out->planes[4]);
	plane_posnormal(pf, dir, out->planes[5]);
	return 0;
}

/* ---- culling ------------------------------------------------------------------------------------------------- */

/* SceneNode::UpdateInternalBoundsData sphere (Source/Scene/RNSceneNode.cpp:579-583, Source/Math/RNSphere.h:62-73),
 * SceneNode::CanRenderUtil (RNSceneNode.cpp:405-417), Camera::InFrustum (Source/Scene/RNCamera.cpp:440-473).
 * Node arrays are in render-list (rank) order; out_list receives the ranks of the visible nodes in that order. */
uint32_t rn_oracle_cull(uint32_t n, const float *wpos, const float *wrot, const float *wscale, const float *sphere,
                        const uint32_t *flags, const uint32_t *group, const oracle_frustum *cam, uint32_t *out_list,
                        float *out_sphere /* optional n*4 */)
{
	uint32_t count = 0;
	for(uint32_t i = 0; i < n; i++)
	{
		vec3 ws = {wscale[i * 3], wscale[i * 3 + 1], wscale[i * 3 + 2]};
		quat wr = {wrot[i * 4], wrot[i * 4 + 1], wrot[i * 4 + 2], wrot[i * 4 + 3]};
		float s = fmax_std(fmax_std(ws.x, ws.y), ws.z);
		float radius = sphere[i * 4 + 3] * s;
		vec3 off = {sphere[i * 4] * s, sphere[i * 4 + 1] * s, sphere[i * 4 + 2] * s};
		off = quat_rotate(wr, off);
		vec3 p = {wpos[i * 3] + off.x, wpos[i * 3 + 1] + off.y, wpos[i * 3 + 2] + off.z};
		if(out_sphere) { out_sphere[i * 4] = p.x; out_sphere[i * 4 + 1] = p.y; out_sphere[i * 4 + 2] = p.z; out_sphere[i * 4 + 3] = radius; }
		if(flags[i] & NODE_HIDDEN) continue;
		if(((group[i] & 0xffu) & (cam->render_group & 0xffu)) == 0) continue;
		if(!(flags[i] & NODE_NOCULLING))
		{
			float dx = cam->center[0] - p.x, dy = cam->center[1] - p.y, dz = cam->center[2] - p.z;
			if(sqrtf(dx * dx + dy * dy + dz * dz) > cam->radius + radius) continue;
			if(!(cam->flags & CAM_SIMPLE))
			{
				int inside = 1;
				for(int k = 0; k < 6 && inside; k++)
				{
					float d = (p.x * cam->planes[k][0] + p.y * cam->planes[k][1] + p.z * cam->planes[k][2]) - cam->planes[k][3];
					if(d > radius) inside = 0;
				}
				if(!inside) continue;
			}
		}
		out_list[count++] = i;
	}
	return count;
}

/* ---- A3: inverse world matrices ------------------------------------------------------------------------------ */

/* SceneNode::UpdateInternalInverseTransformData, Source/Scene/RNSceneNode.cpp:548-568:
 *   invLocal = WithScaling(scale != 0 ? 1/scale : 0); .Rotate(conj(rot)); .Translate(pos * -1)
 *   invWorld = invLocal * parent.invWorld
 * `scale != 0.0f` is Vector3::operator!=(Vector3(0)) - an epsilon compare of all three components
 * (Source/Math/RNVector.h:403-409, k::EpsilonFloat = FLT_EPSILON); Matrix::Translate Source/Math/RNMatrix.h:514-527. */
static mat4 inverse_local_matrix(vec3 p, quat q, vec3 s)
{
	vec3 is = {0.0f, 0.0f, 0.0f};
	if(!(fabsf(s.x - 0.0f) <= FLT_EPSILON && fabsf(s.y - 0.0f) <= FLT_EPSILON && fabsf(s.z - 0.0f) <= FLT_EPSILON))
	{
		is.x = 1.0f / s.x; is.y = 1.0f / s.y; is.z = 1.0f / s.z;
	}
	mat4 m = mat_identity();
	m.m[0] = is.x; m.m[5] = is.y; m.m[10] = is.z;
	quat c = {-q.x, -q.y, -q.z, q.w};
	mat4 r = quat_matrix(c);
	m = mat_mul(&m, &r);
	vec3 t = {p.x * -1.0f, p.y * -1.0f, p.z * -1.0f};
	float tmp[4];
	for(int i = 0; i < 4; i++)
		tmp[i] = m.m[i] * t.x + m.m[4 + i] * t.y + m.m[8 + i] * t.z + m.m[12 + i];
	for(int i = 0; i < 4; i++) m.m[12 + i] = tmp[i];
	return m;
}

int rn_oracle_inverse(uint32_t n, const uint32_t *parent, const float *pos, const float *scale, const float *rot, float *winv)
{
	uint8_t *done = calloc(n ? n : 1, 1);
	uint32_t *stack = malloc(4 * (n + 1));
	for(uint32_t start = 0; start < n; start++)
	{
		uint32_t sp = 0, u = start;
		while(!done[u]) { stack[sp++] = u; if(parent[u] == RN_NO_PARENT) break; u = parent[u]; }
		while(sp)
		{
			u = stack[--sp];
			vec3 p = {pos[u * 3], pos[u * 3 + 1], pos[u * 3 + 2]};
			vec3 s = {scale[u * 3], scale[u * 3 + 1], scale[u * 3 + 2]};
			quat q = {rot[u * 4], rot[u * 4 + 1], rot[u * 4 + 2], rot[u * 4 + 3]};
			mat4 il = inverse_local_matrix(p, q, s);
			if(parent[u] != RN_NO_PARENT)
			{
				mat4 pm;
				memcpy(pm.m, winv + (size_t)parent[u] * 16, 64);
				mat4 w = mat_mul(&il, &pm); /* :559 */
				memcpy(winv + (size_t)u * 16, w.m, 64);
			}
			else
				memcpy(winv + (size_t)u * 16, il.m, 64);
			done[u] = 1;
		}
	}
	free(done); free(stack);
	return 0;
}

/* ---- A4: transformed bounding box ---------------------------------------------------------------------------- */

/* Matrix::operator*(Vector3), Source/Math/RNMatrix.h:145-151 */
static vec3 mat_mul_vec(const mat4 *m, vec3 v)
{
	vec3 r;
	r.x = m->m[0] * v.x + m->m[4] * v.y + m->m[8] * v.z + m->m[12];
	r.y = m->m[1] * v.x + m->m[5] * v.y + m->m[9] * v.z + m->m[13];
	r.z = m->m[2] * v.x + m->m[6] * v.y + m->m[10] * v.z + m->m[14];
	return r;
}

/* SceneNode::UpdateInternalBoundsData box part (Source/Scene/RNSceneNode.cpp:576-579): position = world position;
 * AABB::operator*=(Vector3) Source/Math/RNAABB.h:108-114; AABB::SetRotation :152-177.
 * box: local minExtend xyz, maxExtend xyz (NULL = the default AABB(-1, 1), RNSceneNode.cpp:34); out: 6 floats. */
static void world_aabb(const float *box, vec3 ws, quat wr, float *out)
{
	vec3 mn = {-1.0f, -1.0f, -1.0f}, mx = {1.0f, 1.0f, 1.0f};
	if(box) { mn.x = box[0]; mn.y = box[1]; mn.z = box[2]; mx.x = box[3]; mx.y = box[4]; mx.z = box[5]; }
	mn.x *= ws.x; mn.y *= ws.y; mn.z *= ws.z;
	mx.x *= ws.x; mx.y *= ws.y; mx.z *= ws.z;
	mat4 m = quat_matrix(wr);
	vec3 corners[4];
	vec3 c0 = {mn.x, mn.y, mx.z}, c1 = {mn.x, mx.y, mx.z}, c2 = {mx.x, mx.y, mn.z}, c3 = {mx.x, mn.y, mn.z};
	corners[0] = mat_mul_vec(&m, c0); corners[1] = mat_mul_vec(&m, c1);
	corners[2] = mat_mul_vec(&m, c2); corners[3] = mat_mul_vec(&m, c3);
	mn = mat_mul_vec(&m, mn);
	mx = mat_mul_vec(&m, mx);
	for(int i = 0; i < 4; i++)
	{
		mn.x = (corners[i].x < mn.x) ? corners[i].x : mn.x;
		mn.y = (corners[i].y < mn.y) ? corners[i].y : mn.y;
		mn.z = (corners[i].z < mn.z) ? corners[i].z : mn.z;
		mx.x = (corners[i].x > mx.x) ? corners[i].x : mx.x;
		mx.y = (corners[i].y > mx.y) ? corners[i].y : mx.y;
		mx.z = (corners[i].z > mx.z) ? corners[i].z : mx.z;
	}
	out[0] = mn.x; out[1] = mn.y; out[2] = mn.z; out[3] = mx.x; out[4] = mx.y; out[5] = mx.z;
}

int rn_oracle_bounds(uint32_t n, const float *wrot, const float *wscale, const float *box /* n*6 or NULL */, float *out_aabb /* n*6 */)
{
	for(uint32_t i = 0; i < n; i++)
	{
		vec3 ws = {wscale[i * 3], wscale[i * 3 + 1], wscale[i * 3 + 2]};
		quat wr = {wrot[i * 4], wrot[i * 4 + 1], wrot[i * 4 + 2], wrot[i * 4 + 3]};
		world_aabb(box ? box + (size_t)i * 6 : NULL, ws, wr, out_aabb + (size_t)i * 6);
	}
	return 0;
}

/* Extension restated from the commented-out body of Camera::InFrustum(const AABB&), Source/Scene/RNCamera.cpp:479-496
 * (NO live reference behaviour: the live function returns true).  The comment tests, per plane, the box centre
 * against the plane with the box's projected half extent `width . |normal|`.  Restated on the LIVE plane convention
 * (Plane::GetDistance = p.n - d, a point is outside when the distance is positive, RNCamera.cpp:456-464):
 *     centre = position + (min + max) * 0.5 ; half = (max - min) * 0.5
 *     outside plane k  <=>  GetDistance_k(centre) > half.x*|n.x| + half.y*|n.y| + half.z*|n.z|
 * applied AFTER the reference's sphere predicate (a node must pass both).  Lists produced with it are a subset of
 * the reference's lists. */
uint32_t rn_oracle_cull_aabb(uint32_t n, const float *wpos, const float *aabb, const uint32_t *flags, const oracle_frustum *cam,
                             const uint32_t *in_list, uint32_t in_count, uint32_t *out_list)
{
	uint32_t count = 0;
	(void)n;
	for(uint32_t q = 0; q < in_count; q++)
	{
		uint32_t i = in_list[q];
		int inside = 1;
		if(!(flags[i] & NODE_NOCULLING) && !(cam->flags & CAM_SIMPLE))
		{
			const float *b = aabb + (size_t)i * 6;
			vec3 c = {wpos[i * 3] + (b[0] + b[3]) * 0.5f, wpos[i * 3 + 1] + (b[1] + b[4]) * 0.5f, wpos[i * 3 + 2] + (b[2] + b[5]) * 0.5f};
			vec3 h = {(b[3] - b[0]) * 0.5f, (b[4] - b[1]) * 0.5f, (b[5] - b[2]) * 0.5f};
			for(int k = 0; k < 6 && inside; k++)
			{
				float d = (c.x * cam->planes[k][0] + c.y * cam->planes[k][1] + c.z * cam->planes[k][2]) - cam->planes[k][3];
				float r = h.x * fabsf(cam->planes[k][0]) + h.y * fabsf(cam->planes[k][1]) + h.z * fabsf(cam->planes[k][2]);
				if(d > r) inside = 0;
			}
		}
		if(inside) out_list[count++] = i;
	}
	return count;
}


/* ---- particles ----------------------------------------------------------------------------------------------- */

typedef struct
{
	uint32_t seed, max_particles;
	float spawn_rate;
	float life_span[2];
	float start_color[4], end_color[4];
	float start_size[2], end_size[2], start_rotation[2], end_rotation[2];
	float gravity[3], velocity[3], vel_min[3], vel_max[3], pos_min[3], pos_max[3];
	uint32_t flags;                /* bit 0: !GetIsLocal(); bit 1: max_particles_soft set through SetMaxParticlesSoft */
	uint32_t max_particles_soft;
	float world_position[3], world_scale[3];
} oracle_emitter;

typedef struct
{
	uint32_t mt[624];
	uint32_t offset;
} oracle_mt;

/* MersenneTwister::Seed, Source/Math/RNRandom.cpp:313-323 */
static void mt_seed(oracle_mt *g, uint32_t seed)
{
	g->offset = 0;
	g->mt[0] = seed;
	for(uint32_t i = 1; i < 624; i++) g->mt[i] = 1812433253u * (g->mt[i - 1] ^ (g->mt[i - 1] >> 30)) + i;
}
/* MersenneTwister::GetRandomInt32, Source/Math/RNRandom.cpp:325-358 */
static int32_t mt_next(oracle_mt *g)
{
	uint32_t y, a;
	if(g->offset >= 624)
	{
		g->offset = 0;
		for(uint32_t i = 0; i < 623; i++)
		{
			y = (g->mt[i] & 0x80000000u) | (g->mt[i + 1] & 0x7fffffffu);
			a = (y & 1u) ? 0x9908b0dfu : 0u;
			g->mt[i] = g->mt[(i + 397) % 624] ^ (y >> 1) ^ a;
		}
		y = (g->mt[623] & 0x80000000u) | (g->mt[0] & 0x7fffffffu);
		a = (y & 1u) ? 0x9908b0dfu : 0u;
		g->mt[623] = g->mt[396] ^ (y >> 1) ^ a;
	}
	y = g->mt[g->offset++];
	y ^= (y >> 11);
	y ^= (y << 7) & 0x9d2c5680u;
	y ^= (y << 15) & 0xefc60000u;
	y ^= (y >> 18);
	return (int32_t)(y & 0x7fffffffu);
}
/* Generator::GetRandomFloatRange / GetUniformDeviate, Source/Math/RNRandom.cpp:213-222 */
static float rand_range(oracle_mt *g, float lo, float hi)
{
	float range = hi - lo;
	return (float)(((double)mt_next(g) * (1.0 / (2147483647 + 1.0))) * range + lo);
}

/* particle record: 24 floats, same layout as the harness dump (oracle/refio.py) */
enum { PX = 0, VX = 3, PT = 6, PL = 7, PD = 8, S0 = 9, S1 = 10, R0 = 11, R1 = 12, ROT = 13, SZ = 14, COL = 16, GR = 20, REC = 24 };

/* GenericParticle::Update, Source/Scene/RNParticle.cpp:31-53; Interpolator Linear Source/Math/RNInterpolation.h:101-103 */
static void particle_update(float *p, const oracle_emitter *e, float delta)
{
	p[PT] += delta;
	p[PL] -= delta;
	float t = p[PT], d = p[PD];
	p[ROT] = p[R0] + ((p[R1] - p[R0]) * t / d);
	for(int c = 0; c < 4; c++) p[COL + c] = e->start_color[c] + ((e->end_color[c] - e->start_color[c]) * t / d);
	p[SZ] = p[S0] + ((p[S1] - p[S0]) * t / d);
	p[SZ + 1] = p[SZ];
	for(int c = 0; c < 3; c++) p[VX + c] += p[GR + c] * delta;
	for(int c = 0; c < 3; c++) p[PX + c] += p[VX + c] * delta;
}

typedef struct
{
	oracle_emitter e;
	oracle_mt rng;
	float time;
	uint32_t count, cap;
	int grow;
	float *rec; /* cap * REC */
} oracle_emitter_state;

/* ParticleEmitter::SetMaxParticlesSoft, Source/Scene/RNParticleEmitter.cpp:128-132 */
void rn_oracle_emitter_set_soft_max(oracle_emitter_state *s, uint32_t k, uint32_t soft)
{
	s[k].e.flags |= 2u;
	s[k].e.max_particles_soft = soft;
}
/* the emitter node's world transform (read by CreateParticle when !GetIsLocal()) */
void rn_oracle_emitter_set_transform(oracle_emitter_state *s, uint32_t k, const float *world_position, const float *world_scale)
{
	if(world_position) memcpy(s[k].e.world_position, world_position, 12);
	if(world_scale) memcpy(s[k].e.world_scale, world_scale, 12);
}

oracle_emitter_state *rn_oracle_emitters_create(const oracle_emitter *e, uint32_t n)
{
	oracle_emitter_state *s = calloc(n, sizeof(*s));
	for(uint32_t i = 0; i < n; i++)
	{
		s[i].e = e[i];
		mt_seed(&s[i].rng, e[i].seed);
		s[i].cap = e[i].max_particles + 1;
		s[i].grow = 1;
		s[i].rec = malloc(sizeof(float) * REC * s[i].cap);
	}
	return s;
}
void rn_oracle_emitters_destroy(oracle_emitter_state *s, uint32_t n)
{
	for(uint32_t i = 0; i < n; i++) free(s[i].rec);
	free(s);
}

/* ParticleEmitter::UpdateParticles, Source/Scene/RNParticleEmitter.cpp:180-211;
 * GenericParticleEmitter::CreateParticle :416-454 (both branches of GetIsLocal()) */
void rn_oracle_emitters_step(oracle_emitter_state *s, uint32_t n, float delta)
{
	for(uint32_t k = 0; k < n; k++)
	{
		oracle_emitter_state *em = &s[k];
		const oracle_emitter *e = &em->e;
		uint32_t out = 0;
		for(uint32_t i = 0; i < em->count; i++)
		{
			float *p = em->rec + (size_t)i * REC;
			particle_update(p, e, delta);
			if(p[PL] <= 0.0f) continue; /* delete + erase, order preserved */
			if(out != i) memcpy(em->rec + (size_t)out * REC, p, sizeof(float) * REC);
			out++;
		}
		em->count = out;
		if(e->spawn_rate < FLT_EPSILON) continue;
		uint32_t spawn = (uint32_t)floorf((em->time + delta) / e->spawn_rate);
		em->time = fmodf((em->time + delta), e->spawn_rate);
		if(spawn > 0)
		{
			/* :204-205  if(_maxParticles > 0 && _maxParticlesSoft > 0) spawn = std::min(_maxParticlesSoft - size, spawn), uint32 */
			uint32_t soft = (e->flags & 2u) ? e->max_particles_soft : e->max_particles;
			if(e->max_particles > 0 && soft > 0)
			{
				uint32_t room = soft - em->count;
				spawn = room < spawn ? room : spawn;
			}
			for(uint32_t j = 0; j < spawn; j++)
			{
				if(em->count >= em->cap)
				{
					/* only reachable through the unsigned wrap above (count > soft): the vector just grows */
					em->cap *= 2;
					em->rec = realloc(em->rec, sizeof(float) * REC * em->cap);
				}
				float *p = em->rec + (size_t)em->count * REC;
				memset(p, 0, sizeof(float) * REC);
				for(int c = 0; c < 3; c++) p[PX + c] = rand_range(&em->rng, fmin_std(e->pos_min[c], e->pos_max[c]), fmax_std(e->pos_min[c], e->pos_max[c]));
				float lifespan = rand_range(&em->rng, e->life_span[0], e->life_span[1]);
				p[PL] = lifespan;
				for(int c = 0; c < 3; c++) p[GR + c] = e->gravity[c];
				for(int c = 0; c < 3; c++) p[VX + c] = e->velocity[c] + rand_range(&em->rng, fmin_std(e->vel_min[c], e->vel_max[c]), fmax_std(e->vel_min[c], e->vel_max[c]));
				float sizeScale = 1.0f;
				if(e->flags & 1u)
				{
					/* :428-437  !GetIsLocal() */
					const float *sc = e->world_scale;
					sizeScale = fmax_std(fmax_std(sc[0], sc[1]), sc[2]);
					for(int c = 0; c < 3; c++) p[PX + c] *= sc[c];
					for(int c = 0; c < 3; c++) p[PX + c] += e->world_position[c];
					for(int c = 0; c < 3; c++) p[GR + c] *= sc[c];
					for(int c = 0; c < 3; c++) p[VX + c] *= sc[c];
				}
				p[S0] = rand_range(&em->rng, e->start_size[0], e->start_size[1]) * sizeScale;
				p[S1] = rand_range(&em->rng, e->end_size[0], e->end_size[1]) * sizeScale;
				p[PD] = lifespan;
				p[R0] = rand_range(&em->rng, e->start_rotation[0], e->start_rotation[1]) * 0.01745329251994329577f;
				p[R1] = rand_range(&em->rng, e->end_rotation[0], e->end_rotation[1]) * 0.01745329251994329577f;
				p[PT] = 0.0f;
				particle_update(p, e, 0.0f);
				em->count++;
			}
		}
	}
}

uint32_t rn_oracle_emitter_read(const oracle_emitter_state *s, uint32_t k, float *rec_out, uint32_t cap, float *time_out)
{
	if(time_out) *time_out = s[k].time;
	uint32_t c = s[k].count < cap ? s[k].count : cap;
	if(rec_out) memcpy(rec_out, s[k].rec, sizeof(float) * REC * c);
	return s[k].count;
}

/* Extension P7 (no live reference behaviour): the comparator of the commented-out std::sort in
 * ParticleEmitter::Render, Source/Scene/RNParticleEmitter.cpp:364 - ascending distance to the camera position,
 * made stable (ties keep vector order).  idx_out[i] = index of the i-th nearest particle, key_out = distances. */
typedef struct { float key; uint32_t idx; } sort_item;
static int sort_cmp(const void *a, const void *b)
{
	const sort_item *x = a, *y = b;
	if(x->key < y->key) return -1;
	if(x->key > y->key) return 1;
	return x->idx < y->idx ? -1 : (x->idx > y->idx);
}
void rn_oracle_emitter_sort(const oracle_emitter_state *s, uint32_t k, const float cam[3], uint32_t *idx_out, float *key_out)
{
	uint32_t n = s[k].count;
	sort_item *it = malloc(sizeof(sort_item) * (n ? n : 1));
	for(uint32_t i = 0; i < n; i++)
	{
		const float *p = s[k].rec + (size_t)i * REC;
		float dx = p[PX] - cam[0], dy = p[PX + 1] - cam[1], dz = p[PX + 2] - cam[2];
		it[i].key = sqrtf(dx * dx + dy * dy + dz * dz);
		it[i].idx = i;
		if(key_out) key_out[i] = it[i].key;
	}
	qsort(it, n, sizeof(sort_item), sort_cmp);
	for(uint32_t i = 0; i < n; i++) idx_out[i] = it[i].idx;
	free(it);
}

/* ---- A10: SceneWithVisibilityLists default-volume order ------------------------------------------------------- */

/* SceneWithVisibilityLists::AddNode (Source/Scene/RNSceneWithVisibilityLists.cpp:210-236): UpdateSceneInfo adds the
 * children first (recursively, in child order, Source/Scene/RNSceneNode.cpp:286-291), then the node itself is
 * push_back-ed into _defaultVolume->nodes: post-order, no priority sort.  Nodes in creation order, parent[i] < i or -1,
 * children of a node in creation order, roots added in creation order. */
int rn_oracle_vislist_order(uint32_t n, const int32_t *parent, uint32_t *order)
{
	uint32_t *first = malloc(4 * (n + 1)), *next = malloc(4 * (n + 1)), *last = malloc(4 * (n + 1));
	uint32_t *stack = malloc(4 * (n + 1)), *cursor = malloc(4 * (n + 1));
	const uint32_t NIL = 0xffffffffu;
	for(uint32_t i = 0; i < n; i++) first[i] = next[i] = last[i] = NIL;
	for(uint32_t i = 0; i < n; i++)
	{
		if(parent[i] < 0) continue;
		uint32_t p = (uint32_t)parent[i];
		if(first[p] == NIL) first[p] = i; else next[last[p]] = i;
		last[p] = i;
	}
	uint32_t out = 0;
	for(uint32_t r = 0; r < n; r++)
	{
		if(parent[r] >= 0) continue;
		uint32_t sp = 0;
		stack[sp] = r; cursor[sp] = first[r]; sp++;
		while(sp)
		{
			uint32_t u = stack[sp - 1], c = cursor[sp - 1];
			if(c != NIL) { cursor[sp - 1] = next[c]; stack[sp] = c; cursor[sp] = first[c]; sp++; }
			else { order[out++] = u; sp--; }
		}
	}
	free(first); free(next); free(last); free(stack); free(cursor);
	return out == n ? 0 : -1;
}

/* ---- A11: post-cull sorts of SceneBasic::Render ---------------------------------------------------------------- */

/* Source/Scene/RNSceneBasic.cpp:576-616.  `list` holds indices into the per-node arrays (prio, wpos) in the order
 * the cull loop pushed them.  flags: bit 5 = Camera::Flags::SortFrontToBack... the caller passes the two booleans.
 * firstTransparentIndex == 0 doubles as "unset" (:583-587): a transparent node at list index 0 never becomes
 * `first`.  std::sort is unstable: equal keys may come out in any order; this restatement keeps them in list order
 * (a stable sort), which is one of the permitted outcomes.  Vector3::GetSquaredDistance Source/Math/RNVector.h:543-547. */
#define PRIO_SKY 200000
#define PRIO_TRANSPARENT 300000
typedef struct { float key; int32_t prio; uint32_t pos; uint32_t id; } list_item;
static float sq_dist(const float *wpos, uint32_t i, const float cam[3])
{
	float dx = wpos[i * 3] - cam[0], dy = wpos[i * 3 + 1] - cam[1], dz = wpos[i * 3 + 2] - cam[2];
	return dx * dx + dy * dy + dz * dz;
}
static int cmp_front_to_back(const void *a, const void *b)
{
	const list_item *x = a, *y = b;
	if(x->prio == y->prio && y->prio < PRIO_SKY) { if(x->key < y->key) return -1; if(y->key < x->key) return 1; }
	else { if(x->prio < y->prio) return -1; if(y->prio < x->prio) return 1; }
	return x->pos < y->pos ? -1 : (x->pos > y->pos ? 1 : 0);
}
static int cmp_back_to_front(const void *a, const void *b)
{
	const list_item *x = a, *y = b;
	if(x->key > y->key) return -1;
	if(y->key > x->key) return 1;
	return x->pos < y->pos ? -1 : (x->pos > y->pos ? 1 : 0);
}
int rn_oracle_sort_list(uint32_t *list, uint32_t count, const int32_t *prio, const float *wpos, const float cam[3],
                        int front_to_back, int transparent_back_to_front)
{
	size_t first = 0, last = 0;
	for(uint32_t q = 0; q < count; q++)
	{
		if(prio[list[q]] == PRIO_TRANSPARENT)
		{
			if(first == 0) first = q;
			last = q;
		}
	}
	list_item *it = malloc(sizeof(list_item) * (count ? count : 1));
	if(front_to_back)
	{
		for(uint32_t q = 0; q < count; q++) { it[q].id = list[q]; it[q].pos = q; it[q].prio = prio[list[q]]; it[q].key = sq_dist(wpos, list[q], cam); }
		qsort(it, count, sizeof(list_item), cmp_front_to_back);
		for(uint32_t q = 0; q < count; q++) list[q] = it[q].id;
	}
	if(transparent_back_to_front && first < last)
	{
		uint32_t m = (uint32_t)(last - first + 1);
		for(uint32_t q = 0; q < m; q++) { it[q].id = list[first + q]; it[q].pos = q; it[q].prio = 0; it[q].key = sq_dist(wpos, list[first + q], cam); }
		qsort(it, m, sizeof(list_item), cmp_back_to_front);
		for(uint32_t q = 0; q < m; q++) list[first + q] = it[q].id;
	}
	free(it);
	return 0;
}

/* ---- P6: ParticleEmitter::UpdateMesh --------------------------------------------------------------------------- */

/* Source/Scene/RNParticleEmitter.cpp:213-336.  Outputs (mesh capacity max = e.max_particles): pos[4*max*3],
 * color[4*max*4], uv0[4*max*2], uv1[4*max*2], idx[6*max] (uint16).  Quad k (k-th written) belongs to particle
 * i = k (or stop-1-k when rendered inversed) but its six indices are i*4 + {0,1,2,2,1,3} - the reference indexes
 * with the particle index, not with the write position (:283-288).  Unused index slots are degenerate (:322-330);
 * unused vertices are left untouched (the caller passes zeroed arrays, as the reference's calloc'ed buffers are).
 * world_scale_x = GetWorldScale().x of the emitter node (:223). */
void rn_oracle_emitter_mesh(const oracle_emitter_state *s, uint32_t k, float world_scale_x, int roll, int inversed,
                            float *pos, float *color, float *uv0, float *uv1, uint16_t *idx)
{
	const oracle_emitter_state *st = s + k;
	const int maxp = (int)st->e.max_particles;
	int stop = (int)st->count < maxp ? (int)st->count : maxp;
	int start = 0, inc = 1;
	if(inversed) { start = stop - 1; inc = -1; }
	static const float tex[8] = {0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 1.0f, 1.0f, 1.0f};
	uint32_t w = 0;
	for(int i = start; i >= 0 && i < stop; i += inc, w++)
	{
		const float *p = st->rec + (size_t)i * REC;
		for(int v = 0; v < 4; v++)
		{
			for(int c = 0; c < 3; c++) pos[(w * 4 + v) * 3 + c] = p[c];
			for(int c = 0; c < 4; c++) color[(w * 4 + v) * 4 + c] = p[16 + c];
			uv0[(w * 4 + v) * 2] = tex[v * 2]; uv0[(w * 4 + v) * 2 + 1] = tex[v * 2 + 1];
		}
		/* particle->size / 2.0f * scale */
		float hx = p[14] / 2.0f * world_scale_x, hy = p[15] / 2.0f * world_scale_x;
		float tx, ty, bx, by;
		if(!roll) { tx = hx; ty = hy; bx = hx; by = -hy; }
		else
		{
			float cs = cosf(p[13]), sn = sinf(p[13]);
			tx = cs * hx - sn * hy; ty = sn * hx + cs * hy;
			bx = cs * hx + sn * hy; by = sn * hx - cs * hy;
		}
		float *o = uv1 + (size_t)w * 8;
		o[0] = -tx; o[1] = -ty; o[2] = bx; o[3] = by; o[4] = -bx; o[5] = -by; o[6] = tx; o[7] = ty;
		static const int q[6] = {0, 1, 2, 2, 1, 3};
		for(int c = 0; c < 6; c++) idx[w * 6 + c] = (uint16_t)(i * 4 + q[c]);
	}
	for(int i = stop; i < maxp; i++)
		for(int c = 0; c < 6; c++) idx[(size_t)i * 6 + c] = (uint16_t)(i * 4);
}

/* ---- next row: per-drawable matrices --------------------------------------------------------------------------- */
/* (LOD selection, the other per-drawable step of Entity::Render, is compiled out in this reference:
 * RN_MODEL_LOD_DISABLED 1, Source/Rendering/RNModel.h:9 - there is no live behaviour to restate.) */

/* Drawable::Update (Source/Rendering/RNRenderer.h:65-76) + the matrix uniforms a renderer derives per drawable
 * (Modules/Vulkan/Sources/RNVulkanRenderer.cpp:1112-1168): modelMatrix = world, inverseModelMatrix = inverse world,
 * modelView = view * model, modelViewProjection = projectionView * model (Matrix::operator*).  64 floats per entry. */
void rn_oracle_instances(uint32_t count, const uint32_t *ids, const float *wmat, const float *winv, const float *view,
                         const float *projview, float *out)
{
	mat4 v, pv;
	memcpy(v.m, view, 64); memcpy(pv.m, projview, 64);
	for(uint32_t q = 0; q < count; q++)
	{
		uint32_t i = ids[q];
		mat4 m;
		memcpy(m.m, wmat + (size_t)i * 16, 64);
		mat4 a = mat_mul(&v, &m), p = mat_mul(&pv, &m);
		float *o = out + (size_t)q * 64;
		memcpy(o, m.m, 64);
		memcpy(o + 16, winv + (size_t)i * 16, 64);
		memcpy(o + 32, a.m, 64);
		memcpy(o + 48, p.m, 64);
	}
}

/* Vertex stage of the reference's particle shader, Shaders/Particles.hlsl:52-71 (particles_vertex), restated:
 *   clip = modelViewProjection * (float4(position, 1) + inverseModelView * float4(texCoords2, 0, 0)); colour * ambient.
 * HLSL's mul(M, v) with Rayne's column-major upload is the ordinary M * v; the order of the additions inside a
 * compiled shader is not defined by the source - this restatement sums left to right.  UNPINNED: the shader cannot
 * be executed here (no graphics stack), so no reference output exists for it. */
void rn_oracle_billboards(uint32_t n_vertices, const float *pos, const float *color, const float *uv1, const float *inv_mv,
                          const float *mvp, const float *ambient, float *clip_out, float *color_out)
{
	for(uint32_t v = 0; v < n_vertices; v++)
	{
		float tx = uv1[v * 2], ty = uv1[v * 2 + 1];
		float p[4];
		for(int i = 0; i < 4; i++)
		{
			float base = i < 3 ? pos[v * 3 + i] : 1.0f;
			p[i] = base + (inv_mv[i] * tx + inv_mv[4 + i] * ty);
		}
		for(int i = 0; i < 4; i++)
			clip_out[v * 4 + i] = mvp[i] * p[0] + mvp[4 + i] * p[1] + mvp[8 + i] * p[2] + mvp[12 + i] * p[3];
		for(int i = 0; i < 4; i++)
			color_out[v * 4 + i] = color[v * 4 + i] * ambient[i];
	}
}

/* ---- LOD selection (EXTENSION: compiled out in the reference, RN_MODEL_LOD_DISABLED, Source/Rendering/RNModel.h:9) ---- */
/* The #else branch of Entity::Render, Source/Scene/RNEntity.cpp:98-106:
 *     lodDistance = GetWorldPosition().GetDistance(distanceCamera->GetWorldPosition()) / distanceCamera->GetClipFar()
 * and Model::GetLODStageForDistance, Source/Rendering/RNModel.cpp:260-281.  Parity UNPINNED by nature (no live
 * reference behaviour); this restatement is what the CUDA kernel is checked against.
 * ids[e] indexes wpos / model_of_node; model 0xffff = not an Entity -> stage 0. */
void rn_oracle_select_lod(const uint32_t *ids, uint32_t n_entries, const float *wpos, const uint16_t *model_of_node,
                          const uint32_t *stage_offsets, const float *stage_distances, uint32_t n_models, const float cam[3],
                          float clip_far, uint8_t *out)
{
	for(uint32_t e = 0; e < n_entries; e++)
	{
		uint32_t id = ids[e], model = model_of_node[id], stage = 0;
		if(model < n_models)
		{
			uint32_t first = stage_offsets[model], count = stage_offsets[model + 1] - first;
			float dx = wpos[id * 3] - cam[0], dy = wpos[id * 3 + 1] - cam[1], dz = wpos[id * 3 + 2] - cam[2];
			float d = sqrtf(dx * dx + dy * dy + dz * dz);
			d /= clip_far;
			if(!(count <= 1 || d <= FLT_EPSILON))
			{
				stage = count - 1;
				for(uint32_t i = 0; i < count; i++)
					if(d <= stage_distances[first + i]) { stage = i; break; }
			}
		}
		out[e] = (uint8_t)stage;
	}
}

/* ---- C8: software occlusion culling of SceneBasic::Render ----------------------------------------------------- */
/* Source/Scene/RNSceneBasic.cpp:152-208 RasterizeClipSpaceTriangle, :210-318 RasterizeMesh, :320-405 TestBoundingBox,
 * :445-573 the occluder selection and the per-node test with its 50-frame hysteresis.  Matrices column-major
 * (RN::Matrix::m), products summed left to right (Source/Math/RNMatrix.h:90-142, 226-232).
 * The 40 x 40 depth buffer holds reverse-Z depth (1 = near), cleared to 0. */
enum { OCC_W = 40, OCC_H = 40 };

typedef struct { float x, y, z, w; } ovec4;

static ovec4 omat_vec(const float *m, ovec4 v)
{
	ovec4 r;
	r.x = m[0] * v.x + m[4] * v.y + m[8] * v.z + m[12] * v.w;
	r.y = m[1] * v.x + m[5] * v.y + m[9] * v.z + m[13] * v.w;
	r.z = m[2] * v.x + m[6] * v.y + m[10] * v.z + m[14] * v.w;
	r.w = m[3] * v.x + m[7] * v.y + m[11] * v.z + m[15] * v.w;
	return r;
}
static void omat_mul(const float *a, const float *b, float *o)
{
	for(int c = 0; c < 4; c++)
		for(int r = 0; r < 4; r++)
			o[c * 4 + r] = a[r] * b[c * 4] + a[4 + r] * b[c * 4 + 1] + a[8 + r] * b[c * 4 + 2] + a[12 + r] * b[c * 4 + 3];
}
/* edgeFunction, RNSceneBasic.cpp:27-30 */
static float oedge(float ax, float ay, float bx, float by, float cx, float cy)
{
	return (cx - ax) * (by - ay) - (cy - ay) * (bx - ax);
}
static ovec4 ovadd(ovec4 a, ovec4 b) { ovec4 r = {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; return r; }
static ovec4 ovsub(ovec4 a, ovec4 b) { ovec4 r = {a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; return r; }
static ovec4 ovscale(ovec4 a, float n) { ovec4 r = {a.x * n, a.y * n, a.z * n, a.w * n}; return r; }

/* RasterizeClipSpaceTriangle, :152-208 */
static void occ_raster_triangle(float *depthbuf, ovec4 A, ovec4 B, ovec4 C)
{
	ovec4 *v[3] = {&A, &B, &C};
	for(int k = 0; k < 3; k++)
	{
		ovec4 *p = v[k];
		float w = p->w;
		p->x /= w; p->y /= w; p->z /= w; p->w /= w;
		p->x = p->x * 0.5f + 0.5f;
		p->x *= (float)OCC_W;
		p->y = p->y * 0.5f + 0.5f;
		p->y *= (float)OCC_H;
	}
	float area = oedge(A.x, A.y, B.x, B.y, C.x, C.y);
	if(area < 0) return;
	uint16_t minX = (uint16_t)fmin_std(fmax_std(fmin_std(A.x, fmin_std(B.x, C.x)), 0.0f), (float)(OCC_W - 1));
	uint16_t minY = (uint16_t)fmin_std(fmax_std(fmin_std(A.y, fmin_std(B.y, C.y)), 0.0f), (float)(OCC_H - 1));
	uint16_t maxX = (uint16_t)fmin_std(fmax_std(fmax_std(A.x, fmax_std(B.x, C.x)), 0.0f), (float)(OCC_W - 1));
	uint16_t maxY = (uint16_t)fmin_std(fmax_std(fmax_std(A.y, fmax_std(B.y, C.y)), 0.0f), (float)(OCC_H - 1));
	for(uint16_t y = minY; y <= maxY; y++)
	{
		for(uint16_t x = minX; x <= maxX; x++)
		{
			float px = (float)x, py = (float)y;
			float w0 = oedge(B.x, B.y, C.x, C.y, px, py);
			float w1 = oedge(C.x, C.y, A.x, A.y, px, py);
			float w2 = oedge(A.x, A.y, B.x, B.y, px, py);
			if(w0 >= 0 && w1 >= 0 && w2 >= 0)
			{
				w0 /= area; w1 /= area; w2 /= area;
				float depth = w0 * A.z + w1 * B.z + w2 * C.z - 0.000001f;
				if(depth <= 1.0f)
				{
					float *cell = &depthbuf[OCC_W * (OCC_H - y - 1) + x];
					if(depth > *cell) *cell = depth;
				}
			}
		}
	}
}

/* RasterizeMesh, :210-318: tris = 9 floats per triangle as the mesh's triangle chunk iterator yields them (posA, posB, posC) */
static void occ_raster_mesh(float *depthbuf, const float *mvp, const float *tris, uint32_t ntris)
{
	for(uint32_t i = 0; i < ntris; i++)
	{
		const float *t = tris + (size_t)i * 9;
		ovec4 pa = {t[0], t[1], t[2], 1.0f}, pb = {t[3], t[4], t[5], 1.0f}, pc = {t[6], t[7], t[8], 1.0f};
		ovec4 A = omat_vec(mvp, pb);      /* the reference swaps A and B on purpose (:227-229) */
		ovec4 B = omat_vec(mvp, pa);
		ovec4 C = omat_vec(mvp, pc);
		if(A.z < -A.w && B.z < -B.w && C.z < -C.w) continue;
		if(A.z > A.w && B.z > B.w && C.z > C.w) continue;
		float dA = (A.z - (A.w)), dB = (B.z - (B.w)), dC = (C.z - (C.w));
		if(A.z > A.w && B.z > B.w)
		{
			float tA = dA / (dA - dC), tB = dB / (dB - dC);
			A = ovadd(A, ovscale(ovsub(C, A), tA));
			B = ovadd(B, ovscale(ovsub(C, B), tB));
		}
		else if(A.z > A.w && C.z > C.w)
		{
			float tA = dA / (dA - dB), tC = dC / (dC - dB);
			A = ovadd(A, ovscale(ovsub(B, A), tA));
			C = ovadd(C, ovscale(ovsub(B, C), tC));
		}
		else if(B.z > B.w && C.z > C.w)
		{
			float tB = dB / (dB - dA), tC = dC / (dC - dA);
			B = ovadd(B, ovscale(ovsub(A, B), tB));
			C = ovadd(C, ovscale(ovsub(A, C), tC));
		}
		else if(A.z > A.w)
		{
			float tB = dA / (dA - dB), tC = dA / (dA - dC);
			ovec4 AB = ovadd(A, ovscale(ovsub(B, A), tB));
			ovec4 AC = ovadd(A, ovscale(ovsub(C, A), tC));
			A = AB;
			occ_raster_triangle(depthbuf, C, AC, AB);
		}
		else if(B.z > B.w)
		{
			float tA = dB / (dB - dA), tC = dB / (dB - dC);
			ovec4 BA = ovadd(B, ovscale(ovsub(A, B), tA));
			ovec4 BC = ovadd(B, ovscale(ovsub(C, B), tC));
			B = BA;
			occ_raster_triangle(depthbuf, C, BA, BC);
		}
		else if(C.z > C.w)
		{
			float tA = dC / (dC - dA), tB = dC / (dC - dB);
			ovec4 CA = ovadd(C, ovscale(ovsub(A, C), tA));
			ovec4 CB = ovadd(C, ovscale(ovsub(B, C), tB));
			C = CA;
			occ_raster_triangle(depthbuf, B, CB, CA);
		}
		occ_raster_triangle(depthbuf, A, B, C);
	}
}

/* TestBoundingBox, :320-405.  aabb: world position xyz (the node's), minExtend xyz, maxExtend xyz */
static int occ_test_box(const float *depthbuf, const float *vp, const float *pos, const float *mn, const float *mx, const float pixel[2])
{
	const float ex[8][3] = {{mx[0], mx[1], mx[2]}, {mx[0], mx[1], mn[2]}, {mx[0], mn[1], mn[2]}, {mx[0], mn[1], mx[2]},
	                        {mn[0], mx[1], mx[2]}, {mn[0], mx[1], mn[2]}, {mn[0], mn[1], mn[2]}, {mn[0], mn[1], mx[2]}};
	float maxx = 0, maxy = 0, maxz = 0, minx = 0, miny = 0;
	for(int i = 0; i < 8; i++)
	{
		ovec4 c = {pos[0] + ex[i][0], pos[1] + ex[i][1], pos[2] + ex[i][2], 1.0f + 0.0f};
		c = omat_vec(vp, c);
		if(c.z > c.w) return 1;
		float w = c.w;
		c.x /= w; c.y /= w; c.z /= w; c.w /= w;
		c.x = c.x * 0.5f + 0.5f;
		c.y = c.y * 0.5f + 0.5f;
		if(i == 0)
		{
			maxx = minx = c.x; maxy = miny = c.y; maxz = c.z;
		}
		else
		{
			maxx = fmax_std(maxx, c.x); maxy = fmax_std(maxy, c.y); maxz = fmax_std(maxz, c.z);
			minx = fmin_std(minx, c.x); miny = fmin_std(miny, c.y);
		}
	}
	if((maxx - minx) < pixel[0] && (maxy - miny) < pixel[1]) return 0;
	minx *= (float)OCC_W; miny *= (float)OCC_H;
	maxx *= (float)OCC_W; maxy *= (float)OCC_H;
	uint16_t x0 = (uint16_t)fmax_std(fmin_std(floorf(minx) - 1.0f, (float)(OCC_W - 1)), 0.0f);
	uint16_t x1 = (uint16_t)fmax_std(fmin_std(ceilf(maxx) + 1.0f, (float)(OCC_W - 1)), 0.0f);
	uint16_t y0 = (uint16_t)fmax_std(fmin_std(floorf(miny) - 1.0f, (float)(OCC_H - 1)), 0.0f);
	uint16_t y1 = (uint16_t)fmax_std(fmin_std(ceilf(maxy) + 1.0f, (float)(OCC_H - 1)), 0.0f);
	for(uint16_t y = y0; y <= y1; y++)
		for(uint16_t x = x0; x <= x1; x++)
			if(maxz > depthbuf[OCC_W * (OCC_H - y - 1) + x]) return 1;
	return 0;
}

typedef struct
{
	float view[16], projection[16];   /* Camera::GetViewMatrix / GetProjectionMatrix */
	float offset[3];                  /* the random camera offset of this frame (:488) */
	float pixel[2];                   /* 1 / frame width, 1 / frame height (:486) */
	float cam_pos[3];                 /* camera->GetWorldPosition() */
} oracle_occlusion_params;

typedef struct { uint32_t id; float score, dist; } occ_cand;

/* The camera loop body of SceneBasic::Render with occluders present (:445-573) for ONE camera.
 *   n, order[n]        : render list, order[rank] = node id
 *   can_render[n]      : node->CanRender(renderer, camera) per id (hidden / group / frustum), as the cull computed it
 *   flags, priority    : per id (bit 3 = Occluder)
 *   wpos[n*3], radius[n] (GetBoundingSphere().radius), aabb[n*6] (GetBoundingBox(): min / max extents relative to wpos),
 *   wmat[n*16]         : world state per id
 *   mesh_of[n]         : mesh index of an Entity, -1 = no mesh;  tri_ofs / tris: triangle soups in model space
 *   counter[n]         : SceneBasicInfo::occludedFrameCounter, in/out
 * Writes the ids submitted for rendering in order to out_list (capacity n), the depth buffer (1600 floats) if asked,
 * returns the count; *used = 0 if no occluder passed CanRender (then the caller takes the plain cull list).
 * std::sort leaves equal keys in unspecified order; ties are broken by render-list rank here (stable). */
uint32_t rn_oracle_occlusion(uint32_t n, const uint32_t *order, const uint8_t *can_render, const uint32_t *flags, const int32_t *priority,
                             const float *wpos, const float *radius, const float *aabb, const float *wmat, const int32_t *mesh_of,
                             const uint32_t *tri_ofs, const float *tris, uint32_t *counter, const oracle_occlusion_params *p,
                             uint32_t *out_list, float *out_depth, int *used)
{
	occ_cand *occ = malloc(sizeof(occ_cand) * (n ? n : 1));
	uint32_t nocc = 0;
	for(uint32_t r = 0; r < n; r++)
	{
		uint32_t id = order[r];
		if((flags[id] & 8u) && can_render[id])
		{
			float dx = wpos[id * 3] - p->cam_pos[0], dy = wpos[id * 3 + 1] - p->cam_pos[1], dz = wpos[id * 3 + 2] - p->cam_pos[2];
			float d2 = dx * dx + dy * dy + dz * dz;       /* GetSquaredDistance: difference.GetDotProduct(difference) */
			occ[nocc].id = id;
			occ[nocc].dist = d2;
			occ[nocc].score = radius[id] * radius[id] / fmax_std(d2, 1.0f);
			nocc++;
		}
	}
	*used = nocc > 0;
	if(nocc == 0) { free(occ); return 0; }
	/* biggest 30 by approximated screen size (descending score), then front to back; insertion sorts keep ties in list order */
	for(uint32_t i = 1; i < nocc; i++)
	{
		occ_cand c = occ[i];
		uint32_t j = i;
		while(j > 0 && occ[j - 1].score < c.score) { occ[j] = occ[j - 1]; j--; }
		occ[j] = c;
	}
	if(nocc > 30) nocc = 30;
	for(uint32_t i = 1; i < nocc; i++)
	{
		occ_cand c = occ[i];
		uint32_t j = i;
		while(j > 0 && occ[j - 1].dist > c.dist) { occ[j] = occ[j - 1]; j--; }
		occ[j] = c;
	}
	float depthbuf[OCC_W * OCC_H];
	for(int i = 0; i < OCC_W * OCC_H; i++) depthbuf[i] = 0.0f;
	/* matViewProj = projection * WithTranslation(offset) * view */
	float tr[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, p->offset[0], p->offset[1], p->offset[2], 1};
	float pt[16], vp[16];
	omat_mul(p->projection, tr, pt);
	omat_mul(pt, p->view, vp);
	uint8_t *visible_occluder = calloc(n ? n : 1, 1);
	for(uint32_t k = 0; k < nocc; k++)
	{
		uint32_t id = occ[k].id;
		int res = occ_test_box(depthbuf, vp, wpos + id * 3, aabb + id * 6, aabb + id * 6 + 3, p->pixel);
		if(!res && counter[id] < 1000) counter[id] += 1;
		if(res || counter[id] < 50)
		{
			if(res) counter[id] = 0;
			visible_occluder[id] = 1;
		}
		if(res && mesh_of[id] >= 0)
		{
			float mvp[16];
			omat_mul(vp, wmat + (size_t)id * 16, mvp);
			uint32_t m = (uint32_t)mesh_of[id];
			occ_raster_mesh(depthbuf, mvp, tris + (size_t)tri_ofs[m] * 9, tri_ofs[m + 1] - tri_ofs[m]);
		}
	}
	uint32_t count = 0;
	for(uint32_t r = 0; r < n; r++)
	{
		uint32_t id = order[r];
		if(!can_render[id]) continue;
		if(priority[id] >= 200000) { out_list[count++] = id; continue; }          /* RenderSky and above are never tested */
		if((flags[id] & 8u) && visible_occluder[id]) { out_list[count++] = id; continue; }
		int res = occ_test_box(depthbuf, vp, wpos + id * 3, aabb + id * 6, aabb + id * 6 + 3, p->pixel);
		if(!res && counter[id] < 1000) counter[id] += 1;
		if(res || counter[id] < 50)
		{
			if(res) counter[id] = 0;
			out_list[count++] = id;
		}
	}
	if(out_depth) memcpy(out_depth, depthbuf, sizeof(depthbuf));
	free(occ);
	free(visible_occluder);
	return count;
}

/* The camera jitter of :488: GetRandomVector3Range(Vector3(-0.15, -0.15, 0), Vector3(0.15, 0.15, 0)) drawn n times from a
 * MersenneTwister seeded with `seed` (Source/Math/RNRandom.cpp:246-255: x, y, z in that order; the z draw is consumed
 * although its range is empty). */
void rn_oracle_jitter(uint32_t seed, uint32_t n_draws, float *out)
{
	oracle_mt g;
	mt_seed(&g, seed);
	for(uint32_t i = 0; i < n_draws; i++)
	{
		out[i * 3 + 0] = rand_range(&g, -0.15f, 0.15f);
		out[i * 3 + 1] = rand_range(&g, -0.15f, 0.15f);
		out[i * 3 + 2] = rand_range(&g, 0.0f, 0.0f);
	}
}
