/* oracle/drt_oracle.c -- TEST INFRASTRUCTURE ONLY: CPU restatement of the reference's hot path in plain C.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
 * library, and only as the CHECKER (or the timed CPU baseline) -- never on the product path.
 *
 * Parity status: PINNED. Every function below is checked bit-for-bit against the reference's own sources
 * compiled host-side (oracle/_ref/libdrt_ref.so, built by oracle/Makefile from /root/reference) in
 * tests/test_oracle.py, and against the golden vectors under tests/golden/ generated from that library
 * by tools/make_golden.py. The reference itself ships no tests or golden vectors.
 *
 * Each function cites the reference code it follows (paths relative to DustRayTracer/src/).
 * Arithmetic: fp32, one rounding per operation (compile with -ffp-contract=off, no fast-math), which is
 * what g++ produces for the reference's expressions. Where the reference leaves the evaluation order of
 * function ARGUMENTS open (Core/CudaMath/Random.cu:42-48,62-63) the `rtl` flag selects g++'s right-to-left
 * order (the one libdrt_ref.so exhibits) or nvcc's left-to-right order.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { float x, y, z; } f3;
typedef struct { float x, y; } f2;

static f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
static f3 add3(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
static f3 sub3(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
static f3 mul3(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
static f3 scale3(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
/* Core/CudaMath/helper_math.cuh:1264-1267 */
static float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
/* helper_math.cuh:1436-1439 */
static f3 cross3(f3 a, f3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
/* helper_math.cuh:1325-1329 with the HOST rsqrtf of :78-81 (1.0f / sqrtf(x)) */
static f3 normalize3(f3 v) { float inv = 1.0f / sqrtf(dot3(v, v)); return scale3(v, inv); }
/* helper_math.cuh:1307-1310 */
static float length3(f3 v) { return sqrtf(dot3(v, v)); }

/* ---------------- RNG: Core/CudaMath/Random.cu ---------------- */
/* Random.cu:6-11 */
uint32_t orc_pcg_hash(uint32_t input)
{
	uint32_t state = input * 747796405u + 2891336453u;
	uint32_t word = ((state >> ((state >> 28u) + 4u)) ^ state) * 277803737u;
	return (word >> 22u) ^ word;
}
/* Random.cu:13-17: (float)seed / (float)UINT32_MAX; the divisor rounds to 2^32 */
float orc_random_float(uint32_t* seed)
{
	*seed = orc_pcg_hash(*seed);
	return (float)*seed / (float)UINT32_MAX;
}
/* Random.cu:42-48 */
static f3 random_unit_vec3(uint32_t* seed, int rtl)
{
	float a = orc_random_float(seed) * 2.f - 1.f;
	float b = orc_random_float(seed) * 2.f - 1.f;
	float c = orc_random_float(seed) * 2.f - 1.f;
	return normalize3(rtl ? mk3(c, b, a) : mk3(a, b, c));
}
void orc_random_unit_vec3(uint32_t* seed, int rtl, float* out3)
{
	f3 r = random_unit_vec3(seed, rtl);
	out3[0] = r.x; out3[1] = r.y; out3[2] = r.z;
}
/* Random.cu:50-58 */
static f3 random_unit_sphere_vec3(uint32_t* seed, int rtl)
{
	for (;;) {
		f3 p = random_unit_vec3(seed, rtl);
		float len = length3(p);
		if ((len * len) < 1) return p;
	}
}
void orc_random_unit_sphere_vec3(uint32_t* seed, int rtl, float* out3)
{
	f3 r = random_unit_sphere_vec3(seed, rtl);
	out3[0] = r.x; out3[1] = r.y; out3[2] = r.z;
}
/* Random.cu:60-66 */
static f2 random_in_unit_disk(uint32_t* seed, int rtl)
{
	for (;;) {
		float a = orc_random_float(seed) * 2 - 1;
		float b = orc_random_float(seed) * 2 - 1;
		f2 p;
		p.x = rtl ? b : a; p.y = rtl ? a : b;
		if (p.x * p.x + p.y * p.y < 1.0f) return p;
	}
}

/* ---------------- data model (layouts measured from the reference headers) ---------------- */
typedef struct { f3 position; f3 normal; f2 uv; } Vertex;                 /* Core/Scene/Vertex.cuh:4-12, 32 B */
typedef struct { f3 centroid; float pad_; Vertex v0, v1, v2; f3 face_normal; int material; } Triangle; /* Triangle.cuh:7-19, 128 B */
typedef struct { f3 albedo; f3 emissive; int albedo_tex; float roughness; int transmission; float ior; int metallic; } Material; /* Material.cuh:4-23, 44 B */
typedef struct { int is_leaf; f3 pmin, pmax; int child1, child2, count, start; } BVHNode; /* BVH/BVHNode.cuh:14-44, 44 B (bool@0) */
typedef struct { int width, height, comps; unsigned char* data; } Texture;  /* Scene/Texture.cuh:4-20 */
typedef struct {
	float exposure, vfov_rad, zfar, znear, aspect, defocus_angle, focus_dist, speed;
	f3 position, forward, up, right; uint32_t viewheight, viewwidth;
} Camera; /* Scene/Camera.cuh:32-46, 88 B */
typedef struct {
	uint8_t gamma_correction, tone_mapping, enable_sunlight, pad0_;
	int max_samples, ray_bounce_limit, render_mode, debug_mode, pad1_;
	f2 sunlight_dir; f3 sunlight_color; float sunlight_intensity; f3 sky_color; float sky_intensity;
} Settings; /* Scene/RendererSettings.h:22-34, 64 B */

typedef char assert_tri[(sizeof(Triangle) == 128) ? 1 : -1];
typedef char assert_mat[(sizeof(Material) == 44) ? 1 : -1];
typedef char assert_node[(sizeof(BVHNode) == 44) ? 1 : -1];
typedef char assert_cam[(sizeof(Camera) == 88) ? 1 : -1];
typedef char assert_set[(sizeof(Settings) == 64) ? 1 : -1];

typedef struct {
	Triangle* prims; int* canonical; int n_prims;
	Material* materials; int n_materials;
	Texture* textures; int n_textures;
	BVHNode* nodes; int n_nodes, cap_nodes;
	int rtl;
	int bin_count, leaf_target;
} OrcScene;

typedef struct { f3 o, d, inv; float imin, imax; } Ray; /* Core/Ray.cuh:5-24 */
static Ray make_ray(f3 o, f3 d)
{
	Ray r; r.o = o; r.d = d;
	r.inv = mk3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
	r.imin = FLT_MAX; r.imax = -FLT_MAX; /* Core/Interval.cuh:9: default interval is empty */
	return r;
}

/* ---------------- Core/Bounds.cu:18-41 ---------------- */
static float bounds_intersect(f3 pmin, f3 pmax, const Ray* ray)
{
	f3 t0 = mul3(sub3(pmin, ray->o), ray->inv);
	f3 t1 = mul3(sub3(pmax, ray->o), ray->inv);
	f3 tmin = mk3(fminf(t0.x, t1.x), fminf(t0.y, t1.y), fminf(t0.z, t1.z));
	f3 tmax = mk3(fmaxf(t1.x, t0.x), fmaxf(t1.y, t0.y), fmaxf(t1.z, t0.z));
	float tenter = fmaxf(fmaxf(tmin.x, tmin.y), tmin.z);
	float texit = fminf(fminf(tmax.x, tmax.y), tmax.z);
	if (tenter < 0.0f) tenter = 0.0f;
	if (tenter > texit || texit < 0) return -1;
	return tenter;
}
float orc_bounds_intersect(const float* box6, const float* o3, const float* d3)
{
	Ray r = make_ray(mk3(o3[0], o3[1], o3[2]), mk3(d3[0], d3[1], d3[2]));
	return bounds_intersect(mk3(box6[0], box6[1], box6[2]), mk3(box6[3], box6[4], box6[5]), &r);
}

/* ---------------- Core/Kernel/Shaders/Intersection.cu:4-36 ---------------- */
#define TRIANGLE_EPSILON 0.000001f /* Common/physical_units.hpp:12 */
static int intersection(const Ray* ray, f3 p0, f3 p1, f3 p2, float* t_out, f3* uvw)
{
	f3 v0v1 = sub3(p1, p0), v0v2 = sub3(p2, p0);
	f3 pvec = cross3(ray->d, v0v2);
	float det = dot3(v0v1, pvec);
	if (det > -TRIANGLE_EPSILON && det < TRIANGLE_EPSILON) return 0;
	float inv_det = 1.0f / det;
	f3 tvec = sub3(ray->o, p0);
	float u = inv_det * dot3(tvec, pvec);
	if (u < 0.0f || u > 1.0f) return 0;
	f3 qvec = cross3(tvec, v0v1);
	float v = inv_det * dot3(ray->d, qvec);
	if (v < 0.0f || u + v > 1.0f) return 0;
	float t = inv_det * dot3(v0v2, qvec);
	if (t > TRIANGLE_EPSILON) { *t_out = t; *uvw = mk3(1.0f - u - v, u, v); return 1; }
	return 0;
}
int orc_intersection(const float* tri9, const float* o3, const float* d3, float* t, float* uvw3)
{
	Ray r = make_ray(mk3(o3[0], o3[1], o3[2]), mk3(d3[0], d3[1], d3[2]));
	f3 uvw;
	int hit = intersection(&r, mk3(tri9[0], tri9[1], tri9[2]), mk3(tri9[3], tri9[4], tri9[5]), mk3(tri9[6], tri9[7], tri9[8]), t, &uvw);
	if (hit) { uvw3[0] = uvw.x; uvw3[1] = uvw.y; uvw3[2] = uvw.z; }
	return hit;
}

/* ---------------- Core/Scene/Texture.cu:33-75 ---------------- */
static f3 texture_get_pixel(const Texture* tx, f2 uv)
{
	int x = (uv.x - floorf(uv.x)) * tx->width;
	int y = (uv.y - floorf(uv.y)) * tx->height;
	unsigned char r = 0, g = 0, b = 255;
	long idx = (long)y * tx->width + x, last = (long)tx->width * tx->height - 1;
	if (idx > last) idx = last; /* the reference reads one texel past the image here */
	if (idx < 0) idx = 0;
	if (tx->comps == 3) { const unsigned char* p = tx->data + 3 * idx; r = p[0]; g = p[1]; b = p[2]; }
	else if (tx->comps == 4) { const unsigned char* p = tx->data + 4 * idx; r = p[0]; g = p[1]; b = p[2]; }
	f3 c = mk3(r / (float)255, g / (float)255, b / (float)255);
	return mk3(powf(c.x, 2), powf(c.y, 2), powf(c.z, 2));
}
static float texture_get_alpha(const Texture* tx, f2 uv)
{
	if (tx->comps < 4) return 1;
	int x = (uv.x - floorf(uv.x)) * tx->width;
	int y = (uv.y - floorf(uv.y)) * tx->height;
	long idx = (long)y * tx->width + x, last = (long)tx->width * tx->height - 1;
	if (idx > last) idx = last;
	if (idx < 0) idx = 0;
	return tx->data[4 * idx + 3] / (float)255;
}
static f2 interp_uv(const Triangle* t, f3 uvw)
{
	f2 r;
	r.x = uvw.x * t->v0.uv.x + uvw.y * t->v1.uv.x + uvw.z * t->v2.uv.x;
	r.y = uvw.x * t->v0.uv.y + uvw.y * t->v1.uv.y + uvw.z * t->v2.uv.y;
	return r;
}

/* ---------------- Core/Kernel/Shaders/AnyHit.cuh:8-28 ---------------- */
static int any_hit(const OrcScene* sc, const Triangle* tri, f3 uvw)
{
	const Material* m = &sc->materials[tri->material];
	if (m->albedo_tex < 0) return 1;
	const Texture* tx = &sc->textures[m->albedo_tex];
	if (tx->comps < 4) return 1;
	float alpha = texture_get_alpha(tx, interp_uv(tri, uvw));
	return !(alpha < 1);
}

typedef struct { const Triangle* prim; float t; f3 uvw; f3 normal, position; int front_face; } Payload;

/* ---------------- Core/BVH/BVHTraversal.cuh:14-73 ---------------- */
static void traverse_bvh(const OrcScene* sc, const Ray* ray, int root, Payload* closest)
{
	if (root < 0) return;
	int idx_stack[64]; float dist_stack[64];
	uint8_t sp = 0;
	idx_stack[sp] = root;
	dist_stack[sp++] = bounds_intersect(sc->nodes[root].pmin, sc->nodes[root].pmax, ray);
	while (sp > 0) {
		const BVHNode* node = &sc->nodes[idx_stack[--sp]];
		float d = dist_stack[sp];
		if (!(ray->imin < d && d < ray->imax)) continue;          /* Interval::surrounds, :38 */
		if (closest->prim != NULL && closest->t < d) continue;      /* :41 */
		if (node->is_leaf & 0xff) {
			for (int i = node->start; i < node->start + node->count; i++) {
				const Triangle* tri = &sc->prims[i];
				float t; f3 uvw;
				if (intersection(ray, tri->v0.position, tri->v1.position, tri->v2.position, &t, &uvw) && t < closest->t) {
					if (!any_hit(sc, tri, uvw)) continue;
					closest->t = t; closest->prim = tri; closest->uvw = uvw;
				}
			}
		}
		else {
			float d1 = bounds_intersect(sc->nodes[node->child1].pmin, sc->nodes[node->child1].pmax, ray);
			float d2 = bounds_intersect(sc->nodes[node->child2].pmin, sc->nodes[node->child2].pmax, ray);
			if (d1 > d2) {
				if (d1 >= 0 && d1 < closest->t) { dist_stack[sp] = d1; idx_stack[sp++] = node->child1; }
				if (d2 >= 0 && d2 < closest->t) { dist_stack[sp] = d2; idx_stack[sp++] = node->child2; }
			}
			else {
				if (d2 >= 0 && d2 < closest->t) { dist_stack[sp] = d2; idx_stack[sp++] = node->child2; }
				if (d1 >= 0 && d1 < closest->t) { dist_stack[sp] = d1; idx_stack[sp++] = node->child1; }
			}
		}
	}
}

/* ---------------- Core/BVH/BVHTraversal.cuh:76-134 ---------------- */
static int traverse_bvh_raytest(const OrcScene* sc, const Ray* ray, int root)
{
	if (root < 0) return 0;
	int idx_stack[64];
	uint8_t sp = 0;
	idx_stack[sp++] = root;
	while (sp > 0) {
		int ni = idx_stack[--sp];
		const BVHNode* node = &sc->nodes[ni];
		if (ni == root) {
			if (bounds_intersect(node->pmin, node->pmax, ray) < 0) continue;
		}
		if (node->is_leaf & 0xff) {
			for (int i = node->start; i < node->start + node->count; i++) {
				const Triangle* tri = &sc->prims[i];
				float t; f3 uvw;
				if (intersection(ray, tri->v0.position, tri->v1.position, tri->v2.position, &t, &uvw)) {
					if (any_hit(sc, tri, uvw)) return 1;
				}
			}
		}
		else {
			float h1 = bounds_intersect(sc->nodes[node->child1].pmin, sc->nodes[node->child1].pmax, ray);
			float h2 = bounds_intersect(sc->nodes[node->child2].pmin, sc->nodes[node->child2].pmax, ray);
			if (h1 > h2) {
				if (h1 >= 0) idx_stack[sp++] = node->child1;
				if (h2 >= 0) idx_stack[sp++] = node->child2;
			}
			else {
				if (h2 >= 0) idx_stack[sp++] = node->child2;
				if (h1 >= 0) idx_stack[sp++] = node->child1;
			}
		}
	}
	return 0;
}

/* ---------------- Core/Kernel/TraceRay.cu:15-38 + Shaders/ClosestHit.cuh:4-28 + Miss.cuh ---------------- */
static Payload trace_ray(const OrcScene* sc, const Ray* ray)
{
	Payload p;
	memset(&p, 0, sizeof(p));
	p.prim = NULL;
	p.t = ray->imax;
	traverse_bvh(sc, ray, sc->n_nodes - 1, &p);
	if (p.prim == NULL) return p;
	p.position = add3(ray->o, scale3(ray->d, p.t));
	if (dot3(p.prim->face_normal, normalize3(ray->d)) > 0.f) { p.front_face = 0; p.normal = scale3(p.prim->face_normal, -1.f); }
	else { p.normal = p.prim->face_normal; p.front_face = 1; }
	return p;
}
static int ray_test(const OrcScene* sc, const Ray* ray) { return traverse_bvh_raytest(sc, ray, sc->n_nodes - 1); }

/* ---------------- Core/Scene/Camera.cu:82-129 ---------------- */
static float deg2rad(float degree) { const float PI = 3.14159265359f; return degree * (PI / 180.f); }
static Ray camera_get_ray(const Camera* c, f2 uv, float width, float height, uint32_t* seed, int rtl)
{
	float theta = c->vfov_rad / 2;
	float fov_factor = tanf(theta / 2.0f);
	float aspect_ratio = width / height;
	float plane_h = 2.0f * fov_factor * c->focus_dist;
	float plane_w = plane_h * aspect_ratio;
	f3 forward = normalize3(c->forward);
	f3 right = normalize3(cross3(forward, mk3(0, 1, 0)));
	f3 up = cross3(right, forward);
	f3 horiz = scale3(right, plane_w), vert = scale3(up, plane_h);
	float ox = orc_random_float(seed) - 0.5f;
	float oy = orc_random_float(seed) - 0.5f;
	ox *= 0.0035f; oy *= 0.0035f;
	float defocus_radius = c->focus_dist * tanf(deg2rad(c->defocus_angle) / 2.0f);
	f3 du = scale3(right, defocus_radius), dv = scale3(up, defocus_radius);
	f3 rorig;
	if (c->defocus_angle <= 0) rorig = c->position;
	else {
		f2 p = random_in_unit_disk(seed, rtl);
		rorig = add3(add3(c->position, scale3(du, p.x)), scale3(dv, p.y));
	}
	f3 dir = normalize3(add3(sub3(add3(add3(scale3(forward, c->focus_dist), scale3(horiz, uv.x + ox)), scale3(vert, uv.y + oy)), rorig), c->position));
	return make_ray(rorig, dir);
}

/* ---------------- Core/BVH/BVHBuilder.cu:100-346 + BVHBuilder.cuh:48-95 + Bounds.cu:4-10 ---------------- */
typedef struct { f3 pmin, pmax; int count, start; } ShallowNode;

static f3 absolute_extent(const Triangle* prims, const int* subset, int start, int end, f3* min_extent)
{
	/* BVHBuilder.cuh:48-71 (range form) and :72-95 (pointer-list form: subset != NULL) */
	f3 mn = mk3(FLT_MAX, FLT_MAX, FLT_MAX), mx = mk3(-FLT_MAX, -FLT_MAX, -FLT_MAX);
	for (int i = start; i < end; i++) {
		const Triangle* t = &prims[subset ? subset[i] : i];
		f3 p[3] = { t->v0.position, t->v1.position, t->v2.position };
		for (int k = 0; k < 3; k++) {
			mn.x = fminf(mn.x, p[k].x); mn.y = fminf(mn.y, p[k].y); mn.z = fminf(mn.z, p[k].z);
			mx.x = fmaxf(mx.x, p[k].x); mx.y = fmaxf(mx.y, p[k].y); mx.z = fmaxf(mx.z, p[k].z);
		}
	}
	*min_extent = mn;
	return mk3(mx.x - mn.x, mx.y - mn.y, mx.z - mn.z);
}
static float surface_area(f3 pmin, f3 pmax)
{
	/* Bounds.cu:4-10 */
	float planex = 2 * (pmax.z - pmin.z) * (pmax.y - pmin.y);
	float planey = 2 * (pmax.z - pmin.z) * (pmax.x - pmin.x);
	float planez = 2 * (pmax.x - pmin.x) * (pmax.y - pmin.y);
	return planex + planey + planez;
}
static float node_surface_area(const ShallowNode* n) { return n->count == 0 ? 0 : surface_area(n->pmin, n->pmax); } /* BVHNode.cuh:29-35 */
static float centroid_axis(const Triangle* t, int axis) { return axis == 0 ? t->centroid.x : (axis == 1 ? t->centroid.y : t->centroid.z); }

/* BVHBuilder.cu:216-255 */
static void bin_to_shallow_nodes(const OrcScene* sc, ShallowNode* left, ShallowNode* right, float bin, int axis, int start, int end, int* scratch)
{
	int nl = 0, nr = 0, n = end - start;
	int* lp = scratch; int* rp = scratch + n;
	for (int i = start; i < end; i++) {
		if (centroid_axis(&sc->prims[i], axis) < bin) lp[nl++] = i; else rp[nr++] = i;
	}
	f3 mn, ext;
	left->count = nl; ext = absolute_extent(sc->prims, lp, 0, nl, &mn); left->pmin = mn; left->pmax = add3(mn, ext);
	right->count = nr; ext = absolute_extent(sc->prims, rp, 0, nr, &mn); right->pmin = mn; right->pmax = add3(mn, ext);
}

/* std::partition as libstdc++ implements it for bidirectional iterators (the permutation matters:
 * it defines the order of triangles inside the reference's leaves). BVHBuilder.cu:175-214 */
static int partition_range(OrcScene* sc, int first, int last, int axis, float bin)
{
	for (;;) {
		for (;;) {
			if (first == last) return first;
			else if (centroid_axis(&sc->prims[first], axis) < bin) ++first;
			else break;
		}
		--last;
		for (;;) {
			if (first == last) return first;
			else if (!(centroid_axis(&sc->prims[last], axis) < bin)) --last;
			else break;
		}
		Triangle t = sc->prims[first]; sc->prims[first] = sc->prims[last]; sc->prims[last] = t;
		int c = sc->canonical[first]; sc->canonical[first] = sc->canonical[last]; sc->canonical[last] = c;
		++first;
	}
}

/* BVHBuilder.cu:257-346 (makePartition) + :175-214 (binToNodes) */
static void make_partition(OrcScene* sc, int start, int end, ShallowNode* leftnode, ShallowNode* rightnode, int* scratch)
{
	float best_bin = 0; int best_axis = 0; int lowest = 0x7fffffff;
	f3 mn; f3 extent = absolute_extent(sc->prims, NULL, start, end, &mn);
	f3 pmax = add3(mn, extent);
	float parent_sa = surface_area(mn, pmax);
	for (int axis = 0; axis < 3; axis++) {
		float ext = axis == 0 ? extent.x : (axis == 1 ? extent.y : extent.z);
		float base = axis == 0 ? mn.x : (axis == 1 ? mn.y : mn.z);
		float delta = ext / sc->bin_count;
		for (int i = 1; i < sc->bin_count; i++) {
			float bin = base + (i * delta);
			ShallowNode l, r;
			bin_to_shallow_nodes(sc, &l, &r, bin, axis, start, end, scratch);
			int cost = 1 + ((node_surface_area(&l) / parent_sa) * l.count * 2) + ((node_surface_area(&r) / parent_sa) * r.count * 2);
			if (cost < lowest) { lowest = cost; best_axis = axis; best_bin = bin; }
		}
	}
	int mid = partition_range(sc, start, end, best_axis, best_bin);
	f3 m2, e2;
	leftnode->start = start; leftnode->count = mid - start;
	e2 = absolute_extent(sc->prims, NULL, leftnode->start, leftnode->start + leftnode->count, &m2);
	leftnode->pmin = m2; leftnode->pmax = add3(m2, e2);
	rightnode->start = mid; rightnode->count = end - mid;
	e2 = absolute_extent(sc->prims, NULL, rightnode->start, rightnode->start + rightnode->count, &m2);
	rightnode->pmin = m2; rightnode->pmax = add3(m2, e2);
}

static int push_node(OrcScene* sc, const BVHNode* n)
{
	if (sc->n_nodes == sc->cap_nodes) {
		sc->cap_nodes = sc->cap_nodes ? sc->cap_nodes * 2 : 64;
		sc->nodes = (BVHNode*)realloc(sc->nodes, (size_t)sc->cap_nodes * sizeof(BVHNode));
	}
	sc->nodes[sc->n_nodes] = *n;
	return sc->n_nodes++;
}
static BVHNode node_from_shallow(const ShallowNode* s)
{
	BVHNode n; memset(&n, 0, sizeof(n));
	n.is_leaf = 0; n.pmin = s->pmin; n.pmax = s->pmax; n.child1 = -1; n.child2 = -1; n.count = s->count; n.start = s->start;
	return n;
}
/* BVHBuilder.cu:147-173; returns 0, or -1 when a partition left one side empty deeper than `depth` allows
 * (the reference recurses forever there) */
static int recursive_build(OrcScene* sc, BVHNode* node, int* scratch, int depth)
{
	if (node->count <= sc->leaf_target) { node->child1 = -1; node->child2 = -1; node->is_leaf = 1; return 0; }
	if (depth > 512) { node->is_leaf = 1; return -1; }
	ShallowNode l, r;
	make_partition(sc, node->start, node->start + node->count, &l, &r, scratch);
	BVHNode ln = node_from_shallow(&l), rn = node_from_shallow(&r);
	int rc = recursive_build(sc, &ln, scratch, depth + 1);
	rc |= recursive_build(sc, &rn, scratch, depth + 1);
	node->child1 = push_node(sc, &ln);
	node->child2 = push_node(sc, &rn);
	return rc;
}

/* ---------------- Core/Kernel/Shaders/RayGen.cuh:23-172 ---------------- */
static f3 tonemap_partial(f3 x)
{
	float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
	f3 r;
	r.x = ((x.x * (A * x.x + C * B) + D * E) / (x.x * (A * x.x + B) + D * F)) - E / F;
	r.y = ((x.y * (A * x.y + C * B) + D * E) / (x.y * (A * x.y + B) + D * F)) - E / F;
	r.z = ((x.z * (A * x.z + C * B) + D * E) / (x.z * (A * x.z + B) + D * F)) - E / F;
	return r;
}
static f3 uncharted2_filmic(f3 v, float exposure)
{
	f3 curr = tonemap_partial(scale3(v, exposure));
	f3 w = tonemap_partial(mk3(11.2f, 11.2f, 11.2f));
	f3 white_scale = mk3(1.0f / w.x, 1.0f / w.y, 1.0f / w.z);
	return mul3(curr, white_scale);
}
static f3 sky_model(const Ray* ray, const Settings* s)
{
	float g = 0.5 * (1 + (normalize3(ray->d)).y);
	f3 col1 = s->sky_color;
	f3 c = add3(scale3(mk3(1, 1, 1), (float)(1 - g)), scale3(col1, g));
	return mk3(powf(c.x, 2), powf(c.y, 2), powf(c.z, 2));
}
static f3 ray_gen(const OrcScene* sc, uint32_t x, uint32_t y, uint32_t max_x, uint32_t max_y, const Camera* cam,
	uint32_t frameidx, const Settings* S, unsigned long long* n_closest, unsigned long long* n_shadow)
{
	f2 uv; uv.x = ((float)x / max_x); uv.y = ((float)y / max_y);
	uv.x = uv.x * 2 - 1; uv.y = uv.y * 2 - 1;
	f3 sunpos = scale3(mk3(sinf(S->sunlight_dir.x) * (1 - sinf(S->sunlight_dir.y)), sinf(S->sunlight_dir.y),
		cosf(S->sunlight_dir.x) * (1 - sinf(S->sunlight_dir.y))), 100);
	f3 suncol = scale3(S->sunlight_color, S->sunlight_intensity);
	uint32_t seed = x + y * max_x;
	seed *= frameidx;
	Ray ray = camera_get_ray(cam, uv, (float)max_x, (float)max_y, &seed, sc->rtl);
	ray.imin = -1; ray.imax = FLT_MAX;
	f3 light = mk3(0, 0, 0), throughput = mk3(1, 1, 1);
	f2 tex_uv; tex_uv.x = 0; tex_uv.y = 1;
	for (int i = 0; i <= S->ray_bounce_limit; i++) {
		Payload p = trace_ray(sc, &ray);
		++*n_closest;
		seed += i;
		if (p.prim == NULL) {
			if (S->debug_mode == 4 && S->render_mode == 1) light = mk3(0, 0, 0); /* BVH heat colour: reference-tree specific, not restated */
			else light = add3(light, scale3(mul3(sky_model(&ray, S), throughput), S->sky_intensity));
			break;
		}
		const Material* m = &sc->materials[p.prim->material];
		if (m->albedo_tex < 0) throughput = mul3(throughput, m->albedo);
		else {
			tex_uv = interp_uv(p.prim, p.uvw);
			throughput = mul3(throughput, texture_get_pixel(&sc->textures[m->albedo_tex], tex_uv));
		}
		f3 new_o = add3(p.position, scale3(p.normal, 0.001f));
		if (S->enable_sunlight && S->render_mode == 0) {
			Ray sr = make_ray(new_o, add3(sunpos, scale3(random_unit_vec3(&seed, sc->rtl), 1.5f)));
			++*n_shadow;
			if (!ray_test(sc, &sr)) light = add3(light, mul3(suncol, throughput));
		}
		ray.o = new_o;
		ray.d = add3(p.normal, random_unit_sphere_vec3(&seed, sc->rtl));
		ray.inv = mk3(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
		if (S->render_mode == 1) {
			switch (S->debug_mode) {
			case 0: light = throughput; break;
			case 1: light = p.normal; break;
			case 2: light = p.uvw; break;
			case 3: light = mk3(tex_uv.x, tex_uv.y, 0); break;
			case 4: light = mk3(0, 0.1f, 0.1f); break;
			default: break;
			}
			break;
		}
	}
	if (S->render_mode == 0 || S->debug_mode == 0) {
		if (S->tone_mapping) light = uncharted2_filmic(light, cam->exposure);
		if (S->gamma_correction) light = mk3(sqrtf(light.x), sqrtf(light.y), sqrtf(light.z));
	}
	return light;
}

/* ---------------- C API (ctypes) ---------------- */
void* orc_scene_create(const void* triangles128, int n, const void* materials44, int nm, int rtl)
{
	OrcScene* sc = (OrcScene*)calloc(1, sizeof(OrcScene));
	sc->n_prims = n; sc->n_materials = nm; sc->rtl = rtl;
	sc->prims = (Triangle*)malloc((size_t)(n ? n : 1) * sizeof(Triangle));
	memcpy(sc->prims, triangles128, (size_t)n * sizeof(Triangle));
	sc->canonical = (int*)malloc((size_t)(n ? n : 1) * sizeof(int));
	for (int i = 0; i < n; i++) sc->canonical[i] = i;
	sc->materials = (Material*)malloc((size_t)(nm ? nm : 1) * sizeof(Material));
	memcpy(sc->materials, materials44, (size_t)nm * sizeof(Material));
	return sc;
}
void orc_scene_add_texture(void* h, int w, int hgt, int comps, const unsigned char* pixels)
{
	OrcScene* sc = (OrcScene*)h;
	sc->textures = (Texture*)realloc(sc->textures, (size_t)(sc->n_textures + 1) * sizeof(Texture));
	Texture* t = &sc->textures[sc->n_textures++];
	t->width = w; t->height = hgt; t->comps = comps;
	size_t n = (size_t)w * hgt * comps;
	t->data = (unsigned char*)malloc(n ? n : 1);
	memcpy(t->data, pixels, n);
}
void orc_scene_free(void* h)
{
	OrcScene* sc = (OrcScene*)h;
	if (!sc) return;
	for (int i = 0; i < sc->n_textures; i++) free(sc->textures[i].data);
	free(sc->textures); free(sc->prims); free(sc->canonical); free(sc->materials); free(sc->nodes); free(sc);
}
/* BVHBuilder::build (BVHBuilder.cu:100-145). Returns the node count, negative if the depth guard fired. */
int orc_scene_build_bvh(void* h, int bin_count, int leaf_target)
{
	OrcScene* sc = (OrcScene*)h;
	sc->bin_count = bin_count; sc->leaf_target = leaf_target; sc->n_nodes = 0;
	int n = sc->n_prims;
	BVHNode root; memset(&root, 0, sizeof(root));
	f3 mn; f3 ext = absolute_extent(sc->prims, NULL, 0, n, &mn);
	root.pmin = mn; root.pmax = add3(mn, ext); root.child1 = -1; root.child2 = -1; root.count = n; root.start = -1;
	if (n <= leaf_target) { root.is_leaf = 1; root.start = 0; push_node(sc, &root); return sc->n_nodes; }
	int* scratch = (int*)malloc((size_t)2 * n * sizeof(int));
	ShallowNode l, r;
	make_partition(sc, 0, n, &l, &r, scratch);
	BVHNode ln = node_from_shallow(&l), rn = node_from_shallow(&r);
	int rc = recursive_build(sc, &ln, scratch, 1);
	rc |= recursive_build(sc, &rn, scratch, 1);
	root.child1 = push_node(sc, &ln);
	root.child2 = push_node(sc, &rn);
	push_node(sc, &root);
	free(scratch);
	return rc ? -sc->n_nodes : sc->n_nodes;
}
int orc_scene_node_count(void* h) { return ((OrcScene*)h)->n_nodes; }
void orc_scene_get_nodes(void* h, void* out) { OrcScene* sc = (OrcScene*)h; memcpy(out, sc->nodes, (size_t)sc->n_nodes * sizeof(BVHNode)); }
void orc_scene_get_triangles(void* h, void* out) { OrcScene* sc = (OrcScene*)h; memcpy(out, sc->prims, (size_t)sc->n_prims * sizeof(Triangle)); }
void orc_scene_get_canonical_ids(void* h, int* out) { OrcScene* sc = (OrcScene*)h; memcpy(out, sc->canonical, (size_t)sc->n_prims * sizeof(int)); }

/* Primary rays as RayGen issues them (RayGen.cuh:65-78,90): canonical id (-1 miss), t, u = UVW.y, v = UVW.z */
void orc_trace_primary(void* h, const void* cam88, int W, int H, unsigned frameidx, int* prim_id, float* t, float* u, float* v, int nthreads)
{
	const OrcScene* sc = (const OrcScene*)h;
	Camera cam; memcpy(&cam, cam88, sizeof(cam));
#ifdef _OPENMP
	if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 4)
	for (int y = 0; y < H; y++)
		for (int x = 0; x < W; x++) {
			f2 uv; uv.x = ((float)x / (uint32_t)W) * 2 - 1; uv.y = ((float)y / (uint32_t)H) * 2 - 1;
			uint32_t seed = (uint32_t)x + (uint32_t)y * (uint32_t)W;
			seed *= frameidx;
			Ray ray = camera_get_ray(&cam, uv, (float)(uint32_t)W, (float)(uint32_t)H, &seed, sc->rtl);
			ray.imin = -1; ray.imax = FLT_MAX;
			Payload p = trace_ray(sc, &ray);
			size_t i = (size_t)x + (size_t)y * W;
			if (!p.prim) { prim_id[i] = -1; t[i] = -1; u[i] = 0; v[i] = 0; }
			else { prim_id[i] = sc->canonical[p.prim - sc->prims]; t[i] = p.t; u[i] = p.uvw.y; v[i] = p.uvw.z; }
		}
}

/* Camera rays as RayGen issues them for `frameidx` (RayGen.cuh:65-78 -> Camera.cu:82-123): out_od6[p] = origin xyz, direction xyz,
 * p = x + y*W. rtl = argument evaluation order of the g++-compiled reference (1) or nvcc's (0). */
void orc_camera_rays(const void* cam88, int W, int H, unsigned frameidx, int rtl, float* out_od6)
{
	Camera cam; memcpy(&cam, cam88, sizeof(cam));
	for (int y = 0; y < H; y++)
		for (int x = 0; x < W; x++) {
			f2 uv; uv.x = ((float)x / (uint32_t)W) * 2 - 1; uv.y = ((float)y / (uint32_t)H) * 2 - 1;
			uint32_t seed = (uint32_t)x + (uint32_t)y * (uint32_t)W;
			seed *= frameidx;
			Ray ray = camera_get_ray(&cam, uv, (float)(uint32_t)W, (float)(uint32_t)H, &seed, rtl);
			float* o = out_od6 + 6 * ((size_t)x + (size_t)y * W);
			o[0] = ray.o.x; o[1] = ray.o.y; o[2] = ray.o.z; o[3] = ray.d.x; o[4] = ray.d.y; o[5] = ray.d.z;
		}
}

/* Tree-free checker for big soups: closest hit of every ray over ALL triangles with the reference's Moller-Trumbore
 * (Intersection.cu:4-36, `intersection` above), no BVH, no box test. Among equal t the lowest triangle index wins (the rule the
 * CUDA path documents; the reference itself keeps the first one its own tree order meets, BVHTraversal.cuh:51).
 * tris9: n x (v0, v1, v2); rays: origin xyz + direction xyz. prim = -1 / t = -1 on a miss; u = UVW.y, v = UVW.z. */
void orc_brute_force_closest(const float* tris9, int n, const float* rays_od6, int nrays, int* prim, float* t, float* u, float* v, int nthreads)
{
#ifdef _OPENMP
	if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 16)
	for (int r = 0; r < nrays; r++) {
		const float* od = rays_od6 + 6 * (size_t)r;
		Ray ray = make_ray(mk3(od[0], od[1], od[2]), mk3(od[3], od[4], od[5]));
		int best = -1;
		float bt = FLT_MAX;
		f3 buvw = mk3(0, 0, 0);
		for (int i = 0; i < n; i++) {
			const float* p = tris9 + 9 * (size_t)i;
			float ti;
			f3 uvw;
			if (!intersection(&ray, mk3(p[0], p[1], p[2]), mk3(p[3], p[4], p[5]), mk3(p[6], p[7], p[8]), &ti, &uvw)) continue;
			if (ti < bt) { bt = ti; best = i; buvw = uvw; } /* ascending i: an equal t keeps the lower index */
		}
		prim[r] = best;
		t[r] = best >= 0 ? bt : -1.f;
		u[r] = best >= 0 ? buvw.y : 0.f;
		v[r] = best >= 0 ? buvw.z : 0.f;
	}
}

/* RenderKernel.cu:20-35 + Renderer.cu:80-117: accumulates frames [begin, begin+count) into acc (W*H*3).
 * Returns seconds in the pixel loops; counters[0] = TraceRay calls, counters[1] = RayTest calls. */
double orc_render(void* h, const void* cam88, const void* settings64, int W, int H, unsigned frame_begin, unsigned frame_count,
	float* acc, int nthreads, unsigned long long* counters)
{
	const OrcScene* sc = (const OrcScene*)h;
	Camera cam; memcpy(&cam, cam88, sizeof(cam));
	Settings S; memcpy(&S, settings64, sizeof(S));
	unsigned long long closest = 0, shadow = 0;
#ifdef _OPENMP
	if (nthreads > 0) omp_set_num_threads(nthreads);
	double t0 = omp_get_wtime();
#endif
	for (unsigned f = 0; f < frame_count; f++) {
		const uint32_t frameidx = frame_begin + f;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : closest, shadow)
		for (int j = 0; j < H; j++)
			for (int i = 0; i < W; i++) {
				f3 c = ray_gen(sc, (uint32_t)i, (uint32_t)j, (uint32_t)W, (uint32_t)H, &cam, frameidx, &S, &closest, &shadow);
				float* a = acc + 3 * ((size_t)i + (size_t)j * W);
				a[0] += c.x; a[1] += c.y; a[2] += c.z;
			}
	}
	if (counters) { counters[0] = closest; counters[1] = shadow; }
#ifdef _OPENMP
	return omp_get_wtime() - t0;
#else
	return 0.0;
#endif
}
int orc_max_threads(void)
{
#ifdef _OPENMP
	return omp_get_max_threads();
#else
	return 1;
#endif
}
