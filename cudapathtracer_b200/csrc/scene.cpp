// Host scene container: the reference's globals `verts, tris, mats, lights, totalLightArea`
// (modelLoader.h:43-47) behind the C-ABI, filled with the arithmetic of loadOBJ (modelLoader.h:125-210).
// Compiled with -ffp-contract=off: vertex transform, normal and light-area values are bit-identical to the
// reference's (checked against oracle/_ref by tests/test_scene_ingest.py).
#include "host_common.h"
#include <cstring>

namespace cupt {

static thread_local std::string g_last_error;

int set_error(int code, const char* fmt, ...)
{
	char buf[1024];
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(buf, sizeof(buf), fmt, ap);
	va_end(ap);
	g_last_error = buf;
	return code;
}

// One shape of loadOBJ: modelLoader.h:162-206.  `face_mats` (per triangle) may be NULL, then `material` is
// used for every triangle, which is what the reference does with material_ids[0] (modelLoader.h:189).
static int append_shape(cupt_scene* s, const float* positions, int nverts, const uint32_t* indices, int ntris,
                        int material, const int32_t* face_mats, const float origin[3], float scale, int flip)
{
	const float ox = origin ? origin[0] : 0.f, oy = origin ? origin[1] : 0.f, oz = origin ? origin[2] : 0.f;
	const int32_t indexBufferOffset = (int32_t)s->verts.size();
	s->verts.reserve(s->verts.size() + (size_t)nverts);
	for (int v = 0; v < nverts; ++v) {  // modelLoader.h:167-174: p * scale + origin, two roundings
		cupt_vec3 p;
		p.x = positions[3 * v + 0] * scale + ox;
		p.y = positions[3 * v + 1] * scale + oy;
		p.z = positions[3 * v + 2] * scale + oz;
		s->verts.push_back(p);
	}
	s->tris.reserve(s->tris.size() + (size_t)ntris);
	for (int t = 0; t < ntris; ++t) {
		cupt_triangle tri;
		tri.v0 = (int32_t)indices[3 * t + 0] + indexBufferOffset;  // modelLoader.h:181-183
		tri.v1 = (int32_t)indices[3 * t + 1] + indexBufferOffset;
		tri.v2 = (int32_t)indices[3 * t + 2] + indexBufferOffset;
		const int nv = (int)s->verts.size();
		if (tri.v0 < 0 || tri.v1 < 0 || tri.v2 < 0 || tri.v0 >= nv || tri.v1 >= nv || tri.v2 >= nv)
			return set_error(CUPT_ERR_INVALID, "shape triangle %d references a vertex outside the shape", t);
		tri.mat = face_mats ? face_mats[t] : material;
		if (tri.mat < 0 || tri.mat >= (int)s->mats.size())
			return set_error(CUPT_ERR_INVALID, "triangle %d: material %d outside the table of %d "
			                 "(the reference silently indexes out of range here, modelLoader.h:189-191)",
			                 t, tri.mat, (int)s->mats.size());
		const uint32_t ind = (uint32_t)s->tris.size();
		V3 v0 = toV3(s->verts[tri.v0]), v1 = toV3(s->verts[tri.v1]), v2 = toV3(s->verts[tri.v2]);
		if (s->mats[tri.mat].emission[0] != 0) {  // modelLoader.h:191-199
			s->lights.push_back((int32_t)ind);
			float area = length3(cross3(v1 - v0, v2 - v0)) / 2;
			s->totalLightArea += area;
		}
		V3 n = normalized3(cross3(v1 - v0, v2 - v0));  // modelLoader.h:201
		if (flip)
			n = n * -1.f;  // modelLoader.h:202-203
		tri.norm = fromV3(n);
		s->tris.push_back(tri);
	}
	return CUPT_OK;
}

}  // namespace cupt

using namespace cupt;

extern "C" {

const char* cupt_last_error(void) { return g_last_error.c_str(); }
int cupt_version(void) { return CUPT_VERSION; }

int cupt_struct_sizes(int* out7)
{
	if (!out7)
		return set_error(CUPT_ERR_INVALID, "null output");
	out7[0] = (int)sizeof(cupt_vec3);
	out7[1] = (int)sizeof(cupt_triangle);
	out7[2] = (int)sizeof(cupt_material);
	out7[3] = (int)sizeof(cupt_camera);
	out7[4] = (int)sizeof(cupt_bvh2_node);
	out7[5] = (int)sizeof(cupt_sphere);
	out7[6] = 0;
	return CUPT_OK;
}

int cupt_scene_create(cupt_scene** out)
{
	if (!out)
		return set_error(CUPT_ERR_INVALID, "null output");
	*out = new (std::nothrow) cupt_scene();
	return *out ? CUPT_OK : set_error(CUPT_ERR_NOMEM, "out of memory");
}
void cupt_scene_destroy(cupt_scene* s) { delete s; }

int cupt_scene_clear(cupt_scene* s)
{
	if (!s)
		return set_error(CUPT_ERR_INVALID, "null scene");
	s->verts.clear(); s->tris.clear(); s->mats.clear(); s->lights.clear(); s->spheres.clear(); s->mat_specular.clear();
	s->totalLightArea = 0.f;
	return CUPT_OK;
}

int cupt_scene_add_materials(cupt_scene* s, const cupt_material* mats, int n, int duplicate_table, int* first_index)
{
	if (!s || (n > 0 && !mats) || n < 0)
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	// modelLoader.h:138-159: the table is appended twice and triangles reference the second copy
	if (duplicate_table)
		s->mats.insert(s->mats.end(), mats, mats + n);
	if (first_index)
		*first_index = (int)s->mats.size();  // matsOffset, modelLoader.h:148
	s->mats.insert(s->mats.end(), mats, mats + n);
	return CUPT_OK;
}

int cupt_scene_add_shape(cupt_scene* s, const float* positions, int nverts, const uint32_t* indices, int ntris,
                         int material, const float origin[3], float scale, int flip_normals)
{
	if (!s || nverts < 0 || ntris < 0 || (nverts && !positions) || (ntris && !indices))
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	return append_shape(s, positions, nverts, indices, ntris, material, nullptr, origin, scale, flip_normals);
}

int cupt_scene_add_shape_facemats(cupt_scene* s, const float* positions, int nverts, const uint32_t* indices,
                                  int ntris, const int32_t* face_materials, const float origin[3], float scale,
                                  int flip_normals)
{
	if (!s || nverts < 0 || ntris < 0 || (nverts && !positions) || (ntris && (!indices || !face_materials)))
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	return append_shape(s, positions, nverts, indices, ntris, 0, face_materials, origin, scale, flip_normals);
}

int cupt_scene_load_obj_ex(cupt_scene* s, const char* path, const char* mtl_dir, const float origin[3], float scale,
                           int flip_normals, int flags)
{
	if (!s || !path)
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	std::vector<ObjShape> shapes;
	std::vector<ObjMaterial> materials;
	std::string err;
	int rc = parse_obj(path, mtl_dir ? mtl_dir : "models/", shapes, materials, err);
	if (rc != CUPT_OK)
		return set_error(rc, "%s", err.c_str());
	std::vector<cupt_material> table(materials.size());
	for (size_t i = 0; i < materials.size(); ++i)  // modelLoader.h:140-145: float -> double widening
		for (int k = 0; k < 3; ++k) {
			table[i].albedo[k] = materials[i].diffuse[k];
			table[i].emission[k] = materials[i].emission[k];
		}
	// file material index -> scene material index
	std::vector<int> remap(table.size());
	if (flags & CUPT_OBJ_DEDUP_MATERIALS) {
		// one entry per distinct material instead of the reference's table appended twice per file
		for (size_t i = 0; i < table.size(); ++i) {
			int found = -1;
			for (size_t j = 0; j < s->mats.size() && found < 0; ++j)
				if (memcmp(&s->mats[j], &table[i], sizeof(cupt_material)) == 0) found = (int)j;
			if (found < 0) { found = (int)s->mats.size(); s->mats.push_back(table[i]); }
			remap[i] = found;
		}
	} else {
		int matsOffset = 0;
		rc = cupt_scene_add_materials(s, table.empty() ? nullptr : &table[0], (int)table.size(), 1, &matsOffset);
		if (rc != CUPT_OK)
			return rc;
		for (size_t i = 0; i < table.size(); ++i) remap[i] = (int)i + matsOffset;
	}
	std::vector<int32_t> face;
	for (size_t i = 0; i < shapes.size(); ++i) {
		const ObjShape& sh = shapes[i];
		const int ntris = (int)(sh.indices.size() / 3);
		if (ntris == 0)
			continue;
		const int32_t* face_mats = nullptr;
		int mat = 0;
		if (flags & CUPT_OBJ_FACE_MATERIALS) {
			// material_ids[k] per triangle (the reference takes material_ids[0] for the whole shape, modelLoader.h:189)
			face.resize((size_t)ntris);
			for (int t = 0; t < ntris; ++t) {
				const int32_t m = sh.material_ids[(size_t)t];
				if (m < 0 || m >= (int32_t)remap.size())
					return set_error(CUPT_ERR_PARSE, "shape '%s' face %d has no material (usemtl missing or unknown)", sh.name.c_str(), t);
				face[(size_t)t] = remap[(size_t)m];
			}
			face_mats = &face[0];
		} else {
			// ONE material per shape: material_ids[0] + matsOffset (modelLoader.h:189).  tinyobj reports -1 for
			// "no material"; the reference then indexes the slot before its table.  That is rejected here.
			if (sh.material_ids[0] < 0 || sh.material_ids[0] >= (int32_t)remap.size())
				return set_error(CUPT_ERR_PARSE, "shape '%s' has no material (usemtl missing or unknown)", sh.name.c_str());
			mat = remap[(size_t)sh.material_ids[0]];
		}
		rc = append_shape(s, sh.positions.empty() ? nullptr : &sh.positions[0], (int)(sh.positions.size() / 3),
		                  &sh.indices[0], ntris, mat, face_mats, origin, scale, flip_normals);
		if (rc != CUPT_OK)
			return rc;
	}
	return CUPT_OK;
}

int cupt_scene_load_obj(cupt_scene* s, const char* path, const char* mtl_dir, const float origin[3], float scale,
                        int flip_normals)
{
	return cupt_scene_load_obj_ex(s, path, mtl_dir, origin, scale, flip_normals, 0);   // the reference's semantics
}

// ---- binary scene cache: the five arrays verbatim + FNV-1a checksum ----------------------------
namespace {
struct CacheHeader {
	char magic[8];            // "CUPTSCN1"
	uint64_t nverts, ntris, nmats, nlights, nspheres, nflags;
	float totalLightArea;
	uint32_t pod_sizes;       // sizeof(vec3) | sizeof(triangle) << 8 | sizeof(material) << 16 | sizeof(sphere) << 24
};
uint64_t fnv1a(uint64_t h, const void* p, size_t n)
{
	const unsigned char* b = static_cast<const unsigned char*>(p);
	for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; }
	return h;
}
const uint32_t kPodSizes = (uint32_t)sizeof(cupt_vec3) | ((uint32_t)sizeof(cupt_triangle) << 8) | ((uint32_t)sizeof(cupt_material) << 16) | ((uint32_t)sizeof(cupt_sphere) << 24);
}  // namespace

int cupt_scene_save(const cupt_scene* s, const char* path)
{
	if (!s || !path)
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	FILE* fp = fopen(path, "wb");
	if (!fp)
		return set_error(CUPT_ERR_IO, "cannot open %s for writing", path);
	CacheHeader h;
	memset(&h, 0, sizeof(h));
	memcpy(h.magic, "CUPTSCN1", 8);
	h.nverts = s->verts.size(); h.ntris = s->tris.size(); h.nmats = s->mats.size(); h.nlights = s->lights.size(); h.nspheres = s->spheres.size(); h.nflags = s->mat_specular.size();
	h.totalLightArea = s->totalLightArea; h.pod_sizes = kPodSizes;
	uint64_t sum = fnv1a(14695981039346656037ull, &h, sizeof(h));
	bool ok = fwrite(&h, sizeof(h), 1, fp) == 1;
	auto put = [&](const void* p, size_t bytes) {
		if (bytes == 0) return;
		sum = fnv1a(sum, p, bytes);
		ok = ok && fwrite(p, 1, bytes, fp) == bytes;
	};
	put(s->verts.data(), s->verts.size() * sizeof(cupt_vec3));
	put(s->tris.data(), s->tris.size() * sizeof(cupt_triangle));
	put(s->mats.data(), s->mats.size() * sizeof(cupt_material));
	put(s->lights.data(), s->lights.size() * sizeof(int32_t));
	put(s->spheres.data(), s->spheres.size() * sizeof(cupt_sphere));
	put(s->mat_specular.data(), s->mat_specular.size());
	ok = ok && fwrite(&sum, sizeof(sum), 1, fp) == 1;
	ok = (fclose(fp) == 0) && ok;
	return ok ? CUPT_OK : set_error(CUPT_ERR_IO, "write to %s failed", path);
}

int cupt_scene_load(cupt_scene* s, const char* path)
{
	if (!s || !path)
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	FILE* fp = fopen(path, "rb");
	if (!fp)
		return set_error(CUPT_ERR_IO, "cannot open %s", path);
	CacheHeader h;
	if (fread(&h, sizeof(h), 1, fp) != 1 || memcmp(h.magic, "CUPTSCN1", 8) != 0 || h.pod_sizes != kPodSizes) {
		fclose(fp);
		return set_error(CUPT_ERR_PARSE, "%s is not a scene cache of this library version", path);
	}
	const uint64_t lim = 0x7fffffffull;
	if (h.nverts > lim || h.ntris > lim || h.nmats > lim || h.nlights > h.ntris || h.nspheres > lim || h.nflags > h.nmats) {
		fclose(fp);
		return set_error(CUPT_ERR_PARSE, "%s: implausible counts in the header", path);
	}
	cupt_scene t;
	uint64_t sum = fnv1a(14695981039346656037ull, &h, sizeof(h));
	bool ok = true;
	auto get = [&](void* p, size_t bytes) {
		if (bytes == 0 || !ok) return;
		ok = fread(p, 1, bytes, fp) == bytes;
		if (ok) sum = fnv1a(sum, p, bytes);
	};
	try {
		t.verts.resize(h.nverts); t.tris.resize(h.ntris); t.mats.resize(h.nmats); t.lights.resize(h.nlights); t.spheres.resize(h.nspheres); t.mat_specular.resize(h.nflags);
	} catch (...) {
		fclose(fp);
		return set_error(CUPT_ERR_NOMEM, "out of memory reading %s", path);
	}
	get(t.verts.data(), t.verts.size() * sizeof(cupt_vec3));
	get(t.tris.data(), t.tris.size() * sizeof(cupt_triangle));
	get(t.mats.data(), t.mats.size() * sizeof(cupt_material));
	get(t.lights.data(), t.lights.size() * sizeof(int32_t));
	get(t.spheres.data(), t.spheres.size() * sizeof(cupt_sphere));
	get(t.mat_specular.data(), t.mat_specular.size());
	uint64_t stored = 0;
	ok = ok && fread(&stored, sizeof(stored), 1, fp) == 1;
	fclose(fp);
	if (!ok) return set_error(CUPT_ERR_PARSE, "%s is truncated", path);
	if (stored != sum) return set_error(CUPT_ERR_PARSE, "%s: checksum mismatch", path);
	for (const cupt_triangle& tr : t.tris)
		if (tr.v0 < 0 || tr.v1 < 0 || tr.v2 < 0 || (uint64_t)tr.v0 >= h.nverts || (uint64_t)tr.v1 >= h.nverts || (uint64_t)tr.v2 >= h.nverts ||
		    tr.mat < 0 || (uint64_t)tr.mat >= h.nmats)
			return set_error(CUPT_ERR_PARSE, "%s: triangle references outside the arrays", path);
	for (int32_t l : t.lights)
		if (l < 0 || (uint64_t)l >= h.ntris) return set_error(CUPT_ERR_PARSE, "%s: light index outside the triangle array", path);
	t.totalLightArea = h.totalLightArea;
	s->verts.swap(t.verts); s->tris.swap(t.tris); s->mats.swap(t.mats); s->lights.swap(t.lights); s->spheres.swap(t.spheres); s->mat_specular.swap(t.mat_specular);
	s->totalLightArea = t.totalLightArea;
	return CUPT_OK;
}

int cupt_scene_set_specular(cupt_scene* s, int material, int specular)
{
	if (!s || material < 0 || material >= (int)s->mats.size())
		return set_error(CUPT_ERR_INVALID, "material %d outside the table", material);
	if (s->mat_specular.size() < s->mats.size()) s->mat_specular.resize(s->mats.size(), 0);
	s->mat_specular[(size_t)material] = specular ? 1 : 0;
	return CUPT_OK;
}

int cupt_scene_add_spheres(cupt_scene* s, const cupt_sphere* spheres, int n)
{
	if (!s || n < 0 || (n && !spheres))
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	s->spheres.insert(s->spheres.end(), spheres, spheres + n);
	return CUPT_OK;
}

int cupt_scene_set_arrays(cupt_scene* s, const cupt_vec3* verts, int nverts, const cupt_triangle* tris, int ntris,
                          const cupt_material* mats, int nmats, const int32_t* lights, int nlights,
                          float total_light_area)
{
	if (!s || nverts < 0 || ntris < 0 || nmats < 0 || nlights < 0)
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	for (int t = 0; t < ntris; ++t) {
		const cupt_triangle& tr = tris[t];
		if (tr.v0 < 0 || tr.v1 < 0 || tr.v2 < 0 || tr.v0 >= nverts || tr.v1 >= nverts || tr.v2 >= nverts ||
		    tr.mat < 0 || tr.mat >= nmats)
			return set_error(CUPT_ERR_INVALID, "triangle %d has an index out of range", t);
	}
	for (int l = 0; l < nlights; ++l)
		if (lights[l] < 0 || lights[l] >= ntris)
			return set_error(CUPT_ERR_INVALID, "light %d out of range", l);
	s->verts.assign(verts, verts + nverts);
	s->tris.assign(tris, tris + ntris);
	s->mats.assign(mats, mats + nmats);
	s->lights.assign(lights, lights + nlights);
	s->totalLightArea = total_light_area;
	return CUPT_OK;
}

int cupt_scene_counts(const cupt_scene* s, int* out5)
{
	if (!s || !out5)
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	out5[0] = (int)s->verts.size(); out5[1] = (int)s->tris.size(); out5[2] = (int)s->mats.size();
	out5[3] = (int)s->lights.size(); out5[4] = (int)s->spheres.size();
	return CUPT_OK;
}

int cupt_scene_get_arrays(const cupt_scene* s, cupt_vec3* verts, cupt_triangle* tris, cupt_material* mats,
                          int32_t* lights)
{
	if (!s)
		return set_error(CUPT_ERR_INVALID, "null scene");
	if (verts && !s->verts.empty()) memcpy(verts, &s->verts[0], s->verts.size() * sizeof(cupt_vec3));
	if (tris && !s->tris.empty()) memcpy(tris, &s->tris[0], s->tris.size() * sizeof(cupt_triangle));
	if (mats && !s->mats.empty()) memcpy(mats, &s->mats[0], s->mats.size() * sizeof(cupt_material));
	if (lights && !s->lights.empty()) memcpy(lights, &s->lights[0], s->lights.size() * sizeof(int32_t));
	return CUPT_OK;
}

float cupt_scene_total_light_area(const cupt_scene* s) { return s ? s->totalLightArea : 0.f; }

}  // extern "C"
