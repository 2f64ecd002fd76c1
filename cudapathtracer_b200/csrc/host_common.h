// Shared host-side declarations of libcupt (not part of the public ABI).
#pragma once
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cmath>
#include <string>
#include <vector>
#include "../../include/cupt.h"

namespace cupt {

int set_error(int code, const char* fmt, ...);

// Float helpers with the reference's operation order (vec3.h:58-84).  This file is compiled with
// -ffp-contract=off so that every product and sum below is individually rounded, exactly like the
// reference's host pass.
struct V3 {
	float x, y, z;
};
inline V3 v3(float x, float y, float z) { V3 r = {x, y, z}; return r; }
inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
inline float dot3(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }  // vec3.h:63-66
inline V3 cross3(V3 a, V3 b)                                                 // vec3.h:67-70
{
	return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
inline float length3(V3 v) { return sqrtf(v.x * v.x + v.y * v.y + v.z * v.z); }  // vec3.h:72-75
inline V3 normalized3(V3 v)                                                      // vec3.h:58-62
{
	float len = sqrtf(v.x * v.x + v.y * v.y + v.z * v.z);
	return v3(v.x / len, v.y / len, v.z / len);
}
inline V3 toV3(const cupt_vec3& a) { return v3(a.x, a.y, a.z); }
inline cupt_vec3 fromV3(V3 a) { cupt_vec3 r = {a.x, a.y, a.z}; return r; }

}  // namespace cupt

// The host scene: the reference's process-wide globals (modelLoader.h:43-47) as one object.
struct cupt_scene {
	std::vector<cupt_vec3> verts;
	std::vector<cupt_triangle> tris;
	std::vector<cupt_material> mats;
	std::vector<int32_t> lights;
	float totalLightArea = 0.f;
	std::vector<cupt_sphere> spheres;
	std::vector<uint8_t> mat_specular;   // optional, per material: 1 = perfect mirror (not in the reference; cupt_scene_set_specular)
};

namespace cupt {
// OBJ/MTL parser (obj_loader.cpp): shapes with the semantics of the loader the reference vendors
struct ObjMaterial {
	std::string name;
	float diffuse[3];
	float emission[3];
};
struct ObjShape {
	std::string name;
	std::vector<float> positions;        // xyz per de-duplicated vertex
	std::vector<uint32_t> indices;       // 3 per triangle
	std::vector<int32_t> material_ids;   // per triangle
};
int parse_obj(const char* path, const char* mtl_dir, std::vector<ObjShape>& shapes,
              std::vector<ObjMaterial>& materials, std::string& err);
}  // namespace cupt
