/* cupt_compat.h — source-level shim: keeps a driver written like the reference's main() (kernel.cu:589-778) compiling
 * against libcupt.so.  Include it INSTEAD of modelLoader.h / BVH.h, with the reference's own vec3.h, color.h and camera.h
 * reachable by #include "..." (compile with `-iquote <reference dir>`, not -I: the reference's limits.h would shadow the
 * system <limits.h>).  It provides, under the reference's names:
 *   triangle, materialDesc (modelLoader.h:14-25), BVH_array_node, BVH_array (BVH.h:111-121), the limits.h constants
 *   the globals verts, tris, mats, lights, totalLightArea (modelLoader.h:43-47)
 *   void loadOBJ(std::string filename, vec3 origin, float scale, bool flipNormals = false)   (modelLoader.h:125)
 *   BVH_array buildBVH()                                                                       (BVH.h:443)
 * and cupt_compat_scene(), the scene object behind the globals, for cupt_set_scene / cupt_bvh_build.
 * Like the reference's headers it DEFINES globals and functions: include it in exactly one translation unit.
 * A host C++ compiler is enough (the CUDA qualifiers of the reference's headers are defined away); nvcc works too.
 * Compiled and run by tests (examples/reference_main.cpp). */
#ifndef CUPT_COMPAT_H
#define CUPT_COMPAT_H

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>
#include "cupt.h"

#ifndef __CUDACC__
#ifndef __host__
#define __host__
#endif
#ifndef __device__
#define __device__
#endif
/* the reference's headers call the global min / max / sqrt / pow overloads the CUDA toolchain provides (vec3.h:77-84, color.h:68-71) */
using std::max;
using std::min;
using std::pow;
using std::sqrt;
#endif

#include "vec3.h"
#include "color.h"
#include "camera.h"   /* the reference's own camera: layout-identical to cupt_camera (pinned by tests/test_abi.py) */

static_assert(sizeof(vec3) == sizeof(cupt_vec3), "vec3 layout");
static_assert(sizeof(camera) == sizeof(cupt_camera), "camera layout");

typedef cupt_triangle triangle;         /* modelLoader.h:14-19 (field names v0, v1, v2, norm, mat) */
typedef cupt_material materialDesc;     /* modelLoader.h:21-25 (albedo, emission as double[3]) */
typedef cupt_bvh2_node BVH_array_node;  /* BVH.h:111-115 */
struct BVH_array {                      /* BVH.h:116-121 */
	BVH_array_node* root;
	int size;
	int depth;
};
const float MAX_FLOAT_host = CUPT_MAX_FLOAT;          /* limits.h:3-8 */
const uint32_t MAX_BVH_INDEX = CUPT_MAX_BVH_INDEX;
const uint32_t BVH_LEAF_FLAG = CUPT_BVH_LEAF_FLAG;

/* modelLoader.h:43-47 */
std::vector<vec3> verts;
std::vector<triangle> tris;
std::vector<materialDesc> mats;
std::vector<int32_t> lights;
float totalLightArea = 0;

inline cupt_scene* cupt_compat_scene()
{
	static cupt_scene* s = [] { cupt_scene* p = 0; cupt_scene_create(&p); return p; }();
	return s;
}

/* mirror the scene object into the reference's globals */
inline void cupt_compat_refresh_globals()
{
	int n[5];
	cupt_scene_counts(cupt_compat_scene(), n);
	verts.resize((size_t)n[0]); tris.resize((size_t)n[1]); mats.resize((size_t)n[2]); lights.resize((size_t)n[3]);
	cupt_scene_get_arrays(cupt_compat_scene(), (cupt_vec3*)verts.data(), tris.data(), mats.data(), lights.data());
	totalLightArea = cupt_scene_total_light_area(cupt_compat_scene());
}

inline void loadOBJ(std::string filename, vec3 origin, float scale, bool flipNormals = false)   /* modelLoader.h:125 */
{
	const float o[3] = {origin.x, origin.y, origin.z};
	if (cupt_scene_load_obj(cupt_compat_scene(), filename.c_str(), "models/", o, scale, flipNormals) != 0)
		printf("%s\n", cupt_last_error());   /* the reference prints and carries on (modelLoader.h:134-135) */
	cupt_compat_refresh_globals();
}

inline BVH_array buildBVH()   /* BVH.h:443: reads the globals' scene, returns the reference-layout array (bit-identical tree) */
{
	BVH_array b = {0, 0, 0};
	if (cupt_bvh2_build(cupt_compat_scene(), &b.root, &b.size, &b.depth) != 0) printf("%s\n", cupt_last_error());
	return b;
}

#endif /* CUPT_COMPAT_H */
