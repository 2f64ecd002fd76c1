// Device data layout shared by the host builders and the CUDA kernels (DESIGN.md "Data layout in HBM").
#pragma once
#include <cstdint>

namespace cupt {

// ---- wide BVH node: 80 bytes = five 128-bit ld.global.nc fetches -------------------------------------------
// Replaces the reference's 32-byte {lo, hi, left, right} record (BVH.h:111-115): eight children per fetch
// instead of two, child boxes quantised to 8 bits on a per-node power-of-two grid, children addressed by
// base + popcount instead of explicit links.  Compressed-wide-BVH idea after Ylitie, Karras, Laine,
// "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs" (HPG 2017); layout below is ours.
//
//   word 0-2   origin p = node lo (f32 x3)
//   word 3     ex | ey<<8 | ez<<16 | imask<<24      scale_a = 2^(e_a - 127); imask bit s = slot s is an inner child
//   word 4     child_base   index of the first inner child node (inner children are consecutive, in slot order)
//   word 5     0
//   word 6     tri_base     index of the first 48-byte primitive record of this node (records are consecutive)
//   word 7     pvalid       one nibble per slot: leaf = unary count of its records (1 / 3 / 7 for 1 / 2 / 3), inner or empty = 0.
//                           The records of a node are consecutive in slot order, so the record of bit b of this word is
//                           tri_base + popcount(pvalid below b): the traversal can keep a leaf hit mask in this sparse form
//                           (pvalid & hit nibbles, no per-child shifts) until a primitive is actually listed for testing;
//                           words 6-7 are one aligned 8-byte fetch
//   word 8-9   qlo_x[8]     child box = p + q * scale, lo rounded down, hi rounded up (conservative in exact arithmetic)
//   word 10-11 qlo_y[8]
//   word 12-13 qlo_z[8]
//   word 14-15 qhi_x[8]
//   word 16-17 qhi_y[8]
//   word 18-19 qhi_z[8]
// Slot s sits on the +x side if s&4, +y if s&2, +z if s&1 (builder assigns children to slots so that visiting
// slots in the order (s ^ oct_inv) descending is front-to-back for a ray whose non-negative axes are oct_inv).
struct alignas(16) Node8 {
	float p[3];
	uint8_t e[3];
	uint8_t imask;
	uint32_t child_base;
	uint32_t reserved;
	uint32_t tri_base;
	uint32_t pvalid;
	uint8_t qlo[3][8];
	uint8_t qhi[3][8];
};
static_assert(sizeof(Node8) == 80, "Node8 must be 80 bytes");

// ---- primitive record: 48 bytes = three 128-bit fetches ----------------------------------------------------
// Replaces the reference's 28-byte indexed triangle + three dependent 12-byte vertex gathers (64 B, two round
// trips, modelLoader.h:51-53).  e1 = v1 - v0 and e2 = v2 - v0 are single f32 subtractions, i.e. exactly the
// values triIntersect computes first (modelLoader.h:57-58), so the test stays bit-identical.
//   kind 0 (triangle): a = v0, b = e1, c = e2
//   kind 1 (sphere):   a = centre, b.x = radius
struct alignas(16) Prim48 {
	float a[3];
	float b[3];
	float c[3];
	uint32_t id;     // triangle index in the scene (sphere: index in the sphere table)
	uint32_t rank;   // tie-break rank: the reference's left-first leaf order, or = id
	uint32_t kind;
};
static_assert(sizeof(Prim48) == 48, "Prim48 must be 48 bytes");

// ---- shading tables (f32; the reference keeps colours in f64, color.h:6) -------------------------------------
struct alignas(16) TriShade {   // per triangle: geometric normal + material (modelLoader.h:17-18)
	float n[3];
	int32_t mat;
};
struct alignas(16) MatF {       // per material (modelLoader.h:21-25)
	float albedo[3];
	float emissive_flag;        // 1 when emission.r != 0 in f64 (modelLoader.h:191, kernel.cu:453)
	float emission[3];
	float specular_flag;        // 1 = perfect mirror (an extension: the reference has diffuse materials only)
};
struct alignas(16) LightTri {   // per light triangle (+ one fallback entry for triangle 0, kernel.cu:469)
	float v0[3];
	float area;                 // length(cross(a1, a2)) / 2, kernel.cu:477
	float a1[3];
	float pad0;
	float a2[3];
	float pad1;
};

struct alignas(16) SphereShade {   // per sphere (sphere.h:7-13), f32
	float centre[3];
	float radius;
	float diffuse[3];
	float emissive_flag;          // 1 when emm.r != 0 in f64 (same rule as materials, kernel.cu:453)
	float emm[3];
	float pad;
};

}  // namespace cupt

// host image of the wide BVH
#include <vector>
struct cupt_bvh {
	std::vector<cupt::Node8> nodes;
	std::vector<cupt::Prim48> prims;
	int64_t num_tris = 0, num_spheres = 0;
	int depth = 0;
	double sah_cost = 0, build_seconds = 0;
};
