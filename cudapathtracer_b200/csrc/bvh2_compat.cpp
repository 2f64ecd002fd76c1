// cupt_bvh2_build — the reference-layout binary BVH, reproducing buildBVH() (BVH.h:443-474) bit for bit:
// one triangle per leaf, 3x3x3 centroid grid, nine candidate planes scored by a surface-area heuristic,
// positional halving when a side is empty, breadth-first flattening into 32-byte {lo, hi, left, right}
// records with the leaf flag in the child link.  It exists for two reasons:
//   * drop-in compatibility: a driver written against the reference gets the same BVH_array;
//   * the reference breaks ties between equal hit distances by its traversal order (kernel.cu:127,144-145);
//     cupt_bvh2_leaf_order turns that order into a per-triangle rank the wide BVH can apply.
// Design differences from the reference (behaviour identical): index-based arena instead of heap nodes,
// no per-call new[]/delete[] (two ping-pong index arrays), errors instead of exit()/unbounded recursion.
// Compiled with -ffp-contract=off.
#include "host_common.h"
#include <algorithm>
#include <cfloat>
#include <climits>
#include <cstdlib>
#include <cstring>

namespace cupt {
namespace {

struct Box {
	float lo[3], hi[3];
};
inline Box negative_box()  // AABB::makeNegative, BVH.h:21-25
{
	Box b;
	for (int a = 0; a < 3; ++a) { b.lo[a] = 10000.f; b.hi[a] = -10000.f; }
	return b;
}
inline void grow(Box& b, const Box& o)  // AABBUnion, BVH.h:33-37 (fminf/fmaxf like CUDA's host min/max)
{
	for (int a = 0; a < 3; ++a) { b.lo[a] = fminf(b.lo[a], o.lo[a]); b.hi[a] = fmaxf(b.hi[a], o.hi[a]); }
}
inline float area_weight(const Box& b)  // AABB::weight, BVH.h:27-31
{
	float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
	return 2 * (dx * dy + dx * dz + dy * dz);
}
// (int)float as the x86-64 host code of the reference performs it (cvttss2si: NaN / out of range -> INT_MIN)
inline int trunc_to_int(float f)
{
	if (!(f > -2147483904.0f && f < 2147483648.0f))
		return INT_MIN;
	return (int)f;
}
inline int grid_cell(float c, float unit)  // BVH.h:186-188
{
	return std::min(2, std::max(0, trunc_to_int(c / unit)));
}

struct TreeNode {
	Box box;
	int32_t left, right;  // arena indices; leaves are arena[0..N) with left = right = -1
	int32_t numChildNodes, depth;
};

struct Builder {
	std::vector<TreeNode> arena;  // [0,N) leaves (target = index), then inner nodes
	int ntris;
	int error;

	int make_inner(const Box& box, int l, int r, int numChild, int depth)
	{
		TreeNode n;
		n.box = box; n.left = l; n.right = r; n.numChildNodes = numChild; n.depth = depth;
		arena.push_back(n);
		return (int)arena.size() - 1;
	}

	// buildBVHRecurse, BVH.h:133-328.  `list` holds `count` leaf indices; it may be permuted in place.
	int recurse(int* list, int count, int level)
	{
		if (error)
			return -1;
		if (count <= 0 || level > 4096) {  // the reference recurses without bound here (BVH.h:263-276)
			error = CUPT_ERR_LIMIT;
			return -1;
		}
		if (count == 1)
			return list[0];  // BVH.h:151-154
		if (count == 2) {    // BVH.h:136-150
			Box total = negative_box();
			grow(total, arena[list[0]].box);
			grow(total, arena[list[1]].box);
			return make_inner(total, list[0], list[1], 2, 2);
		}
		Box total = negative_box();
		for (int i = 0; i < count; ++i)
			grow(total, arena[list[i]].box);
		const float totalWeight = area_weight(total);

		Box grid[3][3][3];
		int counts[3][3][3];
		for (int i = 0; i < 3; ++i)
			for (int j = 0; j < 3; ++j)
				for (int k = 0; k < 3; ++k) { grid[i][j][k] = negative_box(); counts[i][j][k] = 0; }
		float unit[3];
		for (int a = 0; a < 3; ++a)
			unit[a] = (total.hi[a] - total.lo[a]) / (float)3;  // BVH.h:182
		for (int i = 0; i < count; ++i) {
			const Box& b = arena[list[i]].box;
			int c[3];
			for (int a = 0; a < 3; ++a)
				c[a] = grid_cell((b.hi[a] + b.lo[a]) / (float)2 - total.lo[a], unit[a]);
			// BVH.h:190 unions the box of leaf number `i` (the loop counter), not of list[i]; kept.
			grow(grid[c[0]][c[1]][c[2]], arena[i].box);
			counts[c[0]][c[1]][c[2]] += 1;
		}

		int bestSlice = 0, bestAxis = 0, bestLeft = 0, bestRight = 0;
		double bestScore = DBL_MAX;
		for (int axis = 0; axis < 3; ++axis)
			for (int slice = 0; slice < 3; ++slice) {  // BVH.h:200-259
				Box L = negative_box(), R = negative_box();
				int nL = 0, nR = 0;
				for (int i = 0; i < 3; ++i)
					for (int j = 0; j < 3; ++j)
						for (int k = 0; k < 3; ++k) {
							if (counts[i][j][k] <= 0)
								continue;
							const int c = axis == 0 ? i : (axis == 1 ? j : k);
							if (c < slice) { nL += counts[i][j][k]; grow(L, grid[i][j][k]); }
							else { nR += counts[i][j][k]; grow(R, grid[i][j][k]); }
						}
				double leftProb = area_weight(L) / totalWeight;
				double rightProb = area_weight(R) / totalWeight;
				double score = nL * leftProb + nR * rightProb;
				if (score < bestScore) {
					bestSlice = slice; bestAxis = axis; bestScore = score; bestLeft = nL; bestRight = nR;
				}
			}

		int nLeft, nRight;
		if (bestLeft == 0 || bestRight == 0) {  // BVH.h:263-288: halve the list by position
			nLeft = bestRight / 2;
			nRight = bestRight - nLeft;
		} else {  // BVH.h:290-313: stable partition by the recomputed cell on bestAxis
			std::vector<int> rightTmp;
			rightTmp.reserve(bestRight);
			int w = 0;
			for (int i = 0; i < count; ++i) {
				const Box& b = arena[list[i]].box;
				float cell = (float)grid_cell((b.hi[bestAxis] + b.lo[bestAxis]) / (float)2 - total.lo[bestAxis], unit[bestAxis]);
				if (cell < bestSlice)
					list[w++] = list[i];
				else
					rightTmp.push_back(list[i]);
			}
			nLeft = w;
			nRight = (int)rightTmp.size();
			if (nRight)
				memcpy(list + nLeft, &rightTmp[0], sizeof(int) * nRight);
			if (nLeft != bestLeft || nRight != bestRight) {
				error = CUPT_ERR_LIMIT;
				return -1;
			}
		}
		int l = recurse(list, nLeft, level + 1);
		int r = recurse(list + nLeft, nRight, level + 1);
		if (error)
			return -1;
		return make_inner(total, l, r, arena[l].numChildNodes + arena[r].numChildNodes + 2,
		                  std::max(arena[l].depth, arena[r].depth) + 1);
	}
};

}  // namespace
}  // namespace cupt

using namespace cupt;

extern "C" {

int cupt_bvh2_build(const cupt_scene* s, cupt_bvh2_node** nodes, int* size, int* depth)
{
	if (!s || !nodes || !size || !depth)
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	const int N = (int)s->tris.size();
	if (N < 2)
		return set_error(CUPT_ERR_LIMIT, "the reference-layout BVH needs at least 2 triangles (got %d): "
		                 "buildBVH() has no representation for a tree without inner nodes (BVH.h:336,375)", N);
	Builder B;
	B.ntris = N;
	B.error = 0;
	B.arena.reserve((size_t)2 * N);
	B.arena.resize(N);
	std::vector<int> list(N);
	for (int i = 0; i < N; ++i) {  // BVH.h:451-462
		const cupt_triangle& t = s->tris[i];
		const cupt_vec3 &a = s->verts[t.v0], &b = s->verts[t.v1], &c = s->verts[t.v2];
		TreeNode& n = B.arena[i];
		n.box.lo[0] = fminf(fminf(a.x, b.x), c.x); n.box.lo[1] = fminf(fminf(a.y, b.y), c.y); n.box.lo[2] = fminf(fminf(a.z, b.z), c.z);
		n.box.hi[0] = fmaxf(fmaxf(a.x, b.x), c.x); n.box.hi[1] = fmaxf(fmaxf(a.y, b.y), c.y); n.box.hi[2] = fmaxf(fmaxf(a.z, b.z), c.z);
		n.left = n.right = -1; n.numChildNodes = 0; n.depth = 0;
		list[i] = i;
	}
	int root = B.recurse(&list[0], N, 0);
	if (B.error || root < 0)
		return set_error(CUPT_ERR_LIMIT, "degenerate geometry: the reference's builder does not terminate on this input "
		                 "(empty split side with no triangles to halve, BVH.h:263-276)");
	const long long arraySize = (long long)B.arena[root].numChildNodes + 1 - N;  // BVH.h:336
	if (arraySize > (long long)CUPT_MAX_BVH_INDEX || arraySize != (long long)B.arena.size() - N)
		return set_error(CUPT_ERR_LIMIT, "too many elements in BVH (BVH.h:337-341)");
	cupt_bvh2_node* out = (cupt_bvh2_node*)malloc(sizeof(cupt_bvh2_node) * (size_t)arraySize);
	if (!out)
		return set_error(CUPT_ERR_NOMEM, "out of memory");
	// BVHTreeToArrayBreadthFirst, BVH.h:331-382: inner nodes numbered in BFS order, left before right
	std::vector<int> queue;
	queue.reserve((size_t)arraySize);
	queue.push_back(root);
	size_t head = 0;
	while (head < queue.size()) {
		const TreeNode& cur = B.arena[queue[head]];
		cupt_bvh2_node& o = out[head];
		const int kids[2] = {cur.left, cur.right};
		uint32_t link[2];
		for (int k = 0; k < 2; ++k) {
			if (kids[k] < N) {
				link[k] = (uint32_t)kids[k] | CUPT_BVH_LEAF_FLAG;
			} else {
				link[k] = (uint32_t)queue.size();
				queue.push_back(kids[k]);
			}
		}
		o.lo.x = cur.box.lo[0]; o.lo.y = cur.box.lo[1]; o.lo.z = cur.box.lo[2];
		o.hi.x = cur.box.hi[0]; o.hi.y = cur.box.hi[1]; o.hi.z = cur.box.hi[2];
		o.left = link[0]; o.right = link[1];
		++head;
	}
	*nodes = out;
	*size = (int)arraySize;
	*depth = B.arena[root].depth;
	return CUPT_OK;
}

void cupt_bvh2_free(cupt_bvh2_node* nodes) { free(nodes); }

int cupt_bvh2_leaf_order(const cupt_bvh2_node* nodes, int size, int ntris, uint32_t* rank)
{
	if (!nodes || size < 1 || ntris < 1 || !rank)
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	for (int i = 0; i < ntris; ++i)
		rank[i] = 0xffffffffu;
	// trace() pushes right then left, so leaves are met left-first (kernel.cu:139-151)
	std::vector<uint32_t> stack;
	stack.push_back(0);
	uint32_t next = 0;
	size_t guard = 0;
	while (!stack.empty()) {
		uint32_t e = stack.back();
		stack.pop_back();
		if (++guard > (size_t)4 * ((size_t)size + (size_t)ntris))
			return set_error(CUPT_ERR_INVALID, "BVH links form a cycle");
		if (e & CUPT_BVH_LEAF_FLAG) {
			uint32_t id = e ^ CUPT_BVH_LEAF_FLAG;
			if (id >= (uint32_t)ntris)
				return set_error(CUPT_ERR_INVALID, "leaf %u outside the triangle array", id);
			if (rank[id] == 0xffffffffu)
				rank[id] = next++;
		} else {
			if (e >= (uint32_t)size)
				return set_error(CUPT_ERR_INVALID, "inner link %u outside the node array", e);
			stack.push_back(nodes[e].right);
			stack.push_back(nodes[e].left);
		}
	}
	for (int i = 0; i < ntris; ++i)
		if (rank[i] == 0xffffffffu)
			rank[i] = next++;  // unreachable triangles (none for a tree built here) keep a defined order
	return CUPT_OK;
}

}  // extern "C"
