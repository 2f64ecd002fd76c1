// cupt_bvh_build — host builder of the wide quantised BVH (replaces buildBVH, reference BVH.h:133-474).
//
//   1. binned-SAH binary BVH, top-down, OpenMP tasks over subtrees (the reference: 3x3x3 centroid grid,
//      nine candidate planes, one triangle per leaf, single thread — BVH.h:166-259)
//   2. collapse to 8-wide by repeatedly opening the child with the largest surface area
//   3. children assigned to octant slots so that a ray can order them with one XOR (layout.h)
//   4. child boxes quantised to 8 bits on a per-node power-of-two grid, conservatively in exact arithmetic
//   5. breadth-first emission: inner children of a node and its primitive records are consecutive
#include "host_common.h"
#include "layout.h"
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstring>
#include <cstdlib>
#include <omp.h>

namespace cupt {
namespace {

struct Box3 {
	float lo[3], hi[3];
	void reset() { for (int a = 0; a < 3; ++a) { lo[a] = 3.0e38f; hi[a] = -3.0e38f; } }
	void grow(const Box3& o) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], o.lo[a]); hi[a] = std::max(hi[a], o.hi[a]); } }
	void grow(const float* p) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); } }
	double area() const
	{
		double dx = (double)hi[0] - lo[0], dy = (double)hi[1] - lo[1], dz = (double)hi[2] - lo[2];
		if (dx < 0 || dy < 0 || dz < 0) return 0.0;
		return 2.0 * (dx * dy + dx * dz + dy * dz);
	}
};

struct Node2 {
	Box3 box;
	int32_t left, right;   // inner: children; leaf: left = -1
	int32_t first, count;  // leaf: range in the primitive index array
};

struct Build2 {
	const Box3* pbox;
	const float* cent;  // 3 per primitive
	std::vector<uint32_t> idx;
	std::vector<Node2> nodes;
	std::atomic<int32_t> next;
	int bins, maxLeaf;

	int32_t alloc() { return next.fetch_add(1); }

	void build(int32_t ni, int32_t begin, int32_t end, int depth)
	{
		Node2& n = nodes[ni];
		const int32_t count = end - begin;
		Box3 box, cb;
		box.reset(); cb.reset();
		for (int32_t i = begin; i < end; ++i) { box.grow(pbox[idx[i]]); cb.grow(cent + 3 * (size_t)idx[i]); }
		n.box = box;
		if (count <= maxLeaf) { n.left = n.right = -1; n.first = begin; n.count = count; return; }

		// binned SAH over the three axes
		const int K = bins;
		int bestAxis = -1, bestSplit = 0;
		double bestCost = 1e300;
		Box3 bb[256];
		int32_t bc[256];
		double rightArea[256];
		for (int axis = 0; axis < 3; ++axis) {
			const float lo = cb.lo[axis], ext = cb.hi[axis] - cb.lo[axis];
			if (!(ext > 0.f)) continue;
			const float k1 = (float)K * (1.0f - 1e-6f) / ext;
			for (int b = 0; b < K; ++b) { bb[b].reset(); bc[b] = 0; }
			for (int32_t i = begin; i < end; ++i) {
				const uint32_t p = idx[i];
				int b = (int)((cent[3 * (size_t)p + axis] - lo) * k1);
				b = b < 0 ? 0 : (b >= K ? K - 1 : b);
				bb[b].grow(pbox[p]); bc[b]++;
			}
			Box3 acc; acc.reset();
			for (int b = K - 1; b > 0; --b) { acc.grow(bb[b]); rightArea[b] = acc.area(); }
			acc.reset();
			int32_t nl = 0;
			for (int b = 0; b < K - 1; ++b) {
				acc.grow(bb[b]); nl += bc[b];
				if (nl == 0 || nl == count) continue;
				double cost = acc.area() * nl + rightArea[b + 1] * (count - nl);
				if (cost < bestCost) { bestCost = cost; bestAxis = axis; bestSplit = b + 1; }
			}
		}
		int32_t mid;
		if (bestAxis < 0) {
			mid = begin + count / 2;  // all centroids coincide: split by position
		} else {
			const float lo = cb.lo[bestAxis], ext = cb.hi[bestAxis] - cb.lo[bestAxis];
			const float k1 = (float)K * (1.0f - 1e-6f) / ext;
			const float* c = cent;
			const int axis = bestAxis, split = bestSplit;
			uint32_t* m = std::partition(&idx[begin], &idx[0] + end, [=](uint32_t p) {
				int b = (int)((c[3 * (size_t)p + axis] - lo) * k1);
				b = b < 0 ? 0 : (b >= K ? K - 1 : b);
				return b < split;
			});
			mid = (int32_t)(m - &idx[0]);
			if (mid == begin || mid == end) mid = begin + count / 2;
		}
		const int32_t l = alloc(), r = alloc();
		nodes[ni].left = l; nodes[ni].right = r; nodes[ni].first = begin; nodes[ni].count = count;
		if (count > 8192) {
#pragma omp task default(shared) firstprivate(l, begin, mid, depth)
			build(l, begin, mid, depth + 1);
#pragma omp task default(shared) firstprivate(r, mid, end, depth)
			build(r, mid, end, depth + 1);
#pragma omp taskwait
		} else {
			build(l, begin, mid, depth + 1);
			build(r, mid, end, depth + 1);
		}
	}
};

// ---- insertion-based optimisation of the binary tree (after Bittner, Hapala, Havran, "Fast Insertion-Based Optimization of
// Bounding Volume Hierarchies", CGF 2013; branch-and-bound search as in later open-source builders) ---------------------------
// A top-down SAH build fixes every split with local knowledge only.  Here every subtree is taken out of the tree in turn (its
// parent node disappears, the sibling moves up) and re-inserted next to the node where the total surface area of the inner
// nodes — the SAH cost, leaves unchanged — decreases most, found by a search that walks up from the old position and down into
// the siblings it passes, pruned by the best gain so far.
struct Reinserter {
	std::vector<Node2>& nodes;
	std::vector<int32_t> parent;
	int32_t root = 0;

	Reinserter(std::vector<Node2>& n, int32_t count) : nodes(n), parent((size_t)count, -1)
	{
		for (int32_t i = 0; i < count; ++i)
			if (nodes[(size_t)i].left >= 0) { parent[(size_t)nodes[(size_t)i].left] = i; parent[(size_t)nodes[(size_t)i].right] = i; }
	}
	int32_t sibling(int32_t n) const { const Node2& p = nodes[(size_t)parent[(size_t)n]]; return p.left == n ? p.right : p.left; }
	static Box3 merged(const Box3& a, const Box3& b) { Box3 r = a; r.grow(b); return r; }
	void refit_from(int32_t n)
	{
		for (; n >= 0; n = parent[(size_t)n]) {
			Node2& m = nodes[(size_t)n];
			const Box3 b = merged(nodes[(size_t)m.left].box, nodes[(size_t)m.right].box);
			if (memcmp(&b, &m.box, sizeof(Box3)) == 0) break;
			m.box = b;
		}
	}
	// best place for the subtree `in`: the node to pair it with and the surface area saved (first == in: leave it where it is)
	std::pair<int32_t, double> find(int32_t in, std::vector<std::pair<double, int32_t>>& stack) const
	{
		const Box3 bin = nodes[(size_t)in].box;
		const double ain = bin.area();
		const int32_t par = parent[(size_t)in];
		double area_diff = nodes[(size_t)par].box.area();     // the parent node disappears
		int32_t sib = sibling(in), pivot = par;
		Box3 pivot_box = nodes[(size_t)sib].box;               // box of the pivot once `in` is gone
		std::pair<int32_t, double> best(in, 0.0);
		for (;;) {
			stack.clear();
			stack.emplace_back(area_diff, sib);
			while (!stack.empty()) {
				const std::pair<double, int32_t> top = stack.back();
				stack.pop_back();
				if (top.first - ain <= best.second) continue;   // even a zero-growth insertion below this node cannot win
				const Node2& dst = nodes[(size_t)top.second];
				const double gain = top.first - merged(dst.box, bin).area();
				if (gain > best.second) best = std::make_pair(top.second, gain);
				if (dst.left >= 0) {
					const double child = gain + dst.box.area();   // descending: dst becomes the merged box, the new node appears below it
					stack.emplace_back(child, dst.left);
					stack.emplace_back(child, dst.right);
				}
			}
			if (pivot == root) break;
			if (pivot != par) {                                  // the pivot shrinks once `in` is gone
				pivot_box = merged(pivot_box, nodes[(size_t)sib].box);
				area_diff += nodes[(size_t)pivot].box.area() - pivot_box.area();
			}
			sib = sibling(pivot);
			pivot = parent[(size_t)pivot];
		}
		if (best.first == sibling(in) || best.first == par) best = std::make_pair(in, 0.0);
		return best;
	}
	void reinsert(int32_t in, int32_t out)
	{
		const int32_t par = parent[(size_t)in], sib = sibling(in), grand = parent[(size_t)par];
		// take `in` out: the sibling replaces the parent
		if (grand >= 0) { Node2& g = nodes[(size_t)grand]; (g.left == par ? g.left : g.right) = sib; }
		else root = sib;
		parent[(size_t)sib] = grand;
		// the freed parent node becomes the parent of (out, in)
		const int32_t op = parent[(size_t)out];
		if (op >= 0) { Node2& o = nodes[(size_t)op]; (o.left == out ? o.left : o.right) = par; }
		else root = par;
		parent[(size_t)par] = op;
		nodes[(size_t)par].left = out; nodes[(size_t)par].right = in;
		parent[(size_t)out] = par; parent[(size_t)in] = par;
		nodes[(size_t)par].box = merged(nodes[(size_t)out].box, nodes[(size_t)in].box);
		if (op >= 0) refit_from(op);
		if (grand >= 0) refit_from(grand);
	}
	// one sweep over every subtree, large ones first; returns the number of moves
	int64_t pass(int32_t count)
	{
		std::vector<std::pair<float, int32_t>> order;
		order.reserve((size_t)count);
		for (int32_t i = 0; i < count; ++i)
			if (i != root && parent[(size_t)i] >= 0 && parent[(size_t)i] != root) order.emplace_back((float)nodes[(size_t)i].box.area(), i);
		std::sort(order.begin(), order.end(), [](const std::pair<float, int32_t>& a, const std::pair<float, int32_t>& b) { return a.first > b.first || (a.first == b.first && a.second < b.second); });
		std::vector<std::pair<double, int32_t>> stack;
		int64_t moves = 0;
		for (const auto& c : order) {
			const int32_t in = c.second;
			if (in == root || parent[(size_t)in] < 0 || parent[(size_t)in] == root) continue;
			const std::pair<int32_t, double> b = find(in, stack);
			if (b.first != in && b.second > 0.0) { reinsert(in, b.first); ++moves; }
		}
		return moves;
	}
};

inline int float_exponent(float v)  // floor(log2(|v|)) for normal values; -127 for 0/denormal
{
	uint32_t u;
	memcpy(&u, &v, 4);
	return (int)((u >> 23) & 0xff) - 127;
}

}  // namespace
}  // namespace cupt

using namespace cupt;

extern "C" {

int cupt_bvh_build(const cupt_scene* s, const cupt_bvh_opts* opts, cupt_bvh** out)
{
	if (!s || !out)
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	const int64_t nt = (int64_t)s->tris.size(), ns = (int64_t)s->spheres.size();
	const int64_t N = nt + ns;
	if (N < 1)
		return set_error(CUPT_ERR_INVALID, "scene has no primitives");
	if (N > 0x7fffffff)
		return set_error(CUPT_ERR_LIMIT, "more than 2^31-1 primitives (limits.h:6)");
	const int bins = (opts && opts->sah_bins > 1) ? std::min(opts->sah_bins, 256) : 32;
	const int maxLeaf = (opts && opts->max_leaf_tris >= 1) ? std::min(opts->max_leaf_tris, 3) : 3;
	const int threads = (opts && opts->threads > 0) ? opts->threads : omp_get_max_threads();
	const uint32_t* tie_rank = opts ? opts->tie_rank : nullptr;
	auto t0 = std::chrono::steady_clock::now();

	// primitive boxes, padded by a few ulps of the coordinate magnitude: the reference's triangle test is
	// plain f32 (modelLoader.h:49-82) and may accept a ray that misses the exact triangle by rounding
	std::vector<Box3> pbox((size_t)N);
	std::vector<float> cent((size_t)N * 3);
#pragma omp parallel for num_threads(threads) schedule(static)
	for (int64_t i = 0; i < N; ++i) {
		Box3 b; b.reset();
		if (i < nt) {
			const cupt_triangle& t = s->tris[(size_t)i];
			b.grow(&s->verts[t.v0].x); b.grow(&s->verts[t.v1].x); b.grow(&s->verts[t.v2].x);
		} else {
			const cupt_sphere& sp = s->spheres[(size_t)(i - nt)];
			const float c[3] = {sp.pos.x, sp.pos.y, sp.pos.z};
			for (int a = 0; a < 3; ++a) { b.lo[a] = c[a] - sp.rad; b.hi[a] = c[a] + sp.rad; }
		}
		for (int a = 0; a < 3; ++a) {
			cent[3 * (size_t)i + a] = 0.5f * b.lo[a] + 0.5f * b.hi[a];
			float m = std::max(std::max(fabsf(b.lo[a]), fabsf(b.hi[a])), 1e-30f);
			float pad = m * 4.8e-7f;  // 4 ulp
			b.lo[a] -= pad; b.hi[a] += pad;
		}
		pbox[(size_t)i] = b;
	}

	Build2 B;
	B.pbox = &pbox[0]; B.cent = &cent[0]; B.bins = bins; B.maxLeaf = maxLeaf;
	B.idx.resize((size_t)N);
	for (int64_t i = 0; i < N; ++i) B.idx[(size_t)i] = (uint32_t)i;
	B.nodes.resize((size_t)2 * N + 1);
	B.next.store(1);
#pragma omp parallel num_threads(threads)
#pragma omp single
	B.build(0, 0, (int32_t)N, 0);

	int32_t root2 = 0;
	// default: two sweeps (config 2: SAH cost 20.9 -> 20.3, node visits per ray 11.4 -> 10.7, extend 33.5 -> 31.0 ms per frame; more
	// sweeps change nothing measurable) up to 2^20 primitives, where the single-threaded search stays within a few seconds (6 us
	// per primitive and sweep); none above.  The multi-million soups of config 4 would not profit anyway: uniformly random
	// triangles leave the sweeps nothing to find (1.5 M: SAH cost 129.20 -> 129.04 for 8 s).  A batched parallel search (moves
	// found concurrently on a frozen tree, applied in order when their nodes were untouched) was tried and rejected: even at 64
	// candidates per batch the stale gains cost half of the improvement (SAH 20.49 instead of 20.27, visits 10.95 instead of 10.67).
	int reinsert_passes = (opts && opts->reinsert_passes != 0) ? std::min(opts->reinsert_passes, 16) : (N <= ((int64_t)1 << 20) ? 2 : 0);
	reinsert_passes = std::max(reinsert_passes, 0);
	if (reinsert_passes > 0 && B.next.load() > 3) {
		Reinserter R(B.nodes, B.next.load());
		for (int it = 0; it < reinsert_passes; ++it)
			if (R.pass(B.next.load()) == 0) break;
		root2 = R.root;
	}

	// ---- collapse + emit ---------------------------------------------------------------------------------
	cupt_bvh* bvh = new (std::nothrow) cupt_bvh();
	if (!bvh)
		return set_error(CUPT_ERR_NOMEM, "out of memory");
	bvh->num_tris = nt; bvh->num_spheres = ns;
	// ---- cost-optimal collapse (after Ylitie, Karras, Laine 2017, section 3.1) --------------------------------
	// c[i] = cheapest way to hang the subtree of a binary node under a wide node using at most i of its slots:
	//   leaf:      A * P * c_prim                                   (binary leaves hold <= max_leaf primitives)
	//   c[1]:      A * c_node + D[8]      the subtree becomes a wide node of its own
	//   c[i>1]:    min(D[i], c[i-1]),     D[j] = min_k c_left[k] + c_right[j-k]
	// The greedy rule (open the largest child until there are eight) left 45 % of config 2's nodes with two children.
	struct Dp { float c[8]; uint8_t split[9]; };
	const bool use_dp = !(opts && opts->greedy_collapse);
	const float c_prim = 0.3f, c_node = 1.0f;   // (c_prim 0.15 .. 0.8 give the same config 2 tree to within one node)
	std::vector<Dp> dp;
	if (use_dp) {
		const int32_t n2count = B.next.load();
		dp.resize((size_t)n2count);
		std::vector<int32_t> topo;   // parents before children (the optimisation pass moves subtrees, so index order is not enough)
		topo.reserve((size_t)n2count);
		topo.push_back(root2);
		for (size_t q = 0; q < topo.size(); ++q) {
			const Node2& n = B.nodes[(size_t)topo[q]];
			if (n.left >= 0) { topo.push_back(n.left); topo.push_back(n.right); }
		}
		for (size_t q = topo.size(); q-- > 0;) {       // bottom-up
			const int32_t i = topo[q];
			const Node2& n = B.nodes[(size_t)i];
			Dp& d = dp[(size_t)i];
			const float A = (float)n.box.area();
			if (n.left < 0) {
				for (int k = 0; k < 8; ++k) d.c[k] = A * (float)n.count * c_prim;
				memset(d.split, 0, sizeof(d.split));
				continue;
			}
			const Dp& l = dp[(size_t)n.left];
			const Dp& r = dp[(size_t)n.right];
			float D[9];
			for (int j = 2; j <= 8; ++j) {
				float best = 3.0e38f; int bk = 1;
				for (int k = 1; k < j; ++k) {
					const float v = l.c[std::min(k, 7)] + r.c[std::min(j - k, 7)];
					if (v < best) { best = v; bk = k; }
				}
				D[j] = best; d.split[j] = (uint8_t)bk;
			}
			d.c[0] = 0.f;
			d.c[1] = A * c_node + D[8];
			for (int k = 2; k <= 7; ++k) d.c[k] = std::min(D[k], d.c[k - 1]);
		}
	}
	struct Pending { int32_t n2; int32_t depth; };
	std::vector<Pending> queue;
	queue.push_back({root2, 1});
	bvh->nodes.reserve((size_t)(N / 3 + 16));
	bvh->prims.reserve((size_t)N);
	const double rootArea = std::max(B.nodes[(size_t)root2].box.area(), 1e-300);
	double sah = 0;
	int maxDepth = 0;
	for (size_t qi = 0; qi < queue.size(); ++qi) {
		const Pending cur = queue[qi];
		maxDepth = std::max(maxDepth, (int)cur.depth);
		const Node2& n2 = B.nodes[cur.n2];
		// gather up to 8 children
		int32_t kids[8];
		int nk = 0;
		if (n2.left < 0) {
			kids[nk++] = cur.n2;  // root is a leaf: one leaf child
		} else if (use_dp) {
			// cost-optimal cut below this node (dp[] above): distribute the eight slots over the two subtrees
			struct Frame { int32_t n; int budget; };
			Frame st[64]; int sp = 0;
			const int k8 = dp[(size_t)cur.n2].split[8];
			st[sp++] = {n2.right, 8 - k8}; st[sp++] = {n2.left, k8};
			while (sp > 0) {
				const Frame f = st[--sp];
				const Node2& m = B.nodes[f.n];
				if (m.left < 0) { kids[nk++] = f.n; continue; }
				int i = std::min(f.budget, 7);
				const Dp& d = dp[(size_t)f.n];
				while (i > 1 && d.c[i] == d.c[i - 1]) --i;
				if (i == 1) { kids[nk++] = f.n; continue; }     // a wide node of its own
				const int k = d.split[i];
				st[sp++] = {m.right, i - k}; st[sp++] = {m.left, k};
			}
		} else {
			kids[nk++] = n2.left; kids[nk++] = n2.right;
			while (nk < 8) {
				int best = -1; double bestA = -1;
				for (int k = 0; k < nk; ++k)
					if (B.nodes[kids[k]].left >= 0) {
						double a = B.nodes[kids[k]].box.area();
						if (a > bestA) { bestA = a; best = k; }
					}
				if (best < 0) break;
				const int32_t open = kids[best];
				kids[best] = B.nodes[open].left;
				kids[nk++] = B.nodes[open].right;
			}
		}
		// octant slots: greedy assignment maximising dot(child centre - node centre, slot direction)
		float nc[3];
		for (int a = 0; a < 3; ++a) nc[a] = 0.5f * n2.box.lo[a] + 0.5f * n2.box.hi[a];
		int slotOf[8], childIn[8];
		for (int k = 0; k < 8; ++k) { slotOf[k] = -1; childIn[k] = -1; }
		float cost[8][8];
		for (int k = 0; k < nk; ++k) {
			const Box3& cb = B.nodes[kids[k]].box;
			float d[3];
			for (int a = 0; a < 3; ++a) d[a] = 0.5f * cb.lo[a] + 0.5f * cb.hi[a] - nc[a];
			for (int sl = 0; sl < 8; ++sl)
				cost[k][sl] = ((sl & 4) ? d[0] : -d[0]) + ((sl & 2) ? d[1] : -d[1]) + ((sl & 1) ? d[2] : -d[2]);
		}
		for (int it = 0; it < nk; ++it) {
			int bk = -1, bs = -1; float bc = -3.0e38f;
			for (int k = 0; k < nk; ++k) {
				if (slotOf[k] >= 0) continue;
				for (int sl = 0; sl < 8; ++sl)
					if (childIn[sl] < 0 && cost[k][sl] > bc) { bc = cost[k][sl]; bk = k; bs = sl; }
			}
			slotOf[bk] = bs; childIn[bs] = bk;
		}
		// quantisation grid
		Node8 nd;
		memset(&nd, 0, sizeof(nd));
		double scale[3];
		for (int a = 0; a < 3; ++a) {
			nd.p[a] = n2.box.lo[a];
			const double ext = (double)n2.box.hi[a] - (double)n2.box.lo[a];
			int e = -126;
			if (ext > 0) { e = (int)ceil(log2(ext / 255.0)); while (ldexp(255.0, e) < ext) ++e; }
			const float m = std::max(fabsf(n2.box.lo[a]), fabsf(n2.box.hi[a]));
			e = std::max(e, float_exponent(m) - 24);
			e = std::max(-126, std::min(127, e));
			nd.e[a] = (uint8_t)(e + 127);
			scale[a] = ldexp(1.0, e);
		}
		nd.child_base = (uint32_t)queue.size();       // BFS: index in queue == index of the emitted node
		nd.tri_base = (uint32_t)bvh->prims.size();
		uint32_t triOfs = 0;
		for (int sl = 0; sl < 8; ++sl) {
			const int k = childIn[sl];
			if (k < 0) continue;
			const Node2& c = B.nodes[kids[k]];
			for (int a = 0; a < 3; ++a) {
				const double p = nd.p[a];
				double ql = floor(((double)c.box.lo[a] - p) / scale[a]);
				double qh = ceil(((double)c.box.hi[a] - p) / scale[a]);
				while (ql > 0 && p + ql * scale[a] > (double)c.box.lo[a]) ql -= 1;
				while (p + qh * scale[a] < (double)c.box.hi[a]) qh += 1;
				ql = std::max(0.0, std::min(255.0, ql));
				if (qh > 255.0) { delete bvh; return set_error(CUPT_ERR_LIMIT, "internal: quantised box overflow"); }
				qh = std::max(qh, ql);
				nd.qlo[a][sl] = (uint8_t)ql; nd.qhi[a][sl] = (uint8_t)qh;
			}
			sah += c.box.area() / rootArea * (c.left < 0 ? c.count : 1.0);
			if (c.left >= 0) {
				nd.imask |= (uint8_t)(1u << sl);
				queue.push_back({kids[k], cur.depth + 1});
			} else {
				const uint32_t unary = c.count == 1 ? 1u : (c.count == 2 ? 3u : 7u);
				nd.pvalid |= unary << (4 * sl);
				for (int32_t i = 0; i < c.count; ++i) {
					const uint32_t p = B.idx[(size_t)c.first + i];
					Prim48 r;
					memset(&r, 0, sizeof(r));
					if ((int64_t)p < nt) {
						const cupt_triangle& t = s->tris[p];
						const V3 v0 = toV3(s->verts[t.v0]), e1 = toV3(s->verts[t.v1]) - v0, e2 = toV3(s->verts[t.v2]) - v0;
						r.a[0] = v0.x; r.a[1] = v0.y; r.a[2] = v0.z;
						r.b[0] = e1.x; r.b[1] = e1.y; r.b[2] = e1.z;
						r.c[0] = e2.x; r.c[1] = e2.y; r.c[2] = e2.z;
						r.id = p; r.rank = tie_rank ? tie_rank[p] : p; r.kind = 0;
					} else {
						const cupt_sphere& sp = s->spheres[p - (uint32_t)nt];
						r.a[0] = sp.pos.x; r.a[1] = sp.pos.y; r.a[2] = sp.pos.z;
						r.b[0] = sp.rad;
						r.id = p - (uint32_t)nt; r.rank = p; r.kind = 1;
					}
					bvh->prims.push_back(r);
				}
				triOfs += (uint32_t)c.count;
			}
		}
		bvh->nodes.push_back(nd);
	}
	bvh->depth = maxDepth;
	bvh->sah_cost = sah;
	bvh->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	*out = bvh;
	return CUPT_OK;
}

void cupt_bvh_destroy(cupt_bvh* b) { delete b; }

int cupt_bvh_get_info(const cupt_bvh* b, cupt_bvh_info* info)
{
	if (!b || !info)
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	info->num_nodes = (int64_t)b->nodes.size();
	info->num_tri_records = (int64_t)b->prims.size();
	info->num_tris = b->num_tris; info->num_spheres = b->num_spheres;
	info->depth = b->depth;
	info->node_bytes = (int32_t)sizeof(Node8); info->tri_bytes = (int32_t)sizeof(Prim48);
	info->sah_cost = b->sah_cost; info->build_seconds = b->build_seconds;
	return CUPT_OK;
}

int cupt_bvh_get_arrays(const cupt_bvh* b, void* nodes80, void* tris48)
{
	if (!b)
		return set_error(CUPT_ERR_INVALID, "null bvh");
	if (nodes80) memcpy(nodes80, b->nodes.data(), b->nodes.size() * sizeof(Node8));
	if (tris48) memcpy(tris48, b->prims.data(), b->prims.size() * sizeof(Prim48));
	return CUPT_OK;
}

}  // extern "C"
