// TEST INFRASTRUCTURE — see oracle.h.  CPU restatement of simple-knn's distCUDA2.
//
// The reference (simple_knn.cu:136-172) returns, per point, the mean of the three smallest squared distances to
// OTHER points (index != own; coincident points count with distance 0), found by an exact pruned search
// (strict `>` pruning at :161), so any exact 3-NN search reproduces it.  The squared distance is
// d.x*d.x + d.y*d.y + d.z*d.z with d = other - self (updateKBest, :120-134), which the reference build rounds as
// fma(d.z,d.z, fma(d.x,d.x, d.y*d.y)); the mean is ((b0 + b1) + b2) / 3.0f with b ascending (:171).  Both are
// spelled out below so results are bit-comparable with the GPU.
#include "oracle.h"
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <vector>

namespace {

inline float sqdist(const float* a, const float* ref) {
	const float dx = a[0] - ref[0], dy = a[1] - ref[1], dz = a[2] - ref[2];
	return fmaf(dz, dz, fmaf(dx, dx, dy * dy));
}

inline void update3(float dist, float* knn) {   // updateKBest<3> (simple_knn.cu:120-134)
	for (int j = 0; j < 3; j++) if (knn[j] > dist) { float t = knn[j]; knn[j] = dist; dist = t; }
}

inline float mean3(const float* b) { return ((b[0] + b[1]) + b[2]) / 3.0f; }

struct Grid {
	float mn[3], cell; int dim[3];
	std::vector<uint32_t> start, order;
	int cell_of(const float* p, int a) const { int c = (int)((p[a] - mn[a]) / cell); return std::min(std::max(c, 0), dim[a] - 1); }
	void build(int P, const float* pts) {
		float mx[3];
		for (int a = 0; a < 3; a++) { mn[a] = FLT_MAX; mx[a] = -FLT_MAX; }
		for (int i = 0; i < P; i++) for (int a = 0; a < 3; a++) { mn[a] = std::min(mn[a], pts[3 * (size_t)i + a]); mx[a] = std::max(mx[a], pts[3 * (size_t)i + a]); }
		double vol = 1; for (int a = 0; a < 3; a++) vol *= std::max((double)mx[a] - mn[a], 1e-6);
		cell = (float)std::cbrt(vol / std::max(1.0, P / 4.0));
		for (int a = 0; a < 3; a++) {
			dim[a] = std::min(512, std::max(1, (int)(((double)mx[a] - mn[a]) / cell) + 1));
		}
		// widen the cell if the 512 clamp kicked in
		for (int a = 0; a < 3; a++) cell = std::max(cell, (float)(((double)mx[a] - mn[a]) / dim[a]) * 1.0001f);
		const size_t nc = (size_t)dim[0] * dim[1] * dim[2];
		start.assign(nc + 1, 0); order.resize(P);
		std::vector<uint32_t> cid(P);
		for (int i = 0; i < P; i++) {
			const float* p = &pts[3 * (size_t)i];
			cid[i] = (uint32_t)(((size_t)cell_of(p, 2) * dim[1] + cell_of(p, 1)) * dim[0] + cell_of(p, 0));
			start[cid[i] + 1]++;
		}
		for (size_t c = 0; c < nc; c++) start[c + 1] += start[c];
		std::vector<uint32_t> fill(start.begin(), start.end() - 1);
		for (int i = 0; i < P; i++) order[fill[cid[i]]++] = (uint32_t)i;
	}
	float query(int P, const float* pts, int qi) const {
		(void)P;
		const float* q = &pts[3 * (size_t)qi];
		const int c0[3] = { cell_of(q, 0), cell_of(q, 1), cell_of(q, 2) };
		float best[3] = { FLT_MAX, FLT_MAX, FLT_MAX };
		const int rmax = std::max(dim[0], std::max(dim[1], dim[2]));
		for (int r = 0; r <= rmax; r++) {
			if (r > 0) {
				// every point outside the (2r-1)^3 block is at least (r-1)*cell away
				const double lim = (double)(r - 1) * cell;
				if ((double)best[2] < lim * lim * (1.0 - 1e-5)) break;
			}
			for (int z = c0[2] - r; z <= c0[2] + r; z++) {
				if (z < 0 || z >= dim[2]) continue;
				for (int y = c0[1] - r; y <= c0[1] + r; y++) {
					if (y < 0 || y >= dim[1]) continue;
					const bool shell_zy = (std::abs(z - c0[2]) == r) || (std::abs(y - c0[1]) == r);
					for (int x = c0[0] - r; x <= c0[0] + r; x += (shell_zy ? 1 : std::max(1, 2 * r))) {
						if (x < 0 || x >= dim[0]) continue;
						const size_t c = ((size_t)z * dim[1] + y) * dim[0] + x;
						for (uint32_t k = start[c]; k < start[c + 1]; k++) {
							const uint32_t j = order[k];
							if ((int)j == qi) continue;
							update3(sqdist(&pts[3 * (size_t)j], q), best);
						}
					}
				}
			}
		}
		return mean3(best);
	}
};

}  // namespace

extern "C" {

void orc_dist2_brute(int P, const float* points, float* out) {
	#pragma omp parallel for schedule(static)
	for (int i = 0; i < P; i++) {
		float best[3] = { FLT_MAX, FLT_MAX, FLT_MAX };
		for (int j = 0; j < P; j++) { if (j == i) continue; update3(sqdist(&points[3 * (size_t)j], &points[3 * (size_t)i]), best); }
		out[i] = mean3(best);
	}
}

void orc_dist2(int P, const float* points, float* out) {
	if (P <= 0) return;
	Grid g; g.build(P, points);
	#pragma omp parallel for schedule(dynamic, 1024)
	for (int i = 0; i < P; i++) out[i] = g.query(P, points, i);
}

void orc_dist2_queries(int P, const float* points, int Q, const int* query_idx, float* out) {
	if (P <= 0) return;
	Grid g; g.build(P, points);
	#pragma omp parallel for schedule(dynamic, 64)
	for (int k = 0; k < Q; k++) out[k] = g.query(P, points, query_idx[k]);
}

// prepMorton / coord2Morton (simple_knn.cu:34-50)
static uint32_t prep_morton(uint32_t x) {
	x = (x | (x << 16)) & 0x030000FF;
	x = (x | (x << 8)) & 0x0300F00F;
	x = (x | (x << 4)) & 0x030C30C3;
	x = (x | (x << 2)) & 0x09249249;
	return x;
}
uint32_t orc_morton(float cx, float cy, float cz, const float* minn, const float* maxx) {
	const uint32_t x = prep_morton((uint32_t)(((cx - minn[0]) / (maxx[0] - minn[0])) * 1023.0f));
	const uint32_t y = prep_morton((uint32_t)(((cy - minn[1]) / (maxx[1] - minn[1])) * 1023.0f));
	const uint32_t z = prep_morton((uint32_t)(((cz - minn[2]) / (maxx[2] - minn[2])) * 1023.0f));
	return x | (y << 1) | (z << 2);
}

}
