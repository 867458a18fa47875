// C ABI of librainrast.so (include/rain_rast.h): host orchestration of the forward / backward pipelines.
// Replaces CudaRasterizer::Rasterizer::forward/backward/markVisible (reference rasterizer_impl.cu:187-430) and the
// torch binding RasterizeGaussiansCUDA / RasterizeGaussiansBackwardCUDA / markVisible (rasterize_points.cu) minus
// tensor allocation, which stays in Python.
#include <cstdlib>
#include <cstdarg>
#include <cstring>
#include <atomic>
#include <mutex>
#include <string>
#include <vector>
#include "../../include/rain_rast.h"
#include "common.cuh"
#include "kernels.cuh"
#include "radix_sort.cuh"

namespace rr {

static thread_local std::string g_error;
static std::atomic<long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

void set_error(const char* fmt, ...)
{
	char buf[512];
	va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
	g_error = buf;
}

int check_launch(const char* what, cudaStream_t stream, bool sync)
{
	cudaError_t e = cudaGetLastError();
	if (e == cudaSuccess && sync) e = cudaStreamSynchronize(stream);
	if (e != cudaSuccess) { set_error("%s: %s", what, cudaGetErrorString(e)); return -1; }
	return 0;
}

#ifdef RR_CROSSCHECK
// cross-check build only (make crosscheck): RAIN_XC_STAGING=tma stages the tile renderers with per-record TMA bulk copies
int crosscheck_staging() { static const int m = []() { const char* e = getenv("RAIN_XC_STAGING"); return (e && e[0] == 't') ? 1 : 0; }(); return m; }
// work counters of the tile renderers (render_common.cuh); the cross-check build may allocate, the product library never does
static unsigned long long* xc_counters_dev() {
	static unsigned long long* p = []() { unsigned long long* q = nullptr; if (cudaMalloc(&q, 16 * sizeof(unsigned long long)) != cudaSuccess) return (unsigned long long*)nullptr; cudaMemset(q, 0, 16 * sizeof(unsigned long long)); return q; }();
	return p;
}
#define XC_COUNTERS xc_counters_dev()
#else
#define XC_COUNTERS nullptr
#endif

namespace {

// ---- per-stage device timing (CUDA events on the caller's stream), off by default; used by bench.py -----------------
enum StageId { ST_PREPROCESS = 0, ST_TILE_PLAN, ST_DEPTH_SORT, ST_SCAN, ST_DUPLICATE, ST_TILE_SORT, ST_RANGES, ST_RENDER_FWD,
	ST_RENDER_BWD, ST_PREPROCESS_BWD, ST_BACKWARD_VIEWS, ST_DIST2, ST_COUNT };
const char* const kStageNames[ST_COUNT] = { "preprocess_fwd", "tile_plan", "depth_sort", "scan", "duplicate", "tile_sort", "ranges",
	"render_fwd", "render_bwd", "preprocess_bwd", "backward_views", "dist2" };
struct Pending { int stage; cudaEvent_t a, b; };
struct Prof {
	std::mutex mu;
	bool on = false;
	std::vector<cudaEvent_t> pool;
	std::vector<Pending> pending;
	double ms[ST_COUNT] = {};
	int64_t calls[ST_COUNT] = {};
	cudaEvent_t get() {
		if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
		cudaEvent_t e; cudaEventCreate(&e); return e;
	}
	void drain() {   // caller holds mu
		for (const Pending& p : pending) {
			float t = 0.f;
			if (cudaEventSynchronize(p.b) == cudaSuccess && cudaEventElapsedTime(&t, p.a, p.b) == cudaSuccess) { ms[p.stage] += t; calls[p.stage]++; }
			pool.push_back(p.a); pool.push_back(p.b);
		}
		pending.clear();
	}
};
Prof g_prof;
struct Stage {
	cudaStream_t stream; int id; cudaEvent_t a = nullptr; bool live = false;
	Stage(cudaStream_t s, int id_) : stream(s), id(id_) {
		if (!g_prof.on) return;
		std::lock_guard<std::mutex> lk(g_prof.mu);
		a = g_prof.get(); live = true;
		cudaEventRecord(a, stream);
	}
	~Stage() {
		if (!live) return;
		std::lock_guard<std::mutex> lk(g_prof.mu);
		cudaEvent_t b = g_prof.get();
		cudaEventRecord(b, stream);
		g_prof.pending.push_back({ id, a, b });
	}
};

// bit length, the reference's getHigherMsb (rasterizer_impl.cu:24-39)
int tile_bits(int num_tiles) { int b = 0; while ((num_tiles >> b) != 0) b++; return b < 1 ? 1 : b; }

struct GeomLayout {
	SplatRec* rec; float* cov3D; uint8_t* clamped; uint32_t* tiles_touched; uint2* rect;
	uint32_t *key_a, *key_b, *val_a, *val_b, *offsets, *block_sums, *num_rendered; void* sort_ws;
	unsigned long long* tile_mask;
	float* acc;                 // [P, ACC_FLOATS] accumulators of this view's backward tile renderer; rr_forward_render leaves them zero
	size_t total;
	int64_t off[9];
	int64_t acc_off;
};
// opt-in tile culling (preprocess.cu): RAIN_TILE_CULL=1, read once per process — the one run-time switch of the library
bool tile_cull_wanted() { static const bool on = []() { const char* e = getenv("RAIN_TILE_CULL"); return e && e[0] == '1'; }(); return on; }

GeomLayout geom_layout(void* blob, int P)
{
	GeomLayout L; Carver c(blob);
	const size_t Pz = (size_t)(P < 1 ? 1 : P);
	L.rec = c.take<SplatRec>(Pz);
	L.cov3D = c.take<float>(6 * Pz);
	L.clamped = c.take<uint8_t>(Pz);
	L.tiles_touched = c.take<uint32_t>(Pz);
	L.rect = c.take<uint2>(Pz);
	L.key_a = c.take<uint32_t>(Pz); L.key_b = c.take<uint32_t>(Pz);
	L.val_a = c.take<uint32_t>(Pz); L.val_b = c.take<uint32_t>(Pz);
	L.offsets = c.take<uint32_t>(Pz);
	L.block_sums = c.take<uint32_t>(scan_ws_elems((int)Pz));
	L.num_rendered = c.take<uint32_t>(4);
	L.sort_ws = c.take<char>(sort_ws_bytes<uint32_t>(Pz, 0, 32));
	L.tile_mask = tile_cull_wanted() ? c.take<unsigned long long>(Pz) : nullptr;
	L.acc = c.take<float>(Pz * ACC_FLOATS);
	L.total = c.used();
	char* b = reinterpret_cast<char*>(blob);
	L.acc_off = (char*)L.acc - b;
	// depth sort: 4 passes -> result back in the *_a buffers
	L.off[0] = (char*)L.rec - b; L.off[1] = (char*)L.cov3D - b; L.off[2] = (char*)L.clamped - b; L.off[3] = (char*)L.tiles_touched - b;
	L.off[4] = (char*)L.rect - b; L.off[5] = (char*)L.val_a - b; L.off[6] = (char*)L.offsets - b; L.off[7] = L.off[6];
	L.off[8] = (char*)L.num_rendered - b;
	return L;
}

struct BinLayout {
	uint32_t *key_a, *key_b, *val_a, *val_b, *warp_flags; void* sort_ws;
	uint32_t *keys_sorted, *point_list;
	int bits; size_t flag_words; size_t total; int64_t off[5];
};
BinLayout bin_layout(void* blob, int64_t N, int W, int H)
{
	BinLayout L; Carver c(blob);
	const size_t Nz = (size_t)(N < 1 ? 1 : N);
	const int tiles = ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
	L.bits = tile_bits(tiles);
	L.key_a = c.take<uint32_t>(Nz); L.key_b = c.take<uint32_t>(Nz);
	L.val_a = c.take<uint32_t>(Nz); L.val_b = c.take<uint32_t>(Nz);
	L.flag_words = warp_flag_words(Nz, (size_t)tiles);
	L.warp_flags = c.take<uint32_t>(L.flag_words);
	L.sort_ws = c.take<char>(sort_ws_bytes<uint32_t>(Nz, 0, L.bits));
	L.total = c.used();
	const bool in_b = (((L.bits + 7) / 8) & 1) != 0;
	L.keys_sorted = in_b ? L.key_b : L.key_a;
	L.point_list = in_b ? L.val_b : L.val_a;
	char* b = reinterpret_cast<char*>(blob);
	L.off[0] = (char*)L.keys_sorted - b; L.off[1] = (char*)L.point_list - b; L.off[2] = (char*)L.warp_flags - b;
	L.off[3] = (char*)L.point_list - b; L.off[4] = 0;
	return L;
}

struct ImgLayout { float* final_T; uint32_t* n_contrib; uint2* ranges; uint32_t* tile_max; int* plan_ws; uint32_t* plan_bases; size_t total; int64_t off[4]; };
ImgLayout img_layout(void* blob, int W, int H)
{
	ImgLayout L; Carver c(blob);
	const size_t X = (size_t)W * H;
	const size_t tiles = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
	L.final_T = c.take<float>(X < 1 ? 1 : X);
	L.n_contrib = c.take<uint32_t>(X < 1 ? 1 : X);
	L.ranges = c.take<uint2>(tiles < 1 ? 1 : tiles);
	L.tile_max = c.take<uint32_t>(tiles < 1 ? 1 : tiles);
	{	// difference grid of the tile plan (binning.cu)
		const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
		L.plan_ws = tile_plan_supported(gx, gy) ? c.take<int>(tile_plan_ws_bytes(gx, gy) / sizeof(int)) : nullptr;
		L.plan_bases = c.take<uint32_t>(2 * 256);            // exclusive digit bases of the tile sort's passes
	}
	L.total = c.used();
	char* b = reinterpret_cast<char*>(blob);
	L.off[0] = (char*)L.final_T - b; L.off[1] = (char*)L.n_contrib - b; L.off[2] = (char*)L.ranges - b; L.off[3] = (char*)L.tile_max - b;
	return L;
}

// the tile plan (ranges + digit bases from the rects, run in the geometry half) covers everything but the opt-in tile culling
// and grids beyond its shared-memory budget / 16-bit tile ids
bool plan_in_use(const GeomLayout& G, const ImgLayout& I, int W, int H)
{
	const int tiles = ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
	return I.plan_ws && !G.tile_mask && tile_bits(tiles) <= 16;
}

bool bad_ptr(const void* p, const char* name) { if (!p) { set_error("%s must not be null", name); return true; } return false; }

}  // namespace
}  // namespace rr

using namespace rr;

extern "C" {

const char* rr_last_error(void) { return g_error.c_str(); }
int rr_version(void) { return 2; }

int64_t rr_launch_count(void) { return (int64_t)g_launches.load(); }
int rr_tile_cull_enabled(void) { return tile_cull_wanted() ? 1 : 0; }

int rr_prof_enable(int on)
{
	std::lock_guard<std::mutex> lk(g_prof.mu);
	g_prof.on = on != 0;
	return 0;
}
int rr_prof_num_stages(void) { return ST_COUNT; }
const char* rr_prof_stage_name(int i) { return (i >= 0 && i < ST_COUNT) ? kStageNames[i] : ""; }
int rr_prof_read(double* ms, int64_t* calls, int reset)
{
	std::lock_guard<std::mutex> lk(g_prof.mu);
	g_prof.drain();
	for (int i = 0; i < ST_COUNT; i++) { if (ms) ms[i] = g_prof.ms[i]; if (calls) calls[i] = g_prof.calls[i]; }
	if (reset) for (int i = 0; i < ST_COUNT; i++) { g_prof.ms[i] = 0; g_prof.calls[i] = 0; }
	return 0;
}

size_t rr_geom_bytes(int P) { return geom_layout(nullptr, P).total + 256; }
size_t rr_image_bytes(int W, int H) { return img_layout(nullptr, W, H).total + 256; }
size_t rr_binning_bytes(int64_t N, int W, int H) { return bin_layout(nullptr, N, W, H).total + 256; }
size_t rr_geom_acc_offset(int P) { return (size_t)geom_layout(nullptr, P).acc_off; }
size_t rr_backward_ws_bytes(int P) { return (size_t)(P < 1 ? 1 : P) * ACC_FLOATS * sizeof(float) + 256; }

int rr_geom_layout(int P, int64_t o[9]) { const GeomLayout L = geom_layout(nullptr, P); memcpy(o, L.off, sizeof(L.off)); return 0; }
int rr_binning_layout(int64_t N, int W, int H, int64_t o[5]) { const BinLayout L = bin_layout(nullptr, N, W, H); memcpy(o, L.off, sizeof(L.off)); return 0; }
int rr_image_layout(int W, int H, int64_t o[4]) { const ImgLayout L = img_layout(nullptr, W, H); memcpy(o, L.off, sizeof(L.off)); return 0; }

int rr_forward_geom(void* stream_, int P, int D, int M, int W, int H,
	const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tan_fovx, float tan_fovy, float low_pass, int prefiltered, int debug,
	int* radii, void* geom, size_t geom_bytes, void* image, size_t image_bytes, int64_t* num_rendered_host)
{
	(void)prefiltered;   // the reference only uses it to trap on an inconsistent caller (auxiliary.h:145-149)
	cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
	if (bad_ptr(num_rendered_host, "num_rendered_host")) return -1;
	*num_rendered_host = 0;
	if (P == 0) return 0;                                                  // rasterize_points.cu:72
	if (P < 0 || W <= 0 || H <= 0) { set_error("bad sizes P=%d W=%d H=%d", P, W, H); return -1; }
	if (bad_ptr(means3D, "means3D") || bad_ptr(opacities, "opacities") || bad_ptr(viewmatrix, "viewmatrix") || bad_ptr(projmatrix, "projmatrix") ||
		bad_ptr(campos, "campos") || bad_ptr(radii, "radii") || bad_ptr(geom, "geom")) return -1;
	if (!shs && !colors_precomp) { set_error("For non-RGB, provide precomputed Gaussian colors!"); return -1; }   // rasterizer_impl.cu:233-236
	if (!cov3D_precomp && !(scales && rotations)) { set_error("provide scales+rotations or cov3D_precomp"); return -1; }
	if (shs && !colors_precomp && (D < 0 || D > 3 || (D + 1) * (D + 1) > M)) { set_error("SH degree %d needs %d coefficients, M=%d", D, (D + 1) * (D + 1), M); return -1; }
	const GeomLayout L = geom_layout(geom, P);
	if (L.total > geom_bytes) { set_error("geometry blob too small: %zu < %zu", geom_bytes, L.total); return -1; }
	if (bad_ptr(image, "image")) return -1;
	const ImgLayout I = img_layout(image, W, H);
	if (I.total > image_bytes) { set_error("image blob too small: %zu < %zu", image_bytes, I.total); return -1; }

	PreprocessArgs a;
	a.P = P; a.D = D; a.M = M; a.W = W; a.H = H;
	a.grid_x = (W + TILE_X - 1) / TILE_X; a.grid_y = (H + TILE_Y - 1) / TILE_Y;
	a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
	a.focal_y = H / (2.0f * tan_fovy); a.focal_x = W / (2.0f * tan_fovx);       // rasterizer_impl.cu:213-214
	a.scale_modifier = scale_modifier; a.low_pass = low_pass;
	a.means3D = means3D; a.scales = scales; a.rotations = rotations; a.opacities = opacities; a.shs = colors_precomp ? nullptr : shs;
	a.cov3D_precomp = cov3D_precomp; a.colors_precomp = colors_precomp; a.view = viewmatrix; a.proj = projmatrix; a.campos = campos;
	a.radii = radii; a.rec = L.rec; a.cov3D = L.cov3D; a.clamped = L.clamped; a.tiles_touched = L.tiles_touched; a.depth_key = L.key_a; a.rect = L.rect;
	a.tile_mask = L.tile_mask; a.tile_cull = L.tile_mask != nullptr;
	{ Stage st(stream, ST_PREPROCESS); if (launch_preprocess_fwd(stream, a)) return -1; }
	if (debug && check_launch("preprocess_fwd", stream, true)) return -1;
	if (plan_in_use(L, I, W, H)) {
		// tile plan: ranges and the tile sort's digit bases from the rects (needs nothing but the preprocess outputs)
		Stage st(stream, ST_TILE_PLAN);
		PlanSegs ps; memset(&ps, 0, sizeof(ps));
		ps.tiles_touched[0] = L.tiles_touched; ps.rect[0] = L.rect; ps.gdiff[0] = I.plan_ws; ps.ranges[0] = I.ranges; ps.ghist[0] = I.plan_bases;
		int shift[2] = { 0, 0 }; uint32_t mask[2] = { 0, 0 };
		const int npass = tile_sort_passes(tile_bits(a.grid_x * a.grid_y), shift, mask);
		if (launch_tile_plan(stream, P, a.grid_x, a.grid_y, 1, ps, npass, shift, mask)) return -1;
	}
	// depth order of the Gaussians: 32-bit keys, 4 passes, result in the *_a buffers
	int which;
	{ Stage st(stream, ST_DEPTH_SORT); which = sort_pairs<uint32_t>(stream, L.key_a, L.key_b, L.val_a, L.val_b, (size_t)P, 0, 32, L.sort_ws, true); }
	if (which != 0) { if (which > 0) set_error("internal: depth sort landed in the wrong buffer"); return -1; }
	if (debug && check_launch("depth sort", stream, true)) return -1;
	// num_rendered reaches the host as a u32 (the upper half of the int64 was zeroed above).  If the word is pinned host
	// memory (device-addressable under unified addressing) the scan kernel stores it there directly: no copy-engine
	// transfer that could queue behind the caller's own host->device traffic; otherwise a 4-byte async copy.
	uint32_t* mapped = nullptr;
	{
		cudaPointerAttributes pa;
		if (cudaPointerGetAttributes(&pa, num_rendered_host) == cudaSuccess && pa.type == cudaMemoryTypeHost && pa.devicePointer)
			mapped = reinterpret_cast<uint32_t*>(pa.devicePointer);
		else (void)cudaGetLastError();
	}
	{ Stage st(stream, ST_SCAN); if (launch_scan_tiles(stream, P, L.val_a, L.tiles_touched, L.offsets, L.block_sums, L.num_rendered, mapped)) return -1; }
	if (!mapped) cudaMemcpyAsync(num_rendered_host, L.num_rendered, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream);
	return check_launch("forward_geom", stream, debug != 0);
}

int rr_forward_geom_views(void* stream_, int P, int D, int M, int W, int H, int nviews,
	const float* const* viewmatrix, const float* const* projmatrix, const float* const* campos, const float* tan_fovx, const float* tan_fovy,
	const float* means3D, const float* shs, const float* f_rest, const float* opacities, const float* scales, float scale_modifier, const float* rotations, float low_pass,
	int* const* radii, void* const* geom, size_t geom_bytes, void* const* image, size_t image_bytes, int64_t* const* num_rendered_host, int flags, int debug)
{
	cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
	if (nviews == 0) return 0;
	if (nviews < 0 || nviews > PV_MAX_VIEWS) { set_error("rr_forward_geom_views: %d views (1..%d per call)", nviews, PV_MAX_VIEWS); return -1; }
	if (bad_ptr(num_rendered_host, "num_rendered_host")) return -1;
	for (int v = 0; v < nviews; v++) { if (bad_ptr(num_rendered_host[v], "num_rendered_host[v]")) return -1; *num_rendered_host[v] = 0; }
	if (P == 0) return 0;
	if (P < 0 || W <= 0 || H <= 0) { set_error("bad sizes P=%d W=%d H=%d", P, W, H); return -1; }
	if (M != 16 || D < 0 || D > 3) { set_error("rr_forward_geom_views needs M = 16 SH coefficients and degree 0..3 (M=%d, D=%d)", M, D); return -1; }
	if (tile_cull_wanted()) { set_error("rr_forward_geom_views does not support RAIN_TILE_CULL=1 (use rr_forward_geom per view)"); return -1; }
	if (bad_ptr(viewmatrix, "viewmatrix") || bad_ptr(projmatrix, "projmatrix") || bad_ptr(campos, "campos") || bad_ptr(tan_fovx, "tan_fovx") || bad_ptr(tan_fovy, "tan_fovy") ||
		bad_ptr(means3D, "means3D") || bad_ptr(shs, "shs") || bad_ptr(opacities, "opacities") || bad_ptr(scales, "scales") || bad_ptr(rotations, "rotations") ||
		bad_ptr(radii, "radii") || bad_ptr(geom, "geom") || bad_ptr(image, "image")) return -1;
	if (img_layout(nullptr, W, H).total > image_bytes) { set_error("image blob too small"); return -1; }
	const bool leaf = (flags & RR_GEOM_LEAF) != 0;
	if (leaf && bad_ptr(f_rest, "f_rest")) return -1;
	const GeomLayout L0 = geom_layout(nullptr, P);
	if (L0.total > geom_bytes) { set_error("geometry blob too small: %zu < %zu", geom_bytes, L0.total); return -1; }
	PreprocessViewsArgs a; memset(&a, 0, sizeof(a));
	a.P = P; a.D = D; a.W = W; a.H = H; a.grid_x = (W + TILE_X - 1) / TILE_X; a.grid_y = (H + TILE_Y - 1) / TILE_Y; a.nviews = nviews;
	a.write_cov3D = (flags & RR_GEOM_KEEP_COV3D) ? 1 : 0;
	a.scale_modifier = scale_modifier; a.low_pass = low_pass;
	a.means3D = means3D; a.scales = scales; a.rotations = rotations; a.opacities = opacities;
	a.leaf = leaf ? 1 : 0; a.shs = leaf ? nullptr : shs; a.f_dc = leaf ? shs : nullptr; a.f_rest = f_rest;
	// byte offsets of the arrays inside a geometry blob (the layout carved from a null base)
	a.off_rec = reinterpret_cast<uintptr_t>(L0.rec); a.off_cov3D = reinterpret_cast<uintptr_t>(L0.cov3D); a.off_clamped = reinterpret_cast<uintptr_t>(L0.clamped);
	a.off_tiles = reinterpret_cast<uintptr_t>(L0.tiles_touched); a.off_rect = reinterpret_cast<uintptr_t>(L0.rect); a.off_key = reinterpret_cast<uintptr_t>(L0.key_a);
	uint32_t *ka[PV_MAX_VIEWS], *kb[PV_MAX_VIEWS], *va[PV_MAX_VIEWS], *vb[PV_MAX_VIEWS];
	void* sws[PV_MAX_VIEWS];
	ScanSegs sg; memset(&sg, 0, sizeof(sg));
	PlanSegs ps; memset(&ps, 0, sizeof(ps));
	bool use_plan = true;
	for (int v = 0; v < nviews; v++) {
		if (bad_ptr(viewmatrix[v], "viewmatrix[v]") || bad_ptr(projmatrix[v], "projmatrix[v]") || bad_ptr(campos[v], "campos[v]") || bad_ptr(radii[v], "radii[v]") ||
			bad_ptr(geom[v], "geom[v]")) return -1;
		a.view[v] = viewmatrix[v]; a.proj[v] = projmatrix[v]; a.campos[v] = campos[v]; a.tan_fovx[v] = tan_fovx[v]; a.tan_fovy[v] = tan_fovy[v];
		a.radii[v] = radii[v]; a.geom[v] = reinterpret_cast<char*>(geom[v]);
		const GeomLayout L = geom_layout(geom[v], P);
		ka[v] = L.key_a; kb[v] = L.key_b; va[v] = L.val_a; vb[v] = L.val_b; sws[v] = L.sort_ws;
		if (bad_ptr(image[v], "image[v]")) return -1;
		const ImgLayout I = img_layout(image[v], W, H);
		use_plan = use_plan && plan_in_use(L, I, W, H);
		ps.tiles_touched[v] = L.tiles_touched; ps.rect[v] = L.rect; ps.gdiff[v] = I.plan_ws; ps.ranges[v] = I.ranges; ps.ghist[v] = I.plan_bases;
		uint32_t* mapped = nullptr;
		cudaPointerAttributes pa;
		if (cudaPointerGetAttributes(&pa, num_rendered_host[v]) == cudaSuccess && pa.type == cudaMemoryTypeHost && pa.devicePointer) mapped = reinterpret_cast<uint32_t*>(pa.devicePointer);
		else (void)cudaGetLastError();
		sg.order[v] = L.val_a; sg.tiles_touched[v] = L.tiles_touched; sg.status[v] = L.block_sums; sg.offsets[v] = L.offsets; sg.total[v] = L.num_rendered; sg.total_host[v] = mapped;
	}
	{ Stage st(stream, ST_PREPROCESS); if (launch_preprocess_views(stream, a)) return -1; }
	if (debug && check_launch("preprocess_views", stream, true)) return -1;
	if (use_plan) {
		Stage st(stream, ST_TILE_PLAN);
		int shift[2] = { 0, 0 }; uint32_t mask[2] = { 0, 0 };
		const int npass = tile_sort_passes(tile_bits(a.grid_x * a.grid_y), shift, mask);
		if (launch_tile_plan(stream, P, a.grid_x, a.grid_y, nviews, ps, npass, shift, mask)) return -1;
	}
	int which;
	{ Stage st(stream, ST_DEPTH_SORT); which = sort_pairs_segments<uint32_t>(stream, nviews, ka, kb, va, vb, (size_t)P, 0, 32, sws, true); }
	if (which != 0) { if (which > 0) set_error("internal: depth sort landed in the wrong buffer"); return -1; }
	if (debug && check_launch("depth sort", stream, true)) return -1;
	{ Stage st(stream, ST_SCAN); if (launch_scan_tiles_segments(stream, P, nviews, sg)) return -1; }
	for (int v = 0; v < nviews; v++)
		if (!sg.total_host[v]) cudaMemcpyAsync(num_rendered_host[v], sg.total[v], sizeof(uint32_t), cudaMemcpyDeviceToHost, stream);
	return check_launch("forward_geom_views", stream, debug != 0);
}

int rr_forward_bin(void* stream_, int P, int W, int H, int64_t N,
	void* geom, size_t geom_bytes, void* binning, size_t binning_bytes, void* image, size_t image_bytes, int debug)
{
	cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
	if (P == 0) return 0;
	if (bad_ptr(geom, "geom") || bad_ptr(binning, "binning") || bad_ptr(image, "image")) return -1;
	if (N < 0 || N >= (1ll << 30) - 1) { set_error("num_rendered=%lld out of range (the tile-duplicate count saturates at 2^30 - 1)", (long long)N); return -1; }
	const GeomLayout G = geom_layout(geom, P);
	const BinLayout B = bin_layout(binning, N, W, H);
	const ImgLayout I = img_layout(image, W, H);
	if (G.total > geom_bytes || B.total > binning_bytes || I.total > image_bytes) { set_error("scratch blob too small"); return -1; }
	const int grid_x = (W + TILE_X - 1) / TILE_X, grid_y = (H + TILE_Y - 1) / TILE_Y;

	// Default: 16-bit tile ids, digit bases (and ranges) from the geometry half's tile plan, value-only last pass (tile_sort16).  The
	// 32-bit-key kernels below remain for what it does not cover: the opt-in tile culling (per-Gaussian tile masks) and grids
	// beyond the tile plan's shared-memory budget.
	if (plan_in_use(G, I, W, H)) {
		Stage st(stream, ST_TILE_SORT);
		const int which = tile_sort16(stream, P, (size_t)N, B.bits, grid_x, G.val_a, G.offsets, G.rect, I.plan_bases,
			reinterpret_cast<uint16_t*>(B.key_a), reinterpret_cast<uint16_t*>(B.key_b), B.val_a, B.val_b, B.sort_ws);
		if (which < 0) return -1;
		if ((which ? B.val_b : B.val_a) != B.point_list) { set_error("internal: tile sort landed in the wrong buffer"); return -1; }
		return check_launch("forward_bin", stream, debug != 0);
	}
	if (N > 0) {
		{ Stage st(stream, ST_DUPLICATE); if (launch_duplicate(stream, P, G.val_a, G.offsets, G.rect, G.tile_mask, grid_x, B.key_a, B.val_a)) return -1; }
		if (debug && check_launch("duplicate", stream, true)) return -1;
		int which;
		{ Stage st(stream, ST_TILE_SORT); which = sort_pairs<uint32_t>(stream, B.key_a, B.key_b, B.val_a, B.val_b, (size_t)N, 0, B.bits, B.sort_ws, false); }
		if (which < 0) return -1;
		if ((which ? B.key_b : B.key_a) != B.keys_sorted) { set_error("internal: tile sort landed in the wrong buffer"); return -1; }
		if (debug && check_launch("tile sort", stream, true)) return -1;
	}
	{ Stage st(stream, ST_RANGES); if (launch_tile_ranges(stream, (size_t)N, B.keys_sorted, I.ranges, grid_x * grid_y)) return -1; }
	return check_launch("forward_bin", stream, debug != 0);
}

int rr_forward_render(void* stream_, int P, int W, int H, int64_t N, const float* background,
	void* geom, size_t geom_bytes, void* binning, size_t binning_bytes, void* image, size_t image_bytes,
	float* out_color, float* out_depth, int debug)
{
	cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
	if (P == 0) return 0;
	if (bad_ptr(background, "background") || bad_ptr(geom, "geom") || bad_ptr(binning, "binning") || bad_ptr(image, "image") ||
		bad_ptr(out_color, "out_color") || bad_ptr(out_depth, "out_depth")) return -1;
	if (N < 0 || N >= (1ll << 30) - 1) { set_error("num_rendered=%lld out of range (the tile-duplicate count saturates at 2^30 - 1)", (long long)N); return -1; }
	const GeomLayout G = geom_layout(geom, P);
	const BinLayout B = bin_layout(binning, N, W, H);
	const ImgLayout I = img_layout(image, W, H);
	if (G.total > geom_bytes || B.total > binning_bytes || I.total > image_bytes) { set_error("scratch blob too small"); return -1; }
	const int grid_x = (W + TILE_X - 1) / TILE_X, grid_y = (H + TILE_Y - 1) / TILE_Y;

	RenderArgs r; memset(&r, 0, sizeof(r));
	r.W = W; r.H = H; r.grid_x = grid_x; r.grid_y = grid_y;
	r.ranges = I.ranges; r.rec = G.rec; r.point_list = B.point_list; r.bg = background; r.final_T = I.final_T; r.n_contrib = I.n_contrib; r.warp_flags = B.warp_flags;
	r.out_color = out_color; r.out_depth = out_depth; r.xc = XC_COUNTERS;
	r.zero = reinterpret_cast<float4*>(G.acc); r.zero_n16 = (uint32_t)((size_t)P * ACC_FLOATS / 4);
	{ Stage st(stream, ST_RENDER_FWD); if (launch_render_fwd(stream, r)) return -1; }
	return check_launch("forward_render", stream, debug != 0);
}

int rr_forward_raster(void* stream, int P, int W, int H, int64_t N, const float* background,
	void* geom, size_t geom_bytes, void* binning, size_t binning_bytes, void* image, size_t image_bytes,
	float* out_color, float* out_depth, int debug)
{
	if (P == 0) return 0;
	if (bad_ptr(background, "background") || bad_ptr(out_color, "out_color") || bad_ptr(out_depth, "out_depth")) return -1;
	if (rr_forward_bin(stream, P, W, H, N, geom, geom_bytes, binning, binning_bytes, image, image_bytes, debug)) return -1;
	return rr_forward_render(stream, P, W, H, N, background, geom, geom_bytes, binning, binning_bytes, image, image_bytes, out_color, out_depth, debug);
}

// B1: zero the per-Gaussian accumulators and run the backward tile renderer (shared by rr_backward and rr_backward_render)
static int backward_render_impl(cudaStream_t stream, int P, int W, int H, int64_t N, const float* background,
	void* geom, size_t geom_bytes, void* binning, size_t binning_bytes, void* image, size_t image_bytes,
	const float* dL_dpix, void* ws, size_t ws_bytes, int debug)
{
	const GeomLayout G = geom_layout(geom, P);
	const BinLayout B = bin_layout(binning, N, W, H);
	const ImgLayout I = img_layout(image, W, H);
	if (G.total > geom_bytes || B.total > binning_bytes || I.total > image_bytes || (ws && rr_backward_ws_bytes(P) > ws_bytes + 256)) { set_error("scratch blob too small"); return -1; }
	// ws == null: the accumulator region of the geometry blob, which rr_forward_render of this view zero-filled while it rendered
	// (valid for ONE backward per forward); otherwise the caller's scratch, zero-filled here
	float* grad_acc = ws ? reinterpret_cast<float*>(ws) : G.acc;
	RenderArgs r; memset(&r, 0, sizeof(r));
	r.W = W; r.H = H; r.grid_x = (W + TILE_X - 1) / TILE_X; r.grid_y = (H + TILE_Y - 1) / TILE_Y;
	r.ranges = I.ranges; r.rec = G.rec; r.point_list = B.point_list; r.bg = background; r.final_T = I.final_T; r.n_contrib = I.n_contrib; r.warp_flags = B.warp_flags;
	r.dL_dpix = dL_dpix; r.grad_acc = grad_acc; r.xc = XC_COUNTERS;
	// the accumulators start from zero (the fill is part of the render_bwd stage time)
	Stage st(stream, ST_RENDER_BWD);
	if (ws) cudaMemsetAsync(grad_acc, 0, (size_t)P * ACC_FLOATS * sizeof(float), stream);
	if (N > 0 && launch_render_bwd(stream, r)) return -1;
	if (debug && check_launch("render_bwd", stream, true)) return -1;
	return 0;
}

int rr_backward_render(void* stream_, int P, int W, int H, int64_t N, const float* background,
	void* geom, size_t geom_bytes, void* binning, size_t binning_bytes, void* image, size_t image_bytes,
	const float* dL_dpix, void* ws, size_t ws_bytes, int debug)
{
	cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
	if (P == 0) return 0;
	if (bad_ptr(background, "background") || bad_ptr(geom, "geom") || bad_ptr(binning, "binning") || bad_ptr(image, "image") ||
		bad_ptr(dL_dpix, "dL_dpix")) return -1;
	if (backward_render_impl(stream, P, W, H, N, background, geom, geom_bytes, binning, binning_bytes, image, image_bytes, dL_dpix, ws, ws_bytes, debug)) return -1;
	return check_launch("backward_render", stream, debug != 0);
}

int rr_backward(void* stream_, int P, int D, int M, int W, int H, int64_t N, const float* background,
	const float* means3D, const float* shs, const float* colors_precomp,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tan_fovx, float tan_fovy, float low_pass, const int* radii,
	void* geom, size_t geom_bytes, void* binning, size_t binning_bytes, void* image, size_t image_bytes,
	const float* dL_dpix, void* ws, size_t ws_bytes,
	float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
	float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
	int debug, int flags)
{
	cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
	if (P == 0) return 0;                                                  // rasterize_points.cu:155
	if (bad_ptr(background, "background") || bad_ptr(means3D, "means3D") || bad_ptr(radii, "radii") || bad_ptr(geom, "geom") ||
		bad_ptr(binning, "binning") || bad_ptr(image, "image") || bad_ptr(dL_dpix, "dL_dpix") ||
		bad_ptr(dL_dmean2D, "dL_dmean2D") || bad_ptr(dL_dopacity, "dL_dopacity") || bad_ptr(dL_dmean3D, "dL_dmean3D")) return -1;
	const bool use_sh = shs && !colors_precomp;
	// the gradient of an input the caller did not pass is an intermediate: it may be omitted
	if (!use_sh && bad_ptr(dL_dcolor, "dL_dcolor")) return -1;
	if (cov3D_precomp && bad_ptr(dL_dcov3D, "dL_dcov3D")) return -1;
	if (use_sh && bad_ptr(dL_dsh, "dL_dsh")) return -1;
	if (backward_render_impl(stream, P, W, H, N, background, geom, geom_bytes, binning, binning_bytes, image, image_bytes, dL_dpix, ws, ws_bytes, debug)) return -1;
	const GeomLayout G = geom_layout(geom, P);

	PreprocessBwdArgs a; memset(&a, 0, sizeof(a));
	a.P = P; a.D = D; a.M = M; a.W = W; a.H = H;
	a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy; a.focal_y = H / (2.0f * tan_fovy); a.focal_x = W / (2.0f * tan_fovx);
	a.scale_modifier = scale_modifier; a.low_pass = low_pass;
	a.means3D = means3D; a.scales = cov3D_precomp ? nullptr : scales; a.rotations = cov3D_precomp ? nullptr : rotations;
	a.shs = use_sh ? shs : nullptr; a.cov3D = cov3D_precomp ? cov3D_precomp : G.cov3D;        // rasterizer_impl.cu:406
	a.view = viewmatrix; a.proj = projmatrix; a.campos = campos; a.has_colors_precomp = colors_precomp != nullptr; a.accumulate = (flags & RR_BWD_ACCUMULATE) ? 1 : 0;
	a.radii = radii; a.clamped = G.clamped; a.rec = G.rec; a.grad_acc = ws ? reinterpret_cast<const float*>(ws) : G.acc;      // (ws == null: see rr_backward_render)
	a.dL_dmean2D = dL_dmean2D; a.dL_dconic = dL_dconic; a.dL_dopacity = dL_dopacity; a.dL_dcolor = dL_dcolor; a.dL_dmean3D = dL_dmean3D;
	a.dL_dcov3D = dL_dcov3D; a.dL_dsh = dL_dsh; a.dL_dscale = dL_dscale; a.dL_drot = dL_drot;
	{ Stage st(stream, ST_PREPROCESS_BWD); if (launch_preprocess_bwd(stream, a)) return -1; }
	return check_launch("backward", stream, debug != 0);
}

int rr_backward_views(void* stream_, int P, int D, int M, int W, int H, int nviews,
	const float* const* viewmatrix, const float* const* projmatrix, const float* const* campos, const float* tan_fovx, const float* tan_fovy,
	const int* const* radii, void* const* geom, const float* const* grad_acc, float* const* dL_dmean2D,
	const float* means3D, const float* shs_or_f_dc, const float* f_rest, const float* scales, float scale_modifier, const float* rotations, float low_pass,
	float* dL_dmean3D, float* dL_dsh_or_f_dc, float* dL_df_rest, float* dL_dopacity, float* dL_dscale, float* dL_drot,
	float* stat_max_radii2D, float* stat_grad_accum, float* stat_denom, int first, int count, int flags, int debug)
{
	cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
	if (P == 0 || nviews == 0 || count == 0) return 0;
	if (first < 0 || count < 0 || (int64_t)first + count > P) { set_error("bad Gaussian range [%d, %d + %d) of %d", first, first, count, P); return -1; }
	if (P < 0 || nviews < 0 || W <= 0 || H <= 0) { set_error("bad sizes P=%d nviews=%d W=%d H=%d", P, nviews, W, H); return -1; }
	const bool leaf = (flags & RR_BWD_LEAF) != 0;
	if (M != 16 || D < 0 || D > 3) { set_error("rr_backward_views needs M = 16 SH coefficients and degree 0..3 (M=%d, D=%d)", M, D); return -1; }
	if (bad_ptr(viewmatrix, "viewmatrix") || bad_ptr(projmatrix, "projmatrix") || bad_ptr(campos, "campos") || bad_ptr(tan_fovx, "tan_fovx") ||
		bad_ptr(tan_fovy, "tan_fovy") || bad_ptr(radii, "radii") || bad_ptr(geom, "geom") || bad_ptr(grad_acc, "grad_acc") ||
		bad_ptr(means3D, "means3D") || bad_ptr(shs_or_f_dc, leaf ? "f_dc" : "shs") || bad_ptr(scales, "scales") || bad_ptr(rotations, "rotations") ||
		bad_ptr(dL_dmean3D, "dL_dmean3D") || bad_ptr(dL_dsh_or_f_dc, "dL_dsh") || bad_ptr(dL_dopacity, "dL_dopacity") || bad_ptr(dL_dscale, "dL_dscale") ||
		bad_ptr(dL_drot, "dL_drot")) return -1;
	if (leaf && (bad_ptr(f_rest, "f_rest") || bad_ptr(dL_df_rest, "dL_df_rest"))) return -1;
	if ((stat_max_radii2D || stat_grad_accum || stat_denom) && !(stat_max_radii2D && stat_grad_accum && stat_denom)) { set_error("pass all three statistics or none"); return -1; }
	Stage st(stream, ST_BACKWARD_VIEWS);
	for (int v0 = 0; v0 < nviews; v0 += BV_MAX_VIEWS) {
		BackwardViewsArgs a; memset(&a, 0, sizeof(a));
		a.P = P; a.D = D; a.M = M; a.W = W; a.H = H; a.nviews = nviews - v0 < BV_MAX_VIEWS ? nviews - v0 : BV_MAX_VIEWS;
		a.first = first; a.count = count; a.leave_room = (flags & RR_BWD_LEAVE_ROOM) ? 1 : 0;
		a.scale_modifier = scale_modifier; a.low_pass = low_pass; a.leaf = leaf ? 1 : 0;
		a.accumulate = ((flags & RR_BWD_ACCUMULATE) || v0 > 0) ? 1 : 0;      // later chunks of a long batch add to the first one's rows
		a.means3D = means3D; a.scales = scales; a.rotations = rotations; a.shs = leaf ? nullptr : shs_or_f_dc; a.f_dc = leaf ? shs_or_f_dc : nullptr; a.f_rest = f_rest;
		for (int j = 0; j < a.nviews; j++) {
			const int v = v0 + j;
			if (bad_ptr(viewmatrix[v], "viewmatrix[v]") || bad_ptr(projmatrix[v], "projmatrix[v]") || bad_ptr(campos[v], "campos[v]") || bad_ptr(radii[v], "radii[v]") ||
				bad_ptr(geom[v], "geom[v]") || bad_ptr(grad_acc[v], "grad_acc[v]")) return -1;
			const GeomLayout G = geom_layout(geom[v], P);
			a.view[j] = viewmatrix[v]; a.proj[j] = projmatrix[v]; a.campos[j] = campos[v]; a.tan_fovx[j] = tan_fovx[v]; a.tan_fovy[j] = tan_fovy[v];
			a.radii[j] = radii[v]; a.rec[j] = G.rec; a.clamped[j] = G.clamped; a.grad_acc[j] = grad_acc[v];
			a.dL_dmean2D[j] = dL_dmean2D ? dL_dmean2D[v] : nullptr;
		}
		a.dL_dmean3D = dL_dmean3D; a.dL_dsh = dL_dsh_or_f_dc; a.dL_dfrest = dL_df_rest; a.dL_dopacity = dL_dopacity; a.dL_dscale = dL_dscale; a.dL_drot = dL_drot;
		a.stat_max_radii = stat_max_radii2D; a.stat_grad_accum = stat_grad_accum; a.stat_denom = stat_denom;
		if (launch_backward_views(stream, a)) return -1;
	}
	return check_launch("backward_views", stream, debug != 0);
}

#ifdef RR_CROSSCHECK
/* cross-check build only (not declared in include/rain_rast.h): copy the 16 work counters to the host, optionally reset them */
int rr_xc_counters(unsigned long long* out_host, int reset)
{
	unsigned long long* d = xc_counters_dev();
	if (!d) { set_error("rr_xc_counters: no counter buffer"); return -1; }
	if (cudaDeviceSynchronize() != cudaSuccess || cudaMemcpy(out_host, d, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost) != cudaSuccess) { set_error("rr_xc_counters: copy failed"); return -1; }
	if (reset) cudaMemset(d, 0, 16 * sizeof(unsigned long long));
	return 0;
}
#endif

int rr_selftest_exp(void* stream_, uint32_t first_bits, uint32_t count, unsigned long long* mismatches_dev)
{
	if (bad_ptr(mismatches_dev, "mismatches_dev")) return -1;
	if (count == 0) return 0;
	return launch_exp_selftest(reinterpret_cast<cudaStream_t>(stream_), first_bits, count, mismatches_dev);
}

int rr_mark_visible(void* stream_, int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present)
{
	(void)projmatrix;   // in_frustum computes p_proj but only tests view-space z (auxiliary.h:137-143)
	if (P == 0) return 0;
	if (bad_ptr(means3D, "means3D") || bad_ptr(viewmatrix, "viewmatrix") || bad_ptr(present, "present")) return -1;
	return launch_mark_visible(reinterpret_cast<cudaStream_t>(stream_), P, means3D, viewmatrix, present);
}

size_t rr_loss_ws_bytes(int C, int H, int W) { return (C > 0 && H > 0 && W > 0) ? loss_ws_bytes(C, H, W) : 256; }

int rr_loss_l1_ssim(void* stream_, int C, int H, int W, const float* image, const float* gt, float lambda_dssim,
	void* ws, size_t ws_bytes, float* out3, float* dL_dimage)
{
	if (C <= 0 || H <= 0 || W <= 0) { set_error("bad image size C=%d H=%d W=%d", C, H, W); return -1; }
	if (bad_ptr(image, "image") || bad_ptr(gt, "gt") || bad_ptr(ws, "ws")) return -1;
	if (!out3 && !dL_dimage) { set_error("nothing to compute: out3 and dL_dimage are both null"); return -1; }
	if (ws_bytes < loss_ws_bytes(C, H, W)) { set_error("loss workspace too small: %zu < %zu", ws_bytes, loss_ws_bytes(C, H, W)); return -1; }
	return launch_loss_l1_ssim(reinterpret_cast<cudaStream_t>(stream_), C, H, W, image, gt, lambda_dssim, ws, out3, dL_dimage);
}

int rr_loss_l2(void* stream_, int64_t n, const float* image, const float* gt, float* dL_dimage, float* loss_accum)
{
	if (n < 0) { set_error("bad element count %lld", (long long)n); return -1; }
	if (n == 0) return 0;
	if (bad_ptr(image, "image") || bad_ptr(gt, "gt") || bad_ptr(dL_dimage, "dL_dimage")) return -1;
	return launch_loss_l2(reinterpret_cast<cudaStream_t>(stream_), (size_t)n, image, gt, 0, 1.0f, dL_dimage, loss_accum);
}

int rr_loss_l2_u8(void* stream_, int64_t n, const float* image, const uint8_t* gt_u8, float scale, float* dL_dimage, float* loss_accum)
{
	if (n < 0) { set_error("bad element count %lld", (long long)n); return -1; }
	if (n == 0) return 0;
	if (bad_ptr(image, "image") || bad_ptr(gt_u8, "gt_u8") || bad_ptr(dL_dimage, "dL_dimage")) return -1;
	return launch_loss_l2(reinterpret_cast<cudaStream_t>(stream_), (size_t)n, image, gt_u8, 1, scale, dL_dimage, loss_accum);
}

int rr_param_step(void* stream_, int P, int M, float* xyz, float* f_dc, float* f_rest, float* opacity, float* scaling, float* rotation,
	float* exp_avg, float* exp_avg_sq, const float* grads, const float* hyper, int step,
	float* shs_out, float* opacities_out, float* scales_out, float* rotations_out)
{
	if (P == 0) return 0;
	if (P < 0 || M < 1) { set_error("bad sizes P=%d M=%d", P, M); return -1; }
	if (step < 1) { set_error("Adam step count must be >= 1, got %d", step); return -1; }
	if (bad_ptr(xyz, "xyz") || bad_ptr(f_dc, "f_dc") || bad_ptr(opacity, "opacity") || bad_ptr(scaling, "scaling") || bad_ptr(rotation, "rotation") ||
		bad_ptr(exp_avg, "exp_avg") || bad_ptr(exp_avg_sq, "exp_avg_sq") || bad_ptr(grads, "grads") || bad_ptr(hyper, "hyper")) return -1;
	if (M > 1 && bad_ptr(f_rest, "f_rest")) return -1;
	return launch_param_step(reinterpret_cast<cudaStream_t>(stream_), P, M, xyz, f_dc, f_rest, opacity, scaling, rotation, exp_avg, exp_avg_sq, grads, hyper, step,
		shs_out, opacities_out, scales_out, rotations_out);
}

int rr_param_step_leaf(void* stream_, int P, int M, float* xyz, float* f_dc, float* f_rest, float* opacity, float* scaling, float* rotation,
	float* exp_avg, float* exp_avg_sq, const float* grads, const float* hyper, int step)
{
	if (P == 0) return 0;
	if (P < 0 || M < 1) { set_error("bad sizes P=%d M=%d", P, M); return -1; }
	if (step < 1) { set_error("Adam step count must be >= 1, got %d", step); return -1; }
	if (bad_ptr(xyz, "xyz") || bad_ptr(f_dc, "f_dc") || bad_ptr(opacity, "opacity") || bad_ptr(scaling, "scaling") || bad_ptr(rotation, "rotation") ||
		bad_ptr(exp_avg, "exp_avg") || bad_ptr(exp_avg_sq, "exp_avg_sq") || bad_ptr(grads, "grads") || bad_ptr(hyper, "hyper")) return -1;
	if (M > 1 && bad_ptr(f_rest, "f_rest")) return -1;
	return launch_param_step_leaf(reinterpret_cast<cudaStream_t>(stream_), P, M, xyz, f_dc, f_rest, opacity, scaling, rotation, exp_avg, exp_avg_sq, grads, hyper, step);
}

int rr_activate(void* stream_, int P, const float* opacity, const float* scaling, const float* rotation, float* opacities_out, float* scales_out, float* rotations_out)
{
	if (P == 0) return 0;
	if (P < 0) { set_error("bad size P=%d", P); return -1; }
	if (bad_ptr(opacity, "opacity") || bad_ptr(scaling, "scaling") || bad_ptr(rotation, "rotation") || bad_ptr(opacities_out, "opacities_out") ||
		bad_ptr(scales_out, "scales_out") || bad_ptr(rotations_out, "rotations_out")) return -1;
	return launch_activate(reinterpret_cast<cudaStream_t>(stream_), P, opacity, scaling, rotation, opacities_out, scales_out, rotations_out);
}

int rr_densify_stats(void* stream_, int P, const int* radii, const float* dL_dmean2D, float* max_radii2D, float* xyz_gradient_accum, float* denom)
{
	if (P == 0) return 0;
	if (P < 0) { set_error("bad size P=%d", P); return -1; }
	if (bad_ptr(radii, "radii") || bad_ptr(dL_dmean2D, "dL_dmean2D") || bad_ptr(max_radii2D, "max_radii2D") ||
		bad_ptr(xyz_gradient_accum, "xyz_gradient_accum") || bad_ptr(denom, "denom")) return -1;
	return launch_densify_stats(reinterpret_cast<cudaStream_t>(stream_), P, radii, dL_dmean2D, max_radii2D, xyz_gradient_accum, denom);
}

size_t rr_allreduce_signal_bytes(void) { return sizeof(uint32_t) * AR_MAX_CTAS * AR_MAX_WORLD + 256; }

int rr_allreduce_sum_f32(void* stream_, int rank, int world, int64_t n, void* const* buffers, void* multicast, void* const* signals,
	uint32_t* error_flag, uint32_t epoch, int ctas)
{
	const int64_t off = 0;
	return rr_allreduce_sum_f32_ranges(stream_, rank, world, 1, &off, &n, buffers, multicast, signals, error_flag, epoch, ctas, 0);
}

int rr_allreduce_sum_f32_ranges(void* stream_, int rank, int world, int nranges, const int64_t* offsets, const int64_t* lengths, void* const* buffers, void* multicast,
	void* const* signals, uint32_t* error_flag, uint32_t epoch, int ctas, int threads)
{
	if (world <= 1 || nranges == 0) return 0;
	if (world > AR_MAX_WORLD || rank < 0 || rank >= world || nranges < 0 || nranges > AR_MAX_RANGES) {
		set_error("rr_allreduce_sum_f32: bad rank/world/ranges %d/%d/%d (at most %d ranks, %d ranges)", rank, world, nranges, AR_MAX_WORLD, AR_MAX_RANGES); return -1;
	}
	if (bad_ptr(offsets, "offsets") || bad_ptr(lengths, "lengths") || bad_ptr(buffers, "buffers") || bad_ptr(signals, "signals") || bad_ptr(error_flag, "error_flag")) return -1;
	if (ctas <= 0) ctas = 148;
	if (ctas > AR_MAX_CTAS) ctas = AR_MAX_CTAS;
	AllReduceArgs a; memset(&a, 0, sizeof(a));
	a.multicast = reinterpret_cast<float*>(multicast); a.error_flag = error_flag; a.rank = rank; a.world = world; a.epoch = epoch;
	a.nranges = nranges;
	int64_t total = 0;
	for (int r = 0; r < nranges; r++) {
		if (offsets[r] < 0 || lengths[r] < 0 || (nranges > 1 && ((offsets[r] | lengths[r]) & 3))) { set_error("rr_allreduce_sum_f32: ranges must be non-negative and, if several, multiples of 4 floats"); return -1; }
		if (offsets[r] & 3) { set_error("rr_allreduce_sum_f32: range offsets must be multiples of 4 floats"); return -1; }
		a.off[r] = offsets[r]; a.len[r] = lengths[r]; total += lengths[r];
	}
	if (total == 0) return 0;
	for (int r = 0; r < world; r++) {
		if (bad_ptr(buffers[r], "buffers[r]") || bad_ptr(signals[r], "signals[r]")) return -1;
		if (reinterpret_cast<uintptr_t>(buffers[r]) & 15) { set_error("rr_allreduce_sum_f32: buffers must be 16-byte aligned"); return -1; }
		a.buffers[r] = reinterpret_cast<float*>(buffers[r]); a.signal[r] = reinterpret_cast<uint32_t*>(signals[r]);
	}
	return launch_allreduce(reinterpret_cast<cudaStream_t>(stream_), a, ctas, threads);
}

size_t rr_dist2_ws_bytes(int P) { return knn_ws_bytes(P); }
int rr_dist2(void* stream_, int P, const float* points, float* out, void* ws, size_t ws_bytes)
{
	if (P == 0) return 0;
	if (bad_ptr(points, "points") || bad_ptr(out, "out") || bad_ptr(ws, "ws")) return -1;
	Stage st(reinterpret_cast<cudaStream_t>(stream_), ST_DIST2);
	return launch_dist2(reinterpret_cast<cudaStream_t>(stream_), P, points, out, ws, ws_bytes);
}

size_t rr_sort_ws_bytes(int64_t n, int key_bytes, int begin_bit, int end_bit)
{
	return key_bytes == 8 ? sort_ws_bytes<uint64_t>((size_t)n, begin_bit, end_bit) : sort_ws_bytes<uint32_t>((size_t)n, begin_bit, end_bit);
}

}  // extern "C"

template <class K>
static int sort_entry(void* stream_, int64_t n, K* keys_in, K* keys_out, uint32_t* vals_in, uint32_t* vals_out, int begin_bit, int end_bit, void* ws)
{
	cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
	if (n == 0) return 0;
	if (bad_ptr(keys_in, "keys_in") || bad_ptr(keys_out, "keys_out") || bad_ptr(vals_out, "vals_out") || bad_ptr(ws, "ws")) return -1;
	uint32_t* va = vals_in;
	const bool iota = vals_in == nullptr;
	if (iota) { set_error("vals_in must be a scratch buffer even for identity values"); return -1; }
	const int which = sort_pairs<K>(stream, keys_in, keys_out, va, vals_out, (size_t)n, begin_bit, end_bit, ws, false);
	if (which < 0) return -1;
	if (which == 0) {
		cudaMemcpyAsync(keys_out, keys_in, sizeof(K) * (size_t)n, cudaMemcpyDeviceToDevice, stream);
		cudaMemcpyAsync(vals_out, va, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToDevice, stream);
	}
	return check_launch("sort_pairs", stream, false);
}

extern "C" {

int rr_sort_pairs_u32(void* stream, int64_t n, uint32_t* keys_in, uint32_t* keys_out, uint32_t* vals_in, uint32_t* vals_out, int begin_bit, int end_bit, void* ws)
{ return sort_entry<uint32_t>(stream, n, keys_in, keys_out, vals_in, vals_out, begin_bit, end_bit, ws); }
int rr_sort_pairs_u64(void* stream, int64_t n, uint64_t* keys_in, uint64_t* keys_out, uint32_t* vals_in, uint32_t* vals_out, int begin_bit, int end_bit, void* ws)
{ return sort_entry<uint64_t>(stream, n, keys_in, keys_out, vals_in, vals_out, begin_bit, end_bit, ws); }

int rr_export_sorted_keys(void* stream_, int P, int64_t N, int W, int H, void* geom, void* binning, void* image, uint64_t* keys_out)
{
	if (N == 0) return 0;
	const GeomLayout G = geom_layout(geom, P);
	const BinLayout B = bin_layout(binning, N, W, H);
	const ImgLayout I = img_layout(image, W, H);
	// tile ids from the tile ranges, depth bits from rec.c.y (float index 9 of the 12-float record)
	const int tiles = ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
	return launch_export_keys(reinterpret_cast<cudaStream_t>(stream_), tiles, I.ranges, B.point_list, reinterpret_cast<const uint32_t*>(G.rec), keys_out);
}

}  // extern "C"
