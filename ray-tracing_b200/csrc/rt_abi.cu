// rt_abi.cu — host side of librt_b200.so: the C ABI declared in include/rt_abi.h.
// Validates the flattened scene, packs it into the device layout of rt_device.cuh (fp32 or fp64), launches the
// kernels and moves framebuffers.  There is no CPU rendering path in this library: without a CUDA device every
// compute entry point returns RT_ERR_CUDA.
#include "../../include/rt_abi.h"
#include "rt_device.cuh"
#include "rt_launch.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

using namespace rtb;

namespace {

thread_local std::string g_err;
int fail(int code, const std::string& msg) { g_err = msg; return code; }
#define CU(call)                                                                                             \
  do {                                                                                                       \
    cudaError_t e_ = (call);                                                                                 \
    if (e_ != cudaSuccess) return fail(RT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));     \
  } while (0)

// deep copy of the caller's RtSceneDesc (pointers are only borrowed during rt_scene_create)
struct HostScene {
  std::vector<double> node_box, sphere_cr, moving, rect, cuboid, translate_off, rotate_sc, medium_nid, mat_param, tex_rgb, tex_scale, perlin_ranvec;
  std::vector<uint32_t> node_left, node_right, node_size, sphere_mat, moving_mat, rect_mat, cuboid_mat, translate_child, rotate_axis, rotate_child,
      medium_mat, medium_boundary, mat_kind, mat_tex, tex_kind, tex_a, tex_b, image_w, image_h;
  std::vector<int32_t> perlin_perm;
  std::vector<std::vector<uint8_t>> images;
  uint32_t n_nodes = 0, n_spheres = 0, n_moving = 0, n_rects = 0, n_cuboids = 0, n_translates = 0, n_rotates = 0, n_media = 0, n_materials = 0,
           n_textures = 0, n_perlin = 0, n_images = 0;
  uint32_t world_root = 0, lights_root = 0;
  double background[3] = {0, 0, 0}, camera[24] = {0};
  std::vector<uint32_t> node_lsize, node_rsize;  // htblSize of the children (Lib.hs:662-671)
  int stackDepth = 1;
  uint32_t lightCap = 0;  // entries of the light-pdf walk stack (rt_traverse.cuh)
};

template <class T> struct DevBuffers {
  std::vector<void*> allocs;
  DScene<T> sc{};
  bool ready = false;
  ~DevBuffers() { for (void* p : allocs) cudaFree(p); }
};

struct SceneEntry {
  std::mutex mu;  // calls on one handle are serialised: they share the device scene, the work counter and the staging buffers
  HostScene hs;
  int device = 0, numSMs = 148;
  DevBuffers<float> f32;
  DevBuffers<double> f64;
  unsigned long long* d_counter = nullptr;  // [0] path counter, [1..16] work counters, [17..32] scheduler statistics
  double* d_fb = nullptr; size_t fbElems = 0;
  double* h_fb = nullptr; size_t hfbElems = 0;  // pinned staging
  ~SceneEntry() {
    if (d_counter) cudaFree(d_counter);
    if (d_fb) cudaFree(d_fb);
    if (h_fb) cudaFreeHost(h_fb);
  }
};

std::mutex g_mu;
std::map<uint64_t, std::shared_ptr<SceneEntry>> g_scenes;
uint64_t g_nextId = 1;

template <class V, class P> void copyIn(V& dst, const P* src, size_t n) {
  dst.assign(n, typename V::value_type{});
  if (n) std::memcpy(dst.data(), src, n * sizeof(typename V::value_type));
}

bool refOk(const HostScene& h, uint32_t ref) {
  uint32_t k = ref >> 28, i = ref & 0x0fffffffu;  // full 28-bit field: counts are limited to 27 bits, so a set bit 27 is out of range
  switch (k) {
    case K_UNHITTABLE: return true;
    case K_SPHERE: return i < h.n_spheres;
    case K_MOVING: return i < h.n_moving;
    case K_RECT_XY: case K_RECT_XZ: case K_RECT_YZ: return i < h.n_rects;
    case K_CUBOID: return i < h.n_cuboids;
    case K_BVH: return i < h.n_nodes;
    case K_TRANSLATE: return i < h.n_translates;
    case K_ROTATE: return i < h.n_rotates;
    case K_MEDIUM: return i < h.n_media;
    default: return false;
  }
}

// Structural analysis of the hittable DAG: traversal stack need (push/pop discipline of rtb::traverse), open
// frames, medium nesting, htblSize; detects cycles.  Memoised per ref.
struct Analysis {
  const HostScene& h;
  std::map<uint32_t, int> stackNeed, frames, media, size;
  std::map<uint32_t, int> state;  // 1 = in progress, 2 = done
  bool cycle = false;
  explicit Analysis(const HostScene& hs) : h(hs) {}
  void go(uint32_t ref) {
    if (cycle) return;
    int& st = state[ref];
    if (st == 2) return;
    if (st == 1) { cycle = true; return; }
    st = 1;
    uint32_t k = ref >> 28, i = ref & kIdxMask;
    int sn = 0, fr = 0, md = 0, sz = 1;
    if (k == K_BVH) {
      uint32_t l = h.node_left[i], r = h.node_right[i];
      go(l); go(r);
      if (cycle) return;
      sn = std::max(2, std::max(1 + stackNeed[l], stackNeed[r]));
      fr = std::max(frames[l], frames[r]); md = std::max(media[l], media[r]); sz = (int)h.node_size[i];
    } else if (k == K_TRANSLATE || k == K_ROTATE) {
      uint32_t c = k == K_TRANSLATE ? h.translate_child[i] : h.rotate_child[i];
      go(c);
      if (cycle) return;
      sn = std::max(2, 1 + stackNeed[c]); fr = 1 + frames[c]; md = media[c]; sz = size[c];
    } else if (k == K_MEDIUM) {
      uint32_t c = h.medium_boundary[i];
      go(c);
      if (cycle) return;
      sn = std::max(2, 1 + stackNeed[c]); fr = frames[c]; md = 1 + media[c]; sz = 1;
    } else if (k == K_UNHITTABLE) {
      sz = 0;
    }
    stackNeed[ref] = sn; frames[ref] = fr; media[ref] = md; size[ref] = sz;
    state[ref] = 2;
  }
};
int lightDepth(const HostScene& h, uint32_t ref, int d = 0) {
  if (d > 64) return d;
  if ((ref >> 28) != K_BVH) return 1;
  uint32_t i = ref & kIdxMask;
  return 1 + std::max(lightDepth(h, h.node_left[i], d + 1), lightDepth(h, h.node_right[i], d + 1));
}

template <class T> T fromBits(uint32_t b);
template <> float fromBits<float>(uint32_t b) { float f; std::memcpy(&f, &b, 4); return f; }
template <> double fromBits<double>(uint32_t b) { uint64_t w = b; double d; std::memcpy(&d, &w, 8); return d; }

template <class T, class V> int upload(DevBuffers<T>& db, const std::vector<V>& host, const V** out) {
  *out = nullptr;
  if (host.empty()) return RT_OK;
  void* p = nullptr;
  CU(cudaMalloc(&p, host.size() * sizeof(V)));
  db.allocs.push_back(p);
  CU(cudaMemcpy(p, host.data(), host.size() * sizeof(V), cudaMemcpyHostToDevice));
  *out = static_cast<const V*>(p);
  return RT_OK;
}

// which leaf classes are tested as pairs (bit 0 sphere-like, bit 1 rect / cuboid): sphere pairs always pay; rect / cuboid pairs pay
// in cuboid-heavy scenes (cfg3 +3 %) and cost in rect scenes (cfg4 -9 %).  RTB_PAIR overrides for experiments.
static int pairMask(const HostScene& h) {
  const char* v = std::getenv("RTB_PAIR");
  return v ? std::atoi(v) : (h.n_cuboids > h.n_rects ? 3 : 1);
}

// pack the flattened scene into the device layout (rt_device.cuh DScene<T>); doubles are rounded to T here
template <class T> int buildDevice(const HostScene& h, DevBuffers<T>& db) {
  if (db.ready) return RT_OK;
  using Q = Q4<T>;
  auto B = [](uint32_t b) { return fromBits<T>(b); };
  std::vector<Q> nodes(2 * (size_t)h.n_nodes), spheres(h.n_spheres), moving(3 * (size_t)h.n_moving), rects(2 * (size_t)h.n_rects), cuboids(2 * (size_t)h.n_cuboids),
      translates(h.n_translates), rotates(h.n_rotates), media(h.n_media), mats(h.n_materials), texs(2 * (size_t)h.n_textures), pvec(256 * (size_t)h.n_perlin);
  std::vector<uint4> nsz(h.n_nodes);
  std::vector<uint8_t> pperm(768 * (size_t)h.n_perlin);
  const int pmask = pairMask(h);
  for (size_t i = 0; i < h.n_nodes; ++i) {
    const double* b = &h.node_box[i * 6];
    // both children leaves of one step class (sphere-like, or rect / cuboid): mark the pair (rt_device.cuh kPairBit)
    auto leafClass = [](uint32_t r) { uint32_t k = r >> 28; return (k == K_SPHERE || k == K_MOVING) ? 1 : ((k >= K_RECT_XY && k <= K_CUBOID) ? 2 : 0); };
    const int lcl = leafClass(h.node_left[i]);
    const uint32_t pair = (lcl != 0 && ((pmask >> (lcl - 1)) & 1) && lcl == leafClass(h.node_right[i])) ? kPairBit : 0u;
    nodes[2 * i] = Q{(T)b[0], (T)b[1], (T)b[2], B(h.node_left[i] | pair)};
    nodes[2 * i + 1] = Q{(T)b[3], (T)b[4], (T)b[5], B(h.node_right[i])};
    nsz[i] = make_uint4(h.node_size[i], h.node_lsize[i], h.node_rsize[i], 0);
  }
  for (size_t i = 0; i < h.n_spheres; ++i) { const double* s = &h.sphere_cr[i * 4]; spheres[i] = Q{(T)s[0], (T)s[1], (T)s[2], (T)s[3]}; }
  for (size_t i = 0; i < h.n_moving; ++i) {
    const double* m = &h.moving[i * 9];
    moving[3 * i] = Q{(T)m[0], (T)m[1], (T)m[2], (T)m[8]};
    moving[3 * i + 1] = Q{(T)m[3], (T)m[4], (T)m[5], (T)m[6]};
    moving[3 * i + 2] = Q{(T)m[7], B(h.moving_mat[i]), (T)0, (T)0};
  }
  for (size_t i = 0; i < h.n_rects; ++i) {
    const double* r = &h.rect[i * 5];
    rects[2 * i] = Q{(T)r[0], (T)r[1], (T)r[2], (T)r[3]};
    rects[2 * i + 1] = Q{(T)r[4], B(h.rect_mat[i]), (T)0, (T)0};
  }
  for (size_t i = 0; i < h.n_cuboids; ++i) {
    const double* c = &h.cuboid[i * 6];
    cuboids[2 * i] = Q{(T)c[0], (T)c[1], (T)c[2], B(h.cuboid_mat[i])};
    cuboids[2 * i + 1] = Q{(T)c[3], (T)c[4], (T)c[5], (T)0};
  }
  for (size_t i = 0; i < h.n_translates; ++i) { const double* t = &h.translate_off[i * 3]; translates[i] = Q{(T)t[0], (T)t[1], (T)t[2], B(h.translate_child[i])}; }
  for (size_t i = 0; i < h.n_rotates; ++i) rotates[i] = Q{(T)h.rotate_sc[i * 2], (T)h.rotate_sc[i * 2 + 1], B(h.rotate_axis[i]), B(h.rotate_child[i])};
  for (size_t i = 0; i < h.n_media; ++i) media[i] = Q{(T)h.medium_nid[i], B(h.medium_mat[i]), B(h.medium_boundary[i]), (T)0};
  for (size_t i = 0; i < h.n_materials; ++i) mats[i] = Q{B(h.mat_kind[i]), B(h.mat_tex[i]), (T)h.mat_param[i], (T)0};
  for (size_t i = 0; i < h.n_textures; ++i) {
    texs[2 * i] = Q{(T)h.tex_rgb[i * 3], (T)h.tex_rgb[i * 3 + 1], (T)h.tex_rgb[i * 3 + 2], (T)h.tex_scale[i]};
    texs[2 * i + 1] = Q{B(h.tex_kind[i]), B(h.tex_a[i]), B(h.tex_b[i]), (T)0};
  }
  for (size_t i = 0; i < (size_t)h.n_perlin * 256; ++i) pvec[i] = Q{(T)h.perlin_ranvec[i * 3], (T)h.perlin_ranvec[i * 3 + 1], (T)h.perlin_ranvec[i * 3 + 2], (T)0};
  for (size_t i = 0; i < (size_t)h.n_perlin * 768; ++i) pperm[i] = (uint8_t)(h.perlin_perm[i] & 255);

  DScene<T>& sc = db.sc;
  int rc;
  if ((rc = upload(db, nodes, &sc.nodes))) return rc;
  if ((rc = upload(db, nsz, &sc.node_sz))) return rc;
  if ((rc = upload(db, spheres, &sc.spheres))) return rc;
  if ((rc = upload(db, h.sphere_mat, &sc.sphere_mat))) return rc;
  if ((rc = upload(db, moving, &sc.moving))) return rc;
  if ((rc = upload(db, rects, &sc.rects))) return rc;
  if ((rc = upload(db, cuboids, &sc.cuboids))) return rc;
  if ((rc = upload(db, translates, &sc.translates))) return rc;
  if ((rc = upload(db, rotates, &sc.rotates))) return rc;
  if ((rc = upload(db, media, &sc.media))) return rc;
  if ((rc = upload(db, mats, &sc.mats))) return rc;
  if ((rc = upload(db, texs, &sc.texs))) return rc;
  if ((rc = upload(db, pvec, &sc.perlin_vec))) return rc;
  if ((rc = upload(db, pperm, &sc.perlin_perm))) return rc;
  std::vector<DImage> imgs(h.n_images);
  for (size_t i = 0; i < h.n_images; ++i) {
    size_t npx = (size_t)h.image_w[i] * h.image_h[i];
    std::vector<uchar4> px(npx);
    for (size_t p = 0; p < npx; ++p) px[p] = make_uchar4(h.images[i][p * 3], h.images[i][p * 3 + 1], h.images[i][p * 3 + 2], 255);
    const uchar4* dp = nullptr;
    if ((rc = upload(db, px, &dp))) return rc;
    imgs[i] = DImage{dp, h.image_w[i], h.image_h[i], 0};
  }
  if ((rc = upload(db, imgs, &sc.images))) return rc;
  sc.world_root = h.world_root; sc.lights_root = h.lights_root;
  sc.bigCoords = 0; sc.pad_ = 0;
  for (double v : h.node_box) if (!(std::fabs(v) <= 1099511627776.0)) sc.bigCoords = 1;  // beyond 2^40 or non-finite (rt_device.cuh kDivHi)
  for (int i = 0; i < 3; ++i) sc.bg[i] = (T)h.background[i];
  for (int i = 0; i < 24; ++i) sc.cam[i] = (T)h.camera[i];
  db.ready = true;
  return RT_OK;
}

// the entry stays alive while a call holds the pointer, even if another thread destroys the handle meanwhile
std::shared_ptr<SceneEntry> lookup(rt_scene_t id) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_scenes.find(id);
  return it == g_scenes.end() ? nullptr : it->second;
}

// CUDA events that do not outlive an early error return
struct Events {
  cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
  int n = 0;
  cudaError_t create(int count) { for (; n < count; ++n) { cudaError_t e = cudaEventCreate(&ev[n]); if (e != cudaSuccess) return e; } return cudaSuccess; }
  ~Events() { for (int i = 0; i < n; ++i) cudaEventDestroy(ev[i]); }
};

int deviceScene(SceneEntry* e, uint32_t precision, const void** out) {
  int rc;
  if (precision == RT_PREC_F64) { if ((rc = buildDevice<double>(e->hs, e->f64))) return rc; *out = &e->f64.sc; }
  else if (precision == RT_PREC_F32 || precision == RT_PREC_MIXED) { if ((rc = buildDevice<float>(e->hs, e->f32))) return rc; *out = &e->f32.sc; }
  else return fail(RT_ERR_ARG, "bad precision");
  return RT_OK;
}

int normaliseParams(const RtRenderParams* p, RtRenderParams& q) {
  if (!p) return fail(RT_ERR_ARG, "params is null");
  q = *p;
  if (q.width == 0 || q.height == 0) return fail(RT_ERR_ARG, "width/height must be > 0");
  if ((uint64_t)q.width * q.height > 0x0fffffffu) return fail(RT_ERR_ARG, "image too large");
  if (q.row0 == 0 && q.row1 == 0) q.row1 = q.height;
  if (q.col0 == 0 && q.col1 == 0) q.col1 = q.width;
  if (q.row1 > q.height || q.col1 > q.width || q.row0 >= q.row1 || q.col0 >= q.col1) return fail(RT_ERR_ARG, "bad region");
  if (q.precision > RT_PREC_F64) return fail(RT_ERR_ARG, "bad precision");
  return RT_OK;
}

// scheduler knobs of renderKernel; environment overrides are for tuning experiments only
uint32_t tuneParam(const char* name, uint32_t dflt) {
  const char* v = std::getenv(name);
  return v ? (uint32_t)std::strtoul(v, nullptr, 10) : dflt;
}

// launches the render kernel into d_out (accumulating); counters optional
int renderLaunch(SceneEntry* e, const RtRenderParams& q, double* d_out, cudaStream_t st, bool counted, int* gridOut) {
  const void* dsc = nullptr;
  int rc = deviceScene(e, q.precision, &dsc);
  if (rc) return rc;
  if (!e->d_counter) { CU(cudaMalloc(&e->d_counter, 33 * sizeof(unsigned long long))); }
  CU(cudaMemsetAsync(e->d_counter, 0, 33 * sizeof(unsigned long long), st));
  if (q.sample_count == 0) return RT_OK;
  RenderArgs a{};
  a.W = q.width; a.H = q.height; a.s0 = q.sample_offset; a.ns = q.sample_count; a.maxDepth = q.max_depth;
  a.row0 = q.row0; a.row1 = q.row1; a.col0 = q.col0; a.col1 = q.col1;
  a.seed = q.seed;
  a.total_paths = (unsigned long long)(q.row1 - q.row0) * (q.col1 - q.col0) * q.sample_count;
  // chunk: ~1024 paths, halved until there are at least 8 chunks per warp slot for balance, down to 4 paths: launches with few, very
  // long paths (the cfg5 sample: seconds per path) are tail-bound, and a warp that refills 4 lanes at a time keeps its lanes busier than
  // one that waits for a 32-path chunk to drain (tools/chunk_probe.py: 90 k spheres, 480x270x1: 1424 -> 1242 ms; results are identical)
  unsigned long long warps = (unsigned long long)e->numSMs * 16ull;
  unsigned long long chunk = 1024;
  const unsigned long long chunkMin = std::max(1u, tuneParam("RTB_CHUNK_MIN", 4));
  while (chunk > chunkMin && a.total_paths / chunk < warps * 8ull) chunk /= 2;
  a.chunk = (uint32_t)chunk;
  // scheduler thresholds (rt_render_kernel.inl): single rects are cheap steps and worth running for 3 lanes, cuboid-heavy scenes batch
  // 6; media make the rare class frequent (3 rare steps per medium visit) and worth running for 4 lanes.  Measured on the
  // reference scenes (tools/knob_sweep.py); results never depend on these.
  const HostScene& hs = e->hs;
  a.shadeMin = tuneParam("RTB_SHADE_MIN", 16);
  a.rectMin = std::max(1u, tuneParam("RTB_RECT_MIN", hs.n_cuboids > hs.n_rects ? 6u : 3u));
  a.rareMin = std::max(1u, tuneParam("RTB_RARE_MIN", hs.n_media > 0 ? 4u : 8u));
  a.keepNode = std::max(1u, tuneParam("RTB_KEEP_NODE", 12)); a.keepLeaf = std::max(1u, tuneParam("RTB_KEEP_LEAF", 8));
  a.lightCap = hs.lightCap;
  a.counter = e->d_counter; a.ctr = counted ? e->d_counter + 1 : nullptr; a.out = d_out;
  cudaError_t ce = (q.flags & RT_FLAG_FAST_MATH) ? launchRender<true>((int)q.precision, counted ? 1 : 0, dsc, a, e->hs.stackDepth, e->numSMs, st, gridOut)
                                                 : launchRender<false>((int)q.precision, counted ? 1 : 0, dsc, a, e->hs.stackDepth, e->numSMs, st, gridOut);
  if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("render launch: ") + cudaGetErrorString(ce));
  return RT_OK;
}

int renderHost(rt_scene_t scene, const RtRenderParams* params, double* out, RtStats* stats, bool counted) {
  const std::shared_ptr<SceneEntry> ep = lookup(scene);
  if (!ep) return fail(RT_ERR_HANDLE, "unknown scene handle");
  SceneEntry* e = ep.get();
  std::lock_guard<std::mutex> lk(e->mu);
  if (!out) return fail(RT_ERR_ARG, "out_rgb_sum is null");
  RtRenderParams q;
  int rc = normaliseParams(params, q);
  if (rc) return rc;
  CU(cudaSetDevice(e->device));
  const size_t elems = (size_t)q.width * q.height * 3;
  if (e->fbElems < elems) {
    if (e->d_fb) cudaFree(e->d_fb);
    e->d_fb = nullptr; e->fbElems = 0;
    CU(cudaMalloc(&e->d_fb, elems * sizeof(double)));
    e->fbElems = elems;
  }
  if (e->hfbElems < elems) {
    if (e->h_fb) cudaFreeHost(e->h_fb);
    e->h_fb = nullptr; e->hfbElems = 0;
    CU(cudaHostAlloc(&e->h_fb, elems * sizeof(double), cudaHostAllocDefault));
    e->hfbElems = elems;
  }
  Events evs;
  CU(evs.create(3));
  const cudaEvent_t e0 = evs.ev[0], e1 = evs.ev[1], e2 = evs.ev[2];
  CU(cudaMemsetAsync(e->d_fb, 0, elems * sizeof(double), 0));
  CU(cudaEventRecord(e0, 0));
  int grid = 0;
  rc = renderLaunch(e, q, e->d_fb, 0, counted, &grid);
  if (rc) return rc;
  CU(cudaEventRecord(e1, 0));
  // copy back only the rows of the region
  const size_t rowElems = (size_t)q.width * 3;
  CU(cudaMemcpyAsync(e->h_fb + q.row0 * rowElems, e->d_fb + q.row0 * rowElems, (size_t)(q.row1 - q.row0) * rowElems * sizeof(double), cudaMemcpyDeviceToHost, 0));
  CU(cudaEventRecord(e2, 0));
  CU(cudaEventSynchronize(e2));
  for (uint32_t r = q.row0; r < q.row1; ++r)
    std::memcpy(out + r * rowElems + (size_t)q.col0 * 3, e->h_fb + r * rowElems + (size_t)q.col0 * 3, (size_t)(q.col1 - q.col0) * 3 * sizeof(double));
  if (stats) {
    std::memset(stats, 0, sizeof(*stats));
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1); stats->kernel_ms = ms;
    cudaEventElapsedTime(&ms, e1, e2); stats->d2h_ms = ms;
    stats->samples = (uint64_t)(q.row1 - q.row0) * (q.col1 - q.col0) * q.sample_count;
    stats->launches = q.sample_count ? 1 : 0;
    if (counted) {
      unsigned long long c[33];
      CU(cudaMemcpy(c, e->d_counter, sizeof(c), cudaMemcpyDeviceToHost));
      for (int i = 0; i < 16; ++i) { stats->counters[i] = c[i + 1]; stats->sched[i] = c[i + 17]; }
    }
  }
  return RT_OK;
}

}  // namespace

namespace rtb { void setLastError(const std::string& s) { g_err = s; } }  // for the other translation units of the library

extern "C" {

int rt_abi_version(void) { return RT_ABI_VERSION; }
const char* rt_last_error(void) { return g_err.c_str(); }
int rt_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}
int rt_get_limits(RtLimits* out) {
  if (!out) return fail(RT_ERR_ARG, "out is null");
  out->max_stack = kMaxStack; out->max_frames = kMaxFrames; out->max_medium_depth = 1; out->max_light_depth = kMaxLightDepth;
  return RT_OK;
}

int rt_scene_create(const RtSceneDesc* d, int device, rt_scene_t* out) {
  if (!d || !out) return fail(RT_ERR_ARG, "desc/out is null");
  if (d->abi_version != RT_ABI_VERSION) return fail(RT_ERR_ARG, "abi_version mismatch");
  auto need = [](uint32_t n, const void* p) { return n == 0 || p != nullptr; };
  if (!need(d->n_nodes, d->node_box) || !need(d->n_nodes, d->node_left) || !need(d->n_nodes, d->node_right) || !need(d->n_nodes, d->node_size) ||
      !need(d->n_spheres, d->sphere_cr) || !need(d->n_spheres, d->sphere_mat) || !need(d->n_moving, d->moving) || !need(d->n_moving, d->moving_mat) ||
      !need(d->n_rects, d->rect) || !need(d->n_rects, d->rect_mat) || !need(d->n_cuboids, d->cuboid) || !need(d->n_cuboids, d->cuboid_mat) ||
      !need(d->n_translates, d->translate_off) || !need(d->n_translates, d->translate_child) || !need(d->n_rotates, d->rotate_sc) ||
      !need(d->n_rotates, d->rotate_axis) || !need(d->n_rotates, d->rotate_child) || !need(d->n_media, d->medium_neg_inv_density) ||
      !need(d->n_media, d->medium_mat) || !need(d->n_media, d->medium_boundary) || !need(d->n_materials, d->mat_kind) || !need(d->n_materials, d->mat_tex) ||
      !need(d->n_materials, d->mat_param) || !need(d->n_textures, d->tex_kind) || !need(d->n_textures, d->tex_rgb) || !need(d->n_textures, d->tex_a) ||
      !need(d->n_textures, d->tex_b) || !need(d->n_textures, d->tex_scale) || !need(d->n_perlin, d->perlin_ranvec) || !need(d->n_perlin, d->perlin_perm) ||
      !need(d->n_images, d->image_rgb) || !need(d->n_images, d->image_w) || !need(d->n_images, d->image_h))
    return fail(RT_ERR_ARG, "a non-empty array is null");
  const uint32_t lim = kIdxMask;  // 27-bit indices: bit 27 of a ref is the pair mark
  if (d->n_nodes > lim || d->n_spheres > lim || d->n_moving > lim || d->n_rects > lim || d->n_cuboids > lim) return fail(RT_ERR_ARG, "too many objects");

  auto e = std::make_shared<SceneEntry>();
  HostScene& h = e->hs;
  h.n_nodes = d->n_nodes; h.n_spheres = d->n_spheres; h.n_moving = d->n_moving; h.n_rects = d->n_rects; h.n_cuboids = d->n_cuboids;
  h.n_translates = d->n_translates; h.n_rotates = d->n_rotates; h.n_media = d->n_media; h.n_materials = d->n_materials; h.n_textures = d->n_textures;
  h.n_perlin = d->n_perlin; h.n_images = d->n_images;
  copyIn(h.node_box, d->node_box, 6 * (size_t)h.n_nodes); copyIn(h.node_left, d->node_left, h.n_nodes); copyIn(h.node_right, d->node_right, h.n_nodes);
  copyIn(h.node_size, d->node_size, h.n_nodes);
  copyIn(h.sphere_cr, d->sphere_cr, 4 * (size_t)h.n_spheres); copyIn(h.sphere_mat, d->sphere_mat, h.n_spheres);
  copyIn(h.moving, d->moving, 9 * (size_t)h.n_moving); copyIn(h.moving_mat, d->moving_mat, h.n_moving);
  copyIn(h.rect, d->rect, 5 * (size_t)h.n_rects); copyIn(h.rect_mat, d->rect_mat, h.n_rects);
  copyIn(h.cuboid, d->cuboid, 6 * (size_t)h.n_cuboids); copyIn(h.cuboid_mat, d->cuboid_mat, h.n_cuboids);
  copyIn(h.translate_off, d->translate_off, 3 * (size_t)h.n_translates); copyIn(h.translate_child, d->translate_child, h.n_translates);
  copyIn(h.rotate_sc, d->rotate_sc, 2 * (size_t)h.n_rotates); copyIn(h.rotate_axis, d->rotate_axis, h.n_rotates); copyIn(h.rotate_child, d->rotate_child, h.n_rotates);
  copyIn(h.medium_nid, d->medium_neg_inv_density, h.n_media); copyIn(h.medium_mat, d->medium_mat, h.n_media); copyIn(h.medium_boundary, d->medium_boundary, h.n_media);
  copyIn(h.mat_kind, d->mat_kind, h.n_materials); copyIn(h.mat_tex, d->mat_tex, h.n_materials); copyIn(h.mat_param, d->mat_param, h.n_materials);
  copyIn(h.tex_kind, d->tex_kind, h.n_textures); copyIn(h.tex_rgb, d->tex_rgb, 3 * (size_t)h.n_textures); copyIn(h.tex_a, d->tex_a, h.n_textures);
  copyIn(h.tex_b, d->tex_b, h.n_textures); copyIn(h.tex_scale, d->tex_scale, h.n_textures);
  copyIn(h.perlin_ranvec, d->perlin_ranvec, 768 * (size_t)h.n_perlin); copyIn(h.perlin_perm, d->perlin_perm, 768 * (size_t)h.n_perlin);
  copyIn(h.image_w, d->image_w, h.n_images); copyIn(h.image_h, d->image_h, h.n_images);
  h.images.resize(h.n_images);
  for (uint32_t i = 0; i < h.n_images; ++i) {
    if (!d->image_rgb[i] || h.image_w[i] == 0 || h.image_h[i] == 0) return fail(RT_ERR_ARG, "bad image");
    h.images[i].assign(d->image_rgb[i], d->image_rgb[i] + (size_t)h.image_w[i] * h.image_h[i] * 3);
  }
  h.world_root = d->world_root; h.lights_root = d->lights_root;
  std::memcpy(h.background, d->background, sizeof(h.background)); std::memcpy(h.camera, d->camera, sizeof(h.camera));

  // ---- validation
  auto matOk = [&](uint32_t m) { return m < h.n_materials; };
  for (uint32_t i = 0; i < h.n_nodes; ++i) if (!refOk(h, h.node_left[i]) || !refOk(h, h.node_right[i])) return fail(RT_ERR_ARG, "BVH child ref out of range");
  for (uint32_t i = 0; i < h.n_translates; ++i) if (!refOk(h, h.translate_child[i])) return fail(RT_ERR_ARG, "translate child out of range");
  for (uint32_t i = 0; i < h.n_rotates; ++i) if (!refOk(h, h.rotate_child[i]) || h.rotate_axis[i] > 2) return fail(RT_ERR_ARG, "rotate child/axis out of range");
  for (uint32_t i = 0; i < h.n_media; ++i) if (!refOk(h, h.medium_boundary[i]) || !matOk(h.medium_mat[i])) return fail(RT_ERR_ARG, "medium boundary/material out of range");
  for (uint32_t i = 0; i < h.n_spheres; ++i) if (!matOk(h.sphere_mat[i])) return fail(RT_ERR_ARG, "sphere material out of range");
  for (uint32_t i = 0; i < h.n_moving; ++i) if (!matOk(h.moving_mat[i])) return fail(RT_ERR_ARG, "moving sphere material out of range");
  for (uint32_t i = 0; i < h.n_rects; ++i) if (!matOk(h.rect_mat[i])) return fail(RT_ERR_ARG, "rect material out of range");
  for (uint32_t i = 0; i < h.n_cuboids; ++i) if (!matOk(h.cuboid_mat[i])) return fail(RT_ERR_ARG, "cuboid material out of range");
  for (uint32_t i = 0; i < h.n_materials; ++i) {
    if (h.mat_kind[i] > M_ISOTROPIC) return fail(RT_ERR_ARG, "bad material kind");
    if (h.mat_kind[i] != M_DIELECTRIC && h.mat_tex[i] >= h.n_textures) return fail(RT_ERR_ARG, "material texture out of range");
  }
  for (uint32_t i = 0; i < h.n_textures; ++i) {
    uint32_t k = h.tex_kind[i];
    if (k > TX_IMAGE) return fail(RT_ERR_ARG, "bad texture kind");
    if (k == TX_CHECKER && (h.tex_a[i] >= h.n_textures || h.tex_b[i] >= h.n_textures)) return fail(RT_ERR_ARG, "checker child out of range");
    if (k == TX_PERLIN && h.tex_a[i] >= h.n_perlin) return fail(RT_ERR_ARG, "perlin table out of range");
    if (k == TX_IMAGE && h.tex_a[i] != RT_NONE && h.tex_a[i] >= h.n_images) return fail(RT_ERR_ARG, "image out of range");
  }
  if (!refOk(h, h.world_root) || !refOk(h, h.lights_root)) return fail(RT_ERR_ARG, "root ref out of range");
  Analysis an(h);
  an.go(h.world_root); an.go(h.lights_root);
  if (!an.cycle) for (uint32_t i = 0; i < h.n_nodes; ++i) { an.go(h.node_left[i]); an.go(h.node_right[i]); }
  if (an.cycle) return fail(RT_ERR_ARG, "hittable graph has a cycle");
  int stackNeed = 1, frames = 0, media = 0;
  for (auto& kv : an.stackNeed) stackNeed = std::max(stackNeed, kv.second);
  for (auto& kv : an.frames) frames = std::max(frames, kv.second);
  for (auto& kv : an.media) media = std::max(media, kv.second);
  if (stackNeed > kMaxStack) return fail(RT_ERR_UNSUPPORTED, "scene needs a traversal stack of " + std::to_string(stackNeed) + " entries (limit " + std::to_string(kMaxStack) + ")");
  if (frames > kMaxFrames) return fail(RT_ERR_UNSUPPORTED, "more than " + std::to_string(kMaxFrames) + " nested Translate/Rotate frames");
  if (media > 1) return fail(RT_ERR_UNSUPPORTED, "ConstantMedium nested inside a ConstantMedium boundary");
  if (lightDepth(h, h.lights_root) + 1 > kMaxLightDepth) return fail(RT_ERR_UNSUPPORTED, "lights BVH too deep");
  h.stackDepth = stackNeed;
  h.lightCap = (h.lights_root >> 28) == K_BVH ? (uint32_t)lightDepth(h, h.lights_root) : 0u;  // DFS over L levels holds at most L entries
  h.node_lsize.resize(h.n_nodes); h.node_rsize.resize(h.n_nodes);
  for (uint32_t i = 0; i < h.n_nodes; ++i) { h.node_lsize[i] = (uint32_t)an.size[h.node_left[i]]; h.node_rsize[i] = (uint32_t)an.size[h.node_right[i]]; }

  // ---- device
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(RT_ERR_CUDA, "no CUDA device available (this library has no CPU path)"); }
  if (device < 0) { CU(cudaGetDevice(&device)); }
  if (device >= ndev) return fail(RT_ERR_ARG, "device ordinal out of range");
  CU(cudaSetDevice(device));
  e->device = device;
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  e->numSMs = prop.multiProcessorCount;
  int rc = buildDevice<float>(h, e->f32);  // the production layout is uploaded eagerly; fp64 on first use
  if (rc) return rc;
  std::lock_guard<std::mutex> lk(g_mu);
  *out = g_nextId++;
  g_scenes[*out] = std::move(e);
  return RT_OK;
}

int rt_scene_destroy(rt_scene_t scene) {
  std::shared_ptr<SceneEntry> e;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_scenes.find(scene);
    if (it == g_scenes.end()) return fail(RT_ERR_HANDLE, "unknown scene handle");
    e = std::move(it->second);
    g_scenes.erase(it);
  }
  {
    std::lock_guard<std::mutex> lk(e->mu);  // a call in flight on another thread finishes first
    cudaSetDevice(e->device);
    cudaDeviceSynchronize();                // and so does asynchronous work of rt_render_device
  }
  e.reset();  // the buffers go with the last holder
  return RT_OK;
}

int rt_render(rt_scene_t scene, const RtRenderParams* params, double* out, RtStats* stats) { return renderHost(scene, params, out, stats, false); }
int rt_render_counted(rt_scene_t scene, const RtRenderParams* params, double* out, RtStats* stats) { return renderHost(scene, params, out, stats, true); }

int rt_render_device(rt_scene_t scene, const RtRenderParams* params, double* d_out, void* cuda_stream, RtStats* stats) {
  const std::shared_ptr<SceneEntry> ep = lookup(scene);
  if (!ep) return fail(RT_ERR_HANDLE, "unknown scene handle");
  SceneEntry* e = ep.get();
  std::lock_guard<std::mutex> lk(e->mu);
  if (!d_out) return fail(RT_ERR_ARG, "d_out_rgb_sum is null");
  RtRenderParams q;
  int rc = normaliseParams(params, q);
  if (rc) return rc;
  CU(cudaSetDevice(e->device));
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  Events evs;
  if (stats) { CU(evs.create(2)); CU(cudaEventRecord(evs.ev[0], st)); }
  const cudaEvent_t e0 = evs.ev[0], e1 = evs.ev[1];
  rc = renderLaunch(e, q, d_out, st, false, nullptr);
  if (rc) return rc;
  if (stats) {
    CU(cudaEventRecord(e1, st));
    CU(cudaEventSynchronize(e1));
    std::memset(stats, 0, sizeof(*stats));
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    stats->kernel_ms = ms; stats->samples = (uint64_t)(q.row1 - q.row0) * (q.col1 - q.col0) * q.sample_count; stats->launches = q.sample_count ? 1 : 0;
  }
  return RT_OK;
}

int rt_trace(rt_scene_t scene, uint32_t precision, uint32_t flags, uint32_t n, const double* rays, double t_min, double t_max, uint64_t seed,
             uint32_t log_stride, int32_t* hit, double* rec, uint32_t* prim, uint32_t* mat, uint8_t* front, uint64_t* nvisit, uint64_t* vhash, uint32_t* vlog) {
  const std::shared_ptr<SceneEntry> ep = lookup(scene);
  if (!ep) return fail(RT_ERR_HANDLE, "unknown scene handle");
  SceneEntry* e = ep.get();
  std::lock_guard<std::mutex> lk(e->mu);
  if (!rays && n) return fail(RT_ERR_ARG, "rays is null");
  if (n == 0) return RT_OK;
  CU(cudaSetDevice(e->device));
  const void* dsc = nullptr;
  int rc = deviceScene(e, precision, &dsc);
  if (rc) return rc;
  const uint32_t stride = log_stride ? log_stride : 1;
  struct Buf { void* p = nullptr; ~Buf() { if (p) cudaFree(p); } };
  Buf bRays, bHit, bRec, bPrim, bMat, bFront, bNv, bVh, bLog;
  CU(cudaMalloc(&bRays.p, (size_t)n * 7 * sizeof(double)));
  CU(cudaMalloc(&bHit.p, (size_t)n * 4)); CU(cudaMalloc(&bRec.p, (size_t)n * 9 * sizeof(double))); CU(cudaMalloc(&bPrim.p, (size_t)n * 4));
  CU(cudaMalloc(&bMat.p, (size_t)n * 4)); CU(cudaMalloc(&bFront.p, (size_t)n)); CU(cudaMalloc(&bNv.p, (size_t)n * 8)); CU(cudaMalloc(&bVh.p, (size_t)n * 8));
  CU(cudaMalloc(&bLog.p, (size_t)n * stride * 4));
  CU(cudaMemcpy(bRays.p, rays, (size_t)n * 7 * sizeof(double), cudaMemcpyHostToDevice));
  TraceArgs a{};
  a.n = n; a.logStride = log_stride; a.tmin = t_min; a.tmax = t_max; a.seed = seed; a.rays = (const double*)bRays.p;
  a.hit = (int32_t*)bHit.p; a.rec = (double*)bRec.p; a.prim = (uint32_t*)bPrim.p; a.mat = (uint32_t*)bMat.p; a.front = (uint8_t*)bFront.p;
  a.nvisit = (unsigned long long*)bNv.p; a.vhash = (unsigned long long*)bVh.p; a.vlog = (uint32_t*)bLog.p;
  cudaError_t ce = (flags & RT_FLAG_FAST_MATH) ? launchTrace<true>((int)precision, dsc, a, e->hs.stackDepth, 0) : launchTrace<false>((int)precision, dsc, a, e->hs.stackDepth, 0);
  if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("trace launch: ") + cudaGetErrorString(ce));
  CU(cudaDeviceSynchronize());
  if (hit) CU(cudaMemcpy(hit, bHit.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
  if (rec) CU(cudaMemcpy(rec, bRec.p, (size_t)n * 9 * sizeof(double), cudaMemcpyDeviceToHost));
  if (prim) CU(cudaMemcpy(prim, bPrim.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
  if (mat) CU(cudaMemcpy(mat, bMat.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
  if (front) CU(cudaMemcpy(front, bFront.p, (size_t)n, cudaMemcpyDeviceToHost));
  if (nvisit) CU(cudaMemcpy(nvisit, bNv.p, (size_t)n * 8, cudaMemcpyDeviceToHost));
  if (vhash) CU(cudaMemcpy(vhash, bVh.p, (size_t)n * 8, cudaMemcpyDeviceToHost));
  if (vlog && log_stride) CU(cudaMemcpy(vlog, bLog.p, (size_t)n * log_stride * 4, cudaMemcpyDeviceToHost));
  return RT_OK;
}

int rt_texture_value(rt_scene_t scene, uint32_t tex, uint32_t precision, uint32_t n, const double* q, double* out) {
  const std::shared_ptr<SceneEntry> ep = lookup(scene);
  if (!ep) return fail(RT_ERR_HANDLE, "unknown scene handle");
  SceneEntry* e = ep.get();
  std::lock_guard<std::mutex> lk(e->mu);
  if (tex >= e->hs.n_textures) return fail(RT_ERR_ARG, "texture index out of range");
  if (n == 0) return RT_OK;
  if (!q || !out) return fail(RT_ERR_ARG, "q/out is null");
  CU(cudaSetDevice(e->device));
  const void* dsc = nullptr;
  int rc = deviceScene(e, precision, &dsc);
  if (rc) return rc;
  double *dq = nullptr, *dout = nullptr;
  CU(cudaMalloc(&dq, (size_t)n * 5 * sizeof(double))); CU(cudaMalloc(&dout, (size_t)n * 3 * sizeof(double)));
  CU(cudaMemcpy(dq, q, (size_t)n * 5 * sizeof(double), cudaMemcpyHostToDevice));
  cudaError_t ce = launchTexture((int)precision, dsc, tex, n, dq, dout, 0);
  if (ce == cudaSuccess) ce = cudaMemcpy(out, dout, (size_t)n * 3 * sizeof(double), cudaMemcpyDeviceToHost);
  cudaFree(dq); cudaFree(dout);
  if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("texture probe: ") + cudaGetErrorString(ce));
  return RT_OK;
}

int rt_tonemap(const double* rgb_sum, uint32_t n, uint32_t total_samples, uint8_t* out_rgb8) {
  if (n == 0) return RT_OK;
  if (!rgb_sum || !out_rgb8) return fail(RT_ERR_ARG, "null buffer");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(RT_ERR_CUDA, "no CUDA device available (this library has no CPU path)"); }
  double* din = nullptr; uint8_t* dout = nullptr;
  CU(cudaMalloc(&din, (size_t)n * sizeof(double))); CU(cudaMalloc(&dout, n));
  CU(cudaMemcpy(din, rgb_sum, (size_t)n * sizeof(double), cudaMemcpyHostToDevice));
  cudaError_t ce = launchTonemap(din, n, (double)total_samples, dout, 0);
  if (ce == cudaSuccess) ce = cudaMemcpy(out_rgb8, dout, n, cudaMemcpyDeviceToHost);
  cudaFree(din); cudaFree(dout);
  if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("tonemap: ") + cudaGetErrorString(ce));
  return RT_OK;
}

int rt_selftest_division(uint64_t n, uint64_t seed, uint64_t* checked, uint64_t* mismatches) {
  if (!checked || !mismatches) return fail(RT_ERR_ARG, "null output");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(RT_ERR_CUDA, "no CUDA device available (this library has no CPU path)"); }
  unsigned long long* d = nullptr;
  CU(cudaMalloc(&d, 2 * sizeof(unsigned long long)));
  CU(cudaMemset(d, 0, 2 * sizeof(unsigned long long)));
  cudaError_t ce = launchDivSelftest(n, seed, d, 0);
  unsigned long long h[2] = {0, 0};
  if (ce == cudaSuccess) ce = cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("division selftest: ") + cudaGetErrorString(ce));
  *checked = h[0]; *mismatches = h[1];
  return RT_OK;
}

void rt_shutdown(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  g_scenes.clear();
}

}  // extern "C"
