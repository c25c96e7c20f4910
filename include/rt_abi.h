/* rt_abi.h — C ABI of librt_b200.so: the B200 path-tracing core behind the reference's render entry point.
 *
 * What it replaces (paths relative to the reference checkout):
 *   runRender :: RenderStaticEnv -> [RandGen] -> [Vector RGB]        src/Lib.hs:1491-1523
 *   and everything runRender reaches per sample: uniformRandomUVs (1352-1371), renderPos/sampleColor (1335-1350),
 *   getRay (1253-1267), rayColor (1298-1333), hit (970-1109), boxRayIntersect (798-814), scatter/scatteringPdf/
 *   emitted (819-903), pdfGenerate/pdfValue (362-382), htblPdfValue/htblRandom (662-724), textureValue/noise/turb
 *   (441-513), the samplers (1119-1228).
 * The reference has no FFI today (package.yaml:25-35 has no c-sources / extra-libraries); INTEGRATION.md shows the
 * `foreign import ccall` stubs and the flattener a maintainer adds inside Lib.  Scene construction (Scenes.hs, the
 * Lib smart constructors, makeBVH, makePerlin, newCamera) stays on the host and arrives here as the flat
 * struct-of-arrays RtSceneDesc below.  All reals cross the boundary as the host's Double; the library rounds
 * them to fp32 (round-to-nearest-even) when it uploads for the fp32 kernels.
 *
 * Conventions: every function returns 0 on success and a negative RT_ERR_* code on failure; rt_last_error()
 * returns a thread-local message.  Host pointers are borrowed for the duration of the call only.  The library
 * never falls back to a CPU implementation: without a usable CUDA device every compute entry point fails with
 * RT_ERR_CUDA.
 */
#ifndef RT_ABI_H
#define RT_ABI_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 2

/* error codes */
#define RT_OK 0
#define RT_ERR_ARG (-1)        /* null pointer, bad size, bad enum, index out of range in the scene description */
#define RT_ERR_UNSUPPORTED (-2) /* nesting deeper than the compiled traversal budget (see rt_limits) */
#define RT_ERR_CUDA (-3)       /* no device, allocation failure, launch failure (message has the CUDA error string) */
#define RT_ERR_HANDLE (-4)     /* unknown / destroyed scene handle */

/* Hittable reference = (kind << 28) | index-within-kind.  Kinds follow the constructors of
 * `data Hittable` (Lib.hs:521-585) with `Rect` split by its Rectangle constructor (Lib.hs:607-647); the three rect
 * kinds index one shared rect array.  Every array may hold at most 2^27 entries (rt_scene_create rejects more). */
enum RtHittableKind {
  RT_UNHITTABLE = 0,
  RT_SPHERE = 1,
  RT_MOVING_SPHERE = 2,
  RT_RECT_XY = 3,
  RT_RECT_XZ = 4,
  RT_RECT_YZ = 5,
  RT_CUBOID = 6,
  RT_BVH_NODE = 7,
  RT_TRANSLATE = 8,
  RT_ROTATE = 9,
  RT_CONSTANT_MEDIUM = 10
};
#define RT_REF(kind, index) ((((uint32_t)(kind)) << 28) | ((uint32_t)(index) & 0x0fffffffu))
#define RT_REF_KIND(ref) ((ref) >> 28)
#define RT_REF_INDEX(ref) ((ref) & 0x0fffffffu)
#define RT_NONE 0xffffffffu

/* `data Material` (Lib.hs:339-345) and `data Texture` (Lib.hs:394-419) */
enum RtMaterialKind { RT_LAMBERTIAN = 0, RT_METAL = 1, RT_DIELECTRIC = 2, RT_DIFFUSE_LIGHT = 3, RT_ISOTROPIC = 4 };
enum RtTextureKind { RT_TEX_CONSTANT = 0, RT_TEX_CHECKER = 1, RT_TEX_PERLIN = 2, RT_TEX_IMAGE = 3 };
enum RtAxis { RT_AXIS_X = 0, RT_AXIS_Y = 1, RT_AXIS_Z = 2 }; /* data Axis, Lib.hs:515 */

/* Flattened scene: `type Scene = (Hittable, Hittable, Albedo)` (Lib.hs:84) + `data Camera` (Lib.hs:1230-1251).
 * Arrays may be NULL when their count is 0. */
typedef struct RtSceneDesc {
  uint32_t abi_version; /* RT_ABI_VERSION */

  /* BVHNode left right box size (Lib.hs:552-560) */
  uint32_t n_nodes;
  const double* node_box;     /* [n][6] min.xyz max.xyz */
  const uint32_t* node_left;  /* [n] ref */
  const uint32_t* node_right; /* [n] ref */
  const uint32_t* node_size;  /* [n] bvh_size (number of objects below; used by htblRandom / htblPdfValue weights) */

  /* Sphere center radius material (Lib.hs:522-528) */
  uint32_t n_spheres;
  const double* sphere_cr;    /* [n][4] cx cy cz r */
  const uint32_t* sphere_mat; /* [n] material index */

  /* MovingSphere c0 c1 t0 t1 duration r material (Lib.hs:529-543); t1 is not used by `hit` */
  uint32_t n_moving;
  const double* moving;       /* [n][9] c0.xyz c1.xyz t0 duration r */
  const uint32_t* moving_mat; /* [n] */

  /* Rect (XYRect|XZRect|YZRect) a0 a1 b0 b1 k material (Lib.hs:607-647); the plane is the ref's kind */
  uint32_t n_rects;
  const double* rect;         /* [n][5] a0 a1 b0 b1 k */
  const uint32_t* rect_mat;   /* [n] */

  /* Cuboid min max [six Rects] (Lib.hs:545-551); the rect list is always the one `cuboid` builds (Lib.hs:594-605) */
  uint32_t n_cuboids;
  const double* cuboid;       /* [n][6] min.xyz max.xyz */
  const uint32_t* cuboid_mat; /* [n] */

  /* Translate offset child (Lib.hs:561-565) */
  uint32_t n_translates;
  const double* translate_off;     /* [n][3] */
  const uint32_t* translate_child; /* [n] ref */

  /* Rotate axis sinTheta cosTheta box child (Lib.hs:566-576); the box is only used by the parent BVH */
  uint32_t n_rotates;
  const double* rotate_sc;       /* [n][2] sin cos */
  const uint32_t* rotate_axis;   /* [n] RtAxis */
  const uint32_t* rotate_child;  /* [n] ref */

  /* ConstantMedium negInvDensity material boundary (Lib.hs:577-583) */
  uint32_t n_media;
  const double* medium_neg_inv_density; /* [n] */
  const uint32_t* medium_mat;           /* [n] material index (Isotropic) */
  const uint32_t* medium_boundary;      /* [n] ref */

  uint32_t n_materials;
  const uint32_t* mat_kind;  /* [n] RtMaterialKind */
  const uint32_t* mat_tex;   /* [n] texture index or RT_NONE (Dielectric) */
  const double* mat_param;   /* [n] Fuzz / RefractiveIdx, else 0 */

  uint32_t n_textures;
  const uint32_t* tex_kind;  /* [n] RtTextureKind */
  const double* tex_rgb;     /* [n][3] ConstantColor albedo */
  const uint32_t* tex_a;     /* [n] checker: odd texture; perlin: table index; image: image index or RT_NONE (= ImageTexture Nothing) */
  const uint32_t* tex_b;     /* [n] checker: even texture */
  const double* tex_scale;   /* [n] perlinScale */

  /* Perlin ranvec permX permY permZ (Lib.hs:401-410) */
  uint32_t n_perlin;
  const double* perlin_ranvec; /* [n][256][3] */
  const int32_t* perlin_perm;  /* [n][3][256] X, Y, Z */

  /* ImageTexture pixels (JP.Image PixelRGB8, Lib.hs:384-389), tightly packed RGB8 rows, top row first */
  uint32_t n_images;
  const uint8_t* const* image_rgb; /* [n] -> [h][w][3] */
  const uint32_t* image_w;
  const uint32_t* image_h;

  uint32_t world_root;  /* ref: fst of Scene */
  uint32_t lights_root; /* ref: snd of Scene (RT_REF(RT_UNHITTABLE,0) for `Unhittable`) */
  double background[3]; /* thd of Scene */
  double camera[24];    /* origin llc horiz vert u v w (7 x xyz), lensRadius, t0, t1 (Lib.hs:1230-1251) */
} RtSceneDesc;

/* arithmetic of the device path */
enum RtPrecision {
  RT_PREC_F32 = 0,   /* fp32 everywhere.  Opt-in: with t_min = 1e-4 (Lib.hs:77) an fp32 sphere quadratic re-hits a radius-1000
                        sphere from its own surface, which biases every Book-1/2 scene against the reference's Double */
  RT_PREC_MIXED = 1, /* THE PRODUCTION ARITHMETIC (what runRender and bench.py use): fp32, with the sphere quadratic
                        (Lib.hs:1091-1095) in fp64 — meets the radiance bar against the reference's Double at the speed of F32 on B200 */
  RT_PREC_F64 = 2    /* fp64 everywhere: the reference's Double arithmetic, for validation */
};
#define RT_PREC_DEFAULT RT_PREC_MIXED
#define RT_FLAG_FAST_MATH 1u   /* allow FMA contraction, reciprocal slab test and fast transcendentals (not bit-exact vs the fp32 oracle) */
#define RT_FLAG_STRICT_MATH 2u /* parity instrument: sin / cos / atan2 / asin / log from the shared plain-arithmetic implementations
                                  (ray-tracing_b200/csrc/rt_strict_math.h) instead of libdevice, so that a path is bit-identical to the
                                  test oracle's; same kernels, same step and draw order.  Excludes RT_FLAG_FAST_MATH. */

#define RT_FLAG_COOP 4u        /* force the warp-cooperative scan-stream traversal (one path per warp; ray-tracing_b200/csrc/rt_coop.cuh);
                                  RT_ERR_UNSUPPORTED when the scene is not eligible (world = a BVH of spheres / rects only, light list =
                                  one primitive or Unhittable).  Default: chosen automatically for eligible scenes with a large BVH,
                                  where per-path latency, not throughput, bounds a frame (cfg5).  Same visits, same order, same image. */
#define RT_FLAG_NO_COOP 8u     /* never use it */
#define RT_FLAG_WAVEFRONT 16u  /* force the split traversal / shade kernels over path-state slots in HBM (rt_wavefront.cuh; fp32 / mixed,
                                  literal math; otherwise RT_ERR_UNSUPPORTED).  Default: chosen per scene (shading-heavy scenes).  Same
                                  steps, draws and arithmetic per path as renderKernel: the same image. */
#define RT_FLAG_NO_WAVEFRONT 32u /* never use it */

typedef struct RtRenderParams {
  uint32_t width, height;     /* full image size (imageWidth, imageHeight of mkRenderStaticEnv, Lib.hs:92-108) */
  uint32_t sample_offset;     /* first sample index rendered by this call */
  uint32_t sample_count;      /* samples per pixel rendered by this call (numSamples when offset = 0) */
  uint32_t total_samples;     /* informational: the frame's numSamples (the caller divides the sums by it) */
  uint32_t max_depth;         /* maxDepth */
  uint32_t row0, row1;        /* rows [row0,row1) rendered by this call (row 0 = top of the image); 0,0 = all */
  uint32_t col0, col1;        /* columns [col0,col1); 0,0 = all */
  uint64_t seed;              /* stream seed: sample s of pixel p uses streamGen(seed, p, s) (DESIGN.md "RNG") */
  uint32_t precision;         /* RtPrecision */
  uint32_t flags;             /* RT_FLAG_* */
  int32_t device;             /* CUDA device ordinal, -1 = current */
  uint32_t n_gpus;            /* rt_render only: devices the call shards the frame over.  0 = the library default (1 until rt_init
                                 changes it), k = devices 0..k-1 plus the scene's own.  The scene is replicated onto them on first use,
                                 samples are split into contiguous slices (interleaved rows when sample_count < devices), the partial
                                 framebuffers are summed onto the scene's device over NVLink and copied to the host once. */
} RtRenderParams;

typedef struct RtStats {
  double kernel_ms;        /* device time of the render kernel(s), CUDA events on the launching stream */
  double h2d_ms, d2h_ms;   /* copies inside this call (0 for the *_device entry points) */
  uint64_t samples;        /* camera paths traced by this call */
  uint64_t launches;       /* kernels launched by this call */
  uint64_t counters[24];   /* RT_CTR_* when the call was made with counters enabled, else 0 */
  uint64_t sched[16];      /* warp-scheduler statistics of the instrumented kernel: [0..4] iterations that ran class c (node, sphere,
                              rect, rare, shade), [5..9] lanes served by them, [10..14] lanes waiting per class summed over all
                              iterations, [15] iterations */
  double reduce_ms;        /* multi-GPU rt_render: device time of the partial-framebuffer reduction onto the scene's device */
  double replicate_ms;     /* multi-GPU rt_render: host time spent replicating the scene onto other devices in THIS call (0 once cached) */
  uint32_t gpus;           /* devices that rendered a share of this call (1 for single-device calls) */
  uint32_t reduce_kind;    /* RT_REDUCE_* actually used */
  double gpu_kernel_ms[16];/* per device share: kernel time (kernel_ms is their maximum) */
  uint32_t kernel_kind;    /* 0 = renderKernel (one path per lane), 1 = renderKernelCoop (one path per warp, scan stream), 2 = wavefront (wfTraverse + wfShade rounds) */
  uint32_t rounds;         /* wavefront: rounds of (shade, traverse) launched */
  double tail_ms;          /* the persistent kernel's drain: from the first warp that found the path index space empty to the exit of
                              the last warp (device globaltimer; the maximum over the devices of a multi-GPU call) */
} RtStats;
enum RtReduceKind { RT_REDUCE_NONE = 0, RT_REDUCE_PEER = 1 /* one kernel on the scene's device reading the peers' framebuffers over NVLink */,
                    RT_REDUCE_NCCL = 2 /* ncclReduce(sum) onto the scene's device */, RT_REDUCE_STAGED = 3 /* no peer access: cudaMemcpyPeer + add */ };
/* counters (algorithmic work, same definitions as the oracle's): */
enum RtCounter { RT_CTR_PATHS = 0, RT_CTR_BOUNCES = 1, RT_CTR_BOX_TESTS = 2, RT_CTR_SPHERE_TESTS = 3, RT_CTR_MOVING_TESTS = 4,
                 RT_CTR_RECT_TESTS = 5, RT_CTR_CUBOID_TESTS = 6, RT_CTR_TRANSLATE = 7, RT_CTR_ROTATE = 8, RT_CTR_MEDIUM = 9,
                 RT_CTR_SPHERE_HITS = 10, RT_CTR_NOISE_CALLS = 11, RT_CTR_TEXELS = 12, RT_CTR_QUERIES = 13, RT_CTR_VISITS = 14,
                 RT_CTR_UNHITTABLE = 15,
                 /* shading events (the "+ shading" term of SURVEY 8d's flop count): scatter by material (Lib.hs:822-865), emitted
                    (880-885), htblRandom on the light list (707-724), htblPdfValue leaf formulas (673-705), checker levels (498-501) */
                 RT_CTR_SCATTER_LAMBERTIAN = 16, RT_CTR_SCATTER_METAL = 17, RT_CTR_SCATTER_DIELECTRIC = 18, RT_CTR_SCATTER_ISOTROPIC = 19,
                 RT_CTR_EMITTED = 20, RT_CTR_LIGHT_SAMPLES = 21, RT_CTR_LIGHT_PDF_LEAVES = 22, RT_CTR_CHECKER_EVALS = 23 };

typedef struct RtLimits {
  uint32_t max_stack;        /* traversal stack entries per path the kernels can be launched with */
  uint32_t max_frames;       /* nested Translate/Rotate frames above one primitive */
  uint32_t max_medium_depth; /* ConstantMedium inside a ConstantMedium boundary */
  uint32_t max_light_depth;  /* depth of the lights BVH walked by htblPdfValue */
  uint32_t max_tex_depth;    /* nested CheckerTexture levels */
} RtLimits;

/* Threading: every entry point may be called from any OS thread (Haskell `safe` foreign calls migrate).  Calls on ONE scene handle
 * are serialised inside the library (they share the device copy of the scene, its work counter and its staging buffers); calls on
 * different handles run concurrently.  A handle destroyed while another thread is inside a call on it stays valid until that call
 * returns.  Asynchronous rt_render_device launches of one scene must be issued on ONE stream at a time (same work counter).
 * rt_last_error() is thread-local. */
typedef uint64_t rt_scene_t;

int rt_abi_version(void);
const char* rt_last_error(void);
int rt_device_count(void);             /* number of usable CUDA devices (0 if none) */
int rt_get_limits(RtLimits* out);

/* Multi-GPU set-up (optional).  The reference parallelises INSIDE its one pure entry point (runRender sparks, Lib.hs:1519-1520),
 * so the drop-in shards inside rt_render: rt_init(n) makes n devices the default for every later rt_render whose params.n_gpus is 0
 * (n = 0: all visible devices), enables peer access between them and, with RT_INIT_NCCL in flags, creates one NCCL communicator per
 * device (ncclCommInitAll, libnccl.so.2 loaded at run time) so that the reduction is an ncclReduce instead of the library's own
 * peer-memory kernel.  Returns the number of devices in use, or a negative error.  Without rt_init the library uses one device. */
#define RT_INIT_NCCL 1u
int rt_init(int n_gpus, uint32_t flags);

/* Validates and uploads a scene to `device` (-1 = current).  Replaces nothing in the reference: it is the
 * receiving end of the flattener. */
int rt_scene_create(const RtSceneDesc* desc, int device, rt_scene_t* out);
int rt_scene_destroy(rt_scene_t scene);

/* runRender's numeric core (Lib.hs:1491-1523 minus albedoToColor): adds, for every pixel of the requested region,
 * the radiance of samples [sample_offset, sample_offset+sample_count) to out_rgb_sum.
 *   out_rgb_sum: HOST double[height*width*3], row 0 = top row (y = height-1, Lib.hs:1488-1489); the call
 *   OVERWRITES the region it renders with that call's sums (other pixels are left untouched).
 * The caller divides by numSamples (Lib.hs:1349) and applies albedoToColor (Lib.hs:287-292) — or calls rt_tonemap. */
int rt_render(rt_scene_t scene, const RtRenderParams* params, double* out_rgb_sum, RtStats* stats);

/* Same, with out_rgb_sum a DEVICE pointer on the scene's device (double[height*width*3]); the kernel ACCUMULATES
 * into it (the caller zeroes it), on `cuda_stream` (a cudaStream_t passed as void*, NULL = default stream), and
 * the call returns without synchronising unless stats != NULL.  Used by the multi-GPU host to reduce partial
 * framebuffers with NCCL without host staging. */
int rt_render_device(rt_scene_t scene, const RtRenderParams* params, double* d_out_rgb_sum, void* cuda_stream, RtStats* stats);

/* NO COUNTERPART IN THE REFERENCE — added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  rt_render's contract with a DEVICE output buffer and
 * NO allocation: renderKernel's scratch — its work-counter block and the double framebuffer it accumulates into — lives in a
 * static __device__ array of the library, RT_STATIC_MAX_PIXELS * 24 + 512 bytes = 6 MiB + 512 B, held by every GPU that loads the
 * library until the process exits.  The call clears the part of the scratch it uses, launches renderKernel on `cuda_stream`
 * (cudaStream_t as void*, NULL = default stream), then copies the rendered region onto d_out_rgb_sum (double[height*width*3] on
 * the scene's device; the region is OVERWRITTEN, other pixels are left untouched) on the same stream, and returns without
 * synchronising.  It calls no cudaMalloc / cudaFree / event or stream creation.
 *   - width*height > RT_STATIC_MAX_PIXELS  -> RT_ERR_UNSUPPORTED (message names both numbers); use rt_render_device.
 *   - renderKernel only: RT_FLAG_COOP / RT_FLAG_WAVEFRONT / n_gpus > 1 -> RT_ERR_UNSUPPORTED.
 *   - the scene must already be on the device in the requested arithmetic (rt_scene_create uploads fp32 / mixed; the first fp64
 *     call of another entry point uploads fp64), else RT_ERR_UNSUPPORTED.
 *   - ALL calls on one GPU share the one array (whatever the scene handle) and it keeps the last call's contents: calls whose
 *     GPU work overlaps (different streams, or another thread before the first call's stream work has finished) RACE.  Issue them
 *     on one stream, or order the streams with events.
 *   - the first call on a device resolves the array's device address (cudaGetSymbolAddress; may load the module): make one
 *     warm-up call outside graph capture. */
#define RT_STATIC_MAX_PIXELS 262144u /* 512 x 512: covers the reference's 600x400 and 500x500 frames */
int rt_render_device_static(rt_scene_t scene, const RtRenderParams* params, double* d_out_rgb_sum, void* cuda_stream);

/* Same as rt_render with the algorithmic-work counters of RtStats filled (slower instrumented kernel). */
int rt_render_counted(rt_scene_t scene, const RtRenderParams* params, double* out_rgb_sum, RtStats* stats);

/* `hit world ray t_min t_max` (Lib.hs:970-1109) for n explicit rays — the parity probe for hit ids and BVH
 * traversal order.  rays: [n][7] origin dir time.  Media draw from streamGen(seed, ray index, 0).
 * Outputs (host arrays, any may be NULL): hit[n] 0/1; rec[n][9] = t p.xyz normal.xyz u v; prim[n] ref of the
 * primitive (sphere / rect / cuboid / medium) that produced the hit; mat[n] material index; front[n] frontFace;
 * nvisit[n] number of `hit` invocations; vhash[n] FNV-1a-style hash of the visited refs in order;
 * vlog[n][log_stride] first log_stride visited refs (RT_NONE padded). */
int rt_trace(rt_scene_t scene, uint32_t precision, uint32_t flags, uint32_t n, const double* rays, double t_min, double t_max,
             uint64_t seed, uint32_t log_stride, int32_t* hit, double* rec, uint32_t* prim, uint32_t* mat, uint8_t* front,
             uint64_t* nvisit, uint64_t* vhash, uint32_t* vlog);

/* Same as rt_render with a per-path log of the main queries (`hit world`, Lib.hs:1309) in bounce order: for path i (region pixels in
 * row-major order, samples innermost) and bounce b < max_bounces, prim_log[i*max_bounces+b] = ref of the primitive hit (RT_NONE = miss,
 * untouched entries are left as the caller initialised them) and t_log[...] = its t.  tools/parity_classify.py compares it with the
 * oracle's log to find where and why a path separates.  Test instrument (slower kernel). */
int rt_render_pathlog(rt_scene_t scene, const RtRenderParams* params, double* out_rgb_sum, uint32_t max_bounces, uint32_t* prim_log, double* t_log);

/* textureValue (Lib.hs:496-513) probe: q[n][5] = u v p.xyz -> out[n][3]; flags: 0 or RT_FLAG_STRICT_MATH */
int rt_texture_value(rt_scene_t scene, uint32_t tex, uint32_t precision, uint32_t flags, uint32_t n, const double* q, double* out);

/* Device evaluation of the shared strict-math functions (fn: 0 sin, 1 cos, 2 atan2(x, y = second operand), 3 asin, 4 log) on host
 * arrays x[n], y[n] -> out[n]; as_float != 0 evaluates the fp32 entry points.  For the host/device bit-identity test. */
int rt_strict_math_eval(uint32_t fn, uint32_t n, const double* x, const double* y, double* out, int as_float);

/* albedoToColor (Lib.hs:287-292) on the device: mean = sum / total_samples, sqrt, clamp [0,0.999], *256, floor.
 * in: host double[n] sums; out: host uint8[n].  NaN maps to 0 (GHC's floor NaN :: Word8 is platform-defined). */
int rt_tonemap(const double* rgb_sum, uint32_t n, uint32_t total_samples, uint8_t* out_rgb8);

/* Device self-test of the literal fp32 build's slab-test division (the hoisted-reciprocal form of the compiler's own IEEE
 * sequence, rt_device.cuh): n random operand pairs of the validated range against __fdiv_rn; bit-exactness requires
 * *mismatches == 0.  Not part of the reference path. */
int rt_selftest_division(uint64_t n, uint64_t seed, uint64_t* checked, uint64_t* mismatches);

/* Device self-test of the interval filter in front of that division (slabFiltered, rt_device.cuh: boxRayIntersect, Lib.hs:798-814,
 * decided from one-instruction approximate quotients whenever the comparison margin exceeds their error bound): n random and
 * adversarial (box, ray, [t_min, t_max]) triples; out3 = {triples checked, of those decided by the filter alone, filter decisions
 * that differ from the decision of the three IEEE divisions}.  Bit-exact traversal requires out3[2] == 0.  Not part of the
 * reference path. */
int rt_selftest_box_filter(uint64_t n, uint64_t seed, uint64_t out3[3]);

/* makeBVH (Lib.hs:941-961) built on the device, tree-identical to the host recursion: the node's objects are sorted stably by
 * box_min[axis] (Data.Sequence.sortBy of boxCompare, Lib.hs:963-968), split at n `div` 2; one object -> BVHNode h h (948),
 * two objects -> ordered by `comparator h1 h2 == LT`, ties swapped (949-952); box = surroundingBox of the children (929-939).
 *   n, leaf_box[n][6]   boundingBox (Lib.hs:905-927) of each object of the sequence, in sequence order: min.xyz max.xyz
 *   node_axis[n_nodes]  `floor <$> randomDoubleRM 0 3` of every node in PRE-ORDER (the order the RandomState monad draws
 *                       them: a node, then its whole left subtree, then the right one); n_nodes = rt_bvh_node_count(n)
 *   outputs, indexed by pre-order node id (root = 0): node_box[n_nodes][6], node_size[n_nodes] (bvh_size), and the children:
 *                       bit 31 set = a node (low bits: its pre-order id), clear = an object (index into leaf_box)
 *   build_ms            optional: device time of the build including its copies
 * NaN box coordinates are rejected (`compare` is not an order on them).  Replaces the host-side makeBVH for large scenes
 * (SURVEY 8(f) rank 3); the hot path never calls it. */
uint32_t rt_bvh_node_count(uint32_t n);
int rt_bvh_build(uint32_t n, const double* leaf_box, const uint8_t* node_axis, uint32_t n_nodes, double* node_box, uint32_t* node_left,
                 uint32_t* node_right, uint32_t* node_size, int device, double* build_ms);

void rt_shutdown(void);

#ifdef __cplusplus
}
#endif
#endif /* RT_ABI_H */
