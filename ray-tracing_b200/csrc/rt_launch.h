// rt_launch.h — argument blocks and launcher declarations shared by the kernel TUs and the C-ABI host code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtb {

struct RenderArgs {
  uint32_t W, H;                   // full image
  uint32_t s0, ns;                 // sample range [s0, s0+ns)
  uint32_t maxDepth;
  uint32_t row0, row1, col0, col1; // region (rows counted from the top)
  uint32_t rowStep;                // rows row0, row0 + rowStep, ... < row1 (1 = every row; n = the interleaved row share of one of n devices)
  uint32_t chunk;                  // path indices a warp takes per atomic (at most; guided: fewer towards the end of the index space)
  uint32_t chunkMin, chunkDiv;     // smallest chunk; a warp asks for (indices left) / chunkDiv
  uint32_t stackCap;               // traversal stack entries per thread (set by the launcher)
  uint32_t lightCap;               // entries of htblPdfValue's walk stack per thread: levels of the lights BVH, 0 when the light list is no BVH
  uint32_t shadeMin, rectMin, rareMin;  // full vote: run the shade / rect+cuboid / rare class as soon as this many lanes wait for it
  uint32_t keepNode, keepLeaf;     // run a node / sphere step without a full vote when at least this many lanes wait for one
  unsigned long long seed;
  unsigned long long total_paths;  // region pixels * ns
  unsigned long long* counter;     // device: [0] next free path index ... [kCtrTailBase], [+1] globaltimer stamps of the tail (rt_device.cuh kCtr*)
  unsigned long long* ctr;         // device: kWorkCounters algorithmic-work counters + kSchedCounters scheduler statistics (INSTR=1) or null
  double* out;                     // device: double[H*W*3] radiance sums, accumulated with atomicAdd
  uint32_t* plogPrim; double* plogT; uint32_t plogStride;  // INSTR=3: per-path bounce log [total_paths][plogStride]
  uint32_t wfChunk;                // wavefront: consecutive path indices (= consecutive samples of a pixel) a slot takes at a time
  float wfGuideInv;                // wavefront: a slot asks for (indices left) * wfGuideInv, at least 1 and at most wfChunk (guided: short drain at the end of the frame)
  uint32_t wfSort;                 // wavefront: shade blocks sort their slots by material (RTB_WF_SORT=0 turns it off for A/B runs)
  uint32_t wfMedia;                // wavefront: the scene has ConstantMedium (the traversal draws from the path's stream)
};

// path-state slots of the wavefront kernels (rt_wavefront.cuh): struct of arrays over P slots, one device allocation
struct WfState {
  uint32_t P, pad;
  float *ox, *oy, *oz, *dx, *dy, *dz, *time;  // the item's ray (world frame) and the path's shutter time
  float *bx, *by, *bz;                        // throughput
  unsigned long long *rngS, *rngG;
  uint32_t *depth, *phase;
  float *v2, *spdf;                           // WF_LPDF: cosine-pdf value and scatteringPdf of the sampled direction (Lib.hs:374-382, 874-878)
  float *rT, *rW; uint32_t *rPrim, *rFlags, *rChain;  // result: t, light-pdf weight, primitive, Trav flags, frames of the hit [4][P]
  double* acc; uint32_t* accPix;              // per-slot radiance accumulator of its current pixel [3][P]
  unsigned long long* chunkNext; uint32_t* chunkLeft;  // the slot's chunk of consecutive path indices
  unsigned long long* ctl;                    // [0] traversal cursor over slots, [1] live slots of the shade in flight, [2] done, [3] next free path index, [4] shade blocks finished
};
size_t wfLayout(uint32_t P, char* base, WfState* out);  // fills the pointers for a blob at `base` (may be null: size only); returns bytes
// `rounds` rounds of (shade, traverse) on `st`, preceded by the slot initialisation when `init`; prec 0 / 1 only
cudaError_t launchWavefront(int prec, const void* dscene, const RenderArgs& a, const WfState& w, int stackDepth, int numSMs, int rounds, bool init, cudaStream_t st, int* gridOut);

struct TraceArgs {
  uint32_t n, logStride;
  uint32_t stackCap;
  double tmin, tmax;
  unsigned long long seed;
  const double* rays;  // device [n][7]
  int32_t* hit; double* rec; uint32_t* prim; uint32_t* mat; uint8_t* front;
  unsigned long long* nvisit; unsigned long long* vhash; uint32_t* vlog;
};

// prec: 0 fp32, 1 mixed, 2 fp64 ; dscene points at a host copy of DScene<float> (prec 0/1) or DScene<double> (2)
// MODE: 0 literal arithmetic (rt_launch_exact.cu), 1 fast math (rt_launch_fast.cu), 2 strict math (rt_launch_strict.cu); instr: 0, 1 (counters), 3 (path log)
template <int MODE> cudaError_t launchRender(int prec, int instr, const void* dscene, const RenderArgs& a, int stackDepth, int numSMs, cudaStream_t st, int* gridOut);
template <int MODE> cudaError_t launchTrace(int prec, const void* dscene, const TraceArgs& a, int stackDepth, cudaStream_t st);
// scan-stream (warp-cooperative, rt_coop.cuh) variants for eligible scenes; MODE 0 and 2 only
template <int MODE> cudaError_t launchRenderCoop(int prec, int instr, const void* dscene, const RenderArgs& a, int numSMs, cudaStream_t st, int* gridOut);
template <int MODE> cudaError_t launchTraceCoop(int prec, const void* dscene, const TraceArgs& a, cudaStream_t st);
template <int MODE> cudaError_t launchTexture(int prec, const void* dscene, uint32_t tex, uint32_t n, const double* q, double* out, cudaStream_t st);
cudaError_t launchStrictMathEval(uint32_t fn, uint32_t n, const double* x, const double* y, double* out, int as_float, cudaStream_t st);
cudaError_t launchDivSelftest(unsigned long long n, unsigned long long seed, unsigned long long* out2, cudaStream_t st);
cudaError_t launchBoxFilterSelftest(unsigned long long n, unsigned long long seed, unsigned long long* out3, cudaStream_t st);
// dst[i] += src[0][i] + ... + src[n-1][i] (left to right), src may be peer-device memory (NVLink loads)
cudaError_t launchReducePeer(double* dst, const double* const* src, int n, size_t elems, int numSMs, cudaStream_t st);
cudaError_t launchTonemap(const double* sum, uint32_t n, double ns, uint8_t* out, cudaStream_t st);

}  // namespace rtb
