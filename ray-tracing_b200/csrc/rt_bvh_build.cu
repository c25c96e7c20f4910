// makeBVH (reference src/Lib.hs:941-961) as a level-synchronous device build that reproduces the host-built tree exactly
// (SURVEY 8(f) rank 3).  The reference draws an axis at every node (one generator draw per node, pre-order), sorts the node's
// objects STABLY by the minimum corner of their bounding box along that axis (Data.Sequence.sortBy), splits at n `div` 2 and
// recurses; one object becomes BVHNode h h, two objects are ordered by `comparator h1 h2 == LT` (ties swap, unlike the sort).
//
// The SHAPE of the tree (segment starts / counts, pre-order numbering) depends on n alone, so the host walks it once; only the
// permutation of the objects depends on the data.  Per level the device
//   1. keys every object by box_min[axis of its segment] (finished segments get a constant key),
//   2. stable radix sort of (key, position) over ALL objects, then a stable radix sort by segment start: objects are again
//      grouped by segment, each segment stably ordered by key -- exactly sortBy, for every node of the level at once,
// then assigns leaf children and folds the boxes bottom-up, level by level (surroundingBox, Lib.hs:929-939).
// Doubles throughout (the reference's precision); -0.0 and +0.0 compare equal as in Haskell's `compare`.
#include <cub/device/device_radix_sort.cuh>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/rt_abi.h"

namespace rtb {
void setLastError(const std::string& s);  // rt_abi.cu

namespace {
constexpr uint32_t kNodeBit = 0x80000000u;  // child encoding of rt_bvh_build's outputs: node (pre-order index) vs object index

struct ShapeNode { uint32_t start, count, left, right, level; };

// total order of doubles as unsigned integers; both zeros map to the same key (compare (-0.0) 0.0 == EQ)
__device__ __forceinline__ unsigned long long sortableKey(double d) {
  if (d == 0.0) d = 0.0;
  const unsigned long long u = (unsigned long long)__double_as_longlong(d);
  return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}

// segStart[]: starts of this level's segments (active nodes of the level and every finished shallower node), ascending
__global__ void keyKernel(uint32_t n, const double* __restrict__ box, const uint32_t* __restrict__ perm, const uint32_t* __restrict__ segStart,
                          const uint8_t* __restrict__ segAxis, uint32_t nSeg, unsigned long long* __restrict__ key, uint32_t* __restrict__ pos,
                          uint32_t* __restrict__ segOf) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  uint32_t lo = 0, hi = nSeg;  // last segment whose start <= p
  while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (segStart[mid] <= p) lo = mid; else hi = mid; }
  const uint8_t ax = segAxis[lo];  // 0..2, or 3 = finished segment (keeps its order)
  key[p] = ax < 3 ? sortableKey(box[(size_t)perm[p] * 6 + ax]) : 0ull;
  pos[p] = p;
  segOf[p] = segStart[lo];
}
__global__ void gatherSegKernel(uint32_t n, const uint32_t* __restrict__ posSorted, const uint32_t* __restrict__ segOf, uint32_t* __restrict__ segSorted) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) segSorted[i] = segOf[posSorted[i]];
}
__global__ void permuteKernel(uint32_t n, const uint32_t* __restrict__ posFinal, const uint32_t* __restrict__ permIn, uint32_t* __restrict__ permOut) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) permOut[i] = permIn[posFinal[i]];
}

__device__ __forceinline__ double hmin(double x, double y) { return x <= y ? x : y; }  // Prelude min / max on Double
__device__ __forceinline__ double hmax(double x, double y) { return x <= y ? y : x; }

// one level of nodes, deepest first: children (leaf assignment for count <= 2) and the surrounding box
__global__ void foldKernel(uint32_t nIds, const uint32_t* __restrict__ ids, const ShapeNode* __restrict__ shape, const uint8_t* __restrict__ axis,
                           const uint32_t* __restrict__ perm, const double* __restrict__ leafBox, double* __restrict__ nodeBox,
                           uint32_t* __restrict__ nodeLeft, uint32_t* __restrict__ nodeRight, uint32_t* __restrict__ nodeSize) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nIds) return;
  const uint32_t id = ids[t];
  const ShapeNode s = shape[id];
  uint32_t l, r;
  const double *bl, *br;
  if (s.count == 1) {
    l = r = perm[s.start]; bl = br = leafBox + (size_t)l * 6;
  } else if (s.count == 2) {
    const uint32_t h1 = perm[s.start], h2 = perm[s.start + 1];
    const bool lt = leafBox[(size_t)h1 * 6 + axis[id]] < leafBox[(size_t)h2 * 6 + axis[id]];  // comparator h1 h2 == LT, else swapped
    l = lt ? h1 : h2; r = lt ? h2 : h1;
    bl = leafBox + (size_t)l * 6; br = leafBox + (size_t)r * 6;
  } else {
    l = s.left | kNodeBit; r = s.right | kNodeBit;
    bl = nodeBox + (size_t)s.left * 6; br = nodeBox + (size_t)s.right * 6;
  }
  double* o = nodeBox + (size_t)id * 6;
  o[0] = hmin(bl[0], br[0]); o[1] = hmin(bl[1], br[1]); o[2] = hmin(bl[2], br[2]);
  o[3] = hmax(bl[3], br[3]); o[4] = hmax(bl[4], br[4]); o[5] = hmax(bl[5], br[5]);
  nodeLeft[id] = l; nodeRight[id] = r; nodeSize[id] = s.count;
}

uint32_t nodeCount(uint32_t n) {  // N(1) = N(2) = 1, N(n) = 1 + N(n / 2) + N(n - n / 2)
  if (n <= 2) return n ? 1u : 0u;
  return 1u + nodeCount(n / 2) + nodeCount(n - n / 2);
}

struct DevBuf {
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1); }
  template <class T> T* as() { return static_cast<T*>(p); }
};
}  // namespace
}  // namespace rtb

extern "C" uint32_t rt_bvh_node_count(uint32_t n) { return rtb::nodeCount(n); }

#define BVH_CU(call)                                                                                  \
  do {                                                                                                \
    cudaError_t e_ = (call);                                                                          \
    if (e_ != cudaSuccess) { rtb::setLastError(std::string(#call) + ": " + cudaGetErrorString(e_)); return RT_ERR_CUDA; } \
  } while (0)

extern "C" int rt_bvh_build(uint32_t n, const double* leaf_box, const uint8_t* node_axis, uint32_t n_nodes, double* node_box, uint32_t* node_left,
                            uint32_t* node_right, uint32_t* node_size, int device, double* build_ms) {
  using namespace rtb;
  if (!n || !leaf_box || !node_axis || !node_box || !node_left || !node_right || !node_size) { setLastError("rt_bvh_build: null argument or empty input"); return RT_ERR_ARG; }
  if (n >= kNodeBit) { setLastError("rt_bvh_build: too many objects"); return RT_ERR_ARG; }
  if (n_nodes != nodeCount(n)) { setLastError("rt_bvh_build: n_nodes != rt_bvh_node_count(n)"); return RT_ERR_ARG; }
  for (uint32_t i = 0; i < n_nodes; ++i) if (node_axis[i] > 2) { setLastError("rt_bvh_build: axis out of range"); return RT_ERR_ARG; }
  for (size_t i = 0; i < (size_t)n * 6; ++i) if (!(leaf_box[i] == leaf_box[i])) { setLastError("rt_bvh_build: NaN box coordinate (compare is not an order)"); return RT_ERR_ARG; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); setLastError("no CUDA device available (this library has no CPU path)"); return RT_ERR_CUDA; }
  if (device < 0) BVH_CU(cudaGetDevice(&device));
  if (device >= ndev) { setLastError("device ordinal out of range"); return RT_ERR_ARG; }
  BVH_CU(cudaSetDevice(device));

  // ---- the shape: pre-order walk; depends on n only
  std::vector<ShapeNode> shape(n_nodes);
  uint32_t maxLevel = 0;
  {
    uint32_t next = 0;
    // explicit pre-order: a node takes the next id when it is visited; left subtree first
    struct Frame { uint32_t start, count, level, parent; bool isRight; };
    std::vector<Frame> st; st.push_back({0, n, 0, 0xffffffffu, false});
    while (!st.empty()) {
      const Frame f = st.back(); st.pop_back();
      const uint32_t id = next++;
      shape[id] = ShapeNode{f.start, f.count, 0, 0, f.level};
      if (f.parent != 0xffffffffu) { if (f.isRight) shape[f.parent].right = id; else shape[f.parent].left = id; }
      maxLevel = std::max(maxLevel, f.level);
      if (f.count > 2) {
        const uint32_t half = f.count / 2;
        st.push_back({f.start + half, f.count - half, f.level + 1, id, true});  // popped after the whole left subtree
        st.push_back({f.start, half, f.level + 1, id, false});
      }
    }
  }
  std::vector<std::vector<uint32_t>> byLevel(maxLevel + 1);
  for (uint32_t i = 0; i < n_nodes; ++i) byLevel[shape[i].level].push_back(i);

  // ---- device buffers
  DevBuf dBox, dAxis, dShape, dPermA, dPermB, dKeyA, dKeyB, dPosA, dPosB, dSegOf, dSegA, dSegB, dSegStart, dSegAxis, dIds, dNodeBox, dLeft, dRight, dSize, dTmp;
  BVH_CU(dBox.alloc((size_t)n * 6 * 8)); BVH_CU(dAxis.alloc(n_nodes)); BVH_CU(dShape.alloc((size_t)n_nodes * sizeof(ShapeNode)));
  BVH_CU(dPermA.alloc((size_t)n * 4)); BVH_CU(dPermB.alloc((size_t)n * 4)); BVH_CU(dKeyA.alloc((size_t)n * 8)); BVH_CU(dKeyB.alloc((size_t)n * 8));
  BVH_CU(dPosA.alloc((size_t)n * 4)); BVH_CU(dPosB.alloc((size_t)n * 4)); BVH_CU(dSegOf.alloc((size_t)n * 4)); BVH_CU(dSegA.alloc((size_t)n * 4)); BVH_CU(dSegB.alloc((size_t)n * 4));
  BVH_CU(dSegStart.alloc((size_t)n_nodes * 4)); BVH_CU(dSegAxis.alloc(n_nodes)); BVH_CU(dIds.alloc((size_t)n_nodes * 4));
  BVH_CU(dNodeBox.alloc((size_t)n_nodes * 6 * 8)); BVH_CU(dLeft.alloc((size_t)n_nodes * 4)); BVH_CU(dRight.alloc((size_t)n_nodes * 4)); BVH_CU(dSize.alloc((size_t)n_nodes * 4));
  size_t tmp1 = 0, tmp2 = 0;
  int segBits = 1; while ((1ull << segBits) < (unsigned long long)n) ++segBits;
  BVH_CU(cub::DeviceRadixSort::SortPairs(nullptr, tmp1, dKeyA.as<unsigned long long>(), dKeyB.as<unsigned long long>(), dPosA.as<uint32_t>(), dPosB.as<uint32_t>(), (int)n));
  BVH_CU(cub::DeviceRadixSort::SortPairs(nullptr, tmp2, dSegA.as<uint32_t>(), dSegB.as<uint32_t>(), dPosB.as<uint32_t>(), dPosA.as<uint32_t>(), (int)n, 0, segBits));
  BVH_CU(dTmp.alloc(std::max(tmp1, tmp2)));
  size_t tmpBytes = std::max(tmp1, tmp2);

  struct Ev { cudaEvent_t e = nullptr; ~Ev() { if (e) cudaEventDestroy(e); } } ev0, ev1;  // destroyed on every early error return too
  BVH_CU(cudaEventCreate(&ev0.e)); BVH_CU(cudaEventCreate(&ev1.e));
  const cudaEvent_t e0 = ev0.e, e1 = ev1.e;
  BVH_CU(cudaEventRecord(e0));
  BVH_CU(cudaMemcpy(dBox.p, leaf_box, (size_t)n * 6 * 8, cudaMemcpyHostToDevice));
  BVH_CU(cudaMemcpy(dAxis.p, node_axis, n_nodes, cudaMemcpyHostToDevice));
  BVH_CU(cudaMemcpy(dShape.p, shape.data(), (size_t)n_nodes * sizeof(ShapeNode), cudaMemcpyHostToDevice));
  {
    std::vector<uint32_t> iota(n); for (uint32_t i = 0; i < n; ++i) iota[i] = i;
    BVH_CU(cudaMemcpy(dPermA.p, iota.data(), (size_t)n * 4, cudaMemcpyHostToDevice));
  }
  uint32_t* perm = dPermA.as<uint32_t>(); uint32_t* permAlt = dPermB.as<uint32_t>();
  const uint32_t tpb = 256, gridN = (n + tpb - 1) / tpb;

  // ---- top-down: one segmented stable sort per level that still has a node with more than two objects.  The level's segment
  // list (ascending starts) is derived from the previous one: an active node is replaced by its two children, a finished one stays.
  constexpr uint32_t kFinished = 0xffffffffu;
  std::vector<std::pair<uint32_t, uint32_t>> cur{{0u, 0u}}, nxt;  // (start, node id | kFinished)
  std::vector<uint32_t> ss; std::vector<uint8_t> sa;
  for (uint32_t L = 0; L <= maxLevel; ++L) {
    bool any = false;
    ss.resize(cur.size()); sa.resize(cur.size());
    nxt.clear(); nxt.reserve(cur.size() * 2);
    for (size_t i = 0; i < cur.size(); ++i) {
      const uint32_t id = cur[i].second;
      const bool active = id != kFinished && shape[id].count > 2;
      any |= active;
      ss[i] = cur[i].first; sa[i] = active ? node_axis[id] : (uint8_t)3;
      if (active) { nxt.push_back({shape[shape[id].left].start, shape[id].left}); nxt.push_back({shape[shape[id].right].start, shape[id].right}); }
      else nxt.push_back({cur[i].first, kFinished});
    }
    cur.swap(nxt);
    if (!any) break;
    BVH_CU(cudaMemcpyAsync(dSegStart.p, ss.data(), ss.size() * 4, cudaMemcpyHostToDevice));
    BVH_CU(cudaMemcpyAsync(dSegAxis.p, sa.data(), sa.size(), cudaMemcpyHostToDevice));
    keyKernel<<<gridN, tpb>>>(n, dBox.as<double>(), perm, dSegStart.as<uint32_t>(), dSegAxis.as<uint8_t>(), (uint32_t)ss.size(), dKeyA.as<unsigned long long>(),
                              dPosA.as<uint32_t>(), dSegOf.as<uint32_t>());
    BVH_CU(cub::DeviceRadixSort::SortPairs(dTmp.p, tmpBytes, dKeyA.as<unsigned long long>(), dKeyB.as<unsigned long long>(), dPosA.as<uint32_t>(), dPosB.as<uint32_t>(), (int)n));
    gatherSegKernel<<<gridN, tpb>>>(n, dPosB.as<uint32_t>(), dSegOf.as<uint32_t>(), dSegA.as<uint32_t>());
    BVH_CU(cub::DeviceRadixSort::SortPairs(dTmp.p, tmpBytes, dSegA.as<uint32_t>(), dSegB.as<uint32_t>(), dPosB.as<uint32_t>(), dPosA.as<uint32_t>(), (int)n, 0, segBits));
    permuteKernel<<<gridN, tpb>>>(n, dPosA.as<uint32_t>(), perm, permAlt);
    std::swap(perm, permAlt);
    BVH_CU(cudaStreamSynchronize(0));  // ss / sa are rewritten for the next level
  }
  // ---- bottom-up: leaf assignment and boxes, deepest level first
  for (int L = (int)maxLevel; L >= 0; --L) {
    const std::vector<uint32_t>& ids = byLevel[L];
    BVH_CU(cudaMemcpyAsync(dIds.p, ids.data(), ids.size() * 4, cudaMemcpyHostToDevice));
    foldKernel<<<((uint32_t)ids.size() + tpb - 1) / tpb, tpb>>>((uint32_t)ids.size(), dIds.as<uint32_t>(), dShape.as<ShapeNode>(), dAxis.as<uint8_t>(), perm, dBox.as<double>(),
                                                              dNodeBox.as<double>(), dLeft.as<uint32_t>(), dRight.as<uint32_t>(), dSize.as<uint32_t>());
    BVH_CU(cudaStreamSynchronize(0));
  }
  BVH_CU(cudaGetLastError());
  BVH_CU(cudaMemcpy(node_box, dNodeBox.p, (size_t)n_nodes * 6 * 8, cudaMemcpyDeviceToHost));
  BVH_CU(cudaMemcpy(node_left, dLeft.p, (size_t)n_nodes * 4, cudaMemcpyDeviceToHost));
  BVH_CU(cudaMemcpy(node_right, dRight.p, (size_t)n_nodes * 4, cudaMemcpyDeviceToHost));
  BVH_CU(cudaMemcpy(node_size, dSize.p, (size_t)n_nodes * 4, cudaMemcpyDeviceToHost));
  BVH_CU(cudaEventRecord(e1)); BVH_CU(cudaEventSynchronize(e1));
  float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
  if (build_ms) *build_ms = ms;
  return RT_OK;
}
