// rt_wavefront.cuh — the hot path split into a TRAVERSAL kernel and a SHADE kernel over path-state slots in HBM.
// Included by rt_kernels.cuh.
//
// Why: shading-heavy scenes (cfg4, the Cornell box) run renderKernel at 28 % issue-active: one sample walks 64 KB of SASS against a
// 32 KB L1.5 instruction cache, and inside the shade class the materials / samplers diverge down to 3-8 lanes per instruction
// (profiles/r1_ncu_renderKernel_cornell_spp32_v15.txt).  Splitting gives a traversal kernel whose whole code is the ~10 KB of vote +
// step functions (no spills at 64 registers, 8 blocks per SM) and a shade kernel in which every thread of a warp shades — no lanes
// parked in traversal classes — with its own instruction-cache footprint.
//
// One SLOT = one in-flight camera path.  A frame is rendered in rounds of (wfTraverse, wfShade) over P slots:
//   wfShade    one thread per slot: consume the finished item (apply the light-pdf weight, build the hit record, texture, scatter —
//              the SAME shadeStep as renderKernel), or end the path (accumulate, then take the next path index of the slot's chunk:
//              path regeneration keeps every slot busy until the frame's index space is empty), and write the next ITEM:
//              the scattered ray, and for a Lambertian bounce under a light list the mixture-pdf inputs;
//   wfTraverse persistent warps with the class-voting scheduler of renderKernel (node / sphere / rect / rare classes, same step
//              functions): a lane takes an item, runs htblPdfValue's light queries (lpdfAdvance, unchanged) and then the bounce's
//              `hit world`, writes (t, primitive, frame chain, flags, weight) and takes the next item.
// Per path the order of steps, of RNG draws and of arithmetic operations is renderKernel's, so the image is the same up to the order
// of the double additions into a pixel (tests/test_gpu_wavefront.py).  State per slot ~150 B; a bounce moves ~200 B through HBM:
// at 10^9 bounces/s that is 0.2 TB/s, 3 % of the measured HBM bandwidth.  fp32 / mixed arithmetic, literal math only; counters,
// path logs, strict and fast math stay on renderKernel.
#pragma once

namespace rtb {

enum : uint32_t { WF_DEAD = 0u, WF_NEW = 1u, WF_MAIN = 2u, WF_LPDF = 3u, WF_LPDF_NOMAIN = 4u };  // slot phase = what the next kernel finds
constexpr uint32_t WF_R_NOMAIN = 1u << 30;  // result flag: the item ended after its light queries (depth exhausted)

constexpr int kWfCtlWords = 8;

// ---- traversal ---------------------------------------------------------------------------------------------------------------------
template <class T, class S>
__global__ void __launch_bounds__(kBlock, 8) wfTraverseKernel(const __grid_constant__ DScene<T> sc, const __grid_constant__ RenderArgs a, const __grid_constant__ WfState w) {
  extern __shared__ uint32_t smem[];
  if (w.ctl[2]) return;  // frame finished: the remaining launches of the batch are no-ops
  const unsigned lane = threadIdx.x & 31u;
  const unsigned ltMask = (1u << lane) - 1u;
  Instr<0> ins;
  Trav<T> tr;
  tr.col = smem + threadIdx.x;
  tr.cur = kEmpty; tr.sp = 0; tr.flags = 0; tr.bprim = 0; tr.tmin = 0; tr.tbest = 0;
  tr.o = mk<T>(0, 0, 0); tr.d = mk<T>(0, 0, 1); tr.inv = mk<T>(0, 0, 0);
  PathState<T> ps;
  ps.phase = PH_IDLE; ps.beta = mk<T>(1, 1, 1); ps.time = 0; ps.depthLeft = 0; ps.rng = Rng{0, 1};
  uint32_t slot = kNone;  // the item this lane is working on
  volatile uint32_t* ws = smem + (size_t)columnWords<T>(a.stackCap, a.lightCap) * blockDim.x + (threadIdx.x >> 5) * 4;
  if (lane == 0) { ws[2] = 0; ws[3] = 0; }
  __syncwarp();
  bool exhausted = false;
  const uint32_t chunk = a.chunk;

  for (;;) {
    const uint32_t kind = tr.cur >> 28;
    const bool isNode = kind == K_BVH, isLeaf = kind == K_SPHERE || kind == K_MOVING;
    const int cls = stepClass(tr.cur);
    int pick = 0;
    const int cN = __popc(__ballot_sync(0xffffffffu, isNode));
    if (cN < (int)a.keepNode) {
      const int cL = __popc(__ballot_sync(0xffffffffu, isLeaf));
      pick = 1;
      if (cL < (int)a.keepLeaf) {
        const int c2 = __popc(__ballot_sync(0xffffffffu, cls == 2)), c3 = __popc(__ballot_sync(0xffffffffu, cls == 3)),
                  c4 = __popc(__ballot_sync(0xffffffffu, cls == 4));
        if ((cN | cL | c2 | c3 | c4) == 0) break;
        int best = cN;
        pick = 0;
        if (cL > best) { pick = 1; best = cL; }
        if (c4 > best || c4 >= (int)a.shadeMin) { pick = 4; best = 32; }
        if (c2 > best || c2 >= (int)a.rectMin) { pick = 2; best = 32; }
        if (c3 > best || c3 >= (int)a.rareMin) pick = 3;
      }
    }
    if (pick == 0) {
      if (isNode) stepNode<T, 0, MATH_LITERAL>(sc, tr, ins);
    } else if (pick == 1) {
      if (isLeaf) stepSphere<T, S, 0>(sc, tr, (ps.phase == PH_MAIN) ? ps.time : (T)0, ins);  // htblPdfValue traces with time 0 (Lib.hs:677)
    } else if (pick == 2) {
      if (cls == 2) stepRect<T, 0>(sc, tr, ins);
    } else if (pick == 3) {
      if (cls == 3) stepRare<T, 0, MATH_LITERAL>(sc, tr, ps.rng, ins);
    }
    if (pick != 4) {
      if (ps.phase == PH_LPDF && tr.cur == kEmpty) lpdfAdvance<T, 0, MATH_LITERAL>(sc, tr, ps, ins, a.stackCap + (uint32_t)extraWords<T>(), a.lightCap);
    } else {
      // ---- item class: (1) write the finished item back
      if (cls == 4 && slot != kNone) {
        uint32_t fl = tr.flags & 0x0000ffffu;  // F_QHIT, F_NANT, best-hit depth / aux (division modes are per query)
        if (ps.phase == PH_IDLE) fl |= WF_R_NOMAIN;
        w.rT[slot] = tr.tbest; w.rW[slot] = ps.beta.x; w.rPrim[slot] = tr.bprim; w.rFlags[slot] = fl;
        const uint32_t bd = (fl >> F_BDEPTH_SHIFT) & 7u;
        for (uint32_t i = 0; i < bd; ++i) w.rChain[(size_t)i * w.P + slot] = tr.xu(X_BCHAIN + (int)i);
        if (a.wfMedia) w.rngS[slot] = ps.rng.s;  // ConstantMedium draws inside `hit` (Lib.hs:1068): the stream moved
        slot = kNone; ps.phase = PH_IDLE;
      }
      // ---- (2) idle lanes take the next slot indices of the warp's chunk, in lane order; dead slots are skipped
      bool start = false;
      unsigned idle = __ballot_sync(0xffffffffu, cls == 4 && slot == kNone);
      while (idle) {
        uint32_t cNext = ws[2], cEnd = ws[3];
        if (cNext >= cEnd) {
          if (exhausted) break;
          unsigned long long base = 0;
          if (lane == 0) base = atomicAdd(w.ctl, (unsigned long long)chunk);
          base = __shfl_sync(0xffffffffu, base, 0);
          if (base >= w.P) { exhausted = true; break; }
          cNext = (uint32_t)base;
          cEnd = (base + chunk < w.P) ? (uint32_t)base + chunk : w.P;
          __syncwarp();
          if (lane == 0) ws[3] = cEnd;
          __syncwarp();
        }
        const unsigned rank = __popc(idle & ltMask);
        if (cls == 4 && slot == kNone && !start && cNext + rank < cEnd) {
          const uint32_t s = cNext + rank;
          const uint32_t ph = w.phase[s];
          if (ph == WF_MAIN || ph == WF_LPDF || ph == WF_LPDF_NOMAIN) {
            slot = s; start = true;
            tr.o = mk<T>(w.ox[s], w.oy[s], w.oz[s]); tr.d = mk<T>(w.dx[s], w.dy[s], w.dz[s]);
            ps.time = w.time[s];
            ps.beta = mk<T>(1, 1, 1);
            if (a.wfMedia) ps.rng = Rng{w.rngS[s], w.rngG[s]};
            if (ph == WF_MAIN) { ps.phase = PH_MAIN; ps.depthLeft = 2; }
            else {  // the light queries of the mixture pdf first (shadeStep's hand-over to lpdfAdvance, rt_render_kernel.inl)
              ps.phase = PH_LPDF; ps.depthLeft = (ph == WF_LPDF) ? 2 : 1;
              xst<T>(tr, X_LV2, w.v2[s]); xst<T>(tr, X_LSPDF, w.spdf[s]); xst<T>(tr, X_LV1, (T)0); xst<T>(tr, X_CURW, (T)1);
              tr.xu(X_LSP) = 0; tr.xu(X_CURLIGHT) = sc.lights_root;
            }
          }
        }
        const uint32_t avail = cEnd - cNext; const uint32_t n = (uint32_t)__popc(idle);
        __syncwarp();
        if (lane == 0) ws[2] = cNext + ((n < avail) ? n : avail);
        __syncwarp();
        idle = __ballot_sync(0xffffffffu, cls == 4 && slot == kNone && !start);
      }
      if (start) {
        travBegin<T, MATH_LITERAL>(sc, tr, ps.phase == PH_MAIN ? sc.world_root : tr.xu(X_CURLIGHT), tr.o, tr.d, kEps<T>(), kInf<T>());
      } else if (cls == 4 && slot == kNone) {
        tr.cur = kDead;  // only reached when the slot space is exhausted
      }
    }
  }
}

// ---- shade ---------------------------------------------------------------------------------------------------------------------------
constexpr int kShadeBlock = 512;  // slots per shade block = the group that is sorted by material (4 column groups of kBlock threads)

// what a slot's thread will execute: 0 = no hit to shade (new slot, miss, depth exhausted: accumulate / regenerate), 1 + material kind
// for a hit, 7 = dead slot
template <class T> RTB_D uint32_t wfShadeBin(const DScene<T>& sc, const WfState& w, uint32_t s) {
  if (s >= w.P) return 7u;
  const uint32_t ph = w.phase[s];
  if (ph == WF_DEAD) return 7u;
  if (ph == WF_NEW) return 0u;
  const uint32_t fl = w.rFlags[s];
  if ((fl & WF_R_NOMAIN) || !(fl & F_QHIT)) return 0u;
  const uint32_t prim = w.rPrim[s], kind = prim >> 28, idx = prim & kIdxMask;
  uint32_t mat;
  if (kind == K_SPHERE) mat = __ldg(sc.sphere_mat + idx);
  else if (kind == K_MOVING) mat = bitsOf(ldq(sc.moving + 3 * (size_t)idx + 2).y);
  else if (kind == K_CUBOID) mat = bitsOf(ldq(sc.cuboids + 2 * (size_t)idx).w);
  else if (kind == K_MEDIUM) mat = bitsOf(ldq(sc.media + idx).y);
  else mat = bitsOf(ldq(sc.rects + 2 * (size_t)idx + 1).y);
  return 1u + bitsOf(ldq(sc.mats + mat).x);
}

template <class T>
__global__ void __launch_bounds__(kShadeBlock) wfShadeKernel(const __grid_constant__ DScene<T> sc, const __grid_constant__ RenderArgs a, const __grid_constant__ WfState w) {
  extern __shared__ uint32_t smem[];
  if (w.ctl[2]) return;
  if (blockIdx.x == 0 && threadIdx.x == 0) w.ctl[0] = 0;  // the next traversal starts at slot 0 (the previous traversal has finished: stream order)
  // ---- the block's slots sorted by what their threads will execute (a counting sort in shared memory): warps shade one material
  //      instead of running every material's code with a handful of lanes each.  Which thread takes which slot changes nothing per path.
  const uint32_t colWords = (uint32_t)extraWords<T>() * kBlock;
  uint32_t* perm = smem + (kShadeBlock / kBlock) * colWords;
  uint32_t* cnt = perm + kShadeBlock;
  const uint32_t sBase = blockIdx.x * (uint32_t)kShadeBlock;
  if (threadIdx.x < 8) cnt[threadIdx.x] = 0;
  __syncthreads();
  uint32_t s = sBase + threadIdx.x;
  if (a.wfSort) {
    const uint32_t myBin = wfShadeBin<T>(sc, w, s);
    const uint32_t myRank = atomicAdd(&cnt[myBin], 1u);
    __syncthreads();
    uint32_t off = 0;
    for (uint32_t b = 0; b < myBin; ++b) off += cnt[b];
    perm[off + myRank] = threadIdx.x;
    __syncthreads();
    s = sBase + perm[threadIdx.x];
  }
  bool live = false;
  if (s < w.P) {
    uint32_t ph = w.phase[s];
    if (ph != WF_DEAD) {
      Instr<0> ins;
      Trav<T> tr;
      tr.col = smem + (threadIdx.x / kBlock) * colWords + (threadIdx.x % kBlock);
      tr.sp = 0; tr.tmin = 0; tr.inv = mk<T>(0, 0, 0); tr.cur = kEmpty;
      stD(&tr.xw(XW_ACC + 0), tr.stride, 0.0); stD(&tr.xw(XW_ACC + 2), tr.stride, 0.0); stD(&tr.xw(XW_ACC + 4), tr.stride, 0.0);
      PathState<T> ps;
      ps.phase = PH_IDLE; ps.time = 0; ps.depthLeft = 0; ps.beta = mk<T>(1, 1, 1); ps.rng = Rng{0, 1};
      bool start = false; V3<T> qo = mk<T>(0, 0, 0), qd = mk<T>(0, 0, 1);
      if (ph != WF_NEW) {
        // the finished item: weight of the light queries first (Lib.hs:1329-1333), then the bounce's hit
        ps.beta = scale(w.rW[s], mk<T>(w.bx[s], w.by[s], w.bz[s]));
        ps.time = w.time[s]; ps.depthLeft = (int)w.depth[s]; ps.rng = Rng{w.rngS[s], w.rngG[s]};
        const uint32_t fl = w.rFlags[s];
        if (fl & WF_R_NOMAIN) {
          accAdd<T>(tr, ps.beta * mk<T>(0, 0, 0));  // rayColor with depth <= 0 (Lib.hs:1303-1304) under the accumulated weight
        } else {
          tr.o = mk<T>(w.ox[s], w.oy[s], w.oz[s]); tr.d = mk<T>(w.dx[s], w.dy[s], w.dz[s]);
          tr.tbest = w.rT[s]; tr.bprim = w.rPrim[s]; tr.flags = fl & 0x0000ffffu;
          const uint32_t bd = (fl >> F_BDEPTH_SHIFT) & 7u;
          for (uint32_t i = 0; i < bd; ++i) tr.xu(X_BCHAIN + (int)i) = w.rChain[(size_t)i * w.P + s];
          ps.phase = PH_MAIN;
          start = shadeStep<T, 0, MATH_LITERAL>(sc, tr, ps, ins, qo, qd);
        }
        if (!start) {  // the path ended: its radiance goes to the slot's accumulator
          w.acc[s] += ldD(&tr.xw(XW_ACC + 0), tr.stride);
          w.acc[(size_t)w.P + s] += ldD(&tr.xw(XW_ACC + 2), tr.stride);
          w.acc[2 * (size_t)w.P + s] += ldD(&tr.xw(XW_ACC + 4), tr.stride);
        }
      }
      if (!start) {
        // ---- path regeneration: the next path index of the slot's chunk (consecutive indices = consecutive samples of one pixel)
        for (;;) {
          uint32_t left = w.chunkLeft[s];
          unsigned long long next = w.chunkNext[s];
          if (left == 0) {
            // guided: full chunks while the index space is large, shorter ones towards its end, so that the slots run dry together
            // (a 32-path chunk taken last is ~200 more rounds at falling occupancy; any split of the index space gives the same frame)
            const unsigned long long seen = *reinterpret_cast<volatile unsigned long long*>(w.ctl + 3);
            const float rem = seen < a.total_paths ? __ull2float_rn(a.total_paths - seen) : 0.0f;
            const uint32_t want = (uint32_t)fminf(fmaxf(rem * a.wfGuideInv, 1.0f), (float)a.wfChunk);
            next = atomicAdd(w.ctl + 3, (unsigned long long)want);
            if (next >= a.total_paths) { ph = WF_DEAD; break; }
            left = (next + want <= a.total_paths) ? want : (uint32_t)(a.total_paths - next);
          }
          w.chunkNext[s] = next + 1; w.chunkLeft[s] = left - 1;
          const uint32_t regionW = a.col1 - a.col0;
          const uint32_t pl = (uint32_t)(next / a.ns), smp = a.s0 + (uint32_t)(next % a.ns);
          const uint32_t row = a.row0 + (pl / regionW) * a.rowStep, col = a.col0 + pl % regionW;
          const uint32_t pixel = row * a.W + col;
          const uint32_t apx = w.accPix[s];
          if (pixel != apx) {
            if (apx != kNone) {
              atomicAdd(a.out + (size_t)apx * 3, w.acc[s]); atomicAdd(a.out + (size_t)apx * 3 + 1, w.acc[(size_t)w.P + s]);
              atomicAdd(a.out + (size_t)apx * 3 + 2, w.acc[2 * (size_t)w.P + s]);
            }
            w.accPix[s] = pixel; w.acc[s] = 0.0; w.acc[(size_t)w.P + s] = 0.0; w.acc[2 * (size_t)w.P + s] = 0.0;
          }
          ps.rng = streamGen(a.seed, pixel, smp);
          const T ru = rnd<T>(ps.rng), rv = rnd<T>(ps.rng);
          const T u = ((T)col + ru) / (T)a.W, v = ((T)(a.H - 1 - row) + rv) / (T)a.H;
          const V3<T> rd = scale(sc.cam[21], randomInUnitDisk<T>(ps.rng));
          const V3<T> cu = mk<T>(sc.cam[12], sc.cam[13], sc.cam[14]), cv = mk<T>(sc.cam[15], sc.cam[16], sc.cam[17]);
          const V3<T> offset = scale(rd.x, cu) + scale(rd.y, cv);
          const T r = rnd<T>(ps.rng);
          ps.time = sc.cam[22] + (sc.cam[23] - sc.cam[22]) * r;
          const V3<T> org = mk<T>(sc.cam[0], sc.cam[1], sc.cam[2]);
          qo = org + offset;
          qd = mk<T>(sc.cam[3], sc.cam[4], sc.cam[5]) + scale(u, mk<T>(sc.cam[6], sc.cam[7], sc.cam[8])) + scale(v, mk<T>(sc.cam[9], sc.cam[10], sc.cam[11])) - org - offset;
          ps.beta = mk<T>(1, 1, 1); ps.depthLeft = (int)a.maxDepth;
          ps.phase = PH_MAIN;
          if (ps.depthLeft > 0) { start = true; break; }  // rayColor with depth <= 0 is black (Lib.hs:1303-1304): take the next index
        }
      }
      if (start) {
        live = true;
        uint32_t nph = WF_MAIN;
        if (ps.phase == PH_LPDF) {  // shadeStep left the mixture-pdf inputs in the column; the depth step of Lib.hs:1321-1333 is taken here
          w.v2[s] = xld<T>(tr, X_LV2); w.spdf[s] = xld<T>(tr, X_LSPDF);
          --ps.depthLeft;
          nph = ps.depthLeft > 0 ? WF_LPDF : WF_LPDF_NOMAIN;
        }
        w.ox[s] = qo.x; w.oy[s] = qo.y; w.oz[s] = qo.z; w.dx[s] = qd.x; w.dy[s] = qd.y; w.dz[s] = qd.z; w.time[s] = ps.time;
        w.bx[s] = ps.beta.x; w.by[s] = ps.beta.y; w.bz[s] = ps.beta.z;
        w.rngS[s] = ps.rng.s; w.rngG[s] = ps.rng.g; w.depth[s] = (uint32_t)ps.depthLeft;
        w.phase[s] = nph;
      } else {  // WF_DEAD: no path index left for this slot — flush its accumulator
        const uint32_t apx = w.accPix[s];
        if (apx != kNone) {
          atomicAdd(a.out + (size_t)apx * 3, w.acc[s]); atomicAdd(a.out + (size_t)apx * 3 + 1, w.acc[(size_t)w.P + s]);
          atomicAdd(a.out + (size_t)apx * 3 + 2, w.acc[2 * (size_t)w.P + s]);
          w.accPix[s] = kNone;
        }
        w.phase[s] = WF_DEAD;
      }
    }
  }
  // live slots of this round, block-aggregated; the LAST block to finish closes the round: no live slot left -> done (the later
  // launches of the batch return at once), else the live count is reset for the next round
  const int nLive = __syncthreads_count(live);
  if (threadIdx.x == 0) {
    if (nLive) atomicAdd(w.ctl + 1, (unsigned long long)nLive);
    __threadfence();
    if (atomicAdd(w.ctl + 4, 1ull) == (unsigned long long)gridDim.x - 1ull) {
      __threadfence();
      if (atomicExch(w.ctl + 1, 0ull) == 0ull) w.ctl[2] = 1ull;
      w.ctl[4] = 0ull;
    }
  }
}

// first launch of a frame: every slot is new
static __global__ void wfInitKernel(const WfState w) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s < w.P) { w.phase[s] = WF_NEW; w.accPix[s] = kNone; w.chunkLeft[s] = 0; w.chunkNext[s] = 0; }
  if (s < kWfCtlWords) w.ctl[s] = 0;
}

}  // namespace rtb
