// rt_render_kernel.inl — K2 (renderKernel) and K1 (traceKernel).  Included by rt_kernels.cuh inside namespace rtb.
//
// renderKernel: persistent warps, one camera path per lane.  Every lane is a small state machine whose pending work falls in
// one of five classes: a BVH-node step, a sphere-leaf step, a rect/cuboid-leaf step, a rare traversal step, or "shade"
// (consume a finished query: build the hit record, scatter, start the next query, or end the path and take the next path
// index of the warp's chunk).  Each warp iteration ballots the classes and runs the class MOST lanes are waiting for; lanes of
// the other classes wait and accumulate, so divergent paths are regrouped by what they need to do next instead of executing
// every branch with a handful of lanes.  The node and sphere classes then stay in a tight inner loop (one ballot per step)
// for as long as enough lanes remain in the class.  Per lane the order of steps, and of RNG draws, is unchanged — results do
// not depend on the schedule.
//
// Register discipline: the hot state (ray, t-interval, cursor, stack pointer, throughput, RNG) is scalars in registers; all
// rarely touched state (frame chains, parked medium query, light-pdf walk, per-pixel accumulator) lives in the thread's
// shared-memory column (rt_traverse.cuh) so the hot loop neither spills nor drags cold registers along.

enum : int { PH_IDLE = 0, PH_MAIN = 1, PH_LPDF = 2 };

template <class T> struct PathState {
  V3<T> beta; T time; int depthLeft; int phase; Rng rng;
};

RTB_D unsigned long long globalTimerNs() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
RTB_D double ldD(const uint32_t* p, uint32_t stride) { return __longlong_as_double((long long)((unsigned long long)p[0] | ((unsigned long long)p[stride] << 32))); }
RTB_D void stD(uint32_t* p, uint32_t stride, double v) { unsigned long long b = (unsigned long long)__double_as_longlong(v); p[0] = (uint32_t)b; p[stride] = (uint32_t)(b >> 32); }

// add a finished path's radiance to the lane's accumulator (current pixel)
template <class T> RTB_D void accAdd(Trav<T>& tr, const V3<T>& L) {
  stD(&tr.xw(XW_ACC + 0), tr.stride, ldD(&tr.xw(XW_ACC + 0), tr.stride) + (double)L.x);
  stD(&tr.xw(XW_ACC + 2), tr.stride, ldD(&tr.xw(XW_ACC + 2), tr.stride) + (double)L.y);
  stD(&tr.xw(XW_ACC + 4), tr.stride, ldD(&tr.xw(XW_ACC + 4), tr.stride) + (double)L.z);
}
template <class T> RTB_D void accFlush(Trav<T>& tr, double* out) {
  const uint32_t pix = tr.xw(XW_ACCPIX);
  if (pix != kNone) {
    atomicAdd(out + (size_t)pix * 3, ldD(&tr.xw(XW_ACC + 0), tr.stride));
    atomicAdd(out + (size_t)pix * 3 + 1, ldD(&tr.xw(XW_ACC + 2), tr.stride));
    atomicAdd(out + (size_t)pix * 3 + 2, ldD(&tr.xw(XW_ACC + 4), tr.stride));
  }
}

// consume the finished query of this lane.  Returns true when a new query must be started on ray (qo, qd); its root is
// world_root in phase PH_MAIN and the light in slot X_CURLIGHT in phase PH_LPDF.
template <class T, int INSTR, int MODE>
__device__ __forceinline__ bool shadeStep(const DScene<T>& sc, Trav<T>& tr, PathState<T>& ps, Instr<INSTR>& ins, V3<T>& qo, V3<T>& qd) {
  bool scattered = false;  // the path continues with (qo, qd) as the next bounce ray
  qo = tr.o; qd = tr.d;    // all frames are closed when a query ends: this is the query's world ray
  {
    if (!tr.qhit()) { ins.bounce(kNone, 0.0); accAdd<T>(tr, ps.beta * mk<T>(sc.bg[0], sc.bg[1], sc.bg[2])); ps.phase = PH_IDLE; return false; }  // Lib.hs:1310-1313
    ins.bounce(tr.bprim, (double)tr.tbest);
    HitRec<T> rec;
    evalHit<T, MODE>(sc, tr.o, tr.d, ps.time, tr.tbest, tr.bprim, tr.baux(), tr.bdepth(), &tr.xu(X_BCHAIN), tr.stride * (uint32_t)(sizeof(T) / 4), rec);
    const Q4<T> m = ldq(sc.mats + rec.mat);
    const uint32_t mkind = bitsOf(m.x), tex = bitsOf(m.y);
    // the material's texture, evaluated at ONE call site (textureValue draws no random numbers, so its position relative to the
    // samplers below is free): not needed by Dielectric (white) and by a light seen from its front face (black)
    V3<T> tv = mk<T>(0, 0, 0);
    if (mkind != M_DIELECTRIC && !(mkind == M_LIGHT && rec.front)) tv = textureValue<T, INSTR, MODE>(sc, tex, rec.u, rec.v, rec.p, ins);
    if (mkind == M_LIGHT) {  // scatter = Nothing -> emitted, back face only (Lib.hs:860, 880-885)
      ins.count(20);
      accAdd<T>(tr, ps.beta * tv); ps.phase = PH_IDLE;
      return false;
    } else if (mkind == M_LAMBERTIAN) {  // Lib.hs:823-836 with pdfGenerate/pdfValue of the mixture (367-382)
      ins.count(16);
      const Onb<T> uvw = onbFromW(rec.n);
      const T rd = rnd<T>(ps.rng);
      V3<T> pdfD;
      if (rd < (T)0.5) { ins.count(21); pdfD = htblRandom<T, MODE>(sc, sc.lights_root, rec.p, ps.rng); }
      else pdfD = onbLocal(uvw, randomCosineDirection<T, MODE>(ps.rng));
      const V3<T> dir = unit(pdfD);
      const T cosine = dot(unit(dir), uvw.w);
      const T v2 = (cosine <= 0) ? (T)0 : cosine / kPi<T>();
      const T c2 = dot(rec.n, dir);  // scatteringPdf (Lib.hs:874-878)
      const T spdf = (c2 < 0) ? (T)0 : c2 / kPi<T>();
      ps.beta = ps.beta * tv;
      qo = rec.p; qd = dir;
      if ((sc.lights_root >> 28) == K_UNHITTABLE) {  // htblPdfValue Unhittable = 0 without any traversal (Lib.hs:677-680, 1109)
        ins.count(13); ins.visit(sc.lights_root);
        const T pdfVal = (T)0.5 * ((T)0 + v2);
        ps.beta = scale(spdf / pdfVal, ps.beta);
        scattered = true;
      } else {
        ps.phase = PH_LPDF;
        xst<T>(tr, X_LV2, v2); xst<T>(tr, X_LSPDF, spdf); xst<T>(tr, X_LV1, (T)0); xst<T>(tr, X_CURW, (T)1);
        tr.xu(X_LSP) = 0; tr.xu(X_CURLIGHT) = sc.lights_root;
        return true;
      }
    } else if (mkind == M_METAL) {  // Lib.hs:837-841: never absorbs, fuzz unclamped
      ins.count(17);
      const V3<T> rUnit = randomUnitVector<T, MODE>(ps.rng);
      const V3<T> reflected = reflectV(unit(tr.d), rec.n);
      ps.beta = ps.beta * tv;
      qo = rec.p; qd = reflected + scale(m.z, rUnit);
      scattered = true;
    } else if (mkind == M_DIELECTRIC) {  // Lib.hs:842-859: xi is always drawn
      ins.count(18);
      const T eta = rec.front ? (T)1.0 / m.z : m.z;
      const V3<T> ud = unit(tr.d);
      const T cosTheta = hmin(dot(neg(ud), rec.n), (T)1.0);
      const T sinTheta = msqrt((T)1.0 - cosTheta * cosTheta);
      const T rd = rnd<T>(ps.rng);
      if ((eta * sinTheta > (T)1.0) || rd < schlick<T>(cosTheta, eta)) qd = reflectV(ud, rec.n);
      else qd = refractV(ud, rec.n, eta);
      qo = rec.p;
      scattered = true;
    } else {  // Isotropic (Lib.hs:861-865) weighted by the generic scatteringPdf / 1.0 (867-878, 1329-1333)
      ins.count(19);
      const V3<T> randDr = randomInUnitSphere<T>(ps.rng);
      const T c2 = dot(rec.n, randDr);
      const T spdf = (c2 < 0) ? (T)0 : c2 / kPi<T>();
      ps.beta = scale(spdf / (T)1.0, ps.beta * tv);
      qo = rec.p; qd = randDr;
      scattered = true;
    }
  }
  if (scattered) {  // rayColorHelp newRay (d - 1) (Lib.hs:1321-1333)
    --ps.depthLeft;
    if (ps.depthLeft <= 0) { accAdd<T>(tr, ps.beta * mk<T>(0, 0, 0)); ps.phase = PH_IDLE; return false; }
    return true;
  }
  return false;
}

// A finished query of the Lambertian mixture pdf's light term, htblPdfValue (Lib.hs:673-705): account the light in X_CURLIGHT
// (walked depth-first, left first), then start the query of the next light on the same ray, or — all lights done — weight the
// throughput (Lib.hs:1329-1333) and start the bounce's `hit world`.  Runs right after the traversal step that finished the query
// (not through the shade class): the lane stays in the traversal classes.
template <class T, int INSTR, int MODE>
RTB_D void lpdfAdvance(const DScene<T>& sc, Trav<T>& tr, PathState<T>& ps, Instr<INSTR>& ins, uint32_t lbase, uint32_t lightCap) {
  const uint32_t curLight = tr.xu(X_CURLIGHT);
  const T curW = xld<T>(tr, X_CURW);
  uint32_t lsp = tr.xu(X_LSP);
  if (tr.qhit()) {
    const uint32_t lk = curLight >> 28, li = curLight & kIdxMask;
    if (lk == K_RECT_XZ) {
      ins.count(22);
      const Q4<T> ra = ldq(sc.rects + 2 * (size_t)li);
      const T area = (ra.y - ra.x) * (ra.w - ra.z);
      const T distanceSquared = tr.tbest * tr.tbest * sqlen(tr.d);
      const T cosine = mabs(tr.d.y / len(tr.d));  // |v . (0,+-1,0)| / |v|
      xst<T>(tr, X_LV1, xld<T>(tr, X_LV1) + curW * (distanceSquared / (cosine * area)));
    } else if (lk == K_SPHERE) {
      ins.count(22);
      const Q4<T> q = ldq(sc.spheres + li);
      const T cosThetaMax = msqrt(1 - q.w * q.w / sqlen(mk<T>(q.x, q.y, q.z) - tr.o));
      const T solidAngle = 2 * kPi<T>() * (1 - cosThetaMax);
      xst<T>(tr, X_LV1, xld<T>(tr, X_LV1) + curW * (1 / solidAngle));
    } else if (lk == K_BVH) {
      const uint4 sz = __ldg(sc.node_sz + li);
      const uint32_t lc = bitsOf(ldq(sc.nodes + 2 * (size_t)li).w) & ~kPairBit, rc = bitsOf(ldq(sc.nodes + 2 * (size_t)li + 1).w);
      if (lsp + 2 <= lightCap) {
        uint32_t* e = lightEntry<T>(tr, lbase, lsp);
        e[0] = rc; stT(e + tr.stride, tr.stride, curW * ((T)sz.z / (T)sz.x)); ++lsp;
        e = lightEntry<T>(tr, lbase, lsp);
        e[0] = lc; stT(e + tr.stride, tr.stride, curW * ((T)sz.y / (T)sz.x)); ++lsp;
      } else {
        __trap();  // cannot happen: rt_scene_create sizes lightCap from the depth of the lights BVH (a dropped child would bias the pdf silently)
      }
    }
  }
  const V3<T> qo = tr.o, qd = tr.d;  // all frames are closed: the query's world ray = the scattered ray
  uint32_t root;
  if (lsp > 0) {
    --lsp;
    const uint32_t* e = lightEntry<T>(tr, lbase, lsp);
    root = e[0];
    T wt; ldT(e + tr.stride, tr.stride, wt);
    tr.xu(X_CURLIGHT) = root; xst<T>(tr, X_CURW, wt); tr.xu(X_LSP) = lsp;
  } else {
    const T pdfVal = (T)0.5 * (xld<T>(tr, X_LV1) + xld<T>(tr, X_LV2));
    ps.beta = scale(xld<T>(tr, X_LSPDF) / pdfVal, ps.beta);
    ps.phase = PH_MAIN;
    --ps.depthLeft;  // rayColorHelp newRay (d - 1) (Lib.hs:1321-1333)
    if (ps.depthLeft <= 0) { accAdd<T>(tr, ps.beta * mk<T>(0, 0, 0)); ps.phase = PH_IDLE; return; }  // cur stays kEmpty: regenerated by the shade class
    root = sc.world_root;
    ins.count(1);
  }
  ins.count(13);
  travBegin<T, MODE>(sc, tr, root, qo, qd, kEps<T>(), kInf<T>());
}

template <class T, class S, int INSTR, int MODE>
__global__ void __launch_bounds__(kBlock, MinBlocks<T>::v) renderKernel(const __grid_constant__ DScene<T> sc, const __grid_constant__ RenderArgs a) {
  extern __shared__ uint32_t smem[];
  const unsigned lane = threadIdx.x & 31u;
  const unsigned ltMask = (1u << lane) - 1u;
  Instr<INSTR> ins;
  if constexpr (INSTR == 1) ins.init();
  if constexpr (INSTR == 3) ins.init(a.plogPrim, a.plogT, a.plogStride);

  Trav<T> tr;
  tr.col = smem + threadIdx.x;
  tr.cur = kEmpty; tr.sp = 0; tr.flags = 0; tr.bprim = 0; tr.tmin = 0; tr.tbest = 0;
  tr.o = mk<T>(0, 0, 0); tr.d = mk<T>(0, 0, 1); tr.inv = mk<T>(0, 0, 0);
  PathState<T> ps;
  ps.phase = PH_IDLE; ps.beta = mk<T>(1, 1, 1); ps.time = 0; ps.depthLeft = 0; ps.rng = Rng{0, 1};
  tr.xw(XW_ACCPIX) = kNone;
  // the warp's chunk of the path index space: {first pixel of the region, first sample, next offset, end offset}
  volatile uint32_t* ws = smem + (size_t)columnWords<T>(a.stackCap, a.lightCap) * blockDim.x + (threadIdx.x >> 5) * 4;
  if (lane == 0) { ws[0] = 0; ws[1] = 0; ws[2] = 0; ws[3] = 0; }
  __syncwarp();
  bool exhausted = false;
  unsigned long long sch[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // INSTR == 1 only
  const uint32_t regionW = a.col1 - a.col0;

  for (;;) {
    // ---- fast votes: a node step when at least keepNode lanes wait for one, else a sphere step when keepLeaf lanes do;
    //      otherwise the full vote: the class most lanes wait for, but rare steps as soon as rareMin lanes wait for one, rect /
    //      cuboid steps from rectMin and shade from shadeMin (lanes parked in a minority class are lanes the majority classes
    //      run without; the thresholds are tuned per scene family)
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

    if constexpr (INSTR == 1) {  // scheduler statistics (RtStats::sched)
      const int sc_ = cls;
      const int n0 = __popc(__ballot_sync(0xffffffffu, sc_ == 0)), n1 = __popc(__ballot_sync(0xffffffffu, sc_ == 1)), n2 = __popc(__ballot_sync(0xffffffffu, sc_ == 2)),
                n3 = __popc(__ballot_sync(0xffffffffu, sc_ == 3)), n4 = __popc(__ballot_sync(0xffffffffu, sc_ == 4));
      if (lane == 0) {
        const int served = pick == 0 ? n0 : pick == 1 ? n1 : pick == 2 ? n2 : pick == 3 ? n3 : n4;
        ++sch[pick]; sch[5 + pick] += (unsigned)served; sch[10] += n0; sch[11] += n1; sch[12] += n2; sch[13] += n3; sch[14] += n4; ++sch[15];
      }
    }
    if (pick == 0) {
      if (isNode) stepNode<T, INSTR, MODE>(sc, tr, ins);
    } else if (pick == 1) {
      if (isLeaf) stepSphere<T, S, INSTR>(sc, tr, (ps.phase == PH_MAIN) ? ps.time : (T)0, ins);  // htblPdfValue traces with time 0 (Lib.hs:677)
    } else if (pick == 2) {
      if (cls == 2) stepRect<T, INSTR, MODE>(sc, tr, ins);
    } else if (pick == 3) {
      if (cls == 3) stepRare<T, INSTR, MODE>(sc, tr, ps.rng, ins);
    }
    if (pick != 4) {
      if (ps.phase == PH_LPDF && tr.cur == kEmpty) lpdfAdvance<T, INSTR, MODE>(sc, tr, ps, ins, a.stackCap + (uint32_t)extraWords<T>(), a.lightCap);
    } else {
      // ---- shade class: (1) consume finished queries
      bool start = false; V3<T> qo = tr.o, qd = tr.d;
      if (cls == 4 && ps.phase != PH_IDLE) start = shadeStep<T, INSTR, MODE>(sc, tr, ps, ins, qo, qd);
      // ---- (2) path regeneration: idle lanes take the next path indices of the warp's chunk, in lane order
      unsigned idle = __ballot_sync(0xffffffffu, cls == 4 && ps.phase == PH_IDLE);
      while (idle) {
        uint32_t cNext = ws[2], cEnd = ws[3];
        if (cNext >= cEnd) {
          if (exhausted) break;
          // guided chunks: a.chunk path indices while plenty are left, fewer towards the end of the index space (what is left divided
          // among the resident warps, never below a.chunkMin) so that the kernel's tail is the time of a SMALL chunk, not of a full one
          unsigned long long base = 0; uint32_t csize = a.chunk;
          if (lane == 0) {
            const unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(a.counter);
            const unsigned long long left = cur < a.total_paths ? a.total_paths - cur : 0ull;
            const unsigned long long want = left / a.chunkDiv;
            csize = (uint32_t)(want < a.chunkMin ? a.chunkMin : (want > a.chunk ? a.chunk : want));
            base = atomicAdd(a.counter, (unsigned long long)csize);
          }
          base = __shfl_sync(0xffffffffu, base, 0); csize = __shfl_sync(0xffffffffu, csize, 0);
          if (base >= a.total_paths) {  // first warp to find the index space empty stamps the start of the kernel's tail (RtStats::tail_ms)
            if (lane == 0) atomicCAS(a.counter + kCtrTailBase, 0ull, globalTimerNs());
            exhausted = true; break;
          }
          cNext = 0;
          cEnd = (base + csize < a.total_paths) ? csize : (uint32_t)(a.total_paths - base);
          if (lane == 0) { ws[0] = (uint32_t)(base / a.ns); ws[1] = (uint32_t)(base % a.ns); ws[3] = cEnd; }
          __syncwarp();
        }
        const unsigned rank = __popc(idle & ltMask);
        if (cls == 4 && ps.phase == PH_IDLE && cNext + rank < cEnd) {
          // path index -> (pixel of the region, sample); consecutive indices = consecutive samples of one pixel
          const uint32_t off = ws[1] + cNext + rank;
          const uint32_t pl = ws[0] + off / a.ns, s = a.s0 + off % a.ns;
          const uint32_t row = a.row0 + (pl / regionW) * a.rowStep, col = a.col0 + pl % regionW;
          const uint32_t pixel = row * a.W + col;
          if (pixel != tr.xw(XW_ACCPIX)) {
            accFlush<T>(tr, a.out);
            tr.xw(XW_ACCPIX) = pixel;
            stD(&tr.xw(XW_ACC + 0), tr.stride, 0.0); stD(&tr.xw(XW_ACC + 2), tr.stride, 0.0); stD(&tr.xw(XW_ACC + 4), tr.stride, 0.0);
          }
          ps.rng = streamGen(a.seed, pixel, s);
          ins.count(0);
          ins.beginPath((unsigned long long)pl * a.ns + off % a.ns);
          // uniformRandomUVs (Lib.hs:1362-1365): u = (x + ru)/W, v = (y + rv)/H with y counted from the bottom row
          const T ru = rnd<T>(ps.rng), rv = rnd<T>(ps.rng);
          const T u = ((T)col + ru) / (T)a.W, v = ((T)(a.H - 1 - row) + rv) / (T)a.H;
          // getRay (Lib.hs:1253-1267): lens sample and shutter time are drawn for every ray
          const V3<T> rd = scale(sc.cam[21], randomInUnitDisk<T>(ps.rng));
          const V3<T> cu = mk<T>(sc.cam[12], sc.cam[13], sc.cam[14]), cv = mk<T>(sc.cam[15], sc.cam[16], sc.cam[17]);
          const V3<T> offset = scale(rd.x, cu) + scale(rd.y, cv);
          const T r = rnd<T>(ps.rng);
          ps.time = sc.cam[22] + (sc.cam[23] - sc.cam[22]) * r;
          const V3<T> org = mk<T>(sc.cam[0], sc.cam[1], sc.cam[2]);
          qo = org + offset;
          qd = mk<T>(sc.cam[3], sc.cam[4], sc.cam[5]) + scale(u, mk<T>(sc.cam[6], sc.cam[7], sc.cam[8])) + scale(v, mk<T>(sc.cam[9], sc.cam[10], sc.cam[11])) - org - offset;
          ps.beta = mk<T>(1, 1, 1); ps.depthLeft = (int)a.maxDepth;
          ps.phase = PH_MAIN; start = true;
          if (ps.depthLeft <= 0) { ps.phase = PH_IDLE; start = false; }  // rayColor with depth <= 0 is black (Lib.hs:1303-1304)
        }
        const uint32_t avail = cEnd - cNext; const uint32_t n = (uint32_t)__popc(idle);
        __syncwarp();
        if (lane == 0) ws[2] = cNext + ((n < avail) ? n : avail);
        __syncwarp();
        idle = __ballot_sync(0xffffffffu, cls == 4 && ps.phase == PH_IDLE);
      }
      // ---- (3) start the next query: the bounce's `hit world` (Lib.hs:1309) or one `hit htbl (Ray o v 0)` of htblPdfValue (677)
      if (start) {
        ins.count(13);
        if (ps.phase == PH_MAIN) ins.count(1);
        travBegin<T, MODE>(sc, tr, ps.phase == PH_MAIN ? sc.world_root : tr.xu(X_CURLIGHT), qo, qd, kEps<T>(), kInf<T>());
      } else if (cls == 4) {
        tr.cur = kDead;  // only reached when the index space is exhausted
      }
    }
  }
  accFlush<T>(tr, a.out);
  if (lane == 0) atomicMax(a.counter + kCtrTailBase + 1, globalTimerNs());  // the last warp to leave ends the tail
  if constexpr (INSTR == 1) {
    for (int i = 0; i < kWorkCounters; ++i) if (ins.c[i]) atomicAdd(a.ctr + i, (unsigned long long)ins.c[i]);
    if (lane == 0) for (int i = 0; i < kSchedCounters; ++i) if (sch[i]) atomicAdd(a.ctr + kWorkCounters + i, sch[i]);
  }
}

// K1: one ray per thread, the same step functions run to completion, with the visit log
template <class T, class S, int MODE>
__global__ void __launch_bounds__(kBlock) traceKernel(const __grid_constant__ DScene<T> sc, const __grid_constant__ TraceArgs a) {
  extern __shared__ uint32_t smem[];
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  const double* r = a.rays + (size_t)i * 7;
  const V3<T> o = mk<T>((T)r[0], (T)r[1], (T)r[2]), d = mk<T>((T)r[3], (T)r[4], (T)r[5]);
  const T time = (T)r[6];
  Rng rng = streamGen(a.seed, i, 0);
  Instr<2> ins; ins.init(a.vlog + (size_t)i * a.logStride, a.logStride);
  for (uint32_t k = 0; k < a.logStride; ++k) a.vlog[(size_t)i * a.logStride + k] = kNone;
  Trav<T> tr;
  tr.col = smem + threadIdx.x;
  travBegin<T, MODE>(sc, tr, sc.world_root, o, d, (T)a.tmin, (T)a.tmax);
  travRun<T, S, 2, MODE>(sc, tr, time, rng, ins);
  a.nvisit[i] = ins.n; a.vhash[i] = ins.h;
  a.hit[i] = tr.qhit() ? 1 : 0;
  double* out = a.rec + (size_t)i * 9;
  if (tr.qhit()) {
    HitRec<T> rec;
    evalHit<T, MODE>(sc, o, d, time, tr.tbest, tr.bprim, tr.baux(), tr.bdepth(), &tr.xu(X_BCHAIN), tr.stride * (uint32_t)(sizeof(T) / 4), rec);
    out[0] = rec.t; out[1] = rec.p.x; out[2] = rec.p.y; out[3] = rec.p.z; out[4] = rec.n.x; out[5] = rec.n.y; out[6] = rec.n.z; out[7] = rec.u; out[8] = rec.v;
    a.prim[i] = tr.bprim; a.mat[i] = rec.mat; a.front[i] = rec.front ? 1 : 0;
  } else {
    for (int k = 0; k < 9; ++k) out[k] = 0;
    a.prim[i] = 0; a.mat[i] = kNone; a.front[i] = 0;
  }
}
