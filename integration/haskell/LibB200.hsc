{-# LANGUAGE ForeignFunctionInterface #-}
-- | Drop-in body for `Lib.runRender` (reference src/Lib.hs:1491-1523) on top of librt_b200.so.
--
-- NOT COMPILED in the build image (no ghc / stack / hsc2hs there or on the GPU box); it is the binding a maintainer adds, kept
-- next to include/rt_abi.h so that the two change together.  The same ABI is exercised by the C++ mirror of Lib / Scenes
-- (ray-tracing_b200/host/, whose Arena::flatten is this file's flattenScene) and by Python ctypes.
--
-- How to wire it in (see INTEGRATION.md):
--   * package.yaml, library: `extra-libraries: rt_b200`, `extra-lib-dirs`, `include-dirs: <repo>/include`,
--     `build-tools: hsc2hs`, dependencies += vector (already there) ;
--   * src/Lib.hs export list (Lib.hs:4-51) already has Albedo(..), Axis(..), Box(..), Fuzz(..), Hittable(..), Image(..),
--     Material(..), RefractiveIdx(..), Texture(..); it additionally exports what this module pattern-matches on or calls:
--     Rectangle(..), Camera(..), Vec3(..), RenderStaticEnv(..), RGB, divide, albedoToColor
--     (or this file becomes Lib.Internal's neighbour and Lib re-exports runRenderB200 as runRender);
--   * src/Random.hs exports `randomWord64 (RandGen g) = let (w, g') = Random.genWord64 g in (w, RandGen g')`;
--   * app/Main.hs imports `runRenderB200 as runRender` -- nothing else changes: printRow forces the rows as before.
module LibB200
  ( runRenderB200
  , b200DeviceCount
  ) where

#include "rt_abi.h"

import qualified Codec.Picture        as JP (Image (..))
import           Control.Exception    (bracket, evaluate, finally)
import           Control.Monad        (forM, forM_, when)
import           Data.IORef
import qualified Data.IntMap.Strict   as IM
import           Data.Int             (Int32)
import qualified Data.Vector          as VV
import qualified Data.Vector.Storable as SV
import qualified Data.Vector.Unboxed  as V
import           Data.Word            (Word32, Word64, Word8)
import           Foreign
import           Foreign.C.String     (CString, peekCString)
import           Foreign.C.Types
import           Lib                  hiding (length) -- Lib exports its own Vec3 `length` (Lib.hs:13)
import           Random               (RandGen, randomWord64)
import           System.IO.Unsafe     (unsafePerformIO)
import           System.Mem.StableName (StableName, eqStableName, hashStableName, makeStableName)

-- ---------------------------------------------------------------------------------------------------------------------------
-- foreign imports: one per entry point of include/rt_abi.h that a renderer needs
-- ---------------------------------------------------------------------------------------------------------------------------
data RtSceneDesc
data RtRenderParams
data RtStats
type RtScene = Word64

foreign import ccall unsafe "rt_abi_version"   c_rt_abi_version   :: IO CInt
foreign import ccall unsafe "rt_last_error"    c_rt_last_error    :: IO CString
foreign import ccall unsafe "rt_device_count"  c_rt_device_count  :: IO CInt
foreign import ccall safe   "rt_init"          c_rt_init          :: CInt -> Word32 -> IO CInt
-- `safe`: these block for milliseconds .. minutes and must not hold an RTS capability
foreign import ccall safe   "rt_scene_create"  c_rt_scene_create  :: Ptr RtSceneDesc -> CInt -> Ptr RtScene -> IO CInt
foreign import ccall safe   "rt_scene_destroy" c_rt_scene_destroy :: RtScene -> IO CInt
foreign import ccall safe   "rt_render"        c_rt_render        :: RtScene -> Ptr RtRenderParams -> Ptr Double -> Ptr RtStats -> IO CInt

b200DeviceCount :: IO Int
b200DeviceCount = fromIntegral <$> c_rt_device_count

check :: CInt -> IO ()
check rc = when (rc /= 0) $ do
  msg <- c_rt_last_error >>= peekCString
  error ("rt_b200: " ++ msg) -- the reference reports impossible cases with `error` too (Lib.hs:462-464, 868-873, 927)

-- ---------------------------------------------------------------------------------------------------------------------------
-- flat tables: one growable list per array of RtSceneDesc; `push` returns the index of the new row
-- ---------------------------------------------------------------------------------------------------------------------------
newtype Tbl a = Tbl (IORef (Word32, [a]))

newTbl :: IO (Tbl a)
newTbl = Tbl <$> newIORef (0, [])

push :: Tbl a -> a -> IO Word32
push (Tbl r) x = do
  (n, xs) <- readIORef r
  writeIORef r (n + 1, x : xs)
  return n

rowsOf :: Tbl a -> IO [a]
rowsOf (Tbl r) = reverse . snd <$> readIORef r

data Flat = Flat
  { fNodes      :: Tbl (Box, Word32, Word32, Int)               -- box, left ref, right ref, bvh_size
  , fSpheres    :: Tbl (Vec3, Double, Word32)                    -- centre, radius, material
  , fMoving     :: Tbl (Vec3, Vec3, Double, Double, Double, Word32) -- c0 c1 t0 duration r material
  , fRects      :: Tbl ((Double, Double, Double, Double, Double), Word32)
  , fCuboids    :: Tbl (Vec3, Vec3, Word32)
  , fTranslates :: Tbl (Vec3, Word32)
  , fRotates    :: Tbl (Word32, Double, Double, Word32)          -- axis, sin, cos, child
  , fMedia      :: Tbl (Double, Word32, Word32)                  -- negInvDensity, material, boundary
  , fMats       :: Tbl (Word32, Word32, Double)                  -- kind, texture | RT_NONE, parameter
  , fTexs       :: Tbl (Word32, Vec3, Word32, Word32, Double)    -- kind, rgb, a, b, scale
  , fPerlin     :: Tbl (VV.Vector Vec3, V.Vector Int, V.Vector Int, V.Vector Int)
  , fImages     :: Tbl (SV.Vector Word8, Int, Int)
  , fMemoT      :: Memo Texture   -- a Texture / Material / image shared by many objects (one earth map, one glass) is pushed ONCE
  , fMemoM      :: Memo Material
  , fMemoI      :: Memo Image
  }

-- Sharing by heap identity: StableName of the evaluated constructor, bucketed by its hash.  Two structurally equal but separately built
-- values are simply pushed twice (same semantics, a few more rows).
type Memo a = IORef (IM.IntMap [(StableName a, Word32)])

memoised :: Memo a -> a -> IO Word32 -> IO Word32
memoised memo x build = do
  x' <- evaluate x
  sn <- makeStableName x'
  seen <- readIORef memo
  case [i | (s, i) <- IM.findWithDefault [] (hashStableName sn) seen, eqStableName s sn] of
    (i : _) -> return i
    []      -> do
      i <- build
      modifyIORef' memo (IM.insertWith (++) (hashStableName sn) [(sn, i)])
      return i

newFlat :: IO Flat
newFlat = Flat <$> newTbl <*> newTbl <*> newTbl <*> newTbl <*> newTbl <*> newTbl <*> newTbl <*> newTbl <*> newTbl <*> newTbl
               <*> newTbl <*> newTbl <*> newIORef IM.empty <*> newIORef IM.empty <*> newIORef IM.empty

-- ref = kind << 28 | index (include/rt_abi.h RT_REF); the kinds are enum RtHittableKind
ref :: Word32 -> Word32 -> Word32
ref kind i = (kind `shiftL` 28) .|. i

rtNone :: Word32
rtNone = #{const RT_NONE}

-- ---------------------------------------------------------------------------------------------------------------------------
-- the flattener: a depth-first walk of Hittable (Lib.hs:521-585), Material (339-345) and Texture (394-419)
-- ---------------------------------------------------------------------------------------------------------------------------
flattenT :: Flat -> Texture -> IO Word32
flattenT f t = memoised (fMemoT f) t (flattenT' f t)

flattenT' :: Flat -> Texture -> IO Word32
flattenT' f (ConstantColor (Albedo rgb)) = push (fTexs f) (#{const RT_TEX_CONSTANT}, rgb, 0, 0, 0)
flattenT' f (CheckerTexture odd' even') = do
  o <- flattenT f odd'
  e <- flattenT f even'
  push (fTexs f) (#{const RT_TEX_CHECKER}, Vec3 0 0 0, o, e, 0)
flattenT' f (Perlin ranvec px py pz sc) = do
  t <- push (fPerlin f) (ranvec, px, py, pz)
  push (fTexs f) (#{const RT_TEX_PERLIN}, Vec3 0 0 0, t, 0, sc)
flattenT' f (ImageTexture Nothing _ _) = push (fTexs f) (#{const RT_TEX_IMAGE}, Vec3 0 0 0, rtNone, 0, 0) -- renders (0,1,1), Lib.hs:504-505
flattenT' f (ImageTexture (Just img@(Image im)) w h) = do
  i <- memoised (fMemoI f) img (push (fImages f) (JP.imageData im, w, h)) -- tightly packed RGB8, top row first (JuicyPixels)
  push (fTexs f) (#{const RT_TEX_IMAGE}, Vec3 0 0 0, i, 0, 0)

flattenM :: Flat -> Material -> IO Word32
flattenM f m = memoised (fMemoM f) m (flattenM' f m)

flattenM' :: Flat -> Material -> IO Word32
flattenM' f (Lambertian t)                   = flattenT f t >>= \ti -> push (fMats f) (#{const RT_LAMBERTIAN}, ti, 0)
flattenM' f (Metal t (Fuzz fz))              = flattenT f t >>= \ti -> push (fMats f) (#{const RT_METAL}, ti, fz)
flattenM' f (Dielectric (RefractiveIdx ior)) = push (fMats f) (#{const RT_DIELECTRIC}, rtNone, ior)
flattenM' f (DiffuseLight t)                 = flattenT f t >>= \ti -> push (fMats f) (#{const RT_DIFFUSE_LIGHT}, ti, 0)
flattenM' f (Isotropic t)                    = flattenT f t >>= \ti -> push (fMats f) (#{const RT_ISOTROPIC}, ti, 0)

flattenR :: Flat -> Rectangle -> IO Word32
flattenR f (XYRect a0 a1 b0 b1 k m) = flattenM f m >>= \mi -> ref #{const RT_RECT_XY} <$> push (fRects f) ((a0, a1, b0, b1, k), mi)
flattenR f (XZRect a0 a1 b0 b1 k m) = flattenM f m >>= \mi -> ref #{const RT_RECT_XZ} <$> push (fRects f) ((a0, a1, b0, b1, k), mi)
flattenR f (YZRect a0 a1 b0 b1 k m) = flattenM f m >>= \mi -> ref #{const RT_RECT_YZ} <$> push (fRects f) ((a0, a1, b0, b1, k), mi)

rectMaterial :: Rectangle -> Material
rectMaterial (XYRect _ _ _ _ _ m) = m
rectMaterial (XZRect _ _ _ _ _ m) = m
rectMaterial (YZRect _ _ _ _ _ m) = m

axisCode :: Axis -> Word32
axisCode XAxis = #{const RT_AXIS_X}
axisCode YAxis = #{const RT_AXIS_Y}
axisCode ZAxis = #{const RT_AXIS_Z}

-- Children first, then the parent row: no patching, and the numbering is a post-order (the library accepts any numbering;
-- shared sub-objects such as the light that sits in both the world and the light list are simply flattened twice).
flattenH :: Flat -> Hittable -> IO Word32
flattenH _ Unhittable = return (ref #{const RT_UNHITTABLE} 0)
flattenH f (Sphere c r m) = flattenM f m >>= \mi -> ref #{const RT_SPHERE} <$> push (fSpheres f) (c, r, mi)
flattenH f (MovingSphere c0 c1 t0 _t1 dur r m) =
  flattenM f m >>= \mi -> ref #{const RT_MOVING_SPHERE} <$> push (fMoving f) (c0, c1, t0, dur, r, mi)
flattenH f (Rect r) = flattenR f r
flattenH f (Cuboid mn mx sides) = case sides of
  -- the six sides are always the ones `cuboid` builds from min / max / material (Lib.hs:594-605): min, max, material suffice
  (s : _) -> flattenM f (rectMaterial s) >>= \mi -> ref #{const RT_CUBOID} <$> push (fCuboids f) (mn, mx, mi)
  []      -> error "rt_b200: Cuboid without sides"
flattenH f (BVHNode l r box size) = do
  li <- flattenH f l
  ri <- flattenH f r -- a 1-element makeBVH stores the same child twice (Lib.hs:948): it is flattened twice, same semantics
  ref #{const RT_BVH_NODE} <$> push (fNodes f) (box, li, ri, size)
flattenH f (Translate off h) = flattenH f h >>= \c -> ref #{const RT_TRANSLATE} <$> push (fTranslates f) (off, c)
flattenH f (Rotate ax s c _box h) = flattenH f h >>= \ch -> ref #{const RT_ROTATE} <$> push (fRotates f) (axisCode ax, s, c, ch)
flattenH f (ConstantMedium nid m h) = do
  mi <- flattenM f m
  b <- flattenH f h
  ref #{const RT_CONSTANT_MEDIUM} <$> push (fMedia f) (nid, mi, b)

v3 :: Vec3 -> [Double]
v3 (Vec3 x y z) = [x, y, z]

cameraBlock :: Camera -> [Double]
cameraBlock (Camera o llc hz vt u v w lr t0 t1) = concatMap v3 [o, llc, hz, vt, u, v, w] ++ [lr, t0, t1] -- Lib.hs:1230-1251

-- ---------------------------------------------------------------------------------------------------------------------------
-- marshalling: every table becomes one C array; RtSceneDesc / RtRenderParams are filled at the offsets hsc2hs reads from
-- include/rt_abi.h, so the layout cannot drift from the header
-- ---------------------------------------------------------------------------------------------------------------------------
-- | Allocate C arrays for the scene, run the action with the filled descriptor, free everything (the library copies the
-- scene to the device inside rt_scene_create and never keeps a host pointer).
withSceneDesc :: Scene -> Camera -> (Ptr RtSceneDesc -> IO a) -> IO a
withSceneDesc (world, lights, Albedo bg) cam act = do
  f <- newFlat
  worldRef <- flattenH f world
  lightsRef <- flattenH f lights
  nodes <- rowsOf (fNodes f); sph <- rowsOf (fSpheres f); mov <- rowsOf (fMoving f); rcts <- rowsOf (fRects f)
  cubs <- rowsOf (fCuboids f); trs <- rowsOf (fTranslates f); rots <- rowsOf (fRotates f); meds <- rowsOf (fMedia f)
  mats <- rowsOf (fMats f); texs <- rowsOf (fTexs f); pers <- rowsOf (fPerlin f); imgs <- rowsOf (fImages f)
  frees <- newIORef ([] :: [IO ()])
  let arr :: Storable e => [e] -> IO (Ptr e)
      arr xs = do p <- newArray xs; modifyIORef frees (free p :); return p
      dbl = arr :: [Double] -> IO (Ptr Double)
      u32 = arr :: [Word32] -> IO (Ptr Word32)
      boxD (Box mn mx) = v3 mn ++ v3 mx
      n :: [e] -> Word32
      n = fromIntegral . length
  (`finally` (readIORef frees >>= sequence_)) $ allocaBytes #{size RtSceneDesc} $ \d -> do
    fillBytes d 0 #{size RtSceneDesc}
    #{poke RtSceneDesc, abi_version} d (#{const RT_ABI_VERSION} :: Word32)
    -- BVH nodes
    #{poke RtSceneDesc, n_nodes} d (n nodes)
    #{poke RtSceneDesc, node_box} d =<< dbl (concat [boxD b | (b, _, _, _) <- nodes])
    #{poke RtSceneDesc, node_left} d =<< u32 [l | (_, l, _, _) <- nodes]
    #{poke RtSceneDesc, node_right} d =<< u32 [r | (_, _, r, _) <- nodes]
    #{poke RtSceneDesc, node_size} d =<< u32 [fromIntegral s | (_, _, _, s) <- nodes]
    -- leaves
    #{poke RtSceneDesc, n_spheres} d (n sph)
    #{poke RtSceneDesc, sphere_cr} d =<< dbl (concat [v3 c ++ [r] | (c, r, _) <- sph])
    #{poke RtSceneDesc, sphere_mat} d =<< u32 [m | (_, _, m) <- sph]
    #{poke RtSceneDesc, n_moving} d (n mov)
    #{poke RtSceneDesc, moving} d =<< dbl (concat [v3 c0 ++ v3 c1 ++ [t0, dur, r] | (c0, c1, t0, dur, r, _) <- mov])
    #{poke RtSceneDesc, moving_mat} d =<< u32 [m | (_, _, _, _, _, m) <- mov]
    #{poke RtSceneDesc, n_rects} d (n rcts)
    #{poke RtSceneDesc, rect} d =<< dbl (concat [[a0, a1, b0, b1, k] | ((a0, a1, b0, b1, k), _) <- rcts])
    #{poke RtSceneDesc, rect_mat} d =<< u32 (map snd rcts)
    #{poke RtSceneDesc, n_cuboids} d (n cubs)
    #{poke RtSceneDesc, cuboid} d =<< dbl (concat [v3 mn ++ v3 mx | (mn, mx, _) <- cubs])
    #{poke RtSceneDesc, cuboid_mat} d =<< u32 [m | (_, _, m) <- cubs]
    -- instances and media
    #{poke RtSceneDesc, n_translates} d (n trs)
    #{poke RtSceneDesc, translate_off} d =<< dbl (concatMap (v3 . fst) trs)
    #{poke RtSceneDesc, translate_child} d =<< u32 (map snd trs)
    #{poke RtSceneDesc, n_rotates} d (n rots)
    #{poke RtSceneDesc, rotate_sc} d =<< dbl (concat [[s, c] | (_, s, c, _) <- rots])
    #{poke RtSceneDesc, rotate_axis} d =<< u32 [a | (a, _, _, _) <- rots]
    #{poke RtSceneDesc, rotate_child} d =<< u32 [c | (_, _, _, c) <- rots]
    #{poke RtSceneDesc, n_media} d (n meds)
    #{poke RtSceneDesc, medium_neg_inv_density} d =<< dbl [x | (x, _, _) <- meds]
    #{poke RtSceneDesc, medium_mat} d =<< u32 [m | (_, m, _) <- meds]
    #{poke RtSceneDesc, medium_boundary} d =<< u32 [b | (_, _, b) <- meds]
    -- materials and textures
    #{poke RtSceneDesc, n_materials} d (n mats)
    #{poke RtSceneDesc, mat_kind} d =<< u32 [k | (k, _, _) <- mats]
    #{poke RtSceneDesc, mat_tex} d =<< u32 [t | (_, t, _) <- mats]
    #{poke RtSceneDesc, mat_param} d =<< dbl [p | (_, _, p) <- mats]
    #{poke RtSceneDesc, n_textures} d (n texs)
    #{poke RtSceneDesc, tex_kind} d =<< u32 [k | (k, _, _, _, _) <- texs]
    #{poke RtSceneDesc, tex_rgb} d =<< dbl (concat [v3 c | (_, c, _, _, _) <- texs])
    #{poke RtSceneDesc, tex_a} d =<< u32 [a | (_, _, a, _, _) <- texs]
    #{poke RtSceneDesc, tex_b} d =<< u32 [b | (_, _, _, b, _) <- texs]
    #{poke RtSceneDesc, tex_scale} d =<< dbl [s | (_, _, _, _, s) <- texs]
    #{poke RtSceneDesc, n_perlin} d (n pers)
    #{poke RtSceneDesc, perlin_ranvec} d =<< dbl (concat [concatMap v3 (VV.toList rv) | (rv, _, _, _) <- pers])
    #{poke RtSceneDesc, perlin_perm} d =<<
      (arr (concat [map fromIntegral (V.toList px ++ V.toList py ++ V.toList pz) | (_, px, py, pz) <- pers]) :: IO (Ptr Int32))
    -- images: one RGB8 block each; the pointer table is itself a C array
    imgPtrs <- forM imgs $ \(px, _, _) -> do
      p <- mallocBytes (SV.length px) :: IO (Ptr Word8)
      modifyIORef frees (free p :)
      SV.unsafeWith px $ \src -> copyBytes p src (SV.length px)
      return p
    #{poke RtSceneDesc, n_images} d (n imgs)
    #{poke RtSceneDesc, image_rgb} d =<< (arr imgPtrs :: IO (Ptr (Ptr Word8)))
    #{poke RtSceneDesc, image_w} d =<< u32 [fromIntegral w | (_, w, _) <- imgs]
    #{poke RtSceneDesc, image_h} d =<< u32 [fromIntegral h | (_, _, h) <- imgs]
    -- roots, background, camera
    #{poke RtSceneDesc, world_root} d worldRef
    #{poke RtSceneDesc, lights_root} d lightsRef
    pokeArray (#{ptr RtSceneDesc, background} d) (v3 bg)
    pokeArray (#{ptr RtSceneDesc, camera} d) (cameraBlock cam)
    act d

-- ---------------------------------------------------------------------------------------------------------------------------
-- runRender with the reference's signature (Lib.hs:1491)
-- ---------------------------------------------------------------------------------------------------------------------------
-- | Same type and same result shape as `runRender`: H rows (top row first, as pixelPositions, Lib.hs:1488-1489) of W `RGB`.
-- The device RNG is counter-based -- sample s of pixel p draws from streamGen(seed, p, s) -- so one 64-bit seed taken from the
-- first generator replaces the per-column generator chain of Lib.hs:1499-1523 (DESIGN.md section 3); pixel values are therefore
-- not those of the CPU build sample for sample, they agree in distribution (tests/test_gpu_parity.py).
-- NOINLINE: the unsafePerformIO body uploads a scene and renders; GHC must not duplicate or float it (ADVICE r1).
{-# NOINLINE runRenderB200 #-}
runRenderB200 :: RenderStaticEnv -> [RandGen] -> [VV.Vector RGB]
runRenderB200 (RenderStaticEnv (scene, cam, (w, h), ns, depth, _numThreads, _)) gens = unsafePerformIO $ do
  ver <- c_rt_abi_version
  when (ver /= #{const RT_ABI_VERSION}) $ error "rt_b200: ABI version mismatch between rt_abi.h and librt_b200.so"
  -- the reference parallelises inside runRender (sparks, Lib.hs:1519-1520): so does the drop-in -- every visible GPU renders a
  -- slice of the samples and the partial frames are summed over NVLink inside rt_render (n_gpus = 0: the rt_init default)
  _ <- c_rt_init 0 0
  let seed = case gens of
        (g : _) -> fst (randomWord64 g)
        []      -> error "runRender: no generators"
  sums <- mallocForeignPtrArray (w * h * 3) :: IO (ForeignPtr Double)
  withSceneDesc scene cam $ \desc ->
    bracket
      (alloca $ \hp -> c_rt_scene_create desc (-1) hp >>= check >> peek hp)
      c_rt_scene_destroy $ \sc ->
        allocaBytes #{size RtRenderParams} $ \p -> do
          fillBytes p 0 #{size RtRenderParams} -- row0 = row1 = col0 = col1 = 0: the whole image; flags 0; reserved 0
          #{poke RtRenderParams, width} p (fromIntegral w :: Word32)
          #{poke RtRenderParams, height} p (fromIntegral h :: Word32)
          #{poke RtRenderParams, sample_offset} p (0 :: Word32)
          #{poke RtRenderParams, sample_count} p (fromIntegral ns :: Word32)
          #{poke RtRenderParams, total_samples} p (fromIntegral ns :: Word32)
          #{poke RtRenderParams, max_depth} p (fromIntegral depth :: Word32)
          #{poke RtRenderParams, seed} p (seed :: Word64)
          #{poke RtRenderParams, precision} p (#{const RT_PREC_DEFAULT} :: Word32)
          #{poke RtRenderParams, device} p (-1 :: Int32)
          #{poke RtRenderParams, n_gpus} p (0 :: Word32)
          withForeignPtr sums $ \out -> c_rt_render sc p out nullPtr >>= check
  -- rt_render returns per-pixel radiance SUMS (double, row 0 = top); divide by ns and albedoToColor exactly as renderPos /
  -- runRender do (Lib.hs:1349, 287-292), so the 8-bit conversion -- NaN behaviour included -- stays the reference's own code
  forM [0 .. h - 1] $ \r ->
    VV.generateM w $ \x -> withForeignPtr sums $ \p -> do
      let at c = peekElemOff p ((r * w + x) * 3 + c)
      cr <- at 0; cg <- at 1; cb <- at 2
      return (albedoToColor (Albedo (divide (Vec3 cr cg cb) (fromIntegral ns))))
