{-# LANGUAGE ForeignFunctionInterface, ScopedTypeVariables, RecordWildCards #-}
-- | Language.CUDA — the B200 sampling backend of Stochaskell, sitting where "Language.CC" sits today.
--
-- Same front end (the P / Expression DSL, the IR, `lpdf`, `rjmc`, ...); what changes is the layer that EXECUTES a
-- probabilistic program:
--
-- >  compileCC  :: (s -> P t) -> IO (s -> IO t)            -- Language/CC.hs:465   g++ + one process per call
-- >  compileCUDA :: (s -> P t) -> IO (Int -> [s] -> IO [t]) -- here: NVRTC once, then n transitions of many chains per call
--
-- The lowering itself (ccProgram/ccDAG/ccNode/ccPNode, CC.hs:123-432) lives in libstk_cuda (`stk_ir_to_cuda`,
-- csrc/stkc.cpp): this module hands it the text `showPBlock` prints — exactly what `ccProgram` embeds at the top of
-- every generated file (CC.hs:400-402) — plus the input types, and drives the C-ABI of include/stk_cuda.h.
--
-- NOTE: GHC is not available in the build image of this repository, so this module is written against the reference's
-- sources (Data.Program, Data.Expression, Data.Expression.Const) but has NOT been compiled; every call it makes, the
-- state layout (flattenInput / unflattenOutput = stochaskell_b200/layout.py, tests/test_layout_cpu.py) and both entry
-- shapes (compileCUDA's closure on values = api.CompiledKernel; hmcCUDA on a conditioned model = api.hmcCUDA with IR text)
-- are exercised from Python by the test suite.  INTEGRATION.md lists what a maintainer adds.
module Language.CUDA
  ( compileCUDA, runCUDA, hmcCUDA, hmcCUDAModule, CudaMethod(..), CudaEngine(..), defaultCudaMethod, stanDefaultMethod
  , StkCtx, StkModule, StkChains, withCtx
  ) where

import Control.Monad
import Data.Array.Abstract (toList)
import Data.Either.Utils (fromRight')
import Data.Expression
import Data.Expression.Const
import Data.Expression.Eval (evalTuple, emptyEnv)
import Data.List (intercalate, isPrefixOf)
import qualified Data.Map as Map
import Data.Maybe (fromJust)
import Data.Program
import Foreign
import Foreign.C.String
import Foreign.C.Types
import System.Directory (getCurrentDirectory)

------------------------------------------------------------------------------
-- FFI: include/stk_cuda.h                                                  --
------------------------------------------------------------------------------

data StkCtx_; data StkModule_; data StkChains_
type StkCtx = Ptr StkCtx_; type StkModule = Ptr StkModule_; type StkChains = Ptr StkChains_

foreign import ccall unsafe "stk_last_error"   c_last_error   :: IO CString
foreign import ccall unsafe "stk_ctx_create"   c_ctx_create   :: CInt -> Ptr StkCtx -> IO CInt
foreign import ccall unsafe "stk_ctx_destroy"  c_ctx_destroy  :: StkCtx -> IO CInt
foreign import ccall unsafe "stk_ir_to_cuda"   c_ir_to_cuda   :: CString -> CString -> CString -> CString -> CSize -> Ptr CSize -> IO CInt
foreign import ccall safe   "stk_module_load"  c_module_load  :: StkCtx -> CString -> CString -> Ptr StkModule -> IO CInt
foreign import ccall unsafe "stk_module_field" c_module_field :: StkModule -> CString -> IO CInt
foreign import ccall unsafe "stk_module_kernel" c_module_kernel :: StkModule -> CString -> IO CInt
foreign import ccall unsafe "stk_chains_create" c_chains_create :: StkModule -> Int64 -> Word64 -> Ptr StkChains -> IO CInt
foreign import ccall unsafe "stk_chains_destroy" c_chains_destroy :: StkChains -> IO CInt
foreign import ccall unsafe "stk_set_data"     c_set_data     :: StkChains -> CInt -> Ptr () -> CSize -> IO CInt
foreign import ccall unsafe "stk_set_state"    c_set_state    :: StkChains -> CInt -> Ptr () -> CSize -> CInt -> IO CInt
foreign import ccall unsafe "stk_get_state"    c_get_state    :: StkChains -> CInt -> Ptr () -> CSize -> IO CInt
foreign import ccall unsafe "stk_device_ptr"   c_device_ptr   :: StkChains -> CInt -> Ptr (Ptr ()) -> Ptr CSize -> IO CInt
foreign import ccall unsafe "stk_num_params"   c_num_params   :: StkChains -> IO CInt
foreign import ccall unsafe "stk_set_param"    c_set_param    :: StkChains -> CString -> CDouble -> IO CInt
foreign import ccall safe   "stk_step"         c_step         :: StkChains -> CInt -> Int64 -> Ptr () -> IO CInt
foreign import ccall safe   "stk_simulate"     c_simulate     :: StkChains -> Ptr () -> IO CInt
foreign import ccall unsafe "stk_record"       c_record       :: StkChains -> Int64 -> Int64 -> IO CInt
foreign import ccall safe   "stk_get_draws"    c_get_draws    :: StkChains -> Ptr CDouble -> CSize -> Ptr Int64 -> IO CInt
foreign import ccall safe   "stk_moments"      c_moments      :: StkChains -> Ptr CDouble -> Ptr CDouble -> Ptr CDouble -> Ptr CDouble -> Ptr CDouble -> IO CInt
foreign import ccall safe   "stk_sync"         c_sync         :: StkChains -> IO CInt

-- | every entry point returns 0 or a negative status; the message is thread-local (stk_last_error).
-- Mirrors the reference's habit of failing with `error` (Util.hs:118-120) — no exception crosses the C boundary.
check :: String -> IO CInt -> IO ()
check what act = do
  rc <- act
  when (rc < 0) $ do
    msg <- c_last_error >>= peekCString
    error $ "Language.CUDA: "++ what ++": "++ msg

withCtx :: Int -> (StkCtx -> IO a) -> IO a
withCtx dev f = alloca $ \p -> do
  check "stk_ctx_create" $ c_ctx_create (fromIntegral dev) p
  ctx <- peek p
  r <- f ctx
  _ <- c_ctx_destroy ctx
  return r

------------------------------------------------------------------------------
-- IR -> module                                                             --
------------------------------------------------------------------------------

-- | the CUDA text of a kernel  s -> P t  (ccProgram's counterpart, CC.hs:398-432)
cudaProgram :: forall s t. (ExprTuple s, ExprTuple t) => [Type] -> (s -> P t) -> IO String
cudaProgram ts prog = do
  let (rets, pb) = runProgExprs "cuda" . prog $ toExprTuple
        [DExpr . return $ Var (Dummy 0 i) t | (i,t) <- zip [0..] ts]   -- same Dummy 0 i inputs as CC.hs:430-431
      ir = showPBlock pb $ show rets
      tys = intercalate ";" (show <$> ts)
  withCString ir $ \cir -> withCString tys $ \ctys -> withCString "" $ \copt -> alloca $ \pn -> do
    check "stk_ir_to_cuda" $ c_ir_to_cuda cir ctys copt nullPtr 0 pn
    n <- peek pn
    allocaBytes (fromIntegral n) $ \buf -> do
      check "stk_ir_to_cuda" $ c_ir_to_cuda cir ctys copt buf n nullPtr
      peekCString buf

-- | flatten one ConstVal into the state fields the lowering declares for input k (csrc/stkc.cpp "state layout"):
-- scalars -> in<k>; vectors -> in<k> + in<k>_len; unions -> in<k>_tag + one slot per (kind, occurrence).
-- Replaces printCC (CC.hs:434-438): the state goes to the GPU as arrays [len][n_chains], not as text.
data Field = FieldZ String [[Int]] | FieldR String [[Double]]   -- name, [element][chain]

flattenInput :: Int -> Type -> [ConstVal] -> [Field]
flattenInput k t cs = case t of
  IntT  -> [FieldZ nm [fromIntegral . toInteger <$> cs]]
  RealT -> [FieldR nm [toDouble <$> cs]]
  SubrangeT t' _ _ -> flattenInput k t' cs
  ArrayT _ [_] RealT -> [FieldR nm (transposeTo cap $ map toDouble . toList <$> cs), lenField nm]
  ArrayT _ [_] _     -> [FieldZ nm (transposeTo cap $ map (fromIntegral . toInteger) . toList <$> cs), lenField nm]
  UnionT tss -> unionFields nm tss cs
  _ -> error $ "Language.CUDA: input of type "++ show t ++" is not supported"
  where nm = "in"++ show k
        cap = maximum (length . toList <$> cs)
        lenField n = FieldZ (n ++"_len") [[length (toList c) | c <- cs]]
        transposeTo n rows = [ [ if i < length r then r !! i else 0 | r <- rows ] | i <- [0..n-1] ]

-- | kind of a union field as the lowering classifies it (csrc/stkc.cpp union_layout): r / z scalars, rv / zv vectors
slotKind :: Type -> String
slotKind RealT = "r"
slotKind IntT  = "z"
slotKind (SubrangeT t _ _) = slotKind t
slotKind (ArrayT _ [_] t) = slotKind t ++ "v"
slotKind t = error $ "Language.CUDA: union field of type "++ show t ++" is not supported"

-- | the slots of a union, in first-occurrence order, and for every constructor the slot of each of its fields:
-- field j of constructor c that is the i-th field of its kind IN THAT CONSTRUCTOR uses slot <kind><i>; constructors share
-- slots (stochaskell_b200/layout.py union_slots is the same function; tests/test_layout_cpu.py pins it to the lowering)
unionSlots :: [[Type]] -> ([(String, Type)], [[Int]])
unionSlots tss = foldl ctor ([], []) tss
  where ctor (slots, cmap) ts =
          let (slots', fmap, _) = foldl field (slots, [], []) ts in (slots', cmap ++ [fmap])
        field (slots, fmap, used) t =
          let k = slotKind t
              i = length (filter (== k) used)
              name = k ++ show i
          in case lookup name (zip (map fst slots) [0..]) of
               Just si -> (slots, fmap ++ [si], used ++ [k])
               Nothing -> (slots ++ [(name, t)], fmap ++ [length slots], used ++ [k])

-- | one Field (two for a vector: data + length) per slot; a chain whose constructor does not use the slot holds 0 / empty
slotFields :: String -> Type -> [Maybe ConstVal] -> [Field]
slotFields name t ms = case slotKind t of
  "r"  -> [FieldR name [[maybe 0 toDouble m | m <- ms]]]
  "z"  -> [FieldZ name [[maybe 0 (fromIntegral . toInteger) m | m <- ms]]]
  "rv" -> [FieldR name (pad [maybe [] (map toDouble . toList) m | m <- ms]), lens]
  _    -> [FieldZ name (pad [maybe [] (map (fromIntegral . toInteger) . toList) m | m <- ms]), lens]
  where lens = FieldZ (name ++"_len") [[maybe 0 (length . toList) m | m <- ms]]
        pad :: Num a => [[a]] -> [[a]]
        pad rows = let n = maximum (1 : map length rows)
                   in [ [ if i < length r then r !! i else 0 | r <- rows ] | i <- [0..n-1] ]

-- | fields of a union-typed input: tag + slots
unionFields :: String -> [[Type]] -> [ConstVal] -> [Field]
unionFields nm tss cs =
  FieldZ (nm ++"_tag") [[c | Tagged c _ <- cs]] :
  concat [ slotFields (nm ++"_"++ sname) sty
             [ case [fs !! j | (j, s) <- zip [0..] (cmap !! c), s == si] of { (v:_) -> Just v; [] -> Nothing } | Tagged c fs <- cs ]
         | (si, (sname, sty)) <- zip [0..] slots ]
  where (slots, cmap) = unionSlots tss

-- | compile a probabilistic program to a resident GPU module (compileCC's counterpart, CC.hs:465-498).
-- The closure runs `n` fused transitions of every chain in `xs` and returns the new states; with a single chain and
-- n = 1 it has exactly compileCC's shape.  The SHA1-keyed cubin cache lives in ./cache/cuda (CC.hs:471-479).
compileCUDA :: forall s t. (ExprTuple s, ExprTuple t) => (s -> P t) -> IO (Int -> [s] -> IO [t])
compileCUDA prog = do
  let TypesIs ts = typesOf :: TypesOf s
  code <- cudaProgram ts prog
  pwd <- getCurrentDirectory
  return $ \n xs -> withCtx 0 $ \ctx ->
    withCString code $ \ccode -> withCString (pwd ++"/cache/cuda") $ \cdir ->
    alloca $ \pm -> alloca $ \pc -> do
      check "stk_module_load" $ c_module_load ctx ccode cdir pm
      m <- peek pm
      check "stk_chains_create" $ c_chains_create m (fromIntegral $ length xs) 1 pc
      ch <- peek pc
      let inputs = fromRight' . evalTuple emptyEnv <$> xs :: [[ConstVal]]
      forM_ (zip3 [0..] ts (transpose' inputs)) $ \(k,t,cs) ->
        forM_ (flattenInput k t cs) $ \f -> uploadField m ch f
      kid <- withCString "step" $ c_module_kernel m
      check "stk_step" $ c_step ch kid (fromIntegral n) nullPtr
      check "stk_sync" $ c_sync ch
      outs <- downloadStates m ch ts (length xs)          -- readCC's counterpart (CC.hs:446-456)
      _ <- c_chains_destroy ch
      return $ fromConstVals <$> outs
  where transpose' = foldr (zipWith (:)) (repeat [])
        uploadField m ch f = case f of
          FieldZ name rows -> withCString name $ \cn -> do
            i <- c_module_field m cn
            withArrayLen (fromIntegral <$> concat rows :: [Int32]) $ \len p ->
              check ("stk_set_state "++ name) $ c_set_state ch i (castPtr p) (fromIntegral $ 4 * len) 0
          FieldR name rows -> withCString name $ \cn -> do
            i <- c_module_field m cn
            withArrayLen (realToFrac <$> concat rows :: [CDouble]) $ \len p ->
              check ("stk_set_state "++ name) $ c_set_state ch i (castPtr p) (fromIntegral $ 8 * len) 0

-- | readCC's counterpart (CC.hs:446-456): the state fields of output k of type t, downloaded with stk_get_state, back to
-- one ConstVal per chain — the inverse of 'flattenInput' (stochaskell_b200/layout.py unflatten_output)
downloadStates :: StkModule -> StkChains -> [Type] -> Int -> IO [[ConstVal]]
downloadStates m ch ts nChains = do
  cols <- forM (zip [0..] ts) $ \(k,t) -> unflattenOutput m ch nChains ("in"++ show (k :: Int)) t
  return $ foldr (zipWith (:)) (repeat []) cols        -- [output][chain] -> [chain][output]

getZ :: StkModule -> StkChains -> Int -> String -> IO [[Int]]          -- [element][chain]
getZ m ch nChains name = withCString name $ \cn -> do
  i <- c_module_field m cn
  len <- fieldLength ch i nChains 4
  allocaArray (len * nChains) $ \(p :: Ptr Int32) -> do
    check ("stk_get_state "++ name) $ c_get_state ch i (castPtr p) (fromIntegral $ 4 * len * nChains)
    chunk nChains . map fromIntegral <$> peekArray (len * nChains) p

getR :: StkModule -> StkChains -> Int -> String -> IO [[Double]]
getR m ch nChains name = withCString name $ \cn -> do
  i <- c_module_field m cn
  len <- fieldLength ch i nChains 8
  allocaArray (len * nChains) $ \(p :: Ptr CDouble) -> do
    check ("stk_get_state "++ name) $ c_get_state ch i (castPtr p) (fromIntegral $ 8 * len * nChains)
    chunk nChains . map realToFrac <$> peekArray (len * nChains) p

-- | per-chain length of a state field: its byte size (stk_device_ptr) / element size / chains
fieldLength :: StkChains -> CInt -> Int -> Int -> IO Int
fieldLength ch i nChains esize = alloca $ \pp -> alloca $ \pb -> do
  check "stk_device_ptr" $ c_device_ptr ch i pp pb
  bytes <- peek pb
  return $ fromIntegral bytes `div` (esize * nChains)

chunk :: Int -> [a] -> [[a]]
chunk _ [] = []
chunk k xs = let (a,b) = splitAt k xs in a : chunk k b

unflattenOutput :: StkModule -> StkChains -> Int -> String -> Type -> IO [ConstVal]
unflattenOutput m ch n nm t = case t of
  IntT  -> map fromIntegral . head <$> getZ m ch n nm
  RealT -> map fromDouble . head <$> getR m ch n nm
  SubrangeT t' _ _ -> unflattenOutput m ch n nm t'
  ArrayT _ [_] RealT -> do
    rows <- getR m ch n nm; [lens] <- getZ m ch n (nm ++"_len")
    return [ fromList [ fromDouble (rows !! i !! c) | i <- [0 .. l-1] ] | (c,l) <- zip [0..] lens ]
  ArrayT _ [_] _ -> do
    rows <- getZ m ch n nm; [lens] <- getZ m ch n (nm ++"_len")
    return [ fromList [ fromIntegral (rows !! i !! c) | i <- [0 .. l-1] ] | (c,l) <- zip [0..] lens ]
  UnionT tss -> do
    [tags] <- getZ m ch n (nm ++"_tag")
    let (slots, cmap) = unionSlots tss
    slotVals <- forM slots $ \(sname, sty) -> unflattenOutput m ch n (nm ++"_"++ sname) sty
    return [ Tagged c [ slotVals !! si !! chain | si <- cmap !! c ] | (chain, c) <- zip [0..] tags ]
  _ -> error $ "Language.CUDA: output of type "++ show t ++" is not supported"

-- | sample a probabilistic program via 'compileCUDA' (runCC's counterpart, CC.hs:459-462)
runCUDA :: (ExprTuple s, ExprTuple t) => (s -> P t) -> s -> IO t
runCUDA prog x = do
  p <- compileCUDA prog
  head <$> p 1 [x]

------------------------------------------------------------------------------
-- HMC (hmcStan's counterpart)                                              --
------------------------------------------------------------------------------

-- | mirrors StanMethod / StanHMCEngine (Language/Stan.hs:378-440).  Engines the kernels implement:
--   CudaStatic  static HMC, unit_e metric, no adaptation (= benchStanHMC, app/benchmark/Main.hs:248-262): kernel "hmc"
--   CudaStaticAdapt  Stan's adapt_diag_e_static_hmc per chain (dual averaging + windowed diagonal metric): "hmc_adapt"
--   CudaNuts    hmcStan's DEFAULT (Stan.hs:428-440: NUTS, diag_e, adaptation, max_depth 10, stepsize 1): kernel "nuts"
data CudaEngine = CudaStatic | CudaStaticAdapt | CudaNuts deriving (Eq, Show)

data CudaMethod = CudaSample
  { numChains  :: Int
  , numSamples :: Int
  , numWarmup  :: Int
  , thin       :: Int
  , hmcStepSize :: Double
  , hmcNumLeapfrog :: Int   -- int_time = hmcNumLeapfrog * hmcStepSize
  , seed       :: Word64
  , engine     :: CudaEngine
  , adaptDelta :: Double    -- 0.8; gamma 0.05, kappa 0.75, t0 10, windows 75/50/25 are the module's defaults (= CmdStan's)
  , maxDepth   :: Int
  }

defaultCudaMethod :: CudaMethod
defaultCudaMethod = CudaSample
  { numChains = 4096, numSamples = 1000, numWarmup = 1000, thin = 1
  , hmcStepSize = 0.01, hmcNumLeapfrog = 10, seed = 1
  , engine = CudaStatic, adaptDelta = 0.8, maxDepth = 10 }

-- | what `hmcStan n` runs (Stan.hs:493-496): NUTS with adaptation, n warm-up + n sampling transitions
stanDefaultMethod :: Int -> CudaMethod
stanDefaultMethod n = defaultCudaMethod { numSamples = n, numWarmup = n, hmcStepSize = 1, engine = CudaNuts }

-- | hmcStan's counterpart in the reference's own shape (runStan :: StanMethod -> P t -> Maybe t -> IO [t], Stan.hs:443;
-- hmcStan / hmcStanInit, Stan.hs:493-500): the conditioned model goes in, chains x draws of `t` come out.
--   * the program is printed with showPBlock and lowered by stk_ir_to_cuda in `mode=density` (csrc/stkc_grad.inc): data =
--     the PNodes of the PBlock's `given` map, parameters = the others (stanProgram's split, Stan.hs:325-344);
--   * the observed values are uploaded as the data fields named after their PNodes (stanDump's role, Stan.hs:346-363);
--   * the unconstrained draws come back through the module's `qmap` lines (the constraining transform per parameter
--     block) and are turned into `t` by evaluating the program's return expressions (stanRead + evalPBlock, Stan.hs:365-371, 487).
hmcCUDA :: forall t. (ExprTuple t) => CudaMethod -> P t -> Maybe t -> IO [[t]]
hmcCUDA method prog initial = do
  let (rets, pb@(PBlock _ _ given _)) = runProgExprs "stan" prog
      ir = showPBlock pb $ show rets
      givenIds = [ show i | (LVar i, _) <- Map.toAscList given ]
      opts = "mode=density given="++ intercalate "," givenIds
  code <- withCString ir $ \cir -> withCString "" $ \ctys -> withCString opts $ \copt -> alloca $ \pn -> do
    check "stk_ir_to_cuda" $ c_ir_to_cuda cir ctys copt nullPtr 0 pn
    n <- peek pn
    allocaBytes (fromIntegral n) $ \buf -> check "stk_ir_to_cuda" (c_ir_to_cuda cir ctys copt buf n nullPtr) >> peekCString buf
  let dat = [ (show i, toDouble <$> elems' c) | (LVar i, (c,_)) <- Map.toAscList given ]   -- row-major, as the data fields expect
      qmaps = [ words l | l <- lines code, take 8 (dropWhile (`elem` " \"") l) == "qmap x_s" || "qmap " `isPrefixOf` dropWhile (`elem` " \"") l ]
  -- initial values (hmcStanInit): the constrained value of every parameter block, pushed through the inverse transforms
  let initq = unconstrainInit qmaps pb rets <$> initial
  draws <- hmcCUDAModule method code dat initq
  -- draws [sample][param][chain] -> per chain, per sample: the parameter values on the model's scale -> t
  return [ [ fromRight' . evalPBlock pb rets $ constrainEnv qmaps pb [ draws !! sIdx !! j !! c | j <- [0 .. length (head (head draws)) - 1] ]
           | sIdx <- [0 .. length draws - 1] ]
         | c <- [0 .. numChains method - 1] ]
  where elems' = map fromRational . toList . toRationalArray
        -- qmap <PNode> <offset> <rows> <cols> <lower|-> <upper|->  (descriptor lines of a generated module)
        constrainEnv qmaps pb q = Map.fromList
          [ (LVar (pnodeId pb name), reshape rows cols [ con lo hi (q !! (off + k)) | k <- [0 .. max 1 rows * max 1 cols - 1] ])
          | (_:name:offS:rowsS:colsS:lo:hi:_) <- qmaps, let off = read offS; rows = read rowsS; cols = read colsS ]
        con "-" "-" x = x
        con lo  "-" x = read lo + exp x
        con "-" hi  x = read hi - exp x
        con lo  hi  x = read lo + (read hi - read lo) / (1 + exp (negate x))
        unc "-" "-" v = v
        unc lo  "-" v = log (v - read lo)
        unc "-" hi  v = log (read hi - v)
        unc lo  hi  v = let u = (v - read lo) / (read hi - read lo) in log (u / (1 - u))
        unconstrainInit qmaps pb rets t0 =
          let env = unifyTuple (definitions pb) rets t0 emptyEnv          -- the values the caller gave, by PNode (as runStan does, Stan.hs:449-450)
          in concat [ [ unc lo hi v | v <- toDouble <$> elems' (fromJust $ Map.lookup (LVar (pnodeId pb name)) env) ]
                    | (_:name:_:_:_:lo:hi:_) <- qmaps ]
        reshape 0 _ [x] = fromDouble x
        reshape r 0 xs  = fromList (fromDouble <$> xs)
        reshape r c xs  = fromMatrix' r c xs
        pnodeId pb name = head [ i | i <- Map.keys (pnodes pb), show i == name ]

-- | static HMC / adaptive HMC / NUTS on a module given by name ("builtin:linreg", ...) or as generated CUDA text:
-- draws[sample][param][chain] on the unconstrained scale of the module's `q` field.
hmcCUDAModule :: CudaMethod -> String -> [(String, [Double])] -> Maybe [Double] -> IO [[[Double]]]
hmcCUDAModule CudaSample{..} spec dat initq = withCtx 0 $ \ctx ->
  withCString spec $ \cspec -> withCString "cache/cuda" $ \cdir -> alloca $ \pm -> alloca $ \pc -> do
    check "stk_module_load" $ c_module_load ctx cspec cdir pm
    m <- peek pm
    check "stk_chains_create" $ c_chains_create m (fromIntegral numChains) seed pc
    ch <- peek pc
    forM_ dat $ \(name, xs) -> withCString name $ \cn -> do
      i <- c_module_field m cn
      withArrayLen (realToFrac <$> xs :: [CDouble]) $ \len p ->
        check ("stk_set_data "++ name) $ c_set_data ch i (castPtr p) (fromIntegral $ 8 * len)
    withCString "eps" $ \cn -> check "eps" $ c_set_param ch cn (realToFrac hmcStepSize)
    withCString "L"   $ \cn -> check "L"   $ c_set_param ch cn (fromIntegral hmcNumLeapfrog)
    forM_ initq $ \q -> withCString "q" $ \cn -> do        -- hmcStanInit: one initial state, broadcast to all chains
      i <- c_module_field m cn
      withArrayLen (realToFrac <$> q :: [CDouble]) $ \len p ->
        check "stk_set_state q" $ c_set_state ch i (castPtr p) (fromIntegral $ 8 * len) 1
    let setp name v = withCString name $ \cn -> check name $ c_set_param ch cn (realToFrac (v :: Double))
    when (engine /= CudaStatic) $ do     -- the ten adaptation params of csrc/stk_adapt.cuh; the rest keep CmdStan's defaults
      setp "stepsize" hmcStepSize
      setp "int_time" (fromIntegral hmcNumLeapfrog * hmcStepSize)
      setp "num_warmup" (fromIntegral numWarmup)
      setp "delta" adaptDelta
      when (engine == CudaNuts) $ setp "max_depth" (fromIntegral maxDepth)
    kid <- withCString (case engine of { CudaStatic -> "hmc"; CudaStaticAdapt -> "hmc_adapt"; CudaNuts -> "nuts" }) $ c_module_kernel m
    when (numWarmup > 0) $ check "warmup" $ c_step ch kid (fromIntegral numWarmup) nullPtr
    check "stk_record" $ c_record ch (fromIntegral numSamples) (fromIntegral thin)
    check "stk_step" $ c_step ch kid (fromIntegral $ numSamples * thin) nullPtr
    p <- fromIntegral <$> c_num_params ch
    let n = numSamples * p * numChains
    draws <- allocaArray n $ \buf -> alloca $ \pn -> do
      check "stk_get_draws" $ c_get_draws ch buf (fromIntegral $ 8 * n) pn
      map realToFrac <$> peekArray n buf
    _ <- c_chains_destroy ch
    return $ chunk p . chunk numChains $ draws
