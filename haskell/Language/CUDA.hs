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
-- sources (Data.Program, Data.Expression, Data.Expression.Const) but has NOT been compiled; the same calls are
-- exercised from Python (stochaskell_b200/api.py) by the test suite.  INTEGRATION.md lists what a maintainer adds.
module Language.CUDA
  ( compileCUDA, runCUDA, hmcCUDA, CudaMethod(..), CudaEngine(..), defaultCudaMethod, stanDefaultMethod
  , StkCtx, StkModule, StkChains, withCtx
  ) where

import Control.Monad
import Data.Array.Abstract (toList)
import Data.Either.Utils (fromRight')
import Data.Expression
import Data.Expression.Const
import Data.Expression.Eval (evalTuple, emptyEnv)
import Data.List (intercalate)
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
  UnionT tss ->
    FieldZ (nm ++"_tag") [[c | Tagged c _ <- cs]] :
    concat [ slotFields (nm ++"_"++ slot) ty [ if c == ctor then Just (fs !! j) else Nothing | Tagged c fs <- cs ]
           | (slot, ty, ctor, j) <- unionSlots tss ]
  _ -> error $ "Language.CUDA: input of type "++ show t ++" is not supported"
  where nm = "in"++ show k
        cap = maximum (length . toList <$> cs)
        lenField n = FieldZ (n ++"_len") [[length (toList c) | c <- cs]]
        transposeTo n rows = [ [ if i < length r then r !! i else 0 | r <- rows ] | i <- [0..n-1] ]
        slotFields = error "slotFields: one Field per slot, Nothing -> 0 (see flattenInput in stochaskell_b200/api.py)"
        unionSlots = error "unionSlots: (slot name, type, constructor, field index) in first-occurrence order"

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
        downloadStates = error "downloadStates: stk_get_state per field, inverse of flattenInput"

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

-- | static HMC on a builtin module ("builtin:linreg", ...): draws[sample][param][chain] on the unconstrained scale.
-- hmcStan / hmcStanInit (Stan.hs:493-500) return [t]; here every chain's draws come back, chains x draws.
hmcCUDA :: CudaMethod -> String -> [(String, [Double])] -> Maybe [Double] -> IO [[[Double]]]
hmcCUDA CudaSample{..} spec dat initq = withCtx 0 $ \ctx ->
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
    let p = length (maybe [] id initq)   -- callers normally query stk_num_params
        n = numSamples * p * numChains
    draws <- allocaArray n $ \buf -> alloca $ \pn -> do
      check "stk_get_draws" $ c_get_draws ch buf (fromIntegral $ 8 * n) pn
      map realToFrac <$> peekArray n buf
    _ <- c_chains_destroy ch
    return $ chunk p . chunk numChains $ draws
  where chunk k [] = []
        chunk k xs = let (a,b) = splitAt k xs in a : chunk k b
