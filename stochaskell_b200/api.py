"""Host-side mirror of the reference's entry points for the sampling hot path.

Reference interface (davidar/stochaskell)          ->  here
  compileCC :: (s -> P t) -> IO (s -> IO t)   CC.hs:465   compileCUDA(model) -> Sampler
  runCC                                       CC.hs:459   runCUDA
  runStep / simulate          Program.hs:1350, 1038       Sampler.step / Sampler.simulate
  hmcStan n / hmcStanInit n p x0     Stan.hs:493-500      hmcCUDA(method, model, init)
  StanMethod record                  Stan.hs:378-440      CudaMethod dataclass

Everything below is plumbing over the C-ABI in include/stk_cuda.h: numpy arrays
in HOST memory go in and out; all arithmetic happens in the CUDA kernels.  A
missing library / device raises StkError (no fallback).
"""
from __future__ import annotations

import ctypes as C
import dataclasses
import os
from typing import Dict, Optional, Sequence

import numpy as np

from ._lib import CSRC_DIR, StkError, check, lib

DEFAULT_CACHE = os.environ.get("STOCHASKELL_DIRECTORY", os.getcwd()) + "/cache/cuda"  # Stochaskell.hs:160-162


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class Context:
    """One per GPU (stk_ctx)."""

    def __init__(self, device: int = 0):
        h = C.c_void_p()
        check(lib().stk_ctx_create(device, C.byref(h)))
        self._h = h
        self.device = device

    @property
    def stream(self) -> int:
        s = C.c_void_p()
        check(lib().stk_ctx_stream(self._h, C.byref(s)))
        return s.value or 0

    def close(self):
        if self._h:
            lib().stk_ctx_destroy(self._h)
            self._h = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


def compile_source(cuda_src: str, cache_dir: Optional[str] = None, include_dir: Optional[str] = None) -> str:
    """NVRTC-compile generated CUDA text to an sm_100a cubin (no GPU needed). Returns the cubin path."""
    buf = C.create_string_buffer(4096)
    check(lib().stk_compile(cuda_src.encode(), (cache_dir or DEFAULT_CACHE).encode(),
                            (include_dir or CSRC_DIR).encode(), buf, len(buf)))
    return buf.value.decode()


def ir_to_cuda(ir_text: str, input_types: Sequence[str] = (), **options) -> str:
    """The Language.CUDA lowering (stkc, csrc/stkc.cpp): IR text as printed by showPBlock -> CUDA module source."""
    need = C.c_size_t()
    opts = " ".join(f"{k}={v}" for k, v in options.items()).encode()
    types = ";".join(input_types).encode()
    check(lib().stk_ir_to_cuda(ir_text.encode(), types, opts, None, 0, C.byref(need)))
    buf = C.create_string_buffer(need.value)
    check(lib().stk_ir_to_cuda(ir_text.encode(), types, opts, buf, len(buf), None))
    return buf.value.decode()


class Module:
    """A compiled model (stk_module): 'builtin:<name>', a .cubin path, or CUDA source text."""

    def __init__(self, ctx: Context, spec: str, cache_dir: Optional[str] = None):
        self.ctx = ctx
        if not spec.startswith("builtin:") and not spec.endswith(".cubin"):
            spec = compile_source(spec, cache_dir)  # python side passes the in-tree include dir
        h = C.c_void_p()
        check(lib().stk_module_load(ctx._h, spec.encode(), (cache_dir or DEFAULT_CACHE).encode(), C.byref(h)))
        self._h = h
        buf = C.create_string_buffer(1 << 16)
        check(lib().stk_module_describe(h, buf, len(buf)))
        self.descriptor = buf.value.decode()
        self.fields: Dict[str, dict] = {}
        self.kernels = []
        self.params = {}
        self.nparam = 0
        self.name = ""
        for line in self.descriptor.splitlines():
            t = line.split()
            if not t:
                continue
            if t[0] == "field":
                dt = {"i32": np.int32, "f32": np.float32, "bf16": np.uint16, "f16": np.float16}.get(t[3], np.float64)
                self.fields[t[1]] = dict(index=len(self.fields), state=(t[2] == "state"), kind=t[2], dtype=dt, len=t[4])
            elif t[0] == "kernel" or (t[0] == "plan" and t[2] == "kernel"):
                self.kernels.append(t[1])
            elif t[0] == "param":
                self.params[t[1]] = float(t[2])
            elif t[0] == "nparam":
                self.nparam = int(t[1]) if t[1].isdigit() else 0   # an expression: ask stk_num_params(chains)
            elif t[0] == "name":
                self.name = t[1]

    def kernel_id(self, name: str) -> int:
        return check(lib().stk_module_kernel(self._h, name.encode()))

    def close(self):
        if self._h:
            lib().stk_module_destroy(self._h)
            self._h = None


class Chains:
    """Batched chain state resident on the GPU (stk_chains)."""

    def __init__(self, module: Module, n_chains: int, seed: int = 0, first_global_chain: int = 0):
        self.module = module
        self.n = int(n_chains)
        h = C.c_void_p()
        check(lib().stk_chains_create(module._h, self.n, C.c_uint64(seed), C.byref(h)))
        self._h = h
        if first_global_chain:
            check(lib().stk_chains_set_offset(h, first_global_chain))

    # -- data / state ---------------------------------------------------------
    def _field(self, name: str) -> dict:
        try:
            return self.module.fields[name]
        except KeyError:
            raise StkError(-1, f"module {self.module.name} has no field {name!r}") from None

    def set_data(self, name: str, value) -> None:
        f = self._field(name)
        a = np.ascontiguousarray(value, dtype=f["dtype"])
        check(lib().stk_set_data(self._h, f["index"], _ptr(a), a.nbytes))

    def set_state(self, name: str, value, broadcast: bool = False) -> None:
        """value: [L][n_chains] (or [L] with broadcast=True, the hmcStanInit analogue)."""
        f = self._field(name)
        a = np.ascontiguousarray(value, dtype=f["dtype"])
        check(lib().stk_set_state(self._h, f["index"], _ptr(a), a.nbytes, 1 if broadcast else 0))

    def get_state(self, name: str, out: Optional[np.ndarray] = None) -> np.ndarray:
        """download a field; `out` (flat, right dtype, e.g. a pinned buffer) avoids the allocation"""
        f = self._field(name)
        if out is None:
            d = C.c_void_p(); nb = C.c_size_t()
            check(lib().stk_device_ptr(self._h, f["index"], C.byref(d), C.byref(nb)))
            out = np.empty(nb.value // np.dtype(f["dtype"]).itemsize, dtype=f["dtype"])
        check(lib().stk_get_state(self._h, f["index"], _ptr(out), out.nbytes))
        return out.reshape(-1, self.n) if f["state"] else out

    def device_ptr(self, name: str):
        f = self._field(name)
        d = C.c_void_p(); nb = C.c_size_t()
        check(lib().stk_device_ptr(self._h, f["index"], C.byref(d), C.byref(nb)))
        return d.value, nb.value

    def set_param(self, name: str, value: float) -> None:
        check(lib().stk_set_param(self._h, name.encode(), float(value)))

    # -- compute --------------------------------------------------------------
    def logp(self, host: bool = True) -> Optional[np.ndarray]:
        if not host:                                 # evaluate on the device only (the result stays in the `lp` field / out_lp)
            check(lib().stk_logp(self._h, None))
            return None
        out = np.empty(self.n)
        check(lib().stk_logp(self._h, _ptr(out)))
        return out

    def logp_grad(self, out_lp: Optional[np.ndarray] = None, out_grad: Optional[np.ndarray] = None, host: bool = True):
        """lp [chains] and gradient [P][chains] at the resident state.  `out_*`: caller-owned (e.g. pinned) host buffers;
        `host=False` evaluates on the device only (results stay in the library's buffers, nothing is copied back)."""
        if not host:
            check(lib().stk_logp_grad(self._h, None, None))
            return None, None
        lp = np.empty(self.n) if out_lp is None else out_lp
        g = np.empty((lib().stk_num_params(self._h), self.n)) if out_grad is None else out_grad
        check(lib().stk_logp_grad(self._h, _ptr(lp), _ptr(g)))
        return lp, g

    def step(self, kernel: str | int, n_transitions: int = 1, stream: int = 0) -> None:
        kid = kernel if isinstance(kernel, int) else self.module.kernel_id(kernel)
        check(lib().stk_step(self._h, kid, int(n_transitions), C.c_void_p(stream)))

    def simulate(self, stream: int = 0) -> None:
        check(lib().stk_simulate(self._h, C.c_void_p(stream)))

    def record(self, n_slots: int, thin: int = 1) -> None:
        check(lib().stk_record(self._h, int(n_slots), int(thin)))

    def draws(self) -> np.ndarray:
        """[n_recorded][P_rec][n_chains]"""
        n = C.c_int64()
        check(lib().stk_get_draws(self._h, None, 0, C.byref(n)))
        p = lib().stk_num_recorded_values(self._h)
        out = np.empty((n.value, p, self.n))
        check(lib().stk_get_draws(self._h, _ptr(out), out.nbytes, C.byref(n)))
        return out

    def moments(self):
        p = lib().stk_num_recorded_values(self._h)
        mean, var, rhat, ess = (np.empty(p) for _ in range(4))
        m4 = np.empty((4, p))
        check(lib().stk_moments(self._h, _ptr(mean), _ptr(var), _ptr(rhat), _ptr(ess), _ptr(m4)))
        return dict(mean=mean, var=var, rhat=rhat, ess=ess, moments4=m4)

    def moments_nccl(self, comm: "NcclComm"):
        """R-hat / ESS over the chains of ALL ranks: the library all-reduces the additive buffer with NCCL (collective call)"""
        p = lib().stk_num_recorded_values(self._h)
        mean, var, rhat, ess = (np.empty(p) for _ in range(4))
        check(lib().stk_moments_nccl(self._h, comm._h, _ptr(mean), _ptr(var), _ptr(rhat), _ptr(ess)))
        return dict(mean=mean, var=var, rhat=rhat, ess=ess)

    def diag_partial(self, max_lag: int) -> np.ndarray:
        p = lib().stk_num_recorded_values(self._h)
        buf = np.empty(p * (4 + max_lag + 1))
        check(lib().stk_diag_partial(self._h, int(max_lag), _ptr(buf), buf.nbytes))
        return buf

    def sync(self) -> None:
        check(lib().stk_sync(self._h))

    def close(self):
        if self._h:
            lib().stk_chains_destroy(self._h)
            self._h = None


class NcclComm:
    """An NCCL communicator owned by the library (stk_nccl_comm_create): one per rank of a one-process-per-GPU job.  Rank 0
    makes the unique id (`NcclComm.unique_id()`), the host program ships the 128 bytes to the other ranks."""

    @staticmethod
    def unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        check(lib().stk_nccl_unique_id(buf))
        return buf.raw

    def __init__(self, ctx: Context, n_ranks: int, rank: int, unique_id: bytes):
        h = C.c_void_p()
        check(lib().stk_nccl_comm_create(ctx._h, n_ranks, rank, C.create_string_buffer(unique_id, 128), C.byref(h)))
        self._h = h

    def close(self):
        if self._h:
            lib().stk_nccl_comm_destroy(self._h)
            self._h = None


def diag_finish(buf: np.ndarray, P: int, n_draws: int, max_lag: int):
    mean, var, rhat, ess = (np.empty(P) for _ in range(4))
    buf = np.ascontiguousarray(buf, dtype=np.float64)
    check(lib().stk_diag_finish(_ptr(buf), P, n_draws, max_lag, _ptr(mean), _ptr(var), _ptr(rhat), _ptr(ess)))
    return dict(mean=mean, var=var, rhat=rhat, ess=ess)


# ---------------------------------------------------------------------------
# reference-shaped entry points
# ---------------------------------------------------------------------------
@dataclasses.dataclass
class CudaMethod:
    """Mirrors StanMethod / StanHMCEngine (Stan.hs:378-440) for the static-HMC engine of
    benchStanHMC (app/benchmark/Main.hs:248-262): num_samples draws, num_warmup discarded,
    int_time = num_leapfrog * step_size, unit_e metric, no adaptation."""
    num_chains: int = 4096
    num_samples: int = 1000
    num_warmup: int = 0
    thin: int = 1
    step_size: float = 0.01
    num_leapfrog: int = 10
    seed: int = 1
    device: int = 0
    devices: Optional[Sequence[int]] = None   # several GPUs of one box driven by this process: chains sharded over them (MultiSampler)
    kernel: str = "hmc"
    # adapt=True: Stan's `adapt engaged=1` for the static engine with a diagonal metric (adapt_diag_e_static_hmc):
    # every chain tunes its own step size (dual averaging towards `delta`) and metric over num_warmup transitions,
    # the trajectory length is int_time; kernel `hmc_adapt` of the module (csrc/stk_adapt.cuh).  Defaults are
    # CmdStan's (Stan.hs:428-440 leaves them untouched): delta 0.8, gamma 0.05, kappa 0.75, t0 10, windows 75/50/25.
    adapt: bool = False
    # engine="nuts" (+ adapt=True, num_warmup=1000, step_size=1): hmcStan's default, kernel `nuts` (max_depth as below)
    engine: str = "static"
    max_depth: int = 10
    int_time: float = 0.0            # 0 = num_leapfrog * step_size
    delta: float = 0.8
    gamma: float = 0.05
    kappa: float = 0.75
    t0: float = 10.0
    init_buffer: int = 75
    term_buffer: int = 50
    window: int = 25


class Sampler:
    """What compileCC returns (a closure state -> IO state), batched over chains."""

    def __init__(self, module_spec: str, n_chains: int, data: Optional[dict] = None, seed: int = 1,
                 device: int = 0, first_global_chain: int = 0, ctx: Optional[Context] = None):
        self.ctx = ctx or Context(device)
        self.module = Module(self.ctx, module_spec)
        self.chains = Chains(self.module, n_chains, seed, first_global_chain)
        for k, v in (data or {}).items():
            self.chains.set_data(k, v)

    def __call__(self, state: Optional[dict] = None, kernel: str | int = 0, n_transitions: int = 1) -> dict:
        """runCC semantics: upload `state` (if given), run transitions, download the new state."""
        if state:
            for k, v in state.items():
                self.chains.set_state(k, v, broadcast=(np.ndim(v) < 2 and np.size(v) != self.chains.n))
        self.chains.step(kernel, n_transitions)
        return {k: self.chains.get_state(k) for k, f in self.module.fields.items()
                if f["state"] and not str(f["len"]).startswith("@")}

    def simulate(self) -> dict:
        self.chains.simulate()
        self.chains.sync()
        return {k: self.chains.get_state(k) for k, f in self.module.fields.items() if f["state"]}


class CompiledKernel:
    """What `compileCUDA :: (s -> P t) -> IO (Int -> [s] -> IO [t])` returns (haskell/Language/CUDA.hs; compileCC's closure,
    CC.hs:465-498, batched): call it with n and one input tuple per chain, get one output tuple per chain back.  The values
    travel as SoA state fields (stochaskell_b200/layout.py), not as text; the module and its chains stay resident."""

    def __init__(self, ir_text: str, input_types: Sequence[str], device: int = 0, seed: int = 1, ctx: Optional[Context] = None,
                 cache_dir: Optional[str] = None, **options):
        self.types = list(input_types)
        self.ctx = ctx or Context(device)
        self.module = Module(self.ctx, ir_to_cuda(ir_text, self.types, **options), cache_dir=cache_dir)
        self.seed = seed
        self.chains: Optional[Chains] = None
        self._caps: Dict[str, int] = {}

    def __call__(self, n: int, xs: Sequence[Sequence]) -> list:
        from .layout import flatten_input, unflatten_output
        C = len(xs)
        if self.chains is None or self.chains.n != C:
            self.chains = Chains(self.module, C, self.seed)
            self._caps = {}
        fields = {}
        for k, t in enumerate(self.types):
            f = flatten_input(k, t, [x[k] for x in xs])
            for name, a in f.items():
                # a vector field keeps the capacity of its first upload: pad later (shorter) uploads up to it
                cap = self._caps.setdefault(name, a.shape[0])
                if a.shape[0] < cap:
                    a = np.concatenate([a, np.zeros((cap - a.shape[0], a.shape[1]), dtype=a.dtype)], axis=0)
                fields[name] = a
        for name, a in fields.items():
            self.chains.set_state(name, np.ascontiguousarray(a))
        self.chains.step("step", n)
        got = {name: self.chains.get_state(name) for name in fields}
        outs = [unflatten_output(k, t, got) for k, t in enumerate(self.types)]
        return [tuple(o[c] for o in outs) for c in range(C)]


class MultiSampler:
    """One job over several devices (CudaMethod.devices): the chains are sharded contiguously, shard d owning the
    global chains [d C/D, (d+1) C/D) on its own context, the data is replicated, launches are issued to every device before
    any is waited for, and the only exchange is the diagnostics buffer (stk_moments_multi).  Philox counters carry the
    GLOBAL chain index, so the draws are bit-identical to a single-device run of the same chain count."""

    def __init__(self, module_spec: str, n_chains: int, data: Optional[dict] = None, seed: int = 1,
                 devices: Sequence[int] = (0,), first_global_chain: int = 0):
        D = len(devices)
        if D < 1 or n_chains < D:
            raise StkError(-1, "need at least one device and one chain per device")
        self.devices = list(devices)
        self.bounds = [n_chains * d // D for d in range(D + 1)]
        self.shards = [Sampler(module_spec, self.bounds[d + 1] - self.bounds[d], data, seed=seed, device=dev,
                               first_global_chain=first_global_chain + self.bounds[d]) for d, dev in enumerate(self.devices)]
        self.module = self.shards[0].module
        self.n = n_chains

    def set_param(self, name, v):
        for s in self.shards:
            s.chains.set_param(name, v)

    def set_state(self, name, value, broadcast=False):
        a = np.asarray(value)
        for d, s in enumerate(self.shards):
            s.chains.set_state(name, a if broadcast else np.ascontiguousarray(a[..., self.bounds[d]:self.bounds[d + 1]]), broadcast=broadcast)

    def get_state(self, name):
        return np.concatenate([s.chains.get_state(name) for s in self.shards], axis=-1)

    def step(self, kernel, n=1):
        for s in self.shards:            # stk_step returns once the launches are queued: the devices run concurrently
            s.chains.step(kernel, n)

    def simulate(self):
        for s in self.shards:
            s.chains.simulate()

    def sync(self):
        for s in self.shards:
            s.chains.sync()

    def record(self, n_slots, thin=1):
        for s in self.shards:
            s.chains.record(n_slots, thin)

    def draws(self):
        return np.concatenate([s.chains.draws() for s in self.shards], axis=-1)

    def logp(self):
        return np.concatenate([s.chains.logp() for s in self.shards])

    def moments(self):
        """R-hat / ESS over ALL shards (stk_moments_multi)"""
        P = lib().stk_num_recorded_values(self.shards[0].chains._h)
        mean, var, rhat, ess = (np.empty(P) for _ in range(4))
        hs = (C.c_void_p * len(self.shards))(*[s.chains._h for s in self.shards])
        check(lib().stk_moments_multi(hs, len(self.shards), _ptr(mean), _ptr(var), _ptr(rhat), _ptr(ess)))
        return dict(mean=mean, var=var, rhat=rhat, ess=ess)


def compileCUDA(module_spec: str, n_chains: int = 1, data: Optional[dict] = None, input_types: Optional[Sequence[str]] = None, **kw):
    """compileCC analogue (CC.hs:465-498).  Given a kernel `s -> P t` as IR text and the types of its inputs it returns the
    closure `n -> [s] -> [t]` (CompiledKernel: values in, values out, n fused transitions, one chain per input); given a
    module ('builtin:<name>', CUDA text, a .cubin) it returns a Sampler over `n_chains` chains."""
    if input_types is not None and is_ir_text(module_spec):
        return CompiledKernel(module_spec, input_types, **kw)
    return Sampler(module_spec, n_chains, data, **kw)


def runCUDA(module_spec: str, state: dict, n_chains: int = 1, data: Optional[dict] = None, **kw) -> dict:
    """runCC analogue (CC.hs:459-462): compile, run one transition, return the new state."""
    return compileCUDA(module_spec, n_chains, data, **kw)(state)


def is_ir_text(spec: str) -> bool:
    """a program in showPBlock syntax (as opposed to 'builtin:<name>', a .cubin path or CUDA source text)"""
    if spec.startswith("builtin:") or spec.endswith(".cubin") or "__global__" in spec:
        return False
    t = spec.lstrip()
    return t.startswith(("let", "do ", "\\", "[runStep]", "/*")) and ("<-" in spec or " in " in spec)


def simulateCUDA(program_ir: str, n_draws: int = 1, seed: int = 1, device: int = 0, ctx: Optional[Context] = None,
                 input_types: Sequence[str] = (), **options) -> list:
    """`simulate :: P t -> IO t` (Program.hs:1038) for a program given as IR text, `n_draws` independent draws at once (one chain
    each; chain c draws from the Philox streams of global chain c).  Returns the program's returned values in order: a scalar as
    an array [n_draws], a vector as a list of n_draws 1-D arrays, a matrix as a list of n_draws 2-D arrays.  The same model text
    lowers to the HMC module through `model_to_module` / `hmcCUDA`: data and initial values of the reference's benchmarks
    (`simulate (birats n t)`, app/benchmark/Main.hs:220-226) come from the model itself."""
    c = ctx or Context(device)
    m = Module(c, ir_to_cuda(program_ir, input_types, **options))
    ch = Chains(m, n_draws, seed)
    ch.simulate(); ch.sync()
    out = []
    k = 0
    while f"out{k}" in m.fields:
        a = ch.get_state(f"out{k}")
        if f"out{k}_rows" in m.fields:
            r, cc = ch.get_state(f"out{k}_rows")[0], ch.get_state(f"out{k}_cols")[0]
            out.append([a[:r[i] * cc[i], i].reshape(r[i], cc[i]) for i in range(n_draws)])
        elif f"out{k}_len" in m.fields:
            ln = ch.get_state(f"out{k}_len")[0]
            out.append([a[:ln[i], i].copy() for i in range(n_draws)])
        else:
            out.append(a[0].copy())
        k += 1
    ch.close()
    return out


def densityCUDA(program_ir: str, points: Dict[str, np.ndarray], device: int = 0, ctx: Optional[Context] = None, **options) -> np.ndarray:
    """`density'` (Program.hs:940-944) of a program at MANY points at once: `points` maps every PNode id to an array whose first
    axis indexes the points (scalar PNode: [n]; vector: [n][len]; matrix: [n][rows][cols]); returns the joint log-density of each
    point [n].  The program is lowered with every PNode given per chain (`joint=1`, csrc/stkc_grad.inc): one chain = one point,
    long vectors are summed by the sliced sum loops of the gradient lowering, the values are read from HBM coalesced over points."""
    ids = list(points.keys())
    n = len(np.asarray(points[ids[0]]))
    src = ir_to_cuda(program_ir, [], mode="density", given=",".join(ids), perchain=",".join(ids), joint=1, **options)
    c = ctx or Context(device)
    m = Module(c, src)
    ch = Chains(m, n, 1)
    for k in ids:
        v = np.asarray(points[k])
        if v.shape[0] != n:
            raise StkError(-1, f"{k}: {v.shape[0]} points, expected {n}")
        ch.set_state(k, np.ascontiguousarray(v.reshape(n, -1).T, dtype=m.fields[k]["dtype"]))      # [len][n]: element-major, point-minor
    lp = ch.logp()
    ch.close()
    return lp


def model_to_module(model_ir: str, given: Sequence[str], name: str = "ir_model", **options) -> str:
    """`stanProgram` analogue (Stan.hs:325-344): a conditioned model `P t` -> CUDA source of a module with logp / logp_grad
    entries and a static-HMC kernel (csrc/stkc_grad.inc); `given` = the ids of the PNodes that are conditioned on"""
    return ir_to_cuda(model_ir, [], name=name, mode="density", given=",".join(given), **options)


def constrain_draws(descriptor: str, draws: np.ndarray) -> Dict[str, np.ndarray]:
    """draws [num_samples][P][num_chains] on the unconstrained scale -> {PNode id: [num_samples][shape..][num_chains]} on the
    model's own scale, by the `qmap` lines of a generated module's descriptor (lb: lo + exp q; ub: hi - exp q; lub: logistic)"""
    out = {}
    for line in descriptor.splitlines():
        t = line.split()
        if not t or t[0] != "qmap":
            continue
        ident, off, rows, cols, lo, hi = t[1], int(t[2]), int(t[3]), int(t[4]), t[5], t[6]
        if lo == "corr":
            out[ident] = _corr_matrix_constrain(draws[:, off:off + rows * (rows - 1) // 2, :], rows)
            continue
        if lo == "cov":            # cov_matrix_constrain: K(K+1)/2 values row by row, the diagonal through exp, W = L L'
            K = rows
            y = draws[:, off:off + K * (K + 1) // 2, :]
            L = np.zeros((y.shape[0], K, K, y.shape[2]))
            i = 0
            for m in range(K):
                for n in range(m):
                    L[:, m, n] = y[:, i]; i += 1
                L[:, m, m] = np.exp(y[:, i]); i += 1
            out[ident] = np.einsum("sikc,sjkc->sijc", L, L)
            continue
        n = max(1, rows) * max(1, cols)
        x = draws[:, off:off + n, :]
        if lo != "-" and hi == "-":
            x = float(lo) + np.exp(x)
        elif lo == "-" and hi != "-":
            x = float(hi) - np.exp(x)
        elif lo != "-" and hi != "-":
            x = float(lo) + (float(hi) - float(lo)) / (1.0 + np.exp(-x))
        if cols:            # a matrix block is column-major in q (Stan's order): [S][cols][rows][C] -> [S][rows][cols][C]
            out[ident] = x.reshape(draws.shape[0], cols, rows, draws.shape[2]).transpose(0, 2, 1, 3)
        else:
            out[ident] = x.reshape((draws.shape[0],) + ((rows,) if rows else ()) + (draws.shape[2],))
    return out


def unconstrain_values(descriptor: str, values: Dict[str, np.ndarray]) -> np.ndarray:
    """the inverse of `constrain_draws` for ONE point: {PNode id: value on the model's scale} -> q[P] (what CmdStan does with the
    `init=` file of hmcStanInit, Stan.hs:498-500: lb / ub / lub free transforms, matrices column-major, a corr_matrix through
    corr_matrix_free: Cholesky factor -> canonical partial correlations -> atanh)"""
    out = []
    for line in descriptor.splitlines():
        t = line.split()
        if not t or t[0] != "qmap":
            continue
        ident, off, rows, cols, lo, hi = t[1], int(t[2]), int(t[3]), int(t[4]), t[5], t[6]
        if ident not in values:
            raise StkError(-1, f"initial value of {ident} missing")
        x = np.asarray(values[ident], dtype=np.float64)
        if lo == "corr":
            if x.shape != (rows, rows):
                raise StkError(-1, f"initial value of {ident} must be a {rows} x {rows} correlation matrix")
            L = np.linalg.cholesky(x)
            z = []
            for j in range(rows - 1):
                for i in range(j + 1, rows):
                    z.append(L[i, j] / np.sqrt(1.0 - np.sum(L[i, :j] ** 2)))
            q = np.arctanh(np.array(z))
        elif lo == "cov":
            if x.shape != (rows, rows):
                raise StkError(-1, f"initial value of {ident} must be a {rows} x {rows} covariance matrix")
            L = np.linalg.cholesky(x)
            q = np.array([np.log(L[m, n]) if m == n else L[m, n] for m in range(rows) for n in range(m + 1)])
        else:
            want = (rows, cols) if cols else (rows,) if rows else ()
            if x.shape != want:
                raise StkError(-1, f"initial value of {ident} has shape {x.shape}, the model says {want}")
            q = (x.T if cols else x).reshape(-1)
            if lo != "-" and hi == "-":
                q = np.log(q - float(lo))
            elif lo == "-" and hi != "-":
                q = np.log(float(hi) - q)
            elif lo != "-" and hi != "-":
                u = (q - float(lo)) / (float(hi) - float(lo))
                q = np.log(u / (1.0 - u))
        assert off == sum(len(o) for o in out)
        out.append(np.atleast_1d(q))
    return np.concatenate(out)


def _corr_matrix_constrain(y: np.ndarray, K: int) -> np.ndarray:
    """[S][K(K-1)/2][C] unconstrained values -> [S][K][K][C] correlation matrices: tanh -> canonical partial correlations ->
    Cholesky factor -> L L' (Stan's corr_matrix_constrain / read_corr_L, the transform lkj_node of csrc/stkc_grad.inc applies)"""
    z = np.tanh(y)
    S, _, C = y.shape
    L = np.zeros((S, K, K, C))
    acc = np.ones((S, K - 1, C))
    L[:, 0, 0] = 1.0
    L[:, 1:, 0] = z[:, :K - 1]
    acc[:] = 1.0 - z[:, :K - 1] ** 2
    position, pull = 0, K - 1
    for i in range(1, K - 1):
        position += pull
        pull -= 1
        cpc = z[:, position:position + pull]
        L[:, i, i] = np.sqrt(acc[:, i - 1])
        L[:, K - pull:, i] = cpc * np.sqrt(acc[:, K - 1 - pull:])
        acc[:, K - 1 - pull:] *= 1.0 - cpc ** 2
    L[:, K - 1, K - 1] = np.sqrt(acc[:, K - 2])
    return np.einsum("sikc,sjkc->sijc", L, L)


def hmcCUDA(method: CudaMethod, module_spec: str, data: dict, init: Optional[Sequence[float]] = None,
            first_global_chain: int = 0) -> np.ndarray:
    """hmcStan / hmcStanInit analogue (Stan.hs:493-500).  `module_spec` is either a module ('builtin:<name>', CUDA text, a
    .cubin) or — the reference's own shape `hmcStan :: Int -> P t -> IO [t]` — the conditioned model itself as IR text
    (showPBlock syntax): `data` then maps the conditioned PNodes' ids to their observed values, and the module is generated
    (density + gradient + HMC kernel, csrc/stkc_grad.inc) and compiled on the fly.  Returns draws [num_samples][P][num_chains]
    on the unconstrained scale of the module's `q` field (`constrain_draws` maps them back for a generated module)."""
    if is_ir_text(module_spec):
        module_spec = model_to_module(module_spec, list(data.keys()))
    if method.devices and len(method.devices) > 1:
        return _hmc_multi(method, module_spec, data, init, first_global_chain)
    s = Sampler(module_spec, method.num_chains, data, seed=method.seed, device=(method.devices[0] if method.devices else method.device),
                first_global_chain=first_global_chain)
    ch = s.chains
    if "eps" in s.module.params:
        ch.set_param("eps", method.step_size)
        ch.set_param("L", method.num_leapfrog)
    kernel = method.kernel
    if method.adapt:
        if "stepsize" not in s.module.params:
            raise StkError(-1, f"module {module_spec} has no adaptive HMC kernel")
        kernel = "nuts" if method.engine == "nuts" else "hmc_adapt"
        if method.engine == "nuts":
            ch.set_param("max_depth", float(method.max_depth))
        for k, v in dict(stepsize=method.step_size, int_time=method.int_time or method.num_leapfrog * method.step_size,
                         num_warmup=method.num_warmup, delta=method.delta, gamma=method.gamma, kappa=method.kappa,
                         t0=method.t0, init_buffer=method.init_buffer, term_buffer=method.term_buffer,
                         window=method.window).items():
            ch.set_param(k, float(v))
    elif method.engine == "nuts":
        # no adaptation: the nuts kernel with num_warmup = 0 would re-tune the step size (init_stepsize) and then use
        # exp(x_bar) = 1 like the reference's `adapt engaged=0 ... engine=nuts` does NOT; keep the caller's step size
        raise StkError(-1, "engine='nuts' needs adapt=True (the reference's default); fixed-step NUTS is not exposed")
    method = dataclasses.replace(method, kernel=kernel)
    if isinstance(init, dict):                     # hmcStanInit's shape: values of the unobserved PNodes on the model's scale
        init = unconstrain_values(s.module.descriptor, init)
    if init is not None:
        ch.set_state("q", np.asarray(init, dtype=np.float64), broadcast=True)
    if method.num_warmup:
        ch.step(method.kernel, method.num_warmup)
    ch.record(method.num_samples, method.thin)
    ch.step(method.kernel, method.num_samples * method.thin)
    out = ch.draws()
    s.last_chains = ch
    hmcCUDA.last_sampler = s
    return out


def _hmc_multi(method: CudaMethod, module_spec: str, data: dict, init, first_global_chain: int) -> np.ndarray:
    """hmcCUDA over several devices of one box (static engine): the chains sharded over CudaMethod.devices"""
    if method.adapt or method.engine != "static":
        raise StkError(-1, "multi-device hmcCUDA supports the static engine; run adaptive engines one process per GPU")
    ms = MultiSampler(module_spec, method.num_chains, data, seed=method.seed, devices=method.devices, first_global_chain=first_global_chain)
    ms.set_param("eps", method.step_size); ms.set_param("L", method.num_leapfrog)
    if init is not None:
        ms.set_state("q", np.asarray(init, dtype=np.float64), broadcast=True)
    if method.num_warmup:
        ms.step(method.kernel, method.num_warmup)
    ms.record(method.num_samples, method.thin)
    ms.step(method.kernel, method.num_samples * method.thin)
    out = ms.draws()
    hmcCUDA.last_sampler = ms
    return out


# ---------------------------------------------------------------------------
# runStan-shaped entry point: R-dump files in, Stan CSV files out (Stan.hs:443-490)
# ---------------------------------------------------------------------------
# How a builtin module's fields map onto the identifiers of the Stan program Language.Stan generates for the same
# model (data block / parameters block), and the constraining transform of its unconstrained `q`.
STAN_BINDINGS = {
    # linreg.ipynb cell 7 (vector w): data x_stan_0_0 (X), x_stan_0_4 (y); parameters x_stan_0_1 (b), x_stan_0_2 (w),
    # x_stan_0_3 (sigma, <lower=0>: q holds log sigma)
    "builtin:linreg": dict(
        data={"X": "x_stan_0_0", "y": "x_stan_0_4"},
        params=lambda d: [("x_stan_0_1", (), slice(0, 1)), ("x_stan_0_2", (d["X"].shape[1],), slice(1, 1 + d["X"].shape[1])),
                          ("x_stan_0_3", (), slice(1 + d["X"].shape[1], 2 + d["X"].shape[1]))],
        constrain=lambda q: np.concatenate([q[:-1], np.exp(q[-1:])], axis=0),
        unconstrain=lambda v: np.concatenate([v[:-1], np.log(v[-1:])], axis=0),
    ),
}


def runStanCUDA(method: CudaMethod, module_spec: str, data_file: str, init_file: Optional[str] = None,
                output_file: Optional[str] = None, csv_chains: int = 4) -> np.ndarray:
    """`runStan` with the process boundary of the reference kept (Stan.hs:443-490): data and initial values come from the
    R-dump files `stanDump` writes, the draws of the first `csv_chains` chains go to Stan CSV files
    (`<output_file stem>_<chain>.csv`, CmdStan's per-chain layout: lp__, the constrained parameters) that `stanRead`
    and stansummary read; returns the constrained draws [num_samples][P][num_chains]."""
    from .io import rdump_load, stan_columns, write_stan_csv
    if is_ir_text(module_spec):
        return _run_stan_ir(method, module_spec, data_file, init_file, output_file, csv_chains)
    b = STAN_BINDINGS.get(module_spec)
    if b is None:
        raise StkError(-1, f"no Stan identifier binding for {module_spec}")
    env = rdump_load(data_file)
    data = {}
    for field, ident in b["data"].items():
        if ident not in env:
            raise StkError(-1, f"{data_file}: variable {ident} (field {field}) is missing")
        data[field] = np.ascontiguousarray(env[ident], dtype=np.float64)
    layout = b["params"](data)
    init = None
    if init_file:
        ienv = rdump_load(init_file)
        P = layout[-1][2].stop
        v = np.zeros(P)
        for ident, shape, sl in layout:
            if ident not in ienv:
                raise StkError(-1, f"{init_file}: parameter {ident} is missing")
            v[sl] = np.asarray(ienv[ident], dtype=np.float64).flatten(order="F")
        init = b["unconstrain"](v)
    draws = hmcCUDA(method, module_spec, data, init=init)
    ch = hmcCUDA.last_sampler.chains
    cons = np.stack([b["constrain"](draws[t]) for t in range(draws.shape[0])]) if draws.size else draws
    if output_file:
        names = ["lp__"] + [c for ident, shape, _ in layout for c in stan_columns(ident, shape)]
        stem = output_file[:-4] if output_file.endswith(".csv") else output_file
        eps = ch.get_state("ad")[0] if method.adapt else None
        minv = ch.get_state("minv") if method.adapt else None
        # lp__ per draw: the module's logp entry on the recorded draws (joint log-density on the unconstrained scale with
        # all constants kept, as Program.hs `lpdf`; Stan's own lp__ drops the constants of `~` statements)
        q_last = ch.get_state("q")
        lps = np.stack([(ch.set_state("q", np.ascontiguousarray(draws[t])), ch.logp())[1] for t in range(draws.shape[0])]) \
            if draws.shape[0] else np.zeros((0, draws.shape[2]))
        ch.set_state("q", q_last)
        for c in range(min(csv_chains, draws.shape[2])):
            lp = lps[:, c:c + 1]
            comment = f"model = {module_spec}\nalgorithm = hmc, engine = {method.engine}, chains = {method.num_chains}, chain = {c + 1}"
            if eps is not None:
                comment += (f"\nAdaptation terminated\nStep size = {eps[c]:.6g}\nDiagonal elements of inverse mass matrix:\n"
                            + ", ".join(f"{x:.6g}" for x in minv[:, c]))
            write_stan_csv(f"{stem}_{c + 1}.csv", names, np.concatenate([lp, cons[:, :, c]], axis=1), comment=comment)
    return cons


def _run_stan_ir(method: CudaMethod, model_ir: str, data_file: str, init_file: Optional[str], output_file: Optional[str],
                 csv_chains: int) -> np.ndarray:
    """runStanCUDA for a model given as IR text: the identifiers of the Stan program `stanProgram` generates ARE the PNode ids
    (stanId, Stan.hs:43-57), so no binding table is needed — the PNodes that appear in the data file are the conditioned ones, the
    init file holds the others on the model's scale, and the CSV columns are the qmap blocks in program order, elements column-major
    (a corr_matrix as its K x K entries), exactly CmdStan's layout."""
    import re as _re
    from .io import rdump_load, stan_columns, write_stan_csv
    env = rdump_load(data_file)
    ids = _re.findall(r"^\s*(?:in do\s+)?(\w+)\s*<-", model_ir, _re.M)
    given = [i for i in ids if i in env]
    if not given:
        raise StkError(-1, f"{data_file}: none of the program's PNodes {ids} is in the data file")
    data = {i: np.ascontiguousarray(env[i], dtype=np.float64) for i in given}
    init = None
    if init_file:
        ienv = rdump_load(init_file)
        init = {i: np.asarray(ienv[i], dtype=np.float64) for i in ids if i in ienv}
    draws = hmcCUDA(method, model_ir, data, init=init)
    s = hmcCUDA.last_sampler
    ch = s.chains
    vals = constrain_draws(s.module.descriptor, draws)
    names, cols = ["lp__"], []
    for ident, v in vals.items():                      # qmap order = program order = the parameters block
        shape = v.shape[1:-1]
        names += stan_columns(ident, shape)
        S, Cn = v.shape[0], v.shape[-1]
        cols.append(v.reshape(S, -1, Cn) if len(shape) < 2 else v.transpose(0, 2, 1, 3).reshape(S, -1, Cn))      # column-major elements
    cons = np.concatenate(cols, axis=1) if cols else draws
    if output_file:
        stem = output_file[:-4] if output_file.endswith(".csv") else output_file
        eps = ch.get_state("ad")[0] if method.adapt else None
        minv = ch.get_state("minv") if method.adapt else None
        q_last = ch.get_state("q")
        lps = np.stack([(ch.set_state("q", np.ascontiguousarray(draws[t])), ch.logp())[1] for t in range(draws.shape[0])]) \
            if draws.shape[0] else np.zeros((0, draws.shape[2]))
        ch.set_state("q", q_last)
        for c in range(min(csv_chains, draws.shape[2])):
            comment = f"model = {s.module.name} (generated from IR)\nalgorithm = hmc, engine = {method.engine}, chains = {method.num_chains}, chain = {c + 1}"
            if eps is not None:
                comment += (f"\nAdaptation terminated\nStep size = {eps[c]:.6g}\nDiagonal elements of inverse mass matrix:\n"
                            + ", ".join(f"{x:.6g}" for x in minv[:, c]))
            write_stan_csv(f"{stem}_{c + 1}.csv", names, np.concatenate([lps[:, c:c + 1], cons[:, :, c]], axis=1), comment=comment)
    return cons
