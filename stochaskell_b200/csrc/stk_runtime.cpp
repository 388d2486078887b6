// stk_runtime.cpp — host runtime of libstk_cuda.so (see include/stk_cuda.h).
//
// Plays the role of compileCC's closure (stochaskell/src/Language/CC.hs:465-498)
// for the CUDA backend: compile-once (NVRTC, SHA1-keyed cubin cache like
// CC.hs:471-479), then keep the module and the chains' state resident on the
// GPU and run thousands of transitions per call instead of one process per
// transition.
//
// CUDA runtime API only (static cudart: the library loads on a machine with
// no driver; every compute entry point then fails loudly with STK_ERR_CUDA).
// No torch types, no CPU fallback.
#include "../../include/stk_cuda.h"
#include "stk_abi.h"

#include <cuda_runtime_api.h>
#include <nvrtc.h>

#include <dlfcn.h>
#include <execinfo.h>
#include <signal.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

namespace {

thread_local std::string g_err;
std::atomic<long long> g_launches{0};

void segv_handler(int sig) {
  void* bt[64]; int n = backtrace(bt, 64);
  fprintf(stderr, "[stk] signal %d, native backtrace:\n", sig);
  backtrace_symbols_fd(bt, n, 2);
  _exit(139);
}
bool init_trace() {
  bool on = getenv("STK_TRACE") != nullptr;
  if (on) signal(SIGSEGV, segv_handler);
  return on;
}
bool g_trace = init_trace();
#define TRACE(...) do { if (g_trace) { fprintf(stderr, "[stk] " __VA_ARGS__); fputc('\n', stderr); fflush(stderr); } } while (0)

int fail(int code, const char* fmt, ...) {
  char buf[4096];
  va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
  g_err = buf;
  return code;
}
#define CU(call)                                                                   \
  do {                                                                             \
    cudaError_t e_ = (call);                                                       \
    if (e_ != cudaSuccess)                                                         \
      return fail(STK_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

// ---- SHA1 (cache key, as CC.hs:472 `SHA1.hash code`) ------------------------
struct Sha1 {
  uint32_t h[5] = {0x67452301u, 0xEFCDAB89u, 0x98BADCFEu, 0x10325476u, 0xC3D2E1F0u};
  static uint32_t rol(uint32_t v, int b) { return (v << b) | (v >> (32 - b)); }
  void block(const uint8_t* p) {
    uint32_t w[80];
    for (int i = 0; i < 16; ++i) w[i] = (p[4 * i] << 24) | (p[4 * i + 1] << 16) | (p[4 * i + 2] << 8) | p[4 * i + 3];
    for (int i = 16; i < 80; ++i) w[i] = rol(w[i - 3] ^ w[i - 8] ^ w[i - 14] ^ w[i - 16], 1);
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4];
    for (int i = 0; i < 80; ++i) {
      uint32_t f, k;
      if (i < 20) { f = (b & c) | (~b & d); k = 0x5A827999u; }
      else if (i < 40) { f = b ^ c ^ d; k = 0x6ED9EBA1u; }
      else if (i < 60) { f = (b & c) | (b & d) | (c & d); k = 0x8F1BBCDCu; }
      else { f = b ^ c ^ d; k = 0xCA62C1D6u; }
      uint32_t t = rol(a, 5) + f + e + k + w[i];
      e = d; d = c; c = rol(b, 30); b = a; a = t;
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e;
  }
  static std::string hex(const std::string& s) {
    Sha1 x;
    std::vector<uint8_t> m(s.begin(), s.end());
    uint64_t bits = (uint64_t)m.size() * 8;
    m.push_back(0x80);
    while (m.size() % 64 != 56) m.push_back(0);
    for (int i = 7; i >= 0; --i) m.push_back((uint8_t)(bits >> (8 * i)));
    for (size_t i = 0; i < m.size(); i += 64) x.block(&m[i]);
    char out[41];
    for (int i = 0; i < 5; ++i) snprintf(out + 8 * i, 9, "%08x", x.h[i]);
    return std::string(out, 40);
  }
};

bool file_exists(const std::string& p) { struct stat st; return stat(p.c_str(), &st) == 0 && S_ISREG(st.st_mode); }
bool read_file(const std::string& p, std::string* out) {
  std::ifstream f(p, std::ios::binary);
  if (!f) return false;
  std::ostringstream ss; ss << f.rdbuf(); *out = ss.str();
  return true;
}
int mkdirs(const std::string& p) {
  std::string cur;
  for (size_t i = 0; i <= p.size(); ++i) {
    if (i == p.size() || p[i] == '/') {
      if (!cur.empty() && cur != "/") { if (mkdir(cur.c_str(), 0777) != 0 && errno != EEXIST) return -1; }
    }
    if (i < p.size()) cur.push_back(p[i]);
  }
  return 0;
}
std::string lib_dir() {
  Dl_info info;
  if (dladdr((void*)&stk_last_error, &info) && info.dli_fname) {
    std::string p(info.dli_fname);
    size_t s = p.rfind('/');
    return s == std::string::npos ? std::string(".") : p.substr(0, s);
  }
  return ".";
}

struct FieldDesc {
  std::string name; bool is_state; bool is_i32; long long len; std::string len_ref; bool sorted = false;
  bool is_work = false; int esize = 8; std::string len_expr;   // len_expr: set when <len> is not a plain integer
};
struct StageDesc { int kind = 0; std::string symbol; int block = 0; std::string gx, gy; long long smem = 0; bool coop = false; cudaKernel_t k = nullptr; };
struct PlanDesc { std::string name, role, loop_param; std::vector<StageDesc> stages; };
struct KernelDesc {
  std::string name, symbol;
  int block = 0, cpb = 1; long long smem_fixed = 0; std::string smem_field; long long smem_per = 0;
  cudaKernel_t k = nullptr;
  std::string pre_symbol; int pre_block = 0; long long pre_per = 0; cudaKernel_t pre_k = nullptr;
  std::string ph_symbol, ph_field; int ph_block = 0, ph_cpb = 1; long long ph_items = 1, ph_smem = 0; cudaKernel_t ph_k = nullptr;
  int plan = -1;   // index into stk_module::plans when this transition kernel / entry is a plan
};

}  // namespace

struct stk_ctx {
  int device; cudaStream_t stream; int sm_count = 0;
  cudaLibrary_t diag_lib = nullptr; cudaKernel_t k_moments = nullptr, k_acov = nullptr;
};

struct stk_module {
  stk_ctx* ctx = nullptr;
  cudaLibrary_t lib = nullptr;
  std::string desc, name;
  int nparam = 0;
  std::vector<FieldDesc> fields;
  std::vector<std::pair<std::string, double>> params;
  std::vector<int> record;
  std::vector<KernelDesc> kernels;
  KernelDesc entry[3];  // logp, logp_grad, simulate
  bool has_entry[3] = {false, false, false};
  std::vector<PlanDesc> plans;
  int prepare_plan = -1;
  std::string nparam_expr;
};

struct stk_chains {
  stk_module* mod = nullptr;
  long long n = 0, chain0 = 0;
  unsigned long long seed = 0, step = 0;
  std::vector<void*> dptr;
  std::vector<long long> flen;
  std::vector<double> par;
  double* draws = nullptr; long long slots = 0, thin = 1, recorded = 0, nrec = 0;
  double* out_lp = nullptr; double* out_grad = nullptr;
  unsigned long long* accept = nullptr;
  double* scratch = nullptr; size_t scratch_bytes = 0;
  bool dirty_data = true, dirty_state = true;   // plans: run `prepare` / `init` stages before the next launch
  long long grad_cap = 0;
  cudaStream_t last = nullptr;                  // caller stream of the most recent stk_step / stk_simulate (nullptr = the ctx stream)
};

namespace {

int field_index(const stk_module* m, const std::string& n) {
  for (size_t i = 0; i < m->fields.size(); ++i) if (m->fields[i].name == n) return (int)i;
  return -1;
}

int parse_kernel_line(std::istringstream& ls, KernelDesc* k) {
  std::string sf;
  if (!(ls >> k->name >> k->symbol >> k->block >> k->cpb >> k->smem_fixed >> sf >> k->smem_per)) return -1;
  k->smem_field = (sf == "-") ? "" : sf;
  std::string opt;
  while (ls >> opt) {
    if (opt == "pre") { if (!(ls >> k->pre_symbol >> k->pre_block >> k->pre_per) || k->pre_block <= 0 || k->pre_per <= 0) return -1; }
    else if (opt == "phase") {
      if (!(ls >> k->ph_symbol >> k->ph_block >> k->ph_cpb >> k->ph_field >> k->ph_items >> k->ph_smem) || k->ph_block <= 0 || k->ph_cpb <= 0 || k->ph_items <= 0) return -1;
    }
    else return -1;
  }
  return 0;
}

int parse_desc(stk_module* m) {
  std::istringstream in(m->desc);
  std::string line;
  bool header = false;
  while (std::getline(in, line)) {
    std::istringstream ls(line);
    std::string kw;
    TRACE("desc line: %s", line.c_str());
    if (!(ls >> kw)) continue;
    if (kw == "stk-module") { int v; ls >> v; if (v != 1) return fail(STK_ERR_ARG, "descriptor version %d unsupported", v); header = true; }
    else if (kw == "name") ls >> m->name;
    else if (kw == "nparam") {
      std::string v; ls >> v;
      if (v.find_first_not_of("0123456789") == std::string::npos) m->nparam = atoi(v.c_str()); else m->nparam_expr = v;
    }
    else if (kw == "field") {
      FieldDesc f; std::string kind, ty, len;
      if (!(ls >> f.name >> kind >> ty >> len)) return fail(STK_ERR_ARG, "bad field line: %s", line.c_str());
      f.is_state = (kind == "state"); f.is_work = (kind == "work"); f.is_i32 = (ty == "i32");
      f.esize = (ty == "i32" || ty == "f32") ? 4 : (ty == "bf16" || ty == "f16") ? 2 : 8;
      if (len.find_first_not_of("0123456789") == std::string::npos) f.len = atoll(len.c_str());
      else {
        f.len = 0; f.len_expr = len;
        if (len[0] == '@' && len.find_first_of("*^") == std::string::npos) f.len_ref = len.substr(1);
      }
      { std::string attr; while (ls >> attr) if (attr == "sorted") f.sorted = true; }
      if ((int)m->fields.size() >= STK_MAX_FIELDS) return fail(STK_ERR_ARG, "too many fields");
      m->fields.push_back(f);
    } else if (kw == "param") {
      std::string n; double v; ls >> n >> v;
      if ((int)m->params.size() >= STK_MAX_PARAMS) return fail(STK_ERR_ARG, "too many params");
      m->params.push_back({n, v});
    } else if (kw == "record") {
      std::string n;
      while (ls >> n) { int i = field_index(m, n); if (i < 0) return fail(STK_ERR_ARG, "record: unknown field %s", n.c_str()); m->record.push_back(i); }
    } else if (kw == "kernel") {
      KernelDesc k; if (parse_kernel_line(ls, &k)) return fail(STK_ERR_ARG, "bad kernel line: %s", line.c_str());
      m->kernels.push_back(k);
    } else if (kw == "plan") {
      PlanDesc p;
      if (!(ls >> p.name >> p.role >> p.loop_param)) return fail(STK_ERR_ARG, "bad plan line: %s", line.c_str());
      if (p.loop_param == "-") p.loop_param.clear();
      int pi = (int)m->plans.size();
      KernelDesc k; k.name = p.name; k.plan = pi;
      if (p.role == "kernel") m->kernels.push_back(k);
      else if (p.role == "prepare") m->prepare_plan = pi;
      else {
        int e = p.role == "logp" ? 0 : p.role == "logp_grad" ? 1 : p.role == "simulate" ? 2 : -1;
        if (e < 0) return fail(STK_ERR_ARG, "plan %s: unknown role %s", p.name.c_str(), p.role.c_str());
        m->entry[e] = k; m->has_entry[e] = true;
      }
      m->plans.push_back(p);
    } else if (kw == "stage") {
      std::string pn, kind, opt; StageDesc sd;
      if (!(ls >> pn >> kind >> sd.symbol >> sd.block >> sd.gx >> sd.gy >> sd.smem) || sd.block <= 0) return fail(STK_ERR_ARG, "bad stage line: %s", line.c_str());
      sd.kind = kind == "init" ? 0 : kind == "pre" ? 1 : kind == "body" ? 2 : kind == "post" ? 3 : -1;
      if (sd.kind < 0) return fail(STK_ERR_ARG, "stage: unknown kind %s", kind.c_str());
      while (ls >> opt) { if (opt == "coop") sd.coop = true; else return fail(STK_ERR_ARG, "stage: unknown option %s", opt.c_str()); }
      PlanDesc* pl = nullptr;
      for (auto& p : m->plans) if (p.name == pn) pl = &p;
      if (!pl) return fail(STK_ERR_ARG, "stage names unknown plan %s", pn.c_str());
      pl->stages.push_back(sd);
    } else if (kw == "entry") {
      KernelDesc k; if (parse_kernel_line(ls, &k)) return fail(STK_ERR_ARG, "bad entry line: %s", line.c_str());
      int e = k.name == "logp" ? 0 : k.name == "logp_grad" ? 1 : k.name == "simulate" ? 2 : -1;
      if (e < 0) return fail(STK_ERR_ARG, "unknown entry %s", k.name.c_str());
      m->entry[e] = k; m->has_entry[e] = true;
    }
  }
  if (!header) return fail(STK_ERR_ARG, "module has no stk-module header");
  return STK_OK;
}

// <len> / <gx> expressions: product (or one ceiling quotient) of terms <int> | sms | chains[^k] | @field[^k];
// 0 when a referenced field has no size yet, -1 on a syntax error
long long eval_term(const stk_chains* c, const std::string& t) {
  std::string base = t; long long round = 0;
  size_t h = t.find('^');
  if (h != std::string::npos) { base = t.substr(0, h); round = atoll(t.c_str() + h + 1); if (round <= 0) return -1; }
  long long v;
  if (base == "chains") v = c->n;
  else if (base == "sms") v = c->mod->ctx->sm_count;
  else if (!base.empty() && base[0] == '@') { int r = field_index(c->mod, base.substr(1)); if (r < 0) return -1; v = c->flen[r]; }
  else if (!base.empty() && base.find_first_not_of("0123456789") == std::string::npos) v = atoll(base.c_str());
  else return -1;
  if (round > 0 && v > 0) v = (v + round - 1) / round * round;
  return v;
}
long long eval_expr(const stk_chains* c, const std::string& e) {
  size_t sl = e.find('/');
  if (sl != std::string::npos) {
    long long a = eval_expr(c, e.substr(0, sl)), b = atoll(e.c_str() + sl + 1);
    if (a < 0 || b <= 0) return -1;
    return (a + b - 1) / b;
  }
  long long v = 1; size_t i = 0;
  while (i <= e.size()) {
    size_t j = e.find('*', i);
    if (j == std::string::npos) j = e.size();
    long long t = eval_term(c, e.substr(i, j - i));
    if (t < 0) return -1;
    v *= t; i = j + 1;
  }
  return v;
}
long long state_len(const stk_chains* c, int f) {
  const FieldDesc& fd = c->mod->fields[f];
  if (fd.len > 0) return fd.len;
  if (!fd.len_expr.empty()) { long long v = eval_expr(c, fd.len_expr); return v > 0 ? v : 0; }
  return c->flen[f];
}
size_t elem_size(const FieldDesc& f) { return (size_t)f.esize; }
long long num_params(const stk_chains* c) {
  if (c->mod->nparam_expr.empty()) return c->mod->nparam;
  long long v = eval_expr(c, c->mod->nparam_expr);
  return v > 0 ? v : 0;
}

// allocate deferred-size state fields whose size is now known
int ensure_fields(stk_chains* c) {
  for (size_t f = 0; f < c->mod->fields.size(); ++f) {
    const FieldDesc& fd = c->mod->fields[f];
    if (!(fd.is_state || fd.is_work) || c->dptr[f]) continue;
    long long L = state_len(c, (int)f);
    if (L <= 0) continue;
    size_t bytes = (size_t)L * (fd.is_state ? (size_t)c->n : 1) * elem_size(fd);
    CU(cudaMalloc(&c->dptr[f], bytes));
    CU(cudaMemsetAsync(c->dptr[f], 0, bytes, c->mod->ctx->stream));
    // allocation is rare (once per field): finish the zero-fill here, so that a launch on a caller's stream cannot race it
    CU(cudaStreamSynchronize(c->mod->ctx->stream));
    c->flen[f] = L;
  }
  return STK_OK;
}

// every stream that may still be running this chain set's kernels: the context's own and the caller stream of the last launch
int sync_chains(stk_chains* c) {
  CU(cudaStreamSynchronize(c->mod->ctx->stream));
  if (c->last && c->last != c->mod->ctx->stream) CU(cudaStreamSynchronize(c->last));
  return STK_OK;
}

// generated (stkc) modules carry an i32 work field `err` that a kernel sets when a per-thread vector would overflow its
// storage (STK_VECCAP) or a state field's capacity: reported, after the launch has finished, by every call that synchronises
int check_err(stk_chains* c) {
  int fi = field_index(c->mod, "err");
  if (fi < 0 || !c->mod->fields[fi].is_work || !c->mod->fields[fi].is_i32 || !c->dptr[fi]) return STK_OK;
  int code = 0;
  CU(cudaMemcpy(&code, c->dptr[fi], sizeof code, cudaMemcpyDeviceToHost));
  if (code == 0) return STK_OK;
  CU(cudaMemset(c->dptr[fi], 0, sizeof code));
  return fail(STK_ERR_ARG, "module %s: a kernel reported error %d (%s); the chain state is not valid — rebuild the module with a larger "
              "veccap / upload the state field with more rows", c->mod->name.c_str(), code,
              code == 1 ? "a vector outgrew the per-thread storage STK_VECCAP" : code == 2 ? "a vector outgrew the capacity of its state field" :
              code == 3 ? "a matrix outgrew the per-thread storage STK_MATCAP" :
              code == 4 ? "a program's conditions (guard) were not met in 65536 attempts" : "unknown");
}

// data fields declared with a length expression (e.g. `field X data f64 @y*10`) must have exactly that many elements
int check_data_shapes(stk_chains* c) {
  for (size_t f = 0; f < c->mod->fields.size(); ++f) {
    const FieldDesc& fd = c->mod->fields[f];
    if (fd.is_state || fd.is_work || fd.len_expr.empty() || !c->dptr[f]) continue;
    long long want = eval_expr(c, fd.len_expr);
    if (want > 0 && want != c->flen[f])
      return fail(STK_ERR_ARG, "module %s: data field %s holds %lld elements, the descriptor requires %s = %lld", c->mod->name.c_str(),
                  fd.name.c_str(), c->flen[f], fd.len_expr.c_str(), want);
  }
  return STK_OK;
}

void fill_args(stk_chains* c, long long n_steps, StkArgs* ap) {
  stk_module* m = c->mod;
  StkArgs& a = *ap;
  memset(&a, 0, sizeof a);
  a.n_chains = c->n; a.chain0 = c->chain0; a.seed = c->seed; a.step0 = c->step; a.n_steps = n_steps;
  for (size_t f = 0; f < m->fields.size(); ++f) { a.field[f] = c->dptr[f]; a.flen[f] = c->flen[f]; }
  for (size_t p = 0; p < c->par.size(); ++p) a.par[p] = c->par[p];
  a.draws = c->draws; a.draw_slots = c->slots; a.draw_thin = c->thin; a.draw_base = c->recorded; a.n_rec = c->nrec;
  a.out_lp = c->out_lp; a.out_grad = c->out_grad; a.accept = c->accept; a.scratch = c->scratch;
  a.sm_count = m->ctx->sm_count; a.stage_kind = 4;
}

// one stage of a plan
int launch_stage(stk_chains* c, StageDesc& sd, StkArgs& a, cudaStream_t st, int kind = -1) {
  long long gx = eval_expr(c, sd.gx), gy = eval_expr(c, sd.gy);
  if (gx <= 0 || gy <= 0 || gy > 65535) return fail(STK_ERR_ARG, "stage %s: grid %s x %s evaluates to %lld x %lld", sd.symbol.c_str(), sd.gx.c_str(), sd.gy.c_str(), gx, gy);
  if (sd.smem > 227 * 1024) return fail(STK_ERR_UNSUPPORTED, "stage %s needs %lld B shared memory (> 227 KB)", sd.symbol.c_str(), sd.smem);
  if (sd.smem > 48 * 1024)
    CU(cudaFuncSetAttribute((const void*)sd.k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sd.smem));
  a.stage_kind = kind >= 0 ? kind : sd.kind;
  void* params[1] = {&a};
  TRACE("stage %s grid=%lld x %lld block=%d smem=%lld coop=%d iter=%lld/%lld", sd.symbol.c_str(), gx, gy, sd.block, sd.smem, (int)sd.coop, a.iter, a.n_iter);
  if (sd.coop) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)gx, (unsigned)gy); cfg.blockDim = dim3((unsigned)sd.block);
    cfg.dynamicSmemBytes = (size_t)sd.smem; cfg.stream = st;
    cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeCooperative; at[0].val.cooperative = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    CU(cudaLaunchKernelExC(&cfg, (const void*)sd.k, params));
  } else {
    CU(cudaLaunchKernel((const void*)sd.k, dim3((unsigned)gx, (unsigned)gy), dim3((unsigned)sd.block), params, (size_t)sd.smem, st));
  }
  g_launches.fetch_add(1);
  return STK_OK;
}

// per-data-set precomputation: runs when a data field changed since the last launch
int run_prepare(stk_chains* c, cudaStream_t st) {
  stk_module* m = c->mod;
  if (!c->dirty_data) return STK_OK;
  int rc = ensure_fields(c);
  if (rc) return rc;
  if ((rc = check_data_shapes(c))) return rc;
  if (m->prepare_plan >= 0) {
    StkArgs a; fill_args(c, 0, &a);
    for (auto& sd : m->plans[m->prepare_plan].stages) if ((rc = launch_stage(c, sd, a, st, 4))) return rc;
    int fi = field_index(m, "info");
    if (fi >= 0 && m->fields[fi].is_work && c->dptr[fi]) {
      int info = 0;
      CU(cudaMemcpyAsync(&info, c->dptr[fi], sizeof info, cudaMemcpyDeviceToHost, st));
      CU(cudaStreamSynchronize(st));
      if (info != 0) return fail(STK_ERR_ARG, "module %s: prepare failed, info = %d (e.g. covariance not positive definite at pivot %d)", m->name.c_str(), info, info);
    }
  }
  c->dirty_data = false;
  return STK_OK;
}

int ensure_grad(stk_chains* c) {
  long long P = num_params(c);
  if (P > 0 && P * c->n > c->grad_cap) {
    CU(cudaStreamSynchronize(c->mod->ctx->stream));
    if (c->out_grad) { CU(cudaFree(c->out_grad)); c->out_grad = nullptr; c->grad_cap = 0; }
    CU(cudaMalloc((void**)&c->out_grad, sizeof(double) * (size_t)P * (size_t)c->n));
    c->grad_cap = P * c->n;
  }
  return STK_OK;
}

// a plan used as an entry (logp / logp_grad / simulate): its stages once, in order
int run_entry_plan(stk_chains* c, int pi, cudaStream_t st) {
  int rc = run_prepare(c, st);
  if (rc) return rc;
  if ((rc = ensure_fields(c))) return rc;
  StkArgs a; fill_args(c, 0, &a);
  for (auto& sd : c->mod->plans[pi].stages) if ((rc = launch_stage(c, sd, a, st, 4))) return rc;
  return STK_OK;
}

// a plan used as a transition kernel: n_steps transitions, host-sequenced
int run_kernel_plan(stk_chains* c, int pi, long long n_steps, cudaStream_t st) {
  stk_module* m = c->mod;
  PlanDesc& pl = m->plans[pi];
  int rc = run_prepare(c, st);
  if (rc) return rc;
  if ((rc = ensure_fields(c))) return rc;
  long long n_iter = 1;
  if (!pl.loop_param.empty()) {
    int pidx = -1;
    for (size_t i = 0; i < m->params.size(); ++i) if (m->params[i].first == pl.loop_param) pidx = (int)i;
    if (pidx < 0) return fail(STK_ERR_ARG, "plan %s: unknown loop parameter %s", pl.name.c_str(), pl.loop_param.c_str());
    n_iter = (long long)c->par[pidx];
    if (n_iter < 1) return fail(STK_ERR_ARG, "plan %s: %s must be >= 1", pl.name.c_str(), pl.loop_param.c_str());
  }
  StkArgs a; fill_args(c, 1, &a);
  a.n_iter = n_iter;
  if (c->dirty_state) {
    for (auto& sd : pl.stages) if (sd.kind == 0 && (rc = launch_stage(c, sd, a, st))) return rc;
    c->dirty_state = false;
  }
  for (long long t = 0; t < n_steps; ++t) {
    a.step0 = c->step; a.draw_base = c->recorded; a.iter = 0;
    for (auto& sd : pl.stages) if (sd.kind == 1 && (rc = launch_stage(c, sd, a, st))) return rc;
    for (long long it = 0; it < n_iter; ++it) {
      a.iter = it;
      for (auto& sd : pl.stages) if (sd.kind == 2 && (rc = launch_stage(c, sd, a, st))) return rc;
    }
    a.iter = n_iter;
    for (auto& sd : pl.stages) if (sd.kind == 3 && (rc = launch_stage(c, sd, a, st))) return rc;
    if (c->draws && c->slots > 0) {
      unsigned long long s0 = c->step, s1 = c->step + 1;
      long long add = (long long)(s1 / (unsigned long long)c->thin - s0 / (unsigned long long)c->thin);
      c->recorded = std::min(c->slots, c->recorded + add);
    }
    c->step += 1;
  }
  return STK_OK;
}

int launch(stk_chains* c, KernelDesc& k, long long n_steps, cudaStream_t st) {
  stk_module* m = c->mod;
  int rc = run_prepare(c, st);
  if (rc) return rc;
  if ((rc = ensure_fields(c))) return rc;
  StkArgs a;
  fill_args(c, n_steps, &a);
  if (k.pre_k) {   // pre-pass: one thread per (chain, transition) of this launch
    long long items = c->n * n_steps;
    long long pgrid = (items + k.pre_block - 1) / k.pre_block;
    void* pparams[1] = {&a};
    TRACE("launch pre %s grid=%lld block=%d", k.pre_symbol.c_str(), pgrid, k.pre_block);
    CU(cudaLaunchKernel((const void*)k.pre_k, dim3((unsigned)pgrid), dim3((unsigned)k.pre_block), pparams, 0, st));
    g_launches.fetch_add(1);
  }
  if (k.ph_k) {    // per-transition phase kernel on a (chain groups) x (item tiles) grid
    int f = field_index(m, k.ph_field);
    long long items = c->flen[f];
    if (items <= 0) return fail(STK_ERR_ARG, "kernel %s: field %s not set", k.name.c_str(), k.ph_field.c_str());
    dim3 pg((unsigned)((c->n + k.ph_cpb - 1) / k.ph_cpb), (unsigned)((items + k.ph_items - 1) / k.ph_items));
    if (k.ph_smem > 48 * 1024)
      CU(cudaFuncSetAttribute((const void*)k.ph_k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)k.ph_smem));
    void* pparams[1] = {&a};
    TRACE("launch phase %s grid=%u x %u block=%d", k.ph_symbol.c_str(), pg.x, pg.y, k.ph_block);
    CU(cudaLaunchKernel((const void*)k.ph_k, pg, dim3((unsigned)k.ph_block), pparams, (size_t)k.ph_smem, st));
    g_launches.fetch_add(1);
  }
  long long smem = k.smem_fixed;
  if (!k.smem_field.empty()) {
    int f = field_index(m, k.smem_field);
    if (f < 0) return fail(STK_ERR_ARG, "kernel %s: unknown smem field %s", k.name.c_str(), k.smem_field.c_str());
    if (c->flen[f] <= 0) return fail(STK_ERR_ARG, "kernel %s: data field %s not set", k.name.c_str(), k.smem_field.c_str());
    smem += c->flen[f] * k.smem_per;
  }
  if (smem > 227 * 1024) return fail(STK_ERR_UNSUPPORTED, "kernel %s needs %lld B shared memory (> 227 KB)", k.name.c_str(), smem);
  int block = k.block;
  if (smem > 48 * 1024)
    CU(cudaFuncSetAttribute((const void*)k.k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  long long grid = (c->n + k.cpb - 1) / k.cpb;
  TRACE("launch %s grid=%lld block=%d smem=%lld steps=%lld", k.symbol.c_str(), grid, block, smem, n_steps);
  void* params[1] = {&a};
  CU(cudaLaunchKernel((const void*)k.k, dim3((unsigned)grid), dim3((unsigned)block), params, (size_t)smem, st));
  g_launches.fetch_add(1);
  return STK_OK;
}

int need_device(stk_ctx* ctx) {
  if (!ctx) return fail(STK_ERR_ARG, "null ctx");
  CU(cudaSetDevice(ctx->device));
  return STK_OK;
}

}  // namespace

extern "C" {

const char* stk_last_error(void) { return g_err.c_str(); }
int stk_set_error_(int code, const char* msg) { return fail(code, "%s", msg ? msg : ""); }  // for stkc.cpp
int stk_abi_version(void) { return STK_ABI_VERSION; }
int64_t stk_launch_count(void) { return g_launches.load(); }

int stk_ctx_create(int device, stk_ctx** out) {
  if (!out) return fail(STK_ERR_ARG, "null out");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0)
    return fail(STK_ERR_CUDA, "no CUDA device available (%s); libstk_cuda has no CPU fallback",
                e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (device < 0 || device >= n) return fail(STK_ERR_ARG, "device %d out of range (0..%d)", device, n - 1);
  CU(cudaSetDevice(device));
  stk_ctx* c = new stk_ctx; c->device = device; c->stream = nullptr;
  cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device);
  cudaError_t e2 = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
  if (e2 != cudaSuccess) { delete c; return fail(STK_ERR_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(e2)); }
  *out = c;
  return STK_OK;
}

int stk_ctx_destroy(stk_ctx* ctx) {
  if (!ctx) return STK_OK;
  cudaSetDevice(ctx->device);
  cudaStreamDestroy(ctx->stream);
  if (ctx->diag_lib) cudaLibraryUnload(ctx->diag_lib);
  delete ctx;
  return STK_OK;
}

int stk_ctx_stream(stk_ctx* ctx, stk_stream* out) {
  if (!ctx || !out) return fail(STK_ERR_ARG, "null argument");
  *out = (stk_stream)ctx->stream;
  return STK_OK;
}

int stk_compile(const char* cuda_src, const char* cache_dir, const char* include_dir,
                char* path_out, size_t path_cap) {
  if (!cuda_src || !path_out) return fail(STK_ERR_ARG, "null argument");
  std::string src(cuda_src);
  std::string dir = cache_dir && *cache_dir ? cache_dir : "cache/cuda";
  std::string inc = include_dir && *include_dir ? include_dir : lib_dir();
  // the device library is part of the program text: hash it too, so that a
  // changed stk_device.cuh never resurrects a stale cubin
  std::string devlib, abi;
  read_file(inc + "/stk_device.cuh", &devlib);
  read_file(inc + "/stk_abi.h", &abi);
  std::string key = Sha1::hex(src + "\n//--\n" + devlib + "\n//--\n" + abi + "\n//--\n" + (getenv("STK_NVRTC_FLAGS") ? getenv("STK_NVRTC_FLAGS") : ""));
  std::string path = dir + "/stk_" + key + ".cubin";
  if ((size_t)path.size() + 1 > path_cap) return fail(STK_ERR_ARG, "path buffer too small");
  if (!file_exists(path)) {
    if (mkdirs(dir) != 0) return fail(STK_ERR_IO, "cannot create cache dir %s", dir.c_str());
    nvrtcProgram prog;
    if (nvrtcCreateProgram(&prog, src.c_str(), "stk_model.cu", 0, nullptr, nullptr) != NVRTC_SUCCESS)
      return fail(STK_ERR_COMPILE, "nvrtcCreateProgram failed");
    std::string iopt = "-I" + inc;
    std::vector<std::string> extra;                     // STK_NVRTC_FLAGS: extra compiler options (debugging knob, e.g. "-G")
    if (const char* xf = getenv("STK_NVRTC_FLAGS")) { std::istringstream xs(xf); std::string t; while (xs >> t) extra.push_back(t); }
    std::vector<const char*> opts = {"--gpu-architecture=sm_100a", "--std=c++17", "--fmad=true",
                                     "-default-device", iopt.c_str(), "-I/usr/local/cuda/include"};
    bool debug = false; for (auto& t : extra) if (t == "-G" || t == "--device-debug") debug = true;
    if (!debug) { if (getenv("STK_NVRTC_OLDORDER")) opts.insert(opts.begin() + 2, "-lineinfo"); else opts.push_back("-lineinfo"); }
    for (auto& t : extra) opts.push_back(t.c_str());
    nvrtcResult r = nvrtcCompileProgram(prog, (int)opts.size(), opts.data());
    if (r != NVRTC_SUCCESS) {
      size_t n = 0; nvrtcGetProgramLogSize(prog, &n);
      std::string log(n, '\0'); if (n) nvrtcGetProgramLog(prog, &log[0]);
      nvrtcDestroyProgram(&prog);
      return fail(STK_ERR_COMPILE, "NVRTC: %s\n%s", nvrtcGetErrorString(r), log.c_str());
    }
    size_t n = 0;
    if (nvrtcGetCUBINSize(prog, &n) != NVRTC_SUCCESS || n == 0) { nvrtcDestroyProgram(&prog); return fail(STK_ERR_COMPILE, "NVRTC produced no cubin"); }
    std::string bin(n, '\0');
    nvrtcGetCUBIN(prog, &bin[0]);
    nvrtcDestroyProgram(&prog);
    std::string tmp = path + ".tmp." + std::to_string((long long)getpid());
    { std::ofstream f(tmp, std::ios::binary); if (!f) return fail(STK_ERR_IO, "cannot write %s", tmp.c_str()); f.write(bin.data(), (std::streamsize)bin.size()); }
    { std::ofstream f(dir + "/stk_" + key + ".cu"); f << src; }  // keep the text beside it, like sampler_<hash>.cc
    if (rename(tmp.c_str(), path.c_str()) != 0) return fail(STK_ERR_IO, "rename %s failed", tmp.c_str());
  }
  memcpy(path_out, path.c_str(), path.size() + 1);
  return STK_OK;
}

int stk_module_load(stk_ctx* ctx, const char* spec, const char* cache_dir, stk_module** out) {
  if (!spec || !out) return fail(STK_ERR_ARG, "null argument");
  int rc = need_device(ctx);
  if (rc) return rc;
  std::string s(spec), path;
  if (s.rfind("builtin:", 0) == 0) {
    path = lib_dir() + "/cubin/" + s.substr(8) + ".cubin";
    if (!file_exists(path)) return fail(STK_ERR_IO, "builtin module %s not built (%s missing; run __graft_entry__.build())", s.c_str() + 8, path.c_str());
  } else if (s.size() > 6 && s.compare(s.size() - 6, 6, ".cubin") == 0 && s.find('\n') == std::string::npos) {
    path = s;
    if (!file_exists(path)) return fail(STK_ERR_IO, "cubin %s not found", path.c_str());
  } else {
    char buf[4096];
    rc = stk_compile(spec, cache_dir, nullptr, buf, sizeof buf);
    if (rc) return rc;
    path = buf;
  }
  std::string bin;
  if (!read_file(path, &bin)) return fail(STK_ERR_IO, "cannot read %s", path.c_str());
  stk_module* m = new stk_module;
  m->ctx = ctx;
  TRACE("module_load %s (%zu bytes)", path.c_str(), bin.size());
  cudaError_t e = cudaLibraryLoadData(&m->lib, bin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0);
  if (e != cudaSuccess) { delete m; return fail(STK_ERR_CUDA, "cudaLibraryLoadData(%s): %s", path.c_str(), cudaGetErrorString(e)); }
  TRACE("library loaded");
  void* dptr = nullptr; size_t bytes = 0;
  e = cudaLibraryGetGlobal(&dptr, &bytes, m->lib, "stk_desc");
  if (e != cudaSuccess) { cudaLibraryUnload(m->lib); delete m; return fail(STK_ERR_ARG, "module %s exports no stk_desc: %s", path.c_str(), cudaGetErrorString(e)); }
  TRACE("stk_desc at %p, %zu bytes", dptr, bytes);
  m->desc.assign(bytes, '\0');
  e = cudaMemcpy(&m->desc[0], dptr, bytes, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) { cudaLibraryUnload(m->lib); delete m; return fail(STK_ERR_CUDA, "read stk_desc: %s", cudaGetErrorString(e)); }
  while (!m->desc.empty() && m->desc.back() == '\0') m->desc.pop_back();
  TRACE("descriptor read:\n%s", m->desc.c_str());
  rc = parse_desc(m);
  if (rc) { cudaLibraryUnload(m->lib); delete m; return rc; }
  TRACE("descriptor parsed: %zu fields, %zu kernels", m->fields.size(), m->kernels.size());
  auto resolve = [&](KernelDesc& k) -> int {
    if (k.plan >= 0) return STK_OK;
    TRACE("resolve %s", k.symbol.c_str());
    cudaError_t e3 = cudaLibraryGetKernel(&k.k, m->lib, k.symbol.c_str());
    if (e3 != cudaSuccess) return fail(STK_ERR_ARG, "module %s: kernel symbol %s: %s", m->name.c_str(), k.symbol.c_str(), cudaGetErrorString(e3));
    if (!k.pre_symbol.empty()) {
      e3 = cudaLibraryGetKernel(&k.pre_k, m->lib, k.pre_symbol.c_str());
      if (e3 != cudaSuccess) return fail(STK_ERR_ARG, "module %s: pre-pass symbol %s: %s", m->name.c_str(), k.pre_symbol.c_str(), cudaGetErrorString(e3));
    }
    if (!k.ph_symbol.empty()) {
      e3 = cudaLibraryGetKernel(&k.ph_k, m->lib, k.ph_symbol.c_str());
      if (e3 != cudaSuccess) return fail(STK_ERR_ARG, "module %s: phase symbol %s: %s", m->name.c_str(), k.ph_symbol.c_str(), cudaGetErrorString(e3));
      if (field_index(m, k.ph_field) < 0) return fail(STK_ERR_ARG, "module %s: phase items field %s unknown", m->name.c_str(), k.ph_field.c_str());
    }
    return STK_OK;
  };
  for (auto& k : m->kernels) if ((rc = resolve(k))) { cudaLibraryUnload(m->lib); delete m; return rc; }
  for (int i = 0; i < 3; ++i) if (m->has_entry[i] && (rc = resolve(m->entry[i]))) { cudaLibraryUnload(m->lib); delete m; return rc; }
  for (auto& pl : m->plans)
    for (auto& sd : pl.stages) {
      cudaError_t e3 = cudaLibraryGetKernel(&sd.k, m->lib, sd.symbol.c_str());
      if (e3 != cudaSuccess) {
        rc = fail(STK_ERR_ARG, "module %s: stage symbol %s: %s", m->name.c_str(), sd.symbol.c_str(), cudaGetErrorString(e3));
        cudaLibraryUnload(m->lib); delete m; return rc;
      }
    }
  *out = m;
  return STK_OK;
}

int stk_module_destroy(stk_module* m) {
  if (!m) return STK_OK;
  cudaSetDevice(m->ctx->device);
  cudaLibraryUnload(m->lib);
  delete m;
  return STK_OK;
}

int stk_module_describe(stk_module* m, char* buf, size_t cap) {
  if (!m || !buf) return fail(STK_ERR_ARG, "null argument");
  if (m->desc.size() + 1 > cap) return fail(STK_ERR_ARG, "buffer too small (%zu needed)", m->desc.size() + 1);
  memcpy(buf, m->desc.c_str(), m->desc.size() + 1);
  return STK_OK;
}

int stk_module_field(stk_module* m, const char* name) {
  if (!m || !name) return fail(STK_ERR_ARG, "null argument");
  int i = field_index(m, name);
  return i >= 0 ? i : fail(STK_ERR_ARG, "unknown field %s", name);
}

int stk_module_kernel(stk_module* m, const char* name) {
  if (!m || !name) return fail(STK_ERR_ARG, "null argument");
  for (size_t i = 0; i < m->kernels.size(); ++i) if (m->kernels[i].name == name) return (int)i;
  return fail(STK_ERR_ARG, "unknown kernel %s", name);
}

int stk_chains_create(stk_module* m, int64_t n_chains, uint64_t seed, stk_chains** out) {
  if (!m || !out) return fail(STK_ERR_ARG, "null argument");
  if (n_chains <= 0) return fail(STK_ERR_ARG, "n_chains must be positive");
  int rc = need_device(m->ctx);
  if (rc) return rc;
  stk_chains* c = new stk_chains;
  c->mod = m; c->n = n_chains; c->seed = seed;
  c->dptr.assign(m->fields.size(), nullptr);
  c->flen.assign(m->fields.size(), 0);
  for (auto& p : m->params) c->par.push_back(p.second);
  for (size_t f = 0; f < m->fields.size(); ++f) if (m->fields[f].is_state) c->flen[f] = m->fields[f].len;
  rc = ensure_fields(c);
  if (rc) { stk_chains_destroy(c); return rc; }
  cudaError_t e = cudaMalloc((void**)&c->accept, sizeof(unsigned long long) * (size_t)n_chains);
  if (e == cudaSuccess) e = cudaMemset(c->accept, 0, sizeof(unsigned long long) * (size_t)n_chains);
  if (e == cudaSuccess) e = cudaMalloc((void**)&c->out_lp, sizeof(double) * (size_t)n_chains);
  if (e == cudaSuccess && m->nparam > 0) {
    e = cudaMalloc((void**)&c->out_grad, sizeof(double) * (size_t)n_chains * (size_t)m->nparam);
    if (e == cudaSuccess) c->grad_cap = (long long)n_chains * m->nparam;
  }
  if (e != cudaSuccess) { stk_chains_destroy(c); return fail(STK_ERR_CUDA, "cudaMalloc: %s", cudaGetErrorString(e)); }
  for (int f : m->record) c->nrec += 0 * f;  // sized at stk_record time (capacities may be deferred)
  *out = c;
  return STK_OK;
}

int stk_chains_destroy(stk_chains* c) {
  if (!c) return STK_OK;
  cudaSetDevice(c->mod->ctx->device);
  cudaStreamSynchronize(c->mod->ctx->stream);
  if (c->last && c->last != c->mod->ctx->stream) cudaStreamSynchronize(c->last);
  for (void* p : c->dptr) if (p) cudaFree(p);
  if (c->draws) cudaFree(c->draws);
  if (c->out_lp) cudaFree(c->out_lp);
  if (c->out_grad) cudaFree(c->out_grad);
  if (c->accept) cudaFree(c->accept);
  if (c->scratch) cudaFree(c->scratch);
  delete c;
  return STK_OK;
}

int stk_chains_set_offset(stk_chains* c, int64_t first) {
  if (!c || first < 0) return fail(STK_ERR_ARG, "bad argument");
  c->chain0 = first;
  return STK_OK;
}

int stk_set_data(stk_chains* c, int field, const void* host, size_t bytes) {
  if (!c || !host) return fail(STK_ERR_ARG, "null argument");
  stk_module* m = c->mod;
  if (field < 0 || field >= (int)m->fields.size() || m->fields[field].is_state) return fail(STK_ERR_ARG, "field %d is not a data field", field);
  int rc = need_device(m->ctx);
  if (rc) return rc;
  const FieldDesc& fd = m->fields[field];
  size_t es = elem_size(fd);
  if (bytes % es) return fail(STK_ERR_ARG, "data field %s: %zu bytes is not a multiple of %zu", fd.name.c_str(), bytes, es);
  long long len = (long long)(bytes / es);
  if (fd.len > 0 && len != fd.len) return fail(STK_ERR_ARG, "data field %s expects %lld elements, got %lld", fd.name.c_str(), fd.len, len);
  if ((rc = sync_chains(c))) return rc;            // no kernel of these chains may still be reading the old buffer
  if (c->dptr[field] && c->flen[field] != len) {
    CU(cudaFree(c->dptr[field])); c->dptr[field] = nullptr;
    // fields sized by an expression follow the data: drop them, ensure_fields re-creates them (zeroed)
    for (size_t f = 0; f < m->fields.size(); ++f)
      if (!m->fields[f].len_expr.empty() && (m->fields[f].is_state || m->fields[f].is_work) && c->dptr[f]) { CU(cudaFree(c->dptr[f])); c->dptr[f] = nullptr; c->flen[f] = 0; }
  }
  if (!c->dptr[field]) CU(cudaMalloc(&c->dptr[field], bytes ? bytes : 8));
  c->flen[field] = len;
  c->dirty_data = true; c->dirty_state = true;
  // `sorted` data fields (kernels that binary-search them; a sum over the observations does not depend on
  // their order): upload an ascending copy.
  std::vector<double> tmp;
  if (fd.sorted && !fd.is_i32 && !std::is_sorted((const double*)host, (const double*)host + len)) {
    tmp.assign((const double*)host, (const double*)host + len);
    std::sort(tmp.begin(), tmp.end());
    host = tmp.data();
  }
  CU(cudaMemcpyAsync(c->dptr[field], host, bytes, cudaMemcpyHostToDevice, m->ctx->stream));
  CU(cudaStreamSynchronize(m->ctx->stream));
  return STK_OK;
}

int stk_set_state(stk_chains* c, int field, const void* host, size_t bytes, int broadcast) {
  if (!c || !host) return fail(STK_ERR_ARG, "null argument");
  stk_module* m = c->mod;
  if (field < 0 || field >= (int)m->fields.size() || !m->fields[field].is_state) return fail(STK_ERR_ARG, "field %d is not a state field", field);
  int rc = need_device(m->ctx);
  if (rc) return rc;
  if ((rc = ensure_fields(c))) return rc;
  if (c->last && c->last != m->ctx->stream) CU(cudaStreamSynchronize(c->last));   // a launch on a caller stream may still use the field
  const FieldDesc& fd = m->fields[field];
  size_t es = elem_size(fd);
  if (c->flen[field] <= 0 && fd.len == 0 && fd.len_expr.empty()) {
    // a state field declared with length 0 (vectors of stkc-generated modules) takes its capacity from the first upload
    size_t per = broadcast ? es : es * (size_t)c->n;
    if (bytes == 0 || bytes % per) return fail(STK_ERR_ARG, "state field %s: %zu bytes is not a whole number of %s", fd.name.c_str(), bytes, broadcast ? "elements" : "[.][n_chains] rows");
    c->flen[field] = (long long)(bytes / per);
    if ((rc = ensure_fields(c))) return rc;
  }
  long long L = c->flen[field];
  if (L <= 0 || !c->dptr[field]) return fail(STK_ERR_ARG, "state field %s has no size yet", fd.name.c_str());
  c->dirty_state = true;
  if (broadcast) {
    if (bytes != (size_t)L * es) return fail(STK_ERR_ARG, "state field %s (broadcast) expects %zu bytes, got %zu", fd.name.c_str(), (size_t)L * es, bytes);
    std::vector<char> tmp((size_t)L * (size_t)c->n * es);
    for (long long e = 0; e < L; ++e)
      for (long long k = 0; k < c->n; ++k) memcpy(&tmp[((size_t)e * c->n + k) * es], (const char*)host + e * es, es);
    CU(cudaMemcpyAsync(c->dptr[field], tmp.data(), tmp.size(), cudaMemcpyHostToDevice, m->ctx->stream));
    CU(cudaStreamSynchronize(m->ctx->stream));
  } else {
    size_t want = (size_t)L * (size_t)c->n * es;
    if (bytes != want) return fail(STK_ERR_ARG, "state field %s expects %zu bytes ([%lld][%lld]), got %zu", fd.name.c_str(), want, L, c->n, bytes);
    CU(cudaMemcpyAsync(c->dptr[field], host, bytes, cudaMemcpyHostToDevice, m->ctx->stream));
    CU(cudaStreamSynchronize(m->ctx->stream));
  }
  return STK_OK;
}

int stk_get_state(stk_chains* c, int field, void* host, size_t bytes) {
  if (!c || !host) return fail(STK_ERR_ARG, "null argument");
  stk_module* m = c->mod;
  if (field < 0 || field >= (int)m->fields.size()) return fail(STK_ERR_ARG, "bad field %d", field);
  int rc = need_device(m->ctx);
  if (rc) return rc;
  const FieldDesc& fd = m->fields[field];
  size_t want = (size_t)c->flen[field] * elem_size(fd) * (fd.is_state ? (size_t)c->n : 1);
  if (!c->dptr[field] || want == 0) return fail(STK_ERR_ARG, "field %s is empty", fd.name.c_str());
  if (bytes != want) return fail(STK_ERR_ARG, "field %s holds %zu bytes, buffer is %zu", fd.name.c_str(), want, bytes);
  if (c->last && c->last != m->ctx->stream) CU(cudaStreamSynchronize(c->last));
  CU(cudaMemcpyAsync(host, c->dptr[field], bytes, cudaMemcpyDeviceToHost, m->ctx->stream));
  CU(cudaStreamSynchronize(m->ctx->stream));
  return check_err(c);
}

int stk_device_ptr(stk_chains* c, int field, void** dptr, size_t* bytes) {
  if (!c) return fail(STK_ERR_ARG, "null argument");
  stk_module* m = c->mod;
  if (field < 0 || field >= (int)m->fields.size()) return fail(STK_ERR_ARG, "bad field %d", field);
  int rc = ensure_fields(c);
  if (rc) return rc;
  const FieldDesc& fd = m->fields[field];
  if (dptr) *dptr = c->dptr[field];
  if (bytes) *bytes = (size_t)c->flen[field] * elem_size(fd) * (fd.is_state ? (size_t)c->n : 1);
  return STK_OK;
}

int stk_set_param(stk_chains* c, const char* name, double v) {
  if (!c || !name) return fail(STK_ERR_ARG, "null argument");
  for (size_t i = 0; i < c->mod->params.size(); ++i) if (c->mod->params[i].first == name) { c->par[i] = v; return STK_OK; }
  return fail(STK_ERR_ARG, "module %s has no parameter %s", c->mod->name.c_str(), name);
}

int stk_get_param(stk_chains* c, const char* name, double* v) {
  if (!c || !name || !v) return fail(STK_ERR_ARG, "null argument");
  for (size_t i = 0; i < c->mod->params.size(); ++i) if (c->mod->params[i].first == name) { *v = c->par[i]; return STK_OK; }
  return fail(STK_ERR_ARG, "module %s has no parameter %s", c->mod->name.c_str(), name);
}

static int run_entry(stk_chains* c, int e, const char* what) {
  if (!c) return fail(STK_ERR_ARG, "null chains");
  if (!c->mod->has_entry[e]) return fail(STK_ERR_UNSUPPORTED, "module %s has no %s entry", c->mod->name.c_str(), what);
  int rc = need_device(c->mod->ctx);
  if (rc) return rc;
  if (e == 1 && (rc = ensure_grad(c))) return rc;
  if (c->last && c->last != c->mod->ctx->stream) { CU(cudaStreamSynchronize(c->last)); c->last = nullptr; }
  if (c->mod->entry[e].plan >= 0) {
    if (e == 2) c->dirty_state = true;
    return run_entry_plan(c, c->mod->entry[e].plan, c->mod->ctx->stream);
  }
  return launch(c, c->mod->entry[e], 0, c->mod->ctx->stream);
}

int stk_logp(stk_chains* c, double* lp_host) {
  int rc = run_entry(c, 0, "logp");
  if (rc) return rc;
  if (lp_host) CU(cudaMemcpyAsync(lp_host, c->out_lp, sizeof(double) * (size_t)c->n, cudaMemcpyDeviceToHost, c->mod->ctx->stream));
  CU(cudaStreamSynchronize(c->mod->ctx->stream));
  return check_err(c);
}

int stk_logp_grad(stk_chains* c, double* lp_host, double* grad_host) {
  int rc = run_entry(c, 1, "logp_grad");
  if (rc) return rc;
  cudaStream_t st = c->mod->ctx->stream;
  if (lp_host) CU(cudaMemcpyAsync(lp_host, c->out_lp, sizeof(double) * (size_t)c->n, cudaMemcpyDeviceToHost, st));
  if (grad_host) CU(cudaMemcpyAsync(grad_host, c->out_grad, sizeof(double) * (size_t)c->n * (size_t)num_params(c), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return check_err(c);
}

int stk_step(stk_chains* c, int kernel_id, int64_t n_transitions, stk_stream stream) {
  if (!c) return fail(STK_ERR_ARG, "null chains");
  if (kernel_id < 0 || kernel_id >= (int)c->mod->kernels.size()) return fail(STK_ERR_ARG, "kernel_id %d out of range", kernel_id);
  if (n_transitions < 0) return fail(STK_ERR_ARG, "negative n_transitions");
  if (n_transitions == 0) return STK_OK;
  int rc = need_device(c->mod->ctx);
  if (rc) return rc;
  cudaStream_t st = stream ? (cudaStream_t)stream : c->mod->ctx->stream;
  if (c->last && c->last != st) CU(cudaStreamSynchronize(c->last));   // launches of one chain set are ordered, whatever the stream
  c->last = st;
  KernelDesc& kd = c->mod->kernels[kernel_id];
  if (kd.plan >= 0) return run_kernel_plan(c, kd.plan, n_transitions, st);
  // a kernel with a pre-pass runs in chunks whose scratch ([transitions][k][chains] doubles) fits the budget
  long long chunk = n_transitions;
  if (kd.ph_k) chunk = 1;
  if (kd.pre_k) {
    size_t per_tr = sizeof(double) * (size_t)kd.pre_per * (size_t)c->n;
    size_t budget = (size_t)2048 << 20;
    if (const char* mb = getenv("STK_SCRATCH_MB")) budget = (size_t)std::max(1, atoi(mb)) << 20;
    chunk = std::max<long long>(1, std::min<long long>(n_transitions, (long long)(budget / per_tr)));
    size_t need = per_tr * (size_t)chunk;
    if (need > c->scratch_bytes) {
      CU(cudaStreamSynchronize(st));
      if (c->scratch) { CU(cudaFree(c->scratch)); c->scratch = nullptr; c->scratch_bytes = 0; }
      CU(cudaMalloc((void**)&c->scratch, need));
      c->scratch_bytes = need;
    }
  }
  for (long long done = 0; done < n_transitions; done += chunk) {
    long long nt = std::min(chunk, n_transitions - done);
    rc = launch(c, kd, nt, st);
    if (rc) return rc;
    if (c->draws && c->slots > 0) {
      unsigned long long a = c->step, b = c->step + (unsigned long long)nt;
      long long add = (long long)(b / (unsigned long long)c->thin - a / (unsigned long long)c->thin);
      c->recorded = std::min(c->slots, c->recorded + add);
    }
    c->step += (unsigned long long)nt;
  }
  return STK_OK;
}

int stk_simulate(stk_chains* c, stk_stream stream) {
  if (!c) return fail(STK_ERR_ARG, "null chains");
  if (!c->mod->has_entry[2]) return fail(STK_ERR_UNSUPPORTED, "module %s has no simulate entry", c->mod->name.c_str());
  int rc = need_device(c->mod->ctx);
  if (rc) return rc;
  cudaStream_t st = stream ? (cudaStream_t)stream : c->mod->ctx->stream;
  if (c->last && c->last != st) CU(cudaStreamSynchronize(c->last));
  c->last = st;
  c->dirty_state = true;
  if (c->mod->entry[2].plan >= 0) return run_entry_plan(c, c->mod->entry[2].plan, st);
  return launch(c, c->mod->entry[2], 0, st);
}

int stk_record(stk_chains* c, int64_t n_slots, int64_t thin) {
  if (!c || n_slots < 0 || thin < 1) return fail(STK_ERR_ARG, "bad argument");
  int rc = need_device(c->mod->ctx);
  if (rc) return rc;
  if ((rc = sync_chains(c))) return rc;
  if (c->draws) { CU(cudaFree(c->draws)); c->draws = nullptr; }
  c->slots = 0; c->recorded = 0; c->thin = thin;
  if (n_slots == 0) return STK_OK;
  if ((rc = ensure_fields(c))) return rc;
  long long nrec = 0;
  for (int f : c->mod->record) nrec += c->flen[f];
  if (nrec <= 0) return fail(STK_ERR_UNSUPPORTED, "module %s records nothing", c->mod->name.c_str());
  c->nrec = nrec;
  CU(cudaMalloc((void**)&c->draws, sizeof(double) * (size_t)n_slots * (size_t)nrec * (size_t)c->n));
  c->slots = n_slots;
  return STK_OK;
}

int stk_get_draws(stk_chains* c, double* host, size_t bytes, int64_t* n_recorded) {
  if (!c) return fail(STK_ERR_ARG, "null chains");
  if (n_recorded) *n_recorded = c->recorded;
  if (!host) return STK_OK;
  int rc = need_device(c->mod->ctx);
  if (rc) return rc;
  size_t want = sizeof(double) * (size_t)c->recorded * (size_t)c->nrec * (size_t)c->n;
  if (bytes < want) return fail(STK_ERR_ARG, "draws need %zu bytes, buffer is %zu", want, bytes);
  if (c->last && c->last != c->mod->ctx->stream) CU(cudaStreamSynchronize(c->last));
  if (want) CU(cudaMemcpyAsync(host, c->draws, want, cudaMemcpyDeviceToHost, c->mod->ctx->stream));
  CU(cudaStreamSynchronize(c->mod->ctx->stream));
  return STK_OK;
}

int stk_sync(stk_chains* c) {
  if (!c) return fail(STK_ERR_ARG, "null chains");
  int rc = need_device(c->mod->ctx);
  if (rc) return rc;
  if ((rc = sync_chains(c))) return rc;
  return check_err(c);
}

int stk_num_params(stk_chains* c) { return c ? (int)num_params(c) : fail(STK_ERR_ARG, "null chains"); }
int64_t stk_num_chains(stk_chains* c) { return c ? c->n : (int64_t)fail(STK_ERR_ARG, "null chains"); }

}  // extern "C"

// ---------------------------------------------------------------------------
// diagnostics (R-hat / ESS); kernels in models/diag.cu
// ---------------------------------------------------------------------------
static int load_diag(stk_ctx* ctx) {
  if (ctx->diag_lib) return STK_OK;
  std::string path = lib_dir() + "/cubin/diag.cubin", bin;
  if (!read_file(path, &bin)) return fail(STK_ERR_IO, "%s missing; run __graft_entry__.build()", path.c_str());
  CU(cudaLibraryLoadData(&ctx->diag_lib, bin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0));
  CU(cudaLibraryGetKernel(&ctx->k_moments, ctx->diag_lib, "diag_moments"));
  CU(cudaLibraryGetKernel(&ctx->k_acov, ctx->diag_lib, "diag_acov"));
  return STK_OK;
}

// additive diagnostics buffer of this shard: [4][P] moments then [P][max_lag+1]
// mean-over-chains autocovariance SUMS.  bytes = 8 * P * (4 + max_lag + 1).
// Left on the device (*d_out, caller frees) so that a collective can reduce it in place before it is read.
static int diag_partial_device(stk_chains* c, int64_t* max_lag_io, double** d_out, size_t* bytes_out) {
  if (!c->draws || c->recorded < 2) return fail(STK_ERR_ARG, "need at least 2 recorded draws (stk_record + stk_step)");
  int rc = need_device(c->mod->ctx);
  if (rc) return rc;
  stk_ctx* ctx = c->mod->ctx;
  if ((rc = load_diag(ctx))) return rc;
  if (c->last && c->last != ctx->stream) CU(cudaStreamSynchronize(c->last));
  long long N = c->recorded, C = c->n; int P = (int)c->nrec;
  long long max_lag = *max_lag_io;
  if (max_lag < 0 || max_lag > N - 1) max_lag = N - 1;
  *max_lag_io = max_lag;
  size_t want = sizeof(double) * (size_t)P * (size_t)(4 + max_lag + 1);
  double *d_buf = nullptr, *d_mean = nullptr;
  CU(cudaMalloc((void**)&d_buf, want));
  cudaError_t e = cudaMalloc((void**)&d_mean, sizeof(double) * (size_t)P * (size_t)C);
  if (e != cudaSuccess) { cudaFree(d_buf); return fail(STK_ERR_CUDA, "cudaMalloc: %s", cudaGetErrorString(e)); }
  cudaStream_t st = ctx->stream;
  cudaMemsetAsync(d_buf, 0, want, st);
  const double* draws = c->draws;
  {
    void* args[] = {(void*)&draws, &N, &P, &C, &d_mean, &d_buf};
    dim3 grid((unsigned)((C + 255) / 256), (unsigned)P);
    e = cudaLaunchKernel((const void*)ctx->k_moments, grid, dim3(256), args, 0, st);
    g_launches.fetch_add(1);
  }
  if (e == cudaSuccess) {
    int CB = 32;
    while (CB > 1 && (size_t)N * CB * 8 > 200 * 1024) CB >>= 1;
    size_t smem = (size_t)N * CB * 8;
    if (smem > 227 * 1024) { cudaFree(d_buf); cudaFree(d_mean); return fail(STK_ERR_UNSUPPORTED, "too many draws (%lld) for the ESS kernel; thin the record", N); }
    if (smem > 48 * 1024) e = cudaFuncSetAttribute((const void*)ctx->k_acov, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    double* d_acov = d_buf + 4 * (size_t)P;
    long long ml = max_lag;
    void* args[] = {(void*)&draws, &d_mean, &N, &P, &C, &CB, &ml, &d_acov};
    dim3 grid((unsigned)((C + CB - 1) / CB), (unsigned)P);
    if (e == cudaSuccess) e = cudaLaunchKernel((const void*)ctx->k_acov, grid, dim3(256), args, smem, st);
    g_launches.fetch_add(1);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);      // d_mean is only needed by the two kernels
  cudaFree(d_mean);
  if (e != cudaSuccess) { cudaFree(d_buf); return fail(STK_ERR_CUDA, "diagnostics: %s", cudaGetErrorString(e)); }
  *d_out = d_buf; *bytes_out = want;
  return STK_OK;
}

extern "C" {

int stk_diag_partial(stk_chains* c, int64_t max_lag, double* buf, size_t bytes) {
  if (!c || !buf) return fail(STK_ERR_ARG, "null argument");
  double* d_buf = nullptr; size_t want = 0;
  {
    long long N = c->recorded; int P = (int)c->nrec;
    long long ml = (max_lag < 0 || max_lag > N - 1) ? N - 1 : max_lag;
    if (c->draws && N >= 2 && bytes != sizeof(double) * (size_t)P * (size_t)(4 + ml + 1))
      return fail(STK_ERR_ARG, "diag buffer must be %zu bytes, got %zu", sizeof(double) * (size_t)P * (size_t)(4 + ml + 1), bytes);
  }
  int rc = diag_partial_device(c, &max_lag, &d_buf, &want);
  if (rc) return rc;
  cudaError_t e = cudaMemcpy(buf, d_buf, want, cudaMemcpyDeviceToHost);
  cudaFree(d_buf);
  if (e != cudaSuccess) return fail(STK_ERR_CUDA, "diagnostics: %s", cudaGetErrorString(e));
  return STK_OK;
}

// ---- the one collective of the path behind the C-ABI: NCCL through dlopen (no link-time dependency: the library loads
// and every other entry point works on a machine without NCCL) -------------------------------------------------------------
struct Id128 { char b[128]; };           // ncclUniqueId (NCCL_UNIQUE_ID_BYTES = 128), passed BY VALUE to ncclCommInitRank
namespace {
struct NcclApi {
  void* h = nullptr;
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, Id128, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
}  // namespace
static NcclApi g_nccl;
static int nccl_load() {
  if (g_nccl.h) return STK_OK;
  const char* names[] = {"libnccl.so.2", "libnccl.so", nullptr};
  void* h = nullptr;
  for (int i = 0; names[i] && !h; ++i) h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
  if (!h) return fail(STK_ERR_UNSUPPORTED, "NCCL is not available (dlopen libnccl.so.2: %s)", dlerror());
  *(void**)&g_nccl.GetUniqueId = dlsym(h, "ncclGetUniqueId");
  *(void**)&g_nccl.CommInitRank = dlsym(h, "ncclCommInitRank");
  *(void**)&g_nccl.AllReduce = dlsym(h, "ncclAllReduce");
  *(void**)&g_nccl.CommDestroy = dlsym(h, "ncclCommDestroy");
  *(void**)&g_nccl.GetErrorString = dlsym(h, "ncclGetErrorString");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.CommDestroy) return fail(STK_ERR_UNSUPPORTED, "libnccl lacks an expected symbol");
  g_nccl.h = h;
  return STK_OK;
}
struct stk_comm { void* comm; stk_ctx* ctx; int n_ranks, rank; };
#define NC(call) do { int r_ = (call); if (r_ != 0) return fail(STK_ERR_CUDA, "%s: %s", #call, g_nccl.GetErrorString ? g_nccl.GetErrorString(r_) : "nccl error"); } while (0)

int stk_nccl_unique_id(void* id128) {
  if (!id128) return fail(STK_ERR_ARG, "null id");
  int rc = nccl_load(); if (rc) return rc;
  NC(g_nccl.GetUniqueId(id128));
  return STK_OK;
}
int stk_nccl_comm_create(stk_ctx* ctx, int n_ranks, int rank, const void* id128, stk_comm** out) {
  if (!ctx || !id128 || !out || n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(STK_ERR_ARG, "bad argument");
  int rc = nccl_load(); if (rc) return rc;
  if ((rc = need_device(ctx))) return rc;
  Id128 id; memcpy(&id, id128, sizeof id);
  void* comm = nullptr;
  NC(g_nccl.CommInitRank(&comm, n_ranks, id, rank));
  *out = new stk_comm{comm, ctx, n_ranks, rank};
  return STK_OK;
}
int stk_nccl_comm_destroy(stk_comm* c) {
  if (!c) return STK_OK;
  if (g_nccl.h && c->comm) g_nccl.CommDestroy(c->comm);
  delete c;
  return STK_OK;
}
// R-hat / ESS over the chains of ALL ranks: the local additive buffer is all-reduced (sum, fp64) in place on the device over
// NVLink, then finished on the host of every rank ("stk_moments: local + NCCL-reduced diagnostics")
int stk_moments_nccl(stk_chains* c, stk_comm* comm, double* mean, double* var, double* rhat, double* ess) {
  if (!c || !comm) return fail(STK_ERR_ARG, "null argument");
  int64_t max_lag = -1; double* d_buf = nullptr; size_t bytes = 0;
  int rc = diag_partial_device(c, &max_lag, &d_buf, &bytes);
  if (rc) return rc;
  cudaStream_t st = c->mod->ctx->stream;
  int r = g_nccl.AllReduce(d_buf, d_buf, bytes / sizeof(double), /* ncclFloat64 */ 8, /* ncclSum */ 0, comm->comm, st);
  std::vector<double> buf(bytes / sizeof(double));
  cudaError_t e = r == 0 ? cudaMemcpyAsync(buf.data(), d_buf, bytes, cudaMemcpyDeviceToHost, st) : cudaSuccess;
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(d_buf);
  if (r != 0) return fail(STK_ERR_CUDA, "ncclAllReduce: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "nccl error");
  if (e != cudaSuccess) return fail(STK_ERR_CUDA, "diagnostics: %s", cudaGetErrorString(e));
  return stk_diag_finish(buf.data(), (int)c->nrec, c->recorded, max_lag, mean, var, rhat, ess);
}

// host-side finish on an (all-reduced) additive buffer.  Pure arithmetic on
// O(P * max_lag) numbers: Gelman-Rubin R-hat and Geyer initial-monotone ESS,
// compute_effective_sample_size.hpp:47-118.
int stk_diag_finish(const double* buf, int P, int64_t n_draws, int64_t max_lag,
                    double* mean, double* var, double* rhat, double* ess) {
  if (!buf || P <= 0 || n_draws < 2 || max_lag < 1) return fail(STK_ERR_ARG, "bad argument");
  const double N = (double)n_draws;
  for (int r = 0; r < P; ++r) {
    double M = buf[0 * P + r], sm = buf[1 * P + r], sm2 = buf[2 * P + r], sv = buf[3 * P + r];
    if (M < 1) return fail(STK_ERR_ARG, "empty diagnostics buffer");
    const double* ac = buf + 4 * (size_t)P + (size_t)r * (size_t)(max_lag + 1);  // sums over chains
    double mean_var = sv / M;
    double var_plus = mean_var * (N - 1) / N;
    if (M > 1) var_plus += (sm2 - sm * sm / M) / (M - 1);
    if (mean) mean[r] = sm / M;
    if (var) var[r] = var_plus;
    if (rhat) rhat[r] = std::sqrt(var_plus / mean_var);
    if (ess) {
      std::vector<double> rho((size_t)max_lag + 3, 0.0);
      auto rho_at = [&](long long t) { return 1.0 - (mean_var - ac[t] / M) / var_plus; };
      double even = 1.0, odd = rho_at(1);
      rho[0] = even; rho[1] = odd;
      long long s = 1;
      while (s < n_draws - 4 && s + 2 <= max_lag && even + odd > 0) {
        even = rho_at(s + 1); odd = rho_at(s + 2);
        if (even + odd >= 0) { rho[s + 1] = even; rho[s + 2] = odd; }
        s += 2;
      }
      long long max_s = s;
      if (even > 0) rho[max_s + 1] = even;
      for (long long k = 1; k <= max_s - 3; k += 2)
        if (rho[k + 1] + rho[k + 2] > rho[k - 1] + rho[k]) { rho[k + 1] = (rho[k - 1] + rho[k]) / 2; rho[k + 2] = rho[k + 1]; }
      double tau = -1.0;
      for (long long k = 0; k < max_s; ++k) tau += 2 * rho[k];
      tau += rho[max_s + 1];
      double tot = M * N;
      ess[r] = std::min(tot / tau, tot * std::log10(tot));
    }
  }
  return STK_OK;
}

int stk_moments(stk_chains* c, double* mean, double* var, double* rhat, double* ess, double* moments4) {
  if (!c) return fail(STK_ERR_ARG, "null chains");
  long long N = c->recorded; int P = (int)c->nrec;
  if (N < 4) return fail(STK_ERR_ARG, "need at least 4 recorded draws");
  long long max_lag = N - 1;
  std::vector<double> buf((size_t)P * (size_t)(4 + max_lag + 1));
  int rc = stk_diag_partial(c, max_lag, buf.data(), buf.size() * sizeof(double));
  if (rc) return rc;
  if (moments4) memcpy(moments4, buf.data(), sizeof(double) * 4 * (size_t)P);
  return stk_diag_finish(buf.data(), P, N, max_lag, mean, var, rhat, ess);
}

int stk_moments_multi(stk_chains* const* shards, int n_shards, double* mean, double* var, double* rhat, double* ess) {
  if (!shards || n_shards <= 0) return fail(STK_ERR_ARG, "no shards");
  for (int i = 0; i < n_shards; ++i) if (!shards[i]) return fail(STK_ERR_ARG, "null shard %d", i);
  const long long N = shards[0]->recorded; const int P = (int)shards[0]->nrec;
  if (N < 4) return fail(STK_ERR_ARG, "need at least 4 recorded draws");
  for (int i = 1; i < n_shards; ++i)
    if (shards[i]->recorded != N || (int)shards[i]->nrec != P)
      return fail(STK_ERR_ARG, "shard %d recorded %lld draws of %d values, shard 0 %lld of %d", i, shards[i]->recorded, (int)shards[i]->nrec, N, P);
  const long long max_lag = N - 1;
  std::vector<double> sum((size_t)P * (size_t)(4 + max_lag + 1), 0.0), buf(sum.size());
  for (int i = 0; i < n_shards; ++i) {
    int rc = stk_diag_partial(shards[i], max_lag, buf.data(), buf.size() * sizeof(double));
    if (rc) return rc;
    for (size_t k = 0; k < sum.size(); ++k) sum[k] += buf[k];          // shard order is fixed: the result does not depend on timing
  }
  return stk_diag_finish(sum.data(), P, N, max_lag, mean, var, rhat, ess);
}

}  // extern "C"

extern "C" int stk_num_recorded_values(stk_chains* c) { return c ? (int)c->nrec : fail(STK_ERR_ARG, "null chains"); }

// hooks for stk_noalloc.cu (a separate nvcc translation unit): the thread-local error text and the launch counter live here
extern "C" int stk_fail_(int code, const char* msg) { return fail(code, "%s", msg ? msg : ""); }
extern "C" void stk_count_launches_(int n) { g_launches.fetch_add(n); }
