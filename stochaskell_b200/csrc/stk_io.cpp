// stk_io.cpp — the wire formats either side of the HMC path, part of libstk_cuda's C-ABI:
//   * R-dump text: what `stanDump` writes for data and initial values (Stan.hs:346-363: `name <- scalar`,
//     `name <- c(v,...)`, `name <- structure(c(v,...), .Dim = c(d,...))`, column-major) and CmdStan reads;
//   * Stan CSV: what `runStan` gets back and `stanRead` parses (Stan.hs:365-371, 466-484): `#` comment lines, one
//     header line of column names (`lp__`, sampler columns ending in `__`, then `name.i.j`), one row per draw.
// With these, data prepared for `hmcStan` can be fed to the CUDA backend unchanged, and its draws can be post-processed
// by the reference's own `stanRead` / CmdStan's `stansummary`.
#include <cctype>
#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/stk_cuda.h"

extern "C" int stk_set_error_(int code, const char* msg);   // stk_runtime.cpp

namespace {
struct Var { std::string name; std::vector<long long> dims; std::vector<double> v; bool is_int; };
struct Csv { std::vector<std::string> cols; std::vector<double> v; long long rows; std::string comments; };

struct Lexer {
  const char* p; const char* e; std::string err;
  void ws() { while (p < e && (isspace((unsigned char)*p))) ++p; }
  bool eat(char c) { ws(); if (p < e && *p == c) { ++p; return true; } return false; }
  bool lit(const char* s) { ws(); size_t n = strlen(s); if ((size_t)(e - p) >= n && !strncmp(p, s, n)) { p += n; return true; } return false; }
  // number; R integer suffix L accepted; NaN / Inf / -Inf / NA accepted
  bool number(double& out, bool& integral) {
    ws();
    const char* q = p;
    if (lit("NaN") || lit("NA")) { out = NAN; integral = false; return true; }
    if (lit("Inf")) { out = INFINITY; integral = false; return true; }
    if (lit("-Inf")) { out = -INFINITY; integral = false; return true; }
    p = q;
    char* end = nullptr;
    errno = 0;
    out = strtod(p, &end);
    if (end == p) return false;
    integral = true;
    for (const char* c = p; c < end; ++c) if (*c == '.' || *c == 'e' || *c == 'E' || *c == 'n' || *c == 'i') integral = false;
    p = end;
    if (p < e && *p == 'L') ++p;
    return true;
  }
  // c(v, ...) | a:b | a single number
  bool vec(std::vector<double>& v, bool& all_int) {
    ws();
    if (lit("c(")) {
      if (eat(')')) return true;
      for (;;) {
        double x; bool ii;
        if (!number(x, ii)) return false;
        v.push_back(x); all_int = all_int && ii;
        if (eat(',')) continue;
        return eat(')');
      }
    }
    double a; bool ii;
    if (!number(a, ii)) return false;
    all_int = all_int && ii;
    ws();
    if (p < e && *p == ':') {   // integer range a:b (R writes .Dim = 3:4 on occasion)
      ++p; double b; bool jj;
      if (!number(b, jj)) return false;
      for (double x = a; x <= b; x += 1) v.push_back(x);
      return true;
    }
    v.push_back(a);
    return true;
  }
};
}  // namespace

struct stk_rdump { std::vector<Var> vars; };
struct stk_csv { Csv c; };

extern "C" int stk_rdump_parse(const char* text, stk_rdump** out) {
  if (!text || !out) return stk_set_error_(STK_ERR_ARG, "stk_rdump_parse: null argument");
  stk_rdump* d = new stk_rdump();
  Lexer lx{text, text + strlen(text), ""};
  for (;;) {
    lx.ws();
    if (lx.p >= lx.e) break;
    if (*lx.p == '#') { while (lx.p < lx.e && *lx.p != '\n') ++lx.p; continue; }
    Var v; v.is_int = true;
    bool quoted = lx.eat('"');
    while (lx.p < lx.e && (isalnum((unsigned char)*lx.p) || *lx.p == '_' || *lx.p == '.')) v.name.push_back(*lx.p++);
    if (quoted) lx.eat('"');
    if (v.name.empty() || !lx.lit("<-")) { delete d; return stk_set_error_(STK_ERR_ARG, ("R dump: expected `name <-` near: " + std::string(lx.p, std::min<size_t>(24, lx.e - lx.p))).c_str()); }
    bool ok;
    if (lx.lit("structure(")) {
      ok = lx.vec(v.v, v.is_int) && lx.eat(',') && lx.lit(".Dim") && lx.eat('=');
      std::vector<double> dm; bool di = true;
      ok = ok && lx.vec(dm, di) && lx.eat(')');
      long long tot = 1;
      for (double x : dm) { v.dims.push_back((long long)x); tot *= (long long)x; }
      if (ok && tot != (long long)v.v.size()) { delete d; return stk_set_error_(STK_ERR_ARG, ("R dump: .Dim of " + v.name + " does not match its length").c_str()); }
    } else {
      const bool is_c = (lx.ws(), lx.p + 1 < lx.e && lx.p[0] == 'c' && lx.p[1] == '(');
      ok = lx.vec(v.v, v.is_int);
      if (is_c || v.v.size() != 1) v.dims.push_back((long long)v.v.size());
    }
    if (!ok) { delete d; return stk_set_error_(STK_ERR_ARG, ("R dump: cannot parse the value of " + v.name).c_str()); }
    d->vars.push_back(std::move(v));
  }
  *out = d;
  return STK_OK;
}
extern "C" int stk_rdump_count(const stk_rdump* d) { return d ? (int)d->vars.size() : 0; }
extern "C" const char* stk_rdump_name(const stk_rdump* d, int i) { return d->vars[i].name.c_str(); }
extern "C" int stk_rdump_ndim(const stk_rdump* d, int i) { return (int)d->vars[i].dims.size(); }
extern "C" const long long* stk_rdump_dims(const stk_rdump* d, int i) { return d->vars[i].dims.data(); }
extern "C" int stk_rdump_is_int(const stk_rdump* d, int i) { return d->vars[i].is_int ? 1 : 0; }
extern "C" long long stk_rdump_size(const stk_rdump* d, int i) { return (long long)d->vars[i].v.size(); }
extern "C" const double* stk_rdump_values(const stk_rdump* d, int i) { return d->vars[i].v.data(); }
extern "C" void stk_rdump_free(stk_rdump* d) { delete d; }

static void put_num(std::string& s, double x, bool is_int) {
  char b[40];
  if (is_int) snprintf(b, sizeof b, "%lld", (long long)x);
  else if (std::isnan(x)) snprintf(b, sizeof b, "NaN");
  else if (std::isinf(x)) snprintf(b, sizeof b, x > 0 ? "Inf" : "-Inf");
  else {
    snprintf(b, sizeof b, "%.17g", x);
    if (!strpbrk(b, ".eEn")) strcat(b, ".0");    // keep reals recognisable as reals (Haskell's show does the same)
  }
  s += b;
}

// One `name <- value` line in stanDump's shape.  values are column-major for ndim >= 2.  Writes at most cap bytes
// (NUL-terminated) and returns the full length in *need; call with cap = 0 to size the buffer.
extern "C" int stk_rdump_format(const char* name, const double* values, int is_int, int ndim, const long long* dims,
                                char* buf, size_t cap, size_t* need) {
  if (!name || (!values && ndim >= 0 && !(ndim >= 1 && dims && dims[0] == 0)) || !need) return stk_set_error_(STK_ERR_ARG, "stk_rdump_format: null argument");
  long long tot = 1;
  for (int k = 0; k < ndim; ++k) tot *= dims[k];
  std::string s = std::string(name) + " <- ";
  auto cvec = [&](long long n) { s += "c("; for (long long i = 0; i < n; ++i) { if (i) s += ","; put_num(s, values[i], is_int != 0); } s += ")"; };
  if (ndim == 0) put_num(s, values[0], is_int != 0);
  else if (ndim == 1) cvec(tot);
  else {
    s += "structure("; cvec(tot); s += ", .Dim = c(";
    for (int k = 0; k < ndim; ++k) { if (k) s += ","; s += std::to_string(dims[k]); }
    s += "))";
  }
  s += "\n";
  *need = s.size();
  if (buf && cap) { const size_t n = std::min(cap - 1, s.size()); memcpy(buf, s.data(), n); buf[n] = 0; }
  return STK_OK;
}

// ---- Stan CSV ---------------------------------------------------------------------------------------------------------
// draws[(d * stride_draw) + (c * stride_col)] is column c of draw d; `comment` (may be NULL) is written as leading
// `# ` lines (CmdStan puts the configuration and, after warm-up, the step size / metric there).
extern "C" int stk_stan_csv_write(const char* path, int n_cols, const char* const* col_names, const double* draws,
                                  long long n_draws, long long stride_draw, long long stride_col, const char* comment) {
  if (!path || n_cols <= 0 || !col_names || (!draws && n_draws > 0)) return stk_set_error_(STK_ERR_ARG, "stk_stan_csv_write: bad argument");
  FILE* f = fopen(path, "w");
  if (!f) return stk_set_error_(STK_ERR_ARG, (std::string("cannot open ") + path + ": " + strerror(errno)).c_str());
  if (comment) {
    const char* p = comment;
    while (*p) { const char* nl = strchr(p, '\n'); size_t n = nl ? (size_t)(nl - p) : strlen(p); fprintf(f, "# %.*s\n", (int)n, p); p += n + (nl ? 1 : 0); }
  }
  for (int c = 0; c < n_cols; ++c) fprintf(f, "%s%s", c ? "," : "", col_names[c]);
  fputc('\n', f);
  std::string row;
  for (long long d = 0; d < n_draws; ++d) {
    row.clear();
    for (int c = 0; c < n_cols; ++c) {
      char b[40];
      const double x = draws[d * stride_draw + (long long)c * stride_col];
      if (std::isnan(x)) snprintf(b, sizeof b, "nan"); else if (std::isinf(x)) snprintf(b, sizeof b, x > 0 ? "inf" : "-inf");
      else snprintf(b, sizeof b, "%.17g", x);
      if (c) row += ',';
      row += b;
    }
    row += '\n';
    fwrite(row.data(), 1, row.size(), f);
  }
  if (fclose(f) != 0) return stk_set_error_(STK_ERR_ARG, "stk_stan_csv_write: write failed");
  return STK_OK;
}

extern "C" int stk_stan_csv_read(const char* path, stk_csv** out) {
  if (!path || !out) return stk_set_error_(STK_ERR_ARG, "stk_stan_csv_read: null argument");
  FILE* f = fopen(path, "r");
  if (!f) return stk_set_error_(STK_ERR_ARG, (std::string("cannot open ") + path + ": " + strerror(errno)).c_str());
  stk_csv* r = new stk_csv(); r->c.rows = 0;
  std::string line; char buf[1 << 16];
  bool have_header = false, bad = false;
  auto handle = [&](const std::string& ln) {
    if (ln.empty()) return;
    if (ln[0] == '#') { r->c.comments += ln; r->c.comments += '\n'; return; }
    if (!have_header) {
      size_t a = 0;
      for (;;) { size_t b = ln.find(',', a); r->c.cols.push_back(ln.substr(a, b == std::string::npos ? b : b - a)); if (b == std::string::npos) break; a = b + 1; }
      have_header = true; return;
    }
    const char* p = ln.c_str(); size_t n0 = r->c.v.size();
    for (;;) {
      char* end; double x = strtod(p, &end);
      if (end == p) { bad = true; break; }
      r->c.v.push_back(x); p = end;
      if (*p == ',') { ++p; continue; }
      break;
    }
    if (r->c.v.size() - n0 != r->c.cols.size()) bad = true;
    ++r->c.rows;
  };
  while (fgets(buf, sizeof buf, f)) {
    line += buf;
    if (!line.empty() && line.back() == '\n') { line.pop_back(); if (!line.empty() && line.back() == '\r') line.pop_back(); handle(line); line.clear(); }
  }
  if (!line.empty()) handle(line);
  fclose(f);
  // stanRead: `error "CSV file contains no data"` (Stan.hs:371)
  if (!have_header) { delete r; return stk_set_error_(STK_ERR_ARG, "CSV file contains no data"); }
  if (bad) { delete r; return stk_set_error_(STK_ERR_ARG, "Stan CSV: a row does not match the header"); }
  *out = r;
  return STK_OK;
}
extern "C" int stk_csv_cols(const stk_csv* c) { return (int)c->c.cols.size(); }
extern "C" long long stk_csv_rows(const stk_csv* c) { return c->c.rows; }
extern "C" const char* stk_csv_col_name(const stk_csv* c, int i) { return c->c.cols[i].c_str(); }
extern "C" const double* stk_csv_values(const stk_csv* c) { return c->c.v.data(); }   // row-major [rows][cols]
extern "C" const char* stk_csv_comments(const stk_csv* c) { return c->c.comments.c_str(); }
extern "C" void stk_csv_free(stk_csv* c) { delete c; }

// ---- .npy and headerless CSV matrices (Data/Expression/Const/IO.hs) -----------------------------------------------------
// writeNPy (IO.hs:28-45): NumPy format 1.0, '<f8', C order; the dictionary is padded with spaces so that
// 10 + len(header) is a multiple of 16, exactly as the reference computes it (pad = (10 + len(dic) + 1) mod 16).
extern "C" int stk_npy_write(const char* path, const double* values, int ndim, const long long* dims) {
  if (!path || ndim < 0 || (ndim > 0 && !dims)) return stk_set_error_(STK_ERR_ARG, "stk_npy_write: bad argument");
  long long n = 1;
  std::string sh;
  for (int i = 0; i < ndim; ++i) {
    if (dims[i] < 0) return stk_set_error_(STK_ERR_ARG, "stk_npy_write: negative dimension");
    n *= dims[i];
    if (i) sh += ",";
    sh += std::to_string(dims[i]);
  }
  if (ndim == 1) sh += ",";
  if (n > 0 && !values) return stk_set_error_(STK_ERR_ARG, "stk_npy_write: null values");
  std::string dic = "{'descr': '<f8', 'fortran_order': False, 'shape': (" + sh + "), }";
  const size_t pad = (10 + dic.size() + 1) % 16;
  std::string header = dic + (pad ? std::string(16 - pad, ' ') : std::string()) + "\n";
  FILE* f = fopen(path, "wb");
  if (!f) return stk_set_error_(STK_ERR_ARG, (std::string("cannot open ") + path + ": " + strerror(errno)).c_str());
  const unsigned char magic[10] = {0x93, 'N', 'U', 'M', 'P', 'Y', 1, 0, (unsigned char)(header.size() & 0xff), (unsigned char)(header.size() >> 8)};
  bool ok = fwrite(magic, 1, 10, f) == 10 && fwrite(header.data(), 1, header.size(), f) == header.size();
  // putFloat64le: the image's hosts are little-endian, so the in-memory doubles are the file's bytes
  if (ok && n > 0) ok = fwrite(values, sizeof(double), (size_t)n, f) == (size_t)n;
  if (fclose(f) != 0 || !ok) return stk_set_error_(STK_ERR_ARG, "stk_npy_write: write failed");
  return STK_OK;
}

// readRealMatrix (IO.hs:22-26): CSV.decode NoHeader into rows of doubles.  Returned as an stk_csv without column names
// (stk_csv_col_name gives ""); every row must have the first row's length.
extern "C" int stk_real_matrix_read(const char* path, stk_csv** out) {
  if (!path || !out) return stk_set_error_(STK_ERR_ARG, "stk_real_matrix_read: null argument");
  FILE* f = fopen(path, "r");
  if (!f) return stk_set_error_(STK_ERR_ARG, (std::string("cannot open ") + path + ": " + strerror(errno)).c_str());
  stk_csv* r = new stk_csv(); r->c.rows = 0;
  std::string line; char buf[1 << 16];
  bool bad = false;
  auto handle = [&](const std::string& ln) {
    if (ln.empty()) return;
    const char* p = ln.c_str(); size_t n0 = r->c.v.size();
    for (;;) {
      while (*p == ' ' || *p == '"') ++p;
      char* end; double x = strtod(p, &end);
      if (end == p) { bad = true; break; }
      r->c.v.push_back(x); p = end;
      while (*p == ' ' || *p == '"') ++p;
      if (*p == ',') { ++p; continue; }
      if (*p) bad = true;
      break;
    }
    const size_t w = r->c.v.size() - n0;
    if (r->c.rows == 0) r->c.cols.assign(w, std::string());
    else if (w != r->c.cols.size()) bad = true;
    ++r->c.rows;
  };
  while (fgets(buf, sizeof buf, f)) {
    line += buf;
    if (!line.empty() && line.back() == '\n') { line.pop_back(); if (!line.empty() && line.back() == '\r') line.pop_back(); handle(line); line.clear(); }
  }
  if (!line.empty()) handle(line);
  fclose(f);
  if (bad) { delete r; return stk_set_error_(STK_ERR_ARG, "readRealMatrix: a field is not a number or the rows are ragged"); }
  *out = r;
  return STK_OK;
}
