// stkc.cpp — the `Language.CUDA` lowering: Stochaskell IR (a probabilistic program printed by
// `showPBlock`, Data/Program.hs:181-187) -> CUDA source text of a module for libstk_cuda
// (descriptor + kernels over stk_device.cuh), compiled by stk_compile (NVRTC, sm_100a).
//
// This is the counterpart of Language.CC's code generator, construct by construct:
//   ccProgram  CC.hs:398-432   -> emit_module   (prelude, state load, transition loop, state store)
//   ccDAG      CC.hs:327-346   -> Gen::body     (one statement per node, dependency order)
//   ccNode     CC.hs:123-274   -> Gen::def      (Apply / Array / FoldScan / Case templates)
//   ccPNode'   CC.hs:276-321   -> Gen::pnode    (draws, SubrangeT rejection loop, Switch over a union)
//   ccRead/ccPrint :353-396    -> the SoA state fields of the descriptor (no text round trip)
// The input is the text the reference itself embeds at the top of every generated C++ file
// (CC.hs:400-402) and prints from `runStep` (Program.hs:1351): grammar in SURVEY.md Appendix A
// (showLet' Expression.hs:207-213, showNode :221-258, showNodeRef :141-161, showType :437-453,
// Show PNode Program.hs:133-161).  Kernel inputs are the Dummy ids `i_<L>_<k>` (L = 0 for
// compileCC, 9 for runStep); their types are not part of the printed IR and are passed beside it.
//
// Execution shape of the generated module: ONE THREAD PER CHAIN, the whole transition loop inside
// the kernel, state in registers/local memory between transitions.  Philox counter of a draw of PNode x_<ns>_<L>_<P>:
//   c0 = global chain, c1 = transition (low word), c2 = transition (high word) ^ (L << 24 | P << 12 | enclosing chain PNode + 1),
//   c3 = element index of a joint << 16 | block inside the draw;   key = seed, its high word ^ iteration of the enclosing chain
// so every PNode (and every element, and every iteration of a `chain`) owns a counter range of 2^16 blocks that no other
// draw can reach — a draw is a pure function of (seed, chain, transition, PNode), whatever the sharding.
// Hand-tuned builtins (models/*.cu) exist for the benchmark configs; this is the generic path.
#include "../../include/stk_cuda.h"

#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <functional>
#include <map>
#include <memory>
#include <set>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace stkc {

struct Error : std::runtime_error { using std::runtime_error::runtime_error; };
[[noreturn]] static void die(const std::string& m) { throw Error(m); }

// ---------------------------------------------------------------------------------------------
// types (showType, Expression.hs:437-453)
// ---------------------------------------------------------------------------------------------
struct Type {
  enum K { Z, R, B, Vec, Mat, Tuple, Union, Unknown } k = Unknown;
  std::vector<Type> sub;  // Vec/Mat: element; Tuple: members; Union: one Tuple per constructor
  bool cov = false;       // printed as `cov_matrix T` (ArrayT (Just "cov_matrix"), Expression.hs:449): Stan declares it cov_matrix[K]
  bool scalar() const { return k == Z || k == R || k == B; }
  bool integral() const { return k == Z || k == B; }
  std::string str() const {
    switch (k) {
      case Z: return "Z"; case R: return "R"; case B: return "B";
      case Vec: return sub[0].str() + "Vec"; case Mat: return sub[0].str() + "Mat";
      case Tuple: { std::string s = "("; for (size_t i = 0; i < sub.size(); ++i) s += (i ? ", " : "") + sub[i].str(); return s + ")"; }
      case Union: { std::string s = "Union["; for (size_t i = 0; i < sub.size(); ++i) s += (i ? ", " : "") + sub[i].str(); return s + "]"; }
      default: return "???";
    }
  }
  bool operator==(const Type& o) const { return str() == o.str(); }
};
static Type scalarT(Type::K k) { Type t; t.k = k; return t; }
static Type vecT(Type e) { Type t; t.k = Type::Vec; t.sub.push_back(e); return t; }

// ---------------------------------------------------------------------------------------------
// tokens
// ---------------------------------------------------------------------------------------------
struct Tok { std::string s; int line, col; bool first; enum K { Id, Num, Sym, End } k; };

static std::vector<Tok> lex(const std::string& src) {
  std::vector<Tok> out;
  int line = 1, col = 0; bool first = true;
  size_t i = 0, n = src.size();
  auto push = [&](const std::string& s, Tok::K k, int c) { out.push_back({s, line, c, first, k}); first = false; };
  while (i < n) {
    char ch = src[i];
    if (ch == '\n') { ++line; col = 0; first = true; ++i; continue; }
    if (ch == ' ' || ch == '\t' || ch == '\r') { ++col; ++i; continue; }
    if (ch == '/' && i + 1 < n && src[i + 1] == '*' && first) { i += 2; col += 2; continue; }   // the CC comment frame
    if (ch == '*' && i + 1 < n && src[i + 1] == '/' && first) break;
    int c0 = col;
    auto adv = [&](size_t k) { i += k; col += (int)k; };
    bool prev_space = i == 0 || strchr(" \t\n[,(", src[i - 1]) != nullptr;
    if (isdigit((unsigned char)ch) || (ch == '-' && prev_space && i + 1 < n && isdigit((unsigned char)src[i + 1]))) {
      size_t j = i + 1;
      while (j < n && isdigit((unsigned char)src[j])) ++j;
      // a fraction, but not the `...` of a range and not the `.T_F` of an Extract
      if (j + 1 < n && src[j] == '.' && isdigit((unsigned char)src[j + 1])) {
        size_t k = j + 1; while (k < n && isdigit((unsigned char)src[k])) ++k;
        if (!(k < n && src[k] == '_')) j = k;
      }
      if (j < n && (src[j] == 'e' || src[j] == 'E')) {
        size_t k = j + 1; if (k < n && (src[k] == '-' || src[k] == '+')) ++k;
        if (k < n && isdigit((unsigned char)src[k])) { while (k < n && isdigit((unsigned char)src[k])) ++k; j = k; }
      }
      std::string s = src.substr(i, j - i); adv(j - i); push(s, Tok::Num, c0); continue;
    }
    if (isalpha((unsigned char)ch) || ch == '_') {
      size_t j = i + 1;
      while (j < n && (isalnum((unsigned char)src[j]) || src[j] == '_' || src[j] == '\'')) ++j;
      std::string s = src.substr(i, j - i); adv(j - i);
      if (ch == '-' ) {}
      push(s, Tok::Id, c0); continue;
    }
    if (ch == '-' && prev_space && src.compare(i, 9, "-Infinity") == 0) { adv(9); push("-Infinity", Tok::Num, c0); continue; }
    // punctuation and operators
    static const char* multi[] = {"...", "::", "<-", "->", "=>", "==", "/=", "<=", ">=", "&&s", "&&", "||", "**", "<>", "#>", "*>", "<.>", "<\\>", "+s", "*s", ".?", nullptr};
    bool done = false;
    for (int m = 0; multi[m]; ++m) {
      size_t L = strlen(multi[m]);
      if (src.compare(i, L, multi[m]) == 0) {
        // "+s" / "*s" only when a blank follows (otherwise `+` then an identifier starting with s)
        if ((!strcmp(multi[m], "+s") || !strcmp(multi[m], "*s") || !strcmp(multi[m], "&&s")) && !(i + L < n && (src[i + L] == ' ' || src[i + L] == '\n'))) continue;
        adv(L); push(multi[m], Tok::Sym, c0); done = true; break;
      }
    }
    if (done) continue;
    adv(1); push(std::string(1, ch), Tok::Sym, c0);
  }
  out.push_back({"", line + 1, 0, true, Tok::End});
  return out;
}

// ---------------------------------------------------------------------------------------------
// IR (NodeRef Expression.hs:67-78, Node :175-205, PNode Program.hs:44-86)
// ---------------------------------------------------------------------------------------------
struct Ref;
using RefP = std::shared_ptr<Ref>;
struct Body;
struct PCData;
struct Ref {
  // Cond: `{ c => v, .. }` (mutually exclusive guards, Expression.hs:79-80); BlockVec: `(block array (lo,hi) [(i,ref),..] :: T)`
  // (a 1-D BlockArray: the concatenation of its scalar / vector blocks); BlockMat: `block[[a,b],[c,d]]` (isBlockMatrix,
  // Expression.hs:100-104); PC: PartiallyConstrained `( key -> value | i <- lo...hi )` (:82, :158-160); Let: a whole
  // let-expression printed where a reference is expected (the DExpr bounds / keys / values inside a PartiallyConstrained)
  enum K { Var, Num, Index, Extract, TagOf, Data, List, Cond, BlockVec, BlockMat, PC, Let } k = Var;
  std::string name;          // Var: id; Num: literal text
  RefP base;                 // Index / Extract / TagOf
  std::vector<RefP> args;    // Index: subscripts (outermost first); Data: fields; List: elements; Cond: c1, v1, c2, v2, ..;
                             // BlockVec: blocks in order; BlockMat: blocks row-major (`field` rows x `tag` columns)
  int tag = 0, field = 0;    // Extract (.tag_field), Data (C<tag>), BlockMat (columns, rows)
  Type type;                 // Data: the annotated union type; BlockVec: its annotation
  std::shared_ptr<Body> let; // Let
  std::shared_ptr<PCData> pc;
};
struct Def;
struct PNode;
struct Body {                // `let defs in (do pnodes return rets | rets)`
  std::vector<Def> defs;
  std::vector<PNode> pnodes;
  std::vector<RefP> rets;
};
struct PCData {              // one `key -> value` entry over a 1-D index range
  std::shared_ptr<Body> key, val;
  std::string param; RefP lo, hi;
};
struct Alt { std::vector<std::string> inputs; Body body; int tag = -1; /* constructor alternative `C<tag> fields ->` */ };
struct Node {
  enum K { Apply, Array, Fold, Case, Function } k = Apply;   // Function: `\i.. -> body` (Expression.hs:204-205; interpreter only in the reference)
  std::string f;                                  // Apply name; Fold: foldl/foldr/scanl/scanr
  std::vector<RefP> args;                         // Apply args; Fold: seed, list; Case: head
  std::vector<std::string> params;                // Array: index ids; Fold: acc, elem
  std::vector<std::pair<RefP, RefP>> shape;       // Array bounds
  std::shared_ptr<Body> lam;                      // Array / Fold body
  std::vector<Alt> alts;                          // Case
  Type type;
};
struct Def { std::string id; Node node; };
struct PNode {
  enum K { Dist, Switch, Loop, Chain, Ordered } k = Dist;
  std::string id, dist;
  std::vector<RefP> args;     // Dist args; Switch: head; Chain: lo, hi, initial state
  std::vector<Alt> alts;      // Switch; Chain: one alternative, inputs = index and carried state
  std::vector<std::string> params; std::vector<std::pair<RefP, RefP>> shape; std::shared_ptr<PNode> inner; std::shared_ptr<Body> lam;  // Loop
  Type type;
  RefP sub_lo, sub_hi;        // SubrangeT bounds of the PNode's type (`truncated`): the draw is a rejection loop
};

// ---------------------------------------------------------------------------------------------
// parser
// ---------------------------------------------------------------------------------------------
struct Parser {
  std::vector<Tok> t; size_t p = 0;
  explicit Parser(const std::string& src) : t(lex(src)) {}
  const Tok& peek(size_t k = 0) const { return t[std::min(p + k, t.size() - 1)]; }
  bool is(const char* s, size_t k = 0) const { return peek(k).k != Tok::End && peek(k).s == s; }
  [[noreturn]] void fail(const std::string& m) const {
    die("IR parse error at line " + std::to_string(peek().line) + " col " + std::to_string(peek().col + 1) + " near `" + peek().s + "`: " + m);
  }
  Tok next() { Tok x = peek(); if (x.k == Tok::End) fail("unexpected end of input"); ++p; return x; }
  void expect(const char* s) { if (!is(s)) fail(std::string("expected `") + s + "`"); ++p; }
  bool accept(const char* s) { if (is(s)) { ++p; return true; } return false; }

  static bool ctor_name(const std::string& s, int* tag) {
    if (s.size() < 2 || s[0] != 'C') return false;
    for (size_t i = 1; i < s.size(); ++i) if (!isdigit((unsigned char)s[i])) return false;
    if (tag) *tag = atoi(s.c_str() + 1);
    return true;
  }

  Type type() {
    if (accept("(")) {
      Type tu; tu.k = Type::Tuple;
      if (!accept(")")) { do tu.sub.push_back(type()); while (accept(",")); expect(")"); }
      if (tu.sub.size() == 1) return tu.sub[0];
      return tu;
    }
    Tok x = next();
    if (x.s == "Z") return scalarT(Type::Z);
    if (x.s == "R") return scalarT(Type::R);
    if (x.s == "B") return scalarT(Type::B);
    if (x.s == "ZVec" || x.s == "RVec" || x.s == "BVec") return vecT(scalarT(x.s[0] == 'Z' ? Type::Z : x.s[0] == 'R' ? Type::R : Type::B));
    if (x.s == "ZMat" || x.s == "RMat") { Type m; m.k = Type::Mat; m.sub.push_back(scalarT(x.s[0] == 'Z' ? Type::Z : Type::R)); return m; }
    if (x.s == "vector") return vecT(type());
    if (x.s == "matrix" || x.s == "corr_matrix" || x.s == "cov_matrix") { Type m; m.k = Type::Mat; m.cov = x.s == "cov_matrix"; m.sub.push_back(type()); return m; }
    if (x.s == "Array" && is("[")) {  // `Array [(lo,hi)] T`: how the 2019 build behind rj-coal.ipynb printed a 1-D array type
      next(); expect("("); ref(); expect(","); ref(); expect(")");
      if (accept(",")) fail("multi-dimensional Array types are not supported by the CUDA lowering");
      expect("]");
      return vecT(type());
    }
    if (x.s == "Array") {            // Array lo...hi[, lo...hi] T   (showType of ArrayT Nothing, Expression.hs:444-445: what `joint` builds, Program.hs:711-721)
      ref(); expect("..."); ref();
      if (accept(",")) {
        ref(); expect("..."); ref();
        if (accept(",")) fail("Array types of more than two dimensions are not supported by the CUDA lowering");
        Type m; m.k = Type::Mat; m.sub.push_back(type()); return m;
      }
      return vecT(type());
    }
    if (x.s == "Subrange") {       // Subrange T lo hi (showType, Expression.hs:442-443); `Nothing` = unbounded
      Type b = type(); RefP lo = ref(), hi = ref();
      auto some = [](const RefP& r) { return !(r->k == Ref::Var && r->name == "Nothing"); };
      last_sub_lo = some(lo) ? lo : nullptr; last_sub_hi = some(hi) ? hi : nullptr;
      return b;
    }
    if (x.s == "Union") {
      expect("[");
      Type u; u.k = Type::Union;
      do {
        expect("(");
        Type tu; tu.k = Type::Tuple;
        if (!accept(")")) { do tu.sub.push_back(type()); while (accept(",")); expect(")"); }
        u.sub.push_back(tu);
      } while (accept(","));
      expect("]");
      return u;
    }
    --p; fail("unknown type");
  }

  // primary [.T_F | .?]* [! primary[.T_F]*]*
  RefP atom() {
    RefP r = std::make_shared<Ref>();
    if (is("{")) {                                                   // Cond: { c => v, c => v }
      next(); r->k = Ref::Cond;
      if (!accept("}")) {
        do { r->args.push_back(ref()); expect("=>"); r->args.push_back(ref()); } while (accept(","));
        expect("}");
      }
    } else if (is("block") && is("[", 1)) {                          // block[[a,b],[c,d]]
      next(); next(); r->k = Ref::BlockMat;
      int rows = 0, cols = -1;
      do {
        expect("["); int c = 0;
        do { r->args.push_back(ref()); ++c; } while (accept(","));
        expect("]");
        if (cols >= 0 && c != cols) fail("ragged block matrix");
        cols = c; ++rows;
      } while (accept(","));
      expect("]");
      r->field = rows; r->tag = cols;
    } else if (is("(") && is("let", 1)) {                            // PartiallyConstrained: ( key -> value | i <- lo...hi )
      next(); r->k = Ref::PC; r->pc = std::make_shared<PCData>();
      r->pc->key = std::make_shared<Body>(letexpr()); expect("->");
      r->pc->val = std::make_shared<Body>(letexpr());
      if (is(",")) fail("PartiallyConstrained references with several entries are not supported by the CUDA lowering");
      expect("|");
      Tok i = next(); if (i.k != Tok::Id) fail("expected an index id");
      r->pc->param = i.s; expect("<-"); r->pc->lo = ref_or_let(); expect("..."); r->pc->hi = ref_or_let();
      if (is(",")) fail("multi-dimensional PartiallyConstrained references are not supported by the CUDA lowering");
      expect(")");
    } else if (is("(") && is("block", 1) && is("array", 2)) {        // (block array ([1],[n]) [([1],a),([2],b),..] :: T)
      next(); next(); next(); r->k = Ref::BlockVec;
      expect("("); expect("["); next(); if (is(",")) fail("only 1-D block arrays are supported"); expect("]"); expect(",");
      expect("["); next(); expect("]"); expect(")");
      expect("[");
      do { expect("("); expect("["); next(); expect("]"); expect(","); r->args.push_back(ref()); expect(")"); } while (accept(","));
      expect("]");
      expect("::"); r->type = type(); expect(")");
    } else if (accept("(")) {
      int tag;
      if (peek().k == Tok::Id && ctor_name(peek().s, &tag)) {       // (C<k> fields :: Union[..])
        next(); r->k = Ref::Data; r->tag = tag;
        while (!is("::")) r->args.push_back(ref());
        expect("::"); r->type = type(); expect(")");
      } else { r = ref(); expect(")"); }
    } else if (accept("[")) {
      r->k = Ref::List;
      if (!accept("]")) { do r->args.push_back(ref()); while (accept(",")); expect("]"); }
    } else {
      Tok x = next();
      if (x.k == Tok::Num || x.s == "Infinity" || x.s == "NaN") { r->k = Ref::Num; r->name = x.s; }
      else if (x.k == Tok::Id) { r->k = Ref::Var; r->name = x.s; }
      else { --p; fail("expected a reference"); }
    }
    for (;;) {
      if (is(".") && peek(1).k == Tok::Num && peek(1).col == peek().col + 1) {   // .T_F  (lexed as `.`, `T`, then `_F` glued: handle both spellings)
        next(); Tok a = next();
        RefP e = std::make_shared<Ref>(); e->k = Ref::Extract; e->base = r; e->tag = atoi(a.s.c_str());
        if (peek().k == Tok::Id && peek().s[0] == '_' && peek().col == a.col + (int)a.s.size()) { e->field = atoi(next().s.c_str() + 1); }
        else fail("malformed Extract");
        r = e; continue;
      }
      if (is(".?")) { next(); RefP e = std::make_shared<Ref>(); e->k = Ref::TagOf; e->base = r; r = e; continue; }
      break;
    }
    return r;
  }
  // a let-expression where a reference is expected (DExpr printed inside a PartiallyConstrained reference)
  Body letexpr() {
    Body b;
    if (accept("let")) { while (!is("in")) b.defs.push_back(def()); expect("in"); }
    b.rets.push_back(ref());
    return b;
  }
  RefP ref_or_let() {
    if (!is("let")) return ref();
    RefP r = std::make_shared<Ref>(); r->k = Ref::Let; r->let = std::make_shared<Body>(letexpr());
    return r;
  }
  RefP ref() {
    RefP r = atom();
    if (is("!")) {
      RefP ix = std::make_shared<Ref>(); ix->k = Ref::Index; ix->base = r;
      while (accept("!")) ix->args.push_back(atom());
      return ix;
    }
    return r;
  }

  static bool binop(const std::string& s) {
    static const char* ops[] = {"+", "-", "*", "/", "**", "==", "/=", "<", "<=", ">", ">=", "&&", "||", "<>", "#>", "*>", "<.>", "<\\>", nullptr};
    for (int i = 0; ops[i]; ++i) if (s == ops[i]) return true;
    return false;
  }

  // let-block body: [let defs in] (do ... | ret | [rets])
  Body body(bool list_rets) {
    Body b;
    if (accept("let")) {
      while (!is("in")) b.defs.push_back(def());
      expect("in");
    }
    if (accept("do")) {
      while (!is("return")) b.pnodes.push_back(pnode());
      expect("return");
      if (accept("[")) { if (!accept("]")) { do b.rets.push_back(ref()); while (accept(",")); expect("]"); } }
      else b.rets.push_back(ref());                  // a Chain body returns one NodeRef, printed bare
    } else if (list_rets && is("[")) {
      next();
      if (!accept("]")) { do b.rets.push_back(ref()); while (accept(",")); expect("]"); }
    } else {
      b.rets.push_back(ref());
    }
    return b;
  }

  void array_params(std::vector<std::string>* ids, std::vector<std::pair<RefP, RefP>>* sh) {
    do {
      Tok i = next(); if (i.k != Tok::Id) fail("expected an index id");
      expect("<-"); RefP lo = ref(); expect("..."); RefP hi = ref();
      ids->push_back(i.s); sh->push_back({lo, hi});
    } while (accept(","));
  }

  std::vector<Alt> alts(bool probabilistic) {
    std::vector<Alt> out;
    for (;;) {
      const Tok& h = peek();
      int tag;
      bool int_alt = h.k == Tok::Num && is("->", 1);
      bool ctor_alt = h.k == Tok::Id && ctor_name(h.s, &tag) && h.first;
      if (!int_alt && !ctor_alt) break;
      Alt a; next();
      if (ctor_alt) a.tag = tag;
      if (ctor_alt) while (!is("->")) { Tok i = next(); if (i.k != Tok::Id) fail("expected a constructor field id"); a.inputs.push_back(i.s); }
      expect("->");
      a.body = body(!probabilistic);
      out.push_back(std::move(a));
    }
    if (out.empty()) fail("case/switch without alternatives");
    return out;
  }

  Def def() {
    Def d; Tok id = next();
    if (id.k != Tok::Id) { --p; fail("expected a definition"); }
    d.id = id.s; expect("=");
    Node& n = d.node;
    if (is("[") ) {                                  // Array comprehension
      next(); n.k = Node::Array;
      n.lam = std::make_shared<Body>(body(false));
      expect("|"); array_params(&n.params, &n.shape); expect("]");
    } else if (is("foldl") || is("foldr") || is("scanl") || is("scanr") || is("scan")) {
      n.k = Node::Fold; n.f = next().s;
      n.args.push_back(ref()); n.args.push_back(ref());
      expect("$"); expect("\\");
      n.params.push_back(next().s); n.params.push_back(next().s);   // printed (acc elem) for foldl, showNode Expression.hs:228-241
      expect("->");
      n.lam = std::make_shared<Body>(body(false));
    } else if (accept("case")) {
      n.k = Node::Case; n.args.push_back(ref()); expect("of");
      n.alts = alts(false);
    } else if (is("\\")) {                           // Function: \i_1_1 .. -> body
      next(); n.k = Node::Function;
      while (!is("->")) { Tok i = next(); if (i.k != Tok::Id) fail("expected a lambda parameter"); n.params.push_back(i.s); }
      expect("->");
      n.lam = std::make_shared<Body>(body(false));
    } else {
      n.k = Node::Apply;
      size_t save = p;
      bool infix = false;
      if (!(peek().k == Tok::Sym && (is("+s") || is("*s") || is("&&s")))) {
        if (peek().k != Tok::Id || peek().s.find('_') != std::string::npos || is("(") || peek(1).k == Tok::Sym) {
          // try `a op b`
          try {
            RefP a = ref();
            if (peek().k == Tok::Sym && binop(peek().s)) {
              n.f = next().s; n.args.push_back(a); n.args.push_back(ref()); infix = true;
            }
          } catch (const Error&) {}
        }
      }
      if (!infix) {
        p = save;
        Tok f = next(); n.f = f.s;
        int line = f.line;
        while (!is("::") && peek().k != Tok::End && peek().line == line) n.args.push_back(ref());
      }
    }
    if (accept("::")) n.type = type();
    return d;
  }

  RefP last_sub_lo, last_sub_hi;   // bounds of the most recently parsed Subrange type
  std::string input_name;          // `\input ->`: the kernel's inputs are input.0_0, input.0_1, ..

  PNode pnode() {
    PNode q; Tok id = next();
    if (id.k != Tok::Id) { --p; fail("expected `x <- ...`"); }
    q.id = id.s; expect("<-");
    if (accept("switch")) {
      q.k = PNode::Switch; q.args.push_back(ref()); expect("of");
      q.alts = alts(true);
    } else if (accept("[")) {                          // joint: [ [let .. in] dist args :: P T | i <- lo...hi ] :: T
      q.k = PNode::Loop;
      q.lam = std::make_shared<Body>();
      if (accept("let")) { while (!is("in")) q.lam->defs.push_back(def()); expect("in"); }
      q.inner = std::make_shared<PNode>();
      q.inner->id = q.id;
      dist(q.inner.get());
      expect("|"); array_params(&q.params, &q.shape); expect("]");
      expect("::"); q.type = type();
    } else if (accept("chain")) {                      // chain (lo,hi) x0 $ \i j -> [let .. in] do .. return r   (Program.hs:154-157)
      q.k = PNode::Chain;
      expect("("); q.args.push_back(ref()); expect(","); q.args.push_back(ref()); expect(")");
      q.args.push_back(ref());
      expect("$"); expect("\\");
      Alt a; a.inputs.push_back(next().s); a.inputs.push_back(next().s);
      expect("->");
      a.body = body(false);
      q.alts.push_back(std::move(a));
    } else if (is("unfoldP") || is("foldP") || is("AnamorphismP") || is("CatamorphismP")) {
      fail("RecursionScheme PNodes are not supported by the CUDA lowering");
    } else {
      dist(&q);
    }
    return q;
  }
  void dist(PNode* q) {
    Tok f = next(); if (f.k != Tok::Id) { --p; fail("expected a distribution name"); }
    q->k = PNode::Dist; q->dist = f.s;
    if (is("(") && peek(1).k == Tok::Id && !ctor_name(peek(1).s, nullptr) && peek(1).s.find('_') == std::string::npos &&
        !is("block", 1) && !is("let", 1)) {
      // HODist: `d (inner dist args :: P T) args :: P T'` (Show PNode, Program.hs:146)
      if (f.s != "orderedSample") fail("higher-order distributions (" + f.s + ") are not supported by the CUDA lowering");
      next();
      q->k = PNode::Ordered;
      q->inner = std::make_shared<PNode>();
      dist(q->inner.get());
      expect(")");
    }
    while (!is("::")) q->args.push_back(ref());
    expect("::");
    if (!(peek().k == Tok::Id && peek().s == "P")) fail("expected `P <type>`");
    next();
    last_sub_lo = last_sub_hi = nullptr;
    q->type = type();
    q->sub_lo = last_sub_lo; q->sub_hi = last_sub_hi;
  }

  Body program() {
    if (is("[") && peek(1).s == "runStep") { next(); next(); expect("]"); }
    if (is("\\")) { next(); input_name = next().s; expect("->"); }      // `\input ->` header of Show (a -> P b)
    Body b = body(true);
    if (peek().k != Tok::End) fail("trailing input after the program");
    return b;
  }
};

// ---------------------------------------------------------------------------------------------
// state layout: kernel input k of type T -> descriptor fields (ccRead/ccPrint, CC.hs:353-396)
//   Z/B  -> in<k>           i32 [1]
//   R    -> in<k>           f64 [1]
//   vector -> in<k> f64|i32 [cap] + in<k>_len i32 [1]
//   Union[..] -> in<k>_tag i32 + one slot per (scalar kind | vector kind, occurrence) shared by all
//              constructors: in<k>_r<j>, in<k>_z<j>, in<k>_rv<j>(+_len), in<k>_zv<j>(+_len)
// ---------------------------------------------------------------------------------------------
struct Slot { std::string name; Type type; };
struct UnionLayout { std::vector<Slot> slots; std::vector<std::vector<int>> ctor; };  // ctor[c][field] -> slot

static UnionLayout union_layout(const std::string& prefix, const Type& u) {
  UnionLayout L;
  std::map<std::string, std::vector<int>> byKind;
  for (const Type& c : u.sub) {
    std::map<std::string, int> used;
    std::vector<int> f;
    for (const Type& t : c.sub) {
      std::string kind;
      if (t.k == Type::R) kind = "r"; else if (t.integral()) kind = "z";
      else if (t.k == Type::Vec && t.sub[0].k == Type::R) kind = "rv";
      else if (t.k == Type::Vec && t.sub[0].integral()) kind = "zv";
      else die("union field of type " + t.str() + " is not supported by the CUDA lowering");
      int j = used[kind]++;
      auto& v = byKind[kind];
      if ((int)v.size() <= j) {
        Slot s; s.name = prefix + "_" + kind + std::to_string(j);
        s.type = (kind == "r") ? scalarT(Type::R) : (kind == "z") ? scalarT(Type::Z) : vecT(scalarT(kind == "rv" ? Type::R : Type::Z));
        v.push_back((int)L.slots.size()); L.slots.push_back(s);
      }
      f.push_back(v[j]);
    }
    L.ctor.push_back(f);
  }
  return L;
}

// ---------------------------------------------------------------------------------------------
// code generator
// ---------------------------------------------------------------------------------------------
// a matrix kept as an expression (Eigen-style lazy evaluation): dimensions and an element accessor over C index expressions
// (1-based); `prepare` emits, at the point of use, whatever the accessor needs in storage (products are materialised)
struct Mat {
  std::string rows, cols;
  std::function<std::string(const std::string&, const std::string&)> at;
  std::function<void()> prepare;
  bool zeros = false, eye = false;
};
struct VA {                // anything that can be read like a vector: length + element accessor (1-based index expression)
  std::string len;
  std::function<std::string(const std::string&)> at;
};
struct Val {               // what an IR id denotes in the emitted code
  Type type;
  std::string expr;        // scalar: C expression; vector: name of a VecD/VecI view; union: name of a struct variable
  const Node* array = nullptr;   // un-materialised Array node (inlined at its use)
  std::vector<std::string> tuple;  // Case with several returns: one expr per member
  std::shared_ptr<Mat> mat;        // matrix-typed definition
  std::shared_ptr<VA> lazy;        // vector-typed definition kept as an accessor (a row of a matrix)
  const Node* fn = nullptr;        // Function node
  int mat_scope = -1;              // materialised array: id of the C scope that holds its storage
};
struct NotInlinable {};    // thrown when a lambda body that is being turned into one C expression needs a statement

struct Gen {
  std::ostringstream out, decls;   // decls: per-thread vector buffers, hoisted to function scope
  int ind = 2;
  int veccap = 64;
  int uniq = 0;
  std::vector<std::map<std::string, Val>> scopes;
  std::map<std::string, UnionLayout> unions;   // by type string
  std::map<std::string, std::string> union_cname;

  int lanes = 1;                   // lanes per chain (option `lanes`): additive folds are split over them and combined by shuffles
  bool inline_mode = false;        // lowering a lambda body into ONE expression: no statements may be emitted
  int rename_depth = 0;            // > 0 inside a nested let-expression (its ids v_0_k clash with the enclosing block's)
  std::vector<int> scope_ids; int next_scope_id = 0;
  void line(const std::string& s) { if (inline_mode) throw NotInlinable(); out << std::string(ind, ' ') << s << "\n"; }
  void push() { scopes.emplace_back(); scope_ids.push_back(next_scope_id++); }
  void pop() { scopes.pop_back(); scope_ids.pop_back(); }
  bool scope_live(int id) const { for (int s : scope_ids) if (s == id) return true; return false; }
  void bind(const std::string& id, const Val& v) { scopes.back()[id] = v; }
  const Val& lookup(const std::string& id) const {
    for (size_t i = scopes.size(); i-- > 0;) { auto it = scopes[i].find(id); if (it != scopes[i].end()) return it->second; }
    die("unbound identifier " + id + " (is it a kernel input whose type was not given?)");
  }
  std::string fresh(const std::string& b) { return b + "_" + std::to_string(uniq++); }

  static std::string ctype(const Type& t) {
    if (t.k == Type::R) return "double";
    if (t.integral()) return "int";
    if (t.k == Type::Vec) return t.sub[0].k == Type::R ? "VecD" : "VecI";
    die("no C type for " + t.str());
  }
  std::string union_struct(const Type& u) {
    std::string key = u.str();
    auto it = union_cname.find(key);
    if (it != union_cname.end()) return it->second;
    std::string nm = "U" + std::to_string(union_cname.size());
    union_cname[key] = nm; unions[key] = union_layout("f", u);
    return nm;
  }

  static std::string num(const std::string& s, bool as_real) {
    if (s == "Infinity") return "stk::pos_inf()";
    if (s == "-Infinity") return "stk::neg_inf()";
    if (s == "NaN") return "(0.0/0.0)";
    bool isreal = s.find('.') != std::string::npos || s.find('e') != std::string::npos || s.find('E') != std::string::npos;
    if (as_real && !isreal) return s + ".0";
    return s;
  }

  Type type_of(const RefP& r) {
    switch (r->k) {
      case Ref::Num: { bool isreal = r->name.find_first_of(".eEIN") != std::string::npos; return scalarT(isreal ? Type::R : Type::Z); }
      case Ref::Var: return lookup(r->name).type;
      case Ref::Index: {
        Type b = type_of(r->base);
        if (b.k == Type::Mat) { if (r->args.size() == 1) return vecT(b.sub[0]); return b.sub[0]; }
        if (b.k != Type::Vec) die("indexing a non-vector (" + b.str() + ")");
        return b.sub[0];
      }
      case Ref::Extract: { Type b = type_of(r->base); if (b.k == Type::Tuple) return b.sub.at(r->field); if (r->field == 0) return b; die("Extract of a non-tuple"); }
      case Ref::TagOf: return scalarT(Type::Z);
      case Ref::Data: return r->type;
      case Ref::List: {
        if (!r->args.empty() && r->args[0]->k == Ref::List) { Type m; m.k = Type::Mat; m.sub.push_back(scalarT(Type::R)); return m; }
        return vecT(r->args.empty() ? scalarT(Type::R) : type_of(r->args[0]));
      }
      case Ref::Cond: return r->args.size() >= 2 ? type_of(r->args[1]) : scalarT(Type::R);
      case Ref::BlockVec: return r->type.k == Type::Vec ? r->type : vecT(scalarT(Type::R));
      case Ref::BlockMat: { Type m; m.k = Type::Mat; m.sub.push_back(scalarT(Type::R)); return m; }
      case Ref::PC: return vecT(scalarT(Type::R));
      case Ref::Let: return scalarT(Type::Z);
    }
    return Type();
  }

  Val* find_val(const std::string& name) {
    for (size_t i = scopes.size(); i-- > 0;) { auto it = scopes[i].find(name); if (it != scopes[i].end()) return &it->second; }
    return nullptr;
  }

  // scalar / view expression for a reference (ccNodeRef, CC.hs:47-60)
  // An Array comprehension used as a value (returned, stored, passed to a vector operation) is materialised once, where
  // it is first needed, into per-thread storage — the resize + loops of CC.hs:231-235; its other uses stay inlined.
  std::string materialise(const std::string& name) {
    if (inline_mode) throw NotInlinable();
    Val* vp = find_val(name);
    if (!vp->array && vp->mat_scope >= 0) return vp->expr;
    const Node& an = *vp->array;
    if (an.params.size() != 1) die("array " + name + " has " + std::to_string(an.params.size()) + " dimensions; only vectors can be materialised");
    Type et = vp->type.k == Type::Vec ? vp->type.sub[0] : scalarT(Type::R);
    if (!et.scalar()) die("array " + name + " of " + et.str() + " cannot be materialised");
    bool real = et.k == Type::R;
    prematerialise(*an.lam);                        // arrays its body reads by index get their storage before the loop
    std::string buf = fresh(name + "_buf"), lo = fresh("lo"), n = fresh("n"), vec = fresh(name + "_m");
    const std::string& iv = an.params[0];
    decls << "  " << (real ? "double" : "int") << " " << buf << "[STK_VECCAP];\n";
    line("const int " + lo + " = " + rv(an.shape[0].first) + ";");
    line("const int " + n + " = stk::cap_len(" + rv(an.shape[0].second) + " - " + lo + " + 1, STK_VECCAP, err_);");
    line("for (int " + iv + " = " + lo + "; " + iv + " < " + lo + " + " + n + "; ++" + iv + ") {");
    ind += 2; push();
    Val kv; kv.type = scalarT(Type::Z); kv.expr = iv; bind(iv, kv);
    for (const Def& ad : an.lam->defs) def(ad);
    line(buf + "[" + iv + " - " + lo + "] = " + rv(an.lam->rets.at(0), real) + ";");
    pop(); ind -= 2; line("}");
    Type vt = vecT(et);
    line("const " + ctype(vt) + " " + vec + "(" + buf + ", 1, " + n + ");");
    // remember the storage only when it was filled in the C scope the array was defined in (storage filled inside a loop
    // or branch body is out of scope afterwards: the array stays symbolic for later uses)
    for (size_t i = scopes.size(); i-- > 0;) {
      auto it = scopes[i].find(name);
      if (it == scopes[i].end()) continue;
      if (i + 1 == scopes.size()) { it->second.array = nullptr; it->second.type = vt; it->second.expr = vec; it->second.mat_scope = scope_ids[i]; }
      break;
    }
    return vec;
  }

  // Before a loop is emitted: every comprehension that the loop body reads (by index, as a whole vector, through a block
  // vector or a Function it calls) and that cannot be folded into one expression is given storage HERE, outside the loop.
  std::set<const Body*> premat_seen;
  void prematerialise_ref(const RefP& r) {
    if (!r) return;
    if (r->k == Ref::Var) {
      Val* v = find_val(r->name);
      if (!v) return;
      if (v->fn && v->fn->lam) { prematerialise(*v->fn->lam); return; }
      if (v->array) {
        prematerialise(*v->array->lam);
        if (v->array->params.size() == 1 && !pure_inlinable(*v->array)) {
          // only arrays defined in the current C scope keep their storage; others are materialised where they are read
          for (size_t i = scopes.size(); i-- > 0;) {
            auto it = scopes[i].find(r->name);
            if (it == scopes[i].end()) continue;
            if (i + 1 == scopes.size()) materialise(r->name);
            break;
          }
        }
      }
      return;
    }
    prematerialise_ref(r->base);
    for (auto& a : r->args) prematerialise_ref(a);
    if (r->let) prematerialise(*r->let);
  }
  void prematerialise(const Body& b) {
    if (inline_mode || premat_seen.count(&b)) return;
    premat_seen.insert(&b);
    for (const Def& d : b.defs) {
      for (auto& a : d.node.args) prematerialise_ref(a);
      if (d.node.lam && d.node.k != Node::Array && d.node.k != Node::Function) prematerialise(*d.node.lam);
      for (auto& al : d.node.alts) prematerialise(al.body);
    }
    for (auto& r : b.rets) prematerialise_ref(r);
    premat_seen.erase(&b);
  }

  // can element `idx` of the comprehension be written as ONE C expression (no statement needed)?
  bool pure_inlinable(const Node& an) {
    if (an.params.size() != 1) return false;
    std::string e;
    return try_inline_array_at(an, "0", &e);
  }
  bool try_inline_array_at(const Node& a, const std::string& idx, std::string* out_e, bool want_real = false) {
    if (a.params.size() != 1) return false;
    const bool saved = inline_mode; const size_t depth = scopes.size();
    inline_mode = true;
    try {
      push(); Val v; v.type = scalarT(Type::Z); v.expr = "(" + idx + ")"; bind(a.params[0], v);
      for (const Def& ad : a.lam->defs) def(ad);
      *out_e = rv(a.lam->rets.at(0), want_real);
      pop(); inline_mode = saved;
      return true;
    } catch (const NotInlinable&) {
      while (scopes.size() > depth) pop();
      inline_mode = saved;
      return false;
    }
  }

  // a 2-D comprehension `[ body | i <- a...b, j <- c...d ] :: matrix` (CC.hs:231-235 with two loops): a matrix expression whose
  // element is the body with the indices substituted when the body needs no statement, else filled into per-thread storage
  bool try_inline_array2_at(const Node& a, const std::string& i, const std::string& j, std::string* out_e) {
    const bool saved = inline_mode; const size_t depth = scopes.size();
    inline_mode = true;
    try {
      push();
      Val vi; vi.type = scalarT(Type::Z); vi.expr = "(" + i + ")"; bind(a.params[0], vi);
      Val vj; vj.type = scalarT(Type::Z); vj.expr = "(" + j + ")"; bind(a.params[1], vj);
      for (const Def& ad : a.lam->defs) def(ad);
      *out_e = rv(a.lam->rets.at(0), true);
      pop(); inline_mode = saved;
      return true;
    } catch (const NotInlinable&) {
      while (scopes.size() > depth) pop();
      inline_mode = saved;
      return false;
    }
  }
  void array2_def(const Def& d) {
    const Node& n = d.node;
    const Node* np = &n;
    if (!inline_mode) { for (auto& sh : n.shape) { need(sh.first); need(sh.second); } need_body(*n.lam); }
    std::string lo0 = rv(n.shape[0].first), hi0 = rv(n.shape[0].second), lo1 = rv(n.shape[1].first), hi1 = rv(n.shape[1].second);
    auto m = std::make_shared<Mat>();
    m->rows = "(" + hi0 + " - " + lo0 + " + 1)"; m->cols = "(" + hi1 + " - " + lo1 + " + 1)"; m->prepare = [] {};
    std::string probe;
    if (try_inline_array2_at(n, "1", "1", &probe)) {
      m->at = [this, np, lo0, lo1](const std::string& i, const std::string& j) {
        std::string e;
        if (!try_inline_array2_at(*np, lo0 == "1" ? i : "(" + i + ") + " + lo0 + " - 1", lo1 == "1" ? j : "(" + j + ") + " + lo1 + " - 1", &e)) die("internal: 2-D array stopped being inlinable");
        return e;
      };
    } else {
      if (inline_mode) throw NotInlinable();
      std::string buf = fresh(d.id + "_mat"), R = fresh("R"), C = fresh("C");
      decls << "  double " << buf << "[STK_MATCAP * STK_MATCAP];\n";
      line("const int " + R + " = stk::cap_len(" + m->rows + ", STK_MATCAP, err_, 3), " + C + " = stk::cap_len(" + m->cols + ", STK_MATCAP, err_, 3);");
      const std::string &iv = n.params[0], &jv = n.params[1];
      line("for (int " + iv + " = " + lo0 + "; " + iv + " < " + lo0 + " + " + R + "; ++" + iv + ") for (int " + jv + " = " + lo1 + "; " + jv + " < " + lo1 + " + " + C + "; ++" + jv + ") {");
      ind += 2; push();
      Val vi; vi.type = scalarT(Type::Z); vi.expr = iv; bind(iv, vi);
      Val vj; vj.type = scalarT(Type::Z); vj.expr = jv; bind(jv, vj);
      for (const Def& ad : n.lam->defs) def(ad);
      line(buf + "[(" + iv + " - " + lo0 + ") * " + C + " + " + jv + " - " + lo1 + "] = " + rv(n.lam->rets.at(0), true) + ";");
      pop(); ind -= 2; line("}");
      m = stored_mat(buf, R, C);
    }
    Val v; v.type = n.type; if (v.type.k != Type::Mat) { v.type = Type(); v.type.k = Type::Mat; v.type.sub.push_back(scalarT(Type::R)); }
    v.mat = m; bind(d.id, v);
  }

  // vector-like access to a reference: a state / buffer view, a comprehension (inlined or materialised), a row of a
  // matrix, a literal list
  VA va(const RefP& r) {
    VA a;
    if (r->k == Ref::Var) {
      const Val& v = lookup(r->name);
      if (v.lazy) return *v.lazy;
      if (v.array) {
        const Node* an = v.array;
        if (an->params.size() != 1) die("multi-dimensional array used as a vector");
        if (pure_inlinable(*an)) {
          a.len = "(" + rv(an->shape[0].second) + " - " + rv(an->shape[0].first) + " + 1)";
          std::string lo = rv(an->shape[0].first);
          a.at = [this, an, lo](const std::string& i) {
            std::string e;
            if (!try_inline_array_at(*an, lo == "1" ? i : "(" + i + ") + " + lo + " - 1", &e)) die("internal: array stopped being inlinable");
            return e;
          };
          return a;
        }
        std::string m = materialise(r->name);
        a.len = m + ".len"; a.at = [m](const std::string& i) { return m + ".at(" + i + ")"; };
        return a;
      }
    }
    if (r->k == Ref::List) {
      std::vector<std::string> es; for (auto& e : r->args) es.push_back(rv(e, true));
      a.len = std::to_string(es.size());
      a.at = [es](const std::string& i) {
        if (es.size() == 1) return es[0];
        std::string e = es.back();
        for (size_t k = es.size() - 1; k-- > 0;) e = "((" + i + ") == " + std::to_string(k + 1) + " ? " + es[k] + " : " + e + ")";
        return e;
      };
      return a;
    }
    if (r->k == Ref::Index && type_of(r->base).k == Type::Mat && r->args.size() == 1) {     // M!i: row i of a matrix
      std::shared_ptr<Mat> m = mat_of(r->base);
      if (!inline_mode) m->prepare();
      std::string row = rv(r->args[0]);
      a.len = m->cols; a.at = [m, row](const std::string& j) { return m->at(row, j); };
      return a;
    }
    if (r->k == Ref::BlockVec) {
      std::vector<VA> parts; std::vector<bool> scalar;
      for (auto& e : r->args) {
        Type t = type_of(e);
        if (t.scalar()) { VA p; p.len = "1"; std::string x = rv(e, true); p.at = [x](const std::string&) { return x; }; parts.push_back(p); }
        else if (t.k == Type::Vec) parts.push_back(va(e));
        else die("block vector element of type " + t.str() + " is not supported");
      }
      std::string tot;
      for (auto& p : parts) tot += (tot.empty() ? "" : " + ") + p.len;
      a.len = "(" + tot + ")";
      a.at = [parts](const std::string& i) {
        // walk the blocks: element i lives in the first block whose running end reaches i
        std::string e = parts.back().at("(" + i + ")" + offset_of(parts, parts.size() - 1));
        for (size_t k = parts.size() - 1; k-- > 0;) {
          std::string end; for (size_t j = 0; j <= k; ++j) end += (j ? " + " : "") + parts[j].len;
          e = "((" + i + ") <= " + end + " ? " + parts[k].at("(" + i + ")" + offset_of(parts, k)) + " : " + e + ")";
        }
        return e;
      };
      return a;
    }
    Type t = type_of(r);
    if (t.k != Type::Vec) die("a vector was expected, got " + t.str());
    std::string m = rv(r);
    a.len = m + ".len"; a.at = [m](const std::string& i) { return m + ".at(" + i + ")"; };
    return a;
  }
  static std::string offset_of(const std::vector<VA>& parts, size_t k) {
    std::string o; for (size_t j = 0; j < k; ++j) o += " - " + parts[j].len;
    return o;
  }

  // value of a guarded reference { c => v, .. } (mutually exclusive guards): nested selects; no guard true -> NaN / 0
  std::string cond_expr(const RefP& r, bool want_real) {
    Type t = type_of(r);
    bool real = want_real || t.k == Type::R;
    std::string e = real ? "stk::oob_value<double>()" : "0";
    for (size_t i = r->args.size(); i >= 2; i -= 2)
      e = "(" + rv(r->args[i - 2]) + " ? " + rv(r->args[i - 1], real) + " : " + e + ")";
    return e;
  }

  // element `idx` of a PartiallyConstrained array ( key -> value | i <- lo...hi ): the value whose key equals idx
  // (Expression.hs:82: an array known only at the positions its keys name; elsewhere unknown = NaN)
  std::string pc_lookup(const Ref& pr, const std::string& idx) {
    if (inline_mode) throw NotInlinable();
    const PCData& pc = *pr.pc;
    std::string res = fresh("pc"), i = fresh("pi"), ix = fresh("px");
    line("double " + res + " = stk::oob_value<double>();");
    line("{ const int " + ix + " = " + idx + ";");
    ind += 2; push(); ++rename_depth;
    std::string lo = rv(pc.lo), hi = rv(pc.hi);
    line("for (int " + i + " = " + lo + "; " + i + " <= " + hi + "; ++" + i + ") {");
    ind += 2; push();
    Val iv; iv.type = scalarT(Type::Z); iv.expr = i; bind(pc.param, iv);
    for (const Def& d : pc.key->defs) def(d);
    line("if (" + rv(pc.key->rets.at(0)) + " == " + ix + ") {");
    ind += 2; push();
    for (const Def& d : pc.val->defs) def(d);
    line(res + " = " + rv(pc.val->rets.at(0), true) + "; break;");
    pop(); ind -= 2; line("}");
    pop(); ind -= 2; line("}");
    --rename_depth; pop(); ind -= 2; line("}");
    return res;
  }

  std::string rv(const RefP& r, bool want_real = false) {
    switch (r->k) {
      case Ref::Num: return num(r->name, want_real);
      case Ref::Var: {
        const Val& v = lookup(r->name);
        if (v.array) return materialise(r->name);
        if (v.mat || v.fn) die(r->name + " is a matrix / function: it cannot be used as a scalar or vector value here");
        if (v.lazy) die(r->name + " is a lazily evaluated vector: it cannot be used as a stored vector here");
        if (want_real && v.type.integral()) return "(double)" + v.expr;
        return v.expr;
      }
      case Ref::Index: {
        RefP b = r->base;
        Type bt = type_of(b);
        if (bt.k == Type::Mat) {
          if (r->args.size() != 2) die("a matrix row used as a scalar");
          std::shared_ptr<Mat> m = mat_of(b);
          m->prepare();                                          // a product among its operands gets its storage here
          return m->at(rv(r->args[1]), rv(r->args[0]));       // printed innermost index first (showNodeRef reverses)
        }
        if (r->args.size() != 1) die("only 1-D indexing of vectors is supported");
        std::string e;
        if (b->k == Ref::PC) e = pc_lookup(*b, rv(r->args[0]));
        else if (b->k == Ref::BlockVec || b->k == Ref::List) e = va(b).at(rv(r->args[0]));
        else if (b->k == Ref::Var && lookup(b->name).lazy) e = lookup(b->name).lazy->at(rv(r->args[0]));
        else if (b->k == Ref::Var && lookup(b->name).array) {
          const Node& an = *lookup(b->name).array;
          std::string idx = rv(r->args[0]);
          // a comprehension read at one index: folded into the expression when its body needs no statement, else read
          // from its storage
          if (!try_inline_array_at(an, idx, &e, want_real && type_of(r).integral()))
            e = materialise(b->name) + ".at(" + idx + ")";
        } else e = rv(b) + ".at(" + rv(r->args[0]) + ")";
        if (want_real && type_of(r).integral()) return "(double)" + e;
        return e;
      }
      case Ref::Extract: {
        if (r->base->k == Ref::Var) {
          const Val& v = lookup(r->base->name);
          std::string e;
          if (!v.tuple.empty()) e = v.tuple.at(r->field);
          else if (r->field == 0) e = v.expr;
          else die("unsupported Extract");
          if (want_real && type_of(r).integral()) return "(double)" + e;
          return e;
        }
        die("unsupported Extract");
      }
      case Ref::TagOf: return rv(r->base) + ".tag";
      case Ref::Data: die("union construction is only supported as an ifThenElse branch or a return value");
      case Ref::List: die("list literal is only supported as a pmf argument");
      case Ref::Cond: return cond_expr(r, want_real);
      case Ref::BlockVec: die("a block vector can only be indexed or used as a probability table");
      case Ref::BlockMat: die("a block matrix can only be the argument of log_det");
      case Ref::PC: die("a partially constrained array can only be indexed");
      case Ref::Let: {
        // a whole let-expression in reference position: its definitions are value-numbered against the enclosing block
        // (they restate, under their own ids, nodes the enclosing DAG already holds) and emitted under fresh C names.
        // They are bound in the CURRENT scope: the caller (pc_lookup) opens one for them.
        ++rename_depth;
        for (const Def& d : r->let->defs) def(d);
        std::string e = rv(r->let->rets.at(0), want_real);
        --rename_depth;
        return e;
      }
    }
    return "";
  }

  // Array node evaluated at one index without materialising it (only for bodies that are a bare reference)
  std::string inline_array_at(const Node& a, const std::string& idx) {
    std::string e;
    if (!try_inline_array_at(a, idx, &e)) die("indexing a computed array is not supported here");
    return e;
  }

  // ---- union values ------------------------------------------------------------------------
  // assign `dst` (a union struct variable) from a Data ref or a union-typed variable
  void assign_union(const std::string& dst, const Type& ut, const RefP& src) {
    if (src->k == Ref::Data) {
      const UnionLayout& L = unions[ut.str()];
      if (src->tag >= (int)L.ctor.size() || L.ctor[src->tag].size() != src->args.size()) die("constructor C" + std::to_string(src->tag) + " does not fit " + ut.str());
      line(dst + ".tag = " + std::to_string(src->tag) + ";");
      for (size_t j = 0; j < src->args.size(); ++j) {
        const Slot& s = L.slots[L.ctor[src->tag][j]];
        line(dst + "." + s.name + " = " + rv(src->args[j], s.type.k == Type::R) + ";");
      }
    } else {
      line(dst + " = " + rv(src) + ";");
    }
  }

  // ---- nodes (ccNode, CC.hs:123-274) ----------------------------------------------------------
  static const char* c_binop(const std::string& f) {
    static const std::pair<const char*, const char*> ops[] = {{"+", "+"}, {"-", "-"}, {"*", "*"}, {"==", "=="}, {"/=", "!="},
      {">", ">"}, {">=", ">="}, {"<", "<"}, {"<=", "<="}, {"&&", "&&"}, {"||", "||"}};   // ccOperators, CC.hs:92-107
    for (auto& o : ops) if (f == o.first) return o.second;
    return nullptr;
  }

  // scalar log-densities: the log-space forms of stk_device.cuh behind the names of CC.hs:198-228
  std::string lpdf_call(const std::string& f, const std::vector<RefP>& args) {
    static const std::set<std::string> intx = {"poisson_lpdf", "neg_binomial_lpdf", "geometric_lpdf", "bernoulli_lpdf", "bernoulliLogit_lpdf", "discreteUniform_lpdf"};
    static const std::map<std::string, std::string> rename = {{"bernoulliLogit_lpdf", "bernoulli_logit_lpdf"}, {"discreteUniform_lpdf", "discrete_uniform_lpdf"}};
    static const std::set<std::string> known = {"normal_lpdf", "gamma_lpdf", "inv_gamma_lpdf", "geometric_lpdf", "neg_binomial_lpdf", "poisson_lpdf", "uniform_lpdf",
      "discreteUniform_lpdf", "bernoulli_lpdf", "bernoulliLogit_lpdf", "cauchy_lpdf", "logistic_lpdf", "beta_lpdf"};
    if (!known.count(f)) die("log-density `" + f + "` is not in the device library");
    auto it = rename.find(f);
    std::string e = "stk::" + (it == rename.end() ? f : it->second) + "(" + rv(args.at(0), !intx.count(f));
    for (size_t i = 1; i < args.size(); ++i) e += ", " + rv(args[i], f != "discreteUniform_lpdf");
    return e + ")";
  }

  // ---- value numbering -------------------------------------------------------------------------------------------------
  // A let-expression printed in reference position (the DExprs inside a PartiallyConstrained reference) restates, under
  // its own ids v_0_0.., nodes that the enclosing block already defines — the reference hash-conses each DAG separately
  // (Expression.hs:301-330).  Lowering it, every definition is first looked up by (function, lowered operands, type)
  // among the definitions already emitted in live scopes; a hit aliases the existing C variable instead of recomputing.
  std::string key_of(const RefP& r) {
    switch (r->k) {
      case Ref::Num: return r->name;
      case Ref::Var: {
        const Val& v = lookup(r->name);
        if (v.array || v.mat || v.fn || v.lazy) { char b[32]; snprintf(b, sizeof b, "@%p", (const void*)(v.array ? (const void*)v.array : v.fn ? (const void*)v.fn : v.mat ? (const void*)v.mat.get() : (const void*)v.lazy.get())); return b; }
        if (!v.tuple.empty()) { std::string k = "<"; for (auto& t : v.tuple) k += t + ","; return k + ">"; }
        return v.expr;
      }
      case Ref::Index: { std::string k = key_of(r->base); for (auto& a : r->args) k += "!" + key_of(a); return k; }
      case Ref::Extract: {
        if (r->base->k == Ref::Var) { const Val& v = lookup(r->base->name); if (!v.tuple.empty()) return v.tuple.at(r->field); if (r->field == 0) return v.expr; }
        return key_of(r->base) + "." + std::to_string(r->tag) + "_" + std::to_string(r->field);
      }
      case Ref::TagOf: return key_of(r->base) + ".?";
      case Ref::Cond: case Ref::List: case Ref::BlockVec: case Ref::BlockMat: case Ref::Data: {
        std::string k = std::string("#") + std::to_string((int)r->k) + "[";
        for (auto& a : r->args) k += key_of(a) + ",";
        return k + "]";
      }
      default: return "";
    }
  }
  std::string cse_key(const Node& n) {
    if (n.k == Node::Apply) {
      if (n.f == "getExternal") return "";
      std::string k = "#cse:" + n.f + "(";
      for (auto& a : n.args) { std::string ka = key_of(a); if (ka.empty()) return ""; k += ka + ";"; }
      return k + ")::" + n.type.str();
    }
    if (n.k == Node::Case) {
      std::string k = "#cse:case(" + key_of(n.args.at(0)) + ")";
      for (auto& al : n.alts) {
        if (!al.body.defs.empty() || !al.inputs.empty()) return "";
        k += "{";
        for (auto& r : al.body.rets) { std::string kr = key_of(r); if (kr.empty()) return ""; k += kr + ";"; }
        k += "}";
      }
      return k;
    }
    return "";
  }

  void def(const Def& d) {
    const Node& n = d.node;
    if (inline_mode || n.k == Node::Array || n.k == Node::Function || n.k == Node::Fold ||
        (n.k == Node::Apply && n.f == "getExternal")) { def_impl(d); return; }
    std::string key;
    try { key = cse_key(n); } catch (const Error&) { key.clear(); }
    if (rename_depth > 0) {
      if (!key.empty()) if (Val* hit = find_val(key)) { Val v = *hit; bind(d.id, v); return; }
      Def dd = d; dd.id = fresh(d.id);              // nested let: its ids clash with the enclosing block's, C names must not
      def_impl(dd);
      Val v = lookup(dd.id); bind(d.id, v);
    } else def_impl(d);
    if (!key.empty()) { Val v = lookup(d.id); bind(key, v); }
  }

  void def_impl(const Def& d) {
    const Node& n = d.node;
    Val v; v.type = n.type;
    if (n.k == Node::Array && n.params.size() == 2) { array2_def(d); return; }
    if (n.k == Node::Array) {                      // kept symbolic; consumers inline it (CC.hs:231-235 materialises)
      if (v.type.k == Type::Unknown) v.type = vecT(scalarT(Type::R));
      v.array = &n; bind(d.id, v); return;
    }
    if (n.k == Node::Function) { v.fn = &n; bind(d.id, v); return; }     // applied where it is used (poissonProcess_lpdf)
    if (n.k == Node::Fold) { fold(d); return; }
    if (n.k == Node::Case) { case_node(d); return; }
    const std::string& f = n.f;
    auto arg = [&](size_t i, bool real = false) { return rv(n.args.at(i), real); };
    bool rt = n.type.k == Type::R;
    if (n.type.k == Type::Mat) { mat_def(d); return; }
    if (f == "getExternal") {                      // alias of a PNode's value (CC.hs:124-125 emits the draw here)
      ensure_drawn(n.args.at(0)->name);
      bind(d.id, lookup(n.args.at(0)->name)); return;
    }
    if (n.type.k == Type::Union) {                 // only ifThenElse / id over unions
      std::string st = union_struct(n.type);
      line(st + " " + d.id + " = " + st + "();");
      if (f == "ifThenElse") {
        line("if (" + arg(0) + ") {"); ind += 2; assign_union(d.id, n.type, n.args[1]); ind -= 2;
        line("} else {"); ind += 2; assign_union(d.id, n.type, n.args[2]); ind -= 2; line("}");
      } else if (f == "id") assign_union(d.id, n.type, n.args[0]);
      else die("node `" + f + "` producing a union is not supported");
      v.expr = d.id; bind(d.id, v); return;
    }
    if (n.type.k == Type::Vec && n.type.sub[0].k == Type::R && (f == "+" || f == "-" || f == "*" || f == "/" || f == "*>" || f == "+s" || f == "negate") && !n.args.empty()) {
      // element-wise vector arithmetic (scalars broadcast; CC.hs:92-107 on Eigen vectors): kept as an accessor, evaluated by its consumer
      std::vector<VA> vs; std::vector<std::string> sc; std::string len;
      for (auto& a : n.args) {
        if (type_of(a).k == Type::Vec) { vs.push_back(va(a)); sc.push_back(""); if (len.empty()) len = vs.back().len; }
        else { vs.push_back(VA()); sc.push_back("(" + rv(a, true) + ")"); }
      }
      if (len.empty()) die("vector node `" + f + "` without a vector operand");
      const std::string op = f == "-" ? " - " : f == "/" ? " / " : (f == "*" || f == "*>") ? " * " : " + ";
      auto lz = std::make_shared<VA>(); lz->len = len;
      const bool neg = f == "negate";
      lz->at = [vs, sc, op, neg](const std::string& i) {
        std::string e;
        for (size_t k = 0; k < vs.size(); ++k) { std::string t = sc[k].empty() ? vs[k].at(i) : sc[k]; e = e.empty() ? t : "(" + e + op + t + ")"; }
        return neg ? "(-" + e + ")" : e;
      };
      v.lazy = lz; bind(d.id, v); return;
    }
    {
      static const std::map<std::string, std::string> vun = {{"square", "stk_sq"}, {"exp", "exp"}, {"log", "log"}, {"sqrt", "sqrt"}, {"abs", "fabs"}, {"tanh", "tanh"}, {"log1p", "log1p"}};
      auto vu = vun.find(f);
      if (n.type.k == Type::Vec && n.type.sub[0].k == Type::R && vu != vun.end() && n.args.size() == 1 && type_of(n.args[0]).k == Type::Vec) {
        VA src = va(n.args[0]);                       // element-wise unary function of a vector: an accessor, evaluated by its consumer
        const std::string fn = vu->second;
        auto lz = std::make_shared<VA>(); lz->len = src.len;
        lz->at = [src, fn](const std::string& i) { return fn == "stk_sq" ? "((" + src.at(i) + ") * (" + src.at(i) + "))" : fn + "(" + src.at(i) + ")"; };
        v.lazy = lz; bind(d.id, v); return;
      }
    }
    if (n.type.k == Type::Vec) { vector_node(d); return; }
    if (!n.type.scalar()) die("definition " + d.id + " has unsupported type " + n.type.str());
    std::string e;
    if (f == "+s" || f == "*s" || f == "&&s") {    // CC.hs:129-132: left-to-right in operand order
      const char* op = f == "+s" ? " + " : f == "*s" ? " * " : " && ";
      for (size_t i = 0; i < n.args.size(); ++i) e = i ? "(" + e + op + arg(i, rt) + ")" : arg(i, rt);
    } else if (f == "/" && rt) e = "(double)" + arg(0) + " / (double)" + arg(1);          // CC.hs:133-134
    else if (f == "/") e = arg(0) + " / " + arg(1);
    else if ((f == "==" || f == "/=") && type_of(n.args.at(0)).k == Type::Vec) {          // vector equality
      Type t0 = type_of(n.args.at(0));
      std::string a0 = vec_value(n.args.at(0)), a1 = vec_value(n.args.at(1));
      e = std::string(f == "==" ? "" : "!") + "stk::vec_eq(" + a0 + ", " + a1 + ")";
    }
    else if (const char* op = c_binop(f)) {
      bool anyreal = type_of(n.args.at(0)).k == Type::R || type_of(n.args.at(1)).k == Type::R;
      e = arg(0, anyreal) + " " + op + " " + arg(1, anyreal);
    }
    else if (f == "**") e = "pow(" + arg(0, true) + ", " + arg(1, true) + ")";             // CC.hs:155-156
    else if (f == "ifThenElse") e = arg(0) + " ? " + arg(1, rt) + " : " + arg(2, rt);      // CC.hs:141-143
    else if (f == "negate") e = "-(" + arg(0) + ")";
    else if (f == "id") e = arg(0, rt);
    else if (f == "not") e = "!" + arg(0);
    else if (f == "min" || f == "max") {
      // interpreter semantics (Const.hs:81-82): min a b = if a < b then a else b.  On NaN this keeps b, so an
      // acceptance ratio `min 1 NaN` stays NaN and the bernoulli rejects; the C++ template min<double>(1, NaN)
      // (CC.hs:146-147) would yield 1 and accept.  DESIGN.md 3 lists the divergence.
      std::string a0 = arg(0, rt), a1 = arg(1, rt);
      e = "(" + a0 + (f == "min" ? " < " : " > ") + a1 + " ? " + a0 + " : " + a1 + ")";
    }
    else if (f == "vectorSize") {                                                          // CC.hs:144-145
      const RefP& a = n.args.at(0);
      if (a->k == Ref::Var && lookup(a->name).array) {
        const Node& an = *lookup(a->name).array;
        e = "(" + rv(an.shape[0].second) + " - " + rv(an.shape[0].first) + " + 1)";
      } else if (a->k == Ref::Var && lookup(a->name).lazy) e = lookup(a->name).lazy->len;
      else e = arg(0) + ".len";
    }
    else if (f == "matrixCols") e = mat_of(n.args.at(0))->cols;
    else if (f == "matrixRows") e = mat_of(n.args.at(0))->rows;
    else if (f == "log_det") { log_det_node(d); return; }
    else if (f == "det") {
      std::string R, C, mb = mat_store(mat_of(n.args.at(0)), d.id, &R, &C);
      e = "stk::det_lu(" + mb + ", " + R + ")";
    }
    else if (f == "<.>") {                                                                // dot product
      VA x = va(n.args.at(0)), y = va(n.args.at(1));
      std::string k = fresh("k");
      line("double " + d.id + " = 0.0;");
      line("for (int " + k + " = 1; " + k + " <= " + x.len + "; ++" + k + ") " + d.id + " += " + x.at(k) + " * " + y.at(k) + ";");
      v.expr = d.id; bind(d.id, v); return;
    }
    else if (f == "multi_normal_lpdf") {
      // ccNode's template (CC.hs:221-225): -0.5 * ((L^-1 (z - m)).squaredNorm() + log(2 pi) n + LOG_DET(C)), L = chol C
      VA z = va(n.args.at(0)), mu = va(n.args.at(1));
      std::string R, C, cb = mat_store(mat_of(n.args.at(2)), d.id, &R, &C), rb = fresh(d.id + "_r"), i = fresh("i"), ss = fresh("ss"), ld = fresh("ld");
      decls << "  double " << rb << "[STK_MATCAP];\n";
      line("stk::chol_lower(" + cb + ", " + R + ");");
      line("for (int " + i + " = 1; " + i + " <= " + R + "; ++" + i + ") " + rb + "[" + i + " - 1] = " + z.at(i) + " - " + mu.at(i) + ";");
      line("stk::tri_solve_lower(" + cb + ", " + R + ", " + rb + ", 1);");
      line("double " + ss + " = 0.0, " + ld + " = 0.0;");
      line("for (int " + i + " = 0; " + i + " < " + R + "; ++" + i + ") { " + ss + " += " + rb + "[" + i + "] * " + rb + "[" + i + "]; " + ld + " += log(" + cb + "[" + i + " * " + R + " + " + i + "]); }");
      e = "-0.5 * (" + ss + " + stk::kLog2Pi * (double)" + R + " + 2.0 * " + ld + ")";
    }
    else if (f == "poissonProcess_lpdf") { poisson_process_node(d); return; }
    else if (f == "findSortedInsertIndex") e = "stk::find_sorted_insert_index_v(" + arg(0, true) + ", " + vec_value(n.args.at(1)) + ")";   // CC.hs:192-197
    else if (f == "logFactorial") e = "stk::log_factorial(" + arg(0) + ")";               // CC.hs:167-168
    else if (f == "pmf_lpdf" || f == "categorical_lpdf") {                                 // Const.hs:94
      const RefP& ps = n.args.at(1);
      if (ps->k != Ref::List && ps->k != Ref::BlockVec) die(f + " needs a literal probability list");
      for (auto& pe : ps->args) if (!type_of(pe).scalar()) die(f + ": probability table with a non-scalar block");
      // log ps[k] (Const.hs:94); k is a literal in the IRs the rjmc builder emits, so the table folds into a select
      VA tab = va(ps);
      e = "log(" + tab.at(arg(0)) + ")";
    }
    else if (f == "normals_lpdf" || f == "uniforms_lpdf" || f == "cauchys_lpdf" || f == "bernoulliLogits_lpdf" || f == "poissons_lpdf" || f == "bernoullis_lpdf") {
      // vectorised log-densities: the loop of CC.hs:214-220 (`normals_lpdf`; 1-D only, as there), summed in index order
      std::string sf = f.substr(0, f.size() - 6) + "_lpdf", k = fresh("k"), len;
      const RefP& x0 = n.args.at(0);
      if (x0->k == Ref::Var && lookup(x0->name).array) { const Node& an = *lookup(x0->name).array; len = "(" + rv(an.shape[0].second) + " - " + rv(an.shape[0].first) + " + 1)"; }
      else len = rv(x0) + ".len";
      line("double " + d.id + " = 0.0;");
      line("for (int " + k + " = 1; " + k + " <= " + len + "; ++" + k + ") {");
      ind += 2; push();
      Val kv; kv.type = scalarT(Type::Z); kv.expr = k; bind(k, kv);
      std::vector<RefP> el;
      for (auto& a : n.args) {
        auto ix = std::make_shared<Ref>(); ix->k = Ref::Index; ix->base = a;
        auto kr = std::make_shared<Ref>(); kr->k = Ref::Var; kr->name = k; ix->args.push_back(kr);
        el.push_back(ix);
      }
      line(d.id + " += " + lpdf_call(sf, el) + ";");
      pop(); ind -= 2; line("}");
      v.expr = d.id; bind(d.id, v); return;
    }
    else if (f.size() > 5 && f.compare(f.size() - 5, 5, "_lpdf") == 0) e = lpdf_call(f, n.args);   // CC.hs:198-228
    else if (f == "uniform_cdf") e = "stk::uniform_cdf(" + arg(0, true) + ", " + arg(1, true) + ", " + arg(2, true) + ")";
    else {
      static const std::set<std::string> math1 = {"exp", "log", "sqrt", "abs", "sin", "cos", "tan", "tanh", "sinh", "cosh", "floor", "ceil", "lgamma", "log1p", "expm1", "asin", "acos", "atan"};
      if (math1.count(f)) {
        e = (f == "abs" ? std::string(rt ? "fabs" : "abs") : f) + "(" + arg(0, rt) + ")";    // CC.hs:229-230 (generic call)
      } else die("node `" + f + "` is not supported by the CUDA lowering");
    }
    if (inline_mode) { v.expr = "(" + e + ")"; bind(d.id, v); return; }     // folded into the consuming expression
    line("const " + ctype(n.type) + " " + d.id + " = " + e + ";");
    v.expr = d.id; bind(d.id, v);
  }

  // a vector operand as a stored view (VecD / VecI): state fields, buffers, materialised comprehensions
  std::string vec_value(const RefP& r) {
    if (r->k == Ref::Var && lookup(r->name).array) return materialise(r->name);
    return rv(r);
  }

  // vectors that must exist as storage: replaceIndex / insertIndex / deleteIndex (CC.hs:169-191)
  void vector_node(const Def& d) {
    if (inline_mode) throw NotInlinable();
    const Node& n = d.node; const std::string& f = n.f;
    bool real = n.type.sub[0].k == Type::R;
    if (f == "ifThenElse") {                        // a select between two views: no copy
      std::string a1 = vec_value(n.args.at(1)), a2 = vec_value(n.args.at(2));
      line("const " + ctype(n.type) + " " + d.id + " = (" + rv(n.args.at(0)) + ") ? " + a1 + " : " + a2 + ";");
      Val v; v.type = n.type; v.expr = d.id; bind(d.id, v);
      return;
    }
    std::string ct = real ? "double" : "int", buf = fresh(d.id + "_buf");
    decls << "  " << ct << " " << buf << "[STK_VECCAP];\n";
    if (f == "#>" || (f == "<\\>" && type_of(n.args.at(0)).k == Type::Mat)) {
      // matrix #> vector (CC.hs:157-158) and matrix <\> vector (CC.hs:159-160: the solution of M x = v)
      std::shared_ptr<Mat> m = mat_of(n.args.at(0));
      VA x = va(n.args.at(1));
      std::string ln = fresh("n"), i = fresh("i"), k = fresh("k"), acc = fresh("acc");
      if (f == "#>") {
        m->prepare();
        line("const int " + ln + " = stk::cap_len(" + m->rows + ", STK_VECCAP, err_);");
        line("for (int " + i + " = 1; " + i + " <= " + ln + "; ++" + i + ") { double " + acc + " = 0.0; for (int " + k + " = 1; " + k + " <= " + m->cols + "; ++" + k + ") " +
             acc + " += " + m->at(i, k) + " * " + x.at(k) + "; " + buf + "[" + i + " - 1] = " + acc + "; }");
      } else {
        std::string R, C, mb = mat_store(m, d.id, &R, &C);
        line("const int " + ln + " = stk::cap_len(" + R + ", STK_VECCAP, err_);");
        line("for (int " + i + " = 1; " + i + " <= " + ln + "; ++" + i + ") " + buf + "[" + i + " - 1] = " + x.at(i) + ";");
        line("stk::lu_solve(" + mb + ", " + R + ", " + buf + ", 1);");
      }
      line("const " + ctype(n.type) + " " + d.id + "(" + buf + ", 1, " + ln + ");");
      Val v; v.type = n.type; v.expr = d.id; bind(d.id, v);
      return;
    }
    std::string src = vec_value(n.args.at(0)), i = fresh("i"), ln = fresh("n");
    // every copy into per-thread storage is bounded by STK_VECCAP; an overflow sets the module's err flag
    if (f == "replaceIndex") {
      std::string j = fresh("j");
      line("const int " + ln + " = stk::cap_len(" + src + ".len, STK_VECCAP, err_);");
      line("for (int " + i + " = 1; " + i + " <= " + ln + "; ++" + i + ") " + buf + "[" + i + " - 1] = " + src + ".at(" + i + ");");
      line("const int " + j + " = " + rv(n.args.at(1)) + ";");
      line("if (" + j + " >= 1 && " + j + " <= " + ln + ") " + buf + "[" + j + " - 1] = " + rv(n.args.at(2), real) + ";");
      line("const " + ctype(n.type) + " " + d.id + "(" + buf + ", 1, " + ln + ");");
    } else if (f == "insertIndex") {
      std::string j = rv(n.args.at(1)), e = rv(n.args.at(2), real);
      line("const int " + ln + " = stk::cap_len(" + src + ".len + 1, STK_VECCAP, err_);");
      line("for (int " + i + " = 1; " + i + " <= " + ln + "; ++" + i + ") " + buf + "[" + i + " - 1] = " + i + " < " + j + " ? " + src + ".at(" + i + ") : (" + i + " == " + j + " ? " + e + " : " + src + ".at(" + i + " - 1));");
      line("const " + ctype(n.type) + " " + d.id + "(" + buf + ", 1, " + ln + ");");
    } else if (f == "deleteIndex") {
      std::string j = rv(n.args.at(1));
      line("const int " + ln + " = stk::cap_len(" + src + ".len - 1, STK_VECCAP, err_);");
      line("for (int " + i + " = 1; " + i + " <= " + ln + "; ++" + i + ") " + buf + "[" + i + " - 1] = " + i + " < " + j + " ? " + src + ".at(" + i + ") : " + src + ".at(" + i + " + 1);");
      line("const " + ctype(n.type) + " " + d.id + "(" + buf + ", 1, " + ln + ");");
    } else if (f == "id") {
      line("const " + ctype(n.type) + " " + d.id + " = " + src + ";");
    } else die("vector node `" + f + "` is not supported by the CUDA lowering");
    Val v; v.type = n.type; v.expr = d.id; bind(d.id, v);
  }

  static bool mentions(const RefP& r, const std::string& id) {
    if (!r) return false;
    if (r->k == Ref::Var && r->name == id) return true;
    if (mentions(r->base, id)) return true;
    for (auto& a : r->args) if (mentions(a, id)) return true;
    if (r->let) return mentions_body(*r->let, id);
    if (r->pc) return mentions_body(*r->pc->key, id) || mentions_body(*r->pc->val, id) || mentions(r->pc->lo, id) || mentions(r->pc->hi, id);
    return false;
  }
  static bool mentions_body(const Body& b, const std::string& id) {
    for (const Def& d : b.defs) {
      for (auto& a : d.node.args) if (mentions(a, id)) return true;
      for (auto& sh : d.node.shape) if (mentions(sh.first, id) || mentions(sh.second, id)) return true;
      if (d.node.lam && mentions_body(*d.node.lam, id)) return true;
      for (auto& al : d.node.alts) if (mentions_body(al.body, id)) return true;
    }
    for (auto& r : b.rets) if (mentions(r, id)) return true;
    return false;
  }
  // an ADDITIVE fold: the body returns `acc + term` (or `term + acc`) and reads the accumulator nowhere else — a sum, whose
  // terms can be evaluated in any order.  Returns the index of the final definition and of its non-accumulator operand.
  static bool additive_fold(const Node& n, const std::string& acc, int* def_ix, int* term_arg) {
    const Body& b = *n.lam;
    if (b.defs.empty() || b.rets.size() != 1 || b.rets[0]->k != Ref::Var) return false;
    const Def& last = b.defs.back();
    if (last.id != b.rets[0]->name || last.node.k != Node::Apply || last.node.f != "+" || last.node.args.size() != 2) return false;
    int ai = -1;
    for (int k = 0; k < 2; ++k) if (last.node.args[k]->k == Ref::Var && last.node.args[k]->name == acc) ai = k;
    if (ai < 0 || mentions(last.node.args[1 - ai], acc)) return false;
    for (size_t k = 0; k + 1 < b.defs.size(); ++k) {
      Body one; one.defs.push_back(b.defs[k]);
      if (mentions_body(one, acc)) return false;
    }
    *def_ix = (int)b.defs.size() - 1; *term_arg = 1 - ai;
    return true;
  }

  // FoldScan (CC.hs:236-262): sequential loop, left/right; the list is a vector view or an inlined Array.
  // With `lanes` > 1 lanes per chain an additive fold over reals (a sum of log-density terms over the observations — the
  // `foldl (+)` that `lpdf` of a joint builds) is split over the lanes: lane s takes the elements s, s + lanes, ..., the partial
  // sums are combined by an xor-butterfly of shuffles over the chain's lane group (every lane ends with the same value, fixed order:
  // deterministic) and the seed is added once.  Everything else runs redundantly on the lanes of a chain.
  // Scans keep every partial result (interpreter semantics, Eval.hs:213-221 = Haskell's scanl / scanr: n + 1 elements,
  // the seed first for scanl, last for scanr) in per-thread storage.  The CC template indexes `name[idx-1-1]` for a left
  // scan (CC.hs:256-259 builds "loc-1" from the string "idx-1"), i.e. reads one element before the seed: the interpreter
  // is the reference here.
  void fold(const Def& d) {
    if (inline_mode) throw NotInlinable();
    const Node& n = d.node;
    prematerialise(*n.lam);
    if (n.args.at(1)->k == Ref::Var) if (Val* lv = find_val(n.args[1]->name)) if (lv->array) prematerialise(*lv->array->lam);
    const bool scan = n.f == "scanl" || n.f == "scanr";
    const bool left = n.f == "foldl" || n.f == "scanl";
    if (n.f != "foldl" && n.f != "foldr" && !scan) die("`" + n.f + "` is not supported by the CUDA lowering");
    Type at;
    if (scan) {
      if (n.type.k == Type::Vec && n.type.sub[0].scalar()) at = n.type.sub[0];
      else if (n.type.k == Type::Unknown) at = type_of(n.args[0]);
      else die("scan producing " + n.type.str() + " is not supported");
    } else {
      if (!n.type.scalar() && n.type.k != Type::Unknown) die("fold producing " + n.type.str() + " is not supported");
      at = n.type.k == Type::Unknown ? type_of(n.args[0]) : n.type;
    }
    bool rt = at.k == Type::R;
    const std::string cur = scan ? fresh(d.id + "_acc") : d.id;
    line(ctype(at) + " " + cur + " = " + rv(n.args[0], rt) + ";");
    const RefP& ls = n.args[1];
    std::string it = fresh("k");
    const Node* arr = (ls->k == Ref::Var && lookup(ls->name).array) ? lookup(ls->name).array : nullptr;
    // showNode prints the lambda inputs reversed for a left fold / scan: \acc elem
    std::string acc = left ? n.params[0] : n.params[1], elem = left ? n.params[1] : n.params[0];
    std::string lo, hi;
    if (arr) { if (arr->params.size() != 1) die("fold over a multi-dimensional array"); lo = rv(arr->shape[0].first); hi = rv(arr->shape[0].second); }
    else { lo = "1"; hi = rv(ls) + ".len"; }
    std::string buf, len;
    if (scan) {
      buf = fresh(d.id + "_buf"); len = fresh("n");
      decls << "  " << (rt ? "double" : "int") << " " << buf << "[STK_VECCAP];\n";
      line("const int " + len + " = stk::cap_len((" + hi + ") - (" + lo + ") + 1, STK_VECCAP - 1, err_);");
      hi = "(" + lo + ") + " + len + " - 1";
      line(buf + "[" + (left ? "0" : len) + "] = " + cur + ";");
    }
    int add_def = -1, add_arg = -1;
    const bool par = lanes > 1 && !scan && rt && additive_fold(n, acc, &add_def, &add_arg);
    std::string part;
    if (par) {
      part = fresh("part");
      line("double " + part + " = 0.0;");
      line("for (int " + it + " = (" + lo + ") + sub_; " + it + " <= " + hi + "; " + it + " += STK_LANES) {");
    }
    else if (left) line("for (int " + it + " = " + lo + "; " + it + " <= " + hi + "; ++" + it + ") {");
    else line("for (int " + it + " = " + hi + "; " + it + " >= " + lo + "; --" + it + ") {");
    ind += 2; push();
    Val ev;
    if (arr) {
      Val iv; iv.type = scalarT(Type::Z); iv.expr = it; bind(arr->params[0], iv);
      ev.type = lookup(ls->name).type.sub[0];
      if (ev.type.k != Type::Unknown && !arr->lam->defs.empty()) {
        // the comprehension's definitions live at the same DAG level as the fold body's (both are lambdas of this
        // block: v_<L+1>_k in either), so they get a C scope of their own
        line(ctype(ev.type) + " " + elem + ";");
        line("{"); ind += 2; push();
        for (const Def& ad : arr->lam->defs) def(ad);
        line(elem + " = " + rv(arr->lam->rets[0], ev.type.k == Type::R) + ";");
        pop(); ind -= 2; line("}");
      } else {
        for (const Def& ad : arr->lam->defs) def(ad);
        if (ev.type.k == Type::Unknown) ev.type = type_of(arr->lam->rets[0]);
        line("const " + ctype(ev.type) + " " + elem + " = " + rv(arr->lam->rets[0], ev.type.k == Type::R) + ";");
      }
    } else {
      ev.type = type_of(ls).sub[0];
      line("const " + ctype(ev.type) + " " + elem + " = " + rv(ls) + ".at(" + it + ");");
    }
    ev.expr = elem; bind(elem, ev);
    if (par) {
      for (int k = 0; k < add_def; ++k) def(n.lam->defs[k]);
      line(part + " += " + rv(n.lam->defs[add_def].node.args[add_arg], true) + ";");
      pop(); ind -= 2; line("}");
      line(cur + " += stk::group_sum_masked<STK_LANES>(" + part + ", gmask_);");
    } else {
    Val av; av.type = at; av.expr = acc;
    line("const " + ctype(at) + " " + acc + " = " + cur + ";");
    bind(acc, av);
    for (const Def& bd : n.lam->defs) def(bd);
    line(cur + " = " + rv(n.lam->rets[0], rt) + ";");
    if (scan) line(buf + "[" + it + " - (" + lo + ")" + (left ? " + 1" : "") + "] = " + cur + ";");
    pop(); ind -= 2; line("}");
    }
    Val v;
    if (scan) {
      v.type = vecT(at);
      line("const " + ctype(v.type) + " " + d.id + "(" + buf + ", 1, " + len + " + 1);");
    } else v.type = at;
    v.expr = d.id; bind(d.id, v);
  }

  // Case on an integer head, alternatives 1..n (CC.hs:263-273)
  void case_node(const Def& d) {
    if (inline_mode) throw NotInlinable();
    const Node& n = d.node;
    std::vector<Type> ts;
    if (n.type.k == Type::Tuple) ts = n.type.sub; else if (n.type.k != Type::Unknown) ts.push_back(n.type);
    else {
      // no annotation: take the member types from the first alternative whose returns are all typed in this scope
      // (alternatives without definitions of their own); scalars default to R
      const Alt& a0 = n.alts[0];
      for (auto& r : a0.body.rets) {
        Type t = scalarT(Type::R);
        if (r->k == Ref::Num) t = type_of(r);
        else if (a0.body.defs.empty() && a0.inputs.empty()) { try { t = type_of(r); } catch (const Error&) {} }
        ts.push_back(t);
      }
    }
    Val v; v.type = n.type.k == Type::Unknown ? (ts.size() == 1 ? ts[0] : Type()) : n.type;
    if (v.type.k == Type::Unknown) { v.type.k = Type::Tuple; v.type.sub = ts; }
    for (size_t j = 0; j < ts.size(); ++j) {
      std::string nm = ts.size() == 1 ? d.id : d.id + "_" + std::to_string(j);
      if (ts[j].k == Type::Vec && ts[j].sub[0].scalar()) line(ctype(ts[j]) + " " + nm + ";");      // a view, re-pointed per alternative
      else if (!ts[j].scalar()) die("case alternatives returning " + ts[j].str() + " are not supported");
      else line(ctype(ts[j]) + " " + nm + " = 0;");
      v.tuple.push_back(nm);
    }
    v.expr = v.tuple[0];
    if (ts.size() == 1) v.tuple.clear();
    // Case over an algebraic data type (`caseD`, what pdfPNode emits for a Switch, Program.hs:1003-1012; the C++ template
    // refuses it, CC.hs:268): the head is a tagged union, alternative C<k> binds the constructor's fields from its slots
    const bool adt = n.alts[0].tag >= 0;
    Type ut; std::string hd = rv(n.args[0]);
    if (adt) {
      ut = type_of(n.args[0]);
      if (ut.k != Type::Union) die("case over constructors needs a union-typed head, got " + ut.str());
      union_struct(ut);
    }
    line("switch (" + hd + (adt ? ".tag" : "") + ") {");
    for (size_t a = 0; a < n.alts.size(); ++a) {
      if ((n.alts[a].tag >= 0) != adt) die("case mixing integer and constructor alternatives");
      line("case " + std::to_string(adt ? n.alts[a].tag : (int)a + 1) + ": {"); ind += 2; push();
      if (adt) {
        const UnionLayout& L = unions[ut.str()];
        const int c = n.alts[a].tag;
        if (c >= (int)L.ctor.size() || n.alts[a].inputs.size() != L.ctor[c].size()) die("case alternative C" + std::to_string(c) + " does not fit " + ut.str());
        for (size_t j = 0; j < n.alts[a].inputs.size(); ++j) {
          const Slot& sl = L.slots[L.ctor[c][j]];
          std::string cn = fresh(n.alts[a].inputs[j]);
          line("const " + ctype(sl.type) + " " + cn + " = " + hd + "." + sl.name + ";");
          Val iv; iv.type = sl.type; iv.expr = cn; bind(n.alts[a].inputs[j], iv);
        }
      }
      for (const Def& bd : n.alts[a].body.defs) def(bd);
      if (n.alts[a].body.rets.size() != ts.size()) die("case alternative arity mismatch");
      for (size_t j = 0; j < ts.size(); ++j)
        line((ts.size() == 1 ? d.id : d.id + "_" + std::to_string(j)) + " = " +
             (ts[j].k == Type::Vec ? vec_value(n.alts[a].body.rets[j]) : rv(n.alts[a].body.rets[j], ts[j].k == Type::R)) + ";");
      line("break;"); pop(); ind -= 2; line("}");
    }
    line("}");
    bind(d.id, v);
  }

  // ---- matrices (ccNode's Eigen one-liners, CC.hs:155-166; the interpreter's hmatrix primitives, Const.hs:56-66) ------------
  // One chain = one thread, so a matrix is not an Eigen object but an EXPRESSION: dimensions + an element accessor, composed
  // node by node (zeros / eye / asRow / scalar and element-wise arithmetic / row replace-insert-delete / select) and
  // evaluated where a consumer needs elements.  Products `<>` are the one node whose elements are re-read many times:
  // they are materialised into per-thread storage, lazily, at the first consumer (so a product that only one branch of a
  // guarded log_det uses is computed only when that branch runs).
  std::shared_ptr<Mat> mat_of(const RefP& r) {
    if (r->k == Ref::Var) {
      const Val& v = lookup(r->name);
      if (!v.mat) die(r->name + " is not a matrix");
      return v.mat;
    }
    if (r->k == Ref::List && !r->args.empty() && r->args[0]->k == Ref::List) {        // literal [[a,b],[c,d]]
      auto m = std::make_shared<Mat>();
      size_t R = r->args.size(), C = r->args[0]->args.size();
      std::vector<std::string> es;
      for (auto& row : r->args) { if (row->k != Ref::List || row->args.size() != C) die("ragged matrix literal"); for (auto& e : row->args) es.push_back(rv(e, true)); }
      m->rows = std::to_string(R); m->cols = std::to_string(C);
      m->at = [es, C](const std::string& i, const std::string& j) {
        if (es.size() == 1) return es[0];
        std::string k = "((" + i + ") - 1) * " + std::to_string(C) + " + (" + j + ")", e = es.back();
        for (size_t t = es.size() - 1; t-- > 0;) e = "(" + k + " == " + std::to_string(t + 1) + " ? " + es[t] + " : " + e + ")";
        return e;
      };
      m->prepare = [] {};
      return m;
    }
    die("a matrix was expected");
  }

  // write a matrix expression into per-thread storage (row-major, pitch = its column count); returns the buffer and the names of
  // the int variables holding its dimensions
  std::string mat_store(const std::shared_ptr<Mat>& m, const std::string& hint, std::string* R, std::string* C) {
    if (inline_mode) throw NotInlinable();
    m->prepare();
    std::string buf = fresh(hint + "_mat"), i = fresh("i"), j = fresh("j");
    *R = fresh("R"); *C = fresh("C");
    decls << "  double " << buf << "[STK_MATCAP * STK_MATCAP];\n";
    line("const int " + *R + " = stk::cap_len(" + m->rows + ", STK_MATCAP, err_, 3), " + *C + " = stk::cap_len(" + m->cols + ", STK_MATCAP, err_, 3);");
    if (m->zeros) line("for (int " + i + " = 0; " + i + " < " + *R + " * " + *C + "; ++" + i + ") " + buf + "[" + i + "] = 0.0;");
    else line("for (int " + i + " = 1; " + i + " <= " + *R + "; ++" + i + ") for (int " + j + " = 1; " + j + " <= " + *C + "; ++" + j + ") " + buf + "[(" + i + " - 1) * " + *C + " + " + j + " - 1] = " + m->at(i, j) + ";");
    return buf;
  }
  static std::shared_ptr<Mat> stored_mat(const std::string& buf, const std::string& R, const std::string& C) {
    auto m = std::make_shared<Mat>();
    m->rows = R; m->cols = C; m->prepare = [] {};
    m->at = [buf, C](const std::string& i, const std::string& j) { return buf + "[((" + i + ") - 1) * " + C + " + (" + j + ") - 1]"; };
    return m;
  }

  void mat_def(const Def& d) {
    const Node& n = d.node; const std::string& f = n.f;
    if (f == "chol" || f == "inv" || (f == "<\\>" && n.args.size() == 2)) {
      // factorisations: the operand is written into per-thread storage at the definition and factorised / solved in place
      // (`chol` CC.hs:161-162, `inv` :163-164, `<\>` :159-160)
      std::string R, C, buf = mat_store(mat_of(n.args.at(0)), d.id, &R, &C);
      if (f == "chol") { line("stk::chol_lower(" + buf + ", " + R + ");"); Val v; v.type = n.type; v.mat = stored_mat(buf, R, R); bind(d.id, v); return; }
      std::string R2, C2, rhs;
      if (f == "inv") {
        auto eye = std::make_shared<Mat>(); eye->rows = eye->cols = R; eye->eye = true; eye->prepare = [] {};
        eye->at = [](const std::string& i, const std::string& j) { return "((" + i + ") == (" + j + ") ? 1.0 : 0.0)"; };
        rhs = mat_store(eye, d.id + "_rhs", &R2, &C2);
      } else rhs = mat_store(mat_of(n.args.at(1)), d.id + "_rhs", &R2, &C2);
      line("stk::lu_solve(" + buf + ", " + R + ", " + rhs + ", " + C2 + ");");
      Val v; v.type = n.type; v.mat = stored_mat(rhs, R2, C2); bind(d.id, v); return;
    }
    if (f == "diag") {
      VA dv = va(n.args.at(0));
      auto m = std::make_shared<Mat>(); m->rows = m->cols = dv.len; m->prepare = [] {};
      m->at = [dv](const std::string& i, const std::string& j) { return "((" + i + ") == (" + j + ") ? " + dv.at(i) + " : 0.0)"; };
      Val v; v.type = n.type; v.mat = m; bind(d.id, v); return;
    }
    if (f == "quad_form_diag" || f == "qfDiag") {          // diag(v) M diag(v)  (covPrior, app/benchmark/Main.hs:112)
      auto a = mat_of(n.args.at(0)); VA dv = va(n.args.at(1));
      auto m = std::make_shared<Mat>(); m->rows = a->rows; m->cols = a->cols; m->prepare = [a] { a->prepare(); };
      m->at = [a, dv](const std::string& i, const std::string& j) { return "(" + dv.at(i) + " * " + a->at(i, j) + " * " + dv.at(j) + ")"; };
      Val v; v.type = n.type; v.mat = m; bind(d.id, v); return;
    }
    if (f == "outer") {
      VA a = va(n.args.at(0)), b = va(n.args.at(1));
      auto m = std::make_shared<Mat>(); m->rows = a.len; m->cols = b.len; m->prepare = [] {};
      m->at = [a, b](const std::string& i, const std::string& j) { return "(" + a.at(i) + " * " + b.at(j) + ")"; };
      Val v; v.type = n.type; v.mat = m; bind(d.id, v); return;
    }
    auto m = std::make_shared<Mat>();
    m->prepare = [] {};
    auto is_mat = [&](size_t i) { return type_of(n.args.at(i)).k == Type::Mat; };
    auto sc = [&](size_t i) { return "(" + rv(n.args.at(i), true) + ")"; };
    auto both_prepare = [](std::shared_ptr<Mat> a, std::shared_ptr<Mat> b) { return [a, b] { a->prepare(); if (b) b->prepare(); }; };
    if (f == "zeros") { m->rows = "(" + rv(n.args.at(0)) + ")"; m->cols = "(" + rv(n.args.at(1)) + ")"; m->zeros = true; m->at = [](const std::string&, const std::string&) { return std::string("0.0"); }; }
    else if (f == "eye") { m->rows = m->cols = "(" + rv(n.args.at(0)) + ")"; m->eye = true; m->at = [](const std::string& i, const std::string& j) { return "((" + i + ") == (" + j + ") ? 1.0 : 0.0)"; }; }
    else if (f == "asRow" || f == "asColumn") {
      VA v = va(n.args.at(0));
      const bool row = f == "asRow";
      m->rows = row ? "1" : v.len; m->cols = row ? v.len : "1";
      m->at = [v, row](const std::string& i, const std::string& j) { return v.at(row ? j : i); };
    }
    else if (f == "negate") { auto a = mat_of(n.args.at(0)); *m = *a; m->eye = false; m->at = [a](const std::string& i, const std::string& j) { return "(-" + a->at(i, j) + ")"; }; }
    else if (f == "id") { *m = *mat_of(n.args.at(0)); }
    else if (f == "tr'" || f == "tr") { auto a = mat_of(n.args.at(0)); *m = *a; m->rows = a->cols; m->cols = a->rows; m->at = [a](const std::string& i, const std::string& j) { return a->at(j, i); }; }
    else if ((f == "exp" || f == "log" || f == "sqrt" || f == "square") && n.args.size() == 1) {      // element-wise unary function
      auto a = mat_of(n.args.at(0)); const std::string fn = f;
      *m = *a; m->eye = false; m->zeros = false;
      m->at = [a, fn](const std::string& i, const std::string& j) { return fn == "square" ? "((" + a->at(i, j) + ") * (" + a->at(i, j) + "))" : fn + "(" + a->at(i, j) + ")"; };
    }
    else if ((f == "*" || f == "/" || f == "*>") && n.args.size() == 2 && is_mat(0) != is_mat(1)) {       // matrix (*|/) scalar, scalar * matrix
      const size_t mi = is_mat(0) ? 0 : 1;
      auto a = mat_of(n.args.at(mi)); std::string s = sc(1 - mi);
      *m = *a; m->eye = false;
      if (f == "/" && mi != 0) die("scalar / matrix is not supported by the CUDA lowering");
      const std::string op = f == "/" ? " / " : " * ";
      if (!a->zeros) m->at = [a, s, op, mi](const std::string& i, const std::string& j) { return mi == 0 ? "(" + a->at(i, j) + op + s + ")" : "(" + s + op + a->at(i, j) + ")"; };
    }
    else if (f == "+" || f == "-" || f == "+s" || f == "*") {                                 // element-wise
      std::vector<std::shared_ptr<Mat>> ms;
      for (auto& a : n.args) {
        if (type_of(a).k == Type::Mat) { ms.push_back(mat_of(a)); continue; }
        // a scalar operand (`cast lsv - M`): a 1 x 1 matrix, replicated by the broadcast rule below
        auto sm = std::make_shared<Mat>(); sm->rows = sm->cols = "1"; sm->prepare = [] {};
        const std::string e = "(" + rv(a, true) + ")";
        sm->at = [e](const std::string&, const std::string&) { return e; };
        ms.push_back(sm);
      }
      // a column (cols == 1) or a row against a full matrix is replicated (rep_matrix in the reference's Stan output)
      m->rows = ms[0]->rows; m->cols = ms[0]->cols;
      for (auto& a : ms) { if (m->rows == "1" && a->rows != "1") m->rows = a->rows; if (m->cols == "1" && a->cols != "1") m->cols = a->cols; }
      {
        const std::string RR = m->rows, CC = m->cols;
        for (auto& a : ms) {
          const bool br = a->rows == "1" && RR != "1", bc = a->cols == "1" && CC != "1";
          if (!br && !bc) continue;
          auto src = a; auto b = std::make_shared<Mat>(); *b = *src; b->rows = RR; b->cols = CC; b->eye = false;
          b->at = [src, br, bc](const std::string& i, const std::string& j) { return src->at(br ? "1" : i, bc ? "1" : j); };
          a = b;
        }
      }
      const std::string op = f == "-" ? " - " : f == "*" ? " * " : " + ";
      if (f != "*") {                                  // a + 0 = a: static zeros drop out of sums (not out of a - b's first operand)
        bool all_zero = true; for (auto& a : ms) all_zero = all_zero && a->zeros;
        m->zeros = all_zero;
      }
      m->prepare = [ms] { for (auto& a : ms) a->prepare(); };
      m->at = [ms, op, f](const std::string& i, const std::string& j) {
        std::string e;
        for (size_t k = 0; k < ms.size(); ++k) {
          if (f != "*" && ms[k]->zeros && (k > 0 || ms.size() > 1) && !(f == "-" && k == 0 && ms.size() == 2 && !ms[1]->zeros && false)) {
            if (!(f == "-" && k == 0)) continue;       // keep `0 - b` explicit
          }
          std::string t = ms[k]->at(i, j);
          e = e.empty() ? t : "(" + e + op + t + ")";
        }
        return e.empty() ? std::string("0.0") : e;
      };
    }
    else if (f == "<>") {                               // matrix product: materialised at its first consumer
      auto a = mat_of(n.args.at(0)), b = mat_of(n.args.at(1));
      m->rows = a->rows; m->cols = b->cols;
      m->zeros = a->zeros || b->zeros;
      if (m->zeros) m->at = [](const std::string&, const std::string&) { return std::string("0.0"); };
      else {
        const bool one_row = a->rows == "1";
        std::string buf = fresh(d.id + "_prod"), pitch = fresh("pc");
        decls << "  double " << buf << (one_row ? "[STK_MATCAP];\n" : "[STK_MATCAP * STK_MATCAP];\n");
        decls << "  int " << pitch << " = 0;\n";
        auto state = std::make_shared<int>(-1);       // id of the scope the product was last materialised in
        m->prepare = [this, a, b, buf, pitch, state, one_row] {
          if (*state >= 0 && scope_live(*state)) return;
          if (inline_mode) throw NotInlinable();
          a->prepare(); b->prepare();
          std::string i = fresh("i"), j = fresh("j"), k = fresh("k"), acc = fresh("acc"), R = fresh("R"), C = fresh("C"), K = fresh("K");
          line("{ const int " + R + " = stk::cap_len(" + a->rows + ", " + (one_row ? "1" : "STK_MATCAP") + ", err_, 3), " + C + " = stk::cap_len(" + b->cols +
               ", STK_MATCAP, err_, 3), " + K + " = " + a->cols + "; " + pitch + " = " + C + ";");
          line("  for (int " + i + " = 1; " + i + " <= " + R + "; ++" + i + ") for (int " + j + " = 1; " + j + " <= " + C + "; ++" + j + ") {");
          line("    double " + acc + " = 0.0;");
          line("    for (int " + k + " = 1; " + k + " <= " + K + "; ++" + k + ") " + acc + " += " + a->at(i, k) + " * " + b->at(k, j) + ";");
          line("    " + buf + "[(" + i + " - 1) * " + C + " + " + j + " - 1] = " + acc + ";");
          line("  }");
          line("}");
          *state = scope_ids.back();
        };
        m->at = [buf, pitch](const std::string& i, const std::string& j) { return buf + "[((" + i + ") - 1) * " + pitch + " + (" + j + ") - 1]"; };
      }
    }
    else if (f == "replaceIndex" || f == "insertIndex" || f == "deleteIndex") {              // row operations (CC.hs:169-191 on a matrix)
      auto a = mat_of(n.args.at(0));
      std::string at = "(" + rv(n.args.at(1)) + ")";
      VA row; if (f != "deleteIndex") row = va(n.args.at(2));
      m->cols = a->cols;
      m->rows = f == "replaceIndex" ? a->rows : f == "insertIndex" ? "(" + a->rows + " + 1)" : "(" + a->rows + " - 1)";
      m->prepare = [a] { a->prepare(); };
      if (f == "replaceIndex") m->at = [a, at, row](const std::string& i, const std::string& j) { return "((" + i + ") == " + at + " ? " + row.at(j) + " : " + a->at(i, j) + ")"; };
      else if (f == "insertIndex") m->at = [a, at, row](const std::string& i, const std::string& j) {
        return "((" + i + ") < " + at + " ? " + a->at(i, j) + " : ((" + i + ") == " + at + " ? " + row.at(j) + " : " + a->at("(" + i + ") - 1", j) + "))"; };
      else m->at = [a, at](const std::string& i, const std::string& j) { return "((" + i + ") < " + at + " ? " + a->at(i, j) + " : " + a->at("(" + i + ") + 1", j) + ")"; };
    }
    else if (f == "ifThenElse") {
      std::string c = "(" + rv(n.args.at(0)) + ")";
      auto a = mat_of(n.args.at(1)), b = mat_of(n.args.at(2));
      m->rows = a->rows; m->cols = a->cols;
      m->prepare = both_prepare(a, b);
      m->at = [c, a, b](const std::string& i, const std::string& j) { return "(" + c + " ? " + a->at(i, j) + " : " + b->at(i, j) + ")"; };
    }
    else die("matrix node `" + f + "` is not supported by the CUDA lowering");
    Val v; v.type = n.type; v.mat = m; bind(d.id, v);
  }

  // log_det (CC.hs:165-166 / Const.hs:65; log |det| by LU as the interpreter's invlndet).  The argument may be a guarded
  // choice of block matrices (the Jacobian of a reversible-jump move, rjmcTransRatio', Program.hs:1306-1347): one branch per
  // guard.  A (nested) block matrix is flattened to one grid of blocks, then reduced SYMBOLICALLY before anything is
  // evaluated: a block row (column) whose only non-zero block is an identity contributes a factor +-1 to the determinant and
  // is removed together with that block's column (row) — the same kind of rule the reference's own simplifier applies
  // (log_det (eye) = 0, Expression.hs:1180-1181).  What is left is written into per-thread storage and LU-factorised.
  struct Grid { int R = 0, C = 0; std::vector<std::shared_ptr<Mat>> b; std::shared_ptr<Mat>& at(int r, int c) { return b[(size_t)r * C + c]; } };
  Grid flatten(const RefP& r) {
    Grid g;
    if (r->k != Ref::BlockMat) { g.R = g.C = 1; g.b.push_back(mat_of(r)); return g; }
    const int R = r->field, C = r->tag;
    std::vector<Grid> sub;
    for (auto& e : r->args) sub.push_back(flatten(e));
    std::vector<int> rh(R, 0), cw(C, 0);
    for (int i = 0; i < R; ++i) for (int j = 0; j < C; ++j) {
      const Grid& s = sub[(size_t)i * C + j];
      if ((rh[i] && rh[i] != s.R) || (cw[j] && cw[j] != s.C)) die("nested block matrix whose inner partitions do not line up");
      rh[i] = s.R; cw[j] = s.C;
    }
    for (int i = 0; i < R; ++i) g.R += rh[i];
    for (int j = 0; j < C; ++j) g.C += cw[j];
    g.b.resize((size_t)g.R * g.C);
    int ro = 0;
    for (int i = 0; i < R; ++i) {
      int co = 0;
      for (int j = 0; j < C; ++j) {
        Grid& s = sub[(size_t)i * C + j];
        for (int a = 0; a < s.R; ++a) for (int b = 0; b < s.C; ++b) g.at(ro + a, co + b) = s.at(a, b);
        co += cw[j];
      }
      ro += rh[i];
    }
    return g;
  }
  static void reduce(Grid& g) {
    for (bool changed = true; changed;) {
      changed = false;
      for (int pass = 0; pass < 2 && !changed; ++pass) {         // pass 0: rows, pass 1: columns
        const int N = pass == 0 ? g.R : g.C, M = pass == 0 ? g.C : g.R;
        for (int i = 0; i < N && !changed; ++i) {
          int eye_at = -1; bool ok = true;
          for (int j = 0; j < M && ok; ++j) {
            const Mat& m = pass == 0 ? *g.at(i, j) : *g.at(j, i);
            if (m.eye) { if (eye_at >= 0) ok = false; eye_at = j; }
            else if (!m.zeros) ok = false;
          }
          if (!ok || eye_at < 0) continue;
          const int dr = pass == 0 ? i : eye_at, dc = pass == 0 ? eye_at : i;
          Grid h; h.R = g.R - 1; h.C = g.C - 1;
          for (int r = 0; r < g.R; ++r) for (int c = 0; c < g.C; ++c) if (r != dr && c != dc) h.b.push_back(g.at(r, c));
          g = h; changed = true;
        }
      }
    }
  }
  void log_det_branch(const std::string& out_var, const RefP& r) {
    Grid g = flatten(r);
    reduce(g);
    if (g.R == 0 && g.C == 0) { line(out_var + " = 0.0;"); return; }            // nothing but identity blocks
    if (g.R == 0 || g.C == 0) die("log_det of a block matrix that is not square in blocks after reduction");
    need_matbuf = true;
    for (auto& m : g.b) m->prepare();
    std::vector<std::string> ro(g.R + 1), co(g.C + 1);
    std::string decl = "int ";
    for (int i = 0; i <= g.R; ++i) { ro[i] = fresh("ro"); decl += ro[i] + " = " + (i ? ro[i - 1] + " + " + g.at(i - 1, 0)->rows : std::string("0")) + ", "; }
    for (int j = 0; j <= g.C; ++j) { co[j] = fresh("co"); decl += co[j] + " = " + (j ? co[j - 1] + " + " + g.at(0, j - 1)->cols : std::string("0")) + (j < g.C ? ", " : ";"); }
    line(decl);
    const std::string& dim = ro[g.R];
    line("if (" + dim + " != " + co[g.C] + " || " + dim + " > STK_MATCAP) { *err_ = 3; } else {");
    ind += 2;
    std::string t = fresh("t");
    line("for (int " + t + " = 0; " + t + " < " + dim + " * " + dim + "; ++" + t + ") mat_[" + t + "] = 0.0;");
    for (int i = 0; i < g.R; ++i) for (int j = 0; j < g.C; ++j) {
      const Mat& m = *g.at(i, j);
      if (m.zeros) continue;
      std::string a = fresh("i"), b = fresh("j");
      if (m.eye) line("for (int " + a + " = 1; " + a + " <= " + m.rows + "; ++" + a + ") mat_[(" + ro[i] + " + " + a + " - 1) * " + dim + " + " + co[j] + " + " + a + " - 1] = 1.0;");
      else line("for (int " + a + " = 1; " + a + " <= " + m.rows + "; ++" + a + ") for (int " + b + " = 1; " + b + " <= " + m.cols + "; ++" + b + ") mat_[(" + ro[i] + " + " + a + " - 1) * " + dim + " + " + co[j] + " + " + b + " - 1] = " + m.at(a, b) + ";");
    }
    line(out_var + " = stk::log_abs_det(mat_, " + dim + ");");
    ind -= 2; line("}");
  }
  bool need_matbuf = false;
  void log_det_node(const Def& d) {
    if (inline_mode) throw NotInlinable();
    const RefP& a = d.node.args.at(0);
    line("double " + d.id + " = stk::oob_value<double>();");
    if (a->k == Ref::Cond) {
      for (size_t i = 0; i + 1 < a->args.size(); i += 2) {
        line(std::string(i ? "else " : "") + "if (" + rv(a->args[i]) + ") {");
        ind += 2; push(); log_det_branch(d.id, a->args[i + 1]); pop(); ind -= 2;
        line("}");
      }
    } else { line("{"); ind += 2; push(); log_det_branch(d.id, a); pop(); ind -= 2; line("}"); }
    Val v; v.type = scalarT(Type::R); v.expr = d.id; bind(d.id, v);
  }

  // poissonProcess_lpdf y t rate mean (evalNode, Eval.hs:185-190): sum_i log (rate y_i) - mean, the rate being a Function
  // node of the block, applied to each observed point in order
  void poisson_process_node(const Def& d) {
    if (inline_mode) throw NotInlinable();
    const Node& n = d.node;
    const RefP& fr = n.args.at(2);
    if (fr->k != Ref::Var || !lookup(fr->name).fn) die("poissonProcess_lpdf: the rate must be a Function node");
    const Node& fn = *lookup(fr->name).fn;
    if (fn.params.size() != 1) die("poissonProcess_lpdf: the rate takes one argument");
    prematerialise(*fn.lam);
    std::string y = vec_value(n.args.at(0)), k = fresh("k"), acc = fresh("acc");
    line("double " + acc + " = 0.0;");
    if (lanes > 1) line("for (int " + k + " = 1 + sub_; " + k + " <= " + y + ".len; " + k + " += STK_LANES) {");      // the events split over the chain's lanes
    else line("for (int " + k + " = 1; " + k + " <= " + y + ".len; ++" + k + ") {");
    ind += 2; push();
    std::string p = fresh(fn.params[0]);
    line("const double " + p + " = " + y + ".at(" + k + ");");
    Val pv; pv.type = scalarT(Type::R); pv.expr = p; bind(fn.params[0], pv);
    for (const Def& bd : fn.lam->defs) def(bd);
    line(acc + " += log(" + rv(fn.lam->rets.at(0), true) + ");");
    pop(); ind -= 2; line("}");
    if (lanes > 1) line(acc + " = stk::group_sum_masked<STK_LANES>(" + acc + ", gmask_);");
    line("const double " + d.id + " = " + acc + " - " + rv(n.args.at(3), true) + ";");
    Val v; v.type = scalarT(Type::R); v.expr = d.id; bind(d.id, v);
  }

  // ---- PNodes (ccPNode', CC.hs:276-321) ------------------------------------------------------------
  std::map<std::string, const PNode*> pending;       // not yet emitted draws of the current do-block chain
  std::vector<std::map<std::string, const PNode*>> pending_stack;

  static void id_level_ptr(const std::string& id, int* level, int* ptr) {
    // x_<ns>_<L>_<P>
    size_t b = id.rfind('_'); size_t a = id.rfind('_', b - 1);
    if (b == std::string::npos || a == std::string::npos) die("malformed PNode id " + id);
    *ptr = atoi(id.c_str() + b + 1); *level = atoi(id.c_str() + a + 1);
  }

  // definitions of the body being lowered that have not been reached yet: a PNode may read a definition that is printed after the
  // one that needs the PNode (the interpreter evaluates by need, Program.hs:1046-1091), so such a definition is lowered on demand —
  // only while no inner scope is open, where its statements would land inside a loop
  struct PendingDefs { std::map<std::string, const Def*> defs; size_t depth; };
  std::vector<PendingDefs> pending_defs;
  void run_def(const Def& d) {
    for (auto& a : d.node.args) need(a);
    for (auto& sh : d.node.shape) { need(sh.first); need(sh.second); }
    if (d.node.k != Node::Array && d.node.lam) need_body(*d.node.lam);
    for (auto& al : d.node.alts) need_body(al.body);
    if (d.node.k == Node::Fold && d.node.args[1]->k == Ref::Var) {
      const Val& lv = lookup(d.node.args[1]->name);
      if (lv.array) { need_body(*lv.array->lam); for (auto& sh : lv.array->shape) { need(sh.first); need(sh.second); } }
    }
    def(d);
  }
  void ensure_drawn(const std::string& id) {
    for (size_t i = pending_stack.size(); i-- > 0;) {
      auto it = pending_stack[i].find(id);
      if (it != pending_stack[i].end()) { const PNode* q = it->second; pending_stack[i].erase(it); pnode(*q); return; }
    }
    if (!pending_defs.empty() && pending_defs.back().depth == scopes.size() && !inline_mode) {
      auto it = pending_defs.back().defs.find(id);
      if (it != pending_defs.back().defs.end()) { const Def* d = it->second; pending_defs.back().defs.erase(it); run_def(*d); }
    }
  }
  void need(const RefP& r) {       // draw any PNode the reference depends on before it is used
    if (!r) return;
    if (r->k == Ref::Var) ensure_drawn(r->name);
    need(r->base);
    for (auto& a : r->args) need(a);
  }

  void need_body(const Body& b) {  // draws referenced from inside a lambda body must precede the loop
    for (const Def& d : b.defs) {
      for (auto& a : d.node.args) need(a);
      for (auto& sh : d.node.shape) { need(sh.first); need(sh.second); }
      if (d.node.lam) need_body(*d.node.lam);
      for (auto& al : d.node.alts) need_body(al.body);
    }
    for (auto& r : b.rets) need(r);
  }

  // the sampler call of a Dist PNode on the generator variable `g` (ccPNode, CC.hs:276-297)
  std::string draw_expr(const PNode& q, const std::string& g) {
    auto arg = [&](size_t i, bool real = true) { return rv(q.args.at(i), real); };
    const std::string& f = q.dist;
    std::string e;
    if (f == "normal") e = "stk::draw_normal(" + g + ", " + arg(0) + ", " + arg(1) + ")";                 // CC.hs:278
    else if (f == "uniform") e = "stk::draw_uniform(" + g + ", " + arg(0) + ", " + arg(1) + ")";          // :280
    else if (f == "bernoulli") e = "stk::draw_bernoulli(" + g + ", " + arg(0) + ")";                      // :290-291
    else if (f == "bernoulliLogit") e = "stk::draw_bernoulli(" + g + ", 1.0 / (1.0 + exp(-(" + arg(0) + "))))";
    else if (f == "gamma") e = "stk::draw_gamma(" + g + ", " + arg(0) + ", " + arg(1) + ")";              // :294-296
    else if (f == "poisson") e = "stk::draw_poisson(" + g + ", " + arg(0) + ")";                          // :279
    else if (f == "discreteUniform") e = "stk::draw_discrete_uniform(" + g + ", " + arg(0, false) + ", " + arg(1, false) + ")";   // :277
    else if (f == "cauchy") e = "stk::draw_cauchy(" + g + ", " + arg(0) + ", " + arg(1) + ")";
    // distributions only the interpreter samples in the reference (samplePNode, Program.hs:1093-1170)
    else if (f == "inv_gamma") e = "stk::draw_inv_gamma(" + g + ", " + arg(0) + ", " + arg(1) + ")";
    else if (f == "geometric") e = "stk::draw_geometric(" + g + ", " + arg(0) + ")";
    else if (f == "logistic") e = "stk::draw_logistic(" + g + ", " + arg(0) + ", " + arg(1) + ")";
    else if (f == "neg_binomial") e = "stk::draw_neg_binomial(" + g + ", " + arg(0) + ", " + arg(1) + ")";
    else if (f == "beta") e = "stk::draw_beta(" + g + ", " + arg(0) + ", " + arg(1) + ")";
    else if (f == "pmf" || f == "categorical") {                                                          // :292-293
      const RefP& ps = q.args.at(0);
      if (ps->k != Ref::List && ps->k != Ref::BlockVec) die("pmf needs a literal probability list");
      std::string tab = fresh("ps"), init;
      for (size_t i = 0; i < ps->args.size(); ++i) {
        if (!type_of(ps->args[i]).scalar()) die("pmf: probability table with a non-scalar block");
        init += (i ? ", " : "") + rv(ps->args[i], true);
      }
      line("const double " + tab + "[] = {" + init + "};");
      e = "stk::draw_pmf(" + g + ", " + tab + ", " + std::to_string(ps->args.size()) + ")";
    } else die("distribution `" + f + "` is not supported by the CUDA lowering");
    return e;
  }

  // joint (Loop) PNode (ccPNode (Loop ..), CC.hs:299-302: for-loops around the element's DAG and draw; samplePNode,
  // Program.hs:1171-1181): the elements are drawn in index order into per-thread storage; element k (0-based) of the
  // PNode with slot s uses Philox slot s + (k + 1) * 2^16, so every element has its own counter range and the stream
  // does not depend on the order of evaluation.
  // A 2-D joint over a scalar distribution fills a per-thread matrix, element (i, j) from the element index (i - 1) cols + j of the
  // PNode's counter range; a 1-D joint over a vector-valued multivariate normal (`joint (designMatrix k) [ normal m S | i <- 1...n ]`,
  // app/benchmark/Main.hs:119) fills row i with m + chol(S) #> w, w_k from element index (i - 1) k + k'.
  void loop_node(const PNode& q, const std::string& slot) {
    if (q.params.size() > 2) die("joint (Loop) PNodes of more than two dimensions are not supported by the CUDA lowering");
    const PNode& in = *q.inner;
    if (in.k != PNode::Dist) die("joint (Loop) over a non-primitive distribution is not supported");
    if (q.params.size() == 2 || !in.type.scalar()) { loop_node_matrix(q, slot); return; }
    need(q.shape[0].first); need(q.shape[0].second);
    need_body(*q.lam);
    for (auto& a : in.args) need(a);
    bool real = in.type.k == Type::R;
    std::string buf = fresh(q.id + "_buf"), lo = fresh("lo"), hi = fresh("hi");
    decls << "  " << (real ? "double" : "int") << " " << buf << "[STK_VECCAP];\n";
    line("const int " + lo + " = " + rv(q.shape[0].first) + ";");
    line("const int " + hi + " = " + lo + " + stk::cap_len(" + rv(q.shape[0].second) + " - " + lo + " + 1, STK_VECCAP, err_) - 1;");
    const std::string& iv = q.params[0];
    line("for (int " + iv + " = " + lo + "; " + iv + " <= " + hi + "; ++" + iv + ") {");
    ind += 2; push();
    Val v; v.type = scalarT(Type::Z); v.expr = iv; bind(iv, v);
    for (const Def& d : q.lam->defs) def(d);
    std::string g = fresh("g");
    line(gen_decl(g, slot, "(" + iv + " - " + lo + " + 1)"));
    elem_draw(in, buf + "[" + iv + " - " + lo + "]", g);
    pop(); ind -= 2; line("}");
    Type vt = vecT(in.type);
    line("const " + ctype(vt) + " " + q.id + "(" + buf + ", 1, " + hi + " < " + lo + " ? 0 : " + hi + " - " + lo + " + 1);");
    Val r; r.type = vt; r.expr = q.id; bind(q.id, r);
  }

  // `dst = draw` of one element of a joint; an element distribution on a SubrangeT is redrawn until it is in range
  // (samplePNode on a SubrangeT, Program.hs:1081-1090, reached through the Loop's element), at most 2^16 times
  void elem_draw(const PNode& in, const std::string& dst, const std::string& g) {
    std::string e = draw_expr(in, g);
    if (!in.sub_lo && !in.sub_hi) { line(dst + " = " + e + ";"); return; }
    need(in.sub_lo); need(in.sub_hi);
    const bool real = in.type.k == Type::R;
    std::string cond, tries = fresh("tries");
    if (in.sub_lo) cond = dst + " < " + rv(in.sub_lo, real);
    if (in.sub_hi) cond += (cond.empty() ? "" : " || ") + dst + " > " + rv(in.sub_hi, real);
    line("{ int " + tries + " = 0; do { " + dst + " = " + e + "; } while ((" + cond + ") && ++" + tries + " < 65536); }");
  }
  void loop_node_matrix(const PNode& q, const std::string& slot) {
    const PNode& in = *q.inner;
    const bool two = q.params.size() == 2;
    for (auto& sh : q.shape) { need(sh.first); need(sh.second); }
    need_body(*q.lam);
    for (auto& a : in.args) need(a);
    std::string buf = fresh(q.id + "_mat"), R = fresh("R"), C = fresh("C"), lo0 = fresh("lo"), lo1 = fresh("lo");
    decls << "  double " << buf << "[STK_MATCAP * STK_MATCAP];\n";
    line("const int " + lo0 + " = " + rv(q.shape[0].first) + ", " + R + " = stk::cap_len(" + rv(q.shape[0].second) + " - " + lo0 + " + 1, STK_MATCAP, err_, 3);");
    const std::string& iv = q.params[0];
    if (two) {
      if (!in.type.scalar()) die("2-D joint over a non-scalar distribution is not supported");
      const std::string& jv = q.params[1];
      line("const int " + lo1 + " = " + rv(q.shape[1].first) + ", " + C + " = stk::cap_len(" + rv(q.shape[1].second) + " - " + lo1 + " + 1, STK_MATCAP, err_, 3);");
      line("if ((long long)" + R + " * " + C + " > 65535) *err_ = 3;");        // the element index has 16 bits of the block counter
      line("for (int " + iv + " = " + lo0 + "; " + iv + " < " + lo0 + " + " + R + "; ++" + iv + ") for (int " + jv + " = " + lo1 + "; " + jv + " < " + lo1 + " + " + C + "; ++" + jv + ") {");
      ind += 2; push();
      Val vi; vi.type = scalarT(Type::Z); vi.expr = iv; bind(iv, vi);
      Val vj; vj.type = scalarT(Type::Z); vj.expr = jv; bind(jv, vj);
      for (const Def& d : q.lam->defs) def(d);
      std::string g = fresh("g");
      line(gen_decl(g, slot, "((" + iv + " - " + lo0 + ") * " + C + " + " + jv + " - " + lo1 + " + 1)"));
      elem_draw(in, buf + "[(" + iv + " - " + lo0 + ") * " + C + " + " + jv + " - " + lo1 + "]", g);
      pop(); ind -= 2; line("}");
    } else {
      const bool mvn = (in.dist == "multi_normal" || in.dist == "normal" || in.dist == "normals") && in.args.size() == 2 && type_of(in.args[1]).k == Type::Mat;
      if (!mvn || in.type.k != Type::Vec) die("joint (Loop) over the vector-valued distribution `" + in.dist + "` is not supported");
      // the covariance and mean must not depend on the row index: factorised once
      for (const Def& d : q.lam->defs) def(d);
      VA mean = va(in.args.at(0));
      std::string RS, CS, L = mat_store(mat_of(in.args.at(1)), q.id + "_S", &RS, &CS);
      std::string wb = fresh(q.id + "_w"), k = fresh("k"), t = fresh("t"), g = fresh("g"), acc = fresh("acc");
      decls << "  double " << wb << "[STK_MATCAP];\n";
      line("stk::chol_lower(" + L + ", " + RS + ");");
      line("const int " + C + " = " + RS + ";");
      line("if ((long long)" + R + " * " + C + " > 65535) *err_ = 3;");
      line("for (int " + iv + " = " + lo0 + "; " + iv + " < " + lo0 + " + " + R + "; ++" + iv + ") {");
      ind += 2;
      line("for (int " + k + " = 1; " + k + " <= " + C + "; ++" + k + ") { " + gen_decl(g, slot, "((" + iv + " - " + lo0 + ") * " + C + " + " + k + ")") + " " + wb + "[" + k + " - 1] = stk::draw_normal(" + g + ", 0.0, 1.0); }");
      line("for (int " + k + " = 1; " + k + " <= " + C + "; ++" + k + ") { double " + acc + " = " + mean.at(k) + "; for (int " + t + " = 1; " + t + " <= " + k + "; ++" + t + ") " +
           acc + " += " + L + "[(" + k + " - 1) * " + C + " + " + t + " - 1] * " + wb + "[" + t + " - 1]; " + buf + "[(" + iv + " - " + lo0 + ") * " + C + " + " + k + " - 1] = " + acc + "; }");
      ind -= 2; line("}");
    }
    Type mt; mt.k = Type::Mat; mt.sub.push_back(scalarT(Type::R));
    Val r; r.type = mt; r.mat = stored_mat(buf, R, C); bind(q.id, r);
  }

  // Chain PNode (ccPNode (Chain ..), CC.hs:314-320: `name = x; for i in lo..hi { j = name; <body>; name = ret; }`;
  // chainRange', Program.hs:304-316).  The body's PNodes keep their own slots; iteration k (1-based) of the chain PNode
  // with slot s draws from a generator whose third counter word is XOR-ed with (k << 12) ^ s, i.e. from the streams of
  // "transition step ^ (((k << 12) ^ s) << 32)": two chains of one program, and two iterations of one chain, never share
  // a Philox block.
  std::string rng_name = "rng";
  void chain_node(const PNode& q) {
    if (rng_name != "rng") die("nested chain PNodes are not supported by the CUDA lowering");
    for (auto& a : q.args) need(a);
    const Alt& a = q.alts.at(0);
    Type st = type_of(q.args[2]);
    if (!st.scalar()) die("chain carrying a state of type " + st.str() + " is not supported by the CUDA lowering");
    bool real = st.k == Type::R;
    need_body(a.body);
    std::string lo = fresh("lo"), hi = fresh("hi"), rc = fresh("rngc");
    line(ctype(st) + " " + q.id + " = " + rv(q.args[2], real) + ";");
    line("const int " + lo + " = " + rv(q.args[0]) + ", " + hi + " = " + rv(q.args[1]) + ";");
    const std::string &iv = a.inputs[0], &jv = a.inputs[1];
    line("for (int " + iv + " = " + lo + "; " + iv + " <= " + hi + "; ++" + iv + ") {");
    ind += 2; push();
    Val v; v.type = scalarT(Type::Z); v.expr = iv; bind(iv, v);
    line("const " + ctype(st) + " " + jv + " = " + q.id + ";");
    Val j; j.type = st; j.expr = jv; bind(jv, j);
    int level, ptr; id_level_ptr(q.id, &level, &ptr);
    if (ptr >= 4095) die("PNode pointer " + std::to_string(ptr) + " of a chain exceeds the 4094 the Philox layout reserves");
    line("stk::Rng " + rc + " = rng; " + rc + ".k1 ^= (unsigned)(" + iv + " - " + lo + " + 1); " + rc + ".c2 ^= " + std::to_string(ptr + 1) + "u;");
    rng_name = rc;
    body(a.body);
    if (a.body.rets.size() != 1) die("chain body must return one value");
    line(q.id + " = " + rv(a.body.rets[0], real) + ";");
    rng_name = "rng";
    pop(); ind -= 2; line("}");
    Val r; r.type = st; r.expr = q.id; bind(q.id, r);
  }

  // orderedSample n d (HODist, samplePNode Program.hs:1174-1176: `sort <$> sequence [sample d | _ <- [1..n]]`): n draws
  // of the scalar distribution d, element k from Philox slot s + (k + 1) 2^16 like a joint, insertion-sorted ascending
  // into per-thread storage.
  void ordered_node(const PNode& q, const std::string& slot) {
    const PNode& in = *q.inner;
    if (in.k != PNode::Dist || !in.type.scalar() || in.sub_lo || in.sub_hi) die("orderedSample over `" + in.dist + "` is not supported by the CUDA lowering");
    if (q.args.size() != 1) die("orderedSample takes one count");
    need(q.args[0]);
    for (auto& a : in.args) need(a);
    bool real = in.type.k == Type::R;
    std::string ct = real ? "double" : "int", buf = fresh(q.id + "_buf"), n = fresh("n"), i = fresh("i"), j = fresh("j"), v = fresh("v"), g = fresh("g");
    decls << "  " << ct << " " << buf << "[STK_VECCAP];\n";
    line("const int " + n + " = stk::cap_len(" + rv(q.args[0]) + ", STK_VECCAP, err_);");
    line("for (int " + i + " = 0; " + i + " < " + n + "; ++" + i + ") {");
    ind += 2;
    line(gen_decl(g, slot, "(" + i + " + 1)"));
    line("const " + ct + " " + v + " = " + draw_expr(in, g) + ";");
    line("int " + j + " = " + i + ";");
    line("while (" + j + " > 0 && " + buf + "[" + j + " - 1] > " + v + ") { " + buf + "[" + j + "] = " + buf + "[" + j + " - 1]; --" + j + "; }");
    line(buf + "[" + j + "] = " + v + ";");
    ind -= 2; line("}");
    Type vt = vecT(in.type);
    line("const " + ctype(vt) + " " + q.id + "(" + buf + ", 1, " + n + ");");
    Val r; r.type = vt; r.expr = q.id; bind(q.id, r);
  }

  // vectorised Dist PNodes `normals m s`, `uniforms a b`, `cauchys m s`, `bernoulliLogits l`, `bernoullis p`, `poissons l`
  // (samplePNode, Program.hs:1093-1170; normalChol / gpclas / gpClassifier use them): element k ~ the scalar distribution
  // at the k-th elements of the argument vectors (views, or comprehensions with a bare body such as `[ 0.0 | _ <- 1...n ]`),
  // drawn like a joint: Philox slot s + (k + 1) 2^16, per-thread storage.
  bool vectorised_node(const PNode& q, const std::string& slot) {
    static const std::map<std::string, std::string> scalar_of = {{"normals", "normal"}, {"uniforms", "uniform"}, {"cauchys", "cauchy"},
      {"bernoulliLogits", "bernoulliLogit"}, {"bernoullis", "bernoulli"}, {"poissons", "poisson"}};
    auto it = scalar_of.find(q.dist);
    if (it == scalar_of.end()) return false;
    if (q.type.k == Type::Mat && q.type.sub[0].k == Type::R) {       // `normals M S :: P RMat`: element (i, j) at the (i, j) elements of the argument matrices
      for (auto& a : q.args) need(a);
      std::vector<std::shared_ptr<Mat>> ms;
      for (auto& a : q.args) { ms.push_back(mat_of(a)); ms.back()->prepare(); }
      std::string buf = fresh(q.id + "_mat"), R = fresh("R"), C = fresh("C"), i = fresh("i"), j = fresh("j"), g = fresh("g");
      decls << "  double " << buf << "[STK_MATCAP * STK_MATCAP];\n";
      line("const int " + R + " = stk::cap_len(" + ms[0]->rows + ", STK_MATCAP, err_, 3), " + C + " = stk::cap_len(" + ms[0]->cols + ", STK_MATCAP, err_, 3);");
      line("if ((long long)" + R + " * " + C + " > 65535) *err_ = 3;");
      line("for (int " + i + " = 1; " + i + " <= " + R + "; ++" + i + ") for (int " + j + " = 1; " + j + " <= " + C + "; ++" + j + ") {");
      ind += 2; push();
      PNode el; el.k = PNode::Dist; el.dist = it->second; el.type = scalarT(Type::R); el.id = q.id;
      for (size_t a = 0; a < ms.size(); ++a) {                     // the element's arguments as fresh scalar bindings
        std::string an = fresh("arg");
        line("const double " + an + " = " + ms[a]->at(i, j) + ";");
        Val av; av.type = scalarT(Type::R); av.expr = an; bind(an, av);
        auto ar = std::make_shared<Ref>(); ar->k = Ref::Var; ar->name = an; el.args.push_back(ar);
      }
      line(gen_decl(g, slot, "((" + i + " - 1) * " + C + " + " + j + ")"));
      line(buf + "[(" + i + " - 1) * " + C + " + " + j + " - 1] = " + draw_expr(el, g) + ";");
      pop(); ind -= 2; line("}");
      Val r; r.type = q.type; r.mat = stored_mat(buf, R, C); bind(q.id, r);
      return true;
    }
    if (q.type.k != Type::Vec || !q.type.sub[0].scalar()) die("`" + q.dist + "` of type " + q.type.str() + " is not supported by the CUDA lowering (vectors only)");
    for (auto& a : q.args) need(a);
    bool real = q.type.sub[0].k == Type::R;
    std::string buf = fresh(q.id + "_buf"), n = fresh("n"), k = fresh("k"), g = fresh("g");
    decls << "  " << (real ? "double" : "int") << " " << buf << "[STK_VECCAP];\n";
    auto length = [&](const RefP& a) -> std::string {
      if (a->k == Ref::Var && lookup(a->name).array) { const Node& an = *lookup(a->name).array; return "(" + rv(an.shape[0].second) + " - " + rv(an.shape[0].first) + " + 1)"; }
      return "(" + va(a).len + ")";
    };
    line("const int " + n + " = stk::cap_len(" + length(q.args.at(0)) + ", STK_VECCAP, err_);");
    line("for (int " + k + " = 1; " + k + " <= " + n + "; ++" + k + ") {");
    ind += 2;
    // the scalar PNode of one element: arguments are the vectors indexed at k
    PNode el; el.k = PNode::Dist; el.dist = it->second; el.type = q.type.sub[0]; el.id = q.id;
    push(); Val kv; kv.type = scalarT(Type::Z); kv.expr = k; bind(k, kv);
    for (auto& a : q.args) {
      auto ix = std::make_shared<Ref>(); ix->k = Ref::Index; ix->base = a;
      auto kr = std::make_shared<Ref>(); kr->k = Ref::Var; kr->name = k; ix->args.push_back(kr);
      el.args.push_back(ix);
    }
    line(gen_decl(g, slot, k));
    line(buf + "[" + k + " - 1] = " + draw_expr(el, g) + ";");
    pop(); ind -= 2; line("}");
    line("const " + ctype(q.type) + " " + q.id + "(" + buf + ", 1, " + n + ");");
    Val r; r.type = q.type; r.expr = q.id; bind(q.id, r);
    return true;
  }

  // generator of one draw: the PNode's tag goes into counter word c2, the element index of a joint into the high half of
  // the block counter; the low 16 bits count the blocks the draw itself consumes (rejection loops included)
  std::string gen_decl(const std::string& g, const std::string& tag, const std::string& elem) {
    return "stk::Rng " + g + " = " + rng_name + "; " + g + ".c2 ^= " + tag + "; " + g + ".slot = " +
           (elem.empty() ? std::string("0u") : "(unsigned)" + elem + " << 16") + ";";
  }

  // draws whose arguments or values are matrices (only the reference's interpreter samples them: samplePNode,
  // Program.hs:1114-1157; definitions Data/Random/Distribution/Abstract.hs:91-98, 172-189), on per-thread storage:
  //   multi_normal m S   x = m + chol(S) #> w, w_k ~ normal 0 1 (element k of the PNode's counter range)
  //   wishart n V        Bartlett as the reference writes it: A lower triangular, A_ii = chiSquare (n - i + 1) (the draw itself,
  //                      not its square root — restated faithfully), A_ij ~ normal 0 1 (i > j), L = chol V, L A (L A)^T
  //   inv_wishart n W    inv (wishart n (inv W))
  //   lkj_corr v         2 x 2 only, as the reference: [[1, r], [r, 1]], r = 2 beta(v, v) - 1
  bool matrix_dist_node(const PNode& q, const std::string& slot) {
    const std::string& f = q.dist;
    if (f == "multi_normal" || ((f == "normal" || f == "normals") && q.type.k == Type::Vec && q.args.size() == 2 && type_of(q.args[1]).k == Type::Mat)) {
      VA m = va(q.args.at(0));
      std::string R, C, L = mat_store(mat_of(q.args.at(1)), q.id + "_S", &R, &C);
      std::string buf = fresh(q.id + "_buf"), wb = fresh(q.id + "_w"), n = fresh("n"), i = fresh("i"), k = fresh("k"), g = fresh("g"), acc = fresh("acc");
      decls << "  double " << buf << "[STK_VECCAP];\n  double " << wb << "[STK_VECCAP];\n";
      line("stk::chol_lower(" + L + ", " + R + ");");
      line("const int " + n + " = stk::cap_len(" + R + ", STK_VECCAP, err_);");
      line("for (int " + k + " = 1; " + k + " <= " + n + "; ++" + k + ") { " + gen_decl(g, slot, k) + " " + wb + "[" + k + " - 1] = stk::draw_normal(" + g + ", 0.0, 1.0); }");
      line("for (int " + i + " = 1; " + i + " <= " + n + "; ++" + i + ") { double " + acc + " = " + m.at(i) + "; for (int " + k + " = 1; " + k + " <= " + i + "; ++" + k + ") " +
           acc + " += " + L + "[(" + i + " - 1) * " + R + " + " + k + " - 1] * " + wb + "[" + k + " - 1]; " + buf + "[" + i + " - 1] = " + acc + "; }");
      Type vt = vecT(scalarT(Type::R));
      line("const VecD " + q.id + "(" + buf + ", 1, " + n + ");");
      Val v; v.type = vt; v.expr = q.id; bind(q.id, v);
      return true;
    }
    if (f == "wishart" || f == "inv_wishart") {
      const bool inv = f == "inv_wishart";
      std::string df = rv(q.args.at(0));
      std::string R, C, V = mat_store(mat_of(q.args.at(1)), q.id + "_V", &R, &C);
      std::string A = fresh(q.id + "_A"), B = fresh(q.id + "_B"), W = fresh(q.id + "_W"), i = fresh("i"), j = fresh("j"), k = fresh("k"), g = fresh("g"), acc = fresh("acc");
      decls << "  double " << A << "[STK_MATCAP * STK_MATCAP];\n  double " << B << "[STK_MATCAP * STK_MATCAP];\n  double " << W << "[STK_MATCAP * STK_MATCAP];\n";
      auto ident = [&](const std::string& buf) { line("for (int " + i + " = 0; " + i + " < " + R + "; ++" + i + ") for (int " + j + " = 0; " + j + " < " + R + "; ++" + j + ") " + buf + "[" + i + " * " + R + " + " + j + "] = " + i + " == " + j + " ? 1.0 : 0.0;"); };
      if (inv) { ident(B); line("stk::lu_solve(" + V + ", " + R + ", " + B + ", " + R + ");"); line("for (int " + i + " = 0; " + i + " < " + R + " * " + R + "; ++" + i + ") " + V + "[" + i + "] = " + B + "[" + i + "];"); }
      line("stk::chol_lower(" + V + ", " + R + ");");
      // element (i, j) of A draws with element index (i - 1) p + j of the PNode's counter range
      line("for (int " + i + " = 1; " + i + " <= " + R + "; ++" + i + ") for (int " + j + " = 1; " + j + " <= " + R + "; ++" + j + ") { " + gen_decl(g, slot, "((" + i + " - 1) * " + R + " + " + j + ")") +
           " " + A + "[(" + i + " - 1) * " + R + " + " + j + " - 1] = " + i + " == " + j + " ? stk::draw_gamma(" + g + ", 0.5 * (double)(" + df + " - " + i + " + 1), 0.5) : (" + i + " > " + j + " ? stk::draw_normal(" + g + ", 0.0, 1.0) : 0.0); }");
      auto matmul = [&](const std::string& out, const std::string& X, const std::string& Y, bool yt) {      // out = X Y or X Y^T
        line("for (int " + i + " = 0; " + i + " < " + R + "; ++" + i + ") for (int " + j + " = 0; " + j + " < " + R + "; ++" + j + ") { double " + acc + " = 0.0; for (int " + k + " = 0; " + k + " < " + R + "; ++" + k + ") " +
             acc + " += " + X + "[" + i + " * " + R + " + " + k + "] * " + Y + (yt ? "[" + j + " * " + R + " + " + k + "]" : "[" + k + " * " + R + " + " + j + "]") + "; " + out + "[" + i + " * " + R + " + " + j + "] = " + acc + "; }");
      };
      matmul(B, V, A, false);              // B = L A
      matmul(W, B, B, true);               // W = B B^T = L A A^T L^T
      if (inv) { ident(B); line("stk::lu_solve(" + W + ", " + R + ", " + B + ", " + R + ");"); line("for (int " + i + " = 0; " + i + " < " + R + " * " + R + "; ++" + i + ") " + W + "[" + i + "] = " + B + "[" + i + "];"); }
      Val v; v.type = q.type; v.mat = stored_mat(W, R, R); bind(q.id, v);
      return true;
    }
    if (f == "lkj_corr") {
      std::string g = fresh("g"), r = fresh(q.id + "_r");
      line(gen_decl(g, slot, ""));
      line("const double " + r + " = stk::draw_lkj_corr2(" + g + ", " + rv(q.args.at(0), true) + ");");
      auto m = std::make_shared<Mat>(); m->rows = m->cols = "2"; m->prepare = [] {};
      m->at = [r](const std::string& i, const std::string& j) { return "((" + i + ") == (" + j + ") ? 1.0 : " + r + ")"; };
      Val v; v.type = q.type; v.mat = m; bind(q.id, v);
      return true;
    }
    return false;
  }

  void pnode(const PNode& q) {
    int level, ptr; id_level_ptr(q.id, &level, &ptr);
    // the PNode's own word of the Philox counter (slots of different PNodes must not be reachable from one another)
    if (level < 0 || level >= 256 || ptr < 0 || ptr >= 4096)
      die("PNode " + q.id + ": level/pointer outside the 256 x 4096 ids the Philox counter layout reserves");
    std::string slot = std::to_string(((unsigned)level << 24) | ((unsigned)ptr << 12)) + "u";
    if (q.k == PNode::Switch) { switch_node(q); return; }
    if (q.k == PNode::Loop) { loop_node(q, slot); return; }
    if (q.k == PNode::Chain) { chain_node(q); return; }
    if (q.k == PNode::Ordered) { ordered_node(q, slot); return; }
    if (vectorised_node(q, slot)) return;
    for (auto& a : q.args) need(a);
    if (matrix_dist_node(q, slot)) return;
    std::string g = fresh("g");
    line(gen_decl(g, slot, ""));
    std::string e = draw_expr(q, g);
    if (!q.type.scalar()) die("draw of type " + q.type.str() + " is not supported");
    if (q.sub_lo || q.sub_hi) {
      // ccPNode' on a SubrangeT (`truncated`, CC.hs:283-287): do { draw } while (out of range); the generator keeps
      // counting blocks across the retries.  The reference loops for ever on an impossible range; here the loop gives
      // up after 2^16 tries and keeps the last draw, so that a kernel cannot hang the device.
      need(q.sub_lo); need(q.sub_hi);
      bool real = q.type.k == Type::R;
      std::string cond, tries = fresh("tries");
      if (q.sub_lo) cond = q.id + " < " + rv(q.sub_lo, real);
      if (q.sub_hi) cond += (cond.empty() ? "" : " || ") + q.id + " > " + rv(q.sub_hi, real);
      line(ctype(q.type) + " " + q.id + "; int " + tries + " = 0;");
      line("do { " + q.id + " = " + e + "; } while ((" + cond + ") && ++" + tries + " < 65536);");
      Val v; v.type = q.type; v.expr = q.id; bind(q.id, v);
      return;
    }
    line("const " + ctype(q.type) + " " + q.id + " = " + e + ";");
    Val v; v.type = q.type; v.expr = q.id; bind(q.id, v);
  }

  // Switch over a union-typed head (CC.hs:306-316): one case per constructor, fields tied to ids
  void switch_node(const PNode& q) {
    need(q.args[0]);
    Type ut = type_of(q.args[0]);
    if (ut.k != Type::Union) die("switch head is not a union");
    // `Show PNode` prints no type for a Switch (Program.hs:142-153): take it from the first alternative's
    // return when that is a constructor, else the head's own type (a kernel t -> P t)
    Type qt = q.type;
    if (qt.k == Type::Unknown) {
      const RefP& r0 = q.alts.at(0).body.rets.at(0);
      qt = ut;
      if (r0->k == Ref::Data) qt = r0->type;
      else for (const Def& d0 : q.alts[0].body.defs) if (r0->k == Ref::Var && d0.id == r0->name && d0.node.type.k == Type::Union) qt = d0.node.type;
    }
    std::string st = union_struct(qt), hd = rv(q.args[0]);
    const UnionLayout& L = unions[ut.str()];
    line(st + " " + q.id + " = " + (qt == ut ? hd : st + "()") + ";");
    line("switch (" + hd + ".tag) {");
    for (size_t c = 0; c < q.alts.size(); ++c) {
      const Alt& a = q.alts[c];
      if (a.inputs.size() != L.ctor[c].size()) die("switch alternative C" + std::to_string(c) + " arity mismatch");
      line("case " + std::to_string(c) + ": {"); ind += 2; push();
      for (size_t j = 0; j < a.inputs.size(); ++j) {
        const Slot& s = L.slots[L.ctor[c][j]];
        line("const " + ctype(s.type) + " " + a.inputs[j] + " = " + hd + "." + s.name + ";");
        Val v; v.type = s.type; v.expr = a.inputs[j]; bind(a.inputs[j], v);
      }
      body(a.body);
      if (a.body.rets.size() != 1) die("switch alternative must return one value");
      assign_union(q.id, qt, a.body.rets[0]);
      line("break;"); pop(); ind -= 2; line("}");
    }
    line("}");
    Val v; v.type = qt; v.expr = q.id; bind(q.id, v);
  }

  // ccDAG (CC.hs:327-346): definitions in pointer order; a PNode is drawn right before its first use
  // (the reference places the draw at the node's `getExternal`), the rest at the end of the block.
  void body(const Body& b) {
    pending_stack.emplace_back();
    for (const PNode& q : b.pnodes) pending_stack.back()[q.id] = &q;
    pending_defs.push_back({{}, scopes.size()});
    if (rename_depth == 0) for (const Def& d : b.defs) pending_defs.back().defs[d.id] = &d;
    for (const Def& d : b.defs) {
      if (rename_depth == 0) { auto it = pending_defs.back().defs.find(d.id); if (it == pending_defs.back().defs.end()) continue; pending_defs.back().defs.erase(it); }
      run_def(d);
    }
    for (const PNode& q : b.pnodes) ensure_drawn(q.id);
    for (auto& r : b.rets) need(r);
    pending_defs.pop_back();
    pending_stack.pop_back();
  }
};

// ---------------------------------------------------------------------------------------------
// whole module (ccProgram, CC.hs:398-432)
// ---------------------------------------------------------------------------------------------
struct Field { std::string name; bool i32; std::string len; const char* kind = "state"; };

static std::vector<Type> parse_types(const std::string& s) {
  std::vector<Type> ts;
  Parser p(s);
  while (p.peek().k != Tok::End) { ts.push_back(p.type()); p.accept(";"); p.accept(","); }
  return ts;
}

static std::map<std::string, std::string> parse_opts(const char* o) {
  std::map<std::string, std::string> m;
  std::istringstream in(o ? o : "");
  std::string kv;
  while (in >> kv) { size_t e = kv.find('='); if (e == std::string::npos) m[kv] = "1"; else m[kv.substr(0, e)] = kv.substr(e + 1); }
  return m;
}

#include "stkc_grad.inc"

static std::string generate(const std::string& ir, const std::string& input_types, const char* options) {
  auto opt = parse_opts(options);
  Parser P(ir);
  Body prog = P.program();
  if (opt.count("mode") && opt["mode"] == "density") {     // log-density + gradient + HMC module of a model `P t`
    GradGen gg;
    if (opt.count("exact_div") && opt["exact_div"] != "0") gg.G.recip_div = false;      // keep every quotient a correctly rounded division
    return gg.generate(prog, opt);
  }
  std::vector<Type> its = parse_types(input_types);
  int level = opt.count("level") ? atoi(opt["level"].c_str()) : -1;
  std::string name = opt.count("name") ? opt["name"] : "ir";
  int veccap = opt.count("veccap") ? atoi(opt["veccap"].c_str()) : 64;
  // one warp per CTA by default: a generated kernel is a latency-bound instruction stream per chain, and 4096 chains are only
  // 128 warps — as CTAs of 128 threads they would sit on 32 of the 148 SMs (measured: profiles/r02_ncu_facts.json, grid 32)
  int block = opt.count("block") ? atoi(opt["block"].c_str()) : 32;
  int matcap = opt.count("matcap") ? atoi(opt["matcap"].c_str()) : 64;
  // entry=logp: the program is a log-density `s -> R` (what `lpdf` / `lpdfAux` build, Program.hs:893-1030, and what one runCC
  // call of it returns): its inputs are the state fields, its single real return goes to the module's `logp` entry
  const bool lp_entry = opt.count("entry") && opt["entry"] == "logp";
  bool kernel = !its.empty() && !lp_entry;
  if (level < 0) {                                  // find the Dummy level of the inputs: i_0_k (compileCC) or i_9_k (runStep)
    level = ir.find("i_9_0") != std::string::npos ? 9 : 0;
  }
  Gen g; g.veccap = veccap;
  g.lanes = opt.count("lanes") ? atoi(opt["lanes"].c_str()) : 1;
  if (g.lanes < 1 || g.lanes > 32 || (g.lanes & (g.lanes - 1)) || block % g.lanes) die("lanes must be a power of two <= 32 that divides the block size");
  g.push();
  // ---- state fields ----
  std::vector<Field> fields;
  std::ostringstream load, store, rec;
  std::vector<std::string> in_names;
  for (size_t k = 0; k < its.size(); ++k) {
    const Type& t = its[k];
    std::string id = "i_" + std::to_string(level) + "_" + std::to_string(k), fn = "in" + std::to_string(k);
    Val v; v.type = t;
    auto F = [&](const std::string& nm) { return "F_" + nm; };
    auto add_scalar = [&](const std::string& nm, bool i32, const std::string& cvar) {
      fields.push_back({nm, i32, "1"});
      load << "  " << (i32 ? "int " : "double ") << cvar << " = ((const " << (i32 ? "int" : "double") << "*)a.field[" << F(nm) << "])[chain];\n";
      store << "  ((" << (i32 ? "int" : "double") << "*)a.field[" << F(nm) << "])[chain] = " << cvar << ";\n";
      rec << " " << nm;
    };
    auto add_vec = [&](const std::string& nm, bool i32, const std::string& cvar) {
      fields.push_back({nm, i32, "0"}); fields.push_back({nm + "_len", true, "1"});
      load << "  " << (i32 ? "VecI " : "VecD ") << cvar << "((" << (i32 ? "int" : "double") << "*)a.field[" << F(nm) << "] + chain, a.n_chains, ((const int*)a.field[" << F(nm + "_len") << "])[chain]);\n";
      store << "  store_vec(" << cvar << ", (" << (i32 ? "int" : "double") << "*)a.field[" << F(nm) << "] + chain, a.n_chains, (int)a.flen[" << F(nm) << "], (int*)a.field[" << F(nm + "_len") << "] + chain, err_);\n";
    };
    if (t.k == Type::R) { add_scalar(fn, false, id); v.expr = id; }
    else if (t.integral()) { add_scalar(fn, true, id); v.expr = id; }
    else if (t.k == Type::Vec) { add_vec(fn, t.sub[0].k != Type::R, id); v.expr = id; }
    else if (t.k == Type::Union) {
      std::string st = g.union_struct(t);
      const UnionLayout& L = g.unions[t.str()];
      load << "  " << st << " " << id << ";\n";
      fields.push_back({fn + "_tag", true, "1"});
      load << "  " << id << ".tag = ((const int*)a.field[F_" << fn << "_tag])[chain];\n";
      store << "  ((int*)a.field[F_" << fn << "_tag])[chain] = " << id << ".tag;\n";
      rec << " " << fn << "_tag";
      for (const Slot& s : L.slots) {
        std::string nm = fn + "_" + s.name.substr(2);
        if (s.type.k == Type::Vec) {
          bool i32 = s.type.sub[0].k != Type::R;
          fields.push_back({nm, i32, "0"}); fields.push_back({nm + "_len", true, "1"});
          load << "  " << id << "." << s.name << " = " << (i32 ? "VecI" : "VecD") << "((" << (i32 ? "int" : "double") << "*)a.field[F_" << nm << "] + chain, a.n_chains, ((const int*)a.field[F_" << nm << "_len])[chain]);\n";
          store << "  store_vec(" << id << "." << s.name << ", (" << (i32 ? "int" : "double") << "*)a.field[F_" << nm << "] + chain, a.n_chains, (int)a.flen[F_" << nm << "], (int*)a.field[F_" << nm << "_len] + chain, err_);\n";
        } else {
          bool i32 = s.type.k != Type::R;
          fields.push_back({nm, i32, "1"});
          load << "  " << id << "." << s.name << " = ((const " << (i32 ? "int" : "double") << "*)a.field[F_" << nm << "])[chain];\n";
          store << "  ((" << (i32 ? "int" : "double") << "*)a.field[F_" << nm << "])[chain] = " << id << "." << s.name << ";\n";
          rec << " " << nm;
        }
      }
      v.expr = id;
    } else die("kernel input of type " + t.str() + " is not supported by the CUDA lowering");
    g.bind(id, v); in_names.push_back(id);
  }
  if (!P.input_name.empty() && !its.empty()) {      // `\input ->`: the inputs are the members of one tuple, input.0_k
    Val iv; iv.type.k = Type::Tuple; iv.type.sub = its; iv.tuple = in_names;
    g.bind(P.input_name, iv);
  }
  // ---- the transition body ----
  g.ind = 4;
  g.body(prog);
  if (kernel && prog.rets.size() != its.size()) die("a transition kernel must return as many values as it takes (t -> P t): " + std::to_string(its.size()) + " inputs, " + std::to_string(prog.rets.size()) + " returns");
  std::ostringstream feed;
  std::vector<Field> out_fields; std::ostringstream out_store;
  for (size_t k = 0; k < prog.rets.size(); ++k) {
    Type rt = g.type_of(prog.rets[k]);
    if (kernel) {
      if (!(rt == its[k]) && !(rt.integral() && its[k].integral())) die("return " + std::to_string(k) + " has type " + rt.str() + ", input has " + its[k].str());
      std::string e = g.rv(prog.rets[k], its[k].k == Type::R);
      std::string nx = "nx" + std::to_string(k);
      g.line((its[k].k == Type::Union ? g.union_struct(its[k]) : Gen::ctype(its[k])) + " " + nx + " = " + e + ";");
      if (its[k].k == Type::Vec) feed << "    assign_vec(" << in_names[k] << ", " << nx << ", (int)a.flen[F_in" << k << "], err_);\n";
      else if (its[k].k == Type::Union) {
        const UnionLayout& L = g.unions[its[k].str()];
        feed << "    " << in_names[k] << ".tag = " << nx << ".tag;\n";
        for (const Slot& s : L.slots) {
          if (s.type.k == Type::Vec) feed << "    assign_vec(" << in_names[k] << "." << s.name << ", " << nx << "." << s.name << ", (int)a.flen[F_in" << k << "_" << s.name.substr(2) << "], err_);\n";
          else feed << "    " << in_names[k] << "." << s.name << " = " << nx << "." << s.name << ";\n";
        }
      } else feed << "    " << in_names[k] << " = " << nx << ";\n";
    } else if (lp_entry) {
      if (prog.rets.size() != 1 || !rt.scalar()) die("entry=logp: the program must return one scalar (a log-density), got " + rt.str());
      g.line("a.out_lp[chain] = " + g.rv(prog.rets[k], true) + ";");
    } else {
      std::string nm = "out" + std::to_string(k);
      if (rt.k == Type::Vec && rt.sub[0].scalar()) {   // a joint's draws: [veccap][chains] + the length
        bool i32 = rt.sub[0].k != Type::R;
        out_fields.push_back({nm, i32, std::to_string(veccap)}); out_fields.push_back({nm + "_len", true, "1"});
        g.line("store_vec(" + g.rv(prog.rets[k]) + ", (" + (i32 ? "int" : "double") + "*)a.field[F_" + nm + "] + chain, a.n_chains, " + std::to_string(veccap) +
               ", (int*)a.field[F_" + nm + "_len] + chain, err_);");
        continue;
      }
      if (rt.k == Type::Mat) {                         // a matrix value: [matcap * matcap][chains] row-major + its dimensions
        std::shared_ptr<Mat> m = g.mat_of(prog.rets[k]);
        m->prepare();
        out_fields.push_back({nm, false, std::to_string(matcap * matcap)}); out_fields.push_back({nm + "_rows", true, "1"}); out_fields.push_back({nm + "_cols", true, "1"});
        std::string R = g.fresh("R"), C = g.fresh("C"), i = g.fresh("i"), j = g.fresh("j");
        g.line("{ const int " + R + " = stk::cap_len(" + m->rows + ", STK_MATCAP, err_, 3), " + C + " = stk::cap_len(" + m->cols + ", STK_MATCAP, err_, 3);");
        g.line("  for (int " + i + " = 1; " + i + " <= " + R + "; ++" + i + ") for (int " + j + " = 1; " + j + " <= " + C + "; ++" + j + ") ((double*)a.field[F_" + nm + "])[(long long)((" + i + " - 1) * " + C + " + " + j + " - 1) * a.n_chains + chain] = " + m->at(i, j) + ";");
        g.line("  ((int*)a.field[F_" + nm + "_rows])[chain] = " + R + "; ((int*)a.field[F_" + nm + "_cols])[chain] = " + C + "; }");
        continue;
      }
      if (!rt.scalar()) die("simulate: return of type " + rt.str() + " is not supported");
      out_fields.push_back({nm, rt.k != Type::R, "1"});
      g.line("((" + std::string(rt.k != Type::R ? "int" : "double") + "*)a.field[F_" + nm + "])[chain] = " + g.rv(prog.rets[k], rt.k == Type::R) + ";");
      rec << " " << nm;
    }
  }
  // probe=<id>,<id>..: the value of these top-level definitions after the last transition of a launch, as state fields
  // probe<k> — how the parity tests read intermediate quantities (log acceptance ratio, densities) of a generated kernel
  if (opt.count("probe")) {
    std::istringstream ps(opt["probe"]); std::string id; int k = 0;
    while (std::getline(ps, id, ',')) {
      if (id.empty()) continue;
      auto r = std::make_shared<Ref>(); r->k = Ref::Var; r->name = id;
      std::string nm = "probe" + std::to_string(k++);
      out_fields.push_back({nm, false, "1"});
      g.line("((double*)a.field[F_" + nm + "])[chain] = " + g.rv(r, true) + ";");
    }
  }
  std::string guard_cond;
  if (opt.count("guard")) {
    if (kernel) die("guard= applies to simulate programs (a kernel's conditions are part of its acceptance ratio)");
    std::istringstream gs(opt["guard"]); std::string id;
    while (std::getline(gs, id, ',')) {
      if (id.empty()) continue;
      auto r = std::make_shared<Ref>(); r->k = Ref::Var; r->name = id;
      guard_cond += (guard_cond.empty() ? "(" : " && (") + g.rv(r) + ")";
    }
  }
  for (auto& f : out_fields) fields.push_back(f);
  { Field ef{"err", true, "1"}; ef.kind = "work"; fields.push_back(ef); }   // overflow flag, checked by the runtime after the launch
  if ((int)fields.size() > 48) die("too many state fields (the module ABI carries 48, stk_abi.h)");

  std::ostringstream o;
  o << "// generated by stkc (libstk_cuda, Language.CUDA lowering) — do not edit\n";
  o << "#define STK_VECCAP " << veccap << "\n#define STK_MATCAP " << matcap << "\n#define STK_LANES " << g.lanes << "\n";
  o << "#include \"stk_device.cuh\"\nusing namespace stk;\n";
  o << "enum {";
  for (size_t i = 0; i < fields.size(); ++i) o << (i ? ", " : " ") << "F_" << fields[i].name << " = " << i;
  o << " };\n";
  for (auto& kv : g.union_cname) {
    const UnionLayout& L = g.unions[kv.first];
    o << "struct " << kv.second << " { int tag;";
    for (const Slot& s : L.slots) o << " " << Gen::ctype(s.type) << " " << s.name << ";";
    o << " };   // " << kv.first << "\n";
  }
  const char* sym = kernel ? "stk_ir_step" : lp_entry ? "stk_ir_logp" : "stk_ir_simulate";
  o << "extern \"C\" __global__ void __launch_bounds__(" << block << ") " << sym << "(const __grid_constant__ StkArgs a) {\n";
  if (g.lanes > 1) {
    // STK_LANES consecutive lanes per chain: they run the same instruction stream on the same state and draws, and split the
    // terms of additive folds between them (Gen::fold)
    o << "  const long long gtid_ = (long long)blockIdx.x * blockDim.x + threadIdx.x;\n";
    o << "  const long long chain = gtid_ / STK_LANES; const int sub_ = (int)(gtid_ % STK_LANES);\n";
    o << "  const unsigned gmask_ = (STK_LANES == 32 ? 0xffffffffu : ((1u << STK_LANES) - 1u) << ((threadIdx.x & 31) & ~(STK_LANES - 1)));\n";
    o << "  (void)sub_; (void)gmask_;\n  if (chain >= a.n_chains) return;\n";
  } else o << "  const long long chain = (long long)blockIdx.x * blockDim.x + threadIdx.x;\n  if (chain >= a.n_chains) return;\n";
  o << "  int* const err_ = (int*)a.field[F_err];\n";
  o << load.str() << g.decls.str();
  if (g.need_matbuf) o << "  double mat_[STK_MATCAP * STK_MATCAP];\n";
  if (kernel) {
    o << "  for (long long t_ = 0; t_ < a.n_steps; ++t_) {\n";
    o << "    Rng rng; rng.init(a.seed, (u64)(a.chain0 + chain), a.step0 + (u64)t_);\n";
    o << g.out.str() << feed.str();
    o << "  }\n" << store.str() << "}\n";
  } else {
    if (!guard_cond.empty()) {
      // conditions of the program (`guard`, Program.hs:745-760): samplePBlock re-draws the WHOLE program until every condition holds
      // (Program.hs:1046-1058).  Attempt k draws from the streams of transition 2^64 - 1 - k; the loop gives up after 2^16 attempts
      // and reports it through the err flag (code 4) instead of hanging the device.
      o << "  unsigned try_ = 0;\n  for (; try_ < 65536u; ++try_) {\n    Rng rng; rng.init(a.seed, (u64)(a.chain0 + chain), ~0ULL - (u64)try_);\n"
        << g.out.str() << "    if (" << guard_cond << ") break;\n  }\n  if (try_ == 65536u) *err_ = 4;\n}\n";
    } else
    o << "  {\n    Rng rng; rng.init(a.seed, (u64)(a.chain0 + chain), ~0ULL);\n" << g.out.str() << "  }\n}\n";
  }
  o << "extern \"C\" __device__ const char stk_desc[] =\n  \"stk-module 1\\n\"\n  \"name " << name << "\\n\"\n  \"nparam 0\\n\"\n";
  for (auto& f : fields) o << "  \"field " << f.name << " " << f.kind << " " << (f.i32 ? "i32" : "f64") << " " << f.len << "\\n\"\n";
  if (!rec.str().empty()) o << "  \"record" << rec.str() << "\\n\"\n";
  if (kernel) o << "  \"kernel step " << sym << " " << block << " " << block / g.lanes << " 0 - 0\\n\";\n";
  else o << "  \"entry " << (lp_entry ? "logp " : "simulate ") << sym << " " << block << " " << block / g.lanes << " 0 - 0\\n\";\n";
  return o.str();
}

}  // namespace stkc

extern "C" int stk_set_error_(int code, const char* msg);   // stk_runtime.cpp

extern "C" int stk_ir_to_cuda(const char* ir_text, const char* input_types, const char* options,
                              char* out, size_t cap, size_t* needed) {
  if (!ir_text) return stk_set_error_(STK_ERR_ARG, "null IR text");
  try {
    std::string s = stkc::generate(ir_text, input_types ? input_types : "", options);
    if (needed) *needed = s.size() + 1;
    if (!out || cap == 0) return STK_OK;
    if (s.size() + 1 > cap) return stk_set_error_(STK_ERR_ARG, "output buffer too small");
    memcpy(out, s.c_str(), s.size() + 1);
    return STK_OK;
  } catch (const stkc::Error& e) {
    return stk_set_error_(STK_ERR_UNSUPPORTED, e.what());
  } catch (const std::exception& e) {
    return stk_set_error_(STK_ERR_ARG, e.what());
  }
}
