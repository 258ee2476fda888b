// Miniature of the type passes that run after static_kernel_pick (reference:
// lite/core/optimizer.h:136-168 -- type_target_cast_pass, type_precision_cast_pass,
// type_layout_cast_pass, each followed by variable_place_inference), plus the extension
// SURVEY 8f rank 1 asks for: ops that only MOVE data (concat, split, shuffle_channel, relu, relu6)
// carry no enable_int8 attribute in a quantised model, so the reference picks their fp32 NCHW
// (or host-only) kernels and surrounds them with calib / layout / io_copy pairs.  With the
// extension such an op inherits int8 when everything around it is int8, picks the
// {op}/int8_nhwc kernel, and the activations stay int8-NHWC-kCUDA between convs.
//
// Rules restated:
//   * kernel pick per op: valid places in preference order; int8 kernels only for an int8 op; the
//     int8 candidate whose DECLARED output precision matches "every consumer is int8"
//     (lite/core/mir/static_kernel_pick_pass.cc:93-166);
//   * a variable takes the declared output type of its producer's picked kernel
//     (lite/core/mir/variable_place_inference_pass.h);
//   * for every (variable -> consumer input): target mismatch -> io_copy
//     (type_target_cast_pass.cc), then precision mismatch -> calib (type_precision_cast_pass.cc:
//     126-160, PrecisionCompatibleTo of lite/core/type_system.h:203-210), then layout mismatch ->
//     layout (type_layout_cast_pass.cc, DataLayoutCompatibleTo of type_system.h:191-195); one cast
//     node per (variable, kind) is shared by all consumers (the passes' cast_nodes map).
// The planner works on DECLARATIONS only (ParamTypeRegistry + a caller-supplied table for the
// reference's own fp32 kernels), so it runs without a GPU.
#pragma once
#include <map>
#include <set>
#include <sstream>
#include <string>
#include <vector>

#include "lite/core/type_system.h"

namespace paddle {
namespace lite {
namespace mir {

struct MiniOp {
  std::string type;
  std::vector<std::string> inputs, outputs;  // activation variables only
  bool enable_int8{false};                   // the op desc attribute (set by quant_dequant_fuse)
};

struct MiniGraph {
  std::vector<MiniOp> ops;  // topological order; "feed" / "fetch" ops mark the graph boundary
};

struct KernelDeclLite {
  std::string op, alias;
  Place place;
  Type in, out;  // declared type of the activation inputs / outputs
};

struct InsertedCast {
  std::string kind;  // "io_copy" | "calib" | "layout"
  std::string var, consumer;
  Type from, to;
  int bytes_per_elem{0};  // read + write of the cast kernel
};

struct TypePlan {
  std::vector<std::string> picked;  // "op/alias/place" per op of the graph ("" = none found)
  std::vector<InsertedCast> casts;
  int casts_between_int8_ops{0};    // casts that are neither at a feed nor at a fetch
};

inline bool IsTransparentOp(const std::string& t) {
  return t == "concat" || t == "split" || t == "shuffle_channel" || t == "relu" || t == "relu6";
}

inline bool IsHostTarget(TargetType t) { return t == TARGET(kHost) || t == TARGET(kX86) || t == TARGET(kARM); }

// declarations of every kernel registered in this library
inline std::vector<KernelDeclLite> RegisteredDecls() {
  std::vector<KernelDeclLite> out;
  for (const auto& kv : ParamTypeRegistry::Global().all()) {
    KernelDeclLite d;
    const std::string& key = kv.first;  // op/alias/place
    const size_t a = key.find('/'), b = key.find('/', a + 1);
    if (a == std::string::npos || b == std::string::npos) continue;
    d.op = key.substr(0, a);
    d.alias = key.substr(a + 1, b - a - 1);
    d.place = kv.second.place;
    bool have_in = false, have_out = false;
    for (const char* n : {"Input", "X", "x"}) {
      auto it = kv.second.inputs.find(n);
      if (it != kv.second.inputs.end()) { d.in = it->second; have_in = true; break; }
    }
    for (const char* n : {"Output", "Out"}) {
      auto it = kv.second.outputs.find(n);
      if (it != kv.second.outputs.end()) { d.out = it->second; have_out = true; break; }
    }
    if (have_in && have_out) out.push_back(d);
  }
  return out;
}

// the reference's own fp32 registrations for the ops a quantised CNN uses besides the int8 ones
// (kCUDA where it has a kernel: lite/kernels/cuda/{conv,concat,relu,pool,elementwise,softmax,scale}_compute.*;
// split / shuffle_channel exist for kARM only: lite/kernels/arm/{split,shuffle_channel}_compute.cc,
// i.e. a host kernel from the GPU's point of view; feed / fetch: lite/kernels/host/{feed,fetch}_compute.cc)
inline std::vector<KernelDeclLite> ReferenceFp32Decls() {
  const Type cuda_f{TARGET(kCUDA), PRECISION(kFloat), DATALAYOUT(kNCHW)};
  const Type host_f{TARGET(kHost), PRECISION(kFloat), DATALAYOUT(kNCHW)};
  const Type host_any{TARGET(kHost), PRECISION(kAny), DATALAYOUT(kAny)};
  std::vector<KernelDeclLite> v;
  for (const char* op : {"conv2d", "depthwise_conv2d", "concat", "relu", "pool2d", "elementwise_add", "softmax", "scale"})
    v.push_back({op, "def", Place(TARGET(kCUDA), PRECISION(kFloat), DATALAYOUT(kNCHW)), cuda_f, cuda_f});
  for (const char* op : {"split", "shuffle_channel", "relu6"})
    v.push_back({op, "def", Place(TARGET(kHost), PRECISION(kFloat), DATALAYOUT(kNCHW)), host_f, host_f});
  v.push_back({"feed", "def", Place(TARGET(kHost), PRECISION(kAny), DATALAYOUT(kAny)), host_any, host_f});
  v.push_back({"fetch", "def", Place(TARGET(kHost), PRECISION(kAny), DATALAYOUT(kAny)), host_any, host_any});
  return v;
}

inline int ElemBytes(PrecisionType p) { return p == PRECISION(kInt8) ? 1 : 4; }

inline std::string TypeStr(const Type& t) {
  return std::string(lite_api::TargetToStr(t.target)) + "/" + lite_api::PrecisionToStr(t.precision) + "/" +
         lite_api::DataLayoutToStr(t.layout);
}

inline TypePlan PlanTypes(const MiniGraph& g, const std::vector<Place>& valid_places,
                          const std::vector<KernelDeclLite>& decls, bool int8_resident_extension) {
  const size_t n = g.ops.size();
  std::map<std::string, int> producer;              // var -> op index
  std::map<std::string, std::vector<int>> consumers;  // var -> op indices
  for (size_t i = 0; i < n; ++i) {
    for (const auto& v : g.ops[i].outputs) producer[v] = static_cast<int>(i);
    for (const auto& v : g.ops[i].inputs) consumers[v].push_back(static_cast<int>(i));
  }
  // ---- which ops are int8: the attribute, or (extension) a transparent op inside an int8 region.
  // Fixpoint from "every transparent op is int8": one loses it when a producer or a consumer is not.
  std::vector<char> is_int8(n, 0);
  for (size_t i = 0; i < n; ++i)
    is_int8[i] = g.ops[i].enable_int8 || (int8_resident_extension && IsTransparentOp(g.ops[i].type));
  for (bool changed = int8_resident_extension; changed;) {
    changed = false;
    for (size_t i = 0; i < n; ++i) {
      if (!is_int8[i] || g.ops[i].enable_int8) continue;
      bool ok = true;
      for (const auto& v : g.ops[i].inputs) {
        auto it = producer.find(v);
        ok = ok && it != producer.end() && is_int8[it->second];
      }
      for (const auto& v : g.ops[i].outputs) {
        const auto& cs = consumers[v];
        ok = ok && !cs.empty();
        for (int c : cs) ok = ok && is_int8[c];
      }
      if (!ok) { is_int8[i] = 0; changed = true; }
    }
  }
  // ---- static kernel pick on declarations
  TypePlan plan;
  std::vector<const KernelDeclLite*> picked(n, nullptr);
  for (size_t i = 0; i < n; ++i) {
    const MiniOp& op = g.ops[i];
    bool out_int8 = is_int8[i];
    bool any_consumer = false;
    for (const auto& v : op.outputs)
      for (int c : consumers[v]) { any_consumer = true; out_int8 = out_int8 && is_int8[c]; }
    out_int8 = out_int8 && any_consumer;
    const PrecisionType want_out = out_int8 ? PRECISION(kInt8) : PRECISION(kFloat);
    std::vector<Place> places = valid_places;
    if (op.type == "feed" || op.type == "fetch") places = {Place(TARGET(kHost), PRECISION(kAny), DATALAYOUT(kAny))};
    for (const auto& place : places) {
      if (place.precision == PRECISION(kInt8) && !is_int8[i]) continue;
      for (const auto& d : decls) {
        if (d.op != op.type || !(d.place == place)) continue;
        if (place.precision == PRECISION(kInt8) && d.out.precision != want_out) continue;
        picked[i] = &d;
        break;
      }
      if (picked[i]) break;
    }
    plan.picked.push_back(picked[i] ? picked[i]->op + "/" + picked[i]->alias + "/" + picked[i]->place.DebugString() : "");
  }
  // ---- variable types, then the three cast passes per consumer input
  std::map<std::string, Type> var_type;
  std::map<std::string, Type> cast_out;  // "var|kind" -> type after that cast (shared by consumers)
  for (size_t i = 0; i < n; ++i) {
    const MiniOp& op = g.ops[i];
    if (!picked[i]) continue;
    for (const auto& v : op.inputs) {
      auto it = var_type.find(v);
      if (it == var_type.end()) continue;
      Type cur = it->second;
      const Type& want = picked[i]->in;
      const bool boundary = op.type == "fetch" || (producer.count(v) && g.ops[producer[v]].type == "feed");
      auto add = [&](const char* kind, const Type& to, int bytes) {
        const std::string key = v + "|" + kind + "|" + TypeStr(to);
        if (!cast_out.count(key)) {
          cast_out[key] = to;
          plan.casts.push_back({kind, v, op.type, cur, to, bytes});
          if (!boundary) ++plan.casts_between_int8_ops;
        }
        cur = to;
      };
      const bool target_ok = IsHostTarget(cur.target) ? IsHostTarget(want.target) : cur.target == want.target;
      if (!target_ok) add("io_copy", Type{want.target, cur.precision, cur.layout}, 2 * ElemBytes(cur.precision));
      if (!(cur.precision == want.precision || want.precision == PRECISION(kAny) || cur.precision == PRECISION(kAny)))
        add("calib", Type{cur.target, want.precision, cur.layout}, ElemBytes(cur.precision) + ElemBytes(want.precision));
      if (!(cur.layout == want.layout || want.layout == DATALAYOUT(kAny)))
        add("layout", Type{cur.target, cur.precision, want.layout}, 2 * ElemBytes(cur.precision));
    }
    for (const auto& v : op.outputs) var_type[v] = picked[i]->out;
  }
  return plan;
}

// text form used by the test harness: one op per line "type;in1,in2;out1,out2;enable_int8"
inline MiniGraph ParseMiniGraph(const std::string& text) {
  MiniGraph g;
  std::istringstream ss(text);
  std::string line;
  auto split = [](const std::string& s, char sep) {
    std::vector<std::string> out;
    std::string cur;
    for (char c : s) {
      if (c == sep) { if (!cur.empty()) out.push_back(cur); cur.clear(); }
      else cur.push_back(c);
    }
    if (!cur.empty()) out.push_back(cur);
    return out;
  };
  while (std::getline(ss, line)) {
    if (line.empty() || line[0] == '#') continue;
    std::vector<std::string> f;
    std::string cur;
    for (char c : line) {
      if (c == ';') { f.push_back(cur); cur.clear(); }
      else cur.push_back(c);
    }
    f.push_back(cur);
    if (f.size() < 4) continue;
    MiniOp op;
    op.type = f[0];
    op.inputs = split(f[1], ',');
    op.outputs = split(f[2], ',');
    op.enable_int8 = f[3] == "1";
    g.ops.push_back(op);
  }
  return g;
}

inline std::string PlanToText(const MiniGraph& g, const TypePlan& p) {
  std::ostringstream os;
  for (size_t i = 0; i < g.ops.size(); ++i) os << "op " << g.ops[i].type << " -> " << (p.picked[i].empty() ? "NONE" : p.picked[i]) << "\n";
  int bytes = 0;
  for (const auto& c : p.casts) {
    os << "cast " << c.kind << " " << c.var << " before " << c.consumer << " " << TypeStr(c.from) << " -> " << TypeStr(c.to)
       << " bytes_per_elem " << c.bytes_per_elem << "\n";
    bytes += c.bytes_per_elem;
  }
  os << "casts " << p.casts.size() << " inside " << p.casts_between_int8_ops << " bytes_per_elem_total " << bytes << "\n";
  return os.str();
}

}  // namespace mir
}  // namespace lite
}  // namespace paddle
