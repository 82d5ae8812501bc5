// hs_ingest.cc — the data formats on the INPUT side of the hot path: the node register
// annotations that describe a node's devices and the pod annotations that record a committed allocation.
// Plain host C++ (no CUDA); the entry points that change the HBM inventory (hs_node_register, hs_pod_ingest,
// hs_pod_shrink) live in hs_engine.cu and call the decoders here.
//
//   decode_node_csv     device.DecodeNodeDevices      pkg/device/devices.go:317-381   (DCU, iluvatar: "id,count,mem,core,type,numa,health[,index,mode]:")
//   decode_node_json    device.UnMarshalNodeDevices   pkg/device/devices.go:445-449   (NVIDIA, Ascend, …: JSON array of DeviceInfo, tags devices.go:128-140)
//   decode_pair_scores  device.DecodePairScores       pkg/device/devices.go:383-389 + the fit loop nvidia/device.go:319-340
//   decode_pod_devices  device.DecodePodDevices       pkg/device/devices.go:526-556 (+ DecodeContainerDevices :496-524)
//   collapse            device.CollapseInitContainerUsage / AppContainersOnlyDeviceUsage   pkg/device/initContainer.go:36-157
#include "hs_ingest.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <memory>

namespace hs_ingest {
namespace {

// ---- strconv look-alikes: same accept/reject set and the same error text as Go's ----------------------
bool all_digits(const std::string& s, size_t from) {
  if (from >= s.size()) return false;
  for (size_t i = from; i < s.size(); i++)
    if (s[i] < '0' || s[i] > '9') return false;
  return true;
}
std::string quote(const std::string& s) {   // strconv.Quote for the plain-ASCII strings that occur here
  std::string o = "\"";
  for (char c : s) {
    if (c == '"' || c == '\\') { o += '\\'; o += c; }
    else if (c == '\n') o += "\\n";
    else if (c == '\t') o += "\\t";
    else if (c == '\r') o += "\\r";
    else o += c;
  }
  return o + "\"";
}
// strconv.ParseInt(s, 10, bits) / strconv.Atoi (fn names the function in the error, Atoi parses as int64)
bool parse_int(const std::string& s, int bits, const char* fn, int64_t& out, std::string& err) {
  size_t i = (!s.empty() && (s[0] == '+' || s[0] == '-')) ? 1 : 0;
  if (!all_digits(s, i)) { err = std::string("strconv.") + fn + ": parsing " + quote(s) + ": invalid syntax"; return false; }
  const bool neg = s[0] == '-';
  unsigned long long v = 0;
  bool over = false;
  for (; i < s.size(); i++) {
    if (v > (~0ull - 9) / 10) { over = true; break; }
    v = v * 10 + (unsigned)(s[i] - '0');
  }
  const unsigned long long lim = 1ull << (bits - 1);
  if (over || (neg ? v > lim : v > lim - 1)) { err = std::string("strconv.") + fn + ": parsing " + quote(s) + ": value out of range"; return false; }
  out = neg ? -(int64_t)v : (int64_t)v;
  return true;
}
bool parse_bool(const std::string& s, bool& out, std::string& err) {
  static const char* t[] = {"1", "t", "T", "TRUE", "true", "True"};
  static const char* f[] = {"0", "f", "F", "FALSE", "false", "False"};
  for (auto x : t) if (s == x) { out = true; return true; }
  for (auto x : f) if (s == x) { out = false; return true; }
  err = "strconv.ParseBool: parsing " + quote(s) + ": invalid syntax";
  return false;
}
std::vector<std::string> split(const std::string& s, char sep) {   // strings.Split
  std::vector<std::string> out;
  size_t b = 0;
  for (;;) {
    size_t e = s.find(sep, b);
    if (e == std::string::npos) { out.push_back(s.substr(b)); break; }
    out.push_back(s.substr(b, e - b));
    b = e + 1;
  }
  return out;
}

// ---- a small strict JSON reader (RFC 8259) --------------------------------------------------------------

struct JV {
  enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
  bool b = false;
  std::string s;                                   // Str: decoded text; Num: the literal as written
  std::vector<JV> a;
  std::vector<std::pair<std::string, JV>> o;       // members in document order (a later duplicate wins, like Go)
};

struct JParser {
  const std::string& t;
  size_t i = 0;
  std::string err;
  int depth = 0;
  explicit JParser(const std::string& text) : t(text) {}
  void ws() { while (i < t.size() && (t[i] == ' ' || t[i] == '\t' || t[i] == '\n' || t[i] == '\r')) i++; }
  bool fail(const std::string& m) { if (err.empty()) err = m; return false; }
  bool lit(const char* w) {
    size_t n = strlen(w);
    if (t.compare(i, n, w) != 0) return fail(std::string("invalid character near offset ") + std::to_string(i));
    i += n;
    return true;
  }
  static void utf8(std::string& o, uint32_t c) {
    if (c < 0x80) o += (char)c;
    else if (c < 0x800) { o += (char)(0xC0 | (c >> 6)); o += (char)(0x80 | (c & 63)); }
    else if (c < 0x10000) { o += (char)(0xE0 | (c >> 12)); o += (char)(0x80 | ((c >> 6) & 63)); o += (char)(0x80 | (c & 63)); }
    else { o += (char)(0xF0 | (c >> 18)); o += (char)(0x80 | ((c >> 12) & 63)); o += (char)(0x80 | ((c >> 6) & 63)); o += (char)(0x80 | (c & 63)); }
  }
  bool hex4(uint32_t& v) {
    if (i + 4 > t.size()) return fail("unexpected end of JSON input");
    v = 0;
    for (int k = 0; k < 4; k++) {
      char c = t[i++];
      v <<= 4;
      if (c >= '0' && c <= '9') v |= (uint32_t)(c - '0');
      else if (c >= 'a' && c <= 'f') v |= (uint32_t)(c - 'a' + 10);
      else if (c >= 'A' && c <= 'F') v |= (uint32_t)(c - 'A' + 10);
      else return fail("invalid character in \\u escape");
    }
    return true;
  }
  bool str(std::string& out) {
    i++;  // opening quote
    for (;;) {
      if (i >= t.size()) return fail("unexpected end of JSON input");
      unsigned char c = (unsigned char)t[i++];
      if (c == '"') return true;
      if (c < 0x20) return fail("invalid character in string literal");
      if (c != '\\') { out += (char)c; continue; }
      if (i >= t.size()) return fail("unexpected end of JSON input");
      char e = t[i++];
      switch (e) {
        case '"': out += '"'; break;
        case '\\': out += '\\'; break;
        case '/': out += '/'; break;
        case 'b': out += '\b'; break;
        case 'f': out += '\f'; break;
        case 'n': out += '\n'; break;
        case 'r': out += '\r'; break;
        case 't': out += '\t'; break;
        case 'u': {
          uint32_t u;
          if (!hex4(u)) return false;
          if (u >= 0xD800 && u < 0xDC00 && i + 6 <= t.size() && t[i] == '\\' && t[i + 1] == 'u') {
            size_t save = i;
            i += 2;
            uint32_t lo;
            if (!hex4(lo)) return false;
            if (lo >= 0xDC00 && lo < 0xE000) u = 0x10000 + ((u - 0xD800) << 10) + (lo - 0xDC00);
            else { i = save; u = 0xFFFD; }
          } else if (u >= 0xD800 && u < 0xE000) u = 0xFFFD;
          utf8(out, u);
          break;
        }
        default: return fail("invalid character in string escape code");
      }
    }
  }
  bool num(JV& v) {
    size_t b = i;
    if (i < t.size() && t[i] == '-') i++;
    if (i >= t.size()) return fail("unexpected end of JSON input");
    if (t[i] == '0') i++;
    else if (t[i] >= '1' && t[i] <= '9') { while (i < t.size() && t[i] >= '0' && t[i] <= '9') i++; }
    else return fail("invalid character in numeric literal");
    if (i < t.size() && t[i] == '.') {
      i++;
      if (i >= t.size() || t[i] < '0' || t[i] > '9') return fail("invalid character after decimal point in numeric literal");
      while (i < t.size() && t[i] >= '0' && t[i] <= '9') i++;
    }
    if (i < t.size() && (t[i] == 'e' || t[i] == 'E')) {
      i++;
      if (i < t.size() && (t[i] == '+' || t[i] == '-')) i++;
      if (i >= t.size() || t[i] < '0' || t[i] > '9') return fail("invalid character in exponent of numeric literal");
      while (i < t.size() && t[i] >= '0' && t[i] <= '9') i++;
    }
    v.kind = JV::Num;
    v.s = t.substr(b, i - b);
    return true;
  }
  bool value(JV& v) {
    ws();
    if (i >= t.size()) return fail("unexpected end of JSON input");
    if (++depth > 200) return fail("exceeded max depth");
    bool ok = false;
    char c = t[i];
    if (c == '{') {
      v.kind = JV::Obj;
      i++;
      ws();
      if (i < t.size() && t[i] == '}') { i++; ok = true; }
      else for (;;) {
        ws();
        if (i >= t.size() || t[i] != '"') { fail("invalid character looking for beginning of object key string"); break; }
        std::string k;
        if (!str(k)) break;
        ws();
        if (i >= t.size() || t[i] != ':') { fail("invalid character after object key"); break; }
        i++;
        JV m;
        if (!value(m)) break;
        v.o.emplace_back(std::move(k), std::move(m));
        ws();
        if (i < t.size() && t[i] == ',') { i++; continue; }
        if (i < t.size() && t[i] == '}') { i++; ok = true; break; }
        fail("invalid character after object key:value pair");
        break;
      }
    } else if (c == '[') {
      v.kind = JV::Arr;
      i++;
      ws();
      if (i < t.size() && t[i] == ']') { i++; ok = true; }
      else for (;;) {
        JV m;
        if (!value(m)) break;
        v.a.push_back(std::move(m));
        ws();
        if (i < t.size() && t[i] == ',') { i++; continue; }
        if (i < t.size() && t[i] == ']') { i++; ok = true; break; }
        fail("invalid character after array element");
        break;
      }
    } else if (c == '"') { v.kind = JV::Str; ok = str(v.s); }
    else if (c == 't') { v.kind = JV::Bool; v.b = true; ok = lit("true"); }
    else if (c == 'f') { v.kind = JV::Bool; v.b = false; ok = lit("false"); }
    else if (c == 'n') { v.kind = JV::Null; ok = lit("null"); }
    else if (c == '-' || (c >= '0' && c <= '9')) ok = num(v);
    else fail("invalid character looking for beginning of value");
    depth--;
    return ok;
  }
  bool document(JV& v) {
    if (!value(v)) return false;
    ws();
    if (i != t.size()) return fail("invalid character after top-level value");
    return true;
  }
};

// encoding/json matches object keys to struct fields exactly first, then case-insensitively
bool key_is(const std::string& k, const char* name) {
  size_t n = strlen(name);
  if (k.size() != n) return false;
  for (size_t i = 0; i < n; i++) {
    char a = k[i], b = name[i];
    if (a >= 'A' && a <= 'Z') a = (char)(a - 'A' + 'a');
    if (b >= 'A' && b <= 'Z') b = (char)(b - 'A' + 'a');
    if (a != b) return false;
  }
  return true;
}
const char* kind_name(const JV& v) {
  switch (v.kind) {
    case JV::Null: return "null"; case JV::Bool: return "bool"; case JV::Num: return "number";
    case JV::Str: return "string"; case JV::Arr: return "array"; default: return "object";
  }
}
// a JSON number into a Go integer field: the literal must be a plain integer that fits (else UnmarshalTypeError)
bool num_to_int(const JV& v, int bits, bool is_unsigned, int64_t& out) {
  if (v.s.find_first_of(".eE") != std::string::npos) return false;
  if (is_unsigned && !v.s.empty() && v.s[0] == '-') return false;
  std::string e;
  return parse_int(v.s, is_unsigned ? 64 : bits, "ParseInt", out, e);
}
std::string type_err(const JV& v, const char* field, const char* gotype) {
  return std::string("json: cannot unmarshal ") + kind_name(v) + " into Go struct field DeviceInfo." + field + " of type " + gotype;
}

}  // namespace

// ---- device.DecodeNodeDevices (devices.go:317-381) -------------------------------------------------------
bool decode_node_csv(const std::string& s, std::vector<Device>& out, std::string& err) {
  out.clear();
  if (s.find(':') == std::string::npos) { err = "node annotation missing device separator"; return false; }
  for (const std::string& val : split(s, ':')) {
    if (val.empty()) continue;
    if (val.find(',') == std::string::npos) { err = "malformed node annotation segment: " + quote(val); return false; }
    const std::vector<std::string> it = split(val, ',');
    if (it.size() != 7 && it.size() != 9) {
      err = "unexpected field count " + std::to_string(it.size()) + " in node annotation";
      return false;
    }
    Device d;
    int64_t v;
    std::string e;
    bool hb;
    if (!parse_int(it[1], 32, "ParseInt", v, e)) { err = "invalid count field: " + e; return false; }
    d.count = (int32_t)v;
    if (!parse_int(it[2], 32, "ParseInt", v, e)) { err = "invalid memory field: " + e; return false; }
    d.devmem = (int32_t)v;
    if (!parse_int(it[3], 32, "ParseInt", v, e)) { err = "invalid core field: " + e; return false; }
    d.devcore = (int32_t)v;
    if (!parse_bool(it[6], hb, e)) { err = "invalid health field: " + e; return false; }
    d.health = hb;
    if (!parse_int(it[5], 64, "Atoi", v, e)) { err = "invalid numa field: " + e; return false; }
    if (v != (int32_t)v) { err = "numa field does not fit the engine's int32: " + it[5]; return false; }   // Go's int is 64-bit; the HBM record is not
    d.numa = (int32_t)v;
    d.mode = "hami-core";
    if (it.size() == 9) {
      if (!parse_int(it[7], 64, "Atoi", v, e)) { err = "invalid index field: " + e; return false; }
      if (v < 0) { err = "index field must not be negative: " + std::to_string(v); return false; }
      if (v > 255) { err = "index field does not fit the engine's 8-bit device index: " + it[7]; return false; }
      d.index = (uint32_t)v;
      d.mode = it[8];
    }
    d.id = it[0];
    d.type = it[4];
    out.push_back(d);
  }
  return true;
}

// ---- device.UnMarshalNodeDevices (devices.go:445-449): json.Unmarshal into []*DeviceInfo --------------
bool decode_node_json(const std::string& s, std::vector<Device>& out, std::string& err) {
  out.clear();
  JParser p(s);
  JV doc;
  if (!p.document(doc)) { err = p.err; return false; }
  if (doc.kind == JV::Null) return true;
  if (doc.kind != JV::Arr) { err = std::string("json: cannot unmarshal ") + kind_name(doc) + " into Go value of type []*device.DeviceInfo"; return false; }
  std::string first_err;   // encoding/json keeps decoding after a type error and reports the first one
  auto bad = [&](const std::string& m) { if (first_err.empty()) first_err = m; };
  for (const JV& el : doc.a) {
    Device d;
    if (el.kind == JV::Null) { err = "null device entry in node annotation"; return false; }   // a nil *DeviceInfo: the reference would dereference it
    if (el.kind != JV::Obj) { bad(std::string("json: cannot unmarshal ") + kind_name(el) + " into Go value of type device.DeviceInfo"); out.push_back(d); continue; }
    for (const auto& kv : el.o) {
      const std::string& k = kv.first;
      const JV& v = kv.second;
      if (v.kind == JV::Null) continue;       // null leaves the field at its zero value
      int64_t n;
      if (key_is(k, "id")) { if (v.kind == JV::Str) d.id = v.s; else bad(type_err(v, "id", "string")); }
      else if (key_is(k, "type")) { if (v.kind == JV::Str) d.type = v.s; else bad(type_err(v, "type", "string")); }
      else if (key_is(k, "mode")) { if (v.kind == JV::Str) d.mode = v.s; else bad(type_err(v, "mode", "string")); }
      else if (key_is(k, "devicevendor")) { if (v.kind != JV::Str) bad(type_err(v, "devicevendor", "string")); }
      else if (key_is(k, "health")) { if (v.kind == JV::Bool) d.health = v.b; else bad(type_err(v, "health", "bool")); }
      else if (key_is(k, "index")) {
        if (v.kind == JV::Num && num_to_int(v, 64, true, n)) { if (n > 255) bad("index does not fit the engine's 8-bit device index"); d.index = (uint32_t)n; }
        else bad(type_err(v, "index", "uint"));
      }
      else if (key_is(k, "count")) { if (v.kind == JV::Num && num_to_int(v, 32, false, n)) d.count = (int32_t)n; else bad(type_err(v, "count", "int32")); }
      else if (key_is(k, "devmem")) { if (v.kind == JV::Num && num_to_int(v, 32, false, n)) d.devmem = (int32_t)n; else bad(type_err(v, "devmem", "int32")); }
      else if (key_is(k, "devcore")) { if (v.kind == JV::Num && num_to_int(v, 32, false, n)) d.devcore = (int32_t)n; else bad(type_err(v, "devcore", "int32")); }
      else if (key_is(k, "numa")) {
        if (v.kind == JV::Num && num_to_int(v, 64, false, n)) { if (n != (int32_t)n) bad("numa does not fit the engine's int32"); d.numa = (int32_t)n; }
        else bad(type_err(v, "numa", "int"));
      }
      else if (key_is(k, "migProfiles")) {
        if (v.kind == JV::Null) { d.has_mig_profiles = false; d.mig_profiles.clear(); continue; }
        if (v.kind != JV::Arr) { bad(type_err(v, "migProfiles", "[]device.MigProfile")); continue; }
        d.has_mig_profiles = !v.a.empty();
        d.mig_profiles.clear();
        for (const JV& pj : v.a) {      // device.MigProfile, devices.go:79-90
          MigProfileRec pr;
          if (pj.kind != JV::Obj) { if (pj.kind != JV::Null) bad(type_err(pj, "migProfiles", "device.MigProfile")); d.mig_profiles.push_back(pr); continue; }
          for (const auto& f : pj.o) {
            int64_t x = 0;
            if (key_is(f.first, "name")) { if (f.second.kind == JV::Str) pr.name = f.second.s; else if (f.second.kind != JV::Null) bad(type_err(f.second, "migProfiles.name", "string")); }
            else if (key_is(f.first, "memoryMB")) { if (f.second.kind == JV::Num && num_to_int(f.second, 32, false, x)) pr.memory_mb = (int32_t)x; else if (f.second.kind != JV::Null) bad(type_err(f.second, "migProfiles.memoryMB", "int32")); }
            else if (key_is(f.first, "core")) { if (f.second.kind == JV::Num && num_to_int(f.second, 32, false, x)) pr.core = (int32_t)x; else if (f.second.kind != JV::Null) bad(type_err(f.second, "migProfiles.core", "int32")); }
            else if (key_is(f.first, "sliceCount")) { if (f.second.kind == JV::Num && num_to_int(f.second, 32, true, x)) pr.slice_count = (uint32_t)x; else if (f.second.kind != JV::Null) bad(type_err(f.second, "migProfiles.sliceCount", "uint32")); }
            else if (key_is(f.first, "placements")) {
              if (f.second.kind == JV::Null) { pr.placements.clear(); continue; }
              if (f.second.kind != JV::Arr) { bad(type_err(f.second, "migProfiles.placements", "[]device.MigPlacement")); continue; }
              pr.placements.clear();
              for (const JV& qj : f.second.a) {
                MigPlacementRec q;
                if (qj.kind == JV::Obj)
                  for (const auto& g : qj.o) {
                    if (key_is(g.first, "start")) { if (g.second.kind == JV::Num && num_to_int(g.second, 32, true, x)) q.start = (uint32_t)x; else if (g.second.kind != JV::Null) bad(type_err(g.second, "migProfiles.placements.start", "uint32")); }
                    else if (key_is(g.first, "size")) { if (g.second.kind == JV::Num && num_to_int(g.second, 32, true, x)) q.size = (uint32_t)x; else if (g.second.kind != JV::Null) bad(type_err(g.second, "migProfiles.placements.size", "uint32")); }
                  }
                else if (qj.kind != JV::Null) bad(type_err(qj, "migProfiles.placements", "device.MigPlacement"));
                pr.placements.push_back(q);
              }
            }
          }
          d.mig_profiles.push_back(pr);
        }
      }
      else if (key_is(k, "migtemplate")) { if (v.kind != JV::Arr) bad(type_err(v, "migtemplate", "[]device.Geometry")); }
      else if (key_is(k, "custominfo")) {
        if (v.kind != JV::Obj) { bad(type_err(v, "custominfo", "map[string]interface {}")); continue; }
        // map[string]any: numbers arrive as float64; Ascend reads CustomInfo["NetworkID"] (ascend/device.go:340-352)
        d.die_strategy = false;      // a later duplicate "custominfo" replaces the earlier map
        d.has_aic = false; d.aic.clear();
        for (const auto& ci : v.o) {
          if (ci.first == "AIC") { d.has_aic = ci.second.kind == JV::Str; d.aic = d.has_aic ? ci.second.s : std::string(); }
          if (ci.first == "NetworkID" && ci.second.kind == JV::Num) d.network_id = (int32_t)std::strtod(ci.second.s.c_str(), nullptr);
          // vastai's isDieMode (vastai/device.go:389-403)
          if (ci.first == "DeviceStrategy") d.die_strategy = ci.second.kind == JV::Str && ci.second.s == "die";
        }
      }
      // unknown keys are ignored
    }
    out.push_back(d);
  }
  if (!first_err.empty()) { err = first_err; out.clear(); return false; }
  return true;
}

// ---- json.Unmarshal into map[int]int (parseMetaxAnnos, metax/device.go:168-183) --------------------------
// false on ANY error, as the caller there turns every error into 0: a syntax error, a document that is not an object
// (null is fine and leaves the map empty), a key strconv.ParseInt rejects, a value that is not an integer number
// (null is fine and stores 0).  A later duplicate key wins.
// metax sGPU: json.Unmarshal into []*MetaxSDeviceInfo, then convertMetaxSDeviceToHAMIDevice (protocol.go:96-116)
bool metax_sgpu_devices(const std::string& json, std::vector<Device>& out, std::string& err) {
  out.clear();
  JParser p(json);
  JV doc;
  if (!p.document(doc)) { err = p.err; return false; }
  if (doc.kind != JV::Arr && doc.kind != JV::Null) { err = std::string("json: cannot unmarshal ") + kind_name(doc) + " into Go value of type []*metax.MetaxSDeviceInfo"; return false; }
  size_t idx = 0;
  for (const JV& el : doc.a) {
    Device d;
    if (el.kind == JV::Null) { err = "null device entry in node annotation"; return false; }
    if (el.kind != JV::Obj) { err = std::string("json: cannot unmarshal ") + kind_name(el) + " into Go value of type metax.MetaxSDeviceInfo"; return false; }
    d.index = (uint32_t)idx;
    d.type = "Metax-SGPU";
    d.has_sgpu = true;
    for (const auto& f : el.o) {
      const JV& v = f.second;
      int64_t x = 0;
      auto i32 = [&](const char* field, int32_t& dst) {
        if (v.kind == JV::Null) return true;
        if (v.kind == JV::Num && num_to_int(v, 32, false, x)) { dst = (int32_t)x; return true; }
        err = std::string("json: cannot unmarshal ") + kind_name(v) + " into Go struct field MetaxSDeviceInfo." + field + " of type int32";
        return false;
      };
      if (key_is(f.first, "uuid")) { if (v.kind == JV::Str) d.id = v.s; else if (v.kind != JV::Null) { err = "json: cannot unmarshal into Go struct field MetaxSDeviceInfo.uuid of type string"; return false; } }
      else if (key_is(f.first, "totalDevCount")) { if (!i32("totalDevCount", d.count)) return false; }
      else if (key_is(f.first, "totalCompute")) { if (!i32("totalCompute", d.devcore)) return false; }
      else if (key_is(f.first, "totalVRam")) { if (!i32("totalVRam", d.devmem)) return false; }
      else if (key_is(f.first, "numa")) { if (!i32("numa", d.numa)) return false; }
      else if (key_is(f.first, "linkZone")) { if (!i32("linkZone", d.sgpu_link_zone)) return false; }
      else if (key_is(f.first, "healthy")) { if (v.kind == JV::Bool) d.health = v.b; else if (v.kind != JV::Null) { err = "json: cannot unmarshal into Go struct field MetaxSDeviceInfo.healthy of type bool"; return false; } }
      else if (key_is(f.first, "qosPolicy")) { if (v.kind == JV::Str) d.sgpu_qos = v.s; else if (v.kind != JV::Null) { err = "json: cannot unmarshal into Go struct field MetaxSDeviceInfo.qosPolicy of type string"; return false; } }
    }
    out.push_back(d);
    idx++;
  }
  if (out.empty()) { err = "no sdevice found on node"; return false; }
  return true;
}

// enflame GetNodeDevices, enflame/device.go:198-252 (parseProfile :621-643, parseDRSCapacity :659-683)
static void enf_profile_parts(const std::string& name, int64_t& size, int64_t& mem_gb) {
  size = 0; mem_gb = 0;
  std::string n;
  {   // strings.ToLower(strings.TrimSpace(name))
    size_t a = 0, b = name.size();
    while (a < b && (name[a] == ' ' || (name[a] >= '\t' && name[a] <= '\r'))) a++;
    while (b > a && (name[b - 1] == ' ' || (name[b - 1] >= '\t' && name[b - 1] <= '\r'))) b--;
    n = name.substr(a, b - a);
    for (auto& c : n) if (c >= 'A' && c <= 'Z') c = (char)(c + 32);
  }
  const size_t dot = n.find('.');
  if (dot == std::string::npos) return;
  const size_t dot2 = n.find('.', dot + 1);
  auto trim = [](std::string t) {
    size_t a = 0, b = t.size();
    while (a < b && (t[a] == ' ' || (t[a] >= '\t' && t[a] <= '\r'))) a++;
    while (b > a && (t[b - 1] == ' ' || (t[b - 1] >= '\t' && t[b - 1] <= '\r'))) b--;
    return t.substr(a, b - a);
  };
  std::string sp = trim(n.substr(0, dot));
  if (!sp.empty() && sp.back() == 'g') sp.pop_back();
  std::string e;
  int64_t v;
  if (!parse_int(sp, 64, "Atoi", v, e)) return;
  size = v;
  std::string mp = trim(dot2 == std::string::npos ? n.substr(dot + 1) : n.substr(dot + 1, dot2 - dot - 1));
  if (mp.size() >= 2 && mp.compare(mp.size() - 2, 2, "gb") == 0) mp.erase(mp.size() - 2);
  if (!parse_int(mp, 64, "Atoi", v, e)) return;
  mem_gb = v;
}
bool enflame_devices_from_drs(const std::string& json, const std::string& node, std::vector<Device>& out, std::string& err) {
  out.clear();
  JParser p(json);
  JV doc;
  if (!p.document(doc)) { err = "failed to parse enflame.com/gcu-drs-capacity: " + p.err; return false; }
  std::vector<std::pair<std::string, std::string>> profiles;
  const JV* devs = nullptr;
  if (doc.kind == JV::Obj)
    for (const auto& kv : doc.o) {
      if (key_is(kv.first, "devices")) {
        if (kv.second.kind == JV::Arr) devs = &kv.second;
        else if (kv.second.kind != JV::Null) { err = "failed to parse enflame.com/gcu-drs-capacity: json: cannot unmarshal " + std::string(kind_name(kv.second)) + " into Go struct field drsCapacitySpec.devices of type []enflame.drsDeviceSpec"; return false; }
        else devs = nullptr;
      } else if (key_is(kv.first, "profiles")) {
        profiles.clear();
        if (kv.second.kind == JV::Obj) {
          for (const auto& pr : kv.second.o) {
            if (pr.second.kind != JV::Str && pr.second.kind != JV::Null) { err = "failed to parse enflame.com/gcu-drs-capacity: json: cannot unmarshal " + std::string(kind_name(pr.second)) + " into Go struct field drsCapacitySpec.profiles of type string"; return false; }
            bool found = false;
            for (auto& q : profiles) if (q.first == pr.first) { q.second = pr.second.s; found = true; }
            if (!found) profiles.emplace_back(pr.first, pr.second.kind == JV::Str ? pr.second.s : std::string());
          }
        } else if (kv.second.kind != JV::Null) { err = "failed to parse enflame.com/gcu-drs-capacity: json: cannot unmarshal " + std::string(kind_name(kv.second)) + " into Go struct field drsCapacitySpec.profiles of type map[string]string"; return false; }
      }
    }
  else if (doc.kind != JV::Null) { err = "failed to parse enflame.com/gcu-drs-capacity: json: cannot unmarshal " + std::string(kind_name(doc)) + " into Go value of type enflame.drsCapacitySpec"; return false; }
  int64_t maxSlice = 0, maxMemGB = 0;
  for (const auto& pr : profiles) {
    int64_t s, m;
    enf_profile_parts(pr.first, s, m);
    if (s > maxSlice) maxSlice = s;
    if (m > maxMemGB) maxMemGB = m;
  }
  if (maxSlice <= 0) { err = "no valid drs profiles found on node " + node; return false; }
  if (maxMemGB <= 0) maxMemGB = maxSlice;
  if (maxSlice > 2147483647LL) { err = "drs profile slice count outside int32"; return false; }
  size_t idx = 0;
  if (devs)
    for (const JV& d : devs->a) {
      std::string index, minor;
      const JV* cap = nullptr;
      if (d.kind == JV::Obj)
        for (const auto& f : d.o) {
          if (key_is(f.first, "index")) { if (f.second.kind == JV::Str) index = f.second.s; else if (f.second.kind != JV::Null) { err = "failed to parse enflame.com/gcu-drs-capacity: json: cannot unmarshal " + std::string(kind_name(f.second)) + " into Go struct field drsDeviceSpec.devices.index of type string"; return false; } }
          else if (key_is(f.first, "minor")) { if (f.second.kind == JV::Str) minor = f.second.s; else if (f.second.kind != JV::Null) { err = "failed to parse enflame.com/gcu-drs-capacity: json: cannot unmarshal " + std::string(kind_name(f.second)) + " into Go struct field drsDeviceSpec.devices.minor of type string"; return false; } }
          else if (key_is(f.first, "capacity")) cap = &f.second;
        }
      else if (d.kind != JV::Null) { err = "failed to parse enflame.com/gcu-drs-capacity: json: cannot unmarshal " + std::string(kind_name(d)) + " into Go value of type enflame.drsDeviceSpec"; return false; }
      int64_t devIndex = (int64_t)idx;
      {
        std::string e; int64_t v;
        if (parse_int(index, 64, "Atoi", v, e)) devIndex = v;
      }
      int64_t capacity = 0;
      bool cap_ok = false;
      if (cap && cap->kind == JV::Num) {                      // float64: int32(typed) after the range check
        char* endp = nullptr;
        const double x = strtod(cap->s.c_str(), &endp);
        if (x == x && x <= 2147483647.0 && x >= -2147483648.0) { capacity = (int64_t)x; cap_ok = true; }
      } else if (cap && cap->kind == JV::Str) {
        std::string t = cap->s, e;
        size_t a = 0, b = t.size();
        while (a < b && (t[a] == ' ' || (t[a] >= '\t' && t[a] <= '\r'))) a++;
        while (b > a && (t[b - 1] == ' ' || (t[b - 1] >= '\t' && t[b - 1] <= '\r'))) b--;
        int64_t v;
        if (parse_int(t.substr(a, b - a), 64, "Atoi", v, e)) { capacity = v > 2147483647LL ? 2147483647LL : (v < -2147483648LL ? -2147483648LL : v); cap_ok = true; }
      }
      if (!cap_ok || capacity <= 0) { err = "invalid drs capacity on node " + node; out.clear(); return false; }
      if (devIndex < 0 || devIndex > 255) { err = "drs device index does not fit the engine's 8-bit device index"; out.clear(); return false; }
      Device dv;
      dv.index = (uint32_t)devIndex;
      dv.id = node + "-enflame-drs-" + std::to_string(devIndex);
      dv.count = (int32_t)capacity;
      dv.devmem = (int32_t)(uint32_t)((uint64_t)maxMemGB * 1024u);      // int32(maxMemGB * 1024)
      dv.devcore = 100; dv.type = "Enflame"; dv.numa = 0; dv.health = true;
      dv.has_enflame = true;
      {
        size_t a = 0, b = minor.size();
        while (a < b && (minor[a] == ' ' || (minor[a] >= '\t' && minor[a] <= '\r'))) a++;
        while (b > a && (minor[b - 1] == ' ' || (minor[b - 1] >= '\t' && minor[b - 1] <= '\r'))) b--;
        dv.enf_minor = minor.substr(a, b - a);
      }
      if (dv.enf_minor.empty()) dv.enf_minor = std::to_string(devIndex);
      dv.enf_index = std::to_string(devIndex);
      dv.enf_profiles = profiles;
      dv.enf_max_slice = (int32_t)maxSlice;
      out.push_back(dv);
      idx++;
    }
  if (out.empty()) { err = "no drs devices found on node " + node; return false; }
  return true;
}

// nvidia.DecodeMigAllocations, nvidia/mig_allocations.go:78-112
bool decode_mig_allocations(const std::string& json, std::vector<MigReservationRec>& out, std::string& err) {
  out.clear();
  if (json.empty()) return true;
  JParser p(json);
  JV doc;
  if (!p.document(doc)) { err = p.err; return false; }
  if (doc.kind == JV::Null) return true;
  if (doc.kind != JV::Arr) { err = std::string("json: cannot unmarshal ") + kind_name(doc) + " into Go value of type []nvidia.MigAllocation"; return false; }
  std::vector<std::pair<int64_t, int64_t>> seen;
  size_t i = 0;
  for (const JV& el : doc.a) {
    MigReservationRec r;
    int64_t ci = 0, di = 0, x = 0;
    bool has_uuid = false, has_gi = false, has_ci = false;
    if (el.kind == JV::Obj)
      for (const auto& f : el.o) {
        const JV& v = f.second;
        auto fail = [&](const char* field, const char* gotype) {
          err = std::string("json: cannot unmarshal ") + kind_name(v) + " into Go struct field MigAllocation." + field + " of type " + gotype;
        };
        if (key_is(f.first, "containerIndex")) { if (v.kind == JV::Num && num_to_int(v, 64, false, x)) ci = x; else if (v.kind != JV::Null) { fail("containerIndex", "int"); return false; } }
        else if (key_is(f.first, "deviceIndex")) { if (v.kind == JV::Num && num_to_int(v, 64, false, x)) di = x; else if (v.kind != JV::Null) { fail("deviceIndex", "int"); return false; } }
        else if (key_is(f.first, "gpuUUID")) { if (v.kind == JV::Str) r.gpu_uuid = v.s; else if (v.kind != JV::Null) { fail("gpuUUID", "string"); return false; } }
        else if (key_is(f.first, "profile")) { if (v.kind == JV::Str) r.profile = v.s; else if (v.kind != JV::Null) { fail("profile", "string"); return false; } }
        else if (key_is(f.first, "migUUID")) { if (v.kind == JV::Str) has_uuid = !v.s.empty(); else if (v.kind != JV::Null) { fail("migUUID", "string"); return false; } }
        else if (key_is(f.first, "gpuInstanceID")) { if (v.kind == JV::Num && num_to_int(v, 32, true, x)) has_gi = true; else if (v.kind == JV::Null) has_gi = false; else { fail("gpuInstanceID", "uint32"); return false; } }
        else if (key_is(f.first, "computeInstanceID")) { if (v.kind == JV::Num && num_to_int(v, 32, true, x)) has_ci = true; else if (v.kind == JV::Null) has_ci = false; else { fail("computeInstanceID", "uint32"); return false; } }
        else if (key_is(f.first, "placement")) {
          if (v.kind == JV::Obj) {
            for (const auto& g : v.o) {
              if (key_is(g.first, "start")) { if (g.second.kind == JV::Num && num_to_int(g.second, 32, true, x)) r.start = (uint32_t)x; else if (g.second.kind != JV::Null) { err = "json: cannot unmarshal into Go struct field MigAllocation.placement.start of type uint32"; return false; } }
              else if (key_is(g.first, "size")) { if (g.second.kind == JV::Num && num_to_int(g.second, 32, true, x)) r.size = (uint32_t)x; else if (g.second.kind != JV::Null) { err = "json: cannot unmarshal into Go struct field MigAllocation.placement.size of type uint32"; return false; } }
            }
          } else if (v.kind != JV::Null) { fail("placement", "device.MigPlacement"); return false; }
        }
      }
    else if (el.kind != JV::Null) { err = std::string("json: cannot unmarshal ") + kind_name(el) + " into Go value of type nvidia.MigAllocation"; return false; }
    if (ci < 0 || di < 0 || r.gpu_uuid.empty() || r.profile.empty() || r.size == 0) { err = "MIG allocation " + std::to_string(i) + " is incomplete"; return false; }
    const int runtime = (has_uuid ? 1 : 0) + (has_gi ? 1 : 0) + (has_ci ? 1 : 0);
    if (runtime != 0 && runtime != 3) { err = "MIG allocation " + std::to_string(i) + " has partial runtime identity"; return false; }
    for (auto& k : seen)
      if (k.first == ci && k.second == di) { err = "duplicate MIG allocation for container " + std::to_string(ci) + " device " + std::to_string(di); return false; }
    seen.emplace_back(ci, di);
    out.push_back(r);
    i++;
  }
  return true;
}

bool decode_int_map(const std::string& s, std::vector<std::pair<int64_t, int64_t>>& out) {
  out.clear();
  JParser p(s);
  JV doc;
  if (!p.document(doc)) return false;
  if (doc.kind == JV::Null) return true;
  if (doc.kind != JV::Obj) return false;
  bool ok = true;
  for (auto& kv : doc.o) {
    int64_t k = 0, v = 0;
    std::string e;
    if (!parse_int(kv.first, 64, "ParseInt", k, e)) { ok = false; continue; }
    if (kv.second.kind == JV::Num) { if (!num_to_int(kv.second, 64, false, v)) { ok = false; continue; } }
    else if (kv.second.kind != JV::Null) { ok = false; continue; }
    bool found = false;
    for (auto& o : out) if (o.first == k) { o.second = v; found = true; }
    if (!found) out.emplace_back(k, v);
  }
  if (!ok) out.clear();
  return ok;
}

// ---- AWS Neuron GetNodeDevices, awsneuron/device.go:156-198 ---------------------------------------------------
bool neuron_devices_from_capacity(const std::string& node, int64_t counts, int64_t cores_total, int64_t cores_per, const char* instance_type,
                                  std::vector<Device>& out, std::string& err) {
  out.clear();
  if (counts == 0) { err = "device not found aws.amazon.com/neuron"; return false; }
  uint64_t per = (uint64_t)cores_per;
  if (per == 0) per = (uint64_t)cores_total / (uint64_t)counts;      // uint(coresTotal) / uint(counts)
  uint64_t mask = 0;
  const int64_t bits = (int64_t)(int)per;                              // for i < int(coresPerAWSNeuron); all ones after 64 rounds
  for (int64_t i = 0; i < (bits > 64 ? 64 : bits); i++) mask = mask * 2 + 1;
  const std::string type = instance_type ? instance_type : "";          // a missing label reads "" from the map
  const bool ring = type.find("inf") != std::string::npos || type.find("Inf") != std::string::npos;
  for (int64_t i = 0; i < counts; i++) {
    Device d;
    d.index = (uint32_t)i;
    d.id = node + "-AWSNeuron-" + std::to_string(i);
    d.count = (int32_t)per; d.devmem = 0; d.devcore = (int32_t)mask; d.type = "AWSNeuron"; d.numa = 0; d.health = true;
    d.network_id = ring ? 1 : 0;
    out.push_back(d);
    if (out.size() > 4096) { err = "capacity yields more than 4096 cards"; out.clear(); return false; }
  }
  return true;
}

// ---- vendors without a register annotation: devices derived from the node's capacity -----------------------
bool devices_from_capacity(int kind, const std::string& node, int64_t cores, int64_t memory, std::vector<Device>& out, std::string& err) {
  out.clear();
  if (kind == 3) {                                         // kunlun/device.go:64-91: `cores` carries the XPU count; the upper half sits on NUMA 1
    if (cores == 0) { err = "device not found kunlunxin.com/xpu"; return false; }
    for (int64_t i = 0; i < cores; i++) {
      Device d;
      d.index = (uint32_t)i;
      d.id = node + "-kunlun-" + std::to_string(i);
      d.count = 100; d.devmem = 98304; d.devcore = 100; d.type = "kunlun"; d.health = true;
      d.numa = i >= cores / 2 ? 1 : 0;
      out.push_back(d);
      if (out.size() > 4096) { err = "capacity yields more than 4096 cards"; out.clear(); return false; }
    }
    return true;
  }
  if (kind == 4) {                                         // enflame/gcu.go:67-89: `cores` carries the GCU count; whole cards, 100 / 100
    if (cores == 0) { err = "device not found enflame.com/gcu"; return false; }
    for (int64_t i = 0; i < cores; i++) {
      Device d;
      d.index = (uint32_t)i;
      d.id = node + "-GCU-" + std::to_string(i);
      d.count = 1; d.devmem = 100; d.devcore = 100; d.type = "GCU"; d.numa = 0; d.health = true;
      out.push_back(d);
      if (out.size() > 4096) { err = "capacity yields more than 4096 cards"; out.clear(); return false; }
    }
    return true;
  }
  if (kind == 2) {                                         // metax/device.go:67-89: `cores` carries the GPU count
    if (cores == 0) { err = "device not found metax-tech.com/gpu"; return false; }
    for (int64_t i = 0; i < cores; i++) {
      Device d;
      d.index = (uint32_t)i;
      d.id = node + "-metax-" + std::to_string(i);
      d.count = 1; d.devmem = 65536; d.devcore = 100; d.type = "Metax-GPU"; d.numa = 0; d.health = true;
      out.push_back(d);
      if (out.size() > 4096) { err = "capacity yields more than 4096 cards"; out.clear(); return false; }
    }
    return true;
  }
  const bool mlu = kind == 0;
  if (cores == 0) { err = std::string("device not found ") + (mlu ? "cambricon.com/mlu.smlu.vcore" : "mthreads.com/sgpu-core"); return false; }
  const int64_t per_card = mlu ? 100 : 16;                 // cores of one card
  const int64_t factor = mlu ? 256 : 512;                  // MemoryFactor (cambricon/device.go:60, mthreads/device.go:57)
  for (int64_t i = 0; i * per_card < cores; i++) {
    Device d;
    d.index = (uint32_t)i;
    d.id = node + (mlu ? "-cambricon-mlu-" : "-mthreads-") + std::to_string(i);
    d.count = 100;
    d.devmem = (int32_t)((int64_t)((uint64_t)memory * (uint64_t)factor * (uint64_t)per_card) / cores);   // Go: wrapping int64, then int32()
    d.devcore = (int32_t)per_card;
    d.type = mlu ? "MLU" : "Mthreads";
    d.numa = 0;
    d.health = true;
    out.push_back(d);
    if (out.size() > 4096) { err = "capacity yields more than 4096 cards"; out.clear(); return false; }
  }
  return true;
}

// ---- device.DecodePairScores (devices.go:383-389) + the fit loop of GetNodeDevices (nvidia/device.go:319-340)
// DevicePairScores = []DevicePairScore{ID `json:"uuid"`, Scores map[string]int `json:"score"`}; the FIRST record whose
// uuid equals a device's ID is attached to it; a device without a record keeps an empty score map (all 0).
bool decode_pair_scores(const std::string& json, const std::vector<Device>& devs, std::vector<int32_t>& scores, std::string& err) {
  const size_t n = devs.size();
  scores.assign(n * n, 0);
  JParser p(json);
  JV doc;
  if (!p.document(doc)) { err = p.err; return false; }
  if (doc.kind == JV::Null) return true;
  if (doc.kind != JV::Arr) { err = std::string("json: cannot unmarshal ") + kind_name(doc) + " into Go value of type device.DevicePairScores"; return false; }
  struct Rec { std::string uuid; std::map<std::string, int64_t> sc; };
  std::vector<Rec> recs;
  for (const JV& el : doc.a) {
    Rec r;
    if (el.kind == JV::Null) { recs.push_back(r); continue; }
    if (el.kind != JV::Obj) { err = std::string("json: cannot unmarshal ") + kind_name(el) + " into Go value of type device.DevicePairScore"; return false; }
    for (const auto& kv : el.o) {
      if (kv.second.kind == JV::Null) continue;
      if (key_is(kv.first, "uuid")) {
        if (kv.second.kind != JV::Str) { err = std::string("json: cannot unmarshal ") + kind_name(kv.second) + " into Go struct field DevicePairScore.uuid of type string"; return false; }
        r.uuid = kv.second.s;
      } else if (key_is(kv.first, "score")) {
        if (kv.second.kind != JV::Obj) { err = std::string("json: cannot unmarshal ") + kind_name(kv.second) + " into Go struct field DevicePairScore.score of type map[string]int"; return false; }
        for (const auto& m : kv.second.o) {
          int64_t v = 0;
          if (m.second.kind == JV::Null) { r.sc[m.first] = 0; continue; }
          if (m.second.kind != JV::Num || !num_to_int(m.second, 64, false, v)) {
            err = std::string("json: cannot unmarshal ") + kind_name(m.second) + " into Go struct field DevicePairScore.score of type int";
            return false;
          }
          r.sc[m.first] = v;
        }
      }
    }
    recs.push_back(r);
  }
  for (size_t i = 0; i < n; i++)
    for (const Rec& r : recs) {
      if (r.uuid != devs[i].id) continue;
      for (size_t j = 0; j < n; j++) {
        auto it = r.sc.find(devs[j].id);
        if (it != r.sc.end()) scores[i * n + j] = (int32_t)it->second;
      }
      break;
    }
  return true;
}

// ---- device.DecodePodDevices / DecodeContainerDevices (devices.go:496-556) -------------------------------
bool decode_pod_devices(const std::vector<std::pair<std::string, std::string>>& checklist,
                        const std::vector<std::pair<std::string, std::string>>& annos, std::vector<PodEntry>& out, std::string& err) {
  out.clear();
  if (annos.empty()) return true;
  std::map<std::string, std::string> keys(checklist.begin(), checklist.end());   // the result is a map: key order is not observable, sort it
  for (const auto& ck : keys) {
    const std::string* str = nullptr;
    for (const auto& a : annos) if (a.first == ck.second) str = &a.second;       // a later duplicate annotation wins, like a map
    if (!str) continue;
    uint32_t cidx = 0;
    for (const std::string& seg : split(*str, ';')) {    // every segment is a container, empty ones included (index mapping)
      size_t before = out.size();
      if (!seg.empty())
        for (const std::string& val : split(seg, ':')) {
          if (val.find(',') == std::string::npos) continue;
          const std::vector<std::string> f = split(val, ',');
          if (f.size() < 4) { err = "pod annotation format error, missing fields, do not use nodeName in task spec"; out.clear(); return false; }
          PodEntry e;
          e.key = ck.first; e.container = cidx; e.uuid = f[0]; e.type = f[1];
          int64_t v;
          std::string pe;
          if (!parse_int(f[2], 32, "ParseInt", v, pe)) { err = "invalid memory field: " + pe; out.clear(); return false; }
          e.usedmem = (int32_t)v;
          if (!parse_int(f[3], 32, "ParseInt", v, pe)) { err = "invalid core field: " + pe; out.clear(); return false; }
          e.usedcores = (int32_t)v;
          out.push_back(e);
        }
      if (out.size() == before) {
        PodEntry e;
        e.key = ck.first; e.container = cidx; e.empty = true;
        out.push_back(e);
      }
      cidx++;
    }
  }
  return true;
}

// ---- device.CollapseInitContainerUsage / AppContainersOnlyDeviceUsage (initContainer.go:36-157) --------
void collapse(int mode, uint32_t n_init, const std::vector<PodEntry>& raw, std::vector<PodEntry>& out) {
  out.clear();
  struct U { int32_t mem = 0, cores = 0, slots = 0; };
  std::map<std::string, std::map<std::string, std::pair<U, U>>> acc;   // key → uuid → (init peak, app sum); std::map = sorted by UUID
  for (const PodEntry& e : raw) {
    auto& per_key = acc[e.key];                    // a key with only empty containers still yields an (empty) collapsed list
    if (e.empty) continue;
    if (e.container < n_init) {
      if (mode == 1) continue;                     // app-only: init containers are skipped
      U& u = per_key[e.uuid].first;
      u.mem = std::max(u.mem, e.usedmem);
      u.cores = std::max(u.cores, e.usedcores);
      u.slots = 1;                                 // init containers run one at a time
    } else {
      U& u = per_key[e.uuid].second;
      u.mem += e.usedmem; u.cores += e.usedcores; u.slots++;
    }
  }
  for (const auto& k : acc) {
    size_t before = out.size();
    for (const auto& d : k.second) {
      const U& i = d.second.first;
      const U& a = d.second.second;
      PodEntry e;
      e.key = k.first; e.type = k.first; e.uuid = d.first; e.container = 0;
      if (mode == 1) { e.usedmem = a.mem; e.usedcores = a.cores; e.slots = a.slots; }
      else { e.usedmem = std::max(i.mem, a.mem); e.usedcores = std::max(i.cores, a.cores); e.slots = std::max(std::max(i.slots, a.slots), 1); }
      out.push_back(e);
    }
    if (out.size() == before) {
      PodEntry e;
      e.key = k.first; e.container = 0; e.empty = true;
      out.push_back(e);
    }
  }
}

}  // namespace hs_ingest
