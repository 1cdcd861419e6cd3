#include "klm_wire.h"

#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <initializer_list>
#include <deque>
#include <string_view>
#include <vector>
#include <stdexcept>

namespace klmhost {
namespace {

// The document is one flat array of nodes (children linked as first-child / next-sibling, member
// names kept in the member's node) over the caller's text: strings without escapes are views into
// it, the rare escaped ones are decoded into a side pool. No allocation per value, no map per
// object (the objects of this path have a handful of members; a lookup is a short scan, the LAST
// member of a name winning as in a Go map built by encoding/json).
constexpr uint32_t kNil = 0xffffffffu;
struct JValue {
  enum Kind : uint8_t { Null, Bool, Int, Float, String, Array, Object } kind = Null;
  bool b = false;
  uint32_t first = kNil, next = kNil;  // first child, next sibling (indices into the arena)
  double num = 0;                      // Int and Float both carry the value here
  std::string_view sv;                 // String: its (unescaped) text
  std::string_view key;                // member name, when the parent is an object
  const JValue* base = nullptr;        // the arena (set once parsing is over)
  const JValue* child() const { return first == kNil ? nullptr : base + first; }
  const JValue* sibling() const { return next == kNil ? nullptr : base + next; }
  std::string str() const { return std::string(sv); }
  const JValue* get(std::string_view k) const {
    if (kind != Object) return nullptr;
    const JValue* hit = nullptr;
    for (const JValue* c = child(); c; c = c->sibling())
      if (c->key == k) hit = c;
    return hit;
  }
};

class Doc {
 public:
  explicit Doc(const std::string& s) : s_(s) {
    nodes_.reserve(s.size() / 6 + 16);
    const uint32_t root = value();
    ws();
    if (i_ != s_.size()) fail("trailing characters");
    for (JValue& v : nodes_) v.base = nodes_.data();
    root_ = root;
  }
  const JValue& root() const { return nodes_[root_]; }

 private:
  const std::string& s_;
  size_t i_ = 0;
  uint32_t root_ = 0;
  std::vector<JValue> nodes_;
  std::deque<std::string> pool_;  // decoded text of strings that carry escapes (stable addresses)
  [[noreturn]] void fail(const char* what) const {
    throw std::runtime_error("JSON: " + std::string(what) + " at offset " + std::to_string(i_));
  }
  void ws() {
    while (i_ < s_.size() && (s_[i_] == ' ' || s_[i_] == '\n' || s_[i_] == '\t' || s_[i_] == '\r')) ++i_;
  }
  bool eat(const char* lit) {
    size_t n = std::char_traits<char>::length(lit);
    if (s_.compare(i_, n, lit) == 0) { i_ += n; return true; }
    return false;
  }
  static void utf8(std::string& out, unsigned cp) {
    if (cp < 0x80) out += static_cast<char>(cp);
    else if (cp < 0x800) { out += static_cast<char>(0xC0 | (cp >> 6)); out += static_cast<char>(0x80 | (cp & 0x3F)); }
    else if (cp < 0x10000) {
      out += static_cast<char>(0xE0 | (cp >> 12)); out += static_cast<char>(0x80 | ((cp >> 6) & 0x3F));
      out += static_cast<char>(0x80 | (cp & 0x3F));
    } else {
      out += static_cast<char>(0xF0 | (cp >> 18)); out += static_cast<char>(0x80 | ((cp >> 12) & 0x3F));
      out += static_cast<char>(0x80 | ((cp >> 6) & 0x3F)); out += static_cast<char>(0x80 | (cp & 0x3F));
    }
  }
  unsigned hex4() {
    if (i_ + 4 > s_.size()) fail("short \\u escape");
    unsigned v = 0;
    for (int k = 0; k < 4; ++k) {
      char c = s_[i_++];
      v <<= 4;
      if (c >= '0' && c <= '9') v |= c - '0';
      else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
      else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
      else fail("bad \\u escape");
    }
    return v;
  }
  std::string_view string() {
    if (i_ >= s_.size() || s_[i_] != '"') fail("expected string");
    ++i_;
    const size_t st = i_;
    const char* p = s_.data();
    const size_t n = s_.size();
    while (i_ < n && p[i_] != '"' && p[i_] != '\\') ++i_;
    if (i_ < n && p[i_] == '"') {  // the usual case: no escapes, a view into the text
      ++i_;
      return std::string_view(p + st, i_ - 1 - st);
    }
    std::string out(p + st, i_ - st);
    while (i_ < n && p[i_] != '"') {
      char c = p[i_++];
      if (c != '\\') { out += c; continue; }
      if (i_ >= n) fail("short escape");
      char e = p[i_++];
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
          unsigned cp = hex4();
          if (cp >= 0xD800 && cp < 0xDC00 && s_.compare(i_, 2, "\\u") == 0) {
            i_ += 2;
            unsigned lo = hex4();
            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
          }
          utf8(out, cp);
          break;
        }
        default: fail("bad escape");
      }
    }
    if (i_ >= n) fail("unterminated string");
    ++i_;
    pool_.push_back(std::move(out));
    return pool_.back();
  }
  uint32_t make(JValue::Kind k) {
    nodes_.emplace_back();
    nodes_.back().kind = k;
    return static_cast<uint32_t>(nodes_.size() - 1);
  }
  uint32_t value() {
    ws();
    if (i_ >= s_.size()) fail("unexpected end");
    const char c = s_[i_];
    if (c == '{') {
      const uint32_t v = make(JValue::Object);
      ++i_;
      ws();
      if (i_ < s_.size() && s_[i_] == '}') { ++i_; return v; }
      uint32_t last = kNil;
      for (;;) {
        ws();
        const std::string_view k = string();
        ws();
        if (i_ >= s_.size() || s_[i_] != ':') fail("expected ':'");
        ++i_;
        const uint32_t m = value();
        nodes_[m].key = k;
        if (last == kNil) nodes_[v].first = m; else nodes_[last].next = m;
        last = m;
        ws();
        if (i_ < s_.size() && s_[i_] == ',') { ++i_; continue; }
        if (i_ < s_.size() && s_[i_] == '}') { ++i_; return v; }
        fail("expected ',' or '}'");
      }
    }
    if (c == '[') {
      const uint32_t v = make(JValue::Array);
      ++i_;
      ws();
      if (i_ < s_.size() && s_[i_] == ']') { ++i_; return v; }
      uint32_t last = kNil;
      for (;;) {
        const uint32_t m = value();
        if (last == kNil) nodes_[v].first = m; else nodes_[last].next = m;
        last = m;
        ws();
        if (i_ < s_.size() && s_[i_] == ',') { ++i_; continue; }
        if (i_ < s_.size() && s_[i_] == ']') { ++i_; return v; }
        fail("expected ',' or ']'");
      }
    }
    if (c == '"') {
      const std::string_view t = string();
      const uint32_t v = make(JValue::String);
      nodes_[v].sv = t;
      return v;
    }
    if (eat("true")) { const uint32_t v = make(JValue::Bool); nodes_[v].b = true; return v; }
    if (eat("false")) return make(JValue::Bool);
    if (eat("null")) return make(JValue::Null);
    // number: integral text (no '.', 'e', 'E') is an int64 for apimachinery's decoder
    const size_t st = i_;
    bool integral = true;
    if (s_[i_] == '-') ++i_;
    while (i_ < s_.size()) {
      char d = s_[i_];
      if (d >= '0' && d <= '9') { ++i_; continue; }
      if (d == '.' || d == 'e' || d == 'E' || d == '+' || d == '-') { integral = false; ++i_; continue; }
      break;
    }
    if (i_ == st) fail("unexpected character");
    const uint32_t v = make(integral ? JValue::Int : JValue::Float);
    const size_t len = i_ - st;
    double x;
    if (integral && len <= 15) {  // exact in a double: no strtod needed
      int64_t a = 0;
      size_t k = st + (s_[st] == '-' ? 1 : 0);
      for (; k < i_; ++k) a = a * 10 + (s_[k] - '0');
      x = s_[st] == '-' ? -static_cast<double>(a) : static_cast<double>(a);
    } else if (len < 64) {        // the token alone (strtod must not read past it: "0x1p3", "1e5e5")
      char buf[64];
      s_.copy(buf, len, st);
      buf[len] = 0;
      x = std::strtod(buf, nullptr);
    } else {
      x = std::strtod(s_.substr(st, len).c_str(), nullptr);
    }
    nodes_[v].num = x;
    return v;
  }
};

// the elements of a list document: the root array, or the root object's "items"
const JValue* first_item(const JValue& root) {
  if (root.kind == JValue::Array) return root.child();
  const JValue* it = root.get("items");
  if (!it || it->kind != JValue::Array) throw std::runtime_error("JSON: no \"items\" array");
  return it->child();
}

}  // namespace

std::vector<UAVMetricItem> DecodeUAVMetricList(const std::string& json) {
  const Doc doc(json);
  std::vector<UAVMetricItem> out;
  for (const JValue* it = first_item(doc.root()); it; it = it->sibling()) {
    out.emplace_back();
    UAVMetricItem& u = out.back();
    const JValue* spec = it->get("spec");      // NestedMap(item, "spec"), controller.go:182
    const JValue* status = it->get("status");  // :183
    if (spec) {
      if (const JValue* v = spec->get("node_name"); v && v->kind == JValue::String) {  // :185
        u.HasNodeName = true;
        u.NodeName.assign(v->sv);
      }
      if (const JValue* v = spec->get("uav_id"); v && v->kind == JValue::String) u.UAVID.assign(v->sv);  // :186
      if (const JValue* b = spec->get("battery"))  // readFloat(spec, "battery", "remaining_percent"), :187
        if (const JValue* v = b->get("remaining_percent"); v && (v->kind == JValue::Int || v->kind == JValue::Float)) {
          u.HasBattery = true;  // float64 / int64 cases of the type switch, :262-271
          u.Battery = v->num;
        }
    }
    if (status)
      if (const JValue* v = status->get("collection_status"); v && v->kind == JValue::String) {  // :188
        u.HasStatus = true;
        u.CollectionStatus.assign(v->sv);
      }
  }
  return out;
}

std::vector<SchedulingRequest> DecodeSchedulingRequestList(const std::string& json, bool apimachinery_ints) {
  const Doc doc(json);
  std::vector<SchedulingRequest> out;
  for (const JValue* it = first_item(doc.root()); it; it = it->sibling()) {
    out.emplace_back();
    SchedulingRequest& q = out.back();
    if (const JValue* st = it->get("status"))
      if (const JValue* v = st->get("phase"); v && v->kind == JValue::String) {  // NestedString, :113
        q.HasPhase = true;
        q.Phase.assign(v->sv);
      }
    if (const JValue* spec = it->get("spec")) {
      if (const JValue* wl = spec->get("workload"); wl && wl->kind == JValue::Object) {  // :127-131
        if (const JValue* v = wl->get("name"); v && v->kind == JValue::String) q.Spec.Workload.Name.assign(v->sv);
        if (const JValue* v = wl->get("namespace"); v && v->kind == JValue::String) q.Spec.Workload.Namespace.assign(v->sv);
        if (const JValue* v = wl->get("type"); v && v->kind == JValue::String) q.Spec.Workload.Type.assign(v->sv);
      }
      if (const JValue* v = spec->get("minBatteryPercent"))  // .(float64) only, :133
        if (v->kind == JValue::Float || (v->kind == JValue::Int && !apimachinery_ints)) q.Spec.MinBatteryPercent = v->num;
      if (const JValue* v = spec->get("preferredNodes"); v && v->kind == JValue::Array)  // :137-143
        for (const JValue* e = v->child(); e; e = e->sibling())
          if (e->kind == JValue::String) q.Spec.PreferredNodes.emplace_back(e->sv);
    }
  }
  return out;
}

// ---------------------------------------------------------------- UAVReport (A11)

namespace {
[[noreturn]] void bad_type(const char* field, const char* want) {
  throw std::runtime_error(std::string("JSON: cannot unmarshal field ") + field + " (want " + want + ")");
}
std::string str_field(const JValue* o, const char* k) {
  const JValue* v = o ? o->get(k) : nullptr;
  if (!v || v->kind == JValue::Null) return "";
  if (v->kind != JValue::String) bad_type(k, "string");
  return v->str();
}
double num_field(const JValue* o, const char* k) {  // float64 field: any JSON number
  const JValue* v = o ? o->get(k) : nullptr;
  if (!v || v->kind == JValue::Null) return 0.0;
  if (v->kind != JValue::Int && v->kind != JValue::Float) bad_type(k, "number");
  return v->num;
}
double int_field(const JValue* o, const char* k) {  // int field: Go rejects 3.5 and 1e2 alike
  const JValue* v = o ? o->get(k) : nullptr;
  if (!v || v->kind == JValue::Null) return 0.0;
  if (v->kind != JValue::Int) bad_type(k, "integer");
  return v->num;
}
double bool_field(const JValue* o, const char* k) {
  const JValue* v = o ? o->get(k) : nullptr;
  if (!v || v->kind == JValue::Null) return 0.0;
  if (v->kind != JValue::Bool) bad_type(k, "bool");
  return v->b ? 1.0 : 0.0;
}
const JValue* obj_field(const JValue* o, const char* k) {
  const JValue* v = o ? o->get(k) : nullptr;
  if (!v || v->kind == JValue::Null) return nullptr;
  if (v->kind != JValue::Object) bad_type(k, "object");
  return v;
}
double enum_code(const std::string& s, std::initializer_list<const char*> names) {
  int i = 0;
  for (const char* n : names) {
    if (s == n) return i;
    ++i;
  }
  return -1.0;
}
}  // namespace

UAVReport DecodeUAVReport(const std::string& json) {
  const Doc doc(json);
  if (doc.root().kind != JValue::Object) throw std::runtime_error("JSON: a UAVReport is an object");
  const JValue* o = &doc.root();
  UAVReport r;
  r.NodeName = str_field(o, "node_name");     // models.go:183-187
  r.NodeIP = str_field(o, "node_ip");
  r.UAVID = str_field(o, "uav_id");
  r.Source = str_field(o, "source");
  r.Status = str_field(o, "status");
  r.Timestamp = str_field(o, "timestamp");    // :188 (time.Time travels as RFC3339 text)
  if (r.Timestamp == "0001-01-01T00:00:00Z") r.Timestamp.clear();  // Go's zero time
  r.HeartbeatIntervalSeconds = static_cast<int64_t>(int_field(o, "heartbeat_interval_seconds"));  // :189
  const JValue* st = obj_field(o, "state");   // :190 *uav.UAVState
  if (!st) return r;
  r.HasState = true;
  float* c = r.Channels;
  const JValue* gps = obj_field(st, "gps");   // mavlink_simulator.go:39-50
  c[0] = (float)num_field(gps, "latitude"); c[1] = (float)num_field(gps, "longitude");
  c[2] = (float)num_field(gps, "altitude"); c[3] = (float)num_field(gps, "relative_altitude");
  c[4] = (float)num_field(gps, "hdop"); c[5] = (float)int_field(gps, "satellite_count");
  c[6] = (float)int_field(gps, "fix_type"); c[7] = (float)num_field(gps, "ground_speed");
  c[8] = (float)num_field(gps, "course_over_ground");
  const JValue* at = obj_field(st, "attitude");  // :53-61
  c[9] = (float)num_field(at, "roll"); c[10] = (float)num_field(at, "pitch"); c[11] = (float)num_field(at, "yaw");
  c[12] = (float)num_field(at, "roll_rate"); c[13] = (float)num_field(at, "pitch_rate");
  c[14] = (float)num_field(at, "yaw_rate");
  const JValue* fl = obj_field(st, "flight");    // :64-72
  c[15] = (float)num_field(fl, "airspeed"); c[16] = (float)num_field(fl, "ground_speed");
  c[17] = (float)num_field(fl, "vertical_speed"); c[18] = (float)num_field(fl, "throttle_percent");
  const JValue* bt = obj_field(st, "battery");   // :75-85
  r.RemainingPercent = num_field(bt, "remaining_percent");
  c[19] = (float)num_field(bt, "voltage"); c[20] = (float)num_field(bt, "current");
  c[21] = (float)r.RemainingPercent; c[22] = (float)num_field(bt, "remaining_capacity");
  c[23] = (float)num_field(bt, "total_capacity"); c[24] = (float)num_field(bt, "temperature");
  c[25] = (float)int_field(bt, "cell_count"); c[26] = (float)int_field(bt, "time_remaining");
  const JValue* ms = obj_field(st, "mission");   // :88-95
  c[27] = (float)int_field(ms, "current_waypoint"); c[28] = (float)int_field(ms, "total_waypoints");
  c[29] = (float)num_field(ms, "distance_to_wp"); c[30] = (float)int_field(ms, "eta_to_wp");
  const JValue* he = obj_field(st, "health");    // :98-106
  c[31] = (float)int_field(he, "error_count"); c[32] = (float)int_field(he, "warning_count");
  c[33] = (float)bool_field(fl, "armed");
  const JValue* sh = obj_field(he, "sensors_health");
  static const char* kSensors[6] = {"accelerometer", "barometer", "battery", "compass", "gps", "gyroscope"};
  for (int k = 0; k < 6; ++k) c[34 + k] = (float)bool_field(sh, kSensors[k]);
  r.SystemStatus = str_field(he, "system_status");
  c[40] = (float)enum_code(str_field(fl, "mode"), {"MANUAL", "STABILIZE", "LOITER", "AUTO", "RTL", "LAND"});
  c[41] = (float)enum_code(str_field(ms, "mission_state"), {"IDLE", "ACTIVE", "PAUSED", "COMPLETED"});
  c[42] = (float)enum_code(r.SystemStatus, {"OK", "WARNING", "CRITICAL", "ERROR"});
  return r;
}

bool ApplyUAVReportDefaults(UAVReport& r, const std::string& now_rfc3339) {
  if (r.NodeName.empty()) return false;                     // manager.go:392-394
  if (r.Timestamp.empty()) r.Timestamp = now_rfc3339;       // :396-399
  if (r.Status.empty()) r.Status = "active";                // :401-404
  if (r.Source.empty()) r.Source = "agent";                 // :406-409
  return true;
}

static void quote_into(std::string& o, const std::string& s) {
  o += '"';
  for (unsigned char c : s) {
    switch (c) {
      case '"': o += "\\\""; break;
      case '\\': o += "\\\\"; break;
      case '\n': o += "\\n"; break;
      case '\r': o += "\\r"; break;
      case '\t': o += "\\t"; break;
      default:
        if (c < 0x20) { char b[8]; std::snprintf(b, sizeof(b), "\\u%04x", c); o += b; }
        else o += static_cast<char>(c);
    }
  }
  o += '"';
}

// Go's encoding/json float64 encoder (encode.go floatEncoder): the shortest digits that
// round-trip (strconv 'g' precision -1), written in plain decimal ('f') unless the exponent is
// below -6 or at least 21, where it uses 'e' and trims "e-0X" to "e-X". 90 -> "90", 85.5 -> "85.5",
// 1e21 -> "1e+21", 1e-7 -> "1e-7". (Non-finite values are an encoder error in Go; they print as 0.)
std::string FormatJSONFloat(double v) {
  if (!(v == v) || v == HUGE_VAL || v == -HUGE_VAL) return "0";
  if (v == 0.0) return std::signbit(v) ? "-0" : "0";
  char e[48];
  {  // shortest digits that round-trip, as d[.ddd]e[+-]XX
    const std::to_chars_result r = std::to_chars(e, e + sizeof(e) - 1, v, std::chars_format::scientific);
    *r.ptr = 0;
  }
  // e = [-]d[.ddd]e[+-]XX: digits and decimal exponent
  std::string digits;
  const char* p = e;
  const bool neg = *p == '-';
  if (neg) ++p;
  for (; *p && *p != 'e'; ++p)
    if (*p != '.') digits += *p;
  const int x = std::atoi(p + 1);
  while (digits.size() > 1 && digits.back() == '0') digits.pop_back();
  std::string out = neg ? "-" : "";
  if (x < -6 || x >= 21) {
    out += digits[0];
    if (digits.size() > 1) out += "." + digits.substr(1);
    char t[16];
    if (x < 0 && x > -10) std::snprintf(t, sizeof(t), "e-%d", -x);  // json trims e-07 to e-7
    else std::snprintf(t, sizeof(t), "e%c%02d", x < 0 ? '-' : '+', x < 0 ? -x : x);
    return out + t;
  }
  if (x < 0) {
    out += "0." + std::string(static_cast<size_t>(-x - 1), '0') + digits;
  } else if (static_cast<size_t>(x) + 1 >= digits.size()) {
    out += digits + std::string(static_cast<size_t>(x) + 1 - digits.size(), '0');
  } else {
    out += digits.substr(0, static_cast<size_t>(x) + 1) + "." + digits.substr(static_cast<size_t>(x) + 1);
  }
  return out;
}

std::string EncodeStatus(const SchedulingRequestStatus& st, const std::string& last_updated) {
  static const std::string kPending = "Pending";
  std::string o;
  o.reserve(128 + st.AssignedNode.size() + st.AssignedUAV.size() + st.Message.size() + last_updated.size());
  o += "{\"phase\":";
  quote_into(o, st.Phase.empty() ? kPending : st.Phase);  // controller.go:238-240
  o += ",\"assignedNode\":";
  quote_into(o, st.AssignedNode);
  o += ",\"assignedUAV\":";
  quote_into(o, st.AssignedUAV);
  o += ",\"score\":";
  o += FormatJSONFloat(st.Score);
  o += ",\"message\":";
  quote_into(o, st.Message);
  o += ",\"lastUpdated\":";
  quote_into(o, last_updated);
  o += '}';
  return o;
}

}  // namespace klmhost
