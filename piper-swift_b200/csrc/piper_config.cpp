#include "piper_config.h"

#include <cstdlib>
#include <fstream>
#include <sstream>
#include <stdexcept>

namespace piper {
namespace {

struct P {
  const std::string& s;
  size_t i = 0;
  [[noreturn]] void fail(const char* what) {
    throw std::runtime_error(std::string("json: ") + what + " at offset " + std::to_string(i));
  }
  void ws() {
    while (i < s.size() && (s[i] == ' ' || s[i] == '\n' || s[i] == '\t' || s[i] == '\r')) ++i;
  }
  bool eat(char c) {
    ws();
    if (i < s.size() && s[i] == c) {
      ++i;
      return true;
    }
    return false;
  }
  static void utf8(std::string& out, uint32_t cp) {
    if (cp < 0x80) out += char(cp);
    else if (cp < 0x800) { out += char(0xC0 | (cp >> 6)); out += char(0x80 | (cp & 0x3F)); }
    else if (cp < 0x10000) { out += char(0xE0 | (cp >> 12)); out += char(0x80 | ((cp >> 6) & 0x3F)); out += char(0x80 | (cp & 0x3F)); }
    else { out += char(0xF0 | (cp >> 18)); out += char(0x80 | ((cp >> 12) & 0x3F)); out += char(0x80 | ((cp >> 6) & 0x3F)); out += char(0x80 | (cp & 0x3F)); }
  }
  uint32_t hex4() {
    if (i + 4 > s.size()) fail("truncated \\u escape");
    uint32_t v = 0;
    for (int k = 0; k < 4; ++k) {
      char c = s[i++];
      v <<= 4;
      if (c >= '0' && c <= '9') v |= c - '0';
      else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
      else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
      else fail("bad hex digit");
    }
    return v;
  }
  std::string string() {
    ws();
    if (i >= s.size() || s[i] != '"') fail("expected string");
    ++i;
    std::string out;
    while (true) {
      if (i >= s.size()) fail("unterminated string");
      char c = s[i++];
      if (c == '"') break;
      if (c != '\\') {
        out += c;
        continue;
      }
      if (i >= s.size()) fail("unterminated escape");
      char e = s[i++];
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
          uint32_t cp = hex4();
          if (cp >= 0xD800 && cp < 0xDC00 && i + 1 < s.size() && s[i] == '\\' && s[i + 1] == 'u') {
            i += 2;
            uint32_t lo = hex4();
            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
          }
          utf8(out, cp);
          break;
        }
        default: fail("bad escape");
      }
    }
    return out;
  }
  JsonValue value() {
    ws();
    if (i >= s.size()) fail("unexpected end");
    JsonValue v;
    char c = s[i];
    if (c == '{') {
      ++i;
      v.kind = JsonValue::Object;
      if (eat('}')) return v;
      do {
        std::string k = string();
        if (!eat(':')) fail("expected ':'");
        v.obj.emplace_back(std::move(k), value());
      } while (eat(','));
      if (!eat('}')) fail("expected '}'");
    } else if (c == '[') {
      ++i;
      v.kind = JsonValue::Array;
      if (eat(']')) return v;
      do v.arr.push_back(value());
      while (eat(','));
      if (!eat(']')) fail("expected ']'");
    } else if (c == '"') {
      v.kind = JsonValue::String;
      v.str = string();
    } else if (s.compare(i, 4, "true") == 0) {
      v.kind = JsonValue::Bool; v.b = true; i += 4;
    } else if (s.compare(i, 5, "false") == 0) {
      v.kind = JsonValue::Bool; i += 5;
    } else if (s.compare(i, 4, "null") == 0) {
      i += 4;
    } else {
      const char* b = s.c_str() + i;
      char* e = nullptr;
      v.num = std::strtod(b, &e);
      if (e == b) fail("unexpected token");
      v.kind = JsonValue::Number;
      i += size_t(e - b);
    }
    return v;
  }
};

const JsonValue& need(const JsonValue& o, const char* key, JsonValue::Kind kind) {
  const JsonValue* v = o.get(key);
  if (!v || v->kind != kind) throw std::runtime_error(std::string("config: missing or mistyped key '") + key + "'");
  return *v;
}

}  // namespace

JsonValue parse_json(const std::string& text) {
  P p{text};
  JsonValue v = p.value();
  p.ws();
  if (p.i != text.size()) p.fail("trailing characters");
  return v;
}

PiperConfig PiperConfig::from_json(const std::string& text) {
  JsonValue root = parse_json(text);
  if (root.kind != JsonValue::Object) throw std::runtime_error("config: top level is not an object");
  PiperConfig c;
  const JsonValue& audio = need(root, "audio", JsonValue::Object);
  c.sample_rate = int(need(audio, "sample_rate", JsonValue::Number).num);
  if (auto q = audio.get("quality"); q && q->kind == JsonValue::String) c.quality = q->str;
  if (auto e = root.get("espeak"); e && e->kind == JsonValue::Object)
    if (auto v = e->get("voice"); v && v->kind == JsonValue::String) c.espeak_voice = v->str;
  const JsonValue& inf = need(root, "inference", JsonValue::Object);
  c.noise_scale = float(need(inf, "noise_scale", JsonValue::Number).num);
  c.length_scale = float(need(inf, "length_scale", JsonValue::Number).num);
  c.noise_w = float(need(inf, "noise_w", JsonValue::Number).num);
  c.phoneme_type = need(root, "phoneme_type", JsonValue::String).str;
  for (auto& kv : need(root, "phoneme_id_map", JsonValue::Object).obj) {
    std::vector<int64_t> ids;
    for (auto& x : kv.second.arr) ids.push_back(int64_t(x.num));
    c.phoneme_id_map[kv.first] = std::move(ids);
  }
  c.num_symbols = int(need(root, "num_symbols", JsonValue::Number).num);
  c.num_speakers = int(need(root, "num_speakers", JsonValue::Number).num);
  if (auto m = root.get("speaker_id_map"); m && m->kind == JsonValue::Object)
    for (auto& kv : m->obj) c.speaker_id_map[kv.first] = int64_t(kv.second.num);
  return c;
}

PiperConfig PiperConfig::load(const std::string& path) {
  std::ifstream f(path);
  if (!f) throw std::runtime_error("io: cannot open '" + path + "'");
  std::stringstream ss;
  ss << f.rdbuf();
  return from_json(ss.str());
}

}  // namespace piper
