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

namespace {
// Unicode general category Cf (format), Unicode 15 ranges
bool is_format_scalar(uint32_t c) {
  static const uint32_t r[][2] = {
      {0x00AD, 0x00AD}, {0x0600, 0x0605}, {0x061C, 0x061C}, {0x06DD, 0x06DD}, {0x070F, 0x070F}, {0x0890, 0x0891},
      {0x08E2, 0x08E2}, {0x180E, 0x180E}, {0x200B, 0x200F}, {0x202A, 0x202E}, {0x2060, 0x2064}, {0x2066, 0x206F},
      {0xFEFF, 0xFEFF}, {0xFFF9, 0xFFFB}, {0x110BD, 0x110BD}, {0x110CD, 0x110CD}, {0x13430, 0x1343F},
      {0x1BCA0, 0x1BCA3}, {0x1D173, 0x1D17A}, {0xE0001, 0xE0001}, {0xE0020, 0xE007F}};
  for (auto& x : r)
    if (c >= x[0] && c <= x[1]) return true;
  return false;
}
}  // namespace

std::vector<int64_t> PiperConfig::ipa_to_ids(const std::string& s) const {
  auto first_id = [&](const char* sym) -> int64_t {
    auto it = phoneme_id_map.find(sym);
    if (it == phoneme_id_map.end() || it->second.empty())
      throw std::runtime_error("invalid argument: ^/$/_ missing from phoneme_id_map");
    return it->second.front();
  };
  const int64_t bos = first_id("^"), eos = first_id("$"), blank = first_id("_");
  std::vector<int64_t> ids;
  ids.reserve(2 * s.size() + 2);
  ids.push_back(bos);
  size_t i = 0;
  while (i < s.size()) {
    const unsigned char b0 = (unsigned char)s[i];
    size_t len = b0 < 0x80 ? 1 : (b0 >> 5) == 0x6 ? 2 : (b0 >> 4) == 0xE ? 3 : (b0 >> 3) == 0x1E ? 4 : 0;
    if (len == 0 || i + len > s.size()) throw std::runtime_error("invalid argument: malformed UTF-8 in the IPA string");
    uint32_t cp = len == 1 ? b0 : b0 & (0xFF >> (len + 1));
    for (size_t k = 1; k < len; ++k) {
      const unsigned char bk = (unsigned char)s[i + k];
      if ((bk & 0xC0) != 0x80) throw std::runtime_error("invalid argument: malformed UTF-8 in the IPA string");
      cp = (cp << 6) | (bk & 0x3F);
    }
    const std::string sym = s.substr(i, len);
    i += len;
    if (cp == '\n' || cp == '\r') continue;
    if (cp == 0x200B || cp == 0x200C || cp == 0x200D || cp == 0xFE0E || cp == 0xFE0F || is_format_scalar(cp)) continue;
    auto it = phoneme_id_map.find(sym);
    if (it == phoneme_id_map.end() || it->second.empty())
      throw std::runtime_error("invalid argument: unknown phoneme symbol '" + sym + "'");
    ids.push_back(it->second.front());
    ids.push_back(blank);
  }
  ids.push_back(eos);
  return ids;
}

}  // namespace piper
