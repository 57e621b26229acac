// C++ twin of PiperConfig (Sources/PiperCore/PiperConfig.swift:3-46): the voice's
// `*.onnx.json`.  Same required keys as the Swift Codable (audio.sample_rate, inference.*,
// phoneme_type, phoneme_id_map, num_symbols, num_speakers); a missing required key is an error.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace piper {

struct JsonValue {
  enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
  bool b = false;
  double num = 0;
  std::string str;
  std::vector<JsonValue> arr;
  std::vector<std::pair<std::string, JsonValue>> obj;
  const JsonValue* get(const std::string& key) const {
    for (auto& kv : obj)
      if (kv.first == key) return &kv.second;
    return nullptr;
  }
};

JsonValue parse_json(const std::string& text);   // throws std::runtime_error

struct PiperConfig {
  int sample_rate = 22050;
  std::string quality;
  std::string espeak_voice;
  float noise_scale = 0.667f, length_scale = 1.0f, noise_w = 0.8f;
  std::string phoneme_type;
  std::map<std::string, std::vector<int64_t>> phoneme_id_map;   // UTF-8 phoneme -> ids
  int num_symbols = 0;
  int num_speakers = 1;
  std::map<std::string, int64_t> speaker_id_map;

  // IPA string -> phoneme ids with the rule of ESpeakPhonemizer.phonemeIDs (Sources/PiperCLI/ESpeakPhonemizer.swift:73-100)
  // and idsFromIPA (PiperCLI.swift:115-136): BOS "^", then per Unicode scalar its id followed by the blank "_",
  // EOS "$" without a trailing blank; CR / LF, format characters (general category Cf) and U+200B/C/D,
  // U+FE0E/F are skipped; a scalar missing from phoneme_id_map is an error naming the symbol.
  std::vector<int64_t> ipa_to_ids(const std::string& ipa_utf8) const;

  static PiperConfig load(const std::string& path);             // PiperConfig.load(from:) :41-45
  static PiperConfig from_json(const std::string& text);
};

}  // namespace piper
