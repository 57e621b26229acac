// C++ twin of the reference's PiperONNX module (Sources/PiperONNX): a protobuf wire reader
// (Wire.swift:38-150) and the ONNX subset decoder (ONNXLoader.swift:25-385) producing an IR
// equivalent to ONNXIR.swift:3-93.  Written from the field tables, not translated.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace piper {

struct OnnxTensor {
  std::string name;
  int32_t dtype = 0;                 // ONNX data_type (1 f32, 6 i32, 7 i64, 9 bool)
  std::vector<int64_t> dims;
  std::vector<float> f32;            // dtype 1
  std::vector<int64_t> i64;          // dtype 6 (widened), 7, 9
  int64_t count() const {
    int64_t n = 1;
    for (auto d : dims) n *= d;
    return n;
  }
};

struct OnnxAttr {
  std::string name;
  int32_t type = 0;
  float f = 0.f;
  int64_t i = 0;
  std::string s;
  std::vector<int64_t> ints;
  std::vector<float> floats;
  bool has_t = false;                // attribute `t` (field 5): the value of a Constant node
  OnnxTensor t;
};

struct OnnxNode {
  std::string op_type, name, domain;
  std::vector<std::string> inputs, outputs;
  std::vector<OnnxAttr> attrs;
  const OnnxAttr* attr(const std::string& n) const {
    for (auto& a : attrs)
      if (a.name == n) return &a;
    return nullptr;
  }
};

struct OnnxModel {
  int64_t ir_version = 0;
  int64_t opset = 0;
  std::string graph_name;
  std::vector<OnnxNode> nodes;
  std::vector<OnnxTensor> initializers;
  std::vector<std::string> inputs, outputs;   // graph feeds that are not initializers
};

// Throws std::runtime_error with an ONNXDecodingError-style message.
OnnxModel load_onnx(const std::string& path);
OnnxModel parse_onnx(const uint8_t* data, size_t size);

}  // namespace piper
