#include "onnx_reader.h"

#include <cstring>
#include <fstream>
#include <set>
#include <stdexcept>

namespace piper {
namespace {

struct Reader {
  const uint8_t* p;
  const uint8_t* end;
  bool done() const { return p >= end; }
  uint64_t varint() {
    uint64_t v = 0;
    int shift = 0;
    while (true) {
      if (p >= end) throw std::runtime_error("onnx: truncated varint");
      uint8_t b = *p++;
      v |= uint64_t(b & 0x7F) << shift;
      if (!(b & 0x80)) return v;
      shift += 7;
      if (shift > 63) throw std::runtime_error("onnx: varint too long");
    }
  }
  Reader sub() {
    uint64_t n = varint();
    if (n > uint64_t(end - p)) throw std::runtime_error("onnx: truncated length-delimited field");
    Reader r{p, p + n};
    p += n;
    return r;
  }
  void skip(int wt) {
    switch (wt) {
      case 0: varint(); break;
      case 1: need(8); p += 8; break;
      case 2: sub(); break;
      case 5: need(4); p += 4; break;
      default: throw std::runtime_error("onnx: unsupported wire type " + std::to_string(wt));
    }
  }
  void need(size_t n) {
    if (size_t(end - p) < n) throw std::runtime_error("onnx: truncated fixed-width field");
  }
  float f32() {
    need(4);
    float v;
    std::memcpy(&v, p, 4);
    p += 4;
    return v;
  }
  std::string str() {
    Reader r = sub();
    return std::string(reinterpret_cast<const char*>(r.p), r.end - r.p);
  }
};

void read_ints(Reader& r, int wt, std::vector<int64_t>& out) {
  if (wt == 2) {
    Reader s = r.sub();
    while (!s.done()) out.push_back(int64_t(s.varint()));
  } else {
    out.push_back(int64_t(r.varint()));
  }
}

void read_floats(Reader& r, int wt, std::vector<float>& out) {
  if (wt == 2) {
    Reader s = r.sub();
    while (!s.done()) out.push_back(s.f32());
  } else {
    out.push_back(r.f32());
  }
}

OnnxTensor read_tensor(Reader r) {
  OnnxTensor t;
  const uint8_t* raw = nullptr;
  size_t raw_n = 0;
  std::vector<float> fdata;
  std::vector<int64_t> idata;
  while (!r.done()) {
    uint64_t key = r.varint();
    int f = int(key >> 3), wt = int(key & 7);
    switch (f) {
      case 1: read_ints(r, wt, t.dims); break;
      case 2: t.dtype = int32_t(r.varint()); break;
      case 4: read_floats(r, wt, fdata); break;
      case 7: read_ints(r, wt, idata); break;
      case 8: t.name = r.str(); break;
      case 9: {
        Reader s = r.sub();
        raw = s.p;
        raw_n = s.end - s.p;
        break;
      }
      default: r.skip(wt);
    }
  }
  // extents must be non-negative and their product must not overflow: the binders index f32 / i64 from dims
  int64_t n = 1;
  for (int64_t d : t.dims) {
    if (d < 0 || (d > 0 && n > ((int64_t)1 << 40) / d))
      throw std::runtime_error("onnx: bad dims for tensor '" + t.name + "'");
    n *= d;
  }
  auto check = [&](size_t elt) {
    if (raw && raw_n != size_t(n) * elt)
      throw std::runtime_error("onnx: raw_data size mismatch for '" + t.name + "'");
  };
  auto check_typed = [&](size_t have) {   // float_data / int64_data path: one value per element
    if (have != size_t(n))
      throw std::runtime_error("onnx: tensor '" + t.name + "' holds " + std::to_string(have) + " values for " +
                               std::to_string(n) + " elements");
  };
  switch (t.dtype) {
    case 1:
      if (raw) {
        check(4);
        t.f32.resize(n);
        std::memcpy(t.f32.data(), raw, raw_n);   // little-endian host assumed (x86-64 / aarch64)
      } else {
        check_typed(fdata.size());
        t.f32 = std::move(fdata);
      }
      break;
    case 7:
      if (raw) {
        check(8);
        t.i64.resize(n);
        std::memcpy(t.i64.data(), raw, raw_n);
      } else {
        check_typed(idata.size());
        t.i64 = std::move(idata);
      }
      break;
    case 6:   // int32 widens to int64 (TensorValue.swift:84-100)
      if (raw) {
        check(4);
        t.i64.resize(n);
        for (int64_t i = 0; i < n; ++i) {
          int32_t v;
          std::memcpy(&v, raw + 4 * i, 4);
          t.i64[i] = v;
        }
      } else {
        check_typed(idata.size());
        t.i64 = std::move(idata);
      }
      t.dtype = 7;
      break;
    case 9:
      if (raw) {
        check(1);
        t.i64.resize(n);
        for (int64_t i = 0; i < n; ++i) t.i64[i] = raw[i] != 0;
      } else {
        check_typed(idata.size());
        t.i64 = std::move(idata);
      }
      break;
    default:
      throw std::runtime_error("onnx: unsupported tensor data_type " + std::to_string(t.dtype) +
                               " for '" + t.name + "'");
  }
  return t;
}

OnnxAttr read_attr(Reader r) {
  OnnxAttr a;
  while (!r.done()) {
    uint64_t key = r.varint();
    int f = int(key >> 3), wt = int(key & 7);
    switch (f) {
      case 1: a.name = r.str(); break;
      case 2: a.f = r.f32(); break;
      case 3: a.i = int64_t(r.varint()); break;
      case 4: a.s = r.str(); break;
      case 5:   // tensor-valued attribute: the value of a Constant node (a float / int64 tensor; other types are skipped)
        try {
          a.t = read_tensor(r.sub());
          a.has_t = true;
        } catch (const std::runtime_error&) {
          a.has_t = false;
        }
        break;
      case 7: read_floats(r, wt, a.floats); break;
      case 8: read_ints(r, wt, a.ints); break;
      case 20: a.type = int32_t(r.varint()); break;
      default: r.skip(wt);   // graphs, sparse tensors: not needed on this path
    }
  }
  return a;
}

OnnxNode read_node(Reader r) {
  OnnxNode n;
  while (!r.done()) {
    uint64_t key = r.varint();
    int f = int(key >> 3), wt = int(key & 7);
    switch (f) {
      case 1: n.inputs.push_back(r.str()); break;
      case 2: n.outputs.push_back(r.str()); break;
      case 3: n.name = r.str(); break;
      case 4: n.op_type = r.str(); break;
      case 5: n.attrs.push_back(read_attr(r.sub())); break;
      case 7: n.domain = r.str(); break;
      default: r.skip(wt);
    }
  }
  return n;
}

std::string read_value_info_name(Reader r) {
  std::string name;
  while (!r.done()) {
    uint64_t key = r.varint();
    int f = int(key >> 3), wt = int(key & 7);
    if (f == 1) name = r.str();
    else r.skip(wt);
  }
  return name;
}

}  // namespace

OnnxModel parse_onnx(const uint8_t* data, size_t size) {
  OnnxModel m;
  Reader r{data, data + size};
  bool have_graph = false;
  Reader graph{nullptr, nullptr};
  while (!r.done()) {
    uint64_t key = r.varint();
    int f = int(key >> 3), wt = int(key & 7);
    if (f == 1 && wt == 0) {
      m.ir_version = int64_t(r.varint());
    } else if (f == 7 && wt == 2) {
      graph = r.sub();
      have_graph = true;
    } else if (f == 8 && wt == 2) {
      Reader o = r.sub();
      while (!o.done()) {
        uint64_t k2 = o.varint();
        if ((k2 >> 3) == 2 && (k2 & 7) == 0) {
          int64_t v = int64_t(o.varint());
          if (v > m.opset) m.opset = v;
        } else {
          o.skip(int(k2 & 7));
        }
      }
    } else {
      r.skip(wt);
    }
  }
  if (!have_graph) throw std::runtime_error("onnx: model has no graph");
  std::vector<std::string> raw_inputs;
  while (!graph.done()) {
    uint64_t key = graph.varint();
    int f = int(key >> 3), wt = int(key & 7);
    if (wt != 2) {
      graph.skip(wt);
      continue;
    }
    switch (f) {
      case 1: m.nodes.push_back(read_node(graph.sub())); break;
      case 2: m.graph_name = graph.str(); break;
      case 5: m.initializers.push_back(read_tensor(graph.sub())); break;
      case 11: raw_inputs.push_back(read_value_info_name(graph.sub())); break;
      case 12: m.outputs.push_back(read_value_info_name(graph.sub())); break;
      default: graph.skip(wt);
    }
  }
  std::set<std::string> init_names;
  for (auto& t : m.initializers) init_names.insert(t.name);
  for (auto& n : raw_inputs)
    if (!init_names.count(n)) m.inputs.push_back(n);
  return m;
}

OnnxModel load_onnx(const std::string& path) {
  std::ifstream f(path, std::ios::binary | std::ios::ate);
  if (!f) throw std::runtime_error("io: cannot open '" + path + "'");
  std::streamsize n = f.tellg();
  f.seekg(0);
  std::vector<uint8_t> buf(static_cast<size_t>(n));
  if (n > 0 && !f.read(reinterpret_cast<char*>(buf.data()), n))
    throw std::runtime_error("io: short read on '" + path + "'");
  return parse_onnx(buf.data(), buf.size());
}

}  // namespace piper
