// Generic ONNX-op executor on the GPU: the B200 counterpart of the reference's node-by-node
// interpreter, GraphExecutor.executeNode (Sources/PiperMetal/Execution/GraphExecutor.swift:591-2663) over MetalBackend /
// Kernels/*.metal - for graphs the fused Piper-VITS pipeline does not recognise, and as an independent second
// implementation of the recognised ones.  Every node runs as one or a few plain CUDA kernels (fp32 / int64 / bool
// tensors, TensorValue.swift:4-8; general broadcasting; one thread per output element like the reference's Metal
// kernels).  It is the fallback, not the hot path: no tensor cores, no fusion.  Data-dependent shapes (NonZero, the
// shape-arithmetic chains of an export) are read back to the host where a node needs them as a shape.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <map>
#include <string>
#include <vector>

#include "kernels.h"
#include "onnx_reader.h"

namespace piper {

struct GTensor {
  int dtype = 1;                  // ONNX data_type: 1 float32, 7 int64, 9 bool (one byte per element)
  std::vector<int64_t> shape;
  void* d = nullptr;              // device memory (owned by the executor's pools)
  int64_t numel() const {
    int64_t n = 1;
    for (int64_t v : shape) n *= v;
    return n;
  }
  size_t esize() const { return dtype == 1 ? 4 : (dtype == 7 ? 8 : 1); }
  size_t bytes() const { return (size_t)numel() * esize(); }
};

class GraphExec {
 public:
  explicit GraphExec(const OnnxModel& m);   // uploads initializers and Constant values once
  ~GraphExec();
  GraphExec(const GraphExec&) = delete;
  GraphExec& operator=(const GraphExec&) = delete;

  // Runs every node in graph order.  feeds: graph inputs by name (host data, uploaded here).  Afterwards `fetch` reads
  // any value of the run (node outputs stay alive until the next run).
  struct Feed { std::string name; int dtype; std::vector<int64_t> shape; const void* data; };
  void run(Exec& ex, const std::vector<Feed>& feeds, uint32_t seed);
  const GTensor* find(const std::string& name) const;
  void fetch(Exec& ex, const GTensor& t, void* dst_host) const;

  const OnnxModel& model() const { return m_; }
  int64_t nodes_run = 0;

 private:
  const OnnxModel& m_;
  std::map<std::string, GTensor> consts_;   // initializers + Constant nodes (persistent pool)
  std::map<std::string, GTensor> vals_;     // node outputs and feeds of the current run (run pool)
  struct Pool {
    std::vector<std::pair<char*, size_t>> chunks;
    size_t cur = 0, off = 0;
    void* alloc(size_t bytes);
    void reset() { cur = 0; off = 0; }
    void release();
  } persist_, runp_;
  cudaStream_t st_ = nullptr;
  Exec* ex_ = nullptr;
  uint32_t seed_ = 1234;

  GTensor make(int dtype, const std::vector<int64_t>& shape, bool persistent = false);
  GTensor upload(int dtype, const std::vector<int64_t>& shape, const void* host, bool persistent = false);
  const GTensor& in(const OnnxNode& n, size_t i) const;
  bool has_in(const OnnxNode& n, size_t i) const { return i < n.inputs.size() && !n.inputs[i].empty(); }
  std::vector<int64_t> host_ints(const GTensor& t) const;   // small tensor -> host (stream sync)
  double host_scalar(const GTensor& t) const;
  void exec(const OnnxNode& n);
  // op families
  void op_binary(const OnnxNode& n);
  void op_unary(const OnnxNode& n);
  void op_movement(const OnnxNode& n);
  void op_index(const OnnxNode& n);
  void op_reduce(const OnnxNode& n);
  void op_linear(const OnnxNode& n);
  GTensor strided_copy(const GTensor& src, const std::vector<int64_t>& out_shape, const std::vector<int64_t>& src_strides,
                       int64_t src_off);
  void copy_into(const GTensor& src, GTensor& dst, const std::vector<int64_t>& dst_strides, int64_t dst_off);
};

}  // namespace piper
