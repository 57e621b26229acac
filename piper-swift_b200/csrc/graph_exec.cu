// Generic ONNX-op executor (see graph_exec.h).  Reference: GraphExecutor.executeNode,
// Sources/PiperMetal/Execution/GraphExecutor.swift:591-2663; kernels Sources/PiperMetal/Kernels/*.metal.
#include "graph_exec.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <stdexcept>

namespace piper {

namespace {

#define G_CUDA_OK(x)                                                                         \
  do {                                                                                       \
    cudaError_t e_ = (x);                                                                    \
    if (e_ != cudaSuccess)                                                                   \
      throw std::runtime_error(std::string("cuda: ") + cudaGetErrorString(e_) + " at " #x);  \
  } while (0)

constexpr int GR = 8;   // maximum tensor rank
struct Dims {
  int rank;
  int64_t dim[GR];      // iteration space (the output's shape)
  int64_t s0[GR], s1[GR], s2[GR];   // element strides of up to three operands in that space (0 = broadcast)
};

__device__ __forceinline__ void decompose(const Dims& d, int64_t i, int64_t& o0, int64_t& o1, int64_t& o2) {
  o0 = o1 = o2 = 0;
  for (int a = d.rank - 1; a >= 0; --a) {
    const int64_t c = i % d.dim[a];
    i /= d.dim[a];
    o0 += c * d.s0[a]; o1 += c * d.s1[a]; o2 += c * d.s2[a];
  }
}

enum BinOp { B_ADD, B_SUB, B_MUL, B_DIV, B_POW, B_EQ, B_LT, B_LE, B_GE, B_AND };

template <typename T> __device__ __forceinline__ T bin_arith(int op, T a, T b);
template <> __device__ __forceinline__ float bin_arith<float>(int op, float a, float b) {
  switch (op) {
    case B_ADD: return a + b;
    case B_SUB: return a - b;
    case B_MUL: return a * b;
    case B_DIV: return a / b;
    default: return powf(a, b);
  }
}
template <> __device__ __forceinline__ int64_t bin_arith<int64_t>(int op, int64_t a, int64_t b) {
  switch (op) {
    case B_ADD: return a + b;
    case B_SUB: return a - b;
    case B_MUL: return a * b;
    case B_DIV: return b ? a / b : 0;      // C++ integer division truncates toward zero, like the reference
    default: return (int64_t)pow((double)a, (double)b);
  }
}

// elementwise.metal broadcasting binary ops (GraphExecutor.swift:1237-1530)
template <typename T>
__global__ void k_binary(int op, Dims d, const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ out, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int64_t oa, ob, oc;
  decompose(d, i, oa, ob, oc);
  out[i] = bin_arith<T>(op, a[oa], b[ob]);
}
template <typename T>
__global__ void k_compare(int op, Dims d, const T* __restrict__ a, const T* __restrict__ b, uint8_t* __restrict__ out, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int64_t oa, ob, oc;
  decompose(d, i, oa, ob, oc);
  const T x = a[oa], y = b[ob];
  bool r;
  switch (op) {
    case B_EQ: r = x == y; break;
    case B_LT: r = x < y; break;
    case B_LE: r = x <= y; break;
    case B_GE: r = x >= y; break;
    default: r = (x != 0) && (y != 0); break;
  }
  out[i] = r ? 1 : 0;
}
template <typename T>
__global__ void k_where(Dims d, const uint8_t* __restrict__ c, const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ out,
                        int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int64_t oc, oa, ob;
  decompose(d, i, oc, oa, ob);
  out[i] = c[oc] ? a[oa] : b[ob];
}

enum UnOp { U_NEG, U_EXP, U_SQRT, U_TANH, U_SIGMOID, U_SOFTPLUS, U_RELU, U_ERF, U_CEIL, U_LRELU, U_CLIP };
__global__ void k_unary_f32(int op, const float* __restrict__ x, float* __restrict__ out, int64_t n, float p0, float p1) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float v = x[i];
  float r;
  switch (op) {
    case U_NEG: r = -v; break;
    case U_EXP: r = expf(v); break;
    case U_SQRT: r = sqrtf(v); break;
    case U_TANH: r = tanhf(v); break;
    case U_SIGMOID: r = 1.0f / (1.0f + expf(-v)); break;
    case U_SOFTPLUS: r = v > 20.f ? v : log1pf(expf(v)); break;
    case U_RELU: r = fmaxf(v, 0.f); break;
    case U_ERF: r = erff(v); break;
    case U_CEIL: r = ceilf(v); break;
    case U_LRELU: r = v >= 0.f ? v : v * p0; break;
    default: r = fminf(fmaxf(v, p0), p1); break;
  }
  out[i] = r;
}
__global__ void k_neg_i64(const int64_t* __restrict__ x, int64_t* __restrict__ out, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = -x[i];
}
__global__ void k_clip_i64(const int64_t* __restrict__ x, int64_t* __restrict__ out, int64_t n, int64_t lo, int64_t hi) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = x[i] < lo ? lo : (x[i] > hi ? hi : x[i]);
}
__global__ void k_not(const uint8_t* __restrict__ x, uint8_t* __restrict__ out, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = x[i] ? 0 : 1;
}
// cast.metal:8-29: float -> int64 truncates toward zero, anything -> bool is != 0
template <typename TI, typename TO>
__global__ void k_cast(const TI* __restrict__ x, TO* __restrict__ out, int64_t n, int to_bool) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (to_bool) out[i] = (TO)(x[i] != (TI)0 ? 1 : 0);
  else out[i] = (TO)x[i];
}
template <typename T>
__global__ void k_fill(T* __restrict__ out, int64_t n, T v) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = v;
}
// out[doff + c . s1] = src[soff + c . s0] over the index space d.dim: Transpose, Slice (negative steps), Expand (stride 0),
// Concat / Pad / Split (a sub-block of the destination)
template <typename T>
__global__ void k_copy_strided(Dims d, const T* __restrict__ src, T* __restrict__ dst, int64_t soff, int64_t doff, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int64_t os, od, o2;
  decompose(d, i, os, od, o2);
  dst[doff + od] = src[soff + os];
}
// Gather (GraphExecutor.swift:780-860): out[o][j][r] = data[o][idx[j]][r]
template <typename T>
__global__ void k_gather(const T* __restrict__ data, const int64_t* __restrict__ idx, T* __restrict__ out, int64_t outer, int64_t axis_dim,
                         int64_t inner, int64_t n_idx) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= outer * n_idx * inner) return;
  const int64_t r = i % inner, j = (i / inner) % n_idx, o = i / (inner * n_idx);
  int64_t k = idx[j];
  if (k < 0) k += axis_dim;
  out[i] = data[(o * axis_dim + k) * inner + r];
}
// GatherElements along `axis` (d.s0 = data strides with the axis stride zeroed, d.s1 unused)
template <typename T>
__global__ void k_gather_elements(Dims d, const T* __restrict__ data, const int64_t* __restrict__ idx, T* __restrict__ out,
                                  int64_t axis_stride, int64_t axis_dim, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int64_t o0, o1, o2;
  decompose(d, i, o0, o1, o2);
  int64_t k = idx[i];
  if (k < 0) k += axis_dim;
  out[i] = data[o0 + k * axis_stride];
}
// GatherND / ScatterND (GraphExecutor.swift:2120-2189): idx [n_idx][r] addresses a slice of `slice` elements
template <typename T>
__global__ void k_gathernd(const T* __restrict__ data, const int64_t* __restrict__ idx, T* __restrict__ out, Dims d, int r, int64_t slice,
                           int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t s = i % slice, j = i / slice;
  int64_t off = 0;
  for (int a = 0; a < r; ++a) {
    int64_t k = idx[j * r + a];
    if (k < 0) k += d.dim[a];
    off += k * d.s0[a];
  }
  out[i] = data[off + s];
}
template <typename T>
__global__ void k_scatternd(T* __restrict__ out, const int64_t* __restrict__ idx, const T* __restrict__ upd, Dims d, int r, int64_t slice,
                            int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t s = i % slice, j = i / slice;
  int64_t off = 0;
  for (int a = 0; a < r; ++a) {
    int64_t k = idx[j * r + a];
    if (k < 0) k += d.dim[a];
    off += k * d.s0[a];
  }
  out[off + s] = upd[i];
}

// ReduceMean / ReduceSum / ReduceMax over a contiguous block of axes: [outer][red][inner] (GraphExecutor.swift:1930-2015)
enum RedOp { R_MEAN, R_SUM, R_MAX };
template <typename T>
__global__ void k_reduce(int op, const T* __restrict__ x, T* __restrict__ out, int64_t outer, int64_t red, int64_t inner) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= outer * inner) return;
  const int64_t o = i / inner, r = i % inner;
  const T* p = x + o * red * inner + r;
  T acc = op == R_MAX ? p[0] : (T)0;
  for (int64_t k = 0; k < red; ++k) {
    const T v = p[k * inner];
    if (op == R_MAX) acc = v > acc ? v : acc;
    else acc += v;
  }
  if (op == R_MEAN) acc = acc / (T)red;
  out[i] = acc;
}
// softmax.metal:13-41: max, exp, sum, divide along the axis
__global__ void k_softmax(const float* __restrict__ x, float* __restrict__ out, int64_t outer, int64_t n, int64_t inner) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= outer * inner) return;
  const int64_t o = i / inner, r = i % inner;
  const float* p = x + o * n * inner + r;
  float* q = out + o * n * inner + r;
  float mx = -FLT_MAX;
  for (int64_t k = 0; k < n; ++k) mx = fmaxf(mx, p[k * inner]);
  float sum = 0.f;
  for (int64_t k = 0; k < n; ++k) {
    const float e = expf(p[k * inner] - mx);
    q[k * inner] = e;
    sum += e;
  }
  for (int64_t k = 0; k < n; ++k) q[k * inner] = q[k * inner] / sum;
}
// cumsum.metal:10-32 (inclusive, forward)
template <typename T>
__global__ void k_cumsum(const T* __restrict__ x, T* __restrict__ out, int64_t outer, int64_t n, int64_t inner) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= outer * inner) return;
  const int64_t o = i / inner, r = i % inner;
  T acc = (T)0;
  for (int64_t k = 0; k < n; ++k) {
    acc += x[(o * n + k) * inner + r];
    out[(o * n + k) * inner + r] = acc;
  }
}
// matmul.metal:22-49: one thread per output element; d.dim = batch dims, s0 / s1 = batch strides of A / B
__global__ void k_matmul(Dims d, const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out, int64_t batch, int M,
                         int N, int K) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= batch * M * N) return;
  const int col = (int)(i % N), row = (int)((i / N) % M);
  const int64_t bi = i / ((int64_t)M * N);
  int64_t oa, ob, oc;
  decompose(d, bi, oa, ob, oc);
  const float* ap = a + oa + (int64_t)row * K;
  const float* bp = b + ob + col;
  float acc = 0.f;
  for (int k = 0; k < K; ++k) acc = fmaf(ap[k], bp[(int64_t)k * N], acc);
  out[i] = acc;
}
// conv1d.metal:28-71 / 97-142: x [N][Cin][L], one thread per output sample, zero padding, groups
__global__ void k_conv1d(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias, float* __restrict__ out,
                         int N, int Cin, int L, int Cout, int K, int stride, int dil, int pad, int group, int Lout) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)N * Cout * Lout) return;
  const int t = (int)(i % Lout), co = (int)((i / Lout) % Cout), nb = (int)(i / ((int64_t)Lout * Cout));
  const int cg = Cin / group, og = Cout / group, g = co / og;
  float acc = bias ? bias[co] : 0.f;
  for (int c = 0; c < cg; ++c) {
    const float* xp = x + ((int64_t)nb * Cin + g * cg + c) * L;
    const float* wp = w + ((int64_t)co * cg + c) * K;
    for (int k = 0; k < K; ++k) {
      const int p = t * stride - pad + k * dil;
      if (p >= 0 && p < L) acc = fmaf(xp[p], wp[k], acc);
    }
  }
  out[i] = acc;
}
__global__ void k_convtr1d(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias, float* __restrict__ out,
                           int N, int Cin, int L, int Cout, int K, int stride, int dil, int pad, int group, int Lout) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)N * Cout * Lout) return;
  const int t = (int)(i % Lout), co = (int)((i / Lout) % Cout), nb = (int)(i / ((int64_t)Lout * Cout));
  const int cg = Cin / group, og = Cout / group, g = co / og, col = co - g * og;
  float acc = bias ? bias[co] : 0.f;
  for (int k = 0; k < K; ++k) {
    const int num = t + pad - k * dil;
    if (num < 0 || num % stride) continue;
    const int p = num / stride;
    if (p >= L) continue;
    for (int c = 0; c < cg; ++c)
      acc = fmaf(x[((int64_t)nb * Cin + g * cg + c) * L + p], w[(((int64_t)(g * cg + c)) * og + col) * K + k], acc);
  }
  out[i] = acc;
}
__global__ void k_scale_shift(float* __restrict__ x, int64_t n, float scale, float mean) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) x[i] = x[i] * scale + mean;
}

inline unsigned blocks(int64_t n) { return (unsigned)((n + 255) / 256); }

std::vector<int64_t> contiguous_strides(const std::vector<int64_t>& shape) {
  std::vector<int64_t> s(shape.size(), 1);
  for (int a = (int)shape.size() - 2; a >= 0; --a) s[a] = s[a + 1] * shape[a + 1];
  return s;
}
// numpy-style broadcast of several shapes
std::vector<int64_t> broadcast_shape(const std::vector<std::vector<int64_t>>& shapes) {
  size_t r = 0;
  for (auto& s : shapes) r = std::max(r, s.size());
  std::vector<int64_t> out(r, 1);
  for (auto& s : shapes)
    for (size_t a = 0; a < s.size(); ++a) {
      const size_t o = r - s.size() + a;
      if (s[a] != 1) {
        if (out[o] != 1 && out[o] != s[a]) throw std::runtime_error("graph: shapes do not broadcast");
        out[o] = s[a];
      }
    }
  return out;
}
// strides of `shape` viewed in the broadcast space `out` (0 where it is stretched)
void bstrides(const std::vector<int64_t>& shape, const std::vector<int64_t>& out, int64_t* dst) {
  const std::vector<int64_t> cs = contiguous_strides(shape);
  const size_t r = out.size();
  for (size_t a = 0; a < r; ++a) dst[a] = 0;
  for (size_t a = 0; a < shape.size(); ++a) {
    const size_t o = r - shape.size() + a;
    dst[o] = shape[a] == 1 ? 0 : cs[a];
  }
}
Dims make_dims(const std::vector<int64_t>& out) {
  if (out.size() > GR) throw std::runtime_error("graph: tensor rank above 8");
  Dims d{};
  d.rank = (int)out.size();
  for (size_t a = 0; a < out.size(); ++a) d.dim[a] = out[a];
  return d;
}
int64_t attr_i(const OnnxNode& n, const char* name, int64_t dflt) {
  const OnnxAttr* a = n.attr(name);
  if (!a) return dflt;
  return a->ints.empty() ? a->i : a->ints[0];
}
float attr_f(const OnnxNode& n, const char* name, float dflt) {
  const OnnxAttr* a = n.attr(name);
  return a ? a->f : dflt;
}
std::vector<int64_t> attr_ints(const OnnxNode& n, const char* name) {
  const OnnxAttr* a = n.attr(name);
  if (!a) return {};
  if (!a->ints.empty()) return a->ints;
  return {};
}

}  // namespace

// ---- pools, tensors ---------------------------------------------------------------------------------------------
void* GraphExec::Pool::alloc(size_t bytes) {
  bytes = (bytes + 255) & ~size_t(255);
  if (bytes == 0) bytes = 256;
  while (cur < chunks.size()) {
    if (off + bytes <= chunks[cur].second) {
      void* p = chunks[cur].first + off;
      off += bytes;
      return p;
    }
    ++cur;
    off = 0;
  }
  const size_t cap = std::max(bytes, (size_t)64 << 20);
  char* p = nullptr;
  G_CUDA_OK(cudaMalloc(&p, cap));
  chunks.push_back({p, cap});
  cur = chunks.size() - 1;
  off = bytes;
  return p;
}
void GraphExec::Pool::release() {
  for (auto& c : chunks) cudaFree(c.first);
  chunks.clear();
  reset();
}

GTensor GraphExec::make(int dtype, const std::vector<int64_t>& shape, bool persistent) {
  if (dtype != 1 && dtype != 7 && dtype != 9) throw std::runtime_error("unsupported: tensor data_type " + std::to_string(dtype));
  if (shape.size() > GR) throw std::runtime_error("unsupported: tensor rank above 8");
  GTensor t;
  t.dtype = dtype;
  t.shape = shape;
  for (int64_t v : shape)
    if (v < 0) throw std::runtime_error("graph: negative dimension");
  t.d = (persistent ? persist_ : runp_).alloc(t.bytes());
  return t;
}
GTensor GraphExec::upload(int dtype, const std::vector<int64_t>& shape, const void* host, bool persistent) {
  GTensor t = make(dtype, shape, persistent);
  if (t.bytes()) G_CUDA_OK(cudaMemcpyAsync(t.d, host, t.bytes(), cudaMemcpyHostToDevice, st_));
  // the source may be a temporary: make the copy complete before it goes away (small tensors; not the hot path)
  G_CUDA_OK(cudaStreamSynchronize(st_));
  return t;
}

static GTensor from_onnx(GraphExec* self, const OnnxTensor& t, bool persistent,
                         GTensor (GraphExec::*up)(int, const std::vector<int64_t>&, const void*, bool)) {
  if (t.dtype == 1) return (self->*up)(1, t.dims, t.f32.data(), persistent);
  if (t.dtype == 7 || t.dtype == 6) return (self->*up)(7, t.dims, t.i64.data(), persistent);
  if (t.dtype == 9) {
    std::vector<uint8_t> b(t.i64.size());
    for (size_t i = 0; i < b.size(); ++i) b[i] = t.i64[i] ? 1 : 0;
    return (self->*up)(9, t.dims, b.data(), persistent);
  }
  throw std::runtime_error("unsupported: initializer '" + t.name + "' has data_type " + std::to_string(t.dtype));
}

GraphExec::GraphExec(const OnnxModel& m) : m_(m) {
  for (const OnnxTensor& t : m.initializers) {
    if ((t.dtype == 1 && (int64_t)t.f32.size() != t.count()) || (t.dtype != 1 && (int64_t)t.i64.size() != t.count()))
      throw std::runtime_error("onnx: initializer '" + t.name + "' element count does not match its dims");
    consts_[t.name] = from_onnx(this, t, true, &GraphExec::upload);
  }
  for (const OnnxNode& n : m.nodes) {
    if (n.op_type != "Constant" || n.outputs.empty()) continue;
    const OnnxAttr* a = n.attr("value");
    if (!a || !a->has_t) throw std::runtime_error("unsupported: Constant node '" + n.name + "' without a tensor value");
    consts_[n.outputs[0]] = from_onnx(this, a->t, true, &GraphExec::upload);
  }
}
GraphExec::~GraphExec() {
  persist_.release();
  runp_.release();
}

const GTensor* GraphExec::find(const std::string& name) const {
  auto it = vals_.find(name);
  if (it != vals_.end()) return &it->second;
  auto ic = consts_.find(name);
  return ic == consts_.end() ? nullptr : &ic->second;
}
const GTensor& GraphExec::in(const OnnxNode& n, size_t i) const {
  if (i >= n.inputs.size() || n.inputs[i].empty()) throw std::runtime_error("graph: node '" + n.name + "' misses input " + std::to_string(i));
  const GTensor* t = find(n.inputs[i]);
  if (!t) throw std::runtime_error("graph: value '" + n.inputs[i] + "' is not defined before node '" + n.name + "'");
  return *t;
}
void GraphExec::fetch(Exec& ex, const GTensor& t, void* dst_host) const {
  if (!t.bytes()) return;
  G_CUDA_OK(cudaMemcpyAsync(dst_host, t.d, t.bytes(), cudaMemcpyDeviceToHost, ex.st));
  G_CUDA_OK(cudaStreamSynchronize(ex.st));
}
std::vector<int64_t> GraphExec::host_ints(const GTensor& t) const {
  const int64_t n = t.numel();
  if (n > (1 << 20)) throw std::runtime_error("graph: shape-like tensor too large");
  std::vector<int64_t> out((size_t)n);
  if (!n) return out;
  if (t.dtype == 7) {
    G_CUDA_OK(cudaMemcpyAsync(out.data(), t.d, (size_t)n * 8, cudaMemcpyDeviceToHost, st_));
    G_CUDA_OK(cudaStreamSynchronize(st_));
  } else if (t.dtype == 1) {
    std::vector<float> f((size_t)n);
    G_CUDA_OK(cudaMemcpyAsync(f.data(), t.d, (size_t)n * 4, cudaMemcpyDeviceToHost, st_));
    G_CUDA_OK(cudaStreamSynchronize(st_));
    for (int64_t i = 0; i < n; ++i) out[i] = (int64_t)f[i];
  } else {
    std::vector<uint8_t> b((size_t)n);
    G_CUDA_OK(cudaMemcpyAsync(b.data(), t.d, (size_t)n, cudaMemcpyDeviceToHost, st_));
    G_CUDA_OK(cudaStreamSynchronize(st_));
    for (int64_t i = 0; i < n; ++i) out[i] = b[i];
  }
  return out;
}
double GraphExec::host_scalar(const GTensor& t) const {
  if (t.numel() < 1) throw std::runtime_error("graph: empty scalar");
  if (t.dtype == 1) {
    float f = 0.f;
    G_CUDA_OK(cudaMemcpyAsync(&f, t.d, 4, cudaMemcpyDeviceToHost, st_));
    G_CUDA_OK(cudaStreamSynchronize(st_));
    return f;
  }
  return (double)host_ints(t)[0];
}

void GraphExec::run(Exec& ex, const std::vector<Feed>& feeds, uint32_t seed) {
  ex_ = &ex;
  st_ = ex.st;
  seed_ = seed;
  vals_.clear();
  runp_.reset();
  nodes_run = 0;
  for (const Feed& f : feeds) vals_[f.name] = upload(f.dtype, f.shape, f.data, false);
  for (const std::string& name : m_.inputs)
    if (!vals_.count(name)) throw std::runtime_error("invalid argument: graph input '" + name + "' was not fed");
  for (const OnnxNode& n : m_.nodes) {
    if (n.op_type == "Constant") continue;
    try {
      exec(n);
    } catch (const std::exception& e) {
      const std::string msg = e.what();
      if (msg.rfind("unsupported", 0) == 0 || msg.rfind("cuda", 0) == 0) throw;
      throw std::runtime_error(msg + " [node '" + n.name + "' (" + n.op_type + ")]");
    }
    ++nodes_run;
  }
  G_CUDA_OK(cudaGetLastError());
}

void GraphExec::exec(const OnnxNode& n) {
  const std::string& op = n.op_type;
  static const char* bin[] = {"Add", "Sub", "Mul", "Div", "Pow", "Equal", "Less", "LessOrEqual", "GreaterOrEqual", "And", "Where"};
  static const char* un[] = {"Neg", "Exp", "Sqrt", "Tanh", "Sigmoid", "Softplus", "Relu", "Erf", "Ceil", "LeakyRelu", "Clip", "Not", "Cast"};
  static const char* mv[] = {"Shape", "ConstantOfShape", "Range", "Reshape", "Transpose", "Unsqueeze", "Squeeze", "Concat", "Split",
                             "Slice", "Expand", "Pad"};
  static const char* ix[] = {"Gather", "GatherElements", "GatherND", "ScatterND", "NonZero"};
  static const char* rd[] = {"ReduceMean", "ReduceSum", "ReduceMax", "Softmax", "CumSum"};
  static const char* ln[] = {"MatMul", "Conv", "ConvTranspose", "RandomNormalLike"};
  for (const char* s : bin) if (op == s) return op_binary(n);
  for (const char* s : un) if (op == s) return op_unary(n);
  for (const char* s : mv) if (op == s) return op_movement(n);
  for (const char* s : ix) if (op == s) return op_index(n);
  for (const char* s : rd) if (op == s) return op_reduce(n);
  for (const char* s : ln) if (op == s) return op_linear(n);
  throw std::runtime_error("unsupported: op '" + op + "' (node '" + n.name + "')");   // ExecutionError.unsupportedOp
}

// ---- elementwise ------------------------------------------------------------------------------------------------
void GraphExec::op_binary(const OnnxNode& n) {
  const std::string& op = n.op_type;
  if (op == "Where") {
    const GTensor &c = in(n, 0), &a = in(n, 1), &b = in(n, 2);
    if (c.dtype != 9 || a.dtype != b.dtype) throw std::runtime_error("graph: Where operand types");
    const std::vector<int64_t> os = broadcast_shape({c.shape, a.shape, b.shape});
    GTensor o = make(a.dtype, os);
    Dims d = make_dims(os);
    bstrides(c.shape, os, d.s0); bstrides(a.shape, os, d.s1); bstrides(b.shape, os, d.s2);
    const int64_t cnt = o.numel();
    if (cnt) {
      if (a.dtype == 1) k_where<float><<<blocks(cnt), 256, 0, st_>>>(d, (const uint8_t*)c.d, (const float*)a.d, (const float*)b.d, (float*)o.d, cnt);
      else if (a.dtype == 7) k_where<int64_t><<<blocks(cnt), 256, 0, st_>>>(d, (const uint8_t*)c.d, (const int64_t*)a.d, (const int64_t*)b.d, (int64_t*)o.d, cnt);
      else k_where<uint8_t><<<blocks(cnt), 256, 0, st_>>>(d, (const uint8_t*)c.d, (const uint8_t*)a.d, (const uint8_t*)b.d, (uint8_t*)o.d, cnt);
      ++ex_->launches;
    }
    vals_[n.outputs[0]] = o;
    return;
  }
  const GTensor &a = in(n, 0), &b = in(n, 1);
  if (a.dtype != b.dtype) throw std::runtime_error("graph: " + op + " on mixed types " + std::to_string(a.dtype) + " / " + std::to_string(b.dtype));
  const std::vector<int64_t> os = broadcast_shape({a.shape, b.shape});
  Dims d = make_dims(os);
  bstrides(a.shape, os, d.s0); bstrides(b.shape, os, d.s1);
  int code;
  bool cmp = false;
  if (op == "Add") code = B_ADD; else if (op == "Sub") code = B_SUB; else if (op == "Mul") code = B_MUL;
  else if (op == "Div") code = B_DIV; else if (op == "Pow") code = B_POW;
  else { cmp = true; code = op == "Equal" ? B_EQ : op == "Less" ? B_LT : op == "LessOrEqual" ? B_LE : op == "GreaterOrEqual" ? B_GE : B_AND; }
  GTensor o = make(cmp ? 9 : a.dtype, os);
  const int64_t cnt = o.numel();
  if (cnt) {
    if (cmp) {
      if (a.dtype == 1) k_compare<float><<<blocks(cnt), 256, 0, st_>>>(code, d, (const float*)a.d, (const float*)b.d, (uint8_t*)o.d, cnt);
      else if (a.dtype == 7) k_compare<int64_t><<<blocks(cnt), 256, 0, st_>>>(code, d, (const int64_t*)a.d, (const int64_t*)b.d, (uint8_t*)o.d, cnt);
      else k_compare<uint8_t><<<blocks(cnt), 256, 0, st_>>>(code, d, (const uint8_t*)a.d, (const uint8_t*)b.d, (uint8_t*)o.d, cnt);
    } else if (a.dtype == 1) {
      k_binary<float><<<blocks(cnt), 256, 0, st_>>>(code, d, (const float*)a.d, (const float*)b.d, (float*)o.d, cnt);
    } else if (a.dtype == 7) {
      k_binary<int64_t><<<blocks(cnt), 256, 0, st_>>>(code, d, (const int64_t*)a.d, (const int64_t*)b.d, (int64_t*)o.d, cnt);
    } else {
      throw std::runtime_error("graph: arithmetic on bool tensors");
    }
    ++ex_->launches;
  }
  vals_[n.outputs[0]] = o;
}

void GraphExec::op_unary(const OnnxNode& n) {
  const std::string& op = n.op_type;
  const GTensor& a = in(n, 0);
  const int64_t cnt = a.numel();
  if (op == "Cast") {
    const int64_t to_raw = attr_i(n, "to", 1);
    const int to = to_raw == 6 ? 7 : (to_raw == 11 ? 1 : (int)to_raw);   // int32 widens, double narrows (TensorValue.swift:84-100)
    GTensor o = make(to, a.shape);
    if (cnt) {
      const unsigned g = blocks(cnt);
      if (a.dtype == 1 && to == 1) k_cast<float, float><<<g, 256, 0, st_>>>((const float*)a.d, (float*)o.d, cnt, 0);
      else if (a.dtype == 1 && to == 7) k_cast<float, int64_t><<<g, 256, 0, st_>>>((const float*)a.d, (int64_t*)o.d, cnt, 0);
      else if (a.dtype == 1 && to == 9) k_cast<float, uint8_t><<<g, 256, 0, st_>>>((const float*)a.d, (uint8_t*)o.d, cnt, 1);
      else if (a.dtype == 7 && to == 1) k_cast<int64_t, float><<<g, 256, 0, st_>>>((const int64_t*)a.d, (float*)o.d, cnt, 0);
      else if (a.dtype == 7 && to == 7) k_cast<int64_t, int64_t><<<g, 256, 0, st_>>>((const int64_t*)a.d, (int64_t*)o.d, cnt, 0);
      else if (a.dtype == 7 && to == 9) k_cast<int64_t, uint8_t><<<g, 256, 0, st_>>>((const int64_t*)a.d, (uint8_t*)o.d, cnt, 1);
      else if (a.dtype == 9 && to == 1) k_cast<uint8_t, float><<<g, 256, 0, st_>>>((const uint8_t*)a.d, (float*)o.d, cnt, 0);
      else if (a.dtype == 9 && to == 7) k_cast<uint8_t, int64_t><<<g, 256, 0, st_>>>((const uint8_t*)a.d, (int64_t*)o.d, cnt, 0);
      else k_cast<uint8_t, uint8_t><<<g, 256, 0, st_>>>((const uint8_t*)a.d, (uint8_t*)o.d, cnt, 1);
      ++ex_->launches;
    }
    vals_[n.outputs[0]] = o;
    return;
  }
  GTensor o = make(a.dtype, a.shape);
  if (op == "Not") {
    if (a.dtype != 9) throw std::runtime_error("graph: Not on a non-bool tensor");
    if (cnt) { k_not<<<blocks(cnt), 256, 0, st_>>>((const uint8_t*)a.d, (uint8_t*)o.d, cnt); ++ex_->launches; }
    vals_[n.outputs[0]] = o;
    return;
  }
  if (a.dtype == 7) {
    if (op == "Neg") {
      if (cnt) k_neg_i64<<<blocks(cnt), 256, 0, st_>>>((const int64_t*)a.d, (int64_t*)o.d, cnt);
    } else if (op == "Clip") {
      const int64_t lo = has_in(n, 1) ? host_ints(in(n, 1))[0] : INT64_MIN, hi = has_in(n, 2) ? host_ints(in(n, 2))[0] : INT64_MAX;
      if (cnt) k_clip_i64<<<blocks(cnt), 256, 0, st_>>>((const int64_t*)a.d, (int64_t*)o.d, cnt, lo, hi);
    } else {
      throw std::runtime_error("graph: " + op + " on an int64 tensor");
    }
    ++ex_->launches;
    vals_[n.outputs[0]] = o;
    return;
  }
  if (a.dtype != 1) throw std::runtime_error("graph: " + op + " on a bool tensor");
  int code;
  float p0 = 0.f, p1 = 0.f;
  if (op == "Neg") code = U_NEG; else if (op == "Exp") code = U_EXP; else if (op == "Sqrt") code = U_SQRT;
  else if (op == "Tanh") code = U_TANH; else if (op == "Sigmoid") code = U_SIGMOID; else if (op == "Softplus") code = U_SOFTPLUS;
  else if (op == "Relu") code = U_RELU; else if (op == "Erf") code = U_ERF; else if (op == "Ceil") code = U_CEIL;
  else if (op == "LeakyRelu") { code = U_LRELU; p0 = attr_f(n, "alpha", 0.01f); }
  else {
    code = U_CLIP;
    p0 = has_in(n, 1) ? (float)host_scalar(in(n, 1)) : -FLT_MAX;
    p1 = has_in(n, 2) ? (float)host_scalar(in(n, 2)) : FLT_MAX;
  }
  if (cnt) { k_unary_f32<<<blocks(cnt), 256, 0, st_>>>(code, (const float*)a.d, (float*)o.d, cnt, p0, p1); ++ex_->launches; }
  vals_[n.outputs[0]] = o;
}

// ---- data movement ----------------------------------------------------------------------------------------------
template <typename T>
static void launch_copy(cudaStream_t st, const Dims& d, const void* src, void* dst, int64_t soff, int64_t doff, int64_t n) {
  k_copy_strided<T><<<blocks(n), 256, 0, st>>>(d, (const T*)src, (T*)dst, soff, doff, n);
}
static void copy_any(cudaStream_t st, size_t esize, const Dims& d, const void* src, void* dst, int64_t soff, int64_t doff, int64_t n) {
  if (!n) return;
  if (esize == 4) launch_copy<uint32_t>(st, d, src, dst, soff, doff, n);
  else if (esize == 8) launch_copy<uint64_t>(st, d, src, dst, soff, doff, n);
  else launch_copy<uint8_t>(st, d, src, dst, soff, doff, n);
}
// a new contiguous tensor of `out_shape` whose element c reads src[src_off + c . src_strides]
GTensor GraphExec::strided_copy(const GTensor& src, const std::vector<int64_t>& out_shape, const std::vector<int64_t>& src_strides,
                                int64_t src_off) {
  GTensor o = make(src.dtype, out_shape);
  Dims d = make_dims(out_shape);
  const std::vector<int64_t> cs = contiguous_strides(out_shape);
  for (size_t a = 0; a < out_shape.size(); ++a) { d.s0[a] = src_strides[a]; d.s1[a] = cs[a]; }
  copy_any(st_, src.esize(), d, src.d, o.d, src_off, 0, o.numel());
  ++ex_->launches;
  return o;
}
// the whole of `src` (contiguous) into a block of `dst` with the given destination strides
void GraphExec::copy_into(const GTensor& src, GTensor& dst, const std::vector<int64_t>& dst_strides, int64_t dst_off) {
  Dims d = make_dims(src.shape);
  const std::vector<int64_t> cs = contiguous_strides(src.shape);
  for (size_t a = 0; a < src.shape.size(); ++a) { d.s0[a] = cs[a]; d.s1[a] = dst_strides[a]; }
  copy_any(st_, src.esize(), d, src.d, dst.d, 0, dst_off, src.numel());
  ++ex_->launches;
}

void GraphExec::op_movement(const OnnxNode& n) {
  const std::string& op = n.op_type;
  if (op == "Shape") {
    const GTensor& a = in(n, 0);
    vals_[n.outputs[0]] = upload(7, {(int64_t)a.shape.size()}, a.shape.data());
    return;
  }
  if (op == "ConstantOfShape") {
    const std::vector<int64_t> shp = host_ints(in(n, 0));
    const OnnxAttr* v = n.attr("value");
    const int dt = (v && v->has_t) ? (v->t.dtype == 1 ? 1 : (v->t.dtype == 9 ? 9 : 7)) : 1;
    GTensor o = make(dt, shp);
    const int64_t cnt = o.numel();
    if (cnt) {
      if (dt == 1) k_fill<float><<<blocks(cnt), 256, 0, st_>>>((float*)o.d, cnt, (v && v->has_t && !v->t.f32.empty()) ? v->t.f32[0] : 0.f);
      else if (dt == 7) k_fill<int64_t><<<blocks(cnt), 256, 0, st_>>>((int64_t*)o.d, cnt, v->t.i64.empty() ? 0 : v->t.i64[0]);
      else k_fill<uint8_t><<<blocks(cnt), 256, 0, st_>>>((uint8_t*)o.d, cnt, (uint8_t)(!v->t.i64.empty() && v->t.i64[0] ? 1 : 0));
      ++ex_->launches;
    }
    vals_[n.outputs[0]] = o;
    return;
  }
  if (op == "Range") {
    const GTensor& s = in(n, 0);
    if (s.dtype == 7) {
      const int64_t a = host_ints(s)[0], b = host_ints(in(n, 1))[0], st = host_ints(in(n, 2))[0];
      if (st == 0) throw std::runtime_error("graph: Range with delta 0");
      std::vector<int64_t> v;
      for (int64_t x = a; st > 0 ? x < b : x > b; x += st) v.push_back(x);
      vals_[n.outputs[0]] = upload(7, {(int64_t)v.size()}, v.data());
    } else {
      const double a = host_scalar(s), b = host_scalar(in(n, 1)), st = host_scalar(in(n, 2));
      const int64_t cnt = std::max<int64_t>(0, (int64_t)std::ceil((b - a) / st));
      std::vector<float> v((size_t)cnt);
      for (int64_t i = 0; i < cnt; ++i) v[i] = (float)(a + i * st);
      vals_[n.outputs[0]] = upload(1, {cnt}, v.data());
    }
    return;
  }
  const GTensor& a = in(n, 0);
  const int r = (int)a.shape.size();
  auto norm_axis = [](int64_t ax, int rank) {
    if (ax < 0) ax += rank;
    if (ax < 0 || ax >= rank) throw std::runtime_error("graph: axis out of range");
    return (int)ax;
  };
  if (op == "Reshape" || op == "Unsqueeze" || op == "Squeeze") {   // views: same device memory, new shape
    GTensor o = a;
    if (op == "Reshape") {
      std::vector<int64_t> shp = host_ints(in(n, 1));
      int64_t known = 1;
      int infer = -1;
      for (size_t i = 0; i < shp.size(); ++i) {
        if (shp[i] == 0) shp[i] = i < a.shape.size() ? a.shape[i] : 0;
        if (shp[i] == -1) infer = (int)i; else known *= shp[i];
      }
      if (infer >= 0) shp[infer] = known ? a.numel() / known : 0;
      o.shape = shp;
      if (o.numel() != a.numel()) throw std::runtime_error("graph: Reshape changes the element count");
    } else if (op == "Unsqueeze") {
      std::vector<int64_t> axes = has_in(n, 1) ? host_ints(in(n, 1)) : attr_ints(n, "axes");
      const int rank = r + (int)axes.size();
      for (auto& ax : axes) ax = norm_axis(ax, rank);
      std::sort(axes.begin(), axes.end());
      std::vector<int64_t> shp = a.shape;
      for (int64_t ax : axes) shp.insert(shp.begin() + ax, 1);
      o.shape = shp;
    } else {
      std::vector<int64_t> axes = has_in(n, 1) ? host_ints(in(n, 1)) : attr_ints(n, "axes");
      std::vector<int64_t> shp;
      for (int i = 0; i < r; ++i) {
        bool drop = axes.empty() ? a.shape[i] == 1 : false;
        for (int64_t ax : axes) if (norm_axis(ax, r) == i) drop = true;
        if (!drop) shp.push_back(a.shape[i]);
      }
      o.shape = shp;
    }
    vals_[n.outputs[0]] = o;
    return;
  }
  const std::vector<int64_t> cs = contiguous_strides(a.shape);
  if (op == "Transpose") {
    std::vector<int64_t> perm = attr_ints(n, "perm");
    if (perm.empty()) for (int i = r - 1; i >= 0; --i) perm.push_back(i);
    std::vector<int64_t> shp(r), ss(r);
    for (int i = 0; i < r; ++i) { shp[i] = a.shape[perm[i]]; ss[i] = cs[perm[i]]; }
    vals_[n.outputs[0]] = strided_copy(a, shp, ss, 0);
    return;
  }
  if (op == "Expand") {
    const std::vector<int64_t> os = broadcast_shape({a.shape, host_ints(in(n, 1))});
    std::vector<int64_t> ss(os.size());
    bstrides(a.shape, os, ss.data());
    vals_[n.outputs[0]] = strided_copy(a, os, ss, 0);
    return;
  }
  if (op == "Slice") {   // GraphExecutor.swift:940-1098: clamped starts / ends, negative steps
    const std::vector<int64_t> starts = host_ints(in(n, 1)), ends = host_ints(in(n, 2));
    std::vector<int64_t> axes, steps;
    if (has_in(n, 3)) axes = host_ints(in(n, 3)); else for (size_t i = 0; i < starts.size(); ++i) axes.push_back((int64_t)i);
    if (has_in(n, 4)) steps = host_ints(in(n, 4)); else steps.assign(starts.size(), 1);
    std::vector<int64_t> shp = a.shape, ss = cs;
    int64_t off = 0;
    for (size_t i = 0; i < starts.size(); ++i) {
      const int ax = norm_axis(axes[i], r);
      const int64_t dim = a.shape[ax], st = steps[i];
      int64_t s = starts[i], e = ends[i];
      if (st == 0) throw std::runtime_error("graph: Slice with step 0");
      if (s < 0) s += dim;
      if (e < 0) e += dim;
      int64_t cnt;
      if (st > 0) {
        s = std::min(std::max<int64_t>(s, 0), dim);
        e = std::min(std::max<int64_t>(e, 0), dim);
        cnt = e > s ? (e - s + st - 1) / st : 0;
      } else {
        s = std::min(std::max<int64_t>(s, -1), dim - 1);
        e = std::min(std::max<int64_t>(e, -1), dim - 1);
        cnt = s > e ? (s - e + (-st) - 1) / (-st) : 0;
      }
      shp[ax] = cnt;
      off += s * cs[ax];
      ss[ax] = cs[ax] * st;
    }
    vals_[n.outputs[0]] = strided_copy(a, shp, ss, off);
    return;
  }
  if (op == "Concat") {
    const int ax = norm_axis(attr_i(n, "axis", 0), r);
    std::vector<int64_t> shp = a.shape;
    shp[ax] = 0;
    for (size_t i = 0; i < n.inputs.size(); ++i) {
      const GTensor& t = in(n, i);
      if (t.dtype != a.dtype || (int)t.shape.size() != r) throw std::runtime_error("graph: Concat operands differ in type or rank");
      shp[ax] += t.shape[ax];
    }
    GTensor o = make(a.dtype, shp);
    const std::vector<int64_t> os = contiguous_strides(shp);
    int64_t pos = 0;
    for (size_t i = 0; i < n.inputs.size(); ++i) {
      const GTensor& t = in(n, i);
      if (t.numel()) copy_into(t, o, os, pos * os[ax]);
      pos += t.shape[ax];
    }
    vals_[n.outputs[0]] = o;
    return;
  }
  if (op == "Split") {
    const int ax = norm_axis(attr_i(n, "axis", 0), r);
    std::vector<int64_t> sizes = has_in(n, 1) ? host_ints(in(n, 1)) : attr_ints(n, "split");
    if (sizes.empty()) sizes.assign(n.outputs.size(), a.shape[ax] / (int64_t)n.outputs.size());
    int64_t pos = 0;
    for (size_t i = 0; i < n.outputs.size() && i < sizes.size(); ++i) {
      std::vector<int64_t> shp = a.shape;
      shp[ax] = sizes[i];
      if (!n.outputs[i].empty()) vals_[n.outputs[i]] = strided_copy(a, shp, cs, pos * cs[ax]);
      pos += sizes[i];
    }
    return;
  }
  if (op == "Pad") {   // constant mode (GraphExecutor.swift:1602-1700)
    const std::vector<int64_t> pads = host_ints(in(n, 1));
    if ((int)pads.size() != 2 * r) throw std::runtime_error("graph: Pad needs 2 * rank pads");
    std::vector<int64_t> shp(r);
    for (int i = 0; i < r; ++i) shp[i] = a.shape[i] + pads[i] + pads[i + r];
    GTensor o = make(a.dtype, shp);
    const int64_t cnt = o.numel();
    if (cnt) {
      if (a.dtype == 1) k_fill<float><<<blocks(cnt), 256, 0, st_>>>((float*)o.d, cnt, has_in(n, 2) && in(n, 2).numel() ? (float)host_scalar(in(n, 2)) : 0.f);
      else if (a.dtype == 7) k_fill<int64_t><<<blocks(cnt), 256, 0, st_>>>((int64_t*)o.d, cnt, has_in(n, 2) && in(n, 2).numel() ? host_ints(in(n, 2))[0] : 0);
      else k_fill<uint8_t><<<blocks(cnt), 256, 0, st_>>>((uint8_t*)o.d, cnt, 0);
      ++ex_->launches;
    }
    const std::vector<int64_t> os = contiguous_strides(shp);
    int64_t off = 0;
    bool negative = false;
    for (int i = 0; i < r; ++i) { off += pads[i] * os[i]; negative |= pads[i] < 0 || pads[i + r] < 0; }
    if (negative) throw std::runtime_error("unsupported: Pad with negative pads");
    if (a.numel()) copy_into(a, o, os, off);
    vals_[n.outputs[0]] = o;
    return;
  }
  throw std::runtime_error("unsupported: op '" + op + "'");
}

// ---- indexing ---------------------------------------------------------------------------------------------------
#define DISPATCH_T(dt, CALL)                       \
  do {                                             \
    if ((dt) == 1) { using T = float; CALL; }      \
    else if ((dt) == 7) { using T = int64_t; CALL; } \
    else { using T = uint8_t; CALL; }              \
  } while (0)

void GraphExec::op_index(const OnnxNode& n) {
  const std::string& op = n.op_type;
  const GTensor& a = in(n, 0);
  const int r = (int)a.shape.size();
  const std::vector<int64_t> cs = contiguous_strides(a.shape);
  if (op == "NonZero") {   // nonzero.metal: row-major scan; the count is data-dependent, so the (small) tensor goes to the host
    const std::vector<int64_t> v = host_ints(a);
    std::vector<std::vector<int64_t>> coords((size_t)r);
    for (int64_t i = 0; i < (int64_t)v.size(); ++i) {
      if (!v[i]) continue;
      int64_t rem = i;
      for (int ax = 0; ax < r; ++ax) { coords[ax].push_back(rem / cs[ax]); rem %= cs[ax]; }
    }
    const int64_t cnt = r ? (int64_t)coords[0].size() : 0;
    std::vector<int64_t> flat;
    for (int ax = 0; ax < r; ++ax) flat.insert(flat.end(), coords[ax].begin(), coords[ax].end());
    vals_[n.outputs[0]] = upload(7, {(int64_t)r, cnt}, flat.data());
    return;
  }
  const GTensor& idx = in(n, 1);
  if (idx.dtype != 7) throw std::runtime_error("graph: " + op + " indices must be int64");
  if (op == "Gather") {
    int64_t ax = attr_i(n, "axis", 0);
    if (ax < 0) ax += r;
    if (ax < 0 || ax >= r) throw std::runtime_error("graph: Gather axis out of range");
    int64_t outer = 1, inner = 1;
    for (int i = 0; i < ax; ++i) outer *= a.shape[i];
    for (int i = (int)ax + 1; i < r; ++i) inner *= a.shape[i];
    std::vector<int64_t> shp(a.shape.begin(), a.shape.begin() + ax);
    shp.insert(shp.end(), idx.shape.begin(), idx.shape.end());
    shp.insert(shp.end(), a.shape.begin() + ax + 1, a.shape.end());
    GTensor o = make(a.dtype, shp);
    const int64_t cnt = o.numel();
    if (cnt) {
      DISPATCH_T(a.dtype, (k_gather<T><<<blocks(cnt), 256, 0, st_>>>((const T*)a.d, (const int64_t*)idx.d, (T*)o.d, outer, a.shape[ax], inner, idx.numel())));
      ++ex_->launches;
    }
    vals_[n.outputs[0]] = o;
    return;
  }
  if (op == "GatherElements") {
    int64_t ax = attr_i(n, "axis", 0);
    if (ax < 0) ax += r;
    if ((int)idx.shape.size() != r) throw std::runtime_error("graph: GatherElements rank mismatch");
    GTensor o = make(a.dtype, idx.shape);
    Dims d = make_dims(idx.shape);
    for (int i = 0; i < r; ++i) d.s0[i] = i == ax ? 0 : cs[i];
    const int64_t cnt = o.numel();
    if (cnt) {
      DISPATCH_T(a.dtype, (k_gather_elements<T><<<blocks(cnt), 256, 0, st_>>>(d, (const T*)a.d, (const int64_t*)idx.d, (T*)o.d, cs[ax], a.shape[ax], cnt)));
      ++ex_->launches;
    }
    vals_[n.outputs[0]] = o;
    return;
  }
  // GatherND / ScatterND: the last axis of idx addresses the first k axes of the data
  if (idx.shape.empty()) throw std::runtime_error("graph: " + op + " indices need a last axis");
  const int k = (int)idx.shape.back();
  if (k > r) throw std::runtime_error("graph: " + op + " index depth exceeds the data rank");
  int64_t slice = 1;
  for (int i = k; i < r; ++i) slice *= a.shape[i];
  const int64_t n_idx = k ? idx.numel() / k : 0;
  Dims d = make_dims(a.shape);
  for (int i = 0; i < r; ++i) d.s0[i] = cs[i];
  if (op == "GatherND") {
    std::vector<int64_t> shp(idx.shape.begin(), idx.shape.end() - 1);
    shp.insert(shp.end(), a.shape.begin() + k, a.shape.end());
    GTensor o = make(a.dtype, shp);
    const int64_t cnt = n_idx * slice;
    if (cnt) {
      DISPATCH_T(a.dtype, (k_gathernd<T><<<blocks(cnt), 256, 0, st_>>>((const T*)a.d, (const int64_t*)idx.d, (T*)o.d, d, k, slice, cnt)));
      ++ex_->launches;
    }
    vals_[n.outputs[0]] = o;
    return;
  }
  const GTensor& upd = in(n, 2);
  if (upd.dtype != a.dtype) throw std::runtime_error("graph: ScatterND updates differ in type from the data");
  GTensor o = make(a.dtype, a.shape);
  if (a.bytes()) G_CUDA_OK(cudaMemcpyAsync(o.d, a.d, a.bytes(), cudaMemcpyDeviceToDevice, st_));
  const int64_t cnt = n_idx * slice;
  if (cnt) {
    if (upd.numel() != cnt) throw std::runtime_error("graph: ScatterND updates do not match the indices");
    DISPATCH_T(a.dtype, (k_scatternd<T><<<blocks(cnt), 256, 0, st_>>>((T*)o.d, (const int64_t*)idx.d, (const T*)upd.d, d, k, slice, cnt)));
    ++ex_->launches;
  }
  vals_[n.outputs[0]] = o;
}

// ---- reductions -------------------------------------------------------------------------------------------------
void GraphExec::op_reduce(const OnnxNode& n) {
  const std::string& op = n.op_type;
  const GTensor& a = in(n, 0);
  const int r = (int)a.shape.size();
  auto split3 = [&](int lo, int hi, int64_t& outer, int64_t& mid, int64_t& inner) {   // axes [lo, hi] form the middle block
    outer = mid = inner = 1;
    for (int i = 0; i < lo; ++i) outer *= a.shape[i];
    for (int i = lo; i <= hi; ++i) mid *= a.shape[i];
    for (int i = hi + 1; i < r; ++i) inner *= a.shape[i];
  };
  if (op == "Softmax" || op == "CumSum") {
    int64_t ax = op == "Softmax" ? attr_i(n, "axis", -1) : host_ints(in(n, 1))[0];
    if (ax < 0) ax += r;
    if (ax < 0 || ax >= r) throw std::runtime_error("graph: " + op + " axis out of range");
    int64_t outer, mid, inner;
    split3((int)ax, (int)ax, outer, mid, inner);
    GTensor o = make(a.dtype, a.shape);
    if (a.numel()) {
      if (op == "Softmax") {
        if (a.dtype != 1) throw std::runtime_error("graph: Softmax on a non-float tensor");
        k_softmax<<<blocks(outer * inner), 256, 0, st_>>>((const float*)a.d, (float*)o.d, outer, mid, inner);
      } else {
        if (attr_i(n, "exclusive", 0) || attr_i(n, "reverse", 0)) throw std::runtime_error("unsupported: CumSum exclusive / reverse");
        if (a.dtype == 1) k_cumsum<float><<<blocks(outer * inner), 256, 0, st_>>>((const float*)a.d, (float*)o.d, outer, mid, inner);
        else if (a.dtype == 7) k_cumsum<int64_t><<<blocks(outer * inner), 256, 0, st_>>>((const int64_t*)a.d, (int64_t*)o.d, outer, mid, inner);
        else throw std::runtime_error("graph: CumSum on a bool tensor");
      }
      ++ex_->launches;
    }
    vals_[n.outputs[0]] = o;
    return;
  }
  std::vector<int64_t> axes = has_in(n, 1) ? host_ints(in(n, 1)) : attr_ints(n, "axes");
  const bool keep = attr_i(n, "keepdims", 1) != 0;
  if (axes.empty()) for (int i = 0; i < r; ++i) axes.push_back(i);
  for (auto& ax : axes) { if (ax < 0) ax += r; if (ax < 0 || ax >= r) throw std::runtime_error("graph: reduce axis out of range"); }
  std::sort(axes.begin(), axes.end());
  for (size_t i = 1; i < axes.size(); ++i)
    if (axes[i] != axes[i - 1] + 1) throw std::runtime_error("unsupported: reduction over non-adjacent axes");
  int64_t outer, mid, inner;
  split3((int)axes.front(), (int)axes.back(), outer, mid, inner);
  std::vector<int64_t> shp;
  for (int i = 0; i < r; ++i) {
    const bool red = i >= axes.front() && i <= axes.back();
    if (!red) shp.push_back(a.shape[i]); else if (keep) shp.push_back(1);
  }
  GTensor o = make(a.dtype, shp);
  const int code = op == "ReduceMean" ? R_MEAN : (op == "ReduceSum" ? R_SUM : R_MAX);
  if (mid == 0) throw std::runtime_error("graph: reduction over an empty axis");
  if (o.numel()) {
    if (a.dtype == 1) k_reduce<float><<<blocks(outer * inner), 256, 0, st_>>>(code, (const float*)a.d, (float*)o.d, outer, mid, inner);
    else if (a.dtype == 7) k_reduce<int64_t><<<blocks(outer * inner), 256, 0, st_>>>(code, (const int64_t*)a.d, (int64_t*)o.d, outer, mid, inner);
    else throw std::runtime_error("graph: reduction of a bool tensor");
    ++ex_->launches;
  }
  vals_[n.outputs[0]] = o;
}

// ---- MatMul, Conv, ConvTranspose, RandomNormalLike ------------------------------------------------------------------------
void GraphExec::op_linear(const OnnxNode& n) {
  const std::string& op = n.op_type;
  const GTensor& a = in(n, 0);
  if (a.dtype != 1) throw std::runtime_error("graph: " + op + " needs float32 input");
  if (op == "RandomNormalLike") {   // elementwise.metal:132-163, both noise nodes of a run share the seed (GraphExecutor.swift:2657)
    GTensor o = make(1, a.shape);
    const int64_t cnt = o.numel();
    if (cnt) {
      launch_random_normal(*ex_, (float*)o.d, cnt, seed_);
      const float scale = attr_f(n, "scale", 1.f), mean = attr_f(n, "mean", 0.f);
      if (scale != 1.f || mean != 0.f) { k_scale_shift<<<blocks(cnt), 256, 0, st_>>>((float*)o.d, cnt, scale, mean); ++ex_->launches; }
    }
    vals_[n.outputs[0]] = o;
    return;
  }
  const GTensor& b = in(n, 1);
  if (b.dtype != 1) throw std::runtime_error("graph: " + op + " needs float32 weights");
  if (op == "MatMul") {   // matmul.metal:22-49 with numpy batch broadcasting
    if (a.shape.size() < 2 || b.shape.size() < 2) throw std::runtime_error("unsupported: MatMul on rank-1 operands");
    const int M = (int)a.shape[a.shape.size() - 2], K = (int)a.shape.back(), N = (int)b.shape.back();
    if (b.shape[b.shape.size() - 2] != K) throw std::runtime_error("graph: MatMul inner dimensions differ");
    const std::vector<int64_t> ba(a.shape.begin(), a.shape.end() - 2), bb(b.shape.begin(), b.shape.end() - 2);
    const std::vector<int64_t> bo = broadcast_shape({ba, bb});
    Dims d = make_dims(bo);
    bstrides(ba, bo, d.s0); bstrides(bb, bo, d.s1);
    for (size_t i = 0; i < bo.size(); ++i) { d.s0[i] *= (int64_t)M * K; d.s1[i] *= (int64_t)K * N; }
    std::vector<int64_t> shp = bo;
    shp.push_back(M); shp.push_back(N);
    GTensor o = make(1, shp);
    int64_t batch = 1;
    for (int64_t v : bo) batch *= v;
    if (o.numel()) {
      k_matmul<<<blocks(o.numel()), 256, 0, st_>>>(d, (const float*)a.d, (const float*)b.d, (float*)o.d, batch, M, N, K);
      ++ex_->launches;
    }
    vals_[n.outputs[0]] = o;
    return;
  }
  if (a.shape.size() != 3 || b.shape.size() != 3) throw std::runtime_error("unsupported: " + op + " on non-1D data");
  const float* bias = has_in(n, 2) ? (const float*)in(n, 2).d : nullptr;
  const int N = (int)a.shape[0], Cin = (int)a.shape[1], L = (int)a.shape[2], K = (int)b.shape[2];
  const int stride = (int)attr_i(n, "strides", 1), dil = (int)attr_i(n, "dilations", 1), group = (int)attr_i(n, "group", 1);
  const std::vector<int64_t> pads = attr_ints(n, "pads");
  const int pl = pads.size() > 0 ? (int)pads[0] : 0, pr = pads.size() > 1 ? (int)pads[1] : pl;
  if (op == "Conv") {
    const int Cout = (int)b.shape[0];
    if ((int)b.shape[1] * group != Cin) throw std::runtime_error("graph: Conv channels do not match the weight");
    const int Lout = (L + pl + pr - dil * (K - 1) - 1) / stride + 1;
    GTensor o = make(1, {N, Cout, std::max(Lout, 0)});
    if (o.numel()) {
      k_conv1d<<<blocks(o.numel()), 256, 0, st_>>>((const float*)a.d, (const float*)b.d, bias, (float*)o.d, N, Cin, L, Cout, K, stride, dil, pl,
                                                   group, Lout);
      ++ex_->launches;
    }
    vals_[n.outputs[0]] = o;
    return;
  }
  if ((int)b.shape[0] != Cin) throw std::runtime_error("graph: ConvTranspose channels do not match the weight");
  const int Cout = (int)b.shape[1] * group;
  const int opad = (int)attr_i(n, "output_padding", 0);
  const int Lout = (L - 1) * stride + dil * (K - 1) + 1 + opad - pl - pr;
  GTensor o = make(1, {N, Cout, std::max(Lout, 0)});
  if (o.numel()) {
    k_convtr1d<<<blocks(o.numel()), 256, 0, st_>>>((const float*)a.d, (const float*)b.d, bias, (float*)o.d, N, Cin, L, Cout, K, stride, dil, pl,
                                                   group, Lout);
    ++ex_->launches;
  }
  vals_[n.outputs[0]] = o;
}

}  // namespace piper
