// STUB of the slice of the TensorFlow 1.15 C++ op API that
// integration/tf_ops/b200_range_coding_kernels.cc uses.  TensorFlow cannot be installed in this
// image, so this header exists ONLY so that the kernel-registration shell is
// compiled by a test (tests/test_cabi_cpu.py) instead of rotting in a markdown block.  It
// declares names and signatures, no behaviour; build the shell against the real headers
// (`tf.sysconfig.get_compile_flags()`) to use it.
#ifndef TFC_B200_TF_STUB_OP_KERNEL_H_
#define TFC_B200_TF_STUB_OP_KERNEL_H_

#include <cstdint>
#include <initializer_list>
#include <string>
#include <vector>

namespace tensorflow {

typedef int64_t int64;
typedef int32_t int32;
typedef int16_t int16;
typedef uint8_t uint8;
typedef std::string tstring;

enum DataType { DT_FLOAT, DT_INT16, DT_INT32, DT_UINT8, DT_STRING };

class Status {
 public:
  Status() {}
  explicit Status(const std::string& m) : msg_(m), ok_(false) {}
  bool ok() const { return ok_; }
  static Status OK() { return Status(); }
 private:
  std::string msg_;
  bool ok_ = true;
};

namespace errors {
template <typename... Args>
Status InvalidArgument(const Args&...) { return Status("invalid argument"); }
template <typename... Args>
Status Internal(const Args&...) { return Status("internal"); }
}  // namespace errors

class TensorShape {
 public:
  TensorShape() {}
  TensorShape(std::initializer_list<int64> d) : dims_(d) {}
  int dims() const { return (int)dims_.size(); }
  int64 dim_size(int i) const { return dims_[i]; }
  void AddDim(int64 d) { dims_.push_back(d); }
  int64 num_elements() const { int64 n = 1; for (int64 d : dims_) n *= d; return n; }
 private:
  std::vector<int64> dims_;
};

template <typename T>
struct FlatView {
  T* data() const { return ptr; }
  T& operator()(int64 i) const { return ptr[i]; }
  T& operator()() const { return ptr[0]; }
  T* ptr;
};

class Tensor {
 public:
  const TensorShape& shape() const { return shape_; }
  int dims() const { return shape_.dims(); }
  int64 dim_size(int i) const { return shape_.dim_size(i); }
  int64 NumElements() const { return shape_.num_elements(); }
  template <typename T> FlatView<T> flat() { return FlatView<T>{static_cast<T*>(buf_)}; }
  template <typename T> FlatView<const T> flat() const {
    return FlatView<const T>{static_cast<const T*>(buf_)};
  }
  template <typename T> FlatView<T> scalar() { return flat<T>(); }
  template <typename T> FlatView<const T> scalar() const { return flat<T>(); }
 private:
  TensorShape shape_;
  void* buf_ = nullptr;
};

struct GpuDevice {
  void* stream() const { return nullptr; }   // cudaStream_t in the real header
};

class OpKernelConstruction {
 public:
  template <typename T> Status GetAttr(const char*, T*) const { return Status::OK(); }
  void SetStatus(const Status&) {}
};

class OpKernelContext {
 public:
  const Tensor& input(int) { return dummy_; }
  Status allocate_output(int, const TensorShape&, Tensor** out) { *out = &dummy_; return Status::OK(); }
  Status allocate_temp(DataType, const TensorShape&, Tensor* out) { (void)out; return Status::OK(); }
  const GpuDevice& eigen_gpu_device() const { return dev_; }
  void SetStatus(const Status&) {}
 private:
  Tensor dummy_;
  GpuDevice dev_;
};

class OpKernel {
 public:
  explicit OpKernel(OpKernelConstruction*) {}
  virtual ~OpKernel() {}
  virtual void Compute(OpKernelContext* ctx) = 0;
};

constexpr const char* DEVICE_GPU = "GPU";

class Name {
 public:
  explicit Name(const char*) {}
  Name& Device(const char*) { return *this; }
  Name& HostMemory(const char*) { return *this; }
};

namespace stub {
struct Registrar { Registrar(const Name&, OpKernel* (*)(OpKernelConstruction*)) {} };
}  // namespace stub

#define TFC_STUB_CAT_(a, b) a##b
#define TFC_STUB_CAT(a, b) TFC_STUB_CAT_(a, b)
#define REGISTER_KERNEL_BUILDER(builder, ...)                                        \
  static ::tensorflow::stub::Registrar TFC_STUB_CAT(tfc_stub_registrar_, __LINE__)(  \
      builder, [](::tensorflow::OpKernelConstruction* c) -> ::tensorflow::OpKernel* { \
        return new __VA_ARGS__(c); })

#define OP_REQUIRES(CTX, EXP, STATUS) \
  do { if (!(EXP)) { (CTX)->SetStatus(STATUS); return; } } while (0)
#define OP_REQUIRES_OK(CTX, ...) \
  do { ::tensorflow::Status _s(__VA_ARGS__); if (!_s.ok()) { (CTX)->SetStatus(_s); return; } } while (0)

}  // namespace tensorflow

#endif  // TFC_B200_TF_STUB_OP_KERNEL_H_
