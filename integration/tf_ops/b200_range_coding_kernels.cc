// GPU kernel registrations of the reference's five entropy-coding ops on top of libtfc_b200.so
// (include/tfc_b200.h).  The op SIGNATURES are the reference's own
// (cc/ops/range_coding_ops.cc:29-35 RangeEncode, :91-98 RangeDecode, :125-135
// UnboundedIndexRangeEncode, :203-213 UnboundedIndexRangeDecode, :244-248 PmfToQuantizedCdf) and
// are not re-registered here: a maintainer adds this file to the op library next to the CPU
// kernels (cc/kernels/range_coding_kernels.cc, unbounded_index_range_coding_kernels.cc,
// range_coding_helper_kernels.cc) and the existing Python wrappers pick the GPU kernel when
// the inputs live on a GPU.
//
// Every op is one stream of the batched C entry points (S = 1).  The `encoded` string is a host
// scalar in the reference's signature, so the encode kernels copy length and bytes back on the
// op's stream and the decode kernels upload the string; the layer classes of this repo call the
// batched device-resident forms instead and never cross PCIe (python/ops/range_coding_ops.py).
//
// Compiled in this repo only against integration/tf_stub (TensorFlow 1.15 cannot be installed in
// the image); tests/test_cabi_cpu.py::test_tf_op_shell_compiles keeps it compiling.
#include <cuda_runtime_api.h>

#include <cstring>
#include <vector>

#include "tensorflow/core/framework/op_kernel.h"
#include "tensorflow/core/framework/tensor.h"
#include "tensorflow/core/framework/tensor_shape.h"
#include "tensorflow/core/lib/core/errors.h"
#include "tfc_b200.h"

namespace tensorflow_compression {
namespace b200 {

using tensorflow::DEVICE_GPU;
using tensorflow::OpKernel;
using tensorflow::OpKernelConstruction;
using tensorflow::OpKernelContext;
using tensorflow::Tensor;
using tensorflow::TensorShape;
using tensorflow::int16;
using tensorflow::int32;
using tensorflow::int64;
using tensorflow::tstring;
using tensorflow::uint8;
namespace errors = tensorflow::errors;

namespace {

// device scratch of one single-stream encode: [cap bytes | nbytes | status]
struct EncodeScratch {
  Tensor bytes, words;
  uint8* out() { return bytes.flat<uint8>().data(); }
  int32* nbytes() { return words.flat<int32>().data(); }
  int32* status() { return words.flat<int32>().data() + 1; }
};

tensorflow::Status AllocateScratch(OpKernelContext* ctx, int64 cap, EncodeScratch* s) {
  tensorflow::Status st = ctx->allocate_temp(tensorflow::DT_UINT8, TensorShape({cap}), &s->bytes);
  if (!st.ok()) return st;
  return ctx->allocate_temp(tensorflow::DT_INT32, TensorShape({2}), &s->words);
}

// length + status to the host (one small copy), then the string itself
template <typename Launch>
void EncodeToHostString(OpKernelContext* ctx, int64 symbols, Launch launch) {
  cudaStream_t stream = static_cast<cudaStream_t>(ctx->eigen_gpu_device().stream());
  Tensor* encoded = nullptr;
  OP_REQUIRES_OK(ctx, ctx->allocate_output(0, TensorShape({}), &encoded));
  int64 cap = 2 * symbols + 64;            // enough unless escapes dominate; retried once below
  for (int attempt = 0; attempt < 2; ++attempt) {
    EncodeScratch s;
    OP_REQUIRES_OK(ctx, AllocateScratch(ctx, cap, &s));
    const int rc = launch(s.out(), cap, s.nbytes(), s.status(), stream);
    OP_REQUIRES(ctx, rc == TFC_OK, errors::InvalidArgument(tfc_last_error()));
    int32 host[2] = {0, 0};
    OP_REQUIRES(ctx,
                cudaMemcpyAsync(host, s.nbytes(), sizeof(host), cudaMemcpyDeviceToHost, stream) ==
                        cudaSuccess &&
                    cudaStreamSynchronize(stream) == cudaSuccess,
                errors::Internal("copy of the encoded length failed"));
    if (host[1] == TFC_STREAM_CAPACITY && attempt == 0) {
      cap = (int64)host[0] + 16;           // the kernel reports the size it needed
      continue;
    }
    OP_REQUIRES(ctx, host[1] == TFC_STREAM_OK, errors::Internal("range encoder status ", host[1]));
    tstring* dst = &encoded->scalar<tstring>()();
    dst->resize(host[0]);
    if (host[0] > 0) {
      OP_REQUIRES(ctx,
                  cudaMemcpyAsync(&(*dst)[0], s.out(), host[0], cudaMemcpyDeviceToHost, stream) ==
                          cudaSuccess &&
                      cudaStreamSynchronize(stream) == cudaSuccess,
                  errors::Internal("copy of the encoded string failed"));
    }
    return;
  }
}

// the host string onto the device, padded to 16 bytes: [bytes | nbytes | status]
struct DecodeScratch {
  Tensor bytes, words;
  int64 cap = 0;
};

tensorflow::Status UploadString(OpKernelContext* ctx, const tstring& encoded, DecodeScratch* s) {
  cudaStream_t stream = static_cast<cudaStream_t>(ctx->eigen_gpu_device().stream());
  s->cap = ((int64)encoded.size() + 15) / 16 * 16 + 16;
  tensorflow::Status st = ctx->allocate_temp(tensorflow::DT_UINT8, TensorShape({s->cap}), &s->bytes);
  if (!st.ok()) return st;
  st = ctx->allocate_temp(tensorflow::DT_INT32, TensorShape({2}), &s->words);
  if (!st.ok()) return st;
  const int32 host[2] = {(int32)encoded.size(), 0};
  if (cudaMemcpyAsync(s->words.flat<int32>().data(), host, sizeof(host), cudaMemcpyHostToDevice,
                      stream) != cudaSuccess ||
      (!encoded.empty() &&
       cudaMemcpyAsync(s->bytes.flat<uint8>().data(), encoded.data(), encoded.size(),
                       cudaMemcpyHostToDevice, stream) != cudaSuccess) ||
      cudaStreamSynchronize(stream) != cudaSuccess)   // `host` and the string must outlive the copy
    return errors::Internal("upload of the encoded string failed");
  return tensorflow::Status::OK();
}

// a decoder that ran past the end of its table: the reference CHECK-aborts
// (cc/kernels/range_coder.cc:347); here it is an op error
tensorflow::Status CheckDecodeStatus(OpKernelContext* ctx, DecodeScratch* s) {
  cudaStream_t stream = static_cast<cudaStream_t>(ctx->eigen_gpu_device().stream());
  int32 status = 0;
  if (cudaMemcpyAsync(&status, s->words.flat<int32>().data() + 1, sizeof(status),
                      cudaMemcpyDeviceToHost, stream) != cudaSuccess ||
      cudaStreamSynchronize(stream) != cudaSuccess)
    return errors::Internal("copy of the decoder status failed");
  if (status != TFC_STREAM_OK) return errors::InvalidArgument("corrupt range-coded string");
  return tensorflow::Status::OK();
}

std::vector<int64_t> Dims(const TensorShape& shape) {
  std::vector<int64_t> d(shape.dims());
  for (int i = 0; i < shape.dims(); ++i) d[i] = shape.dim_size(i);
  return d;
}

}  // namespace

// ---- UnboundedIndexRangeEncode (cc/kernels/unbounded_index_range_coding_kernels.cc:130-251) ----
class B200UnboundedIndexRangeEncodeOp : public OpKernel {
 public:
  explicit B200UnboundedIndexRangeEncodeOp(OpKernelConstruction* c) : OpKernel(c) {
    OP_REQUIRES_OK(c, c->GetAttr("precision", &precision_));
    OP_REQUIRES_OK(c, c->GetAttr("overflow_width", &overflow_width_));
    OP_REQUIRES_OK(c, c->GetAttr("debug_level", &debug_level_));
  }
  void Compute(OpKernelContext* ctx) override {
    const Tensor& data = ctx->input(0);
    const Tensor& index = ctx->input(1);
    const Tensor& cdf = ctx->input(2);
    const Tensor& cdf_size = ctx->input(3);
    const Tensor& offset = ctx->input(4);
    OP_REQUIRES(ctx, data.NumElements() == index.NumElements(),
                errors::InvalidArgument("`data` and `index` should have the same shape"));
    OP_REQUIRES(ctx, cdf.dims() == 2, errors::InvalidArgument("`cdf` should be 2-D."));
    OP_REQUIRES(ctx, cdf_size.dims() == 1 && cdf_size.dim_size(0) == cdf.dim_size(0),
                errors::InvalidArgument("`cdf_size` should be 1-D and match `cdf`"));
    OP_REQUIRES(ctx, offset.dims() == 1 && offset.dim_size(0) == cdf.dim_size(0),
                errors::InvalidArgument("`offset` should be 1-D and match `cdf`"));
    const int64 n = data.NumElements();
    EncodeToHostString(ctx, n, [&](uint8* out, int64 cap, int32* nbytes, int32* status,
                                   cudaStream_t stream) {
      return tfc_unbounded_index_range_encode(
          data.flat<int32>().data(), index.flat<int32>().data(), /*index_stride=*/n, /*S=*/1, n,
          cdf.flat<int32>().data(), (int32_t)cdf.dim_size(0), (int32_t)cdf.dim_size(1),
          cdf_size.flat<int32>().data(), offset.flat<int32>().data(), precision_, overflow_width_,
          debug_level_, out, cap, nbytes, status, stream);
    });
  }
 private:
  int precision_ = 0, overflow_width_ = 0, debug_level_ = 1;
};

// ---- UnboundedIndexRangeDecode (cc/kernels/unbounded_index_range_coding_kernels.cc:256-368) ----
class B200UnboundedIndexRangeDecodeOp : public OpKernel {
 public:
  explicit B200UnboundedIndexRangeDecodeOp(OpKernelConstruction* c) : OpKernel(c) {
    OP_REQUIRES_OK(c, c->GetAttr("precision", &precision_));
    OP_REQUIRES_OK(c, c->GetAttr("overflow_width", &overflow_width_));
    OP_REQUIRES_OK(c, c->GetAttr("debug_level", &debug_level_));
  }
  void Compute(OpKernelContext* ctx) override {
    const Tensor& encoded = ctx->input(0);           // HostMemory
    const Tensor& index = ctx->input(1);
    const Tensor& cdf = ctx->input(2);
    const Tensor& cdf_size = ctx->input(3);
    const Tensor& offset = ctx->input(4);
    OP_REQUIRES(ctx, encoded.dims() == 0, errors::InvalidArgument("`encoded` should be a scalar."));
    OP_REQUIRES(ctx, cdf.dims() == 2, errors::InvalidArgument("`cdf` should be 2-D."));
    OP_REQUIRES(ctx, cdf_size.dims() == 1 && cdf_size.dim_size(0) == cdf.dim_size(0),
                errors::InvalidArgument("`cdf_size` should be 1-D and match `cdf`"));
    OP_REQUIRES(ctx, offset.dims() == 1 && offset.dim_size(0) == cdf.dim_size(0),
                errors::InvalidArgument("`offset` should be 1-D and match `cdf`"));
    Tensor* decoded = nullptr;
    OP_REQUIRES_OK(ctx, ctx->allocate_output(0, index.shape(), &decoded));
    DecodeScratch s;
    OP_REQUIRES_OK(ctx, UploadString(ctx, encoded.scalar<tstring>()(), &s));
    const int64 n = index.NumElements();
    int32* words = s.words.flat<int32>().data();
    const int rc = tfc_unbounded_index_range_decode(
        s.bytes.flat<uint8>().data(), s.cap, words, index.flat<int32>().data(), n, 1, n,
        cdf.flat<int32>().data(), (int32_t)cdf.dim_size(0), (int32_t)cdf.dim_size(1),
        cdf_size.flat<int32>().data(), offset.flat<int32>().data(), precision_, overflow_width_,
        debug_level_, decoded->flat<int32>().data(), words + 1, ctx->eigen_gpu_device().stream());
    OP_REQUIRES(ctx, rc == TFC_OK, errors::InvalidArgument(tfc_last_error()));
    OP_REQUIRES_OK(ctx, CheckDecodeStatus(ctx, &s));
  }
 private:
  int precision_ = 0, overflow_width_ = 0, debug_level_ = 1;
};

// ---- RangeEncode (cc/kernels/range_coding_kernels.cc:178-270) ------------------------------------
class B200RangeEncodeOp : public OpKernel {
 public:
  explicit B200RangeEncodeOp(OpKernelConstruction* c) : OpKernel(c) {
    OP_REQUIRES_OK(c, c->GetAttr("precision", &precision_));
    OP_REQUIRES_OK(c, c->GetAttr("debug_level", &debug_level_));
  }
  void Compute(OpKernelContext* ctx) override {
    const Tensor& data = ctx->input(0);
    const Tensor& cdf = ctx->input(1);
    const std::vector<int64_t> ds = Dims(data.shape()), cs = Dims(cdf.shape());
    EncodeToHostString(ctx, data.NumElements(), [&](uint8* out, int64 cap, int32* nbytes,
                                                    int32* status, cudaStream_t stream) {
      // the shape checks of CheckCdfShape / MergeAxes (range_coding_kernels_util.cc:36-92) are
      // made by the library and come back as TFC_ERR_INVALID_ARGUMENT with the same wording
      return tfc_range_encode(data.flat<int16>().data(), /*S=*/1, ds.data(), (int)ds.size(),
                              cdf.flat<int32>().data(), cs.data(), (int)cs.size(),
                              /*cdf_stream_stride=*/0, precision_, debug_level_, out, cap, nbytes,
                              status, stream);
    });
  }
 private:
  int precision_ = 0, debug_level_ = 1;
};

// ---- RangeDecode (cc/kernels/range_coding_kernels.cc:272-380) ------------------------------------
class B200RangeDecodeOp : public OpKernel {
 public:
  explicit B200RangeDecodeOp(OpKernelConstruction* c) : OpKernel(c) {
    OP_REQUIRES_OK(c, c->GetAttr("precision", &precision_));
    OP_REQUIRES_OK(c, c->GetAttr("debug_level", &debug_level_));
  }
  void Compute(OpKernelContext* ctx) override {
    const Tensor& encoded = ctx->input(0);           // HostMemory
    const Tensor& shape = ctx->input(1);             // HostMemory
    const Tensor& cdf = ctx->input(2);
    OP_REQUIRES(ctx, encoded.dims() == 0, errors::InvalidArgument("Invalid `encoded` shape"));
    OP_REQUIRES(ctx, shape.dims() == 1, errors::InvalidArgument("Invalid `shape` shape"));
    TensorShape out_shape;
    std::vector<int64_t> ds(shape.NumElements());
    for (int64 i = 0; i < shape.NumElements(); ++i) {
      ds[i] = shape.flat<int32>()(i);
      OP_REQUIRES(ctx, ds[i] >= 0, errors::InvalidArgument("`shape` has a negative entry"));
      out_shape.AddDim(ds[i]);
    }
    const std::vector<int64_t> cs = Dims(cdf.shape());
    Tensor* decoded = nullptr;
    OP_REQUIRES_OK(ctx, ctx->allocate_output(0, out_shape, &decoded));
    DecodeScratch s;
    OP_REQUIRES_OK(ctx, UploadString(ctx, encoded.scalar<tstring>()(), &s));
    int32* words = s.words.flat<int32>().data();
    const int rc = tfc_range_decode(s.bytes.flat<uint8>().data(), s.cap, words, /*S=*/1, ds.data(),
                                    (int)ds.size(), cdf.flat<int32>().data(), cs.data(),
                                    (int)cs.size(), /*cdf_stream_stride=*/0, precision_,
                                    debug_level_, decoded->flat<int16>().data(), words + 1,
                                    ctx->eigen_gpu_device().stream());
    OP_REQUIRES(ctx, rc == TFC_OK, errors::InvalidArgument(tfc_last_error()));
    OP_REQUIRES_OK(ctx, CheckDecodeStatus(ctx, &s));
  }
 private:
  int precision_ = 0, debug_level_ = 1;
};

// ---- PmfToQuantizedCdf (cc/kernels/range_coding_helper_kernels.cc:53-204) ------------------------
class B200PmfToQuantizedCdfOp : public OpKernel {
 public:
  explicit B200PmfToQuantizedCdfOp(OpKernelConstruction* c) : OpKernel(c) {
    OP_REQUIRES_OK(c, c->GetAttr("precision", &precision_));
  }
  void Compute(OpKernelContext* ctx) override {
    const Tensor& pmf = ctx->input(0);
    OP_REQUIRES(ctx, pmf.dims() >= 1, errors::InvalidArgument("`pmf` should be at least 1-D."));
    const int64 n = pmf.dim_size(pmf.dims() - 1);
    OP_REQUIRES(ctx, n > 1, errors::InvalidArgument("`pmf` size should be at least 2 in the last axis."));
    TensorShape out_shape;
    for (int i = 0; i + 1 < pmf.dims(); ++i) out_shape.AddDim(pmf.dim_size(i));
    out_shape.AddDim(n + 1);
    Tensor* cdf = nullptr;
    OP_REQUIRES_OK(ctx, ctx->allocate_output(0, out_shape, &cdf));
    const int64 rows = pmf.NumElements() / n;
    const size_t bytes = tfc_pmf_to_quantized_cdf_workspace(rows, (int32_t)n);
    Tensor workspace;
    OP_REQUIRES_OK(ctx, ctx->allocate_temp(tensorflow::DT_UINT8, TensorShape({(int64)bytes + 16}),
                                           &workspace));
    const int rc = tfc_pmf_to_quantized_cdf(pmf.flat<float>().data(), rows, (int32_t)n, precision_,
                                            cdf->flat<int32>().data(),
                                            workspace.flat<uint8>().data(), bytes,
                                            ctx->eigen_gpu_device().stream());
    OP_REQUIRES(ctx, rc == TFC_OK, errors::InvalidArgument(tfc_last_error()));
  }
 private:
  int precision_ = 0;
};

REGISTER_KERNEL_BUILDER(tensorflow::Name("UnboundedIndexRangeEncode").Device(DEVICE_GPU).HostMemory("encoded"),
                        B200UnboundedIndexRangeEncodeOp);
REGISTER_KERNEL_BUILDER(tensorflow::Name("UnboundedIndexRangeDecode").Device(DEVICE_GPU).HostMemory("encoded"),
                        B200UnboundedIndexRangeDecodeOp);
REGISTER_KERNEL_BUILDER(tensorflow::Name("RangeEncode").Device(DEVICE_GPU).HostMemory("encoded"),
                        B200RangeEncodeOp);
REGISTER_KERNEL_BUILDER(
    tensorflow::Name("RangeDecode").Device(DEVICE_GPU).HostMemory("encoded").HostMemory("shape"),
    B200RangeDecodeOp);
REGISTER_KERNEL_BUILDER(tensorflow::Name("PmfToQuantizedCdf").Device(DEVICE_GPU),
                        B200PmfToQuantizedCdfOp);

}  // namespace b200
}  // namespace tensorflow_compression
