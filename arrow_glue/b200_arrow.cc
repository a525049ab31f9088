// b200_arrow.cc — Arrow FunctionRegistry glue over libarrow1_b200.so.  See b200_arrow.h.
//
// Host code only: unwraps device addresses out of CudaBuffer-resident ArrayData into b200_col
// descriptors, sizes and allocates outputs through CudaContext::Allocate, maps b200_status onto
// arrow::Status (the codes are equal by construction, include/arrow1_b200.h), and reproduces the
// reference's host-side rules (argument checks, output schema, error messages) for the path.
#include "b200_arrow.h"

#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <utility>

#include "arrow/array.h"
#include "arrow/array/data.h"
#include "arrow/array/util.h"
#include "arrow/buffer.h"
#include "arrow/chunked_array.h"
#include "arrow/compute/api_vector.h"
#include "arrow/compute/function.h"
#include "arrow/compute/kernel.h"
#include "arrow/gpu/cuda_context.h"
#include "arrow/gpu/cuda_memory.h"
#include "arrow/record_batch.h"
#include "arrow/scalar.h"
#include "arrow/table.h"
#include "arrow/type.h"
#include "arrow/util/bit_util.h"
#include "arrow/util/checked_cast.h"
#include "arrow/util/make_unique.h"

#include <cuda.h>

#include "arrow/c/bridge.h"
#include "arrow1_b200.h"

namespace arrow {
namespace compute {
namespace cuda {

using ::arrow::internal::checked_cast;

namespace {

// ---- device context ---------------------------------------------------------------------------
// One b200_ctx per device, created on first use, living for the process.  It runs on the legacy
// default stream so that it is ordered with CudaContext's synchronous cuMemcpy calls
// (gpu/cuda_context.cc CopyHostToDevice / CopyDeviceToHost).  A b200_ctx is single-stream and
// not re-entrant, so calls are serialised per device.
struct DeviceState {
  b200_ctx* ctx = nullptr;
  std::shared_ptr<::arrow::cuda::CudaContext> arrow_ctx;
  std::mutex mutex;
  // Device blocks handed back by results and scratch buffers that went out of scope, reused by later calls:
  // cuMemAlloc / cuMemFree of the gigabyte-sized outputs and scratch of a 100 M-row call cost more than its kernels
  // (bench_glue, round 2: SortIndices 9.1 ms per call against 5.8 ms of kernels).
  std::mutex pool_mutex;
  std::multimap<int64_t, std::shared_ptr<::arrow::cuda::CudaBuffer>> pool;
  int64_t pool_bytes = 0;
  // pinned host blocks of results that went out of scope (cuMemHostAlloc of an 800 MB result costs ~0.3 s)
  std::multimap<int64_t, std::shared_ptr<::arrow::cuda::CudaHostBuffer>> host_pool;
  int64_t host_pool_bytes = 0;
  // ingest / egress: a copy stream and two pinned staging blocks for pageable host buffers
  CUstream copy_stream = nullptr;
  void* staging[2] = {nullptr, nullptr};
  CUevent staging_done[2] = {nullptr, nullptr};
};

constexpr int64_t kPoolLimitBytes = int64_t(48) << 30;   // idle bytes kept per device
constexpr int64_t kStagingBytes = int64_t(32) << 20;
constexpr int64_t kHostPoolLimitBytes = int64_t(8) << 30;

// A pinned host buffer over a pooled block (see PooledCudaBuffer)
class PooledHostBuffer : public MutableBuffer {
 public:
  PooledHostBuffer(struct DeviceState* dev, std::shared_ptr<::arrow::cuda::CudaHostBuffer> block, int64_t size);
  ~PooledHostBuffer() override;

 private:
  struct DeviceState* dev_;
  std::shared_ptr<::arrow::cuda::CudaHostBuffer> block_;
};

// A CudaBuffer over a pooled block: when the last reference goes, the block returns to its device's pool.
class PooledCudaBuffer : public ::arrow::cuda::CudaBuffer {
 public:
  PooledCudaBuffer(DeviceState* dev, std::shared_ptr<::arrow::cuda::CudaBuffer> block, int64_t size)
      : ::arrow::cuda::CudaBuffer(block->address(), size, block->context(), /*own_data=*/false), dev_(dev), block_(std::move(block)) {}
  ~PooledCudaBuffer() override {
    std::lock_guard<std::mutex> lock(dev_->pool_mutex);
    if (dev_->pool_bytes + block_->size() <= kPoolLimitBytes) {
      dev_->pool_bytes += block_->size();
      dev_->pool.emplace(block_->size(), std::move(block_));
    }   // else: the block is freed by its own destructor
  }

 private:
  DeviceState* dev_;
  std::shared_ptr<::arrow::cuda::CudaBuffer> block_;
};

Status FromB200(b200_ctx* ctx, int code) {
  if (code == B200_OK) return Status::OK();
  const char* msg = ctx ? b200_last_error(ctx) : "";
  return Status(static_cast<StatusCode>(code), msg && *msg ? msg : "arrow1_b200 call failed");
}

Result<DeviceState*> GetDevice(int device) {
  static std::mutex registry_mutex;
  static std::map<int, std::unique_ptr<DeviceState>> devices;
  std::lock_guard<std::mutex> lock(registry_mutex);
  auto it = devices.find(device);
  if (it != devices.end()) return it->second.get();
  auto st = ::arrow::internal::make_unique<DeviceState>();
  ARROW_ASSIGN_OR_RAISE(auto manager, ::arrow::cuda::CudaDeviceManager::Instance());
  ARROW_ASSIGN_OR_RAISE(st->arrow_ctx, manager->GetContext(device));
  const int code = b200_ctx_create(device, /*cu_stream=*/nullptr, &st->ctx);
  if (code != B200_OK) {
    return Status(static_cast<StatusCode>(code), "b200_ctx_create failed for device " + std::to_string(device) +
                                                     " (no CPU fallback exists)");
  }
  DeviceState* raw = st.get();
  devices.emplace(device, std::move(st));
  return raw;
}

#define B200_OK_OR_RAISE(dev, expr)                          \
  do {                                                       \
    const int _code = (expr);                                \
    if (_code != B200_OK) return FromB200((dev)->ctx, _code); \
  } while (0)

PooledHostBuffer::PooledHostBuffer(DeviceState* dev, std::shared_ptr<::arrow::cuda::CudaHostBuffer> block, int64_t size)
    : MutableBuffer(block->mutable_data(), size), dev_(dev), block_(std::move(block)) {}
PooledHostBuffer::~PooledHostBuffer() {
  std::lock_guard<std::mutex> lock(dev_->pool_mutex);
  if (dev_->host_pool_bytes + block_->size() <= kHostPoolLimitBytes) {
    dev_->host_pool_bytes += block_->size();
    dev_->host_pool.emplace(block_->size(), std::move(block_));
  }
}

Result<std::shared_ptr<Buffer>> AllocatePinnedPooled(DeviceState* dev, int64_t nbytes) {
  const int64_t padded = ((nbytes < 1 ? 1 : nbytes) + 63) / 64 * 64;
  std::shared_ptr<::arrow::cuda::CudaHostBuffer> block;
  {
    std::lock_guard<std::mutex> lock(dev->pool_mutex);
    auto it = dev->host_pool.lower_bound(padded);
    if (it != dev->host_pool.end() && it->first <= padded + padded / 4 + (int64_t(1) << 20)) {
      block = std::move(it->second);
      dev->host_pool_bytes -= it->first;
      dev->host_pool.erase(it);
    }
  }
  if (!block) {
    const int64_t cap = padded < (int64_t(1) << 20) ? padded : (padded + (int64_t(2) << 20) - 1) / (int64_t(2) << 20) * (int64_t(2) << 20);
    ARROW_ASSIGN_OR_RAISE(block, ::arrow::cuda::AllocateCudaHostBuffer(dev->arrow_ctx->device_number(), cap));
  }
  return std::static_pointer_cast<Buffer>(std::make_shared<PooledHostBuffer>(dev, std::move(block), nbytes));
}

// Device allocation through the reference's own allocator; zero-filled when asked (bitmaps must
// come back zero-padded, compute/kernel.cc:51-59).  cuMemAlloc rejects 0 bytes, and the C ABI
// wants 8-byte multiples, so sizes are rounded up to 64 like Arrow's host buffers.
Result<std::shared_ptr<Buffer>> AllocateDevice(DeviceState* dev, int64_t nbytes) {
  const int64_t padded = ((nbytes < 1 ? 1 : nbytes) + 63) / 64 * 64;
  if (padded < (int64_t(1) << 20)) {   // small buffers: not worth pooling
    ARROW_ASSIGN_OR_RAISE(auto buf, dev->arrow_ctx->Allocate(padded));
    return std::static_pointer_cast<Buffer>(buf);
  }
  std::shared_ptr<::arrow::cuda::CudaBuffer> block;
  {
    std::lock_guard<std::mutex> lock(dev->pool_mutex);
    auto it = dev->pool.lower_bound(padded);
    if (it != dev->pool.end() && it->first <= padded + padded / 4 + (int64_t(1) << 20)) {
      block = std::move(it->second);
      dev->pool_bytes -= it->first;
      dev->pool.erase(it);
    }
  }
  if (!block) {
    const int64_t cap = (padded + (int64_t(2) << 20) - 1) / (int64_t(2) << 20) * (int64_t(2) << 20);
    auto fresh = dev->arrow_ctx->Allocate(cap);
    if (!fresh.ok()) {   // give the idle blocks back and try once more
      {
        std::lock_guard<std::mutex> lock(dev->pool_mutex);
        dev->pool.clear();
        dev->pool_bytes = 0;
      }
      fresh = dev->arrow_ctx->Allocate(cap);
    }
    ARROW_ASSIGN_OR_RAISE(block, std::move(fresh));
  }
  return std::static_pointer_cast<Buffer>(std::make_shared<PooledCudaBuffer>(dev, std::move(block), padded));
}

int64_t BitmapBytes(int64_t bits) { return (bits + 7) / 8; }

// ---- ArrayData -> b200_col ----------------------------------------------------------------------
Status CheckOnDevice(const std::shared_ptr<Buffer>& buf, const char* what) {
  if (buf && buf->is_cpu()) {
    return Status::Invalid("CUDA-backed kernel got a host-resident ", what,
                           " buffer; inputs must be CudaBuffer-resident ArrayData");
  }
  return Status::OK();
}

// physical b200_type of a fixed-width column: filter and group keys only move bits
// (PrimitiveFilter dispatches on bit_width, vector_selection.cc:813-832)
Result<int32_t> PhysicalType(const DataType& type) {
  if (type.id() == Type::BOOL) return B200_BOOL;
  if (!is_primitive(type.id())) {
    return Status::NotImplemented("CUDA-backed kernels handle fixed-width primitive columns, got ",
                                  type.ToString());
  }
  switch (checked_cast<const FixedWidthType&>(type).bit_width()) {
    case 8: return B200_UINT8;
    case 16: return B200_UINT16;
    case 32: return B200_UINT32;
    case 64: return B200_UINT64;
    default: break;
  }
  return Status::NotImplemented("bit width of ", type.ToString());
}

// group keys: fixed-width columns, and dictionary columns through their index column (DictionaryKeyEncoder,
// hash_aggregate.cc:198-233: the indices are the key, the dictionary rides along)
Result<int32_t> KeyPhysicalType(const DataType& type) {
  if (type.id() == Type::DICTIONARY) return PhysicalType(*checked_cast<const DictionaryType&>(type).index_type());
  // VarLengthKeyEncoder<BinaryType / StringType> (hash_aggregate.cc:344-347); the 64-bit-offset forms are not bound
  if (type.id() == Type::BINARY || type.id() == Type::STRING) return B200_BINARY;
  if (type.id() != Type::BOOL && !is_primitive(type.id())) return Status::NotImplemented("Keys of type ", type.ToString());
  return PhysicalType(type);
}
int KeyBitWidth(const DataType& type) {
  if (type.id() == Type::DICTIONARY) return checked_cast<const FixedWidthType&>(*checked_cast<const DictionaryType&>(type).index_type()).bit_width();
  return checked_cast<const FixedWidthType&>(type).bit_width();
}
// the buffers of one key column of `num_groups` groups, sized for b200_groupby_finalize
struct KeyOut {
  std::shared_ptr<Buffer> data, validity, bytes;   // bytes: variable-length keys only (data = int32 offsets)
};
Result<KeyOut> AllocateKeyOut(DeviceState* dev, b200_groupby* gb, int key_index, int32_t key_type, const DataType& type,
                              uint32_t num_groups, b200_out_col* out) {
  KeyOut k;
  std::memset(out, 0, sizeof(*out));
  if (key_type == B200_BINARY) {
    int64_t nbytes = 0;
    B200_OK_OR_RAISE(dev, b200_groupby_key_bytes(gb, key_index, &nbytes));
    ARROW_ASSIGN_OR_RAISE(k.data, AllocateDevice(dev, (int64_t(num_groups) + 1) * 4));
    ARROW_ASSIGN_OR_RAISE(k.bytes, AllocateDevice(dev, nbytes));
    out->extra = reinterpret_cast<void*>(k.bytes->address());
  } else {
    const int bits = key_type == B200_BOOL ? 1 : KeyBitWidth(type);
    ARROW_ASSIGN_OR_RAISE(k.data, AllocateDevice(dev, bits == 1 ? BitmapBytes(num_groups) : int64_t(num_groups) * (bits / 8)));
  }
  ARROW_ASSIGN_OR_RAISE(k.validity, AllocateDevice(dev, BitmapBytes(num_groups)));
  out->data = reinterpret_cast<void*>(k.data->address());
  out->validity = reinterpret_cast<void*>(k.validity->address());
  return k;
}
std::shared_ptr<ArrayData> KeyArrayData(const std::shared_ptr<DataType>& type, uint32_t num_groups, const KeyOut& k, int64_t nulls) {
  if (k.bytes) return ArrayData::Make(type, num_groups, {nulls ? k.validity : nullptr, k.data, k.bytes}, nulls);
  return ArrayData::Make(type, num_groups, {nulls ? k.validity : nullptr, k.data}, nulls);
}
// every batch of a dictionary key must carry the first batch's dictionary ("Unifying differing dictionaries" is
// NotImplemented upstream too, hash_aggregate.cc:205-210); device dictionaries are compared by identity
Status NoteKeyDictionary(const ArrayData& key, std::shared_ptr<ArrayData>* seen) {
  if (key.type->id() != Type::DICTIONARY) return Status::OK();
  if (!*seen) {
    *seen = key.dictionary;
  } else if (seen->get() != key.dictionary.get()) {
    return Status::NotImplemented("Unifying differing dictionaries");
  }
  return Status::OK();
}

// logical b200_type: ordering and arithmetic depend on signedness / floating point
Result<int32_t> NumericType(const DataType& type) {
  switch (type.id()) {
    case Type::UINT8: return B200_UINT8;
    case Type::INT8: return B200_INT8;
    case Type::UINT16: return B200_UINT16;
    case Type::INT16: return B200_INT16;
    case Type::UINT32: return B200_UINT32;
    case Type::INT32: case Type::DATE32: case Type::TIME32: return B200_INT32;
    case Type::UINT64: return B200_UINT64;
    case Type::INT64: case Type::DATE64: case Type::TIME64: case Type::TIMESTAMP:
    case Type::DURATION: return B200_INT64;
    case Type::FLOAT: return B200_FLOAT;
    case Type::DOUBLE: return B200_DOUBLE;
    default: break;
  }
  return Status::NotImplemented("no CUDA-backed kernel for type ", type.ToString());
}

Status ToCol(const ArrayData& a, int32_t type, b200_col* c) {
  if (a.buffers.size() < 2) return Status::Invalid("expected a primitive array layout");
  RETURN_NOT_OK(CheckOnDevice(a.buffers[0], "validity"));
  RETURN_NOT_OK(CheckOnDevice(a.buffers[1], "data"));
  std::memset(c, 0, sizeof(*c));
  if (type == B200_BINARY) {   // offsets in buffers[1], value bytes in buffers[2]
    if (a.buffers.size() < 3) return Status::Invalid("expected a binary array layout");
    RETURN_NOT_OK(CheckOnDevice(a.buffers[2], "value"));
    c->extra = a.buffers[2] ? reinterpret_cast<const void*>(a.buffers[2]->address()) : nullptr;
  }
  c->validity = a.buffers[0] ? reinterpret_cast<const void*>(a.buffers[0]->address()) : nullptr;
  c->data = a.buffers[1] ? reinterpret_cast<const void*>(a.buffers[1]->address()) : nullptr;
  c->type = type;
  c->length = a.length;
  c->offset = a.offset;
  // may be kUnknownNullCount, like PrimitiveArg (util_internal.cc:45-58); never computed on the host
  c->null_count = a.buffers[0] ? a.null_count.load() : 0;
  return Status::OK();
}

// ---- Filter --------------------------------------------------------------------------------------
// == PrimitiveFilter (vector_selection.cc:786-835): GetFilterOutputSize -> PreallocateData -> Exec, for every
// column that shares the mask: ONE plan (the scan of the mask) serves all of them, which is what
// FilterRecordBatch / FilterTable get from their shared take-index vector (vector_selection.cc:1776-1866)
Result<std::vector<std::shared_ptr<ArrayData>>> FilterColumns(DeviceState* dev, const std::vector<const ArrayData*>& columns,
                                                              const ArrayData& mask, int null_selection) {
  if (mask.type->id() != Type::BOOL) {
    return Status::NotImplemented("Filter argument must be boolean type");
  }
  b200_col m;
  RETURN_NOT_OK(ToCol(mask, B200_BOOL, &m));
  std::vector<b200_col> v(columns.size());
  std::vector<int32_t> vtypes(columns.size());
  for (size_t c = 0; c < columns.size(); ++c) {
    if (columns[c]->length != mask.length) {
      // exec.cc:186-189 (ExecBatchIterator) — the code the reference's tests assert
      return Status::Invalid("Array arguments must all be the same length");
    }
    ARROW_ASSIGN_OR_RAISE(vtypes[c], PhysicalType(*columns[c]->type));
    RETURN_NOT_OK(ToCol(*columns[c], vtypes[c], &v[c]));
  }

  std::lock_guard<std::mutex> lock(dev->mutex);
  size_t plan_bytes = 0;
  B200_OK_OR_RAISE(dev, b200_filter_plan_bytes(m.length, &plan_bytes));
  ARROW_ASSIGN_OR_RAISE(auto plan, AllocateDevice(dev, static_cast<int64_t>(plan_bytes)));
  void* plan_ptr = reinterpret_cast<void*>(plan->address());
  int64_t out_len = 0;
  B200_OK_OR_RAISE(dev, b200_filter_plan(dev->ctx, &m, null_selection, plan_ptr, &out_len));

  std::vector<std::shared_ptr<ArrayData>> out(columns.size());
  for (size_t c = 0; c < columns.size(); ++c) {
    int allocate_validity = 0;
    int64_t out_null_count = 0;
    B200_OK_OR_RAISE(dev, b200_filter_out_spec(&v[c], &m, null_selection, &allocate_validity, &out_null_count));
    const int bits = vtypes[c] == B200_BOOL ? 1 : checked_cast<const FixedWidthType&>(*columns[c]->type).bit_width();
    const int64_t data_bytes = bits == 1 ? BitmapBytes(out_len) : out_len * (bits / 8);
    ARROW_ASSIGN_OR_RAISE(auto out_data, AllocateDevice(dev, data_bytes));
    std::shared_ptr<Buffer> out_validity;
    if (allocate_validity) {
      ARROW_ASSIGN_OR_RAISE(out_validity, AllocateDevice(dev, BitmapBytes(out_len)));
    }
    B200_OK_OR_RAISE(dev, b200_filter_apply(dev->ctx, &v[c], &m, null_selection, plan_ptr,
                                            reinterpret_cast<void*>(out_data->address()),
                                            out_validity ? reinterpret_cast<void*>(out_validity->address()) : nullptr,
                                            out_len));
    out[c] = ArrayData::Make(columns[c]->type, out_len, {std::move(out_validity), std::move(out_data)}, out_null_count);
  }
  // the plan buffer is released when this returns: its last reader must have finished
  B200_OK_OR_RAISE(dev, b200_ctx_synchronize(dev->ctx));
  return out;
}

Result<std::shared_ptr<ArrayData>> FilterArray(DeviceState* dev, const ArrayData& values,
                                               const ArrayData& mask, int null_selection) {
  if (values.length != mask.length && mask.type->id() == Type::BOOL) {
    return Status::Invalid("Array arguments must all be the same length");
  }
  ARROW_ASSIGN_OR_RAISE(auto out, FilterColumns(dev, {&values}, mask, null_selection));
  return out[0];
}

// An output whose null count PrimitiveFilter leaves unknown (vector_selection.cc:799-812) must have it
// settled before it becomes a chunk: ChunkedArray's constructor sums chunk->null_count()
// (chunked_array.cc:46-63), which would scan the bitmap on the host.  The set bits of the validity bitmap are
// counted on the device instead — GetFilterOutputSize of the bitmap read as a boolean column.
Status SettleNullCount(DeviceState* dev, ArrayData* a) {
  if (a->null_count.load() != kUnknownNullCount) return Status::OK();
  if (!a->buffers[0] || a->length == 0) {
    a->null_count = 0;
    return Status::OK();
  }
  b200_col bits;
  std::memset(&bits, 0, sizeof(bits));
  bits.data = reinterpret_cast<const void*>(a->buffers[0]->address());
  bits.type = B200_BOOL;
  bits.length = a->length;
  bits.offset = a->offset;
  int64_t valid = 0;
  std::lock_guard<std::mutex> lock(dev->mutex);
  B200_OK_OR_RAISE(dev, b200_filter_count(dev->ctx, &bits, 0, &valid));
  a->null_count = a->length - valid;
  return Status::OK();
}

// ---- Take ----------------------------------------------------------------------------------------
// == PrimitiveTake (vector_selection.cc:493-526): boundscheck -> PreallocateData (validity always) -> Exec
Result<std::shared_ptr<ArrayData>> TakeArray(DeviceState* dev, const ArrayData& values, const ArrayData& indices,
                                             bool boundscheck) {
  ARROW_ASSIGN_OR_RAISE(const int32_t vtype, PhysicalType(*values.type));
  if (!is_integer(indices.type->id())) return Status::NotImplemented("Take indices must be integer type");
  ARROW_ASSIGN_OR_RAISE(const int32_t itype, NumericType(*indices.type));
  b200_col v, i;
  RETURN_NOT_OK(ToCol(values, vtype, &v));
  RETURN_NOT_OK(ToCol(indices, itype, &i));
  std::lock_guard<std::mutex> lock(dev->mutex);
  const int bits = vtype == B200_BOOL ? 1 : checked_cast<const FixedWidthType&>(*values.type).bit_width();
  const int64_t n = indices.length;
  // bitmaps are written as whole 32-bit words
  const int64_t bitmap_bytes = (n + 31) / 32 * 4;
  ARROW_ASSIGN_OR_RAISE(auto out_data, AllocateDevice(dev, bits == 1 ? bitmap_bytes : n * (bits / 8)));
  ARROW_ASSIGN_OR_RAISE(auto out_validity, AllocateDevice(dev, bitmap_bytes));
  int64_t null_count = 0;
  B200_OK_OR_RAISE(dev, b200_take(dev->ctx, &v, &i, boundscheck ? 1 : 0, reinterpret_cast<void*>(out_data->address()),
                                  reinterpret_cast<void*>(out_validity->address()), &null_count));
  return ArrayData::Make(values.type, n, {std::move(out_validity), std::move(out_data)}, null_count);
}

// == TakeCA (vector_selection.cc:1921-1944) without its Concatenate: the chunks are gathered from where
// they lie (b200_take_chunked); one output array
Result<std::shared_ptr<ArrayData>> TakeChunks(DeviceState* dev, const ChunkedArray& values, const ArrayData& indices,
                                              bool boundscheck) {
  if (values.num_chunks() == 0) return Status::Invalid("Must pass at least one array");   // Concatenate's
  if (values.num_chunks() == 1) return TakeArray(dev, *values.chunk(0)->data(), indices, boundscheck);
  ARROW_ASSIGN_OR_RAISE(const int32_t vtype, PhysicalType(*values.type()));
  if (!is_integer(indices.type->id())) return Status::NotImplemented("Take indices must be integer type");
  ARROW_ASSIGN_OR_RAISE(const int32_t itype, NumericType(*indices.type));
  std::vector<b200_col> v(values.num_chunks());
  for (int c = 0; c < values.num_chunks(); ++c) RETURN_NOT_OK(ToCol(*values.chunk(c)->data(), vtype, &v[c]));
  b200_col i;
  RETURN_NOT_OK(ToCol(indices, itype, &i));
  std::lock_guard<std::mutex> lock(dev->mutex);
  const int bits = vtype == B200_BOOL ? 1 : checked_cast<const FixedWidthType&>(*values.type()).bit_width();
  const int64_t n = indices.length;
  const int64_t bitmap_bytes = (n + 31) / 32 * 4;
  ARROW_ASSIGN_OR_RAISE(auto out_data, AllocateDevice(dev, bits == 1 ? bitmap_bytes : n * (bits / 8)));
  ARROW_ASSIGN_OR_RAISE(auto out_validity, AllocateDevice(dev, bitmap_bytes));
  int64_t null_count = 0;
  B200_OK_OR_RAISE(dev, b200_take_chunked(dev->ctx, v.data(), int(v.size()), &i, boundscheck ? 1 : 0,
                                          reinterpret_cast<void*>(out_data->address()),
                                          reinterpret_cast<void*>(out_validity->address()), &null_count));
  return ArrayData::Make(values.type(), n, {std::move(out_validity), std::move(out_data)}, null_count);
}

struct TakeStateB200 : public KernelState {
  explicit TakeStateB200(bool boundscheck) : boundscheck(boundscheck) {}
  bool boundscheck;
};

const FunctionDoc array_take_doc(
    "Select values from an array by integer indices (CUDA)",
    "Same contract as the CPU array_take; inputs and output are CudaBuffer-resident.", {"values", "indices"},
    "TakeOptions");
const FunctionDoc take_doc("Select values by integer indices (CUDA)", "Array x Array inputs; see array_take.",
                           {"input", "indices"}, "TakeOptions");
const TakeOptions kDefaultTakeOptions = TakeOptions::Defaults();

// == TakeMetaFunction (vector_selection.cc:2028-2081) and its routes TakeAA / CA / CC / AC / RA / TA / TC
// (:1914-2017): same dispatch, same output shapes (CA -> one chunk; CC, AC -> one chunk per index chunk)
class CudaTakeMetaFunction : public MetaFunction {
 public:
  explicit CudaTakeMetaFunction(int device)
      : MetaFunction("take", Arity::Binary(), &take_doc, &kDefaultTakeOptions), device_(device) {}
  Result<Datum> ExecuteImpl(const std::vector<Datum>& args, const FunctionOptions* options,
                            ExecContext* ctx) const override {
    const Datum::Kind index_kind = args[1].kind();
    const bool boundscheck = options ? static_cast<const TakeOptions*>(options)->boundscheck : true;
    ARROW_ASSIGN_OR_RAISE(DeviceState * dev, GetDevice(device_));
    switch (args[0].kind()) {
      case Datum::ARRAY:
        if (index_kind == Datum::ARRAY) return CallFunction("array_take", args, options, ctx);
        if (index_kind == Datum::CHUNKED_ARRAY) {   // TakeAC
          ArrayVector out;
          for (const auto& ix : args[1].chunked_array()->chunks()) {
            ARROW_ASSIGN_OR_RAISE(auto piece, TakeArray(dev, *args[0].array(), *ix->data(), boundscheck));
            out.push_back(MakeArray(piece));
          }
          return Datum(std::make_shared<ChunkedArray>(std::move(out), args[0].type()));
        }
        break;
      case Datum::CHUNKED_ARRAY:
        if (index_kind == Datum::ARRAY) return TakeCA(dev, *args[0].chunked_array(), *args[1].array(), boundscheck);
        if (index_kind == Datum::CHUNKED_ARRAY) return TakeCC(dev, *args[0].chunked_array(), *args[1].chunked_array(), boundscheck);
        break;
      case Datum::RECORD_BATCH:
        if (index_kind == Datum::ARRAY) {   // TakeRA
          const auto& batch = *args[0].record_batch();
          std::vector<std::shared_ptr<ArrayData>> columns(batch.num_columns());
          for (int j = 0; j < batch.num_columns(); ++j) {
            ARROW_ASSIGN_OR_RAISE(columns[j], TakeArray(dev, *batch.column_data(j), *args[1].array(), boundscheck));
          }
          return Datum(RecordBatch::Make(batch.schema(), args[1].length(), std::move(columns)));
        }
        break;
      case Datum::TABLE: {
        const auto& table = *args[0].table();
        std::vector<std::shared_ptr<ChunkedArray>> columns(table.num_columns());
        if (index_kind == Datum::ARRAY) {   // TakeTA
          for (int j = 0; j < table.num_columns(); ++j) {
            ARROW_ASSIGN_OR_RAISE(columns[j], TakeCA(dev, *table.column(j), *args[1].array(), boundscheck));
          }
          return Datum(Table::Make(table.schema(), std::move(columns)));
        }
        if (index_kind == Datum::CHUNKED_ARRAY) {   // TakeTC
          for (int j = 0; j < table.num_columns(); ++j) {
            ARROW_ASSIGN_OR_RAISE(columns[j], TakeCC(dev, *table.column(j), *args[1].chunked_array(), boundscheck));
          }
          return Datum(Table::Make(table.schema(), std::move(columns)));
        }
        break;
      }
      default:
        break;
    }
    return Status::NotImplemented("Unsupported types for take operation: values=", args[0].ToString(),
                                  "indices=", args[1].ToString());
  }

 private:
  static Result<std::shared_ptr<ChunkedArray>> TakeCA(DeviceState* dev, const ChunkedArray& values, const ArrayData& indices,
                                                      bool boundscheck) {
    ARROW_ASSIGN_OR_RAISE(auto piece, TakeChunks(dev, values, indices, boundscheck));
    return std::make_shared<ChunkedArray>(ArrayVector{MakeArray(piece)}, values.type());
  }
  static Result<std::shared_ptr<ChunkedArray>> TakeCC(DeviceState* dev, const ChunkedArray& values, const ChunkedArray& indices,
                                                      bool boundscheck) {
    ArrayVector out;
    for (const auto& ix : indices.chunks()) {
      ARROW_ASSIGN_OR_RAISE(auto piece, TakeChunks(dev, values, *ix->data(), boundscheck));
      out.push_back(MakeArray(piece));
    }
    return std::make_shared<ChunkedArray>(std::move(out), values.type());
  }
  int device_;
};

int NullSelectionOf(const FunctionOptions* options) {
  if (!options) return 0;
  return static_cast<const FilterOptions*>(options)->null_selection_behavior == FilterOptions::EMIT_NULL ? 1 : 0;
}

const FilterOptions kDefaultFilterOptions = FilterOptions::Defaults();
const ArraySortOptions kDefaultArraySortOptions = ArraySortOptions::Defaults();
const SortOptions kDefaultSortOptions = SortOptions::Defaults();

const FunctionDoc array_filter_doc(
    "Filter with a boolean selection filter (CUDA, sm_100a)",
    "Same contract as the CPU array_filter; inputs and output are CudaBuffer-resident.",
    {"array", "selection_filter"}, "FilterOptions");
const FunctionDoc filter_doc(
    "Filter with a boolean selection filter (CUDA, sm_100a)",
    "Same contract as the CPU filter; inputs and output are CudaBuffer-resident.",
    {"input", "selection_filter"}, "FilterOptions");
const FunctionDoc array_sort_indices_doc(
    "Return the indices that would sort an array (CUDA, sm_100a)",
    "Stable; nulls last, NaNs before nulls; input and output are CudaBuffer-resident.",
    {"array"}, "ArraySortOptions");
const FunctionDoc sort_indices_doc(
    "Return the indices that would sort an array (CUDA, sm_100a)",
    "Array and single-chunk ChunkedArray inputs; see array_sort_indices.", {"input"}, "SortOptions");
const FunctionDoc hash_count_doc("Count the number of null / non-null values per group (CUDA)", "", {"array", "group_id_array", "group_count"}, "CountOptions");
const FunctionDoc hash_sum_doc("Sum values per group (CUDA)", "", {"array", "group_id_array", "group_count"});
const FunctionDoc hash_min_max_doc("Min and max per group (CUDA)", "", {"array", "group_id_array", "group_count"}, "MinMaxOptions");
const FunctionDoc hash_mean_doc("Mean per group (CUDA); MeanImpl::Finalize over the grouped sum state", "", {"array", "group_id_array", "group_count"});

struct FilterState : public KernelState {
  explicit FilterState(int null_selection) : null_selection(null_selection) {}
  int null_selection;
};
struct SortState : public KernelState {
  explicit SortState(SortOrder order) : order(order) {}
  SortOrder order;
};

// Datum of chunks -> list of ArrayData
Result<ArrayVector> ChunksOf(const Datum& d) {
  if (d.kind() == Datum::ARRAY) return ArrayVector{d.make_array()};
  if (d.kind() == Datum::CHUNKED_ARRAY) return d.chunked_array()->chunks();
  return Status::NotImplemented("CUDA-backed kernels take Array or ChunkedArray, got ", d.ToString());
}

// == FilterMetaFunction (vector_selection.cc:1877-1903) + FilterRecordBatch / FilterTable (:1776-1866):
// the table forms are the same mask applied column by column (pinned by vector_selection_test.cc:208-236)
class CudaFilterMetaFunction : public MetaFunction {
 public:
  explicit CudaFilterMetaFunction(int device)
      : MetaFunction("filter", Arity::Binary(), &filter_doc, &kDefaultFilterOptions), device_(device) {}

  Result<Datum> ExecuteImpl(const std::vector<Datum>& args, const FunctionOptions* options,
                            ExecContext* ctx) const override {
    if (args[1].type()->id() != Type::BOOL) {
      return Status::NotImplemented("Filter argument must be boolean type");
    }
    ARROW_ASSIGN_OR_RAISE(DeviceState * dev, GetDevice(device_));
    const int null_selection = NullSelectionOf(options);
    if (args[0].kind() == Datum::RECORD_BATCH) {
      // FilterRecordBatch (vector_selection.cc:1776-1797): one scan of the mask, every column compacted by it
      if (args[1].kind() != Datum::ARRAY) return Status::NotImplemented("Filter should be array-like");
      const auto& batch = *args[0].record_batch();
      if (batch.num_rows() != args[1].length()) return Status::Invalid("Filter inputs must all be the same length");
      std::vector<const ArrayData*> in(batch.num_columns());
      for (int i = 0; i < batch.num_columns(); ++i) in[i] = batch.column_data(i).get();
      ARROW_ASSIGN_OR_RAISE(auto columns, FilterColumns(dev, in, *args[1].array(), null_selection));
      int64_t out_len = 0;
      if (columns.empty()) {   // no columns: the row count still follows the mask
        std::lock_guard<std::mutex> lock(dev->mutex);
        b200_col m;
        RETURN_NOT_OK(ToCol(*args[1].array(), B200_BOOL, &m));
        B200_OK_OR_RAISE(dev, b200_filter_count(dev->ctx, &m, null_selection, &out_len));
      } else {
        out_len = columns[0]->length;
      }
      return Datum(RecordBatch::Make(batch.schema(), out_len, std::move(columns)));
    }
    if (args[0].kind() == Datum::TABLE) {
      // FilterTable (vector_selection.cc:1799-1866): columns and mask re-chunked consistently (slices only move
      // offsets; no buffer is touched), then every mask chunk is planned once and applied to all columns
      const auto& table = *args[0].table();
      if (args[1].kind() != Datum::ARRAY && args[1].kind() != Datum::CHUNKED_ARRAY) return Status::NotImplemented("Filter should be array-like");
      if (table.num_rows() != args[1].length()) return Status::Invalid("Filter inputs must all be the same length");
      if (table.num_rows() == 0) return Datum(Table::Make(table.schema(), table.columns(), 0));
      const int num_columns = table.num_columns();
      std::vector<ArrayVector> inputs(num_columns + 1);
      for (int i = 0; i < num_columns; ++i) inputs[i] = table.column(i)->chunks();
      ARROW_ASSIGN_OR_RAISE(inputs.back(), ChunksOf(args[1]));
      inputs = arrow::internal::RechunkArraysConsistently(inputs);
      std::vector<ArrayVector> out_columns(num_columns);
      int64_t out_rows = 0;
      for (size_t b = 0; b < inputs.back().size(); ++b) {
        std::vector<const ArrayData*> in(num_columns);
        for (int c = 0; c < num_columns; ++c) in[c] = inputs[c][b]->data().get();
        ARROW_ASSIGN_OR_RAISE(auto pieces, FilterColumns(dev, in, *inputs.back()[b]->data(), null_selection));
        if (num_columns == 0 || pieces[0]->length == 0) continue;   // like upstream, empty pieces are not kept
        for (int c = 0; c < num_columns; ++c) {
          RETURN_NOT_OK(SettleNullCount(dev, pieces[c].get()));
          out_columns[c].push_back(MakeArray(pieces[c]));
        }
        out_rows += pieces[0]->length;
      }
      std::vector<std::shared_ptr<ChunkedArray>> out_chunks(num_columns);
      for (int c = 0; c < num_columns; ++c) {
        out_chunks[c] = std::make_shared<ChunkedArray>(std::move(out_columns[c]), table.column(c)->type());
      }
      return Datum(Table::Make(table.schema(), std::move(out_chunks), out_rows));
    }
    if (args[0].kind() == Datum::CHUNKED_ARRAY || args[1].kind() == Datum::CHUNKED_ARRAY) {
      // upstream: array_filter through ExecBatchIterator, which slices both inputs to common chunk boundaries
      // (exec.cc:150-215); here the same slicing, one output chunk per common piece
      ARROW_ASSIGN_OR_RAISE(auto vchunks, ChunksOf(args[0]));
      ARROW_ASSIGN_OR_RAISE(auto mchunks, ChunksOf(args[1]));
      if (args[0].length() != args[1].length()) return Status::Invalid("Array arguments must all be the same length");
      if (args[0].length() == 0) return Datum(std::make_shared<ChunkedArray>(ArrayVector{}, args[0].type()));
      auto groups = arrow::internal::RechunkArraysConsistently({vchunks, mchunks});
      ArrayVector out;
      for (size_t i = 0; i < groups[0].size(); ++i) {
        ARROW_ASSIGN_OR_RAISE(auto piece, FilterArray(dev, *groups[0][i]->data(), *groups[1][i]->data(), null_selection));
        RETURN_NOT_OK(SettleNullCount(dev, piece.get()));
        out.push_back(MakeArray(piece));
      }
      return Datum(std::make_shared<ChunkedArray>(std::move(out), args[0].type()));
    }
    return CallFunction("array_filter", args, options, ctx);
  }

 private:
  int device_;
};

// ---- SortIndices ---------------------------------------------------------------------------------
// == ArraySortIndices::Exec (vector_sort.cc:559-575); output uint64, never null (:1747)
Result<std::shared_ptr<ArrayData>> SortIndicesArray(DeviceState* dev, const ArrayData& values, SortOrder order) {
  ARROW_ASSIGN_OR_RAISE(const int32_t vtype, NumericType(*values.type));
  b200_col v;
  RETURN_NOT_OK(ToCol(values, vtype, &v));
  std::lock_guard<std::mutex> lock(dev->mutex);
  size_t scratch_bytes = 0;
  B200_OK_OR_RAISE(dev, b200_sort_indices_scratch_bytes(v.length, &scratch_bytes));
  ARROW_ASSIGN_OR_RAISE(auto scratch, AllocateDevice(dev, static_cast<int64_t>(scratch_bytes)));
  ARROW_ASSIGN_OR_RAISE(auto out, AllocateDevice(dev, v.length * 8));
  B200_OK_OR_RAISE(dev, b200_sort_indices(dev->ctx, &v, order == SortOrder::Descending ? 1 : 0, /*index_base=*/0,
                                          reinterpret_cast<uint64_t*>(out->address()),
                                          reinterpret_cast<void*>(scratch->address()), scratch_bytes));
  B200_OK_OR_RAISE(dev, b200_ctx_synchronize(dev->ctx));  // scratch is released on return
  return ArrayData::Make(uint64(), v.length, {nullptr, std::move(out)}, 0);
}

// == SortIndicesMetaFunction (vector_sort.cc:1565-1711), Array / single-chunk ChunkedArray routes
class CudaSortIndicesMetaFunction : public MetaFunction {
 public:
  explicit CudaSortIndicesMetaFunction(int device)
      : MetaFunction("sort_indices", Arity::Unary(), &sort_indices_doc, &kDefaultSortOptions), device_(device) {}

  Result<Datum> ExecuteImpl(const std::vector<Datum>& args, const FunctionOptions* options,
                            ExecContext* ctx) const override {
    const SortOptions& sort_options = static_cast<const SortOptions&>(*options);
    SortOrder order = SortOrder::Ascending;  // vector_sort.cc:1600-1606
    if (!sort_options.sort_keys.empty()) order = sort_options.sort_keys[0].order;
    ArraySortOptions array_options(order);
    switch (args[0].kind()) {
      case Datum::ARRAY:
        return CallFunction("array_sort_indices", {args[0]}, &array_options, ctx);
      case Datum::CHUNKED_ARRAY: {
        const auto& chunked = *args[0].chunked_array();
        if (chunked.num_chunks() == 1) {
          return CallFunction("array_sort_indices", {Datum(chunked.chunk(0))}, &array_options, ctx);
        }
        if (chunked.num_chunks() == 0) {
          ARROW_ASSIGN_OR_RAISE(DeviceState * dev, GetDevice(device_));
          ARROW_ASSIGN_OR_RAISE(auto out, AllocateDevice(dev, 0));
          return Datum(ArrayData::Make(uint64(), 0, {nullptr, std::move(out)}, 0));
        }
        // ChunkedArraySorter semantics (vector_sort.cc:657-832): global indices over the logical concatenation
        ARROW_ASSIGN_OR_RAISE(DeviceState * dev, GetDevice(device_));
        ARROW_ASSIGN_OR_RAISE(const int32_t vtype, NumericType(*chunked.type()));
        std::vector<b200_col> cols(chunked.num_chunks());
        for (int c = 0; c < chunked.num_chunks(); ++c) RETURN_NOT_OK(ToCol(*chunked.chunk(c)->data(), vtype, &cols[c]));
        std::lock_guard<std::mutex> lock(dev->mutex);
        size_t scratch_bytes = 0;
        B200_OK_OR_RAISE(dev, b200_sort_indices_chunked_scratch_bytes(chunked.length(), &scratch_bytes));
        ARROW_ASSIGN_OR_RAISE(auto scratch, AllocateDevice(dev, static_cast<int64_t>(scratch_bytes)));
        ARROW_ASSIGN_OR_RAISE(auto out, AllocateDevice(dev, chunked.length() * 8));
        B200_OK_OR_RAISE(dev, b200_sort_indices_chunked(dev->ctx, cols.data(), int(cols.size()), order == SortOrder::Descending ? 1 : 0,
                                                        reinterpret_cast<uint64_t*>(out->address()),
                                                        reinterpret_cast<void*>(scratch->address()), scratch_bytes));
        B200_OK_OR_RAISE(dev, b200_ctx_synchronize(dev->ctx));
        return Datum(ArrayData::Make(uint64(), chunked.length(), {nullptr, std::move(out)}, 0));
      }
      case Datum::RECORD_BATCH: {
        // SortIndices(RecordBatch) (vector_sort.cc:1632-1670)
        const auto& batch = *args[0].record_batch();
        const auto& sort_keys = sort_options.sort_keys;
        if (sort_keys.empty()) return Status::Invalid("Must specify one or more sort keys");
        std::vector<std::shared_ptr<ArrayData>> keys;
        for (const auto& key : sort_keys) {
          auto array = batch.GetColumnByName(key.name);
          if (!array) return Status::Invalid("Nonexistent sort key column: ", key.name);
          keys.push_back(array->data());
        }
        if (keys.size() == 1) return ExecuteImpl({Datum(keys[0])}, options, ctx);
        return SortMulti(keys, sort_keys, batch.num_rows());
      }
      case Datum::TABLE: {
        // SortIndices(Table) (vector_sort.cc:1672-1711)
        const auto& table = *args[0].table();
        const auto& sort_keys = sort_options.sort_keys;
        if (sort_keys.empty()) return Status::Invalid("Must specify one or more sort keys");
        std::vector<std::shared_ptr<ChunkedArray>> columns;
        for (const auto& key : sort_keys) {
          auto chunked = table.GetColumnByName(key.name);
          if (!chunked) return Status::Invalid("Nonexistent sort key column: ", key.name);
          columns.push_back(chunked);
        }
        if (columns.size() == 1) return ExecuteImpl({Datum(columns[0])}, options, ctx);
        if (table.num_rows() == 0) {
          ARROW_ASSIGN_OR_RAISE(DeviceState * dev, GetDevice(device_));
          ARROW_ASSIGN_OR_RAISE(auto out, AllocateDevice(dev, 0));
          return Datum(ArrayData::Make(uint64(), 0, {nullptr, std::move(out)}, 0));
        }
        // key columns sliced to common chunk boundaries (offsets move, buffers stay), then one call
        std::vector<ArrayVector> groups;
        for (const auto& c : columns) groups.push_back(c->chunks());
        groups = arrow::internal::RechunkArraysConsistently(groups);
        std::vector<std::shared_ptr<ArrayData>> keys;
        for (const auto& g : groups)
          for (const auto& piece : g) keys.push_back(piece->data());
        return SortMulti(keys, sort_keys, table.num_rows(), int(groups[0].size()));
      }
      default:
        break;
    }
    return Status::NotImplemented("Unsupported types for sort_indices operation: values=", args[0].ToString());
  }

 private:
  // several sort keys: b200_sort_indices_multi[_chunked] (one stable radix sort per key, least significant first).
  // keys: key-major, n_chunks pieces per sort key
  Result<Datum> SortMulti(const std::vector<std::shared_ptr<ArrayData>>& keys, const std::vector<SortKey>& sort_keys,
                          int64_t length, int n_chunks = 1) const {
    ARROW_ASSIGN_OR_RAISE(DeviceState * dev, GetDevice(device_));
    std::vector<b200_col> cols(keys.size());
    std::vector<int> orders(sort_keys.size());
    for (size_t i = 0; i < keys.size(); ++i) {
      ARROW_ASSIGN_OR_RAISE(const int32_t t, NumericType(*keys[i]->type));
      RETURN_NOT_OK(ToCol(*keys[i], t, &cols[i]));
    }
    for (size_t k = 0; k < sort_keys.size(); ++k) orders[k] = sort_keys[k].order == SortOrder::Descending ? 1 : 0;
    std::lock_guard<std::mutex> lock(dev->mutex);
    size_t scratch_bytes = 0;
    B200_OK_OR_RAISE(dev, b200_sort_indices_multi_chunked_scratch_bytes(length, &scratch_bytes));
    ARROW_ASSIGN_OR_RAISE(auto scratch, AllocateDevice(dev, static_cast<int64_t>(scratch_bytes)));
    ARROW_ASSIGN_OR_RAISE(auto out, AllocateDevice(dev, length * 8));
    B200_OK_OR_RAISE(dev, b200_sort_indices_multi_chunked(dev->ctx, cols.data(), orders.data(), int(sort_keys.size()), n_chunks,
                                                          reinterpret_cast<uint64_t*>(out->address()),
                                                          reinterpret_cast<void*>(scratch->address()), scratch_bytes));
    B200_OK_OR_RAISE(dev, b200_ctx_synchronize(dev->ctx));  // scratch is released on return
    return Datum(ArrayData::Make(uint64(), length, {nullptr, std::move(out)}, 0));
  }

  int device_;
};

// ---- grouped aggregation -------------------------------------------------------------------------
struct AggInfo {
  int32_t kind;
  std::shared_ptr<DataType> out_type;
};

// accumulator / output types: GroupedCountImpl -> int64 (hash_aggregate.cc:507-509), GroupedSumImpl
// -> int64 / uint64 / double (aggregate_internal.h:30-50), GroupedMinMaxImpl -> struct<min,max> (:759-761)
Result<AggInfo> ResolveAggregate(const std::string& function, const std::shared_ptr<DataType>& arg_type_ptr) {
  const DataType& arg_type = *arg_type_ptr;
  AggInfo info;
  if (function == "hash_count") {
    info.kind = B200_HASH_COUNT;
    info.out_type = int64();
    return info;
  }
  if (function == "hash_sum" || function == "hash_mean") {
    info.kind = function == "hash_sum" ? B200_HASH_SUM : B200_HASH_MEAN;
    if (!is_integer(arg_type.id()) && !is_floating(arg_type.id()) ) {
      return Status::NotImplemented("Summing data of type ", arg_type.ToString());
    }
    if (arg_type.id() == Type::HALF_FLOAT) return Status::NotImplemented("Summing data of type ", arg_type.ToString());
    if (function == "hash_mean") info.out_type = float64();
    else if (is_floating(arg_type.id())) info.out_type = float64();
    else if (is_signed_integer(arg_type.id())) info.out_type = int64();
    else info.out_type = uint64();
    return info;
  }
  if (function == "hash_min_max") {
    info.kind = B200_HASH_MIN_MAX;
    if ((!is_integer(arg_type.id()) && !is_floating(arg_type.id())) || arg_type.id() == Type::HALF_FLOAT) {
      return Status::NotImplemented("Grouped MinMax data of type ", arg_type.ToString());
    }
    info.out_type = struct_({field("min", arg_type_ptr), field("max", arg_type_ptr)});
    return info;
  }
  return Status::KeyError("No function registered with name: ", function);
}

int32_t OptionOf(int32_t kind, const FunctionOptions* options) {
  if (!options) return 0;
  if (kind == B200_HASH_COUNT) return static_cast<const CountOptions*>(options)->count_mode == CountOptions::COUNT_NULL ? 1 : 0;
  if (kind == B200_HASH_MIN_MAX) return static_cast<const MinMaxOptions*>(options)->null_handling == MinMaxOptions::EMIT_NULL ? 1 : 0;
  return 0;
}

Result<int32_t> AggArgType(int32_t kind, const DataType& type) {
  if (kind == B200_HASH_COUNT) {
    if (type.id() == Type::BOOL) return B200_BOOL;
    return PhysicalType(type);  // only validity is read
  }
  return NumericType(type);
}

struct GroupByHandle {
  DeviceState* dev = nullptr;
  b200_groupby* gb = nullptr;
  ~GroupByHandle() {
    if (gb) {
      std::lock_guard<std::mutex> lock(dev->mutex);
      b200_groupby_destroy(gb);
    }
  }
};

// allocates and fills the aggregate outputs of a handle; returns one ArrayData per aggregate
Result<ArrayDataVector> FinalizeAggregates(GroupByHandle* h, const std::vector<AggInfo>& infos,
                                           const std::vector<std::shared_ptr<DataType>>& arg_types, uint32_t num_groups,
                                           std::vector<b200_out_col>* out_keys) {
  DeviceState* dev = h->dev;
  std::vector<b200_out_col> outs;
  std::vector<std::shared_ptr<Buffer>> data_bufs, validity_bufs;
  for (size_t j = 0; j < infos.size(); ++j) {
    const int n_out = infos[j].kind == B200_HASH_MIN_MAX ? 2 : 1;
    int width = 8;
    if (infos[j].kind == B200_HASH_MIN_MAX) width = checked_cast<const FixedWidthType&>(*arg_types[j]).bit_width() / 8;
    std::shared_ptr<Buffer> validity;
    ARROW_ASSIGN_OR_RAISE(validity, AllocateDevice(dev, BitmapBytes(num_groups)));
    for (int q = 0; q < n_out; ++q) {
      ARROW_ASSIGN_OR_RAISE(auto data, AllocateDevice(dev, int64_t(num_groups) * width));
      b200_out_col o;
      std::memset(&o, 0, sizeof(o));
      o.data = reinterpret_cast<void*>(data->address());
      o.validity = reinterpret_cast<void*>(validity->address());
      outs.push_back(o);
      data_bufs.push_back(std::move(data));
      validity_bufs.push_back(validity);
    }
  }
  B200_OK_OR_RAISE(dev, b200_groupby_finalize(h->gb, outs.data(), out_keys ? out_keys->data() : nullptr));
  ArrayDataVector result;
  size_t oi = 0;
  for (size_t j = 0; j < infos.size(); ++j) {
    if (infos[j].kind == B200_HASH_MIN_MAX) {
      // struct<min,max>, children share one validity (hash_aggregate.cc:741-757)
      auto mins = ArrayData::Make(arg_types[j], num_groups, {validity_bufs[oi], data_bufs[oi]}, outs[oi].null_count);
      auto maxes = ArrayData::Make(arg_types[j], num_groups, {validity_bufs[oi], data_bufs[oi + 1]}, outs[oi].null_count);
      result.push_back(ArrayData::Make(infos[j].out_type, num_groups, {nullptr}, {std::move(mins), std::move(maxes)}, 0));
      oi += 2;
    } else if (infos[j].kind == B200_HASH_COUNT) {
      result.push_back(ArrayData::Make(infos[j].out_type, num_groups, {nullptr, data_bufs[oi]}, 0));  // never null
      oi += 1;
    } else {
      // validity only if some group is empty (hash_aggregate.cc:595-605)
      const int64_t nulls = outs[oi].null_count;
      result.push_back(ArrayData::Make(infos[j].out_type, num_groups,
                                       {nulls ? validity_bufs[oi] : nullptr, data_bufs[oi]}, nulls));
      oi += 1;
    }
  }
  return result;
}

// ---- HashAggregateKernel{init, consume, merge, finalize} (kernel.h:681-721) over dense group ids
struct CudaGroupedAggregator : public KernelState {
  GroupByHandle handle;
  AggInfo info;
  std::shared_ptr<DataType> arg_type;
  uint32_t num_groups = 0;
};

HashAggregateKernel MakeHashKernel(const std::string& function, int device) {
  HashAggregateKernel kernel;
  kernel.init = [function, device](KernelContext*, const KernelInitArgs& args) -> Result<std::unique_ptr<KernelState>> {
    auto state = ::arrow::internal::make_unique<CudaGroupedAggregator>();
    state->arg_type = args.inputs[0].type;
    ARROW_ASSIGN_OR_RAISE(state->info, ResolveAggregate(function, state->arg_type));
    ARROW_ASSIGN_OR_RAISE(state->handle.dev, GetDevice(device));
    b200_agg_spec spec;
    std::memset(&spec, 0, sizeof(spec));
    spec.kind = state->info.kind;
    ARROW_ASSIGN_OR_RAISE(spec.arg_type, AggArgType(spec.kind, *state->arg_type));
    spec.option = OptionOf(spec.kind, args.options);
    const int32_t dummy_key = B200_UINT32;  // ids arrive dense: the key side of the handle is unused
    std::lock_guard<std::mutex> lock(state->handle.dev->mutex);
    B200_OK_OR_RAISE(state->handle.dev, b200_groupby_create(state->handle.dev->ctx, 1, &dummy_key, 1, &spec, 1024, &state->handle.gb));
    return std::move(state);
  };
  kernel.signature = KernelSignature::Make(
      {InputType(ValueDescr::ARRAY), InputType::Array(Type::UINT32), InputType::Scalar(Type::UINT32)},
      OutputType([](KernelContext* ctx, const std::vector<ValueDescr>&) -> Result<ValueDescr> {
        return ValueDescr::Array(checked_cast<CudaGroupedAggregator*>(ctx->state())->info.out_type);
      }));
  kernel.consume = [](KernelContext* ctx, const ExecBatch& batch) -> Status {
    auto* state = checked_cast<CudaGroupedAggregator*>(ctx->state());
    DeviceState* dev = state->handle.dev;
    const ArrayData& arg = *batch[0].array();
    const ArrayData& ids = *batch[1].array();
    const uint32_t num_groups = checked_cast<const UInt32Scalar&>(*batch[2].scalar()).value;
    b200_col c;
    ARROW_ASSIGN_OR_RAISE(const int32_t t, AggArgType(state->info.kind, *arg.type));
    RETURN_NOT_OK(ToCol(arg, t, &c));
    RETURN_NOT_OK(CheckOnDevice(ids.buffers[1], "group id"));
    const uint32_t* id_ptr = reinterpret_cast<const uint32_t*>(ids.buffers[1]->address()) + ids.offset;
    std::lock_guard<std::mutex> lock(dev->mutex);
    B200_OK_OR_RAISE(dev, b200_hash_agg_consume(state->handle.gb, 0, &c, id_ptr, num_groups));
    state->num_groups = std::max(state->num_groups, num_groups);
    return Status::OK();
  };
  kernel.merge = [](KernelContext* ctx, KernelState&& src, KernelState* dst) -> Status {
    // the reference leaves this NotImplemented (hash_aggregate.cc:795-798); here two dense states over the
    // same id space fold slot by slot
    auto* d = checked_cast<CudaGroupedAggregator*>(dst);
    auto& s = checked_cast<CudaGroupedAggregator&>(src);
    std::lock_guard<std::mutex> lock(d->handle.dev->mutex);
    B200_OK_OR_RAISE(d->handle.dev, b200_hash_agg_merge(d->handle.gb, s.handle.gb));
    d->num_groups = std::max(d->num_groups, s.num_groups);
    return Status::OK();
  };
  kernel.finalize = [](KernelContext* ctx, Datum* out) -> Status {
    auto* state = checked_cast<CudaGroupedAggregator*>(ctx->state());
    std::lock_guard<std::mutex> lock(state->handle.dev->mutex);
    ARROW_ASSIGN_OR_RAISE(auto arrays, FinalizeAggregates(&state->handle, {state->info}, {state->arg_type}, state->num_groups, nullptr));
    *out = Datum(arrays[0]);
    return Status::OK();
  };
  return kernel;
}

// ---- Grouper twin ----------------------------------------------------------------------------
class CudaGrouper : public internal::Grouper {
 public:
  static Result<std::unique_ptr<internal::Grouper>> Make(const std::vector<ValueDescr>& descrs, int device) {
    auto g = ::arrow::internal::make_unique<CudaGrouper>();
    ARROW_ASSIGN_OR_RAISE(g->handle_.dev, GetDevice(device));
    for (const auto& d : descrs) {
      // GrouperImpl::Make: "Keys of type ..." (hash_aggregate.cc:325-352)
      ARROW_ASSIGN_OR_RAISE(const int32_t t, KeyPhysicalType(*d.type));
      g->key_types_.push_back(t);
      g->types_.push_back(d.type);
      g->dictionaries_.emplace_back();
    }
    std::lock_guard<std::mutex> lock(g->handle_.dev->mutex);
    B200_OK_OR_RAISE(g->handle_.dev, b200_groupby_create(g->handle_.dev->ctx, int(g->key_types_.size()), g->key_types_.data(), 0, nullptr, 0, &g->handle_.gb));
    return std::unique_ptr<internal::Grouper>(std::move(g));
  }

  Result<Datum> Consume(const ExecBatch& batch) override {
    DeviceState* dev = handle_.dev;
    std::vector<b200_col> keys(batch.num_values());
    if (size_t(batch.num_values()) != key_types_.size()) return Status::Invalid("expected ", key_types_.size(), " key columns");
    for (int k = 0; k < batch.num_values(); ++k) {
      RETURN_NOT_OK(NoteKeyDictionary(*batch[k].array(), &dictionaries_[k]));
      RETURN_NOT_OK(ToCol(*batch[k].array(), key_types_[k], &keys[k]));
    }
    std::lock_guard<std::mutex> lock(dev->mutex);
    ARROW_ASSIGN_OR_RAISE(auto ids, AllocateDevice(dev, batch.length * 4));
    B200_OK_OR_RAISE(dev, b200_grouper_consume(handle_.gb, keys.data(), reinterpret_cast<uint32_t*>(ids->address())));
    B200_OK_OR_RAISE(dev, b200_groupby_num_groups(handle_.gb, &num_groups_));
    return Datum(ArrayData::Make(uint32(), batch.length, {nullptr, std::move(ids)}, 0));
  }

  Result<ExecBatch> GetUniques() override {
    DeviceState* dev = handle_.dev;
    std::lock_guard<std::mutex> lock(dev->mutex);
    ARROW_ASSIGN_OR_RAISE(auto keys, EmitKeys(&handle_, types_, key_types_, num_groups_));
    ExecBatch out({}, num_groups_);
    for (size_t k = 0; k < keys.size(); ++k) {
      if (types_[k]->id() == Type::DICTIONARY) {   // DictionaryKeyEncoder::Decode (hash_aggregate.cc:215-230)
        if (dictionaries_[k]) {
          keys[k]->dictionary = dictionaries_[k];
        } else {
          ARROW_ASSIGN_OR_RAISE(auto empty, MakeArrayOfNull(checked_cast<const DictionaryType&>(*types_[k]).value_type(), 0));
          keys[k]->dictionary = empty->data();
        }
      }
      out.values.emplace_back(std::move(keys[k]));
    }
    return out;
  }

  uint32_t num_groups() const override { return num_groups_; }

  // key columns of a handle in group order: GetUniques (hash_aggregate.cc:412-429); validity only if a key is null
  static Result<ArrayDataVector> EmitKeys(GroupByHandle* h, const std::vector<std::shared_ptr<DataType>>& types,
                                          const std::vector<int32_t>& key_types, uint32_t num_groups) {
    DeviceState* dev = h->dev;
    std::vector<b200_out_col> outs(types.size());
    std::vector<KeyOut> bufs(types.size());
    for (size_t k = 0; k < types.size(); ++k) {
      ARROW_ASSIGN_OR_RAISE(bufs[k], AllocateKeyOut(dev, h->gb, int(k), key_types[k], *types[k], num_groups, &outs[k]));
    }
    B200_OK_OR_RAISE(dev, b200_groupby_finalize(h->gb, nullptr, outs.data()));
    ArrayDataVector result;
    for (size_t k = 0; k < types.size(); ++k) result.push_back(KeyArrayData(types[k], num_groups, bufs[k], outs[k].null_count));
    return result;
  }

 private:
  GroupByHandle handle_;
  std::vector<int32_t> key_types_;
  std::vector<std::shared_ptr<DataType>> types_;
  std::vector<std::shared_ptr<ArrayData>> dictionaries_;
  uint32_t num_groups_ = 0;
};


// ---- unique / value_counts / dictionary_encode (kernels/vector_hash.cc:723-782) ---------------------------------
// Upstream these are VectorKernels whose state (a memo table) is fed chunk by chunk (HashExec) and flushed by a
// finalize step; here one device group-by handle plays the memo table: its group ids ARE the memo indices
// (both are handed out in first-occurrence order), its key column the dictionary.  MetaFunctions, because a
// device ChunkedArray must never be re-sliced or concatenated by the host executor.
const FunctionDoc unique_doc("Compute unique elements (CUDA)", "Distinct values in first-occurrence order.", {"array"});
const FunctionDoc value_counts_doc("Compute counts of unique elements (CUDA)",
                                   "struct<values, counts: int64> of the distinct values in first-occurrence order.", {"array"});
const FunctionDoc dictionary_encode_doc("Dictionary-encode array (CUDA)", "int32 indices into the distinct values.", {"array"},
                                        "DictionaryEncodeOptions");
const auto kDefaultDictionaryEncodeOptions = DictionaryEncodeOptions::Defaults();

enum class HashOp { kUnique, kValueCounts, kDictionaryEncode };

class CudaHashMetaFunction : public MetaFunction {
 public:
  CudaHashMetaFunction(const char* name, HashOp op, const FunctionDoc* doc, const FunctionOptions* defaults, int device)
      : MetaFunction(name, Arity::Unary(), doc, defaults), op_(op), device_(device) {}

  Result<Datum> ExecuteImpl(const std::vector<Datum>& args, const FunctionOptions* options, ExecContext*) const override {
    ARROW_ASSIGN_OR_RAISE(auto chunks, ChunksOf(args[0]));
    const auto type = args[0].type();
    // the index column of a dictionary array is hashed like any integer column (DictionaryHashKernel,
    // vector_hash.cc:520-603); dictionary_encode of a dictionary is not supported upstream either (:775-777)
    const bool is_dict = type->id() == Type::DICTIONARY;
    if (is_dict && op_ == HashOp::kDictionaryEncode) return Status::NotImplemented("dictionary_encode of a dictionary array");
    const std::shared_ptr<DataType> value_type = is_dict ? checked_cast<const DictionaryType&>(*type).index_type() : type;
    const bool is_var = value_type->id() == Type::BINARY || value_type->id() == Type::STRING;
    if (!is_var && value_type->id() != Type::BOOL && !is_primitive(value_type->id())) {
      return Status::NotImplemented("CUDA-backed ", name(), " handles fixed-width primitive and binary / utf8 columns, got ", type->ToString());
    }
    ARROW_ASSIGN_OR_RAISE(const int32_t ktype, KeyPhysicalType(*value_type));
    std::shared_ptr<ArrayData> dictionary;
    if (is_dict) {
      for (const auto& c : chunks) {
        if (!dictionary) dictionary = c->data()->dictionary;
        // "Only hashing for data with equal dictionaries currently supported" (vector_hash.cc:549-552); device
        // dictionaries are compared by identity: their values cannot be read here
        else if (c->data()->dictionary.get() != dictionary.get()) {
          return Status::Invalid("Only hashing for data with equal dictionaries currently supported");
        }
      }
    }
    GroupByHandle handle;
    ARROW_ASSIGN_OR_RAISE(handle.dev, GetDevice(device_));
    DeviceState* dev = handle.dev;
    std::unique_lock<std::mutex> lock(dev->mutex);
    b200_agg_spec count_all;
    std::memset(&count_all, 0, sizeof(count_all));
    count_all.kind = B200_HASH_COUNT;
    count_all.arg_type = is_var ? B200_INT32 : ktype;   // (a binary column is counted through its offsets buffer)
    const bool counts = op_ == HashOp::kValueCounts;
    B200_OK_OR_RAISE(dev, b200_groupby_create(dev->ctx, 1, &ktype, counts ? 1 : 0, counts ? &count_all : nullptr, 0, &handle.gb));
    std::vector<std::shared_ptr<Buffer>> id_chunks;
    uint64_t row_base = 0;
    for (const auto& c : chunks) {
      b200_col key;
      RETURN_NOT_OK(ToCol(*c->data(), ktype, &key));
      if (op_ == HashOp::kDictionaryEncode) {
        ARROW_ASSIGN_OR_RAISE(auto ids, AllocateDevice(dev, c->length() * 4));
        if (c->length() > 0) B200_OK_OR_RAISE(dev, b200_grouper_consume(handle.gb, &key, reinterpret_cast<uint32_t*>(ids->address())));
        id_chunks.push_back(std::move(ids));
      } else if (c->length() > 0) {
        // value_counts counts every row of a group, the null group's too (ValueCountsAction::ObserveNullFound,
        // vector_hash.cc:120-163): the counted argument is the key column read without its validity
        b200_col all_rows = key;
        all_rows.validity = nullptr;
        all_rows.null_count = 0;
        all_rows.type = count_all.arg_type;
        all_rows.extra = nullptr;
        B200_OK_OR_RAISE(dev, b200_groupby_consume(handle.gb, &key, counts ? &all_rows : nullptr, row_base));
      }
      row_base += uint64_t(c->length());
    }
    uint32_t num_groups = 0;
    B200_OK_OR_RAISE(dev, b200_groupby_num_groups(handle.gb, &num_groups));
    std::shared_ptr<ArrayData> uniques, count_data;
    if (counts) {
      std::vector<b200_out_col> key_outs(1);
      ARROW_ASSIGN_OR_RAISE(auto kbuf, AllocateKeyOut(dev, handle.gb, 0, ktype, *value_type, num_groups, &key_outs[0]));
      AggInfo info;
      info.kind = B200_HASH_COUNT;
      info.out_type = int64();
      ARROW_ASSIGN_OR_RAISE(auto aggs, FinalizeAggregates(&handle, {info}, {value_type}, num_groups, &key_outs));
      count_data = aggs[0];
      uniques = KeyArrayData(type, num_groups, kbuf, key_outs[0].null_count);
    } else {
      ARROW_ASSIGN_OR_RAISE(auto keys, CudaGrouper::EmitKeys(&handle, {type}, {ktype}, num_groups));
      uniques = keys[0];
    }
    if (is_dict) uniques->dictionary = dictionary;
    if (op_ == HashOp::kUnique) {
      lock.unlock();
      return Datum(uniques);
    }
    if (op_ == HashOp::kValueCounts) {
      lock.unlock();
      auto out_type = struct_({field(kValuesFieldName, type), field(kCountsFieldName, int64())});
      return Datum(ArrayData::Make(std::move(out_type), num_groups, {nullptr}, {uniques, count_data}, 0));
    }
    // ---- dictionary_encode: DictEncodeAction + DictEncodeFinalize (vector_hash.cc:165-226, 624-636)
    const auto* encode_options = static_cast<const DictionaryEncodeOptions*>(options);
    const bool mask_nulls = !encode_options || encode_options->null_encoding_behavior == DictionaryEncodeOptions::MASK;
    int64_t null_group = -1;
    std::shared_ptr<ArrayData> dict_values = uniques;
    if (mask_nulls && uniques->null_count.load() != 0) {
      // where the one null entry sits: num_groups validity bits to the host; the dictionary is the key column without it
      std::vector<uint8_t> bits(size_t(BitmapBytes(num_groups)));
      B200_OK_OR_RAISE(dev, b200_ctx_synchronize(dev->ctx));
      ARROW_ASSIGN_OR_RAISE(auto cuda_validity, ::arrow::cuda::CudaBuffer::FromBuffer(uniques->buffers[0]));
      RETURN_NOT_OK(cuda_validity->CopyToHost(0, int64_t(bits.size()), bits.data()));
      for (uint32_t g = 0; g < num_groups; ++g)
        if (!BitUtil::GetBit(bits.data(), g)) {
          null_group = g;
          break;
        }
      if (is_var) {
        // the null entry holds no bytes: the value buffer stays, the offsets lose entry null_group + 1
        // (== entry null_group).  Two device-to-device copies.
        ARROW_ASSIGN_OR_RAISE(auto offsets, AllocateDevice(dev, int64_t(num_groups) * 4));
        const uintptr_t src = uniques->buffers[1]->address();
        const int64_t head = (null_group + 1) * 4, tail = (int64_t(num_groups) - 1 - null_group) * 4;
        ARROW_ASSIGN_OR_RAISE(auto cuda_offsets, ::arrow::cuda::CudaBuffer::FromBuffer(offsets));
        RETURN_NOT_OK(cuda_offsets->CopyFromDevice(0, reinterpret_cast<const void*>(src), head));
        if (tail > 0) RETURN_NOT_OK(cuda_offsets->CopyFromDevice(head, reinterpret_cast<const void*>(src + head + 4), tail));
        dict_values = ArrayData::Make(type, int64_t(num_groups) - 1, {nullptr, std::move(offsets), uniques->buffers[2]}, 0);
      } else {
        auto as_mask = ArrayData::Make(boolean(), num_groups, {nullptr, uniques->buffers[0]}, 0);
        lock.unlock();
        ARROW_ASSIGN_OR_RAISE(dict_values, FilterArray(dev, *uniques, *as_mask, 0 /* DROP */));
        lock.lock();
        dict_values->buffers[0] = nullptr;   // every entry left is valid
        dict_values->null_count = 0;
      }
    }
    auto dict_type = ::arrow::dictionary(int32(), type);
    ArrayVector out_chunks;
    for (size_t i = 0; i < chunks.size(); ++i) {
      const int64_t n = chunks[i]->length();
      // zero-length chunks of a ChunkedArray yield no output chunk (ExecBatchIterator skips them, exec.cc:221-229)
      if (n == 0 && args[0].kind() == Datum::CHUNKED_ARRAY) continue;
      std::shared_ptr<Buffer> validity;
      int64_t nulls = 0;
      if (null_group >= 0) {
        ARROW_ASSIGN_OR_RAISE(validity, AllocateDevice(dev, (n + 31) / 32 * 4));
        ARROW_ASSIGN_OR_RAISE(auto indices, AllocateDevice(dev, n * 4));
        B200_OK_OR_RAISE(dev, b200_dictionary_indices(dev->ctx, reinterpret_cast<const uint32_t*>(id_chunks[i]->address()), n, null_group,
                                                      reinterpret_cast<int32_t*>(indices->address()), reinterpret_cast<void*>(validity->address())));
        id_chunks[i] = std::move(indices);
        nulls = chunks[i]->data()->buffers[0] ? chunks[i]->data()->null_count.load() : 0;   // may stay unknown, like the input's
        if (nulls == 0) validity = nullptr;
      }
      // ENCODE, or no null anywhere: the ids are the indices (uint32 ids < 2^31 read as int32)
      auto data = ArrayData::Make(dict_type, n, {std::move(validity), id_chunks[i]}, nulls);
      data->dictionary = dict_values;
      out_chunks.push_back(MakeArray(std::move(data)));
    }
    B200_OK_OR_RAISE(dev, b200_ctx_synchronize(dev->ctx));
    lock.unlock();
    if (args[0].kind() == Datum::ARRAY) return Datum(out_chunks[0]);
    for (auto& c : out_chunks) RETURN_NOT_OK(SettleNullCount(dev, c->data().get()));
    return Datum(std::make_shared<ChunkedArray>(std::move(out_chunks), dict_type));
  }

 private:
  HashOp op_;
  int device_;
};

}  // namespace

// ==================================================================================================
Status RegisterCudaKernels(FunctionRegistry* registry, int device) {
  if (!registry) return Status::Invalid("null registry");
  // array_filter: VectorFunction, one kernel over every primitive type (vector_selection.cc:2133)
  {
    auto func = std::make_shared<VectorFunction>("array_filter", Arity::Binary(), &array_filter_doc, &kDefaultFilterOptions);
    VectorKernel kernel;
    // outputs are device buffers the executor cannot allocate or pre-fill
    kernel.mem_allocation = MemAllocation::NO_PREALLOCATE;
    kernel.null_handling = NullHandling::COMPUTED_NO_PREALLOCATE;
    kernel.can_execute_chunkwise = false;  // never re-slice device arrays on the host
    kernel.init = [](KernelContext*, const KernelInitArgs& args) -> Result<std::unique_ptr<KernelState>> {
      return std::unique_ptr<KernelState>(new FilterState(NullSelectionOf(args.options)));
    };
    kernel.signature = KernelSignature::Make(
        {InputType(match::Primitive(), ValueDescr::ARRAY), InputType::Array(boolean())},
        OutputType([](KernelContext*, const std::vector<ValueDescr>& descrs) -> Result<ValueDescr> { return descrs[0]; }));
    kernel.exec = [device](KernelContext* ctx, const ExecBatch& batch, Datum* out) -> Status {
      ARROW_ASSIGN_OR_RAISE(DeviceState * dev, GetDevice(device));
      const int null_selection = checked_cast<const FilterState*>(ctx->state())->null_selection;
      ARROW_ASSIGN_OR_RAISE(auto result, FilterArray(dev, *batch[0].array(), *batch[1].array(), null_selection));
      out->value = std::move(result);
      return Status::OK();
    };
    RETURN_NOT_OK(func->AddKernel(std::move(kernel)));
    RETURN_NOT_OK(registry->AddFunction(std::move(func)));
  }
  RETURN_NOT_OK(registry->AddFunction(std::make_shared<CudaFilterMetaFunction>(device)));
  // array_take: VectorFunction over (primitive array, integer array) (vector_selection.cc:2154-2176)
  {
    auto func = std::make_shared<VectorFunction>("array_take", Arity::Binary(), &array_take_doc, &kDefaultTakeOptions);
    VectorKernel kernel;
    kernel.mem_allocation = MemAllocation::NO_PREALLOCATE;
    kernel.null_handling = NullHandling::COMPUTED_NO_PREALLOCATE;
    kernel.can_execute_chunkwise = false;
    kernel.init = [](KernelContext*, const KernelInitArgs& args) -> Result<std::unique_ptr<KernelState>> {
      const bool boundscheck = args.options ? static_cast<const TakeOptions*>(args.options)->boundscheck : true;
      return std::unique_ptr<KernelState>(new TakeStateB200(boundscheck));
    };
    kernel.signature = KernelSignature::Make(
        {InputType(match::Primitive(), ValueDescr::ARRAY), InputType(match::Integer(), ValueDescr::ARRAY)},
        OutputType([](KernelContext*, const std::vector<ValueDescr>& descrs) -> Result<ValueDescr> { return descrs[0]; }));
    kernel.exec = [device](KernelContext* ctx, const ExecBatch& batch, Datum* out) -> Status {
      ARROW_ASSIGN_OR_RAISE(DeviceState * dev, GetDevice(device));
      const bool boundscheck = checked_cast<const TakeStateB200*>(ctx->state())->boundscheck;
      ARROW_ASSIGN_OR_RAISE(auto result, TakeArray(dev, *batch[0].array(), *batch[1].array(), boundscheck));
      out->value = std::move(result);
      return Status::OK();
    };
    RETURN_NOT_OK(func->AddKernel(std::move(kernel)));
    RETURN_NOT_OK(registry->AddFunction(std::move(func)));
  }
  RETURN_NOT_OK(registry->AddFunction(std::make_shared<CudaTakeMetaFunction>(device)));
  // array_sort_indices: one kernel per numeric / temporal type (vector_sort.cc:582-595)
  {
    auto func = std::make_shared<VectorFunction>("array_sort_indices", Arity::Unary(), &array_sort_indices_doc, &kDefaultArraySortOptions);
    VectorKernel base;
    base.mem_allocation = MemAllocation::NO_PREALLOCATE;
    base.null_handling = NullHandling::OUTPUT_NOT_NULL;
    base.can_execute_chunkwise = false;
    base.init = [](KernelContext*, const KernelInitArgs& args) -> Result<std::unique_ptr<KernelState>> {
      const auto* options = static_cast<const ArraySortOptions*>(args.options);
      return std::unique_ptr<KernelState>(new SortState(options ? options->order : SortOrder::Ascending));
    };
    base.exec = [device](KernelContext* ctx, const ExecBatch& batch, Datum* out) -> Status {
      ARROW_ASSIGN_OR_RAISE(DeviceState * dev, GetDevice(device));
      ARROW_ASSIGN_OR_RAISE(auto result, SortIndicesArray(dev, *batch[0].array(), checked_cast<const SortState*>(ctx->state())->order));
      out->value = std::move(result);
      return Status::OK();
    };
    std::vector<std::shared_ptr<DataType>> types = {uint8(), int8(), uint16(), int16(), uint32(), int32(), uint64(), int64(), float32(), float64(), date32(), date64()};
    for (const auto& ty : types) {
      base.signature = KernelSignature::Make({InputType::Array(ty)}, uint64());
      RETURN_NOT_OK(func->AddKernel(base));
    }
    for (auto id : {Type::TIMESTAMP, Type::TIME32, Type::TIME64, Type::DURATION}) {
      base.signature = KernelSignature::Make({InputType::Array(id)}, uint64());
      RETURN_NOT_OK(func->AddKernel(base));
    }
    RETURN_NOT_OK(registry->AddFunction(std::move(func)));
  }
  RETURN_NOT_OK(registry->AddFunction(std::make_shared<CudaSortIndicesMetaFunction>(device)));
  // hash aggregates (hash_aggregate.cc:1039-1062) + hash_mean
  {
    static auto default_count_options = CountOptions::Defaults();
    static auto default_minmax_options = MinMaxOptions::Defaults();
    struct Entry { const char* name; const FunctionDoc* doc; const FunctionOptions* defaults; };
    const Entry entries[] = {{"hash_count", &hash_count_doc, &default_count_options},
                             {"hash_sum", &hash_sum_doc, nullptr},
                             {"hash_min_max", &hash_min_max_doc, &default_minmax_options},
                             {"hash_mean", &hash_mean_doc, nullptr}};
    for (const auto& e : entries) {
      auto func = std::make_shared<HashAggregateFunction>(e.name, Arity::Ternary(), e.doc, e.defaults);
      RETURN_NOT_OK(func->AddKernel(MakeHashKernel(e.name, device)));
      RETURN_NOT_OK(registry->AddFunction(std::move(func)));
    }
  }
  // unique / value_counts / dictionary_encode (vector_hash.cc:723-782)
  RETURN_NOT_OK(registry->AddFunction(std::make_shared<CudaHashMetaFunction>("unique", HashOp::kUnique, &unique_doc, nullptr, device)));
  RETURN_NOT_OK(registry->AddFunction(std::make_shared<CudaHashMetaFunction>("value_counts", HashOp::kValueCounts, &value_counts_doc, nullptr, device)));
  RETURN_NOT_OK(registry->AddFunction(std::make_shared<CudaHashMetaFunction>("dictionary_encode", HashOp::kDictionaryEncode,
                                                                             &dictionary_encode_doc, &kDefaultDictionaryEncodeOptions, device)));
  return Status::OK();
}

Result<std::unique_ptr<FunctionRegistry>> MakeCudaRegistry(int device) {
  auto registry = FunctionRegistry::Make();
  RETURN_NOT_OK(RegisterCudaKernels(registry.get(), device));
  return std::move(registry);
}


// ---- ingest / egress ----------------------------------------------------------------------------
// The reference's own copies (CudaBuffer::CopyFromHost / CopyToHost, gpu/cuda_memory.h:64-71) are synchronous cuMemcpy
// calls from pageable memory: 3.4 GB/s measured on the B200 box, 240 ms for the 812 MB of a 100 M-row int64 column.
// Here a host buffer reaches the device (a) straight by DMA when it is pinned (a CudaHostBuffer, or anything
// cuMemHostRegister'ed: cuPointerGetAttribute tells), (b) through two pinned staging blocks otherwise, the CPU filling
// one while the copy engine drains the other.  Everything runs on one private copy stream and is complete on return.
namespace {
#define CU_OK_OR_RAISE(expr)                                                                  \
  do {                                                                                        \
    CUresult _r = (expr);                                                                     \
    if (_r != CUDA_SUCCESS) {                                                                 \
      const char* _m = nullptr;                                                               \
      cuGetErrorString(_r, &_m);                                                              \
      return Status::IOError("CUDA driver call failed: " #expr ": ", _m ? _m : "unknown");   \
    }                                                                                         \
  } while (0)

Status EnsureCopyEngine(DeviceState* dev) {
  if (dev->copy_stream) return Status::OK();
  CU_OK_OR_RAISE(cuCtxPushCurrent(reinterpret_cast<CUcontext>(dev->arrow_ctx->handle())));
  CUresult r = cuStreamCreate(&dev->copy_stream, CU_STREAM_NON_BLOCKING);
  for (int i = 0; i < 2 && r == CUDA_SUCCESS; ++i) {
    r = cuMemHostAlloc(&dev->staging[i], size_t(kStagingBytes), CU_MEMHOSTALLOC_PORTABLE);
    if (r == CUDA_SUCCESS) r = cuEventCreate(&dev->staging_done[i], CU_EVENT_DISABLE_TIMING);
  }
  CUcontext popped;
  cuCtxPopCurrent(&popped);
  CU_OK_OR_RAISE(r);
  return Status::OK();
}

bool IsPinned(const void* p) {
  unsigned int type = 0;
  return cuPointerGetAttribute(&type, CU_POINTER_ATTRIBUTE_MEMORY_TYPE, reinterpret_cast<CUdeviceptr>(p)) == CUDA_SUCCESS &&
         type == CU_MEMORYTYPE_HOST;
}

// host -> device (to_device) or device -> host of `nbytes`; the caller holds dev->mutex
Status CopyStaged(DeviceState* dev, bool to_device, uintptr_t device_addr, uint8_t* host, int64_t nbytes) {
  if (nbytes == 0) return Status::OK();
  RETURN_NOT_OK(EnsureCopyEngine(dev));
  CU_OK_OR_RAISE(cuCtxPushCurrent(reinterpret_cast<CUcontext>(dev->arrow_ctx->handle())));
  CUresult r = CUDA_SUCCESS;
  if (IsPinned(host)) {
    r = to_device ? cuMemcpyHtoDAsync(device_addr, host, size_t(nbytes), dev->copy_stream)
                  : cuMemcpyDtoHAsync(host, device_addr, size_t(nbytes), dev->copy_stream);
  } else if (!to_device) {
    r = cuMemcpyDtoH(host, device_addr, size_t(nbytes));   // (CopyToHost always lands in pinned memory; kept for completeness)
  } else {
    int64_t done = 0;
    for (int i = 0; done < nbytes && r == CUDA_SUCCESS; ++i) {
      const int b = i & 1;
      const int64_t len = std::min(kStagingBytes, nbytes - done);
      if (i >= 2) r = cuEventSynchronize(dev->staging_done[b]);   // the block's previous copy has left it
      if (r != CUDA_SUCCESS) break;
      std::memcpy(dev->staging[b], host + done, size_t(len));
      r = cuMemcpyHtoDAsync(device_addr + done, dev->staging[b], size_t(len), dev->copy_stream);
      if (r == CUDA_SUCCESS) r = cuEventRecord(dev->staging_done[b], dev->copy_stream);
      done += len;
    }
  }
  if (r == CUDA_SUCCESS) r = cuStreamSynchronize(dev->copy_stream);
  CUcontext popped;
  cuCtxPopCurrent(&popped);
  CU_OK_OR_RAISE(r);
  return Status::OK();
}

Result<std::shared_ptr<ArrayData>> MoveArrayData(DeviceState* dev, const ArrayData& a, bool to_device) {
  auto out = a.Copy();
  for (auto& buf : out->buffers) {
    if (!buf) continue;
    if (to_device) {
      if (!buf->is_cpu()) continue;   // already there
      ARROW_ASSIGN_OR_RAISE(auto dbuf, AllocateDevice(dev, buf->size()));
      RETURN_NOT_OK(CopyStaged(dev, true, dbuf->address(), const_cast<uint8_t*>(buf->data()), buf->size()));
      buf = std::move(dbuf);
    } else {
      if (buf->is_cpu()) continue;
      ARROW_ASSIGN_OR_RAISE(auto hbuf, AllocatePinnedPooled(dev, buf->size()));
      RETURN_NOT_OK(CopyStaged(dev, false, buf->address(), hbuf->mutable_data(), buf->size()));
      buf = std::move(hbuf);
    }
  }
  for (auto& child : out->child_data) { ARROW_ASSIGN_OR_RAISE(child, MoveArrayData(dev, *child, to_device)); }
  return out;
}
}  // namespace

Result<std::shared_ptr<ArrayData>> CopyToDevice(const ArrayData& host, int device) {
  ARROW_ASSIGN_OR_RAISE(DeviceState * dev, GetDevice(device));
  std::lock_guard<std::mutex> lock(dev->mutex);
  // a device-resident ArrayData must carry its null count (array/data.cc:119-131 would scan the bitmap on the host)
  auto settled = host.Copy();
  (void)settled->GetNullCount();
  return MoveArrayData(dev, *settled, true);
}

Result<std::shared_ptr<ArrayData>> CopyToHost(const ArrayData& device_data, int device) {
  ARROW_ASSIGN_OR_RAISE(DeviceState * dev, GetDevice(device));
  std::lock_guard<std::mutex> lock(dev->mutex);
  return MoveArrayData(dev, device_data, false);
}

Result<std::shared_ptr<ArrayData>> ImportToDevice(struct ArrowArray* array, struct ArrowSchema* schema, int device) {
  ARROW_ASSIGN_OR_RAISE(auto host, ImportArray(array, schema));   // c/bridge.cc: takes ownership of both structs
  return CopyToDevice(*host->data(), device);
}

Status ExportFromDevice(const ArrayData& device_data, struct ArrowArray* out, struct ArrowSchema* out_schema, int device) {
  ARROW_ASSIGN_OR_RAISE(auto host, CopyToHost(device_data, device));
  return ExportArray(*MakeArray(host), out, out_schema);
}

Result<std::shared_ptr<Buffer>> AllocatePinned(int64_t nbytes, int device) {
  ARROW_ASSIGN_OR_RAISE(DeviceState * dev, GetDevice(device));
  return AllocatePinnedPooled(dev, nbytes);
}

Result<std::unique_ptr<internal::Grouper>> MakeGrouper(const std::vector<ValueDescr>& descrs, ExecContext*, int device) {
  return CudaGrouper::Make(descrs, device);
}

// == Grouper::MakeGroupings (hash_aggregate.cc:980-1018): the counting sort of the ids runs on the device
Result<std::shared_ptr<ListArray>> MakeGroupings(const ArrayData& ids, uint32_t num_groups, int device) {
  if (ids.type->id() != Type::UINT32) return Status::TypeError("MakeGroupings takes uint32 group ids, got ", ids.type->ToString());
  if (ids.buffers[0] && ids.null_count.load() != 0) return Status::Invalid("MakeGroupings with null ids");
  ARROW_ASSIGN_OR_RAISE(DeviceState * dev, GetDevice(device));
  b200_col c;
  RETURN_NOT_OK(ToCol(ids, B200_UINT32, &c));
  std::lock_guard<std::mutex> lock(dev->mutex);
  ARROW_ASSIGN_OR_RAISE(auto offsets, AllocateDevice(dev, (int64_t(num_groups) + 1) * 4));
  ARROW_ASSIGN_OR_RAISE(auto indices, AllocateDevice(dev, ids.length * 4));
  size_t scratch_bytes = 0;
  B200_OK_OR_RAISE(dev, b200_make_groupings_scratch_bytes(ids.length, &scratch_bytes));
  ARROW_ASSIGN_OR_RAISE(auto scratch, AllocateDevice(dev, int64_t(scratch_bytes)));
  B200_OK_OR_RAISE(dev, b200_make_groupings(dev->ctx, reinterpret_cast<const uint32_t*>(ids.buffers[1]->address()) + ids.offset, ids.length,
                                            num_groups, reinterpret_cast<int32_t*>(offsets->address()),
                                            reinterpret_cast<int32_t*>(indices->address()), reinterpret_cast<void*>(scratch->address()),
                                            scratch_bytes));
  auto values = ArrayData::Make(int32(), ids.length, {nullptr, std::move(indices)}, 0);
  auto data = ArrayData::Make(list(int32()), num_groups, {nullptr, std::move(offsets)}, {std::move(values)}, 0);
  return std::make_shared<ListArray>(std::move(data));
}

// == Grouper::ApplyGroupings (hash_aggregate.cc:968-978): Take without bounds check under the same offsets
Result<std::shared_ptr<ListArray>> ApplyGroupings(const ListArray& groupings, const ArrayData& array, int device) {
  ARROW_ASSIGN_OR_RAISE(DeviceState * dev, GetDevice(device));
  ARROW_ASSIGN_OR_RAISE(auto sorted, TakeArray(dev, array, *groupings.data()->child_data[0], /*boundscheck=*/false));
  auto data = ArrayData::Make(list(array.type), groupings.length(), {nullptr, groupings.data()->buffers[1]}, {std::move(sorted)}, 0,
                              groupings.offset());
  return std::make_shared<ListArray>(std::move(data));
}

Result<Datum> GroupBy(const std::vector<Datum>& arguments, const std::vector<Datum>& keys,
                      const std::vector<internal::Aggregate>& aggregates, ExecContext* ctx) {
  return GroupBy(arguments, keys, aggregates, Datum(), ctx);
}

Result<Datum> GroupBy(const std::vector<Datum>& arguments, const std::vector<Datum>& keys,
                      const std::vector<internal::Aggregate>& aggregates, const Datum& selection, ExecContext* ctx) {
  if (!ctx) return Status::Invalid("cuda::GroupBy needs an ExecContext carrying the CUDA registry");
  // GetKernels (hash_aggregate.cc:809-830)
  if (aggregates.size() != arguments.size()) {
    return Status::Invalid(aggregates.size(), " aggregate functions were specified but ", arguments.size(), " arguments were provided.");
  }
  if (keys.empty()) return Status::Invalid("GroupBy needs at least one key");
  std::vector<AggInfo> infos(aggregates.size());
  std::vector<b200_agg_spec> specs(aggregates.size());
  std::vector<std::shared_ptr<DataType>> arg_types(aggregates.size());
  for (size_t i = 0; i < aggregates.size(); ++i) {
    ARROW_ASSIGN_OR_RAISE(auto function, ctx->func_registry()->GetFunction(aggregates[i].function));  // KeyError if absent
    if (function->kind() != Function::HASH_AGGREGATE) return Status::Invalid(aggregates[i].function, " is not a hash aggregate function");
    arg_types[i] = arguments[i].type();
    ARROW_ASSIGN_OR_RAISE(infos[i], ResolveAggregate(aggregates[i].function, arg_types[i]));
    const FunctionOptions* options = aggregates[i].options ? aggregates[i].options : function->default_options();
    std::memset(&specs[i], 0, sizeof(b200_agg_spec));
    specs[i].kind = infos[i].kind;
    ARROW_ASSIGN_OR_RAISE(specs[i].arg_type, AggArgType(infos[i].kind, *arg_types[i]));
    specs[i].option = OptionOf(infos[i].kind, options);
  }
  std::vector<int32_t> key_types;
  std::vector<std::shared_ptr<DataType>> key_dtypes;
  for (const auto& k : keys) {
    ARROW_ASSIGN_OR_RAISE(const int32_t t, KeyPhysicalType(*k.type()));
    key_types.push_back(t);
    key_dtypes.push_back(k.type());
  }
  // every argument and key shares one chunk layout (the reference re-chunks on the host with
  // ExecBatchIterator; device chunks are consumed as they are, batch by batch)
  std::vector<ArrayVector> arg_chunks(arguments.size()), key_chunks(keys.size());
  for (size_t i = 0; i < arguments.size(); ++i) { ARROW_ASSIGN_OR_RAISE(arg_chunks[i], ChunksOf(arguments[i])); }
  for (size_t i = 0; i < keys.size(); ++i) { ARROW_ASSIGN_OR_RAISE(key_chunks[i], ChunksOf(keys[i])); }
  const size_t n_batches = key_chunks[0].size();
  const int64_t total = keys[0].length();
  for (size_t i = 0; i < arguments.size(); ++i)
    if (arguments[i].length() != total) return Status::Invalid("Array arguments must all be the same length");
  for (size_t i = 0; i < keys.size(); ++i)
    if (keys[i].length() != total) return Status::Invalid("Array arguments must all be the same length");
  for (const auto& c : arg_chunks) if (c.size() != n_batches) return Status::NotImplemented("GroupBy: arguments and keys must share one chunk layout");
  for (const auto& c : key_chunks) if (c.size() != n_batches) return Status::NotImplemented("GroupBy: arguments and keys must share one chunk layout");

  // the selection: a device-resident boolean column in the keys' chunk layout, no null slots (a
  // filter with nulls drops or emits them by FilterOptions, vector_selection.cc:786-835 — that is
  // Filter's business), every chunk's bitmap at a 32-bit word boundary
  const bool selected = selection.kind() != Datum::NONE;
  ArrayVector sel_chunks;
  if (selected) {
    if (selection.type()->id() != Type::BOOL) return Status::TypeError("GroupBy selection must be boolean, got ", selection.type()->ToString());
    if (selection.length() != total) return Status::Invalid("Array arguments must all be the same length");
    ARROW_ASSIGN_OR_RAISE(sel_chunks, ChunksOf(selection));
    if (sel_chunks.size() != n_batches) return Status::NotImplemented("GroupBy: selection and keys must share one chunk layout");
    for (const auto& c : sel_chunks) {
      const ArrayData& d = *c->data();
      if (d.buffers[0] && d.null_count.load() != 0) return Status::NotImplemented("GroupBy: selection with null slots (run Filter first)");
      if (d.offset % 32 != 0 || (d.buffers[1] && d.buffers[1]->address() % 4 != 0))
        return Status::NotImplemented("GroupBy: selection bitmap must start at a 32-bit word boundary");
    }
  }

  GroupByHandle handle;
  ARROW_ASSIGN_OR_RAISE(handle.dev, GetDevice(0));
  DeviceState* dev = handle.dev;
  std::unique_lock<std::mutex> lock(dev->mutex);
  B200_OK_OR_RAISE(dev, b200_groupby_create(dev->ctx, int(key_types.size()), key_types.data(), int(specs.size()), specs.data(), 0, &handle.gb));
  uint64_t row_base = 0;
  std::vector<std::shared_ptr<ArrayData>> key_dictionaries(keys.size());
  for (size_t b = 0; b < n_batches; ++b) {
    const int64_t len = key_chunks[0][b]->length();
    if (len == 0) continue;
    std::vector<b200_col> kcols(keys.size()), acols(arguments.size());
    for (size_t k = 0; k < keys.size(); ++k) {
      if (key_chunks[k][b]->length() != len) return Status::NotImplemented("GroupBy: arguments and keys must share one chunk layout");
      RETURN_NOT_OK(NoteKeyDictionary(*key_chunks[k][b]->data(), &key_dictionaries[k]));
      RETURN_NOT_OK(ToCol(*key_chunks[k][b]->data(), key_types[k], &kcols[k]));
    }
    for (size_t i = 0; i < arguments.size(); ++i) {
      if (arg_chunks[i][b]->length() != len) return Status::NotImplemented("GroupBy: arguments and keys must share one chunk layout");
      RETURN_NOT_OK(ToCol(*arg_chunks[i][b]->data(), specs[i].arg_type, &acols[i]));
    }
    if (selected) {
      const ArrayData& d = *sel_chunks[b]->data();
      if (d.length != len) return Status::NotImplemented("GroupBy: selection and keys must share one chunk layout");
      const void* bits = reinterpret_cast<const void*>(d.buffers[1]->address() + d.offset / 8);
      B200_OK_OR_RAISE(dev, b200_groupby_consume_selected(handle.gb, kcols.data(), acols.data(), row_base, bits));
    } else {
      B200_OK_OR_RAISE(dev, b200_groupby_consume(handle.gb, kcols.data(), acols.data(), row_base));
    }
    row_base += uint64_t(len);
  }
  uint32_t num_groups = 0;
  B200_OK_OR_RAISE(dev, b200_groupby_num_groups(handle.gb, &num_groups));

  // outputs: aggregates then keys (hash_aggregate.cc:946-965)
  std::vector<b200_out_col> key_outs(keys.size());
  std::vector<KeyOut> key_bufs(keys.size());
  for (size_t k = 0; k < keys.size(); ++k) {
    ARROW_ASSIGN_OR_RAISE(key_bufs[k], AllocateKeyOut(dev, handle.gb, int(k), key_types[k], *key_dtypes[k], num_groups, &key_outs[k]));
  }
  ARROW_ASSIGN_OR_RAISE(auto out_data, FinalizeAggregates(&handle, infos, arg_types, num_groups, &key_outs));
  FieldVector out_fields;
  for (size_t i = 0; i < aggregates.size(); ++i) out_fields.push_back(field(aggregates[i].function, infos[i].out_type));
  for (size_t k = 0; k < keys.size(); ++k) {
    out_data.push_back(KeyArrayData(key_dtypes[k], num_groups, key_bufs[k], key_outs[k].null_count));
    if (key_dtypes[k]->id() == Type::DICTIONARY) {
      if (key_dictionaries[k]) {
        out_data.back()->dictionary = key_dictionaries[k];
      } else {
        ARROW_ASSIGN_OR_RAISE(auto empty, MakeArrayOfNull(checked_cast<const DictionaryType&>(*key_dtypes[k]).value_type(), 0));
        out_data.back()->dictionary = empty->data();
      }
    }
    out_fields.push_back(field("key_" + std::to_string(k), key_dtypes[k]));
  }
  lock.unlock();  // ~GroupByHandle takes the device mutex again
  return Datum(ArrayData::Make(struct_(std::move(out_fields)), num_groups, {nullptr}, std::move(out_data), 0));
}

}  // namespace cuda
}  // namespace compute
}  // namespace arrow
