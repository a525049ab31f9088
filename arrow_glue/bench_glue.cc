// bench_glue — timing of the hot path THROUGH THE ARROW REGISTRY (compute::CallFunction / cuda::GroupBy of
// arrow_glue/b200_arrow.cc over CudaBuffer-resident ArrayData), the path a C++ user of the reference takes.
//
//   bench_glue [workload] [rows] [steps]       workload: sort_i64 | filter | groupby_c3       (one JSON line on stdout)
//
// Two numbers per workload, both with CUDA-event-free wall clocks around device-synchronising calls (every registry
// call returns after its results are on the device and the context is synchronised):
//   device   inputs already CudaBuffers; one registry call per step
//   e2e      host Arrow arrays in, host Arrow array out, per step: cuda::CopyToDevice of every input column, the
//            registry call, cuda::CopyToHost of the result (arrow_glue's ingest / egress: DMA from / to pinned
//            memory, pooled device blocks).  The input columns live in pinned host buffers (cuda::AllocatePinned), as
//            a pipeline that feeds the GPU would keep them.
//   e2e_ref  the same with the reference's own copies, CudaBuffer::CopyFromHost / CopyToHost (gpu/cuda_memory.h:64-71:
//            synchronous cuMemcpy, here from the same pinned buffers), for comparison
// The columns are the generator of arrow1_b200/synth.py restated in C++ (splitmix64 of seed * GOLDEN + row).
// Built by arrow_glue/build.sh where /root/reference exists; shipped prebuilt to the GPU box.
#include <chrono>
#include <cmath>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "arrow/api.h"
#include "arrow/compute/api.h"
#include "arrow/gpu/cuda_api.h"
#include "arrow/util/bit_util.h"
#include "b200_arrow.h"

using namespace arrow;  // NOLINT
namespace cp = arrow::compute;

#define GET(expr) Unwrap((expr), #expr)
template <typename T>
T Unwrap(Result<T> r, const char* what) {
  if (!r.ok()) {
    std::cerr << "ABORT " << what << ": " << r.status().ToString() << std::endl;
    std::exit(2);
  }
  return r.MoveValueUnsafe();
}
#define OK(expr)                                                                \
  do {                                                                          \
    Status _s = (expr);                                                         \
    if (!_s.ok()) {                                                             \
      std::cerr << "ABORT " #expr ": " << _s.ToString() << std::endl;           \
      std::exit(2);                                                             \
    }                                                                           \
  } while (0)

static std::shared_ptr<cuda::CudaContext> g_cuda;

std::shared_ptr<Buffer> PinnedBuffer(int64_t nbytes) {
  auto r = cp::cuda::AllocatePinned(nbytes, 0);
  if (!r.ok()) {
    std::cerr << "ABORT AllocatePinned: " << r.status().ToString() << std::endl;
    std::exit(2);
  }
  return r.MoveValueUnsafe();
}

static inline uint64_t R(uint64_t seed, uint64_t i) {
  uint64_t x = i + seed * 0x9E3779B97F4A7C15ull + 0x9E3779B97F4A7C15ull;
  uint64_t z = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

std::shared_ptr<Buffer> ValidityOf(uint64_t seed, int64_t n, int64_t* nulls) {
  auto buf = PinnedBuffer(BitUtil::BytesForBits(n));
  std::memset(buf->mutable_data(), 0, buf->size());
  int64_t zeros = 0;
  for (int64_t i = 0; i < n; ++i) {
    const bool v = ((R(seed, i) >> 33) % 100) != 0;
    if (v) BitUtil::SetBit(buf->mutable_data(), i);
    else ++zeros;
  }
  *nulls = zeros;
  return buf;
}

std::shared_ptr<ArrayData> ToDevice(const ArrayData& a) {
  auto out = a.Copy();
  for (auto& buf : out->buffers) {
    if (!buf) continue;
    auto dbuf = GET(g_cuda->Allocate(std::max<int64_t>(buf->size(), 64)));
    if (buf->size() > 0) OK(dbuf->CopyFromHost(0, buf->data(), buf->size()));
    buf = dbuf;
  }
  for (auto& child : out->child_data) child = ToDevice(*child);
  return out;
}
int64_t BytesOf(const ArrayData& a) {
  int64_t b = 0;
  for (auto& buf : a.buffers)
    if (buf) b += buf->size();
  for (auto& c : a.child_data) b += BytesOf(*c);
  return b;
}
std::shared_ptr<ArrayData> ToHost(const ArrayData& a) {
  auto out = a.Copy();
  for (auto& buf : out->buffers) {
    if (!buf) continue;
    auto hbuf = GET(AllocateBuffer(buf->size()));
    if (buf->size() > 0) OK(GET(cuda::CudaBuffer::FromBuffer(buf))->CopyToHost(0, buf->size(), hbuf->mutable_data()));
    buf = std::move(hbuf);
  }
  for (auto& child : out->child_data) child = ToHost(*child);
  return out;
}

double Now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main(int argc, char** argv) {
  const std::string wl = argc > 1 ? argv[1] : "sort_i64";
  const int64_t n = argc > 2 ? std::atoll(argv[2]) : 100000000;
  const int steps = argc > 3 ? std::atoi(argv[3]) : 5;
  auto manager = GET(cuda::CudaDeviceManager::Instance());
  g_cuda = GET(manager->GetContext(0));
  auto registry = GET(cp::cuda::MakeCudaRegistry(0));
  cp::ExecContext ctx(default_memory_pool(), registry.get());

  std::vector<std::shared_ptr<ArrayData>> host;   // the workload's input columns
  if (wl == "sort_i64") {
    auto data = PinnedBuffer(n * 8);
    auto* v = reinterpret_cast<int64_t*>(data->mutable_data());
    for (int64_t i = 0; i < n; ++i) v[i] = int64_t(R(4, i));
    int64_t nulls = 0;
    auto valid = ValidityOf(6, n, &nulls);
    host.push_back(ArrayData::Make(int64(), n, {valid, std::move(data)}, nulls));
  } else if (wl == "filter") {
    auto data = PinnedBuffer(n * 8);
    auto* v = reinterpret_cast<int64_t*>(data->mutable_data());
    for (int64_t i = 0; i < n; ++i) v[i] = int64_t(R(1, i));
    int64_t nulls = 0;
    auto valid = ValidityOf(2, n, &nulls);
    host.push_back(ArrayData::Make(int64(), n, {valid, std::move(data)}, nulls));
    auto mask = PinnedBuffer(BitUtil::BytesForBits(n));
    std::memset(mask->mutable_data(), 0, mask->size());
    for (int64_t i = 0; i < n; ++i)
      if (R(3, i) & 1) BitUtil::SetBit(mask->mutable_data(), i);
    host.push_back(ArrayData::Make(boolean(), n, {nullptr, std::move(mask)}, 0));
  } else if (wl == "groupby_c3") {
    auto keys = PinnedBuffer(n * 4);
    auto* k = reinterpret_cast<int32_t*>(keys->mutable_data());
    for (int64_t i = 0; i < n; ++i) k[i] = int32_t((R(7, i) >> 20) % 1000);
    host.push_back(ArrayData::Make(int32(), n, {nullptr, std::move(keys)}, 0));
    auto data = PinnedBuffer(n * 8);
    auto* v = reinterpret_cast<double*>(data->mutable_data());
    for (int64_t i = 0; i < n; ++i) v[i] = double(R(8, i) >> 11) * std::ldexp(1.0, -53) * 1e14;
    int64_t nulls = 0;
    auto valid = ValidityOf(11, n, &nulls);
    host.push_back(ArrayData::Make(float64(), n, {valid, std::move(data)}, nulls));
  } else {
    std::cerr << "unknown workload " << wl << std::endl;
    return 2;
  }

  auto run = [&](const std::vector<std::shared_ptr<ArrayData>>& cols) -> Datum {
    if (wl == "sort_i64") {
      cp::ArraySortOptions opts(cp::SortOrder::Ascending);
      return GET(cp::CallFunction("array_sort_indices", {Datum(cols[0])}, &opts, &ctx));
    }
    if (wl == "filter") {
      cp::FilterOptions opts(cp::FilterOptions::DROP);
      return GET(cp::CallFunction("filter", {Datum(cols[0]), Datum(cols[1])}, &opts, &ctx));
    }
    std::vector<cp::internal::Aggregate> aggs = {{"hash_sum", nullptr}, {"hash_count", nullptr}, {"hash_min_max", nullptr}};
    return GET(cp::cuda::GroupBy({Datum(cols[1]), Datum(cols[1]), Datum(cols[1])}, {Datum(cols[0])}, aggs, &ctx));
  };

  // ---- device-resident
  std::vector<std::shared_ptr<ArrayData>> dev;
  for (auto& h : host) dev.push_back(ToDevice(*h));
  for (int i = 0; i < 3; ++i) run(dev);
  OK(g_cuda->Synchronize());
  double t0 = Now();
  for (int i = 0; i < steps; ++i) run(dev);
  OK(g_cuda->Synchronize());
  const double dev_ms = (Now() - t0) * 1e3 / steps;
  dev.clear();

  // ---- end to end: host arrays in, host array out
  int64_t h2d = 0, d2h = 0;
  auto e2e = [&](bool glue_copies) {
    std::vector<std::shared_ptr<ArrayData>> d;
    h2d = 0;
    for (auto& h : host) {
      d.push_back(glue_copies ? GET(cp::cuda::CopyToDevice(*h, 0)) : ToDevice(*h));
      h2d += BytesOf(*h);
    }
    Datum out = run(d);
    auto back = glue_copies ? GET(cp::cuda::CopyToHost(*out.array(), 0)) : ToHost(*out.array());
    d2h = BytesOf(*back);
  };
  e2e(true);
  e2e(true);
  t0 = Now();
  for (int i = 0; i < steps; ++i) e2e(true);
  const double e2e_ms = (Now() - t0) * 1e3 / steps;
  e2e(false);
  t0 = Now();
  for (int i = 0; i < 2; ++i) e2e(false);
  const double e2e_ref_ms = (Now() - t0) * 1e3 / 2;

  std::cout << "{\"workload\": \"" << wl << "\", \"rows\": " << n << ", \"steps\": " << steps
            << ", \"api\": \"compute::CallFunction / cuda::GroupBy over the CUDA FunctionRegistry (arrow_glue)\""
            << ", \"device_ms_per_step\": " << dev_ms << ", \"device_rows_per_s\": " << double(n) / (dev_ms * 1e-3)
            << ", \"e2e_ms_per_step\": " << e2e_ms << ", \"e2e_rows_per_s\": " << double(n) / (e2e_ms * 1e-3)
            << ", \"h2d_bytes_per_step\": " << h2d << ", \"d2h_bytes_per_step\": " << d2h
            << ", \"e2e_with_reference_copies_ms_per_step\": " << e2e_ref_ms
            << ", \"copies\": \"cuda::CopyToDevice / CopyToHost (DMA from / to pinned host buffers, pooled device blocks); "
               "reference copies = CudaBuffer::CopyFromHost / CopyToHost\"}" << std::endl;
  return 0;
}
