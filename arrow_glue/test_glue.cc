// test_glue.cc — GPU parity driver for the Arrow glue: every check runs the SAME call twice through
// the reference's public API — once with the reference's default (CPU) registry on host arrays,
// once with the CUDA registry on CudaBuffer-resident copies — and compares the results on the host
// with the reference's own Array::Equals.  Cases follow the reference's tests
// (vector_selection_test.cc:296-321, vector_sort_test.cc:423-464, hash_aggregate_test.cc:480-605).
// Test infrastructure: run by tests/test_glue_gpu.py on the GPU box; exits non-zero on any mismatch.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <random>
#include <string>
#include <vector>

#include "arrow/api.h"
#include "arrow/compute/api.h"
#include "arrow/compute/exec.h"
#include "arrow/compute/kernel.h"
#include "arrow/gpu/cuda_api.h"
#include "arrow/util/checked_cast.h"

#include "arrow/c/bridge.h"
#include "b200_arrow.h"

using namespace arrow;  // NOLINT
namespace cp = arrow::compute;

static int g_failures = 0, g_checks = 0;
#define CHECK_TRUE(cond, what)                                         \
  do {                                                                 \
    ++g_checks;                                                        \
    if (!(cond)) {                                                     \
      ++g_failures;                                                    \
      std::cout << "FAIL " << what << " (" #cond ")" << std::endl;     \
    } else {                                                           \
      std::cout << "ok   " << what << std::endl;                       \
    }                                                                  \
  } while (0)
#define ABORT_NOT_OK(expr)                                                         \
  do {                                                                             \
    ::arrow::Status _st = (expr);                                                  \
    if (!_st.ok()) {                                                               \
      std::cout << "ABORT " #expr ": " << _st.ToString() << std::endl;             \
      std::exit(2);                                                                \
    }                                                                              \
  } while (0)
template <typename T>
T Unwrap(Result<T> r, const char* what) {
  if (!r.ok()) {
    std::cout << "ABORT " << what << ": " << r.status().ToString() << std::endl;
    std::exit(2);
  }
  return r.MoveValueUnsafe();
}
#define GET(expr) Unwrap((expr), #expr)

static std::shared_ptr<cuda::CudaContext> g_cuda;

// ---- host <-> device copies of ArrayData (buffers copied whole; offsets / null counts kept)
std::shared_ptr<ArrayData> ToDevice(const ArrayData& a) {
  auto out = a.Copy();
  for (auto& buf : out->buffers) {
    if (!buf) continue;
    auto dbuf = GET(g_cuda->Allocate(std::max<int64_t>(buf->size(), 64)));
    if (buf->size() > 0) ABORT_NOT_OK(dbuf->CopyFromHost(0, buf->data(), buf->size()));
    buf = dbuf;
  }
  for (auto& child : out->child_data) child = ToDevice(*child);
  if (out->dictionary) out->dictionary = ToDevice(*out->dictionary);
  return out;
}
std::shared_ptr<ArrayData> ToHost(const ArrayData& a) {
  auto out = a.Copy();
  for (auto& buf : out->buffers) {
    if (!buf) continue;
    auto hbuf = GET(AllocateBuffer(buf->size()));
    if (buf->size() > 0) {
      auto cbuf = GET(cuda::CudaBuffer::FromBuffer(buf));
      ABORT_NOT_OK(cbuf->CopyToHost(0, buf->size(), hbuf->mutable_data()));
    }
    buf = std::move(hbuf);
  }
  for (auto& child : out->child_data) child = ToHost(*child);
  if (out->dictionary) out->dictionary = ToHost(*out->dictionary);
  return out;
}
Datum DeviceDatum(const std::shared_ptr<Array>& a) { return Datum(ToDevice(*a->data())); }
std::shared_ptr<Array> HostArray(const Datum& d) { return MakeArray(ToHost(*d.array())); }

// host ChunkedArray -> device-resident twin.  Null counts are settled on the host first: ChunkedArray's
// constructor sums chunk->null_count() (chunked_array.cc:46-63), which must not scan a device bitmap.
std::shared_ptr<ChunkedArray> DeviceChunked(const ChunkedArray& h) {
  ArrayVector d;
  for (const auto& c : h.chunks()) { (void)c->null_count(); d.push_back(MakeArray(ToDevice(*c->data()))); }
  return std::make_shared<ChunkedArray>(d, h.type());
}
// slices of one host array as a ChunkedArray (cuts: boundaries, first 0, last length)
std::shared_ptr<ChunkedArray> Chunked(const std::shared_ptr<Array>& a, const std::vector<int64_t>& cuts) {
  ArrayVector out;
  for (size_t i = 0; i + 1 < cuts.size(); ++i) out.push_back(a->Slice(cuts[i], cuts[i + 1] - cuts[i]));
  return std::make_shared<ChunkedArray>(out, a->type());
}
// all chunks as one host array
std::shared_ptr<Array> Flatten(const ChunkedArray& c, bool on_device) {
  ArrayVector host;
  for (const auto& k : c.chunks()) host.push_back(on_device ? MakeArray(ToHost(*k->data())) : k);
  if (host.empty()) return GET(MakeArrayOfNull(c.type(), 0));
  return GET(Concatenate(host));
}

// ---- builders
template <typename BuilderT, typename V>
std::shared_ptr<Array> Make(const std::vector<V>& values, const std::vector<bool>& valid = {}) {
  BuilderT b;
  for (size_t i = 0; i < values.size(); ++i) {
    if (!valid.empty() && !valid[i]) ABORT_NOT_OK(b.AppendNull());
    else ABORT_NOT_OK(b.Append(static_cast<typename BuilderT::value_type>(values[i])));
  }
  return GET(b.Finish());
}
template <typename BuilderT, typename V>
std::shared_ptr<Array> MakeRandom(int64_t n, std::mt19937_64& rng, double null_p, V lo, V hi, bool floating) {
  BuilderT b;
  std::uniform_real_distribution<double> u(0, 1);
  for (int64_t i = 0; i < n; ++i) {
    if (u(rng) < null_p) { ABORT_NOT_OK(b.AppendNull()); continue; }
    const double x = u(rng);
    V v = floating ? static_cast<V>(lo + (hi - lo) * x) : static_cast<V>(lo + static_cast<V>(x * (double(hi) - double(lo))));
    ABORT_NOT_OK(b.Append(static_cast<typename BuilderT::value_type>(v)));
  }
  return GET(b.Finish());
}

void ExpectEqual(const std::shared_ptr<Array>& got, const std::shared_ptr<Array>& want, const std::string& what) {
  const bool same = got->type()->Equals(want->type()) && got->Equals(want);
  CHECK_TRUE(same, what);
  if (!same && want->length() < 40) std::cout << "  want " << want->ToString() << "\n  got  " << got->ToString() << std::endl;
  auto st = got->ValidateFull();
  CHECK_TRUE(st.ok(), what + " ValidateFull");
}

// sums / means: |gpu - ref| <= 1e-12 * |ref| (north_star's float64 tolerance); everything else exact
void ExpectStructClose(const std::shared_ptr<Array>& got, const std::shared_ptr<Array>& want, const std::string& what) {
  bool ok = got->type()->Equals(want->type()) && got->length() == want->length();
  if (ok) {
    const auto& g = internal::checked_cast<const StructArray&>(*got);
    const auto& w = internal::checked_cast<const StructArray&>(*want);
    for (int f = 0; f < g.num_fields() && ok; ++f) {
      auto gf = g.field(f), wf = w.field(f);
      if (gf->type_id() == Type::DOUBLE) {
        const auto& gd = internal::checked_cast<const DoubleArray&>(*gf);
        const auto& wd = internal::checked_cast<const DoubleArray&>(*wf);
        for (int64_t i = 0; i < gd.length() && ok; ++i) {
          if (gd.IsNull(i) != wd.IsNull(i)) ok = false;
          else if (gd.IsValid(i)) {
            const double a = gd.Value(i), b = wd.Value(i);
            if (!(a == b || std::fabs(a - b) <= 1e-12 * std::fabs(b) || (std::isnan(a) && std::isnan(b)))) ok = false;
          }
        }
      } else {
        ok = gf->Equals(wf);
      }
      if (!ok) std::cout << "  field " << f << " (" << got->type()->field(f)->name() << ") differs" << std::endl;
    }
  }
  CHECK_TRUE(ok, what);
  if (!ok && want->length() < 40) std::cout << "  want " << want->ToString() << "\n  got  " << got->ToString() << std::endl;
}

int main() {
  auto manager = GET(cuda::CudaDeviceManager::Instance());
  g_cuda = GET(manager->GetContext(0));
  auto registry = GET(cp::cuda::MakeCudaRegistry(0));
  cp::ExecContext gpu_ctx(default_memory_pool(), registry.get());
  cp::ExecContext cpu_ctx;  // the reference's own registry and kernels
  std::mt19937_64 rng(42);

  {  // the registry holds exactly the reference's names for the path
    for (const char* name : {"filter", "array_filter", "sort_indices", "array_sort_indices", "take", "array_take", "hash_count", "hash_sum", "hash_min_max", "hash_mean"})
      CHECK_TRUE(registry->GetFunction(name).ok(), std::string("registered: ") + name);
    CHECK_TRUE(registry->GetFunction("cumulative_sum").status().IsKeyError(), "unknown function -> KeyError (registry.cc:62-64)");
  }

  // ---------------- ingest / egress: cuda::CopyToDevice / CopyToHost, C Data Interface ----------------
  {
    const int64_t n = 3000001;   // several staging blocks, ragged tail
    auto vals = MakeRandom<Int64Builder, int64_t>(n, rng, 0.05, -(1LL << 60), 1LL << 60, false);
    auto sliced = vals->Slice(12345, n - 20000);
    auto dev = GET(cp::cuda::CopyToDevice(*sliced->data(), 0));   // pageable source: staged through pinned blocks
    CHECK_TRUE(!dev->buffers[1]->is_cpu() && dev->offset == sliced->data()->offset && dev->null_count.load() == sliced->null_count(),
               "CopyToDevice keeps offset and settles the null count");
    auto back = MakeArray(GET(cp::cuda::CopyToHost(*dev, 0)));
    ExpectEqual(back, sliced, "CopyToDevice -> CopyToHost round trip (pageable source, sliced)");
    // pinned source: straight DMA
    auto pinned = GET(cp::cuda::AllocatePinned(n * 8, 0));
    std::memcpy(pinned->mutable_data(), vals->data()->buffers[1]->data(), n * 8);
    auto parr = ArrayData::Make(int64(), n, {vals->data()->buffers[0], pinned}, vals->null_count());
    auto pdev = GET(cp::cuda::CopyToDevice(*parr, 0));
    ExpectEqual(MakeArray(GET(cp::cuda::CopyToHost(*pdev, 0))), vals, "CopyToDevice from a pinned buffer");
    // C Data Interface: export on the host, import onto the device, compute, export back
    struct ArrowArray c_array;
    struct ArrowSchema c_schema;
    ABORT_NOT_OK(ExportArray(*vals, &c_array, &c_schema));
    auto imported = GET(cp::cuda::ImportToDevice(&c_array, &c_schema, 0));
    auto mask = MakeRandom<BooleanBuilder, int>(n, rng, 0.0, 0, 2, false);
    auto want = GET(cp::Filter(vals, mask, cp::FilterOptions::Defaults(), &cpu_ctx)).make_array();
    auto got = GET(cp::Filter(Datum(imported), DeviceDatum(mask), cp::FilterOptions::Defaults(), &gpu_ctx));
    struct ArrowArray c_out;
    struct ArrowSchema c_out_schema;
    ABORT_NOT_OK(cp::cuda::ExportFromDevice(*got.array(), &c_out, &c_out_schema, 0));
    auto round = GET(ImportArray(&c_out, &c_out_schema));
    ExpectEqual(round, want, "C Data Interface: ImportToDevice -> filter -> ExportFromDevice");
  }

  // ---------------- Filter ----------------
  {
    auto values = Make<Int32Builder, int32_t>({7, 8, 9});
    auto mask = Make<BooleanBuilder, bool>({false, true, false}, {false, true, true});
    for (auto ns : {cp::FilterOptions::DROP, cp::FilterOptions::EMIT_NULL}) {
      cp::FilterOptions opts(ns);
      auto want = GET(cp::Filter(values, mask, opts, &cpu_ctx)).make_array();
      auto got = HostArray(GET(cp::Filter(DeviceDatum(values), DeviceDatum(mask), opts, &gpu_ctx)));
      ExpectEqual(got, want, std::string("filter [7,8,9]x[null,1,0] ") + (ns ? "EMIT_NULL" : "DROP"));
    }
    // sliced mask / values (vector_selection_test.cc:312-315)
    auto v2 = Make<Int64Builder, int64_t>({1, 2, 3, 4, 5, 6, 7, 8, 9, 10}, {true, true, false, true, true, true, true, false, true, true});
    auto m2 = Make<BooleanBuilder, bool>({true, false, true, true, false, true, false, true, true, true, false, true});
    auto vs = v2->Slice(3, 6), ms = m2->Slice(5, 6);
    auto want = GET(cp::Filter(vs, ms, cp::FilterOptions::Defaults(), &cpu_ctx)).make_array();
    auto got = HostArray(GET(cp::Filter(DeviceDatum(vs), DeviceDatum(ms), cp::FilterOptions::Defaults(), &gpu_ctx)));
    ExpectEqual(got, want, "filter sliced values + sliced mask");
    // length mismatch: the reference's test asserts Invalid (vector_selection_test.cc:317-320)
    auto bad = cp::Filter(DeviceDatum(values), DeviceDatum(m2), cp::FilterOptions::Defaults(), &gpu_ctx);
    CHECK_TRUE(bad.status().IsInvalid(), "filter length mismatch -> Invalid");
    auto notbool = cp::Filter(DeviceDatum(values), DeviceDatum(values), cp::FilterOptions::Defaults(), &gpu_ctx);
    CHECK_TRUE(notbool.status().IsNotImplemented(), "filter non-boolean mask -> NotImplemented (vector_selection.cc:1885-1887)");
    auto hostin = cp::Filter(Datum(values), Datum(mask), cp::FilterOptions::Defaults(), &gpu_ctx);
    CHECK_TRUE(hostin.status().IsInvalid(), "filter on host-resident input -> Invalid (no CPU fallback)");
  }
  {
    const int64_t n = 200003;
    std::vector<std::shared_ptr<Array>> columns = {
        MakeRandom<Int64Builder, int64_t>(n, rng, 0.01, -(1LL << 60), 1LL << 60, false),
        MakeRandom<Int8Builder, int8_t>(n, rng, 0.05, -100, 100, false),
        MakeRandom<UInt16Builder, uint16_t>(n, rng, 0.0, 0, 60000, false),
        MakeRandom<FloatBuilder, float>(n, rng, 0.1, -1e3f, 1e3f, true),
        MakeRandom<DoubleBuilder, double>(n, rng, 0.5, -1e9, 1e9, true),
        MakeRandom<BooleanBuilder, int>(n, rng, 0.02, 0, 2, false),
        MakeRandom<Date32Builder, int32_t>(n, rng, 0.02, 8000, 11000, false)};
    auto mask_nonull = MakeRandom<BooleanBuilder, int>(n, rng, 0.0, 0, 2, false);
    auto mask_nulls = MakeRandom<BooleanBuilder, int>(n, rng, 0.05, 0, 2, false);
    for (auto& col : columns) {
      for (auto& mask : {mask_nonull, mask_nulls}) {
        for (auto ns : {cp::FilterOptions::DROP, cp::FilterOptions::EMIT_NULL}) {
          cp::FilterOptions opts(ns);
          auto want = GET(cp::CallFunction("filter", {col, mask}, &opts, &cpu_ctx)).make_array();
          auto got = HostArray(GET(cp::CallFunction("filter", {DeviceDatum(col), DeviceDatum(mask)}, &opts, &gpu_ctx)));
          ExpectEqual(got, want, "filter random " + col->type()->ToString() + (mask.get() == mask_nulls.get() ? " mask+nulls " : " mask ") + (ns ? "EMIT_NULL" : "DROP"));
        }
      }
    }
    // RecordBatch form == per-column array_filter (vector_selection_test.cc:208-236)
    auto schema_ = schema({field("a", columns[0]->type()), field("b", columns[4]->type())});
    auto batch = RecordBatch::Make(schema_, n, {columns[0], columns[4]});
    auto dbatch = RecordBatch::Make(schema_, n, {ToDevice(*columns[0]->data()), ToDevice(*columns[4]->data())});
    cp::FilterOptions opts(cp::FilterOptions::EMIT_NULL);
    auto want = GET(cp::CallFunction("filter", {Datum(batch), Datum(mask_nulls)}, &opts, &cpu_ctx)).record_batch();
    auto got = GET(cp::CallFunction("filter", {Datum(dbatch), DeviceDatum(mask_nulls)}, &opts, &gpu_ctx)).record_batch();
    for (int c = 0; c < 2; ++c) ExpectEqual(MakeArray(ToHost(*got->column_data(c))), want->column(c), "filter RecordBatch column " + std::to_string(c));
  }

  // ---------------- SortIndices ----------------
  {
    auto values = Make<Int64Builder, int64_t>({10, 12, 4, 50, 50, 32, 11});
    for (auto order : {cp::SortOrder::Ascending, cp::SortOrder::Descending}) {
      auto want = GET(cp::SortIndices(*values, order, &cpu_ctx));
      auto dvalues = MakeArray(ToDevice(*values->data()));
      auto got = GET(cp::SortIndices(*dvalues, order, &gpu_ctx));
      ExpectEqual(MakeArray(ToHost(*got->data())), want, std::string("sort_indices ties ") + (order == cp::SortOrder::Ascending ? "asc" : "desc"));
    }
    const int64_t n = 300007;
    DoubleBuilder fb;
    std::uniform_real_distribution<double> u(0, 1);
    for (int64_t i = 0; i < n; ++i) {
      const double x = u(rng);
      if (x < 0.02) ABORT_NOT_OK(fb.AppendNull());
      else if (x < 0.04) ABORT_NOT_OK(fb.Append(std::nan("")));
      else if (x < 0.05) ABORT_NOT_OK(fb.Append(-0.0));
      else if (x < 0.06) ABORT_NOT_OK(fb.Append(0.0));
      else if (x < 0.07) ABORT_NOT_OK(fb.Append(x < 0.065 ? INFINITY : -INFINITY));
      else ABORT_NOT_OK(fb.Append(std::floor((u(rng) - 0.5) * 2000.0)));  // heavy duplicates
    }
    std::vector<std::shared_ptr<Array>> columns = {
        GET(fb.Finish()),
        MakeRandom<Int64Builder, int64_t>(n, rng, 0.01, -(1LL << 62), 1LL << 62, false),
        MakeRandom<Int32Builder, int32_t>(n, rng, 0.01, -1000, 1000, false),
        MakeRandom<UInt8Builder, uint8_t>(n, rng, 0.1, 0, 255, false),
        MakeRandom<FloatBuilder, float>(n, rng, 0.0, -5.f, 5.f, true),
        MakeRandom<UInt64Builder, uint64_t>(n, rng, 0.0, 0, ~0ull >> 1, false)};
    for (auto& col : columns) {
      for (auto order : {cp::SortOrder::Ascending, cp::SortOrder::Descending}) {
        cp::ArraySortOptions opts(order);
        auto want = GET(cp::CallFunction("array_sort_indices", {col}, &opts, &cpu_ctx)).make_array();
        auto got = HostArray(GET(cp::CallFunction("array_sort_indices", {DeviceDatum(col)}, &opts, &gpu_ctx)));
        ExpectEqual(got, want, "array_sort_indices random " + col->type()->ToString() + (order == cp::SortOrder::Ascending ? " asc" : " desc"));
      }
    }
    // sliced input, through the meta function with SortOptions
    auto sl = columns[0]->Slice(1001, 50000);
    cp::SortOptions sopts({cp::SortKey("", cp::SortOrder::Descending)});
    auto want = GET(cp::CallFunction("sort_indices", {sl}, &sopts, &cpu_ctx)).make_array();
    auto got = HostArray(GET(cp::CallFunction("sort_indices", {DeviceDatum(sl)}, &sopts, &gpu_ctx)));
    ExpectEqual(got, want, "sort_indices sliced float64 desc (meta function)");
    // ChunkedArray: indices over the logical concatenation, NaNs / nulls of every chunk last (vector_sort_test.cc:629-647)
    for (auto& col : {columns[0], columns[1]}) {
      ArrayVector hchunks = {col->Slice(0, 70001), col->Slice(70001, 0), col->Slice(70001, 130000), col->Slice(200001)};
      ArrayVector dchunks;
      // ChunkedArray's constructor sums chunk->null_count() (chunked_array.cc:46-63); for a slice that is
      // kUnknownNullCount, i.e. a host scan of the bitmap, so the count has to be settled before the buffers
      // are device pointers
      for (auto& c : hchunks) { (void)c->null_count(); dchunks.push_back(MakeArray(ToDevice(*c->data()))); }
      auto hca = std::make_shared<ChunkedArray>(hchunks, col->type());
      auto dca = std::make_shared<ChunkedArray>(dchunks, col->type());
      for (auto order : {cp::SortOrder::Ascending, cp::SortOrder::Descending}) {
        cp::SortOptions o({cp::SortKey("", order)});
        auto want_c = GET(cp::CallFunction("sort_indices", {Datum(hca)}, &o, &cpu_ctx)).make_array();
        auto got_c = HostArray(GET(cp::CallFunction("sort_indices", {Datum(dca)}, &o, &gpu_ctx)));
        ExpectEqual(got_c, want_c, "sort_indices ChunkedArray (4 chunks) " + col->type()->ToString() + (order == cp::SortOrder::Ascending ? " asc" : " desc"));
      }
    }
  }

  // ---------------- Take ----------------
  {
    const int64_t n = 60000;
    auto values = MakeRandom<DoubleBuilder, double>(n, rng, 0.1, -1e6, 1e6, true);
    auto small = MakeRandom<Int16Builder, int16_t>(200, rng, 0.2, -300, 300, false);
    auto ix32 = MakeRandom<Int32Builder, int32_t>(100000, rng, 0.05, 0, int32_t(n - 1), false);
    auto ix8 = MakeRandom<UInt8Builder, uint8_t>(5000, rng, 0.0, 0, 199, false);
    struct Case { std::shared_ptr<Array> v, i; const char* what; };
    for (const Case& c : {Case{values, ix32, "take double by nullable int32"}, Case{small, ix8, "take int16 by uint8"},
                          Case{values->Slice(100, 50000), ix32->Slice(7, 9000)->View(int32()).ValueOrDie(), "take sliced values"}}) {
      // the sliced case needs in-range indices: reduce them on the host first
      std::shared_ptr<Array> idx = c.i;
      if (std::string(c.what) == "take sliced values") {
        Int32Builder b;
        const auto& src = internal::checked_cast<const Int32Array&>(*c.i);
        for (int64_t k = 0; k < src.length(); ++k) {
          if (src.IsNull(k)) ABORT_NOT_OK(b.AppendNull());
          else ABORT_NOT_OK(b.Append(src.Value(k) % 50000));
        }
        idx = GET(b.Finish());
      }
      auto want = GET(cp::CallFunction("take", {c.v, idx}, &cpu_ctx)).make_array();
      auto got = HostArray(GET(cp::CallFunction("take", {DeviceDatum(c.v), DeviceDatum(idx)}, &gpu_ctx)));
      ExpectEqual(got, want, c.what);
      CHECK_TRUE(got->null_count() == want->null_count(), std::string(c.what) + " null_count");
    }
    // SortIndices -> Take through both registries: the sorted column itself
    cp::ArraySortOptions asc(cp::SortOrder::Ascending);
    auto widx = GET(cp::CallFunction("array_sort_indices", {values}, &asc, &cpu_ctx));
    auto wsorted = GET(cp::CallFunction("take", {Datum(values), widx}, &cpu_ctx)).make_array();
    auto dvalues = DeviceDatum(values);
    auto gidx = GET(cp::CallFunction("array_sort_indices", {dvalues}, &asc, &gpu_ctx));
    auto gsorted = HostArray(GET(cp::CallFunction("take", {dvalues, gidx}, &gpu_ctx)));
    ExpectEqual(gsorted, wsorted, "take(values, sort_indices(values)) on the device");
    // out of bounds -> IndexError from both (vector_selection_test.cc: TakeTest index errors)
    auto bad = Make<Int32Builder, int32_t>({0, 200, 1});
    auto st_cpu = cp::CallFunction("take", {small, bad}, &cpu_ctx).status();
    auto st_gpu = cp::CallFunction("take", {DeviceDatum(small), DeviceDatum(bad)}, &gpu_ctx).status();
    CHECK_TRUE(st_cpu.IsIndexError() && st_gpu.IsIndexError(), "take out of bounds -> IndexError");
    CHECK_TRUE(st_gpu.message().find("Index 200 out of bounds") != std::string::npos, "take IndexError message names the index");
  }

  // ---------------- chunked / table forms of Filter and Take (vector_selection.cc:1776-1866, 1914-2017) ------
  {
    std::mt19937_64 rng(99);
    const int64_t n = 50000;
    auto a = MakeRandom<Int64Builder, int64_t>(n, rng, 0.1, -1000, 1000, false);
    auto b = MakeRandom<DoubleBuilder, double>(n, rng, 0.0, -1.0, 1.0, true);
    auto c = MakeRandom<BooleanBuilder, int>(n, rng, 0.2, 0, 2, false);
    auto mask = MakeRandom<BooleanBuilder, int>(n, rng, 0.05, 0, 2, false);
    const std::vector<int64_t> cuts_a = {0, 0, 1, 4097, 20000, 20000, n}, cuts_b = {0, 33, 30001, n}, cuts_m = {0, 777, 40000, 40000, n};
    auto ca = Chunked(a, cuts_a), cb = Chunked(b, cuts_b), cc = Chunked(c, cuts_a), cm = Chunked(mask, cuts_m);
    auto da = DeviceChunked(*ca), db = DeviceChunked(*cb), dc = DeviceChunked(*cc), dm = DeviceChunked(*cm);
    for (auto ns : {cp::FilterOptions::DROP, cp::FilterOptions::EMIT_NULL}) {
      cp::FilterOptions opts(ns);
      const std::string tag = ns ? " EMIT_NULL" : " DROP";
      // ChunkedArray x ChunkedArray with different chunk layouts, ChunkedArray x Array, Array x ChunkedArray
      auto want = GET(cp::CallFunction("filter", {Datum(ca), Datum(cm)}, &opts, &cpu_ctx)).chunked_array();
      auto got = GET(cp::CallFunction("filter", {Datum(da), Datum(dm)}, &opts, &gpu_ctx)).chunked_array();
      ExpectEqual(Flatten(*got, true), Flatten(*want, false), "filter(ChunkedArray, ChunkedArray), layouts differ" + tag);
      CHECK_TRUE(got->null_count() == want->null_count(), "filter(ChunkedArray, ChunkedArray) null_count" + tag);
      got = GET(cp::CallFunction("filter", {Datum(dc), DeviceDatum(mask)}, &opts, &gpu_ctx)).chunked_array();
      want = GET(cp::CallFunction("filter", {Datum(cc), Datum(mask)}, &opts, &cpu_ctx)).chunked_array();
      ExpectEqual(Flatten(*got, true), Flatten(*want, false), "filter(ChunkedArray bool, Array)" + tag);
      got = GET(cp::CallFunction("filter", {DeviceDatum(b), Datum(dm)}, &opts, &gpu_ctx)).chunked_array();
      want = GET(cp::CallFunction("filter", {Datum(b), Datum(cm)}, &opts, &cpu_ctx)).chunked_array();
      ExpectEqual(Flatten(*got, true), Flatten(*want, false), "filter(Array, ChunkedArray)" + tag);
      // Table x ChunkedArray and Table x Array: one plan per mask chunk shared by the three columns
      auto sch = schema({field("a", a->type()), field("b", b->type()), field("c", c->type())});
      auto htable = Table::Make(sch, {ca, cb, cc});
      auto dtable = Table::Make(sch, {da, db, dc});
      for (int form = 0; form < 2; ++form) {
        auto wt = GET(cp::CallFunction("filter", {Datum(htable), form ? Datum(mask) : Datum(cm)}, &opts, &cpu_ctx)).table();
        auto gt = GET(cp::CallFunction("filter", {Datum(dtable), form ? DeviceDatum(mask) : Datum(dm)}, &opts, &gpu_ctx)).table();
        CHECK_TRUE(gt->num_rows() == wt->num_rows() && gt->schema()->Equals(*wt->schema()), "filter(Table) rows and schema" + tag);
        for (int k = 0; k < 3; ++k)
          ExpectEqual(Flatten(*gt->column(k), true), Flatten(*wt->column(k), false),
                      std::string("filter(Table, ") + (form ? "Array" : "ChunkedArray") + ") column " + std::to_string(k) + tag);
      }
    }
    auto short_mask = MakeRandom<BooleanBuilder, int>(n - 1, rng, 0.0, 0, 2, false);
    auto sch = schema({field("a", a->type())});
    auto st = cp::CallFunction("filter", {Datum(Table::Make(sch, {da})), DeviceDatum(short_mask)}, &gpu_ctx).status();
    CHECK_TRUE(st.IsInvalid() && st.message().find("Filter inputs must all be the same length") != std::string::npos,
               "filter(Table) length mismatch -> Invalid");

    // Take: every route of TakeMetaFunction
    auto idx = MakeRandom<Int32Builder, int32_t>(70000, rng, 0.1, 0, int32_t(n), false);
    auto cidx = Chunked(idx, {0, 1, 1, 30000, 70000});
    auto didx = DeviceChunked(*cidx);
    {
      auto want = GET(cp::CallFunction("take", {Datum(ca), Datum(idx)}, &cpu_ctx)).chunked_array();
      auto got = GET(cp::CallFunction("take", {Datum(da), DeviceDatum(idx)}, &gpu_ctx)).chunked_array();
      CHECK_TRUE(got->num_chunks() == want->num_chunks(), "take(ChunkedArray, Array) -> one chunk");
      ExpectEqual(Flatten(*got, true), Flatten(*want, false), "take(ChunkedArray, Array): gathered from the chunks where they lie");
      CHECK_TRUE(got->null_count() == want->null_count(), "take(ChunkedArray, Array) null_count");
      want = GET(cp::CallFunction("take", {Datum(cc), Datum(cidx)}, &cpu_ctx)).chunked_array();
      got = GET(cp::CallFunction("take", {Datum(dc), Datum(didx)}, &gpu_ctx)).chunked_array();
      CHECK_TRUE(got->num_chunks() == want->num_chunks(), "take(ChunkedArray, ChunkedArray) -> one chunk per index chunk");
      ExpectEqual(Flatten(*got, true), Flatten(*want, false), "take(ChunkedArray bool, ChunkedArray)");
      want = GET(cp::CallFunction("take", {Datum(b), Datum(cidx)}, &cpu_ctx)).chunked_array();
      got = GET(cp::CallFunction("take", {DeviceDatum(b), Datum(didx)}, &gpu_ctx)).chunked_array();
      CHECK_TRUE(got->num_chunks() == want->num_chunks(), "take(Array, ChunkedArray) chunk count");
      ExpectEqual(Flatten(*got, true), Flatten(*want, false), "take(Array, ChunkedArray)");
    }
    {
      auto sch2 = schema({field("a", a->type()), field("c", c->type())});
      auto hb = RecordBatch::Make(sch2, n, {a, c});
      auto dbt = RecordBatch::Make(sch2, n, {ToDevice(*a->data()), ToDevice(*c->data())});
      auto want = GET(cp::CallFunction("take", {Datum(hb), Datum(idx)}, &cpu_ctx)).record_batch();
      auto got = GET(cp::CallFunction("take", {Datum(dbt), DeviceDatum(idx)}, &gpu_ctx)).record_batch();
      CHECK_TRUE(got->num_rows() == want->num_rows(), "take(RecordBatch, Array) rows");
      for (int k = 0; k < 2; ++k) ExpectEqual(MakeArray(ToHost(*got->column_data(k))), want->column(k), "take(RecordBatch, Array) column " + std::to_string(k));
      auto sch3 = schema({field("a", a->type()), field("b", b->type()), field("c", c->type())});
      auto htable = Table::Make(sch3, {ca, cb, cc});
      auto dtable = Table::Make(sch3, {da, db, dc});
      for (int form = 0; form < 2; ++form) {
        auto wt = GET(cp::CallFunction("take", {Datum(htable), form ? Datum(cidx) : Datum(idx)}, &cpu_ctx)).table();
        auto gt = GET(cp::CallFunction("take", {Datum(dtable), form ? Datum(didx) : DeviceDatum(idx)}, &gpu_ctx)).table();
        CHECK_TRUE(gt->num_rows() == wt->num_rows(), "take(Table) rows");
        for (int k = 0; k < 3; ++k)
          ExpectEqual(Flatten(*gt->column(k), true), Flatten(*wt->column(k), false),
                      std::string("take(Table, ") + (form ? "ChunkedArray" : "Array") + ") column " + std::to_string(k));
      }
      // the bound is the table's row count
      auto bad = Make<Int64Builder, int64_t>({0, n, 1});
      auto st_cpu = cp::CallFunction("take", {Datum(htable), Datum(bad)}, &cpu_ctx).status();
      auto st_gpu = cp::CallFunction("take", {Datum(dtable), DeviceDatum(bad)}, &gpu_ctx).status();
      CHECK_TRUE(st_cpu.IsIndexError() && st_gpu.IsIndexError(), "take(Table) out of bounds -> IndexError");
    }
  }

  // ---------------- SortIndices with several sort keys (vector_sort.cc:1632-1711) ----------------
  {
    // vector_sort_test.cc:783-842 golden cases: a ascending, b descending
    {
      auto a = Make<FloatBuilder, float>({0, 1, 3, 0, NAN, NAN, NAN, 1}, {false, true, true, false, true, true, true, true});
      auto b = Make<DoubleBuilder, double>({5, 3, 0, 0, 0, NAN, 5, 5}, {true, true, false, false, false, true, true, true});
      auto sch = schema({field("a", a->type()), field("b", b->type())});
      auto dbatch = RecordBatch::Make(sch, 8, {ToDevice(*a->data()), ToDevice(*b->data())});
      cp::SortOptions opts({cp::SortKey("a", cp::SortOrder::Ascending), cp::SortKey("b", cp::SortOrder::Descending)});
      auto got = HostArray(GET(cp::CallFunction("sort_indices", {Datum(dbatch)}, &opts, &gpu_ctx)));
      ExpectEqual(got, Make<UInt64Builder, uint64_t>({7, 1, 2, 6, 5, 4, 0, 3}), "sort_indices RecordBatch NaNAndNull golden");
    }
    std::mt19937_64 rng(4242);
    const int64_t n = 60000;
    auto k0 = MakeRandom<Int8Builder, int8_t>(n, rng, 0.1, 0, 4, false);
    auto k1 = MakeRandom<DoubleBuilder, double>(n, rng, 0.1, 0.0, 3.0, false);   // few distinct values: floors of [0, 3)
    auto k2 = MakeRandom<Int64Builder, int64_t>(n, rng, 0.0, -100000, 100000, false);
    auto sch = schema({field("x", k0->type()), field("y", k1->type()), field("z", k2->type())});
    auto hbatch = RecordBatch::Make(sch, n, {k0, k1, k2});
    auto dbatch = RecordBatch::Make(sch, n, {ToDevice(*k0->data()), ToDevice(*k1->data()), ToDevice(*k2->data())});
    for (int variant = 0; variant < 3; ++variant) {
      std::vector<cp::SortKey> keys;
      if (variant == 0) keys = {cp::SortKey("x", cp::SortOrder::Ascending), cp::SortKey("y", cp::SortOrder::Descending), cp::SortKey("z", cp::SortOrder::Ascending)};
      if (variant == 1) keys = {cp::SortKey("y", cp::SortOrder::Ascending), cp::SortKey("x", cp::SortOrder::Descending)};
      if (variant == 2) keys = {cp::SortKey("z", cp::SortOrder::Descending)};
      cp::SortOptions opts(keys);
      auto want = GET(cp::CallFunction("sort_indices", {Datum(hbatch)}, &opts, &cpu_ctx)).make_array();
      auto got = HostArray(GET(cp::CallFunction("sort_indices", {Datum(dbatch)}, &opts, &gpu_ctx)));
      ExpectEqual(got, want, "sort_indices RecordBatch, " + std::to_string(keys.size()) + " sort keys");
      auto htable = GET(Table::FromRecordBatches({hbatch}));
      auto dtable = GET(Table::FromRecordBatches({dbatch}));
      auto want_t = GET(cp::CallFunction("sort_indices", {Datum(htable)}, &opts, &cpu_ctx)).make_array();
      auto got_t = HostArray(GET(cp::CallFunction("sort_indices", {Datum(dtable)}, &opts, &gpu_ctx)));
      ExpectEqual(got_t, want_t, "sort_indices Table, " + std::to_string(keys.size()) + " sort keys");
      // the same table in chunks, each column cut at different places (vector_sort_test.cc:886-980 "several chunks")
      auto hx = Chunked(k0, {0, 7, 30000, n}), hy = Chunked(k1, {0, 20000, 20000, n}), hz = Chunked(k2, {0, 1, 59999, n});
      auto htable_c = Table::Make(sch, {hx, hy, hz});
      auto dtable_c = Table::Make(sch, {DeviceChunked(*hx), DeviceChunked(*hy), DeviceChunked(*hz)});
      auto want_c = GET(cp::CallFunction("sort_indices", {Datum(htable_c)}, &opts, &cpu_ctx)).make_array();
      auto got_c = HostArray(GET(cp::CallFunction("sort_indices", {Datum(dtable_c)}, &opts, &gpu_ctx)));
      ExpectEqual(got_c, want_c, "sort_indices multi-chunk Table, " + std::to_string(keys.size()) + " sort keys");
    }
    cp::SortOptions none;
    auto st0 = cp::CallFunction("sort_indices", {Datum(dbatch)}, &none, &gpu_ctx).status();
    CHECK_TRUE(st0.IsInvalid() && st0.message().find("Must specify one or more sort keys") != std::string::npos, "sort_indices RecordBatch without keys -> Invalid");
    cp::SortOptions missing({cp::SortKey("x", cp::SortOrder::Ascending), cp::SortKey("nope", cp::SortOrder::Ascending)});
    auto st1 = cp::CallFunction("sort_indices", {Datum(dbatch)}, &missing, &gpu_ctx).status();
    CHECK_TRUE(st1.IsInvalid() && st1.message().find("Nonexistent sort key column: nope") != std::string::npos, "sort_indices nonexistent key -> Invalid");
  }

  // ---------------- GroupBy ----------------
  {
    // hash_aggregate_test.cc:555-605 (CountAndSum shape): null key group sits at its first-occurrence position
    auto arg = Make<DoubleBuilder, double>({1.0, 0.0, 0.0, 3.25, 0.125, -0.25, 0.75, 0.0}, {true, false, true, true, true, true, true, false});
    auto key = Make<Int64Builder, int64_t>({1, 1, 2, 3, 0, 2, 0, 3}, {true, true, true, true, false, true, false, true});
    cp::CountOptions count_nulls(cp::CountOptions::COUNT_NULL);
    cp::MinMaxOptions emit_null(cp::MinMaxOptions::EMIT_NULL);
    std::vector<cp::internal::Aggregate> aggs = {{"hash_count", nullptr}, {"hash_sum", nullptr}, {"hash_min_max", nullptr},
                                                 {"hash_count", &count_nulls}, {"hash_min_max", &emit_null}};
    std::vector<Datum> host_args(aggs.size(), Datum(arg)), dev_args;
    for (size_t i = 0; i < aggs.size(); ++i) dev_args.push_back(DeviceDatum(arg));
    auto want = GET(cp::internal::GroupBy(host_args, {Datum(key)}, aggs, &cpu_ctx)).make_array();
    auto got = HostArray(GET(cp::cuda::GroupBy(dev_args, {DeviceDatum(key)}, aggs, &gpu_ctx)));
    ExpectStructClose(got, want, "GroupBy count/sum/min_max, null key, options (golden shape)");

    auto bad = cp::cuda::GroupBy({DeviceDatum(arg)}, {DeviceDatum(key)}, {{"hash_sum", nullptr}, {"hash_count", nullptr}}, &gpu_ctx);
    CHECK_TRUE(bad.status().IsInvalid(), "GroupBy aggregates/arguments count mismatch -> Invalid (hash_aggregate.cc:812-815)");
    auto unknown = cp::cuda::GroupBy({DeviceDatum(arg)}, {DeviceDatum(key)}, {{"hash_nope", nullptr}}, &gpu_ctx);
    CHECK_TRUE(unknown.status().IsKeyError(), "GroupBy unknown aggregate -> KeyError");
    // direct execution of a hash aggregate: code + message substring (hash_aggregate_test.cc:474-477)
    auto gid = Make<UInt32Builder, uint32_t>({0, 0, 1, 2, 3, 1, 3, 2});
    auto direct = cp::CallFunction("hash_sum", {DeviceDatum(arg), DeviceDatum(gid), Datum(uint32_t(4))}, &gpu_ctx);
    CHECK_TRUE(direct.status().IsNotImplemented() && direct.status().message().find("Direct execution of HASH_AGGREGATE functions") != std::string::npos,
               "direct hash_sum execution -> NotImplemented");
  }
  {
    const int64_t n = 400000;
    auto keys32 = MakeRandom<Int32Builder, int32_t>(n, rng, 0.01, 0, 1000, false);
    auto keys8 = MakeRandom<UInt8Builder, uint8_t>(n, rng, 0.0, 0, 3, false);
    auto keysb = MakeRandom<BooleanBuilder, int>(n, rng, 0.1, 0, 2, false);
    auto keys64 = MakeRandom<Int64Builder, int64_t>(n, rng, 0.0, -50000, 50000, false);
    auto vals = MakeRandom<DoubleBuilder, double>(n, rng, 0.01, 0.0, 1e14, true);
    auto ivals = MakeRandom<Int64Builder, int64_t>(n, rng, 0.02, -(1LL << 40), 1LL << 40, false);
    auto i32vals = MakeRandom<Int32Builder, int32_t>(n, rng, 0.02, -100000, 100000, false);
    std::vector<cp::internal::Aggregate> aggs = {{"hash_sum", nullptr}, {"hash_count", nullptr}, {"hash_min_max", nullptr}};
    struct Case { std::string name; std::vector<std::shared_ptr<Array>> keys; std::shared_ptr<Array> arg; };
    std::vector<Case> cases = {{"int32 key (1K groups, C3 shape) x float64", {keys32}, vals},
                               {"int64 key (100K groups) x int64", {keys64}, ivals},
                               {"(uint8, bool) keys (Q1 shape) x float64", {keys8, keysb}, vals},
                               {"(int32, uint8) keys x int32", {keys32, keys8}, i32vals},
                               // keys wider than 64 packed bits (FixedWidthKeyEncoder takes any, hash_aggregate.cc:142-196)
                               {"(int64, int64) keys x float64", {keys64, ivals}, vals},
                               {"(int64, int32, uint8) keys x int64", {keys64, keys32, keys8}, ivals},
                               {"(uint8, bool, int32, uint8, int32) keys x float64", {keys8, keysb, keys32, keys8, i32vals}, vals}};
    for (auto& c : cases) {
      std::vector<Datum> hk, dk, ha, da;
      for (auto& k : c.keys) { hk.push_back(k); dk.push_back(DeviceDatum(k)); }
      for (size_t i = 0; i < aggs.size(); ++i) { ha.push_back(c.arg); da.push_back(DeviceDatum(c.arg)); }
      auto want = GET(cp::internal::GroupBy(ha, hk, aggs, &cpu_ctx)).make_array();
      auto got = HostArray(GET(cp::cuda::GroupBy(da, dk, aggs, &gpu_ctx)));
      ExpectStructClose(got, want, "GroupBy random " + c.name);
    }
    // selection folded into the aggregation == the reference's Filter on every column, then its GroupBy
    {
      auto sel = MakeRandom<BooleanBuilder, int>(n, rng, 0.0, 0, 2, false);
      for (auto& c : cases) {
        std::vector<Datum> hk, dk, ha, da;
        for (auto& k : c.keys) { hk.push_back(GET(cp::CallFunction("filter", {Datum(k), Datum(sel)}, &cpu_ctx))); dk.push_back(DeviceDatum(k)); }
        auto farg = GET(cp::CallFunction("filter", {Datum(c.arg), Datum(sel)}, &cpu_ctx));
        for (size_t i = 0; i < aggs.size(); ++i) { ha.push_back(farg); da.push_back(DeviceDatum(c.arg)); }
        auto want = GET(cp::internal::GroupBy(ha, hk, aggs, &cpu_ctx)).make_array();
        auto got = HostArray(GET(cp::cuda::GroupBy(da, dk, aggs, DeviceDatum(sel), &gpu_ctx)));
        ExpectStructClose(got, want, "GroupBy with selection == GroupBy(Filter(...)) " + c.name);
      }
      auto with_nulls = MakeRandom<BooleanBuilder, int>(n, rng, 0.1, 0, 2, false);
      auto st = cp::cuda::GroupBy({DeviceDatum(vals)}, {DeviceDatum(keys32)}, {{"hash_sum", nullptr}}, DeviceDatum(with_nulls), &gpu_ctx).status();
      CHECK_TRUE(st.IsNotImplemented(), "GroupBy selection with null slots -> NotImplemented");
      auto st2 = cp::cuda::GroupBy({DeviceDatum(vals)}, {DeviceDatum(keys32)}, {{"hash_sum", nullptr}}, DeviceDatum(keys8), &gpu_ctx).status();
      CHECK_TRUE(st2.IsTypeError(), "GroupBy non-boolean selection -> TypeError");
    }
    // hash_mean == MeanImpl::Finalize over hash_sum / hash_count of the reference
    {
      auto want = GET(cp::internal::GroupBy({vals, vals}, {keys32}, {{"hash_sum", nullptr}, {"hash_count", nullptr}}, &cpu_ctx)).make_array();
      auto got = HostArray(GET(cp::cuda::GroupBy({DeviceDatum(vals)}, {DeviceDatum(keys32)}, {{"hash_mean", nullptr}}, &gpu_ctx)));
      const auto& w = internal::checked_cast<const StructArray&>(*want);
      const auto& g = internal::checked_cast<const StructArray&>(*got);
      const auto& sums = internal::checked_cast<const DoubleArray&>(*w.field(0));
      const auto& counts = internal::checked_cast<const Int64Array&>(*w.field(1));
      const auto& means = internal::checked_cast<const DoubleArray&>(*g.field(0));
      bool ok = g.length() == w.length() && g.type()->field(0)->name() == "hash_mean" && g.field(1)->Equals(w.field(2));
      for (int64_t i = 0; ok && i < g.length(); ++i) {
        if (counts.Value(i) == 0) ok = means.IsNull(i);
        else ok = means.IsValid(i) && std::fabs(means.Value(i) - sums.Value(i) / double(counts.Value(i))) <= 1e-12 * std::fabs(sums.Value(i) / double(counts.Value(i)));
      }
      CHECK_TRUE(ok, "GroupBy hash_mean == sum / count of the reference (1e-12)");
    }

    // Grouper twin: ids and uniques over two batches
    {
      auto k1 = keys64->Slice(0, n / 2), k2 = keys64->Slice(n / 2);
      auto cpu_grouper = GET(cp::internal::Grouper::Make({ValueDescr::Array(int64())}, &cpu_ctx));
      auto gpu_grouper = GET(cp::cuda::MakeGrouper({ValueDescr::Array(int64())}, &gpu_ctx));
      for (auto& k : {k1, k2}) {
        auto want = GET(cpu_grouper->Consume(GET(cp::ExecBatch::Make({Datum(k)})))).make_array();
        auto got = HostArray(GET(gpu_grouper->Consume(GET(cp::ExecBatch::Make({DeviceDatum(k)})))));
        ExpectEqual(got, want, "Grouper::Consume ids, batch of " + std::to_string(k->length()));
      }
      CHECK_TRUE(cpu_grouper->num_groups() == gpu_grouper->num_groups(), "Grouper::num_groups");
      auto want = GET(cpu_grouper->GetUniques());
      auto got = GET(gpu_grouper->GetUniques());
      ExpectEqual(HostArray(got[0]), want[0].make_array(), "Grouper::GetUniques");

      // the reference's GroupBy loop (hash_aggregate.cc:928-960) driven by hand over the CUDA registry's
      // HashAggregateKernels: ids -> consume -> finalize
      auto ids_want_grouper = GET(cp::internal::Grouper::Make({ValueDescr::Array(int64())}, &cpu_ctx));
      auto gpu_grouper2 = GET(cp::cuda::MakeGrouper({ValueDescr::Array(int64())}, &gpu_ctx));
      auto dkeys = DeviceDatum(keys64);
      auto dvals = DeviceDatum(ivals);
      auto ids = GET(gpu_grouper2->Consume(GET(cp::ExecBatch::Make({dkeys}))));
      for (const char* fname : {"hash_sum", "hash_count", "hash_min_max"}) {
        auto function = GET(registry->GetFunction(fname));
        auto kernel = static_cast<const cp::HashAggregateKernel*>(GET(function->DispatchExact({ValueDescr::Array(int64()), ValueDescr::Array(uint32()), ValueDescr::Scalar(uint32())})));
        cp::KernelContext kctx(&gpu_ctx);
        auto state = GET(kernel->init(&kctx, cp::KernelInitArgs{kernel, {int64(), uint32(), uint32()}, function->default_options()}));
        kctx.SetState(state.get());
        ABORT_NOT_OK(kernel->consume(&kctx, GET(cp::ExecBatch::Make({dvals, ids, Datum(gpu_grouper2->num_groups())}))));
        Datum out;
        ABORT_NOT_OK(kernel->finalize(&kctx, &out));
        auto want_all = GET(cp::internal::GroupBy({ivals}, {keys64}, {{fname, nullptr}}, &cpu_ctx)).make_array();
        ExpectEqual(HostArray(out), internal::checked_cast<const StructArray&>(*want_all).field(0), std::string("HashAggregateKernel consume/finalize over device ids: ") + fname);
      }
    }
  }

  // ---- unique / value_counts / dictionary_encode (vector_hash.cc:723-782) and the Grouper's other consumers
  {
    const int64_t n = 50000;
    std::vector<std::pair<std::string, std::shared_ptr<Array>>> columns = {
        {"int32 with nulls", MakeRandom<Int32Builder, int32_t>(n, rng, 0.05, -200, 200, false)},
        {"int64 no nulls", MakeRandom<Int64Builder, int64_t>(n, rng, 0.0, 0, 3000, false)},
        {"uint8 with nulls", MakeRandom<UInt8Builder, int>(n, rng, 0.2, 0, 255, false)},
        {"double with nulls", MakeRandom<DoubleBuilder, double>(n, rng, 0.1, 0.0, 50.0, false)},
        {"bool with nulls", MakeRandom<BooleanBuilder, int>(n, rng, 0.3, 0, 2, false)},
        {"tiny", Make<Int16Builder, int>({5, 7, 5, 0, 9, 7, 0, 3}, {true, true, true, false, true, true, false, true})},
        {"all null", Make<Int32Builder, int>({0, 0, 0}, {false, false, false})},
        {"empty", Make<Int32Builder, int>({})}};
    cp::DictionaryEncodeOptions mask_opts(cp::DictionaryEncodeOptions::MASK), encode_opts(cp::DictionaryEncodeOptions::ENCODE);
    for (auto& c : columns) {
      auto dev = DeviceDatum(c.second);
      ExpectEqual(HostArray(GET(cp::CallFunction("unique", {dev}, &gpu_ctx))), GET(cp::CallFunction("unique", {Datum(c.second)}, &cpu_ctx)).make_array(),
                  "unique " + c.first);
      ExpectEqual(HostArray(GET(cp::CallFunction("value_counts", {dev}, &gpu_ctx))),
                  GET(cp::CallFunction("value_counts", {Datum(c.second)}, &cpu_ctx)).make_array(), "value_counts " + c.first);
      for (auto* opts : {&mask_opts, &encode_opts}) {
        auto want = GET(cp::CallFunction("dictionary_encode", {Datum(c.second)}, opts, &cpu_ctx)).make_array();
        auto got = HostArray(GET(cp::CallFunction("dictionary_encode", {dev}, opts, &gpu_ctx)));
        ExpectEqual(got, want, std::string("dictionary_encode ") + (opts == &mask_opts ? "MASK " : "ENCODE ") + c.first);
      }
    }
    // chunked input: one memo table across the chunks (HashExec per chunk, vector_hash.cc:605-610)
    {
      auto col = columns[0].second;
      auto chunked = Chunked(col, {0, 1000, 1000, 20000, n});
      auto dchunked = DeviceChunked(*chunked);
      ExpectEqual(HostArray(GET(cp::CallFunction("unique", {Datum(dchunked)}, &gpu_ctx))),
                  GET(cp::CallFunction("unique", {Datum(chunked)}, &cpu_ctx)).make_array(), "unique chunked");
      ExpectEqual(HostArray(GET(cp::CallFunction("value_counts", {Datum(dchunked)}, &gpu_ctx))),
                  GET(cp::CallFunction("value_counts", {Datum(chunked)}, &cpu_ctx)).make_array(), "value_counts chunked");
      auto want = GET(cp::CallFunction("dictionary_encode", {Datum(chunked)}, &mask_opts, &cpu_ctx)).chunked_array();
      auto got = GET(cp::CallFunction("dictionary_encode", {Datum(dchunked)}, &mask_opts, &gpu_ctx)).chunked_array();
      CHECK_TRUE(got->num_chunks() == want->num_chunks(), "dictionary_encode chunked: chunk count");
      for (int i = 0; i < std::min(got->num_chunks(), want->num_chunks()); ++i) {
        ExpectEqual(MakeArray(ToHost(*got->chunk(i)->data())), want->chunk(i), "dictionary_encode chunked, chunk " + std::to_string(i));
      }
    }
    // dictionary columns: hashed / grouped through their indices, the dictionary rides along (DictionaryHashKernel
    // vector_hash.cc:520-603; DictionaryKeyEncoder hash_aggregate.cc:198-233)
    {
      auto indices = MakeRandom<Int8Builder, int>(n, rng, 0.1, 0, 5, false);
      auto dict_values = Make<Int64Builder, int64_t>({100, 200, 300, 400, 500});
      auto dict_type = dictionary(int8(), int64());
      auto darr = GET(DictionaryArray::FromArrays(dict_type, indices, dict_values));
      auto dev = DeviceDatum(darr);
      ExpectEqual(HostArray(GET(cp::CallFunction("unique", {dev}, &gpu_ctx))), GET(cp::CallFunction("unique", {Datum(darr)}, &cpu_ctx)).make_array(),
                  "unique of a dictionary array");
      ExpectEqual(HostArray(GET(cp::CallFunction("value_counts", {dev}, &gpu_ctx))),
                  GET(cp::CallFunction("value_counts", {Datum(darr)}, &cpu_ctx)).make_array(), "value_counts of a dictionary array");
      auto vals = MakeRandom<DoubleBuilder, double>(n, rng, 0.05, 0.0, 1.0, true);
      std::vector<cp::internal::Aggregate> aggs = {{"hash_sum", nullptr}, {"hash_count", nullptr}};
      auto want = GET(cp::internal::GroupBy({vals, vals}, {darr}, aggs, &cpu_ctx)).make_array();
      auto got = HostArray(GET(cp::cuda::GroupBy({DeviceDatum(vals), DeviceDatum(vals)}, {dev}, aggs, &gpu_ctx)));
      ExpectStructClose(got, want, "GroupBy with a dictionary key");
      auto other = GET(DictionaryArray::FromArrays(dict_type, indices, Make<Int64Builder, int64_t>({1, 2, 3, 4, 5})));
      auto two = std::make_shared<ChunkedArray>(ArrayVector{MakeArray(dev.array()), MakeArray(ToDevice(*other->data()))}, dict_type);
      auto vals2 = std::make_shared<ChunkedArray>(ArrayVector{MakeArray(ToDevice(*vals->data())), MakeArray(ToDevice(*vals->data()))}, float64());
      auto st = cp::cuda::GroupBy({Datum(vals2)}, {Datum(two)}, {{"hash_sum", nullptr}}, &gpu_ctx).status();
      CHECK_TRUE(st.IsNotImplemented(), "GroupBy: batches with differing key dictionaries -> NotImplemented");
    }
    // binary / utf8 keys (VarLengthKeyEncoder, hash_aggregate.cc:236-311) and the hash kernels on strings
    {
      std::vector<std::string> vocab = {"", "a", std::string("a\0", 2), std::string("a\0b", 3), "ab", "abc", "abcdefgh", "abcdefghi"};
      std::uniform_int_distribution<int> len(0, 20), byte(0, 255);
      while (vocab.size() < 400) {
        std::string w(size_t(len(rng)), '\0');
        for (auto& ch : w) ch = char(byte(rng) & 0x7F);   // ASCII: utf8 arrays must pass ValidateFull
        vocab.push_back(w);
      }
      std::uniform_int_distribution<size_t> pick(0, vocab.size() - 1);
      std::uniform_real_distribution<double> u(0, 1);
      StringBuilder sb;
      BinaryBuilder bb;
      for (int64_t i = 0; i < n; ++i) {
        if (u(rng) < 0.05) { ABORT_NOT_OK(sb.AppendNull()); ABORT_NOT_OK(bb.AppendNull()); continue; }
        const std::string& w = vocab[pick(rng)];
        ABORT_NOT_OK(sb.Append(w));
        ABORT_NOT_OK(bb.Append(w));
      }
      auto strings = GET(sb.Finish());
      auto binaries = GET(bb.Finish());
      auto vals = MakeRandom<DoubleBuilder, double>(n, rng, 0.05, 0.0, 1.0, true);
      auto ikeys = MakeRandom<Int32Builder, int32_t>(n, rng, 0.05, 0, 4, false);
      std::vector<cp::internal::Aggregate> aggs = {{"hash_sum", nullptr}, {"hash_count", nullptr}, {"hash_min_max", nullptr}};
      for (auto& kc : std::vector<std::pair<std::string, std::vector<std::shared_ptr<Array>>>>{
               {"utf8", {strings}}, {"binary", {binaries}}, {"utf8 + int32", {strings, ikeys}}, {"int32 + binary + utf8", {ikeys, binaries, strings}}}) {
        std::vector<Datum> hk, dk;
        for (auto& k : kc.second) { hk.push_back(k); dk.push_back(DeviceDatum(k)); }
        auto want = GET(cp::internal::GroupBy({vals, vals, vals}, hk, aggs, &cpu_ctx)).make_array();
        auto got = HostArray(GET(cp::cuda::GroupBy({DeviceDatum(vals), DeviceDatum(vals), DeviceDatum(vals)}, dk, aggs, &gpu_ctx)));
        ExpectStructClose(got, want, "GroupBy keyed by " + kc.first);
      }
      // chunked: later chunks bring new strings
      {
        auto ck = Chunked(strings, {0, 100, 100, n / 2, n}), cv = Chunked(vals, {0, 100, 100, n / 2, n});
        auto want = GET(cp::internal::GroupBy({Datum(cv)}, {Datum(ck)}, {{"hash_sum", nullptr}}, &cpu_ctx)).make_array();
        auto got = HostArray(GET(cp::cuda::GroupBy({Datum(DeviceChunked(*cv))}, {Datum(DeviceChunked(*ck))}, {{"hash_sum", nullptr}}, &gpu_ctx)));
        ExpectStructClose(got, want, "GroupBy keyed by utf8, chunked");
      }
      auto dev = DeviceDatum(strings);
      ExpectEqual(HostArray(GET(cp::CallFunction("unique", {dev}, &gpu_ctx))), GET(cp::CallFunction("unique", {Datum(strings)}, &cpu_ctx)).make_array(),
                  "unique utf8");
      ExpectEqual(HostArray(GET(cp::CallFunction("value_counts", {dev}, &gpu_ctx))),
                  GET(cp::CallFunction("value_counts", {Datum(strings)}, &cpu_ctx)).make_array(), "value_counts utf8");
      cp::DictionaryEncodeOptions mask_opts(cp::DictionaryEncodeOptions::MASK), encode_opts(cp::DictionaryEncodeOptions::ENCODE);
      for (auto* opts : {&mask_opts, &encode_opts}) {
        auto want = GET(cp::CallFunction("dictionary_encode", {Datum(strings)}, opts, &cpu_ctx)).make_array();
        auto got = HostArray(GET(cp::CallFunction("dictionary_encode", {dev}, opts, &gpu_ctx)));
        ExpectEqual(got, want, std::string("dictionary_encode utf8 ") + (opts == &mask_opts ? "MASK" : "ENCODE"));
      }
      // the Grouper twin on (utf8, int32)
      auto cpu_grouper = GET(cp::internal::Grouper::Make({ValueDescr::Array(utf8()), ValueDescr::Array(int32())}, &cpu_ctx));
      auto gpu_grouper = GET(cp::cuda::MakeGrouper({ValueDescr::Array(utf8()), ValueDescr::Array(int32())}, &gpu_ctx));
      auto want_ids = GET(cpu_grouper->Consume(GET(cp::ExecBatch::Make({Datum(strings), Datum(ikeys)})))).make_array();
      auto got_ids = HostArray(GET(gpu_grouper->Consume(GET(cp::ExecBatch::Make({dev, DeviceDatum(ikeys)})))));
      ExpectEqual(got_ids, want_ids, "Grouper::Consume ids keyed by (utf8, int32)");
      auto want_u = GET(cpu_grouper->GetUniques());
      auto got_u = GET(gpu_grouper->GetUniques());
      ExpectEqual(HostArray(got_u[0]), want_u[0].make_array(), "Grouper::GetUniques utf8 column");
      ExpectEqual(HostArray(got_u[1]), want_u[1].make_array(), "Grouper::GetUniques int32 column");
    }
    // MakeGroupings / ApplyGroupings (hash_aggregate.cc:968-1018) over the Grouper twin's ids
    {
      auto keys = MakeRandom<Int32Builder, int32_t>(n, rng, 0.05, 0, 700, false);
      auto payload = MakeRandom<DoubleBuilder, double>(n, rng, 0.1, 0.0, 1.0, true);
      auto cpu_grouper = GET(cp::internal::Grouper::Make({ValueDescr::Array(int32())}, &cpu_ctx));
      auto gpu_grouper = GET(cp::cuda::MakeGrouper({ValueDescr::Array(int32())}, &gpu_ctx));
      auto want_ids = GET(cpu_grouper->Consume(GET(cp::ExecBatch::Make({Datum(keys)})))).make_array();
      auto got_ids = GET(gpu_grouper->Consume(GET(cp::ExecBatch::Make({DeviceDatum(keys)}))));
      const uint32_t groups = cpu_grouper->num_groups();
      auto want = GET(cp::internal::Grouper::MakeGroupings(internal::checked_cast<const UInt32Array&>(*want_ids), groups + 3, &cpu_ctx));
      auto got = GET(cp::cuda::MakeGroupings(*got_ids.array(), groups + 3));
      ExpectEqual(MakeArray(ToHost(*got->data())), want, "MakeGroupings (with empty trailing groups)");
      auto want_applied = GET(cp::internal::Grouper::ApplyGroupings(*want, *payload, &cpu_ctx));
      auto got_applied = GET(cp::cuda::ApplyGroupings(*got, *ToDevice(*payload->data())));
      ExpectEqual(MakeArray(ToHost(*got_applied->data())), want_applied, "ApplyGroupings");
      auto st = cp::cuda::MakeGroupings(*got_ids.array(), groups - 1).status();
      CHECK_TRUE(st.IsIndexError(), "MakeGroupings: id >= num_groups -> IndexError");
    }
  }

  std::cout << (g_failures ? "GLUE FAILED: " : "GLUE OK: ") << g_checks - g_failures << "/" << g_checks << " checks passed" << std::endl;
  return g_failures ? 1 : 0;
}
