// b200_arrow.h — the Arrow-side host glue over the C ABI (include/arrow1_b200.h).
//
// This is the reference-facing half of the drop-in boundary: C++ that links the
// UNMODIFIED reference libarrow + arrow_cuda and registers CUDA-backed Functions under the
// reference's own names in a compute::FunctionRegistry, so that
//
//     auto registry = compute::cuda::MakeCudaRegistry().ValueOrDie();
//     compute::ExecContext ctx(default_memory_pool(), registry.get());
//     compute::CallFunction("filter", {device_values, device_mask}, &options, &ctx);
//
// runs the sm_100a kernels on CudaBuffer-resident ArrayData (gpu/cuda_memory.h:37-109) and hands
// back CudaBuffer-resident results.  Nothing here dereferences a data buffer on the host, and
// nothing here computes: every operation is a call into libarrow1_b200.so.
//
// Registered names (exactly the reference's, compute/registry.cc + the kernels' Register*):
//   filter, array_filter                 kernels/vector_selection.cc:1877-1903, 2133-2152
//   sort_indices, array_sort_indices     kernels/vector_sort.cc:1565-1711, 1741-1755
//   hash_count, hash_sum, hash_min_max   kernels/hash_aggregate.cc:1039-1062
//   hash_mean                            absent upstream; MeanImpl::Finalize over the sum state
//   take, array_take                     kernels/vector_selection.cc:1905-2081, 2154-2176
//   unique, value_counts, dictionary_encode   kernels/vector_hash.cc:723-782
#pragma once

#include <memory>
#include <vector>

#include "arrow/array/array_nested.h"
#include "arrow/c/abi.h"
#include "arrow/compute/api_aggregate.h"
#include "arrow/compute/exec.h"
#include "arrow/compute/registry.h"
#include "arrow/datum.h"
#include "arrow/result.h"
#include "arrow/status.h"

namespace arrow {
namespace compute {
namespace cuda {

/// Adds the CUDA-backed functions to `registry` (a registry made with FunctionRegistry::Make(),
/// registry.h:51).  `device` is the CUDA device the kernels run on; inputs must live there.
/// Returns KeyError if a function of the same name is already registered (registry.cc:43-47).
Status RegisterCudaKernels(FunctionRegistry* registry, int device = 0);

/// A fresh registry holding only the CUDA-backed functions; select it with
/// ExecContext(pool, registry) (exec.h:61-62).
Result<std::unique_ptr<FunctionRegistry>> MakeCudaRegistry(int device = 0);

/// compute::internal::GroupBy (api_aggregate.h:400-403, hash_aggregate.cc:890-966) over
/// device-resident arguments and keys: same signature, same output (a StructArray with one field
/// per aggregate named after its function, then key_0..key_n, groups in first-occurrence order),
/// all buffers CudaBuffers.  Aggregates are resolved by name through ctx->func_registry(), like
/// GetKernels (hash_aggregate.cc:809-830), so the registry must hold the functions above.
/// Grouping and every aggregate run fused in one pass over the rows (no group-id column is
/// materialised).
Result<Datum> GroupBy(const std::vector<Datum>& arguments, const std::vector<Datum>& keys,
                      const std::vector<internal::Aggregate>& aggregates, ExecContext* ctx);

/// GroupBy over the rows selected by a device-resident boolean column (no null slots):
/// == GroupBy over Filter(column, selection) of every argument and key (FilterOptions::DROP), group
/// order included, without materialising the filtered columns (b200_groupby_consume_selected).
/// No counterpart upstream, where this is Filter on every column followed by GroupBy.
Result<Datum> GroupBy(const std::vector<Datum>& arguments, const std::vector<Datum>& keys,
                      const std::vector<internal::Aggregate>& aggregates, const Datum& selection,
                      ExecContext* ctx);

/// compute::internal::Grouper (api_aggregate.h:313-331) over device-resident key batches.
/// The reference's Grouper::Make is a hard-wired static factory (hash_aggregate.cc:885-888) and
/// cannot be swapped through the registry, hence this twin.  Consume() returns a device-resident
/// UInt32Array of group ids (ids in first-occurrence order), GetUniques() device-resident keys.
Result<std::unique_ptr<internal::Grouper>> MakeGrouper(const std::vector<ValueDescr>& descrs,
                                                       ExecContext* ctx, int device = 0);

/// Grouper::MakeGroupings / ApplyGroupings (api_aggregate.h:333-386, hash_aggregate.cc:968-1018) over device-resident
/// group ids (the output of the Grouper twin's Consume) and arrays: the reference's are static functions that walk
/// the ids on the host.  Results are device-resident ListArrays (offsets and values in CudaBuffers).
Result<std::shared_ptr<ListArray>> MakeGroupings(const ArrayData& ids, uint32_t num_groups, int device = 0);
Result<std::shared_ptr<ListArray>> ApplyGroupings(const ListArray& groupings, const ArrayData& array, int device = 0);

/// Ingest / egress of columns.  CopyToDevice returns the same ArrayData with every host buffer
/// replaced by a pooled CudaBuffer holding its bytes (offsets kept, the null count settled on the host first);
/// CopyToHost is the way back, into pinned CudaHostBuffers.  Pinned sources / destinations (CudaHostBuffer,
/// AllocatePinned, registered memory) move by DMA at PCIe speed; pageable ones go through two pinned staging blocks,
/// the CPU copy of one chunk overlapping the DMA of the previous one.  Replaces CudaBuffer::CopyFromHost /
/// CopyToHost (gpu/cuda_memory.h:64-71), which are synchronous copies from pageable memory.
Result<std::shared_ptr<ArrayData>> CopyToDevice(const ArrayData& host, int device = 0);
Result<std::shared_ptr<ArrayData>> CopyToHost(const ArrayData& device_data, int device = 0);
/// A pinned host buffer (CudaHostBuffer) for columns that are produced on the host and consumed on the device.
Result<std::shared_ptr<Buffer>> AllocatePinned(int64_t nbytes, int device = 0);
/// Arrow C Data Interface (arrow/c/abi.h, c/bridge.h): an exported host array -> device-resident ArrayData (the
/// structs are consumed, like ImportArray), and a device-resident array exported as a host array.
Result<std::shared_ptr<ArrayData>> ImportToDevice(struct ArrowArray* array, struct ArrowSchema* schema, int device = 0);
Status ExportFromDevice(const ArrayData& device_data, struct ArrowArray* out, struct ArrowSchema* out_schema, int device = 0);

}  // namespace cuda
}  // namespace compute
}  // namespace arrow
