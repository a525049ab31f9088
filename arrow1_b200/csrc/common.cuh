// Shared host/device plumbing for the sm_100a kernels: context, error capture, Arrow bitmap access.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "arrow1_b200.h"

struct b200_ctx {
  int device = 0;
  int sort_high_word_first = 0;   // b200_ctx_set_option: 0 = decided per call from the digit histograms, 1 = forced, -1 = off
  cudaStream_t stream = nullptr;
  int num_sms = 148;
  int64_t launches = 0;
  std::string error;
  // small device scratch owned by the context (flags, counters); grown on demand
  void* scratch = nullptr;
  size_t scratch_bytes = 0;
  // pinned host word block for scalar read-backs
  int64_t* host_words = nullptr;
  // optional per-kernel CUDA-event timing (bench.py's roofline leg)
  bool timing = false;
  struct TimedLaunch {
    const char* name;
    cudaEvent_t start, stop;
  };
  std::vector<TimedLaunch> timed;
  // device blocks handed back by b200_dev_free, reused by b200_dev_alloc (all work of a context is
  // ordered on its one stream, so reuse needs no synchronisation)
  struct Block {
    void* ptr;
    size_t bytes;
    bool in_use;
  };
  std::vector<Block> blocks;
  // kernels whose >48 KB dynamic shared memory opt-in has been set on this context's device (one bit per site:
  // cudaFuncSetAttribute applies per device, so the opt-in is tracked per context, not per process)
  uint32_t attr_done = 0;
  // device word a kernel sets when it had to give up (the sort's look-back bail-out); read back by
  // b200_ctx_synchronize, which fails with B200_UNKNOWN_ERROR when it is non-zero
  uint32_t* async_err = nullptr;
};

// first call for `bit` on this context -> true
static inline bool b200_ctx_once(b200_ctx* ctx, int bit) {
  if (ctx->attr_done & (1u << bit)) return false;
  ctx->attr_done |= 1u << bit;
  return true;
}

// Every entry point runs on its context's device whatever device is current in the calling thread (a process that
// holds contexts on several GPUs, or a fresh thread whose current device is 0), and puts the caller's device back.
struct b200_device_guard {
  int prev = -1;
  bool switched = false;
  explicit b200_device_guard(const b200_ctx* ctx) {
    if (ctx && cudaGetDevice(&prev) == cudaSuccess && prev != ctx->device) switched = cudaSetDevice(ctx->device) == cudaSuccess;
  }
  ~b200_device_guard() {
    if (switched) cudaSetDevice(prev);
  }
  b200_device_guard(const b200_device_guard&) = delete;
  b200_device_guard& operator=(const b200_device_guard&) = delete;
};
#define B200_ENTER(ctx) b200_device_guard _b200_device_guard(ctx)

// caching device allocator of a context: cudaMalloc / cudaFree stay out of the steady state
int b200_dev_alloc(b200_ctx* ctx, void** out, size_t bytes);
void b200_dev_free(b200_ctx* ctx, void* ptr);

void b200_time_begin(b200_ctx* ctx, const char* name);
void b200_time_end(b200_ctx* ctx);

// a kernel launch bracketed by CUDA events on the launching stream when timing is enabled
#define B200_TIMED(ctx, name, launch)   \
  do {                                  \
    b200_time_begin((ctx), (name));     \
    launch;                             \
    b200_time_end((ctx));               \
    B200_LAUNCHED(ctx);                 \
  } while (0)

int b200_fail(b200_ctx* ctx, int code, const char* fmt, ...);
int b200_ctx_reserve_scratch(b200_ctx* ctx, size_t bytes);

#define B200_CUDA(ctx, expr)                                                              \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess)                                                                \
      return b200_fail((ctx), _e == cudaErrorMemoryAllocation ? B200_OUT_OF_MEMORY        \
                                                              : B200_UNKNOWN_ERROR,       \
                       "CUDA error %s at %s:%d: %s", cudaGetErrorName(_e), __FILE__,      \
                       __LINE__, cudaGetErrorString(_e));                                 \
  } while (0)

#define B200_RETURN_NOT_OK(expr) \
  do {                           \
    int _s = (expr);             \
    if (_s != B200_OK) return _s; \
  } while (0)

// launch bookkeeping + error check after every kernel launch
#define B200_LAUNCHED(ctx)                     \
  do {                                         \
    (ctx)->launches++;                         \
    B200_CUDA((ctx), cudaGetLastError());      \
  } while (0)

static inline int b200_type_bits(int t) {
  switch (t) {
    case B200_BOOL: return 1;
    case B200_UINT8: case B200_INT8: return 8;
    case B200_UINT16: case B200_INT16: return 16;
    case B200_UINT32: case B200_INT32: case B200_FLOAT: return 32;
    case B200_UINT64: case B200_INT64: case B200_DOUBLE: return 64;
    default: return 0;
  }
}

static inline size_t b200_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ------------------------------------------------------------------------------------------
// Arrow bitmaps on the device.  A (pointer, bit offset) pair is normalised to an 8-byte aligned
// word pointer plus a bit offset; reading the aligned words that cover the wanted bits never
// leaves the allocation (CUDA allocations are >= 256-byte aligned and sized).
struct BitmapRef {
  const uint64_t* words;  // 8-byte aligned; nullptr = "all ones"
  int64_t bit_offset;     // >= 0
};

static inline BitmapRef make_bitmap_ref(const void* p, int64_t offset_bits) {
  BitmapRef r;
  if (p == nullptr) {
    r.words = nullptr;
    r.bit_offset = 0;
    return r;
  }
  uintptr_t a = reinterpret_cast<uintptr_t>(p);
  r.words = reinterpret_cast<const uint64_t*>(a & ~uintptr_t(7));
  r.bit_offset = offset_bits + int64_t(a & 7) * 8;
  return r;
}

#ifdef __CUDACC__
// 64 bits of the bitmap starting at logical bit `pos` (LSB = bit pos).  Bits at or beyond
// `length` are returned as 0.  A null bitmap reads as all ones (inside the length).
__device__ __forceinline__ uint64_t bitmap_load64(const BitmapRef& b, int64_t pos, int64_t length) {
  if (pos >= length) return 0;
  uint64_t w;
  if (b.words == nullptr) {
    w = ~uint64_t(0);
  } else {
    int64_t bit = b.bit_offset + pos;
    int64_t wi = bit >> 6;
    int sh = int(bit & 63);
    uint64_t lo = __ldg(b.words + wi);
    w = lo >> sh;
    int64_t last_bit = b.bit_offset + (length - 1);
    if (sh != 0 && (wi + 1) <= (last_bit >> 6)) {
      uint64_t hi = __ldg(b.words + wi + 1);
      w |= hi << (64 - sh);
    }
  }
  int64_t remain = length - pos;
  if (remain < 64) w &= (uint64_t(1) << remain) - 1;
  return w;
}

__device__ __forceinline__ uint32_t bitmap_load32(const BitmapRef& b, int64_t pos, int64_t length) {
  if (pos >= length) return 0;
  uint32_t w;
  if (b.words == nullptr) {
    w = ~uint32_t(0);
  } else {
    int64_t bit = b.bit_offset + pos;
    const uint32_t* w32 = reinterpret_cast<const uint32_t*>(b.words);
    int64_t wi = bit >> 5;
    int sh = int(bit & 31);
    uint32_t lo = __ldg(w32 + wi);
    uint32_t hi = 0;
    int64_t last_bit = b.bit_offset + (length - 1);
    if (sh != 0 && (wi + 1) <= (last_bit >> 5)) hi = __ldg(w32 + wi + 1);
    w = __funnelshift_r(lo, hi, sh);
  }
  int64_t remain = length - pos;
  if (remain < 32) w &= (uint32_t(1) << remain) - 1;
  return w;
}

__device__ __forceinline__ bool bitmap_get(const BitmapRef& b, int64_t pos) {
  if (b.words == nullptr) return true;
  int64_t bit = b.bit_offset + pos;
  const uint8_t* bytes = reinterpret_cast<const uint8_t*>(b.words);
  return (__ldg(bytes + (bit >> 3)) >> (bit & 7)) & 1;
}

__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}
#endif
