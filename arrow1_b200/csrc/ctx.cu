// Context: one device + one stream, error message storage, small scratch (SURVEY §8b "Threading").
#include <stdarg.h>

#include <cstring>

#include "common.cuh"

int b200_fail(b200_ctx* ctx, int code, const char* fmt, ...) {
  B200_ENTER(ctx);
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (ctx) ctx->error = buf;
  return code;
}

int b200_ctx_reserve_scratch(b200_ctx* ctx, size_t bytes) {
  B200_ENTER(ctx);
  if (bytes <= ctx->scratch_bytes) return B200_OK;
  if (ctx->scratch) {
    B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    B200_CUDA(ctx, cudaFree(ctx->scratch));
    ctx->scratch = nullptr;
    ctx->scratch_bytes = 0;
  }
  bytes = b200_align_up(bytes, 1 << 20);
  B200_CUDA(ctx, cudaMalloc(&ctx->scratch, bytes));
  ctx->scratch_bytes = bytes;
  return B200_OK;
}

int b200_dev_alloc(b200_ctx* ctx, void** out, size_t bytes) {
  B200_ENTER(ctx);
  bytes = b200_align_up(bytes ? bytes : 1, 512);
  int best = -1;
  for (size_t i = 0; i < ctx->blocks.size(); ++i) {
    const b200_ctx::Block& b = ctx->blocks[i];
    if (!b.in_use && b.bytes >= bytes && b.bytes <= 2 * bytes + (1 << 20) && (best < 0 || b.bytes < ctx->blocks[best].bytes)) best = int(i);
  }
  if (best >= 0) {
    ctx->blocks[best].in_use = true;
    *out = ctx->blocks[best].ptr;
    return B200_OK;
  }
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess) {
    // give cached blocks back and retry once
    cudaGetLastError();
    cudaStreamSynchronize(ctx->stream);
    for (size_t i = 0; i < ctx->blocks.size();) {
      if (!ctx->blocks[i].in_use) {
        cudaFree(ctx->blocks[i].ptr);
        ctx->blocks.erase(ctx->blocks.begin() + i);
      } else {
        ++i;
      }
    }
    e = cudaMalloc(&p, bytes);
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    return b200_fail(ctx, B200_OUT_OF_MEMORY, "device allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
  }
  ctx->blocks.push_back({p, bytes, true});
  *out = p;
  return B200_OK;
}

void b200_dev_free(b200_ctx* ctx, void* ptr) {
  if (!ptr) return;
  size_t idle = 0;
  for (auto& b : ctx->blocks) {
    if (b.ptr == ptr) b.in_use = false;
    if (!b.in_use) ++idle;
  }
  if (idle > 24) {  // bound the cache: drop the largest idle block
    int big = -1;
    for (size_t i = 0; i < ctx->blocks.size(); ++i)
      if (!ctx->blocks[i].in_use && (big < 0 || ctx->blocks[i].bytes > ctx->blocks[big].bytes)) big = int(i);
    if (big >= 0) {
      cudaStreamSynchronize(ctx->stream);
      cudaFree(ctx->blocks[big].ptr);
      ctx->blocks.erase(ctx->blocks.begin() + big);
    }
  }
}

void b200_time_begin(b200_ctx* ctx, const char* name) {
  if (!ctx->timing) return;
  b200_ctx::TimedLaunch t;
  t.name = name;
  cudaEventCreate(&t.start);
  cudaEventCreate(&t.stop);
  cudaEventRecord(t.start, ctx->stream);
  ctx->timed.push_back(t);
}

void b200_time_end(b200_ctx* ctx) {
  if (!ctx->timing || ctx->timed.empty()) return;
  cudaEventRecord(ctx->timed.back().stop, ctx->stream);
}

static void clear_timed(b200_ctx* ctx) {
  for (auto& t : ctx->timed) {
    cudaEventDestroy(t.start);
    cudaEventDestroy(t.stop);
  }
  ctx->timed.clear();
}

extern "C" {

int b200_ctx_set_kernel_timing(b200_ctx* ctx, int enabled) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  cudaStreamSynchronize(ctx->stream);
  clear_timed(ctx);
  ctx->timing = enabled != 0;
  return B200_OK;
}

int b200_ctx_kernel_time(b200_ctx* ctx, const char* name, double* total_ms, int64_t* launches) {
  B200_ENTER(ctx);
  if (!ctx || !name) return B200_INVALID;
  B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  double total = 0;
  int64_t n = 0;
  for (auto& t : ctx->timed) {
    if (strcmp(t.name, name) != 0) continue;
    float ms = 0;
    if (cudaEventElapsedTime(&ms, t.start, t.stop) == cudaSuccess) {
      total += ms;
      ++n;
    }
  }
  if (total_ms) *total_ms = total;
  if (launches) *launches = n;
  return B200_OK;
}

const char* b200_version(void) { return "arrow1_b200 0.1 (sm_100a)"; }

int b200_ctx_create(int device, void* cu_stream, b200_ctx** out) {
  if (!out) return B200_INVALID;
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    // no CPU fallback: fail loudly
    fprintf(stderr, "arrow1_b200: no CUDA device available (%s)\n",
            e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    return B200_UNKNOWN_ERROR;
  }
  if (device < 0 || device >= n) return B200_INVALID;
  b200_ctx* ctx = new b200_ctx();
  ctx->device = device;
  ctx->stream = reinterpret_cast<cudaStream_t>(cu_stream);
  if (cudaSetDevice(device) != cudaSuccess) {
    delete ctx;
    return B200_UNKNOWN_ERROR;
  }
  cudaDeviceProp prop = {};
  if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) ctx->num_sms = prop.multiProcessorCount;
  if (prop.major < 10) {
    fprintf(stderr, "arrow1_b200: device %d is sm_%d%d; this library is built for sm_100a only\n",
            device, prop.major, prop.minor);
    delete ctx;
    return B200_NOT_IMPLEMENTED;
  }
  if (cudaMallocHost(reinterpret_cast<void**>(&ctx->host_words), 64 * sizeof(int64_t)) != cudaSuccess) {
    delete ctx;
    return B200_OUT_OF_MEMORY;
  }
  if (cudaMalloc(reinterpret_cast<void**>(&ctx->async_err), 256) != cudaSuccess || cudaMemset(ctx->async_err, 0, 256) != cudaSuccess) {
    cudaFreeHost(ctx->host_words);
    delete ctx;
    return B200_OUT_OF_MEMORY;
  }
  *out = ctx;
  return B200_OK;
}

int b200_ctx_destroy(b200_ctx* ctx) {
  if (!ctx) return B200_OK;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  clear_timed(ctx);
  if (ctx->scratch) cudaFree(ctx->scratch);
  for (auto& b : ctx->blocks) cudaFree(b.ptr);
  ctx->blocks.clear();
  if (ctx->host_words) cudaFreeHost(ctx->host_words);
  if (ctx->async_err) cudaFree(ctx->async_err);
  delete ctx;
  return B200_OK;
}

int b200_ctx_set_stream(b200_ctx* ctx, void* cu_stream) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  cudaStreamSynchronize(ctx->stream);  // cached device blocks are reused in stream order
  ctx->stream = reinterpret_cast<cudaStream_t>(cu_stream);
  return B200_OK;
}

int b200_ctx_synchronize(b200_ctx* ctx) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  // the word kernels set when they had to give up rides back with the synchronisation
  B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words + 63, ctx->async_err, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
  B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  const uint32_t err = *reinterpret_cast<const uint32_t*>(ctx->host_words + 63);
  if (err != 0) {
    B200_CUDA(ctx, cudaMemsetAsync(ctx->async_err, 0, sizeof(uint32_t), ctx->stream));
    return b200_fail(ctx, B200_UNKNOWN_ERROR, "a kernel on this context gave up (code %u: %s); results since the last synchronisation are invalid",
                     err,
                     err == 1 ? "sort look-back timed out waiting for a preceding tile"
                              : (err == 2 ? "filter output buffers smaller than the plan's output length" : "injected"));
  }
  return B200_OK;
}

int b200_ctx_inject_async_error(b200_ctx* ctx, uint32_t code) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  B200_CUDA(ctx, cudaMemcpyAsync(ctx->async_err, &code, sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
  B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return B200_OK;
}

int b200_ctx_set_option(b200_ctx* ctx, const char* name, int64_t value) {
  if (!ctx || !name) return B200_INVALID;
  if (!strcmp(name, "sort_high_word_first")) ctx->sort_high_word_first = value > 0 ? 1 : (value < 0 ? -1 : 0);
  else return b200_fail(ctx, B200_KEY_ERROR, "context: no option named %s", name);
  return B200_OK;
}

const char* b200_last_error(b200_ctx* ctx) { return ctx ? ctx->error.c_str() : "null context"; }

int64_t b200_ctx_launch_count(b200_ctx* ctx) { return ctx ? ctx->launches : 0; }

}  // extern "C"
