// m2k_scan_flags_noalloc: the byte scan for callers that must not allocate GPU memory.
//
// Who that is: a host that drives this library from inside a region where cudaMalloc is off limits --
// a plugin called while the application is capturing or replaying work it laid out beforehand, a
// service whose device memory is handed out once at start-up by its own arena and then locked, or a
// process sharing the GPU under a fixed memory quota.  cudaMalloc there either fails, or stalls the
// whole device while the driver maps memory.  Those callers own every buffer they pass in; the only
// memory the scan needs beyond its inputs and outputs is the per-chunk record table (8 bytes per
// 64 KiB of content), and here that table lives in a fixed array inside the library image.
//
// The array costs kNoAllocMaxChunks * 8 bytes (2 MiB) on EVERY GPU that loads the library, until the
// process exits, whether or not this entry point is ever called.  That bounds one call to
// kNoAllocMaxChunks chunks (16 GiB when every chunk is full); larger inputs are refused.
//
// The first call on a device resolves the array's device address (cudaGetSymbolAddress) and reads
// the SM count; make that first call before starting any stream capture.  All calls share the one
// array: it keeps whatever the previous call left, which is harmless because every record a call
// reads is written by that same call first -- nothing in it has to start at zero.  The one word that
// must start at zero, the counter the scan draws its chunks from, is cleared on the caller's stream
// by every call.  Two calls whose
// kernels overlap on the GPU (different streams, or different threads) would overwrite each other's
// records: run them one after the other.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

#include <mutex>

#include "m2k_internal.h"

namespace m2k {

constexpr uint32_t kNoAllocMaxChunks = 262144;  // 2 MiB of records

// In host code this name is only a handle for the runtime, not an address a kernel can use.
static __device__ __align__(256) ChunkRec g_noalloc_records[kNoAllocMaxChunks];
// the scan hands out chunks from a counter; it lives next to the table and IS cleared by every call
static __device__ uint32_t g_noalloc_ticket;

constexpr int kMaxDevices = 64;
static std::mutex g_setup_mutex;
static ChunkRec* g_records_ptr[kMaxDevices];  // device address of g_noalloc_records, per device
static uint32_t* g_ticket_ptr[kMaxDevices];   // ... and of g_noalloc_ticket
static int g_sm_count[kMaxDevices];

// Per-file flags from the chunk records: one thread per file walks its records in order and keeps
// the bytes since the last newline; the records are exact to the byte, so the file itself is not
// looked at again.  Device code may name the array directly.
// `n_chunks` is what the host checked against the table and what the scan wrote: a prefix array that
// runs past it (it must not: n_chunks == chunk_first[n_files]) is cut off there, so this kernel never
// reads a record this call did not write, nor past the table.
__global__ void __launch_bounds__(128) k_noalloc_file_flags(uint32_t n_files, uint32_t n_chunks, const uint32_t* __restrict__ chunk_first,
                                                             const uint64_t* __restrict__ file_len, uint8_t* __restrict__ file_flags) {
  uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n_files) return;
  const uint64_t len = file_len[f];
  uint32_t has_main = 0, is_long = 0;
  uint64_t run = 0;  // bytes since the last newline (or since the start of the file)
  const uint32_t c_end = chunk_first[f + 1] < n_chunks ? chunk_first[f + 1] : n_chunks;
  for (uint32_t c = chunk_first[f], k = 0; c < c_end; c++, k++) {
    const ChunkRec rec = g_noalloc_records[c];
    const uint64_t b0 = (uint64_t)k * kChunkBytes;
    const uint64_t nb = len - b0 < kChunkBytes ? len - b0 : kChunkBytes;
    has_main |= rec_main(rec);
    if (rec_lead(rec) == nb) { run += nb; continue; }  // no newline in this chunk
    if (run + rec_lead(rec) >= kLongLine) is_long = 1;
    run = rec_trail(rec);
  }
  if (run >= kLongLine) is_long = 1;
  file_flags[f] = (uint8_t)((has_main ? M2K_FILE_HAS_MAIN : 0) | (is_long ? M2K_FILE_LONG_LINE : 0));
}

}  // namespace m2k

using namespace m2k;

// messages go where every other entry point's go: m2k_global_error()
static int noalloc_fail(int code, const char* fmt, ...) {
  char buf[320];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  set_global_error(buf);
  return code;
}

extern "C" int m2k_scan_flags_noalloc(int device, void* stream, const uint8_t* d_bytes, const uint64_t* d_file_off,
                                      const uint64_t* d_file_len, const uint32_t* d_chunk_first, uint32_t n_files,
                                      uint32_t n_chunks, uint8_t* d_file_flags) {
  if (device < 0 || device >= kMaxDevices || !d_bytes || !d_file_off || !d_file_len || !d_chunk_first || !d_file_flags)
    return noalloc_fail(M2K_ERR_ARG, "m2k_scan_flags_noalloc: bad argument");
  if (n_chunks > kNoAllocMaxChunks)
    return noalloc_fail(M2K_ERR_LIMIT,
                        "m2k_scan_flags_noalloc: %u chunks of 64 KiB exceed the built-in record table (%u chunks, 16 GiB of "
                        "content): split the input or use m2k_scan", n_chunks, kNoAllocMaxChunks);
  // the caller is in the middle of its own CUDA work: its current device is put back on every way out
  struct DeviceGuard {
    int prev = -1;
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
  } guard;
  int cur = -1;
  cudaError_t e = cudaGetDevice(&cur);
  if (e != cudaSuccess) return noalloc_fail(M2K_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
  if (cur != device) {
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return noalloc_fail(M2K_ERR_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    guard.prev = cur;
  }
  {
    // set-up, once per device, with that device current
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_records_ptr[device]) {
      ChunkRec* ptr = nullptr;
      e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_records);
      if (e != cudaSuccess || !ptr)
        return noalloc_fail(M2K_ERR_CUDA, "cudaGetSymbolAddress(record table) on device %d: %s", device, cudaGetErrorString(e));
      uint32_t* tptr = nullptr;
      e = cudaGetSymbolAddress((void**)&tptr, g_noalloc_ticket);
      if (e != cudaSuccess || !tptr)
        return noalloc_fail(M2K_ERR_CUDA, "cudaGetSymbolAddress(chunk counter) on device %d: %s", device, cudaGetErrorString(e));
      g_ticket_ptr[device] = tptr;
      int sms = 0;
      e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
      if (e != cudaSuccess || sms <= 0)
        return noalloc_fail(M2K_ERR_CUDA, "cudaDeviceGetAttribute(SM count) on device %d: %s", device, cudaGetErrorString(e));
      g_sm_count[device] = sms;
      g_records_ptr[device] = ptr;
    }
  }
  if (n_files == 0) return M2K_OK;
  // the existing scan kernel, pointed at the caller's buffers and at the built-in record table
  Ctx c;
  c.device = device;
  c.stream = (cudaStream_t)stream;
  c.sm_count = g_sm_count[device];
  DevTree t;
  t.n_files = n_files;
  t.file_begin = 0;
  t.file_end = n_files;
  t.n_chunks = n_chunks;
  t.chunk_first = const_cast<uint32_t*>(d_chunk_first);
  t.file_off = const_cast<uint64_t*>(d_file_off);
  t.file_len = const_cast<uint64_t*>(d_file_len);
  t.bytes = const_cast<uint8_t*>(d_bytes);
  t.shard_byte0 = 0;
  t.shard_bytes = ~0ull >> 1;  // the caller's buffer: its size is not known here (checked builds compare against this)
  e = cudaMemsetAsync(g_ticket_ptr[device], 0, sizeof(uint32_t), c.stream);  // a memset node, not an allocation
  if (e != cudaSuccess) return noalloc_fail(M2K_ERR_CUDA, "m2k_scan_flags_noalloc: clearing the chunk counter: %s", cudaGetErrorString(e));
  launch_scan_bytes(&c, &t, g_records_ptr[device], g_ticket_ptr[device]);
  k_noalloc_file_flags<<<(n_files + 127) / 128, 128, 0, c.stream>>>(n_files, n_chunks, d_chunk_first, d_file_len, d_file_flags);
  e = cudaGetLastError();
  if (e != cudaSuccess) return noalloc_fail(M2K_ERR_CUDA, "m2k_scan_flags_noalloc: kernel launch: %s", cudaGetErrorString(e));
  return M2K_OK;
}
