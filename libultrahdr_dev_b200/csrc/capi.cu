// capi.cu — the C ABI of include/uhdr_b200.h: argument validation in the reference's order
// (lib/src/ultrahdr.cpp:189-202, 245-247, 364-406, 518-523; lib/src/jpegr.cpp:1134-1140), plane
// staging for the host-pointer entry points, and kernel launches on CUDA streams (which replace
// the reference's JobQueue + <=4 worker threads, ultrahdr.cpp:131-183, 304-354, 500-513).
//
// There is NO CPU fallback: without a usable CUDA device every compute entry point returns
// UHDRB200_ERROR_INSUFFICIENT_RESOURCE and uhdrb200_last_error() says why.
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <thread>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#if defined(__x86_64__) || defined(_M_X64)
#include <emmintrin.h>
#endif
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/uhdr_b200.h"
#include "../../include/uhdr_b200_test.h"
#include "uhdr_kernels.cuh"
#include "uhdr_math.cuh"

namespace {
using namespace uhdr;

std::atomic<uint64_t> g_launches{0};
std::atomic<int> g_sync_mode{-1};  // -1: from the environment (UHDRB200_BLOCKING_SYNC), 0 spin, 1 blocking
// test hook (include/uhdr_b200_test.h), per calling thread: 0 auto, 1 analytic kernel, 2 two-table kernel, 3 the TMA-fed
// pipeline variant of the q-domain kernel
thread_local int t_apply_mode = 0;
// numeric mode of the calling thread (include/uhdr_b200.h): UHDRB200_MODE_ANALYTIC (the reference as built) or
// UHDRB200_MODE_LUT (the reference's -DUSE_*_LUT=1 build, jpegr.cpp:33-38)
thread_local int t_numeric_mode = UHDRB200_MODE_ANALYTIC;
// programmatic dependent launch of the generate / apply kernels (uhdr_io.cuh); UHDRB200_NO_PDL=1 turns it off
bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("UHDRB200_NO_PDL");
    return !(e && e[0] == '1');
  }();
  return on;
}

// ---------------------------------------------------------------- per-thread context
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
};
constexpr int kRing = 16;
struct BatchSlot {
  Planes* host = nullptr;  // pinned
  Planes* dev = nullptr;
  int cap = 0;
  cudaEvent_t done = nullptr;
};
struct Ctx {
  int device = -1;  // -1: not bound yet
  bool ready = false;
  cudaStream_t stream = nullptr;
  cudaEvent_t sync_ev = nullptr;  // cudaEventBlockingSync: the joining host thread sleeps instead of spinning
  // page-locked staging chunks for callers that pass pageable planes (what an unmodified libultrahdr caller does)
  static constexpr int kSlots = 4;
  static constexpr size_t kChunk = size_t(4) << 20;
  uint8_t* pin[kSlots] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t pin_ev[kSlots] = {nullptr, nullptr, nullptr, nullptr};
  bool pin_busy[kSlots] = {false, false, false, false};
  int pin_next = 0;
  DevBuf buf[8];  // staging planes of the *_host entry points
  BatchSlot ring[kRing];
  int ring_next = 0;
  int ring_device = -1;  // device the ring slots were allocated on (the *_batch_dev entry points follow the caller's device)
  std::string err;
  void release_ring();
  void release();
  // A host thread that used the library owns a stream, events, page-locked chunks and device staging buffers; they go
  // away with the thread (or with uhdrb200_shutdown). At process exit the CUDA runtime may already be unloading: every
  // call below then fails harmlessly and is ignored.
  ~Ctx() { release(); }
};
thread_local Ctx t_ctx;

void Ctx::release_ring() {
  for (auto& s : ring) {
    if (s.done) cudaEventSynchronize(s.done);
    if (s.host) cudaFreeHost(s.host);
    if (s.dev) cudaFree(s.dev);
    if (s.done) cudaEventDestroy(s.done);
    s = BatchSlot();
  }
  ring_next = 0;
  ring_device = -1;
  cudaGetLastError();
}
void Ctx::release() {
  int cur = -1;
  const bool have_cur = cudaGetDevice(&cur) == cudaSuccess;
  if (ring_device >= 0 && cudaSetDevice(ring_device) == cudaSuccess) release_ring();
  if (ready && cudaSetDevice(device) == cudaSuccess) {
    if (stream) cudaStreamSynchronize(stream);
    for (auto& b : buf) {
      if (b.p) cudaFree(b.p);
      b = DevBuf();
    }
    for (int i = 0; i < kSlots; i++) {
      if (pin[i]) cudaFreeHost(pin[i]);
      if (pin_ev[i]) cudaEventDestroy(pin_ev[i]);
      pin[i] = nullptr;
      pin_ev[i] = nullptr;
      pin_busy[i] = false;
    }
    pin_next = 0;
    if (stream) cudaStreamDestroy(stream);
    if (sync_ev) cudaEventDestroy(sync_ev);
  }
  stream = nullptr;
  sync_ev = nullptr;
  ready = false;
  if (have_cur && cur >= 0) cudaSetDevice(cur);
  cudaGetLastError();
}

int fail(cudaError_t e, const char* what) {
  char msg[256];
  snprintf(msg, sizeof msg, "%s: %s", what, cudaGetErrorString(e));
  t_ctx.err = msg;
  cudaGetLastError();  // clear the sticky-less error state
  return UHDRB200_ERROR_INSUFFICIENT_RESOURCE;
}
#define CU(call)                                   \
  do {                                             \
    cudaError_t e__ = (call);                      \
    if (e__ != cudaSuccess) return fail(e__, #call); \
  } while (0)

int ensure_ctx() {
  Ctx& c = t_ctx;
  if (c.ready) {
    CU(cudaSetDevice(c.device));
    return 0;
  }
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    t_ctx.err = std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    cudaGetLastError();
    return UHDRB200_ERROR_INSUFFICIENT_RESOURCE;
  }
  if (c.device < 0) {
    int cur = 0;
    if (cudaGetDevice(&cur) != cudaSuccess) cur = 0;
    c.device = cur;
  }
  CU(cudaSetDevice(c.device));
  CU(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
  CU(cudaEventCreateWithFlags(&c.sync_ev, cudaEventBlockingSync | cudaEventDisableTiming));
  c.ready = true;
  return 0;
}

// The *_dev and *_batch_dev entry points run on the caller's stream and device pointers: they follow the caller's CURRENT
// device (never switch it) and key the cached IDW / transfer-function tables and the descriptor ring on it. The *_host
// entry points bind the thread to its uhdrb200_init device first (ensure_ctx), so both agree there.
int core_device(int* dev) {
  cudaError_t e = cudaGetDevice(dev);
  if (e != cudaSuccess) {
    t_ctx.err = std::string("no CUDA device: ") + cudaGetErrorString(e);
    cudaGetLastError();
    return UHDRB200_ERROR_INSUFFICIENT_RESOURCE;
  }
  return 0;
}

// The join of a synchronous *_host entry point (the reference joins its worker threads, ultrahdr.cpp:354, 513).
// Spinning (cudaStreamSynchronize) has the lowest latency; when more host threads call into the library than there are
// cores (several processes x several threads per GPU) the spinners steal the cores from each other, so the caller can
// ask for a sleeping wait on a cudaEventBlockingSync event (uhdrb200_set_sync_mode, or UHDRB200_BLOCKING_SYNC=1).
cudaError_t join_stream(Ctx& c) {
  int mode = g_sync_mode.load();
  if (mode < 0) {
    const char* e = getenv("UHDRB200_BLOCKING_SYNC");
    mode = (e && e[0] == '1') ? 1 : 0;
    g_sync_mode = mode;
  }
  if (mode == 0) return cudaStreamSynchronize(c.stream);
  cudaError_t e = cudaEventRecord(c.sync_ev, c.stream);
  if (e != cudaSuccess) return e;
  return cudaEventSynchronize(c.sync_ev);
}

// ---- host <-> device copies of the *_host entry points -------------------------------------------------------------
// Page-locked caller memory (uhdrb200_host_alloc, cudaHostRegister) is handed to the copy engine as it is. Pageable
// memory (plain malloc / new, what JpegR passes) would make cudaMemcpyAsync stage through the driver's single bounce
// buffer under a lock shared by all threads (measured 11 GB/s for 4 calling threads); instead every calling thread
// owns kSlots page-locked chunks and overlaps its own memcpy into / out of them with the DMA of the previous chunk.
bool is_pageable(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return true;
  }
  return a.type == cudaMemoryTypeUnregistered;
}
int ensure_pins(Ctx& c) {
  if (c.pin[0]) return 0;
  for (int i = 0; i < Ctx::kSlots; i++) {
    CU(cudaHostAlloc((void**)&c.pin[i], Ctx::kChunk, cudaHostAllocDefault));
    CU(cudaEventCreateWithFlags(&c.pin_ev[i], cudaEventDisableTiming));
    c.pin_busy[i] = false;
  }
  return 0;
}
// Copy into a page-locked chunk that only the copy engine will read: streaming (non-temporal) stores keep the chunk out
// of the cache hierarchy and skip the read-for-ownership of a plain memcpy (SSE2: baseline x86-64; memcpy elsewhere).
#ifndef UHDR_STAGE_NT
#define UHDR_STAGE_NT 1
#endif
void stage_copy(uint8_t* dst, const uint8_t* src, size_t n) {
#if UHDR_STAGE_NT && (defined(__x86_64__) || defined(_M_X64))
  size_t i = 0;  // dst is page-aligned (cudaHostAlloc)
  for (; i + 64 <= n; i += 64) {
    const __m128i a = _mm_loadu_si128((const __m128i*)(src + i)), b = _mm_loadu_si128((const __m128i*)(src + i + 16));
    const __m128i c = _mm_loadu_si128((const __m128i*)(src + i + 32)), d = _mm_loadu_si128((const __m128i*)(src + i + 48));
    _mm_stream_si128((__m128i*)(dst + i), a);
    _mm_stream_si128((__m128i*)(dst + i + 16), b);
    _mm_stream_si128((__m128i*)(dst + i + 32), c);
    _mm_stream_si128((__m128i*)(dst + i + 48), d);
  }
  _mm_sfence();
  if (i < n) memcpy(dst + i, src + i, n - i);
#else
  memcpy(dst, src, n);
#endif
}
// The reference spends up to four host threads on one call (ultrahdr.cpp:304, 500); for a caller with pageable planes the
// host-side cost here is the copy into / out of the page-locked chunks, so that copy gets the same budget: a chunk is cut
// into slices which the calling thread and up to three helpers of a process-wide pool move concurrently (one core's copy
// rate, ~10 GB/s, is a fifth of what the PCIe link takes). UHDRB200_STAGE_THREADS sets the pool size (0 = caller only).
class StagePool {
 public:
  static StagePool& get() {
    static StagePool* p = new StagePool();  // never destroyed: its threads sleep on the queue until the process exits
    return *p;
  }
  // copies n bytes; `nt` = streaming stores (dst 16-byte aligned, only read by the copy engine afterwards)
  void copy(uint8_t* dst, const uint8_t* src, size_t n, bool nt) {
    const size_t kSlice = size_t(1) << 20;
    if (workers_ == 0 || n < 2 * kSlice) {
      move(dst, src, n, nt);
      return;
    }
    const int parts = (int)((n + kSlice - 1) / kSlice < (size_t)kPerCall ? (n + kSlice - 1) / kSlice : (size_t)kPerCall);
    const size_t per = ((n / parts) + 63) & ~size_t(63);
    std::atomic<int> pending{parts - 1};
    {
      std::lock_guard<std::mutex> lk(mu_);
      for (int i = 1; i < parts; i++) {
        const size_t off = per * i, len = i == parts - 1 ? n - off : per;
        q_.push_back(Task{dst + off, src + off, len, nt, &pending});
      }
    }
    cv_.notify_all();
    move(dst, src, per, nt);
    Task t;
    while (pending.load(std::memory_order_acquire) > 0) {  // help with whatever is queued (mine or another caller's), then wait
      if (pop(&t)) run(t);
      else std::this_thread::yield();
    }
  }

 private:
  static constexpr int kPerCall = 4;
  struct Task {
    uint8_t* dst;
    const uint8_t* src;
    size_t n;
    bool nt;
    std::atomic<int>* pending;
  };
  StagePool() {
    int n = (int)std::thread::hardware_concurrency() / 2;
    n = n < 2 ? 2 : (n > 12 ? 12 : n);
    if (const char* e = getenv("UHDRB200_STAGE_THREADS")) n = atoi(e);
    workers_ = n < 0 ? 0 : n;
    for (int i = 0; i < workers_; i++) std::thread([this] { loop(); }).detach();
  }
  static void move(uint8_t* dst, const uint8_t* src, size_t n, bool nt) {
    if (nt) stage_copy(dst, src, n);
    else memcpy(dst, src, n);
  }
  static void run(const Task& t) {
    move(t.dst, t.src, t.n, t.nt);
    t.pending->fetch_sub(1, std::memory_order_release);
  }
  bool pop(Task* t) {
    std::lock_guard<std::mutex> lk(mu_);
    if (q_.empty()) return false;
    *t = q_.front();
    q_.pop_front();
    return true;
  }
  void loop() {
    for (;;) {
      Task t;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [this] { return !q_.empty(); });
        t = q_.front();
        q_.pop_front();
      }
      run(t);
    }
  }
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<Task> q_;
  int workers_ = 0;
};

int copy_h2d(Ctx& c, void* dst, const void* src, size_t bytes) {
  if (!bytes) return 0;
  if (!is_pageable(src)) {
    CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c.stream));
    return 0;
  }
  int st = ensure_pins(c);
  if (st) return st;
  for (size_t off = 0; off < bytes; off += Ctx::kChunk) {
    const size_t n = bytes - off < Ctx::kChunk ? bytes - off : Ctx::kChunk;
    const int k = c.pin_next;
    c.pin_next = (k + 1) % Ctx::kSlots;
    if (c.pin_busy[k]) CU(cudaEventSynchronize(c.pin_ev[k]));  // the chunk's previous DMA has drained
    StagePool::get().copy(c.pin[k], (const uint8_t*)src + off, n, true);
    CU(cudaMemcpyAsync((uint8_t*)dst + off, c.pin[k], n, cudaMemcpyHostToDevice, c.stream));
    CU(cudaEventRecord(c.pin_ev[k], c.stream));
    c.pin_busy[k] = true;
  }
  return 0;
}
// On return from the pageable path the bytes are in dst; the page-locked path only enqueues (the caller joins the stream).
int copy_d2h(Ctx& c, void* dst, const void* src, size_t bytes) {
  if (!bytes) return 0;
  if (!is_pageable(dst)) {
    CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c.stream));
    return 0;
  }
  int st = ensure_pins(c);
  if (st) return st;
  const size_t nchunks = (bytes + Ctx::kChunk - 1) / Ctx::kChunk;
  size_t issued = 0, done = 0;
  int slot_of[Ctx::kSlots];
  while (done < nchunks) {
    while (issued < nchunks && issued - done < (size_t)Ctx::kSlots) {
      const int k = c.pin_next;
      c.pin_next = (k + 1) % Ctx::kSlots;
      if (c.pin_busy[k]) CU(cudaEventSynchronize(c.pin_ev[k]));
      const size_t off = issued * Ctx::kChunk, n = bytes - off < Ctx::kChunk ? bytes - off : Ctx::kChunk;
      CU(cudaMemcpyAsync(c.pin[k], (const uint8_t*)src + off, n, cudaMemcpyDeviceToHost, c.stream));
      CU(cudaEventRecord(c.pin_ev[k], c.stream));
      c.pin_busy[k] = true;
      slot_of[issued % Ctx::kSlots] = k;
      issued++;
    }
    const int k = slot_of[done % Ctx::kSlots];
    const size_t off = done * Ctx::kChunk, n = bytes - off < Ctx::kChunk ? bytes - off : Ctx::kChunk;
    CU(cudaEventSynchronize(c.pin_ev[k]));
    StagePool::get().copy((uint8_t*)dst + off, c.pin[k], n, false);
    c.pin_busy[k] = false;
    done++;
  }
  return 0;
}

int reserve(DevBuf& b, size_t bytes) {
  if (bytes <= b.cap) return 0;
  if (b.p) cudaFree(b.p);
  b.p = nullptr;
  b.cap = 0;
  size_t cap = (bytes + (size_t(1) << 20)) & ~((size_t(1) << 20) - 1);
  CU(cudaMalloc(&b.p, cap));
  b.cap = cap;
  return 0;
}

// ---------------------------------------------------------------- IDW tables (gainmapmath.h:184-228, gainmapmath.cpp:69-110)
// Built on the host with the reference's float arithmetic, cached on the device per (device, scale).
float idw_dist(float x1, float x2, float y1, float y2) {
  return sqrtf(((y2 - y1) * (y2 - y1)) + (x2 - x1) * (x2 - x1));
}
void idw_fill(float* w, int s, int inc_r, int inc_b) {
  for (int y = 0; y < s; y++)
    for (int x = 0; x < s; x++) {
      float px = ((float)x) / s, py = ((float)y) / s;
      int cx = (int)floorf(px), cy = (int)floorf(py);
      int nx = cx + inc_r, ny = cy + inc_b;
      float* o = w + (y * s + x) * 4;
      float d1 = idw_dist(px, (float)cx, py, (float)cy);
      if (d1 == 0) {
        o[0] = 1.f; o[1] = 0.f; o[2] = 0.f; o[3] = 0.f;
        continue;
      }
      float w1 = 1.f / d1;
      float w2 = 1.f / idw_dist(px, (float)cx, py, (float)ny);
      float w3 = 1.f / idw_dist(px, (float)nx, py, (float)cy);
      float w4 = 1.f / idw_dist(px, (float)nx, py, (float)ny);
      float tot = w1 + w2 + w3 + w4;
      o[0] = w1 / tot; o[1] = w2 / tot; o[2] = w3 / tot; o[3] = w4 / tot;
    }
}
std::mutex g_idw_mu;
std::map<std::pair<int, int>, float*> g_idw;  // (device, scale) -> device tables [4][s][s][4]

int get_idw(int device, int scale, const float** out) {
  std::lock_guard<std::mutex> lk(g_idw_mu);
  auto it = g_idw.find({device, scale});
  if (it != g_idw.end()) {
    *out = it->second;
    return 0;
  }
  const size_t n = (size_t)scale * scale * 4;
  std::vector<float> h(4 * n);
  idw_fill(h.data() + 0 * n, scale, 1, 1);  // mWeights
  idw_fill(h.data() + 1 * n, scale, 0, 1);  // mWeightsNR
  idw_fill(h.data() + 2 * n, scale, 1, 0);  // mWeightsNB
  idw_fill(h.data() + 3 * n, scale, 0, 0);  // mWeightsC
  float* d = nullptr;
  CU(cudaMalloc(&d, h.size() * sizeof(float)));
  CU(cudaMemcpy(d, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice));
  g_idw[{device, scale}] = d;
  *out = d;
  return 0;
}

// ---------------------------------------------------------------- transfer-function tables of the fast apply path
std::map<int, ApplyTables> g_tables;  // device -> device-resident tables (guarded by g_idw_mu)

int get_apply_tables(int device, ApplyTables* out) {
  std::lock_guard<std::mutex> lk(g_idw_mu);
  auto it = g_tables.find(device);
  if (it != g_tables.end()) {
    *out = it->second;
    return 0;
  }
  const size_t n2 = (size_t)kSEntries + (size_t)kSEntriesHi + 2 * (size_t)kOEntries + 2 * (size_t)kPEntries + (size_t)kLEntries;
  std::vector<float2> h2(n2);
  std::vector<float> e255(256);
  float2* hs = h2.data();
  float2* hs_hi = hs + kSEntries;
  float2* ho_hlg = hs_hi + kSEntriesHi;
  float2* ho_pq = ho_hlg + kOEntries;
  float2* hp_hlg = ho_pq + kOEntries;
  float2* hp_pq = hp_hlg + kPEntries;
  float2* hl_pow = hp_pq + kPEntries;
  build_apply_tables(hs, hs_hi, ho_hlg, ho_pq, hp_hlg, hp_pq, hl_pow, e255.data());
  char* d = nullptr;
  CU(cudaMalloc((void**)&d, n2 * sizeof(float2) + 256 * sizeof(float)));
  CU(cudaMemcpy(d, h2.data(), n2 * sizeof(float2), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(d + n2 * sizeof(float2), e255.data(), 256 * sizeof(float), cudaMemcpyHostToDevice));
  ApplyTables t;
  t.s = (const float2*)d;
  t.s_hi = t.s + kSEntries;
  t.o_hlg = t.s_hi + kSEntriesHi;
  t.o_pq = t.o_hlg + kOEntries;
  t.p_hlg = t.o_pq + kOEntries;
  t.p_pq = t.p_hlg + kPEntries;
  t.l_pow = t.p_pq + kPEntries;
  t.e255 = (const float*)(d + n2 * sizeof(float2));
  g_tables[device] = t;
  *out = t;
  return 0;
}

// ---------------------------------------------------------------- LUT numeric mode: device copies of the reference's tables
std::map<int, LutTables> g_lut_tables;  // device -> the five transfer-function tables (guarded by g_idw_mu)

int get_lut_tables(int device, LutTables* out) {
  std::lock_guard<std::mutex> lk(g_idw_mu);
  auto it = g_lut_tables.find(device);
  if (it != g_lut_tables.end()) {
    *out = it->second;
    return 0;
  }
  const size_t n = (size_t)kLutSrgbInv + kLutHlgInv + kLutPqInv + kLutHlgOetf + kLutPqOetf;
  std::vector<float> h(n);
  float* h_srgb = h.data();
  float* h_hlg_inv = h_srgb + kLutSrgbInv;
  float* h_pq_inv = h_hlg_inv + kLutHlgInv;
  float* h_hlg = h_pq_inv + kLutPqInv;
  float* h_pq = h_hlg + kLutHlgOetf;
  build_lut_tables(h_srgb, h_hlg_inv, h_pq_inv, h_hlg, h_pq);
  float* d = nullptr;
  CU(cudaMalloc((void**)&d, n * sizeof(float)));
  CU(cudaMemcpy(d, h.data(), n * sizeof(float), cudaMemcpyHostToDevice));
  LutTables t;
  t.srgb_inv = d;
  t.hlg_inv = t.srgb_inv + kLutSrgbInv;
  t.pq_inv = t.hlg_inv + kLutHlgInv;
  t.hlg_oetf = t.pq_inv + kLutPqInv;
  t.pq_oetf = t.hlg_oetf + kLutHlgOetf;
  g_lut_tables[device] = t;
  *out = t;
  return 0;
}

// GainLUT (gainmapmath.h:150-183) depends on the metadata and the display boost of the call: one 4 KB device table per
// distinct (device, minContentBoost, maxContentBoost, displayBoost), built once with the reference's own host arithmetic
// and kept (a decoder sees a handful of distinct boosts). The cache is bounded: beyond kGainLutMax tables the device is
// drained and the tables are dropped before the new one is built, so no kernel in flight can lose its table.
struct GainLutKey {
  int device;
  uint32_t mn, mx, db;  // the floats' bit patterns
  bool operator<(const GainLutKey& o) const { return std::tie(device, mn, mx, db) < std::tie(o.device, o.mn, o.mx, o.db); }
};
constexpr size_t kGainLutMax = 256;
std::map<GainLutKey, float*> g_gain_luts;  // guarded by g_idw_mu

int get_gain_lut(int device, float mn, float mx, float db, const float** out) {
  GainLutKey key;
  key.device = device;
  memcpy(&key.mn, &mn, 4);
  memcpy(&key.mx, &mx, 4);
  memcpy(&key.db, &db, 4);
  std::lock_guard<std::mutex> lk(g_idw_mu);
  auto it = g_gain_luts.find(key);
  if (it != g_gain_luts.end()) {
    *out = it->second;
    return 0;
  }
  if (g_gain_luts.size() >= kGainLutMax) {
    int cur = 0;
    CU(cudaGetDevice(&cur));
    for (auto& kv : g_gain_luts) {
      if (cudaSetDevice(kv.first.device) == cudaSuccess) {
        cudaDeviceSynchronize();
        cudaFree(kv.second);
      }
    }
    g_gain_luts.clear();
    CU(cudaSetDevice(cur));
  }
  float h[kLutGain];
  build_gain_lut(h, mn, mx, db);
  float* d = nullptr;
  CU(cudaMalloc((void**)&d, sizeof h));
  CU(cudaMemcpy(d, h, sizeof h, cudaMemcpyHostToDevice));
  g_gain_luts[key] = d;
  *out = d;
  return 0;
}

// ---------------------------------------------------------------- saturated-white pixels under a saturated map
// The reference's own float / double chain for a channel with c == 1 in an interior scale-4 cell whose four map samples
// are 255: sampleMap (gainmapmath.cpp:700-719, e = 255 / 255.0f = 1, left-to-right float sum of the four weights),
// applyGain (:550-555, double log2 / exp2 narrowed to float), `/ display_boost` (ultrahdr.cpp:451), srgbInvOetf(1)
// (:149-155), hlgOetf / pqOetf (:259-267, 309-314, double) and the truncating `* 1023.0f` (:722-727).
// The 64 evaluations cost ~5 us of libm per call, a third of a single 4K image's kernel time: the last few (metadata,
// display boost, format) combinations of the calling thread are kept (UHDRB200_NO_SATCACHE=1 turns the cache off).
struct SatCellEntry {
  uint32_t mn, mx, db;
  int fmt;
  bool valid;
  float vsat[4][16], rsat[4][16];
};
thread_local SatCellEntry t_sat_cache[4];
thread_local int t_sat_next = 0;

void fill_sat_cell_uncached(ApplyFastParams& f, float mn, float mx, float db, int fmt);
void fill_sat_cell(ApplyFastParams& f, float mn, float mx, float db, int fmt) {
  static const bool off = [] {
    const char* e = getenv("UHDRB200_NO_SATCACHE");
    return e && e[0] == '1';
  }();
  if (off) return fill_sat_cell_uncached(f, mn, mx, db, fmt);
  uint32_t k[3];
  memcpy(&k[0], &mn, 4);
  memcpy(&k[1], &mx, 4);
  memcpy(&k[2], &db, 4);
  for (auto& c : t_sat_cache)
    if (c.valid && c.mn == k[0] && c.mx == k[1] && c.db == k[2] && c.fmt == fmt) {
      memcpy(f.vsat, c.vsat, sizeof f.vsat);
      memcpy(f.rsat, c.rsat, sizeof f.rsat);
      return;
    }
  fill_sat_cell_uncached(f, mn, mx, db, fmt);
  SatCellEntry& c = t_sat_cache[t_sat_next];
  t_sat_next = (t_sat_next + 1) % 4;
  c.mn = k[0]; c.mx = k[1]; c.db = k[2];
  c.fmt = fmt;
  memcpy(c.vsat, f.vsat, sizeof f.vsat);
  memcpy(c.rsat, f.rsat, sizeof f.rsat);
  c.valid = true;
}
void fill_sat_cell_uncached(ApplyFastParams& f, float mn, float mx, float db, int fmt) {
  const float lin1 = (float)pow((double)((1.0f + 0.055f) / 1.055f), 2.4);
  const int inc[4][2] = {{1, 1}, {0, 1}, {1, 0}, {0, 0}};  // mWeights, NR, NB, C (gainmapmath.h:184-228)
  for (int t = 0; t < 4; t++) {
   float w[16][4];
   idw_fill(&w[0][0], 4, inc[t][0], inc[t][1]);
   for (int i = 0; i < 16; i++) {
    const float e = 255.0f / 255.0f;
    const float gain = e * w[i][0] + e * w[i][1] + e * w[i][2] + e * w[i][3];
    const float log_boost = (float)(log2((double)mn) * (double)(1.0f - gain) + log2((double)mx) * (double)gain);
    const float gf = (float)exp2((double)(log_boost * db / mx));
    const float r = (lin1 * gf) / db;
    f.rsat[t][i] = r;
    float o = r;
    if (fmt == UHDRB200_OUTPUT_HDR_HLG) {
      const float A = 0.17883277f, B = 0.28466892f, C = (float)0.55991073;
      o = r <= 1.0f / 12.0f ? sqrtf(3.0f * r) : (float)((double)A * log((double)(12.0f * r - B)) + (double)C);
    } else if (fmt == UHDRB200_OUTPUT_HDR_PQ) {
      const float m1 = 2610.0f / 16384.0f, m2 = 2523.0f / 4096.0f * 128.0f;
      const float c1 = 3424.0f / 4096.0f, c2 = 2413.0f / 4096.0f * 32.0f, c3 = 2392.0f / 4096.0f * 32.0f;
      const double pw = pow((double)r, (double)m1);
      o = r <= 0.0f ? 0.0f : (float)pow(((double)c1 + (double)c2 * pw) / (1 + (double)c3 * pw), (double)m2);
    }
    f.vsat[t][i] = (float)(uint32_t)(o * 1023.0f) + 0.5f;
   }
  }
}

// ---------------------------------------------------------------- batch descriptor staging
// A launch over n>1 images reads its per-image plane pointers from a small device array. Slots
// come from a ring guarded by events so an async caller can have kRing launches in flight.
int stage_batch(const std::vector<Planes>& v, cudaStream_t s, const Planes** dev, BatchSlot** slot_out) {
  Ctx& c = t_ctx;
  int cur = 0;
  CU(cudaGetDevice(&cur));
  if (c.ring_device != cur) {  // first use, or the calling thread moved to another GPU: the slots belong to one device
    if (c.ring_device >= 0) {
      CU(cudaSetDevice(c.ring_device));
      c.release_ring();
      CU(cudaSetDevice(cur));
    }
    c.ring_device = cur;
  }
  BatchSlot& sl = c.ring[c.ring_next];
  c.ring_next = (c.ring_next + 1) % kRing;
  if (sl.done) CU(cudaEventSynchronize(sl.done));
  else CU(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
  if ((int)v.size() > sl.cap) {
    if (sl.host) cudaFreeHost(sl.host);
    if (sl.dev) cudaFree(sl.dev);
    sl.host = nullptr; sl.dev = nullptr; sl.cap = 0;
    int cap = 64;
    while (cap < (int)v.size()) cap *= 2;
    CU(cudaHostAlloc((void**)&sl.host, sizeof(Planes) * cap, cudaHostAllocDefault));
    CU(cudaMalloc((void**)&sl.dev, sizeof(Planes) * cap));
    sl.cap = cap;
  }
  memcpy(sl.host, v.data(), sizeof(Planes) * v.size());
  CU(cudaMemcpyAsync(sl.dev, sl.host, sizeof(Planes) * v.size(), cudaMemcpyHostToDevice, s));
  *dev = sl.dev;
  *slot_out = &sl;
  return 0;
}

bool same_layout(const uhdrb200_image& a, const uhdrb200_image& b) {
  return a.width == b.width && a.height == b.height && a.luma_stride == b.luma_stride &&
         a.chroma_stride == b.chroma_stride && a.colorGamut == b.colorGamut;
}

}  // namespace

// ================================================================ validation (no CUDA involved)
namespace {

// ultrahdr.cpp:189-202, then :222-248 for the transfer function, :267-302 for gamut ranges.
int validate_generate(const uhdrb200_image* sdr, const uhdrb200_image* hdr, int hdr_tf,
                      const uhdrb200_metadata* md, const uhdrb200_image* dest) {
  if (!sdr || !hdr || !md || !dest || !sdr->data || !sdr->chroma_data || !hdr->data || !hdr->chroma_data)
    return UHDRB200_ERROR_BAD_PTR;
  if (sdr->width != hdr->width || sdr->height != hdr->height) return UHDRB200_ERROR_RESOLUTION_MISMATCH;
  if (sdr->colorGamut == UHDRB200_COLORGAMUT_UNSPECIFIED || hdr->colorGamut == UHDRB200_COLORGAMUT_UNSPECIFIED)
    return UHDRB200_ERROR_INVALID_COLORGAMUT;
  if (hdr_tf != UHDRB200_TF_LINEAR && hdr_tf != UHDRB200_TF_HLG && hdr_tf != UHDRB200_TF_PQ)
    return UHDRB200_ERROR_INVALID_TRANS_FUNC;
  if (sdr->colorGamut < 0 || sdr->colorGamut > 2 || hdr->colorGamut < 0 || hdr->colorGamut > 2)
    return UHDRB200_ERROR_INVALID_COLORGAMUT;
  return UHDRB200_NO_ERROR;
}

// ultrahdr.cpp:250-257
void fill_metadata(uhdrb200_metadata* md, int hdr_tf) {
  const float white = hdr_tf == UHDRB200_TF_PQ ? 10000.0f : 1000.0f;  // kPqMaxNits / kHlgMaxNits
  memset(md->version, 0, sizeof md->version);
  memcpy(md->version, "1.0", 3);  // kGainMapVersion
  md->maxContentBoost = white / 203.0f;  // kSdrWhiteNits
  md->minContentBoost = 1.0f;
  md->gamma = 1.0f;
  md->offsetSdr = 0.0f;
  md->offsetHdr = 0.0f;
  md->hdrCapacityMin = 1.0f;
  md->hdrCapacityMax = md->maxContentBoost;
}

// ultrahdr.cpp:209-216 (data pointer is set by the caller of this helper)
void fill_map_desc(uhdrb200_image* dest, size_t mw, size_t mh) {
  dest->width = mw;
  dest->height = mh;
  dest->colorGamut = UHDRB200_COLORGAMUT_UNSPECIFIED;
  dest->luma_stride = mw;
  dest->chroma_data = nullptr;
  dest->chroma_stride = 0;
  dest->pixelFormat = UHDRB200_PIX_FMT_MONOCHROME;
}

// ultrahdr.cpp:364-406
int validate_apply(const uhdrb200_image* sdr, const uhdrb200_image* map, const uhdrb200_metadata* md,
                   const uhdrb200_image* dest) {
  if (!sdr || !map || !md || !dest || !sdr->data || !sdr->chroma_data || !map->data) return UHDRB200_ERROR_BAD_PTR;
  if (strncmp(md->version, "1.0", sizeof md->version) != 0) return UHDRB200_ERROR_BAD_METADATA;
  if (md->gamma != 1.0f) return UHDRB200_ERROR_BAD_METADATA;
  if (md->offsetSdr != 0.0f || md->offsetHdr != 0.0f) return UHDRB200_ERROR_BAD_METADATA;
  if (md->hdrCapacityMin != md->minContentBoost || md->hdrCapacityMax != md->maxContentBoost)
    return UHDRB200_ERROR_BAD_METADATA;
  if (map->width == 0 || map->height == 0) return UHDRB200_ERROR_UNSUPPORTED_MAP_SCALE_FACTOR;  // the reference would divide by zero
  if (sdr->width % map->width != 0 || sdr->height % map->height != 0) return UHDRB200_ERROR_UNSUPPORTED_MAP_SCALE_FACTOR;
  if (sdr->width * map->height != sdr->height * map->width) return UHDRB200_ERROR_UNSUPPORTED_MAP_SCALE_FACTOR;
  return UHDRB200_NO_ERROR;
}

// ultrahdr.cpp:518-523 (+ null plane pointers, which the reference would dereference)
int validate_tonemap(const uhdrb200_image* src, const uhdrb200_image* dest) {
  if (!src || !dest) return UHDRB200_ERROR_BAD_PTR;
  if (src->width != dest->width || src->height != dest->height) return UHDRB200_ERROR_RESOLUTION_MISMATCH;
  if (!src->data || !src->chroma_data || !dest->data || !dest->chroma_data) return UHDRB200_ERROR_BAD_PTR;
  return UHDRB200_NO_ERROR;
}

bool fits_int(size_t v) { return v <= (size_t)0x7fffffff; }

// encodeGain's constants (gainmapmath.cpp:524-541, ultrahdr.cpp:259-260): float-rounded log2 bounds, and the codes of a
// gain clamped to min / max evaluated in double like the reference
GainCode gain_code_of(float min_boost, float max_boost) {
  GainCode k;
  k.min_boost = min_boost;
  k.max_boost = max_boost;
  k.log2_min = (float)log2((double)min_boost);
  const float log2_max = (float)log2((double)max_boost);
  k.log2_max = log2_max;
  const double den = (double)(log2_max - k.log2_min);
  k.code_at_min = (int)(uint8_t)((log2((double)min_boost) - (double)k.log2_min) / den * 255.0);
  k.code_at_max = (int)(uint8_t)((log2((double)max_boost) - (double)k.log2_min) / den * 255.0);
  k.code_scale = (float)(255.0 / den);
  return k;
}

Geometry geometry_of(const uhdrb200_image* sdr, const uhdrb200_image* hdr) {
  Geometry g;
  memset(&g, 0, sizeof g);
  const uhdrb200_image* any = sdr ? sdr : hdr;
  g.width = (int)any->width;
  g.height = (int)any->height;
  if (sdr) {
    g.sdr_luma_stride = (int)sdr->luma_stride;
    g.sdr_chroma_stride = (int)sdr->chroma_stride;
  }
  if (hdr) {
    g.hdr_luma_stride = (int)(hdr->luma_stride == 0 ? hdr->width : hdr->luma_stride);  // gainmapmath.cpp:585
    g.hdr_chroma_stride = (int)hdr->chroma_stride;
  }
  return g;
}
void set_sdr_planes(Planes& pl, const uhdrb200_image& im) {
  pl.sdr_y = im.data;
  pl.sdr_u = im.chroma_data;
  pl.sdr_v = (const uint8_t*)im.chroma_data + im.chroma_stride * (im.height / 2);  // gainmapmath.cpp:568
}

// ================================================================ device-pointer cores
int generate_core(int n, const uhdrb200_image* sdr, const uhdrb200_image* hdr, int hdr_tf,
                  uhdrb200_metadata* md, uhdrb200_image* dest, int sdr_is_601, cudaStream_t s, bool fused = false,
                  const float* conv = nullptr) {
  if (n <= 0) return UHDRB200_NO_ERROR;
  if (!sdr || !hdr || !md || !dest) return UHDRB200_ERROR_BAD_PTR;
  for (int i = 0; i < n; i++) {
    int st = validate_generate(&sdr[i], &hdr[i], hdr_tf, &md[i], &dest[i]);
    if (st) return st;
    if (!dest[i].data) return UHDRB200_ERROR_BAD_PTR;
    if (i && (!same_layout(sdr[i], sdr[0]) || !same_layout(hdr[i], hdr[0]))) return UHDRB200_ERROR_RESOLUTION_MISMATCH;
  }
  if (!fits_int(sdr[0].width) || !fits_int(sdr[0].height) || !fits_int(sdr[0].luma_stride) || !fits_int(hdr[0].luma_stride))
    return UHDRB200_ERROR_UNSUPPORTED_WIDTH_HEIGHT;
  int dev = 0;
  int st = core_device(&dev);
  if (st) return st;
  GenerateParams p;
  memset(&p, 0, sizeof p);
  p.g = geometry_of(&sdr[0], &hdr[0]);
  p.g.map_scale = 4;  // kMapDimensionScaleFactor, ultrahdr.h:213
  p.g.map_width = p.g.width / 4;
  p.g.map_height = p.g.height / 4;
  p.sdr_k = yuv_coefs(sdr_is_601 ? 1 : sdr[0].colorGamut);  // ultrahdr.cpp:267-286
  p.hdr_k = yuv_coefs(hdr[0].colorGamut);                   // ultrahdr.cpp:288-302
  p.luma = luma_coefs(sdr[0].colorGamut);
  p.gamut = gamut_matrix(sdr[0].colorGamut, hdr[0].colorGamut);  // ultrahdr.cpp:262-263
  p.use_gamut = sdr[0].colorGamut != hdr[0].colorGamut;
  p.hdr_tf = hdr_tf;
  p.hdr_white_nits = hdr_tf == UHDRB200_TF_PQ ? 10000.0f : 1000.0f;
  std::vector<Planes> v(n);
  bool vec = true;
  for (int i = 0; i < n; i++) {
    fill_metadata(&md[i], hdr_tf);
    fill_map_desc(&dest[i], (size_t)p.g.map_width, (size_t)p.g.map_height);
    memset(&v[i], 0, sizeof(Planes));
    set_sdr_planes(v[i], sdr[i]);
    v[i].hdr_y = hdr[i].data;
    v[i].hdr_uv = hdr[i].chroma_data;
    v[i].out = dest[i].data;
    vec = vec && generate_vec_ok(p.g, v[i]);
  }
  p.gc = gain_code_of(md[0].minContentBoost, md[0].maxContentBoost);
  p.pdl = pdl_enabled() ? 1 : 0;
  if (t_numeric_mode == UHDRB200_MODE_LUT) {
    p.lut = 1;
    if ((st = get_lut_tables(dev, &p.lt))) return st;
  }
  if (conv) {
    p.convert = 1;
    memcpy(p.conv, conv, sizeof p.conv);
  }
  BatchSlot* slot = nullptr;
  if (n == 1) p.single = v[0];
  else if ((st = stage_batch(v, s, &p.batch, &slot))) return st;
  cudaError_t e = fused ? launch_api0_fused_impl(p, n, s) : launch_generate_impl(p, n, vec, s);
  if (slot) cudaEventRecord(slot->done, s);
  if (e != cudaSuccess) return fail(e, "generate kernel launch");
  g_launches++;
  return UHDRB200_NO_ERROR;
}

int apply_core(int n, const uhdrb200_image* sdr, const uhdrb200_image* map, const uhdrb200_metadata* md,
               int fmt, float max_display_boost, uhdrb200_image* dest, cudaStream_t s) {
  if (n <= 0) return UHDRB200_NO_ERROR;
  if (!sdr || !map || !md || !dest) return UHDRB200_ERROR_BAD_PTR;
  for (int i = 0; i < n; i++) {
    int st = validate_apply(&sdr[i], &map[i], md, &dest[i]);
    if (st) return st;
    if (!dest[i].data) return UHDRB200_ERROR_BAD_PTR;
    if (i && (!same_layout(sdr[i], sdr[0]) || map[i].width != map[0].width || map[i].height != map[0].height))
      return UHDRB200_ERROR_RESOLUTION_MISMATCH;
  }
  if (!fits_int(sdr[0].width) || !fits_int(sdr[0].height) || !fits_int(sdr[0].luma_stride))
    return UHDRB200_ERROR_UNSUPPORTED_WIDTH_HEIGHT;
  for (int i = 0; i < n; i++) {  // ultrahdr.cpp:411-413
    dest[i].width = sdr[i].width;
    dest[i].height = sdr[i].height;
    dest[i].colorGamut = sdr[i].colorGamut;
  }
  if (fmt != UHDRB200_OUTPUT_HDR_LINEAR && fmt != UHDRB200_OUTPUT_HDR_PQ && fmt != UHDRB200_OUTPUT_HDR_HLG &&
      fmt != UHDRB200_OUTPUT_HDR_LINEAR_RGB_10BIT)
    return UHDRB200_NO_ERROR;  // ultrahdr.cpp:491-493: nothing is written, NO_ERROR
  int dev = 0;
  int st = core_device(&dev);
  if (st) return st;
  ApplyParams p;
  memset(&p, 0, sizeof p);
  p.g = geometry_of(&sdr[0], nullptr);
  p.g.map_width = (int)map[0].width;
  p.g.map_height = (int)map[0].height;
  p.g.map_scale = (int)(sdr[0].width / map[0].width);  // ultrahdr.cpp:409
  p.output_format = fmt;
  p.log2_min = (float)log2((double)md->minContentBoost);
  p.log2_max = (float)log2((double)md->maxContentBoost);
  p.max_boost = md->maxContentBoost;
  p.display_boost = max_display_boost < md->maxContentBoost ? max_display_boost : md->maxContentBoost;  // :415
  p.inv_display_boost = 1.0f / p.display_boost;
  if ((st = get_idw(dev, p.g.map_scale, &p.idw))) return st;
  ApplyTables tabs;
  if ((st = get_apply_tables(dev, &tabs))) return st;
  p.e255 = tabs.e255;
  std::vector<Planes> v(n);
  bool vec = true;
  for (int i = 0; i < n; i++) {
    memset(&v[i], 0, sizeof(Planes));
    set_sdr_planes(v[i], sdr[i]);
    v[i].map_in = map[i].data;
    v[i].out = dest[i].data;
    vec = vec && apply_vec_ok(p.g, v[i]);
  }
  // Fast path (apply_fast.cu): map scale 4, 4-aligned planes, and a gain factor
  // f = 2^(gain*A + B) (gain in [0,1]) small enough for the OETF table range.
  const bool lut = t_numeric_mode == UHDRB200_MODE_LUT;  // LUT mode: the per-pixel kernel with the reference's tables
  if (lut) {
    p.lut = 1;
    if ((st = get_lut_tables(dev, &p.lt))) return st;
    if ((st = get_gain_lut(dev, md->minContentBoost, md->maxContentBoost, p.display_boost, &p.gain_lut))) return st;
  }
  bool fast = t_apply_mode != 1 && !lut;
  for (int i = 0; i < n && fast; i++) fast = apply_fast_ok(p.g, v[i]);
  // other integer scale factors (or scale 4 with a ragged map) keep the q-domain tables through apply_qs_kernel
  bool any_scale = !fast && !lut && (t_apply_mode == 0 || t_apply_mode == 3);
  for (int i = 0; i < n && any_scale; i++) any_scale = apply_qs_ok(p.g, v[i]);
  fast = fast || any_scale;
  const double k = (double)p.display_boost / (double)p.max_boost;
  const double A = ((double)p.log2_max - (double)p.log2_min) * k;
  const double B = (double)p.log2_min * k - log2((double)p.display_boost);
  const double fmax = exp2(A > 0 ? A + B : B);
  if (!(fmax < 3.9) || !(p.display_boost > 0.0f) || !std::isfinite(A) || !std::isfinite(B)) fast = false;
  BatchSlot* slot = nullptr;
  const Planes* batch = nullptr;
  if (n > 1 && (st = stage_batch(v, s, &batch, &slot))) return st;
  cudaError_t e;
  if (fast) {
    ApplyFastParams f;
    memset(&f, 0, sizeof f);
    f.g = p.g;
    f.output_format = fmt;
    f.A = (float)A;
    f.B = (float)B;
    f.e0 = (float)ldexp(1.0, kOLoExp);
    f.s_bias = 0x4B000000u << 3;
    f.need_mask = fmax > 1.0005 ? 1 : 0;  // at f <= 1.0005 no 10-bit code can reach 1024
    idw_fill(&f.w[0][0], 4, 1, 1);  // ShepardsIDW::mWeights for scale 4 ...
    fill_sat_cell(f, md->minContentBoost, md->maxContentBoost, p.display_boost, fmt);
    for (int i = 0; i < 16; i++)
      for (int c = 0; c < 4; c++) f.w[i][c] = (float)((double)f.w[i][c] * A);  // ... pre-multiplied by A
    f.Aq = (float)(A / 2.4 / 255.0);  // the q kernel keeps the map samples as 0..255 floats
    f.Bq = (float)(B / 2.4);
    f.q0 = (float)ldexp(1.0, kPLoExp);
    f.two_table = t_apply_mode == 2;
    f.pipe = (t_apply_mode == 3 && !any_scale) ? 1 : 0;
    for (int i = 0; i < n && f.pipe; i++) f.pipe = apply_pipe_ok(p.g, v[i]) ? 1 : 0;
    f.pdl = pdl_enabled() ? 1 : 0;
    f.any_scale = any_scale ? 1 : 0;
    if (any_scale && (p.g.map_scale == 1 || p.g.map_scale == 2)) {
      bool al = true;  // the specialised kernels read the map as 32-bit (scale 1) / 16-bit (scale 2) words
      for (int i = 0; i < n; i++) al = al && (reinterpret_cast<uintptr_t>(v[i].map_in) % (p.g.map_scale == 1 ? 4 : 2)) == 0;
      if (al) f.scale_special = p.g.map_scale;
      idw_fill(&f.w2[0][0], 2, 1, 1);
    } else if (any_scale && p.g.map_scale % 4 == 0) {
      f.scale_special = 4;  // a 4x4 pixel block never straddles a map cell: one set of map samples per block
    }
    // linear outputs: every non-toe channel has Phi > 0.0904, so q >= 0.0904 * f'min must reach the L table's first
    // octave, and q^2.4 must be a normal half with margin (the F16 kernel then skips the subnormal/flush checks)
    {
      const double qmin = 0.0904 * exp2((A > 0 ? B : A + B) / 2.4);
      f.lin_table_ok = (qmin >= ldexp(1.0, kLLoExp) && pow(qmin, 2.4) >= ldexp(1.0, -14) * 1.01) ? 1 : 0;
    }
    for (int i = 0; i < 16; i++)
      for (int c = 1; c < 4; c++) f.wq[c - 1][i] = (float)((double)f.w[i][c] / 2.4 / 255.0);
    f.idw = p.idw;
    f.tab = tabs;
    f.single = v[0];
    f.batch = batch;
    e = launch_apply_fast_impl(f, n, s);
  } else {
    p.single = v[0];
    p.batch = batch;
    e = launch_apply_impl(p, n, vec, s);
  }
  if (slot) cudaEventRecord(slot->done, s);
  if (e != cudaSuccess) return fail(e, "apply kernel launch");
  g_launches++;
  return UHDRB200_NO_ERROR;
}

int tonemap_core(int n, const uhdrb200_image* src, uhdrb200_image* dest, cudaStream_t s) {
  if (n <= 0) return UHDRB200_NO_ERROR;
  if (!src || !dest) return UHDRB200_ERROR_BAD_PTR;
  for (int i = 0; i < n; i++) {
    int st = validate_tonemap(&src[i], &dest[i]);
    if (st) return st;
    if (i && (!same_layout(src[i], src[0]) || dest[i].luma_stride != dest[0].luma_stride ||
              dest[i].chroma_stride != dest[0].chroma_stride))
      return UHDRB200_ERROR_RESOLUTION_MISMATCH;
  }
  if (!fits_int(src[0].width) || !fits_int(src[0].height) || !fits_int(src[0].luma_stride) || !fits_int(dest[0].luma_stride))
    return UHDRB200_ERROR_UNSUPPORTED_WIDTH_HEIGHT;
  int dev = 0;
  int st = core_device(&dev);
  if (st) return st;
  ToneMapParams p;
  memset(&p, 0, sizeof p);
  p.g = geometry_of(nullptr, &src[0]);
  p.g.hdr_luma_stride = (int)src[0].luma_stride;  // toneMap uses the stride as given, ultrahdr.cpp:527
  p.g.dst_luma_stride = (int)dest[0].luma_stride;
  p.g.dst_chroma_stride = (int)dest[0].chroma_stride;
  std::vector<Planes> v(n);
  bool vec = true;
  for (int i = 0; i < n; i++) {
    memset(&v[i], 0, sizeof(Planes));
    v[i].hdr_y = src[i].data;
    v[i].hdr_uv = src[i].chroma_data;
    v[i].out = dest[i].data;
    v[i].out_u = dest[i].chroma_data;
    v[i].out_v = (uint8_t*)dest[i].chroma_data + dest[i].chroma_stride * dest[i].height / 2;  // ultrahdr.cpp:539-540
    vec = vec && tonemap_vec_ok(p.g, v[i]);
    dest[i].colorGamut = src[i].colorGamut;  // ultrahdr.cpp:556
  }
  BatchSlot* slot = nullptr;
  if (n == 1) p.single = v[0];
  else if ((st = stage_batch(v, s, &p.batch, &slot))) return st;
  cudaError_t e = launch_tonemap_impl(p, n, vec, s);
  if (slot) cudaEventRecord(slot->done, s);
  if (e != cudaSuccess) return fail(e, "toneMap kernel launch");
  g_launches++;
  return UHDRB200_NO_ERROR;
}

// jpegr.cpp:1134-1195: selector; identity conversions return NO_ERROR without touching the image
// YUV->YUV matrices, gainmapmath.cpp:447-481; [src][dst]. All have the first column (1, 0, 0), which
// transform_quad (uhdr_convert.cuh) relies on.
const float kYuvToYuv[3][3][9] = {
    {{0}, {1.0f, 0.101579f, 0.196076f, 0.0f, 0.989854f, -0.110653f, 0.0f, -0.072453f, 0.983398f},
     {1.0f, -0.016969f, 0.096312f, 0.0f, 0.995306f, -0.051192f, 0.0f, 0.011507f, 1.002637f}},
    {{1.0f, -0.118188f, -0.212685f, 0.0f, 1.018640f, 0.114618f, 0.0f, 0.075049f, 1.025327f}, {0},
     {1.0f, -0.128245f, -0.115879f, 0.0f, 1.010016f, 0.061592f, 0.0f, 0.086969f, 1.029350f}},
    {{1.0f, 0.018149f, -0.095132f, 0.0f, 1.004123f, 0.051267f, 0.0f, -0.011524f, 0.996782f},
     {1.0f, 0.117887f, 0.105521f, 0.0f, 0.995211f, -0.059549f, 0.0f, -0.084085f, 0.976518f}, {0}}};

int convert_yuv_core(uhdrb200_image* im, int src_enc, int dst_enc, cudaStream_t s) {
  if (!im) return UHDRB200_ERROR_BAD_PTR;
  if (src_enc == UHDRB200_COLORGAMUT_UNSPECIFIED || dst_enc == UHDRB200_COLORGAMUT_UNSPECIFIED)
    return UHDRB200_ERROR_INVALID_COLORGAMUT;
  if (src_enc < 0 || src_enc > 2 || dst_enc < 0 || dst_enc > 2) return UHDRB200_ERROR_INVALID_COLORGAMUT;
  if (src_enc == dst_enc) return UHDRB200_NO_ERROR;
  if (!im->data || !im->chroma_data) return UHDRB200_ERROR_BAD_PTR;
  int dev = 0;
  int st = core_device(&dev);
  if (st) return st;
  ConvertYuvParams p;
  memset(&p, 0, sizeof p);
  p.g = geometry_of(im, nullptr);
  memcpy(p.m, kYuvToYuv[src_enc][dst_enc], sizeof p.m);
  set_sdr_planes(p.single, *im);
  p.single.sdr_v = (const uint8_t*)im->chroma_data + im->chroma_stride * im->height / 2;  // gainmapmath.cpp:507
  cudaError_t e = launch_convert_yuv_impl(p, 1, convert_yuv_vec_ok(p.g, p.single), s);
  if (e != cudaSuccess) return fail(e, "convertYuv kernel launch");
  g_launches++;
  return UHDRB200_NO_ERROR;
}

// ---- row N3: SDR output pack (examples/ultrahdr_app.cpp:682-734) and copy-with-restride (jpegr.cpp:315-359)
int sdr_pack_core(int n, const uhdrb200_image* sdr, uhdrb200_image* dest, cudaStream_t s) {
  if (n <= 0) return UHDRB200_NO_ERROR;
  if (!sdr || !dest) return UHDRB200_ERROR_BAD_PTR;
  for (int i = 0; i < n; i++) {
    if (!sdr[i].data || !sdr[i].chroma_data || !dest[i].data) return UHDRB200_ERROR_BAD_PTR;
    if (sdr[i].width % 2 || sdr[i].height % 2) return UHDRB200_ERROR_UNSUPPORTED_WIDTH_HEIGHT;
    if (sdr[i].luma_stride < sdr[i].width || sdr[i].chroma_stride < sdr[i].width / 2) return UHDRB200_ERROR_INVALID_STRIDE;
    if (i && (!same_layout(sdr[i], sdr[0]) || dest[i].luma_stride != dest[0].luma_stride)) return UHDRB200_ERROR_RESOLUTION_MISMATCH;
  }
  if (!fits_int(sdr[0].width) || !fits_int(sdr[0].height) || !fits_int(sdr[0].luma_stride) || !fits_int(dest[0].luma_stride))
    return UHDRB200_ERROR_UNSUPPORTED_WIDTH_HEIGHT;
  int dev = 0;
  int st = core_device(&dev);
  if (st) return st;
  SdrPackParams p;
  memset(&p, 0, sizeof p);
  p.g = geometry_of(&sdr[0], nullptr);
  p.g.dst_luma_stride = (int)(dest[0].luma_stride >= sdr[0].width ? dest[0].luma_stride : sdr[0].width);  // in pixels
  std::vector<Planes> v(n);
  bool vec = true;
  for (int i = 0; i < n; i++) {
    memset(&v[i], 0, sizeof(Planes));
    set_sdr_planes(v[i], sdr[i]);
    v[i].out = dest[i].data;
    vec = vec && sdr_pack_vec_ok(p.g, v[i]);
    dest[i].width = sdr[i].width;
    dest[i].height = sdr[i].height;
    dest[i].colorGamut = sdr[i].colorGamut;
  }
  BatchSlot* slot = nullptr;
  if (n == 1) p.single = v[0];
  else if ((st = stage_batch(v, s, &p.batch, &slot))) return st;
  cudaError_t e = launch_sdr_pack_impl(p, n, vec, s);
  if (slot) cudaEventRecord(slot->done, s);
  if (e != cudaSuccess) return fail(e, "yuv420_to_rgba8888 kernel launch");
  g_launches++;
  return UHDRB200_NO_ERROR;
}

int copy_yuv420_core(const uhdrb200_image* src, uhdrb200_image* dst, cudaStream_t s) {
  if (!src || !dst || !src->data || !src->chroma_data || !dst->data || !dst->chroma_data) return UHDRB200_ERROR_BAD_PTR;
  if (src->width % 2 || src->height % 2) return UHDRB200_ERROR_UNSUPPORTED_WIDTH_HEIGHT;
  if (src->luma_stride < src->width || src->chroma_stride < src->width / 2 || dst->luma_stride < src->width ||
      dst->chroma_stride < src->width / 2)
    return UHDRB200_ERROR_INVALID_STRIDE;
  if (!fits_int(src->width) || !fits_int(src->height) || !fits_int(src->luma_stride) || !fits_int(dst->luma_stride))
    return UHDRB200_ERROR_UNSUPPORTED_WIDTH_HEIGHT;
  int dev = 0;
  int st = core_device(&dev);
  if (st) return st;
  SdrPackParams p;
  memset(&p, 0, sizeof p);
  p.g = geometry_of(src, nullptr);
  p.g.dst_luma_stride = (int)dst->luma_stride;
  p.g.dst_chroma_stride = (int)dst->chroma_stride;
  set_sdr_planes(p.single, *src);
  p.single.out = dst->data;
  p.single.out_u = dst->chroma_data;
  p.single.out_v = (uint8_t*)dst->chroma_data + dst->chroma_stride * (src->height / 2);
  cudaError_t e = launch_copy_yuv420_impl(p, 1, s);
  if (e != cudaSuccess) return fail(e, "copy_yuv420 kernel launch");
  g_launches++;
  dst->width = src->width;
  dst->height = src->height;
  dst->colorGamut = src->colorGamut;
  return UHDRB200_NO_ERROR;
}

// API-0 (jpegr.cpp:167-225): toneMap(hdr -> sdr) then generateGainMap(sdr, hdr). One fused launch when the
// layout allows it (width, height % 4 == 0, vector-aligned planes), else the two kernels back to back.
int api0_core(int n, const uhdrb200_image* hdr, int hdr_tf, int sdr_to_601, uhdrb200_image* sdr, uhdrb200_metadata* md,
              uhdrb200_image* dest, cudaStream_t s) {
  if (n <= 0) return UHDRB200_NO_ERROR;
  if (!hdr || !sdr || !md || !dest) return UHDRB200_ERROR_BAD_PTR;
  for (int i = 0; i < n; i++) {
    int st = validate_tonemap(&hdr[i], &sdr[i]);
    if (st) return st;
    uhdrb200_image tmp = sdr[i];
    tmp.colorGamut = hdr[i].colorGamut;  // what toneMap would have set (ultrahdr.cpp:556)
    if ((st = validate_generate(&tmp, &hdr[i], hdr_tf, &md[i], &dest[i]))) return st;
    if (!dest[i].data) return UHDRB200_ERROR_BAD_PTR;
  }
  // jpegr.cpp:223-225: the SDR image is re-encoded to BT.601 for the JPEG unless its gamut is P3 already
  const int gamut = hdr[0].colorGamut;
  const bool conv = sdr_to_601 && gamut != UHDRB200_COLORGAMUT_P3;
  bool fused = true;
  for (int i = 0; i < n && fused; i++) {
    Geometry g = geometry_of(&sdr[i], &hdr[i]);
    Planes pl;
    memset(&pl, 0, sizeof pl);
    set_sdr_planes(pl, sdr[i]);
    pl.hdr_y = hdr[i].data;
    pl.hdr_uv = hdr[i].chroma_data;
    fused = api0_fused_ok(g, pl) && hdr[i].luma_stride != 0 && (i == 0 || same_layout(sdr[i], sdr[0]));
  }
  if (!fused) {
    int st = tonemap_core(n, hdr, sdr, s);
    if (st) return st;
    if ((st = generate_core(n, sdr, hdr, hdr_tf, md, dest, 0, s))) return st;
    for (int i = 0; i < n && conv; i++)
      if ((st = convert_yuv_core(&sdr[i], gamut, UHDRB200_COLORGAMUT_P3, s))) return st;
    return UHDRB200_NO_ERROR;
  }
  for (int i = 0; i < n; i++) sdr[i].colorGamut = hdr[i].colorGamut;
  return generate_core(n, sdr, hdr, hdr_tf, md, dest, 0, s, true, conv ? kYuvToYuv[gamut][UHDRB200_COLORGAMUT_P3] : nullptr);
}

}  // namespace

// ================================================================ no-allocation batched generateGainMap
// (no counterpart in the reference, see include/uhdr_b200.h)
// The batched launches read their per-image plane pointers from a small device array. The regular entry point takes that
// array from a cudaMalloc'ed ring (stage_batch); this variant keeps it in a file-scope __device__ array instead, so
// that it never allocates. Every GPU that loads the library holds the array (18 KB) until exit; all calls share it.
static __device__ __align__(256) unsigned char g_noalloc_batch[UHDRB200_NOALLOC_MAX_IMAGES * sizeof(uhdr::Planes)];

namespace {

std::mutex g_noalloc_mu;
Planes* g_noalloc_dev[64];  // per device: the DEVICE address of g_noalloc_batch (in host code the symbol is only a stand-in)

// once per device, with that device current
int noalloc_scratch(Planes** out) {
  int dev = 0;
  CU(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) {
    t_ctx.err = "no-alloc entry point: device index beyond 63";
    return UHDRB200_ERROR_UNSUPPORTED_FEATURE;
  }
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_dev[dev]) {
    void* ptr = nullptr;
    CU(cudaGetSymbolAddress((void**)&ptr, g_noalloc_batch));
    g_noalloc_dev[dev] = static_cast<Planes*>(ptr);
  }
  *out = g_noalloc_dev[dev];
  return 0;
}

int generate_noalloc(int n, const uhdrb200_image* sdr, const uhdrb200_image* hdr, int hdr_tf, uhdrb200_metadata* md,
                     uhdrb200_image* dest, int sdr_is_601, cudaStream_t s) {
  if (n <= 0) return UHDRB200_NO_ERROR;
  if (n > UHDRB200_NOALLOC_MAX_IMAGES) {
    char msg[160];
    snprintf(msg, sizeof msg, "uhdrb200_generate_gainmap_batch_dev_noalloc: n = %d exceeds UHDRB200_NOALLOC_MAX_IMAGES (%d)", n,
             UHDRB200_NOALLOC_MAX_IMAGES);
    t_ctx.err = msg;
    return UHDRB200_ERROR_UNSUPPORTED_FEATURE;
  }
  if (t_numeric_mode != UHDRB200_MODE_ANALYTIC) {  // the LUT mode's tables live in allocated device memory
    t_ctx.err = "uhdrb200_generate_gainmap_batch_dev_noalloc: only UHDRB200_MODE_ANALYTIC is supported";
    return UHDRB200_ERROR_UNSUPPORTED_FEATURE;
  }
  if (!sdr || !hdr || !md || !dest) return UHDRB200_ERROR_BAD_PTR;
  for (int i = 0; i < n; i++) {  // the checks of generate_core, in the reference's order (ultrahdr.cpp:189-202, 245-247)
    int st = validate_generate(&sdr[i], &hdr[i], hdr_tf, &md[i], &dest[i]);
    if (st) return st;
    if (!dest[i].data) return UHDRB200_ERROR_BAD_PTR;
    if (i && (!same_layout(sdr[i], sdr[0]) || !same_layout(hdr[i], hdr[0]))) return UHDRB200_ERROR_RESOLUTION_MISMATCH;
  }
  if (!fits_int(sdr[0].width) || !fits_int(sdr[0].height) || !fits_int(sdr[0].luma_stride) || !fits_int(hdr[0].luma_stride))
    return UHDRB200_ERROR_UNSUPPORTED_WIDTH_HEIGHT;
  Planes* scratch = nullptr;
  int st = noalloc_scratch(&scratch);  // runs on the caller's current device and stream: no per-thread context, no stream creation
  if (st) return st;
  GenerateParams p;
  memset(&p, 0, sizeof p);
  p.g = geometry_of(&sdr[0], &hdr[0]);
  p.g.map_scale = 4;  // kMapDimensionScaleFactor, ultrahdr.h:213
  p.g.map_width = p.g.width / 4;
  p.g.map_height = p.g.height / 4;
  p.sdr_k = yuv_coefs(sdr_is_601 ? 1 : sdr[0].colorGamut);
  p.hdr_k = yuv_coefs(hdr[0].colorGamut);
  p.luma = luma_coefs(sdr[0].colorGamut);
  p.gamut = gamut_matrix(sdr[0].colorGamut, hdr[0].colorGamut);
  p.use_gamut = sdr[0].colorGamut != hdr[0].colorGamut;
  p.hdr_tf = hdr_tf;
  p.hdr_white_nits = hdr_tf == UHDRB200_TF_PQ ? 10000.0f : 1000.0f;
  static thread_local Planes host_planes[UHDRB200_NOALLOC_MAX_IMAGES];  // fixed storage: no heap either
  bool vec = true;
  for (int i = 0; i < n; i++) {
    fill_metadata(&md[i], hdr_tf);
    fill_map_desc(&dest[i], (size_t)p.g.map_width, (size_t)p.g.map_height);
    memset(&host_planes[i], 0, sizeof(Planes));
    set_sdr_planes(host_planes[i], sdr[i]);
    host_planes[i].hdr_y = hdr[i].data;
    host_planes[i].hdr_uv = hdr[i].chroma_data;
    host_planes[i].out = dest[i].data;
    vec = vec && generate_vec_ok(p.g, host_planes[i]);
  }
  p.gc = gain_code_of(md[0].minContentBoost, md[0].maxContentBoost);
  // the n descriptors overwrite the head of the shared array (nothing in it has to start at zero; entries beyond n keep
  // the previous call's contents and are never read: the grid's z extent is n). Pageable source: the runtime has staged
  // the bytes when this returns, so host_planes may be reused by the next call.
  CU(cudaMemcpyAsync(scratch, host_planes, sizeof(Planes) * (size_t)n, cudaMemcpyHostToDevice, s));
  p.batch = scratch;
  cudaError_t e = launch_generate_impl(p, n, vec, s);
  if (e != cudaSuccess) return fail(e, "generate kernel launch (no-alloc)");
  g_launches++;
  return UHDRB200_NO_ERROR;
}

}  // namespace

// ================================================================ host-pointer staging helpers
namespace {

size_t yuv420_luma_bytes(const uhdrb200_image& im) { return im.height ? im.luma_stride * (im.height - 1) + im.width : 0; }
size_t yuv420_chroma_bytes(const uhdrb200_image& im) {  // U rows followed by V rows at +chroma_stride*(h/2)
  size_t ch = im.height / 2;
  return ch ? im.chroma_stride * ch + im.chroma_stride * (ch - 1) + im.width / 2 : 0;
}
size_t p010_luma_stride(const uhdrb200_image& im) { return im.luma_stride == 0 ? im.width : im.luma_stride; }
size_t p010_luma_bytes(const uhdrb200_image& im) { return im.height ? 2 * (p010_luma_stride(im) * (im.height - 1) + im.width) : 0; }
size_t p010_chroma_bytes(const uhdrb200_image& im) {
  size_t ch = im.height / 2;
  return ch ? 2 * (im.chroma_stride * (ch - 1) + (im.width & ~(size_t)1)) : 0;
}

int upload(DevBuf& b, const void* host, size_t bytes, cudaStream_t s) {
  int st = reserve(b, bytes ? bytes : 1);
  if (st) return st;
  (void)s;  // always the calling thread's stream
  return copy_h2d(t_ctx, b.p, host, bytes);
}

int generate_host(const uhdrb200_image* sdr, const uhdrb200_image* hdr, int hdr_tf, uhdrb200_metadata* md,
                  uhdrb200_image* dest, int sdr_is_601, void* map_buf, size_t map_buf_bytes, bool own_alloc) {
  int st = validate_generate(sdr, hdr, hdr_tf, md, dest);
  if (st) return st;
  if ((st = ensure_ctx())) return st;
  Ctx& c = t_ctx;
  const size_t mw = sdr->width / 4, mh = sdr->height / 4, mbytes = mw * mh;
  if (!own_alloc && (!map_buf || map_buf_bytes < mbytes)) return -20004;  // ERROR_ULTRAHDR_BUFFER_TOO_SMALL
  if ((st = upload(c.buf[0], sdr->data, yuv420_luma_bytes(*sdr), c.stream))) return st;
  if ((st = upload(c.buf[1], sdr->chroma_data, yuv420_chroma_bytes(*sdr), c.stream))) return st;
  if ((st = upload(c.buf[2], hdr->data, p010_luma_bytes(*hdr), c.stream))) return st;
  if ((st = upload(c.buf[3], hdr->chroma_data, p010_chroma_bytes(*hdr), c.stream))) return st;
  if ((st = reserve(c.buf[4], mbytes ? mbytes : 1))) return st;
  uhdrb200_image dsdr = *sdr, dhdr = *hdr, ddst;
  memset(&ddst, 0, sizeof ddst);
  dsdr.data = c.buf[0].p; dsdr.chroma_data = c.buf[1].p;
  dhdr.data = c.buf[2].p; dhdr.chroma_data = c.buf[3].p;
  ddst.data = c.buf[4].p;
  if ((st = generate_core(1, &dsdr, &dhdr, hdr_tf, md, &ddst, sdr_is_601, c.stream))) return st;
  void* out = own_alloc ? malloc(mbytes ? mbytes : 1) : map_buf;
  if (!out) return UHDRB200_ERROR_INSUFFICIENT_RESOURCE;
  st = copy_d2h(c, out, c.buf[4].p, mbytes);
  cudaError_t e = st ? cudaErrorUnknown : join_stream(c);
  if (e != cudaSuccess) {
    if (own_alloc) free(out);
    return st ? st : fail(e, "generateGainMap device->host");
  }
  fill_map_desc(dest, mw, mh);
  dest->data = out;
  return UHDRB200_NO_ERROR;
}

size_t out_bpp(int fmt) {
  switch (fmt) {
    case UHDRB200_OUTPUT_HDR_LINEAR: return 8;
    case UHDRB200_OUTPUT_HDR_PQ:
    case UHDRB200_OUTPUT_HDR_HLG: return 4;
    case UHDRB200_OUTPUT_HDR_LINEAR_RGB_10BIT: return 6;
    default: return 0;
  }
}

int apply_host(const uhdrb200_image* sdr, const uhdrb200_image* map, const uhdrb200_metadata* md, int fmt,
               float max_display_boost, uhdrb200_image* dest) {
  int st = validate_apply(sdr, map, md, dest);
  if (st) return st;
  const size_t obytes = sdr->width * sdr->height * out_bpp(fmt);
  if (obytes == 0) {  // unknown format: reference sets the dims and writes nothing (ultrahdr.cpp:411-413, 491-493)
    dest->width = sdr->width; dest->height = sdr->height; dest->colorGamut = sdr->colorGamut;
    return UHDRB200_NO_ERROR;
  }
  if (!dest->data) return UHDRB200_ERROR_BAD_PTR;
  if ((st = ensure_ctx())) return st;
  Ctx& c = t_ctx;
  uhdrb200_image dsdr = *sdr, dmap = *map, ddst = *dest;
  if ((sdr->luma_stride % 4 != 0 || sdr->chroma_stride % 2 != 0) && sdr->width % 4 == 0 && sdr->height % 4 == 0) {
    // Row strides the table-driven kernels cannot take (they read 32-bit luma and 16-bit chroma words): the planes are
    // re-strided by the copy engine on their way up, so the image still gets those kernels instead of the per-pixel one
    const size_t w = sdr->width, h = sdr->height, ls2 = w, cs2 = w / 2;  // w % 4 == 0
    if ((st = reserve(c.buf[0], ls2 * h))) return st;
    if ((st = reserve(c.buf[1], cs2 * (h / 2) * 2))) return st;
    const uint8_t* ch = (const uint8_t*)sdr->chroma_data;
    CU(cudaMemcpy2DAsync(c.buf[0].p, ls2, sdr->data, sdr->luma_stride, w, h, cudaMemcpyHostToDevice, c.stream));
    CU(cudaMemcpy2DAsync(c.buf[1].p, cs2, ch, sdr->chroma_stride, w / 2, h / 2, cudaMemcpyHostToDevice, c.stream));
    CU(cudaMemcpy2DAsync((uint8_t*)c.buf[1].p + cs2 * (h / 2), cs2, ch + sdr->chroma_stride * (h / 2), sdr->chroma_stride, w / 2, h / 2,
                         cudaMemcpyHostToDevice, c.stream));
    dsdr.luma_stride = ls2;
    dsdr.chroma_stride = cs2;
  } else {
    if ((st = upload(c.buf[0], sdr->data, yuv420_luma_bytes(*sdr), c.stream))) return st;
    if ((st = upload(c.buf[1], sdr->chroma_data, yuv420_chroma_bytes(*sdr), c.stream))) return st;
  }
  if ((st = upload(c.buf[5], map->data, map->width * map->height, c.stream))) return st;
  if ((st = reserve(c.buf[4], obytes))) return st;
  dsdr.data = c.buf[0].p; dsdr.chroma_data = c.buf[1].p;
  dmap.data = c.buf[5].p;
  ddst.data = c.buf[4].p;
  if ((st = apply_core(1, &dsdr, &dmap, md, fmt, max_display_boost, &ddst, c.stream))) return st;
  if ((st = copy_d2h(c, dest->data, c.buf[4].p, obytes))) return st;
  CU(join_stream(c));
  dest->width = sdr->width; dest->height = sdr->height; dest->colorGamut = sdr->colorGamut;
  return UHDRB200_NO_ERROR;
}

int tonemap_host(const uhdrb200_image* src, uhdrb200_image* dest) {
  int st = validate_tonemap(src, dest);
  if (st) return st;
  if ((st = ensure_ctx())) return st;
  Ctx& c = t_ctx;
  uhdrb200_image s2 = *src;  // toneMap reads luma_stride as given (no 0 -> width fallback, ultrahdr.cpp:527)
  const size_t ly = src->height ? 2 * (src->luma_stride * (src->height - 1) + src->width) : 0;
  const size_t dl = dest->luma_stride * dest->height;               // rows are written out to the stride
  const size_t dc = 2 * (dest->chroma_stride * (dest->height / 2));  // U rows then V rows
  if ((st = upload(c.buf[2], src->data, ly, c.stream))) return st;
  if ((st = upload(c.buf[3], src->chroma_data, p010_chroma_bytes(*src), c.stream))) return st;
  if ((st = reserve(c.buf[6], dl ? dl : 1)) || (st = reserve(c.buf[7], dc ? dc : 1))) return st;
  uhdrb200_image dd = *dest;
  s2.data = c.buf[2].p; s2.chroma_data = c.buf[3].p;
  dd.data = c.buf[6].p; dd.chroma_data = c.buf[7].p;
  if ((st = tonemap_core(1, &s2, &dd, c.stream))) return st;
  // stride tails of the last rows are part of the reference's writes (memset to the stride)
  if ((st = copy_d2h(c, dest->data, c.buf[6].p, dl))) return st;
  if ((st = copy_d2h(c, dest->chroma_data, c.buf[7].p, dc))) return st;
  CU(join_stream(c));
  dest->colorGamut = src->colorGamut;
  return UHDRB200_NO_ERROR;
}

int convert_yuv_host(uhdrb200_image* im, int src_enc, int dst_enc) {
  if (!im) return UHDRB200_ERROR_BAD_PTR;
  if (src_enc == UHDRB200_COLORGAMUT_UNSPECIFIED || dst_enc == UHDRB200_COLORGAMUT_UNSPECIFIED)
    return UHDRB200_ERROR_INVALID_COLORGAMUT;
  if (src_enc < 0 || src_enc > 2 || dst_enc < 0 || dst_enc > 2) return UHDRB200_ERROR_INVALID_COLORGAMUT;
  if (src_enc == dst_enc) return UHDRB200_NO_ERROR;
  if (!im->data || !im->chroma_data) return UHDRB200_ERROR_BAD_PTR;
  int st = ensure_ctx();
  if (st) return st;
  Ctx& c = t_ctx;
  const size_t lb = yuv420_luma_bytes(*im), cb = yuv420_chroma_bytes(*im);
  if ((st = upload(c.buf[0], im->data, lb, c.stream))) return st;
  if ((st = upload(c.buf[1], im->chroma_data, cb, c.stream))) return st;
  uhdrb200_image d = *im;
  d.data = c.buf[0].p; d.chroma_data = c.buf[1].p;
  if ((st = convert_yuv_core(&d, src_enc, dst_enc, c.stream))) return st;
  if ((st = copy_d2h(c, im->data, c.buf[0].p, lb))) return st;
  if ((st = copy_d2h(c, im->chroma_data, c.buf[1].p, cb))) return st;
  CU(join_stream(c));
  return UHDRB200_NO_ERROR;
}

int sdr_pack_host(const uhdrb200_image* sdr, uhdrb200_image* dest) {
  if (!sdr || !dest || !sdr->data || !sdr->chroma_data || !dest->data) return UHDRB200_ERROR_BAD_PTR;
  if (sdr->luma_stride < sdr->width || sdr->chroma_stride < sdr->width / 2) return UHDRB200_ERROR_INVALID_STRIDE;
  int st = ensure_ctx();
  if (st) return st;
  Ctx& c = t_ctx;
  const size_t lb = yuv420_luma_bytes(*sdr), cb = yuv420_chroma_bytes(*sdr);
  const size_t os = dest->luma_stride >= sdr->width ? dest->luma_stride : sdr->width;
  const size_t ob = sdr->height ? 4 * (os * (sdr->height - 1) + sdr->width) : 0;
  if ((st = upload(c.buf[0], sdr->data, lb, c.stream))) return st;
  if ((st = upload(c.buf[1], sdr->chroma_data, cb, c.stream))) return st;
  if ((st = reserve(c.buf[6], ob ? ob : 1))) return st;
  uhdrb200_image s2 = *sdr, d2 = *dest;
  s2.data = c.buf[0].p; s2.chroma_data = c.buf[1].p;
  d2.data = c.buf[6].p;
  if ((st = sdr_pack_core(1, &s2, &d2, c.stream))) return st;
  if (os == sdr->width) {
    if ((st = copy_d2h(c, dest->data, c.buf[6].p, ob))) return st;
  } else if (ob) {  // only the pixels: bytes between width and the stride stay the caller's
    CU(cudaMemcpy2DAsync(dest->data, os * 4, c.buf[6].p, os * 4, sdr->width * 4, sdr->height, cudaMemcpyDeviceToHost, c.stream));
  }
  CU(join_stream(c));
  dest->width = sdr->width;
  dest->height = sdr->height;
  dest->colorGamut = sdr->colorGamut;
  return UHDRB200_NO_ERROR;
}

int copy_yuv420_host(const uhdrb200_image* src, uhdrb200_image* dst) {
  if (!src || !dst || !src->data || !src->chroma_data || !dst->data || !dst->chroma_data) return UHDRB200_ERROR_BAD_PTR;
  if (src->luma_stride < src->width || src->chroma_stride < src->width / 2 || dst->luma_stride < src->width ||
      dst->chroma_stride < src->width / 2)
    return UHDRB200_ERROR_INVALID_STRIDE;
  int st = ensure_ctx();
  if (st) return st;
  Ctx& c = t_ctx;
  const size_t lb = yuv420_luma_bytes(*src), cb = yuv420_chroma_bytes(*src);
  const size_t dl = dst->luma_stride * src->height, dc = 2 * (dst->chroma_stride * (src->height / 2));
  if ((st = upload(c.buf[0], src->data, lb, c.stream))) return st;
  if ((st = upload(c.buf[1], src->chroma_data, cb, c.stream))) return st;
  if ((st = reserve(c.buf[6], dl ? dl : 1)) || (st = reserve(c.buf[7], dc ? dc : 1))) return st;
  uhdrb200_image s2 = *src, d2 = *dst;
  s2.data = c.buf[0].p; s2.chroma_data = c.buf[1].p;
  d2.data = c.buf[6].p; d2.chroma_data = c.buf[7].p;
  if ((st = copy_yuv420_core(&s2, &d2, c.stream))) return st;
  if ((st = copy_d2h(c, dst->data, c.buf[6].p, dl))) return st;
  if ((st = copy_d2h(c, dst->chroma_data, c.buf[7].p, dc))) return st;
  CU(join_stream(c));
  dst->width = src->width;
  dst->height = src->height;
  dst->colorGamut = src->colorGamut;
  return UHDRB200_NO_ERROR;
}

int api0_host(const uhdrb200_image* hdr, int hdr_tf, int sdr_to_601, uhdrb200_image* sdr, uhdrb200_metadata* md, uhdrb200_image* dest,
              void* map_buf, size_t map_buf_bytes) {
  int st = validate_tonemap(hdr, sdr);
  if (st) return st;
  uhdrb200_image tmp = *sdr;
  tmp.colorGamut = hdr->colorGamut;
  if ((st = validate_generate(&tmp, hdr, hdr_tf, md, dest))) return st;
  const size_t mw = hdr->width / 4, mh = hdr->height / 4, mbytes = mw * mh;
  if (!map_buf || map_buf_bytes < mbytes) return -20004;  // ERROR_ULTRAHDR_BUFFER_TOO_SMALL
  if ((st = ensure_ctx())) return st;
  Ctx& c = t_ctx;
  // toneMap reads the luma stride as given (ultrahdr.cpp:527), generateGainMap reads 0 as `width` (gainmapmath.cpp:585):
  // stage whichever extent is larger
  const size_t ly_lit = hdr->height ? 2 * (hdr->luma_stride * (hdr->height - 1) + hdr->width) : 0;
  const size_t ly = ly_lit > p010_luma_bytes(*hdr) ? ly_lit : p010_luma_bytes(*hdr);
  const size_t dl = sdr->luma_stride * sdr->height, dc = 2 * (sdr->chroma_stride * (sdr->height / 2));
  if ((st = upload(c.buf[2], hdr->data, ly, c.stream))) return st;
  if ((st = upload(c.buf[3], hdr->chroma_data, p010_chroma_bytes(*hdr), c.stream))) return st;
  if ((st = reserve(c.buf[6], dl ? dl : 1)) || (st = reserve(c.buf[7], dc ? dc : 1)) || (st = reserve(c.buf[4], mbytes ? mbytes : 1))) return st;
  uhdrb200_image dh = *hdr, ds = *sdr, dm;
  memset(&dm, 0, sizeof dm);
  dh.data = c.buf[2].p; dh.chroma_data = c.buf[3].p;
  ds.data = c.buf[6].p; ds.chroma_data = c.buf[7].p;
  dm.data = c.buf[4].p;
  if ((st = api0_core(1, &dh, hdr_tf, sdr_to_601, &ds, md, &dm, c.stream))) return st;
  if ((st = copy_d2h(c, sdr->data, c.buf[6].p, dl))) return st;
  if ((st = copy_d2h(c, sdr->chroma_data, c.buf[7].p, dc))) return st;
  if ((st = copy_d2h(c, map_buf, c.buf[4].p, mbytes))) return st;
  CU(join_stream(c));
  sdr->colorGamut = hdr->colorGamut;
  fill_map_desc(dest, mw, mh);
  dest->data = map_buf;
  return UHDRB200_NO_ERROR;
}

}  // namespace

// ================================================================ row N4: JPEG coefficient planes -> sample planes
namespace {

int validate_jpeg(int n, const uhdrb200_jpeg_component* comps, int max_n) {
  if (n <= 0) return UHDRB200_NO_ERROR;
  if (!comps) return UHDRB200_ERROR_BAD_PTR;
  if (n > max_n) return UHDRB200_ERROR_UNSUPPORTED_FEATURE;
  for (int i = 0; i < n; i++) {
    const uhdrb200_jpeg_component& c = comps[i];
    if (c.blocks_w == 0 || c.blocks_h == 0) continue;
    if (!c.coeffs || !c.quant || !c.plane) return UHDRB200_ERROR_BAD_PTR;
    if (c.width > c.blocks_w * 8 || c.height > c.blocks_h * 8) return UHDRB200_ERROR_RESOLUTION_MISMATCH;
    if (c.stride < c.width) return UHDRB200_ERROR_INVALID_STRIDE;
    if (!fits_int(c.blocks_w * 8) || !fits_int(c.blocks_h * 8)) return UHDRB200_ERROR_UNSUPPORTED_WIDTH_HEIGHT;
  }
  return UHDRB200_NO_ERROR;
}

int jpeg_idct_core(int n, const uhdrb200_jpeg_component* comps, cudaStream_t s) {
  int st = validate_jpeg(n, comps, 4);
  if (st || n <= 0) return st;
  int dev = 0;
  if ((st = core_device(&dev))) return st;
  JpegIdctParams p;
  memset(&p, 0, sizeof p);
  p.n_components = n;
  for (int i = 0; i < n; i++) {
    if ((reinterpret_cast<uintptr_t>(comps[i].coeffs) & 1) != 0 || (reinterpret_cast<uintptr_t>(comps[i].quant) & 1) != 0) {
      t_ctx.err = "uhdrb200_jpeg_idct: coeffs and quant must be 2-byte aligned";
      return UHDRB200_ERROR_UNSUPPORTED_FEATURE;
    }
    p.comp[i] = JpegIdctComponent{comps[i].coeffs, comps[i].quant, comps[i].blocks_w, comps[i].blocks_h, (uint8_t*)comps[i].plane,
                                  comps[i].width, comps[i].height, comps[i].stride};
  }
  cudaError_t e = launch_jpeg_idct_impl(p, s);
  if (e != cudaSuccess) return fail(e, "jpeg_idct kernel launch");
  g_launches++;
  return UHDRB200_NO_ERROR;
}

int jpeg_idct_host(int n, const uhdrb200_jpeg_component* comps) {
  int st = validate_jpeg(n, comps, 3);
  if (st || n <= 0) return st;
  if ((st = ensure_ctx())) return st;
  Ctx& c = t_ctx;
  uhdrb200_jpeg_component d[3];
  if ((st = reserve(c.buf[3], 3 * 64 * sizeof(uint16_t)))) return st;
  for (int i = 0; i < n; i++) {
    d[i] = comps[i];
    const size_t cb = comps[i].blocks_w * comps[i].blocks_h * 64 * sizeof(int16_t);
    const size_t pb = comps[i].height ? comps[i].stride * (comps[i].height - 1) + comps[i].width : 0;
    if (cb == 0) continue;
    if ((st = upload(c.buf[i], comps[i].coeffs, cb, c.stream))) return st;
    if ((st = copy_h2d(c, (uint8_t*)c.buf[3].p + i * 128, comps[i].quant, 128))) return st;
    if ((st = reserve(c.buf[4 + i], pb ? pb : 1))) return st;
    d[i].coeffs = (const int16_t*)c.buf[i].p;
    d[i].quant = (const uint16_t*)((uint8_t*)c.buf[3].p + i * 128);
    d[i].plane = c.buf[4 + i].p;
  }
  if ((st = jpeg_idct_core(n, d, c.stream))) return st;
  for (int i = 0; i < n; i++) {
    const uhdrb200_jpeg_component& h = comps[i];
    if (h.blocks_w == 0 || h.blocks_h == 0 || h.height == 0 || h.width == 0) continue;
    if (h.stride == h.width) {
      if ((st = copy_d2h(c, h.plane, d[i].plane, h.stride * (h.height - 1) + h.width))) return st;
    } else {  // only the samples: bytes between width and the stride stay the caller's
      CU(cudaMemcpy2DAsync(h.plane, h.stride, d[i].plane, h.stride, h.width, h.height, cudaMemcpyDeviceToHost, c.stream));
    }
  }
  CU(join_stream(c));
  return UHDRB200_NO_ERROR;
}

// ---------------------------------------------------------------- N4: whole baseline-JPEG decode (jpeg_huff.cu + jpeg_idct.cu)
// JpegDecoderHelper::decompressImage(image, length, DECODE_TO_YCBCR) (lib/src/jpegdecoderhelper.cpp:190-320, 380-470):
// 4:2:0 YCbCr -> planar Y | U | V (strides width and width / 2, floor), greyscale -> one plane; anything else is refused
// there ("decoding to YUV only supports 4:2:0 subsampling") and here. Sizes above kMaxWidth / kMaxHeight = 8192 are
// refused like jpegdecoderhelper.cpp:254-258.
int jpeg_status(int parse_status, const char* why) {
  if (parse_status == 0) return UHDRB200_NO_ERROR;
  t_ctx.err = std::string("uhdrb200_jpeg: ") + why;
  return parse_status == 2 ? UHDRB200_ERROR_UNSUPPORTED_FEATURE : UHDRB200_ERROR_DECODE;
}
int jpeg_check_layout(const JpegHeader& h) {
  if (h.width > 8192 || h.height > 8192) {
    t_ctx.err = "uhdrb200_jpeg: image larger than 8192 x 8192 (kMaxWidth / kMaxHeight)";
    return UHDRB200_ERROR_UNSUPPORTED_WIDTH_HEIGHT;
  }
  if (h.n_comp == 3 && !(h.comp_h[0] == 2 && h.comp_v[0] == 2 && h.comp_h[1] == 1 && h.comp_v[1] == 1 && h.comp_h[2] == 1 && h.comp_v[2] == 1)) {
    t_ctx.err = "uhdrb200_jpeg: decoding to YUV only supports 4:2:0 subsampling";  // jpegdecoderhelper.cpp:286-292
    return UHDRB200_ERROR_UNSUPPORTED_FEATURE;
  }
  return UHDRB200_NO_ERROR;
}

// UHDRB200_JPEG_TIMING=1: wall-clock phase times of every decode on stderr (each phase joined; a debugging aid, not a mode)
struct JpegPhaseTimer {
  bool on;
  cudaStream_t s = nullptr;
  std::chrono::steady_clock::time_point t0;
  std::string line;
  JpegPhaseTimer() {
    static const bool env = [] { const char* e = getenv("UHDRB200_JPEG_TIMING"); return e && e[0] == '1'; }();
    on = env;
    t0 = std::chrono::steady_clock::now();
  }
  void mark(const char* what) {
    if (!on) return;
    if (s) cudaStreamSynchronize(s);
    const auto t1 = std::chrono::steady_clock::now();
    char b[64];
    snprintf(b, sizeof b, " %s %.0f us", what, std::chrono::duration<double, std::micro>(t1 - t0).count());
    line += b;
    t0 = t1;
  }
  ~JpegPhaseTimer() { if (on) fprintf(stderr, "[uhdrb200 jpeg]%s\n", line.c_str()); }
};

int jpeg_decode(const void* jpeg, size_t bytes, uhdrb200_image* dest, bool to_device, cudaStream_t user_stream) {
  if (!jpeg || !dest || !dest->data) return UHDRB200_ERROR_BAD_PTR;
  JpegPhaseTimer tm;
  static thread_local JpegHeader h;  // 11 KB with its tables
  const char* why = "";
  const int parsed = jpeg_parse((const uint8_t*)jpeg, bytes, &h, &why);
  int st = jpeg_status(parsed, why);
  if (st) return st;
  if ((st = jpeg_check_layout(h))) return st;
  if ((st = ensure_ctx())) return st;
  Ctx& c = t_ctx;
  cudaStream_t s = to_device ? user_stream : c.stream;  // _dev: the caller's stream (0 = the default stream), ordered after its producers
  tm.s = s;
  tm.mark("parse");
  const size_t w = (size_t)h.width, ht = (size_t)h.height;
  // ---- entropy-coded segment: stuffing removed on the host, zero padded to whole 32-bit words + slack for the bit reader
  static thread_local std::vector<uint8_t> seg;
  static thread_local std::vector<uint32_t> ivals;  // [interval_first (n_intervals) | interval_of (n_subseq)] once laid out
  const uint64_t total_mcus = h.n_comp == 1 ? (uint64_t)((w + 7) / 8) * ((ht + 7) / 8)
                                            : (uint64_t)((w + 8 * h.hmax - 1) / (8 * h.hmax)) * ((ht + 8 * h.vmax - 1) / (8 * h.vmax));
  uint32_t ri = (uint32_t)h.restart_interval;
  if (ri >= total_mcus) ri = 0;  // one interval: the prediction never restarts
  const uint32_t max_intervals = ri ? (uint32_t)((total_mcus + ri - 1) / ri) : 1;
  constexpr size_t kSubBytes = kJpegSubseqBits / 8;
  seg.resize(h.scan_end - h.scan_begin + (size_t)max_intervals * kSubBytes + 64);
  ivals.resize(max_intervals);
  int marker = 0;
  uint32_t n_intervals = 1;
  const size_t nseg = jpeg_destuff((const uint8_t*)jpeg, h.scan_begin, h.scan_end, seg.data(), &marker, max_intervals, ivals.data(), &n_intervals);
  if (marker >= 0xD0 && marker <= 0xD7) {
    t_ctx.err = ri ? "uhdrb200_jpeg: more restart markers than restart intervals" : "uhdrb200_jpeg: restart markers without a restart interval (DRI)";
    return UHDRB200_ERROR_DECODE;
  }
  const size_t padded = (nseg + 63) & ~(size_t)31;
  memset(seg.data() + nseg, 0, padded - nseg);
  tm.mark("destuff");
  JpegScan sc;
  memset(&sc, 0, sizeof sc);
  sc.n_comp = h.n_comp;
  sc.total_bits = (uint64_t)nseg * 8;
  sc.n_subseq = (uint32_t)((sc.total_bits + kJpegSubseqBits - 1) / kJpegSubseqBits);
  sc.restart_interval = ri;
  sc.n_intervals = n_intervals;
  if (ri) {  // interval_first in subsequences, then interval_of per subsequence
    ivals.resize((size_t)n_intervals + sc.n_subseq);
    for (uint32_t i = 0; i < n_intervals; i++) ivals[i] = (uint32_t)(ivals[i] / kSubBytes);
    for (uint32_t i = 0; i < n_intervals; i++) {
      const uint32_t t0 = ivals[i], t1 = i + 1 < n_intervals ? ivals[i + 1] : sc.n_subseq;
      for (uint32_t t = t0; t < t1 && t < sc.n_subseq; t++) ivals[n_intervals + t] = i;
    }
  }
  if (h.n_comp == 1) {  // a single-component scan is not interleaved: one block per MCU whatever the sampling factors say (A.2.2)
    sc.blocks_per_mcu = 1;
    sc.mcus_x = (uint32_t)((w + 7) / 8);
    sc.mcus_y = (uint32_t)((ht + 7) / 8);
    sc.comp_h[0] = sc.comp_v[0] = 1;
    sc.comp_dc[0] = (uint8_t)h.comp_td[0];
    sc.comp_ac[0] = (uint8_t)h.comp_ta[0];
  } else {
    sc.mcus_x = (uint32_t)((w + 8 * h.hmax - 1) / (8 * h.hmax));
    sc.mcus_y = (uint32_t)((ht + 8 * h.vmax - 1) / (8 * h.vmax));
    int b = 0;
    for (int ci = 0; ci < h.n_comp; ci++) {
      sc.comp_h[ci] = (uint8_t)h.comp_h[ci];
      sc.comp_v[ci] = (uint8_t)h.comp_v[ci];
      sc.comp_dc[ci] = (uint8_t)h.comp_td[ci];
      sc.comp_ac[ci] = (uint8_t)h.comp_ta[ci];
      for (int y = 0; y < h.comp_v[ci]; y++)
        for (int x = 0; x < h.comp_h[ci]; x++, b++) {
          sc.blk_comp[b] = (uint8_t)ci;
          sc.blk_dx[b] = (uint8_t)x;
          sc.blk_dy[b] = (uint8_t)y;
        }
    }
    sc.blocks_per_mcu = b;
  }
  for (int b = 0; b < sc.blocks_per_mcu; b++) {
    sc.dc_mask |= (uint32_t)sc.comp_dc[sc.blk_comp[b]] << b;
    sc.ac_mask |= (uint32_t)sc.comp_ac[sc.blk_comp[b]] << b;
  }
  sc.total_blocks = sc.mcus_x * sc.mcus_y * (uint32_t)sc.blocks_per_mcu;
  size_t coeff_off[4] = {0, 0, 0, 0}, coeff_bytes = 0;
  for (int ci = 0; ci < h.n_comp; ci++) {
    sc.comp_blocks_w[ci] = sc.mcus_x * sc.comp_h[ci];
    sc.comp_blocks_h[ci] = sc.mcus_y * sc.comp_v[ci];
    coeff_off[ci] = coeff_bytes;
    coeff_bytes += (size_t)sc.comp_blocks_w[ci] * sc.comp_blocks_h[ci] * 64 * sizeof(int16_t);
  }
  // ---- device buffers of the calling thread: segment, tables, workspace, coefficients (+ planes for the host variant)
  if ((st = upload(c.buf[0], seg.data(), padded, s))) return st;
  if ((st = upload(c.buf[1], &h.tables, sizeof h.tables, s))) return st;
  if (ri) {
    if ((st = upload(c.buf[5], ivals.data(), ivals.size() * sizeof(uint32_t), s))) return st;
    sc.interval_first = (const uint32_t*)c.buf[5].p;
    sc.interval_of = sc.interval_first + n_intervals;
  }
  if ((st = reserve(c.buf[2], jpeg_huff_workspace_bytes(sc.n_subseq)))) return st;
  if ((st = reserve(c.buf[3], coeff_bytes))) return st;
  if (s != c.stream) CU(cudaStreamSynchronize(c.stream));  // the uploads ran on the thread's own stream: done before `s` reads them
  CU(cudaMemsetAsync(c.buf[3].p, 0, coeff_bytes, s));
  sc.data = (const uint32_t*)c.buf[0].p;
  sc.tables = (const JpegDeviceTables*)c.buf[1].p;
  for (int ci = 0; ci < h.n_comp; ci++) sc.coeffs[ci] = (int16_t*)((uint8_t*)c.buf[3].p + coeff_off[ci]);
  tm.mark("upload+memset");
  int iters = 0;
  cudaError_t e = jpeg_huff_decode(sc, c.buf[2].p, s, &iters);
  if (e != cudaSuccess) return fail(e, "jpeg entropy decode");
  tm.mark("huffman");
  g_launches += 6 + (uint64_t)iters;
  // ---- planes: Y w x h, then U and V (w / 2) x (h / 2), packed (decompressYUV, jpegdecoderhelper.cpp:380-399)
  const size_t cw = w / 2, chh = ht / 2;
  const size_t plane_bytes = h.n_comp == 1 ? w * ht : w * ht + 2 * cw * chh;
  uint8_t* planes = (uint8_t*)dest->data;
  if (!to_device) {
    if ((st = reserve(c.buf[4], plane_bytes))) return st;
    planes = (uint8_t*)c.buf[4].p;
  }
  uhdrb200_jpeg_component comp[3];
  const JpegDeviceTables* dt = sc.tables;
  for (int ci = 0; ci < h.n_comp; ci++) {
    comp[ci].coeffs = sc.coeffs[ci];
    comp[ci].quant = dt->quant[h.comp_tq[ci]];
    comp[ci].blocks_w = sc.comp_blocks_w[ci];
    comp[ci].blocks_h = sc.comp_blocks_h[ci];
    comp[ci].plane = ci == 0 ? planes : planes + w * ht + (size_t)(ci - 1) * cw * chh;
    comp[ci].width = ci == 0 ? w : cw;
    comp[ci].height = ci == 0 ? ht : chh;
    comp[ci].stride = comp[ci].width;
  }
  if (h.n_comp == 3 && (cw == 0 || chh == 0)) {
    t_ctx.err = "uhdrb200_jpeg: 4:2:0 image narrower or lower than 2 pixels";
    return UHDRB200_ERROR_UNSUPPORTED_WIDTH_HEIGHT;
  }
  {
    JpegIdctParams p;
    memset(&p, 0, sizeof p);
    p.n_components = h.n_comp;
    for (int i = 0; i < h.n_comp; i++)
      p.comp[i] = JpegIdctComponent{comp[i].coeffs, comp[i].quant, comp[i].blocks_w, comp[i].blocks_h, (uint8_t*)comp[i].plane,
                                    comp[i].width, comp[i].height, comp[i].stride};
    if ((e = launch_jpeg_idct_impl(p, s)) != cudaSuccess) return fail(e, "jpeg_idct kernel launch");
    g_launches++;
  }
  if (!to_device) {
    if ((st = copy_d2h(c, dest->data, planes, plane_bytes))) return st;
    CU(join_stream(c));
  } else {
    CU(cudaStreamSynchronize(s));  // the scratch buffers above belong to the calling thread: nothing may still read them
  }
  tm.mark("idct+join");
  dest->width = w;
  dest->height = ht;
  dest->luma_stride = w;
  dest->colorGamut = UHDRB200_COLORGAMUT_UNSPECIFIED;
  if (h.n_comp == 1) {
    dest->pixelFormat = UHDRB200_PIX_FMT_MONOCHROME;
    dest->chroma_data = nullptr;
    dest->chroma_stride = 0;
  } else {
    dest->pixelFormat = UHDRB200_PIX_FMT_YUV420;
    dest->chroma_data = (uint8_t*)dest->data + w * ht;
    dest->chroma_stride = cw;
  }
  return UHDRB200_NO_ERROR;
}

// ---------------------------------------------------------------- N4: baseline-JPEG encode (jpeg_enc.cu)
// JpegEncoderHelper::compressImage(yBuffer, uvBuffer, width, height, lumaStride, chromaStride, quality, iccBuffer, iccSize)
// + getCompressedImagePtr / Size (lib/src/jpegencoderhelper.cpp:35-58, 86-130): uvBuffer == nullptr means greyscale.
int jpeg_encode(const uhdrb200_image* img, int quality, const void* icc, size_t icc_bytes, void* out, size_t cap, size_t* out_bytes,
                bool planes_on_device, cudaStream_t user_stream) {
  if (!img || !img->data || !out || !out_bytes) return UHDRB200_ERROR_BAD_PTR;
  *out_bytes = 0;
  const bool grey = img->chroma_data == nullptr || img->pixelFormat == UHDRB200_PIX_FMT_MONOCHROME;
  const size_t w = img->width, ht = img->height;
  if (w == 0 || ht == 0 || w > 65500 || ht > 65500 || (!grey && (w < 2 || ht < 2))) return UHDRB200_ERROR_UNSUPPORTED_WIDTH_HEIGHT;  // JPEG_MAX_DIMENSION
  if (quality < 0 || quality > 100) return UHDRB200_ERROR_INVALID_QUALITY_FACTOR;
  const size_t ls = img->luma_stride ? img->luma_stride : w, cs = grey ? 0 : (img->chroma_stride ? img->chroma_stride : w / 2);
  if (ls < w || (!grey && cs < w / 2)) return UHDRB200_ERROR_INVALID_STRIDE;
  int st = ensure_ctx();
  if (st) return st;
  Ctx& c = t_ctx;
  cudaStream_t s = planes_on_device ? user_stream : c.stream;  // _dev: the caller's stream (0 = the default stream), ordered after the producer of the planes
  static thread_local JpegEncTables tabs;
  jpeg_enc_tables(quality, grey, &tabs);
  // ---- geometry: MCUs of 16 x 16 (4:2:0) or 8 x 8 (grey) samples
  JpegScan sc;
  memset(&sc, 0, sizeof sc);
  sc.n_comp = grey ? 1 : 3;
  const uint32_t mcu = grey ? 8 : 16;
  sc.mcus_x = (uint32_t)((w + mcu - 1) / mcu);
  sc.mcus_y = (uint32_t)((ht + mcu - 1) / mcu);
  int nb = 0;
  for (int ci = 0; ci < sc.n_comp; ci++) {
    const int hv = (!grey && ci == 0) ? 2 : 1;
    sc.comp_h[ci] = sc.comp_v[ci] = (uint8_t)hv;
    for (int y = 0; y < hv; y++)
      for (int x = 0; x < hv; x++, nb++) {
        sc.blk_comp[nb] = (uint8_t)ci;
        sc.blk_dx[nb] = (uint8_t)x;
        sc.blk_dy[nb] = (uint8_t)y;
      }
    sc.comp_blocks_w[ci] = sc.mcus_x * hv;
    sc.comp_blocks_h[ci] = sc.mcus_y * hv;
  }
  sc.blocks_per_mcu = nb;
  sc.total_blocks = sc.mcus_x * sc.mcus_y * (uint32_t)nb;
  size_t coeff_off[3] = {0, 0, 0}, coeff_bytes = 0;
  for (int ci = 0; ci < sc.n_comp; ci++) {
    coeff_off[ci] = coeff_bytes;
    coeff_bytes += (size_t)sc.comp_blocks_w[ci] * sc.comp_blocks_h[ci] * 64 * sizeof(int16_t);
  }
  // ---- planes: the caller's device memory, or staged copies of the host planes (U rows then V rows at chroma_stride * (h / 2))
  const uint8_t* d_y = (const uint8_t*)img->data;
  const uint8_t* d_u = (const uint8_t*)img->chroma_data;
  const size_t cw = w / 2, chh = ht / 2;
  if (!planes_on_device) {
    if ((st = upload(c.buf[0], img->data, ls * (ht - 1) + w, s))) return st;
    d_y = (const uint8_t*)c.buf[0].p;
    if (!grey) {
      if ((st = upload(c.buf[1], img->chroma_data, cs * (chh * 2 - 1) + cw, s))) return st;
      d_u = (const uint8_t*)c.buf[1].p;
    }
  }
  if ((st = upload(c.buf[5], &tabs, sizeof tabs, s))) return st;
  if (s != c.stream) CU(cudaStreamSynchronize(c.stream));  // the uploads ran on the thread's own stream: done before `s` reads them
  const JpegEncTables* d_tabs = (const JpegEncTables*)c.buf[5].p;
  if ((st = reserve(c.buf[3], coeff_bytes))) return st;
  if ((st = reserve(c.buf[2], jpeg_enc_workspace_bytes(sc.total_blocks)))) return st;
  JpegEncParams ep;
  memset(&ep, 0, sizeof ep);
  ep.n_components = sc.n_comp;
  for (int ci = 0; ci < sc.n_comp; ci++) {
    sc.coeffs[ci] = (int16_t*)((uint8_t*)c.buf[3].p + coeff_off[ci]);
    JpegEncComponent& e = ep.comp[ci];
    e.plane = ci == 0 ? d_y : d_u + (size_t)(ci - 1) * cs * chh;  // compressYuv, jpegencoderhelper.cpp:134-138
    e.width = ci == 0 ? w : cw;
    e.height = ci == 0 ? ht : chh;
    e.stride = ci == 0 ? ls : cs;
    e.blocks_w = sc.comp_blocks_w[ci];
    e.blocks_h = sc.comp_blocks_h[ci];
    e.quant = d_tabs->quant[ci == 0 ? 0 : 1];
    e.coeffs = sc.coeffs[ci];
  }
  cudaError_t e = launch_jpeg_fdct_impl(ep, s);
  if (e != cudaSuccess) return fail(e, "jpeg fdct kernel launch");
  uint32_t* d_total = (uint32_t*)((uint8_t*)c.buf[2].p + jpeg_enc_workspace_bytes(sc.total_blocks) - 64);
  uint32_t* d_offsets = nullptr;
  if ((e = jpeg_huff_encode(sc, d_tabs, c.buf[2].p, d_total, s, &d_offsets)) != cudaSuccess) return fail(e, "jpeg huffman sizing");
  uint32_t total_bits = 0;
  CU(cudaMemcpyAsync(&total_bits, d_total, 4, cudaMemcpyDeviceToHost, s));
  CU(cudaStreamSynchronize(s));
  const size_t nbytes = ((size_t)total_bits + 7) / 8, nwords = (nbytes + 3) / 4 + 1;
  if ((st = reserve(c.buf[4], nwords * 4))) return st;
  CU(cudaMemsetAsync(c.buf[4].p, 0, nwords * 4, s));
  if ((e = jpeg_huff_emit(sc, d_tabs, d_offsets, (uint32_t*)c.buf[4].p, s)) != cudaSuccess) return fail(e, "jpeg huffman emit");
  g_launches += 6;
  static thread_local std::vector<uint8_t> raw;
  raw.resize(nbytes + 8);
  if (s != c.stream) CU(cudaStreamSynchronize(s));  // the copy below runs on the thread's own stream
  if ((st = copy_d2h(c, raw.data(), c.buf[4].p, nbytes))) return st;
  CU(join_stream(c));
  if (total_bits & 7u) raw[nbytes - 1] |= (uint8_t)(0xFFu >> (total_bits & 7u));  // flush_bits: the last byte is filled with ones
  uint8_t* o = (uint8_t*)out;
  size_t n = jpeg_write_headers(o, cap, (int)w, (int)ht, grey, tabs, icc, icc_bytes);
  if (n == 0) return icc_bytes > 65533 ? UHDRB200_ERROR_UNSUPPORTED_FEATURE : UHDRB200_ERROR_BUFFER_TOO_SMALL;
  const size_t body = jpeg_stuff(raw.data(), nbytes, o + n, cap - n);
  if ((body == 0 && nbytes != 0) || n + body + 2 > cap) return UHDRB200_ERROR_BUFFER_TOO_SMALL;
  n += body;
  o[n++] = 0xFF;
  o[n++] = 0xD9;
  *out_bytes = n;
  return UHDRB200_NO_ERROR;
}

// the codec entry points size host vectors from the caller's data: an allocation failure must not unwind through the C ABI
template <class F>
int no_throw(F f) {
  try {
    return f();
  } catch (const std::exception& e) {
    t_ctx.err = std::string("uhdrb200_jpeg: ") + e.what();
    return UHDRB200_ERROR_INSUFFICIENT_RESOURCE;
  }
}

int jpeg_get_info(const void* jpeg, size_t bytes, size_t* width, size_t* height, int* pixel_format) {
  if (!jpeg) return UHDRB200_ERROR_BAD_PTR;
  static thread_local JpegHeader h;
  const char* why = "";
  const int parsed = jpeg_parse((const uint8_t*)jpeg, bytes, &h, &why);
  int st = jpeg_status(parsed, why);
  if (st) return st;
  if (width) *width = (size_t)h.width;
  if (height) *height = (size_t)h.height;
  if (pixel_format) *pixel_format = h.n_comp == 1 ? UHDRB200_PIX_FMT_MONOCHROME : UHDRB200_PIX_FMT_YUV420;
  return jpeg_check_layout(h);
}

}  // namespace

// ================================================================ extern "C"
extern "C" {

int uhdrb200_abi_version(void) { return UHDRB200_ABI_VERSION; }
int uhdrb200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}
int uhdrb200_init(int device) {
  Ctx& c = t_ctx;
  if (c.ready && c.device != device) c.release();  // rebind: everything the old device owns goes (stream, events, pins, buffers, ring)
  c.device = device;
  return ensure_ctx();
}
void uhdrb200_shutdown(void) { t_ctx.release(); }
const char* uhdrb200_last_error(void) { return t_ctx.err.c_str(); }
void* uhdrb200_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (ensure_ctx()) return nullptr;
  if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}
void uhdrb200_host_free(void* p) { if (p) cudaFreeHost(p); }
void uhdrb200_free(void* p) { free(p); }
uint64_t uhdrb200_launch_count(void) { return g_launches.load(); }
int uhdrb200_test_set_apply_mode(int mode) {
  if (mode < UHDRB200_APPLY_AUTO || mode > UHDRB200_APPLY_Q_PIPE) return UHDRB200_ERROR_UNSUPPORTED_FEATURE;
  t_apply_mode = mode;
  return UHDRB200_NO_ERROR;
}
int uhdrb200_test_eval(int fn, const float* in_dev, float* out_dev, size_t n, void* stream) {
  if (fn < 0 || fn >= UHDRB200_FN_COUNT) return UHDRB200_ERROR_UNSUPPORTED_FEATURE;
  if (n && (!in_dev || !out_dev)) return UHDRB200_ERROR_BAD_PTR;
  int dev = 0;
  int st = core_device(&dev);
  if (st) return st;
  ApplyTables tab;
  if ((st = get_apply_tables(dev, &tab))) return st;
  const GainCode gc = gain_code_of(1.0f, 1000.0f / 203.0f);  // what generateGainMap writes for HLG (ultrahdr.cpp:250-257)
  cudaError_t e = launch_kat_impl(fn, in_dev, out_dev, n, tab, gc, (cudaStream_t)stream);
  if (e != cudaSuccess) return fail(e, "test_eval kernel launch");
  return UHDRB200_NO_ERROR;
}
size_t uhdrb200_test_jpeg_headers(int quality, int grey, int width, int height, const void* icc, size_t icc_bytes, void* out, size_t cap) {
  JpegEncTables t;
  jpeg_enc_tables(quality, grey != 0, &t);
  return jpeg_write_headers((uint8_t*)out, cap, width, height, grey != 0, t, icc, icc_bytes);
}
size_t uhdrb200_test_jpeg_stuff(const void* src, size_t n, void* out, size_t cap) { return jpeg_stuff((const uint8_t*)src, n, (uint8_t*)out, cap); }
size_t uhdrb200_test_jpeg_destuff(const void* src, size_t n, void* out, int* marker, unsigned max_intervals, unsigned* interval_start,
                                  unsigned* n_intervals) {
  if (!src || !out || !marker || !interval_start || !n_intervals || max_intervals == 0) return 0;
  return jpeg_destuff((const uint8_t*)src, 0, n, (uint8_t*)out, marker, max_intervals, interval_start, n_intervals);
}
size_t uhdrb200_test_jpeg_subseq_bytes(void) { return kJpegSubseqBits / 8; }
int uhdrb200_set_numeric_mode(int mode) {
  if (mode != UHDRB200_MODE_ANALYTIC && mode != UHDRB200_MODE_LUT) return UHDRB200_ERROR_UNSUPPORTED_FEATURE;
  t_numeric_mode = mode;
  return UHDRB200_NO_ERROR;
}
int uhdrb200_get_numeric_mode(void) { return t_numeric_mode; }
int uhdrb200_set_sync_mode(int mode) {
  if (mode != UHDRB200_SYNC_SPIN && mode != UHDRB200_SYNC_BLOCKING) return UHDRB200_ERROR_UNSUPPORTED_FEATURE;
  g_sync_mode = mode;
  return UHDRB200_NO_ERROR;
}

int uhdrb200_generate_gainmap_host(const uhdrb200_image* yuv420, const uhdrb200_image* p010, int hdr_tf,
                                   uhdrb200_metadata* metadata, uhdrb200_image* dest, int sdr_is_601) {
  return generate_host(yuv420, p010, hdr_tf, metadata, dest, sdr_is_601, nullptr, 0, true);
}
int uhdrb200_generate_gainmap_host_into(const uhdrb200_image* yuv420, const uhdrb200_image* p010, int hdr_tf,
                                        uhdrb200_metadata* metadata, uhdrb200_image* dest, int sdr_is_601,
                                        void* map_buf, size_t map_buf_bytes) {
  return generate_host(yuv420, p010, hdr_tf, metadata, dest, sdr_is_601, map_buf, map_buf_bytes, false);
}
int uhdrb200_generate_gainmap_dev(const uhdrb200_image* yuv420, const uhdrb200_image* p010, int hdr_tf,
                                  uhdrb200_metadata* metadata, uhdrb200_image* dest, int sdr_is_601, void* stream) {
  return generate_core(1, yuv420, p010, hdr_tf, metadata, dest, sdr_is_601, (cudaStream_t)stream);
}
int uhdrb200_generate_gainmap_batch_dev(int n, const uhdrb200_image* yuv420, const uhdrb200_image* p010, int hdr_tf,
                                        uhdrb200_metadata* metadata, uhdrb200_image* dest, int sdr_is_601,
                                        void* stream) {
  return generate_core(n, yuv420, p010, hdr_tf, metadata, dest, sdr_is_601, (cudaStream_t)stream);
}
int uhdrb200_generate_gainmap_batch_dev_noalloc(int n, const uhdrb200_image* yuv420, const uhdrb200_image* p010, int hdr_tf,
                                                uhdrb200_metadata* metadata, uhdrb200_image* dest, int sdr_is_601, void* stream) {
  return generate_noalloc(n, yuv420, p010, hdr_tf, metadata, dest, sdr_is_601, (cudaStream_t)stream);
}
int uhdrb200_apply_gainmap_host(const uhdrb200_image* yuv420, const uhdrb200_image* gainmap,
                                const uhdrb200_metadata* metadata, int output_format, float max_display_boost,
                                uhdrb200_image* dest) {
  return apply_host(yuv420, gainmap, metadata, output_format, max_display_boost, dest);
}
int uhdrb200_apply_gainmap_dev(const uhdrb200_image* yuv420, const uhdrb200_image* gainmap,
                               const uhdrb200_metadata* metadata, int output_format, float max_display_boost,
                               uhdrb200_image* dest, void* stream) {
  return apply_core(1, yuv420, gainmap, metadata, output_format, max_display_boost, dest, (cudaStream_t)stream);
}
int uhdrb200_apply_gainmap_batch_dev(int n, const uhdrb200_image* yuv420, const uhdrb200_image* gainmap,
                                     const uhdrb200_metadata* metadata, int output_format, float max_display_boost,
                                     uhdrb200_image* dest, void* stream) {
  return apply_core(n, yuv420, gainmap, metadata, output_format, max_display_boost, dest, (cudaStream_t)stream);
}
int uhdrb200_tonemap_host(const uhdrb200_image* src, uhdrb200_image* dest) { return tonemap_host(src, dest); }
int uhdrb200_tonemap_dev(const uhdrb200_image* src, uhdrb200_image* dest, void* stream) {
  return tonemap_core(1, src, dest, (cudaStream_t)stream);
}
int uhdrb200_tonemap_batch_dev(int n, const uhdrb200_image* src, uhdrb200_image* dest, void* stream) {
  return tonemap_core(n, src, dest, (cudaStream_t)stream);
}
int uhdrb200_tonemap_generate_gainmap_host_into(const uhdrb200_image* p010, int hdr_tf, int sdr_to_601, uhdrb200_image* sdr_dest,
                                                uhdrb200_metadata* metadata, uhdrb200_image* map_dest, void* map_buf,
                                                size_t map_buf_bytes) {
  return api0_host(p010, hdr_tf, sdr_to_601, sdr_dest, metadata, map_dest, map_buf, map_buf_bytes);
}
int uhdrb200_tonemap_generate_gainmap_dev(const uhdrb200_image* p010, int hdr_tf, int sdr_to_601, uhdrb200_image* sdr_dest,
                                          uhdrb200_metadata* metadata, uhdrb200_image* map_dest, void* stream) {
  return api0_core(1, p010, hdr_tf, sdr_to_601, sdr_dest, metadata, map_dest, (cudaStream_t)stream);
}
int uhdrb200_tonemap_generate_gainmap_batch_dev(int n, const uhdrb200_image* p010, int hdr_tf, int sdr_to_601,
                                                uhdrb200_image* sdr_dest, uhdrb200_metadata* metadata, uhdrb200_image* map_dest,
                                                void* stream) {
  return api0_core(n, p010, hdr_tf, sdr_to_601, sdr_dest, metadata, map_dest, (cudaStream_t)stream);
}
int uhdrb200_convert_yuv_host(uhdrb200_image* image, int src_encoding, int dest_encoding) {
  return convert_yuv_host(image, src_encoding, dest_encoding);
}
int uhdrb200_convert_yuv_dev(uhdrb200_image* image, int src_encoding, int dest_encoding, void* stream) {
  return convert_yuv_core(image, src_encoding, dest_encoding, (cudaStream_t)stream);
}

int uhdrb200_yuv420_to_rgba8888_host(const uhdrb200_image* yuv420, uhdrb200_image* dest) { return sdr_pack_host(yuv420, dest); }
int uhdrb200_yuv420_to_rgba8888_dev(const uhdrb200_image* yuv420, uhdrb200_image* dest, void* stream) {
  return sdr_pack_core(1, yuv420, dest, (cudaStream_t)stream);
}
int uhdrb200_yuv420_to_rgba8888_batch_dev(int n, const uhdrb200_image* yuv420, uhdrb200_image* dest, void* stream) {
  return sdr_pack_core(n, yuv420, dest, (cudaStream_t)stream);
}
int uhdrb200_jpeg_idct_host(int n_components, const uhdrb200_jpeg_component* comps) { return jpeg_idct_host(n_components, comps); }
int uhdrb200_jpeg_idct_dev(int n_components, const uhdrb200_jpeg_component* comps, void* stream) {
  return jpeg_idct_core(n_components, comps, (cudaStream_t)stream);
}
int uhdrb200_jpeg_get_info(const void* jpeg, size_t bytes, size_t* width, size_t* height, int* pixel_format) {
  return jpeg_get_info(jpeg, bytes, width, height, pixel_format);
}
int uhdrb200_jpeg_decode_host(const void* jpeg, size_t bytes, uhdrb200_image* dest) {
  return no_throw([&] { return jpeg_decode(jpeg, bytes, dest, false, nullptr); });
}
int uhdrb200_jpeg_decode_dev(const void* jpeg, size_t bytes, uhdrb200_image* dest, void* stream) {
  return no_throw([&] { return jpeg_decode(jpeg, bytes, dest, true, (cudaStream_t)stream); });
}
int uhdrb200_jpeg_encode_host(const uhdrb200_image* image, int quality, const void* icc, size_t icc_bytes, void* out, size_t out_capacity,
                              size_t* out_bytes) {
  return no_throw([&] { return jpeg_encode(image, quality, icc, icc_bytes, out, out_capacity, out_bytes, false, nullptr); });
}
int uhdrb200_jpeg_encode_dev(const uhdrb200_image* image, int quality, const void* icc, size_t icc_bytes, void* out, size_t out_capacity,
                             size_t* out_bytes, void* stream) {
  return no_throw([&] { return jpeg_encode(image, quality, icc, icc_bytes, out, out_capacity, out_bytes, true, (cudaStream_t)stream); });
}
int uhdrb200_copy_yuv420_host(const uhdrb200_image* src, uhdrb200_image* dest) { return copy_yuv420_host(src, dest); }
int uhdrb200_copy_yuv420_dev(const uhdrb200_image* src, uhdrb200_image* dest, void* stream) {
  return copy_yuv420_core(src, dest, (cudaStream_t)stream);
}

}  // extern "C"
