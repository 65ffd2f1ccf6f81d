// runtime.cu — host side of libtritonhash.so: contexts, pinned rings, streams, tickets.
//
// What it stands in for: nothing in tritonmedia/downloader allocates or schedules for hashing —
// [dep] webtorrent hands a piece Buffer to sha1(buf, cb) and [dep] minio pipes a part range through
// getHashSummer, both synchronously on Node's one event-loop thread.  Here a
// batch of such buffers becomes: pinned ring slot -> one cudaMemcpyAsync H2D -> one digest kernel per
// requested algorithm (each on its own stream, so MD5 and SHA-256 of the same parts overlap) ->
// cudaMemcpyAsync D2H of the 16..32-byte digests.  Everything is enqueued by th_submit and never
// blocks; rings are double-buffered simply by holding two or more slots in flight.
//
// Threading: every GPU of a context has its own lock, rings, streams, job buffers, tickets and
// statistics, so threads feeding different GPUs never meet; ticket and slot ids carry their GPU's index.  One
// library thread per context (the "completer") delivers completion callbacks with no lock held and doubles as
// the watchdog that reports a faulted device to callers who would otherwise wait for a callback for ever.
//
// There is no CPU fallback: every path ends in a kernel of kernels.cu or fails with an error code.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <numeric>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/tritonhash.h"

namespace {

constexpr size_t kRingAlign = 4096;  // slots start on a page: O_DIRECT reads can land in them (reader.py)
constexpr int kGpuBits = 6;  // ids = (sequence << 6) | index of the GPU in its context
constexpr int kMaxGpus = 1 << kGpuBits;

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Per-submit staging: one pinned block + one device block, carved identically.
//   [offsets n*8][lengths n*8][order n*4][expected n*20][sha1 n*20][md5 n*16][sha256 n*32][ok words]
struct JobBuf {
  size_t cap_n = 0;
  uint8_t* h = nullptr;
  uint8_t* d = nullptr;
  size_t bytes = 0;
  bool from_arena = false;     // carved out of the per-GPU staging arena (no CUDA allocation on the submit path)
  size_t arena_off = 0;
  static size_t bytes_for(size_t n) { return align_up(n * (8 + 8 + 4 + 20 + 20 + 16 + 32) + ((n + 31) / 32) * 4 + 9 * 256, 256); }
  size_t o_off(size_t) const { return 0; }
  size_t o_len(size_t n) const { return align_up(n * 8, 256); }
  size_t o_ord(size_t n) const { return o_len(n) + align_up(n * 8, 256); }
  size_t o_exp(size_t n) const { return o_ord(n) + align_up(n * 4, 256); }
  size_t o_sha1(size_t n) const { return o_exp(n) + align_up(n * 20, 256); }
  size_t o_md5(size_t n) const { return o_sha1(n) + align_up(n * 20, 256); }
  size_t o_sha256(size_t n) const { return o_md5(n) + align_up(n * 16, 256); }
  size_t o_ok(size_t n) const { return o_sha256(n) + align_up(n * 32, 256); }
};

struct Slot {
  uint64_t id = 0;
  int gpu = 0;
  size_t off = 0, bytes = 0;   // region of the pinned ring
  size_t doff = 0;             // region of the device ring (same length)
  size_t used = 0;             // bytes the caller asked for (what the H2D copy moves)
  bool submitted = false;
  bool host_free = false;      // H2D copy completed (or slot released) -> pinned region reusable
  bool done = false;           // every ticket completed (or slot released) -> device region reusable
  std::vector<uint64_t> tickets;  // a slot may be submitted several times (pieces AND parts of the same bytes)
  cudaEvent_t ev_copied = nullptr;  // H2D of the slot, issued by its first submit
};

struct Ticket {
  uint64_t id = 0;
  int gpu = 0;
  size_t n = 0;
  uint32_t mask = 0;
  bool has_expected = false;
  bool finished = false;       // every tail's `done` event observed
  bool timed = false;
  int status = TH_OK;
  JobBuf* jb = nullptr;
  uint64_t slot_id = 0;        // 0 = device-resident submit
  // One entry per launch item: the end of its kernel (timing) and of its digest copy-back.  A ticket is joined
  // on the HOST over these events — never by one stream waiting for another: a stream parked on an event of a
  // 0.25..1 s chain also parks every stream that shares its hardware queue (120 streams share 32 queues).
  struct Tail { cudaEvent_t k_end; cudaEvent_t done; int alg; bool also_md5; };  // also_md5: a fused MD5 + SHA-256 launch
  std::vector<Tail> tails;
  std::atomic<int> cb_left{0};  // launch items that have not reached their host function yet
  cudaEvent_t ev_ready = nullptr;  // inputs (payload + descriptors) on the device; timing-enabled when timing is on
  bool ev_ready_pooled = false;
  double times_ms[7] = {-1, -1, -1, -1, -1, -1, -1};  // since th_init: ready, then start / end per digest
  cudaEvent_t ev_k0[3] = {nullptr, nullptr, nullptr};  // start of each digest's first launch item
  th_done_cb cb = nullptr;
  void* user = nullptr;
  int cb_state = 0;            // 0 not delivered yet, 1 running on the completer thread, 2 returned
  struct th_ctx* ctx = nullptr;
};

// Kernels of different submits must be able to overlap: a slot of large parts holds only a few message
// groups, and a chain's latency (0.26 s for a 16 MiB SHA-256 part) is far longer than the slot's H2D copy.
// Tickets therefore rotate over a pool of stream sets; only the H2D copies share one stream (one PCIe link).
constexpr int kStreamSets = 40;      // default: 120 digest streams (the GPU runs up to 128 kernels at once)
constexpr int kMaxStreamSets = 40;   // TRITONHASH_STREAM_SETS narrows the rotation
constexpr int kTimedStreamSets = 10; // with ticket timing on: 30 streams + copy + misc = the 32 hardware queues
// Streams beyond the 32 hardware work queues (CUDA_DEVICE_MAX_CONNECTIONS, set in th_init when CUDA is not
// initialised yet) share queues; that is harmless as long as nothing queued behind a long kernel waits for it.
struct StreamSet {
  cudaStream_t alg[3] = {nullptr, nullptr, nullptr};
};

struct Gpu {
  int dev = -1;
  int index = 0;               // position in th_ctx::gpus (the low bits of this GPU's ticket and slot ids)
  std::mutex mu;               // guards everything below; threads feeding different GPUs never contend
  std::unordered_map<uint64_t, Ticket*> tickets;
  cudaStream_t s_copy = nullptr, s_misc = nullptr;  // H2D copies; odd jobs (ETag combine)
  cudaEvent_t ev_epoch = nullptr;                   // recorded in th_init: origin of th_ticket_timing
  StreamSet sets[kMaxStreamSets];
  int n_sets = kStreamSets;     // TRITONHASH_STREAM_SETS (1..40) overrides
  uint64_t next_stream[3] = {0, 0, 0};  // each digest rotates over its own streams, whatever mix of submits arrives
  // Two FIFO rings with separate lifetimes: a slot's pinned bytes are needed only until its H2D copy has run,
  // its device bytes until the last kernel reading them has finished.  A chain over a 16 MiB part runs for
  // ~0.25 s, several times longer than the copy of the whole slot, so the device ring may be made (much) larger
  // than the pinned one (th_init_staged): HBM, not pinned host memory, holds the bytes in flight.
  uint8_t* ring_h = nullptr;
  uint8_t* ring_d = nullptr;
  size_t ring_bytes = 0, dring_bytes = 0;
  size_t head = 0, tail = 0;   // pinned ring allocator: live region is [head, tail) modulo wrap
  size_t dhead = 0, dtail = 0; // device ring allocator, same scheme
  std::deque<Slot> slots;      // FIFO, oldest first; a slot leaves when its device region is free
  std::vector<JobBuf*> free_jobs;
  // Staging arena: one pinned + one device block allocated in th_init, carved first-fit into job buffers.
  // cudaMalloc / cudaHostAlloc on the submit path would wait for every running kernel (they synchronise the
  // device), which serialises exactly the long kernels that must overlap; they remain only as the overflow path
  // (counted in th_stats.overflow_allocs).
  uint8_t* arena_h = nullptr;
  uint8_t* arena_d = nullptr;
  size_t arena_bytes = 0;
  std::map<size_t, size_t> arena_free;  // offset -> length, coalesced
  std::vector<cudaEvent_t> ev_pool;     // untimed events, recycled: a submit records one per launch item
  th_stats st{};
};

}  // namespace

struct th_ctx {
  std::vector<std::unique_ptr<Gpu>> gpus;
  std::atomic<uint64_t> next_seq{1};
  std::mutex err_mu;
  char last_error[512] = {0};
  std::atomic<int> sticky{TH_OK};
  // Per-ticket device timeline (th_ticket_timing, th_stats.kernel_ms).  OFF by default: a timing-enabled event
  // recorded behind a kernel holds its hardware queue until the kernel ends (measured, tools/microbench/queues.cu:
  // 128 streams of 50 ms kernels take 50 ms with plain events and 200 ms with timing events), which caps the
  // long chains of large parts at 32 at a time.  When on, the rotation keeps to 10 stream sets (one queue each).
  std::atomic<bool> timing{false};
  // Completion delivery.  The CUDA host function of a ticket's last launch item only queues the ticket id here;
  // the completer thread calls the user's callback with no library lock held, so the callback may call th_*.
  std::thread completer;
  std::thread::id completer_id;
  std::mutex cq_mu;
  std::condition_variable cq_cv;     // work for the completer
  std::condition_variable cb_cv;     // a callback has returned (th_result waits for its ticket's)
  std::deque<uint64_t> cq_ready;
  std::atomic<int> cb_outstanding{0};  // tickets submitted with a callback that has not been delivered
  bool stop = false;
};

namespace {

int cuda_fail(th_ctx* c, cudaError_t e, const char* what) {
  {
    std::lock_guard<std::mutex> lk(c->err_mu);
    snprintf(c->last_error, sizeof(c->last_error), "%s: %s", what, cudaGetErrorString(e));
  }
  c->sticky.store(TH_ECUDA);
  return TH_ECUDA;
}
#define CU(ctx, call)                                             \
  do {                                                            \
    cudaError_t e__ = (call);                                     \
    if (e__ != cudaSuccess) return cuda_fail((ctx), e__, #call);  \
  } while (0)

inline int gpu_of_id(uint64_t id) { return (int)(id & (kMaxGpus - 1)); }
inline uint64_t new_id(th_ctx* c, const Gpu& g) { return (c->next_seq.fetch_add(1) << kGpuBits) | (uint64_t)g.index; }
Gpu* gpu_of(th_ctx* c, uint64_t id) {
  const int gi = gpu_of_id(id);
  return gi < (int)c->gpus.size() ? c->gpus[(size_t)gi].get() : nullptr;
}

cudaError_t get_event(Gpu& g, cudaEvent_t* ev) {
  if (!g.ev_pool.empty()) { *ev = g.ev_pool.back(); g.ev_pool.pop_back(); return cudaSuccess; }
  return cudaEventCreateWithFlags(ev, cudaEventDisableTiming);
}
void put_event(Gpu& g, cudaEvent_t ev) {
  if (!ev) return;
  if (g.ev_pool.size() < 4096) g.ev_pool.push_back(ev);
  else cudaEventDestroy(ev);
}

void put_job(Gpu& g, JobBuf* jb) {
  if (!jb->from_arena) { g.free_jobs.push_back(jb); return; }
  size_t off = jb->arena_off, len = jb->bytes;
  delete jb;
  auto next = g.arena_free.lower_bound(off);
  if (next != g.arena_free.begin()) {
    auto prev = std::prev(next);
    if (prev->first + prev->second == off) { off = prev->first; len += prev->second; g.arena_free.erase(prev); }
  }
  if (next != g.arena_free.end() && off + len == next->first) { len += next->second; g.arena_free.erase(next); }
  g.arena_free[off] = len;
}

// `need` bytes of pinned + device staging (same offsets in both).  cap_n records how many messages the block was
// sized for (0 for odd jobs such as the ETag combine).
JobBuf* get_buf(th_ctx* c, Gpu& g, size_t need, size_t cap_n, int* rc) {
  for (auto it = g.arena_free.begin(); it != g.arena_free.end(); ++it) {
    if (it->second >= need) {
      JobBuf* jb = new JobBuf();
      jb->cap_n = cap_n; jb->bytes = need; jb->from_arena = true; jb->arena_off = it->first;
      jb->h = g.arena_h + it->first; jb->d = g.arena_d + it->first;
      const size_t rest = it->second - need, off = it->first;
      g.arena_free.erase(it);
      if (rest) g.arena_free[off + need] = rest;
      return jb;
    }
  }
  for (size_t i = 0; i < g.free_jobs.size(); i++) {
    if (g.free_jobs[i]->bytes >= need) {
      JobBuf* jb = g.free_jobs[i];
      g.free_jobs.erase(g.free_jobs.begin() + (long)i);
      jb->cap_n = cap_n;
      return jb;
    }
  }
  // Overflow: the arena is full (very many small messages in flight).  These allocations synchronise the device,
  // which stalls every running chain — callers can see that it happened in th_stats.overflow_allocs.
  JobBuf* jb = new JobBuf();
  jb->cap_n = cap_n;
  jb->bytes = std::max(need, JobBuf::bytes_for(1024));
  cudaError_t e = cudaHostAlloc((void**)&jb->h, jb->bytes, cudaHostAllocPortable);
  if (e == cudaSuccess) e = cudaMalloc((void**)&jb->d, jb->bytes);
  if (e != cudaSuccess) {
    if (jb->h) cudaFreeHost(jb->h);
    delete jb;
    if (e == cudaErrorMemoryAllocation) { cudaGetLastError(); *rc = TH_ENOMEM; }
    else *rc = cuda_fail(c, e, "job buffer allocation");
    return nullptr;
  }
  g.st.overflow_allocs += 1;
  return jb;
}

JobBuf* get_job(th_ctx* c, Gpu& g, size_t n, int* rc) { return get_buf(c, g, JobBuf::bytes_for(n), n, rc); }

// order[0..n) = indices 0..n-1 by descending 64-byte block count, stable: LSD radix sort, 3 passes of 11 bits over
// the complemented count (a comparison sort of 300 k lengths cost 10 ms on the submit path; this is < 2 ms).
void sort_by_blocks_desc(const uint64_t* len, size_t n, uint32_t* order) {
  std::vector<uint32_t> key(n), key2(n), tmp(n);
  for (size_t i = 0; i < n; i++) {
    key[i] = ~(uint32_t)std::min<uint64_t>(len[i] / 64, 0xffffffffu);
    order[i] = (uint32_t)i;
  }
  uint32_t* src = order;
  uint32_t* dst = tmp.data();
  uint32_t* ks = key.data();
  uint32_t* kd = key2.data();
  for (int shift = 0; shift < 33; shift += 11) {
    size_t count[2049] = {0};
    for (size_t i = 0; i < n; i++) count[((ks[i] >> shift) & 2047u) + 1]++;
    if (count[((ks[0] >> shift) & 2047u) + 1] == n) continue;  // every key has the same digit
    for (int b = 0; b < 2048; b++) count[b + 1] += count[b];
    for (size_t i = 0; i < n; i++) {
      const size_t p = count[(ks[i] >> shift) & 2047u]++;
      dst[p] = src[i];
      kd[p] = ks[i];
    }
    std::swap(src, dst);
    std::swap(ks, kd);
  }
  if (src != order) std::memcpy(order, src, n * sizeof(uint32_t));
}

// Oldest slot (in acquisition order) whose pinned bytes are still needed; nullptr when the pinned ring is idle.
Slot* oldest_host_busy(Gpu& g) {
  for (auto& s : g.slots)
    if (!s.host_free) return &s;
  return nullptr;
}

Slot* find_slot(Gpu& g, uint64_t id) {
  for (auto& s : g.slots)
    if (s.id == id) return &s;
  return nullptr;
}

// Where `need` bytes fit at the tail of a FIFO ring whose live region is [head, tail) modulo one wrap;
// (size_t)-1 when they do not.
size_t ring_fit(size_t cap, size_t head, size_t tail, bool idle, size_t need) {
  if (idle) return 0;
  if (tail > head) {                       // live region does not wrap
    if (cap - tail >= need) return tail;
    if (head >= need) return 0;
  } else if (tail < head) {                // wrapped: the free gap is [tail, head)
    if (head - tail >= need) return tail;
  }
  return (size_t)-1;
}

// cudaEventQuery / cudaEventSynchronize over every tail of a ticket.
cudaError_t ticket_poll(Ticket* t, bool block) {
  for (auto& tl : t->tails) {
    cudaError_t e = block ? cudaEventSynchronize(tl.done) : cudaEventQuery(tl.done);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

// Marks copied slots' pinned bytes and finished slots' device bytes reusable and advances both ring heads.
// Called with g.mu held.
void reap(th_ctx* c, Gpu& g, bool block_on_oldest) {
  for (auto& s : g.slots) {
    if (s.done || !s.submitted) continue;
    bool all = true;
    for (uint64_t id : s.tickets) {
      auto it = g.tickets.find(id);
      if (it == g.tickets.end()) continue;  // already retired by th_result
      Ticket* t = it->second;
      if (!t->finished) {
        cudaError_t e = ticket_poll(t, block_on_oldest && &s == &g.slots.front());
        if (e == cudaSuccess) t->finished = true;
        else if (e != cudaErrorNotReady) { t->finished = true; t->status = cuda_fail(c, e, "ticket event"); }
      }
      all &= t->finished;
    }
    if (all) s.done = s.host_free = true;
    if (!s.host_free && s.ev_copied && cudaEventQuery(s.ev_copied) == cudaSuccess) s.host_free = true;
  }
  while (!g.slots.empty() && g.slots.front().done) {
    put_event(g, g.slots.front().ev_copied);
    g.slots.pop_front();
    g.dhead = g.slots.empty() ? g.dtail : g.slots.front().doff;
  }
  if (g.slots.empty()) g.dhead = g.dtail = 0;
  const Slot* busy = oldest_host_busy(g);
  if (busy) g.head = busy->off;
  else g.head = g.tail = 0;
}

// Runs on CUDA's callback thread once per launch item.  It may not call CUDA and must not take long: the item that
// finishes last only hands the ticket's id to the completer thread.
void CUDART_CB host_item_done(void* p) {
  Ticket* t = static_cast<Ticket*>(p);
  if (t->cb_left.fetch_sub(1) == 1 && t->cb) {
    th_ctx* c = t->ctx;
    std::lock_guard<std::mutex> lk(c->cq_mu);
    c->cq_ready.push_back(t->id);
    c->cq_cv.notify_one();
  }
}

// Completer thread: delivers the callback of ticket `id` with `status` unless it has been delivered already.
void deliver(th_ctx* c, uint64_t id, int status) {
  Gpu* g = gpu_of(c, id);
  if (!g) return;
  th_done_cb cb = nullptr;
  void* user = nullptr;
  {
    std::lock_guard<std::mutex> lk(g->mu);
    auto it = g->tickets.find(id);
    if (it == g->tickets.end() || !it->second->cb || it->second->cb_state != 0) return;
    Ticket* t = it->second;
    t->cb_state = 1;
    if (status != TH_OK) { t->status = status; t->finished = true; }
    cb = t->cb;
    user = t->user;
  }
  cb(user, id, status);  // no lock held: the callback may call th_result / th_wait / th_submit ...
  {
    std::lock_guard<std::mutex> lk(g->mu);
    auto it = g->tickets.find(id);          // gone if the callback retired its own ticket
    if (it != g->tickets.end()) it->second->cb_state = 2;
  }
  c->cb_outstanding.fetch_sub(1);
  std::lock_guard<std::mutex> lk(c->cq_mu);
  c->cb_cv.notify_all();
}

// A faulted device never runs the host functions queued behind the fault, so a caller that only waits for its
// callback (the N-API shim's Promises) would hang.  Every 100 ms without a completion the completer looks at the
// events of the tickets that still owe a callback; an event in error state settles all of that GPU's with TH_ECUDA.
void watchdog(th_ctx* c) {
  for (auto& gp : c->gpus) {
    Gpu& g = *gp;
    std::vector<uint64_t> owed;
    bool fault = false;
    {
      std::lock_guard<std::mutex> lk(g.mu);
      cudaSetDevice(g.dev);
      for (auto& kv : g.tickets) {
        Ticket* t = kv.second;
        if (!t->cb || t->cb_state != 0) continue;
        owed.push_back(t->id);
        if (!fault) {
          const cudaError_t e = ticket_poll(t, false);
          if (e != cudaSuccess && e != cudaErrorNotReady) {
            cuda_fail(c, e, "device fault seen by the completion watchdog");
            fault = true;
          }
        }
      }
    }
    if (fault)
      for (uint64_t id : owed) deliver(c, id, TH_ECUDA);
  }
}

void completer_main(th_ctx* c) {
  std::unique_lock<std::mutex> lk(c->cq_mu);
  for (;;) {
    if (c->cq_ready.empty()) {
      if (c->stop) return;
      if (c->cb_outstanding.load() > 0) {
        if (c->cq_cv.wait_for(lk, std::chrono::milliseconds(100)) == std::cv_status::timeout && c->cq_ready.empty() && !c->stop) {
          lk.unlock();
          watchdog(c);
          lk.lock();
        }
      } else {
        c->cq_cv.wait(lk);
      }
      continue;
    }
    const uint64_t id = c->cq_ready.front();
    c->cq_ready.pop_front();
    lk.unlock();
    deliver(c, id, TH_OK);
    lk.lock();
  }
}

int enqueue(th_ctx* c, Gpu& g, Ticket* t, const uint8_t* d_base, Slot* slot, const th_msg* msgs, size_t n,
            uint32_t mask, const uint8_t* expected) {
  JobBuf* jb = t->jb;
  uint64_t* h_off = reinterpret_cast<uint64_t*>(jb->h + jb->o_off(n));
  uint64_t* h_len = reinterpret_cast<uint64_t*>(jb->h + jb->o_len(n));
  uint32_t* h_ord = reinterpret_cast<uint32_t*>(jb->h + jb->o_ord(n));
  uint64_t total = 0;
  for (size_t i = 0; i < n; i++) {
    h_off[i] = msgs[i].offset;
    h_len[i] = msgs[i].length;
    total += msgs[i].length;
  }
  // Equal-length prefix at a fixed stride (all pieces of a torrent / parts of an object but the last):
  // goes to the tensor-map kernels.  m is kept a multiple of 32 so the verify bitfield splits on a word.
  size_t m = 0;
  uint64_t ulen = msgs[0].length, ustride = n > 1 ? msgs[1].offset - msgs[0].offset : ((msgs[0].length + 15) & ~15ull);
  if (ulen >= 1024 && ulen < (1ull << 31) && (ustride & 15) == 0 && ustride >= ulen && (n == 1 || msgs[1].offset > msgs[0].offset)) {
    m = 1;
    while (m < n && msgs[m].length == ulen && msgs[m].offset == msgs[0].offset + m * ustride) m++;
    if (m < n && expected) m &= ~(size_t)31;  // keep the verify bitfield split on a word boundary
    if (m < n && m < 64 && m * ulen < (8u << 20)) m = 0;  // a short prefix of small messages is not worth a launch
  }
  const size_t rest = n - m;
  // Length bucketing for the remainder: a warp runs as long as its longest message, so lanes are dealt
  // messages in descending block-count order (stable).  Equal-length remainders keep the identity mapping.
  bool rest_uniform = true;
  for (size_t i = m; i < n; i++) rest_uniform &= (h_len[i] / 64 == h_len[m] / 64);
  const bool use_order = rest > 0 && !rest_uniform;
  if (use_order) {
    sort_by_blocks_desc(h_len + m, rest, h_ord + m);  // indices relative to the remainder
  }
  if (expected) std::memcpy(jb->h + jb->o_exp(n), expected, n * 20);

  const bool timing = c->timing.load();
  CU(c, cudaSetDevice(g.dev));
  // H2D: payload (the whole ring slot, once — later submits on the same slot reuse it) then descriptors.
  if (slot && !slot->ev_copied) {
    CU(c, cudaMemcpyAsync(g.ring_d + slot->doff, g.ring_h + slot->off, slot->used, cudaMemcpyHostToDevice, g.s_copy));
    g.st.h2d_bytes += slot->used;
    CU(c, get_event(g, &slot->ev_copied));
    CU(c, cudaEventRecord(slot->ev_copied, g.s_copy));
  }
  const size_t desc_bytes = expected ? jb->o_sha1(n) : (use_order ? jb->o_exp(n) : jb->o_ord(n));
  CU(c, cudaMemcpyAsync(jb->d, jb->h, desc_bytes, cudaMemcpyHostToDevice, g.s_copy));
  g.st.h2d_bytes += desc_bytes;
  cudaEvent_t ev_in = nullptr;
  if (timing) CU(c, cudaEventCreateWithFlags(&ev_in, cudaEventDefault));
  else CU(c, get_event(g, &ev_in));
  t->ev_ready = ev_in;
  t->ev_ready_pooled = !timing;
  CU(c, cudaEventRecord(ev_in, g.s_copy));

  const uint64_t* d_off = reinterpret_cast<const uint64_t*>(jb->d + jb->o_off(n)) + m;
  const uint64_t* d_len = reinterpret_cast<const uint64_t*>(jb->d + jb->o_len(n)) + m;
  const uint32_t* d_ord = use_order ? reinterpret_cast<const uint32_t*>(jb->d + jb->o_ord(n)) + m : nullptr;
  const size_t out_off[3] = {jb->o_sha1(n), jb->o_md5(n), jb->o_sha256(n)};
  const size_t dlen[3] = {20, 16, 32};
  const size_t ok_bytes = ((n + 31) / 32) * 4;
  const uint8_t* d_uni = d_base + msgs[0].offset;

  // Launch items: the equal-length prefix (tensor-map kernels), then the remainder.  A remainder of a few LARGE
  // messages (the short last piece / part, or a handful of whole-file bodies) is launched message by message as
  // "equal-length batches of one": each gets the TMA path and, for SHA-1 / SHA-256, the schedule-helper warp,
  // which halves its chain latency compared with the per-lane kernel.  Every item after a digest's first runs on
  // another stream of the rotation; each item copies back the digests it produced on its own stream.
  struct Item { size_t first, cnt; int kind; };  // kind 0 prefix, 1 lone, 2 per-lane remainder
  struct ItemList { std::vector<Item> items; size_t skip = 0; };
  auto build_items = [&](bool verify) {
    ItemList il;
    std::vector<Item>& items = il.items;
    if (m) items.push_back({0, m, 0});
    if (rest) {
      bool lone_ok = rest <= 64 && !(verify && rest > 1);
      for (size_t i = m; i < n && lone_ok; i++)
        lone_ok = msgs[i].length < (1ull << 31) && msgs[i].length >= (rest == 1 ? 1024u : (256u << 10));
      if (lone_ok) {
        for (size_t i = m; i < n; i++) items.push_back({i, 1, 1});
      } else {
        // Heavy-tailed remainders: the batch cannot finish before its longest message's chain does, and that chain
        // runs about twice as fast through the TMA / helper-warp path as in a lane of the per-lane kernel.  The
        // few outliers (>= 256 KiB and >= 8 x the median; they lead the descending order; at most 8) are peeled into lone
        // launches; the per-lane launch takes the rest of the order.
        size_t peel = 0;
        if (use_order && !verify) {
          const uint64_t median = h_len[m + h_ord[m + rest / 2]];
          static const size_t peel_max = [] { const char* e = getenv("TRITONHASH_PEEL_MAX"); return (size_t)(e ? std::max(0, atoi(e)) : 8); }();
          while (peel < peel_max && peel + 32 < rest) {
            const uint64_t l = h_len[m + h_ord[m + peel]];
            if (l < (256u << 10) || l < 8 * median || l >= (1ull << 31)) break;
            peel++;
          }
        }
        items.push_back({m, rest - peel, 2});
        for (size_t k = 0; k < peel; k++) items.push_back({m + h_ord[m + k], 1, 1});
        il.skip = peel;
      }
    }
    return il;
  };
  // the item list depends on the digest only through the fused verify (SHA-1 with an expected table)
  const bool sha1_verify = (mask & 1u) && expected;
  const ItemList plain = build_items(false);
  const ItemList verified = sha1_verify ? build_items(true) : ItemList{};
  // MD5 and SHA-256 of the same few large messages (parts of an upload) share ONE kernel: the helper warp of the
  // SHA-256 chain runs the MD5 chain too (kernels.cu, MD5_TOO).  Such an item is launched in the SHA-256 pass only.
  const bool both = (mask & TH_ALG_MD5) && (mask & TH_ALG_SHA256);
  auto item_len = [&](const Item& it) { return it.kind == 1 ? msgs[it.first].length : ulen; };
  auto item_stride = [&](const Item& it) { return it.kind == 1 ? ((msgs[it.first].length + 15) & ~15ull) : ustride; };
  auto fused = [&](const Item& it) {
    return both && it.kind != 2 && th_md5_sha256_fused(item_stride(it), item_len(it), (uint32_t)it.cnt) == 1;
  };
  const uint64_t sets = (uint64_t)(timing ? std::min(g.n_sets, kTimedStreamSets) : g.n_sets);
  int n_items = 0;
  for (int a = 0; a < 3; a++) {
    if (!(mask & (1u << a))) continue;
    for (const Item& it : ((a == 0 && sha1_verify) ? verified : plain).items)
      if (!(a == 1 && fused(it))) n_items++;
  }
  t->cb_left.store(n_items);
  t->tails.reserve((size_t)n_items);

  for (int a = 0; a < 3; a++) {
    if (!(mask & (1u << a))) continue;
    cudaStream_t st = g.sets[g.next_stream[a]++ % sets].alg[a];
    CU(c, cudaStreamWaitEvent(st, ev_in, 0));
    int rc = TH_OK;
    if (timing) CU(c, cudaEventCreate(&t->ev_k0[a]));
    uint8_t* d_out = jb->d + out_off[a];
    const uint8_t* d_exp = (a == 0 && expected) ? jb->d + jb->o_exp(n) : nullptr;
    uint32_t* d_ok = (a == 0 && expected) ? reinterpret_cast<uint32_t*>(jb->d + jb->o_ok(n)) : nullptr;
    if (d_ok && use_order) CU(c, cudaMemsetAsync(d_ok + m / 32, 0, ok_bytes - (m / 32) * 4, st));
    if (timing) CU(c, cudaEventRecord(t->ev_k0[a], st));
    const ItemList& il = (a == 0 && sha1_verify) ? verified : plain;
    const size_t ragged_skip = il.skip;  // leading entries of the order that are launched alone
    const std::vector<Item>& items = il.items;
    bool first_launch = true;
    for (size_t k = 0; k < items.size() && rc == TH_OK; k++) {
      const Item& it = items[k];
      if (a == 1 && fused(it)) continue;  // its MD5 comes out of the SHA-256 pass's fused launch
      const bool with_md5 = a == 2 && fused(it);
      cudaStream_t ps = st;
      if (!first_launch && !(it.kind == 2 && d_ok && use_order)) {  // (the bitfield memset above is ordered on st only)
        ps = g.sets[g.next_stream[a]++ % sets].alg[a];
        CU(c, cudaStreamWaitEvent(ps, ev_in, 0));
      }
      const bool uni = it.kind == 0, lone = it.kind == 1;
      const size_t first = it.first;
      const uint8_t* pb = uni ? d_uni : (lone ? d_base + msgs[first].offset : d_base);
      const uint64_t* po = it.kind == 2 ? d_off : nullptr;
      const uint64_t* pl = it.kind == 2 ? d_len : nullptr;
      const uint32_t* pr = it.kind == 2 && d_ord ? d_ord + ragged_skip : nullptr;
      first_launch = false;
      const uint64_t pstride = lone ? ((msgs[first].length + 15) & ~15ull) : ustride;
      const uint64_t plen = lone ? msgs[first].length : ulen;
      uint8_t* o = d_out + first * dlen[a];
      if (with_md5)
        rc = th_md5_sha256_batch(pb, pstride, plen, (uint32_t)it.cnt, jb->d + out_off[1] + first * 16, o, ps);
      else if (a == 0)
        rc = th_sha1_batch(pb, po, pl, pr, pstride, plen, (uint32_t)it.cnt, o, d_exp ? d_exp + first * 20 : nullptr,
                           d_ok ? d_ok + first / 32 : nullptr, ps);
      else if (a == 1)
        rc = th_md5_batch(pb, po, pl, pr, pstride, plen, (uint32_t)it.cnt, o, ps);
      else
        rc = th_sha256_batch(pb, po, pl, pr, pstride, plen, (uint32_t)it.cnt, o, ps);
      if (rc != TH_OK) break;
      g.st.launches += 1;
      Ticket::Tail tl{nullptr, nullptr, a, with_md5};
      if (timing) {
        CU(c, cudaEventCreate(&tl.k_end));
        CU(c, cudaEventRecord(tl.k_end, ps));
      }
      // Copy back what this item wrote.  The per-lane remainder writes at original indices all over [m, n) except
      // the peeled messages', which their own items copy: its range is cut around those (two copies of one host
      // byte from different streams must not be in flight together).
      const size_t span = it.kind == 2 ? rest : it.cnt;
      std::vector<size_t> holes;
      if (it.kind == 2)
        for (size_t q = 0; q < ragged_skip; q++) holes.push_back(m + h_ord[m + q]);
      std::sort(holes.begin(), holes.end());
      holes.push_back(first + span);
      size_t lo = first;
      for (size_t hole : holes) {
        if (hole > lo) {
          CU(c, cudaMemcpyAsync(jb->h + out_off[a] + lo * dlen[a], jb->d + out_off[a] + lo * dlen[a], (hole - lo) * dlen[a],
                                cudaMemcpyDeviceToHost, ps));
          g.st.d2h_bytes += (hole - lo) * dlen[a];
        }
        lo = hole + 1;
      }
      if (with_md5) {  // the MD5 digests the same launch produced
        CU(c, cudaMemcpyAsync(jb->h + out_off[1] + first * 16, jb->d + out_off[1] + first * 16, it.cnt * 16, cudaMemcpyDeviceToHost, ps));
        g.st.d2h_bytes += it.cnt * 16;
      }
      if (d_ok) {  // the bitfield words this item produced (items split on word boundaries)
        const size_t w0 = first / 32, w1 = (first + span + 31) / 32;
        CU(c, cudaMemcpyAsync(jb->h + jb->o_ok(n) + w0 * 4, jb->d + jb->o_ok(n) + w0 * 4, (w1 - w0) * 4, cudaMemcpyDeviceToHost, ps));
        g.st.d2h_bytes += (w1 - w0) * 4;
      }
      // the host function is enqueued before `done`: once a ticket's `done` events have completed, no host
      // function of it is still to run, so th_result may free it
      if (t->cb) CU(c, cudaLaunchHostFunc(ps, host_item_done, t));
      CU(c, get_event(g, &tl.done));
      t->tails.push_back(tl);  // (before the record, so that a failing record still leaves the event with its ticket)
      CU(c, cudaEventRecord(tl.done, ps));
    }
    if (rc != TH_OK) {
      if (rc == TH_ECUDA) return cuda_fail(c, cudaGetLastError(), "digest kernel launch");
      return rc;
    }
    g.st.messages += n;
  }
  g.st.bytes += total;
  g.st.submits += 1;
  return TH_OK;
}

int validate(const th_msg* msgs, size_t n, uint32_t mask, size_t limit) {
  if (!msgs && n) return TH_EINVAL;
  if (mask == 0 || (mask & ~7u)) return TH_EINVAL;
  if (n == 0 || n > 0x7fffffffu) return TH_EINVAL;
  for (size_t i = 0; i < n; i++) {
    if (msgs[i].offset & 15) return TH_EALIGN;
    if (msgs[i].length >= (1ull << 37)) return TH_ETOOBIG;
    if (limit && (msgs[i].offset > limit || msgs[i].length > limit - msgs[i].offset)) return TH_EINVAL;
  }
  return TH_OK;
}

void free_ticket_events(Gpu& g, Ticket* t) {
  for (auto& tl : t->tails) {
    if (tl.k_end) cudaEventDestroy(tl.k_end);
    put_event(g, tl.done);
  }
  t->tails.clear();
  if (t->ev_ready) {
    if (t->ev_ready_pooled) put_event(g, t->ev_ready);
    else cudaEventDestroy(t->ev_ready);
    t->ev_ready = nullptr;
  }
  for (int a = 0; a < 3; a++) {
    if (t->ev_k0[a]) cudaEventDestroy(t->ev_k0[a]);
    t->ev_k0[a] = nullptr;
  }
}

}  // namespace

extern "C" {

int th_init(const int* gpu_ids, int n_gpus, size_t ring_bytes_per_gpu, th_ctx** out) {
  return th_init_staged(gpu_ids, n_gpus, ring_bytes_per_gpu, 0, out);
}

int th_init_staged(const int* gpu_ids, int n_gpus, size_t ring_bytes_per_gpu, size_t device_bytes_per_gpu, th_ctx** out) {
  if (!out || n_gpus <= 0 || n_gpus > kMaxGpus) return TH_EINVAL;
  if (device_bytes_per_gpu == 0) device_bytes_per_gpu = ring_bytes_per_gpu;
  if (ring_bytes_per_gpu == 0) device_bytes_per_gpu = 0;
  *out = nullptr;
  setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);  // read when the CUDA context is created; no effect if it already exists
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return TH_ENODEVICE;
  th_ctx* c = new th_ctx();
  for (int i = 0; i < n_gpus; i++) c->gpus.emplace_back(new Gpu());
  for (int i = 0; i < n_gpus; i++) {
    Gpu& g = *c->gpus[(size_t)i];
    g.index = i;
    g.dev = gpu_ids ? gpu_ids[i] : i;
    if (const char* e = getenv("TRITONHASH_STREAM_SETS")) {
      const int v = atoi(e);
      if (v >= 1 && v <= kMaxStreamSets) g.n_sets = v;
    }
    cudaDeviceProp prop{};
    if (g.dev < 0 || g.dev >= count || cudaGetDeviceProperties(&prop, g.dev) != cudaSuccess || prop.major < 10) {
      g.dev = -1;
      th_shutdown(c);
      return TH_ENODEVICE;  // kernels are sm_100a only; there is no other path
    }
    cudaError_t e = cudaSetDevice(g.dev);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&g.s_copy, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&g.s_misc, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreate(&g.ev_epoch);
    if (e == cudaSuccess) e = cudaEventRecord(g.ev_epoch, g.s_copy);
    for (int k = 0; k < g.n_sets && e == cudaSuccess; k++)
      for (int a = 0; a < 3 && e == cudaSuccess; a++) e = cudaStreamCreateWithFlags(&g.sets[k].alg[a], cudaStreamNonBlocking);
    if (e == cudaSuccess && ring_bytes_per_gpu) {
      g.ring_bytes = align_up(ring_bytes_per_gpu, kRingAlign);
      e = cudaHostAlloc((void**)&g.ring_h, g.ring_bytes, cudaHostAllocPortable);
      g.dring_bytes = align_up(device_bytes_per_gpu, kRingAlign);
      if (e == cudaSuccess) e = cudaMalloc((void**)&g.ring_d, g.dring_bytes);
    }
    if (e == cudaSuccess) {  // ~108 bytes of descriptors + digests per message; 1/128 of the ring covers 16 KiB messages
      g.arena_bytes = align_up(std::max<size_t>(64u << 20, std::max(g.ring_bytes, g.dring_bytes) / 128), 256);
      e = cudaHostAlloc((void**)&g.arena_h, g.arena_bytes, cudaHostAllocPortable);
      if (e == cudaSuccess) e = cudaMalloc((void**)&g.arena_d, g.arena_bytes);
      if (e == cudaSuccess) g.arena_free[0] = g.arena_bytes;
    }
    if (e != cudaSuccess) {
      int rc = (e == cudaErrorMemoryAllocation) ? TH_ENOMEM : TH_ECUDA;
      cudaGetLastError();
      th_shutdown(c);
      return rc;
    }
  }
  if (const char* e = getenv("TRITONHASH_TICKET_TIMING")) c->timing.store(atoi(e) != 0);
  c->completer = std::thread(completer_main, c);
  c->completer_id = c->completer.get_id();
  *out = c;
  return TH_OK;
}

int th_set_ticket_timing(th_ctx* c, int on) {
  if (!c) return TH_EINVAL;
  c->timing.store(on != 0);
  return TH_OK;
}

int th_shutdown(th_ctx* c) {
  if (!c) return TH_EINVAL;
  for (auto& gp : c->gpus) {
    if (gp->dev < 0) continue;
    cudaSetDevice(gp->dev);
    cudaDeviceSynchronize();  // every host function has run: the completer's queue holds all that will ever come
  }
  if (c->completer.joinable()) {
    {
      std::lock_guard<std::mutex> lk(c->cq_mu);
      c->stop = true;
      c->cq_cv.notify_all();
    }
    c->completer.join();  // delivers what is queued, then returns
  }
  for (auto& gp : c->gpus) {
    Gpu& g = *gp;
    if (g.dev < 0) continue;
    cudaSetDevice(g.dev);
    for (auto& kv : g.tickets) {
      Ticket* t = kv.second;
      free_ticket_events(g, t);
      if (t->jb) { if (!t->jb->from_arena) { cudaFreeHost(t->jb->h); cudaFree(t->jb->d); } delete t->jb; }
      delete t;
    }
    g.tickets.clear();
    for (auto& s : g.slots) put_event(g, s.ev_copied);
    for (auto* jb : g.free_jobs) { cudaFreeHost(jb->h); cudaFree(jb->d); delete jb; }
    for (auto ev : g.ev_pool) cudaEventDestroy(ev);
    if (g.arena_h) cudaFreeHost(g.arena_h);
    if (g.arena_d) cudaFree(g.arena_d);
    if (g.ring_h) cudaFreeHost(g.ring_h);
    if (g.ring_d) cudaFree(g.ring_d);
    if (g.s_copy) cudaStreamDestroy(g.s_copy);
    if (g.s_misc) cudaStreamDestroy(g.s_misc);
    if (g.ev_epoch) cudaEventDestroy(g.ev_epoch);
    for (auto& ss : g.sets)
      for (auto& s : ss.alg) if (s) cudaStreamDestroy(s);
  }
  delete c;
  return TH_OK;
}

int th_gpu_count(const th_ctx* c) { return c ? (int)c->gpus.size() : TH_EINVAL; }

static int ring_acquire(th_ctx* c, int gpu, size_t bytes, void** host_ptr, th_slot* slot, bool wait);

int th_ring_acquire(th_ctx* c, int gpu, size_t bytes, void** host_ptr, th_slot* slot) {
  return ring_acquire(c, gpu, bytes, host_ptr, slot, true);
}

int th_ring_try_acquire(th_ctx* c, int gpu, size_t bytes, void** host_ptr, th_slot* slot) {
  return ring_acquire(c, gpu, bytes, host_ptr, slot, false);
}

static int ring_acquire(th_ctx* c, int gpu, size_t bytes, void** host_ptr, th_slot* slot, bool wait) {
  if (!c || !host_ptr || !slot || gpu < 0 || gpu >= (int)c->gpus.size() || bytes == 0) return TH_EINVAL;
  if (c->sticky.load()) return c->sticky.load();
  Gpu& g = *c->gpus[(size_t)gpu];
  std::lock_guard<std::mutex> lk(g.mu);
  const size_t need = align_up(bytes, kRingAlign);
  if (need > g.ring_bytes || need > g.dring_bytes) return TH_ETOOBIG;
  cudaSetDevice(g.dev);
  for (;;) {
    reap(c, g, false);
    Slot* busy = oldest_host_busy(g);
    const size_t off = ring_fit(g.ring_bytes, g.head, g.tail, busy == nullptr, need);
    const size_t doff = ring_fit(g.dring_bytes, g.dhead, g.dtail, g.slots.empty(), need);
    if (off != (size_t)-1 && doff != (size_t)-1) {
      Slot s;
      s.id = new_id(c, g);
      s.gpu = gpu;
      s.off = off;
      s.doff = doff;
      s.bytes = need;
      s.used = (bytes + 15) & ~(size_t)15;
      if (!busy) g.head = off;
      g.tail = off + need;
      if (g.slots.empty()) g.dhead = doff;
      g.dtail = doff + need;
      g.slots.push_back(s);
      *host_ptr = g.ring_h + off;
      *slot = s.id;
      return TH_OK;
    }
    if (off == (size_t)-1) {  // pinned ring full: wait for the oldest outstanding H2D copy
      if (!busy->submitted || !wait) return TH_EBUSY;
      cudaError_t e = cudaEventSynchronize(busy->ev_copied);
      if (e != cudaSuccess) return cuda_fail(c, e, "cudaEventSynchronize(slot copy)");
      busy->host_free = true;
      continue;
    }
    // device ring full: wait for the oldest slot's tickets
    if (!g.slots.front().submitted || !wait) return TH_EBUSY;
    reap(c, g, true);
    if (c->sticky.load()) return c->sticky.load();
  }
}

int th_ring_release(th_ctx* c, th_slot slot) {
  if (!c) return TH_EINVAL;
  Gpu* g = gpu_of(c, slot);
  if (!g) return TH_ESTATE;
  std::lock_guard<std::mutex> lk(g->mu);
  Slot* s = find_slot(*g, slot);
  if (!s || s->submitted) return TH_ESTATE;
  s->done = s->host_free = true;
  cudaSetDevice(g->dev);
  reap(c, *g, false);
  return TH_OK;
}

// g.mu held.
static int submit_common(th_ctx* c, Gpu& g, Slot* slot, const void* d_base_user, const th_msg* msgs, size_t n,
                         uint32_t mask, const uint8_t* expected, th_done_cb cb, void* user, th_ticket* out) {
  int rc = TH_OK;
  cudaSetDevice(g.dev);
  JobBuf* jb = get_job(c, g, n, &rc);
  if (!jb) return rc;
  Ticket* t = new Ticket();
  t->id = new_id(c, g);
  t->gpu = g.index;
  t->n = n;
  t->mask = mask;
  t->has_expected = expected != nullptr && (mask & TH_ALG_SHA1);
  t->jb = jb;
  t->cb = cb;
  t->user = user;
  t->ctx = c;
  const uint8_t* d_base = slot ? g.ring_d + slot->doff : static_cast<const uint8_t*>(d_base_user);
  if (slot) t->slot_id = slot->id;
  rc = enqueue(c, g, t, d_base, slot, msgs, n, mask, t->has_expected ? expected : nullptr);
  if (rc != TH_OK) {
    // items launched before the failure still reference the ticket from their host functions; they cannot queue
    // it for delivery (fewer items were enqueued than cb_left counts), but they must have run before it is freed
    cudaDeviceSynchronize();
    free_ticket_events(g, t);
    put_job(g, jb);
    delete t;
    return rc;
  }
  if (slot) { slot->submitted = true; slot->tickets.push_back(t->id); }
  g.tickets[t->id] = t;
  if (cb) {
    c->cb_outstanding.fetch_add(1);
    std::lock_guard<std::mutex> lk(c->cq_mu);
    c->cq_cv.notify_one();  // the completer now waits with the watchdog's time-out
  }
  *out = t->id;
  return TH_OK;
}

int th_submit(th_ctx* c, th_slot slot_id, const th_msg* msgs, size_t n, uint32_t mask, const uint8_t* expected,
              th_done_cb cb, void* user, th_ticket* out) {
  if (!c || !out) return TH_EINVAL;
  if (c->sticky.load()) return c->sticky.load();
  Gpu* g = gpu_of(c, slot_id);
  if (!g) return TH_ESTATE;
  std::lock_guard<std::mutex> lk(g->mu);
  Slot* s = find_slot(*g, slot_id);
  if (!s || s->done) return TH_ESTATE;
  int rc = validate(msgs, n, mask, s->used);
  if (rc != TH_OK) return rc;
  return submit_common(c, *g, s, nullptr, msgs, n, mask, expected, cb, user, out);
}

int th_submit_device(th_ctx* c, int gpu, const void* d_base, const th_msg* msgs, size_t n, uint32_t mask,
                     const uint8_t* expected, th_done_cb cb, void* user, th_ticket* out) {
  if (!c || !out || gpu < 0 || gpu >= (int)c->gpus.size()) return TH_EINVAL;
  if (reinterpret_cast<uintptr_t>(d_base) & 15) return TH_EALIGN;
  if (c->sticky.load()) return c->sticky.load();
  int rc = validate(msgs, n, mask, 0);
  if (rc != TH_OK) return rc;
  if (!d_base) return TH_EINVAL;
  Gpu& g = *c->gpus[(size_t)gpu];
  std::lock_guard<std::mutex> lk(g.mu);
  return submit_common(c, g, nullptr, d_base, msgs, n, mask, expected, cb, user, out);
}

// Waits for ticket `id` (lk holds g.mu on entry and on return; it is released while blocking, so other threads can
// keep submitting to this GPU).  Returns the ticket, or nullptr with *rc set.
static Ticket* wait_ticket(th_ctx* c, Gpu& g, std::unique_lock<std::mutex>& lk, th_ticket id, bool block, int* rc) {
  auto it = g.tickets.find(id);
  if (it == g.tickets.end()) { *rc = TH_ESTATE; return nullptr; }
  Ticket* t = it->second;
  if (!t->finished) {
    cudaSetDevice(g.dev);
    if (!block) {
      const cudaError_t e = ticket_poll(t, false);
      if (e == cudaErrorNotReady) { *rc = TH_EBUSY; return nullptr; }
      t->finished = true;
      if (e != cudaSuccess) t->status = cuda_fail(c, e, "cudaEventQuery(ticket)");
    } else {
      std::vector<cudaEvent_t> evs;
      for (auto& tl : t->tails) evs.push_back(tl.done);
      lk.unlock();  // the events belong to the ticket, and only its owner (this caller) retires it
      cudaError_t e = cudaSuccess;
      for (cudaEvent_t ev : evs) {
        e = cudaEventSynchronize(ev);
        if (e != cudaSuccess) break;
      }
      lk.lock();
      it = g.tickets.find(id);
      if (it == g.tickets.end()) { *rc = TH_ESTATE; return nullptr; }  // retired by another thread meanwhile
      t = it->second;
      if (!t->finished) {
        t->finished = true;
        if (e != cudaSuccess) t->status = cuda_fail(c, e, "cudaEventSynchronize(ticket)");
      }
    }
  }
  if (!t->timed && t->status == TH_OK && t->ev_k0[0] == nullptr && t->ev_k0[1] == nullptr && t->ev_k0[2] == nullptr) {
    t->timed = true;  // submitted with ticket timing off: times stay -1
  }
  if (!t->timed && t->status == TH_OK) {
    float ms = 0.f;
    cudaSetDevice(g.dev);
    if (t->ev_ready && cudaEventElapsedTime(&ms, g.ev_epoch, t->ev_ready) == cudaSuccess) t->times_ms[0] = ms;
    for (int a = 0; a < 3; a++) {
      if (!t->ev_k0[a]) continue;
      if (cudaEventElapsedTime(&ms, g.ev_epoch, t->ev_k0[a]) == cudaSuccess) t->times_ms[1 + 2 * a] = ms;
      for (auto& tl : t->tails)  // a digest ends when the last of its launch items does
        if ((tl.alg == a || (a == 1 && tl.also_md5)) && tl.k_end && cudaEventElapsedTime(&ms, g.ev_epoch, tl.k_end) == cudaSuccess)
          t->times_ms[2 + 2 * a] = std::max(t->times_ms[2 + 2 * a], (double)ms);
      if (t->times_ms[1 + 2 * a] >= 0 && t->times_ms[2 + 2 * a] >= 0) g.st.kernel_ms += t->times_ms[2 + 2 * a] - t->times_ms[1 + 2 * a];
    }
    t->timed = true;
  }
  *rc = t->status;
  return t;
}

int th_ticket_timing(th_ctx* c, th_ticket id, double* ms7) {
  if (!c || !ms7) return TH_EINVAL;
  Gpu* g = gpu_of(c, id);
  if (!g) return TH_ESTATE;
  std::unique_lock<std::mutex> lk(g->mu);
  int rc = TH_OK;
  Ticket* t = wait_ticket(c, *g, lk, id, true, &rc);
  if (!t) return rc;
  for (int i = 0; i < 7; i++) ms7[i] = t->times_ms[i];
  return rc;
}

int th_wait(th_ctx* c, th_ticket id) {
  if (!c) return TH_EINVAL;
  Gpu* g = gpu_of(c, id);
  if (!g) return TH_ESTATE;
  std::unique_lock<std::mutex> lk(g->mu);
  int rc = TH_OK;
  wait_ticket(c, *g, lk, id, true, &rc);
  return rc;
}

int th_poll(th_ctx* c, th_ticket id) {
  if (!c) return TH_EINVAL;
  Gpu* g = gpu_of(c, id);
  if (!g) return TH_ESTATE;
  std::unique_lock<std::mutex> lk(g->mu);
  int rc = TH_OK;
  wait_ticket(c, *g, lk, id, false, &rc);
  return rc;
}

int th_result(th_ctx* c, th_ticket id, uint8_t* sha1, uint8_t* md5, uint8_t* sha256, uint32_t* ok_bits) {
  if (!c) return TH_EINVAL;
  Gpu* gp = gpu_of(c, id);
  if (!gp) return TH_ESTATE;
  Gpu& g = *gp;
  std::unique_lock<std::mutex> lk(g.mu);
  int rc = TH_OK;
  Ticket* t = wait_ticket(c, g, lk, id, true, &rc);
  if (!t) return rc;
  // A ticket submitted with a callback is retired only after the callback has returned (its `user` pointer and
  // the caller's thunk must outlive it) — unless this IS the callback, running on the completer thread.
  if (t->cb && t->cb_state != 2 && std::this_thread::get_id() != c->completer_id) {
    for (;;) {
      lk.unlock();
      {
        std::unique_lock<std::mutex> ql(c->cq_mu);
        c->cb_cv.wait_for(ql, std::chrono::milliseconds(5));
      }
      lk.lock();
      auto it = g.tickets.find(id);
      if (it == g.tickets.end()) return TH_ESTATE;
      t = it->second;
      if (t->cb_state == 2) break;
    }
    rc = t->status;
  }
  const size_t n = t->n;
  JobBuf* jb = t->jb;
  if (rc == TH_OK) {
    if (sha1) { if (!(t->mask & TH_ALG_SHA1)) rc = TH_EINVAL; else std::memcpy(sha1, jb->h + jb->o_sha1(n), n * 20); }
    if (md5) { if (!(t->mask & TH_ALG_MD5)) rc = TH_EINVAL; else std::memcpy(md5, jb->h + jb->o_md5(n), n * 16); }
    if (sha256) { if (!(t->mask & TH_ALG_SHA256)) rc = TH_EINVAL; else std::memcpy(sha256, jb->h + jb->o_sha256(n), n * 32); }
    if (ok_bits) { if (!t->has_expected) rc = TH_EINVAL; else std::memcpy(ok_bits, jb->h + jb->o_ok(n), ((n + 31) / 32) * 4); }
  }
  cudaSetDevice(g.dev);
  free_ticket_events(g, t);
  put_job(g, jb);
  g.tickets.erase(id);
  delete t;
  reap(c, g, false);
  return rc;
}

int th_etag_combine(th_ctx* c, int gpu, const uint8_t* md5s, const uint32_t* counts, size_t n_objects,
                    uint8_t* out16) {
  if (!c || gpu < 0 || gpu >= (int)c->gpus.size() || !counts || !out16 || n_objects > 0x7fffffffu) return TH_EINVAL;
  if (n_objects == 0) return TH_OK;
  if (c->sticky.load()) return c->sticky.load();
  Gpu& g = *c->gpus[(size_t)gpu];
  size_t parts = 0;
  for (size_t i = 0; i < n_objects; i++) parts += counts[i];
  if (parts && !md5s) return TH_EINVAL;
  // One block of the staging arena holds the part digests, the prefix sums and the result, pinned and device side:
  // no allocation (a device-wide synchronisation) happens here, so a combine does not disturb running chains.
  const size_t b_md5 = align_up(parts * 16 + 16, 256), b_pre = align_up((n_objects + 1) * 8, 256), b_out = align_up(n_objects * 16, 256);
  JobBuf* jb = nullptr;
  cudaEvent_t ev = nullptr;
  int rc = TH_OK;
  {
    std::lock_guard<std::mutex> lk(g.mu);
    CU(c, cudaSetDevice(g.dev));
    jb = get_buf(c, g, b_md5 + b_pre + b_out, 0, &rc);
    if (!jb) return rc;
    if (parts) std::memcpy(jb->h, md5s, parts * 16);
    uint64_t* prefix = reinterpret_cast<uint64_t*>(jb->h + b_md5);
    prefix[0] = 0;
    for (size_t i = 0; i < n_objects; i++) prefix[i + 1] = prefix[i] + counts[i];
    cudaStream_t st = g.s_misc;
    cudaError_t e = cudaMemcpyAsync(jb->d, jb->h, b_md5 + b_pre, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) rc = th_etag_combine_device(jb->d, reinterpret_cast<uint64_t*>(jb->d + b_md5), (uint32_t)n_objects, jb->d + b_md5 + b_pre, st);
    if (e == cudaSuccess && rc == TH_OK) e = cudaMemcpyAsync(jb->h + b_md5 + b_pre, jb->d + b_md5 + b_pre, n_objects * 16, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = get_event(g, &ev);
    if (e == cudaSuccess) e = cudaEventRecord(ev, st);
    if (e != cudaSuccess || rc != TH_OK) {
      cudaStreamSynchronize(st);
      put_event(g, ev);
      put_job(g, jb);
      return e != cudaSuccess ? cuda_fail(c, e, "th_etag_combine") : rc;
    }
    g.st.launches += 1;
    g.st.h2d_bytes += b_md5 + b_pre;
    g.st.d2h_bytes += n_objects * 16;
  }
  const cudaError_t e = cudaEventSynchronize(ev);  // outside the lock: submits to this GPU go on meanwhile
  std::lock_guard<std::mutex> lk(g.mu);
  if (e == cudaSuccess) std::memcpy(out16, jb->h + b_md5 + b_pre, n_objects * 16);
  put_event(g, ev);
  put_job(g, jb);
  return e == cudaSuccess ? TH_OK : cuda_fail(c, e, "th_etag_combine");
}

int th_stats_get(th_ctx* c, int gpu, th_stats* out) {
  if (!c || !out || gpu >= (int)c->gpus.size()) return TH_EINVAL;
  std::memset(out, 0, sizeof(*out));
  for (size_t i = 0; i < c->gpus.size(); i++) {
    if (gpu >= 0 && (int)i != gpu) continue;
    std::lock_guard<std::mutex> lk(c->gpus[i]->mu);
    const th_stats& s = c->gpus[i]->st;
    out->messages += s.messages; out->bytes += s.bytes; out->launches += s.launches;
    out->h2d_bytes += s.h2d_bytes; out->d2h_bytes += s.d2h_bytes; out->submits += s.submits;
    out->kernel_ms += s.kernel_ms; out->overflow_allocs += s.overflow_allocs;
  }
  return TH_OK;
}

const char* th_strerror(int code) {
  switch (code) {
    case TH_OK: return "ok";
    case TH_EINVAL: return "invalid argument";
    case TH_ENODEVICE: return "no usable CUDA device (sm_100 required; there is no CPU fallback)";
    case TH_ECUDA: return "CUDA runtime error (see th_last_error)";
    case TH_ENOMEM: return "out of host or device memory";
    case TH_EALIGN: return "message offset or base pointer is not 16-byte aligned";
    case TH_EBUSY: return "not ready: ring full and its oldest slot has not been submitted, or ticket still running (th_poll)";
    case TH_ETOOBIG: return "request exceeds the ring size or a fixed limit";
    case TH_ESTATE: return "unknown ticket/slot or call out of order";
    default: return "unknown error";
  }
}

const char* th_last_error(const th_ctx* c) { return c ? c->last_error : ""; }

void th_hex(const uint8_t* in, size_t n, char* out) {
  static const char d[] = "0123456789abcdef";
  for (size_t i = 0; i < n; i++) { out[2 * i] = d[in[i] >> 4]; out[2 * i + 1] = d[in[i] & 15]; }
  out[2 * n] = 0;
}

size_t th_base64(const uint8_t* in, size_t n, char* out) {
  static const char a[] = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
  size_t o = 0;
  for (size_t i = 0; i + 2 < n; i += 3) {
    const uint32_t v = ((uint32_t)in[i] << 16) | ((uint32_t)in[i + 1] << 8) | in[i + 2];
    out[o++] = a[v >> 18]; out[o++] = a[(v >> 12) & 63]; out[o++] = a[(v >> 6) & 63]; out[o++] = a[v & 63];
  }
  const size_t rem = n % 3;
  if (rem == 1) {
    const uint32_t v = (uint32_t)in[n - 1] << 4;
    out[o++] = a[v >> 6]; out[o++] = a[v & 63]; out[o++] = '='; out[o++] = '=';
  } else if (rem == 2) {
    const uint32_t v = (((uint32_t)in[n - 2] << 8) | in[n - 1]) << 2;
    out[o++] = a[v >> 12]; out[o++] = a[(v >> 6) & 63]; out[o++] = a[v & 63]; out[o++] = '=';
  }
  out[o] = 0;
  return o;
}

}  // extern "C"
