// tritonhash_napi.cc — N-API binding of include/tritonhash.h.  The image has no node_api.h and no Node.js
// (js/README.md): a CPU test type-checks it against N-API's documented prototypes (tests/stubs/node_api.h) and links
// it against libtritonhash.so, and a GPU test executes it under a miniature N-API runtime (tests/napi_mock/).  One th_ctx per process; every call that touches the GPU is
// asynchronous: th_submit enqueues, the library's completion callback (delivered on the context's completion
// thread, also when the device has faulted — then with TH_ECUDA) hops to the JS thread through a
// napi_threadsafe_function, and th_result fills the Buffers the Promise resolves with (or the Promise rejects).
#include <node_api.h>

#include <cstring>
#include <string>
#include <vector>

#include "tritonhash.h"

namespace {

th_ctx* g_ctx = nullptr;

#define NAPI_OK(call)                                   \
  do {                                                  \
    if ((call) != napi_ok) {                            \
      napi_throw_error(env, nullptr, "N-API: " #call);  \
      return nullptr;                                   \
    }                                                   \
  } while (0)

napi_value throw_th(napi_env env, int code, const char* where) {
  std::string msg = std::string(where) + ": " + th_strerror(code);
  if (code == TH_ECUDA && g_ctx) msg += std::string(" — ") + th_last_error(g_ctx);
  napi_throw_error(env, "ERRTRITONHASH", msg.c_str());
  return nullptr;
}

struct Job {
  napi_deferred deferred = nullptr;
  napi_threadsafe_function tsfn = nullptr;
  napi_ref keepalive = nullptr;  // the caller's Buffer when the data was not in the ring
  th_ticket ticket = 0;
  size_t n = 0;
  uint32_t algs = 0;
  bool verify = false;
};

// Library completion thread: no napi_env here, only the thread-safe hop.
void on_done(void* user, th_ticket, int status) {
  Job* j = static_cast<Job*>(user);
  napi_call_threadsafe_function(j->tsfn, reinterpret_cast<void*>(static_cast<intptr_t>(status)), napi_tsfn_nonblocking);
}

napi_value make_buffer(napi_env env, size_t bytes, void** data) {
  napi_value b;
  if (napi_create_buffer(env, bytes, data, &b) != napi_ok) return nullptr;
  return b;
}

// JS thread: collect the digests and settle the Promise.
void finish(napi_env env, napi_value, void* context, void* data) {
  Job* j = static_cast<Job*>(context);
  int status = static_cast<int>(reinterpret_cast<intptr_t>(data));
  void *p1 = nullptr, *p5 = nullptr, *p256 = nullptr, *pok = nullptr;
  napi_value result, v;
  napi_create_object(env, &result);
  if (j->algs & TH_ALG_SHA1) { v = make_buffer(env, j->n * 20, &p1); napi_set_named_property(env, result, "sha1", v); }
  if (j->algs & TH_ALG_MD5) { v = make_buffer(env, j->n * 16, &p5); napi_set_named_property(env, result, "md5", v); }
  if (j->algs & TH_ALG_SHA256) { v = make_buffer(env, j->n * 32, &p256); napi_set_named_property(env, result, "sha256", v); }
  if (j->verify) { v = make_buffer(env, ((j->n + 31) / 32) * 4, &pok); napi_set_named_property(env, result, "okBits", v); }
  // th_result is called whatever the status: it is what retires the ticket and gives its buffers back
  const int rc = th_result(g_ctx, j->ticket, static_cast<uint8_t*>(p1), static_cast<uint8_t*>(p5),
                           static_cast<uint8_t*>(p256), static_cast<uint32_t*>(pok));
  if (status == TH_OK) status = rc;
  if (status == TH_OK) {
    napi_resolve_deferred(env, j->deferred, result);
  } else {
    napi_value msg, err;
    std::string text = std::string("hashBatch: ") + th_strerror(status);
    if (status == TH_ECUDA && g_ctx) text += std::string(" — ") + th_last_error(g_ctx);
    napi_create_string_utf8(env, text.c_str(), NAPI_AUTO_LENGTH, &msg);
    napi_create_error(env, nullptr, msg, &err);
    napi_reject_deferred(env, j->deferred, err);  // the stage throws -> lib/main.js nacks, as today
  }
  if (j->keepalive) napi_delete_reference(env, j->keepalive);
  napi_release_threadsafe_function(j->tsfn, napi_tsfn_release);
  delete j;
}

bool read_msgs(napi_env env, napi_value arr, std::vector<th_msg>* out) {
  uint32_t n = 0;
  if (napi_get_array_length(env, arr, &n) != napi_ok) return false;
  out->resize(n);
  for (uint32_t i = 0; i < n; i++) {
    napi_value e, o, l;
    double off = 0, len = 0;
    if (napi_get_element(env, arr, i, &e) != napi_ok || napi_get_named_property(env, e, "offset", &o) != napi_ok ||
        napi_get_named_property(env, e, "length", &l) != napi_ok || napi_get_value_double(env, o, &off) != napi_ok ||
        napi_get_value_double(env, l, &len) != napi_ok)
      return false;
    (*out)[i].offset = static_cast<uint64_t>(off);
    (*out)[i].length = static_cast<uint64_t>(len);
  }
  return true;
}

// init(gpuIds: number[], ringBytesPerGpu: number)
napi_value Init(napi_env env, napi_callback_info info) {
  size_t argc = 3;  // init(gpuIds, pinnedRingBytes[, deviceRingBytes])
  napi_value argv[3];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr));
  uint32_t n = 0;
  NAPI_OK(napi_get_array_length(env, argv[0], &n));
  std::vector<int> ids(n);
  for (uint32_t i = 0; i < n; i++) {
    napi_value e;
    NAPI_OK(napi_get_element(env, argv[0], i, &e));
    NAPI_OK(napi_get_value_int32(env, e, &ids[i]));
  }
  double ring = 0;
  NAPI_OK(napi_get_value_double(env, argv[1], &ring));
  double dring = 0;  // 0 = same as the pinned ring
  if (argc > 2) NAPI_OK(napi_get_value_double(env, argv[2], &dring));
  int rc = th_init_staged(ids.data(), static_cast<int>(n), static_cast<size_t>(ring), static_cast<size_t>(dring), &g_ctx);
  if (rc != TH_OK) return throw_th(env, rc, "init");  // TH_ENODEVICE: there is no CPU fallback
  napi_value u;
  napi_get_undefined(env, &u);
  return u;
}

// ringAcquire(gpu, bytes) -> { slot: bigint, buffer: Buffer over the pinned slot (fs.read target) } | null (ring full)
napi_value RingAcquire(napi_env env, napi_callback_info info) {
  size_t argc = 2;
  napi_value argv[2];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr));
  int32_t gpu = 0;
  double bytes = 0;
  NAPI_OK(napi_get_value_int32(env, argv[0], &gpu));
  NAPI_OK(napi_get_value_double(env, argv[1], &bytes));
  void* host = nullptr;
  th_slot slot = 0;
  // never block Node's event loop: when the ring is full the caller gets null and awaits its oldest Promise
  int rc = th_ring_try_acquire(g_ctx, gpu, static_cast<size_t>(bytes), &host, &slot);
  if (rc == TH_EBUSY) {
    napi_value nul;
    NAPI_OK(napi_get_null(env, &nul));
    return nul;
  }
  if (rc != TH_OK) return throw_th(env, rc, "ringAcquire");
  napi_value out, s, b;
  NAPI_OK(napi_create_object(env, &out));
  NAPI_OK(napi_create_bigint_uint64(env, slot, &s));
  NAPI_OK(napi_create_external_buffer(env, static_cast<size_t>(bytes), host, nullptr, nullptr, &b));
  NAPI_OK(napi_set_named_property(env, out, "slot", s));
  NAPI_OK(napi_set_named_property(env, out, "buffer", b));
  return out;
}

// ringRelease(slot: bigint): gives back a slot that will not be submitted (the read that was to fill it failed)
napi_value RingRelease(napi_env env, napi_callback_info info) {
  size_t argc = 1;
  napi_value argv[1];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr));
  uint64_t slot = 0;
  bool lossless = false;
  NAPI_OK(napi_get_value_bigint_uint64(env, argv[0], &slot, &lossless));
  int rc = th_ring_release(g_ctx, slot);
  if (rc != TH_OK) return throw_th(env, rc, "ringRelease");
  napi_value u;
  napi_get_undefined(env, &u);
  return u;
}

// hashBatch(algMask, slot: bigint, msgs: {offset,length}[], expectedSha1?: Buffer) -> Promise<{sha1?,md5?,sha256?,okBits?}>
napi_value HashBatch(napi_env env, napi_callback_info info) {
  size_t argc = 4;
  napi_value argv[4];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr));
  uint32_t algs = 0;
  uint64_t slot = 0;
  bool lossless = false;
  NAPI_OK(napi_get_value_uint32(env, argv[0], &algs));
  NAPI_OK(napi_get_value_bigint_uint64(env, argv[1], &slot, &lossless));
  std::vector<th_msg> msgs;
  if (!read_msgs(env, argv[2], &msgs)) {
    napi_throw_type_error(env, nullptr, "hashBatch: msgs must be [{offset, length}]");
    return nullptr;
  }
  const uint8_t* expected = nullptr;
  size_t exp_len = 0;
  bool is_buf = false;
  if (argc > 3 && napi_is_buffer(env, argv[3], &is_buf) == napi_ok && is_buf) {
    void* p = nullptr;
    NAPI_OK(napi_get_buffer_info(env, argv[3], &p, &exp_len));
    if (exp_len != msgs.size() * 20) {
      napi_throw_range_error(env, nullptr, "hashBatch: expectedSha1 must be 20 bytes per message");
      return nullptr;
    }
    expected = static_cast<const uint8_t*>(p);  // copied by th_submit before it returns
  }
  Job* j = new Job();
  j->n = msgs.size();
  j->algs = algs;
  j->verify = expected != nullptr && (algs & TH_ALG_SHA1);
  napi_value promise, name;
  NAPI_OK(napi_create_promise(env, &j->deferred, &promise));
  NAPI_OK(napi_create_string_utf8(env, "tritonhash.done", NAPI_AUTO_LENGTH, &name));
  NAPI_OK(napi_create_threadsafe_function(env, nullptr, nullptr, name, 0, 1, nullptr, nullptr, j, finish, &j->tsfn));
  int rc = th_submit(g_ctx, slot, msgs.data(), msgs.size(), algs, expected, on_done, j, &j->ticket);
  if (rc != TH_OK) {
    napi_release_threadsafe_function(j->tsfn, napi_tsfn_abort);
    delete j;
    return throw_th(env, rc, "hashBatch");
  }
  return promise;
}

// etagCombine(md5s: Buffer, counts: number[]) -> Buffer (16 bytes per object)
napi_value EtagCombine(napi_env env, napi_callback_info info) {
  size_t argc = 2;
  napi_value argv[2];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr));
  void* md5s = nullptr;
  size_t len = 0;
  NAPI_OK(napi_get_buffer_info(env, argv[0], &md5s, &len));
  uint32_t n = 0;
  NAPI_OK(napi_get_array_length(env, argv[1], &n));
  std::vector<uint32_t> counts(n);
  for (uint32_t i = 0; i < n; i++) {
    napi_value e;
    NAPI_OK(napi_get_element(env, argv[1], i, &e));
    NAPI_OK(napi_get_value_uint32(env, e, &counts[i]));
  }
  void* out = nullptr;
  napi_value buf = make_buffer(env, static_cast<size_t>(n) * 16, &out);
  int rc = th_etag_combine(g_ctx, 0, static_cast<const uint8_t*>(md5s), counts.data(), n, static_cast<uint8_t*>(out));
  if (rc != TH_OK) return throw_th(env, rc, "etagCombine");
  return buf;
}

napi_value Stats(napi_env env, napi_callback_info) {
  th_stats s;
  int rc = th_stats_get(g_ctx, -1, &s);
  if (rc != TH_OK) return throw_th(env, rc, "stats");
  napi_value o, v;
  napi_create_object(env, &o);
  const struct { const char* k; double val; } f[] = {{"messages", (double)s.messages}, {"bytes", (double)s.bytes},
      {"launches", (double)s.launches}, {"h2dBytes", (double)s.h2d_bytes}, {"d2hBytes", (double)s.d2h_bytes},
      {"submits", (double)s.submits}, {"kernelMs", s.kernel_ms}, {"overflowAllocs", (double)s.overflow_allocs}};
  for (const auto& e : f) { napi_create_double(env, e.val, &v); napi_set_named_property(env, o, e.k, v); }
  return o;
}

napi_value Shutdown(napi_env env, napi_callback_info) {
  if (g_ctx) { th_shutdown(g_ctx); g_ctx = nullptr; }
  napi_value u;
  napi_get_undefined(env, &u);
  return u;
}

}  // namespace

NAPI_MODULE_INIT() {
  const struct { const char* name; napi_callback fn; } fns[] = {{"init", Init}, {"ringAcquire", RingAcquire}, {"ringRelease", RingRelease},
      {"hashBatch", HashBatch}, {"etagCombine", EtagCombine}, {"stats", Stats}, {"shutdown", Shutdown}};
  for (const auto& f : fns) {
    napi_value v;
    if (napi_create_function(env, f.name, NAPI_AUTO_LENGTH, f.fn, nullptr, &v) != napi_ok) return nullptr;
    napi_set_named_property(env, exports, f.name, v);
  }
  napi_value c;
  napi_create_uint32(env, TH_ALG_SHA1, &c); napi_set_named_property(env, exports, "SHA1", c);
  napi_create_uint32(env, TH_ALG_MD5, &c); napi_set_named_property(env, exports, "MD5", c);
  napi_create_uint32(env, TH_ALG_SHA256, &c); napi_set_named_property(env, exports, "SHA256", c);
  return exports;
}
