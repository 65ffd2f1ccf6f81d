// napi_mock.cc — TEST INFRASTRUCTURE: a miniature stand-in for the N-API runtime, just enough to EXECUTE
// js/addon/tritonhash_napi.cc on a GPU box that has no Node.js: values are tagged structs, a Promise is a
// settled-flag + value, a thread-safe function is a queue the "event loop" below drains on the main thread.
// The driver at the bottom plays the JavaScript side of js/lib/tritonhash.js: init -> ringAcquire -> fill the
// pinned Buffer -> hashBatch (Promise) -> etagCombine -> stats -> shutdown, and checks every digest against
// libcrypto (the checker; nothing of this file is shipped).  Semantics follow Node's n-api.md for the calls used.
#include <node_api.h>

#include <openssl/evp.h>

#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <mutex>
#include <string>
#include <vector>

struct napi_value__ {
  enum T { UNDEF, NUL, NUM, BIGINT, STR, OBJ, ARR, BUF, FUNC, ERR, PROMISE } t = UNDEF;
  double num = 0;
  uint64_t big = 0;
  std::string str;
  std::map<std::string, napi_value> props;
  std::vector<napi_value> arr;
  void* data = nullptr;
  size_t len = 0;
  std::vector<uint8_t> own;
  napi_callback fn = nullptr;
  napi_deferred deferred = nullptr;
};
struct napi_deferred__ { bool settled = false, rejected = false; napi_value value = nullptr; };
struct napi_env__ { bool pending = false; std::string code, message; };
struct napi_callback_info__ { std::vector<napi_value> args; };
struct napi_threadsafe_function__ { napi_threadsafe_function_call_js cb; void* ctx; bool released = false; };

namespace {
std::mutex g_mu;
std::condition_variable g_cv;
std::deque<std::pair<napi_threadsafe_function, void*>> g_queue;
napi_value mk(napi_value__::T t) { napi_value v = new napi_value__(); v->t = t; return v; }
napi_status thrown(napi_env env, const char* code, const char* msg) {
  env->pending = true; env->code = code ? code : ""; env->message = msg ? msg : ""; return napi_ok;
}
}  // namespace

extern "C" {
napi_status napi_get_cb_info(napi_env, napi_callback_info info, size_t* argc, napi_value* argv, napi_value*, void**) {
  const size_t cap = *argc;
  for (size_t i = 0; i < cap; i++) argv[i] = i < info->args.size() ? info->args[i] : mk(napi_value__::UNDEF);
  *argc = info->args.size();  // as Node: the actual count, even if larger or smaller than the capacity
  return napi_ok;
}
napi_status napi_get_array_length(napi_env, napi_value v, uint32_t* r) { if (v->t != napi_value__::ARR) return napi_invalid_arg; *r = (uint32_t)v->arr.size(); return napi_ok; }
napi_status napi_get_element(napi_env, napi_value v, uint32_t i, napi_value* r) { if (v->t != napi_value__::ARR || i >= v->arr.size()) return napi_invalid_arg; *r = v->arr[i]; return napi_ok; }
napi_status napi_get_value_int32(napi_env, napi_value v, int32_t* r) { if (v->t != napi_value__::NUM) return napi_invalid_arg; *r = (int32_t)v->num; return napi_ok; }
napi_status napi_get_value_uint32(napi_env, napi_value v, uint32_t* r) { if (v->t != napi_value__::NUM) return napi_invalid_arg; *r = (uint32_t)v->num; return napi_ok; }
napi_status napi_get_value_double(napi_env, napi_value v, double* r) { if (v->t != napi_value__::NUM) return napi_invalid_arg; *r = v->num; return napi_ok; }
napi_status napi_get_value_bigint_uint64(napi_env, napi_value v, uint64_t* r, bool* lossless) { if (v->t != napi_value__::BIGINT) return napi_invalid_arg; *r = v->big; *lossless = true; return napi_ok; }
napi_status napi_get_undefined(napi_env, napi_value* r) { *r = mk(napi_value__::UNDEF); return napi_ok; }
napi_status napi_get_null(napi_env, napi_value* r) { *r = mk(napi_value__::NUL); return napi_ok; }
napi_status napi_get_named_property(napi_env, napi_value o, const char* k, napi_value* r) {
  if (o->t != napi_value__::OBJ) return napi_object_expected;
  auto it = o->props.find(k); *r = it == o->props.end() ? mk(napi_value__::UNDEF) : it->second; return napi_ok;
}
napi_status napi_set_named_property(napi_env, napi_value o, const char* k, napi_value v) { if (o->t != napi_value__::OBJ) return napi_object_expected; o->props[k] = v; return napi_ok; }
napi_status napi_get_buffer_info(napi_env, napi_value v, void** d, size_t* l) { if (v->t != napi_value__::BUF) return napi_invalid_arg; *d = v->data; *l = v->len; return napi_ok; }
napi_status napi_is_buffer(napi_env, napi_value v, bool* r) { *r = v->t == napi_value__::BUF; return napi_ok; }
napi_status napi_create_uint32(napi_env, uint32_t x, napi_value* r) { *r = mk(napi_value__::NUM); (*r)->num = x; return napi_ok; }
napi_status napi_create_double(napi_env, double x, napi_value* r) { *r = mk(napi_value__::NUM); (*r)->num = x; return napi_ok; }
napi_status napi_create_object(napi_env, napi_value* r) { *r = mk(napi_value__::OBJ); return napi_ok; }
napi_status napi_create_string_utf8(napi_env, const char* s, size_t n, napi_value* r) { *r = mk(napi_value__::STR); (*r)->str = n == NAPI_AUTO_LENGTH ? std::string(s) : std::string(s, n); return napi_ok; }
napi_status napi_create_error(napi_env, napi_value, napi_value msg, napi_value* r) { *r = mk(napi_value__::ERR); (*r)->str = msg ? msg->str : ""; return napi_ok; }
napi_status napi_create_bigint_uint64(napi_env, uint64_t x, napi_value* r) { *r = mk(napi_value__::BIGINT); (*r)->big = x; return napi_ok; }
napi_status napi_create_buffer(napi_env, size_t n, void** d, napi_value* r) {
  *r = mk(napi_value__::BUF); (*r)->own.assign(n ? n : 1, 0); (*r)->data = (*r)->own.data(); (*r)->len = n; if (d) *d = (*r)->data; return napi_ok;
}
napi_status napi_create_external_buffer(napi_env, size_t n, void* d, napi_finalize, void*, napi_value* r) { *r = mk(napi_value__::BUF); (*r)->data = d; (*r)->len = n; return napi_ok; }
napi_status napi_create_function(napi_env, const char*, size_t, napi_callback cb, void*, napi_value* r) { *r = mk(napi_value__::FUNC); (*r)->fn = cb; return napi_ok; }
napi_status napi_create_promise(napi_env, napi_deferred* d, napi_value* p) { *d = new napi_deferred__(); *p = mk(napi_value__::PROMISE); (*p)->deferred = *d; return napi_ok; }
napi_status napi_resolve_deferred(napi_env, napi_deferred d, napi_value v) { d->settled = true; d->value = v; return napi_ok; }
napi_status napi_reject_deferred(napi_env, napi_deferred d, napi_value v) { d->settled = d->rejected = true; d->value = v; return napi_ok; }
napi_status napi_throw_error(napi_env env, const char* code, const char* msg) { return thrown(env, code, msg); }
napi_status napi_throw_type_error(napi_env env, const char* code, const char* msg) { return thrown(env, code ? code : "TypeError", msg); }
napi_status napi_throw_range_error(napi_env env, const char* code, const char* msg) { return thrown(env, code ? code : "RangeError", msg); }
napi_status napi_delete_reference(napi_env, napi_ref) { return napi_ok; }
napi_status napi_create_threadsafe_function(napi_env, napi_value, napi_value, napi_value, size_t, size_t, void*, napi_finalize, void* ctx,
                                            napi_threadsafe_function_call_js cb, napi_threadsafe_function* r) {
  *r = new napi_threadsafe_function__{cb, ctx};
  return napi_ok;
}
napi_status napi_call_threadsafe_function(napi_threadsafe_function f, void* data, napi_threadsafe_function_call_mode) {
  std::lock_guard<std::mutex> lk(g_mu);  // called from the library's CUDA callback thread
  g_queue.emplace_back(f, data);
  g_cv.notify_one();
  return napi_ok;
}
napi_status napi_release_threadsafe_function(napi_threadsafe_function f, napi_threadsafe_function_release_mode) { f->released = true; return napi_ok; }
napi_value napi_register_module_v1(napi_env env, napi_value exports);  // defined by the shim (NAPI_MODULE_INIT)
}

// ------------------------------------------------------------------------------------------ the "JavaScript" side
namespace {
napi_env__ g_env;
napi_value num(double x) { napi_value v = mk(napi_value__::NUM); v->num = x; return v; }
napi_value call(napi_value exports, const char* name, std::vector<napi_value> args) {
  napi_callback_info__ info{std::move(args)};
  g_env.pending = false;
  return exports->props.at(name)->fn(&g_env, &info);
}
#define CHECK(cond, what)                                                     \
  do {                                                                        \
    if (!(cond)) { std::fprintf(stderr, "NAPI MOCK FAIL: %s (line %d)\n", what, __LINE__); std::exit(1); } \
  } while (0)
// the event loop: run queued thread-safe-function calls on this (the "JS") thread until the promise settles
void await_promise(napi_value p) {
  while (!p->deferred->settled) {
    std::pair<napi_threadsafe_function, void*> item;
    {
      std::unique_lock<std::mutex> lk(g_mu);
      g_cv.wait(lk, [] { return !g_queue.empty(); });
      item = g_queue.front();
      g_queue.pop_front();
    }
    item.first->cb(&g_env, nullptr, item.first->ctx, item.second);
  }
}
std::vector<uint8_t> evp(const EVP_MD* md, const uint8_t* p, size_t n) {
  std::vector<uint8_t> out(EVP_MD_size(md));
  unsigned int len = 0;
  EVP_Digest(p, n, out.data(), &len, md, nullptr);
  return out;
}
}  // namespace

int main() {
  napi_value exports = mk(napi_value__::OBJ);
  CHECK(napi_register_module_v1(&g_env, exports) == exports, "module init");
  CHECK(exports->props.count("hashBatch") && exports->props.at("SHA256")->num == 4, "exports");
  napi_value ids = mk(napi_value__::ARR);
  ids->arr.push_back(num(0));
  call(exports, "init", {ids, num(64 << 20), num(128 << 20)});
  CHECK(!g_env.pending, g_env.message.c_str());

  // 40 equal pieces + a short last one, as download.js cuts them; one expected digest corrupted
  const size_t piece = 256 << 10, n = 41, total = 40 * piece + 12345;
  napi_value acq = call(exports, "ringAcquire", {num(0), num((double)total)});
  CHECK(!g_env.pending && acq && acq->props.at("buffer")->len == total, "ringAcquire");
  uint8_t* buf = static_cast<uint8_t*>(acq->props.at("buffer")->data);
  uint64_t x = 0x9e3779b97f4a7c15ull;
  for (size_t i = 0; i < total; i++) { x = x * 6364136223846793005ull + 1442695040888963407ull; buf[i] = (uint8_t)(x >> 56); }
  napi_value msgs = mk(napi_value__::ARR);
  std::vector<uint8_t> exp1, want1, want5, want256;
  for (size_t i = 0; i < n; i++) {
    const size_t off = i * piece, len = i + 1 < n ? piece : total - off;
    napi_value m = mk(napi_value__::OBJ);
    m->props["offset"] = num((double)off);
    m->props["length"] = num((double)len);
    msgs->arr.push_back(m);
    auto a = evp(EVP_sha1(), buf + off, len), b = evp(EVP_md5(), buf + off, len), c = evp(EVP_sha256(), buf + off, len);
    want1.insert(want1.end(), a.begin(), a.end());
    want5.insert(want5.end(), b.begin(), b.end());
    want256.insert(want256.end(), c.begin(), c.end());
  }
  exp1 = want1;
  exp1[7 * 20 + 3] ^= 0x40;  // piece 7 "fails verification"
  napi_value expected = mk(napi_value__::BUF);
  expected->data = exp1.data();
  expected->len = exp1.size();
  napi_value p = call(exports, "hashBatch", {num(7), acq->props.at("slot"), msgs, expected});
  CHECK(!g_env.pending && p && p->t == napi_value__::PROMISE, g_env.message.c_str());
  await_promise(p);
  CHECK(!p->deferred->rejected, p->deferred->value ? p->deferred->value->str.c_str() : "rejected");
  napi_value r = p->deferred->value;
  CHECK(r->props.at("sha1")->len == n * 20 && !std::memcmp(r->props.at("sha1")->data, want1.data(), n * 20), "sha1 digests");
  CHECK(r->props.at("md5")->len == n * 16 && !std::memcmp(r->props.at("md5")->data, want5.data(), n * 16), "md5 digests");
  CHECK(r->props.at("sha256")->len == n * 32 && !std::memcmp(r->props.at("sha256")->data, want256.data(), n * 32), "sha256 digests");
  const uint32_t* ok = static_cast<const uint32_t*>(r->props.at("okBits")->data);
  for (size_t i = 0; i < n; i++) CHECK(((ok[i / 32] >> (i % 32)) & 1u) == (i == 7 ? 0u : 1u), "verify bitfield");

  // multipart ETag over the part digests
  napi_value md5s = mk(napi_value__::BUF);
  md5s->data = want5.data();
  md5s->len = want5.size();
  napi_value counts = mk(napi_value__::ARR);
  counts->arr.push_back(num((double)n));
  napi_value etag = call(exports, "etagCombine", {md5s, counts});
  CHECK(!g_env.pending && etag && etag->len == 16, "etagCombine");
  CHECK(!std::memcmp(etag->data, evp(EVP_md5(), want5.data(), want5.size()).data(), 16), "etag digest");

  // errors are exceptions / rejections, never crashes: a misaligned offset throws synchronously
  napi_value acq2 = call(exports, "ringAcquire", {num(0), num(4096)});
  napi_value bad = mk(napi_value__::ARR), m = mk(napi_value__::OBJ);
  m->props["offset"] = num(8);
  m->props["length"] = num(64);
  bad->arr.push_back(m);
  napi_value q = call(exports, "hashBatch", {num(1), acq2->props.at("slot"), bad});
  CHECK(q == nullptr && g_env.pending && g_env.code == "ERRTRITONHASH", "misaligned offset must throw ERRTRITONHASH");
  q = call(exports, "hashBatch", {num(1), acq2->props.at("slot"), num(3)});
  CHECK(q == nullptr && g_env.pending && g_env.code == "TypeError", "non-array msgs must throw a TypeError");

  // a full ring is not an error and must not block the event loop: ringAcquire returns null
  napi_value big = call(exports, "ringAcquire", {num(0), num((double)((64 << 20) - 2048))});
  CHECK(!g_env.pending && big && big->t == napi_value__::NUL, "ringAcquire on a full ring must return null");
  call(exports, "ringRelease", {acq2->props.at("slot")});  // the slot that was never submitted goes back to the ring
  CHECK(!g_env.pending, g_env.message.c_str());
  call(exports, "ringRelease", {acq2->props.at("slot")});
  CHECK(g_env.pending && g_env.code == "ERRTRITONHASH", "releasing a slot twice must throw");

  napi_value st = call(exports, "stats", {});
  CHECK(st && st->props.at("submits")->num == 1 && st->props.at("messages")->num == 3 * n, "stats");
  call(exports, "shutdown", {});
  std::printf("NAPI MOCK OK: %zu messages x 3 digests + bitfield + etag through tritonhash_napi.cc\n", n);
  return 0;
}
