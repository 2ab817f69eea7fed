// Host runtime pieces of libsmpb200_core.so: gradient-readiness counter, gradient bucket planner,
// pipeline timeline recorder, memory / allocation metrics.
//
// Counterparts of the reference's native GradCounter and Reducer planning helpers (API reconstructed
// from reference smdistributed/modelparallel/torch/model.py:401-403, worker.py:406-417,
// patches/execution.py:156,220,420, server.py:410,455, allreduce/reducer.py:92 and
// ddp_model.py:336-341) and of its timeline / metrics accessors (reference backend/core.py:290,
// 524-562).  Parameters are addressed by dense integer ids (the Python side owns the name map).
#include <cuda_runtime.h>

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#define SMPC_API extern "C" __attribute__((visibility("default")))

// ------------------------------------------------------------------------------------ GradCounter
namespace {

struct GradCounter {
  int num_params, num_mb;
  int cur_mb = 0;
  std::vector<int> expected;  // [mb][param]
  std::vector<int> seen;      // [mb][param]
  std::vector<char> fwd_done; // [mb]
  std::vector<int> outstanding;  // [param] sum over mb of (expected - seen)
  std::vector<char> used;        // [param] expected > 0 in some microbatch
  int fwd_done_count = 0;
  std::mutex mu;
  GradCounter(int p, int m)
      : num_params(p), num_mb(m), expected((size_t)p * m, 0), seen((size_t)p * m, 0), fwd_done(m, 0),
        outstanding(p, 0), used(p, 0) {}
  bool ready(int id) const { return fwd_done_count == num_mb && used[id] && outstanding[id] == 0; }
};

}  // namespace

SMPC_API void* smpc_gc_create(int num_params, int num_microbatches) {
  if (num_params < 0 || num_microbatches <= 0) return nullptr;
  return new GradCounter(num_params, num_microbatches);
}
SMPC_API void smpc_gc_destroy(void* h) { delete static_cast<GradCounter*>(h); }
SMPC_API void smpc_gc_set_microbatch(void* h, int mb) { static_cast<GradCounter*>(h)->cur_mb = mb; }
SMPC_API int smpc_gc_get_microbatch(void* h) { return static_cast<GradCounter*>(h)->cur_mb; }

// A parameter was used once more in the forward pass of microbatch mb: one more gradient
// contribution will arrive for it during that microbatch's backward.
SMPC_API int smpc_gc_increment_expected(void* h, int mb, const int* ids, int n) {
  auto* g = static_cast<GradCounter*>(h);
  if (mb < 0 || mb >= g->num_mb) return -1;
  std::lock_guard<std::mutex> lk(g->mu);
  for (int i = 0; i < n; ++i) {
    const int id = ids[i];
    if (id < 0 || id >= g->num_params) return -1;
    g->expected[(size_t)mb * g->num_params + id]++;
    g->outstanding[id]++;
    g->used[id] = 1;
  }
  return 0;
}
SMPC_API int smpc_gc_mark_fwd_done(void* h, int mb) {
  auto* g = static_cast<GradCounter*>(h);
  if (mb < 0 || mb >= g->num_mb) return -1;
  std::lock_guard<std::mutex> lk(g->mu);
  if (!g->fwd_done[mb]) {
    g->fwd_done[mb] = 1;
    g->fwd_done_count++;
  }
  return 0;
}
// One gradient contribution for `id` landed in the current microbatch's backward.  Returns 1 when
// this was the last contribution of the last microbatch (the gradient is final for the step),
// 0 otherwise, -1 on a bad id, -2 when more contributions arrived than were announced.
SMPC_API int smpc_gc_mark_grad_computed(void* h, int id) {
  auto* g = static_cast<GradCounter*>(h);
  if (id < 0 || id >= g->num_params) return -1;
  std::lock_guard<std::mutex> lk(g->mu);
  const size_t at = (size_t)g->cur_mb * g->num_params + id;
  if (g->seen[at] >= g->expected[at]) return -2;
  g->seen[at]++;
  g->outstanding[id]--;
  return g->ready(id) ? 1 : 0;
}
SMPC_API int smpc_gc_is_ready(void* h, int id) {
  auto* g = static_cast<GradCounter*>(h);
  if (id < 0 || id >= g->num_params) return -1;
  std::lock_guard<std::mutex> lk(g->mu);
  return g->ready(id) ? 1 : 0;
}
SMPC_API int smpc_gc_expected(void* h, int mb, int id) {
  auto* g = static_cast<GradCounter*>(h);
  if (mb < 0 || mb >= g->num_mb || id < 0 || id >= g->num_params) return -1;
  return g->expected[(size_t)mb * g->num_params + id];
}
SMPC_API int smpc_gc_seen(void* h, int mb, int id) {
  auto* g = static_cast<GradCounter*>(h);
  if (mb < 0 || mb >= g->num_mb || id < 0 || id >= g->num_params) return -1;
  return g->seen[(size_t)mb * g->num_params + id];
}
// total expected contributions per parameter over the minibatch -> out[num_params]
SMPC_API void smpc_gc_expected_totals(void* h, int* out) {
  auto* g = static_cast<GradCounter*>(h);
  std::lock_guard<std::mutex> lk(g->mu);
  for (int p = 0; p < g->num_params; ++p) {
    int s = 0;
    for (int m = 0; m < g->num_mb; ++m) s += g->expected[(size_t)m * g->num_params + p];
    out[p] = s;
  }
}
SMPC_API void smpc_gc_clear(void* h) {
  auto* g = static_cast<GradCounter*>(h);
  std::lock_guard<std::mutex> lk(g->mu);
  std::fill(g->expected.begin(), g->expected.end(), 0);
  std::fill(g->seen.begin(), g->seen.end(), 0);
  std::fill(g->fwd_done.begin(), g->fwd_done.end(), 0);
  std::fill(g->outstanding.begin(), g->outstanding.end(), 0);
  std::fill(g->used.begin(), g->used.end(), 0);
  g->fwd_done_count = 0;
  g->cur_mb = 0;
}

// ---------------------------------------------------------------------------------- bucket planner
// Assign parameters (in the order given, normally reverse registration order = gradient arrival
// order) to buckets: a bucket holds one dtype key, the first bucket is capped at first_cap bytes,
// later ones at cap bytes.  bucket_of[i] receives the bucket index; returns the bucket count.
SMPC_API int smpc_plan_buckets(int n, const long long* nbytes, const int* dtype_key, long long first_cap,
                               long long cap, int* bucket_of) {
  struct Open {
    int key;
    int bucket;
    long long bytes;
  };
  std::vector<Open> open;
  int next = 0;
  bool first_closed = false;
  for (int i = 0; i < n; ++i) {
    Open* o = nullptr;
    for (auto& c : open)
      if (c.key == dtype_key[i]) o = &c;
    if (!o) {
      open.push_back(Open{dtype_key[i], next++, 0});
      o = &open.back();
    }
    const long long limit = (!first_closed && o->bucket == 0) ? first_cap : cap;
    if (o->bytes > 0 && o->bytes + nbytes[i] > limit) {
      if (o->bucket == 0) first_closed = true;
      o->bucket = next++;
      o->bytes = 0;
    }
    bucket_of[i] = o->bucket;
    o->bytes += nbytes[i];
  }
  return next;
}

// ---------------------------------------------------------------------------------------- timeline
namespace {

struct Event {
  int step, microbatch;
  int64_t t_us;
  char phase;  // 'B' step begin, 'E' step end, 'i' instant
  std::string label;
};
struct Timeline {
  std::string path;
  int rank = 0;
  int step = -1;
  std::vector<Event> events;
  std::mutex mu;
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  int64_t now() const {
    return std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t0).count();
  }
};
Timeline* g_tl = nullptr;

void json_escape(const std::string& s, std::string& out) {
  for (char c : s) {
    if (c == '"' || c == '\\') {
      out.push_back('\\');
      out.push_back(c);
    } else if (c == '\n') {
      out += "\\n";
    } else if ((unsigned char)c < 0x20) {
      out.push_back(' ');
    } else {
      out.push_back(c);
    }
  }
}

}  // namespace

SMPC_API int smpc_timeline_create(const char* path, int rank) {
  delete g_tl;
  g_tl = new Timeline();
  g_tl->path = path ? path : "";
  g_tl->rank = rank;
  return 0;
}
SMPC_API int smpc_timeline_enabled() { return g_tl != nullptr; }
SMPC_API void smpc_timeline_start_step() {
  if (!g_tl) return;
  std::lock_guard<std::mutex> lk(g_tl->mu);
  g_tl->step++;
  g_tl->events.push_back(Event{g_tl->step, -1, g_tl->now(), 'B', "step"});
}
SMPC_API void smpc_timeline_end_step() {
  if (!g_tl) return;
  std::lock_guard<std::mutex> lk(g_tl->mu);
  g_tl->events.push_back(Event{g_tl->step, -1, g_tl->now(), 'E', "step"});
}
SMPC_API void smpc_timeline_record_pipeline_event(int microbatch, const char* label) {
  if (!g_tl) return;
  std::lock_guard<std::mutex> lk(g_tl->mu);
  g_tl->events.push_back(Event{g_tl->step, microbatch, g_tl->now(), 'i', label ? label : ""});
}
SMPC_API int smpc_timeline_num_events() {
  if (!g_tl) return 0;
  std::lock_guard<std::mutex> lk(g_tl->mu);
  return (int)g_tl->events.size();
}
// Chrome trace-event JSON (chrome://tracing, Perfetto): pid = rank, tid = microbatch + 1 (0 = step lane).
SMPC_API int smpc_timeline_flush() {
  if (!g_tl || g_tl->path.empty()) return -1;
  std::lock_guard<std::mutex> lk(g_tl->mu);
  FILE* f = fopen(g_tl->path.c_str(), "w");
  if (!f) return -1;
  fputs("{\"traceEvents\":[\n", f);
  bool first = true;
  for (const auto& e : g_tl->events) {
    std::string name;
    json_escape(e.label, name);
    fprintf(f, "%s{\"name\":\"%s\",\"ph\":\"%c\",\"ts\":%lld,\"pid\":%d,\"tid\":%d,\"args\":{\"step\":%d,\"microbatch\":%d}%s}",
            first ? "" : ",\n", name.c_str(), e.phase, (long long)e.t_us, g_tl->rank, e.microbatch + 1, e.step,
            e.microbatch, e.phase == 'i' ? ",\"s\":\"t\"" : "");
    first = false;
  }
  fputs("\n]}\n", f);
  fclose(f);
  return 0;
}
SMPC_API void smpc_timeline_destroy() {
  delete g_tl;
  g_tl = nullptr;
}

// ------------------------------------------------------------------------------------------ metrics
namespace {
struct PoolStats {
  std::mutex mu;
  long long cur_alloc = 0, peak_alloc = 0, cur_reserved = 0, peak_reserved = 0;
  long long success = 0, failure = 0;
} g_pool;
}  // namespace

// The pipeline p2p layer reports its staging-buffer use here (the successor of the reference's
// receiver-side D2D pool): delta_alloc / delta_reserved in bytes, ok = allocation succeeded.
SMPC_API void smpc_pool_record(long long delta_alloc, long long delta_reserved, int ok) {
  std::lock_guard<std::mutex> lk(g_pool.mu);
  g_pool.cur_alloc += delta_alloc;
  g_pool.cur_reserved += delta_reserved;
  if (g_pool.cur_alloc > g_pool.peak_alloc) g_pool.peak_alloc = g_pool.cur_alloc;
  if (g_pool.cur_reserved > g_pool.peak_reserved) g_pool.peak_reserved = g_pool.cur_reserved;
  if (delta_alloc > 0 || delta_reserved > 0) (ok ? g_pool.success : g_pool.failure)++;
}
// out = {peak_allocated_mb, peak_reserved_mb, gpu_free_mb, gpu_total_mb}; peaks reset to current.
SMPC_API int smpc_get_and_reset_memory_metrics(double* out) {
  {
    std::lock_guard<std::mutex> lk(g_pool.mu);
    out[0] = g_pool.peak_alloc / 1048576.0;
    out[1] = g_pool.peak_reserved / 1048576.0;
    g_pool.peak_alloc = g_pool.cur_alloc;
    g_pool.peak_reserved = g_pool.cur_reserved;
  }
  size_t free_b = 0, total_b = 0;
  int count = 0;
  if (cudaGetDeviceCount(&count) == cudaSuccess && count > 0 && cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
    out[2] = free_b / 1048576.0;
    out[3] = total_b / 1048576.0;
  } else {
    (void)cudaGetLastError();
    out[2] = out[3] = 0.0;
  }
  return 0;
}
SMPC_API int smpc_get_and_reset_alloc_metrics(long long* out) {
  std::lock_guard<std::mutex> lk(g_pool.mu);
  out[0] = g_pool.success;
  out[1] = g_pool.failure;
  g_pool.success = g_pool.failure = 0;
  return 0;
}

SMPC_API int smpc_device_count() {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0;
  }
  return n;
}

// ---- in-library self-test (the reference ships native self-tests inside its library as well: test_d2d(),
// from_blob_tensor_test(), test_rdma_p2p() - SURVEY 2.2 N1e).  Exercises the gradient counter, the bucket
// planner and the pool metrics through the same C ABI the Python side uses.  Returns 0, or the number of the
// first failed check with a description in `msg`.
SMPC_API int smpc_selftest_runtime(char* msg, int cap) {
  int check = 0;
  auto bad = [&](const char* what) {
    if (msg && cap > 0) snprintf(msg, (size_t)cap, "check %d failed: %s", check, what);
    return check;
  };
#define SMPC_EXPECT(cond) do { ++check; if (!(cond)) return bad(#cond); } while (0)
  // gradient counter: 3 parameters, 2 microbatches; parameter 2 is used twice per microbatch
  void* g = smpc_gc_create(3, 2);
  SMPC_EXPECT(g != nullptr);
  const int ids[4] = {0, 1, 2, 2};
  for (int mb = 0; mb < 2; ++mb) {
    SMPC_EXPECT(smpc_gc_increment_expected(g, mb, ids, 4) == 0);
    SMPC_EXPECT(smpc_gc_mark_fwd_done(g, mb) == 0);
  }
  SMPC_EXPECT(smpc_gc_expected(g, 1, 2) == 2 && smpc_gc_expected(g, 0, 0) == 1);
  smpc_gc_set_microbatch(g, 0);
  SMPC_EXPECT(smpc_gc_get_microbatch(g) == 0);
  SMPC_EXPECT(smpc_gc_mark_grad_computed(g, 0) == 0);
  SMPC_EXPECT(smpc_gc_mark_grad_computed(g, 0) == -2);      // more than announced
  SMPC_EXPECT(smpc_gc_mark_grad_computed(g, 2) == 0 && smpc_gc_mark_grad_computed(g, 2) == 0);
  SMPC_EXPECT(smpc_gc_is_ready(g, 0) == 0 && smpc_gc_mark_grad_computed(g, 7) == -1);
  smpc_gc_set_microbatch(g, 1);
  SMPC_EXPECT(smpc_gc_mark_grad_computed(g, 0) == 1);       // last contribution of the last microbatch
  SMPC_EXPECT(smpc_gc_mark_grad_computed(g, 2) == 0 && smpc_gc_mark_grad_computed(g, 2) == 1);
  SMPC_EXPECT(smpc_gc_is_ready(g, 2) == 1 && smpc_gc_is_ready(g, 1) == 0);
  SMPC_EXPECT(smpc_gc_seen(g, 1, 2) == 2);
  int totals[3] = {0, 0, 0};
  smpc_gc_expected_totals(g, totals);
  SMPC_EXPECT(totals[0] == 2 && totals[1] == 2 && totals[2] == 4);
  smpc_gc_clear(g);
  SMPC_EXPECT(smpc_gc_expected(g, 0, 0) == 0 && smpc_gc_seen(g, 1, 2) == 0);
  smpc_gc_destroy(g);
  // bucket planner: a small first bucket, a size cap, and a key change always starts a new bucket
  const long long nbytes[6] = {600, 600, 3000, 3000, 100, 100};
  const int keys[6] = {0, 0, 0, 0, 0, 1};
  int bucket_of[6] = {-1, -1, -1, -1, -1, -1};
  const int nb = smpc_plan_buckets(6, nbytes, keys, 1000, 5000, bucket_of);
  SMPC_EXPECT(nb >= 3);
  SMPC_EXPECT(bucket_of[0] == 0 && bucket_of[5] == nb - 1 && bucket_of[4] != bucket_of[5]);
  for (int i = 1; i < 6; ++i) SMPC_EXPECT(bucket_of[i] >= bucket_of[i - 1]);
  SMPC_EXPECT(bucket_of[2] != bucket_of[0]);                // the first bucket closes at first_cap
  // pool metrics accumulate and reset
  long long am[2];
  smpc_get_and_reset_alloc_metrics(am);
  smpc_pool_record(1 << 20, 2 << 20, 1);
  smpc_pool_record(0, 4 << 20, 0);                       // a reservation that failed
  smpc_pool_record(0, -(4 << 20), 1);
  smpc_get_and_reset_alloc_metrics(am);
  SMPC_EXPECT(am[0] == 1 && am[1] == 1);
  smpc_get_and_reset_alloc_metrics(am);
  SMPC_EXPECT(am[0] == 0 && am[1] == 0);
  smpc_pool_record(-(1 << 20), -(2 << 20), 1);
  smpc_get_and_reset_alloc_metrics(am);
#undef SMPC_EXPECT
  if (msg && cap > 0) snprintf(msg, (size_t)cap, "ok: %d checks", check);
  return 0;
}
