// Host-side DDP reducer core: bucket planning and the ready/launch state machine.
//
// Native counterpart of c10d::Reducer + compute_bucket_assignment_by_size that the reference gets from
// torch.nn.parallel.DistributedDataParallel (reference resnet/main.py:80).  Pure C++ (no CUDA), so it is
// also what the CPU/gloo configuration runs.  The Python wrapper (parallel/ddp.py) owns tensors and streams;
// this object answers two questions: "which bucket does parameter i live in" and "which buckets may be
// launched now, in the fixed order every rank agrees on".
#include "b200ddp.h"

#include <algorithm>
#include <cstring>
#include <mutex>
#include <vector>

// Greedy size-capped assignment over parameters given in gradient-ready order.
// Semantics match torch's: append until the running size reaches the current limit, then close the bucket
// (so a bucket may overshoot by one tensor); the first bucket uses `first_cap_bytes` (1 MiB by default) unless
// it is <= 0, every later bucket `cap_bytes`.
extern "C" int b2_plan_buckets(const long long* sizes_bytes, int n, long long first_cap_bytes, long long cap_bytes,
                               int* bucket_of_param) {
  int bucket = 0;
  long long acc = 0;
  long long limit = first_cap_bytes > 0 ? first_cap_bytes : cap_bytes;
  bool open = false;
  for (int i = 0; i < n; ++i) {
    bucket_of_param[i] = bucket;
    acc += sizes_bytes[i];
    open = true;
    if (acc >= limit) {
      ++bucket;
      acc = 0;
      limit = cap_bytes;
      open = false;
    }
  }
  return bucket + (open ? 1 : 0);
}

struct b2_reducer {
  std::vector<int> bucket_of_param;
  std::vector<int> bucket_size;     // params per bucket
  std::vector<int> pending;         // params not yet ready, per bucket
  std::vector<char> param_ready;
  std::vector<int> ready_order;     // parameter indices in the order they became ready this iteration
  int next_bucket = 0;              // first bucket not yet launched
  int n_buckets = 0;
  long long iterations = 0;
  std::mutex mu;                    // autograd hooks can fire from the engine thread
};

extern "C" b2_reducer* b2_reducer_create(int n_params, const int* bucket_of_param, int n_buckets) {
  auto* r = new b2_reducer();
  r->bucket_of_param.assign(bucket_of_param, bucket_of_param + n_params);
  r->n_buckets = n_buckets;
  r->bucket_size.assign(n_buckets, 0);
  for (int i = 0; i < n_params; ++i) {
    if (bucket_of_param[i] < 0 || bucket_of_param[i] >= n_buckets) {
      delete r;
      return nullptr;
    }
    r->bucket_size[bucket_of_param[i]]++;
  }
  r->pending = r->bucket_size;
  r->param_ready.assign(n_params, 0);
  return r;
}

extern "C" void b2_reducer_release_launch_state(b2_reducer* r);
extern "C" void b2_reducer_destroy(b2_reducer* r) {
  b2_reducer_release_launch_state(r);
  delete r;
}

// prepare_for_backward: every bucket waits for all of its parameters again.
extern "C" void b2_reducer_reset(b2_reducer* r) {
  std::lock_guard<std::mutex> g(r->mu);
  r->pending = r->bucket_size;
  std::fill(r->param_ready.begin(), r->param_ready.end(), 0);
  r->ready_order.clear();
  r->next_bucket = 0;
}

// Marks parameter `idx` ready. Writes to `launch` the buckets that may be launched now (ascending, every rank
// produces the same sequence because launches are strictly in bucket order) and returns how many.
// Returns -1 if the parameter was already marked in this iteration (a second backward without reset).
extern "C" int b2_reducer_mark_ready(b2_reducer* r, int idx, int* launch) {
  std::lock_guard<std::mutex> g(r->mu);
  if (idx < 0 || idx >= (int)r->param_ready.size()) return -2;
  if (r->param_ready[idx]) return -1;
  r->param_ready[idx] = 1;
  r->ready_order.push_back(idx);
  r->pending[r->bucket_of_param[idx]]--;
  int n = 0;
  while (r->next_bucket < r->n_buckets && r->pending[r->next_bucket] == 0) launch[n++] = r->next_bucket++;
  if (r->next_bucket == r->n_buckets) r->iterations++;
  return n;
}

extern "C" int b2_reducer_unlaunched(const b2_reducer* r) { return r->n_buckets - r->next_bucket; }
extern "C" long long b2_reducer_iterations(const b2_reducer* r) { return r->iterations; }
extern "C" int b2_reducer_ready_order(const b2_reducer* r, int* out) {
  std::memcpy(out, r->ready_order.data(), r->ready_order.size() * sizeof(int));
  return (int)r->ready_order.size();
}

// ---------------------------------------------------------------------------------------------------------------
// Native launch path (P2P backend): per-bucket launch descriptors are registered once; marking a parameter ready then
// launches, from C++, the fused all-reduce(+SGD) kernel of every bucket that became launchable - event record on the
// compute stream, wait on the communication stream, kernel launch - and the end-of-backward join is one call as well.
// (c10d::Reducer::mark_variable_ready -> mark_bucket_ready -> all_reduce_bucket and finalize_backward of the reference
// stack, reference resnet/main.py:80,123; Python keeps the tensors, C++ does the sequencing and the launches.)
#include <cuda_runtime.h>

struct b2_bucket_launch {
  b2_bucket_desc d;
  cudaEvent_t ready = nullptr;
  bool set = false;
};
struct b2_reducer_launch_state {
  std::vector<b2_bucket_launch> buckets;
  cudaEvent_t joined = nullptr;
};
static std::mutex g_ls_mu;
static std::vector<std::pair<const b2_reducer*, b2_reducer_launch_state*>> g_ls;

static b2_reducer_launch_state* launch_state(const b2_reducer* r, bool create) {
  std::lock_guard<std::mutex> g(g_ls_mu);
  for (auto& kv : g_ls)
    if (kv.first == r) return kv.second;
  if (!create) return nullptr;
  auto* s = new b2_reducer_launch_state();
  s->buckets.resize(r->n_buckets);
  g_ls.emplace_back(r, s);
  return s;
}

extern "C" int b2_reducer_set_bucket(b2_reducer* r, int bucket, const b2_bucket_desc* desc) {
  if (!r || !desc || bucket < 0 || bucket >= r->n_buckets) return cudaErrorInvalidValue;
  b2_reducer_launch_state* s = launch_state(r, true);
  b2_bucket_launch& b = s->buckets[bucket];
  b.d = *desc;
  if (!b.ready) {
    cudaError_t e = cudaEventCreateWithFlags(&b.ready, cudaEventDisableTiming);
    if (e != cudaSuccess) return e;
  }
  if (!s->joined) {
    cudaError_t e = cudaEventCreateWithFlags(&s->joined, cudaEventDisableTiming);
    if (e != cudaSuccess) return e;
  }
  b.set = true;
  return 0;
}

extern "C" void b2_reducer_release_launch_state(b2_reducer* r) {
  std::lock_guard<std::mutex> g(g_ls_mu);
  for (size_t i = 0; i < g_ls.size(); ++i) {
    if (g_ls[i].first != r) continue;
    for (auto& b : g_ls[i].second->buckets)
      if (b.ready) cudaEventDestroy(b.ready);
    if (g_ls[i].second->joined) cudaEventDestroy(g_ls[i].second->joined);
    delete g_ls[i].second;
    g_ls.erase(g_ls.begin() + i);
    return;
  }
}

// mark parameter idx ready; launch what became launchable. launched[] receives the bucket indices, the return value
// their number (< 0: -1 marked twice, -2 bad index, -3 a bucket without a descriptor, -100 - cudaError otherwise).
extern "C" int b2_reducer_mark_ready_launch(b2_reducer* r, int idx, int fused, float lr, float momentum,
                                            float weight_decay, int first_step, const float* hyper_dev,
                                            void* compute_stream, void* comm_stream, int* launched) {
  const int n = b2_reducer_mark_ready(r, idx, launched);
  if (n <= 0) return n;
  b2_reducer_launch_state* s = launch_state(r, false);
  if (!s) return -3;
  cudaStream_t cs = (cudaStream_t)compute_stream, ms = (cudaStream_t)comm_stream;
  for (int i = 0; i < n; ++i) {
    b2_bucket_launch& b = s->buckets[launched[i]];
    if (!b.set) return -3;
    cudaError_t e = cudaEventRecord(b.ready, cs);          // the kernels that produced this bucket's gradients ...
    if (e == cudaSuccess) e = cudaStreamWaitEvent(ms, b.ready, 0);   // ... before the reduction reads them
    if (e != cudaSuccess) return -100 - (int)e;
    const b2_bucket_desc& d = b.d;
    int rc;
    if (d.wire_bf16)
      rc = b2_allreduce_bf16(&d.ctx, d.mc_base, d.wire_offset_bytes, d.n, d.algo, d.scale, d.g32,
                             (d.write_grad || !fused) ? d.g32 : nullptr, fused ? d.p32 : nullptr, d.mom, d.shadow, lr,
                             momentum, weight_decay, first_step, hyper_dev, d.nblocks, comm_stream);
    else
      rc = b2_allreduce(&d.ctx, d.mc_base, d.offset_bytes, d.n, d.algo, d.scale, fused ? d.p32 : nullptr, d.mom,
                        d.shadow, lr, momentum, weight_decay, first_step, hyper_dev, d.write_grad, d.nblocks,
                        comm_stream);
    if (rc) return -100 - rc;
  }
  return n;
}

// finalize_backward: everything queued on the compute stream from here on waits for the reductions
extern "C" int b2_reducer_join(b2_reducer* r, void* compute_stream, void* comm_stream) {
  b2_reducer_launch_state* s = launch_state(r, false);
  if (!s || !s->joined) return cudaErrorInvalidValue;
  cudaError_t e = cudaEventRecord(s->joined, (cudaStream_t)comm_stream);
  if (e == cudaSuccess) e = cudaStreamWaitEvent((cudaStream_t)compute_stream, s->joined, 0);
  return (int)e;
}
