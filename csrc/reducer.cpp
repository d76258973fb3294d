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

extern "C" void b2_reducer_destroy(b2_reducer* r) { delete r; }

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
