// Continuous-batching scheduler core (host side).  See include/llmctl_host.h.
//
// Budgets keep the reference's meaning (reference llmctl/serve/server.py:102-125): at most
// `max_batch_size` sequences run concurrently and the sum over running sequences of
// (prompt_len + max_tokens) never exceeds `max_batch_tokens`.  What changes:
//   * admission is continuous: finished sequences free their budget and waiting ones join the running
//     set between decode steps (the reference pops one batch, runs one forward and never re-queues);
//   * no head-of-line blocking: a waiting request that does not fit right now is skipped, not pushed back
//     to the head to block everyone behind it; a request that can *never* fit is rejected at add();
//   * KV-aware (policy vllm-like, or whenever a block manager is attached): admission also needs free KV
//     blocks above the watermark, decode growth allocates a block when a sequence crosses a block boundary
//     and, if the pool is exhausted, the most recently admitted sequence is preempted (its blocks are freed
//     and it is re-queued at the front for re-prefill);
//   * policy simple = static batching: a new batch is admitted only when the running set is empty.
#include "llmctl_host.h"

#include <algorithm>
#include <deque>
#include <mutex>
#include <unordered_map>
#include <vector>

#define API extern "C" __attribute__((visibility("default")))

namespace {

struct Req {
  long long id;
  int prompt_len;
  int max_tokens;
  int generated = 0;
  bool prefilled = false;
  long long admit_order = 0;
};

struct Sched {
  int max_batch_size, max_batch_tokens, policy;
  void* kv;
  std::deque<Req> waiting;
  std::vector<Req> running;
  std::vector<long long> preempted;
  int reserved_tokens = 0;
  long long admit_counter = 0;
  std::mutex mu;

  int reserve(const Req& r) const { return r.prompt_len + r.max_tokens; }

  bool fits(const Req& r) const {
    if (static_cast<int>(running.size()) >= max_batch_size) return false;
    if (reserved_tokens + reserve(r) > max_batch_tokens) return false;
    if (kv && !llmctl_kv_can_allocate(kv, r.prompt_len + r.generated + 1)) {
      // the watermark keeps head-room for running sequences to grow; with nothing running it would only starve a
      // request that needs more than (total - watermark) blocks, so an empty engine admits on free blocks alone
      const int bs = llmctl_kv_block_size(kv);
      const int need = (r.prompt_len + r.generated + 1 + bs - 1) / bs;
      if (!(running.empty() && llmctl_kv_free_blocks(kv) >= need)) return false;
    }
    return true;
  }

  void admit() {
    if (policy == LLMCTL_SCHED_SIMPLE && !running.empty()) return;
    for (auto it = waiting.begin(); it != waiting.end();) {
      if (static_cast<int>(running.size()) >= max_batch_size) break;
      if (fits(*it)) {
        Req r = *it;
        if (kv && llmctl_kv_allocate(kv, r.id, r.prompt_len + r.generated) != 0) {
          ++it;
          continue;
        }
        r.admit_order = admit_counter++;
        r.prefilled = false;
        reserved_tokens += reserve(r);
        running.push_back(r);
        it = waiting.erase(it);
      } else {
        ++it;  // skip: no head-of-line blocking
      }
    }
  }

  // Evict the most recently admitted running sequence other than `keep`.  Returns false if none.
  bool preempt_one(long long keep) {
    int victim = -1;
    for (int i = 0; i < static_cast<int>(running.size()); ++i) {
      if (running[i].id == keep) continue;
      if (victim < 0 || running[i].admit_order > running[victim].admit_order) victim = i;
    }
    if (victim < 0) return false;
    Req r = running[victim];
    running.erase(running.begin() + victim);
    reserved_tokens -= reserve(r);
    if (kv) llmctl_kv_free(kv, r.id);
    r.prefilled = false;  // prompt + generated tokens are recomputed on re-admission
    preempted.push_back(r.id);
    waiting.push_front(r);
    return true;
  }
};

}  // namespace

API void* llmctl_sched_create(int max_batch_size, int max_batch_tokens, int policy, void* kv) {
  if (max_batch_size <= 0 || max_batch_tokens <= 0) return nullptr;
  Sched* s = new Sched();
  s->max_batch_size = max_batch_size;
  s->max_batch_tokens = max_batch_tokens;
  s->policy = policy;
  s->kv = kv;
  return s;
}
API void llmctl_sched_destroy(void* h) { delete static_cast<Sched*>(h); }

API int llmctl_sched_add(void* h, long long id, int prompt_len, int max_tokens) {
  Sched* s = static_cast<Sched*>(h);
  std::lock_guard<std::mutex> lk(s->mu);
  if (prompt_len < 0 || max_tokens < 0) return -2;
  if (prompt_len + max_tokens > s->max_batch_tokens) return -2;
  if (s->kv) {
    const int bs = llmctl_kv_block_size(s->kv);
    if ((prompt_len + max_tokens + bs - 1) / bs > llmctl_kv_total_blocks(s->kv)) return -2;
  }
  Req r;
  r.id = id;
  r.prompt_len = prompt_len;
  r.max_tokens = max_tokens;
  s->waiting.push_back(r);
  return 0;
}

API int llmctl_sched_step(void* h, long long* ids, int* phases, int* num_tokens, int max_out) {
  Sched* s = static_cast<Sched*>(h);
  std::lock_guard<std::mutex> lk(s->mu);
  s->admit();
  int n = 0;
  // prefills first so that new sequences produce their first token this step, then decodes
  for (auto& r : s->running) {
    if (n >= max_out) break;
    if (!r.prefilled) {
      ids[n] = r.id;
      phases[n] = LLMCTL_PHASE_PREFILL;
      num_tokens[n] = r.prompt_len + r.generated;
      r.prefilled = true;
      ++n;
    }
  }
  for (auto& r : s->running) {
    if (n >= max_out) break;
    bool listed = false;
    for (int i = 0; i < n; ++i) listed |= (ids[i] == r.id);
    if (listed) continue;
    ids[n] = r.id;
    phases[n] = LLMCTL_PHASE_DECODE;
    num_tokens[n] = 1;
    ++n;
  }
  return n;
}

API int llmctl_sched_on_token(void* h, long long id) {
  Sched* s = static_cast<Sched*>(h);
  std::lock_guard<std::mutex> lk(s->mu);
  for (auto& r : s->running) {
    if (r.id != id) continue;
    r.generated += 1;
    if (s->kv) {
      const int need = r.prompt_len + r.generated;
      while (llmctl_kv_allocate(s->kv, id, need) != 0) {
        if (!s->preempt_one(id)) {
          // nothing else to evict: this sequence itself cannot grow -> preempt it
          for (int i = 0; i < static_cast<int>(s->running.size()); ++i) {
            if (s->running[i].id == id) {
              Req me = s->running[i];
              s->running.erase(s->running.begin() + i);
              s->reserved_tokens -= s->reserve(me);
              llmctl_kv_free(s->kv, id);
              me.prefilled = false;
              s->preempted.push_back(id);
              s->waiting.push_front(me);
              break;
            }
          }
          return -1;
        }
      }
    }
    return 0;
  }
  // evicted earlier in this step (by another sequence's growth) after its logits were computed: the token it just
  // produced is valid and must be counted, or re-admission would allocate one token too few
  for (auto& r : s->waiting) {
    if (r.id == id) {
      r.generated += 1;
      break;
    }
  }
  return -1;
}

API void llmctl_sched_finish(void* h, long long id) {
  Sched* s = static_cast<Sched*>(h);
  std::lock_guard<std::mutex> lk(s->mu);
  for (int i = 0; i < static_cast<int>(s->running.size()); ++i) {
    if (s->running[i].id == id) {
      s->reserved_tokens -= s->reserve(s->running[i]);
      s->running.erase(s->running.begin() + i);
      if (s->kv) llmctl_kv_free(s->kv, id);
      return;
    }
  }
  for (auto it = s->waiting.begin(); it != s->waiting.end(); ++it) {
    if (it->id == id) {
      s->waiting.erase(it);
      return;
    }
  }
}

API int llmctl_sched_num_waiting(void* h) {
  Sched* s = static_cast<Sched*>(h);
  std::lock_guard<std::mutex> lk(s->mu);
  return static_cast<int>(s->waiting.size());
}
API int llmctl_sched_num_running(void* h) {
  Sched* s = static_cast<Sched*>(h);
  std::lock_guard<std::mutex> lk(s->mu);
  return static_cast<int>(s->running.size());
}
API int llmctl_sched_pop_preempted(void* h, long long* ids, int max_out) {
  Sched* s = static_cast<Sched*>(h);
  std::lock_guard<std::mutex> lk(s->mu);
  const int n = std::min<int>(max_out, static_cast<int>(s->preempted.size()));
  for (int i = 0; i < n; ++i) ids[i] = s->preempted[i];
  s->preempted.erase(s->preempted.begin(), s->preempted.begin() + n);
  return n;
}
