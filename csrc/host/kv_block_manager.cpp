// Paged KV-cache block manager (host side).  See include/llmctl_host.h.
//
// Blocks are fixed-size token slots; a sequence owns an ordered list of block ids (its block table, which
// the paged-attention kernel K12 reads on the device).  Blocks are reference counted so forked sequences
// share a prefix; the free list is LIFO so recently used blocks (still warm in L2) are reused first.
//
// Reference: reference llmctl/serve/server.py:57-87 `KVCacheManager` is a request_id -> (k, v) dict with
// LRU eviction whose set_cache() is never called.  Nothing of it is reused.
#include "llmctl_host.h"

#include <algorithm>
#include <mutex>
#include <unordered_map>
#include <vector>

#define API extern "C" __attribute__((visibility("default")))

namespace {

struct Seq {
  std::vector<int> blocks;
  int num_tokens = 0;
};

struct KV {
  int num_blocks, block_size, watermark_blocks;
  std::vector<int> free_list;
  std::vector<int> refcnt;
  std::unordered_map<long long, Seq> seqs;
  std::mutex mu;

  KV(int n, int bs, float wm) : num_blocks(n), block_size(bs), refcnt(n, 0) {
    watermark_blocks = static_cast<int>(wm * n);
    free_list.reserve(n);
    for (int i = n - 1; i >= 0; --i) free_list.push_back(i);  // pop_back hands out 0, 1, 2, ...
  }
  int blocks_for(int tokens) const { return (tokens + block_size - 1) / block_size; }
};

}  // namespace

API void* llmctl_kv_create(int num_blocks, int block_size, float watermark) {
  if (num_blocks <= 0 || block_size <= 0 || watermark < 0.f || watermark >= 1.f) return nullptr;
  return new KV(num_blocks, block_size, watermark);
}
API void llmctl_kv_destroy(void* h) { delete static_cast<KV*>(h); }

API int llmctl_kv_allocate(void* h, long long seq_id, int num_tokens) {
  KV* kv = static_cast<KV*>(h);
  std::lock_guard<std::mutex> lk(kv->mu);
  Seq& s = kv->seqs[seq_id];
  const int need = kv->blocks_for(num_tokens) - static_cast<int>(s.blocks.size());
  if (need > static_cast<int>(kv->free_list.size())) {
    if (s.blocks.empty() && s.num_tokens == 0) kv->seqs.erase(seq_id);
    return -1;
  }
  for (int i = 0; i < need; ++i) {
    const int b = kv->free_list.back();
    kv->free_list.pop_back();
    kv->refcnt[b] = 1;
    s.blocks.push_back(b);
  }
  if (num_tokens > s.num_tokens) s.num_tokens = num_tokens;
  return 0;
}

API int llmctl_kv_free(void* h, long long seq_id) {
  KV* kv = static_cast<KV*>(h);
  std::lock_guard<std::mutex> lk(kv->mu);
  auto it = kv->seqs.find(seq_id);
  if (it == kv->seqs.end()) return 0;
  int released = 0;
  for (auto rit = it->second.blocks.rbegin(); rit != it->second.blocks.rend(); ++rit) {
    if (--kv->refcnt[*rit] == 0) {
      kv->free_list.push_back(*rit);
      ++released;
    }
  }
  kv->seqs.erase(it);
  return released;
}

API int llmctl_kv_num_tokens(void* h, long long seq_id) {
  KV* kv = static_cast<KV*>(h);
  std::lock_guard<std::mutex> lk(kv->mu);
  auto it = kv->seqs.find(seq_id);
  return it == kv->seqs.end() ? -1 : it->second.num_tokens;
}

API int llmctl_kv_block_table(void* h, long long seq_id, int* out, int max_out) {
  KV* kv = static_cast<KV*>(h);
  std::lock_guard<std::mutex> lk(kv->mu);
  auto it = kv->seqs.find(seq_id);
  if (it == kv->seqs.end()) return 0;
  const int n = static_cast<int>(it->second.blocks.size());
  for (int i = 0; i < n && i < max_out; ++i) out[i] = it->second.blocks[i];
  return n;
}

API int llmctl_kv_free_blocks(void* h) {
  KV* kv = static_cast<KV*>(h);
  std::lock_guard<std::mutex> lk(kv->mu);
  return static_cast<int>(kv->free_list.size());
}
API int llmctl_kv_total_blocks(void* h) { return static_cast<KV*>(h)->num_blocks; }
API int llmctl_kv_block_size(void* h) { return static_cast<KV*>(h)->block_size; }

API int llmctl_kv_can_allocate(void* h, int num_tokens) {
  KV* kv = static_cast<KV*>(h);
  std::lock_guard<std::mutex> lk(kv->mu);
  return static_cast<int>(kv->free_list.size()) - kv->blocks_for(num_tokens) >= kv->watermark_blocks ? 1 : 0;
}

API int llmctl_kv_fork(void* h, long long parent, long long child, int* cow_src, int* cow_dst) {
  KV* kv = static_cast<KV*>(h);
  std::lock_guard<std::mutex> lk(kv->mu);
  auto it = kv->seqs.find(parent);
  if (it == kv->seqs.end() || kv->seqs.count(child)) return -1;
  Seq c = it->second;  // copy the table
  const bool partial = (c.num_tokens % kv->block_size) != 0 && !c.blocks.empty();
  if (partial && kv->free_list.empty()) return -1;
  for (int b : c.blocks) ++kv->refcnt[b];
  int ret = 0;
  if (partial) {
    const int src = c.blocks.back();
    const int dst = kv->free_list.back();
    kv->free_list.pop_back();
    kv->refcnt[dst] = 1;
    --kv->refcnt[src];  // the child does not share the partial block
    c.blocks.back() = dst;
    if (cow_src) *cow_src = src;
    if (cow_dst) *cow_dst = dst;
    ret = 1;
  }
  kv->seqs.emplace(child, std::move(c));
  return ret;
}
