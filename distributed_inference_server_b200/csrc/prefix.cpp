// prefix.cpp -- the page index of prefix_index.hpp behind a C ABI of its own (include/dis_engine.h,
// dis_prefix_index_*), so that Property 9 (prefix reuse) and Property 10 (LRU eviction) of the reference's spec
// (.kiro/specs/distributed-inference-server/design.md:734-744) can be property-tested without a GPU.  The engine
// embeds the same class (engine.cu: seq_add / prefill / seq_remove).
#include <mutex>
#include <vector>

#include "../../include/dis_engine.h"
#include "prefix_index.hpp"

struct dis_prefix_index {
  dis::PrefixIndex ix;
  int page_size = 0;
  std::mutex mu;
};

static void fill_stats(const dis::PrefixStats &s, dis_prefix_stats *o) {
  o->cached_pages = s.cached_pages; o->evictable_pages = s.evictable_pages; o->free_pages = s.free_pages;
  o->lookups = s.lookups; o->hits = s.hits; o->lookup_tokens = s.lookup_tokens; o->hit_tokens = s.hit_tokens;
  o->evictions = s.evictions;
}

namespace dis { void prefix_fill_stats(const PrefixStats &s, dis_prefix_stats *o) { fill_stats(s, o); } }

extern "C" {

int dis_prefix_index_create(int32_t page_size, int64_t n_pages, dis_prefix_index **out) {
  if (!out || page_size <= 0 || n_pages <= 0) return DIS_ERR_INVALID_ARGUMENT;
  dis_prefix_index *p = new dis_prefix_index();
  p->page_size = page_size;
  p->ix.init(page_size, n_pages, true);
  *out = p;
  return DIS_OK;
}

int dis_prefix_index_destroy(dis_prefix_index *ix) {
  if (!ix) return DIS_ERR_INVALID_ARGUMENT;
  delete ix;
  return DIS_OK;
}

int dis_prefix_index_admit(dis_prefix_index *ix, const uint32_t *tokens, int32_t n_tokens, int32_t *pages,
                           int32_t pages_cap, int32_t *n_pages, int32_t *matched_tokens) {
  if (!ix || !tokens || n_tokens <= 0 || !pages || !n_pages || !matched_tokens) return DIS_ERR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lk(ix->mu);
  const int need = (n_tokens + 1 + ix->page_size - 1) / ix->page_size;
  if (need > pages_cap) return DIS_ERR_INVALID_ARGUMENT;
  std::vector<int> got;
  const int matched = ix->ix.match(tokens, n_tokens, &got);
  if (need - (int)got.size() > ix->ix.available()) { ix->ix.release(got); return DIS_ERR_OUT_OF_MEMORY; }  // nothing is evicted for a request that cannot be admitted
  while ((int)got.size() < need) {
    int p = -1;
    ix->ix.alloc(&p);
    got.push_back(p);
  }
  for (int i = 0; i < need; i++) pages[i] = got[(size_t)i];
  *n_pages = need;
  *matched_tokens = matched * ix->page_size;
  return DIS_OK;
}

int dis_prefix_index_commit(dis_prefix_index *ix, const uint32_t *tokens, int32_t n_tokens, const int32_t *pages,
                            int32_t n_pages) {
  if (!ix || !tokens || n_tokens < 0 || !pages || n_pages < 0) return DIS_ERR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lk(ix->mu);
  ix->ix.commit(tokens, n_tokens / ix->page_size, std::vector<int>(pages, pages + n_pages));
  return DIS_OK;
}

int dis_prefix_index_release(dis_prefix_index *ix, const int32_t *pages, int32_t n_pages) {
  if (!ix || !pages || n_pages < 0) return DIS_ERR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lk(ix->mu);
  ix->ix.release(std::vector<int>(pages, pages + n_pages));
  return DIS_OK;
}

int dis_prefix_index_stats(dis_prefix_index *ix, dis_prefix_stats *out) {
  if (!ix || !out) return DIS_ERR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lk(ix->mu);
  fill_stats(ix->ix.stats(), out);
  return DIS_OK;
}

}  // extern "C"
