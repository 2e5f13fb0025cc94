// prefix_index.hpp -- KV page allocator with prefix reuse, host side, header only.
//
// The reference specifies a KVCacheManager whose get_prefix(tokens) returns cached K/V for the longest cached
// prefix of a request (.kiro/specs/distributed-inference-server/design.md:363-413; Property 9 `:734-738`,
// Property 10 LRU eviction) and stores them as host Vec<f32> copies (`:386-390`).  Here the cache IS the paged
// device pool: a full page of page_size tokens is immutable once written, so a later request whose prompt starts
// with the same tokens maps the same physical pages into its page table and prefills only the rest.  Nothing is
// copied; "eviction" is handing a page that nobody references back to the allocator.
//
//   key of page i of a sequence = 128-bit chain hash over tokens[0 .. (i+1)*page_size)  (+ exact compare of the
//   page's own tokens on lookup)
//   page states: FREE -> OWNED (private to one sequence) -> CACHED (registered; refcount = sequences mapping it;
//   refcount 0 = evictable, kept in LRU order) -> FREE
//   a sequence's pages are released last-to-first, so a chain's deeper pages sit nearer the eviction end than
//   their parents: what remains cached is always a usable prefix.
#pragma once

#include <cstdint>
#include <cstring>
#include <list>
#include <unordered_map>
#include <vector>

namespace dis {

struct PrefixStats {
  int64_t cached_pages = 0, evictable_pages = 0, free_pages = 0;
  uint64_t lookups = 0, hits = 0, lookup_tokens = 0, hit_tokens = 0, evictions = 0;
};

class PrefixIndex {
 public:
  void init(int page_size, int64_t n_pages, bool enabled) {
    ps_ = page_size; enabled_ = enabled;
    meta_.assign((size_t)n_pages, Page());
    free_.clear(); free_.reserve((size_t)n_pages);
    for (int64_t p = n_pages - 1; p >= 0; p--) free_.push_back((int)p);  // page 0 is handed out first
    map_.clear(); lru_.clear();
    stats_ = PrefixStats();
  }
  bool enabled() const { return enabled_; }
  int64_t available() const { return (int64_t)free_.size() + (int64_t)lru_.size(); }

  // A page for private use: a free one, else the least recently used unreferenced cached page.
  bool alloc(int *out) {
    if (!free_.empty()) { *out = free_.back(); free_.pop_back(); meta_[(size_t)*out].state = OWNED; return true; }
    if (lru_.empty()) return false;
    const int p = lru_.front();
    unregister(p);
    stats_.evictions++;
    meta_[(size_t)p].state = OWNED;
    *out = p;
    return true;
  }

  // Longest cached prefix of tokens[0..n) in whole pages, leaving at least one token to compute (its logits
  // are what the caller is after).  Takes a reference on every matched page.  Returns the number of pages.
  int match(const uint32_t *tokens, int n, std::vector<int> *pages) {
    if (!enabled_) return 0;
    stats_.lookups++;
    stats_.lookup_tokens += (uint64_t)n;
    const int limit = (n - 1) / ps_;
    Key k{0x9E3779B97F4A7C15ull, 0xC2B2AE3D27D4EB4Full};
    int got = 0;
    for (int i = 0; i < limit; i++) {
      k = chain(k, tokens + (size_t)i * ps_);
      const int p = find(k, tokens + (size_t)i * ps_);
      if (p < 0) break;
      Page &m = meta_[(size_t)p];
      if (m.refs++ == 0) lru_.erase(m.lru_it);
      pages->push_back(p);
      got++;
    }
    if (got) { stats_.hits++; stats_.hit_tokens += (uint64_t)got * ps_; }
    return got;
  }

  // Register the full pages of a sequence (tokens[0 .. n_full*page_size) live in pages[0..n_full)).  Pages that
  // are already registered (matched earlier, or registered by a previous call) are skipped; if another sequence
  // registered the same content first, ours stays private.
  void commit(const uint32_t *tokens, int n_full, const std::vector<int> &pages) {
    if (!enabled_) return;
    Key k{0x9E3779B97F4A7C15ull, 0xC2B2AE3D27D4EB4Full};
    for (int i = 0; i < n_full && i < (int)pages.size(); i++) {
      k = chain(k, tokens + (size_t)i * ps_);
      Page &m = meta_[(size_t)pages[(size_t)i]];
      if (m.state == CACHED) continue;
      if (find(k, tokens + (size_t)i * ps_) >= 0) continue;
      m.state = CACHED; m.refs = 1; m.key = k;
      m.tokens.assign(tokens + (size_t)i * ps_, tokens + (size_t)(i + 1) * ps_);
      map_[k.a].push_back(pages[(size_t)i]);
    }
  }

  // A sequence lets go of its pages (in sequence order).
  void release(const std::vector<int> &pages) {
    for (size_t i = pages.size(); i-- > 0;) {
      const int p = pages[i];
      Page &m = meta_[(size_t)p];
      if (m.state == CACHED) {
        if (--m.refs == 0) { lru_.push_back(p); m.lru_it = std::prev(lru_.end()); }
      } else {
        m.state = FREE;
        free_.push_back(p);
      }
    }
  }

  PrefixStats stats() const {
    PrefixStats s = stats_;
    s.free_pages = (int64_t)free_.size();
    s.evictable_pages = (int64_t)lru_.size();
    int64_t c = 0;
    for (auto &b : map_) c += (int64_t)b.second.size();
    s.cached_pages = c;
    return s;
  }

 private:
  enum State : uint8_t { FREE = 0, OWNED, CACHED };
  struct Key { uint64_t a, b; };
  struct Page {
    State state = FREE;
    int refs = 0;
    Key key{0, 0};
    std::vector<uint32_t> tokens;
    std::list<int>::iterator lru_it;
  };
  static uint64_t mix(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  Key chain(Key k, const uint32_t *t) const {
    for (int i = 0; i < ps_; i++) {
      k.a = mix(k.a + 0x9E3779B97F4A7C15ull + t[i]);
      k.b = mix((k.b ^ t[i]) * 0xD6E8FEB86659FD93ull + (uint64_t)i);
    }
    return k;
  }
  int find(const Key &k, const uint32_t *t) const {
    auto it = map_.find(k.a);
    if (it == map_.end()) return -1;
    for (int p : it->second) {
      const Page &m = meta_[(size_t)p];
      if (m.key.b == k.b && memcmp(m.tokens.data(), t, (size_t)ps_ * sizeof(uint32_t)) == 0) return p;
    }
    return -1;
  }
  void unregister(int p) {
    Page &m = meta_[(size_t)p];
    lru_.erase(m.lru_it);
    auto it = map_.find(m.key.a);
    if (it != map_.end()) {
      auto &v = it->second;
      for (size_t i = 0; i < v.size(); i++) if (v[i] == p) { v[i] = v.back(); v.pop_back(); break; }
      if (v.empty()) map_.erase(it);
    }
    m.tokens.clear(); m.tokens.shrink_to_fit();
    m.refs = 0;
  }

  int ps_ = 16;
  bool enabled_ = false;
  std::vector<Page> meta_;
  std::vector<int> free_;
  std::list<int> lru_;  // front = next to evict
  std::unordered_map<uint64_t, std::vector<int>> map_;
  PrefixStats stats_;
};

}  // namespace dis
