// caches.h — domain caches: which cache block (slot of the surface buffer) holds which domain.
// Reference: LruCache (src/render/lru_cache.h:34-81, lru_cache.cc:44-171) and InfiniteCache
// (src/render/infinite_cache.h, infinite_cache.cc:31-60).  Same interface and the same hit/miss and
// eviction sequences; the LRU list is an intrusive doubly linked list over the blocks (O(1), no
// std::list / std::map nodes).
#pragma once
#include <vector>

#include "spray/check.h"

namespace spray {

struct CacheBlock {
  int block;
  int domain;
};

class LruCache {
 public:
  // cache_size < 0 or >= num_domains: every domain gets its own block (lru_cache.cc:50-54)
  void init(int num_domains, int cache_size, bool insitu_mode = false) {
    (void)insitu_mode;
    SPRAY_CHECK(cache_size != 0, "cache size 0");
    capacity_ = (cache_size < 0 || cache_size >= num_domains) ? num_domains : cache_size;
    SPRAY_CHECK(capacity_ <= num_domains && capacity_ > 0, "");
    ndomains_ = num_domains;
    SPRAY_CHECK(ndomains_ > 0, "");
    size_ = 0;
    block_of_.assign((std::size_t)ndomains_, -1);
    domain_of_.assign((std::size_t)capacity_, -1);
    prev_.assign((std::size_t)capacity_, -1);
    next_.assign((std::size_t)capacity_, -1);
    lru_ = mru_ = -1;
  }

  // returns true on a hit; *cache_block_id = the block holding (hit) or now assigned to (miss) domid.
  // On a miss with a full cache the least recently used block is recycled (lru_cache.cc:124-150).
  bool load(int domid, int* cache_block_id) {
    SPRAY_CHECK(domid >= 0 && domid < ndomains_, "domain id " << domid);
    int b = block_of_[(std::size_t)domid];
    const bool hit = b >= 0;
    if (hit) {
      unlink(b);
    } else if (size_ < capacity_) {
      b = size_++;
    } else {
      b = lru_;
      unlink(b);
      block_of_[(std::size_t)domain_of_[(std::size_t)b]] = -1;
    }
    domain_of_[(std::size_t)b] = domid;
    block_of_[(std::size_t)domid] = b;
    pushMru(b);
    *cache_block_id = b;
    return hit;
  }

  int getCacheSize() const { return capacity_; }
  int getSize() const { return size_; }
  bool isLoaded(int domid) const { return block_of_[(std::size_t)domid] >= 0; }
  // blocks from least to most recently used (the reference's std::list order, front = LRU)
  std::vector<CacheBlock> getBlocks() const {
    std::vector<CacheBlock> out;
    for (int b = lru_; b >= 0; b = next_[(std::size_t)b]) out.push_back({b, domain_of_[(std::size_t)b]});
    return out;
  }

 private:
  void unlink(int b) {
    const int p = prev_[(std::size_t)b], n = next_[(std::size_t)b];
    if (p >= 0) next_[(std::size_t)p] = n; else lru_ = n;
    if (n >= 0) prev_[(std::size_t)n] = p; else mru_ = p;
    prev_[(std::size_t)b] = next_[(std::size_t)b] = -1;
  }
  void pushMru(int b) {
    prev_[(std::size_t)b] = mru_;
    next_[(std::size_t)b] = -1;
    if (mru_ >= 0) next_[(std::size_t)mru_] = b; else lru_ = b;
    mru_ = b;
  }

  int size_ = 0, capacity_ = 0, ndomains_ = 0;
  int lru_ = -1, mru_ = -1;
  std::vector<int> block_of_;   // domain -> block or -1
  std::vector<int> domain_of_;  // block -> domain
  std::vector<int> prev_, next_;
};

class InfiniteCache {
 public:
  void init(int num_domains, int cache_size, bool insitu_mode = false) {
    SPRAY_CHECK(cache_size < 0 || cache_size >= num_domains || insitu_mode, "InfiniteCache needs room for every domain");
    capacity_ = num_domains;
    loaded_.assign((std::size_t)capacity_, 0);
  }
  // block id == domain id (infinite_cache.cc:47-60)
  bool load(int domid, int* cache_block_id) {
    SPRAY_CHECK(domid >= 0 && domid < capacity_, "domain id " << domid);
    *cache_block_id = domid;
    if (loaded_[(std::size_t)domid]) return true;
    loaded_[(std::size_t)domid] = 1;
    return false;
  }
  int getCacheSize() const { return capacity_; }
  bool isLoaded(int domid) const { return loaded_[(std::size_t)domid] != 0; }

 private:
  int capacity_ = 0;
  std::vector<char> loaded_;
};

}  // namespace spray
