// cluster_image.hpp — one cluster's informer caches as the host layer keeps them, and the packer
// that lays them out as an arena (include/ncdrift.h "arenas").
//
// Stands in for the four listers of a cluster (controller.go:268-278; shards.Shard's listers): an
// insertion-ordered store keyed by (kind, "namespace/name").  Informer events map onto three costs:
//   value of an existing object changed (UpdateFunc)   -> ncd_arena_sync_records: that record only
//   object added / deleted / re-owned (Add/DeleteFunc) -> ncd_arena_recommit: small arrays + new records
//   first scan, or the arena ran out of room           -> ncd_arena_commit of a fresh, larger arena
// Every put / remove that changes something is stamped with the checker's event counter; the
// stamps are what Lookup's freshness rule reads (include/ncdhost.h).
#pragma once
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <condition_variable>
#include <functional>
#include <iterator>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../../include/ncdrift.h"

namespace ncdh {

struct OwnerRef {
  std::string kind, uid;
  bool operator==(const OwnerRef& o) const { return kind == o.kind && uid == o.uid; }
};

// Everything about an object except its compared value.  Immutable once made: a put that changes
// any of it makes a new one, so snapshots can share it with the live image without copying.
struct ObjInfo {
  uint32_t kind = 0;
  std::string key;  // "namespace/name" (cache.ObjectName.String())
  std::string name, uid;
  std::vector<OwnerRef> owners;
  std::vector<std::string> secrets, configmaps;  // GetSecretNames() / GetConfigMapNames() (templates)
  bool same_structure(const ObjInfo& o) const {
    return uid == o.uid && owners == o.owners && secrets == o.secrets && configmaps == o.configmaps;
  }
};

// The compared value's canonical bytes live in exactly one place: in memory the event brought
// (`pending`, kept alive by `hold`: a batch's slab or a single buffer) until the next Scan lays the
// record out in the arena's pinned payload, and only there afterwards.  The pinned arena IS the
// cache of values: no second host copy is kept.  A batch may also write a new value that fits the
// record's place straight into the pinned payload (`direct`: no pending bytes, nothing for the
// packer to copy); `claimed` is the right to read or write the record's pinned bytes outside the
// exclusive cache lock, taken and released with the helpers below.
struct CachedObject {
  std::shared_ptr<const ObjInfo> info;
  uint32_t len = 0;                  // canonical length of the current value
  const uint8_t* pending = nullptr;  // its bytes while they are not in the arena yet
  std::shared_ptr<const void> hold;  // owner of `pending`
  bool live = true;
  uint64_t gen = 0;  // event counter value of the last put that changed this object
  // where the record lives in the resident arena's payload (-1: not placed yet)
  int64_t rec_off = -1;
  uint64_t rec_cap = 0;  // padded bytes reserved at rec_off
  bool direct = false;   // the pinned record was rewritten in place since the last Scan: the device copy is behind
  uint8_t claimed = 0;   // accessed with __atomic builtins only
};
inline bool try_claim(const CachedObject& o) {
  return __atomic_exchange_n(const_cast<uint8_t*>(&o.claimed), (uint8_t)1, __ATOMIC_ACQUIRE) == 0;
}
inline void release_claim(const CachedObject& o) { __atomic_store_n(const_cast<uint8_t*>(&o.claimed), (uint8_t)0, __ATOMIC_RELEASE); }

inline std::string store_key(uint32_t kind, const std::string& key) { return std::string(1, (char)('0' + kind)) + key; }
inline uint64_t pad16(uint64_t n) { return (n + 15) & ~15ull; }
inline std::string ncd_err(const char* op) { return std::string(op) + ": " + ncd_last_error(); }

struct ArenaHandle {
  ncd_arena* a = nullptr;
  ArenaHandle() = default;
  ArenaHandle(const ArenaHandle&) = delete;
  ArenaHandle& operator=(const ArenaHandle&) = delete;
  ArenaHandle(ArenaHandle&& o) noexcept : a(o.a) { o.a = nullptr; }
  void reset() {
    if (a) ncd_arena_destroy(a);
    a = nullptr;
  }
  ~ArenaHandle() { reset(); }
};

// A small pool of worker threads shared by every checker of the process: the canonicaliser, the
// packer's record copies and the per-cluster cache updates are short parallel loops run many times
// a second, and creating threads for each of them cost more than the loops themselves.  One loop at
// a time (callers queue on job_mu; lock order is always cache lock, then job_mu); the calling thread
// works too.  Never destroyed: the workers sleep on a condition variable until the process exits.
class Workers {
 public:
  static Workers& get() {
    static Workers* w = new Workers();
    return *w;
  }
  int size() const { return (int)threads_.size() + 1; }

  // fn(chunk) for chunk in [0, chunks), on up to `width` threads including the caller.
  template <class Fn>
  void run(size_t chunks, int width, Fn&& fn) {
    if (chunks == 0) return;
    if (chunks == 1 || width <= 1 || threads_.empty()) {
      for (size_t k = 0; k < chunks; ++k) fn(k);
      return;
    }
    std::lock_guard<std::mutex> job(job_mu_);
    std::exception_ptr err;
    std::mutex err_mu;
    std::function<void(size_t)> body = [&](size_t k) {
      try {
        fn(k);
      } catch (...) {
        std::lock_guard<std::mutex> lk(err_mu);
        if (!err) err = std::current_exception();
      }
    };
    {
      std::lock_guard<std::mutex> lk(mu_);
      body_ = &body;
      chunks_ = chunks;
      next_ = 0;
      done_ = 0;
      helpers_ = std::min<size_t>((size_t)width - 1, std::min(threads_.size(), chunks - 1));
      woken_ = 0;
      ++epoch_;
    }
    cv_.notify_all();
    work(body, chunks);
    {
      std::unique_lock<std::mutex> lk(mu_);
      idle_.wait(lk, [&] { return done_ == chunks_ && active_ == 0; });
      body_ = nullptr;
    }
    if (err) std::rethrow_exception(err);
  }

 private:
  Workers() {
    unsigned hc = std::thread::hardware_concurrency();
    if (hc == 0) hc = 4;
    if (hc > 32) hc = 32;
    for (unsigned i = 1; i < hc; ++i) threads_.emplace_back([this] { loop(); });
    for (auto& t : threads_) t.detach();
  }
  void work(std::function<void(size_t)>& body, size_t chunks) {
    for (;;) {
      size_t k;
      {
        std::lock_guard<std::mutex> lk(mu_);
        if (next_ >= chunks) return;
        k = next_++;
      }
      body(k);
      std::lock_guard<std::mutex> lk(mu_);
      if (++done_ == chunks_) idle_.notify_all();
    }
  }
  void loop() {
    uint64_t seen = 0;
    for (;;) {
      std::function<void(size_t)>* body;
      size_t chunks;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return epoch_ != seen; });
        seen = epoch_;
        if (!body_ || woken_ >= helpers_) continue;  // this loop needs fewer helpers than there are workers
        ++woken_;
        ++active_;
        body = body_;
        chunks = chunks_;
      }
      work(*body, chunks);
      std::lock_guard<std::mutex> lk(mu_);
      if (--active_ == 0) idle_.notify_all();
    }
  }
  std::vector<std::thread> threads_;
  std::mutex job_mu_, mu_;
  std::condition_variable cv_, idle_;
  std::function<void(size_t)>* body_ = nullptr;
  size_t chunks_ = 0, next_ = 0, done_ = 0, helpers_ = 0, woken_ = 0, active_ = 0;
  uint64_t epoch_ = 0;
};

// Runs fn(i) for i in [0, n) on up to `threads` threads (the packer's record copies and the
// canonicaliser are memory-bound loops over independent records), in contiguous slices.
template <class Fn>
void parallel_for(size_t n, int threads, size_t min_per_thread, Fn fn) {
  size_t t = threads < 1 ? 1 : (size_t)threads;
  if (min_per_thread && n / min_per_thread < t) t = n / min_per_thread;
  if (t <= 1) {
    for (size_t i = 0; i < n; ++i) fn(i);
    return;
  }
  const size_t slices = t * 4;  // finer than the thread count: a slow thread does not hold the loop up
  Workers::get().run(slices, (int)t, [&](size_t k) {
    for (size_t i = n * k / slices; i < n * (k + 1) / slices; ++i) fn(i);
  });
}

// What one Scan has to send to the GPU for a cluster, decided while the cache lock is held and
// carried out after it is released.
enum class Upload { None, Values, Structural, Full };

struct ClusterImage {
  std::vector<CachedObject> objs;
  std::unordered_map<std::string, size_t> index;   // store_key -> position in objs (live objects)
  std::unordered_map<std::string, uint64_t> tomb;  // store_key -> stamp of the removal
  std::vector<uint32_t> dirty;                     // positions whose value changed since the last Scan
  uint64_t last_change = 0;                        // stamp of the latest put / remove that changed anything
  uint64_t layout = 0;                             // bumped whenever positions in objs are reassigned
  ArenaHandle arena;
  const uint8_t* host_payload = nullptr;           // the arena's pinned payload (where placed records are read back from)
  uint32_t* host_rec_len = nullptr;                // ... and its pinned rec_len array
  uint64_t payload_used = 0, payload_cap = 0, nobj_cap = 0, aux_cap = 0, nowner_cap = 0, ndep_cap = 0;
  bool structural = true;   // keys / owners / dependents changed since the last commit
  bool force_full = false;  // the resident device copy is not trusted: upload everything again
  uint64_t uploads_full = 0, uploads_structural = 0, uploads_values = 0, compactions = 0;
  uint64_t in_place_last = 0;  // records of the last prepared Values upload that needed no copy
  uint64_t pulled_early = 0;   // records handed to the device as their batch was applied, since the last Scan
  uint64_t payload_in_use() const { return payload_used; }
  std::vector<uint32_t> pending_changed;  // records a prepared Values upload still has to push

  const uint8_t* bytes_of(const CachedObject& o) const { return o.pending ? o.pending : host_payload + o.rec_off; }

  // What a batch found out about an event ahead of the update, in parallel and under the SHARED lock
  // (looking the object up and comparing the value's bytes are the expensive parts of a put): valid
  // only while positions were not reassigned (same layout) and the object has not changed (same gen).
  struct PutHint {
    bool found = false, same_value = false, same_structure = false;
    uint64_t seen_gen = 0, layout = 0;
    size_t pos = 0;
    uint32_t first_diff = 0;  // the record's pinned bytes equal the new value up to here (a multiple of 256)
  };
  // Caller holds the claim on the object at `pos` (its bytes may be in the pinned payload, which
  // another batch could otherwise be rewriting).
  PutHint precheck(size_t pos, bool same_structure, const uint8_t* bytes, uint32_t len) const {
    PutHint h;
    const CachedObject& o = objs[pos];
    h.found = true;
    h.pos = pos;
    h.layout = layout;
    h.seen_gen = o.gen;
    h.same_value = false;
    if (o.len == len) {  // block by block: where the first difference is tells write_in_place what it can skip
      const uint8_t* have = bytes_of(o);
      uint32_t at = 0;
      while (at < len) {
        const uint32_t n = len - at < 256u ? len - at : 256u;
        if (memcmp(have + at, bytes + at, n) != 0) break;
        at += n;
      }
      h.same_value = at == len;
      if (!o.pending) h.first_diff = at & ~255u;  // only the pinned record itself can be kept in part
    }
    h.same_structure = same_structure;
    return h;
  }
  // True when a value of up to `len` bytes can be written over the object's record in the pinned
  // payload: it has a place there and the place is large enough.
  bool fits_in_place(const CachedObject& o, uint64_t len) const {
    return arena.a && host_payload && o.rec_off >= 0 && pad16(len) <= o.rec_cap;
  }
  // Caller holds the claim and the pinned-memory lock (no Scan is reading the payload).
  // from: the pinned bytes before it already equal the new value (PutHint::first_diff).
  void write_in_place(const CachedObject& o, const uint8_t* bytes, uint32_t len, uint32_t from = 0) const {
    uint8_t* dst = const_cast<uint8_t*>(host_payload) + o.rec_off;
    memcpy(dst + from, bytes + from, len - from);
    memset(dst + len, 0, pad16(len) - len);
  }

  // ---- events (caller holds the cache lock exclusively)
  // Returns true when the put changed anything.  k = store_key(kind, key); bytes / len = the canonical
  // value, owned by `hold`.
  // force: the value counts as changed whatever the cache holds (its pinned bytes were overwritten).
  bool put(const std::string& k, std::shared_ptr<const ObjInfo> info, const uint8_t* bytes, uint32_t len,
           const std::shared_ptr<const void>& hold, uint64_t stamp, const PutHint* hint = nullptr, bool force = false) {
    size_t pos = 0;
    bool hinted = hint && hint_valid(*hint);
    if (hinted) {
      pos = hint->pos;
    } else {
      auto it = index.find(k);
      if (it == index.end()) {
        index.emplace(k, objs.size());
        CachedObject o;
        o.info = std::move(info);
        o.len = len;
        o.pending = bytes;
        o.hold = hold;
        o.gen = stamp;
        objs.push_back(std::move(o));
        tomb.erase(k);
        structural = true;
        last_change = stamp;
        return true;
      }
      pos = it->second;
    }
    CachedObject& old = objs[pos];
    bool changed = false;
    if (!(hinted ? hint->same_structure : old.info->same_structure(*info))) {
      structural = true;
      changed = true;
      old.info = std::move(info);
    }
    if (force || !(hinted ? hint->same_value : (old.len == len && memcmp(bytes_of(old), bytes, len) == 0))) {
      if (!old.pending && !old.direct) dirty.push_back((uint32_t)pos);
      old.pending = bytes;
      old.hold = hold;
      old.len = len;
      changed = true;
    }
    if (changed) old.gen = last_change = stamp;
    return changed;
  }
  // The put of a value that phase one of a batch already wrote over the record's pinned bytes
  // (write_in_place).  False when the hint no longer describes the object — something else was
  // applied to it in between — and the caller has to put the value the ordinary way, with force.
  bool hint_valid(const PutHint& hint) const {
    return hint.found && hint.layout == layout && hint.pos < objs.size() && objs[hint.pos].live && objs[hint.pos].gen == hint.seen_gen;
  }
  // pull_now != null: a record that was clean until now is not left for the next Scan's lay-out — its new
  // length goes into the pinned rec_len array here and its position onto *pull_now, for the caller to have
  // it pulled to the device right away (the record stays clean).
  bool put_direct(std::shared_ptr<const ObjInfo>& info, uint32_t len, uint64_t stamp, const PutHint& hint,
                  std::vector<uint32_t>* pull_now = nullptr) {
    if (!hint_valid(hint)) return false;
    CachedObject& o = objs[hint.pos];
    if (!hint.same_structure) {
      structural = true;
      o.info = std::move(info);
    }
    const bool clean = !o.pending && !o.direct;
    o.pending = nullptr;  // an older value still waiting for the packer is superseded
    o.hold.reset();
    o.len = len;
    o.gen = last_change = stamp;
    if (clean && pull_now && !force_full && host_rec_len) {
      host_rec_len[hint.pos] = len;
      pull_now->push_back((uint32_t)hint.pos);
      ++pulled_early;
      return true;
    }
    if (clean) dirty.push_back((uint32_t)hint.pos);
    o.direct = true;
    return true;
  }
  // The same for a caller that applies a cluster's in-place events on SEVERAL threads at once (the objects are
  // distinct and each is claimed by the caller): only the object itself is touched here; what belongs to the
  // image as a whole is collected in `acc` and added by merge_direct, on one thread.
  struct DirectAcc {
    std::vector<uint32_t> pull, dirty;
    uint64_t last_change = 0, pulled = 0;
    bool structural = false;
  };
  bool put_direct_mt(std::shared_ptr<const ObjInfo>& info, uint32_t len, uint64_t stamp, const PutHint& hint, bool pull_now,
                     DirectAcc& acc) {
    if (!hint_valid(hint)) return false;
    CachedObject& o = objs[hint.pos];
    if (!hint.same_structure) {
      acc.structural = true;
      o.info = std::move(info);
    }
    const bool clean = !o.pending && !o.direct;
    o.pending = nullptr;
    o.hold.reset();
    o.len = len;
    o.gen = stamp;
    if (stamp > acc.last_change) acc.last_change = stamp;
    if (clean && pull_now && !force_full && host_rec_len) {
      host_rec_len[hint.pos] = len;
      acc.pull.push_back((uint32_t)hint.pos);
      ++acc.pulled;
      return true;
    }
    if (clean) acc.dirty.push_back((uint32_t)hint.pos);
    o.direct = true;
    return true;
  }
  void merge_direct(const DirectAcc& acc) {
    dirty.insert(dirty.end(), acc.dirty.begin(), acc.dirty.end());
    if (acc.last_change > last_change) last_change = acc.last_change;
    pulled_early += acc.pulled;
    structural = structural || acc.structural;
  }
  bool drop(uint32_t kind, const std::string& key, uint64_t stamp) { return drop(store_key(kind, key), stamp); }
  bool drop(const std::string& k, uint64_t stamp) {
    auto it = index.find(k);
    if (it == index.end()) return false;
    objs[it->second].live = false;
    index.erase(it);
    tomb[k] = stamp;
    structural = true;
    last_change = stamp;
    return true;
  }
  const CachedObject* get(uint32_t kind, const std::string& key) const {
    auto it = index.find(store_key(kind, key));
    return it == index.end() ? nullptr : &objs[it->second];
  }
  // Stamp of the last event that touched this key (0: none the image remembers).
  uint64_t changed_at(const std::string& skey) const {
    auto it = index.find(skey);
    if (it != index.end()) return objs[it->second].gen;
    auto t = tomb.find(skey);
    return t == tomb.end() ? 0 : t->second;
  }
  // After a failed Scan: the resident DEVICE copy is not trusted any more.  The pinned arena keeps
  // the values (it is their only copy), so the next Scan commits the same arena again in full.
  void invalidate() {
    structural = true;
    force_full = true;
    pending_changed.clear();
  }

  // deferred != null: only the cheap case is handled (values rewritten inside a resident arena that has
  // room for them) — anything that makes or compacts an arena or rebuilds the per-object arrays sets
  // *deferred and returns without having changed anything, for the caller to run alone later.
  Upload prepare(ncd_ctx* ctx, int threads, uint64_t tomb_floor, uint64_t* h2d, bool* deferred = nullptr);
  void upload(Upload what);
  // A prepared Full upload, for the caller that commits several clusters in one ncd_arenas_commit.
  void full_counts(uint64_t out[5]) const {
    out[0] = objs_committed;
    out[1] = payload_used;
    out[2] = aux_used;
    out[3] = own_used;
    out[4] = dep_used;
  }
  void full_done() { force_full = false; }

 private:
  uint64_t aux_used = 0, own_used = 0, dep_used = 0, objs_committed = 0, dirty_from = 0;
  ncd_arena_view host_view();
  Upload prepare_structural(ncd_ctx* ctx, int threads, uint64_t* h2d, bool regrow = false);
  void assign(CachedObject& o, bool force_append);
  static void copy_record(ncd_arena_view& v, CachedObject& o, const uint8_t* src);
  void bind_arena();
  bool reclaim(uint64_t append);
  void write_all(ncd_arena_view& v, int threads, const uint8_t* old_payload, bool replace_all);
};

}  // namespace ncdh
