// drift_checker.cpp — DriftChecker (Scan / Lookup) and handler replay over libncdrift.so.
//
// This is the C++ stand-in for the Go layer of INTEGRATION.md.  It decides nothing about drift:
// it lays the informer caches out as arenas (canonicalising the compared values), calls the
// scan on the GPU, and reads decision codes back.  The replay functions walk a template's /
// workgroup's slice of those codes in the order the reference's handlers walk the shards
// (controller.go:790-831 with :504-626; :720-745) and stop where the handler would return.
//
// Concurrency (the reference runs its handlers on `workers` goroutines, controller.go:875-877,
// while informer callbacks update the caches, :286-359):
//   cache_mu   shared_mutex over the cluster images.  Event writers take it exclusively for the
//              update; a Scan takes it exclusively only while it lays changed records out in
//              pinned memory; readers take it shared for the freshness check.
//   scan_mu    one Scan at a time; also serialises every call on the ncd_ctx.
//   snap       the result of the last finished Scan, an immutable object swapped in atomically:
//              Lookup / plan never wait for a Scan in flight and never see half a result.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../../include/ncdhost.h"
#include "../../../include/ncdrift.h"
#include "cluster_image.hpp"

namespace ncdh {
std::string canonicalise(const uint8_t* value, uint64_t len);
uint64_t canonicalise_into(const uint8_t* value, uint64_t len, uint8_t* out, uint64_t cap);
}
using namespace ncdh;

namespace {

thread_local std::string g_err;
int fail(const std::string& m) {
  g_err = m;
  return -1;
}
double ms_since(std::chrono::steady_clock::time_point t0) {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

// The controller objects the rows of a scan were made from, in arena order.  Shared between
// snapshots while the controller's structure does not change (value updates keep it).
struct RowDirectory {
  std::vector<std::shared_ptr<const ObjInfo>> objs;  // by controller arena index
  std::vector<uint64_t> row0;                        // first row of object i (templates / workgroups)
  std::unordered_map<std::string, uint32_t> by_key;  // store_key -> arena index, templates and workgroups
  uint64_t rows = 0;
};

// What one finished Scan published.
struct Snapshot {
  uint64_t clock = 0;  // event counter when the Scan began: everything stamped later is not in it
  size_t nshards = 0;
  uint64_t rows = 0;
  std::shared_ptr<const RowDirectory> dir;
  std::vector<ncd_entry> entries;       // the diff list, ordered by (obj, dep, shard); SKIP is absent
  std::vector<ncd_adoption> unadopted;  // ordered by (obj, dep)

  uint32_t code(uint32_t obj, uint32_t dep, uint32_t shard) const {
    auto lt = [](const ncd_entry& e, const ncd_entry& k) {
      return e.obj != k.obj ? e.obj < k.obj : (e.dep != k.dep ? e.dep < k.dep : e.shard < k.shard);
    };
    const ncd_entry key{obj, dep, shard, 0};
    auto it = std::lower_bound(entries.begin(), entries.end(), key, lt);
    return (it != entries.end() && it->obj == obj && it->dep == dep && it->shard == shard) ? it->code : (uint32_t)NCD_SKIP;
  }
  bool needs_adoption(uint32_t obj, uint32_t dep) const {
    auto lt = [](const ncd_adoption& e, const ncd_adoption& k) { return e.obj != k.obj ? e.obj < k.obj : e.dep < k.dep; };
    const ncd_adoption key{obj, dep};
    auto it = std::lower_bound(unadopted.begin(), unadopted.end(), key, lt);
    return it != unadopted.end() && it->obj == obj && it->dep == dep;
  }
};

}  // namespace

// Pulls records to the device while the batch that rewrote them is still being applied: a batch is
// worked off in chunks, and while the host canonicalises chunk k + 1 this thread has the GPU fetch
// the records chunk k wrote in place (ncd_arenas_sync_records, synchronous in here only).  One job at
// a time, in order.  A Scan drains it before it touches the arenas.
struct Puller {
  // The batch hands over its claims on the records with them: nobody writes a record's pinned bytes
  // while the GPU reads them; the claims are released when the pull is done.
  struct Job {
    std::vector<ClusterImage*> images;
    std::vector<ncd_arena*> arenas;
    std::vector<std::vector<uint32_t>> lists;  // positions in images[q]->objs == object numbers in arenas[q]
    std::vector<uint64_t> payload_bytes;
  };
  std::mutex mu;
  std::condition_variable cv, idle;
  std::deque<Job> jobs;
  bool busy = false, stop = false, started = false;
  std::string error;           // first failure since the last drain
  uint64_t records = 0, bytes = 0, calls = 0;
  double busy_ms = 0;
  std::thread th;
  std::mutex* gpu_mu = nullptr;  // library calls on the checker's context are made one at a time
  std::shared_mutex* cache_mu = nullptr;

  void submit(Job&& j) {
    std::lock_guard<std::mutex> lk(mu);
    if (!started) {
      started = true;
      th = std::thread([this] { loop(); });
    }
    jobs.push_back(std::move(j));
    cv.notify_one();
  }
  // Waits until everything submitted has been pulled; returns the first error since the last drain.
  std::string drain() {
    std::unique_lock<std::mutex> lk(mu);
    idle.wait(lk, [&] { return jobs.empty() && !busy; });
    std::string e;
    e.swap(error);
    return e;
  }
  void shutdown() {
    {
      std::lock_guard<std::mutex> lk(mu);
      stop = true;
      cv.notify_all();
    }
    if (th.joinable()) th.join();
  }
  void loop() {
    for (;;) {
      Job j;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return stop || !jobs.empty(); });
        if (jobs.empty()) return;  // stop, and nothing left
        j = std::move(jobs.front());
        jobs.pop_front();
        busy = true;
      }
      const auto t0 = std::chrono::steady_clock::now();
      std::vector<uint64_t> ns;
      std::vector<const uint32_t*> idx;
      uint64_t recs = 0;
      for (const auto& l : j.lists) {
        ns.push_back(l.size());
        idx.push_back(l.data());
        recs += l.size();
      }
      int rc;
      {
        std::lock_guard<std::mutex> g(*gpu_mu);
        rc = ncd_arenas_sync_records((int)j.arenas.size(), j.arenas.data(), ns.data(), idx.data(), j.payload_bytes.data());
      }
      const std::string err = rc == NCD_OK ? std::string() : ncd_err("ncd_arenas_sync_records (early pull)");
      // objs does not move while the cache lock is held shared; in slices, so that a batch waiting to
      // update (exclusive) is not held up by a long list
      for (size_t q = 0; q < j.images.size(); ++q)
        for (size_t at = 0; at < j.lists[q].size(); at += 256) {
          std::shared_lock<std::shared_mutex> rd(*cache_mu);
          const size_t end = std::min(j.lists[q].size(), at + 256);
          for (size_t x = at; x < end; ++x) release_claim(j.images[q]->objs[j.lists[q][x]]);
        }
      std::lock_guard<std::mutex> lk(mu);
      if (!err.empty() && error.empty()) error = err;
      records += recs;
      ++calls;
      busy_ms += ms_since(t0);
      busy = false;
      if (jobs.empty()) idle.notify_all();
    }
  }
};

struct ncdh_checker {
  ncd_ctx* ctx = nullptr;
  uint32_t mode = 0;
  int threads = 1;
  std::shared_mutex cache_mu;
  std::mutex scan_mu;
  // Who may touch record bytes in the pinned payloads outside the exclusive cache lock: batches
  // that write values in place hold it shared from their first write until their update is applied;
  // a Scan holds it exclusively from laying records out until the GPU has pulled them.  Taken
  // before cache_mu.  Batches only ever TRY it (they have the slab path when a Scan is at work).
  std::shared_mutex pinned_mu;
  std::atomic<int> scan_wants_pinned{0};
  std::mutex gpu_mu;  // one library call at a time on ctx (a Scan's uploads and scan; the early pulls)
  Puller puller;
  std::atomic<uint64_t> early_h2d{0};  // bytes of the records handed to the puller since the last Scan
  std::atomic<uint64_t> clock{0};
  std::atomic<uint64_t> events_since_scan{0};
  std::atomic<uint64_t> canon_ns_since_scan{0};
  ClusterImage controller;
  std::vector<ClusterImage> shards;
  std::shared_ptr<const Snapshot> snap;  // read / written with std::atomic_load / atomic_store
  // scan thread only
  std::shared_ptr<const RowDirectory> dir;
  std::vector<ncd_adoption> unadopted;  // of the controller arena as last committed
  bool unadopted_valid = false;
  std::mutex timing_mu;
  ncdh_scan_timing timing{};

  ClusterImage* image(int cluster) {
    if (cluster == -1) return &controller;
    if (cluster >= 0 && (size_t)cluster < shards.size()) return &shards[cluster];
    return nullptr;
  }
  std::shared_ptr<const Snapshot> snapshot() const { return std::atomic_load(&snap); }
};

namespace {

struct Action {
  int shard;
  const char* client;  // "nexus" (templates, workgroups) or "kube" (secrets, configmaps)
  const char* verb;
  const char* resource;
  std::string name;
};

int64_t render(const std::vector<Action>& acts, const std::string& error, char* out, uint64_t cap) {
  std::string s;
  for (const Action& a : acts)
    s += std::to_string(a.shard) + "\t" + a.client + "\t" + a.verb + "\t" + a.resource + "\t" + a.name + "\n";
  if (!error.empty()) s += "error\t" + error + "\n";
  if (out && s.size() <= cap) memcpy(out, s.data(), s.size());
  return (int64_t)s.size();
}

// One event, canonicalised and ready to be put under the lock.
struct Prepared {
  int cluster = 0;
  uint32_t op = NCDH_EV_PUT;
  std::shared_ptr<const ObjInfo> info;
  std::string skey;
  const uint8_t* bytes = nullptr;  // canonical value, inside the batch's slab
  uint32_t len = 0;
  ClusterImage::PutHint hint;
  bool hinted = false;
  bool direct = false;  // the value is already in the record's place in the pinned payload (claim still held)
};

// Memory the canonical values of a batch of events are written into, one buffer per batch.  The
// cached objects keep it alive (shared_ptr) until the next Scan has laid their records out in the
// pinned arena; then it returns to the pool and the next batch reuses it — already mapped, so the
// canonicaliser does not pay a page fault per 4 KiB of output the way fresh allocations do.
struct Slab {
  uint8_t* p = nullptr;
  size_t cap = 0;
};
struct SlabPool {
  std::mutex mu;
  std::vector<Slab> free_list;
  static constexpr size_t kKeepBytes = 4ull << 30;  // buffers beyond this in total go back to the system
  // At least `bytes`, with headroom so that the next batch of about the same size fits too.  A fresh
  // buffer is NOT cleared (the canonicaliser's threads touch it first, in parallel) and asks for
  // huge pages.
  Slab take(size_t bytes) {
    {
      std::lock_guard<std::mutex> lk(mu);
      int best = -1;
      for (size_t i = 0; i < free_list.size(); ++i)
        if (free_list[i].cap >= bytes && (best < 0 || free_list[i].cap < free_list[best].cap)) best = (int)i;
      if (best >= 0) {
        Slab out = free_list[best];
        free_list.erase(free_list.begin() + best);
        return out;
      }
    }
    Slab out;
    out.cap = (bytes + bytes / 8 + (2u << 20)) & ~((size_t)(2u << 20) - 1);
    void* mem = nullptr;
    if (posix_memalign(&mem, 2u << 20, out.cap) != 0) throw std::bad_alloc();
#ifdef MADV_HUGEPAGE
    madvise(mem, out.cap, MADV_HUGEPAGE);
#endif
    out.p = (uint8_t*)mem;
    return out;
  }
  void give(Slab s) {
    {
      std::lock_guard<std::mutex> lk(mu);
      size_t kept = 0;
      for (const Slab& b : free_list) kept += b.cap;
      if (kept + s.cap <= kKeepBytes) {
        free_list.push_back(s);
        return;
      }
    }
    free(s.p);
  }
};
SlabPool g_pool;

// An event checked and keyed; the object's description (make_info) is only built when the cache
// does not already hold an equal one — per event it is a dozen small allocations, freed on another
// thread, which cost more than canonicalising a small value.
void prepare_meta(const ncdh_event& e, Prepared* p) {
  if (!e.ns || !e.name) throw std::runtime_error("null namespace / name");
  if (e.kind > NCD_KIND_CONFIGMAP) throw std::runtime_error("unknown kind");
  p->cluster = e.cluster;
  p->op = e.op;
  if (e.op == NCDH_EV_PUT) {
    if (!e.uid || (!e.value && e.value_len)) throw std::runtime_error("null uid / value");
    if (e.nowners < 0 || e.nsecrets < 0 || e.nconfigmaps < 0) throw std::runtime_error("negative count");
    if (e.value_len >= 0xFFFFFFFFull) throw std::runtime_error("value longer than 4 GiB");
  } else if (e.op != NCDH_EV_DELETE) {
    throw std::runtime_error("unknown event op");
  }
  const size_t ln = strlen(e.ns), lm = strlen(e.name);
  p->skey.reserve(ln + lm + 2);
  p->skey.push_back((char)('0' + e.kind));  // store_key(kind, "namespace/name")
  p->skey.append(e.ns, ln);
  p->skey.push_back('/');
  p->skey.append(e.name, lm);
}

std::shared_ptr<const ObjInfo> make_info(const ncdh_event& e) {
  auto info = std::make_shared<ObjInfo>();
  info->kind = e.kind;
  info->key = std::string(e.ns) + "/" + e.name;  // cache.ObjectName.String()
  info->name = e.name;
  if (e.op == NCDH_EV_PUT) {
    info->uid = e.uid;
    for (int i = 0; i < e.nowners; ++i) info->owners.push_back({e.owner_kinds[i], e.owner_uids[i]});
    for (int i = 0; i < e.nsecrets; ++i) info->secrets.push_back(e.secret_names[i]);
    for (int i = 0; i < e.nconfigmaps; ++i) info->configmaps.push_back(e.configmap_names[i]);
  }
  return info;
}

// ObjInfo::same_structure against an event, without building the event's ObjInfo.
bool same_structure(const ObjInfo& o, const ncdh_event& e) {
  if (o.uid != e.uid || o.owners.size() != (size_t)e.nowners || o.secrets.size() != (size_t)e.nsecrets ||
      o.configmaps.size() != (size_t)e.nconfigmaps)
    return false;
  for (int i = 0; i < e.nowners; ++i)
    if (o.owners[i].kind != e.owner_kinds[i] || o.owners[i].uid != e.owner_uids[i]) return false;
  for (int i = 0; i < e.nsecrets; ++i)
    if (o.secrets[i] != e.secret_names[i]) return false;
  for (int i = 0; i < e.nconfigmaps; ++i)
    if (o.configmaps[i] != e.configmap_names[i]) return false;
  return true;
}

// out / cap: where the canonical value goes (cap >= e.value_len always suffices).
void prepare_value(const ncdh_event& e, Prepared* p, uint8_t* out, uint64_t cap) {
  if (e.op != NCDH_EV_PUT) return;
  p->len = (uint32_t)canonicalise_into(e.value, e.value_len, out, cap);
  p->bytes = out;
}

void prepare_event(const ncdh_event& e, Prepared* p, uint8_t* out, uint64_t cap) {
  prepare_meta(e, p);
  p->info = make_info(e);
  prepare_value(e, p, out, cap);
}

// Applies prepared events in order.  Caller holds cache_mu exclusively.
int apply_locked(ncdh_checker* c, std::vector<Prepared>& evs, size_t count, const std::shared_ptr<const void>& hold,
                 int* removed_missing) {
  for (size_t i = 0; i < count; ++i) {
    Prepared& p = evs[i];
    ClusterImage* img = c->image(p.cluster);
    if (!img) return fail("no such cluster");
    const uint64_t stamp = c->clock.fetch_add(1) + 1;
    bool changed;
    if (p.op == NCDH_EV_PUT) {
      changed = img->put(p.skey, std::move(p.info), p.bytes, p.len, hold, stamp, p.hinted ? &p.hint : nullptr);
    } else {
      changed = img->drop(p.skey, stamp);
      if (!changed && removed_missing) ++*removed_missing;
    }
    if (changed) c->events_since_scan.fetch_add(1, std::memory_order_relaxed);
  }
  return 0;
}

}  // namespace

extern "C" {

const char* ncdh_last_error(void) { return g_err.c_str(); }

int64_t ncdh_canon(const uint8_t* value, uint64_t value_len, uint8_t* out, uint64_t cap) {
  try {
    const std::string c = ncdh::canonicalise(value, value_len);
    if (out && c.size() <= cap) memcpy(out, c.data(), c.size());
    return (int64_t)c.size();
  } catch (const std::exception& e) {
    return fail(e.what());
  }
}

int ncdh_checker_create(int device, uint32_t mode, int nshards, ncdh_checker** out) {
  if (!out || nshards < 0) return fail("ncdh_checker_create: bad argument");
  std::unique_ptr<ncdh_checker> c(new ncdh_checker);
  if (ncd_init(device, &c->ctx) != NCD_OK) return fail(ncd_err("ncd_init"));  // no GPU: an error, not a fallback
  c->mode = mode;
  c->puller.gpu_mu = &c->gpu_mu;
  c->puller.cache_mu = &c->cache_mu;
  c->shards.resize(nshards);
  const unsigned hc = std::thread::hardware_concurrency();
  c->threads = hc ? (int)(hc > 32 ? 32 : hc) : 4;
  *out = c.release();
  return 0;
}

void ncdh_checker_destroy(ncdh_checker* c) {
  if (!c) return;
  c->puller.shutdown();         // pulls what was still queued, then ends the thread
  c->controller.arena.reset();  // arenas before the context they live in
  for (ClusterImage& sh : c->shards) sh.arena.reset();
  if (c->ctx) ncd_destroy(c->ctx);
  delete c;
}

int ncdh_apply_events(ncdh_checker* c, const ncdh_event* events, uint64_t n, int threads) {
  if (!c || (!events && n)) return fail("ncdh_apply_events: null argument");
  if (n == 0) return 0;
  if (threads <= 0) threads = c->threads;
  const auto t0 = std::chrono::steady_clock::now();
  constexpr uint32_t kSkipped = 0xFFFFFFFFu;  // Prepared::op of an event that could not be prepared
  std::vector<Prepared> evs(n);
  std::vector<std::string> errs(n);
  // claims a failed call would otherwise leave behind
  struct ClaimGuard {
    ncdh_checker* c;
    std::vector<Prepared>& evs;
    ~ClaimGuard() {
      for (Prepared& p : evs)
        if (p.direct) release_claim(c->image(p.cluster)->objs[p.hint.pos]);
    }
  };
  try {
    // one slab for the canonical values that are not written in place (a canonical form is as long as its transport form)
    std::vector<uint64_t> off(n + 1, 0);
    for (uint64_t i = 0; i < n; ++i) off[i + 1] = off[i] + (events[i].op == NCDH_EV_PUT ? events[i].value_len : 0);
    const Slab slab = g_pool.take(off[n] ? off[n] : 1);
    const std::shared_ptr<const void> hold(slab.p, [slab](const void*) { g_pool.give(slab); });  // back to the pool when the
                                                                                                 // last cached object lets go
    // In place when no Scan is laying out or pulling records; held until the update below is done,
    // so no Scan can see a record whose bytes are ahead of its cached length.
    std::shared_lock<std::shared_mutex> pinned(c->pinned_mu, std::defer_lock);
    static const bool no_in_place = getenv("NCDH_NO_IN_PLACE") != nullptr;  // measurement knob: every value through the slab
    const bool in_place = !no_in_place && c->scan_wants_pinned.load(std::memory_order_acquire) == 0 && pinned.try_lock();
    ClaimGuard guard{c, evs};
    // A large batch is worked off in chunks: the records one chunk wrote in place are pulled to the
    // device (Puller) while the next chunk is canonicalised.
    static const bool no_early = getenv("NCDH_NO_EARLY_PULL") != nullptr;  // measurement knob
    const bool early = in_place && !no_early;
    const uint64_t chunk = early ? std::max<uint64_t>(2048, std::min<uint64_t>(65536, (n + 3) / 4)) : n;  // four chunks, within bounds
    double t_phase1 = 0, t_phase2 = 0;
    uint64_t sliced = 0;  // events applied in slices on several workers (a cluster's chunk written in place throughout)
    for (uint64_t c0 = 0; c0 < n; c0 += chunk) {
    const uint64_t c1 = std::min(n, c0 + chunk);
    const auto t_chunk = std::chrono::steady_clock::now();
    // Canonicalise in parallel, outside the exclusive lock, and look every object up and compare its
    // value with what the cache holds under the SHARED lock: readers and other batches go on, only the
    // short update below is exclusive.  A new value that fits its record's place goes straight into
    // the pinned payload (through a cache-resident scratch buffer, so that an unchanged value — a
    // resync — writes nothing); any other goes to the slab and waits for the packer.
    {
      std::shared_lock<std::shared_mutex> rd(c->cache_mu);
      // In blocks of eight events: first the keys and the lookups of all eight (each lookup ends in a few
      // dependent cache misses: bucket, node, object, its description, its pinned record), with the lines
      // they lead to requested, then the events themselves — the misses of eight events overlap instead
      // of queueing behind one another on the thread.
      constexpr size_t B = 8;
      constexpr size_t kNone = ~(size_t)0;
      parallel_for((c1 - c0 + B - 1) / B, threads, 1, [&](size_t bk) {
        const size_t b0 = c0 + bk * B, b1 = std::min<size_t>(c1, b0 + B);
        size_t pos[B];
        ClusterImage* imgs[B];
        for (size_t i = b0; i < b1; ++i) {
          const size_t k = i - b0;
          Prepared& p = evs[i];
          pos[k] = kNone;
          imgs[k] = nullptr;
          try {
            const ncdh_event& e = events[i];
            prepare_meta(e, &p);
            imgs[k] = c->image(p.cluster);
            if (!imgs[k]) throw std::runtime_error("no such cluster");
            if (p.op != NCDH_EV_PUT) continue;
            auto it = imgs[k]->index.find(p.skey);
            if (it != imgs[k]->index.end()) {
              pos[k] = it->second;
              __builtin_prefetch(&imgs[k]->objs[pos[k]]);
              __builtin_prefetch(reinterpret_cast<const char*>(&imgs[k]->objs[pos[k]]) + 64);
            }
            if (e.value) __builtin_prefetch(e.value);
          } catch (const std::exception& e) {
            errs[i] = e.what();
            p.op = kSkipped;
          }
        }
        for (size_t i = b0; i < b1; ++i) {
          const size_t k = i - b0;
          if (pos[k] == kNone) continue;
          const CachedObject& o = imgs[k]->objs[pos[k]];
          __builtin_prefetch(o.info.get());
          if (!o.pending && o.rec_off >= 0 && imgs[k]->host_payload) __builtin_prefetch(imgs[k]->host_payload + o.rec_off);
        }
        for (size_t i = b0; i < b1; ++i) {
          const size_t k = i - b0;
          Prepared& p = evs[i];
          if (p.op != NCDH_EV_PUT) continue;  // deletions and skipped events have nothing more to prepare
          ClusterImage* img = imgs[k];
          const ncdh_event& e = events[i];
          try {
            const CachedObject* o = pos[k] == kNone ? nullptr : &img->objs[pos[k]];
            if (!o || !try_claim(*o)) {  // new object, or another batch is at its bytes: compared under the exclusive lock
              p.info = make_info(e);
              prepare_value(e, &p, slab.p + off[i], off[i + 1] - off[i]);
              continue;
            }
            try {
              const bool same = same_structure(*o->info, e);
              if (!same) p.info = make_info(e);
              if (in_place && img->fits_in_place(*o, e.value_len)) {
                thread_local std::vector<uint8_t> scratch;
                if (scratch.size() < e.value_len) scratch.resize(e.value_len + e.value_len / 2 + 64);
                p.len = (uint32_t)canonicalise_into(e.value, e.value_len, scratch.data(), e.value_len);
                p.hint = img->precheck(pos[k], same, scratch.data(), p.len);
                p.hinted = true;
                if (!p.hint.same_value) {
                  img->write_in_place(*o, scratch.data(), p.len, p.hint.first_diff);
                  p.direct = true;  // the claim stays until the update is applied
                  continue;
                }
              } else {
                prepare_value(e, &p, slab.p + off[i], off[i + 1] - off[i]);
                p.hint = img->precheck(pos[k], same, p.bytes, p.len);
                p.hinted = true;
              }
            } catch (...) {
              release_claim(*o);
              throw;
            }
            release_claim(*o);
          } catch (const std::exception& ex) {
            errs[i] = ex.what();
            p.op = kSkipped;
          }
        }
      });
    }
    c->canon_ns_since_scan.fetch_add((uint64_t)(ms_since(t_chunk) * 1e6), std::memory_order_relaxed);
    t_phase1 += ms_since(t_chunk);
    const auto t_update = std::chrono::steady_clock::now();
    // The update itself: the events of different clusters touch different images, so each cluster's
    // events are applied, in order, by their own worker.
    std::vector<std::vector<uint32_t>> by_cluster(c->shards.size() + 1);
    for (uint64_t i = c0; i < c1; ++i)
      if (evs[i].op != kSkipped) by_cluster[(size_t)(evs[i].cluster + 1)].push_back((uint32_t)i);
    std::vector<std::vector<uint32_t>> pull_now(by_cluster.size());  // per cluster: records to hand to the device at once
    std::vector<uint64_t> pull_bytes(by_cluster.size(), 0);
    Puller::Job job;
    {
      std::unique_lock<std::shared_mutex> lk(c->cache_mu);
      // One event, the ordinary way (in order, on the cluster's one worker).
      auto apply_one = [&](size_t k, uint32_t i, uint64_t stamp) -> bool {
        ClusterImage* img = c->image((int)k - 1);
        Prepared& p = evs[i];
        bool changed;
        if (p.op == NCDH_EV_PUT && !p.info && !(p.hinted && img->hint_valid(p.hint) && p.hint.same_structure))
          p.info = make_info(events[i]);  // the object phase one compared against is not there any more
        if (p.op != NCDH_EV_PUT) {
          changed = img->drop(p.skey, stamp);
        } else if (p.direct || (!p.bytes && !img->hint_valid(p.hint))) {
          // written in place, or found equal in phase one without keeping the bytes
          const bool was_direct = p.direct;
          const size_t before = pull_now[k].size();
          if (was_direct && img->put_direct(p.info, p.len, stamp, p.hint, early ? &pull_now[k] : nullptr)) {
            changed = true;
            if (pull_now[k].size() != before) pull_bytes[k] += pad16(p.len) + 4;
          } else {
            // something else was applied to the object since phase one looked at it: the value is
            // canonicalised again and put the ordinary way (after an in-place write the pinned
            // bytes already are the new value, so it counts as changed without comparing)
            auto buf = std::make_shared<std::vector<uint8_t>>(events[i].value_len ? events[i].value_len : 1);
            const uint32_t len = (uint32_t)canonicalise_into(events[i].value, events[i].value_len, buf->data(), events[i].value_len);
            changed = img->put(p.skey, std::move(p.info), buf->data(), len, std::shared_ptr<const void>(buf, buf->data()), stamp,
                               nullptr, /*force=*/was_direct);
          }
          if (was_direct) {
            if (pull_now[k].size() == before) release_claim(img->objs[p.hint.pos]);  // else the claim goes with the pull
            p.direct = false;
          }
        } else {
          changed = img->put(p.skey, std::move(p.info), p.bytes, p.len, hold, stamp, p.hinted ? &p.hint : nullptr);
        }
        return changed;
      };
      // Work items: a cluster's events in order on one worker — or, when every one of them was written in place
      // (the steady state: distinct objects, each claimed by this batch, nothing to order), the cluster's list
      // in several slices on several workers, which touch the objects only and leave what belongs to the
      // image to be merged afterwards.
      struct Item {
        size_t k, from, to;
        bool fast;
        ClusterImage::DirectAcc acc;
        uint64_t bytes = 0, changed = 0;
        std::vector<uint32_t> retry;
      };
      std::vector<Item> items;
      for (size_t k = 0; k < by_cluster.size(); ++k) {
        const std::vector<uint32_t>& mine = by_cluster[k];
        if (mine.empty()) continue;
        bool all_direct = mine.size() >= 512;
        for (size_t at = 0; at < mine.size() && all_direct; ++at) all_direct = evs[mine[at]].op == NCDH_EV_PUT && evs[mine[at]].direct;
        const size_t parts = all_direct ? std::min<size_t>(4, mine.size() / 256) : 1;
        for (size_t q = 0; q < parts; ++q) {
          Item it;
          it.k = k;
          it.from = mine.size() * q / parts;
          it.to = mine.size() * (q + 1) / parts;
          it.fast = all_direct;
          items.push_back(std::move(it));
        }
      }
      Workers::get().run(items.size(), threads, [&](size_t q) {
        Item& it = items[q];
        ClusterImage* img = c->image((int)it.k - 1);
        const std::vector<uint32_t>& mine = by_cluster[it.k];
        // one block of stamps per item (the clock is shared by every worker): an object's history only
        // needs its own stamps to rise, and all of them to lie after earlier Scans
        uint64_t stamp = c->clock.fetch_add(it.to - it.from);
        for (size_t at = it.from; at < it.to; ++at) {
          const uint32_t i = mine[at];
          ++stamp;
          if (at + 4 < it.to) {  // the next objects' lines: each is a cache miss otherwise
            const Prepared& nx = evs[mine[at + 4]];
            if (nx.hinted && nx.hint.pos < img->objs.size()) {
              __builtin_prefetch(&img->objs[nx.hint.pos], 1);
              if (img->host_rec_len) __builtin_prefetch(img->host_rec_len + nx.hint.pos, 1);
            }
          }
          if (!it.fast) {
            it.changed += apply_one(it.k, i, stamp) ? 1 : 0;
            continue;
          }
          Prepared& p = evs[i];
          if (!p.info && !(img->hint_valid(p.hint) && p.hint.same_structure)) p.info = make_info(events[i]);
          const size_t before = it.acc.pull.size();
          if (img->put_direct_mt(p.info, p.len, stamp, p.hint, early, it.acc)) {
            ++it.changed;
            if (it.acc.pull.size() != before) it.bytes += pad16(p.len) + 4;  // the claim goes with the pull
            else release_claim(img->objs[p.hint.pos]);
            p.direct = false;
          } else {
            it.retry.push_back(i);  // something else was applied to the object meanwhile: the ordinary way, below
          }
        }
      });
      for (Item& it : items) {
        ClusterImage* img = c->image((int)it.k - 1);
        if (it.fast) {
          sliced += it.to - it.from;
          img->merge_direct(it.acc);
          pull_now[it.k].insert(pull_now[it.k].end(), it.acc.pull.begin(), it.acc.pull.end());
          pull_bytes[it.k] += it.bytes;
          for (uint32_t i : it.retry) it.changed += apply_one(it.k, i, c->clock.fetch_add(1) + 1) ? 1 : 0;
        }
        c->events_since_scan.fetch_add(it.changed, std::memory_order_relaxed);
      }
      for (size_t k = 0; k < pull_now.size(); ++k)  // arena and payload size as they are under the lock
        if (!pull_now[k].empty()) {
          ClusterImage* img = c->image((int)k - 1);
          job.images.push_back(img);
          job.arenas.push_back(img->arena.a);
          job.payload_bytes.push_back(img->payload_in_use());
          job.lists.push_back(std::move(pull_now[k]));
          c->early_h2d.fetch_add(pull_bytes[k], std::memory_order_relaxed);
        }
    }
    if (!job.arenas.empty()) c->puller.submit(std::move(job));
    t_phase2 += ms_since(t_update);
    }  // chunks
    const auto t_release = std::chrono::steady_clock::now();
    parallel_for(n, threads, 256, [&](size_t i) {  // what the events still own is released in parallel too
      evs[i].info.reset();
      std::string().swap(evs[i].skey);
    });
    static const bool profile = getenv("NCDH_PROFILE") != nullptr;
    if (profile)
      fprintf(stderr, "ncdh_apply_events: %llu events in chunks of %llu, lookup + canonicalise %.3f ms, update %.3f ms (%llu events in slices), release %.3f ms, total %.3f ms%s\n",
              (unsigned long long)n, (unsigned long long)chunk, t_phase1, t_phase2, (unsigned long long)sliced, ms_since(t_release), ms_since(t0),
              in_place ? "" : " (slab path: a Scan held the pinned payloads)");
  } catch (const std::exception& e) {
    return fail(std::string("ncdh_apply_events: ") + e.what());
  }
  for (uint64_t i = 0; i < n; ++i)
    if (evs[i].op == kSkipped) return fail("ncdh_apply_events: event " + std::to_string(i) + ": " + errs[i] + " (skipped; every other event of the batch was applied)");
  return 0;
}

int ncdh_cluster_add(ncdh_checker* c, int cluster, uint32_t kind, const char* ns, const char* name, const char* uid,
                     const uint8_t* value, uint64_t value_len, const char* const* owner_kinds,
                     const char* const* owner_uids, int nowners, const char* const* secret_names, int nsecrets,
                     const char* const* configmap_names, int nconfigmaps) {
  if (!c || !ns || !name || !uid || (!value && value_len)) return fail("ncdh_cluster_add: null argument");
  if (!c->image(cluster)) return fail("ncdh_cluster_add: no such cluster");
  if (kind > NCD_KIND_CONFIGMAP) return fail("ncdh_cluster_add: unknown kind");
  ncdh_event e;
  memset(&e, 0, sizeof(e));
  e.cluster = cluster;
  e.kind = kind;
  e.op = NCDH_EV_PUT;
  e.ns = ns;
  e.name = name;
  e.uid = uid;
  e.value = value;
  e.value_len = value_len;
  e.owner_kinds = owner_kinds;
  e.owner_uids = owner_uids;
  e.nowners = nowners;
  e.secret_names = secret_names;
  e.nsecrets = nsecrets;
  e.configmap_names = configmap_names;
  e.nconfigmaps = nconfigmaps;
  try {
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<Prepared> one(1);
    auto buf = std::make_shared<std::vector<uint8_t>>(value_len ? value_len : 1);
    prepare_event(e, &one[0], buf->data(), value_len);  // canonicalised on the caller's thread, outside the lock
    c->canon_ns_since_scan.fetch_add((uint64_t)(ms_since(t0) * 1e6), std::memory_order_relaxed);
    std::unique_lock<std::shared_mutex> lk(c->cache_mu);
    return apply_locked(c, one, 1, std::shared_ptr<const void>(buf, buf->data()), nullptr);
  } catch (const std::exception& ex) {
    return fail(std::string("ncdh_cluster_add: ") + ex.what());
  }
}

int ncdh_cluster_remove(ncdh_checker* c, int cluster, uint32_t kind, const char* ns, const char* name) {
  if (!c || !c->image(cluster) || !ns || !name || kind > NCD_KIND_CONFIGMAP) return fail("ncdh_cluster_remove: bad argument");
  std::vector<Prepared> one(1);
  one[0].cluster = cluster;
  one[0].op = NCDH_EV_DELETE;
  one[0].skey = store_key(kind, std::string(ns) + "/" + name);
  int missing = 0;
  std::unique_lock<std::shared_mutex> lk(c->cache_mu);
  if (apply_locked(c, one, 1, nullptr, &missing) != 0) return -1;
  return missing ? 1 : 0;
}

int ncdh_scan(ncdh_checker* c) {
  if (!c) return fail("ncdh_scan: null checker");
  std::lock_guard<std::mutex> scan_lock(c->scan_mu);
  const auto t_all = std::chrono::steady_clock::now();
  ncdh_scan_timing tm;
  memset(&tm, 0, sizeof(tm));
  // record bytes in pinned memory are this Scan's until the GPU has pulled them: batches in flight
  // finish first, later ones use the slab path meanwhile
  c->scan_wants_pinned.fetch_add(1, std::memory_order_release);
  std::unique_lock<std::shared_mutex> pinned(c->pinned_mu);
  c->scan_wants_pinned.fetch_sub(1, std::memory_order_release);
  try {
    // records that batches handed to the device as they were applied: all of them are there now
    const auto t_drain = std::chrono::steady_clock::now();
    const std::string early_error = c->puller.drain();
    if (!early_error.empty()) throw std::runtime_error(early_error);
    const double drain_ms = ms_since(t_drain);
    const size_t S = c->shards.size();
    std::vector<ClusterImage*> imgs;
    imgs.push_back(&c->controller);
    for (ClusterImage& sh : c->shards) imgs.push_back(&sh);
    std::vector<Upload> ups(imgs.size(), Upload::None);
    auto snap = std::make_shared<Snapshot>();
    const std::shared_ptr<const Snapshot> prev = c->snapshot();
    {
      // ---- under the cache lock: stamp the scan, lay the changes out in pinned memory
      std::unique_lock<std::shared_mutex> lk(c->cache_mu);
      const auto t0 = std::chrono::steady_clock::now();
      snap->clock = c->clock.load();
      tm.events = c->events_since_scan.exchange(0);
      tm.canon_ms = (double)c->canon_ns_since_scan.exchange(0) / 1e6;
      // the common case — values rewritten inside resident arenas — for all clusters at once, one worker
      // each; whatever needs a new arena, a compaction or new per-object arrays runs alone afterwards
      const uint64_t floor = prev ? prev->clock : 0;
      std::vector<uint64_t> h2d(imgs.size(), 0);
      std::vector<char> later(imgs.size(), 0);
      Workers::get().run(imgs.size(), c->threads, [&](size_t k) {
        bool deferred = false;
        ups[k] = imgs[k]->prepare(c->ctx, 1, floor, &h2d[k], &deferred);
        later[k] = deferred;
      });
      for (size_t k = 0; k < imgs.size(); ++k) {
        if (later[k]) ups[k] = imgs[k]->prepare(c->ctx, c->threads, floor, &h2d[k]);
        tm.h2d_bytes += h2d[k];
      }
      if (ups[0] == Upload::Full || ups[0] == Upload::Structural || !c->dir) {
        // row numbering: arena order, one row per template + one per dependent, one per workgroup
        auto dir = std::make_shared<RowDirectory>();
        const std::vector<CachedObject>& objs = c->controller.objs;
        dir->objs.reserve(objs.size());
        dir->row0.assign(objs.size(), 0);
        uint64_t row = 0;
        for (size_t i = 0; i < objs.size(); ++i) {
          const ObjInfo& f = *objs[i].info;
          dir->objs.push_back(objs[i].info);
          if (f.kind == NCD_KIND_TEMPLATE || f.kind == NCD_KIND_WORKGROUP) {
            dir->row0[i] = row;
            dir->by_key.emplace(store_key(f.kind, f.key), (uint32_t)i);
            row += 1 + (f.kind == NCD_KIND_TEMPLATE ? f.secrets.size() + f.configmaps.size() : 0);
          }
        }
        dir->rows = row;
        c->dir = std::move(dir);
        c->unadopted_valid = false;
      }
      for (ClusterImage* img : imgs) {
        tm.compactions += img->compactions;
        tm.records_synced += img->pulled_early;
        tm.records_in_place += img->pulled_early;
        img->pulled_early = 0;
      }
      tm.h2d_bytes += c->early_h2d.exchange(0);
      tm.pack_ms = ms_since(t0);
    }
    // ---- GPU: uploads, then the scan (the cache lock is free: events and lookups go on)
    std::lock_guard<std::mutex> gpu(c->gpu_mu);
    const auto t_up = std::chrono::steady_clock::now();
    std::vector<ncd_arena*> as, full;
    std::vector<uint64_t> ns, pbs, fc[5];
    std::vector<const uint32_t*> idx;
    for (size_t k = 0; k < imgs.size(); ++k) {
      ClusterImage* img = imgs[k];
      if (ups[k] == Upload::Full) ++tm.full_uploads;
      if (ups[k] == Upload::Structural) ++tm.structural_uploads;
      if (ups[k] == Upload::Full) {  // whole clusters (first scan, a larger arena): their uploads overlap each other's index build
        uint64_t c5[5];
        img->full_counts(c5);
        full.push_back(img->arena.a);
        for (int q = 0; q < 5; ++q) fc[q].push_back(c5[q]);
      } else if (ups[k] == Upload::Values) {  // the value rewrites of every cluster: one round trip
        as.push_back(img->arena.a);
        ns.push_back(img->pending_changed.size());
        idx.push_back(img->pending_changed.data());
        pbs.push_back(img->payload_used);
        tm.records_synced += img->pending_changed.size();
        tm.records_in_place += img->in_place_last;
      } else {
        img->upload(ups[k]);
      }
    }
    if (!full.empty()) {
      if (ncd_arenas_commit((int)full.size(), full.data(), fc[0].data(), fc[1].data(), fc[2].data(), fc[3].data(), fc[4].data()) != NCD_OK)
        throw std::runtime_error(ncd_err("ncd_arenas_commit"));
      for (size_t k = 0; k < imgs.size(); ++k)
        if (ups[k] == Upload::Full) imgs[k]->full_done();
    }
    if (!as.empty()) {
      const int rc = ncd_arenas_sync_records((int)as.size(), as.data(), ns.data(), idx.data(), pbs.data());
      for (ClusterImage* img : imgs) img->pending_changed.clear();
      if (rc != NCD_OK) throw std::runtime_error(ncd_err("ncd_arenas_sync_records"));
      tm.d2h_bytes += 8 * as.size();
    }
    if (!c->unadopted_valid) {  // ownerReferences only change with a structural upload of the controller
      uint64_t n = 0;
      c->unadopted.resize(1024);
      if (ncd_find_unadopted(c->ctx, c->controller.arena.a, c->unadopted.data(), c->unadopted.size(), &n) != NCD_OK)
        throw std::runtime_error(ncd_err("ncd_find_unadopted"));
      if (n > c->unadopted.size()) {
        c->unadopted.resize(n);
        if (ncd_find_unadopted(c->ctx, c->controller.arena.a, c->unadopted.data(), c->unadopted.size(), &n) != NCD_OK)
          throw std::runtime_error(ncd_err("ncd_find_unadopted"));
      }
      c->unadopted.resize(n);
      c->unadopted_valid = true;
      tm.d2h_bytes += 8 + n * sizeof(ncd_adoption);
    }
    tm.upload_ms = ms_since(t_up) + drain_ms;  // what this Scan still had to wait for
    pinned.unlock();
    static const bool profile = getenv("NCDH_PROFILE") != nullptr;
    if (profile) {
      std::string line = "ncdh_scan: uploads";
      for (size_t k = 0; k < imgs.size(); ++k) {
        const char* names[] = {"none", "values", "structural", "full"};
        line += std::string(" ") + names[(int)ups[k]];
        if (ups[k] == Upload::Values) line += "(" + std::to_string(imgs[k]->in_place_last) + " in place)";
      }
      fprintf(stderr, "%s; pack %.3f ms, upload %.3f ms\n", line.c_str(), tm.pack_ms, tm.upload_ms);
    }
    const auto t_scan = std::chrono::steady_clock::now();
    std::vector<const ncd_arena*> ptrs;
    for (ClusterImage& sh : c->shards) ptrs.push_back(sh.arena.a);
    const ncd_entry* ents = nullptr;
    uint64_t nent = 0;
    if (ncd_scan_host(c->ctx, c->controller.arena.a, ptrs.data(), (int)S, c->mode, &ents, &nent) != NCD_OK)
      throw std::runtime_error(ncd_err("ncd_scan_host"));
    ncd_scan_stats st;
    if (ncd_scan_last_stats(c->ctx, &st) != NCD_OK) throw std::runtime_error(ncd_err("ncd_scan_last_stats"));
    tm.scan_ms = ms_since(t_scan);
    tm.gpu_scan_ms = st.total_ms;
    tm.d2h_bytes += 16 + nent * sizeof(ncd_entry);
    // ---- publish
    const auto t_pub = std::chrono::steady_clock::now();
    if (st.rows != c->dir->rows) throw std::runtime_error("row count disagrees with the library");
    snap->nshards = S;
    snap->rows = st.rows;
    snap->dir = c->dir;
    snap->entries.assign(ents, ents + nent);
    snap->unadopted = c->unadopted;
    for (const ncd_entry& e : snap->entries)
      if (e.obj >= snap->dir->objs.size()) throw std::runtime_error("diff entry names an object the controller does not hold");
    std::atomic_store(&c->snap, std::shared_ptr<const Snapshot>(std::move(snap)));
    tm.publish_ms = ms_since(t_pub);
  } catch (const std::exception& e) {
    // whatever was laid out for this Scan may not have reached the GPU: the next Scan uploads
    // every cluster in full again rather than trust a half-updated resident copy
    const std::string what = e.what();
    std::unique_lock<std::shared_mutex> lk(c->cache_mu);
    c->controller.invalidate();
    for (ClusterImage& sh : c->shards) sh.invalidate();
    c->dir.reset();
    return fail("ncdh_scan: " + what);
  }
  tm.total_ms = ms_since(t_all);
  std::lock_guard<std::mutex> tl(c->timing_mu);
  c->timing = tm;
  return 0;
}

}  // extern "C"

namespace {

// Freshness of everything a decision about template / workgroup `f` on `shards` [s0, s1) rests on:
// the controller's copy of the object and of each dependent it names, and every listed shard's
// copies of the same keys (the shard template's UID decides ownership, controller.go:529).
// True when any of them was put or removed after the snapshot's Scan began.  Caller holds
// cache_mu (shared).
bool stale_for(ncdh_checker* c, const Snapshot& s, const std::string& skey, const ObjInfo* f, size_t s0, size_t s1) {
  bool any = c->controller.last_change > s.clock;
  for (size_t k = s0; k < s1 && !any; ++k) any = c->shards[k].last_change > s.clock;
  if (!any) return false;  // nothing at all changed in the clusters involved: the common case
  std::vector<std::string> keys;
  keys.push_back(skey);
  if (f && f->kind == NCD_KIND_TEMPLATE) {
    const std::string ns = f->key.substr(0, f->key.find('/') + 1);
    for (const std::string& n : f->secrets) keys.push_back(store_key(NCD_KIND_SECRET, ns + n));
    for (const std::string& n : f->configmaps) keys.push_back(store_key(NCD_KIND_CONFIGMAP, ns + n));
  }
  for (const std::string& k : keys) {
    if (c->controller.changed_at(k) > s.clock) return true;
    for (size_t sh = s0; sh < s1; ++sh)
      if (c->shards[sh].changed_at(k) > s.clock) return true;
  }
  return false;
}

}  // namespace

extern "C" {

int ncdh_lookup(ncdh_checker* c, uint32_t kind, const char* ns, const char* name, int dep, int shard, uint32_t* code) {
  if (!c || !ns || !name || !code) return fail("ncdh_lookup: null argument");
  const std::shared_ptr<const Snapshot> s = c->snapshot();
  if (!s) return fail("ncdh_lookup: no scan yet");
  if (shard < 0 || (size_t)shard >= s->nshards) return fail("ncdh_lookup: no such shard");
  const std::string k = store_key(kind, std::string(ns) + "/" + name);
  auto it = s->dir->by_key.find(k);
  std::shared_lock<std::shared_mutex> lk(c->cache_mu);
  if (it == s->dir->by_key.end())  // not in the snapshot: gone for good, or added since
    return c->controller.changed_at(k) > s->clock ? NCDH_STALE : NCDH_NOT_FOUND;
  const ObjInfo& f = *s->dir->objs[it->second];
  const size_t ndeps = f.kind == NCD_KIND_TEMPLATE ? f.secrets.size() + f.configmaps.size() : 0;
  if (dep < 0 || (size_t)dep > ndeps) return fail("ncdh_lookup: no such dependent");
  if (stale_for(c, *s, k, &f, (size_t)shard, (size_t)shard + 1)) return NCDH_STALE;
  *code = s->code(it->second, (uint32_t)dep, (uint32_t)shard);
  return 0;
}

int64_t ncdh_plan_template(ncdh_checker* c, const char* ns, const char* name, char* out, uint64_t cap) {
  if (!c || !ns || !name) return fail("ncdh_plan_template: null argument");
  const std::shared_ptr<const Snapshot> s = c->snapshot();
  if (!s) return fail("ncdh_plan_template: no scan yet");
  std::vector<Action> acts;
  std::string error;
  const std::string k = store_key(NCD_KIND_TEMPLATE, std::string(ns) + "/" + name);
  auto it = s->dir->by_key.find(k);
  const size_t S = s->nshards;
  {
    std::shared_lock<std::shared_mutex> lk(c->cache_mu);
    if (it == s->dir->by_key.end()) {
      if (c->controller.changed_at(k) > s->clock) return -NCDH_STALE;
      return render(acts, error, out, cap);  // :769-772: gone from the cache, nil
    }
    if (stale_for(c, *s, k, s->dir->objs[it->second].get(), 0, S)) return -NCDH_STALE;
  }
  const uint32_t obj = it->second;
  const ObjInfo& t = *s->dir->objs[obj];
  const size_t nsec = t.secrets.size(), ndeps = nsec + t.configmaps.size();
  auto dep_name = [&](size_t j) { return j < nsec ? t.secrets[j] : t.configmaps[j - nsec]; };
  auto dep_res = [&](size_t j) { return j < nsec ? "secrets" : "configmaps"; };
  // adoptReferences (:783 -> :647-676) fails the handler before any shard when a referenced
  // object is missing on the controller; ERR_MISSING_REF does not depend on the shard
  if (S > 0) {
    for (size_t j = 0; j < ndeps; ++j)
      if (s->code(obj, (uint32_t)j + 1, 0) == NCD_ERR_MISSING_REF) {
        error = std::string("missing ") + dep_res(j) + " " + dep_name(j);
        return render(acts, error, out, cap);
      }
  } else {
    std::shared_lock<std::shared_mutex> lk(c->cache_mu);
    for (size_t j = 0; j < ndeps; ++j)
      if (!c->controller.get(j < nsec ? NCD_KIND_SECRET : NCD_KIND_CONFIGMAP, std::string(ns) + "/" + dep_name(j))) {
        error = std::string("missing ") + dep_res(j) + " " + dep_name(j);
        return render(acts, error, out, cap);
      }
  }
  for (size_t sh = 0; sh < S && error.empty(); ++sh) {  // :790
    const uint32_t tc = s->code(obj, 0, (uint32_t)sh);
    if (tc & NCD_UPDATE_DATA) acts.push_back({(int)sh, "nexus", "update", "NexusAlgorithmTemplates", t.name});  // :795-797
    if (tc == NCD_CREATE) acts.push_back({(int)sh, "nexus", "create", "NexusAlgorithmTemplates", t.name});      // :805-808
    for (size_t j = 0; j < ndeps; ++j) {  // :817 then :825
      const uint32_t code = s->code(obj, (uint32_t)j + 1, (uint32_t)sh);
      if (code == NCD_ERR_ROGUE) {  // :529-535 / :591-597
        error = "Resource \"" + dep_name(j) + "\" already exists and is not managed by any Algorithm";
        break;
      }
      if (code == NCD_CREATE) {  // :518-520 / :580-582
        acts.push_back({(int)sh, "kube", "create", dep_res(j), dep_name(j)});
        continue;
      }
      if (code & NCD_UPDATE_DATA) acts.push_back({(int)sh, "kube", "update-data", dep_res(j), dep_name(j)});    // :539-541 / :600-602
      if (code & NCD_UPDATE_OWNER) acts.push_back({(int)sh, "kube", "update-owner", dep_res(j), dep_name(j)});  // :550-552 / :612-614
    }
  }
  return render(acts, error, out, cap);
}

int64_t ncdh_plan_workgroup(ncdh_checker* c, const char* ns, const char* name, char* out, uint64_t cap) {
  if (!c || !ns || !name) return fail("ncdh_plan_workgroup: null argument");
  const std::shared_ptr<const Snapshot> s = c->snapshot();
  if (!s) return fail("ncdh_plan_workgroup: no scan yet");
  std::vector<Action> acts;
  const std::string k = store_key(NCD_KIND_WORKGROUP, std::string(ns) + "/" + name);
  auto it = s->dir->by_key.find(k);
  const size_t S = s->nshards;
  {
    std::shared_lock<std::shared_mutex> lk(c->cache_mu);
    if (it == s->dir->by_key.end()) {
      if (c->controller.changed_at(k) > s->clock) return -NCDH_STALE;
      return render(acts, "", out, cap);  // :705-708
    }
    if (stale_for(c, *s, k, s->dir->objs[it->second].get(), 0, S)) return -NCDH_STALE;
  }
  const std::string& wname = s->dir->objs[it->second]->name;
  for (size_t sh = 0; sh < S; ++sh) {  // :720
    const uint32_t code = s->code(it->second, 0, (uint32_t)sh);
    if (code & NCD_UPDATE_DATA) acts.push_back({(int)sh, "nexus", "update", "NexusAlgorithmWorkgroups", wname});  // :725-727
    if (code == NCD_CREATE) acts.push_back({(int)sh, "nexus", "create", "NexusAlgorithmWorkgroups", wname});      // :735-738
  }
  return render(acts, "", out, cap);
}

int64_t ncdh_plan_adoption(ncdh_checker* c, const char* ns, const char* name, char* out, uint64_t cap) {
  if (!c || !ns || !name) return fail("ncdh_plan_adoption: null argument");
  const std::shared_ptr<const Snapshot> s = c->snapshot();
  if (!s) return fail("ncdh_plan_adoption: no scan yet");
  const std::string k = store_key(NCD_KIND_TEMPLATE, std::string(ns) + "/" + name);
  auto it = s->dir->by_key.find(k);
  std::string text;
  {
    std::shared_lock<std::shared_mutex> lk(c->cache_mu);
    if (it == s->dir->by_key.end()) return c->controller.changed_at(k) > s->clock ? -NCDH_STALE : 0;
    if (stale_for(c, *s, k, s->dir->objs[it->second].get(), 0, 0)) return -NCDH_STALE;
  }
  const ObjInfo& t = *s->dir->objs[it->second];
  const size_t nsec = t.secrets.size(), ndeps = nsec + t.configmaps.size();
  for (size_t j = 0; j < ndeps; ++j) {  // :648 secrets, then :671 configmaps
    // a reference the controller does not hold ends adoptReferences with an error (:649-653, :672-676)
    if (s->nshards > 0 && s->code(it->second, (uint32_t)j + 1, 0) == NCD_ERR_MISSING_REF) break;
    if (s->needs_adoption(it->second, (uint32_t)j + 1))
      text += std::string("kube\tupdate-owner\t") + (j < nsec ? "secrets" : "configmaps") + "\t" +
              (j < nsec ? t.secrets[j] : t.configmaps[j - nsec]) + "\n";
  }
  if (out && text.size() <= cap) memcpy(out, text.data(), text.size());
  return (int64_t)text.size();
}

int64_t ncdh_sweep_queue(ncdh_checker* c, char* out, uint64_t cap) {
  if (!c) return fail("ncdh_sweep_queue: null checker");
  const std::shared_ptr<const Snapshot> s = c->snapshot();
  if (!s) return fail("ncdh_sweep_queue: no scan yet");
  // objects with a shard decision != SKIP, merged with templates that still have references to
  // adopt on the controller cluster; both lists are ordered by controller index
  std::vector<uint32_t> objs;  // the merged, de-duplicated object numbers (a sequential walk)
  objs.reserve(s->entries.size() / (s->nshards ? s->nshards : 1) + s->unadopted.size() + 1);
  size_t i = 0, j = 0;
  uint32_t last = 0xFFFFFFFFu;
  while (i < s->entries.size() || j < s->unadopted.size()) {
    uint32_t obj;
    if (j >= s->unadopted.size() || (i < s->entries.size() && s->entries[i].obj <= s->unadopted[j].obj)) obj = s->entries[i++].obj;
    else obj = s->unadopted[j++].obj;
    if (obj == last) continue;
    last = obj;
    objs.push_back(obj);
  }
  // their names: every one is two or three cache misses away (directory -> object -> key), so long lists are
  // rendered in slices on the worker pool
  auto render = [&](size_t from, size_t to, std::string& text) {
    text.reserve(64 * (to - from));
    for (size_t k = from; k < to; ++k) {
      const ObjInfo& f = *s->dir->objs[objs[k]];
      text.append(f.kind == NCD_KIND_TEMPLATE ? "template\t" : "workgroup\t");
      text.append(f.key);
      text.push_back('\n');
    }
  };
  std::string text;
  if (objs.size() < 2048 || c->threads < 2) {
    render(0, objs.size(), text);
  } else {
    const size_t slices = std::min<size_t>((size_t)c->threads, objs.size() / 512);
    std::vector<std::string> parts(slices);
    Workers::get().run(slices, c->threads, [&](size_t k) { render(objs.size() * k / slices, objs.size() * (k + 1) / slices, parts[k]); });
    size_t total = 0;
    for (const std::string& p : parts) total += p.size();
    text.reserve(total);
    for (const std::string& p : parts) text.append(p);
  }
  if (out && text.size() <= cap) memcpy(out, text.data(), text.size());
  return (int64_t)text.size();
}

int ncdh_upload_counts(ncdh_checker* c, int cluster, uint64_t* full, uint64_t* structural, uint64_t* values) {
  ClusterImage* img = c ? c->image(cluster) : nullptr;
  if (!img) return fail("ncdh_upload_counts: no such cluster");
  std::shared_lock<std::shared_mutex> lk(c->cache_mu);
  if (full) *full = img->uploads_full;
  if (structural) *structural = img->uploads_structural;
  if (values) *values = img->uploads_values;
  return 0;
}

int ncdh_last_counts(ncdh_checker* c, uint64_t* ndiff, uint64_t* decisions) {
  if (!c) return fail("ncdh_last_counts: null checker");
  const std::shared_ptr<const Snapshot> s = c->snapshot();
  if (!s) return fail("ncdh_last_counts: no scan yet");
  if (ndiff) *ndiff = s->entries.size();
  if (decisions) *decisions = s->rows * s->nshards;
  return 0;
}

int ncdh_diff_digest(ncdh_checker* c, uint64_t* digest, uint64_t* dirty_objects) {
  if (!c || !digest) return fail("ncdh_diff_digest: null argument");
  const std::shared_ptr<const Snapshot> s = c->snapshot();
  if (!s) return fail("ncdh_diff_digest: no scan yet");
  auto fmix = [](uint64_t x) {
    x ^= x >> 33;
    x *= 0xff51afd7ed558ccdull;
    x ^= x >> 33;
    x *= 0xc4ceb9fe1a85ec53ull;
    return x ^ (x >> 33);
  };
  uint64_t sum = 0, dirty = 0, kh = 0;
  uint32_t last = 0xFFFFFFFFu;
  for (const ncd_entry& e : s->entries) {
    if (e.obj != last) {  // the key hash is per object: entries of one object are adjacent
      const ObjInfo& f = *s->dir->objs[e.obj];
      kh = 0xcbf29ce484222325ull;
      for (unsigned char ch : store_key(f.kind, f.key)) kh = (kh ^ ch) * 0x100000001b3ull;
      last = e.obj;
      ++dirty;
    }
    sum += fmix(kh ^ ((uint64_t)e.dep * 0x9E3779B185EBCA87ull) ^ ((uint64_t)e.shard * 0xC2B2AE3D27D4EB4Full) ^
                ((uint64_t)e.code * 0x165667B19E3779F9ull));
  }
  *digest = sum;
  if (dirty_objects) *dirty_objects = dirty;
  return 0;
}

int ncdh_last_timing(ncdh_checker* c, ncdh_scan_timing* out) {
  if (!c || !out) return fail("ncdh_last_timing: null argument");
  std::lock_guard<std::mutex> tl(c->timing_mu);
  *out = c->timing;
  return 0;
}

}  // extern "C"
