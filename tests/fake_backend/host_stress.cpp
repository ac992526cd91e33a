// host_stress.cpp — the host layer (csrc/host) hammered from several threads, the way the reference
// uses its caches: informer callbacks write (controller.go:286-359) while `workers` handler
// goroutines read (controller.go:875-877).  Linked against the CPU test double of libncdrift
// (fake_ncdrift.cpp) and built with -fsanitize=thread by tests/fake_backend/build.py; run by
// tests/test_host_threads.py.  Checks, beside the absence of data races:
//   - a Secret rotated after a scan is answered NCDH_STALE, never SKIP, until the next scan
//   - after that scan the same lookup answers UPDATE_DATA
//   - readers only ever see 0 / NOT_FOUND / STALE from lookups and >= 0 / -STALE from plans
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include "../../include/ncdhost.h"
#include "../../include/ncdrift.h"

static std::string wire_str(const std::string& s) {
  std::string w(1, (char)3);
  uint32_t n = (uint32_t)s.size();
  w.append((const char*)&n, 4);
  return w + s;
}
static std::string wire_bytes_map(const std::string& key, const std::string& val) {
  std::string w(1, (char)8);
  uint32_t one = 1, n = (uint32_t)val.size();
  w.append((const char*)&one, 4);
  w += wire_str(key);
  w.push_back((char)4);
  w.append((const char*)&n, 4);
  return w + val;
}

static std::atomic<int> failures{0};
#define CHECK(cond, ...)                    \
  do {                                      \
    if (!(cond)) {                          \
      fprintf(stderr, "FAIL %s:%d: ", __FILE__, __LINE__); \
      fprintf(stderr, __VA_ARGS__);         \
      fprintf(stderr, "\n");                \
      failures.fetch_add(1);                \
    }                                       \
  } while (0)

static void add(ncdh_checker* c, int cluster, uint32_t kind, const std::string& name, const std::string& uid,
                const std::string& value, const char* owner_uid, const char* secret, const char* configmap) {
  const char* ok[] = {"NexusAlgorithmTemplate"};
  const char* ou[] = {owner_uid};
  const char* sn[] = {secret};
  const char* cn[] = {configmap};
  const int rc = ncdh_cluster_add(c, cluster, kind, "default", name.c_str(), uid.c_str(), (const uint8_t*)value.data(), value.size(), ok,
                                  ou, owner_uid ? 1 : 0, sn, secret ? 1 : 0, cn, configmap ? 1 : 0);
  CHECK(rc == 0, "ncdh_cluster_add: %s", ncdh_last_error());
}

int main(int argc, char** argv) {
  const int rounds = argc > 1 ? atoi(argv[1]) : 60;
  const int T = 24, S = 2;
  ncdh_checker* c = nullptr;
  if (ncdh_checker_create(0, NCD_MODE_DIRECT, S, &c) != 0) {
    fprintf(stderr, "create: %s\n", ncdh_last_error());
    return 2;
  }
  for (int cluster = -1; cluster < S; ++cluster)
    for (int t = 0; t < T; ++t) {
      const std::string n = std::to_string(t), tuid = (cluster < 0 ? "c-" : "s" + std::to_string(cluster) + "-") + n;
      add(c, cluster, NCD_KIND_TEMPLATE, "algo-" + n, tuid, wire_str("spec-" + n), nullptr, ("sec-" + n).c_str(), ("cm-" + n).c_str());
      add(c, cluster, NCD_KIND_SECRET, "sec-" + n, "", wire_bytes_map("token", "v0-" + n), tuid.c_str(), nullptr, nullptr);
      add(c, cluster, NCD_KIND_CONFIGMAP, "cm-" + n, "", wire_bytes_map("file", "c0-" + n), tuid.c_str(), nullptr, nullptr);
    }
  CHECK(ncdh_scan(c) == 0, "first scan: %s", ncdh_last_error());
  uint64_t nd = 0, dec = 0;
  ncdh_last_counts(c, &nd, &dec);
  CHECK(nd == 0 && dec == (uint64_t)T * 3 * S, "converged fleet: %llu entries, %llu decisions", (unsigned long long)nd, (unsigned long long)dec);

  std::atomic<bool> stop{false};
  std::atomic<uint64_t> lookups{0}, stale_seen{0};
  // readers: every template, every dependent, every shard, over and over
  std::vector<std::thread> readers;
  for (int r = 0; r < 4; ++r)
    readers.emplace_back([&, r] {
      char buf[4096];
      for (uint64_t i = r; !stop.load(); ++i) {
        const std::string n = "algo-" + std::to_string(i % (T + 2));  // two names that may not exist
        uint32_t code = 99;
        const int rc = ncdh_lookup(c, NCD_KIND_TEMPLATE, "default", n.c_str(), (int)(i % 3), (int)(i % S), &code);
        CHECK(rc == 0 || rc == NCDH_NOT_FOUND || rc == NCDH_STALE, "lookup rc %d: %s", rc, ncdh_last_error());
        if (rc == 0) CHECK(code == NCD_SKIP || code == NCD_UPDATE_DATA || code == NCD_CREATE || code == NCD_UPDATE_OWNER ||
                               code == (NCD_UPDATE_DATA | NCD_UPDATE_OWNER),
                           "lookup code %u", code);
        if (rc == NCDH_STALE) stale_seen.fetch_add(1);
        const int64_t pl = ncdh_plan_template(c, "default", n.c_str(), buf, sizeof(buf));
        CHECK(pl >= 0 || pl == -NCDH_STALE, "plan rc %lld: %s", (long long)pl, ncdh_last_error());
        const int64_t ad = ncdh_plan_adoption(c, "default", n.c_str(), buf, sizeof(buf));
        CHECK(ad >= 0 || ad == -NCDH_STALE, "adoption rc %lld", (long long)ad);
        CHECK(ncdh_sweep_queue(c, buf, sizeof(buf)) >= 0, "sweep queue: %s", ncdh_last_error());
        lookups.fetch_add(1);
      }
    });
  // a second writer on its own templates (the upper half), with batches and structural churn, and its own scans
  std::thread writer2([&] {
    for (int k = 0; !stop.load(); ++k) {
      const int t = T / 2 + k % (T / 2);
      const std::string n = std::to_string(t), v = wire_bytes_map("file", "c" + std::to_string(k) + "-" + n);
      const std::string name = "cm-" + n, tuid = "c-" + n;
      const char* ok[] = {"NexusAlgorithmTemplate"};
      const char* ou[] = {tuid.c_str()};
      ncdh_event ev[2];
      memset(ev, 0, sizeof(ev));
      for (int q = 0; q < 2; ++q) {
        ev[q].cluster = q == 0 ? -1 : k % S;
        ev[q].kind = NCD_KIND_CONFIGMAP;
        ev[q].op = NCDH_EV_PUT;
        ev[q].ns = "default";
        ev[q].name = name.c_str();
        ev[q].uid = "";
        ev[q].value = (const uint8_t*)v.data();
        ev[q].value_len = v.size();
        ev[q].owner_kinds = ok;
        ev[q].owner_uids = ou;
        ev[q].nowners = 1;
      }
      CHECK(ncdh_apply_events(c, ev, 2, 2) == 0, "apply_events: %s", ncdh_last_error());
      if (k % 5 == 0) {  // structural: an extra workgroup comes and goes on the controller
        add(c, -1, NCD_KIND_WORKGROUP, "wg-extra", "wg", wire_str("wg" + std::to_string(k)), nullptr, nullptr, nullptr);
      } else if (k % 5 == 3) {
        ncdh_cluster_remove(c, -1, NCD_KIND_WORKGROUP, "default", "wg-extra");
      }
      if (k % 3 == 0) CHECK(ncdh_scan(c) == 0, "scan (writer2): %s", ncdh_last_error());
    }
  });
  // a third writer at the SAME ConfigMaps as the second, in larger batches: both write values in place
  // (same length most of the time), so they meet at the same records' pinned bytes
  std::thread writer3([&] {
    for (int k = 0; !stop.load(); ++k) {
      std::vector<std::string> names, vals, uids;
      for (int q = 0; q < 6; ++q) {
        const std::string n = std::to_string(T / 2 + (k + q) % (T / 2));
        names.push_back("cm-" + n);
        vals.push_back(wire_bytes_map("file", "d" + std::to_string(k % 10) + "-" + n));
        uids.push_back("c-" + n);
      }
      const char* ok[] = {"NexusAlgorithmTemplate"};
      std::vector<const char*> ou(6);
      ncdh_event ev[6];
      memset(ev, 0, sizeof(ev));
      for (int q = 0; q < 6; ++q) {
        ou[q] = uids[q].c_str();
        ev[q].cluster = q % 2 == 0 ? -1 : (k + q) % S;
        ev[q].kind = NCD_KIND_CONFIGMAP;
        ev[q].op = NCDH_EV_PUT;
        ev[q].ns = "default";
        ev[q].name = names[q].c_str();
        ev[q].uid = "";
        ev[q].value = (const uint8_t*)vals[q].data();
        ev[q].value_len = vals[q].size();
        ev[q].owner_kinds = ok;
        ev[q].owner_uids = &ou[q];
        ev[q].nowners = 1;
      }
      CHECK(ncdh_apply_events(c, ev, 6, 3) == 0, "apply_events (writer3): %s", ncdh_last_error());
    }
  });
  // the checked writer: rotates secrets of the lower half
  for (int k = 0; k < rounds; ++k) {
    const int t = k % (T / 2);
    const std::string n = std::to_string(t), algo = "algo-" + n;
    uint32_t code = 99;
    add(c, -1, NCD_KIND_SECRET, "sec-" + n, "", wire_bytes_map("token", "v" + std::to_string(k + 1) + "-" + n), ("c-" + n).c_str(), nullptr,
        nullptr);
    int rc = ncdh_lookup(c, NCD_KIND_TEMPLATE, "default", algo.c_str(), 1, 0, &code);
    // STALE — or already UPDATE_DATA when the other writer's scan got in between; never SKIP
    CHECK(rc == NCDH_STALE || (rc == 0 && code == NCD_UPDATE_DATA),
          "round %d: a Secret rotated after the scan was answered rc=%d code=%u instead of STALE", k, rc, code);
    char buf[2048];
    const int64_t pl = ncdh_plan_template(c, "default", algo.c_str(), buf, sizeof(buf));
    CHECK(pl == -NCDH_STALE || (pl > 0 && (size_t)pl < sizeof(buf) && std::string(buf, (size_t)pl).find("update-data") != std::string::npos),
          "round %d: plan neither stale nor an update (%lld)", k, (long long)pl);
    // an untouched template of the lower half is still answered from the snapshot
    const std::string other = "algo-" + std::to_string((t + 1) % (T / 2));
    rc = ncdh_lookup(c, NCD_KIND_TEMPLATE, "default", other.c_str(), 0, 1, &code);
    CHECK(rc == 0 && code == NCD_SKIP, "round %d: untouched template answered rc=%d code=%u", k, rc, code);
    CHECK(ncdh_scan(c) == 0, "scan: %s", ncdh_last_error());
    rc = ncdh_lookup(c, NCD_KIND_TEMPLATE, "default", algo.c_str(), 1, 0, &code);
    CHECK(rc == 0 && code == NCD_UPDATE_DATA, "round %d: after the scan rc=%d code=%u, expected UPDATE_DATA", k, rc, code);
    // the shards catch up (what UpdateSecret would do); checked next time this template comes round
    for (int s = 0; s < S; ++s)
      add(c, s, NCD_KIND_SECRET, "sec-" + n, "", wire_bytes_map("token", "v" + std::to_string(k + 1) + "-" + n),
          ("s" + std::to_string(s) + "-" + n).c_str(), nullptr, nullptr);
  }
  stop.store(true);
  writer2.join();
  writer3.join();
  for (auto& r : readers) r.join();
  CHECK(ncdh_scan(c) == 0, "last scan: %s", ncdh_last_error());
  // whatever the writers left behind, one more batch of equal values everywhere converges the upper half:
  // nothing written in place may have been lost on the way to the arenas
  {
    std::vector<std::string> names, vals, uids;
    std::vector<ncdh_event> ev;
    const char* ok[] = {"NexusAlgorithmTemplate"};
    for (int t = T / 2; t < T; ++t)
      for (int cluster = -1; cluster < S; ++cluster) {
        const std::string n = std::to_string(t);
        names.push_back("cm-" + n);
        vals.push_back(wire_bytes_map("file", "z9-" + n));
        uids.push_back((cluster < 0 ? "c-" : "s" + std::to_string(cluster) + "-") + n);
      }
    std::vector<const char*> ou(names.size());
    for (size_t q = 0; q < names.size(); ++q) {
      ncdh_event e;
      memset(&e, 0, sizeof(e));
      ou[q] = uids[q].c_str();
      e.cluster = (int)(q % (S + 1)) - 1;
      e.kind = NCD_KIND_CONFIGMAP;
      e.op = NCDH_EV_PUT;
      e.ns = "default";
      e.name = names[q].c_str();
      e.uid = "";
      e.value = (const uint8_t*)vals[q].data();
      e.value_len = vals[q].size();
      e.owner_kinds = ok;
      e.owner_uids = &ou[q];
      e.nowners = 1;
      ev.push_back(e);
    }
    CHECK(ncdh_apply_events(c, ev.data(), ev.size(), 4) == 0, "final batch: %s", ncdh_last_error());
    CHECK(ncdh_scan(c) == 0, "scan after the final batch: %s", ncdh_last_error());
    for (int t = T / 2; t < T; ++t)
      for (int s = 0; s < S; ++s) {
        uint32_t code = 99;
        const int rc = ncdh_lookup(c, NCD_KIND_TEMPLATE, "default", ("algo-" + std::to_string(t)).c_str(), 2, s, &code);
        CHECK(rc == 0 && code == NCD_SKIP, "converged ConfigMap of algo-%d on shard %d answered rc=%d code=%u", t, s, rc, code);
      }
  }
  ncdh_checker_destroy(c);
  printf("host_stress: %d rounds, %llu reader iterations, %llu stale answers seen by readers, %d failures\n", rounds,
         (unsigned long long)lookups.load(), (unsigned long long)stale_seen.load(), failures.load());
  return failures.load() ? 1 : 0;
}
