// cluster_image.cpp — the packer: a cluster image laid out in the pinned host arrays of its arena.
#include "cluster_image.hpp"

namespace ncdh {

ncd_arena_view ClusterImage::host_view() {
  ncd_arena_view v;
  if (ncd_arena_host_view(arena.a, &v) != NCD_OK) throw std::runtime_error(ncd_err("ncd_arena_host_view"));
  return v;
}

// Brings the pinned host arrays up to date with the image and says what has to be uploaded.
// Caller holds the cache lock exclusively; no GPU work here except creating a larger arena.
// h2d: bytes that upload will move.  tomb_floor: removals at or before this stamp are part of the
// published snapshot and can be forgotten.
Upload ClusterImage::prepare(ncd_ctx* ctx, int threads, uint64_t tomb_floor, uint64_t* h2d) {
  for (auto it = tomb.begin(); it != tomb.end();)
    it = it->second <= tomb_floor ? tomb.erase(it) : std::next(it);
  if (structural) {  // forget dead objects (their payload becomes unreferenced until the next full commit)
    bool any_dead = false;
    for (const CachedObject& o : objs) any_dead = any_dead || !o.live;
    if (any_dead) {
      std::vector<CachedObject> keep;
      keep.reserve(objs.size());
      for (CachedObject& o : objs)
        if (o.live) keep.push_back(std::move(o));
      objs.swap(keep);
      index.clear();
      for (size_t i = 0; i < objs.size(); ++i) index.emplace(store_key(objs[i].info->kind, objs[i].info->key), i);
    }
    dirty.clear();
  }
  if (!arena.a || structural) return prepare_structural(ctx, threads, h2d);
  if (dirty.empty()) return Upload::None;

  // same objects, some values changed: rewrite those records, to be pushed alone
  uint64_t append = 0;
  for (uint32_t i : dirty)
    if (pad16(objs[i].canon.size()) > objs[i].rec_cap) append += pad16(objs[i].canon.size());
  if (payload_used + append > payload_cap) {  // no room to append: start over with a larger arena
    arena.reset();
    dirty.clear();
    return prepare_structural(ctx, threads, h2d);
  }
  ncd_arena_view v = host_view();
  for (uint32_t i : dirty) {  // offsets first (serial), bytes after (parallel)
    CachedObject& o = objs[i];
    assign(o, /*force_append=*/false);
    v.rec_off[i] = (uint64_t)o.rec_off;
    v.rec_len[i] = (uint32_t)o.canon.size();
    *h2d += pad16(o.canon.size()) + 4;
  }
  parallel_for(dirty.size(), threads, 64, [&](size_t k) { copy_record(v, objs[dirty[k]]); });
  pending_changed.swap(dirty);
  dirty.clear();
  ++uploads_values;
  return Upload::Values;
}

// No cache lock held; one scan at a time.  Values uploads are batched by the caller.
void ClusterImage::upload(Upload what) {
  if (what == Upload::Full) {
    if (ncd_arena_commit(arena.a, objs_committed, payload_used, aux_used, own_used, dep_used) != NCD_OK)
      throw std::runtime_error(ncd_err("ncd_arena_commit"));
  } else if (what == Upload::Structural) {
    if (ncd_arena_recommit(arena.a, objs_committed, payload_used, aux_used, own_used, dep_used, dirty_from) != NCD_OK)
      throw std::runtime_error(ncd_err("ncd_arena_recommit"));
  }
}

Upload ClusterImage::prepare_structural(ncd_ctx* ctx, int threads, uint64_t* h2d) {
  uint64_t aux = 0, nowner = 0, ndep = 0, append = 0, compact = 0;
  for (const CachedObject& o : objs) {
    const ObjInfo& f = *o.info;
    aux += f.key.size() + f.uid.size();
    for (const OwnerRef& r : f.owners) aux += r.uid.size();
    nowner += f.owners.size();
    for (const auto& s : f.secrets) aux += f.key.size() + s.size();  // "namespace/" + name <= key + name
    for (const auto& s : f.configmaps) aux += f.key.size() + s.size();
    ndep += f.secrets.size() + f.configmaps.size();
    compact += pad16(o.canon.size());
    if (o.rec_off < 0 || o.value_dirty) append += pad16(o.canon.size());
  }
  const bool fits = arena.a && objs.size() <= nobj_cap && aux <= aux_cap && nowner <= nowner_cap && ndep <= ndep_cap &&
                    payload_used + append <= payload_cap;
  if (!fits) {
    // a fresh arena with room to grow; every record is placed again, compactly
    arena.reset();
    payload_cap = compact + compact / 4 + (64 << 10);  // pinned and device memory both: a quarter of spare room
    nobj_cap = objs.size() + objs.size() / 2 + 16;
    aux_cap = aux + aux / 2 + 1024;
    nowner_cap = nowner + nowner / 2 + 16;
    ndep_cap = ndep + ndep / 2 + 16;
    if (ncd_arena_create(ctx, payload_cap, nobj_cap, aux_cap, nowner_cap, ndep_cap, NCD_ARENA_HOST_PINNED, &arena.a) != NCD_OK)
      throw std::runtime_error(ncd_err("ncd_arena_create"));
    payload_used = 0;
    for (CachedObject& o : objs) o.rec_off = -1;
  }
  ncd_arena_view v = host_view();
  dirty_from = payload_used;
  write_all(v, threads);
  objs_committed = objs.size();
  *h2d += (payload_used - dirty_from) + aux_used + objs_committed * 29 + 8 + own_used * 9 + dep_used * 9;
  ++(fits ? uploads_structural : uploads_full);
  structural = false;
  dirty.clear();
  return fits ? Upload::Structural : Upload::Full;
}

// Gives o a place in the payload: where it already is if it fits, else at the end.
void ClusterImage::assign(CachedObject& o, bool force_append) {
  const uint64_t need = pad16(o.canon.size());
  if (o.rec_off < 0 || need > o.rec_cap || force_append) {
    o.rec_off = (int64_t)payload_used;
    o.rec_cap = need;
    payload_used += need;
  }
}

void ClusterImage::copy_record(ncd_arena_view& v, CachedObject& o) {
  const uint64_t need = pad16(o.canon.size());
  memcpy(v.payload + o.rec_off, o.canon.data(), o.canon.size());
  memset(v.payload + o.rec_off + o.canon.size(), 0, need - o.canon.size());
  o.value_dirty = false;
}

// Every per-object array from scratch (arena layout: ncdrift.h "arenas"); records that are
// already on the device and did not change keep their place, the others are appended.
void ClusterImage::write_all(ncd_arena_view& v, int threads) {
  uint64_t ax = 0, oi = 0, di = 0, i = 0;
  auto put_aux = [&](const std::string& s, uint32_t* o, uint32_t* l) {
    memcpy(v.aux + ax, s.data(), s.size());
    *o = (uint32_t)ax;
    *l = (uint32_t)s.size();
    ax += s.size();
  };
  std::vector<uint32_t> to_copy;
  for (CachedObject& o : objs) {
    const ObjInfo& f = *o.info;
    if (o.rec_off < 0 || o.value_dirty) {  // only appended payload is re-uploaded
      assign(o, /*force_append=*/true);
      to_copy.push_back((uint32_t)i);
    }
    v.rec_off[i] = (uint64_t)o.rec_off;
    v.rec_len[i] = (uint32_t)o.canon.size();
    v.kind[i] = (uint8_t)f.kind;
    put_aux(f.key, &v.key_off[i], &v.key_len[i]);
    put_aux(f.uid, &v.uid_off[i], &v.uid_len[i]);
    v.own_start[i] = (uint32_t)oi;
    for (const OwnerRef& r : f.owners) {
      v.own_is_template[oi] = r.kind == "NexusAlgorithmTemplate";  // the Kind test of controller.go:490
      put_aux(r.uid, &v.own_uid_off[oi], &v.own_uid_len[oi]);
      ++oi;
    }
    v.dep_start[i] = (uint32_t)di;
    const std::string ns = f.key.substr(0, f.key.find('/') + 1);
    for (int pass = 0; pass < 2; ++pass)  // secrets first, then configmaps (:817 before :825)
      for (const std::string& name : pass == 0 ? f.secrets : f.configmaps) {
        v.dep_kind[di] = pass == 0 ? NCD_KIND_SECRET : NCD_KIND_CONFIGMAP;
        put_aux(ns + name, &v.dep_key_off[di], &v.dep_key_len[di]);
        ++di;
      }
    ++i;
  }
  v.own_start[i] = (uint32_t)oi;
  v.dep_start[i] = (uint32_t)di;
  aux_used = ax;
  own_used = oi;
  dep_used = di;
  parallel_for(to_copy.size(), threads, 64, [&](size_t k) { copy_record(v, objs[to_copy[k]]); });
}

}  // namespace ncdh
