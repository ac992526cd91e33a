// cluster_image.cpp — the packer: a cluster image laid out in the pinned host arrays of its arena.
#include "cluster_image.hpp"

namespace ncdh {

ncd_arena_view ClusterImage::host_view() {
  ncd_arena_view v;
  if (ncd_arena_host_view(arena.a, &v) != NCD_OK) throw std::runtime_error(ncd_err("ncd_arena_host_view"));
  return v;
}

void ClusterImage::bind_arena() {
  host_payload = nullptr;
  host_rec_len = nullptr;
  if (arena.a) {
    const ncd_arena_view v = host_view();
    host_payload = v.payload;
    host_rec_len = v.rec_len;
  }
}

// Brings the pinned host arrays up to date with the image and says what has to be uploaded.
// Caller holds the cache lock exclusively; no GPU work here except creating a larger arena.
// h2d: bytes that upload will move.  tomb_floor: removals at or before this stamp are part of the
// published snapshot and can be forgotten.
Upload ClusterImage::prepare(ncd_ctx* ctx, int threads, uint64_t tomb_floor, uint64_t* h2d, bool* deferred) {
  for (auto it = tomb.begin(); it != tomb.end();)
    it = it->second <= tomb_floor ? tomb.erase(it) : std::next(it);
  if (deferred && (!arena.a || structural || force_full)) {
    *deferred = true;
    return Upload::None;
  }
  if (structural) {  // forget dead objects (their payload becomes unreferenced until the arena is re-made)
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
      ++layout;
    }
    dirty.clear();
  }
  if (!arena.a || structural || force_full) return prepare_structural(ctx, threads, h2d);
  if (dirty.empty()) return Upload::None;

  // same objects, some values changed: rewrite those records, to be pushed alone
  uint64_t append = 0;
  ncd_arena_view v = host_view();
  for (uint32_t i : dirty) {
    if (pad16(objs[i].len) > objs[i].rec_cap) append += pad16(objs[i].len);
    if (!objs[i].pending) v.rec_len[i] = objs[i].len;  // rewritten in place: a compaction below must move its new length
  }
  if (deferred && payload_used + append > payload_cap) {
    *deferred = true;
    return Upload::None;
  }
  if (payload_used + append > payload_cap && !reclaim(append)) {  // no room to append: move to a larger arena
    dirty.clear();
    return prepare_structural(ctx, threads, h2d, /*regrow=*/true);
  }
  for (uint32_t i : dirty) {  // offsets first (serial), bytes after (parallel)
    CachedObject& o = objs[i];
    assign(o, /*force_append=*/false);
    v.rec_off[i] = (uint64_t)o.rec_off;
    v.rec_len[i] = o.len;
    *h2d += pad16(o.len) + 4;
  }
  size_t to_copy = 0;
  for (uint32_t i : dirty) to_copy += objs[i].pending ? 1 : 0;
  if (to_copy)  // none when every value was written in place by its batch: not worth waking the workers
    parallel_for(dirty.size(), threads, 64, [&](size_t k) {
      CachedObject& o = objs[dirty[k]];
      if (o.pending) copy_record(v, o, o.pending);
    });
  in_place_last = 0;
  for (uint32_t i : dirty) {  // the arena holds the bytes now: the events' memory is released
    in_place_last += objs[i].pending ? 0 : 1;
    objs[i].pending = nullptr;
    objs[i].hold.reset();
    objs[i].direct = false;
  }
  pending_changed.swap(dirty);
  dirty.clear();
  ++uploads_values;
  return Upload::Values;
}

// Out of payload room: if at least a quarter of what is used is dead (old places of records that
// grew, records of deleted objects), ncd_arena_compact moves the live records together on both
// sides without any PCIe traffic, and the arena lives on.  True when `append` more bytes fit afterwards.
bool ClusterImage::reclaim(uint64_t append) {
  if (!arena.a || force_full) return false;
  uint64_t live = 0;
  for (const CachedObject& o : objs)
    if (o.live && o.rec_off >= 0) live += o.rec_cap;
  if (live + append > payload_cap || live * 4 > payload_used * 3) return false;
  uint64_t used = 0;
  if (ncd_arena_compact(arena.a, &used) != NCD_OK) return false;  // the arena is untouched on failure: regrow instead
  ncd_arena_view v = host_view();
  // positions follow the arena's object numbering of the last commit, which is objs order
  for (size_t i = 0; i < objs.size() && i < objs_committed; ++i) {
    CachedObject& o = objs[i];
    if (o.rec_off < 0) continue;
    o.rec_off = (int64_t)v.rec_off[i];
    o.rec_cap = pad16(v.rec_len[i]);  // the spare room a shrunken record kept is given back too
  }
  payload_used = used;
  ++compactions;
  return payload_used + append <= payload_cap;
}

// No cache lock held; one scan at a time.  Values uploads are batched by the caller.
void ClusterImage::upload(Upload what) {
  if (what == Upload::Full) {
    if (ncd_arena_commit(arena.a, objs_committed, payload_used, aux_used, own_used, dep_used) != NCD_OK)
      throw std::runtime_error(ncd_err("ncd_arena_commit"));
    force_full = false;
  } else if (what == Upload::Structural) {
    if (ncd_arena_recommit(arena.a, objs_committed, payload_used, aux_used, own_used, dep_used, dirty_from) != NCD_OK)
      throw std::runtime_error(ncd_err("ncd_arena_recommit"));
  }
}

Upload ClusterImage::prepare_structural(ncd_ctx* ctx, int threads, uint64_t* h2d, bool regrow) {
  uint64_t aux = 0, nowner = 0, ndep = 0, append = 0, compact = 0;
  for (const CachedObject& o : objs) {
    const ObjInfo& f = *o.info;
    aux += f.key.size() + f.uid.size();
    for (const OwnerRef& r : f.owners) aux += r.uid.size();
    nowner += f.owners.size();
    for (const auto& s : f.secrets) aux += f.key.size() + s.size();  // "namespace/" + name <= key + name
    for (const auto& s : f.configmaps) aux += f.key.size() + s.size();
    ndep += f.secrets.size() + f.configmaps.size();
    compact += pad16(o.len);
    if (o.pending || o.direct) append += pad16(o.len);
  }
  const bool fits = arena.a && !regrow && objs.size() <= nobj_cap && aux <= aux_cap && nowner <= nowner_cap && ndep <= ndep_cap &&
                    payload_used + append <= payload_cap;
  ArenaHandle old;  // the arena being left: the only copy of the records that are not pending
  const uint8_t* old_payload = host_payload;
  if (!fits) {
    // a fresh arena with room to grow; every record is placed again, compactly
    ncd_arena* fresh = nullptr;
    const uint64_t p_cap = compact + compact / 4 + (64 << 10);  // pinned and device memory both: a quarter of spare room
    const uint64_t n_cap = objs.size() + objs.size() / 2 + 16, a_cap = aux + aux / 2 + 1024, o_cap = nowner + nowner / 2 + 16,
                   d_cap = ndep + ndep / 2 + 16;
    if (ncd_arena_create(ctx, p_cap, n_cap, a_cap, o_cap, d_cap, NCD_ARENA_HOST_PINNED, &fresh) != NCD_OK)
      throw std::runtime_error(ncd_err("ncd_arena_create"));  // the old arena, if any, is untouched
    old.a = arena.a;
    arena.a = fresh;
    payload_cap = p_cap;
    nobj_cap = n_cap;
    aux_cap = a_cap;
    nowner_cap = o_cap;
    ndep_cap = d_cap;
    payload_used = 0;
    bind_arena();
  }
  ncd_arena_view v = host_view();
  dirty_from = fits ? payload_used : 0;
  write_all(v, threads, old.a ? old_payload : nullptr, !fits);
  old.reset();
  objs_committed = objs.size();
  const bool full = !fits || force_full;
  *h2d += (full ? payload_used : payload_used - dirty_from) + aux_used + objs_committed * 29 + 8 + own_used * 9 + dep_used * 9;
  ++(full ? uploads_full : uploads_structural);
  structural = false;
  dirty.clear();
  return full ? Upload::Full : Upload::Structural;
}

// Gives o a place in the payload: where it already is if it fits, else at the end.
void ClusterImage::assign(CachedObject& o, bool force_append) {
  const uint64_t need = pad16(o.len);
  if (o.rec_off < 0 || need > o.rec_cap || force_append) {
    o.rec_off = (int64_t)payload_used;
    o.rec_cap = need;
    payload_used += need;
  }
}

void ClusterImage::copy_record(ncd_arena_view& v, CachedObject& o, const uint8_t* src) {
  const uint64_t need = pad16(o.len);
  memcpy(v.payload + o.rec_off, src, o.len);
  memset(v.payload + o.rec_off + o.len, 0, need - o.len);
}

// Every per-object array from scratch (arena layout: ncdrift.h "arenas").  Same arena: records
// already in it keep their place and only pending ones are appended (only appended payload is
// re-uploaded); a record a batch rewrote in place is behind on the device, so it moves to the
// appended part as well.  New arena (replace_all): every record is placed again, pending ones from their
// events' memory, the others from the arena being left (old_payload).
void ClusterImage::write_all(ncd_arena_view& v, int threads, const uint8_t* old_payload, bool replace_all) {
  uint64_t ax = 0, oi = 0, di = 0, i = 0;
  auto put_aux = [&](const std::string& s, uint32_t* o, uint32_t* l) {
    memcpy(v.aux + ax, s.data(), s.size());
    *o = (uint32_t)ax;
    *l = (uint32_t)s.size();
    ax += s.size();
  };
  struct Copy {
    uint32_t obj;
    const uint8_t* src;
  };
  std::vector<Copy> to_copy;
  for (CachedObject& o : objs) {
    const ObjInfo& f = *o.info;
    if (o.pending || replace_all || o.direct) {
      const uint8_t* src = o.pending;
      if (!src) {
        const uint8_t* from = replace_all ? old_payload : v.payload;
        if (!from || o.rec_off < 0) throw std::runtime_error("packer: a record has no bytes anywhere");
        src = from + o.rec_off;
      }
      assign(o, /*force_append=*/true);
      to_copy.push_back({(uint32_t)i, src});
    }
    v.rec_off[i] = (uint64_t)o.rec_off;
    v.rec_len[i] = o.len;
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
  parallel_for(to_copy.size(), threads, 64, [&](size_t k) { copy_record(v, objs[to_copy[k].obj], to_copy[k].src); });
  for (const Copy& c : to_copy) {
    objs[c.obj].pending = nullptr;
    objs[c.obj].hold.reset();
    objs[c.obj].direct = false;
  }
}

}  // namespace ncdh
