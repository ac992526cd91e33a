// Package drift — ncdrift.go (UNCOMPILED, see go/README.md): cgo binding of include/ncdrift.h.
package drift

/*
#cgo LDFLAGS: -lncdrift
#include <stdlib.h>
#include <ncdrift.h>
*/
import "C"

import (
	"errors"
	"sync"
	"unsafe"
)

// Decision codes and modes (ncdrift.h).
const (
	Skip           = uint32(C.NCD_SKIP)
	Create         = uint32(C.NCD_CREATE)
	UpdateData     = uint32(C.NCD_UPDATE_DATA)
	UpdateOwner    = uint32(C.NCD_UPDATE_OWNER)
	ErrRogue       = uint32(C.NCD_ERR_ROGUE)
	ErrMissingRef  = uint32(C.NCD_ERR_MISSING_REF)
	KindTemplate   = uint32(C.NCD_KIND_TEMPLATE)
	KindWorkgroup  = uint32(C.NCD_KIND_WORKGROUP)
	KindSecret     = uint32(C.NCD_KIND_SECRET)
	KindConfigMap  = uint32(C.NCD_KIND_CONFIGMAP)
	ModeDirect     = uint32(C.NCD_MODE_DIRECT)
	ModeHash64     = uint32(C.NCD_MODE_HASH64)
	ModeSHA256     = uint32(C.NCD_MODE_SHA256)
	ModeXXH64      = uint32(C.NCD_MODE_XXH64)
	ModeConfirm    = uint32(C.NCD_MODE_CONFIRM)
	arenaHostPinned = C.uint32_t(C.NCD_ARENA_HOST_PINNED)
)

func lastErr(op string) error { return errors.New(op + ": " + C.GoString(C.ncd_last_error())) }

// Ctx wraps ncd_ctx; calls on one context are serialised.
type Ctx struct {
	mu sync.Mutex
	h  *C.ncd_ctx
}

func NewCtx(device int) (*Ctx, error) {
	c := &Ctx{}
	if C.ncd_init(C.int(device), &c.h) != 0 {
		return nil, lastErr("ncd_init") // no GPU is an error; there is no CPU fallback
	}
	return c, nil
}

func (c *Ctx) Close() { C.ncd_destroy(c.h) }

// Arena wraps ncd_arena.  Its pinned arrays are C memory: Go writes through the slices below
// and C keeps no Go pointer (cgo rule).
type Arena struct {
	h *C.ncd_arena
	v C.ncd_arena_view

	Payload                       []byte
	RecOff                        []uint64
	RecLen, KeyOff, KeyLen        []uint32
	UIDOff, UIDLen, OwnStart      []uint32
	OwnUIDOff, OwnUIDLen          []uint32
	DepStart, DepKeyOff, DepKeyLen []uint32
	Kind, OwnIsTemplate, DepKind  []uint8
	Aux                           []byte
}

func u32s(p *C.uint32_t, n C.uint64_t) []uint32 { return unsafe.Slice((*uint32)(unsafe.Pointer(p)), int(n)) }
func u8s(p *C.uint8_t, n C.uint64_t) []uint8   { return unsafe.Slice((*uint8)(unsafe.Pointer(p)), int(n)) }

func (c *Ctx) NewArena(payload, nobj, aux, nowner, ndep uint64) (*Arena, error) {
	a := &Arena{}
	if C.ncd_arena_create(c.h, C.uint64_t(payload), C.uint64_t(nobj), C.uint64_t(aux), C.uint64_t(nowner),
		C.uint64_t(ndep), arenaHostPinned, &a.h) != 0 {
		return nil, lastErr("ncd_arena_create")
	}
	if C.ncd_arena_host_view(a.h, &a.v) != 0 {
		return nil, lastErr("ncd_arena_host_view")
	}
	v := &a.v
	a.Payload = u8s(v.payload, v.payload_bytes)
	a.RecOff = unsafe.Slice((*uint64)(unsafe.Pointer(v.rec_off)), int(v.nobj))
	a.RecLen, a.Kind = u32s(v.rec_len, v.nobj), u8s(v.kind, v.nobj)
	a.KeyOff, a.KeyLen = u32s(v.key_off, v.nobj), u32s(v.key_len, v.nobj)
	a.UIDOff, a.UIDLen = u32s(v.uid_off, v.nobj), u32s(v.uid_len, v.nobj)
	a.OwnStart, a.DepStart = u32s(v.own_start, v.nobj+1), u32s(v.dep_start, v.nobj+1)
	a.OwnIsTemplate, a.OwnUIDOff, a.OwnUIDLen = u8s(v.own_is_template, v.nowner), u32s(v.own_uid_off, v.nowner), u32s(v.own_uid_len, v.nowner)
	a.DepKind, a.DepKeyOff, a.DepKeyLen = u8s(v.dep_kind, v.ndep), u32s(v.dep_key_off, v.ndep), u32s(v.dep_key_len, v.ndep)
	a.Aux = u8s(v.aux, v.aux_bytes)
	return a, nil
}

func (a *Arena) Close() { C.ncd_arena_destroy(a.h) }

// Commit uploads everything and builds the index (first scan, or after the arena was re-created).
func (a *Arena) Commit(nobj, payload, aux, nowner, ndep uint64) error {
	if C.ncd_arena_commit(a.h, C.uint64_t(nobj), C.uint64_t(payload), C.uint64_t(aux), C.uint64_t(nowner), C.uint64_t(ndep)) != 0 {
		return lastErr("ncd_arena_commit")
	}
	return nil
}

// Recommit uploads the per-object arrays and payload[dirtyFrom:payload] only (Add/Delete/re-own).
func (a *Arena) Recommit(nobj, payload, aux, nowner, ndep, dirtyFrom uint64) error {
	if C.ncd_arena_recommit(a.h, C.uint64_t(nobj), C.uint64_t(payload), C.uint64_t(aux), C.uint64_t(nowner),
		C.uint64_t(ndep), C.uint64_t(dirtyFrom)) != 0 {
		return lastErr("ncd_arena_recommit")
	}
	return nil
}

// SyncRecords pushes the records of the listed objects, rewritten in Payload/RecOff/RecLen (UpdateFunc).
func (a *Arena) SyncRecords(objs []uint32, payload uint64) error {
	var p *C.uint32_t
	if len(objs) > 0 {
		p = (*C.uint32_t)(unsafe.Pointer(&objs[0])) // read during the call only
	}
	if C.ncd_arena_sync_records(a.h, C.uint64_t(len(objs)), p, C.uint64_t(payload)) != 0 {
		return lastErr("ncd_arena_sync_records")
	}
	return nil
}

// Save writes the committed arena to a snapshot file; LoadArena brings one back, checked and indexed
// (offline diffing, SURVEY section 8 row f-4).
func (a *Arena) Save(path string) error {
	cp := C.CString(path)
	defer C.free(unsafe.Pointer(cp))
	if C.ncd_arena_save(a.h, cp) != 0 {
		return lastErr("ncd_arena_save")
	}
	return nil
}

func (c *Ctx) LoadArena(path string, deviceOnly bool, payloadSlack uint64) (*Arena, error) {
	cp := C.CString(path)
	defer C.free(unsafe.Pointer(cp))
	flags := C.uint32_t(C.NCD_ARENA_HOST_PINNED)
	if deviceOnly {
		flags = C.uint32_t(C.NCD_ARENA_DEVICE_ONLY)
	}
	a := &Arena{}
	if C.ncd_arena_load(c.h, cp, flags, C.uint64_t(payloadSlack), &a.h) != 0 {
		return nil, lastErr("ncd_arena_load")
	}
	return a, nil
}

// Entry == ncd_entry.
type Entry struct{ Obj, Dep, Shard, Code uint32 }

// Result of one scan, copied into Go memory.
type Result struct {
	Rows, Shards int
	Codes        []uint8  // Rows x Shards, row-major
	Entries      []Entry  // codes != Skip, (row, shard) order
	Dirty        []uint32 // controller indices of templates / workgroups with work to do
}

// Scan runs the drift check (ncd_scan) and copies the result out.
func (c *Ctx) Scan(ctrl *Arena, shards []*Arena, mode uint32) (*Result, error) {
	c.mu.Lock()
	defer c.mu.Unlock()
	hs := make([]*C.ncd_arena, len(shards))
	for i, s := range shards {
		hs[i] = s.h
	}
	// the handle array lives in C memory for the duration of the call (no Go pointers to Go pointers)
	n := len(hs)
	var arr **C.ncd_arena
	if n > 0 {
		arr = (**C.ncd_arena)(C.malloc(C.size_t(n) * C.size_t(unsafe.Sizeof(uintptr(0)))))
		defer C.free(unsafe.Pointer(arr))
		copy(unsafe.Slice(arr, n), hs)
	}
	var d *C.ncd_diff
	if C.ncd_scan(c.h, ctrl.h, arr, C.int(n), C.uint32_t(mode), &d) != 0 {
		return nil, lastErr("ncd_scan") // the handler returns it: AddRateLimited requeue (controller.go:405-411)
	}
	defer C.ncd_diff_destroy(d)
	var rows C.uint64_t
	if C.ncd_arena_rows(ctrl.h, &rows) != 0 {
		return nil, lastErr("ncd_arena_rows")
	}
	r := &Result{Rows: int(rows), Shards: n, Codes: make([]uint8, int(rows)*n)}
	if len(r.Codes) > 0 && C.ncd_diff_codes(d, (*C.uint8_t)(unsafe.Pointer(&r.Codes[0])), C.uint64_t(len(r.Codes))) != 0 {
		return nil, lastErr("ncd_diff_codes")
	}
	var ents *C.ncd_entry
	var ne C.uint64_t
	if C.ncd_diff_view(d, &ents, &ne) != 0 {
		return nil, lastErr("ncd_diff_view")
	}
	r.Entries = make([]Entry, int(ne))
	copy(r.Entries, unsafe.Slice((*Entry)(unsafe.Pointer(ents)), int(ne)))
	var nd C.uint64_t
	if C.ncd_diff_dirty_objects(d, nil, 0, &nd) != 0 {
		return nil, lastErr("ncd_diff_dirty_objects")
	}
	r.Dirty = make([]uint32, int(nd))
	if nd > 0 && C.ncd_diff_dirty_objects(d, (*C.uint32_t)(unsafe.Pointer(&r.Dirty[0])), nd, &nd) != 0 {
		return nil, lastErr("ncd_diff_dirty_objects")
	}
	return r, nil
}

// Adoption == ncd_adoption: a (template, dependent) reference whose controller-side object does not
// carry the template's ownerReference yet (adoptReferences, controller.go:647-695).
type Adoption struct{ Obj, Dep uint32 }

// Unadopted lists the references adoptReferences would still adopt (ncd_find_unadopted).
func (c *Ctx) Unadopted(ctrl *Arena) ([]Adoption, error) {
	c.mu.Lock()
	defer c.mu.Unlock()
	var n C.uint64_t
	out := make([]Adoption, 1024)
	for {
		if C.ncd_find_unadopted(c.h, ctrl.h, (*C.ncd_adoption)(unsafe.Pointer(&out[0])), C.uint64_t(len(out)), &n) != 0 {
			return nil, lastErr("ncd_find_unadopted")
		}
		if int(n) <= len(out) {
			return out[:n], nil
		}
		out = make([]Adoption, int(n))
	}
}

// KeyHash is the multi-GPU partition function: gpu = KeyHash(kind, "ns/name") % nGPUs.
func KeyHash(kind uint32, key string) uint64 {
	return uint64(C.ncd_keyhash(C.uint32_t(kind), unsafe.Pointer(unsafe.StringData(key)), C.uint32_t(len(key))))
}

// Exchange is one GPU's end of the peer-memory exchange of diff lists (ncd_xchg_*): after a
// partitioned scan every GPU holds every GPU's list.
type Exchange struct {
	c     *Ctx
	h     *C.ncd_xchg
	world int
	cap   uint64
}

func (c *Ctx) NewExchange(world, rank int, capEntries uint64) (*Exchange, error) {
	x := &Exchange{c: c, world: world, cap: capEntries}
	if C.ncd_xchg_create(c.h, C.int(world), C.int(rank), C.uint64_t(capEntries), &x.h) != 0 {
		return nil, lastErr("ncd_xchg_create")
	}
	return x, nil
}

func (x *Exchange) Close() { C.ncd_xchg_destroy(x.h) }

// Handle returns the 64 bytes every other rank needs to map this rank's buffer.
func (x *Exchange) Handle() ([64]byte, error) {
	var h [64]byte
	if C.ncd_xchg_handle(x.h, unsafe.Pointer(&h[0])) != 0 {
		return h, lastErr("ncd_xchg_handle")
	}
	return h, nil
}

// Open maps the buffers of all ranks (handles in rank order; the own entry is ignored).
func (x *Exchange) Open(handles [][64]byte) error {
	if len(handles) != x.world {
		return errors.New("ncd_xchg_open: one handle per rank")
	}
	flat := make([]byte, 0, 64*len(handles))
	for i := range handles {
		flat = append(flat, handles[i][:]...)
	}
	if C.ncd_xchg_open(x.h, unsafe.Pointer(&flat[0])) != 0 {
		return lastErr("ncd_xchg_open")
	}
	return nil
}

// ScanExchange enqueues the drift check and the exchange of its diff list (asynchronous).
func (c *Ctx) ScanExchange(ctrl *Arena, shards []*Arena, mode uint32, x *Exchange) error {
	c.mu.Lock()
	defer c.mu.Unlock()
	n := len(shards)
	var arr **C.ncd_arena
	if n > 0 {
		arr = (**C.ncd_arena)(C.malloc(C.size_t(n) * C.size_t(unsafe.Sizeof(uintptr(0)))))
		defer C.free(unsafe.Pointer(arr))
		for i, s := range shards {
			unsafe.Slice(arr, n)[i] = s.h
		}
	}
	if C.ncd_scan_exchange(c.h, ctrl.h, arr, C.int(n), C.uint32_t(mode), x.h) != 0 {
		return lastErr("ncd_scan_exchange")
	}
	return nil
}

// Fetch waits for the latest exchange and returns every rank's entries, in rank order.
func (x *Exchange) Fetch() ([][]Entry, error) {
	x.c.mu.Lock()
	defer x.c.mu.Unlock()
	words := 4 + 4*int(x.cap)
	buf := make([]uint32, x.world*words)
	if C.ncd_xchg_fetch(x.h, unsafe.Pointer(&buf[0]), C.uint64_t(4*len(buf))) != 0 {
		return nil, lastErr("ncd_xchg_fetch")
	}
	out := make([][]Entry, x.world)
	for r := 0; r < x.world; r++ {
		slot := buf[r*words : (r+1)*words]
		if slot[3] != 0 {
			return nil, errors.New("ncd_xchg_fetch: a diff list exceeded the exchange capacity")
		}
		out[r] = make([]Entry, int(slot[0]))
		copy(out[r], unsafe.Slice((*Entry)(unsafe.Pointer(&slot[4])), int(slot[0])))
	}
	return out, nil
}
