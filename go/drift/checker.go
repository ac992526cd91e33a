// Package drift — checker.go (UNCOMPILED, see go/README.md): the DriftChecker the patched handlers
// consult.  It reads the SAME informer caches the reference reads (listers, controller.go:268-278;
// shards.Shard.*Lister) and decides nothing itself: every code comes from ncd_scan on the GPU.
package drift

import (
	"fmt"
	"reflect"
	"strings"
	"sync"
	"sync/atomic"

	v1 "github.com/SneaksAndData/nexus-core/pkg/apis/science/v1"
	nexuslisters "github.com/SneaksAndData/nexus-core/pkg/generated/listers/science/v1"
	"github.com/SneaksAndData/nexus-core/pkg/shards"
	"k8s.io/apimachinery/pkg/labels"
	corelisters "k8s.io/client-go/listers/core/v1"
)

// object is one cache entry reduced to what the drift check reads.
type object struct {
	kind      uint32
	key, uid  string // key = "namespace/name" (cache.ObjectName.String())
	version   string // metadata.resourceVersion of the object that was packed
	value     reflect.Value // .Spec or .Data — the operand of reflect.DeepEqual in the reference
	owners    []owner
	secrets   []string // templates: GetSecretNames(), then GetConfigMapNames() (controller.go:505,567)
	configMap []string
}
type owner struct {
	isTemplate bool // Kind == "NexusAlgorithmTemplate" (controller.go:490)
	uid        string
}

// snapshot is what Lookup reads; swapped in atomically after every Scan, never modified afterwards:
// handlers on other goroutines (controller.go:875-877) read it without waiting for a Scan in flight.
type snapshot struct {
	res       *Result
	firstRow  map[string]int // kind + key -> first row of a template / workgroup
	ndeps     map[string]int
	order     []object            // controller objects in arena order (Dirty indexes into it)
	versions  []map[string]string // [0] controller, [1+s] shard s: kind + key -> resourceVersion that was packed
	unadopted []Adoption          // references adoptReferences (controller.go:647-695) would still adopt
}

// Freshness of an answer.  The reference's handlers read LIVE caches (controller.go:766, 792, 507, 516);
// a snapshot is only as fresh as its Scan, so Lookup compares the resourceVersions that were packed with
// what the listers hold NOW for everything the decision rests on, and says Stale instead of answering
// from a past the caches have left.  A caller that gets Stale takes the reference's own path for that
// object (lister.Get + reflect.DeepEqual + isMissingOwnership) or runs Scan and asks again; it never
// treats a missing answer as Skip.
type Freshness int

const (
	Fresh    Freshness = iota // code is valid for the caches as they are now
	NotFound                  // the object is not in the controller cache (controller.go:769-772)
	Stale                     // something the decision rests on changed after the Scan
)

type Checker struct {
	ctx       *Ctx
	mode      uint32
	templates nexuslisters.NexusAlgorithmTemplateLister
	workgroup nexuslisters.NexusAlgorithmWorkgroupLister
	secrets   corelisters.SecretLister
	configs   corelisters.ConfigMapLister
	shards    []*shards.Shard
	scanMu    sync.Mutex // one Scan at a time; the ncd_ctx is used under it
	snap      atomic.Pointer[snapshot]
}

func NewChecker(device int, mode uint32, t nexuslisters.NexusAlgorithmTemplateLister, w nexuslisters.NexusAlgorithmWorkgroupLister,
	s corelisters.SecretLister, c corelisters.ConfigMapLister, sh []*shards.Shard) (*Checker, error) {
	ctx, err := NewCtx(device)
	if err != nil {
		return nil, err
	}
	return &Checker{ctx: ctx, mode: mode, templates: t, workgroup: w, secrets: s, configs: c, shards: sh}, nil
}

// list turns one cluster's four listers into objects.
func list(t nexuslisters.NexusAlgorithmTemplateLister, w nexuslisters.NexusAlgorithmWorkgroupLister,
	s corelisters.SecretLister, c corelisters.ConfigMapLister) ([]object, error) {
	var out []object
	ts, err := t.List(labels.Everything())
	if err != nil {
		return nil, err
	}
	for _, x := range ts {
		out = append(out, object{kind: KindTemplate, key: x.Namespace + "/" + x.Name, uid: string(x.UID), version: x.ResourceVersion,
			value: reflect.ValueOf(x.Spec), secrets: x.GetSecretNames(), configMap: x.GetConfigMapNames()})
	}
	ws, err := w.List(labels.Everything())
	if err != nil {
		return nil, err
	}
	for _, x := range ws {
		out = append(out, object{kind: KindWorkgroup, key: x.Namespace + "/" + x.Name, uid: string(x.UID), version: x.ResourceVersion, value: reflect.ValueOf(x.Spec)})
	}
	ss, err := s.List(labels.Everything())
	if err != nil {
		return nil, err
	}
	for _, x := range ss {
		o := object{kind: KindSecret, key: x.Namespace + "/" + x.Name, uid: string(x.UID), version: x.ResourceVersion, value: reflect.ValueOf(x.Data)}
		for _, r := range x.OwnerReferences {
			o.owners = append(o.owners, owner{r.Kind == "NexusAlgorithmTemplate", string(r.UID)})
		}
		out = append(out, o)
	}
	cs, err := c.List(labels.Everything())
	if err != nil {
		return nil, err
	}
	for _, x := range cs {
		o := object{kind: KindConfigMap, key: x.Namespace + "/" + x.Name, uid: string(x.UID), version: x.ResourceVersion, value: reflect.ValueOf(x.Data)}
		for _, r := range x.OwnerReferences {
			o.owners = append(o.owners, owner{r.Kind == "NexusAlgorithmTemplate", string(r.UID)})
		}
		out = append(out, o)
	}
	return out, nil
}

// pack lays objects out as a committed arena (layout: ncdrift.h "arenas").  This is the simple
// form: a fresh arena per scan; csrc/host/drift_checker.cpp shows the resident form (SyncRecords /
// Recommit) a production packer would use.
func (k *Checker) pack(objs []object) (*Arena, error) {
	canon := make([][]byte, len(objs))
	var payload, aux, nowner, ndep uint64
	for i, o := range objs {
		b, err := Canon(nil, o.value)
		if err != nil {
			return nil, fmt.Errorf("%s: %w", o.key, err)
		}
		canon[i] = b
		payload += (uint64(len(b)) + 15) &^ 15
		aux += uint64(len(o.key) + len(o.uid))
		for _, r := range o.owners {
			aux += uint64(len(r.uid))
		}
		nowner += uint64(len(o.owners))
		for _, n := range append(append([]string{}, o.secrets...), o.configMap...) {
			aux += uint64(len(o.key) + len(n))
		}
		ndep += uint64(len(o.secrets) + len(o.configMap))
	}
	a, err := k.ctx.NewArena(payload, uint64(len(objs)), aux, nowner, ndep)
	if err != nil {
		return nil, err
	}
	var off, ax uint64
	var oi, di uint32
	put := func(s string) (uint32, uint32) {
		copy(a.Aux[ax:], s)
		o := uint32(ax)
		ax += uint64(len(s))
		return o, uint32(len(s))
	}
	for i, o := range objs {
		b := canon[i]
		a.RecOff[i], a.RecLen[i], a.Kind[i] = off, uint32(len(b)), uint8(o.kind)
		copy(a.Payload[off:], b)
		pad := (uint64(len(b)) + 15) &^ 15
		clear(a.Payload[off+uint64(len(b)) : off+pad])
		off += pad
		a.KeyOff[i], a.KeyLen[i] = put(o.key)
		a.UIDOff[i], a.UIDLen[i] = put(o.uid)
		a.OwnStart[i] = oi
		for _, r := range o.owners {
			if r.isTemplate {
				a.OwnIsTemplate[oi] = 1
			} else {
				a.OwnIsTemplate[oi] = 0
			}
			a.OwnUIDOff[oi], a.OwnUIDLen[oi] = put(r.uid)
			oi++
		}
		a.DepStart[i] = di
		ns := o.key[:strings.IndexByte(o.key, '/')+1] // "namespace/"
		for pass, names := range [][]string{o.secrets, o.configMap} {
			for _, n := range names {
				if pass == 0 {
					a.DepKind[di] = uint8(KindSecret)
				} else {
					a.DepKind[di] = uint8(KindConfigMap)
				}
				a.DepKeyOff[di], a.DepKeyLen[di] = put(ns + n)
				di++
			}
		}
	}
	a.OwnStart[len(objs)], a.DepStart[len(objs)] = oi, di
	if err := a.Commit(uint64(len(objs)), off, ax, uint64(oi), uint64(di)); err != nil {
		a.Close()
		return nil, err
	}
	return a, nil
}

// Scan: one drift check over the controller cluster and every shard.  Called once per resync
// sweep (main.go:70-71) or on demand; handlers then consult Lookup.
func (k *Checker) Scan() error {
	k.scanMu.Lock()
	defer k.scanMu.Unlock()
	ctrlObjs, err := list(k.templates, k.workgroup, k.secrets, k.configs)
	if err != nil {
		return err
	}
	ctrl, err := k.pack(ctrlObjs)
	if err != nil {
		return err
	}
	defer ctrl.Close()
	versions := []map[string]string{versionsOf(ctrlObjs)}
	arenas := make([]*Arena, 0, len(k.shards))
	defer func() {
		for _, a := range arenas {
			a.Close()
		}
	}()
	for _, sh := range k.shards {
		objs, err := list(sh.TemplateLister, sh.WorkgroupLister, sh.SecretLister, sh.ConfigMapLister)
		if err != nil {
			return err
		}
		a, err := k.pack(objs)
		if err != nil {
			return err
		}
		arenas = append(arenas, a)
		versions = append(versions, versionsOf(objs))
	}
	unadopted, err := k.ctx.Unadopted(ctrl)
	if err != nil {
		return err
	}
	res, err := k.ctx.Scan(ctrl, arenas, k.mode)
	if err != nil {
		return err
	}
	s := &snapshot{res: res, firstRow: map[string]int{}, ndeps: map[string]int{}, order: ctrlObjs, versions: versions, unadopted: unadopted}
	row := 0
	for _, o := range ctrlObjs {
		id := string(rune('0'+o.kind)) + o.key
		switch o.kind {
		case KindTemplate:
			s.firstRow[id], s.ndeps[id] = row, len(o.secrets)+len(o.configMap)
			row += 1 + len(o.secrets) + len(o.configMap)
		case KindWorkgroup:
			s.firstRow[id] = row
			row++
		}
	}
	k.snap.Store(s)
	return nil
}

func versionsOf(objs []object) map[string]string {
	m := make(map[string]string, len(objs))
	for _, o := range objs {
		m[string(rune('0'+o.kind))+o.key] = o.version
	}
	return m
}

// liveVersion is the resourceVersion a cluster's lister holds NOW for (kind, namespace, name); "" = absent.
func liveVersion(t nexuslisters.NexusAlgorithmTemplateLister, w nexuslisters.NexusAlgorithmWorkgroupLister,
	s corelisters.SecretLister, c corelisters.ConfigMapLister, kind uint32, namespace, name string) string {
	switch kind {
	case KindTemplate:
		if x, err := t.NexusAlgorithmTemplates(namespace).Get(name); err == nil {
			return x.ResourceVersion
		}
	case KindWorkgroup:
		if x, err := w.NexusAlgorithmWorkgroups(namespace).Get(name); err == nil {
			return x.ResourceVersion
		}
	case KindSecret:
		if x, err := s.Secrets(namespace).Get(name); err == nil {
			return x.ResourceVersion
		}
	case KindConfigMap:
		if x, err := c.ConfigMaps(namespace).Get(name); err == nil {
			return x.ResourceVersion
		}
	}
	return ""
}

// unchanged: does cluster (0 = controller, 1+s = shard s) still hold the version of (kind, key) that was packed?
func (k *Checker) unchanged(s *snapshot, cluster int, kind uint32, namespace, name string) bool {
	var live string
	if cluster == 0 {
		live = liveVersion(k.templates, k.workgroup, k.secrets, k.configs, kind, namespace, name)
	} else {
		sh := k.shards[cluster-1]
		live = liveVersion(sh.TemplateLister, sh.WorkgroupLister, sh.SecretLister, sh.ConfigMapLister, kind, namespace, name)
	}
	return s.versions[cluster][string(rune('0'+kind))+namespace+"/"+name] == live // "" == "": absent then and now
}

// Lookup: the decision of the last Scan, with its freshness.  dep = 0 the object itself, j+1 its j-th
// dependent (secrets first, then configmaps).  The decision for (template, dependent, shard) rests on the
// controller's template and dependent and on the shard's copies of both (the shard template's UID decides
// ownership, controller.go:529); if any of the four changed since the Scan the answer is Stale.
func (k *Checker) Lookup(kind uint32, namespace, name string, dep, shard int) (code uint32, f Freshness) {
	s := k.snap.Load()
	if s == nil {
		return 0, Stale
	}
	id := string(rune('0'+kind)) + namespace + "/" + name
	row, found := s.firstRow[id]
	if !found {
		if liveVersion(k.templates, k.workgroup, k.secrets, k.configs, kind, namespace, name) != "" {
			return 0, Stale // created after the Scan: not in the snapshot yet, which is not the same as absent
		}
		return 0, NotFound
	}
	if dep < 0 || dep > s.ndeps[id] || shard < 0 || shard >= s.res.Shards {
		return 0, Stale
	}
	if !k.unchanged(s, 0, kind, namespace, name) || !k.unchanged(s, 1+shard, kind, namespace, name) {
		return 0, Stale
	}
	if dep > 0 {
		o := s.order[s.objOf(id)]
		dk, dn := uint32(KindSecret), ""
		if dep <= len(o.secrets) {
			dn = o.secrets[dep-1]
		} else {
			dk, dn = KindConfigMap, o.configMap[dep-1-len(o.secrets)]
		}
		if !k.unchanged(s, 0, dk, namespace, dn) || !k.unchanged(s, 1+shard, dk, namespace, dn) {
			return 0, Stale
		}
	}
	return uint32(s.res.Codes[(row+dep)*s.res.Shards+shard]), Fresh
}

// objOf: index in order of the template / workgroup with this id.
func (s *snapshot) objOf(id string) int {
	for i, o := range s.order { // (a map in production; kept short here)
		if string(rune('0'+o.kind))+o.key == id {
			return i
		}
	}
	return -1
}

// Dirty: the templates / workgroups a resync still has to enqueue — those with a shard decision other
// than Skip, and the templates whose controller-side dependents adoptReferences (controller.go:647-695)
// would still adopt.  Everything else is converged as far as the shards and the ownerReferences go; the
// caller additionally enqueues objects whose status lacks the ready condition (reportTemplateInitCondition
// / reportTemplateSyncedCondition, controller.go:428-480, read nothing this package packs).
func (k *Checker) Dirty() (templates, workgroups []string) {
	s := k.snap.Load()
	if s == nil {
		return nil, nil
	}
	seen := map[uint32]bool{}
	add := func(i uint32) {
		if seen[i] {
			return
		}
		seen[i] = true
		o := s.order[i]
		if o.kind == KindTemplate {
			templates = append(templates, o.key)
		} else {
			workgroups = append(workgroups, o.key)
		}
	}
	for _, i := range s.res.Dirty {
		add(i)
	}
	for _, a := range s.unadopted {
		add(a.Obj)
	}
	return templates, workgroups
}

var _ = v1.SchemeGroupVersion // the Spec types come from nexus-core v1.4.1
