// The remaining hot-path bindings: Service, VirtualService, culler, the sparse StatefulSet form, the one-call
// reconcile step, the namespace -> shard table and the resident store.
//
// NOT COMPILED in the build image (no Go toolchain there), like nbgraft.go; the C entry points are built and
// tested (tests/c/abi_client.c, tests/test_*.py).  Call sites in the reference: INTEGRATION.md section 2.
package nbgraft

/*
#include "nbgraft.h"
*/
import "C"

import (
	"fmt"
	"runtime"
	"unsafe"
)

func status(where string, rc C.int) error {
	if rc == C.NBG_OK {
		return nil
	}
	return fmt.Errorf("%s: %s", where, C.GoString(C.nbg_strerror(rc)))
}

// ptr returns the address of a slice's first element (nil for an empty slice) and pins it for the call.
func ptr[T any](pin *runtime.Pinner, s []T) unsafe.Pointer {
	if len(s) == 0 {
		return nil
	}
	pin.Pin(&s[0])
	return unsafe.Pointer(&s[0])
}

func (m KvMap) c(pin *runtime.Pinner) C.nbg_kvmap {
	return C.nbg_kvmap{off: (*C.uint32_t)(ptr(pin, m.Off)), ent: (*C.nbg_kv)(ptr(pin, m.Ent))}
}

func outputs(n int) ([]byte, []uint64) { return make([]byte, n), make([]uint64, (n+63)/64) }

// ---- CopyServiceFields (util.go:166-195) --------------------------------------------------------------
type SvcBatch struct {
	Flags              []uint8  // n
	PortsFrom, PortsTo []uint64 // n * NBG_SVC_SLOTS (8)
	LabelsFrom, LabelsTo, AnnFrom, AnnTo, SelFrom, SelTo KvMap
}

func (c *Context) DiffServices(b *SvcBatch) (reason []byte, bitmap []uint64, err error) {
	n := len(b.Flags)
	reason, bitmap = outputs(n)
	if n == 0 {
		return
	}
	var pin runtime.Pinner
	defer pin.Unpin()
	in := C.nbg_svc_batch{
		flags: (*C.uint8_t)(ptr(&pin, b.Flags)), ports_from: (*C.uint64_t)(ptr(&pin, b.PortsFrom)), ports_to: (*C.uint64_t)(ptr(&pin, b.PortsTo)),
		labels_from: b.LabelsFrom.c(&pin), labels_to: b.LabelsTo.c(&pin), ann_from: b.AnnFrom.c(&pin), ann_to: b.AnnTo.c(&pin),
		sel_from: b.SelFrom.c(&pin), sel_to: b.SelTo.c(&pin),
	}
	err = status("nbg_diff_svc", C.nbg_diff_svc(c.h, &in, C.uint64_t(n), (*C.uint8_t)(ptr(&pin, reason)), (*C.uint64_t)(ptr(&pin, bitmap))))
	return
}

// ---- CopyVirtualService (util.go:199-219) -------------------------------------------------------------
type VsBatch struct {
	Flags            []uint8  // n
	SpecFrom, SpecTo []uint64 // n * 8
	HdrFrom, HdrTo   KvMap
}

func (c *Context) DiffVirtualServices(b *VsBatch) (reason []byte, bitmap []uint64, err error) {
	n := len(b.Flags)
	reason, bitmap = outputs(n)
	if n == 0 {
		return
	}
	var pin runtime.Pinner
	defer pin.Unpin()
	in := C.nbg_vs_batch{flags: (*C.uint8_t)(ptr(&pin, b.Flags)), spec_from: (*C.uint64_t)(ptr(&pin, b.SpecFrom)),
		spec_to: (*C.uint64_t)(ptr(&pin, b.SpecTo)), hdr_from: b.HdrFrom.c(&pin), hdr_to: b.HdrTo.c(&pin)}
	err = status("nbg_diff_vs", C.nbg_diff_vs(c.h, &in, C.uint64_t(n), (*C.uint8_t)(ptr(&pin, reason)), (*C.uint64_t)(ptr(&pin, bitmap))))
	return
}

// ---- the culler's decision (culling_controller.go:86-241, 315-428) ------------------------------------
type CullBatch struct {
	KOff         []uint32 // n+1
	KSec         []int64
	KNsf         []uint32
	TOff         []uint32 // n+1
	TSec         []int64
	TNsf         []uint32
	AnnSec, ChkSec []int64  // n
	AnnNsf, ChkNsf []uint32 // n
	NbFlags      []uint8  // n
}

func (b *CullBatch) c(pin *runtime.Pinner) C.nbg_cull_batch {
	return C.nbg_cull_batch{
		k_off: (*C.uint32_t)(ptr(pin, b.KOff)), k_sec: (*C.int64_t)(ptr(pin, b.KSec)), k_nsf: (*C.uint32_t)(ptr(pin, b.KNsf)),
		t_off: (*C.uint32_t)(ptr(pin, b.TOff)), t_sec: (*C.int64_t)(ptr(pin, b.TSec)), t_nsf: (*C.uint32_t)(ptr(pin, b.TNsf)),
		ann_sec: (*C.int64_t)(ptr(pin, b.AnnSec)), ann_nsf: (*C.uint32_t)(ptr(pin, b.AnnNsf)),
		chk_sec: (*C.int64_t)(ptr(pin, b.ChkSec)), chk_nsf: (*C.uint32_t)(ptr(pin, b.ChkNsf)),
		nb_flags: (*C.uint8_t)(ptr(pin, b.NbFlags)),
	}
}

type CullParams struct {
	NowSec                        int64
	NowNsec                       int32
	CullIdleMin, CheckPeriodMin   int32 // CULL_IDLE_TIME, IDLENESS_CHECK_PERIOD
	Reconcile                     bool  // true: the decision flow of CullingReconciler.Reconcile; false: the bare predicates
}

func (p CullParams) c() C.nbg_cull_params {
	mode := C.int32_t(C.NBG_CULL_PREDICATES)
	if p.Reconcile {
		mode = C.NBG_CULL_RECONCILE
	}
	return C.nbg_cull_params{now_sec: C.int64_t(p.NowSec), now_nsec: C.int32_t(p.NowNsec),
		cull_idle_min: C.int32_t(p.CullIdleMin), check_period_min: C.int32_t(p.CheckPeriodMin), mode: mode}
}

type CullResult struct {
	Action                   []uint8  // NBG_ACT_* bits per notebook
	Argmax                   []uint32 // record that supplies the new last-activity (src 2/3)
	CullBitmap, PeriodBitmap []uint64
}

func (c *Context) Cull(b *CullBatch, p CullParams) (r CullResult, err error) {
	n := len(b.NbFlags)
	r = CullResult{make([]uint8, n), make([]uint32, n), make([]uint64, (n+63)/64), make([]uint64, (n+63)/64)}
	if n == 0 {
		return
	}
	var pin runtime.Pinner
	defer pin.Unpin()
	in, params := b.c(&pin), p.c()
	err = status("nbg_cull", C.nbg_cull(c.h, &in, C.uint64_t(n), &params, (*C.uint8_t)(ptr(&pin, r.Action)),
		(*C.uint32_t)(ptr(&pin, r.Argmax)), (*C.uint64_t)(ptr(&pin, r.CullBitmap)), (*C.uint64_t)(ptr(&pin, r.PeriodBitmap))))
	return
}

// ---- sparse StatefulSet form and device-generated desired Service / VirtualService --------------------
type PodSparse struct {
	Mask []uint64 // n * 3
	Off  []uint32 // n + 1, even
	Val  []uint64 // allocate over C.nbg_alloc_pinned memory: 16-byte aligned and fastest to copy
}

func (s PodSparse) c(pin *runtime.Pinner) C.nbg_podsparse {
	return C.nbg_podsparse{mask: (*C.uint64_t)(ptr(pin, s.Mask)), off: (*C.uint32_t)(ptr(pin, s.Off)), val: (*C.uint64_t)(ptr(pin, s.Val))}
}

type StsSparseBatch struct {
	ReplicasFrom, ReplicasTo []int32
	TmplFrom, TmplTo         []uint64
	PodFrom, PodTo           PodSparse
	LabelsFrom, LabelsTo     KvMap
	AnnFrom, AnnTo           KvMap
}

func (b *StsSparseBatch) c(pin *runtime.Pinner) C.nbg_sts_sparse_batch {
	return C.nbg_sts_sparse_batch{
		replicas_from: (*C.int32_t)(ptr(pin, b.ReplicasFrom)), replicas_to: (*C.int32_t)(ptr(pin, b.ReplicasTo)),
		tmpl_from: (*C.uint64_t)(ptr(pin, b.TmplFrom)), tmpl_to: (*C.uint64_t)(ptr(pin, b.TmplTo)),
		pod_from: b.PodFrom.c(pin), pod_to: b.PodTo.c(pin),
		labels_from: b.LabelsFrom.c(pin), labels_to: b.LabelsTo.c(pin), ann_from: b.AnnFrom.c(pin), ann_to: b.AnnTo.c(pin),
	}
}

type SvcGenBatch struct {
	NameHash   []uint64 // hash64(notebook name)
	TargetPort []int32
	Flags      []uint8
	PortsTo    []uint64 // n * NBG_SVC_SLOTS (8)
	LabelsTo, AnnTo, SelTo KvMap
}

func (b *SvcGenBatch) c(pin *runtime.Pinner) C.nbg_svc_gen_batch {
	return C.nbg_svc_gen_batch{name_hash: (*C.uint64_t)(ptr(pin, b.NameHash)), target_port: (*C.int32_t)(ptr(pin, b.TargetPort)),
		flags: (*C.uint8_t)(ptr(pin, b.Flags)), ports_to: (*C.uint64_t)(ptr(pin, b.PortsTo)),
		labels_to: b.LabelsTo.c(pin), ann_to: b.AnnTo.c(pin), sel_to: b.SelTo.c(pin)}
}

type VsGenBatch struct {
	PrefixHash, RewriteHash, HostHash []uint64
	HdrFrom                           KvMap
	Flags                             []uint8
	SpecTo                            []uint64 // n * 8
	HdrTo                             KvMap
}

func (b *VsGenBatch) c(pin *runtime.Pinner) C.nbg_vs_gen_batch {
	return C.nbg_vs_gen_batch{prefix_hash: (*C.uint64_t)(ptr(pin, b.PrefixHash)), rewrite_hash: (*C.uint64_t)(ptr(pin, b.RewriteHash)),
		host_hash: (*C.uint64_t)(ptr(pin, b.HostHash)), hdr_from: b.HdrFrom.c(pin), flags: (*C.uint8_t)(ptr(pin, b.Flags)),
		spec_to: (*C.uint64_t)(ptr(pin, b.SpecTo)), hdr_to: b.HdrTo.c(pin)}
}

// VsConsts: hash64(ISTIO_HOST), hash64(ISTIO_GATEWAY) and the packer's residual digest of the generated spec.
type VsConsts struct{ IstioHostHash, GatewayHash, Residual uint64 }

// StepResult holds what one pass of both reconcilers decided for every notebook of the batch.
type StepResult struct {
	StsReason, SvcReason, VsReason []uint8
	StsBitmap, SvcBitmap, VsBitmap []uint64
	Cull                           CullResult
}

// Step is one pass of NotebookReconciler.Reconcile's three comparisons (notebook_controller.go:197,228,689) and
// CullingReconciler.Reconcile's decision (culling_controller.go:156-195) over n notebooks, as one copy/compute
// pipeline.  Any batch may be nil.
func (c *Context) Step(sts *StsSparseBatch, svc *SvcGenBatch, vs *VsGenBatch, k VsConsts, cull *CullBatch, p CullParams, n int) (r StepResult, err error) {
	words := (n + 63) / 64
	var pin runtime.Pinner
	defer pin.Unpin()
	var out C.nbg_step_outputs
	var cs *C.nbg_sts_sparse_batch
	var cv *C.nbg_svc_gen_batch
	var cw *C.nbg_vs_gen_batch
	var cc *C.nbg_cull_batch
	if sts != nil {
		v := sts.c(&pin)
		cs = &v
		r.StsReason, r.StsBitmap = make([]uint8, n), make([]uint64, words)
		out.sts_reason, out.sts_bitmap = (*C.uint8_t)(ptr(&pin, r.StsReason)), (*C.uint64_t)(ptr(&pin, r.StsBitmap))
	}
	if svc != nil {
		v := svc.c(&pin)
		cv = &v
		r.SvcReason, r.SvcBitmap = make([]uint8, n), make([]uint64, words)
		out.svc_reason, out.svc_bitmap = (*C.uint8_t)(ptr(&pin, r.SvcReason)), (*C.uint64_t)(ptr(&pin, r.SvcBitmap))
	}
	if vs != nil {
		v := vs.c(&pin)
		cw = &v
		r.VsReason, r.VsBitmap = make([]uint8, n), make([]uint64, words)
		out.vs_reason, out.vs_bitmap = (*C.uint8_t)(ptr(&pin, r.VsReason)), (*C.uint64_t)(ptr(&pin, r.VsBitmap))
	}
	if cull != nil {
		v := cull.c(&pin)
		cc = &v
		r.Cull = CullResult{make([]uint8, n), make([]uint32, n), make([]uint64, words), make([]uint64, words)}
		out.action, out.argmax = (*C.uint8_t)(ptr(&pin, r.Cull.Action)), (*C.uint32_t)(ptr(&pin, r.Cull.Argmax))
		out.cull_bitmap, out.period_bitmap = (*C.uint64_t)(ptr(&pin, r.Cull.CullBitmap)), (*C.uint64_t)(ptr(&pin, r.Cull.PeriodBitmap))
	}
	consts := C.nbg_vs_gen_consts{istio_host_hash: C.uint64_t(k.IstioHostHash), gateway_hash: C.uint64_t(k.GatewayHash), residual: C.uint64_t(k.Residual)}
	params := p.c()
	err = status("nbg_step", C.nbg_step(c.h, cs, cv, cw, &consts, cc, &params, C.uint64_t(n), &out))
	return
}

// ---- namespace -> shard (notebook_controller.go:159: the namespace is the unit) ------------------------
func ShardOf(nsHash uint64, shards uint32) uint32 { return uint32(C.nbg_shard_of(C.uint64_t(nsHash), C.uint32_t(shards))) }

// BalanceShards is the size-aware table: whole namespaces, largest first, each to the lightest shard.
func BalanceShards(nsHash, nsSize []uint64, shards uint32) (table []uint32, load []uint64, err error) {
	table, load = make([]uint32, len(nsHash)), make([]uint64, shards)
	var pin runtime.Pinner
	defer pin.Unpin()
	err = status("nbg_balance_shards", C.nbg_balance_shards((*C.uint64_t)(ptr(&pin, nsHash)), (*C.uint64_t)(ptr(&pin, nsSize)),
		C.uint64_t(len(nsHash)), C.uint32_t(shards), (*C.uint32_t)(ptr(&pin, table)), (*C.uint64_t)(ptr(&pin, load))))
	return
}

// ---- the resident fleet: pack an object when the informer delivers it, put its row, step every tick -----
type Store struct{ h *C.nbg_store }

func (c *Context) NewStore(device int, capacity uint64) (*Store, error) {
	var h *C.nbg_store
	if err := status("nbg_store_create", C.nbg_store_create(c.h, C.int(device), C.uint64_t(capacity), &h)); err != nil {
		return nil, err
	}
	return &Store{h}, nil
}

func (s *Store) Close() { C.nbg_store_destroy(s.h); s.h = nil }

// Put replaces the rows of the named slots (each slot once per call).  Any batch may be nil.
func (s *Store) Put(slots []uint64, sts *StsSparseBatch, svc *SvcGenBatch, vs *VsGenBatch, cull *CullBatch) error {
	var pin runtime.Pinner
	defer pin.Unpin()
	var cs *C.nbg_sts_sparse_batch
	var cv *C.nbg_svc_gen_batch
	var cw *C.nbg_vs_gen_batch
	var cc *C.nbg_cull_batch
	if sts != nil {
		v := sts.c(&pin)
		cs = &v
	}
	if svc != nil {
		v := svc.c(&pin)
		cv = &v
	}
	if vs != nil {
		v := vs.c(&pin)
		cw = &v
	}
	if cull != nil {
		v := cull.c(&pin)
		cc = &v
	}
	return status("nbg_store_put", C.nbg_store_put(s.h, (*C.uint64_t)(ptr(&pin, slots)), C.uint64_t(len(slots)), cs, cv, cw, cc))
}

// Step runs all four predicates over slots [0, n) and returns the results in host memory.
func (s *Store) Step(n int, k VsConsts, p CullParams) (r StepResult, err error) {
	words := (n + 63) / 64
	r = StepResult{StsReason: make([]uint8, n), SvcReason: make([]uint8, n), VsReason: make([]uint8, n),
		StsBitmap: make([]uint64, words), SvcBitmap: make([]uint64, words), VsBitmap: make([]uint64, words),
		Cull: CullResult{make([]uint8, n), make([]uint32, n), make([]uint64, words), make([]uint64, words)}}
	if n == 0 {
		return
	}
	var pin runtime.Pinner
	defer pin.Unpin()
	out := C.nbg_step_outputs{
		sts_reason: (*C.uint8_t)(ptr(&pin, r.StsReason)), sts_bitmap: (*C.uint64_t)(ptr(&pin, r.StsBitmap)),
		svc_reason: (*C.uint8_t)(ptr(&pin, r.SvcReason)), svc_bitmap: (*C.uint64_t)(ptr(&pin, r.SvcBitmap)),
		vs_reason: (*C.uint8_t)(ptr(&pin, r.VsReason)), vs_bitmap: (*C.uint64_t)(ptr(&pin, r.VsBitmap)),
		action: (*C.uint8_t)(ptr(&pin, r.Cull.Action)), argmax: (*C.uint32_t)(ptr(&pin, r.Cull.Argmax)),
		cull_bitmap: (*C.uint64_t)(ptr(&pin, r.Cull.CullBitmap)), period_bitmap: (*C.uint64_t)(ptr(&pin, r.Cull.PeriodBitmap)),
	}
	consts := C.nbg_vs_gen_consts{istio_host_hash: C.uint64_t(k.IstioHostHash), gateway_hash: C.uint64_t(k.GatewayHash), residual: C.uint64_t(k.Residual)}
	params := p.c()
	err = status("nbg_store_step", C.nbg_store_step(s.h, C.uint64_t(n), &consts, &params, &out))
	return
}
