// Package nbgraft binds libnbgraft.so (include/nbgraft.h) for the notebook controllers.
//
// NOT COMPILED in the build image (no Go toolchain there); kept in the tree as the binding a maintainer
// would start from — INTEGRATION.md explains where the calls go in the reference.  The C side it binds is
// built and tested (tests/c/abi_client.c drives the same entry points from C).
package nbgraft

/*
#cgo CFLAGS: -I${SRCDIR}/../../include
#cgo LDFLAGS: -L${SRCDIR}/../../odh-kubeflow_b200 -lnbgraft
#include "nbgraft.h"
*/
import "C"

import (
	"fmt"
	"runtime"
	"unsafe"
)

type Context struct{ h *C.nbg_ctx }

func NewContext(devices []int32) (*Context, error) {
	var h *C.nbg_ctx
	var p *C.int
	if len(devices) > 0 {
		p = (*C.int)(unsafe.Pointer(&devices[0]))
	}
	if rc := C.nbg_init(p, C.int(len(devices)), &h); rc != C.NBG_OK {
		return nil, fmt.Errorf("nbg_init: %s", C.GoString(C.nbg_strerror(rc))) // no CPU fallback
	}
	return &Context{h}, nil
}

// StsBatch mirrors nbg_sts_batch with Go slices (flat, pointer-free => legal to pass to C).
type StsBatch struct {
	ReplicasFrom, ReplicasTo []int32
	TmplFrom, TmplTo         []uint64
	PodFrom, PodTo           []uint64 // n * 192
	LabelsFrom, LabelsTo     KvMap
	AnnFrom, AnnTo           KvMap
}
type KvMap struct {
	Off []uint32 // n+1
	Ent []C.nbg_kv
}

// DiffStatefulSets is the batch form of reconcilehelper.CopyStatefulSetFields
// (+ the template-label pre-step of notebook_controller.go:188-194).
func (c *Context) DiffStatefulSets(b *StsBatch) (reason []byte, bitmap []uint64, err error) {
	n := len(b.ReplicasFrom)
	reason, bitmap = make([]byte, n), make([]uint64, (n+63)/64)
	if n == 0 {
		return
	}
	var pin runtime.Pinner // Go 1.21+: keep every slice fixed for the duration of the call
	defer pin.Unpin()
	kv := func(m KvMap) C.nbg_kvmap {
		pin.Pin(&m.Off[0])
		out := C.nbg_kvmap{off: (*C.uint32_t)(unsafe.Pointer(&m.Off[0]))}
		if len(m.Ent) > 0 {
			pin.Pin(&m.Ent[0])
			out.ent = &m.Ent[0]
		}
		return out
	}
	for _, p := range []unsafe.Pointer{unsafe.Pointer(&b.ReplicasFrom[0]), unsafe.Pointer(&b.ReplicasTo[0]),
		unsafe.Pointer(&b.TmplFrom[0]), unsafe.Pointer(&b.TmplTo[0]), unsafe.Pointer(&b.PodFrom[0]), unsafe.Pointer(&b.PodTo[0])} {
		pin.Pin(p)
	}
	in := C.nbg_sts_batch{
		replicas_from: (*C.int32_t)(unsafe.Pointer(&b.ReplicasFrom[0])), replicas_to: (*C.int32_t)(unsafe.Pointer(&b.ReplicasTo[0])),
		tmpl_from: (*C.uint64_t)(unsafe.Pointer(&b.TmplFrom[0])), tmpl_to: (*C.uint64_t)(unsafe.Pointer(&b.TmplTo[0])),
		pod_from: (*C.uint64_t)(unsafe.Pointer(&b.PodFrom[0])), pod_to: (*C.uint64_t)(unsafe.Pointer(&b.PodTo[0])),
		labels_from: kv(b.LabelsFrom), labels_to: kv(b.LabelsTo), ann_from: kv(b.AnnFrom), ann_to: kv(b.AnnTo),
	}
	rc := C.nbg_diff_sts(c.h, &in, C.uint64_t(n), (*C.uint8_t)(unsafe.Pointer(&reason[0])), (*C.uint64_t)(unsafe.Pointer(&bitmap[0])))
	if rc != C.NBG_OK {
		err = fmt.Errorf("nbg_diff_sts: %s", C.GoString(C.nbg_strerror(rc)))
	}
	return
}
