//go:build cuda

package provider

/*
#include <stdlib.h>
#include "llmc.h"
*/
import "C"

import (
	"context"
	"fmt"
	"runtime"
	"sync"
	"sync/atomic"
	"time"
	"unsafe"
)

type LocalConfig struct {
	Shape        int     // C.LLMC_SHAPE_LLAMA_1B / _8B
	Seed         uint64  // --seed (+ the -sN suffix of the model name)
	Device       int     // --device; the runner maps model i to GPU i
	WeightsPath  string  // --weights: .safetensors file or shard directory, HF Llama names ("" = seeded random init)
	MaxNewTokens int
	MaxCtx       int
	MaxSlots     int     // queries in flight on this instance (its entries in --models, +1 if it is the judge); 0 -> 4
	StopIDs      []int32 // generation ends before the first of these ids (128001 / 128009 for Llama-3)
}

// job is one QueryStream call in the hands of the scheduler goroutine.
type job struct {
	slot C.int32_t
	left int           // tokens still to decode
	out  chan []int32  // decoded chunks; closed when left reaches 0 or on error (err set before the close)
	err  error
	quit chan struct{} // closed by the caller when it gives up (ctx done, stop id)
	gone chan struct{} // closed by the scheduler once the job is in no running llmc_decode call
}

type Local struct {
	e     *C.llmc_engine
	cfg   LocalConfig
	vocab int
	slots chan struct{} // counting semaphore over the engine's sequence slots
	join  chan *job
	once  sync.Once

	admitting atomic.Int32 // callers inside prefill(): the scheduler pauses for them between chunks

	cacheMu   sync.Mutex // the prompt cache and the prefill that refreshes it
	cacheSlot C.int32_t  // slot holding the KV cache of the last prompt prefilled here, -1 = none
	cacheIDs  []int32
}

const minPrefix = 32 // shorter common prefixes are not worth a fork

func NewLocal(cfg LocalConfig) (*Local, error) {
	if cfg.MaxSlots <= 0 {
		cfg.MaxSlots = 4
	}
	var c C.llmc_config
	c.device, c.shape, c.seed = C.int32_t(cfg.Device), C.int32_t(cfg.Shape), C.uint64_t(cfg.Seed)
	c.max_slots, c.max_ctx = C.int32_t(cfg.MaxSlots+1), C.int32_t(cfg.MaxCtx) // + the slot that holds the cached prompt
	if cfg.WeightsPath != "" {
		p := C.CString(cfg.WeightsPath)
		defer C.free(unsafe.Pointer(p))
		c.weights_path = p // read during llmc_engine_create only
	}
	var e *C.llmc_engine
	if rc := C.llmc_engine_create(&c, &e); rc != C.LLMC_OK {
		return nil, fmt.Errorf("local provider: %s", C.GoString(C.llmc_last_error(nil)))
	}
	l := &Local{e: e, cfg: cfg, vocab: 128256, slots: make(chan struct{}, cfg.MaxSlots), join: make(chan *job, cfg.MaxSlots), cacheSlot: -1}
	for i := 0; i < cfg.MaxSlots; i++ {
		l.slots <- struct{}{}
	}
	return l, nil
}

func (l *Local) Query(ctx context.Context, req Request) (Response, error) {
	return l.QueryStream(ctx, req, nil) // never called by the CLI path (runner.go:92, judge.go:96)
}

// QueryStream prefills its own sequence slot, then hands the decode loop to the scheduler goroutine, which
// advances every live sequence of this instance together: one llmc_decode(slots[], n, k) per chunk, i.e. one pass
// over the weights per step however many --models entries name this model (runner.go:62 starts one goroutine per
// entry, duplicates included).  Chunks come back over a channel and are emitted here, on the caller's goroutine.
func (l *Local) QueryStream(ctx context.Context, req Request, cb StreamCallback) (Response, error) {
	start := time.Now()
	// Every llmc_* call selects its device itself, so goroutine migration is harmless; locking the
	// thread just avoids re-binding the CUDA context between the calls of one query.
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()

	ids := Tokenize(req.Prompt, l.vocab) // the codec of host/host.cpp, ported 1:1
	if len(ids) == 0 {
		ids = []int32{'\n'}
	}
	if l.cfg.MaxNewTokens <= 0 {
		return Response{Model: req.Model, Provider: "local", Latency: time.Since(start)}, nil
	}
	select { // a sequence slot: wait for one rather than fail when callers outnumber slots
	case <-l.slots:
	case <-ctx.Done():
		return Response{}, ctx.Err()
	}
	defer func() { l.slots <- struct{}{} }()
	// The scheduler holds the engine for a whole chunk per llmc_decode call and would take it again at once; while a
	// caller is being admitted it pauses between chunks, so the newcomer joins at the next chunk boundary.
	l.admitting.Add(1)
	slot, first, err := l.prefill(ids)
	l.admitting.Add(-1)
	if slot >= 0 {
		defer C.llmc_seq_free(l.e, slot)
	}
	if err != nil {
		return Response{}, err
	}
	content, stopped := "", false
	emit := func(t []int32) {
		for i, id := range t {
			for _, s := range l.cfg.StopIDs {
				if id == s && !stopped {
					stopped, t = true, t[:i] // end-of-text: the chunk is cut before it
				}
			}
			if stopped {
				break
			}
		}
		if len(t) == 0 {
			return
		}
		chunk := Detokenize(t)
		content += chunk
		if cb != nil {
			cb(chunk) // synchronously, on the caller's goroutine, like openai.go:192-196
		}
	}
	emit([]int32{int32(first)})
	if left := l.cfg.MaxNewTokens - 1; left > 0 && !stopped {
		const chunk = 16
		j := &job{slot: slot, left: left, out: make(chan []int32, left/chunk+2), quit: make(chan struct{}), gone: make(chan struct{})}
		l.once.Do(func() { go l.schedule(chunk) })
		l.join <- j
		leave := func() { // the deferred llmc_seq_free must not run while a decode call still holds the slot
			close(j.quit)
			<-j.gone
		}
	loop:
		for !stopped {
			select {
			case t, ok := <-j.out:
				if !ok {
					break loop
				}
				emit(t)
			case <-ctx.Done(): // per-model timeout (runner.go:65) and SIGINT (main.go:90)
				leave()
				return Response{}, ctx.Err() // lands in warnings as "<model>: context deadline exceeded" (runner.go:101)
			}
		}
		leave()
		if j.err != nil {
			return Response{}, j.err
		}
	}
	return Response{Model: req.Model, Content: content, Provider: "local", Latency: time.Since(start)}, nil
}

// prefill puts the prompt into a fresh slot.  The instance keeps the KV cache of the LAST prompt it prefilled in a
// slot of its own (llmc_seq_fork: full pages shared, the tail page copied).  The runner sends the same prompt once
// per duplicate --models entry (runner.go:62-95) and every judge prompt starts with the template's fixed head
// (judge.go:12-19), so an identical prompt is a clone of the cached sequence, pending token included (no prefill,
// bit-identical answer), and a prompt sharing >= minPrefix leading tokens forks the prefix and prefills the rest.
func (l *Local) prefill(ids []int32) (C.int32_t, C.int32_t, error) {
	l.cacheMu.Lock()
	defer l.cacheMu.Unlock()
	n, common := len(ids), 0
	if l.cacheSlot >= 0 {
		for common < n && common < len(l.cacheIDs) && ids[common] == l.cacheIDs[common] {
			common++
		}
	}
	slot, first := C.int32_t(-1), C.int32_t(-1)
	if l.cacheSlot >= 0 && common == n && n == len(l.cacheIDs) {
		if C.llmc_seq_fork(l.e, l.cacheSlot, C.int32_t(n), &slot) != C.LLMC_OK {
			return -1, -1, l.err("cloning the cached prompt")
		}
		if C.llmc_seq_get_next_token(l.e, slot, &first) != C.LLMC_OK {
			return slot, -1, l.err("reading the cached prompt's first token")
		}
		return slot, first, nil
	}
	have := 0
	if l.cacheSlot >= 0 && common >= minPrefix {
		have = min(common, n-1) // at least one token goes through the model: its logits pick the first answer token
		if C.llmc_seq_fork(l.e, l.cacheSlot, C.int32_t(have), &slot) != C.LLMC_OK {
			return -1, -1, l.err("forking the cached prefix")
		}
	} else if C.llmc_seq_alloc(l.e, &slot) != C.LLMC_OK {
		return -1, -1, l.err("allocating a sequence slot")
	}
	if C.llmc_prefill(l.e, slot, (*C.int32_t)(unsafe.Pointer(&ids[have])), C.int32_t(n-have), &first, nil) != C.LLMC_OK {
		return slot, -1, l.err("prefill")
	}
	if l.cacheSlot >= 0 { // this prompt becomes the cached one
		C.llmc_seq_free(l.e, l.cacheSlot)
	}
	l.cacheSlot = -1
	if C.llmc_seq_fork(l.e, slot, C.int32_t(n), &l.cacheSlot) == C.LLMC_OK {
		l.cacheIDs = append(l.cacheIDs[:0], ids...)
	} else {
		l.cacheSlot = -1 // no room: carry on without a cache
	}
	return slot, first, nil
}

// schedule runs on its own goroutine for the life of the provider.
func (l *Local) schedule(chunk int) {
	runtime.LockOSThread()
	var live []*job
	for {
		if len(live) == 0 {
			j, ok := <-l.join
			if !ok {
				return
			}
			live = append(live, j)
		}
		for l.admitting.Load() > 0 { // a caller is prefilling: let it have the engine
			time.Sleep(50 * time.Microsecond)
		}
		for more := true; more; { // whoever else has finished its prefill joins at this chunk boundary
			select {
			case j := <-l.join:
				live = append(live, j)
			default:
				more = false
			}
		}
		keep := live[:0]
		for _, j := range live { // callers that gave up leave here, between two decode calls
			select {
			case <-j.quit:
				close(j.gone)
			default:
				keep = append(keep, j)
			}
		}
		live = keep
		if len(live) == 0 {
			continue
		}
		k := chunk
		slots := make([]C.int32_t, len(live))
		for i, j := range live {
			k, slots[i] = min(k, j.left), j.slot
		}
		toks := make([]int32, len(live)*k)
		rc := C.llmc_decode(l.e, &slots[0], C.int32_t(len(live)), C.int32_t(k), (*C.int32_t)(unsafe.Pointer(&toks[0])), nil)
		keep = live[:0]
		for i, j := range live {
			if rc != C.LLMC_OK {
				j.err = l.err("decode")
				j.left = 0
			} else {
				j.out <- toks[i*k : (i+1)*k : (i+1)*k]
				j.left -= k
			}
			if j.left > 0 {
				keep = append(keep, j)
				continue
			}
			close(j.out)
			go func(j *job) { // finished: acknowledge its leave() whenever it comes
				<-j.quit
				close(j.gone)
			}(j)
		}
		live = keep
	}
}

func (l *Local) err(what string) error {
	return fmt.Errorf("local provider: %s: %s", what, C.GoString(C.llmc_last_error(l.e)))
}

func (l *Local) Close() {
	close(l.join)
	C.llmc_engine_destroy(l.e)
}

// LocalDeviceCount is what createProvider uses to deal models to GPUs round-robin.
func LocalDeviceCount() int { return int(C.llmc_device_count()) }
