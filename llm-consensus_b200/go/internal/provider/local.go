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
	"time"
	"unsafe"
)

type LocalConfig struct {
	Shape        int    // C.LLMC_SHAPE_LLAMA_1B / _8B
	Seed         uint64 // --seed (+ the -sN suffix of the model name)
	Device       int    // --device; the runner maps model i to GPU i
	WeightsPath  string // --weights: .safetensors file or shard directory, HF Llama names ("" = seeded random init)
	MaxNewTokens int
	MaxCtx       int
}

type Local struct {
	e     *C.llmc_engine
	cfg   LocalConfig
	vocab int
}

func NewLocal(cfg LocalConfig) (*Local, error) {
	var c C.llmc_config
	c.device, c.shape, c.seed = C.int32_t(cfg.Device), C.int32_t(cfg.Shape), C.uint64_t(cfg.Seed)
	c.max_slots, c.max_ctx = 2, C.int32_t(cfg.MaxCtx)
	if cfg.WeightsPath != "" {
		p := C.CString(cfg.WeightsPath)
		defer C.free(unsafe.Pointer(p))
		c.weights_path = p // read during llmc_engine_create only
	}
	var e *C.llmc_engine
	if rc := C.llmc_engine_create(&c, &e); rc != C.LLMC_OK {
		return nil, fmt.Errorf("local provider: %s", C.GoString(C.llmc_last_error(nil)))
	}
	return &Local{e: e, cfg: cfg, vocab: 128256}, nil
}

func (l *Local) Query(ctx context.Context, req Request) (Response, error) {
	return l.QueryStream(ctx, req, nil) // never called by the CLI path (runner.go:92, judge.go:96)
}

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
	var slot C.int32_t
	if C.llmc_seq_alloc(l.e, &slot) != C.LLMC_OK {
		return Response{}, l.err("allocating a sequence slot")
	}
	defer C.llmc_seq_free(l.e, slot)

	var first C.int32_t
	if C.llmc_prefill(l.e, slot, (*C.int32_t)(unsafe.Pointer(&ids[0])), C.int32_t(len(ids)), &first, nil) != C.LLMC_OK {
		return Response{}, l.err("prefill")
	}
	content := ""
	emit := func(t []int32) {
		chunk := Detokenize(t)
		content += chunk
		if cb != nil {
			cb(chunk) // synchronously, on the caller's goroutine, like openai.go:192-196
		}
	}
	emit([]int32{int32(first)})
	const chunk = 16
	toks := make([]int32, chunk)
	for left := l.cfg.MaxNewTokens - 1; left > 0; {
		if err := ctx.Err(); err != nil { // per-model timeout (runner.go:65) and SIGINT (main.go:90)
			return Response{}, err // lands in warnings as "<model>: context deadline exceeded" (runner.go:101)
		}
		k := min(left, chunk)
		if C.llmc_decode(l.e, &slot, 1, C.int32_t(k), (*C.int32_t)(unsafe.Pointer(&toks[0])), nil) != C.LLMC_OK {
			return Response{}, l.err("decode")
		}
		emit(toks[:k])
		left -= k
	}
	return Response{Model: req.Model, Content: content, Provider: "local", Latency: time.Since(start)}, nil
}

func (l *Local) err(what string) error {
	return fmt.Errorf("local provider: %s: %s", what, C.GoString(C.llmc_last_error(l.e)))
}

func (l *Local) Close() { C.llmc_engine_destroy(l.e) }

// LocalDeviceCount is what createProvider uses to deal models to GPUs round-robin.
func LocalDeviceCount() int { return int(C.llmc_device_count()) }
