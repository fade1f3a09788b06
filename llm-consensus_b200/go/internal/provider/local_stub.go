//go:build !cuda

package provider

import (
	"context"
	"errors"
)

// LocalConfig mirrors the cuda build's type so that cmd/llm-consensus compiles without cgo.
type LocalConfig struct {
	Shape        int
	Seed         uint64
	Device       int
	WeightsPath  string
	MaxNewTokens int
	MaxCtx       int
}

// Local is never constructed in this build.
type Local struct{}

// NewLocal fails in the default (CGO_ENABLED=0) build; local models need `-tags cuda`.
func NewLocal(LocalConfig) (*Local, error) {
	return nil, errors.New("local provider: built without cuda (rebuild with CGO_ENABLED=1 -tags cuda)")
}

// Query and QueryStream keep *Local a Provider in this build as well.
func (*Local) Query(context.Context, Request) (Response, error) {
	return Response{}, errors.New("local provider: built without cuda")
}

func (*Local) QueryStream(context.Context, Request, StreamCallback) (Response, error) {
	return Response{}, errors.New("local provider: built without cuda")
}

// Close is a no-op here.
func (*Local) Close() {}

// LocalDeviceCount reports no GPUs in this build.
func LocalDeviceCount() int { return 0 }
