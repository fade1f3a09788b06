package llamaref

import (
	"math"
	"testing"
)

// Values below were produced by the C++ oracle (oracle/llama_ref.cpp through oracle/ref.py, shape "tiny-untied",
// seed 2, max_ctx 64) -- the oracle that is pinned against HuggingFace LlamaForCausalLM fixtures and that the GPU
// parity tests use.  Run: cd oracle/go && go test ./llamaref

func model() *Model { return New(TinyUntied.WithDefaults(2, 64)) }

func TestWeightsFollowTheRNGSpecification(t *testing.T) {
	m := model()
	cases := []struct {
		layer, kind, n int
		head        [4]uint16
		last        uint16
	}{
		{-1, KEmb, 256000, [4]uint16{48117, 48364, 15554, 15234}, 15326},
		{-1, KLMHead, 256000, [4]uint16{47874, 15376, 15547, 15514}, 15494},
		{-1, KFinalNorm, 256, [4]uint16{16256, 16239, 16244, 16266}, 16264},
		{0, KAttnNorm, 256, [4]uint16{16259, 16266, 16253, 16243}, 16257},
		{1, KWD, 196608, [4]uint16{15509, 48314, 15530, 48021}, 15619},
	}
	for _, c := range cases {
		w := m.Tensor(c.layer, c.kind)
		if len(w) != c.n {
			t.Fatalf("tensor (%d,%d): %d elements, want %d", c.layer, c.kind, len(w), c.n)
		}
		for i, want := range c.head {
			if w[i] != want {
				t.Errorf("tensor (%d,%d)[%d] = %d, want %d", c.layer, c.kind, i, w[i], want)
			}
		}
		if w[len(w)-1] != c.last {
			t.Errorf("tensor (%d,%d) last = %d, want %d", c.layer, c.kind, w[len(w)-1], c.last)
		}
	}
}

func TestGreedyDecodeMatchesTheCppOracle(t *testing.T) {
	m := model()
	prompt := make([]int32, 12)
	for i := range prompt {
		prompt[i] = int32((37*i + 11) % 1000)
	}
	wantTok := []int32{472, 176, 487, 175, 175, 335, 90, 378}
	wantTop := []float32{1.05163, 0.99654, 1.01538, 1.21736, 1.17485, 1.17136, 1.03998, 1.00611}
	toks, logits, err := m.Generate(prompt, len(wantTok))
	if err != nil {
		t.Fatal(err)
	}
	// the smallest top-2 margin on this path is 1.1e-2; the two oracles differ by summation order only (~1e-5)
	for s := range wantTok {
		if toks[s] != wantTok[s] {
			t.Errorf("step %d: token %d, want %d", s, toks[s], wantTok[s])
		}
		if d := math.Abs(float64(logits[s][toks[s]] - wantTop[s])); d > 2e-4 {
			t.Errorf("step %d: top logit %.5f, want %.5f", s, logits[s][toks[s]], wantTop[s])
		}
	}
	if m.Len() != len(prompt)+len(wantTok)-1 {
		t.Errorf("context holds %d tokens", m.Len())
	}
}

func TestChunkedForwardEqualsSinglePass(t *testing.T) {
	a, b := model(), model()
	toks := make([]int32, 20)
	for i := range toks {
		toks[i] = int32((101*i + 7) % 1000)
	}
	la, _ := a.Forward(toks, LastLogits)
	b.Forward(toks[:9], NoLogits)
	lb, _ := b.Forward(toks[9:], LastLogits)
	for i := range la {
		if la[i] != lb[i] { // same operations per position in the same order: bit-identical
			t.Fatalf("logit %d: %g vs %g", i, la[i], lb[i])
		}
	}
}

func TestArgmaxTakesTheLowestIndexOnTies(t *testing.T) {
	if Argmax([]float32{1, 3, 3, 2}) != 1 {
		t.Fatal("tie rule")
	}
}
