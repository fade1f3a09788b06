// Package llamaref is the Go twin of oracle/llama_ref.cpp: a pure-Go fp32 CPU forward of the random-init
// Llama-shaped model that the local provider runs on the GPU -- the form of the oracle that BASELINE.json's
// north_star names ("a pure-Go fp32 CPU forward of the same random-init architecture").
//
// TEST INFRASTRUCTURE ONLY: nothing under llm-consensus_b200/ imports it.  It exists so that a maintainer of
// johnayoung/llm-consensus can check the cgo provider (internal/provider/local.go, INTEGRATION.md) against an
// oracle inside `go test`, with no Python or C++ in the loop.
//
// STATUS: SOURCE ONLY.  There is no Go toolchain in the image this repository was built in, so this package has not
// been compiled or run there; it restates oracle/llama_ref.cpp function by function (same RNG specification, same
// operation order per element except where noted) and its test holds values produced by that C++ oracle
// (llamaref_test.go).  The boundary the forward sits behind is provider.Provider.QueryStream
// (/root/reference/internal/provider/provider.go:13-21); the reference itself contains no inference arithmetic.
//
// Architecture (Llama-3 family):
//
//	x = E[tok]
//	per layer: h = rmsnorm(x, g1); q,k,v = Wq h, Wk h, Wv h; rope(q,k) rotate-half; causal GQA attention with
//	           fp32 softmax; x += Wo a; h = rmsnorm(x, g2); x += Wd (silu(Wg h) * (Wu h))
//	logits = LMHead rmsnorm(x, gf)   (LMHead = E when tied); greedy = lowest-index argmax
//
// Weights are the bf16-rounded values of a counter-based RNG, widened to fp32; activations and sums are fp32.
package llamaref

import (
	"errors"
	"math"
)

// Config mirrors oref_config (oracle/llama_ref.cpp).
type Config struct {
	Hidden, Layers, QHeads, KVHeads, HeadDim, Inter, Vocab int
	Tied                                                   bool
	RopeTheta, RMSEps, InitStd                             float32
	Seed                                                   uint64
	MaxCtx                                                 int
	// RoPE scaling as in Hugging Face's rope_type "llama3" (0 factor = off)
	RopeFactor, RopeLowFreqFactor, RopeHighFreqFactor float32
	RopeOriginalCtx                                   int
}

// Shapes named by BASELINE.json's configs and the small test shapes of oracle/ref.py.
var (
	Llama1B    = Config{Hidden: 2048, Layers: 16, QHeads: 32, KVHeads: 8, HeadDim: 64, Inter: 8192, Vocab: 128256, Tied: true}
	Llama8B    = Config{Hidden: 4096, Layers: 32, QHeads: 32, KVHeads: 8, HeadDim: 128, Inter: 14336, Vocab: 128256}
	Tiny       = Config{Hidden: 256, Layers: 2, QHeads: 8, KVHeads: 2, HeadDim: 32, Inter: 512, Vocab: 1024, Tied: true}
	TinyUntied = Config{Hidden: 256, Layers: 2, QHeads: 4, KVHeads: 4, HeadDim: 64, Inter: 768, Vocab: 1000}
)

// WithDefaults fills the numeric defaults of oracle/ref.py's make_config.
func (c Config) WithDefaults(seed uint64, maxCtx int) Config {
	c.RopeTheta, c.RMSEps, c.InitStd, c.Seed, c.MaxCtx = 500000, 1e-5, 0.02, seed, maxCtx
	return c
}

// ---------------------------------------------------------------- RNG specification (shared with csrc/init.cu)
//
//	key  = mix64(seed ^ (GOLDEN * (tensor_id + 1)))
//	z    = mix64(key + GOLDEN * (index + 1))
//	u    = (int32(z >> 40) - 2^23) * 2^-23            in [-1, 1), exact in fp32
//	W[i] = bf16_rne(u * amp), amp = float32(init_std * sqrt(3))     (matrices)
//	g[i] = bf16_rne(1 + 0.1f * u)  (two separately rounded fp32 operations)   (norm gains)
//
// tensor_id = (layer + 1) * 16 + kind; globals use layer = -1.
const golden = 0x9E3779B97F4A7C15

func mix64(z uint64) uint64 {
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9
	z = (z ^ (z >> 27)) * 0x94D049BB133111EB
	return z ^ (z >> 31)
}

func rngKey(seed uint64, tensorID uint32) uint64 {
	return mix64(seed ^ (golden * uint64(tensorID+1)))
}

func rngUnit(key, idx uint64) float32 {
	z := mix64(key + golden*(idx+1))
	v := int32(z>>40) - (1 << 23)
	return float32(v) * (1.0 / 8388608.0)
}

// BF16RNE rounds an fp32 value to bf16 bits, ties to even.
func BF16RNE(f float32) uint16 {
	b := math.Float32bits(f)
	if b&0x7fffffff > 0x7f800000 {
		return uint16(b>>16) | 0x40 // NaN
	}
	b += 0x7fff + ((b >> 16) & 1)
	return uint16(b >> 16)
}

// BF16ToF32 widens bf16 bits.
func BF16ToF32(h uint16) float32 { return math.Float32frombits(uint32(h) << 16) }

// tensor kinds
const (
	KEmb = iota
	KLMHead
	KFinalNorm
)
const (
	KAttnNorm = iota
	KWQ
	KWK
	KWV
	KWO
	KMLPNorm
	KWG
	KWU
	KWD
)

func tensorID(layer, kind int) uint32 { return uint32((layer+1)*16 + kind) }

func fillMatrix(n int, seed uint64, tid uint32, amp float32) []uint16 {
	w := make([]uint16, n)
	key := rngKey(seed, tid)
	for i := range w {
		w[i] = BF16RNE(rngUnit(key, uint64(i)) * amp)
	}
	return w
}

func fillGain(n int, seed uint64, tid uint32) []uint16 {
	w := make([]uint16, n)
	key := rngKey(seed, tid)
	for i := range w {
		p := float32(0.1 * rngUnit(key, uint64(i))) // explicit conversion: a rounded product, never fused into the add
		w[i] = BF16RNE(float32(1.0 + p))
	}
	return w
}

type layer struct {
	g1, wq, wk, wv, wo, g2, wg, wu, wd []uint16
}

// Model holds the weights, the dense fp32 KV cache [layer][pos][kv_heads*head_dim] and the RoPE tables.
type Model struct {
	C                Config
	emb, lmhead, gf  []uint16
	l                []layer
	kc, vc           []float32
	ropeCos, ropeSin []float32
	n                int // tokens in the context
}

// New builds the model of cfg (oref_create).
func New(cfg Config) *Model {
	c := cfg
	H, D := c.Hidden, c.HeadDim
	QD, KD, I := c.QHeads*D, c.KVHeads*D, c.Inter
	amp := float32(float64(c.InitStd) * 1.7320508075688772)
	m := &Model{C: c}
	m.emb = fillMatrix(c.Vocab*H, c.Seed, tensorID(-1, KEmb), amp)
	if !c.Tied {
		m.lmhead = fillMatrix(c.Vocab*H, c.Seed, tensorID(-1, KLMHead), amp)
	}
	m.gf = fillGain(H, c.Seed, tensorID(-1, KFinalNorm))
	m.l = make([]layer, c.Layers)
	for i := range m.l {
		y := &m.l[i]
		y.g1 = fillGain(H, c.Seed, tensorID(i, KAttnNorm))
		y.wq = fillMatrix(QD*H, c.Seed, tensorID(i, KWQ), amp)
		y.wk = fillMatrix(KD*H, c.Seed, tensorID(i, KWK), amp)
		y.wv = fillMatrix(KD*H, c.Seed, tensorID(i, KWV), amp)
		y.wo = fillMatrix(H*QD, c.Seed, tensorID(i, KWO), amp)
		y.g2 = fillGain(H, c.Seed, tensorID(i, KMLPNorm))
		y.wg = fillMatrix(I*H, c.Seed, tensorID(i, KWG), amp)
		y.wu = fillMatrix(I*H, c.Seed, tensorID(i, KWU), amp)
		y.wd = fillMatrix(H*I, c.Seed, tensorID(i, KWD), amp)
	}
	m.kc = make([]float32, c.Layers*c.MaxCtx*KD)
	m.vc = make([]float32, c.Layers*c.MaxCtx*KD)
	m.ropeCos = make([]float32, c.MaxCtx*(D/2))
	m.ropeSin = make([]float32, c.MaxCtx*(D/2))
	for p := 0; p < c.MaxCtx; p++ {
		for i := 0; i < D/2; i++ {
			inv := math.Pow(float64(c.RopeTheta), -2.0*float64(i)/float64(D))
			if c.RopeFactor != 0 {
				// wavelengths longer than original_ctx / low_freq_factor are stretched by factor, those shorter than
				// original_ctx / high_freq_factor are kept, the band between is interpolated (HF _compute_llama3_parameters)
				wavelen, oc := 2*math.Pi/inv, float64(c.RopeOriginalCtx)
				lo, hi := oc/float64(c.RopeLowFreqFactor), oc/float64(c.RopeHighFreqFactor)
				if wavelen > lo {
					inv /= float64(c.RopeFactor)
				} else if !(wavelen < hi) {
					smooth := (oc/wavelen - float64(c.RopeLowFreqFactor)) / (float64(c.RopeHighFreqFactor) - float64(c.RopeLowFreqFactor))
					inv = (1-smooth)*inv/float64(c.RopeFactor) + smooth*inv
				}
			}
			a := float64(p) * inv
			m.ropeCos[p*(D/2)+i] = float32(math.Cos(a))
			m.ropeSin[p*(D/2)+i] = float32(math.Sin(a))
		}
	}
	return m
}

// Reset empties the context; Len is the number of cached tokens.
func (m *Model) Reset()   { m.n = 0 }
func (m *Model) Len() int { return m.n }

// Tensor returns the bf16 bits of a weight tensor (layer -1: globals), for pinning the GPU initialiser.
func (m *Model) Tensor(layer, kind int) []uint16 {
	if layer < 0 {
		switch kind {
		case KEmb:
			return m.emb
		case KLMHead:
			if m.C.Tied {
				return m.emb
			}
			return m.lmhead
		case KFinalNorm:
			return m.gf
		}
		return nil
	}
	if layer >= len(m.l) || kind < 0 || kind > KWD {
		return nil
	}
	y := &m.l[layer]
	return [][]uint16{y.g1, y.wq, y.wk, y.wv, y.wo, y.g2, y.wg, y.wu, y.wd}[kind]
}

// linear: Y[n][N] = X[n][K] * W[N][K]^T with W widened from bf16, fp32 accumulation in index order.
// (The C++ oracle lets the compiler vectorise this sum, so the two agree to rounding, not bit for bit.)
func linear(x []float32, n, K int, w []uint16, N int, y []float32, ldy int) {
	wrow := make([]float32, K)
	for j := 0; j < N; j++ {
		wr := w[j*K : (j+1)*K]
		for k := range wrow {
			wrow[k] = BF16ToF32(wr[k])
		}
		for t := 0; t < n; t++ {
			xr := x[t*K : (t+1)*K]
			var acc float32
			for k := 0; k < K; k++ {
				acc += float32(xr[k] * wrow[k])
			}
			y[t*ldy+j] = acc
		}
	}
}

func rmsnorm(x []float32, g []uint16, eps float32, y []float32) {
	var ss float32
	for _, v := range x {
		ss += float32(v * v)
	}
	r := 1 / float32(math.Sqrt(float64(ss/float32(len(x))+eps)))
	for i, v := range x {
		y[i] = float32(float32(v*r) * BF16ToF32(g[i]))
	}
}

// Logits modes of Forward.
const (
	NoLogits = iota
	LastLogits
	AllLogits
)

// Forward appends tokens to the context and runs them through the model (oref_forward).
// mode LastLogits returns [vocab], AllLogits [len(tokens)*vocab].
func (m *Model) Forward(tokens []int32, mode int) ([]float32, error) {
	c := m.C
	H, D, NQ, NKV := c.Hidden, c.HeadDim, c.QHeads, c.KVHeads
	QD, KD, I, V, G := NQ*D, NKV*D, c.Inter, c.Vocab, NQ/NKV
	n := len(tokens)
	if n == 0 {
		return nil, nil
	}
	if m.n+n > c.MaxCtx {
		return nil, errors.New("llamaref: context overflow")
	}
	for _, t := range tokens {
		if t < 0 || int(t) >= V {
			return nil, errors.New("llamaref: token id out of range")
		}
	}
	p0 := m.n
	x, hn := make([]float32, n*H), make([]float32, n*H)
	q, att := make([]float32, n*QD), make([]float32, n*QD)
	gate, up, tmp := make([]float32, n*I), make([]float32, n*I), make([]float32, n*H)
	for t, tok := range tokens {
		for i := 0; i < H; i++ {
			x[t*H+i] = BF16ToF32(m.emb[int(tok)*H+i])
		}
	}
	scale := 1 / float32(math.Sqrt(float64(D)))
	for l := range m.l {
		y := &m.l[l]
		kc := m.kc[l*c.MaxCtx*KD : (l+1)*c.MaxCtx*KD]
		vc := m.vc[l*c.MaxCtx*KD : (l+1)*c.MaxCtx*KD]
		for t := 0; t < n; t++ {
			rmsnorm(x[t*H:(t+1)*H], y.g1, c.RMSEps, hn[t*H:(t+1)*H])
		}
		linear(hn, n, H, y.wq, QD, q, QD)
		linear(hn, n, H, y.wk, KD, kc[p0*KD:], KD)
		linear(hn, n, H, y.wv, KD, vc[p0*KD:], KD)
		for t := 0; t < n; t++ { // RoPE (rotate-half) on q and the new k rows
			cs := m.ropeCos[(p0+t)*(D/2) : (p0+t+1)*(D/2)]
			sn := m.ropeSin[(p0+t)*(D/2) : (p0+t+1)*(D/2)]
			for hh := 0; hh < NQ+NKV; hh++ {
				var v []float32
				if hh < NQ {
					v = q[t*QD+hh*D : t*QD+(hh+1)*D]
				} else {
					v = kc[(p0+t)*KD+(hh-NQ)*D : (p0+t)*KD+(hh-NQ+1)*D]
				}
				for i := 0; i < D/2; i++ {
					a, b := v[i], v[i+D/2]
					v[i] = float32(a*cs[i]) - float32(b*sn[i])
					v[i+D/2] = float32(b*cs[i]) + float32(a*sn[i])
				}
			}
		}
		for t := 0; t < n; t++ { // causal GQA attention, fp32 softmax
			for hh := 0; hh < NQ; hh++ {
				ctx, kvh := p0+t+1, hh/G
				qv := q[t*QD+hh*D : t*QD+(hh+1)*D]
				s := make([]float32, ctx)
				mx := float32(math.Inf(-1))
				for p := 0; p < ctx; p++ {
					kr := kc[p*KD+kvh*D : p*KD+(kvh+1)*D]
					var acc float32
					for i := 0; i < D; i++ {
						acc += float32(qv[i] * kr[i])
					}
					s[p] = acc * scale
					if s[p] > mx {
						mx = s[p]
					}
				}
				var den float32
				for p := range s {
					s[p] = float32(math.Exp(float64(s[p] - mx)))
					den += s[p]
				}
				o := att[t*QD+hh*D : t*QD+(hh+1)*D]
				for i := range o {
					o[i] = 0
				}
				for p := 0; p < ctx; p++ {
					vr := vc[p*KD+kvh*D : p*KD+(kvh+1)*D]
					w := s[p] / den
					for i := 0; i < D; i++ {
						o[i] += float32(w * vr[i])
					}
				}
			}
		}
		linear(att, n, QD, y.wo, H, tmp, H)
		for i := range x {
			x[i] += tmp[i]
		}
		for t := 0; t < n; t++ {
			rmsnorm(x[t*H:(t+1)*H], y.g2, c.RMSEps, hn[t*H:(t+1)*H])
		}
		linear(hn, n, H, y.wg, I, gate, I)
		linear(hn, n, H, y.wu, I, up, I)
		for i := range gate {
			g := gate[i]
			gate[i] = float32(g/(1+float32(math.Exp(float64(-g))))) * up[i]
		}
		linear(gate, n, I, y.wd, H, tmp, H)
		for i := range x {
			x[i] += tmp[i]
		}
	}
	m.n += n
	if mode == NoLogits {
		return nil, nil
	}
	head := m.lmhead
	if c.Tied {
		head = m.emb
	}
	if mode == LastLogits {
		rmsnorm(x[(n-1)*H:n*H], m.gf, c.RMSEps, hn[:H])
		out := make([]float32, V)
		linear(hn[:H], 1, H, head, V, out, V)
		return out, nil
	}
	for t := 0; t < n; t++ {
		rmsnorm(x[t*H:(t+1)*H], m.gf, c.RMSEps, hn[t*H:(t+1)*H])
	}
	out := make([]float32, n*V)
	linear(hn, n, H, head, V, out, V)
	return out, nil
}

// Argmax is the greedy rule: the lowest index wins ties.
func Argmax(v []float32) int32 {
	best := 0
	for i := 1; i < len(v); i++ {
		if v[i] > v[best] {
			best = i
		}
	}
	return int32(best)
}

// Generate prefills prompt, then greedy-decodes nNew tokens; logits[s] is the row token s was the argmax of.
func (m *Model) Generate(prompt []int32, nNew int) (tokens []int32, logits [][]float32, err error) {
	lg, err := m.Forward(prompt, LastLogits)
	if err != nil {
		return nil, nil, err
	}
	for s := 0; s < nNew; s++ {
		t := Argmax(lg)
		tokens = append(tokens, t)
		logits = append(logits, lg)
		if s+1 < nNew {
			if lg, err = m.Forward([]int32{t}, LastLogits); err != nil {
				return nil, nil, err
			}
		}
	}
	return tokens, logits, nil
}
