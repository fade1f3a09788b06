package provider

// Token codec of the local backend (same rules as llm-consensus_b200/host/host.cpp, which the tests exercise):
// a token id that is '\n' or printable ASCII other than '[' is written as that byte, every other id as "[[id]]"
// (decimal, no leading zeros); Tokenize inverts Detokenize exactly and maps any other byte to its own value.
// It stands in for a real tokenizer: the random-init models of the parity configs have no vocabulary file.

import "strconv"

func plainID(t int32) bool { return t == 0x0A || (t >= 0x20 && t <= 0x7E && t != '[') }

// Detokenize renders token ids as text.
func Detokenize(ids []int32) string {
	b := make([]byte, 0, len(ids))
	for _, t := range ids {
		if plainID(t) {
			b = append(b, byte(t))
		} else {
			b = append(b, '[', '[')
			b = strconv.AppendInt(b, int64(t), 10)
			b = append(b, ']', ']')
		}
	}
	return string(b)
}

// Tokenize maps text to token ids below vocab.
func Tokenize(text string, vocab int) []int32 {
	b := []byte(text)
	n := len(b)
	ids := make([]int32, 0, n)
	for i := 0; i < n; {
		if b[i] == '[' && i+1 < n && b[i+1] == '[' {
			j := i + 2
			for j < n && j-(i+2) < 7 && b[j] >= '0' && b[j] <= '9' {
				j++
			}
			nd := j - (i + 2)
			if nd >= 1 && j+1 < n && b[j] == ']' && b[j+1] == ']' {
				v := 0
				for k := i + 2; k < j; k++ {
					v = v*10 + int(b[k]-'0')
				}
				// canonical escapes only: no leading zero, inside the vocabulary, not an id that prints as itself
				if (nd == 1 || b[i+2] != '0') && v < vocab && !plainID(int32(v)) {
					ids = append(ids, int32(v))
					i = j + 2
					continue
				}
			}
		}
		ids = append(ids, int32(b[i]))
		i++
	}
	return ids
}
