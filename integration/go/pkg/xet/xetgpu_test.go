package xet

import (
	"context"
	"os"
	"path/filepath"
	"testing"
)

// Golden values from hf_xet.hash_files (SURVEY.md Appendix B).
func TestGpuHasherGolden(t *testing.T) {
	h, err := NewGpuHasher(-1)
	if err != nil {
		t.Skipf("no B200 available: %v", err)
	}
	defer h.Close()
	dir := t.TempDir()
	cases := map[string]struct {
		data []byte
		hash string
	}{
		"empty":       {nil, "0000000000000000000000000000000000000000000000000000000000000000"},
		"hello_world": {[]byte("hello world"), "f15edd43b7e85cfd25e2758ffe13b99fdd09dfa817e8d9f6548722a7a126db7f"},
		"zeros_8192":  {make([]byte, 8192), "711574865581cce65f5d06a1818a37a1dd4cfe3f65e3f4aaae2b1bacbfc253db"},
		"zeros_1MiB":  {make([]byte, 1<<20), "1e671fe124cea35586b1d1c30b9d4fc6b4e05ee60c93406986444f7c23d54056"},
	}
	for name, c := range cases {
		p := filepath.Join(dir, name)
		if err := os.WriteFile(p, c.data, 0o600); err != nil {
			t.Fatal(err)
		}
		got, err := h.HashFiles(context.Background(), []string{p})
		if err != nil {
			t.Fatalf("%s: %v", name, err)
		}
		if got[0].Hash != c.hash {
			t.Errorf("%s: got %s want %s", name, got[0].Hash, c.hash)
		}
		if err := h.VerifyFile(p, c.hash); err != nil {
			t.Errorf("%s: verify: %v", name, err)
		}
		if err := h.VerifyFile(p, "ff"+c.hash[2:]); !IsChecksumMismatch(err) {
			t.Errorf("%s: expected checksum mismatch, got %v", name, err)
		}
	}
}

// A cancelled context must reach the library: the bridge reused from xet.go type-asserts *cancellationBridge
// (xet.go:68-85), which is what cancelToken wraps the context in.
func TestGpuHasherCancellation(t *testing.T) {
	h, err := NewGpuHasher(-1)
	if err != nil {
		t.Skipf("no B200 available: %v", err)
	}
	defer h.Close()
	p := filepath.Join(t.TempDir(), "big")
	if err := os.WriteFile(p, make([]byte, 64<<20), 0o600); err != nil {
		t.Fatal(err)
	}
	ctx, cancel := context.WithCancel(context.Background())
	cancel()
	if _, err := h.HashFiles(ctx, []string{p}); err == nil {
		t.Fatal("expected a cancellation error")
	} else if xe, ok := err.(*XetError); !ok || xe.Code != 7 {
		t.Fatalf("expected XetError code 7 (Cancelled), got %v", err)
	}
	if _, err := h.VerifyFiles(ctx, []string{p}, []string{"00"}); err == nil {
		t.Fatal("expected an error")
	}
}

// The replica agent's GetFileChecksum shape for algorithm "xet".
func TestXetFileChecksum(t *testing.T) {
	h, err := NewGpuHasher(-1)
	if err != nil {
		t.Skipf("no B200 available: %v", err)
	}
	defer h.Close()
	p := filepath.Join(t.TempDir(), "hello")
	if err := os.WriteFile(p, []byte("hello world"), 0o600); err != nil {
		t.Fatal(err)
	}
	got, err := h.XetFileChecksum(p)
	if err != nil || got != "f15edd43b7e85cfd25e2758ffe13b99fdd09dfa817e8d9f6548722a7a126db7f" {
		t.Fatalf("got %q, %v", got, err)
	}
	if XetChecksumAlgorithm != "xet" || XetChecksumMetadataKey != "opc-meta-xet" {
		t.Fatal("constants changed")
	}
}
