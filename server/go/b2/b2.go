// Package b2 binds the C ABI of the B200 GGUF engine (include/b2gguf.h) through cgo.
//
// UNCOMPILED in this repository: the build image has no Go toolchain.  It is the binding a maintainer adds next to the
// operator's `cmd/` to produce the `serve` binary the pod template of pkg/model/pod.go starts; the C++ stand-in that is
// built and tested here (server/b2_server.cpp) makes the same calls in the same order.
package b2

/*
#cgo CFLAGS: -I${SRCDIR}/../../../include
#cgo LDFLAGS: -L${SRCDIR}/../../../ollama-operator_b200 -lb200gguf
#include <stdlib.h>
#include "b2gguf.h"
*/
import "C"

import (
	"errors"
	"runtime"
	"unsafe"
)

func lastError(rc C.int) error {
	if rc >= 0 {
		return nil
	}
	return errors.New(C.GoString(C.b2_strerror()))
}

// Engine owns a b2_ctx and the resident model.  All methods must be called from the goroutine that created it: the
// library allows one OS thread per model, so New locks that goroutine to its thread.
type Engine struct {
	ctx   *C.b2_ctx
	model *C.b2_model
	tok   *C.b2_tokenizer
	Info  C.b2_model_info
	TInfo C.b2_tokenizer_info
}

// New initialises the engine on devices 0..nDev-1 (nDev > 1: tensor parallel).
func New(nDev int) (*Engine, error) {
	runtime.LockOSThread()
	devs := make([]C.int, nDev)
	for i := range devs {
		devs[i] = C.int(i)
	}
	e := &Engine{}
	if err := lastError(C.b2_init(C.int(nDev), &devs[0], &e.ctx)); err != nil {
		return nil, err
	}
	return e, nil
}

// Resolve maps a model name to the blobs of the mounted store (/root/.ollama).
func Resolve(store, name string) (C.b2_store_entry, error) {
	cs, cn := C.CString(store), C.CString(name)
	defer C.free(unsafe.Pointer(cs))
	defer C.free(unsafe.Pointer(cn))
	var ent C.b2_store_entry
	err := lastError(C.b2_store_resolve(cs, cn, &ent))
	return ent, err
}

// Load makes the GGUF file the resident model (weights to HBM, KV cache, tokenizer).
func (e *Engine) Load(path string, maxBatch, maxCtx, tp int) error {
	cp := C.CString(path)
	defer C.free(unsafe.Pointer(cp))
	opts := C.b2_load_opts{max_batch: C.int32_t(maxBatch), max_ctx: C.int32_t(maxCtx), tp_size: C.int32_t(tp)}
	if err := lastError(C.b2_model_load(e.ctx, cp, &opts, &e.model)); err != nil {
		return err
	}
	if err := lastError(C.b2_tokenizer_load(cp, &e.tok)); err != nil {
		return err
	}
	C.b2_model_get_info(e.model, &e.Info)
	C.b2_tokenizer_get_info(e.tok, &e.TInfo)
	return nil
}

// Tokenize turns text into ids (BOS per the file's default, special tokens parsed).
func (e *Engine) Tokenize(text string) []int32 {
	ct := C.CString(text)
	defer C.free(unsafe.Pointer(ct))
	ids := make([]int32, len(text)+16)
	n := int(C.b2_tokenize(e.tok, ct, -1, 1, (*C.int32_t)(unsafe.Pointer(&ids[0])), C.int(len(ids))))
	return ids[:n]
}

// Piece renders one token as text.
func (e *Engine) Piece(tok int32) string {
	buf := make([]byte, 256)
	n := int(C.b2_token_piece(e.tok, C.int32_t(tok), 0, (*C.char)(unsafe.Pointer(&buf[0])), C.int(len(buf))))
	if n <= 0 || n >= len(buf) {
		return ""
	}
	return string(buf[:n])
}

// NewSeq, Prefill, Decode and FreeSeq are the per-request calls of the admit / step / retire loop.
func (e *Engine) NewSeq() (int, error) {
	var s C.int
	err := lastError(C.b2_seq_new(e.model, &s))
	return int(s), err
}

func (e *Engine) Prefill(seq int, toks []int32, logits []float32) error {
	return lastError(C.b2_prefill(e.model, C.int(seq), (*C.int32_t)(unsafe.Pointer(&toks[0])), C.int(len(toks)), (*C.float)(unsafe.Pointer(&logits[0]))))
}

// Decode advances every listed sequence by one token in ONE batched step; next receives the greedy tokens.
func (e *Engine) Decode(seqs []int32, last, next []int32, logits []float32) error {
	var lp *C.float
	if logits != nil {
		lp = (*C.float)(unsafe.Pointer(&logits[0]))
	}
	return lastError(C.b2_decode(e.model, C.int(len(seqs)), (*C.int)(unsafe.Pointer(&seqs[0])), (*C.int32_t)(unsafe.Pointer(&last[0])),
		(*C.int32_t)(unsafe.Pointer(&next[0])), lp))
}

func (e *Engine) FreeSeq(seq int) { C.b2_seq_free(e.model, C.int(seq)) }

// Close releases the model and the context.
func (e *Engine) Close() {
	if e.tok != nil {
		C.b2_tokenizer_free(e.tok)
	}
	if e.model != nil {
		C.b2_model_free(e.model)
	}
	if e.ctx != nil {
		C.b2_shutdown(e.ctx)
	}
}
