// Command server is the Go form of the `serve` process (UNCOMPILED here, see package b2).  It keeps the structure of the
// tested C++ stand-in: net/http handlers per connection, ONE engine goroutine locked to its OS thread that admits
// requests into sequence slots, advances all of them with one batched Decode per step, and retires them.
package main

import (
	"encoding/json"
	"log"
	"net/http"
	"os"
	"strconv"
	"strings"
	"time"

	"github.com/nekomeowww/ollama-operator/server/go/b2"
)

type request struct {
	prompt     []int32
	numPredict int
	out        chan string // detokenised pieces; closed when the sequence retires
}

type running struct {
	req   *request
	seq   int
	last  int32
	count int
}

func engineLoop(store string, jobs <-chan *request, model string) {
	tp, _ := strconv.Atoi(os.Getenv("B2_TP_SIZE"))
	if tp < 1 {
		tp = 1
	}
	eng, err := b2.New(tp) // locks this goroutine to its OS thread
	if err != nil {
		log.Fatal(err)
	}
	defer eng.Close()
	ent, err := b2.Resolve(store, model)
	if err != nil {
		log.Fatal(err)
	}
	if err := eng.Load(cGo(ent), 8, 4096, tp); err != nil {
		log.Fatal(err)
	}
	vocab := int(eng.Info.n_vocab)
	logits := make([]float32, vocab)
	var active []*running
	for {
		// admit: block for work only when nothing is running
		var next *request
		if len(active) == 0 {
			next = <-jobs
		} else {
			select {
			case next = <-jobs:
			default:
			}
		}
		if next != nil && len(active) < int(eng.Info.max_batch) {
			seq, err := eng.NewSeq()
			if err == nil && eng.Prefill(seq, next.prompt, logits) == nil {
				first := argmax(logits)
				next.out <- eng.Piece(first)
				active = append(active, &running{req: next, seq: seq, last: first, count: 1})
			} else {
				close(next.out)
			}
			continue
		}
		// step: one batched decode for every running sequence
		seqs, last, nxt := make([]int32, len(active)), make([]int32, len(active)), make([]int32, len(active))
		for i, r := range active {
			seqs[i], last[i] = int32(r.seq), r.last
		}
		if err := eng.Decode(seqs, last, nxt, nil); err != nil {
			log.Print(err)
		}
		keep := active[:0]
		for i, r := range active {
			r.last, r.count = nxt[i], r.count+1
			done := nxt[i] == int32(eng.TInfo.eos) || r.count >= r.req.numPredict
			if !done {
				r.req.out <- eng.Piece(nxt[i])
				keep = append(keep, r)
				continue
			}
			eng.FreeSeq(r.seq) // retire
			close(r.req.out)
		}
		active = keep
	}
}

func argmax(v []float32) int32 {
	best := 0
	for i := range v {
		if v[i] > v[best] {
			best = i
		}
	}
	return int32(best)
}

func main() {
	store := os.Getenv("OLLAMA_MODELS")
	if store == "" {
		store = "/root/.ollama"
	}
	host := os.Getenv("OLLAMA_HOST")
	if host == "" {
		host = "127.0.0.1"
	}
	if !strings.Contains(host, ":") {
		host += ":11434"
	}
	jobs := make(chan *request, 64)
	go engineLoop(store, jobs, os.Getenv("B2_MODEL"))

	http.HandleFunc("/", func(w http.ResponseWriter, r *http.Request) { w.Write([]byte("Ollama is running")) })
	http.HandleFunc("/api/version", func(w http.ResponseWriter, r *http.Request) { json.NewEncoder(w).Encode(map[string]string{"version": "0.1.32-b200"}) })
	// /api/tags is both probes of the pod template: it reads the store listing only (b2_store_list), never the engine.
	http.HandleFunc("/api/tags", func(w http.ResponseWriter, r *http.Request) { w.Write(listStore(store)) })
	http.HandleFunc("/api/generate", func(w http.ResponseWriter, r *http.Request) {
		var body struct {
			Model, Prompt string
			Options       struct{ NumPredict int `json:"num_predict"` }
		}
		if json.NewDecoder(r.Body).Decode(&body) != nil {
			http.Error(w, `{"error":"invalid JSON body"}`, http.StatusBadRequest)
			return
		}
		if body.Options.NumPredict <= 0 {
			body.Options.NumPredict = 128
		}
		req := &request{prompt: tokenize(body.Prompt), numPredict: body.Options.NumPredict, out: make(chan string, 16)}
		jobs <- req
		w.Header().Set("Content-Type", "application/x-ndjson")
		enc, fl := json.NewEncoder(w), w.(http.Flusher)
		for piece := range req.out {
			enc.Encode(map[string]any{"model": body.Model, "created_at": time.Now().UTC(), "response": piece, "done": false})
			fl.Flush()
		}
		enc.Encode(map[string]any{"model": body.Model, "created_at": time.Now().UTC(), "response": "", "done": true})
	})
	log.Fatal(http.ListenAndServe(host, nil))
}
