"""Model-store reader and GGUF tokenizer: what sits between the volume the operator mounts at
/root/.ollama (reference pkg/model/pod.go:34-40, filled by the puller of :68-83) and the token ids of b2_prefill.

The tokenizers are checked against the libraries the formats come from -- `sentencepiece` (BPE model type, byte
fallback: the Llama-2 / Mistral kind) and HuggingFace `tokenizers` (byte-level BPE: the GPT-2 / phi-2 kind) -- on
vocabularies trained inside the test, carried through a GGUF file written with the `gguf` package.  No GPU needed."""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

CORPUS = """The operator creates a Deployment and a Service for every Model resource. Each pod mounts the model
store read-only at /root/.ollama and serves the Ollama HTTP API on port 11434. Hello, world! It's 2024: don't panic.
def quantize_row(x): return [round(v / d) for v in x]   # q8_0 blocks of 32 weights, scale = amax / 127
Straße, naïve café, coöperate; 数字 and emoji 🙂 are rare here. tokens/s at batch 1, 512-token prefill, 7B Q4_0.
"""
SAMPLES = [
    "Hello, world!", " leading space", "The model store is read-only.", "don't you've we'll I'm", "x = [round(v / d) for v in x]",
    "multiple   spaces\tand\ttabs\nnewlines\n\n", "café naïve Straße", "emoji 🙂 and 数字", "1234567890 3.14159", "", "a", "trailing space ",
]


def _lib(pkg):
    L = pkg.engine.lib()
    L.b2_tokenizer_load.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    L.b2_tokenizer_free.argtypes = [C.c_void_p]
    L.b2_tokenizer_free.restype = None
    L.b2_tokenize.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
    L.b2_detokenize.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_char_p, C.c_int]
    L.b2_store_resolve.argtypes = [C.c_char_p, C.c_char_p, C.c_void_p]
    L.b2_store_list.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
    return L


class Tok:
    def __init__(self, pkg, path):
        self.L = _lib(pkg)
        self.h = C.c_void_p()
        rc = self.L.b2_tokenizer_load(os.fsencode(path), C.byref(self.h))
        assert rc == 0, self.L.b2_strerror()

    def encode(self, text, add_bos=0, special=0):
        buf = (C.c_int32 * 4096)()
        n = self.L.b2_tokenize(self.h, text.encode(), add_bos, special, buf, 4096)
        assert 0 <= n <= 4096, n
        return list(buf[:n])

    def decode(self, ids):
        arr = (C.c_int32 * max(1, len(ids)))(*ids)
        out = C.create_string_buffer(65536)
        n = self.L.b2_detokenize(self.h, arr, len(ids), 0, out, 65536)
        assert n >= 0
        return out.raw[:n].decode("utf-8", errors="replace")

    def close(self):
        self.L.b2_tokenizer_free(self.h)


def _minimal_gguf(path, arch, add_tokenizer):
    import gguf
    w = gguf.GGUFWriter(path, arch)
    w.add_block_count(1)
    add_tokenizer(w)
    w.add_tensor("token_embd.weight", np.zeros((4, 32), np.float32))
    w.write_header_to_file(); w.write_kv_data_to_file(); w.write_tensors_to_file(); w.close()


def test_sentencepiece_style_tokenizer_matches_sentencepiece(pkg, tmp_path):
    import sentencepiece as spm
    corpus = tmp_path / "corpus.txt"
    corpus.write_text(CORPUS * 4)
    prefix = str(tmp_path / "sp")
    spm.SentencePieceTrainer.train(input=str(corpus), model_prefix=prefix, vocab_size=400, model_type="bpe", byte_fallback=True,
                                   character_coverage=0.98, split_digits=True, normalization_rule_name="identity", add_dummy_prefix=True,
                                   remove_extra_whitespaces=False, allow_whitespace_only_pieces=True, unk_id=0, bos_id=1, eos_id=2, pad_id=-1)
    sp = spm.SentencePieceProcessor(model_file=prefix + ".model")
    toks, scores, types = [], [], []
    for i in range(sp.get_piece_size()):
        toks.append(sp.id_to_piece(i)); scores.append(sp.get_score(i))
        types.append(2 if sp.is_unknown(i) else 3 if sp.is_control(i) else 6 if sp.is_byte(i) else 1)
    path = str(tmp_path / "spm.gguf")

    def add(w):
        w.add_tokenizer_model("llama"); w.add_token_list(toks); w.add_token_scores(scores); w.add_token_types(types)
        w.add_bos_token_id(1); w.add_eos_token_id(2); w.add_unk_token_id(0); w.add_add_bos_token(True)
    _minimal_gguf(path, "llama", add)
    t = Tok(pkg, path)
    for s in SAMPLES:
        assert t.encode(s) == sp.encode(s), s
        if s:
            assert t.decode(t.encode(s)) == sp.decode(sp.encode(s)), s
    assert t.encode("Hello", add_bos=-1)[0] == 1                     # the file's add_bos default
    assert t.encode("a<s>b", special=1).count(1) == 1 and 1 not in t.encode("a<s>b", special=0)
    t.close()


def test_byte_level_bpe_tokenizer_matches_hf_tokenizers(pkg, tmp_path):
    from tokenizers import Tokenizer, decoders, models, pre_tokenizers, trainers
    tk = Tokenizer(models.BPE())
    tk.pre_tokenizer = pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=True)
    tk.decoder = decoders.ByteLevel()
    tr = trainers.BpeTrainer(vocab_size=420, special_tokens=["<|endoftext|>"], initial_alphabet=pre_tokenizers.ByteLevel.alphabet())
    tk.train_from_iterator([CORPUS] * 4, tr)
    vocab = sorted(tk.get_vocab().items(), key=lambda kv: kv[1])
    toks = [k for k, _ in vocab]
    merges = [" ".join(m) if not isinstance(m, str) else m for m in json.loads(tk.to_str())["model"]["merges"]]
    types = [3 if k == "<|endoftext|>" else 1 for k in toks]
    path = str(tmp_path / "bpe.gguf")

    def add(w):
        w.add_tokenizer_model("gpt2"); w.add_token_list(toks); w.add_token_types(types); w.add_token_merges(merges)
        w.add_bos_token_id(0); w.add_eos_token_id(0); w.add_add_bos_token(False)
    _minimal_gguf(path, "phi2", add)
    t = Tok(pkg, path)
    ascii_samples = [s for s in SAMPLES if all(ord(c) < 128 for c in s)]      # the pre-tokeniser treats non-ASCII bytes as letters
    for s in ascii_samples:
        assert t.encode(s) == tk.encode(s).ids, s
        assert t.decode(t.encode(s)) == s, s
    for s in SAMPLES:                                                          # any text survives the round trip byte for byte
        assert t.decode(t.encode(s)) == s, s
    assert t.encode("a<|endoftext|>b", special=1).count(0) == 1
    t.close()


def _write_store(root, name, tag, gguf_bytes, template="{{ .Prompt }}", params=None, old_blob_names=False):
    def blob(data):
        d = "sha256:" + hashlib.sha256(data).hexdigest()
        os.makedirs(os.path.join(root, "blobs"), exist_ok=True)
        with open(os.path.join(root, "blobs", d if old_blob_names else d.replace(":", "-")), "wb") as f:
            f.write(data)
        return d, len(data)
    layers = []
    for mt, data in (("application/vnd.ollama.image.model", gguf_bytes), ("application/vnd.ollama.image.template", template.encode()),
                     ("application/vnd.ollama.image.params", json.dumps(params or {"stop": ["</s>"]}).encode())):
        d, n = blob(data)
        layers.append({"mediaType": mt, "digest": d, "size": n})
    man = {"schemaVersion": 2, "mediaType": "application/vnd.docker.distribution.manifest.v2+json",
           "config": {"mediaType": "application/vnd.docker.container.image.v1+json", "digest": layers[0]["digest"], "size": 2}, "layers": layers}
    mdir = os.path.join(root, "manifests", "registry.ollama.ai", *name.split("/"))
    os.makedirs(mdir, exist_ok=True)
    with open(os.path.join(mdir, tag), "w") as f:
        json.dump(man, f)
    return layers


class StoreEntry(C.Structure):
    _fields_ = [("name", C.c_char * 256), ("model_path", C.c_char * 1024), ("template_path", C.c_char * 1024), ("system_path", C.c_char * 1024),
                ("params_path", C.c_char * 1024), ("digest", C.c_char * 80), ("model_size", C.c_uint64), ("modified_unix", C.c_int64)]


def test_model_store_resolves_names_to_blobs_and_loads(pkg, tmp_path, model_dir):
    L = _lib(pkg)
    home = tmp_path / "ollama_home"
    root = str(home / "models")
    gguf_path = pkg.synth.cached_model("tiny-llama", model_dir)
    data = open(gguf_path, "rb").read()
    layers = _write_store(root, "library/phi", "latest", data)
    _write_store(root, "library/llama2", "7b", data[: len(data)], old_blob_names=True)
    _write_store(root, "someone/custom", "v1", data)
    e = StoreEntry()
    for models_arg in (root, str(home)):                     # the models directory itself, or the OLLAMA home above it
        assert L.b2_store_resolve(models_arg.encode(), b"phi", C.byref(e)) == 0, L.b2_strerror()
        assert e.name == b"phi:latest" and e.digest.decode() == layers[0]["digest"] and e.model_size == len(data)
        assert open(e.model_path, "rb").read() == data
        assert open(e.template_path).read() == "{{ .Prompt }}" and json.load(open(e.params_path))["stop"] == ["</s>"]
    assert L.b2_store_resolve(root.encode(), b"llama2:7b", C.byref(e)) == 0 and e.name == b"llama2:7b"
    assert L.b2_store_resolve(root.encode(), b"registry.ollama.ai/someone/custom:v1", C.byref(e)) == 0 and e.name == b"someone/custom:v1"
    assert L.b2_store_resolve(root.encode(), b"phi:nosuchtag", C.byref(e)) == -3 and b"not pulled" in L.b2_strerror()
    assert L.b2_store_resolve(root.encode(), b"../../etc/passwd", C.byref(e)) < 0
    buf = C.create_string_buffer(4096)
    n = L.b2_store_list(root.encode(), buf, 4096)
    names = sorted(line.split("\t")[0] for line in buf.raw[:n].decode().splitlines())
    assert names == ["llama2:7b", "phi:latest", "someone/custom:v1"]
    # a manifest whose blob is gone is reported, not followed
    assert L.b2_store_resolve(root.encode(), b"llama2:7b", C.byref(e)) == 0
    os.remove(e.model_path)                                   # (the three models share one blob: same bytes, same digest ... but llama2's uses the old file name)
    assert L.b2_store_resolve(root.encode(), b"llama2:7b", C.byref(e)) in (0, -3)
    for f in os.listdir(os.path.join(root, "blobs")):
        if os.path.getsize(os.path.join(root, "blobs", f)) == len(data):
            os.remove(os.path.join(root, "blobs", f))
    assert L.b2_store_resolve(root.encode(), b"someone/custom:v1", C.byref(e)) == -3 and b"missing" in L.b2_strerror()
