"""Host logic: the synthetic GGUF writer emits files both readers agree on."""
import numpy as np
import pytest


@pytest.mark.parametrize("preset", ["tiny-llama", "tiny-gqa", "tiny-phi2"])
def test_written_files_read_back(pkg, model_dir, preset):
    import gguf
    path = pkg.synth.cached_model(preset, model_dir)
    cfg = pkg.synth.PRESETS[preset]
    r = gguf.GGUFReader(path)
    names = {t.name: t for t in r.tensors}
    types = pkg.synth.tensor_types(cfg)
    for n, t in types.items():
        assert int(names[n + ".weight"].tensor_type) == int(t), n
    assert r.fields["general.architecture"].contents() == cfg.arch
    assert all(t.data_offset % 32 == 0 for t in r.tensors)


def test_random_blocks_have_requested_scale(pkg):
    import gguf
    from gguf import GGMLQuantizationType as T
    rng = np.random.default_rng(0)
    for t in (T.Q4_0, T.Q8_0, T.Q4_K, T.Q6_K):
        b = pkg.synth.random_blocks(t, 16, 1024, 0.02, rng)
        w = gguf.quants.dequantize(b, t)
        assert np.isfinite(w).all()
        assert 0.012 < w.std() < 0.03, (t, w.std())
        assert abs(w.mean()) < 0.004, (t, w.mean())


def test_q4_k_m_mix_follows_llama_cpp_rule(pkg):
    from gguf import GGMLQuantizationType as T
    cfg = pkg.synth.PRESETS["mistral-7b"]
    types = pkg.synth.tensor_types(cfg)
    hi = [i for i in range(32) if types[f"blk.{i}.attn_v"] == T.Q6_K]
    assert hi == [0, 1, 2, 3, 6, 9, 12, 15, 18, 21, 24, 27, 28, 29, 30, 31]
    assert types["blk.5.attn_q"] == T.Q4_K and types["output"] == T.Q6_K
