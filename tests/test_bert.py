"""Description front end: BertPooler through the C ABI against transformers.BertModel -- the library call the
reference makes (run/say.py:99-106) -- with the same (seeded, random-init: no pretrained weights offline) parameters on the host.
Bar: pooler_output max-abs <= 1e-3 (tanh-bounded), last_hidden_state <= 2e-3 (LayerNorm-normalised, O(1))."""
import pytest
import torch

from tacotron2_b200.bert import BERT_BASE, BertPooler, DescriptionEmbeddingCache

SMALL = dict(BERT_BASE, hidden_size=256, num_hidden_layers=2, num_attention_heads=4, intermediate_size=512, vocab_size=1000)


def _hf(config, seed):
    tr = pytest.importorskip("transformers")
    torch.manual_seed(seed)
    cfg = tr.BertConfig(**{k: v for k, v in config.items()})
    m = tr.BertModel(cfg).eval()
    with torch.no_grad():   # the default init (std 0.02, zero biases, unit LayerNorm) is too tame: perturb everything
        for p in m.parameters():
            p.add_(torch.randn_like(p) * 0.02)
    return m


def test_state_dict_matches_hf_cpu():
    hf = _hf(SMALL, 0)
    m = BertPooler(SMALL)
    want = {k: tuple(v.shape) for k, v in hf.state_dict().items() if not k.endswith("position_ids")}
    assert {k: tuple(v.shape) for k, v in m.state_dict().items()} == want
    m2 = BertPooler.from_hf(hf)
    assert torch.equal(m2.pooler.dense.weight, hf.pooler.dense.weight)


def test_description_cache_roundtrip_cpu(tmp_path):
    c = DescriptionEmbeddingCache(str(tmp_path))
    e = torch.randn(2, 768)
    c.save_batch(["a/x.pt", "y.pt"], e)
    assert tuple(torch.load(c.path("a/x.pt")).shape) == (768,)      # the dataset unsqueezes it to (1, 768)
    got = c.load_batch(["a/x.pt", None, "y.pt"], 768)
    assert torch.equal(got[0], e[0]) and torch.equal(got[2], e[1]) and float(got[1].abs().sum()) == 0.0


@pytest.mark.gpu
@pytest.mark.parametrize("config,B,S", [(SMALL, 3, 17), (SMALL, 1, 1), (BERT_BASE, 4, 24), (BERT_BASE, 2, 130)])
def test_pooler_vs_transformers(config, B, S):
    hf = _hf(config, 1)
    dev = torch.device("cuda:0")
    m = BertPooler.from_hf(hf).to(dev).eval()
    g = torch.Generator().manual_seed(3)
    ids = torch.randint(0, config["vocab_size"], (B, S), generator=g)
    tt = torch.randint(0, 2, (B, S), generator=g)
    lens = torch.randint(max(1, S // 2), S + 1, (B,), generator=g)
    lens[0] = S
    am = (torch.arange(S)[None, :] < lens[:, None]).long()
    with torch.no_grad():
        want = hf(input_ids=ids, token_type_ids=tt, attention_mask=am)
        got = m(ids.to(dev), tt.to(dev), am.to(dev))
    assert float((got.pooler_output.cpu() - want.pooler_output).abs().max()) <= 1e-3
    valid = am.bool()[:, :, None]
    err = ((got.last_hidden_state.cpu() - want.last_hidden_state) * valid).abs().max()   # padded query rows are unspecified in both
    assert float(err) <= 2e-3, float(err)
    with torch.no_grad():   # defaults: token types 0, no padding
        w2 = hf(input_ids=ids).pooler_output
        g2 = m(ids.to(dev)).pooler_output
    assert float((g2.cpu() - w2).abs().max()) <= 1e-3
