"""CPU: host-side logic of the drop-in (constructors, state_dict layout, error conventions, masks,
sequence bookkeeping).  No kernel is launched here."""
import pytest
import torch

from acai_omr_b200.layers import Packed, lens_from_mask
from acai_omr_b200.models import kv_caching as K
from acai_omr_b200.models import models as M

VOCAB = M.default_vocab_path()


def test_vitomr_state_dict_layout_matches_reference(golden):
    g = golden["vitomr"]
    enc = M.FineTuneOMREncoder(4, 6, 10, 2, num_layers=2, hidden_dim=32, num_heads=4, mlp_dim=64)
    dec = M.OMRDecoder(20, VOCAB, num_layers=2, hidden_dim=48, num_heads=4, mlp_dim=96, use_caching=True,
                       max_batch_size=4, cache_dtype=torch.float)
    vit = M.TeacherForcedViTOMR(enc, None, dec, transition_head_dim=40)
    assert {k: tuple(v.shape) for k, v in vit.state_dict().items()} == golden["vitomr_keys"]
    vit.load_state_dict(g["sd"])  # strict
    # KV caches are non-persistent buffers (kv_caching.py:36-44)
    assert not any("cache" in k for k in vit.state_dict())
    # full fine-tune: nothing frozen (models.py:677)
    enc_full = M.FineTuneOMREncoder(4, 6, 10, 2, num_layers=2, hidden_dim=32, num_heads=4, mlp_dim=64)
    assert enc_full.frozen_blocks is None
    assert all(p.requires_grad for p in M.TeacherForcedViTOMR(enc_full, None, dec, 40).parameters())


def test_freezing_semantics():
    """models.py:667-676."""
    dec = M.OMRDecoder(20, VOCAB, num_layers=1, hidden_dim=16, num_heads=2, mlp_dim=16)
    enc = M.FineTuneOMREncoder(4, 6, 10, 1, num_layers=3, hidden_dim=16, num_heads=2, mlp_dim=16)
    vit = M.TeacherForcedViTOMR(enc, None, dec, 8)
    assert not any(p.requires_grad for p in vit.encoder.frozen_blocks.parameters())
    assert not vit.encoder.pos_embedding.requires_grad and not vit.encoder.projection.weight.requires_grad
    assert all(p.requires_grad for p in vit.encoder.fine_tune_blocks.parameters())
    vit2 = M.TeacherForcedViTOMR(M.OMREncoder(4, 6, 10, num_layers=1, hidden_dim=16, num_heads=2, mlp_dim=16), None, dec, 8)
    assert not any(p.requires_grad for p in vit2.encoder.parameters())
    groups, lrs = vit.create_fine_tune_param_groups(1e-3, 1e-4, 0.5)
    assert len(groups) == 2 + 1 + 3 and lrs == [1e-4]


def test_mae_to_encoder_key_mapping():
    """models.py:679-713."""
    dec = M.OMRDecoder(20, VOCAB, num_layers=1, hidden_dim=16, num_heads=2, mlp_dim=16)
    enc = M.FineTuneOMREncoder(4, 6, 10, 1, num_layers=3, hidden_dim=16, num_heads=2, mlp_dim=16)
    src = M.Encoder(4, 6, 10, num_layers=3, hidden_dim=16, num_heads=2, mlp_dim=16)
    mae_sd = {"encoder." + k: v for k, v in src.state_dict().items()}
    mae_sd["mask_token"] = torch.zeros(1)
    vit = M.TeacherForcedViTOMR(enc, mae_sd, dec, 8)
    assert torch.equal(vit.encoder.frozen_blocks.layers[1].linear1.weight, src.encoder_blocks.layers[1].linear1.weight)
    assert torch.equal(vit.encoder.fine_tune_blocks.layers[0].linear1.weight, src.encoder_blocks.layers[2].linear1.weight)
    assert torch.equal(vit.encoder.fine_tune_blocks.norm.weight, src.encoder_blocks.norm.weight)


def test_decoder_attributes_and_tokeniser():
    dec = M.OMRDecoder(15, VOCAB, num_layers=2, hidden_dim=24, num_heads=4, mlp_dim=48)
    assert (dec.bos_idx, dec.pad_idx, dec.eos_idx, dec.vocab_size) == (0, 1, 2, 227)
    # reference tests/test_omr_teacher_force_train.py:22-23
    assert [dec.tokens_to_idxs[t] for t in "measure key:fifths:-7 time".split()] == [3, 4, 19]
    assert dec.head_dim == 6.0 and dec.decoder_blocks.layers[0].self_attn.head_dim == 6
    assert dec.vocab_embedding.weight[dec.pad_idx].abs().sum() == 0  # padding_idx row
    cached = dec.to_cached_version(4, torch.float32)
    assert isinstance(cached.decoder_blocks, K.CachedTransformerDecoder)
    cached.load_state_dict(dec.state_dict())
    assert cached.decoder_blocks.self_attn_caches[0].k_cache.shape == (4, 4, 15, 6)
    with pytest.raises(RuntimeError):
        dec.prepare_caches(torch.zeros(1, 2, 24))
    with pytest.raises(RuntimeError):
        cached.cached_generate(torch.zeros(1, 1, dtype=torch.long), 15)
    with pytest.raises(ValueError):
        dec.generate(torch.zeros(1, 16, dtype=torch.long), torch.zeros(1, 2, 24))
    with pytest.raises(AssertionError):
        K.CachedTransformerDecoder(M.TransformerDecoderLayer(8, 2), 1, 1, 4, torch.float32)
    with pytest.raises(AssertionError):
        M.FineTuneOMREncoder(4, 6, 10, 0)


def test_encoder_grid_errors_fire_before_any_launch():
    enc = M.Encoder(2, 3, 3, num_layers=1, hidden_dim=8, num_heads=1, mlp_dim=8)
    with pytest.raises(ValueError):
        enc.batchify([torch.ones(1, 8, 4)])  # 4 x 2 patches > 3 x 3 grid   (models.py:45-46)
    assert enc.create_attention_mask([4, 8], 8).tolist() == [[False] * 4 + [True] * 4, [False] * 8]


def test_kvcache_host_logic():
    c = K.KVCache(2, 3, 2, 6, dtype=torch.float)
    assert c.size == 0 and c.k_cache.shape == (2, 2, 3, 6) and c.cache_pos.tolist() == [0, 1, 2]
    with pytest.raises(ValueError):
        c.update(torch.zeros(3, 2, 1, 6), torch.zeros(3, 2, 1, 6))  # kv_caching.py:84-88
    with pytest.raises(AssertionError):
        c.update(torch.zeros(2, 2, 4, 6), torch.zeros(2, 2, 4, 6))  # kv_caching.py:90
    assert "k_cache" not in c.state_dict()


def test_packed_and_masks():
    pk = Packed([3, 5, 1], "cpu")
    assert pk.cu.tolist() == [0, 3, 8, 9] and pk.total == 9 and pk.max_len == 5
    m = pk.mask("cpu")
    assert m.tolist() == [[False] * 3 + [True] * 2, [False] * 5, [False] + [True] * 4]
    assert lens_from_mask(m, 3, 5) == [3, 5, 1]
    assert lens_from_mask(m.clone(), 3, 5) == [3, 5, 1]  # without the attached lengths (read back)
    assert lens_from_mask(None, 2, 7) == [7, 7]
    bad = torch.tensor([[False, True, False]])
    with pytest.raises(ValueError):
        lens_from_mask(bad, 1, 3)


def test_split_lmx_and_inference_mask():
    """reference tests/test_vitomr.py:151-172; models.py:550-559."""
    seqs = [torch.tensor([0, 2, 3, 226]), torch.tensor([0, 2, 2, 3, 4, 226])]
    inp, tgt, mask = M.batchify_and_split_lmx_seqs(seqs, 1, "cpu")
    assert torch.equal(inp, torch.tensor([[0, 2, 3, 226, 1], [0, 2, 2, 3, 4]]))
    assert torch.equal(tgt, torch.tensor([[2, 3, 226, 1, 1], [2, 2, 3, 4, 226]]))
    assert mask.tolist() == [[False] * 4 + [True], [False] * 5]
    dec = M.OMRDecoder(8, VOCAB, num_layers=1, hidden_dim=8, num_heads=1, mlp_dim=8)
    vit = M.ViTOMR(None, None, dec)
    s = torch.tensor([[0, 5, 2, 7, 2], [0, 5, 6, 7, 8]])
    assert vit.create_inference_mask(s).tolist() == [[True, True, True, False, False], [True] * 5]
    a, b, c = vit.mask_and_clip_seqs(s, torch.ones(2, 5))
    assert a.tolist() == [[0, 5, 2, 1, 1], [0, 5, 6, 7, 8]] and b[0].tolist() == [1, 1, 1, 0, 0]


def test_mae_state_dict_matches_reference_fixture(golden):
    """debug_pretrained_mae.pth (the reference's tiny MAE fixture, tests/test_pre_train.py:29): same keys and shapes."""
    keys = golden["debug_mae_keys"]
    mae = M.MAE(0.75, 16, 60, 200, encoder_hidden_dim=10, decoder_hidden_dim=10,
                encoder_kwargs=dict(num_layers=2, num_heads=1, mlp_dim=1), decoder_kwargs=dict(num_layers=2, num_heads=1, mlp_dim=1))
    ours = {k: tuple(v.shape) for k, v in mae.state_dict().items()}
    assert set(ours) == set(keys)
    mismatched = {k: (ours[k], keys[k]) for k in keys if ours[k] != keys[k] and "linear" not in k}
    assert not mismatched, mismatched


def test_mae_mask_indices_roundtrip():
    """tests/test_mae.py:36-55: unshuffling the shuffled order restores the sequence; 25 % kept."""
    enc = M.MAEEncoder(0.75, 4, 6, 10, num_layers=1, num_heads=1, hidden_dim=8, mlp_dim=8)
    keep, restore, seq_mask, n = enc.mask_indices(40, "cpu")
    assert n == 10 and keep.shape == (10,) and int(seq_mask.sum()) == 30
    shuffled = torch.argsort(restore)
    assert torch.equal(shuffled[:n], keep) and torch.equal(shuffled[restore], torch.arange(40))
    assert torch.equal(seq_mask[keep], torch.zeros(10, dtype=torch.int))


def test_weight_cache_entries_die_with_their_parameter():
    """The low-precision weight copies are keyed by id(parameter): a freed parameter's id, data pointer and version
    counter all come back for the next allocation, so an entry must never outlive (or be served to a successor of) its
    parameter.  CPython reuses ids immediately in a loop like this one."""
    import gc

    import torch
    from acai_omr_b200.layers import WEIGHTS
    WEIGHTS.clear()
    for i in range(200):
        p = torch.nn.Parameter(torch.full((4, 8), float(i)))
        c = WEIGHTS.get(p, torch.bfloat16)
        assert c.dtype == torch.bfloat16 and float(c[0, 0]) == float(i), i
        assert WEIGHTS.get(p, torch.bfloat16) is c                        # cached while the parameter lives
        r = WEIGHTS.get(p, torch.bfloat16, rows=(1, 3))
        assert r.shape == (2, 8) and float(r[0, 0]) == float(i)
        with torch.no_grad():
            p.add_(1.0)                                                   # in-place update: version bump -> refreshed
        assert float(WEIGHTS.get(p, torch.bfloat16)[0, 0]) == float(i) + 1.0
        del p, c, r
    gc.collect()
    assert len(WEIGHTS._store) == 0


# ---- attention-probability dropout: the bit-parallel mask function of csrc/common.cuh, restated in numpy ---------------
def _attn_drop_word(seed, rowkey, kb, thr):
    """attn_drop_word(attn_drop_rowconst(seed, rowkey), attn_drop_rowbase(rowkey), kb, thr) of common.cuh: the 32 keep
    decisions of keys [32 kb, 32 kb + 32) of probability row `rowkey`; key 4 s + j is bit 8 j + 7 - s."""
    import numpy as np
    m32, u = np.uint64(0xFFFFFFFF), np.uint64
    k3 = [0xC2B2AE3D, 0x27D4EB2F, 0x165667B1, 0xD3A2646D, 0xFD7046C5, 0xB55A4F09, 0x7FEB352D, 0x846CA68B,
          0xCC9E2D51, 0x1B873593, 0xE6546B65, 0x2C1B3C6D, 0x297A2D39, 0x5BD1E995, 0x68E31DA5, 0xB5297A4D]

    def mix32(x):
        x = x & m32
        x = x ^ (x >> u(16)); x = (x * u(0x7FEB352D)) & m32
        x = x ^ (x >> u(15)); x = (x * u(0x846CA68B)) & m32
        return x ^ (x >> u(16))

    def fold32(x, k):
        m = (x & m32) * u(k)
        return (m ^ (m >> u(32))) & m32

    rowkey, kb = np.asarray(rowkey, dtype=np.uint64), np.asarray(kb, dtype=np.uint64)
    sh = rowkey << u(11)
    rc = mix32((sh >> u(32)) ^ u(seed & 0xFFFFFFFF)) ^ u(seed >> 32)
    h = fold32(fold32((((sh & m32) + kb) & m32) ^ rc, 0x9E3779B1), 0x85EBCA77)
    ge = np.full(h.shape, 0xFFFFFFFF, dtype=np.uint64)
    for i in range(16):                                        # bit-sliced u >= thr, least significant slice first
        w = fold32(h, k3[i])
        ge = (w & ge) if (thr >> i) & 1 else (w | ge)
    return ge


def test_attention_dropout_mask_function_known_answers():
    """Values printed by a HOST build of common.cuh (the functions are __host__ __device__): pins the numpy restatement
    below to the kernels' definition; tests/test_training_gpu.py pins every kernel to acai_attn_dropout_mask."""
    import numpy as np
    want = {(0, 0): 4294918143, (0, 1): 3200253695, (0, 37): 3154116575, (1, 0): 4290772991, (1, 1): 4026507263,
            (1, 37): 4278188031, (5000000, 0): 4294696927, (5000000, 1): 4294964987, (5000000, 37): 3220699135,
            ((1 << 33) + 17, 0): 2885681151, ((1 << 33) + 17, 1): 4294959039, ((1 << 33) + 17, 37): 4269785023}
    thr = 6554                                                 # dropout_threshold(0.1f)
    for (rk, kb), v in want.items():
        assert int(_attn_drop_word(0xACA100000042, np.array([rk]), np.array([kb]), thr)[0]) == v


@pytest.mark.parametrize("p", [0.05, 0.1, 0.25])
def test_attention_dropout_mask_statistics(p):
    """Drop rate, neighbour correlations along keys / rows / the diagonal, and the variance of per-row, per-column and
    per-32-key-block drop counts against the binomial, on 2 048 rows x 1 024 keys of one head."""
    import numpy as np
    thr = int(p * 65536 + 0.5)
    rows, nkeys = 2048, 1024
    rk = np.arange(rows, dtype=np.uint64) + np.uint64(123457)
    w = _attn_drop_word(99, rk[:, None], np.arange(nkeys // 32, dtype=np.uint64)[None, :], thr)
    kk = np.arange(32)
    bit = (8 * (kk & 3) + 7 - (kk >> 2)).astype(np.uint64)
    drop = 1.0 - ((w[:, :, None] >> bit[None, None, :]) & np.uint64(1)).reshape(rows, nkeys).astype(np.float64)
    n = drop.size
    assert abs(drop.mean() - thr / 65536) < 4 * (p * (1 - p) / n) ** 0.5

    def corr(a, b):
        a, b = a - a.mean(), b - b.mean()
        return float((a * b).mean() / ((a * a).mean() * (b * b).mean()) ** 0.5)

    lim = 5 / n ** 0.5                                          # five standard errors of a zero correlation
    for a, b in ((drop[:, :-1], drop[:, 1:]), (drop[:, :-2], drop[:, 2:]), (drop[:, :-4], drop[:, 4:]),
                 (drop[:, :-32], drop[:, 32:]), (drop[:-1], drop[1:]), (drop[:-2], drop[2:]), (drop[:-1, :-1], drop[1:, 1:])):
        assert abs(corr(a, b)) < lim
    assert 0.85 < drop.mean(axis=1).var() / (p * (1 - p) / nkeys) < 1.15
    assert 0.85 < drop.mean(axis=0).var() / (p * (1 - p) / rows) < 1.15
    assert 0.95 < drop.reshape(rows, nkeys // 32, 32).sum(axis=2).var() / (32 * p * (1 - p)) < 1.05
