"""Host-side logic of the drop-in that needs no GPU, held to the oracle (which is pinned to the reference):
input-independent encodings (paper.py:46-80, transformer.py:35-48), inverse_sigmoid (util/misc.py:477-481) and the
synthetic clip generator's dataset-dict contract (datasets/utils.py:19-38)."""
import torch

from fod_b200 import model as M
from fod_b200.synthetic import make_clip
from oracle import fod_oracle as O


def test_spatial_encoding_matches_the_oracle():
    pe = M.PositionalEncoder(no_temporal=True)
    for h, w in ((1, 1), (3, 4), (14, 25), (28, 50)):
        got = pe.spatial_tokens(h, w, 256, torch.device("cpu"))            # (h*w, c), tokens in (h w) order
        want = O.image_pos_encoding(h, w, 256).flatten(1).t()               # (c, h, w) -> (h*w, c)
        assert got.shape == (h * w, 256)
        assert torch.equal(got, want)
    assert pe.spatial_tokens(3, 4, 256, torch.device("cpu")) is pe.spatial_tokens(3, 4, 256, torch.device("cpu"))  # cached


def test_temporal_code_matches_the_oracle():
    """paper.py:66-73: one vector per (clip, frame) — offsets given (+ extra offset) or the frame counter."""
    dev = torch.device("cpu")
    offs = torch.tensor([[-1.0, -0.5, -0.25], [-0.9, -0.3, -0.1]])
    for extra in (0.0, 0.37):
        pe = M.PositionalEncoder(no_temporal=False, extra_temporal_offset=extra)
        got = pe.temporal_channels(offs, 2, 3, 256, dev)
        want = O.temporal_encoding(offs, 2, 3, 2, 2, extra=extra)[:, :, :, 0, 0]
        assert got.shape == (2, 3, 256)
        assert torch.allclose(got, want, atol=1e-6, rtol=0)
    pe = M.PositionalEncoder(no_temporal=False)
    got = pe.temporal_channels(None, 2, 4, 256, dev)
    want = O.temporal_encoding(None, 2, 4, 1, 1)[:, :, :, 0, 0]
    assert torch.allclose(got, want, atol=1e-6, rtol=0)


def test_query_sine_embedding_and_inverse_sigmoid_match_the_oracle():
    g = torch.Generator().manual_seed(3)
    ref = torch.rand(128, 2, generator=g)
    got = M.gen_sineembed_for_position(ref)
    want = O.query_sine_embedding(ref[:, None, :])[:, 0]
    assert torch.equal(got, want)
    x = torch.tensor([-0.5, 0.0, 1e-7, 1e-5, 0.3, 0.5, 1 - 1e-6, 1.0, 1.5])
    assert torch.equal(M.inverse_sigmoid(x), O.inverse_sigmoid(x))


def test_synthetic_clip_follows_the_dataset_dict_contract():
    d = make_clip(3, 4, 64, 96, seed=5)
    assert d["video"].shape == (3, 4, 3, 64, 96)
    for k, c in (("translation", 3), ("acceleration", 3), ("rotation", 4), ("rotation_rate", 3), ("speed", 1)):
        assert d[k].shape == (3, 4, c)
    assert torch.all(d["translation"][:, 0] == 0)                            # frame 0 is the origin (nu_scenes.py:276-279)
    assert torch.allclose(d["rotation"].norm(dim=-1), torch.ones(3, 4), atol=1e-5)
    assert torch.equal(d["rotation"][:, 0], torch.tensor([1.0, 0, 0, 0]).expand(3, 4))
    assert d["boxes"].shape == (3, 256, 4) and d["classes"].shape == (3, 256) and d["active"].shape == (3, 256)
    act = d["active"].bool()
    b = d["boxes"][act]
    assert torch.all(b[:, 2] > b[:, 0]) and torch.all(b[:, 3] > b[:, 1])     # never degenerate (box_ops.py:59-60)
    assert torch.all(d["boxes"][~act] == 0) and int(d["classes"].max()) < 8
    assert torch.equal(d["annotated_frame_idx"], torch.full((3,), 3))
    e = make_clip(3, 4, 64, 96, seed=5)
    assert all(torch.equal(d[k], e[k]) for k in d)                           # seeded: identical on every host
    assert not torch.equal(d["video"], make_clip(3, 4, 64, 96, seed=6)["video"])
    assert int(make_clip(2, 3, 64, 96, seed=1, fixed_targets=0)["active"].sum()) == 0
