"""Teaching models of the reference's notebooks (models/teaching.py): recurrent cells against torch.nn, masks and causality of
the encoder-decoder transformer, and that each model trains on a toy task."""
import pytest
import torch
from torch import nn

from llm_in_practise_b200.models import teaching as T


@pytest.mark.parametrize("cell,bidir", [("rnn", False), ("gru", False), ("lstm", False), ("gru", True), ("lstm", True)])
def test_recurrent_layer_matches_torch(cell, bidir):
    torch.manual_seed(0)
    ref = {"rnn": nn.RNN, "gru": nn.GRU, "lstm": nn.LSTM}[cell](12, 16, batch_first=True, bidirectional=bidir)
    mine = T.RecurrentLayer(12, 16, cell, bidir).load_torch(ref)
    x = torch.randn(3, 7, 12, requires_grad=True)
    x2 = x.detach().clone().requires_grad_()
    out_ref, h_ref = ref(x)
    h_ref = h_ref[0] if cell == "lstm" else h_ref
    out, h = mine(x2)
    assert torch.allclose(out, out_ref, atol=1e-5) and torch.allclose(h, h_ref.transpose(0, 1).reshape(3, -1), atol=1e-5)
    out.square().sum().backward()
    out_ref.square().sum().backward()
    assert torch.allclose(x2.grad, x.grad, atol=1e-4)
    assert torch.allclose(mine.w_hh[0].grad, ref.weight_hh_l0.grad, atol=1e-4)
    # variable lengths: same final state as pack_padded_sequence
    lengths = torch.tensor([7, 4, 2])
    packed = nn.utils.rnn.pack_padded_sequence(x.detach(), lengths, batch_first=True, enforce_sorted=True)
    _, hp = ref(packed)
    hp = hp[0] if cell == "lstm" else hp
    _, h_len = mine(x.detach(), lengths)
    assert torch.allclose(h_len, hp.transpose(0, 1).reshape(3, -1), atol=1e-5)


def _train(model, batch, steps, lr=1e-2):
    opt = torch.optim.Adam(model.parameters(), lr=lr)
    losses = []
    for _ in range(steps):
        loss = model(*batch)["loss"]
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss.detach()))
    return losses


def test_recurrent_classifier_learns_which_token_came_first():
    torch.manual_seed(1)
    ids = torch.randint(1, 20, (64, 9))
    ids[:, 6:] = 0                                            # padded tail
    labels = (ids[:, 0] > 10).long()
    for cell, bidir in (("gru", True), ("lstm", False)):
        m = T.RecurrentClassifier(20, 16, 32, 2, cell, bidir)
        losses = _train(m, (ids, labels), 60)
        assert losses[-1] < 0.3 * losses[0]
        acc = (m(ids)["logits"].argmax(-1) == labels).float().mean()
        assert acc > 0.9
    # the padded tail does not change the prediction
    m.eval()
    assert torch.allclose(m(ids)["logits"], m(ids[:, :6])["logits"], atol=1e-5)


def test_encoder_decoder_masks_and_causality():
    torch.manual_seed(2)
    m = T.EncoderDecoderTransformer(30, 30, d_model=32, n_head=4, n_layer=2, d_ff=64, dropout=0.0, max_len=32).eval()
    src = torch.randint(1, 30, (2, 6))
    tgt = torch.randint(1, 30, (2, 5))
    base = m(src, tgt)["logits"]
    padded = m(torch.cat([src, torch.zeros(2, 3, dtype=torch.long)], 1), tgt)["logits"]
    assert torch.allclose(base, padded, atol=1e-5)            # source padding is invisible
    tgt2 = tgt.clone()
    tgt2[:, 3:] = torch.randint(1, 30, (2, 2))
    assert torch.allclose(base[:, :3], m(src, tgt2)["logits"][:, :3], atol=1e-5)      # no look-ahead in the decoder
    assert not torch.allclose(base[:, 3:], m(src, tgt2)["logits"][:, 3:], atol=1e-3)
    src2 = src.clone()
    src2[:, 0] = (src2[:, 0] % 29) + 1
    assert not torch.allclose(base, m(src2, tgt)["logits"], atol=1e-4)                # cross-attention reads the encoder


def test_encoder_decoder_learns_to_copy():
    torch.manual_seed(3)
    BOS, EOS = 1, 2
    body = torch.randint(3, 12, (32, 5))
    src = body
    tgt = torch.cat([torch.full((32, 1), BOS), body, torch.full((32, 1), EOS)], 1)
    m = T.EncoderDecoderTransformer(12, 12, d_model=32, n_head=4, n_layer=1, d_ff=64, dropout=0.0, max_len=16)
    losses = _train(m, (src, tgt[:, :-1], tgt[:, 1:]), 150, lr=5e-3)
    assert losses[-1] < 0.1 * losses[0]
    out = m.greedy_decode(src[:4], BOS, EOS, max_new_tokens=8)
    assert torch.equal(out[:, 1:6], body[:4]) and bool((out[:, 6] == EOS).all())


def test_mlp_and_convnets():
    torch.manual_seed(4)
    x = torch.randn(8, 1, 32, 32)
    net = T.lenet5()
    assert net(x).shape == (8, 10) and net.head.layers[0].in_features == 16 * 5 * 5
    maps = net.features(x)
    assert [tuple(f.shape[1:]) for f in maps] == [(6, 28, 28), (16, 10, 10)]
    gap = T.ConvNet(3, (8, 16), 5, 16, gap=True)
    y = gap(torch.randn(4, 3, 16, 16))
    assert y.shape == (4, 5)
    y.sum().backward()
    assert gap.convs[0].weight.grad is not None
    mlp = T.MLP(20, (32, 16), 3, activation="tanh", batch_norm=True, dropout=0.1)
    xs = torch.randn(64, 20)
    ys = (xs[:, 0] > 0).long() + (xs[:, 1] > 0).long()
    opt = torch.optim.Adam(mlp.parameters(), lr=1e-2)
    first = last = None
    for i in range(80):
        loss = torch.nn.functional.cross_entropy(mlp(xs), ys)
        opt.zero_grad()
        loss.backward()
        opt.step()
        first = float(loss) if first is None else first
        last = float(loss)
    assert last < 0.5 * first


def test_seq2seq_with_additive_attention_learns_to_reverse():
    torch.manual_seed(5)
    BOS, EOS = 1, 2
    body = torch.randint(3, 10, (48, 4))
    src = torch.cat([body, torch.zeros(48, 2, dtype=torch.long)], 1)                  # two pad positions
    tgt = torch.cat([torch.full((48, 1), BOS), body.flip(1), torch.full((48, 1), EOS)], 1)
    m = T.Seq2SeqRNN(10, 10, embed_dim=16, hidden_size=32, attn_dim=16)
    losses = _train(m, (src, tgt), 200, lr=1e-2)
    assert losses[-1] < 0.1 * losses[0]
    toks, maps = m.translate(src[:6], BOS, EOS, max_new_tokens=6)
    assert torch.equal(toks[:, :4], body[:6].flip(1))
    assert torch.allclose(maps.sum(-1), torch.ones_like(maps.sum(-1)), atol=1e-5)    # attention weights are a distribution
    assert float(maps[..., 4:].abs().max()) == 0.0                                     # ... with no mass on padding
    mixed = m(src, tgt, teacher_forcing=0.5)["loss"]                                   # scheduled sampling path
    assert torch.isfinite(mixed)
