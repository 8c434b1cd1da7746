"""The wide classifier head on the tensor cores (csrc/head_wide.cu: TMA + tcgen05.mma + TMEM, online log-sum-exp epilogue)
against the oracle: F.cross_entropy(F.linear(features.float(), weight.float(), bias), labels, reduction='none') on the
bf16-rounded operands (fp32 accumulation either way), rtol/atol 2e-5.  The reference's Bottleneck ResNets end in
nn.Linear(2048, num_classes) (pytorch_model.py:39-64, 78, 94-96) followed by the per-sample loss (pytorch_collab.py:102)."""
import numpy as np
import pytest
import torch

from oracle import mercury_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from mercury_b200 import ops as _ops

    return _ops


def _case(n, d, c, seed, scale=0.05):
    g = torch.Generator().manual_seed(seed)
    feats = torch.randn(n, d, generator=g).to(torch.bfloat16)
    w = (torch.randn(c, d, generator=g) * scale).to(torch.bfloat16)
    b = torch.randn(c, generator=g) * 0.1
    y = torch.randint(0, c, (n,), generator=g)
    return feats, w, b, y


@pytest.mark.parametrize("n,d,c,bias", [
    (128, 64, 256, True),        # one row tile, one class tile, one K block
    (1000, 2048, 1000, True),    # the ImageNet-shaped head: 4 class tiles (the last one ragged), 32 K blocks, ragged rows
    (333, 512, 17, False),       # just above the small-head limit, one ragged class tile
    (5000, 2048, 1000, True),    # several row tiles per CTA is not reached at this size; many CTAs
    (40000, 1024, 1500, True),   # persistent loop: more row tiles than SMs, 6 class tiles
    (257, 72, 300, True),        # feature width not a multiple of the K block (zero-filled by TMA)
    (70000, 512, 10, True),      # narrow shape (class tiles of 64, 8 stages): the reference's 512 -> 10 head with bf16 weights
    (3000, 512, 100, True),      # narrow shape, two class tiles (the second one ragged)
    (1500, 256, 128, False),     # narrow shape, exactly two full class tiles
    (1500, 256, 129, False),     # one class more: back to the 256-wide shape
])
def test_wide_head_matches_linear_then_cross_entropy(ops, n, d, c, bias):
    feats, w, b, y = _case(n, d, c, seed=n + c)
    if n > 10:
        y[3] = -100                     # ignore_index scores 0
    got = ops.score_head(feats.to(DEV), w.to(DEV), b.to(DEV) if bias else None, y.to(DEV))
    torch.cuda.synchronize()
    ref = orc.score_head(feats.float(), w.float(), b if bias else None, y)
    np.testing.assert_allclose(got.cpu().numpy(), ref.numpy(), rtol=2e-5, atol=2e-5)
    # and against materialised logits scored by K1 (the unfused path the head replaces)
    logits = torch.nn.functional.linear(feats.float(), w.float(), b if bias else None)
    k1 = ops.score(logits.to(DEV), y.to(DEV))
    np.testing.assert_allclose(got.cpu().numpy(), k1.cpu().numpy(), rtol=2e-5, atol=2e-5)


def test_wide_head_bad_label_raises_like_torch(ops):
    feats, w, b, y = _case(300, 128, 40, seed=1)
    y[7] = 40
    st = ops.new_status(DEV)
    ops.score_head(feats.to(DEV), w.to(DEV), b.to(DEV), y.to(DEV), status=st)
    with pytest.raises(IndexError):
        ops.check_status(st)


def test_wide_head_on_the_reference_resnet50_head(ops):
    """The shape of the reference's ResNet50(num_classes=1000) classifier (pytorch_model.py: Bottleneck, expansion 4 ->
    nn.Linear(2048, 1000), torch's own initialisation) on pooled post-ReLU features, bf16 operands."""
    torch.manual_seed(0)
    lin = torch.nn.Linear(2048, 1000)   # the layer ResNet50(num_classes=1000).linear is (pytorch_model.py:78 with expansion 4)
    g = torch.Generator().manual_seed(9)
    feats = torch.relu(torch.randn(2000, 2048, generator=g)).to(torch.bfloat16)   # pooled post-ReLU features
    y = torch.randint(0, 1000, (2000,), generator=g)
    w, b = lin.weight.detach().to(torch.bfloat16), lin.bias.detach()
    got = ops.score_head(feats.to(DEV), w.to(DEV), b.to(DEV), y.to(DEV))
    ref = orc.score_head(feats.float(), w.float(), b, y)
    np.testing.assert_allclose(got.cpu().numpy(), ref.numpy(), rtol=2e-5, atol=2e-5)
