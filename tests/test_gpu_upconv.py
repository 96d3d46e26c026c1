"""GPU parity of the fused ConvTranspose3d kernel (kernel = stride = 2, upconv2_sm100.cu: every input voxel writes its
2x2x2 output voxels from one GEMM) against torch's conv_transpose3d on the same bf16-rounded operands."""
import pytest
import torch
import torch.nn.functional as TF

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _q(t):
    return t.bfloat16().float()


def _rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    assert a.shape == b.shape
    return float((a - b).abs().max()) / max(1e-6, float(b.abs().max()))


CASES = [
    # cin, cout, input spatial, batch
    (64, 64, (5, 9, 11), 2),          # ragged tiles, two 256-column chunks
    (128, 128, (4, 16, 8), 1),        # the U-Net level-1 -> 0 up-sampling shape class (double-buffered input tile)
    (256, 256, (3, 8, 8), 1),         # level 2 -> 1: four K chunks, single-buffered input tile, 8 column chunks
    (128, 32, (3, 7, 9), 1),          # 8 * Cout = 256: one chunk holds all 8 parity classes
    (64, 96, (2, 16, 16), 1),         # chunk boundaries inside a parity class
]


@pytest.mark.parametrize("case", CASES)
def test_upconv2_matches_conv_transpose3d(case):
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    cin, cout, sp, n = case
    torch.manual_seed(41)
    x = _q(torch.randn(n, cin, *sp))
    m = torch.nn.ConvTranspose3d(cin, cout, 2, stride=2)
    w, b = _q(m.weight.detach()), m.bias.detach()
    ref = TF.leaky_relu(TF.conv_transpose3d(x, w, b, stride=2), 0.2)
    with rb.precision("bf16"), torch.no_grad():
        xg = x.to(DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
        wg = w.to(DEV)
        assert CV.upconv2_ok(xg, wg, (2, 2, 2), (0, 0, 0), (0, 0, 0), False)
        y = CV.upconv2_fprop(xg, wg, b.to(DEV), 2, 0.2)
    assert y.dtype == torch.bfloat16 and tuple(y.shape) == tuple(ref.shape)
    assert _rel(y, ref) <= 1e-2, _rel(y, ref)


def test_upconv2_many_tiles_strided_input_and_channel_slice_output():
    """More tiles than SMs (persistent loop, barrier phases wrap), the input an interior view of a padded buffer and
    the output the upper channel slice of a wider concat buffer whose other channels stay untouched."""
    import raygun_b200 as rb
    from raygun_b200 import conv as CV
    torch.manual_seed(42)
    cin, cout = 128, 128
    big = _q(torch.randn(1, cin, 14, 42, 42))
    x = big[:, :, 1:13, 1:41, 1:41]
    w = _q(torch.randn(cin, cout, 2, 2, 2) * 0.05)
    ref = TF.conv_transpose3d(x, w, None, stride=2)
    with rb.precision("bf16"), torch.no_grad():
        bg = big.to(DEV).bfloat16().contiguous(memory_format=torch.channels_last_3d)
        xg = bg[:, :, 1:13, 1:41, 1:41]
        full = torch.zeros(1, 64 + cout, 24, 80, 80, device=DEV, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last_3d)
        out = full[:, 64:]
        y = CV.upconv2_fprop(xg, w.to(DEV), None, 0, 0.0, out=out)
        assert y.data_ptr() == out.data_ptr()
    assert _rel(y, ref) <= 1e-2
    assert float(full[:, :64].abs().max()) == 0.0


def test_conv_transpose_module_forward_uses_the_fused_kernel_and_backward_matches():
    import raygun_b200 as rb
    from raygun_b200 import conv as CV, nn as L
    torch.manual_seed(43)
    m = L.ConvTranspose3d(64, 64, 2, stride=2).to(DEV)
    ref = torch.nn.ConvTranspose3d(64, 64, 2, stride=2)
    ref.load_state_dict({k: v.cpu() for k, v in m.state_dict().items()})
    x = _q(torch.randn(1, 64, 4, 8, 8))
    xr = x.clone().requires_grad_(True)
    yr = ref(xr)
    g = _q(torch.randn_like(yr))
    yr.backward(g)
    with rb.precision("bf16"):
        before = CV.stats.get("upconv2", 0)
        xg = x.to(DEV).requires_grad_(True)
        y = m(xg)
        assert CV.stats.get("upconv2", 0) == before + 1
        y.backward(g.to(DEV).to(y.dtype))
    assert _rel(y, yr) <= 2e-2
    assert _rel(xg.grad, xr.grad) <= 2e-2
    assert _rel(m.weight.grad, ref.weight.grad) <= 2e-2
    assert _rel(m.bias.grad, ref.bias.grad) <= 2e-2
