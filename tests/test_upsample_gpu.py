"""K4 convex upsampling: CUDA kernel vs the reference fixture and vs the oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import ppm_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def P():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    import ppmstereo_b200
    return ppmstereo_b200


@pytest.mark.parametrize("kind", ["fp32", "bf16"])
def test_vs_reference_golden(P, golden_dir, kind):
    g = np.load(os.path.join(golden_dir, "convex_upsample.npz"))
    flow = torch.from_numpy(g["flow"])
    if kind == "fp32":
        mask, ref, tol = torch.from_numpy(g["mask"]), torch.from_numpy(g["out_fp32_mask"]), 1e-5
    else:
        mask = torch.from_numpy(g["mask_bf16"].astype(np.int16)).view(torch.bfloat16)
        ref, tol = torch.from_numpy(g["out_bf16_mask"]), 1e-5      # bf16 is storage only: fp32 softmax weights
    got = P.convex_upsample(flow.cuda(), mask.cuda(), 4).cpu()
    assert got.shape == ref.shape and got.dtype == torch.float32
    d = (got - ref).abs()
    assert float(d.max()) <= tol * float(ref.abs().max())
    assert float(d.mean()) <= 1e-5 * float(ref.abs().mean()) + 1e-7


def _torch_sequence(flow, mask, rate=4):
    """The reference's op sequence (models/core/ppmstereo.py:185-197) restated with torch ops, run on the GPU."""
    import torch.nn.functional as F
    n, _, h, w = flow.shape
    m = torch.softmax(mask.view(n, 1, 9, rate, rate, h, w).contiguous(), dim=2)
    up = F.unfold(rate * flow, [3, 3], padding=1).view(n, 2, 9, 1, 1, h, w)
    up = torch.sum(m * up, dim=2).permute(0, 1, 4, 2, 5, 3).contiguous()
    return up.reshape(n, 2, rate * h, rate * w).float()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_vs_torch_cuda_sequence(P, dtype):
    """Pins the kernel to the reference's literal op sequence run by torch on the GPU UNDER AUTOCAST, as
    PPMStereo.forward runs it (ppmstereo.py:617): softmax is an autocast-to-fp32 op, so a bf16 mask yields fp32
    weights (no bf16 rounding of the weights)."""
    g = torch.Generator().manual_seed(77)
    flow = (torch.randn(2, 2, 23, 37, generator=g) * 10).cuda()
    mask = (torch.randn(2, 144, 23, 37, generator=g) * 3).to(dtype).cuda()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        want = _torch_sequence(flow, mask)
        got = P.convex_upsample(flow, mask, 4)
    d = (got - want).abs()
    assert float(d.max()) <= 1e-5 * float(want.abs().max())
    assert float(d.mean()) <= 1e-5 * float(want.abs().mean())


@pytest.mark.parametrize("shape", [(1, 1, 1), (3, 5, 130), (2, 46, 80)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_vs_oracle_shapes(P, shape, dtype):
    N, H, W = shape
    g = torch.Generator().manual_seed(N * 1000 + H * 10 + W)
    flow = torch.randn(N, 2, H, W, generator=g) * 10
    mask = (torch.randn(N, 144, H, W, generator=g) * 3).to(dtype)
    want = O.convex_upsample(flow, mask, 4)
    got = P.convex_upsample(flow.cuda(), mask.cuda(), 4).cpu()
    d = (got - want).abs()
    assert float(d.max()) <= 1e-5 * float(want.abs().max())
    assert float(d.mean()) <= 1e-5 * float(want.abs().mean()) + 1e-7


def test_partition_of_unity_full_size(P):
    """Size-independent property at the 720p 1/4-res shape: the weights of every fine pixel sum to one, so a
    constant flow c upsamples to 4c away from the border, and the op is linear in the flow."""
    N, H, W = 20, 184, 320
    g = torch.Generator(device="cuda").manual_seed(3)
    mask = torch.randn(N, 144, H, W, generator=g, device="cuda")
    const = torch.full((N, 2, H, W), 1.5, device="cuda")
    up = P.convex_upsample(const, mask, 4)
    assert float((up[:, :, 4:-4, 4:-4] - 6.0).abs().max()) <= 1e-5
    a = torch.randn(N, 2, H, W, generator=g, device="cuda")
    b = torch.randn(N, 2, H, W, generator=g, device="cuda")
    lin = P.convex_upsample(a + 2 * b, mask, 4) - (P.convex_upsample(a, mask, 4) + 2 * P.convex_upsample(b, mask, 4))
    assert float(lin.abs().max()) <= 1e-4


def test_rejects_cpu_tensors_and_bad_shapes(P):
    with pytest.raises(RuntimeError):
        P.convex_upsample(torch.zeros(1, 2, 4, 4), torch.zeros(1, 144, 4, 4))
    with pytest.raises(RuntimeError):
        P.convex_upsample(torch.zeros(1, 2, 4, 4).cuda(), torch.zeros(1, 100, 4, 4).cuda())


@pytest.mark.parametrize("shape", [(2, 46, 80), (1, 7, 6), (3, 5, 130)])
def test_bf16_pixel_pair_kernel_equals_one_pixel_kernel(P, shape):
    """bf16 masks with an even W take the pixel-pair kernel (4-byte mask loads); a mask whose storage is only
    2-byte aligned takes the one-pixel kernel.  Same arithmetic per pixel -> bit-identical outputs."""
    N, H, W = shape
    g = torch.Generator(device="cuda").manual_seed(H * 100 + W)
    flow = torch.randn(N, 2, H, W, generator=g, device="cuda") * 10
    n = N * 144 * H * W
    buf = (torch.randn(n + 1, generator=g, device="cuda") * 3).bfloat16()
    odd = buf[1:].view(N, 144, H, W)                   # data_ptr % 4 == 2 -> one-pixel kernel
    even = odd.clone()                                 # fresh allocation -> pixel-pair kernel
    assert odd.data_ptr() % 4 == 2 and even.data_ptr() % 4 == 0
    assert torch.equal(P.convex_upsample(flow, even, 4), P.convex_upsample(flow, odd, 4))
