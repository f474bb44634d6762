"""Output stage: oracle known answers on CPU, kernel == oracle byte for byte on the GPU."""
import numpy as np
import pytest
import torch

from oracle.output_stage import convert_pt_to_numpy as oracle_convert


def test_oracle_known_answers():
    # -1 -> 0, 1 -> 255, 0 -> 127 (127.5 truncates), out-of-range clamps; per dtype the intermediates round differently
    x = torch.tensor([-1.0, 1.0, 0.0, -3.0, 2.5, 0.5, -0.5]).view(1, 1, 1, 7)
    for dt in (torch.float32, torch.float16, torch.bfloat16):
        out = oracle_convert(x.to(dt))[0].reshape(-1)
        assert out.dtype == np.uint8
        assert out.tolist()[:5] == [0, 255, 127, 0, 255]
    assert oracle_convert(x)[0].reshape(-1).tolist() == [0, 255, 127, 0, 255, 191, 63]
    # the intermediates are rounded to the tensor's dtype, so the byte depends on it: (fp32, fp16, bf16)
    for v, want in ((-0.984, [2, 2, 1]), (-0.906, [11, 12, 11]), (-0.992, [1, 0, 0])):
        y = torch.tensor([v]).view(1, 1, 1, 1)
        assert [oracle_convert(y.to(dt))[0].item() for dt in (torch.float32, torch.float16, torch.bfloat16)] == want


def test_oracle_layout():
    x = torch.rand(2, 3, 4, 5) * 2 - 1
    out = oracle_convert(x)
    assert len(out) == 2 and out[0].shape == (4, 5, 3)


def test_product_refuses_cpu_tensor():
    from stablediffusion_b200.serving import convert_pt_to_numpy
    with pytest.raises(RuntimeError):
        convert_pt_to_numpy(torch.zeros(1, 3, 8, 8))


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("shape", [(1, 3, 8, 8), (3, 3, 33, 17), (2, 3, 512, 512), (2, 1, 5, 7)])
def test_kernel_bytes_equal_reference(cuda_device, dtype, shape):
    from stablediffusion_b200.serving import convert_pt_to_numpy
    g = torch.Generator().manual_seed(hash((str(dtype), shape)) % 2 ** 31)
    x = (torch.randn(shape, generator=g) * 0.8).to(dtype)  # ~20% of the values fall outside [-1, 1]
    got = convert_pt_to_numpy(x.to(cuda_device))
    ref = oracle_convert(x)
    assert len(got) == len(ref)
    for a, b in zip(got, ref):
        assert a.dtype == np.uint8 and a.shape == b.shape
        assert np.array_equal(a, b)


@pytest.mark.gpu
def test_every_bf16_value_in_range(cuda_device):
    """Exhaustive over the bf16 grid in [-1.5, 1.5]: every representable pixel value gives the reference's byte."""
    bits = torch.arange(0, 1 << 16, dtype=torch.int32).to(torch.int16)
    x = bits.view(torch.bfloat16)
    x = x[torch.isfinite(x.float()) & (x.float().abs() <= 1.5)]
    n = x.numel() // 3 * 3
    x = x[:n].view(1, 3, 1, n // 3)
    from stablediffusion_b200.serving import convert_pt_to_numpy
    assert np.array_equal(convert_pt_to_numpy(x.to(cuda_device))[0], oracle_convert(x)[0])
